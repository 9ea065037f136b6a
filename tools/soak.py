#!/usr/bin/env python
"""Long parity soak (not part of the suite): GPU vs the det oracle, phases of seek policy (combat) and random
actions (or the stress stream), auto-reset, forked graph (no full-team tensors).
usage: python tools/soak.py [worlds] [steps] [seed] [sim_flags] [random|stress] [pillars]"""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import simlib as S
from test_gpu_parity import assert_same
arg = lambda i, d: type(d)(sys.argv[i]) if len(sys.argv) > i else d
W, steps, seed, flags, mode, pillars = arg(1, 64), arg(2, 2000), arg(3, 10), arg(4, 0), arg(5, "random"), arg(6, 0)
scene = S.SCENE
if pillars:
    scene = S.make_scene(os.path.join(tempfile.mkdtemp(), "pillars"), extra_boxes=S.pillar_boxes(pillars))
gpu = S.GpuSim(W, rand_seed=seed, sim_flags=flags, auto_reset=True, full_team=False, scene=scene)
orc = S.Oracle(W, rand_seed=seed, sim_flags=flags, auto_reset=True, det=True, num_threads=0, scene=scene)
gpu.init(); orc.init()
for s in range(steps):
    if (s // 100) % 2 == 0:
        d, a = S.seek_policy(orc, s, W, seed=77 + seed)
    elif mode == "stress":
        d, a = S.random_actions(s, W, stress=True, magazine=orc.export("magazine"))
    else:
        d, a = S.random_actions(s, W, seed=77 + seed)
    for sim in (gpu, orc):
        sim.set_actions(d, a); sim.step()
    if s % 50 == 49:
        assert_same(gpu, orc, W, f"step {s}", full_team=False)
print(f"soak ok: {W} worlds x {steps} steps (seed {seed}, flags {flags}, {mode}, {pillars} pillars, kernels {gpu.kernel_variants()}) "
      f"bit-identical, guards damaged = {gpu.damaged_guards()}")
