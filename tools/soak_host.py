#!/usr/bin/env python
"""CPU-only parity soak: the host build of the stage logic (tests/hostsim) vs the det oracle.
usage: python tools/soak_host.py [worlds] [steps] [seed] [sim_flags] [random|stress] [world_id_offset] [pillars]"""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import simlib as S
from test_hostsim_parity import assert_same
arg = lambda i, d: type(d)(sys.argv[i]) if len(sys.argv) > i else d
W, steps, seed, flags, mode, off, pillars = arg(1, 16), arg(2, 3100), arg(3, 21), arg(4, 0), arg(5, "random"), arg(6, 0), arg(7, 0)
scene = S.SCENE
if pillars:
    scene = S.make_scene(os.path.join(tempfile.mkdtemp(), "pillars"), extra_boxes=S.pillar_boxes(pillars, seed=seed))
host = S.HostSim(W, rand_seed=seed, sim_flags=flags, auto_reset=True, full_team=False, world_id_offset=off, scene=scene)
orc = S.Oracle(W, rand_seed=seed, sim_flags=flags, auto_reset=True, det=True, num_threads=2, world_id_offset=off, scene=scene)
host.init(); orc.init()
for s in range(steps):
    if (s // 100) % 2 == 0:
        d, a = S.seek_policy(orc, s, W, world_id_offset=off, seed=77 + seed)
    elif mode == "stress":
        d, a = S.random_actions(s, W, off, stress=True, magazine=orc.export("magazine"))
    else:
        d, a = S.random_actions(s, W, off, seed=77 + seed)
    for sim in (host, orc):
        sim.set_actions(d, a); sim.step()
    if s % 25 == 24:
        assert_same(host, orc, W, f"step {s}", full_team=False)
print(f"host soak ok: {W} worlds (offset {off}) x {steps} steps (seed {seed}, flags {flags}, {mode}, {pillars} pillars) bit-identical")
