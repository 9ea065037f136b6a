#!/usr/bin/env python
"""Long parity soak (not part of the suite): GPU vs the det oracle, seek policy (combat), auto-reset,
forked graph (no full-team tensors).  usage: python tools/soak.py [worlds] [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import simlib as S
from test_gpu_parity import assert_same
W = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
gpu = S.GpuSim(W, auto_reset=True, full_team=False)
orc = S.Oracle(W, auto_reset=True, det=True, num_threads=0)
gpu.init(); orc.init()
kills = 0
for s in range(steps):
    d, a = S.seek_policy(orc, s, W) if (s // 100) % 2 == 0 else S.random_actions(s, W)
    for sim in (gpu, orc):
        sim.set_actions(d, a); sim.step()
    if s % 50 == 49:
        assert_same(gpu, orc, W, f"step {s}", full_team=False)
        kills = int(orc.export("match_result")[:, 1:3].sum())
print(f"soak ok: {W} worlds x {steps} steps bit-identical, guards damaged = {gpu.damaged_guards()}, kills in the last episodes = {kills}")
