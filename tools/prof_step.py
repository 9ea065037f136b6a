#!/usr/bin/env python
"""Short driver for ncu: W worlds, `warm` random-action steps, then `steps` more.
usage: python tools/prof_step.py [worlds] [warm] [steps] [stress]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import simlib as S
W = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
warm = int(sys.argv[2]) if len(sys.argv) > 2 else 60
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
stress = len(sys.argv) > 4 and sys.argv[4] == "stress"
sim = S.GpuSim(W, auto_reset=True, full_team=False)
sim.init()
for s in range(warm + steps):
    mag = sim.export("magazine") if stress else None
    d, a = S.random_actions(s, W, stress=stress, magazine=mag)
    sim.set_actions(d, a)
    sim.step()
print("prof_step done", sim.launches())
sim.close()
