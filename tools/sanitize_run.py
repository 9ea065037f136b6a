#!/usr/bin/env python
"""Small driver for compute-sanitizer (memcheck / racecheck): smoke() plus a 37-world, 40-step
combat rollout with auto-reset, full-team tensors and the event log on (every kernel of the step
runs, incl. the warp-synchronous pool / stream loops and the self re-arming work counters)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import simlib as S
import __graft_entry__ as G
G.smoke()
W = 37
sim = S.GpuSim(W, auto_reset=True, full_team=True, event_log=True)
sim.init()
for s in range(40):
    d, a = S.seek_policy(sim, s, W)
    sim.set_actions(d, a)
    sim.step()
print("sanitize_run done: launches/step", sim.launches(), "stats", sim.episode_stats()[:4],
      "damaged guard bands", sim.damaged_guards())
sim.close()
