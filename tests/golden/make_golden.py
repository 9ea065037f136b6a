#!/usr/bin/env python
"""Generates tests/golden/rollouts.json and tests/golden/state_w2.npz from the
oracle (the reference's own simulation code compiled from /root/reference by
oracle/Makefile, deterministic-transcendental build).  Run in the container
that has /root/reference:

    python tests/golden/make_golden.py

The files pin the oracle build itself: the GPU box has no /root/reference, so
there the prebuilt oracle/_ref/*.so is checked against these digests before
it is trusted as the checker for the CUDA path.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import simlib as S  # noqa: E402

SPECS = {
    # BASELINE.json configs[0]: 32 worlds, random actions, 1000 steps
    "random_w32": dict(worlds=32, seed=10, auto_reset=False, policy="random", steps=1000,
                       checkpoints=[1, 10, 100, 500, 1000]),
    # scripted seek-and-fire policy: exercises hits, kills, respawn, zone points, episode resets
    "combat_w8": dict(worlds=8, seed=10, auto_reset=True, policy="seek", steps=600,
                      checkpoints=[1, 50, 200, 400, 600]),
}


def digest(sim):
    hs = hashlib.sha256()
    for name in S.STEP_OUTPUTS + S.FULL_TEAM_OUTPUTS:
        hs.update(sim.export(name).tobytes())
    return hs.hexdigest()


def main():
    S.build_oracle()
    out = {}
    for name, spec in SPECS.items():
        W = spec["worlds"]
        o = S.Oracle(W, rand_seed=spec["seed"], auto_reset=spec["auto_reset"], det=True, num_threads=0)
        o.init()
        dg = {"init": digest(o)}
        pol = S.seek_policy if spec["policy"] == "seek" else (lambda sim, s, W: S.random_actions(s, W))
        for s in range(spec["steps"]):
            d, a = pol(o, s, W)
            o.set_actions(d, a)
            o.step()
            if (s + 1) in spec["checkpoints"]:
                dg[str(s + 1)] = digest(o)
        mr = o.export("match_result")
        out[name] = dict(worlds=W, seed=spec["seed"], auto_reset=spec["auto_reset"], policy=spec["policy"],
                         steps=spec["steps"], digests=dg, match_result_sum=int(mr[:, 1:5].sum()))
        if name == "combat_w8":
            # small raw fixture: the first 2 worlds' exported tensors after the rollout
            np.savez_compressed(os.path.join(HERE, "state_combat_w8_first2.npz"),
                                **{n: o.export(n)[: S.rows_of(S.EXPORT_SHAPES[n][0], 2)]
                                   for n in S.STEP_OUTPUTS})
        o.close()
        print(name, out[name]["match_result_sum"], dg)
    json.dump(out, open(os.path.join(HERE, "rollouts.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
