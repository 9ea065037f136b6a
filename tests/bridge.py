"""Parity bridge to the UNMODIFIED reference arithmetic (test helper).

Every bit-exact test diffs against the det oracle (the reference sim code with
sinf/cosf/logf/atan2f/asinf routed to csrc/mp_math.h).  This module compares an
implementation with the libm oracle (oracle/_ref/libmpenv_ref.so = the reference's
own arithmetic) STEP BY STEP under the north-star bar:

  * integer exports bit-identical;
  * floats within 1e-4 * max(1, |row|_inf) -- except discontinuity flips, which
    are counted per tensor against a committed budget: a <= 2 ulp difference in a
    transcendental moves a ray that grazes a silhouette edge onto another surface
    (reference sim.cpp:3439-3500 writes that depth; utils.cpp:10-72 picks the
    surface), so a per-sample tolerance cannot hold for a discontinuous query;
  * lidar class flags identical on every sample whose depth agrees.
"""
import numpy as np

import simlib as S

INT_EXPORTS = ("hp", "alive", "magazine", "done", "match_result", "opponent_masks")
LIDAR_EXPORTS = ("fwd_lidar", "rear_lidar")
FLOAT_EXPORTS = ("self_obs", "teammate_obs", "opponent_obs", "opponent_last_known_obs", "self_pos",
                 "teammate_pos", "opponent_pos", "opponent_last_known_pos", "filters_state_obs", "reward")
REL_TOL = 1e-4

# Fraction of samples (lidar: rays; other tensors: elements) allowed outside REL_TOL over a
# whole run.  Measured det-vs-libm, 32 worlds x 1000 steps of the random-action stream
# (`python tests/bridge.py` prints them): fwd_lidar 7.8e-4 of rays (19,236 of 24.6 M; first at
# step 38), rear_lidar 1.6e-4 (960 of 6.2 M), every other tensor <= 5.2e-6 of elements (reward:
# 2 of 384 k).  Budgets are ~2.5x the measurement.
FLIP_BUDGET = {"fwd_lidar": 2e-3, "rear_lidar": 4e-4}
DEFAULT_FLIP_BUDGET = 1.5e-5


class BridgeStats:
    def __init__(self):
        self.samples = {}
        self.flips = {}
        self.first_flip = {}
        self.max_ok_err = {}

    def add(self, name, n, bad, step, err):
        self.samples[name] = self.samples.get(name, 0) + int(n)
        self.flips[name] = self.flips.get(name, 0) + int(bad)
        if bad and name not in self.first_flip:
            self.first_flip[name] = step
        self.max_ok_err[name] = max(self.max_ok_err.get(name, 0.0), float(err))

    def rates(self):
        return {k: self.flips[k] / max(1, self.samples[k]) for k in self.samples}

    def check_budgets(self):
        for k, r in self.rates().items():
            budget = FLIP_BUDGET.get(k, DEFAULT_FLIP_BUDGET)
            assert r <= budget, (f"{k}: {self.flips[k]} of {self.samples[k]} samples outside {REL_TOL} "
                                 f"relative (rate {r:.2e} > budget {budget:.1e}; first at step {self.first_flip.get(k)})")

    def summary(self):
        return {k: {"flip_rate": self.flips[k] / max(1, self.samples[k]), "flips": self.flips[k],
                    "first_flip_step": self.first_flip.get(k), "max_rel_err_within_tol": self.max_ok_err[k]}
                for k in sorted(self.samples)}


def compare_step(ours, ref, step, stats, exact_ints=True):
    """One step's exported tensors of `ours` against `ref` (the libm oracle)."""
    if exact_ints:
        for name in INT_EXPORTS:
            x, y = ours.export(name), ref.export(name)
            assert np.array_equal(x.view(np.int32), y.view(np.int32)), \
                f"step {step}: integer export {name} differs from the libm reference"
    for name in LIDAR_EXPORTS:
        xr, yr = ours.export(name), ref.export(name)
        # tolerance per agent row, as for the other tensors: 1e-4 * max(1, |row|_inf); a miss is
        # depth -1 and depth is clamped to the world diagonal (sim.cpp:3480-3490)
        rays = xr.shape[1] // 4
        scale = np.repeat(np.maximum(1.0, np.abs(yr).max(axis=1)), rays)
        x, y = xr.reshape(-1, 4), yr.reshape(-1, 4)
        err = np.abs(x[:, 0] - y[:, 0]) / scale
        ok = err <= REL_TOL
        assert np.array_equal(x[ok, 1:].view(np.int32), y[ok, 1:].view(np.int32)), \
            f"step {step}: {name} class flags differ on a sample whose depth agrees"
        stats.add(name, len(x), (~ok).sum(), step, err[ok].max() if ok.any() else 0.0)
    for name in FLOAT_EXPORTS:
        x, y = ours.export(name), ref.export(name)
        scale = np.maximum(1.0, np.abs(y).max(axis=1, keepdims=True))
        err = np.abs(x - y) / scale
        ok = err <= REL_TOL
        stats.add(name, x.size, (~ok).sum(), step, err[ok].max() if ok.any() else 0.0)


def run_bridge(make_ours, W=32, steps=1000, seed=10):
    """`make_ours()` -> a _SimBase; rolled out beside Oracle(det=False) on BASELINE.json
    configs[0] (W worlds, random-action stream, autoReset off), compared every step."""
    ours = make_ours()
    ref = S.Oracle(W, rand_seed=seed, det=False, num_threads=0)
    stats = BridgeStats()
    try:
        ours.init(); ref.init()
        compare_step(ours, ref, -1, stats)
        for s in range(steps):
            d, a = S.random_actions(s, W)
            for sim in (ours, ref):
                sim.set_actions(d, a)
                sim.step()
            compare_step(ours, ref, s, stats)
    finally:
        ours.close(); ref.close()
    return stats


if __name__ == "__main__":
    import json
    S.build_oracle()
    st = run_bridge(lambda: S.Oracle(32, det=True, num_threads=0), 32, 1000)
    print(json.dumps(st.summary(), indent=1))
    st.check_budgets()
