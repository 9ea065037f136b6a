"""mpenv_episode_stats_static: the episode-statistics reduction for callers that must not allocate
(static device scratch instead of a cudaMalloc'ed accumulator; include/mpenv.h).  No counterpart in
the reference.  Collected last (file name) so that a failure here cannot hide another test under -x.

The values are integer sums: the tolerance is the one the other statistics tests use -- none
(np.array_equal with mpenv_episode_stats, and with the host-side sum of the exported tensors)."""
import ctypes as C

import numpy as np
import pytest

import simlib as S

pytestmark = pytest.mark.gpu


def _static_stats(sim, out):
    """One call through the C ABI on the simulator's own stream; `out` is a CUDA int64[16] tensor."""
    L = sim.lib
    L.mpenv_episode_stats_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.mpenv_episode_stats_static.restype = C.c_int
    L.mpenv_stream.restype = C.c_void_p
    L.mpenv_stream.argtypes = [C.c_void_p]
    rc = L.mpenv_episode_stats_static(sim.h, C.c_void_p(L.mpenv_stream(sim.h)), C.c_void_p(out.data_ptr()))
    sim._check(rc, "episode_stats_static")
    sim._check(L.mpenv_sync(sim.h), "sync")
    return out.cpu().numpy().copy()


@pytest.mark.parametrize("W", [37, 700])
def test_static_scratch_stats_match_the_allocating_entry_point(W):
    """Two (and more) calls in a row on the shared static scratch, with and without a step in
    between, one CTA (37 worlds) and several (700): each must equal mpenv_episode_stats."""
    import torch
    sim = S.GpuSim(W, auto_reset=True)
    try:
        out = torch.full((16,), -7, dtype=torch.int64, device="cuda:0")     # stale contents must not survive
        torch.cuda.synchronize()
        sim.init()
        nonzero = 0
        for s in range(30):
            d, a = S.seek_policy(sim, s, W)
            sim.set_actions(d, a)
            sim.step()
            want = sim.episode_stats()
            first = _static_stats(sim, out)
            second = _static_stats(sim, out)                                 # same state: same answer
            assert np.array_equal(first, want), f"step {s}: first call {first} != {want}"
            assert np.array_equal(second, want), f"step {s}: second call {second} != {want}"
            assert np.array_equal(want, S.episode_stats_host(sim)), f"step {s}"
            nonzero += int(np.count_nonzero(want))
        assert nonzero > 30 * 3, "statistics all zero: nothing was compared"
        assert sim.damaged_guards() == 0
    finally:
        sim.close()


def test_static_scratch_two_handles_back_to_back():
    """Two simulators on one GPU share the scratch; calls ordered by the host do not disturb each other."""
    import torch
    a, b = S.GpuSim(64, auto_reset=True), S.GpuSim(300, auto_reset=True, rand_seed=11)
    try:
        out = torch.zeros(16, dtype=torch.int64, device="cuda:0")
        torch.cuda.synchronize()
        a.init(); b.init()
        for s in range(5):
            for sim, W in ((a, 64), (b, 300)):
                d, act = S.random_actions(s, W)
                sim.set_actions(d, act)
                sim.step()
            for sim in (a, b, a):
                assert np.array_equal(_static_stats(sim, out), sim.episode_stats()), f"step {s}"
    finally:
        a.close(); b.close()


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    """More worlds than MPENV_STATIC_STATS_MAX_WORLDS (65536): a clear error, nothing launched."""
    import torch
    sim = S.GpuSim(65536 + 12, auto_reset=True, full_team=False)
    try:
        out = torch.full((16,), -7, dtype=torch.int64, device="cuda:0")
        torch.cuda.synchronize()
        with pytest.raises(RuntimeError, match="exceed the static scratch's maximum of 65536"):
            _static_stats(sim, out)
        assert (out.cpu().numpy() == -7).all()
    finally:
        sim.close()
