"""world_size-2 test of the N>1 path on CPU (gloo): each rank owns a shard of the
world batch (global world ids via world_id_offset, no data-path collective) and
the ONLY exchange is the sum-allreduce of the 16 episode statistics
(madrona_mp_env.dist).  The shards run the hostsim harness (the kernels' stage
functions compiled for the host); the sum must equal the statistics of one big
batch, and each shard must equal its slice of the big batch bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

import simlib as S

W_PER_RANK = 3
STEPS = 60


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, os.path.join(S.ROOT, "madrona-mp-env_b200"))
    import torch
    import torch.distributed as dist
    from madrona_mp_env import dist as D
    dist.init_process_group("gloo", rank=rank, world_size=world)
    r, n, _ = D.rank_info()
    off = D.shard_world_offset(r, W_PER_RANK)
    sim = S.HostSim(W_PER_RANK, auto_reset=True, world_id_offset=off)
    sim.init()
    total = W_PER_RANK * n
    for s in range(STEPS):
        d, a = S.random_actions(s, total)
        lo, hi = off * S.AGENTS, (off + W_PER_RANK) * S.AGENTS
        sim.set_actions(d[lo:hi], a[lo:hi])
        sim.step()
    local = torch.from_numpy(S.episode_stats_host(sim))
    glob = D.allreduce_episode_stats(local.clone())
    np.save(os.path.join(out_dir, f"stats_{rank}.npy"), glob.numpy())
    np.save(os.path.join(out_dir, f"local_{rank}.npy"), local.numpy())
    np.save(os.path.join(out_dir, f"selfobs_{rank}.npy"), sim.export("self_obs"))
    sim.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_and_stats_allreduce(built, tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    big = S.HostSim(2 * W_PER_RANK, auto_reset=True)
    big.init()
    for s in range(STEPS):
        d, a = S.random_actions(s, 2 * W_PER_RANK)
        big.set_actions(d, a)
        big.step()
    want = S.episode_stats_host(big)
    g0, g1 = np.load(tmp_path / "stats_0.npy"), np.load(tmp_path / "stats_1.npy")
    assert np.array_equal(g0, g1), "allreduce result differs between ranks"
    assert np.array_equal(g0, want), f"sum over shards {g0} != one big batch {want}"
    assert g0[0] == 2 * W_PER_RANK
    l0, l1 = np.load(tmp_path / "local_0.npy"), np.load(tmp_path / "local_1.npy")
    assert np.array_equal(l0 + l1, g0)
    so = big.export("self_obs")
    half = len(so) // 2
    assert np.array_equal(so[:half].view(np.int32), np.load(tmp_path / "selfobs_0.npy").view(np.int32))
    assert np.array_equal(so[half:].view(np.int32), np.load(tmp_path / "selfobs_1.npy").view(np.int32))
    big.close()


def test_allreduce_shape_check():
    import torch
    sys.path.insert(0, os.path.join(S.ROOT, "madrona-mp-env_b200"))
    from madrona_mp_env import dist as D
    with pytest.raises(ValueError):
        D.allreduce_episode_stats(torch.zeros(4, dtype=torch.int64))
    t = torch.arange(16, dtype=torch.int64)
    assert D.allreduce_episode_stats(t) is t            # no process group: identity
    assert D.stats_dict(t)["episodes_finished"] == 1
