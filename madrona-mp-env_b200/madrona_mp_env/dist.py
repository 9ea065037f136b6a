"""Multi-GPU plumbing for the step path: one process per GPU, the
world batch sharded by rank, and ONE collective -- a sum-allreduce of the 16
int64 episode statistics -- over torch.distributed (NCCL on NVLink on the GPU
box, gloo in the CPU tests).  The reference has no multi-GPU path at all
(Manager::Config::gpuID selects one device, mgr.hpp:35); worlds never read each
other (sim.cpp), so nothing else is exchanged.
"""
import os

NUM_STATS = 16
STAT_NAMES = ("worlds", "episodes_finished", "wins_team0", "wins_team1", "draws", "kills_team0",
              "kills_team1", "points_team0", "points_team1", "agents_alive", "sum_done",
              "sum_reward_micro", "sum_cur_step")


def rank_info():
    """(rank, world_size, local_rank) from the torchrun environment."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def shard_world_offset(rank, worlds_per_gpu):
    """Global id of this rank's local world 0.  The episode RNG key is
    split_i(initKey, episode, GLOBAL world id) (sim.cpp:743-746), so rank r of an
    N x worlds_per_gpu job reproduces worlds [r*W, (r+1)*W) of one big batch."""
    if rank < 0 or worlds_per_gpu <= 0:
        raise ValueError("rank must be >= 0 and worlds_per_gpu > 0")
    off = rank * worlds_per_gpu
    if off + worlds_per_gpu > (1 << 32):
        raise ValueError("global world id exceeds 32 bits")
    return off


def allreduce_episode_stats(stats, group=None):
    """In-place sum over ranks of the int64[16] statistics vector (the device buffer
    SimManager.episode_stats() returns, or any int64 tensor of that shape).
    Returns it as a {name: int} dict only on request via stats_dict()."""
    import torch.distributed as dist
    if tuple(stats.shape) != (NUM_STATS,):
        raise ValueError(f"episode statistics must have shape ({NUM_STATS},), got {tuple(stats.shape)}")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def stats_dict(stats):
    v = stats.detach().cpu().tolist()
    return {n: int(v[i]) for i, n in enumerate(STAT_NAMES)}
