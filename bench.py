#!/usr/bin/env python
"""bench.py -- throughput of the MadronaMPEnv per-step simulation loop.

    python bench.py --gpus N --steps K --warmup W            (our arm: libmpenv_b200.so)
    python bench.py --impl reference --steps K --warmup W    (reference sim code on host cores)
    python bench.py --workload stress --worlds 16384         (BASELINE.json configs[3])
    python bench.py --worlds 4096                            (configs[4] batch, DLPack-aliased tensors)

Metric (BASELINE.json): agent-steps/sec at 8192 worlds per GPU on data/simple_map,
random actions, all train-interface tensors written every step.  One JSON line
on stdout (rank 0).  See DESIGN.md "Measurement" for the definitions.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "madrona-mp-env_b200"))

WORLDS_PER_GPU = 8192
AGENTS = 12
METRIC = "agent-steps/sec at 8192 worlds/GPU (simple_map, random actions, full obs export)"

# Algorithmic bytes per agent-step of each kernel: what THAT kernel must read and write once
# (DESIGN.md "Kernels").  They sum to 6,136 B, more than the step's 5,369 B,
# because state rows that two kernels both need (positions, aim, combat state, visibility)
# are counted by each: a per-kernel roofline charges a kernel for its own traffic.
KERNEL_BYTES = {"k_move": 252, "k_fire": 230, "k_world": 600, "k_los": 110, "k_obs": 3600,
                "k_fullteam": 1700, "k_lidar": 1344}
STEP_BYTES = 5369           # every state row / export once per step


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def csrc_digest():
    """sha256 over the kernel sources: ties profiles/ncu_traffic.json to the code it was captured on
    (there is no .git on the GPU box)."""
    d = os.path.join(ROOT, "madrona-mp-env_b200", "csrc")
    h = hashlib.sha256()
    for f in sorted(os.listdir(d)):
        if f.endswith((".h", ".cuh", ".cu", ".cpp")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


class ClockSampler(threading.Thread):
    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.stop_flag = gpu, [], False
        self.max_mhz = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                f = [x.strip() for x in out.split(",")]
                self.samples.append((float(f[0]), f[2:]))
                self.max_mhz = float(f[1])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for _, r in self.samples for i, v in enumerate(r) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median([s for s, _ in self.samples])), "sm_max_mhz": self.max_mhz,
                "reasons": reasons}


def gen_actions_torch(torch, steps, n, device, seed):
    """Synthetic random-action stream (distribution of SURVEY 8d), generated on the device."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    u = lambda: torch.rand((steps, n), generator=g, device=device)
    move_amount = torch.randint(0, 3, (steps, n), generator=g, device=device, dtype=torch.int32)
    move_angle = torch.randint(0, 8, (steps, n), generator=g, device=device, dtype=torch.int32)
    uf, us = u(), u()
    fire = torch.where(uf < 0.6, 0, torch.where(uf < 0.95, 1, 2)).to(torch.int32)
    stand = torch.where(us < 0.8, 0, torch.where(us < 0.9, 1, 2)).to(torch.int32)
    yaw = torch.randint(0, 13, (steps, n), generator=g, device=device, dtype=torch.int32)
    pitch = torch.randint(0, 7, (steps, n), generator=g, device=device, dtype=torch.int32)
    return (torch.stack([move_amount, move_angle, fire, stand], dim=2).contiguous(),
            torch.stack([yaw, pitch], dim=2).contiguous())


def workload_name(args):
    acts = ("random actions" if args.workload == "random" else
            "raycast stress: every agent fires every step, reloads when its magazine reads 0")
    cfg = {("random", 8192): "BASELINE.json configs[1]", ("stress", 16384): "BASELINE.json configs[3]",
           ("random", 4096): "BASELINE.json configs[4] batch size"}.get((args.workload, args.worlds), "variant")
    return (f"data/simple_map, {args.worlds} worlds/GPU, Task::Zone 6v6, {acts}, autoReset, "
            f"all train-interface tensors written every step ({cfg})")


def run_ours(args):
    import torch
    import torch.distributed as dist
    import madrona_mp_env as mp
    from madrona_mp_env import exports as X

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    W = args.worlds
    N = W * AGENTS
    stress = args.workload == "stress"

    mgr = mp.SimManager(mp.ExecMode.CUDA, local, W, 10, True, 0, mp.Task.Zone, 6, 0, 1,
                        X.SCENE, world_id_offset=rank * W, enable_full_team=False)
    lib, h = mgr._lib, mgr._h
    # zero-copy aliases of the simulator's tensors (DLPack / __cuda_array_interface__), as the
    # reference's training loop uses them (scripts/policy.py:105-113)
    act_d = mgr.pvp_discrete_action_tensor().to_torch()
    act_a = mgr.pvp_discrete_aim_action_tensor().to_torch()
    magazine = mgr.magazine_tensor().to_torch()
    total = args.steps + args.warmup
    ring = min(total, 256)
    all_d, all_a = gen_actions_torch(torch, ring, N, dev, 1234 + rank)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    sp = stream.cuda_stream

    mgr.init_async(sp)
    # the one collective of the path (episode statistics): at least 4 inside any timed region
    stats_every = max(1, min(100, args.steps // 4))
    coll = {"calls": 0, "events": []}
    one = torch.ones((), dtype=torch.int32, device=dev)
    two = one * 2

    def set_actions(i):
        act_d.copy_(all_d[i % ring], non_blocking=True)
        act_a.copy_(all_a[i % ring], non_blocking=True)
        if stress:
            # configs[3]: fire = 1; fire = 2 (reload) when the magazine tensor reads 0 (sim.cpp:1443-1615)
            act_d[:, 2] = torch.where(magazine[:, 0] == 0, two, one)

    def one_step(i, timed=False):
        set_actions(i)
        mgr.step_async(sp)
        if (i + 1) % stats_every == 0:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) if timed else None
            if ev:
                ev[0].record()
            st = mgr.episode_stats(sp)          # k_stats: int64[16] per device
            if world > 1:
                dist.all_reduce(st)             # NCCL sum over NVLink
            if ev:
                ev[1].record()
                coll["calls"] += 1
                coll["events"].append(ev)

    kms = (C.c_float * 8)()
    knames = (C.c_char_p * 8)()
    cnt = C.c_uint32()
    lib.mpenv_step_timed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
    acc, acc_n, timed_kernel_steps = {}, [0], []

    def timed_kernel_step(i):
        """One step as individual launches with an event pair per kernel (mpenv_step_timed)."""
        set_actions(i)
        torch.cuda.synchronize()
        assert lib.mpenv_step_timed(h, kms, knames, 8, C.byref(cnt)) == 0
        for k in range(cnt.value):
            acc[knames[k].decode()] = acc.get(knames[k].decode(), 0.0) + kms[k]
        acc_n[0] += 1
        timed_kernel_steps.append(i)

    # warm-up; its last steps are per-kernel timed, so kernel_ms brackets the timed region
    # (the step cost drifts with episode phase: agents spread out, more line-of-sight rays)
    kreps = min(5, args.warmup)
    for i in range(args.warmup - kreps):
        one_step(i)
    for i in range(args.warmup - kreps, args.warmup):
        timed_kernel_step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(args.steps):
        one_step(args.warmup + i, timed=True)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    sampler.stop_flag = True
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * N * args.steps / (ms * 1e-3)
    coll_us = [a.elapsed_time(b) * 1e3 for a, b in coll["events"]]
    collective = {"calls": coll["calls"], "every_steps": stats_every, "bytes": 16 * 8,
                  "us_per_call": round(float(np.mean(coll_us)), 1) if coll_us else None,
                  "what": "k_stats (episode statistics, int64[16] per device)" +
                          (" + NCCL all_reduce(sum)" if world > 1 else "; no all_reduce at 1 GPU")}

    # ---- per-kernel durations: the steps right after the timed region too -------------------
    for i in range(args.warmup + args.steps, args.warmup + args.steps + kreps):
        timed_kernel_step(i)
    acc = {k: v / acc_n[0] for k, v in acc.items()}
    dom = max(acc, key=acc.get)
    peak, peak_src = peaks()
    achieved = KERNEL_BYTES[dom] * N / (acc[dom] * 1e-3) / 1e9
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if tj.get("worlds") == W and tj.get("csrc_sha16") == csrc_digest():
            traffic = tj["bytes_per_launch"].get(dom)
            traffic_src = tj.get("source")
        else:
            traffic_src = ("profiles/ncu_traffic.json was captured on other kernel sources / batch size "
                           f"(csrc_sha16 {tj.get('csrc_sha16')} vs {csrc_digest()}): not reported")
    step_gbs = value / world * STEP_BYTES / 1e9
    roofline = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 2), "peak": peak,
                "peak_source": peak_src, "unit": "GB/s", "frac": round(achieved / peak, 5),
                "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": KERNEL_BYTES[dom] * N,
                "bytes_per_agent_step": KERNEL_BYTES[dom],
                "kernel_ms": {k: round(v, 4) for k, v in acc.items()},
                "kernel_ms_steps": f"mpenv_step_timed on steps {timed_kernel_steps[0]}..{timed_kernel_steps[kreps - 1]} and "
                                   f"{timed_kernel_steps[kreps]}..{timed_kernel_steps[-1]} after reset "
                                   f"(the timed region is steps {args.warmup}..{args.warmup + args.steps - 1})",
                "step": {"bytes_per_agent_step": STEP_BYTES, "achieved_gbs": round(step_gbs, 2),
                         "frac": round(step_gbs / peak, 5),
                         "note": "SURVEY 8(d): per-GPU agent-steps/s x 5,369 B / HBM peak; the path is "
                                 "instruction-issue bound in BVH traversal, not HBM bound"}}

    # ---- e2e: C ABI with HOST buffers, copies inside the timed region --------
    e2e = run_e2e(args, torch, dist, lib, h, X, all_d, all_a, rank, world, W, N, dev)

    cpu_baseline = None
    census = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import simlib as S
        cpu_baseline = run_cpu_sample(S, args, cores=0)
        try:
            census = ray_census(S, local, stress=stress)
        except Exception as exc:           # the census is a side measurement: never fail the bench line
            census = {"unavailable": str(exc)[:200]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": "agent-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": workload_name(args),
                       "worlds_per_gpu": W, "agents_per_world": AGENTS, "parallelism": f"worlds sharded x{world}",
                       "timed_steps": [args.warmup, args.warmup + args.steps],
                       "coupling": "actions written and magazine read through torch tensors aliasing the "
                                   "simulator's buffers (DLPack zero-copy), stream-ordered mpenv_step_async",
                       "l2": f"per-step working set ~{W * AGENTS * 6136 / 1e6:.0f} MB (> 126 MB L2): inputs larger than L2, "
                             "no explicit flush"},
            "clocks": sampler.summary() if rank == 0 else None,
            "e2e": e2e, "gpu_launches": mgr_launches(lib, h) * args.steps + coll["calls"],
            "collective": collective,
            "roofline": roofline, "cpu_baseline": cpu_baseline,
            "rays_per_agent_step": census,
            "bvh_rays_per_s": (round(value * census["rays"], 1) if census and "rays" in census
                               else None),
            "sphere_casts_per_s": (round(value * census["sphere_casts"], 1) if census and "rays" in census
                                   else None),
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_e2e(args, torch, dist, lib, h, X, all_d, all_a, rank, world, W, N, dev):
    """Same metric through the C ABI with HOST buffers: every step, pinned host actions in
    (mpenv_write_tensor_async), mpenv_step_async, and the 18 train-interface outputs back to
    pinned host memory (mpenv_read_tensors_pipelined: device snapshot, then D2H on a copy
    stream overlapping the next step).  Also measures the box's pinned D2H ceiling with every
    rank copying at once, so the number can be read against what PCIe allows."""
    out_names = [n for n in X.STEP_OUTPUTS if n != "reward_hyper_params"]
    lib.mpenv_host_alloc.argtypes = [C.c_uint64, C.POINTER(C.c_void_p)]
    lib.mpenv_host_free.argtypes = [C.c_void_p]
    lib.mpenv_write_tensor_async.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
    lib.mpenv_read_tensors_pipelined.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
    lib.mpenv_read_wait.argtypes = [C.c_void_p]
    lib.mpenv_stream.restype = C.c_void_p
    lib.mpenv_stream.argtypes = [C.c_void_p]
    lib.mpenv_sync.argtypes = [C.c_void_p]
    sizes = [X.rows_of(X.EXPORT_SHAPES[n][0], W) * X.EXPORT_SHAPES[n][1] * 4 for n in out_names]
    d2h = sum(sizes)
    # one pinned allocation (cudaMallocHost through the C ABI) carved into the output buffers
    base = C.c_void_p()
    assert lib.mpenv_host_alloc(d2h + 256 * len(sizes), C.byref(base)) == 0
    offs, o = [], 0
    for s in sizes:
        offs.append(o)
        o += (s + 255) & ~255
    hd, ha = all_d[:8].cpu().pin_memory(), all_a[:8].cpu().pin_memory()
    h2d = hd[0].numel() * 4 + ha[0].numel() * 4
    ids = (C.c_uint32 * len(sizes))(*[X.EXPORT_IDS[n] for n in out_names])
    dsts = (C.c_void_p * len(sizes))(*[base.value + x for x in offs])
    hs = C.c_void_p(lib.mpenv_stream(h))

    lib.mpenv_snapshot_tensors_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    lib.mpenv_read_snapshot_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]

    def e2e_step(i):
        # host actions in, step, host observations out.  The device snapshot of step i is queued behind
        # the step at once (double-buffered), so it is taken while step i-1 is still crossing PCIe; its
        # own transfer starts as soon as the host buffers are free again.
        lib.mpenv_write_tensor_async(h, X.EXPORT_IDS["pvp_discrete_action"], hd[i % 8].data_ptr(), hd[0].numel() * 4)
        lib.mpenv_write_tensor_async(h, X.EXPORT_IDS["pvp_discrete_aim_action"], ha[i % 8].data_ptr(), ha[0].numel() * 4)
        lib.mpenv_step_async(h, hs)
        assert lib.mpenv_snapshot_tensors_async(h, ids, len(sizes)) == 0
        lib.mpenv_read_wait(h)              # step i-1's tensors are now on the host
        assert lib.mpenv_read_snapshot_async(h, dsts, len(sizes)) == 0
    k2 = max(5, min(args.steps, 50))
    for i in range(3):
        e2e_step(i)
    lib.mpenv_read_wait(h); lib.mpenv_sync(h)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(k2):
        e2e_step(i)
    lib.mpenv_read_wait(h); lib.mpenv_sync(h)
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt = float(tt.item())

    # PCIe ceiling: the same 18 tensors into the same pinned buffers through the C ABI's plain
    # read (mpenv_read_tensor_async), all ranks at once, no kernels running (outside every
    # timed region)
    lib.mpenv_read_tensor_async.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
    reps = 8

    def copy_all():
        for j in range(len(sizes)):
            assert lib.mpenv_read_tensor_async(h, ids[j], dsts[j], sizes[j]) == 0
    copy_all(); copy_all()
    lib.mpenv_sync(h)
    if world > 1:
        dist.barrier()
    c0 = time.perf_counter()
    for _ in range(reps):
        copy_all()
    lib.mpenv_sync(h)
    cdt = time.perf_counter() - c0
    ct = torch.tensor([cdt], device=dev)
    if world > 1:
        dist.all_reduce(ct, op=dist.ReduceOp.MAX)
    ceil_gbs = d2h * reps / float(ct.item()) / 1e9          # per rank, all ranks copying at once
    lib.mpenv_host_free(base)
    got_gbs = d2h * k2 / dt / 1e9
    return {"value": round(world * N * k2 / dt, 1), "unit": "agent-steps/s",
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": k2,
            "d2h_gbs_per_rank": round(got_gbs, 2),
            "pcie_gbs": round(ceil_gbs, 2),
            "frac_of_ceiling": round(got_gbs / ceil_gbs, 3),
            "ceiling_agent_steps_per_s": round(world * ceil_gbs * 1e9 / (d2h / N), 1),
            "note": "C ABI, pinned host buffers (mpenv_host_alloc), every step: mpenv_write_tensor_async(actions) + "
                    "mpenv_step_async + mpenv_snapshot_tensors_async + mpenv_read_snapshot_async (18 output tensors, "
                    "double-buffered device snapshot, D2H on a copy stream overlapping the next step); host wall clock around k steps incl. the last read. "
                    "pcie_gbs = pinned D2H rate per rank of the same bytes with all ranks copying at once and no "
                    "kernels running; the e2e number is bound by it, not by the kernels"}


def ray_census(S, gpu_id, worlds=2048, warmup=60, steps=40, stress=False):
    """Rays per agent-step from the INSTRUMENTED twin library (device counters where the
    reference calls traceRayAgainstWorld / sphereCast), on the same workload at a
    smaller batch.  Never timed: the rate is this count x the timed agent-steps/s."""
    sim = S.GpuSim(worlds, auto_reset=True, full_team=False, gpu_id=gpu_id, instrumented=True)
    try:
        sim.init()
        for s in range(warmup + steps):
            if s == warmup:
                sim.ray_counts(reset=True)
            d, a = S.random_actions(s, worlds, stress=stress, magazine=sim.export("magazine") if stress else None)
            sim.set_actions(d, a)
            sim.step()
        c = sim.ray_counts().astype(float) / (worlds * AGENTS * steps)
    finally:
        sim.close()
    return {"lidar": round(c[0], 3), "line_of_sight": round(c[1], 3), "hitscan": round(c[2], 3),
            "sphere_casts": round(c[3], 3), "rays": round(c[0] + c[1] + c[2], 3),
            "sample": f"{worlds} worlds x {steps} steps after {warmup} warm-up, libmpenv_b200_count.so"}


def mgr_launches(lib, h):
    a, b = C.c_uint32(), C.c_uint32()
    lib.mpenv_launches_per_step.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.mpenv_launches_per_step(h, C.byref(a), C.byref(b))
    return a.value


REF_BUILD = ("reference sim.cpp/utils.cpp/level_gen.cpp/mesh_bvh.inl compiled in place under oracle/shim, "
             "g++ -O3 -march=x86-64-v3 -ffp-contract=off, libm arithmetic (oracle/_ref/libmpenv_ref_fast.so), "
             "one world per task over all host threads")


def cpu_actions(S, args, orc, s, worlds):
    stress = args.workload == "stress"
    return S.random_actions(s, worlds, stress=stress, magazine=orc.export("magazine") if stress else None)


def run_cpu_sample(S, args, cores=0, steps=6, warmup=2):
    """The reference's own sim code (-O3 build, libm arithmetic) under the shim executor on the
    bench's own batch size; a bounded sample of steps."""
    worlds = args.worlds
    orc = S.Oracle(worlds, auto_reset=True, det=False, num_threads=cores, fast=True)
    orc.init()
    # the random stream is generated before the clock starts (the stress stream reads the magazine)
    pre = [S.random_actions(s, worlds) for s in range(warmup + steps)] if args.workload == "random" else None
    act = lambda s: pre[s] if pre else cpu_actions(S, args, orc, s, worlds)
    for s in range(warmup):
        orc.set_actions(*act(s)); orc.step()
    t0 = time.perf_counter()
    for s in range(warmup, warmup + steps):
        orc.set_actions(*act(s)); orc.step()
    dt = time.perf_counter() - t0
    n = orc.num_threads()
    orc.close()
    return {"value": round(worlds * AGENTS * steps / dt, 1), "unit": "agent-steps/s", "cores": int(n),
            "kind": "reference",
            "sample": f"{worlds} worlds x {steps} steps of the same workload after {warmup} warm-up steps "
                      f"({REF_BUILD}; {n} threads, {dt:.1f} s)"}


def run_reference(args):
    """bench.py --impl reference: the reference's CPU implementation of the step on the host
    cores, on OUR arm's config (same worlds per GPU, same workload); each bench step is one
    simulation step of the whole batch."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import simlib as S
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    worlds = args.worlds
    orc = S.Oracle(worlds, auto_reset=True, det=False, num_threads=0, fast=True)
    orc.init()
    acts = [S.random_actions(s, worlds) for s in range(8)] if args.workload == "random" else None
    k = 0

    def step():
        nonlocal k
        orc.set_actions(*(acts[k % 8] if acts else cpu_actions(S, args, orc, k, worlds)))
        orc.step()
        k += 1
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    n = int(orc.num_threads())
    v = round(worlds * AGENTS * args.steps / dt, 1)
    sample = f"each step = 1 simulation step of {worlds} worlds; {REF_BUILD}; {n} host threads"
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "agent-steps/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(dt * 1e3 / args.steps, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "worlds_per_gpu": worlds, "agents_per_world": AGENTS,
                   "note": "host cores only (rank 0); the same batch as one GPU of our arm"},
        "cpu_baseline": {"value": v, "unit": "agent-steps/s", "cores": n, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


_JSON_OUT = None


def emit(line):
    """The ONE JSON line, on the process's original stdout."""
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


def main():
    # Libraries write to fd 1 behind Python's back (NCCL prints its version banner there when
    # NCCL_DEBUG is set): keep the original stdout for the JSON line only, send the rest to stderr.
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--worlds", type=int, default=WORLDS_PER_GPU)
    ap.add_argument("--workload", default="random", choices=["random", "stress"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
