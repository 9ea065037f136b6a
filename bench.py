#!/usr/bin/env python
"""bench.py -- throughput of the MadronaMPEnv per-step simulation loop.

    python bench.py --gpus N --steps K --warmup W            (our arm: libmpenv_b200.so)
    python bench.py --impl reference --steps K --warmup W    (reference sim code on host cores)

Metric (BASELINE.json): agent-steps/sec at 8192 worlds per GPU on data/simple_map,
random actions, all train-interface tensors written every step.  One JSON line
on stdout (rank 0).  See DESIGN.md "Measurement" for the definitions.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORLDS_PER_GPU = 8192
AGENTS = 12
METRIC = "agent-steps/sec at 8192 worlds/GPU (simple_map, random actions, full obs export)"

# Algorithmic bytes per agent-step of each kernel (state read + state / export
# written once; derivation in DESIGN.md "Kernels").
KERNEL_BYTES = {"k_move": 252, "k_fire": 230, "k_world": 600, "k_los": 110, "k_obs": 3600,
                "k_fullteam": 1700, "k_lidar": 1344}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


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


def run_ours(args):
    import torch
    import torch.distributed as dist
    import madrona_mp_env as mp
    import simlib as S

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    W = args.worlds
    N = W * AGENTS

    mgr = mp.SimManager(mp.ExecMode.CUDA, local, W, 10, True, 0, mp.Task.Zone, 6, 0, 1,
                        S.SCENE, world_id_offset=rank * W, enable_full_team=False)
    lib, h = mgr._lib, mgr._h
    act_d = mgr.pvp_discrete_action_tensor().to_torch()
    act_a = mgr.pvp_discrete_aim_action_tensor().to_torch()
    total = args.steps + args.warmup
    ring = min(total, 256)
    all_d, all_a = gen_actions_torch(torch, ring, N, dev, 1234 + rank)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    sp = stream.cuda_stream

    mgr.init_async(sp)
    stats_every = 100

    def one_step(i):
        act_d.copy_(all_d[i % ring], non_blocking=True)
        act_a.copy_(all_a[i % ring], non_blocking=True)
        mgr.step_async(sp)
        if world > 1 and (i + 1) % stats_every == 0:
            st = mgr.episode_stats(sp)
            dist.all_reduce(st)         # the only collective: episode statistics over NVLink

    for i in range(args.warmup):
        one_step(i)
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
        one_step(args.warmup + i)
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

    # ---- per-kernel durations (CUDA events on the launching stream) ----------
    kms = (C.c_float * 8)()
    knames = (C.c_char_p * 8)()
    cnt = C.c_uint32()
    lib.mpenv_step_timed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
    acc = {}
    reps = 10
    for _ in range(reps):
        assert lib.mpenv_step_timed(h, kms, knames, 8, C.byref(cnt)) == 0
        for k in range(cnt.value):
            acc[knames[k].decode()] = acc.get(knames[k].decode(), 0.0) + kms[k] / reps
    dom = max(acc, key=acc.get)
    peak, peak_src = peaks()
    achieved = KERNEL_BYTES[dom] * N / (acc[dom] * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if tj.get("worlds") == W:
            traffic = tj["bytes_per_launch"].get(dom)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 2), "peak": peak,
                "peak_source": peak_src, "unit": "GB/s", "frac": round(achieved / peak, 5),
                "traffic": traffic, "algorithmic_bytes_per_launch": KERNEL_BYTES[dom] * N, "kernel_ms": {k: round(v, 4) for k, v in acc.items()},
                "bytes_per_agent_step": KERNEL_BYTES[dom]}

    # ---- e2e: C ABI with HOST buffers, copies inside the timed region --------
    e2e = None
    if rank == 0 or world > 1:
        out_names = [n for n in S.STEP_OUTPUTS if n != "reward_hyper_params"]
        hd, ha = all_d[:8].cpu().pin_memory(), all_a[:8].cpu().pin_memory()
        outs = {n: torch.empty((S.rows_of(S.EXPORT_SHAPES[n][0], W), S.EXPORT_SHAPES[n][1]),
                               dtype=torch.int32 if S.EXPORT_SHAPES[n][2] == np.int32 else torch.float32).pin_memory()
                for n in out_names}
        lib.mpenv_write_tensor_async.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        lib.mpenv_read_tensor_async.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        h2d = hd[0].numel() * 4 + ha[0].numel() * 4
        d2h = sum(t.numel() * 4 for t in outs.values())

        ids = (C.c_uint32 * len(outs))(*[S.EXPORT_IDS[n] for n in outs])
        dsts = (C.c_void_p * len(outs))(*[t.data_ptr() for t in outs.values()])
        lib.mpenv_read_tensors_pipelined.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        lib.mpenv_read_wait.argtypes = [C.c_void_p]
        lib.mpenv_stream.restype = C.c_void_p
        lib.mpenv_stream.argtypes = [C.c_void_p]
        hs = C.c_void_p(lib.mpenv_stream(h))

        def e2e_step(i):
            # host actions in, step, host observations out; the read-back of step i overlaps step i+1
            lib.mpenv_write_tensor_async(h, S.EXPORT_IDS["pvp_discrete_action"], hd[i % 8].data_ptr(), hd[0].numel() * 4)
            lib.mpenv_write_tensor_async(h, S.EXPORT_IDS["pvp_discrete_aim_action"], ha[i % 8].data_ptr(), ha[0].numel() * 4)
            lib.mpenv_step_async(h, hs)
            lib.mpenv_read_wait(h)              # step i-1's tensors are now on the host
            assert lib.mpenv_read_tensors_pipelined(h, ids, dsts, len(outs)) == 0
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
        e2e = {"value": round(world * N * k2 / float(tt.item()), 1), "unit": "agent-steps/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": k2,
               "note": "C ABI, pinned host buffers, every step: mpenv_write_tensor_async(actions) + mpenv_step_async + "
                       "mpenv_read_tensors_pipelined(18 output tensors, device snapshot then D2H on a copy stream, "
                       "overlapping the next step) ; host wall clock around k steps incl. the last read"}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_baseline = run_cpu_sample(S, cores=0)

    census = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            census = ray_census(S, local)
        except Exception as exc:           # the census is a side measurement: never fail the bench line
            census = {"unavailable": str(exc)[:200]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": "agent-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "data/simple_map, 8192 worlds/GPU, Task::Zone 6v6, random actions, autoReset, "
                                   "all train-interface tensors written every step (BASELINE.json configs[1])",
                       "worlds_per_gpu": W, "agents_per_world": AGENTS, "parallelism": f"worlds sharded x{world}",
                       "l2": "per-step working set ~600 MB (> 126 MB L2): inputs larger than L2, no explicit flush"},
            "clocks": sampler.summary() if rank == 0 else None,
            "e2e": e2e, "gpu_launches": mgr_launches(lib, h) * args.steps,
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


def ray_census(S, gpu_id, worlds=2048, warmup=60, steps=40):
    """Rays per agent-step from the INSTRUMENTED twin library (device counters where the
    reference calls traceRayAgainstWorld / sphereCast), on the same workload at a
    smaller batch.  Never timed: the rate is this count x the timed agent-steps/s."""
    sim = S.GpuSim(worlds, auto_reset=True, full_team=False, gpu_id=gpu_id, instrumented=True)
    try:
        sim.init()
        for s in range(warmup + steps):
            if s == warmup:
                sim.ray_counts(reset=True)
            d, a = S.random_actions(s, worlds)
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


def run_cpu_sample(S, cores=0, worlds=256, steps=40, warmup=5):
    """The reference's own sim code (oracle/_ref/libmpenv_ref.so, libm arithmetic) under the
    shim executor, one world per task over `cores` host threads (0 = all)."""
    orc = S.Oracle(worlds, auto_reset=True, det=False, num_threads=cores)
    orc.init()
    acts = [S.random_actions(s, worlds) for s in range(8)]
    for s in range(warmup):
        orc.set_actions(*acts[s % 8]); orc.step()
    t0 = time.perf_counter()
    for s in range(steps):
        orc.set_actions(*acts[s % 8]); orc.step()
    dt = time.perf_counter() - t0
    n = orc.num_threads()
    orc.close()
    return {"value": round(worlds * AGENTS * steps / dt, 1), "unit": "agent-steps/s", "cores": int(n),
            "kind": "reference",
            "sample": f"{worlds} worlds x {steps} steps of the same workload (reference sim.cpp/utils.cpp/"
                      f"level_gen.cpp/mesh_bvh.inl under oracle/shim, {n} threads, {dt:.1f} s)"}


def run_reference(args):
    import simlib as S
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    worlds = 256
    steps_per = 8            # each bench "step" = a bounded sample: 8 sim steps of 256 worlds
    orc = S.Oracle(worlds, auto_reset=True, det=False, num_threads=0)
    orc.init()
    acts = [S.random_actions(s, worlds) for s in range(8)]
    k = 0
    for _ in range(args.warmup * steps_per):
        orc.set_actions(*acts[k % 8]); orc.step(); k += 1
    t0 = time.perf_counter()
    for _ in range(args.steps * steps_per):
        orc.set_actions(*acts[k % 8]); orc.step(); k += 1
    dt = time.perf_counter() - t0
    n = int(orc.num_threads())
    v = round(worlds * AGENTS * args.steps * steps_per / dt, 1)
    sample = (f"each step = {steps_per} sim steps of {worlds} worlds of the same workload; reference sim code under "
              f"oracle/shim (oracle/_ref/libmpenv_ref.so), {n} host threads")
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "agent-steps/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(dt * 1e3 / args.steps, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "data/simple_map, Task::Zone 6v6, random actions, autoReset (bounded sample of "
                               "BASELINE.json configs[1])"},
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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        sys.path.insert(0, os.path.join(ROOT, "madrona-mp-env_b200"))
        run_ours(args)


if __name__ == "__main__":
    main()
