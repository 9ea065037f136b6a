"""EXPERIMENT: union of visited nodes / tested triangles over warp-sized ray groups
(fixed-order packet traversal).  Run: python tools/exp/exp_packet.py"""
import sys, os, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import simlib as S
S.build_hostsim = lambda: "/tmp/libexp.so"
W = 64
sim = S.HostSim(W, auto_reset=True, full_team=False)
L = sim.lib
L.exp_lidar_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
L.exp_trace.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 9
sim.init()
N = W * 12 * 80
FMAX = np.float32(3.4028234663852886e38)
def popc(x):
    x = x.astype(np.uint64); c = np.zeros(x.shape, np.int64)
    for i in range(64): c += ((x >> np.uint64(i)) & np.uint64(1)).astype(np.int64)
    return c
for step in range(101):
    d_, a_ = S.random_actions(step, W)
    sim.set_actions(d_, a_); sim.step()
    if step not in (60, 100): continue
    o = np.zeros((N, 3), np.float32); d = np.zeros((N, 3), np.float32)
    L.exp_lidar_rays(sim.h, o.ctypes.data, d.ctypes.data)
    U = np.full(N, FMAX)
    t = np.zeros(N, np.float32); hit = np.zeros(N, np.int32); nd = np.zeros(N, np.int32); tr = np.zeros(N, np.int32)
    nm = np.zeros(N, np.uint64); tm = np.zeros((N, 4), np.uint64)
    L.exp_trace(sim.h, N, o.ctypes.data, d.ctypes.data, U.ctypes.data, t.ctypes.data, hit.ctypes.data, nd.ctypes.data, tr.ctypes.data, nm.ctypes.data, tm.ctypes.data)
    print(f"step {step}: per-ray nodes {nd.mean():.2f} tris {tr.mean():.2f}")
    idx = np.arange(N).reshape(W * 12, 80)
    fw = idx[:, :64].reshape(-1, 2, 32); rr = idx[:, 64:].reshape(-1, 2, 8)
    def report(name, groups):
        # groups: [G, k] ray indices
        un = np.bitwise_or.reduce(nm[groups], axis=1)
        ut = np.bitwise_or.reduce(tm[groups], axis=1)   # [G,4]
        nn = popc(un); tt = popc(ut).sum(axis=1)
        k = groups.shape[1]
        print(f"   {name}: group {k} rays: union nodes {nn.mean():.2f} (per ray {nn.mean()/k:.3f}; util {nd[groups].sum(axis=1).mean()/ (nn.mean()*32):.2f}) union tris {tt.mean():.2f} (per ray {tt.mean()/k:.3f}, util {tr[groups].sum(axis=1).mean()/(tt.mean()*32):.2f})")
        return nn.mean(), tt.mean()
    A = fw.shape[0]
    report("fwd 1 height x 32 cols", fw.reshape(-1, 32))
    g = fw.reshape(A, 2, 2, 16).transpose(0, 2, 1, 3).reshape(-1, 32)
    report("fwd 2 heights x 16 cols", g)
    g8 = fw.reshape(A, 2, 4, 8).transpose(0, 2, 1, 3).reshape(-1, 16)
    report("fwd 2 heights x 8 cols (16 lanes)", g8)
    report("rear 2 heights x 8 cols (16 lanes)", rr.reshape(-1, 16))
    # two agents' rear in one warp
    report("rear of 2 agents (32 lanes)", rr.reshape(-1, 32))
    g4 = fw.reshape(A, 2, 8, 4).transpose(0, 2, 1, 3).reshape(-1, 8)
    report("fwd 2 heights x 4 cols (8 lanes)", g4)
