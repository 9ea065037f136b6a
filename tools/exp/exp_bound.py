"""EXPERIMENT: how much traversal work a finite initial t_max saves (host build of the
stage code, reference-order traversal).  Run: python tools/exp/exp_bound.py"""
import sys, os, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import simlib as S
S.HOSTSIM_SO = "/tmp/libexp.so"
S.build_hostsim = lambda: "/tmp/libexp.so"
W = 64
sim = S.HostSim(W, auto_reset=True, full_team=False)
L = sim.lib
L.exp_lidar_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
L.exp_trace.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 7
L.exp_num_nodes.argtypes = [C.c_void_p]
print("nodes in tree:", L.exp_num_nodes(sim.h))
sim.init()
N = W * 12 * 80


def rays():
    o = np.zeros((N, 3), np.float32); d = np.zeros((N, 3), np.float32)
    L.exp_lidar_rays(sim.h, o.ctypes.data, d.ctypes.data)
    return o, d


def trace(o, d, U):
    n = len(o)
    o = np.ascontiguousarray(o); d = np.ascontiguousarray(d); U = np.ascontiguousarray(U, np.float32)
    t = np.zeros(n, np.float32); hit = np.zeros(n, np.int32); nd = np.zeros(n, np.int32); tr = np.zeros(n, np.int32)
    L.exp_trace(sim.h, n, o.ctypes.data, d.ctypes.data, U.ctypes.data, t.ctypes.data, hit.ctypes.data,
                nd.ctypes.data, tr.ctypes.data)
    return t, hit, nd, tr


FMAX = np.float32(3.4028234663852886e38)
prev_t = None
for step in range(120):
    d_, a_ = S.random_actions(step, W)
    sim.set_actions(d_, a_); sim.step()
    if step < 60 or step % 10:
        if step >= 59:
            o, d = rays(); prev_t, prev_hit, _, _ = trace(o, d, np.full(N, FMAX))
        continue
    o, d = rays()
    t0, hit0, nd0, tr0 = trace(o, d, np.full(N, FMAX))
    print(f"step {step}: base nodes/ray {nd0.mean():.2f} tris/ray {tr0.mean():.2f} hit {hit0.mean():.3f}")
    U = np.where(hit0 > 0, t0 * np.float32(1.001), FMAX).astype(np.float32)
    t1, hit1, nd1, tr1 = trace(o, d, U)
    assert np.array_equal(t1[hit0 > 0], t0[hit0 > 0]) and np.array_equal(hit1, hit0)
    print(f"   ideal bound: nodes {nd1.mean():.2f} tris {tr1.mean():.2f}")
    for m in (1.05, 1.2, 1.5, 2.0):
        U = np.where(hit0 > 0, t0 * np.float32(m), FMAX).astype(np.float32)
        t1, hit1, nd1, tr1 = trace(o, d, U)
        print(f"   bound x{m}: nodes {nd1.mean():.2f} tris {tr1.mean():.2f}")
    idx = np.arange(N).reshape(W * 12, 80)
    fw = idx[:, :64].reshape(-1, 2, 32); rr = idx[:, 64:].reshape(-1, 2, 8)
    low = np.concatenate([fw[:, 0, :].ravel(), rr[:, 0, :].ravel()])
    high = np.concatenate([fw[:, 1, :].ravel(), rr[:, 1, :].ravel()])
    Ul = np.where(hit0[high] > 0, t0[high] * np.float32(1.001), FMAX).astype(np.float32)
    tl, hl, ndl, trl = trace(o[low], d[low], Ul)
    ok = (hl > 0) & (tl < Ul * np.float32(0.999)) | (Ul == FMAX)
    fail = ~ok
    assert np.array_equal(tl[ok & (hl > 0)], t0[low][ok & (hl > 0)])
    nodes_tw = (nd0[high].sum() + ndl.sum() + nd0[low][fail].sum()) / N
    tris_tw = (tr0[high].sum() + trl.sum() + tr0[low][fail].sum()) / N
    print(f"   twin(high->low): fail {fail.mean():.4f} nodes {nodes_tw:.2f} tris {tris_tw:.2f}")
    if prev_t is not None:
        for (mm, add) in ((1.1, 40.), (1.3, 60.), (1.5, 100.)):
            Up = np.where(prev_hit > 0, prev_t * np.float32(mm) + np.float32(add), FMAX).astype(np.float32)
            tp, hp, ndp, trp = trace(o, d, Up)
            ok = ((hp > 0) & (tp < Up * np.float32(0.999))) | (Up == FMAX)
            fail = ~ok
            print(f"   prev-step x{mm}+{add}: fail {fail.mean():.4f} nodes {(ndp.sum()+nd0[fail].sum())/N:.2f} "
                  f"tris {(trp.sum()+tr0[fail].sum())/N:.2f}")

    def chain(fan, m, add):
        A, H, Wc = fan.shape
        nodes = 0; tris = 0; fails = 0
        c0 = fan[:, :, 0].ravel(); nodes += nd0[c0].sum(); tris += tr0[c0].sum()
        for c in range(1, Wc):
            cur = fan[:, :, c].ravel(); prv = fan[:, :, c - 1].ravel()
            Uc = np.where(hit0[prv] > 0, t0[prv] * np.float32(m) + np.float32(add), FMAX).astype(np.float32)
            tc, hc, ndc, trc = trace(o[cur], d[cur], Uc)
            ok = ((hc > 0) & (tc < Uc * np.float32(0.999))) | (Uc == FMAX)
            f = ~ok
            nodes += ndc.sum() + nd0[cur][f].sum(); tris += trc.sum() + tr0[cur][f].sum(); fails += f.sum()
        return nodes, tris, fails
    for (m, add) in ((1.1, 0.), (1.25, 0.), (1.5, 0.)):
        n1, t1_, f1 = chain(fw, m, add); n2, t2_, f2 = chain(rr, m, add)
        print(f"   column chain x{m}: fail {(f1+f2)/N:.4f} nodes {(n1+n2)/N:.2f} tris {(t1_+t2_)/N:.2f}")
    prev_t, prev_hit = t0, hit0
