"""the scripted A* bot (policy id -1): planAStarAISystem
(sim.cpp:5041-5172) + applyBotActionsSystem (sim.cpp:2057-2091) over the navmesh
and the A* next-triangle table (mgr.cpp:946-1211).

The system itself is the reference's own code in the oracle.  The navmesh
adjacency and the table are built by csrc/navmesh.cpp (madrona::Navmesh and
meshoptimizer are not vendored; mgr.cpp cannot be compiled) and handed to the
oracle like the BVH, so this file also pins the table with independent checks.
"""
import ctypes as C
import heapq

import numpy as np

import simlib as S
from test_hostsim_parity import assert_same


def _nav(built):
    return S.load_navmesh(C.CDLL(S.build_hostsim()))


def test_navmesh_and_astar_table_properties(built):
    verts, tris, adj, lookup = _nav(built)
    nt = len(tris)
    assert nt == 155 and len(verts) == 109            # simple_map: 212 file vertices, 109 distinct
    # adjacency is symmetric and each pair shares exactly two vertices
    for t in range(nt):
        for e in range(3):
            n = adj[t, e]
            if n < 0:
                continue
            # simple_map's navmesh holds 8 coincident duplicate triangles: a duplicate links to the
            # real neighbour, which links back to the first copy only (directed half-edges)
            assert t in adj[n] or any((sorted(tris[t]) == sorted(tris[k])) for k in adj[n] if k >= 0), (t, n)
            assert len(set(tris[t]) & set(tris[n])) >= 2
            assert {tris[t, e], tris[t, (e + 1) % 3]} <= set(tris[n]), "neighbour e lies across edge (v_e, v_e+1)"
    look = lookup.astype(np.int64)
    look[look == 0xFFFFFFFF] = -1
    assert (np.diag(look) == np.arange(nt)).all(), "start == goal returns the goal"
    centers = verts[tris].astype(np.float64).mean(axis=1)

    def dijkstra(src):
        dist = np.full(nt, np.inf); dist[src] = 0.0
        pq = [(0.0, src)]
        while pq:
            d, u = heapq.heappop(pq)
            if d > dist[u]:
                continue
            for n in adj[u]:
                if n >= 0:
                    nd = d + np.linalg.norm(centers[u] - centers[n])
                    if nd < dist[n]:
                        dist[n] = nd; heapq.heappush(pq, (nd, n))
        return dist
    dist = np.stack([dijkstra(s) for s in range(nt)])
    reach = np.isfinite(dist)
    assert reach[look >= 0].all(), "an entry only for reachable pairs"
    # the reference's open list (a std::set ordered by live scores) drops candidates whose score
    # ties an element already present, so a few reachable pairs end up without a path
    assert (look >= 0)[reach].mean() > 0.97, (look >= 0)[reach].mean()
    # the first hop is a neighbour of the start, and following the table always arrives
    for s in range(nt):
        for g in range(nt):
            if s == g or look[s, g] < 0:
                continue
            assert look[s, g] in adj[s], (s, g, look[s, g])
    for s, g in [(0, nt - 1), (17, 120), (88, 3), (150, 42)]:
        cur, hops = s, 0
        while cur != g:
            cur = look[cur, g]; hops += 1
            assert cur >= 0 and hops <= nt
    # A* with a consistent heuristic: the hop lies on a path no longer than 1.0001 x the optimum
    # for the overwhelming majority (the reference's std::set open list drops score ties)
    s_idx, g_idx = np.nonzero((look >= 0) & ~np.eye(nt, dtype=bool))
    hop = look[s_idx, g_idx]
    via = np.linalg.norm(centers[s_idx] - centers[hop], axis=1) + dist[hop, g_idx]
    assert (via <= dist[s_idx, g_idx] * 1.0001 + 1e-3).mean() > 0.95


def _bot_rollout(W, steps, bot_mask, seed=10):
    host = S.HostSim(W, rand_seed=seed, auto_reset=True)
    orc = S.Oracle(W, rand_seed=seed, auto_reset=True, det=True)
    try:
        host.init(); orc.init()
        pol = np.where(bot_mask, -1, 0).astype(np.int32).reshape(-1, 1)
        for sim in (host, orc):
            sim.set_export("agent_policy", pol)
        moved = 0.0
        start = orc.export("self_pos").copy()
        for s in range(steps):
            d, a = S.seek_policy(orc, s, W)
            for sim in (host, orc):
                sim.set_actions(d, a)
                sim.step()
            assert_same(host, orc, W, f"step {s}")
            # the bot's applied action rows are part of the tensor API
            for name in ("pvp_discrete_action", "pvp_aim_action"):
                x, y = host.export(name), orc.export(name)
                assert np.array_equal(x.view(np.int32), y.view(np.int32)), f"step {s}: {name}"
        moved = np.abs(orc.export("self_pos") - start).reshape(W * 12, -1)[bot_mask.reshape(-1)].max()
        return moved, orc.export("match_result")
    finally:
        host.close(); orc.close()


def test_all_bots_match_oracle(built):
    W = 4
    moved, mr = _bot_rollout(W, 400, np.ones((W, 12), bool))
    assert moved > 0.05, "bots never moved"


def test_bots_vs_policy_team(built):
    """Team B scripted, team A driven by the action tensors (the PBT baseline setup, jax_train.py:282-283)."""
    W = 4
    mask = np.zeros((W, 12), bool)
    mask[:, 6:] = True
    _bot_rollout(W, 250, mask)
