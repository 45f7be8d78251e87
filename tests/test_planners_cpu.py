"""Host logic of the kPRM* front-end (hpp-core_b200/planners.py) on the CPU oracle: the four-batch form builds the
roadmap of the item-by-item form that follows the reference state machine (src/path-planner/k-prm-star.cc:59-222)."""
import numpy as np
import pytest

from oracle_backend import OracleBackend


@pytest.mark.parametrize("n,method,param", [(60, 1, 1e-3), (250, 2, 0.0)])
def test_kprm_star_batch_equals_sequential(pkg, ob, ur5, n, method, param):
    o = ob.Oracle(ur5)
    cand = pkg.uniform_configs(ur5, 400, seed=5)
    free = cand[o.validate_configs(cand, check_bounds=True)[0] == 1]
    seq = pkg.KPrmStar(OracleBackend(ob.Oracle(ur5), method, param), ur5, free[0], free[1], n, seed=3)
    bat = pkg.KPrmStar(OracleBackend(ob.Oracle(ur5), method, param), ur5, free[0], free[1], n, seed=3)
    assert seq.k == int(np.floor(2 * np.e * np.log(n) + 0.5))  # k-prm-star.cc:45,68-69
    ss, sb = seq.solve_sequential(), bat.solve_batch(shoot_batch=64)
    assert ss == sb and seq.shots == bat.shots
    assert np.array_equal(seq.roadmap.nodes, bat.roadmap.nodes) and len(seq.roadmap.cc) == n
    assert seq.roadmap.edges == bat.roadmap.edges and len(seq.roadmap.edges) > n
    assert np.array_equal(seq.roadmap.cc, bat.roadmap.cc)
    # every node is collision free, every edge was validated from its first end
    assert o.validate_configs(bat.roadmap.nodes, check_bounds=True)[0].all()
    e = np.array(bat.roadmap.edges)
    a, b = bat.roadmap.nodes[e[:, 0]], bat.roadmap.nodes[e[:, 1]]
    assert o.validate_edges(method, param, a, b, o.straight_lengths(a, b))["valid"].all()


def test_kprm_star_rejects_zero_nodes(pkg, ob, ur5):
    with pytest.raises(RuntimeError):
        pkg.KPrmStar(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, np.zeros(6), np.ones(6), 0)


def _free_pair(pkg, ob, desc, seed=5):
    o = ob.Oracle(desc)
    cand = pkg.uniform_configs(desc, 400, seed=seed)
    free = cand[o.validate_configs(cand, check_bounds=True)[0] == 1]
    return free[0], free[1]


@pytest.mark.parametrize("method,param", [(1, 1e-3), (2, 0.0)])
def test_diffusing_planner_batch_equals_sequential(pkg, ob, ur5, method, param):
    """DiffusingPlanner::oneStep (src/diffusing-planner.cc:180-318): the batched inner loops (all extensions in one edge
    batch, all connection attempts in another, replayed in the reference's order) build the roadmap of the path-by-path form."""
    q0, q1 = _free_pair(pkg, ob, ur5)
    seq = pkg.DiffusingPlanner(OracleBackend(ob.Oracle(ur5), method, param), ur5, q0, q1, seed=11)
    bat = pkg.DiffusingPlanner(OracleBackend(ob.Oracle(ur5), method, param), ur5, q0, q1, seed=11)
    for _ in range(25):
        seq.one_step()
        bat.one_step_batch()
        assert np.array_equal(seq.roadmap.nodes, bat.roadmap.nodes)
        assert seq.roadmap.edges == bat.roadmap.edges and np.array_equal(seq.roadmap.cc, bat.roadmap.cc)
    assert seq.roadmap.n > 20 and len(seq.roadmap.edges) >= seq.roadmap.n - len(set(seq.roadmap.cc.tolist()))
    assert bat.validations >= seq.validations  # speculative attempts are validated too, their results discarded
    o = ob.Oracle(ur5)
    e = np.array(bat.roadmap.edges)
    a, b = bat.roadmap.nodes[e[:, 0]], bat.roadmap.nodes[e[:, 1]]
    assert o.validate_edges(method, param, a, b, o.straight_lengths(a, b))["valid"].all()  # every roadmap edge is a valid path


def test_visibility_prm_batch_equals_sequential(pkg, ob, ur5):
    """VisibilityPrmPlanner::oneStep / visibleFromCC (src/visibility-prm-planner.cc:73-186): one configuration against
    every guard of every component as ONE edge batch gives the guards, connection nodes and edges of the one-by-one form."""
    q0, q1 = _free_pair(pkg, ob, ur5)
    seq = pkg.VisibilityPrmPlanner(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, q0, q1, seed=21)
    bat = pkg.VisibilityPrmPlanner(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, q0, q1, seed=21)
    for _ in range(40):
        seq.one_step()
        bat.one_step_batch()
        assert seq.shots == bat.shots
        assert np.array_equal(seq.roadmap.nodes, bat.roadmap.nodes) and seq.guard == bat.guard
        assert seq.roadmap.edges == bat.roadmap.edges and np.array_equal(seq.roadmap.cc, bat.roadmap.cc)
    guards = [n for n, g in bat.guard.items() if g]
    assert len(guards) > 2 and any(not g for g in bat.guard.values())  # guards and connection nodes both appear
    # visibility roadmap property: no two guards of different components see each other when they were inserted;
    # every connection node has >= 2 edges
    deg = {}
    for a, b in bat.roadmap.edges:
        deg[a] = deg.get(a, 0) + 1
        deg[b] = deg.get(b, 0) + 1
    assert all(deg.get(n, 0) >= 2 for n, g in bat.guard.items() if not g)


@pytest.mark.parametrize("loops", [5, 12])
def test_random_shortcut_speculative_equals_sequential(pkg, ob, ur5, loops):
    """RandomShortcut::optimize (src/path-optimization/random-shortcut.cc:73-181): the speculative batch form returns the
    path of the path-by-path loop, from the same random numbers, with fewer validation calls."""
    q0, q1 = _free_pair(pkg, ob, ur5)
    o = ob.Oracle(ur5)
    planner = pkg.BiRRTPlanner(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, q0, q1, seed=4)
    planner.solve(max_steps=400)
    assert planner.solved()
    way = planner.roadmap.nodes[planner.path()]
    assert len(way) > 3
    seq = pkg.RandomShortcut(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, loops, seed=9)
    bat = pkg.RandomShortcut(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, loops, seed=9)
    ws, wb = seq.optimize(way), bat.optimize(way, batch=True)
    assert np.array_equal(ws, wb) and seq.draws == bat.draws
    assert bat.validation_calls < seq.validation_calls
    L0, L1 = o.straight_lengths(way[:-1], way[1:]).sum(), o.straight_lengths(wb[:-1], wb[1:]).sum()
    assert L1 < L0 and np.array_equal(wb[0], way[0]) and np.array_equal(wb[-1], way[-1])
    assert o.validate_edges(1, 1e-3, wb[:-1], wb[1:], o.straight_lengths(wb[:-1], wb[1:]))["valid"].all()  # still collision free


def _planned_path(pkg, ob, ur5, seed=4):
    q0, q1 = _free_pair(pkg, ob, ur5)
    planner = pkg.BiRRTPlanner(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, q0, q1, seed=seed)
    planner.solve(max_steps=400)
    assert planner.solved()
    way = planner.roadmap.nodes[planner.path()]
    assert len(way) > 3
    return way


def test_simple_shortcut_all_pairs_batch_and_astar(pkg, ob, ur5):
    """SimpleShortcut::optimize (src/path-optimization/simple-shortcut.cc:51-83): every non-adjacent waypoint pair
    validated in one batch, then the A* of src/astar.hh.  The result is checked against the definition: it is the
    shortest path of the graph {own edges + valid shortcuts} (exhaustive dynamic programme over the DAG), its edges are
    valid, and it starts / ends where the input does."""
    way = _planned_path(pkg, ob, ur5)
    o = ob.Oracle(ur5)
    opt = pkg.SimpleShortcut(OracleBackend(o, 1, 1e-3))
    res = opt.optimize(way)
    n = len(way)
    assert opt.validated == (n - 1) * (n - 2) // 2
    assert np.array_equal(res[0], way[0]) and np.array_equal(res[-1], way[-1])
    Lr = o.straight_lengths(res[:-1], res[1:])
    assert o.validate_edges(1, 1e-3, res[:-1], res[1:], Lr)["valid"].all()
    # brute force: best[j] = min over i < j with an edge i -> j
    best, INF = [0.0] + [np.inf] * (n - 1), np.inf
    for j in range(1, n):
        for i in range(j):
            L = o.straight_lengths(way[i:i + 1], way[j:j + 1])
            ok = j == i + 1 or bool(o.validate_edges(1, 1e-3, way[i:i + 1], way[j:j + 1], L)["valid"][0])
            if ok and best[i] + float(L[0]) < best[j]:
                best[j] = best[i] + float(L[0])
    assert abs(float(Lr.sum()) - best[-1]) <= 1e-12 * best[-1]
    assert float(Lr.sum()) <= float(o.straight_lengths(way[:-1], way[1:]).sum())
    # every waypoint of the result is one of the input waypoints, in order
    idx = [int(np.nonzero((way == r).all(axis=1))[0][0]) for r in res]
    assert idx == sorted(idx) and idx[0] == 0 and idx[-1] == n - 1
    # two waypoints: nothing to do
    assert np.array_equal(opt.optimize(way[:2]), way[:2])


def test_simple_shortcut_astar_replay_tie_rule():
    """the A* replay keeps the reference's containers: with two routes of equal cost the one found first stays
    (strict '<' on the cost from start, src/astar.hh:121), and the open list is re-sorted stably"""
    from hpp_core_b200.planners import SimpleShortcut
    out = [[(1, 1.0), (2, 1.0)], [(3, 1.0)], [(3, 1.0)], []]
    assert SimpleShortcut._astar(out, [0.0, 0.0, 0.0, 0.0], 3) == [0, 1, 3]
    out = [[(1, 1.0), (2, 0.5)], [(3, 1.0)], [(3, 1.0)], []]
    assert SimpleShortcut._astar(out, [0.0, 0.0, 0.0, 0.0], 3) == [0, 2, 3]


def test_partial_shortcut_speculative_equals_sequential(pkg, ob, ur5):
    """PartialShortcut::optimize, default parameters (src/path-optimization/partial-shortcut.cc:84-98,186-223): the
    speculative batch form adopts the same joints in the same order as the joint-by-joint loop, with fewer calls;
    an adopted joint moves linearly in arc length between the end configurations, the others keep their waypoints."""
    way = _planned_path(pkg, ob, ur5, seed=6)
    o = ob.Oracle(ur5)
    seq = pkg.PartialShortcut(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5)
    bat = pkg.PartialShortcut(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5)
    ws, wb = seq.optimize(way), bat.optimize(way, batch=True)
    assert np.array_equal(ws, wb) and seq.failed_joints == bat.failed_joints
    assert seq.validation_calls == ur5.njoints and bat.validation_calls <= seq.validation_calls
    assert ws.shape == way.shape and np.array_equal(ws[0], way[0]) and np.array_equal(ws[-1], way[-1])
    L = o.straight_lengths(ws[:-1], ws[1:])
    assert o.validate_edges(1, 1e-3, ws[:-1], ws[1:], L)["valid"].all()
    for j in range(ur5.njoints):
        if j in seq.failed_joints:
            continue
        d = np.diff(ws[:, j])
        assert np.all(d >= -1e-12) or np.all(d <= 1e-12)  # monotone between the two ends
    for j in seq.failed_joints:
        assert np.array_equal(ws[:, j], way[:, j])
    assert len(seq.failed_joints) < ur5.njoints  # the planned path is not already optimal
    # a single straight path: generatePath returns null for every joint (rkAtP2 == rkAtP1)
    one = pkg.PartialShortcut(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5)
    assert np.array_equal(one.optimize(way[:2]), way[:2]) and one.failed_joints == list(range(ur5.njoints))


def test_kprm_link_helpers_match_the_array_form(pkg):
    """hppb200_kprm_link_candidates / hppb200_roadmap_merge_edges (host C++, csrc/planner_link.cpp) against the plain numpy
    statement of the same rules: `second` = (i, j) with j < i and i among j's neighbours, its mirror the slot of (j, i);
    edge merges = ConnectedComponent::merge edge by edge (the second component joins the first)."""
    rng = np.random.default_rng(3)
    n, k = 300, 9
    pts = rng.normal(size=(n, 3))
    pts[17] = pts[40]  # exact distance ties
    d = ((pts[:, None, :] - pts[None, :, :]) ** 2).sum(-1)
    order = np.lexsort((np.broadcast_to(np.arange(n), (n, n)), d), axis=1)  # by (distance, index); the node itself first
    nbrs = order[:, :k].astype(np.int64)
    dist = np.take_along_axis(d, nbrs, axis=1)
    nbrs[5, -2:] = -1  # padding
    for symmetric in (True, False):
        kind, mirror = pkg.kprm_link_candidates(nbrs, dist if symmetric else None, symmetric)
        for i in range(n):
            for s in range(k):
                j = nbrs[i, s]
                if j < 0 or j == i:
                    assert kind[i, s] == 0 and mirror[i, s] == -1
                    continue
                pos = np.nonzero(nbrs[j] == i)[0]
                if j < i and len(pos):
                    assert kind[i, s] == 2 and mirror[i, s] == j * k + pos[0], (i, s)
                else:
                    assert kind[i, s] == 1 and mirror[i, s] == -1, (i, s)
    lab = np.arange(n, dtype=np.int64) + 1000
    lab[7] = lab[3]; lab[250] = lab[3]
    pairs = rng.integers(0, n, (400, 2))
    cc = lab.copy()
    for a, b in pairs:
        ca, cb = cc[a], cc[b]
        if ca != cb:
            cc[cc == cb] = ca
    assert np.array_equal(pkg.roadmap_merge_edges(lab, pairs), cc)
    assert np.array_equal(pkg.roadmap_merge_edges(lab, np.zeros((0, 2), np.int64)), lab)
