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
