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
