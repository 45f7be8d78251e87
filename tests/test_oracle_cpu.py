"""CPU tests (-m "not gpu"): the oracle against the reference's own fixtures/invariants, and
oracle self-checks."""
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- Intervals: pinned against the reference --------------------------------------------
def test_intervals_known_answers(ob):
    """tests/test-intervals.cc:55-233 known answers."""
    cases = json.load(open(os.path.join(ROOT, "tests", "golden", "intervals_known_answers.json")))
    assert len(cases) == 5
    for c in cases:
        u = np.array(c["unions"])
        got = ob.intervals_union(u[:, 0], u[:, 1])
        assert np.array_equal(got, np.array(c["expected"])), c["name"]


def test_intervals_touching_merge(ob):
    """tests/test-intervals.cc interval_6: [0,1] u [2,3] u [1,2] -> one interval."""
    got = ob.intervals_union(np.array([0., 2., 1.]), np.array([1., 3., 2.]))
    assert np.array_equal(got, np.array([[0., 3.]]))


def test_intervals_vs_reference_build(ob):
    """oracle::Intervals == the reference's own class (oracle/_ref) on random sequences,
    including exact ties (grid-valued end points) and zero-length intervals."""
    if ob.ref_intervals() is None:
        pytest.skip("oracle/_ref not built (reference absent)")
    rng = np.random.default_rng(3)
    for trial in range(400):
        n = int(rng.integers(1, 40))
        if trial % 2:
            lo = rng.integers(0, 30, n).astype(float)
            hi = lo + rng.integers(0, 5, n)
        else:
            t = rng.uniform(0, 1, n)
            l = rng.uniform(0, .05, n)
            lo, hi = t - l, t + l
        a = ob.intervals_union(lo, hi, "oracle")
        b = ob.intervals_union(lo, hi, "ref")
        assert np.array_equal(a, b), (trial, lo, hi, a, b)
        assert np.all(a[1:, 0] > a[:-1, 1])  # sorted, disjoint (test-intervals.cc checkIntervals)


# ---- numerics self-checks ----------------------------------------------------------------
def test_portable_sincos(ob):
    x = np.concatenate([np.random.default_rng(0).uniform(-20, 20, 200000), np.linspace(-np.pi, np.pi, 10001),
                        [0.0, np.pi / 2, -np.pi / 2, np.pi, 1e-300, 1e4]])
    s, c = ob.sincos(x)
    assert np.abs(s - np.sin(x)).max() < 2.3e-16
    assert np.abs(c - np.cos(x)).max() < 2.3e-16


def test_fk_vs_dh_end_effector(ob, ur5, ur5_oracle):
    """FK against an independent numpy chain product (the URDF chain of Appendix A.1)."""
    q = np.random.default_rng(1).uniform(-np.pi, np.pi, (50, 6))
    oMg = ur5_oracle.geometry_placements(q)

    def rot(axis, a):
        c, s = np.cos(a), np.sin(a)
        return {0: np.array([[1, 0, 0], [0, c, -s], [0, s, c]]), 1: np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]]),
                2: np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])}[axis]
    for i in range(q.shape[0]):
        R, t = np.eye(3), np.zeros(3)
        for j in range(6):
            P = ur5.joint_placement[j]
            Rp, tp = P[:9].reshape(3, 3), P[9:]
            t = R @ tp + t
            R = R @ Rp @ rot(int(ur5.joint_type[j]), q[i, j])
        g = ur5.ngeoms - 1  # tool box on wrist_3
        Pg = ur5.geom_placement[g]
        assert np.allclose(oMg[i, g, :9].reshape(3, 3), R @ Pg[:9].reshape(3, 3), atol=1e-13)
        assert np.allclose(oMg[i, g, 9:], R @ Pg[9:] + t, atol=1e-13)


def _rand_T(rng, scale=1.0):
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    x, y, z, w = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    return np.concatenate([R.reshape(-1), rng.uniform(-scale, scale, 3)])


def _box_verts(h):
    return np.array([[sx * h[0], sy * h[1], sz * h[2]] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)])


def test_gjk_vs_closed_forms(ob):
    """GJK (via convex point sets) agrees with the closed forms on sphere/capsule cases."""
    rng = np.random.default_rng(5)
    for _ in range(300):
        Ta, Tb = _rand_T(rng), _rand_T(rng)
        ra, rb, ha, hb = rng.uniform(.05, .3, 4)
        # capsule-capsule closed form vs GJK on 2-vertex convex cores
        d_closed = ob.shape_distance(1, [ra, ha], Ta, 1, [rb, hb], Tb)
        va = np.array([[0, 0, -ha], [0, 0, ha]])
        vb = np.array([[0, 0, -hb], [0, 0, hb]])
        d_gjk = ob.shape_distance(3, [0], Ta, 3, [0], Tb, va, vb) - ra - rb
        if d_closed > -ra - rb + 1e-9:  # cores separated
            assert abs(d_closed - d_gjk) <= 1e-9 * max(1.0, abs(d_closed)), (d_closed, d_gjk)
        # sphere-box closed form vs GJK(point, box vertices)
        hbx = rng.uniform(.05, .4, 3)
        d_sb = ob.shape_distance(0, [ra], Ta, 2, list(hbx), Tb)
        d_sb_gjk = ob.shape_distance(3, [0], Ta, 3, [0], Tb, np.zeros((1, 3)), _box_verts(hbx)) - ra
        if d_sb > -ra + 1e-9:
            assert abs(d_sb - d_sb_gjk) <= 1e-9 * max(1.0, abs(d_sb)), (d_sb, d_sb_gjk)
        # box primitive support == box as convex hull
        hax = rng.uniform(.05, .4, 3)
        d_bb = ob.shape_distance(2, list(hax), Ta, 2, list(hbx), Tb)
        d_bb_c = ob.shape_distance(3, [0], Ta, 3, [0], Tb, _box_verts(hax), _box_verts(hbx))
        assert abs(d_bb - d_bb_c) <= 1e-9 * max(1.0, d_bb), (d_bb, d_bb_c)


def test_gjk_box_box_vs_sampling(ob):
    """Box-box / capsule-box distance vs brute-force optimisation (scipy) on separated cases."""
    from scipy.optimize import minimize
    rng = np.random.default_rng(11)
    for _ in range(40):
        Ta, Tb = _rand_T(rng, 0.8), _rand_T(rng, 0.8)
        ha, hb = rng.uniform(.05, .3, 3), rng.uniform(.05, .3, 3)
        d = ob.shape_distance(2, list(ha), Ta, 2, list(hb), Tb)
        Ra, ta, Rb, tb = Ta[:9].reshape(3, 3), Ta[9:], Tb[:9].reshape(3, 3), Tb[9:]

        def f(x):
            pa = Ra @ (np.clip(x[:3], -1, 1) * ha) + ta
            pb = Rb @ (np.clip(x[3:], -1, 1) * hb) + tb
            return np.sum((pa - pb) ** 2)
        best = min(minimize(f, rng.uniform(-1, 1, 6), bounds=[(-1, 1)] * 6, method="L-BFGS-B",
                            options=dict(ftol=1e-16, gtol=1e-12)).fun for _ in range(6))
        ref = np.sqrt(max(best, 0))
        if ref > 1e-3:
            assert abs(d - ref) < 2e-6 * max(1, ref), (d, ref)
        else:
            assert d < 2e-3


# ---- model tables ---------------------------------------------------------------------------
def test_pair_order(ur5, ur5_oracle):
    """cPairs_ order: self pairs first, then obstacle-major / joint-major / inner object
    (src/obstacle-user.cc:53-76,222-235); universe geometry gets no obstacle pair."""
    a, b, m = ur5_oracle.config_pairs()
    ns = ur5.self_pairs.shape[0]
    assert np.array_equal(np.stack([a[:ns], b[:ns]], 1), ur5.self_pairs)
    moving = [g for g in range(ur5.ngeoms) if ur5.geom_joint[g] != 0]
    exp = [(g, -(o + 1)) for o in range(ur5.nobstacles) for g in sorted(moving, key=lambda g: (ur5.geom_joint[g], g))]
    assert list(zip(a[ns:], b[ns:])) == exp
    assert len(a) == ns + ur5.nobstacles * len(moving)


def test_body_pairs_and_coefficients(ur5, ur5_oracle):
    bp = ur5_oracle.body_pairs()
    # 15 self body pairs (non-adjacent link pairs) + 8 obstacles x 6 joints with a body
    assert len(bp["joint_a"]) == 15 + 8 * 6
    # Initialize::doExecute swaps so that joint_a is the moving joint of the second geometry
    assert bp["joint_a"][0] == 2 and bp["joint_b"][0] == 0
    # obstacle pair of joint 6: chain 6->...->0, coefficient_i = cumulative length (revolute)
    last = len(bp["joint_a"]) - 1
    c0, c1 = bp["coef_begin"][last], bp["coef_begin"][last + 1]
    assert list(bp["coef_joint"][c0:c1]) == [6, 5, 4, 3, 2, 1]
    radii = ur5_oracle.body_radii()
    dists = [np.linalg.norm(ur5.joint_placement[j - 1][9:]) for j in (6, 5, 4, 3, 2, 1)]
    exp = radii[6] + np.concatenate([[0], np.cumsum(dists[:-1])])
    assert np.allclose(bp["coef_value"][c0:c1], exp, rtol=1e-15)
    # obstacle pairs have no reverse coefficients
    assert bp["rcoef_begin"][last] == bp["rcoef_begin"][last + 1]


# ---- the reference's invariants (tests/test-continuous-validation.cc) ---------------------
def _edges_from(configs):
    return configs[0::2], configs[1::2]


@pytest.mark.parametrize("source", ["golden", "random"])
def test_reference_invariants(pkg, ur5, ur5_oracle, golden_configs, source):
    """I1: PathValidation::validate(q) == CollisionValidation::validate(q) (:200-202,242-244)
       I2: Discretized(0.05) invalid => Progressive(1e-3) and Dichotomy(0) invalid (:203-205,251-253)
       I3: forward, Discretized valid => both continuous valid is NOT guaranteed in general; the
           reference asserts the converse direction only where it holds (:217-219): continuous
           valid => discretized valid."""
    o = ur5_oracle
    cfg = golden_configs[:200] if source == "golden" else pkg.uniform_configs(ur5, 600, seed=99)
    q1, q2 = _edges_from(cfg)
    L = o.straight_lengths(q1, q2)
    v, _, _ = o.validate_configs(q1)
    zero = np.zeros(len(q1))
    for meth, par in ((0, 0.05), (1, 1e-3), (2, 0.0)):
        r = o.validate_edges(meth, par, q1, q1, zero)
        assert np.array_equal(r["valid"], v), "I1 method %d" % meth
    for rev in (False, True):
        d = o.validate_edges(0, 0.05, q1, q2, L, reverse=rev)
        p = o.validate_edges(1, 1e-3, q1, q2, L, reverse=rev)
        c = o.validate_edges(2, 0.0, q1, q2, L, reverse=rev)
        assert not np.any((d["valid"] == 0) & (p["valid"] == 1)), "I2 progressive"
        assert not np.any((d["valid"] == 0) & (c["valid"] == 1)), "I2 dichotomy"
        assert np.all(d["valid"][c["valid"] == 1] == 1)
        assert np.all(d["valid"][p["valid"] == 1] == 1)
        # valid part ends before (forward) / after (reverse) the reported collision parameter
        bad = p["valid"] == 0
        if rev:
            assert np.all(p["t_last_valid"][bad] >= p["t_fail"][bad])
        else:
            assert np.all(p["t_last_valid"][bad] <= p["t_fail"][bad])


def test_progressive_single_interval_property(ur5, ur5_oracle, pkg):
    """Persistent pair order (the reference's pooled vector) vs per-edge reset: flags may differ
    only inside the tolerance band; measured here on random edges."""
    cfg = pkg.uniform_configs(ur5, 400, seed=5)
    q1, q2 = _edges_from(cfg)
    L = ur5_oracle.straight_lengths(q1, q2)
    a = ur5_oracle.validate_edges(1, 1e-3, q1, q2, L, order_mode=0)
    b = ur5_oracle.validate_edges(1, 1e-3, q1, q2, L, order_mode=1)
    assert np.mean(a["valid"] != b["valid"]) < 0.02


def test_edge_cases(ur5, ur5_oracle):
    o = ur5_oracle
    q = np.zeros((1, 6))
    # zero-length edge: exactly one sample for Progressive (progressive.cc:100), two for Discretized
    assert o.validate_edges(1, 1e-3, q, q, np.zeros(1))["samples"][0] == 1
    assert o.validate_edges(0, 0.05, q, q, np.zeros(1))["samples"][0] == 2
    # constructor checks
    with pytest.raises(RuntimeError):
        o.validate_edges(1, 0.0, q, q, np.zeros(1))
    with pytest.raises(RuntimeError):
        o.validate_edges(2, 1e-3, q, q, np.zeros(1))
    # joint bounds: first violated rank
    qq = np.array([[0, 4.0, 0, -5.0, 0, 0]])
    v, idx, reason = o.validate_configs(qq, check_bounds=True)
    assert v[0] == 0 and idx[0] == 1 and reason[0] == 2
    # empty batch
    v, _, _ = o.validate_configs(np.zeros((0, 6)))
    assert v.shape == (0,)


def test_nn_basic(ur5, ob, pkg):
    o = ob.Oracle(ur5)
    nodes = pkg.uniform_configs(ur5, 500, seed=1)
    nodes[100] = nodes[50]  # exact duplicate: the earlier index must win (strict '<')
    cc = (np.arange(500) % 3).astype(np.int32)
    o.knn_append(nodes, cc)
    w = np.array(ur5.weights)
    qs = pkg.uniform_configs(ur5, 20, seed=2)
    qs[0] = nodes[50]
    idx, dist = o.knn_query(qs, 1)
    assert idx[0, 0] == 50 and dist[0, 0] == 0
    d = np.sqrt((((nodes[None] - qs[:, None]) * w) ** 2).sum(-1))
    assert np.array_equal(idx[:, 0], d.argmin(1))
    idx5, dist5 = o.knn_query(qs, 5, cc=1)
    dm = np.where(cc[None] == 1, d, np.inf)
    assert np.array_equal(idx5[1:], np.argsort(dm, 1, kind="stable")[1:, :5])
    assert np.allclose(dist5, np.sort(dm, 1)[:, :5], rtol=1e-14)
    ball = o.knn_within_ball(qs[3], 1.0)
    assert np.array_equal(ball, np.nonzero(d[3] < 1.0)[0])


def test_humanoid_tree_oracle_invariants(pkg, ob):
    """The 30-joint TREE model with convex meshes (BASELINE configs[3]-like, fixed pelvis): FK of a branch does not
    depend on the other branches' joints, GJK distances of convex meshes are lower-bounded by their bounding-sphere
    distance, and the reference invariants I1/I2 hold for the three path validations."""
    desc = pkg.ModelDescription.load("humanoid_obst6.json")
    o = ob.Oracle(desc)
    names = desc.geom_names
    q = pkg.uniform_configs(desc, 200, seed=7)
    P = o.geometry_placements(q).reshape(len(q), len(names), 12)
    q2 = q.copy()
    q2[:, 18:] = pkg.uniform_configs(desc, 200, seed=8)[:, 18:]  # joints 19..30 = the two legs
    P2 = o.geometry_placements(q2).reshape(len(q), len(names), 12)
    upper = [i for i, n in enumerate(names) if not any(t in n for t in ("thigh", "shin", "foot", "toe"))]
    lower = [i for i, n in enumerate(names) if any(t in n for t in ("thigh", "shin", "foot", "toe"))]
    assert np.array_equal(P[:, upper], P2[:, upper]) and not np.array_equal(P[:, lower], P2[:, lower])
    R = P[:, :, :9].reshape(len(q), len(names), 3, 3)
    assert np.allclose(np.einsum("nkij,nklj->nkil", R, R), np.eye(3), atol=1e-12)
    # I1 / I2 on local edges
    q1 = pkg.uniform_configs(desc, 600, seed=9)
    d = np.random.default_rng(9).normal(size=q1.shape)
    qb = np.clip(q1 + 0.4 * d / np.linalg.norm(d, axis=1, keepdims=True), desc.q_lower, desc.q_upper)
    L = o.straight_lengths(q1, qb)
    v = o.validate_configs(q1)[0]
    for meth, par in ((0, 0.05), (1, 1e-3), (2, 0.0)):
        assert np.array_equal(o.validate_edges(meth, par, q1, q1, np.zeros(len(q1)))["valid"], v)
    dd = o.validate_edges(0, 0.05, q1, qb, L)["valid"]
    pp = o.validate_edges(1, 1e-3, q1, qb, L)["valid"]
    cc = o.validate_edges(2, 0.0, q1, qb, L)["valid"]
    assert not np.any((dd == 0) & (cc == 1)) and np.mean((dd == 0) & (pp == 1)) < 0.01
    assert 0.05 < cc.mean() < 0.95


def test_portable_atan_accuracy(ob):
    """The arctangent behind the SO(3) logarithm stays within 1 ulp of libm over its whole domain."""
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(0, 4, 100000), 10.0 ** rng.uniform(-12, 25, 100000),
                        [0.0, np.inf, 0.4375, 0.6875, 1.1875, 2.4375, 2.0 ** -27, 2.0 ** 66]])
    y, r = ob.atan(x), np.arctan(x)
    assert np.all(np.abs(y - r) <= np.spacing(np.maximum(r, 1e-300)))


def test_freeflyer_lie_group_ops(pkg, ob):
    """Free-flyer = R^3 x SO(3): interpolation returns the stored end points exactly,
    stays on the unit sphere, is a geodesic of WeighedDistance (d(q1,q) + d(q,q2) = L), the difference has the
    rotation angle of the relative quaternion (shorter arc), and -q2 describes the same path end."""
    desc = pkg.ModelDescription.load("humanoid_ff_obst6.json")
    o = ob.Oracle(desc)
    assert (o.nq, o.nv) == (37, 36) and pkg.freeflyer_ranks(desc) == [0]
    q1, q2 = pkg.uniform_configs(desc, 2000, seed=3), pkg.uniform_configs(desc, 2000, seed=4)
    L = o.straight_lengths(q1, q2)
    assert np.array_equal(o.interpolate(q1, q2, L, 0 * L), q1) and np.array_equal(o.interpolate(q1, q2, L, L), q2)
    v = o.difference(q1, q2)
    assert np.array_equal(v[:, :3], q2[:, :3] - q1[:, :3]) and np.array_equal(v[:, 6:], q2[:, 7:] - q1[:, 7:])
    dots = np.abs(np.sum(q1[:, 3:7] * q2[:, 3:7], axis=1))
    assert np.allclose(np.linalg.norm(v[:, 3:6], axis=1), 2 * np.arccos(np.clip(dots, 0, 1)), atol=1e-12)
    for u in (1e-9, 0.25, 0.5, 0.9, 1 - 1e-9):
        q = o.interpolate(q1, q2, L, u * L)
        assert np.abs(np.linalg.norm(q[:, 3:7], axis=1) - 1).max() < 1e-15
        assert np.abs(o.straight_lengths(q1, q) + o.straight_lengths(q, q2) - L).max() < 1e-13
        assert np.allclose(o.straight_lengths(q1, q), u * L, rtol=1e-9, atol=1e-13)
    flipped = q2.copy()
    flipped[:, 3:7] *= -1
    assert np.allclose(o.straight_lengths(q1, flipped), L, rtol=1e-14)
    assert np.allclose(o.geometry_placements(q2[:50]), o.geometry_placements(flipped[:50]), atol=1e-15)
    # zero rotation: the difference vanishes (to rounding), interpolation keeps the quaternion
    same = q2.copy()
    same[:, 3:7] = q1[:, 3:7]
    assert np.abs(o.difference(q1, same)[:, 3:6]).max() < 1e-15
    assert np.abs(o.interpolate(q1, same, o.straight_lengths(q1, same), 0.3 * o.straight_lengths(q1, same))[:, 3:7]
                  - q1[:, 3:7]).max() < 1e-15


def test_freeflyer_humanoid_oracle_invariants(pkg, ob):
    """Reference invariants I1 / I2 (tests/test-continuous-validation.cc:200-205,251-253) on the free-flyer
    humanoid, and FK of the floating base: every geometry placement moves rigidly with the root joint."""
    desc = pkg.ModelDescription.load("humanoid_ff_obst6.json")
    o = ob.Oracle(desc)
    q = pkg.uniform_configs(desc, 300, seed=5)
    ng = len(desc.geom_names)
    nrobot = ng  # obstacles are not in geometry_placements
    P = o.geometry_placements(q).reshape(len(q), nrobot, 12)
    base = q.copy()
    base[:, :3] = 0
    base[:, 3:7] = [0, 0, 0, 1]
    P0 = o.geometry_placements(base).reshape(len(q), nrobot, 12)
    x, y, z, w = q[:, 3], q[:, 4], q[:, 5], q[:, 6]
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                  2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                  2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=1).reshape(-1, 3, 3)
    t_exp = np.einsum("nij,nkj->nki", R, P0[:, :, 9:]) + q[:, None, :3]
    assert np.allclose(P[:, :, 9:], t_exp, atol=1e-12)
    q1 = pkg.uniform_configs(desc, 500, seed=6)
    d = np.random.default_rng(6).normal(size=q1.shape)
    qb = pkg.normalize_quaternions(desc, np.clip(q1 + 0.4 * d / np.linalg.norm(d, axis=1, keepdims=True),
                                                 desc.q_lower, desc.q_upper))
    L = o.straight_lengths(q1, qb)
    v = o.validate_configs(q1)[0]
    for meth, par in ((0, 0.05), (1, 1e-3), (2, 0.0)):
        assert np.array_equal(o.validate_edges(meth, par, q1, q1, np.zeros(len(q1)))["valid"], v)
    dd = o.validate_edges(0, 0.05, q1, qb, L)["valid"]
    pp = o.validate_edges(1, 1e-3, q1, qb, L)["valid"]
    cc = o.validate_edges(2, 0.0, q1, qb, L)["valid"]
    assert not np.any((dd == 0) & (cc == 1)) and np.mean((dd == 0) & (pp == 1)) < 0.01
    assert 0.05 < cc.mean() < 0.95


def _random_splines(pkg, desc, n, seed):
    """steeringMethod::Spline<BernsteinBasis, 3>::steer(q1, {1}, v1, q2, {1}, v2) with length = distance(q1, q2), as in
    tests/test-continuous-validation.cc:288-345 (random end configurations and end velocities)"""
    from oracle import binding as ob
    cfg = pkg.uniform_configs(desc, 2 * n, seed=seed)
    q1, q2 = np.ascontiguousarray(cfg[0::2]), np.ascontiguousarray(cfg[1::2])
    rng = np.random.default_rng(seed)
    v1, v2 = rng.uniform(-1, 1, q1.shape), rng.uniform(-1, 1, q1.shape)
    L = ob.Oracle(desc).straight_lengths(q1, q2)
    base, P = pkg.spline_steer(q1, v1, q2, v2, L)
    return q1, v1, q2, v2, L, base, P


def test_spline_path_eval_and_velocity_bound(pkg, ob, ur5):
    """path::Spline<BernsteinBasis, 3> (src/path/spline.cc): end points and end velocities of the steered spline, the
    Bernstein basis against its textbook form, and impl_velocityBound(t0, t1) really bounding |dq/dt| on [t0, t1]."""
    o = ob.Oracle(ur5)
    n = 200
    q1, v1, q2, v2, L, base, P = _random_splines(pkg, ur5, n, 11)
    assert np.array_equal(o.spline_eval(base, P, L, np.zeros(n)), q1)
    assert np.allclose(o.spline_eval(base, P, L, L), q2, rtol=0, atol=1e-14)
    h = 1e-6
    assert np.allclose((o.spline_eval(base, P, L, np.full(n, h)) - q1) / h, v1, atol=2e-5)
    assert np.allclose((q2 - o.spline_eval(base, P, L, L - h)) / h, v2, atol=2e-5)
    rng = np.random.default_rng(3)
    t = rng.uniform(0, 1, n) * L
    u = (t / L)[:, None]
    B = np.stack([(1 - u) ** 3, 3 * u * (1 - u) ** 2, 3 * u ** 2 * (1 - u), u ** 3], axis=1)  # [n, 4, 1]
    assert np.allclose(o.spline_eval(base, P, L, t), base + (P * B).sum(axis=1), rtol=0, atol=1e-13)
    # the bound on sub-intervals (what the refinement of BodyPairCollision::collisionFreeInterval calls)
    a, b = np.sort(rng.uniform(0, 1, (2, n)), axis=0) * L
    vb = o.spline_velocity_bound(base, P, L, a, b)
    for s in np.linspace(0, 1, 33):
        ts = a + s * (b - a)
        num = (o.spline_eval(base, P, L, np.minimum(ts + h, L)) - o.spline_eval(base, P, L, np.maximum(ts - h, 0))) / (np.minimum(ts + h, L) - np.maximum(ts - h, 0))[:, None]
        assert np.all(np.abs(num) <= vb * (1 + 1e-6) + 1e-6)
    whole = o.spline_velocity_bound(base, P, L, np.zeros(n), L)
    assert np.all(vb <= whole * (1 + 1e-12))  # a sub-interval never needs a larger bound than the whole path


def test_spline_validation_consistency_like_the_reference_test(pkg, ob, ur5):
    """tests/test-continuous-validation.cc:288-427 on the oracle: where Discretized (0.05) finds a collision on a random
    spline, Progressive (0.01) must find one too, forward and reverse; a spline that IS a straight line (end velocities
    = (q2 - q1) / L) validates like the straight path."""
    o = ob.Oracle(ur5)
    n = 600
    q1, v1, q2, v2, L, base, P = _random_splines(pkg, ur5, n, 12)
    for rev in (False, True):
        d = o.validate_spline_edges(0, 0.05, base, P, L, reverse=rev)
        p = o.validate_spline_edges(1, 0.01, base, P, L, reverse=rev)
        start_free = o.validate_configs(q2 if rev else q1)[0] == 1
        # (the tolerance lets Progressive accept penetrations of that order, e.g. an end configuration that barely
        #  touches: a handful of such edges is what the tolerance means, many would be a wrong velocity bound)
        assert np.sum(start_free & (d["valid"] == 0) & (p["valid"] == 1)) <= n // 100
        assert 0.05 < p["valid"].mean() < 0.95
        # Dichotomy has no tolerance: a path it accepts has no colliding sample at all, and it is at least as strict
        # as Progressive
        z = o.validate_spline_edges(2, 0.0, base, P, L, reverse=rev)
        assert not np.any((z["valid"] == 1) & (d["valid"] == 0))
        assert not np.any((z["valid"] == 1) & (p["valid"] == 0))
        fine = o.validate_spline_edges(0, 0.002, base[:150], P[:150], L[:150], reverse=rev)
        assert not np.any((z["valid"][:150] == 1) & (fine["valid"] == 0))
    vs = (q2 - q1) / L[:, None]
    sb, sp = pkg.spline_steer(q1, vs, q2, vs, L)
    s = o.validate_spline_edges(1, 1e-3, sb, sp, L)
    e = o.validate_edges(1, 1e-3, q1, q2, L)
    assert (s["valid"] == e["valid"]).mean() > 0.99
    both = (s["valid"] == 0) & (e["valid"] == 0)
    assert np.allclose(s["t_fail"][both], e["t_fail"][both], atol=0.05)
    # zero-length spline: one sample, the configuration decides
    z = o.validate_spline_edges(1, 1e-3, q1[:50], np.zeros((50, 4, ur5.nq)), np.zeros(50))
    assert np.array_equal(z["valid"], o.validate_configs(q1[:50])[0])
    assert np.all(z["samples"] <= 2)
