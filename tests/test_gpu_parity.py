"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle on
the same inputs.  Bar: bit-exact flags, pair indices, edge parameters and distances (both sides
follow the same numerics contract); any deviation must sit within 1e-9 of contact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng(pkg, ur5):
    m = pkg.Model(ur5)
    return m, pkg.Workspace(m, 0)


def test_fk_bit_exact(pkg, eng, ur5, ur5_oracle, golden_configs):
    _, ws = eng
    q = np.concatenate([golden_configs, pkg.uniform_configs(ur5, 5000, seed=3), np.zeros((1, 6))])
    assert np.array_equal(ws.geometry_placements(q), ur5_oracle.geometry_placements(q))


def test_pair_distances_bit_exact(pkg, eng, ur5, ur5_oracle, golden_configs):
    _, ws = eng
    q = np.concatenate([golden_configs, pkg.uniform_configs(ur5, 3000, seed=4)])
    d, o = ws.pair_distances(q), ur5_oracle.pair_distances(q)
    assert np.array_equal(d, o)


def _check_flags(valid, first, ov, op, dist):
    """flags/pair identical; a mismatch is tolerated only within 1e-9 of contact (north_star)."""
    bad = np.nonzero((valid != ov) | (first != op))[0]
    for i in bad:
        assert np.abs(dist[i]).min() < 1e-9, i
    return len(bad)


def test_config_validation_golden_and_random(pkg, eng, ur5, ur5_oracle, golden_configs):
    _, ws = eng
    q = np.concatenate([golden_configs, pkg.uniform_configs(ur5, 200000, seed=12345)])
    valid, first, reason = ws.validate_configs(q)
    ov, op, orr = ur5_oracle.validate_configs(q)
    assert np.array_equal(valid, ov)
    assert np.array_equal(first, op)
    assert np.array_equal(reason, orr)
    assert 0.2 < valid.mean() < 0.9


def test_config_validation_joint_bounds_and_ragged(pkg, eng, ur5, ur5_oracle):
    _, ws = eng
    q = pkg.uniform_configs(ur5, 1000 + 37, seed=8) * 1.2  # some out of bounds; ragged last block
    for cb in (False, True):
        got = ws.validate_configs(q, check_bounds=cb)
        exp = ur5_oracle.validate_configs(q, check_bounds=cb)
        for a, b in zip(got, exp):
            assert np.array_equal(a, b)
    assert (ws.validate_configs(q, check_bounds=True)[2] == 2).any()
    v, f, r = ws.validate_configs(np.zeros((0, 6)))
    assert v.shape == (0,)
    v, f, r = ws.validate_configs(np.zeros((1, 6)))
    assert v[0] == 1 and f[0] == -1


@pytest.mark.parametrize("name", ["ur5_boxes8.json", "humanoid_obst16.json", "humanoid_ff_obst6.json"])
def test_config_pass_shadow_cores_exact_and_zero_copy(pkg, ob, monkeypatch, name):
    """configPassKernel keeps 32-bit shadows of the bounding cores: every form must give the oracle's
    bits: shadow (default), FP64 cores (0), shadow kernel with every pair redone from the FP64 placements (2), and
    the zero-copy host-pointer form reading a pinned buffer over PCIe.  Hostile rows (NaN, infinities, huge angles)
    take the exact path."""
    import torch
    desc = pkg.ModelDescription.load(name)
    model = pkg.Model(desc)
    o = ob.Oracle(desc)
    n = 60000 + 37
    q = pkg.uniform_configs(desc, n, seed=77)
    q[5, 0] = np.nan; q[6, :] = np.nan; q[7, 1] = np.inf; q[8, 2] = 9.9e4; q[9, 0] = -np.inf; q[10, -1] = 1e300
    exp = o.validate_configs(q)
    for mode in ("1", "0", "2"):
        monkeypatch.setenv("HPPB200_CFG_SHADOW", mode)
        ws = pkg.Workspace(model, 0)
        for a, b in zip(ws.validate_configs(q), exp):
            assert np.array_equal(a, b), mode
    monkeypatch.setenv("HPPB200_CFG_SHADOW", "1")
    monkeypatch.setenv("HPPB200_CFG_ZEROCOPY", "1")
    ws = pkg.Workspace(model, 0)
    qp = torch.from_numpy(q).pin_memory()
    v = torch.empty(n, dtype=torch.uint8).pin_memory(); f = torch.empty(n, dtype=torch.int32).pin_memory()
    r = torch.empty(n, dtype=torch.uint8).pin_memory()
    for m in (n, 1000 + 37, 1):  # sliced, single slice with a ragged last block, one row
        v.zero_(); f.zero_(); r.zero_()
        ws.validate_configs_ptr(qp.data_ptr(), m, v.data_ptr(), f.data_ptr(), r.data_ptr(), 0)
        for a, b in zip((v.numpy()[:m], f.numpy()[:m], r.numpy()[:m]), exp):
            assert np.array_equal(a, b[:m]), m


@pytest.mark.parametrize("name", ["ur5_boxes0.json", "ur5_boxes1.json", "ur5_boxes32.json"])
def test_config_validation_other_scenes(pkg, ob, name):
    desc = pkg.ModelDescription.load(name)
    ws = pkg.Workspace(pkg.Model(desc), 0)
    o = ob.Oracle(desc)
    q = pkg.uniform_configs(desc, 20000, seed=21)
    for a, b in zip(ws.validate_configs(q), o.validate_configs(q)):
        assert np.array_equal(a, b)


def test_security_margins(pkg, ob):
    desc = pkg.ModelDescription.load("ur5_boxes8.json")
    desc.security_margins = np.full((7, 7), 0.02)
    ws = pkg.Workspace(pkg.Model(desc), 0)
    o = ob.Oracle(desc)
    q = pkg.uniform_configs(desc, 20000, seed=22)
    got, exp = ws.validate_configs(q), o.validate_configs(q)
    for a, b in zip(got, exp):
        assert np.array_equal(a, b)
    base = ob.Oracle(pkg.ModelDescription.load("ur5_boxes8.json")).validate_configs(q)[0]
    assert got[0].sum() < base.sum()  # margins only remove valid configurations


def test_straight_lengths(pkg, eng, ur5, ur5_oracle):
    _, ws = eng
    q = pkg.uniform_configs(ur5, 20000, seed=6)
    assert np.array_equal(ws.straight_lengths(q[0::2], q[1::2]), ur5_oracle.straight_lengths(q[0::2], q[1::2]))


def _edges(pkg, ur5, o, n, seed, short=False):
    q = pkg.uniform_configs(ur5, 2 * n, seed=seed)
    q1, q2 = q[0::2].copy(), q[1::2].copy()
    if short:
        d = q2 - q1
        q2 = np.clip(q1 + 0.3 * d / np.linalg.norm(d, axis=1, keepdims=True), -np.pi, np.pi)
    return q1, q2, o.straight_lengths(q1, q2)


@pytest.mark.parametrize("method,param", [(0, 0.05), (1, 1e-3), (2, 0.0)])
@pytest.mark.parametrize("reverse", [False, True])
def test_edge_validation_parity(pkg, eng, ur5, ur5_oracle, golden_configs, method, param, reverse):
    """C1-style edges (q1 = conf[2i], q2 = conf[2i+1], tests/test-continuous-validation.cc:165-176):
    the 100 golden edges + random long and short edges."""
    _, ws = eng
    o = ur5_oracle
    g1, g2 = golden_configs[0:200:2], golden_configs[1:200:2]
    a1, a2, _ = _edges(pkg, ur5, o, 4000, 31)
    b1, b2, _ = _edges(pkg, ur5, o, 4000, 32, short=True)
    q1, q2 = np.concatenate([g1, a1, b1]), np.concatenate([g2, a2, b2])
    L = o.straight_lengths(q1, q2)
    r = ws.validate_edges(method, param, q1, q2, L, reverse=reverse)
    e = o.validate_edges(method, param, q1, q2, L, reverse=reverse)
    st = ws.stats()
    assert st["iter_cap_hits"] == 0
    assert np.array_equal(r["valid"], e["valid"])
    assert np.array_equal(r["reason"], e["reason"])
    assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
    assert np.array_equal(r["t_last_valid"], e["t_last_valid"])
    assert np.array_equal(r["pair"], e["pair"])
    assert 0.02 < r["valid"].mean() < 0.98


def test_edge_zero_length_and_endpoints(pkg, eng, ur5, ur5_oracle):
    """PathValidation::validate(q) == zero-length StraightPath (src/path-validation.cc:36-49): I1."""
    _, ws = eng
    q = pkg.uniform_configs(ur5, 3000, seed=41)
    v = ws.validate_configs(q)[0]
    z = np.zeros(len(q))
    for method, param in ((0, 0.05), (1, 1e-3), (2, 0.0)):
        r = ws.validate_edges(method, param, q, q, z)
        assert np.array_equal(r["valid"], v)
        assert np.all(r["t_fail"][v == 0] == 0.0)


def test_edge_invariants_on_gpu(pkg, eng, ur5, ur5_oracle):
    """I2: Discretized(0.05) invalid => Dichotomy(0) invalid, strictly.  Progressive(tol > 0)
    tolerates penetrations up to its tolerance, so a discretized sample may catch a graze it
    accepts: rare, and then the GPU must still agree with the oracle."""
    _, ws = eng
    q1, q2, L = _edges(pkg, ur5, ur5_oracle, 20000, 51)
    for rev in (False, True):
        d = ws.validate_edges(0, 0.05, q1, q2, L, reverse=rev)["valid"]
        p = ws.validate_edges(1, 1e-3, q1, q2, L, reverse=rev)["valid"]
        c = ws.validate_edges(2, 0.0, q1, q2, L, reverse=rev)["valid"]
        assert not np.any((d == 0) & (c == 1))
        odd = np.nonzero((d == 0) & (p == 1))[0]
        assert len(odd) < 0.002 * len(L)
        if len(odd):
            e = ur5_oracle.validate_edges(1, 1e-3, q1[odd], q2[odd], L[odd], reverse=rev)
            assert np.all(e["valid"] == 1)


def test_edge_iteration_cap_is_flagged(pkg, eng, ur5, ur5_oracle):
    m, _ = eng
    ws = pkg.Workspace(m, 0)
    ws.set_max_iterations(3)
    q1, q2, L = _edges(pkg, ur5, ur5_oracle, 500, 61)
    r = ws.validate_edges(1, 1e-3, q1, q2, L)
    e = ur5_oracle.validate_edges(1, 1e-3, q1, q2, L, max_iter=3)
    assert np.array_equal(r["reason"], e["reason"])
    assert (r["reason"] == 3).any() and np.all(r["valid"][r["reason"] == 3] == 0)


def test_error_codes(pkg, eng):
    _, ws = eng
    q = np.zeros((1, 6))
    with pytest.raises(pkg.EngineError):   # std::runtime_error in the reference constructors
        ws.validate_edges(1, 0.0, q, q, np.zeros(1))
    with pytest.raises(pkg.EngineError):
        ws.validate_edges(2, 1e-3, q, q, np.zeros(1))
    with pytest.raises(ValueError):
        ws.validate_edges(7, 1e-3, q, q, np.zeros(1))


def test_validator_mirror_classes(pkg, eng, ur5, ur5_oracle):
    """The reference-shaped interface: validate(q) / validate(path, reverse)."""
    _, ws = eng
    cv = pkg.CollisionValidation.create(ws)
    prog = pkg.Progressive.create(ws, 1e-3)
    q = pkg.uniform_configs(ur5, 40, seed=71)
    ov = ur5_oracle.validate_configs(q)
    for i in range(20):
        ok, rep = cv.validate(q[i])
        assert ok == bool(ov[0][i])
        if not ok:
            assert rep.reason == 1 and rep.index == ov[1][i]
        ok5, _ = prog.validate_config(q[i])
        assert ok5 == ok
    L = ur5_oracle.straight_lengths(q[:20], q[20:])
    e = ur5_oracle.validate_edges(1, 1e-3, q[:20], q[20:], L)
    for i in range(20):
        path = pkg.StraightPath(q[i], q[20 + i], L[i])
        ok, part, rep = prog.validate(path, False)
        assert ok == bool(e["valid"][i])
        if ok:
            assert part is path
        else:
            assert rep.parameter == e["t_fail"][i] and part.length == e["t_last_valid"][i]


def test_knn_parity(pkg, eng, ur5, ob):
    m, _ = eng
    o = ob.Oracle(ur5)
    nn = pkg.NearestNeighbor(m, 0)
    nodes = pkg.uniform_configs(ur5, 30000, seed=81)
    nodes[20000] = nodes[123]  # exact duplicate
    cc = (np.arange(len(nodes)) % 4).astype(np.int32)
    for lo in range(0, len(nodes), 7000):  # incremental appends, as a growing roadmap
        nn.append(nodes[lo:lo + 7000], cc[lo:lo + 7000])
        o.knn_append(nodes[lo:lo + 7000], cc[lo:lo + 7000])
    assert len(nn) == len(nodes)
    qs = pkg.uniform_configs(ur5, 64, seed=82)
    qs[0] = nodes[123]
    # k <= 32: single-pass warp top-k; k = 33: selection passes.  The k-nearest SET is fixed by the reference's
    # strict '>' (earlier node kept); the order of exact ties in its output is std::priority_queue's, so ties
    # are canonicalised by node index (the engine's documented key is (distance, node index)).
    for k, c in ((1, -1), (1, 2), (2, 1), (10, -1), (7, 3), (32, -1), (33, -1)):
        gi, gd = nn.query(qs, k, c)
        oi, od = o.knn_query(qs, k, c)
        srt = np.lexsort((oi, od), axis=1)
        oi, od = np.take_along_axis(oi, srt, axis=1), np.take_along_axis(od, srt, axis=1)
        assert np.array_equal(gi, oi), (k, c)
        assert np.array_equal(gd, od)
    assert nn.query(qs[:1], 1)[0][0, 0] == 123
    nn.relabel(3, 0)
    o.knn_relabel(3, 0)
    assert np.array_equal(nn.query(qs, 5, 0)[0], o.knn_query(qs, 5, 0)[0])
    balls, count = nn.within_ball(qs[:8], 0.9)
    for i in range(8):
        assert np.array_equal(balls[i], o.knn_within_ball(qs[i], 0.9))
    # k larger than the component: padded with -1
    small = pkg.NearestNeighbor(m, 0)
    small.append(nodes[:3])
    gi, gd = small.query(qs[:2], 5)
    assert np.all(gi[:, 3:] == -1) and np.all(np.isinf(gd[:, 3:]))


def _exact_nn(pkg, m, monkeypatch):
    """A mirror with the tensor-core candidate pass switched off (exact FP64 scan only)."""
    monkeypatch.setenv("HPPB200_KNN_TC", "0")
    nn = pkg.NearestNeighbor(m, 0)
    monkeypatch.delenv("HPPB200_KNN_TC")
    return nn


def test_knn_tensor_core_tile_error_bound(pkg, eng, ur5):
    """The TMEM accumulator of the tcgen05 TF32 GEMM IS the approximate squared WeighedDistance: its error against
    FP64 must stay far below the bound eps = 2^-16 (|a|^2 + max |b|^2) the thresholds are inflated by."""
    m, _ = eng
    nn = pkg.NearestNeighbor(m, 0)
    nodes = pkg.uniform_configs(ur5, 20000, seed=91)
    qs = pkg.uniform_configs(ur5, 100, seed=92)
    nn.append(nodes)
    w = np.asarray(ur5.weights)
    ref = (((qs[:, None, :] - nodes[None, :256, :]) * w) ** 2).sum(-1)
    tile = nn.tc_debug_tile(qs)[:100]
    na, nb = ((qs * w) ** 2).sum(-1), ((nodes * w) ** 2).sum(-1).max()
    eps = (na + nb) * 2.0 ** -15
    assert np.all(np.abs(tile - ref) < 0.125 * eps[:, None])
    assert np.all(nn.tc_debug_tile(qs)[100:] > 1e29)  # padding rows never pass a threshold


def _prismatic_chain(pkg, nj, lo, hi):
    """a chain of nj prismatic joints with configuration values in [lo, hi]: norms far larger than distances"""
    eye = [1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0.1]
    return pkg.ModelDescription(dict(name="prismatic%d" % nj, joint_parent=list(range(nj)), joint_type=[3 + (k % 3) for k in range(nj)],
                                     joint_placement=[eye] * nj, q_lower=[lo] * nj, q_upper=[hi] * nj, geom_joint=[nj],
                                     geom_placement=[eye], geom_shape=[0], geom_params=[[0.05, 0, 0, 0]], self_pairs=[],
                                     weights=[1.0 + 0.1 * k for k in range(nj)]))


@pytest.mark.parametrize("nj,lo,hi", [(14, 90.0, 110.0), (10, -3.0, 3.0), (14, -1.0, 1.0)])
def test_knn_tensor_core_error_bound_wide_models_far_from_origin(pkg, nj, lo, hi):
    """the bound the thresholds carry must hold for the widest models the tensor-core pass takes (kc = 6: 48 K
    slots at nq = 14) and for nodes far from the origin (norms >> distances), and a miss would drop a true neighbour
    silently: raw accumulators against FP64 (|error| < eps / 2), then full queries against the exact scan, exact
    duplicates (ties) included."""
    desc = _prismatic_chain(pkg, nj, lo, hi)
    m = pkg.Model(desc)
    nn = pkg.NearestNeighbor(m, 0)
    nodes = pkg.uniform_configs(desc, 40000, seed=191)
    nodes[20000:20040] = nodes[100:140]  # exact duplicates: ties decided by node index
    qs = pkg.uniform_configs(desc, 128, seed=192)
    qs[:40] = nodes[100:140] + 1e-7
    nn.append(nodes)
    w = np.asarray(desc.weights)  # prismatic joints: translation part is unweighted in WeighedDistance
    ref = ((qs[:, None, :] - nodes[None, :256, :]) ** 2).sum(-1)
    tile = nn.tc_debug_tile(qs)
    na, nb = (qs ** 2).sum(-1), (nodes ** 2).sum(-1).max()
    eps = (na + nb) * 2.0 ** (-14 if 3 * nj + 6 > 32 else -15)
    assert np.all(np.abs(tile[:128] - ref) < 0.5 * eps[:, None]), float((np.abs(tile[:128] - ref) / eps[:, None]).max())
    import os
    for k in (1, 10):
        gi, gd = nn.query(qs, k)
        assert nn.tc_stats()["used"]
        os.environ["HPPB200_KNN_TC"] = "0"
        try:
            ex = pkg.NearestNeighbor(m, 0)
        finally:
            del os.environ["HPPB200_KNN_TC"]
        ex.append(nodes)
        ei, ed = ex.query(qs, k)
        assert not ex.tc_stats()["used"]
        assert np.array_equal(gi, ei) and np.array_equal(gd, ed)
    assert np.array_equal(gi[:40, 0], np.arange(100, 140))  # the earlier of two identical nodes wins


def test_knn_tensor_core_pass_matches_exact_scan_and_oracle(pkg, eng, ur5, ob, monkeypatch):
    """k-NN indices and distances of the tensor-core candidate pass + FP64 re-rank: identical to the exact scan
    (1 <= k <= 32, ragged query counts, roadmap growing between queries, duplicates) and to the oracle."""
    m, _ = eng
    nn, ex = pkg.NearestNeighbor(m, 0), _exact_nn(pkg, m, monkeypatch)
    o = ob.Oracle(ur5)
    nodes = pkg.uniform_configs(ur5, 300000, seed=93)
    nodes[250000] = nodes[77]
    qs = pkg.uniform_configs(ur5, 333, seed=94)
    qs[5] = nodes[77]
    qs[6] = nodes[299999]
    done = 0
    for upto in (20000, 20001, 70000, 300000):  # appends between queries: the operand tiles follow the mirror
        nn.append(nodes[done:upto]); ex.append(nodes[done:upto]); o.knn_append(nodes[done:upto], None)
        done = upto
        for k, nqry in ((1, 333), (10, 333), (32, 40), (7, 129), (33, 64), (75, 100)):
            gi, gd = nn.query(qs[:nqry], k)
            st = nn.tc_stats()
            assert st["used"] and st["redone"] == 0, st
            ei, ed = ex.query(qs[:nqry], k)
            assert not ex.tc_stats()["used"]
            assert np.array_equal(gi, ei), (upto, k)
            assert np.array_equal(gd, ed), (upto, k)
            assert st["candidates"] < 400 * k * nqry  # the filter does filter
        gi, gd = nn.query(qs[:64], 10)
        oi, od = o.knn_query(qs[:64], 10, -1)
        srt = np.lexsort((oi, od), axis=1)
        assert np.array_equal(gi, np.take_along_axis(oi, srt, axis=1))
        assert np.array_equal(gd, np.take_along_axis(od, srt, axis=1))
    assert nn.query(qs[:32], 1)[0][5, 0] == 77  # the earlier of two identical nodes
    # below the query threshold the exact scan runs
    nn.query(qs[:8], 3)
    assert not nn.tc_stats()["used"]
    # a connected-component filter is applied by the exact re-rank (thresholds = k-th distance INSIDE the component);
    # a component with fewer than k nodes in every prefix ends in the exact scan: same results either way
    lab = (np.arange(len(nodes)) % 3).astype(np.int32)
    lab[1000:1010] = 7
    nn2, ex2 = pkg.NearestNeighbor(m, 0), _exact_nn(pkg, m, monkeypatch)
    nn2.append(nodes, lab); ex2.append(nodes, lab)
    for k, c in ((1, 0), (5, 2), (12, 7), (40, 1)):
        gi, gd = nn2.query(qs[:100], k, cc=c)
        assert nn2.tc_stats()["used"]
        ei, ed = ex2.query(qs[:100], k, cc=c)
        assert np.array_equal(gi, ei) and np.array_equal(gd, ed), (k, c)


def test_knn_tensor_core_pass_clustered_roadmap(pkg, eng, ur5, monkeypatch):
    """Roadmaps are not uniform: a dense cluster appended AFTER the bootstrap prefix makes the prefix thresholds
    loose, candidate lists overflow, and those queries are redone by the exact scan: same results either way."""
    m, _ = eng
    nn, ex = pkg.NearestNeighbor(m, 0), _exact_nn(pkg, m, monkeypatch)
    rng = np.random.default_rng(5)
    base = pkg.uniform_configs(ur5, 50000, seed=95)
    centre = base[11]
    cluster = centre + 1e-3 * rng.standard_normal((120000, 6))
    nodes = np.concatenate([base, cluster, cluster[:100]])  # + exact duplicates inside the cluster
    nn.append(nodes); ex.append(nodes)
    qs = np.concatenate([pkg.uniform_configs(ur5, 100, seed=96), centre + 1e-3 * rng.standard_normal((60, 6))])
    for k in (1, 5, 32):
        gi, gd = nn.query(qs, k)
        st = nn.tc_stats()
        ei, ed = ex.query(qs, k)
        assert st["used"]
        assert np.array_equal(gi, ei) and np.array_equal(gd, ed), k
    assert st["redone"] > 0  # the cluster queries overflowed their lists (3750 chunks of 32 nodes over 148 CTAs)


def test_kprm_star_batch_same_roadmap_as_oracle(pkg, ob):
    """kPRM* (src/path-planner/k-prm-star.cc) from four batches on the engine -- configurations, ONE k-NN call over
    all nodes (k = 53 > 32: tensor-core pass + sorted exact re-rank), two edge batches -- next to the same planner on
    the oracle: same nodes, same edges in the same order, same components."""
    desc = pkg.ModelDescription.load("ur5_boxes8.json")
    o = ob.Oracle(desc)
    cand = pkg.uniform_configs(desc, 400, seed=5)
    free = cand[o.validate_configs(cand, check_bounds=True)[0] == 1]
    n = 17000
    eb = pkg.EngineBackend(pkg, pkg.Model(desc), 0, 1, 1e-3)
    gpu = pkg.KPrmStar(eb, desc, free[0], free[1], n, seed=21)
    cpu = pkg.KPrmStar(_OracleBackend(ob.Oracle(desc), 1, 1e-3), desc, free[0], free[1], n, seed=21)
    assert gpu.k == 53
    sg, sc = gpu.solve_batch(), cpu.solve_batch()
    assert eb.nn.tc_stats()["used"] and eb.nn.tc_stats()["redone"] == 0
    assert sg == sc and gpu.shots == cpu.shots and gpu.validated_edges == cpu.validated_edges
    assert np.array_equal(gpu.roadmap.nodes, cpu.roadmap.nodes)
    assert gpu.roadmap.edges == cpu.roadmap.edges and np.array_equal(gpu.roadmap.cc, cpu.roadmap.cc)
    assert len(gpu.roadmap.edges) > n
    # the device mirror received the final component labels (hppb200_knn_set_labels): component-restricted searches agree
    qs = pkg.uniform_configs(desc, 16, seed=77)
    for lab in sorted(set(gpu.roadmap.cc.tolist()))[:3]:
        assert np.array_equal(eb.nearest(qs, lab)[0], cpu.b.nearest(qs, lab)[0])


def test_knn_tensor_core_full_size(pkg, eng, ur5, monkeypatch):
    """BASELINE config[2] size (1 M-node roadmap): 1024 queries, k = 10, against the exact scan."""
    m, _ = eng
    nn, ex = pkg.NearestNeighbor(m, 0), _exact_nn(pkg, m, monkeypatch)
    nodes = pkg.uniform_configs(ur5, 1 << 20, seed=97)
    nn.append(nodes); ex.append(nodes)
    qs = pkg.uniform_configs(ur5, 1024, seed=98)
    for k in (1, 10):
        gi, gd = nn.query(qs, k)
        st = nn.tc_stats()
        ei, ed = ex.query(qs, k)
        assert st["used"] and st["redone"] == 0 and st["stages"] == 3
        assert np.array_equal(gi, ei) and np.array_equal(gd, ed)


def test_full_size_properties(pkg, eng, ur5, ur5_oracle):
    """BASELINE config[1] size (1M configurations): size-independent properties —
    idempotence (same batch twice -> same flags), permutation equivariance, consistency of a
    random 20k subsample with the oracle, and flag/pair coherence."""
    _, ws = eng
    n = 1 << 20
    q = pkg.uniform_configs(ur5, n, seed=12345)
    v1, f1, r1 = ws.validate_configs(q)
    v2, f2, _ = ws.validate_configs(q)
    assert np.array_equal(v1, v2) and np.array_equal(f1, f2)
    perm = np.random.default_rng(0).permutation(n)
    v3, f3, _ = ws.validate_configs(q[perm])
    assert np.array_equal(v3, v1[perm]) and np.array_equal(f3, f1[perm])
    assert np.all((f1 >= 0) == (v1 == 0)) and np.all((r1 == 1) == (v1 == 0))
    sub = np.random.default_rng(1).choice(n, 20000, replace=False)
    ov, op, _ = ur5_oracle.validate_configs(q[sub])
    assert np.array_equal(v1[sub], ov) and np.array_equal(f1[sub], op)


@pytest.mark.parametrize("reverse", [False, True])
@pytest.mark.parametrize("slots", [0, 1024])
def test_wavefront_engine_parity(pkg, eng, ur5, ur5_oracle, golden_configs, monkeypatch, reverse, slots):
    """Progressive through the round-synchronous wavefront engine (kernels_wavefront.cu), forced for a small
    batch; slots=1024 < n exercises the refill of finished slots.  Same bar as the other engines: bit-exact."""
    m, _ = eng
    monkeypatch.setenv("HPPB200_WF_MIN", "1")
    monkeypatch.setenv("HPPB200_WF_SLOTS", str(slots))
    ws = pkg.Workspace(m, 0)
    o = ur5_oracle
    g1, g2 = golden_configs[0:200:2], golden_configs[1:200:2]
    a1, a2, _ = _edges(pkg, ur5, o, 3000, 33)
    b1, b2, _ = _edges(pkg, ur5, o, 2000, 34, short=True)
    z = pkg.uniform_configs(ur5, 300, seed=35)  # zero-length paths
    q1, q2 = np.concatenate([g1, a1, b1, z]), np.concatenate([g2, a2, b2, z])
    L = o.straight_lengths(q1, q2)
    s0 = ws.stats()
    r = ws.validate_edges(1, 1e-3, q1, q2, L, reverse=reverse)
    s1 = ws.stats()
    e = o.validate_edges(1, 1e-3, q1, q2, L, reverse=reverse)
    assert s1["engine_ticks"] - s0["engine_ticks"] < 2000  # rounds, not CTA ticks: the wavefront engine ran
    assert s1["iter_cap_hits"] == 0
    for k in ("valid", "reason", "pair", "t_last_valid"):
        assert np.array_equal(r[k], e[k]), k
    assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)


def test_wavefront_engine_iteration_cap(pkg, eng, ur5, ur5_oracle, monkeypatch):
    m, _ = eng
    monkeypatch.setenv("HPPB200_WF_MIN", "1")
    ws = pkg.Workspace(m, 0)
    ws.set_max_iterations(3)
    q1, q2, L = _edges(pkg, ur5, ur5_oracle, 500, 62)
    r = ws.validate_edges(1, 1e-3, q1, q2, L)
    e = ur5_oracle.validate_edges(1, 1e-3, q1, q2, L, max_iter=3)
    assert np.array_equal(r["reason"], e["reason"]) and np.array_equal(r["valid"], e["valid"])
    assert (r["reason"] == 3).any()


@pytest.mark.parametrize("name", ["ur5_boxes0.json", "ur5_boxes1.json", "ur5_boxes32.json"])
def test_wavefront_engine_other_scenes(pkg, ob, monkeypatch, name):
    """15, 21 and 207 body pairs: the last one takes the multi-word position masks (more than 64 pairs)."""
    monkeypatch.setenv("HPPB200_WF_MIN", "1")
    desc = pkg.ModelDescription.load(name)
    ws = pkg.Workspace(pkg.Model(desc), 0)
    o = ob.Oracle(desc)
    q1, q2, L = _edges(pkg, desc, o, 3000, 36)
    for rev in (False, True):
        s0 = ws.stats()
        r = ws.validate_edges(1, 1e-3, q1, q2, L, reverse=rev)
        assert ws.stats()["engine_ticks"] - s0["engine_ticks"] < 4000
        e = o.validate_edges(1, 1e-3, q1, q2, L, reverse=rev)
        for k in ("valid", "reason", "pair", "t_last_valid"):
            assert np.array_equal(r[k], e[k]), (name, rev, k)
        assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)


def _local_edges(pkg, desc, n, seed, step):
    """Edges from a random configuration to a nearby one (|dq|_2 = step), inside the joint limits."""
    q1 = pkg.uniform_configs(desc, n, seed=seed)
    d = np.random.default_rng(seed).normal(size=q1.shape)
    q2 = np.clip(q1 + step * d / np.linalg.norm(d, axis=1, keepdims=True), desc.q_lower, desc.q_upper)
    return q1, pkg.normalize_quaternions(desc, q2)


@pytest.mark.parametrize("name", ["humanoid_obst6.json", "humanoid_obst16.json"])
def test_humanoid_tree_convex_parity(pkg, ob, monkeypatch, name):
    """BASELINE configs[3]-like model (fixed pelvis): 30 revolute joints in a TREE, convex meshes through GJK,
    133 / 273 body pairs.  FK, pair distances, configuration validation and the three path validations
    (Progressive through the wavefront engine when the model fits it, else CTA engine / walker) against the
    oracle, bit-exact."""
    monkeypatch.setenv("HPPB200_WF_MIN", "1")
    desc = pkg.ModelDescription.load(name)
    ws = pkg.Workspace(pkg.Model(desc), 0)
    o = ob.Oracle(desc)
    q = pkg.uniform_configs(desc, 20000, seed=81)
    assert np.array_equal(ws.geometry_placements(q[:2000]), o.geometry_placements(q[:2000]))
    assert np.array_equal(ws.pair_distances(q[:2000]), o.pair_distances(q[:2000]))
    got, exp = ws.validate_configs(q), o.validate_configs(q)
    for a, b in zip(got, exp):
        assert np.array_equal(a, b)
    assert 0.02 < got[0].mean() < 0.9
    a1, a2 = _local_edges(pkg, desc, 3000, 82, 0.4)
    b1, b2, _ = _edges(pkg, desc, o, 1000, 83)
    q1, q2 = np.concatenate([a1, b1]), np.concatenate([a2, b2])
    L = o.straight_lengths(q1, q2)
    assert np.array_equal(ws.straight_lengths(q1, q2), L)
    for method, param in ((0, 0.05), (1, 1e-3), (2, 0.0)):
        for rev in (False, True):
            r = ws.validate_edges(method, param, q1, q2, L, reverse=rev)
            e = o.validate_edges(method, param, q1, q2, L, reverse=rev)
            assert ws.stats()["iter_cap_hits"] == 0
            for k in ("valid", "reason", "pair", "t_last_valid"):
                assert np.array_equal(r[k], e[k]), (name, method, rev, k)
            assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
            assert 0.02 < r["valid"].mean() < 0.98


@pytest.mark.parametrize("name", ["ur5_boxes8.json", "humanoid_obst16.json", "humanoid_ff_obst6.json"])
def test_first_sample_screen_same_results(pkg, ob, monkeypatch, name):
    """First-sample screen of the edge path (kernels_screen.cu): q(t0) of every edge through the configuration kernels on
    the continuous pair table, the edges that start in collision finished there, the rest gathered for the edge engines.
    Same results as the unscreened engines and as the oracle: random edges (most start in collision on the humanoid),
    edges that start at VALID configurations, zero-length edges, forward and reverse, a batch that is screened away
    entirely and one nothing is removed from."""
    desc = pkg.ModelDescription.load(name)
    o = ob.Oracle(desc)
    monkeypatch.setenv("HPPB200_FS_MIN", "1")
    monkeypatch.setenv("HPPB200_FS_CHUNK", "1000")  # (read at every call) the screen in several chunks, the last one partial
    ws_on = pkg.Workspace(pkg.Model(desc), 0)
    monkeypatch.setenv("HPPB200_FIRST_SAMPLE_SCREEN", "0")
    ws_off = pkg.Workspace(pkg.Model(desc), 0)
    a1, a2 = _local_edges(pkg, desc, 2500, 301, 0.4)
    b1, b2, _ = _edges(pkg, desc, o, 1500, 302)
    z = pkg.uniform_configs(desc, 64, seed=303)
    q1, q2 = np.concatenate([a1, b1, z]), np.concatenate([a2, b2, z])
    L = o.straight_lengths(q1, q2)
    ok1 = o.validate_configs(q1)[0].astype(bool)
    ok2 = o.validate_configs(q2)[0].astype(bool)
    assert 0.02 < ok1.mean() < 0.98
    subsets = {"all": np.arange(len(q1)), "valid start": np.flatnonzero(ok1 & ok2), "colliding start": np.flatnonzero(~ok1 & ~ok2)}
    for method, param in ((1, 1e-3), (2, 0.0)):
        for rev in (False, True):
            for label, idx in subsets.items():
                if len(idx) == 0:
                    continue
                r = ws_on.validate_edges(method, param, q1[idx], q2[idx], L[idx], reverse=rev)
                u = ws_off.validate_edges(method, param, q1[idx], q2[idx], L[idx], reverse=rev)
                e = o.validate_edges(method, param, q1[idx], q2[idx], L[idx], reverse=rev)
                for k in ("valid", "reason", "pair", "t_last_valid"):
                    assert np.array_equal(r[k], e[k]), (name, method, rev, label, k)
                    assert np.array_equal(r[k], u[k]), (name, method, rev, label, k)
                assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
                if label == "colliding start":
                    assert not r["valid"].any()


@pytest.mark.parametrize("path", ["cta", "walker"])
@pytest.mark.parametrize("reverse", [False, True])
def test_dichotomy_cta_per_edge_parity(pkg, eng, ur5, ur5_oracle, golden_configs, monkeypatch, path, reverse):
    """Dichotomy below the wavefront's batch size: the CTA-per-edge kernel (kernels_tail.cu: threads share one
    sample, two-stage narrow phase, replay in the reference order) and, with HPPB200_TAIL=0, the lane-per-edge
    walker.  Golden edges, long and short random edges, zero-length edges; all outputs bit-exact, and the
    iteration cap is reported at the same sample."""
    m, _ = eng
    monkeypatch.setenv("HPPB200_WF_MIN", "100000000")
    monkeypatch.setenv("HPPB200_TAIL", "1" if path == "cta" else "0")
    monkeypatch.setenv("HPPB200_TAIL_MAX", "100000")
    ws = pkg.Workspace(m, 0)
    o = ur5_oracle
    g1, g2 = golden_configs[0:200:2], golden_configs[1:200:2]
    a1, a2, _ = _edges(pkg, ur5, o, 5000, 41)
    b1, b2, _ = _edges(pkg, ur5, o, 3000, 42, short=True)
    z = pkg.uniform_configs(ur5, 300, seed=43)
    q1, q2 = np.concatenate([g1, a1, b1, z]), np.concatenate([g2, a2, b2, z])
    L = o.straight_lengths(q1, q2)
    s0 = ws.stats()
    r = ws.validate_edges(2, 0.0, q1, q2, L, reverse=reverse)
    s1 = ws.stats()
    e = o.validate_edges(2, 0.0, q1, q2, L, reverse=reverse)
    for k in ("valid", "reason", "pair", "t_last_valid"):
        assert np.array_equal(r[k], e[k]), (path, reverse, k)
    assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
    assert s1["iter_cap_hits"] == s0["iter_cap_hits"]
    if path == "cta":  # one launch for the batch (+ at most the second pass for lists beyond 64 entries), behind the
        # first-sample screen (kernels_screen.cu: start configurations, 2 configuration kernels, 2 of the select, finish, gather, scatter)
        assert s1["kernel_launches"] - s0["kernel_launches"] <= 3 + 8
    ws.set_max_iterations(7)
    r = ws.validate_edges(2, 0.0, q1[:2000], q2[:2000], L[:2000], reverse=reverse)
    e = o.validate_edges(2, 0.0, q1[:2000], q2[:2000], L[:2000], reverse=reverse, max_iter=7)
    for k in ("valid", "reason", "pair", "t_last_valid"):
        assert np.array_equal(r[k], e[k]), (path, reverse, k, "cap")
    assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
    assert (r["reason"] == 3).any()


@pytest.mark.parametrize("reverse", [False, True])
def test_cta_per_edge_lists_outgrown_in_kernel_retry(pkg, eng, ur5, ur5_oracle, monkeypatch, reverse):
    """CTA-per-edge kernel with SHORT shared-memory lists (HPPB200_TAIL_K=3; whole batches on large models run with 8 entries
    per pair): an edge that outgrows them is validated again by the same CTA with 64-entry lists in global memory, inside the
    launch.  Same results as the oracle, and no edge is left for a second pass."""
    m, _ = eng
    monkeypatch.setenv("HPPB200_WF_MIN", "100000000")
    monkeypatch.setenv("HPPB200_TAIL_MAX", "100000")
    monkeypatch.setenv("HPPB200_TAIL_K", "3")
    monkeypatch.setenv("HPPB200_FIRST_SAMPLE_SCREEN", "0")
    ws = pkg.Workspace(m, 0)
    o = ur5_oracle
    q1, q2, L = _edges(pkg, ur5, o, 3000, 47)
    s0 = ws.stats()
    r = ws.validate_edges(2, 0.0, q1, q2, L, reverse=reverse)
    s1 = ws.stats()
    e = o.validate_edges(2, 0.0, q1, q2, L, reverse=reverse)
    for k in ("valid", "reason", "pair", "t_last_valid"):
        assert np.array_equal(r[k], e[k]), (reverse, k)
    assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
    assert s1["kernel_launches"] - s0["kernel_launches"] == 1
    monkeypatch.delenv("HPPB200_TAIL_K")  # (read at every launch) 64-entry lists in shared memory: nothing is redone
    u = ws.validate_edges(2, 0.0, q1, q2, L, reverse=reverse)
    s2 = ws.stats()
    assert np.array_equal(u["valid"], r["valid"]) and np.array_equal(u["t_fail"], r["t_fail"], equal_nan=True)
    assert s1["edge_samples"] - s0["edge_samples"] > s2["edge_samples"] - s1["edge_samples"]  # the redone edges' first attempts


@pytest.mark.parametrize("wf_min", ["1", "100000000"])
def test_freeflyer_humanoid_parity(pkg, ob, monkeypatch, wf_min):
    """BASELINE configs[3] model WITH its free-flyer root (R^3 x SO(3), nq = 37, nv = 36): FK from the quaternion,
    Lie-group interpolation along the edge (SO(3) log / exp), velocity bounds from q2 (-) q1, WeighedDistance with
    the rotation logarithm.  Through the wavefront engine (wf_min = 1) and through the CTA engine / walker
    (wavefront off).  Bit-exact against the oracle; quaternions of either sign."""
    monkeypatch.setenv("HPPB200_WF_MIN", wf_min)
    desc = pkg.ModelDescription.load("humanoid_ff_obst6.json")
    ws = pkg.Workspace(pkg.Model(desc), 0)
    o = ob.Oracle(desc)
    q = pkg.uniform_configs(desc, 20000, seed=91)
    assert np.array_equal(ws.geometry_placements(q[:2000]), o.geometry_placements(q[:2000]))
    assert np.array_equal(ws.pair_distances(q[:2000]), o.pair_distances(q[:2000]))
    got, exp = ws.validate_configs(q, check_bounds=True), o.validate_configs(q, check_bounds=True)
    for a, b in zip(got, exp):
        assert np.array_equal(a, b)
    assert 0.02 < got[0].mean() < 0.9
    a1, a2 = _local_edges(pkg, desc, 3000, 92, 0.4)
    b1, b2, _ = _edges(pkg, desc, o, 1000, 93)
    z = pkg.uniform_configs(desc, 100, seed=94)  # zero-length edges
    q1, q2 = np.concatenate([a1, b1, z]), np.concatenate([a2, b2, z])
    L = o.straight_lengths(q1, q2)
    assert np.array_equal(ws.straight_lengths(q1, q2), L)
    for method, param in ((0, 0.05), (1, 1e-3), (2, 0.0)):
        for rev in (False, True):
            r = ws.validate_edges(method, param, q1, q2, L, reverse=rev)
            e = o.validate_edges(method, param, q1, q2, L, reverse=rev)
            assert ws.stats()["iter_cap_hits"] == 0
            for k in ("valid", "reason", "pair", "t_last_valid"):
                assert np.array_equal(r[k], e[k]), (method, rev, k)
            assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
            assert 0.02 < r["valid"].mean() < 0.98


@pytest.mark.parametrize("name", ["ur5_boxes8.json", "humanoid_ff_obst6.json"])
def test_path_eval_and_valid_part(pkg, ob, name):
    """Path::eval / StraightPath::impl_compute for many edges (hppb200_path_eval) against the oracle, end points
    included, and the adapter's validPart = extract(0, t_last_valid) built from it (src/straight-path.cc:142-158)."""
    desc = pkg.ModelDescription.load(name)
    ws = pkg.Workspace(pkg.Model(desc), 0)
    o = ob.Oracle(desc)
    q1, q2 = pkg.uniform_configs(desc, 3000, seed=97), pkg.uniform_configs(desc, 3000, seed=98)
    L = o.straight_lengths(q1, q2)
    L[::7] = 0.0
    u = np.random.default_rng(99).uniform(size=len(L))
    u[::5] = 0.0
    u[1::5] = 1.0
    t = u * L
    assert np.array_equal(ws.path_eval(q1, q2, L, t), o.interpolate(q1, q2, L, t))
    L = o.straight_lengths(q1, q2)
    pv = pkg.Progressive.create(ws, 1e-3)
    e = o.validate_edges(1, 1e-3, q1[:200], q2[:200], L[:200])
    checked = 0
    for i in np.nonzero(e["valid"] == 0)[0][:20]:
        ok, part, rep = pv.validate(pkg.StraightPath(q1[i], q2[i], L[i]))
        assert not ok and rep.parameter == e["t_fail"][i] and part.length == e["t_last_valid"][i]
        assert np.array_equal(part.initial, q1[i])
        assert np.array_equal(part.end, o.interpolate(q1[i], q2[i], L[i:i + 1], e["t_last_valid"][i:i + 1])[0])
        checked += 1
    assert checked > 0


def test_freeflyer_knn_parity(pkg, ob):
    """WeighedDistance k-NN over free-flyer configurations: translation unweighted + w^2 |log3(q_node^-1 q)|^2."""
    desc = pkg.ModelDescription.load("humanoid_ff_obst6.json")
    m, o = pkg.Model(desc), ob.Oracle(desc)
    nn = pkg.NearestNeighbor(m, 0)
    nodes = pkg.uniform_configs(desc, 20000, seed=95)
    nodes[5000] = nodes[77]
    cc = (np.arange(len(nodes)) % 3).astype(np.int32)
    nn.append(nodes, cc)
    o.knn_append(nodes, cc)
    qs = pkg.uniform_configs(desc, 37, seed=96)
    qs[0] = nodes[77]
    for k, c in ((1, -1), (1, 2), (10, -1), (7, 1), (33, -1)):
        gi, gd = nn.query(qs, k, c)
        oi, od = o.knn_query(qs, k, c)
        srt = np.lexsort((oi, od), axis=1)
        oi, od = np.take_along_axis(oi, srt, axis=1), np.take_along_axis(od, srt, axis=1)
        assert np.array_equal(gi, oi), (k, c)
        assert np.array_equal(gd, od)
    balls, _ = nn.within_ball(qs[:4], 2.0)
    for i in range(4):
        assert np.array_equal(balls[i], o.knn_within_ball(qs[i], 2.0))


@pytest.mark.parametrize("reverse", [False, True])
@pytest.mark.parametrize("slots,tail", [(0, -1), (1024, 0), (1024, 64)])
def test_wavefront_dichotomy_parity(pkg, eng, ur5, ur5_oracle, golden_configs, monkeypatch, reverse, slots, tail):
    """Dichotomy through the wavefront engine (interval LISTS per body pair, valid subset per edge), forced for a
    small batch: slot refill, tail hand-off to the walker (tail = 64 slots) and list overflow -> walker are all on
    the path.  Bit-exact against the oracle."""
    m, _ = eng
    monkeypatch.setenv("HPPB200_WF_MIN", "1")
    monkeypatch.setenv("HPPB200_WF_SLOTS", str(slots))
    monkeypatch.setenv("HPPB200_WF_TAIL", str(tail))
    ws = pkg.Workspace(m, 0)
    o = ur5_oracle
    g1, g2 = golden_configs[0:200:2], golden_configs[1:200:2]
    a1, a2, _ = _edges(pkg, ur5, o, 3000, 37)
    b1, b2, _ = _edges(pkg, ur5, o, 2000, 38, short=True)
    z = pkg.uniform_configs(ur5, 300, seed=39)
    q1, q2 = np.concatenate([g1, a1, b1, z]), np.concatenate([g2, a2, b2, z])
    L = o.straight_lengths(q1, q2)
    s0 = ws.stats()
    r = ws.validate_edges(2, 0.0, q1, q2, L, reverse=reverse)
    s1 = ws.stats()
    e = o.validate_edges(2, 0.0, q1, q2, L, reverse=reverse)
    assert 0 < s1["engine_ticks"] - s0["engine_ticks"] < 20000  # wavefront rounds were run
    assert s1["iter_cap_hits"] == 0
    for k in ("valid", "reason", "pair", "t_last_valid"):
        assert np.array_equal(r[k], e[k]), k
    assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)


@pytest.mark.parametrize("name", ["ur5_boxes1.json", "ur5_boxes32.json"])
def test_wavefront_dichotomy_other_scenes(pkg, ob, monkeypatch, name):
    monkeypatch.setenv("HPPB200_WF_MIN", "1")
    desc = pkg.ModelDescription.load(name)
    ws = pkg.Workspace(pkg.Model(desc), 0)
    o = ob.Oracle(desc)
    q1, q2, L = _edges(pkg, desc, o, 2500, 40)
    for rev in (False, True):
        r = ws.validate_edges(2, 0.0, q1, q2, L, reverse=rev)
        e = o.validate_edges(2, 0.0, q1, q2, L, reverse=rev)
        for k in ("valid", "reason", "pair", "t_last_valid"):
            assert np.array_equal(r[k], e[k]), (name, rev, k)
        assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)


from oracle_backend import OracleBackend as _OracleBackend  # noqa: E402  (tests/oracle_backend.py)


@pytest.mark.parametrize("name,method,param,batch", [("ur5_boxes8.json", 1, 1e-3, 0), ("ur5_boxes8.json", 2, 0.0, 0),
                                                     ("ur5_boxes32.json", 1, 1e-3, 16), ("humanoid_ff_obst6.json", 1, 1e-3, 8)])
def test_birrt_end_to_end_same_roadmap(pkg, ob, name, method, param, batch):
    """BASELINE configs[2]: BiRRT (control flow of src/bi-rrt-planner.cc:109-170) on top of the engine's path
    validation, straight-path lengths, path evaluation and nearest neighbour, next to the same planner on the
    oracle: same seed => the same roadmap node for node and edge for edge, the same number of steps, and a
    solution whose every edge the oracle validates."""
    desc = pkg.ModelDescription.load(name)
    o = ob.Oracle(desc)
    cand = pkg.uniform_configs(desc, 400, seed=5)
    free = cand[o.validate_configs(cand)[0] == 1]
    q_init, q_goal = free[0], free[1]
    gpu = pkg.BiRRTPlanner(pkg.EngineBackend(pkg, pkg.Model(desc), 0, method, param), desc, q_init, q_goal, seed=11)
    cpu = pkg.BiRRTPlanner(_OracleBackend(ob.Oracle(desc), method, param), desc, q_init, q_goal, seed=11)
    sg, sc = gpu.solve(max_steps=300, batch=batch), cpu.solve(max_steps=300, batch=batch)
    assert sg == sc and gpu.solved() == cpu.solved()
    assert np.array_equal(gpu.roadmap.nodes, cpu.roadmap.nodes)
    assert gpu.roadmap.edges == cpu.roadmap.edges and np.array_equal(gpu.roadmap.cc, cpu.roadmap.cc)
    assert len(gpu.roadmap.cc) > 4
    if gpu.solved():
        path = gpu.path()
        assert path[0] == gpu.init and path[-1] == gpu.goal and path == cpu.path()
        a, b = gpu.roadmap.nodes[path[:-1]], gpu.roadmap.nodes[path[1:]]
        assert o.validate_edges(0, 0.05, a, b, o.straight_lengths(a, b))["valid"].all()


@pytest.mark.parametrize("check_parameterized", [False, True])
def test_pair_table_edits_gpu_parity(pkg, ob, monkeypatch, check_parameterized):
    """A model edited through the ObstacleUserInterface call log (tests/test_abi.py::pair_op_model): configuration
    validation (flags, first pair in the edited table, distances) and the three path validations agree with the oracle."""
    from test_abi import pair_op_model
    desc = pair_op_model(pkg)
    desc.check_parameterized = check_parameterized
    o = ob.Oracle(desc)
    m = pkg.Model(desc)
    ws = pkg.Workspace(m, 0)
    q = pkg.uniform_configs(desc, 30000, seed=91)
    for a, b in zip(ws.validate_configs(q), o.validate_configs(q)):
        assert np.array_equal(a, b)
    assert np.array_equal(ws.pair_distances(q[:500]), o.pair_distances(q[:500]))
    base = pkg.Workspace(pkg.Model(pkg.ModelDescription.load("ur5_boxes8.json")), 0)
    assert not np.array_equal(base.validate_configs(q)[0], ws.validate_configs(q)[0])  # the edits change flags
    q1, q2, L = _edges(pkg, desc, o, 3000, 92)
    for wf in ("100000000", "1"):
        monkeypatch.setenv("HPPB200_WF_MIN", wf)
        w2 = pkg.Workspace(m, 0)
        for meth, par in ((0, 0.05), (1, 1e-3), (2, 0.0)):
            r = w2.validate_edges(meth, par, q1, q2, L)
            e = o.validate_edges(meth, par, q1, q2, L)
            for k in ("valid", "reason", "pair", "t_last_valid"):
                assert np.array_equal(r[k], e[k]), (wf, meth, k)
            assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)


@pytest.mark.parametrize("method,param", [(1, 1e-3), (2, 0.0)])
@pytest.mark.parametrize("reverse", [False, True])
def test_edge_chains_pathvector_recursion(pkg, eng, ur5, ur5_oracle, method, param, reverse):
    """PathVectors of straight segments (what RandomShortcut hands to pathValidation()->validate): the engine validates
    all leaves as one batch and cuts each chain at its first failing leaf on the device; the oracle walks the
    sub-paths one by one like ContinuousValidation::validate (src/continuous-validation.cc:153-192)."""
    m, ws = eng
    rng = np.random.default_rng(17)
    way = pkg.uniform_configs(ur5, 4000, seed=71)
    # chains through consecutive way points, short steps so that many chains are valid for several leaves
    lens = rng.integers(0, 7, size=600)  # includes empty PathVectors
    q1, q2, cb = [], [], [0]
    k = 0
    for n in lens:
        a = way[k]
        for j in range(n):
            d = way[k + j + 1] - a
            b = np.clip(a + 0.4 * d / np.linalg.norm(d), -np.pi, np.pi)
            q1.append(a); q2.append(b)
            a = b
        k += n + 1
        cb.append(len(q1))
    q1, q2 = np.array(q1), np.array(q2)
    L = ur5_oracle.straight_lengths(q1, q2)
    r = ws.validate_edge_chains(method, param, q1, q2, L, cb, reverse=reverse)
    e = ur5_oracle.validate_edge_chains(method, param, q1, q2, L, cb, reverse=reverse)
    for key in ("valid", "reason", "pair", "first_bad_edge", "t_last_valid"):
        assert np.array_equal(r[key], e[key]), key
    assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
    assert 0.05 < r["valid"].mean() < 0.95 and (r["first_bad_edge"] > np.array(cb[:-1])).any()  # failures beyond the first leaf
    empty = np.nonzero(lens == 0)[0]
    assert r["valid"][empty].all() and (r["t_last_valid"][empty] == 0).all()
    with pytest.raises(pkg.EngineError):
        ws.validate_edge_chains(0, 0.05, q1, q2, L, cb)


def test_multi_gpu_device_list_entry(pkg, eng, ur5, ur5_oracle):
    """hppb200_multi_*: a device list sharded internally.  Three shards (on one GPU box all on device 0, which exercises the
    same splitting / threading / placement; on a multi-GPU box every visible device too): same outputs as the
    single-GPU entry points and the oracle, ragged sizes, errors reported with the device."""
    m, ws = eng
    ndev = pkg.lib().hppb200_device_count()
    for devices in ([0, 0, 0], list(range(ndev)) if ndev > 1 else None):
        mg = pkg.MultiGpu(m, devices)
        assert mg.ndevices == (len(devices) if devices else ndev)
        q = pkg.uniform_configs(ur5, 100003, seed=81) * 1.03
        got = mg.validate_configs(q, check_bounds=True)
        ref = ws.validate_configs(q, check_bounds=True)
        for a, b in zip(got, ref):
            assert np.array_equal(a, b)
        sub = slice(0, 20000)
        for a, b in zip(got, ur5_oracle.validate_configs(q[sub], check_bounds=True)):
            assert np.array_equal(a[sub], b)
        assert all(mg.shard_stats(r)["configs"] > 0 for r in range(mg.ndevices))
        q1, q2, L = _edges(pkg, ur5, ur5_oracle, 5001, 82)
        for meth, par in ((1, 1e-3), (2, 0.0), (0, 0.05)):
            r = mg.validate_edges(meth, par, q1, q2, L, reverse=(meth == 2))
            e = ur5_oracle.validate_edges(meth, par, q1, q2, L, reverse=(meth == 2))
            for k in ("valid", "reason", "pair", "t_last_valid"):
                assert np.array_equal(r[k], e[k]), (meth, k)
            assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
        v, f, r = mg.validate_configs(q[:2])  # fewer units than shards
        assert np.array_equal(v, ref[0][:2])
        with pytest.raises(RuntimeError, match="device"):
            mg.validate_edges(1, 0.0, q1, q2, L)  # Progressive needs tol > 0: the shard's error comes back with its device
    with pytest.raises(ValueError):
        pkg.MultiGpu(m, [99])


def test_exchange_helpers_pack_mask_and_knn_shard_merge(pkg, eng, ur5):
    """Device-side helpers of the sharded form: bit-packed validity masks (n not a multiple of 32) and the W-way merge of
    per-shard k-NN candidates against the unsharded query, duplicates across shards included (the earlier GLOBAL index wins)."""
    import torch
    from hpp_core_b200 import sharding as sh
    m, ws = eng
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(3)
    v = torch.from_numpy((rng.random(100003) < 0.4).astype(np.uint8)).to(dev)
    out, bits = sh.all_gather_mask_packed(v)
    assert torch.equal(out, v) and bits.numel() == (100003 + 31) // 32
    nodes = pkg.uniform_configs(ur5, 40000, seed=91)
    nodes[30000] = nodes[17]      # exact duplicates in different shards
    nodes[39999] = nodes[25000]
    qs = pkg.uniform_configs(ur5, 257, seed=92)
    qs[0], qs[1] = nodes[17], nodes[25000]
    whole = pkg.NearestNeighbor(m, 0)
    whole.append(nodes)
    for k in (1, 10, 33):
        ri, rd = whole.query(qs, k)
        W = 3
        gi, gd = [], []
        for b, e in sh.node_owner_ranges(len(nodes), W):
            part = pkg.NearestNeighbor(m, 0)
            part.append(nodes[b:e])
            li, ld = part.query(qs, k)
            li = np.where(li >= 0, li + b, li)
            gi.append(li); gd.append(ld)
        gi = torch.from_numpy(np.stack(gi)).to(dev).contiguous()
        gd = torch.from_numpy(np.stack(gd)).to(dev).contiguous()
        oi, od = sh.merge_topk_device(gi, gd, k)
        torch.cuda.synchronize()
        assert np.array_equal(oi.cpu().numpy(), ri) and np.array_equal(od.cpu().numpy(), rd)
        hi, hd = sh.merge_topk([x for x in gd.cpu().numpy()], [x for x in gi.cpu().numpy()], k)
        assert np.array_equal(hi, ri) and np.array_equal(hd, rd)


@pytest.mark.parametrize("name,check_parameterized", [("ur5_boxes8.json", False), ("ur5_boxes32.json", False), ("edited", True)])
def test_all_collisions_and_distance_between_objects(pkg, ob, name, check_parameterized):
    """CollisionValidation with computeAllContacts (AllCollisionsValidationReport: the first colliding pair and
    every later colliding pair of the same table, src/collision-validation.cc:72-89) and DistanceBetweenObjects
    (src/distance-between-objects.cc:41-77), against the oracle, bit for bit."""
    if name == "edited":
        from test_abi import pair_op_model
        desc = pair_op_model(pkg)
        desc.check_parameterized = check_parameterized
    else:
        desc = pkg.ModelDescription.load(name)
    o = ob.Oracle(desc)
    ws = pkg.Workspace(pkg.Model(desc), 0)
    q = pkg.uniform_configs(desc, 4000, seed=123)
    gv, gf, gl, gc = ws.all_collisions(q)
    ov, of, ol, oc = o.all_collisions(q)
    assert np.array_equal(gv, ov) and np.array_equal(gf, of) and np.array_equal(gc, oc)
    assert all(np.array_equal(a, b) for a, b in zip(gl, ol))
    assert gc.max() > 1 and np.array_equal(gv, ws.validate_configs(q)[0])  # several contacts at once do occur; flags as validate
    d, e = ws.distance_between_objects(q[:1500]), o.distance_between_objects(q[:1500])
    assert d.shape[1] == desc.nobstacles * int((desc.geom_joint > 0).sum()) and np.array_equal(d, e)


def test_planner_front_ends_engine_equals_oracle(pkg, ob):
    """SURVEY 8f-1/8f-2 on the engine: DiffusingPlanner and VisibilityPrmPlanner steps (batched inner loops) and the
    speculative RandomShortcut build, on the GPU, exactly the roadmap / path they build on the CPU oracle."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from oracle_backend import OracleBackend
    desc = pkg.ModelDescription.load("ur5_boxes8.json")
    o = ob.Oracle(desc)
    cand = pkg.uniform_configs(desc, 400, seed=5)
    free = cand[o.validate_configs(cand, check_bounds=True)[0] == 1]
    q0, q1 = free[0], free[1]
    model = pkg.Model(desc)
    for cls, steps in ((pkg.DiffusingPlanner, 30), (pkg.VisibilityPrmPlanner, 60)):
        gpu = cls(pkg.EngineBackend(pkg, model, 0, 1, 1e-3), desc, q0, q1, seed=31)
        cpu = cls(OracleBackend(ob.Oracle(desc), 1, 1e-3), desc, q0, q1, seed=31)
        for _ in range(steps):
            gpu.one_step_batch()
            cpu.one_step_batch()
        assert np.array_equal(gpu.roadmap.nodes, cpu.roadmap.nodes) and gpu.roadmap.n > 10
        assert gpu.roadmap.edges == cpu.roadmap.edges and np.array_equal(gpu.roadmap.cc, cpu.roadmap.cc)
    planner = pkg.BiRRTPlanner(pkg.EngineBackend(pkg, model, 0, 1, 1e-3), desc, q0, q1, seed=4)
    planner.solve(max_steps=400)
    way = planner.roadmap.nodes[planner.path()]
    g = pkg.RandomShortcut(pkg.EngineBackend(pkg, model, 0, 1, 1e-3), desc, 8, seed=9).optimize(way, batch=True)
    c = pkg.RandomShortcut(OracleBackend(ob.Oracle(desc), 1, 1e-3), desc, 8, seed=9).optimize(way, batch=True)
    assert np.array_equal(g, c) and len(g) >= 2
    # SimpleShortcut (all waypoint pairs in one edge batch + A*) and PartialShortcut (joint-wise straightening, speculative)
    g = pkg.SimpleShortcut(pkg.EngineBackend(pkg, model, 0, 1, 1e-3)).optimize(way)
    c = pkg.SimpleShortcut(OracleBackend(ob.Oracle(desc), 1, 1e-3)).optimize(way)
    assert np.array_equal(g, c) and 2 <= len(g) <= len(way)
    gp = pkg.PartialShortcut(pkg.EngineBackend(pkg, model, 0, 1, 1e-3), desc)
    cp = pkg.PartialShortcut(OracleBackend(ob.Oracle(desc), 1, 1e-3), desc)
    g, c = gp.optimize(way, batch=True), cp.optimize(way, batch=False)
    assert np.array_equal(g, c) and gp.failed_joints == cp.failed_joints


def test_validate_edges_indexed_over_the_roadmap_mirror(pkg, eng, ur5, ur5_oracle):
    """hppb200_validate_edges_indexed: edges between nodes of the device mirror by index == the same edges given by their
    end configurations (and the oracle); bad indices are rejected."""
    m, ws = eng
    nn = pkg.NearestNeighbor(m, 0)
    nodes = pkg.uniform_configs(ur5, 5000, seed=141)
    nn.append(nodes[:3000])
    nn.append(nodes[3000:])
    rng = np.random.default_rng(5)
    a, b = rng.integers(0, 5000, 4000), rng.integers(0, 5000, 4000)
    a[:5] = b[:5]  # zero-length edges too
    for meth, par in ((1, 1e-3), (2, 0.0), (0, 0.05)):
        r = ws.validate_edges_indexed(nn, meth, par, a, b)
        L = ur5_oracle.straight_lengths(nodes[a], nodes[b])
        e = ur5_oracle.validate_edges(meth, par, nodes[a], nodes[b], L)
        assert np.array_equal(r["length"], L)
        for k in ("valid", "reason", "pair", "t_last_valid"):
            assert np.array_equal(r[k], e[k]), (meth, k)
        assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True)
    with pytest.raises(ValueError):
        ws.validate_edges_indexed(nn, 1, 1e-3, [0, 5000], [1, 2])
    with pytest.raises(ValueError):
        ws.validate_edges_indexed(nn, 1, 1e-3, [0, -1], [1, 2])


def test_knn_single_query_latency_path(pkg, eng, ur5, ob):
    """One query per call (BiRRT's nearestNode): the one-kernel path (query in the parameter block, last CTA merges and
    writes the result to mapped pinned memory) == the oracle == the tiled scan it replaces, ties by node index, component
    filter, fewer than k candidates, repeated calls (the ticket counter is reset by the kernel)."""
    import os
    humanoid = pkg.ModelDescription.load("humanoid_ff_obst6.json")
    for desc, n in ((ur5, 70001), (humanoid, 9000)):  # revolute chain / free-flyer root (SO(3) distance)
        m = pkg.Model(desc)
        o = ob.Oracle(desc)
        nn = pkg.NearestNeighbor(m, 0)
        nodes = pkg.uniform_configs(desc, n, seed=301)
        nodes[n - 1] = nodes[17]  # exact duplicate: the earlier node wins
        cc = (np.arange(n) % 3).astype(np.int32)
        nn.append(nodes, cc)
        o.knn_append(nodes, cc)
        os.environ["HPPB200_KNN_LATENCY"] = "0"
        try:
            tiled = pkg.NearestNeighbor(m, 0)
        finally:
            del os.environ["HPPB200_KNN_LATENCY"]
        tiled.append(nodes, cc)
        qs = pkg.uniform_configs(desc, 12, seed=302)
        qs[0] = nodes[17]
        for k, c in ((1, -1), (1, 1), (3, 2), (10, -1), (32, -1), (32, 0)):
            for q in qs:
                gi, gd = nn.query(q[None, :], k, c)
                oi, od = o.knn_query(q[None, :], k, c)
                srt = np.lexsort((oi, od), axis=1)
                oi, od = np.take_along_axis(oi, srt, axis=1), np.take_along_axis(od, srt, axis=1)
                ti, td = tiled.query(q[None, :], k, c)
                assert np.array_equal(gi, oi) and np.array_equal(gd, od), (desc.name, k, c)
                assert np.array_equal(gi, ti) and np.array_equal(gd, td)
        assert nn.query(qs[:1], 1)[0][0, 0] == 17
    small = pkg.NearestNeighbor(pkg.Model(ur5), 0)
    small.append(pkg.uniform_configs(ur5, 3, seed=303))
    gi, gd = small.query(pkg.uniform_configs(ur5, 1, seed=304), 10)
    assert np.array_equal(gi[0, 3:], -np.ones(7, np.int64)) and np.all(np.isinf(gd[0, 3:])) and sorted(gi[0, :3]) == [0, 1, 2]


@pytest.mark.parametrize("name", ["ur5_boxes8.json", "humanoid_obst6.json"])
def test_spline_edges_parity(pkg, ob, name):
    """path::Spline<BernsteinBasis, 3> edges (tests/test-continuous-validation.cc:288-427) through
    hppb200_validate_spline_edges == the oracle (Spline eval, impl_velocityBound on sub-intervals, the two-round
    refinement of BodyPairCollision::collisionFreeInterval), all three methods, forward and reverse, bit for bit;
    zero-length splines, joint-bound failures of Discretized, the iteration cap."""
    desc = pkg.ModelDescription.load(name)
    assert desc.nq == sum(7 if int(t) == 6 else 1 for t in desc.joint_type)
    o = ob.Oracle(desc)
    m = pkg.Model(desc)
    ws = pkg.Workspace(m, 0)
    n = 3000 if name.startswith("ur5") else 600
    cfg = pkg.uniform_configs(desc, 2 * n, seed=77)
    q1, q2 = np.ascontiguousarray(cfg[0::2]), np.ascontiguousarray(cfg[1::2])
    q2[:40] = q1[:40]  # zero-length: L = 0
    rng = np.random.default_rng(78)
    v1, v2 = rng.uniform(-1, 1, q1.shape), rng.uniform(-1, 1, q1.shape)
    v1[:40] = 0
    v2[:40] = 0
    L = o.straight_lengths(q1, q2)
    base, P = pkg.spline_steer(q1, v1, q2, v2, L)
    for meth, par in ((1, 1e-2), (1, 1e-3), (2, 0.0), (0, 0.05)):
        for rev in (False, True):
            r = ws.validate_spline_edges(meth, par, base, P, L, reverse=rev)
            e = o.validate_spline_edges(meth, par, base, P, L, reverse=rev)
            for k in ("valid", "reason", "pair", "t_last_valid"):
                assert np.array_equal(r[k], e[k]), (name, meth, par, rev, k, int((r[k] != e[k]).sum()))
            assert np.array_equal(r["t_fail"], e["t_fail"], equal_nan=True), (name, meth, par, rev)
            assert 0.02 < r["valid"].mean() < 0.98
    # Discretized with the joint-bound check: random end velocities push splines out of the joint range
    r = ws.validate_spline_edges(0, 0.05, base, P, L, check_bounds=True)
    e = o.validate_spline_edges(0, 0.05, base, P, L, check_bounds=True)
    for k in ("valid", "reason", "pair", "t_last_valid"):
        assert np.array_equal(r[k], e[k]), k
    assert np.any(r["reason"] == pkg.REASON_JOINT_BOUND)
    # iteration cap
    ws.set_max_iterations(3)
    try:
        r = ws.validate_spline_edges(1, 1e-3, base[:500], P[:500], L[:500])
        e = o.validate_spline_edges(1, 1e-3, base[:500], P[:500], L[:500], max_iter=3)
        for k in ("valid", "reason", "t_last_valid"):
            assert np.array_equal(r[k], e[k]), k
    finally:
        ws.set_max_iterations(1000000)
    # free-flyer models are refused (vector-space configuration spaces only)
    ff = pkg.ModelDescription.load("humanoid_ff_obst6.json")
    wf = pkg.Workspace(pkg.Model(ff), 0)
    with pytest.raises(ValueError):
        wf.validate_spline_edges(1, 1e-3, np.zeros((1, ff.nq)), np.zeros((1, 4, ff.nq)), np.ones(1))


def test_peer_mask_exchange_single_rank_matches_pack_mask(pkg, eng, ur5):
    """hppb200_exchange_mask with world = 1 (the fused pack + peer-store + flag kernel writing into its own buffer): the gathered
    words equal hppb200_pack_mask's on three calls in a row (both slot parities, ragged n); the world > 1 form is checked against
    the NCCL all-gather by bench.py on every multi-GPU run."""
    import ctypes as C
    import torch
    _, ws = eng
    lib = pkg.lib()
    dev = torch.device("cuda", 0)
    h = C.c_void_p()
    handle = (C.c_ubyte * 64)()
    assert lib.hppb200_exchange_create(C.c_int(0), C.c_int(0), C.c_int(1), C.c_size_t(100000), C.byref(h), handle) == 0
    assert lib.hppb200_exchange_open_peers(h, C.c_char_p(bytes(handle))) == 0
    try:
        for n, seed in ((100000, 1), (99999, 2), (77, 3)):
            valid = (torch.rand(n, device=dev, generator=torch.Generator(device=dev).manual_seed(seed)) < 0.4).to(torch.uint8)
            bits = torch.zeros((n + 31) // 32, dtype=torch.int32, device=dev)
            stream = torch.cuda.current_stream().cuda_stream
            assert lib.hppb200_pack_mask(C.c_void_p(stream), C.c_void_p(valid.data_ptr()), C.c_size_t(n), C.c_void_p(bits.data_ptr())) == 0
            out = C.c_void_p()
            assert lib.hppb200_exchange_mask(h, C.c_void_p(stream), C.c_void_p(valid.data_ptr()), C.c_size_t(n), C.byref(out)) == 0
            torch.cuda.synchronize()
            from hpp_core_b200 import sharding as sh
            got = torch.as_tensor(sh.PeerMaskExchange._Raw(out.value, bits.numel()), device=dev)
            assert torch.equal(got, bits), n
        out = C.c_void_p()
        assert lib.hppb200_exchange_mask(h, C.c_void_p(0), C.c_void_p(valid.data_ptr()), C.c_size_t(1 << 22), C.byref(out)) != 0  # beyond max_units
    finally:
        lib.hppb200_exchange_destroy(h)
