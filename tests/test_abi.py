"""CPU tests: the C-ABI library loads, exports every symbol include/hppb200.h declares, and its
host-side model compiler (pair order, body pairs, velocity coefficients) agrees with the oracle.
No compute calls (no GPU here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_header_symbols(pkg):
    hdr = open(os.path.join(ROOT, "include", "hppb200.h")).read()
    declared = sorted(set(re.findall(r"\b(hppb200_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 30
    assert sorted(pkg.ABI_SYMBOLS) == declared
    L = pkg.lib()
    for s in declared:
        assert hasattr(L, s), s
    assert L.hppb200_abi_version() == 2


def test_struct_layout_matches_header(pkg):
    """ctypes mirror of hppb200_model_desc has the same field order as the header."""
    hdr = open(os.path.join(ROOT, "include", "hppb200.h")).read()
    body = re.search(r"typedef struct hppb200_model_desc \{(.*?)\} hppb200_model_desc;", hdr, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = re.findall(r"(\w+);", body)
    assert names == [f[0] for f in pkg.hppb200_model_desc._fields_]


@pytest.mark.parametrize("name", ["ur5_boxes0.json", "ur5_boxes1.json", "ur5_boxes8.json", "ur5_boxes32.json",
                                  "humanoid_obst6.json", "humanoid_obst16.json", "humanoid_ff_obst6.json"])
def test_model_tables_match_oracle(pkg, ob, name):
    desc = pkg.ModelDescription.load(name)
    m = pkg.Model(desc)
    o = ob.Oracle(desc)
    assert m.nq == o.nq == desc.nq and m.nv == o.nv
    for x, y in zip(m.config_pairs(), o.config_pairs()):
        assert np.array_equal(x, y)
    bm, bo = m.body_pairs(), o.body_pairs()
    assert bm.keys() == bo.keys()
    for k in bm:
        assert np.array_equal(bm[k], bo[k]), k
    assert np.array_equal(m.body_radii(), o.body_radii())


def test_security_margins_and_errors(pkg, ob):
    desc = pkg.ModelDescription.load("ur5_boxes1.json")
    desc.security_margins = np.arange(49, dtype=np.float64).reshape(7, 7) * 1e-3
    m, o = pkg.Model(desc), ob.Oracle(desc)
    for x, y in zip(m.config_pairs(), o.config_pairs()):
        assert np.array_equal(x, y)
    assert m.config_pairs()[2].max() > 0
    bad = pkg.ModelDescription.load("ur5_boxes1.json")
    bad.joint_parent = bad.joint_parent.copy()
    bad.joint_parent[2] = 5  # parent after child
    with pytest.raises(ValueError):
        pkg.Model(bad)


def pair_op_model(pkg, name="ur5_boxes8.json"):
    """A model whose validators received every ObstacleUserInterface call after construction (include/hpp/core/
    obstacle-user.hh:47-88): obstacle 7 only attached to joint 2 and its children, obstacle 0 removed from joint 3,
    a relative-motion filter (joints 1-2 constrained, 4-6 parameterized), a margin matrix, one margin between two bodies."""
    from hpp_core_b200 import model_io as mio
    desc = pkg.ModelDescription.load(name)
    nj = desc.njoints
    og = np.ones(desc.nobstacles, np.uint8)
    og[7] = 0
    og[6] = 0   # obstacle 6 is added by addObstacle() AFTER the margin matrix and the filter: margin 0, not filtered
    desc.obst_global = og
    rel = np.full((nj + 1, nj + 1), 2, np.int32)
    np.fill_diagonal(rel, 0)
    rel[1, 2] = rel[2, 1] = 0
    rel[4, 6] = rel[6, 4] = 1
    rel[5, 0] = rel[0, 5] = 0   # joint 5 rigidly fixed to the environment: its obstacle pairs go away
    mar = (np.arange((nj + 1) ** 2, dtype=np.float64).reshape(nj + 1, nj + 1) % 5) * 2e-3
    mar = np.minimum(mar, mar.T)
    geom_j3 = int(np.nonzero(desc.geom_joint == 3)[0][0])
    desc.pair_ops = [(mio.OP_ADD_OBSTACLE_TO_JOINT, 7, 2, 1, 0.0, None),
                     (mio.OP_REMOVE_OBSTACLE_FROM_JOINT, 0, 3, 0, 0.0, None),
                     (mio.OP_SET_SECURITY_MARGINS, 0, 0, 0, 0.0, mar),
                     (mio.OP_FILTER_COLLISION_PAIRS, 0, 0, 0, 0.0, rel),
                     (mio.OP_ADD_OBSTACLE, 6, 0, 0, 0.0, None),
                     (mio.OP_SET_MARGIN_BETWEEN_BODIES, geom_j3, -2, 0, 0.0123, None)]
    return desc


@pytest.mark.parametrize("check_parameterized", [False, True])
def test_pair_table_edits_match_oracle(pkg, ob, check_parameterized):
    """addObstacleToJoint / removeObstacleFromJoint / filterCollisionPairs / setSecurityMargins /
    setSecurityMarginBetweenBodies (reference src/obstacle-user.cc:60-220, src/continuous-validation.cc:224-367):
    the engine's model compiler and the oracle apply the same call log to both pair tables."""
    desc = pair_op_model(pkg)
    desc.check_parameterized = check_parameterized
    base = pkg.Model(pkg.ModelDescription.load("ur5_boxes8.json"))
    m, o = pkg.Model(desc), ob.Oracle(desc)
    a, b, mg = m.config_pairs()
    for x, y in zip((a, b, mg), o.config_pairs()):
        assert np.array_equal(x, y)
    assert len(a) != len(base.config_pairs()[0])
    npar = pkg.lib().hppb200_model_num_parameterized_pairs(m.h)
    assert (npar > 0) == check_parameterized
    gj = desc.geom_joint
    jb = np.where(b >= 0, gj[np.maximum(b, 0)], 0)
    # obstacle 7 only on joints >= 2; obstacle 0 not on joint 3; constrained joint pairs gone; one edited margin
    assert not np.any((b == -8) & (gj[a] < 2)) and np.any(b == -8)
    assert not np.any((b == -1) & (gj[a] == 3))
    assert not np.any((gj[a] == 5) & (jb == 0) & (b != -7))  # (-7: the obstacle added after the filter)
    assert 0.0123 in mg
    assert np.any((b == -7) & (gj[a] == 5)) and np.all(mg[b == -7] == 0)  # late obstacle: unfiltered, default margin
    bm, bo = m.body_pairs(), o.body_pairs()
    for k in bm:
        assert np.array_equal(bm[k], bo[k]), k
    late = np.zeros(len(bm["joint_a"]), bool)
    for i in range(len(late)):
        late[i] = np.all(bm["geomB"][bm["pair_begin"][i]:bm["pair_begin"][i + 1]] == -7)
    assert not np.any((bm["joint_a"] == 5) & (bm["joint_b"] == 0) & ~late) and np.any((bm["joint_a"] == 5) & late)
    assert not np.any(bm["geomB"] == -8)  # ContinuousValidation has no addObstacleToJoint
    assert 0.0123 in bm["margin"]


def test_pair_table_edit_errors(pkg, ob):
    from hpp_core_b200 import model_io as mio
    desc = pkg.ModelDescription.load("ur5_boxes8.json")
    desc.pair_ops = [(mio.OP_REMOVE_OBSTACLE_FROM_JOINT, 0, 3, 0, 0.0, None)] * 2  # second removal: not registered any more
    with pytest.raises(pkg.EngineError):  # std::runtime_error in the reference (obstacle-user.cc:104-110)
        pkg.Model(desc)
    desc.pair_ops = [(mio.OP_SET_MARGIN_BETWEEN_BODIES, 0, 0, 0, 0.1, None)]  # no such pair: std::invalid_argument (:216-219)
    with pytest.raises(ValueError):
        pkg.Model(desc)
    desc.pair_ops = [(99, 0, 0, 0, 0.0, None)]
    with pytest.raises(ValueError):
        pkg.Model(desc)


def test_no_cpu_fallback(pkg):
    """Without a CUDA device every compute entry point must fail loudly."""
    if pkg.lib().hppb200_device_count() > 0:
        pytest.skip("GPU present")
    m = pkg.Model("ur5_boxes1.json")
    with pytest.raises(pkg.EngineError):
        pkg.Workspace(m, 0)
    with pytest.raises(pkg.EngineError):
        pkg.NearestNeighbor(m, 0)


def test_validator_constructor_checks(pkg):
    """Progressive needs tol > 0 (progressive.cc:138-145), Dichotomy tol == 0 (dichotomy.cc:139-148)."""
    with pytest.raises(RuntimeError):
        pkg.Progressive(None, 0.0)
    with pytest.raises(RuntimeError):
        pkg.Dichotomy(None, 1e-3)
    p = pkg.StraightPath([0.0, 0.0], [1.0, 2.0], 2.0)
    e = p.extract(0.0, 1.0)
    assert np.allclose(e.end, [0.5, 1.0]) and e.length == 1.0
