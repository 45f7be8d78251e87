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
    assert L.hppb200_abi_version() == 1


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
