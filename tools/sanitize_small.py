"""Small invocation of every kernel family, meant to run under compute-sanitizer
(memcheck / racecheck / synccheck / initcheck): sizes are picked so that a ~100x slowdown stays
within minutes.  Usage: compute-sanitizer --tool racecheck python tools/sanitize_small.py [family ...]
Families: config edges wavefront dich tail knn knn_tc humanoid.  No oracle, no parity: this is a
memory/race check only (parity is tests/test_gpu_parity.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load()
fams = sys.argv[1:] or ["config", "edges", "wavefront", "dich", "tail", "knn", "knn_tc", "humanoid"]


def edges(desc, n, seed, short=False):
    q = pkg.uniform_configs(desc, 2 * n, seed=seed)
    q1, q2 = q[0::2].copy(), q[1::2].copy()
    if short:
        q2 = q1 + 0.1 * (q2 - q1)
    return q1, q2


desc = pkg.ModelDescription.load("ur5_boxes8.json")
model = pkg.Model(desc)
ws = pkg.Workspace(model, 0)

if "config" in fams:
    q = pkg.uniform_configs(desc, 3000 + 37, seed=1)
    v, f, r = ws.validate_configs(q, check_bounds=True)
    ws.geometry_placements(q[:200]); ws.pair_distances(q[:200])
    print("config ok", v.mean())

if "edges" in fams:
    q1, q2 = edges(desc, 300, 2)
    L = ws.straight_lengths(q1, q2)
    for m, p in ((0, 0.05), (1, 1e-3), (2, 0.0)):
        os.environ["HPPB200_WF_MIN"] = "100000000"
        w = pkg.Workspace(model, 0)
        r = w.validate_edges(m, p, q1, q2, L)
        print("edges method", m, r["valid"].mean())
    ws.path_eval(q1, q2, L, 0.5 * L)

if "wavefront" in fams:
    os.environ["HPPB200_WF_MIN"] = "1"
    os.environ["HPPB200_WF_SLOTS"] = "256"
    w = pkg.Workspace(model, 0)
    q1, q2 = edges(desc, 600, 3)
    L = w.straight_lengths(q1, q2)
    r = w.validate_edges(1, 1e-3, q1, q2, L)
    print("wavefront progressive", r["valid"].mean())

if "dich" in fams:
    os.environ["HPPB200_WF_MIN"] = "1"
    os.environ["HPPB200_WF_SLOTS"] = "256"
    w = pkg.Workspace(model, 0)
    q1, q2 = edges(desc, 400, 4)
    L = w.straight_lengths(q1, q2)
    r = w.validate_edges(2, 0.0, q1, q2, L)
    print("wavefront dichotomy", r["valid"].mean())

if "tail" in fams:
    os.environ["HPPB200_WF_MIN"] = "100000000"
    w = pkg.Workspace(model, 0)
    q1, q2 = edges(desc, 64, 5)
    L = w.straight_lengths(q1, q2)
    r = w.validate_edges(2, 0.0, q1, q2, L)
    print("dichotomy tail", r["valid"].mean())

if "knn" in fams:
    nn = pkg.NearestNeighbor(model, 0)
    nodes = pkg.uniform_configs(desc, 5000, seed=6)
    nn.append(nodes, np.arange(5000, dtype=np.int32) % 3)
    qs = pkg.uniform_configs(desc, 9, seed=7)
    for k in (1, 10, 40):
        i, d = nn.query(qs, k)
    nn.query(qs, 5, cc=1)
    nn.within_ball(qs, 1.5)
    print("knn ok")

if "knn_tc" in fams:
    nn = pkg.NearestNeighbor(model, 0)
    nodes = pkg.uniform_configs(desc, 20000, seed=8)
    nn.append(nodes)
    qs = pkg.uniform_configs(desc, 64, seed=9)
    for k in (1, 10, 40):
        i, d = nn.query(qs, k)
    print("knn_tc", nn.tc_stats())

if "humanoid" in fams:
    for name in ("humanoid_obst6.json", "humanoid_ff_obst6.json"):
        d2 = pkg.ModelDescription.load(name)
        m2 = pkg.Model(d2)
        os.environ["HPPB200_WF_MIN"] = "1"
        os.environ["HPPB200_WF_SLOTS"] = "128"
        w = pkg.Workspace(m2, 0)
        q = pkg.uniform_configs(d2, 400, seed=10)
        v, f, r = w.validate_configs(q)
        q1 = q[:150]
        q2 = q1 + 0.05 * (q[150:300] - q1)
        if "ff" in name:
            q2 = pkg.normalize_quaternions(d2, q2)
        L = w.straight_lengths(q1, q2)
        for mth, p in ((1, 1e-3), (2, 0.0)):
            r = w.validate_edges(mth, p, q1, q2, L)
            print(name, mth, r["valid"].mean())
print("sanitize_small done")
