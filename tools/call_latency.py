#!/usr/bin/env python3
"""Per-call latency of the one-item calls a planner makes through the C ABI (host pointers): one configuration, one nearest-node
   query over 1 M nodes, one straight length, one edge (Progressive / Dichotomy, short and long).   python tools/call_latency.py"""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load()
desc = pkg.ModelDescription.load("ur5_boxes8.json")
model = pkg.Model(desc)
ws = pkg.Workspace(model, 0)
nn = pkg.NearestNeighbor(model, 0)
nn.append(pkg.uniform_configs(desc, 1 << 20, seed=3))
q = pkg.uniform_configs(desc, 4096, seed=5)
valid = ws.validate_configs(q)[0].astype(bool)
free = q[valid]
def bench(name, fn, reps=300):
    for _ in range(20): fn()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    print("%-46s %8.1f us per call" % (name, (time.perf_counter() - t0) / reps * 1e6), flush=True)
bench("validate_configs, 1 configuration", lambda: ws.validate_configs(q[:1]))
bench("knn query, 1 query, k = 1, 1 M nodes", lambda: nn.query(q[:1], 1))
a, b = free[:1], free[:1] + 0.05
bench("straight_lengths, 1 edge", lambda: ws.straight_lengths(a, b))
L = ws.straight_lengths(a, b)
for meth, par, nm in ((1, 1e-3, "Progressive"), (2, 0.0, "Dichotomy"), (0, 0.05, "Discretized")):
    r = ws.validate_edges(meth, par, a, b, L)
    s0 = ws.stats()["edge_samples"]; ws.validate_edges(meth, par, a, b, L); ns = ws.stats()["edge_samples"] - s0
    bench("validate_edges %s, 1 short edge (%d samples, valid %d)" % (nm, ns, r["valid"][0]), lambda: ws.validate_edges(meth, par, a, b, L))
a2, b2 = free[1:2], free[2:3]
L2 = ws.straight_lengths(a2, b2)
for meth, par, nm in ((1, 1e-3, "Progressive"), (2, 0.0, "Dichotomy")):
    s0 = ws.stats()["edge_samples"]; r = ws.validate_edges(meth, par, a2, b2, L2); ns = ws.stats()["edge_samples"] - s0
    bench("validate_edges %s, 1 long edge (%d samples, valid %d)" % (nm, ns, r["valid"][0]), lambda: ws.validate_edges(meth, par, a2, b2, L2), reps=100)
