#!/usr/bin/env python3
"""The longest chains of a batch, validated alone: per-sample latency of the CTA-per-edge kernel without neighbours.
   python tools/monster_edges.py [MODEL.json] [n] [local_step]   (oracle sample counts pick the edges)"""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load()
from oracle import binding as ob
name = sys.argv[1] if len(sys.argv) > 1 else "humanoid_obst16.json"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
step = float(sys.argv[3]) if len(sys.argv) > 3 else 0.2
desc = pkg.ModelDescription.load(name)
ws = pkg.Workspace(pkg.Model(desc), 0)
o = ob.Oracle(desc)
q1 = pkg.uniform_configs(desc, n, seed=101)
d = np.random.default_rng(5).normal(size=q1.shape)
q2 = pkg.normalize_quaternions(desc, np.clip(q1 + step * d / np.linalg.norm(d, axis=1, keepdims=True), desc.q_lower, desc.q_upper))
L = o.straight_lengths(q1, q2)
e = o.validate_edges(2, 0.0, q1, q2, L)
s = e["samples"].astype(np.int64)
top = np.argsort(-s)[:6]
for i in top:
    a, b, l = q1[i:i + 1].copy(), q2[i:i + 1].copy(), L[i:i + 1].copy()
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter(); r = ws.validate_edges(2, 0.0, a, b, l); best = min(best, time.perf_counter() - t0)
    ok = r["valid"][0] == e["valid"][i] and (r["t_fail"][0] == e["t_fail"][i] or (r["t_fail"][0] != r["t_fail"][0] and e["t_fail"][i] != e["t_fail"][i]))
    print("edge %d: %d samples, valid %d, alone %.2f ms = %.1f us per sample, same result %s" % (i, s[i], e["valid"][i], best * 1e3, best * 1e6 / s[i], ok), flush=True)
