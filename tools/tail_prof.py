#!/usr/bin/env python3
"""One launch of the CTA-per-edge Dichotomy kernel on n random UR5 edges (for ncu)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["HPPB200_WF_MIN"] = "1000000000"
os.environ["HPPB200_TAIL_MAX"] = "1000000000"
import numpy as np
import __graft_entry__ as ge
pkg = ge.load()
name = sys.argv[1] if len(sys.argv) > 1 else "ur5_boxes8.json"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
desc = pkg.ModelDescription.load(name)
ws = pkg.Workspace(pkg.Model(desc), 0)
q1, q2 = pkg.uniform_configs(desc, n, seed=101), pkg.uniform_configs(desc, n, seed=102)
L = ws.straight_lengths(q1, q2)
r = ws.validate_edges(2, 0.0, q1, q2, L)
print("valid", r["valid"].mean())
