#!/usr/bin/env python3
"""ncu target: one Dichotomy call on the 30-joint tree model (BASELINE configs[3] shape: local edges), all of it in edgeTailKernel.
   python tools/tail_prof.py [n]"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
desc = pkg.ModelDescription.load("humanoid_obst16.json")
ws = pkg.Workspace(pkg.Model(desc), 0)
q1 = pkg.uniform_configs(desc, n, seed=101)
d = np.random.default_rng(5).normal(size=q1.shape)
q2 = pkg.normalize_quaternions(desc, np.clip(q1 + 0.2 * d / np.linalg.norm(d, axis=1, keepdims=True), desc.q_lower, desc.q_upper))
L = ws.straight_lengths(q1, q2)
dev = torch.device("cuda", 0)
a, b, l = (torch.from_numpy(x).to(dev) for x in (q1, q2, L))
v = torch.empty(n, dtype=torch.uint8, device=dev); tf = torch.empty(n, dtype=torch.float64, device=dev); tl = torch.empty(n, dtype=torch.float64, device=dev)
pr = torch.empty(n, dtype=torch.int32, device=dev); rs = torch.empty(n, dtype=torch.uint8, device=dev)
for _ in range(2):
    ws.validate_edges_ptr(2, 0.0, False, a.data_ptr(), b.data_ptr(), l.data_ptr(), n, v.data_ptr(), tf.data_ptr(), tl.data_ptr(), pr.data_ptr(), rs.data_ptr(), pkg.DEVICE_POINTERS)
print("done", ws.stats()["last_kernel_ms"], float(v.float().mean()))
