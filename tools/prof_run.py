#!/usr/bin/env python3
"""Short program for ncu captures: one configuration batch and one edge batch, device resident.
   python tools/prof_run.py [configs|prog|disc|dich] [n]"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load()
what = sys.argv[1] if len(sys.argv) > 1 else "configs"
n = int(sys.argv[2]) if len(sys.argv) > 2 else (1 << 20 if what == "configs" else 1 << 16)
desc = pkg.ModelDescription.load("ur5_boxes8.json")
model = pkg.Model(desc)
ws = pkg.Workspace(model, 0)
dev = torch.device("cuda", 0)
if what == "configs":
    q = torch.from_numpy(pkg.uniform_configs(desc, n, seed=12345)).to(dev)
    v = torch.empty(n, dtype=torch.uint8, device=dev); f = torch.empty(n, dtype=torch.int32, device=dev); r = torch.empty(n, dtype=torch.uint8, device=dev)
    for _ in range(3):
        ws.validate_configs_ptr(q.data_ptr(), n, v.data_ptr(), f.data_ptr(), r.data_ptr(), pkg.DEVICE_POINTERS)
    print("configs", n, "ms", ws.stats()["last_kernel_ms"], "valid", float(v.float().mean()))
else:
    meth, par = {"prog": (1, 1e-3), "disc": (0, 0.05), "dich": (2, 0.0)}[what]
    cfg = pkg.uniform_configs(desc, 2 * n, seed=12346)
    q1 = np.ascontiguousarray(cfg[0::2]); q2 = np.ascontiguousarray(cfg[1::2])
    L = ws.straight_lengths(q1, q2)
    a, b, l = (torch.from_numpy(x).to(dev) for x in (q1, q2, L))
    v = torch.empty(n, dtype=torch.uint8, device=dev); tf = torch.empty(n, dtype=torch.float64, device=dev); tl = torch.empty(n, dtype=torch.float64, device=dev)
    pr = torch.empty(n, dtype=torch.int32, device=dev); rs = torch.empty(n, dtype=torch.uint8, device=dev)
    for _ in range(3):
        ws.validate_edges_ptr(meth, par, False, a.data_ptr(), b.data_ptr(), l.data_ptr(), n, v.data_ptr(), tf.data_ptr(), tl.data_ptr(), pr.data_ptr(), rs.data_ptr(), pkg.DEVICE_POINTERS)
    st = ws.stats()
    print(what, n, "ms", st["last_kernel_ms"], "valid", float(v.float().mean()), st)
