#!/usr/bin/env python3
"""Wavefront-engine experiment: parity against the oracle on a sample, then device timing.
   python tools/wf_run.py [n_time] [n_parity]   (HPPB200_ENGINE / HPPB200_WF_SLOTS / HPPB200_ENGINE_ALPHA from env)"""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load()
from oracle import binding as ob
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 17
npar = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
meth = int(sys.argv[3]) if len(sys.argv) > 3 else 1   # 1 Progressive(1e-3), 2 Dichotomy(0)
par = 1e-3 if meth == 1 else 0.0
desc = pkg.ModelDescription.load("ur5_boxes8.json")
model = pkg.Model(desc)
ws = pkg.Workspace(model, 0)
dev = torch.device("cuda", 0)
cfg = pkg.uniform_configs(desc, 2 * n, seed=12346)
q1 = np.ascontiguousarray(cfg[0::2]); q2 = np.ascontiguousarray(cfg[1::2])
L = ws.straight_lengths(q1, q2)
if npar:
    o = ob.Oracle(desc)
    for rev in (False, True):
        r = ws.validate_edges(meth, par, q1[:npar], q2[:npar], L[:npar], reverse=rev)
        e = o.validate_edges(meth, par, q1[:npar], q2[:npar], L[:npar], reverse=rev)
        ok = all(np.array_equal(r[k], e[k], equal_nan=True) for k in ("valid", "reason", "t_fail", "t_last_valid", "pair"))
        print("parity reverse=%d: %s (valid %.3f)" % (rev, ok, r["valid"].mean()), flush=True)
        if not ok:
            for k in ("valid", "reason", "t_fail", "t_last_valid", "pair"):
                bad = np.nonzero(~((r[k] == e[k]) | ((r[k] != r[k]) & (e[k] != e[k]))))[0]
                print("  ", k, len(bad), bad[:5], r[k][bad[:5]], e[k][bad[:5]])
a, b, l = (torch.from_numpy(x).to(dev) for x in (q1, q2, L))
v = torch.empty(n, dtype=torch.uint8, device=dev); tf = torch.empty(n, dtype=torch.float64, device=dev); tl = torch.empty(n, dtype=torch.float64, device=dev)
pr = torch.empty(n, dtype=torch.int32, device=dev); rs = torch.empty(n, dtype=torch.uint8, device=dev)
for it in range(3):
    s0 = ws.stats()
    t0 = time.perf_counter()
    ws.validate_edges_ptr(meth, par, False, a.data_ptr(), b.data_ptr(), l.data_ptr(), n, v.data_ptr(), tf.data_ptr(), tl.data_ptr(), pr.data_ptr(), rs.data_ptr(), pkg.DEVICE_POINTERS)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    st = ws.stats()
    d = {k: st[k] - s0[k] for k in st if isinstance(st[k], int) and k in s0}
    print("n=%d wall %.2f ms  kernel_ms %.2f  %.2f M edges/s  valid %.3f  %s" % (n, wall * 1e3, st["last_kernel_ms"], n / wall / 1e6, float(v.float().mean()), d), flush=True)
