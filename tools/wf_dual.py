#!/usr/bin/env python3
"""Experiment: S wavefront engines on S streams of ONE GPU, each with 1/S of the edges (separate workspaces, one host
   thread each), against one engine over the whole batch.   python tools/wf_dual.py [n] [method]"""
import sys, os, time, threading
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
meth = int(sys.argv[2]) if len(sys.argv) > 2 else 1
par = 1e-3 if meth == 1 else 0.0
desc = pkg.ModelDescription.load("ur5_boxes8.json")
model = pkg.Model(desc)
dev = torch.device("cuda", 0)
cfg = pkg.uniform_configs(desc, 2 * n, seed=12346)
q1 = np.ascontiguousarray(cfg[0::2]); q2 = np.ascontiguousarray(cfg[1::2])
ws0 = pkg.Workspace(model, 0)
L = ws0.straight_lengths(q1, q2)
a, b, l = (torch.from_numpy(x).to(dev) for x in (q1, q2, L))
nq = q1.shape[1]
def outs():
    return (torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.float64, device=dev),
            torch.empty(n, dtype=torch.float64, device=dev), torch.empty(n, dtype=torch.int32, device=dev),
            torch.empty(n, dtype=torch.uint8, device=dev))
ref = None
for S in (1, 2, 3, 4, 2, 1):
    wss = [pkg.Workspace(model, 0) for _ in range(S)]
    o = outs()
    bounds = [n * i // S for i in range(S + 1)]
    def run(i):
        lo, hi = bounds[i], bounds[i + 1]
        v, tf, tl, pr, rs = o
        wss[i].validate_edges_ptr(meth, par, False, a.data_ptr() + lo * nq * 8, b.data_ptr() + lo * nq * 8, l.data_ptr() + lo * 8, hi - lo,
                                  v.data_ptr() + lo, tf.data_ptr() + lo * 8, tl.data_ptr() + lo * 8, pr.data_ptr() + lo * 4, rs.data_ptr() + lo,
                                  pkg.DEVICE_POINTERS)
    for it in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        th = [threading.Thread(target=run, args=(i,)) for i in range(S)]
        for t in th: t.start()
        for t in th: t.join()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        print("S=%d wall %.2f ms  %.2f M edges/s" % (S, wall * 1e3, n / wall / 1e6), flush=True)
    res = [x.cpu().numpy() for x in o]
    if ref is None: ref = res
    else: print("  same bits as S=1:", all(np.array_equal(x, y, equal_nan=True) for x, y in zip(res, ref)), flush=True)
    del wss
