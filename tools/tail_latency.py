#!/usr/bin/env python3
"""Per-sample latency of the CTA-per-edge Dichotomy kernel (kernels_tail.cu): the edges with the most samples of a
   random batch, validated one at a time, and the rate of a batch that fills the GPU.
   python tools/tail_latency.py MODEL.json N [local_step]"""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["HPPB200_WF_MIN"] = "1000000000"
os.environ.setdefault("HPPB200_TAIL_MAX", "1000000000")
import torch
import __graft_entry__ as ge
pkg = ge.load()
from oracle import binding as ob
name, n = sys.argv[1], int(sys.argv[2])
step = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
desc = pkg.ModelDescription.load(name)
ws = pkg.Workspace(pkg.Model(desc), 0)
o = ob.Oracle(desc)
q1 = pkg.uniform_configs(desc, n, seed=101)
if step > 0:
    d = np.random.default_rng(5).normal(size=q1.shape)
    q2 = pkg.normalize_quaternions(desc, np.clip(q1 + step * d / np.linalg.norm(d, axis=1, keepdims=True), desc.q_lower, desc.q_upper))
else:
    q2 = pkg.uniform_configs(desc, n, seed=102)
L = ws.straight_lengths(q1, q2)
e = o.validate_edges(2, 0.0, q1, q2, L)
smp = e["samples"].astype(np.int64)
top = np.argsort(-smp)[:6]
dev = torch.device("cuda", 0)
def run(idx):
    a, b, l = (torch.from_numpy(np.ascontiguousarray(x[idx])).to(dev) for x in (q1, q2, L))
    m = len(idx)
    v = torch.empty(m, dtype=torch.uint8, device=dev); tf = torch.empty(m, dtype=torch.float64, device=dev); tl = torch.empty(m, dtype=torch.float64, device=dev)
    pr = torch.empty(m, dtype=torch.int32, device=dev); rs = torch.empty(m, dtype=torch.uint8, device=dev)
    best = 1e9
    for it in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ws.validate_edges_ptr(2, 0.0, False, a.data_ptr(), b.data_ptr(), l.data_ptr(), m, v.data_ptr(), tf.data_ptr(), tl.data_ptr(), pr.data_ptr(), rs.data_ptr(), pkg.DEVICE_POINTERS)
        torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    ok = np.array_equal(v.cpu().numpy(), e["valid"][idx]) and np.array_equal(tf.cpu().numpy(), e["t_fail"][idx], equal_nan=True)
    return best, ok
for i in top:
    t, ok = run(np.array([i]))
    print("edge %d: %d samples, %.2f ms alone = %.1f us per sample, parity %s" % (i, smp[i], t * 1e3, t * 1e6 / smp[i], ok), flush=True)
t, ok = run(np.arange(n))
print("%s: batch of %d edges (%.1f samples/edge, max %d): %.2f ms = %.3f M edges/s, %.1f M samples/s, parity %s" % (
    name, n, smp.mean(), smp.max(), t * 1e3, n / t / 1e6, smp.sum() / t / 1e6, ok), flush=True)
