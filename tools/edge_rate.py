#!/usr/bin/env python3
"""Device edges/s (inputs resident in HBM) next to the oracle on all host threads, any model / method.
   python tools/edge_rate.py MODEL.json METHOD(0|1|2) N [local_step]"""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load()
from oracle import binding as ob
name, meth, n = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
step = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
par = {0: 0.05, 1: 1e-3, 2: 0.0}[meth]
desc = pkg.ModelDescription.load(name)
ws = pkg.Workspace(pkg.Model(desc), 0)
o = ob.Oracle(desc)
q1 = pkg.uniform_configs(desc, n, seed=101)
if os.environ.get("EDGE_RATE_VALID_START"):  # edges that start at VALID configurations (what a planner extends from)
    pool = pkg.uniform_configs(desc, 20 * n, seed=103)
    ok = np.flatnonzero(ws.validate_configs(pool)[0])
    q1 = np.ascontiguousarray(pool[ok[np.arange(n) % len(ok)]])
if step > 0:
    d = np.random.default_rng(5).normal(size=q1.shape)
    q2 = pkg.normalize_quaternions(desc, np.clip(q1 + step * d / np.linalg.norm(d, axis=1, keepdims=True), desc.q_lower, desc.q_upper))
else:
    q2 = pkg.uniform_configs(desc, n, seed=102)
L = ws.straight_lengths(q1, q2)
dev = torch.device("cuda", 0)
a, b, l = (torch.from_numpy(x).to(dev) for x in (q1, q2, L))
v = torch.empty(n, dtype=torch.uint8, device=dev); tf = torch.empty(n, dtype=torch.float64, device=dev); tl = torch.empty(n, dtype=torch.float64, device=dev)
pr = torch.empty(n, dtype=torch.int32, device=dev); rs = torch.empty(n, dtype=torch.uint8, device=dev)
best = 1e9
for it in range(3):
    s0 = ws.stats(); torch.cuda.synchronize(); t0 = time.perf_counter()
    ws.validate_edges_ptr(meth, par, False, a.data_ptr(), b.data_ptr(), l.data_ptr(), n, v.data_ptr(), tf.data_ptr(), tl.data_ptr(), pr.data_ptr(), rs.data_ptr(), pkg.DEVICE_POINTERS)
    torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0); s1 = ws.stats()
m = min(n, 20000)
t0 = time.perf_counter(); e = o.validate_edges(meth, par, q1[:m], q2[:m], L[:m]); dt = time.perf_counter() - t0
ok = np.array_equal(v.cpu().numpy()[:m], e["valid"]) and np.array_equal(tf.cpu().numpy()[:m], e["t_fail"], equal_nan=True)
print("%s method %d n %d step %g: device %.3f M edges/s (%.1f ms), oracle %.3f M edges/s on %d threads (%d edges), valid %.3f, samples/edge %.1f, pair tests/edge %.1f, parity %s" % (
    name, meth, n, step, n / best / 1e6, best * 1e3, m / dt / 1e6, os.cpu_count(), m, float(v.float().mean()),
    (s1["edge_samples"] - s0["edge_samples"]) / n, (s1["pair_tests"] - s0["pair_tests"]) / n, ok), flush=True)
