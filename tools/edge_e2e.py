import sys, os, time
import numpy as np
sys.path.insert(0, os.getcwd())
import torch
import __graft_entry__ as ge
pkg = ge.load()
desc = pkg.ModelDescription.load("humanoid_obst16.json")
ws = pkg.Workspace(pkg.Model(desc), 0)
n = 1 << 20
q1 = pkg.uniform_configs(desc, n, seed=101)
d = np.random.default_rng(5).normal(size=q1.shape)
q2 = pkg.normalize_quaternions(desc, np.clip(q1 + 0.2 * d / np.linalg.norm(d, axis=1, keepdims=True), desc.q_lower, desc.q_upper))
L = ws.straight_lengths(q1, q2)
a, b, l = (torch.from_numpy(np.ascontiguousarray(x)).pin_memory() for x in (q1, q2, L))
v = torch.empty(n, dtype=torch.uint8).pin_memory(); tf = torch.empty(n, dtype=torch.float64).pin_memory(); tl = torch.empty(n, dtype=torch.float64).pin_memory()
pr = torch.empty(n, dtype=torch.int32).pin_memory(); rs = torch.empty(n, dtype=torch.uint8).pin_memory()
for meth, par in ((2, 0.0), (1, 1e-3)):
    best = 1e9
    for it in range(5):
        t0 = time.perf_counter()
        ws.validate_edges_ptr(meth, par, False, a.data_ptr(), b.data_ptr(), l.data_ptr(), n, v.data_ptr(), tf.data_ptr(), tl.data_ptr(), pr.data_ptr(), rs.data_ptr(), 0)
        best = min(best, time.perf_counter() - t0)
    print("method %d host buffers: %.1f ms = %.2f M edges/s (valid %.3f)" % (meth, best * 1e3, n / best / 1e6, float(v.float().mean())), flush=True)
