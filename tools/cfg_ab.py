#!/usr/bin/env python3
"""configPassKernel A/B: HPPB200_CFG_SHADOW=1 (32-bit shadow cores, default) / 0 (FP64 cores) / 2 (shadow kernel, every pair exact),
   device-resident timing and bit comparison with the oracle.   python tools/cfg_ab.py [n] [model.json]"""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load()
from oracle import binding as ob
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
name = sys.argv[2] if len(sys.argv) > 2 else "ur5_boxes8.json"
desc = pkg.ModelDescription.load(name)
model = pkg.Model(desc)
dev = torch.device("cuda", 0)
q = pkg.uniform_configs(desc, n, seed=12345)
# a few hostile rows: NaN, infinities, huge angles
q[5, 0] = np.nan; q[6, :] = np.nan; q[7, 1] = np.inf; q[8, 2] = 9.9e4; q[9, 0] = -np.inf
orc = ob.Oracle(desc)
t0 = time.perf_counter(); ov, op, orr = orc.validate_configs(q); print("oracle %.2f s" % (time.perf_counter() - t0), flush=True)
qd = torch.from_numpy(q).to(dev)
for mode in ("1", "0", "2", "1"):
    os.environ["HPPB200_CFG_SHADOW"] = mode
    ws = pkg.Workspace(model, 0)
    v = torch.empty(n, dtype=torch.uint8, device=dev); f = torch.empty(n, dtype=torch.int32, device=dev); r = torch.empty(n, dtype=torch.uint8, device=dev)
    ts = []
    for it in range(6):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        ws.set_stream(torch.cuda.current_stream().cuda_stream)
        e0.record()
        ws.validate_configs_ptr(qd.data_ptr(), n, v.data_ptr(), f.data_ptr(), r.data_ptr(), pkg.DEVICE_POINTERS)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    same = np.array_equal(v.cpu().numpy(), ov) and np.array_equal(f.cpu().numpy(), op) and np.array_equal(r.cpu().numpy(), orr)
    st = ws.stats()
    print("shadow=%s  ms %s  best %.3f  %.1f M configs/s  == oracle: %s  valid %.4f  jobs %s" % (
        mode, " ".join("%.3f" % t for t in ts), min(ts), n / min(ts) / 1e3, same, float(v.float().mean()), st.get("last_jobs")), flush=True)
    if not same:
        vv, ff = v.cpu().numpy(), f.cpu().numpy()
        bad = np.nonzero((vv != ov) | (ff != op))[0]
        print("   differing:", len(bad), bad[:10], vv[bad[:10]], ov[bad[:10]], ff[bad[:10]], op[bad[:10]])
    del ws
