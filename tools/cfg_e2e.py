#!/usr/bin/env python3
"""Host-pointer hppb200_validate_configs (pinned buffers, H2D + kernels + D2H inside): slice schedule sweep.
   python tools/cfg_e2e.py [n]"""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
desc = pkg.ModelDescription.load("ur5_boxes8.json")
model = pkg.Model(desc)
ws = pkg.Workspace(model, 0)
q = torch.from_numpy(pkg.uniform_configs(desc, n, seed=12345)).pin_memory()
v = torch.empty(n, dtype=torch.uint8).pin_memory(); f = torch.empty(n, dtype=torch.int32).pin_memory(); r = torch.empty(n, dtype=torch.uint8).pin_memory()
ref = None
for zc, sw, tw in ((0, 1, 2), (0, 1, 2), (0, 2, 2), (0, 4, 2), (1, 4, 0), (0, 1, 2)):
    os.environ["HPPB200_CFG_ZEROCOPY"] = str(zc); os.environ["HPPB200_CFG_SLICE_WAVES"] = str(sw); os.environ["HPPB200_CFG_TAIL_WAVES"] = str(tw)
    ts = []
    for it in range(12):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ws.validate_configs_ptr(q.data_ptr(), n, v.data_ptr(), f.data_ptr(), r.data_ptr(), 0)
        ts.append((time.perf_counter() - t0) * 1e3)
    ts = sorted(ts[2:])
    out = (v.numpy().copy(), f.numpy().copy(), r.numpy().copy())
    if ref is None: ref = out
    same = all(np.array_equal(a, b) for a, b in zip(out, ref))
    print("zero_copy %d slice_waves %d tail_waves %d: median %.3f ms best %.3f ms  %.0f M configs/s  same=%s" % (zc, sw, tw, ts[len(ts) // 2], ts[0], n / ts[len(ts) // 2] / 1e3, same), flush=True)
os.environ["HPPB200_CFG_TRACE"] = "1"
ws.validate_configs_ptr(q.data_ptr(), n, v.data_ptr(), f.data_ptr(), r.data_ptr(), 0)
