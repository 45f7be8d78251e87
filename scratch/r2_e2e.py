"""e2e configs/s through the host-pointer C ABI (pinned buffers) vs slices; parity with the device path."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load()
desc = pkg.ModelDescription.load("ur5_boxes8.json")
model = pkg.Model(desc)
n = 1 << 20
q_host = torch.from_numpy(pkg.uniform_configs(desc, n, seed=12345)).pin_memory()
dev = torch.device("cuda", 0)
q_dev = q_host.to(dev)
vd = torch.empty(n, dtype=torch.uint8, device=dev); fd = torch.empty(n, dtype=torch.int32, device=dev); rd = torch.empty(n, dtype=torch.uint8, device=dev)
vh = torch.empty(n, dtype=torch.uint8).pin_memory(); fh = torch.empty(n, dtype=torch.int32).pin_memory(); rh = torch.empty(n, dtype=torch.uint8).pin_memory()
for sl in sys.argv[1:] or ["1", "4", "8", "16"]:
    os.environ["HPPB200_CFG_SLICES"] = sl
    ws = pkg.Workspace(model, 0)
    ws.validate_configs_ptr(q_dev.data_ptr(), n, vd.data_ptr(), fd.data_ptr(), rd.data_ptr(), pkg.DEVICE_POINTERS)
    for _ in range(3):
        ws.validate_configs_ptr(q_host.data_ptr(), n, vh.data_ptr(), fh.data_ptr(), rh.data_ptr(), 0)
    ok = torch.equal(vh, vd.cpu()) and torch.equal(fh, fd.cpu()) and torch.equal(rh, rd.cpu())
    ts = []
    for _ in range(10):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ws.validate_configs_ptr(q_host.data_ptr(), n, vh.data_ptr(), fh.data_ptr(), rh.data_ptr(), 0)
        ts.append(time.perf_counter() - t0)
    tm = sorted(ts)[len(ts) // 2]
    print("slices %s: e2e %.3f ms = %.0f M configs/s (last_kernel_ms %.3f) identical to device path: %s" % (sl, tm * 1e3, n / tm / 1e6, ws.stats()["last_kernel_ms"], ok), flush=True)
# ragged / small sizes through the host path vs device path
ws = pkg.Workspace(model, 0)
for m in (1, 127, 32768, 32769, 100003, 300001):
    a = ws.validate_configs(q_host.numpy()[:m])
    ws.validate_configs_ptr(q_dev.data_ptr(), m, vd.data_ptr(), fd.data_ptr(), rd.data_ptr(), pkg.DEVICE_POINTERS)
    okm = np.array_equal(a[0], vd.cpu().numpy()[:m]) and np.array_equal(a[1], fd.cpu().numpy()[:m]) and np.array_equal(a[2], rd.cpu().numpy()[:m])
    print("n=%d host==device: %s" % (m, okm))
