"""hppb200_validate_configs_noalloc (include/hppb200.h): the configuration-validation kernels with their scratch in the
library's static device array instead of cudaMalloc'ed workspace buffers.  Collected last on purpose (file name).
Checked through the C ABI against the existing entry point hppb200_validate_configs on the
same device-resident inputs, bit for bit (flags, first colliding pair, reason: integer outputs, the tolerance of every
other configuration test here), on two calls in a row with different inputs and sizes (the array keeps the previous
call's contents)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class _Cuda:
    """Plain cudart through ctypes for the test's own buffers (the entry point itself allocates nothing)."""

    def __init__(self):
        self.rt = None
        for name in ("libcudart.so", "libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so"):
            try:
                self.rt = C.CDLL(name)
                break
            except OSError:
                continue
        if self.rt is None:
            pytest.skip("libcudart not found")
        self.bufs = []

    def malloc(self, nbytes):
        p = C.c_void_p()
        assert self.rt.cudaMalloc(C.byref(p), C.c_size_t(max(nbytes, 1))) == 0
        self.bufs.append(p)
        return p.value

    def h2d(self, dptr, arr):
        assert self.rt.cudaMemcpy(C.c_void_p(dptr), C.c_void_p(arr.ctypes.data), C.c_size_t(arr.nbytes), C.c_int(1)) == 0

    def d2h(self, arr, dptr):
        assert self.rt.cudaMemcpy(C.c_void_p(arr.ctypes.data), C.c_void_p(dptr), C.c_size_t(arr.nbytes), C.c_int(2)) == 0

    def fill(self, dptr, byte, nbytes):
        assert self.rt.cudaMemset(C.c_void_p(dptr), C.c_int(byte), C.c_size_t(nbytes)) == 0

    def free_bytes(self):
        free, total = C.c_size_t(), C.c_size_t()
        assert self.rt.cudaMemGetInfo(C.byref(free), C.byref(total)) == 0
        return free.value

    def close(self):
        for p in self.bufs:
            self.rt.cudaFree(p)
        self.bufs = []


def _run(pkg, ws, cu, q, entry):
    """one call of `entry` ('noalloc' or 'reference') on device copies of q; outputs pre-filled with garbage"""
    n = q.shape[0]
    dq, dv, df, dr = cu.malloc(q.nbytes), cu.malloc(n), cu.malloc(4 * n), cu.malloc(n)
    cu.h2d(dq, q)
    for ptr, nb in ((dv, n), (df, 4 * n), (dr, n)):
        cu.fill(ptr, 0x5A, nb)
    free0 = cu.free_bytes()
    if entry == "noalloc":
        ws.validate_configs_noalloc_ptr(dq, n, dv, df, dr, pkg.DEVICE_POINTERS | pkg.CHECK_JOINT_BOUNDS)
        assert cu.free_bytes() == free0, "device memory was allocated or freed inside the no-alloc entry point"
    else:
        ws.validate_configs_ptr(dq, n, dv, df, dr, pkg.DEVICE_POINTERS | pkg.CHECK_JOINT_BOUNDS)
    v, f, r = np.zeros(n, np.uint8), np.zeros(n, np.int32), np.zeros(n, np.uint8)
    cu.d2h(v, dv); cu.d2h(f, df); cu.d2h(r, dr)
    return v, f, r


@pytest.mark.parametrize("name", ["ur5_boxes8.json", "humanoid_obst6.json"])
def test_validate_configs_noalloc_matches_existing_entry_point(pkg, name):
    desc = pkg.ModelDescription.load(name)
    model = pkg.Model(desc)
    ws = pkg.Workspace(model, 0)
    cu = _Cuda()
    try:
        big = 60000 if name.startswith("ur5") else 12000
        q1 = np.ascontiguousarray(pkg.uniform_configs(desc, big, seed=501) * 1.05)   # some joint-bound failures too
        q2 = np.ascontiguousarray(pkg.uniform_configs(desc, big // 3 + 7, seed=502))  # smaller, ragged: stale scratch behind it
        ref1, ref2 = _run(pkg, ws, cu, q1, "reference"), _run(pkg, ws, cu, q2, "reference")
        # two calls in a row on the shared static array
        got1 = _run(pkg, ws, cu, q1, "noalloc")
        got2 = _run(pkg, ws, cu, q2, "noalloc")
        for got, ref in ((got1, ref1), (got2, ref2)):
            for a, b in zip(got, ref):
                assert np.array_equal(a, b)
        assert 0.02 < got2[0].mean() < 0.98 and (got1[2] == pkg.REASON_JOINT_BOUND).any() and (got1[2] == pkg.REASON_COLLISION).any()
        # and again, same inputs: same bits
        for a, b in zip(_run(pkg, ws, cu, q1, "noalloc"), ref1):
            assert np.array_equal(a, b)
    finally:
        cu.close()


def test_validate_configs_noalloc_limits(pkg):
    desc = pkg.ModelDescription.load("ur5_boxes8.json")
    ws = pkg.Workspace(pkg.Model(desc), 0)
    cu = _Cuda()
    try:
        n = 200000  # 200000 * (8 * 96 + 16) B > 64 MiB
        dq, dv, df, dr = cu.malloc(n * desc.nq * 8), cu.malloc(n), cu.malloc(4 * n), cu.malloc(n)
        with pytest.raises(ValueError, match="static array"):
            ws.validate_configs_noalloc_ptr(dq, n, dv, df, dr)
        with pytest.raises(ValueError, match="device pointers"):
            ws.validate_configs_noalloc_ptr(dq, 10, dv, df, dr, flags=0)
        with pytest.raises(ValueError):
            ws.validate_configs_noalloc_ptr(dq, 10, dv, 0, dr)
        ws.validate_configs_noalloc_ptr(dq, 0, dv, df, dr)  # empty batch: nothing to do
    finally:
        cu.close()
