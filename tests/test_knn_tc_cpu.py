"""Host-side checks of the tensor-core k-NN candidate pass (hpp-core_b200/csrc/kernels_knn_tc.cu): the TF32
three-term split that turns the squared WeighedDistance (reference src/weighed-distance.cc:272-289) into ONE dot
product, emulated in numpy, against the error bound the thresholds are inflated by; and the operand tile layout."""
import numpy as np


def tf32(x):
    """round-to-nearest (ties away) to 10 explicit mantissa bits, as cvt.rna.tf32.f32"""
    u = np.asarray(x, np.float32).view(np.uint32).astype(np.uint64)
    u = ((u + 0x1000) & 0xFFFFE000).astype(np.uint32)
    return u.view(np.float32)


def operand_rows(x, side):
    """x [n][nq] already scaled by the weights (FP64) -> [n][3 nq + 4] TF32-representable FP32 slots"""
    h = tf32(x)
    l = tf32(x - h.astype(np.float64))
    norm = (x * x).sum(-1)
    nh = tf32(norm)
    nl = tf32(norm - nh.astype(np.float64))
    one = np.ones_like(nh)
    if side == 0:
        return np.concatenate([h, h, l, nh[:, None], nl[:, None], one[:, None], one[:, None]], axis=1)
    return np.concatenate([-2 * h, -2 * l, -2 * h, one[:, None], one[:, None], nh[:, None], nl[:, None]], axis=1)


def test_tf32_split_dot_is_the_squared_distance_within_the_bound():
    rng = np.random.default_rng(1)
    w = np.array([0.82, 0.83, 0.44, 0.15, 0.14, 0.13])
    a = rng.uniform(-np.pi, np.pi, (200, 6)) * w
    b = rng.uniform(-np.pi, np.pi, (300, 6)) * w
    A, B = operand_rows(a, 0), operand_rows(b, 1)
    assert np.array_equal(A, tf32(A)) and np.array_equal(B, tf32(B))  # every slot is exactly a TF32 value
    approx = A.astype(np.float64) @ B.astype(np.float64).T  # products of TF32 values are exact in FP32 / FP64
    exact = ((a[:, None, :] - b[None, :, :]) ** 2).sum(-1)
    eps = ((a * a).sum(-1)[:, None] + (b * b).sum(-1).max()) * 2.0 ** -16
    # the split alone (no accumulation error) is two orders of magnitude inside the bound
    assert np.all(np.abs(approx - exact) < eps / 64)
    # FP32 accumulation in any order stays inside it too
    acc = np.zeros((200, 300), np.float32)
    for s in rng.permutation(A.shape[1]):
        acc = (acc + A[:, s:s + 1] * B[None, :, s].astype(np.float32)).astype(np.float32)
    assert np.all(np.abs(acc - exact) < eps / 4)


def test_operand_tile_layout_is_the_umma_k_major_canonical_layout():
    """tcIndex: 8-row x 16-byte core matrices, K halves 128 B apart, 8-row groups kc * 256 B apart, no overlap."""
    def tc_index(r, s, kc):
        return (r >> 3) * (kc * 64) + (s >> 2) * 32 + (r & 7) * 4 + (s & 3)
    for kc, rows in ((3, 256), (3, 128), (6, 256)):
        idx = np.array([[tc_index(r, s, kc) for s in range(8 * kc)] for r in range(rows)])
        assert sorted(idx.ravel()) == list(range(rows * 8 * kc))  # a bijection onto the tile
        assert tc_index(0, 4, kc) * 4 == 128 and tc_index(8, 0, kc) * 4 == kc * 256 and tc_index(1, 0, kc) * 4 == 16
