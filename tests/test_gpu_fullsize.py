"""Parity at BASELINE.json's full sizes, through size-independent properties (the oracle would
take minutes to hours there): sampled entries against the f64/integer truth on the host,
involution of transpose, checksum-of-checksums for integer elementwise ops, Freivalds' check for
the integer GEMM (exact mod 2^32) and sampled f64 dot products for the f32 GEMM."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def la():
    import linear_algebra_b200 as la
    la.init(0)
    return la


def test_elementwise_256M_f32(la):             # configs[1]
    n = 1 << 28
    rng = np.random.default_rng(100)
    a = rng.random(n, dtype=np.float32); b = rng.random(n, dtype=np.float32) + np.float32(1.0)
    A = la.Vector.from_vec(a); B = la.Vector.from_vec(b)
    idx = rng.integers(0, n, 1 << 20); idx[:4] = (0, 1, n - 2, n - 1)
    for got, expect in ((A + B, a[idx] + b[idx]), (A - B, a[idx] - b[idx]), (A * B, a[idx] * b[idx]), (A / B, a[idx] / b[idx]),
                        (A * np.float32(1.5), a[idx] * np.float32(1.5))):
        h = got.to_vec()
        assert h[idx].tobytes() == expect.tobytes()
        del got
    C = A.clone(); C += B; C -= B      # not exact in floating point, but close; checks the in-place forms run at size
    h = C.to_vec()
    assert np.max(np.abs(h[idx] - a[idx])) <= 2.5e-7
    # dot/sum at 2^28 against f64 (pairwise-style accumulation keeps f32 error tiny)
    d = float(la.dot(A, B)); t = float(np.dot(a.astype(np.float64), b.astype(np.float64)))
    assert abs(d - t) <= 2e-6 * t
    s = float(A.sum()); t = float(a.sum(dtype=np.float64))
    assert abs(s - t) <= 2e-6 * t


def test_integer_8192_square(la):              # configs[4]: i32 elementwise + GEMM + transpose at 8192^2
    n = 8192
    rng = np.random.default_rng(101)
    a = rng.integers(-2**31, 2**31 - 1, n * n, dtype=np.int32, endpoint=True)
    b = rng.integers(-2**31, 2**31 - 1, n * n, dtype=np.int32, endpoint=True)
    A = la.Matrix.from_vec(a, n, n); B = la.Matrix.from_vec(b, n, n)
    au = a.view(np.uint32).astype(np.uint64); bu = b.view(np.uint32).astype(np.uint64)
    M32 = np.uint64(0xFFFFFFFF)
    # elementwise: checksum of checksums, exact mod 2^32
    sa = int(au.sum() & M32); sb = int(bu.sum() & M32)
    assert int(np.uint32((A + B).data.sum())) == (sa + sb) & 0xFFFFFFFF
    assert int(np.uint32((A - B).data.sum())) == (sa - sb) & 0xFFFFFFFF
    assert int(np.uint32((A.data * B.data).sum())) == int(np.uint32(la.dot(A.data, B.data))) == int((au * bu & M32).sum() & M32)
    # transpose: involution + sampled definition
    T = A.transpose()
    assert T.transpose() == A
    t = T.to_vec()
    idx = rng.integers(0, n, (4096, 2))
    assert np.array_equal(t[idx[:, 1] * n + idx[:, 0]], a[idx[:, 0] * n + idx[:, 1]])
    del T, t
    # GEMM: Freivalds, exact mod 2^32: C x == A (B x) for random x (device GEMV cross-checks the GEMM) ...
    C = A * B
    for seed in range(32):                      # 32 random vectors, exact mod 2^32
        x = la.Vector.from_vec(np.random.default_rng(seed).integers(-2**31, 2**31 - 1, n, dtype=np.int32))
        assert la.mul_transpose_mat(x, C) == la.mul_transpose_mat(la.mul_transpose_mat(x, B), A)
    # ... and 64 sampled entries against the host
    c = C.to_vec()
    A2 = a.reshape(n, n); B2 = b.reshape(n, n)
    for r, col in rng.integers(0, n, (64, 2)):
        expect = int((A2[r].view(np.uint32).astype(np.uint64) * B2[:, col].view(np.uint32).astype(np.uint64) & M32).sum() & M32)
        assert int(np.uint32(c[r * n + col])) == expect


@pytest.mark.parametrize("mode", ["simt", "3xtf32"])
def test_f32_gemm_8192(la, mode):              # the headline config, on both f32 engines
    n = 8192
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_SIMT if mode == "simt" else la._lib.GEMM_3XTF32)
    rng = np.random.default_rng(102)
    a = rng.random(n * n, dtype=np.float32) * np.float32(2) - np.float32(1); b = rng.random(n * n, dtype=np.float32) * np.float32(2) - np.float32(1)
    try:
        C = (la.Matrix.from_vec(a, n, n) * la.Matrix.from_vec(b, n, n)).to_vec()
        assert la._lib.lib.la_gemm_last_engine() == (1 if mode == "simt" else 2)
    finally:
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)
    A2 = a.reshape(n, n).astype(np.float64); B2 = b.reshape(n, n)
    worst = 0.0
    for r, col in rng.integers(0, n, (256, 2)):
        bc = B2[:, col].astype(np.float64)
        truth = float(A2[r] @ bc); scale = float(np.abs(A2[r]) @ np.abs(bc))
        # reference order error bound for a serial f32 sum of K terms is ~K * 2^-24 * scale; ours must be well inside
        worst = max(worst, abs(float(C[r * n + col]) - truth) / scale)
    assert worst <= 64 * 2.0 ** -24, worst
    # two full rows, normwise
    for r in (0, n - 1):
        truth = A2[r] @ B2.astype(np.float64)
        assert np.linalg.norm(C[r * n:(r + 1) * n] - truth) / np.linalg.norm(truth) <= 1e-5
    # ... and through the gate of tests/test_gpu_parity.py at factor 1, at the headline size: four full rows (32768 outputs)
    # against the reference's own k-ascending order (oracle, both contraction modes) on the same rows
    import oracle
    P = oracle.port()
    rows = [0, 1, n // 2, n - 1]
    ar = np.ascontiguousarray(a.reshape(n, n)[rows]).reshape(-1)
    truth = (A2[rows] @ B2.astype(np.float64)).reshape(-1)
    scale = (np.abs(A2[rows]) @ np.abs(B2.astype(np.float64))).reshape(-1)
    e_ref = max(float((np.abs(P.mul_mat_mat(ar, b, len(rows), n, n, fma=f).astype(np.float64) - truth) / scale).max()) for f in (False, True))
    ours = np.concatenate([C[r * n:(r + 1) * n] for r in rows]).astype(np.float64)
    e_ours = float((np.abs(ours - truth) / scale).max())
    assert e_ours <= max(e_ref, 2.0 ** -22), (mode, e_ours, e_ref)


def test_gemm_16384_banded_tile_order(la):     # C3's largest size; above M = 8192 the tensor engines walk the tiles in bands
    n = 16384
    A = la.Matrix(la.Vector.synth(0x5EED0003, n * n, "f32"), n, n); B = la.Matrix(la.Vector.synth(0x5EED0004, n * n, "f32"), n, n)
    C = A * B
    assert la._lib.lib.la_gemm_last_engine() == 2
    rng = np.random.default_rng(104)
    Bt = B.transpose()                          # rows of Bt = columns of B, so a sampled entry needs two contiguous downloads
    worst = 0.0
    for r, col in rng.integers(0, n, (48, 2)):
        ar = A.data.slice_to_vec(int(r) * n, n).astype(np.float64); bc = Bt.data.slice_to_vec(int(col) * n, n).astype(np.float64)
        got = float(C.data.slice_to_vec(int(r) * n + int(col), 1)[0])
        worst = max(worst, abs(got - float(ar @ bc)) / float(np.abs(ar) @ np.abs(bc)))
    assert worst <= 64 * 2.0 ** -24, worst
    del Bt
    # every entry at once: C x = A (B x) with the library's own M.v (verified independently), x a +-1 vector, normwise
    x = la.Vector.from_vec((rng.integers(0, 2, n) * 2 - 1).astype(np.float32))
    lhs = la.mul_transpose_mat(x, C).to_vec().astype(np.float64)
    rhs = la.mul_transpose_mat(la.mul_transpose_mat(x, B), A).to_vec().astype(np.float64)
    assert np.linalg.norm(lhs - rhs) / np.linalg.norm(rhs) <= 1e-5
    # the same product in m-major order is bit-identical (with the K-split of the last wave off: which tiles it splits
    # depends on the order)
    la._lib.lib.la_gemm_set_stream_tail(0)
    try:
        banded = A * B
        la._lib.lib.la_gemm_set_raster(0)
        identical = ((A * B) == banded)
        assert identical
    finally:
        la._lib.lib.la_gemm_set_raster(-1); la._lib.lib.la_gemm_set_stream_tail(-1)


def test_matvec_32768(la):                     # configs[2], one GPU
    n = 32768
    rng = np.random.default_rng(103)
    m = rng.random(n * n, dtype=np.float32) * np.float32(2) - np.float32(1)
    v = rng.random(n, dtype=np.float32) * np.float32(2) - np.float32(1)
    M = la.Matrix.from_vec(m, n, n); V = la.Vector.from_vec(v)
    M2 = m.reshape(n, n)
    y = la.mul_transpose_mat(V, M).to_vec()             # M x v
    rows = rng.integers(0, n, 64)
    for r in rows:
        truth = float(M2[r].astype(np.float64) @ v.astype(np.float64)); scale = float(np.abs(M2[r]).astype(np.float64) @ np.abs(v).astype(np.float64))
        assert abs(float(y[r]) - truth) <= 2.0 ** -20 * scale
    z = (V * M).to_vec()                                 # v x M
    for c in rng.integers(0, n, 64):
        col = M2[:, c].astype(np.float64)
        truth = float(col @ v.astype(np.float64)); scale = float(np.abs(col) @ np.abs(v).astype(np.float64))
        assert abs(float(z[c]) - truth) <= 2.0 ** -20 * scale


def test_more_than_2_pow_32_elements(la):
    """Sizes the reference cannot even express (it passes counts as i32, vector.rs:277): 2^32 + 7 one-byte
    elements.  Operands are generated on the device (la_synth) and spot-checked against the host mirror of
    the generator, so nothing of that size crosses PCIe.  Exercises 64-bit indexing, the ragged tail and
    the wrap-around identities of sum/dot."""
    from linear_algebra_b200 import synth
    n = (1 << 32) + 7
    A = la.Vector.synth(0xA, n, "u8"); B = la.Vector.synth(0xB, n, "u8")
    C = A + B
    D = A * B
    spots = [0, (1 << 31) - 64, (1 << 31), (1 << 32) - 64, n - 128]
    for s in spots:
        a = synth.values("u8", 0xA, s, 128 if s + 128 <= n else n - s); b = synth.values("u8", 0xB, s, a.size)
        assert A.slice_to_vec(s, a.size).tobytes() == a.tobytes()
        with np.errstate(over="ignore"):
            assert C.slice_to_vec(s, a.size).tobytes() == (a + b).astype(np.uint8).tobytes()
            assert D.slice_to_vec(s, a.size).tobytes() == (a * b).astype(np.uint8).tobytes()
    # reductions: exact mod 2^8 whatever the order -> checksum identities across independent kernels
    sa, sb = int(A.sum()), int(B.sum())
    assert int(C.sum()) == (sa + sb) & 0xFF
    assert int(D.sum()) == int(la.dot(A, B))
    # and one independent anchor: the sum of a 2^24-element slice against the host mirror
    m = 1 << 24
    part = la.Vector(A.dtype, A.ptr + ((1 << 32) - m + 7), m, owner=False)       # the last 2^24 elements, unaligned start
    assert int(part.sum()) == int(synth.values("u8", 0xA, (1 << 32) - m + 7, m).sum(dtype=np.uint64) & 0xFF)
    assert A == A and not (A == B)
    del C, D
    # transpose of a 65536 x 65543 byte matrix (> 2^32 elements, ragged columns -> element kernel)
    R, Cc = 65536, 65543
    X = la.Matrix(la.Vector.synth(0xC, R * Cc, "u8"), R, Cc)
    T = X.transpose()
    for r, c in ((0, 0), (65535, 65542), (40000, 123), (7, 65000)):
        assert T.data.slice_to_vec(c * R + r, 1)[0] == synth.values("u8", 0xC, r * Cc + c, 1)[0]
    del X, T
    # same size with columns a multiple of 16: the register-permute sub-word kernel, 64-bit offsets
    Cc = 65552
    X = la.Matrix(la.Vector.synth(0xD, R * Cc, "u8"), R, Cc)
    T = X.transpose()
    for r, c in ((0, 0), (65535, 65551), (40000, 123), (7, 65000), (65535, 0), (0, 65551)):
        assert T.data.slice_to_vec(c * R + r, 1)[0] == synth.values("u8", 0xD, r * Cc + c, 1)[0]
    row = T.data.slice_to_vec(65551 * R, R)                       # one whole output row = last input column
    assert row[:4096].tobytes() == np.array([synth.values("u8", 0xD, r * Cc + 65551, 1)[0] for r in range(4096)], np.uint8).tobytes()


def test_empty_operands(la):
    """Zero-length vectors and matrices are legal in the reference (Vector::from_vec(vec![]))."""
    for name in ("u8", "i32", "f32", "f64"):
        e = la.Vector.from_vec([], name)
        assert (e + e).len() == 0 and (e * 3).len() == 0 and e.sum() == 0 and la.dot(e, e) == 0 and e == e
        e += e
        m0 = la.Matrix.from_vec([], 0, 5, name); m1 = la.Matrix.from_vec([], 5, 0, name)
        assert (m0 * la.Matrix.from_vec(np.zeros(15, la.parameter(name).np), 5, 3, name)).len() == (0, 3)
        z = (m1 * m0)                                  # 5x0 times 0x5 = 5x5 of zeros
        assert z.len() == (5, 5) and not z.to_vec().any()
        assert m0.transpose().len() == (5, 0)
        assert (la.Vector.from_vec([], name) * m0).to_vec().tolist() == [0] * 5      # v(0) x M(0x5) = zeros(5)
        assert la.mul_transpose_mat(la.Vector.from_vec([], name), m1).to_vec().tolist() == [0] * 5
