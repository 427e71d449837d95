"""CPU model of the operand format of the f32 tensor engine's f16-split form (csrc/gemm_tc.cu: f16_split_one), in numpy: the
properties DESIGN.md section 3.2 states for it -- what is exact, what the per-product error is, where the 2^-8 threshold
comes from.  No GPU and no library call: this pins the arithmetic the kernels rely on (IEEE fp16 round-to-nearest, which the
device's cvt.rn.f16.f32 and numpy's float16 share)."""
import numpy as np

KEEP_MIN = 2.0 ** -8            # F16_KEEP_MIN: smaller scaled elements are left out of the planes
LIFT = 64.0                     # the residual is lifted by 2^6, hi lowered by 2^-6


def split(xs):
    """xs: scaled f32 values.  Returns hi, hs, ls as float64 views of the fp16 planes."""
    hi = xs.astype(np.float16)
    h = hi.astype(np.float32)
    hs = (h * np.float32(1.0 / LIFT)).astype(np.float16)
    ls = ((xs - h) * np.float32(LIFT)).astype(np.float16)
    return hi.astype(np.float64), hs.astype(np.float64), ls.astype(np.float64)


def scaled_values(n, seed):
    """Magnitudes log-uniform over [2^-8, 2^15): the range a kept element of a row scaled to [2^14, 2^15) can have."""
    rng = np.random.default_rng(seed)
    mag = np.exp2(rng.uniform(-8.0, 15.0, n)).astype(np.float32)
    mag = np.minimum(mag, np.float32(32767.0))
    return (mag * rng.choice(np.float32([-1.0, 1.0]), n)).astype(np.float32)


def test_hs_is_exact_and_nothing_overflows_or_goes_subnormal():
    xs = np.concatenate([scaled_values(1 << 20, 1), np.float32([KEEP_MIN, -KEEP_MIN, 32767.0, 2.0 ** 14, 1.0, 0.0])])
    hi, hs, ls = split(xs)
    assert np.all(np.isfinite(hi)) and np.all(np.isfinite(hs)) and np.all(np.isfinite(ls))
    assert np.array_equal(hs * LIFT, hi), "hs = hi * 2^-6 must be exact: it stays a normal fp16 down to |hi| = 2^-8"
    tiny = np.finfo(np.float16).tiny                        # 2^-14, the smallest normal fp16
    assert np.all((np.abs(hs) >= tiny) | (hs == 0))
    assert np.abs(ls).max() <= 2.0 ** 10                    # |x - hi| <= 2^-11 |x| <= 2^4, lifted by 2^6


def test_the_split_keeps_22_bits_of_every_kept_element():
    xs = scaled_values(1 << 20, 2)
    hi, hs, ls = split(xs)
    err = np.abs(xs.astype(np.float64) - (hi + ls / LIFT))
    assert (err / np.abs(xs)).max() <= 2.0 ** -22           # 11 bits in hi, 11 of the residual's 13 in ls
    # below the threshold hs would lose bits of hi: that is why such elements are left out and added in f32 instead
    small = np.float32([2.0 ** -8.5, 2.0 ** -11 * 1.2345, 2.0 ** -13 * 1.7])
    hi_s, hs_s, _ = split(small)
    assert not np.array_equal(hs_s * LIFT, hi_s)


def test_per_product_error_is_about_2_pow_minus_22():
    a = scaled_values(1 << 20, 3); b = scaled_values(1 << 20, 4)
    ah, as_, al = split(a); bh, bs, bl = split(b)
    approx = ah * bh + as_ * bl + al * bs                   # the three MMAs; every product of two fp16 values is exact in f32
    exact = a.astype(np.float64) * b.astype(np.float64)
    rel = np.abs(approx - exact) / np.abs(exact)
    assert rel.max() <= 2.0 ** -21                          # dropped lo*lo (<= 2^-22) + the two roundings of ls (<= 2^-23 each)
    assert np.median(rel) <= 2.0 ** -24
    # with bf16 planes for the cross terms (bf16(hi) x bf16(x - hi), round 1's form) a lone product errs ~2^-20: above the gate
    def bf16(v):
        u = v.astype(np.float32).view(np.uint32)
        return (((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16).astype(np.uint32).view(np.float32).astype(np.float64)
    a_hi = ah; b_hi = bh
    a_lo = a.astype(np.float64) - a_hi; b_lo = b.astype(np.float64) - b_hi
    approx_bf = a_hi * b_hi + bf16(a_hi) * bf16(b_lo) + bf16(a_lo) * bf16(b_hi)
    rel_bf = np.abs(approx_bf - exact) / np.abs(exact)
    assert rel_bf.max() > 2.0 ** -21 > rel.max() * 0.999


def test_power_of_two_scaling_is_exact():
    rng = np.random.default_rng(5)
    x = rng.standard_normal(1 << 16).astype(np.float32) * np.float32(2.0 ** -17)
    m = np.abs(x).max()
    e = int(np.floor(np.log2(m))) - 14                      # f16_exponent_of_max
    xs = x * np.float32(2.0 ** -e)
    assert 2.0 ** 14 <= np.abs(xs).max() < 2.0 ** 15
    assert np.array_equal((xs * np.float32(2.0 ** e)).astype(np.float32), x)
