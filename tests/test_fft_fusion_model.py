"""Host model of csrc/fft.cuh: the fused pass pairs (stockham_pass_pair) against the separate Stockham passes.

The device code fuses two consecutive passes (radix R1 at sequence size Ns, radix R2 at R1 Ns) into one R1 R2-point
step per thread and claims the result is bit-identical because every butterfly sees the same operands.  This test
restates both forms in numpy with ONE butterfly / twiddle routine (complex64, so any difference in which operands meet
shows up as a bit difference), follows the same index formulas as the CUDA source (including the i + i/16 padding of
the first 4.4 pair's output and the greedy pairing schedule of fft_smem_fused), and asserts exact equality for the
lengths the temperature path transforms (3840 columns, 2160 rows) and for lengths where nothing or only a part fuses.
It also pins the factorisation order of fft1d_create.  No GPU involved: it guards the algebra, the GPU parity tests
guard the kernels."""
import numpy as np
import pytest


def factorise(n):                       # fft.cu: fft1d_create
    radix, m = [], n
    for r in (4, 2, 3, 5):
        while m % r == 0:
            radix.append(r)
            m //= r
    f = 7
    while m > 1:
        while m % f == 0:
            radix.append(f)
            m //= f
        f += 2
    return radix


def twiddles(n):
    k = np.arange(n)
    a = -2.0 * np.pi * k / n
    return (np.cos(a).astype(np.float32) + 1j * np.sin(a).astype(np.float32)).astype(np.complex64)


def butterfly(x):
    """Any fixed function of the R inputs does for an equality test; a complex64 DFT keeps it a real transform."""
    r = x.shape[0]
    w = np.exp(-2j * np.pi * np.outer(np.arange(r), np.arange(r)) / r).astype(np.complex64)
    out = np.zeros_like(x)
    for p in range(r):
        acc = x[0] * w[p, 0]
        for q in range(1, r):
            acc = (acc + x[q] * w[p, q]).astype(np.complex64)
        out[p] = acc
    return out


def single_pass(src, n, ns, r, tw):     # fft.cuh: stockham_pass
    nb, stride = n // r, n // (ns * r)
    j = np.arange(nb)
    k = j % ns
    x = np.stack([src[j + q * nb] for q in range(r)])
    if ns > 1:
        for q in range(1, r):
            x[q] = (x[q] * tw[q * k * stride]).astype(np.complex64)
    y = butterfly(x)
    out = np.zeros(n, np.complex64)
    for q in range(r):
        out[(j - k) * r + k + q * ns] = y[q]
    return out


def pad(i):
    return i + (i >> 4)


def pair_pass(src, n, ns, r1, r2, tw, pad_in, pad_out):     # fft.cuh: stockham_pass_pair
    rr = r1 * r2
    ng, stride1, stride2 = n // rr, n // (ns * r1), n // (ns * rr)
    assert ng % ns == 0
    a = np.arange(ng)
    k = a % ns
    x = [src[pad(a + m * ng) if pad_in else a + m * ng] for m in range(rr)]
    for qp in range(r2):
        X = np.stack([x[qp + r2 * q] for q in range(r1)])
        if ns > 1:
            for q in range(1, r1):
                X[q] = (X[q] * tw[q * k * stride1]).astype(np.complex64)
        Y = butterfly(X)
        for r in range(r1):
            x[qp + r2 * r] = Y[r]
    out = np.zeros(n + n // 16 + 1, np.complex64)
    o0 = rr * (a - k) + k
    for r in range(r1):
        Z = np.stack([x[qp + r2 * r] for qp in range(r2)])
        k2 = r * ns + k
        for qp in range(1, r2):
            Z[qp] = (Z[qp] * tw[qp * k2 * stride2]).astype(np.complex64)
        W = butterfly(Z)
        for qq in range(r2):
            i = o0 + r * ns + qq * r1 * ns
            out[pad(i) if pad_out else i] = W[qq]
    return out


def fft_separate(x, n):
    tw, ns, cur = twiddles(n), 1, x.astype(np.complex64)
    for r in factorise(n):
        cur = single_pass(cur, n, ns, r, tw)
        ns *= r
    return cur


def fft_fused(x, n, odd):               # fft.cuh: fft_smem_fused<SIGN, ODD>
    tw, radix = twiddles(n), factorise(n)
    ns, f, in_pad, cur, steps = 1, 0, False, x.astype(np.complex64), []
    while f < len(radix):
        r1 = radix[f]
        r2 = radix[f + 1] if f + 1 < len(radix) else 0
        can = r2 != 0 and n % (r1 * r2) == 0 and (n // (r1 * r2)) % ns == 0
        p44 = can and (r1, r2) == (4, 4)
        p33 = odd and can and (r1, r2) == (3, 3)
        p35 = odd and can and (r1, r2) == (3, 5)
        if p44 or p33 or p35:
            out_pad = False
            if p44 and ns == 1 and f + 3 < len(radix):
                r3, r4 = radix[f + 2], radix[f + 3]
                out_pad = ((r3, r4) == (4, 4) or (odd and (r3, r4) == (3, 3))) and n % (16 * r3 * r4) == 0 and (n // (r3 * r4)) % 16 == 0
            assert not (in_pad and p35)
            cur = pair_pass(cur, n, ns, r1, r2, tw, in_pad, out_pad)
            steps.append((r1, r2, in_pad, out_pad))
            in_pad = out_pad
            ns *= r1 * r2
            f += 2
        else:
            assert not in_pad, "a padded buffer must be read by a pair"
            cur = single_pass(cur[:n], n, ns, r1, tw)
            steps.append((r1,))
            ns *= r1
            f += 1
    assert not in_pad
    return cur[:n], steps


def test_factorisation_order_of_the_plans():
    assert factorise(3840) == [4, 4, 4, 4, 3, 5]
    assert factorise(2160) == [4, 4, 3, 3, 3, 5]
    assert factorise(1374) == [2, 3, 229]


@pytest.mark.parametrize("n,odd,want_steps", [
    (3840, False, [(4, 4, False, True), (4, 4, True, False), (3,), (5,)]),                      # row transforms
    (2160, True, [(4, 4, False, True), (3, 3, True, False), (3, 5, False, False)]),            # column transform
    (3840, True, [(4, 4, False, True), (4, 4, True, False), (3, 5, False, False)]),
    (2160, False, [(4, 4, False, False), (3,), (3,), (3,), (5,)]),
    (1374, True, [(2,), (3,), (229,)]),                                                         # nothing to fuse
    (4096, True, [(4, 4, False, True), (4, 4, True, False), (4, 4, False, False)]),            # configs[4] grid
    (720, True, [(4, 4, False, True), (3, 3, True, False), (5,)]),
    (240, True, [(4, 4, False, False), (3, 5, False, False)]),                                 # a 3.5 pair never reads a padded buffer
])
def test_fused_pairs_equal_the_separate_passes_bit_for_bit(n, odd, want_steps):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    ref = fft_separate(x, n)
    got, steps = fft_fused(x, n, odd)
    assert steps == want_steps
    assert np.array_equal(ref.view(np.uint32), got.view(np.uint32)), "fused pairs changed a bit"
    exact = np.fft.fft(x.astype(np.complex128))
    assert float(np.max(np.abs(ref - exact)) / np.max(np.abs(exact))) < 5e-6          # and it is the DFT


def test_padding_map_spreads_the_sixteen_outputs_of_a_thread():
    """At Ns = 1 thread a stores elements 16 a .. 16 a + 15; with i + i/16 the 16 lanes of a half warp hit 16 different
    bank pairs for every one of the 16 stores (64-bit words, 32 banks of 4 bytes)."""
    for e in range(16):
        lanes = np.arange(16)
        i = 16 * lanes + e
        banks = (2 * pad(i)) % 32
        assert len(set(banks.tolist())) == 16
        assert len(set(((2 * i) % 32).tolist())) == 1                                   # unpadded: one bank pair
