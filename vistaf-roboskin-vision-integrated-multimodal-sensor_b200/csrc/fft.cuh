// Mixed-radix Stockham FFT staged through shared memory (device-side building block).
// Radix 2/3/4/5 butterflies are hand-written; any other prime factor (e.g. 229 for the
// reference's 1374-point FTP transform, 1374 = 2*3*229) goes through a direct small-DFT pass.
// Twiddles come from a float2 table exp(-2*pi*i*k/n), k<n, computed in double on the host.
#pragma once
#include "common.cuh"

struct Fft1D {
  int n;
  int nf;
  int radix[16];
  const float2* tw;  // device pointer, n entries
};

template <int SIGN>
__device__ __forceinline__ float2 tw_get(const float2* __restrict__ tw, int i) {
  float2 w = __ldg(tw + i);
  if (SIGN > 0) w.y = -w.y;
  return w;
}
// multiply by -i (forward) or +i (inverse)
template <int SIGN>
__device__ __forceinline__ float2 rotq(float2 v) {
  return SIGN < 0 ? make_float2(v.y, -v.x) : make_float2(-v.y, v.x);
}

template <int SIGN, int R>
__device__ __forceinline__ void butterfly(float2 (&x)[R]) {
  if (R == 2) {
    float2 a = x[0], b = x[1];
    x[0] = cadd(a, b);
    x[1] = csub(a, b);
  } else if (R == 4) {
    float2 t0 = cadd(x[0], x[2]), t1 = csub(x[0], x[2]);
    float2 t2 = cadd(x[1], x[3]), t3 = rotq<SIGN>(csub(x[1], x[3]));
    x[0] = cadd(t0, t2);
    x[2] = csub(t0, t2);
    x[1] = cadd(t1, t3);
    x[3] = csub(t1, t3);
  } else if (R == 3) {
    const float s = 0.86602540378443864676f;
    float2 t = cadd(x[1], x[2]);
    float2 m = make_float2(x[0].x - 0.5f * t.x, x[0].y - 0.5f * t.y);
    float2 d = csub(x[1], x[2]);
    d = rotq<SIGN>(make_float2(d.x * s, d.y * s));
    x[0] = cadd(x[0], t);
    x[1] = cadd(m, d);
    x[2] = csub(m, d);
  } else if (R == 5) {
    const float c1 = 0.30901699437494742410f, c2 = -0.80901699437494742410f;
    const float s1 = 0.95105651629515357212f, s2 = 0.58778525229247312917f;
    float2 a1 = cadd(x[1], x[4]), a2 = cadd(x[2], x[3]);
    float2 b1 = csub(x[1], x[4]), b2 = csub(x[2], x[3]);
    float2 m1 = make_float2(x[0].x + c1 * a1.x + c2 * a2.x, x[0].y + c1 * a1.y + c2 * a2.y);
    float2 m2 = make_float2(x[0].x + c2 * a1.x + c1 * a2.x, x[0].y + c2 * a1.y + c1 * a2.y);
    float2 n1 = rotq<SIGN>(make_float2(s1 * b1.x + s2 * b2.x, s1 * b1.y + s2 * b2.y));
    float2 n2 = rotq<SIGN>(make_float2(s2 * b1.x - s1 * b2.x, s2 * b1.y - s1 * b2.y));
    x[0] = make_float2(x[0].x + a1.x + a2.x, x[0].y + a1.y + a2.y);
    x[1] = cadd(m1, n1);
    x[4] = csub(m1, n1);
    x[2] = cadd(m2, n2);
    x[3] = csub(m2, n2);
  }
}

// Index arithmetic of a pass without per-butterfly divisions: (sequence, butterfly) advance incrementally with the
// thread stride, and k = j mod Ns is a mask when Ns is a power of two, else a multiply-shift by ceil(2^32 / Ns)
// (exact for j < 2^16, Ns < 2^16).
template <int SIGN, int R>
__device__ __forceinline__ void stockham_pass(const float2* __restrict__ in, float2* __restrict__ out, int n,
                                              int Ns, int nseq, int ld, const float2* __restrict__ tw, int ld_out = -1) {
  if (ld_out < 0) ld_out = ld;
  const int nb = n / R;              // butterflies per sequence
  const int stride = n / (Ns * R);   // twiddle index step
  const bool pow2 = (Ns & (Ns - 1)) == 0;
  const unsigned int magic = (unsigned int)((0x100000000ull + (unsigned)Ns - 1) / (unsigned)Ns);
  int s = 0, j = threadIdx.x;
  while (j >= nb) { j -= nb; s++; }
  for (int t = threadIdx.x; t < nb * nseq; t += blockDim.x) {
    const int k = pow2 ? (j & (Ns - 1)) : j - (int)__umulhi((unsigned)j, magic) * Ns;
    const float2* src = in + s * ld;
    float2 x[R];
#pragma unroll
    for (int q = 0; q < R; q++) {
      x[q] = src[j + q * nb];
      if (q > 0 && Ns > 1) x[q] = cmul(x[q], tw_get<SIGN>(tw, q * k * stride));
    }
    butterfly<SIGN, R>(x);
    float2* dst = out + s * ld_out + (j - k) * R + k;
#pragma unroll
    for (int q = 0; q < R; q++) dst[q * Ns] = x[q];
    j += blockDim.x;
    while (j >= nb) { j -= nb; s++; }
  }
}

// Arbitrary radix r: one thread per output element, r complex MACs each.
template <int SIGN>
__device__ __forceinline__ void stockham_pass_generic(const float2* __restrict__ in, float2* __restrict__ out, int n,
                                                      int r, int Ns, int nseq, int ld,
                                                      const float2* __restrict__ tw, int ld_out = -1) {
  if (ld_out < 0) ld_out = ld;
  const int nb = n / r;
  const int stride = n / (Ns * r);
  for (int t = threadIdx.x; t < n * nseq; t += blockDim.x) {
    const int s = t / n, e = t - s * n;
    const int q = e / nb, j = e - q * nb;
    const int k = j % Ns;
    const float2* src = in + s * ld;
    float2 acc = make_float2(0.f, 0.f);
    for (int p = 0; p < r; p++) {
      float2 v = src[j + p * nb];
      // twiddle W_{Ns*r}^{p*k} times root W_r^{p*q}  (both exact table entries)
      float2 w = tw_get<SIGN>(tw, (int)(((long long)p * k * stride + (long long)((p * q) % r) * nb) % n));
      acc = cadd(acc, cmul(v, w));
    }
    out[s * ld_out + (j - k) * r + k + q * Ns] = acc;
  }
}

// In-shared-memory FFT of `nseq` sequences of length p.n, sequence stride `ld` (float2 units).
// a holds the input, b is scratch of the same size; returns the buffer holding the result.
// Unnormalised in both directions.  All threads of the block must call.
template <int SIGN>
__device__ float2* fft_smem(float2* a, float2* b, const Fft1D& p, int nseq, int ld) {
  int Ns = 1;
  for (int f = 0; f < p.nf; f++) {
    const int r = p.radix[f];
    __syncthreads();
    switch (r) {
      case 2: stockham_pass<SIGN, 2>(a, b, p.n, Ns, nseq, ld, p.tw); break;
      case 3: stockham_pass<SIGN, 3>(a, b, p.n, Ns, nseq, ld, p.tw); break;
      case 4: stockham_pass<SIGN, 4>(a, b, p.n, Ns, nseq, ld, p.tw); break;
      case 5: stockham_pass<SIGN, 5>(a, b, p.n, Ns, nseq, ld, p.tw); break;
      default: stockham_pass_generic<SIGN>(a, b, p.n, r, Ns, nseq, ld, p.tw); break;
    }
    Ns *= r;
    float2* t = a;
    a = b;
    b = t;
  }
  __syncthreads();
  return a;
}

// Two consecutive passes (radix R1 at sequence size Ns, then radix R2 at R1 Ns) as one R1 R2-point step per thread.
// With nG = n / (R1 R2) and a = Ns c' + k < nG (k = a mod Ns; needs nG % Ns == 0), the R2 first-pass butterflies
// j = a + q' nG (q' < R2) read the inputs a + (q' + R2 q) nG (q < R1) and produce exactly the inputs of the R1
// second-pass butterflies j' = R1 Ns c' + r Ns + k (r < R1): output r of butterfly q' is input q' of butterfly r.  So
// the intermediate values stay in registers.  Every product and sum is the one the two separate passes compute --
// same operands, same order, same twiddle table entries -- so the result is bit-identical; one shared-memory round
// trip and one barrier fewer per pair.
// PAD_IN / PAD_OUT: the buffer read / written holds element i at i + i/16.  At Ns = 1 a thread owns R1 R2 = 16
// consecutive outputs; without the padding a whole warp would store to one pair of banks.
#ifndef VISTAF_FFT_FUSE
#define VISTAF_FFT_FUSE 1
#endif
template <int SIGN, int R1, int R2, bool PAD_IN, bool PAD_OUT>
__device__ __forceinline__ void stockham_pass_pair(const float2* __restrict__ in, float2* __restrict__ out, int n, int Ns,
                                                   int nseq, int ld_in, int ld_out, const float2* __restrict__ tw) {
  constexpr int RR = R1 * R2;
  const int nG = n / RR, stride1 = n / (Ns * R1), stride2 = n / (Ns * RR);
  const bool pow2 = (Ns & (Ns - 1)) == 0;
  for (int t = threadIdx.x; t < nG * nseq; t += blockDim.x) {
    int s = 0, a = t;
    while (a >= nG) { a -= nG; s++; }
    const int k = pow2 ? (a & (Ns - 1)) : a % Ns;
    const float2* src = in + s * ld_in;
    float2* dst = out + s * ld_out;
    float2 x[RR];
#pragma unroll
    for (int m = 0; m < RR; m++) { const int i = a + m * nG; x[m] = src[PAD_IN ? i + (i >> 4) : i]; }
    // first pass: butterfly q' reads x[q' + R2 q] (q < R1); inputs q > 0 times W^(q k stride1) when Ns > 1
    float2 w1[R1 - 1];
    if (Ns > 1) {
#pragma unroll
      for (int q = 1; q < R1; q++) w1[q - 1] = tw_get<SIGN>(tw, q * k * stride1);
    }
#pragma unroll
    for (int qp = 0; qp < R2; qp++) {
      float2 X[R1];
#pragma unroll
      for (int q = 0; q < R1; q++) X[q] = x[qp + R2 * q];
      if (Ns > 1) {
#pragma unroll
        for (int q = 1; q < R1; q++) X[q] = cmul(X[q], w1[q - 1]);
      }
      butterfly<SIGN, R1>(X);
#pragma unroll
      for (int r = 0; r < R1; r++) x[qp + R2 * r] = X[r];          // output r of butterfly q'
    }
    // second pass (sequence size R1 Ns > 1): butterfly r reads x[q' + R2 r] (q' < R2); inputs q' > 0 times
    // W^(q' (r Ns + k) stride2)
    const int o0 = RR * (a - k) + k;                   // R1 R2 Ns c' + k
#pragma unroll
    for (int r = 0; r < R1; r++) {
      float2 Z[R2];
#pragma unroll
      for (int qp = 0; qp < R2; qp++) Z[qp] = x[qp + R2 * r];
      const int k2 = r * Ns + k;
#pragma unroll
      for (int qp = 1; qp < R2; qp++) Z[qp] = cmul(Z[qp], tw_get<SIGN>(tw, qp * k2 * stride2));
      butterfly<SIGN, R2>(Z);
#pragma unroll
      for (int qq = 0; qq < R2; qq++) { const int i = o0 + r * Ns + qq * R1 * Ns; dst[PAD_OUT ? i + (i >> 4) : i] = Z[qq]; }
    }
  }
}

// padded sequence length of the second buffer of fft_smem_fused
__host__ __device__ inline int fft_fused_ldb(int n) { return n + n / 16 + 1; }

// fft_smem with pairs of passes fused (stockham_pass_pair: 4.4, 3.3, 3.5).  a: nseq sequences at stride ld_a (>= n);
// b: scratch, nseq sequences at stride ld_b >= fft_fused_ldb(n) (the padded intermediate of a first 4.4 pair that a
// second pair reads lives there).  Returns the buffer that holds the result and its sequence stride in ld_res.  Same
// result as fft_smem, bit for bit.
// ODD = false keeps the 3.3 / 3.5 pairs (up to 15 values per thread next to the radix-5 butterfly's temporaries: ~100
// registers) out of the kernel: the row transforms stay at 78 registers and three blocks per SM with the 4.4 pairs only.
template <int SIGN, bool ODD>
__device__ float2* fft_smem_fused(float2* a, float2* b, const Fft1D& p, int nseq, int ld_a, int ld_b, int& ld_res) {
  int Ns = 1, f = 0;
  bool in_pad = false;
  const int n = p.n;
  while (f < p.nf) {
    const int r1 = p.radix[f], r2 = f + 1 < p.nf ? p.radix[f + 1] : 0;
    __syncthreads();
    const bool can = VISTAF_FFT_FUSE && r2 != 0 && n % (r1 * r2) == 0 && (n / (r1 * r2)) % Ns == 0;
    const bool p44 = can && r1 == 4 && r2 == 4, p33 = ODD && can && r1 == 3 && r2 == 3, p35 = ODD && can && r1 == 3 && r2 == 5;
    if (p44 || p33 || p35) {
      // pad the output only at Ns = 1 of a 4.4 pair (first pass: a -> b) and only when a 4.4 or 3.3 pair reads it
      bool out_pad = false;
      if (p44 && Ns == 1 && f + 3 < p.nf) {
        const int r3 = p.radix[f + 2], r4 = p.radix[f + 3];
        out_pad = ((r3 == 4 && r4 == 4) || (ODD && r3 == 3 && r4 == 3)) && n % (16 * r3 * r4) == 0 && (n / (r3 * r4)) % 16 == 0;
      }
      if (p44) {
        if (in_pad) stockham_pass_pair<SIGN, 4, 4, true, false>(a, b, n, Ns, nseq, ld_a, ld_b, p.tw);
        else if (out_pad) stockham_pass_pair<SIGN, 4, 4, false, true>(a, b, n, Ns, nseq, ld_a, ld_b, p.tw);
        else stockham_pass_pair<SIGN, 4, 4, false, false>(a, b, n, Ns, nseq, ld_a, ld_b, p.tw);
      } else if constexpr (ODD) {
        if (p33) {
          if (in_pad) stockham_pass_pair<SIGN, 3, 3, true, false>(a, b, n, Ns, nseq, ld_a, ld_b, p.tw);
          else stockham_pass_pair<SIGN, 3, 3, false, false>(a, b, n, Ns, nseq, ld_a, ld_b, p.tw);
        } else {
          stockham_pass_pair<SIGN, 3, 5, false, false>(a, b, n, Ns, nseq, ld_a, ld_b, p.tw);    // never reads a padded buffer (see out_pad)
        }
      }
      in_pad = out_pad;
      Ns *= r1 * r2;
      f += 2;
    } else {
      // single pass; a padded input cannot reach here (out_pad requires a pair as the reader), strides may differ
      switch (r1) {
        case 2: stockham_pass<SIGN, 2>(a, b, n, Ns, nseq, ld_a, p.tw, ld_b); break;
        case 3: stockham_pass<SIGN, 3>(a, b, n, Ns, nseq, ld_a, p.tw, ld_b); break;
        case 4: stockham_pass<SIGN, 4>(a, b, n, Ns, nseq, ld_a, p.tw, ld_b); break;
        case 5: stockham_pass<SIGN, 5>(a, b, n, Ns, nseq, ld_a, p.tw, ld_b); break;
        default: stockham_pass_generic<SIGN>(a, b, n, r1, Ns, nseq, ld_a, p.tw, ld_b); break;
      }
      Ns *= r1;
      f += 1;
    }
    float2* t = a; a = b; b = t;
    const int tl = ld_a; ld_a = ld_b; ld_b = tl;
  }
  __syncthreads();
  ld_res = ld_a;
  return a;
}

// Host: factorise n (prefer radix 4, then 2, 3, 5, then remaining primes) and build the table.
int fft1d_create(Fft1D* p, int n);   // allocates p->tw on the device
void fft1d_destroy(Fft1D* p);
