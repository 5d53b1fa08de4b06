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
                                              int Ns, int nseq, int ld, const float2* __restrict__ tw) {
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
    float2* dst = out + s * ld + (j - k) * R + k;
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
                                                      const float2* __restrict__ tw) {
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
    out[s * ld + (j - k) * r + k + q * Ns] = acc;
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

// Two consecutive radix-4 passes (sequence sizes Ns and 4 Ns) as one 16-point step per thread.  The four first-pass
// butterflies j = a + q' n/16 (q' = 0..3; a = Ns c' + k < n/16) produce exactly the inputs of the four second-pass
// butterflies j' = 4 Ns c' + r Ns + k (r = 0..3), so the intermediate values stay in registers.  Every product and sum
// is the one the two separate passes compute -- same operands, same order, same twiddle table entries -- so the result
// is bit-identical; one shared-memory round trip and one barrier fewer per pair.
// PAD_IN / PAD_OUT: the buffer read / written holds element i at i + i/16.  At Ns = 1 a thread owns 16 consecutive
// outputs; without the padding a whole warp would store to one pair of banks.
#ifndef VISTAF_FFT_FUSE44
#define VISTAF_FFT_FUSE44 1
#endif
template <int SIGN, bool PAD_IN, bool PAD_OUT>
__device__ __forceinline__ void stockham_pass_4x4(const float2* __restrict__ in, float2* __restrict__ out, int n, int Ns,
                                                  const float2* __restrict__ tw) {
  const int n16 = n >> 4, stride1 = n / (Ns * 4), stride2 = stride1 >> 2;
  for (int a = threadIdx.x; a < n16; a += blockDim.x) {
    const int k = a & (Ns - 1);                       // Ns is a power of two here
    float2 x[16];
#pragma unroll
    for (int m = 0; m < 16; m++) { const int i = a + m * n16; x[m] = in[PAD_IN ? i + (i >> 4) : i]; }
    // first pass: butterfly q' reads x[q' + 4 q] (q = 0..3), inputs q > 0 times W^(q k stride1) when Ns > 1
    float2 w1[3];
    if (Ns > 1) {
#pragma unroll
      for (int q = 1; q < 4; q++) w1[q - 1] = tw_get<SIGN>(tw, q * k * stride1);
    }
#pragma unroll
    for (int qp = 0; qp < 4; qp++) {
      float2 X[4] = {x[qp], x[qp + 4], x[qp + 8], x[qp + 12]};
      if (Ns > 1) {
#pragma unroll
        for (int q = 1; q < 4; q++) X[q] = cmul(X[q], w1[q - 1]);
      }
      butterfly<SIGN, 4>(X);
      x[qp] = X[0]; x[qp + 4] = X[1]; x[qp + 8] = X[2]; x[qp + 12] = X[3];      // output r of butterfly q' -> x[q' + 4 r]
    }
    // second pass (sequence size 4 Ns > 1): butterfly r reads x[q' + 4 r] (q' = 0..3), inputs q' > 0 times W^(q' (r Ns + k) stride2)
    const int o0 = 16 * (a - k) + k;                 // 16 Ns c' + k
#pragma unroll
    for (int r = 0; r < 4; r++) {
      float2 Z[4] = {x[4 * r], x[4 * r + 1], x[4 * r + 2], x[4 * r + 3]};
      const int k2 = r * Ns + k;
#pragma unroll
      for (int qp = 1; qp < 4; qp++) Z[qp] = cmul(Z[qp], tw_get<SIGN>(tw, qp * k2 * stride2));
      butterfly<SIGN, 4>(Z);
#pragma unroll
      for (int qq = 0; qq < 4; qq++) { const int i = o0 + r * Ns + qq * 4 * Ns; out[PAD_OUT ? i + (i >> 4) : i] = Z[qq]; }
    }
  }
}

// fft_smem for ONE sequence with radix-4 pairs fused (stockham_pass_4x4).  b must hold n + n/16 entries (the padded
// intermediate of a first pair that is followed by a second pair lives there); a holds n.  Same result, bit for bit.
template <int SIGN>
__device__ float2* fft_smem_fused(float2* a, float2* b, const Fft1D& p) {
  int Ns = 1, f = 0;
  bool in_pad = false;
  while (f < p.nf) {
    const int r = p.radix[f];
    __syncthreads();
    const bool pair = VISTAF_FFT_FUSE44 && r == 4 && f + 1 < p.nf && p.radix[f + 1] == 4 && (Ns & (Ns - 1)) == 0 &&
                      ((p.n >> 4) % Ns) == 0;
    if (pair) {
      // pad the output only at Ns = 1 (first pass: a -> b) and only when the reader is a fused pair as well
      const bool out_pad = Ns == 1 && f + 3 < p.nf && p.radix[f + 2] == 4 && p.radix[f + 3] == 4 && ((p.n >> 4) % 16) == 0;
      if (in_pad) stockham_pass_4x4<SIGN, true, false>(a, b, p.n, Ns, p.tw);
      else if (out_pad) stockham_pass_4x4<SIGN, false, true>(a, b, p.n, Ns, p.tw);
      else stockham_pass_4x4<SIGN, false, false>(a, b, p.n, Ns, p.tw);
      in_pad = out_pad;
      Ns *= 16;
      f += 2;
    } else {
      switch (r) {
        case 2: stockham_pass<SIGN, 2>(a, b, p.n, Ns, 1, p.n, p.tw); break;
        case 3: stockham_pass<SIGN, 3>(a, b, p.n, Ns, 1, p.n, p.tw); break;
        case 4: stockham_pass<SIGN, 4>(a, b, p.n, Ns, 1, p.n, p.tw); break;
        case 5: stockham_pass<SIGN, 5>(a, b, p.n, Ns, 1, p.n, p.tw); break;
        default: stockham_pass_generic<SIGN>(a, b, p.n, r, Ns, 1, p.n, p.tw); break;
      }
      Ns *= r;
      f += 1;
    }
    float2* t = a;
    a = b;
    b = t;
  }
  __syncthreads();
  return a;
}

// Host: factorise n (prefer radix 4, then 2, 3, 5, then remaining primes) and build the table.
int fft1d_create(Fft1D* p, int n);   // allocates p->tw on the device
void fft1d_destroy(Fft1D* p);
