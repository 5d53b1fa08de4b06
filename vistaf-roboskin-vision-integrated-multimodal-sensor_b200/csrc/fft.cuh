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

// Host: factorise n (prefer radix 4, then 2, 3, 5, then remaining primes) and build the table.
int fft1d_create(Fft1D* p, int n);   // allocates p->tw on the device
void fft1d_destroy(Fft1D* p);
