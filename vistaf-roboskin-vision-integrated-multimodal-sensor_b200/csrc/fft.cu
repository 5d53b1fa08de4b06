// Dense batched 2-D FFT (C2C, optional real input) built on fft_smem: a row pass and a column
// pass, each staging whole sequences in shared memory.  Used for the FTP reference frame
// (1374 x 1374, shape_ftp.py:867) and as the dense contract path for deformed frames.
#include "fft.cuh"
#include <vector>

int fft1d_create(Fft1D* p, int n) {
  p->n = n;
  p->nf = 0;
  int m = n;
  while (m % 4 == 0) { p->radix[p->nf++] = 4; m /= 4; }
  while (m % 2 == 0) { p->radix[p->nf++] = 2; m /= 2; }
  while (m % 3 == 0) { p->radix[p->nf++] = 3; m /= 3; }
  while (m % 5 == 0) { p->radix[p->nf++] = 5; m /= 5; }
  for (int f = 7; m > 1; f += 2)
    while (m % f == 0) { if (p->nf >= 16) return 1; p->radix[p->nf++] = f; m /= f; }
  std::vector<float2> h(n);
  for (int k = 0; k < n; k++) {
    double a = -2.0 * VISTAF_PI * (double)k / (double)n;
    h[k] = make_float2((float)cos(a), (float)sin(a));
  }
  float2* d = nullptr;
  CK(cudaMalloc(&d, sizeof(float2) * n));
  CK(cudaMemcpy(d, h.data(), sizeof(float2) * n, cudaMemcpyHostToDevice));
  p->tw = d;
  return 0;
}

void fft1d_destroy(Fft1D* p) {
  if (p->tw) cudaFree((void*)p->tw);
  p->tw = nullptr;
}

// One block = `nseq` rows.  in: real (REAL_IN) or complex rows of length n, row stride ld_in
// elements; out: complex rows, stride ld_out.  plane strides in elements; grid.y = batch.
template <int SIGN, bool REAL_IN>
__global__ void k_fft_rows(const void* __restrict__ in_, float2* __restrict__ out, Fft1D p, int rows, int nseq,
                           long long ld_in, long long ld_out, long long plane_in, long long plane_out, float scale) {
  extern __shared__ float2 sm[];
  const int n = p.n;
  float2* a = sm;
  float2* b = sm + (size_t)nseq * n;
  const int r0 = blockIdx.x * nseq;
  const int ns = min(nseq, rows - r0);
  const long long bi = blockIdx.y;
  for (int t = threadIdx.x; t < ns * n; t += blockDim.x) {
    int s = t / n, x = t - s * n;
    if (REAL_IN) {
      const float* in = (const float*)in_ + bi * plane_in + (long long)(r0 + s) * ld_in;
      a[s * n + x] = make_float2(in[x], 0.f);
    } else {
      const float2* in = (const float2*)in_ + bi * plane_in + (long long)(r0 + s) * ld_in;
      a[s * n + x] = in[x];
    }
  }
  float2* res = fft_smem<SIGN>(a, b, p, ns, n);
  for (int t = threadIdx.x; t < ns * n; t += blockDim.x) {
    int s = t / n, x = t - s * n;
    float2 v = res[s * n + x];
    out[bi * plane_out + (long long)(r0 + s) * ld_out + x] = make_float2(v.x * scale, v.y * scale);
  }
}

// One block = `ncol` adjacent columns of a (rows x cols) complex plane, transformed in place
// along y.  Shared layout: column-major with stride n+1 to avoid bank conflicts on the transpose.
template <int SIGN>
__global__ void k_fft_cols(const float2* __restrict__ in, float2* __restrict__ out, Fft1D p, int cols, int ncol,
                           long long ld, long long plane, float scale) {
  extern __shared__ float2 sm[];
  const int n = p.n, lds = n + 1;
  float2* a = sm;
  float2* b = sm + (size_t)ncol * lds;
  const int c0 = blockIdx.x * ncol;
  const int nc = min(ncol, cols - c0);
  const long long base = (long long)blockIdx.y * plane;
  for (int t = threadIdx.x; t < nc * n; t += blockDim.x) {
    int y = t / nc, c = t - y * nc;
    a[c * lds + y] = in[base + (long long)y * ld + c0 + c];
  }
  float2* res = fft_smem<SIGN>(a, b, p, nc, lds);
  for (int t = threadIdx.x; t < nc * n; t += blockDim.x) {
    int y = t / nc, c = t - y * nc;
    float2 v = res[c * lds + y];
    out[base + (long long)y * ld + c0 + c] = make_float2(v.x * scale, v.y * scale);
  }
}

static int smem_opt_in(const void* f, size_t bytes) {
  if (bytes > 48 * 1024) CK(cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

// Dense 2-D transform of `batch` planes.  sign<0 forward, >0 inverse (scaled by 1/(h*w)).
// real_in: `in` is float32 (h x w, stride w); otherwise complex.  out: complex h x w.
int fft2d_dense(const void* in, float2* out, const Fft1D& pw, const Fft1D& ph, int h, int w, int batch, int sign,
                bool real_in, cudaStream_t st) {
  const int nseq_r = 1;
  size_t sm_r = sizeof(float2) * 2 * nseq_r * (size_t)w;
  int ncol = 4;
  while (ncol > 1 && sizeof(float2) * 2 * ncol * (size_t)(h + 1) > 200 * 1024) ncol >>= 1;
  size_t sm_c = sizeof(float2) * 2 * ncol * (size_t)(h + 1);
  if (sm_r > 227 * 1024 || sm_c > 227 * 1024) return vistaf_set_err("fft2d_dense: sequence too long for smem", cudaErrorInvalidValue, __FILE__, __LINE__);
  dim3 gr(cdiv(h, nseq_r), batch), gc(cdiv(w, ncol), batch);
  const long long plane = (long long)h * w;
  const float sc = sign > 0 ? 1.0f / ((float)h * (float)w) : 1.0f;
  if (sign < 0) {
    if (real_in) {
      if (smem_opt_in((const void*)k_fft_rows<-1, true>, sm_r)) return 1;
      VL(), k_fft_rows<-1, true><<<gr, 256, sm_r, st>>>(in, out, pw, h, nseq_r, w, w, plane, plane, 1.f);
    } else {
      if (smem_opt_in((const void*)k_fft_rows<-1, false>, sm_r)) return 1;
      VL(), k_fft_rows<-1, false><<<gr, 256, sm_r, st>>>(in, out, pw, h, nseq_r, w, w, plane, plane, 1.f);
    }
    CKL();
    if (smem_opt_in((const void*)k_fft_cols<-1>, sm_c)) return 1;
    VL(), k_fft_cols<-1><<<gc, 256, sm_c, st>>>(out, out, ph, w, ncol, w, plane, sc);
  } else {
    if (smem_opt_in((const void*)k_fft_rows<1, false>, sm_r)) return 1;
    VL(), k_fft_rows<1, false><<<gr, 256, sm_r, st>>>(in, out, pw, h, nseq_r, w, w, plane, plane, 1.f);
    CKL();
    if (smem_opt_in((const void*)k_fft_cols<1>, sm_c)) return 1;
    VL(), k_fft_cols<1><<<gc, 256, sm_c, st>>>(out, out, ph, w, ncol, w, plane, sc);
  }
  CKL();
  return 0;
}
