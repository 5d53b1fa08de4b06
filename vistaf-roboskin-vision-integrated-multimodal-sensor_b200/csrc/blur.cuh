// Separable float32 Gaussian blur with OpenCV semantics (cv2.GaussianBlur, ksize = round(8*sigma+1)|1,
// BORDER_REFLECT_101), templated on a load functor (row pass prologue) and a store functor
// (column pass epilogue) so normalisation / masking / clamping fuse into the two sweeps.
// Both passes stage a tile + halo in shared memory and give each thread 8 consecutive outputs
// along the filter axis with a sliding register window (1 LDS per 8 FMA).
#pragma once
#include "common.cuh"
#include <stdlib.h>

#define BLUR_MAXK 368          // >= 361 taps (sigma 45), multiple of 8
#define BLUR_ROW_SEG 1280      // outputs per row block (160 threads x 8)
#define BLUR_ROW_THREADS 160
#define BLUR_ROW_PITCH 228     // (1280 + 368 + 8)/8 = 207 -> next value == 4 (mod 32)

struct GaussKernel {
  int ksize;                 // odd
  int kpad;                  // ksize rounded up to a multiple of 8
  float w[BLUR_MAXK];        // zero padded
};

static inline GaussKernel make_gauss(double sigma, int ksize_override = 0) {
  GaussKernel g;
  int k = ksize_override > 0 ? ksize_override : (((int)lrint(sigma * 8.0 + 1.0)) | 1);
  if (k > 361) k = 361;
  g.ksize = k;
  g.kpad = (k + 7) & ~7;
  double sum = 0.0, t[BLUR_MAXK];
  const double s2 = -0.5 / (sigma * sigma);
  for (int i = 0; i < k; i++) {
    double x = i - (k - 1) * 0.5;
    t[i] = exp(s2 * x * x);
    sum += t[i];
  }
  for (int i = 0; i < BLUR_MAXK; i++) g.w[i] = i < k ? (float)(t[i] / sum) : 0.f;
  return g;
}

// Optional active window.  The column pass computes outputs only in blocks that overlap [x0,x1) x [y0,y1); elsewhere
// it hands the Store the per-frame constant *fill (0 when fill is null).  The row pass computes only rows [ry0,ry1)
// (and row segments near the window); the column pass reads every other row as the per-frame constant *rowfill.
// Two uses: (a) a Store that ignores the value outside a ROI (fill = rowfill = null, [ry0,ry1) = the rows the active
// column blocks read); (b) an input that is a per-frame constant c outside a bounding box: rows outside the box blur
// to rowfill = R(c) and pixels farther than the kernel radius from it to fill = C(R(c)), both computed once per
// frame with the same fmaf sequence as the kernels below (gauss_constant), so the result is bit-identical.
struct BlurWindow {
  int x0, x1, y0, y1;
  int ry0, ry1;
  const float* fill; const float* rowfill; size_t stride;   // per-frame values: *(float*)((char*)fill + b * stride)
  const unsigned int* gate; unsigned int gate_mask;         // optional: frame b runs only if (gate[b] (byte stride) & mask) != 0
};
__device__ __forceinline__ bool blur_gated_off(const BlurWindow& win, int b) {
  return win.gate && !(*(const unsigned int*)((const char*)win.gate + (size_t)b * win.stride) & win.gate_mask);
}
__device__ __forceinline__ float blur_frame_const(const float* p, size_t stride, int b) {
  return p ? *(const float*)((const char*)p + (size_t)b * stride) : 0.f;
}
// blur of a constant row / column: the kernels' accumulation order (taps 0..kpad-1, zero-padded)
__device__ __forceinline__ float gauss_constant(const float* w, int kpad, float c) {
  float acc = 0.f;
  for (int k = 0; k < kpad; k++) acc = fmaf(w[k], c, acc);
  return acc;
}

// Stores may provide `void outside(int b, int y, int x, float v)` for pixels of blocks outside the window (a cheaper
// path that need not look at per-pixel masks); otherwise the normal operator() is used there too.
template <class S>
__device__ __forceinline__ auto blur_outside(S& s, int b, int y, int x, float v, int) -> decltype(s.outside(b, y, x, v), void()) {
  s.outside(b, y, x, v);
}
template <class S>
__device__ __forceinline__ void blur_outside(S& s, int b, int y, int x, float v, long) { s(b, y, x, v); }

// Row pass.  grid = (ceil(w/1280), h, batch), block = 160.  Load: float operator()(int b,int y,int x).
template <class Load>
__global__ void __launch_bounds__(BLUR_ROW_THREADS, 12) k_blur_rows(Load ld, float* __restrict__ tmp, int h, int w,
                                                                 long long plane, GaussKernel g, BlurWindow win) {
  __shared__ float s[8 * BLUR_ROW_PITCH];
  __shared__ float wk[BLUR_MAXK];
  const int b = blockIdx.z, y = blockIdx.y, x0 = blockIdx.x * BLUR_ROW_SEG;
  if (y < win.ry0 || y >= win.ry1 || x0 >= win.x1 + 32 || x0 + BLUR_ROW_SEG <= win.x0 - 32) return;
  if (blur_gated_off(win, b)) return;
  if (blur_gated_off(win, b)) return;
  const int rad = g.ksize >> 1;
  const int need = BLUR_ROW_SEG + g.kpad + 8;
  // inputs of outputs nobody reads (farther than the kernel radius from the window grown by one column block) stay 0
  const int xlo = win.x0 - 40 - rad, xhi = win.x1 + 40 + rad;
  for (int e = threadIdx.x; e < need; e += BLUR_ROW_THREADS) {
    int x = x0 - rad + e;
    float v = 0.f;
    if (e < BLUR_ROW_SEG + g.ksize - 1 && x < w + rad && x >= xlo && x < xhi) v = ld(b, y, reflect101(x, w));
    s[(e & 7) * BLUR_ROW_PITCH + (e >> 3)] = v;
  }
  for (int e = threadIdx.x; e < g.kpad; e += BLUR_ROW_THREADS) wk[e] = g.w[e];
  __syncthreads();
  const int t = threadIdx.x;
  // outputs no active column block reads (column blocks are 32 wide, aligned to their own grid) are not computed
  if (x0 + t * 8 + 8 <= win.x0 - 32 || x0 + t * 8 >= win.x1 + 32) return;
  float acc[8], wreg[16];
#pragma unroll
  for (int j = 0; j < 8; j++) { acc[j] = 0.f; wreg[j] = s[j * BLUR_ROW_PITCH + t]; }
  for (int kb = 0; kb < g.kpad; kb += 8) {
    const int c = t + (kb >> 3) + 1;
#pragma unroll
    for (int j = 0; j < 8; j++) wreg[8 + j] = s[j * BLUR_ROW_PITCH + c];
#pragma unroll
    for (int kk = 0; kk < 8; kk++) {
      const float wv = wk[kb + kk];
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] = fmaf(wv, wreg[kk + j], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < 8; j++) wreg[j] = wreg[8 + j];
  }
  float* o = tmp + (long long)b * plane + (long long)y * w;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    int x = x0 + t * 8 + j;
    if (x < w) o[x] = acc[j];
  }
}

// Column pass.  grid = (ceil(w/32), ceil(h/(8*TYT)), batch), block = (32, TYT).
// dynamic smem = (8*TYT + kpad + 8) * 32 floats.  Store: void operator()(int b,int y,int x,float v); void finish(int b).
#ifndef VISTAF_BLUR_COLS_MINB
#define VISTAF_BLUR_COLS_MINB 2     // resident column blocks per SM at 1024 threads: one block stages its tile while the other computes
#endif
template <class Store>
__global__ void __launch_bounds__(1024, VISTAF_BLUR_COLS_MINB)
k_blur_cols(const float* __restrict__ tmp, Store st, int h, int w, long long plane, GaussKernel g,
                            BlurWindow win) {
  extern __shared__ float sc[];
  __shared__ float wk[BLUR_MAXK];
  const int TY = 8 * blockDim.y;
  const int b = blockIdx.z, x = blockIdx.x * 32 + threadIdx.x, y0 = blockIdx.y * TY;
  if (blur_gated_off(win, b)) return;
  if ((int)blockIdx.x * 32 >= win.x1 || (int)blockIdx.x * 32 + 32 <= win.x0 || y0 >= win.y1 || y0 + TY <= win.y0) {
    Store sto = st;                // block outside the window: nothing computed
    if (x < w) {
      const int r0o = threadIdx.y * 8;
      const float fv = blur_frame_const(win.fill, win.stride, b);
#pragma unroll
      for (int j = 0; j < 8; j++) { const int y = y0 + r0o + j; if (y < h) blur_outside(sto, b, y, x, fv, 0); }
    }
    sto.finish(b);
    return;
  }
  const int rad = g.ksize >> 1;
  const int need = TY + g.kpad + 8;
  const float* src = tmp + (long long)b * plane;
  const float rowfill = blur_frame_const(win.rowfill, win.stride, b);
  for (int e = threadIdx.y; e < need; e += blockDim.y) {
    int y = y0 - rad + e;
    float v = 0.f;
    if (e < TY + g.ksize - 1 && x < w && y < h + rad) {
      const int yr = reflect101(y, h);
      v = (yr >= win.ry0 && yr < win.ry1) ? src[(long long)yr * w + x] : rowfill;
    }
    sc[e * 32 + threadIdx.x] = v;
  }
  const int tid = threadIdx.y * 32 + threadIdx.x;
  for (int e = tid; e < g.kpad; e += 32 * blockDim.y) wk[e] = g.w[e];
  __syncthreads();
  const int r0 = threadIdx.y * 8;
  float acc[8], wreg[16];
#pragma unroll
  for (int j = 0; j < 8; j++) { acc[j] = 0.f; wreg[j] = sc[(r0 + j) * 32 + threadIdx.x]; }
  for (int kb = 0; kb < g.kpad; kb += 8) {
#pragma unroll
    for (int j = 0; j < 8; j++) wreg[8 + j] = sc[(r0 + kb + 8 + j) * 32 + threadIdx.x];
#pragma unroll
    for (int kk = 0; kk < 8; kk++) {
      const float wv = wk[kb + kk];
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] = fmaf(wv, wreg[kk + j], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < 8; j++) wreg[j] = wreg[8 + j];
  }
  Store stl = st;                 // per-thread copy: stores may accumulate (reductions fused in the epilogue)
  if (x < w) {
#pragma unroll
    for (int j = 0; j < 8; j++) {
      int y = y0 + r0 + j;
      if (y < h) stl(b, y, x, acc[j]);
    }
  }
  stl.finish(b);                  // block-wide; every thread reaches this
}

// Host launcher: dst/epilogue through Store, source/prologue through Load, tmp = scratch plane(s).
template <class Load, class Store>
static int gauss_blur(Load ld, Store st, float* tmp, int h, int w, int batch, const GaussKernel& gx,
                      const GaussKernel& gy, cudaStream_t stream, const BlurWindow* window = nullptr) {
  const long long plane = (long long)h * w;
  static int tyt_big = 0;                      // column-block height for long kernels: 8 * tyt rows (VISTAF_BLUR_TYT, default 32)
  if (!tyt_big) { const char* e = getenv("VISTAF_BLUR_TYT"); tyt_big = e ? atoi(e) : 32; if (tyt_big != 16 && tyt_big != 32) tyt_big = 32; }
  const int tyt = gy.ksize > 200 ? tyt_big : 16;         // measured: 161 taps run best at 128-row blocks, 361 taps at 256
  BlurWindow win = window ? *window : BlurWindow{0, w, 0, h, 0, h, nullptr, nullptr, 0, nullptr, 0u};
  if (window && !window->rowfill) {           // rows the active column blocks (aligned to their own grid) read
    win.ry0 = win.y0 - (gy.ksize >> 1) - 8 * tyt;
    win.ry1 = win.y1 + (gy.ksize >> 1) + 8 * tyt;
  }
  dim3 g1(cdiv(w, BLUR_ROW_SEG), h, batch);
  VL(), k_blur_rows<Load><<<g1, BLUR_ROW_THREADS, 0, stream>>>(ld, tmp, h, w, plane, gx, win);
  CKL();
  dim3 blk(32, tyt), g2(cdiv(w, 32), cdiv(h, 8 * tyt), batch);
  size_t sm = sizeof(float) * 32 * (size_t)(8 * tyt + gy.kpad + 8);
  static bool attr_done = false;   // per template instantiation
  if (!attr_done && sm > 48 * 1024) {
    CK(cudaFuncSetAttribute((const void*)k_blur_cols<Store>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CK(cudaFuncSetAttribute((const void*)k_blur_cols<Store>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_done = true;
  }
  VL(), k_blur_cols<Store><<<g2, blk, sm, stream>>>(tmp, st, h, w, plane, gy, win);
  CKL();
  return 0;
}
