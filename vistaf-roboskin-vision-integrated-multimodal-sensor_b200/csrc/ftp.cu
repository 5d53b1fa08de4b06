// Fourier-Transform-Profilometry shape path + force model on sm_100a (Code/shape_ftp.py, Code/force_sensor.py).
// Deformed frames are processed as a batch against one cached reference frame; every stage is a
// kernel over the whole batch (grid.z / grid.y = frame).  Reference file:line cited per stage.
//
// Stage map (shape_ftp.py):
//   A  demodulation :810-1040   Sobel + exact percentiles -> bad pixels -> Telea fill -> illumination
//                               normalisation (sigma 45) -> pre-blur -> apodise -> median -> reflect pad ->
//                               carrier patch of the 1374^2 spectrum (pruned DFT: 21 columns x 21 rows are
//                               all the dense transform's consumer reads) -> windowed inverse -> fractional
//                               ramp -> ratio with the cached reference field -> wrapped phase, amplitude
//   B  reliable mask :739-776   quality blur -> p25 -> close -> largest 8-CC -> chamfer erosion
//   C  unwrap :1043-1080        integer wrap-count field by weighted union-find (labels.cuh)
//   D  detrend :1357-1420, :1716-1751   IRLS plane, two-pass IRLS quadric, contact / background split
//   E  smoothing, sign :1755-1768
//   F  frontier / assembly / phase->mm :1803-1851, :682-705
//   G  blob filter :1215-1281, force reduction force_sensor.py:93-167
#include "../../include/vistaf.h"
#include "common.cuh"
#include "fft.cuh"
#include "blur.cuh"
#include "select.cuh"
#include "morph.cuh"
#include "labels.cuh"
#include "telea.cuh"
#include <cuda_pipeline.h>
#include <string.h>
#include <mutex>
#include <new>
#include <vector>
#include <algorithm>

int fft2d_dense(const void* in, float2* out, const Fft1D& pw, const Fft1D& ph, int h, int w, int batch, int sign,
                bool real_in, cudaStream_t st);

namespace vftp {
#define NPATCH_MAX 41             // 2*PATCH_HALF_WIDTH_BINS + 1 <= 41
#define FNAN __int_as_float(0x7fc00000)
#define IRLS_C 4.685f

struct FtpScal {                  // per-frame scalars in the workspace
  double ne[28];                  // normal equations: 21 upper-triangular + 6 rhs + count
  float coef[6];
  float sigma, med;               // IRLS scale / location
  int irls_n, irls_skip;
  unsigned int irls_ticket, nc;     // nc: samples in the compacted list of the current fit
  float hi_thr, grad_thr, mu, amp_thr, contact_thr, bg_med, core_thr, core_med, q2;
  ull seed_key;                   // (quality key << 32) | ~index
  ull best_cc;                    // (area << 32) | ~root
  unsigned int n_rel, n_bg, n_contact, n_contact_d, n_bad, n_kept, residues, uf_bad, status;
  unsigned int n_holes, n_cand, n_outrel;      // F14: reliable pixels without a height, hole candidates, output_reliable pixels
  float hole_med, hole_fill;
  int seed_root, seed_k;
  int bg_is_rel, flip, detrend1;
  unsigned int gmax_bits;         // blob filter: max depth (float bits, depth >= 0)
  double vol_sum; unsigned int vol_cnt; unsigned int vol_max_bits;
  double pos_sum, neg_sum;
};

struct FtpWs {                    // byte offsets into the chunk workspace
  size_t img, grad, bad0, bad, tmp, I, Iw, R, Pw, wrapped, qprod, quality, m1, m2, reliable, uf, area, unw,
      contact, contact_d, hnum, hm, dist_in, dist_out, dtmp, snum, sm, peak, scal, sel, telea, active, total;
};

}  // namespace vftp
struct VistafFtpPlan {
  VistafFtpParams P;
  int h, w, N, np;                // crop size, padded FFT size, patch size
  uint8_t *roi, *circ, *valid;    // device [h][w]
  float *apo, *den9;              // device [h][w]
  float *xnt, *ynt;               // device [w], [h]: normalised fit coordinates (x - cx) / cx, float32 (shape_ftp.py:1107)
  Fft1D fN;
  GaussKernel g_illum, g_pre, g_qual, g_smooth, g_unrel;
  MorphSE se_bad, se_close, se_dil, se_edge;
  bool have_ref;
  float2* cref;                   // reference-frame field (shape_ftp.py:1632-1639), device [h][w]
  float* amp_ref;
  float2 *rampx, *rampy;          // fractional demodulation ramps over padded coordinates, device [N]
  float* hann;                    // device [np]
  int* kxs; int* kys;             // device [np]: unsigned frequency index of every patch column / row
  int debug_flags;                // bit 0: run the sequential flood unwrap on every frame (parity tests)
  ull* peak_keys;                 // device [32]: top-n spectrum peaks of the reference frame
  void* peak_result;              // device PeakResult
};

namespace vftp {
// ------------------------------------------------------------------ calibrated curves (shape_ftp.py:682-700, force_sensor.py:129-167)
__host__ __device__ inline double curve_eval(const VistafCurveModel& m, double x) {
  const double a = m.p[0], b = m.p[1], c = m.p[2];
  switch (m.type) {
    case VISTAF_MODEL_LINEAR0: return a * x;
    case VISTAF_MODEL_LINEAR: return a * x + b;
    case VISTAF_MODEL_POLY2: return a * x * x + b * x + c;
    case VISTAF_MODEL_SAT_EXP: return a * (1.0 - exp(-b * fmax(x, 0.0)));
    case VISTAF_MODEL_GROWTH: return a * (exp(b * fmax(x, 0.0)) - 1.0);
    case VISTAF_MODEL_HINGE_SAT:
      return a * ((1.0 - exp(-b * fmax(x - c, 0.0))) - (1.0 - exp(-b * fmax(0.0 - c, 0.0))));
  }
  return nan("");
}
}  // namespace vftp
extern "C" double vistaf_curve_predict(const VistafCurveModel* model, double x) { return vftp::curve_eval(*model, x); }
namespace vftp {

// ------------------------------------------------------------------ F1: geometry (shape_ftp.py:383-403)
__global__ void k_ftp_geometry(uint8_t* roi, uint8_t* circ, uint8_t* valid, float* apo, int h, int w, int cx, int cy,
                               int r, int r_valid, int taper) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  long long d2 = (long long)(x - cx) * (x - cx) + (long long)(y - cy) * (y - cy);
  size_t i = (size_t)y * w + x;
  circ[i] = d2 <= (long long)r * r;
  roi[i] = d2 <= (long long)r_valid * r_valid;
  double d = sqrt((double)d2);
  double r_in = fmax(0.0, (double)(r - taper));
  float a = 0.f;
  if (d <= r_in) a = 1.0f;
  else if (d <= (double)r && taper > 0) {
    double t = (d - r_in) / fmax(1e-6, (double)taper);
    a = (float)(0.5 * (1.0 + cos(VISTAF_PI * t)));
  }
  apo[i] = a;
  valid[i] = a > 1e-6f;
}

// ------------------------------------------------------------------ crop + gray (shape_ftp.py:1508-1511)
__global__ void k_crop_gray(const uint8_t* __restrict__ frames, uint8_t* __restrict__ gray, int H, int W, int x1,
                            int y1, int h, int w) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
  if (x >= w) return;
  const uint8_t* p = frames + ((size_t)b * H * W + (size_t)(y1 + y) * W + (x1 + x)) * 3;
  gray[((size_t)b * h + y) * w + x] = (uint8_t)((9798 * p[2] + 19235 * p[1] + 3735 * p[0] + 16384) >> 15);
}

}  // namespace vftp
extern "C" int vistaf_crop_gray(const uint8_t* frames, int batch, int H, int W, int x1, int y1, int h, int w,
                                uint8_t* gray, void* stream) {
  if (x1 < 0 || y1 < 0 || x1 + w > W || y1 + h > H) return vistaf_set_err("crop outside frame", cudaErrorInvalidValue, __FILE__, __LINE__);
  vistaf_stage_mark("F.crop_gray", (cudaStream_t)stream);
  VL(), vftp::k_crop_gray<<<dim3(cdiv(w, 256), h, batch), 256, 0, (cudaStream_t)stream>>>(frames, gray, H, W, x1, y1, h, w);
  CKL();
  vistaf_stage_mark("", (cudaStream_t)stream);
  return 0;
}

namespace vftp {
// ------------------------------------------------------------------ generic small helpers
__global__ void k_ftp_init(FtpScal* sc, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  FtpScal z;
  memset(&z, 0, sizeof(z));
  sc[i] = z;
}

struct PlaneLoad {
  const float* p; long long plane; int w;
  __device__ float operator()(int b, int y, int x) const { return p[(long long)b * plane + (long long)y * w + x]; }
};
struct PlaneStore {
  float* p; long long plane; int w;
  __device__ void operator()(int b, int y, int x, float v) { p[(long long)b * plane + (long long)y * w + x] = v; }
  __device__ void finish(int) {}
};

enum { G_HI = 0, G_GRAD, G_MU, G_AMP, G_CONTACT, G_BGMED, G_CORETHR, G_COREMED, G_IRLS_MED, G_IRLS_MAD, G_CONTACT2 };
__device__ __forceinline__ void ftp_grab(FtpScal& s, int which, float r, unsigned int cnt) {   // a finished quantile -> its slot
  switch (which) {
    case G_HI: s.hi_thr = cnt ? r : 255.0f; break;                           // shape_ftp.py:630
    case G_GRAD: s.grad_thr = cnt ? r : 0.0f; break;                         // :640
    case G_MU: s.mu = cnt ? r : 0.0f; break;                                 // :848 (fallback 0)
    case G_AMP: s.amp_thr = r; break;
    case G_CONTACT: s.contact_thr = (cnt && r == r) ? r : 0.0f; break;
    case G_CONTACT2: if (s.q2 == s.q2 && cnt) s.contact_thr = r; break;
    case G_BGMED: s.bg_med = (cnt && r == r) ? r : 0.0f; break;
    case G_CORETHR: s.core_thr = r; break;
    case G_COREMED: s.core_med = r; s.flip = (cnt > 0 && r > 0.f) ? 1 : 0; break;
    case G_IRLS_MED: s.med = r; break;
    case G_IRLS_MAD: s.sigma = __fmul_rn(1.4826f, __fadd_rn(r, 1e-6f)); break;
  }
}
struct GrabOut {                 // output functor of select_quantile_out: runs on thread 0 of the finishing block
  FtpScal* sc; int which;
  __device__ void operator()(int b, float r, unsigned int cnt) const { ftp_grab(sc[b], which, r, cnt); }
};
// select + grab in one call (lists behind the `batch` states of `sel`)
template <class Val>
static int select_grab(Val val, FtpScal* sc, int which, SelectState* sel, long long n, int batch, float q, cudaStream_t st,
                       const float* qv = nullptr, size_t qstride = 0, bool heavy = false) {
  return select_quantile_out2(val, val, GrabOut{sc, which}, sel, sel_lists(sel, batch), n, batch, q, st, qv, qstride, SEL_CAP,
                              heavy);
}

// ------------------------------------------------------------------ F2: bad pixels (shape_ftp.py:625-649)
// img = float(gray); |Sobel| with BORDER_REFLECT_101; both thresholds are exact percentiles over `valid`.
__global__ void k_ftp_sobel(const uint8_t* __restrict__ gray, float* __restrict__ img, float* __restrict__ grad, int h,
                            int w) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
  if (x >= w) return;
  const uint8_t* g = gray + (size_t)b * h * w;
  int xm = reflect101(x - 1, w), xp = reflect101(x + 1, w), ym = reflect101(y - 1, h), yp = reflect101(y + 1, h);
  int a00 = g[(size_t)ym * w + xm], a01 = g[(size_t)ym * w + x], a02 = g[(size_t)ym * w + xp];
  int a10 = g[(size_t)y * w + xm], a12 = g[(size_t)y * w + xp];
  int a20 = g[(size_t)yp * w + xm], a21 = g[(size_t)yp * w + x], a22 = g[(size_t)yp * w + xp];
  int gx = (a02 + 2 * a12 + a22) - (a00 + 2 * a10 + a20);
  int gy = (a20 + 2 * a21 + a22) - (a00 + 2 * a01 + a02);
  size_t i = (size_t)b * h * w + (size_t)y * w + x;
  img[i] = (float)g[(size_t)y * w + x];
  grad[i] = sqrtf((float)(gx * gx + gy * gy));
}

// four mask bytes, then (only where one is set) four values: bit k of the result = element k is a finite masked sample
__device__ __forceinline__ unsigned int masked_vec4(const float* p4, const uint8_t* m4, long long i, long long n, float v[4]) {
  const uchar4 mm = *reinterpret_cast<const uchar4*>(m4);
  unsigned int m = ((mm.x ? 1u : 0u) | (mm.y ? 2u : 0u) | (mm.z ? 4u : 0u) | (mm.w ? 8u : 0u)) & sel_tail_mask(i, n);
  if (!m) return 0u;
  const float4 f = *reinterpret_cast<const float4*>(p4);
  v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
#pragma unroll
  for (int k = 0; k < 4; k++)
    if (!(fabsf(v[k]) <= 3.4e38f)) m &= ~(1u << k);
  return m;
}
struct MaskedVal {              // value stream: plane[b][i] where mask[i] (mask shared across frames)
  const float* p; const uint8_t* m; long long plane;
  __device__ bool operator()(int b, long long i, float& v) const {
    if (!m[i]) return false;
    v = p[(long long)b * plane + i];
    return fabsf(v) <= 3.4e38f;
  }
  __device__ bool vec_ok() const { return (plane & 3) == 0 && sel_aligned16(p) && ((unsigned long long)m & 3ull) == 0; }
  __device__ unsigned int vec4(int b, long long i, long long n, float v[4]) const {
    return masked_vec4(p + (long long)b * plane + i, m + i, i, n, v);
  }
};
struct MaskedValB {             // per-frame mask
  const float* p; const uint8_t* m; long long plane;
  __device__ bool operator()(int b, long long i, float& v) const {
    if (!m[(long long)b * plane + i]) return false;
    v = p[(long long)b * plane + i];
    return fabsf(v) <= 3.4e38f;
  }
  __device__ bool vec_ok() const { return (plane & 3) == 0 && sel_aligned16(p) && ((unsigned long long)m & 3ull) == 0; }
  __device__ unsigned int vec4(int b, long long i, long long n, float v[4]) const {
    return masked_vec4(p + (long long)b * plane + i, m + (long long)b * plane + i, i, n, v);
  }
};

struct BadLoad {                // (img >= hi) | (grad >= gthr), within valid
  const float* img; const float* grad; const uint8_t* valid; const FtpScal* sc; long long plane; int w;
  __device__ int operator()(int b, int y, int x) const {
    long long i = (long long)y * w + x;
    if (!valid[i]) return 0;
    return img[(long long)b * plane + i] >= sc[b].hi_thr || grad[(long long)b * plane + i] >= sc[b].grad_thr;
  }
};
__global__ void k_ftp_badraw(BadLoad ld, uint8_t* __restrict__ dst, int h, int w) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
  if (x >= w) return;
  dst[(size_t)b * h * w + (size_t)y * w + x] = (uint8_t)ld(b, y, x);
}
__global__ void k_ftp_count_mask(const uint8_t* __restrict__ m, FtpScal* sc, int npix, int which) {
  const int b = blockIdx.y;
  int c = 0;
  const uint8_t* mb = m + (size_t)b * npix;
  if ((npix & 3) == 0 && ((unsigned long long)m & 3ull) == 0) {       // four bytes per load, four loads in flight per thread
    const unsigned int* m4 = reinterpret_cast<const unsigned int*>(mb);
    const int n4 = npix >> 2, step = gridDim.x * blockDim.x;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += 4 * step) {
      unsigned int u[4];
#pragma unroll
      for (int j = 0; j < 4; j++) u[j] = i0 + j * step < n4 ? m4[i0 + j * step] : 0u;
#pragma unroll
      for (int j = 0; j < 4; j++)
        c += ((u[j] & 0xffu) != 0) + ((u[j] & 0xff00u) != 0) + ((u[j] & 0xff0000u) != 0) + ((u[j] & 0xff000000u) != 0);
    }
  } else {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) c += mb[i] != 0;
  }
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0 && c) {
    unsigned int* dst = which == 0 ? &sc[b].n_bad : which == 1 ? &sc[b].n_rel : which == 2 ? &sc[b].n_contact
                      : which == 3 ? &sc[b].n_contact_d : &sc[b].n_kept;
    atomicAdd(dst, (unsigned)c);
  }
}
struct TeleaMask {
  const uint8_t* m; long long plane; int w;
  __device__ bool operator()(int b, int y, int x) const { return m[(long long)b * plane + (long long)y * w + x] != 0; }
};

// ------------------------------------------------------------------ F4: normalisation (shape_ftp.py:831-848)
struct IllumStoreF {            // I = img / (blur + 1e-6) - 1
  const float* img; float* I; long long plane; int w;
  __device__ void operator()(int b, int y, int x, float blur) {
    long long i = (long long)b * plane + (long long)y * w + x;
    I[i] = __fsub_rn(__fdiv_rn(img[i], __fadd_rn(blur, 1e-6f)), 1.0f);
  }
  __device__ void finish(int) {}
};
struct ApoStore {               // Iw = blur(I) * apodization
  const float* apo; float* Iw; long long plane; int w;
  __device__ void operator()(int b, int y, int x, float v) {
    long long i = (long long)y * w + x;
    Iw[(long long)b * plane + i] = __fmul_rn(v, apo[i]);
  }
  __device__ void finish(int) {}
};

// ------------------------------------------------------------------ F5-F7: carrier patch by pruned DFT
// Row part: R[y][j] = sum_X Ip[y][X] e^{-2 pi i kx_j X / N} over the reflect-padded row (BORDER_REFLECT, :857)
// of Iw - mu.  Padded rows are mirror images of inner rows, so only the h inner rows are transformed.
__global__ void __launch_bounds__(256) k_ftp_rowdft(const float* __restrict__ Iw, const FtpScal* sc, float2* __restrict__ R,
                                                     const float2* __restrict__ tw, const int* __restrict__ kxs, int h, int w,
                                                     int N, int pad, int np) {
  extern __shared__ float row[];                   // N floats
  const int b = blockIdx.y, y = blockIdx.x;
  const float mu = sc[b].mu;
  const float* src = Iw + ((size_t)b * h + y) * w;
  for (int X = threadIdx.x; X < N; X += blockDim.x) row[X] = __fsub_rn(src[reflect(X - pad, w)], mu);
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int j = wid; j < np; j += nw) {
    const int k = kxs[j];
    int idx = (int)(((long long)k * lane) % N);
    const int step = (int)(((long long)k * 32) % N);
    float ar = 0.f, ai = 0.f;
    for (int X = lane; X < N; X += 32) {
      const float2 t = __ldg(tw + idx);
      const float v = row[X];
      ar = fmaf(v, t.x, ar);
      ai = fmaf(v, t.y, ai);
      idx += step;
      if (idx >= N) idx -= N;
    }
    ar = warp_sum(ar); ai = warp_sum(ai);
    if (lane == 0) R[((size_t)b * h + y) * np + j] = make_float2(ar, ai);
  }
}
// Column part + window: Pw[i][j] = hann[i] hann[j] sum_Y R[refl(Y - pad)][j] e^{-2 pi i ky_i Y / N}  (:929-947)
__global__ void __launch_bounds__(256) k_ftp_coldft(const float2* __restrict__ R, float2* __restrict__ Pw,
                                                     const float2* __restrict__ tw, const int* __restrict__ kys,
                                                     const float* __restrict__ hann, int h, int N, int pad, int np) {
  extern __shared__ float2 col[];                  // h entries
  const int b = blockIdx.y, j = blockIdx.x;
  for (int y = threadIdx.x; y < h; y += blockDim.x) col[y] = R[((size_t)b * h + y) * np + j];
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int i = wid; i < np; i += nw) {
    const int k = kys[i];
    int idx = (int)(((long long)k * lane) % N);
    const int step = (int)(((long long)k * 32) % N);
    float2 acc = make_float2(0.f, 0.f);
    for (int Y = lane; Y < N; Y += 32) {
      const float2 t = __ldg(tw + idx);
      const float2 v = col[reflect(Y - pad, h)];
      acc.x = fmaf(v.x, t.x, fmaf(-v.y, t.y, acc.x));
      acc.y = fmaf(v.x, t.y, fmaf(v.y, t.x, acc.y));
      idx += step;
      if (idx >= N) idx -= N;
    }
    acc.x = warp_sum(acc.x); acc.y = warp_sum(acc.y);
    if (lane == 0) {
      const float wgt = __fmul_rn(hann[i], hann[j]);
      Pw[((size_t)b * np + i) * np + j] = make_float2(acc.x * wgt, acc.y * wgt);
    }
  }
}
// Inverse of the re-centred patch on the un-padded rows/columns, fractional ramp (:955-961), and either
//   REF: store the reference field and its amplitude (:990-997), or
//   DEF: ratio with the reference field -> wrapped phase (:1681-1689) and amp_ref * amp_def (:741).
// The twiddle e^{+2 pi i f X / N} of tap f at position X is the forward-table entry (|f| X) mod N, conjugated for f >= 0: walking f = -half ..
// half, the index moves by -(X mod N) down to 0 and by +(X mod N) up again, so one modulo per pixel replaces one 64-bit
// multiply + modulo per tap; the entries read are the same ones.
struct TwWalk {
  int idx, s, N, f;
  __device__ TwWalk(int half, int X, int N_) : s(X % N_), N(N_), f(-half) { idx = (int)(((long long)half * s) % N_); }
  __device__ __forceinline__ float2 next(const float2* __restrict__ tw) {
    float2 t = __ldg(tw + idx);
    if (f >= 0) t.y = -t.y;
    if (f < 0) { idx -= s; if (idx < 0) idx += N; }
    else { idx += s; if (idx >= N) idx -= N; }
    f++;
    return t;
  }
};
template <bool REF>
__global__ void __launch_bounds__(256) k_ftp_field(const float2* __restrict__ Pw, const float2* __restrict__ tw,
                                                    const float2* __restrict__ rampx, const float2* __restrict__ rampy,
                                                    float2* __restrict__ cref, float* __restrict__ amp_ref,
                                                    float* __restrict__ wrapped, float* __restrict__ qprod, int h, int w,
                                                    int N, int pad, int np) {
  __shared__ float2 Tj[NPATCH_MAX];
  const int b = blockIdx.y, y = blockIdx.x, Y = y + pad, half = np / 2;
  if (threadIdx.x < np) {
    const int j = threadIdx.x;
    float2 acc = make_float2(0.f, 0.f);
    TwWalk wk(half, Y, N);
    for (int i = 0; i < np; i++) {
      const float2 p = Pw[((size_t)b * np + i) * np + j];
      const float2 t = wk.next(tw);
      acc.x = fmaf(p.x, t.x, fmaf(-p.y, t.y, acc.x));
      acc.y = fmaf(p.x, t.y, fmaf(p.y, t.x, acc.y));
    }
    Tj[j] = acc;
  }
  __syncthreads();
  const float inv = 1.0f / ((float)N * (float)N);
  const float2 ry = rampy[Y];
  for (int x = threadIdx.x; x < w; x += blockDim.x) {
    const int X = x + pad;
    float2 acc = make_float2(0.f, 0.f);
    TwWalk wk(half, X, N);
    for (int j = 0; j < np; j++) {
      const float2 p = Tj[j];
      const float2 t = wk.next(tw);
      acc.x = fmaf(p.x, t.x, fmaf(-p.y, t.y, acc.x));
      acc.y = fmaf(p.x, t.y, fmaf(p.y, t.x, acc.y));
    }
    float2 c = cmul(make_float2(acc.x * inv, acc.y * inv), cmul(rampx[X], ry));
    const size_t i = (size_t)y * w + x;
    const float amp = sqrtf(c.x * c.x + c.y * c.y);
    if (REF) {
      cref[i] = c;
      amp_ref[i] = amp;
    } else {
      const float2 r = cmulc(c, cref[i]);
      wrapped[(size_t)b * h * w + i] = atan2f(r.y, r.x);
      qprod[(size_t)b * h * w + i] = __fmul_rn(amp_ref[i], amp);
    }
  }
}
// dense path input (reference frame): reflect-padded, median-removed plane (:857-862)
__global__ void k_ftp_pad(const float* __restrict__ Iw, const FtpScal* sc, float* __restrict__ Ip, int h, int w, int N, int pad) {
  int X = blockIdx.x * blockDim.x + threadIdx.x, Y = blockIdx.y;
  if (X >= N) return;
  Ip[(size_t)Y * N + X] = __fsub_rn(Iw[(size_t)reflect(Y - pad, h) * w + reflect(X - pad, w)], sc[0].mu);
}

// ------------------------------------------------------------------ F6: carrier peak (shape_ftp.py:420-483) on the device
// find_top_peaks: the n largest |F| outside the DC box of the fftshift-ed spectrum, ordered by (magnitude desc, raster
// index asc).  One sweep per peak: round r takes the largest key strictly below the key of round r-1.
struct PeakResult {             // device-resident carrier of the reference frame (no host round trip inside a step)
  float pxf, pyf; int bx, by; float mag; int ok;
  int pxi, pyi;                 // nearest integer bin (patch centre)
  double kx, ky;                // signed carrier frequency in bins
};
__global__ void __launch_bounds__(256) k_ftp_peak_round(const float2* __restrict__ F, ull* __restrict__ keys, int round, int N,
                                                        int dc) {
  const ull limit = round ? keys[round - 1] : ~0ull;
  const int c = N / 2;
  ull best = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N * N; i += gridDim.x * blockDim.x) {
    const int sy = i / N, sx = i - sy * N;                       // fftshift-ed coordinates
    if (sy >= c - dc && sy < c + dc && sx >= c - dc && sx < c + dc) continue;
    const int ky = (sy - c + N) % N, kx = (sx - c + N) % N;
    const float2 v = F[(size_t)ky * N + kx];
    const ull k = ((ull)__float_as_uint(hypotf(v.x, v.y)) << 32) | (ull)(0xffffffffu - (unsigned int)i);
    if (k < limit && k > best) best = k;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { const ull t = __shfl_xor_sync(0xffffffffu, best, o); best = t > best ? t : best; }
  if ((threadIdx.x & 31) == 0 && best) atomicMax(keys + round, best);
}
// choose_carrier_peak (:444) + refine_peak_parabolic_log (:473)
__global__ void k_ftp_peak_choose(const float2* __restrict__ F, const ull* __restrict__ keys, int n, int N, float max_dy_frac,
                                  PeakResult* out) {
  const int c = N / 2, md = (int)(max_dy_frac * (float)N);
  int bx = -1, by = -1;
  float bm = -1.f;
  bool any_right = false, any_right_near = false, any_near = false;
  for (int q = 0; q < n; q++) {
    if (!keys[q]) continue;
    const unsigned int iq = 0xffffffffu - (unsigned int)(keys[q] & 0xffffffffull);
    const int yq = (int)(iq / (unsigned)N), xq = (int)(iq % (unsigned)N);
    any_right |= xq > c; any_near |= abs(yq - c) <= md; any_right_near |= xq > c && abs(yq - c) <= md;
  }
  // the reference keeps the right-half peaks (if any), then of those the ones near the centre row (if any), and
  // takes the strongest (the first of equals in peak order)
  for (int r = 0; r < n; r++) {
    if (!keys[r]) continue;
    const unsigned int i = 0xffffffffu - (unsigned int)(keys[r] & 0xffffffffull);
    const int y = (int)(i / (unsigned)N), x = (int)(i % (unsigned)N);
    const float m = __uint_as_float((unsigned int)(keys[r] >> 32));
    const bool right = x > c, near = abs(y - c) <= md;
    bool in_set;
    if (any_right) in_set = any_right_near ? (right && near) : right;
    else in_set = any_near ? near : true;
    if (in_set && m > bm) { bm = m; bx = x; by = y; }
  }
  PeakResult r;
  r.bx = bx; r.by = by; r.mag = bm; r.ok = bx >= 0;
  r.pxf = (float)bx; r.pyf = (float)by;
  if (bx > 0 && bx < N - 1 && by > 0 && by < N - 1) {
    auto lm = [&](int sy, int sx) {
      const float2 v = F[(size_t)((sy - c + N) % N) * N + (sx - c + N) % N];
      return logf(hypotf(v.x, v.y) + 1e-12f);
    };
    auto parab = [](float fm, float f0, float fp) {
      const float den = fm - 2.0f * f0 + fp;
      return fabsf(den) < 1e-12f ? 0.0f : 0.5f * (fm - fp) / den;
    };
    r.pxf = (float)bx + parab(lm(by, bx - 1), lm(by, bx), lm(by, bx + 1));
    r.pyf = (float)by + parab(lm(by - 1, bx), lm(by, bx), lm(by + 1, bx));
  }
  *out = r;
}

// Patch frequencies and fractional demodulation ramps from the refined peak (shape_ftp.py:929-961), on the device.
__global__ void k_ftp_carrier_setup(PeakResult* pr, int* __restrict__ kxs, int* __restrict__ kys, float2* __restrict__ rampx,
                                    float2* __restrict__ rampy, int N, int np) {
  const int bw = np / 2, c = N / 2;
  const float pxf = pr->pxf, pyf = pr->pyf;
  const int pxi = __double2int_rn((double)pxf), pyi = __double2int_rn((double)pyf);
  const bool fits = pr->ok && pxi - bw >= 0 && pxi + bw + 1 <= N && pyi - bw >= 0 && pyi + bw + 1 <= N;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) {
    pr->pxi = pxi; pr->pyi = pyi;
    pr->kx = (double)pxf - (double)c; pr->ky = (double)pyf - (double)c;
    if (!fits) pr->ok = 0;        // carrier patch touches the spectrum border (unsupported): frames report NO_PEAK
  }
  if (i < np) {
    kxs[i] = fits ? ((pxi - bw + i - c) % N + N) % N : 0;       // fftshift-ed column -> unsigned frequency index
    kys[i] = fits ? ((pyi - bw + i - c) % N + N) % N : 0;
  }
  if (i < N) {
    const double dpx = (double)pxf - (double)pxi, dpy = (double)pyf - (double)pyi;
    const bool on = fabs(dpx) > 1e-6 || fabs(dpy) > 1e-6;
    const double t = (double)i / (double)N;
    const double ax = on ? __dmul_rn(__dmul_rn(-2.0 * VISTAF_PI, dpx), t) : 0.0;
    const double ay = on ? __dmul_rn(__dmul_rn(-2.0 * VISTAF_PI, dpy), t) : 0.0;
    rampx[i] = make_float2((float)cos(ax), (float)sin(ax));
    rampy[i] = make_float2((float)cos(ay), (float)sin(ay));
  }
}

// ------------------------------------------------------------------ F8: reliable mask (shape_ftp.py:739-776)
struct RelRawLoad {             // roi & (q >= thr) & finite
  const float* q; const uint8_t* roi; const FtpScal* sc; long long plane; int w;
  __device__ int operator()(int b, int y, int x) const {
    long long i = (long long)y * w + x;
    if (!roi[i]) return 0;
    const float v = q[(long long)b * plane + i];
    return v >= sc[b].amp_thr && fabsf(v) <= 3.4e38f;
  }
};
struct MaskFg {                 // union-find foreground from a 0/1 plane
  const uint8_t* m; long long plane; int w;
  __device__ bool operator()(int b, int y, int x) const { return m[(long long)b * plane + (long long)y * w + x] != 0; }
};
struct LargestLoad {            // (component == largest) & roi, read from the flattened union-find plane
  const ull* uf; const uint8_t* roi; const FtpScal* sc; long long plane; int w;
  __device__ int operator()(int b, int y, int x) const {
    long long i = (long long)y * w + x;
    const unsigned int root = uf_parent(uf[(long long)b * plane + i]);
    if (root == UF_BG || !roi[i]) return 0;
    const ull best = sc[b].best_cc;
    return best != 0 && root == 0xffffffffu - (unsigned int)(best & 0xffffffffull);
  }
};
__global__ void k_ftp_seed(const float* __restrict__ q, const uint8_t* __restrict__ rel, FtpScal* sc, int npix) {
  const int b = blockIdx.y;
  ull best = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x)
    if (rel[(size_t)b * npix + i]) {
      ull k = ((ull)f2key(q[(size_t)b * npix + i]) << 32) | (ull)(0xffffffffu - (unsigned int)i);
      best = k > best ? k : best;
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { ull t = __shfl_xor_sync(0xffffffffu, best, o); best = t > best ? t : best; }
  if ((threadIdx.x & 31) == 0 && best) atomicMax(&sc[b].seed_key, best);
}

// ------------------------------------------------------------------ F10: unwrap (shape_ftp.py:1043-1080)
// u[n] = u[p] + wrap(w[n] - w[p]): wrap adds -2 pi when the step exceeds pi, +2 pi below -pi, so the wrap
// count obeys k[n] = k[p] + m(p, n).  A residue-free field makes k path independent.
struct PhaseEdge {
  const float* wr; long long plane;
  __device__ int operator()(int b, int p, int n) const {
    // np.angle(np.exp(1j * d)) in complex64: sin(float32(pi)) < 0, so a step of exactly +-float32(pi) wraps as well
    const float d = __fsub_rn(wr[(long long)b * plane + n], wr[(long long)b * plane + p]);
    return d >= 3.14159274f ? -1 : (d <= -3.14159274f ? 1 : 0);
  }
};
__global__ void k_ftp_residues(const float* __restrict__ wr, const uint8_t* __restrict__ rel, FtpScal* sc, int h, int w) {
  const int b = blockIdx.z, y = blockIdx.y, x = blockIdx.x * blockDim.x + threadIdx.x;
  int bad = 0;
  if (x + 1 < w && y + 1 < h) {
    const size_t o = (size_t)b * h * w;
    const int p00 = y * w + x, p01 = p00 + 1, p10 = p00 + w, p11 = p10 + 1;
    const bool m00 = rel[o + p00], m01 = rel[o + p01], m10 = rel[o + p10], m11 = rel[o + p11];
    // the four phases are fetched whatever the masks say: eight independent loads in flight instead of two dependent rounds
    const float w00 = wr[o + p00], w01 = wr[o + p01], w10 = wr[o + p10], w11 = wr[o + p11];
    auto e = [](float from, float to) {               // PhaseEdge on values: wrap count of the step from -> to
      const float d = __fsub_rn(to, from);
      return d >= 3.14159274f ? -1 : (d <= -3.14159274f ? 1 : 0);
    };
    if (m00 && m01 && m11) bad += (e(w00, w01) + e(w01, w11) + e(w11, w00)) != 0;
    if (m00 && m10 && m11) bad += (e(w00, w10) + e(w10, w11) + e(w11, w00)) != 0;
    if (m00 && m01 && m10) bad += (e(w00, w01) + e(w01, w10) + e(w10, w00)) != 0;
    if (m01 && m10 && m11) bad += (e(w01, w11) + e(w11, w10) + e(w10, w01)) != 0;
  }
  bad = warp_sum(bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicAdd(&sc[b].residues, (unsigned)bad);
}
__global__ void k_ftp_seed_root(const ull* __restrict__ uf, FtpScal* sc, int npix) {
  FtpScal& s = sc[blockIdx.x];
  if (s.seed_key == 0) { s.seed_root = -1; s.seed_k = 0; return; }
  const unsigned int idx = 0xffffffffu - (unsigned int)(s.seed_key & 0xffffffffull);
  const ull v = uf[(size_t)blockIdx.x * npix + idx];
  s.seed_root = (int)uf_parent(v);
  s.seed_k = uf_delta(v);
}
// F10 on frames whose wrapped field is NOT path independent (phase residues inside the reliable mask, or a loop around
// a hole of the mask that does not close): the reference's quality-guided flood itself (shape_ftp.py:1043-1080), run
// sequentially by one thread per flagged frame.  The Python heap holds tuples (-q, y, x, py, px): a pixel is fixed at
// its first pop, which happens when it is the best (largest q, then smallest raster index) pixel on the frontier, and
// the tuple popped first is the one pushed by its lexicographically smallest (py, px) already-fixed neighbour.  So one
// heap entry per frontier pixel, keyed (~key(q) << 32 | index), and a parent scan in (py, px) order reproduce the
// flood exactly.  The wrap count of a pixel is its parent's plus the wrap of the step (PhaseEdge), the field is
// written in the union-find format k_ftp_unwrap_apply reads: (root = seed pixel, delta = k).  The binary heap lives in
// the frame's union-find plane (free at this point), its top UH_TOP entries in shared memory.  Seconds per flagged
// frame -- the reference spends 32-42 s here on every frame.
#define UH_TOP 4096
__global__ void __launch_bounds__(32) k_ftp_unwrap_heap(const float* __restrict__ wr_all, const float* __restrict__ q_all,
                                                         const uint8_t* __restrict__ rel_all, FtpScal* sc, ull* __restrict__ uf_all,
                                                         unsigned int* __restrict__ aux_all, int h, int w, int force) {
  __shared__ ull hs[UH_TOP];
  const int b = blockIdx.x, lane = threadIdx.x, npix = h * w;
  FtpScal& s = sc[b];
  if (!(force || s.residues || s.uf_bad) || s.seed_key == 0) return;
  const float* wr = wr_all + (size_t)b * npix;
  const float* q = q_all + (size_t)b * npix;
  const uint8_t* rel = rel_all + (size_t)b * npix;
  ull* H = uf_all + (size_t)b * npix;
  unsigned int* A = aux_all + (size_t)b * npix;        // (k << 2) | state: 0 untouched, 1 on the frontier, 2 fixed
  for (int i = lane; i < npix; i += 32) A[i] = 0u;
  __syncwarp();
  const unsigned int seed = 0xffffffffu - (unsigned int)(s.seed_key & 0xffffffffull);
  if (lane == 0) {
    int n = 0;
    auto hget = [&](int i) -> ull { return i < UH_TOP ? hs[i] : H[i]; };
    auto hset = [&](int i, ull v) { if (i < UH_TOP) hs[i] = v; else H[i] = v; };
    auto push = [&](ull key) {
      int i = n++;
      while (i > 0) {
        const int p = (i - 1) >> 1;
        const ull pv = hget(p);
        if (pv <= key) break;
        hset(i, pv);
        i = p;
      }
      hset(i, key);
    };
    auto pop = [&]() -> ull {
      const ull top = hs[0];
      const ull last = hget(--n);
      int i = 0;
      for (;;) {
        int c = 2 * i + 1;
        if (c >= n) break;
        ull cv = hget(c);
        if (c + 1 < n) { const ull c2 = hget(c + 1); if (c2 < cv) { cv = c2; c++; } }
        if (cv >= last) break;
        hset(i, cv);
        i = c;
      }
      if (n) hset(i, last);
      return top;
    };
    auto visit = [&](int y, int x) {                     // push the untouched reliable neighbours of a fixed pixel
      for (int dy = -1; dy <= 1; dy++)
        for (int dx = -1; dx <= 1; dx++) {
          const int ny = y + dy, nx = x + dx;
          if ((dy | dx) == 0 || ny < 0 || ny >= h || nx < 0 || nx >= w) continue;
          const int j = ny * w + nx;
          if (!rel[j] || (A[j] & 3u)) continue;
          A[j] = 1u;
          push(((ull)(~f2key(__fadd_rn(q[j], 0.0f))) << 32) | (ull)(unsigned int)j);    // -0.0 and 0.0 compare equal in Python
        }
    };
    PhaseEdge e{wr, 0};
    A[seed] = 2u;
    visit((int)(seed / (unsigned)w), (int)(seed % (unsigned)w));
    while (n > 0) {
      const int i = (int)(unsigned int)(pop() & 0xffffffffull);
      const int y = i / w, x = i - y * w;
      int kp = 0, p = -1;
      for (int dy = -1; dy <= 1 && p < 0; dy++)          // first fixed neighbour in (py, px) order
        for (int dx = -1; dx <= 1; dx++) {
          const int ny = y + dy, nx = x + dx;
          if ((dy | dx) == 0 || ny < 0 || ny >= h || nx < 0 || nx >= w) continue;
          const unsigned int a = A[ny * w + nx];
          if ((a & 3u) == 2u) { p = ny * w + nx; kp = (int)a >> 2; break; }
        }
      const int k = kp + e(0, p, i);
      A[i] = ((unsigned int)k << 2) | 2u;
      visit(y, x);
    }
  }
  __syncwarp();
  for (int i = lane; i < npix; i += 32) {
    ull v = uf_pack(UF_BG, 0);
    if (rel[i]) v = (A[i] & 3u) == 2u ? uf_pack(seed, (int)A[i] >> 2) : uf_pack((unsigned int)i, 0);   // unreached: own root
    H[i] = v;
  }
}
__global__ void k_ftp_unwrap_apply(const ull* __restrict__ uf, const float* __restrict__ wr, FtpScal* sc,
                                   float* __restrict__ unw, int npix) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const ull v = uf[(size_t)b * npix + i];
  float u = FNAN;
  if (uf_parent(v) != UF_BG) {
    if ((int)uf_parent(v) == sc[b].seed_root) {
      const int dk = uf_delta(v) - sc[b].seed_k;
      u = __fadd_rn(wr[(size_t)b * npix + i], (float)(6.283185307179586 * (double)dk));
    } else if (!(*(volatile unsigned int*)&sc[b].status & VISTAF_ST_SPLIT_RELIABLE)) {
      // the erosion (:721) split the reliable mask: the flood never reaches this component (NaN, as in the reference),
      // but the internal-hole fill and output_reliable bookkeeping of such pixels (F14, :1776-1801) are not carried out
      atomicOr(&sc[b].status, VISTAF_ST_SPLIT_RELIABLE);
    }
  }
  unw[(size_t)b * npix + i] = u;
}

// ------------------------------------------------------------------ F11/F12: IRLS polynomial surface (shape_ftp.py:1100-1136)
__device__ __forceinline__ float poly_fit(const float* c, float xn, float yn, bool order2) {
  float f = __fadd_rn(__fadd_rn(__fmul_rn(c[0], xn), __fmul_rn(c[1], yn)), c[2]);
  if (order2)
    f = __fadd_rn(f, __fadd_rn(__fadd_rn(__fmul_rn(__fmul_rn(c[3], xn), xn), __fmul_rn(__fmul_rn(c[4], xn), yn)),
                               __fmul_rn(__fmul_rn(c[5], yn), yn)));
  return f;
}
struct FitGeom {
  int w; float cx, cy;           // (w-1)/2, (h-1)/2
  unsigned int magic;            // ceil(2^40 / w): i / w == (i * magic) >> 40 for i < 2^24, 256 <= w < 65536
  const float* xnt; const float* ynt;   // tables of (x - cx) / cx and (y - cy) / cy, rounded like the float32 division
  __device__ void norm(int i, float& xn, float& yn) const {
    const int y = (int)(((unsigned long long)(unsigned int)i * magic) >> 40), x = i - y * w;
    xn = __ldg(xnt + x);
    yn = __ldg(ynt + y);
  }
  __device__ unsigned int pack(int i) const {                     // (y << 16) | x
    const unsigned int y = (unsigned int)(((unsigned long long)(unsigned int)i * magic) >> 40);
    return (y << 16) | ((unsigned int)i - y * (unsigned int)w);
  }
  __device__ void norm_packed(unsigned int pk, float& xn, float& yn) const {
    xn = __ldg(xnt + (pk & 0xffffu));
    yn = __ldg(ynt + (pk >> 16));
  }
};
static FitGeom make_fitgeom(const VistafFtpPlan* p) {
  return FitGeom{p->w, (float)((p->w - 1) / 2.0), (float)((p->h - 1) / 2.0), (unsigned int)(((1ull << 40) + p->w - 1) / p->w),
                 p->xnt, p->ynt};
}
// mask selector for the three fits: 0 = reliable, 1 = background (reliable & ~contact_dilated, or reliable)
struct FitMask {
  const uint8_t* rel; const uint8_t* contact_d; const FtpScal* sc; long long plane; int use_bg;
  __device__ bool operator()(int b, long long i) const {
    if (!rel[(long long)b * plane + i]) return false;
    if (use_bg && !sc[b].bg_is_rel && contact_d[(long long)b * plane + i]) return false;
    return true;
  }
};

// Compacted sample list of a fit: the pixels that pass the fit's mask and are finite, as (value, pixel index) pairs
// in arbitrary order (every consumer is order independent: float64 sums by atomics, exact order statistics).  The
// ~30 sweeps of a fit then stream two dense arrays instead of chasing mask -> sample through a 43 % empty plane.
__global__ void __launch_bounds__(256) k_irls_compact(const float* __restrict__ z, FitMask fm, FtpScal* sc, FitGeom g,
                                                       float* __restrict__ zc, unsigned int* __restrict__ pc, int npix) {
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  if (sc[b].irls_skip) return;
  const size_t o = (size_t)b * npix;
  const int stride = gridDim.x * 1024;
  for (int i0 = blockIdx.x * 1024 + (threadIdx.x & ~31) * 4; i0 < npix; i0 += stride) {
    bool ok[4];
    float zz[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int i = i0 + j * 32 + lane;
      zz[j] = i < npix ? z[o + i] : 0.f;
      ok[j] = i < npix && fm(b, i) && fabsf(zz[j]) <= 3.4e38f;
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const unsigned int bits = __ballot_sync(0xffffffffu, ok[j]);
      if (!bits) continue;
      unsigned int base = 0;
      if (lane == 0) base = atomicAdd(&sc[b].nc, (unsigned int)__popc(bits));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (ok[j]) {
        const unsigned int at = base + (unsigned int)__popc(bits & ((1u << lane) - 1u));
        zc[o + at] = zz[j];
        pc[o + at] = g.pack(i0 + j * 32 + lane);
      }
    }
  }
}

// Solve the normal equations (Gaussian elimination with partial pivoting, float64), store float32 coefficients.
template <bool ORDER2>
__device__ void irls_solve_dev(FtpScal& s, int iter) {
  constexpr int NC = ORDER2 ? 6 : 3, NT = NC * (NC + 1) / 2;
  double A[NC][NC + 1];
  int k = 0;
  for (int p = 0; p < NC; p++)
    for (int q = p; q < NC; q++) { const double v = __ldcg(&s.ne[k]); A[p][q] = v; A[q][p] = v; k++; }
  for (int p = 0; p < NC; p++) A[p][NC] = __ldcg(&s.ne[NT + p]);
  const int n = (int)(__ldcg(&s.ne[NT + NC]) + 0.5);
  for (int i = 0; i < 28; i++) s.ne[i] = 0.0;
  if (iter == 0) {
    s.irls_n = n;
    if (n < 200) {                                   // shape_ftp.py:1103: zero surface
      for (int i = 0; i < 6; i++) s.coef[i] = 0.f;
      s.irls_skip = 1;
      s.status |= VISTAF_ST_FEW_RELIABLE;
      return;
    }
  }
  for (int c = 0; c < NC; c++) {
    int piv = c;
    for (int r = c + 1; r < NC; r++) if (fabs(A[r][c]) > fabs(A[piv][c])) piv = r;
    if (piv != c) for (int q = 0; q <= NC; q++) { double t = A[c][q]; A[c][q] = A[piv][q]; A[piv][q] = t; }
    const double d = A[c][c];
    if (fabs(d) < 1e-300) continue;
    for (int r = c + 1; r < NC; r++) {
      const double f = A[r][c] / d;
      for (int q = c; q <= NC; q++) A[r][q] -= f * A[c][q];
    }
  }
  double x[NC];
  for (int c = NC - 1; c >= 0; c--) {
    double v = A[c][NC];
    for (int q = c + 1; q < NC; q++) v -= A[c][q] * x[q];
    x[c] = fabs(A[c][c]) < 1e-300 ? 0.0 : v / A[c][c];
  }
  for (int i = 0; i < 6; i++) s.coef[i] = i < NC ? (float)x[i] : 0.f;
}

// Each block streams a contiguous range of the frame's compacted samples (value, packed pixel, previous residual)
// through a three-stage ring of 1024-sample tiles in shared memory filled by 16-byte asynchronous copies
// (cp.async), so ~36 KB per block are in flight whatever the register budget of the 28 float64 accumulators.
#define ACC_TILE 1024
#define ACC_STAGES 3
template <bool ORDER2>
__global__ void __launch_bounds__(256, 3) k_irls_accum(const float* __restrict__ zc, const unsigned int* __restrict__ pc,
                                                     const float* __restrict__ rc, FtpScal* sc, FitGeom g, int npix, int iter) {
  constexpr int NC = ORDER2 ? 6 : 3, NV = NC * (NC + 1) / 2 + NC + 1;
  __shared__ double sh[NV * 32];
  __shared__ __align__(16) float s_z[ACC_STAGES][ACC_TILE];
  __shared__ __align__(16) unsigned int s_p[ACC_STAGES][ACC_TILE];
  __shared__ __align__(16) float s_r[ACC_STAGES][ACC_TILE];
  const int b = blockIdx.y;
  if (sc[b].irls_skip) return;
  const float cs = __fmul_rn(IRLS_C, sc[b].sigma);
  const int n = (int)sc[b].nc;
  const size_t o = (size_t)b * npix;
  double acc[NV];
#pragma unroll
  for (int i = 0; i < NV; i++) acc[i] = 0.0;
  const int ntiles = (n + ACC_TILE - 1) / ACC_TILE;
  const int per = (ntiles + gridDim.x - 1) / gridDim.x;
  const int t0 = blockIdx.x * per, t1 = min(ntiles, t0 + per);
  const bool use_r = iter > 0;
  // tile t of the frame -> ring slot; every thread copies 4 consecutive samples of each array (the arrays are
  // 16-byte aligned per frame and padded to the plane size, so a tile may over-read past n inside the plane)
  auto issue = [&](int t) {
    if (t < t1 && (size_t)t * ACC_TILE + threadIdx.x * 4 + 4 <= (size_t)npix) {
      const int slot = (t - t0) % ACC_STAGES;
      const size_t e = o + (size_t)t * ACC_TILE + threadIdx.x * 4;
      __pipeline_memcpy_async(&s_z[slot][threadIdx.x * 4], zc + e, 16);
      __pipeline_memcpy_async(&s_p[slot][threadIdx.x * 4], pc + e, 16);
      if (use_r) __pipeline_memcpy_async(&s_r[slot][threadIdx.x * 4], rc + e, 16);
    }
    __pipeline_commit();
  };
  issue(t0);
  issue(t0 + 1);
  for (int t = t0; t < t1; t++) {
    issue(t + 2);
    __pipeline_wait_prior(2);
    __syncthreads();
    const int slot = (t - t0) % ACC_STAGES;
    const int jbase = t * ACC_TILE;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int e = threadIdx.x + j * 256;
      if (jbase + e >= n) continue;
      const float zz = s_z[slot][e];
      float xn, yn;
      g.norm_packed(s_p[slot][e], xn, yn);
      float wt = 1.0f;
      if (use_r) {
        const float u = __fdiv_rn(s_r[slot][e], cs);
        wt = __fdiv_rn(1.0f, __fadd_rn(1.0f, __fmul_rn(u, u)));
      }
      // products and sums in float64: the sums then agree to ~1e-16 whatever the grid / batch layout, so the float32
      // coefficients (and every mask derived from them) do not depend on how a batch is chunked
      const double w2 = (double)wt * (double)wt;
      double a[NC];
      a[0] = xn; a[1] = yn; a[2] = 1.0;
      if (ORDER2) { a[3] = (double)__fmul_rn(xn, xn); a[4] = (double)__fmul_rn(xn, yn); a[5] = (double)__fmul_rn(yn, yn); }
      int k = 0;
#pragma unroll
      for (int p = 0; p < NC; p++) {
        const double wa = w2 * a[p];
#pragma unroll
        for (int q = p; q < NC; q++) acc[k++] += wa * a[q];
      }
#pragma unroll
      for (int p = 0; p < NC; p++) acc[k++] += w2 * a[p] * (double)zz;
      acc[k] += 1.0;
    }
    __syncthreads();                                   // slot is refilled by the next issue()
  }
  __pipeline_wait_prior(0);
  block_sum<NV>(acc, sh);
  if (threadIdx.x == 0) {
    for (int i = 0; i < NV; i++)
      if (acc[i] != 0.0) atomicAdd(&sc[b].ne[i], acc[i]);
    __threadfence();
    if (atomicAdd(&sc[b].irls_ticket, 1u) == gridDim.x - 1) {      // last block of this frame: solve
      __threadfence();
      sc[b].irls_ticket = 0;
      irls_solve_dev<ORDER2>(sc[b], iter);
    }
  }
}

__global__ void k_irls_begin(FtpScal* sc, int mode) {    // mode 0: plane on reliable, 1: quadric on reliable, 2: quadric on background
  FtpScal& s = sc[blockIdx.x];
  for (int i = 0; i < 28; i++) s.ne[i] = 0.0;
  for (int i = 0; i < 6; i++) s.coef[i] = 0.f;
  s.sigma = 1.f; s.med = 0.f;
  s.irls_skip = 0;
  s.irls_ticket = 0;
  s.nc = 0;
  if (mode == 0) {
    s.detrend1 = s.n_rel >= 500 ? 1 : 0;                 // shape_ftp.py:1379
    if (!s.detrend1) s.irls_skip = 1;
  }
}
// Residual streams of the current fit over its compacted samples.  RFirst evaluates r = z - fit once per solve (first
// sweep of the median selection) and stores it; the other sweeps of the iteration (RPlain: r, RAbs: |r - med|, and the
// next accumulate) re-read the 4-byte residual instead of the sample, its pixel index and the polynomial.
template <bool ORDER2>
struct RFirst {
  const float* zc; const unsigned int* pc; float* rc; const FtpScal* sc; FitGeom g; long long plane;
  __device__ long long limit(int b) const { return (long long)sc[b].nc; }
  __device__ bool operator()(int b, long long j, float& v) const {
    const long long o = (long long)b * plane + j;
    float xn, yn;
    g.norm_packed(pc[o], xn, yn);
    v = __fsub_rn(zc[o], poly_fit(sc[b].coef, xn, yn, ORDER2));
    rc[o] = v;
    return true;
  }
  // 16-byte path (select.cuh): the compacted arrays are padded to the plane, so a group may over-read past n inside it
  __device__ bool vec_ok() const { return (plane & 3) == 0 && sel_aligned16(zc) && sel_aligned16(pc) && sel_aligned16(rc); }
  __device__ unsigned int vec4(int b, long long j, long long n, float v[4]) const {
    const long long o = (long long)b * plane + j;
    const float4 z4 = *reinterpret_cast<const float4*>(zc + o);
    const uint4 p4 = *reinterpret_cast<const uint4*>(pc + o);
    const unsigned int m = sel_tail_mask(j, n);
    const float zz[4] = {z4.x, z4.y, z4.z, z4.w};
    const unsigned int pp[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
      v[k] = 0.f;
      if ((m >> k) & 1u) {
        float xn, yn;
        g.norm_packed(pp[k], xn, yn);
        v[k] = __fsub_rn(zz[k], poly_fit(sc[b].coef, xn, yn, ORDER2));
      }
    }
    *reinterpret_cast<float4*>(rc + o) = make_float4(v[0], v[1], v[2], v[3]);
    return m;
  }
};
struct RPlain {
  const float* rc; const FtpScal* sc; long long plane;
  __device__ long long limit(int b) const { return (long long)sc[b].nc; }
  __device__ bool operator()(int b, long long j, float& v) const { v = rc[(long long)b * plane + j]; return true; }
  __device__ bool vec_ok() const { return (plane & 3) == 0 && sel_aligned16(rc); }
  __device__ unsigned int vec4(int b, long long j, long long n, float v[4]) const {
    const float4 r4 = *reinterpret_cast<const float4*>(rc + (long long)b * plane + j);
    v[0] = r4.x; v[1] = r4.y; v[2] = r4.z; v[3] = r4.w;
    return sel_tail_mask(j, n);
  }
};
struct RAbs {
  const float* rc; const FtpScal* sc; long long plane;
  __device__ long long limit(int b) const { return (long long)sc[b].nc; }
  __device__ bool operator()(int b, long long j, float& v) const {
    v = fabsf(__fsub_rn(rc[(long long)b * plane + j], sc[b].med));
    return true;
  }
  __device__ bool vec_ok() const { return (plane & 3) == 0 && sel_aligned16(rc); }
  __device__ unsigned int vec4(int b, long long j, long long n, float v[4]) const {
    const float4 r4 = *reinterpret_cast<const float4*>(rc + (long long)b * plane + j);
    const float med = sc[b].med;
    v[0] = fabsf(__fsub_rn(r4.x, med)); v[1] = fabsf(__fsub_rn(r4.y, med));
    v[2] = fabsf(__fsub_rn(r4.z, med)); v[3] = fabsf(__fsub_rn(r4.w, med));
    return sel_tail_mask(j, n);
  }
};
__global__ void k_ftp_sub_fit1(float* __restrict__ unw, const FtpScal* sc, FitGeom g, int npix) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix || !sc[b].detrend1) return;
  float xn, yn;
  g.norm(i, xn, yn);
  const size_t o = (size_t)b * npix + i;
  unw[o] = __fsub_rn(unw[o], poly_fit(sc[b].coef, xn, yn, false));
}

// ------------------------------------------------------------------ F12: contact / background (shape_ftp.py:1716-1751)
struct AbsResVal {              // |unw - fit0| over reliable & finite
  const float* z; const uint8_t* rel; const FtpScal* sc; FitGeom g; long long plane; int need_q2;
  __device__ bool operator()(int b, long long i, float& v) const {
    if (need_q2 && !(sc[b].q2 == sc[b].q2)) return false;
    if (!rel[(long long)b * plane + i]) return false;
    const float zz = z[(long long)b * plane + i];
    float xn, yn;
    g.norm((int)i, xn, yn);
    v = fabsf(__fsub_rn(zz, poly_fit(sc[b].coef, xn, yn, true)));
    return fabsf(v) <= 3.4e38f;
  }
};
__global__ void k_ftp_contact(const float* __restrict__ z, const uint8_t* __restrict__ rel, FtpScal* sc, FitGeom g,
                              uint8_t* __restrict__ contact, int npix, int count) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  int c = 0;
  if (i < npix) {
    const size_t o = (size_t)b * npix + i;
    const uint8_t r = rel[o];
    const float zz = z[o];                             // fetched beside the mask byte, not after it
    if (r) {
      float xn, yn;
      g.norm(i, xn, yn);
      const float v = fabsf(__fsub_rn(zz, poly_fit(sc[b].coef, xn, yn, true)));
      c = (v >= sc[b].contact_thr && fabsf(v) <= 3.4e38f) ? 1 : 0;
    }
    contact[o] = (uint8_t)c;
  }
  if (count) {
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&sc[b].n_contact, (unsigned)c);
  }
}
__global__ void k_ftp_contact_frac(FtpScal* sc, float min_frac, float max_frac) {
  FtpScal& s = sc[blockIdx.x];
  const double frac = (double)s.n_contact / (double)(s.n_rel > 1u ? s.n_rel : 1u);
  s.q2 = FNAN;
  if (frac < (double)min_frac) { s.q2 = 95.f; s.status |= VISTAF_ST_CONTACT_P95; }
  else if (frac > (double)max_frac) { s.q2 = 98.f; s.status |= VISTAF_ST_CONTACT_P95; }
}
__global__ void k_ftp_bg_decide(FtpScal* sc) {
  FtpScal& s = sc[blockIdx.x];
  const unsigned int nbg = s.n_rel - s.n_contact_d;
  s.n_bg = nbg;
  s.bg_is_rel = nbg < (unsigned int)(int)(0.15 * (double)s.n_rel) ? 1 : 0;     // shape_ftp.py:1744
  if (s.bg_is_rel) s.status |= VISTAF_ST_BG_FALLBACK;
}
struct DetVal {                 // det = unw - fit over background (median, :1749)
  const float* z; FitMask fm; const FtpScal* sc; FitGeom g; long long plane;
  __device__ bool operator()(int b, long long i, float& v) const {
    if (!fm(b, i)) return false;
    float xn, yn;
    g.norm((int)i, xn, yn);
    v = __fsub_rn(z[(long long)b * plane + i], poly_fit(sc[b].coef, xn, yn, true));
    return fabsf(v) <= 3.4e38f;
  }
};

// ------------------------------------------------------------------ F13: smoothing, sign (shape_ftp.py:1139-1147, 1755-1768)
__device__ __forceinline__ float hm_at(const float* z, const FtpScal& s, const FitGeom& g, long long o, int i) {
  float xn, yn;
  g.norm(i, xn, yn);
  return __fsub_rn(__fsub_rn(z[o], poly_fit(s.coef, xn, yn, true)), s.bg_med);
}
struct HmNumLoad {              // z0 = hm where reliable & finite, else 0
  const float* z; const uint8_t* rel; const FtpScal* sc; FitGeom g; long long plane; int w;
  __device__ float operator()(int b, int y, int x) const {
    const int i = y * w + x;
    const long long o = (long long)b * plane + i;
    if (!rel[o]) return 0.f;
    const float v = hm_at(z, sc[b], g, o, i);
    return fabsf(v) <= 3.4e38f ? v : 0.f;
  }
};
struct HmDenLoad {
  const float* z; const uint8_t* rel; long long plane; int w;
  __device__ float operator()(int b, int y, int x) const {
    const long long o = (long long)b * plane + (long long)y * w + x;
    return (rel[o] && fabsf(z[o]) <= 3.4e38f) ? 1.f : 0.f;
  }
};
struct RatioStore {             // out = num / (den + 1e-6)
  const float* num; float* out; long long plane; int w;
  __device__ void operator()(int b, int y, int x, float den) {
    const long long o = (long long)b * plane + (long long)y * w + x;
    out[o] = __fdiv_rn(num[o], __fadd_rn(den, 1e-6f));
  }
  __device__ void finish(int) {}
};
struct CoreVal {                // hm over reliable & hm <= core_thr
  const float* hm; const uint8_t* rel; const FtpScal* sc; long long plane;
  __device__ bool operator()(int b, long long i, float& v) const {
    if (!rel[(long long)b * plane + i]) return false;
    if (!(sc[b].core_thr == sc[b].core_thr)) return false;
    v = hm[(long long)b * plane + i];
    return v <= sc[b].core_thr;
  }
};

// ------------------------------------------------------------------ F14: internal holes of the reliable mask (shape_ftp.py:1153-1203, 1776-1801)
// holes = reliable & ~isfinite(height).  A hole pixel is a candidate when >= frac_thr of the reliable pixels in its k x k
// box (cv2.boxFilter, BORDER_REFLECT_101, un-normalised) have a height and it lies >= min_dist (3-4 chamfer) inside the
// reliable mask.  Candidates are Telea-filled (r = 5) from tmp = height with the other holes set to the median of the
// known heights, on a canvas that is the median of the non-candidate reliable pixels elsewhere (inpaint_only_mask).
// Every kernel returns at once for frames without holes (all frames of the shipped configuration: with the reliable
// smoothing on, masked_gaussian_smooth leaves no NaN inside the mask), so the stage costs a few empty launches.
__global__ void k_holes_count(const float* __restrict__ hm, const uint8_t* __restrict__ rel, FtpScal* sc, int npix) {
  const int b = blockIdx.y;
  int c = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
    const size_t o = (size_t)b * npix + i;
    c += rel[o] && !(fabsf(hm[o]) <= 3.4e38f);
  }
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&sc[b].n_holes, (unsigned)c);
}
struct HoleGate {               // frames with holes (chamfer / union-find `active` convention: < 0 = skip)
  const FtpScal* sc;
  __device__ bool on(int b) const { return sc[b].n_holes != 0; }
};
__global__ void k_holes_active(const FtpScal* sc, int* active, int which) {
  const FtpScal& s = sc[blockIdx.x];
  active[blockIdx.x] = (which == 0 ? s.n_holes : s.n_cand) ? (int)blockIdx.x : -1;
}
__global__ void k_holes_candidates(const float* __restrict__ hm, const uint8_t* __restrict__ rel, const float* __restrict__ dist,
                                   FtpScal* sc, uint8_t* __restrict__ cand, int h, int w, int k, float frac_thr, float min_dist) {
  const int b = blockIdx.z, y = blockIdx.y, x = blockIdx.x * blockDim.x + threadIdx.x;
  if (sc[b].n_holes == 0) return;
  int c = 0;
  if (x < w) {
    const size_t o = (size_t)b * h * w, i = (size_t)y * w + x;
    if (rel[o + i] && !(fabsf(hm[o + i]) <= 3.4e38f) && dist[o + i] >= min_dist) {
      float ck = 0.f, cc = 0.f;
      const int r = k / 2;
      for (int dy = -r; dy <= r; dy++) {
        const size_t row = (size_t)reflect101(y + dy, h) * w;
        for (int dx = -r; dx <= r; dx++) {
          const size_t j = o + row + reflect101(x + dx, w);
          if (rel[j]) { cc += 1.f; if (fabsf(hm[j]) <= 3.4e38f) ck += 1.f; }
        }
      }
      c = __fdiv_rn(ck, __fadd_rn(cc, 1e-6f)) >= frac_thr;
    }
    cand[o + i] = (uint8_t)c;
  }
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&sc[b].n_cand, (unsigned)c);
}
struct HoleKnownVal {           // heights of the known pixels (median #1), frames with candidates only
  const float* hm; const uint8_t* rel; const FtpScal* sc; long long plane;
  __device__ long long limit(int b) const { return sc[b].n_cand ? plane : 0; }
  __device__ bool operator()(int b, long long i, float& v) const {
    if (sc[b].n_cand == 0 || !rel[(long long)b * plane + i]) return false;
    v = hm[(long long)b * plane + i];
    return fabsf(v) <= 3.4e38f;
  }
};
struct HoleTmpVal {             // tmp over reliable & ~candidate: known heights, other holes = hole_med (median #2)
  const float* hm; const uint8_t* rel; const uint8_t* cand; const FtpScal* sc; long long plane;
  __device__ long long limit(int b) const { return sc[b].n_cand ? plane : 0; }
  __device__ bool operator()(int b, long long i, float& v) const {
    const long long o = (long long)b * plane + i;
    if (sc[b].n_cand == 0 || !rel[o] || cand[o]) return false;
    v = hm[o];
    if (!(fabsf(v) <= 3.4e38f)) v = sc[b].hole_med;
    return true;
  }
};
struct HoleGrab { FtpScal* sc; int which;
  __device__ void operator()(int b, float r, unsigned int cnt) const {
    if (which == 0) sc[b].hole_med = (cnt && r == r) ? r : 0.f; else sc[b].hole_fill = (cnt && r == r) ? r : 0.f;
  }
};
__global__ void k_holes_canvas(const float* __restrict__ hm, const uint8_t* __restrict__ rel, const uint8_t* __restrict__ cand,
                               const FtpScal* sc, float* __restrict__ zin, int npix) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (sc[b].n_cand == 0 || i >= npix) return;
  const size_t o = (size_t)b * npix + i;
  float v = sc[b].hole_fill;
  if (rel[o] && !cand[o]) { v = hm[o]; if (!(fabsf(v) <= 3.4e38f)) v = sc[b].hole_med; }
  zin[o] = v;
}
struct HoleMask {
  const uint8_t* cand; const FtpScal* sc; long long plane; int w;
  __device__ bool operator()(int b, int y, int x) const { return sc[b].n_cand != 0 && cand[(long long)b * plane + (long long)y * w + x] != 0; }
};
// height[cand] = filled; output_reliable = reliable & isfinite(height)
__global__ void k_holes_apply(float* __restrict__ hm, const uint8_t* __restrict__ rel, const uint8_t* __restrict__ cand,
                              const float* __restrict__ zin, FtpScal* sc, uint8_t* __restrict__ outrel, int npix) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  int c = 0;
  if (i < npix) {
    const size_t o = (size_t)b * npix + i;
    if (sc[b].n_cand && cand[o]) hm[o] = zin[o];
    c = rel[o] && fabsf(hm[o]) <= 3.4e38f;
    outrel[o] = (uint8_t)c;
  }
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&sc[b].n_outrel, (unsigned)c);
  if (i == 0 && sc[b].n_cand) atomicOr(&sc[b].status, VISTAF_ST_HOLES_FILLED);
}
// sigma <= 0: masked_gaussian_smooth returns its input (shape_ftp.py:1140): hm = unw - fit - median on reliable, NaN elsewhere
__global__ void k_ftp_hm_plain(const float* __restrict__ z, const uint8_t* __restrict__ rel, const FtpScal* sc, FitGeom g,
                               float* __restrict__ hm, int npix) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const size_t o = (size_t)b * npix + i;
  hm[o] = rel[o] ? hm_at(z, sc[b], g, (long long)o, i) : FNAN;
}

// ------------------------------------------------------------------ F15: frontier, assembly (shape_ftp.py:1287-1330, 1803-1841)
struct NotMaskZero {            // chamfer zero set = pixels outside the mask
  const uint8_t* m; long long plane; int w;
  __device__ bool operator()(int b, int y, int x) const { return m[(long long)b * plane + (long long)y * w + x] == 0; }
};
struct MaskZero {               // chamfer zero set = the mask
  const uint8_t* m; long long plane; int w;
  __device__ bool operator()(int b, int y, int x) const { return m[(long long)b * plane + (long long)y * w + x] != 0; }
};
__device__ __forceinline__ float filled_val(float v, float d, const FtpScal& s, float band) {
  if (s.flip) v = __fmul_rn(v, -1.0f);
  float t = __fdiv_rn(fmaxf(__fsub_rn(d, 1.0f), 0.f), fmaxf(1e-6f, band));
  t = fminf(fmaxf(t, 0.f), 1.f);
  const float wgt = __fmul_rn(__fmul_rn(t, t), __fsub_rn(3.0f, __fmul_rn(2.0f, t)));
  return __fmul_rn(v, wgt);
}
__device__ __forceinline__ float filled_at(const float* hm, const float* dist_in, const FtpScal& s, long long o, float band) {
  const float v = hm[o], d = dist_in[o];
  return filled_val(v, d, s, band);
}
struct AsmLoad {                // hf_ on the ROI before the sigma-9 smoothing: reliable -> filled, else 0
  const float* hm; const float* dist_in; const uint8_t* rel; const uint8_t* roi; const FtpScal* sc; long long plane; int w;
  float band;
  __device__ float operator()(int b, int y, int x) const {
    const long long i = (long long)y * w + x, o = (long long)b * plane + i;
    const uint8_t in_roi = roi[i], in_rel = rel[o];
    if (!in_roi || !in_rel) return 0.f;
    return filled_at(hm, dist_in, sc[b], o, band);
  }
};
struct Den9Store {              // sm = num / den9 (den9 = blur(roi) + 1e-6, frame independent)
  const float* den9; float* out; long long plane; int w;
  __device__ void operator()(int b, int y, int x, float num) {
    const long long i = (long long)y * w + x;
    out[(long long)b * plane + i] = __fdiv_rn(num, den9[i]);
  }
  __device__ void finish(int) {}
};
struct RoiLoad {
  const uint8_t* roi; int w;
  __device__ float operator()(int, int y, int x) const { return roi[(long long)y * w + x] ? 1.f : 0.f; }
};
struct AddEpsStore {
  float* out; int w;
  __device__ void operator()(int, int y, int x, float v) { out[(long long)y * w + x] = __fadd_rn(v, 1e-6f); }
  __device__ void finish(int) {}
};

// final per-pixel assembly + phase -> mm (shape_ftp.py:1816-1851) + candidate statistics for the blob filter
__global__ void __launch_bounds__(256, 8) k_ftp_assemble(const float* __restrict__ hm, const float* __restrict__ dist_in,
                               const float* __restrict__ dist_out, const float* __restrict__ sm,
                               const uint8_t* __restrict__ rel, const uint8_t* __restrict__ roi, FtpScal* sc,
                               float* __restrict__ out, int npix, float band, VistafCurveModel model, int use_neg) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  float mm = FNAN;
  if (i < npix) {
    const size_t o = (size_t)b * npix + i;
    const uint8_t in_roi = roi[i], in_rel = rel[o];       // both mask bytes in flight, then the pixel's pair of planes together
    if (in_roi && !(sc[b].status & VISTAF_ST_EMPTY_ROI)) {
      const float a = (in_rel ? hm : sm)[o], d = (in_rel ? dist_in : dist_out)[o];
      float v;
      if (in_rel) v = filled_val(a, d, sc[b], band);
      else {
        v = a;
        if (fmaxf(__fsub_rn(d, 1.0f), 0.f) <= band) v = 0.f;
      }
      if (fabsf(v) <= 3.4e38f) v = fminf(v, 0.f);
      const double x = fmax((double)(use_neg ? -v : v), 0.0);
      mm = (float)curve_eval(model, x);
    }
    out[o] = mm;
  }
  float cand = (mm > 0.f) ? mm : 0.f;
  cand = warp_max(cand);
  if ((threadIdx.x & 31) == 0 && cand > 0.f) atomicMax(&sc[b].gmax_bits, __float_as_uint(cand));
}

// ------------------------------------------------------------------ F17: blob filter (shape_ftp.py:1215-1281)
struct CandFg {
  const float* mm; long long plane; int w;
  __device__ bool operator()(int b, int y, int x) const { return mm[(long long)b * plane + (long long)y * w + x] > 0.f; }
};
__global__ void k_ftp_blob_peak(const ull* __restrict__ uf, const float* __restrict__ mm, unsigned int* __restrict__ peak,
                                int npix) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned int root = UF_BG, bits = 0;
  if (i < npix) {
    const ull u = uf[(size_t)b * npix + i];
    const float depth = mm[(size_t)b * npix + i];                             // fetched beside the label, not after it
    root = uf_parent(u);
    if (root != UF_BG) bits = __float_as_uint(depth);                         // depth > 0: bit order == value order
  }
  // lanes of one blob elect a leader that carries the group's maximum: one atomic per blob per warp
  const unsigned int grp = __match_any_sync(0xffffffffu, root);
  const unsigned int mx = __reduce_max_sync(grp, bits);
  if (root != UF_BG && (int)(__ffs(grp) - 1) == (int)(threadIdx.x & 31)) atomicMax(peak + (size_t)b * npix + root, mx);
}
__global__ void k_ftp_blob_apply(const ull* __restrict__ uf, const unsigned int* __restrict__ peak, FtpScal* sc,
                                 float* __restrict__ mm, uint8_t* __restrict__ kept, int npix, double min_peak, double rel_frac) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  int k = 0;
  if (i < npix) {
    const unsigned int root = uf_parent(uf[(size_t)b * npix + i]);
    if (root != UF_BG) {
      const double gmax = (double)__uint_as_float(sc[b].gmax_bits);
      const double thr = fmax(min_peak, rel_frac * gmax);
      k = (double)__uint_as_float(peak[(size_t)b * npix + root]) >= thr;
      if (!k) mm[(size_t)b * npix + i] = 0.f;
    }
    if (kept) kept[(size_t)b * npix + i] = (uint8_t)k;
  }
  k = warp_sum(k);
  if ((threadIdx.x & 31) == 0 && k) atomicAdd(&sc[b].n_kept, (unsigned)k);
}

// ------------------------------------------------------------------ P1: force reduction (force_sensor.py:93-123)
__global__ void __launch_bounds__(256) k_depth_sign(const float* __restrict__ z, const uint8_t* roi, double* pos, double* neg,
                                                     size_t stride, long long npix) {
  __shared__ double sh[64];
  const int b = blockIdx.y;
  double v[2] = {0.0, 0.0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (long long)gridDim.x * blockDim.x) {
    const float x = z[(size_t)b * npix + i];
    if (x == x) { if (x > 0.f) v[0] += (double)x; else v[1] += (double)(-x); }   // nansum over the whole map
  }
  (void)roi;
  block_sum<2>(v, sh);
  if (threadIdx.x == 0) {
    if (v[0] != 0.0) atomicAdd((double*)((char*)pos + (size_t)b * stride), v[0]);
    if (v[1] != 0.0) atomicAdd((double*)((char*)neg + (size_t)b * stride), v[1]);
  }
}
__global__ void __launch_bounds__(256) k_depth_reduce(const float* __restrict__ z, const uint8_t* roi, const double* pos,
                                                       const double* neg, size_t stride, double* vsum, unsigned int* vcnt,
                                                       unsigned int* vmax, long long npix, float eps) {
  __shared__ double sh[32];
  __shared__ int shc[32];
  __shared__ float shm[32];
  const int b = blockIdx.y;
  const bool use_neg = *(const double*)((const char*)neg + (size_t)b * stride) > *(const double*)((const char*)pos + (size_t)b * stride);
  double v[1] = {0.0};
  int c = 0;
  float mx = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (long long)gridDim.x * blockDim.x) {
    float x = z[(size_t)b * npix + i];
    if (roi && !roi[i]) continue;
    x = use_neg ? -x : x;
    if (!(x > 0.f) || !(fabsf(x) <= 3.4e38f)) continue;      // clip at 0, non-finite -> 0
    if (x > eps) { v[0] += (double)x; c++; mx = fmaxf(mx, x); }
  }
  block_sum<1>(v, sh);
  c = warp_sum(c); mx = warp_max(mx);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) { shc[threadIdx.x >> 5] = c; shm[threadIdx.x >> 5] = mx; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int ct = 0; float m = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); i++) { ct += shc[i]; m = fmaxf(m, shm[i]); }
    if (v[0] != 0.0) atomicAdd((double*)((char*)vsum + (size_t)b * stride), v[0]);
    if (ct) atomicAdd((unsigned int*)((char*)vcnt + (size_t)b * stride), (unsigned)ct);
    if (m > 0.f) atomicMax((unsigned int*)((char*)vmax + (size_t)b * stride), __float_as_uint(m));
  }
}

__global__ void k_ftp_records(const FtpScal* sc, VistafFtpRecord* rec, int n, int w, const PeakResult* pr, int N, double pitch_mm,
                              VistafCurveModel force_model, int count_outrel) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const FtpScal& s = sc[i];
  const double period_px = fabs(pr->kx) > 1e-9 ? (double)N / fabs(pr->kx) : 0.0;
  VistafFtpRecord r;
  memset(&r, 0, sizeof(r));
  const double mm_per_px = pitch_mm / period_px;
  const double a = mm_per_px * mm_per_px;
  r.mm_per_px = mm_per_px; r.period_px = period_px;
  if (s.vol_cnt) {
    const float v32 = __fmul_rn((float)s.vol_sum, (float)a);        // np.sum(float32) * float -> float32 (force_sensor.py:120)
    r.volume_cm3 = (double)v32 / 1000.0;
    r.contact_area_mm2 = (double)s.vol_cnt * a;
    r.max_depth_mm = (double)__uint_as_float(s.vol_max_bits);
  }
  r.force_N = curve_eval(force_model, r.volume_cm3);
  r.amp_thr = s.amp_thr; r.contact_thr = s.contact_thr; r.bg_median = s.bg_med; r.core_median = s.core_med;
  r.reliable_px = (int)s.n_rel; r.output_reliable_px = (int)(count_outrel ? s.n_outrel : s.n_rel); r.contact_dilated_px = (int)s.n_contact_d;
  r.contact_kept_px = (int)s.n_kept; r.bad_px = (int)s.n_bad; r.residues = (int)s.residues;
  if (s.seed_key) {
    const unsigned int idx = 0xffffffffu - (unsigned int)(s.seed_key & 0xffffffffull);
    r.seed_x = (int)(idx % (unsigned)w); r.seed_y = (int)(idx / (unsigned)w);
  } else { r.seed_x = r.seed_y = -1; }
  unsigned int st = s.status;
  if (s.n_rel == 0) st |= VISTAF_ST_EMPTY_ROI;
  if (!pr->ok) st |= VISTAF_ST_NO_PEAK;
  if (s.residues || s.uf_bad) st |= VISTAF_ST_RESIDUES;
  if (s.flip) st |= VISTAF_ST_SIGN_FLIPPED;
  r.status = (int)st;
  rec[i] = r;
}
__global__ void k_ftp_mark_empty(FtpScal* sc) {
  FtpScal& s = sc[blockIdx.x];
  if (s.n_rel == 0) s.status |= VISTAF_ST_EMPTY_ROI;
}

}  // namespace vftp
using namespace vftp;

// ================================================================== host side
static void ftp_layout(const VistafFtpPlan* p, int chunk, FtpWs& L) {
  const size_t S = (size_t)p->h * p->w, c = (size_t)chunk;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o += align256(bytes); return r; };
  L.img = take(c * S * 4); L.grad = take(c * S * 4); L.bad0 = take(c * S); L.bad = take(c * S);
  L.tmp = take(c * S * 4); L.I = take(c * S * 4); L.Iw = take(c * S * 4);
  L.R = take(c * (size_t)p->h * p->np * 8); L.Pw = take(c * (size_t)p->np * p->np * 8);
  L.wrapped = take(c * S * 4); L.qprod = take(c * S * 4); L.quality = take(c * S * 4);
  L.m1 = take(c * S); L.m2 = take(c * S); L.reliable = take(c * S);
  L.uf = take(c * S * 8); L.area = take(c * S * 4); L.unw = take(c * S * 4);
  L.contact = take(c * S); L.contact_d = take(c * S);
  L.hnum = take(c * S * 4); L.hm = take(c * S * 4); L.dist_in = take(c * S * 4); L.dist_out = take(c * S * 4);
  L.dtmp = take(c * S * 4); L.snum = take(c * S * 4); L.sm = take(c * S * 4); L.peak = take(c * S * 4);
  L.scal = take(c * sizeof(FtpScal)); L.sel = take(sel_ws_bytes((int)c));
  L.telea = take(telea_ws_bytes(p->h, p->w, chunk));
  L.active = take(c * 4 * 2);
  L.total = o;
}

static MorphSE se_chamfer_ball(float margin) {       // {(dx,dy): chamfer distance <= margin} as per-row half widths
  MorphSE s;
  int R = (int)floorf(margin / 0.955f) + 1;
  if (R > MORPH_MAXROWS / 2) R = MORPH_MAXROWS / 2;
  s.rows = 2 * R + 1;
  for (int i = 0; i < MORPH_MAXROWS; i++) s.half[i] = -1;
  for (int i = 0; i < s.rows; i++) {
    const int dy = abs(i - R);
    int best = -1;
    for (int dx = 0; dx <= R + 1; dx++) {
      const unsigned mn = dx < dy ? dx : dy, mx = dx < dy ? dy : dx;
      const float d = (float)(CH_HV * (mx - mn) + CH_DG * mn) * (1.0f / 65536.0f);
      if (d <= margin) best = dx;
    }
    s.half[i] = best;
  }
  return s;
}

extern "C" int vistaf_ftp_plan_create(const VistafFtpParams* P, VistafFtpPlan** out) {
  if (!P || !out) return vistaf_set_err("null argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  const int np = 2 * (P->patch_half < 3 ? 3 : P->patch_half) + 1;
  if (np > NPATCH_MAX || P->crop_w > CH_THREADS * CH_MAXPT || (P->dilate_kernel | 1) > MORPH_MAXROWS ||
      (P->close_kernel | 1) > MORPH_MAXROWS || P->edge_margin > 15 || P->crop_w < 256 ||
      (long long)P->crop_w * P->crop_h >= (1ll << 24) || P->bad_inpaint_radius < 1 || P->bad_inpaint_radius > 8 ||
      (P->fill_internal_holes && (P->hole_inpaint_radius < 1 || P->hole_inpaint_radius > 8 || P->hole_neighborhood > 31)))
    return vistaf_set_err("unsupported FTP parameters", cudaErrorInvalidValue, __FILE__, __LINE__);
  VistafFtpPlan* p = new (std::nothrow) VistafFtpPlan();
  if (!p) return vistaf_set_err("out of host memory", cudaErrorMemoryAllocation, __FILE__, __LINE__);
  memset(p, 0, sizeof(*p));
  p->P = *P;
  p->h = P->crop_h; p->w = P->crop_w; p->np = np;
  p->N = P->crop_w + 2 * P->fft_pad;
  if (P->crop_h != P->crop_w) return vistaf_set_err("FTP crop must be square", cudaErrorInvalidValue, __FILE__, __LINE__);
  if (fft1d_create(&p->fN, p->N)) return 1;
  {
    std::vector<float> tx(p->w), ty(p->h);
    volatile float cxv = (float)((p->w - 1) / 2.0), cyv = (float)((p->h - 1) / 2.0);
    for (int x = 0; x < p->w; x++) { volatile float d = (float)x - cxv; volatile float q = d / cxv; tx[x] = q; }
    for (int y = 0; y < p->h; y++) { volatile float d = (float)y - cyv; volatile float q = d / cyv; ty[y] = q; }
    CK(cudaMalloc(&p->xnt, sizeof(float) * p->w));
    CK(cudaMalloc(&p->ynt, sizeof(float) * p->h));
    CK(cudaMemcpy(p->xnt, tx.data(), sizeof(float) * p->w, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(p->ynt, ty.data(), sizeof(float) * p->h, cudaMemcpyHostToDevice));
  }
  p->g_illum = make_gauss(P->illum_sigma);
  p->g_pre = make_gauss(P->pre_blur_sigma);
  p->g_qual = make_gauss(P->quality_sigma);
  p->g_smooth = make_gauss(P->smooth_sigma > 0.f ? P->smooth_sigma : 1.0);
  p->g_unrel = make_gauss(P->unrel_sigma);
  p->se_bad = se_ellipse(P->bad_dilate < 3 ? 3 : (P->bad_dilate | 1));
  p->se_close = se_ellipse(P->close_kernel < 3 ? 3 : (P->close_kernel | 1));
  p->se_dil = se_ellipse(P->dilate_kernel);
  p->se_edge = se_chamfer_ball((float)P->edge_margin);
  const size_t S = (size_t)p->h * p->w;
  CK(cudaMalloc(&p->roi, S)); CK(cudaMalloc(&p->circ, S)); CK(cudaMalloc(&p->valid, S));
  CK(cudaMalloc(&p->apo, S * 4)); CK(cudaMalloc(&p->den9, S * 4));
  CK(cudaMalloc(&p->cref, S * 8)); CK(cudaMalloc(&p->amp_ref, S * 4));
  CK(cudaMalloc(&p->rampx, sizeof(float2) * p->N)); CK(cudaMalloc(&p->rampy, sizeof(float2) * p->N));
  CK(cudaMalloc(&p->hann, sizeof(float) * np)); CK(cudaMalloc(&p->kxs, sizeof(int) * np)); CK(cudaMalloc(&p->kys, sizeof(int) * np));
  CK(cudaMalloc(&p->peak_keys, sizeof(ull) * 32)); CK(cudaMalloc(&p->peak_result, 64));
  int r_valid = P->r - P->roi_erode_px;
  if (r_valid < 0) r_valid = 0;
  VL(), k_ftp_geometry<<<dim3(cdiv(p->w, 256), p->h), 256>>>(p->roi, p->circ, p->valid, p->apo, p->h, p->w, P->cx, P->cy, P->r,
                                                       r_valid, P->apod_taper);
  CKL();
  // den9 = GaussianBlur(roi, sigma 9) + 1e-6 (shape_ftp.py:1146 with mask = roi_eroded, :1820)
  float* tmp = nullptr;
  CK(cudaMalloc(&tmp, S * 4));
  if (gauss_blur(RoiLoad{p->roi, p->w}, AddEpsStore{p->den9, p->w}, tmp, p->h, p->w, 1, p->g_unrel, p->g_unrel, 0)) return 1;
  CK(cudaDeviceSynchronize());
  cudaFree(tmp);
  std::vector<float> hn(np);
  for (int i = 0; i < np; i++) hn[i] = (float)(0.5 - 0.5 * cos(2.0 * VISTAF_PI * i / (double)(np - 1)));   // np.hanning
  CK(cudaMemcpy(p->hann, hn.data(), sizeof(float) * np, cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute((const void*)k_ftp_coldft, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(p->h * 8)));
  *out = p;
  return 0;
}

extern "C" void vistaf_ftp_plan_destroy(VistafFtpPlan* p) {
  if (!p) return;
  vistaf_graph_drop(p);
  fft1d_destroy(&p->fN);
  cudaFree(p->roi); cudaFree(p->circ); cudaFree(p->valid); cudaFree(p->apo); cudaFree(p->den9);
  cudaFree(p->cref); cudaFree(p->amp_ref); cudaFree(p->rampx); cudaFree(p->rampy); cudaFree(p->hann);
  cudaFree(p->xnt); cudaFree(p->ynt);
  cudaFree(p->kxs); cudaFree(p->kys); cudaFree(p->peak_keys); cudaFree(p->peak_result);
  delete p;
}

extern "C" size_t vistaf_ftp_workspace_bytes(const VistafFtpPlan* p, int chunk) {
  FtpWs L;
  ftp_layout(p, chunk < 1 ? 1 : chunk, L);
  // the reference pass additionally needs the dense padded plane + spectrum
  size_t dense = align256((size_t)p->N * p->N * 4) + align256((size_t)p->N * p->N * 8);
  return L.total + dense;
}

extern "C" int vistaf_ftp_masks(const VistafFtpPlan* p, uint8_t* roi, uint8_t* circ, float* apo, void* stream) {
  const size_t S = (size_t)p->h * p->w;
  if (roi) CK(cudaMemcpyAsync(roi, p->roi, S, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  if (circ) CK(cudaMemcpyAsync(circ, p->circ, S, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  if (apo) CK(cudaMemcpyAsync(apo, p->apo, S * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}

extern "C" int vistaf_ftp_workspace_plane(const VistafFtpPlan* p, int chunk, const char* name, size_t* off, size_t* bytes) {
  FtpWs L;
  ftp_layout(p, chunk, L);
  const size_t S = (size_t)p->h * p->w;
  struct { const char* n; size_t o, b; } tab[] = {
      {"img", L.img, S * 4}, {"grad", L.grad, S * 4}, {"bad", L.bad, S}, {"I", L.I, S * 4}, {"Iw", L.Iw, S * 4},
      {"R", L.R, (size_t)p->h * p->np * 8}, {"Pw", L.Pw, (size_t)p->np * p->np * 8}, {"wrapped", L.wrapped, S * 4},
      {"qprod", L.qprod, S * 4}, {"quality", L.quality, S * 4}, {"m1", L.m1, S}, {"m2", L.m2, S},
      {"reliable", L.reliable, S}, {"uf", L.uf, S * 8}, {"unw", L.unw, S * 4}, {"contact", L.contact, S},
      {"contact_d", L.contact_d, S}, {"hm", L.hm, S * 4}, {"dist_in", L.dist_in, S * 4}, {"dist_out", L.dist_out, S * 4},
      {"sm", L.sm, S * 4}, {"scal", L.scal, sizeof(FtpScal)}};
  for (auto& t : tab)
    if (!strcmp(t.n, name)) { *off = t.o; *bytes = t.b; return 0; }
  return vistaf_set_err("unknown workspace plane", cudaErrorInvalidValue, __FILE__, __LINE__);
}

// Stage A up to the apodised, normalised image Iw and its median (shared by reference and deformed frames).
static int ftp_preprocess(VistafFtpPlan* p, const uint8_t* gray, int nb, char* ws, const FtpWs& L, cudaStream_t st,
                          bool ref = false) {       // ref: the reference frame's pass reports under R.* stage names
  const VistafFtpParams& P = p->P;
  const int h = p->h, w = p->w, npix = h * w;
  const long long plane = npix;
  float* img = (float*)(ws + L.img); float* grad = (float*)(ws + L.grad);
  uint8_t* bad0 = (uint8_t*)(ws + L.bad0); uint8_t* bad = (uint8_t*)(ws + L.bad);
  float* tmp = (float*)(ws + L.tmp); float* I = (float*)(ws + L.I); float* Iw = (float*)(ws + L.Iw);
  FtpScal* sc = (FtpScal*)(ws + L.scal);
  SelectState* sel = (SelectState*)(ws + L.sel);
  if (sel_reset(sel, nb, st)) return 1;
  vistaf_stage_mark(ref ? "R.sobel_percentiles_bad" : "F.sobel_percentiles_bad", st);
  VL(), k_ftp_init<<<cdiv(nb, 128), 128, 0, st>>>(sc, nb);
  VL(), k_ftp_sobel<<<dim3(cdiv(w, 256), h, nb), 256, 0, st>>>(gray, img, grad, h, w);
  CKL();
  if (select_grab(MaskedVal{img, p->valid, plane}, sc, G_HI, sel, npix, nb, P.bad_int_pct, st, nullptr, 0, true)) return 1;   // 8-bit levels
  if (select_grab(MaskedVal{grad, p->valid, plane}, sc, G_GRAD, sel, npix, nb, P.bad_grad_pct, st, nullptr, 0, true)) return 1;
  BadLoad bl{img, grad, p->valid, sc, plane, w};
  if (P.bad_dilate > 1) {
    if (morph<true>(bl, PostId(), bad, h, w, nb, p->se_bad, st)) return 1;
  } else {
    VL(), k_ftp_badraw<<<dim3(cdiv(w, 256), h, nb), 256, 0, st>>>(bl, bad, h, w);
  }
  (void)bad0;
  VL(), k_ftp_count_mask<<<dim3(64, nb), 256, 0, st>>>(bad, sc, npix, 0);
  CKL();
  vistaf_stage_mark(ref ? "R.telea_inpaint" : "F.telea_inpaint", st);
  if (telea_inpaint(TeleaMask{bad, plane, w}, img, (size_t)npix, &sc[0].status, sizeof(FtpScal), ws + L.telea, h, w, nb,
                    P.bad_inpaint_radius, st))
    return 1;
  vistaf_stage_mark(ref ? "R.illum_blur45" : "F.illum_blur45", st);
  if (gauss_blur(PlaneLoad{img, plane, w}, IllumStoreF{img, I, plane, w}, tmp, h, w, nb, p->g_illum, p->g_illum, st)) return 1;
  vistaf_stage_mark(ref ? "R.preblur_apo_median" : "F.preblur_apo_median", st);
  if (gauss_blur(PlaneLoad{I, plane, w}, ApoStore{p->apo, Iw, plane, w}, tmp, h, w, nb, p->g_pre, p->g_pre, st)) return 1;
  if (select_grab(MaskedVal{Iw, p->valid, plane}, sc, G_MU, sel, npix, nb, -1.0f, st)) return 1;
  CKL();
  return 0;
}

static int ftp_patch_field(VistafFtpPlan* p, int nb, char* ws, const FtpWs& L, bool ref, cudaStream_t st) {
  const int h = p->h, w = p->w, N = p->N, pad = p->P.fft_pad, np = p->np;
  float* Iw = (float*)(ws + L.Iw);
  float2* R = (float2*)(ws + L.R); float2* Pw = (float2*)(ws + L.Pw);
  FtpScal* sc = (FtpScal*)(ws + L.scal);
  vistaf_stage_mark(ref ? "R.carrier_dft_field" : "F.carrier_dft_field", st);
  VL(), k_ftp_rowdft<<<dim3(h, nb), 256, (size_t)N * 4, st>>>(Iw, sc, R, p->fN.tw, p->kxs, h, w, N, pad, np);
  VL(), k_ftp_coldft<<<dim3(np, nb), 256, (size_t)h * 8, st>>>(R, Pw, p->fN.tw, p->kys, p->hann, h, N, pad, np);
  if (ref)
    VL(), k_ftp_field<true><<<dim3(h, nb), 256, 0, st>>>(Pw, p->fN.tw, p->rampx, p->rampy, p->cref, p->amp_ref, nullptr, nullptr, h, w, N, pad, np);
  else
    VL(), k_ftp_field<false><<<dim3(h, nb), 256, 0, st>>>(Pw, p->fN.tw, p->rampx, p->rampy, p->cref, p->amp_ref,
                                                    (float*)(ws + L.wrapped), (float*)(ws + L.qprod), h, w, N, pad, np);
  CKL();
  return 0;
}

static int ftp_reference_issue(VistafFtpPlan* p, const uint8_t* ref_gray, char* ws, const FtpWs& L, cudaStream_t st) {
  const int N = p->N;
  if (ftp_preprocess(p, ref_gray, 1, ws, L, st, true)) return 1;
  float* Ip = (float*)(ws + L.total);
  float2* F = (float2*)(ws + L.total + align256((size_t)N * N * 4));
  vistaf_stage_mark("R.dense_fft_carrier", st);
  VL(), k_ftp_pad<<<dim3(cdiv(N, 256), N), 256, 0, st>>>((float*)(ws + L.Iw), (FtpScal*)(ws + L.scal), Ip, p->h, p->w, N, p->P.fft_pad);
  CKL();
  if (fft2d_dense(Ip, F, p->fN, p->fN, N, N, 1, -1, true, st)) return 1;
  // carrier search on the device: n_peaks sweeps of the spectrum, then choice + log-parabolic refinement
  const int npk = p->P.n_peaks < 1 ? 1 : (p->P.n_peaks > 32 ? 32 : p->P.n_peaks);
  CK(cudaMemsetAsync(p->peak_keys, 0, sizeof(ull) * 32, st));
  for (int r = 0; r < npk; r++) VL(), k_ftp_peak_round<<<296, 256, 0, st>>>(F, p->peak_keys, r, N, p->P.dc_exclusion);
  VL(), k_ftp_peak_choose<<<1, 1, 0, st>>>(F, p->peak_keys, npk, N, p->P.peak_max_dy_frac, (PeakResult*)p->peak_result);
  CKL();
  VL(), k_ftp_carrier_setup<<<cdiv(N > p->np ? N : p->np, 256), 256, 0, st>>>((PeakResult*)p->peak_result, p->kxs, p->kys, p->rampx,
                                                                        p->rampy, N, p->np);
  CKL();
  if (ftp_patch_field(p, 1, ws, L, true, st)) return 1;
  vistaf_stage_mark("", st);
  return 0;
}

extern "C" int vistaf_ftp_set_reference(VistafFtpPlan* p, const uint8_t* ref_gray, void* workspace, size_t ws_bytes,
                                        VistafFtpReferenceInfo* info, void* stream) {
  if (!p || !ref_gray || !workspace) return vistaf_set_err("vistaf_ftp_set_reference: null argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  cudaStream_t st = (cudaStream_t)stream;
  FtpWs L;
  ftp_layout(p, 1, L);
  const int N = p->N;
  const size_t dense = align256((size_t)N * N * 4) + align256((size_t)N * N * 8);
  if (L.total + dense > ws_bytes) return vistaf_set_err("vistaf_ftp_set_reference: workspace too small", cudaErrorInvalidValue, __FILE__, __LINE__);
  // the carrier stays on the device, so the call is asynchronous (and replayable) unless the caller asks for `info`
  const VistafGraphKey key = vistaf_graph_key(1, p, ref_gray, workspace, ws_bytes);
  if (vistaf_graph_run(key, st, [&](cudaStream_t s) { return ftp_reference_issue(p, ref_gray, (char*)workspace, L, s); })) return 1;
  p->have_ref = true;
  if (info) {
    PeakResult pr;
    FtpScal hs;
    CK(cudaMemcpyAsync(&pr, p->peak_result, sizeof(pr), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&hs, (char*)workspace + L.scal, sizeof(FtpScal), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (!pr.ok) return vistaf_set_err("no carrier peak (or its patch touches the spectrum border)", cudaErrorInvalidValue, __FILE__, __LINE__);
    info->peak_x = pr.pxf; info->peak_y = pr.pyf; info->fft_h = N; info->fft_w = N;
    info->kx = (float)pr.kx; info->ky = (float)pr.ky;
    info->period_px = fabs(pr.kx) > 1e-9 ? (float)((double)N / fabs(pr.kx)) : 0.f;
    info->bad_px = (int)hs.n_bad;
  }
  return 0;
}

// The fit runs over groups of IRLS_GROUP frames: one group's samples and masks (7 MB per frame) stay resident in the
// 126 MB L2 across the ~30 sweeps of the six iterations, so only the first sweep of a group reads HBM.  Per
// iteration: accumulate (+ solve in the last block), median of the residuals, median of |r - med| (3 launches each).
#define IRLS_GROUP_DEFAULT 64
static int irls_group() {
  static int g = 0;
  if (!g) { const char* e = getenv("VISTAF_IRLS_GROUP"); g = e ? atoi(e) : IRLS_GROUP_DEFAULT; if (g < 1) g = 1; }
  return g;
}
template <bool ORDER2>
static int irls_fit(VistafFtpPlan* p, const float* z, FitMask fm, FtpScal* sc, SelectState* sel, float* zc, unsigned int* pc,
                    float* rc, int nb, int mode, cudaStream_t st) {
  const int npix = p->h * p->w;
  const FitGeom g = make_fitgeom(p);
  unsigned int* lists = sel_lists(sel, nb);
  VL(), k_irls_begin<<<nb, 1, 0, st>>>(sc, mode);
  const int G = irls_group();
  for (int g0 = 0; g0 < nb; g0 += G) {
    const int ng = nb - g0 < G ? nb - g0 : G;
    const float* zg = z + (size_t)g0 * npix;
    float* zcg = zc + (size_t)g0 * npix;
    unsigned int* pcg = pc + (size_t)g0 * npix;
    float* rcg = rc + (size_t)g0 * npix;
    FtpScal* scg = sc + g0;
    SelectState* selg = sel + g0;
    unsigned int* listg = lists + (size_t)g0 * SEL_CAP;
    FitMask fmg = fm;
    fmg.rel += (size_t)g0 * fm.plane; fmg.contact_d += (size_t)g0 * fm.plane; fmg.sc = scg;
    int gx = (148 * 16 + ng - 1) / ng;           // ~8 waves of blocks: the last partial wave costs little
    if (gx > 128) gx = 128;
    VL(), k_irls_compact<<<dim3(gx, ng), 256, 0, st>>>(zg, fmg, scg, g, zcg, pcg, npix);
    for (int it = 0; it < 6; it++) {
      VL(), k_irls_accum<ORDER2><<<dim3(gx, ng), 256, 0, st>>>(zcg, pcg, rcg, scg, g, npix, it);
      if (it == 5) break;
      RFirst<ORDER2> r1{zcg, pcg, rcg, scg, g, (long long)npix};
      RPlain r2{rcg, scg, (long long)npix};
      RAbs r3{rcg, scg, (long long)npix};
      if (select_quantile_out2(r1, r2, GrabOut{scg, G_IRLS_MED}, selg, listg, npix, ng, -1.0f, st)) return 1;
      if (select_quantile_out(r3, GrabOut{scg, G_IRLS_MAD}, selg, listg, npix, ng, -1.0f, st)) return 1;
    }
  }
  CKL();
  return 0;
}

// F14 on a batch: hm [nb][h][w] in / out (NaN = no height), rel in, outrel = output_reliable out, cand = hole candidates
// (written for every frame only when cand_is_output).  Scratch: zin, dist (float planes), dtmp (uint plane), the Telea
// workspace, two int arrays of nb entries.
static int ftp_fill_holes(VistafFtpPlan* p, float* hm, const uint8_t* rel, uint8_t* outrel, uint8_t* cand, bool cand_is_output,
                          float* zin, float* dist, unsigned int* dtmp, char* telea_ws, int* active, FtpScal* sc, SelectState* sel,
                          int nb, cudaStream_t st, bool search = true) {   // search = false: the caller knows there is no hole
  const VistafFtpParams& P = p->P;
  const int h = p->h, w = p->w, npix = h * w;
  const long long plane = npix;
  const dim3 gpix(cdiv(npix, 256), nb);
  int* act_holes = active;
  int* act_cand = active + nb;
  if (P.fill_internal_holes && search) {
    VL(), k_holes_count<<<dim3(64, nb), 256, 0, st>>>(hm, rel, sc, npix);
    VL(), k_holes_active<<<nb, 1, 0, st>>>(sc, act_holes, 0);
    if (cand_is_output) CK(cudaMemsetAsync(cand, 0, (size_t)npix * nb, st));
    if (chamfer_dt(NotMaskZero{rel, plane, w}, dist, dtmp, h, w, nb, st, act_holes)) return 1;
    const int k = (P.hole_neighborhood < 3 ? 3 : P.hole_neighborhood) | 1;
    VL(), k_holes_candidates<<<dim3(cdiv(w, 256), h, nb), 256, 0, st>>>(hm, rel, dist, sc, cand, h, w, k, P.hole_known_frac,
                                                                    P.hole_min_dist);
    VL(), k_holes_active<<<nb, 1, 0, st>>>(sc, act_cand, 1);
    unsigned int* lists = sel_lists(sel, nb);
    if (select_quantile_out(HoleKnownVal{hm, rel, sc, plane}, HoleGrab{sc, 0}, sel, lists, npix, nb, -1.0f, st)) return 1;
    if (select_quantile_out(HoleTmpVal{hm, rel, cand, sc, plane}, HoleGrab{sc, 1}, sel, lists, npix, nb, -1.0f, st)) return 1;
    VL(), k_holes_canvas<<<gpix, 256, 0, st>>>(hm, rel, cand, sc, zin, npix);
    if (telea_inpaint(HoleMask{cand, sc, plane, w}, zin, (size_t)npix, &sc[0].status, sizeof(FtpScal), telea_ws, h, w, nb,
                      P.hole_inpaint_radius, st, 0, act_cand))
      return 1;
  }
  VL(), k_holes_apply<<<gpix, 256, 0, st>>>(hm, rel, cand, zin, sc, outrel, npix);
  CKL();
  return 0;
}

static int ftp_chunk(VistafFtpPlan* p, const uint8_t* gray_all, int nb, const VistafFtpOutputs& O, size_t f0, char* ws,
                     const FtpWs& L, cudaStream_t st, bool preprocessed = false) {
  const VistafFtpParams& P = p->P;
  const int h = p->h, w = p->w, npix = h * w;
  const long long plane = npix;
  const uint8_t* gray = gray_all + f0 * (size_t)npix;
  float* tmp = (float*)(ws + L.tmp);
  float* wrapped = (float*)(ws + L.wrapped); float* qprod = (float*)(ws + L.qprod); float* quality = (float*)(ws + L.quality);
  uint8_t* m1 = (uint8_t*)(ws + L.m1); uint8_t* m2 = (uint8_t*)(ws + L.m2);
  uint8_t* rel = O.reliable ? O.reliable + f0 * (size_t)npix : (uint8_t*)(ws + L.reliable);
  ull* uf = (ull*)(ws + L.uf); unsigned int* area = (unsigned int*)(ws + L.area);
  float* unw = (float*)(ws + L.unw);
  uint8_t* contact = (uint8_t*)(ws + L.contact);
  uint8_t* contact_d = O.contact_dilated ? O.contact_dilated + f0 * (size_t)npix : (uint8_t*)(ws + L.contact_d);
  float* hnum = (float*)(ws + L.hnum); float* hm = (float*)(ws + L.hm);
  float* dist_in = (float*)(ws + L.dist_in); float* dist_out = (float*)(ws + L.dist_out);
  unsigned int* dtmp = (unsigned int*)(ws + L.dtmp);
  float* sm = (float*)(ws + L.sm);
  unsigned int* peak = (unsigned int*)(ws + L.peak);
  FtpScal* sc = (FtpScal*)(ws + L.scal);
  SelectState* sel = (SelectState*)(ws + L.sel);
  float* out = O.height_mm + f0 * (size_t)npix;
  const FitGeom g = make_fitgeom(p);
  const dim3 gpix(cdiv(npix, 256), nb);

  // ---- A: demodulation
  if (!preprocessed && ftp_preprocess(p, gray, nb, ws, L, st)) return 1;
  if (ftp_patch_field(p, nb, ws, L, false, st)) return 1;
  vistaf_stage_mark("F.quality_reliable_mask", st);
  // ---- B: reliable mask
  if (gauss_blur(PlaneLoad{qprod, plane, w}, PlaneStore{quality, plane, w}, tmp, h, w, nb, p->g_qual, p->g_qual, st)) return 1;
  if (select_grab(MaskedVal{quality, p->roi, plane}, sc, G_AMP, sel, npix, nb, P.amp_percentile, st)) return 1;
  RelRawLoad rl{quality, p->roi, sc, plane, w};
  if (morph<true>(rl, PostId(), m1, h, w, nb, p->se_close, st)) return 1;                       // close = dilate, erode
  if (morph<false>(MaskLoad{m1, plane, w}, PostAnd{p->roi, 0, w}, m2, h, w, nb, p->se_close, st)) return 1;
  if (uf_label(MaskFg{m2, plane, w}, EdgeZero(), uf, area, (unsigned int*)nullptr, 0, h, w, nb, st)) return 1;
  VL(), k_uf_largest<<<gpix, 256, 0, st>>>(uf, area, &sc[0].best_cc, sizeof(FtpScal), npix);
  LargestLoad ll{uf, p->roi, sc, plane, w};
  if (P.edge_margin > 0) {
    if (morph<false>(ll, PostId(), rel, h, w, nb, p->se_edge, st)) return 1;                    // chamfer distance > margin
  } else {
    MorphSE one = se_rect(1, 1);
    if (morph<false>(ll, PostId(), rel, h, w, nb, one, st)) return 1;
  }
  VL(), k_ftp_count_mask<<<dim3(64, nb), 256, 0, st>>>(rel, sc, npix, 1);
  VL(), k_ftp_mark_empty<<<nb, 1, 0, st>>>(sc);
  VL(), k_ftp_seed<<<dim3(64, nb), 256, 0, st>>>(quality, rel, sc, npix);
  CKL();
  vistaf_stage_mark("F.unwrap_unionfind", st);
  // ---- C: unwrap
  VL(), k_ftp_residues<<<dim3(cdiv(w, 256), h, nb), 256, 0, st>>>(wrapped, rel, sc, h, w);
  if (uf_label(MaskFg{rel, plane, w}, PhaseEdge{wrapped, plane}, uf, (unsigned int*)nullptr, &sc[0].uf_bad, sizeof(FtpScal), h, w, nb, st))
    return 1;
  // frames with residues (or a loop around a mask hole that does not close): the reference's sequential flood
  VL(), k_ftp_unwrap_heap<<<nb, 32, 0, st>>>(wrapped, quality, rel, sc, uf, area, h, w, p->debug_flags & 1);
  VL(), k_ftp_seed_root<<<nb, 1, 0, st>>>(uf, sc, npix);
  VL(), k_ftp_unwrap_apply<<<gpix, 256, 0, st>>>(uf, wrapped, sc, unw, npix);
  CKL();
  vistaf_stage_mark("F.irls_plane", st);
  // ---- D: plane removal, two-pass quadric detrend
  FitMask fm_rel{rel, contact_d, sc, plane, 0}, fm_bg{rel, contact_d, sc, plane, 1};
  if (irls_fit<false>(p, unw, fm_rel, sc, sel, hnum, (unsigned int*)hm, dist_in, nb, 0, st)) return 1;
  vistaf_stage_mark("F.irls_quadric_contact", st);
  VL(), k_ftp_sub_fit1<<<gpix, 256, 0, st>>>(unw, sc, g, npix);
  if (irls_fit<true>(p, unw, fm_rel, sc, sel, hnum, (unsigned int*)hm, dist_in, nb, 1, st)) return 1;
  if (select_grab(AbsResVal{unw, rel, sc, g, plane, 0}, sc, G_CONTACT, sel, npix, nb, P.contact_pct, st)) return 1;
  VL(), k_ftp_contact<<<gpix, 256, 0, st>>>(unw, rel, sc, g, contact, npix, 1);
  VL(), k_ftp_contact_frac<<<nb, 1, 0, st>>>(sc, P.min_contact_frac, P.max_contact_frac);
  if (select_grab(AbsResVal{unw, rel, sc, g, plane, 1}, sc, G_CONTACT2, sel, npix, nb, 0.f, st, &sc[0].q2, sizeof(FtpScal))) return 1;
  VL(), k_ftp_contact<<<gpix, 256, 0, st>>>(unw, rel, sc, g, contact, npix, 0);
  {
    const uint8_t* src = contact;
    for (int it = 0; it < P.dilate_iters; it++) {
      uint8_t* dst = (it == P.dilate_iters - 1) ? contact_d : ((it & 1) ? m2 : m1);
      if (it == P.dilate_iters - 1) { if (morph<true>(MaskLoad{src, plane, w}, PostAnd{rel, plane, w}, dst, h, w, nb, p->se_dil, st)) return 1; }
      else if (morph<true>(MaskLoad{src, plane, w}, PostId(), dst, h, w, nb, p->se_dil, st)) return 1;
      src = dst;
    }
    if (P.dilate_iters <= 0) {
      MorphSE one = se_rect(1, 1);
      if (morph<true>(MaskLoad{contact, plane, w}, PostAnd{rel, plane, w}, contact_d, h, w, nb, one, st)) return 1;
    }
  }
  VL(), k_ftp_count_mask<<<dim3(64, nb), 256, 0, st>>>(contact_d, sc, npix, 3);
  VL(), k_ftp_bg_decide<<<nb, 1, 0, st>>>(sc);
  vistaf_stage_mark("F.irls_quadric_background", st);
  if (irls_fit<true>(p, unw, fm_bg, sc, sel, hnum, (unsigned int*)hm, dist_in, nb, 2, st)) return 1;
  if (select_grab(DetVal{unw, fm_bg, sc, g, plane}, sc, G_BGMED, sel, npix, nb, -1.0f, st)) return 1;
  CKL();
  vistaf_stage_mark("F.smooth_sign", st);
  // ---- E: masked smoothing (sigma 2.5) and sign
  if (P.smooth_sigma > 0.f) {
    if (gauss_blur(HmNumLoad{unw, rel, sc, g, plane, w}, PlaneStore{hnum, plane, w}, tmp, h, w, nb, p->g_smooth, p->g_smooth, st)) return 1;
    if (gauss_blur(HmDenLoad{unw, rel, plane, w}, RatioStore{hnum, hm, plane, w}, tmp, h, w, nb, p->g_smooth, p->g_smooth, st)) return 1;
  } else {
    VL(), k_ftp_hm_plain<<<gpix, 256, 0, st>>>(unw, rel, sc, g, hm, npix);
  }
  if (select_grab(MaskedValB{hm, rel, plane}, sc, G_CORETHR, sel, npix, nb, P.core_pct, st)) return 1;
  if (select_grab(CoreVal{hm, rel, sc, plane}, sc, G_COREMED, sel, npix, nb, -1.0f, st)) return 1;
  CKL();
  vistaf_stage_mark("F.internal_holes", st);
  // ---- F14: internal holes (gated: no frame of the shipped configuration has any), output_reliable
  // with the reliable smoothing on, masked_gaussian_smooth (num / (den + 1e-6)) is finite on every pixel: no hole can
  // exist, output_reliable IS reliable and the stage is one copy of the mask when the caller asked for it
  const uint8_t* outrel = rel;
  if (P.smooth_sigma > 0.f) {
    if (O.hole_candidates) CK(cudaMemsetAsync(O.hole_candidates + f0 * (size_t)npix, 0, (size_t)npix * nb, st));
    if (O.output_reliable && O.output_reliable + f0 * (size_t)npix != rel)
      CK(cudaMemcpyAsync(O.output_reliable + f0 * (size_t)npix, rel, (size_t)npix * nb, cudaMemcpyDeviceToDevice, st));
  } else {
    uint8_t* orl = O.output_reliable ? O.output_reliable + f0 * (size_t)npix : m2;
    uint8_t* cand = O.hole_candidates ? O.hole_candidates + f0 * (size_t)npix : contact;
    if (ftp_fill_holes(p, hm, rel, orl, cand, O.hole_candidates != nullptr, hnum, dist_in, dtmp, ws + L.telea, (int*)(ws + L.active),
                       sc, sel, nb, st, true))
      return 1;
    outrel = orl;
  }
  vistaf_stage_mark("F.chamfer_frontier_assemble", st);
  // ---- F: frontier taper, assembly, sigma-9 fill of the unreliable part, phase -> mm (on output_reliable, :1803-1841)
  if (chamfer_dt2(NotMaskZero{outrel, plane, w}, dist_in, dtmp, MaskZero{outrel, plane, w}, dist_out, (unsigned int*)sm, h, w, nb, st))
    return 1;                                     // sm (written by the sigma-9 blur below) doubles as the second scratch plane
  if (gauss_blur(AsmLoad{hm, dist_in, outrel, p->roi, sc, plane, w, P.frontier_band}, Den9Store{p->den9, sm, plane, w}, tmp, h, w,
                 nb, p->g_unrel, p->g_unrel, st))
    return 1;
  VL(), k_ftp_assemble<<<gpix, 256, 0, st>>>(hm, dist_in, dist_out, sm, outrel, p->roi, sc, out, npix, P.frontier_band,
                                       P.height_model, P.use_negated_height);
  CKL();
  vistaf_stage_mark("F.blob_filter_force", st);
  // ---- G: blob filter + force reduction
  if (uf_label(CandFg{out, plane, w}, EdgeZero(), uf, (unsigned int*)nullptr, (unsigned int*)nullptr, 0, h, w, nb, st)) return 1;
  CK(cudaMemsetAsync(peak, 0, sizeof(unsigned int) * (size_t)npix * nb, st));
  VL(), k_ftp_blob_peak<<<gpix, 256, 0, st>>>(uf, out, peak, npix);
  uint8_t* kept = O.contact_kept ? O.contact_kept + f0 * (size_t)npix : nullptr;
  VL(), k_ftp_blob_apply<<<gpix, 256, 0, st>>>(uf, peak, sc, out, kept, npix, P.blob_min_peak_mm, P.blob_rel_frac);
  VL(), k_depth_sign<<<dim3(64, nb), 256, 0, st>>>(out, nullptr, &sc[0].pos_sum, &sc[0].neg_sum, sizeof(FtpScal), npix);
  VL(), k_depth_reduce<<<dim3(64, nb), 256, 0, st>>>(out, nullptr, &sc[0].pos_sum, &sc[0].neg_sum, sizeof(FtpScal), &sc[0].vol_sum,
                                               &sc[0].vol_cnt, &sc[0].vol_max_bits, npix, (float)P.depth_eps_mm);
  VL(), k_ftp_records<<<cdiv(nb, 128), 128, 0, st>>>(sc, O.records + f0, nb, w, (const PeakResult*)p->peak_result, p->N,
                                                P.grating_pitch_mm, P.force_model, P.smooth_sigma > 0.f ? 0 : 1);
  CKL();
  vistaf_stage_mark("", st);
  return 0;
}

static unsigned long long out_hash(const VistafFtpOutputs& o) {       // FNV-1a over the output pointers
  unsigned long long h = 1469598103934665603ull;
  const unsigned char* c = (const unsigned char*)&o;
  for (size_t i = 0; i < sizeof(o); i++) { h ^= c[i]; h *= 1099511628211ull; }
  return h;
}

extern "C" int vistaf_ftp_batch(VistafFtpPlan* p, const uint8_t* def_gray, int batch, const VistafFtpOutputs* out,
                                void* workspace, size_t ws_bytes, void* stream) {
  if (!p || !def_gray || !out || !out->height_mm || !out->records || !workspace)
    return vistaf_set_err("vistaf_ftp_batch: null argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  if (!p->have_ref) return vistaf_set_err("vistaf_ftp_batch: call vistaf_ftp_set_reference first", cudaErrorInvalidValue, __FILE__, __LINE__);
  FtpWs L;
  int chunk = batch;
  while (chunk >= 1) { ftp_layout(p, chunk, L); if (L.total <= ws_bytes) break; chunk = chunk > 8 ? chunk - chunk / 8 : chunk - 1; }
  if (chunk < 1) return vistaf_set_err("vistaf_ftp_batch: workspace too small for one frame", cudaErrorInvalidValue, __FILE__, __LINE__);
  const VistafFtpOutputs O = *out;
  const VistafGraphKey key = vistaf_graph_key(2, p, def_gray, batch, out_hash(O), O.height_mm, O.records, workspace, ws_bytes, p->debug_flags);
  return vistaf_graph_run(key, (cudaStream_t)stream, [&](cudaStream_t st) {
    FtpWs Lc = L;
    for (int f0 = 0; f0 < batch; f0 += chunk) {
      int nb = batch - f0 < chunk ? batch - f0 : chunk;
      if (nb != chunk) ftp_layout(p, nb, Lc);
      if (ftp_chunk(p, def_gray, nb, O, (size_t)f0, (char*)workspace, Lc, st)) return 1;
    }
    return 0;
  });
}

// Split form of vistaf_ftp_batch for ONE chunk: _begin runs the part of the chain that does not need the reference
// frame (bad pixels, Telea fill, illumination normalisation, apodisation, median), so a caller can issue it before /
// beside vistaf_ftp_set_reference; _finish runs the rest on the same workspace.
extern "C" int vistaf_ftp_batch_begin(VistafFtpPlan* p, const uint8_t* def_gray, int batch, void* workspace, size_t ws_bytes,
                                      void* stream) {
  if (!p || !def_gray || !workspace || batch < 1)
    return vistaf_set_err("vistaf_ftp_batch_begin: bad argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  FtpWs L;
  ftp_layout(p, batch, L);
  if (L.total > ws_bytes) return vistaf_set_err("vistaf_ftp_batch_begin: workspace too small for the batch in one chunk", cudaErrorInvalidValue, __FILE__, __LINE__);
  const VistafGraphKey key = vistaf_graph_key(3, p, def_gray, batch, workspace, ws_bytes);
  return vistaf_graph_run(key, (cudaStream_t)stream,
                          [&](cudaStream_t st) { return ftp_preprocess(p, def_gray, batch, (char*)workspace, L, st); });
}
extern "C" int vistaf_ftp_batch_finish(VistafFtpPlan* p, const uint8_t* def_gray, int batch, const VistafFtpOutputs* out,
                                       void* workspace, size_t ws_bytes, void* stream) {
  if (!p || !def_gray || !out || !out->height_mm || !out->records || !workspace || batch < 1)
    return vistaf_set_err("vistaf_ftp_batch_finish: bad argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  if (!p->have_ref) return vistaf_set_err("vistaf_ftp_batch_finish: call vistaf_ftp_set_reference first", cudaErrorInvalidValue, __FILE__, __LINE__);
  FtpWs L;
  ftp_layout(p, batch, L);
  if (L.total > ws_bytes) return vistaf_set_err("vistaf_ftp_batch_finish: workspace too small for the batch in one chunk", cudaErrorInvalidValue, __FILE__, __LINE__);
  const VistafFtpOutputs O = *out;
  const VistafGraphKey key = vistaf_graph_key(4, p, def_gray, batch, out_hash(O), O.height_mm, O.records, workspace, ws_bytes, p->debug_flags);
  return vistaf_graph_run(key, (cudaStream_t)stream,
                          [&](cudaStream_t st) { return ftp_chunk(p, def_gray, batch, O, 0, (char*)workspace, L, st, true); });
}

extern "C" int vistaf_ftp_set_debug(VistafFtpPlan* p, int flags) {
  if (!p) return vistaf_set_err("vistaf_ftp_set_debug: null plan", cudaErrorInvalidValue, __FILE__, __LINE__);
  p->debug_flags = flags;
  return 0;
}

// F14 on caller planes (see include/vistaf.h)
extern "C" int vistaf_ftp_fill_holes(VistafFtpPlan* p, float* height, const uint8_t* reliable, int batch, uint8_t* output_reliable,
                                     uint8_t* hole_candidates, void* workspace, size_t ws_bytes, void* stream) {
  if (!p || !height || !reliable || !output_reliable || !hole_candidates || !workspace || batch < 1)
    return vistaf_set_err("vistaf_ftp_fill_holes: bad argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  FtpWs L;
  ftp_layout(p, batch, L);
  if (L.total > ws_bytes) return vistaf_set_err("vistaf_ftp_fill_holes: workspace too small", cudaErrorInvalidValue, __FILE__, __LINE__);
  char* ws = (char*)workspace;
  cudaStream_t st = (cudaStream_t)stream;
  FtpScal* sc = (FtpScal*)(ws + L.scal);
  SelectState* sel = (SelectState*)(ws + L.sel);
  if (sel_reset(sel, batch, st)) return 1;
  VL(), k_ftp_init<<<cdiv(batch, 128), 128, 0, st>>>(sc, batch);
  return ftp_fill_holes(p, height, reliable, output_reliable, hole_candidates, true, (float*)(ws + L.hnum), (float*)(ws + L.dist_in),
                        (unsigned int*)(ws + L.dtmp), ws + L.telea, (int*)(ws + L.active), sc, sel, batch, st, true);
}

// F10 on caller planes (see include/vistaf.h)
__global__ void k_ftp_residue_out(const FtpScal* sc, int* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (int)(sc[i].residues + sc[i].uf_bad);
}
extern "C" int vistaf_ftp_unwrap(VistafFtpPlan* p, const float* wrapped, const float* quality, const uint8_t* reliable, int batch,
                                 float* unwrapped, int32_t* residues, int mode, void* workspace, size_t ws_bytes, void* stream) {
  if (!p || !wrapped || !quality || !reliable || !unwrapped || !workspace || batch < 1)
    return vistaf_set_err("vistaf_ftp_unwrap: bad argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  FtpWs L;
  ftp_layout(p, batch, L);
  if (L.total > ws_bytes) return vistaf_set_err("vistaf_ftp_unwrap: workspace too small", cudaErrorInvalidValue, __FILE__, __LINE__);
  char* ws = (char*)workspace;
  cudaStream_t st = (cudaStream_t)stream;
  const int h = p->h, w = p->w, npix = h * w;
  const long long plane = npix;
  FtpScal* sc = (FtpScal*)(ws + L.scal);
  ull* uf = (ull*)(ws + L.uf);
  unsigned int* area = (unsigned int*)(ws + L.area);
  VL(), k_ftp_init<<<cdiv(batch, 128), 128, 0, st>>>(sc, batch);
  VL(), k_ftp_seed<<<dim3(64, batch), 256, 0, st>>>(quality, reliable, sc, npix);
  VL(), k_ftp_residues<<<dim3(cdiv(w, 256), h, batch), 256, 0, st>>>(wrapped, reliable, sc, h, w);
  if (uf_label(MaskFg{reliable, plane, w}, PhaseEdge{wrapped, plane}, uf, (unsigned int*)nullptr, &sc[0].uf_bad, sizeof(FtpScal), h, w,
               batch, st))
    return 1;
  VL(), k_ftp_unwrap_heap<<<batch, 32, 0, st>>>(wrapped, quality, reliable, sc, uf, area, h, w, mode == 1);
  VL(), k_ftp_seed_root<<<batch, 1, 0, st>>>(uf, sc, npix);
  VL(), k_ftp_unwrap_apply<<<dim3(cdiv(npix, 256), batch), 256, 0, st>>>(uf, wrapped, sc, unwrapped, npix);
  if (residues) VL(), k_ftp_residue_out<<<cdiv(batch, 128), 128, 0, st>>>(sc, residues, batch);
  CKL();
  return 0;
}

// force_sensor.depth_map_to_volume_cm3 on device maps: out3[b] = (volume_cm3, area_mm2, max_depth_mm)
struct VolScratch { double pos, neg, vsum; unsigned int vcnt, vmax; };
__global__ void k_vol_finish(const VolScratch* s, double* out3, int n, double mm_per_px) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double a = mm_per_px * mm_per_px;
  double V = 0, A = 0, D = 0;
  if (s[i].vcnt) {
    V = (double)__fmul_rn((float)s[i].vsum, (float)a) / 1000.0;
    A = (double)s[i].vcnt * a;
    D = (double)__uint_as_float(s[i].vmax);
  }
  out3[3 * i] = V; out3[3 * i + 1] = A; out3[3 * i + 2] = D;
}
extern "C" int vistaf_depth_to_volume(const float* height_mm, const uint8_t* roi, int batch, long long npix, double mm_per_px,
                                      double depth_eps_mm, double* out3, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  VolScratch* s = nullptr;
  CK(cudaMallocAsync(&s, sizeof(VolScratch) * batch, st));
  CK(cudaMemsetAsync(s, 0, sizeof(VolScratch) * batch, st));
  VL(), k_depth_sign<<<dim3(64, batch), 256, 0, st>>>(height_mm, roi, &s[0].pos, &s[0].neg, sizeof(VolScratch), npix);
  VL(), k_depth_reduce<<<dim3(64, batch), 256, 0, st>>>(height_mm, roi, &s[0].pos, &s[0].neg, sizeof(VolScratch), &s[0].vsum,
                                                  &s[0].vcnt, &s[0].vmax, npix, (float)depth_eps_mm);
  VL(), k_vol_finish<<<cdiv(batch, 128), 128, 0, st>>>(s, out3, batch, mm_per_px);
  CKL();
  CK(cudaFreeAsync(s, st));
  return 0;
}

// ------------------------------------------------------------------ allocation-free variant of vistaf_depth_to_volume
// Same three kernels, but the per-frame accumulators live in a file-scope device array instead of a cudaMallocAsync
// block: nothing is allocated or freed, so the call is legal under stream capture and in loops that forbid cudaMalloc.
// Every GPU that loads the library holds the array (VISTAF_VOL_STATIC_MAX_BATCH * 40 bytes = 160 KiB) until exit.  All
// calls on a device share it: it is cleared on the caller's stream before use, and two calls that overlap on the GPU
// (different streams, no ordering between them) race on it.
static __device__ __align__(256) VolScratch g_vol_static[VISTAF_VOL_STATIC_MAX_BATCH];

namespace {
constexpr int kVolMaxDev = 64;
std::mutex g_vol_static_mu;
VolScratch* g_vol_static_ptr[kVolMaxDev] = {};
// g_vol_static in HOST code is a stand-in symbol, not a device address: resolve it once per device, with that device current
int vol_static_addr(VolScratch** out) {
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kVolMaxDev)
    return vistaf_set_err("vistaf_depth_to_volume_static: device ordinal out of range", cudaErrorInvalidDevice, __FILE__, __LINE__);
  std::lock_guard<std::mutex> lk(g_vol_static_mu);
  if (!g_vol_static_ptr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_vol_static);
    if (e != cudaSuccess || !p) return vistaf_set_err("vistaf_depth_to_volume_static: cudaGetSymbolAddress", e, __FILE__, __LINE__);
    g_vol_static_ptr[dev] = (VolScratch*)p;
  }
  *out = g_vol_static_ptr[dev];
  return 0;
}
}  // namespace

extern "C" int vistaf_depth_to_volume_static_max_batch(void) { return VISTAF_VOL_STATIC_MAX_BATCH; }

extern "C" int vistaf_depth_to_volume_static(const float* height_mm, const uint8_t* roi, int batch, long long npix, double mm_per_px,
                                             double depth_eps_mm, double* out3, void* stream) {
  if (!height_mm || !out3 || batch < 1 || npix < 1)
    return vistaf_set_err("vistaf_depth_to_volume_static: bad argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  if (batch > VISTAF_VOL_STATIC_MAX_BATCH) {
    char msg[160];
    snprintf(msg, sizeof(msg), "vistaf_depth_to_volume_static: batch %d exceeds the static scratch (max %d frames per call)", batch,
             VISTAF_VOL_STATIC_MAX_BATCH);
    return vistaf_set_err(msg, cudaErrorInvalidValue, __FILE__, __LINE__);
  }
  cudaStream_t st = (cudaStream_t)stream;
  VolScratch* s = nullptr;
  if (vol_static_addr(&s)) return 1;
  CK(cudaMemsetAsync(s, 0, sizeof(VolScratch) * batch, st));          // the array keeps the last call's sums
  VL(), k_depth_sign<<<dim3(64, batch), 256, 0, st>>>(height_mm, roi, &s[0].pos, &s[0].neg, sizeof(VolScratch), npix);
  VL(), k_depth_reduce<<<dim3(64, batch), 256, 0, st>>>(height_mm, roi, &s[0].pos, &s[0].neg, sizeof(VolScratch), &s[0].vsum,
                                                  &s[0].vcnt, &s[0].vmax, npix, (float)depth_eps_mm);
  VL(), k_vol_finish<<<cdiv(batch, 128), 128, 0, st>>>(s, out3, batch, mm_per_px);
  CKL();
  return 0;
}
