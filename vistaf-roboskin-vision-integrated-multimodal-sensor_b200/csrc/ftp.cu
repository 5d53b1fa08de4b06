// Fourier-Transform-Profilometry shape path + force model on sm_100a (Code/shape_ftp.py, Code/force_sensor.py).
// Deformed frames are processed as a batch against one cached reference frame; every stage is a
// kernel over the whole batch (grid.z / grid.y = frame).  Reference file:line cited per stage.
#include "../../include/vistaf.h"
#include "common.cuh"
#include "fft.cuh"
#include "blur.cuh"
#include "select.cuh"
#include "morph.cuh"
#include "labels.cuh"
#include <string.h>
#include <new>
#include <vector>

int fft2d_dense(const void* in, float2* out, const Fft1D& pw, const Fft1D& ph, int h, int w, int batch, int sign,
                bool real_in, cudaStream_t st);

#define NPATCH 21                 // 2*PATCH_HALF_WIDTH_BINS + 1 (shape_ftp.py:25)
#define FNAN __int_as_float(0x7fc00000)

struct FtpScal {                  // per-frame scalars in the workspace
  SelectState sel[2];
  double ne[28];                  // normal equations: 21 upper-triangular + 6 rhs + count
  float coef[6];
  float sigma, med;               // IRLS scale / location
  float hi_thr, grad_thr, mu, amp_thr, contact_thr, bg_med, core_thr, core_med;
  unsigned long long seed_key;    // (quality key << 32) | ~index
  unsigned int n_rel, n_bg, n_contact, n_contact_d, n_outrel, n_bad, n_kept, residues, status;
  int bg_is_rel, flip, few;
  float gmax;                     // blob filter
  double vol_sum; unsigned int vol_cnt; float vol_max;
  double pos_sum, neg_sum;
};

struct FtpWs {
  size_t img, grad, bad, I, tmp, Iw, T, G, wrapped, amp, quality, mA, mB, reliable, lab, unw, krel, runs, nruns,
      runoff, z0, z1, den, dist, contact, bg, scal, total;
};

struct VistafFtpPlan {
  VistafFtpParams P;
  int h, w, N;                    // crop size, padded FFT size
  uint8_t *roi, *circ, *valid;    // device [h][w]
  float* apo;                     // device [h][w]
  Fft1D fN;
  GaussKernel g_illum, g_pre, g_qual, g_smooth, g_unrel;
  MorphSE se_bad, se_close, se_dil, se_edge;
  // reference-frame products (shape_ftp.py:1632-1639)
  bool have_ref;
  float2* cref;                   // device [h][w]
  float* amp_ref;
  float peak_xf, peak_yf;
  int pxi, pyi;
  double kx, ky;
  float2 *rampx, *rampy;          // fractional demodulation ramps, device [N]
  float* hann;                    // device [NPATCH]
  int roi_count;
};

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

extern "C" int vistaf_crop_gray(const uint8_t* frames, int batch, int H, int W, int x1, int y1, int h, int w,
                                uint8_t* gray, void* stream) {
  if (x1 < 0 || y1 < 0 || x1 + w > W || y1 + h > H) return vistaf_set_err("crop outside frame", cudaErrorInvalidValue, __FILE__, __LINE__);
  k_crop_gray<<<dim3(cdiv(w, 256), h, batch), 256, 0, (cudaStream_t)stream>>>(frames, gray, H, W, x1, y1, h, w);
  CKL();
  return 0;
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

struct MaskedVal {              // value stream: plane[b][i] where mask[i] (mask shared across frames)
  const float* p; const uint8_t* m; long long plane;
  __device__ bool operator()(int b, long long i, float& v) const {
    if (!m[i]) return false;
    v = p[(long long)b * plane + i];
    return v == v;
  }
};
struct MaskedValB {             // per-frame mask
  const float* p; const uint8_t* m; long long plane;
  __device__ bool operator()(int b, long long i, float& v) const {
    if (!m[(long long)b * plane + i]) return false;
    v = p[(long long)b * plane + i];
    return v == v;
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

__global__ void k_ftp_grab_thr(FtpScal* sc, int which) {     // copy a finished quantile into its named slot
  FtpScal& s = sc[blockIdx.x];
  float r = s.sel[0].result;
  switch (which) {
    case 0: s.hi_thr = r; break;
    case 1: s.grad_thr = r; break;
    case 2: s.mu = r; break;
    case 3: s.amp_thr = r; break;
    case 4: s.contact_thr = r; break;
    case 5: s.bg_med = r; break;
    case 6: s.core_thr = r; break;
    case 7: s.core_med = r; s.flip = (s.sel[0].count > 0 && r > 0.f) ? 1 : 0; break;
    case 8: s.med = r; break;
    case 9: s.sigma = 1.4826f * (r + 1e-6f); break;
  }
}
