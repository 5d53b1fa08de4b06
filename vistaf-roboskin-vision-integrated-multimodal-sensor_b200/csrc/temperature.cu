// Temperature path on sm_100a: BGR frame -> stripe segmentation (FFT sideband) -> blurred CIELAB
// features -> two polynomial regressors -> fusion -> oriented smoothing -> statistics.
// Mirrors Code/temperature_sensor.py (reference file:line cited per kernel); batched over frames.
#include "../../include/vistaf.h"
#include "common.cuh"
#include "fft.cuh"
#include "blur.cuh"
#include "select.cuh"
#include "morph.cuh"
#include "labels.cuh"
#include "telea.cuh"
#include <string.h>
#include <new>
#include <vector>
#include <algorithm>

#define TOPK 16
// row transforms: two buffers of W float2, the second with W/16 entries of padding for fft_smem_fused
#define ROWS_SMEM(W) (((size_t)(W) + (size_t)fft_fused_ldb(W)) * sizeof(float2))
// column transform: COLS_PER_BLOCK sequences of H + 1 float2 and the scratch at the padded stride
#define COLS_SMEM(H) ((size_t)COLS_PER_BLOCK * ((size_t)(H) + 1 + (size_t)fft_fused_ldb(H)) * sizeof(float2))
#define COLS_PER_BLOCK 2      // 2 columns x 2 buffers x 2161 float2 = 69 KB: three blocks per SM overlap each other's barriers
#define COLS_THREADS 256
#define COLS_PT 18             // bins per thread held in registers for the top-16 search: COLS_PER_BLOCK * H <= 18 * 256
#define MAXBAND 45            // 2*22+1
#define TEMP_SEL_CAP (1u << 20) // key-list capacity per frame for the median of the final map: temperatures of a frame
                                // crowd into one or two top-11-bit float bins (values in [20, 75])

struct Cand { float mag; int ky, kx; int pad; };

struct TempWs {               // per-chunk workspace layout (byte offsets), planes are [chunk][...]
  size_t gray8, satraw, sat, norm, tmp, A, G, maskA, m1, m2, cand, scal, sel, total;
  size_t t10_frame, t10_stat, t10_wide, t10_color, t10_telea;      // T10 slots (saturated frames only)
  size_t t12_frame;                                                // T12 slots (frames with a tilted carrier)
  int t10_slots;
};

struct TempScal {             // per-frame scalars living in the workspace
  unsigned int hist[256];
  unsigned int sat_raw_count, sat_count, n_eff, status;
  float median;
  float mu_f;
  double norm_sum;
  double csum[2];
  double sumA, sumB;
  unsigned int cntA, cntB;
  int peak_x, peak_y;
  int a_is_dark;
  unsigned int dark_count, color_ok_count;
  double t_sum, t_sumsq;
  float t_min, t_max;
  unsigned int t_count, pad;
  double rot_c, rot_s;          // cos / sin of phi0 (temperature_sensor.py:470-474), set by k_temp_phi0
  float blur_rc, blur_cc;       // sigma-20 blur of the constant background: row pass R(median), column pass C(R(median))
  int t10_slot;                 // T10 slot of a saturated frame (-1: none free, the fill is skipped and flagged)
  unsigned int wmin_key, wmax_key, pad2;      // ordered keys of min / max of the raw wide-model map over roi_eff
  double aff_f[6], aff_b[6];    // T12 on a tilted carrier: dst -> src maps of the forward / backward warpAffine
  unsigned int peak_floor, pad3;  // bits of a lower bound of the frame's 16th largest |F| (raised by the column blocks as they finish)
};

struct VistafTempPlan {
  VistafTempParams P;
  int H, W, Wh;
  uint8_t* roi;               // device [H][W]
  int y0, y1, x0, x1;         // ROI bounding box (inclusive-exclusive)
  Fft1D fw, fh;
  GaussKernel g_illum, g_across, g_along;
  MorphSE se_sat;
  VistafPolyModel* models;    // device [2]: colour, wide
  int* lab_gamma;             // device
  int* lab_cbrt;
  int nblk_cols;
};

// ------------------------------------------------------------------ ROI (temperature_sensor.py:180-184)
__global__ void k_roi_mask(uint8_t* roi, int h, int w, double cx, double cy, double r2) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  double dx = __dsub_rn((double)x, cx), dy = __dsub_rn((double)y, cy);
  double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
  roi[(size_t)y * w + x] = d2 <= r2 ? 1 : 0;
}

// ------------------------------------------------------------------ T2/T3 front: gray, saturation
// cv2.cvtColor(BGR2GRAY) 8-bit: (9798 R + 19235 G + 3735 B + 2^14) >> 15   (temperature_sensor.py:379,439)
__device__ __forceinline__ int gray8_of(int b, int g, int r) { return (9798 * r + 19235 * g + 3735 * b + 16384) >> 15; }

__global__ void k_temp_gray(const uint8_t* __restrict__ frames, const uint8_t* __restrict__ roi,
                            uint8_t* __restrict__ gray8, uint8_t* __restrict__ satraw, TempScal* sc, int npix,
                            int sat_thresh) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;      // group of 4 pixels
  int nsat = 0;
  if (q * 4 < npix) {
    const uint32_t* src = (const uint32_t*)(frames + (size_t)b * npix * 3) + (size_t)q * 3;
    uint32_t w0 = __ldg(src), w1 = __ldg(src + 1), w2 = __ldg(src + 2);
    const uint32_t wv[3] = {w0, w1, w2};
    int px[12];
#pragma unroll
    for (int i = 0; i < 12; i++) px[i] = (int)((wv[i >> 2] >> (8 * (i & 3))) & 255u);
    uchar4 rm = *(const uchar4*)(roi + (size_t)q * 4);
    uint8_t r4[4] = {rm.x, rm.y, rm.z, rm.w};
    uint8_t g4[4], s4[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      int g = gray8_of(px[3 * i], px[3 * i + 1], px[3 * i + 2]);
      g4[i] = (uint8_t)g;
      s4[i] = (g >= sat_thresh && r4[i]) ? 1 : 0;
      nsat += s4[i];
    }
    *(uchar4*)(gray8 + (size_t)b * npix + (size_t)q * 4) = make_uchar4(g4[0], g4[1], g4[2], g4[3]);
    *(uchar4*)(satraw + (size_t)b * npix + (size_t)q * 4) = make_uchar4(s4[0], s4[1], s4[2], s4[3]);
  }
  nsat = warp_sum(nsat);
  if ((threadIdx.x & 31) == 0 && nsat) atomicAdd(&sc[b].sat_raw_count, (unsigned)nsat);
}

// dilate(sat, ellipse k) & roi — only when the frame has saturated pixels (temperature_sensor.py:381-386)
#define SATD_ROWS 8
__global__ void k_temp_sat_dilate(const uint8_t* __restrict__ satraw, const uint8_t* __restrict__ roi,
                                  uint8_t* __restrict__ sat, TempScal* sc, int h, int w, MorphSE se, int y0, int y1) {
  const int b = blockIdx.z;
  if (sc[b].sat_raw_count == 0) return;             // the plane is never read for such a frame (sat_at)
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const bool inb = x < w;
  const uint8_t* s = satraw + (size_t)b * h * w;
  int c = 0;
  // rows [y0, y1) = the ROI bounding box: the mask is `& roi`, and its readers test the ROI first
  for (int y = y0 + (int)blockIdx.y * SATD_ROWS; y < min(y1, y0 + ((int)blockIdx.y + 1) * SATD_ROWS); y++) {
    int v = 0;
    if (inb && roi[(size_t)y * w + x]) {
      for (int i = 0; i < se.rows && !v; i++) {
        int yy = y + i - se.rows / 2;
        if (yy < 0 || yy >= h || se.half[i] < 0) continue;
        for (int xx = max(0, x - se.half[i]); xx <= min(w - 1, x + se.half[i]); xx++)
          if (s[(size_t)yy * w + xx]) { v = 1; break; }
      }
    }
    if (inb) sat[(size_t)b * h * w + (size_t)y * w + x] = (uint8_t)v;
    c += v;
  }
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&sc[b].sat_count, (unsigned)c);
}

__device__ __forceinline__ bool sat_at(const TempScal& s, const uint8_t* sat, size_t i) {
  return s.sat_raw_count != 0 && sat[i] != 0;
}

// histogram of gray over roi_eff -> exact median of integer-valued float32 (temperature_sensor.py:446-447)
__global__ void k_temp_hist(const uint8_t* __restrict__ gray8, const uint8_t* __restrict__ roi,
                            const uint8_t* __restrict__ sat, TempScal* sc, int npix, int i0, int i1) {   // pixels [i0, i1): the ROI rows
  __shared__ unsigned int h[256];
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < 256; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const uint8_t* g = gray8 + (size_t)b * npix;
  const uint8_t* s = sat + (size_t)b * npix;
  const bool has_sat = sc[b].sat_raw_count != 0;
  for (int i = i0 + blockIdx.x * blockDim.x + threadIdx.x; i < i1; i += gridDim.x * blockDim.x)
    if (roi[i] && !(has_sat && s[i])) atomicAdd(&h[g[i]], 1u);
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += blockDim.x)
    if (h[i]) atomicAdd(&sc[b].hist[i], h[i]);
}

__global__ void k_temp_median(TempScal* sc) {
  TempScal& s = sc[blockIdx.x];
  if (threadIdx.x != 0) return;
  unsigned int n = 0;
  for (int i = 0; i < 256; i++) n += s.hist[i];
  s.n_eff = n;
  if (n == 0) { s.status |= VISTAF_ST_EMPTY_ROI; s.median = 0.f; return; }
  unsigned int k0 = (n - 1) / 2, k1 = n / 2, acc = 0;
  int v0 = -1, v1 = -1;
  for (int i = 0; i < 256; i++) {
    acc += s.hist[i];
    if (v0 < 0 && acc > k0) v0 = i;
    if (v1 < 0 && acc > k1) { v1 = i; break; }
  }
  s.median = ((float)v0 + (float)v1) / 2.0f;
}

// ------------------------------------------------------------------ T3: illumination normalisation
// g[~roi] = median; norm = g / GaussianBlur(g, sigma 20) (blur < 1e-6 -> 1); mu = mean(norm[roi_eff])
// (temperature_sensor.py:446-450, 363-375).  Load / Store functors for gauss_blur().
struct IllumLoad {
  const uint8_t* gray8; const uint8_t* roi; const TempScal* sc; long long plane; int w;
  __device__ float operator()(int b, int y, int x) const {
    long long i = (long long)y * w + x;
    return roi[i] ? (float)gray8[(long long)b * plane + i] : sc[b].median;
  }
};
struct IllumStore {
  const uint8_t* gray8; const uint8_t* roi; const uint8_t* sat; TempScal* sc; float* norm; long long plane; int w;
  double acc;
  __device__ void operator()(int b, int y, int x, float blur) {
    long long i = (long long)y * w + x;
    const bool in = roi[i] != 0;
    float g = in ? (float)gray8[(long long)b * plane + i] : sc[b].median;
    if (blur < 1e-6f) blur = 1.0f;
    float n = __fdiv_rn(g, blur);
    norm[(long long)b * plane + i] = n;
    if (in && !sat_at(sc[b], sat + (long long)b * plane, i)) acc += (double)n;
  }
  __device__ void outside(int b, int y, int x, float blur) {     // window blocks lie outside the ROI: g = median
    if (blur < 1e-6f) blur = 1.0f;
    norm[(long long)b * plane + (long long)y * w + x] = __fdiv_rn(sc[b].median, blur);
  }
  __device__ void finish(int b) {
    __shared__ double sh[32];
    double v[1] = {acc};
    block_sum<1>(v, sh);
    if (threadIdx.x == 0 && threadIdx.y == 0 && v[0] != 0.0) atomicAdd(&sc[b].norm_sum, v[0]);
  }
};

// blur of the median-filled background far from the ROI: the exact value both passes of gauss_blur produce there
__global__ void k_temp_blur_consts(TempScal* sc, GaussKernel g) {
  TempScal& s = sc[blockIdx.x];
  s.blur_rc = gauss_constant(g.w, g.kpad, s.median);
  s.blur_cc = gauss_constant(g.w, g.kpad, s.blur_rc);
}

__global__ void k_temp_mu(TempScal* sc) {
  TempScal& s = sc[blockIdx.x];
  double mu = s.n_eff ? s.norm_sum / (double)s.n_eff : 1.0;
  if (fabs(mu) <= 1e-9) mu = 1.0;
  s.mu_f = (float)mu;
}

// ------------------------------------------------------------------ T4: forward FFT
// Row pass, two real rows per complex transform (z = row0 + i*row1), unpacked by Hermitian symmetry
// into the half spectrum A[y][kx], kx = 0..W/2.  I_norm = norm / mu is applied on load.
__global__ void __launch_bounds__(256) k_temp_fft_rows(const float* __restrict__ norm, float2* __restrict__ A,
                                                        const TempScal* sc, Fft1D p, int h, int wh) {
  extern __shared__ float2 sm[];
  const int n = p.n, b = blockIdx.y, r = blockIdx.x * 2;
  float2* a = sm;
  float2* bb = sm + n;
  const float mu = sc[b].mu_f;
  const float* r0 = norm + ((size_t)b * h + r) * n;
  const float* r1 = r0 + n;
  const bool has1 = r + 1 < h;
  for (int x0 = threadIdx.x; x0 < n; x0 += blockDim.x * 4) {       // all loads of four trips issued before use
    float u0[4], u1[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int x = x0 + j * blockDim.x;
      u0[j] = x < n ? r0[x] : 0.f;
      u1[j] = (x < n && has1) ? r1[x] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int x = x0 + j * blockDim.x;
      if (x < n) a[x] = make_float2(__fdiv_rn(u0[j], mu), has1 ? __fdiv_rn(u1[j], mu) : 0.f);
    }
  }
  int ldz;
  float2* z = fft_smem_fused<-1, false>(a, bb, p, 1, n, fft_fused_ldb(n), ldz);
  float2* o0 = A + ((size_t)b * h + r) * wh;
  float2* o1 = o0 + wh;
  for (int k = threadIdx.x; k < wh; k += blockDim.x) {
    float2 zk = z[k], zn = z[(n - k) % n];
    o0[k] = make_float2(0.5f * (zk.x + zn.x), 0.5f * (zk.y - zn.y));
    if (has1) o1[k] = make_float2(0.5f * (zk.y + zn.y), 0.5f * (zn.x - zk.x));
  }
}

// Is the fftshift-ed bin (x, y) inside the DC exclusion box? (temperature_sensor.py:323-327)
__device__ __forceinline__ bool dc_excluded(int x, int y, int h, int w, int dc) {
  int cy = h / 2, cx = w / 2;
  return y >= max(0, cy - dc) && y < min(h, cy + dc) && x >= max(0, cx - dc) && x < min(w, cx + dc);
}

// Column pass over the half spectrum + fused |F| and per-block top-16 (the dense spectrum is never
// written: only the band around the carrier is needed later and is recomputed from A).
__global__ void __launch_bounds__(COLS_THREADS, 2) k_temp_fft_cols_peaks(const float2* __restrict__ A, Cand* __restrict__ cand,
                                                              Fft1D p, int w, int wh, int dc, TempScal* sc) {
  extern __shared__ float2 sm[];
  const int n = p.n, lds = n + 1, b = blockIdx.y, c0 = blockIdx.x * COLS_PER_BLOCK;
  const int nc = min(COLS_PER_BLOCK, wh - c0);
  float2* a = sm;
  float2* bb = sm + (size_t)COLS_PER_BLOCK * lds;
  const float2* src = A + (size_t)b * n * wh;
  // eight independent 8-byte loads per thread in flight (the staging loop is latency bound: one block per SM)
  for (int t0 = threadIdx.x; t0 < nc * n; t0 += blockDim.x * 8) {
    float2 v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int t = t0 + j * blockDim.x;
      if (t < nc * n) { const int y = t / nc, c = t - y * nc; v[j] = src[(size_t)y * wh + c0 + c]; }
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int t = t0 + j * blockDim.x;
      if (t < nc * n) { const int y = t / nc, c = t - y * nc; a[c * lds + y] = v[j]; }
    }
  }
  int ldf;
  float2* F = fft_smem_fused<-1, true>(a, bb, p, nc, lds, fft_fused_ldb(n), ldf);
  float* mag = (float*)(F == a ? bb : a);          // scratch buffer is free now
  __syncthreads();
  // the DC exclusion box only reaches columns |kx| <= dc: every other block skips the (modulo-heavy) box test
  const bool near_dc = c0 <= dc + 1 || c0 + COLS_PER_BLOCK >= w / 2;
  for (int t = threadIdx.x; t < nc * n; t += blockDim.x) {
    int c = 0, ky = t;
#pragma unroll
    for (int q = 1; q < COLS_PER_BLOCK; q++) if (ky >= n) { ky -= n; c++; }
    const int kx = c0 + c;
    float2 v = F[c * ldf + ky];
    float m = sqrtf(v.x * v.x + v.y * v.y);
    if (near_dc) {
      // entry 1: (kx, ky); entry 2 (twin, same magnitude): (-kx, -ky).  Drop the bin if both are excluded.
      int x1 = (kx + w / 2) % w, y1 = (ky + n / 2) % n;
      bool ex1 = dc_excluded(x1, y1, n, w, dc);
      bool ex2 = true;
      if (kx != 0 && kx != w / 2) {
        int x2 = (w - kx + w / 2) % w, y2 = ((n - ky) % n + n / 2) % n;
        ex2 = dc_excluded(x2, y2, n, w, dc);
      }
      if (ex1 && ex2) m = -1.f;
    }
    mag[t] = m;
  }
  __syncthreads();
  // Top-16 of the block's bins without block-wide rounds: every lane keeps its bins in registers, each warp extracts
  // its own 16 largest with shuffles only (ties -> smaller index, as a full-block argmax would), and warp 0 merges
  // the 8 x 16 survivors.
  __shared__ float wm[COLS_THREADS / 32][TOPK];
  __shared__ int wi[COLS_THREADS / 32][TOPK];
  Cand* out = cand + ((size_t)b * gridDim.x + blockIdx.x) * TOPK;
  const int tot = nc * n, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float mv[COLS_PT];
#pragma unroll
  for (int j = 0; j < COLS_PT; j++) { const int t = threadIdx.x + j * COLS_THREADS; mv[j] = t < tot ? mag[t] : -3.f; }
  // A block that has found 16 bins raises the frame's floor to its 16th largest magnitude: the frame's final 16th
  // largest is at least that.  A block none of whose bins reaches the floor cannot contribute to the top 16 and skips
  // the extraction (the low-|kx| blocks, scheduled first, hold the large magnitudes: almost every later block skips).
  // The RESULT does not depend on the order the blocks run in, only the amount of work does.
  {
    const float floor_m = __uint_as_float(*(volatile unsigned int*)&sc[b].peak_floor);
    int any = 0;
#pragma unroll
    for (int j = 0; j < COLS_PT; j++) any |= mv[j] >= floor_m;
    if (!__syncthreads_or(any)) {
      if (threadIdx.x < TOPK) { Cand cd; cd.mag = -1.f; cd.kx = -1; cd.ky = -1; cd.pad = 0; out[threadIdx.x] = cd; }
      return;
    }
  }
  for (int round = 0; round < TOPK; round++) {
    float best = -4.f;
    int bt = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < COLS_PT; j++)
      if (mv[j] > best) { best = mv[j]; bt = threadIdx.x + j * COLS_THREADS; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int ot = __shfl_xor_sync(0xffffffffu, bt, o);
      if (ob > best || (ob == best && ot < bt)) { best = ob; bt = ot; }
    }
#pragma unroll
    for (int j = 0; j < COLS_PT; j++)
      if (threadIdx.x + j * COLS_THREADS == bt) mv[j] = -4.f;       // the winner leaves its lane's set
    if (lane == 0) { wm[wid][round] = best; wi[wid][round] = bt; }
  }
  __syncthreads();
  if (wid == 0) {
    constexpr int NE = (COLS_THREADS / 32) * TOPK / 32;             // survivors per lane
    float em[NE];
    int ei[NE];
#pragma unroll
    for (int q = 0; q < NE; q++) {
      const int e = lane + 32 * q;
      em[q] = wm[e / TOPK][e % TOPK]; ei[q] = wi[e / TOPK][e % TOPK];
    }
    for (int round = 0; round < TOPK; round++) {
      float best = -4.f;
      int bt = 0x7fffffff;
#pragma unroll
      for (int q = 0; q < NE; q++)
        if (em[q] > best || (em[q] == best && ei[q] < bt)) { best = em[q]; bt = ei[q]; }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int ot = __shfl_xor_sync(0xffffffffu, bt, o);
        if (ob > best || (ob == best && ot < bt)) { best = ob; bt = ot; }
      }
#pragma unroll
      for (int q = 0; q < NE; q++)
        if (ei[q] == bt && em[q] == best) em[q] = -4.f;
      if (lane == 0) {
        Cand cd;
        cd.mag = best; cd.pad = 0;
        if (best >= 0.f) { const int c = bt / n; cd.kx = c0 + c; cd.ky = bt - c * n; }
        else { cd.kx = -1; cd.ky = -1; cd.mag = -1.f; }
        out[round] = cd;
        if (round == TOPK - 1 && best > 0.f) atomicMax(&sc[b].peak_floor, __float_as_uint(best));
      }
    }
  }
}

// ------------------------------------------------------------------ T4: carrier peak
// Merge the per-block candidates of one frame: expand each half-plane bin into its (up to) two
// full-plane entries, keep the 16 largest outside the DC box, then apply the selection rule of
// temperature_sensor.py:341-360 (right half plane, |dy| <= 0.14 h, largest magnitude).
__global__ void __launch_bounds__(256) k_temp_peak_select(const Cand* __restrict__ cand, TempScal* sc, int ncand,
                                                           int h, int w, int dc, float max_dy_frac) {
  extern __shared__ float emag[];                  // 2*ncand entry magnitudes
  __shared__ float smax[8];
  __shared__ int sidx[8];
  __shared__ int top_x[TOPK], top_y[TOPK];
  __shared__ float top_m[TOPK];
  const int b = blockIdx.x;
  const Cand* c = cand + (size_t)b * ncand;
  for (int e = threadIdx.x; e < 2 * ncand; e += blockDim.x) {
    Cand cd = c[e >> 1];
    float m = -1.f;
    if (cd.kx >= 0) {
      int x, y;
      bool ok = true;
      if ((e & 1) == 0) { x = (cd.kx + w / 2) % w; y = (cd.ky + h / 2) % h; }
      else {
        ok = cd.kx != 0 && cd.kx != w / 2;
        x = (w - cd.kx + w / 2) % w; y = ((h - cd.ky) % h + h / 2) % h;
      }
      if (ok && !dc_excluded(x, y, h, w, dc)) m = cd.mag;
    }
    emag[e] = m;
  }
  __syncthreads();
  int ntop = 0;
  for (int round = 0; round < TOPK; round++) {
    float best = -2.f;
    int bi = -1;
    for (int e = threadIdx.x; e < 2 * ncand; e += blockDim.x)
      if (emag[e] > best) { best = emag[e]; bi = e; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ob = __shfl_xor_sync(0xffffffffu, best, o);
      int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ob > best || (ob == best && oi >= 0 && (bi < 0 || oi < bi))) { best = ob; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { smax[threadIdx.x >> 5] = best; sidx[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int i = 1; i < 8; i++)
        if (smax[i] > best || (smax[i] == best && sidx[i] >= 0 && (bi < 0 || sidx[i] < bi))) { best = smax[i]; bi = sidx[i]; }
      if (bi >= 0 && best >= 0.f) {
        Cand cd = c[bi >> 1];
        int x, y;
        if ((bi & 1) == 0) { x = (cd.kx + w / 2) % w; y = (cd.ky + h / 2) % h; }
        else { x = (w - cd.kx + w / 2) % w; y = ((h - cd.ky) % h + h / 2) % h; }
        top_x[round] = x; top_y[round] = y; top_m[round] = best;
        emag[bi] = -1.f;
      } else top_m[round] = -1.f;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const int cy = h / 2, cx = w / 2, max_dy = (int)(max_dy_frac * (float)h);
    for (int i = 0; i < TOPK; i++) if (top_m[i] >= 0.f) ntop++;
    if (ntop == 0) { sc[b].status |= VISTAF_ST_NO_PEAK; sc[b].peak_x = cx; sc[b].peak_y = cy; return; }
    bool any_right = false, any_near = false;
    for (int i = 0; i < TOPK; i++) if (top_m[i] >= 0.f && top_x[i] > cx) any_right = true;
    for (int i = 0; i < TOPK; i++)
      if (top_m[i] >= 0.f && (!any_right || top_x[i] > cx) && abs(top_y[i] - cy) <= max_dy) any_near = true;
    int sel = -1;
    for (int i = 0; i < TOPK; i++) {
      if (top_m[i] < 0.f) continue;
      if (any_right && !(top_x[i] > cx)) continue;
      if (any_near && !(abs(top_y[i] - cy) <= max_dy)) continue;
      if (sel < 0 || top_m[i] > top_m[sel]) sel = i;
    }
    sc[b].peak_x = top_x[sel];
    sc[b].peak_y = top_y[sel];
    if (top_y[sel] != cy) sc[b].status |= VISTAF_ST_ROTATED_CARRIER;
  }
}

// ------------------------------------------------------------------ T5: band-pass + inverse, column part
// Block (j, b): column j of the (2R+1)^2 box around the peak.  F[ky,kx] on the disc is recomputed by
// a direct DFT of column |kx| of A (Hermitian twin for kx < 0), then G[y][j] = sum_ky F e^{+2 pi i ky y/H}/(HW).
__global__ void __launch_bounds__(256) k_temp_band(const float2* __restrict__ A, float2* __restrict__ G,
                                                    const TempScal* sc, Fft1D ph, int w, int wh, int R) {
  extern __shared__ float2 col[];                  // H entries
  __shared__ float2 Fd[MAXBAND];
  const int h = ph.n, b = blockIdx.y, j = blockIdx.x, nb = 2 * R + 1;
  const int px = sc[b].peak_x, py = sc[b].peak_y;
  const int xs = px + j - R;                        // fftshift-ed column of this block
  const bool col_ok = xs >= 0 && xs < w;
  int kx = xs - w / 2;                              // signed frequency
  const bool twin = kx < 0;
  const int ka = twin ? -kx : kx;                   // column of A (0..W/2)
  const float2* src = A + (size_t)b * h * wh;
  for (int y = threadIdx.x; y < h; y += blockDim.x) col[y] = col_ok ? src[(size_t)y * wh + ka] : make_float2(0.f, 0.f);
  __syncthreads();
  // step 1: the nb bins of this column, one warp per bin (strided), lanes split the y sum
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int i = wid; i < nb; i += nw) {
    const int ys = py + i - R;
    const int dx = j - R, dy = i - R;
    bool in = col_ok && ys >= 0 && ys < h && (dx * dx + dy * dy) <= R * R;
    float2 acc = make_float2(0.f, 0.f);
    if (in) {
      int ky = ((ys - h / 2) % h + h) % h;
      int kq = twin ? (h - ky) % h : ky;            // F[ky,kx] = conj(F[-ky,-kx])
      int idx = (int)(((long long)kq * lane) % h);  // (kq y) mod h, advanced by (32 kq) mod h per trip: the same entries
      const int step = (int)(((long long)kq * 32) % h);
      for (int y = lane; y < h; y += 32) {
        float2 t = __ldg(ph.tw + idx);
        acc = cadd(acc, cmul(col[y], t));
        idx += step;
        if (idx >= h) idx -= h;
      }
    }
    acc.x = warp_sum(acc.x); acc.y = warp_sum(acc.y);
    if (lane == 0) Fd[i] = twin ? make_float2(acc.x, -acc.y) : acc;
  }
  __syncthreads();
  // step 2: inverse along y for this column
  const float scale = 1.0f / ((float)h * (float)w);
  for (int y = threadIdx.x; y < h; y += blockDim.x) {
    float2 acc = make_float2(0.f, 0.f);
    // ky of consecutive rows of the box differ by one, so (ky y) mod h advances by y mod h per row
    const int ky0 = ((py - R - h / 2) % h + h) % h;
    int idx = (int)(((long long)ky0 * y) % h);
    for (int i = 0; i < nb; i++) {
      int ys = py + i - R;
      if (ys >= 0 && ys < h) {
        float2 t = __ldg(ph.tw + idx);
        t.y = -t.y;                                 // e^{+i...}
        acc = cadd(acc, cmul(Fd[i], t));
      }
      idx += y;                                     // y < h
      if (idx >= h) idx -= h;
    }
    G[((size_t)b * h + y) * MAXBAND + j] = make_float2(acc.x * scale, acc.y * scale);
  }
}

// ------------------------------------------------------------------ T5/T6: inverse rows
// z[y][:] = IFFT_W(sparse spectrum of <= 45 bins).  PASS 1 accumulates c = sum_roi_eff z * (I_norm - 1)
// (temperature_sensor.py:469-470); PASS 2 rotates by phi0 = angle(c), thresholds s = Re(z e^{-i phi0}) >= 0
// and accumulates the mean gray of both classes (:471-480).  Only ROI rows are launched.
template <int PASS>
__global__ void __launch_bounds__(256) k_temp_ifft_rows(const float2* __restrict__ G, const float* __restrict__ norm,
                                                         const uint8_t* __restrict__ gray8,
                                                         const uint8_t* __restrict__ roi, const uint8_t* __restrict__ sat,
                                                         uint8_t* __restrict__ maskA, TempScal* sc, Fft1D p, int h,
                                                         int ybeg, int R, float2* __restrict__ Z, int zx0, int zw, int zrows) {
  extern __shared__ float2 sm[];
  __shared__ double red[4 * 32];
  const int n = p.n, b = blockIdx.y, y = ybeg + blockIdx.x, nb = 2 * R + 1;
  float2* a = sm;
  float2* bb = sm + n;
  TempScal& S = sc[b];
  for (int x = threadIdx.x; x < n; x += blockDim.x) a[x] = make_float2(0.f, 0.f);
  __syncthreads();
  if ((int)threadIdx.x < nb) {
    int xs = S.peak_x + (int)threadIdx.x - R;
    if (xs >= 0 && xs < n) a[((xs - n / 2) % n + n) % n] = G[((size_t)b * h + y) * MAXBAND + threadIdx.x];
  }
  int ldz;
  float2* z = fft_smem_fused<1, false>(a, bb, p, 1, n, fft_fused_ldb(n), ldz);
  const size_t row = (size_t)y * n, fr = (size_t)b * h * n;
  const bool has_sat = S.sat_raw_count != 0;
  if (PASS == 1) {
    const float mu = S.mu_f;
    double v[2] = {0.0, 0.0};
    for (int x = threadIdx.x; x < n; x += blockDim.x) {
      if (roi[row + x] && !(has_sat && sat[fr + row + x])) {
        float m = __fsub_rn(__fdiv_rn(norm[fr + row + x], mu), 1.0f);
        v[0] += (double)(z[x].x * m);
        v[1] += (double)(z[x].y * m);
      }
    }
    block_sum<2>(v, red);
    if (threadIdx.x == 0) { atomicAdd(&S.csum[0], v[0]); atomicAdd(&S.csum[1], v[1]); }
    // keep the analytic signal over the ROI bounding box: the sign split (k_temp_split) re-reads it instead of
    // transforming every row a second time
    if (Z) {
      float2* zr = Z + ((size_t)b * zrows + blockIdx.x) * zw;
      for (int x = threadIdx.x; x < zw; x += blockDim.x) zr[x] = z[zx0 + x];
    }
  } else {
    const double phi0 = atan2(S.csum[1], S.csum[0]);
    const double cs = cos(phi0), sn = sin(phi0);
    double v[4] = {0.0, 0.0, 0.0, 0.0};
    for (int x = threadIdx.x; x < n; x += blockDim.x) {
      int mA = 0;
      if (roi[row + x] && !(has_sat && sat[fr + row + x])) {
        float s = (float)((double)z[x].x * cs + (double)z[x].y * sn);
        float g = (float)gray8[fr + row + x];
        if (s >= 0.f) { mA = 1; v[0] += g; v[1] += 1.0; } else { v[2] += g; v[3] += 1.0; }
      }
      maskA[fr + row + x] = (uint8_t)mA;
    }
    block_sum<4>(v, red);
    if (threadIdx.x == 0) {
      atomicAdd(&S.sumA, v[0]); atomicAdd(&S.cntA, (unsigned)v[1]);
      atomicAdd(&S.sumB, v[2]); atomicAdd(&S.cntB, (unsigned)v[3]);
    }
  }
}

// A / B split by the sign of s = Re(z e^{-i phi0}) over roi_eff, and the mean gray of both classes
// (temperature_sensor.py:475-489), from the analytic signal stored by k_temp_ifft_rows<1>.  maskA is written over
// the ROI bounding box only (its readers test the ROI first).
#define SPLIT_ROWS 8            // rows per block: one float64 block reduction per 8 x 256 pixels instead of per 256
__global__ void __launch_bounds__(256, 8) k_temp_split(const float2* __restrict__ Z, const uint8_t* __restrict__ gray8,
                                                     const uint8_t* __restrict__ roi, const uint8_t* __restrict__ sat,
                                                     uint8_t* __restrict__ maskA, TempScal* sc, int h, int w, int ybeg, int zx0,
                                                     int zw, int zrows) {
  __shared__ double red[4 * 32];
  const int b = blockIdx.z, xx = blockIdx.x * 256 + threadIdx.x;
  TempScal& S = sc[b];
  const bool has_sat = S.sat_raw_count != 0;
  const double cs = S.rot_c, sn = S.rot_s;
  double v[4] = {0.0, 0.0, 0.0, 0.0};
  if (xx < zw) {
    const int x = zx0 + xx;
    const size_t fr = (size_t)b * h * w;
    const int r1 = min(zrows, ((int)blockIdx.y + 1) * SPLIT_ROWS);
    for (int ry = blockIdx.y * SPLIT_ROWS; ry < r1; ry++) {
      const size_t row = (size_t)(ybeg + ry) * w;
      int mA = 0;
      if (roi[row + x] && !(has_sat && sat[fr + row + x])) {
        const float2 z = Z[((size_t)b * zrows + ry) * zw + xx];
        const float s = (float)((double)z.x * cs + (double)z.y * sn);
        const float g = (float)gray8[fr + row + x];
        if (s >= 0.f) { mA = 1; v[0] += g; v[1] += 1.0; } else { v[2] += g; v[3] += 1.0; }
      }
      maskA[fr + row + x] = (uint8_t)mA;
    }
  }
  block_sum<4>(v, red);
  if (threadIdx.x == 0 && (v[1] != 0.0 || v[3] != 0.0)) {
    atomicAdd(&S.sumA, v[0]); atomicAdd(&S.cntA, (unsigned)v[1]);
    atomicAdd(&S.sumB, v[2]); atomicAdd(&S.cntB, (unsigned)v[3]);
  }
}

__global__ void k_temp_phi0(TempScal* sc) {       // one float64 atan2 / sincos per frame
  TempScal& s = sc[blockIdx.x];
  const double phi0 = atan2(s.csum[1], s.csum[0]);
  s.rot_c = cos(phi0); s.rot_s = sin(phi0);
}
__global__ void k_temp_choose_dark(TempScal* sc) {
  TempScal& s = sc[blockIdx.x];
  double ma = s.cntA ? s.sumA / s.cntA : 1e9, mb = s.cntB ? s.sumB / s.cntB : 1e9;
  s.a_is_dark = ma <= mb ? 1 : 0;
}

// dark candidate before morphology: A or (roi_eff & ~A)  (temperature_sensor.py:482-489)
struct DarkLoad {
  const uint8_t* maskA; const uint8_t* roi; const uint8_t* sat; const TempScal* sc; long long plane; int w; int ybeg, yend;
  __device__ int operator()(int b, int y, int x) const {
    if (y < ybeg || y >= yend) return 0;             // maskA is only written for ROI rows
    long long i = (long long)y * w + x;
    if (!roi[i] || sat_at(sc[b], sat + (long long)b * plane, i)) return 0;
    int a = maskA[(long long)b * plane + i] != 0;
    return sc[b].a_is_dark ? a : !a;
  }
};
struct PostRoiEff {
  const uint8_t* roi; const uint8_t* sat; const TempScal* sc; long long plane; int w;
  __device__ int operator()(int b, int y, int x, int v) const {
    long long i = (long long)y * w + x;
    return v && roi[i] && !sat_at(sc[b], sat + (long long)b * plane, i);
  }
};

// ------------------------------------------------------------------ T7-T11: features, models, fusion
struct DevModel {
  int nfeat, nterms, niso, canon;       // canon = degree (2 / 3) when the terms are in canonical sklearn order, else 0
  float mean[4], scale[4];
  int prev[VISTAF_MAX_TERMS], feat[VISTAF_MAX_TERMS];   // XP[t] = XP[prev[t]] * z[feat[t]]  (sklearn build order)
  double coef[VISTAF_MAX_TERMS];
  double intercept;
  double iso_x[VISTAF_MAX_ISO], iso_y[VISTAF_MAX_ISO];
};

// StandardScaler (float32) -> PolynomialFeatures (float32 products) -> dot (float64) -> float32
// [-> isotonic clip + linear interpolation in float64 -> float32]   (temperature_sensor.py:236-240)
// Canonical PolynomialFeatures(degree DEG, include_bias) order over NF features, fully unrolled: every coefficient
// index is a compile-time constant and the term products stay in registers.  Same products and the same
// accumulation order as the table-driven loop below (term (i<=j<=k) = (z_j * z_k) * z_i).
template <int NF, int DEG>
__device__ __forceinline__ double poly_dot(const DevModel* __restrict__ M, const float* z) {
  double acc = M->intercept + M->coef[0];
  int t = 1;
#pragma unroll
  for (int i = 0; i < NF; i++) acc = fma(M->coef[t++], (double)z[i], acc);
  float p2[NF][NF];
#pragma unroll
  for (int i = 0; i < NF; i++)
#pragma unroll
    for (int j = i; j < NF; j++) {
      p2[i][j] = __fmul_rn(z[j], z[i]);
      acc = fma(M->coef[t++], (double)p2[i][j], acc);
    }
  if (DEG >= 3) {
#pragma unroll
    for (int i = 0; i < NF; i++)
#pragma unroll
      for (int j = i; j < NF; j++)
#pragma unroll
        for (int k = j; k < NF; k++) acc = fma(M->coef[t++], (double)__fmul_rn(p2[j][k], z[i]), acc);
  }
  return acc;
}

__device__ float model_predict(const DevModel* __restrict__ M, const float* x) {
  float z[4];
  const int nf = M->nfeat, nt = M->nterms;
  for (int i = 0; i < 4; i++) z[i] = i < nf ? __fdiv_rn(__fsub_rn(x[i], M->mean[i]), M->scale[i]) : 0.f;
  double acc;
  const int canon = M->canon;
  if (canon == 3 && nf == 4) acc = poly_dot<4, 3>(M, z);
  else if (canon == 2 && nf == 3) acc = poly_dot<3, 2>(M, z);
  else if (canon == 2 && nf == 4) acc = poly_dot<4, 2>(M, z);
  else if (canon == 3 && nf == 3) acc = poly_dot<3, 3>(M, z);
  else {
    float XP[VISTAF_MAX_TERMS];
    XP[0] = 1.0f;
    acc = M->intercept + M->coef[0];
    for (int t = 1; t < nt; t++) {
      float v = M->prev[t] == 0 ? z[M->feat[t]] : __fmul_rn(XP[M->prev[t]], z[M->feat[t]]);
      XP[t] = v;
      acc = fma(M->coef[t], (double)v, acc);
    }
  }
  float pred = (float)acc;
  if (M->niso > 0) {
    const int n = M->niso;
    double v = fmin(fmax((double)pred, M->iso_x[0]), M->iso_x[n - 1]);
    int lo = 0, hi = n - 1;                       // largest j with iso_x[j] <= v
    while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (M->iso_x[mid] <= v) lo = mid; else hi = mid; }
    if (v >= M->iso_x[n - 1]) pred = (float)M->iso_y[n - 1];
    else if (v == M->iso_x[lo]) pred = (float)M->iso_y[lo];
    else {
      double slope = (M->iso_y[lo + 1] - M->iso_y[lo]) / (M->iso_x[lo + 1] - M->iso_x[lo]);
      pred = (float)(slope * (v - M->iso_x[lo]) + M->iso_y[lo]);
    }
  }
  return pred;
}

struct FeatArgs {
  const uint8_t* frames; const uint8_t* roi; const uint8_t* sat; const uint8_t* dark;
  TempScal* sc; const DevModel* models; const int* gam; const int* cbr;
  float* fused; uint8_t *o_sat, *o_light, *o_support, *o_colorok, *o_source;
  int h, w;
  int bx0, bx1, by0, by1;        // ROI bounding box: tiles outside it only write the NaN / zero background
  float color_lo, color_hi, ok_lo, ok_hi, blend_lo, blend_hi, final_lo, final_hi, chroma_min;
  // T10 (saturated frames that own a slot): raw wide-model map and clamped colour map on the ROI window
  float *t10_wide, *t10_color; const int* t10_frame; unsigned int* t10_stat;
  int wx0, wy0, ww, wh;
  int tile_x0, tile_y0;          // first tile of the launch grid (only the tiles over the ROI window are launched then)
  int skip_background;           // the output planes already hold NaN / 0 outside the ROI bounding box
};

// fusion of the wide and colour maps at one pixel (temperature_sensor.py:594-641); wide / colour are clamped already
__device__ __forceinline__ float fuse_px(const FeatArgs& a, float wide, float color, bool support, int& cok, int& source) {
  cok = support && color >= a.ok_lo && color <= a.ok_hi;
  source = 0;
  float fused = wide;
  if (cok) { fused = color; source = 255; }
  if (wide == wide && cok && wide > a.blend_lo && wide < a.blend_hi) {
    float wg = __fdiv_rn(__fsub_rn(a.blend_hi, wide), __fsub_rn(a.blend_hi, a.blend_lo));
    wg = fminf(fmaxf(wg, 0.f), 1.f);
    fused = __fadd_rn(__fmul_rn(wg, color), __fmul_rn(__fsub_rn(1.0f, wg), wide));
    source = 128;
  }
  if (fused == fused) fused = fminf(fmaxf(fused, a.final_lo), a.final_hi);
  return fused;
}
__device__ __forceinline__ unsigned int f32_key(float f) {       // order-preserving uint key of a float
  const unsigned int u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_f32(unsigned int k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

#define FT_W 32
#define FT_H 16
// T10 = false: every frame of the chunk (grid.z = frame), all outputs.  T10 = true: the saturated frames only (grid.z =
// T10 slot): the same per-pixel features and models, but the only products are the raw wide map and the clamped colour
// map on the slot's ROI window plus min / max of the raw wide map -- the inputs of k_t10_quant / k_t10_refuse.
template <bool T10>
__global__ void __launch_bounds__(FT_W * FT_H) k_temp_features(FeatArgs a) {
  // planar BGR tile with a 2-pixel halo, and its horizontal (1 4 6 4 1) sums: the 5x5 kernel is an outer product of
  // integers, so rows-then-columns gives the same integer as the 25-tap sum with 10 taps per channel
  __shared__ uint8_t tile[3][(FT_H + 4) * (FT_W + 4)];
  __shared__ unsigned short hs[3][(FT_H + 4) * FT_W];
  const int b = T10 ? a.t10_frame[blockIdx.z] : (int)blockIdx.z;
  if (T10 && b < 0) return;
  const int x0 = ((int)blockIdx.x + a.tile_x0) * FT_W, y0 = ((int)blockIdx.y + a.tile_y0) * FT_H, h = a.h, w = a.w;
  const size_t plane = (size_t)h * w;
  const uint8_t* img = a.frames + (size_t)b * plane * 3;
  const int tid = threadIdx.y * FT_W + threadIdx.x;
  if (x0 >= a.bx1 || x0 + FT_W <= a.bx0 || y0 >= a.by1 || y0 + FT_H <= a.by0) {
    if (T10 || a.skip_background) return;
    const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
    if (x < w && y < h) {
      const size_t bi = (size_t)b * plane + (size_t)y * w + x;
      a.fused[bi] = __int_as_float(0x7fc00000);
      if (a.o_sat) a.o_sat[bi] = 0;
      if (a.o_light) a.o_light[bi] = 0;
      if (a.o_support) a.o_support[bi] = 0;
      if (a.o_colorok) a.o_colorok[bi] = 0;
      if (a.o_source) a.o_source[bi] = 0;
    }
    return;
  }
  // 5x5 Gaussian on uint8 uses BORDER_REFLECT_101 (cv2.GaussianBlur default)
  for (int e = tid; e < (FT_H + 4) * (FT_W + 4); e += FT_W * FT_H) {
    int ty = e / (FT_W + 4), tx = e - ty * (FT_W + 4);
    int y = reflect101(y0 - 2 + ty, h), x = reflect101(x0 - 2 + tx, w);
    const uint8_t* p = img + ((size_t)y * w + x) * 3;
    tile[0][e] = p[0]; tile[1][e] = p[1]; tile[2][e] = p[2];
  }
  __syncthreads();
  for (int e = tid; e < (FT_H + 4) * FT_W; e += FT_W * FT_H) {
    const int ty = e / FT_W, tx = e - ty * FT_W;
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const uint8_t* r = &tile[c][ty * (FT_W + 4) + tx];
      hs[c][e] = (unsigned short)(r[0] + 4 * r[1] + 6 * r[2] + 4 * r[3] + r[4]);
    }
  }
  __syncthreads();
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  int n_ok = 0, n_dark = 0, slot = -1;
  float wlo = 3.0e38f, whi = -3.0e38f;
  if (x < w && y < h) {
    const size_t i = (size_t)y * w + x, bi = (size_t)b * plane + i;
    const TempScal& S = a.sc[b];
    const bool in_roi = a.roi[i] != 0;
    const bool is_sat = in_roi && sat_at(S, a.sat + (size_t)b * plane, i);
    slot = (T10 || S.sat_count) ? S.t10_slot : -1;
    float fused = __int_as_float(0x7fc00000);
    int light = 0, support = 0, cok = 0, source = 0;
    if (in_roi) {
      const bool roi_eff = !is_sat;
      const int is_dark = a.dark[bi] != 0;
      light = roi_eff && !is_dark;
      n_dark += is_dark;
      // blurred BGR: ((1 4 6 4 1)x(1 4 6 4 1) * p + 128) >> 8
      int s[3];
#pragma unroll
      for (int c = 0; c < 3; c++) {
        const unsigned short* q = &hs[c][threadIdx.y * FT_W + threadIdx.x];
        s[c] = (int)q[0] + 4 * (int)q[FT_W] + 6 * (int)q[2 * FT_W] + 4 * (int)q[3 * FT_W] + (int)q[4 * FT_W];
      }
      const int B = (s[0] + 128) >> 8, Gc = (s[1] + 128) >> 8, Rc = (s[2] + 128) >> 8;
      // OpenCV 8-bit BGR2Lab, integer path (tables verified on all 2^24 colours, tools/make_lab_lut.py)
      const int R = __ldg(a.gam + Rc), G = __ldg(a.gam + Gc), Bl = __ldg(a.gam + B);
      const int fX = __ldg(a.cbr + ((R * 1777 + G * 1541 + Bl * 778 + 2048) >> 12));
      const int fY = __ldg(a.cbr + ((R * 871 + G * 2929 + Bl * 296 + 2048) >> 12));
      const int fZ = __ldg(a.cbr + ((R * 73 + G * 448 + Bl * 3575 + 2048) >> 12));
      int L = (296 * fY - 1336934 + 16384) >> 15;
      int A = (500 * (fX - fY) + 128 * 32768 + 16384) >> 15;
      int Bb = (200 * (fY - fZ) + 128 * 32768 + 16384) >> 15;
      L = min(max(L, 0), 255); A = min(max(A, 0), 255); Bb = min(max(Bb, 0), 255);
      const int gray = gray8_of(B, Gc, Rc);
      const float feat[4] = {(float)L, (float)A, (float)Bb, (float)gray};
      // colour support: dilate(light, 3x3 cross) & roi_eff & ~sat & chroma >= 10  (temperature_sensor.py:795-799)
      int light_d = light;
      if (!light_d && roi_eff) {
        const int nx[4] = {x - 1, x + 1, x, x}, ny[4] = {y, y, y - 1, y + 1};
        for (int q = 0; q < 4 && !light_d; q++) {
          if (nx[q] < 0 || nx[q] >= w || ny[q] < 0 || ny[q] >= h) continue;
          size_t j = (size_t)ny[q] * w + nx[q];
          if (a.roi[j] && !sat_at(S, a.sat + (size_t)b * plane, j) && !a.dark[(size_t)b * plane + j]) light_d = 1;
        }
      }
      const int ca = A - 128, cb = Bb - 128;
      support = light_d && roi_eff && (sqrtf((float)(ca * ca + cb * cb)) >= a.chroma_min);
      float wide = __int_as_float(0x7fc00000), color = wide, wraw = wide;
      if (roi_eff) { wraw = model_predict(a.models + 1, feat); wide = fminf(fmaxf(wraw, a.final_lo), a.final_hi); }
      if (support) color = fminf(fmaxf(model_predict(a.models + 0, feat), a.color_lo), a.color_hi);
      if (T10) {
        const size_t wi = ((size_t)slot * a.wh + (y - a.wy0)) * a.ww + (x - a.wx0);
        a.t10_wide[wi] = wraw;
        a.t10_color[wi] = color;
        if (roi_eff) { wlo = wraw; whi = wraw; }
      } else {
        fused = fuse_px(a, wide, color, support, cok, source);
        if (slot >= 0) cok = 0;     // saturated frame: k_t10_refuse redoes the fusion from the Telea-filled wide map
        n_ok += cok;
      }
    }
    if (!T10) {
      a.fused[bi] = fused;
      if (a.o_sat) a.o_sat[bi] = is_sat;
      if (a.o_light) a.o_light[bi] = light;
      if (a.o_support) a.o_support[bi] = support;
      if (a.o_colorok) a.o_colorok[bi] = cok;
      if (a.o_source) a.o_source[bi] = source;
    }
  }
  if (T10) {                    // min / max of the raw wide map over roi_eff (temperature_sensor.py:557)
    wlo = warp_min(wlo); whi = warp_max(whi);
    if ((tid & 31) == 0 && wlo <= whi) {
      atomicMin(&a.sc[b].wmin_key, f32_key(wlo));
      atomicMax(&a.sc[b].wmax_key, f32_key(whi));
    }
    return;
  }
  n_ok = warp_sum(n_ok);
  n_dark = warp_sum(n_dark);
  if ((tid & 31) == 0) {
    if (n_ok) atomicAdd(&a.sc[b].color_ok_count, (unsigned)n_ok);
    if (n_dark) atomicAdd(&a.sc[b].dark_count, (unsigned)n_dark);
  }
}

// ------------------------------------------------------------------ T10: Telea fill of the wide map on saturated frames
// temperature_sensor.py:546-577: when a masked-in pixel is non-finite (= saturated pixels exist) the known values are
// quantised to 8 bits over [vmin, vmax], the holes are filled by cv2.inpaint(TELEA, r = 7) and EVERY inside value is
// replaced by the de-quantised byte.  Saturated frames are rare, so the step runs on `slots` compacted frames per chunk.
__global__ void k_t10_slots(TempScal* sc, int nb, int* slot_frame, unsigned int* slot_stat, int slots) {
  const int lane = threadIdx.x;                  // one warp: frames in order, 32 at a time
  int n = 0;
  for (int b0 = 0; b0 < nb; b0 += 32) {
    const int b = b0 + lane;
    const bool sat = b < nb && sc[b].sat_count != 0;
    const unsigned int bits = __ballot_sync(0xffffffffu, sat);
    if (sat) {
      const int at = n + __popc(bits & ((1u << lane) - 1u));
      if (at < slots) { slot_frame[at] = b; sc[b].t10_slot = at; }
      else sc[b].status |= VISTAF_ST_RERUN_ALONE;
    }
    n += __popc(bits);
  }
  for (int s = lane; s < slots; s += 32) { slot_stat[s] = 0; if (s >= n) slot_frame[s] = -1; }
}
// mode 0: quantise + Telea; 1: flat map (hole pixels take vmin, :562-565); 2: no known pixel (pass-through)
__device__ __forceinline__ int t10_mode(const TempScal& S, float& vmin, float& rng) {
  if (S.wmin_key > S.wmax_key) { vmin = 0.f; rng = 0.f; return 2; }
  vmin = key_f32(S.wmin_key);
  const double d = (double)key_f32(S.wmax_key) - (double)vmin;
  rng = (float)d;
  return d < 1e-6 ? 1 : 0;
}
__global__ void k_t10_quant(FeatArgs a) {
  const int s = blockIdx.z, b = a.t10_frame[s];
  if (b < 0) return;
  const int wx = blockIdx.x * blockDim.x + threadIdx.x, wy = blockIdx.y;
  if (wx >= a.ww) return;
  const TempScal& S = a.sc[b];
  float vmin, rng;
  const int mode = t10_mode(S, vmin, rng);
  const size_t i = (size_t)(wy + a.wy0) * a.w + wx + a.wx0, wi = ((size_t)s * a.wh + wy) * a.ww + wx;
  const bool in_roi = a.roi[i] != 0;
  const bool known = in_roi && !sat_at(S, a.sat + (size_t)b * a.h * a.w, i);
  float v = known ? a.t10_wide[wi] : 0.f;
  if (mode == 0) {
    if (known) v = truncf(fminf(fmaxf(__fmul_rn(__fdiv_rn(__fsub_rn(v, vmin), rng), 255.0f), 0.f), 255.f));
  } else if (mode == 1 && in_roi && !known) v = vmin;
  a.t10_wide[wi] = v;
}
struct T10Mask {                 // holes of a slot's window: roi & saturated
  const int* slot_frame; const uint8_t* roi; const uint8_t* sat; const TempScal* sc; size_t plane; int W, wx0, wy0;
  __device__ bool operator()(int s, int y, int x) const {
    const int b = slot_frame[s];
    if (b < 0) return false;
    const TempScal& S = sc[b];
    float vmin, rng;
    if (t10_mode(S, vmin, rng) != 0) return false;
    const size_t i = (size_t)(y + wy0) * W + x + wx0;
    return roi[i] && sat_at(S, sat + (size_t)b * plane, i);
  }
};
__global__ void k_t10_refuse(FeatArgs a) {
  const int s = blockIdx.z, b = a.t10_frame[s];
  if (b < 0) return;
  const int wx = blockIdx.x * blockDim.x + threadIdx.x, wy = blockIdx.y;
  TempScal& S = a.sc[b];
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && a.t10_stat[s]) atomicOr(&S.status, a.t10_stat[s]);
  int cok = 0;
  if (wx < a.ww) {
    float vmin, rng;
    const int mode = t10_mode(S, vmin, rng);
    const size_t plane = (size_t)a.h * a.w;
    const size_t i = (size_t)(wy + a.wy0) * a.w + wx + a.wx0, wi = ((size_t)s * a.wh + wy) * a.ww + wx;
    if (a.roi[i]) {
      float wide = a.t10_wide[wi];
      if (mode == 0) wide = __fadd_rn(__fmul_rn(__fdiv_rn(wide, 255.0f), rng), vmin);
      else if (mode == 2) wide = __int_as_float(0x7fc00000);
      else if (sat_at(S, a.sat + (size_t)b * plane, i)) wide = vmin;
      if (wide == wide) wide = fminf(fmaxf(wide, a.final_lo), a.final_hi);
      const float color = a.t10_color[wi];
      int source;
      const float fused = fuse_px(a, wide, color, color == color, cok, source);
      const size_t bi = (size_t)b * plane + i;
      a.fused[bi] = fused;
      if (a.o_colorok) a.o_colorok[bi] = cok;
      if (a.o_source) a.o_source[bi] = source;
    }
  }
  cok = warp_sum(cok);
  if ((threadIdx.x & 31) == 0 && cok) atomicAdd(&S.color_ok_count, (unsigned)cok);
}

// ------------------------------------------------------------------ T12: oriented smoothing + M1 statistics
// angle == 0 (all demo and synthetic frames): the rotations are identity warps, so the step reduces to an
// anisotropic Gaussian (sigma_x = across, sigma_y = along) of the map with NaN -> 0, then NaN outside the
// ROI and a clamp (temperature_sensor.py:705-742, 853-859).
struct FusedLoad {
  const float* fused; long long plane; int w;
  __device__ float operator()(int b, int y, int x) const {
    float v = fused[(long long)b * plane + (long long)y * w + x];
    return v == v ? v : 0.f;
  }
};
struct FinalStore {
  const uint8_t* roi; float* out; TempScal* sc; long long plane; int w; float lo, hi;
  double s1, s2; float mn, mx; unsigned int cnt; int skip_bg;
  __device__ void operator()(int b, int y, int x, float v) {
    long long i = (long long)y * w + x;
    float r = __int_as_float(0x7fc00000);
    // a frame with a tilted carrier gets its ROI values (and statistics) from the warpAffine path below (k_t12_back)
    if (roi[i] && (sc[b].status & (VISTAF_ST_ROTATED_CARRIER | VISTAF_ST_RERUN_ALONE)) != VISTAF_ST_ROTATED_CARRIER) {
      r = fminf(fmaxf(v, lo), hi);
      s1 += (double)r; s2 += (double)r * (double)r;
      mn = fminf(mn, r); mx = fmaxf(mx, r); cnt++;
    }
    out[(long long)b * plane + i] = r;
  }
  __device__ void outside(int b, int y, int x, float) {
    if (!skip_bg) out[(long long)b * plane + (long long)y * w + x] = __int_as_float(0x7fc00000);
  }
  __device__ void finish(int b) {
    __shared__ double sh[3 * 32];
    __shared__ float shm[2 * 32];
    double v[3] = {s1, s2, (double)cnt};
    block_sum<3>(v, sh);
    const int tid = threadIdx.y * blockDim.x + threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nw = (blockDim.x * blockDim.y + 31) >> 5;
    float a = warp_min(mn), c = warp_max(mx);
    __syncthreads();
    if (lane == 0) { shm[wid] = a; shm[32 + wid] = c; }
    __syncthreads();
    if (tid == 0) {
      for (int i = 1; i < nw; i++) { a = fminf(a, shm[i]); c = fmaxf(c, shm[32 + i]); }
      if (v[2] > 0.0) {
        atomicAdd(&sc[b].t_sum, v[0]); atomicAdd(&sc[b].t_sumsq, v[1]); atomicAdd(&sc[b].t_count, (unsigned)v[2]);
        atomicMin((int*)&sc[b].t_min, __float_as_int(a));      // temperatures are positive: int order == float order
        atomicMax((int*)&sc[b].t_max, __float_as_int(c));
      }
    }
  }
};
struct FinalVal {              // value stream for the median: final map over the ROI (rows of its bounding box only)
  const float* fin; const uint8_t* roi; long long plane; long long off;     // off = first bounding-box row * width
  __device__ bool operator()(int b, long long i, float& v) const {
    if (!roi[off + i]) return false;
    v = fin[(long long)b * plane + off + i];
    return v == v;
  }
  __device__ bool vec_ok() const {
    return ((plane | off) & 3) == 0 && sel_aligned16(fin) && ((unsigned long long)roi & 3ull) == 0;
  }
  __device__ unsigned int vec4(int b, long long i, long long n, float v[4]) const {
    const uchar4 mm = *reinterpret_cast<const uchar4*>(roi + off + i);
    unsigned int m = ((mm.x ? 1u : 0u) | (mm.y ? 2u : 0u) | (mm.z ? 4u : 0u) | (mm.w ? 8u : 0u)) & sel_tail_mask(i, n);
    if (!m) return 0u;
    const float4 f = *reinterpret_cast<const float4*>(fin + (long long)b * plane + off + i);
    v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (!(v[k] == v[k])) m &= ~(1u << k);
    return m;
  }
};

// ------------------------------------------------------------------ T12 on a tilted carrier (temperature_sensor.py:705-742)
// angle = atan2(dy, dx) of the integer carrier peak; the map is rotated by cv2.warpAffine (INTER_LINEAR, BORDER_REFLECT)
// so the stripes' normal lies along +x, blurred with sigma (across, along), rotated back; the ROI takes the same two
// warps with INTER_NEAREST / constant 0 and blanks what falls outside.  cv2.warpAffine arithmetic (imgwarp.cpp,
// WarpAffineInvoker), reproduced here: the 2x3 matrix is inverted in float64, source coordinates are 22.10 fixed point
// (X0 = rint((M1 y + M2) 1024) + delta, adelta = rint(M0 x 1024)), the bilinear weights are the 1/32-pixel table
// (1 - a/32, a/32) multiplied in float32, the four taps summed left to right.  Verified bit-exact against cv2 4.13 on
// the host restatement (tests/test_temperature_gpu.py compares the device result with cv2 through the oracle).
__device__ __forceinline__ void t12_rotation(double cx, double cy, double angle_deg, double* m) {
  const double a = __dmul_rn(angle_deg, VISTAF_PI / 180.0);
  const double alpha = cos(a), beta = sin(a);
  double M0 = alpha, M1 = beta, M2 = __dsub_rn(__dmul_rn(__dsub_rn(1.0, alpha), cx), __dmul_rn(beta, cy));
  double M3 = -beta, M4 = alpha, M5 = __dadd_rn(__dmul_rn(beta, cx), __dmul_rn(__dsub_rn(1.0, alpha), cy));
  // warpAffine without WARP_INVERSE_MAP inverts the matrix first
  double D = __dsub_rn(__dmul_rn(M0, M4), __dmul_rn(M1, M3));
  D = D != 0.0 ? __ddiv_rn(1.0, D) : 0.0;
  const double A11 = __dmul_rn(M4, D), A22 = __dmul_rn(M0, D);
  M0 = A11; M1 = __dmul_rn(M1, -D); M3 = __dmul_rn(M3, -D); M4 = A22;
  const double b1 = __dsub_rn(__dmul_rn(-M0, M2), __dmul_rn(M1, M5));
  const double b2 = __dsub_rn(__dmul_rn(-M3, M2), __dmul_rn(M4, M5));
  m[0] = M0; m[1] = M1; m[2] = b1; m[3] = M3; m[4] = M4; m[5] = b2;
}
// tilted frames of a chunk -> slots (frame list + 0/1 gate per slot), like k_t10_slots
__global__ void k_t12_slots(TempScal* sc, int nb, int* slot_frame, unsigned int* slot_on, int slots) {
  const int lane = threadIdx.x;
  int n = 0;
  for (int b0 = 0; b0 < nb; b0 += 32) {
    const int b = b0 + lane;
    const bool rot = b < nb && (sc[b].status & VISTAF_ST_ROTATED_CARRIER) != 0;
    const unsigned int bits = __ballot_sync(0xffffffffu, rot);
    if (rot) {
      const int at = n + __popc(bits & ((1u << lane) - 1u));
      if (at < slots) slot_frame[at] = b;
      else sc[b].status |= VISTAF_ST_RERUN_ALONE;      // its final map keeps the axis-aligned blur until it is re-run
    }
    n += __popc(bits);
  }
  for (int s = lane; s < slots; s += 32) { slot_on[s] = s < n ? 1u : 0u; if (s >= n) slot_frame[s] = -1; }
}
__global__ void k_t12_setup(TempScal* sc, int h, int w) {
  TempScal& s = sc[blockIdx.x];
  if (!(s.status & VISTAF_ST_ROTATED_CARRIER)) return;
  const double dx = (double)(s.peak_x - w / 2), dy = (double)(s.peak_y - h / 2);
  const double angle_rad = atan2(dy, dx);
  const double deg = __ddiv_rn(__dmul_rn(-angle_rad, 180.0), VISTAF_PI);
  t12_rotation((double)((float)w / 2.0f), (double)((float)h / 2.0f), deg, s.aff_f);
  t12_rotation((double)((float)w / 2.0f), (double)((float)h / 2.0f), -deg, s.aff_b);
}
// fixed-point source coordinates of destination pixel (x, y): `shift` = 5 (bilinear, 1/32 px) or 10 (nearest)
__device__ __forceinline__ void t12_src(const double* m, int x, int y, int shift, int delta, int& X, int& Y) {
  const int ad = __double2int_rn(__dmul_rn(__dmul_rn(m[0], (double)x), 1024.0));
  const int bd = __double2int_rn(__dmul_rn(__dmul_rn(m[3], (double)x), 1024.0));
  const int X0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(m[1], (double)y), m[2]), 1024.0)) + delta;
  const int Y0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(m[4], (double)y), m[5]), 1024.0)) + delta;
  X = (X0 + ad) >> shift; Y = (Y0 + bd) >> shift;
}
template <class Src>
__device__ __forceinline__ float t12_bilinear(const Src& src, const double* m, int x, int y, int h, int w) {
  int X, Y;
  t12_src(m, x, y, 5, 16, X, Y);
  const int sx = X >> 5, sy = Y >> 5;
  const float fx = (float)(X & 31) * (1.0f / 32.0f), fy = (float)(Y & 31) * (1.0f / 32.0f);
  const float tx0 = __fsub_rn(1.0f, fx), ty0 = __fsub_rn(1.0f, fy);
  const int xa = reflect(sx, w), xb = reflect(sx + 1, w), ya = reflect(sy, h), yb = reflect(sy + 1, h);
  float v = __fmul_rn(src(ya, xa), __fmul_rn(ty0, tx0));
  v = __fadd_rn(v, __fmul_rn(src(ya, xb), __fmul_rn(ty0, fx)));
  v = __fadd_rn(v, __fmul_rn(src(yb, xa), __fmul_rn(fy, tx0)));
  v = __fadd_rn(v, __fmul_rn(src(yb, xb), __fmul_rn(fy, fx)));
  return v;
}
struct T12Fused {               // map0 = fused with non-finite -> 0
  const float* p; int w;
  __device__ float operator()(int y, int x) const { const float v = p[(size_t)y * w + x]; return fabsf(v) <= 3.4e38f ? v : 0.f; }
};
struct T12Plane { const float* p; int w; __device__ float operator()(int y, int x) const { return p[(size_t)y * w + x]; } };
__global__ void __launch_bounds__(256) k_t12_forward(const float* __restrict__ fused, float* __restrict__ rot, const TempScal* sc,
                                                      const int* __restrict__ slot_frame, int h, int w) {
  const int s = blockIdx.z, b = slot_frame[s], y = blockIdx.y, x = blockIdx.x * 256 + threadIdx.x;
  if (b < 0 || x >= w) return;
  rot[(size_t)s * h * w + (size_t)y * w + x] = t12_bilinear(T12Fused{fused + (size_t)b * h * w, w}, sc[b].aff_f, x, y, h, w);
}
// nearest-neighbour warp of the ROI, twice (forward then back), evaluated on the fly
__device__ __forceinline__ bool t12_back_roi(const uint8_t* __restrict__ roi, const TempScal& S, int x, int y, int h, int w) {
  int X, Y;
  t12_src(S.aff_b, x, y, 10, 512, X, Y);            // pixel of rot_roi read by back_roi(x, y)
  if (X < 0 || X >= w || Y < 0 || Y >= h) return false;
  int X2, Y2;
  t12_src(S.aff_f, X, Y, 10, 512, X2, Y2);          // pixel of roi read by rot_roi(X, Y)
  if (X2 < 0 || X2 >= w || Y2 < 0 || Y2 >= h) return false;
  return roi[(size_t)Y2 * w + X2] != 0;
}
__global__ void __launch_bounds__(256) k_t12_back(const float* __restrict__ blurred, const uint8_t* __restrict__ roi,
                                                   float* __restrict__ fin, TempScal* sc, const int* __restrict__ slot_frame, int h,
                                                   int w, int bx0, int bx1, int by0, float lo, float hi) {
  __shared__ double sh[3 * 32];
  __shared__ float shm[2 * 32];
  const int slot = blockIdx.z, b = slot_frame[slot], y = by0 + blockIdx.y, x = bx0 + blockIdx.x * 256 + threadIdx.x;
  if (b < 0) return;
  TempScal& S = sc[b];
  double v[3] = {0.0, 0.0, 0.0};
  float mn = 3.0e38f, mx = 0.f;
  if (x < bx1) {
    const size_t i = (size_t)y * w + x, fr = (size_t)b * h * w;
    float r = __int_as_float(0x7fc00000);
    if (roi[i] && t12_back_roi(roi, S, x, y, h, w)) {
      r = fminf(fmaxf(t12_bilinear(T12Plane{blurred + (size_t)slot * h * w, w}, S.aff_b, x, y, h, w), lo), hi);
      v[0] = (double)r; v[1] = (double)r * (double)r; v[2] = 1.0;
      mn = r; mx = r;
    }
    fin[fr + i] = r;
  }
  block_sum<3>(v, sh);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float a = warp_min(mn), c = warp_max(mx);
  __syncthreads();
  if (lane == 0) { shm[wid] = a; shm[32 + wid] = c; }
  __syncthreads();
  if (threadIdx.x == 0 && v[2] > 0.0) {
    for (int k = 1; k < 8; k++) { a = fminf(a, shm[k]); c = fmaxf(c, shm[32 + k]); }
    atomicAdd(&S.t_sum, v[0]); atomicAdd(&S.t_sumsq, v[1]); atomicAdd(&S.t_count, (unsigned)v[2]);
    atomicMin((int*)&S.t_min, __float_as_int(a));
    atomicMax((int*)&S.t_max, __float_as_int(c));
  }
}
struct T12Load { const float* p; long long plane; int w;
  __device__ float operator()(int b, int y, int x) const { return p[(long long)b * plane + (long long)y * w + x]; } };
struct T12Store { float* p; long long plane; int w;
  __device__ void operator()(int b, int y, int x, float v) { p[(long long)b * plane + (long long)y * w + x] = v; }
  __device__ void finish(int) {} };

__global__ void k_temp_init(TempScal* sc, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  TempScal z;
  memset(&z, 0, sizeof(z));
  z.t_min = 3.0e38f; z.t_max = 0.f;
  z.t10_slot = -1; z.wmin_key = 0xffffffffu; z.wmax_key = 0u;
  sc[i] = z;
}

__global__ void k_temp_records(const TempScal* sc, const SelectState* sel, VistafTempRecord* rec, int n, int cx, int cy) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const TempScal& s = sc[i];
  VistafTempRecord r;
  memset(&r, 0, sizeof(r));
  const double cnt = (double)s.t_count;
  if (s.t_count) {
    double mean = s.t_sum / cnt;
    double var = fmax(s.t_sumsq / cnt - mean * mean, 0.0);
    r.mean_C = (float)mean; r.std_C = (float)sqrt(var); r.min_C = s.t_min; r.max_C = s.t_max;
    r.median_C = sel[i].result;
  } else {
    r.mean_C = r.std_C = r.min_C = r.max_C = r.median_C = __int_as_float(0x7fc00000);
  }
  r.valid_pixels = (int)s.t_count;
  r.peak_x = s.peak_x; r.peak_y = s.peak_y;
  r.phi0 = (float)atan2(s.csum[1], s.csum[0]);
  r.mean_gray_a = s.cntA ? (float)(s.sumA / s.cntA) : 1e9f;
  r.mean_gray_b = s.cntB ? (float)(s.sumB / s.cntB) : 1e9f;
  r.median_gray = s.median;
  r.sat_pixels = (int)s.sat_count; r.roi_eff_pixels = (int)s.n_eff;
  r.dark_pixels = (int)s.dark_count; r.color_ok_pixels = (int)s.color_ok_count;
  r.status = (int)s.status | (s.sat_count ? VISTAF_ST_SATURATED : 0);      // SAT_UNFILLED comes from k_t10_slots
  rec[i] = r;
}

// ================================================================== host side
struct T10Win { int x0, y0, w, h; };
#define T10_RADIUS 7            // temperature_sensor.py:829 (wide map); the colour map never has holes inside its support
static T10Win t10_window(const VistafTempPlan* p) {       // ROI bounding box grown past the Telea ring
  const int m = T10_RADIUS + 5;
  const int x0 = std::max(0, p->x0 - m), x1 = std::min(p->W, p->x1 + m);
  const int y0 = std::max(0, p->y0 - m), y1 = std::min(p->H, p->y1 + m);
  return T10Win{x0, y0, x1 - x0, y1 - y0};
}
static void temp_layout(const VistafTempPlan* p, int chunk, TempWs& L) {
  const size_t npix = (size_t)p->H * p->W, c = (size_t)chunk;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o += align256(bytes); return r; };
  L.gray8 = take(c * npix);
  L.satraw = take(c * npix);
  L.sat = take(c * npix);
  L.norm = take(c * npix * 4);
  L.tmp = take(c * npix * 4);
  L.A = take(c * (size_t)p->H * p->Wh * 8);
  L.G = take(c * (size_t)p->H * MAXBAND * 8);
  L.maskA = take(c * npix);
  L.m1 = take(c * npix);
  L.m2 = take(c * npix);
  L.cand = take(c * (size_t)p->nblk_cols * TOPK * sizeof(Cand));
  L.scal = take(c * sizeof(TempScal));
  L.sel = take(sel_ws_bytes((int)c, TEMP_SEL_CAP));
  // T10 (and the warpAffine path of T12) work on at most an eighth of the chunk's frames at a time, compacted into slots
  // (saturated / tilted frames are rare; every slot costs a round of early-exit blocks on the other frames).  A frame
  // beyond the slots comes back with VISTAF_ST_RERUN_ALONE.
  const T10Win wn = t10_window(p);
  const size_t ts = (size_t)((chunk + 7) / 8), wpix = (size_t)wn.h * wn.w;
  L.t10_slots = (int)ts;
  L.t10_frame = take(ts * 4);
  L.t10_stat = take(ts * 4);
  L.t10_wide = take(ts * wpix * 4);
  L.t10_color = take(ts * wpix * 4);
  L.t10_telea = take(telea_ws_bytes(wn.h, wn.w, (int)ts));
  L.t12_frame = take(ts * 4 * 2);
  L.total = o;
}

static int build_dev_model(const VistafPolyModel& m, DevModel& d) {
  memset(&d, 0, sizeof(d));
  d.nfeat = m.n_features; d.nterms = m.n_terms; d.niso = m.n_iso;
  if (d.nfeat < 1 || d.nfeat > 4 || d.nterms < 1 || d.nterms > VISTAF_MAX_TERMS || d.niso > VISTAF_MAX_ISO)
    return vistaf_set_err("model shape out of range", cudaErrorInvalidValue, __FILE__, __LINE__);
  for (int i = 0; i < d.nfeat; i++) { d.mean[i] = (float)m.mean[i]; d.scale[i] = (float)m.scale[i]; }
  for (int t = 0; t < d.nterms; t++) {
    d.coef[t] = m.coef[t];
    int deg = 0, low = -1;
    for (int f = 0; f < d.nfeat; f++) { deg += m.powers[t][f]; if (low < 0 && m.powers[t][f] > 0) low = f; }
    if (deg == 0) { if (t != 0) return vistaf_set_err("bias term must be first", cudaErrorInvalidValue, __FILE__, __LINE__); continue; }
    int want[4] = {0, 0, 0, 0};
    for (int f = 0; f < d.nfeat; f++) want[f] = m.powers[t][f] - (f == low ? 1 : 0);
    int prev = -1;
    for (int u = 0; u < t; u++) {
      bool eq = true;
      for (int f = 0; f < d.nfeat; f++) eq = eq && m.powers[u][f] == want[f];
      if (eq) { prev = u; break; }
    }
    if (prev < 0) return vistaf_set_err("polynomial terms not in PolynomialFeatures order", cudaErrorInvalidValue, __FILE__, __LINE__);
    d.prev[t] = prev; d.feat[t] = low;
  }
  // canonical order check (enables the unrolled evaluation): bias, degree 1, degree 2 (i <= j), degree 3 (i <= j <= k)
  for (int deg = 2; deg <= 3 && !d.canon; deg++) {
    const int nf = d.nfeat;
    std::vector<int> cp(1, 0), cf(1, 0);
    std::vector<int> idx2(16, -1);
    for (int i = 0; i < nf; i++) { cp.push_back(0); cf.push_back(i); }
    for (int i = 0; i < nf; i++)
      for (int j = i; j < nf; j++) { idx2[i * 4 + j] = (int)cp.size(); cp.push_back(1 + j); cf.push_back(i); }
    if (deg == 3)
      for (int i = 0; i < nf; i++)
        for (int j = i; j < nf; j++)
          for (int k = j; k < nf; k++) { cp.push_back(idx2[j * 4 + k]); cf.push_back(i); }
    if ((int)cp.size() != d.nterms) continue;
    bool same = true;
    for (int t = 1; t < d.nterms; t++) same = same && cp[t] == d.prev[t] && cf[t] == d.feat[t];
    if (same) d.canon = deg;
  }
  d.intercept = m.intercept;
  for (int i = 0; i < d.niso; i++) { d.iso_x[i] = m.iso_x[i]; d.iso_y[i] = m.iso_y[i]; }
  return 0;
}

extern "C" int vistaf_temp_plan_create(const VistafTempParams* P, VistafTempPlan** out) {
  if (!P || !out) return vistaf_set_err("null argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  if (P->width % 4 || P->band_radius * 2 + 1 > MAXBAND || P->sat_dilate > MORPH_MAXROWS || P->close_ky > MORPH_MAXROWS)
    return vistaf_set_err("unsupported temperature parameters", cudaErrorInvalidValue, __FILE__, __LINE__);
  VistafTempPlan* p = new (std::nothrow) VistafTempPlan();
  if (!p) return vistaf_set_err("out of host memory", cudaErrorMemoryAllocation, __FILE__, __LINE__);
  memset(p, 0, sizeof(*p));
  p->P = *P;
  p->H = P->height; p->W = P->width; p->Wh = P->width / 2 + 1;
  p->nblk_cols = cdiv(p->Wh, COLS_PER_BLOCK);
  if (fft1d_create(&p->fw, p->W) || fft1d_create(&p->fh, p->H)) return 1;
  if (ROWS_SMEM(p->W) > 200 * 1024 || COLS_SMEM(p->H) > 220 * 1024 ||
      COLS_PER_BLOCK * p->H > COLS_PT * COLS_THREADS)
    return vistaf_set_err("frame too large for the shared-memory FFT", cudaErrorInvalidValue, __FILE__, __LINE__);
  p->g_illum = make_gauss(P->illum_sigma);
  p->g_across = make_gauss(P->sigma_across);
  p->g_along = make_gauss(P->sigma_along);
  p->se_sat = se_ellipse(P->sat_dilate | 1);
  const size_t npix = (size_t)p->H * p->W;
  CK(cudaMalloc(&p->roi, npix));
  VL(), k_roi_mask<<<dim3(cdiv(p->W, 256), p->H), 256>>>(p->roi, p->H, p->W, P->roi_cx, P->roi_cy, P->roi_r * P->roi_r);
  CKL();
  uint8_t* hroi = new uint8_t[npix];
  CK(cudaMemcpy(hroi, p->roi, npix, cudaMemcpyDeviceToHost));
  p->y0 = p->H; p->y1 = 0; p->x0 = p->W; p->x1 = 0;
  for (int y = 0; y < p->H; y++)
    for (int x = 0; x < p->W; x++)
      if (hroi[(size_t)y * p->W + x]) {
        if (y < p->y0) p->y0 = y;
        if (y + 1 > p->y1) p->y1 = y + 1;
        if (x < p->x0) p->x0 = x;
        if (x + 1 > p->x1) p->x1 = x + 1;
      }
  delete[] hroi;
  if (p->y1 <= p->y0) { p->y0 = 0; p->y1 = 1; p->x0 = 0; p->x1 = 1; }
  DevModel dm[2];
  if (build_dev_model(P->color_model, dm[0]) || build_dev_model(P->wide_model, dm[1])) return 1;
  CK(cudaMalloc(&p->models, sizeof(dm)));
  CK(cudaMemcpy(p->models, dm, sizeof(dm), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&p->lab_gamma, sizeof(int) * 256));
  CK(cudaMalloc(&p->lab_cbrt, sizeof(int) * 3072));
  CK(cudaMemcpy(p->lab_gamma, P->lab_gamma, sizeof(int) * 256, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(p->lab_cbrt, P->lab_cbrt, sizeof(int) * 3072, cudaMemcpyHostToDevice));
  // opt-in shared memory for the FFT kernels
  CK(cudaFuncSetAttribute((const void*)k_temp_fft_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ROWS_SMEM(p->W)));
  CK(cudaFuncSetAttribute((const void*)k_temp_ifft_rows<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ROWS_SMEM(p->W)));
  CK(cudaFuncSetAttribute((const void*)k_temp_ifft_rows<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ROWS_SMEM(p->W)));
  CK(cudaFuncSetAttribute((const void*)k_temp_fft_cols_peaks, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          (int)COLS_SMEM(p->H)));
  CK(cudaFuncSetAttribute((const void*)k_temp_peak_select, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          2 * p->nblk_cols * TOPK * 4));
  CK(cudaDeviceSynchronize());
  *out = p;
  return 0;
}

extern "C" void vistaf_temp_plan_destroy(VistafTempPlan* p) {
  if (!p) return;
  vistaf_graph_drop(p);
  fft1d_destroy(&p->fw); fft1d_destroy(&p->fh);
  cudaFree(p->roi); cudaFree(p->models); cudaFree(p->lab_gamma); cudaFree(p->lab_cbrt);
  delete p;
}

extern "C" size_t vistaf_temp_workspace_bytes(const VistafTempPlan* p, int chunk) {
  TempWs L;
  temp_layout(p, chunk < 1 ? 1 : chunk, L);
  return L.total;
}

extern "C" int vistaf_temp_roi_mask(const VistafTempPlan* p, uint8_t* dst, void* stream) {
  CK(cudaMemcpyAsync(dst, p->roi, (size_t)p->H * p->W, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}

extern "C" int vistaf_temp_workspace_plane(const VistafTempPlan* p, int chunk, const char* name, size_t* off, size_t* bytes) {
  TempWs L;
  temp_layout(p, chunk, L);
  const size_t npix = (size_t)p->H * p->W;
  struct { const char* n; size_t o, b; } tab[] = {
      {"gray8", L.gray8, npix}, {"satraw", L.satraw, npix}, {"sat", L.sat, npix}, {"norm", L.norm, npix * 4},
      {"A", L.A, (size_t)p->H * p->Wh * 8}, {"G", L.G, (size_t)p->H * MAXBAND * 8}, {"maskA", L.maskA, npix},
      {"scal", L.scal, sizeof(TempScal)}, {"cand", L.cand, (size_t)p->nblk_cols * TOPK * sizeof(Cand)}};
  for (auto& t : tab)
    if (!strcmp(t.n, name)) { *off = t.o; *bytes = t.b; return 0; }
  return vistaf_set_err("unknown workspace plane", cudaErrorInvalidValue, __FILE__, __LINE__);
}

static int temp_chunk(VistafTempPlan* p, const uint8_t* frames, int nb, const VistafTempOutputs& O, size_t f0,
                      char* ws, const TempWs& L, cudaStream_t st) {
  const VistafTempParams& P = p->P;
  const int H = p->H, W = p->W, Wh = p->Wh;
  const int npix = H * W;
  const long long plane = (long long)npix;
  uint8_t* gray8 = (uint8_t*)(ws + L.gray8);
  uint8_t* satraw = (uint8_t*)(ws + L.satraw);
  uint8_t* sat = (uint8_t*)(ws + L.sat);
  float* norm = (float*)(ws + L.norm);
  float* tmp = (float*)(ws + L.tmp);
  float2* A = (float2*)(ws + L.A);
  float2* G = (float2*)(ws + L.G);
  uint8_t* maskA = (uint8_t*)(ws + L.maskA);
  uint8_t* m1 = (uint8_t*)(ws + L.m1);
  uint8_t* m2 = (uint8_t*)(ws + L.m2);
  Cand* cand = (Cand*)(ws + L.cand);
  TempScal* sc = (TempScal*)(ws + L.scal);
  SelectState* sel = (SelectState*)(ws + L.sel);
  const uint8_t* fr = frames + f0 * (size_t)npix * 3;

  vistaf_stage_mark("T.gray_sat_median", st);
  VL(), k_temp_init<<<cdiv(nb, 128), 128, 0, st>>>(sc, nb);
  VL(), k_temp_gray<<<dim3(cdiv(npix / 4, 256), nb), 256, 0, st>>>(fr, p->roi, gray8, satraw, sc, npix, P.sat_thresh);
  VL(), k_temp_sat_dilate<<<dim3(cdiv(W, 256), cdiv(p->y1 - p->y0, SATD_ROWS), nb), 256, 0, st>>>(satraw, p->roi, sat, sc, H, W, p->se_sat,
                                                                                         p->y0, p->y1);
  VL(), k_temp_hist<<<dim3(296, nb), 256, 0, st>>>(gray8, p->roi, sat, sc, npix, p->y0 * W, p->y1 * W);
  VL(), k_temp_median<<<nb, 32, 0, st>>>(sc);
  CKL();
  vistaf_stage_mark("T.illum_blur20", st);
  // illumination normalisation
  IllumLoad il{gray8, p->roi, sc, plane, W};
  IllumStore is{gray8, p->roi, sat, sc, norm, plane, W, 0.0};
  // outside the ROI the image is the per-frame median: only the ROI bounding box grown by the kernel radius is blurred,
  // everything else takes the constants R(median) / C(R(median)) (bit-identical to blurring the whole frame)
  VL(), k_temp_blur_consts<<<nb, 1, 0, st>>>(sc, p->g_illum);
  const int irad = p->g_illum.ksize >> 1;
  const BlurWindow iwin{p->x0 - irad, p->x1 + irad, p->y0 - irad, p->y1 + irad, p->y0, p->y1, &sc[0].blur_cc, &sc[0].blur_rc,
                        sizeof(TempScal)};
  if (gauss_blur(il, is, tmp, H, W, nb, p->g_illum, p->g_illum, st, &iwin)) return 1;
  VL(), k_temp_mu<<<nb, 1, 0, st>>>(sc);
  vistaf_stage_mark("T.fft_rows", st);
  // forward FFT + carrier peak
  VL(), k_temp_fft_rows<<<dim3(cdiv(H, 2), nb), 256, ROWS_SMEM(W), st>>>(norm, A, sc, p->fw, H, Wh);
  vistaf_stage_mark("T.fft_cols_peaks", st);
  VL(), k_temp_fft_cols_peaks<<<dim3(p->nblk_cols, nb), COLS_THREADS, COLS_SMEM(H), st>>>(A, cand, p->fh, W, Wh,
                                                                                                  P.dc_exclusion, sc);
  const int ncand = p->nblk_cols * TOPK;
  VL(), k_temp_peak_select<<<nb, 256, (size_t)2 * ncand * 4, st>>>(cand, sc, ncand, H, W, P.dc_exclusion, P.peak_max_dy_frac);
  CKL();
  vistaf_stage_mark("T.band_ifft", st);
  // band-pass + inverse
  const int R = P.band_radius, nbnd = 2 * R + 1;
  VL(), k_temp_band<<<dim3(nbnd, nb), 256, (size_t)H * 8, st>>>(A, G, sc, p->fh, W, Wh, R);
  const int nrows = p->y1 - p->y0;
  // the half spectrum A has been consumed by k_temp_band: its plane now holds the analytic signal over the ROI box
  float2* Z = A;
  const int zw = p->x1 - p->x0;
  if ((size_t)zw * nrows > (size_t)H * Wh) return vistaf_set_err("temperature: ROI box larger than the spectrum plane", cudaErrorInvalidValue, __FILE__, __LINE__);
  VL(), k_temp_ifft_rows<1><<<dim3(nrows, nb), 256, ROWS_SMEM(W), st>>>(G, norm, gray8, p->roi, sat, maskA, sc, p->fw, H, p->y0, R,
                                                                      Z, p->x0, zw, nrows);
  VL(), k_temp_phi0<<<nb, 1, 0, st>>>(sc);
  VL(), k_temp_split<<<dim3(cdiv(zw, 256), cdiv(nrows, SPLIT_ROWS), nb), 256, 0, st>>>(Z, gray8, p->roi, sat, maskA, sc, H, W, p->y0, p->x0, zw, nrows);
  VL(), k_temp_choose_dark<<<nb, 1, 0, st>>>(sc);
  CKL();
  vistaf_stage_mark("T.morph_close_open", st);
  // close RECT(kx x ky) then open RECT(kx x ky), then & roi_eff  (temperature_sensor.py:390-406, 490)
  DarkLoad dl{maskA, p->roi, sat, sc, plane, W, p->y0, p->y1};
  MorphSE se_c = se_rect(P.close_kx | 1, P.close_ky | 1), se_o = se_rect(P.open_kx | 1, P.open_ky | 1);
  uint8_t* dark = O.dark ? O.dark + f0 * (size_t)npix : maskA;      // final dark mask destination
  // every plane of this chain is 0 outside the ROI bounding box grown by the structuring elements (<= 15 + 3 rows)
  const int4 mwin = make_int4(p->x0 - 24, p->x1 + 24, p->y0 - 24, p->y1 + 24);
  // (the scratch planes m1 / m2 are never read outside the window: their far tiles are not written at all; the output
  // plane keeps its zeros there when the caller says its background is valid)
  const bool bg_ok = O.background_valid != 0;
  if (morph<true>(dl, PostId(), m1, H, W, nb, se_c, st, &mwin, true)) return 1;
  if (morph<false>(MaskLoad{m1, plane, W}, PostId(), m2, H, W, nb, se_c, st, &mwin, true)) return 1;
  if (morph<false>(MaskLoad{m2, plane, W}, PostId(), m1, H, W, nb, se_o, st, &mwin, true)) return 1;
  if (morph<true>(MaskLoad{m1, plane, W}, PostRoiEff{p->roi, sat, sc, plane, W}, dark, H, W, nb, se_o, st, &mwin, O.dark ? bg_ok : true)) return 1;
  vistaf_stage_mark("T.features_models_fuse", st);
  // features, models, fusion
  FeatArgs fa;
  fa.frames = fr; fa.roi = p->roi; fa.sat = sat; fa.dark = dark; fa.sc = sc;
  fa.models = (const DevModel*)p->models; fa.gam = p->lab_gamma; fa.cbr = p->lab_cbrt;
  fa.fused = O.fused + f0 * (size_t)npix;
  fa.o_sat = O.sat ? O.sat + f0 * (size_t)npix : nullptr;
  fa.o_light = O.light ? O.light + f0 * (size_t)npix : nullptr;
  fa.o_support = O.color_support ? O.color_support + f0 * (size_t)npix : nullptr;
  fa.o_colorok = O.color_ok ? O.color_ok + f0 * (size_t)npix : nullptr;
  fa.o_source = O.source ? O.source + f0 * (size_t)npix : nullptr;
  fa.h = H; fa.w = W;
  fa.bx0 = p->x0; fa.bx1 = p->x1; fa.by0 = p->y0; fa.by1 = p->y1;
  fa.color_lo = P.color_t_min - 5.0f; fa.color_hi = P.color_t_max + 5.0f;
  fa.ok_lo = P.color_t_min - P.guard_band; fa.ok_hi = P.color_t_max + P.guard_band;
  fa.blend_lo = P.color_t_max - P.switch_margin; fa.blend_hi = P.color_t_max + P.switch_margin;
  fa.final_lo = P.final_t_min; fa.final_hi = P.final_t_max; fa.chroma_min = P.chroma_min;
  const T10Win wn = t10_window(p);
  const int ts = L.t10_slots;
  fa.t10_wide = (float*)(ws + L.t10_wide); fa.t10_color = (float*)(ws + L.t10_color);
  int* t10_frame = (int*)(ws + L.t10_frame);
  fa.t10_frame = t10_frame; fa.t10_stat = (unsigned int*)(ws + L.t10_stat);
  fa.wx0 = wn.x0; fa.wy0 = wn.y0; fa.ww = wn.w; fa.wh = wn.h;
  fa.tile_x0 = 0; fa.tile_y0 = 0; fa.skip_background = bg_ok ? 1 : 0;
  VL(), k_t10_slots<<<1, 32, 0, st>>>(sc, nb, t10_frame, fa.t10_stat, ts);
  if (bg_ok) {                  // only the tiles over the ROI bounding box
    fa.tile_x0 = p->x0 / FT_W; fa.tile_y0 = p->y0 / FT_H;
    VL(), k_temp_features<false><<<dim3(cdiv(p->x1, FT_W) - fa.tile_x0, cdiv(p->y1, FT_H) - fa.tile_y0, nb), dim3(FT_W, FT_H), 0, st>>>(fa);
  } else {
    VL(), k_temp_features<false><<<dim3(cdiv(W, FT_W), cdiv(H, FT_H), nb), dim3(FT_W, FT_H), 0, st>>>(fa);
  }
  CKL();
  vistaf_stage_mark("T.sat_telea_fill", st);
  {
    const dim3 gw(cdiv(wn.w, 256), wn.h, ts);
    fa.tile_x0 = std::max(wn.x0, p->x0) / FT_W; fa.tile_y0 = std::max(wn.y0, p->y0) / FT_H;
    const int tx1 = cdiv(std::min(wn.x0 + wn.w, p->x1), FT_W), ty1 = cdiv(std::min(wn.y0 + wn.h, p->y1), FT_H);
    VL(), k_temp_features<true><<<dim3(tx1 - fa.tile_x0, ty1 - fa.tile_y0, ts), dim3(FT_W, FT_H), 0, st>>>(fa);
    VL(), k_t10_quant<<<gw, 256, 0, st>>>(fa);
    CKL();
    if (telea_inpaint(T10Mask{t10_frame, p->roi, sat, sc, (size_t)npix, W, wn.x0, wn.y0}, fa.t10_wide, (size_t)wn.h * wn.w,
                      fa.t10_stat, sizeof(unsigned int), ws + L.t10_telea, wn.h, wn.w, ts, T10_RADIUS, st, 1, t10_frame)) return 1;
    VL(), k_t10_refuse<<<gw, 256, 0, st>>>(fa);
    CKL();
  }
  vistaf_stage_mark("T.final_blur", st);
  // oriented smoothing (axis aligned) + statistics
  float* fin = O.final_map + f0 * (size_t)npix;
  FusedLoad fl{fa.fused, plane, W};
  FinalStore fs{p->roi, fin, sc, plane, W, P.final_t_min, P.final_t_max, 0.0, 0.0, 3.0e38f, 0.f, 0u, bg_ok ? 1 : 0};
  const BlurWindow fwin{p->x0, p->x1, p->y0, p->y1, 0, 0, nullptr, nullptr, 0};      // the final map is NaN outside the ROI whatever the blur gives
  // frames with a tilted carrier (status bit set by k_temp_peak_select) are compacted into slots first: the axis-aligned
  // pass below skips their statistics, the warpAffine pass after it redoes their ROI
  int* t12_frame = (int*)(ws + L.t12_frame);
  unsigned int* t12_on = (unsigned int*)(t12_frame + ts);
  VL(), k_t12_slots<<<1, 32, 0, st>>>(sc, nb, t12_frame, t12_on, ts);
  if (gauss_blur(fl, fs, tmp, H, W, nb, p->g_across, p->g_along, st, &fwin)) return 1;
  {
    // warp, blur, warp back over the full frame on slot planes carved from the illumination planes `norm` / `tmp`, free by
    // now.  Every block of an empty slot returns at once.
    VL(), k_t12_setup<<<nb, 1, 0, st>>>(sc, H, W);
    VL(), k_t12_forward<<<dim3(cdiv(W, 256), H, ts), 256, 0, st>>>(fa.fused, norm, sc, t12_frame, H, W);
    const BlurWindow gwin{0, W, 0, H, 0, H, nullptr, nullptr, sizeof(unsigned int), t12_on, 1u};
    if (gauss_blur(T12Load{norm, plane, W}, T12Store{norm, plane, W}, tmp, H, W, ts, p->g_across, p->g_along, st, &gwin)) return 1;
    VL(), k_t12_back<<<dim3(cdiv(p->x1 - p->x0, 256), p->y1 - p->y0, ts), 256, 0, st>>>(norm, p->roi, fin, sc, t12_frame, H, W, p->x0,
                                                                                   p->x1, p->y0, P.final_t_min, P.final_t_max);
    CKL();
  }
  vistaf_stage_mark("T.stats_median", st);
  FinalVal fv{fin, p->roi, plane, (long long)p->y0 * W};
  const long long nsel = (long long)(p->y1 - p->y0) * W;
  if (sel_reset(sel, nb, st)) return 1;
  if (select_quantile(fv, sel, nsel, nb, -1.0f, true, st, nullptr, 0, TEMP_SEL_CAP, true)) return 1;
  VL(), k_temp_records<<<cdiv(nb, 128), 128, 0, st>>>(sc, sel, O.records + f0, nb, W / 2, H / 2);
  CKL();
  vistaf_stage_mark("", st);
  return 0;
}

extern "C" int vistaf_temp_batch(VistafTempPlan* p, const uint8_t* frames, int batch, const VistafTempOutputs* out,
                                 void* workspace, size_t ws_bytes, void* stream) {
  if (!p || !frames || !out || !out->fused || !out->final_map || !out->records || !workspace)
    return vistaf_set_err("vistaf_temp_batch: null argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  TempWs L;
  int chunk = batch;                                   // largest chunk the workspace holds, searched downward from the batch
  while (chunk >= 1) { temp_layout(p, chunk, L); if (L.total <= ws_bytes) break; chunk = chunk > 8 ? chunk - (chunk + 15) / 16 : chunk - 1; }
  if (chunk < 1) return vistaf_set_err("vistaf_temp_batch: workspace too small for one frame", cudaErrorInvalidValue, __FILE__, __LINE__);
  const VistafTempOutputs O = *out;
  unsigned long long oh = 1469598103934665603ull;
  for (size_t i = 0; i < sizeof(O); i++) { oh ^= ((const unsigned char*)&O)[i]; oh *= 1099511628211ull; }
  const VistafGraphKey key = vistaf_graph_key(10, p, frames, batch, oh, O.fused, O.final_map, O.records, workspace, ws_bytes);
  return vistaf_graph_run(key, (cudaStream_t)stream, [&](cudaStream_t st) {
    for (int f0 = 0; f0 < batch; f0 += chunk) {
      const int nb = batch - f0 < chunk ? batch - f0 : chunk;
      TempWs Lc = L;
      if (nb != chunk) temp_layout(p, nb, Lc);
      if (temp_chunk(p, frames, nb, O, (size_t)f0, (char*)workspace, Lc, st)) return 1;
    }
    return 0;
  });
}
