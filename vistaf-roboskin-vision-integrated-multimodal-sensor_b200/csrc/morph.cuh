// Binary morphology on 0/1 uint8 masks with OpenCV semantics (cv2.dilate / erode / morphologyEx with
// the default constant border: pixels outside the image never set a dilation and never block an erosion).
// The structuring element is described per row by a half-width (ellipse: getStructuringElement
// MORPH_ELLIPSE formula; rectangle: constant), which covers every element the reference uses.
#pragma once
#include "common.cuh"

#define MORPH_MAXROWS 33
struct MorphSE {
  int rows;                   // odd
  int half[MORPH_MAXROWS];    // half-width of row i (-1: empty row)
};

// cv::getStructuringElement(MORPH_ELLIPSE, (k,k)):  dx = cvRound(c*sqrt((r^2-dy^2)/r^2)), r=c=k/2
static inline MorphSE se_ellipse(int k) {
  MorphSE s;
  s.rows = k;
  int r = k / 2, c = k / 2;
  double inv_r2 = r ? 1.0 / ((double)r * r) : 0.0;
  for (int i = 0; i < k; i++) {
    int dy = i - r;
    int dx = (int)lrint(c * sqrt((double)(r * r - dy * dy) * inv_r2));
    s.half[i] = dx;           // columns [c-dx, c+dx]
  }
  for (int i = k; i < MORPH_MAXROWS; i++) s.half[i] = -1;
  return s;
}
static inline MorphSE se_rect(int kx, int ky) {
  MorphSE s;
  s.rows = ky;
  for (int i = 0; i < MORPH_MAXROWS; i++) s.half[i] = i < ky ? kx / 2 : -1;
  return s;
}

// Bit-packed tile kernel.  A block of 128 threads produces a 32-column x 128-row tile.  Phase 1 packs every tile
// row with its halo (64 columns: x0-16 .. x0+47, half-widths are <= 16) into one 64-bit word through warp
// ballots; phase 2 gives one output row to each thread: for every element row the horizontal span
// [-hw, hw] is folded into a word by log-step shift-OR (shift-AND for erosion), rows are combined by OR / AND;
// phase 3 unpacks the 32 result bits of a row with one coalesced byte store per warp.
//   Load: int operator()(int b, int y, int x) -> 0/1 (only called in-bounds).
//   Post: int operator()(int b, int y, int x, int v) -> value stored (e.g. & roi).
#define MORPH_TW 32
#define MORPH_TH 128
template <bool DILATE, class Load, class Post>
__global__ void __launch_bounds__(128) k_morph(Load ld, Post post, uint8_t* __restrict__ dst, int h, int w, MorphSE se,
                                               int4 win, int skip_out) {   // win = (x0, x1, y0, y1): the input is known to be 0 outside it
  // skip_out: tiles whose whole neighbourhood lies outside the window are not written (their zeros are either never read:
  // the loads below substitute 0 outside the window, or already in place: an output plane with a valid background)
  __shared__ unsigned long long Wd[MORPH_TH + MORPH_MAXROWS - 1];
  __shared__ unsigned int Rw[MORPH_TH];
  const int ry = se.rows / 2, nrow = MORPH_TH + se.rows - 1;
  const int b = blockIdx.z, x0 = blockIdx.x * MORPH_TW, y0 = blockIdx.y * MORPH_TH;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (x0 - 16 >= win.y || x0 + MORPH_TW + 16 <= win.x || y0 - ry >= win.w || y0 + MORPH_TH + ry <= win.z) {
    if (skip_out) return;
    for (int ty = wid; ty < MORPH_TH; ty += 4) {     // all-zero neighbourhood: dilation and erosion both give 0
      const int y = y0 + ty, x = x0 + lane;
      if (y < h && x < w) dst[(size_t)b * h * w + (size_t)y * w + x] = (uint8_t)post(b, y, x, 0);
    }
    return;
  }
  for (int ty = wid; ty < nrow; ty += 4) {
    const int y = y0 - ry + ty;
    const bool yin = y >= 0 && y < h;
    unsigned int word[2];
#pragma unroll
    for (int hf = 0; hf < 2; hf++) {
      const int x = x0 - 16 + hf * 32 + lane;
      int v = DILATE ? 0 : 1;                       // border value
      if (yin && x >= 0 && x < w) v = (x >= win.x && x < win.y && y >= win.z && y < win.w) ? ld(b, y, x) : 0;
      word[hf] = __ballot_sync(0xffffffffu, v != 0);
    }
    if (lane == 0) Wd[ty] = (unsigned long long)word[0] | ((unsigned long long)word[1] << 32);
  }
  __syncthreads();
  {
    const int ly = threadIdx.x;
    unsigned int acc = DILATE ? 0u : 0xffffffffu;
    for (int i = 0; i < se.rows; i++) {
      const int hw = se.half[i];
      if (hw < 0) continue;
      unsigned long long a = Wd[ly + i] >> (16 - hw);   // bit c <-> column x0 + c - hw
      const int span = 2 * hw + 1;
      for (int L = 1; L < span;) {
        const int sft = min(L, span - L);
        a = DILATE ? (a | (a >> sft)) : (a & (a >> sft));
        L += sft;
      }
      acc = DILATE ? (acc | (unsigned int)a) : (acc & (unsigned int)a);
    }
    Rw[ly] = acc;
  }
  __syncthreads();
  for (int ty = wid; ty < MORPH_TH; ty += 4) {
    const int y = y0 + ty, x = x0 + lane;
    if (y < h && x < w) dst[(size_t)b * h * w + (size_t)y * w + x] = (uint8_t)post(b, y, x, (int)((Rw[ty] >> lane) & 1u));
  }
}

template <bool DILATE, class Load, class Post>
static int morph(Load ld, Post post, uint8_t* dst, int h, int w, int batch, const MorphSE& se, cudaStream_t st,
                 const int4* window = nullptr, bool skip_outside = false) {
  for (int i = 0; i < se.rows; i++)
    if (se.half[i] > 16) return vistaf_set_err("morph: structuring element wider than 33", cudaErrorInvalidValue, __FILE__, __LINE__);
  if (se.rows > MORPH_MAXROWS) return vistaf_set_err("morph: structuring element taller than 33", cudaErrorInvalidValue, __FILE__, __LINE__);
  dim3 grid(cdiv(w, MORPH_TW), cdiv(h, MORPH_TH), batch);
  const int4 win = window ? *window : make_int4(0, w, 0, h);
  VL(), k_morph<DILATE, Load, Post><<<grid, 128, 0, st>>>(ld, post, dst, h, w, se, win, window && skip_outside ? 1 : 0);
  CKL();
  return 0;
}

struct MaskLoad {               // plain 0/1 plane
  const uint8_t* p; long long plane; int w;
  __device__ int operator()(int b, int y, int x) const { return p[(long long)b * plane + (long long)y * w + x] != 0; }
};
struct PostId { __device__ int operator()(int, int, int, int v) const { return v; } };
struct PostAnd {                // & another mask (shared across the batch when plane == 0)
  const uint8_t* m; long long plane; int w;
  __device__ int operator()(int b, int y, int x, int v) const { return v & (m[(long long)b * plane + (long long)y * w + x] != 0); }
};
