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

// Tile kernel: block (32 x 8) threads computes a 32 x 32 tile (4 rows per thread) from a shared
// tile with halo.  Load: int operator()(int b, int y, int x) -> 0/1 (only called in-bounds).
// Post: int operator()(int b, int y, int x, int v) -> value stored (e.g. & roi).
#define MORPH_TW 32
#define MORPH_TH 32
template <bool DILATE, class Load, class Post>
__global__ void k_morph(Load ld, Post post, uint8_t* __restrict__ dst, int h, int w, MorphSE se) {
  extern __shared__ uint8_t tile[];
  const int ry = se.rows / 2;
  int rx = 0;
  for (int i = 0; i < se.rows; i++) rx = max(rx, se.half[i]);
  const int tw = MORPH_TW + 2 * rx, th = MORPH_TH + 2 * ry;
  const int b = blockIdx.z, x0 = blockIdx.x * MORPH_TW, y0 = blockIdx.y * MORPH_TH;
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, nthr = blockDim.x * blockDim.y;
  for (int e = tid; e < tw * th; e += nthr) {
    int ty = e / tw, tx = e - ty * tw;
    int y = y0 - ry + ty, x = x0 - rx + tx;
    int v = DILATE ? 0 : 1;                       // border value
    if (y >= 0 && y < h && x >= 0 && x < w) v = ld(b, y, x);
    tile[e] = (uint8_t)v;
  }
  __syncthreads();
  for (int j = 0; j < MORPH_TH / 8; j++) {
    int ly = threadIdx.y + 8 * j, lx = threadIdx.x;
    int y = y0 + ly, x = x0 + lx;
    if (y >= h || x >= w) continue;
    int v = DILATE ? 0 : 1;
    for (int i = 0; i < se.rows; i++) {
      int hw = se.half[i];
      if (hw < 0) continue;
      const uint8_t* row = tile + (ly + i) * tw + lx + rx;
      for (int d = -hw; d <= hw; d++) {
        if (DILATE) v |= row[d]; else v &= row[d];
      }
    }
    dst[(size_t)b * h * w + (size_t)y * w + x] = (uint8_t)post(b, y, x, v);
  }
}

template <bool DILATE, class Load, class Post>
static int morph(Load ld, Post post, uint8_t* dst, int h, int w, int batch, const MorphSE& se, cudaStream_t st) {
  int rx = 0;
  for (int i = 0; i < se.rows; i++) rx = rx > se.half[i] ? rx : se.half[i];
  size_t sm = (size_t)(MORPH_TW + 2 * rx) * (MORPH_TH + se.rows - 1);
  dim3 grid(cdiv(w, MORPH_TW), cdiv(h, MORPH_TH), batch), blk(32, 8);
  VL(), k_morph<DILATE, Load, Post><<<grid, blk, sm, st>>>(ld, post, dst, h, w, se);
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
