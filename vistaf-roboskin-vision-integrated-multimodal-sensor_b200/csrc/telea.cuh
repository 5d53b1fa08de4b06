// Telea fast-marching inpainting (cv2.inpaint(..., INPAINT_TELEA); reached from shape_ftp.py:652-666 on
// float32 crops with radius 3 and from temperature_sensor.py:546-577).  OpenCV's processing order is a
// global priority queue keyed by (T, insertion order); that order only matters between pixels that can
// read each other's state.  The hole mask dilated by the (2r+1)-square (hole + narrow band + outer ring)
// falls apart into 8-connected CLUSTERS that never read each other, so every cluster is marched by its
// own warp with the same relative order as the global queue:
//   1. k_telea_prep     flags (hole / band / ring) and T on the 1-pixel padded frame
//   2. uf_label         clusters = connected components of the dilated mask   (labels.cuh)
//   3. k_telea_slots/_bbox   compact cluster list, bounding boxes, cell counts
//   4. k_telea_march    per-cluster warp: outward distance march over the ring, then the inward march
//                       that fills pixels; the (2r+1)^2 neighbourhood of a pixel is evaluated one tap per
//                       lane and summed in raster order, so float results do not depend on the lane count.
#pragma once
#include "common.cuh"
#include "labels.cuh"
#include <stdlib.h>

#define TL_MASK 1
#define TL_BAND 2
#define TL_RING 4
#define TL_OUT 8          // ring cell reached by the outward march
#define TL_IN 16          // hole cell filled by the inward march
#define TL_MAXCLUSTERS 65536
#define TL_WARPS 4
#define TL_BLOCKS_PER_SM 2  // resident march blocks per SM (persistent grid): leaves the SM to co-running kernels
#define TL_NCLASS 24      // size classes (log2 of the cell count) for largest-first scheduling

struct TeleaFrame {       // per-frame bookkeeping (device)
  unsigned int nclusters, next, overflow, pad_;
};
struct TeleaCluster {     // per cluster slot
  unsigned int root, count, nband, bfill;   // cells, band cells, band-list fill cursor
  unsigned int off, boff;                   // offsets of the heap / log region and of the band list
  unsigned int pad_[2];
};

struct TeleaWs {          // device pointers, dense [batch][...] planes
  uint8_t* flag;          // [Hp*Wp]
  float* T;               // [Hp*Wp]
  ull* uf;                // [Hp*Wp]
  unsigned int* slot;     // [Hp*Wp]  slot index stored at root cells
  TeleaFrame* fr;         // [1]
  TeleaCluster* cl;       // [TL_MAXCLUSTERS]
  unsigned int* order;    // [TL_MAXCLUSTERS] cluster slots, largest size class first
  ull* keys;              // [cap]  heap keys (T bits << 32) | seq
  int* pos;               // [cap]  heap payload: cell
  int* log;               // [cap]  cells popped by the outward march (to negate their T)
  int* band;              // [cap]  band cells per cluster (unordered; keys restore raster order)
  size_t sp, cap;         // padded cells per frame, capacity per frame
};

static inline size_t telea_ws_bytes(int h, int w, int batch) {
  const size_t sp = (size_t)(h + 2) * (w + 2), cap = sp / 2, B = (size_t)batch;
  return align256(B * sp * 8) + align256(B * sp * 4) * 2 + align256(B * sp) + align256(B * sizeof(TeleaFrame)) +
         align256(B * sizeof(TeleaCluster) * TL_MAXCLUSTERS) + align256(B * 4 * TL_MAXCLUSTERS) + align256(B * cap * 8) +
         align256(B * cap * 4) * 3;
}
static inline void telea_ws_carve(char* base, int h, int w, int batch, TeleaWs& W) {
  const size_t sp = (size_t)(h + 2) * (w + 2), cap = sp / 2, B = (size_t)batch;
  size_t o = 0;
  auto take = [&](size_t bytes) { char* r = base + o; o += align256(bytes); return r; };
  W.uf = (ull*)take(B * sp * 8);
  W.T = (float*)take(B * sp * 4);
  W.slot = (unsigned int*)take(B * sp * 4);
  W.flag = (uint8_t*)take(B * sp);
  W.fr = (TeleaFrame*)take(B * sizeof(TeleaFrame));
  W.cl = (TeleaCluster*)take(B * sizeof(TeleaCluster) * TL_MAXCLUSTERS);
  W.order = (unsigned int*)take(B * 4 * TL_MAXCLUSTERS);
  W.keys = (ull*)take(B * cap * 8);
  W.pos = (int*)take(B * cap * 4);
  W.log = (int*)take(B * cap * 4);
  W.band = (int*)take(B * cap * 4);
  W.sp = sp; W.cap = cap;
}

// ------------------------------------------------------------------ 1. flags on the padded frame
// Mask: bool operator()(int b, int y, int x) on image coordinates.  block (32, 8), tile 32 x 32.
template <class Mask>
__global__ void k_telea_prep(Mask mask, uint8_t* __restrict__ flag, float* __restrict__ T, int h, int w, int range,
                             const int* __restrict__ active) {
  if (active && active[blockIdx.z] < 0) return;
  extern __shared__ uint8_t tl_prep_sm[];         // m[(32+2R)^2], rowhit[(32+2R) x 32]
  const int R = range, TW = 32 + 2 * R;
  uint8_t* m = tl_prep_sm;
  uint8_t* rh = tl_prep_sm + TW * TW;
  const int Hp = h + 2, Wp = w + 2, b = blockIdx.z;
  const int X0 = blockIdx.x * 32, Y0 = blockIdx.y * 32;      // padded coordinates of the tile
  const int tid = threadIdx.y * 32 + threadIdx.x;
  for (int e = tid; e < TW * TW; e += 256) {
    int ty = e / TW, tx = e - ty * TW;
    int Y = Y0 - R + ty, X = X0 - R + tx;                    // padded coords; image = (Y-1, X-1)
    int v = 0;
    if (Y >= 1 && Y <= h && X >= 1 && X <= w) v = mask(b, Y - 1, X - 1) ? 1 : 0;
    m[e] = (uint8_t)v;
  }
  __syncthreads();
  for (int e = tid; e < TW * 32; e += 256) {                 // horizontal OR over [x-R, x+R]
    int ty = e >> 5, tx = e & 31;
    int v = 0;
    for (int d = 0; d <= 2 * R; d++) v |= m[ty * TW + tx + d];
    rh[e] = (uint8_t)v;
  }
  __syncthreads();
  for (int j = 0; j < 4; j++) {
    const int ly = threadIdx.y + 8 * j, lx = threadIdx.x;
    const int Y = Y0 + ly, X = X0 + lx;
    if (Y >= Hp || X >= Wp) continue;
    int near = 0;
    for (int d = 0; d <= 2 * R; d++) near |= rh[(ly + d) * 32 + lx];
    const uint8_t* c = m + (ly + R) * TW + lx + R;
    int f = 0;
    float t = 1.0e6f;
    if (c[0]) f = TL_MASK;
    else if ((c[-1] | c[1] | c[-TW] | c[TW]) && Y >= 1 && Y <= h && X >= 1 && X <= w) { f = TL_BAND; t = 0.f; }
    else if (near) f = TL_RING;
    flag[((size_t)b * Hp + Y) * Wp + X] = (uint8_t)f;
    T[((size_t)b * Hp + Y) * Wp + X] = t;
  }
}

struct TeleaFg {          // cluster foreground on the padded frame
  const uint8_t* flag; size_t stride; int Wp;
  __device__ bool operator()(int b, int y, int x) const { return flag[(size_t)b * stride + (size_t)y * Wp + x] != 0; }
};


// ------------------------------------------------------------------ 3. cluster table
// slots: one slot per union-find root.
static __global__ void k_telea_slots(const ull* __restrict__ uf, unsigned int* __restrict__ slot, TeleaFrame* fr,
                              TeleaCluster* cl, int Hp, int Wp, const int* __restrict__ active) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (active && active[b] < 0) return;
  const size_t sp = (size_t)Hp * Wp;
  if (i >= Hp * Wp) return;
  if (uf_parent(uf[(size_t)b * sp + i]) != (unsigned int)i) return;
  TeleaFrame* f = fr + b;
  unsigned int s = atomicAdd(&f->nclusters, 1u);
  slot[(size_t)b * sp + i] = s;
  if (s < TL_MAXCLUSTERS) {
    TeleaCluster* c = cl + (size_t)b * TL_MAXCLUSTERS + s;
    c->root = (unsigned int)i; c->count = 0; c->nband = 0; c->bfill = 0; c->off = 0; c->boff = 0;
  }
}
// counts: cells and band cells per cluster (warp-aggregated by root).
static __global__ void k_telea_count(const ull* __restrict__ uf, const unsigned int* __restrict__ slot,
                              const uint8_t* __restrict__ flag, TeleaCluster* cl, int Hp, int Wp,
                              const int* __restrict__ active) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (active && active[b] < 0) return;
  const size_t sp = (size_t)Hp * Wp;
  unsigned int root = UF_BG;
  bool band = false;
  if (i < Hp * Wp) {
    root = uf_parent(uf[(size_t)b * sp + i]);
    band = root != UF_BG && (flag[(size_t)b * sp + i] & TL_BAND);
  }
  const unsigned int grp = __match_any_sync(0xffffffffu, root);
  const unsigned int bandbits = __ballot_sync(0xffffffffu, band);
  if (root == UF_BG || (int)(__ffs(grp) - 1) != (int)(threadIdx.x & 31)) return;
  const unsigned int s = slot[(size_t)b * sp + root];
  if (s >= TL_MAXCLUSTERS) return;
  TeleaCluster* c = cl + (size_t)b * TL_MAXCLUSTERS + s;
  atomicAdd(&c->count, (unsigned int)__popc(grp));
  const unsigned int nb = (unsigned int)__popc(grp & bandbits);
  if (nb) atomicAdd(&c->nband, nb);
}
// schedule: order[] = cluster slots sorted by size class (largest first); off / boff = exclusive prefix sums of
// count / nband in that order.  One block of 256 threads per frame.
static __global__ void __launch_bounds__(256) k_telea_schedule(TeleaFrame* fr, TeleaCluster* cl, unsigned int* order,
                                                               unsigned int cap, unsigned int* status, size_t status_stride) {
  __shared__ unsigned int ncls[TL_NCLASS], base[TL_NCLASS], fill[TL_NCLASS];
  __shared__ unsigned int wsum[8][2], carry[2];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  TeleaFrame* f = fr + b;
  TeleaCluster* c = cl + (size_t)b * TL_MAXCLUSTERS;
  unsigned int* ord = order + (size_t)b * TL_MAXCLUSTERS;
  const unsigned int n = min(f->nclusters, (unsigned int)TL_MAXCLUSTERS);
  if (tid < TL_NCLASS) { ncls[tid] = 0; fill[tid] = 0; }
  if (tid < 2) carry[tid] = 0;
  __syncthreads();
  for (unsigned int s = tid; s < n; s += 256) atomicAdd(&ncls[min(TL_NCLASS - 1, 31 - __clz(max(1u, c[s].count)))], 1u);
  __syncthreads();
  if (tid == 0) {
    unsigned int a = 0;
    for (int k = TL_NCLASS - 1; k >= 0; k--) { base[k] = a; a += ncls[k]; }
  }
  __syncthreads();
  for (unsigned int s = tid; s < n; s += 256) {
    const int k = min(TL_NCLASS - 1, 31 - __clz(max(1u, c[s].count)));
    ord[base[k] + atomicAdd(&fill[k], 1u)] = s;
  }
  __syncthreads();
  // exclusive scans over the ordered list, 256 clusters per round
  for (unsigned int i0 = 0; i0 < n; i0 += 256) {
    const unsigned int i = i0 + tid;
    const unsigned int s = i < n ? ord[i] : 0;
    unsigned int v0 = i < n ? c[s].count : 0, v1 = i < n ? c[s].nband : 0;
    unsigned int s0 = v0, s1 = v1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned int t0 = __shfl_up_sync(0xffffffffu, s0, o), t1 = __shfl_up_sync(0xffffffffu, s1, o);
      if (lane >= o) { s0 += t0; s1 += t1; }
    }
    if (lane == 31) { wsum[wid][0] = s0; wsum[wid][1] = s1; }
    __syncthreads();
    unsigned int p0 = carry[0], p1 = carry[1];
    for (int k = 0; k < wid; k++) { p0 += wsum[k][0]; p1 += wsum[k][1]; }
    if (i < n) {
      c[s].off = p0 + s0 - v0;
      c[s].boff = p1 + s1 - v1;
      if (p0 + s0 > cap || p1 + s1 > cap) {           // does not fit the workspace: skipped and reported
        c[s].count = 0xffffffffu;
        atomicAdd(&f->overflow, 1u);
      }
    }
    __syncthreads();
    if (tid == 255) { carry[0] = p0 + s0; carry[1] = p1 + s1; }
    __syncthreads();
  }
  if (tid == 0 && (f->nclusters > TL_MAXCLUSTERS || f->overflow))
    atomicOr((unsigned int*)((char*)status + (size_t)b * status_stride), VISTAF_ST_TELEA_OVERFLOW);
}
// band lists (unordered within a cluster; the heap keys carry the raster order).
static __global__ void k_telea_bandlist(const ull* __restrict__ uf, const unsigned int* __restrict__ slot,
                                 const uint8_t* __restrict__ flag, TeleaCluster* cl, int* __restrict__ band, size_t cap,
                                 int Hp, int Wp, const int* __restrict__ active) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (active && active[b] < 0) return;
  const size_t sp = (size_t)Hp * Wp;
  if (i >= Hp * Wp) return;
  if (!(flag[(size_t)b * sp + i] & TL_BAND)) return;
  const unsigned int root = uf_parent(uf[(size_t)b * sp + i]);
  if (root == UF_BG) return;
  const unsigned int s = slot[(size_t)b * sp + root];
  if (s >= TL_MAXCLUSTERS) return;
  TeleaCluster* c = cl + (size_t)b * TL_MAXCLUSTERS + s;
  if (c->count == 0xffffffffu) return;
  band[(size_t)b * cap + c->boff + atomicAdd(&c->bfill, 1u)] = i;
}

// ------------------------------------------------------------------ 4. the march
// One warp per cluster; the march is a chain of dependent steps, so everything on that chain is kept at
// shared-memory / register latency:
//  * the priority queue is a 32-ary min-heap on (T, seq) keys whose first TL_HCAP entries (levels 0-2) live in
//    shared memory; a sift-down step examines the 32 children of a node with one access and a warp reduction;
//  * per popped cell ONE gather brings T and flags of the 13 cells within Manhattan distance 2 (lane k loads
//    cell k); the four neighbours' distance solves then run in registers on lanes 0..3 (they are mutually
//    independent: no neighbour of the popped cell is a 4-neighbour of another);
//  * the (2r+1)^2 neighbourhood of a filled pixel is evaluated one tap per lane with all loads issued up front,
//    and summed in raster order by lanes 0..3 (one accumulator each) from shared memory.
// Band cells enter with seq = their raster cell index (OpenCV pushes them in raster order, all with T = 0);
// later pushes take seq from a counter above any cell index.
#define TL_HCAP 512
#define TL_MAXTAPS 289        // (2 * 8 + 1)^2

struct TeleaHeap {
  ull* sk; int* sp;        // shared: entries [0, TL_HCAP)
  ull* gk; int* gp;        // global: entries [TL_HCAP, ...) (indexed by the entry number)
  int n; unsigned int seq;
};
__device__ __forceinline__ ull th_key(const TeleaHeap& H, int i) { return i < TL_HCAP ? H.sk[i] : H.gk[i]; }
__device__ __forceinline__ int th_pos(const TeleaHeap& H, int i) { return i < TL_HCAP ? H.sp[i] : H.gp[i]; }
__device__ __forceinline__ void th_set(TeleaHeap& H, int i, ull k, int p) {
  if (i < TL_HCAP) { H.sk[i] = k; H.sp[i] = p; } else { H.gk[i] = k; H.gp[i] = p; }
}
// children of node i are 32 i + 1 .. 32 i + 32.  All lanes call these together.
__device__ __forceinline__ void th_sift_down(TeleaHeap& H, int i, ull key, int cell, int lane) {
  for (;;) {
    const int first = 32 * i + 1;
    if (first >= H.n) break;
    const int idx = first + lane;
    ull k = ~0ull;
    int pp = 0;
    if (idx < H.n) { k = th_key(H, idx); pp = th_pos(H, idx); }
    const unsigned int hi = (unsigned int)(k >> 32), lo = (unsigned int)k;
    const unsigned int mhi = __reduce_min_sync(0xffffffffu, hi);
    const unsigned int mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
    const ull kmin = ((ull)mhi << 32) | (ull)mlo;
    if (kmin >= key) break;
    const int src = __ffs(__ballot_sync(0xffffffffu, k == kmin)) - 1;
    const int cpos = __shfl_sync(0xffffffffu, pp, src);
    if (lane == 0) th_set(H, i, kmin, cpos);
    i = first + src;
  }
  if (lane == 0) th_set(H, i, key, cell);
  __syncwarp();
}
__device__ __forceinline__ void th_push(TeleaHeap& H, int cell, float t, int lane) {
  const ull key = ((ull)__float_as_uint(t) << 32) | (ull)H.seq;
  H.seq++;
  int i = H.n++;
  while (i > 0) {
    const int p = (i - 1) >> 5;
    const ull pk = th_key(H, p);
    if (pk <= key) break;
    const int pp = th_pos(H, p);
    if (lane == 0) th_set(H, i, pk, pp);
    i = p;
  }
  if (lane == 0) th_set(H, i, key, cell);
  __syncwarp();
}
__device__ __forceinline__ int th_pop(TeleaHeap& H, int lane) {      // removes the minimum (T, seq); returns its cell
  const int cell = th_pos(H, 0);
  H.n--;
  if (H.n > 0) {
    const ull key = th_key(H, H.n);
    const int last = th_pos(H, H.n);
    __syncwarp();
    th_sift_down(H, 0, key, last, lane);
  }
  return cell;
}
// initial queue: the band cells of this cluster, T = 0, ordered by raster index through the key
__device__ __forceinline__ void th_fill_band(TeleaHeap& H, const int* band, int nband, int ncells, int lane) {
  for (int e = lane; e < nband; e += 32) {
    const int cell = band[e];
    th_set(H, e, (ull)(unsigned int)cell, cell);
  }
  H.n = nband;
  H.seq = (unsigned int)ncells;
  __syncwarp();
  for (int node = (nband - 2) >> 5; node >= 0; node--) {            // heapify: sift every internal node down
    const ull key = th_key(H, node);
    const int cell = th_pos(H, node);
    __syncwarp();
    th_sift_down(H, node, key, cell, lane);
  }
}

// Eikonal update from two 4-neighbours (T value, "still unknown" flag each); OpenCV's FastMarching_solve.
__device__ __forceinline__ float tl_solve2(float t1, bool in1, float t2, bool in2) {
  const double a11 = (double)t1, a22 = (double)t2, m12 = a11 < a22 ? a11 : a22;
  double sol;
  if (!in1) {
    if (!in2) {
      if (fabs(a11 - a22) >= 1.0) sol = 1.0 + m12;
      else sol = (a11 + a22 + sqrt(2.0 - (a11 - a22) * (a11 - a22))) * 0.5;
    } else sol = 1.0 + a11;
  } else if (!in2) sol = 1.0 + a22;
  else sol = 1.0 + m12;
  return (float)sol;
}
template <bool OUTWARD>
__device__ __forceinline__ bool tl_in(int v) {
  return OUTWARD ? ((v & TL_RING) && !(v & TL_OUT)) : ((v & TL_MASK) && !(v & TL_IN));
}
// The 13 cells within Manhattan distance 2 of the popped cell: index -> (dy, dx).  1..4 are the neighbours in
// OpenCV's visiting order (up, left, down, right).
__device__ __forceinline__ void tl_diamond(int k, int& dy, int& dx) {
  dy = (int)((0x364a88268aull >> (3 * k)) & 7ull) - 2;
  dx = (int)((0x32cc413452ull >> (3 * k)) & 7ull) - 2;
}
// diamond indices of the (up, left, down, right) neighbours of neighbour q
__device__ __forceinline__ int tl_nbr(int q, int dir) { return (int)((0x8c0ac7b00b69a095ull >> (4 * (4 * q + dir))) & 15ull); }

// gather of the diamond around `cell` (lane k < 13 loads cell k; out-of-frame cells read as "known, far away")
__device__ __forceinline__ void tl_gather(const uint8_t* f, const float* T, int ii, int jj, int Hp, int Wp, int lane,
                                          float& t, int& v) {
  t = 1.0e6f; v = 0;
  if (lane < 13) {
    int dy, dx;
    tl_diamond(lane, dy, dx);
    const int y = ii + dy, x = jj + dx;
    if (y >= 0 && x >= 0 && y < Hp && x < Wp) { t = T[y * Wp + x]; v = f[y * Wp + x]; }
  }
}

static __global__ void __launch_bounds__(32 * TL_WARPS) k_telea_march(TeleaWs Wsp, float* __restrict__ img,
                                                                size_t img_stride, int h, int w, int range, int batch,
                                                                int quant) {
  extern __shared__ unsigned char tl_smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  ull* sk = (ull*)tl_smem + (size_t)wid * TL_HCAP;
  int* sp = (int*)(tl_smem + (size_t)TL_WARPS * TL_HCAP * 8) + (size_t)wid * TL_HCAP;
  float (*tp)[4] = (float (*)[4])(tl_smem + (size_t)TL_WARPS * TL_HCAP * 12) + (size_t)wid * 32;
  // per-tap constants of the (2r+1)^2 neighbourhood: offsets (0x7fff = outside the radius-r disc) and 1 / |r|^3
  short2* tapd = (short2*)(tl_smem + (size_t)TL_WARPS * (TL_HCAP * 12 + 32 * 16));
  float* tapw = (float*)(tapd + TL_MAXTAPS);
  const int Hp = h + 2, Wp = w + 2;
  const int r2 = range * range, side = 2 * range + 1, ntaps = side * side;
  for (int tt = threadIdx.x; tt < ntaps; tt += blockDim.x) {
    const int dk = tt / side - range, dl = tt - (tt / side) * side - range;
    const bool in_disc = dk * dk + dl * dl <= r2 && (dk | dl) != 0;        // the centre is the pixel being filled
    const float ry = (float)(-dk), rx = (float)(-dl);
    const float len2 = __fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry));
    tapd[tt] = in_disc ? make_short2((short)dk, (short)dl) : make_short2(0x7fff, 0x7fff);
    tapw[tt] = in_disc ? (float)(1.0 / (double)__fmul_rn(len2, __fsqrt_rn(len2))) : 0.f;
  }
  __syncthreads();
  const int q = lane & 3;                                   // lanes 0..3 own neighbour q; other lanes mirror them
  int qdy, qdx;
  tl_diamond(q + 1, qdy, qdx);
  // Persistent grid: every warp serves one frame's largest-first cluster queue until it runs dry, then moves on to
  // the next frame.  Warps start spread over the frames, so the few giant clusters of EVERY frame (the critical
  // path: a serial march of 10^4 cells) begin at once.
  int b = (int)((blockIdx.x * TL_WARPS + wid) % (unsigned int)batch);
  int dry = 0;
  for (;;) {
    TeleaFrame* fr = Wsp.fr + b;
    const unsigned int ncl = min(fr->nclusters, (unsigned int)TL_MAXCLUSTERS);
    unsigned int s = 0xffffffffu;
    if (lane == 0 && *(volatile unsigned int*)&fr->next < ncl) s = atomicAdd(&fr->next, 1u);
    s = __shfl_sync(0xffffffffu, s, 0);
    if (s >= ncl) {
      if (++dry >= batch) break;
      b = b + 1 == batch ? 0 : b + 1;
      continue;
    }
    dry = 0;
    const TeleaCluster* clt = Wsp.cl + (size_t)b * TL_MAXCLUSTERS;
    const unsigned int* ord = Wsp.order + (size_t)b * TL_MAXCLUSTERS;
    uint8_t* f = Wsp.flag + (size_t)b * Wsp.sp;
    float* T = Wsp.T + (size_t)b * Wsp.sp;
    float* im = img + (size_t)b * img_stride;
    const TeleaCluster cl = clt[ord[s]];
    if (cl.count == 0xffffffffu) continue;                  // did not fit the workspace (status bit already set)
    TeleaHeap H;
    H.sk = sk; H.sp = sp;
    H.gk = Wsp.keys + (size_t)b * Wsp.cap + cl.off;
    H.gp = Wsp.pos + (size_t)b * Wsp.cap + cl.off;
    int* log = Wsp.log + (size_t)b * Wsp.cap + cl.off;
    const int* band = Wsp.band + (size_t)b * Wsp.cap + cl.boff;
    int nlog = 0;
    // ---- outward: distances over the ring
    __syncwarp();
    th_fill_band(H, band, (int)cl.nband, Hp * Wp, lane);
    while (H.n > 0) {
      const int cell = th_pop(H, lane);
      if (lane == 0) log[nlog] = cell;
      nlog++;
      const int ii = cell / Wp, jj = cell - ii * Wp;
      float t; int v;
      tl_gather(f, T, ii, jj, Hp, Wp, lane, t, v);
      // lane q: neighbour q = (ii, jj) + d_q
      const int i = ii + qdy, j = jj + qdx;
      const int vq = __shfl_sync(0xffffffffu, v, q + 1);
      const bool ok = i > 0 && j > 0 && i < Hp - 1 && j < Wp - 1 && tl_in<true>(vq);
      const float tu = __shfl_sync(0xffffffffu, t, tl_nbr(q, 0)), tl_ = __shfl_sync(0xffffffffu, t, tl_nbr(q, 1));
      const float td = __shfl_sync(0xffffffffu, t, tl_nbr(q, 2)), tr = __shfl_sync(0xffffffffu, t, tl_nbr(q, 3));
      const bool iu = tl_in<true>(__shfl_sync(0xffffffffu, v, tl_nbr(q, 0))), il = tl_in<true>(__shfl_sync(0xffffffffu, v, tl_nbr(q, 1)));
      const bool id = tl_in<true>(__shfl_sync(0xffffffffu, v, tl_nbr(q, 2))), ir = tl_in<true>(__shfl_sync(0xffffffffu, v, tl_nbr(q, 3)));
      float d = 0.f;
      if (ok) d = fminf(fminf(tl_solve2(tu, iu, tl_, il), tl_solve2(td, id, tl_, il)),
                        fminf(tl_solve2(tu, iu, tr, ir), tl_solve2(td, id, tr, ir)));
      if (ok && lane < 4) { T[i * Wp + j] = d; f[i * Wp + j] = (uint8_t)(vq | TL_OUT); }
      const unsigned int okbits = __ballot_sync(0xffffffffu, ok) & 15u;
#pragma unroll
      for (int qq = 0; qq < 4; qq++) {
        if (!((okbits >> qq) & 1u)) continue;
        int dy, dx;
        tl_diamond(qq + 1, dy, dx);
        th_push(H, (ii + dy) * Wp + jj + dx, __shfl_sync(0xffffffffu, d, qq), lane);
      }
    }
    // negate the distances of every cell the outward march visited (band and reached ring)
    __syncwarp();
    for (int e = lane; e < nlog; e += 32) { const int cell = log[e]; T[cell] = -T[cell]; }
    __syncwarp();
    // ---- inward: fill the hole
    th_fill_band(H, band, (int)cl.nband, Hp * Wp, lane);
    while (H.n > 0) {
      const int cell = th_pop(H, lane);
      const int ii = cell / Wp, jj = cell - ii * Wp;
      float t; int v;
      tl_gather(f, T, ii, jj, Hp, Wp, lane, t, v);
      const int iq = ii + qdy, jq = jj + qdx;
      const int vq = __shfl_sync(0xffffffffu, v, q + 1);
      const bool ok = iq > 0 && jq > 0 && iq < Hp - 1 && jq < Wp - 1 && tl_in<false>(vq);
      const float tu = __shfl_sync(0xffffffffu, t, tl_nbr(q, 0)), tl_ = __shfl_sync(0xffffffffu, t, tl_nbr(q, 1));
      const float td = __shfl_sync(0xffffffffu, t, tl_nbr(q, 2)), tr = __shfl_sync(0xffffffffu, t, tl_nbr(q, 3));
      const bool iu = tl_in<false>(__shfl_sync(0xffffffffu, v, tl_nbr(q, 0))), il = tl_in<false>(__shfl_sync(0xffffffffu, v, tl_nbr(q, 1)));
      const bool id = tl_in<false>(__shfl_sync(0xffffffffu, v, tl_nbr(q, 2))), ir = tl_in<false>(__shfl_sync(0xffffffffu, v, tl_nbr(q, 3)));
      float dq = 0.f, gxq = 0.f, gyq = 0.f;
      if (ok) {
        dq = fminf(fminf(tl_solve2(tu, iu, tl_, il), tl_solve2(td, id, tl_, il)),
                   fminf(tl_solve2(tu, iu, tr, ir), tl_solve2(td, id, tr, ir)));
        if (!ir) gxq = !il ? __fmul_rn(__fsub_rn(tr, tl_), 0.5f) : __fsub_rn(tr, dq);
        else gxq = !il ? __fsub_rn(dq, tl_) : 0.f;
        if (!id) gyq = !iu ? __fmul_rn(__fsub_rn(td, tu), 0.5f) : __fsub_rn(td, dq);
        else gyq = !iu ? __fsub_rn(dq, tu) : 0.f;
      }
      const unsigned int okbits = __ballot_sync(0xffffffffu, ok) & 15u;
      for (int qq = 0; qq < 4; qq++) {
        if (!((okbits >> qq) & 1u)) continue;
        int dy, dx;
        tl_diamond(qq + 1, dy, dx);
        const int i = ii + dy, j = jj + dx, p = i * Wp + j;
        const float dist = __shfl_sync(0xffffffffu, dq, qq);
        const float gTx = __shfl_sync(0xffffffffu, gxq, qq), gTy = __shfl_sync(0xffffffffu, gyq, qq);
        const int vp = __shfl_sync(0xffffffffu, vq, qq);
        float Ia = 0.f, Jx = 0.f, Jy = 0.f, sw = 1.0e-20f;        // meaningful on lanes 0..3 (one accumulator each)
        for (int t0 = 0; t0 < ntaps; t0 += 32) {
          const int tap = t0 + lane;
          float wgt = 0.f, ta = 0.f, tx = 0.f, ty = 0.f;
          const short2 dd = tap < ntaps ? tapd[tap] : make_short2(0x7fff, 0x7fff);
          const int dk = dd.x, dl = dd.y;
          const int k = i + dk, l = j + dl;
          if (dk != 0x7fff && k > 0 && l > 0 && k < Hp - 1 && l < Wp - 1) {
            const int pk = k * Wp + l;
            const int km = k - 1 + (k == 1), kp = k - 1 - (k == Hp - 2);
            const int lm = l - 1 + (l == 1), lp = l - 1 - (l == Wp - 2);
            // every load of this tap is issued before any of them is used (one memory round trip)
            const int f0 = f[pk], fl = f[pk - 1], fr_ = f[pk + 1], fu = f[pk - Wp], fd = f[pk + Wp];
            const float Tk = T[pk];
#define TL_O(a, bb) im[(size_t)(a) * w + (bb)]
            const float o_c = TL_O(k - 1, l - 1);
            const float o_r1 = TL_O(km, lp + 1), o_l1 = TL_O(km, lm - 1), o_lm = TL_O(km, lm), o_lp = TL_O(km, lp);
            const float o_d1 = TL_O(kp + 1, lm), o_u1 = TL_O(km - 1, lm), o_kp = TL_O(kp, lm);
#undef TL_O
            if (!tl_in<false>(f0)) {
              const float ry = (float)(-dk), rx = (float)(-dl);
              const float dst = tapw[tap];
              const float lev = (float)(1.0 / (1.0 + fabs((double)__fsub_rn(Tk, dist))));
              float dir = __fadd_rn(__fmul_rn(rx, gTx), __fmul_rn(ry, gTy));
              if (fabs((double)dir) <= 0.01) dir = 0.000001f;
              wgt = fabsf(__fmul_rn(__fmul_rn(dst, lev), dir));
              float gIx, gIy;
              if (!tl_in<false>(fr_)) gIx = !tl_in<false>(fl) ? __fmul_rn(__fsub_rn(o_r1, o_l1), 2.0f) : __fsub_rn(o_r1, o_lm);
              else gIx = !tl_in<false>(fl) ? __fsub_rn(o_lp, o_l1) : 0.f;
              if (!tl_in<false>(fd)) gIy = !tl_in<false>(fu) ? __fmul_rn(__fsub_rn(o_d1, o_u1), 2.0f) : __fsub_rn(o_d1, o_lm);
              else gIy = !tl_in<false>(fu) ? __fsub_rn(o_kp, o_u1) : 0.f;
              ta = __fmul_rn(wgt, o_c);
              tx = __fmul_rn(wgt, __fmul_rn(gIx, rx));
              ty = __fmul_rn(wgt, __fmul_rn(gIy, ry));
            }
          }
          // raster-order accumulation: lane a < 4 sums term a of the 32 taps of this round sequentially (inactive taps
          // contribute +0, which leaves a float sum unchanged), so results do not depend on the lane count.
          __syncwarp();
          tp[lane][0] = ta; tp[lane][1] = -tx; tp[lane][2] = -ty; tp[lane][3] = wgt;
          __syncwarp();
          if (lane < 4) {
            float acc = lane == 0 ? Ia : lane == 1 ? Jx : lane == 2 ? Jy : sw;
#pragma unroll
            for (int tt = 0; tt < 32; tt++) acc = __fadd_rn(acc, tp[tt][lane]);
            if (lane == 0) Ia = acc; else if (lane == 1) Jx = acc; else if (lane == 2) Jy = acc; else sw = acc;
          }
        }
        Ia = __shfl_sync(0xffffffffu, Ia, 0); Jx = __shfl_sync(0xffffffffu, Jx, 1);
        Jy = __shfl_sync(0xffffffffu, Jy, 2); sw = __shfl_sync(0xffffffffu, sw, 3);
        const float nrm = __fadd_rn(__fsqrt_rn(__fadd_rn(__fmul_rn(Jx, Jx), __fmul_rn(Jy, Jy))), 1.0e-20f);
        float val = __fadd_rn(__fdiv_rn(Ia, sw), __fdiv_rn(__fadd_rn(Jx, Jy), nrm));
        // 8-bit images (cv2.inpaint on CV_8UC1): OpenCV stores saturate_cast<uchar>(value + 0.5f), i.e. round-half-even of
        // the biased value clamped to [0, 255]; later taps read the stored byte.
        if (quant) val = fminf(fmaxf(rintf(__fadd_rn(val, 0.5f)), 0.f), 255.f);
        if (lane == 0) {
          im[(size_t)(i - 1) * w + (j - 1)] = val;
          T[p] = dist;
          f[p] = (uint8_t)(vp | TL_IN);
        }
        __syncwarp();
        th_push(H, p, dist, lane);
      }
    }
  }
}

// Host driver.  mask functor on image coordinates; img is modified in place on the hole pixels.
template <class Mask>
static int telea_inpaint(Mask mask, float* img, size_t img_stride, unsigned int* status, size_t status_stride,
                         char* ws, int h, int w, int batch, int range, cudaStream_t st, int quant = 0,
                         const int* active = nullptr) {
  if (range < 1 || range > 8) return vistaf_set_err("telea: radius out of range", cudaErrorInvalidValue, __FILE__, __LINE__);
  TeleaWs W;
  telea_ws_carve(ws, h, w, batch, W);
  const int Hp = h + 2, Wp = w + 2, sp = Hp * Wp;
  const int TW = 32 + 2 * range;
  VL(), k_telea_prep<Mask><<<dim3(cdiv(Wp, 32), cdiv(Hp, 32), batch), dim3(32, 8), (size_t)TW * TW + (size_t)TW * 32, st>>>(
      mask, W.flag, W.T, h, w, range, active);
  CKL();
  TeleaFg fg{W.flag, (size_t)sp, Wp};
  if (uf_label(fg, EdgeZero(), W.uf, (unsigned int*)nullptr, (unsigned int*)nullptr, 0, Hp, Wp, batch, st, active)) return 1;
  CK(cudaMemsetAsync(W.fr, 0, sizeof(TeleaFrame) * (size_t)batch, st));
  const dim3 gcell(cdiv(sp, 256), batch);
  VL(), k_telea_slots<<<gcell, 256, 0, st>>>(W.uf, W.slot, W.fr, W.cl, Hp, Wp, active);
  VL(), k_telea_count<<<gcell, 256, 0, st>>>(W.uf, W.slot, W.flag, W.cl, Hp, Wp, active);
  VL(), k_telea_schedule<<<batch, 256, 0, st>>>(W.fr, W.cl, W.order, (unsigned int)W.cap, status, status_stride);
  VL(), k_telea_bandlist<<<gcell, 256, 0, st>>>(W.uf, W.slot, W.flag, W.cl, W.band, W.cap, Hp, Wp, active);
  const size_t march_smem = (size_t)TL_WARPS * (TL_HCAP * 12 + 32 * 16) + (size_t)TL_MAXTAPS * 8;
  static int n_sm = 0;
  if (!n_sm) {
    int dev = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaFuncSetAttribute((const void*)k_telea_march, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)march_smem));
  }
  VL(), k_telea_march<<<n_sm * TL_BLOCKS_PER_SM, 32 * TL_WARPS, march_smem, st>>>(W, img, img_stride, h, w, range, batch, quant);
  CKL();
  return 0;
}
