// Label kernels: 8-connected components with integer edge potentials (weighted union-find) and the
// 3x3 chamfer distance transform of cv2.distanceTransform(DIST_L2, 3).
//
// Weighted union-find.  Every foreground pixel x owns one 64-bit word  (delta << 32) | parent  where
// delta = k[x] - k[parent] for an integer potential k defined on the component up to a constant
// (k = 0 everywhere gives plain connected-component labelling; for phase unwrapping k is the wrap
// count, shape_ftp.py:1043-1080).  Roots are hooked larger-index-under-smaller by one 64-bit CAS that
// writes parent and delta together, so concurrent unions stay consistent; any (ancestor, delta) pair
// ever written stays true, so stale reads are harmless.  Rows are initialised run-wise (every pixel
// points at the first pixel of its horizontal run, delta = prefix sum along the run), which leaves only
// one union per pair of touching runs.
#pragma once
#include "common.cuh"

typedef unsigned long long ull;
#define UF_BG 0xffffffffu

__device__ __forceinline__ ull uf_pack(unsigned int parent, int delta) {
  return ((ull)(unsigned int)delta << 32) | (ull)parent;
}
__device__ __forceinline__ unsigned int uf_parent(ull v) { return (unsigned int)(v & 0xffffffffull); }
__device__ __forceinline__ int uf_delta(ull v) { return (int)(unsigned int)(v >> 32); }

// root of x and d = k[x] - k[root]
__device__ __forceinline__ void uf_find(const ull* uf, unsigned int x, unsigned int& root, int& d) {
  d = 0;
  ull v = __ldcg(uf + x);
  unsigned int par = uf_parent(v);
  while (par != x) {
    d += uf_delta(v);
    x = par;
    v = __ldcg(uf + x);
    par = uf_parent(v);
  }
  root = x;
}

// declare k[n] = k[p] + m.  Returns false when p and n were already connected with a different offset.
__device__ __forceinline__ bool uf_union(ull* uf, unsigned int p, unsigned int n, int m) {
  for (;;) {
    unsigned int rp, rn;
    int dp, dn;
    uf_find(uf, p, rp, dp);
    uf_find(uf, n, rn, dn);
    if (rp == rn) return (dn - dp) == m;
    const int krn_minus_krp = m - dn + dp;
    if (rp < rn) {
      ull expect = uf_pack(rn, 0);
      if (atomicCAS(uf + rn, expect, uf_pack(rp, krn_minus_krp)) == expect) return true;
    } else {
      ull expect = uf_pack(rp, 0);
      if (atomicCAS(uf + rp, expect, uf_pack(rn, -krn_minus_krp)) == expect) return true;
    }
  }
}

// Functor contracts:
//   Fg:   bool operator()(int b, int y, int x) const           foreground test (only called in bounds)
//   Edge: int  operator()(int b, int p, int n) const           m with k[n] = k[p] + m, p/n = pixel indices
struct EdgeZero { __device__ int operator()(int, int, int) const { return 0; } };

// One warp per row: run-wise initialisation.  grid = (ceil(h/4), batch), block = 128.
template <class Fg, class Edge>
__global__ void k_uf_init_rows(Fg fg, Edge edge, ull* __restrict__ uf, int h, int w, const int* __restrict__ active) {
  const int b = blockIdx.y, y = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (y >= h || (active && active[b] < 0)) return;
  ull* u = uf + (size_t)b * h * w;
  bool carry_run = false;
  int carry_start = 0, carry_delta = 0;
  for (int x0 = 0; x0 < w; x0 += 32) {
    const int x = x0 + lane;
    const bool f = x < w && fg(b, y, x);
    const unsigned int bits = __ballot_sync(0xffffffffu, f);
    const bool left = lane > 0 ? ((bits >> (lane - 1)) & 1u) : carry_run;
    int inc = (f && left) ? edge(b, y * w + x - 1, y * w + x) : 0;
    int s = inc;                       // inclusive warp scan of inc
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += t;
    }
    const unsigned int below = ~bits & ((1u << lane) - 1u);     // zeros strictly below this lane
    const int z = below ? 31 - __clz(below) : 0;                // highest zero below -> run starts at z + 1
    const int sz = __shfl_sync(0xffffffffu, s, z);              // S[start] == S[z] because inc[start] == 0
    int start, delta;
    if (below == 0) {                  // run reaches the chunk start
      if (carry_run) { start = carry_start; delta = carry_delta + s; }
      else { start = x0; delta = s; }  // inc of lane 0 is 0 here
    } else {
      start = x0 + z + 1;
      delta = s - sz;
    }
    if (f) u[(size_t)y * w + x] = uf_pack((unsigned int)(y * w + start), delta);
    else if (x < w) u[(size_t)y * w + x] = uf_pack(UF_BG, 0);
    carry_run = (bits >> 31) & 1u;
    carry_start = __shfl_sync(0xffffffffu, start, 31);
    carry_delta = __shfl_sync(0xffffffffu, delta, 31);
  }
}

// One thread per pixel: unions with the row above (one per pair of touching runs).  `bad` counts
// inconsistent loops (phase residues).  grid = (ceil(w/256), h, batch).
template <class Fg, class Edge>
__global__ void k_uf_merge(Fg fg, Edge edge, ull* __restrict__ uf, unsigned int* __restrict__ bad, size_t bad_stride,
                           int h, int w, const int* __restrict__ active) {
  const int b = blockIdx.z, y = blockIdx.y, x = blockIdx.x * blockDim.x + threadIdx.x;
  if (active && active[b] < 0) return;
  if (y == 0 || x >= w || !fg(b, y, x)) return;
  ull* u = uf + (size_t)b * h * w;
  const int p = y * w + x;
  const bool N = fg(b, y - 1, x);
  const bool W = x > 0 && fg(b, y, x - 1), E = x + 1 < w && fg(b, y, x + 1);
  const bool NW = x > 0 && fg(b, y - 1, x - 1), NE = x + 1 < w && fg(b, y - 1, x + 1);
  bool ok = true;
  if (N) {
    if (!W || !NW) ok = uf_union(u, p, p - w, edge(b, p, p - w));
  } else {
    if (NW && !W) ok = uf_union(u, p, p - w - 1, edge(b, p, p - w - 1));
    if (NE && !E) ok = uf_union(u, p, p - w + 1, edge(b, p, p - w + 1)) && ok;
  }
  if (!ok && bad) atomicAdd((unsigned int*)((char*)bad + (size_t)b * bad_stride), 1u);
}

// Flatten: every foreground word becomes (root, k - k[root]); optionally accumulates component areas
// (area[b][root], pre-zeroed).  grid = (ceil(npix/256), batch).
static __global__ void k_uf_flatten(ull* __restrict__ uf, unsigned int* __restrict__ area, int npix,
                                    const int* __restrict__ active) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (active && active[b] < 0) return;
  ull* u = uf + (size_t)b * npix;
  unsigned int root = UF_BG;
  if (i < npix) {
    ull v = u[i];
    if (uf_parent(v) != UF_BG) {
      int d;
      uf_find(u, (unsigned int)i, root, d);
      u[i] = uf_pack(root, d);
    }
  }
  if (area) {
    // warp-aggregated histogram: lanes with equal roots elect one leader
    const unsigned int grp = __match_any_sync(0xffffffffu, root);
    if (root != UF_BG && (int)(__ffs(grp) - 1) == (int)(threadIdx.x & 31))
      atomicAdd(area + (size_t)b * npix + root, (unsigned int)__popc(grp));
  }
}

// key of the largest component: (area << 32) | ~root  -> atomicMax picks the largest area, smallest root
static __global__ void k_uf_largest(const ull* __restrict__ uf, const unsigned int* __restrict__ area, ull* __restrict__ best,
                             size_t best_stride, int npix) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  if (uf_parent(uf[(size_t)b * npix + i]) != (unsigned int)i) return;
  ull key = ((ull)area[(size_t)b * npix + i] << 32) | (ull)(0xffffffffu - (unsigned int)i);
  atomicMax((ull*)((char*)best + (size_t)b * best_stride), key);
}

template <class Fg, class Edge>
static int uf_label(Fg fg, Edge edge, ull* uf, unsigned int* area, unsigned int* bad, size_t bad_stride, int h, int w,
                    int batch, cudaStream_t st, const int* active = nullptr) {
  // active (optional, device): frames b with active[b] < 0 are skipped by every kernel (their planes stay undefined)
  const int npix = h * w;
  VL(), k_uf_init_rows<Fg, Edge><<<dim3(cdiv(h, 4), batch), 128, 0, st>>>(fg, edge, uf, h, w, active);
  VL(), k_uf_merge<Fg, Edge><<<dim3(cdiv(w, 256), h, batch), 256, 0, st>>>(fg, edge, uf, bad, bad_stride, h, w, active);
  if (area) CK(cudaMemsetAsync(area, 0, sizeof(unsigned int) * (size_t)npix * batch, st));
  VL(), k_uf_flatten<<<dim3(cdiv(npix, 256), batch), 256, 0, st>>>(uf, area, npix, active);
  CKL();
  return 0;
}

// ------------------------------------------------------------------ chamfer distance transform
// cv2.distanceTransform(src, DIST_L2, 3): two raster sweeps with the 3x3 mask a = 0.955, b = 1.3693 in
// 16.16 fixed point; pixels outside the image are infinitely far.  Here: one block per frame walks the
// rows; inside a row the left/right propagation  d[x] = min_{x'} (t[x'] + a|x - x'|)  is a prefix-min
// scan of t[x'] -/+ a x'.  Zero: bool operator()(int b,int y,int x) -> true where the distance is 0.
#define CH_HV 62587u      // round(0.955  * 65536)
#define CH_DG 89739u      // round(1.3693 * 65536)
#define CH_INF 0x1fffffffu
#define CH_THREADS 512
#define CH_MAXPT 8        // pixels per thread -> rows up to 4096 (the 3904^2 streaming geometry of configs[4])

__device__ __forceinline__ int block_prefix_min(int v, int* sh) {   // inclusive, in thread order
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v = min(v, t);
  }
  __syncthreads();
  if (lane == 31) sh[wid] = v;
  __syncthreads();
  if (wid == 0) {
    int t = lane < (CH_THREADS >> 5) ? sh[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int s = __shfl_up_sync(0xffffffffu, t, o);
      if (lane >= o) t = min(t, s);
    }
    if (lane < (CH_THREADS >> 5)) sh[lane] = t;
  }
  __syncthreads();
  if (wid > 0) v = min(v, sh[wid - 1]);
  return v;
}

// out: float distances [batch][h][w]
template <class Zero>
__device__ __forceinline__ void chamfer_body(Zero zero, float* __restrict__ out, unsigned int* __restrict__ tmp, int h, int w);

template <class Zero>
__global__ void __launch_bounds__(CH_THREADS) k_chamfer(Zero zero, float* __restrict__ out, unsigned int* __restrict__ tmp,
                                                         int h, int w, const int* __restrict__ active) {
  if (active && active[blockIdx.x] < 0) return;      // frame skipped (its plane stays undefined)
  chamfer_body(zero, out, tmp, h, w);
}
// two independent transforms in one launch (grid.y selects): the row walk is a serial chain, so running both side by
// side halves the latency of the pair
template <class ZeroA, class ZeroB>
__global__ void __launch_bounds__(CH_THREADS) k_chamfer2(ZeroA za, float* __restrict__ outa, unsigned int* __restrict__ tmpa,
                                                          ZeroB zb, float* __restrict__ outb, unsigned int* __restrict__ tmpb,
                                                          int h, int w) {
  if (blockIdx.y == 0) chamfer_body(za, outa, tmpa, h, w);
  else chamfer_body(zb, outb, tmpb, h, w);
}

template <class Zero>
__device__ __forceinline__ void chamfer_body(Zero zero, float* __restrict__ out, unsigned int* __restrict__ tmp, int h, int w) {
  extern __shared__ unsigned int rows[];          // prev[w + 2], cur[w + 2]
  __shared__ int sh[32];
  const int b = blockIdx.x, W2 = w + 2;
  unsigned int* prev = rows;
  unsigned int* cur = rows + W2;
  unsigned int* T = tmp + (size_t)b * h * w;
  const int pt = (w + CH_THREADS - 1) / CH_THREADS;       // <= CH_MAXPT
  const int xa = threadIdx.x * pt;
  for (int i = threadIdx.x; i < 2 * W2; i += CH_THREADS) rows[i] = CH_INF;
  __syncthreads();
  // forward sweep: from above (three neighbours) then left-to-right.  The zero-set flags of row y + 1 are fetched while
  // row y goes through its barriers (the row walk is a serial chain: an exposed global load per row would dominate it).
  bool zc[CH_MAXPT], zn[CH_MAXPT];
#pragma unroll
  for (int j = 0; j < CH_MAXPT; j++) { const int x = xa + j; zc[j] = (j < pt && x < w && h > 0) ? zero(b, 0, x) : false; }
  for (int y = 0; y < h; y++) {
    unsigned int t[CH_MAXPT];
    int run = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < CH_MAXPT; j++) { const int x = xa + j; zn[j] = (j < pt && x < w && y + 1 < h) ? zero(b, y + 1, x) : false; }
#pragma unroll
    for (int j = 0; j < CH_MAXPT; j++) {
      const int x = xa + j;
      t[j] = CH_INF;
      if (j < pt && x < w) {
        unsigned int v = 0;
        if (!zc[j]) v = min(min(prev[x] + CH_DG, prev[x + 1] + CH_HV), min(prev[x + 2] + CH_DG, CH_INF));
        t[j] = v;
        run = min(run, (int)v - (int)(CH_HV * (unsigned)x));   // t[x'] - a x'   (fits: |a x| < 2^28)
      }
    }
    // exclusive-of-later prefix min over threads, then inside the thread
    int incl = block_prefix_min(run, sh);
    int before = __shfl_up_sync(0xffffffffu, incl, 1);
    if ((threadIdx.x & 31) == 0) before = threadIdx.x == 0 ? 0x7fffffff : sh[(threadIdx.x >> 5) - 1];
    int m = before;
    __syncthreads();                                   // everyone has read prev / sh
#pragma unroll
    for (int j = 0; j < CH_MAXPT; j++) {
      const int x = xa + j;
      if (j < pt && x < w) {
        m = min(m, (int)t[j] - (int)(CH_HV * (unsigned)x));
        long long cand = (long long)m + (long long)(CH_HV * (unsigned)x);
        unsigned int v = (unsigned int)min((long long)t[j], min(cand, (long long)CH_INF));
        cur[x + 1] = v;
        T[(size_t)y * w + x] = v;
      }
    }
#pragma unroll
    for (int j = 0; j < CH_MAXPT; j++) zc[j] = zn[j];
    __syncthreads();
    unsigned int* sw = prev; prev = cur; cur = sw;
  }
  // backward sweep
  for (int i = threadIdx.x; i < 2 * W2; i += CH_THREADS) rows[i] = CH_INF;
  __syncthreads();
  unsigned int tc[CH_MAXPT], tn[CH_MAXPT];
#pragma unroll
  for (int j = 0; j < CH_MAXPT; j++) { const int x = w - 1 - (xa + j); tc[j] = (j < pt && x >= 0 && h > 0) ? T[(size_t)(h - 1) * w + x] : CH_INF; }
  for (int y = h - 1; y >= 0; y--) {
    unsigned int t[CH_MAXPT];
    int run = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < CH_MAXPT; j++) { const int x = w - 1 - (xa + j); tn[j] = (j < pt && x >= 0 && y > 0) ? T[(size_t)(y - 1) * w + x] : CH_INF; }
#pragma unroll
    for (int j = 0; j < CH_MAXPT; j++) {
      const int x = w - 1 - (xa + j);                  // mirrored thread order: scan runs right-to-left
      t[j] = CH_INF;
      if (j < pt && x >= 0) {
        unsigned int v = tc[j];
        v = min(v, min(min(prev[x] + CH_DG, prev[x + 1] + CH_HV), prev[x + 2] + CH_DG));
        v = min(v, CH_INF);
        t[j] = v;
        run = min(run, (int)v + (int)(CH_HV * (unsigned)x) - (int)(CH_HV * (unsigned)w));   // t[x'] + a x' (shifted)
      }
    }
    int incl = block_prefix_min(run, sh);
    int before = __shfl_up_sync(0xffffffffu, incl, 1);
    if ((threadIdx.x & 31) == 0) before = threadIdx.x == 0 ? 0x7fffffff : sh[(threadIdx.x >> 5) - 1];
    int m = before;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < CH_MAXPT; j++) {
      const int x = w - 1 - (xa + j);
      if (j < pt && x >= 0) {
        m = min(m, (int)t[j] + (int)(CH_HV * (unsigned)x) - (int)(CH_HV * (unsigned)w));
        long long cand = (long long)m - (long long)(CH_HV * (unsigned)x) + (long long)(CH_HV * (unsigned)w);
        unsigned int v = (unsigned int)min((long long)t[j], min(cand, (long long)CH_INF));
        cur[x + 1] = v;
        out[(size_t)b * h * w + (size_t)y * w + x] = (float)v * (1.0f / 65536.0f);
      }
    }
#pragma unroll
    for (int j = 0; j < CH_MAXPT; j++) tc[j] = tn[j];
    __syncthreads();
    unsigned int* sw = prev; prev = cur; cur = sw;
  }
}

template <class ZeroA, class ZeroB>
static int chamfer_dt2(ZeroA za, float* outa, unsigned int* tmpa, ZeroB zb, float* outb, unsigned int* tmpb, int h, int w,
                       int batch, cudaStream_t st) {
  if (w > CH_THREADS * CH_MAXPT) return vistaf_set_err("chamfer_dt: row too long", cudaErrorInvalidValue, __FILE__, __LINE__);
  VL(), k_chamfer2<ZeroA, ZeroB><<<dim3(batch, 2), CH_THREADS, sizeof(unsigned int) * 2 * (size_t)(w + 2), st>>>(
      za, outa, tmpa, zb, outb, tmpb, h, w);
  CKL();
  return 0;
}
template <class Zero>
static int chamfer_dt(Zero zero, float* out, unsigned int* tmp, int h, int w, int batch, cudaStream_t st,
                      const int* active = nullptr) {
  if (w > CH_THREADS * CH_MAXPT) return vistaf_set_err("chamfer_dt: row too long", cudaErrorInvalidValue, __FILE__, __LINE__);
  VL(), k_chamfer<Zero><<<batch, CH_THREADS, sizeof(unsigned int) * 2 * (size_t)(w + 2), st>>>(zero, out, tmp, h, w, active);
  CKL();
  return 0;
}
