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

#define TL_MASK 1
#define TL_BAND 2
#define TL_RING 4
#define TL_OUT 8          // ring cell reached by the outward march
#define TL_IN 16          // hole cell filled by the inward march
#define TL_MAXCLUSTERS 65536
#define TL_WARPS 4
#define TL_GRIDX 48

struct TeleaFrame {       // per-frame bookkeeping (device)
  unsigned int nclusters, next, cursor, overflow;
};
struct TeleaCluster {     // per cluster slot
  unsigned int root, count;
  int x0, y0, x1, y1;     // padded-frame bounding box, inclusive
};

struct TeleaWs {          // device pointers, dense [batch][...] planes
  uint8_t* flag;          // [Hp*Wp]
  float* T;               // [Hp*Wp]
  ull* uf;                // [Hp*Wp]
  unsigned int* slot;     // [Hp*Wp]  slot index stored at root cells
  TeleaFrame* fr;         // [1]
  TeleaCluster* cl;       // [TL_MAXCLUSTERS]
  ull* keys;              // [cap]  (T bits << 32) | seq
  int* pos;               // [cap]
  size_t sp, cap;         // padded cells per frame, queue capacity per frame
};

static inline size_t telea_ws_bytes(int h, int w, int batch) {
  const size_t sp = (size_t)(h + 2) * (w + 2), cap = sp / 2, B = (size_t)batch;
  return align256(B * sp * 8) + align256(B * sp * 4) * 2 + align256(B * sp) + align256(B * sizeof(TeleaFrame)) +
         align256(B * sizeof(TeleaCluster) * TL_MAXCLUSTERS) + align256(B * cap * 8) + align256(B * cap * 4);
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
  W.keys = (ull*)take(B * cap * 8);
  W.pos = (int*)take(B * cap * 4);
  W.sp = sp; W.cap = cap;
}

// ------------------------------------------------------------------ 1. flags on the padded frame
// Mask: bool operator()(int b, int y, int x) on image coordinates.  block (32, 8), tile 32 x 32.
template <class Mask>
__global__ void k_telea_prep(Mask mask, uint8_t* __restrict__ flag, float* __restrict__ T, int h, int w, int range) {
  extern __shared__ uint8_t sm[];                 // m[(32+2R)^2], rowhit[(32+2R) x 32]
  const int R = range, TW = 32 + 2 * R;
  uint8_t* m = sm;
  uint8_t* rh = sm + TW * TW;
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
static __global__ void k_telea_slots(const ull* __restrict__ uf, unsigned int* __restrict__ slot, TeleaFrame* fr,
                              TeleaCluster* cl, int Hp, int Wp) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t sp = (size_t)Hp * Wp;
  if (i >= Hp * Wp) return;
  if (uf_parent(uf[(size_t)b * sp + i]) != (unsigned int)i) return;
  TeleaFrame* f = fr + b;
  unsigned int s = atomicAdd(&f->nclusters, 1u);
  slot[(size_t)b * sp + i] = s;
  if (s < TL_MAXCLUSTERS) {
    TeleaCluster* c = cl + (size_t)b * TL_MAXCLUSTERS + s;
    c->root = (unsigned int)i; c->count = 0;
    c->x0 = Wp; c->y0 = Hp; c->x1 = -1; c->y1 = -1;
  }
}
static __global__ void k_telea_bbox(const ull* __restrict__ uf, const unsigned int* __restrict__ slot, TeleaCluster* cl,
                             int Hp, int Wp) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t sp = (size_t)Hp * Wp;
  if (i >= Hp * Wp) return;
  unsigned int root = uf_parent(uf[(size_t)b * sp + i]);
  if (root == UF_BG) return;
  unsigned int s = slot[(size_t)b * sp + root];
  if (s >= TL_MAXCLUSTERS) return;
  TeleaCluster* c = cl + (size_t)b * TL_MAXCLUSTERS + s;
  const int y = i / Wp, x = i - y * Wp;
  atomicAdd(&c->count, 1u);
  atomicMin(&c->x0, x); atomicMax(&c->x1, x);
  atomicMin(&c->y0, y); atomicMax(&c->y1, y);
}

// ------------------------------------------------------------------ 4. the march
// All lanes of the warp execute the scalar parts redundantly (identical stores), so no cross-lane memory
// ordering is needed; only the queue scan and the neighbourhood taps are split across lanes.
struct TeleaCtx {
  uint8_t* f; float* T; float* img; ull* keys; int* pos;
  int Hp, Wp, w, n; unsigned int seq;
};

template <bool OUTWARD>
__device__ __forceinline__ bool tl_inside(const TeleaCtx& c, int cell) {
  const int v = c.f[cell];
  return OUTWARD ? ((v & TL_RING) && !(v & TL_OUT)) : ((v & TL_MASK) && !(v & TL_IN));
}
template <bool OUTWARD>
__device__ __forceinline__ float tl_solve(const TeleaCtx& c, int c1, int c2) {
  const double a11 = (double)c.T[c1], a22 = (double)c.T[c2], m12 = a11 < a22 ? a11 : a22;
  double sol;
  if (!tl_inside<OUTWARD>(c, c1)) {
    if (!tl_inside<OUTWARD>(c, c2)) {
      if (fabs(a11 - a22) >= 1.0) sol = 1.0 + m12;
      else sol = (a11 + a22 + sqrt(2.0 - (a11 - a22) * (a11 - a22))) * 0.5;
    } else sol = 1.0 + a11;
  } else if (!tl_inside<OUTWARD>(c, c2)) sol = 1.0 + a22;
  else sol = 1.0 + m12;
  return (float)sol;
}
template <bool OUTWARD>
__device__ __forceinline__ float tl_dist(const TeleaCtx& c, int i, int j) {
  const int p = i * c.Wp + j, W = c.Wp;
  float a = tl_solve<OUTWARD>(c, p - W, p - 1), b = tl_solve<OUTWARD>(c, p + W, p - 1);
  float d = tl_solve<OUTWARD>(c, p - W, p + 1), e = tl_solve<OUTWARD>(c, p + W, p + 1);
  return fminf(fminf(a, b), fminf(d, e));
}
__device__ __forceinline__ void tl_push(TeleaCtx& c, int cell, float t) {
  c.keys[c.n] = ((ull)__float_as_uint(t) << 32) | (ull)c.seq;
  c.pos[c.n] = cell;
  c.n++; c.seq++;
}
// pops the minimum (T, seq); returns its cell
__device__ __forceinline__ int tl_pop(TeleaCtx& c, int lane) {
  ull best = ~0ull;
  int bi = -1;
  for (int e = lane; e < c.n; e += 32) {
    ull k = c.keys[e];
    if (k < best) { best = k; bi = e; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ull ok = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ok < best) { best = ok; bi = oi; }            // keys are unique (seq), no tie handling needed
  }
  const int cell = c.pos[bi];
  c.n--;
  const ull lk = c.keys[c.n];
  const int lp = c.pos[c.n];
  __syncwarp();
  c.keys[bi] = lk;
  c.pos[bi] = lp;
  __syncwarp();
  return cell;
}
// initial queue: band cells of this cluster in raster order, T = 0
__device__ __forceinline__ void tl_fill_band(TeleaCtx& c, const ull* uf, const TeleaCluster& cl, int lane) {
  c.n = 0; c.seq = 0;
  for (int y = cl.y0; y <= cl.y1; y++)
    for (int xb = cl.x0; xb <= cl.x1; xb += 32) {
      const int x = xb + lane;
      bool is = false;
      if (x <= cl.x1) {
        const int cell = y * c.Wp + x;
        is = (c.f[cell] & TL_BAND) && uf_parent(uf[cell]) == cl.root;
      }
      const unsigned int bits = __ballot_sync(0xffffffffu, is);
      if (is) {
        const int k = c.n + __popc(bits & ((1u << lane) - 1u));
        c.keys[k] = (ull)(c.seq + (unsigned)(k - c.n));     // T = 0
        c.pos[k] = y * c.Wp + x;
      }
      c.n += __popc(bits);
      c.seq += __popc(bits);
    }
  __syncwarp();
}

static __global__ void __launch_bounds__(32 * TL_WARPS) k_telea_march(TeleaWs Wsp, float* __restrict__ img,
                                                                size_t img_stride, unsigned int* status,
                                                                size_t status_stride, int h, int w, int range) {
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  const int Hp = h + 2, Wp = w + 2;
  TeleaFrame* fr = Wsp.fr + b;
  const TeleaCluster* clt = Wsp.cl + (size_t)b * TL_MAXCLUSTERS;
  const ull* uf = Wsp.uf + (size_t)b * Wsp.sp;
  const unsigned int ncl = min(fr->nclusters, (unsigned int)TL_MAXCLUSTERS);
  if (fr->nclusters > TL_MAXCLUSTERS && threadIdx.x == 0 && blockIdx.x == 0)
    atomicOr((unsigned int*)((char*)status + (size_t)b * status_stride), 0x8000u);
  TeleaCtx c;
  c.f = Wsp.flag + (size_t)b * Wsp.sp;
  c.T = Wsp.T + (size_t)b * Wsp.sp;
  c.img = img + (size_t)b * img_stride;
  c.Hp = Hp; c.Wp = Wp; c.w = w;
  const int r2 = range * range, side = 2 * range + 1, ntaps = side * side;
  for (;;) {
    unsigned int s = 0;
    if (lane == 0) s = atomicAdd(&fr->next, 1u);
    s = __shfl_sync(0xffffffffu, s, 0);
    if (s >= ncl) break;
    const TeleaCluster cl = clt[s];
    unsigned int off = 0;
    if (lane == 0) off = atomicAdd(&fr->cursor, cl.count);
    off = __shfl_sync(0xffffffffu, off, 0);
    if ((size_t)off + cl.count > Wsp.cap) {
      if (lane == 0) { atomicAdd(&fr->overflow, 1u); atomicOr((unsigned int*)((char*)status + (size_t)b * status_stride), 0x8000u); }
      continue;
    }
    c.keys = Wsp.keys + (size_t)b * Wsp.cap + off;
    c.pos = Wsp.pos + (size_t)b * Wsp.cap + off;
    // ---- outward: distances over the ring
    tl_fill_band(c, uf, cl, lane);
    while (c.n > 0) {
      const int cell = tl_pop(c, lane);
      const int ii = cell / Wp, jj = cell - ii * Wp;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int i = ii + (q == 0 ? -1 : q == 2 ? 1 : 0), j = jj + (q == 1 ? -1 : q == 3 ? 1 : 0);
        if (i <= 0 || j <= 0 || i >= Hp - 1 || j >= Wp - 1) continue;
        const int p = i * Wp + j;
        if (!tl_inside<true>(c, p)) continue;
        const float d = tl_dist<true>(c, i, j);
        c.T[p] = d;
        c.f[p] = (uint8_t)(c.f[p] | TL_OUT);
        tl_push(c, p, d);
      }
    }
    // negate the distances of every cell the outward march visited (band and reached ring)
    for (int y = max(cl.y0, 1); y <= min(cl.y1, Hp - 2); y++)
      for (int x = max(cl.x0, 1) + lane; x <= min(cl.x1, Wp - 2); x += 32) {
        const int cell = y * Wp + x;
        const int v = c.f[cell];
        if (uf_parent(uf[cell]) == cl.root && ((v & TL_BAND) || ((v & TL_RING) && (v & TL_OUT)))) c.T[cell] = -c.T[cell];
      }
    __syncwarp();
    // ---- inward: fill the hole
    tl_fill_band(c, uf, cl, lane);
    while (c.n > 0) {
      const int cell = tl_pop(c, lane);
      const int ii = cell / Wp, jj = cell - ii * Wp;
      for (int q = 0; q < 4; q++) {
        const int i = ii + (q == 0 ? -1 : q == 2 ? 1 : 0), j = jj + (q == 1 ? -1 : q == 3 ? 1 : 0);
        if (i <= 0 || j <= 0 || i >= Hp - 1 || j >= Wp - 1) continue;
        const int p = i * Wp + j;
        if (!tl_inside<false>(c, p)) continue;
        const float dist = tl_dist<false>(c, i, j);
        c.T[p] = dist;
        float gTx, gTy;
        if (!tl_inside<false>(c, p + 1)) {
          if (!tl_inside<false>(c, p - 1)) gTx = __fmul_rn(__fsub_rn(c.T[p + 1], c.T[p - 1]), 0.5f);
          else gTx = __fsub_rn(c.T[p + 1], dist);
        } else {
          if (!tl_inside<false>(c, p - 1)) gTx = __fsub_rn(dist, c.T[p - 1]);
          else gTx = 0.f;
        }
        if (!tl_inside<false>(c, p + Wp)) {
          if (!tl_inside<false>(c, p - Wp)) gTy = __fmul_rn(__fsub_rn(c.T[p + Wp], c.T[p - Wp]), 0.5f);
          else gTy = __fsub_rn(c.T[p + Wp], dist);
        } else {
          if (!tl_inside<false>(c, p - Wp)) gTy = __fsub_rn(dist, c.T[p - Wp]);
          else gTy = 0.f;
        }
        float Ia = 0.f, Jx = 0.f, Jy = 0.f, sw = 1.0e-20f;
        for (int t0 = 0; t0 < ntaps; t0 += 32) {
          const int tap = t0 + lane;
          bool act = false;
          float wgt = 0.f, ta = 0.f, tx = 0.f, ty = 0.f;
          if (tap < ntaps) {
            const int dk = tap / side - range, dl = tap - (tap / side) * side - range;
            const int k = i + dk, l = j + dl;
            if (k > 0 && l > 0 && k < Hp - 1 && l < Wp - 1 && dk * dk + dl * dl <= r2) {
              const int pk = k * Wp + l;
              if (!tl_inside<false>(c, pk)) {
                act = true;
                const int km = k - 1 + (k == 1), kp = k - 1 - (k == Hp - 2);
                const int lm = l - 1 + (l == 1), lp = l - 1 - (l == Wp - 2);
                const float ry = (float)(i - k), rx = (float)(j - l);
                const float len2 = __fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry));
                const float dst = (float)(1.0 / (double)__fmul_rn(len2, __fsqrt_rn(len2)));
                const float lev = (float)(1.0 / (1.0 + fabs((double)__fsub_rn(c.T[pk], dist))));
                float dir = __fadd_rn(__fmul_rn(rx, gTx), __fmul_rn(ry, gTy));
                if (fabs((double)dir) <= 0.01) dir = 0.000001f;
                wgt = fabsf(__fmul_rn(__fmul_rn(dst, lev), dir));
#define TL_O(a, bb) c.img[(size_t)(a) * c.w + (bb)]
                float gIx, gIy;
                if (!tl_inside<false>(c, pk + 1)) {
                  if (!tl_inside<false>(c, pk - 1)) gIx = __fmul_rn(__fsub_rn(TL_O(km, lp + 1), TL_O(km, lm - 1)), 2.0f);
                  else gIx = __fsub_rn(TL_O(km, lp + 1), TL_O(km, lm));
                } else {
                  if (!tl_inside<false>(c, pk - 1)) gIx = __fsub_rn(TL_O(km, lp), TL_O(km, lm - 1));
                  else gIx = 0.f;
                }
                if (!tl_inside<false>(c, pk + Wp)) {
                  if (!tl_inside<false>(c, pk - Wp)) gIy = __fmul_rn(__fsub_rn(TL_O(kp + 1, lm), TL_O(km - 1, lm)), 2.0f);
                  else gIy = __fsub_rn(TL_O(kp + 1, lm), TL_O(km, lm));
                } else {
                  if (!tl_inside<false>(c, pk - Wp)) gIy = __fsub_rn(TL_O(kp, lm), TL_O(km - 1, lm));
                  else gIy = 0.f;
                }
                ta = __fmul_rn(wgt, TL_O(k - 1, l - 1));
                tx = __fmul_rn(wgt, __fmul_rn(gIx, rx));
                ty = __fmul_rn(wgt, __fmul_rn(gIy, ry));
#undef TL_O
              }
            }
          }
          unsigned int bits = __ballot_sync(0xffffffffu, act);
          while (bits) {                               // raster-order accumulation, identical on every lane
            const int src = __ffs(bits) - 1;
            bits &= bits - 1;
            Ia = __fadd_rn(Ia, __shfl_sync(0xffffffffu, ta, src));
            Jx = __fsub_rn(Jx, __shfl_sync(0xffffffffu, tx, src));
            Jy = __fsub_rn(Jy, __shfl_sync(0xffffffffu, ty, src));
            sw = __fadd_rn(sw, __shfl_sync(0xffffffffu, wgt, src));
          }
        }
        const float nrm = __fadd_rn(__fsqrt_rn(__fadd_rn(__fmul_rn(Jx, Jx), __fmul_rn(Jy, Jy))), 1.0e-20f);
        const float val = __fadd_rn(__fdiv_rn(Ia, sw), __fdiv_rn(__fadd_rn(Jx, Jy), nrm));
        c.img[(size_t)(i - 1) * w + (j - 1)] = val;
        c.f[p] = (uint8_t)(c.f[p] | TL_IN);
        tl_push(c, p, dist);
        __syncwarp();
      }
    }
  }
}

// Host driver.  mask functor on image coordinates; img is modified in place on the hole pixels.
template <class Mask>
static int telea_inpaint(Mask mask, float* img, size_t img_stride, unsigned int* status, size_t status_stride,
                         char* ws, int h, int w, int batch, int range, cudaStream_t st) {
  if (range < 1 || range > 8) return vistaf_set_err("telea: radius out of range", cudaErrorInvalidValue, __FILE__, __LINE__);
  TeleaWs W;
  telea_ws_carve(ws, h, w, batch, W);
  const int Hp = h + 2, Wp = w + 2, sp = Hp * Wp;
  const int TW = 32 + 2 * range;
  VL(), k_telea_prep<Mask><<<dim3(cdiv(Wp, 32), cdiv(Hp, 32), batch), dim3(32, 8), (size_t)TW * TW + (size_t)TW * 32, st>>>(
      mask, W.flag, W.T, h, w, range);
  CKL();
  TeleaFg fg{W.flag, (size_t)sp, Wp};
  if (uf_label(fg, EdgeZero(), W.uf, (unsigned int*)nullptr, (unsigned int*)nullptr, 0, Hp, Wp, batch, st)) return 1;
  CK(cudaMemsetAsync(W.fr, 0, sizeof(TeleaFrame) * (size_t)batch, st));
  VL(), k_telea_slots<<<dim3(cdiv(sp, 256), batch), 256, 0, st>>>(W.uf, W.slot, W.fr, W.cl, Hp, Wp);
  VL(), k_telea_bbox<<<dim3(cdiv(sp, 256), batch), 256, 0, st>>>(W.uf, W.slot, W.cl, Hp, Wp);
  VL(), k_telea_march<<<dim3(TL_GRIDX, batch), 32 * TL_WARPS, 0, st>>>(W, img, img_stride, status, status_stride, h, w, range);
  CKL();
  return 0;
}
