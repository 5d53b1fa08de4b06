// Exact order statistics (np.percentile / np.median semantics, method="linear") on order-preserving float
// keys, all on device, batched over frames (grid.y = frame).  The value stream comes from a functor so
// residuals / masked planes are never materialised:
//     struct Val { __device__ bool operator()(int b, long long i, float& v) const; };   // false = skip
//
// Three launches, two sweeps of the data:
//   S1  2048-bin histogram of the top 11 key bits; the LAST block of a frame to finish (ticket counter) counts
//       the elements, locates the bin b1 of rank k and decides gather (bin fits the list) or radix;
//   S2  keys of bin b1: 2048-bin histogram of key bits [20:10] and (gather mode) warp-aggregated append to the
//       per-frame list; when rank k+1 lies above bin b1, the smallest key above the bin is tracked by atomicMin;
//   P2  one block of 1024 threads per frame: sub-bin of rank k, then the last 10 key bits from one pass over the
//       list (gather) or over the frame's values (radix mode: heavily duplicated data), rank k+1 from the same
//       histograms, numpy's float32 interpolation, result handed to an output functor, state cleaned.
// Invariant between selections: hist = 0, nlist = 0, ticket = 0, min_next = ~0 (sel_reset establishes it once per
// workspace; P2 restores it).
#pragma once
#include "common.cuh"
#include <stdlib.h>

#define SEL_BINS 2048
#define SEL_CAP (1 << 17)      // list capacity per frame (keys)

struct SelectState {           // one per frame, device memory
  unsigned int hist[SEL_BINS];
  unsigned int count;          // number of valid elements
  unsigned int k;              // rank of the lower neighbour (0-based)
  unsigned int mode;           // 0 = empty / inactive, 1 = gather, 2 = radix
  unsigned int b1;             // level-1 bin of rank k
  unsigned int below;          // elements in bins below b1
  unsigned int inbin;          // elements in bin b1
  unsigned int nlist;          // keys gathered
  unsigned int ticket;         // blocks of S1 that have flushed
  unsigned int min_next;       // smallest key in bins above b1 (tracked when rank k+1 is not in b1)
  unsigned int need_next;
  float vk, vk1;               // k-th and (k+1)-th smallest
  float result;                // interpolated quantile (float32, numpy lerp)
  unsigned int pad_[3];
};

static inline size_t sel_ws_bytes(int chunk, unsigned int cap = SEL_CAP) { return (size_t)chunk * (sizeof(SelectState) + (size_t)cap * 4); }
__host__ __device__ inline unsigned int* sel_lists(SelectState* st, int batch) { return (unsigned int*)(st + batch); }

// NumPy >= 2 evaluates the virtual index of a float32 sample in float32:
//   pos = float32(n-1) * (float32(q) / float32(100)),  k = floor(pos), gamma = pos - k   (verified
// against np.percentile / np.nanpercentile for n up to 3e6).  q < 0 selects np.median semantics:
// k = (n-1)/2 and, for even n, the float32 mean of the two middle values.
__device__ __forceinline__ float sel_pos(unsigned int count, float q) {
  if (count == 0) return 0.f;
  if (q < 0.f) return (float)((count - 1) >> 1);
  return (float)(count - 1) * (q / 100.0f);
}
// qv != nullptr: per-frame quantile read from device memory (byte stride qstride); NaN = frame inactive.
__device__ __forceinline__ float sel_q(float q, const float* qv, size_t qstride, int b) {
  return qv ? *(const float*)((const char*)qv + (size_t)b * qstride) : q;
}
// numpy _lerp in float32: a + (b-a)*t, switched to b - (b-a)*(1-t) for t >= 0.5, t = frac(pos).
__device__ __forceinline__ void sel_interpolate(SelectState& s, float q) {
  if (s.count == 0 || q != q) { s.vk = s.vk1 = s.result = __int_as_float(0x7fc00000); return; }
  const float pos = sel_pos(s.count, q);
  const float g = pos - floorf(pos);
  const float a = s.vk, bb = s.vk1;
  if (q < 0.f) { s.result = (s.count & 1u) ? a : __fdiv_rn(__fadd_rn(a, bb), 2.0f); return; }
  const float d = __fsub_rn(bb, a);
  float r = __fadd_rn(a, __fmul_rn(d, g));
  if (g >= 0.5f) r = __fsub_rn(bb, __fmul_rn(d, __fsub_rn(1.0f, g)));
  s.result = r;
}


// Optional per-frame element count of a value functor (`long long limit(int b) const`): sweeps stop there.
template <class V>
__device__ __forceinline__ auto sel_limit(const V& v, int b, long long n, int) -> decltype(v.limit(b)) {
  const long long l = v.limit(b);
  return l < n ? l : n;
}
template <class V>
__device__ __forceinline__ long long sel_limit(const V&, int, long long n, long) { return n; }

// Optional 16-byte path of a value functor over dense, aligned streams:
//     __device__ bool vec_ok() const;                                              // alignment holds for this call
//     __device__ unsigned int vec4(int b, long long i, long long n, float v[4]) const;   // i % 4 == 0; bit j = element i + j is a sample
// The S1 / S2 sweeps then move four elements per load instruction (two such groups in flight per thread): same
// samples, same histogram counts and list contents as the scalar sweep, only the order in which they are met differs.
#ifndef VISTAF_SEL_VEC
#define VISTAF_SEL_VEC 1
#endif
template <class V>
struct sel_has_vec4 {
  template <class U> static auto test(int) -> decltype(((const U*)nullptr)->vec_ok(), char());
  template <class U> static long test(...);
  static constexpr bool value = VISTAF_SEL_VEC && sizeof(test<V>(0)) == sizeof(char);
};
__device__ __forceinline__ unsigned int sel_tail_mask(long long i, long long n) {      // elements i .. i+3 below n
  return n - i >= 4 ? 15u : ((1u << (int)(n - i)) - 1u);
}
__device__ __forceinline__ bool sel_aligned16(const void* a) { return ((unsigned long long)a & 15ull) == 0; }

static __global__ void k_sel_zero(SelectState* st) {
  SelectState& s = st[blockIdx.x];
  for (int i = threadIdx.x; i < SEL_BINS; i += blockDim.x) s.hist[i] = 0;
  if (threadIdx.x == 0) {
    s.count = 0; s.k = 0; s.mode = 0; s.b1 = 0; s.below = 0; s.inbin = 0; s.nlist = 0; s.ticket = 0;
    s.min_next = 0xffffffffu; s.need_next = 0;
  }
}
// establishes the between-selections invariant for `batch` states (once per workspace use)
static inline int sel_reset(SelectState* st, int batch, cudaStream_t stream) {
  VL(), k_sel_zero<<<batch, 256, 0, stream>>>(st);
  CKL();
  return 0;
}

// Block-wide search of the bin holding 0-based rank `r` in hist[0..NB): returns (bin, elements below it)
// to every thread.  NT threads, NT | NB, NT a multiple of 32 and <= 1024.  sh: >= 34 unsigned ints.
// If r >= total, bin = 0xffffffff.
template <int NT, int NB>
__device__ __forceinline__ void sel_find_bin(const unsigned int* hist, unsigned int r, unsigned int* sh, unsigned int& bin,
                                             unsigned int& below) {
  constexpr int PER = NB / NT;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  unsigned int loc[PER], sum = 0;
#pragma unroll
  for (int j = 0; j < PER; j++) { loc[j] = hist[tid * PER + j]; sum += loc[j]; }
  unsigned int inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
  __syncthreads();
  if (lane == 31) sh[wid] = inc;
  if (tid == 0) { sh[32] = 0xffffffffu; sh[33] = 0; }
  __syncthreads();
  unsigned int wbase = 0;
  for (int w = 0; w < wid; w++) wbase += sh[w];
  const unsigned int excl = wbase + inc - sum;
  if (r >= excl && r < excl + sum) {               // exactly one thread (sum > 0 there)
    unsigned int a = excl;
#pragma unroll
    for (int j = 0; j < PER; j++) {
      if (r < a + loc[j]) { sh[32] = (unsigned int)(tid * PER + j); sh[33] = a; break; }
      a += loc[j];
    }
  }
  __syncthreads();
  bin = sh[32]; below = sh[33];
  __syncthreads();
}
// first non-empty bin strictly above `after` (0xffffffff if none).  sh: >= 1 unsigned int.
template <int NT, int NB>
__device__ __forceinline__ unsigned int sel_next_bin(const unsigned int* hist, unsigned int after, unsigned int* sh) {
  __syncthreads();
  if (threadIdx.x == 0) sh[0] = 0xffffffffu;
  __syncthreads();
  unsigned int best = 0xffffffffu;
  for (unsigned int i = threadIdx.x; i < NB; i += NT)
    if (i > after && hist[i] && i < best) best = i;
  best = __reduce_min_sync(0xffffffffu, best);
  if ((threadIdx.x & 31) == 0 && best != 0xffffffffu) atomicMin(&sh[0], best);
  __syncthreads();
  const unsigned int r = sh[0];
  __syncthreads();
  return r;
}

// ---- S1: level-1 histogram (top 11 key bits) + rank location by the last block of each frame.
#define SEL_U 4                 // element fetches in flight per thread in the S1 / S2 sweeps
#define SEL_TILE (256 * SEL_U)  // elements per block per sweep step
template <class Val>
__global__ void __launch_bounds__(256) k_sel_s1(Val val, SelectState* st, long long n, float q0, const float* qv,
                                                size_t qstride, unsigned int cap) {
  __shared__ unsigned int h[SEL_BINS];
  __shared__ unsigned int sh[34];
  __shared__ unsigned int s_last;
  const int b = blockIdx.y;
  SelectState& s = st[b];
  for (int i = threadIdx.x; i < SEL_BINS; i += 256) h[i] = 0;
  __syncthreads();
  n = sel_limit(val, b, n, 0);
  bool swept = false;
  if constexpr (sel_has_vec4<Val>::value) {
    if (val.vec_ok()) {                                  // two 16-byte groups per thread in flight, 2048 elements per block step
      swept = true;
      for (long long base = (long long)blockIdx.x * (2 * SEL_TILE); base < n; base += (long long)gridDim.x * (2 * SEL_TILE)) {
        float v[2][4];
        unsigned int m[2];
#pragma unroll
        for (int u = 0; u < 2; u++) {
          const long long i = base + u * SEL_TILE + threadIdx.x * 4;
          m[u] = i < n ? val.vec4(b, i, n, v[u]) : 0u;
        }
#pragma unroll
        for (int u = 0; u < 2; u++)
#pragma unroll
          for (int j = 0; j < 4; j++)
            if ((m[u] >> j) & 1u) atomicAdd(&h[f2key(v[u][j]) >> 21], 1u);
      }
    }
  }
  // four independent element fetches are in flight before the first histogram update (the value functors are
  // chains of dependent loads: mask byte, then sample)
  for (long long i0 = (long long)blockIdx.x * SEL_TILE + threadIdx.x; i0 < n && !swept; i0 += (long long)gridDim.x * SEL_TILE) {
    float v[SEL_U];
    bool ok[SEL_U];
#pragma unroll
    for (int j = 0; j < SEL_U; j++) { const long long i = i0 + (long long)j * 256; ok[j] = i < n && val(b, i, v[j]); }
#pragma unroll
    for (int j = 0; j < SEL_U; j++) if (ok[j]) atomicAdd(&h[f2key(v[j]) >> 21], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SEL_BINS; i += 256)
    if (h[i]) atomicAdd(&s.hist[i], h[i]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&s.ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  // ---- last block of this frame: every other block's histogram is visible (fence + ticket)
  __threadfence();
  unsigned int c = 0;
  for (int i = threadIdx.x; i < SEL_BINS; i += 256) { const unsigned int x = __ldcg(&s.hist[i]); h[i] = x; c += x; s.hist[i] = 0; }
  c = (unsigned int)warp_sum((int)c);
  if (threadIdx.x == 0) sh[0] = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) atomicAdd(&sh[0], c);
  __syncthreads();
  const unsigned int count = sh[0];
  __syncthreads();
  const float q = sel_q(q0, qv, qstride, b);
  if (count == 0 || q != q) {
    if (threadIdx.x == 0) { s.count = count; s.mode = 0; s.ticket = 0; }
    return;
  }
  const unsigned int k = (unsigned int)floorf(sel_pos(count, q));
  unsigned int b1, below;
  sel_find_bin<256, SEL_BINS>(h, k, sh, b1, below);
  if (threadIdx.x == 0) {
    const unsigned int inbin = h[b1];
    s.count = count; s.k = k; s.b1 = b1; s.below = below; s.inbin = inbin; s.nlist = 0; s.ticket = 0;
    s.mode = inbin <= cap ? 1u : 2u;
    s.need_next = (k + 1 < count && k + 1 - below >= inbin) ? 1u : 0u;
    s.min_next = 0xffffffffu;
  }
}

// ---- S2: keys of bin b1 -> level-2 histogram (bits [20:10]) and, in gather mode, the per-frame list.
template <class Val>
__global__ void __launch_bounds__(256) k_sel_s2(Val val, SelectState* st, unsigned int* lists, long long n, unsigned int cap) {
  __shared__ unsigned int h[SEL_BINS];
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  SelectState& s = st[b];
  const unsigned int mode = s.mode;
  if (mode == 0) return;
  const unsigned int b1 = s.b1;
  const bool need_next = s.need_next != 0, gather = mode == 1;
  unsigned int* list = lists + (size_t)b * cap;
  for (int i = threadIdx.x; i < SEL_BINS; i += 256) h[i] = 0;
  __syncthreads();
  unsigned int mn = 0xffffffffu;
  n = sel_limit(val, b, n, 0);
  // warp-uniform trip count (ballots inside): lanes of a warp walk consecutive elements
  const long long stride = (long long)gridDim.x * SEL_TILE;
  const long long start = (long long)blockIdx.x * SEL_TILE + (threadIdx.x & ~31) * SEL_U;
  bool swept = false;
  if constexpr (sel_has_vec4<Val>::value) {
    if (val.vec_ok()) {                                  // block-uniform trip count (ballots inside)
      swept = true;
      for (long long base = (long long)blockIdx.x * (2 * SEL_TILE); base < n; base += (long long)gridDim.x * (2 * SEL_TILE)) {
        float v[2][4];
        unsigned int m[2];
#pragma unroll
        for (int u = 0; u < 2; u++) {
          const long long i = base + u * SEL_TILE + threadIdx.x * 4;
          m[u] = i < n ? val.vec4(b, i, n, v[u]) : 0u;
        }
        // pass 1: level-2 histogram, smallest key above the bin, and how many keys this warp appends
        unsigned int tot = 0;
#pragma unroll
        for (int u = 0; u < 2; u++)
#pragma unroll
          for (int j = 0; j < 4; j++) {
            bool take = false;
            if ((m[u] >> j) & 1u) {
              const unsigned int key = f2key(v[u][j]);
              const unsigned int bin = key >> 21;
              take = bin == b1;
              if (take) atomicAdd(&h[(key >> 10) & 2047u], 1u);
              if (need_next && bin > b1) mn = min(mn, key);
            }
            if (gather) tot += (unsigned int)__popc(__ballot_sync(0xffffffffu, take));
          }
        if (!gather || !tot) continue;                   // warp-uniform
        // pass 2: ONE list reservation per warp and block step (up to 256 keys), slots filled group by group
        unsigned int at0 = 0;
        if (lane == 0) at0 = atomicAdd(&s.nlist, tot);
        at0 = __shfl_sync(0xffffffffu, at0, 0);
#pragma unroll
        for (int u = 0; u < 2; u++)
#pragma unroll
          for (int j = 0; j < 4; j++) {
            unsigned int key = 0;
            bool take = false;
            if ((m[u] >> j) & 1u) { key = f2key(v[u][j]); take = (key >> 21) == b1; }
            const unsigned int bits = __ballot_sync(0xffffffffu, take);
            if (take) {
              const unsigned int at = at0 + (unsigned int)__popc(bits & ((1u << lane) - 1u));
              if (at < cap) list[at] = key;
            }
            at0 += (unsigned int)__popc(bits);
          }
      }
    }
  }
  for (long long i0 = start; i0 < n && !swept; i0 += stride) {
    float v[SEL_U];
    bool ok[SEL_U];
#pragma unroll
    for (int j = 0; j < SEL_U; j++) { const long long i = i0 + j * 32 + lane; ok[j] = i < n && val(b, i, v[j]); }
#pragma unroll
    for (int j = 0; j < SEL_U; j++) {
      unsigned int key = 0;
      bool take = false;
      if (ok[j]) {
        key = f2key(v[j]);
        const unsigned int bin = key >> 21;
        take = bin == b1;
        if (need_next && bin > b1) mn = min(mn, key);
      }
      if (take) atomicAdd(&h[(key >> 10) & 2047u], 1u);
      if (!gather) continue;
      const unsigned int bits = __ballot_sync(0xffffffffu, take);
      if (!bits) continue;
      unsigned int base = 0;
      if (lane == 0) base = atomicAdd(&s.nlist, (unsigned int)__popc(bits));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (take) {
        const unsigned int at = base + (unsigned int)__popc(bits & ((1u << lane) - 1u));
        if (at < cap) list[at] = key;
      }
    }
  }
  if (need_next) {
    mn = __reduce_min_sync(0xffffffffu, mn);
    if (lane == 0 && mn != 0xffffffffu) atomicMin(&s.min_next, mn);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SEL_BINS; i += 256)
    if (h[i]) atomicAdd(&s.hist[i], h[i]);
}

struct SelNoOut { __device__ void operator()(int, float, unsigned int) const {} };

// ---- P2: one block per frame finishes the selection (see the header).  Out: void operator()(int b, float result,
// unsigned int count), called by thread 0.
template <class Val, class Out>
__global__ void __launch_bounds__(1024) k_sel_p2(Val val, Out out, SelectState* st, unsigned int* lists, long long n, float q0,
                                                 const float* qv, size_t qstride, unsigned int cap, int defer) {
  __shared__ unsigned int h3[1024];
  __shared__ unsigned int sh[34];
  __shared__ unsigned int s_min;
  const int b = blockIdx.x;
  SelectState& s = st[b];
  const unsigned int mode = s.mode;
  const float q = sel_q(q0, qv, qstride, b);
  if (mode == 0) {
    if (threadIdx.x == 0) { sel_interpolate(s, q); out(b, s.result, s.count); s.min_next = 0xffffffffu; s.need_next = 0; }
    return;                                             // S1 left hist / nlist / ticket clean
  }
  n = sel_limit(val, b, n, 0);
  const unsigned int* list = lists + (size_t)b * cap;
  const unsigned int nl = min(s.nlist, cap);
  const unsigned int count = s.count, k = s.k, b1 = s.b1, inbin = s.inbin;
  const unsigned int r0 = k - s.below;                  // rank inside bin b1
  unsigned int sb, below2;
  sel_find_bin<1024, SEL_BINS>(s.hist, r0, sh, sb, below2);
  const unsigned int p22 = (b1 << 11) | sb;
  if (mode == 2 && defer) {
    // radix mode with a grid sweep to follow (S3 / P3): leave the 22-bit prefix, the rank inside the sub-bin and where
    // rank k+1 comes from (0 same sub-bin, 2 first key of the next sub-bin of b1, 3 first key above bin b1)
    const unsigned int insub = s.hist[sb], r = r0 - below2;
    unsigned int src = 0, p2 = 0;
    if (k + 1 < count && r + 1 >= insub) {
      if (r0 + 1 < inbin) { src = 2; p2 = (b1 << 11) | sel_next_bin<1024, SEL_BINS>(s.hist, sb, sh); }
      else src = 3;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < SEL_BINS; i += 1024) s.hist[i] = 0;
    if (threadIdx.x == 0) { s.b1 = p22; s.below = r; s.inbin = insub; s.need_next = src; s.nlist = p2; s.mode = 4; }
    return;
  }
  h3[threadIdx.x] = 0;
  __syncthreads();
  if (mode == 1) {
    for (unsigned int i0 = threadIdx.x; i0 < nl; i0 += 8192) {      // eight key loads in flight per thread
      unsigned int key[8];
#pragma unroll
      for (int j = 0; j < 8; j++) { const unsigned int i = i0 + j * 1024; key[j] = i < nl ? list[i] : 0u; }
#pragma unroll
      for (int j = 0; j < 8; j++)
        if (i0 + j * 1024 < nl && (key[j] >> 10) == p22) atomicAdd(&h3[key[j] & 1023u], 1u);
    }
  } else {
    for (long long i0 = threadIdx.x; i0 < n; i0 += 4096) {          // four element fetches in flight per thread
      float v[4];
      bool ok[4];
#pragma unroll
      for (int j = 0; j < 4; j++) { const long long i = i0 + j * 1024; ok[j] = i < n && val(b, i, v[j]); }
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (ok[j]) { const unsigned int key = f2key(v[j]); if ((key >> 10) == p22) atomicAdd(&h3[key & 1023u], 1u); }
    }
  }
  __syncthreads();
  const unsigned int r = r0 - below2;                   // rank inside the sub-bin
  unsigned int lo, below3;
  sel_find_bin<1024, 1024>(h3, r, sh, lo, below3);
  const unsigned int key0 = (p22 << 10) | lo;
  unsigned int key1 = key0;
  if (k + 1 < count) {
    const unsigned int insub = s.hist[sb];
    if (r + 1 - below3 < h3[lo]) key1 = key0;                                   // duplicate of the same key
    else if (r + 1 < insub) key1 = (p22 << 10) | sel_next_bin<1024, 1024>(h3, lo, sh);
    else if (r0 + 1 < inbin) {                                                  // first key of the next sub-bin of b1
      const unsigned int sb2 = sel_next_bin<1024, SEL_BINS>(s.hist, sb, sh);
      const unsigned int p2 = (b1 << 11) | sb2;
      if (threadIdx.x == 0) s_min = 0xffffffffu;
      __syncthreads();
      unsigned int mn = 0xffffffffu;
      if (mode == 1) {
        for (unsigned int i = threadIdx.x; i < nl; i += 1024) { const unsigned int key = list[i]; if ((key >> 10) == p2) mn = min(mn, key); }
      } else {
        for (long long i = threadIdx.x; i < n; i += 1024) {
          float v;
          if (val(b, i, v)) { const unsigned int key = f2key(v); if ((key >> 10) == p2) mn = min(mn, key); }
        }
      }
      mn = __reduce_min_sync(0xffffffffu, mn);
      if ((threadIdx.x & 31) == 0 && mn != 0xffffffffu) atomicMin(&s_min, mn);
      __syncthreads();
      key1 = s_min;
    } else key1 = s.min_next;                                                   // first key above bin b1
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SEL_BINS; i += 1024) s.hist[i] = 0;
  if (threadIdx.x == 0) {
    s.vk = key2f(key0); s.vk1 = key2f(key1);
    sel_interpolate(s, q);
    out(b, s.result, count);
    s.mode = 0; s.nlist = 0; s.min_next = 0xffffffffu; s.need_next = 0;
  }
}

// ---- S3 (deferred radix mode only): grid sweep for the last 10 key bits under the 22-bit prefix (+ first key of the
// next sub-bin when rank k+1 lives there).  Frames in any other mode cost one early exit per block.
template <class Val>
__global__ void __launch_bounds__(256) k_sel_s3(Val val, SelectState* st, long long n) {
  __shared__ unsigned int h[1024];
  const int b = blockIdx.y;
  SelectState& s = st[b];
  if (s.mode != 4) return;
  const unsigned int p22 = s.b1, p2 = s.nlist;
  const bool want2 = s.need_next == 2;
  for (int i = threadIdx.x; i < 1024; i += 256) h[i] = 0;
  __syncthreads();
  n = sel_limit(val, b, n, 0);
  unsigned int mn = 0xffffffffu;
  for (long long i0 = (long long)blockIdx.x * 1024 + threadIdx.x; i0 < n; i0 += (long long)gridDim.x * 1024) {
    float v[4];
    bool ok[4];
#pragma unroll
    for (int j = 0; j < 4; j++) { const long long i = i0 + (long long)j * 256; ok[j] = i < n && val(b, i, v[j]); }
#pragma unroll
    for (int j = 0; j < 4; j++)
      if (ok[j]) {
        const unsigned int key = f2key(v[j]);
        if ((key >> 10) == p22) atomicAdd(&h[key & 1023u], 1u);
        else if (want2 && (key >> 10) == p2) mn = min(mn, key);
      }
  }
  if (want2) {
    mn = __reduce_min_sync(0xffffffffu, mn);
    if ((threadIdx.x & 31) == 0 && mn != 0xffffffffu) atomicMin(&s.min_next, mn);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 1024; i += 256)
    if (h[i]) atomicAdd(&s.hist[i], h[i]);
}
// ---- P3 (deferred radix mode only): finish from the 1024-bin histogram.
template <class Out>
__global__ void __launch_bounds__(1024) k_sel_p3(Out out, SelectState* st, float q0, const float* qv, size_t qstride) {
  __shared__ unsigned int sh[34];
  const int b = blockIdx.x;
  SelectState& s = st[b];
  if (s.mode != 4) return;
  const float q = sel_q(q0, qv, qstride, b);
  const unsigned int p22 = s.b1, r = s.below, insub = s.inbin, count = s.count, k = s.k;
  unsigned int lo, below3;
  sel_find_bin<1024, 1024>(s.hist, r, sh, lo, below3);
  const unsigned int key0 = (p22 << 10) | lo;
  unsigned int key1 = key0;
  if (k + 1 < count) {
    if (r + 1 - below3 < s.hist[lo]) key1 = key0;
    else if (r + 1 < insub) key1 = (p22 << 10) | sel_next_bin<1024, 1024>(s.hist, lo, sh);
    else key1 = s.min_next;
  }
  __syncthreads();
  s.hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) {
    s.vk = key2f(key0); s.vk1 = key2f(key1);
    sel_interpolate(s, q);
    out(b, s.result, count);
    s.mode = 0; s.nlist = 0; s.min_next = 0xffffffffu; s.need_next = 0;
  }
}

// Host: selection of quantile q (percent; q < 0 = median) for `batch` frames over n candidates each.
// `lists`: batch * SEL_CAP keys.  The result lands in st[b].result / st[b].count and in `out`.
// val1 feeds the first sweep, val2 the second sweep and the finishing block: they must describe the same value
// stream (a first-sweep functor may materialise the values it computes for the second to re-read).
template <class Val1, class Val2, class Out>
static int select_quantile_out2(Val1 val1, Val2 val2, Out out, SelectState* st, unsigned int* lists, long long n, int batch,
                                float q, cudaStream_t stream, const float* qv = nullptr, size_t qstride = 0,
                                unsigned int cap = SEL_CAP, bool heavy = false) {   // heavy: duplicated data expected
  static int per_sm = 0;                        // blocks per SM over the whole grid (VISTAF_SEL_BLOCKS_PER_SM, default 16)
  if (!per_sm) { const char* e = getenv("VISTAF_SEL_BLOCKS_PER_SM"); per_sm = e ? atoi(e) : 16; if (per_sm < 1) per_sm = 16; }
  int gx = (148 * per_sm + batch - 1) / batch;  // several waves of blocks per sweep: a short last wave
  const int gmax = (int)((n + 1023) / 1024);
  if (gx > gmax) gx = gmax;
  if (gx > 592) gx = 592;
  if (gx < 1) gx = 1;
  dim3 grid(gx, batch);
  VL(), k_sel_s1<Val1><<<grid, 256, 0, stream>>>(val1, st, n, q, qv, qstride, cap);
  VL(), k_sel_s2<Val2><<<grid, 256, 0, stream>>>(val2, st, lists, n, cap);
  VL(), k_sel_p2<Val2, Out><<<batch, 1024, 0, stream>>>(val2, out, st, lists, n, q, qv, qstride, cap, heavy ? 1 : 0);
  if (heavy) {      // frames whose bin overflowed the list finish with a grid sweep instead of a one-block sweep
    VL(), k_sel_s3<Val2><<<grid, 256, 0, stream>>>(val2, st, n);
    VL(), k_sel_p3<Out><<<batch, 1024, 0, stream>>>(out, st, q, qv, qstride);
  }
  CKL();
  return 0;
}
template <class Val, class Out>
static int select_quantile_out(Val val, Out out, SelectState* st, unsigned int* lists, long long n, int batch, float q,
                               cudaStream_t stream, const float* qv = nullptr, size_t qstride = 0) {
  return select_quantile_out2(val, val, out, st, lists, n, batch, q, stream, qv, qstride);
}
// call-site compatible form: the key lists live right behind the `batch` states (sel_ws_bytes)
template <class Val>
static int select_quantile(Val val, SelectState* st, long long n, int batch, float q, bool need_count,
                           cudaStream_t stream, const float* qv = nullptr, size_t qstride = 0, unsigned int cap = SEL_CAP,
                           bool heavy = false) {
  (void)need_count;
  return select_quantile_out2(val, val, SelNoOut(), st, sel_lists(st, batch), n, batch, q, stream, qv, qstride, cap, heavy);
}
