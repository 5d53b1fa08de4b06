// Exact order statistics (np.percentile / np.median semantics, method="linear") on order-preserving float
// keys, all on device, batched over frames (grid.y = frame).  The value stream comes from a functor so
// residuals / masked planes are never materialised:
//     struct Val { __device__ bool operator()(int b, long long i, float& v) const; };   // false = skip
//
// Two sweeps of the data in the common case:
//   S1  2048-bin histogram of the top 11 key bits (also yields the element count);
//   P1  locate the bin of rank k and the bin of rank k+1 (numpy interpolates between the two);
//   S2  gather the keys of those one or two bins into a per-frame list (warp-aggregated append);
//   P2  one block per frame selects ranks k, k+1 inside the list exactly and interpolates.
// If the bins hold more keys than the list can take (heavily duplicated data), the same launches fall through to
// MSB radix selection: S2 / S3 become 11- and 10-bit histogram sweeps under the fixed prefix and S4 is a
// neighbour sweep for rank k+1; on the gather path S3 / S4 exit at once.
#pragma once
#include "common.cuh"

#define SEL_BINS 2048
#define SEL_CAP (1 << 17)      // list capacity per frame (keys)

struct SelectState {           // one per frame, device memory
  unsigned int hist[SEL_BINS];
  unsigned int count;          // number of valid elements
  unsigned int k;              // rank of the lower neighbour (0-based)
  unsigned int mode;           // 0 = done / empty, 1 = gather, 2 = radix
  unsigned int b1, b1n;        // level-1 bin of rank k and of rank k+1
  unsigned int below;          // elements in bins below b1
  unsigned int nlist;          // keys gathered
  unsigned int prefix, krem;   // radix path: key bits fixed so far, rank inside the current bucket
  unsigned int cnt_le, min_gt; // radix path, neighbour sweep
  float vk, vk1;               // k-th and (k+1)-th smallest
  float result;                // interpolated quantile (float32, numpy lerp)
  unsigned int pad_[2];
};

static inline size_t sel_ws_bytes(int chunk) { return (size_t)chunk * (sizeof(SelectState) + (size_t)SEL_CAP * 4); }
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

static __global__ void k_sel_zero(SelectState* st) {
  SelectState& s = st[blockIdx.x];
  for (int i = threadIdx.x; i < SEL_BINS; i += blockDim.x) s.hist[i] = 0;
  if (threadIdx.x == 0) {
    s.count = 0; s.k = 0; s.mode = 0; s.b1 = s.b1n = 0; s.below = 0; s.nlist = 0; s.prefix = 0; s.krem = 0;
    s.cnt_le = 0; s.min_gt = 0xffffffffu;
  }
}

// ---- S1: level-1 histogram (top 11 bits).  Consecutive lanes often fall into the same bin; shared-memory atomics
// serialise only on equal addresses, and the per-block histogram is flushed once.
template <class Val>
__global__ void __launch_bounds__(256) k_sel_s1(Val val, SelectState* st, long long n) {
  __shared__ unsigned int h[SEL_BINS];
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < SEL_BINS; i += blockDim.x) h[i] = 0;
  __syncthreads();
  // four independent element fetches are in flight before the first histogram update (the value functors are
  // chains of dependent loads: mask byte, then sample)
  for (long long i0 = (long long)blockIdx.x * blockDim.x * 4 + threadIdx.x; i0 < n; i0 += (long long)gridDim.x * blockDim.x * 4) {
    float v[4];
    bool ok[4];
#pragma unroll
    for (int j = 0; j < 4; j++) { const long long i = i0 + (long long)j * blockDim.x; ok[j] = i < n && val(b, i, v[j]); }
#pragma unroll
    for (int j = 0; j < 4; j++) if (ok[j]) atomicAdd(&h[f2key(v[j]) >> 21], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SEL_BINS; i += blockDim.x)
    if (h[i]) atomicAdd(&st[b].hist[i], h[i]);
}

// Block-wide search of the bin holding 0-based rank `r` in hist[0..SEL_BINS): returns (bin, elements below it)
// to every thread.  NT threads, NT | SEL_BINS, NT a multiple of 32 and <= 1024.  sh: >= 34 unsigned ints.
template <int NT>
__device__ __forceinline__ void sel_find_bin(const unsigned int* hist, unsigned int r, unsigned int* sh, unsigned int& bin,
                                             unsigned int& below) {
  constexpr int PER = SEL_BINS / NT;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  unsigned int loc[PER], sum = 0;
#pragma unroll
  for (int j = 0; j < PER; j++) { loc[j] = hist[tid * PER + j]; sum += loc[j]; }
  unsigned int inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
  __syncthreads();
  if (lane == 31) sh[wid] = inc;
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

// ---- P1: count, ranks, level-1 bins; chooses gather or radix.
static __global__ void __launch_bounds__(256) k_sel_p1(SelectState* st, float q0, const float* qv, size_t qstride) {
  __shared__ unsigned int sh[34];
  __shared__ unsigned int tot;
  SelectState& s = st[blockIdx.x];
  const float q = sel_q(q0, qv, qstride, blockIdx.x);
  unsigned int c = 0;
  for (int i = threadIdx.x; i < SEL_BINS; i += 256) c += s.hist[i];
  c = (unsigned int)warp_sum((int)c);
  if (threadIdx.x == 0) tot = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) atomicAdd(&tot, c);
  __syncthreads();
  const unsigned int count = tot;
  if (count == 0 || q != q) {
    if (threadIdx.x == 0) { s.count = count; s.mode = 0; }
    return;
  }
  const unsigned int k = (unsigned int)floorf(sel_pos(count, q));
  const unsigned int k1 = k + 1 < count ? k + 1 : k;
  unsigned int b1, below, b1n, below_n;
  sel_find_bin<256>(s.hist, k, sh, b1, below);
  sel_find_bin<256>(s.hist, k1, sh, b1n, below_n);
  const unsigned int need = s.hist[b1] + (b1n != b1 ? s.hist[b1n] : 0u);
  const unsigned int mode = need <= (unsigned int)SEL_CAP ? 1u : 2u;
  __syncthreads();
  if (threadIdx.x == 0) {
    s.count = count; s.k = k; s.b1 = b1; s.b1n = b1n; s.below = below; s.nlist = 0; s.mode = mode;
    s.prefix = b1 << 21; s.krem = k - below;
  }
  if (mode == 2) for (int i = threadIdx.x; i < SEL_BINS; i += 256) s.hist[i] = 0;
}

// ---- S2: gather the keys of bins b1 / b1n (mode 1), or histogram key bits [20:10] under the prefix (mode 2).
template <class Val>
__global__ void __launch_bounds__(256) k_sel_s2(Val val, SelectState* st, unsigned int* lists, long long n) {
  __shared__ unsigned int h[SEL_BINS];
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  const unsigned int mode = st[b].mode;
  if (mode == 0) return;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long start = (long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31);
  if (mode == 1) {
    const unsigned int b1 = st[b].b1, b1n = st[b].b1n;
    unsigned int* list = lists + (size_t)b * SEL_CAP;
    for (long long i0 = start; i0 < n; i0 += 4 * stride) {
      float v[4];
      bool ok[4];
#pragma unroll
      for (int j = 0; j < 4; j++) { const long long i = i0 + j * stride + lane; ok[j] = i < n && val(b, i, v[j]); }
#pragma unroll
      for (int j = 0; j < 4; j++) {
        unsigned int key = 0;
        bool take = false;
        if (ok[j]) { key = f2key(v[j]); const unsigned int bin = key >> 21; take = bin == b1 || bin == b1n; }
        const unsigned int bits = __ballot_sync(0xffffffffu, take);
        if (!bits) continue;
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(&st[b].nlist, (unsigned int)__popc(bits));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (take) {
          const unsigned int at = base + (unsigned int)__popc(bits & ((1u << lane) - 1u));
          if (at < (unsigned int)SEL_CAP) list[at] = key;
        }
      }
    }
    return;
  }
  for (int i = threadIdx.x; i < SEL_BINS; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const unsigned int top = st[b].prefix >> 21;
  for (long long i = start + lane; i < n; i += stride) {
    float v;
    if (val(b, i, v)) { const unsigned int key = f2key(v); if ((key >> 21) == top) atomicAdd(&h[(key >> 10) & 2047u], 1u); }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SEL_BINS; i += blockDim.x)
    if (h[i]) atomicAdd(&st[b].hist[i], h[i]);
}

// exact key of 0-based rank r inside list[0..n) by three histogram levels (11 / 11 / 10 bits), one block of 1024
__device__ __forceinline__ unsigned int sel_list_rank(const unsigned int* __restrict__ list, unsigned int n, unsigned int r,
                                                      unsigned int* h, unsigned int* sh) {
  unsigned int prefix = 0, himask = 0;
  const int shifts[3] = {21, 10, 0};
  const unsigned int masks[3] = {2047u, 2047u, 1023u};
#pragma unroll
  for (int lv = 0; lv < 3; lv++) {
    for (int i = threadIdx.x; i < SEL_BINS; i += 1024) h[i] = 0;
    __syncthreads();
    for (unsigned int i = threadIdx.x; i < n; i += 1024) {
      const unsigned int key = list[i];
      if ((key & himask) == prefix) atomicAdd(&h[(key >> shifts[lv]) & masks[lv]], 1u);
    }
    __syncthreads();
    unsigned int bin, below;
    sel_find_bin<1024>(h, r, sh, bin, below);
    r -= below;
    prefix |= bin << shifts[lv];
    himask |= masks[lv] << shifts[lv];
  }
  return prefix;
}

// ---- P2: mode 1: exact ranks k, k+1 inside the gathered list, interpolation (done).  mode 2: level-2 digit.
static __global__ void __launch_bounds__(1024) k_sel_p2(SelectState* st, unsigned int* lists, float q0, const float* qv,
                                                        size_t qstride) {
  __shared__ unsigned int h[SEL_BINS];
  __shared__ unsigned int sh[34];
  SelectState& s = st[blockIdx.x];
  const unsigned int mode = s.mode;
  if (mode == 0) return;
  if (mode == 1) {
    const unsigned int* list = lists + (size_t)blockIdx.x * SEL_CAP;
    const unsigned int n = min(s.nlist, (unsigned int)SEL_CAP);
    const unsigned int r0 = s.k - s.below;
    const unsigned int r1 = s.k + 1 < s.count ? r0 + 1 : r0;
    const unsigned int k0 = sel_list_rank(list, n, r0, h, sh);
    const unsigned int k1 = r1 == r0 ? k0 : sel_list_rank(list, n, r1, h, sh);
    if (threadIdx.x == 0) {
      s.vk = key2f(k0); s.vk1 = key2f(k1);
      sel_interpolate(s, sel_q(q0, qv, qstride, blockIdx.x));
      s.mode = 0;
    }
    return;
  }
  unsigned int bin, below;
  sel_find_bin<1024>(s.hist, s.krem, sh, bin, below);
  __syncthreads();
  if (threadIdx.x == 0) { s.krem -= below; s.prefix |= bin << 10; }
  for (int i = threadIdx.x; i < SEL_BINS; i += 1024) s.hist[i] = 0;
}

// ---- S3 (mode 2 only): histogram of the low 10 key bits under the 22-bit prefix.
template <class Val>
__global__ void __launch_bounds__(256) k_sel_s3(Val val, SelectState* st, long long n) {
  __shared__ unsigned int h[1024];
  const int b = blockIdx.y;
  if (st[b].mode != 2) return;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const unsigned int top = st[b].prefix >> 10;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v;
    if (val(b, i, v)) { const unsigned int key = f2key(v); if ((key >> 10) == top) atomicAdd(&h[key & 1023u], 1u); }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 1024; i += blockDim.x)
    if (h[i]) atomicAdd(&st[b].hist[i], h[i]);
}
// ---- P3 (mode 2 only): final digit -> v_k; rank k+1 is the same key when the final bin holds it too.
static __global__ void __launch_bounds__(1024) k_sel_p3(SelectState* st) {
  __shared__ unsigned int sh[34];
  SelectState& s = st[blockIdx.x];
  if (s.mode != 2) return;
  unsigned int bin, below;
  sel_find_bin<1024>(s.hist, s.krem, sh, bin, below);
  if (threadIdx.x == 0) {
    s.prefix |= bin;
    s.vk = key2f(s.prefix);
    const unsigned int inbin = s.hist[bin], r = s.krem - below;
    if (r + 1 < inbin || s.k + 1 >= s.count) { s.vk1 = s.vk; s.mode = 3; }      // 3 = radix path, neighbour known
  }
}
// ---- S4 (mode 2 only): smallest key above v_k.
template <class Val>
__global__ void __launch_bounds__(256) k_sel_s4(Val val, SelectState* st, long long n) {
  const int b = blockIdx.y;
  if (st[b].mode != 2) return;
  const unsigned int kk = st[b].prefix;
  unsigned int mg = 0xffffffffu;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v;
    if (val(b, i, v)) { const unsigned int key = f2key(v); if (key > kk) mg = min(mg, key); }
  }
  mg = __reduce_min_sync(0xffffffffu, mg);
  if ((threadIdx.x & 31) == 0 && mg != 0xffffffffu) atomicMin(&st[b].min_gt, mg);
}
static __global__ void k_sel_finish(SelectState* st, float q0, const float* qv, size_t qstride) {
  SelectState& s = st[blockIdx.x];
  const float q = sel_q(q0, qv, qstride, blockIdx.x);
  if (s.count == 0 || q != q) { sel_interpolate(s, q); return; }
  if (s.mode == 2) s.vk1 = s.min_gt == 0xffffffffu ? s.vk : key2f(s.min_gt);
  if (s.mode == 2 || s.mode == 3) { sel_interpolate(s, q); s.mode = 0; }
}

// Host: full selection of quantile q (percent; q < 0 = median) for `batch` frames over n candidates each.  The
// per-frame key lists live right behind the `batch` states (sel_ws_bytes).  `need_count` is kept for call-site
// compatibility: the count always comes from the first sweep.
template <class Val>
static int select_quantile(Val val, SelectState* st, long long n, int batch, float q, bool need_count,
                           cudaStream_t stream, const float* qv = nullptr, size_t qstride = 0) {
  (void)need_count;
  const int thr = 256;
  int gx = (int)((n + thr * 8 - 1) / (thr * 8));
  if (gx > 592) gx = 592;
  if (gx < 1) gx = 1;
  dim3 grid(gx, batch);
  unsigned int* lists = sel_lists(st, batch);
  VL(), k_sel_zero<<<batch, 256, 0, stream>>>(st);
  VL(), k_sel_s1<Val><<<grid, thr, 0, stream>>>(val, st, n);
  VL(), k_sel_p1<<<batch, 256, 0, stream>>>(st, q, qv, qstride);
  VL(), k_sel_s2<Val><<<grid, thr, 0, stream>>>(val, st, lists, n);
  VL(), k_sel_p2<<<batch, 1024, 0, stream>>>(st, lists, q, qv, qstride);
  VL(), k_sel_s3<Val><<<grid, thr, 0, stream>>>(val, st, n);
  VL(), k_sel_p3<<<batch, 1024, 0, stream>>>(st);
  VL(), k_sel_s4<Val><<<grid, thr, 0, stream>>>(val, st, n);
  VL(), k_sel_finish<<<batch, 1, 0, stream>>>(st, q, qv, qstride);
  CKL();
  return 0;
}
