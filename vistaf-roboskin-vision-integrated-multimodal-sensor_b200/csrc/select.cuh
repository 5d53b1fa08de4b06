// Exact order statistics (np.percentile / np.median semantics, method="linear") by MSB-first
// radix select on order-preserving float keys: four 8-bit histogram sweeps + one neighbour sweep,
// all on device, batched over frames (grid.y = frame).  The value stream comes from a functor
// so residuals / masked planes are never materialised:
//     struct Val { __device__ bool operator()(int b, long long i, float& v) const; };   // false = skip
#pragma once
#include "common.cuh"

struct SelectState {       // one per frame, device memory
  unsigned int hist[256];
  unsigned int prefix;     // key bits fixed so far
  unsigned int krem;       // rank remaining inside the current bucket
  unsigned int count;      // number of valid elements
  unsigned int cnt_le;     // elements <= v_k          (neighbour sweep)
  unsigned int min_gt;     // min key of elements > v_k (neighbour sweep)
  float vk, vk1;           // k-th and (k+1)-th smallest
  float result;            // interpolated quantile (float32, numpy lerp)
  int pad;
};

template <class Val>
__global__ void k_sel_count(Val val, SelectState* st, long long n) {
  const int b = blockIdx.y;
  unsigned int c = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v;
    if (val(b, i, v)) c++;
  }
  c = warp_sum((int)c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&st[b].count, c);
}

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
static __global__ void k_sel_begin(SelectState* st, float q0, const float* qv, size_t qstride) {
  SelectState& s = st[blockIdx.x];
  const float q = sel_q(q0, qv, qstride, blockIdx.x);
  for (int i = threadIdx.x; i < 256; i += blockDim.x) s.hist[i] = 0;
  if (threadIdx.x == 0) {
    float pos = sel_pos(s.count, q);
    s.krem = (unsigned int)floorf(pos);
    s.prefix = 0;
    s.cnt_le = 0;
    s.min_gt = 0xffffffffu;
  }
}

template <class Val>
__global__ void k_sel_hist(Val val, SelectState* st, long long n, int shift) {
  __shared__ unsigned int h[256];
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < 256; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const unsigned int prefix = st[b].prefix;
  const unsigned int himask = shift == 24 ? 0u : (0xffffffffu << (shift + 8));
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v;
    if (val(b, i, v)) {
      unsigned int k = f2key(v);
      if ((k & himask) == prefix) atomicAdd(&h[(k >> shift) & 255u], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += blockDim.x)
    if (h[i]) atomicAdd(&st[b].hist[i], h[i]);
}

static __global__ void k_sel_pick(SelectState* st, int shift) {
  SelectState& s = st[blockIdx.x];
  if (threadIdx.x == 0) {
    unsigned int k = s.krem, acc = 0;
    int d = 0;
    for (; d < 256; d++) {
      unsigned int c = s.hist[d];
      if (acc + c > k) break;
      acc += c;
    }
    if (d > 255) d = 255;
    s.krem = k - acc;
    s.prefix |= ((unsigned int)d) << shift;
    if (shift == 0) s.vk = key2f(s.prefix);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += blockDim.x) s.hist[i] = 0;
}

template <class Val>
__global__ void k_sel_neighbour(Val val, SelectState* st, long long n) {
  const int b = blockIdx.y;
  const unsigned int kk = st[b].prefix;
  unsigned int le = 0, mg = 0xffffffffu;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v;
    if (val(b, i, v)) {
      unsigned int k = f2key(v);
      if (k <= kk) le++; else mg = min(mg, k);
    }
  }
  le = warp_sum((int)le);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mg = min(mg, __shfl_xor_sync(0xffffffffu, mg, o));
  if ((threadIdx.x & 31) == 0) {
    if (le) atomicAdd(&st[b].cnt_le, le);
    if (mg != 0xffffffffu) atomicMin(&st[b].min_gt, mg);
  }
}

// numpy _lerp in float32: a + (b-a)*t, switched to b - (b-a)*(1-t) for t >= 0.5, t = frac(pos).
static __global__ void k_sel_finish(SelectState* st, float q0, const float* qv, size_t qstride) {
  SelectState& s = st[blockIdx.x];
  const float q = sel_q(q0, qv, qstride, blockIdx.x);
  if (s.count == 0 || q != q) { s.vk = s.vk1 = s.result = __int_as_float(0x7fc00000); return; }
  float pos = sel_pos(s.count, q);
  unsigned int k = (unsigned int)floorf(pos);
  float g = pos - floorf(pos);
  float a = s.vk;
  float bb = (s.cnt_le > k + 1 || s.min_gt == 0xffffffffu) ? a : key2f(s.min_gt);
  if (k + 1 >= s.count) bb = a;
  s.vk1 = bb;
  if (q < 0.f) { s.result = (s.count & 1u) ? a : __fdiv_rn(__fadd_rn(a, bb), 2.0f); return; }
  float t = g, d = __fsub_rn(bb, a);
  float r = __fadd_rn(a, __fmul_rn(d, t));
  if (t >= 0.5f) r = __fsub_rn(bb, __fmul_rn(d, __fsub_rn(1.0f, t)));
  s.result = r;
}

// Host: full selection of quantile q (percent) for `batch` frames over n candidates each.
template <class Val>
static int select_quantile(Val val, SelectState* st, long long n, int batch, float q, bool need_count,
                           cudaStream_t stream, const float* qv = nullptr, size_t qstride = 0) {
  const int thr = 256;
  int gx = (int)((n + thr * 8 - 1) / (thr * 8));
  if (gx > 592) gx = 592;
  if (gx < 1) gx = 1;
  dim3 grid(gx, batch);
  if (need_count) {
    CK(cudaMemsetAsync(st, 0, sizeof(SelectState) * batch, stream));
    VL(), k_sel_count<Val><<<grid, thr, 0, stream>>>(val, st, n);
  }
  VL(), k_sel_begin<<<batch, 256, 0, stream>>>(st, q, qv, qstride);
  for (int shift = 24; shift >= 0; shift -= 8) {
    VL(), k_sel_hist<Val><<<grid, thr, 0, stream>>>(val, st, n, shift);
    VL(), k_sel_pick<<<batch, 256, 0, stream>>>(st, shift);
  }
  VL(), k_sel_neighbour<Val><<<grid, thr, 0, stream>>>(val, st, n);
  VL(), k_sel_finish<<<batch, 1, 0, stream>>>(st, q, qv, qstride);
  CKL();
  return 0;
}
