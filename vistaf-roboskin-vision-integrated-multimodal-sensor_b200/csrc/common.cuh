// Shared device/host helpers for libvistaf (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#define VISTAF_PI 3.14159265358979323846

int vistaf_set_err(const char* what, cudaError_t e, const char* file, int line);   // message kept per calling thread

#define CK(call)                                                              \
  do {                                                                        \
    cudaError_t e_ = (call);                                                  \
    if (e_ != cudaSuccess) return vistaf_set_err(#call, e_, __FILE__, __LINE__); \
  } while (0)
#define CKL() CK(cudaGetLastError())

// launch accounting + stage timers (api.cu).  Every kernel launch in the library is written `VL(), k<<<...>>>(...)`.
#include <atomic>
#include <functional>
extern std::atomic<long long> g_vistaf_launches;
#define VL() ((void)g_vistaf_launches.fetch_add(1, std::memory_order_relaxed))
void vistaf_stage_mark(const char* name, cudaStream_t st);      // "" closes the previous stage without opening one
bool vistaf_profiling(void);

// CUDA-graph replay cache (api.cu): v[0] = entry-point tag, v[1] = plan handle, the rest = the call's arguments.
struct VistafGraphKey {
  unsigned long long v[16];
  bool operator<(const VistafGraphKey& o) const {
    for (int i = 0; i < 16; i++) if (v[i] != o.v[i]) return v[i] < o.v[i];
    return false;
  }
};
template <class... A>
static inline VistafGraphKey vistaf_graph_key(int tag, const void* plan, A... a) {
  static_assert(sizeof...(A) <= 14, "too many key fields");
  VistafGraphKey k;
  for (int i = 0; i < 16; i++) k.v[i] = 0;
  k.v[0] = (unsigned long long)tag; k.v[1] = (unsigned long long)(uintptr_t)plan;
  const unsigned long long t[] = {(unsigned long long)a..., 0ull};
  for (size_t i = 0; i < sizeof...(A); i++) k.v[2 + i] = t[i];
  return k;
}
int vistaf_graph_run(const VistafGraphKey& key, cudaStream_t st, const std::function<int(cudaStream_t)>& body);
void vistaf_graph_drop(const void* plan);                       // forget every graph recorded for a plan (plan_destroy)

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

// ---------------------------------------------------------------- warp / block reductions
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum of NV doubles per thread; result valid in thread 0.  sh: >= NV*32 doubles.
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* sh) {
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  const int lane = tid & 31, wid = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
#pragma unroll
  for (int i = 0; i < NV; i++) v[i] = warp_sum(v[i]);
  __syncthreads();
  if (lane == 0)
#pragma unroll
    for (int i = 0; i < NV; i++) sh[i * 32 + wid] = v[i];
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int i = 0; i < NV; i++) {
      double x = lane < nw ? sh[i * 32 + lane] : 0.0;
      v[i] = warp_sum(x);
    }
  }
}

// ---------------------------------------------------------------- order-preserving float keys
__device__ __forceinline__ uint32_t f2key(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}

__device__ __forceinline__ int reflect101(int i, int n) {
  // BORDER_REFLECT_101: gfedcb|abcdefgh|gfedcba
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
  return i;
}
__device__ __forceinline__ int reflect(int i, int n) {
  // BORDER_REFLECT: fedcba|abcdefgh|hgfedcb
  while (i < 0 || i >= n) i = i < 0 ? -i - 1 : 2 * n - 1 - i;
  return i;
}

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {  // a * conj(b)
  return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
