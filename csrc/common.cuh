// Shared device helpers for the kubedl_b200 kernel library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#define KDL_OK 0
#define KDL_ERR_ARG 1
#define KDL_ERR_CUDA 2
#define KDL_ERR_WORKSPACE 3
#define KDL_ERR_UNSUPPORTED 4

#define KDL_CHECK_LAUNCH()                                                   \
  do {                                                                       \
    cudaError_t e__ = cudaGetLastError();                                    \
    if (e__ != cudaSuccess) {                                                \
      kdl_set_error("%s:%d launch failed: %s", __FILE__, __LINE__,          \
                    cudaGetErrorString(e__));                                \
      return KDL_ERR_CUDA;                                                   \
    }                                                                        \
  } while (0)

#define KDL_CUDA(call)                                                       \
  do {                                                                       \
    cudaError_t e__ = (call);                                                \
    if (e__ != cudaSuccess) {                                                \
      kdl_set_error("%s:%d %s: %s", __FILE__, __LINE__, #call,              \
                    cudaGetErrorString(e__));                                \
      return KDL_ERR_CUDA;                                                   \
    }                                                                        \
  } while (0)

// error.cu
void kdl_set_error(const char* fmt, ...);

namespace kdl {

typedef __nv_bfloat16 bf16;

// Programmatic dependent launch: a kernel launched with kdl_launch_pdl may be scheduled while its predecessor in
// the stream is still draining; it must call pdl_wait() before its first global-memory access (that blocks
// until the predecessor has completed and flushed) and calls pdl_launch() to let ITS successor be scheduled
// early.  Tried on the BatchNorm kernels (half of the ~430 launches of a step): no measurable effect inside the
// replayed CUDA graph, so it is off unless KDL_PDL=1.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }


__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}

// L2 replacement hints.  Every activation is written by one kernel and read by the next; a kernel's own streaming reads
// of data that is dead afterwards should not push the freshly written output of its producer (or its own output) out of
// the 126 MB L2.  l2_policy(true) = evict_first (this line may go first), l2_policy(false) = the default priority.
__device__ __forceinline__ uint64_t l2_policy(bool evict_first) {
  uint64_t pol;
  if (evict_first) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
// 128-bit loads with an L2 policy: read-only / no L1 allocation (data read once), and the coherent variant
__device__ __forceinline__ uint4 ld_stream_hint(const void* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ uint4 ld_plain_hint(const void* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p), "l"(pol)
               : "memory");
  return r;
}

// Streaming 128-bit load that does not allocate in L1 (data read once).
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ uint4 ld_plain(const void* p) {
  return *reinterpret_cast<const uint4*>(p);
}

__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p),
               "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// System-scope flag ops used by the cross-GPU protocols.
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

}  // namespace kdl

#include <stdlib.h>
#include <utility>
namespace kdl {
inline int pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("KDL_PDL");
    v = e ? atoi(e) : 0;     // measured: no effect on the graph-replayed step (20.98 vs 21.00 ms, profiles/pdl_ab.log)
  }
  return v;
}
// <<<grid, block, smem, stream>>>, with the programmatic-stream-serialization attribute when KDL_PDL=1
template <typename... KArgs, typename... Args>
inline cudaError_t kdl_launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                  Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}
}  // namespace kdl

// ---------------------------------------------------------------------------------------------------------
// Sum of `planes` fp32 planes of `MN` elements each (plane stride MN) into a bf16 matrix, for SMALL MN: split-K
// and per-group weight-gradient partials of the narrow layers.  One thread per element walks the planes as a chain
// of dependent-latency loads (a 64 x 64 gradient over 128 planes: 16 rounds of 8 loads, 16 us for 2 MB); here SUB
// threads share an element — plane s belongs to sub-lane s % SUB — and combine through shared memory.  The
// summation order is fixed by (SUB, planes), so results stay bit-reproducible from run to run.
namespace kdl {
template <int SUB>
__global__ void __launch_bounds__(256) plane_reduce_small_kernel(const float* __restrict__ partial,
                                                                 __nv_bfloat16* __restrict__ out, int64_t MN, int N,
                                                                 int ldc, int planes) {
  constexpr int BI = 256 / SUB;                    // elements per block
  __shared__ float sm[SUB][BI + 1];
  const int sub = threadIdx.x / BI, ii = threadIdx.x % BI;
  const int64_t i = (int64_t)blockIdx.x * BI + ii;
  float acc = 0.f;
  if (i < MN) {
    int s = sub;
    for (; s + 3 * SUB < planes; s += 4 * SUB) {   // four independent loads in flight
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldcg(partial + (size_t)(s + u * SUB) * MN + i);
#pragma unroll
      for (int u = 0; u < 4; ++u) acc += v[u];
    }
    for (; s < planes; s += SUB) acc += __ldcg(partial + (size_t)s * MN + i);
  }
  sm[sub][ii] = acc;
  __syncthreads();
  if (sub == 0 && i < MN) {
    float t = 0.f;
#pragma unroll
    for (int u = 0; u < SUB; ++u) t += sm[u][ii];
    const int64_t r = i / N, c = i % N;
    out[r * ldc + c] = __float2bfloat16(t);
  }
}

// Launches the shared-element reduction when it pays (few elements, many planes); false = caller's own kernel.
inline bool launch_plane_reduce_small(const float* partial, __nv_bfloat16* out, int64_t MN, int N, int ldc, int planes,
                                      cudaStream_t stream) {
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("KDL_PLANE_REDUCE_SMALL");
    enabled = e ? atoi(e) : 1;
  }
  if (!enabled) return false;
  const int64_t kThreads = 148LL * 8 * 256;        // one full wave of 256-thread blocks
  int sub = 1;
  while (sub < 32 && MN * sub * 2 <= kThreads && sub * 2 * 4 <= planes) sub *= 2;
  if (sub < 2) return false;
  const int bi = 256 / sub;
  const unsigned grid = (unsigned)((MN + bi - 1) / bi);
  switch (sub) {
    case 2: plane_reduce_small_kernel<2><<<grid, 256, 0, stream>>>(partial, out, MN, N, ldc, planes); break;
    case 4: plane_reduce_small_kernel<4><<<grid, 256, 0, stream>>>(partial, out, MN, N, ldc, planes); break;
    case 8: plane_reduce_small_kernel<8><<<grid, 256, 0, stream>>>(partial, out, MN, N, ldc, planes); break;
    case 16: plane_reduce_small_kernel<16><<<grid, 256, 0, stream>>>(partial, out, MN, N, ldc, planes); break;
    default: plane_reduce_small_kernel<32><<<grid, 256, 0, stream>>>(partial, out, MN, N, ldc, planes); break;
  }
  return true;
}
}  // namespace kdl
