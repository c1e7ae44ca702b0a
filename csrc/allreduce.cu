// N1: fused DDP gradient-bucket all-reduce.
//
//   out = cast( (sum over ranks of in_r) * scale )
//
// One kernel per bucket does: device-side barrier with the peers -> reduce-scatter (each rank
// reduces the tiles it owns by loading them from every peer over NVLink) -> scale + cast ->
// all-gather (the owner stores the result tile into every peer's output) -> device-side
// barrier.  No NCCL, no separate div/cast kernels, no host round trip.  Two data paths:
//   * P2P  : plain 128-bit loads/stores on peer-mapped pointers (cudaIpc / cuMem imports);
//   * NVLS : multimem.ld_reduce / multimem.st on a multicast mapping (reduction in the switch).
// The stock worker this replaces calls ncclAllReduce from torch DDP's reducer and then runs
// div_/copy kernels; the reference operator only wires MASTER_ADDR/RANK for it
// (reference controllers/pytorch/pytorchjob_controller.go:180-233).
#include "common.cuh"
#include "../include/kubedl_b200.h"
#include <cooperative_groups.h>

namespace kdl {

constexpr int kMaxWorld = KDL_MAX_WORLD;
constexpr int kMaxBlocks = KDL_AR_MAX_BLOCKS;
constexpr int kThreads = 256;

struct ArParams {
  const void* in[kMaxWorld];
  void* out[kMaxWorld];
  uint32_t* flags[kMaxWorld];  // per rank: [channels][world][kMaxBlocks]
  const void* mc_in;           // multicast alias of in (NVLS) or null
  void* mc_out;                // multicast alias of out (NVLS) or null
  int world;
  int rank;
  int channel;
  int emulate;  // 1: blockIdx.y is the rank (single-GPU protocol test, cooperative launch)
  int64_t ngroups;  // groups of 8 elements
  float scale;
  uint32_t* epoch_state;  // local device memory [ranks(emulate)|1][channels][kMaxBlocks]
  int channels;
  uint32_t* status;      // optional: written non-zero on watchdog expiry
  uint64_t timeout_ns;   // 0 = wait forever
};

__device__ __forceinline__ uint32_t* flag_slot(const ArParams& p, int owner, int from, int block) {
  return p.flags[owner] + ((size_t)p.channel * kMaxWorld + from) * kMaxBlocks + block;
}

// The barrier epoch lives in device memory (one word per channel and block) so that a captured
// CUDA graph can be replayed: every launch reads its block's epoch, uses epoch+1 / epoch+2 for
// the two barriers and stores epoch+2 back.  Block b takes part in the same launches on every
// rank (same n -> same grid), so the per-block epochs stay in lock step across ranks.
__device__ __forceinline__ uint32_t* epoch_slot(const ArParams& p, int rank) {
  size_t base = p.emulate ? (size_t)rank * p.channels * kMaxBlocks : 0;
  return p.epoch_state + base + (size_t)p.channel * kMaxBlocks + blockIdx.x;
}

// Every rank's block `b` tells every peer "I am at `target`", then waits for all of them.
__device__ __forceinline__ void block_barrier(const ArParams& p, int rank, uint32_t target) {
  __syncthreads();
  const int t = threadIdx.x;
  if (t < p.world) {
    __threadfence_system();
    st_release_sys(flag_slot(p, t, rank, blockIdx.x), target);
    const uint32_t* mine = flag_slot(p, rank, t, blockIdx.x);
    uint64_t t0 = 0;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - target) < 0) {
      if (p.timeout_ns && (++spins & 1023u) == 0) {
        uint64_t now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        if (now - t0 > p.timeout_ns) {
          if (p.status) *((volatile uint32_t*)p.status) = 0x80000000u | (uint32_t)t;
          __threadfence_system();
          __trap();
        }
      }
    }
  }
  __syncthreads();
}

template <typename T>
struct Vec8;
template <>
struct Vec8<bf16> {
  static __device__ __forceinline__ void load(const void* base, int64_t g, float* f) {
    uint4 v;
    const bf16* p = (const bf16*)base + g * 8;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    unpack8(v, f);
  }
  static __device__ __forceinline__ void store(void* base, int64_t g, const float* f) {
    st_stream((bf16*)base + g * 8, pack8(f));
  }
};
template <>
struct Vec8<float> {
  static __device__ __forceinline__ void load(const void* base, int64_t g, float* f) {
    const float* p = (const float*)base + g * 8;
    uint4 a, b;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w)
                 : "l"(p)
                 : "memory");
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(p + 4)
                 : "memory");
    f[0] = __uint_as_float(a.x); f[1] = __uint_as_float(a.y);
    f[2] = __uint_as_float(a.z); f[3] = __uint_as_float(a.w);
    f[4] = __uint_as_float(b.x); f[5] = __uint_as_float(b.y);
    f[6] = __uint_as_float(b.z); f[7] = __uint_as_float(b.w);
  }
  static __device__ __forceinline__ void store(void* base, int64_t g, const float* f) {
    float* p = (float*)base + g * 8;
    uint4 a = make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                         __float_as_uint(f[3]));
    uint4 b = make_uint4(__float_as_uint(f[4]), __float_as_uint(f[5]), __float_as_uint(f[6]),
                         __float_as_uint(f[7]));
    st_stream(p, a);
    st_stream(p + 4, b);
  }
};

// P2P two-shot all-reduce.  Tile t (kThreads groups) is owned by rank t % world.
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(kThreads) allreduce_p2p_kernel(ArParams p) {
  const int rank = p.emulate ? (int)blockIdx.y : p.rank;
  const int W = p.world;
  const uint32_t epoch = *epoch_slot(p, rank);
  block_barrier(p, rank, epoch + 1);

  const int64_t ntiles = (p.ngroups + kThreads - 1) / kThreads;
  // owner tiles of this rank: t = rank + W*k ; distribute k over blocks
  for (int64_t k = blockIdx.x;; k += gridDim.x) {
    int64_t t = rank + (int64_t)W * k;
    if (t >= ntiles) break;
    int64_t g = t * kThreads + threadIdx.x;
    if (g < p.ngroups) {
      float acc[8];
      // fixed summation order (rank 0..W-1) -> bitwise identical result on every rank
      float v[kMaxWorld][8];
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r)
        if (r < W) Vec8<TIn>::load(p.in[r], g, v[r]);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = v[0][j];
#pragma unroll
      for (int r = 1; r < kMaxWorld; ++r)
        if (r < W) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += v[r][j];
        }
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] *= p.scale;
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r)
        if (r < W) Vec8<TOut>::store(p.out[(rank + r) % W], g, acc);
    }
  }
  block_barrier(p, rank, epoch + 2);
  if (threadIdx.x == 0) *epoch_slot(p, rank) = epoch + 2;
}

// NVLS path: the switch performs the reduction (multimem.ld_reduce) and the broadcast
// (multimem.st); each rank touches only the 1/W of the bucket it owns.
__device__ __forceinline__ void mc_ld_reduce_bf16(const void* mc, float* f) {
  uint4 v;
  asm volatile(
      "multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
      : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
      : "l"(mc)
      : "memory");
  unpack8(v, f);
}
__device__ __forceinline__ void mc_st_bf16(void* mc, const float* f) {
  uint4 v = pack8(f);
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc),
               "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_ld_reduce_f32(const void* mc, float* f) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(f[0]), "=f"(f[1]), "=f"(f[2]), "=f"(f[3])
               : "l"(mc)
               : "memory");
}
__device__ __forceinline__ void mc_st_f32(void* mc, const float* f) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc),
               "f"(f[0]), "f"(f[1]), "f"(f[2]), "f"(f[3])
               : "memory");
}

template <typename T>
__global__ void __launch_bounds__(kThreads) allreduce_nvls_kernel(ArParams p) {
  const int rank = p.rank;
  const int W = p.world;
  const uint32_t epoch = *epoch_slot(p, rank);
  block_barrier(p, rank, epoch + 1);
  const int64_t ntiles = (p.ngroups + kThreads - 1) / kThreads;
  for (int64_t k = blockIdx.x;; k += gridDim.x) {
    int64_t t = rank + (int64_t)W * k;
    if (t >= ntiles) break;
    int64_t g = t * kThreads + threadIdx.x;
    if (g < p.ngroups) {
      float f[8];
      if (sizeof(T) == 2) {
        mc_ld_reduce_bf16((const bf16*)p.mc_in + g * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] *= p.scale;
        mc_st_bf16((bf16*)p.mc_out + g * 8, f);
      } else {
        mc_ld_reduce_f32((const float*)p.mc_in + g * 8, f);
        mc_ld_reduce_f32((const float*)p.mc_in + g * 8 + 4, f + 4);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] *= p.scale;
        mc_st_f32((float*)p.mc_out + g * 8, f);
        mc_st_f32((float*)p.mc_out + g * 8 + 4, f + 4);
      }
    }
  }
  block_barrier(p, rank, epoch + 2);
  if (threadIdx.x == 0) *epoch_slot(p, rank) = epoch + 2;
}

// world == 1: only the fused scale + cast remains.
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(kThreads) scale_cast_kernel(const void* in, void* out,
                                                              int64_t ngroups, float scale) {
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += T) {
    float f[8];
    Vec8<TIn>::load(in, g, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= scale;
    Vec8<TOut>::store(out, g, f);
  }
}

// Flag-only barrier across the ranks (no payload): block 0 of every rank announces epoch+1 to every peer
// and waits for all of them.  Used where two kernels of one step must be ordered ACROSS GPUs and no
// collective sits between them (DIN: every rank's forward gather before anybody's remote sparse update).
// emulate=1: gridDim.y ranks inside one cooperative launch (single-GPU protocol test).
__global__ void __launch_bounds__(kThreads) symm_barrier_kernel(ArParams p) {
  const int rank = p.emulate ? (int)blockIdx.y : p.rank;
  uint32_t* ep = epoch_slot(p, rank);
  const uint32_t epoch = *ep;
  block_barrier(p, rank, epoch + 1);
  if (threadIdx.x == 0) *ep = epoch + 1;
}

}  // namespace kdl

using namespace kdl;

extern "C" int kdl_symm_barrier(void* const* flag_ptrs, int world, int rank, void* epoch_state, int channels,
                                int channel, int emulate, void* status, unsigned long long timeout_ns,
                                cudaStream_t stream) {
  if (world < 1 || world > kMaxWorld || epoch_state == nullptr || channel < 0 || channel >= channels) {
    kdl_set_error("symm_barrier: need 1 <= world <= %d, epoch_state and 0 <= channel < channels", kMaxWorld);
    return KDL_ERR_ARG;
  }
  if (world == 1) return KDL_OK;
  ArParams p = {};
  for (int r = 0; r < world; ++r) p.flags[r] = (uint32_t*)flag_ptrs[r];
  p.world = world;
  p.rank = rank;
  p.channel = channel;
  p.emulate = emulate;
  p.epoch_state = (uint32_t*)epoch_state;
  p.channels = channels;
  p.status = (uint32_t*)status;
  p.timeout_ns = timeout_ns;
  if (emulate) {
    void* args[] = {&p};
    cudaError_t e = cudaLaunchCooperativeKernel((void*)symm_barrier_kernel, dim3(1, world), dim3(kThreads), args, 0, stream);
    if (e != cudaSuccess) {
      kdl_set_error("symm_barrier(emulate) cooperative launch failed: %s", cudaGetErrorString(e));
      return KDL_ERR_CUDA;
    }
    return KDL_OK;
  }
  symm_barrier_kernel<<<1, kThreads, 0, stream>>>(p);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" size_t kdl_allreduce_flags_bytes(int channels) {
  return (size_t)channels * kMaxWorld * kMaxBlocks * sizeof(uint32_t);
}

extern "C" int kdl_allreduce_fused(const void* const* in_ptrs, void* const* out_ptrs,
                                   void* const* flag_ptrs, const void* mc_in, void* mc_out,
                                   int world, int rank, int64_t n, int in_dtype, int out_dtype,
                                   float scale, void* epoch_state, int channels, int channel,
                                   int max_blocks,
                                   int emulate, void* status, unsigned long long timeout_ns,
                                   cudaStream_t stream) {
  if (world < 1 || world > kMaxWorld || n <= 0 || (n % 8) != 0) {
    kdl_set_error("allreduce: need 1 <= world <= %d and n %% 8 == 0 (n=%lld)", kMaxWorld,
                  (long long)n);
    return KDL_ERR_ARG;
  }
  const int64_t ngroups = n / 8;
  if (world > 1 && (epoch_state == nullptr || channel < 0 || channel >= channels)) {
    kdl_set_error("allreduce: need epoch_state and 0 <= channel < channels");
    return KDL_ERR_ARG;
  }
  if (world == 1) {
    int64_t g64 = (ngroups + kThreads * 4 - 1) / (kThreads * 4);
    int g = (int)(g64 > 148 * 8 ? 148 * 8 : (g64 < 1 ? 1 : g64));
    if (in_dtype == KDL_BF16 && out_dtype == KDL_BF16) {
      if (in_ptrs[0] == out_ptrs[0] && scale == 1.f) return KDL_OK;
      scale_cast_kernel<bf16, bf16><<<g, kThreads, 0, stream>>>(in_ptrs[0], out_ptrs[0], ngroups, scale);
    } else if (in_dtype == KDL_F32 && out_dtype == KDL_BF16) {
      scale_cast_kernel<float, bf16><<<g, kThreads, 0, stream>>>(in_ptrs[0], out_ptrs[0], ngroups, scale);
    } else if (in_dtype == KDL_F32 && out_dtype == KDL_F32) {
      if (in_ptrs[0] == out_ptrs[0] && scale == 1.f) return KDL_OK;
      scale_cast_kernel<float, float><<<g, kThreads, 0, stream>>>(in_ptrs[0], out_ptrs[0], ngroups, scale);
    } else {
      kdl_set_error("allreduce: unsupported dtype pair %d -> %d", in_dtype, out_dtype);
      return KDL_ERR_UNSUPPORTED;
    }
    KDL_CHECK_LAUNCH();
    return KDL_OK;
  }
  ArParams p = {};
  for (int r = 0; r < world; ++r) {
    p.in[r] = in_ptrs[r];
    p.out[r] = out_ptrs[r];
    p.flags[r] = (uint32_t*)flag_ptrs[r];
  }
  p.mc_in = mc_in;
  p.mc_out = mc_out;
  p.world = world;
  p.rank = rank;
  p.channel = channel;
  p.emulate = emulate;
  p.ngroups = ngroups;
  p.scale = scale;
  p.epoch_state = (uint32_t*)epoch_state;
  p.channels = channels;
  p.status = (uint32_t*)status;
  p.timeout_ns = timeout_ns;

  const int64_t ntiles = (ngroups + kThreads - 1) / kThreads;
  int64_t per_rank = (ntiles + world - 1) / world;
  int blocks = (int)(per_rank < 1 ? 1 : per_rank);
  int cap = max_blocks > 0 ? max_blocks : 148 * 2;
  if (cap > kMaxBlocks) cap = kMaxBlocks;
  if (emulate) {
    int ecap = (148 * 2) / world;
    if (cap > ecap) cap = ecap;
  }
  if (blocks > cap) blocks = cap;

  if (emulate) {
    dim3 grid(blocks, world);
    void* args[] = {&p};
    cudaError_t e;
    if (in_dtype == KDL_BF16 && out_dtype == KDL_BF16)
      e = cudaLaunchCooperativeKernel((void*)allreduce_p2p_kernel<bf16, bf16>, grid, dim3(kThreads), args, 0, stream);
    else if (in_dtype == KDL_F32 && out_dtype == KDL_BF16)
      e = cudaLaunchCooperativeKernel((void*)allreduce_p2p_kernel<float, bf16>, grid, dim3(kThreads), args, 0, stream);
    else if (in_dtype == KDL_F32 && out_dtype == KDL_F32)
      e = cudaLaunchCooperativeKernel((void*)allreduce_p2p_kernel<float, float>, grid, dim3(kThreads), args, 0, stream);
    else {
      kdl_set_error("allreduce: unsupported dtype pair %d -> %d", in_dtype, out_dtype);
      return KDL_ERR_UNSUPPORTED;
    }
    if (e != cudaSuccess) {
      kdl_set_error("allreduce(emulate) cooperative launch failed: %s", cudaGetErrorString(e));
      return KDL_ERR_CUDA;
    }
    return KDL_OK;
  }

  if (mc_in != nullptr && mc_out != nullptr && in_dtype == out_dtype) {
    if (in_dtype == KDL_BF16)
      allreduce_nvls_kernel<bf16><<<blocks, kThreads, 0, stream>>>(p);
    else
      allreduce_nvls_kernel<float><<<blocks, kThreads, 0, stream>>>(p);
    KDL_CHECK_LAUNCH();
    return KDL_OK;
  }
  if (in_dtype == KDL_BF16 && out_dtype == KDL_BF16)
    allreduce_p2p_kernel<bf16, bf16><<<blocks, kThreads, 0, stream>>>(p);
  else if (in_dtype == KDL_F32 && out_dtype == KDL_BF16)
    allreduce_p2p_kernel<float, bf16><<<blocks, kThreads, 0, stream>>>(p);
  else if (in_dtype == KDL_F32 && out_dtype == KDL_F32)
    allreduce_p2p_kernel<float, float><<<blocks, kThreads, 0, stream>>>(p);
  else {
    kdl_set_error("allreduce: unsupported dtype pair %d -> %d", in_dtype, out_dtype);
    return KDL_ERR_UNSUPPORTED;
  }
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}
