// N4 / N5: sharded sparse-embedding exchange fused with the adjacent compute, over NVLink peer
// memory (BASELINE.json config #5 "XDLJob-shaped sparse CTR model ...
// embedding all-to-all dispatch/combine over NVSwitch").
//
// The table E[V, D] (fp32 master rows) is row-sharded: rank r owns rows [r*rows_per_rank, ...),
// every shard lives in the symmetric heap, so any rank can address any row directly.
//
//   din_gather_attend (N4, forward "dispatch"): one warp per sample PULLS the user row, the
//       candidate row and L history rows straight from their owners' shards (no id exchange, no
//       vector all-to-all), computes DIN-style attention a = softmax(<h_i, c>/sqrt(D)), pools
//       p = sum a_i h_i, and writes the MLP input [u, c, p, p*c] in bf16 — gather, pooling and
//       concat in one pass.  History rows and weights are kept locally for the backward pass.
//   din_scatter_update (N5, backward "combine"): one warp per sample back-propagates through the
//       pooling, then pushes -lr * grad for every touched row to its owner with vectorised
//       red.global.add.v4.f32 over NVLink — scatter-add, duplicate coalescing (atomics) and the
//       sparse SGD update in one pass; no gradient all-to-all and no dense gradient table.
//
// Stock equivalent: all_to_all_single(ids) + embedding_bag + all_to_all_single(vectors) forward,
// all_to_all_single(grads) + index_add_ + sparse optimizer backward.  The reference operator only
// provides TASK_NAME / TASK_INDEX / ZK_ADDR for XDL's own PS exchange (reference
// controllers/xdl/xdljob_controller.go:191-217).
#include "common.cuh"
#include "tcgen05.cuh"
#include "../include/kubedl_b200.h"
#include <stdlib.h>

namespace kdl {

constexpr int kD = 64;            // embedding width; half a warp owns a row, lane hl holds dims 4hl..4hl+3
constexpr int kWarpsPerBlock = 8;

struct EmbParams {
  float* shard[KDL_MAX_WORLD];    // peer-mapped shard base pointers
  int world;
  int64_t rows_per_rank;
  int interleaved;                // 1: row id lives on rank id % world at local row id / world; 0: row ranges
  int wshift;                     // log2(world) when world is a power of two, else -1
  const int64_t* user;            // [B]
  const int64_t* cand;            // [B]
  const int64_t* hist;            // [B, L], -1 = padding
  int B, L;
  bf16* x;                        // [B, 4*D] MLP input (fwd out)
  bf16* hsave;                    // [B, L, D] gathered history rows (fwd out, bwd in)
  float* asave;                   // [B, L] attention weights (fwd out, bwd in)
  const bf16* dx;                 // [B, 4*D] gradient of the MLP input (bwd in)
  float lr;
};

// Placement of row `id`.  Interleaved (default): consecutive ids go to consecutive ranks, so a skewed id
// distribution (hot low ids, the usual case for CTR vocabularies sorted by frequency) spreads evenly over the
// owners and every NVLink port carries the same load; NVSwitch gives the same bandwidth to every peer, so
// nothing is gained by keeping neighbours together.  With row RANGES half of the Zipf-skewed lookups of the
// benchmark hit rank 0 and its egress port alone set the kernel time (profiles/r2_din_sharding.md).
__device__ __forceinline__ const float* row_ptr(const EmbParams& p, int64_t id) {
  if (p.interleaved) {
    const uint32_t i = (uint32_t)id;          // host checked: vocab < 2^32
    uint32_t owner, local;
    if (p.wshift >= 0) { owner = i & (uint32_t)(p.world - 1); local = i >> p.wshift; }
    else { local = i / (uint32_t)p.world; owner = i - local * (uint32_t)p.world; }
    return p.shard[owner] + (size_t)local * kD;
  }
  const int owner = (int)(id / p.rows_per_rank);
  return p.shard[owner] + (id - (int64_t)owner * p.rows_per_rank) * kD;
}

// A row is 64 fp32 = 256 B.  In the scatter-update kernel half a warp owns a row: lane hl = lane & 15 holds
// dims 4hl..4hl+3, so every remote reduction is a 16-byte request per lane and one warp instruction covers
// TWO rows.
__device__ __forceinline__ float4 ld_row4(const float* row, int hl) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(row + 4 * hl)
               : "memory");
  return v;
}

__device__ __forceinline__ void red_add4(float* row, int hl, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(row + 4 * hl), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

__device__ __forceinline__ float half_sum(float v) {     // over the 16 lanes of a half warp
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float dot4(const float4& a, const float4& b) {
  return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
}

__device__ __forceinline__ void st_bf16x4(bf16* dst, float a, float b, float c, float d) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
  uint2 o;
  o.x = *reinterpret_cast<uint32_t*>(&lo);
  o.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(dst) = o;
}

__device__ __forceinline__ float4 ld_bf16x4(const bf16* src) {
  const uint2 v = *reinterpret_cast<const uint2*>(src);
  const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.x));
  const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.y));
  return make_float4(lo.x, lo.y, hi.x, hi.y);
}

// Forward gather: one warp per row, 8 bytes per lane (measured faster than the half-warp layout for plain
// loads: 80 vs 99 us at 2 GPUs); the scatter-update below uses the half-warp / 16-byte layout, which is what
// helps the remote reductions (70 -> 62 us at 2 GPUs, 38 -> 25 us locally).
__device__ __forceinline__ float2 ld_row2(const float* row, int lane) {
  float2 v;
  asm volatile("ld.global.L1::no_allocate.v2.f32 {%0,%1}, [%2];"
               : "=f"(v.x), "=f"(v.y)
               : "l"(row + 2 * lane)
               : "memory");
  return v;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32) din_gather_attend_kernel(EmbParams p) {
  const int lane = threadIdx.x & 31;
  const int s = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (s >= p.B) return;
  const float inv_sqrt_d = rsqrtf((float)kD);
  const float2 u = ld_row2(row_ptr(p, p.user[s]), lane);
  const float2 c = ld_row2(row_ptr(p, p.cand[s]), lane);
  const int64_t* hid = p.hist + (size_t)s * p.L;
  bf16* hs = p.hsave + (size_t)s * p.L * kD;
  float* as = p.asave + (size_t)s * p.L;

  // pass 1: pull history rows, 8 remote rows in flight per warp (16 measured slower: 114 vs 80 us at
  // 2 GPUs — occupancy, not round trips, is what matters here), scores -> online softmax
  float m = -1e30f, denom = 0.f;
  float2 acc = make_float2(0.f, 0.f);
  constexpr int U = 8;
  for (int i0 = 0; i0 < p.L; i0 += U) {
    float2 h[U];
    int64_t id[U];
#pragma unroll
    for (int k = 0; k < U; ++k) {
      id[k] = (i0 + k < p.L) ? hid[i0 + k] : -1;
      h[k] = id[k] >= 0 ? ld_row2(row_ptr(p, id[k]), lane) : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int k = 0; k < U; ++k) {
      if (i0 + k >= p.L) break;
      *reinterpret_cast<__nv_bfloat162*>(hs + (size_t)(i0 + k) * kD + 2 * lane) = __floats2bfloat162_rn(h[k].x, h[k].y);
      if (id[k] < 0) {              // padding: weight 0
        if (lane == 0) as[i0 + k] = -1e30f;
        continue;
      }
      float sc = warp_sum(h[k].x * c.x + h[k].y * c.y) * inv_sqrt_d;
      if (lane == 0) as[i0 + k] = sc;
      const float mn = fmaxf(m, sc);
      const float corr = __expf(m - mn), e = __expf(sc - mn);
      denom = denom * corr + e;
      acc.x = acc.x * corr + e * h[k].x;
      acc.y = acc.y * corr + e * h[k].y;
      m = mn;
    }
  }
  const float inv = denom > 0.f ? 1.f / denom : 0.f;
  const float2 pl = make_float2(acc.x * inv, acc.y * inv);
  __syncwarp();
  // normalise the saved scores into weights
  for (int i = lane; i < p.L; i += 32) {
    float sc = as[i];
    as[i] = sc > -1e29f ? __expf(sc - m) * inv : 0.f;
  }
  bf16* x = p.x + (size_t)s * 4 * kD;
  *reinterpret_cast<__nv_bfloat162*>(x + 2 * lane) = __floats2bfloat162_rn(u.x, u.y);
  *reinterpret_cast<__nv_bfloat162*>(x + kD + 2 * lane) = __floats2bfloat162_rn(c.x, c.y);
  *reinterpret_cast<__nv_bfloat162*>(x + 2 * kD + 2 * lane) = __floats2bfloat162_rn(pl.x, pl.y);
  *reinterpret_cast<__nv_bfloat162*>(x + 3 * kD + 2 * lane) = __floats2bfloat162_rn(pl.x * c.x, pl.y * c.y);
}

// ---- N4, second version: the rows come in by TMA bulk copies ----------------------------------------------
// The register version above keeps 8 rows in flight per warp (more costs occupancy) and spends an LSU request
// per row per warp.  Here every lane ISSUES the rows it is responsible for as 256-byte cp.async.bulk copies
// (global -> shared, straight from the owner's shard over NVLink when the row is remote) onto one mbarrier per
// warp: all of a sample's rows (up to 2 + L) are in flight at once, no registers hold them, and the attention
// pooling then reads them from shared memory.  Plain bulk copies, no multicast.
constexpr int kTmaWarps = 6;                       // per block: 6 x 16.9 KB of row buffers, 2 blocks per SM
constexpr int kTmaMaxRows = 66;                    // user + candidate + up to 64 history rows

__global__ void __launch_bounds__(kTmaWarps * 32) din_gather_attend_tma_kernel(EmbParams p) {
  extern __shared__ __align__(128) uint8_t dsm[];
  const int lane = threadIdx.x & 31;
  const int w = threadIdx.x >> 5;
  float* rows = reinterpret_cast<float*>(dsm) + (size_t)w * kTmaMaxRows * kD;     // [row][64] fp32
  const uint32_t rows_u32 = smem_u32(rows);
  const uint32_t bar = smem_u32(dsm + (size_t)kTmaWarps * kTmaMaxRows * kD * 4) + 8u * w;
  if (lane == 0) mbar_init(bar, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  const float inv_sqrt_d = rsqrtf((float)kD);
  uint32_t phase = 0;
  for (int s = blockIdx.x * kTmaWarps + w; s < p.B; s += gridDim.x * kTmaWarps) {
    const int64_t* hid = p.hist + (size_t)s * p.L;
    // slot 0 = user, 1 = candidate, 2 + i = history item i; lane l issues slots l, l + 32, l + 64
    int64_t id[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int slot = lane + 32 * k;
      id[k] = slot == 0 ? p.user[s] : slot == 1 ? p.cand[s] : (slot - 2 < p.L ? hid[slot - 2] : -1);
    }
    // which slots hold a real row (warp-uniform bit masks)
    const uint32_t vm0 = __ballot_sync(0xffffffffu, id[0] >= 0), vm1 = __ballot_sync(0xffffffffu, id[1] >= 0),
                   vm2 = __ballot_sync(0xffffffffu, id[2] >= 0);
    const int nvalid = __popc(vm0) + __popc(vm1) + __popc(vm2);
    if (lane == 0) mbar_expect_tx(bar, (uint32_t)nvalid * kD * 4);       // arrive + expected bytes
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int slot = lane + 32 * k;
      if (id[k] >= 0)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(rows_u32 + (uint32_t)slot * kD * 4), "l"(row_ptr(p, id[k])), "r"(kD * 4), "r"(bar)
                     : "memory");
    }
    mbar_wait(bar, phase);
    phase ^= 1u;

    const float2 u = *reinterpret_cast<const float2*>(rows + 2 * lane);
    const float2 c = *reinterpret_cast<const float2*>(rows + kD + 2 * lane);
    bf16* hs = p.hsave + (size_t)s * p.L * kD;
    float* as = p.asave + (size_t)s * p.L;
    float m = -1e30f, denom = 0.f;
    float2 acc = make_float2(0.f, 0.f);
    for (int i = 0; i < p.L; ++i) {
      const int slot = i + 2;
      const bool valid = (((slot < 32 ? vm0 : slot < 64 ? vm1 : vm2) >> (slot & 31)) & 1u) != 0;
      const float2 h = valid ? *reinterpret_cast<const float2*>(rows + (size_t)slot * kD + 2 * lane) : make_float2(0.f, 0.f);
      *reinterpret_cast<__nv_bfloat162*>(hs + (size_t)i * kD + 2 * lane) = __floats2bfloat162_rn(h.x, h.y);
      if (!valid) {                 // padding: weight 0
        if (lane == 0) as[i] = -1e30f;
        continue;
      }
      const float sc = warp_sum(h.x * c.x + h.y * c.y) * inv_sqrt_d;
      if (lane == 0) as[i] = sc;
      const float mn = fmaxf(m, sc);
      const float corr = __expf(m - mn), e = __expf(sc - mn);
      denom = denom * corr + e;
      acc.x = acc.x * corr + e * h.x;
      acc.y = acc.y * corr + e * h.y;
      m = mn;
    }
    const float inv = denom > 0.f ? 1.f / denom : 0.f;
    const float2 pl = make_float2(acc.x * inv, acc.y * inv);
    __syncwarp();
    for (int i = lane; i < p.L; i += 32) {
      const float sc = as[i];
      as[i] = sc > -1e29f ? __expf(sc - m) * inv : 0.f;
    }
    bf16* x = p.x + (size_t)s * 4 * kD;
    *reinterpret_cast<__nv_bfloat162*>(x + 2 * lane) = __floats2bfloat162_rn(u.x, u.y);
    *reinterpret_cast<__nv_bfloat162*>(x + kD + 2 * lane) = __floats2bfloat162_rn(c.x, c.y);
    *reinterpret_cast<__nv_bfloat162*>(x + 2 * kD + 2 * lane) = __floats2bfloat162_rn(pl.x, pl.y);
    *reinterpret_cast<__nv_bfloat162*>(x + 3 * kD + 2 * lane) = __floats2bfloat162_rn(pl.x * c.x, pl.y * c.y);
    __syncwarp();      // every lane is done with the row buffers before the next sample's copies land in them
  }
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32) din_scatter_update_kernel(EmbParams p) {
  const int lane = threadIdx.x & 31;
  const int hl = lane & 15, half = lane >> 4;
  const int s = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (s >= p.B) return;
  const float inv_sqrt_d = rsqrtf((float)kD);
  const bf16* dx = p.dx + (size_t)s * 4 * kD;
  const float4 du = ld_bf16x4(dx + 4 * hl), dc0 = ld_bf16x4(dx + kD + 4 * hl);
  const float4 dp0 = ld_bf16x4(dx + 2 * kD + 4 * hl), dpc = ld_bf16x4(dx + 3 * kD + 4 * hl);
  const int64_t cid = p.cand[s];
  // the candidate vector exactly as the MLP saw it (the table row may already be receiving
  // updates from other samples in this very launch)
  const float4 c = ld_bf16x4(p.x + (size_t)s * 4 * kD + kD + 4 * hl);
  const bf16* hs = p.hsave + (size_t)s * p.L * kD;
  const float* as = p.asave + (size_t)s * p.L;
  const int64_t* hid = p.hist + (size_t)s * p.L;
  // recompute p = sum a_i h_i (each half sums every other row, then the halves are added)
  float4 pl = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = half; i < p.L; i += 2) {
    const float a = as[i];
    const float4 h = ld_bf16x4(hs + (size_t)i * kD + 4 * hl);
    pl.x += a * h.x; pl.y += a * h.y; pl.z += a * h.z; pl.w += a * h.w;
  }
  pl.x += __shfl_xor_sync(0xffffffffu, pl.x, 16); pl.y += __shfl_xor_sync(0xffffffffu, pl.y, 16);
  pl.z += __shfl_xor_sync(0xffffffffu, pl.z, 16); pl.w += __shfl_xor_sync(0xffffffffu, pl.w, 16);
  const float4 dp = make_float4(dp0.x + dpc.x * c.x, dp0.y + dpc.y * c.y, dp0.z + dpc.z * c.z, dp0.w + dpc.w * c.w);
  float4 dc = make_float4(0.f, 0.f, 0.f, 0.f);
  const float t = half_sum(dot4(dp, pl));   // sum_j a_j da_j  (da_j = <dp, h_j>); same in both halves
  const float nlr = -p.lr;
  const int Lp = (p.L + 1) & ~1;            // both halves run the same number of iterations (shuffles)
  for (int i0 = 0; i0 < Lp; i0 += 2) {
    const int i = i0 + half;
    const bool ok = i < p.L;
    const int64_t id = ok ? hid[i] : -1;
    const float a = ok ? as[i] : 0.f;
    const float4 h = ok ? ld_bf16x4(hs + (size_t)i * kD + 4 * hl) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float da = half_sum(dot4(dp, h));
    if (id < 0) continue;
    const float ds = a * (da - t) * inv_sqrt_d;
    // dh_i = a_i dp + ds_i c ; dc += ds_i h_i
    red_add4(const_cast<float*>(row_ptr(p, id)), hl, nlr * (a * dp.x + ds * c.x), nlr * (a * dp.y + ds * c.y),
             nlr * (a * dp.z + ds * c.z), nlr * (a * dp.w + ds * c.w));
    dc.x += ds * h.x; dc.y += ds * h.y; dc.z += ds * h.z; dc.w += ds * h.w;
  }
  dc.x += __shfl_xor_sync(0xffffffffu, dc.x, 16); dc.y += __shfl_xor_sync(0xffffffffu, dc.y, 16);
  dc.z += __shfl_xor_sync(0xffffffffu, dc.z, 16); dc.w += __shfl_xor_sync(0xffffffffu, dc.w, 16);
  if (half == 0) {
    red_add4(const_cast<float*>(row_ptr(p, cid)), hl, nlr * (dc0.x + dpc.x * pl.x + dc.x), nlr * (dc0.y + dpc.y * pl.y + dc.y),
             nlr * (dc0.z + dpc.z * pl.z + dc.z), nlr * (dc0.w + dpc.w * pl.w + dc.w));
  } else {
    red_add4(const_cast<float*>(row_ptr(p, p.user[s])), hl, nlr * du.x, nlr * du.y, nlr * du.z, nlr * du.w);
  }
}

// ---- the elementwise glue of the DIN MLP, fused ----------------------------------------------------------------
// At B = 4096 every tensor of the MLP is a few megabytes, so the step is bound by the NUMBER of launches; the
// loss and the three "ReLU backward + bias gradient" chains were 29 ATen kernels (profiles/r2_din_step_*.csv).

// Binary cross-entropy with logits on column 0 of z [B, ldz] (bf16), mean over B, and its gradient as the padded
// 8-column bf16 matrix the wgrad / dgrad GEMMs take: dz[:, 0] = (sigmoid(z) - y) / B, dz[:, 1:8] = 0.  One block.
__global__ void __launch_bounds__(1024) din_bce_grad_kernel(const bf16* __restrict__ z, int ldz,
                                                            const float* __restrict__ label, int B,
                                                            float* __restrict__ loss, bf16* __restrict__ dz) {
  __shared__ float red[32];
  const float invB = 1.f / (float)B;
  float acc = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const float x = __bfloat162float(z[(size_t)i * ldz]);
    const float y = label[i];
    acc += fmaxf(x, 0.f) - x * y + log1pf(__expf(-fabsf(x)));
    const float g = (1.f / (1.f + __expf(-x)) - y) * invB;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    o.x = (uint32_t)__bfloat16_as_ushort(__float2bfloat16(g));
    *reinterpret_cast<uint4*>(dz + (size_t)i * 8) = o;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) loss[0] = v * invB;
  }
}

// dh = (h > 0 ? dh : 0) in place (RELU) and dbias[n] = sum_m dh[m, n], fixed summation order.  Grid (N / 64
// column tiles, row chunks); the last block of a column tile to finish adds the chunks' partial rows.
template <bool RELU>
__global__ void __launch_bounds__(256) relu_bwd_colsum_kernel(bf16* __restrict__ dh, const bf16* __restrict__ h, int M,
                                                              int N, float* __restrict__ partial,
                                                              unsigned* __restrict__ counter, bf16* __restrict__ dbias) {
  __shared__ float red[32][65];
  __shared__ int is_last;
  const int cg = threadIdx.x & 7, rl = threadIdx.x >> 3;
  const int col0 = blockIdx.x * 64 + cg * 8;
  const bool active = col0 < N;
  const int rows_per = (M + gridDim.y - 1) / gridDim.y;
  const int r_end = min(M, (int)(blockIdx.y + 1) * rows_per);
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  if (active) {
    for (int r = blockIdx.y * rows_per + rl; r < r_end; r += 32) {
      bf16* q = dh + (size_t)r * N + col0;
      float f[8];
      unpack8(*reinterpret_cast<const uint4*>(q), f);
      if (RELU) {
        float g[8];
        unpack8(*reinterpret_cast<const uint4*>(h + (size_t)r * N + col0), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = g[j] > 0.f ? f[j] : 0.f;
        *reinterpret_cast<uint4*>(q) = pack8(f);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += f[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[rl][cg * 8 + j] = s[j];
  __syncthreads();
  if (threadIdx.x < 64) {
    float a = 0.f;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) a += red[r][threadIdx.x];
    const int col = blockIdx.x * 64 + threadIdx.x;
    if (col < N) partial[(size_t)blockIdx.y * N + col] = a;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(counter + blockIdx.x, 1u) == gridDim.y - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (threadIdx.x < 64) {
    const int col = blockIdx.x * 64 + threadIdx.x;
    if (col < N) {
      float a = 0.f;
      for (unsigned c = 0; c < gridDim.y; ++c) a += __ldcg(partial + (size_t)c * N + col);
      dbias[col] = __float2bfloat16(a);
    }
  }
  if (threadIdx.x == 0) counter[blockIdx.x] = 0u;
}

// ---- generic dispatch / combine over the sharded table (any row width) --------------------------------------------
// The two halves of an embedding all-to-all without the DIN model around them: emb_gather pulls n rows by id from
// their owners (out[i, :] = E[ids[i], :], fp32 or bf16 output), emb_scatter_add pushes scale * grad[i, :] onto row
// ids[i] with vectorised red.global.add over NVLink (duplicates coalesce in the atomics).  ids < 0 are padding:
// gather writes zeros, scatter skips.  A warp takes four rows at a time, 16 bytes per lane and row chunk.
struct GenEmbParams {
  float* shard[KDL_MAX_WORLD];
  int world, interleaved, wshift;
  int64_t rows_per_rank;
  const int64_t* ids;
  int64_t n;
  int D;                 // floats per row, multiple of 4
  void* out;             // gather: [n, D] fp32 or bf16
  int out_bf16;
  const void* grad;      // scatter: [n, D] fp32 or bf16
  int grad_bf16;
  float scale;
};

__device__ __forceinline__ float* gen_row_ptr(const GenEmbParams& p, int64_t id) {
  if (p.interleaved) {
    const uint32_t i = (uint32_t)id;
    uint32_t owner, local;
    if (p.wshift >= 0) { owner = i & (uint32_t)(p.world - 1); local = i >> p.wshift; }
    else { local = i / (uint32_t)p.world; owner = i - local * (uint32_t)p.world; }
    return p.shard[owner] + (size_t)local * p.D;
  }
  const int owner = (int)(id / p.rows_per_rank);
  return p.shard[owner] + (size_t)(id - (int64_t)owner * p.rows_per_rank) * p.D;
}

__global__ void __launch_bounds__(256) emb_gather_kernel(GenEmbParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int chunks = p.D >> 2;
  constexpr int R = 4;
  for (int64_t r0 = warp * R; r0 < p.n; r0 += nwarps * R) {
    const float* src[R];
#pragma unroll
    for (int k = 0; k < R; ++k) {
      const int64_t id = r0 + k < p.n ? p.ids[r0 + k] : -1;
      src[k] = id >= 0 ? gen_row_ptr(p, id) : nullptr;
    }
    for (int c0 = 0; c0 < chunks; c0 += 32) {
      const int c = c0 + lane;
      float4 v[R];
#pragma unroll
      for (int k = 0; k < R; ++k)          // all loads of the four rows in flight before any is used
        v[k] = (src[k] != nullptr && c < chunks) ? ld_row4(src[k], c) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int k = 0; k < R; ++k) {
        if (r0 + k >= p.n || c >= chunks) continue;
        const size_t o = (size_t)(r0 + k) * p.D + 4 * c;
        if (p.out_bf16) st_bf16x4(reinterpret_cast<bf16*>(p.out) + o, v[k].x, v[k].y, v[k].z, v[k].w);
        else *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + o) = v[k];
      }
    }
  }
}

__global__ void __launch_bounds__(256) emb_scatter_add_kernel(GenEmbParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int chunks = p.D >> 2;
  for (int64_t r = warp; r < p.n; r += nwarps) {
    const int64_t id = p.ids[r];
    if (id < 0) continue;
    float* dst = gen_row_ptr(p, id);
    for (int c = lane; c < chunks; c += 32) {
      const size_t o = (size_t)r * p.D + 4 * c;
      float4 g;
      if (p.grad_bf16) g = ld_bf16x4(reinterpret_cast<const bf16*>(p.grad) + o);
      else g = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.grad) + o);
      red_add4(dst, c, p.scale * g.x, p.scale * g.y, p.scale * g.z, p.scale * g.w);
    }
  }
}

}  // namespace kdl

using namespace kdl;

static int gen_fill(GenEmbParams* p, void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved, int D,
                    const int64_t* ids, int64_t n) {
  if (world < 1 || world > KDL_MAX_WORLD || rows_per_rank < 1 || D < 4 || (D & 3) || n < 0) {
    kdl_set_error("embedding: bad arguments world=%d rows_per_rank=%lld D=%d n=%lld (D must be a multiple of 4)", world,
                  (long long)rows_per_rank, D, (long long)n);
    return KDL_ERR_ARG;
  }
  if (interleaved && rows_per_rank * world > (int64_t)1 << 32) {
    kdl_set_error("embedding: interleaved placement supports vocabularies below 2^32 rows");
    return KDL_ERR_UNSUPPORTED;
  }
  for (int r = 0; r < world; ++r) p->shard[r] = (float*)shard_ptrs[r];
  p->world = world; p->interleaved = interleaved ? 1 : 0; p->rows_per_rank = rows_per_rank;
  p->wshift = -1;
  for (int b = 0; b < 8; ++b)
    if ((1 << b) == world) p->wshift = b;
  p->ids = ids; p->n = n; p->D = D;
  return KDL_OK;
}

static int gen_grid(int64_t n, int rows_per_warp) {
  int64_t g = (n + 8 * rows_per_warp - 1) / (8 * rows_per_warp);
  if (g > 148 * 8) g = 148 * 8;
  return (int)(g < 1 ? 1 : g);
}

extern "C" int kdl_emb_gather(void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved, int D,
                              const int64_t* ids, int64_t n, void* out, int out_bf16, cudaStream_t stream) {
  GenEmbParams p = {};
  int rc = gen_fill(&p, shard_ptrs, world, rows_per_rank, interleaved, D, ids, n);
  if (rc) return rc;
  if (n == 0) return KDL_OK;
  p.out = out; p.out_bf16 = out_bf16;
  emb_gather_kernel<<<gen_grid(n, 4), 256, 0, stream>>>(p);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_emb_scatter_add(void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved, int D,
                                   const int64_t* ids, int64_t n, const void* grad, int grad_bf16, float scale,
                                   cudaStream_t stream) {
  GenEmbParams p = {};
  int rc = gen_fill(&p, shard_ptrs, world, rows_per_rank, interleaved, D, ids, n);
  if (rc) return rc;
  if (n == 0) return KDL_OK;
  p.grad = grad; p.grad_bf16 = grad_bf16; p.scale = scale;
  emb_scatter_add_kernel<<<gen_grid(n, 1), 256, 0, stream>>>(p);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_din_bce_grad(const void* z_bf16, int ldz, const float* label, int B, float* loss, void* dz_bf16,
                                cudaStream_t stream) {
  if (B < 1 || ldz < 1) {
    kdl_set_error("din_bce_grad: bad arguments B=%d ldz=%d", B, ldz);
    return KDL_ERR_ARG;
  }
  din_bce_grad_kernel<<<1, 1024, 0, stream>>>((const bf16*)z_bf16, ldz, label, B, loss, (bf16*)dz_bf16);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" size_t kdl_relu_bwd_colsum_workspace_bytes(int N) { return 256 + (size_t)32 * N * sizeof(float); }

// workspace: kdl_relu_bwd_colsum_workspace_bytes(N) bytes, zeroed ONCE (the counters reset themselves).
extern "C" int kdl_relu_bwd_colsum(void* dh_bf16, const void* h_bf16, int M, int N, void* dbias_bf16, void* workspace,
                                   size_t workspace_bytes, cudaStream_t stream) {
  if (M < 1 || N < 8 || (N & 7) || N > 64 * 64) {
    kdl_set_error("relu_bwd_colsum: need N %% 8 == 0 and N <= 4096 (M=%d N=%d)", M, N);
    return KDL_ERR_ARG;
  }
  if (workspace == nullptr || workspace_bytes < kdl_relu_bwd_colsum_workspace_bytes(N)) {
    kdl_set_error("relu_bwd_colsum: workspace too small");
    return KDL_ERR_WORKSPACE;
  }
  int chunks = (M + 255) / 256;
  if (chunks > 32) chunks = 32;
  if (chunks < 1) chunks = 1;
  const dim3 grid((N + 63) / 64, chunks);
  unsigned* counter = (unsigned*)workspace;
  float* partial = (float*)((char*)workspace + 256);
  if (h_bf16 != nullptr)
    relu_bwd_colsum_kernel<true><<<grid, 256, 0, stream>>>((bf16*)dh_bf16, (const bf16*)h_bf16, M, N, partial, counter,
                                                           (bf16*)dbias_bf16);
  else
    relu_bwd_colsum_kernel<false><<<grid, 256, 0, stream>>>((bf16*)dh_bf16, nullptr, M, N, partial, counter,
                                                            (bf16*)dbias_bf16);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

static int fill(EmbParams* p, void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved,
                const int64_t* user, const int64_t* cand, const int64_t* hist, int B, int L, int D) {
  if (D != kD) {
    kdl_set_error("embedding kernels are specialised for D=%d (got %d)", kD, D);
    return KDL_ERR_UNSUPPORTED;
  }
  if (world < 1 || world > KDL_MAX_WORLD || B < 1 || L < 0 || rows_per_rank < 1) {
    kdl_set_error("embedding: bad arguments world=%d B=%d L=%d rows_per_rank=%lld", world, B, L, (long long)rows_per_rank);
    return KDL_ERR_ARG;
  }
  if (interleaved && rows_per_rank * world > (int64_t)1 << 32) {
    kdl_set_error("embedding: interleaved placement supports vocabularies below 2^32 rows");
    return KDL_ERR_UNSUPPORTED;
  }
  for (int r = 0; r < world; ++r) p->shard[r] = (float*)shard_ptrs[r];
  p->world = world; p->rows_per_rank = rows_per_rank;
  p->interleaved = interleaved ? 1 : 0;
  p->wshift = -1;
  for (int b = 0; b < 8; ++b)
    if ((1 << b) == world) p->wshift = b;
  p->user = user; p->cand = cand; p->hist = hist; p->B = B; p->L = L;
  return KDL_OK;
}

extern "C" int kdl_din_gather_attend(void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved,
                                     const int64_t* user, const int64_t* cand, const int64_t* hist, int B,
                                     int L, int D, void* x_bf16, void* hsave_bf16, float* asave,
                                     cudaStream_t stream) {
  EmbParams p = {};
  int rc = fill(&p, shard_ptrs, world, rows_per_rank, interleaved, user, cand, hist, B, L, D);
  if (rc) return rc;
  p.x = (bf16*)x_bf16; p.hsave = (bf16*)hsave_bf16; p.asave = asave;
  static int use_tma = -1;      // KDL_DIN_GATHER=ldg keeps the register version (A/B runs)
  if (use_tma < 0) {
    const char* e = getenv("KDL_DIN_GATHER");
    use_tma = (e && e[0] == 'l') ? 0 : 1;
  }
  if (use_tma && L + 2 <= kTmaMaxRows) {
    const size_t smem = (size_t)kTmaWarps * kTmaMaxRows * kD * 4 + kTmaWarps * 8;
    static bool attr_done = false;
    if (!attr_done) {
      KDL_CUDA(cudaFuncSetAttribute(din_gather_attend_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_done = true;
    }
    int grid = (B + kTmaWarps - 1) / kTmaWarps;
    if (grid > 2 * 148) grid = 2 * 148;           // persistent: two blocks per SM, warps stride over the samples
    din_gather_attend_tma_kernel<<<grid, kTmaWarps * 32, smem, stream>>>(p);
  } else {
    din_gather_attend_kernel<<<(B + kWarpsPerBlock - 1) / kWarpsPerBlock, kWarpsPerBlock * 32, 0, stream>>>(p);
  }
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_din_scatter_update(void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved,
                                      const int64_t* user, const int64_t* cand, const int64_t* hist, int B,
                                      int L, int D, const void* x_bf16, const void* dx_bf16,
                                      const void* hsave_bf16, const float* asave, float lr,
                                      cudaStream_t stream) {
  EmbParams p = {};
  int rc = fill(&p, shard_ptrs, world, rows_per_rank, interleaved, user, cand, hist, B, L, D);
  if (rc) return rc;
  p.x = (bf16*)x_bf16; p.dx = (const bf16*)dx_bf16; p.hsave = (bf16*)hsave_bf16; p.asave = const_cast<float*>(asave); p.lr = lr;
  din_scatter_update_kernel<<<(B + kWarpsPerBlock - 1) / kWarpsPerBlock, kWarpsPerBlock * 32, 0, stream>>>(p);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}
