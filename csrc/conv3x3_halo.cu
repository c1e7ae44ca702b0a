// kubedl_b200 — 3x3 / stride-1 / pad-1 convolution with the input halo tile resident in shared
// memory (kernel N7b; fprop + dgrad here, wgrad below).
//
// Why a second convolution kernel: the TMA-im2col implicit GEMM in gemm_tcgen05.cu fetches one
// 128-pixel x 64-channel operand tile per filter tap, i.e. every input pixel crosses L2 -> shared
// memory nine times, and the im2col mode of the TMA unit delivers only ~1 pixel row per 7 clocks
// (measured: ~870 clk per 128-pixel load, independent of the tile shape;
// profiles/conv3x3_vs_cudnn_2cta.log).  That caps the narrow layers (C = 64, 128) of ResNet-50 at
// a third of cuDNN's speed.  Here a tile is TR whole output rows of one image:
//
//   * ONE tiled TMA box load per 64-channel chunk brings the (TR+2) x (W+2) input halo (out-of-
//     bounds rows / columns zero-filled by the TMA unit = the padding) into shared memory as
//     rows of 128 B, row index  i*(W+2) + j.
//   * Output pixel (oh, ow) of the tile is accumulator row  m = oh*(W+2) + ow, and filter tap (r, s)
//     reads halo row  m + r*(W+2) + s.  So all nine taps are the SAME buffer seen through UMMA
//     shared-memory descriptors whose start address is shifted by (r*(W+2) + s) * 128 B — the
//     128-byte swizzle is a function of the absolute shared-memory address, so a row-shifted
//     window stays consistent with what TMA wrote.  No data is moved or duplicated per tap.
//   * The weights of the CTA's 64 output channels (9 x C x 64 bf16, 72 KB per 64 input channels)
//     are loaded once and stay resident for the whole persistent loop.
//   * Accumulator rows with ow >= W (two per image row) and m >= TR*(W+2) are junk; the epilogue
//     compacts the valid rows into a swizzled staging tile and one TMA store writes the TR*W
//     contiguous output pixels.
//
// Shared-memory fill per tile drops from 9 x 16 KB (+ weights) to ~30 KB per 64 channels, which
// makes the kernel tensor-pipe / HBM bound instead of TMA bound.  dgrad is the same kernel on dy
// with the taps mirrored and the weights read MN-major (no transposed copy).
//
// The reference operator has no convolution at all; this replaces the cuDNN kernels the stock
// ResNet-50 DDP worker image would run (reference example/pytorch/pytorch_job_mnist_mpi.yaml:15
// only names a container image).
#include "common.cuh"
#include "tcgen05.cuh"
#include "../include/kubedl_b200.h"
#include <cuda.h>
#include <stdlib.h>

namespace kdl {

constexpr int kHaloThreads = 256;     // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps 4-7 epilogue
constexpr int kHaloBN = 64;           // output channels per CTA
constexpr int kWSlotBytes = 64 * 128; // one (tap, 64-channel chunk) weight block

struct HaloParams {
  int n_img, H, W, Wp, TR;
  int tiles_per_img, tiles;
  int n_parts;       // output-channel parts of 64
  int chunks;        // 64-channel chunks of the patch operand
  int c_patch;       // channels of the patch operand
  int c_out;         // output channels
  int flip;          // dgrad: mirrored taps
  int nbuf;          // A ring depth
  int a_buf_bytes;   // bytes reserved per A buffer (multiple of 1024)
  int a_box_bytes;   // bytes one halo box delivers
  int stg_bytes;     // bytes per staging buffer (multiple of 1024)
  int relu;
  const bf16* addend;   // optional [pixels][c_out], added in the epilogue
};

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0,
                                            int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

template <int A_MN, int B_MN, int M, int N>
__host__ __device__ constexpr uint32_t make_idesc_mn() {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)A_MN << 15) | ((uint32_t)B_MN << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// descriptor with explicit leading-dimension byte offset (distance between 64-wide MN atoms)
template <int MN_MAJOR>
__device__ __forceinline__ uint64_t make_desc_lbo(uint32_t smem_addr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((1024 >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__device__ __forceinline__ void epi_bar_sync() {   // the 4 epilogue warps only
  asm volatile("bar.sync 1, 128;" ::: "memory");
}

template <int B_MN>
__global__ void __launch_bounds__(kHaloThreads, 1)
conv3x3_halo_kernel(const __grid_constant__ CUtensorMap tma_a,   // patch operand, 4-D {C, W, H, N}
                    const __grid_constant__ CUtensorMap tma_w,   // weights [Cout][9*Cin]
                    const __grid_constant__ CUtensorMap tma_c,   // output [pixels][c_out], box {64, TR*W}
                    const HaloParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const int w_bytes = 9 * p.chunks * kWSlotBytes;
  const uint32_t a_base = smem_base + w_bytes;
  const uint32_t stg_base = a_base + p.nbuf * p.a_buf_bytes;
  const uint32_t bar_base = stg_base + 2 * p.stg_bytes;
  uint8_t* stg_ptr = smem + w_bytes + p.nbuf * p.a_buf_bytes;
  auto wfull_bar = [&]() { return bar_base; };
  auto afull_bar = [&](int s) { return bar_base + 8u * (1 + s); };
  auto aempty_bar = [&](int s) { return bar_base + 8u * (1 + p.nbuf + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (1 + 2 * p.nbuf + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (3 + 2 * p.nbuf + s); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + (bar_base - smem_base) + 8 * (5 + 2 * p.nbuf));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 2 * kHaloBN;

  // the A ring is read beyond the rows a box delivers (junk accumulator rows): keep it finite
  {
    uint4* z = reinterpret_cast<uint4*>(smem + w_bytes);
    const int n16 = p.nbuf * p.a_buf_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += kHaloThreads) z[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_c) : "memory");
  }
  if (warp == 1 && lane == 0) {
    mbar_init(wfull_bar(), 1);
    for (int s = 0; s < p.nbuf; ++s) {
      mbar_init(afull_bar(s), 1);
      mbar_init(aempty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_ptr_smem)),
                 "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  const int part = blockIdx.x % p.n_parts;
  const int n0 = part * kHaloBN;
  const int t_first = blockIdx.x / p.n_parts;
  const int t_step = gridDim.x / p.n_parts;

  if (warp == 0) {
    // ============================== TMA producer ==============================
    if (lane == 0) {
      mbar_expect_tx(wfull_bar(), (uint32_t)w_bytes);
      for (int tap = 0; tap < 9; ++tap) {
        const int gt = p.flip ? 8 - tap : tap;
        for (int ch = 0; ch < p.chunks; ++ch) {
          const uint32_t dst = smem_base + (tap * p.chunks + ch) * kWSlotBytes;
          if (B_MN) tma_load_2d(dst, &tma_w, wfull_bar(), gt * p.c_out + n0, ch * 64);   // rows = patch channel
          else      tma_load_2d(dst, &tma_w, wfull_bar(), gt * p.c_patch + ch * 64, n0); // rows = output channel
        }
      }
      int buf = 0;
      uint32_t phase = 0;
      for (int t = t_first; t < p.tiles; t += t_step) {
        const int img = t / p.tiles_per_img;
        const int oh0 = (t - img * p.tiles_per_img) * p.TR;
        for (int ch = 0; ch < p.chunks; ++ch) {
          mbar_wait(aempty_bar(buf), phase ^ 1);
          mbar_expect_tx(afull_bar(buf), (uint32_t)p.a_box_bytes);
          tma_load_4d(a_base + buf * p.a_buf_bytes, &tma_a, afull_bar(buf), ch * 64, -1, oh0 - 1, img);
          if (++buf == p.nbuf) { buf = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer ==============================
    // The whole warp walks the loop (waits included); one elected lane issues.
    {
      const uint32_t leader = elect_one();
      constexpr uint32_t idesc = make_idesc_mn<0, B_MN, 128, kHaloBN>();
      mbar_wait(wfull_bar(), 0);
      tcgen05_fence_after();
      const int wp16 = p.Wp * 8;                          // one halo row of 128 B, in 16-byte units
      const int wslot16 = p.chunks * (kWSlotBytes >> 4);  // weight slots of consecutive taps
      int buf = 0;
      uint32_t phase = 0;
      int iter = 0;
      for (int t = t_first; t < p.tiles; t += t_step, ++iter) {
        const int as = iter & 1;
        const uint32_t aphase = (iter >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + as * kHaloBN;
        for (int ch = 0; ch < p.chunks; ++ch) {
          mbar_wait(afull_bar(buf), phase);
          tcgen05_fence_after();
          // 36 MMAs of only 32 tensor-pipe clocks each per buffer: the issue path must be a couple of
          // integer adds.  Descriptors are built once per buffer and advanced in their 16-byte-unit
          // address field (taps, K slices and weight slots are compile-time offsets).
          const uint64_t ad0 = make_desc<0>(a_base + buf * p.a_buf_bytes);
          const uint64_t bd0 = make_desc<B_MN>(smem_base + ch * kWSlotBytes);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const int r = tap / 3, s = tap - r * 3;
            const uint64_t ad = ad0 + (uint64_t)(r * wp16 + s * 8);
            const uint64_t bd = bd0 + (uint64_t)(tap * wslot16);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              umma_bf16_pred(tmem_d, ad + (uint64_t)(k * (int)(k_advance_bytes<0>(1) >> 4)),
                             bd + (uint64_t)(k * (int)(k_advance_bytes<B_MN>(1) >> 4)), idesc,
                             (tap > 0 || k > 0) ? 1u : (ch > 0 ? 1u : 0u), leader);
            }
          }
          umma_commit_pred(aempty_bar(buf), leader);
          if (ch == p.chunks - 1) umma_commit_pred(tfull_bar(as), leader);
          if (++buf == p.nbuf) { buf = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ============================== epilogue ==============================
    const int q = warp & 3;
    const int m = q * 32 + lane;                 // accumulator row = TMEM lane
    const int rr = m / p.Wp, cc = m - rr * p.Wp;
    const bool valid = rr < p.TR && cc < p.W;
    const int srow = rr * p.W + cc;              // compacted row inside the output tile
    const bool leader = (warp == 4 && lane == 0);
    int iter = 0;
    for (int t = t_first; t < p.tiles; t += t_step, ++iter) {
      const int as = iter & 1;
      const uint32_t aphase = (iter >> 1) & 1;
      const int img = t / p.tiles_per_img;
      const int oh0 = (t - img * p.tiles_per_img) * p.TR;
      const int pix0 = (img * p.H + oh0) * p.W;
      mbar_wait(tfull_bar(as), aphase);
      tcgen05_fence_after();
      uint32_t r[64];
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * kHaloBN;
      tmem_ld32(taddr, r);
      tmem_ld32(taddr + 32, r + 32);
      tmem_ld_wait();
      float f[64];
#pragma unroll
      for (int j = 0; j < 64; ++j) f[j] = __uint_as_float(r[j]);
      if (p.addend && valid) {
        const bf16* ad = p.addend + (size_t)(pix0 + srow) * p.c_out + n0;
#pragma unroll
        for (int j = 0; j < 64; j += 8) {
          float b[8];
          unpack8(ld_plain(ad + j), b);
#pragma unroll
          for (int i = 0; i < 8; ++i) f[j + i] += b[i];
        }
      }
      if (p.relu) {
#pragma unroll
        for (int j = 0; j < 64; ++j) f[j] = fmaxf(f[j], 0.f);
      }
      // staging buffer `as` was handed to the TMA store two tiles ago
      if (leader) tma_store_wait_read<1>();
      epi_bar_sync();
      if (valid) {
        uint8_t* dst = stg_ptr + as * p.stg_bytes + srow * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(dst + ((j ^ (srow & 7)) << 4)) = pack8(f + 8 * j);
      }
      tcgen05_fence_before();
      fence_proxy_async_smem();
      epi_bar_sync();
      if (leader) {
        tma_store_2d(&tma_c, stg_base + as * p.stg_bytes, n0, pix0);
        tma_store_commit();
        mbar_arrive(tempty_bar(as));
      }
    }
    if (leader) tma_store_wait_read<0>();
    __syncwarp();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols)
                 : "memory");
  }
}

// ------------------------------------------------------------------ 2-CTA fprop / dgrad
// ncu on the 1-CTA kernel (profiles/ncu_conv3x3_halo_raw.csv): 48 tensor-core shared-memory wavefronts
// per 128x64x16 MMA (32 for A, 16 for B) at 76 % of the pipe's peak, tensor pipe 50 % active — operand
// reads from shared memory are the limit, and A is re-read for every 64 output channels.  With
// cta_group::2 a CTA pair issues ONE M = 256 MMA over both CTAs' pixel tiles and all N = Cout (<= 128)
// output channels; each CTA holds only its half of the weight rows, so the same 48 wavefronts per
// CTA now feed four times the math (N = 128) and the loop becomes tensor-pipe bound.
// Weights are K-major here in both directions: dgrad gets a flipped + transposed copy of the filter
// (halo_flip_transpose_kernel, 74-295 KB) instead of MN-major descriptors, because half of a
// 64-channel N tile is only 64 B wide.
struct Halo2Params {
  int n_img, H, W, Wp, TR;
  int tiles_per_img, tiles, pair_tiles;
  int chunks, c_patch, c_out;   // c_out = total output channels
  int n_parts;                  // c_out / NB: pairs alternate over the output-channel parts
  int nbuf, a_buf_bytes, a_box_bytes, stg_bytes;
  int relu;
  const bf16* addend;
  bf16* out;                    // NB = 128 stores rows straight from registers (no staging)
};

__device__ __forceinline__ void tma_load_4d_2sm(uint32_t dst, const CUtensorMap* map, uint32_t leader_bar,
                                                int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

template <int NB>   // output channels of the pair (64 or 128)
__global__ void __launch_bounds__(kHaloThreads, 1)
conv3x3_halo2_kernel(const __grid_constant__ CUtensorMap tma_a,   // patch operand {C, W, H, N}
                     const __grid_constant__ CUtensorMap tma_w,   // weights [NB rows][9*C], box {64, NB/2}
                     const __grid_constant__ CUtensorMap tma_c,   // output [pixels][NB], box {64, TR*W}
                     const Halo2Params p) {
  constexpr int kSlot = (NB / 2) * 128;      // this CTA's rows of one (tap, chunk) weight block
  constexpr int kNChunks = NB / 64;          // 64-column chunks of the output tile
  // NB = 128: the CTA's weights take 144 KB at 128 input channels, and a third halo buffer (needed to
  // cover the TMA latency of the next tile) is worth more than a staging tile — every thread owns one
  // output pixel = 256 contiguous bytes, which it stores directly.
  constexpr bool kDirect = NB == 128;
  constexpr uint32_t kPeerMask = 0xFEFFFFFFu;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const int w_bytes = 9 * p.chunks * kSlot;
  const uint32_t a_base = smem_base + w_bytes;
  const uint32_t stg_base = a_base + p.nbuf * p.a_buf_bytes;
  const uint32_t bar_base = stg_base + (kDirect ? 0 : kNChunks * p.stg_bytes);
  uint8_t* stg_ptr = smem + w_bytes + p.nbuf * p.a_buf_bytes;
  auto wfull_bar = [&](int tap) { return bar_base + 8u * tap; };       // one per filter tap: MMAs start
  auto afull_bar = [&](int s) { return bar_base + 8u * (9 + s); };      // as soon as tap 0 has landed
  auto aempty_bar = [&](int s) { return bar_base + 8u * (9 + p.nbuf + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (9 + 2 * p.nbuf + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (11 + 2 * p.nbuf + s); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + (bar_base - smem_base) + 8 * (13 + 2 * p.nbuf));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader_cta = cta == 0;
  constexpr uint32_t kTmemCols = 2 * NB;

  {
    uint4* z = reinterpret_cast<uint4*>(smem + w_bytes);
    const int n16 = p.nbuf * p.a_buf_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += kHaloThreads) z[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_c) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int tap = 0; tap < 9; ++tap) mbar_init(wfull_bar(tap), 2);
    for (int s = 0; s < p.nbuf; ++s) {
      mbar_init(afull_bar(s), 2);     // one arrive per CTA's producer; bytes of both land on the leader's
      mbar_init(aempty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 2);    // the epilogue of each CTA
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_ptr_smem)),
                 "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  cluster_sync_all();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  const int part = (blockIdx.x / 2) % p.n_parts;          // which NB output channels this pair owns
  const int num_pairs = (gridDim.x / 2) / p.n_parts;      // pairs walking the pixel tiles of one part
  const int pair = (blockIdx.x / 2) / p.n_parts;
  const int n0 = part * NB;
  auto my_tile = [&](int pt) { return min(2 * pt + (int)cta, p.tiles - 1); };

  if (warp == 0) {
    // ============================== TMA producer (both CTAs) ==============================
    if (lane == 0) {
      int buf = 0;
      uint32_t phase = 0;
      bool weights_pending = true;
      for (int pt = pair; pt < p.pair_tiles; pt += num_pairs) {
        const int t = my_tile(pt);
        const int img = t / p.tiles_per_img;
        const int oh0 = (t - img * p.tiles_per_img) * p.TR;
        for (int ch = 0; ch < p.chunks; ++ch) {
          mbar_wait(aempty_bar(buf), phase ^ 1);
          tma_load_4d_2sm(a_base + buf * p.a_buf_bytes, &tma_a, afull_bar(buf) & kPeerMask, ch * 64, -1, oh0 - 1, img);
          if (leader_cta) mbar_expect_tx(afull_bar(buf), 2u * (uint32_t)p.a_box_bytes);
          else mbar_arrive_remote(afull_bar(buf), 0);
          if (++buf == p.nbuf) { buf = 0; phase ^= 1; }
          if (weights_pending) {
            // the resident weights go out right behind the first halo box, tap by tap
            weights_pending = false;
            for (int tap = 0; tap < 9; ++tap) {
              const uint32_t lbar = wfull_bar(tap) & kPeerMask;
              for (int c2 = 0; c2 < p.chunks; ++c2)
                tma_load_2d_2sm(smem_base + (tap * p.chunks + c2) * kSlot, &tma_w, lbar, tap * p.c_patch + c2 * 64,
                                n0 + (int)cta * (NB / 2));
              if (leader_cta) mbar_expect_tx(wfull_bar(tap), 2u * (uint32_t)(p.chunks * kSlot));
              else mbar_arrive_remote(wfull_bar(tap), 0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer (leader CTA) ==============================
    if (leader_cta) {
      const uint32_t el = elect_one();
      constexpr uint32_t idesc = make_idesc_mn<0, 0, 256, NB>();
      bool first = true;
      const int wp16 = p.Wp * 8;
      const int wslot16 = p.chunks * (kSlot >> 4);
      int buf = 0;
      uint32_t phase = 0;
      int iter = 0;
      for (int pt = pair; pt < p.pair_tiles; pt += num_pairs, ++iter) {
        const int as = iter & 1;
        const uint32_t aphase = (iter >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + as * NB;
        for (int ch = 0; ch < p.chunks; ++ch) {
          mbar_wait(afull_bar(buf), phase);
          tcgen05_fence_after();
          const uint64_t ad0 = make_desc<0>(a_base + buf * p.a_buf_bytes);
          const uint64_t bd0 = make_desc<0>(smem_base + ch * kSlot);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            if (first) {
              mbar_wait(wfull_bar(tap), 0);
              tcgen05_fence_after();
            }
            const int r = tap / 3, s = tap - r * 3;
            const uint64_t ad = ad0 + (uint64_t)(r * wp16 + s * 8);
            const uint64_t bd = bd0 + (uint64_t)(tap * wslot16);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_bf16_2sm_pred(tmem_d, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc,
                                 (tap > 0 || k > 0) ? 1u : (ch > 0 ? 1u : 0u), el);
          }
          first = false;
          umma_commit_2sm_pred(aempty_bar(buf), el);
          if (ch == p.chunks - 1) umma_commit_2sm_pred(tfull_bar(as), el);
          if (++buf == p.nbuf) { buf = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ============================== epilogue (each CTA: its own 128 accumulator rows) ==============================
    const int q = warp & 3;
    const int m = q * 32 + lane;
    const int rr = m / p.Wp, cc = m - rr * p.Wp;
    const bool valid = rr < p.TR && cc < p.W;
    const int srow = rr * p.W + cc;
    const bool leader = (warp == 4 && lane == 0);
    int iter = 0;
    for (int pt = pair; pt < p.pair_tiles; pt += num_pairs, ++iter) {
      const int as = iter & 1;
      const uint32_t aphase = (iter >> 1) & 1;
      const int t = my_tile(pt);
      const int img = t / p.tiles_per_img;
      const int oh0 = (t - img * p.tiles_per_img) * p.TR;
      const int pix0 = (img * p.H + oh0) * p.W;
      mbar_wait(tfull_bar(as), aphase);
      tcgen05_fence_after();
      if (!kDirect) {
        if (leader) tma_store_wait_read<0>();    // single staging set: the previous tile's stores have read it
        epi_bar_sync();
      }
#pragma unroll 1
      for (int c = 0; c < kNChunks; ++c) {
        uint32_t r[64];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * NB + c * 64;
        tmem_ld32(taddr, r);
        tmem_ld32(taddr + 32, r + 32);
        tmem_ld_wait();
        float f[64];
#pragma unroll
        for (int j = 0; j < 64; ++j) f[j] = __uint_as_float(r[j]);
        if (p.addend && valid) {
          const bf16* ad = p.addend + (size_t)(pix0 + srow) * p.c_out + n0 + c * 64;
#pragma unroll
          for (int j = 0; j < 64; j += 8) {
            float b[8];
            unpack8(ld_plain(ad + j), b);
#pragma unroll
            for (int i = 0; i < 8; ++i) f[j + i] += b[i];
          }
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 64; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        if (valid) {
          if (kDirect) {
            uint4* dst = reinterpret_cast<uint4*>(p.out + (size_t)(pix0 + srow) * p.c_out + n0 + c * 64);
#pragma unroll
            for (int j = 0; j < 8; ++j) dst[j] = pack8(f + 8 * j);
          } else {
            uint8_t* dst = stg_ptr + c * p.stg_bytes + srow * 128;
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *reinterpret_cast<uint4*>(dst + ((j ^ (srow & 7)) << 4)) = pack8(f + 8 * j);
          }
        }
      }
      tcgen05_fence_before();
      if (!kDirect) fence_proxy_async_smem();
      epi_bar_sync();
      if (leader) {
        if (!kDirect) {
          for (int c = 0; c < kNChunks; ++c) tma_store_2d(&tma_c, stg_base + c * p.stg_bytes, n0 + c * 64, pix0);
          tma_store_commit();
        }
        mbar_arrive_remote(tempty_bar(as), 0);
      }
    }
    if (leader) tma_store_wait_read<0>();
    __syncwarp();
  }

  tcgen05_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols)
                 : "memory");
  }
}

// wt[ci][2-r][2-s][co] = w[co][r][s][ci]: dgrad as an fprop over dy with K-major weights
__global__ void halo_flip_transpose_kernel(const bf16* __restrict__ w, bf16* __restrict__ wt, int Cout, int Cin) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;   // index into wt
  if (i >= Cout * 9 * Cin) return;
  const int co = i % Cout;
  const int tap = (i / Cout) % 9;
  const int ci = i / (9 * Cout);
  wt[i] = w[((size_t)co * 9 + (8 - tap)) * Cin + ci];
}

// ------------------------------------------------------------------ wgrad
// dW[co][r][s][ci] = sum over pixels of dy[p][co] * x[p + (r-1, s-1)][ci].  The pixel index is the
// MMA K dimension, so both operands are MN-major views of NHWC tiles (as in the tn GEMM), and a
// CTA keeps its accumulators in TMEM across ALL of its pixel tiles:
//   * CTA role r in {0,1,2} = filter row: D[co][(s, ci)] for s = 0..2 is 3*Cin <= 384 TMEM columns.
//   * per tile: dy rows are loaded with a (W+2)-wide box, so the two columns past the image edge are
//     zero-filled by TMA and dy row index == halo row index; the x halo of filter row r is TR rows
//     starting one column left of the image; filter column s is the x buffer shifted by s rows.
//   * K = 128 halo rows per tile (rows past TR*(W+2) of the dy buffer stay zero for the whole kernel).
// At the end each CTA writes its fp32 partial slice and a second kernel reduces over the CTAs of
// a role in a fixed order (deterministic) and casts to bf16.
struct HaloWgradParams {
  int n_img, H, W, Wp, TR;
  int tiles_per_img, tiles;
  int cich, coch;         // 64-channel chunks of Cin / Cout one CTA accumulates (1 or 2 each)
  int ci_roles, co_roles; // channel blocks of 64*cich / 64*coch: role = (filter row, ci block, co block)
  int Cin, Cout;
  int NI;                 // images per tile (small feature maps: 2 x 7x7 fill the 128 K-rows)
  int G;                  // CTAs per role
  int stages;
  int box_bytes;          // bytes one (chunk) box delivers
  float* partial;         // [G][Cout][9*Cin]
};

constexpr int kDyBufBytes = 128 * 128;   // 128 K-rows x 64 output channels
constexpr int kXBufBytes = 136 * 128;    // 128 K-rows + 2 shifts, rounded to 8 rows

__global__ void __launch_bounds__(kHaloThreads, 1)
conv3x3_halo_wgrad_kernel(const __grid_constant__ CUtensorMap tma_dy,   // {Cout, W, H, N}, box {64, Wp, TR, 1}
                          const __grid_constant__ CUtensorMap tma_x,    // {Cin, W, H, N}, same box
                          const HaloWgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const int stage_bytes = p.coch * kDyBufBytes + p.cich * kXBufBytes;
  const uint32_t bar_base = smem_base + p.stages * stage_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (p.stages + s); };
  auto done_bar = [&]() { return bar_base + 8u * (2 * p.stages); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + p.stages * stage_bytes + 8 * (2 * p.stages + 1));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int ncols = 3 * 64 * p.cich;
  const uint32_t tmem_cols = ncols <= 256 ? 256u : 512u;

  {
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = p.stages * stage_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += kHaloThreads) z[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_dy) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_x) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(done_bar(), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_ptr_smem)),
                 "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  const int nroles = 3 * p.ci_roles * p.co_roles;
  const int role_id = blockIdx.x % nroles;
  const int role = role_id % 3;                       // filter row
  const int ci0 = ((role_id / 3) % p.ci_roles) * 64 * p.cich;
  const int co0 = (role_id / (3 * p.ci_roles)) * 64 * p.coch;
  const int g = blockIdx.x / nroles;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = g; t < p.tiles; t += p.G) {
        const int ig = t / p.tiles_per_img;
        const int img = ig * p.NI;
        const int oh0 = (t - ig * p.tiles_per_img) * p.TR;
        mbar_wait(empty_bar(stage), phase ^ 1);
        mbar_expect_tx(full_bar(stage), (uint32_t)((p.coch + p.cich) * p.box_bytes));
        const uint32_t sdy = smem_base + stage * stage_bytes;
        const uint32_t sx = sdy + p.coch * kDyBufBytes;
        for (int c = 0; c < p.coch; ++c)
          tma_load_4d(sdy + c * kDyBufBytes, &tma_dy, full_bar(stage), co0 + c * 64, 0, oh0, img);
        for (int c = 0; c < p.cich; ++c)
          tma_load_4d(sx + c * kXBufBytes, &tma_x, full_bar(stage), ci0 + c * 64, -1, oh0 + role - 1, img);
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    const uint32_t el = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    bool first = true;
    for (int t = g; t < p.tiles; t += p.G) {
      mbar_wait(full_bar(stage), phase);
      tcgen05_fence_after();
      const uint32_t sdy = smem_base + stage * stage_bytes;
      const uint32_t sx = sdy + p.coch * kDyBufBytes;
      const uint64_t ad0 = make_desc_lbo<1>(sdy, (uint32_t)kDyBufBytes);
      const uint64_t bd0 = make_desc_lbo<1>(sx, (uint32_t)kXBufBytes);
      // 64 output channels: an M = 64 MMA (accumulator row m lives in TMEM lane 32*(m/16) + m%16) reads
      // half the A bytes of an M = 128 one, and operand reads from shared memory bound these shapes.
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                             ((uint32_t)((64 * p.cich) >> 3) << 17) | ((uint32_t)((64 * p.coch) >> 4) << 24);
      const uint32_t cw = 64u * p.cich;
#pragma unroll
      for (int s = 0; s < 3; ++s)
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_bf16_pred(tmem_base + s * cw, ad0 + (uint64_t)(k * 128), bd0 + (uint64_t)(s * 8 + k * 128), idesc,
                         (k > 0 || !first) ? 1u : 0u, el);
      first = false;
      umma_commit_pred(empty_bar(stage), el);
      if (++stage == p.stages) { stage = 0; phase ^= 1; }
    }
    umma_commit_pred(done_bar(), el);
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int cw = 64 * p.cich;       // accumulator columns per filter tap
    // M = 128: lane = row.  M = 64: each 32-lane quarter holds 16 rows in its first 16 lanes.
    const int co = co0 + (p.coch == 2 ? q * 32 + lane : q * 16 + lane);
    const bool row_ok = p.coch == 2 || lane < 16;
    mbar_wait(done_bar(), 0);
    tcgen05_fence_after();
    float* dst = p.partial + (((size_t)g * p.Cout + co) * 9 + role * 3) * p.Cin + ci0;
    for (int s3 = 0; s3 < 3; ++s3) {
      for (int c0 = 0; c0 < cw; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + s3 * cw + c0, r);
        tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(dst + (size_t)s3 * p.Cin + c0 + 4 * j) =
                make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        }
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols)
                 : "memory");
  }
}

__global__ void halo_wgrad_reduce_kernel(const float* __restrict__ partial, bf16* __restrict__ out, int n, int G) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  int gI = 0;
  for (; gI + 4 <= G; gI += 4) {      // four independent 16-byte loads in flight, fixed summation order
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const float4*>(partial + (size_t)(gI + u) * n + i);
#pragma unroll
    for (int u = 0; u < 4; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
  }
  for (; gI < G; ++gI) {
    const float4 v = *reinterpret_cast<const float4*>(partial + (size_t)gI * n + i);
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  __nv_bfloat162 lo = __floats2bfloat162_rn(acc.x, acc.y), hi = __floats2bfloat162_rn(acc.z, acc.w);
  uint2 o;
  o.x = *reinterpret_cast<uint32_t*>(&lo);
  o.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(out + i) = o;
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn halo_encode() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)ptr;
    else
      cudaGetLastError();
  }
  return fn;
}

static int encode(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims,
                  const cuuint64_t* strides, const cuuint32_t* box, const char* what) {
  EncodeTiledFn enc = halo_encode();
  if (!enc) {
    kdl_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return KDL_ERR_UNSUPPORTED;
  }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    kdl_set_error("conv3x3 halo: cuTensorMapEncodeTiled(%s) failed (%d)", what, (int)r);
    return KDL_ERR_CUDA;
  }
  return KDL_OK;
}

static int halo_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// rows of one image per tile: the largest divisor of H with TR*(W+2) <= 128
static int pick_tr(int H, int W) {
  int best = 0;
  for (int tr = 1; tr <= H; ++tr)
    if (H % tr == 0 && tr * (W + 2) <= 128 && tr * W <= 256 && tr + 2 <= 256) best = tr;
  return best;
}

constexpr int kSmemLimit = 232448;   // 227 KB opt-in maximum per CTA

struct HaloPlan {
  HaloParams p;
  int smem;
};

static bool plan_fwd(int N, int H, int W, int c_patch, int c_out, HaloPlan* out) {
  if (c_patch % 64 || c_out % 64 || W + 2 > 256) return false;
  const int tr = pick_tr(H, W);
  if (tr == 0) return false;
  HaloParams p = {};
  p.n_img = N; p.H = H; p.W = W; p.Wp = W + 2; p.TR = tr;
  p.tiles_per_img = H / tr; p.tiles = N * p.tiles_per_img;
  p.n_parts = c_out / 64; p.chunks = c_patch / 64; p.c_patch = c_patch; p.c_out = c_out;
  const int win_rows = 2 * p.Wp + 2 + 128;
  const int box_rows = (tr + 2) * p.Wp;
  const int rows = win_rows > box_rows ? win_rows : box_rows;
  p.a_buf_bytes = ((rows * 128) + 1023) / 1024 * 1024;
  p.a_box_bytes = box_rows * 128;
  p.stg_bytes = ((tr * W * 128) + 1023) / 1024 * 1024;
  const int fixed = 9 * p.chunks * kWSlotBytes + 2 * p.stg_bytes + 256;
  int nbuf = (kSmemLimit - fixed) / p.a_buf_bytes;
  if (nbuf > 4) nbuf = 4;
  if (nbuf < 2) return false;
  p.nbuf = nbuf;
  out->p = p;
  out->smem = fixed + nbuf * p.a_buf_bytes;
  return true;
}

static bool plan_wgrad(int N, int H, int W, int Cin, int Cout, HaloWgradParams* out, int* smem_out) {
  if (Cin % 64 || Cout % 64 || W + 2 > 256) return false;
  const int tr = pick_tr(H, W);
  if (tr == 0) return false;
  HaloWgradParams p = {};
  p.n_img = N; p.H = H; p.W = W; p.Wp = W + 2; p.TR = tr;
  p.Cin = Cin; p.Cout = Cout;
  // a tile covers whole rows of NI consecutive images when one image's rows leave K-rows unused
  int ni = 1;
  if (tr == H) {
    ni = 128 / (tr * p.Wp);
    if (ni < 1) ni = 1;
    if (ni > N) ni = N;
  }
  p.NI = ni;
  p.tiles_per_img = H / tr;
  p.tiles = ((N + ni - 1) / ni) * p.tiles_per_img;
  // channels per CTA: 128 x 128 when both counts allow it (3 x 128 accumulator columns), else 64
  p.cich = Cin % 128 == 0 ? 2 : 1;
  p.coch = Cout % 128 == 0 ? 2 : 1;
  p.ci_roles = Cin / (64 * p.cich);
  p.co_roles = Cout / (64 * p.coch);
  const int nroles = 3 * p.ci_roles * p.co_roles;
  p.box_bytes = ni * tr * p.Wp * 128;
  int G = halo_sms() / nroles;
  if (G > p.tiles) G = p.tiles;
  if (G < 1) return false;
  p.G = G;
  const int stage_bytes = p.coch * kDyBufBytes + p.cich * kXBufBytes;
  int stages = (kSmemLimit - 256) / stage_bytes;
  if (stages > 4) stages = 4;
  if (stages < 2) return false;
  p.stages = stages;
  *out = p;
  *smem_out = stages * stage_bytes + 256;
  return true;
}

}  // namespace kdl

using namespace kdl;

extern "C" size_t kdl_conv3x3_halo_wgrad_scratch_bytes(int N, int H, int W, int Cin, int Cout) {
  HaloWgradParams p;
  int smem;
  if (!plan_wgrad(N, H, W, Cin, Cout, &p, &smem)) return 0;
  return (size_t)p.G * Cout * 9 * Cin * sizeof(float);
}

// dw [Cout,3,3,Cin] = wgrad(dy [N,H,W,Cout], x [N,H,W,Cin]); scratch of the size above (0 = unsupported)
extern "C" int kdl_conv3x3_halo_wgrad_bf16(const void* dy, const void* x, void* dw, int N, int H, int W, int Cin,
                                           int Cout, void* scratch, size_t scratch_bytes, cudaStream_t stream) {
  HaloWgradParams p;
  int smem;
  if (!plan_wgrad(N, H, W, Cin, Cout, &p, &smem)) {
    kdl_set_error("conv3x3 halo wgrad: unsupported shape N=%d H=%d W=%d Cin=%d Cout=%d", N, H, W, Cin, Cout);
    return KDL_ERR_UNSUPPORTED;
  }
  const size_t need = (size_t)p.G * Cout * 9 * Cin * sizeof(float);
  if (scratch == nullptr || scratch_bytes < need) {
    kdl_set_error("conv3x3 halo wgrad: scratch too small (%zu < %zu)", scratch_bytes, need);
    return KDL_ERR_WORKSPACE;
  }
  p.partial = (float*)scratch;
  CUtensorMap mdy, mx;
  int rc;
  {
    cuuint64_t dims[4] = {(cuuint64_t)Cout, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)Cout * 2, (cuuint64_t)W * Cout * 2, (cuuint64_t)H * W * Cout * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wp, (cuuint32_t)p.TR, (cuuint32_t)p.NI};
    if ((rc = encode(&mdy, dy, 4, dims, strides, box, "wgrad dy"))) return rc;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)Cin, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)Cin * 2, (cuuint64_t)W * Cin * 2, (cuuint64_t)H * W * Cin * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wp, (cuuint32_t)p.TR, (cuuint32_t)p.NI};
    if ((rc = encode(&mx, x, 4, dims, strides, box, "wgrad x"))) return rc;
  }
  static bool attr = false;
  if (!attr) {
    KDL_CUDA(cudaFuncSetAttribute(conv3x3_halo_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    attr = true;
  }
  conv3x3_halo_wgrad_kernel<<<3 * p.ci_roles * p.co_roles * p.G, kHaloThreads, smem, stream>>>(mdy, mx, p);
  KDL_CHECK_LAUNCH();
  const int n = Cout * 9 * Cin;
  if (!launch_plane_reduce_small(p.partial, (bf16*)dw, n, n, n, p.G, stream))
    halo_wgrad_reduce_kernel<<<(n / 4 + 255) / 256, 256, 0, stream>>>(p.partial, (bf16*)dw, n, p.G);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

namespace kdl {

static bool plan_fwd2(int N, int H, int W, int c_patch, int c_out, Halo2Params* out, int* smem_out, int* nb_out) {
  static int two = -1;
  if (two < 0) {
    const char* e = getenv("KDL_HALO_2CTA");
    two = e ? atoi(e) : 1;
  }
  if (!two || c_patch % 64 || c_out % 64 || W + 2 > 256) return false;
  const int tr = pick_tr(H, W);
  if (tr == 0) return false;
  static int nb_pref = -1;
  if (nb_pref < 0) {
    const char* e = getenv("KDL_HALO_NB");
    nb_pref = e ? atoi(e) : 0;
  }
  // N of the pair's MMA: 128 output channels halve the shared-memory operand reads per FLOP, but their
  // weights (144 KB per CTA at 128 input channels) leave room for only two halo buffers; then 64 with
  // a deeper ring is the better trade
  for (int nb = (c_out % 128 == 0 && nb_pref != 64) ? 128 : 64; nb >= 64; nb -= 64) {
    Halo2Params p = {};
    p.n_img = N; p.H = H; p.W = W; p.Wp = W + 2; p.TR = tr;
    p.tiles_per_img = H / tr; p.tiles = N * p.tiles_per_img; p.pair_tiles = (p.tiles + 1) / 2;
    p.chunks = c_patch / 64; p.c_patch = c_patch; p.c_out = c_out; p.n_parts = c_out / nb;
    const int win_rows = 2 * p.Wp + 2 + 128;
    const int box_rows = (tr + 2) * p.Wp;
    const int rows = win_rows > box_rows ? win_rows : box_rows;
    p.a_buf_bytes = ((rows * 128) + 1023) / 1024 * 1024;
    p.a_box_bytes = box_rows * 128;
    p.stg_bytes = ((tr * W * 128) + 1023) / 1024 * 1024;
    const int fixed = 9 * p.chunks * (nb / 2) * 128 + (nb == 128 ? 0 : p.stg_bytes) + 256;
    int nbuf = (kSmemLimit - fixed) / p.a_buf_bytes;
    if (nbuf > 4) nbuf = 4;
    const int want = (nb == 128 && nb_pref != 128) ? 3 : 2;   // 128-wide only if the ring stays deep enough
    if (nbuf < want) continue;
    if (halo_sms() / 2 < p.n_parts) continue;
    p.nbuf = nbuf;
    *out = p;
    *smem_out = fixed + nbuf * p.a_buf_bytes;
    *nb_out = nb;
    return true;
  }
  return false;
}

template <int NB>
static int launch_halo2(const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mc, const Halo2Params& p,
                        int smem, cudaStream_t stream) {
  auto kern = conv3x3_halo2_kernel<NB>;
  static bool attr_done = false;
  if (!attr_done) {
    KDL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    attr_done = true;
  }
  int pairs = halo_sms() / 2;
  pairs -= pairs % p.n_parts;
  if (p.pair_tiles * p.n_parts < pairs) pairs = p.pair_tiles * p.n_parts;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(kHaloThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  KDL_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mw, mc, p));
  return KDL_OK;
}

}  // namespace kdl

using namespace kdl;

extern "C" int kdl_conv3x3_halo_supported(int mode, int N, int H, int W, int Cin, int Cout) {
  HaloPlan pl;
  if (mode == 0) return plan_fwd(N, H, W, Cin, Cout, &pl) ? 1 : 0;
  if (mode == 1) return plan_fwd(N, H, W, Cout, Cin, &pl) ? 1 : 0;
  return 0;
}

// bytes of scratch kdl_conv3x3_halo_bf16 wants for this call (dgrad on the 2-CTA kernel keeps a flipped,
// transposed copy of the filter there); 0 when none is needed
extern "C" size_t kdl_conv3x3_halo_scratch_bytes(int mode, int N, int H, int W, int Cin, int Cout) {
  Halo2Params p2;
  int smem, nb;
  if (mode == 1 && plan_fwd2(N, H, W, Cout, Cin, &p2, &smem, &nb)) return (size_t)Cout * 9 * Cin * sizeof(bf16);
  return 0;
}

// mode 0 fprop: in0 = x [N,H,W,Cin], in1 = w [Cout,3,3,Cin], out = y [N,H,W,Cout]
// mode 1 dgrad: in0 = dy [N,H,W,Cout], in1 = w, out = dx [N,H,W,Cin] (+ addend)
extern "C" int kdl_conv3x3_halo_bf16(int mode, const void* in0, const void* in1, void* out, int N, int H,
                                     int W, int Cin, int Cout, const void* addend, int relu, void* scratch,
                                     size_t scratch_bytes, cudaStream_t stream) {
  if (mode != 0 && mode != 1) {
    kdl_set_error("conv3x3 halo: unknown mode %d", mode);
    return KDL_ERR_ARG;
  }
  const int c_patch = mode == 0 ? Cin : Cout;
  const int c_out = mode == 0 ? Cout : Cin;
  int rc;
  Halo2Params p2;
  int smem2, nb2;
  if (plan_fwd2(N, H, W, c_patch, c_out, &p2, &smem2, &nb2)) {
    // ---- CTA-pair kernel ----
    const void* wsrc = in1;
    if (mode == 1) {
      const size_t need = (size_t)Cout * 9 * Cin * sizeof(bf16);
      if (scratch == nullptr || scratch_bytes < need) {
        kdl_set_error("conv3x3 halo dgrad: scratch too small (%zu < %zu)", scratch_bytes, need);
        return KDL_ERR_WORKSPACE;
      }
      const int n = Cout * 9 * Cin;
      halo_flip_transpose_kernel<<<(n + 255) / 256, 256, 0, stream>>>((const bf16*)in1, (bf16*)scratch, Cout, Cin);
      KDL_CHECK_LAUNCH();
      wsrc = scratch;
    }
    p2.relu = relu;
    p2.addend = (const bf16*)addend;
    p2.out = (bf16*)out;
    CUtensorMap ma, mw, mc;
    {
      cuuint64_t dims[4] = {(cuuint64_t)c_patch, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
      cuuint64_t strides[3] = {(cuuint64_t)c_patch * 2, (cuuint64_t)W * c_patch * 2, (cuuint64_t)H * W * c_patch * 2};
      cuuint32_t box[4] = {64, (cuuint32_t)p2.Wp, (cuuint32_t)(p2.TR + 2), 1};
      if ((rc = encode(&ma, in0, 4, dims, strides, box, "halo"))) return rc;
    }
    {
      cuuint64_t dims[2] = {(cuuint64_t)9 * c_patch, (cuuint64_t)c_out};
      cuuint64_t strides[1] = {(cuuint64_t)9 * c_patch * 2};
      cuuint32_t box[2] = {64, (cuuint32_t)(nb2 / 2)};
      if ((rc = encode(&mw, wsrc, 2, dims, strides, box, "weights"))) return rc;
    }
    {
      cuuint64_t dims[2] = {(cuuint64_t)c_out, (cuuint64_t)N * H * W};
      cuuint64_t strides[1] = {(cuuint64_t)c_out * 2};
      cuuint32_t box[2] = {64, (cuuint32_t)(p2.TR * W)};
      if ((rc = encode(&mc, out, 2, dims, strides, box, "output"))) return rc;
    }
    return nb2 == 128 ? launch_halo2<128>(ma, mw, mc, p2, smem2, stream)
                      : launch_halo2<64>(ma, mw, mc, p2, smem2, stream);
  }
  HaloPlan pl;
  if (!plan_fwd(N, H, W, c_patch, c_out, &pl)) {
    kdl_set_error("conv3x3 halo: unsupported shape N=%d H=%d W=%d Cin=%d Cout=%d mode=%d", N, H, W, Cin, Cout, mode);
    return KDL_ERR_UNSUPPORTED;
  }
  HaloParams& p = pl.p;
  p.flip = mode == 1;
  p.relu = relu;
  p.addend = (const bf16*)addend;
  CUtensorMap ma, mw, mc;
  {
    cuuint64_t dims[4] = {(cuuint64_t)c_patch, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)c_patch * 2, (cuuint64_t)W * c_patch * 2, (cuuint64_t)H * W * c_patch * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wp, (cuuint32_t)(p.TR + 2), 1};
    if ((rc = encode(&ma, in0, 4, dims, strides, box, "halo"))) return rc;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)9 * Cin, (cuuint64_t)Cout};
    cuuint64_t strides[1] = {(cuuint64_t)9 * Cin * 2};
    cuuint32_t box[2] = {64, 64};
    if ((rc = encode(&mw, in1, 2, dims, strides, box, "weights"))) return rc;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)c_out, (cuuint64_t)N * H * W};
    cuuint64_t strides[1] = {(cuuint64_t)c_out * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)(p.TR * W)};
    if ((rc = encode(&mc, out, 2, dims, strides, box, "output"))) return rc;
  }
  int grid = halo_sms();
  grid -= grid % p.n_parts;
  const int max_ctas = p.tiles * p.n_parts;
  if (grid > max_ctas) grid = max_ctas;
  if (mode == 0) {
    static bool attr0 = false;
    if (!attr0) {
      KDL_CUDA(cudaFuncSetAttribute(conv3x3_halo_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
      attr0 = true;
    }
    conv3x3_halo_kernel<0><<<grid, kHaloThreads, pl.smem, stream>>>(ma, mw, mc, p);
  } else {
    static bool attr1 = false;
    if (!attr1) {
      KDL_CUDA(cudaFuncSetAttribute(conv3x3_halo_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
      attr1 = true;
    }
    conv3x3_halo_kernel<1><<<grid, kHaloThreads, pl.smem, stream>>>(ma, mw, mc, p);
  }
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}
