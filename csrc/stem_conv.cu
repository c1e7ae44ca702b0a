// kubedl_b200 — the 7x7 / stride-2 / pad-3 stem convolution of ResNet-50 straight from the uint8 image
// (kernel N7c): forward and weight gradient without a patch matrix in global memory.
//
// The first stem path wrote normalised bf16 patches [N*112*112, 168] (1.08 GB per step at batch 256), read
// them back in the forward GEMM and again in the wgrad GEMM.  Here a CTA owns one output row (n, oh) at a
// time: "builder" warps stage the seven image rows it needs in shared memory (uint8 -> normalised bf16) and
// write the 112 x 168 patch tile directly in the 128-byte-swizzled K-major layout tcgen05.mma reads
// (generic-proxy stores + fence.proxy.async), double-buffered against the MMA / epilogue of the previous
// row.  The tile is [128 pixel rows x 64 columns] x 3 column blocks (K = 168 padded to 192):
//   * forward: A = patch tile (K-major), B = the 64 x 192 filter (resident), D[128 px, 64] -> TMA store,
//     plus the BatchNorm column statistics of the stored values from the epilogue;
//   * wgrad:   the SAME bytes are an MN-major B operand (K = pixel rows, N = 3 x 64 patch columns),
//     A = the dy tile (MN-major, M = 64 output channels); D[64, 192] stays in TMEM for all rows a CTA
//     processes, fp32 partials are reduced in a fixed order afterwards.
// Patch column layout: kh*24 + kw*3 + c (a filter row = 21 values padded to 24), as kdl_stem_im2col_u8.
//
// No reference counterpart (the operator launches pods; cuDNN runs this layer in the stock worker).
#include "common.cuh"
#include "tcgen05.cuh"
#include "../include/kubedl_b200.h"
#include <cuda.h>

namespace kdl {

constexpr int kStemBuilders = 512;  // building a patch tile is the critical path: 128 builder threads made the
                                   // kernels 3x slower than the patch-matrix path they replace
constexpr int kStemThreads = 256 + kStemBuilders;   // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps 4-7 epilogue, warps 8.. builders
constexpr int kStemW = 224;         // input width / height this kernel is specialised for
constexpr int kStemOW = 112;
constexpr int kStemABytes = 3 * 128 * 128;            // patch tile: 3 column blocks of 128 rows x 128 B
constexpr int kStemWBytes = 3 * 64 * 128;             // filter: 3 column blocks of 64 rows x 128 B
constexpr int kStemStgBytes = 112 * 128;              // output staging tile (112 pixels x 64 channels bf16)

struct StemParams {
  const uint8_t* img;    // [N, 224, 224, 3]
  int rows;              // N * 112 output rows
  int rows_per_cta;      // consecutive output rows per CTA (host: ceil(rows / SMs); grid = ceil(rows / rows_per_cta),
                         // so every CTA has at least one row — the wgrad accumulator of an idle CTA would be garbage)
  float m0, m1, m2, s0, s1, s2;   // mean, 1/std per channel
  float* stats;          // fprop: [gridDim.x][2][64] column sum / sum of squares (or null)
  float* partial;        // wgrad: [gridDim.x][64][192]
};

__device__ __forceinline__ void builders_sync() { asm volatile("bar.sync 2, %0;" ::"n"(kStemBuilders) : "memory"); }
__device__ __forceinline__ void stem_epi_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// Raw image rows: the producer warp prefetches the 7 rows (672 B each) an output row needs with 1-D bulk
// async copies, one output row ahead, into slots of kRawSlot bytes (data at byte 16, zero guards either side);
// the builders convert straight from those bytes.  (The first version had the builders load the image with
// ordinary loads and stage bf16 rows: every tile then paid a global-load round trip, ~5k clocks per tile.)
constexpr int kRawSlot = 16 + kStemW * 3 + 16;     // 704
constexpr int kRawBytes = 7 * kRawSlot;            // one output row's worth

__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// First image row (as filter-row index kh) that tile `row` has to bring in: all seven for the first tile of a
// CTA and for the first output row of an image, otherwise the two rows the previous tile did not have.
__device__ __forceinline__ int stem_first_new_kh(int row, int row_lo) {
  return (row == row_lo || row % kStemOW == 0) ? 0 : 5;
}

// Issued by one thread: the NEW image rows of output row `row` that exist (out-of-image rows are skipped;
// the builders write zeros for them).
__device__ __forceinline__ void prefetch_raw_rows(const StemParams& p, int row, int row_lo, uint32_t raw_smem,
                                                  uint32_t bar) {
  const int n = row / kStemOW, oh = row - n * kStemOW;
  const int kh0 = stem_first_new_kh(row, row_lo);
  int valid = 0;
  for (int kh = kh0; kh < 7; ++kh) {
    const int ih = oh * 2 - 3 + kh;
    valid += (ih >= 0 && ih < kStemW);
  }
  if (valid == 0) {            // nothing to fetch: complete the phase without a transfer
    mbar_arrive(bar);
    return;
  }
  mbar_expect_tx(bar, (uint32_t)(valid * kStemW * 3));
  for (int kh = kh0; kh < 7; ++kh) {
    const int ih = oh * 2 - 3 + kh;
    if (ih < 0 || ih >= kStemW) continue;
    bulk_load_1d(raw_smem + kh * kRawSlot + 16, p.img + ((size_t)n * kStemW + ih) * kStemW * 3, kStemW * 3, bar);
  }
}

constexpr int kStemRowElems = (kStemW + 6) * 3 + 8;   // normalised image row: 3 zero pixels either side + slack

// Emit the patch tile of output row `row` into `tile` (swizzled K-major, rows = output pixels).  Two phases by
// the kStemBuilders builder threads: (1) the prefetched raw uint8 rows -> normalised bf16 rows `srow` (every
// image element converted once; converting per patch vector instead was measured 4x more work and slower),
// (2) 112 x 24 sixteen-byte vectors copied out of `srow` with four aligned 32-bit shared loads each.
__device__ __forceinline__ void build_patch_tile(const StemParams& p, int row, int row_lo, const uint8_t* raw,
                                                 bf16 (*srow)[kStemRowElems], uint8_t* tile, int bt) {
  const int oh = row % kStemOW;
  const int kh0 = stem_first_new_kh(row, row_lo);
  const int nnew = 7 - kh0;
  const float k255 = 1.f / 255.f;
  const float sc[3] = {k255 * p.s0, k255 * p.s1, k255 * p.s2};
  const float bi[3] = {-p.m0 * p.s0, -p.m1 * p.s1, -p.m2 * p.s2};
  constexpr int words = kStemW * 3 / 4;   // 168 four-byte words per image row
  for (int i = bt; i < nnew * words; i += kStemBuilders) {
    const int kh = kh0 + i / words, w4 = i % words;
    const int ih = oh * 2 - 3 + kh;
    const bool in = ih >= 0 && ih < kStemW;
    const uint32_t v = in ? *reinterpret_cast<const uint32_t*>(raw + kh * kRawSlot + 16 + 4 * w4) : 0u;
    const int c0 = (4 * w4) % 3;
    float f[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = (c0 + j) % 3;
      f[j] = in ? fmaf((float)((v >> (8 * j)) & 0xffu), sc[c], bi[c]) : 0.f;
    }
    // srow element 9 + 4*w4: odd element index -> 2-byte stores
    bf16* dst = &srow[(ih + 8) & 7][9 + 4 * w4];      // ring slot of image row ih
#pragma unroll
    for (int j = 0; j < 4; ++j) dst[j] = __float2bfloat16(f[j]);
  }
  builders_sync();
  for (int q = bt; q < kStemOW * 24; q += kStemBuilders) {
    const int ow = q / 24, v = q - ow * 24;
    uint4 o = make_uint4(0, 0, 0, 0);
    if (v < 21) {
      const int kh = v / 3, j0 = (v - kh * 3) * 8;
      const uint32_t* src = reinterpret_cast<const uint32_t*>(&srow[(oh * 2 - 3 + kh + 8) & 7][ow * 6 + j0]);
      o = make_uint4(src[0], src[1], src[2], src[3]);
      if (j0 == 16) { o.z &= 0x0000ffffu; o.w = 0u; }    // columns 21..23 of a filter row are padding
    }
    const int kb = v >> 3, c16 = v & 7;
    *reinterpret_cast<uint4*>(tile + kb * (128 * 128) + ow * 128 + ((c16 ^ (ow & 7)) << 4)) = o;
  }
  fence_proxy_async_smem();     // the tile is read by the tensor core (async proxy)
  builders_sync();
}

// shared-memory map (bytes from the 1024-aligned base)
constexpr int kOffW = 0;
constexpr int kOffA = kOffW + kStemWBytes;              // two patch tiles
constexpr int kOffStg = kOffA + 2 * kStemABytes;        // fprop: two output staging tiles; wgrad: two dy tiles
constexpr int kOffSrow = kOffStg + 2 * 16384;
constexpr int kOffNorm = kOffSrow + ((2 * kRawBytes + 1023) / 1024) * 1024;    // after the two raw-row buffers
constexpr int kOffBar = kOffNorm + ((8 * kStemRowElems * 2 + 1023) / 1024) * 1024;     // ring of 8 normalised rows
constexpr int kStemSmem = kOffBar + 256 + 1024;

__global__ void __launch_bounds__(kStemThreads, 1)
stem_fprop_kernel(const __grid_constant__ CUtensorMap tma_w,   // filter [64][168], box {64, 64}
                  const __grid_constant__ CUtensorMap tma_y,   // output [rows*112][64], box {64, 112}
                  const StemParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar = sbase + kOffBar;
  auto wfull = [&]() { return bar; };
  auto afull = [&](int b) { return bar + 8u * (1 + b); };
  auto aempty = [&](int b) { return bar + 8u * (3 + b); };
  auto tfull = [&](int b) { return bar + 8u * (5 + b); };
  auto tempty = [&](int b) { return bar + 8u * (7 + b); };
  auto rawfull = [&](int b) { return bar + 8u * (9 + b); };
  auto rawempty = [&](int b) { return bar + 8u * (11 + b); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + kOffBar + 8 * 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 128;     // two 64-column accumulators

  // staged rows: the pad pixels are zero for the whole kernel; patch-tile rows 112..127 feed junk
  // accumulator rows only, but keep them finite
  for (int i = threadIdx.x; i < (kOffBar - kOffA) / 16; i += kStemThreads)
    reinterpret_cast<uint4*>(smem + kOffA)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_y) : "memory");
  }
  if (warp == 1 && lane == 0) {
    mbar_init(wfull(), 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(afull(b), 1);
      mbar_init(aempty(b), 1);
      mbar_init(tfull(b), 1);
      mbar_init(tempty(b), 1);
      mbar_init(rawfull(b), 1);
      mbar_init(rawempty(b), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)),
                 "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  // a CTA walks CONSECUTIVE output rows: row oh+1 shares five of its seven image rows with row oh, so only
  // two new rows are fetched and normalised per tile (a ring of eight normalised rows)
  const int row_lo = min((int)blockIdx.x * p.rows_per_cta, p.rows), row_hi = min(row_lo + p.rows_per_cta, p.rows);

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(wfull(), kStemWBytes);
      for (int kb = 0; kb < 3; ++kb) tma_load_2d(sbase + kOffW + kb * 8192, &tma_w, wfull(), kb * 64, 0);
      int iter = 0;
      for (int row = row_lo; row < row_hi; ++row, ++iter) {     // raw image rows, one tile ahead
        const int b = iter & 1;
        mbar_wait(rawempty(b), ((iter >> 1) & 1) ^ 1);
        prefetch_raw_rows(p, row, row_lo, sbase + kOffSrow + b * kRawBytes, rawfull(b));
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: 3 column blocks x 4 K-slices per output row
    const uint32_t el = elect_one();
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    mbar_wait(wfull(), 0);
    tcgen05_fence_after();
    int iter = 0;
    for (int row = row_lo; row < row_hi; ++row, ++iter) {
      const int b = iter & 1;
      const uint32_t ph = (iter >> 1) & 1;
      mbar_wait(tempty(b), ph ^ 1);
      mbar_wait(afull(b), ph);
      tcgen05_fence_after();
      const uint64_t ad0 = make_desc<0>(sbase + kOffA + b * kStemABytes);
      const uint64_t bd0 = make_desc<0>(sbase + kOffW);
#pragma unroll
      for (int kb = 0; kb < 3; ++kb)
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16_pred(tmem_base + b * 64, ad0 + (uint64_t)(kb * 1024 + k * 2), bd0 + (uint64_t)(kb * 512 + k * 2), idesc,
                         (kb > 0 || k > 0) ? 1u : 0u, el);
      umma_commit_pred(aempty(b), el);
      umma_commit_pred(tfull(b), el);
    }
  } else if (warp >= 4 && warp < 8) {
    // ---------------- epilogue: TMEM -> bf16 -> swizzled staging -> one TMA store of 112 pixels; BN statistics
    const int q = warp & 3;
    const int m = q * 32 + lane;
    const bool valid = m < kStemOW;
    const bool leader = warp == 4 && lane == 0;
    float st[4] = {0.f, 0.f, 0.f, 0.f};
    int iter = 0;
    for (int row = row_lo; row < row_hi; ++row, ++iter) {
      const int b = iter & 1;
      const uint32_t ph = (iter >> 1) & 1;
      mbar_wait(tfull(b), ph);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + b * 64;
      if (leader) tma_store_wait_read<1>();     // staging tile b was handed to TMA two rows ago
      stem_epi_sync();
      uint8_t* stg = smem + kOffStg + b * 16384;
#pragma unroll
      for (int hc = 0; hc < 2; ++hc) {          // 32 columns at a time: 768 threads leave ~80 registers each
        uint32_t r[32];
        tmem_ld32(taddr + hc * 32, r);
        tmem_ld_wait();
        if (valid) {
          uint8_t* dst = stg + m * 128;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float f[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(r[8 * j + i]);
            *reinterpret_cast<uint4*>(dst + (((hc * 4 + j) ^ (m & 7)) << 4)) = pack8(f);
          }
        }
      }
      tcgen05_fence_before();
      fence_proxy_async_smem();
      stem_epi_sync();
      if (leader) {
        tma_store_2d(&tma_y, sbase + kOffStg + b * 16384, 0, row * kStemOW);
        tma_store_commit();
        mbar_arrive(tempty(b));
      }
      if (p.stats) {      // lane l: columns 2l, 2l+1 over this warp's valid rows of the stored (bf16) tile
        const int nrows = min(32, kStemOW - q * 32);
        for (int rr = 0; rr < nrows; ++rr) {
          const int mm = q * 32 + rr;
          const uint32_t w = *reinterpret_cast<const uint32_t*>(stg + mm * 128 + ((((lane >> 2) ^ (mm & 7))) << 4) + (lane & 3) * 4);
          const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w));
          st[0] += v.x; st[1] += v.y; st[2] = fmaf(v.x, v.x, st[2]); st[3] = fmaf(v.y, v.y, st[3]);
        }
      }
    }
    if (leader) tma_store_wait_read<0>();
    __syncwarp();
    if (p.stats) {     // add the four warps' partials in shared memory: one partial row per CTA
      stem_epi_sync();                                   // every TMA store has read the staging tiles
      float* sm = reinterpret_cast<float*>(smem + kOffStg);
      sm[(q * 2 + 0) * 64 + 2 * lane] = st[0]; sm[(q * 2 + 0) * 64 + 2 * lane + 1] = st[1];
      sm[(q * 2 + 1) * 64 + 2 * lane] = st[2]; sm[(q * 2 + 1) * 64 + 2 * lane + 1] = st[3];
      stem_epi_sync();
      const int t = q * 32 + lane;                       // 128 threads: 64 sums + 64 sums of squares
      const int k = t >> 6, col = t & 63;
      float acc = 0.f;
#pragma unroll
      for (int w = 0; w < 4; ++w) acc += sm[(w * 2 + k) * 64 + col];
      p.stats[(size_t)blockIdx.x * 128 + k * 64 + col] = acc;
    }
  } else if (warp >= 8) {
    // ---------------- builders
    const int bt = threadIdx.x - 256;
    int iter = 0;
    for (int row = row_lo; row < row_hi; ++row, ++iter) {
      const int b = iter & 1;
      const uint32_t ph = (iter >> 1) & 1;
      mbar_wait(aempty(b), ph ^ 1);
      mbar_wait(rawfull(b), ph);
      build_patch_tile(p, row, row_lo, smem + kOffSrow + b * kRawBytes,
                       reinterpret_cast<bf16 (*)[kStemRowElems]>(smem + kOffNorm), smem + kOffA + b * kStemABytes, bt);
      if (bt == 0) {
        mbar_arrive(afull(b));
        mbar_arrive(rawempty(b));
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// ------------------------------------------------------------------ wgrad
__device__ __forceinline__ uint64_t stem_desc_mn(uint32_t smem_addr, uint32_t lbo) {   // MN-major, 128B swizzle
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((1024 >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__global__ void __launch_bounds__(kStemThreads, 1)
stem_wgrad_kernel(const __grid_constant__ CUtensorMap tma_dy,   // dy [rows*112][64], box {64, 112}
                  const StemParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar = sbase + kOffBar;
  auto dyfull = [&](int b) { return bar + 8u * b; };
  auto afull = [&](int b) { return bar + 8u * (2 + b); };
  auto empty = [&](int b) { return bar + 8u * (4 + b); };
  auto done = [&]() { return bar + 8u * 6; };
  auto rawfull = [&](int b) { return bar + 8u * (7 + b); };
  auto rawempty = [&](int b) { return bar + 8u * (9 + b); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + kOffBar + 8 * 11);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 256;     // D[64, 192]

  // rows 112..127 of the patch tiles and of the dy tiles are K rows of the MMA: they stay zero
  for (int i = threadIdx.x; i < (kOffBar - kOffA) / 16; i += kStemThreads)
    reinterpret_cast<uint4*>(smem + kOffA)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (warp == 0 && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_dy) : "memory");
  if (warp == 1 && lane == 0) {
    for (int b = 0; b < 2; ++b) {
      mbar_init(dyfull(b), 1);
      mbar_init(afull(b), 1);
      mbar_init(empty(b), 1);
      mbar_init(rawfull(b), 1);
      mbar_init(rawempty(b), 1);
    }
    mbar_init(done(), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)),
                 "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  // a CTA walks CONSECUTIVE output rows: row oh+1 shares five of its seven image rows with row oh, so only
  // two new rows are fetched and normalised per tile (a ring of eight normalised rows)
  const int row_lo = min((int)blockIdx.x * p.rows_per_cta, p.rows), row_hi = min(row_lo + p.rows_per_cta, p.rows);

  if (warp == 0) {
    if (lane == 0) {
      int iter = 0;
      for (int row = row_lo; row < row_hi; ++row, ++iter) {
        const int b = iter & 1;
        mbar_wait(rawempty(b), ((iter >> 1) & 1) ^ 1);
        prefetch_raw_rows(p, row, row_lo, sbase + kOffSrow + b * kRawBytes, rawfull(b));
        mbar_wait(empty(b), ((iter >> 1) & 1) ^ 1);
        mbar_expect_tx(dyfull(b), kStemStgBytes);
        tma_load_2d(sbase + kOffStg + b * 16384, &tma_dy, dyfull(b), 0, row * kStemOW);
      }
    }
  } else if (warp == 1) {
    const uint32_t el = elect_one();
    // M = 64 output channels (A = dy, MN-major), N = 192 patch columns (B = patch tile, MN-major, three
    // 64-wide atoms 16 KB apart), K = 16 pixel rows per instruction
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                               ((uint32_t)(192 >> 3) << 17) | ((uint32_t)(64 >> 4) << 24);
    int iter = 0;
    for (int row = row_lo; row < row_hi; ++row, ++iter) {
      const int b = iter & 1;
      const uint32_t ph = (iter >> 1) & 1;
      mbar_wait(dyfull(b), ph);
      mbar_wait(afull(b), ph);
      tcgen05_fence_after();
      const uint64_t ad0 = stem_desc_mn(sbase + kOffStg + b * 16384, 0);
      const uint64_t bd0 = stem_desc_mn(sbase + kOffA + b * kStemABytes, 128 * 128);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        umma_bf16_pred(tmem_base, ad0 + (uint64_t)(k * 128), bd0 + (uint64_t)(k * 128), idesc, (k > 0 || iter > 0) ? 1u : 0u, el);
      umma_commit_pred(empty(b), el);
    }
    umma_commit_pred(done(), el);
  } else if (warp >= 4 && warp < 8) {
    const int q = warp & 3;
    mbar_wait(done(), 0);
    tcgen05_fence_after();
    // M = 64: accumulator row m lives in TMEM lane 32*(m/16) + m%16
    const int co = q * 16 + lane;
    float* dst = p.partial + ((size_t)blockIdx.x * 64 + co) * 192;
    for (int c0 = 0; c0 < 192; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + c0, r);
      tmem_ld_wait();
      if (lane < 16) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(dst + c0 + 4 * j) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
      }
    }
  } else if (warp >= 8) {
    const int bt = threadIdx.x - 256;
    int iter = 0;
    for (int row = row_lo; row < row_hi; ++row, ++iter) {
      const int b = iter & 1;
      mbar_wait(empty(b), ((iter >> 1) & 1) ^ 1);
      mbar_wait(rawfull(b), (iter >> 1) & 1);
      build_patch_tile(p, row, row_lo, smem + kOffSrow + b * kRawBytes,
                       reinterpret_cast<bf16 (*)[kStemRowElems]>(smem + kOffNorm), smem + kOffA + b * kStemABytes, bt);
      if (bt == 0) {
        mbar_arrive(afull(b));
        mbar_arrive(rawempty(b));
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// gw[64][168] = sum over CTAs of partial[g][64][192] (columns 168..191 are K padding)
__global__ void stem_wgrad_reduce_kernel(const float* __restrict__ partial, bf16* __restrict__ gw, int G) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * 168) return;
  const int co = i / 168, k = i - co * 168;
  float acc = 0.f;
  for (int g = 0; g < G; ++g) acc += partial[((size_t)g * 64 + co) * 192 + k];
  gw[i] = __float2bfloat16(acc);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*StemEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static StemEncodeFn stem_encode() {
  static StemEncodeFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (StemEncodeFn)ptr;
    else
      cudaGetLastError();
  }
  return fn;
}

// 2-D bf16 map [rows][cols] (row pitch = cols), 128-byte swizzle
static int stem_map(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint32_t box_c, uint32_t box_r,
                    const char* what) {
  StemEncodeFn enc = stem_encode();
  if (!enc) {
    kdl_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return KDL_ERR_UNSUPPORTED;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {box_c, box_r};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    kdl_set_error("stem conv: cuTensorMapEncodeTiled(%s) failed (%d)", what, (int)r);
    return KDL_ERR_CUDA;
  }
  return KDL_OK;
}

static int stem_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

static int stem_fill(StemParams* p, const void* img, int N, int H, int W, const float* mean3, const float* std3) {
  if (H != kStemW || W != kStemW || N < 1) {
    kdl_set_error("stem conv: specialised for %dx%d images (got %dx%d, N=%d)", kStemW, kStemW, H, W, N);
    return KDL_ERR_UNSUPPORTED;
  }
  p->img = (const uint8_t*)img;
  p->rows = N * kStemOW;
  p->rows_per_cta = (p->rows + stem_sms() - 1) / stem_sms();
  p->m0 = mean3[0]; p->m1 = mean3[1]; p->m2 = mean3[2];
  p->s0 = 1.f / std3[0]; p->s1 = 1.f / std3[1]; p->s2 = 1.f / std3[2];
  return KDL_OK;
}

}  // namespace kdl

using namespace kdl;

extern "C" int kdl_stem_conv_supported(int H, int W) { return H == kStemW && W == kStemW; }
extern "C" int kdl_stem_conv_max_parts(void) { return stem_sms(); }
extern "C" size_t kdl_stem_conv_wgrad_scratch_bytes(void) { return (size_t)stem_sms() * 64 * 192 * sizeof(float); }

// y [N,112,112,64] = conv7x7/s2/p3(normalise(img_u8 [N,224,224,3]), w); w_padded is [64][168] in the
// kh*24 + kw*3 + c column layout.  stats_partial (optional): [kdl_stem_conv_max_parts()][2][64] floats,
// *stats_parts_out rows written — feed kdl_bn_fwd_from_stats.
extern "C" int kdl_stem_conv_fwd(const void* img_u8, const void* w_padded, void* y, int N, int H, int W,
                                 const float* mean3, const float* std3, float* stats_partial, int* stats_parts_out,
                                 cudaStream_t stream) {
  StemParams p = {};
  int rc = stem_fill(&p, img_u8, N, H, W, mean3, std3);
  if (rc) return rc;
  p.stats = stats_partial;
  CUtensorMap mw, my;
  if ((rc = stem_map(&mw, w_padded, 168, 64, 64, 64, "filter"))) return rc;
  if ((rc = stem_map(&my, y, 64, (uint64_t)p.rows * kStemOW, 64, kStemOW, "output"))) return rc;
  static bool attr = false;
  if (!attr) {
    KDL_CUDA(cudaFuncSetAttribute(stem_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStemSmem));
    attr = true;
  }
  const int grid = (p.rows + p.rows_per_cta - 1) / p.rows_per_cta;
  if (stats_parts_out) *stats_parts_out = grid;
  stem_fprop_kernel<<<grid, kStemThreads, kStemSmem, stream>>>(mw, my, p);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

// gw [64][168] (same column layout) = wgrad of the stem: sum over pixels of dy[p, co] * patch[p, k]
extern "C" int kdl_stem_conv_wgrad(const void* img_u8, const void* dy, void* gw_padded, int N, int H, int W,
                                   const float* mean3, const float* std3, void* scratch, size_t scratch_bytes,
                                   cudaStream_t stream) {
  StemParams p = {};
  int rc = stem_fill(&p, img_u8, N, H, W, mean3, std3);
  if (rc) return rc;
  const int grid = (p.rows + p.rows_per_cta - 1) / p.rows_per_cta;
  const size_t need = (size_t)grid * 64 * 192 * sizeof(float);
  if (scratch == nullptr || scratch_bytes < need) {
    kdl_set_error("stem wgrad: scratch too small (%zu < %zu)", scratch_bytes, need);
    return KDL_ERR_WORKSPACE;
  }
  p.partial = (float*)scratch;
  CUtensorMap mdy;
  if ((rc = stem_map(&mdy, dy, 64, (uint64_t)p.rows * kStemOW, 64, kStemOW, "dy"))) return rc;
  static bool attr = false;
  if (!attr) {
    KDL_CUDA(cudaFuncSetAttribute(stem_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStemSmem));
    attr = true;
  }
  stem_wgrad_kernel<<<grid, kStemThreads, kStemSmem, stream>>>(mdy, p);
  KDL_CHECK_LAUNCH();
  stem_wgrad_reduce_kernel<<<(64 * 168 + 255) / 256, 256, 0, stream>>>(p.partial, (bf16*)gw_padded, grid);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}
