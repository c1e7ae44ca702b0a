// kubedl_b200 — the ResNet stem (7x7 / stride-2 / pad-3 convolution on a 3-channel image) as a 4x4 / stride-1
// convolution on a 2x2 space-to-depth image (kernel N7d): forward and weight gradient with NO patch building.
//
// The direct stem kernels (stem_conv.cu) build every 112 x 168 patch tile in shared memory with ordinary loads and
// stores, because a 3-channel pixel is 6 bytes and neither TMA nor a UMMA descriptor can step by 6 (or 12) bytes;
// they ended up bound by the shared-memory pipe ("Known gaps").  Here the image is rewritten once per
// step as  S[n][R][C][16] = normalise(img[n][2R+dy][2C+dx][c])  at channel (dy*2+dx)*3 + c  (12 used, 4 zero), with
// a zero border, so that
//     y[oh][ow][co] = sum_{kh', kw' in 0..3} sum_{ch < 16}  S[oh + kh'][ow + kw'][ch] * W2[co][kh'][kw'][ch]
// (padded coordinates; W2[co][kh'][kw'][(dy,dx,c)] = w[co][2kh'+dy-1][2kw'+dx-1][c], zero where that tap does not
// exist).  The four kw' taps x 16 channels of one kh' are 64 CONTIGUOUS bf16 = 128 bytes of S, and the window of the
// next output pixel starts 32 bytes later: a rank-4 tiled tensor map whose pixel stride (32 B) is smaller than its
// inner extent (128 B) hands TMA the whole [112 pixels x 64] K-block of a tile row in the 128-byte-swizzled layout
// tcgen05.mma reads.  A tile (one output row) is four such loads (K = 256); no thread touches the operands.
//   * forward: A = the four blocks (K-major), B = W2 [64][256] (resident), D[128 px, 64] -> bf16 -> TMA store, plus
//     the BatchNorm column statistics of the stored values from the epilogue;
//   * wgrad:   the SAME blocks are an MN-major operand (K = pixels): the gradient is accumulated transposed,
//     D[256, 64] as two M = 128 halves with B = the dy tile (MN-major, N = 64), and stays in TMEM for all rows of a
//     CTA; fp32 partials are reduced in a fixed order and mapped back to the [64][7][7][3] filter layout.
// S costs one 108 MB write per step (256 images) and is read by both kernels, mostly out of L2.
//
// No reference counterpart (the operator launches pods; cuDNN runs this layer in the stock worker).
#include "common.cuh"
#include "tcgen05.cuh"
#include "../include/kubedl_b200.h"
#include <cuda.h>

namespace kdl {

constexpr int kS2dIn = 224;            // input height / width this path is specialised for
constexpr int kS2dOW = 112;            // output height / width
constexpr int kS2dP = 115;             // padded space-to-depth rows / columns: 2 before, 112, 1 after
constexpr int kS2dC = 16;              // channels per space-to-depth pixel (12 used)
constexpr int kS2dK = 256;             // K of the implicit GEMM: 4 kh' x (4 kw' x 16 ch)
constexpr int kS2dThreads = 256;       // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps 4-7 epilogue
constexpr int kS2dBlock = 128 * 128;   // one K-block of a tile: 128 pixel rows x 128 B

// ------------------------------------------------------------------ space-to-depth image
// One thread per padded pixel: 2 x 2 x 3 bytes in, 32 bytes out (zero in the border and in channels 12..15).
__global__ void __launch_bounds__(256) stem_s2d_pack_kernel(const uint8_t* __restrict__ img, bf16* __restrict__ out,
                                                            int N, float m0, float m1, float m2, float s0, float s1,
                                                            float s2) {
  const int64_t total = (int64_t)N * kS2dP * kS2dP;
  const float k255 = 1.f / 255.f;
  const float sc[3] = {k255 * s0, k255 * s1, k255 * s2};
  const float bi[3] = {-m0 * s0, -m1 * s1, -m2 * s2};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t iu = (uint32_t)i;
    const uint32_t t = iu / (uint32_t)kS2dP;
    const int C = (int)(iu - t * (uint32_t)kS2dP) - 2;
    const int n = (int)(t / (uint32_t)kS2dP);
    const int R = (int)(t - (uint32_t)n * (uint32_t)kS2dP) - 2;
    float f[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = 0.f;
    if ((unsigned)R < (unsigned)kS2dOW && (unsigned)C < (unsigned)kS2dOW) {
#pragma unroll
      for (int dy = 0; dy < 2; ++dy) {
        // six contiguous bytes (two pixels) at an even offset: three 16-bit loads
        const uint16_t* src = reinterpret_cast<const uint16_t*>(img + (((size_t)n * kS2dIn + 2 * R + dy) * kS2dIn + 2 * C) * 3);
        const uint32_t a = src[0], b = src[1], c = src[2];
        const uint32_t by[6] = {a & 0xffu, a >> 8, b & 0xffu, b >> 8, c & 0xffu, c >> 8};
#pragma unroll
        for (int j = 0; j < 6; ++j) f[dy * 6 + j] = fmaf((float)by[j], sc[j % 3], bi[j % 3]);
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(out + (size_t)i * kS2dC);
    dst[0] = pack8(f);
    dst[1] = pack8(f + 8);
  }
}

// ------------------------------------------------------------------ filter layouts
// k2 = kh'*64 + kw'*16 + (dy*2+dx)*3 + c   <->   filter tap (kh, kw, c) = (2kh'+dy-1, 2kw'+dx-1, c);
// returns the column kh*24 + kw*3 + c of the padded [64][168] filter, or -1 where k2 is padding.
__device__ __forceinline__ int s2d_k_to_padded(int k2) {
  const int khp = k2 >> 6, kwp = (k2 >> 4) & 3, ch = k2 & 15;
  if (ch >= 12) return -1;
  const int dy = ch / 6, dx = (ch / 3) & 1, c = ch % 3;
  const int kh = 2 * khp + dy - 1, kw = 2 * kwp + dx - 1;
  if (kh < 0 || kw < 0) return -1;       // kh, kw <= 6 always
  return kh * 24 + kw * 3 + c;
}

__global__ void stem_s2d_weight_kernel(const bf16* __restrict__ w_padded, bf16* __restrict__ w2) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * kS2dK) return;
  const int co = i / kS2dK, k2 = i - co * kS2dK;
  const int k = s2d_k_to_padded(k2);
  w2[i] = k < 0 ? __float2bfloat16(0.f) : w_padded[co * 168 + k];
}

// gw_padded [64][168] = sum over CTAs of partial[g][256][64], mapped back (padding columns of gw are zeroed)
__global__ void stem_s2d_wgrad_reduce_kernel(const float* __restrict__ partial, bf16* __restrict__ gw, int G) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * kS2dK) return;
  const int k2 = i >> 6, co = i & 63;          // co fastest: the partial planes are [k2][co]
  if (k2 < 168 && (k2 % 24) >= 21) gw[co * 168 + k2] = __float2bfloat16(0.f);   // the 3 padding columns per filter row
  const int k = s2d_k_to_padded(k2);
  if (k < 0) return;
  float acc = 0.f;
  int g = 0;
  for (; g + 4 <= G; g += 4) {
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = __ldcg(partial + ((size_t)(g + u) * kS2dK + k2) * 64 + co);
#pragma unroll
    for (int u = 0; u < 4; ++u) acc += v[u];
  }
  for (; g < G; ++g) acc += __ldcg(partial + ((size_t)g * kS2dK + k2) * 64 + co);
  gw[co * 168 + k] = __float2bfloat16(acc);
}

// ------------------------------------------------------------------ kernels
struct S2dParams {
  int rows;              // N * 112 output rows (tiles)
  int rows_per_cta;      // consecutive output rows per CTA; grid = ceil(rows / rows_per_cta): no idle CTA
  float* stats;          // fprop: [gridDim.x][2][64] column sum / sum of squares of the stored values (or null)
  float* partial;        // wgrad: [gridDim.x][256][64] (transposed: space-to-depth column major)
};

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

__device__ __forceinline__ void s2d_epi_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// shared-memory map of the forward kernel (bytes from the 1024-aligned base)
constexpr int kFStages = 8;                               // K-blocks in flight: two tiles' worth
constexpr int kFOffW = 0;                                 // W2: 4 K-blocks of 64 rows x 128 B
constexpr int kFOffA = kFOffW + 4 * 64 * 128;             // ring of K-blocks
constexpr int kFOffStg = kFOffA + kFStages * kS2dBlock;   // two output staging tiles
constexpr int kFOffBar = kFOffStg + 2 * 16384;
constexpr int kFSmem = kFOffBar + 256 + 1024;

__global__ void __launch_bounds__(kS2dThreads, 1)
stem_s2d_fprop_kernel(const __grid_constant__ CUtensorMap tma_s,   // S as {64, 112, 115, N}, strides {32 B, row, image}
                      const __grid_constant__ CUtensorMap tma_w,   // W2 [64][256], box {64, 64}
                      const __grid_constant__ CUtensorMap tma_y,   // output [rows*112][64], box {64, 112}
                      const S2dParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar = sbase + kFOffBar;
  auto wfull = [&]() { return bar; };
  auto full = [&](int s) { return bar + 8u * (1 + s); };
  auto empty = [&](int s) { return bar + 8u * (1 + kFStages + s); };
  auto tfull = [&](int b) { return bar + 8u * (1 + 2 * kFStages + b); };
  auto tempty = [&](int b) { return bar + 8u * (3 + 2 * kFStages + b); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + kFOffBar + 8 * (5 + 2 * kFStages));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 128;     // two 64-column accumulators

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_s) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_y) : "memory");
  }
  if (warp == 1 && lane == 0) {
    mbar_init(wfull(), 1);
    for (int s = 0; s < kFStages; ++s) {
      mbar_init(full(s), 1);
      mbar_init(empty(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tfull(b), 1);
      mbar_init(tempty(b), 1);
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
  const int row_lo = min((int)blockIdx.x * p.rows_per_cta, p.rows), row_hi = min(row_lo + p.rows_per_cta, p.rows);

  if (warp == 0) {
    // ---------------- TMA producer: W2 once, then four K-blocks per output row.  Pixels 112..127 of a box are out of
    // the map's bounds and arrive as zeros (they only feed accumulator rows nobody reads).
    if (lane == 0) {
      mbar_expect_tx(wfull(), 4 * 64 * 128);
      for (int kb = 0; kb < 4; ++kb) tma_load_2d(sbase + kFOffW + kb * 8192, &tma_w, wfull(), kb * 64, 0);
      int it = 0;
      for (int row = row_lo; row < row_hi; ++row) {
        const int n = row / kS2dOW, oh = row - n * kS2dOW;
        for (int kh = 0; kh < 4; ++kh, ++it) {
          const int s = it % kFStages;
          mbar_wait(empty(s), ((it / kFStages) & 1) ^ 1);
          mbar_expect_tx(full(s), kS2dBlock);
          tma_load_4d(sbase + kFOffA + s * kS2dBlock, &tma_s, full(s), 0, 0, oh + kh, n);
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: 4 K-blocks x 4 K-slices per output row, M = 128 pixels, N = 64 channels
    const uint32_t el = elect_one();
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    mbar_wait(wfull(), 0);
    tcgen05_fence_after();
    const uint64_t bd0 = make_desc<0>(sbase + kFOffW);
    int it = 0, iter = 0;
    for (int row = row_lo; row < row_hi; ++row, ++iter) {
      const int b = iter & 1;
      mbar_wait(tempty(b), ((iter >> 1) & 1) ^ 1);
      tcgen05_fence_after();
      for (int kh = 0; kh < 4; ++kh, ++it) {
        const int s = it % kFStages;
        mbar_wait(full(s), (it / kFStages) & 1);
        tcgen05_fence_after();
        const uint64_t ad0 = make_desc<0>(sbase + kFOffA + s * kS2dBlock);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16_pred(tmem_base + b * 64, ad0 + (uint64_t)(k * 2), bd0 + (uint64_t)(kh * 512 + k * 2), idesc,
                         (kh > 0 || k > 0) ? 1u : 0u, el);
        umma_commit_pred(empty(s), el);
      }
      umma_commit_pred(tfull(b), el);
    }
  } else if (warp >= 4) {
    // ---------------- epilogue: TMEM -> bf16 -> swizzled staging -> one TMA store of 112 pixels; BN statistics
    const int q = warp & 3;
    const int m = q * 32 + lane;
    const bool valid = m < kS2dOW;
    const bool leader = warp == 4 && lane == 0;
    float st[4] = {0.f, 0.f, 0.f, 0.f};
    int iter = 0;
    for (int row = row_lo; row < row_hi; ++row, ++iter) {
      const int b = iter & 1;
      mbar_wait(tfull(b), (iter >> 1) & 1);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + b * 64;
      if (leader) tma_store_wait_read<1>();     // staging tile b was handed to TMA two rows ago
      s2d_epi_sync();
      uint8_t* stg = smem + kFOffStg + b * 16384;
#pragma unroll
      for (int hc = 0; hc < 2; ++hc) {
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
      s2d_epi_sync();
      if (leader) {
        tma_store_2d(&tma_y, sbase + kFOffStg + b * 16384, 0, row * kS2dOW);
        tma_store_commit();
        mbar_arrive(tempty(b));
      }
      if (p.stats) {      // lane l: columns 2l, 2l+1 over this warp's valid rows of the stored (bf16) tile
        const int nrows = min(32, kS2dOW - q * 32);
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
      s2d_epi_sync();                                    // every TMA store has read the staging tiles
      float* sm = reinterpret_cast<float*>(smem + kFOffStg);
      sm[(q * 2 + 0) * 64 + 2 * lane] = st[0]; sm[(q * 2 + 0) * 64 + 2 * lane + 1] = st[1];
      sm[(q * 2 + 1) * 64 + 2 * lane] = st[2]; sm[(q * 2 + 1) * 64 + 2 * lane + 1] = st[3];
      s2d_epi_sync();
      const int t = q * 32 + lane;                       // 128 threads: 64 sums + 64 sums of squares
      const int k = t >> 6, col = t & 63;
      float acc = 0.f;
#pragma unroll
      for (int w = 0; w < 4; ++w) acc += sm[(w * 2 + k) * 64 + col];
      p.stats[(size_t)blockIdx.x * 128 + k * 64 + col] = acc;
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
__device__ __forceinline__ uint64_t s2d_desc_mn(uint32_t smem_addr, uint32_t lbo) {   // MN-major, 128B swizzle
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((1024 >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// shared-memory map of the wgrad kernel: stages of HALF a tile (64 pixel rows of the four K-blocks, contiguous — the
// 64-wide M atoms of the A operand are one block apart — and the 64 dy rows that go with them).  Whole tiles left room
// for two buffers only, and an 80 KB load could not start before the MMA two tiles back had finished.
constexpr int kWStages = 5;
constexpr int kWHalf = 64 * 128;                          // one K-block of a half tile: 64 pixel rows x 128 B
constexpr int kWStageBytes = 5 * kWHalf;                  // four K-blocks + the dy rows
constexpr int kWOffBar = kWStages * kWStageBytes;
constexpr int kWSmem = kWOffBar + 256 + 1024;

__global__ void __launch_bounds__(kS2dThreads, 1)
stem_s2d_wgrad_kernel(const __grid_constant__ CUtensorMap tma_s,    // S as in the forward kernel, box {64, 64 px}
                      const __grid_constant__ CUtensorMap tma_dy,   // dy [rows*112][64], box {64, 64}
                      const S2dParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar = sbase + kWOffBar;
  auto full = [&](int b) { return bar + 8u * b; };
  auto empty = [&](int b) { return bar + 8u * (kWStages + b); };
  auto done = [&]() { return bar + 8u * (2 * kWStages); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + kWOffBar + 8 * (2 * kWStages + 1));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 128;     // D^T: two [128, 64] halves

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_s) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_dy) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int b = 0; b < kWStages; ++b) {
      mbar_init(full(b), 1);
      mbar_init(empty(b), 1);
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
  const int row_lo = min((int)blockIdx.x * p.rows_per_cta, p.rows), row_hi = min(row_lo + p.rows_per_cta, p.rows);

  if (warp == 0) {
    // Pixels 112..127 of the second half are outside the image map and arrive as zeros; the dy rows that go with them
    // belong to the next output row (or are outside the tensor: zeros) and meet only those zeros.
    if (lane == 0) {
      int it = 0;
      for (int row = row_lo; row < row_hi; ++row) {
        const int n = row / kS2dOW, oh = row - n * kS2dOW;
        for (int half = 0; half < 2; ++half, ++it) {
          const int b = it % kWStages;
          const uint32_t dst = sbase + b * kWStageBytes;
          mbar_wait(empty(b), ((it / kWStages) & 1) ^ 1);
          mbar_expect_tx(full(b), kWStageBytes);
          for (int kh = 0; kh < 4; ++kh) tma_load_4d(dst + kh * kWHalf, &tma_s, full(b), 0, half * 64, oh + kh, n);
          tma_load_2d(dst + 4 * kWHalf, &tma_dy, full(b), 0, row * kS2dOW + half * 64);
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t el = elect_one();
    // The gradient is accumulated TRANSPOSED, D[256 space-to-depth columns, 64 output channels], as two M = 128
    // halves: A = two K-blocks of the stage (MN-major, 64-wide atoms one block apart), B = the dy rows (MN-major,
    // N = 64), K = 16 pixel rows per instruction.  (With M = 64 output channels the tensor core runs at half rate.)
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                               ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    int it = 0;
    for (int row = row_lo; row < row_hi; ++row) {
      for (int half = 0; half < 2; ++half, ++it) {
        const int b = it % kWStages;
        const uint32_t src = sbase + b * kWStageBytes;
        mbar_wait(full(b), (it / kWStages) & 1);
        tcgen05_fence_after();
        const uint64_t bd0 = s2d_desc_mn(src + 4 * kWHalf, 0);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint64_t ad0 = s2d_desc_mn(src + h * 2 * kWHalf, kWHalf);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16_pred(tmem_base + h * 64, ad0 + (uint64_t)(k * 128), bd0 + (uint64_t)(k * 128), idesc,
                           (k > 0 || it > 0) ? 1u : 0u, el);
        }
        umma_commit_pred(empty(b), el);
      }
    }
    umma_commit_pred(done(), el);
  } else if (warp >= 4) {
    const int q = warp & 3;
    mbar_wait(done(), 0);
    tcgen05_fence_after();
    // TMEM lane = space-to-depth column within the half, 64 columns = output channels: partial[g][k2][co]
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float* dst = p.partial + ((size_t)blockIdx.x * kS2dK + h * 128 + q * 32 + lane) * 64;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + h * 64 + c0, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(dst + c0 + 4 * j) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
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

// ------------------------------------------------------------------ host side
typedef CUresult (*S2dEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static S2dEncodeFn s2d_encode() {
  static S2dEncodeFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (S2dEncodeFn)ptr;
    else
      cudaGetLastError();
  }
  return fn;
}

static int s2d_map(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides,
                   const cuuint32_t* box, const char* what) {
  S2dEncodeFn enc = s2d_encode();
  if (!enc) {
    kdl_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return KDL_ERR_UNSUPPORTED;
  }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims, strides, box,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    kdl_set_error("stem s2d: cuTensorMapEncodeTiled(%s) failed (%d)", what, (int)r);
    return KDL_ERR_CUDA;
  }
  return KDL_OK;
}

// S [N][115][115][16] seen as windows of four pixels: {64 elements, 112 window starts (32 B apart), 115 rows, N}
static int s2d_image_map(CUtensorMap* map, const void* s2d, int N, uint32_t box_pixels = 128) {
  const cuuint64_t dims[4] = {64, (cuuint64_t)kS2dOW, (cuuint64_t)kS2dP, (cuuint64_t)N};
  const cuuint64_t strides[3] = {(cuuint64_t)kS2dC * 2, (cuuint64_t)kS2dP * kS2dC * 2,
                                 (cuuint64_t)kS2dP * kS2dP * kS2dC * 2};
  const cuuint32_t box[4] = {64, box_pixels, 1, 1};
  return s2d_map(map, s2d, 4, dims, strides, box, "space-to-depth image");
}

static int s2d_matrix_map(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint32_t box_c,
                          uint32_t box_r, const char* what) {
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {cols * 2};
  const cuuint32_t box[2] = {box_c, box_r};
  return s2d_map(map, base, 2, dims, strides, box, what);
}

static int s2d_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

static int s2d_fill(S2dParams* p, int N, int H, int W) {
  if (H != kS2dIn || W != kS2dIn || N < 1) {
    kdl_set_error("stem s2d: specialised for %dx%d images (got %dx%d, N=%d)", kS2dIn, kS2dIn, H, W, N);
    return KDL_ERR_UNSUPPORTED;
  }
  p->rows = N * kS2dOW;
  p->rows_per_cta = (p->rows + s2d_sms() - 1) / s2d_sms();
  return KDL_OK;
}

}  // namespace kdl

using namespace kdl;

extern "C" size_t kdl_stem_s2d_image_bytes(int N) { return (size_t)N * kS2dP * kS2dP * kS2dC * 2; }
extern "C" size_t kdl_stem_s2d_wgrad_scratch_bytes(void) { return (size_t)s2d_sms() * kS2dK * 64 * sizeof(float); }

// s2d [N][115][115][16] bf16 = zero-bordered 2x2 space-to-depth of normalise(img_u8 [N,224,224,3])
extern "C" int kdl_stem_s2d_pack(const void* img_u8, void* s2d, int N, int H, int W, const float* mean3,
                                 const float* std3, cudaStream_t stream) {
  S2dParams p = {};
  int rc = s2d_fill(&p, N, H, W);
  if (rc) return rc;
  const int64_t total = (int64_t)N * kS2dP * kS2dP;
  if (total >= ((int64_t)1 << 31)) {
    kdl_set_error("stem s2d: batch too large (N=%d)", N);
    return KDL_ERR_UNSUPPORTED;
  }
  stem_s2d_pack_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
      (const uint8_t*)img_u8, (bf16*)s2d, N, mean3[0], mean3[1], mean3[2], 1.f / std3[0], 1.f / std3[1], 1.f / std3[2]);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

// y [N,112,112,64] = conv7x7/s2/p3 of the image behind `s2d`; w_padded is [64][168] in the kh*24 + kw*3 + c column
// layout, w2_scratch receives its [64][256] space-to-depth form.  stats_partial (optional):
// [kdl_stem_conv_max_parts()][2][64] floats, *stats_parts_out rows written — feed kdl_bn_fwd_from_stats.
extern "C" int kdl_stem_s2d_fwd(const void* s2d, const void* w_padded, void* w2_scratch, void* y, int N, int H, int W,
                                float* stats_partial, int* stats_parts_out, cudaStream_t stream) {
  S2dParams p = {};
  int rc = s2d_fill(&p, N, H, W);
  if (rc) return rc;
  p.stats = stats_partial;
  stem_s2d_weight_kernel<<<(64 * kS2dK + 255) / 256, 256, 0, stream>>>((const bf16*)w_padded, (bf16*)w2_scratch);
  KDL_CHECK_LAUNCH();
  CUtensorMap ms, mw, my;
  if ((rc = s2d_image_map(&ms, s2d, N))) return rc;
  if ((rc = s2d_matrix_map(&mw, w2_scratch, kS2dK, 64, 64, 64, "filter"))) return rc;
  if ((rc = s2d_matrix_map(&my, y, 64, (uint64_t)p.rows * kS2dOW, 64, kS2dOW, "output"))) return rc;
  static bool attr = false;
  if (!attr) {
    KDL_CUDA(cudaFuncSetAttribute(stem_s2d_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFSmem));
    attr = true;
  }
  const int grid = (p.rows + p.rows_per_cta - 1) / p.rows_per_cta;
  if (stats_parts_out) *stats_parts_out = grid;
  stem_s2d_fprop_kernel<<<grid, kS2dThreads, kFSmem, stream>>>(ms, mw, my, p);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

// gw [64][168] (padded filter layout) = wgrad of the stem: sum over pixels of dy[p, co] * patch[p, k]
extern "C" int kdl_stem_s2d_wgrad(const void* s2d, const void* dy, void* gw_padded, int N, int H, int W, void* scratch,
                                  size_t scratch_bytes, cudaStream_t stream) {
  S2dParams p = {};
  int rc = s2d_fill(&p, N, H, W);
  if (rc) return rc;
  const int grid = (p.rows + p.rows_per_cta - 1) / p.rows_per_cta;
  const size_t need = (size_t)grid * kS2dK * 64 * sizeof(float);
  if (scratch == nullptr || scratch_bytes < need) {
    kdl_set_error("stem s2d wgrad: scratch too small (%zu < %zu)", scratch_bytes, need);
    return KDL_ERR_WORKSPACE;
  }
  p.partial = (float*)scratch;
  CUtensorMap ms, mdy;
  if ((rc = s2d_image_map(&ms, s2d, N, 64))) return rc;
  if ((rc = s2d_matrix_map(&mdy, dy, 64, (uint64_t)p.rows * kS2dOW, 64, 64, "dy"))) return rc;
  static bool attr = false;
  if (!attr) {
    KDL_CUDA(cudaFuncSetAttribute(stem_s2d_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWSmem));
    attr = true;
  }
  stem_s2d_wgrad_kernel<<<grid, kS2dThreads, kWSmem, stream>>>(ms, mdy, p);
  KDL_CHECK_LAUNCH();
  stem_s2d_wgrad_reduce_kernel<<<(64 * kS2dK + 255) / 256, 256, 0, stream>>>(p.partial, (bf16*)gw_padded, grid);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}
