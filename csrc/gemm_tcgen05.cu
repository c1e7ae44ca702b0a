// N6: bf16 GEMM on the 5th-generation tensor cores.
//
//   C[M,N] = op(A) * op(B)  (+ bias[N]) (+ addend[M,N]) (ReLU?)      bf16 in, fp32 accumulate
//
// Persistent, warp-specialised kernel, one CTA per SM:
//   warp 0   TMA producer   cp.async.bulk.tensor (128B-swizzled tiles) -> smem ring, mbarrier tx
//   warp 1   MMA issuer     one thread issues tcgen05.mma (M=128, N=BN, K=16) from smem
//                           descriptors into a TMEM accumulator; tcgen05.commit frees the stage
//   warp 2   TMEM allocator tcgen05.alloc / dealloc (2 accumulator stages of BN columns)
//   warp 4.. epilogue       4 or 8 warps: tcgen05.ld TMEM -> registers -> bias / addend (a TMA box in the staging
//                           buffer) / ReLU / BN column statistics -> bf16 -> swizzled smem -> TMA store
// so the epilogue of tile i overlaps the loads and MMAs of tile i+1.  Operands may be K-major
// or MN-major (the transposes needed by dgrad / wgrad are done by the UMMA descriptors, not by
// copies): nt = fwd (x W^T), nn = dgrad (dy W), tn = wgrad (dy^T x).  wgrad has a tiny output
// and a huge K, so it is split over K across CTAs; partial tiles go to an fp32 scratch and a
// second kernel reduces them in a fixed order (deterministic) and casts to bf16.
//
// Replaces the cuBLAS/cuDNN GEMMs the stock worker would call for ResNet-50's 1x1 convolutions
// and FC layer and for the DIN MLP; the reference operator itself has no math at all
// (reference example/pytorch/pytorch_job_mnist_mpi.yaml:15 just names a worker image).
#include "common.cuh"
#include "tcgen05.cuh"
#include "../include/kubedl_b200.h"
#include <cuda.h>
#include <string.h>
#include <stdlib.h>

namespace kdl {

constexpr int BM = 128;       // UMMA M (cta_group::1)
// 4 control warps + EPI epilogue warps (template parameter, 4 or 8).  With 8, two warps share a TMEM lane
// quarter and take alternate 64-column chunks.  One epilogue warp per SM sub-partition runs a long serial
// chain per chunk (TMEM load, addend / statistics, 64 conversions, pack, store) with nobody to hide its
// latencies behind, so tiles with little K per output element (K <= 256..512: the memory-bound 1x1
// convolutions, and everything with a fused addend or fused statistics) are EPILOGUE-bound with 4 warps;
// 8 warps cost 32 KB of staging, i.e. one pipeline stage, which only the deep-K compute-bound shapes miss
// (profiles/gemm_vs_cublas_shapes.log, profiles/r2_addend_epilogue.md).  The host picks per launch.
constexpr int kEpiWarp0 = 4;
constexpr int gemm_threads(int epi) { return 128 + 32 * epi; }

struct GemmParams {
  bf16* C;              // bf16 output (mode 0)
  float* partial;       // fp32 split-K partials [splits][M][N] (mode 1)
  const bf16* bias;     // [N] or null
  const bf16* addend;   // [M, ldc] or null
  int add_h, add_w;     // > 0: rows are pixels (n, h, w) of an add_h x add_w map and `addend` is the compact
                        // [n, add_h/2, add_w/2] tensor of a stride-2 branch: added at even (h, w) only
  int add_tma;              // the (full-size) addend is fetched by TMA into the output staging buffers (tensor map tma_d)
  const uint8_t* add_mask;  // optional [M*N/8] bit mask (bit j of byte v <-> element 8v+j): the addend counts only
                            // where the bit is set (gradient through a ReLU whose mask BatchNorm saved)
  int M, N, K;
  int ldc;
  int relu;
  int splits;
  int kb_per_split;
  int num_m, num_n;
  float* stats;   // optional [parts][2][N]: per epilogue-warp column sum / sum-of-squares of the bf16 output
  // implicit-GEMM convolution (N7): one operand is fetched with TMA im2col loads
  int conv_mode;     // 0 plain GEMM; 1 A = activation patches (fprop / dgrad); 2 B = activation patches (wgrad)
  int conv_q, conv_p;        // spatial extent the pixel index runs over (output W, H of this pass)
  int conv_stride, conv_pad; // traversal stride and the (effective) padding of this pass
  int conv_s, conv_r;        // filter W, H
  int conv_c;                // channels of the patch operand (K per tap for mode 1, N per tap for mode 2)
  int conv_flip;             // dgrad: filter taps are visited mirrored
  // parity class of a STRIDE-2 dgrad run as a stride-1 sub-convolution over dy (kdl_conv2d_dgrad_s2):
  int sub_en;                // 1: the "split" index of a tile is its parity class c = 2 ph + pw (all four classes in one
                             // persistent launch): 1 or 2 taps per axis, weight tap of sub-tap t = (sub_wtaps[c] >> 4t) & 15,
                             // and row (n, a, b) is stored at input pixel (2a + ph, 2b + pw)
  int sub_wtaps[4];          // per class: up to four original filter-tap indices, 4 bits each
};

// K-block range of a tile: a split-K slice, or (sub_en) all taps of the tile's parity class
__device__ __forceinline__ void tile_kb_range(const GemmParams& p, int split, int kb_total, int& kb0, int& kb1) {
  if (p.sub_en) {
    const int taps = ((split >> 1) ? 2 : 1) * ((split & 1) ? 2 : 1);
    kb0 = 0;
    kb1 = taps * (p.conv_c / BK);
  } else {
    kb0 = split * p.kb_per_split;
    kb1 = min(kb0 + p.kb_per_split, kb_total);
  }
}

constexpr int kStatsMaxN = 2048;
constexpr int kStatsBytes = 2 * kStatsMaxN * 4;

// Column statistics of one 32-row x 64-column bf16 chunk sitting in a (128B-swizzled) staging
// buffer: lane l owns columns 2l, 2l+1 and keeps their running sum / sum of squares in REGISTERS for
// the whole persistent loop (acc[0..1] sums, acc[2..3] squares).  That works because the grid is
// sized as a multiple of the number of column tiles when statistics are requested, so a CTA sees the
// same columns in every tile it processes.  (A first version accumulated with shared-memory
// atomicAdd(float) — a compare-and-swap loop in SASS — and lost more time than the saved pass.)
// Feeds the BatchNorm that follows a 1x1 convolution: the separate statistics pass over the
// activation disappears.
__device__ __forceinline__ void chunk_col_stats(const uint8_t* buf, int lane, int rows_valid, float* acc) {
  float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
  if (rows_valid >= 32) {
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      const uint32_t w = *reinterpret_cast<const uint32_t*>(buf + r * 128 + ((((lane >> 2) ^ (r & 7))) << 4) + (lane & 3) * 4);
      const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w));
      s0 += v.x; q0 = fmaf(v.x, v.x, q0);
      s1 += v.y; q1 = fmaf(v.y, v.y, q1);
    }
  } else {
    for (int r = 0; r < rows_valid; ++r) {
      const uint32_t w = *reinterpret_cast<const uint32_t*>(buf + r * 128 + ((((lane >> 2) ^ (r & 7))) << 4) + (lane & 3) * 4);
      const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w));
      s0 += v.x; q0 = fmaf(v.x, v.x, q0);
      s1 += v.y; q1 = fmaf(v.y, v.y, q1);
    }
  }
  acc[0] += s0; acc[1] += s1; acc[2] += q0; acc[3] += q1;
}

// part row layout of the statistics buffer: [nparts][2][N]
__device__ __forceinline__ void store_col_stats(float* stats, int part, int N, int col, const float* acc) {
  float* dst = stats + (size_t)part * 2 * N;
  if (col < N) { dst[col] = acc[0]; dst[N + col] = acc[2]; }
  if (col + 1 < N) { dst[col + 1] = acc[1]; dst[N + col + 1] = acc[3]; }
}

__device__ __forceinline__ void tma_prefetch_l2_2d(const CUtensorMap* map, int c0, int c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];" ::"l"(map), "r"(c0), "r"(c1) : "memory");
}

// The residual-branch gradient ("addend") of a dgrad GEMM, one accumulator row per lane, 64 columns per chunk.
// Per-lane loads of "my row" are the wrong shape for the memory system: 16 bytes per lane at a row pitch of
// 512+ bytes makes every LDG.128 touch 32 different lines (256 L1 wavefronts per 4 KB chunk, four warps
// deep), and the memory-bound dgrads ran at half of copy bandwidth even with the loads issued a chunk
// ahead (profiles/r2_addend_epilogue.md).  So a full-size addend comes in by TMA instead (see the
// epilogue); this register pipeline remains for the compact stride-2 addend, whose rows are not a box.
struct AddendPipe {
  uint4 v[8];
  uint2 m;
  const bf16* slow;   // partial N tile: element-wise path at apply time
  int state;          // 0 nothing to add, 1 registers loaded, 2 slow path

  static __device__ __forceinline__ const bf16* row_ptr(const GemmParams& p, int row) {
    if (p.addend == nullptr || row >= p.M) return nullptr;
    size_t idx = (size_t)row;
    if (p.add_h > 0) {      // gradient of a 1x1 / stride-2 branch, kept compact (never up-sampled)
      const int hw = p.add_h * p.add_w;
      const int ni = row / hw, rem = row - ni * hw, hh = rem / p.add_w, ww = rem - hh * p.add_w;
      if (((hh | ww) & 1) != 0) return nullptr;
      idx = ((size_t)ni * (p.add_h >> 1) + (hh >> 1)) * (p.add_w >> 1) + (ww >> 1);
    }
    return p.addend + idx * p.ldc;
  }
  // L2 prefetch of `ncols` columns of this lane's row (128-byte lines)
  static __device__ __forceinline__ void prefetch_l2(const GemmParams& p, int row, int col, int ncols) {
    const bf16* r = row_ptr(p, row);
    if (r == nullptr) return;
    const int end = min(col + ncols, p.N);
    for (int c = col; c < end; c += 64) asm volatile("prefetch.global.L2 [%0];" ::"l"(r + c));
    if (p.add_mask) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.add_mask + (((size_t)row * p.N + col) >> 3)));
  }
  __device__ __forceinline__ void issue(const GemmParams& p, int row, int col) {
    state = 0;
    const bf16* r = row_ptr(p, row);
    if (r == nullptr || col >= p.N) return;
    if (col + 64 <= p.N) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = ld_plain(r + col + 8 * j);
      if (p.add_mask) m = *reinterpret_cast<const uint2*>(p.add_mask + (((size_t)row * p.N + col) >> 3));
      state = 1;
    } else {
      slow = r + col;
      state = 2;
    }
  }
  __device__ __forceinline__ void apply(const GemmParams& p, int col, float* f) const {
    if (state == 1) {
      if (p.add_mask) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float b[8];
          unpack8(v[j], b);
          const uint32_t bits = (j < 4 ? m.x : m.y) >> ((j & 3) * 8);
#pragma unroll
          for (int i = 0; i < 8; ++i) f[8 * j + i] += ((bits >> i) & 1u) ? b[i] : 0.f;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float b[8];
          unpack8(v[j], b);
#pragma unroll
          for (int i = 0; i < 8; ++i) f[8 * j + i] += b[i];
        }
      }
    } else if (state == 2) {
#pragma unroll
      for (int j = 0; j < 64; ++j)
        if (col + j < p.N) f[j] += __bfloat162float(slow[j]);
    }
  }
};

// Full-size addend by TMA.  The 32-row x 64-column addend chunk lands (128B-swizzled, by the same box shape
// the output store uses) in the very staging buffer the output chunk is about to be written into: lane l
// reads ITS row with eight conflict-free LDS.128, adds it to its accumulator row and later overwrites the
// same 128 bytes with the result, so the addend needs no shared memory of its own.  The load for chunk i+1
// is issued into the other staging buffer while chunk i is converted and stored (after the store that last
// used that buffer has finished reading it), and a whole tile is prefetched into L2 one tile ahead.
struct AddendTma {
  uint32_t phases;     // bit b: parity of the next completion of barrier b
  uint2 mnext;         // ReLU-mask bits of the chunk whose load is in flight

  static __device__ __forceinline__ uint2 load_mask(const GemmParams& p, int row, int col) {
    if (p.add_mask == nullptr || row >= p.M || col >= p.N) return make_uint2(0u, 0u);
    return *reinterpret_cast<const uint2*>(p.add_mask + (((size_t)row * p.N + col) >> 3));
  }
  // lane 0 issues; every lane fetches its own mask word
  __device__ __forceinline__ void issue(const GemmParams& p, const CUtensorMap* map, uint32_t bar, uint32_t dst,
                                        int row0, int col, int lane) {
    if (lane == 0) {
      mbar_expect_tx(bar, 4096);
      tma_load_2d(dst, map, bar, col, row0);
    }
    mnext = load_mask(p, row0 + lane, col);
  }
  // wait for buffer `b`'s chunk and add this lane's row (src = its 128 bytes in the staging buffer)
  __device__ __forceinline__ void consume(const GemmParams& p, uint32_t bar, int b, const uint8_t* src, int lane,
                                          float* f) {
    mbar_wait(bar, (phases >> b) & 1u);
    phases ^= 1u << b;
    const uint2 m = mnext;
    const bool masked = p.add_mask != nullptr;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float v[8];
      unpack8(*reinterpret_cast<const uint4*>(src + ((j ^ (lane & 7)) << 4)), v);
      const uint32_t bits = masked ? ((j < 4 ? m.x : m.y) >> ((j & 3) * 8)) : 0xFFu;
#pragma unroll
      for (int i = 0; i < 8; ++i) f[8 * j + i] += ((bits >> i) & 1u) ? v[i] : 0.f;
    }
  }
};

template <int A_MN, int B_MN, int BN>
__host__ __device__ constexpr uint32_t make_idesc() {
  return (1u << 4)            // D format f32
         | (1u << 7)          // A bf16
         | (1u << 10)         // B bf16
         | ((uint32_t)A_MN << 15) | ((uint32_t)B_MN << 16) | ((uint32_t)(BN >> 3) << 17) |
         ((uint32_t)(BM >> 4) << 24);
}

// The four epilogue warps of a CTA hold statistics of the same columns over different rows: add them in
// shared memory (fixed order) and write ONE partial row per CTA, so the finalizer walks a quarter of the rows.
template <int BNT, int EPI>
__device__ __forceinline__ void reduce_store_col_stats(float* sm, float* stats, int part, int N, int col0, int ew,
                                                       int lane, float (*acc)[4]) {
#pragma unroll
  for (int c = 0; c < BNT / 64; ++c) {
    const int col = c * 64 + 2 * lane;
    sm[(ew * 2 + 0) * BNT + col] = acc[c][0]; sm[(ew * 2 + 0) * BNT + col + 1] = acc[c][1];
    sm[(ew * 2 + 1) * BNT + col] = acc[c][2]; sm[(ew * 2 + 1) * BNT + col + 1] = acc[c][3];
  }
  asm volatile("bar.sync 1, %0;" ::"n"(EPI * 32) : "memory");
  float* dst = stats + (size_t)part * 2 * N;
  for (int col = ew * 32 + lane; col < BNT; col += EPI * 32) {
    float s0 = 0.f, q0 = 0.f;
#pragma unroll
    for (int w = 0; w < EPI; ++w) { s0 += sm[(w * 2 + 0) * BNT + col]; q0 += sm[(w * 2 + 1) * BNT + col]; }
    if (col0 + col < N) { dst[col0 + col] = s0; dst[N + col0 + col] = q0; }
  }
}

template <int BN, int STAGES, int EPI>
struct SmemLayout {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagingOffset = STAGES * kStageBytes;   // 1024-aligned
  static constexpr int kStagingPerWarp = 2 * 4096;             // two 32-row x 128-B buffers
  static constexpr int kStatsOffset = kStagingOffset + EPI * kStagingPerWarp;
  static constexpr int kBarOffset = kStatsOffset + kStatsBytes;
  static constexpr int kAddBarOffset = kBarOffset + (2 * STAGES + 4) * 8 + 16;   // 2 per epilogue warp (addend loads)
  static constexpr int kTotal = kAddBarOffset + EPI * 2 * 8;
  static_assert(kTotal + 1024 <= 232448, "shared memory budget");
};

// ============================== epilogue (shared by the 1-CTA and the CTA-pair kernel) ==============================
// TMEM -> registers -> (bias, addend, ReLU) -> bf16 -> 128B-swizzled smem staging -> TMA store.  Each warp
// owns 32 accumulator rows and two 4 KB staging buffers, so the bulk store of chunk i overlaps the TMEM read
// + math of chunk i+1; TMA clips M/N tails.  TWO: the CTA is one half of a cta_group::2 pair (tile rows
// m_idx*256 + cta*128, "accumulator drained" is reported to the leader CTA's barrier).
template <int BNT, int EPI, bool TWO, class SL>
__device__ __forceinline__ void gemm_epilogue(const GemmParams& p, const CUtensorMap* tma_c, const CUtensorMap* tma_d,
                                              uint8_t* smem, uint32_t smem_base, uint32_t tmem_base, uint32_t tfull0,
                                              uint32_t tempty0, int warp, int lane, int cta, int first_tile,
                                              int tile_stride, int total_tiles) {
  constexpr int kRowsPerTile = TWO ? 256 : BM;
  const int q = warp & 3;                   // TMEM lane quarter this warp may access
  const int ew = warp - kEpiWarp0;
  const int half = ew >> 2;                 // the two warps of a quarter take alternate column chunks (EPI == 8)
  const uint32_t stg_base = smem_base + SL::kStagingOffset + ew * SL::kStagingPerWarp;
  uint8_t* stg_ptr = smem + SL::kStagingOffset + ew * SL::kStagingPerWarp;
  int iter = 0;
  int buf = 0;
  float st_acc[BNT / 64][4];
#pragma unroll
  for (int c = 0; c < BNT / 64; ++c) st_acc[c][0] = st_acc[c][1] = st_acc[c][2] = st_acc[c][3] = 0.f;
  constexpr int kChunkStep = 64 * (EPI / 4);
  auto tile_row0 = [&](int tt) { return ((tt / p.num_n) % p.num_m) * kRowsPerTile + cta * BM + q * 32; };
  auto tile_col = [&](int tt) { return (tt % p.num_n) * BNT; };
  const bool add_tma = p.addend != nullptr && p.add_tma != 0;
  const bool add_reg = p.addend != nullptr && p.add_tma == 0;
  const uint32_t abar0 = smem_base + SL::kAddBarOffset + 16u * ew;     // this warp's two addend barriers
  AddendPipe ap;
  ap.state = 0;
  AddendTma at;
  at.phases = 0u;
  at.mnext = make_uint2(0u, 0u);
  if (first_tile < total_tiles) {
    if (add_reg) ap.issue(p, tile_row0(first_tile) + lane, tile_col(first_tile) + half * 64);
    if (add_tma) at.issue(p, tma_d, abar0, stg_base, tile_row0(first_tile), tile_col(first_tile) + half * 64, lane);
  }
  for (int t = first_tile; t < total_tiles; t += tile_stride, ++iter) {
    const int n_idx = t % p.num_n;
    const int split = t / (p.num_n * p.num_m);
    const int as = iter & 1;
    const uint32_t aphase = (iter >> 1) & 1;
    const int t_next = t + tile_stride;
    if (t_next < total_tiles) {      // next tile's addend -> L2 while this tile is worked on
      if (add_reg) AddendPipe::prefetch_l2(p, tile_row0(t_next) + lane, tile_col(t_next), BNT);
      if (add_tma && lane == 0) {
#pragma unroll
        for (int c = half * 64; c < BNT; c += kChunkStep) tma_prefetch_l2_2d(tma_d, tile_col(t_next) + c, tile_row0(t_next));
      }
    }
    mbar_wait(tfull0 + 8u * as, aphase);
    tcgen05_fence_after();
    const int row0 = tile_row0(t);
    const int row = row0 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BNT;
    if (p.partial != nullptr) {
      // split-K: fp32 partial tile, 32 columns (128 B) per staging row
#pragma unroll 1
      for (int c0 = half * 32; c0 < BNT; c0 += 32 * (EPI / 4)) {
        uint32_t r[32];
        tmem_ld32(taddr + c0, r);
        tmem_ld_wait();
        if (lane == 0) tma_store_wait_read<1>();
        __syncwarp();
        uint8_t* dst = stg_ptr + buf * 4096 + lane * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(dst + ((j ^ (lane & 7)) << 4)) =
              make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_3d(tma_c, stg_base + buf * 4096, n_idx * BNT + c0, row0, split);
          tma_store_commit();
        }
        buf ^= 1;
      }
    } else {
#pragma unroll 1
      for (int c0 = half * 64; c0 < BNT; c0 += kChunkStep) {
        uint32_t r[64];
        tmem_ld32(taddr + c0, r);
        tmem_ld32(taddr + c0 + 32, r + 32);
        tmem_ld_wait();
        const int col = n_idx * BNT + c0;
        float f[64];
#pragma unroll
        for (int j = 0; j < 64; ++j) f[j] = __uint_as_float(r[j]);
        if (p.bias) {
          if (col + 64 <= p.N) {
#pragma unroll
            for (int j = 0; j < 64; j += 8) {
              float b[8];
              unpack8(*reinterpret_cast<const uint4*>(p.bias + col + j), b);
#pragma unroll
              for (int i = 0; i < 8; ++i) f[j + i] += b[i];
            }
          } else {
#pragma unroll
            for (int j = 0; j < 64; ++j)
              if (col + j < p.N) f[j] += __bfloat162float(p.bias[col + j]);
          }
        }
        if (add_reg) {
          ap.apply(p, col, f);
          // registers for the next chunk (of this tile, else the first chunk of the next tile) go out now
          if (c0 + kChunkStep < BNT) ap.issue(p, row, col + kChunkStep);
          else if (t_next < total_tiles) ap.issue(p, tile_row0(t_next) + lane, tile_col(t_next) + half * 64);
        }
        if (add_tma) {
          at.consume(p, abar0 + 8u * buf, buf, stg_ptr + buf * 4096 + lane * 128, lane, f);
          // the next chunk's addend goes into the other buffer once the store that last used it has read it
          const bool in_tile = c0 + kChunkStep < BNT;
          if (in_tile || t_next < total_tiles) {
            if (lane == 0) tma_store_wait_read<0>();
            at.issue(p, tma_d, abar0 + 8u * (buf ^ 1), stg_base + (buf ^ 1) * 4096,
                     in_tile ? row0 : tile_row0(t_next), in_tile ? col + kChunkStep : tile_col(t_next) + half * 64, lane);
          }
          __syncwarp();
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 64; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        if (p.sub_en) {
          // parity class of a stride-2 dgrad: GEMM row (n, a, b) is input pixel (2a + ph, 2b + pw) — rows of a
          // chunk are not a TMA box in dx, so every lane stores its own 128 bytes
          if (row < p.M) {
            const int pq = p.conv_p * p.conv_q;
            const int ni = row / pq, rem = row - ni * pq, aa = rem / p.conv_q, bb = rem - aa * p.conv_q;
            bf16* o = p.C + ((((size_t)ni * (2 * p.conv_p) + 2 * aa + (split >> 1)) * (2 * p.conv_q)) + 2 * bb + (split & 1)) * p.ldc + col;
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (col + 8 * j + 8 <= p.N) *reinterpret_cast<uint4*>(o + 8 * j) = pack8(f + 8 * j);
          }
          continue;
        }
        if (lane == 0) tma_store_wait_read<1>();  // the buffer written two chunks ago is free
        __syncwarp();
        uint8_t* dst = stg_ptr + buf * 4096 + lane * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(dst + ((j ^ (lane & 7)) << 4)) = pack8(f + 8 * j);
        if (p.stats) {
          __syncwarp();
          float t4[4] = {0.f, 0.f, 0.f, 0.f};
          chunk_col_stats(stg_ptr + buf * 4096, lane, p.M - row0, t4);
#pragma unroll
          for (int c = 0; c < BNT / 64; ++c)      // compile-time indices keep st_acc in registers
            if (c == (c0 >> 6)) { st_acc[c][0] += t4[0]; st_acc[c][1] += t4[1]; st_acc[c][2] += t4[2]; st_acc[c][3] += t4[3]; }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(tma_c, stg_base + buf * 4096, col, row0);
          tma_store_commit();
        }
        buf ^= 1;
      }
    }
    tcgen05_fence_before();
    __syncwarp();
    if (lane == 0) {
      if (TWO) mbar_arrive_remote(tempty0 + 8u * as, 0);   // the leader's barrier, from both CTAs
      else mbar_arrive(tempty0 + 8u * as);
    }
  }
  if (lane == 0) tma_store_wait_read<0>();
  __syncwarp();
  if (p.stats) {
    // the column tile of this CTA (pair) is first_tile % num_n for every tile it processed; one partial row per CTA
    const int part = TWO ? (first_tile / p.num_n) * 2 + cta : first_tile / p.num_n;
    reduce_store_col_stats<BNT, EPI>(reinterpret_cast<float*>(smem + SL::kStatsOffset), p.stats, part, p.N,
                                     (first_tile % p.num_n) * BNT, ew, lane, st_acc);
  }
}

template <int A_MN, int B_MN, int BN, int STAGES, int EPI>
__global__ void __launch_bounds__(gemm_threads(EPI), 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a,
                    const __grid_constant__ CUtensorMap tma_b,
                    const __grid_constant__ CUtensorMap tma_c,
                    const __grid_constant__ CUtensorMap tma_d, const GemmParams p) {
  using SL = SmemLayout<BN, STAGES, EPI>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_base = smem_base + SL::kBarOffset;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + SL::kBarOffset + (2 * STAGES + 4) * 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;  // power of two for BN in {32,64,128,256}

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_c) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), EPI);  // one arrive per epilogue warp
    }
    for (int s = 0; s < 2 * EPI; ++s) mbar_init(smem_base + SL::kAddBarOffset + 8u * s, 1);
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

  const int total_tiles = p.num_m * p.num_n * p.splits;
  const int kb_total = (p.K + BK - 1) / BK;

  if (warp == 0) {
    // ============================== TMA producer ==============================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int n_idx = t % p.num_n;
        const int m_idx = (t / p.num_n) % p.num_m;
        const int split = t / (p.num_n * p.num_m);
        int kb0, kb1;
        tile_kb_range(p, split, kb_total, kb0, kb1);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t sa = smem_base + stage * SL::kStageBytes;
          const uint32_t sb = sa + SL::kABytes;
          mbar_expect_tx(full_bar(stage), SL::kStageBytes);
          if (p.conv_mode == 1) {
            // K runs over (tap, channel block): A = 128 output pixels x 64 channels of that tap
            const int cblks = p.conv_c / BK;
            const int tap = kb / cblks, cb = kb - tap * cblks;
            const int cs = p.sub_en ? ((split & 1) ? 2 : 1) : p.conv_s;     // taps along W of this tile's class
            int kh = tap / cs, kw = tap - kh * cs;
            const int m0 = m_idx * BM;
            const int pq = p.conv_p * p.conv_q;
            const int n0 = m0 / pq, rem = m0 - n0 * pq, oh0 = rem / p.conv_q, ow0 = rem - oh0 * p.conv_q;
            const int fh = p.conv_flip ? p.conv_r - 1 - kh : kh, fw = p.conv_flip ? p.conv_s - 1 - kw : kw;
            tma_load_im2col(sa, &tma_a, full_bar(stage), cb * BK, ow0 * p.conv_stride - p.conv_pad,
                            oh0 * p.conv_stride - p.conv_pad, n0, fw, fh);
            if (B_MN) {   // dgrad: weights [Cout][taps*Cin] read as K(cout) x N(cin) tiles of tap (kh,kw)
              const int wtap = p.sub_en ? ((p.sub_wtaps[split] >> (4 * tap)) & 15) : tap;
#pragma unroll
              for (int a = 0; a < BN / 64; ++a)
                tma_load_2d(sb + a * (BK * 128), &tma_b, full_bar(stage), wtap * p.N + n_idx * BN + a * 64, cb * BK);
            } else {      // fprop: weights [Cout][taps*Cin] are already K-major in (tap, channel) order
              tma_load_2d(sb, &tma_b, full_bar(stage), kb * BK, n_idx * BN);
            }
          } else {
          if (A_MN) {
#pragma unroll
            for (int a = 0; a < BM / 64; ++a)
              tma_load_2d(sa + a * (BK * 128), &tma_a, full_bar(stage), m_idx * BM + a * 64, kb * BK);
          } else {
            tma_load_2d(sa, &tma_a, full_bar(stage), kb * BK, m_idx * BM);
          }
          if (p.conv_mode == 2) {
            // wgrad: K runs over 64-pixel blocks; the N tile lies inside one filter tap
            const int ncol = n_idx * BN;
            const int tap = ncol / p.conv_c, c0 = ncol - tap * p.conv_c;
            const int kh = tap / p.conv_s, kw = tap - kh * p.conv_s;
            const int m0 = kb * BK;
            const int pq = p.conv_p * p.conv_q;
            const int n0 = m0 / pq, rem = m0 - n0 * pq, oh0 = rem / p.conv_q, ow0 = rem - oh0 * p.conv_q;
#pragma unroll
            for (int a = 0; a < BN / 64; ++a)
              tma_load_im2col(sb + a * (BK * 128), &tma_b, full_bar(stage), c0 + a * 64,
                              ow0 * p.conv_stride - p.conv_pad, oh0 * p.conv_stride - p.conv_pad, n0, kw, kh);
          } else if (B_MN) {
#pragma unroll
            for (int a = 0; a < BN / 64; ++a)
              tma_load_2d(sb + a * (BK * 128), &tma_b, full_bar(stage), n_idx * BN + a * 64, kb * BK);
          } else {
            tma_load_2d(sb, &tma_b, full_bar(stage), kb * BK, n_idx * BN);
          }
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer ==============================
    // The whole warp walks the loop and one elected lane issues (predicate inside the asm): with no
    // divergent branch around it ptxas keeps descriptors in uniform registers, instead of an
    // ELECT / R2UR broadcast loop in front of every UTCHMMA (which made the issue path, not the
    // tensor pipe, the limit for N <= 128 tiles).
    {
      const uint32_t el = elect_one();
      constexpr uint32_t idesc = make_idesc<A_MN, B_MN, BN>();
      int stage = 0;
      uint32_t phase = 0;
      int iter = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++iter) {
        const int split = t / (p.num_n * p.num_m);
        int kb0, kb1;
        tile_kb_range(p, split, kb_total, kb0, kb1);
        const int as = iter & 1;
        const uint32_t aphase = (iter >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_base + stage * SL::kStageBytes;
          const uint64_t ad0 = make_desc<A_MN>(sa);
          const uint64_t bd0 = make_desc<B_MN>(sa + SL::kABytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            umma_bf16_pred(tmem_d, ad0 + (uint64_t)(k_advance_bytes<A_MN>(k) >> 4),
                           bd0 + (uint64_t)(k_advance_bytes<B_MN>(k) >> 4), idesc,
                           (kb > kb0 || k > 0) ? 1u : 0u, el);
          }
          umma_commit_pred(empty_bar(stage), el);  // frees the smem stage once these MMAs retire
          if (kb == kb1 - 1) umma_commit_pred(tfull_bar(as), el);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= kEpiWarp0) {
    gemm_epilogue<BN, EPI, false, SL>(p, &tma_c, &tma_d, smem, smem_base, tmem_base, tfull_bar(0), tempty_bar(0), warp, lane,
                                      0, (int)blockIdx.x, (int)gridDim.x, total_tiles);
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(kTmemCols)
                 : "memory");
  }
}

// ------------------------------------------------------------------ 2-CTA (cta_group::2) variant
// A CTA pair (cluster of 2 = the two SMs of a TPC) computes one 256 x 256 tile with
// tcgen05.mma.cta_group::2: each CTA stages its own 128 rows of A and its own 128 columns' worth
// of B, so per-SM shared-memory traffic per FLOP halves and the MMA runs at the full UMMA_M=256
// rate (a single CTA tops out near 50 % of tensor peak: profiles/ncu_hot_kernels_raw.csv).
// The leader CTA (rank 0) issues every MMA; TMA loads of both CTAs complete on the leader's
// "full" barrier, commits are multicast to both CTAs' "empty" / "tmem full" barriers, and both
// CTAs' epilogue warps arrive on the leader's "tmem empty" barrier.
// 5 pipeline stages with 4 epilogue warps, 4 with 8 (their staging buffers take the fifth stage's 32 KB)
constexpr int stages2(int epi) { return epi == 8 ? 4 : 5; }

template <int BN2, int EPI>
struct SmemLayout2 {
  static constexpr int kStages = stages2(EPI);
  static constexpr int kABytes = BM * BK * 2;            // this CTA's 128 rows of A
  static constexpr int kBBytes = (BN2 / 2) * BK * 2;     // this CTA's half of B
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagingOffset = kStages * kStageBytes;
  static constexpr int kStagingPerWarp = 2 * 4096;
  static constexpr int kStatsOffset = kStagingOffset + EPI * kStagingPerWarp;
  static constexpr int kBarOffset = kStatsOffset + kStatsBytes;
  static constexpr int kAddBarOffset = kBarOffset + (2 * kStages + 4) * 8 + 16;
  static constexpr int kTotal = kAddBarOffset + EPI * 2 * 8;
  static_assert(kTotal + 1024 <= 232448, "shared memory budget");
};

__device__ __forceinline__ void tma_load_im2col_2sm(uint32_t dst, const CUtensorMap* map, uint32_t leader_bar,
                                                    int c, int w, int h, int n, int off_w, int off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.im2col.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
      ::"r"(dst), "l"(map), "r"(leader_bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"((uint16_t)off_w),
      "h"((uint16_t)off_h)
      : "memory");
}

template <int A_MN, int B_MN, int BN2, int EPI>
__global__ void __launch_bounds__(gemm_threads(EPI), 1)
gemm_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tma_a,
                         const __grid_constant__ CUtensorMap tma_b,
                         const __grid_constant__ CUtensorMap tma_c,
                         const __grid_constant__ CUtensorMap tma_d, const GemmParams p) {
  using SL = SmemLayout2<BN2, EPI>;
  constexpr int STAGES = SL::kStages;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_base = smem_base + SL::kBarOffset;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + SL::kBarOffset + (2 * STAGES + 4) * 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  constexpr uint32_t kTmemCols = 2 * BN2;   // two accumulator stages (256 or 512 columns)
  constexpr uint32_t kPeerMask = 0xFEFFFFFFu;   // clears the CTA-rank bit: address of the leader's copy

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_c) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 2);    // one arrive per CTA's producer
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 2 * EPI);  // epilogue warps of both CTAs
    }
    for (int s = 0; s < 2 * EPI; ++s) mbar_init(smem_base + SL::kAddBarOffset + 8u * s, 1);
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

  const int num_pairs = gridDim.x / 2;
  const int pair = blockIdx.x / 2;
  const int total_tiles = p.num_m * p.num_n * p.splits;    // tiles of 256 x BN2 (x K splits)
  const int kb_total = (p.K + BK - 1) / BK;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair; t < total_tiles; t += num_pairs) {
        const int n_idx = t % p.num_n;
        const int m_idx = (t / p.num_n) % p.num_m;
        const int split = t / (p.num_n * p.num_m);
        int kb0, kb1;
        tile_kb_range(p, split, kb_total, kb0, kb1);
        const int m0 = m_idx * 256 + (int)cta * BM;
        const int n0 = n_idx * BN2 + (int)cta * (BN2 / 2);
        const int pq = p.conv_p * p.conv_q;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t sa = smem_base + stage * SL::kStageBytes;
          const uint32_t sb = sa + SL::kABytes;
          const uint32_t lbar = full_bar(stage) & kPeerMask;
          if (p.conv_mode == 1) {
            const int cblks = p.conv_c / BK;
            const int tap = kb / cblks, cb = kb - tap * cblks;
            const int cs = p.sub_en ? ((split & 1) ? 2 : 1) : p.conv_s;
            const int kh = tap / cs, kw = tap - kh * cs;
            const int in0 = m0 / pq, rem = m0 - in0 * pq, oh0 = rem / p.conv_q, ow0 = rem - oh0 * p.conv_q;
            const int fh = p.conv_flip ? p.conv_r - 1 - kh : kh, fw = p.conv_flip ? p.conv_s - 1 - kw : kw;
            tma_load_im2col_2sm(sa, &tma_a, lbar, cb * BK, ow0 * p.conv_stride - p.conv_pad,
                                oh0 * p.conv_stride - p.conv_pad, in0, fw, fh);
            if (B_MN) {
              const int wtap = p.sub_en ? ((p.sub_wtaps[split] >> (4 * tap)) & 15) : tap;
#pragma unroll
              for (int a = 0; a < (BN2 / 2) / 64; ++a)
                tma_load_2d_2sm(sb + a * (BK * 128), &tma_b, lbar, wtap * p.N + n0 + a * 64, cb * BK);
            } else {
              tma_load_2d_2sm(sb, &tma_b, lbar, kb * BK, n0);
            }
          } else {
          if (A_MN) {
#pragma unroll
            for (int a = 0; a < BM / 64; ++a)
              tma_load_2d_2sm(sa + a * (BK * 128), &tma_a, lbar, m0 + a * 64, kb * BK);
          } else {
            tma_load_2d_2sm(sa, &tma_a, lbar, kb * BK, m0);
          }
          if (p.conv_mode == 2) {
            const int tap = n0 / p.conv_c, c0 = n0 - tap * p.conv_c;
            const int kh = tap / p.conv_s, kw = tap - kh * p.conv_s;
            const int px = kb * BK;
            const int in0 = px / pq, rem = px - in0 * pq, oh0 = rem / p.conv_q, ow0 = rem - oh0 * p.conv_q;
#pragma unroll
            for (int a = 0; a < (BN2 / 2) / 64; ++a)
              tma_load_im2col_2sm(sb + a * (BK * 128), &tma_b, lbar, c0 + a * 64, ow0 * p.conv_stride - p.conv_pad,
                                  oh0 * p.conv_stride - p.conv_pad, in0, kw, kh);
          } else if (B_MN) {
#pragma unroll
            for (int a = 0; a < (BN2 / 2) / 64; ++a)
              tma_load_2d_2sm(sb + a * (BK * 128), &tma_b, lbar, n0 + a * 64, kb * BK);
          } else {
            tma_load_2d_2sm(sb, &tma_b, lbar, kb * BK, n0);
          }
          }
          if (leader) mbar_expect_tx(full_bar(stage), 2 * SL::kStageBytes);
          else mbar_arrive_remote(full_bar(stage), 0);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      const uint32_t el = elect_one();   // whole warp walks the loop, one lane issues (see 1-CTA kernel)
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)A_MN << 15) |
                                 ((uint32_t)B_MN << 16) | ((uint32_t)(BN2 >> 3) << 17) |
                                 ((uint32_t)(256 >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int iter = 0;
      for (int t = pair; t < total_tiles; t += num_pairs, ++iter) {
        const int split = t / (p.num_n * p.num_m);
        int kb0, kb1;
        tile_kb_range(p, split, kb_total, kb0, kb1);
        const int as = iter & 1;
        const uint32_t aphase = (iter >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN2;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_base + stage * SL::kStageBytes;
          const uint64_t ad0 = make_desc<A_MN>(sa);
          const uint64_t bd0 = make_desc<B_MN>(sa + SL::kABytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            umma_bf16_2sm_pred(tmem_d, ad0 + (uint64_t)(k_advance_bytes<A_MN>(k) >> 4),
                               bd0 + (uint64_t)(k_advance_bytes<B_MN>(k) >> 4), idesc,
                               (kb > kb0 || k > 0) ? 1u : 0u, el);
          }
          umma_commit_2sm_pred(empty_bar(stage), el);
          if (kb == kb1 - 1) umma_commit_2sm_pred(tfull_bar(as), el);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= kEpiWarp0) {
    gemm_epilogue<BN2, EPI, true, SL>(p, &tma_c, &tma_d, smem, smem_base, tmem_base, tfull_bar(0), tempty_bar(0), warp, lane,
                                      (int)cta, pair, num_pairs, total_tiles);
  }

  tcgen05_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(kTmemCols)
                 : "memory");
  }
}

// Deterministic split-K reduction: out = bf16(sum_s partial[s]) (fixed order).
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ partial,
                                                            bf16* __restrict__ out, int64_t MN,
                                                            int N, int ldc, int splits) {
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < MN; i += T) {
    // eight independent loads in flight per thread (the loop is latency-bound: a few thousand threads walk
    // up to 128 partial planes); the summation order stays fixed, so the result is deterministic
    float acc = 0.f;
    int s = 0;
    for (; s + 8 <= splits; s += 8) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldcg(partial + (size_t)(s + u) * MN + i);
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += v[u];
    }
    for (; s < splits; ++s) acc += __ldcg(partial + (size_t)s * MN + i);
    const int64_t r = i / N, c = i % N;
    out[r * ldc + c] = __float2bfloat16(acc);
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
    else
      cudaGetLastError();
  }
  return fn;
}

// 2-D bf16 tensor map: inner (contiguous) extent d0, outer extent d1, row pitch ld elements.
static int make_map(CUtensorMap* map, const void* base, uint64_t d0, uint64_t d1, uint64_t ld,
                    uint32_t box0, uint32_t box1) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    kdl_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return KDL_ERR_UNSUPPORTED;
  }
  cuuint64_t dims[2] = {d0, d1};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box0, box1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    kdl_set_error("cuTensorMapEncodeTiled failed (%d): base=%p dims=%llu,%llu ld=%llu box=%u,%u",
                  (int)r, base, (unsigned long long)d0, (unsigned long long)d1,
                  (unsigned long long)ld, box0, box1);
    return KDL_ERR_CUDA;
  }
  return KDL_OK;
}

static int g_num_sms = 0;
static int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

// Output maps: bf16 [M][ldc] with box {64, 32}, or fp32 [splits][M][N] with box {32, 32, 1}.
static int make_c_map(CUtensorMap* map, void* base, int fp32_partial, uint64_t M, uint64_t N,
                      uint64_t ldc, uint64_t splits) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    kdl_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return KDL_ERR_UNSUPPORTED;
  }
  CUresult r;
  if (fp32_partial) {
    cuuint64_t dims[3] = {N, M, splits};
    cuuint64_t strides[2] = {N * 4, M * N * 4};
    cuuint32_t box[3] = {32, 32, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    cuuint64_t dims[2] = {N, M};
    cuuint64_t strides[1] = {ldc * 2};
    cuuint32_t box[2] = {64, 32};
    cuuint32_t estr[2] = {1, 1};
    r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) {
    kdl_set_error("cuTensorMapEncodeTiled(C) failed (%d): M=%llu N=%llu ldc=%llu", (int)r,
                  (unsigned long long)M, (unsigned long long)N, (unsigned long long)ldc);
    return KDL_ERR_CUDA;
  }
  return KDL_OK;
}

// CTAs (or CTA pairs) when column statistics are fused: a multiple of the number of column tiles
static int stats_grid(int want, int num_n) {
  int g = want - want % num_n;
  return g < num_n ? num_n : g;
}

// Epilogue warps for a launch: 8 when a tile carries little K (epilogue-bound, see the top of the file),
// 4 for deep K.  KDL_GEMM_EPI=4|8 forces one (A/B runs).
static int pick_epi(int k_per_tile, bool two_cta) {
  static int env = -1;
  if (env < 0) {
    const char* e = getenv("KDL_GEMM_EPI");
    env = e ? atoi(e) : 0;
  }
  if (env == 4 || env == 8) return env;
  static int k1 = -1, k2 = -1;       // thresholds (K per tile), overridable for sweeps
  if (k1 < 0) {
    const char* a = getenv("KDL_GEMM_EPI8_MAXK");
    const char* b = getenv("KDL_GEMM_EPI8_MAXK_2CTA");
    k1 = a ? atoi(a) : 256;
    k2 = b ? atoi(b) : 512;
  }
  return k_per_tile <= (two_cta ? k2 : k1) ? 8 : 4;
}

template <int A_MN, int B_MN, int BN, int STAGES, int EPI>
static int launch_e(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& md,
                    const GemmParams& p, cudaStream_t stream) {
  using SL = SmemLayout<BN, STAGES, EPI>;
  auto kern = gemm_tcgen05_kernel<A_MN, B_MN, BN, STAGES, EPI>;
  static bool attr_done = false;
  if (!attr_done) {
    KDL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SL::kTotal + 1024));
    attr_done = true;
  }
  int tiles = p.num_m * p.num_n * p.splits;
  int grid = tiles < num_sms() ? tiles : num_sms();
  if (p.stats) grid = stats_grid(grid, p.num_n);   // every CTA keeps one column tile (register-held statistics)
  kern<<<grid, gemm_threads(EPI), SL::kTotal + 1024, stream>>>(ma, mb, mc, md, p);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

// BN = 128: 5 stages with 4 epilogue warps, 4 with 8;  BN = 64: 6 / 5.
template <int A_MN, int B_MN, int BN, int STAGES>
static int launch(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& md,
                  const GemmParams& p, cudaStream_t stream) {
  const int k_per_tile = (p.kb_per_split > 0 ? p.kb_per_split : (p.K + BK - 1) / BK) * BK;
  if (pick_epi(k_per_tile, false) == 8) return launch_e<A_MN, B_MN, BN, STAGES - 1, 8>(ma, mb, mc, md, p, stream);
  return launch_e<A_MN, B_MN, BN, STAGES, 4>(ma, mb, mc, md, p, stream);
}

template <int A_MN, int B_MN, int BN2, int EPI>
static int launch_2cta_e(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& md,
                         const GemmParams& p, cudaStream_t stream) {
  using SL = SmemLayout2<BN2, EPI>;
  auto kern = gemm_tcgen05_2cta_kernel<A_MN, B_MN, BN2, EPI>;
  static bool attr_done = false;
  if (!attr_done) {
    KDL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SL::kTotal + 1024));
    attr_done = true;
  }
  int tiles = p.num_m * p.num_n * (p.splits > 0 ? p.splits : 1);
  int pairs = num_sms() / 2;
  if (tiles < pairs) pairs = tiles;
  if (p.stats) pairs = stats_grid(pairs, p.num_n);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(gemm_threads(EPI));
  cfg.dynamicSmemBytes = SL::kTotal + 1024;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  KDL_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, mc, md, p));
  return KDL_OK;
}

template <int A_MN, int B_MN, int BN2 = 256>
static int launch_2cta(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& md,
                       const GemmParams& p, cudaStream_t stream) {
  const int k_per_tile = (p.kb_per_split > 0 ? p.kb_per_split : (p.K + BK - 1) / BK) * BK;
  // the 8-warp epilogue exists for plain / im2col fprop and dgrad tiles; split-K partial tiles (wgrad) are
  // deep-K by construction
  if (p.partial == nullptr && pick_epi(k_per_tile, true) == 8)
    return launch_2cta_e<A_MN, B_MN, BN2, 8>(ma, mb, mc, md, p, stream);
  return launch_2cta_e<A_MN, B_MN, BN2, 4>(ma, mb, mc, md, p, stream);
}

template <int BN, int STAGES>
static int dispatch_major(int a_mn, int b_mn, const CUtensorMap& ma, const CUtensorMap& mb,
                          const CUtensorMap& mc, const CUtensorMap& md, const GemmParams& p, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch<0, 0, BN, STAGES>(ma, mb, mc, md, p, s);
  if (!a_mn && b_mn) return launch<0, 1, BN, STAGES>(ma, mb, mc, md, p, s);
  if (a_mn && b_mn) return launch<1, 1, BN, STAGES>(ma, mb, mc, md, p, s);
  return launch<1, 0, BN, STAGES>(ma, mb, mc, md, p, s);
}

}  // namespace kdl

using namespace kdl;

static int pick_bn(int N) { return N <= 64 ? 64 : 128; }

// 2-CTA tiles (256 x 256) pay off when the problem is compute-heavy enough: wide N, deep-ish K.
static bool want_2cta(int M, int N, int K, int splits, int force) {
  if (force == 0) return false;
  if (splits > 1) return false;
  if (force == 1) return N >= 128 && M >= 256;
  static int env = -1;
  if (env < 0) {
    const char* e = getenv("KDL_GEMM_2CTA");
    env = e ? atoi(e) : 2;   // 0 = never, 1 = whenever legal, 2 = heuristic
  }
  if (env == 0) return false;
  if (env == 1) return N >= 128 && M >= 256;
  return N >= 256 && M >= 1024 && K >= 256;
}

static int pick_splits(int M, int N, int K, int requested) {
  int bn = pick_bn(N);
  int tiles = ((M + BM - 1) / BM) * ((N + bn - 1) / bn);
  int kb_total = (K + BK - 1) / BK;
  int s = requested;
  if (s <= 0) {
    // auto: only split when the output grid cannot fill the machine and K is deep
    s = 1;
    if (tiles < 148 && kb_total >= 16) {
      s = (148 * 2) / tiles;
      if (s > kb_total / 4) s = kb_total / 4;
      if (s < 1) s = 1;
      if (s > 128) s = 128;
    }
  }
  if (s > kb_total) s = kb_total;
  int per = (kb_total + s - 1) / s;
  s = (kb_total + per - 1) / per;
  return s;
}

extern "C" int kdl_gemm_splits(int M, int N, int K, int requested) {
  return pick_splits(M, N, K, requested);
}

extern "C" size_t kdl_gemm_scratch_bytes(int M, int N, int K, int split_k) {
  int s = pick_splits(M, N, K, split_k);
  return s > 1 ? (size_t)s * M * N * sizeof(float) : 0;
}

// rows a statistics buffer needs: one per CTA
extern "C" int kdl_gemm_max_ctas(void) { return num_sms(); }

static int gemm_impl(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major, int b_mn_major,
                     int lda, int ldb, int ldc, const void* bias, const void* addend, int add_h, int add_w,
                     const void* add_mask, int relu, int split_k, void* scratch, size_t scratch_bytes,
                     float* stats_partial, int* stats_parts_out, cudaStream_t stream);

extern "C" int kdl_gemm_bf16_ex(const void* A, const void* B, void* C, int M, int N, int K,
                                int a_mn_major, int b_mn_major, int lda, int ldb, int ldc,
                                const void* bias, const void* addend, int relu, int split_k,
                                void* scratch, size_t scratch_bytes, float* stats_partial,
                                int* stats_parts_out, cudaStream_t stream) {
  return gemm_impl(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, bias, addend, 0, 0, nullptr, relu,
                   split_k, scratch, scratch_bytes, stats_partial, stats_parts_out, stream);
}

// C[M, N] = A B + (mask ? addend : 0): `relu_mask` is the [M*N/8] bit mask a BatchNorm forward saved
// (kdl_bn_fwd_mask), i.e. the addend is a gradient that still has to pass that ReLU.  Lets the residual
// branch of a bottleneck read dOut and the mask instead of a separately written dOut * mask tensor.
extern "C" int kdl_gemm_bf16_maskadd(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                                     int b_mn_major, int lda, int ldb, int ldc, const void* addend,
                                     const void* relu_mask, cudaStream_t stream) {
  if (addend == nullptr || relu_mask == nullptr || (N & 63) || ldc != N || ((uintptr_t)relu_mask & 7)) {
    kdl_set_error("gemm_maskadd: need addend, an 8-byte aligned mask, N %% 64 == 0 and ldc == N (N=%d ldc=%d)", N, ldc);
    return KDL_ERR_ARG;
  }
  return gemm_impl(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, nullptr, addend, 0, 0, relu_mask, 0, 1,
                   nullptr, 0, nullptr, nullptr, stream);
}

// C[M, N] = A B (+ addend): rows of C are the pixels of an [n, H, W] map and `addend_compact` is the
// [n, H/2, W/2, N] gradient of a 1x1 / stride-2 branch reading the same input, added at even (h, w).
extern "C" int kdl_gemm_bf16_subadd(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                                    int b_mn_major, int lda, int ldb, int ldc, const void* addend_compact, int H,
                                    int W, cudaStream_t stream) {
  if (H <= 0 || W <= 0 || (H & 1) || (W & 1) || M % (H * W) != 0) {
    kdl_set_error("gemm_subadd: need even H, W and M a multiple of H*W (M=%d H=%d W=%d)", M, H, W);
    return KDL_ERR_ARG;
  }
  return gemm_impl(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, nullptr, addend_compact, H, W, nullptr,
                   0, 1, nullptr, 0, nullptr, nullptr, stream);
}

extern "C" int kdl_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                             int a_mn_major, int b_mn_major, int lda, int ldb, int ldc,
                             const void* bias, const void* addend, int relu, int split_k,
                             void* scratch, size_t scratch_bytes, cudaStream_t stream) {
  return kdl_gemm_bf16_ex(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, bias, addend, relu,
                          split_k, scratch, scratch_bytes, nullptr, nullptr, stream);
}

static int gemm_impl(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major, int b_mn_major,
                     int lda, int ldb, int ldc, const void* bias, const void* addend, int add_h, int add_w,
                     const void* add_mask, int relu, int split_k, void* scratch, size_t scratch_bytes,
                     float* stats_partial, int* stats_parts_out, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) {
    kdl_set_error("gemm: bad shape %dx%dx%d", M, N, K);
    return KDL_ERR_ARG;
  }
  if ((lda & 7) || (ldb & 7) || (ldc & 7) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15) ||
      ((uintptr_t)C & 15)) {
    kdl_set_error("gemm: operands must be 16-byte aligned with leading dims %% 8 == 0 (lda=%d ldb=%d ldc=%d)",
                  lda, ldb, ldc);
    return KDL_ERR_ARG;
  }
  const int bn = pick_bn(N);
  const int splits = pick_splits(M, N, K, split_k);
  GemmParams p = {};
  p.C = (bf16*)C;
  p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.bias = (const bf16*)bias;
  p.addend = (const bf16*)addend;
  p.add_h = add_h; p.add_w = add_w;
  p.add_mask = (const uint8_t*)add_mask;
  p.relu = relu;
  p.splits = splits;
  const int kb_total = (K + BK - 1) / BK;
  p.kb_per_split = (kb_total + splits - 1) / splits;
  p.num_m = (M + BM - 1) / BM;
  p.num_n = (N + bn - 1) / bn;
  p.stats = stats_partial;
  if (stats_partial != nullptr) {
    if (N > kStatsMaxN || splits > 1) {
      kdl_set_error("gemm: fused column statistics need N <= %d and no split-K", kStatsMaxN);
      return KDL_ERR_UNSUPPORTED;
    }
  }
  if (splits > 1) {
    if (bias || addend || relu) {
      kdl_set_error("gemm: split-K does not support fused bias/addend/relu");
      return KDL_ERR_UNSUPPORTED;
    }
    if (N & 3) {
      kdl_set_error("gemm: split-K needs N %% 4 == 0");
      return KDL_ERR_UNSUPPORTED;
    }
    size_t need = (size_t)splits * M * N * sizeof(float);
    if (scratch == nullptr || scratch_bytes < need) {
      kdl_set_error("gemm: split-K scratch too small (%zu < %zu)", scratch_bytes, need);
      return KDL_ERR_WORKSPACE;
    }
    p.partial = (float*)scratch;
  }
  CUtensorMap ma, mb;
  int rc;
  const bool two = want_2cta(M, N, K, splits, -1);
  // K-major: global [rows=MN extent][K] ; MN-major: global [K][MN extent]
  if (a_mn_major) rc = make_map(&ma, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BK);
  else rc = make_map(&ma, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM);
  if (rc) return rc;
  if (b_mn_major) rc = make_map(&mb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK);
  else rc = make_map(&mb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, two ? (uint32_t)128 : (uint32_t)bn);
  if (rc) return rc;
  // full-size addend: its own map with the output's box shape, so a chunk lands in a staging buffer
  // exactly as the result chunk will be laid out there (the compact stride-2 addend stays on loads)
  CUtensorMap md;
  rc = make_c_map(&md, addend != nullptr && add_h == 0 ? const_cast<void*>(addend) : C, 0, (uint64_t)M, (uint64_t)N,
                  (uint64_t)ldc, 1);
  if (rc) return rc;
  p.add_tma = (addend != nullptr && add_h == 0 && ((uintptr_t)addend & 15) == 0) ? 1 : 0;
  if (two) {
    CUtensorMap mc2;
    rc = make_c_map(&mc2, C, 0, (uint64_t)M, (uint64_t)N, (uint64_t)ldc, 1);
    if (rc) return rc;
    p.num_m = (M + 255) / 256;
    p.num_n = (N + 255) / 256;
    if (stats_parts_out) {
      int tiles2 = p.num_m * p.num_n, pairs = num_sms() / 2;
      *stats_parts_out = (stats_grid(tiles2 < pairs ? tiles2 : pairs, p.num_n) / p.num_n) * 2;
    }
    if (!a_mn_major && !b_mn_major) return launch_2cta<0, 0>(ma, mb, mc2, md, p, stream);
    if (!a_mn_major && b_mn_major) return launch_2cta<0, 1>(ma, mb, mc2, md, p, stream);
    if (a_mn_major && b_mn_major) return launch_2cta<1, 1>(ma, mb, mc2, md, p, stream);
    return launch_2cta<1, 0>(ma, mb, mc2, md, p, stream);
  }
  if (stats_parts_out) {
    int tiles1 = p.num_m * p.num_n * p.splits;
    *stats_parts_out = stats_grid(tiles1 < num_sms() ? tiles1 : num_sms(), p.num_n) / p.num_n;
  }
  CUtensorMap mc;
  if (splits > 1) rc = make_c_map(&mc, p.partial, 1, (uint64_t)M, (uint64_t)N, (uint64_t)N, (uint64_t)splits);
  else rc = make_c_map(&mc, C, 0, (uint64_t)M, (uint64_t)N, (uint64_t)ldc, 1);
  if (rc) return rc;
  if (bn == 64) rc = dispatch_major<64, 6>(a_mn_major, b_mn_major, ma, mb, mc, md, p, stream);
  else rc = dispatch_major<128, 5>(a_mn_major, b_mn_major, ma, mb, mc, md, p, stream);
  if (rc) return rc;
  if (splits > 1) {
    int64_t MN = (int64_t)M * N;
    int64_t g = (MN + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    if (!launch_plane_reduce_small(p.partial, (bf16*)C, MN, N, ldc, splits, stream))
      splitk_reduce_kernel<<<(int)g, 256, 0, stream>>>(p.partial, (bf16*)C, MN, N, ldc, splits);
    KDL_CHECK_LAUNCH();
  }
  return KDL_OK;
}


// ------------------------------------------------------------------ N7: implicit-GEMM convolution
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeIm2colFn get_encode_im2col() {
  static EncodeIm2colFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeIm2colFn)ptr;
    else
      cudaGetLastError();
  }
  return fn;
}

// NHWC bf16 activation [N,H,W,C] as an im2col tensor map: `pixels` output positions x 64 channels per load.
static int make_im2col_map(CUtensorMap* map, const void* base, int N, int H, int W, int C, int R, int S,
                           int stride, int pad, int pixels, const int* corners = nullptr) {
  EncodeIm2colFn enc = get_encode_im2col();
  if (!enc) {
    kdl_set_error("cuTensorMapEncodeIm2col unavailable (no CUDA driver?)");
    return KDL_ERR_UNSUPPORTED;
  }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  int lower[2] = {-pad, -pad};
  int upper[2] = {pad - (S - 1), pad - (R - 1)};
  if (corners) { lower[0] = corners[0]; lower[1] = corners[1]; upper[0] = corners[2]; upper[1] = corners[3]; }
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower,
                   upper, 64, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    kdl_set_error("cuTensorMapEncodeIm2col failed (%d): N=%d H=%d W=%d C=%d R=%d S=%d stride=%d pad=%d", (int)r, N,
                  H, W, C, R, S, stride, pad);
    return KDL_ERR_CUDA;
  }
  return KDL_OK;
}

extern "C" size_t kdl_conv2d_scratch_bytes(int mode, int N, int H, int W, int Cin, int Cout, int R, int S,
                                           int stride, int pad) {
  if (mode != 2) return 0;
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  size_t a = kdl_gemm_scratch_bytes(Cout, R * S * Cin, N * OH * OW, 0);
  size_t b = (size_t)148 * Cout * R * S * Cin * sizeof(float) / 1;   // upper bound for the 2-CTA split choice
  if (b > ((size_t)64 * Cout * R * S * Cin * sizeof(float))) b = (size_t)64 * Cout * R * S * Cin * sizeof(float);
  return a > b ? a : b;
}

// mode 0 fprop : in0 = x [N,H,W,Cin],   in1 = w [Cout,R,S,Cin] -> out = y  [N,OH,OW,Cout]
// mode 1 dgrad : in0 = dy [N,OH,OW,Cout], in1 = w              -> out = dx [N,H,W,Cin] (+addend); stride 1, 2*pad = R-1
// mode 2 wgrad : in0 = dy,               in1 = x               -> out = dw [Cout,R,S,Cin] (split over pixels)
static int conv2d_impl(int mode, const void* in0, const void* in1, void* out, int N, int H, int W,
                       int Cin, int Cout, int R, int S, int stride, int pad, const void* addend,
                       void* scratch, size_t scratch_bytes, float* stats_partial, int* stats_parts_out,
                       cudaStream_t stream);

extern "C" int kdl_conv2d_bf16(int mode, const void* in0, const void* in1, void* out, int N, int H, int W,
                               int Cin, int Cout, int R, int S, int stride, int pad, const void* addend,
                               void* scratch, size_t scratch_bytes, cudaStream_t stream) {
  return conv2d_impl(mode, in0, in1, out, N, H, W, Cin, Cout, R, S, stride, pad, addend, scratch, scratch_bytes, nullptr,
                     nullptr, stream);
}

// fprop that also returns, from the same epilogue, the per-CTA column sum / sum of squares of y (as stored) for the
// BatchNorm that follows: stats_partial [kdl_gemm_max_ctas()][2][Cout] fp32, *stats_parts_out rows are valid.
extern "C" int kdl_conv2d_fprop_stats(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout,
                                      int R, int S, int stride, int pad, float* stats_partial, int* stats_parts_out,
                                      cudaStream_t stream) {
  if (stats_partial == nullptr || stats_parts_out == nullptr || Cout > kStatsMaxN) {
    kdl_set_error("conv2d_fprop_stats: need a statistics buffer and Cout <= %d", kStatsMaxN);
    return KDL_ERR_ARG;
  }
  return conv2d_impl(0, x, w, y, N, H, W, Cin, Cout, R, S, stride, pad, nullptr, nullptr, 0, stats_partial, stats_parts_out,
                     stream);
}

static int conv2d_impl(int mode, const void* in0, const void* in1, void* out, int N, int H, int W,
                       int Cin, int Cout, int R, int S, int stride, int pad, const void* addend,
                       void* scratch, size_t scratch_bytes, float* stats_partial, int* stats_parts_out,
                       cudaStream_t stream) {
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  if (Cin % 64 || Cout % 64) {
    kdl_set_error("conv2d: Cin and Cout must be multiples of 64 (Cin=%d Cout=%d)", Cin, Cout);
    return KDL_ERR_UNSUPPORTED;
  }
  if (mode == 1 && (stride != 1 || 2 * pad != R - 1 || 2 * pad != S - 1)) {
    kdl_set_error("conv2d dgrad: only stride 1 'same' convolutions are implemented");
    return KDL_ERR_UNSUPPORTED;
  }
  GemmParams p = {};
  CUtensorMap ma, mb, mc;
  int rc;
  p.conv_s = S; p.conv_r = R; p.conv_stride = stride; p.conv_pad = pad;
  p.relu = 0; p.splits = 1;
  if (mode == 0 || mode == 1) {
    const int Cpatch = mode == 0 ? Cin : Cout;        // channels of the patch operand
    const int Nout = mode == 0 ? Cout : Cin;
    const int Hs = mode == 0 ? H : OH, Ws = mode == 0 ? W : OW;   // spatial dims of the patch tensor
    const int P = mode == 0 ? OH : H, Q = mode == 0 ? OW : W;     // spatial dims the pixel index runs over
    const int M = N * P * Q, K = R * S * Cpatch;
    const int bn = Nout <= 64 ? 64 : 128;
    p.C = (bf16*)out; p.M = M; p.N = Nout; p.K = K; p.ldc = Nout;
    p.addend = (const bf16*)addend;
    p.kb_per_split = (K + BK - 1) / BK;
    p.num_m = (M + BM - 1) / BM; p.num_n = (Nout + bn - 1) / bn;
    p.conv_mode = 1; p.conv_q = Q; p.conv_p = P; p.conv_c = Cpatch; p.conv_flip = mode == 1;
    const int eff_pad = mode == 0 ? pad : (R - 1 - pad);
    p.conv_pad = eff_pad;
    rc = make_im2col_map(&ma, in0, N, Hs, Ws, Cpatch, R, S, mode == 0 ? stride : 1, eff_pad, BM);
    if (rc) return rc;
    static int conv2 = -1;
    if (conv2 < 0) {
      const char* e = getenv("KDL_CONV_2CTA");
      conv2 = e ? atoi(e) : 1;
    }
    const bool two = conv2 && M >= 1024 && Nout >= 128;
    // N tile of the CTA pair.  (A 128-wide tile would fill the 74 pairs better for Cout = 512 at 7x7,
    // but the loop is bound by the im2col TMA rate per 128-pixel load, not by the MMA, so the wider
    // tile that amortises each load over more columns wins: measured 0.070 vs 0.098 ms.)
    const int bn2 = Nout >= 256 ? 256 : 128;
    const uint32_t brows = two ? (uint32_t)(bn2 / 2) : (uint32_t)bn;
    if (mode == 0) rc = make_map(&mb, in1, (uint64_t)K, (uint64_t)Cout, (uint64_t)K, BK, brows);
    else rc = make_map(&mb, in1, (uint64_t)(R * S * Cin), (uint64_t)Cout, (uint64_t)(R * S * Cin), 64, BK);
    if (rc) return rc;
    rc = make_c_map(&mc, out, 0, (uint64_t)M, (uint64_t)Nout, (uint64_t)Nout, 1);
    if (rc) return rc;
    CUtensorMap md = mc;
    if (addend != nullptr && ((uintptr_t)addend & 15) == 0) {
      rc = make_c_map(&md, const_cast<void*>(addend), 0, (uint64_t)M, (uint64_t)Nout, (uint64_t)Nout, 1);
      if (rc) return rc;
      p.add_tma = 1;
    }
    if (mode == 0 && stats_partial != nullptr) {
      p.stats = stats_partial;        // every CTA (pair) keeps one column tile: see stats_grid()
      if (two) {
        const int nm = (M + 255) / 256, nn = (Nout + bn2 - 1) / bn2, tiles2 = nm * nn, pairs = num_sms() / 2;
        *stats_parts_out = (stats_grid(tiles2 < pairs ? tiles2 : pairs, nn) / nn) * 2;
      } else {
        const int tiles1 = p.num_m * p.num_n;
        *stats_parts_out = stats_grid(tiles1 < num_sms() ? tiles1 : num_sms(), p.num_n) / p.num_n;
      }
    }
    if (two) {
      p.num_m = (M + 255) / 256;
      p.num_n = (Nout + bn2 - 1) / bn2;
      if (mode == 0) return bn2 == 256 ? launch_2cta<0, 0, 256>(ma, mb, mc, md, p, stream) : launch_2cta<0, 0, 128>(ma, mb, mc, md, p, stream);
      return bn2 == 256 ? launch_2cta<0, 1, 256>(ma, mb, mc, md, p, stream) : launch_2cta<0, 1, 128>(ma, mb, mc, md, p, stream);
    }
    if (mode == 0) return bn == 64 ? launch<0, 0, 64, 6>(ma, mb, mc, md, p, stream) : launch<0, 0, 128, 5>(ma, mb, mc, md, p, stream);
    return bn == 64 ? launch<0, 1, 64, 6>(ma, mb, mc, md, p, stream) : launch<0, 1, 128, 5>(ma, mb, mc, md, p, stream);
  }
  if (mode == 2) {
    const int M = Cout, Nn = R * S * Cin, K = N * OH * OW;
    const int bn = Cin <= 64 ? 64 : 128;
    if (Cin % bn) {
      kdl_set_error("conv2d wgrad: Cin must be a multiple of %d", bn);
      return KDL_ERR_UNSUPPORTED;
    }
    static int conv2w = -1;
    if (conv2w < 0) {
      const char* e = getenv("KDL_CONV_2CTA");
      conv2w = e ? atoi(e) : 1;
    }
    const bool two = conv2w && Cout >= 256 && Cin >= 128;
    const int kb_total = (K + BK - 1) / BK;
    int splits;
    if (two) {
      int tiles2 = ((M + 255) / 256) * (Nn / 256 + ((Nn % 256) ? 1 : 0));
      splits = (148 + tiles2 - 1) / tiles2;
      if (splits > kb_total / 4) splits = kb_total / 4;
      if (splits < 1) splits = 1;
      if (splits > 64) splits = 64;
      int per = (kb_total + splits - 1) / splits;
      splits = (kb_total + per - 1) / per;
    } else {
      splits = pick_splits(M, Nn, K, 0);
    }
    p.C = (bf16*)out; p.M = M; p.N = Nn; p.K = K; p.ldc = Nn;
    p.splits = splits; p.kb_per_split = (kb_total + splits - 1) / splits;
    p.num_m = (M + BM - 1) / BM; p.num_n = Nn / bn;
    p.conv_mode = 2; p.conv_q = OW; p.conv_p = OH; p.conv_c = Cin; p.conv_flip = 0;
    if (splits > 1) {
      size_t need = (size_t)splits * M * Nn * sizeof(float);
      if (scratch == nullptr || scratch_bytes < need) {
        kdl_set_error("conv2d wgrad: scratch too small (%zu < %zu)", scratch_bytes, need);
        return KDL_ERR_WORKSPACE;
      }
      p.partial = (float*)scratch;
    }
    rc = make_map(&ma, in0, (uint64_t)Cout, (uint64_t)K, (uint64_t)Cout, 64, BK);   // dy^T: MN-major A
    if (rc) return rc;
    rc = make_im2col_map(&mb, in1, N, H, W, Cin, R, S, stride, pad, BK);
    if (rc) return rc;
    if (splits > 1) rc = make_c_map(&mc, p.partial, 1, (uint64_t)M, (uint64_t)Nn, (uint64_t)Nn, (uint64_t)splits);
    else rc = make_c_map(&mc, out, 0, (uint64_t)M, (uint64_t)Nn, (uint64_t)Nn, 1);
    if (rc) return rc;
    const CUtensorMap md = mc;
    if (two) {
      p.num_m = (M + 255) / 256;
      p.num_n = (Nn + 255) / 256;
      rc = launch_2cta<1, 1, 256>(ma, mb, mc, md, p, stream);
    } else {
      rc = bn == 64 ? launch<1, 1, 64, 6>(ma, mb, mc, md, p, stream) : launch<1, 1, 128, 5>(ma, mb, mc, md, p, stream);
    }
    if (rc) return rc;
    if (splits > 1) {
      int64_t MN = (int64_t)M * Nn;
      int64_t g = (MN + 255) / 256;
      if (g > 148 * 8) g = 148 * 8;
      if (!launch_plane_reduce_small(p.partial, (bf16*)out, MN, Nn, Nn, splits, stream))
        splitk_reduce_kernel<<<(int)g, 256, 0, stream>>>(p.partial, (bf16*)out, MN, Nn, Nn, splits);
      KDL_CHECK_LAUNCH();
    }
    return KDL_OK;
  }
  kdl_set_error("conv2d: unknown mode %d", mode);
  return KDL_ERR_ARG;
}

// ---- 3x3 / stride-2 / pad-1 dgrad ------------------------------------------------------------------------
// dx[n, ih, iw, :] = sum over (kh, kw) with (ih + 1 - kh), (iw + 1 - kw) even of dy[n, (ih+1-kh)/2, (iw+1-kw)/2, :] w[:, kh, kw, :].
// Split by the parity (ph, pw) of the input pixel, each class is a STRIDE-1 correlation over dy with 1 or 2 taps
// per axis (even: kh = 1 at offset 0; odd: kh = 2 at offset 0 and kh = 0 at offset +1, which runs off the
// bottom / right edge and zero-fills), i.e. 1 + 2 + 2 + 4 = 9 tap-GEMMs in total — exactly the useful work, none
// of the multiplications by the zeros of an up-sampled dy.  ONE persistent launch of the TMA-im2col kernel walks the
// tiles of all four classes (the class rides in the tile's "split" index): patch operand = dy with map corners
// {0, 0}, weight tiles picked through a per-class tap map, output rows stored at the class's pixels.
// Replaces the last cuDNN call of the ResNet-50 step (cutlass implicit_gemm_strided_dgrad via ATen
// convolution_backward).  dy [N, H/2, W/2, Cout], w [Cout, 3, 3, Cin], dx [N, H, W, Cin]; H, W even.
extern "C" int kdl_conv2d_dgrad_s2(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                                   cudaStream_t stream) {
  if ((H & 1) || (W & 1) || Cin % 64 || Cout % 64) {
    kdl_set_error("conv2d_dgrad_s2: need even H, W and Cin, Cout multiples of 64 (H=%d W=%d Cin=%d Cout=%d)", H, W, Cin, Cout);
    return KDL_ERR_UNSUPPORTED;
  }
  const int P = H / 2, Q = W / 2, M = N * P * Q;
  const int bn = Cin <= 64 ? 64 : 128;
  static int conv2 = -1;
  if (conv2 < 0) {
    const char* e = getenv("KDL_CONV_2CTA");
    conv2 = e ? atoi(e) : 1;
  }
  const bool two = conv2 && M >= 1024 && Cin >= 128;
  const int bn2 = Cin >= 256 ? 256 : 128;
  CUtensorMap ma, mb, mc;
  const int corners[4] = {0, 0, 0, 0};
  int rc = make_im2col_map(&ma, dy, N, P, Q, Cout, 1, 1, 1, 0, BM, corners);
  if (rc) return rc;
  rc = make_map(&mb, w, (uint64_t)(9 * Cin), (uint64_t)Cout, (uint64_t)(9 * Cin), 64, BK);
  if (rc) return rc;
  rc = make_c_map(&mc, dx, 0, (uint64_t)N * H * W, (uint64_t)Cin, (uint64_t)Cin, 1);   // unused by the strided store
  if (rc) return rc;
  GemmParams p = {};
  p.C = (bf16*)dx; p.M = M; p.N = Cin; p.K = 4 * Cout; p.ldc = Cin;
  p.splits = 4;                                   // the four parity classes, one persistent launch
  p.kb_per_split = 2 * (Cout / BK);               // typical class (only read by the epilogue-warp heuristic)
  p.conv_mode = 1; p.conv_q = Q; p.conv_p = P; p.conv_c = Cout; p.conv_flip = 0;
  p.conv_s = 2; p.conv_r = 2; p.conv_stride = 1; p.conv_pad = 0;
  p.sub_en = 1;
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const int R = ph ? 2 : 1, S = pw ? 2 : 1;
      int taps = 0;
      for (int th = 0; th < R; ++th)
        for (int tw = 0; tw < S; ++tw) {
          const int kh = ph ? (th == 0 ? 2 : 0) : 1, kw = pw ? (tw == 0 ? 2 : 0) : 1;
          taps |= (kh * 3 + kw) << (4 * (th * S + tw));
        }
      p.sub_wtaps[2 * ph + pw] = taps;
    }
  if (two) {
    p.num_m = (M + 255) / 256; p.num_n = (Cin + bn2 - 1) / bn2;
    rc = bn2 == 256 ? launch_2cta<0, 1, 256>(ma, mb, mc, mc, p, stream) : launch_2cta<0, 1, 128>(ma, mb, mc, mc, p, stream);
  } else {
    p.num_m = (M + BM - 1) / BM; p.num_n = (Cin + bn - 1) / bn;
    rc = bn == 64 ? launch<0, 1, 64, 6>(ma, mb, mc, mc, p, stream) : launch<0, 1, 128, 5>(ma, mb, mc, mc, p, stream);
  }
  if (rc) return rc;
  return KDL_OK;
}
