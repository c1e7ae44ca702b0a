// N3: distributed histogram GBDT kernels (BASELINE.json config #3).
//
//   gbdt_grad_logistic        pred, label -> (g, h)
//   gbdt_hist_allreduce       per-level gradient histograms H[node][feat][bin] = sum (g, h) over the
//                             rows of that node, built in shared memory (one node tile per block),
//                             flushed with red.global into a symmetric buffer, and — fused in the
//                             same launch — reduced across GPUs: the last block to finish a node
//                             signals the peers, waits for theirs and sums that node's histogram
//                             from every peer over NVLink (one-shot; the messages are 56 KiB).
//   gbdt_split_eval           best (feature, threshold) per node by exact gain scan over 256 bins
//   gbdt_apply_split          route rows to children
//   gbdt_update_pred          pred += leaf value of the row's leaf
//
// Replaces XGBoost's gpu_hist build kernel + ncclAllReduce of histograms (Rabit on CPU); the
// reference operator only injects the tracker address for them (reference
// controllers/xgboost/pod.go:106-152).
#include "common.cuh"
#include "../include/kubedl_b200.h"

namespace kdl {

constexpr int kBins = 256;
constexpr int kHistThreads = 256;
constexpr int kFlushRows = 8192;     // rows a block accumulates in int32 fixed point between flushes

__global__ void __launch_bounds__(256) grad_logistic_kernel(const float* __restrict__ pred,
                                                            const float* __restrict__ label,
                                                            float2* __restrict__ gh, int64_t n) {
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += T) {
    float p = 1.f / (1.f + __expf(-pred[i]));
    float g = p - label[i];
    float h = fmaxf(p * (1.f - p), 1e-16f);
    gh[i] = make_float2(g, h);
  }
}

struct HistParams {
  const uint8_t* bins;   // [n][fstride] quantised features, row-major
  const float2* gh;      // [n]
  const int32_t* ridx;   // [n] row ids grouped by node
  const int32_t* seg;    // [num_nodes_level + 1] segment offsets into ridx
  const int32_t* build;  // [nbuild] node ids (within the level) to build
  int nbuild;
  int F;                 // features
  int fstride;           // bytes per row in `bins`
  int rows_per_tile;
  float* local;          // [nbuild][F][256][2] this rank's partial (symmetric memory), zero on entry
  float* out;            // [nbuild][F][256][2] global sum (local memory)
  unsigned* node_done;   // [nbuild] local arrival counters, zero on entry, zero on exit
  // cross-GPU part
  const float* peer_local[KDL_MAX_WORLD];  // peers' `local` buffers
  const float* mc_local;                   // NVLS multicast alias of `local` (or null): in-switch sum
  uint32_t* flags[KDL_MAX_WORLD];          // per rank: [world][max nodes] epoch flags
  int world, rank;
  int max_nodes;
  uint32_t epoch;        // value that marks "this level's histogram of node k is complete"
  uint64_t timeout_ns;
  uint32_t* status;
  float qscale, inv_qscale;   // fixed-point scale of the shared-memory accumulators
  uint32_t* epoch_state;      // optional device-resident {epoch, finished nodes}: makes the launch CUDA-graph replayable
};

// Dynamic shared memory: int hist[F][256][2] (fixed point)
__global__ void __launch_bounds__(kHistThreads) hist_allreduce_kernel(HistParams p) {
  // Shared-memory accumulators are 32-bit fixed point: fp32 atomicAdd on shared memory compiles to a
  // compare-and-swap loop (ATOMS.CAST.SPIN, 64 of them per row here) that collapses under same-bin
  // collisions, integer ATOMS.ADD is native.  |g|, |h| <= gh_abs_max are scaled to 17 bits, a block
  // flushes (-> fp32 global atomics) every kFlushRows = 8192 rows, so the int32 sums cannot overflow;
  // the rounding error per row is 2^-18 of the bound.
  // (Packing g and h into one 64-bit shared atomic does not help: ATOMS has no native 64-bit add, it
  // compiles to ATOMS.CAST.SPIN.64.)
  extern __shared__ int hist[];
  __shared__ int s_last;
  const int hist_floats = p.F * kBins * 2;
  const int plane = p.F * kBins;
  const int tid = threadIdx.x;

  // blockIdx.y = index into the build list; blockIdx.x = tile of that node's rows (strided)
  const int bi = blockIdx.y;
  const int node = p.build[bi];
  const int beg = p.seg[node], end = p.seg[node + 1];
  const int ntiles = (end - beg + p.rows_per_tile - 1) / p.rows_per_tile;
  const int my_first = blockIdx.x;
  float* local = p.local + (size_t)bi * hist_floats;

  if (my_first < ntiles) {
    for (int i = tid; i < hist_floats; i += kHistThreads) hist[i] = 0;
    __syncthreads();
    int tiles_since_flush = 0;
    for (int t = my_first; t < ntiles; t += gridDim.x) {
      const int r0 = beg + t * p.rows_per_tile;
      const int r1 = min(r0 + p.rows_per_tile, end);
      for (int r = r0 + tid; r < r1; r += kHistThreads) {
        const int row = p.ridx[r];
        const float2 gh = p.gh[row];
        const int gq = __float2int_rn(fminf(fmaxf(gh.x * p.qscale, -131072.f), 131072.f));
        const int hq = __float2int_rn(fminf(fmaxf(gh.y * p.qscale, -131072.f), 131072.f));
        const uint8_t* b = p.bins + (size_t)row * p.fstride;
        // 32-byte rows: two 128-bit loads cover up to 32 features
        uint4 v0 = *reinterpret_cast<const uint4*>(b);
        uint4 v1 = p.F > 16 ? *reinterpret_cast<const uint4*>(b + 16) : make_uint4(0, 0, 0, 0);
        const uint32_t w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int f = 0; f < 32; ++f) {
          if (f < p.F) {
            const uint32_t bin = (w[f >> 2] >> ((f & 3) * 8)) & 0xffu;
            // g and h live in separate planes: lane addresses are bin-strided by ONE word, so a warp's
            // atomics spread over all 32 banks (interleaved (g,h) pairs only ever hit the even ones)
            atomicAdd(hist + f * kBins + bin, gq);
            atomicAdd(hist + plane + f * kBins + bin, hq);
          }
        }
      }
      const bool last_tile = t + (int)gridDim.x >= ntiles;
      if (++tiles_since_flush == kFlushRows / 4096 || last_tile) {
        tiles_since_flush = 0;
        __syncthreads();
        // flush: one red.global per non-zero entry
        for (int i = tid; i < hist_floats; i += kHistThreads) {
          const int v = hist[i];
          if (v != 0) {
            // plane layout [2][F][256] in shared memory -> interleaved [F][256][2] in global memory
            const int o = i < plane ? 2 * i : 2 * (i - plane) + 1;
            atomicAdd(local + o, (float)v * p.inv_qscale);
            hist[i] = 0;
          }
        }
        __syncthreads();
      }
    }
  }
  // ---- node completion: the last block of this node performs the cross-GPU reduction
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    unsigned prev = atomicAdd(p.node_done + bi, 1u);
    s_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // the epoch either comes by value (a fresh one per launch from the host) or lives on the device and
  // is advanced by the last finishing node of this launch, so a captured launch can be replayed
  const uint32_t epoch = p.epoch_state ? *((volatile uint32_t*)p.epoch_state) + 1u : p.epoch;
  float* out = p.out + (size_t)bi * hist_floats;
  if (p.world == 1) {
    for (int i = tid; i < hist_floats / 4; i += kHistThreads)
      reinterpret_cast<float4*>(out)[i] = __ldcg(reinterpret_cast<const float4*>(local) + i);
  } else {
    // tell every peer "my histogram of build-slot bi is complete", wait for theirs
    if (tid < p.world) {
      __threadfence_system();
      st_release_sys(p.flags[tid] + (size_t)p.rank * p.max_nodes + bi, epoch);
      const uint32_t* mine = p.flags[p.rank] + (size_t)tid * p.max_nodes + bi;
      uint64_t t0 = 0;
      uint32_t spins = 0;
      while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
        if (p.timeout_ns && (++spins & 1023u) == 0) {
          uint64_t now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          if (now - t0 > p.timeout_ns) {
            if (p.status) *((volatile uint32_t*)p.status) = 0x80000000u | (uint32_t)tid;
            __threadfence_system();
            __trap();
          }
        }
      }
    }
    __syncthreads();
    if (p.mc_local != nullptr) {
      // NVLS: one multimem.ld_reduce per 16 bytes sums all ranks' copies inside the NVSwitch;
      // every rank receives the same switch-computed value, so trees stay identical
      const float* src = p.mc_local + (size_t)bi * hist_floats;
      for (int i = tid; i < hist_floats / 4; i += kHistThreads) {
        float4 v;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "l"(src + 4 * (size_t)i)
                     : "memory");
        reinterpret_cast<float4*>(out)[i] = v;
      }
    } else
    // one-shot reduce in rank order (same order everywhere -> identical trees on every rank)
    for (int i = tid; i < hist_floats / 4; i += kHistThreads) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < KDL_MAX_WORLD; ++r) {
        if (r < p.world) {
          const float4* src = reinterpret_cast<const float4*>(p.peer_local[r] + (size_t)bi * hist_floats) + i;
          float4 v;
          asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                       : "l"(src)
                       : "memory");
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      }
      reinterpret_cast<float4*>(out)[i] = acc;
    }
  }
  if (tid == 0) {
    p.node_done[bi] = 0u;
    if (p.epoch_state) {
      __threadfence();
      const unsigned prev = atomicAdd(p.epoch_state + 1, 1u);
      if (prev == gridDim.y - 1) {          // every node of this launch has read the epoch
        p.epoch_state[1] = 0u;
        *((volatile uint32_t*)p.epoch_state) = epoch;
      }
    }
  }
}

// One block per node: gain scan over bins for every feature, arg-max.
// out[node] = {feature, threshold_bin, gain, GL, HL, G, H, valid}
__global__ void __launch_bounds__(kBins) split_eval_kernel(const float* __restrict__ hist, int F,
                                                           float lambda, float gamma,
                                                           float min_child_weight,
                                                           float* __restrict__ out) {
  __shared__ float sg[kBins], sh[kBins];
  __shared__ float best_gain[kBins / 32];
  __shared__ int best_code[kBins / 32];
  __shared__ float best_gl[kBins / 32], best_hl[kBins / 32];
  const int node = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* H = hist + (size_t)node * F * kBins * 2;
  float my_gain = -1.f, my_gl = 0.f, my_hl = 0.f;
  int my_code = -1;
  float G = 0.f, Hs = 0.f;
  for (int f = 0; f < F; ++f) {
    float2 v = reinterpret_cast<const float2*>(H)[f * kBins + tid];
    // inclusive scan over 256 bins: warp scan + warp totals
    float g = v.x, h = v.y;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      float tg = __shfl_up_sync(0xffffffffu, g, o);
      float th = __shfl_up_sync(0xffffffffu, h, o);
      if (lane >= o) { g += tg; h += th; }
    }
    if (lane == 31) { sg[warp] = g; sh[warp] = h; }
    __syncthreads();
    float og = 0.f, oh = 0.f;
    for (int w = 0; w < warp; ++w) { og += sg[w]; oh += sh[w]; }
    float tg = 0.f, th = 0.f;
    for (int w = 0; w < kBins / 32; ++w) { tg += sg[w]; th += sh[w]; }
    __syncthreads();
    g += og; h += oh;   // GL, HL for "bin <= tid goes left"
    if (f == 0) { G = tg; Hs = th; }
    const float gr = tg - g, hr = th - h;
    if (tid < kBins - 1 && h >= min_child_weight && hr >= min_child_weight) {
      float gain = 0.5f * (g * g / (h + lambda) + gr * gr / (hr + lambda) - tg * tg / (th + lambda)) - gamma;
      if (gain > my_gain) { my_gain = gain; my_code = f * kBins + tid; my_gl = g; my_hl = h; }
    }
  }
  // arg-max across the block (ties -> smallest code, deterministic)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float og = __shfl_xor_sync(0xffffffffu, my_gain, o);
    int oc = __shfl_xor_sync(0xffffffffu, my_code, o);
    float ogl = __shfl_xor_sync(0xffffffffu, my_gl, o);
    float ohl = __shfl_xor_sync(0xffffffffu, my_hl, o);
    if (og > my_gain || (og == my_gain && oc >= 0 && (my_code < 0 || oc < my_code))) {
      my_gain = og; my_code = oc; my_gl = ogl; my_hl = ohl;
    }
  }
  if (lane == 0) { best_gain[warp] = my_gain; best_code[warp] = my_code; best_gl[warp] = my_gl; best_hl[warp] = my_hl; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < kBins / 32; ++w) {
      if (best_gain[w] > my_gain || (best_gain[w] == my_gain && best_code[w] >= 0 && (my_code < 0 || best_code[w] < my_code))) {
        my_gain = best_gain[w]; my_code = best_code[w]; my_gl = best_gl[w]; my_hl = best_hl[w];
      }
    }
    float* o = out + (size_t)node * 8;
    const bool valid = my_code >= 0 && my_gain > 0.f;
    o[0] = valid ? (float)(my_code / kBins) : 0.f;
    o[1] = valid ? (float)(my_code % kBins) : 255.f;   // invalid: everything goes left
    o[2] = valid ? my_gain : 0.f;
    o[3] = valid ? my_gl : G;
    o[4] = valid ? my_hl : Hs;
    o[5] = G;
    o[6] = Hs;
    o[7] = valid ? 1.f : 0.f;
  }
}

__global__ void __launch_bounds__(256) apply_split_kernel(int32_t* __restrict__ node_of_row,
                                                          const uint8_t* __restrict__ bins, int fstride,
                                                          const float* __restrict__ splits, int64_t n) {
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += T) {
    const int node = node_of_row[i];
    const float* s = splits + (size_t)node * 8;
    const int f = (int)s[0];
    const int thr = (int)s[1];
    const int bin = bins[i * fstride + f];
    node_of_row[i] = 2 * node + (bin > thr ? 1 : 0);
  }
}

__global__ void __launch_bounds__(256) update_pred_kernel(float* __restrict__ pred,
                                                          const int32_t* __restrict__ node_of_row,
                                                          const float* __restrict__ leaf_value, int64_t n) {
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += T)
    pred[i] += leaf_value[node_of_row[i]];
}

// ---------------------------------------------------------------------------------------------
// Row partition: rows grouped by tree node with ONE counting-sort pass per level (node ids at
// level d take 2^d <= 256 values).  Replaces torch.sort + scatter_add (4-5 ms per level on 11 M
// rows — most of a boosting round) with three launches that read the node array twice and write the
// row order once (~130 MB per level).  Order inside a node is block-major, arbitrary inside a
// block's chunk: rows of a node stay in ascending-chunk order (good locality for the histogram
// gather) and the histogram sums do not depend on the order beyond fp32 rounding.
constexpr int kPartThreads = 256;
constexpr int kPartMaxNodes = 256;

__device__ __forceinline__ int warp_rank_in_group(unsigned mask, int lane) {
  return __popc(mask & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(kPartThreads) part_count_kernel(const int32_t* __restrict__ node_of_row,
                                                                 int64_t n, int nodes, int64_t chunk,
                                                                 int32_t* __restrict__ block_counts) {
  __shared__ int cnt[kPartMaxNodes];
  for (int i = threadIdx.x; i < nodes; i += kPartThreads) cnt[i] = 0;
  __syncthreads();
  const int64_t lo = (int64_t)blockIdx.x * chunk;
  const int64_t hi = min(lo + chunk, n);
  const int lane = threadIdx.x & 31;
  for (int64_t i0 = lo; i0 < hi; i0 += kPartThreads) {
    const int64_t i = i0 + threadIdx.x;
    const bool ok = i < hi;
    const unsigned act = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const int v = node_of_row[i];
      const unsigned m = __match_any_sync(act, v);
      if (warp_rank_in_group(m, lane) == 0) atomicAdd(&cnt[v], __popc(m));
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nodes; i += kPartThreads) block_counts[(size_t)i * gridDim.x + blockIdx.x] = cnt[i];
}

// Exclusive scan of block_counts (node-major: [nodes][B]) in two small launches: every node's B block counts are
// scanned in place by a block of their own (prefix inside the node), then one block scans the <= 256 node totals into
// seg[] / counts[].  (One 1024-thread block used to walk all nodes x B entries — up to 300 k — which took 37 us per
// level whatever the number of rows, a quarter of a boosting round at 8 GPUs.)
__global__ void __launch_bounds__(256) part_scan_node_kernel(int32_t* __restrict__ block_counts, int B,
                                                            int32_t* __restrict__ node_total) {
  __shared__ int wsum[8];
  int32_t* row = block_counts + (size_t)blockIdx.x * B;
  const int per = (B + 255) / 256;
  const int lo = min((int)threadIdx.x * per, B), hi = min(lo + per, B);
  int sum = 0;
  for (int i = lo; i < hi; ++i) sum += row[i];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = sum;                                    // inclusive scan of the 256 per-thread sums
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += v;
  }
  if (lane == 31) wsum[w] = inc;
  __syncthreads();
  int before = 0, total = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (k < w) before += wsum[k];
    total += wsum[k];
  }
  int run = before + inc - sum;
  for (int i = lo; i < hi; ++i) {
    const int c = row[i];
    row[i] = run;
    run += c;
  }
  if (threadIdx.x == 0) node_total[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kPartMaxNodes) part_scan_final_kernel(const int32_t* __restrict__ node_total,
                                                                       int nodes, int32_t* __restrict__ seg,
                                                                       int32_t* __restrict__ counts) {
  __shared__ int wsum[kPartMaxNodes / 32];
  const int v = threadIdx.x;
  const int c = v < nodes ? node_total[v] : 0;
  const int lane = v & 31, w = v >> 5;
  int inc = c;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += u;
  }
  if (lane == 31) wsum[w] = inc;
  __syncthreads();
  int before = 0, total = 0;
#pragma unroll
  for (int k = 0; k < kPartMaxNodes / 32; ++k) {
    if (k < w) before += wsum[k];
    total += wsum[k];
  }
  if (v < nodes) {
    seg[v] = before + inc - c;
    counts[v] = c;
  }
  if (v == 0) seg[nodes] = total;
}

__global__ void __launch_bounds__(kPartThreads) part_scatter_kernel(const int32_t* __restrict__ node_of_row,
                                                                   int64_t n, int nodes, int64_t chunk,
                                                                   const int32_t* __restrict__ block_offsets,
                                                                   const int32_t* __restrict__ seg,
                                                                   int32_t* __restrict__ ridx) {
  __shared__ int offs[kPartMaxNodes];
  for (int i = threadIdx.x; i < nodes; i += kPartThreads)       // node start + this block's prefix inside the node
    offs[i] = seg[i] + block_offsets[(size_t)i * gridDim.x + blockIdx.x];
  __syncthreads();
  const int64_t lo = (int64_t)blockIdx.x * chunk;
  const int64_t hi = min(lo + chunk, n);
  const int lane = threadIdx.x & 31;
  for (int64_t i0 = lo; i0 < hi; i0 += kPartThreads) {
    const int64_t i = i0 + threadIdx.x;
    const bool ok = i < hi;
    const unsigned act = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const int v = node_of_row[i];
      const unsigned m = __match_any_sync(act, v);
      const int leader = __ffs(m) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(&offs[v], __popc(m));
      base = __shfl_sync(m, base, leader);
      ridx[base + warp_rank_in_group(m, lane)] = (int32_t)i;
    }
  }
}

static int ew_grid(int64_t n) {
  int64_t g = (n + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  return (int)(g < 1 ? 1 : g);
}

}  // namespace kdl

using namespace kdl;

extern "C" int kdl_gbdt_grad_logistic(const float* pred, const float* label, void* gh, int64_t n,
                                      cudaStream_t stream) {
  grad_logistic_kernel<<<ew_grid(n), 256, 0, stream>>>(pred, label, (float2*)gh, n);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" size_t kdl_gbdt_hist_bytes(int nodes, int F) { return (size_t)nodes * F * kBins * 2 * sizeof(float); }

extern "C" int kdl_gbdt_hist_allreduce(const void* bins, int fstride, int F, const void* gh,
                                       const int32_t* ridx, const int32_t* seg, const int32_t* build,
                                       int nbuild, int64_t n_rows_hint, void* local_hist, void* out_hist,
                                       void* node_done, void* const* peer_local, void* const* flag_ptrs,
                                       const void* mc_local, int world, int rank, int max_nodes,
                                       unsigned epoch, void* status,
                                       unsigned long long timeout_ns, float gh_abs_max, void* epoch_state,
                                       cudaStream_t stream) {
  if (F < 1 || F > 32 || fstride < F || (fstride % 16) != 0) {
    kdl_set_error("gbdt_hist: need 1 <= F <= 32 and a 16-byte multiple row stride >= F (F=%d stride=%d)", F, fstride);
    return KDL_ERR_ARG;
  }
  if (nbuild < 1 || nbuild > max_nodes || world < 1 || world > KDL_MAX_WORLD) {
    kdl_set_error("gbdt_hist: bad nbuild=%d (max %d) or world=%d", nbuild, max_nodes, world);
    return KDL_ERR_ARG;
  }
  HistParams p = {};
  p.bins = (const uint8_t*)bins; p.gh = (const float2*)gh; p.ridx = ridx; p.seg = seg; p.build = build;
  p.nbuild = nbuild; p.F = F; p.fstride = fstride;
  p.local = (float*)local_hist; p.out = (float*)out_hist; p.node_done = (unsigned*)node_done;
  p.world = world; p.rank = rank; p.max_nodes = max_nodes; p.epoch = epoch;
  p.timeout_ns = timeout_ns; p.status = (uint32_t*)status;
  p.epoch_state = (uint32_t*)epoch_state;
  if (!(gh_abs_max > 0.f)) gh_abs_max = 1.f;
  p.qscale = 131072.f / gh_abs_max; p.inv_qscale = gh_abs_max / 131072.f;
  p.mc_local = world > 1 ? (const float*)mc_local : nullptr;
  for (int r = 0; r < world; ++r) {
    p.peer_local[r] = peer_local ? (const float*)peer_local[r] : (const float*)local_hist;
    p.flags[r] = flag_ptrs ? (uint32_t*)flag_ptrs[r] : nullptr;
  }
  const size_t smem = (size_t)F * kBins * 2 * sizeof(float);
  static bool attr_done = false;
  if (!attr_done) {
    KDL_CUDA(cudaFuncSetAttribute(hist_allreduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * kBins * 2 * 4));
    attr_done = true;
  }
  // blocks per node: enough to fill the machine across the build list; rows per tile 4096
  p.rows_per_tile = 4096;
  int64_t per_node_rows = n_rows_hint / (nbuild > 0 ? nbuild : 1) + 1;
  int64_t want = (per_node_rows + p.rows_per_tile - 1) / p.rows_per_tile;
  int64_t cap = (148 * 3 + nbuild - 1) / nbuild;   // ~3 resident blocks/SM at 56 KiB smem each
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  dim3 grid((unsigned)want, (unsigned)nbuild);
  hist_allreduce_kernel<<<grid, kHistThreads, smem, stream>>>(p);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_gbdt_split_eval(const void* hist, int nodes, int F, float lambda, float gamma,
                                   float min_child_weight, float* out8, cudaStream_t stream) {
  split_eval_kernel<<<nodes, kBins, 0, stream>>>((const float*)hist, F, lambda, gamma, min_child_weight, out8);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_gbdt_apply_split(int32_t* node_of_row, const void* bins, int fstride,
                                    const float* splits8, int64_t n, cudaStream_t stream) {
  apply_split_kernel<<<ew_grid(n), 256, 0, stream>>>(node_of_row, (const uint8_t*)bins, fstride, splits8, n);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_gbdt_update_pred(float* pred, const int32_t* node_of_row, const float* leaf_value,
                                    int64_t n, cudaStream_t stream) {
  update_pred_kernel<<<ew_grid(n), 256, 0, stream>>>(pred, node_of_row, leaf_value, n);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

// Rows grouped by node: ridx[n] (row ids, node-major), seg[nodes+1] (first slot of every node),
// counts[nodes].  workspace: kdl_gbdt_partition_workspace_bytes(nodes) bytes.
static int part_blocks() { return 148 * 8; }
extern "C" size_t kdl_gbdt_partition_workspace_bytes(int nodes) {
  return ((size_t)nodes * part_blocks() + kPartMaxNodes) * sizeof(int32_t);
}
extern "C" int kdl_gbdt_partition(const int32_t* node_of_row, int64_t n, int nodes, int32_t* ridx, int32_t* seg,
                                  int32_t* counts, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  if (nodes < 1 || nodes > kPartMaxNodes) {
    kdl_set_error("gbdt partition: %d nodes per level unsupported (max %d)", nodes, kPartMaxNodes);
    return KDL_ERR_UNSUPPORTED;
  }
  if (n >= ((int64_t)1 << 31)) {
    kdl_set_error("gbdt partition: %lld rows per rank exceed int32 row ids", (long long)n);
    return KDL_ERR_UNSUPPORTED;
  }
  const int B = part_blocks();
  if (workspace == nullptr || workspace_bytes < ((size_t)nodes * B + kPartMaxNodes) * sizeof(int32_t)) {
    kdl_set_error("gbdt partition: workspace too small");
    return KDL_ERR_WORKSPACE;
  }
  const int64_t chunk = ((n + B - 1) / B + kPartThreads - 1) / kPartThreads * kPartThreads;
  int32_t* bc = (int32_t*)workspace;
  part_count_kernel<<<B, kPartThreads, 0, stream>>>(node_of_row, n, nodes, chunk, bc);
  KDL_CHECK_LAUNCH();
  int32_t* node_total = bc + (size_t)nodes * B;
  part_scan_node_kernel<<<nodes, 256, 0, stream>>>(bc, B, node_total);
  KDL_CHECK_LAUNCH();
  part_scan_final_kernel<<<1, kPartMaxNodes, 0, stream>>>(node_total, nodes, seg, counts);
  KDL_CHECK_LAUNCH();
  part_scatter_kernel<<<B, kPartThreads, 0, stream>>>(node_of_row, n, nodes, chunk, bc, seg, ridx);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

// ---------------------------------------------------------------------------------------------
// Small per-level glue kernels.  A boosting round is ~150 launches and, at 8 GPUs, bound by launch latency
// rather than by any kernel; these three replace about ten tiny tensor-library launches per level.
namespace kdl {

// build[i] = the smaller child of parent i (ties -> left): 2i + (count[2i+1] < count[2i])
__global__ void choose_build_kernel(const void* counts, int is_float, int pairs, int32_t* build) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= pairs) return;
  bool right;
  if (is_float) right = ((const float*)counts)[2 * i + 1] < ((const float*)counts)[2 * i];
  else right = ((const int32_t*)counts)[2 * i + 1] < ((const int32_t*)counts)[2 * i];
  build[i] = 2 * i + (right ? 1 : 0);
}

// level histograms from the built (smaller) children and the parents: sibling = parent - built
__global__ void assemble_hist_kernel(const float4* __restrict__ parent, const float4* __restrict__ built,
                                     const int32_t* __restrict__ build, int nodes, int per_node4,
                                     float4* __restrict__ out) {
  const int64_t total = (int64_t)nodes * per_node4;
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += T) {
    const int node = (int)(i / per_node4), e = (int)(i - (int64_t)node * per_node4);
    const int pr = node >> 1;
    const float4 b = built[(size_t)pr * per_node4 + e];
    if (build[pr] == node) {
      out[i] = b;
    } else {
      const float4 p = parent[(size_t)pr * per_node4 + e];
      out[i] = make_float4(p.x - b.x, p.y - b.y, p.z - b.z, p.w - b.w);
    }
  }
}

// leaf values of the deepest level from its split statistics (feat, thr, gain, GL, HL, G, H, valid)
__global__ void leaf_values_kernel(const float* __restrict__ splits8, int nodes, float lambda, float lr,
                                   float* __restrict__ leaf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nodes) return;
  const float* s = splits8 + (size_t)i * 8;
  const float GL = s[3], HL = s[4], G = s[5], H = s[6];
  leaf[2 * i] = -GL / (HL + lambda) * lr;
  leaf[2 * i + 1] = -(G - GL) / ((H - HL) + lambda) * lr;
}

}  // namespace kdl

extern "C" int kdl_gbdt_choose_build(const void* counts, int counts_are_float, int nodes, int32_t* build,
                                     cudaStream_t stream) {
  const int pairs = nodes / 2;
  if (pairs < 1) return KDL_OK;
  kdl::choose_build_kernel<<<(pairs + 127) / 128, 128, 0, stream>>>(counts, counts_are_float, pairs, build);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_gbdt_assemble_hist(const void* parent_hist, const void* built_hist, const int32_t* build, int nodes,
                                      int F, void* out_hist, cudaStream_t stream) {
  const int per_node4 = F * kdl::kBins * 2 / 4;
  const int64_t total = (int64_t)nodes * per_node4;
  int64_t g = (total + 255) / 256;
  if (g > 148 * 8) g = 148 * 8;
  kdl::assemble_hist_kernel<<<(int)g, 256, 0, stream>>>((const float4*)parent_hist, (const float4*)built_hist, build,
                                                        nodes, per_node4, (float4*)out_hist);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_gbdt_leaf_values(const float* splits8, int nodes, float lambda, float lr, float* leaf,
                                    cudaStream_t stream) {
  kdl::leaf_values_kernel<<<(nodes + 127) / 128, 128, 0, stream>>>(splits8, nodes, lambda, lr, leaf);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}
