// Fused training-mode BatchNorm kernels for NHWC bf16 activations viewed as [M, C]
// (M = N*H*W rows, C channels contiguous).
//
//   forward : (1) per-channel sum / sum-of-squares with the finalisation (mean, rstd,
//                 running stats, scale/shift vectors) done by the last-arriving block,
//             (2) y = relu(a*x + b + residual) in one pass.
//   backward: (1) dbeta/dgamma reduction fused with the ReLU mask, finalised in-kernel into
//                 the three per-channel coefficient vectors of the dx formula,
//             (2) dx = A*dz + B*x + D (and dz written out for the residual branch).
//
// Replaces cuDNN BN + separate ReLU + separate residual add of the stock worker
// (SURVEY.md §2.4 row N8).  No counterpart in the reference operator: it only launches
// opaque worker images (reference example/pytorch/pytorch_job_mnist_mpi.yaml:15).
#include "common.cuh"
#include "../include/kubedl_b200.h"
#include <stdlib.h>

namespace kdl {

struct BnReduceParams {
  // MODE 0 (fwd): a = x.  MODE 1 (bwd): a = dy, b = y (relu output, mask), c = x.
  const bf16* a;
  const bf16* b;
  const bf16* c;
  int64_t M;
  int C;
  int CS;        // channels per slice (blockIdx.y)
  int tpr;       // threads per row-slice = CS/8
  int rpb;       // rows per block iteration
  float* partial;     // [nslices][gx][2*CS]
  unsigned* counter;  // [nslices], zero on entry, zero on exit
  // fwd finalisation
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  float eps;
  float momentum;
  float* mean_rstd;  // [2][C] saved for backward (fwd: out, bwd: in)
  float* coef;       // fwd: [2][C] (a,b); bwd: [4][C] (A,B,D, forward shift b)
  // bwd finalisation
  bf16* dgamma;
  bf16* dbeta;
  int relu;          // bwd: 1 = mask from y (b), 2 = mask recomputed from x with the forward affine (no y read),
                     //      3 = 1 bit per element saved by the forward pass (layers with a residual input)
  const uint8_t* mask;   // [M*C/8], bit j of byte v = element 8v+j was positive
  // bwd: dy is not in memory but the gradient of a 3x3/s2 max-pool output (pool_dp, pool_idx): it is
  // gathered on the fly from the <= 4 windows covering each pixel (rows of this BN = pixels of an H x W map)
  const bf16* pool_dp;
  const uint8_t* pool_idx;
  int pool_H, pool_W;
};

// gradient of the max-pool INPUT at pixel `row` (= (n*H + ih)*W + iw), channels [ch0, ch0+8), rounded to bf16
// exactly like the separate kdl_maxpool3x3s2_bwd pass would store it
__device__ __forceinline__ uint4 pool_grad_gather(const bf16* __restrict__ dp, const uint8_t* __restrict__ idx,
                                                  int64_t row, int ch0, int H, int W, int C) {
  const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
  const int iw = (int)(row % W);
  const int64_t t = row / W;
  const int ih = (int)(t % H);
  const int n = (int)(t / H);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int oh = (ih + 1) / 2 - 1; oh <= (ih + 1) / 2; ++oh) {
    if (oh < 0 || oh >= OH) continue;
    const int kh = ih - (oh * 2 - 1);
    if (kh < 0 || kh > 2) continue;
    for (int ow = (iw + 1) / 2 - 1; ow <= (iw + 1) / 2; ++ow) {
      if (ow < 0 || ow >= OW) continue;
      const int kw = iw - (ow * 2 - 1);
      if (kw < 0 || kw > 2) continue;
      const int64_t o = (((int64_t)n * OH + oh) * OW + ow) * C + ch0;
      const uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
      const uint32_t c4 = (uint32_t)(kh * 3 + kw) * 0x01010101u;
      const uint32_t m0 = __vcmpeq4(pk.x, c4), m1 = __vcmpeq4(pk.y, c4);
      if ((m0 | m1) == 0u) continue;
      float f[8];
      unpack8(ld_plain(dp + o), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t hit = (j < 4 ? (m0 >> (8 * j)) : (m1 >> (8 * (j - 4)))) & 1u;
        if (hit) acc[j] += f[j];
      }
    }
  }
  return pack8(acc);
}

// RELU (backward only) is the compile-time copy of p.relu and POOL of (p.pool_dp != nullptr): each variant keeps
// only the per-channel vectors and loads it needs.  The generic version held mean / rstd / forward scale /
// forward shift for every thread (128 registers, 2 blocks per SM, 25 % occupancy) and spent 136 warp
// instructions per 8-element vector: ncu showed DRAM at 53 % with the issue slots 53 % busy
// (profiles/r2_bn_backward.md).  The second sum is accumulated as sum(dz * x) and centred once, in the
// finaliser: sum(dz * xhat) = rstd * (sum(dz * x) - mean * sum(dz)).
template <int MODE, int RELU, bool POOL>
__global__ void __launch_bounds__(256, (MODE == 1 && (RELU == 1 || RELU == 2 || POOL)) ? 3 : 4) bn_reduce_kernel(BnReduceParams p) {
  __shared__ float red[256 * 16];
  __shared__ int is_last;
  pdl_wait();      // programmatic dependent launch: nothing global is touched before the predecessor has finished
  pdl_launch();
  const int tid = threadIdx.x;
  const int nthreads = blockDim.x;
  const int cg = tid % p.tpr;
  const int r0 = tid / p.tpr;
  const int slice = blockIdx.y;
  const int ch0 = slice * p.CS + cg * 8;
  const int64_t step = (int64_t)gridDim.x * p.rpb;

  float s0[8], s1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s0[j] = s1[j] = 0.f;

  float fwa[8], fwb[8];
  if (MODE == 1 && RELU == 2) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      // the forward pass computed y = relu(fma(a, x, b)), a = gamma*rstd, b = beta - mean*a: the same
      // expression gives the ReLU mask back without reading y (layers without a residual input)
      const float mu = p.mean_rstd[ch0 + j], rs = p.mean_rstd[p.C + ch0 + j];
      fwa[j] = (p.gamma ? p.gamma[ch0 + j] : 1.f) * rs;
      fwb[j] = (p.beta ? p.beta[ch0 + j] : 0.f) - mu * fwa[j];
    }
  }

  constexpr int U = 4;
  for (int64_t row = (int64_t)blockIdx.x * p.rpb + r0; row < p.M; row += step * U) {
    uint4 va[U], vb[U], vc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int64_t r = row + u * step;
      if (r < p.M) {
        int64_t off = r * p.C + ch0;
        if (MODE == 1 && POOL) va[u] = pool_grad_gather(p.pool_dp, p.pool_idx, r, ch0, p.pool_H, p.pool_W, p.C);
        else va[u] = ld_stream(p.a + off);
        if (MODE == 1) {
          vc[u] = ld_stream(p.c + off);
          if (RELU == 1) vb[u] = ld_stream(p.b + off);
          else if (RELU == 3) vb[u].x = p.mask[off >> 3];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int64_t r = row + u * step;
      if (r < p.M) {
        float fa[8];
        unpack8(va[u], fa);
        if (MODE == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s0[j] += fa[j];
            s1[j] = fmaf(fa[j], fa[j], s1[j]);
          }
        } else {
          float fx[8];
          unpack8(vc[u], fx);
          if (RELU == 1) {
            float fy[8];
            unpack8(vb[u], fy);
#pragma unroll
            for (int j = 0; j < 8; ++j) fa[j] = fy[j] > 0.f ? fa[j] : 0.f;
          } else if (RELU == 2) {
#pragma unroll
            for (int j = 0; j < 8; ++j) fa[j] = fmaf(fwa[j], fx[j], fwb[j]) > 0.f ? fa[j] : 0.f;
          } else if (RELU == 3) {
#pragma unroll
            for (int j = 0; j < 8; ++j) fa[j] = ((vb[u].x >> j) & 1u) ? fa[j] : 0.f;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s0[j] += fa[j];
            s1[j] = fmaf(fa[j], fx[j], s1[j]);
          }
        }
      }
    }
  }

  // block reduce over the rpb row lanes
  const int W2 = 2 * p.CS;
  float* mine = red + (size_t)r0 * W2;
  // NOTE: rpb*2*CS <= 256*16 by construction (tpr*rpb <= 256, 16 values per thread)
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mine[cg * 8 + j] = s0[j];
    mine[p.CS + cg * 8 + j] = s1[j];
  }
  __syncthreads();
  float* my_partial = p.partial + ((size_t)slice * gridDim.x + blockIdx.x) * W2;
  for (int o = tid; o < W2; o += nthreads) {
    float acc = 0.f;
    for (int r = 0; r < p.rpb; ++r) acc += red[(size_t)r * W2 + o];
    my_partial[o] = acc;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    unsigned prev = atomicAdd(p.counter + slice, 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();

  // ---- finalisation by the last block of this slice (fixed summation order) ----
  const float* base = p.partial + (size_t)slice * gridDim.x * W2;
  const int nf4 = W2 / 4;
  const int gx = gridDim.x;
  float* fin = red;  // [parts][W2] or [W2]
  if (nf4 <= nthreads) {
    const int parts = nthreads / nf4;
    const int part = tid / nf4;
    const int c4 = tid % nf4;
    if (part < parts) {
      // eight independent loads in flight (with several hundred partial rows and few channels this loop was a
      // serial tail of ~20 us: one L2 round trip per row); the order of the additions is unchanged
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      int bx = part;
      for (; bx + 7 * parts < gx; bx += 8 * parts) {
        float4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldcg(reinterpret_cast<const float4*>(base + (size_t)(bx + u * parts) * W2) + c4);
#pragma unroll
        for (int u = 0; u < 8; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
      }
      for (; bx < gx; bx += parts) {
        float4 v = __ldcg(reinterpret_cast<const float4*>(base + (size_t)bx * W2) + c4);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      reinterpret_cast<float4*>(fin + (size_t)part * W2)[c4] = acc;
    }
    __syncthreads();
    float* fin2 = red + 2048;  // parts*W2 <= 1024 floats, so the two regions are disjoint
    for (int o = tid; o < W2; o += nthreads) {
      float acc = 0.f;
      for (int q = 0; q < parts; ++q) acc += fin[(size_t)q * W2 + o];
      fin2[o] = acc;
    }
    fin = fin2;
  } else {
    for (int c4 = tid; c4 < nf4; c4 += nthreads) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      int bx = 0;
      for (; bx + 4 <= gx; bx += 4) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = __ldcg(reinterpret_cast<const float4*>(base + (size_t)(bx + u) * W2) + c4);
#pragma unroll
        for (int u = 0; u < 4; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
      }
      for (; bx < gx; ++bx) {
        float4 v = __ldcg(reinterpret_cast<const float4*>(base + (size_t)bx * W2) + c4);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      reinterpret_cast<float4*>(fin)[c4] = acc;
    }
  }
  __syncthreads();
  const float invM = 1.f / (float)p.M;
  for (int cl = tid; cl < p.CS; cl += nthreads) {
    const int c = slice * p.CS + cl;
    const float S0 = fin[cl], S1 = fin[p.CS + cl];
    if (MODE == 0) {
      float mu = S0 * invM;
      float var = fmaxf(S1 * invM - mu * mu, 0.f);
      float rs = rsqrtf(var + p.eps);
      float g = p.gamma ? p.gamma[c] : 1.f;
      float bt = p.beta ? p.beta[c] : 0.f;
      float a = g * rs;
      p.mean_rstd[c] = mu;
      p.mean_rstd[p.C + c] = rs;
      p.coef[c] = a;
      p.coef[p.C + c] = bt - mu * a;
      if (p.running_mean) {
        float unbiased = p.M > 1 ? var * ((float)p.M / (float)(p.M - 1)) : var;
        p.running_mean[c] = (1.f - p.momentum) * p.running_mean[c] + p.momentum * mu;
        p.running_var[c] = (1.f - p.momentum) * p.running_var[c] + p.momentum * unbiased;
      }
    } else {
      float mu = p.mean_rstd[c], rs = p.mean_rstd[p.C + c];
      float g = p.gamma ? p.gamma[c] : 1.f;
      float a = g * rs;
      const float S1c = rs * (S1 - mu * S0);      // sum(dz * xhat) from sum(dz * x) and sum(dz)
      float c1 = S0 * invM, c2 = S1c * invM;
      if (p.dbeta) p.dbeta[c] = __float2bfloat16(S0);
      if (p.dgamma) p.dgamma[c] = __float2bfloat16(S1c);
      // dx = a*(dz - c1 - xhat*c2) = A*dz + B*x + D
      p.coef[c] = a;
      p.coef[p.C + c] = -a * c2 * rs;
      p.coef[2 * p.C + c] = a * (c2 * rs * mu - c1);
      p.coef[3 * p.C + c] = (p.beta ? p.beta[c] : 0.f) - mu * a;
    }
  }
  if (tid == 0) p.counter[slice] = 0u;
}

// y = relu?(a*x + b (+res)).  Flat vector index; when the total thread count is a multiple
// of C/8 each thread keeps its channel group (and a,b in registers) across iterations.
template <bool FIXED_CG>
__global__ void __launch_bounds__(256) bn_apply_kernel(const bf16* __restrict__ x,
                                                       const bf16* __restrict__ res,
                                                       bf16* __restrict__ y, int64_t nvec, int cvec,
                                                       int C, const float* __restrict__ coef,
                                                       int relu, uint8_t* __restrict__ mask = nullptr,
                                                       int evict_first = 0) {
  // x and the residual are not read again before the backward pass: let them leave L2 first, so that y (read by
  // the very next kernel) stays
  const uint64_t pol = l2_policy(evict_first != 0);
  pdl_wait();
  pdl_launch();
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float a[8], b[8];
  if (FIXED_CG) {
    int cg = (int)(i % cvec);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      a[j] = coef[cg * 8 + j];
      b[j] = coef[C + cg * 8 + j];
    }
  }
  constexpr int U = 4;
  for (; i < nvec; i += T * U) {
    uint4 vx[U], vr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int64_t k = i + u * T;
      if (k < nvec) {
        vx[u] = ld_stream_hint(x + k * 8, pol);
        if (res) vr[u] = ld_stream_hint(res + k * 8, pol);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int64_t k = i + u * T;
      if (k < nvec) {
        if (!FIXED_CG) {
          int cg = (int)(k % cvec);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            a[j] = coef[cg * 8 + j];
            b[j] = coef[C + cg * 8 + j];
          }
        }
        float f[8];
        unpack8(vx[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = fmaf(a[j], f[j], b[j]);
        if (res) {
          float r[8];
          unpack8(vr[u], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] += r[j];
        }
        if (relu) {
          if (mask) {      // 1 bit per element for the backward pass: 1/16 of the bytes of y
            unsigned bits = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) bits |= (f[j] > 0.f ? 1u : 0u) << j;
            mask[k] = (uint8_t)bits;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        *reinterpret_cast<uint4*>(y + k * 8) = pack8(f);
      }
    }
  }
}

// dx = A*dz + B*x + D with dz = dy * relu'(y); optionally dres = dz.  RELU / POOL: compile-time copies of the
// mask mode and of (pool_dp != nullptr), as in bn_reduce_kernel.
template <bool FIXED_CG, int RELU, bool POOL>
__global__ void __launch_bounds__(256, 3) bn_dx_kernel(const bf16* dy,
                                                       const bf16* __restrict__ x,
                                                       const bf16* __restrict__ y, bf16* dx, bf16* dres,
                                                       int64_t nvec, int cvec, int C,
                                                       const float* __restrict__ coef,
                                                       const uint8_t* __restrict__ mask,
                                                       const bf16* __restrict__ pool_dp,
                                                       const uint8_t* __restrict__ pool_idx,
                                                       int pool_H, int pool_W, int evict_first) {
  const uint64_t pol = l2_policy(evict_first != 0);     // dy and x: last use (the reduction read them just before)
  pdl_wait();
  pdl_launch();
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float A[8], B[8], D[8], Fb[8];
  if (FIXED_CG) {
    int cg = (int)(i % cvec);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      A[j] = coef[cg * 8 + j];
      B[j] = coef[C + cg * 8 + j];
      D[j] = coef[2 * C + cg * 8 + j];
      if (RELU == 2) Fb[j] = coef[3 * C + cg * 8 + j];
    }
  }
  constexpr int U = 2;
  for (; i < nvec; i += T * U) {
    uint4 vd[U], vx[U], vy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int64_t k = i + u * T;
      if (k < nvec) {
        if (POOL) vd[u] = pool_grad_gather(pool_dp, pool_idx, k / cvec, (int)(k % cvec) * 8, pool_H, pool_W, C);
        else vd[u] = ld_plain_hint(dy + k * 8, pol);  // may alias dres: plain (coherent) load
        vx[u] = ld_stream_hint(x + k * 8, pol);
        if (RELU == 1) vy[u] = ld_stream(y + k * 8);
        else if (RELU == 3) vy[u].x = mask[k];
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int64_t k = i + u * T;
      if (k < nvec) {
        if (!FIXED_CG) {
          int cg = (int)(k % cvec);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            A[j] = coef[cg * 8 + j];
            B[j] = coef[C + cg * 8 + j];
            D[j] = coef[2 * C + cg * 8 + j];
            if (RELU == 2) Fb[j] = coef[3 * C + cg * 8 + j];
          }
        }
        float fd[8], fx[8];
        unpack8(vd[u], fd);
        unpack8(vx[u], fx);
        if (RELU == 1) {
          float fy[8];
          unpack8(vy[u], fy);
#pragma unroll
          for (int j = 0; j < 8; ++j) fd[j] = fy[j] > 0.f ? fd[j] : 0.f;
        } else if (RELU == 2) {      // A = gamma*rstd is also the forward scale
#pragma unroll
          for (int j = 0; j < 8; ++j) fd[j] = fmaf(A[j], fx[j], Fb[j]) > 0.f ? fd[j] : 0.f;
        } else if (RELU == 3) {
#pragma unroll
          for (int j = 0; j < 8; ++j) fd[j] = ((vy[u].x >> j) & 1u) ? fd[j] : 0.f;
        }
        if (dres) *reinterpret_cast<uint4*>(dres + k * 8) = pack8(fd);
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = fmaf(A[j], fd[j], fmaf(B[j], fx[j], D[j]));
        *reinterpret_cast<uint4*>(dx + k * 8) = pack8(o);
      }
    }
  }
}

struct BnPlan {
  int CS, tpr, rpb, threads, nslices, gx;
  size_t partial_floats;
};

static bool make_plan(int64_t M, int C, BnPlan* pl) {
  if (C % 8 != 0 || C <= 0 || M <= 0) return false;
  if (C % 64 == 0) {
    pl->CS = 64;
  } else {
    if (C > 2048) return false;
    pl->CS = C;
  }
  pl->tpr = pl->CS / 8;
  pl->rpb = 256 / pl->tpr;
  if (pl->rpb < 1) pl->rpb = 1;
  pl->threads = pl->tpr * pl->rpb;
  pl->nslices = C / pl->CS;
  int64_t want = (M + (int64_t)pl->rpb * 8 - 1) / ((int64_t)pl->rpb * 8);  // >= 8 iterations each
  int64_t cap = (148 * 4) / pl->nslices;
  if (cap < 1) cap = 1;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  pl->gx = (int)want;
  pl->partial_floats = (size_t)pl->nslices * pl->gx * 2 * pl->CS;
  return true;
}

static const size_t kCounterBytes = 256;  // up to 64 slice counters

}  // namespace kdl

using namespace kdl;

extern "C" size_t kdl_bn_workspace_bytes(int64_t M, int C) {
  BnPlan pl;
  if (!make_plan(M, C, &pl)) return 0;
  // counters | coef[4][C] | partials
  return kCounterBytes + (size_t)4 * C * sizeof(float) + pl.partial_floats * sizeof(float);
}

static int carve(void* ws, size_t ws_bytes, int64_t M, int C, BnPlan* pl, unsigned** counter,
                 float** coef, float** partial) {
  if (!make_plan(M, C, pl)) {
    kdl_set_error("bn: unsupported shape M=%lld C=%d (need C%%8==0)", (long long)M, C);
    return KDL_ERR_ARG;
  }
  if (pl->nslices > 64) {
    kdl_set_error("bn: C=%d too large", C);
    return KDL_ERR_ARG;
  }
  size_t need = kdl_bn_workspace_bytes(M, C);
  if (ws == nullptr || ws_bytes < need) {
    kdl_set_error("bn: workspace too small (%zu < %zu)", ws_bytes, need);
    return KDL_ERR_WORKSPACE;
  }
  char* b = (char*)ws;
  *counter = (unsigned*)b;
  *coef = (float*)(b + kCounterBytes);
  *partial = *coef + (size_t)4 * C;
  return KDL_OK;
}

static int apply_grid(int64_t nvec) {
  int64_t g = (nvec + 256 * 4 - 1) / (256 * 4);
  if (g > 148 * 8) g = 148 * 8;
  if (g < 1) g = 1;
  return (int)g;
}

// KDL_BN_EVICT_FIRST=0: default L2 priority for the last-use reads of bn_apply / bn_dx (A/B runs)
static int bn_evict_first() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("KDL_BN_EVICT_FIRST");
    v = e ? atoi(e) : 1;
  }
  return v;
}

static int bn_fwd_impl(const void* x, const void* res, void* y, int64_t M, int C, const float* gamma,
                       const float* beta, float* running_mean, float* running_var, float eps, float momentum,
                       int relu, float* save_mean_rstd, uint8_t* relu_mask, void* workspace,
                       size_t workspace_bytes, cudaStream_t stream) {
  BnPlan pl;
  unsigned* counter;
  float *coef, *partial;
  int rc = carve(workspace, workspace_bytes, M, C, &pl, &counter, &coef, &partial);
  if (rc) return rc;
  BnReduceParams p = {};
  p.a = (const bf16*)x;
  p.M = M; p.C = C; p.CS = pl.CS; p.tpr = pl.tpr; p.rpb = pl.rpb;
  p.partial = partial; p.counter = counter;
  p.gamma = gamma; p.beta = beta;
  p.running_mean = running_mean; p.running_var = running_var;
  p.eps = eps; p.momentum = momentum;
  p.mean_rstd = save_mean_rstd; p.coef = coef;
  KDL_CUDA(kdl_launch_pdl(bn_reduce_kernel<0, 0, false>, dim3(pl.gx, pl.nslices), dim3(pl.threads), 0, stream, p));
  int64_t nvec = M * C / 8;
  int cvec = C / 8;
  int g = apply_grid(nvec);
  bool fixed = (256 % cvec) == 0;
  if (fixed)
    KDL_CUDA(kdl_launch_pdl(bn_apply_kernel<true>, dim3(g), dim3(256), 0, stream, (const bf16*)x, (const bf16*)res,
                            (bf16*)y, nvec, cvec, C, (const float*)coef, relu, relu_mask, bn_evict_first()));
  else
    KDL_CUDA(kdl_launch_pdl(bn_apply_kernel<false>, dim3(g), dim3(256), 0, stream, (const bf16*)x, (const bf16*)res,
                            (bf16*)y, nvec, cvec, C, (const float*)coef, relu, relu_mask, bn_evict_first()));
  return KDL_OK;
}

extern "C" int kdl_bn_fwd(const void* x, const void* res, void* y, int64_t M, int C,
                          const float* gamma, const float* beta, float* running_mean,
                          float* running_var, float eps, float momentum, int relu,
                          float* save_mean_rstd, void* workspace, size_t workspace_bytes,
                          cudaStream_t stream) {
  return bn_fwd_impl(x, res, y, M, C, gamma, beta, running_mean, running_var, eps, momentum, relu, save_mean_rstd,
                     nullptr, workspace, workspace_bytes, stream);
}

// Same, and additionally one bit per element (relu_mask, M*C/8 bytes): "was positive before the ReLU".
// kdl_bn_bwd_mask reads that instead of the 16x larger y — for layers whose ReLU follows a residual add,
// where the mask cannot be recomputed from x alone.
extern "C" int kdl_bn_fwd_mask(const void* x, const void* res, void* y, int64_t M, int C,
                               const float* gamma, const float* beta, float* running_mean,
                               float* running_var, float eps, float momentum, int relu,
                               float* save_mean_rstd, void* relu_mask, void* workspace, size_t workspace_bytes,
                               cudaStream_t stream) {
  return bn_fwd_impl(x, res, y, M, C, gamma, beta, running_mean, running_var, eps, momentum, relu, save_mean_rstd,
                     (uint8_t*)relu_mask, workspace, workspace_bytes, stream);
}

template <bool FIXED_CG>
static cudaError_t launch_bn_dx(int relu, bool pool, dim3 g, cudaStream_t stream, const bf16* dy, const bf16* x,
                                const bf16* y, bf16* dx, bf16* dres, int64_t nvec, int cvec, int C, const float* coef,
                                const uint8_t* mask, const bf16* pool_dp, const uint8_t* pool_idx, int pool_H,
                                int pool_W) {
#define KDL_BN_DX(R, P)                                                                                              \
  return kdl_launch_pdl(bn_dx_kernel<FIXED_CG, R, P>, g, dim3(256), 0, stream, dy, x, y, dx, dres, nvec, cvec, C, coef, \
                        mask, pool_dp, pool_idx, pool_H, pool_W, bn_evict_first())
  if (pool) { KDL_BN_DX(2, true); }
  if (relu == 0) { KDL_BN_DX(0, false); }
  if (relu == 1) { KDL_BN_DX(1, false); }
  if (relu == 2) { KDL_BN_DX(2, false); }
  KDL_BN_DX(3, false);
#undef KDL_BN_DX
}

static int bn_bwd_impl(const void* dy, const void* x, const void* y, void* dx, void* dres,
                       int64_t M, int C, const float* gamma, const float* beta, const float* save_mean_rstd,
                       int relu, const uint8_t* relu_mask, void* dgamma_bf16, void* dbeta_bf16, void* workspace,
                       size_t workspace_bytes, cudaStream_t stream, const bf16* pool_dp = nullptr,
                       const uint8_t* pool_idx = nullptr, int pool_H = 0, int pool_W = 0) {
  BnPlan pl;
  unsigned* counter;
  float *coef, *partial;
  int rc = carve(workspace, workspace_bytes, M, C, &pl, &counter, &coef, &partial);
  if (rc) return rc;
  if (relu && relu_mask != nullptr) {
    relu = 3;
  } else if (relu && y == nullptr) {
    if (beta == nullptr) {
      kdl_set_error("bn_bwd: the relu mask needs y, or beta to recompute it from x");
      return KDL_ERR_ARG;
    }
    relu = 2;
  } else if (relu) {
    relu = 1;
  }
  BnReduceParams p = {};
  p.a = (const bf16*)dy; p.b = (const bf16*)y; p.c = (const bf16*)x;
  p.M = M; p.C = C; p.CS = pl.CS; p.tpr = pl.tpr; p.rpb = pl.rpb;
  p.partial = partial; p.counter = counter;
  p.gamma = gamma; p.beta = beta;
  p.mean_rstd = const_cast<float*>(save_mean_rstd);
  p.coef = coef;
  p.dgamma = (bf16*)dgamma_bf16; p.dbeta = (bf16*)dbeta_bf16;
  p.relu = relu; p.mask = relu_mask;
  p.pool_dp = pool_dp; p.pool_idx = pool_idx; p.pool_H = pool_H; p.pool_W = pool_W;
  {
    const dim3 rg(pl.gx, pl.nslices), rb(pl.threads);
    if (pool_dp != nullptr) {
      if (relu != 2) {
        kdl_set_error("bn_bwd: the pooled variant recomputes the ReLU mask from x (relu mode 2), got %d", relu);
        return KDL_ERR_ARG;
      }
      KDL_CUDA(kdl_launch_pdl(bn_reduce_kernel<1, 2, true>, rg, rb, 0, stream, p));
    } else if (relu == 0) KDL_CUDA(kdl_launch_pdl(bn_reduce_kernel<1, 0, false>, rg, rb, 0, stream, p));
    else if (relu == 1) KDL_CUDA(kdl_launch_pdl(bn_reduce_kernel<1, 1, false>, rg, rb, 0, stream, p));
    else if (relu == 2) KDL_CUDA(kdl_launch_pdl(bn_reduce_kernel<1, 2, false>, rg, rb, 0, stream, p));
    else KDL_CUDA(kdl_launch_pdl(bn_reduce_kernel<1, 3, false>, rg, rb, 0, stream, p));
  }
  int64_t nvec = M * C / 8;
  int cvec = C / 8;
  int64_t g64 = (nvec + 256 * 2 - 1) / (256 * 2);
  if (g64 > 148 * 8) g64 = 148 * 8;
  int g = (int)(g64 < 1 ? 1 : g64);
  bool fixed = (256 % cvec) == 0;
  if (fixed)
    KDL_CUDA(launch_bn_dx<true>(relu, pool_dp != nullptr, dim3(g), stream, (const bf16*)dy, (const bf16*)x, (const bf16*)y,
                                (bf16*)dx, (bf16*)dres, nvec, cvec, C, (const float*)coef, relu_mask, pool_dp, pool_idx,
                                pool_H, pool_W));
  else
    KDL_CUDA(launch_bn_dx<false>(relu, pool_dp != nullptr, dim3(g), stream, (const bf16*)dy, (const bf16*)x, (const bf16*)y,
                                 (bf16*)dx, (bf16*)dres, nvec, cvec, C, (const float*)coef, relu_mask, pool_dp, pool_idx,
                                 pool_H, pool_W));
  return KDL_OK;
}

extern "C" int kdl_bn_bwd(const void* dy, const void* x, const void* y, void* dx, void* dres,
                          int64_t M, int C, const float* gamma, const float* beta, const float* save_mean_rstd,
                          int relu, void* dgamma_bf16, void* dbeta_bf16, void* workspace,
                          size_t workspace_bytes, cudaStream_t stream) {
  return bn_bwd_impl(dy, x, y, dx, dres, M, C, gamma, beta, save_mean_rstd, relu, nullptr, dgamma_bf16, dbeta_bf16,
                     workspace, workspace_bytes, stream);
}

// Backward of relu(bn(x)) followed by a 3x3/s2 max-pool, given the gradient of the POOLED output (dp,
// [N, (H+1)/2, (W+1)/2, C]) and the pool's argmax indices: the gradient of the pool input is gathered on
// the fly inside both BatchNorm passes instead of being written (and read twice) as a full-size tensor.
// The ReLU mask is recomputed from x (no residual input).
extern "C" int kdl_bn_bwd_pooled(const void* dp, const void* pool_idx_u8, const void* x, void* dx, int N, int H,
                                 int W, int C, const float* gamma, const float* beta, const float* save_mean_rstd,
                                 void* dgamma_bf16, void* dbeta_bf16, void* workspace, size_t workspace_bytes,
                                 cudaStream_t stream) {
  if (dp == nullptr || pool_idx_u8 == nullptr || beta == nullptr) {
    kdl_set_error("bn_bwd_pooled: dp, pool_idx and beta are required");
    return KDL_ERR_ARG;
  }
  return bn_bwd_impl(nullptr, x, nullptr, dx, nullptr, (int64_t)N * H * W, C, gamma, beta, save_mean_rstd, 1, nullptr,
                     dgamma_bf16, dbeta_bf16, workspace, workspace_bytes, stream, (const bf16*)dp,
                     (const uint8_t*)pool_idx_u8, H, W);
}

// Backward with the ReLU mask taken from the bit mask kdl_bn_fwd_mask / kdl_bn_fwd_from_stats_mask saved.
extern "C" int kdl_bn_bwd_mask(const void* dy, const void* x, const void* relu_mask, void* dx, void* dres,
                               int64_t M, int C, const float* gamma, const float* save_mean_rstd,
                               void* dgamma_bf16, void* dbeta_bf16, void* workspace, size_t workspace_bytes,
                               cudaStream_t stream) {
  if (relu_mask == nullptr) {
    kdl_set_error("bn_bwd_mask: relu_mask is null");
    return KDL_ERR_ARG;
  }
  return bn_bwd_impl(dy, x, nullptr, dx, dres, M, C, gamma, nullptr, save_mean_rstd, 1, (const uint8_t*)relu_mask,
                     dgamma_bf16, dbeta_bf16, workspace, workspace_bytes, stream);
}

// BN forward when the per-channel sums were already produced by the GEMM epilogue
// (kdl_gemm_bf16_ex `stats_partial`): finalise [nparts][2][C] partials in a fixed order, then apply.
// 32 channels x 32 part-lanes per block: the partials of a channel are summed by 32 threads in a fixed
// interleaving and then in a fixed order (deterministic); a single thread walking several hundred
// partial rows took tens of microseconds per BatchNorm layer.
__global__ void __launch_bounds__(1024) bn_finalize_partials_kernel(const float* __restrict__ partial, int nparts,
                                                                    int64_t M, int C, const float* gamma,
                                                                    const float* beta, float* running_mean,
                                                                    float* running_var, float eps, float momentum,
                                                                    float* mean_rstd, float* coef) {
  __shared__ float r0[32][33], r1[32][33];
  pdl_wait();
  pdl_launch();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float a0 = 0.f, a1 = 0.f;
  if (c < C) {
    int i = ty;
    for (; i + 96 < nparts; i += 128) {     // four independent pairs of loads in flight
      float u0[4], u1[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        u0[u] = __ldcg(partial + (size_t)(i + 32 * u) * 2 * C + c);
        u1[u] = __ldcg(partial + (size_t)(i + 32 * u) * 2 * C + C + c);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) { a0 += u0[u]; a1 += u1[u]; }
    }
    for (; i < nparts; i += 32) {
      a0 += __ldcg(partial + (size_t)i * 2 * C + c);
      a1 += __ldcg(partial + (size_t)i * 2 * C + C + c);
    }
  }
  r0[ty][tx] = a0; r1[ty][tx] = a1;
  __syncthreads();
  if (ty != 0 || c >= C) return;
  float S0 = 0.f, S1 = 0.f;
#pragma unroll
  for (int i = 0; i < 32; ++i) { S0 += r0[i][tx]; S1 += r1[i][tx]; }
  const float invM = 1.f / (float)M;
  const float mu = S0 * invM;
  const float var = fmaxf(S1 * invM - mu * mu, 0.f);
  const float rs = rsqrtf(var + eps);
  const float a = (gamma ? gamma[c] : 1.f) * rs;
  mean_rstd[c] = mu;
  mean_rstd[C + c] = rs;
  coef[c] = a;
  coef[C + c] = (beta ? beta[c] : 0.f) - mu * a;
  if (running_mean) {
    const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mu;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

static int bn_fwd_from_stats_impl(const void* x, const void* res, void* y, int64_t M, int C,
                                  const float* gamma, const float* beta, float* running_mean,
                                  float* running_var, float eps, float momentum, int relu,
                                  float* save_mean_rstd, const float* stats_partial, int nparts, uint8_t* relu_mask,
                                  void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  BnPlan pl;
  unsigned* counter;
  float *coef, *partial;
  int rc = carve(workspace, workspace_bytes, M, C, &pl, &counter, &coef, &partial);
  if (rc) return rc;
  if (nparts < 1 || stats_partial == nullptr) {
    kdl_set_error("bn_fwd_from_stats: no partial statistics given");
    return KDL_ERR_ARG;
  }
  KDL_CUDA(kdl_launch_pdl(bn_finalize_partials_kernel, dim3((C + 31) / 32), dim3(1024), 0, stream, stats_partial, nparts,
                          M, C, gamma, beta, running_mean, running_var, eps, momentum, save_mean_rstd, coef));
  int64_t nvec = M * C / 8;
  int cvec = C / 8;
  int g = apply_grid(nvec);
  if ((256 % cvec) == 0)
    KDL_CUDA(kdl_launch_pdl(bn_apply_kernel<true>, dim3(g), dim3(256), 0, stream, (const bf16*)x, (const bf16*)res,
                            (bf16*)y, nvec, cvec, C, (const float*)coef, relu, relu_mask, bn_evict_first()));
  else
    KDL_CUDA(kdl_launch_pdl(bn_apply_kernel<false>, dim3(g), dim3(256), 0, stream, (const bf16*)x, (const bf16*)res,
                            (bf16*)y, nvec, cvec, C, (const float*)coef, relu, relu_mask, bn_evict_first()));
  return KDL_OK;
}

// Only the finalisation: mean / rstd (save_mean_rstd), running statistics and the forward coefficients
// coef_out = [2][C] (a = gamma*rstd, b = beta - mean*a) from epilogue partials — for consumers that apply
// the affine themselves (kdl_maxpool3x3s2_affine_relu_fwd).
extern "C" int kdl_bn_finalize_from_stats(int64_t M, int C, const float* gamma, const float* beta,
                                          float* running_mean, float* running_var, float eps, float momentum,
                                          float* save_mean_rstd, const float* stats_partial, int nparts,
                                          float* coef_out, cudaStream_t stream) {
  if (nparts < 1 || stats_partial == nullptr || coef_out == nullptr || save_mean_rstd == nullptr) {
    kdl_set_error("bn_finalize_from_stats: missing partial statistics or outputs");
    return KDL_ERR_ARG;
  }
  KDL_CUDA(kdl_launch_pdl(bn_finalize_partials_kernel, dim3((C + 31) / 32), dim3(1024), 0, stream, stats_partial, nparts,
                          M, C, gamma, beta, running_mean, running_var, eps, momentum, save_mean_rstd, coef_out));
  return KDL_OK;
}

extern "C" int kdl_bn_fwd_from_stats(const void* x, const void* res, void* y, int64_t M, int C,
                                     const float* gamma, const float* beta, float* running_mean,
                                     float* running_var, float eps, float momentum, int relu,
                                     float* save_mean_rstd, const float* stats_partial, int nparts,
                                     void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  return bn_fwd_from_stats_impl(x, res, y, M, C, gamma, beta, running_mean, running_var, eps, momentum, relu,
                                save_mean_rstd, stats_partial, nparts, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int kdl_bn_fwd_from_stats_mask(const void* x, const void* res, void* y, int64_t M, int C,
                                          const float* gamma, const float* beta, float* running_mean,
                                          float* running_var, float eps, float momentum, int relu,
                                          float* save_mean_rstd, const float* stats_partial, int nparts,
                                          void* relu_mask, void* workspace, size_t workspace_bytes,
                                          cudaStream_t stream) {
  return bn_fwd_from_stats_impl(x, res, y, M, C, gamma, beta, running_mean, running_var, eps, momentum, relu,
                                save_mean_rstd, stats_partial, nparts, (uint8_t*)relu_mask, workspace, workspace_bytes,
                                stream);
}

// ---------------------------------------------------------------------------------------------
// kdl_bn_fwd_static_scratch: the same fused BN forward for callers that must not allocate.
//
// Who that is: a process that has already given every byte of the device to some other owner (a
// serving runtime whose memory pool holds the whole GPU, a pod running right at its memory
// limit), or a caller sitting on a latency-critical path where cudaMalloc would stall it —
// cudaMalloc synchronises the whole device and can fail outright when the pool owner left
// nothing free.  Such a caller cannot size and own the workspace that kdl_bn_fwd asks for.
//
// What it does instead: the scratch (block partials, per-channel coefficients, arrival counters)
// lives in the fixed array below, KDL_BN_STATIC_SCRATCH_BYTES per GPU, resident from the moment
// the library is loaded on that GPU until the process exits.  Shapes whose scratch need is larger
// are refused with an error; nothing is ever grown.  No cudaMalloc, cudaFree or other allocating
// routine is called here.
//
// Rules for the caller:
//   * The first call on each device looks up the array's device address (once, with that device
//     current).  Make that first call during start-up, before any stream capture begins.
//   * All calls on one GPU share the one array and it keeps whatever the previous call left in it.
//     The arrival counters are cleared at the start of every call; everything else is fully
//     rewritten before it is read.  Two calls that overlap on the GPU (different streams, or
//     different host threads) would trample each other's partial sums: serialise them.
static __device__ __align__(256) unsigned char g_bn_static_scratch[KDL_BN_STATIC_SCRATCH_BYTES];

#include <mutex>

static int bn_static_scratch_ptr(void** out) {
  // In host code `g_bn_static_scratch` is only a handle for the runtime, not a device address;
  // the real address differs per device and must be asked for with that device current.
  static std::mutex mu;
  static void* ptrs[64] = {nullptr};
  int dev = -1;
  KDL_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    kdl_set_error("bn_fwd_static_scratch: device index %d out of range", dev);
    return KDL_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_bn_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      kdl_set_error("bn_fwd_static_scratch: cudaGetSymbolAddress failed on device %d: %s", dev,
                    cudaGetErrorString(e));
      return KDL_ERR_CUDA;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return KDL_OK;
}

extern "C" size_t kdl_bn_static_scratch_limit(void) { return (size_t)KDL_BN_STATIC_SCRATCH_BYTES; }

extern "C" int kdl_bn_fwd_static_scratch(const void* x, const void* res, void* y, int64_t M, int C,
                                         const float* gamma, const float* beta,
                                         float* running_mean, float* running_var, float eps,
                                         float momentum, int relu, float* save_mean_rstd,
                                         cudaStream_t stream) {
  const size_t need = kdl_bn_workspace_bytes(M, C);
  if (need == 0) {
    kdl_set_error("bn_fwd_static_scratch: unsupported shape M=%lld C=%d (need C%%8==0)",
                  (long long)M, C);
    return KDL_ERR_ARG;
  }
  if (need > (size_t)KDL_BN_STATIC_SCRATCH_BYTES) {
    kdl_set_error(
        "bn_fwd_static_scratch: M=%lld C=%d needs %zu bytes of scratch but the built-in array "
        "holds %d; use kdl_bn_fwd with a caller-owned workspace for this shape",
        (long long)M, C, need, KDL_BN_STATIC_SCRATCH_BYTES);
    return KDL_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  int rc = bn_static_scratch_ptr(&scratch);
  if (rc) return rc;
  // the arrival counters (first 256 bytes) are the only part that must start at zero
  KDL_CUDA(cudaMemsetAsync(scratch, 0, 256, stream));
  return kdl_bn_fwd(x, res, y, M, C, gamma, beta, running_mean, running_var, eps, momentum, relu,
                    save_mean_rstd, scratch, (size_t)KDL_BN_STATIC_SCRATCH_BYTES, stream);
}
