// Memory-bound fused kernels of the DDP worker: input normalisation,
// softmax cross-entropy forward+backward, momentum-SGD with fp32 master weights, NHWC max-pool /
// global average pool, residual gradient add.  All 128-bit vectorised, bf16 storage, fp32 math.
// The reference operator has no worker code; these replace the ATen/cuDNN kernels of the stock
// image it would launch (reference example/pytorch/pytorch_job_mnist_mpi.yaml:15).
#include "common.cuh"
#include "../include/kubedl_b200.h"
#include <float.h>

namespace kdl {

// uint8 NHWC (C=3) -> bf16 NHWC with per-channel (x/255 - mean)/std, optionally zero-padded to
// Cpad channels (Cpad = 3, 4 or 8) so the stem convolution can run as a tensor-core GEMM.
__global__ void __launch_bounds__(256) u8_normalize_kernel(const uint8_t* __restrict__ in,
                                                           bf16* __restrict__ out, int64_t npix,
                                                           int cpad, float m0, float m1, float m2,
                                                           float s0, float s1, float s2) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  const float k = 1.f / 255.f;
  for (; i < npix; i += T) {
    const uint8_t* p = in + i * 3;
    float r = ((float)p[0] * k - m0) * s0;
    float g = ((float)p[1] * k - m1) * s1;
    float b = ((float)p[2] * k - m2) * s2;
    bf16* o = out + i * cpad;
    o[0] = __float2bfloat16(r);
    o[1] = __float2bfloat16(g);
    o[2] = __float2bfloat16(b);
    for (int c = 3; c < cpad; ++c) o[c] = __float2bfloat16(0.f);
  }
}

// One block per row: loss_i = logsumexp(z) - z[label]; dlogits = (softmax - onehot) * gscale.
// The mean loss is produced by the last-arriving block (fixed order sum over rows).
template <typename TIn>
__global__ void __launch_bounds__(256) softmax_xent_kernel(const TIn* __restrict__ logits,
                                                           const int64_t* __restrict__ labels,
                                                           bf16* __restrict__ dlogits,
                                                           float* __restrict__ row_loss,
                                                           float* __restrict__ loss_out,
                                                           unsigned* counter, int B, int V, int ldv,
                                                           float gscale) {
  __shared__ float sred[32];
  __shared__ float sval;
  __shared__ int is_last;
  const int row = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const TIn* z = logits + (size_t)row * ldv;
  float mx = -FLT_MAX;
  for (int j = tid; j < V; j += blockDim.x) mx = fmaxf(mx, (float)z[j]);
  mx = warp_max(mx);
  if (lane == 0) sred[warp] = mx;
  __syncthreads();
  if (warp == 0) {
    float v = lane < nw ? sred[lane] : -FLT_MAX;
    v = warp_max(v);
    if (lane == 0) sval = v;
  }
  __syncthreads();
  mx = sval;
  __syncthreads();
  float sum = 0.f;
  for (int j = tid; j < V; j += blockDim.x) sum += __expf((float)z[j] - mx);
  sum = warp_sum(sum);
  if (lane == 0) sred[warp] = sum;
  __syncthreads();
  if (warp == 0) {
    float v = lane < nw ? sred[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) sval = v;
  }
  __syncthreads();
  sum = sval;
  const float inv = 1.f / sum;
  const int lab = (int)labels[row];
  if (dlogits) {
    bf16* d = dlogits + (size_t)row * ldv;
    for (int j = tid; j < V; j += blockDim.x) {
      float pj = __expf((float)z[j] - mx) * inv;
      d[j] = __float2bfloat16((pj - (j == lab ? 1.f : 0.f)) * gscale);
    }
    for (int j = V + tid; j < ldv; j += blockDim.x) d[j] = __float2bfloat16(0.f);
  }
  if (tid == 0) {
    float zl = (lab >= 0 && lab < V) ? (float)z[lab] : mx;
    row_loss[row] = logf(sum) + mx - zl;
    __threadfence();
    unsigned prev = atomicAdd(counter, 1u);
    is_last = (prev == (unsigned)B - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  float acc = 0.f;
  for (int j = tid; j < B; j += blockDim.x) acc += __ldcg(row_loss + j);
  acc = warp_sum(acc);
  __syncthreads();
  if (lane == 0) sred[warp] = acc;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += sred[w];
    *loss_out = t / (float)B;
    *counter = 0u;
  }
}

// Momentum SGD over a flat parameter arena: fp32 master + fp32 momentum, bf16 gradient in,
// bf16 compute copy out.  Elements [0, wd_count) get weight decay (conv/fc weights), the rest
// (BN scale/shift, biases) do not.  grad_scale folds in any deferred averaging.
__global__ void __launch_bounds__(256) sgd_momentum_kernel(float* __restrict__ p32,
                                                           float* __restrict__ mom,
                                                           const bf16* __restrict__ g16,
                                                           bf16* __restrict__ p16, int64_t n,
                                                           int64_t wd_count, float lr, float mu,
                                                           float wd, float grad_scale) {
  const int64_t nvec = n / 8;
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += T) {
    uint4 vg = ld_stream(g16 + i * 8);
    float4 pa = reinterpret_cast<const float4*>(p32)[2 * i];
    float4 pb = reinterpret_cast<const float4*>(p32)[2 * i + 1];
    float4 ma = reinterpret_cast<const float4*>(mom)[2 * i];
    float4 mb = reinterpret_cast<const float4*>(mom)[2 * i + 1];
    float g[8], p[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w},
                m[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
    unpack8(vg, g);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float d = (i * 8 + j) < wd_count ? wd : 0.f;
      float gj = fmaf(d, p[j], g[j] * grad_scale);
      m[j] = fmaf(mu, m[j], gj);
      p[j] = fmaf(-lr, m[j], p[j]);
    }
    reinterpret_cast<float4*>(p32)[2 * i] = make_float4(p[0], p[1], p[2], p[3]);
    reinterpret_cast<float4*>(p32)[2 * i + 1] = make_float4(p[4], p[5], p[6], p[7]);
    reinterpret_cast<float4*>(mom)[2 * i] = make_float4(m[0], m[1], m[2], m[3]);
    reinterpret_cast<float4*>(mom)[2 * i + 1] = make_float4(m[4], m[5], m[6], m[7]);
    *reinterpret_cast<uint4*>(p16 + i * 8) = pack8(p);
  }
  // tail (n not a multiple of 8)
  if (blockIdx.x == 0) {
    for (int64_t k = nvec * 8 + threadIdx.x; k < n; k += blockDim.x) {
      float d = k < wd_count ? wd : 0.f;
      float gj = fmaf(d, p32[k], __bfloat162float(g16[k]) * grad_scale);
      float mk = fmaf(mu, mom[k], gj);
      mom[k] = mk;
      float pk = fmaf(-lr, mk, p32[k]);
      p32[k] = pk;
      p16[k] = __float2bfloat16(pk);
    }
  }
}

// out = a + b  (bf16, 128-bit); out may alias a or b.
__global__ void __launch_bounds__(256) add_bf16_kernel(const bf16* a, const bf16* b, bf16* out,
                                                       int64_t nvec) {
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += T) {
    uint4 va = ld_plain(a + i * 8), vb = ld_plain(b + i * 8);
    float fa[8], fb[8];
    unpack8(va, fa);
    unpack8(vb, fb);
#pragma unroll
    for (int j = 0; j < 8; ++j) fa[j] += fb[j];
    *reinterpret_cast<uint4*>(out + i * 8) = pack8(fa);
  }
}

// 3x3 stride-2 pad-1 max-pool on NHWC bf16; saves the argmax (0..8) per output vector lane so
// the backward pass is a gather (each input pixel looks at the <=4 windows that cover it).
// AFFINE: the input is a raw convolution output and y = maxpool(relu(a*x + b)) with per-channel (a, b) —
// BatchNorm apply + ReLU + max-pool of the ResNet stem in one pass; the normalised activation (411 MB at
// batch 256) is never written or re-read.
template <bool AFFINE>
__global__ void __launch_bounds__(256) maxpool3x3s2_fwd_kernel(const bf16* __restrict__ x,
                                                               bf16* __restrict__ y,
                                                               uint8_t* __restrict__ idx, int N,
                                                               int H, int W, int C, int OH, int OW,
                                                               const float* __restrict__ coef) {
  const int cvec = C / 8;
  const int64_t total = (int64_t)N * OH * OW * cvec;
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += T) {
    // 32-bit index decode (the host checks total < 2^31): 64-bit div / mod are ~100-instruction software
    // routines each, six of them cost as much as the nine loads
    const uint32_t iu = (uint32_t)i;
    uint32_t t = iu / (uint32_t)cvec;
    const int cg = (int)(iu - t * (uint32_t)cvec);
    const uint32_t t2 = t / (uint32_t)OW;
    const int ow = (int)(t - t2 * (uint32_t)OW);
    const int n = (int)(t2 / (uint32_t)OH);
    const int oh = (int)(t2 - (uint32_t)n * (uint32_t)OH);
    // Two bf16 values per 32-bit word: the running maximum (HMNMX2), the "new value is larger" mask and the
    // argmax codes (16 bits per value) are all updated with packed instructions — about a third of the
    // instructions of the scalar fp32 version, which was issue-bound at 2.6 TB/s.
    uint32_t best[4], bi[4], sgn[4];
    float ca[8], cb[8];
    if (AFFINE) {
      // relu(a*x + b) is monotonic in x: compare sign(a)*x (a sign-bit flip) now, apply the affine once at the end
#pragma unroll
      for (int j = 0; j < 8; ++j) { ca[j] = coef[cg * 8 + j]; cb[j] = coef[C + cg * 8 + j]; }
#pragma unroll
      for (int k = 0; k < 4; ++k) sgn[k] = (ca[2 * k] < 0.f ? 0x00008000u : 0u) | (ca[2 * k + 1] < 0.f ? 0x80000000u : 0u);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) { best[k] = 0xFF80FF80u; bi[k] = 0x00040004u; }     // -inf, centre tap
    // All nine taps are loaded before the first compare: with a bounds test (and a `continue`) in front of each
    // load the compiler kept load -> compare -> next load in program order, nine dependent memory latencies
    // per thread (ncu: 56 % issue-active, the rest long-scoreboard stalls, 51 % of DRAM throughput).  A tap
    // outside the image reads the window's centre pixel instead (always inside) and is replaced by -inf.
    uint4 tap[9];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = oh * 2 - 1 + kh;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = ow * 2 - 1 + kw;
        const bool ok = (unsigned)ih < (unsigned)H && (unsigned)iw < (unsigned)W;
        const int ihc = ok ? ih : oh * 2, iwc = ok ? iw : ow * 2;
        uint4 v = ld_plain(x + (((int64_t)n * H + ihc) * W + iwc) * C + cg * 8);
        if (!ok) {                       // -inf AFTER the sign flip applied below
          constexpr uint32_t kNegInf2 = 0xFF80FF80u;
          v = AFFINE ? make_uint4(kNegInf2 ^ sgn[0], kNegInf2 ^ sgn[1], kNegInf2 ^ sgn[2], kNegInf2 ^ sgn[3])
                     : make_uint4(kNegInf2, kNegInf2, kNegInf2, kNegInf2);
        }
        tap[kh * 3 + kw] = v;
      }
    }
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const uint4 v = tap[kh * 3 + kw];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        const uint32_t code = (uint32_t)(kh * 3 + kw) * 0x00010001u;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t xw = AFFINE ? (w[k] ^ sgn[k]) : w[k];
          const __nv_bfloat162 nv = *reinterpret_cast<const __nv_bfloat162*>(&xw);
          const __nv_bfloat162 bv = *reinterpret_cast<const __nv_bfloat162*>(&best[k]);
          const uint32_t m = __hgt2_mask(nv, bv);              // 0xFFFF per half where the new value is larger
          const __nv_bfloat162 mx = __hmax2(nv, bv);
          best[k] = *reinterpret_cast<const uint32_t*>(&mx);
          bi[k] = (bi[k] & ~m) | (code & m);
        }
      }
    }
    uint4 o;
    if (AFFINE) {
      float f[8];
      const uint4 raw = make_uint4(best[0] ^ sgn[0], best[1] ^ sgn[1], best[2] ^ sgn[2], best[3] ^ sgn[3]);
      unpack8(raw, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = fmaxf(fmaf(ca[j], f[j], cb[j]), 0.f);
      o = pack8(f);
    } else {
      o = make_uint4(best[0], best[1], best[2], best[3]);
    }
    *reinterpret_cast<uint4*>(y + i * 8) = o;
    uint2 pk;
    pk.x = __byte_perm(bi[0], bi[1], 0x6420);
    pk.y = __byte_perm(bi[2], bi[3], 0x6420);
    *reinterpret_cast<uint2*>(idx + i * 8) = pk;
  }
}

// Backward for even H, W: one thread owns a 2x2 block of input pixels (x 8 channels).  The block is touched by
// exactly four windows, (a,b) (a,b+1) (a+1,b) (a+1,b+1); each window's argmax codes and gradient vector are
// loaded ONCE and matched against the nine (pixel, tap) pairs with byte-wise SIMD compares — no loops, no
// divergent candidates, 1 window load per output vector instead of 2.25.  Accumulation order per pixel is
// the same as in the generic kernel below, so results are bit-identical.
__global__ void __launch_bounds__(256) maxpool3x3s2_bwd_quad_kernel(const bf16* __restrict__ dy,
                                                                    const uint8_t* __restrict__ idx,
                                                                    bf16* __restrict__ dx, int N, int H, int W,
                                                                    int C, int OH, int OW) {
  const int cvec = C / 8;
  const int HB = H / 2, WB = W / 2;
  const int64_t total = (int64_t)N * HB * WB * cvec;
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += T) {
    const uint32_t iu = (uint32_t)i;            // 32-bit decode, as in the forward kernel
    const uint32_t t = iu / (uint32_t)cvec;
    const int cg = (int)(iu - t * (uint32_t)cvec);
    const uint32_t t2 = t / (uint32_t)WB;
    const int b = (int)(t - t2 * (uint32_t)WB);
    const int n = (int)(t2 / (uint32_t)HB);
    const int a = (int)(t2 - (uint32_t)n * (uint32_t)HB);
    float acc[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[q][j] = 0.f;
    // (window dh, dw) -> list of (pixel index 2*pr+pc, tap code); pixel (pr, pc) = (2a+pr, 2b+pc)
    // the four windows' codes and gradients are loaded before any of them is used (the bounds test in front of each
    // load otherwise serialises them); a window outside the output reads window (a, b) and matches no tap (code 0xFF)
    uint2 pks[4];
    uint4 dys[4];
#pragma unroll
    for (int wdx = 0; wdx < 4; ++wdx) {
      const int dh = wdx >> 1, dw = wdx & 1;
      const bool ok = a + dh < OH && b + dw < OW;
      const int oh = ok ? a + dh : a, ow = ok ? b + dw : b;
      const int64_t o = ((((int64_t)n * OH + oh) * OW + ow) * cvec + cg) * 8;
      uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
      if (!ok) pk = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
      pks[wdx] = pk;
      dys[wdx] = ld_plain(dy + o);
    }
#pragma unroll
    for (int wdx = 0; wdx < 4; ++wdx) {
      const int dh = wdx >> 1, dw = wdx & 1;
      const uint2 pk = pks[wdx];
      float f[8];
      unpack8(dys[wdx], f);
#pragma unroll
      for (int pr = dh; pr < 2; ++pr)        // a window one block down only reaches the block's second row
#pragma unroll
        for (int pc = dw; pc < 2; ++pc) {
          const int kh = dh ? 0 : pr + 1, kw = dw ? 0 : pc + 1;
          const uint32_t c4 = (uint32_t)(kh * 3 + kw) * 0x01010101u;
          const uint32_t m0 = __vcmpeq4(pk.x, c4), m1 = __vcmpeq4(pk.y, c4);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t hit = (j < 4 ? (m0 >> (8 * j)) : (m1 >> (8 * (j - 4)))) & 1u;
            acc[pr * 2 + pc][j] += hit ? f[j] : 0.f;
          }
        }
    }
#pragma unroll
    for (int pr = 0; pr < 2; ++pr)
#pragma unroll
      for (int pc = 0; pc < 2; ++pc) {
        const int64_t o = ((((int64_t)n * H + 2 * a + pr) * W + 2 * b + pc) * cvec + cg) * 8;
        *reinterpret_cast<uint4*>(dx + o) = pack8(acc[pr * 2 + pc]);
      }
  }
}

__global__ void __launch_bounds__(256) maxpool3x3s2_bwd_kernel(const bf16* __restrict__ dy,
                                                               const uint8_t* __restrict__ idx,
                                                               bf16* __restrict__ dx, int N, int H,
                                                               int W, int C, int OH, int OW) {
  const int cvec = C / 8;
  const int64_t total = (int64_t)N * H * W * cvec;
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += T) {
    int cg = (int)(i % cvec);
    int64_t t = i / cvec;
    int iw = (int)(t % W); t /= W;
    int ih = (int)(t % H);
    int n = (int)(t / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    // windows (oh, ow) with ih = 2*oh-1+kh, kh in 0..2  ->  oh in {(ih+1)/2, (ih+1)/2 - 1(if valid)}
    for (int oh = (ih + 1) / 2 - 1; oh <= (ih + 1) / 2; ++oh) {
      if (oh < 0 || oh >= OH) continue;
      int kh = ih - (oh * 2 - 1);
      if (kh < 0 || kh > 2) continue;
      for (int ow = (iw + 1) / 2 - 1; ow <= (iw + 1) / 2; ++ow) {
        if (ow < 0 || ow >= OW) continue;
        int kw = iw - (ow * 2 - 1);
        if (kw < 0 || kw > 2) continue;
        int64_t o = (((int64_t)n * OH + oh) * OW + ow) * cvec + cg;
        uint2 pk = *reinterpret_cast<const uint2*>(idx + o * 8);
        // byte-wise SIMD compare of the eight argmax codes; most windows select another position for
        // all eight channels, then the gradient vector is not even loaded
        const uint32_t c4 = (uint32_t)(kh * 3 + kw) * 0x01010101u;
        const uint32_t m0 = __vcmpeq4(pk.x, c4), m1 = __vcmpeq4(pk.y, c4);
        if ((m0 | m1) == 0u) continue;
        uint4 v = ld_plain(dy + o * 8);
        float f[8];
        unpack8(v, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t hit = (j < 4 ? (m0 >> (8 * j)) : (m1 >> (8 * (j - 4)))) & 1u;
          if (hit) acc[j] += f[j];
        }
      }
    }
    *reinterpret_cast<uint4*>(dx + i * 8) = pack8(acc);
  }
}

// Global average pool NHWC [N, HW, C] -> [N, C] and its backward (broadcast / HW).
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const bf16* __restrict__ x,
                                                          bf16* __restrict__ y, int N, int HW,
                                                          int C) {
  const int cvec = C / 8;
  const int64_t total = (int64_t)N * cvec;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int cg = (int)(i % cvec);
  int n = (int)(i / cvec);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int p = 0; p < HW; ++p) {
    uint4 v = ld_stream(x + ((int64_t)n * HW + p) * C + cg * 8);
    float f[8];
    unpack8(v, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += f[j];
  }
  const float inv = 1.f / (float)HW;
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] *= inv;
  *reinterpret_cast<uint4*>(y + i * 8) = pack8(acc);
}

__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const bf16* __restrict__ dy,
                                                          bf16* __restrict__ dx, int N, int HW,
                                                          int C) {
  const int cvec = C / 8;
  const int64_t total = (int64_t)N * HW * cvec;
  const int64_t T = (int64_t)gridDim.x * blockDim.x;
  const float inv = 1.f / (float)HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += T) {
    int cg = (int)(i % cvec);
    int n = (int)(i / ((int64_t)HW * cvec));
    uint4 v = ld_plain(dy + ((int64_t)n * cvec + cg) * 8);
    float f[8];
    unpack8(v, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= inv;
    *reinterpret_cast<uint4*>(dx + i * 8) = pack8(f);
  }
}

// Stem im2col fused with the input normalisation: uint8 NHWC image -> bf16 patch matrix
// [N*OH*OW, KP] for the 7x7 stride-2 pad-3 convolution, column k = (kh*7 + kw)*3 + c (the order of
// the NHWC weight [Cout, 7, 7, 3]), columns 147..KP-1 zero.  The stem then runs as a plain GEMM on
// the tcgen05 kernel (K = 147 is far too thin for an implicit-GEMM pipeline to pay off, and the
// patch matrix is written once and reused by wgrad).  One thread = one 16-byte output vector.
// Layout of a patch row: 7 filter rows x 24 columns; columns 0..20 of filter row kh are the 7 pixels x 3
// channels the filter row covers (one contiguous 21-byte run of the image), 21..23 are zero.  With the
// rows padded to 24 every 16-byte output vector lies inside one filter row, i.e. is 8 consecutive
// normalised values of one image row: a block stages the 7 image rows an output row (n, oh) needs in
// shared memory once (coalesced 4-byte loads, normalisation applied there) and then emits the
// 112 x 21 vectors with four aligned 32-bit shared loads each.  (The first version gathered byte by
// byte from global memory: 1.0 ms per step against 0.16 ms of pure traffic.)
constexpr int kStemKW = 24;             // padded filter-row width
constexpr int kStemKP = 7 * kStemKW;    // 168 patch columns
constexpr int kStemMaxW = 256;          // widest input row the shared-memory stage is sized for

__global__ void __launch_bounds__(256) stem_im2col_kernel(const uint8_t* __restrict__ img,
                                                          bf16* __restrict__ out, int N, int H, int W,
                                                          int OH, int OW, float m0, float m1,
                                                          float m2, float s0, float s1, float s2) {
  // srow[kh][c]: c = (iw + 3) * 3 + channel for iw in [-3, W + 3)  (3 zero pixels either side; +8 slack
  // because the last vector of a row reads three elements past column 20)
  __shared__ __align__(16) bf16 srow[7][(kStemMaxW + 6) * 3 + 8];
  const int rowlen = (W + 6) * 3 + 8;
  const int n = blockIdx.x / OH, oh = blockIdx.x - n * OH;
  const float k255 = 1.f / 255.f;
  const float mean[3] = {m0, m1, m2}, istd[3] = {s0, s1, s2};
  const bf16 zero = __float2bfloat16(0.f);
  for (int i = threadIdx.x; i < 7 * rowlen; i += 256) (&srow[0][0])[(i / rowlen) * ((kStemMaxW + 6) * 3 + 8) + i % rowlen] = zero;
  __syncthreads();
  const int words = W * 3 / 4;     // W * 3 is a multiple of 4 for W % 4 == 0
  for (int i = threadIdx.x; i < 7 * words; i += 256) {
    const int kh = i / words, w4 = i - kh * words;
    const int ih = oh * 2 - 3 + kh;
    if (ih < 0 || ih >= H) continue;
    const uchar4 v = *reinterpret_cast<const uchar4*>(img + ((size_t)n * H + ih) * W * 3 + 4 * w4);
    const unsigned char b[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e = 4 * w4 + j, c = e % 3;
      srow[kh][9 + e] = __float2bfloat16(((float)b[j] * k255 - mean[c]) * istd[c]);
    }
  }
  __syncthreads();
  const int vpr = kStemKP / 8;     // 21 vectors per output pixel
  bf16* orow = out + (size_t)blockIdx.x * OW * kStemKP;
  for (int q = threadIdx.x; q < OW * vpr; q += 256) {
    const int ow = q / vpr, v = q - ow * vpr;
    const int kh = v / 3, j0 = (v - kh * 3) * 8;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&srow[kh][ow * 6 + j0]);   // even element: 4-byte aligned
    uint4 o = make_uint4(src[0], src[1], src[2], src[3]);
    if (j0 == 16) {      // columns 21..23 of a filter row are padding
      o.z &= 0x0000ffffu;
      o.w = 0u;
    }
    *reinterpret_cast<uint4*>(orow + (size_t)q * 8) = o;
  }
}

static int grid_for(int64_t n, int per_block) {
  int64_t g = (n + per_block - 1) / per_block;
  if (g > 148 * 16) g = 148 * 16;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace kdl

using namespace kdl;

extern "C" int kdl_u8_normalize(const void* in_u8, void* out_bf16, int64_t npix, int cpad,
                                const float* mean3, const float* std3, cudaStream_t stream) {
  if (cpad < 3 || cpad > 8) {
    kdl_set_error("u8_normalize: cpad must be in [3,8]");
    return KDL_ERR_ARG;
  }
  u8_normalize_kernel<<<grid_for(npix, 256 * 4), 256, 0, stream>>>(
      (const uint8_t*)in_u8, (bf16*)out_bf16, npix, cpad, mean3[0], mean3[1], mean3[2],
      1.f / std3[0], 1.f / std3[1], 1.f / std3[2]);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" size_t kdl_softmax_xent_workspace_bytes(int B) { return 256 + (size_t)B * sizeof(float); }

extern "C" int kdl_softmax_xent(const void* logits, int logits_is_fp32, const int64_t* labels,
                                void* dlogits_bf16, float* loss_out, int B, int V, int ldv,
                                float grad_scale, void* workspace, size_t workspace_bytes,
                                cudaStream_t stream) {
  if (workspace_bytes < kdl_softmax_xent_workspace_bytes(B)) {
    kdl_set_error("softmax_xent: workspace too small");
    return KDL_ERR_WORKSPACE;
  }
  unsigned* counter = (unsigned*)workspace;
  float* row_loss = (float*)((char*)workspace + 256);
  if (logits_is_fp32)
    softmax_xent_kernel<float><<<B, 256, 0, stream>>>((const float*)logits, labels,
                                                      (bf16*)dlogits_bf16, row_loss, loss_out,
                                                      counter, B, V, ldv, grad_scale);
  else
    softmax_xent_kernel<bf16><<<B, 256, 0, stream>>>((const bf16*)logits, labels,
                                                     (bf16*)dlogits_bf16, row_loss, loss_out,
                                                     counter, B, V, ldv, grad_scale);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_sgd_momentum(float* p32, float* mom, const void* g16, void* p16, int64_t n,
                                int64_t wd_count, float lr, float mu, float wd, float grad_scale,
                                cudaStream_t stream) {
  sgd_momentum_kernel<<<grid_for(n / 8 + 1, 256), 256, 0, stream>>>(
      p32, mom, (const bf16*)g16, (bf16*)p16, n, wd_count, lr, mu, wd, grad_scale);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_add_bf16(const void* a, const void* b, void* out, int64_t n,
                            cudaStream_t stream) {
  if (n % 8) {
    kdl_set_error("add_bf16: n must be a multiple of 8");
    return KDL_ERR_ARG;
  }
  add_bf16_kernel<<<grid_for(n / 8, 256), 256, 0, stream>>>((const bf16*)a, (const bf16*)b,
                                                            (bf16*)out, n / 8);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_maxpool3x3s2_fwd(const void* x, void* y, void* idx_u8, int N, int H, int W,
                                    int C, cudaStream_t stream) {
  if (C % 8) {
    kdl_set_error("maxpool: C must be a multiple of 8");
    return KDL_ERR_ARG;
  }
  int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
  int64_t total = (int64_t)N * OH * OW * (C / 8);
  if (total >= ((int64_t)1 << 31)) {
    kdl_set_error("maxpool: more than 2^31 output vectors (N=%d H=%d W=%d C=%d)", N, H, W, C);
    return KDL_ERR_UNSUPPORTED;
  }
  maxpool3x3s2_fwd_kernel<false><<<grid_for(total, 256), 256, 0, stream>>>(
      (const bf16*)x, (bf16*)y, (uint8_t*)idx_u8, N, H, W, C, OH, OW, nullptr);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

// y = maxpool3x3s2(relu(a*x + b)), coef = [2][C] (a, b) as kdl_bn_finalize_from_stats writes them
extern "C" int kdl_maxpool3x3s2_affine_relu_fwd(const void* x, const float* coef, void* y, void* idx_u8, int N,
                                                int H, int W, int C, cudaStream_t stream) {
  if (C % 8 || coef == nullptr) {
    kdl_set_error("maxpool_affine: C must be a multiple of 8 and coef non-null");
    return KDL_ERR_ARG;
  }
  int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
  int64_t total = (int64_t)N * OH * OW * (C / 8);
  if (total >= ((int64_t)1 << 31)) {
    kdl_set_error("maxpool: more than 2^31 output vectors (N=%d H=%d W=%d C=%d)", N, H, W, C);
    return KDL_ERR_UNSUPPORTED;
  }
  maxpool3x3s2_fwd_kernel<true><<<grid_for(total, 256), 256, 0, stream>>>(
      (const bf16*)x, (bf16*)y, (uint8_t*)idx_u8, N, H, W, C, OH, OW, coef);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_maxpool3x3s2_bwd(const void* dy, const void* idx_u8, void* dx, int N, int H,
                                    int W, int C, cudaStream_t stream) {
  if (C % 8) {
    kdl_set_error("maxpool: C must be a multiple of 8");
    return KDL_ERR_ARG;
  }
  int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
  int64_t total = (int64_t)N * H * W * (C / 8);
  if ((H % 2) == 0 && (W % 2) == 0 && total / 4 < ((int64_t)1 << 31)) {
    maxpool3x3s2_bwd_quad_kernel<<<grid_for(total / 4, 256), 256, 0, stream>>>(
        (const bf16*)dy, (const uint8_t*)idx_u8, (bf16*)dx, N, H, W, C, OH, OW);
    KDL_CHECK_LAUNCH();
    return KDL_OK;
  }
  maxpool3x3s2_bwd_kernel<<<grid_for(total, 256), 256, 0, stream>>>(
      (const bf16*)dy, (const uint8_t*)idx_u8, (bf16*)dx, N, H, W, C, OH, OW);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_avgpool_fwd(const void* x, void* y, int N, int HW, int C, cudaStream_t stream) {
  if (C % 8) {
    kdl_set_error("avgpool: C must be a multiple of 8");
    return KDL_ERR_ARG;
  }
  int64_t total = (int64_t)N * (C / 8);
  avgpool_fwd_kernel<<<(int)((total + 255) / 256), 256, 0, stream>>>((const bf16*)x, (bf16*)y, N,
                                                                     HW, C);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C,
                               cudaStream_t stream) {
  if (C % 8) {
    kdl_set_error("avgpool: C must be a multiple of 8");
    return KDL_ERR_ARG;
  }
  int64_t total = (int64_t)N * HW * (C / 8);
  avgpool_bwd_kernel<<<grid_for(total, 256), 256, 0, stream>>>((const bf16*)dy, (bf16*)dx, N, HW,
                                                               C);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}

extern "C" int kdl_stem_im2col_u8(const void* img_u8, void* patches_bf16, int N, int H, int W, int KP,
                                  const float* mean3, const float* std3, cudaStream_t stream) {
  if (KP != kStemKP) {
    kdl_set_error("stem_im2col: KP must be %d (7 filter rows x 24 padded columns), got %d", kStemKP, KP);
    return KDL_ERR_ARG;
  }
  if (W > kStemMaxW || (W % 4) != 0) {
    kdl_set_error("stem_im2col: need W %% 4 == 0 and W <= %d (got %d)", kStemMaxW, W);
    return KDL_ERR_UNSUPPORTED;
  }
  const int OH = (H + 6 - 7) / 2 + 1, OW = (W + 6 - 7) / 2 + 1;
  stem_im2col_kernel<<<N * OH, 256, 0, stream>>>(
      (const uint8_t*)img_u8, (bf16*)patches_bf16, N, H, W, OH, OW, mean3[0], mean3[1], mean3[2],
      1.f / std3[0], 1.f / std3[1], 1.f / std3[2]);
  KDL_CHECK_LAUNCH();
  return KDL_OK;
}
