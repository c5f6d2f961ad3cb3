// Skinny layers, losses and the optimizer:
//   * g_h0_lin  : z[B,100] @ Matrix[100,F] + bias            (distriubted_model.py:88, 160-173)
//   * d_h3_lin  : h3[B,F] @ Matrix[F,1] + bias -> logits      (distriubted_model.py:122)
//   * the three sigmoid-cross-entropy means and their logit gradients (image_train.py:91-96)
//   * TensorFlow-form Adam over the flat parameter arena     (image_train.py:109-112)
//   * layout helpers (per-tap weight transposes, column sums, f64 decode, sample grid)
#include "common.cuh"
#include "../include/dcgan_b200.h"

namespace dcgan {

// ---------------------------------------------------------------------------------------------
// g_h0_lin forward: out[n, j] = sum_k z[n,k] * W[k,j] + b[j]  -> bf16.  W is the bf16 copy.
// block = 128 threads x 2 adjacent j (one 4-byte weight load per k) x kLinTile rows of z; z is staged
// in smem as [k][kLinTile] so that a k step reads its 16 values with four 16-byte broadcasts.
// ---------------------------------------------------------------------------------------------
constexpr int kLinTile = 16;
__global__ void __launch_bounds__(128)
lin_fwd_kernel(const float* __restrict__ z, const __nv_bfloat16* __restrict__ W, const float* __restrict__ bias,
               __nv_bfloat16* __restrict__ out, int NB, int K, int F) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ __align__(16) float zs[];  // [K][kLinTile]
  const int n0 = blockIdx.y * kLinTile;
  for (int i = threadIdx.x; i < kLinTile * K; i += 128) {
    const int r = i / K, k = i - r * K;  // consecutive threads read consecutive k of one row
    zs[k * kLinTile + r] = (n0 + r < NB) ? z[(size_t)(n0 + r) * K + k] : 0.f;
  }
  __syncthreads();
  const int j = (blockIdx.x * 128 + threadIdx.x) * 2;
  if (j >= F) return;
  float acc[kLinTile][2];
  const float b0 = bias ? bias[j] : 0.f, b1 = bias ? bias[j + 1] : 0.f;
#pragma unroll
  for (int i = 0; i < kLinTile; ++i) { acc[i][0] = b0; acc[i][1] = b1; }
  constexpr int kBatch = 25;  // weight loads in flight per thread (the kernel is latency-bound: few, long threads)
  for (int k0 = 0; k0 < K; k0 += kBatch) {
    uint32_t wv[kBatch];
#pragma unroll
    for (int u = 0; u < kBatch; ++u)
      wv[u] = (k0 + u < K) ? *reinterpret_cast<const uint32_t*>(W + (size_t)(k0 + u) * F + j) : 0u;
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      if (k0 + u < K) {
        const float w0 = __uint_as_float(wv[u] << 16), w1 = __uint_as_float(wv[u] & 0xffff0000u);
        const float4* zr = reinterpret_cast<const float4*>(zs + (k0 + u) * kLinTile);
#pragma unroll
        for (int q = 0; q < kLinTile / 4; ++q) {
          const float4 zz = zr[q];
          acc[4 * q][0] += zz.x * w0; acc[4 * q][1] += zz.x * w1;
          acc[4 * q + 1][0] += zz.y * w0; acc[4 * q + 1][1] += zz.y * w1;
          acc[4 * q + 2][0] += zz.z * w0; acc[4 * q + 2][1] += zz.z * w1;
          acc[4 * q + 3][0] += zz.w * w0; acc[4 * q + 3][1] += zz.w * w1;
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kLinTile; ++i)
    if (n0 + i < NB)
      *reinterpret_cast<__nv_bfloat162*>(out + (size_t)(n0 + i) * F + j) = __floats2bfloat162_rn(acc[i][0], acc[i][1]);
}

// g_h0_lin weight gradient: dW[k, j] = sum_n z[n,k] * dy[n,j]  (f32, overwrites).
__global__ void __launch_bounds__(128)
lin_wgrad_kernel(const float* __restrict__ z, const __nv_bfloat16* __restrict__ dy, float* __restrict__ dW,
                 float* __restrict__ dbias, int NB, int K, int F) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float zs[];  // [NB][kt]
  const int kt = 20;             // rows of dW per block (K = 100 -> 5 blocks in y)
  const int k0 = blockIdx.y * kt;
  for (int i = threadIdx.x; i < NB * kt; i += 128) {
    const int n = i / kt, k = k0 + i % kt;
    zs[i] = k < K ? z[(size_t)n * K + k] : 0.f;
  }
  __syncthreads();
  const int j = blockIdx.x * 128 + threadIdx.x;
  if (j >= F) return;
  float acc[20];
#pragma unroll
  for (int i = 0; i < kt; ++i) acc[i] = 0.f;
  float bsum = 0.f;
  constexpr int kBatch = 32;  // gradient loads in flight per thread
  for (int nb = 0; nb < NB; nb += kBatch) {
    __nv_bfloat16 dv[kBatch];
#pragma unroll
    for (int u = 0; u < kBatch; ++u) dv[u] = (nb + u < NB) ? dy[(size_t)(nb + u) * F + j] : __float2bfloat16(0.f);
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      if (nb + u < NB) {
        const float d = __bfloat162float(dv[u]);
        bsum += d;
#pragma unroll
        for (int i = 0; i < kt; ++i) acc[i] += zs[(nb + u) * kt + i] * d;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kt; ++i)
    if (k0 + i < K) dW[(size_t)(k0 + i) * F + j] = acc[i];
  if (dbias && blockIdx.y == 0) dbias[j] = bsum;
}

// ---------------------------------------------------------------------------------------------
// d_h3_lin forward: logits[n] = h3[n,:] . w + b ; one block per image
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
head_fwd_kernel(const __nv_bfloat16* __restrict__ h, const float* __restrict__ w, const float* __restrict__ b,
                float* __restrict__ logits, int F) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float part[8];
  const int n = blockIdx.x;
  float acc = 0.f;
  const uint4* hp = reinterpret_cast<const uint4*>(h + (size_t)n * F);
  for (int i = threadIdx.x; i < F / 8; i += 256) {
    const uint4 v = __ldg(hp + i);
    const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(u[j]);
      acc += f.x * w[i * 8 + 2 * j] + f.y * w[i * 8 + 2 * j + 1];
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = b[0];
    for (int i = 0; i < 8; ++i) s += part[i];
    logits[n] = s;
  }
}

// ---------------------------------------------------------------------------------------------
// losses (single block).  logits = [real(B); fake(B)].
//   losses[0] = d_loss_real = mean BCE(real, 1)   losses[1] = d_loss_fake = mean BCE(fake, 0)
//   losses[2] = g_loss      = mean BCE(fake, 1)
//   dlogit slabs: [0,B) d(d_loss)/d real, [B,2B) d(d_loss)/d fake, [2B,3B) d(g_loss)/d fake,
//   each already divided by B * grad_div (grad_div = world size when gradients are summed across ranks).
// Also advances the device-side step counters: counters[0] = Adam t (both optimizers step once per
// iteration), counters[1] = global_step (incremented by g_optim only, image_train.py:112).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float bce_logits(float x, float t) {
  return fmaxf(x, 0.f) - x * t + log1pf(__expf(-fabsf(x)));
}
__global__ void __launch_bounds__(256)
gan_loss_kernel(const float* __restrict__ logits, float* __restrict__ losses, float* __restrict__ dlogit,
                float* __restrict__ probs, long long* __restrict__ counters, int B, float grad_div,
                int advance) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[3][8];
  float l0 = 0.f, l1 = 0.f, l2 = 0.f;
  const float scale = 1.f / ((float)B * grad_div);
  for (int i = threadIdx.x; i < B; i += 256) {
    const float xr = logits[i], xf = logits[B + i];
    l0 += bce_logits(xr, 1.f);
    l1 += bce_logits(xf, 0.f);
    l2 += bce_logits(xf, 1.f);
    const float sr = 1.f / (1.f + __expf(-xr)), sf = 1.f / (1.f + __expf(-xf));
    if (dlogit) {
      dlogit[i] = (sr - 1.f) * scale;
      dlogit[B + i] = sf * scale;
      dlogit[2 * B + i] = (sf - 1.f) * scale;
    }
    if (probs) { probs[i] = sr; probs[B + i] = sf; }
  }
  l0 = warp_sum(l0); l1 = warp_sum(l1); l2 = warp_sum(l2);
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = l0; red[1][threadIdx.x >> 5] = l1; red[2][threadIdx.x >> 5] = l2;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += red[threadIdx.x][i];
    losses[threadIdx.x] = s / (float)B;
  }
  if (threadIdx.x == 0 && advance && counters) { counters[0] += 1; counters[1] += 1; }
}

// ---------------------------------------------------------------------------------------------
// d_h3_lin backward: dh[n,k] = dlogit[n]*w[k] for all 3B slabs (bf16);
// dW[k] += sum_{n<2B} dlogit[n]*h[n,k]; db += sum_{n<2B} dlogit[n]  (atomics; zeroed by caller).
// thread = 8 features x kHeadBwdChunk images.
// ---------------------------------------------------------------------------------------------
// kFusedLoss: the logit gradients are recomputed from the logits by every thread that needs them and block
// (0,0) also produces what gan_loss_kernel would (the three means, dlogit, probs, counters), so the step has
// no single-block loss kernel between the two head kernels.
constexpr int kHeadBwdChunk = 8;
struct LossArgs {
  const float* logits;
  float* losses;
  float* dlogit_out;
  float* probs;
  long long* counters;
  float grad_div;
  int advance;
};
template <bool kFusedLoss>
__global__ void __launch_bounds__(128)
head_bwd_kernel(const __nv_bfloat16* __restrict__ h, const float* __restrict__ w,
                const float* __restrict__ dlogit, __nv_bfloat16* __restrict__ dh, float* __restrict__ dW,
                float* __restrict__ db, int B, int F, LossArgs la) {
  pdl_launch_dependents();
  pdl_wait();
  const float lscale = kFusedLoss ? 1.f / ((float)B * la.grad_div) : 0.f;
  if (kFusedLoss && blockIdx.x == 0 && blockIdx.y == 0) {
    __shared__ float red[3][4];
    float l0 = 0.f, l1 = 0.f, l2 = 0.f;
    for (int i = threadIdx.x; i < B; i += 128) {
      const float xr = la.logits[i], xf = la.logits[B + i];
      l0 += bce_logits(xr, 1.f);
      l1 += bce_logits(xf, 0.f);
      l2 += bce_logits(xf, 1.f);
      const float sr = 1.f / (1.f + __expf(-xr)), sf = 1.f / (1.f + __expf(-xf));
      if (la.dlogit_out) {
        la.dlogit_out[i] = (sr - 1.f) * lscale;
        la.dlogit_out[B + i] = sf * lscale;
        la.dlogit_out[2 * B + i] = (sf - 1.f) * lscale;
      }
      if (la.probs) { la.probs[i] = sr; la.probs[B + i] = sf; }
    }
    l0 = warp_sum(l0); l1 = warp_sum(l1); l2 = warp_sum(l2);
    if ((threadIdx.x & 31) == 0) {
      red[0][threadIdx.x >> 5] = l0; red[1][threadIdx.x >> 5] = l1; red[2][threadIdx.x >> 5] = l2;
    }
    __syncthreads();
    if (threadIdx.x < 3) {
      float s = 0.f;
      for (int i = 0; i < 4; ++i) s += red[threadIdx.x][i];
      la.losses[threadIdx.x] = s / (float)B;
    }
    if (threadIdx.x == 0 && la.advance && la.counters) { la.counters[0] += 1; la.counters[1] += 1; }
  }
  const int k8 = blockIdx.x * 128 + threadIdx.x;
  const int n0 = blockIdx.y * kHeadBwdChunk;
  if (k8 * 8 >= F) return;
  float wv[8], acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { wv[j] = w[k8 * 8 + j]; acc[j] = 0.f; }
  uint4 hv[kHeadBwdChunk];
  float dl[kHeadBwdChunk];
#pragma unroll
  for (int i = 0; i < kHeadBwdChunk; ++i) {
    const int n = n0 + i;
    if (kFusedLoss) {
      dl[i] = 0.f;
      if (n < 3 * B) {
        const float x = la.logits[n < 2 * B ? n : n - B];  // slab 2 (g_loss) re-uses the fake logits
        const float sg = 1.f / (1.f + __expf(-x));
        dl[i] = ((n >= B && n < 2 * B) ? sg : sg - 1.f) * lscale;
      }
    } else {
      dl[i] = n < 3 * B ? dlogit[n] : 0.f;
    }
    hv[i] = make_uint4(0, 0, 0, 0);
    if (n < 2 * B) hv[i] = __ldg(reinterpret_cast<const uint4*>(h + (size_t)n * F) + k8);
  }
  float bsum = 0.f;
#pragma unroll
  for (int i = 0; i < kHeadBwdChunk; ++i) {
    const int n = n0 + i;
    if (n >= 3 * B) break;
    const float d = dl[i];
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = pack_bf16x2(d * wv[2 * j], d * wv[2 * j + 1]);
    reinterpret_cast<uint4*>(dh + (size_t)n * F)[k8] = make_uint4(o[0], o[1], o[2], o[3]);
    if (n < 2 * B) {
      const uint32_t u[4] = {hv[i].x, hv[i].y, hv[i].z, hv[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(u[j]);
        acc[2 * j] += d * f.x; acc[2 * j + 1] += d * f.y;
      }
      bsum += d;
    }
  }
  if (n0 < 2 * B) {
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&dW[k8 * 8 + j], acc[j]);
    if (k8 == 0) atomicAdd(db, bsum);
  }
}

// ---------------------------------------------------------------------------------------------
// TF-form Adam on a flat f32 range + bf16 shadow copy of the updated parameters.
//   lr_t = lr*sqrt(1-b2^t)/(1-b1^t);  m = b1*m+(1-b1)*g;  v = b2*v+(1-b2)*g^2;
//   p -= lr_t*m/(sqrt(v)+eps)       (eps outside the bias correction -- SURVEY.md 2.4d)
// t is read from device memory (counters[0], already advanced for this step) so that the launch can
// live in a CUDA graph.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            __nv_bfloat16* __restrict__ p_bf16, const long long* __restrict__ counters, size_t n, float lr,
            float b1, float b2, float eps, float gscale) {
  pdl_launch_dependents();
  pdl_wait();
  // the bias-corrected step size is the same for every thread: one thread evaluates the two powf (a few hundred
  // instructions; ncu showed this kernel at 60 % issue utilisation, which matters while it runs beside G's
  // backward pass) and the block reads it from shared memory
  __shared__ float s_lr_t;
  if (threadIdx.x == 0) {
    const float t = (float)counters[0];
    s_lr_t = lr * sqrtf(1.f - powf(b2, t)) / (1.f - powf(b1, t));
  }
  __syncthreads();
  const float lr_t = s_lr_t;
  // grid-stride: the launch may cap its block count (dcgan_set_adam_blocks) to leave SMs to concurrent kernels
  for (size_t i4 = ((size_t)blockIdx.x * 256 + threadIdx.x) * 4; i4 < n; i4 += (size_t)gridDim.x * 1024) {
  if (i4 + 4 <= n) {
    float4 pv = *reinterpret_cast<float4*>(p + i4);
    const float4 gv = *reinterpret_cast<const float4*>(g + i4);
    float4 mv = *reinterpret_cast<float4*>(m + i4);
    float4 vv = *reinterpret_cast<float4*>(v + i4);
    float* pp = &pv.x; const float* gp = &gv.x; float* mp = &mv.x; float* vp = &vv.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gj = gp[j] * gscale;
      mp[j] = b1 * mp[j] + (1.f - b1) * gj;
      vp[j] = b2 * vp[j] + (1.f - b2) * gj * gj;
      pp[j] -= lr_t * mp[j] / (sqrtf(vp[j]) + eps);
    }
    *reinterpret_cast<float4*>(p + i4) = pv;
    *reinterpret_cast<float4*>(m + i4) = mv;
    *reinterpret_cast<float4*>(v + i4) = vv;
    if (p_bf16) {
      uint2 o = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
      *reinterpret_cast<uint2*>(p_bf16 + i4) = o;
    }
  } else {
    for (size_t i = i4; i < n; ++i) {
      const float gj = g[i] * gscale;
      m[i] = b1 * m[i] + (1.f - b1) * gj;
      v[i] = b2 * v[i] + (1.f - b2) * gj * gj;
      p[i] -= lr_t * m[i] / (sqrtf(v[i]) + eps);
      if (p_bf16) p_bf16[i] = __float2bfloat16(p[i]);
    }
  }
  }
}

__global__ void __launch_bounds__(256)
cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256)
    dst[i] = __float2bfloat16(src[i]);
}

// per-tap transpose of a bf16 weight: src [T][R][C] -> dst [T][C][R]   (32x32 smem tiles)
__global__ void __launch_bounds__(256)
transpose_taps_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, int R, int C) {
  __shared__ __nv_bfloat16 tile[32][34];
  const int t = blockIdx.z;
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const __nv_bfloat16* s = src + (size_t)t * R * C;
  __nv_bfloat16* d = dst + (size_t)t * R * C;
  for (int i = ty; i < 32; i += 8)
    if (r0 + i < R && c0 + tx < C) tile[i][tx] = s[(size_t)(r0 + i) * C + c0 + tx];
  __syncthreads();
  for (int i = ty; i < 32; i += 8)
    if (c0 + i < C && r0 + tx < R) d[(size_t)(c0 + i) * R + r0 + tx] = tile[tx][i];
}

// column sums of a bf16 [rows][C] matrix with a small channel count (C <= 8; the 3-channel image
// gradients): out[c] += sum_r x[r][c]   (out zeroed by the caller).  One row per thread per step.
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, size_t rows, int C,
              size_t rows_per_block) {
  __shared__ float red[8];
  if (threadIdx.x < 8) red[threadIdx.x] = 0.f;
  __syncthreads();
  float s[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) s[c] = 0.f;
  const size_t r0 = (size_t)blockIdx.x * rows_per_block;
  const size_t r1 = min(r0 + rows_per_block, rows);
  for (size_t r = r0 + threadIdx.x; r < r1; r += 256) {
#pragma unroll
    for (int c = 0; c < 8; ++c)
      if (c < C) s[c] += __bfloat162float(x[r * C + c]);
  }
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    if (c < C) {
      const float v = warp_sum(s[c]);
      if ((threadIdx.x & 31) == 0) atomicAdd(&red[c], v);
    }
  }
  __syncthreads();
  if (threadIdx.x < C) atomicAdd(&out[threadIdx.x], red[threadIdx.x]);
}

// input decode: float64 records (image_input.py:48 decode_raw(.., tf.float64)) -> bf16 NHWC
__global__ void __launch_bounds__(256)
decode_f64_kernel(const double* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n, float scale,
                  float shift) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256)
    dst[i] = __float2bfloat16((float)src[i] * scale + shift);
}
// the same cast with float32 output: what the reference's tf.cast(image, tf.float32) (image_input.py:51) does, moved
// to the device so that the host threads only copy bytes
__global__ void __launch_bounds__(256)
decode_f64_f32_kernel(const double* __restrict__ src, float* __restrict__ dst, size_t n, float scale, float shift) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256)
    dst[i] = (float)src[i] * scale + shift;
}
__global__ void __launch_bounds__(256)
f32_to_bf16_affine_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n, float scale,
                          float shift) {
  pdl_launch_dependents();
  pdl_wait();
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256)
    dst[i] = __float2bfloat16(src[i] * scale + shift);
}

// sample grid (image_train.py:197-219): inverse_transform (x+1)/2, tile `cols` images per row into
// one [rows*H, cols*W, C] uint8 image.  imsave's min/max rescale is applied on the host.
__global__ void __launch_bounds__(256)
grid_merge_kernel(const __nv_bfloat16* __restrict__ img, float* __restrict__ out, int N, int H, int W, int C,
                  int cols) {
  const size_t total = (size_t)N * H * W * C;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int c = i % C;
    const int x = (i / C) % W;
    const int y = (i / ((size_t)C * W)) % H;
    const int n = i / ((size_t)C * W * H);
    const int gi = n % cols, gj = n / cols;
    const float v = (__bfloat162float(img[i]) + 1.f) * 0.5f;
    out[(((size_t)gj * H + y) * (cols * W) + (size_t)gi * W + x) * C + c] = v;
  }
}

// zero fraction of a bf16 tensor ("sparsity" summary, distriubted_model.py:80): count[0] += zeros
__global__ void __launch_bounds__(256)
count_zeros_kernel(const __nv_bfloat16* __restrict__ x, unsigned long long* __restrict__ count, size_t n) {
  unsigned int c = 0;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256)
    c += (__bfloat162float(x[i]) == 0.f);
  c = __reduce_add_sync(0xffffffffu, c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, (unsigned long long)c);
}

}  // namespace dcgan

using namespace dcgan;
#define LAUNCH_OK() (cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH)

static int g_adam_blocks = 0;  // 0 = one block per 1024 parameters; > 0 caps the grid (grid-stride loop)

extern "C" {

void dcgan_set_adam_blocks(int v) { g_adam_blocks = v > 0 ? v : 0; }

int dcgan_lin_fwd(const float* z, const void* W, const float* bias, void* out, int NB, int K, int F,
                  cudaStream_t st) {
  if (F % 2) return DCGAN_ERR_SHAPE;
  dim3 grid((F / 2 + 127) / 128, (NB + kLinTile - 1) / kLinTile);
  launch_k_light(lin_fwd_kernel, dim3(grid), dim3(128), kLinTile * K * sizeof(float), st, z, reinterpret_cast<const __nv_bfloat16*>(W),
                                                                  bias, reinterpret_cast<__nv_bfloat16*>(out), NB,
                                                                  K, F);
  return LAUNCH_OK();
}

int dcgan_lin_wgrad(const float* z, const void* dy, float* dW, float* dbias, int NB, int K, int F,
                    cudaStream_t st) {
  if ((size_t)NB * 20 * sizeof(float) > 48 * 1024) return DCGAN_ERR_SHAPE;
  dim3 grid((F + 127) / 128, (K + 19) / 20);
  launch_k_light(lin_wgrad_kernel, dim3(grid), dim3(128), NB * 20 * sizeof(float), st, z, reinterpret_cast<const __nv_bfloat16*>(dy), dW,
                                                               dbias, NB, K, F);
  return LAUNCH_OK();
}

int dcgan_head_fwd(const void* h, const float* w, const float* b, float* logits, int rows, int F,
                   cudaStream_t st) {
  if (F % 8) return DCGAN_ERR_SHAPE;
  launch_k_light(head_fwd_kernel, dim3(rows), dim3(256), 0, st, reinterpret_cast<const __nv_bfloat16*>(h), w, b, logits, F);
  return LAUNCH_OK();
}

int dcgan_gan_loss(const float* logits, float* losses, float* dlogit, float* probs, long long* counters, int B,
                   float grad_div, int advance, cudaStream_t st) {
  launch_k_light(gan_loss_kernel, dim3(1), dim3(256), 0, st, logits, losses, dlogit, probs, counters, B, grad_div, advance);
  return LAUNCH_OK();
}

int dcgan_head_bwd(const void* h, const float* w, const float* dlogit, void* dh, float* dW, float* db, int B,
                   int F, cudaStream_t st) {
  if (F % 8) return DCGAN_ERR_SHAPE;
  const int threads_needed = F / 8;
  dim3 grid((threads_needed + 127) / 128, (3 * B + kHeadBwdChunk - 1) / kHeadBwdChunk);
  launch_k_light(head_bwd_kernel<false>, dim3(grid), dim3(128), 0, st, reinterpret_cast<const __nv_bfloat16*>(h), w,
           dlogit, reinterpret_cast<__nv_bfloat16*>(dh), dW, db, B, F, LossArgs{});
  return LAUNCH_OK();
}

int dcgan_head_loss_bwd(const void* h, const float* w, const float* logits, float* losses, float* dlogit,
                        float* probs, long long* counters, void* dh, float* dW, float* db, int B, int F,
                        float grad_div, int advance, cudaStream_t st) {
  if (F % 8 || logits == nullptr || losses == nullptr) return DCGAN_ERR_SHAPE;
  const int threads_needed = F / 8;
  dim3 grid((threads_needed + 127) / 128, (3 * B + kHeadBwdChunk - 1) / kHeadBwdChunk);
  LossArgs la{logits, losses, dlogit, probs, counters, grad_div, advance};
  launch_k_light(head_bwd_kernel<true>, dim3(grid), dim3(128), 0, st, reinterpret_cast<const __nv_bfloat16*>(h), w,
           static_cast<const float*>(nullptr), reinterpret_cast<__nv_bfloat16*>(dh), dW, db, B, F, la);
  return LAUNCH_OK();
}

int dcgan_adam(float* p, const float* g, float* m, float* v, void* p_bf16, const long long* counters, size_t n,
               float lr, float b1, float b2, float eps, float gscale, cudaStream_t st) {
  if (n == 0) return 0;
  size_t blocks = (n + 1023) / 1024;
  if (g_adam_blocks > 0 && blocks > (size_t)g_adam_blocks) blocks = g_adam_blocks;
  launch_k(adam_kernel, dim3((unsigned)blocks), dim3(256), 0, st, p, g, m, v, reinterpret_cast<__nv_bfloat16*>(p_bf16), counters, n,
                                               lr, b1, b2, eps, gscale);
  return LAUNCH_OK();
}

int dcgan_cast_bf16(const float* src, void* dst, size_t n, cudaStream_t st) {
  if (n == 0) return 0;
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  cast_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  return LAUNCH_OK();
}

int dcgan_transpose_taps(const void* src, void* dst, int taps, int R, int C, cudaStream_t st) {
  dim3 grid((C + 31) / 32, (R + 31) / 32, taps);
  transpose_taps_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(src),
                                              reinterpret_cast<__nv_bfloat16*>(dst), R, C);
  return LAUNCH_OK();
}

int dcgan_colsum(const void* x, float* out, size_t rows, int C, cudaStream_t st) {
  if (C > 8) return DCGAN_ERR_SHAPE;
  size_t blocks = 296;
  size_t rpb = (rows + blocks - 1) / blocks;
  if (rpb < 64) rpb = 64;
  blocks = (rows + rpb - 1) / rpb;
  colsum_kernel<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), out,
                                                                 rows, C, rpb);
  return LAUNCH_OK();
}

int dcgan_decode_f64(const double* src, void* dst, size_t n, float scale, float shift, cudaStream_t st) {
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  decode_f64_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), n, scale, shift);
  return LAUNCH_OK();
}

int dcgan_decode_f64_f32(const double* src, float* dst, size_t n, float scale, float shift, cudaStream_t st) {
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  decode_f64_f32_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, dst, n, scale, shift);
  return LAUNCH_OK();
}

int dcgan_f32_to_bf16(const float* src, void* dst, size_t n, float scale, float shift, cudaStream_t st) {
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  launch_k_light(f32_to_bf16_affine_kernel, dim3((unsigned)blocks), dim3(256), 0, st, src, reinterpret_cast<__nv_bfloat16*>(dst), n, scale,
                                                              shift);
  return LAUNCH_OK();
}

int dcgan_grid_merge(const void* img, float* out, int N, int H, int W, int C, int cols, cudaStream_t st) {
  const size_t total = (size_t)N * H * W * C;
  size_t blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  grid_merge_kernel<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(img), out, N, H, W, C,
                                                      cols);
  return LAUNCH_OK();
}

int dcgan_count_zeros(const void* x, unsigned long long* count, size_t n, cudaStream_t st) {
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 4) blocks = 148 * 4;
  count_zeros_kernel<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), count, n);
  return LAUNCH_OK();
}

}  // extern "C"
