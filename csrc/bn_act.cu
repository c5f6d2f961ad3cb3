// Batch-norm kernels fused with their activations (forward and backward), NHWC bf16 activations.
//
// Reference semantics (distriubted_model.py:15-52): per-channel batch moments over N*H*W with
// *biased* variance, y = gamma*(x-mean)*rsqrt(var+1e-5)+beta, followed by ReLU (generator,
// :91-105) or LeakyReLU(0.2) (discriminator, :119-121); the moments feed an exponential moving
// average with decay 0.9 (:41-45) that the sampler consumes (:131-153).
//
// "Groups": the discriminator runs on a [real; fake] batch but the reference computes separate
// statistics for the two calls, so every kernel takes a group count and treats
// group g = contiguous slab of rows_per_group rows.  The backward kernels take a third slab (the
// g_loss gradient through D(fake)) that shares the statistics of slab 1 (stat_group map).
#include <atomic>
#include <cstdio>

#include "common.cuh"
#include "bn_tail.cuh"
#include "../include/dcgan_b200.h"

namespace dcgan {

constexpr int kBnThreads = 256;
constexpr int kRowBatch = 4;  // rows in flight per thread in the streaming loops

// Column sums of per-thread partials -> global accumulators, without shared-memory atomics: f32 atomicAdd on
// shared memory compiles to a compare-and-swap loop (ATOMS.CAST.SPIN) and all row-lanes of one channel would
// spin on the same word (profiles/ncu_step_v3_summary.md: the backward reduce ran at 10 % of DRAM bandwidth).
// Every row-lane stores its NV x 8 partials in its own slice of `part` ([lanes][NV*C], at most NV * 2048
// floats because lanes * C <= 8 * kBnThreads); then each thread adds one column over the row-lanes.
template <int NV>
__device__ __forceinline__ void block_colsum_to_global(float* part, const float (&v)[NV][8], bool active, int rl,
                                                       int cv, int lanes, int C, float* __restrict__ out) {
  if (active) {
#pragma unroll
    for (int n = 0; n < NV; ++n) {
      float4* dst = reinterpret_cast<float4*>(part + (size_t)(rl * NV + n) * C + cv * 8);
      dst[0] = make_float4(v[n][0], v[n][1], v[n][2], v[n][3]);
      dst[1] = make_float4(v[n][4], v[n][5], v[n][6], v[n][7]);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NV * C; i += kBnThreads) {
    float acc = 0.f;
    for (int l = 0; l < lanes; ++l) acc += part[(size_t)l * NV * C + i];
    atomicAdd(&out[i], acc);
  }
}

// ---------------------------------------------------------------------------------------------
// forward statistics: sums[g][0][c] += sum x, sums[g][1][c] += sum x^2   (zeroed by the caller)
// grid = (row_chunks, groups); thread t handles 8 channels (t % (C/8)) of rows t/(C/8)+k*R
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void bn_stats_body(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums,
                                              int rows_per_group, int C, int rows_per_block, int g, int chunk,
                                              float* part) {
  const int cvec = C >> 3;
  const int lanes = kBnThreads / cvec;  // rows processed in parallel (C <= 2048)
  const int cv = threadIdx.x % cvec;
  const int rl = threadIdx.x / cvec;
  float acc[2][8];
  float (&s)[8] = acc[0];
  float (&ss)[8] = acc[1];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.f; ss[j] = 0.f; }
  const int r0 = chunk * rows_per_block;
  const int r1 = min(r0 + rows_per_block, rows_per_group);
  if (rl < lanes) {
    const __nv_bfloat16* base = x + ((size_t)g * rows_per_group) * C + cv * 8;
    for (int r = r0 + rl; r < r1; r += lanes) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(base + (size_t)r * C));
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        s[2 * j] += f.x; ss[2 * j] += f.x * f.x;
        s[2 * j + 1] += f.y; ss[2 * j + 1] += f.y * f.y;
      }
    }
  }
  block_colsum_to_global<2>(part, acc, rl < lanes, rl, cv, lanes, C, sums + (size_t)g * 2 * C);
}

__global__ void __launch_bounds__(kBnThreads)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums, int rows_per_group, int C,
                int rows_per_block) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ __align__(16) float part[2 * 8 * kBnThreads];
  bn_stats_body(x, sums, rows_per_group, C, rows_per_block, blockIdx.y, blockIdx.x, part);
}

// ---------------------------------------------------------------------------------------------
// forward apply: finalises the moments, saves (mean, rstd) for backward, updates the EMA shadows
// (block 0 of each group) and writes y = act(gamma*(x-mean)*rstd + beta) as bf16.
// use_ema != 0: normalise with the EMA shadows instead (the sampler path, train=False).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBnThreads)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                const float* __restrict__ sums, const float* __restrict__ gamma,
                const float* __restrict__ beta, float* __restrict__ saved /*[G][2][C]*/,
                float* __restrict__ ema_mean /*[G][C] or null*/, float* __restrict__ ema_var,
                int rows_per_group, int C, float inv_count, float eps, float decay, int act, int use_ema,
                int rows_per_block) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float ss[];  // scale[C], shift[C]
  const int g = blockIdx.y;
  for (int c = threadIdx.x; c < C; c += kBnThreads) {
    float mean, var;
    if (use_ema) {
      mean = ema_mean[(size_t)g * C + c];
      var = ema_var[(size_t)g * C + c];
    } else {
      mean = sums[((size_t)g * 2) * C + c] * inv_count;
      var = fmaxf(sums[((size_t)g * 2 + 1) * C + c] * inv_count - mean * mean, 0.f);
    }
    const float rstd = rsqrtf(var + eps);
    const float sc = gamma[c] * rstd;
    ss[c] = sc;
    ss[C + c] = beta[c] - mean * sc;
    if (blockIdx.x == 0 && !use_ema) {
      if (saved) { saved[((size_t)g * 2) * C + c] = mean; saved[((size_t)g * 2 + 1) * C + c] = rstd; }
      if (ema_mean) {  // shadow -= (1-decay) * (shadow - value)
        float* em = ema_mean + (size_t)g * C + c;
        float* ev = ema_var + (size_t)g * C + c;
        *em -= (1.f - decay) * (*em - mean);
        *ev -= (1.f - decay) * (*ev - var);
      }
    }
  }
  __syncthreads();
  // thread -> fixed 8-channel vector, rows strided: the per-channel scale/shift stay in registers
  const int cvec = C >> 3;
  const int lanes = kBnThreads / cvec;
  const int cv = threadIdx.x % cvec;
  const int rl = threadIdx.x / cvec;
  if (rl >= lanes) return;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sc[j] = ss[cv * 8 + j]; sh[j] = ss[C + cv * 8 + j]; }
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(r0 + rows_per_block, rows_per_group);
  const uint4* xb = reinterpret_cast<const uint4*>(x) + (size_t)g * rows_per_group * cvec + cv;
  uint4* yb = reinterpret_cast<uint4*>(y) + (size_t)g * rows_per_group * cvec + cv;
  // kRowBatch rows per trip, all loads issued before the first use: the tensors are L2 resident and the loop is
  // latency-bound otherwise
  for (int rb = r0 + rl; rb < r1; rb += lanes * kRowBatch) {
    uint4 v[kRowBatch];
#pragma unroll
    for (int u = 0; u < kRowBatch; ++u) {
      const int r = rb + u * lanes;
      v[u] = r < r1 ? __ldg(xb + (size_t)r * cvec) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < kRowBatch; ++u) {
      const int r = rb + u * lanes;
      if (r >= r1) break;
      const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        o[j] = pack_bf16x2(apply_act(f.x * sc[2 * j] + sh[2 * j], act),
                           apply_act(f.y * sc[2 * j + 1] + sh[2 * j + 1], act));
      }
      yb[(size_t)r * cvec] = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// backward reduce: for upstream slab g (statistics / pre-activations of slab stat_group(g)):
//   dy = da * act'(gamma*xhat+beta);  red[g][0][c] += sum dy;  red[g][1][c] += sum dy*xhat
// stat group of slab g is min(g, n_stat_groups-1).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float act_grad(float y, int act) {
  if (act == kActRelu) return y > 0.f ? 1.f : 0.f;
  if (act == kActLrelu) return y > 0.f ? 1.f : 0.2f;
  return 1.f;
}

__device__ __forceinline__ void bn_bwd_reduce_body(const __nv_bfloat16* __restrict__ da,
                                                   const __nv_bfloat16* __restrict__ pre,
                                                   const float* __restrict__ saved, const float* __restrict__ gamma,
                                                   const float* __restrict__ beta, float* __restrict__ red_out,
                                                   int rows_per_group, int C, int n_stat_groups, int act,
                                                   int rows_per_block, int g, int chunk, float* part) {
  const int sg = min(g, n_stat_groups - 1);
  const int cvec = C >> 3;
  const int lanes = kBnThreads / cvec;
  const int cv = threadIdx.x % cvec;
  const int rl = threadIdx.x / cvec;
  float acc[2][8];
  float (&s1)[8] = acc[0];
  float (&s2)[8] = acc[1];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
  if (rl < lanes) {
    float mean[8], rstd[8], gm[8], bt[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cv * 8 + j;
      mean[j] = saved[((size_t)sg * 2) * C + c];
      rstd[j] = saved[((size_t)sg * 2 + 1) * C + c];
      gm[j] = gamma[c]; bt[j] = beta[c];
    }
    const int r0 = chunk * rows_per_block;
    const int r1 = min(r0 + rows_per_block, rows_per_group);
    const __nv_bfloat16* dab = da + ((size_t)g * rows_per_group) * C + cv * 8;
    const __nv_bfloat16* prb = pre + ((size_t)sg * rows_per_group) * C + cv * 8;
    for (int rb = r0 + rl; rb < r1; rb += lanes * kRowBatch) {
      uint4 dvv[kRowBatch], pvv[kRowBatch];
#pragma unroll
      for (int u = 0; u < kRowBatch; ++u) {
        const int r = rb + u * lanes;
        dvv[u] = pvv[u] = make_uint4(0, 0, 0, 0);
        if (r < r1) {
          dvv[u] = __ldg(reinterpret_cast<const uint4*>(dab + (size_t)r * C));
          pvv[u] = __ldg(reinterpret_cast<const uint4*>(prb + (size_t)r * C));
        }
      }
#pragma unroll
      for (int u = 0; u < kRowBatch; ++u) {
        if (rb + u * lanes >= r1) break;
        const uint32_t dw[4] = {dvv[u].x, dvv[u].y, dvv[u].z, dvv[u].w};
        const uint32_t pw[4] = {pvv[u].x, pvv[u].y, pvv[u].z, pvv[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 d = unpack_bf16x2(dw[j]);
          const float2 p = unpack_bf16x2(pw[j]);
          const float xh0 = (p.x - mean[2 * j]) * rstd[2 * j];
          const float xh1 = (p.y - mean[2 * j + 1]) * rstd[2 * j + 1];
          const float dy0 = d.x * act_grad(gm[2 * j] * xh0 + bt[2 * j], act);
          const float dy1 = d.y * act_grad(gm[2 * j + 1] * xh1 + bt[2 * j + 1], act);
          s1[2 * j] += dy0; s2[2 * j] += dy0 * xh0;
          s1[2 * j + 1] += dy1; s2[2 * j + 1] += dy1 * xh1;
        }
      }
    }
  }
  block_colsum_to_global<2>(part, acc, rl < lanes, rl, cv, lanes, C, red_out + (size_t)g * 2 * C);
}

__global__ void __launch_bounds__(kBnThreads, 2)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ da, const __nv_bfloat16* __restrict__ pre,
                     const float* __restrict__ saved, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ red_out, int rows_per_group, int C,
                     int n_stat_groups, int act, int rows_per_block) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ __align__(16) float part[2 * 8 * kBnThreads];
  bn_bwd_reduce_body(da, pre, saved, gamma, beta, red_out, rows_per_group, C, n_stat_groups, act, rows_per_block,
                     blockIdx.y, blockIdx.x, part);
}

// ---------------------------------------------------------------------------------------------
// One-launch variants for the layers whose producer is not an implicit-GEMM conv (which would carry the batch norm
// in its own tail, csrc/igemm2_sm100.cu): statistics (or backward reductions) by row chunks, a grid-wide barrier,
// then the elementwise pass by the whole grid.  1-D grid of groups * chunks CTAs, all resident (host checks).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBnThreads)
bn_forward_fused_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, float* sums,
                        const float* __restrict__ gamma, const float* __restrict__ beta, float* saved,
                        float* ema_mean, float* ema_var, int groups, int rows_per_group, int C, float eps, float decay,
                        int act, int rows_per_block, int chunks, int* barrier, int* err_flag) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ __align__(16) float part[2 * 8 * kBnThreads];
  bn_stats_body(x, sums, rows_per_group, C, rows_per_block, blockIdx.x / chunks, blockIdx.x % chunks, part);
  grid_barrier(barrier, err_flag);
  bn_fwd_apply_grid<kBnThreads>(x, y, sums, gamma, beta, saved, ema_mean, ema_var, groups, rows_per_group, C, eps,
                                decay, act);
}

__global__ void __launch_bounds__(kBnThreads)
bn_backward_fused_kernel(const __nv_bfloat16* __restrict__ da, const __nv_bfloat16* __restrict__ pre,
                         __nv_bfloat16* __restrict__ dpre, const float* __restrict__ saved,
                         const float* __restrict__ gamma, const float* __restrict__ beta, float* red, float* dgamma,
                         float* dbeta, int slabs, int n_stat_groups, int param_slabs, int rows_per_group, int C,
                         int act, int rows_per_block, int chunks, int* barrier, int* err_flag) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ __align__(16) float part[2 * 8 * kBnThreads];
  bn_bwd_reduce_body(da, pre, saved, gamma, beta, red, rows_per_group, C, n_stat_groups, act, rows_per_block,
                     blockIdx.x / chunks, blockIdx.x % chunks, part);
  grid_barrier(barrier, err_flag);
  bn_bwd_apply_grid<kBnThreads>(da, pre, dpre, saved, gamma, beta, red, dgamma, dbeta, slabs, n_stat_groups,
                                param_slabs, rows_per_group, C, act);
}

// ---------------------------------------------------------------------------------------------
// backward apply: dpre = gamma*rstd*(dy - mean(dy) - xhat*mean(dy*xhat)) as bf16; block (0,0)
// also emits dgamma = sum over the first n_param_groups slabs of sum(dy*xhat), dbeta likewise.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ da, const __nv_bfloat16* __restrict__ pre,
                    __nv_bfloat16* __restrict__ dpre, const float* __restrict__ saved,
                    const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ red, float* __restrict__ dgamma, float* __restrict__ dbeta,
                    int rows_per_group, int C, int n_stat_groups, int n_param_groups, float inv_count, int act,
                    int rows_per_block) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float sm[];  // mean, rstd, gamma, beta, m1, m2 : 6*C
  const int g = blockIdx.y;
  const int sg = min(g, n_stat_groups - 1);
  float* s_mean = sm; float* s_rstd = sm + C; float* s_g = sm + 2 * C; float* s_b = sm + 3 * C;
  float* s_m1 = sm + 4 * C; float* s_m2 = sm + 5 * C;
  for (int c = threadIdx.x; c < C; c += kBnThreads) {
    s_mean[c] = saved[((size_t)sg * 2) * C + c];
    s_rstd[c] = saved[((size_t)sg * 2 + 1) * C + c];
    s_g[c] = gamma[c]; s_b[c] = beta[c];
    s_m1[c] = red[((size_t)g * 2) * C + c] * inv_count;
    s_m2[c] = red[((size_t)g * 2 + 1) * C + c] * inv_count;
    if (blockIdx.x == 0 && g == 0 && dgamma) {
      float dg = 0.f, db = 0.f;
      for (int q = 0; q < n_param_groups; ++q) {
        db += red[((size_t)q * 2) * C + c];
        dg += red[((size_t)q * 2 + 1) * C + c];
      }
      dgamma[c] = dg; dbeta[c] = db;
    }
  }
  __syncthreads();
  const int cvec = C >> 3;
  const int lanes = kBnThreads / cvec;
  const int cv = threadIdx.x % cvec;
  const int rl = threadIdx.x / cvec;
  if (rl >= lanes) return;
  float mean[8], rstd[8], gm[8], bt[8], m1[8], m2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = cv * 8 + j;
    mean[j] = s_mean[c]; rstd[j] = s_rstd[c]; gm[j] = s_g[c]; bt[j] = s_b[c]; m1[j] = s_m1[c]; m2[j] = s_m2[c];
  }
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(r0 + rows_per_block, rows_per_group);
  const uint4* dab = reinterpret_cast<const uint4*>(da) + (size_t)g * rows_per_group * cvec + cv;
  const uint4* prb = reinterpret_cast<const uint4*>(pre) + (size_t)sg * rows_per_group * cvec + cv;
  uint4* dpb = reinterpret_cast<uint4*>(dpre) + (size_t)g * rows_per_group * cvec + cv;
  for (int rb = r0 + rl; rb < r1; rb += lanes * kRowBatch) {
    uint4 dvv[kRowBatch], pvv[kRowBatch];
#pragma unroll
    for (int u = 0; u < kRowBatch; ++u) {
      const int r = rb + u * lanes;
      dvv[u] = pvv[u] = make_uint4(0, 0, 0, 0);
      if (r < r1) {
        dvv[u] = __ldg(dab + (size_t)r * cvec);
        pvv[u] = __ldg(prb + (size_t)r * cvec);
      }
    }
#pragma unroll
    for (int u = 0; u < kRowBatch; ++u) {
      const int r = rb + u * lanes;
      if (r >= r1) break;
      const uint32_t dw[4] = {dvv[u].x, dvv[u].y, dvv[u].z, dvv[u].w};
      const uint32_t pw[4] = {pvv[u].x, pvv[u].y, pvv[u].z, pvv[u].w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 d = unpack_bf16x2(dw[j]);
        const float2 p = unpack_bf16x2(pw[j]);
        const float xh0 = (p.x - mean[2 * j]) * rstd[2 * j];
        const float xh1 = (p.y - mean[2 * j + 1]) * rstd[2 * j + 1];
        const float dy0 = d.x * act_grad(gm[2 * j] * xh0 + bt[2 * j], act);
        const float dy1 = d.y * act_grad(gm[2 * j + 1] * xh1 + bt[2 * j + 1], act);
        o[j] = pack_bf16x2(gm[2 * j] * rstd[2 * j] * (dy0 - m1[2 * j] - xh0 * m2[2 * j]),
                           gm[2 * j + 1] * rstd[2 * j + 1] * (dy1 - m1[2 * j + 1] - xh1 * m2[2 * j + 1]));
      }
      dpb[(size_t)r * cvec] = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// activation backward without batch norm (d_h0: bias + LeakyReLU only, distriubted_model.py:118):
// dpre = da * act'(a) using the sign of the saved post-activation; also the bias gradient
// dbias[c] += sum over the first bias_rows rows of dpre.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBnThreads)
act_bwd_kernel(const __nv_bfloat16* __restrict__ da, const __nv_bfloat16* __restrict__ a,
               __nv_bfloat16* __restrict__ dpre, float* __restrict__ dbias, int rows, int a_rows, int C, int act,
               int bias_rows, int rows_per_block) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ __align__(16) float part[8 * kBnThreads];
  const int cvec = C >> 3;
  const int lanes = kBnThreads / cvec;
  const int cv = threadIdx.x % cvec;
  const int rl = threadIdx.x / cvec;
  float acc[1][8];
  float (&s)[8] = acc[0];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  if (rl < lanes) {
    const int r0 = blockIdx.x * rows_per_block;
    const int r1 = min(r0 + rows_per_block, rows);
    for (int rb = r0 + rl; rb < r1; rb += lanes * kRowBatch) {
      uint4 dvv[kRowBatch], avv[kRowBatch];
#pragma unroll
      for (int u = 0; u < kRowBatch; ++u) {
        const int r = rb + u * lanes;
        dvv[u] = avv[u] = make_uint4(0, 0, 0, 0);
        if (r < r1) {
          const int ar = r < a_rows ? r : r - (rows - a_rows);  // third slab re-uses the fake slab's activations
          dvv[u] = __ldg(reinterpret_cast<const uint4*>(da + (size_t)r * C + cv * 8));
          avv[u] = __ldg(reinterpret_cast<const uint4*>(a + (size_t)ar * C + cv * 8));
        }
      }
#pragma unroll
      for (int u = 0; u < kRowBatch; ++u) {
        const int r = rb + u * lanes;
        if (r >= r1) break;
        const uint32_t dw[4] = {dvv[u].x, dvv[u].y, dvv[u].z, dvv[u].w};
        const uint32_t aw[4] = {avv[u].x, avv[u].y, avv[u].z, avv[u].w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 d = unpack_bf16x2(dw[j]);
          const float2 p = unpack_bf16x2(aw[j]);
          const float r0v = d.x * act_grad(p.x, act);
          const float r1v = d.y * act_grad(p.y, act);
          if (r < bias_rows) { s[2 * j] += r0v; s[2 * j + 1] += r1v; }
          o[j] = pack_bf16x2(r0v, r1v);
        }
        *reinterpret_cast<uint4*>(dpre + (size_t)r * C + cv * 8) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
  if (dbias) block_colsum_to_global<1>(part, acc, rl < lanes, rl, cv, lanes, C, dbias);
}

static int pick_rows_per_block(int rows, int C, int groups) {
  // aim for ~2 waves of 148 SMs with at least a few KB of work per block
  int blocks_target = 296 / (groups > 0 ? groups : 1);
  if (blocks_target < 1) blocks_target = 1;
  int rpb = (rows + blocks_target - 1) / blocks_target;
  const int min_rows = (8192 / C) > 1 ? (8192 / C) : 1;
  if (rpb < min_rows) rpb = min_rows;
  return rpb;
}

// resident CTA capacity of a kernel on the current device (the grid barrier needs the whole grid resident)
template <typename K>
static int resident_ctas(K kern, int* cache) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  if (cache[dev] == 0) {
    int per_sm = 0, sms = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBnThreads, 0) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      return 0;
    cache[dev] = per_sm * sms;
  }
  return cache[dev];
}

static int fused_chunks(int rows_per_group, int C, int groups, int capacity, int* rpb) {
  *rpb = pick_rows_per_block(rows_per_group, C, groups);
  int chunks = (rows_per_group + *rpb - 1) / *rpb;
  if (chunks * groups > capacity) {
    chunks = capacity / groups;
    if (chunks < 1) return 0;
    *rpb = (rows_per_group + chunks - 1) / chunks;
    chunks = (rows_per_group + *rpb - 1) / *rpb;
  }
  return chunks;
}

}  // namespace dcgan

using namespace dcgan;

extern "C" {

int dcgan_bn_stats(const void* x, float* sums, int groups, int rows_per_group, int C, cudaStream_t st) {
  if (C % 8 || C / 8 > kBnThreads) return DCGAN_ERR_SHAPE;
  const int rpb = pick_rows_per_block(rows_per_group, C, groups);
  dim3 grid((rows_per_group + rpb - 1) / rpb, groups);
  launch_k_light(bn_stats_kernel, dim3(grid), dim3(kBnThreads), 0, st, reinterpret_cast<const __nv_bfloat16*>(x), sums,
                                                                  rows_per_group, C, rpb);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_bn_apply(const void* x, void* y, const float* sums, const float* gamma, const float* beta, float* saved,
                   float* ema_mean, float* ema_var, int groups, int rows_per_group, int C, float eps, float decay,
                   int act, int use_ema, cudaStream_t st) {
  if (C % 8 || C / 8 > kBnThreads) return DCGAN_ERR_SHAPE;
  const int rpb = pick_rows_per_block(rows_per_group, C, groups);
  dim3 grid((rows_per_group + rpb - 1) / rpb, groups);
  launch_k_light(bn_apply_kernel, dim3(grid), dim3(kBnThreads), 2 * C * sizeof(float), st, 
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), sums, gamma, beta, saved,
      ema_mean, ema_var, rows_per_group, C, 1.0f / (float)rows_per_group, eps, decay, act, use_ema, rpb);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_bn_bwd_reduce(const void* da, const void* pre, const float* saved, const float* gamma, const float* beta,
                        float* red, int groups, int stat_groups, int rows_per_group, int C, int act,
                        cudaStream_t st) {
  if (C % 8 || C / 8 > kBnThreads) return DCGAN_ERR_SHAPE;
  const int rpb = pick_rows_per_block(rows_per_group, C, groups);
  dim3 grid((rows_per_group + rpb - 1) / rpb, groups);
  launch_k_light(bn_bwd_reduce_kernel, dim3(grid), dim3(kBnThreads), 0, st, 
      reinterpret_cast<const __nv_bfloat16*>(da), reinterpret_cast<const __nv_bfloat16*>(pre), saved, gamma, beta,
      red, rows_per_group, C, stat_groups, act, rpb);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_bn_bwd_apply(const void* da, const void* pre, void* dpre, const float* saved, const float* gamma,
                       const float* beta, const float* red, float* dgamma, float* dbeta, int groups,
                       int stat_groups, int param_groups, int rows_per_group, int C, int act, cudaStream_t st) {
  if (C % 8 || C / 8 > kBnThreads) return DCGAN_ERR_SHAPE;
  const int rpb = pick_rows_per_block(rows_per_group, C, groups);
  dim3 grid((rows_per_group + rpb - 1) / rpb, groups);
  launch_k_light(bn_bwd_apply_kernel, dim3(grid), dim3(kBnThreads), 6 * C * sizeof(float), st, 
      reinterpret_cast<const __nv_bfloat16*>(da), reinterpret_cast<const __nv_bfloat16*>(pre),
      reinterpret_cast<__nv_bfloat16*>(dpre), saved, gamma, beta, red, dgamma, dbeta, rows_per_group, C,
      stat_groups, param_groups, 1.0f / (float)rows_per_group, act, rpb);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_bn_forward_fused(const void* x, void* y, float* sums, const float* gamma, const float* beta, float* saved,
                           float* ema_mean, float* ema_var, int* barrier, int groups, int rows_per_group, int C,
                           float eps, float decay, int act, cudaStream_t st) {
  if (C % 8 || C / 8 > kBnThreads || groups < 1 || barrier == nullptr) return DCGAN_ERR_SHAPE;
  static int cache[64] = {0};
  const int cap = resident_ctas(bn_forward_fused_kernel, cache);
  int rpb;
  const int chunks = fused_chunks(rows_per_group, C, groups, cap, &rpb);
  if (chunks < 1) return DCGAN_ERR_LAUNCH;
  int* err = error_flag_ptr();
  if (!err) return DCGAN_ERR_DRIVER;
  launch_k_light(bn_forward_fused_kernel, dim3(groups * chunks), dim3(kBnThreads), 0, st,
           reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), sums, gamma, beta, saved,
           ema_mean, ema_var, groups, rows_per_group, C, eps, decay, act, rpb, chunks, barrier, err);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_bn_backward_fused(const void* da, const void* pre, void* dpre, const float* saved, const float* gamma,
                            const float* beta, float* red, float* dgamma, float* dbeta, int* barrier, int groups,
                            int stat_groups, int param_groups, int rows_per_group, int C, int act, cudaStream_t st) {
  if (C % 8 || C / 8 > kBnThreads || groups < 1 || barrier == nullptr) return DCGAN_ERR_SHAPE;
  static int cache[64] = {0};
  const int cap = resident_ctas(bn_backward_fused_kernel, cache);
  int rpb;
  const int chunks = fused_chunks(rows_per_group, C, groups, cap, &rpb);
  if (chunks < 1) return DCGAN_ERR_LAUNCH;
  int* err = error_flag_ptr();
  if (!err) return DCGAN_ERR_DRIVER;
  launch_k_light(bn_backward_fused_kernel, dim3(groups * chunks), dim3(kBnThreads), 0, st,
           reinterpret_cast<const __nv_bfloat16*>(da), reinterpret_cast<const __nv_bfloat16*>(pre),
           reinterpret_cast<__nv_bfloat16*>(dpre), saved, gamma, beta, red, dgamma, dbeta, groups, stat_groups,
           param_groups, rows_per_group, C, act, rpb, chunks, barrier, err);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_act_bwd(const void* da, const void* a, void* dpre, float* dbias, int rows, int a_rows, int C, int act,
                  int bias_rows, cudaStream_t st) {
  if (C % 8 || C / 8 > kBnThreads) return DCGAN_ERR_SHAPE;
  const int rpb = pick_rows_per_block(rows, C, 1);
  dim3 grid((rows + rpb - 1) / rpb);
  launch_k_light(act_bwd_kernel, dim3(grid), dim3(kBnThreads), 0, st, 
      reinterpret_cast<const __nv_bfloat16*>(da), reinterpret_cast<const __nv_bfloat16*>(a),
      reinterpret_cast<__nv_bfloat16*>(dpre), dbias, rows, a_rows, C, act, bias_rows, rpb);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

}  // extern "C"

// =================================================================================================
// Allocation-free batch-norm forward.  No counterpart in the reference.
//
// dcgan_bn_stats + dcgan_bn_apply need a per-channel sum / sum-of-squares buffer between them, and
// normally the caller supplies it.  Some callers have nowhere to get one from: a plain C host that was
// handed its activation buffers by another runtime's memory pool and owns no allocator, or a service
// loop whose latency budget rules out cudaMalloc because that call stalls the whole device.  For them
// the buffer lives in the library image itself, below.
//
//   * Capacity: kStaticBnFloats floats (64 KiB).  groups * 2 * C must fit; larger requests are refused
//     with DCGAN_ERR_LIMIT and a message on stderr.  Every GPU that loads the library carries these
//     64 KiB until the process exits.
//   * `g_static_bn_sums` is a device symbol.  In host code that name is only a stand-in; handing it to a
//     kernel would make the kernel fault (or, where the GPU can reach host memory, scribble on the wrong
//     memory).  The device address is looked up once per device with cudaGetSymbolAddress, with that
//     device current, and that pointer is what the kernels receive.
//   * The lookup happens inside the first call made on each device, so make that first call outside of
//     any stream capture.
//   * All calls share the one array and it still holds the previous call's sums, so each call clears
//     the part it uses first (stream-ordered).  Two calls that overlap on the GPU -- different streams,
//     or different host threads on the same device -- race on it; keep calls on one stream or serialise.
//   * Nothing here calls cudaMalloc, cudaFree, or any routine that allocates.
// =================================================================================================
namespace dcgan {

constexpr int kStaticBnFloats = 16384;
static __device__ __align__(256) float g_static_bn_sums[kStaticBnFloats];

static float* static_bn_scratch_for_current_device(cudaError_t* err) {
  constexpr int kMaxDev = 64;
  // Filled on first use per device.  Host threads may race to fill a slot: they all store the same address (the
  // symbol's address on that device never changes), and the atomic keeps the store / load well defined.
  static std::atomic<float*> per_device[kMaxDev];
  int dev = -1;
  *err = cudaGetDevice(&dev);
  if (*err != cudaSuccess) return nullptr;
  if (dev < 0 || dev >= kMaxDev) { *err = cudaErrorInvalidDevice; return nullptr; }
  float* ptr = per_device[dev].load(std::memory_order_acquire);
  if (ptr == nullptr) {
    *err = cudaGetSymbolAddress((void**)&ptr, g_static_bn_sums);
    if (*err != cudaSuccess) return nullptr;
    per_device[dev].store(ptr, std::memory_order_release);
  }
  return ptr;
}

}  // namespace dcgan

extern "C" int dcgan_bn_forward_static(const void* x, void* y, const float* gamma, const float* beta,
                                       float* saved, float* ema_mean, float* ema_var, int groups,
                                       int rows_per_group, int C, float eps, float decay, int act,
                                       cudaStream_t st) {
  using namespace dcgan;
  if (groups <= 0 || rows_per_group <= 0 || C <= 0 || C % 8 || C / 8 > kBnThreads) return DCGAN_ERR_SHAPE;
  const size_t need = (size_t)groups * 2 * (size_t)C;
  if (need > (size_t)kStaticBnFloats) {
    fprintf(stderr,
            "dcgan_bn_forward_static: groups*2*C = %zu floats exceeds the built-in scratch of %d floats; "
            "use dcgan_bn_stats + dcgan_bn_apply with a caller-provided buffer instead\n",
            need, kStaticBnFloats);
    return DCGAN_ERR_LIMIT;
  }
  cudaError_t err = cudaSuccess;
  float* sums = static_bn_scratch_for_current_device(&err);
  if (sums == nullptr) {
    fprintf(stderr, "dcgan_bn_forward_static: cudaGetSymbolAddress failed: %s\n", cudaGetErrorString(err));
    return DCGAN_ERR_DRIVER;
  }
  if (cudaMemsetAsync(sums, 0, need * sizeof(float), st) != cudaSuccess) return DCGAN_ERR_LAUNCH;
  const int rpb = pick_rows_per_block(rows_per_group, C, groups);
  dim3 grid((rows_per_group + rpb - 1) / rpb, groups);
  launch_k_light(bn_stats_kernel, dim3(grid), dim3(kBnThreads), 0, st, reinterpret_cast<const __nv_bfloat16*>(x), sums,
                                                                  rows_per_group, C, rpb);
  if (cudaGetLastError() != cudaSuccess) return DCGAN_ERR_LAUNCH;
  launch_k_light(bn_apply_kernel, dim3(grid), dim3(kBnThreads), 2 * C * sizeof(float), st, 
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), sums, gamma, beta, saved,
      ema_mean, ema_var, rows_per_group, C, 1.0f / (float)rows_per_group, eps, decay, act, 0, rpb);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}
