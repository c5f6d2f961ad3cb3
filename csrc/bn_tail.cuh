// Batch-norm elementwise passes executed by a whole 1-D grid after a grid-wide barrier (csrc/common.cuh:
// grid_barrier): the tail of the conv kernels that fuse batch norm (csrc/igemm2_sm100.cu) and of the one-launch
// batch-norm kernels (csrc/bn_act.cu).  Same arithmetic as bn_apply_kernel / bn_bwd_apply_kernel.
//
// Thread t of a CTA owns the 8-channel vector t % (C/8) and walks rows; rows are interleaved across CTAs in batches
// of kTailRows{Fwd,Bwd} per thread so that all loads of a batch are in flight together (the tensors are L2-resident).
// Tensors produced earlier in the same launch by other CTAs are read with ld.global.cg.
#pragma once
#include "common.cuh"

namespace dcgan {

// rows in flight per thread: at the conv kernels' occupancy (12 warps per SM) the pass is bound by memory-level
// parallelism -- four 16-byte loads per thread kept ~3.6 MB in flight chip-wide and cost 12.7 us per layer, three
// times a stand-alone bn_apply
constexpr int kTailRowsFwd = 12;
constexpr int kTailRowsBwd = 8;

__device__ __forceinline__ uint4 ldcg_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float ldcg_f(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// y = act(gamma * (x - mean) * rstd + beta) with the moments finalised from `sums` ([groups][2][C]: sum, sum of
// squares over rows_per_group rows); CTA 0 also writes saved = (mean, rstd) and advances the EMA shadows.
template <int kThreads>
__device__ __forceinline__ void bn_fwd_apply_grid(const __nv_bfloat16* x, __nv_bfloat16* y, const float* sums,
                                                  const float* __restrict__ gamma, const float* __restrict__ beta,
                                                  float* saved, float* ema_mean, float* ema_var, int groups,
                                                  int rows_per_group, int C, float eps, float decay, int act) {
  const int cvec = C >> 3;
  const int lanes = kThreads / cvec;
  const int cv = threadIdx.x % cvec;
  const int rl = threadIdx.x / cvec;
  if (rl >= lanes) return;
  const float inv_count = 1.0f / (float)rows_per_group;
  const int stride = gridDim.x * lanes * kTailRowsFwd;
  for (int g = 0; g < groups; ++g) {
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cv * 8 + j;
      const float mean = ldcg_f(sums + ((size_t)g * 2) * C + c) * inv_count;
      const float var = fmaxf(ldcg_f(sums + ((size_t)g * 2 + 1) * C + c) * inv_count - mean * mean, 0.f);
      const float rstd = rsqrtf(var + eps);
      sc[j] = gamma[c] * rstd;
      sh[j] = beta[c] - mean * sc[j];
      if (blockIdx.x == 0 && rl == 0) {
        if (saved) { saved[((size_t)g * 2) * C + c] = mean; saved[((size_t)g * 2 + 1) * C + c] = rstd; }
        if (ema_mean) {  // shadow -= (1 - decay) * (shadow - value)
          float* em = ema_mean + (size_t)g * C + c;
          float* ev = ema_var + (size_t)g * C + c;
          *em -= (1.f - decay) * (*em - mean);
          *ev -= (1.f - decay) * (*ev - var);
        }
      }
    }
    const uint4* xb = reinterpret_cast<const uint4*>(x) + (size_t)g * rows_per_group * cvec + cv;
    uint4* yb = reinterpret_cast<uint4*>(y) + (size_t)g * rows_per_group * cvec + cv;
    for (int rb = blockIdx.x * lanes * kTailRowsFwd + rl; rb < rows_per_group; rb += stride) {
      uint4 v[kTailRowsFwd];
#pragma unroll
      for (int u = 0; u < kTailRowsFwd; ++u) {
        const int r = rb + u * lanes;
        v[u] = r < rows_per_group ? ldcg_u4(xb + (size_t)r * cvec) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < kTailRowsFwd; ++u) {
        const int r = rb + u * lanes;
        if (r >= rows_per_group) break;
        const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
        float f[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 t = unpack_bf16x2(w[j]);
          f[2 * j] = t.x * sc[2 * j] + sh[2 * j];
          f[2 * j + 1] = t.y * sc[2 * j + 1] + sh[2 * j + 1];
        }
        apply_act_chunk(f, act);
        yb[(size_t)r * cvec] = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                                          pack_bf16x2(f[6], f[7]));
      }
    }
  }
}

// dpre = gamma * rstd * (dy - mean(dy) - xhat * mean(dy * xhat)), dy = da * act'(gamma * xhat + beta), for `slabs`
// slabs of rows_per_slab rows; slab g uses the statistics / pre-activations of group min(g, stat_groups - 1) and the
// reductions red[g] ([slabs][2][C]: sum dy, sum dy * xhat).  CTA 0 also writes dgamma / dbeta = the reductions summed
// over the first param_slabs slabs.
template <int kThreads>
__device__ __forceinline__ void bn_bwd_apply_grid(const __nv_bfloat16* da, const __nv_bfloat16* __restrict__ pre,
                                                  __nv_bfloat16* dpre, const float* __restrict__ saved,
                                                  const float* __restrict__ gamma, const float* __restrict__ beta,
                                                  const float* red, float* dgamma, float* dbeta, int slabs,
                                                  int stat_groups, int param_slabs, int rows_per_slab, int C,
                                                  int act) {
  const int cvec = C >> 3;
  const int lanes = kThreads / cvec;
  const int cv = threadIdx.x % cvec;
  const int rl = threadIdx.x / cvec;
  if (rl >= lanes) return;
  const float inv_count = 1.0f / (float)rows_per_slab;
  const int stride = gridDim.x * lanes * kTailRowsBwd;
  const float pos = 1.f;
  const float neg = act == kActRelu ? 0.f : (act == kActLrelu ? 0.2f : 1.f);
  for (int g = 0; g < slabs; ++g) {
    const int sg = min(g, stat_groups - 1);
    float mean[8], rstd[8], gm[8], bt[8], m1[8], m2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cv * 8 + j;
      mean[j] = saved[((size_t)sg * 2) * C + c];
      rstd[j] = saved[((size_t)sg * 2 + 1) * C + c];
      gm[j] = gamma[c];
      bt[j] = beta[c];
      m1[j] = ldcg_f(red + ((size_t)g * 2) * C + c) * inv_count;
      m2[j] = ldcg_f(red + ((size_t)g * 2 + 1) * C + c) * inv_count;
      if (blockIdx.x == 0 && rl == 0 && g == 0 && dgamma != nullptr) {
        float dg = 0.f, db = 0.f;
        for (int q = 0; q < param_slabs; ++q) {
          db += ldcg_f(red + ((size_t)q * 2) * C + c);
          dg += ldcg_f(red + ((size_t)q * 2 + 1) * C + c);
        }
        dgamma[c] = dg;
        dbeta[c] = db;
      }
    }
    const uint4* dab = reinterpret_cast<const uint4*>(da) + (size_t)g * rows_per_slab * cvec + cv;
    const uint4* prb = reinterpret_cast<const uint4*>(pre) + (size_t)sg * rows_per_slab * cvec + cv;
    uint4* dpb = reinterpret_cast<uint4*>(dpre) + (size_t)g * rows_per_slab * cvec + cv;
    for (int rb = blockIdx.x * lanes * kTailRowsBwd + rl; rb < rows_per_slab; rb += stride) {
      uint4 dvv[kTailRowsBwd], pvv[kTailRowsBwd];
#pragma unroll
      for (int u = 0; u < kTailRowsBwd; ++u) {
        const int r = rb + u * lanes;
        dvv[u] = pvv[u] = make_uint4(0, 0, 0, 0);
        if (r < rows_per_slab) {
          dvv[u] = ldcg_u4(dab + (size_t)r * cvec);
          pvv[u] = __ldg(prb + (size_t)r * cvec);
        }
      }
#pragma unroll
      for (int u = 0; u < kTailRowsBwd; ++u) {
        const int r = rb + u * lanes;
        if (r >= rows_per_slab) break;
        const uint32_t dw[4] = {dvv[u].x, dvv[u].y, dvv[u].z, dvv[u].w};
        const uint32_t pw[4] = {pvv[u].x, pvv[u].y, pvv[u].z, pvv[u].w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 d = unpack_bf16x2(dw[j]);
          const float2 p = unpack_bf16x2(pw[j]);
          const float xh0 = (p.x - mean[2 * j]) * rstd[2 * j];
          const float xh1 = (p.y - mean[2 * j + 1]) * rstd[2 * j + 1];
          const float dy0 = d.x * ((gm[2 * j] * xh0 + bt[2 * j]) > 0.f ? pos : neg);
          const float dy1 = d.y * ((gm[2 * j + 1] * xh1 + bt[2 * j + 1]) > 0.f ? pos : neg);
          o[j] = pack_bf16x2(gm[2 * j] * rstd[2 * j] * (dy0 - m1[2 * j] - xh0 * m2[2 * j]),
                             gm[2 * j + 1] * rstd[2 * j + 1] * (dy1 - m1[2 * j + 1] - xh1 * m2[2 * j + 1]));
        }
        dpb[(size_t)r * cvec] = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
}

}  // namespace dcgan
