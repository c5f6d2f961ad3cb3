// In-kernel collectives over NVLink 5 / NVSwitch peer memory -- no NCCL on these paths.
//
// These replace the reference's parameter-server traffic: the gradient *push*
// (worker -> ps, image_train.py:109-112) and the parameter *pull* (ps -> worker, every read of a
// ps-pinned variable, distriubted_model.py:70) that TensorFlow 0.x moved as gRPC Send/Recv.
//
//   rs_adam_ag      sync replicas: ONE kernel = reduce-scatter of the f32 gradients (P2P loads from every
//                   peer, or an in-switch NVLS multimem.ld_reduce) + average + TF-form Adam on this rank's
//                   parameter shard (the "ps shard": f32 master and both Adam slots stay sharded) +
//                   all-gather of the updated bf16 parameters (P2P stores into every peer, or multimem.st).
//   allreduce_small cross-GPU batch-norm statistics (sum, sum-of-squares / backward reductions).
//   async_apply     asynchronous parameter-server mode: a worker applies its own stale gradients to the
//                   owner's shard through peer memory (unlocked, like the reference's use_locking=False)
//                   and keeps what it read as its fresh local copy.
//
// Signalling: every rank owns a small symmetric "signal pad"; a rank announces a phase by a
// release-store of the epoch into its slot of every peer's pad and waits, bounded, for all slots of its
// own pad.  Epochs live in device memory so the launches are CUDA-graph replayable.
#include <cstdio>

#include "common.cuh"
#include "../include/dcgan_b200.h"

namespace dcgan {

constexpr int kMaxRanks = 16;
constexpr int kCommThreads = 256;

struct PeerTable {
  void* buf_a[kMaxRanks];       // rs_adam_ag: f32 gradients        async: f32 master
  void* buf_b[kMaxRanks];       // rs_adam_ag: bf16 parameters      async: bf16 parameters
  void* buf_c[kMaxRanks];       // async: Adam m
  void* buf_d[kMaxRanks];       // async: Adam v
  unsigned* sig[kMaxRanks];     // signal pads
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_b64(void* mc, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(mc), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y))
               : "memory");
}

// bounded wait until every rank's slot in `mine` has reached `epoch`
// polls before a rank gives up on its peers (about 0.7 us each: ~12 s by default); host-settable, because a
// chief that writes a checkpoint or fills its sample queue between two steps keeps its peers waiting for longer
static __device__ unsigned g_peer_spin_limit = 1u << 24;
__device__ __forceinline__ bool wait_all(const unsigned* mine, int world, unsigned epoch, int* err_flag) {
  bool ok = true;
  if ((int)threadIdx.x < world) {
    unsigned spins = 0;
    const unsigned limit = g_peer_spin_limit;
    while ((int)(ld_acquire_sys(mine + threadIdx.x) - epoch) < 0) {
      if (++spins > limit) {
        atomicExch(err_flag, (int)kErrPeerTimeout);
        ok = false;
        break;
      }
      __nanosleep(64);
    }
  }
  return __syncthreads_and((int)ok) != 0;
}

// Shard size in units of 4 elements, rounded up to an even count so that shard boundaries fall on multiples of
// 8 elements (= one 16-byte chunk of the bf16 shadow).  Every kernel and the Python side use this split.
__host__ __device__ inline size_t shard_quads(size_t total4, int world) {
  return (((total4 + world - 1) / world) + 1) & ~(size_t)1;
}

// signal layout inside a pad: [slot][phase 0/1][src rank], kMaxRanks entries per row
__device__ __forceinline__ unsigned* sig_row(unsigned* pad, int slot, int phase) {
  return pad + ((size_t)slot * 2 + phase) * kMaxRanks;
}

// ---------------------------------------------------------------------------------------------------------
// reduce-scatter + average + Adam + all-gather
// state[0] = epoch of this slot, state[1] = finished-block counter
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCommThreads)
rs_adam_ag_kernel(PeerTable peers, float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                  const long long* __restrict__ counters, unsigned* __restrict__ state, int rank, int world,
                  size_t lo, size_t hi, float lr, float b1, float b2, float eps, int slot, int* err_flag,
                  const float* mc_grad, __nv_bfloat16* mc_bf16, float* mc_master, size_t f32_from, size_t pull_lo,
                  size_t pull_hi) {
  // f32_from : elements >= f32_from are read in f32 by their consumers -> their master value is gathered too
  //            (peers.buf_c = every rank's master array)
  // [pull_lo, pull_hi) : not pushed at all; the consumer kernel pulls this range from the owners itself
  const unsigned epoch = state[0] + 1;
  // ---- phase A: every rank's gradients are complete -------------------------------------------------------
  if (blockIdx.x == 0 && (int)threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(sig_row(peers.sig[threadIdx.x], slot, 0) + rank, epoch);
  }
  const bool ok = wait_all(sig_row(peers.sig[rank], slot, 0), world, epoch, err_flag);

  // the bias-corrected step size is evaluated once per block, not by every thread (two powf ~ 300 instructions)
  __shared__ float s_lr_t;
  if (threadIdx.x == 0) {
    const float t = (float)counters[0];
    s_lr_t = lr * sqrtf(1.f - powf(b2, t)) / (1.f - powf(b1, t));
  }
  __syncthreads();
  const float lr_t = s_lr_t;
  const float inv_world = 1.f / (float)world;
  // shard of [lo, hi) owned by this rank (multiples of 4 elements)
  const size_t total4 = (hi - lo) / 4;
  const size_t per4 = shard_quads(total4, world);
  const size_t s0 = lo + min((size_t)rank * per4, total4) * 4;
  const size_t s1 = lo + min((size_t)(rank + 1) * per4, total4) * 4;
  if (ok) {
    // kUnroll quads per thread and trip, every load (peer / in-switch reduction, master, both moments) issued before
    // the first use: one trip costs one NVLink round trip (~2-3 us), so the loop is bound by trips, not by bytes
    constexpr int kUnroll = 4;
    const size_t stride = (size_t)gridDim.x * kCommThreads * 4;
    for (size_t base = s0 + ((size_t)blockIdx.x * kCommThreads + threadIdx.x) * 4; base < s1;
         base += stride * kUnroll) {
      float4 g[kUnroll], pv[kUnroll], mv[kUnroll], vv[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t i = base + (size_t)u * stride;
        g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < s1) {
          if (mc_grad != nullptr) {
            g[u] = multimem_ld_reduce_f4(mc_grad + i);  // summed inside the NVSwitch
          } else {
#pragma unroll 4
            for (int r = 0; r < world; ++r) {
              const float4 x = ld_peer_f4(reinterpret_cast<const float*>(peers.buf_a[(rank + r) % world]) + i);
              g[u].x += x.x; g[u].y += x.y; g[u].z += x.z; g[u].w += x.w;
            }
          }
          pv[u] = *reinterpret_cast<float4*>(master + i);
          mv[u] = *reinterpret_cast<float4*>(m + i);
          vv[u] = *reinterpret_cast<float4*>(v + i);
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t i = base + (size_t)u * stride;
        if (i >= s1) break;
        float* gp = &g[u].x; float* pp = &pv[u].x; float* mp = &mv[u].x; float* vp = &vv[u].x;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float gj = gp[j] * inv_world;
          mp[j] = b1 * mp[j] + (1.f - b1) * gj;
          vp[j] = b2 * vp[j] + (1.f - b2) * gj * gj;
          pp[j] -= lr_t * mp[j] / (sqrtf(vp[j]) + eps);
        }
        *reinterpret_cast<float4*>(master + i) = pv[u];
        *reinterpret_cast<float4*>(m + i) = mv[u];
        *reinterpret_cast<float4*>(v + i) = vv[u];
        const uint2 packed = make_uint2(pack_bf16x2(pv[u].x, pv[u].y), pack_bf16x2(pv[u].z, pv[u].w));
        if (i >= pull_lo && i < pull_hi) {
          // gathered by its first consumer (lin_fwd_gather): only the owner's own shadow is written here
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(peers.buf_b[rank]) + i) = packed;
        } else if (mc_bf16 != nullptr) {
          multimem_st_b64(mc_bf16 + i, packed);  // one store, every rank's copy
        } else {
#pragma unroll 4
          for (int r = 0; r < world; ++r)
            *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(peers.buf_b[(rank + r) % world]) + i) = packed;
        }
        if (i >= f32_from) {
          if (mc_master != nullptr) {
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_master + i),
                         "f"(pv[u].x), "f"(pv[u].y), "f"(pv[u].z), "f"(pv[u].w)
                         : "memory");
          } else {
            for (int r = 1; r < world; ++r)
              *reinterpret_cast<float4*>(reinterpret_cast<float*>(peers.buf_c[(rank + r) % world]) + i) = pv[u];
          }
        }
      }
    }
  }
  // ---- phase B: every owner's parameters have landed everywhere ----------------------------------------------
  __syncthreads();
  __shared__ int is_last;
  if (threadIdx.x == 0) {
    __threadfence_system();
    is_last = (atomicAdd(&state[1], 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    if ((int)threadIdx.x < world) {
      __threadfence_system();
      st_release_sys(sig_row(peers.sig[threadIdx.x], slot, 1) + rank, epoch);
    }
    wait_all(sig_row(peers.sig[rank], slot, 1), world, epoch, err_flag);
    if (threadIdx.x == 0) {
      state[1] = 0;
      state[0] = epoch;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// small all-reduce (mean) for batch-norm statistics: one block.  staging[rank] is this rank's symmetric
// staging area (double-buffered by epoch parity, so one signal phase per call is enough).
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCommThreads)
allreduce_small_kernel(PeerTable peers /* buf_a = staging */, float* __restrict__ data, int n, int staging_stride,
                       unsigned* __restrict__ state, int rank, int world, int slot, int* err_flag) {
  const unsigned epoch = state[0] + 1;
  const int parity = epoch & 1;
  float* mine = reinterpret_cast<float*>(peers.buf_a[rank]) + (size_t)parity * staging_stride;
  for (int i = threadIdx.x; i < n; i += kCommThreads) mine[i] = data[i];
  __syncthreads();
  if ((int)threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(sig_row(peers.sig[threadIdx.x], slot, 0) + rank, epoch);
  }
  const bool ok = wait_all(sig_row(peers.sig[rank], slot, 0), world, epoch, err_flag);
  if (ok) {
    // all peer loads of a thread are issued before the first use (NVLink latency ~2 us: never wait per element)
    const float inv = 1.f / (float)world;
    constexpr int kMaxVec = 8;  // 8 x float4 x 256 threads = 8192 floats = staging capacity
    const int nvec = n >> 2;
    for (int r0 = 0; r0 < world; r0 += 4) {
      float4 v[4][kMaxVec];
#pragma unroll
      for (int rr = 0; rr < 4; ++rr)
#pragma unroll
        for (int k = 0; k < kMaxVec; ++k) {
          const int i4 = threadIdx.x + k * kCommThreads;
          v[rr][k] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (r0 + rr < world && i4 < nvec)
            v[rr][k] = ld_peer_f4(reinterpret_cast<const float*>(peers.buf_a[r0 + rr]) + (size_t)parity * staging_stride + i4 * 4);
        }
#pragma unroll
      for (int k = 0; k < kMaxVec; ++k) {
        const int i4 = threadIdx.x + k * kCommThreads;
        if (i4 < nvec) {
          float4 acc = (r0 == 0) ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<float4*>(data + i4 * 4);
#pragma unroll
          for (int rr = 0; rr < 4; ++rr) {
            acc.x += v[rr][k].x; acc.y += v[rr][k].y; acc.z += v[rr][k].z; acc.w += v[rr][k].w;
          }
          if (r0 + 4 >= world) { acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv; }
          *reinterpret_cast<float4*>(data + i4 * 4) = acc;
        }
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) state[0] = epoch;
}

// ---------------------------------------------------------------------------------------------------------
// asynchronous parameter-server apply: this worker's gradients -> Adam on the owners' shards over P2P.
// No barrier, no lock: concurrent workers may interleave on an element, as in the reference.
// step_counters (on every rank; the owner's copy counts): [0] = Adam t of that shard, rank 0's [1] = global_step.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCommThreads)
async_apply_kernel(PeerTable peers, const float* __restrict__ grad, __nv_bfloat16* __restrict__ local_bf16,
                   float* __restrict__ local_master, long long* const* __restrict__ peer_counters, int rank,
                   int world, size_t n, size_t lo, size_t hi, int advance, float lr, float b1, float b2, float eps) {
  // [lo, hi) = the part of the arena applied by this launch (ownership is always by the whole arena's shards);
  // advance != 0: this launch is the one of the step that advances the owners' step counts and global_step
  const size_t total4 = n / 4;
  const size_t per4 = shard_quads(total4, world);
  __shared__ float s_lr_t[kMaxRanks];
  if ((int)threadIdx.x < world) {
    // every pusher advances the owner's step count (beta powers advance per apply, like TF's shared
    // beta1_power); block 0 does the increment, the others read a possibly older value
    long long* c = peer_counters[threadIdx.x];
    long long t;
    if (blockIdx.x == 0 && advance) {
      t = atomicAdd_system(reinterpret_cast<unsigned long long*>(c), 1ull) + 1;
      if (threadIdx.x == 0) atomicAdd_system(reinterpret_cast<unsigned long long*>(peer_counters[0] + 1), 1ull);
    } else {
      t = *reinterpret_cast<volatile long long*>(c) + 1;
    }
    const float tf = (float)(t < 1 ? 1 : t);
    s_lr_t[threadIdx.x] = lr * sqrtf(1.f - powf(b2, tf)) / (1.f - powf(b1, tf));
  }
  __syncthreads();
  // kUnroll quads per thread and trip, all peer loads issued before the first use (a trip = one NVLink round trip)
  constexpr int kUnroll = 4;
  const size_t stride4 = (size_t)gridDim.x * kCommThreads;
  const size_t last4 = min(hi / 4, total4);
  for (size_t b4 = lo / 4 + (size_t)blockIdx.x * kCommThreads + threadIdx.x; b4 < last4; b4 += stride4 * kUnroll) {
    float4 pv[kUnroll], mv[kUnroll], vv[kUnroll], g[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t i4 = b4 + (size_t)u * stride4;
      if (i4 < last4) {
        const int owner = (int)min(i4 / per4, (size_t)(world - 1));
        const size_t i = i4 * 4;
        pv[u] = ld_peer_f4(reinterpret_cast<float*>(peers.buf_a[owner]) + i);
        mv[u] = ld_peer_f4(reinterpret_cast<float*>(peers.buf_c[owner]) + i);
        vv[u] = ld_peer_f4(reinterpret_cast<float*>(peers.buf_d[owner]) + i);
        g[u] = *reinterpret_cast<const float4*>(grad + i);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t i4 = b4 + (size_t)u * stride4;
      if (i4 >= last4) break;
      const int owner = (int)min(i4 / per4, (size_t)(world - 1));
      const size_t i = i4 * 4;
      const float* gp = &g[u].x; float* pp = &pv[u].x; float* mp = &mv[u].x; float* vp = &vv[u].x;
      const float lr_t = s_lr_t[owner];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        mp[j] = b1 * mp[j] + (1.f - b1) * gp[j];
        vp[j] = b2 * vp[j] + (1.f - b2) * gp[j] * gp[j];
        pp[j] -= lr_t * mp[j] / (sqrtf(vp[j]) + eps);
      }
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(peers.buf_a[owner]) + i) = pv[u];
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(peers.buf_c[owner]) + i) = mv[u];
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(peers.buf_d[owner]) + i) = vv[u];
      const uint2 packed = make_uint2(pack_bf16x2(pv[u].x, pv[u].y), pack_bf16x2(pv[u].z, pv[u].w));
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(peers.buf_b[owner]) + i) = packed;
      if (owner != rank) {  // what we just read and updated is our fresh pull (bf16 shadow + f32 for the f32 readers)
        *reinterpret_cast<uint2*>(local_bf16 + i) = packed;
        *reinterpret_cast<float4*>(local_master + i) = pv[u];
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// g_h0_lin with its parameter all-gather fused in: out[n, j] = sum_k z[n,k] * W[k,j] + b[j], where every
// element of W is loaded from the bf16 shadow of the rank that owns (and just updated) it -- P2P loads over
// NVLink from inside the consumer -- and mirrored into the local shadow on the way.  W occupies
// [w_off, w_off + K*F) of the flat arena; ownership follows the optimizer kernel's shard split of [lo, hi).
// ---------------------------------------------------------------------------------------------------------
constexpr int kGatherLinTile = 16;
__global__ void __launch_bounds__(128)
lin_fwd_gather_kernel(PeerTable peers /* buf_b = bf16 shadows */, const float* __restrict__ z,
                      const float* __restrict__ bias, __nv_bfloat16* __restrict__ out, int NB, int K, int F,
                      size_t w_off, size_t lo, size_t hi, int rank, int world) {
  extern __shared__ uint8_t gsm[];
  float* zs = reinterpret_cast<float*>(gsm);                                       // [kGatherLinTile][K]
  uint4* wsm = reinterpret_cast<uint4*>(gsm + (size_t)kGatherLinTile * K * 4);     // [K][16] x 16 B = W[:, j0:j0+128]
  const int n0 = blockIdx.y * kGatherLinTile;
  const int j0 = blockIdx.x * 128;
  // ---- gather: this block's 128 columns of W, 16-byte peer loads from whichever rank owns each chunk --------
  const size_t total4 = (hi - lo) / 4;
  const size_t per4 = shard_quads(total4, world);
  uint4* local = reinterpret_cast<uint4*>(peers.buf_b[rank]);
  for (int c = threadIdx.x; c < K * 16; c += 128) {
    const int k = c >> 4, ch = c & 15;
    const size_t idx = w_off + (size_t)k * F + j0 + ch * 8;   // first of 8 bf16; a chunk never straddles shards
    const int owner = (int)min((idx - lo) / 4 / per4, (size_t)(world - 1));
    const uint4 v = __ldcg(reinterpret_cast<const uint4*>(peers.buf_b[owner]) + idx / 8);
    wsm[c] = v;
    if (blockIdx.y == 0 && owner != rank) local[idx / 8] = v;  // mirror into the local shadow
  }
  for (int i = threadIdx.x; i < kGatherLinTile * K; i += 128) {
    const int n = n0 + i / K;
    zs[i] = n < NB ? z[(size_t)n * K + (i % K)] : 0.f;
  }
  __syncthreads();
  // ---- GEMM slice from shared memory ------------------------------------------------------------------------
  const int j = j0 + threadIdx.x;
  if (j >= F) return;
  float acc[kGatherLinTile];
  const float b = bias ? bias[j] : 0.f;
#pragma unroll
  for (int i = 0; i < kGatherLinTile; ++i) acc[i] = b;
  const unsigned short* wcol = reinterpret_cast<const unsigned short*>(wsm) + threadIdx.x;
  for (int k = 0; k < K; ++k) {
    const float w = __uint_as_float((unsigned)wcol[k * 128] << 16);
#pragma unroll
    for (int i = 0; i < kGatherLinTile; ++i) acc[i] += zs[i * K + k] * w;
  }
#pragma unroll
  for (int i = 0; i < kGatherLinTile; ++i)
    if (n0 + i < NB) out[(size_t)(n0 + i) * F + j] = __float2bfloat16(acc[i]);
}

static void fill(void** dst, const void* const* src, int world) {
  for (int i = 0; i < kMaxRanks; ++i) dst[i] = i < world ? const_cast<void*>(src[i]) : nullptr;
}

}  // namespace dcgan

using namespace dcgan;

extern "C" {

int dcgan_comm_set_peer_spin_limit(unsigned polls) {
  if (polls < (1u << 10)) polls = 1u << 10;
  return cudaMemcpyToSymbol(g_peer_spin_limit, &polls, sizeof(unsigned)) == cudaSuccess ? 0 : DCGAN_ERR_DRIVER;
}

int dcgan_comm_rs_adam_ag(const void* const* grad_ptrs, const void* const* bf16_ptrs, const void* const* sig_ptrs,
                          const void* const* master_ptrs, float* master, float* m, float* v,
                          const long long* counters, unsigned* state, int rank, int world, size_t lo, size_t hi,
                          float lr, float b1, float b2, float eps, int slot, const float* mc_grad, void* mc_bf16,
                          float* mc_master, size_t f32_from, size_t pull_lo, size_t pull_hi, int blocks,
                          cudaStream_t st) {
  if (world < 1 || world > kMaxRanks || (lo % 4) || (hi % 4) || hi < lo || (f32_from % 4) || (pull_lo % 4) ||
      (pull_hi % 4))
    return DCGAN_ERR_SHAPE;
  PeerTable t{};
  fill(t.buf_a, grad_ptrs, world);
  fill(t.buf_b, bf16_ptrs, world);
  if (master_ptrs != nullptr) fill(t.buf_c, master_ptrs, world);
  else if (f32_from < hi) return DCGAN_ERR_SHAPE;
  fill(reinterpret_cast<void**>(t.sig), sig_ptrs, world);
  int* err = error_flag_ptr();
  if (!err) return DCGAN_ERR_DRIVER;
  if (blocks <= 0) blocks = 32;
  rs_adam_ag_kernel<<<blocks, kCommThreads, 0, st>>>(t, master, m, v, counters, state, rank, world, lo, hi, lr, b1,
                                                     b2, eps, slot, err, mc_grad,
                                                     reinterpret_cast<__nv_bfloat16*>(mc_bf16), mc_master, f32_from,
                                                     pull_lo, pull_hi);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_comm_lin_fwd_gather(const void* const* bf16_ptrs, const float* z, const float* bias, void* out, int NB,
                              int K, int F, size_t w_off, size_t lo, size_t hi, int rank, int world,
                              cudaStream_t st) {
  if (world < 1 || world > kMaxRanks) return DCGAN_ERR_SHAPE;
  PeerTable t{};
  fill(t.buf_b, bf16_ptrs, world);
  dim3 grid((F + 127) / 128, (NB + kGatherLinTile - 1) / kGatherLinTile);
  if (F % 128 || (w_off % 8)) return DCGAN_ERR_SHAPE;
  lin_fwd_gather_kernel<<<grid, 128, kGatherLinTile * K * sizeof(float) + (size_t)K * 256, st>>>(
      t, z, bias, reinterpret_cast<__nv_bfloat16*>(out), NB, K, F, w_off, lo, hi, rank, world);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_comm_allreduce_small(const void* const* staging_ptrs, const void* const* sig_ptrs, float* data, int n,
                               int staging_stride, unsigned* state, int rank, int world, int slot,
                               cudaStream_t st) {
  if (world < 1 || world > kMaxRanks || n > staging_stride || (n % 4) || n > 8192) return DCGAN_ERR_SHAPE;
  PeerTable t{};
  fill(t.buf_a, staging_ptrs, world);
  fill(reinterpret_cast<void**>(t.sig), sig_ptrs, world);
  int* err = error_flag_ptr();
  if (!err) return DCGAN_ERR_DRIVER;
  allreduce_small_kernel<<<1, kCommThreads, 0, st>>>(t, data, n, staging_stride, state, rank, world, slot, err);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_comm_async_apply(const void* const* master_ptrs, const void* const* bf16_ptrs, const void* const* m_ptrs,
                           const void* const* v_ptrs, const float* grad, void* local_bf16, float* local_master,
                           long long* const* peer_counters_dev, int rank, int world, size_t n, size_t lo, size_t hi,
                           int advance, float lr, float b1, float b2, float eps, int blocks, cudaStream_t st) {
  if (world < 1 || world > kMaxRanks || (n % 4) || (lo % 4) || (hi % 4) || lo > hi || hi > n) return DCGAN_ERR_SHAPE;
  PeerTable t{};
  fill(t.buf_a, master_ptrs, world);
  fill(t.buf_b, bf16_ptrs, world);
  fill(t.buf_c, m_ptrs, world);
  fill(t.buf_d, v_ptrs, world);
  if (blocks <= 0) blocks = 148;
  async_apply_kernel<<<blocks, kCommThreads, 0, st>>>(t, grad, reinterpret_cast<__nv_bfloat16*>(local_bf16),
                                                      local_master, peer_counters_dev, rank, world, n, lo, hi,
                                                      advance, lr, b1, b2, eps);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

}  // extern "C"
