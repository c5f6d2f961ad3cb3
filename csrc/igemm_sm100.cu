// Implicit-GEMM convolution kernels for sm_100a: tcgen05.mma with TMEM accumulators, the dense
// operand staged by TMA, the strided / parity-decomposed im2col operand gathered straight into the
// 128B-swizzled UMMA smem layout by the producer warps.
//
// One kernel template, three modes (geometry is TensorFlow's 5x5 / stride 2 / SAME, i.e. one pixel
// of padding before and two after -- SURVEY.md 2.4a/b):
//
//   MODE_F  "forward-strided":  out[n,oh,ow,:] = sum_{kh,kw} x[n,2oh+kh-1,2ow+kw-1,:] . W[kh,kw]
//           = tf.nn.conv2d(.., strides 2, SAME)           (reference distriubted_model.py:183)
//           = data-gradient of conv2d_transpose            (backward of distriubted_model.py:200)
//   MODE_T  "transposed":       out[n,2q+p,:] = sum_{taps of parity p} x[n,q+(p+1-k)/2,:] . W[kh,kw]
//           = tf.nn.conv2d_transpose(.., strides 2)        (reference distriubted_model.py:200)
//           = data-gradient of conv2d                      (backward of distriubted_model.py:183)
//           blockIdx.z enumerates the 4 output parity classes (3x3, 3x2, 2x3, 2x2 taps).
//   MODE_W  "weight-gradient":  dW[tap][cg][cd] += sum_pixels G[n,2q+k-1,cg] * D[n,q,cd]
//           (autodiff of both layer kinds, image_train.py:109-112); K = pixels, both operands
//           MN-major, split-K over pixel ranges with f32 red.global.add.
//
// GEMM view (BLOCK_M = 128, BLOCK_K = 64 bf16 = one 128-byte swizzle row):
//   F/T : M = pixels (gathered A, K-major), N = out channels (TMA B, K-major), K = taps*Cin
//   W   : M = dense channels (TMA A, MN-major), N = gathered channels (B, MN-major), K = pixels
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer (one lane),
// warps 2-5 = im2col gather producers, then the epilogue (one TMEM lane quadrant each).
// The gather producers are bound by instruction issue, so per-row constants (element offset of the unshifted
// tap, a 5x5 bit mask of which shifted taps exist) are computed once and a k-block costs one uniform offset,
// one bit test and one cp.async (ignore-src form = zero fill for padding) per row.
#include <mutex>
#include <unordered_map>
#include <vector>
#include <cstdio>
#include <cstring>

#include "common.cuh"
#include "igemm2.h"
#include "../include/dcgan_b200.h"

namespace dcgan {

struct IgemmParams {
  const __nv_bfloat16* g;  // gathered tensor, NHWC bf16
  int gH, gW, gC;          // gathered tensor spatial dims / channels
  int rH_log2, rW_log2;    // row (pixel) space spatial dims, powers of two
  int rows_total;          // NB * rH * rW
  int Cn;                  // F/T: output channels.  W: gathered channels (GEMM N extent)
  int Cd;                  // W: dense channels (GEMM M extent)
  int num_kb;              // k-blocks per CTA (W: per split)
  // epilogue
  void* out;               // F/T: bf16 NHWC.  W: f32 [25][Cn][Cd]
  const float* bias;       // F/T optional, per out channel
  int act;                 // F/T activation
  int oH, oW;              // F/T output spatial dims
  float* stats;            // F/T optional: [groups][2][Cn] += per-channel (sum, sum of squares) of the
                           // pre-activation output -- the batch-norm moments (distriubted_model.py:37)
  int stat_group_rows;     // rows per statistics group (multiple of 128)
  int b_mn;                // F/T: weights are [25][Ck][Cn] (Cn contiguous) -> MN-major B operand
  int t_long_first;        // T: schedule the parity classes with the longest K loop first
  int ksplit;              // F/T: CTAs of one thread-block cluster that share an output tile, each taking a
                           // slice of K; partial accumulators meet in the leader's shared memory (DSMEM)
  int* err_flag;           // sticky device error flag (csrc/runtime.cu)
  const int* mn_variant;   // MN-major descriptor convention switch (bring-up harness only)
};

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;
constexpr int kATileBytes = kBlockM * kBlockK * 2;  // 16 KiB
constexpr int kNumGatherThreads = 128;
constexpr int kNumThreads = 192;

template <int BLOCK_N, int STAGES>
struct SmemLayout {
  static constexpr int kBTileBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kStageBytes = kATileBytes + kBTileBytes;
  static constexpr int kBarOffset = STAGES * kStageBytes;
  static constexpr int kStatOffset = kBarOffset + 256;     // [4 warps][2][BLOCK_N] f32 epilogue column sums
  static constexpr int kTotal = kStatOffset + 8 * BLOCK_N * 4 + 1024;  // + alignment slack
};

// 16-byte chunk (row, chunk) of a [rows][128 B] tile in the SWIZZLE_128B pattern
__device__ __forceinline__ uint32_t swz128(uint32_t row, uint32_t chunk) {
  return row * 128u + ((chunk ^ (row & 7u)) << 4);
}

// 16-byte asynchronous global->shared copy (L2 only); src_bytes = 0 zero-fills (padding rows)
__device__ __forceinline__ void cp_async_16(uint32_t smem_dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_dst), "l"(src), "r"(src_bytes)
               : "memory");
}
// same copy with the "ignore-src" predicate form: when `valid` is zero nothing is read and 16 zero bytes are written
__device__ __forceinline__ void cp_async_16_or_zero(uint32_t smem_dst, const void* src, uint32_t valid) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %2, 0;\n\tcp.async.cg.shared.global [%0], [%1], 16, p;\n\t}" ::"r"(smem_dst),
      "l"(src), "r"(valid)
      : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// thread-block cluster helpers (split-K over distributed shared memory)
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d)
               : "memory");
}
// every thread of every CTA in the cluster takes part (non-.aligned forms: callers may be diverged)
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}

// k-blocks whose gather copies are in flight per producer thread before the oldest is published
// (kGatherDepth is min(3, STAGES - 1), a kernel-local constant)

template <int MODE, int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(kNumThreads, (SmemLayout<BLOCK_N, STAGES>::kTotal <= 74 * 1024)    ? 3
                                              : (SmemLayout<BLOCK_N, STAGES>::kTotal <= 110 * 1024) ? 2
                                                                                                    : 1)
igemm_kernel(const __grid_constant__ CUtensorMap tmap, const IgemmParams p) {
  pdl_launch_dependents();
  using L = SmemLayout<BLOCK_N, STAGES>;
  constexpr int kGatherDepth = (STAGES >= 8) ? 6 : ((STAGES - 1 < 3) ? STAGES - 1 : 3);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  float* stat_smem = reinterpret_cast<float*>(smem + L::kStatOffset);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;

  // ---- tile coordinates -----------------------------------------------------------------------
  int m0, n0, tap_fixed = 0, ph = 0, pw = 0, kb_base = 0;
  if (MODE == MODE_W) {
    tap_fixed = blockIdx.x;
    const int n_tiles = p.Cn / BLOCK_N;
    m0 = (blockIdx.y / n_tiles) * kBlockM;   // dense channel tile
    n0 = (blockIdx.y % n_tiles) * BLOCK_N;   // gathered channel tile
    kb_base = blockIdx.z * p.num_kb;
  } else {
    m0 = (blockIdx.x / p.ksplit) * kBlockM;  // pixel tile (the ksplit CTAs of a cluster are consecutive in x)
    n0 = blockIdx.y * BLOCK_N;               // out channel tile
    // longest K loops first (z = 0 is the 3x3-tap class, z = 3 the 2x2 one) so the last wave is made of short CTAs
    if (MODE == MODE_T) {
      ph = blockIdx.z >> 1; pw = blockIdx.z & 1;
      if (p.t_long_first) { ph = 1 - ph; pw = 1 - pw; }
    }
  }
  const int ksplit = (MODE == MODE_W) ? 1 : p.ksplit;
  const int krank = (ksplit > 1) ? (int)cluster_ctarank() : 0;
  int num_kb = (MODE == MODE_T) ? ((ph ? 3 : 2) * (pw ? 3 : 2) * (p.gC / kBlockK)) : p.num_kb;
  if (MODE != MODE_W && ksplit > 1) {  // this CTA's slice of K (the host guarantees num_kb >= ksplit)
    kb_base = (krank * num_kb) / ksplit;
    num_kb = ((krank + 1) * num_kb) / ksplit - kb_base;
  }

  // ---- one-time setup -------------------------------------------------------------------------
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], kNumGatherThreads + 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    *abort_flag = 0;
    fence_mbar_init();
    tma_prefetch_desc(&tmap);
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_smem, BLOCK_N);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_wait();  // everything above overlapped the previous kernel's tail; global memory is touched from here on

  if (warp == 0) {
    // =================================== TMA producer ==========================================
    if (lane == 0) {
      const int cpt = p.gC / kBlockK;  // F/T: channel chunks per tap
      const int nkw = pw ? 3 : 2;
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        const uint32_t phase = (kb / STAGES) & 1;
        if (!mbar_wait(&empty_bar[s], phase ^ 1, abort_flag, p.err_flag)) break;
        uint8_t* sA = smem + s * L::kStageBytes;
        uint8_t* sB = sA + kATileBytes;
        if (MODE == MODE_W) {
          mbar_arrive_expect_tx(&full_bar[s], kATileBytes);
          const int px0 = (kb_base + kb) * kBlockK;
          tma_load_2d(sA, &tmap, &full_bar[s], m0, px0);
          tma_load_2d(sA + kATileBytes / 2, &tmap, &full_bar[s], m0 + 64, px0);
        } else {
          int tap;
          const int lt = (kb_base + kb) / cpt;
          const int c0 = (kb_base + kb - lt * cpt) * kBlockK;
          if (MODE == MODE_F) {
            tap = lt;
          } else {
            const int ikh = lt / nkw, ikw = lt - ikh * nkw;
            const int kh = ph ? 2 * ikh : 2 * ikh + 1;
            const int kw = pw ? 2 * ikw : 2 * ikw + 1;
            tap = kh * 5 + kw;
          }
          mbar_arrive_expect_tx(&full_bar[s], L::kBTileBytes);
          if (p.b_mn) {  // native [tap][Ck][Cn] layout: 64(k) x 64(n) boxes, one per 64 output channels
#pragma unroll
            for (int j = 0; j < BLOCK_N / 64; ++j)
              tma_load_2d(sB + j * 8192, &tmap, &full_bar[s], n0 + 64 * j, tap * p.gC + c0);
          } else {
            tma_load_2d(sB, &tmap, &full_bar[s], c0, tap * p.Cn + n0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // =================================== MMA issuer ============================================
    if (lane == 0) {
      constexpr bool kMN = (MODE == MODE_W);
      const bool b_mn = kMN || (p.b_mn != 0);
      const uint32_t idesc = make_idesc_bf16(kBlockM, BLOCK_N, kMN, b_mn);
      const bool swap = kMN && (*p.mn_variant != 0);
      const uint32_t lbo = kMN ? (swap ? 1024u : 8192u) : 16u;
      const uint32_t sbo = kMN ? (swap ? 8192u : 1024u) : 1024u;
      const uint32_t kstep = kMN ? 2048u : 32u;  // bytes per UMMA_K=16 step
      const uint32_t lbo_b = b_mn ? (swap ? 1024u : 8192u) : 16u;
      const uint32_t sbo_b = b_mn ? (swap ? 8192u : 1024u) : 1024u;
      const uint32_t kstep_b = b_mn ? 2048u : 32u;
      bool ok = true;
      for (int kb = 0; kb < num_kb && ok; ++kb) {
        const int s = kb % STAGES;
        const uint32_t phase = (kb / STAGES) & 1;
        ok = mbar_wait(&full_bar[s], phase, abort_flag, p.err_flag);
        if (!ok) break;
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * L::kStageBytes);
        const uint32_t b_addr = a_addr + kATileBytes;
#pragma unroll
        for (int k = 0; k < kBlockK / 16; ++k) {
          const uint64_t da = make_smem_desc_sw128(a_addr + k * kstep, lbo, sbo);
          const uint64_t db = make_smem_desc_sw128(b_addr + k * kstep_b, lbo_b, sbo_b);
          umma_bf16(tmem_base, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);  // frees the smem slot once these MMAs retire
      }
      umma_commit(tmem_full_bar);  // accumulator complete
    }
  }
  if (MODE != MODE_W && ksplit > 1 && warp < 2) {  // the two cluster barriers of the split-K reduction (below)
    __syncwarp();
    cluster_sync_all();
    cluster_sync_all();
  }
  if (warp >= 2) {
    // ============================ gather producers, then epilogue ==============================
    const int gt = tid - 64;       // 0..127
    const int chunk = gt & 7;      // 16-byte chunk inside the 128-byte row
    const int rbase = gt >> 3;     // 0..15
    const int rW = 1 << p.rW_log2, rH = 1 << p.rH_log2;
    const int stride = (MODE == MODE_T) ? 1 : 2;
    bool ok = true;

    if (MODE != MODE_W) {
      // The producer warps are bound by instruction issue, not by L2 (ncu: ~350 warp instructions per warp and
      // k-block, IPC 1.8 with 12 resident warps), so everything that does not change from one k-block to the
      // next is hoisted: per row the element offset of the unshifted tap and a validity bit per (dy, dx) shift,
      // per k-block one uniform offset and one bit index, advanced without divisions
      // (tests/test_gather_math_cpu.py emulates exactly this arithmetic against the per-k-block formula).
      int base_off[8];      // ((n*gH + y0)*gW + x0)*gC + chunk*8, element units (host: numel < 2^31)
      uint32_t vmask[8];    // bit (dy+1)*5 + (dx+1): the shifted pixel exists; 0 for rows past the end
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int R = m0 + rbase + 16 * i;
        base_off[i] = 0;
        vmask[i] = 0;
        if (R < p.rows_total) {
          const int qw = R & (rW - 1);
          const int qh = (R >> p.rW_log2) & (rH - 1);
          const int n = R >> (p.rW_log2 + p.rH_log2);
          const int y0 = stride * qh, x0 = stride * qw;
          base_off[i] = ((n * p.gH + y0) * p.gW + x0) * p.gC + chunk * 8;
          uint32_t my = 0, mx = 0;
#pragma unroll
          for (int sft = 0; sft < 5; ++sft) {
            if ((unsigned)(y0 + sft - 1) < (unsigned)p.gH) my |= 1u << sft;
            if ((unsigned)(x0 + sft - 1) < (unsigned)p.gW) mx |= 1u << sft;
          }
          uint32_t m = 0;
#pragma unroll
          for (int sy = 0; sy < 5; ++sy)
            if ((my >> sy) & 1u) m |= mx << (5 * sy);
          vmask[i] = m;
        }
      }
      const int cpt = p.gC / kBlockK;
      const int nb = (MODE == MODE_F) ? 5 : (pw ? 3 : 2);  // taps per filter row of this class
      // position of this CTA's first k-block (the only divisions); then (ta, tb, cc) advance like an odometer
      const int lt0 = kb_base / cpt;
      int cc = kb_base - lt0 * cpt;
      int ta = lt0 / nb, tb = lt0 - ta * nb;               // F: (kh, kw).  T: (ikh, ikw)
      int s_issue = 0, s_pub = 0;
      uint32_t ph_issue = 0;
      // cp.async pipeline: the copies of kGatherDepth k-blocks are in flight per thread; a k-block is
      // published (proxy fence + mbarrier arrive) once its group has landed.
      for (int it = 0; it < num_kb + kGatherDepth - 1; ++it) {
        if (it < num_kb) {
          const int dy = (MODE == MODE_F) ? ta - 1 : ph - ta;
          const int dx = (MODE == MODE_F) ? tb - 1 : pw - tb;
          const int delta = (dy * p.gW + dx) * p.gC + cc * kBlockK;
          const uint32_t bitmask = 1u << ((dy + 1) * 5 + (dx + 1));
          if (!mbar_wait(&empty_bar[s_issue], ph_issue ^ 1, abort_flag, p.err_flag)) { ok = false; break; }
          const uint32_t sA = smem_u32(smem + s_issue * L::kStageBytes);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t valid = vmask[i] & bitmask;
            const uint32_t off = valid ? (uint32_t)(base_off[i] + delta) : 0u;  // a valid offset is never negative
            cp_async_16_or_zero(sA + swz128(rbase + 16 * i, chunk), p.g + off, valid);
          }
          if (++cc == cpt) { cc = 0; if (++tb == nb) { tb = 0; ++ta; } }
          if (++s_issue == STAGES) { s_issue = 0; ph_issue ^= 1; }
        }
        cp_async_commit();
        if (it >= kGatherDepth - 1) {
          cp_async_wait<kGatherDepth - 1>();
          fence_proxy_async_smem();
          mbar_arrive(&full_bar[s_pub]);
          if (++s_pub == STAGES) s_pub = 0;
        }
      }
    } else {
      // W mode: rows are pixels of the K dimension and change every k-block
      const int kh = tap_fixed / 5, kw = tap_fixed - kh * 5;
      constexpr int kSub = BLOCK_N / 64;  // 64-channel sub-tiles of the gathered operand
      for (int it = 0; it < num_kb + kGatherDepth - 1; ++it) {
        if (it < num_kb) {
          const int kb = it;
          const int s = kb % STAGES;
          const uint32_t phase = (kb / STAGES) & 1;
          const int px0 = (kb_base + kb) * kBlockK;
          if (!mbar_wait(&empty_bar[s], phase ^ 1, abort_flag, p.err_flag)) { ok = false; break; }
          const uint32_t sB = smem_u32(smem + s * L::kStageBytes + kATileBytes);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int R = px0 + rbase + 16 * i;
            const int qw = R & (rW - 1);
            const int qh = (R >> p.rW_log2) & (rH - 1);
            const int n = R >> (p.rW_log2 + p.rH_log2);
            const int y = 2 * qh + kh - 1, x = 2 * qw + kw - 1;
            const uint32_t valid =
                (R < p.rows_total && (unsigned)y < (unsigned)p.gH && (unsigned)x < (unsigned)p.gW) ? 1u : 0u;
            // 32-bit element offsets (host: numel < 2^31); an ignored copy still gets an in-bounds address
            const uint32_t off = valid ? (uint32_t)(((n * p.gH + y) * p.gW + x) * p.gC + n0 + chunk * 8) : 0u;
            const __nv_bfloat16* src = p.g + off;
#pragma unroll
            for (int j = 0; j < kSub; ++j)
              cp_async_16_or_zero(sB + j * 8192 + swz128(rbase + 16 * i, chunk), src + 64 * j, valid);
          }
        }
        cp_async_commit();
        if (it >= kGatherDepth - 1) {
          cp_async_wait<kGatherDepth - 1>();
          fence_proxy_async_smem();
          mbar_arrive(&full_bar[(it - (kGatherDepth - 1)) % STAGES]);
        }
      }
    }

    // ------------------------------------ epilogue ----------------------------------------------
    const bool do_stats = (MODE != MODE_W) && p.stats != nullptr;
    if (ok) ok = mbar_wait(tmem_full_bar, 0, abort_flag, p.err_flag);
    tc_fence_after();
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    const int row_in_tile = q * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    // Split-K across the cluster: once every CTA's ring is idle (barrier 1) the non-leaders store their f32
    // partial tiles into the leader's ring through distributed shared memory, [col/4][row][4] so that a
    // warp's 16-byte stores are contiguous; after barrier 2 the leader adds them in its epilogue.
    constexpr int kPartialBytes = kBlockM * BLOCK_N * 4;
    bool leader = true;
    if (MODE != MODE_W && ksplit > 1) {
      cluster_sync_all();
      if (krank != 0) {
        leader = false;
        const uint32_t dst = mapa_shared(smem_u32(smem) + (uint32_t)(krank - 1) * kPartialBytes, 0) +
                             (uint32_t)row_in_tile * 16u;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_cluster_v4(dst + (uint32_t)((c >> 2) + j) * (kBlockM * 16u), v[4 * j], v[4 * j + 1], v[4 * j + 2],
                          v[4 * j + 3]);
        }
      }
      cluster_sync_all();
    }
    if (MODE != MODE_W && leader) {
      const int R = m0 + row_in_tile;
      const bool valid = ok && R < p.rows_total;
      __nv_bfloat16* orow = nullptr;
      if (valid) {
        const int qw = R & (rW - 1);
        const int qh = (R >> p.rW_log2) & (rH - 1);
        const int n = R >> (p.rW_log2 + p.rH_log2);
        int oh, ow;
        if (MODE == MODE_F) { oh = qh; ow = qw; } else { oh = 2 * qh + ph; ow = 2 * qw + pw; }
        orow = reinterpret_cast<__nv_bfloat16*>(p.out) + ((size_t)((n * p.oH + oh) * p.oW + ow)) * p.Cn + n0;
      }
#pragma unroll 1
      for (int c = 0; c < BLOCK_N; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + c, v);
        tmem_ld_wait();
        for (int r = 1; r < ksplit; ++r) {
          const float4* part = reinterpret_cast<const float4*>(smem + (size_t)(r - 1) * kPartialBytes) +
                               (size_t)(c >> 2) * kBlockM + row_in_tile;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 t = part[(size_t)j * kBlockM];
            v[4 * j] = __float_as_uint(__uint_as_float(v[4 * j]) + t.x);
            v[4 * j + 1] = __float_as_uint(__uint_as_float(v[4 * j + 1]) + t.y);
            v[4 * j + 2] = __float_as_uint(__uint_as_float(v[4 * j + 2]) + t.z);
            v[4 * j + 3] = __float_as_uint(__uint_as_float(v[4 * j + 3]) + t.w);
          }
        }
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float a = __uint_as_float(v[j]);
          if (p.bias != nullptr) a += __ldg(p.bias + n0 + c + j);
          f[j] = valid ? a : 0.f;
        }
        if (valid) {
          uint32_t packed[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            // (p.act is 0 whenever the moments below are wanted: the *_stats entry points have no activation)
            packed[j] = pack_bf16x2(apply_act(f[2 * j], p.act), apply_act(f[2 * j + 1], p.act));
          }
          uint4* dst = reinterpret_cast<uint4*>(orow + c);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            dst[j] = make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
        }
        if (do_stats) {  // batch-norm moments of this tile, reduced over the warp's 32 rows
          float sq[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) sq[j] = f[j] * f[j];
          const float s1 = warp_transpose_reduce32(f);
          const float s2 = warp_transpose_reduce32(sq);
          // one slice per warp: plain stores (f32 atomics on shared memory are compare-and-swap loops)
          stat_smem[(q * 2) * BLOCK_N + c + lane] = s1;
          stat_smem[(q * 2 + 1) * BLOCK_N + c + lane] = s2;
        }
      }
      if (do_stats) {
        asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only
        float* gs = p.stats + (size_t)(m0 / p.stat_group_rows) * 2 * p.Cn + n0;
        for (int i = gt; i < BLOCK_N; i += kNumGatherThreads) {
          float t1 = 0.f, t2 = 0.f;
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            t1 += stat_smem[(w * 2) * BLOCK_N + i];
            t2 += stat_smem[(w * 2 + 1) * BLOCK_N + i];
          }
          atomicAdd(gs + i, t1);
          atomicAdd(gs + p.Cn + i, t2);
        }
      }
    } else if (MODE == MODE_W) {
      float* obase = reinterpret_cast<float*>(p.out) + (size_t)tap_fixed * p.Cn * p.Cd + (m0 + row_in_tile);
#pragma unroll 1
      for (int c = 0; c < BLOCK_N; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + c, v);
        tmem_ld_wait();
        if (ok) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            atomicAdd(obase + (size_t)(n0 + c + j) * p.Cd, __uint_as_float(v[j]));
        }
      }
    }
  }

  // ---- teardown ---------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BLOCK_N);
}


// =================================================================================================
// Parity-fused transposed conv ("T4"): one CTA computes all four output parity classes of its 128
// input locations.  The nine shifted gathers (dy, dx in {-1,0,1}) are the only distinct A tiles -- the
// per-parity kernels above gather 25 -- and each feeds the 1, 2 or 4 parity accumulators whose filter
// tap lands on that shift.  A tiles and filter tiles travel in separate rings (one A tile serves up to
// four B tiles); the four f32 accumulators take 4 x BLOCK_N TMEM columns.
// Used for the large-M layers (many CTAs), where the gather traffic from L2 is the bound
// (profiles/README.md): d_h1/d_h2 data gradients, g_h3 forward.
// =================================================================================================
template <int BLOCK_N, int A_STAGES, int B_SLOTS>
struct SmemLayoutT4 {
  static constexpr int kBTileBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kBOffset = A_STAGES * kATileBytes;
  static constexpr int kBarOffset = kBOffset + B_SLOTS * kBTileBytes;
  static constexpr int kStatOffset = kBarOffset + 512;
  static constexpr int kTotal = kStatOffset + 2 * BLOCK_N * 4 + 1024;
};

template <int BLOCK_N, int A_STAGES, int B_SLOTS>
__global__ void __launch_bounds__(kNumThreads, (SmemLayoutT4<BLOCK_N, A_STAGES, B_SLOTS>::kTotal <= 110 * 1024 &&
                                                BLOCK_N <= 64) ? 2 : 1)
igemm_t4_kernel(const __grid_constant__ CUtensorMap tmap, const IgemmParams p) {
  pdl_launch_dependents();
  using L = SmemLayoutT4<BLOCK_N, A_STAGES, B_SLOTS>;
  constexpr int kGatherDepth = (A_STAGES - 1 < 3) ? A_STAGES - 1 : 3;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* a_empty = a_full + A_STAGES;
  uint64_t* b_full = a_empty + A_STAGES;
  uint64_t* b_empty = b_full + B_SLOTS;
  uint64_t* tmem_full_bar = b_empty + B_SLOTS;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  float* stat_smem = reinterpret_cast<float*>(smem + L::kStatOffset);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * kBlockM, n0 = blockIdx.y * BLOCK_N;
  const int cpt = p.gC / kBlockK;
  const int num_a = 9 * cpt;  // (shift, channel chunk) items

  if (tid == 0) {
    for (int s = 0; s < A_STAGES; ++s) { mbar_init(&a_full[s], kNumGatherThreads); mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < B_SLOTS; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    mbar_init(tmem_full_bar, 1);
    *abort_flag = 0;
    fence_mbar_init();
    tma_prefetch_desc(&tmap);
  }
  if (warp == 1) { tmem_alloc(tmem_ptr_smem, 4 * BLOCK_N); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_wait();  // everything above overlapped the previous kernel's tail; global memory is touched from here on

  if (warp == 0) {
    // ============================ TMA producer: filter tiles =======================================
    if (lane == 0) {
      int b = 0;
      bool ok = true;
      for (int a = 0; a < num_a && ok; ++a) {
        const int sft = a / cpt, c0 = (a - sft * cpt) * kBlockK;
        const int dy = sft / 3 - 1, dx = sft % 3 - 1;
        for (int ph = (dy == 1) ? 1 : 0; ph < 2 && ok; ++ph)
          for (int pw = (dx == 1) ? 1 : 0; pw < 2; ++pw) {
            const int tap = (ph + 1 - 2 * dy) * 5 + (pw + 1 - 2 * dx);
            const int slot = b % B_SLOTS;
            if (!mbar_wait(&b_empty[slot], ((b / B_SLOTS) & 1) ^ 1, abort_flag, p.err_flag)) { ok = false; break; }
            mbar_arrive_expect_tx(&b_full[slot], L::kBTileBytes);
            tma_load_2d(smem + L::kBOffset + slot * L::kBTileBytes, &tmap, &b_full[slot], c0, tap * p.Cn + n0);
            ++b;
          }
      }
    }
  } else if (warp == 1) {
    // ============================ MMA issuer ========================================================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kBlockM, BLOCK_N, 0, 0);
      int b = 0;
      uint32_t started = 0;  // bit p: parity accumulator p already holds a partial sum
      bool ok = true;
      for (int a = 0; a < num_a && ok; ++a) {
        const int st = a % A_STAGES;
        const int sft = a / cpt;
        const int dy = sft / 3 - 1, dx = sft % 3 - 1;
        ok = mbar_wait(&a_full[st], (a / A_STAGES) & 1, abort_flag, p.err_flag);
        if (!ok) break;
        const uint32_t a_addr = smem_u32(smem + st * kATileBytes);
        for (int ph = (dy == 1) ? 1 : 0; ph < 2 && ok; ++ph)
          for (int pw = (dx == 1) ? 1 : 0; pw < 2; ++pw) {
            const int slot = b % B_SLOTS;
            if (!mbar_wait(&b_full[slot], (b / B_SLOTS) & 1, abort_flag, p.err_flag)) { ok = false; break; }
            tc_fence_after();
            const int par = ph * 2 + pw;
            const uint32_t b_addr = smem_u32(smem + L::kBOffset + slot * L::kBTileBytes);
#pragma unroll
            for (int k = 0; k < kBlockK / 16; ++k) {
              const uint64_t da = make_smem_desc_sw128(a_addr + k * 32, 16, 1024);
              const uint64_t db = make_smem_desc_sw128(b_addr + k * 32, 16, 1024);
              umma_bf16(tmem_base + par * BLOCK_N, da, db, idesc, (((started >> par) & 1u) | (k != 0)) ? 1u : 0u);
            }
            started |= 1u << par;
            umma_commit(&b_empty[slot]);
            ++b;
          }
        umma_commit(&a_empty[st]);
      }
      umma_commit(tmem_full_bar);
    }
  } else {
    // ============================ gather producers, then epilogue ====================================
    const int gt = tid - 64, chunk = gt & 7, rbase = gt >> 3;
    const int rW = 1 << p.rW_log2, rH = 1 << p.rH_log2;
    bool ok = true;
    int pix_base[8], yx[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int R = m0 + rbase + 16 * i;
      if (R < p.rows_total) {
        const int qw = R & (rW - 1), qh = (R >> p.rW_log2) & (rH - 1), n = R >> (p.rW_log2 + p.rH_log2);
        pix_base[i] = n * p.gH * p.gW;
        yx[i] = (qh << 16) | qw;
      } else { pix_base[i] = -1; yx[i] = 0; }
    }
    for (int it = 0; it < num_a + kGatherDepth - 1; ++it) {
      if (it < num_a) {
        const int st = it % A_STAGES;
        const int sft = it / cpt, c0 = (it - sft * cpt) * kBlockK;
        const int dy = sft / 3 - 1, dx = sft % 3 - 1;
        if (!mbar_wait(&a_empty[st], ((it / A_STAGES) & 1) ^ 1, abort_flag, p.err_flag)) { ok = false; break; }
        const uint32_t sA = smem_u32(smem + st * kATileBytes);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int y = (yx[i] >> 16) + dy, x = (yx[i] & 0xffff) + dx;
          const bool valid = pix_base[i] >= 0 && (unsigned)y < (unsigned)p.gH && (unsigned)x < (unsigned)p.gW;
          const size_t off = valid ? ((size_t)(pix_base[i] + y * p.gW + x)) * p.gC + c0 + chunk * 8 : 0;
          cp_async_16(sA + swz128(rbase + 16 * i, chunk), p.g + off, valid ? 16u : 0u);
        }
      }
      cp_async_commit();
      if (it >= kGatherDepth - 1) {
        cp_async_wait<kGatherDepth - 1>();
        fence_proxy_async_smem();
        mbar_arrive(&a_full[(it - (kGatherDepth - 1)) % A_STAGES]);
      }
    }
    // ------------------------------------ epilogue: four parity accumulators ---------------------------
    const bool do_stats = p.stats != nullptr;
    if (do_stats) {
      for (int i = gt; i < 2 * BLOCK_N; i += kNumGatherThreads) stat_smem[i] = 0.f;
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
    if (ok) ok = mbar_wait(tmem_full_bar, 0, abort_flag, p.err_flag);
    tc_fence_after();
    const int q = warp & 3;
    const int R = m0 + q * 32 + lane;
    const bool valid = ok && R < p.rows_total;
    const int qw = R & (rW - 1), qh = (R >> p.rW_log2) & (rH - 1), n = R >> (p.rW_log2 + p.rH_log2);
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
    for (int par = 0; par < 4; ++par) {
      const int oh = 2 * qh + (par >> 1), ow = 2 * qw + (par & 1);
      __nv_bfloat16* orow =
          reinterpret_cast<__nv_bfloat16*>(p.out) + ((size_t)((n * p.oH + oh) * p.oW + ow)) * p.Cn + n0;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + par * BLOCK_N + c, v);
        tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float a = __uint_as_float(v[j]);
          if (p.bias != nullptr) a += __ldg(p.bias + n0 + c + j);
          f[j] = valid ? a : 0.f;
        }
        if (valid) {
          uint32_t packed[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            // (p.act is 0 whenever the moments below are wanted: the *_stats entry points have no activation)
            packed[j] = pack_bf16x2(apply_act(f[2 * j], p.act), apply_act(f[2 * j + 1], p.act));
          }
          uint4* dst = reinterpret_cast<uint4*>(orow + c);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            dst[j] = make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
        }
        if (do_stats) {
          float sq[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) sq[j] = f[j] * f[j];
          const float s1 = warp_transpose_reduce32(f);
          const float s2 = warp_transpose_reduce32(sq);
          atomicAdd(&stat_smem[c + lane], s1);
          atomicAdd(&stat_smem[BLOCK_N + c + lane], s2);
        }
      }
    }
    if (do_stats) {
      asm volatile("bar.sync 1, 128;" ::: "memory");
      float* gs = p.stats + (size_t)(m0 / p.stat_group_rows) * 2 * p.Cn + n0;
      for (int i = gt; i < BLOCK_N; i += kNumGatherThreads) {
        atomicAdd(gs + i, stat_smem[i]);
        atomicAdd(gs + p.Cn + i, stat_smem[BLOCK_N + i]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 4 * BLOCK_N);
}

// =================================================================================================
// host side
// =================================================================================================
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

struct MapKey {
  const void* ptr;
  uint64_t cols, rows;
  uint32_t box_c, box_r;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && cols == o.cols && rows == o.rows && box_c == o.box_c && box_r == o.box_r;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    size_t h = std::hash<const void*>()(k.ptr);
    h = h * 1000003u ^ std::hash<uint64_t>()(k.cols * 1315423911u + k.rows);
    h = h * 1000003u ^ (k.box_c * 257u + k.box_r);
    return h;
  }
};

// 2-D row-major bf16 tensor [rows][cols], box [box_r][box_c], SWIZZLE_128B.  Cached per
// (pointer, shape): encoding costs a driver call and the buffers of the engine are static.
static int get_tmap_2d(const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_r, uint32_t box_c,
                       CUtensorMap* out) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  MapKey key{ptr, cols, rows, box_c, box_r};
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return 0; }
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return DCGAN_ERR_DRIVER;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {box_c, box_r};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return DCGAN_ERR_DRIVER;
  cache.emplace(key, m);
  *out = m;
  return 0;
}

static int ilog2_exact(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return ((1 << l) == v) ? l : -1;
}

template <int MODE, int BLOCK_N, int STAGES>
static int launch(const CUtensorMap& tmap, const IgemmParams& p, dim3 grid, cudaStream_t stream) {
  using L = SmemLayout<BLOCK_N, STAGES>;
  auto kern = igemm_kernel<MODE, BLOCK_N, STAGES>;
  static bool attr_done[64] = {false};  // per instantiation and device
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DCGAN_ERR_DRIVER;
  if (!attr_done[dev]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal) != cudaSuccess)
      return DCGAN_ERR_LAUNCH;
    attr_done[dev] = true;
  }
  if (MODE != MODE_W && p.ksplit > 1) {
    // split-K: the ksplit CTAs that share an output tile form one thread-block cluster along x
    if ((size_t)(p.ksplit - 1) * kBlockM * BLOCK_N * 4 > (size_t)STAGES * L::kStageBytes) return DCGAN_ERR_SHAPE;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = L::kTotal;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = p.ksplit;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, tmap, p) == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
  }
  launch_k(kern, dim3(grid), dim3(kNumThreads), L::kTotal, stream, tmap, p);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

// tuning knob (bring-up / profiling): bit 0 = two CTAs per SM with shallower stage rings
static int g_igemm_tuning = 1;  // bit 7 (128): first-generation kernels of this file (cp.async gather) instead of
                                // the TMA-fed persistent kernels of igemm2_sm100.cu; bit 3 (8): parity-fused transposed conv for large-M layers; bit 4: its N=128 tile
                                // bit 5 (32): no cluster split-K; bit 6 (64): transposed conv parity classes in
                                // natural order instead of longest-first
static int g_igemm_ksplit = 0;  // 0 = choose per layer; 1, 2, 4 = forced (tests / profiling)
static int g_wgrad_cta_target = 148;  // filter-gradient launches split K until they have this many CTAs

// Split-K factor for an F/T launch of `ctas` tiles whose shortest K loop has `min_kb` k-blocks: small-M
// layers leave most of the 148 SMs idle while each CTA walks a long serial K loop, so K is cut across a
// cluster until the resident-CTA slots are filled.  The partial tiles must fit in the leader's stage ring.
// Measured (profiles/igemm_splitk_r1.log): splitting pays while every CTA keeps >= `min_kb_per_cta` k-blocks
// (16 for the strided mode, 8 for the shortest parity class of the transposed mode).
static int pick_ksplit(long ctas, int min_kb, int min_kb_per_cta, int block_n) {
  if (g_igemm_ksplit > 0) return (min_kb >= g_igemm_ksplit) ? g_igemm_ksplit : 1;
  if (block_n > 128) return 1;
  if (g_igemm_tuning & (32 | 4 | 2)) return 1;  // off, or a non-default stage ring is being profiled
  const bool occ2 = (g_igemm_tuning & 1) != 0 && block_n <= 128;
  const long slots = 148L * (occ2 ? 2 : 1);
  const int stages = (block_n == 64) ? (occ2 ? 4 : 6) : (block_n == 128 ? (occ2 ? 3 : 5) : 4);
  const long ring = (long)stages * (kATileBytes + block_n * kBlockK * 2);
  int ks = 1;
  while (ks < 4 && ctas * ks * 2 <= slots && min_kb / (ks * 2) >= min_kb_per_cta &&
         (long)(ks * 2 - 1) * kBlockM * block_n * 4 <= ring)
    ks *= 2;
  return ks;
}

template <int MODE>
static int dispatch_n(int block_n, const CUtensorMap& tmap, const IgemmParams& p, dim3 grid, cudaStream_t st) {
  const bool occ2 = (g_igemm_tuning & 1) != 0;
  switch (block_n) {
    case 64:
      if (g_igemm_tuning & 4) return launch<MODE, 64, 8>(tmap, p, grid, st);  // deep ring, 1 CTA/SM
      if (g_igemm_tuning & 2) return launch<MODE, 64, 3>(tmap, p, grid, st);  // 3 CTAs/SM
      return occ2 ? launch<MODE, 64, 4>(tmap, p, grid, st) : launch<MODE, 64, 6>(tmap, p, grid, st);
    case 128: return occ2 ? launch<MODE, 128, 3>(tmap, p, grid, st) : launch<MODE, 128, 5>(tmap, p, grid, st);
    case 256: return launch<MODE, 256, 4>(tmap, p, grid, st);
    default: return DCGAN_ERR_SHAPE;
  }
}

template <int BLOCK_N, int A_STAGES, int B_SLOTS>
static int launch_t4(const CUtensorMap& tmap, const IgemmParams& p, dim3 grid, cudaStream_t stream) {
  using L = SmemLayoutT4<BLOCK_N, A_STAGES, B_SLOTS>;
  auto kern = igemm_t4_kernel<BLOCK_N, A_STAGES, B_SLOTS>;
  static bool attr_done[64] = {false};  // per device
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DCGAN_ERR_DRIVER;
  if (!attr_done[dev]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal) != cudaSuccess)
      return DCGAN_ERR_LAUNCH;
    attr_done[dev] = true;
  }
  launch_k(kern, dim3(grid), dim3(kNumThreads), L::kTotal, stream, tmap, p);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

static int pick_block_n(int Cn, int rows_tiles, int extra_parallel, int min_kb, int min_kb_per_cta) {
  // the kernels are bound by L2 traffic of the im2col gather, so prefer the widest tile (fewest re-reads
  // of the gathered operand) that, together with its split-K factor, still occupies most of the 148 SMs
  // (measured: profiles/igemm_*.log)
  const int cands[3] = {256, 128, 64};
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (Cn % bn) continue;
    const long ctas = (long)rows_tiles * (Cn / bn) * extra_parallel;
    if (ctas * pick_ksplit(ctas, min_kb, min_kb_per_cta, bn) >= 90 || bn == 64) return bn;
  }
  return (Cn % 64 == 0) ? 64 : -1;
}

}  // namespace dcgan

using namespace dcgan;

extern "C" {

void dcgan_set_igemm_tuning(int v) { g_igemm_tuning = v; }
void dcgan_set_igemm_ksplit(int v) {
  g_igemm_ksplit = (v == 1 || v == 2 || v == 4) ? v : 0;
  igemm2_set_ksplit(v);
}
void dcgan_set_igemm_max_ctas(int v) { igemm2_set_max_ctas(v); }
void dcgan_set_igemm_stage_mode(int v) { igemm2_set_stage_mode(v); }
void dcgan_set_igemm_tile(int v) { igemm2_set_tile(v); }
int dcgan_get_igemm_ksplit(void) { return igemm2_forced_ksplit(); }
int dcgan_igemm_plan_set(int kind, int NB, int H, int W, int Cin, int Cout, int block_n, int ksplit, int variant) {
  return igemm2_plan_set(kind, NB, H, W, Cin, Cout, block_n, ksplit, variant);
}
void dcgan_set_igemm_debug(long long* ptr) { igemm2_set_debug(ptr); }
int dcgan_set_workspace(void* ptr, size_t bytes) { return igemm2_set_workspace(ptr, bytes); }
static inline bool use_v2() { return (g_igemm_tuning & 128) == 0; }
void dcgan_set_wgrad_cta_target(int v) { g_wgrad_cta_target = v > 0 ? v : 148; }

// Host-only view of the launch plan of a strided (transposed = 0) or transposed (1) conv: tile width, split-K
// factor, grid and whether the partial tiles fit the leader's stage ring.  No CUDA call; used by the CPU tests.
int dcgan_debug_plan_conv(int transposed, int NB, int H, int W, int Cin, int Cout, int* block_n, int* ksplit,
                          int* grid_xyz, int* partials_fit) {
  if (Cin % 64 || Cout % 64) return DCGAN_ERR_SHAPE;
  const int rows = transposed ? NB * H * W : NB * (H / 2) * (W / 2);
  const int mt = (rows + kBlockM - 1) / kBlockM;
  const int extra = transposed ? 4 : 1;
  const int min_kb = transposed ? 4 * (Cin / 64) : 25 * (Cin / 64);
  const int per_cta = transposed ? 8 : 16;
  const int bn = pick_block_n(Cout, mt, extra, min_kb, per_cta);
  if (bn < 0) return DCGAN_ERR_SHAPE;
  const int ks = pick_ksplit((long)mt * (Cout / bn) * extra, min_kb, per_cta, bn);
  const bool occ2 = (g_igemm_tuning & 1) != 0 && bn <= 128;
  const int stages = (bn == 64) ? (occ2 ? 4 : 6) : (bn == 128 ? (occ2 ? 3 : 5) : 4);
  *block_n = bn;
  *ksplit = ks;
  grid_xyz[0] = mt * ks; grid_xyz[1] = Cout / bn; grid_xyz[2] = extra;
  *partials_fit = (long)(ks - 1) * kBlockM * bn * 4 <= (long)stages * (kATileBytes + bn * kBlockK * 2);
  return 0;
}

static int conv_fprop_impl(const void* x, const void* w_nk, const float* bias, void* out, int NB, int H, int W,
                           int Cin, int Cout, int act, int block_n, int b_mn, cudaStream_t stream,
                           float* stats = nullptr, int stat_groups = 1) {
  if (use_v2()) {
    return igemm2_conv(0, x, w_nk, bias, out, NB, H, W, Cin, Cout, act, block_n, b_mn, stats, stat_groups, stream);
  }
  if ((H & 1) || (W & 1) || Cin % 64 || Cout % 64) return DCGAN_ERR_SHAPE;
  if ((long)NB * H * W * Cin >= (1L << 31)) return DCGAN_ERR_SHAPE;  // the gather uses 32-bit element offsets
  const int oH = H / 2, oW = W / 2;
  const int lh = ilog2_exact(oH), lw = ilog2_exact(oW);
  if (lh < 0 || lw < 0) return DCGAN_ERR_SHAPE;
  IgemmParams p{};
  p.err_flag = error_flag_ptr();
  p.mn_variant = mn_variant_ptr();
  if (!p.err_flag || !p.mn_variant) return DCGAN_ERR_DRIVER;
  p.g = reinterpret_cast<const __nv_bfloat16*>(x);
  p.gH = H; p.gW = W; p.gC = Cin;
  p.rH_log2 = lh; p.rW_log2 = lw;
  p.rows_total = NB * oH * oW;
  p.Cn = Cout; p.Cd = 0;
  p.num_kb = 25 * (Cin / 64);
  p.b_mn = b_mn;
  if (stats != nullptr) {
    if (stat_groups < 1 || p.rows_total % stat_groups || (p.rows_total / stat_groups) % kBlockM) return DCGAN_ERR_SHAPE;
    p.stats = stats;
    p.stat_group_rows = p.rows_total / stat_groups;
  }
  p.out = out; p.bias = bias; p.act = act; p.oH = oH; p.oW = oW;
  const int mt = (p.rows_total + kBlockM - 1) / kBlockM;
  if (block_n <= 0) block_n = pick_block_n(Cout, mt, 1, p.num_kb, 16);
  if (block_n < 0 || Cout % block_n) return DCGAN_ERR_SHAPE;
  CUtensorMap tmap;
  int rc = b_mn ? get_tmap_2d(w_nk, (uint64_t)25 * Cin, Cout, 64, 64, &tmap)
                : get_tmap_2d(w_nk, (uint64_t)25 * Cout, Cin, block_n, 64, &tmap);
  if (rc) return rc;
  p.ksplit = pick_ksplit((long)mt * (Cout / block_n), p.num_kb, 16, block_n);
  dim3 grid(mt * p.ksplit, Cout / block_n, 1);
  return dispatch_n<MODE_F>(block_n, tmap, p, grid, stream);
}

int dcgan_conv_fprop(const void* x, const void* w_nk, const float* bias, void* out, int NB, int H, int W,
                     int Cin, int Cout, int act, int block_n, cudaStream_t stream) {
  return conv_fprop_impl(x, w_nk, bias, out, NB, H, W, Cin, Cout, act, block_n, 0, stream);
}

int dcgan_conv_fprop_kn(const void* x, const void* w_kn, const float* bias, void* out, int NB, int H, int W,
                        int Cin, int Cout, int act, int block_n, cudaStream_t stream) {
  return conv_fprop_impl(x, w_kn, bias, out, NB, H, W, Cin, Cout, act, block_n, 1, stream);
}

int dcgan_conv_fprop_kn_stats(const void* x, const void* w_kn, const float* bias, void* out, float* stats,
                              int stat_groups, int NB, int H, int W, int Cin, int Cout, int block_n,
                              cudaStream_t stream) {
  return conv_fprop_impl(x, w_kn, bias, out, NB, H, W, Cin, Cout, 0, block_n, 1, stream, stats, stat_groups);
}

static int conv_tconv_impl(const void* x, const void* w_nk, const float* bias, void* out, int NB, int H, int W,
                           int Cin, int Cout, int act, int block_n, cudaStream_t stream, float* stats,
                           int stat_groups) {
  if (use_v2()) {
    return igemm2_conv(1, x, w_nk, bias, out, NB, H, W, Cin, Cout, act, block_n < 0 ? 0 : block_n, 0, stats,
                       stat_groups, stream);
  }
  if (Cin % 64 || Cout % 64) return DCGAN_ERR_SHAPE;
  if ((long)NB * H * W * Cin >= (1L << 31)) return DCGAN_ERR_SHAPE;  // the gather uses 32-bit element offsets
  const int lh = ilog2_exact(H), lw = ilog2_exact(W);
  if (lh < 0 || lw < 0) return DCGAN_ERR_SHAPE;
  IgemmParams p{};
  p.err_flag = error_flag_ptr();
  p.mn_variant = mn_variant_ptr();
  if (!p.err_flag || !p.mn_variant) return DCGAN_ERR_DRIVER;
  p.g = reinterpret_cast<const __nv_bfloat16*>(x);
  p.gH = H; p.gW = W; p.gC = Cin;
  p.rH_log2 = lh; p.rW_log2 = lw;
  p.rows_total = NB * H * W;
  p.Cn = Cout; p.Cd = 0;
  p.num_kb = 0;  // per parity class, derived in-kernel
  p.ksplit = 1;
  p.t_long_first = (g_igemm_tuning & 64) ? 0 : 1;
  if (stats != nullptr) {
    if (stat_groups < 1 || p.rows_total % stat_groups || (p.rows_total / stat_groups) % kBlockM) return DCGAN_ERR_SHAPE;
    p.stats = stats;
    p.stat_group_rows = p.rows_total / stat_groups;
  }
  p.out = out; p.bias = bias; p.act = act; p.oH = 2 * H; p.oW = 2 * W;
  const int mt = (p.rows_total + kBlockM - 1) / kBlockM;
  // large-M layers: parity-fused kernel (9 gathers per channel chunk instead of 25); block_n < 0 forces the
  // per-parity kernels (tests / profiling)
  const bool want_t4 = (g_igemm_tuning & 8) && block_n == 0 && mt >= 90 && (Cout == 64 || Cout == 128);
  if (want_t4) {
    const int bn = (Cout == 128 && (g_igemm_tuning & 16)) ? 128 : 64;
    CUtensorMap tmap4;
    int rc4 = get_tmap_2d(w_nk, (uint64_t)25 * Cout, Cin, bn, 64, &tmap4);
    if (rc4) return rc4;
    dim3 grid4(mt, Cout / bn, 1);
    return bn == 64 ? launch_t4<64, 3, 6>(tmap4, p, grid4, stream) : launch_t4<128, 3, 6>(tmap4, p, grid4, stream);
  }
  if (block_n < 0) block_n = 0;
  if (block_n <= 0) block_n = pick_block_n(Cout, mt, 4, 4 * (Cin / 64), 8);
  if (block_n < 0 || Cout % block_n) return DCGAN_ERR_SHAPE;
  CUtensorMap tmap;
  int rc = get_tmap_2d(w_nk, (uint64_t)25 * Cout, Cin, block_n, 64, &tmap);
  if (rc) return rc;
  p.ksplit = pick_ksplit((long)mt * (Cout / block_n) * 4, 4 * (Cin / 64), 8, block_n);
  dim3 grid(mt * p.ksplit, Cout / block_n, 4);
  return dispatch_n<MODE_T>(block_n, tmap, p, grid, stream);
}

int dcgan_conv_tconv(const void* x, const void* w_nk, const float* bias, void* out, int NB, int H, int W,
                     int Cin, int Cout, int act, int block_n, cudaStream_t stream) {
  return conv_tconv_impl(x, w_nk, bias, out, NB, H, W, Cin, Cout, act, block_n, stream, nullptr, 1);
}

int dcgan_conv_tconv_stats(const void* x, const void* w_nk, const float* bias, void* out, float* stats,
                           int stat_groups, int NB, int H, int W, int Cin, int Cout, int block_n,
                           cudaStream_t stream) {
  return conv_tconv_impl(x, w_nk, bias, out, NB, H, W, Cin, Cout, 0, block_n, stream, stats, stat_groups);
}

// Data-gradient convs whose epilogue also forms the batch-norm backward reductions of the layer they feed
// (second-generation kernels only; see Igemm2Params::bn_*).
int dcgan_conv_tconv_bnred(const void* x, const void* w_nk, void* out, const void* pre, const float* saved,
                           const float* gamma, const float* beta, float* red, int slabs, int stat_groups, int bn_act,
                           int NB, int H, int W, int Cin, int Cout, int block_n, cudaStream_t stream) {
  if (!use_v2()) return DCGAN_ERR_SHAPE;
  BnRedArgs a{pre, saved, gamma, beta, red, slabs, stat_groups, bn_act, nullptr, nullptr, nullptr, 0, 0, 0};
  return igemm2_conv(1, x, w_nk, nullptr, out, NB, H, W, Cin, Cout, 0, block_n, 0, nullptr, 1, stream, &a);
}

int dcgan_conv_fprop_kn_bnred(const void* x, const void* w_kn, void* out, const void* pre, const float* saved,
                              const float* gamma, const float* beta, float* red, int slabs, int stat_groups,
                              int bn_act, int NB, int H, int W, int Cin, int Cout, int block_n,
                              cudaStream_t stream) {
  if (!use_v2()) return DCGAN_ERR_SHAPE;
  BnRedArgs a{pre, saved, gamma, beta, red, slabs, stat_groups, bn_act, nullptr, nullptr, nullptr, 0, 0, 0};
  return igemm2_conv(0, x, w_kn, nullptr, out, NB, H, W, Cin, Cout, 0, block_n, 1, nullptr, 1, stream, &a);
}

// The same with the rest of that layer's batch-norm backward pass in the launch: after a grid-wide barrier the CTAs
// turn out (= da) into dpre and CTA 0 writes dgamma / dbeta (what dcgan_bn_bwd_apply does in a launch of its own).
int dcgan_conv_tconv_bnbwd(const void* x, const void* w_nk, void* out, const void* pre, const float* saved,
                           const float* gamma, const float* beta, float* red, void* dpre, float* dgamma, float* dbeta,
                           int slabs, int stat_groups, int param_slabs, int bn_act, int NB, int H, int W, int Cin,
                           int Cout, int block_n, cudaStream_t stream) {
  if (!use_v2() || dpre == nullptr) return DCGAN_ERR_SHAPE;
  BnRedArgs a{pre, saved, gamma, beta, red, slabs, stat_groups, bn_act, dpre, dgamma, dbeta, param_slabs, 0, 0};
  return igemm2_conv(1, x, w_nk, nullptr, out, NB, H, W, Cin, Cout, 0, block_n, 0, nullptr, 1, stream, &a);
}

int dcgan_conv_fprop_kn_bnbwd(const void* x, const void* w_kn, void* out, const void* pre, const float* saved,
                              const float* gamma, const float* beta, float* red, void* dpre, float* dgamma,
                              float* dbeta, int slabs, int stat_groups, int param_slabs, int bn_act, int NB, int H,
                              int W, int Cin, int Cout, int block_n, cudaStream_t stream) {
  if (!use_v2() || dpre == nullptr) return DCGAN_ERR_SHAPE;
  BnRedArgs a{pre, saved, gamma, beta, red, slabs, stat_groups, bn_act, dpre, dgamma, dbeta, param_slabs, 0, 0};
  return igemm2_conv(0, x, w_kn, nullptr, out, NB, H, W, Cin, Cout, 0, block_n, 1, nullptr, 1, stream, &a);
}

// Transposed-conv data gradient into a layer that is bias + activation only (d_h0_conv): the epilogue multiplies by
// act'(saved activation) and accumulates the bias gradient, so `out` is the gradient w.r.t. that layer's
// pre-activation (what dcgan_act_bwd produced in a launch of its own).
int dcgan_conv_tconv_actbwd(const void* x, const void* w_nk, void* out, const void* act_saved, float* dbias, int slabs,
                            int stat_groups, int bias_slabs, int act, int NB, int H, int W, int Cin, int Cout,
                            int block_n, cudaStream_t stream) {
  if (!use_v2() || act_saved == nullptr || dbias == nullptr) return DCGAN_ERR_SHAPE;
  BnRedArgs a{act_saved, nullptr, nullptr, nullptr, dbias, slabs, stat_groups, act, nullptr, nullptr, nullptr, 0, 1,
              bias_slabs};
  return igemm2_conv(1, x, w_nk, nullptr, out, NB, H, W, Cin, Cout, 0, block_n, 0, nullptr, 1, stream, &a);
}

// Forward conv + training-mode batch norm + activation in one launch: pre = conv(x) + bias (bf16, kept for the
// backward pass), sums += its per-group moments, then -- after a grid-wide barrier -- act_out = act(bn(pre)), saved =
// (mean, rstd), EMA shadows advanced (dcgan_conv_*_stats followed by dcgan_bn_apply).
int dcgan_conv_fprop_kn_bn(const void* x, const void* w_kn, const float* bias, void* pre, void* act_out, float* sums,
                           const float* gamma, const float* beta, float* saved, float* ema_mean, float* ema_var,
                           int stat_groups, int NB, int H, int W, int Cin, int Cout, float eps, float decay, int bn_act,
                           int block_n, cudaStream_t stream) {
  if (!use_v2() || act_out == nullptr || sums == nullptr) return DCGAN_ERR_SHAPE;
  BnFwdArgs f{act_out, gamma, beta, saved, ema_mean, ema_var, eps, decay, bn_act};
  return igemm2_conv(0, x, w_kn, bias, pre, NB, H, W, Cin, Cout, 0, block_n, 1, sums, stat_groups, stream, nullptr, &f);
}

int dcgan_conv_tconv_bn(const void* x, const void* w_nk, const float* bias, void* pre, void* act_out, float* sums,
                        const float* gamma, const float* beta, float* saved, float* ema_mean, float* ema_var,
                        int stat_groups, int NB, int H, int W, int Cin, int Cout, float eps, float decay, int bn_act,
                        int block_n, cudaStream_t stream) {
  if (!use_v2() || act_out == nullptr || sums == nullptr) return DCGAN_ERR_SHAPE;
  BnFwdArgs f{act_out, gamma, beta, saved, ema_mean, ema_var, eps, decay, bn_act};
  return igemm2_conv(1, x, w_nk, bias, pre, NB, H, W, Cin, Cout, 0, block_n, 0, sums, stat_groups, stream, nullptr, &f);
}

int dcgan_conv_wgrad(const void* big, const void* small, float* dw, int NB, int Hs, int Ws, int Cg, int Cd,
                     int splits, int block_n, cudaStream_t stream) {
  if (use_v2()) return igemm2_wgrad(big, small, dw, NB, Hs, Ws, Cg, Cd, splits, block_n, stream);
  if (Cg % 64 || Cd % 128) return DCGAN_ERR_SHAPE;
  const int lh = ilog2_exact(Hs), lw = ilog2_exact(Ws);
  if (lh < 0 || lw < 0) return DCGAN_ERR_SHAPE;
  const int rows = NB * Hs * Ws;
  if (rows % kBlockK) return DCGAN_ERR_SHAPE;
  if ((long)rows * 4 * Cg >= (1L << 31)) return DCGAN_ERR_SHAPE;  // the gather uses 32-bit element offsets
  const int total_kb = rows / kBlockK;
  if (block_n <= 0) block_n = (Cg % 256 == 0) ? 256 : (Cg % 128 == 0 ? 128 : 64);
  if (Cg % block_n) return DCGAN_ERR_SHAPE;
  const int tiles = 25 * (Cd / kBlockM) * (Cg / block_n);
  if (splits <= 0) {
    splits = 1;
    while (tiles * splits < g_wgrad_cta_target && splits * 2 <= total_kb && total_kb % (splits * 2) == 0 &&
           total_kb / (splits * 2) >= 4)
      splits *= 2;
  }
  if (total_kb % splits) return DCGAN_ERR_SHAPE;
  IgemmParams p{};
  p.err_flag = error_flag_ptr();
  p.mn_variant = mn_variant_ptr();
  if (!p.err_flag || !p.mn_variant) return DCGAN_ERR_DRIVER;
  p.g = reinterpret_cast<const __nv_bfloat16*>(big);
  p.gH = 2 * Hs; p.gW = 2 * Ws; p.gC = Cg;
  p.rH_log2 = lh; p.rW_log2 = lw;
  p.rows_total = rows;
  p.Cn = Cg; p.Cd = Cd;
  p.num_kb = total_kb / splits;
  p.out = dw; p.bias = nullptr; p.act = 0; p.oH = 0; p.oW = 0;
  CUtensorMap tmap;
  int rc = get_tmap_2d(small, (uint64_t)rows, Cd, 64, 64, &tmap);
  if (rc) return rc;
  dim3 grid(25, (Cd / kBlockM) * (Cg / block_n), splits);
  return dispatch_n<MODE_W>(block_n, tmap, p, grid, stream);
}

}  // extern "C"
