// Implicit-GEMM convolution kernels, second generation: BOTH operands arrive by TMA, the CTAs are persistent,
// the accumulator is double-buffered in TMEM and the epilogue of one tile runs under the main loop of the next.
//
// What changed against csrc/igemm_sm100.cu (kept as the A/B reference, dcgan_set_igemm_tuning bit 7):
//   * The im2col operand is no longer gathered by 128 threads with cp.async.  The NHWC activation tensor is
//     described to the TMA unit as a 4-D tiled tensor map {C, W, H, N}; a 128-row GEMM tile is a box
//     {64 channels, bw pixels, bh rows, bn images}.  The filter tap is a shift of the box origin, TensorFlow's
//     asymmetric SAME padding (1 before / 2 after) is the TMA unit's out-of-bounds zero fill (signed
//     coordinates), and the stride-2 sampling of the forward conv is the tensor map's element stride {1,2,2,1}.
//     One elected lane issues one bulk tensor copy per operand and k-block; no address arithmetic in the SM.
//   * Persistent CTAs loop over work items (tile x K slice) with a static round-robin schedule; the smem ring
//     keeps streaming across tile boundaries, so a short K loop (8-18 k-blocks in the transposed layers) no
//     longer pays barrier init / TMEM allocation / pipeline fill / epilogue per tile.  Tiles of 64 / 128 columns run
//     TWO CTAs per SM (half the ring each, see launch2): two independent MMA-issue streams hide each other's
//     per-slot control flow; measured residency in profiles/igemm2_residency_r2.log.
//   * TMEM holds two accumulators per CTA (2 x BLOCK_N columns): warps 2-5 drain tile i (tcgen05.ld -> bias / activation /
//     batch-norm moments -> bf16 stores) while warp 1 already issues the MMAs of tile i+1.
//   * Split-K is deterministic: every slice stores its f32 partial tile to a scratch buffer, bumps the tile's
//     arrival counter, and the slice that arrives last adds all partials in slice order and runs the epilogue.
//     Nobody waits for anybody (no co-residency requirement), and the filter gradients no longer depend on the
//     order of floating-point atomics.
//
// Modes (same math as igemm_sm100.cu; geometry = TensorFlow 5x5 / stride 2 / SAME):
//   MODE_F  strided conv          = tf.nn.conv2d (distriubted_model.py:183) / data gradient of conv2d_transpose
//   MODE_T  transposed conv       = tf.nn.conv2d_transpose (distriubted_model.py:200) / data gradient of conv2d;
//           the four output parity classes (3x3, 3x2, 2x3, 2x2 taps) are work items of ONE launch, longest first
//   MODE_W  filter gradient       (autodiff of image_train.py:109-112): K = pixels, both operands MN-major
#include <mutex>
#include <unordered_map>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "igemm2.h"
#include "bn_tail.cuh"
#include "../include/dcgan_b200.h"

namespace dcgan {

namespace {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;
constexpr int kATileBytes = kBlockM * kBlockK * 2;  // 16 KiB
// The ring slots can be dealt to several producer warps (a stand-alone TMA-only loop during bring-up moved ~2x the
// bytes with two issuing warps per CTA; that microbenchmark is not kept).  A slot always belongs to the same warp
// (slot % kNumProducers): its waits on that slot's barrier then come in phase order (a warp that skipped a phase
// could not tell "two phases done" from "none" by parity).  In the real kernels a second producer warp gains nothing
// once two CTAs share an SM (profiles/igemm2_two_producers_r2.log vs profiles/igemm2_harness_r2.log), so it is off.
#ifndef IGEMM2_PRODUCERS
#define IGEMM2_PRODUCERS 1
#endif
constexpr int kNumProducers = IGEMM2_PRODUCERS;
// -DIGEMM2_TIMING: per-role cycle counters written to Igemm2Params::dbg ([CTA][16] long long), bring-up only
#ifdef IGEMM2_TIMING
#define T_NOW() clock64()
#define T_ADD(var, t0) (var) += clock64() - (t0)
#else
#define T_NOW() 0ll
#define T_ADD(var, t0) ((void)0)
#endif
constexpr int kMmaWarp = kNumProducers;             // warps [0, P) TMA producers, warp P MMA, then 4 epilogue warps
constexpr int kThreads = (kNumProducers + 5) * 32;
constexpr int kEpiThreads = 128;

template <int BLOCK_N, int STAGES, int KPS = 1>
struct Smem2 {
  static constexpr int kBTileBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kSubBytes = kATileBytes + kBTileBytes;  // one k-block: A tile + B tile
  static constexpr int kStageBytes = KPS * kSubBytes;          // a ring slot holds KPS consecutive k-blocks
  static constexpr int kBarOffset = STAGES * kStageBytes;  // full[STAGES], empty[STAGES], tfull[2], tempty[2]
  static constexpr int kMiscOffset = kBarOffset + 256;     // tmem base, abort flag, split-K "last" flag
  static constexpr int kStatOffset = kMiscOffset + 64;     // [4 warps][2][BLOCK_N] f32 epilogue column sums
  static constexpr int kStoreOffset = kStatOffset + 8 * BLOCK_N * 4;  // [4 warps][32 rows][64 B] output staging
  static constexpr int kTotal = kStoreOffset + 8192 + 1024;           // + slack for the 1024-byte alignment
  static_assert(2 * STAGES + 4 <= 32, "barrier block is 256 bytes");
};

struct WorkItem {
  int m0, n0;    // F/T: first pixel row / first out channel.  W: first dense channel / first gathered channel
  int tile;      // tile id (scratch slot of the split-K partials)
  int ks;        // K slice
  int kb0, nkb;  // k-block range of this slice
  int ph, pw;    // T: output parity class
  int tap;       // W: filter tap
};

// Work item `u` of the static schedule (tile x K slice).
template <int MODE, int BLOCK_N>
__device__ __forceinline__ WorkItem decode_work(int u, const Igemm2Params& p) {
  WorkItem it;
  it.ks = u % p.ksplit;
  const int t = u / p.ksplit;
  it.ph = it.pw = 0;
  it.tap = 0;
  int total_kb;
  if (MODE == MODE_W) {
    it.tile = t;
    it.tap = t % 25;  // neighbouring CTAs share the dense operand tile (L2) and differ in the tap
    const int q = t / 25;
    it.n0 = (q % p.n_tiles) * BLOCK_N;
    it.m0 = (q / p.n_tiles) * kBlockM;
    total_kb = p.num_kb;
  } else if (MODE == MODE_F) {
    it.m0 = (t % p.m_tiles) * kBlockM;  // neighbouring CTAs share the filter tile (L2)
    it.n0 = (t / p.m_tiles) * BLOCK_N;
    it.tile = t;
    total_kb = p.num_kb;
  } else {
    const int per_class = p.m_tiles * p.n_tiles;
    const int rank = t / per_class;  // 0: 3x3 taps, 1: 3x2, 2: 2x3, 3: 2x2 -- the long K loops are scheduled first
    const int q = t - rank * per_class;
    it.m0 = (q % p.m_tiles) * kBlockM;
    it.n0 = (q / p.m_tiles) * BLOCK_N;
    it.tile = t;
    it.ph = 1 - (rank >> 1);
    it.pw = 1 - (rank & 1);
    total_kb = (it.ph ? 3 : 2) * (it.pw ? 3 : 2) * (p.gC / kBlockK);
  }
  it.kb0 = (it.ks * total_kb) / p.ksplit;
  it.nkb = ((it.ks + 1) * total_kb) / p.ksplit - it.kb0;
  return it;
}

__device__ __forceinline__ float4 ldcg_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void stcg_f4(float4* p, float4 v) {
  asm volatile("st.global.cg.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

template <int MODE, int BLOCK_N, int STAGES, int MINB, int KPS>
__global__ void __launch_bounds__(kThreads, MINB)
igemm2_kernel(const __grid_constant__ CUtensorMap tmap_g,  // gathered operand: 4-D NHWC activation tensor
              const __grid_constant__ CUtensorMap tmap_d,  // dense operand: 2-D (filter, or W: the small activation)
              const Igemm2Params p) {
  pdl_launch_dependents();
#ifdef IGEMM2_TIMING
  long long t_kernel_begin;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_kernel_begin));
#endif
  using L = Smem2<BLOCK_N, STAGES, KPS>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + L::kMiscOffset);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  volatile int* last_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 2);
  float* stat_smem = reinterpret_cast<float*>(smem + L::kStatOffset);
  uint4* store_smem = reinterpret_cast<uint4*>(smem + L::kStoreOffset);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;

  const int first_group = blockIdx.x;
  const int group_step = gridDim.x;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4);  // one elected lane of each epilogue warp
    }
    *abort_flag = 0;
    *last_flag = 0;
    fence_mbar_init();
    tma_prefetch_desc(&tmap_g);
    tma_prefetch_desc(&tmap_d);
  }
  if (warp == kMmaWarp) {
    tmem_alloc(tmem_ptr_smem, 2 * BLOCK_N);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_wait();  // set-up above overlaps the previous kernel's tail; global memory is touched from here on

  const int cpt = p.gC / kBlockK;  // F/T: 64-channel chunks per filter tap
  const int rW = 1 << p.rW_log2, rH = 1 << p.rH_log2;

  if (warp < kNumProducers) {
    // =================================== TMA producer ==========================================
    // all 32 lanes walk the schedule (uniform control flow); the elected lane arms the barrier and issues the copies
    {
      int s = 0;
      uint32_t phase = 0;
      bool ok = true;
      long long tp_wait = 0, tp_n = 0;
      const long long tp_begin = T_NOW();
      for (int u = first_group; u < p.total_groups && ok; u += group_step) {
        const WorkItem wi = decode_work<MODE, BLOCK_N>(u, p);
        // origin of the tile's pixel box (F/T); the box always starts at column 0
        const int img0 = wi.m0 >> (p.rW_log2 + p.rH_log2);
        const int h0 = (wi.m0 >> p.rW_log2) & (rH - 1);
        // (ta, tb, cc): filter-tap row / column index inside the class and channel chunk, advanced like an odometer
        const int nb = (MODE == MODE_F) ? 5 : (wi.pw ? 3 : 2);
        int cc = 0, ta = 0, tb = 0;
        if (MODE != MODE_W) {
          const int lt0 = wi.kb0 / cpt;
          cc = wi.kb0 - lt0 * cpt;
          ta = lt0 / nb;
          tb = lt0 - ta * nb;
        }
        const int kh_w = wi.tap / 5, kw_w = wi.tap - 5 * (wi.tap / 5);
        // a ring slot takes KPS consecutive k-blocks (the last slot of an item may be short): with 64-column tiles
        // the MMA warp's fixed cost per slot exceeds the tensor time of one k-block, two per slot halve its share
        for (int kb = 0; kb < wi.nkb; kb += KPS) {
          const int nsub = (wi.nkb - kb) < KPS ? (wi.nkb - kb) : KPS;
          const bool mine = (s % kNumProducers) == warp;  // every warp walks the whole sequence, fills its own slots
          const long long tw0 = T_NOW();
          if (mine && !mbar_wait(&empty_bar[s], phase ^ 1, abort_flag, p.err_flag)) { ok = false; break; }
          T_ADD(tp_wait, tw0); ++tp_n;
          const bool issue = mine && elect_one();
          if (issue) mbar_arrive_expect_tx(&full_bar[s], nsub * L::kSubBytes);
#pragma unroll
          for (int j = 0; j < KPS; ++j) {
            if (j < nsub) {
              uint8_t* sA = smem + s * L::kStageBytes + j * L::kSubBytes;
              uint8_t* sB = sA + kATileBytes;
              if (issue) {
                if (MODE == MODE_W) {
                  const int px0 = (wi.kb0 + kb + j) * kBlockK;
                  tma_load_2d(sA, &tmap_d, &full_bar[s], wi.m0, px0);
                  tma_load_2d(sA + kATileBytes / 2, &tmap_d, &full_bar[s], wi.m0 + 64, px0);
                  const int imgk = px0 >> (p.rW_log2 + p.rH_log2);
                  const int hk = (px0 >> p.rW_log2) & (rH - 1);
#pragma unroll
                  for (int jj = 0; jj < BLOCK_N / 64; ++jj)
                    tma_load_4d(sB + jj * 8192, &tmap_g, &full_bar[s], wi.n0 + 64 * jj, kw_w - 1, 2 * hk + kh_w - 1, imgk);
                } else {
                  int tap;
                  if (MODE == MODE_F) {
                    tma_load_4d(sA, &tmap_g, &full_bar[s], cc * kBlockK, tb - 1, 2 * h0 + ta - 1, img0);
                    tap = ta * 5 + tb;
                  } else {
                    // out[2q+p] takes x[q + (p - i)] through tap k = 2i (p = 1) or 2i+1 (p = 0)
                    tma_load_4d(sA, &tmap_g, &full_bar[s], cc * kBlockK, wi.pw - tb, h0 + wi.ph - ta, img0);
                    const int kh = wi.ph ? 2 * ta : 2 * ta + 1;
                    const int kw = wi.pw ? 2 * tb : 2 * tb + 1;
                    tap = kh * 5 + kw;
                  }
                  if (p.b_mn) {  // TensorFlow layout [tap][Ck][Cn]: 64(k) x 64(n) boxes, one per 64 output channels
#pragma unroll
                    for (int jj = 0; jj < BLOCK_N / 64; ++jj)
                      tma_load_2d(sB + jj * 8192, &tmap_d, &full_bar[s], wi.n0 + 64 * jj, tap * p.gC + cc * kBlockK);
                  } else {
                    tma_load_2d(sB, &tmap_d, &full_bar[s], cc * kBlockK, tap * p.Cn + wi.n0);
                  }
                }
              }
              if (MODE != MODE_W) { if (++cc == cpt) { cc = 0; if (++tb == nb) { tb = 0; ++ta; } } }
            }
          }
          __syncwarp();
          if (++s == STAGES) { s = 0; phase ^= 1; }
        }
      }
#ifdef IGEMM2_TIMING
      if (p.dbg != nullptr && warp == 0 && lane == 0) {
        long long* d = p.dbg + (size_t)blockIdx.x * 16;
        d[0] = clock64() - tp_begin; d[1] = tp_wait; d[2] = tp_n;
      }
#endif
    }
  } else if (warp == kMmaWarp) {
    // =================================== MMA issuer ============================================
    // all 32 lanes walk the schedule; the tcgen05 instructions run on the elected lane
    {
      constexpr bool kMN = (MODE == MODE_W);
      const bool b_mn = kMN || (p.b_mn != 0);
      const uint32_t idesc = make_idesc_bf16(kBlockM, BLOCK_N, kMN, b_mn);
      const bool swap = (kMN || b_mn) && (*p.mn_variant != 0);
      const uint32_t lbo = kMN ? (swap ? 1024u : 8192u) : 16u;
      const uint32_t hi_a = smem_desc_hi_sw128(kMN ? (swap ? 8192u : 1024u) : 1024u);
      const uint32_t kstep = (kMN ? 2048u : 32u) >> 4;  // descriptor address units per UMMA_K = 16 step
      const uint32_t lbo_b = b_mn ? (swap ? 1024u : 8192u) : 16u;
      const uint32_t hi_b = smem_desc_hi_sw128(b_mn ? (swap ? 8192u : 1024u) : 1024u);
      const uint32_t kstep_b = (b_mn ? 2048u : 32u) >> 4;
      int s = 0;
      uint32_t phase = 0;
      int ti = 0;
      bool ok = true;
      long long tm_full = 0, tm_tmem = 0, tm_n = 0;
      const long long tm_begin = T_NOW();
      for (int u = first_group; u < p.total_groups && ok; u += group_step, ++ti) {
        const WorkItem wi = decode_work<MODE, BLOCK_N>(u, p);
        const int a = ti & 1;
        // the epilogue must have drained this accumulator (two tiles ago)
        const long long tt0 = T_NOW();
        if (!mbar_wait(&tempty_bar[a], ((ti >> 1) & 1) ^ 1, abort_flag, p.err_flag)) break;
        T_ADD(tm_tmem, tt0);
        tc_fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(a * BLOCK_N);
        for (int kb = 0; kb < wi.nkb; kb += KPS) {
          const int nsub = (wi.nkb - kb) < KPS ? (wi.nkb - kb) : KPS;
          const long long tf0 = T_NOW();
          if (!mbar_wait(&full_bar[s], phase, abort_flag, p.err_flag)) { ok = false; break; }
          T_ADD(tm_full, tf0); ++tm_n;
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + s * L::kStageBytes);
          if (elect_one()) {
#pragma unroll
            for (int j = 0; j < KPS; ++j) {
              if (j < nsub) {
                const uint32_t a_lo = smem_desc_lo(a_addr + j * L::kSubBytes, lbo);
                const uint32_t b_lo = smem_desc_lo(a_addr + j * L::kSubBytes + kATileBytes, lbo_b);
#pragma unroll
                for (int k = 0; k < kBlockK / 16; ++k)
                  umma_bf16(tacc, smem_desc_join(hi_a, a_lo + k * kstep), smem_desc_join(hi_b, b_lo + k * kstep_b),
                            idesc, (kb | j | k) != 0 ? 1u : 0u);
              }
            }
            umma_commit(&empty_bar[s]);  // frees the smem slot once these MMAs retire
          }
          __syncwarp();
          if (++s == STAGES) { s = 0; phase ^= 1; }
        }
        if (ok) {
          if (elect_one()) umma_commit(&tfull_bar[a]);  // accumulator complete
          __syncwarp();
        }
      }
#ifdef IGEMM2_TIMING
      if (p.dbg != nullptr && lane == 0) {
        long long* d = p.dbg + (size_t)blockIdx.x * 16;
        d[3] = clock64() - tm_begin; d[4] = tm_full; d[5] = tm_n; d[6] = tm_tmem; d[7] = ti;
      }
#endif
    }
  } else {
    // =================================== epilogue warps ========================================
    const int et = tid - (kMmaWarp + 1) * 32;  // 0..127
    const int q = warp & 3;   // TMEM lane quadrant this warp may read
    const int row_in_tile = q * 32 + lane;
    const bool do_bnred = (MODE != MODE_W) && p.bn_red != nullptr;
    const bool do_stats = (MODE != MODE_W) && (p.stats != nullptr || do_bnred);
    int ti = 0;
    bool ok = true;
    long long te_wait = 0, te_ld = 0, te_st = 0;
    const long long te_begin = T_NOW();
    for (int u = first_group; u < p.total_groups && ok; u += group_step, ++ti) {
      const WorkItem wi = decode_work<MODE, BLOCK_N>(u, p);
      const int a = ti & 1;
      const long long te0 = T_NOW();
      if (!mbar_wait(&tfull_bar[a], (ti >> 1) & 1, abort_flag, p.err_flag)) { ok = false; break; }
      T_ADD(te_wait, te0);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t)(a * BLOCK_N) + ((uint32_t)(q * 32) << 16);

      // ---- split-K: park the partial tile, the last slice to arrive finishes the tile ----------------
      bool from_scratch = false;
      if (p.ksplit > 1) {
        float4* part = reinterpret_cast<float4*>(p.partials) +
                       ((size_t)wi.tile * p.ksplit + wi.ks) * (size_t)(BLOCK_N / 4) * kBlockM + row_in_tile;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            stcg_f4(part + (size_t)((c >> 2) + j) * kBlockM,
                    make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                                __uint_as_float(v[4 * j + 3])));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[a]);  // the MMA warp may reuse this accumulator
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) {
          const int old = atomicAdd(p.counters + wi.tile, 1);
          const int last = (old == p.ksplit - 1) ? 1 : 0;
          if (last) p.counters[wi.tile] = 0;  // ready for the next launch (stream order)
          *last_flag = last;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const bool last = (*last_flag != 0);
        asm volatile("bar.sync 1, 128;" ::: "memory");  // everybody has read the flag before it is rewritten
        if (!last) continue;
        __threadfence();
        from_scratch = true;
      }
      const float4* part0 = reinterpret_cast<const float4*>(p.partials) +
                            (size_t)wi.tile * p.ksplit * (size_t)(BLOCK_N / 4) * kBlockM + row_in_tile;
      // one 32-column chunk of this thread's accumulator row, from TMEM or as the ordered sum of the K slices
      auto load_chunk = [&](int c, float (&f)[32]) {
        if (!from_scratch) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = 0.f;
          for (int sidx = 0; sidx < p.ksplit; ++sidx) {
            const float4* ps = part0 + (size_t)sidx * (size_t)(BLOCK_N / 4) * kBlockM + (size_t)(c >> 2) * kBlockM;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 t = ldcg_f4(ps + (size_t)j * kBlockM);
              f[4 * j] += t.x; f[4 * j + 1] += t.y; f[4 * j + 2] += t.z; f[4 * j + 3] += t.w;
            }
          }
        }
      };

      if (MODE != MODE_W) {
        const int R = wi.m0 + row_in_tile;
        const bool valid = R < p.rows_total;
        const __nv_bfloat16* prow = nullptr;  // fused batch-norm backward: this row of the saved pre-activations
        int sg = 0;
        if (valid) {
          size_t Ro;  // output pixel, row-major over the output tensor
          if (MODE == MODE_F) {
            Ro = (size_t)R;
          } else {
            const int qw = R & (rW - 1);
            const int qh = (R >> p.rW_log2) & (rH - 1);
            const int n = R >> (p.rW_log2 + p.rH_log2);
            Ro = (size_t)((n * p.oH + 2 * qh + wi.ph) * p.oW + 2 * qw + wi.pw);
          }
          if (do_bnred) {
            const int g = wi.m0 / p.stat_group_rows;  // slab of this tile (tiles never straddle slabs)
            sg = min(g, p.bn_stat_groups - 1);
            prow = reinterpret_cast<const __nv_bfloat16*>(p.bn_pre) +
                   (Ro - (size_t)(g - sg) * p.bn_slab_out_rows) * p.Cn + wi.n0;
          }
        }
        float* ssm = stat_smem;
        // Output rows leave through a per-warp staging area: a lane holds 64 bytes of ITS row per 32-column chunk,
        // and storing them directly is four 16-byte stores per lane to 32 different lines (half-sector writes; measured
        // 13-16 k cycles per 128x128 tile, as much as the tile's main loop).  Instead the chunk is transposed through
        // shared memory and every store instruction writes whole 64-byte row segments of 8 rows.
        uint4* stg = store_smem + (size_t)q * 128;  // [32 rows][4 x 16 B], 16-byte pieces XOR-swizzled by (row >> 1)
        __nv_bfloat16* srow[4];                       // rows 8 * i + lane / 4 of this warp's 32: piece lane % 4
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int Rr = wi.m0 + q * 32 + 8 * i + (lane >> 2);
          srow[i] = nullptr;
          if (Rr < p.rows_total) {
            size_t Ror;
            if (MODE == MODE_F) {
              Ror = (size_t)Rr;
            } else {
              const int qw = Rr & (rW - 1);
              const int qh = (Rr >> p.rW_log2) & (rH - 1);
              const int n = Rr >> (p.rW_log2 + p.rH_log2);
              Ror = (size_t)((n * p.oH + 2 * qh + wi.ph) * p.oW + 2 * qw + wi.pw);
            }
            srow[i] = reinterpret_cast<__nv_bfloat16*>(p.out) + Ror * p.Cn + wi.n0 + (lane & 3) * 8;
          }
        }
        if (do_stats) asm volatile("bar.sync 1, 128;" ::: "memory");  // the previous tile's sums have been flushed
#pragma unroll 1
        for (int c = 0; c < BLOCK_N; c += 32) {
          float f[32];
          const long long tl0 = T_NOW();
          load_chunk(c, f);
          T_ADD(te_ld, tl0);
          if (p.bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + wi.n0 + c + j));
              f[j] += b.x; f[j + 1] += b.y; f[j + 2] += b.z; f[j + 3] += b.w;
            }
          }
          if (do_bnred && p.bn_act_only) {
            // the layer below has no batch norm (d_h0: bias + LeakyReLU): what is stored is already the gradient
            // w.r.t. its pre-activation, dy = da * act'(a), with the sign taken from the saved activation
            uint4 pv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              pv[j] = valid ? __ldg(reinterpret_cast<const uint4*>(prow + c) + j) : make_uint4(0, 0, 0, 0);
            const uint32_t* pw32 = reinterpret_cast<const uint32_t*>(pv);
            const float neg = (p.bn_act == kActRelu) ? 0.f : (p.bn_act == kActLrelu ? 0.2f : 1.f);
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float2 a2 = unpack_bf16x2(pw32[j >> 1]);
              f[j] = valid ? f[j] * (a2.x > 0.f ? 1.f : neg) : 0.f;
              f[j + 1] = valid ? f[j + 1] * (a2.y > 0.f ? 1.f : neg) : 0.f;
            }
          }
          const long long ts0 = T_NOW();
          {
            uint32_t packed[16];
            // kind selected once per chunk, not per element; the entry points that fuse batch-norm moments or
            // backward reductions have no activation, so `f` still holds the pre-activation values they need
            apply_act_chunk(f, p.act);
#pragma unroll
            for (int j = 0; j < 16; ++j) packed[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
            __syncwarp();  // the previous chunk has been read out of the staging area
#pragma unroll
            for (int j = 0; j < 4; ++j)
              stg[lane * 4 + (j ^ ((lane >> 1) & 3))] =
                  make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int r = 8 * i + (lane >> 2);
              const uint4 v = stg[r * 4 + ((lane & 3) ^ ((r >> 1) & 3))];
              if (srow[i] != nullptr) *reinterpret_cast<uint4*>(srow[i] + c) = v;
            }
          }
          T_ADD(te_st, ts0);
          if (do_stats) {  // batch-norm moments of this tile, reduced over the warp's 32 rows
            float sq[32];
            if (do_bnred && p.bn_act_only) {
#pragma unroll
              for (int j = 0; j < 32; ++j) sq[j] = 0.f;  // only the column sums of dy (the bias gradient) are wanted
            } else if (do_bnred) {
              // dy = da * act'(gamma * xhat + beta) and dy * xhat instead of (x, x^2)
              uint4 pv[4];
#pragma unroll
              for (int j = 0; j < 4; ++j)
                pv[j] = valid ? __ldg(reinterpret_cast<const uint4*>(prow + c) + j) : make_uint4(0, 0, 0, 0);
              const uint32_t* pw32 = reinterpret_cast<const uint32_t*>(pv);
              const float* mean = p.bn_saved + (size_t)sg * 2 * p.Cn + wi.n0 + c;
              const float* rstd = mean + p.Cn;
              const float* gm = p.bn_gamma + wi.n0 + c;
              const float* bt = p.bn_beta + wi.n0 + c;
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                const float4 m4 = __ldg(reinterpret_cast<const float4*>(mean + j));
                const float4 r4 = __ldg(reinterpret_cast<const float4*>(rstd + j));
                const float4 g4 = __ldg(reinterpret_cast<const float4*>(gm + j));
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(bt + j));
                const float2 p01 = unpack_bf16x2(pw32[j >> 1]);
                const float2 p23 = unpack_bf16x2(pw32[(j >> 1) + 1]);
                const float pe[4] = {p01.x, p01.y, p23.x, p23.y};
                const float me[4] = {m4.x, m4.y, m4.z, m4.w}, re[4] = {r4.x, r4.y, r4.z, r4.w};
                const float ge[4] = {g4.x, g4.y, g4.z, g4.w}, be[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float xh = (pe[e] - me[e]) * re[e];
                  const float y = ge[e] * xh + be[e];
                  const float slope = (p.bn_act == kActRelu) ? (y > 0.f ? 1.f : 0.f)
                                                             : (p.bn_act == kActLrelu ? (y > 0.f ? 1.f : 0.2f) : 1.f);
                  const float dy = valid ? f[j + e] * slope : 0.f;
                  f[j + e] = dy;
                  sq[j + e] = dy * xh;
                }
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                f[j] = valid ? f[j] : 0.f;
                sq[j] = f[j] * f[j];
              }
            }
            const float s1 = warp_transpose_reduce32(f);
            const float s2 = warp_transpose_reduce32(sq);
            ssm[(q * 2) * BLOCK_N + c + lane] = s1;
            ssm[(q * 2 + 1) * BLOCK_N + c + lane] = s2;
          }
        }
        if (!from_scratch) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[a]);
        }
        if (do_stats) {
          asm volatile("bar.sync 1, 128;" ::: "memory");
          const int grp = wi.m0 / p.stat_group_rows;
          float* gs = (do_bnred ? p.bn_red : p.stats) + (size_t)grp * 2 * p.Cn + wi.n0;
          const bool act_only = do_bnred && p.bn_act_only;
          if (act_only) gs = (grp < p.bn_bias_slabs) ? p.bn_red + wi.n0 : nullptr;  // bn_red = the bias gradient [Cn]
          for (int i = et; i < BLOCK_N && wi.m0 < p.rows_total && gs != nullptr; i += kEpiThreads) {
            float t1 = 0.f, t2 = 0.f;
#pragma unroll
            for (int ww = 0; ww < 4; ++ww) {
              t1 += ssm[(ww * 2) * BLOCK_N + i];
              t2 += ssm[(ww * 2 + 1) * BLOCK_N + i];
            }
            atomicAdd(gs + i, t1);
            if (!act_only) atomicAdd(gs + p.Cn + i, t2);
          }
        }
      } else {
        // dW[tap][n0 + col][m0 + row] += acc: one writer per element (split-K slices were summed in order above)
        float* obase = reinterpret_cast<float*>(p.out) + (size_t)wi.tap * p.Cn * p.Cd + (wi.m0 + row_in_tile);
#pragma unroll 1
        for (int c = 0; c < BLOCK_N; c += 32) {
          float f[32];
          load_chunk(c, f);
#pragma unroll
          for (int j = 0; j < 32; ++j) atomicAdd(obase + (size_t)(wi.n0 + c + j) * p.Cd, f[j]);
        }
        if (!from_scratch) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[a]);
        }
      }
    }
#ifdef IGEMM2_TIMING
    if (warp == kMmaWarp + 1 && lane == 0 && p.dbg != nullptr) {
      long long* d = p.dbg + (size_t)blockIdx.x * 16;
      d[8] = clock64() - te_begin; d[9] = te_wait; d[10] = te_ld; d[11] = te_st;
    }
#endif
  }

#ifdef IGEMM2_TIMING
  if (tid == 0 && p.dbg != nullptr) {  // residency probe: which SM, when (ns)
    long long* d = p.dbg + (size_t)blockIdx.x * 16;
    unsigned smid; long long t1;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    d[12] = smid; d[13] = t_kernel_begin; d[14] = t1;
  }
#endif
  // ---- teardown ---------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 2 * BLOCK_N);

  // ---- batch-norm tail ----------------------------------------------------------------------------
  // Once every CTA has stored its tiles and flushed its column sums, the whole grid runs the elementwise pass that
  // used to be a kernel of its own (bn_apply / bn_bwd_apply).  The launch keeps the grid within the CTAs that are
  // resident at once (launch2), which the barrier needs.
  if constexpr (MODE != MODE_W) {
    if (p.tail_kind != 0) {
      grid_barrier(p.grid_barrier, p.err_flag);
      const int groups = p.rows_total / p.stat_group_rows;
      if (p.tail_kind == 1) {
        bn_fwd_apply_grid<kThreads>(reinterpret_cast<const __nv_bfloat16*>(p.out),
                                    reinterpret_cast<__nv_bfloat16*>(p.tail_dst), p.stats, p.bn_gamma, p.bn_beta,
                                    p.tail_saved, p.tail_ema_mean, p.tail_ema_var, groups, p.rows_out / groups, p.Cn,
                                    p.tail_eps, p.tail_decay, p.bn_act);
      } else {
        bn_bwd_apply_grid<kThreads>(reinterpret_cast<const __nv_bfloat16*>(p.out),
                                    reinterpret_cast<const __nv_bfloat16*>(p.bn_pre),
                                    reinterpret_cast<__nv_bfloat16*>(p.tail_dst), p.bn_saved, p.bn_gamma, p.bn_beta,
                                    p.bn_red, p.tail_dgamma, p.tail_dbeta, groups, p.bn_stat_groups,
                                    p.tail_param_slabs, p.bn_slab_out_rows, p.Cn, p.bn_act);
      }
    }
  }
}

// =================================================================================================
// host side
// =================================================================================================
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
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
  uint64_t d[4];
  uint32_t box[4];
  uint32_t stride, rank;
  int dev;
  bool operator==(const MapKey& o) const { return std::memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
    return (size_t)h;
  }
};
static_assert(sizeof(MapKey) % 8 == 0, "hashed as 64-bit words");

std::mutex g_map_mu;
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_map_cache;
constexpr size_t kMapCacheLimit = 4096;  // the engine's buffers are static (a few dozen maps); callers that feed
                                         // fresh tensors every call would otherwise grow the cache without bound

int cached_encode(const MapKey& key, const cuuint64_t* dims, const cuuint64_t* strides, const cuuint32_t* box,
                  const cuuint32_t* estr, CUtensorMap* out) {
  std::lock_guard<std::mutex> lock(g_map_mu);
  auto it = g_map_cache.find(key);
  if (it != g_map_cache.end()) { *out = it->second; return 0; }
  EncodeTiledFn fn = encode_fn();
  if (!fn) return DCGAN_ERR_DRIVER;
  CUtensorMap m;
  CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, key.rank, const_cast<void*>(key.ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return DCGAN_ERR_DRIVER;
  if (g_map_cache.size() >= kMapCacheLimit) g_map_cache.clear();
  g_map_cache.emplace(key, m);
  *out = m;
  return 0;
}

int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}

// 2-D row-major bf16 [rows][cols], box [box_r][box_c]
int tmap_2d(const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_r, uint32_t box_c, CUtensorMap* out) {
  MapKey key;
  std::memset(&key, 0, sizeof key);
  key.ptr = ptr; key.d[0] = cols; key.d[1] = rows; key.box[0] = box_c; key.box[1] = box_r;
  key.stride = 1; key.rank = 2; key.dev = current_device();
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {box_c, box_r};
  cuuint32_t estr[2] = {1, 1};
  return cached_encode(key, dims, strides, box, estr, out);
}

// 4-D NHWC bf16 activation tensor {C, W, H, N}; the box takes 64 channels of bw x bh x bn pixels, sampling every
// `stride`-th pixel in x and y (the box extent passed to the driver is count * stride, see cuTensorMapEncodeTiled)
int tmap_nhwc(const void* ptr, int N, int H, int W, int C, int bw, int bh, int bn, int stride, CUtensorMap* out) {
  if (bw * stride > 256 || bh * stride > 256 || bn > 256) return DCGAN_ERR_SHAPE;
  MapKey key;
  std::memset(&key, 0, sizeof key);
  key.ptr = ptr;
  key.d[0] = (uint64_t)C; key.d[1] = (uint64_t)W; key.d[2] = (uint64_t)H; key.d[3] = (uint64_t)N;
  key.box[0] = 64; key.box[1] = (uint32_t)bw; key.box[2] = (uint32_t)bh; key.box[3] = (uint32_t)bn;
  key.stride = (uint32_t)stride; key.rank = 4; key.dev = current_device();
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)(bw * stride), (cuuint32_t)(bh * stride), (cuuint32_t)bn};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  return cached_encode(key, dims, strides, box, estr, out);
}

// `rows` consecutive pixels (x fastest, then y, then image) of a rH x rW grid as a box; every extent a power of two
bool pixel_box(int rows, int rH, int rW, int* bw, int* bh, int* bn) {
  if (rW > rows || rows % rW) return false;
  *bw = rW;
  if (rH * rW >= rows) {
    *bh = rows / rW;
    *bn = 1;
    return rH % *bh == 0;
  }
  *bh = rH;
  *bn = rows / (rH * rW);
  return rows % (rH * rW) == 0;
}

// ---- split-K scratch ------------------------------------------------------------------------------------
constexpr size_t kCounterBytes = 64 * 1024;            // 16384 tile counters (zero between launches)
constexpr size_t kDefaultScratch = 48u * 1024 * 1024;  // default scratch per device, allocated on first need
struct Workspace { void* ptr = nullptr; size_t bytes = 0; };
thread_local Workspace t_ws;                            // set by dcgan_set_workspace
std::mutex g_ws_mu;
Workspace g_default_ws[64];

int get_workspace(Workspace* out) {
  if (t_ws.ptr != nullptr) { *out = t_ws; return 0; }
  const int dev = current_device();
  if (dev < 0 || dev >= 64) return DCGAN_ERR_DRIVER;
  std::lock_guard<std::mutex> lock(g_ws_mu);
  if (g_default_ws[dev].ptr == nullptr) {
    void* ptr = nullptr;
    // not possible inside a stream capture: callers that capture graphs either run one step before capturing
    // (the engine does) or hand over their own buffer with dcgan_set_workspace
    if (cudaMalloc(&ptr, kDefaultScratch) != cudaSuccess) { cudaGetLastError(); return DCGAN_ERR_DRIVER; }
    if (cudaMemset(ptr, 0, kCounterBytes) != cudaSuccess) { cudaFree(ptr); return DCGAN_ERR_DRIVER; }
    g_default_ws[dev].ptr = ptr;
    g_default_ws[dev].bytes = kDefaultScratch;
  }
  *out = g_default_ws[dev];
  return 0;
}

int num_sms() {
  static int sms[64] = {0};
  const int dev = current_device();
  if (dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}

int g_max_ctas = 0;        // 0 = one CTA per SM
int g_force_ksplit2 = 0;   // 0 = chosen per layer
int g_pref_bn = 0;         // 0 = cost model; 64 / 128 / 256 = tile width wherever it divides the layer
int g_stage_override = 0;  // 0 = deepest ring that fits

// MINB = CTAs per SM the configuration is built for.  A single MMA-issuing warp spends ~450 cycles of control flow
// (barrier wait, fences, descriptor set-up, commit) per ring slot, more than the tensor work of a slot for tiles of
// 64 or 128 columns (128 / 256 cycles), so those run two CTAs per SM -- two independent issue streams -- with half
// the ring depth each; 256-column tiles (512 cycles per slot, all of TMEM) run one.
template <int MODE, int BLOCK_N, int STAGES, int MINB, int KPS = 1>
int launch2(const CUtensorMap& tg, const CUtensorMap& td, const Igemm2Params& p, cudaStream_t stream) {
  using L = Smem2<BLOCK_N, STAGES, KPS>;
  auto kern = igemm2_kernel<MODE, BLOCK_N, STAGES, MINB, KPS>;
  static bool attr_done[64] = {false};  // per instantiation and device
  const int dev = current_device();
  if (dev < 0 || dev >= 64) return DCGAN_ERR_DRIVER;
  if (!attr_done[dev]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal) != cudaSuccess)
      return DCGAN_ERR_LAUNCH;
    // all of the SM's unified memory as shared memory: MINB CTAs of this size only fit with the largest carve-out
    // (and the occupancy query below answers for the configured preference, not for what a launch would pick)
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    attr_done[dev] = true;
  }
  int ctas = g_max_ctas > 0 ? g_max_ctas : num_sms() * MINB;
  // The grid barrier of the batch-norm tail needs every CTA resident.  MINB CTAs per SM fit by construction:
  // registers through __launch_bounds__(kThreads, MINB), shared memory through the assertion below, TMEM through
  // 2 * BLOCK_N * MINB <= 512 columns.  (cudaOccupancyMaxActiveBlocksPerMultiprocessor answers 1 for the MINB = 2
  // configurations although two CTAs do run per SM -- measured with the %smid / %globaltimer probe of the
  // IGEMM2_TIMING build, profiles/igemm2_residency_r2.log -- so it is not consulted.)
  static_assert(MINB * (L::kTotal + 1024) <= 233472, "MINB CTAs must fit the SM's shared memory");
  static_assert(MINB * 2 * BLOCK_N <= 512, "MINB CTAs must fit TMEM");
  if (p.tail_kind != 0 && ctas > num_sms() * MINB) ctas = num_sms() * MINB;
  const int grid = p.total_groups < ctas ? p.total_groups : ctas;
  launch_k(kern, dim3(grid), dim3(kThreads), L::kTotal, stream, tg, td, p);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

template <int MODE>
int dispatch2(int block_n, const CUtensorMap& tg, const CUtensorMap& td, const Igemm2Params& p, cudaStream_t st) {
  const bool deep = g_stage_override == 2;  // one CTA per SM with the deepest ring that fits (A/B runs)
  switch (block_n) {
    case 64:
      // variant 1 (or stage mode 3 for A/B runs): two k-blocks per ring slot, two slots, still two CTAs per SM
      if (p.variant == 1 || g_stage_override == 3) return launch2<MODE, 64, 2, 2, 2>(tg, td, p, st);
      return deep ? launch2<MODE, 64, 8, 1>(tg, td, p, st) : launch2<MODE, 64, 4, 2>(tg, td, p, st);
    case 128: return deep ? launch2<MODE, 128, 6, 1>(tg, td, p, st) : launch2<MODE, 128, 3, 2>(tg, td, p, st);
    case 256: return launch2<MODE, 256, 4, 1>(tg, td, p, st);
    default: return DCGAN_ERR_SHAPE;
  }
}

}  // namespace

int igemm2_scratch(void** ptr, size_t* bytes) {  // the partial-tile area of the calling thread's workspace
  Workspace ws;
  int rc = get_workspace(&ws);
  if (rc) return rc;
  *ptr = reinterpret_cast<uint8_t*>(ws.ptr) + kCounterBytes;
  *bytes = ws.bytes - kCounterBytes;
  return 0;
}
int igemm2_num_sms() { return num_sms(); }
int igemm2_max_ctas() { return g_max_ctas; }
int igemm2_forced_ksplit() { return g_force_ksplit2; }
int igemm2_tmap_nhwc(const void* ptr, int N, int H, int W, int C, int bw, int bh, int bn, int stride, void* out_map) {
  return tmap_nhwc(ptr, N, H, W, C, bw, bh, bn, stride, reinterpret_cast<CUtensorMap*>(out_map));
}
int igemm2_tmap_2d(const void* ptr, unsigned long long rows, unsigned long long cols, unsigned box_r, unsigned box_c,
                   void* out_map) {
  return tmap_2d(ptr, rows, cols, box_r, box_c, reinterpret_cast<CUtensorMap*>(out_map));
}
bool igemm2_pixel_box(int rows, int rH, int rW, int* bw, int* bh, int* bn) { return pixel_box(rows, rH, rW, bw, bh, bn); }

static long long* g_dbg = nullptr;
void igemm2_set_debug(long long* ptr) { g_dbg = ptr; }
void igemm2_set_max_ctas(int v) { g_max_ctas = v > 0 ? v : 0; }
void igemm2_set_tile(int v) { g_pref_bn = (v == 64 || v == 128 || v == 256) ? v : 0; }
void igemm2_set_ksplit(int v) { g_force_ksplit2 = (v >= 1 && v <= 16) ? v : 0; }
void igemm2_set_stage_mode(int v) { g_stage_override = v; }

int igemm2_set_workspace(void* ptr, size_t bytes) {
  if (ptr != nullptr && bytes < kCounterBytes + (1u << 20)) return DCGAN_ERR_LIMIT;
  t_ws.ptr = ptr;
  t_ws.bytes = ptr ? bytes : 0;
  return 0;
}

// Cost model behind the (tile width, split-K) choice, in SM cycles.  A k-block costs the slower of its MMAs
// (128 x bn x 64 MACs at 8192 MACs per cycle and SM) and its bytes from L2 at the per-SM share of the fabric;
// a work item adds a fixed cost (epilogue tail, scheduling) plus the partial-tile traffic when K is split.
static double plan_cost(long tiles, double kb_per_tile, int bn, int ks, int ctas) {
  if (bn <= 128) ctas *= 2;  // two CTAs per SM (see launch2)
  const double kBytesPerCycle = 36.0;  // measured-order L2 -> SM rate per SM with every SM loading
  const double t_kb = fmax(2.0 * bn, (kATileBytes + bn * 128.0) / kBytesPerCycle);
  const double part = (ks > 1) ? (128.0 * bn * 4.0 * 2.0) / kBytesPerCycle + 1500.0 : 0.0;
  const double item = (kb_per_tile / ks) * t_kb + 1200.0 + part;
  const long items = tiles * ks;
  const long waves = (items + ctas - 1) / ctas;
  return (double)waves * item;
}

void igemm2_plan(long tiles_bn64, double kb_per_tile, int Cn, int min_kb, int force_bn, int* block_n, int* ksplit) {
  const int ctas = g_max_ctas > 0 ? g_max_ctas : num_sms();
  double best = 1e30;
  *block_n = -1;
  *ksplit = 1;
  const int cands[3] = {256, 128, 64};
  if (force_bn <= 0 && g_pref_bn > 0 && Cn % g_pref_bn == 0) force_bn = g_pref_bn;
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (Cn % bn) continue;
    if (force_bn > 0 && bn != force_bn) continue;
    const long tiles = tiles_bn64 / (bn / 64);
    for (int ks = 1; ks <= 16; ++ks) {
      if (g_force_ksplit2 > 0 && ks != g_force_ksplit2) continue;
      if (g_force_ksplit2 == 0 && ks > 8) break;
      // every slice keeps at least four k-blocks (a forced factor only has to leave one)
      if (ks > 1 && min_kb / ks < (g_force_ksplit2 > 0 ? 1 : 4)) break;
      const double c = plan_cost(tiles, kb_per_tile, bn, ks, ctas);
      if (c < best) { best = c; *block_n = bn; *ksplit = ks; }
    }
  }
  if (*block_n < 0 && g_force_ksplit2 > 0) {  // forced factor too large for this layer: fall back to the free choice
    const int keep = g_force_ksplit2;
    g_force_ksplit2 = 0;
    igemm2_plan(tiles_bn64, kb_per_tile, Cn, min_kb, force_bn, block_n, ksplit);
    g_force_ksplit2 = keep;
  }
}

// Measured plans (dcgan_igemm_plan_set: the Python autotuner times the candidates of a layer shape once and records
// the winner).  Consulted when the caller leaves tile width and split factor to the library and no global override
// (dcgan_set_igemm_tile / _ksplit) is active.
static std::mutex g_plan_mu;
static std::unordered_map<uint64_t, std::pair<int, int>> g_plan_table;
static bool plan_key(int kind, int NB, int H, int W, int Cin, int Cout, uint64_t* key) {
  if (kind < 0 || kind > 2 || NB < 1 || NB >= (1 << 14) || H < 1 || H >= (1 << 10) || W < 1 || W >= (1 << 10) ||
      Cin < 1 || Cin >= (1 << 13) || Cout < 1 || Cout >= (1 << 13))
    return false;
  *key = (uint64_t)kind | ((uint64_t)NB << 2) | ((uint64_t)H << 16) | ((uint64_t)W << 26) | ((uint64_t)Cin << 36) |
         ((uint64_t)Cout << 49);
  return true;
}
int igemm2_plan_set(int kind, int NB, int H, int W, int Cin, int Cout, int block_n, int ksplit, int variant) {
  uint64_t key;
  if (!plan_key(kind, NB, H, W, Cin, Cout, &key)) return DCGAN_ERR_SHAPE;
  std::lock_guard<std::mutex> lock(g_plan_mu);
  if (block_n <= 0) {
    g_plan_table.erase(key);
    return 0;
  }
  if ((block_n != 64 && block_n != 128 && block_n != 256) || ksplit < 1 || ksplit > 16 || variant < 0 ||
      variant > 1 || (variant == 1 && block_n != 64))
    return DCGAN_ERR_SHAPE;
  g_plan_table[key] = std::make_pair(block_n, ksplit | (variant << 8));
  return 0;
}
static bool plan_lookup(int kind, int NB, int H, int W, int Cin, int Cout, int* block_n, int* ksplit, int* variant) {
  uint64_t key;
  if (g_force_ksplit2 > 0 || g_pref_bn > 0 || !plan_key(kind, NB, H, W, Cin, Cout, &key)) return false;
  std::lock_guard<std::mutex> lock(g_plan_mu);
  auto it = g_plan_table.find(key);
  if (it == g_plan_table.end()) return false;
  *block_n = it->second.first;
  *ksplit = it->second.second & 0xff;
  *variant = it->second.second >> 8;
  return true;
}

static int ilog2_exact(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return ((1 << l) == v) ? l : -1;
}

static int attach_scratch(Igemm2Params* p, long tiles, int block_n);
int igemm2_attach_scratch(Igemm2Params* p, long tiles, int cols_per_tile) { return attach_scratch(p, tiles, cols_per_tile); }
static int attach_scratch(Igemm2Params* p, long tiles, int block_n) {
  p->partials = nullptr;
  p->counters = nullptr;
  p->grid_barrier = nullptr;
  if (p->ksplit <= 1 && p->tail_kind == 0) return 0;
  Workspace ws;
  int rc = get_workspace(&ws);
  if (rc) return rc;
  // the last two ints of the (zero) counter region are the grid barrier of the batch-norm tail
  const long max_tiles = (long)(kCounterBytes / 4) - 2;
  if (p->tail_kind != 0) p->grid_barrier = reinterpret_cast<int*>(ws.ptr) + max_tiles;
  if (p->ksplit <= 1) return 0;
  const size_t need = kCounterBytes + (size_t)tiles * p->ksplit * kBlockM * block_n * 4;
  if (tiles > max_tiles || need > ws.bytes) { p->ksplit = 1; return 0; }  // too big: do not split
  p->counters = reinterpret_cast<int*>(ws.ptr);
  p->partials = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws.ptr) + kCounterBytes);
  return 0;
}

int igemm2_conv(int transposed, const void* x, const void* w, const float* bias, void* out, int NB, int H, int W,
                int Cin, int Cout, int act, int block_n, int b_mn, float* stats, int stat_groups,
                cudaStream_t stream, const BnRedArgs* bnred, const BnFwdArgs* bnfwd) {
  if (Cin % 64 || Cout % 64) return DCGAN_ERR_SHAPE;
  if (bnred != nullptr && (stats != nullptr || bias != nullptr || act != 0 || bnfwd != nullptr)) return DCGAN_ERR_SHAPE;
  if (bnfwd != nullptr && (stats == nullptr || act != 0)) return DCGAN_ERR_SHAPE;
  if (!transposed && ((H & 1) || (W & 1))) return DCGAN_ERR_SHAPE;
  if (transposed && b_mn) return DCGAN_ERR_SHAPE;
  const int rH = transposed ? H : H / 2, rW = transposed ? W : W / 2;
  const int lh = ilog2_exact(rH), lw = ilog2_exact(rW);
  if (lh < 0 || lw < 0) return DCGAN_ERR_SHAPE;
  int bw, bh, bn_img;
  if (!pixel_box(kBlockM, rH, rW, &bw, &bh, &bn_img)) return DCGAN_ERR_SHAPE;
  Igemm2Params p{};
  p.err_flag = error_flag_ptr();
  p.mn_variant = mn_variant_ptr();
  if (!p.err_flag || !p.mn_variant) return DCGAN_ERR_DRIVER;
  p.rH_log2 = lh; p.rW_log2 = lw;
  p.rows_total = NB * rH * rW;
  p.gC = Cin; p.Cn = Cout; p.Cd = 0;
  p.b_mn = b_mn;
  p.dbg = g_dbg;
  p.out = out; p.bias = bias; p.act = act;
  p.oH = transposed ? 2 * H : rH; p.oW = transposed ? 2 * W : rW;
  if (stats != nullptr) {
    if (stat_groups < 1 || p.rows_total % stat_groups || (p.rows_total / stat_groups) % kBlockM) return DCGAN_ERR_SHAPE;
    p.stats = stats;
    p.stat_group_rows = p.rows_total / stat_groups;
  }
  if (bnred != nullptr) {
    // slabs are groups of whole images; every 128-row tile must lie inside one slab
    if (bnred->slabs < 1 || NB % bnred->slabs || ((NB / bnred->slabs) * rH * rW) % kBlockM) return DCGAN_ERR_SHAPE;
    p.bn_act_only = bnred->act_only; p.bn_bias_slabs = bnred->bias_slabs;
    if (bnred->act_only && (bnred->dpre != nullptr || bnred->bias_slabs < 0 || bnred->bias_slabs > bnred->slabs))
      return DCGAN_ERR_SHAPE;
    p.bn_pre = bnred->pre; p.bn_saved = bnred->saved; p.bn_gamma = bnred->gamma; p.bn_beta = bnred->beta;
    p.bn_red = bnred->red; p.bn_stat_groups = bnred->stat_groups; p.bn_act = bnred->act;
    p.stat_group_rows = (NB / bnred->slabs) * rH * rW;
    p.bn_slab_out_rows = (NB / bnred->slabs) * p.oH * p.oW;
    if (bnred->dpre != nullptr) {
      if (bnred->param_slabs < 0 || bnred->param_slabs > bnred->slabs) return DCGAN_ERR_SHAPE;
      p.tail_kind = 2;
      p.tail_dst = bnred->dpre;
      p.tail_dgamma = bnred->dgamma; p.tail_dbeta = bnred->dbeta; p.tail_param_slabs = bnred->param_slabs;
    }
  }
  if (bnfwd != nullptr) {
    if (NB % stat_groups) return DCGAN_ERR_SHAPE;  // groups are sets of whole images (contiguous in `out`)
    p.tail_kind = 1;
    p.tail_dst = bnfwd->act_out;
    p.bn_gamma = bnfwd->gamma; p.bn_beta = bnfwd->beta; p.bn_act = bnfwd->act;
    p.tail_saved = bnfwd->saved; p.tail_ema_mean = bnfwd->ema_mean; p.tail_ema_var = bnfwd->ema_var;
    p.tail_eps = bnfwd->eps; p.tail_decay = bnfwd->decay;
  }
  p.rows_out = NB * p.oH * p.oW;
  const int cpt = Cin / 64;
  const int mt = (p.rows_total + kBlockM - 1) / kBlockM;
  const int classes = transposed ? 4 : 1;
  const int min_kb = transposed ? 4 * cpt : 25 * cpt;
  const double avg_kb = transposed ? 6.25 * cpt : 25.0 * cpt;
  p.num_kb = transposed ? 0 : 25 * cpt;
  int ks;
  int tuned_bn, tuned_ks, tuned_var;
  if (block_n <= 0 && plan_lookup(transposed ? MODE_T : MODE_F, NB, H, W, Cin, Cout, &tuned_bn, &tuned_ks, &tuned_var) &&
      Cout % tuned_bn == 0 && tuned_ks <= min_kb) {
    block_n = tuned_bn;
    ks = tuned_ks;
    p.variant = tuned_var;
  } else {
    igemm2_plan((long)mt * (Cout / 64) * classes, avg_kb, Cout, min_kb, block_n > 0 ? block_n : 0, &block_n, &ks);
  }
  if (block_n <= 0 || Cout % block_n) return DCGAN_ERR_SHAPE;
  p.ksplit = ks;
  p.m_tiles = mt;
  p.n_tiles = Cout / block_n;
  const long tiles = (long)mt * p.n_tiles * classes;
  int rc = attach_scratch(&p, tiles, block_n);
  if (rc) return rc;
  p.total_groups = (int)(tiles * p.ksplit);
  CUtensorMap tg, td;
  rc = tmap_nhwc(x, NB, H, W, Cin, bw, bh, bn_img, transposed ? 1 : 2, &tg);
  if (rc) return rc;
  rc = b_mn ? tmap_2d(w, (uint64_t)25 * Cin, Cout, 64, 64, &td)
            : tmap_2d(w, (uint64_t)25 * Cout, Cin, block_n, 64, &td);
  if (rc) return rc;
  return transposed ? dispatch2<MODE_T>(block_n, tg, td, p, stream) : dispatch2<MODE_F>(block_n, tg, td, p, stream);
}

int igemm2_wgrad(const void* big, const void* small, float* dw, int NB, int Hs, int Ws, int Cg, int Cd, int splits,
                 int block_n, cudaStream_t stream) {
  if (Cg % 64 || Cd % 128) return DCGAN_ERR_SHAPE;
  const int lh = ilog2_exact(Hs), lw = ilog2_exact(Ws);
  if (lh < 0 || lw < 0) return DCGAN_ERR_SHAPE;
  const int rows = NB * Hs * Ws;
  if (rows % kBlockK) return DCGAN_ERR_SHAPE;
  int bw, bh, bn_img;
  if (!pixel_box(kBlockK, Hs, Ws, &bw, &bh, &bn_img)) return DCGAN_ERR_SHAPE;
  const int total_kb = rows / kBlockK;
  Igemm2Params p{};
  p.err_flag = error_flag_ptr();
  p.mn_variant = mn_variant_ptr();
  if (!p.err_flag || !p.mn_variant) return DCGAN_ERR_DRIVER;
  p.rH_log2 = lh; p.rW_log2 = lw;
  p.rows_total = rows;
  p.gC = Cg; p.Cn = Cg; p.Cd = Cd;
  p.num_kb = total_kb;
  p.dbg = g_dbg;
  p.out = dw;
  int ks;
  if (block_n <= 0 && splits > 0) block_n = (Cg % 256 == 0) ? 256 : (Cg % 128 == 0 ? 128 : 64);
  if (splits > 0) {
    if (Cg % block_n) return DCGAN_ERR_SHAPE;
    ks = splits < total_kb ? splits : total_kb;
  } else {
    int tuned_bn, tuned_ks, tuned_var;
    if (block_n <= 0 && plan_lookup(MODE_W, NB, Hs, Ws, Cg, Cd, &tuned_bn, &tuned_ks, &tuned_var) && Cg % tuned_bn == 0 &&
        tuned_ks <= total_kb) {
      block_n = tuned_bn;
      ks = tuned_ks;
      p.variant = tuned_var;
    } else {
      igemm2_plan((long)25 * (Cd / kBlockM) * (Cg / 64), (double)total_kb, Cg, total_kb, block_n > 0 ? block_n : 0,
                  &block_n, &ks);
    }
  }
  if (block_n <= 0 || Cg % block_n) return DCGAN_ERR_SHAPE;
  p.ksplit = ks;
  p.m_tiles = Cd / kBlockM;
  p.n_tiles = Cg / block_n;
  const long tiles = (long)25 * p.m_tiles * p.n_tiles;
  int rc = attach_scratch(&p, tiles, block_n);
  if (rc) return rc;
  p.total_groups = (int)(tiles * p.ksplit);
  CUtensorMap tg, td;
  rc = tmap_nhwc(big, NB, 2 * Hs, 2 * Ws, Cg, bw, bh, bn_img, 2, &tg);
  if (rc) return rc;
  rc = tmap_2d(small, (uint64_t)rows, Cd, 64, 64, &td);
  if (rc) return rc;
  return dispatch2<MODE_W>(block_n, tg, td, p, stream);
}

}  // namespace dcgan
