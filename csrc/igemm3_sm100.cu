// Transposed 5x5 / stride-2 convolution (tf.nn.conv2d_transpose, distriubted_model.py:200; also the data gradient
// of tf.nn.conv2d, :183), third generation: built around the fact that the kernels of igemm2_sm100.cu run at the
// L2 -> SM bandwidth limit (about 9.5 TB/s measured: 461 MB in 51 us for the d_h1 data gradient), not at the tensor
// pipe's.  This kernel moves fewer bytes for the same MMAs:
//
//   * Parity fusion.  out[n, 2q+p] = sum_i x[n, q + (p - i)] . W[tap(p, i)]: the four output parity classes of a
//     128-pixel input tile read the SAME nine shifted input tiles (dy, dx in {-1,0,1}).  One CTA computes all four
//     classes into four TMEM accumulators, so an input tile is fetched 9 times per channel chunk instead of 25.
//   * Row halo.  When a tile is bh whole rows of one image (H * W >= 128, W a multiple of 8), the three vertical
//     shifts are the same pixels moved by W rows of the GEMM operand.  One TMA box of (bh + 2) image rows per
//     horizontal shift is loaded, and the MMAs of dy = -1, 0, +1 read it through UMMA descriptors whose start
//     address is advanced by W rows (W * 128 bytes: whole 1024-byte swizzle atoms, so the 128B swizzle phase is
//     unchanged).  3 x (bh + 2) / bh input tiles per chunk instead of 9 -- 6.7x fewer activation bytes than the
//     per-parity kernels for the 16x16 layers.
//   * Filter multicast.  With the activations cheap, the 25 filter tiles per chunk dominate; the CTAs of a cluster
//     take consecutive pixel tiles, fetch 1/C of every filter tile each and the TMA unit multicasts it.
//
// Activation tiles (X ring) and filter tiles (Y ring) travel in separate rings because one X tile feeds 5-10
// (halo) or 1-4 (no halo) filter tiles.  Persistent CTAs, static round-robin over work items, and -- when four
// BLOCK_N-wide accumulators fit twice into the 512 TMEM columns (BLOCK_N = 64) -- the epilogue of one tile overlaps
// the main loop of the next.  Split-K (over channel chunks) uses the ordered scratch reduction of igemm2.
#include <cstdio>
#include <cstring>

#include "common.cuh"
#include "igemm2.h"
#include "../include/dcgan_b200.h"

namespace dcgan {

namespace {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;
constexpr int kThreads = 192;  // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue
constexpr int kEpiThreads = 128;
constexpr int kXSlotBytes = 24 * 1024;  // up to 192 rows: 128 + a two-row halo of at most 32 pixels each

template <int BLOCK_N, int X_STAGES, int Y_STAGES>
struct Smem3 {
  static constexpr int kYTileBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kYOffset = X_STAGES * kXSlotBytes;
  static constexpr int kBarOffset = kYOffset + Y_STAGES * kYTileBytes;
  // x_full[X], x_empty[X], y_full[Y], y_empty[Y], tfull[2], tempty[2]
  static constexpr int kNumBars = 2 * X_STAGES + 2 * Y_STAGES + 4;
  static constexpr int kMiscOffset = kBarOffset + ((kNumBars * 8 + 63) / 64) * 64;
  static constexpr int kStatOffset = kMiscOffset + 64;                  // [2][4 warps][2][BLOCK_N] f32
  static constexpr int kTotal = kStatOffset + 16 * BLOCK_N * 4 + 1024;  // + slack for the 1024-byte alignment
};

struct Item3 {
  int m0, n0, tile, ks, cc0, cc1;
};

// work group u -> this CTA's item: CTAs of a cluster take consecutive pixel tiles of one (n tile, K slice)
template <int BLOCK_N>
__device__ __forceinline__ Item3 decode3(int u, int r, const Igemm2Params& p) {
  Item3 it;
  it.ks = u % p.ksplit;
  const int t = u / p.ksplit;
  const int mt = (t % p.m_groups) * p.cluster + r;
  const int nt = t / p.m_groups;
  it.m0 = mt * kBlockM;
  it.n0 = nt * BLOCK_N;
  it.tile = nt * (p.m_groups * p.cluster) + mt;
  const int cpt = p.gC / kBlockK;
  it.cc0 = (it.ks * cpt) / p.ksplit;  // K slices are ranges of channel chunks: every slice touches all four parities
  it.cc1 = ((it.ks + 1) * cpt) / p.ksplit;
  return it;
}

__device__ __forceinline__ uint32_t cluster_ctarank3() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all3() {
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ float4 ldcg3(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void stcg3(float4* p, float4 v) {
  asm volatile("st.global.cg.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// Schedules.  The producer and the MMA issuer walk the same sequence of X tiles and, per X tile, of filter tiles:
//   SCHED_T4 (transposed conv, four parity accumulators):
//     halo:     X tiles = (dx, chunk); each feeds dy = -1, 0, +1 (row offset (dy + 1) * W) x the parities it reaches
//     no halo:  X tiles = (dy, dx, chunk); each feeds the parities it reaches
//     Parity (ph, pw) is reached from shift dy through tap kh = ph + 1 - 2 dy when that lies in [0, 4]: dy = +1
//     only reaches ph = 1 (likewise dx / pw).
//   SCHED_FH (strided conv, one accumulator, halo only): input rows 2 oh + kh - 1 for kh = 0, 2, 4 are the same
//     stride-2 sampled rows moved by one output row, likewise kh = 1, 3.  X tiles = (kw, kh parity, chunk) of
//     bh + 2 sampled rows; each feeds its 3 (2) taps at row offsets 0, oW, 2 oW.
enum { SCHED_T4 = 0, SCHED_FH = 1 };

template <int SCHED, class FX, class FY, class FD>
__device__ __forceinline__ bool walk_schedule(const Igemm2Params& p, const Item3& it, int h0, FX&& on_x, FY&& on_y,
                                              FD&& on_x_done) {
  if (SCHED == SCHED_T4) {
    const int n_xg = p.halo ? 3 : 9;
    for (int xg = 0; xg < n_xg; ++xg) {
      const int dx = p.halo ? xg - 1 : xg % 3 - 1;
      const int dy_lo = p.halo ? -1 : xg / 3 - 1;
      const int dy_hi = p.halo ? 1 : dy_lo;
      for (int cc = it.cc0; cc < it.cc1; ++cc) {
        if (!on_x(cc, dx, h0 + dy_lo)) return false;
        for (int dy = dy_lo; dy <= dy_hi; ++dy)
          for (int ph = (dy == 1) ? 1 : 0; ph < 2; ++ph)
            for (int pw = (dx == 1) ? 1 : 0; pw < 2; ++pw)
              if (!on_y(cc, (ph + 1 - 2 * dy) * 5 + (pw + 1 - 2 * dx), p.halo ? (dy + 1) * p.halo_row_bytes : 0,
                        ph * 2 + pw))
                return false;
        on_x_done();
      }
    }
  } else {
    for (int kw = 0; kw < 5; ++kw)
      for (int e = 0; e < 2; ++e)
        for (int cc = it.cc0; cc < it.cc1; ++cc) {
          if (!on_x(cc, kw - 1, 2 * h0 + e - 1)) return false;
          for (int i = 0; i < 3 - e; ++i)
            if (!on_y(cc, (e + 2 * i) * 5 + kw, i * p.halo_row_bytes, 0)) return false;
          on_x_done();
        }
  }
  return true;
}

template <int SCHED, int BLOCK_N, int X_STAGES, int Y_STAGES, bool DB>
__global__ void __launch_bounds__(kThreads, 1)
igemm3_kernel(const __grid_constant__ CUtensorMap tmap_x,  // activations, 4-D NHWC box (with or without halo rows)
              const __grid_constant__ CUtensorMap tmap_w,  // filter: T4 [25 * Cout][Cin], FH [25 * Cin][Cout] (TF layout)
              const Igemm2Params p) {
  pdl_launch_dependents();
  using L = Smem3<BLOCK_N, X_STAGES, Y_STAGES>;
  constexpr int kNAcc = (SCHED == SCHED_T4) ? 4 : 1;
  constexpr int kAccCols = kNAcc * BLOCK_N;
  static_assert((DB ? 2 : 1) * kAccCols <= 512, "TMEM has 512 columns");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sX = smem;
  uint8_t* sY = smem + L::kYOffset;
  uint64_t* x_full = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* x_empty = x_full + X_STAGES;
  uint64_t* y_full = x_empty + X_STAGES;
  uint64_t* y_empty = y_full + Y_STAGES;
  uint64_t* tfull_bar = y_empty + Y_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + L::kMiscOffset);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  volatile int* last_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 2);
  float* stat_smem = reinterpret_cast<float*>(smem + L::kStatOffset);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int csize = p.cluster;
  const int crank = (csize > 1) ? (int)cluster_ctarank3() : 0;
  const uint16_t cmask = (uint16_t)((1u << csize) - 1u);
  const int first_group = blockIdx.x / csize;
  const int group_step = gridDim.x / csize;

  if (tid == 0) {
    for (int s = 0; s < X_STAGES; ++s) { mbar_init(&x_full[s], 1); mbar_init(&x_empty[s], 1); }
    for (int s = 0; s < Y_STAGES; ++s) { mbar_init(&y_full[s], 1); mbar_init(&y_empty[s], csize); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], 4); }
    *abort_flag = 0;
    *last_flag = 0;
    fence_mbar_init();
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_smem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  if (csize > 1) cluster_sync_all3();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_wait();

  const int rW = 1 << p.rW_log2, rH = 1 << p.rH_log2;

  if (warp == 0) {
    // =================================== TMA producer ==========================================
    if (lane == 0) {
      int xs = 0, ys = 0;
      uint32_t xph = 0, yph = 0;
      bool ok = true;
      for (int u = first_group; u < p.total_groups && ok; u += group_step) {
        const Item3 it = decode3<BLOCK_N>(u, crank, p);
        const int img0 = it.m0 >> (p.rW_log2 + p.rH_log2);
        const int h0 = (it.m0 >> p.rW_log2) & (rH - 1);
        ok = walk_schedule<SCHED>(
            p, it, h0,
            [&](int cc, int cx, int cy) {
              if (!mbar_wait(&x_empty[xs], xph ^ 1, abort_flag, p.err_flag)) return false;
              mbar_arrive_expect_tx(&x_full[xs], (uint32_t)p.x_tile_bytes);
              tma_load_4d(sX + xs * kXSlotBytes, &tmap_x, &x_full[xs], cc * kBlockK, cx, cy, img0);
              if (++xs == X_STAGES) { xs = 0; xph ^= 1; }
              return true;
            },
            [&](int cc, int tap, int, int) {
              if (!mbar_wait(&y_empty[ys], yph ^ 1, abort_flag, p.err_flag)) return false;
              mbar_arrive_expect_tx(&y_full[ys], L::kYTileBytes);
              uint8_t* dst = sY + ys * L::kYTileBytes;
              if (SCHED == SCHED_T4) {  // filter rows = output channels, K-major
                if (csize == 1) {
                  tma_load_2d(dst, &tmap_w, &y_full[ys], cc * kBlockK, tap * p.Cn + it.n0);
                } else {  // this CTA's slice of the rows, delivered to every CTA of the cluster
                  const int nr = BLOCK_N / csize;
                  tma_load_2d_mc(dst + crank * nr * 128, &tmap_w, &y_full[ys], cc * kBlockK,
                                 tap * p.Cn + it.n0 + crank * nr, cmask);
                }
              } else {  // TensorFlow layout [tap][Ck][Cn]: 64(k) x 64(n) boxes; the cluster splits the k rows
                const int kr = 64 / csize;
#pragma unroll
                for (int j = 0; j < BLOCK_N / 64; ++j) {
                  if (csize == 1)
                    tma_load_2d(dst + j * 8192, &tmap_w, &y_full[ys], it.n0 + 64 * j, tap * p.gC + cc * kBlockK);
                  else
                    tma_load_2d_mc(dst + j * 8192 + crank * kr * 128, &tmap_w, &y_full[ys], it.n0 + 64 * j,
                                   tap * p.gC + cc * kBlockK + crank * kr, cmask);
                }
              }
              if (++ys == Y_STAGES) { ys = 0; yph ^= 1; }
              return true;
            },
            [] {});
      }
    }
  } else if (warp == 1) {
    // =================================== MMA issuer ============================================
    if (lane == 0) {
      constexpr bool kBmn = (SCHED == SCHED_FH);  // FH reads the filter as an MN-major operand
      constexpr uint32_t idesc = make_idesc_bf16(kBlockM, BLOCK_N, 0, kBmn ? 1 : 0);
      constexpr uint32_t lbo_b = kBmn ? 8192u : 16u;
      constexpr uint32_t kstep_b = kBmn ? 2048u : 32u;
      int xs = 0, ys = 0;
      uint32_t xph = 0, yph = 0;
      int ti = 0;
      bool ok = true;
      for (int u = first_group; u < p.total_groups && ok; u += group_step, ++ti) {
        const Item3 it = decode3<BLOCK_N>(u, crank, p);
        const int h0 = (it.m0 >> p.rW_log2) & (rH - 1);
        const int buf = DB ? (ti & 1) : 0;
        const uint32_t tpar = DB ? (((ti >> 1) & 1) ^ 1) : ((ti & 1) ^ 1);
        if (!mbar_wait(&tempty_bar[buf], tpar, abort_flag, p.err_flag)) break;
        tc_fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(buf * kAccCols);
        uint32_t started = 0;  // bit a: accumulator a already holds a partial sum
        uint32_t x_addr = 0;
        ok = walk_schedule<SCHED>(
            p, it, h0,
            [&](int, int, int) {
              if (!mbar_wait(&x_full[xs], xph, abort_flag, p.err_flag)) return false;
              x_addr = smem_u32(sX + xs * kXSlotBytes);
              return true;
            },
            [&](int, int, int x_off, int acc) {
              if (!mbar_wait(&y_full[ys], yph, abort_flag, p.err_flag)) return false;
              tc_fence_after();
              // a vertical shift is a move by whole image rows inside the halo box: whole 1024-byte swizzle atoms
              const uint32_t a_addr = x_addr + (uint32_t)x_off;
              const uint32_t b_addr = smem_u32(sY + ys * L::kYTileBytes);
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k) {
                const uint64_t da = make_smem_desc_sw128(a_addr + k * 32, 16, 1024);
                const uint64_t db = make_smem_desc_sw128(b_addr + k * kstep_b, lbo_b, 1024);
                umma_bf16(tacc + acc * BLOCK_N, da, db, idesc, (((started >> acc) & 1u) | (k != 0)) ? 1u : 0u);
              }
              started |= 1u << acc;
              if (csize == 1) umma_commit(&y_empty[ys]);
              else umma_commit_mc(&y_empty[ys], cmask);
              if (++ys == Y_STAGES) { ys = 0; yph ^= 1; }
              return true;
            },
            [&] {
              umma_commit(&x_empty[xs]);  // every MMA that reads this X tile has been issued
              if (++xs == X_STAGES) { xs = 0; xph ^= 1; }
            });
        if (ok) umma_commit(&tfull_bar[buf]);
      }
    }
  } else {
    // =================================== epilogue warps ========================================
    const int et = tid - 64;
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    const int row_in_tile = q * 32 + lane;
    const bool do_bnred = p.bn_red != nullptr;
    const bool do_stats = p.stats != nullptr || do_bnred;
    int ti = 0;
    bool ok = true;
    for (int u = first_group; u < p.total_groups && ok; u += group_step, ++ti) {
      const Item3 it = decode3<BLOCK_N>(u, crank, p);
      const int buf = DB ? (ti & 1) : 0;
      const uint32_t tpar = DB ? ((ti >> 1) & 1) : (ti & 1);
      if (!mbar_wait(&tfull_bar[buf], tpar, abort_flag, p.err_flag)) { ok = false; break; }
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t)(buf * kAccCols) + ((uint32_t)(q * 32) << 16);

      // ---- split-K: park the four partial accumulators; the slice that arrives last finishes the tile ----
      bool from_scratch = false;
      if (p.ksplit > 1) {
        float4* part = reinterpret_cast<float4*>(p.partials) +
                       ((size_t)it.tile * p.ksplit + it.ks) * (size_t)(kAccCols / 4) * kBlockM + row_in_tile;
#pragma unroll 1
        for (int c = 0; c < kAccCols; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            stcg3(part + (size_t)((c >> 2) + j) * kBlockM,
                  make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                              __uint_as_float(v[4 * j + 3])));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[buf]);
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) {
          const int old = atomicAdd(p.counters + it.tile, 1);
          const int last = (old == p.ksplit - 1) ? 1 : 0;
          if (last) p.counters[it.tile] = 0;
          *last_flag = last;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const bool last = (*last_flag != 0);
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (!last) continue;
        __threadfence();
        from_scratch = true;
      }
      const float4* part0 = reinterpret_cast<const float4*>(p.partials) +
                            (size_t)it.tile * p.ksplit * (size_t)(kAccCols / 4) * kBlockM + row_in_tile;
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
            const float4* ps = part0 + (size_t)sidx * (size_t)(kAccCols / 4) * kBlockM + (size_t)(c >> 2) * kBlockM;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 t = ldcg3(ps + (size_t)j * kBlockM);
              f[4 * j] += t.x; f[4 * j + 1] += t.y; f[4 * j + 2] += t.z; f[4 * j + 3] += t.w;
            }
          }
        }
      };

      const int R = it.m0 + row_in_tile;
      const bool valid = R < p.rows_total;
      const int qw = R & (rW - 1);
      const int qh = (R >> p.rW_log2) & (rH - 1);
      const int n = R >> (p.rW_log2 + p.rH_log2);
      const int g = do_bnred ? it.m0 / p.stat_group_rows : 0;  // slab of this tile (tiles never straddle slabs)
      const int sg = do_bnred ? min(g, p.bn_stat_groups - 1) : 0;
      float* ssm = stat_smem + (ti & 1) * (8 * BLOCK_N);
#pragma unroll 1
      for (int par = 0; par < kNAcc; ++par) {
        const size_t Ro = (SCHED == SCHED_T4) ? (size_t)((n * p.oH + 2 * qh + (par >> 1)) * p.oW + 2 * qw + (par & 1))
                                              : (size_t)R;
        __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.out) + Ro * p.Cn + it.n0;
        const __nv_bfloat16* prow = reinterpret_cast<const __nv_bfloat16*>(p.bn_pre) +
                                    (Ro - (size_t)(g - sg) * p.bn_slab_out_rows) * p.Cn + it.n0;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N; c += 32) {
          float f[32];
          load_chunk(par * BLOCK_N + c, f);
          if (p.bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + it.n0 + c + j));
              f[j] += b.x; f[j + 1] += b.y; f[j + 2] += b.z; f[j + 3] += b.w;
            }
          }
          if (valid) {
            uint32_t packed[16];
#pragma unroll
            for (int j = 0; j < 16; ++j)
              packed[j] = pack_bf16x2(apply_act(f[2 * j], p.act), apply_act(f[2 * j + 1], p.act));
            uint4* dst = reinterpret_cast<uint4*>(orow + c);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              dst[j] = make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
          }
          if (do_stats) {
            float sq[32];
            if (do_bnred) {  // dy = da * act'(gamma * xhat + beta) and dy * xhat instead of (x, x^2)
              uint4 pv[4];
#pragma unroll
              for (int j = 0; j < 4; ++j)
                pv[j] = valid ? __ldg(reinterpret_cast<const uint4*>(prow + c) + j) : make_uint4(0, 0, 0, 0);
              const uint32_t* pw32 = reinterpret_cast<const uint32_t*>(pv);
              const float* mean = p.bn_saved + (size_t)sg * 2 * p.Cn + it.n0 + c;
              const float* rstd = mean + p.Cn;
              const float* gm = p.bn_gamma + it.n0 + c;
              const float* bt = p.bn_beta + it.n0 + c;
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
            float* d1 = &ssm[(q * 2) * BLOCK_N + c + lane];
            float* d2 = &ssm[(q * 2 + 1) * BLOCK_N + c + lane];
            if (par == 0) { *d1 = s1; *d2 = s2; }  // the four parities add up in the slot this lane owns
            else { *d1 += s1; *d2 += s2; }
          }
        }
      }
      if (!from_scratch) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[buf]);
      }
      if (do_stats) {
        asm volatile("bar.sync 1, 128;" ::: "memory");
        float* gs = (do_bnred ? p.bn_red : p.stats) + (size_t)(it.m0 / p.stat_group_rows) * 2 * p.Cn + it.n0;
        for (int i = et; i < BLOCK_N && it.m0 < p.rows_total; i += kEpiThreads) {
          float t1 = 0.f, t2 = 0.f;
#pragma unroll
          for (int ww = 0; ww < 4; ++ww) {
            t1 += ssm[(ww * 2) * BLOCK_N + i];
            t2 += ssm[(ww * 2 + 1) * BLOCK_N + i];
          }
          atomicAdd(gs + i, t1);
          atomicAdd(gs + p.Cn + i, t2);
        }
      }
    }
  }

  // ---- teardown ---------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (csize > 1) cluster_sync_all3();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace

namespace {

template <int SCHED, int BLOCK_N, int X_STAGES, int Y_STAGES, bool DB>
int launch3(const CUtensorMap& tx, const CUtensorMap& tw, const Igemm2Params& p, int grid, cudaStream_t stream) {
  using L = Smem3<BLOCK_N, X_STAGES, Y_STAGES>;
  auto kern = igemm3_kernel<SCHED, BLOCK_N, X_STAGES, Y_STAGES, DB>;
  static bool attr_done[64] = {false};
  static int max_clusters[64][9] = {{0}};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return DCGAN_ERR_DRIVER;
  if (!attr_done[dev]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal) != cudaSuccess)
      return DCGAN_ERR_LAUNCH;
    attr_done[dev] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = L::kTotal;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = p.cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = p.cluster > 1 ? 1 : 0;
  if (p.cluster > 1) {  // persistent clusters: never more than can be resident at once
    if (max_clusters[dev][p.cluster] == 0) {
      int n = 0;
      cfg.gridDim = dim3(igemm2_num_sms() / p.cluster * p.cluster);
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) { cudaGetLastError(); n = 1; }
      max_clusters[dev][p.cluster] = n;
    }
    int clusters = grid / p.cluster;
    if (clusters > max_clusters[dev][p.cluster]) clusters = max_clusters[dev][p.cluster];
    cfg.gridDim = dim3(clusters * p.cluster);
  }
  return cudaLaunchKernelEx(&cfg, kern, tx, tw, p) == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int ilog2_exact3(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return ((1 << l) == v) ? l : -1;
}

}  // namespace

int igemm3_tconv(const void* x, const void* w, const float* bias, void* out, int NB, int H, int W, int Cin, int Cout,
                 int act, int block_n, float* stats, int stat_groups, cudaStream_t stream, const BnRedArgs* bnred) {
  if (Cin % 64 || Cout % 64) return DCGAN_ERR_SHAPE;
  if (bnred != nullptr && (stats != nullptr || bias != nullptr || act != 0)) return DCGAN_ERR_SHAPE;
  const int lh = ilog2_exact3(H), lw = ilog2_exact3(W);
  if (lh < 0 || lw < 0) return DCGAN_ERR_SHAPE;
  int bw, bh, bn_img;
  if (!igemm2_pixel_box(kBlockM, H, W, &bw, &bh, &bn_img)) return DCGAN_ERR_SHAPE;
  Igemm2Params p{};
  p.err_flag = error_flag_ptr();
  p.mn_variant = mn_variant_ptr();
  if (!p.err_flag || !p.mn_variant) return DCGAN_ERR_DRIVER;
  p.rH_log2 = lh; p.rW_log2 = lw;
  p.rows_total = NB * H * W;
  p.gC = Cin; p.Cn = Cout;
  p.out = out; p.bias = bias; p.act = act;
  p.oH = 2 * H; p.oW = 2 * W;
  if (stats != nullptr) {
    if (stat_groups < 1 || p.rows_total % stat_groups || (p.rows_total / stat_groups) % kBlockM) return DCGAN_ERR_SHAPE;
    p.stats = stats;
    p.stat_group_rows = p.rows_total / stat_groups;
  }
  if (bnred != nullptr) {
    if (bnred->slabs < 1 || NB % bnred->slabs || ((NB / bnred->slabs) * H * W) % kBlockM) return DCGAN_ERR_SHAPE;
    p.bn_pre = bnred->pre; p.bn_saved = bnred->saved; p.bn_gamma = bnred->gamma; p.bn_beta = bnred->beta;
    p.bn_red = bnred->red; p.bn_stat_groups = bnred->stat_groups; p.bn_act = bnred->act;
    p.stat_group_rows = (NB / bnred->slabs) * H * W;
    p.bn_slab_out_rows = (NB / bnred->slabs) * p.oH * p.oW;
  }
  // a tile of whole image rows can carry its two halo rows; the vertical shift (W rows of 128 bytes) must be a whole
  // number of 1024-byte swizzle atoms, and the box has to fit the X slot
  p.halo = (bn_img == 1 && W % 8 == 0 && (bh + 2) * bw * 128 <= kXSlotBytes) ? 1 : 0;
  p.halo_row_bytes = W * 128;
  p.x_tile_bytes = p.halo ? (bh + 2) * bw * 128 : kBlockM * 128;
  if (block_n != 64 && block_n != 128) block_n = 64;
  if (Cout % block_n) return DCGAN_ERR_SHAPE;
  const int cpt = Cin / 64;
  const int mt = (p.rows_total + kBlockM - 1) / kBlockM;
  p.m_tiles = mt;
  p.n_tiles = Cout / block_n;
  int cl = igemm2_cluster_pref() > 0 ? igemm2_cluster_pref() : 2;
  while (cl > 1 && cl > mt) cl >>= 1;
  p.cluster = cl;
  p.m_groups = (mt + cl - 1) / cl;
  int ctas = igemm2_max_ctas() > 0 ? igemm2_max_ctas() : igemm2_num_sms();
  if (ctas < cl) ctas = cl;
  ctas = ctas / cl * cl;
  const long tiles = (long)p.m_groups * cl * p.n_tiles;
  int ks = 1;
  if (igemm2_forced_ksplit() > 0) {
    ks = igemm2_forced_ksplit() < cpt ? igemm2_forced_ksplit() : cpt;
  } else {
    while (tiles * ks * 2 <= ctas && ks * 2 <= cpt) ks *= 2;  // few tiles: cut the channel chunks across CTAs
  }
  p.ksplit = ks;
  int rc = igemm2_attach_scratch(&p, tiles, 4 * block_n);
  if (rc) return rc;
  p.total_groups = (int)((long)p.m_groups * p.n_tiles * p.ksplit);
  CUtensorMap tx, tw;
  rc = igemm2_tmap_nhwc(x, NB, H, W, Cin, bw, p.halo ? bh + 2 : bh, bn_img, 1, &tx);
  if (rc) return rc;
  rc = igemm2_tmap_2d(w, (unsigned long long)25 * Cout, Cin, block_n / cl, 64, &tw);
  if (rc) return rc;
  const long want = (long)p.total_groups * cl;
  const int grid = want < ctas ? (int)want : ctas;
  return block_n == 64 ? launch3<SCHED_T4, 64, 4, 12, true>(tx, tw, p, grid, stream)
                       : launch3<SCHED_T4, 128, 4, 6, false>(tx, tw, p, grid, stream);
}

// Strided conv (filter in the TensorFlow layout [25][Cin][Cout]) for layers whose 128-row tiles are whole rows of one
// image: the 25 taps read 10 halo tiles per channel chunk (5 kw x {even, odd kh}) instead of 25 tiles.
int igemm3_fconv(const void* x, const void* w_kn, const float* bias, void* out, int NB, int H, int W, int Cin, int Cout,
                 int act, int block_n, float* stats, int stat_groups, cudaStream_t stream, const BnRedArgs* bnred) {
  if (Cin % 64 || Cout % 64 || (H & 1) || (W & 1)) return DCGAN_ERR_SHAPE;
  if (bnred != nullptr && (stats != nullptr || bias != nullptr || act != 0)) return DCGAN_ERR_SHAPE;
  const int oH = H / 2, oW = W / 2;
  const int lh = ilog2_exact3(oH), lw = ilog2_exact3(oW);
  if (lh < 0 || lw < 0) return DCGAN_ERR_SHAPE;
  int bw, bh, bn_img;
  if (!igemm2_pixel_box(kBlockM, oH, oW, &bw, &bh, &bn_img)) return DCGAN_ERR_SHAPE;
  if (bn_img != 1 || oW % 8 || (bh + 2) * bw * 128 > kXSlotBytes) return DCGAN_ERR_SHAPE;  // no halo possible
  Igemm2Params p{};
  p.err_flag = error_flag_ptr();
  p.mn_variant = mn_variant_ptr();
  if (!p.err_flag || !p.mn_variant) return DCGAN_ERR_DRIVER;
  p.rH_log2 = lh; p.rW_log2 = lw;
  p.rows_total = NB * oH * oW;
  p.gC = Cin; p.Cn = Cout;
  p.b_mn = 1;
  p.out = out; p.bias = bias; p.act = act;
  p.oH = oH; p.oW = oW;
  if (stats != nullptr) {
    if (stat_groups < 1 || p.rows_total % stat_groups || (p.rows_total / stat_groups) % kBlockM) return DCGAN_ERR_SHAPE;
    p.stats = stats;
    p.stat_group_rows = p.rows_total / stat_groups;
  }
  if (bnred != nullptr) {
    if (bnred->slabs < 1 || NB % bnred->slabs || ((NB / bnred->slabs) * oH * oW) % kBlockM) return DCGAN_ERR_SHAPE;
    p.bn_pre = bnred->pre; p.bn_saved = bnred->saved; p.bn_gamma = bnred->gamma; p.bn_beta = bnred->beta;
    p.bn_red = bnred->red; p.bn_stat_groups = bnred->stat_groups; p.bn_act = bnred->act;
    p.stat_group_rows = (NB / bnred->slabs) * oH * oW;
    p.bn_slab_out_rows = p.stat_group_rows;
  }
  p.halo = 1;
  p.halo_row_bytes = oW * 128;
  p.x_tile_bytes = (bh + 2) * bw * 128;
  if (block_n != 64 && block_n != 128) block_n = (Cout % 128 == 0) ? 128 : 64;
  if (Cout % block_n) return DCGAN_ERR_SHAPE;
  const int cpt = Cin / 64;
  const int mt = p.rows_total / kBlockM;  // whole tiles: bn_img == 1 and oH % bh == 0
  p.m_tiles = mt;
  p.n_tiles = Cout / block_n;
  int cl = igemm2_cluster_pref() > 0 ? igemm2_cluster_pref() : 2;
  while (cl > 1 && cl > mt) cl >>= 1;
  p.cluster = cl;
  p.m_groups = (mt + cl - 1) / cl;
  int ctas = igemm2_max_ctas() > 0 ? igemm2_max_ctas() : igemm2_num_sms();
  if (ctas < cl) ctas = cl;
  ctas = ctas / cl * cl;
  const long tiles = (long)p.m_groups * cl * p.n_tiles;
  int ks = 1;
  if (igemm2_forced_ksplit() > 0) {
    ks = igemm2_forced_ksplit() < cpt ? igemm2_forced_ksplit() : cpt;
  } else {
    while (tiles * ks * 2 <= ctas && ks * 2 <= cpt) ks *= 2;
  }
  p.ksplit = ks;
  int rc = igemm2_attach_scratch(&p, tiles, block_n);
  if (rc) return rc;
  p.total_groups = (int)((long)p.m_groups * p.n_tiles * p.ksplit);
  CUtensorMap tx, tw;
  rc = igemm2_tmap_nhwc(x, NB, H, W, Cin, bw, bh + 2, 1, 2, &tx);
  if (rc) return rc;
  rc = igemm2_tmap_2d(w_kn, (unsigned long long)25 * Cin, Cout, 64 / cl, 64, &tw);
  if (rc) return rc;
  const long want = (long)p.total_groups * cl;
  const int grid = want < ctas ? (int)want : ctas;
  return block_n == 64 ? launch3<SCHED_FH, 64, 4, 12, true>(tx, tw, p, grid, stream)
                       : launch3<SCHED_FH, 128, 4, 6, true>(tx, tw, p, grid, stream);
}

}  // namespace dcgan
