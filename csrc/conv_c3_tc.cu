// Tensor-core versions of the 3-channel strided layer and its filter gradient.
//
// d_h0_conv (64x64x3 -> 32x32x64, distriubted_model.py:118) and the data gradient of g_h4 have
// K = 5*5*3 = 75, far too skinny for the generic implicit-GEMM pipeline, and on the CUDA cores they
// are bound by FP32 issue rate (profiles/launches_step_v1.csv: 75 us and 77 us).  Here the im2col
// operand is built in shared memory, once per 128-pixel tile, straight into the 128B-swizzled UMMA
// layout with K padded to 5 x 16 (each filter row kh = 15 taps*channels + 1 zero), and fed to
// tcgen05.mma with the accumulator in TMEM:
//
//   c3_fprop_tc : out[128 px][64 co] = im2col[128 px][80] . W[64 co][80]^T      (K-major A and B)
//   c3_wgrad_tc : dW[80][64 co]     += im2col[px][80]^T   . dy[px][64 co]      (MN-major A and B, K = pixels,
//                 accumulated in TMEM over all tiles of the CTA, one f32 atomic pass at the end)
//
// Both then run at memory speed.  Filter layout is the shared [25][3][C] of csrc/conv_c3.cu.
#include "common.cuh"
#include "igemm2.h"
#include "../include/dcgan_b200.h"

namespace dcgan {

constexpr int kTcThreads = 128;
constexpr int kPatchMaxElems = 4096;       // bf16; covers tiles of 4x32 (2288), 2x64 (2800), 1x128 (3920) pixels

// 16-byte chunk (row, chunk) of a [rows][128 B] tile in the SWIZZLE_128B pattern
__device__ __forceinline__ uint32_t swz(uint32_t row, uint32_t chunk) {
  return row * 128u + ((chunk ^ (row & 7u)) << 4);
}

struct TileGeom {
  int W_out, rows_per_tile, patch_h, patch_w3, patch_pitch;  // patch_w3 = patch width * 3, pitch in bf16
};
__device__ __forceinline__ TileGeom make_geom(int W_out) {
  TileGeom g;
  g.W_out = W_out;
  g.rows_per_tile = 128 / W_out;
  g.patch_h = 2 * g.rows_per_tile + 3;
  g.patch_w3 = (2 * W_out + 3) * 3;
  g.patch_pitch = (g.patch_w3 + 7) & ~7;
  return g;
}

// input patch (3 channels) of one tile -> smem bf16.  A tile spans full image rows, so the patch columns of
// pixel -1 and of pixels >= W are always padding (TF SAME: 1 before, 2 after): they are zeroed once per
// kernel by zero_patch(); per tile only the in-image part of each row is copied, as 32-bit words with all
// loads of a thread in flight before the first store.
__device__ __forceinline__ void zero_patch(__nv_bfloat16* patch) {
  for (int i = threadIdx.x; i < kPatchMaxElems / 8; i += kTcThreads)
    reinterpret_cast<uint4*>(patch)[i] = make_uint4(0, 0, 0, 0);
}

constexpr int kPatchWordsPerThread = 15;  // 5 rows * 384 words / 128 threads (the 1x128 tile)

// Word i = threadIdx.x + k*128 of a tile's patch: rows of one image are contiguous in memory, so the source is
// simply word i after the first patch row and only the row bounds need checking; the destination (row pitch
// differs from the image pitch) needs i / words_per_row, which depends on the thread alone and is computed once
// per kernel (PatchMap) instead of per tile -- ncu showed these kernels bound by instruction issue, ~900 of the
// ~1900 instructions per warp and tile being these divisions.
struct PatchMap {
  int dst[kPatchWordsPerThread];  // element index of the word's low half in the smem patch, -1 = beyond the patch
};
__device__ __forceinline__ PatchMap make_patch_map(int W, const TileGeom& g) {
  PatchMap m;
  const int words_per_row = (3 * W) >> 1;       // 32-bit words of one image row (W even)
  const int total = g.patch_h * words_per_row;  // <= 11 * 96 = 1056, 7 * 192 = 1344, 5 * 384 = 1920
#pragma unroll
  for (int k = 0; k < kPatchWordsPerThread; ++k) {
    const int i = threadIdx.x + k * kTcThreads;
    m.dst[k] = -1;
    if (i < total) {
      const int r = i / words_per_row, wcol = i - r * words_per_row;
      // pixel 0 sits at element 3 of the patch row (element 0..2 = pixel -1): odd offset -> two 16-bit stores
      m.dst[k] = r * g.patch_pitch + 3 + 2 * wcol;
    }
  }
  return m;
}

__device__ __forceinline__ void fetch_patch(const __nv_bfloat16* __restrict__ x, uint32_t (&v)[kPatchWordsPerThread],
                                            int n, int oy0, int H, int W, const TileGeom& g) {
  const int iy0 = 2 * oy0 - 1;
  const int words_per_row = (3 * W) >> 1;
  const int lo = (iy0 < 0 ? -iy0 : 0) * words_per_row;                              // rows above the image
  const int hi = (g.patch_h < H - iy0 ? g.patch_h : H - iy0) * words_per_row;      // rows below it / patch end
  const long first = ((long)n * H + iy0) * words_per_row;                          // word index of patch row 0
  const uint32_t* xw = reinterpret_cast<const uint32_t*>(x);
#pragma unroll
  for (int k = 0; k < kPatchWordsPerThread; ++k) {
    const int i = threadIdx.x + k * kTcThreads;
    v[k] = (i >= lo && i < hi) ? __ldg(xw + (first + i)) : 0u;
  }
}

__device__ __forceinline__ void commit_patch(__nv_bfloat16* patch, const uint32_t (&v)[kPatchWordsPerThread],
                                             const PatchMap& m) {
#pragma unroll
  for (int k = 0; k < kPatchWordsPerThread; ++k) {
    if (m.dst[k] >= 0) {
      unsigned short* dst = reinterpret_cast<unsigned short*>(patch) + m.dst[k];
      dst[0] = (unsigned short)(v[k] & 0xffffu);
      dst[1] = (unsigned short)(v[k] >> 16);
    }
  }
}

// im2col row of pixel `t` of the tile: chunks 0..9 (k = kh*16 + kw*3 + ci) of its 256-byte row, split over the
// two k-block sub-tiles.  Chunks 10..15 stay zero (written once at kernel start).
__device__ __forceinline__ void build_im2col_row(uint8_t* a_tiles, const __nv_bfloat16* patch, int t, const TileGeom& g) {
  const int ty = t / g.W_out, tx = t - ty * g.W_out;
#pragma unroll
  for (int kh = 0; kh < 5; ++kh) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(patch + (2 * ty + kh) * g.patch_pitch + 6 * tx);
    uint32_t w[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = src[j];  // 16 bf16: 15 taps*channels + 1 neighbour element
    w[7] &= 0x0000ffffu;                        // k = kh*16 + 15 is padding
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int c = kh * 2 + half;              // chunk index 0..9 within the 16-chunk row
      uint8_t* dst = a_tiles + (c >> 3) * (128 * 128) + swz(t, c & 7);
      *reinterpret_cast<uint4*>(dst) = make_uint4(w[4 * half], w[4 * half + 1], w[4 * half + 2], w[4 * half + 3]);
    }
  }
}

__device__ __forceinline__ void zero_im2col_padding(uint8_t* a_tiles, int t) {
#pragma unroll
  for (int c = 10; c < 16; ++c)
    *reinterpret_cast<uint4*>(a_tiles + (c >> 3) * (128 * 128) + swz(t, c & 7)) = make_uint4(0, 0, 0, 0);
}

// ---------------------------------------------------------------------------------------------------------
// forward: grid = (CTAs, C/64).  smem: A 32 KiB | B 16 KiB | patch | barrier
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTcThreads)
c3_fprop_tc_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                   __nv_bfloat16* __restrict__ out, int N, int H, int W, int C, int act, int* err_flag) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* a_tiles = smem;                       // 2 x [128][128 B]
  uint8_t* b_tiles = smem + 32768;               // 2 x [64][128 B]
  __nv_bfloat16* patch = reinterpret_cast<__nv_bfloat16*>(smem + 32768 + 16384);
  uint64_t* mma_bar = reinterpret_cast<uint64_t*>(smem + 32768 + 16384 + kPatchMaxElems * 2);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(mma_bar + 1);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);

  const int t = threadIdx.x, warp = t >> 5;
  const int oH = H / 2, oW = W / 2;
  const TileGeom g = make_geom(oW);
  const int c0 = blockIdx.y * 64;

  if (t == 0) { mbar_init(mma_bar, 1); *abort_flag = 0; fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(tmem_ptr_smem, 64); tmem_relinquish(); }
  // B operand: W[co][k] with k = kh*16 + kw*3 + ci, bf16, K-major SW128, zero padded
  for (int i = t; i < 16384 / 16; i += kTcThreads) reinterpret_cast<uint4*>(b_tiles)[i] = make_uint4(0, 0, 0, 0);
  zero_im2col_padding(a_tiles, t);
  zero_patch(patch);
  pdl_wait();
  {
    // thread -> output channel co = t & 63 and every other filter row; all 38 loads in flight first
    const int co = t & 63, r0 = t >> 6;
    float wv[38];
#pragma unroll
    for (int j = 0; j < 38; ++j) {
      const int row = r0 + 2 * j;                   // row = (kh*5+kw)*3+ci = kh*15 + kk
      wv[j] = row < 75 ? __ldg(w + (size_t)row * C + c0 + co) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 38; ++j) {
      const int row = r0 + 2 * j;
      if (row < 75) {
        const int kh = row / 15, kk = row - kh * 15;
        const int k = kh * 16 + kk;
        uint8_t* dst = b_tiles + (k >> 6) * (64 * 128) + swz(co, (k & 63) >> 3) + (k & 7) * 2;
        *reinterpret_cast<__nv_bfloat16*>(dst) = __float2bfloat16(wv[j]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t a_addr = smem_u32(a_tiles), b_addr = smem_u32(b_tiles);
  constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);

  const int tiles_per_img = oH * oW / 128;
  const int tiles = N * tiles_per_img;
  uint32_t phase = 0;
  const PatchMap pmap = make_patch_map(W, g);
  uint32_t pv[kPatchWordsPerThread];  // the next tile's patch, in flight while this tile computes
  if ((int)blockIdx.x < tiles) {
    const int n = blockIdx.x / tiles_per_img;
    fetch_patch(x, pv, n, (blockIdx.x - n * tiles_per_img) * g.rows_per_tile, H, W, g);
  }
  // per-channel bias of this thread's epilogue columns is loop invariant: keep it in shared memory
  float* bias_s = reinterpret_cast<float*>(smem + 32768 + 16384 + kPatchMaxElems * 2 + 64);
  if (t < 64) bias_s[t] = bias != nullptr ? bias[c0 + t] : 0.f;
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int n = tile / tiles_per_img;
    const int oy0 = (tile - n * tiles_per_img) * g.rows_per_tile;
    commit_patch(patch, pv, pmap);
    __syncthreads();
    {
      const int nt = tile + gridDim.x;
      if (nt < tiles) {
        const int nn = nt / tiles_per_img;
        fetch_patch(x, pv, nn, (nt - nn * tiles_per_img) * g.rows_per_tile, H, W, g);
      }
    }
    build_im2col_row(a_tiles, patch, t, g);
    fence_proxy_async_smem();
    __syncthreads();
    if (t == 0) {
      tc_fence_after();
#pragma unroll
      for (int k = 0; k < 5; ++k) {  // K = 80 = 4 steps in k-block 0 + 1 step in k-block 1
        const uint32_t off = (k < 4) ? k * 32u : 0u;
        const uint64_t da = make_smem_desc_sw128(a_addr + (k >> 2) * 16384 + off, 16, 1024);
        const uint64_t db = make_smem_desc_sw128(b_addr + (k >> 2) * 8192 + off, 16, 1024);
        umma_bf16(tmem_base, da, db, idesc, k != 0 ? 1u : 0u);
      }
      umma_commit(mma_bar);
    }
    const bool ok = mbar_wait(mma_bar, phase, abort_flag, err_flag);
    phase ^= 1;
    tc_fence_after();
    // epilogue: thread t = pixel t of the tile = TMEM lane t
    const int ty = t / oW, tx = t - ty * oW;
    __nv_bfloat16* orow = out + (((size_t)n * oH + oy0 + ty) * oW + tx) * C + c0;
#pragma unroll
    for (int c = 0; c < 64; c += 32) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + c, v);
      tmem_ld_wait();
      if (ok) {
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 b = *reinterpret_cast<const float4*>(bias_s + c + j);
          f[j] = __uint_as_float(v[j]) + b.x; f[j + 1] = __uint_as_float(v[j + 1]) + b.y;
          f[j + 2] = __uint_as_float(v[j + 2]) + b.z; f[j + 3] = __uint_as_float(v[j + 3]) + b.w;
        }
        apply_act_chunk(f, act);  // kind selected once per chunk (a per-element switch multiplies the code size)
        uint32_t packed[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) packed[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
        uint4* dst = reinterpret_cast<uint4*>(orow + c);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          dst[j] = make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
      }
    }
    tc_fence_before();
    __syncthreads();  // TMEM drained and the A tile consumed before the next tile overwrites them
  }
  if (warp == 0) tmem_dealloc(tmem_base, 64);
}

// ---------------------------------------------------------------------------------------------------------
// filter gradient: grid = (CTAs, C/64).  smem: A 32 KiB | B 16 KiB (dy tile) | patch | barrier
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTcThreads)
c3_wgrad_tc_kernel(const __nv_bfloat16* __restrict__ x3, const __nv_bfloat16* __restrict__ y, float* __restrict__ dw,
                   float* __restrict__ partial, int N, int Hs, int Ws, int C, int* err_flag) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* a_tiles = smem;                       // 2 x [128 px][128 B]  (MN-major A: M = k index)
  uint8_t* b_tile = smem + 32768;                // [128 px][128 B]      (MN-major B: N = 64 channels)
  __nv_bfloat16* patch = reinterpret_cast<__nv_bfloat16*>(smem + 32768 + 16384);
  uint64_t* mma_bar = reinterpret_cast<uint64_t*>(smem + 32768 + 16384 + kPatchMaxElems * 2);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(mma_bar + 1);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);

  const int t = threadIdx.x, warp = t >> 5;
  const int H = 2 * Hs, W = 2 * Ws;
  const TileGeom g = make_geom(Ws);
  const int c0 = blockIdx.y * 64;

  if (t == 0) { mbar_init(mma_bar, 1); *abort_flag = 0; fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(tmem_ptr_smem, 64); tmem_relinquish(); }
  zero_im2col_padding(a_tiles, t);
  zero_patch(patch);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t a_addr = smem_u32(a_tiles), b_addr = smem_u32(b_tile);
  constexpr uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);

  const int tiles_per_img = Hs * Ws / 128;
  const int tiles = N * tiles_per_img;
  uint32_t phase = 0;
  bool any = false, ok = true;
  const PatchMap pmap = make_patch_map(W, g);
  uint32_t pv[kPatchWordsPerThread];  // next tile's patch and dy row, fetched while this tile's MMAs run
  uint4 yv[8];
  auto fetch = [&](int tile) {
    const int n = tile / tiles_per_img;
    const int qy0 = (tile - n * tiles_per_img) * g.rows_per_tile;
    const uint4* src = reinterpret_cast<const uint4*>(y + ((size_t)(n * Hs + qy0) * Ws + t) * C + c0);
#pragma unroll
    for (int j = 0; j < 8; ++j) yv[j] = __ldg(src + j);
    fetch_patch(x3, pv, n, qy0, H, W, g);
  };
  if ((int)blockIdx.x < tiles) fetch(blockIdx.x);
  for (int tile = blockIdx.x; tile < tiles && ok; tile += gridDim.x) {
    if (any) {  // the previous tile's MMAs must retire before its operands are overwritten
      ok = mbar_wait(mma_bar, phase, abort_flag, err_flag);
      phase ^= 1;
    }
    // dy rows of this tile: pixel t -> 128 contiguous bytes, swizzled into the B tile
#pragma unroll
    for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(b_tile + swz(t, j)) = yv[j];
    commit_patch(patch, pv, pmap);
    __syncthreads();
    if (tile + (int)gridDim.x < tiles) fetch(tile + gridDim.x);
    build_im2col_row(a_tiles, patch, t, g);
    fence_proxy_async_smem();
    __syncthreads();
    if (t == 0) {
      tc_fence_after();
#pragma unroll
      for (int k = 0; k < 8; ++k) {  // K = 128 pixels of the tile, 16 per instruction
        const uint64_t da = make_smem_desc_sw128(a_addr + k * 2048, 16384, 1024);  // 2 chunks of 64 k-indices
        const uint64_t db = make_smem_desc_sw128(b_addr + k * 2048, 16384, 1024);  // one chunk of 64 channels
        umma_bf16(tmem_base, da, db, idesc, (any || k != 0) ? 1u : 0u);
      }
      umma_commit(mma_bar);
    }
    any = true;
  }
  if (any && ok) ok = mbar_wait(mma_bar, phase, abort_flag, err_flag);
  tc_fence_after();
  // epilogue: TMEM lane m = padded k index kh*16 + kk; rows with kk == 15 or m >= 80 are padding
  const int m = t;
  const int kh = m >> 4, kk = m & 15;
  const bool real = m < 80 && kk < 15;
  float* orow = dw + (size_t)(kh * 15 + kk) * C + c0;
  // with `partial`: this CTA's [75][64] tile goes to scratch (plain stores, zero if it had no tile) and
  // c3_wgrad_reduce_kernel adds the tiles up in a fixed order -- no atomics, reproducible
  float* prow = partial != nullptr
                    ? partial + ((size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 75 + (kh * 15 + kk)) * 64
                    : nullptr;
#pragma unroll
  for (int c = 0; c < 64; c += 32) {
    uint32_t v[32];
    tmem_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + c, v);
    tmem_ld_wait();
    if (real && prow != nullptr) {
      const bool have = any && ok;
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(prow + c + j) =
            have ? make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                               __uint_as_float(v[j + 3]))
                 : make_float4(0.f, 0.f, 0.f, 0.f);
    } else if (real && any && ok) {
#pragma unroll
      for (int j = 0; j < 32; ++j) atomicAdd(orow + c + j, __uint_as_float(v[j]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 64);
}


// dw[row][c0 + col] += sum over the CTAs' partial tiles, in CTA order.  One warp = 32 consecutive outputs; the 8
// warps of a CTA take the partials b = warp, warp + 8, ... and are combined through shared memory in warp order.
// grid = (75 * 64 / 32, C / 64), 256 threads.
__global__ void __launch_bounds__(256)
c3_wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int nparts, int C) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float acc_s[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int o = blockIdx.x * 32 + lane;  // element of the [75][64] tile
  const float* src = partial + (size_t)blockIdx.y * nparts * 4800 + o;
  float acc = 0.f;
  int b = warp;
  for (; b + 56 < nparts; b += 64) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldcg(src + (size_t)(b + 8 * u) * 4800);
#pragma unroll
    for (int u = 0; u < 8; ++u) acc += v[u];
  }
  for (; b < nparts; b += 8) acc += __ldcg(src + (size_t)b * 4800);
  acc_s[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += acc_s[w][lane];
    const int row = o >> 6, col = o & 63;
    dw[(size_t)row * C + blockIdx.y * 64 + col] += t;
  }
}

// ---------------------------------------------------------------------------------------------------------
// C -> 3 transposed conv on tensor cores (g_h4 forward, distriubted_model.py:109; d_h0 data gradient).
//   step 1 (tcgen05): Y[q][tap*3+co] = sum_ci x[q][ci] * W[tap][co][ci]       -- a plain GEMM, no im2col:
//           M = input locations (an 8x16 tile plus a one-pixel halo = 180 rows in two 128-row MMAs),
//           N = 80 (75 tap/channel pairs + padding), K = C channels in KC = C / 64 chunks (C = 64, or 128 for the
//           wide configuration).
//   step 2 (col2im in shared memory): out[2q+p][co] = sum over the taps of parity p of Y[q + (p+1-k)/2][tap][co],
//           then bias+tanh, or the tanh backward against `aux`, and the optional per-channel sum.
// grid = (tiles, 1); C = 64 * KC, KC in {1, 2}.
// ---------------------------------------------------------------------------------------------------------
constexpr int kTH2 = 8, kTW2 = 16;                       // interior tile (input locations)
constexpr int kHaloW = kTW2 + 2, kHaloRows = (kTH2 + 2) * (kTW2 + 2);  // 18, 180
constexpr int kYPitch = 77;                              // f32 per Y row (75 used; odd pitch -> conflict-free)
template <int KC>
constexpr int tconv_tc_smem() {
  return KC * 32768 /*A*/ + KC * 128 * 128 /*B (80 rows used)*/ + kHaloRows * kYPitch * 4 + 64 + 1024;
}

template <int KC>
__global__ void __launch_bounds__(kTcThreads)
c3_tconv_tc_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w_bf16 /*[75][C]*/,
                   const float* __restrict__ bias, const __nv_bfloat16* __restrict__ aux,
                   __nv_bfloat16* __restrict__ out, float* __restrict__ colsum, int N, int H, int W, int act,
                   int mode, int* err_flag) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  constexpr int C = 64 * KC;
  uint8_t* a_tiles = smem;                       // KC x 2 x [128 rows][128 B]: halo rows 0..179 (+ zero rows)
  uint8_t* b_tile = smem + KC * 32768;           // KC x [80 rows (tap*3+co)][128 B]
  float* ysm = reinterpret_cast<float*>(smem + KC * (32768 + 16384));
  uint64_t* mma_bar = reinterpret_cast<uint64_t*>(smem + KC * (32768 + 16384) + kHaloRows * kYPitch * 4);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(mma_bar + 1);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  __shared__ float cred[3];

  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) { mbar_init(mma_bar, 1); *abort_flag = 0; fence_mbar_init(); }
  if (t < 3) cred[t] = 0.f;
  if (warp == 0) { tmem_alloc(tmem_ptr_smem, 256); tmem_relinquish(); }
  pdl_wait();
  // B: filter rows (tap*3+co) x C channels, K-major SW128, one tile per 64-channel chunk; rows 75..79 zero
  for (int i = t; i < KC * 80 * 8; i += kTcThreads) {
    const int kc = i / (80 * 8), r8 = i - kc * (80 * 8);
    const int row = r8 >> 3, ch = r8 & 7;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (row < 75) v = __ldg(reinterpret_cast<const uint4*>(w_bf16 + (size_t)row * C + kc * 64) + ch);
    *reinterpret_cast<uint4*>(b_tile + kc * 16384 + swz(row, ch)) = v;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t a_addr = smem_u32(a_tiles), b_addr = smem_u32(b_tile);
  constexpr uint32_t idesc = make_idesc_bf16(128, 80, 0, 0);

  const int tiles_x = W / kTW2, tiles_y = H / kTH2;
  const int tiles = N * tiles_y * tiles_x;
  const int oH = 2 * H, oW = 2 * W;
  uint32_t phase = 0;
  float csum[3] = {0.f, 0.f, 0.f};
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int n = tile / (tiles_y * tiles_x);
    const int tr = tile - n * tiles_y * tiles_x;
    const int qy0 = (tr / tiles_x) * kTH2, qx0 = (tr % tiles_x) * kTW2;
    // ---- gather the halo tile: 256 rows x 8 chunks, zero outside the image / beyond row 179 -----------------
#pragma unroll 1
    for (int kc = 0; kc < KC; ++kc) {
      uint4 v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int idx = t + k * kTcThreads;        // 0..2047 = row * 8 + chunk
        const int row = idx >> 3, ch = idx & 7;
        v[k] = make_uint4(0, 0, 0, 0);
        if (row < kHaloRows) {
          const int ly = row / kHaloW, lx = row - ly * kHaloW;
          const int qy = qy0 - 1 + ly, qx = qx0 - 1 + lx;
          if ((unsigned)qy < (unsigned)H && (unsigned)qx < (unsigned)W)
            v[k] = __ldg(reinterpret_cast<const uint4*>(x + (((size_t)n * H + qy) * W + qx) * C + kc * 64) + ch);
        }
      }
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int idx = t + k * kTcThreads;
        const int row = idx >> 3, ch = idx & 7;
        *reinterpret_cast<uint4*>(a_tiles + kc * 32768 + (row >> 7) * 16384 + swz(row & 127, ch)) = v[k];
      }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (t == 0) {
      tc_fence_after();
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int kc = 0; kc < KC; ++kc)
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t da = make_smem_desc_sw128(a_addr + kc * 32768 + j * 16384 + k * 32, 16, 1024);
            const uint64_t db = make_smem_desc_sw128(b_addr + kc * 16384 + k * 32, 16, 1024);
            umma_bf16(tmem_base + j * 128, da, db, idesc, (kc | k) != 0 ? 1u : 0u);
          }
      umma_commit(mma_bar);
    }
    const bool ok = mbar_wait(mma_bar, phase, abort_flag, err_flag);
    phase ^= 1;
    tc_fence_after();
    // ---- TMEM -> Y in shared memory ----------------------------------------------------------------------------
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int row = j * 128 + t;
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + j * 128;
      uint32_t v0[32], v1[32], v2[16];
      tmem_ld_32x32(taddr, v0);
      tmem_ld_32x32(taddr + 32, v1);
      tmem_ld_32x16(taddr + 64, v2);
      tmem_ld_wait();
      if (row < kHaloRows) {
        float* yr = ysm + row * kYPitch;
#pragma unroll
        for (int c = 0; c < 32; ++c) { yr[c] = __uint_as_float(v0[c]); yr[32 + c] = __uint_as_float(v1[c]); }
#pragma unroll
        for (int c = 0; c < 11; ++c) yr[64 + c] = __uint_as_float(v2[c]);
      }
    }
    tc_fence_before();
    __syncthreads();
    // ---- col2im: thread = interior location (ty, tx) -> its 2x2 output quad ---------------------------------
    if (ok) {
      const int ty = t / kTW2, tx = t - ty * kTW2;
      const int qy = qy0 + ty, qx = qx0 + tx;
#pragma unroll
      for (int ph = 0; ph < 2; ++ph) {
        float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int pw = 0; pw < 2; ++pw)
#pragma unroll
          for (int kh = 1 - ph; kh < 5; kh += 2)      // taps of this row parity: kh = ph+1 (mod 2)
#pragma unroll
            for (int kw = 1 - pw; kw < 5; kw += 2) {
              const int dy = (ph + 1 - kh) / 2, dx = (pw + 1 - kw) / 2;
              const float* yr = ysm + ((ty + 1 + dy) * kHaloW + (tx + 1 + dx)) * kYPitch + (kh * 5 + kw) * 3;
              v[pw * 3 + 0] += yr[0]; v[pw * 3 + 1] += yr[1]; v[pw * 3 + 2] += yr[2];
            }
        const size_t o = (((size_t)n * oH + 2 * qy + ph) * oW + 2 * qx) * 3;  // 6 contiguous bf16
        if (mode == 0) {
#pragma unroll
          for (int j = 0; j < 6; ++j) v[j] = apply_act(v[j] + (bias ? bias[j % 3] : 0.f), act);
        } else {
          const uint32_t* ap = reinterpret_cast<const uint32_t*>(aux + o);
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            const float2 a = unpack_bf16x2(ap[j]);
            v[2 * j] *= (1.f - a.x * a.x);
            v[2 * j + 1] *= (1.f - a.y * a.y);
          }
        }
#pragma unroll
        for (int j = 0; j < 6; ++j) csum[j % 3] += v[j];
        uint32_t* op = reinterpret_cast<uint32_t*>(out + o);
#pragma unroll
        for (int j = 0; j < 3; ++j) op[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
      }
    }
    __syncthreads();  // Y and the A tiles are free for the next tile
  }
  if (colsum != nullptr) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float s = warp_sum(csum[c]);
      if ((t & 31) == 0) atomicAdd(&cred[c], s);
    }
    __syncthreads();
    if (t < 3) atomicAdd(&colsum[t], cred[t]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 256);
}

constexpr int kTcSmem = 32768 + 16384 + kPatchMaxElems * 2 + 64 + 256 + 1024;

}  // namespace dcgan

using namespace dcgan;

static int g_c3_blocks_fprop = 0, g_c3_blocks_wgrad = 0;  // 0 = default; profiling overrides

template <int KC>
static int launch_tconv_tc(const void* x, const void* w_bf16, const float* bias, const void* aux, void* out,
                           float* colsum, int N, int H, int W, int act, int mode, cudaStream_t st) {
  static bool attr_done[64] = {false};  // per device: the attribute belongs to the device's copy of the kernel
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DCGAN_ERR_DRIVER;
  if (!attr_done[dev]) {
    if (cudaFuncSetAttribute(c3_tconv_tc_kernel<KC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             tconv_tc_smem<KC>()) != cudaSuccess)
      return DCGAN_ERR_LAUNCH;
    attr_done[dev] = true;
  }
  int* err = error_flag_ptr();
  if (!err) return DCGAN_ERR_DRIVER;
  const int tiles = N * (H / kTH2) * (W / kTW2);
  const int cap = KC == 1 ? 296 : 148;  // two CTAs per SM fit with 64 channels (106 KB), one with 128 (154 KB)
  dim3 grid(tiles < cap ? tiles : cap);
  launch_k(c3_tconv_tc_kernel<KC>, dim3(grid), dim3(kTcThreads), tconv_tc_smem<KC>(), st,
           reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(w_bf16), bias,
           reinterpret_cast<const __nv_bfloat16*>(aux), reinterpret_cast<__nv_bfloat16*>(out), colsum, N, H, W, act, mode,
           err);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

extern "C" {

void dcgan_set_c3_tuning(int fprop_blocks, int wgrad_blocks) {
  g_c3_blocks_fprop = fprop_blocks;
  g_c3_blocks_wgrad = wgrad_blocks;
}

int dcgan_c3_fprop_tc(const void* x, const float* w, const float* bias, void* out, int N, int H, int W, int C,
                      int act, cudaStream_t st) {
  const int oH = H / 2, oW = W / 2;
  if (C % 64 || (H & 1) || (W & 1) || (oW != 32 && oW != 64 && oW != 128) || (oH * oW) % 128) return DCGAN_ERR_SHAPE;
  static bool attr_done[64] = {false};  // per device: the attribute belongs to the device's copy of the kernel
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DCGAN_ERR_DRIVER;
  if (!attr_done[dev]) {
    if (cudaFuncSetAttribute(c3_fprop_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmem) != cudaSuccess)
      return DCGAN_ERR_LAUNCH;
    attr_done[dev] = true;
  }
  int* err = error_flag_ptr();
  if (!err) return DCGAN_ERR_DRIVER;
  const int tiles = N * oH * oW / 128;
  const int cap_f = g_c3_blocks_fprop > 0 ? g_c3_blocks_fprop : 148 * 3;  // in-step, chain at high priority: 444 CTAs 0.8112, 296 0.8134
  dim3 grid(tiles < cap_f ? tiles : cap_f, C / 64);
  launch_k(c3_fprop_tc_kernel, dim3(grid), dim3(kTcThreads), kTcSmem, st, reinterpret_cast<const __nv_bfloat16*>(x), w, bias,
                                                        reinterpret_cast<__nv_bfloat16*>(out), N, H, W, C, act, err);
  return cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

int dcgan_c3_tconv_tc(const void* x, const void* w_bf16, const float* bias, const void* aux, void* out, float* colsum,
                      int N, int H, int W, int C, int act, int mode, cudaStream_t st) {
  if ((C != 64 && C != 128) || H % kTH2 || W % kTW2) return DCGAN_ERR_SHAPE;
  return C == 64 ? launch_tconv_tc<1>(x, w_bf16, bias, aux, out, colsum, N, H, W, act, mode, st)
                 : launch_tconv_tc<2>(x, w_bf16, bias, aux, out, colsum, N, H, W, act, mode, st);
}

int dcgan_c3_wgrad_tc(const void* x3, const void* y, float* dw, int N, int Hs, int Ws, int C, cudaStream_t st) {
  if (C % 64 || (Ws != 32 && Ws != 64 && Ws != 128) || (Hs * Ws) % 128) return DCGAN_ERR_SHAPE;
  static bool attr_done[64] = {false};  // per device: the attribute belongs to the device's copy of the kernel
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DCGAN_ERR_DRIVER;
  if (!attr_done[dev]) {
    if (cudaFuncSetAttribute(c3_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmem) != cudaSuccess)
      return DCGAN_ERR_LAUNCH;
    attr_done[dev] = true;
  }
  int* err = error_flag_ptr();
  if (!err) return DCGAN_ERR_DRIVER;
  const int tiles = N * Hs * Ws / 128;
  // Few CTAs: every CTA ends with 75 x 64 f32 atomics on the same addresses, and in the training step this kernel
  // runs on the side stream beside the data-gradient chain.  In-step sweeps (profiles/README.md): without stream
  // priorities 444 CTAs 0.931 ms/step, 148 0.874, 64 0.820, 24 0.921; with the chain captured at high priority
  // (engine default) 444 0.895, 296 0.849, 222 0.834, 148 0.811-0.813, 64 0.821.
  // (those sweeps predate the partial-tile path below, which has no atomics: more CTAs now only cost scratch)
  const int cap_w = g_c3_blocks_wgrad > 0 ? g_c3_blocks_wgrad : 296;
  dim3 grid(tiles < cap_w ? tiles : cap_w, C / 64);
  void* scratch = nullptr;
  size_t scratch_bytes = 0;
  const size_t need = (size_t)grid.x * grid.y * 4800 * sizeof(float);
  float* partial = nullptr;
  if (igemm2_scratch(&scratch, &scratch_bytes) == 0 && scratch_bytes >= need) partial = reinterpret_cast<float*>(scratch);
  launch_k(c3_wgrad_tc_kernel, dim3(grid), dim3(kTcThreads), kTcSmem, st, reinterpret_cast<const __nv_bfloat16*>(x3),
           reinterpret_cast<const __nv_bfloat16*>(y), dw, partial, N, Hs, Ws, C, err);
  if (cudaGetLastError() != cudaSuccess) return DCGAN_ERR_LAUNCH;
  if (partial != nullptr) {
    launch_k(c3_wgrad_reduce_kernel, dim3(4800 / 32, C / 64), dim3(256), 0, st, (const float*)partial, dw, (int)grid.x, C);
    if (cudaGetLastError() != cudaSuccess) return DCGAN_ERR_LAUNCH;
  }
  return 0;
}

}  // extern "C"
