// The two 3-channel layers (d_h0_conv: 64x64x3 -> 32x32x64, distriubted_model.py:118; g_h4:
// 32x32x64 -> 64x64x3, :109) and their gradients.  With K = 75 or N = 3 they are memory-bound and
// a 128-wide tensor-core tile would be >90% padding, so they run as direct
// convolutions on the CUDA cores with the filter and an input patch staged in shared memory.
//
// Both layers share one filter memory layout, [25 taps][3][C]: d_h0_conv/w is HWIO = [5,5,3,64]
// and g_h4/w is TF's deconv layout [5,5,Cout=3,Cin=64].  So three kernels cover all six uses:
//   c3_fprop  3 -> C strided conv      : d_h0 forward,  g_h4 data-gradient
//   c3_tconv  C -> 3 transposed conv   : g_h4 forward,  d_h0 data-gradient
//   c3_wgrad  dW[25][3][C]             : d_h0 and g_h4 filter gradients
#include "common.cuh"
#include "../include/dcgan_b200.h"

namespace dcgan {

constexpr int kC3Threads = 256;
constexpr int kTH = 8, kTW = 32;                 // output tile of c3_fprop / q tile of c3_tconv
constexpr int kPatchH = 2 * kTH + 3;              // 19
constexpr int kPatchW = (2 * kTW + 3) * 3;        // 67 pixels * 3 channels = 201
constexpr int kPatchWPad = kPatchW + 1;

// ---------------------------------------------------------------------------------------------
// c3_fprop: out[n,oy,ox,co] = act(bias[co] + sum_{kh,kw,ci} x[n,2oy+kh-1,2ox+kw-1,ci] * w[kh,kw,ci,co])
// grid = (tiles per image, N, C/64); thread = one output pixel x 64 channels
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kC3Threads)
c3_fprop_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                __nv_bfloat16* __restrict__ out, int H, int W, int C, int act) {
  extern __shared__ float smem[];
  float* ws = smem;                         // [75][64]
  float* patch = smem + 75 * 64;            // [kPatchH][kPatchWPad]
  const int oH = H / 2, oW = W / 2;
  const int tiles_x = oW / kTW;
  const int ty0 = (blockIdx.x / tiles_x) * kTH, tx0 = (blockIdx.x % tiles_x) * kTW;
  const int n = blockIdx.y;
  const int c0 = blockIdx.z * 64;
  for (int i = threadIdx.x; i < 75 * 64; i += kC3Threads) ws[i] = w[(size_t)(i / 64) * C + c0 + (i % 64)];
  const int iy0 = 2 * ty0 - 1, ix0 = 2 * tx0 - 1;
  for (int i = threadIdx.x; i < kPatchH * kPatchW; i += kC3Threads) {
    const int r = i / kPatchW, cc = i % kPatchW;
    const int iy = iy0 + r, ix = ix0 + cc / 3;
    float v = 0.f;
    if ((unsigned)iy < (unsigned)H && (unsigned)ix < (unsigned)W)
      v = __bfloat162float(x[(((size_t)n * H + iy) * W + ix) * 3 + cc % 3]);
    patch[r * kPatchWPad + cc] = v;
  }
  __syncthreads();
  const int ty = threadIdx.x / kTW, tx = threadIdx.x % kTW;
  float acc[64];
#pragma unroll
  for (int j = 0; j < 64; ++j) acc[j] = bias ? bias[c0 + j] : 0.f;
#pragma unroll 1
  for (int kh = 0; kh < 5; ++kh) {
    const float* prow = patch + (2 * ty + kh) * kPatchWPad + 2 * tx * 3;
#pragma unroll
    for (int kk = 0; kk < 15; ++kk) {  // (kw, ci) pairs are contiguous in the patch row
      const float xv = prow[kk];
      const float4* wr = reinterpret_cast<const float4*>(ws + (kh * 15 + kk) * 64);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float4 wv = wr[j];
        acc[4 * j] += xv * wv.x; acc[4 * j + 1] += xv * wv.y;
        acc[4 * j + 2] += xv * wv.z; acc[4 * j + 3] += xv * wv.w;
      }
    }
  }
  __nv_bfloat16* op = out + (((size_t)n * oH + ty0 + ty) * oW + tx0 + tx) * C + c0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      o[k] = pack_bf16x2(apply_act(acc[8 * j + 2 * k], act), apply_act(acc[8 * j + 2 * k + 1], act));
    reinterpret_cast<uint4*>(op)[j] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// ---------------------------------------------------------------------------------------------
// c3_tconv: out[n,2q+p,co] = sum over taps of parity p, ci of x[n,q+(p+1-k)/2,ci] * w[kh,kw,co,ci]
// thread = one input location q -> a 2x2 output quad x 3 channels.
// epilogue: mode 0: act(v + bias); mode 1: v * (1 - aux^2)   (tanh backward against aux = tanh output)
// grid = (tiles per image, N)
// ---------------------------------------------------------------------------------------------
constexpr int kTQ = 2;             // input locations per thread (rows qy and qy + kTH/2 of the tile): the
                                   // broadcast filter reads from shared memory are the bottleneck, so each
                                   // filter value fetched feeds kTQ pixels
constexpr int kTconvThreads = kC3Threads / kTQ;
__global__ void __launch_bounds__(kTconvThreads)
c3_tconv_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                const __nv_bfloat16* __restrict__ aux, __nv_bfloat16* __restrict__ out,
                float* __restrict__ colsum /* optional [3]: += per-channel sum of the f32 results */, int H, int W,
                int C, int act, int mode) {
  extern __shared__ float ws[];  // [25][3][C]
  for (int i = threadIdx.x; i < 75 * C; i += kTconvThreads) ws[i] = w[i];
  __syncthreads();
  const int tiles_x = W / kTW;
  const int qy0 = (blockIdx.x / tiles_x) * kTH + threadIdx.x / kTW;  // + q * (kTH / kTQ)
  const int qx = (blockIdx.x % tiles_x) * kTW + threadIdx.x % kTW;
  const int n = blockIdx.y;
  float acc[kTQ][2][2][3];
#pragma unroll
  for (int q = 0; q < kTQ; ++q)
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int c = 0; c < 3; ++c) acc[q][a][b][c] = 0.f;

#pragma unroll
  for (int dy = -1; dy <= 1; ++dy) {
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      const int ix = qx + dx;
      const uint4* xp[kTQ];
      bool valid[kTQ];
#pragma unroll
      for (int q = 0; q < kTQ; ++q) {
        const int iy = qy0 + q * (kTH / kTQ) + dy;
        valid[q] = (unsigned)iy < (unsigned)H && (unsigned)ix < (unsigned)W;
        xp[q] = reinterpret_cast<const uint4*>(x + (((size_t)n * H + (valid[q] ? iy : 0)) * W + (valid[q] ? ix : 0)) * C);
      }
#pragma unroll 1
      for (int cb = 0; cb < C / 8; ++cb) {
        float xf[kTQ][8];
#pragma unroll
        for (int q = 0; q < kTQ; ++q) {
          uint4 v = make_uint4(0, 0, 0, 0);
          if (valid[q]) v = __ldg(xp[q] + cb);
          float2 f;
          f = unpack_bf16x2(v.x); xf[q][0] = f.x; xf[q][1] = f.y; f = unpack_bf16x2(v.y); xf[q][2] = f.x; xf[q][3] = f.y;
          f = unpack_bf16x2(v.z); xf[q][4] = f.x; xf[q][5] = f.y; f = unpack_bf16x2(v.w); xf[q][6] = f.x; xf[q][7] = f.y;
        }
#pragma unroll
        for (int ph = 0; ph < 2; ++ph) {
          const int kh = ph + 1 - 2 * dy;
          if (kh < 0 || kh > 4) continue;
#pragma unroll
          for (int pw = 0; pw < 2; ++pw) {
            const int kw = pw + 1 - 2 * dx;
            if (kw < 0 || kw > 4) continue;
#pragma unroll
            for (int co = 0; co < 3; ++co) {
              const float4* wp = reinterpret_cast<const float4*>(ws + ((kh * 5 + kw) * 3 + co) * C + cb * 8);
              const float4 w0 = wp[0], w1 = wp[1];
#pragma unroll
              for (int q = 0; q < kTQ; ++q)
                acc[q][ph][pw][co] += xf[q][0] * w0.x + xf[q][1] * w0.y + xf[q][2] * w0.z + xf[q][3] * w0.w +
                                      xf[q][4] * w1.x + xf[q][5] * w1.y + xf[q][6] * w1.z + xf[q][7] * w1.w;
            }
          }
        }
      }
    }
  }
  const int oH = 2 * H, oW = 2 * W;
  float csum[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < kTQ; ++q) {
    const int qy = qy0 + q * (kTH / kTQ);
#pragma unroll
    for (int ph = 0; ph < 2; ++ph) {
      const size_t o = (((size_t)n * oH + 2 * qy + ph) * oW + 2 * qx) * 3;  // 6 contiguous bf16
      float v[6];
#pragma unroll
      for (int pw = 0; pw < 2; ++pw)
#pragma unroll
        for (int co = 0; co < 3; ++co) v[pw * 3 + co] = acc[q][ph][pw][co];
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
  if (colsum != nullptr) {  // bias gradient of the 3-channel layer, reduced in f32 before the bf16 rounding
    __shared__ float cred[3];
    if (threadIdx.x < 3) cred[threadIdx.x] = 0.f;
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float s = warp_sum(csum[c]);
      if ((threadIdx.x & 31) == 0) atomicAdd(&cred[c], s);
    }
    __syncthreads();
    if (threadIdx.x < 3) atomicAdd(&colsum[threadIdx.x], cred[threadIdx.x]);
  }
}

// ---------------------------------------------------------------------------------------------
// c3_wgrad: dw[tap][c3][c] += sum_{n,q} x3[n,2q+k-1,c3] * y[n,q,c]        (f32 atomics)
// block walks tiles of 4x32 q positions; thread (cg, tg): 4 channels x 5 (tap,c3) combinations.
// grid = (blocks, 1, C/64)
// ---------------------------------------------------------------------------------------------
constexpr int kWTH = 4, kWTW = 32;
constexpr int kWPatchH = 2 * kWTH + 3;          // 11
constexpr int kWPatchW = (2 * kWTW + 3) * 3;    // 201
__global__ void __launch_bounds__(kC3Threads)
c3_wgrad_kernel(const __nv_bfloat16* __restrict__ x3, const __nv_bfloat16* __restrict__ y, float* __restrict__ dw,
                int N, int Hs, int Ws, int C) {
  extern __shared__ float smem[];
  float* ys = smem;                          // [128][64]
  float* xs = smem + kWTH * kWTW * 64;       // [11][202]
  const int H = 2 * Hs, W = 2 * Ws;
  const int c0 = blockIdx.z * 64;
  const int cg = threadIdx.x & 15, tg = threadIdx.x >> 4;  // tg 0..15, 15 idle
  int off[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int j = min(tg, 14) * 5 + i;  // (tap, c3) combination
    const int t = j / 3, c3 = j % 3;
    off[i] = (t / 5) * (kWPatchW + 1) + (t % 5) * 3 + c3;
  }
  float acc[5][4];
#pragma unroll
  for (int i = 0; i < 5; ++i)
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[i][k] = 0.f;
  const int tiles_x = Ws / kWTW, tiles_y = Hs / kWTH;
  const int tiles = N * tiles_y * tiles_x;
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int n = tile / (tiles_y * tiles_x);
    const int tr = tile % (tiles_y * tiles_x);
    const int qy0 = (tr / tiles_x) * kWTH, qx0 = (tr % tiles_x) * kWTW;
    __syncthreads();  // previous tile fully consumed
    for (int i = threadIdx.x; i < kWTH * kWTW * 8; i += kC3Threads) {  // 8 x uint4 per pixel (64 ch)
      const int pix = i >> 3, ch = i & 7;
      const int qy = qy0 + pix / kWTW, qx = qx0 + pix % kWTW;
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(y + (((size_t)n * Hs + qy) * Ws + qx) * C + c0) + ch);
      float* d = ys + pix * 64 + ch * 8;
      float2 f;
      f = unpack_bf16x2(v.x); d[0] = f.x; d[1] = f.y; f = unpack_bf16x2(v.y); d[2] = f.x; d[3] = f.y;
      f = unpack_bf16x2(v.z); d[4] = f.x; d[5] = f.y; f = unpack_bf16x2(v.w); d[6] = f.x; d[7] = f.y;
    }
    const int iy0 = 2 * qy0 - 1, ix0 = 2 * qx0 - 1;
    for (int i = threadIdx.x; i < kWPatchH * kWPatchW; i += kC3Threads) {
      const int r = i / kWPatchW, cc = i % kWPatchW;
      const int iy = iy0 + r, ix = ix0 + cc / 3;
      float v = 0.f;
      if ((unsigned)iy < (unsigned)H && (unsigned)ix < (unsigned)W)
        v = __bfloat162float(x3[(((size_t)n * H + iy) * W + ix) * 3 + cc % 3]);
      xs[r * (kWPatchW + 1) + cc] = v;
    }
    __syncthreads();
    if (tg < 15) {
#pragma unroll 4
      for (int pix = 0; pix < kWTH * kWTW; ++pix) {
        const int py = pix / kWTW, px = pix % kWTW;
        const float4 yv = *reinterpret_cast<const float4*>(ys + pix * 64 + cg * 4);
        const float* xb = xs + 2 * py * (kWPatchW + 1) + 2 * px * 3;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
          const float xv = xb[off[i]];
          acc[i][0] += xv * yv.x; acc[i][1] += xv * yv.y; acc[i][2] += xv * yv.z; acc[i][3] += xv * yv.w;
        }
      }
    }
  }
  if (tg < 15) {
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      const int j = tg * 5 + i;
#pragma unroll
      for (int k = 0; k < 4; ++k) atomicAdd(&dw[(size_t)j * C + c0 + cg * 4 + k], acc[i][k]);
    }
  }
}

}  // namespace dcgan

using namespace dcgan;
#define LAUNCH_OK() (cudaGetLastError() == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH)

extern "C" {

int dcgan_c3_fprop(const void* x, const float* w, const float* bias, void* out, int N, int H, int W, int C,
                   int act, cudaStream_t st) {
  if (C % 64 || (H / 2) % kTH || (W / 2) % kTW || (H & 1) || (W & 1)) return DCGAN_ERR_SHAPE;
  const size_t smem = (75 * 64 + kPatchH * kPatchWPad) * sizeof(float);
  dim3 grid((H / 2 / kTH) * (W / 2 / kTW), N, C / 64);
  c3_fprop_kernel<<<grid, kC3Threads, smem, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), w, bias,
                                                  reinterpret_cast<__nv_bfloat16*>(out), H, W, C, act);
  return LAUNCH_OK();
}

int dcgan_c3_tconv(const void* x, const float* w, const float* bias, const void* aux, void* out, float* colsum,
                   int N, int H, int W, int C, int act, int mode, cudaStream_t st) {
  if (C % 8 || H % kTH || W % kTW) return DCGAN_ERR_SHAPE;
  const size_t smem = (size_t)75 * C * sizeof(float);
  if (smem > 48 * 1024) {
    static bool done = false;
    if (!done) {
      if (cudaFuncSetAttribute(c3_tconv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024) !=
          cudaSuccess)
        return DCGAN_ERR_LAUNCH;
      done = true;
    }
    if (smem > 100 * 1024) return DCGAN_ERR_SHAPE;
  }
  dim3 grid((H / kTH) * (W / kTW), N);
  c3_tconv_kernel<<<grid, kTconvThreads, smem, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), w, bias,
                                                  reinterpret_cast<const __nv_bfloat16*>(aux),
                                                  reinterpret_cast<__nv_bfloat16*>(out), colsum, H, W, C, act, mode);
  return LAUNCH_OK();
}

int dcgan_c3_wgrad(const void* x3, const void* y, float* dw, int N, int Hs, int Ws, int C, cudaStream_t st) {
  if (C % 64 || Hs % kWTH || Ws % kWTW) return DCGAN_ERR_SHAPE;
  const size_t smem = (kWTH * kWTW * 64 + kWPatchH * (kWPatchW + 1)) * sizeof(float);
  const int tiles = N * (Hs / kWTH) * (Ws / kWTW);
  int blocks = tiles < 296 ? tiles : 296;
  dim3 grid(blocks, 1, C / 64);
  c3_wgrad_kernel<<<grid, kC3Threads, smem, st>>>(reinterpret_cast<const __nv_bfloat16*>(x3),
                                                  reinterpret_cast<const __nv_bfloat16*>(y), dw, N, Hs, Ws, C);
  return LAUNCH_OK();
}

}  // extern "C"
