// Stand-alone bring-up harness for the tcgen05 implicit-GEMM kernels (no torch needed).
// Builds with the library sources, runs each mode against a CPU reference and times the real
// DCGAN layer shapes.  Usage: test_igemm [quick]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "../include/dcgan_b200.h"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)

static uint32_t rng_state = 12345u;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xffff) / 32768.0f - 1.0f;
}
static float bf16_round(float f) { return __bfloat162float(__float2bfloat16(f)); }

struct Buf {
  std::vector<float> h;  // bf16-rounded values in f32
  __nv_bfloat16* d = nullptr;
  void init(size_t n, float scale) {
    h.resize(n);
    std::vector<__nv_bfloat16> tmp(n);
    for (size_t i = 0; i < n; ++i) { tmp[i] = __float2bfloat16(frand() * scale); h[i] = __bfloat162float(tmp[i]); }
    CK(cudaMalloc(&d, n * 2));
    CK(cudaMemcpy(d, tmp.data(), n * 2, cudaMemcpyHostToDevice));
  }
  void free() { if (d) cudaFree(d); d = nullptr; }
};

static int check(const char* name, const std::vector<float>& got, const std::vector<double>& ref, double tol_rel) {
  double max_err = 0, max_ref = 0;
  size_t bad_i = 0;
  for (size_t i = 0; i < ref.size(); ++i) {
    double e = fabs(got[i] - ref[i]);
    if (e > max_err) { max_err = e; bad_i = i; }
    if (fabs(ref[i]) > max_ref) max_ref = fabs(ref[i]);
  }
  int dev_err = dcgan_last_device_error();
  bool pass = (max_err <= tol_rel * max_ref) && dev_err == 0 && max_ref > 0;
  printf("[%s] %s max_err=%.5g max_ref=%.5g (idx %zu got %.5g ref %.5g) dev_err=%d\n", pass ? "PASS" : "FAIL", name,
         max_err, max_ref, bad_i, got[bad_i], ref[bad_i], dev_err);
  fflush(stdout);
  return pass ? 0 : 1;
}

static int test_fprop(int NB, int H, int W, int Cin, int Cout, int block_n) {
  Buf x, w;
  x.init((size_t)NB * H * W * Cin, 1.0f);
  w.init((size_t)25 * Cout * Cin, 0.05f);
  std::vector<float> bias(Cout);
  for (auto& b : bias) b = frand() * 0.1f;
  float* dbias; CK(cudaMalloc(&dbias, Cout * 4)); CK(cudaMemcpy(dbias, bias.data(), Cout * 4, cudaMemcpyHostToDevice));
  const int oH = H / 2, oW = W / 2;
  size_t on = (size_t)NB * oH * oW * Cout;
  __nv_bfloat16* dout; CK(cudaMalloc(&dout, on * 2)); CK(cudaMemset(dout, 0xff, on * 2));
  int rc = dcgan_conv_fprop(x.d, w.d, dbias, dout, NB, H, W, Cin, Cout, DCGAN_ACT_LRELU, block_n, 0);
  CK(cudaDeviceSynchronize());
  std::vector<__nv_bfloat16> ho(on); CK(cudaMemcpy(ho.data(), dout, on * 2, cudaMemcpyDeviceToHost));
  std::vector<float> got(on); for (size_t i = 0; i < on; ++i) got[i] = __bfloat162float(ho[i]);
  std::vector<double> ref(on);
  for (int n = 0; n < NB; ++n) for (int oh = 0; oh < oH; ++oh) for (int ow = 0; ow < oW; ++ow) for (int co = 0; co < Cout; ++co) {
    double acc = bias[co];
    for (int kh = 0; kh < 5; ++kh) for (int kw = 0; kw < 5; ++kw) {
      int ih = 2 * oh + kh - 1, iw = 2 * ow + kw - 1;
      if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
      const float* xp = &x.h[((size_t)(n * H + ih) * W + iw) * Cin];
      const float* wp = &w.h[((size_t)(kh * 5 + kw) * Cout + co) * Cin];
      for (int ci = 0; ci < Cin; ++ci) acc += (double)xp[ci] * wp[ci];
    }
    ref[((size_t)(n * oH + oh) * oW + ow) * Cout + co] = acc > 0 ? acc : 0.2 * acc;
  }
  char name[128]; snprintf(name, sizeof name, "fprop NB=%d H=%d Cin=%d Cout=%d bn=%d rc=%d", NB, H, Cin, Cout, block_n, rc);
  int r = check(name, got, ref, 0.01) | (rc != 0);
  x.free(); w.free(); cudaFree(dbias); cudaFree(dout);
  return r;
}

static int test_tconv(int NB, int H, int W, int Cin, int Cout, int block_n) {
  Buf x, w;
  x.init((size_t)NB * H * W * Cin, 1.0f);
  w.init((size_t)25 * Cout * Cin, 0.05f);
  const int oH = 2 * H, oW = 2 * W;
  size_t on = (size_t)NB * oH * oW * Cout;
  __nv_bfloat16* dout; CK(cudaMalloc(&dout, on * 2)); CK(cudaMemset(dout, 0xff, on * 2));
  int rc = dcgan_conv_tconv(x.d, w.d, nullptr, dout, NB, H, W, Cin, Cout, DCGAN_ACT_NONE, block_n, 0);
  CK(cudaDeviceSynchronize());
  std::vector<__nv_bfloat16> ho(on); CK(cudaMemcpy(ho.data(), dout, on * 2, cudaMemcpyDeviceToHost));
  std::vector<float> got(on); for (size_t i = 0; i < on; ++i) got[i] = __bfloat162float(ho[i]);
  std::vector<double> ref(on, 0.0);
  for (int n = 0; n < NB; ++n) for (int oy = 0; oy < oH; ++oy) for (int ox = 0; ox < oW; ++ox) for (int co = 0; co < Cout; ++co) {
    double acc = 0;
    for (int kh = 0; kh < 5; ++kh) for (int kw = 0; kw < 5; ++kw) {
      int ty = oy + 1 - kh, tx = ox + 1 - kw;
      if ((ty & 1) || (tx & 1)) continue;
      int iy = ty / 2, ix = tx / 2;
      if (ty < 0 || tx < 0 || iy >= H || ix >= W) continue;
      const float* xp = &x.h[((size_t)(n * H + iy) * W + ix) * Cin];
      const float* wp = &w.h[((size_t)(kh * 5 + kw) * Cout + co) * Cin];
      for (int ci = 0; ci < Cin; ++ci) acc += (double)xp[ci] * wp[ci];
    }
    ref[((size_t)(n * oH + oy) * oW + ox) * Cout + co] = acc;
  }
  char name[128]; snprintf(name, sizeof name, "tconv NB=%d H=%d Cin=%d Cout=%d bn=%d rc=%d", NB, H, Cin, Cout, block_n, rc);
  int r = check(name, got, ref, 0.01) | (rc != 0);
  x.free(); w.free(); cudaFree(dout);
  return r;
}

static int test_wgrad(int NB, int Hs, int Ws, int Cg, int Cd, int splits, int block_n, int variant) {
  Buf big, small;
  const int H = 2 * Hs, W = 2 * Ws;
  big.init((size_t)NB * H * W * Cg, 1.0f);
  small.init((size_t)NB * Hs * Ws * Cd, 1.0f);
  size_t on = (size_t)25 * Cg * Cd;
  float* dw; CK(cudaMalloc(&dw, on * 4)); CK(cudaMemset(dw, 0, on * 4));
  dcgan_set_mn_desc_variant(variant);
  int rc = dcgan_conv_wgrad(big.d, small.d, dw, NB, Hs, Ws, Cg, Cd, splits, block_n, 0);
  CK(cudaDeviceSynchronize());
  dcgan_set_mn_desc_variant(0);
  std::vector<float> got(on); CK(cudaMemcpy(got.data(), dw, on * 4, cudaMemcpyDeviceToHost));
  std::vector<double> ref(on, 0.0);
  for (int kh = 0; kh < 5; ++kh) for (int kw = 0; kw < 5; ++kw)
    for (int n = 0; n < NB; ++n) for (int qh = 0; qh < Hs; ++qh) for (int qw = 0; qw < Ws; ++qw) {
      int y = 2 * qh + kh - 1, x = 2 * qw + kw - 1;
      if (y < 0 || y >= H || x < 0 || x >= W) continue;
      const float* bp = &big.h[((size_t)(n * H + y) * W + x) * Cg];
      const float* sp = &small.h[((size_t)(n * Hs + qh) * Ws + qw) * Cd];
      double* rp = &ref[(size_t)(kh * 5 + kw) * Cg * Cd];
      for (int cg = 0; cg < Cg; ++cg) { double b = bp[cg]; for (int cd = 0; cd < Cd; ++cd) rp[(size_t)cg * Cd + cd] += b * sp[cd]; }
    }
  char name[160]; snprintf(name, sizeof name, "wgrad NB=%d Hs=%d Cg=%d Cd=%d splits=%d bn=%d variant=%d rc=%d", NB, Hs, Cg, Cd, splits, block_n, variant, rc);
  int r = check(name, got, ref, 0.01) | (rc != 0);
  big.free(); small.free(); cudaFree(dw);
  return r;
}

template <class F>
static float time_ms(F f, int iters) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) f();
  CK(cudaDeviceSynchronize());
  cudaEventRecord(a);
  for (int i = 0; i < iters; ++i) f();
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms / iters;
}

static void bench_layers() {
  struct LayerF { const char* name; int NB, H, Cin, Cout; };
  LayerF fl[] = {{"d_h1 fprop", 128, 32, 64, 128}, {"d_h2 fprop", 128, 16, 128, 256}, {"d_h3 fprop", 128, 8, 256, 512}};
  for (auto& l : fl) {
    Buf x, w; x.init((size_t)l.NB * l.H * l.H * l.Cin, 1.f); w.init((size_t)25 * l.Cout * l.Cin, 0.05f);
    __nv_bfloat16* o; CK(cudaMalloc(&o, (size_t)l.NB * l.H * l.H / 4 * l.Cout * 2));
    for (int bn : {64, 128, 256}) {
      if (l.Cout % bn) continue;
      float ms = time_ms([&] { dcgan_conv_fprop(x.d, w.d, nullptr, o, l.NB, l.H, l.H, l.Cin, l.Cout, 0, bn, 0); }, 20);
      double fl_ = 2.0 * l.NB * (l.H / 2) * (l.H / 2) * l.Cout * 25.0 * l.Cin;
      printf("[TIME] %s bn=%d %.3f ms %.1f TFLOP/s err=%d\n", l.name, bn, ms, fl_ / ms * 1e-9, dcgan_last_device_error());
    }
    x.free(); w.free(); cudaFree(o);
  }
  LayerF tl[] = {{"g_h1 tconv", 64, 4, 512, 256}, {"g_h2 tconv", 64, 8, 256, 128}, {"g_h3 tconv", 64, 16, 128, 64},
                 {"d_h3 dgrad", 192, 4, 512, 256}, {"d_h2 dgrad", 192, 8, 256, 128}, {"d_h1 dgrad", 192, 16, 128, 64}};
  for (auto& l : tl) {
    Buf x, w; x.init((size_t)l.NB * l.H * l.H * l.Cin, 1.f); w.init((size_t)25 * l.Cout * l.Cin, 0.05f);
    __nv_bfloat16* o; CK(cudaMalloc(&o, (size_t)l.NB * l.H * l.H * 4 * l.Cout * 2));
    for (int bn : {0, 64, 128, 256}) {
      if (bn && l.Cout % bn) continue;
      float ms = time_ms([&] { dcgan_conv_tconv(x.d, w.d, nullptr, o, l.NB, l.H, l.H, l.Cin, l.Cout, 0, bn, 0); }, 20);
      double fl_ = 2.0 * l.NB * l.H * l.H * l.Cout * 25.0 * l.Cin;
      printf("[TIME] %s bn=%d %.3f ms %.1f TFLOP/s err=%d\n", l.name, bn, ms, fl_ / ms * 1e-9, dcgan_last_device_error());
    }
    x.free(); w.free(); cudaFree(o);
  }
  struct LayerW { const char* name; int NB, Hs, Cg, Cd; };
  LayerW wl[] = {{"d_h1 wgrad", 128, 16, 64, 128}, {"d_h2 wgrad", 128, 8, 128, 256}, {"d_h3 wgrad", 128, 4, 256, 512},
                 {"g_h1 wgrad", 64, 4, 256, 512}, {"g_h2 wgrad", 64, 8, 128, 256}, {"g_h3 wgrad", 64, 16, 64, 128}};
  for (auto& l : wl) {
    Buf big, small; big.init((size_t)l.NB * l.Hs * l.Hs * 4 * l.Cg, 1.f); small.init((size_t)l.NB * l.Hs * l.Hs * l.Cd, 1.f);
    float* dw; CK(cudaMalloc(&dw, (size_t)25 * l.Cg * l.Cd * 4));
    for (int bn : {64, 128, 256}) {
      if (l.Cg % bn) continue;
      float ms = time_ms([&] { dcgan_conv_wgrad(big.d, small.d, dw, l.NB, l.Hs, l.Hs, l.Cg, l.Cd, 0, bn, 0); }, 20);
      double fl_ = 2.0 * l.NB * l.Hs * l.Hs * l.Cg * 25.0 * l.Cd;
      printf("[TIME] %s bn=%d %.3f ms %.1f TFLOP/s err=%d\n", l.name, bn, ms, fl_ / ms * 1e-9, dcgan_last_device_error());
    }
    big.free(); small.free(); cudaFree(dw);
  }
}

// cluster split-K (DSMEM reduction) on the small-M layers: forced factors vs the per-layer choice (0)
static void bench_splitk() {
  struct Layer { const char* name; int tconv, NB, H, Cin, Cout; };
  Layer ls[] = {{"g_h1 dgrad (F)", 0, 64, 8, 256, 512},  {"g_h2 dgrad (F)", 0, 64, 16, 128, 256},
                {"g_h3 dgrad (F)", 0, 64, 32, 64, 128},  {"d_h3 fprop (F)", 0, 128, 8, 256, 512},
                {"d_h2 fprop (F)", 0, 128, 16, 128, 256}, {"d_h1 fprop (F)", 0, 128, 32, 64, 128},
                {"g_h1 tconv (T)", 1, 64, 4, 512, 256},  {"g_h2 tconv (T)", 1, 64, 8, 256, 128},
                {"d_h3 dgrad (T)", 1, 192, 4, 512, 256}, {"d_h2 dgrad (T)", 1, 192, 8, 256, 128}};
  for (auto& l : ls) {
    Buf x, w; x.init((size_t)l.NB * l.H * l.H * l.Cin, 1.f); w.init((size_t)25 * l.Cout * l.Cin, 0.05f);
    const size_t on = l.tconv ? (size_t)l.NB * l.H * l.H * 4 * l.Cout : (size_t)l.NB * l.H * l.H / 4 * l.Cout;
    __nv_bfloat16* o; CK(cudaMalloc(&o, on * 2));
    for (int bn : {0, 64, 128}) {
      if (bn && l.Cout % bn) continue;
      for (int ks : {1, 2, 4, 0}) {
        dcgan_set_igemm_ksplit(ks);
        int rc = 0;
        float ms = time_ms([&] {
          rc |= l.tconv ? dcgan_conv_tconv(x.d, w.d, nullptr, o, l.NB, l.H, l.H, l.Cin, l.Cout, 0, bn, 0)
                        : dcgan_conv_fprop(x.d, w.d, nullptr, o, l.NB, l.H, l.H, l.Cin, l.Cout, 0, bn, 0);
        }, 20);
        printf("[SPLITK] %s bn=%d ksplit=%d rc=%d %.1f us err=%d\n", l.name, bn, ks, rc, ms * 1e3, dcgan_last_device_error());
      }
    }
    dcgan_set_igemm_ksplit(0);
    x.free(); w.free(); cudaFree(o);
  }
}

int main(int argc, char** argv) {
  int fails = 0;
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  printf("device: %s sm_%d%d SMs=%d\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  fails += test_fprop(2, 16, 16, 64, 64, 64);
  fails += test_fprop(2, 16, 16, 128, 128, 128);
  fails += test_fprop(3, 8, 8, 64, 256, 256);
  fails += test_fprop(1, 4, 8, 64, 64, 64);  // partial M tile (8 rows)
  fails += test_tconv(2, 8, 8, 64, 64, 64);
  fails += test_tconv(2, 4, 4, 128, 128, 128);
  fails += test_tconv(1, 8, 8, 64, 256, 256);
  // parity-fused transposed conv (needs >= 90 M tiles: 128*... rows)
  dcgan_set_igemm_tuning(1 | 8);
  fails += test_tconv(48, 16, 16, 64, 64, 0);     // 12288 rows = 96 tiles, N=64
  fails += test_tconv(48, 16, 16, 128, 128, 0);   // Cout=128 as two N=64 tiles
  dcgan_set_igemm_tuning(1 | 8 | 16);
  fails += test_tconv(48, 16, 16, 64, 128, 0);    // N=128 tile, 512 TMEM columns
  dcgan_set_igemm_tuning(1);
  // split-K across a thread-block cluster, partial tiles reduced through distributed shared memory
  for (int ks : {2, 4}) {
    dcgan_set_igemm_ksplit(ks);
    fails += test_fprop(2, 16, 16, 64, 64, 64);
    fails += test_fprop(1, 4, 8, 128, 64, 64);    // partial M tile
    fails += test_tconv(2, 8, 8, 64, 64, 64);
    if (ks == 2) {
      fails += test_fprop(2, 16, 16, 128, 128, 128);
      fails += test_tconv(2, 4, 4, 128, 128, 128);
      fails += test_fprop(3, 8, 8, 64, 256, 256);
    }
  }
  dcgan_set_igemm_ksplit(0);
  int w0 = test_wgrad(2, 8, 8, 64, 128, 1, 64, 0);
  int w1 = 0;
  if (w0) { w1 = test_wgrad(2, 8, 8, 64, 128, 1, 64, 1); printf("MN-major variant 1 %s\n", w1 ? "also fails" : "passes"); }
  fails += w0;
  fails += test_wgrad(2, 8, 8, 128, 128, 2, 128, 0);
  fails += test_wgrad(4, 4, 4, 256, 256, 1, 256, 0);
  fails += test_wgrad(4, 8, 8, 128, 256, 0, 0, 0);
  printf("correctness: %d failing tests\n", fails);
  if (argc >= 2 && !strcmp(argv[1], "splitk")) bench_splitk();
  if (argc < 2) {
    for (int tune : {1, 9, 25}) {
      printf("=== igemm tuning %d ===\n", tune);
      dcgan_set_igemm_tuning(tune);
      bench_layers();
    }
  }
  return fails ? 1 : 0;
}
