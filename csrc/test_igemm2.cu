// Stand-alone harness for the TMA-fed persistent implicit-GEMM kernels (csrc/igemm2_sm100.cu), no torch needed.
// Every mode is checked against a CPU reference -- small shapes exhaustively, the real DCGAN layer shapes on a
// sample of output rows / elements -- and then timed against the first-generation kernels on the same box.
// Usage: test_igemm2 [quick | time]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "../include/dcgan_b200.h"

extern "C" void dcgan_set_igemm_max_ctas(int v);
extern "C" void dcgan_set_igemm_stage_mode(int v);
extern "C" void dcgan_set_igemm_debug(long long* ptr);

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)

static uint32_t rng_state = 12345u;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xffff) / 32768.0f - 1.0f;
}

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

static int report(const char* name, double max_err, double max_ref, double tol_rel, int rc) {
  int dev_err = dcgan_last_device_error();
  bool pass = (max_err <= tol_rel * max_ref) && dev_err == 0 && max_ref > 0 && rc == 0;
  printf("[%s] %s max_err=%.5g max_ref=%.5g rc=%d dev_err=%d\n", pass ? "PASS" : "FAIL", name, max_err, max_ref, rc, dev_err);
  fflush(stdout);
  return pass ? 0 : 1;
}

// layout: 0 = w[25][Cout][Cin] (dcgan_conv_fprop), 1 = w[25][Cin][Cout] (dcgan_conv_fprop_kn)
static int test_fprop(int NB, int H, int W, int Cin, int Cout, int block_n, int layout, int row_step, bool with_stats) {
  Buf x, w;
  x.init((size_t)NB * H * W * Cin, 1.0f);
  w.init((size_t)25 * Cout * Cin, 0.05f);
  std::vector<float> bias(Cout);
  for (auto& b : bias) b = frand() * 0.1f;
  float* dbias; CK(cudaMalloc(&dbias, Cout * 4)); CK(cudaMemcpy(dbias, bias.data(), Cout * 4, cudaMemcpyHostToDevice));
  const int oH = H / 2, oW = W / 2;
  const size_t rows = (size_t)NB * oH * oW, on = rows * Cout;
  __nv_bfloat16* dout; CK(cudaMalloc(&dout, on * 2)); CK(cudaMemset(dout, 0xff, on * 2));
  float* dstats = nullptr;
  const int groups = 2;
  int rc;
  if (with_stats) {
    CK(cudaMalloc(&dstats, groups * 2 * Cout * 4)); CK(cudaMemset(dstats, 0, groups * 2 * Cout * 4));
    rc = dcgan_conv_fprop_kn_stats(x.d, w.d, dbias, dout, dstats, groups, NB, H, W, Cin, Cout, block_n, 0);
  } else if (layout) {
    rc = dcgan_conv_fprop_kn(x.d, w.d, dbias, dout, NB, H, W, Cin, Cout, DCGAN_ACT_LRELU, block_n, 0);
  } else {
    rc = dcgan_conv_fprop(x.d, w.d, dbias, dout, NB, H, W, Cin, Cout, DCGAN_ACT_LRELU, block_n, 0);
  }
  CK(cudaDeviceSynchronize());
  std::vector<__nv_bfloat16> ho(on); CK(cudaMemcpy(ho.data(), dout, on * 2, cudaMemcpyDeviceToHost));
  double max_err = 0, max_ref = 0;
  std::vector<double> s1(groups * Cout, 0.0), s2(groups * Cout, 0.0);
  const int step = with_stats ? 1 : row_step;
  for (size_t R = 0; R < rows; R += step) {
    const int ow = R % oW, oh = (R / oW) % oH, n = R / ((size_t)oW * oH);
    for (int co = 0; co < Cout; ++co) {
      double acc = bias[co];
      for (int kh = 0; kh < 5; ++kh) for (int kw = 0; kw < 5; ++kw) {
        int ih = 2 * oh + kh - 1, iw = 2 * ow + kw - 1;
        if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
        const float* xp = &x.h[((size_t)(n * H + ih) * W + iw) * Cin];
        if (layout || with_stats) {
          const float* wp = &w.h[(size_t)(kh * 5 + kw) * Cin * Cout + co];
          for (int ci = 0; ci < Cin; ++ci) acc += (double)xp[ci] * wp[(size_t)ci * Cout];
        } else {
          const float* wp = &w.h[((size_t)(kh * 5 + kw) * Cout + co) * Cin];
          for (int ci = 0; ci < Cin; ++ci) acc += (double)xp[ci] * wp[ci];
        }
      }
      if (with_stats) {
        const int g = R / (rows / groups);
        s1[g * Cout + co] += acc; s2[g * Cout + co] += acc * acc;
      }
      const double ref = with_stats ? acc : (acc > 0 ? acc : 0.2 * acc);
      const double got = __bfloat162float(ho[R * Cout + co]);
      max_err = fmax(max_err, fabs(got - ref)); max_ref = fmax(max_ref, fabs(ref));
    }
  }
  char name[160];
  snprintf(name, sizeof name, "fprop%s NB=%d H=%d Cin=%d Cout=%d bn=%d%s", layout ? "_kn" : "", NB, H, Cin, Cout, block_n, with_stats ? " +stats" : "");
  int r = report(name, max_err, max_ref, 0.01, rc);
  if (with_stats) {
    std::vector<float> hs(groups * 2 * Cout); CK(cudaMemcpy(hs.data(), dstats, hs.size() * 4, cudaMemcpyDeviceToHost));
    double e = 0, m = 0;
    for (int g = 0; g < groups; ++g) for (int c = 0; c < Cout; ++c) {
      e = fmax(e, fabs(hs[(g * 2) * Cout + c] - s1[g * Cout + c])); m = fmax(m, fabs(s1[g * Cout + c]));
      e = fmax(e, fabs(hs[(g * 2 + 1) * Cout + c] - s2[g * Cout + c])); m = fmax(m, fabs(s2[g * Cout + c]));
    }
    r |= report("   ... moments", e, m, 2e-3, 0);
    cudaFree(dstats);
  }
  x.free(); w.free(); cudaFree(dbias); cudaFree(dout);
  return r;
}

static int test_tconv(int NB, int H, int W, int Cin, int Cout, int block_n, int row_step, bool with_stats) {
  Buf x, w;
  x.init((size_t)NB * H * W * Cin, 1.0f);
  w.init((size_t)25 * Cout * Cin, 0.05f);
  const int oH = 2 * H, oW = 2 * W;
  const size_t rows = (size_t)NB * oH * oW, on = rows * Cout;
  __nv_bfloat16* dout; CK(cudaMalloc(&dout, on * 2)); CK(cudaMemset(dout, 0xff, on * 2));
  float* dstats = nullptr;
  int rc;
  if (with_stats) {
    CK(cudaMalloc(&dstats, 2 * Cout * 4)); CK(cudaMemset(dstats, 0, 2 * Cout * 4));
    rc = dcgan_conv_tconv_stats(x.d, w.d, nullptr, dout, dstats, 1, NB, H, W, Cin, Cout, block_n, 0);
  } else {
    rc = dcgan_conv_tconv(x.d, w.d, nullptr, dout, NB, H, W, Cin, Cout, DCGAN_ACT_NONE, block_n, 0);
  }
  CK(cudaDeviceSynchronize());
  std::vector<__nv_bfloat16> ho(on); CK(cudaMemcpy(ho.data(), dout, on * 2, cudaMemcpyDeviceToHost));
  double max_err = 0, max_ref = 0;
  std::vector<double> s1(Cout, 0.0), s2(Cout, 0.0);
  const int step = with_stats ? 1 : row_step;
  for (size_t R = 0; R < rows; R += step) {
    const int ox = R % oW, oy = (R / oW) % oH, n = R / ((size_t)oW * oH);
    for (int co = 0; co < Cout; ++co) {
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
      s1[co] += acc; s2[co] += acc * acc;
      const double got = __bfloat162float(ho[R * Cout + co]);
      max_err = fmax(max_err, fabs(got - acc)); max_ref = fmax(max_ref, fabs(acc));
    }
  }
  char name[160];
  snprintf(name, sizeof name, "tconv NB=%d H=%d Cin=%d Cout=%d bn=%d%s", NB, H, Cin, Cout, block_n, with_stats ? " +stats" : "");
  int r = report(name, max_err, max_ref, 0.01, rc);
  if (with_stats) {
    std::vector<float> hs(2 * Cout); CK(cudaMemcpy(hs.data(), dstats, hs.size() * 4, cudaMemcpyDeviceToHost));
    double e = 0, m = 0;
    for (int c = 0; c < Cout; ++c) {
      e = fmax(e, fabs(hs[c] - s1[c])); m = fmax(m, fabs(s1[c]));
      e = fmax(e, fabs(hs[Cout + c] - s2[c])); m = fmax(m, fabs(s2[c]));
    }
    r |= report("   ... moments", e, m, 2e-3, 0);
    cudaFree(dstats);
  }
  x.free(); w.free(); cudaFree(dout);
  return r;
}

static int test_wgrad(int NB, int Hs, int Ws, int Cg, int Cd, int splits, int block_n, int elem_step, int repeat) {
  Buf big, small;
  const int H = 2 * Hs, W = 2 * Ws;
  big.init((size_t)NB * H * W * Cg, 1.0f);
  small.init((size_t)NB * Hs * Ws * Cd, 1.0f);
  size_t on = (size_t)25 * Cg * Cd;
  float* dw; CK(cudaMalloc(&dw, on * 4));
  std::vector<float> got(on), first;
  int rc = 0, determ_fail = 0;
  for (int it = 0; it < repeat; ++it) {  // a second run must reproduce the first bit for bit (ordered split-K)
    CK(cudaMemset(dw, 0, on * 4));
    rc |= dcgan_conv_wgrad(big.d, small.d, dw, NB, Hs, Ws, Cg, Cd, splits, block_n, 0);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(got.data(), dw, on * 4, cudaMemcpyDeviceToHost));
    if (it == 0) first = got;
    else if (memcmp(first.data(), got.data(), on * 4) != 0) determ_fail = 1;
  }
  double max_err = 0, max_ref = 0;
  for (size_t e = 0; e < on; e += elem_step) {
    const int cd = e % Cd, cg = (e / Cd) % Cg, tap = e / ((size_t)Cd * Cg);
    const int kh = tap / 5, kw = tap % 5;
    double acc = 0;
    for (int n = 0; n < NB; ++n) for (int qh = 0; qh < Hs; ++qh) for (int qw = 0; qw < Ws; ++qw) {
      int y = 2 * qh + kh - 1, xx = 2 * qw + kw - 1;
      if (y < 0 || y >= H || xx < 0 || xx >= W) continue;
      acc += (double)big.h[((size_t)(n * H + y) * W + xx) * Cg + cg] * small.h[((size_t)(n * Hs + qh) * Ws + qw) * Cd + cd];
    }
    max_err = fmax(max_err, fabs(got[e] - acc)); max_ref = fmax(max_ref, fabs(acc));
  }
  char name[160];
  snprintf(name, sizeof name, "wgrad NB=%d Hs=%d Cg=%d Cd=%d splits=%d bn=%d%s", NB, Hs, Cg, Cd, splits, block_n,
           determ_fail ? " NOT-REPRODUCIBLE" : "");
  int r = report(name, max_err, max_ref, 2e-3, rc) | determ_fail;
  big.free(); small.free(); cudaFree(dw);
  return r;
}

template <class F>
static float time_us(F f, int iters) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) f();
  CK(cudaDeviceSynchronize());
  cudaEventRecord(a);
  for (int i = 0; i < iters; ++i) f();
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms; cudaEventElapsedTime(&ms, a, b);
  cudaEventDestroy(a); cudaEventDestroy(b);
  return ms * 1e3f / iters;
}

struct Layer { const char* name; int kind; int NB, H, Cin, Cout; };  // kind 0 = F (kn layout), 1 = T, 2 = W (H = Hs, Cin = Cg, Cout = Cd)
static const Layer kLayers[] = {
    {"d_h1 fwd   (F)", 0, 128, 32, 64, 128},  {"d_h2 fwd   (F)", 0, 128, 16, 128, 256}, {"d_h3 fwd   (F)", 0, 128, 8, 256, 512},
    {"g_h3 dgrad (F)", 0, 64, 32, 64, 128},   {"g_h2 dgrad (F)", 0, 64, 16, 128, 256},  {"g_h1 dgrad (F)", 0, 64, 8, 256, 512},
    {"g_h1 fwd   (T)", 1, 64, 4, 512, 256},   {"g_h2 fwd   (T)", 1, 64, 8, 256, 128},   {"g_h3 fwd   (T)", 1, 64, 16, 128, 64},
    {"d_h3 dgrad (T)", 1, 192, 4, 512, 256},  {"d_h2 dgrad (T)", 1, 192, 8, 256, 128},  {"d_h1 dgrad (T)", 1, 192, 16, 128, 64},
    {"d_h1 wgrad (W)", 2, 128, 16, 64, 128},  {"d_h2 wgrad (W)", 2, 128, 8, 128, 256},  {"d_h3 wgrad (W)", 2, 128, 4, 256, 512},
    {"g_h1 wgrad (W)", 2, 64, 4, 256, 512},   {"g_h2 wgrad (W)", 2, 64, 8, 128, 256},   {"g_h3 wgrad (W)", 2, 64, 16, 64, 128}};

static double layer_flops(const Layer& l) {
  if (l.kind == 0) return 2.0 * l.NB * (l.H / 2) * (l.H / 2) * l.Cout * 25.0 * l.Cin;
  if (l.kind == 1) return 2.0 * l.NB * l.H * l.H * l.Cout * 25.0 * l.Cin;
  return 2.0 * l.NB * l.H * l.H * l.Cin * 25.0 * l.Cout;
}

static void time_layers(bool sweep) {
  double tot_old = 0, tot_new = 0;
  for (const Layer& l : kLayers) {
    Buf a, b;
    void* o = nullptr;
    if (l.kind == 2) {
      a.init((size_t)l.NB * l.H * l.H * 4 * l.Cin, 1.f); b.init((size_t)l.NB * l.H * l.H * l.Cout, 1.f);
      CK(cudaMalloc(&o, (size_t)25 * l.Cin * l.Cout * 4)); CK(cudaMemset(o, 0, (size_t)25 * l.Cin * l.Cout * 4));
    } else {
      a.init((size_t)l.NB * l.H * l.H * l.Cin, 1.f); b.init((size_t)25 * l.Cout * l.Cin, 0.05f);
      CK(cudaMalloc(&o, (size_t)l.NB * l.H * l.H * 4 * l.Cout * 2));
    }
    auto run = [&](int bn, int splits) {
      if (l.kind == 0) return dcgan_conv_fprop_kn(a.d, b.d, nullptr, o, l.NB, l.H, l.H, l.Cin, l.Cout, 0, bn, 0);
      if (l.kind == 1) return dcgan_conv_tconv(a.d, b.d, nullptr, o, l.NB, l.H, l.H, l.Cin, l.Cout, 0, bn, 0);
      return dcgan_conv_wgrad(a.d, b.d, (float*)o, l.NB, l.H, l.H, l.Cin, l.Cout, splits, bn, 0);
    };
    int rc = 0;
    dcgan_set_igemm_tuning(1 | 128);
    const float t_old = time_us([&] { rc |= run(0, 0); }, 20);
    dcgan_set_igemm_tuning(1);
    const float t_new = time_us([&] { rc |= run(0, 0); }, 20);
    dcgan_set_igemm_stage_mode(2);  // one CTA per SM, deepest ring
    const float t_deep = time_us([&] { rc |= run(0, 0); }, 20);
    dcgan_set_igemm_stage_mode(3);  // 64-column tiles, two k-blocks per slot
    const float t_k2 = time_us([&] { rc |= run(64, l.kind == 2 ? 0 : 0); }, 20);
    dcgan_set_igemm_stage_mode(0);
    const float t_64 = time_us([&] { rc |= run(64, 0); }, 20);
    const double fl = layer_flops(l);
    printf("[TIME] %s  gen1 %7.1f us (%6.1f TF/s)   gen2 %7.1f us (%6.1f TF/s)   x%.2f   gen2 with one CTA per SM: %.1f us   bn64: %.1f us, bn64 two k-blocks per slot: %.1f us  rc=%d err=%d\n",
           l.name, t_old, fl / t_old * 1e-6, t_new, fl / t_new * 1e-6, t_old / t_new, t_deep, t_64, t_k2, rc, dcgan_last_device_error());
#ifdef IGEMM2_TIMING
    {  // per-role cycle counters of one launch (bring-up builds only)
      long long* dbg; CK(cudaMalloc(&dbg, 296 * 16 * 8)); CK(cudaMemset(dbg, 0, 296 * 16 * 8));
      dcgan_set_igemm_debug(dbg);
      rc |= run(0, 0);
      CK(cudaDeviceSynchronize());
      dcgan_set_igemm_debug(nullptr);
      std::vector<long long> h(296 * 16); CK(cudaMemcpy(h.data(), dbg, h.size() * 8, cudaMemcpyDeviceToHost));
      for (int b : {0, 1, 150})
        printf("        [cta %3d] producer: %lld cyc, waits %lld (%lld slots) | mma: %lld cyc, wait full %lld (%lld slots), wait tmem %lld, %lld tiles | epilogue: %lld cyc, wait %lld, tmem/scratch loads %lld, stores %lld\n",
               b, h[b * 16], h[b * 16 + 1], h[b * 16 + 2], h[b * 16 + 3], h[b * 16 + 4], h[b * 16 + 5], h[b * 16 + 6], h[b * 16 + 7], h[b * 16 + 8], h[b * 16 + 9], h[b * 16 + 10], h[b * 16 + 11]);
      {  // how many CTAs were resident on one SM at the same time (start / end stamps in ns)
        int max_overlap = 0, ctas = 0;
        for (int b = 0; b < 296; ++b) {
          if (h[b * 16 + 14] == 0) continue;
          ++ctas;
          int ov = 0;
          for (int c = 0; c < 296; ++c)
            if (h[c * 16 + 14] != 0 && h[c * 16 + 12] == h[b * 16 + 12] && h[c * 16 + 13] < h[b * 16 + 14] &&
                h[b * 16 + 13] < h[c * 16 + 14])
              ++ov;
          if (ov > max_overlap) max_overlap = ov;
        }
        printf("        [residency] %d CTAs, at most %d resident on one SM at a time\n", ctas, max_overlap);
      }
      cudaFree(dbg);
    }
#endif
    tot_old += t_old; tot_new += t_new;
    if (sweep) {
      const int nch = (l.kind == 2) ? l.Cin : l.Cout;
      for (int bn : {64, 128, 256}) {
        if (nch % bn) continue;
        for (int ks : {1, 2, 3, 4, 6, 8}) {
          dcgan_set_igemm_ksplit(ks);
          const float t = time_us([&] { rc |= run(bn, l.kind == 2 ? ks : 0); }, 10);
          printf("    [SWEEP] %s bn=%d ksplit=%d  %7.1f us rc=%d err=%d\n", l.name, bn, ks, t, rc, dcgan_last_device_error());
        }
        dcgan_set_igemm_ksplit(0);
      }
    }
    a.free(); b.free(); cudaFree(o);
  }
  printf("[TIME] sum over the 18 layer GEMMs: gen1 %.1f us, gen2 %.1f us\n", tot_old, tot_new);
}

int main(int argc, char** argv) {
  int fails = 0;
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  printf("device: %s sm_%d%d SMs=%d\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  const bool quick = argc >= 2 && !strcmp(argv[1], "quick");
  const bool only_time = argc >= 2 && !strcmp(argv[1], "time");
  dcgan_set_igemm_tuning(1);
  if (!only_time) {
    // ---- small shapes, every output element ---------------------------------------------------------------
    fails += test_fprop(2, 16, 16, 64, 64, 64, 0, 1, false);
    fails += test_fprop(2, 16, 16, 128, 128, 128, 0, 1, false);
    fails += test_fprop(3, 8, 8, 64, 256, 256, 0, 1, false);
    fails += test_fprop(1, 4, 8, 64, 64, 64, 0, 1, false);  // partial M tile (8 rows)
    fails += test_fprop(2, 16, 16, 128, 128, 128, 1, 1, false);  // TensorFlow filter layout (MN-major B)
    fails += test_fprop(2, 16, 16, 64, 256, 256, 1, 1, false);
    fails += test_fprop(4, 16, 16, 64, 128, 0, 1, 1, true);      // fused batch-norm moments, two groups
    fails += test_tconv(2, 8, 8, 64, 64, 64, 1, false);
    fails += test_tconv(2, 4, 4, 128, 128, 128, 1, false);
    fails += test_tconv(1, 8, 8, 64, 256, 256, 1, false);
    fails += test_tconv(2, 8, 8, 64, 128, 0, 1, true);
    // output tiles of whole image rows (oH * oW >= 128)
    fails += test_fprop(2, 32, 32, 64, 128, 0, 1, 1, false);
    fails += test_fprop(1, 32, 32, 128, 64, 64, 1, 1, false);
    fails += test_fprop(1, 64, 64, 64, 64, 0, 1, 3, false);     // oW = 32: 4-row tiles
    fails += test_fprop(4, 32, 32, 64, 128, 0, 1, 1, true);
    // input tiles of whole image rows (H * W >= 128)
    fails += test_tconv(2, 16, 16, 64, 64, 64, 1, false);
    fails += test_tconv(1, 16, 16, 128, 128, 128, 1, false);
    fails += test_tconv(1, 32, 32, 64, 64, 64, 1, false);
    fails += test_tconv(3, 16, 16, 128, 64, 0, 1, true);
    for (int ks : {2, 3, 5, 8}) {  // ordered split-K through the scratch buffer
      dcgan_set_igemm_ksplit(ks);
      fails += test_fprop(2, 16, 16, 64, 64, 64, 0, 1, false);
      fails += test_fprop(1, 4, 8, 128, 64, 64, 0, 1, false);
      fails += test_fprop(2, 16, 16, 128, 128, 128, 1, 1, false);
      fails += test_tconv(2, 8, 8, 128, 64, 64, 1, false);
      fails += test_tconv(2, 4, 4, 128, 128, 128, 1, false);
      fails += test_tconv(1, 16, 16, 256, 64, 64, 1, false);
      fails += test_fprop(4, 16, 16, 64, 128, 0, 1, 1, true);
    }
    dcgan_set_igemm_ksplit(0);
    fails += test_wgrad(2, 8, 8, 64, 128, 1, 64, 1, 1);
    fails += test_wgrad(2, 8, 8, 128, 128, 2, 128, 1, 2);
    fails += test_wgrad(4, 4, 4, 256, 256, 1, 256, 1, 1);
    fails += test_wgrad(4, 8, 8, 128, 256, 0, 0, 1, 2);
    fails += test_wgrad(8, 16, 16, 64, 128, 5, 64, 7, 2);
    // a capped grid: every CTA walks several work items (ring and accumulator phases across tiles)
    dcgan_set_igemm_max_ctas(3);
    fails += test_fprop(2, 16, 16, 128, 128, 64, 1, 1, false);
    fails += test_tconv(2, 8, 8, 64, 128, 64, 1, false);
    fails += test_wgrad(2, 8, 8, 128, 128, 3, 64, 1, 2);
    fails += test_fprop(4, 16, 16, 64, 128, 64, 1, 1, true);
    dcgan_set_igemm_max_ctas(0);
    dcgan_set_igemm_stage_mode(2);  // one CTA per SM, deep rings
    fails += test_fprop(2, 16, 16, 128, 128, 128, 1, 1, false);
    fails += test_tconv(2, 8, 8, 64, 64, 64, 1, false);
    dcgan_set_igemm_stage_mode(3);  // 64-column tiles with two k-blocks per ring slot (odd and even k-block counts,
                                    // split-K slices that end inside a slot, several items per CTA)
    fails += test_fprop(2, 16, 16, 128, 128, 64, 1, 1, false);
    fails += test_fprop(4, 16, 16, 64, 128, 64, 1, 1, true);
    fails += test_tconv(2, 8, 8, 64, 128, 64, 1, false);
    fails += test_tconv(2, 8, 8, 128, 64, 64, 1, false);
    fails += test_wgrad(8, 16, 16, 64, 128, 5, 64, 7, 2);
    fails += test_wgrad(2, 8, 8, 128, 128, 3, 64, 1, 2);
    dcgan_set_igemm_max_ctas(3);
    fails += test_fprop(2, 16, 16, 128, 128, 64, 1, 1, false);
    fails += test_tconv(2, 8, 8, 64, 128, 64, 1, false);
    dcgan_set_igemm_max_ctas(0);
    for (int ks : {2, 3, 5}) {
      dcgan_set_igemm_ksplit(ks);
      fails += test_fprop(2, 16, 16, 128, 128, 64, 1, 1, false);
      fails += test_tconv(2, 8, 8, 128, 64, 64, 1, false);
    }
    dcgan_set_igemm_ksplit(0);
    dcgan_set_igemm_stage_mode(0);
    printf("small shapes: %d failing tests\n", fails);
    if (!quick) {
      // ---- the real layer shapes, sampled ------------------------------------------------------------------
      for (const Layer& l : kLayers) {
        if (l.kind == 0) fails += test_fprop(l.NB, l.H, l.H, l.Cin, l.Cout, 0, 1, 61, false);
        else if (l.kind == 1) fails += test_tconv(l.NB, l.H, l.H, l.Cin, l.Cout, 0, 61, false);
        else fails += test_wgrad(l.NB, l.H, l.H, l.Cin, l.Cout, 0, 0, 997, 2);
      }
    }
    printf("correctness: %d failing tests\n", fails);
  }
  if (!quick) time_layers(argc >= 3 && !strcmp(argv[2], "sweep"));
  return fails ? 1 : 0;
}
