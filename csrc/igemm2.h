// Internal interface between the C entry points (csrc/igemm_sm100.cu) and the TMA-fed persistent
// implicit-GEMM kernels (csrc/igemm2_sm100.cu).  Not part of the public ABI (include/dcgan_b200.h).
#pragma once
#include <cmath>
#include <cstddef>
#include <cuda_runtime.h>

namespace dcgan {

enum { MODE_F = 0, MODE_T = 1, MODE_W = 2 };

struct Igemm2Params {
  int rH_log2, rW_log2;  // pixel-row space (F: output grid, T: input grid, W: the small grid), powers of two
  int rows_total;        // NB * rH * rW
  int gC;                // channels of the gathered (4-D) tensor
  int Cn;                // F/T: output channels.  W: gathered channels (GEMM N extent)
  int Cd;                // W: dense channels (GEMM M extent)
  int num_kb;            // F: 25 * gC / 64.  W: pixels / 64.  T: derived per parity class
  int m_tiles, n_tiles;  // tile grid (T: per parity class; W: per filter tap)
  int ksplit;            // K slices per tile
  int total_groups;      // work items of the static schedule: tiles * ksplit
  int variant;           // kernel variant of the tile width: 0 default, 1 = two k-blocks per ring slot (64 columns)
  int b_mn;              // F: filter is [25][Ck][Cn] (TensorFlow layout) -> MN-major B operand
  // epilogue
  void* out;              // F/T: bf16 NHWC.  W: f32 [25][Cn][Cd]
  const float* bias;      // F/T optional, per out channel
  int act;                // F/T activation
  int oH, oW;             // F/T output spatial dims
  float* stats;           // F/T optional: [groups][2][Cn] += per-channel (sum, sum of squares), pre-activation
  int stat_group_rows;    // rows per statistics group (multiple of 128)
  // optional: batch-norm backward reductions of the layer BELOW, fused into this data-gradient's epilogue.  The tile
  // just computed is da (gradient w.r.t. that layer's activation); with its saved pre-activation the epilogue forms
  // dy = da * act'(gamma * xhat + beta) and accumulates red[slab][0][c] += sum dy, red[slab][1][c] += sum dy * xhat
  // (what dcgan_bn_bwd_reduce computes in a pass of its own).  stat_group_rows then holds the rows per slab.
  const void* bn_pre;     // bf16 [stat_groups * slab_out_rows][Cn], the layer's pre-activations
  const float* bn_saved;  // [stat_groups][2][Cn] (mean, rstd)
  const float* bn_gamma;
  const float* bn_beta;
  float* bn_red;          // [slabs][2][Cn], zeroed by the caller
  int bn_slab_out_rows;   // output rows (pixels) per slab
  int bn_stat_groups;     // slab g uses the statistics / pre-activations of group min(g, bn_stat_groups - 1)
  int bn_act;             // the layer's activation (ReLU / LeakyReLU)
  int bn_act_only;        // the layer has no batch norm: bn_pre = its saved ACTIVATION, `out` receives dy = da * act'(a),
                          // bn_red = its bias gradient [Cn] += column sums of dy over the first bn_bias_slabs slabs
  int bn_bias_slabs;
  // optional elementwise pass after a grid-wide barrier (every tile stored, every sum complete): the batch-norm
  // forward apply of the layer this conv feeds (tail_kind 1; bn_gamma / bn_beta / bn_act are that layer's, `stats`
  // its sums) or the backward apply of the layer `out` is the activation gradient of (tail_kind 2, after bn_red)
  int tail_kind;
  int rows_out;           // rows of `out`
  void* tail_dst;         // 1: post-activation bf16.  2: gradient w.r.t. the pre-activation, bf16.  Shape of `out`
  float* tail_saved;      // 1: [groups][2][Cn] (mean, rstd) kept for the backward pass
  float* tail_ema_mean;   // 1: optional shadows [groups][Cn]
  float* tail_ema_var;
  float tail_eps, tail_decay;
  float* tail_dgamma;     // 2: optional [Cn], = sum over the first tail_param_slabs slabs
  float* tail_dbeta;
  int tail_param_slabs;
  int* grid_barrier;      // two ints (arrivals, departures), zero between launches
  long long* dbg;         // bring-up only (-DIGEMM2_TIMING): per-CTA cycle counters [CTA][16]
  float* partials;        // split-K scratch: [tile][slice][BLOCK_N / 4][128] float4
  int* counters;          // split-K arrival counters per tile (zero between launches)
  int* err_flag;          // sticky device error flag (csrc/runtime.cu)
  const int* mn_variant;  // MN-major descriptor convention switch (bring-up harness only)
};

// Strided (transposed = 0) or transposed (1) 5x5 / stride-2 conv; see dcgan_conv_fprop* / dcgan_conv_tconv*.
struct BnRedArgs {  // see Igemm2Params::bn_*
  const void* pre;
  const float* saved;
  const float* gamma;
  const float* beta;
  float* red;
  int slabs, stat_groups, act;
  // optional: finish the batch-norm backward pass in the same launch (see Igemm2Params::tail_*)
  void* dpre;
  float* dgamma;
  float* dbeta;
  int param_slabs;
  // activation-only layer below (no batch norm): see Igemm2Params::bn_act_only
  int act_only;
  int bias_slabs;
};
struct BnFwdArgs {  // batch-norm forward apply in the conv's launch; `stats` of igemm2_conv must be given
  void* act_out;
  const float* gamma;
  const float* beta;
  float* saved;
  float* ema_mean;
  float* ema_var;
  float eps, decay;
  int act;
};
int igemm2_conv(int transposed, const void* x, const void* w, const float* bias, void* out, int NB, int H, int W,
                int Cin, int Cout, int act, int block_n, int b_mn, float* stats, int stat_groups,
                cudaStream_t stream, const BnRedArgs* bnred = nullptr, const BnFwdArgs* bnfwd = nullptr);
// Filter gradient; see dcgan_conv_wgrad.
int igemm2_wgrad(const void* big, const void* small, float* dw, int NB, int Hs, int Ws, int Cg, int Cd, int splits,
                 int block_n, cudaStream_t stream);
// (tile width, K slices) for `tiles_bn64` 64-wide tiles of `kb_per_tile` k-blocks on the persistent grid
void igemm2_plan(long tiles_bn64, double kb_per_tile, int Cn, int min_kb, int force_bn, int* block_n, int* ksplit);

// helpers of igemm2_sm100.cu (also used by the bring-up harness)
struct CUtensorMap_st;
int igemm2_tmap_nhwc(const void* ptr, int N, int H, int W, int C, int bw, int bh, int bn, int stride, void* out_map);
int igemm2_tmap_2d(const void* ptr, unsigned long long rows, unsigned long long cols, unsigned box_r, unsigned box_c,
                   void* out_map);
bool igemm2_pixel_box(int rows, int rH, int rW, int* bw, int* bh, int* bn);
int igemm2_attach_scratch(Igemm2Params* p, long tiles, int cols_per_tile);
int igemm2_num_sms();
// scratch area (beyond the counters) of the calling thread's workspace; kernels on one stream may share it
int igemm2_scratch(void** ptr, size_t* bytes);
int igemm2_max_ctas();
int igemm2_forced_ksplit();

void igemm2_set_debug(long long* ptr);
void igemm2_set_max_ctas(int v);
void igemm2_set_ksplit(int v);
void igemm2_set_tile(int v);
// measured plan of one layer shape (kind: MODE_F / MODE_T / MODE_W; W: H, W = the small grid, Cin = Cg, Cout = Cd);
// block_n <= 0 forgets it
int igemm2_plan_set(int kind, int NB, int H, int W, int Cin, int Cout, int block_n, int ksplit, int variant);
void igemm2_set_stage_mode(int v);
int igemm2_set_workspace(void* ptr, size_t bytes);

}  // namespace dcgan
