/* C ABI of libdcgan_b200.so -- the hand-written sm_100a kernels of the B200-native DCGAN framework.
 *
 * Every entry point takes raw device pointers plus a CUDA stream, launches asynchronously on that
 * stream and returns 0 on success or a negative DCGAN_ERR_* code.  None of them synchronises.  The
 * Python package binds these with ctypes (distributed_tensorflow_for_dcgan_b200/ops/_lib.py).
 *
 * Tensor conventions: activations are NHWC bf16; master weights / gradients / optimizer state are
 * f32 in TensorFlow layouts (conv HWIO, deconv [kh,kw,Cout,Cin], linear [in,out]) so that the
 * checkpoint maps one-to-one on the reference's variable names.
 * Geometry is TensorFlow's 5x5 / stride 2 / SAME: one pixel of padding before, two after.
 */
#ifndef DCGAN_B200_H_
#define DCGAN_B200_H_

#include <cuda_runtime_api.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DCGAN_ERR_SHAPE (-1)
#define DCGAN_ERR_DRIVER (-2)
#define DCGAN_ERR_LAUNCH (-3)
#define DCGAN_ERR_LIMIT (-4)

#define DCGAN_ACT_NONE 0
#define DCGAN_ACT_RELU 1
#define DCGAN_ACT_LRELU 2
#define DCGAN_ACT_TANH 3
#define DCGAN_ACT_SIGMOID 4

/* Reads and clears the sticky device-side error flag (0 = none, 1 = mbarrier wait timed out,
 * 2 = peer signal wait timed out).  Synchronises the device. */
int dcgan_last_device_error(void);

/* Bring-up switch for the MN-major UMMA descriptor convention (0 = CuTe convention). */
void dcgan_set_mn_desc_variant(int v);

/* ---- implicit-GEMM convolutions on tcgen05 / TMEM / TMA (csrc/igemm_sm100.cu) ------------------ */

/* Strided forward conv 5x5 s2 SAME: out[NB,H/2,W/2,Cout] = act(conv(x[NB,H,W,Cin]) + bias).
 * w_nk is bf16 [25][Cout][Cin] (tap-major, reduction channel contiguous).
 * Same math as tf.nn.conv2d at reference distriubted_model.py:183 and as the data gradient of the
 * transposed conv at :200.  block_n <= 0 picks a tile width. */
int dcgan_conv_fprop(const void* x, const void* w_nk, const float* bias, void* out, int NB, int H, int W,
                     int Cin, int Cout, int act, int block_n, cudaStream_t stream);

/* Transposed conv 5x5 s2 (TF crop 1 before / 2 after): out[NB,2H,2W,Cout] = act(deconv(x) + bias).
 * w_nk is bf16 [25][Cout][Cin].  Same math as tf.nn.conv2d_transpose at reference
 * distriubted_model.py:200 and as the data gradient of conv2d at :183. */
int dcgan_conv_tconv(const void* x, const void* w_nk, const float* bias, void* out, int NB, int H, int W,
                     int Cin, int Cout, int act, int block_n, cudaStream_t stream);

/* Weight gradient of either layer kind: dw[25][Cg][Cd] (f32, accumulated with atomics -- zero it
 * first) += sum over pixels of big[NB,2Hs,2Ws,Cg] (gathered at 2q+k-1) * small[NB,Hs,Ws,Cd].
 * conv2d:  big = layer input, small = output gradient  -> dw is HWIO.
 * deconv:  big = output gradient, small = layer input  -> dw is [kh,kw,Cout,Cin].
 * (autodiff of reference image_train.py:109-112).  splits <= 0 picks a split-K factor. */
int dcgan_conv_wgrad(const void* big, const void* small, float* dw, int NB, int Hs, int Ws, int Cg, int Cd,
                     int splits, int block_n, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DCGAN_B200_H_ */
