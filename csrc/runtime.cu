// Library-wide runtime state: the sticky device error flag that bounded waits set instead of
// hanging, and its host accessors.  No counterpart in the reference (its runtime is TensorFlow's).
#include <cstdlib>
#include <cuda_runtime.h>
#include "common.cuh"
#include "../include/dcgan_b200.h"

namespace dcgan {

static __device__ int g_device_error = 0;
static __device__ int g_mn_desc_variant = 0;

constexpr int kMaxDevices = 64;

int* error_flag_ptr() {
  static int* ptrs[kMaxDevices] = {nullptr};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  if (!ptrs[dev]) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_device_error) != cudaSuccess) return nullptr;
    ptrs[dev] = static_cast<int*>(p);
  }
  return ptrs[dev];
}

void set_pdl(int v);

int* mn_variant_ptr() {
  static int* ptrs[kMaxDevices] = {nullptr};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  if (!ptrs[dev]) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_mn_desc_variant) != cudaSuccess) return nullptr;
    ptrs[dev] = static_cast<int*>(p);
  }
  return ptrs[dev];
}

// 0 = off, 1 = every hot-path kernel, 2 = the light kernels only (batch norm, heads, linear, casts).
// Measured (profiles/README.md): with 1 the early-launched persistent conv CTAs sit on the SM slots the side
// stream's kernels need: 0.665 -> 0.772 ms/step.
#ifndef DCGAN_PDL_DEFAULT
#define DCGAN_PDL_DEFAULT 0
#endif
static int g_use_pdl = -1;
static int pdl_mode() {
  if (g_use_pdl < 0) {
    const char* e = getenv("DCGAN_PDL");
    g_use_pdl = (e != nullptr && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : DCGAN_PDL_DEFAULT;
  }
  return g_use_pdl;
}
bool use_pdl() { return pdl_mode() == 1; }
bool use_pdl_light() { return pdl_mode() >= 1; }
void set_pdl(int v) { g_use_pdl = (v >= 0 && v <= 2) ? v : 0; }

}  // namespace dcgan

extern "C" {

void dcgan_set_pdl(int v) { dcgan::set_pdl(v); }


int dcgan_last_device_error(void) {
  int v = 0;
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(&v, dcgan::g_device_error, sizeof(int)) != cudaSuccess) return -1;
  if (v != 0) {
    int z = 0;
    cudaMemcpyToSymbol(dcgan::g_device_error, &z, sizeof(int));
  }
  return v;
}

int dcgan_copy_error_flag(int* host_dst, cudaStream_t stream) {
  int* flag = dcgan::error_flag_ptr();
  if (!flag || !host_dst) return DCGAN_ERR_DRIVER;
  return cudaMemcpyAsync(host_dst, flag, sizeof(int), cudaMemcpyDeviceToHost, stream) == cudaSuccess
             ? 0 : DCGAN_ERR_LAUNCH;
}

// Stream-ordered zero fill: one memset node when captured into a CUDA graph (the gradient / reduction buffers of the
// training step), instead of a fill kernel launch.
int dcgan_zero_async(void* ptr, size_t bytes, cudaStream_t stream) {
  if (!ptr || bytes == 0) return 0;
  return cudaMemsetAsync(ptr, 0, bytes, stream) == cudaSuccess ? 0 : DCGAN_ERR_LAUNCH;
}

void dcgan_set_mn_desc_variant(int v) { cudaMemcpyToSymbol(dcgan::g_mn_desc_variant, &v, sizeof(int)); }

}  // extern "C"
