// Shared device-side primitives for the sm_100a kernels: mbarrier, TMA, tcgen05/TMEM
// wrappers (inline PTX), bounded waits, and small math helpers.
//
// Everything here is hand-written for Blackwell (sm_100a).  There is no counterpart in the
// reference tree: the reference borrows all device code from the TensorFlow 0.x binary
// (SURVEY.md 2.2 / 2.4).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dcgan {

// ---------------------------------------------------------------------------------------------
// error reporting: kernels never hang.  A wait that exceeds its spin budget sets this flag and
// the CTA drains; the host checks the flag after synchronising (dcgan_last_device_error()).
// ---------------------------------------------------------------------------------------------
// The flag itself lives in csrc/runtime.cu; kernels receive its device address as a parameter
// (dcgan::error_flag_ptr()), so no relocatable device code is needed.
int* error_flag_ptr();       // host: device address of the flag on the current device
int* mn_variant_ptr();       // host: device address of the MN-major descriptor switch

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch: every hot-path kernel lets its successor start early
// (launch_dependents at the top) and waits for its predecessor only right before it first touches
// global memory (pdl_wait), so barrier init / TMEM allocation / shared-memory setup of kernel N+1
// overlap the tail wave of kernel N.  Without the launch attribute both instructions are no-ops.
// ---------------------------------------------------------------------------------------------
bool use_pdl();        // host: DCGAN_PDL=1 / dcgan_set_pdl(1): every hot-path kernel
bool use_pdl_light();  // host: DCGAN_PDL=2 (or 1): only the light kernels (batch norm, heads, linear, casts)
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k_attr(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                 cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                            Args&&... args) {
  return launch_k_attr(use_pdl(), kern, grid, block, smem, st, static_cast<Args&&>(args)...);
}
// Light kernels (a few warps and KB per CTA): launched early they wait inside griddepcontrol.wait without taking
// the SM slots that the side stream's kernels need -- unlike the persistent conv kernels, which would.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k_light(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                  Args&&... args) {
  return launch_k_attr(use_pdl_light(), kern, grid, block, smem, st, static_cast<Args&&>(args)...);
}

enum DeviceError : int {
  kErrNone = 0,
  kErrMbarTimeout = 1,
  kErrPeerTimeout = 2,
  kErrGridBarrierTimeout = 3,
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 1024-byte aligned start of the dynamic shared memory (SWIZZLE_128B atoms).  Pointer arithmetic on the __shared__
// array itself -- a round trip through uintptr_t makes the compiler lose the address space, and every shared-memory
// access of the kernel becomes a generic LD / ST with 64-bit address math instead of LDS / STS.
__device__ __forceinline__ uint8_t* align_smem_1024(uint8_t* raw) {
  return raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
}

// ---------------------------------------------------------------------------------------------
// mbarrier
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// generic-proxy smem writes -> visible to the async proxy (TMA / tcgen05 operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// Bounded wait.  `abort_flag` lives in shared memory; once any role of the CTA gives up, every
// other wait in the CTA returns at its next poll, so a protocol bug costs one time-out, not one
// per wait.  Returns false when aborted.
constexpr uint32_t kSpinLimit = 1u << 22;
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, volatile int* abort_flag,
                                          int* err_flag) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > kSpinLimit || *abort_flag) {
      *abort_flag = 1;
      atomicExch(err_flag, (int)kErrMbarTimeout);
      return false;
    }
  }
  return true;
}

// ---------------------------------------------------------------------------------------------
// TMA (cp.async.bulk.tensor) -- tiled mode, completion on an mbarrier
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// 4-D tiled load (NHWC activations: coordinates {channel, x, y, image}).  Coordinates are signed and may
// lie outside the tensor: out-of-bounds elements arrive as zeros, which is how the convolution padding is
// produced.  With element strides in the tensor map (stride-2 convs) every `stride`-th pixel is fetched.
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0,
                                            int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// tcgen05 / TMEM
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] * B[smem]; bf16 x bf16 -> f32, single CTA.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// One lane of a converged warp.  The MMA / TMA issuing warps run their loops with all 32 lanes (uniform control flow
// lets the compiler keep ring state, addresses and descriptors in uniform registers, which is what UTCHMMA / UTMALDG
// take their operands from); only the issuing instructions are predicated on the elected lane.  A lone lane under
// `if (lane == 0)` measured ~550 cycles of R2UR conversions and dependent integer chains per 4-MMA ring slot.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
  return pred != 0;
}
// SWIZZLE_128B shared-memory descriptor in two words: the high word is constant per operand kind, the low word
// carries the start address (and the leading byte offset) and is what advances from MMA to MMA.
__host__ __device__ constexpr uint32_t smem_desc_hi_sw128(uint32_t sbo_bytes) {
  return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
}
__device__ __forceinline__ uint32_t smem_desc_lo(uint32_t smem_addr, uint32_t lbo_bytes) {
  return ((smem_addr >> 4) & 0x3FFFu) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
__device__ __forceinline__ uint64_t smem_desc_join(uint32_t hi, uint32_t lo) { return ((uint64_t)hi << 32) | lo; }

// all previously issued MMAs of this thread arrive on `bar` when they retire
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// 32 lanes x 32 consecutive columns (one f32 each) -> 32 registers per thread
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32"
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15,"
      " %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31},"
      "[%32];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]),
        "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),
        "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32"
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15},"
      "[%16];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---------------------------------------------------------------------------------------------
// UMMA descriptors (bit layout checked against the CuTe sm100 descriptor headers)
//   smem descriptor: [0,14) start>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1,
//                    [61,64) layout (2 = SWIZZLE_128B)
//   instr descriptor (kind::f16): [4,6) D fmt (1=f32), [7,10) A fmt (1=bf16), [10,13) B fmt,
//                    15 A major (1 = MN), 16 B major, [17,23) N>>3, [24,29) M>>4
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes,
                                                         uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---------------------------------------------------------------------------------------------
// misc
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float bf16_bits_to_f32(uint16_t b) { return __uint_as_float((uint32_t)b << 16); }
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Column sums across the 32 lanes of a warp: lane L ends up with sum over lanes of their v[L].
// Butterfly transpose-reduce: 31 shuffles instead of 32 x 5.
__device__ __forceinline__ float warp_transpose_reduce32(float (&v)[32]) {
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = upper ? v[i] : v[i + off];
      const float keep = upper ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

enum Act : int { kActNone = 0, kActRelu = 1, kActLrelu = 2, kActTanh = 3, kActSigmoid = 4 };

// Per-element activation.  Written for a loop-invariant `act`: identity costs one (uniform) branch, ReLU / LeakyReLU
// are branch-free.  The earlier form -- a switch evaluated per element -- cost four compare-and-branch pairs per
// element on the identity path; in the conv epilogues that was ~11 k cycles per 128x128 tile, as much as the tile's
// main loop (profiles/README.md).  Hot loops should still prefer apply_act_chunk below.
__device__ __forceinline__ float apply_act(float x, int act) {
  if (act == kActNone) return x;
  if (act >= kActTanh) return act == kActTanh ? tanhf(x) : 1.f / (1.f + __expf(-x));
  const float r = fmaxf(x, 0.f), l = fmaxf(x, 0.2f * x);
  return act == kActRelu ? r : l;
}

// Activation of a register chunk with the kind selected ONCE (warp-uniform branch outside the element loop).
template <int N>
__device__ __forceinline__ void apply_act_chunk(float (&f)[N], int act) {
  if (act == kActNone) return;
  if (act == kActRelu) {
#pragma unroll
    for (int j = 0; j < N; ++j) f[j] = fmaxf(f[j], 0.f);
  } else if (act == kActLrelu) {
#pragma unroll
    for (int j = 0; j < N; ++j) f[j] = fmaxf(f[j], 0.2f * f[j]);
  } else if (act == kActTanh) {
#pragma unroll
    for (int j = 0; j < N; ++j) f[j] = tanhf(f[j]);
  } else {
#pragma unroll
    for (int j = 0; j < N; ++j) f[j] = 1.f / (1.f + __expf(-f[j]));
  }
}

// Grid-wide barrier inside one launch.  `bar` = two ints in global memory (arrivals, departures), zero before the
// launch; the last CTA to leave zeroes them again.  Every CTA of the grid must be resident for this to complete:
// launch at most (SMs x CTAs per SM that really fit) CTAs, and do not run two kernels that use it concurrently on one
// device (each could hold SM slots the other is waiting for).  A kernel launched with programmatic dependent launch
// behind this one only starts once every CTA here has started, so it cannot take those slots either.
// Writes made by any thread of any CTA before the barrier are visible to every thread after it (global memory; read
// them with ld.global.cg / atomics-written data likewise, since this SM's L1 may hold older lines it read earlier).
__device__ __forceinline__ void grid_barrier(int* bar, int* err_flag) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const int n = (int)(gridDim.x * gridDim.y * gridDim.z);
    __threadfence();
    atomicAdd(bar, 1);
    const long long t0 = clock64();
    int seen;
    do {
      asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(bar) : "memory");
      if (seen >= n) break;
      __nanosleep(32);
      if (clock64() - t0 > (4LL << 30)) {  // ~2 s: some CTA of the grid never became resident
        if (err_flag != nullptr) atomicExch(err_flag, (int)kErrGridBarrierTimeout);
        break;
      }
    } while (true);
    if (atomicAdd(bar + 1, 1) == n - 1) {  // everybody has seen the full count
      bar[0] = 0;
      bar[1] = 0;
    }
    __threadfence();
  }
  __syncthreads();
}

}  // namespace dcgan
