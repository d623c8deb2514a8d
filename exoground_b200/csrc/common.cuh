// common.cuh — shared helpers for libexg.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/exg.h"

#define EXG_WARP 32

// ---------------------------------------------------------------- error plumbing
void exg_set_error(const char* fmt, ...);
void exg_count_launch(int n = 1);

#define EXG_CHECK_ARG(cond, ...)                       \
  do {                                                 \
    if (!(cond)) {                                     \
      exg_set_error(__VA_ARGS__);                      \
      return EXG_EINVAL;                               \
    }                                                  \
  } while (0)

#define EXG_LAUNCH_CHECK(name)                                              \
  do {                                                                      \
    cudaError_t e__ = cudaGetLastError();                                   \
    if (e__ != cudaSuccess) {                                               \
      exg_set_error("%s: launch failed: %s", name, cudaGetErrorString(e__)); \
      return (int)e__;                                                      \
    }                                                                       \
    exg_count_launch();                                                     \
  } while (0)

#define EXG_CUDA(call)                                                           \
  do {                                                                           \
    cudaError_t e__ = (call);                                                    \
    if (e__ != cudaSuccess) {                                                    \
      exg_set_error("%s failed: %s", #call, cudaGetErrorString(e__));            \
      return (int)e__;                                                           \
    }                                                                            \
  } while (0)

// ---------------------------------------------------------------- per-class launch profiling
// exg_profile_begin()/exg_profile_end(): while active, every launch site brackets its kernel with a CUDA
// event pair on the launching stream (bench.py's roofline numbers come from these).
enum ExgProfClass {
  EXG_PC_GEMM_TC = 0, EXG_PC_GEMM_SIMT, EXG_PC_ATTN_FWD, EXG_PC_ATTN_BWD, EXG_PC_LN_FWD, EXG_PC_LN_BWD,
  EXG_PC_ROWWISE, EXG_PC_NCE_FWD, EXG_PC_NCE_BWD, EXG_PC_OPTIM, EXG_PC_COUNT
};
bool exg_profiling();
void exg_prof_open(int cls, double work, cudaStream_t s);
void exg_prof_close(cudaStream_t s);
struct ExgProfScope {
  cudaStream_t s; bool on;
  ExgProfScope(int cls, double work, void* stream) : s((cudaStream_t)stream), on(exg_profiling()) {
    if (on) exg_prof_open(cls, work, s);
  }
  ~ExgProfScope() { if (on) exg_prof_close(s); }
};

static inline bool exg_valid_dtype(int d) { return d == EXG_F32 || d == EXG_BF16; }
static inline size_t exg_dtype_size(int d) { return d == EXG_F32 ? 4 : 2; }

// ---------------------------------------------------------------- device helpers
// Row r of a dense [rows, C] view -> row of the strided destination (video / text rows inside the joint sequence).
// Row counts fit 32 bits on this path (checked on the host: EXG_MAX_ROWS), so this is ONE 32-bit division; the two
// 64-bit ones it replaces were ~250 of the ~400 instructions a LayerNorm thread issued per row (ncu: issue slots 50 %
// busy in an HBM-bound kernel, profiles/ncu_full_r02v_ln_summary.txt).
__device__ __forceinline__ long exg_map_row(const exg_rowmap& m, long r) {
  if (m.group > r) return r + m.offset;                   // identity maps (group = rows) and the first group
  const unsigned g = (unsigned)m.group, rr = (unsigned)r;
  const unsigned q = rr / g;
  return (long)q * m.group_stride + (long)(rr - q * g) + m.offset;
}
constexpr long EXG_MAX_ROWS = 0x7fffffffL;

__device__ __forceinline__ float bf16_bits_to_float(uint16_t b) {
  return __uint_as_float(((uint32_t)b) << 16);
}
__device__ __forceinline__ uint16_t float_to_bf16_bits(float f) {
  __nv_bfloat16 h = __float2bfloat16_rn(f);
  return *reinterpret_cast<uint16_t*>(&h);
}

// Load 4 consecutive elements (element index `i`, multiple of 4) as float4 from a dtype-tagged buffer.
__device__ __forceinline__ float4 load4(const void* p, int dtype, long i) {
  if (dtype == EXG_F32) {
    return *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + i);
  } else {
    uint2 u = *reinterpret_cast<const uint2*>(reinterpret_cast<const uint16_t*>(p) + i);
    float4 r;
    r.x = __uint_as_float(u.x << 16);
    r.y = __uint_as_float(u.x & 0xffff0000u);
    r.z = __uint_as_float(u.y << 16);
    r.w = __uint_as_float(u.y & 0xffff0000u);
    return r;
  }
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void store4(void* p, int dtype, long i, float4 v) {
  if (dtype == EXG_F32) {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + i) = v;
  } else {
    uint2 u;
    u.x = pack_bf16x2(v.x, v.y);
    u.y = pack_bf16x2(v.z, v.w);
    *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(p) + i) = u;
  }
}
__device__ __forceinline__ float load1(const void* p, int dtype, long i) {
  if (dtype == EXG_F32) return reinterpret_cast<const float*>(p)[i];
  return bf16_bits_to_float(reinterpret_cast<const uint16_t*>(p)[i]);
}
__device__ __forceinline__ void store1(void* p, int dtype, long i, float v) {
  if (dtype == EXG_F32) reinterpret_cast<float*>(p)[i] = v;
  else reinterpret_cast<uint16_t*>(p)[i] = float_to_bf16_bits(v);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// x * sigmoid(1.702 x) (model/tfm_model.py:11-13).  __fdividef (2 ulp) keeps the fp32 paths free of the
// IEEE-division slow path.
__device__ __forceinline__ float quick_gelu(float v) { return __fdividef(v, 1.0f + __expf(-1.702f * v)); }
__device__ __forceinline__ float quick_gelu_grad(float u) {
  float s = __fdividef(1.0f, 1.0f + __expf(-1.702f * u));
  return s + 1.702f * u * s * (1.0f - s);
}
// bf16-output variants for the tensor-core GEMM epilogues: sigmoid(x) = 0.5 tanh(x/2) + 0.5 with the single-MUFU
// tanh.approx.f32 (rel. error ~2^-11, below bf16 resolution).  ex2 + rcp would be two quarter-rate MUFU ops per
// element and make the c_fc / c_proj-backward epilogues MUFU-bound (128x256 tile: 4096 MUFU cycles = the MMA time).
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float quick_gelu_fast(float v) {
  const float h = 0.5f * v;
  return fmaf(h, tanh_fast(0.851f * v), h);
}
__device__ __forceinline__ float quick_gelu_grad_fast(float u) {
  const float s = fmaf(0.5f, tanh_fast(0.851f * u), 0.5f);
  return fmaf(1.702f * u * s, 1.0f - s, s);
}

int exg_persistent_sms();      // exg_num_sms() minus the SMs reserved through exg_reserve_sms (abi.cu)

static inline int exg_num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}
