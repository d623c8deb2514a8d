// optim.cu — fused AdamW over flat fp32 segments (+ bf16 shadow weights) and EMA update.
// Replaces torch.optim.AdamW(params, lr, weight_decay) (train/main.py:418,
// train/main_egoexo4d_distributed.py:721) and TwinTemporalAligner._momentum_update
// (model/tan_model.py:334-338).  HBM-bound: 4 fp32 reads + 3 fp32 writes (+1 bf16 write) per element.
#include "common.cuh"

namespace {

// hyper (optional, device): {lr, 1 - beta1^step, sqrt(1 - beta2^step)} read at run time, so a launch captured into a
// CUDA graph follows the learning-rate schedule and the bias corrections of the step it is replayed for.
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                             float* __restrict__ v, uint16_t* __restrict__ shadow, long n, float lr, float b1,
                             float b2, float eps, float wd, float bc1, float bc2_sqrt, float gscale,
                             const float* __restrict__ hyper) {
  if (hyper) { lr = hyper[0]; bc1 = hyper[1]; bc2_sqrt = hyper[2]; }
  const long n4 = n / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* P = &pp.x; const float* G = &gg.x; float* M = &mm.x; float* V = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gr = G[k] * gscale;
      P[k] *= 1.0f - lr * wd;                       // decoupled weight decay
      M[k] = b1 * M[k] + (1.0f - b1) * gr;
      V[k] = b2 * V[k] + (1.0f - b2) * gr * gr;
      const float denom = sqrtf(V[k]) / bc2_sqrt + eps;
      P[k] -= (lr / bc1) * (M[k] / denom);
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (shadow) {
      uint2 u; u.x = pack_bf16x2(pp.x, pp.y); u.y = pack_bf16x2(pp.z, pp.w);
      reinterpret_cast<uint2*>(shadow)[i] = u;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long i = n4 * 4 + threadIdx.x;
    const float gr = g[i] * gscale;
    float pv = p[i] * (1.0f - lr * wd);
    const float mv = b1 * m[i] + (1.0f - b1) * gr;
    const float vv = b2 * v[i] + (1.0f - b2) * gr * gr;
    pv -= (lr / bc1) * (mv / (sqrtf(vv) / bc2_sqrt + eps));
    p[i] = pv; m[i] = mv; v[i] = vv;
    if (shadow) shadow[i] = float_to_bf16_bits(pv);
  }
}

__global__ void ema_kernel(float* __restrict__ tgt, const float* __restrict__ onl, long n, float mom) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    tgt[i] = tgt[i] * mom + onl[i] * (1.0f - mom);
}

int blocks_for(long n) {
  long b = (n + 255) / 256;
  long cap = 8L * exg_num_sms();
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace

extern "C" int exg_adamw(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                         long n, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                         float grad_scale, void* stream) {
  EXG_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && step >= 1, "adamw: bad argument");
  EXG_CHECK_ARG(((uintptr_t)param & 15) == 0 && ((uintptr_t)grad & 15) == 0 && ((uintptr_t)exp_avg & 15) == 0 &&
                    ((uintptr_t)exp_avg_sq & 15) == 0 && ((uintptr_t)shadow_bf16 & 7) == 0,
                "adamw: segments must be 16-byte aligned");
  if (n <= 0) return 0;
  const float bc1 = 1.0f - powf(beta1, (float)step);
  const float bc2 = 1.0f - powf(beta2, (float)step);
  ExgProfScope prof(EXG_PC_OPTIM, (double)n * (28.0 + (shadow_bf16 ? 2.0 : 0.0)), stream);
  adamw_kernel<<<blocks_for(n / 4 + 1), 256, 0, (cudaStream_t)stream>>>(
      param, grad, exp_avg, exp_avg_sq, reinterpret_cast<uint16_t*>(shadow_bf16), n, lr, beta1, beta2, eps,
      weight_decay, bc1, sqrtf(bc2), grad_scale, nullptr);
  EXG_LAUNCH_CHECK("adamw");
  return 0;
}

extern "C" int exg_adamw_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                             long n, const float* hyper, float beta1, float beta2, float eps, float weight_decay,
                             float grad_scale, void* stream) {
  EXG_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && hyper, "adamw_dev: bad argument");
  EXG_CHECK_ARG(((uintptr_t)param & 15) == 0 && ((uintptr_t)grad & 15) == 0 && ((uintptr_t)exp_avg & 15) == 0 &&
                    ((uintptr_t)exp_avg_sq & 15) == 0 && ((uintptr_t)shadow_bf16 & 7) == 0,
                "adamw_dev: segments must be 16-byte aligned");
  if (n <= 0) return 0;
  ExgProfScope prof(EXG_PC_OPTIM, (double)n * (28.0 + (shadow_bf16 ? 2.0 : 0.0)), stream);
  adamw_kernel<<<blocks_for(n / 4 + 1), 256, 0, (cudaStream_t)stream>>>(
      param, grad, exp_avg, exp_avg_sq, reinterpret_cast<uint16_t*>(shadow_bf16), n, 0.f, beta1, beta2, eps,
      weight_decay, 1.f, 1.f, grad_scale, hyper);
  EXG_LAUNCH_CHECK("adamw_dev");
  return 0;
}

extern "C" int exg_ema(float* target, const float* online, long n, float m, void* stream) {
  EXG_CHECK_ARG(target && online, "ema: null pointer");
  if (n <= 0) return 0;
  ExgProfScope prof(EXG_PC_OPTIM, (double)n * 12.0, stream);
  ema_kernel<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(target, online, n, m);
  EXG_LAUNCH_CHECK("ema");
  return 0;
}
