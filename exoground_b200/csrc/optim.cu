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
                             const float* __restrict__ hyper, const int* __restrict__ blk2seg,
                             const float* __restrict__ seg_sumsq, float clip) {
  if (hyper) { lr = hyper[0]; bc1 = hyper[1]; bc2_sqrt = hyper[2]; }
  const long n4 = n / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* P = &pp.x; const float* G = &gg.x; float* M = &mm.x; float* V = &vv.x;
    float gs = gscale;
    if (blk2seg) {      // utils/train_utils.py:3-13: grad *= clip / (||grad|| + 1e-6) when that is < 1 (per parameter)
      const float nrm = sqrtf(seg_sumsq[blk2seg[i >> 4]]) * gscale;
      gs *= fminf(1.0f, clip / (nrm + 1e-6f));
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gr = G[k] * gs;
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
    float gs = gscale;
    if (blk2seg) gs *= fminf(1.0f, clip / (sqrtf(seg_sumsq[blk2seg[i >> 6]]) * gscale + 1e-6f));
    const float gr = g[i] * gs;
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
      weight_decay, bc1, sqrtf(bc2), grad_scale, nullptr, nullptr, nullptr, 0.f);
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
      weight_decay, 1.f, 1.f, grad_scale, hyper, nullptr, nullptr, 0.f);
  EXG_LAUNCH_CHECK("adamw_dev");
  return 0;
}

// ---------------------------------------------------------------- per-parameter gradient clipping on the flat buffer
// utils/train_utils.py:3-13 (clip_gradients: one .norm() + one .item() + one mul_ per parameter, ~170 parameters) as
// two launches without a host round trip: exg_grad_sumsq accumulates the squared norm of every parameter's gradient
// (parameters start on 64-element boundaries of the flat buffer; blk2seg[b] = parameter of 64-element block b), then
// either exg_grad_clip_scale rescales the gradients in place (the reference's semantics, for callers that read
// .grad) or exg_adamw_clip folds the per-parameter factor into the optimizer's read of the gradient (no extra write).
namespace {
__global__ void grad_sumsq_kernel(const float* __restrict__ g, const int* __restrict__ blk2seg, long n_blocks,
                                  float* __restrict__ seg_sumsq) {
  const int lane = threadIdx.x & 31;
  const long warp = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long n_warps = ((long)gridDim.x * blockDim.x) >> 5;
  // contiguous run of 64-element blocks per warp: consecutive blocks mostly belong to the same parameter
  const long per = (n_blocks + n_warps - 1) / n_warps;
  const long b0 = warp * per, b1 = min(n_blocks, b0 + per);
  float acc = 0.f;
  int cur = -1;
  for (long b = b0; b < b1; ++b) {
    const int seg = blk2seg[b];
    if (seg != cur) {
      if (cur >= 0) { acc = warp_sum(acc); if (lane == 0) atomicAdd(seg_sumsq + cur, acc); }
      cur = seg; acc = 0.f;
    }
    const float2 v = reinterpret_cast<const float2*>(g + b * 64)[lane];
    acc = fmaf(v.x, v.x, fmaf(v.y, v.y, acc));
  }
  if (cur >= 0) { acc = warp_sum(acc); if (lane == 0) atomicAdd(seg_sumsq + cur, acc); }
}

__global__ void grad_clip_scale_kernel(float* __restrict__ g, const int* __restrict__ blk2seg, long n4,
                                       const float* __restrict__ seg_sumsq, float clip, float gscale) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    const float nrm = sqrtf(seg_sumsq[blk2seg[i >> 4]]) * gscale;
    const float c = clip / (nrm + 1e-6f);
    if (c < 1.0f) {
      float4 v = reinterpret_cast<float4*>(g)[i];
      v.x *= c; v.y *= c; v.z *= c; v.w *= c;
      reinterpret_cast<float4*>(g)[i] = v;
    }
  }
}
}  // namespace

extern "C" int exg_grad_sumsq(const float* grad, const int* blk2seg, long n, float* seg_sumsq, int n_seg,
                              void* stream) {
  EXG_CHECK_ARG(grad && blk2seg && seg_sumsq && n_seg > 0, "grad_sumsq: bad argument");
  EXG_CHECK_ARG(n % 64 == 0 && ((uintptr_t)grad & 15) == 0, "grad_sumsq: the flat buffer is a multiple of 64 elements");
  if (n <= 0) return 0;
  EXG_CUDA(cudaMemsetAsync(seg_sumsq, 0, sizeof(float) * (size_t)n_seg, (cudaStream_t)stream));
  ExgProfScope prof(EXG_PC_OPTIM, (double)n * 4.0, stream);
  grad_sumsq_kernel<<<blocks_for(n / 8), 256, 0, (cudaStream_t)stream>>>(grad, blk2seg, n / 64, seg_sumsq);
  EXG_LAUNCH_CHECK("grad_sumsq");
  return 0;
}

extern "C" int exg_grad_clip_scale(float* grad, const int* blk2seg, long n, const float* seg_sumsq, float clip,
                                   float grad_scale, void* stream) {
  EXG_CHECK_ARG(grad && blk2seg && seg_sumsq && clip > 0.f, "grad_clip_scale: bad argument");
  EXG_CHECK_ARG(n % 64 == 0 && ((uintptr_t)grad & 15) == 0, "grad_clip_scale: the flat buffer is a multiple of 64 elements");
  if (n <= 0) return 0;
  ExgProfScope prof(EXG_PC_OPTIM, (double)n * 8.0, stream);
  grad_clip_scale_kernel<<<blocks_for(n / 4), 256, 0, (cudaStream_t)stream>>>(grad, blk2seg, n / 4, seg_sumsq, clip,
                                                                              grad_scale);
  EXG_LAUNCH_CHECK("grad_clip_scale");
  return 0;
}

// AdamW with the per-parameter clip factor applied to the gradient as it is read.  `hyper` (device, optional):
// {lr, 1 - beta1^step, sqrt(1 - beta2^step)}; when null the host values lr / step are used.  blk2seg is indexed from
// the start of THIS segment (pass blk2seg + offset / 64 for a segment that starts at `offset`).
extern "C" int exg_adamw_clip(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                              long n, const float* hyper, float lr, float beta1, float beta2, float eps,
                              float weight_decay, int step, float grad_scale, const int* blk2seg,
                              const float* seg_sumsq, float clip, void* stream) {
  EXG_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && (hyper || step >= 1), "adamw_clip: bad argument");
  EXG_CHECK_ARG(blk2seg && seg_sumsq && clip > 0.f, "adamw_clip: needs the block -> parameter table and the norms");
  EXG_CHECK_ARG(((uintptr_t)param & 15) == 0 && ((uintptr_t)grad & 15) == 0 && ((uintptr_t)exp_avg & 15) == 0 &&
                    ((uintptr_t)exp_avg_sq & 15) == 0 && ((uintptr_t)shadow_bf16 & 7) == 0,
                "adamw_clip: segments must be 16-byte aligned");
  if (n <= 0) return 0;
  float bc1 = 1.f, bc2s = 1.f;
  if (!hyper) { bc1 = 1.0f - powf(beta1, (float)step); bc2s = sqrtf(1.0f - powf(beta2, (float)step)); }
  ExgProfScope prof(EXG_PC_OPTIM, (double)n * (28.0 + (shadow_bf16 ? 2.0 : 0.0)), stream);
  adamw_kernel<<<blocks_for(n / 4 + 1), 256, 0, (cudaStream_t)stream>>>(
      param, grad, exp_avg, exp_avg_sq, reinterpret_cast<uint16_t*>(shadow_bf16), n, lr, beta1, beta2, eps,
      weight_decay, bc1, bc2s, grad_scale, hyper, blk2seg, seg_sumsq, clip);
  EXG_LAUNCH_CHECK("adamw_clip");
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
