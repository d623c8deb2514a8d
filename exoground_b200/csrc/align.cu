// align.cu — device side of the sliding-window alignment evaluation (eval/eval_zeroshot_align.py:200-237):
// stitch the per-window similarity blocks into the [sentence, second] matrix (overlap average of joint and dual
// logits), apply the reference's `sim == 0 -> -6e4` rule, softmax over time, argmax and row maximum - ONE launch
// instead of index_add_ x3, maximum, div x2, add, masked_fill, softmax, argmax, max.
//
// One CTA per sentence k.  A second t of sentence k receives the windows w with l_w <= k <= r_w and
// s_w <= t < s_w + len_w; they are summed in window order (the order of the reference's `logits[l:r+1, s:s+len] += ..`
// loop), so the fp32 result does not depend on scheduling (index_add_ uses atomics).  HBM-bound and tiny: reads
// 2 * W * T * N floats once, writes 2 * K * vlen.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxWin = 1024;          // windows of one video kept in shared memory (HTM-Align: <= ~130)

struct Win { int s, len, l, r; };

__global__ void __launch_bounds__(kThreads)
align_finish_kernel(const float* __restrict__ sj, const float* __restrict__ sd, const int* __restrict__ win, int W,
                    int T, int N, int K, int vlen, float* __restrict__ sim_out, float* __restrict__ prob_out,
                    long long* __restrict__ pred_out, float* __restrict__ max_out) {
  extern __shared__ float row[];                       // [vlen] stitched similarities of this sentence
  __shared__ Win wl[kMaxWin];                          // the windows that contain sentence k
  __shared__ int n_w;
  __shared__ float red_v[kThreads / 32];
  __shared__ int red_i[kThreads / 32];
  __shared__ float bc_v;
  __shared__ int bc_i;
  const int k = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) {                                      // ordered list (window order = the reference's summation order)
    int n = 0;
    for (int w = 0; w < W; ++w) {
      const int l = win[4 * w + 2], r = win[4 * w + 3];
      if (l <= k && k <= r && n < kMaxWin) { wl[n].s = win[4 * w]; wl[n].len = win[4 * w + 1]; wl[n].l = w; wl[n].r = l; ++n; }
    }
    n_w = n;
  }
  __syncthreads();
  const int nw = n_w;
  // ---- stitch + overlap average + zero rule; running (max, first argmax)
  float best = -INFINITY; int besti = 0x7fffffff;
  for (int t = tid; t < vlen; t += kThreads) {
    float aj = 0.f, ad = 0.f, cnt = 0.f;
    for (int i = 0; i < nw; ++i) {
      const int dt = t - wl[i].s;
      if (dt >= 0 && dt < wl[i].len) {
        const size_t off = ((size_t)wl[i].l * T + dt) * N + (k - wl[i].r);      // wl.l = window index, wl.r = its first sentence
        aj += sj[off]; ad += sd[off]; cnt += 1.f;
      }
    }
    const float den = fmaxf(cnt, 1e-5f);
    float v = (aj / den + ad / den) * 0.5f;                                     // (logits + logits_dual) / 2  (:219)
    if (v == 0.f) v = -6e4f;                                                    // sim[sim == 0] = -6e4        (:220)
    row[t] = v;
    sim_out[(size_t)k * vlen + t] = v;
    if (v > best) { best = v; besti = t; }
  }
  // block argmax (first index among equal maxima)
  auto block_argmax = [&](float v, int i) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, v, o);
      const int oi = __shfl_xor_sync(0xffffffffu, i, o);
      if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
    }
    if (lane == 0) { red_v[wid] = v; red_i[wid] = i; }
    __syncthreads();
    if (wid == 0) {
      v = lane < kThreads / 32 ? red_v[lane] : -INFINITY;
      i = lane < kThreads / 32 ? red_i[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, i, o);
        if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
      }
      if (lane == 0) { bc_v = v; bc_i = i; }
    }
    __syncthreads();
  };
  block_argmax(best, besti);
  const float mx = bc_v;
  if (tid == 0) max_out[k] = mx;
  __syncthreads();
  // ---- softmax over time
  float s = 0.f;
  for (int t = tid; t < vlen; t += kThreads) { const float e = expf(row[t] - mx); row[t] = e; s += e; }
  s = warp_sum(s);
  if (lane == 0) red_v[wid] = s;
  __syncthreads();
  if (wid == 0) {
    float v = lane < kThreads / 32 ? red_v[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) bc_v = v;
  }
  __syncthreads();
  const float inv = 1.0f / bc_v;
  __syncthreads();
  // prob and its argmax (what the reference takes: softmax(...).argmax(-1), first index among equal values)
  best = -INFINITY; besti = 0x7fffffff;
  for (int t = tid; t < vlen; t += kThreads) {
    const float pv = row[t] * inv;
    prob_out[(size_t)k * vlen + t] = pv;
    if (pv > best) { best = pv; besti = t; }
  }
  block_argmax(best, besti);
  if (tid == 0) pred_out[k] = (long long)bc_i;
}

}  // namespace

extern "C" int exg_align_finish(const float* sj, const float* sd, const int* win, int W, int T, int N, int K,
                                int vlen, float* sim, float* prob, long long* pred, float* row_max, void* stream) {
  EXG_CHECK_ARG(sj && sd && win && sim && prob && pred && row_max, "align_finish: null pointer");
  EXG_CHECK_ARG(W >= 0 && T > 0 && N > 0 && K > 0 && vlen > 0, "align_finish: bad sizes");
  EXG_CHECK_ARG(W <= kMaxWin, "align_finish: more than %d windows per video", kMaxWin);
  EXG_CHECK_ARG((size_t)vlen * 4 <= 200 * 1024, "align_finish: video longer than 51200 seconds");
  const size_t smem = (size_t)vlen * sizeof(float);
  static size_t attr = 48 * 1024;
  if (smem > attr) {
    EXG_CUDA(cudaFuncSetAttribute(align_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = smem;
  }
  ExgProfScope prof(EXG_PC_ROWWISE, 2.0 * W * T * N * 4.0 + 2.0 * K * vlen * 4.0, stream);
  align_finish_kernel<<<K, kThreads, smem, (cudaStream_t)stream>>>(sj, sd, win, W, T, N, K, vlen, sim, prob, pred,
                                                                   row_max);
  EXG_LAUNCH_CHECK("align_finish");
  return 0;
}
