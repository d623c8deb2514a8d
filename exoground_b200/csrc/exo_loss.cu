// exo_loss.cu — ExoGround loss kernels (train/loss_egoexo4d.py, model/exo_ground_model.py:206-234):
//   * row-wise softmax cross-entropy with the target on the diagonal (the per-clip T x T InfoNCE of
//     compute_info_nce_loss, after the normalised-feature GEMM),
//   * fused cross-view InfoNCE: gather the positive / negative view rows, three normalised dot products over
//     D per (clip, second), 2- or 3-way log-softmax, plus the L1 / cosine monitors
//     (compute_info_nce_loss_cross_view :269-323, compute_l1_cosine_losses :201-233) and its gradient,
//   * the grounding regression loss: L1(center, duration | start, end) + IoU, IoU@theta counts and the gradient
//     w.r.t. the predictions in one launch (get_grounding_loss_reg_head :73-145), no host round trips.
// All HBM-bound: one warp per row, 16-byte loads.
#include "common.cuh"
#include <math.h>

namespace {

constexpr int kWarps = 4;

// ------------------------------------------------------------------ diagonal cross-entropy
__global__ void __launch_bounds__(kWarps * 32)
xent_diag_fwd_kernel(const void* __restrict__ x, int dt, float* __restrict__ loss, float* __restrict__ lse,
                     long rows, int cols, int period) {
  const int lane = threadIdx.x & 31;
  const long r = (long)blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (r >= rows) return;
  float mx = -INFINITY;
  for (int c = lane; c < cols; c += 32) mx = fmaxf(mx, load1(x, dt, r * cols + c));
  mx = warp_max(mx);
  float s = 0.f;
  for (int c = lane; c < cols; c += 32) s += __expf(load1(x, dt, r * cols + c) - mx);
  s = warp_sum(s);
  if (lane == 0) {
    const float l = mx + __logf(s);
    lse[r] = l;
    loss[r] = l - load1(x, dt, r * cols + (int)(r % period));
  }
}

__global__ void __launch_bounds__(kWarps * 32)
xent_diag_bwd_kernel(const void* __restrict__ x, int dt, const float* __restrict__ lse, float scale,
                     void* __restrict__ g, int gdt, long rows, int cols, int period) {
  const int lane = threadIdx.x & 31;
  const long r = (long)blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (r >= rows) return;
  const float l = lse[r];
  const int tgt = (int)(r % period);
  for (int c = lane; c < cols; c += 32) {
    const float p = __expf(load1(x, dt, r * cols + c) - l);
    store1(g, gdt, r * cols + c, scale * (p - (c == tgt ? 1.f : 0.f)));
  }
}

// ------------------------------------------------------------------ cross-view InfoNCE
struct XvArgs {
  const void* o; int odt;                 // [B,T,D] projected exo features
  const float* ego;                       // [B,V,T,D]
  const long* pos_idx; const long* neg_idx;   // [B,T]
  const float* same; const long* same_idx;    // [B,T,D], [B,T] or NULL
  int only_same;
  int B, V, T, D;
  float inv_temp;
  float* loss; float* l1; float* pos_cos; float* neg_cos;   // [B,T]
  float* stats;                           // [B,T,8]: 1/|o|, c_p, c_n, c_s, w_p, w_n, w_s, -
  const float* gl; void* go; int godt;    // backward: upstream per-(b,t) gradient, d o
};

__device__ __forceinline__ const float* view_row(const XvArgs& p, long idx, int b, int t) {
  return p.ego + (((long)b * p.V + idx) * p.T + t) * p.D;
}

__global__ void __launch_bounds__(kWarps * 32) xview_fwd_kernel(XvArgs p) {
  const int lane = threadIdx.x & 31;
  const long bt = (long)blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (bt >= (long)p.B * p.T) return;
  const int b = (int)(bt / p.T), t = (int)(bt % p.T);
  const float* pr = view_row(p, p.pos_idx[bt], b, t);
  const float* nr = view_row(p, p.neg_idx[bt], b, t);
  const float* sr = p.same ? p.same + ((long)b * p.T + p.same_idx[bt]) * p.D : nullptr;
  float oo = 0.f, pp = 0.f, nn = 0.f, ss = 0.f, op = 0.f, on = 0.f, os = 0.f;
  for (int c = lane * 4; c < p.D; c += 128) {
    const float4 o = load4(p.o, p.odt, bt * p.D + c);
    const float4 a = *reinterpret_cast<const float4*>(pr + c);
    const float4 n = *reinterpret_cast<const float4*>(nr + c);
    oo += o.x * o.x + o.y * o.y + o.z * o.z + o.w * o.w;
    pp += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
    nn += n.x * n.x + n.y * n.y + n.z * n.z + n.w * n.w;
    op += o.x * a.x + o.y * a.y + o.z * a.z + o.w * a.w;
    on += o.x * n.x + o.y * n.y + o.z * n.z + o.w * n.w;
    if (sr) {
      const float4 s = *reinterpret_cast<const float4*>(sr + c);
      ss += s.x * s.x + s.y * s.y + s.z * s.z + s.w * s.w;
      os += o.x * s.x + o.y * s.y + o.z * s.z + o.w * s.w;
    }
  }
  oo = warp_sum(oo); pp = warp_sum(pp); nn = warp_sum(nn); ss = warp_sum(ss);
  op = warp_sum(op); on = warp_sum(on); os = warp_sum(os);
  const float io = 1.f / fmaxf(sqrtf(oo), 1e-12f), ip = 1.f / fmaxf(sqrtf(pp), 1e-12f);   // F.normalize eps
  const float in_ = 1.f / fmaxf(sqrtf(nn), 1e-12f), is = 1.f / fmaxf(sqrtf(ss), 1e-12f);
  const float cp = op * io * ip, cn = on * io * in_, cs = os * io * is;
  // L1 monitor needs a second pass over the normalised vectors
  float l1 = 0.f;
  for (int c = lane * 4; c < p.D; c += 128) {
    const float4 o = load4(p.o, p.odt, bt * p.D + c);
    const float4 a = *reinterpret_cast<const float4*>(pr + c);
    l1 += fabsf(o.x * io - a.x * ip) + fabsf(o.y * io - a.y * ip) + fabsf(o.z * io - a.z * ip) +
          fabsf(o.w * io - a.w * ip);
  }
  l1 = warp_sum(l1);
  if (lane == 0) {
    const float sp = cp * p.inv_temp, sn = cn * p.inv_temp, s3 = cs * p.inv_temp;
    const bool use_n = !(sr && p.only_same), use_s = sr != nullptr;
    float mx = sp;
    if (use_n) mx = fmaxf(mx, sn);
    if (use_s) mx = fmaxf(mx, s3);
    const float ep = __expf(sp - mx), en = use_n ? __expf(sn - mx) : 0.f, es = use_s ? __expf(s3 - mx) : 0.f;
    const float z = ep + en + es;
    p.loss[bt] = mx + __logf(z) - sp;
    p.l1[bt] = l1 / (float)p.D;
    p.pos_cos[bt] = cp;
    p.neg_cos[bt] = cn;
    float* st = p.stats + bt * 8;
    st[0] = io; st[1] = cp; st[2] = cn; st[3] = cs;
    st[4] = ep / z - 1.f; st[5] = en / z; st[6] = es / z; st[7] = 0.f;
  }
}

// d o = g * inv_temp / |o| * sum_k w_k (k_hat - o_hat c_k)
__global__ void __launch_bounds__(kWarps * 32) xview_bwd_kernel(XvArgs p) {
  const int lane = threadIdx.x & 31;
  const long bt = (long)blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (bt >= (long)p.B * p.T) return;
  const int b = (int)(bt / p.T), t = (int)(bt % p.T);
  const float* pr = view_row(p, p.pos_idx[bt], b, t);
  const float* nr = view_row(p, p.neg_idx[bt], b, t);
  const float* sr = p.same ? p.same + ((long)b * p.T + p.same_idx[bt]) * p.D : nullptr;
  const float* st = p.stats + bt * 8;
  const float io = st[0], cp = st[1], cn = st[2], cs = st[3], wp = st[4], wn = st[5], ws = st[6];
  float pp = 0.f, nn = 0.f, ss = 0.f;
  for (int c = lane * 4; c < p.D; c += 128) {
    const float4 a = *reinterpret_cast<const float4*>(pr + c);
    const float4 n = *reinterpret_cast<const float4*>(nr + c);
    pp += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
    nn += n.x * n.x + n.y * n.y + n.z * n.z + n.w * n.w;
    if (sr) {
      const float4 s = *reinterpret_cast<const float4*>(sr + c);
      ss += s.x * s.x + s.y * s.y + s.z * s.z + s.w * s.w;
    }
  }
  const float ip = 1.f / fmaxf(sqrtf(warp_sum(pp)), 1e-12f), in_ = 1.f / fmaxf(sqrtf(warp_sum(nn)), 1e-12f);
  const float is = 1.f / fmaxf(sqrtf(warp_sum(ss)), 1e-12f);
  const float k = p.gl[bt] * p.inv_temp * io;
  const float ko = -(wp * cp + wn * cn + ws * cs) * io;       // coefficient of o (o_hat = o * io)
  for (int c = lane * 4; c < p.D; c += 128) {
    const float4 o = load4(p.o, p.odt, bt * p.D + c);
    const float4 a = *reinterpret_cast<const float4*>(pr + c);
    const float4 n = *reinterpret_cast<const float4*>(nr + c);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (sr) s = *reinterpret_cast<const float4*>(sr + c);
    float4 g;
    g.x = k * (wp * ip * a.x + wn * in_ * n.x + ws * is * s.x + ko * o.x);
    g.y = k * (wp * ip * a.y + wn * in_ * n.y + ws * is * s.y + ko * o.y);
    g.z = k * (wp * ip * a.z + wn * in_ * n.z + ws * is * s.z + ko * o.z);
    g.w = k * (wp * ip * a.w + wn * in_ * n.w + ws * is * s.w + ko * o.w);
    store4(p.go, p.godt, bt * p.D + c, g);
  }
}

// ------------------------------------------------------------------ grounding regression loss
// One CTA.  stats = {mean |p0-g0|, mean |p1-g1|, mean IoU, n_valid, count(IoU > th_0), ...}; dpreds = gradient of
// loss = (1 - mean IoU) + L1 terms (center/duration: sum of both means; start/end: their average) w.r.t. preds.
constexpr int kMaxTh = 8;
struct GlArgs {
  const float* preds; const float* ga; const float* gb; const uint8_t* pad;
  long M; int use_cd; float eps; int nth; float th[kMaxTh];
  float* stats; float* iou; float* dpreds;
};

__global__ void __launch_bounds__(256) ground_loss_kernel(GlArgs p) {
  __shared__ float red[8][4 + kMaxTh];
  __shared__ float tot[4 + kMaxTh];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float acc[4 + kMaxTh];
#pragma unroll
  for (int i = 0; i < 4 + kMaxTh; ++i) acc[i] = 0.f;
  for (long m = tid; m < p.M; m += 256) {
    float iou = 0.f;
    if (!p.pad[m]) {
      const float a = p.preds[2 * m], b = p.preds[2 * m + 1], ga = p.ga[m], gb = p.gb[m];
      const float ps = p.use_cd ? a - b * 0.5f : a, pe = p.use_cd ? a + b * 0.5f : b;
      const float gs = p.use_cd ? ga - gb * 0.5f : ga, ge = p.use_cd ? ga + gb * 0.5f : gb;
      const float inter = fmaxf(fminf(pe, ge) - fmaxf(ps, gs), 0.f);
      const float uni = fmaxf(pe, ge) - fminf(ps, gs);
      iou = inter / (uni + p.eps);
      acc[0] += fabsf(a - ga); acc[1] += fabsf(b - gb); acc[2] += iou; acc[3] += 1.f;
      for (int k = 0; k < p.nth; ++k) acc[4 + k] += iou > p.th[k] ? 1.f : 0.f;
    }
    p.iou[m] = iou;
  }
#pragma unroll
  for (int i = 0; i < 4 + kMaxTh; ++i) {
    const float v = warp_sum(acc[i]);
    if (lane == 0) red[warp][i] = v;
  }
  __syncthreads();
  if (tid < 4 + kMaxTh) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w][tid];
    tot[tid] = t;
  }
  __syncthreads();
  const float n = tot[3], inv_n = n > 0.f ? 1.f / n : 0.f;
  if (tid < 4 + p.nth) p.stats[tid] = (tid < 3) ? tot[tid] * inv_n : tot[tid];
  const float wl1 = p.use_cd ? inv_n : 0.5f * inv_n;
  for (long m = tid; m < p.M; m += 256) {
    float da = 0.f, db = 0.f;
    if (!p.pad[m]) {
      const float a = p.preds[2 * m], b = p.preds[2 * m + 1], ga = p.ga[m], gb = p.gb[m];
      const float ps = p.use_cd ? a - b * 0.5f : a, pe = p.use_cd ? a + b * 0.5f : b;
      const float gs = p.use_cd ? ga - gb * 0.5f : ga, ge = p.use_cd ? ga + gb * 0.5f : gb;
      const float raw = fminf(pe, ge) - fmaxf(ps, gs);
      const float inter = fmaxf(raw, 0.f);
      const float u = fmaxf(pe, ge) - fminf(ps, gs) + p.eps;
      const float on = raw >= 0.f ? 1.f : 0.f;
      // d inter / d pe, d ps ; d union / d pe, d ps   (ties split like torch.min / torch.max)
      const float di_e = on * (pe < ge ? 1.f : (pe == ge ? 0.5f : 0.f));
      const float di_s = -on * (ps > gs ? 1.f : (ps == gs ? 0.5f : 0.f));
      const float du_e = pe > ge ? 1.f : (pe == ge ? 0.5f : 0.f);
      const float du_s = -(ps < gs ? 1.f : (ps == gs ? 0.5f : 0.f));
      const float k = -inv_n / (u * u);                       // d(1 - mean iou)
      const float d_e = k * (di_e * u - inter * du_e), d_s = k * (di_s * u - inter * du_s);
      const float sa = a > ga ? 1.f : (a < ga ? -1.f : 0.f), sb = b > gb ? 1.f : (b < gb ? -1.f : 0.f);
      if (p.use_cd) { da = d_s + d_e + wl1 * sa; db = 0.5f * (d_e - d_s) + wl1 * sb; }
      else          { da = d_s + wl1 * sa;       db = d_e + wl1 * sb; }
    }
    p.dpreds[2 * m] = da; p.dpreds[2 * m + 1] = db;
  }
}

}  // namespace

extern "C" int exg_xent_diag_fwd(const void* logits, int dtype, float* loss_rows, float* lse_rows, long rows,
                                 int cols, int period, void* stream) {
  EXG_CHECK_ARG(logits && loss_rows && lse_rows && exg_valid_dtype(dtype) && cols > 0 && period > 0 && period <= cols,
                "xent_diag_fwd: bad argument");
  if (rows <= 0) return 0;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * cols * exg_dtype_size(dtype), stream);
  xent_diag_fwd_kernel<<<(unsigned)((rows + kWarps - 1) / kWarps), kWarps * 32, 0, (cudaStream_t)stream>>>(
      logits, dtype, loss_rows, lse_rows, rows, cols, period);
  EXG_LAUNCH_CHECK("xent_diag_fwd");
  return 0;
}

extern "C" int exg_xent_diag_bwd(const void* logits, int dtype, const float* lse_rows, float scale, void* G,
                                 int g_dtype, long rows, int cols, int period, void* stream) {
  EXG_CHECK_ARG(logits && lse_rows && G && exg_valid_dtype(dtype) && exg_valid_dtype(g_dtype) && cols > 0 &&
                    period > 0, "xent_diag_bwd: bad argument");
  if (rows <= 0) return 0;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * cols * (exg_dtype_size(dtype) + exg_dtype_size(g_dtype)), stream);
  xent_diag_bwd_kernel<<<(unsigned)((rows + kWarps - 1) / kWarps), kWarps * 32, 0, (cudaStream_t)stream>>>(
      logits, dtype, lse_rows, scale, G, g_dtype, rows, cols, period);
  EXG_LAUNCH_CHECK("xent_diag_bwd");
  return 0;
}

static int xview_fill(XvArgs& a, const void* out_feat, int dtype, const float* ego, const long* pos_idx,
                      const long* neg_idx, const float* same_feat, const long* same_idx, int only_same, int B,
                      int V, int T, int D, float inv_temp, float* stats) {
  EXG_CHECK_ARG(out_feat && ego && pos_idx && neg_idx && stats && exg_valid_dtype(dtype), "xview_nce: null / dtype");
  EXG_CHECK_ARG((same_feat == nullptr) == (same_idx == nullptr), "xview_nce: same_feat and same_idx go together");
  EXG_CHECK_ARG(B > 0 && V > 0 && T > 0 && D > 0 && D % 4 == 0, "xview_nce: bad sizes (D must be a multiple of 4)");
  a = XvArgs{};
  a.o = out_feat; a.odt = dtype; a.ego = ego; a.pos_idx = pos_idx; a.neg_idx = neg_idx; a.same = same_feat;
  a.same_idx = same_idx; a.only_same = only_same; a.B = B; a.V = V; a.T = T; a.D = D; a.inv_temp = inv_temp;
  a.stats = stats;
  return 0;
}

extern "C" int exg_xview_nce_fwd(const void* out_feat, int dtype, const float* ego, const long* pos_idx,
                                 const long* neg_idx, const float* same_feat, const long* same_idx, int only_same,
                                 int B, int V, int T, int D, float inv_temp, float* loss, float* l1,
                                 float* pos_cos, float* neg_cos, float* stats, void* stream) {
  XvArgs a;
  int rc = xview_fill(a, out_feat, dtype, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, B, V, T, D,
                      inv_temp, stats);
  if (rc) return rc;
  EXG_CHECK_ARG(loss && l1 && pos_cos && neg_cos, "xview_nce_fwd: null output");
  a.loss = loss; a.l1 = l1; a.pos_cos = pos_cos; a.neg_cos = neg_cos;
  const long rows = (long)B * T;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * D * (2.0 * exg_dtype_size(dtype) + (same_feat ? 12.0 : 8.0) + 4.0),
                    stream);
  xview_fwd_kernel<<<(unsigned)((rows + kWarps - 1) / kWarps), kWarps * 32, 0, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("xview_nce_fwd");
  return 0;
}

extern "C" int exg_xview_nce_bwd(const void* out_feat, int dtype, const float* ego, const long* pos_idx,
                                 const long* neg_idx, const float* same_feat, const long* same_idx, int only_same,
                                 int B, int V, int T, int D, float inv_temp, const float* stats,
                                 const float* grad_loss, void* grad_out, int g_dtype, void* stream) {
  XvArgs a;
  int rc = xview_fill(a, out_feat, dtype, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, B, V, T, D,
                      inv_temp, const_cast<float*>(stats));
  if (rc) return rc;
  EXG_CHECK_ARG(grad_loss && grad_out && exg_valid_dtype(g_dtype), "xview_nce_bwd: bad argument");
  a.gl = grad_loss; a.go = grad_out; a.godt = g_dtype;
  const long rows = (long)B * T;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * D * (exg_dtype_size(dtype) + (same_feat ? 24.0 : 16.0) +
                                                        exg_dtype_size(g_dtype)), stream);
  xview_bwd_kernel<<<(unsigned)((rows + kWarps - 1) / kWarps), kWarps * 32, 0, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("xview_nce_bwd");
  return 0;
}

extern "C" int exg_ground_loss(const float* preds, const float* gt_a, const float* gt_b, const uint8_t* text_pad,
                               long M, int use_center_duration, float eps, const float* thresholds, int n_thresh,
                               float* stats, float* iou, float* dpreds, void* stream) {
  EXG_CHECK_ARG(preds && gt_a && gt_b && text_pad && stats && iou && dpreds, "ground_loss: null pointer");
  EXG_CHECK_ARG(M > 0 && n_thresh >= 0 && n_thresh <= kMaxTh && (n_thresh == 0 || thresholds),
                "ground_loss: bad sizes (at most %d IoU thresholds)", kMaxTh);
  GlArgs a = {};
  a.preds = preds; a.ga = gt_a; a.gb = gt_b; a.pad = text_pad; a.M = M; a.use_cd = use_center_duration;
  a.eps = eps; a.nth = n_thresh;
  for (int i = 0; i < n_thresh; ++i) a.th[i] = thresholds[i];      // HOST array
  a.stats = stats; a.iou = iou; a.dpreds = dpreds;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)M * 29.0, stream);
  ground_loss_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("ground_loss");
  return 0;
}
