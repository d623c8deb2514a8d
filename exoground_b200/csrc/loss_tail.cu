// loss_tail.cu — the scalar tail of the TAN loss (reference train/loss.py:251-360) in two launches.
//
// After exg_nce_fwd has produced the row / column exp-sums of both logit tensors, the reference computes
//   * the masked means of v = LSE_all - LSE_pos over rows / columns with a positive          (loss.py:251-278)
//   * the per-sentence max-over-time logit, z-scored over the valid sentences, its `loss_threshold`
//     quantile, the kept sentences / rows and the re-weighted means                           (loss.py:280-307)
//   * the median split of the dual / joint maxima into {1, 0, ignore} alignability labels, the optional trim by
//     absolute text position, BCE-with-logits with pos_weight = 1/mean(y) - 1 on the stage-3 alignability logits
//     and its top-1 accuracy                                                                  (loss.py:309-360)
// as ~300 tiny elementwise / reduction / sort launches.  Here:
//   exg_nce_logdiff : grid-wide, vr[r] = sum_s (log row_all - log row_pos), tc[c] = sum_s (log col_all - log col_pos)
//   exg_tan_loss_tail: ONE CTA (1024 threads): row / column masks from the packed target bit rows, counts, weighted
//                      sums, three quantiles by 8-bit radix selection (shared-memory histograms), BCE value +
//                      gradient, and the row / column weights exg_nce_bwd needs.
// Latency-bound: (2R + 2Cc) * S * 8 B read by the first, ~(2 R W + 8 Cc + 4 R) * 4 B by the second.
#include <mutex>

#include "common.cuh"

namespace {

constexpr float kTiny = 1e-38f;

__global__ void nce_logdiff_kernel(const float* __restrict__ rows_d, const float* __restrict__ cols_d, int E,
                                   const float* __restrict__ rows_j, const float* __restrict__ cols_j, int D,
                                   int R, int Cc, float* __restrict__ vr_d, float* __restrict__ tc_d,
                                   float* __restrict__ vr_j, float* __restrict__ tc_j) {
  const int total = 2 * (R + Cc);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const bool joint = i >= R + Cc;
    const int k = joint ? i - (R + Cc) : i;
    const bool is_col = k >= R;
    const int e = is_col ? k - R : k;
    const int n = is_col ? Cc : R;
    const int S = joint ? D : E;
    const float* base = joint ? (is_col ? cols_j : rows_j) : (is_col ? cols_d : rows_d);
    float acc = 0.f;
    for (int s = 0; s < S; ++s) {
      float all = base[(size_t)s * n + e];
      float pos = base[(size_t)(S + s) * n + e];
      if (is_col) all = fmaxf(all, kTiny);
      acc += logf(all) - logf(fmaxf(pos, kTiny));
    }
    float* out = joint ? (is_col ? tc_j : vr_j) : (is_col ? tc_d : vr_d);
    out[e] = acc;
  }
}

constexpr int NT = 1024;

__device__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = threadIdx.x < NT / 32 ? red[threadIdx.x] : 0.f;
  if (threadIdx.x < 32) {
    t = warp_sum(t);
    if (threadIdx.x == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}

__device__ float block_min(float v, float* red) {
  v = -warp_max(-v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = threadIdx.x < NT / 32 ? red[threadIdx.x] : INFINITY;
  if (threadIdx.x < 32) {
    t = -warp_max(-t);
    if (threadIdx.x == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}

__device__ __forceinline__ uint32_t sort_key(float x) {      // monotone float -> uint32 (ascending)
  const uint32_t u = __float_as_uint(x);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_value(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// k-th smallest (0-based) of {x[i] : mask[i]} by 4 passes of 8-bit radix selection: shared-memory histogram of the
// candidates that still match the selected prefix, warp-parallel scan for the bin holding rank k.
__device__ float radix_select(const float* x, const uint8_t* mask, int Cc, int k, uint32_t* hist, float* red) {
  uint32_t prefix = 0, pmask = 0;
  __shared__ uint32_t sel[2];
  for (int shift = 24; shift >= 0; shift -= 8) {
    __syncthreads();
    if (threadIdx.x < 256) hist[threadIdx.x] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < Cc; i += NT) {
      if (!mask[i]) continue;
      const uint32_t key = sort_key(x[i]);
      if ((key & pmask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t h[8], tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { h[j] = hist[threadIdx.x * 8 + j]; tot += h[j]; }
      uint32_t incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)threadIdx.x >= o) incl += t;
      }
      uint32_t before = incl - tot;                 // candidates in lower bins
      if ((uint32_t)k >= before && (uint32_t)k < incl) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if ((uint32_t)k < before + h[j]) { sel[0] = threadIdx.x * 8 + j; sel[1] = before; break; }
          before += h[j];
        }
      }
    }
    __syncthreads();
    prefix |= sel[0] << shift;
    pmask |= 255u << shift;
    k -= (int)sel[1];
  }
  (void)red;
  return key_value(prefix);
}

// torch.quantile(x[mask], q) with linear interpolation: pos = (n-1)*q, lo = floor(pos), hi = min(lo+1, n-1),
// v_lo + (v_hi - v_lo)*(pos - lo) over the sorted valid values (train/loss.py:289, :315-316).  v_lo by radix
// selection; v_hi = v_lo if it has a duplicate at rank lo+1, else the smallest value above v_lo.
__device__ float masked_quantile(const float* x, const uint8_t* mask, int Cc, float n, float q, uint32_t* hist,
                                 float* red) {
  const float pos = (n - 1.f) * q;
  int lo = (int)floorf(pos);
  lo = lo < 0 ? 0 : lo;
  int last = (int)n - 1;
  last = last < 0 ? 0 : last;
  lo = min(lo, last);
  const int hi = min(lo + 1, last);
  const float w = pos - (float)lo;
  const float a = radix_select(x, mask, Cc, lo, hist, red);
  float b = a;
  if (hi > lo) {
    float cnt = 0.f, mn = INFINITY;
    for (int i = threadIdx.x; i < Cc; i += NT) {
      if (!mask[i]) continue;
      const float v = x[i];
      if (v <= a) cnt += 1.f; else mn = fminf(mn, v);
    }
    cnt = block_sum(cnt, red);
    mn = block_min(mn, red);
    b = (cnt >= (float)(lo + 2)) ? a : mn;
  }
  return a + (b - a) * w;
}

struct TailArgs {
  const uint32_t* bits;    // exg_nce_pack_targets([B,N,T] targets AND-ed with the valid sentences): [B*T][W]
  const uint8_t* valid;    // [B*N]
  const float *vr_d, *tc_d, *vr_j, *tc_j;
  const float *md, *mj;    // [B*N] raw per-sentence maxima (scaled by inv_temp here)
  const float* la;         // [B*N] alignability logits or null
  const float* abs_c;      // [B*N] mean absolute text position or null
  float* ws;               // [3*Cc] scratch: md*, mj*, metric
  uint8_t* flags;          // [R + 2*Cc] scratch: rowp / rowp_th, colp, keep
  float *wr_d, *wc_d, *wr_j, *wc_j, *dla;   // outputs consumed by the backward
  float* out;              // [16] scalars
  int B, T, N, E, D, W;
  float inv_temp, q;       // q <= 0: no threshold
  int use_head;
  float nce_weight;
};

__global__ void __launch_bounds__(NT, 1) tan_loss_tail_kernel(TailArgs a) {
  extern __shared__ uint32_t cbits[];     // [2][B*W]: per-clip OR of the target bit rows | kept-sentence bits
  __shared__ float red[33];
  __shared__ uint32_t hist[256];
  const int B = a.B, T = a.T, N = a.N, W = a.W, R = B * T, Cc = B * N;
  const int tid = threadIdx.x;
  uint8_t* rowp = a.flags;
  uint8_t* colp = a.flags + R;
  uint8_t* keep = a.flags + R + Cc;
  uint32_t* kbits = cbits + B * W;
  const bool use_th = a.q > 0.f;

  // ---- rows / columns with at least one positive (loss.py:247-250), from the packed target rows
  for (int i = tid; i < 2 * B * W; i += NT) cbits[i] = 0;
  __syncthreads();
  for (int r = tid; r < R; r += NT) {
    uint32_t any = 0;
    for (int w = 0; w < W; ++w) {
      const uint32_t v = a.bits[(size_t)r * W + w];
      any |= v;
      if (v) atomicOr(&cbits[(r / T) * W + w], v);
    }
    rowp[r] = any != 0;
  }
  __syncthreads();
  for (int c = tid; c < Cc; c += NT) {
    const int b = c / N, n = c % N;
    colp[c] = (cbits[b * W + (n >> 5)] >> (n & 31)) & 1u;
  }
  __syncthreads();
  float nr = 0.f, nc = 0.f, nv = 0.f;
  for (int r = tid; r < R; r += NT) nr += rowp[r];
  for (int c = tid; c < Cc; c += NT) { nc += colp[c]; nv += a.valid[c]; }
  nr = block_sum(nr, red); nc = block_sum(nc, red); nv = block_sum(nv, red);

  // ---- unthresholded masked means (loss.py:251-258): sum over stages already folded by exg_nce_logdiff
  float sd = 0.f, sj = 0.f, cd = 0.f, cj = 0.f;
  for (int r = tid; r < R; r += NT) if (rowp[r]) { sd += a.vr_d[r]; sj += a.vr_j[r]; }
  for (int c = tid; c < Cc; c += NT) if (colp[c]) { cd += a.tc_d[c]; cj += a.tc_j[c]; }
  sd = block_sum(sd, red); sj = block_sum(sj, red); cd = block_sum(cd, red); cj = block_sum(cj, red);
  const float loss_d_all = sd / (2.f * a.E * nr) + cd / (2.f * a.E * nc);
  const float loss_j_all = sj / (2.f * a.D * nr) + cj / (2.f * a.D * nc);
  float loss_d = loss_d_all, loss_j = loss_j_all;
  float nr_w = nr, nc_w = nc;              // the counts behind the weights the backward will use
  const uint8_t* colw = colp;

  float* md = a.ws;
  float* mj = a.ws + Cc;
  float* metric = a.ws + 2 * Cc;
  for (int c = tid; c < Cc; c += NT) { md[c] = a.md[c] * a.inv_temp; mj[c] = a.mj[c] * a.inv_temp; }
  __syncthreads();

  if (use_th) {
    // ---- z-score of the per-sentence maxima over the valid sentences, quantile, kept sentences (loss.py:283-298)
    float s1 = 0.f, s2 = 0.f;
    for (int c = tid; c < Cc; c += NT) if (a.valid[c]) { s1 += md[c]; s2 += mj[c]; }
    const float mean_d = block_sum(s1, red) / nv, mean_j = block_sum(s2, red) / nv;
    s1 = s2 = 0.f;
    for (int c = tid; c < Cc; c += NT)
      if (a.valid[c]) { const float u = md[c] - mean_d, v = mj[c] - mean_j; s1 += u * u; s2 += v * v; }
    const float sd_d = sqrtf(block_sum(s1, red) / (nv - 1.f)), sd_j = sqrtf(block_sum(s2, red) / (nv - 1.f));
    for (int c = tid; c < Cc; c += NT) metric[c] = -((md[c] - mean_d) / sd_d + (mj[c] - mean_j) / sd_j);
    __syncthreads();
    const float qv = masked_quantile(metric, a.valid, Cc, nv, a.q, hist, red);
    __syncthreads();
    for (int c = tid; c < Cc; c += NT) {
      const bool k = a.valid[c] && (metric[c] <= qv);
      keep[c] = k;
      if (k) atomicOr(&kbits[(c / N) * W + ((c % N) >> 5)], 1u << ((c % N) & 31));
    }
    __syncthreads();
    for (int r = tid; r < R; r += NT) {
      uint32_t any = 0;
      for (int w = 0; w < W; ++w) any |= a.bits[(size_t)r * W + w] & kbits[(r / T) * W + w];
      rowp[r] = any != 0;                   // rowp now holds the thresholded row mask
    }
    __syncthreads();
    float nrt = 0.f, nk = 0.f;
    for (int r = tid; r < R; r += NT) nrt += rowp[r];
    for (int c = tid; c < Cc; c += NT) nk += keep[c];
    nrt = block_sum(nrt, red); nk = block_sum(nk, red);
    sd = sj = cd = cj = 0.f;
    for (int r = tid; r < R; r += NT) if (rowp[r]) { sd += a.vr_d[r]; sj += a.vr_j[r]; }
    for (int c = tid; c < Cc; c += NT) if (keep[c]) { cd += a.tc_d[c]; cj += a.tc_j[c]; }
    sd = block_sum(sd, red); sj = block_sum(sj, red); cd = block_sum(cd, red); cj = block_sum(cj, red);
    loss_d = sd / (2.f * a.E * nrt) + cd / (2.f * a.E * nk);
    loss_j = sj / (2.f * a.D * nrt) + cj / (2.f * a.D * nk);
    nr_w = nrt; nc_w = nk; colw = keep;
  }

  // ---- alignability labels + BCE (loss.py:309-360)
  float bce = 0.f, top1 = 0.f;
  if (a.use_head) {
    __syncthreads();
    const float qd = masked_quantile(md, a.valid, Cc, nv, 0.5f, hist, red);
    const float qj = masked_quantile(mj, a.valid, Cc, nv, 0.5f, hist, red);
    __syncthreads();
    float nsel = 0.f, ysum = 0.f;
    for (int c = tid; c < Cc; c += NT) {
      float lab = 2.f;
      if (md[c] > qd && mj[c] > qj) lab = 1.f;
      if (md[c] < qd && mj[c] < qj) lab = 0.f;
      if (a.abs_c != nullptr && (a.abs_c[c] < 0.2f || a.abs_c[c] > 0.8f)) lab = 0.f;
      const bool sel = a.valid[c] && lab != 2.f;
      metric[c] = sel ? lab : -1.f;         // metric now carries y (or -1 = not selected)
      nsel += sel; ysum += sel ? lab : 0.f;
    }
    nsel = block_sum(nsel, red); ysum = block_sum(ysum, red);
    const float pw = nsel / ysum - 1.f;
    float lsum = 0.f, hit = 0.f;
    for (int c = tid; c < Cc; c += NT) {
      const float y = metric[c];
      float g = 0.f;
      if (y >= 0.f) {
        const float x = a.la[c];
        const float lw = 1.f + (pw - 1.f) * y;
        lsum += (1.f - y) * x + lw * (log1pf(expf(-fabsf(x))) + fmaxf(-x, 0.f));
        hit += ((x > 0.f ? 1.f : 0.f) == y) ? 1.f : 0.f;
        const float sneg = 1.f / (1.f + expf(x));           // sigmoid(-x)
        g = ((1.f - y) - lw * sneg) / nsel;
      }
      a.dla[c] = g;
    }
    bce = block_sum(lsum, red) / nsel;
    top1 = block_sum(hit, red) / nsel;
  }

  // ---- weights of the backward: d loss / d (row / column log-ratio); the /2 of (dual + joint)/2 and nce_weight
  const float scale = 0.5f * (a.use_head ? a.nce_weight : 1.f);
  const float wrd = scale / (2.f * a.E * nr_w), wcd = scale / (2.f * a.E * nc_w);
  const float wrj = scale / (2.f * a.D * nr_w), wcj = scale / (2.f * a.D * nc_w);
  for (int r = tid; r < R; r += NT) { const float m = rowp[r]; a.wr_d[r] = m * wrd; a.wr_j[r] = m * wrj; }
  for (int c = tid; c < Cc; c += NT) { const float m = colw[c]; a.wc_d[c] = m * wcd; a.wc_j[c] = m * wcj; }
  if (tid == 0) {
    float loss = (loss_d + loss_j) * 0.5f;
    if (a.use_head) loss = loss * a.nce_weight + bce;
    a.out[0] = loss;
    a.out[1] = loss_d;  a.out[2] = loss_j;            // 'loss-dual' / 'loss-joint' (thresholded when q > 0)
    a.out[3] = loss_d_all; a.out[4] = loss_j_all;     // 'loss-dual-all' / 'loss-joint-all'
    a.out[5] = (loss_d_all + loss_j_all) * 0.5f;      // 'loss-total'
    a.out[6] = bce; a.out[7] = top1;
    a.out[8] = nr_w; a.out[9] = nc_w; a.out[10] = nv;
  }
}

}  // namespace

extern "C" int exg_nce_logdiff(const float* rows_d, const float* cols_d, int E, const float* rows_j,
                               const float* cols_j, int D, int R, int Cc, float* vr_d, float* tc_d, float* vr_j,
                               float* tc_j, void* stream) {
  EXG_CHECK_ARG(rows_d && cols_d && rows_j && cols_j && vr_d && tc_d && vr_j && tc_j, "nce_logdiff: null pointer");
  EXG_CHECK_ARG(E > 0 && D > 0 && R > 0 && Cc > 0, "nce_logdiff: bad sizes");
  ExgProfScope prof(EXG_PC_ROWWISE, (double)(R + Cc) * (2.0 * (E + D) + 2.0) * 4.0, stream);
  const int total = 2 * (R + Cc);
  const int grid = min((total + 255) / 256, 4 * exg_num_sms());
  nce_logdiff_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(rows_d, cols_d, E, rows_j, cols_j, D, R, Cc, vr_d, tc_d,
                                                             vr_j, tc_j);
  EXG_LAUNCH_CHECK("nce_logdiff");
  return 0;
}

extern "C" size_t exg_tan_loss_tail_workspace(int B, int T, int N) {
  const size_t R = (size_t)B * T, Cc = (size_t)B * N;
  return 3 * Cc * sizeof(float) + (R + 2 * Cc + 15) / 16 * 16;
}

extern "C" int exg_tan_loss_tail(const uint32_t* tgt_bits, const uint8_t* valid, const float* vr_d, const float* tc_d,
                                 const float* vr_j, const float* tc_j, const float* own_max_d, const float* own_max_j,
                                 const float* la, const float* abs_c, int B, int T, int N, int E, int D,
                                 float inv_temp, float loss_threshold, int use_head, float nce_weight, float* wr_d,
                                 float* wc_d, float* wr_j, float* wc_j, float* dla, float* out, void* ws,
                                 size_t ws_bytes, void* stream) {
  EXG_CHECK_ARG(tgt_bits && valid && vr_d && tc_d && vr_j && tc_j && own_max_d && own_max_j && wr_d && wc_d && wr_j &&
                wc_j && out && ws, "tan_loss_tail: null pointer");
  EXG_CHECK_ARG(B > 0 && T > 0 && N > 0 && E > 0 && D > 0, "tan_loss_tail: bad sizes");
  EXG_CHECK_ARG(!use_head || (la && dla), "tan_loss_tail: the alignability head needs la and dla");
  EXG_CHECK_ARG(ws_bytes >= exg_tan_loss_tail_workspace(B, T, N), "tan_loss_tail: workspace too small");
  const long Cc = (long)B * N;
  const int W = (N + 31) / 32;
  const size_t smem = (size_t)2 * B * W * sizeof(uint32_t);
  EXG_CHECK_ARG(smem <= 160 * 1024, "tan_loss_tail: B*ceil(N/32) = %d too large for the per-clip bit rows", B * W);
  TailArgs a = {};
  a.bits = tgt_bits; a.valid = valid; a.vr_d = vr_d; a.tc_d = tc_d; a.vr_j = vr_j; a.tc_j = tc_j;
  a.md = own_max_d; a.mj = own_max_j; a.la = la; a.abs_c = abs_c;
  a.ws = (float*)ws; a.flags = (uint8_t*)ws + 3 * Cc * sizeof(float);
  a.wr_d = wr_d; a.wc_d = wc_d; a.wr_j = wr_j; a.wc_j = wc_j; a.dla = dla; a.out = out;
  a.B = B; a.T = T; a.N = N; a.E = E; a.D = D; a.W = W;
  a.inv_temp = inv_temp; a.q = loss_threshold; a.use_head = use_head; a.nce_weight = nce_weight;
  if (smem > 40 * 1024)
    EXG_CUDA(cudaFuncSetAttribute(tan_loss_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ExgProfScope prof(EXG_PC_ROWWISE, (double)B * T * W * 8.0 + (double)(B * T + Cc) * 8.0 * 4.0, stream);
  tan_loss_tail_kernel<<<1, NT, smem, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("tan_loss_tail");
  return 0;
}

// ---------------------------------------------------------------- allocation-free variant
// exg_tan_loss_tail_static: the same kernel with its scratch (`ws`: three [B*N] fp32 columns + the row / column /
// kept flags) in a file-scope device array instead of a caller-allocated buffer, for callers that cannot allocate
// (code under CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the
// library holds kTailStaticBytes until exit.  The kernel writes every scratch element before reading it, so nothing
// has to be cleared between calls; all calls on one device share the array, so two calls that overlap on the GPU
// (different streams, no ordering between them) race.
namespace {
constexpr size_t kTailStaticBytes = 1u << 20;             // 1 MiB: e.g. B*T = 128 Ki rows with B*N = 64 Ki sentences
static __device__ __align__(256) unsigned char g_tail_scratch[kTailStaticBytes];
constexpr int kMaxDevices = 64;
void* g_tail_scratch_addr[kMaxDevices] = {};              // per device: the DEVICE address of g_tail_scratch
std::mutex g_tail_scratch_mu;

// In host code `g_tail_scratch` is only a handle for the symbol, not a device address: resolve it once per device,
// with that device current.
int tail_scratch_ptr(void** out) {
  int dev = -1;
  EXG_CUDA(cudaGetDevice(&dev));
  EXG_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "tan_loss_tail_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_tail_scratch_mu);
  if (g_tail_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    EXG_CUDA(cudaGetSymbolAddress(&p, g_tail_scratch));
    EXG_CHECK_ARG(p != nullptr, "tan_loss_tail_static: cudaGetSymbolAddress returned null");
    g_tail_scratch_addr[dev] = p;
  }
  *out = g_tail_scratch_addr[dev];
  return 0;
}
}  // namespace

extern "C" size_t exg_tan_loss_tail_static_max_bytes(void) { return kTailStaticBytes; }

extern "C" int exg_tan_loss_tail_static(const uint32_t* tgt_bits, const uint8_t* valid, const float* vr_d,
                                        const float* tc_d, const float* vr_j, const float* tc_j,
                                        const float* own_max_d, const float* own_max_j, const float* la,
                                        const float* abs_c, int B, int T, int N, int E, int D, float inv_temp,
                                        float loss_threshold, int use_head, float nce_weight, float* wr_d,
                                        float* wc_d, float* wr_j, float* wc_j, float* dla, float* out, void* stream) {
  EXG_CHECK_ARG(tgt_bits && valid && vr_d && tc_d && vr_j && tc_j && own_max_d && own_max_j && wr_d && wc_d && wr_j &&
                wc_j && out, "tan_loss_tail_static: null pointer");
  EXG_CHECK_ARG(B > 0 && T > 0 && N > 0 && E > 0 && D > 0, "tan_loss_tail_static: bad sizes");
  EXG_CHECK_ARG(!use_head || (la && dla), "tan_loss_tail_static: the alignability head needs la and dla");
  const size_t need = exg_tan_loss_tail_workspace(B, T, N);
  EXG_CHECK_ARG(need <= kTailStaticBytes,
                "tan_loss_tail_static: B=%d T=%d N=%d needs %zu scratch bytes, the built-in scratch holds %zu "
                "(use exg_tan_loss_tail with a caller-allocated workspace)", B, T, N, need, kTailStaticBytes);
  const long Cc = (long)B * N;
  const int W = (N + 31) / 32;
  const size_t smem = (size_t)2 * B * W * sizeof(uint32_t);
  EXG_CHECK_ARG(smem <= 160 * 1024, "tan_loss_tail_static: B*ceil(N/32) = %d too large for the per-clip bit rows",
                B * W);
  void* ws = nullptr;
  if (int rc = tail_scratch_ptr(&ws)) return rc;
  TailArgs a = {};
  a.bits = tgt_bits; a.valid = valid; a.vr_d = vr_d; a.tc_d = tc_d; a.vr_j = vr_j; a.tc_j = tc_j;
  a.md = own_max_d; a.mj = own_max_j; a.la = la; a.abs_c = abs_c;
  a.ws = (float*)ws; a.flags = (uint8_t*)ws + 3 * Cc * sizeof(float);
  a.wr_d = wr_d; a.wc_d = wc_d; a.wr_j = wr_j; a.wc_j = wc_j; a.dla = dla; a.out = out;
  a.B = B; a.T = T; a.N = N; a.E = E; a.D = D; a.W = W;
  a.inv_temp = inv_temp; a.q = loss_threshold; a.use_head = use_head; a.nce_weight = nce_weight;
  if (smem > 40 * 1024)
    EXG_CUDA(cudaFuncSetAttribute(tan_loss_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tan_loss_tail_kernel<<<1, NT, smem, (cudaStream_t)stream>>>(a);     // no profiling scope: its events allocate
  EXG_LAUNCH_CHECK("tan_loss_tail_static");
  return 0;
}
