// nce.cu — masked bidirectional MIL-NCE over materialised stage-major logits L[S, R=B*T, Cc=B*N]
// (train/loss.py:234-278): one streaming pass for the four exp-sums, one streaming pass for dL.
// HBM-bound: forward reads L once; backward reads L once and writes G once.
// Because L is a cosine, |L*inv_temp| <= inv_temp, so exp(L*inv_temp - inv_temp) in (0, 1] needs no
// running max; the shift cancels in every log-ratio and in every softmax quotient.
#include "common.cuh"

namespace {

constexpr int SLAB = 32;       // rows per CTA
constexpr int NT = 256;

struct NceArgs {
  const void* L; int dtype;
  const uint8_t* tgt; const uint8_t* col_valid;
  float *row_all, *row_pos, *col_all, *col_pos;
  const float *wr, *wc;
  void* G; int g_dtype;
  int S, B, T, N; float inv_temp;
};

__device__ __forceinline__ float comp(const float4& v, int j) {
  return j == 0 ? v.x : (j == 1 ? v.y : (j == 2 ? v.z : v.w));
}

__global__ void __launch_bounds__(NT) nce_fwd_kernel(NceArgs p) {
  __shared__ float red[NT / 32][2 * SLAB];
  const int s = blockIdx.y;
  const long R = (long)p.B * p.T, Cc = (long)p.B * p.N;
  const long r0 = (long)blockIdx.x * SLAB;
  const int nrows = (int)min((long)SLAB, R - r0);
  const size_t base = (size_t)s * R * Cc;
  float ra[SLAB], rp[SLAB];
#pragma unroll
  for (int i = 0; i < SLAB; ++i) ra[i] = rp[i] = 0.f;

  for (long c4 = (long)threadIdx.x * 4; c4 < Cc; c4 += NT * 4) {
    int cb[4]; bool cv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      cb[j] = (int)((c4 + j) / p.N);
      cv[j] = p.col_valid[c4 + j] != 0;
    }
    float ca[4] = {0.f, 0.f, 0.f, 0.f}, cp[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < SLAB; ++i) {      // fully unrolled: ra/rp must stay in registers
      if (i >= nrows) continue;
      const long r = r0 + i;
      const int a = (int)(r / p.T), t = (int)(r % p.T);
      const float4 v = load4(p.L, p.dtype, base + (size_t)r * Cc + c4);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (!cv[j]) continue;
        const float e = __expf(comp(v, j) * p.inv_temp - p.inv_temp);
        ra[i] += e; ca[j] += e;
        if (cb[j] == a && p.tgt[((size_t)(c4 + j)) * p.T + t]) { rp[i] += e; cp[j] += e; }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (!cv[j]) continue;
      atomicAdd(p.col_all + (size_t)s * Cc + c4 + j, ca[j]);
      if (cp[j] != 0.f) atomicAdd(p.col_pos + (size_t)s * Cc + c4 + j, cp[j]);
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < SLAB; ++i) {
    const float x = warp_sum(ra[i]), y = warp_sum(rp[i]);
    if (lane == 0) { red[warp][i] = x; red[warp][SLAB + i] = y; }
  }
  __syncthreads();
  if (threadIdx.x < 2 * SLAB) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < NT / 32; ++w) t += red[w][threadIdx.x];
    const int i = threadIdx.x % SLAB;
    if (i < nrows) {
      float* dst = threadIdx.x < SLAB ? p.row_all : p.row_pos;
      dst[(size_t)s * R + r0 + i] = t;
    }
  }
}

__global__ void __launch_bounds__(NT) nce_bwd_kernel(NceArgs p) {
  __shared__ float ra_s[SLAB], rp_s[SLAB];
  const int s = blockIdx.y;
  const long R = (long)p.B * p.T, Cc = (long)p.B * p.N;
  const long r0 = (long)blockIdx.x * SLAB;
  const int nrows = (int)min((long)SLAB, R - r0);
  const size_t base = (size_t)s * R * Cc;
  if (threadIdx.x < SLAB) {
    float a = 0.f, b = 0.f;
    if (threadIdx.x < nrows) {
      const long r = r0 + threadIdx.x;
      const float w = p.wr[r];
      if (w != 0.f) {
        a = w / p.row_all[(size_t)s * R + r];
        const float rp = p.row_pos[(size_t)s * R + r];
        b = rp > 0.f ? w / rp : 0.f;
      }
    }
    ra_s[threadIdx.x] = a; rp_s[threadIdx.x] = b;
  }
  __syncthreads();
  for (long c4 = (long)threadIdx.x * 4; c4 < Cc; c4 += NT * 4) {
    int cb[4]; bool cv[4]; float ka[4], kp[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      cb[j] = (int)((c4 + j) / p.N);
      cv[j] = p.col_valid[c4 + j] != 0;
      const float w = p.wc[c4 + j];
      ka[j] = kp[j] = 0.f;
      if (cv[j] && w != 0.f) {
        ka[j] = w / p.col_all[(size_t)s * Cc + c4 + j];
        const float cp = p.col_pos[(size_t)s * Cc + c4 + j];
        kp[j] = cp > 0.f ? w / cp : 0.f;
      }
    }
#pragma unroll 4
    for (int i = 0; i < SLAB; ++i) {
      if (i >= nrows) break;
      const long r = r0 + i;
      const int a = (int)(r / p.T), t = (int)(r % p.T);
      const float4 v = load4(p.L, p.dtype, base + (size_t)r * Cc + c4);
      float g[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        g[j] = 0.f;
        if (!cv[j]) continue;
        const float e = __expf(comp(v, j) * p.inv_temp - p.inv_temp);
        float coef = ra_s[i] + ka[j];
        if (cb[j] == a && p.tgt[((size_t)(c4 + j)) * p.T + t]) coef -= rp_s[i] + kp[j];
        g[j] = p.inv_temp * e * coef;
      }
      store4(p.G, p.g_dtype, base + (size_t)r * Cc + c4, make_float4(g[0], g[1], g[2], g[3]));
    }
  }
}

__global__ void own_max_kernel(const void* __restrict__ L, int dtype, float* __restrict__ out, int stage,
                               int B, int T, int N) {
  const int b = blockIdx.x;
  const long R = (long)B * T, Cc = (long)B * N;
  for (int k = threadIdx.x; k < N; k += blockDim.x) {
    float m = -INFINITY;
    for (int t = 0; t < T; ++t)
      m = fmaxf(m, load1(L, dtype, (size_t)stage * R * Cc + ((size_t)b * T + t) * Cc + (size_t)b * N + k));
    out[(size_t)b * N + k] = m;
  }
}

int nce_check(const void* logits, int dtype, const uint8_t* tgt, const uint8_t* col_valid, int S, int B, int T,
              int N) {
  EXG_CHECK_ARG(logits && tgt && col_valid, "nce: null pointer");
  EXG_CHECK_ARG(exg_valid_dtype(dtype), "nce: bad dtype");
  EXG_CHECK_ARG(S > 0 && B > 0 && T > 0 && N > 0 && S <= 65535, "nce: bad sizes");
  EXG_CHECK_ARG(((long)B * N) % 4 == 0, "nce: B*N must be a multiple of 4");
  return 0;
}

}  // namespace

extern "C" int exg_nce_fwd(const void* logits, int dtype, const uint8_t* tgt, const uint8_t* col_valid,
                           float* row_all, float* row_pos, float* col_all, float* col_pos, int S, int B, int T,
                           int N, float inv_temp, void* stream) {
  int rc = nce_check(logits, dtype, tgt, col_valid, S, B, T, N);
  if (rc) return rc;
  EXG_CHECK_ARG(row_all && row_pos && col_all && col_pos, "nce_fwd: null output");
  NceArgs a = {};
  a.L = logits; a.dtype = dtype; a.tgt = tgt; a.col_valid = col_valid; a.row_all = row_all; a.row_pos = row_pos;
  a.col_all = col_all; a.col_pos = col_pos; a.S = S; a.B = B; a.T = T; a.N = N; a.inv_temp = inv_temp;
  const long R = (long)B * T;
  dim3 grid((unsigned)((R + SLAB - 1) / SLAB), (unsigned)S);
  ExgProfScope prof(EXG_PC_NCE_FWD, (double)S * R * B * N * exg_dtype_size(dtype), stream);
  nce_fwd_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("nce_fwd");
  return 0;
}

extern "C" int exg_nce_bwd(const void* logits, int dtype, const uint8_t* tgt, const uint8_t* col_valid,
                           const float* row_all, const float* row_pos, const float* col_all,
                           const float* col_pos, const float* wr, const float* wc, void* G, int g_dtype, int S,
                           int B, int T, int N, float inv_temp, void* stream) {
  int rc = nce_check(logits, dtype, tgt, col_valid, S, B, T, N);
  if (rc) return rc;
  EXG_CHECK_ARG(row_all && row_pos && col_all && col_pos && wr && wc && G && exg_valid_dtype(g_dtype),
                "nce_bwd: bad argument");
  NceArgs a = {};
  a.L = logits; a.dtype = dtype; a.tgt = tgt; a.col_valid = col_valid;
  a.row_all = const_cast<float*>(row_all); a.row_pos = const_cast<float*>(row_pos);
  a.col_all = const_cast<float*>(col_all); a.col_pos = const_cast<float*>(col_pos);
  a.wr = wr; a.wc = wc; a.G = G; a.g_dtype = g_dtype; a.S = S; a.B = B; a.T = T; a.N = N; a.inv_temp = inv_temp;
  const long R = (long)B * T;
  dim3 grid((unsigned)((R + SLAB - 1) / SLAB), (unsigned)S);
  ExgProfScope prof(EXG_PC_NCE_BWD, (double)S * R * B * N * (exg_dtype_size(dtype) + exg_dtype_size(g_dtype)), stream);
  nce_bwd_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("nce_bwd");
  return 0;
}

extern "C" int exg_own_max(const void* logits, int dtype, float* own_max, int stage, int B, int T, int N,
                           void* stream) {
  EXG_CHECK_ARG(logits && own_max && exg_valid_dtype(dtype) && B > 0 && T > 0 && N > 0 && stage >= 0,
                "own_max: bad argument");
  ExgProfScope prof(EXG_PC_ROWWISE, (double)B * T * N * exg_dtype_size(dtype), stream);
  own_max_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(logits, dtype, own_max, stage, B, T, N);
  EXG_LAUNCH_CHECK("own_max");
  return 0;
}
