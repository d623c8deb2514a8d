// nce.cu — masked bidirectional MIL-NCE over materialised stage-major logits L[S, R=B*T, Cc=B*N]
// (train/loss.py:234-278): one streaming pass for the four exp-sums, one streaming pass for dL; 16-byte
// loads / stores, column partial sums in registers (one atomic per column per 256-row CTA).
// HBM-bound: forward reads L once; backward reads L once and writes G once.
// Because L is a cosine, |L*inv_temp| <= inv_temp, so exp(L*inv_temp - inv_temp) in (0, 1] needs no
// running max; the shift cancels in every log-ratio and in every softmax quotient.
#include "common.cuh"

namespace {

constexpr int NT = 256;        // threads per CTA; a thread owns 8 consecutive columns
constexpr int CPT = 8;
constexpr int CB = NT * CPT;   // 2048 columns per CTA
constexpr int SUB = 16;        // rows per register sub-slab
constexpr int FWD_ROWS = 128;  // rows per CTA, forward (column partial sums stay in registers across them)
constexpr int BWD_ROWS = 64;  // (multiple of LDB)
constexpr int LDB = 8;        // rows whose loads are issued back to back before any is consumed (memory-level parallelism)
  // rows per CTA, backward

struct NceArgs {
  const void* L; int dtype;
  const uint8_t* tgt; const uint8_t* col_valid;
  const uint8_t* own_row_mask;   // optional [R]: rows whose OWN-clip block entries count as -inf (loss.py:101-106 in-place fill)
  float *row_all, *row_pos, *col_all, *col_pos;
  const float *wr, *wc;
  void* G; int g_dtype;
  int S, B, T, N; float inv_temp;
  int Bc, clip0;   // column clips (== B unless the columns are all-gathered over ranks) and the column clip that
                   // row clip 0 owns: logits are [S, B*T, Bc*N] and row clip a's positives sit in column clip a + clip0
};

// 8 consecutive logits of one row as fp32 (16-byte load for bf16, 2 x 16-byte for fp32)
__device__ __forceinline__ void load8(const void* p, int dtype, size_t i, float (&v)[8]) {
  if (dtype == EXG_F32) {
    const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + i);
    const float4 b = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + i + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
    const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p) + i);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[2 * j] = __uint_as_float(w[j] << 16);
      v[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
    }
  }
}
// raw 16-byte row segment (bf16 x 8) and its conversion; the fp32 path keeps load8
__device__ __forceinline__ void cvt8(const uint4& u, float (&v)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[2 * j] = __uint_as_float(w[j] << 16);
    v[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
  }
}
// Cc % 8 != 0: rows are not 16-byte aligned -> element-wise access of the in-range columns
__device__ __forceinline__ void load8_any(const void* p, int dtype, size_t i, float (&v)[8], bool vec, int nvalid) {
  if (vec) { load8(p, dtype, i, v); return; }
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = j < nvalid ? load1(p, dtype, (long)(i + j)) : 0.f;
}
__device__ __forceinline__ void store8(void* p, int dtype, size_t i, const float (&v)[8]);
__device__ __forceinline__ void store8_any(void* p, int dtype, size_t i, const float (&v)[8], bool vec, int nvalid) {
  if (vec) { store8(p, dtype, i, v); return; }
#pragma unroll
  for (int j = 0; j < 8; ++j) if (j < nvalid) store1(p, dtype, (long)(i + j), v[j]);
}
__device__ __forceinline__ void store8(void* p, int dtype, size_t i, const float (&v)[8]) {
  if (dtype == EXG_F32) {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + i) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + i + 4) = make_float4(v[4], v[5], v[6], v[7]);
  } else {
    uint4 u;
    u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
    u.z = pack_bf16x2(v[4], v[5]); u.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p) + i) = u;
  }
}

// Forward: CTA = FWD_ROWS rows x CB columns of one stage.  Column sums live in registers for the whole CTA (one
// atomic per column per CTA); row sums are reduced over the CTA per 32-row sub-slab (one atomic per row).
__global__ void __launch_bounds__(NT, 2) nce_fwd_kernel(NceArgs p) {
  __shared__ float red[NT / 32][2 * SUB];
  const int s = blockIdx.z;
  const long R = (long)p.B * p.T, Cc = (long)p.Bc * p.N;
  const long rbeg = (long)blockIdx.x * FWD_ROWS;
  const long rend = min(R, rbeg + FWD_ROWS);
  const long c0 = (long)blockIdx.y * CB + (long)threadIdx.x * CPT;
  const size_t base = (size_t)s * R * Cc;
  const float k1 = p.inv_temp * 1.4426950408889634f;
  const bool act = c0 < Cc;
  const long coff = (long)p.clip0 * p.N;
  const bool vec = (Cc % CPT) == 0;
  const int nvalid = act ? (int)min((long)CPT, Cc - c0) : 0;
  int cb[CPT]; bool cv[CPT];
  float ca[CPT], cp[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    cb[j] = j < nvalid ? (int)((c0 + j) / p.N) - p.clip0 : -1 - p.clip0;
    cv[j] = j < nvalid && p.col_valid[c0 + j] != 0;
    ca[j] = cp[j] = 0.f;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long r0 = rbeg; r0 < rend; r0 += SUB) {
    float ra[SUB], rp[SUB];
#pragma unroll
    for (int i = 0; i < SUB; ++i) ra[i] = rp[i] = 0.f;
    if (act) {
      const bool fast = vec && p.dtype == EXG_BF16;
#pragma unroll
      for (int i0 = 0; i0 < SUB; i0 += LDB) {
        uint4 raw[LDB];
        if (fast) {       // issue LDB independent 16-byte loads before consuming any of them
#pragma unroll
          for (int u = 0; u < LDB; ++u) {
            const long r = min(r0 + i0 + u, rend - 1);
            raw[u] = __ldcs(reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.L) + base +
                                                           (size_t)r * Cc + c0));
          }
        }
#pragma unroll
        for (int u = 0; u < LDB; ++u) {
          const int i = i0 + u;
          const long r = r0 + i;
          if (r >= rend) continue;
          float v[8];
          if (fast) cvt8(raw[u], v);
          else load8_any(p.L, p.dtype, base + (size_t)r * Cc + c0, v, vec, nvalid);
          const int a = (int)(r / p.T), t = (int)(r % p.T);
          const bool own = cb[0] == a || cb[CPT - 1] == a;     // can any of this thread's columns be positive here?
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            if (!cv[j]) continue;
            float e = exp2f(fmaf(v[j], k1, -k1));
            if (own && cb[j] == a && p.own_row_mask && p.own_row_mask[r]) e = 0.f;
            ra[i] += e; ca[j] += e;
            if (own && cb[j] == a && p.tgt[((size_t)(c0 + j - coff)) * p.T + t]) { rp[i] += e; cp[j] += e; }
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < SUB; ++i) {
      const float x = warp_sum(ra[i]), y = warp_sum(rp[i]);
      if (lane == 0) { red[warp][i] = x; red[warp][SUB + i] = y; }
    }
    __syncthreads();
    if (threadIdx.x < 2 * SUB) {
      float tsum = 0.f;
#pragma unroll
      for (int w = 0; w < NT / 32; ++w) tsum += red[w][threadIdx.x];
      const int i = threadIdx.x % SUB;
      if (r0 + i < rend && tsum != 0.f) {
        float* dst = threadIdx.x < SUB ? p.row_all : p.row_pos;
        atomicAdd(dst + (size_t)s * R + r0 + i, tsum);
      }
    }
    __syncthreads();
  }
  if (act) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      if (!cv[j]) continue;
      atomicAdd(p.col_all + (size_t)s * Cc + c0 + j, ca[j]);
      if (cp[j] != 0.f) atomicAdd(p.col_pos + (size_t)s * Cc + c0 + j, cp[j]);
    }
  }
}

// Backward: G = inv_temp * e * (wr/row_all + wc/col_all - pos * (wr/row_pos + wc/col_pos)); one streaming pass.
__global__ void __launch_bounds__(NT) nce_bwd_kernel(NceArgs p) {
  __shared__ float ra_s[BWD_ROWS], rp_s[BWD_ROWS], rl_s[BWD_ROWS];
  __shared__ int ncols_s;
  const int s = blockIdx.z;
  const long R = (long)p.B * p.T, Cc = (long)p.Bc * p.N;
  const long r0 = (long)blockIdx.x * BWD_ROWS;
  const int nrows = (int)min((long)BWD_ROWS, R - r0);
  const size_t base = (size_t)s * R * Cc;
  const float k1 = p.inv_temp * 1.4426950408889634f;
  // Own-row mask only: positives the mask removed still receive the reference's gradient when NO positive of
  // their weighted row (column) survives -- its positive LSE then runs over -6e4 fills alone and spreads 1/#entries
  // over them (loss.py:101-106 fills under no_grad, so autograd treats the filled logits as the originals).
  if (p.own_row_mask) {
    if (threadIdx.x == 0) ncols_s = 0;
    __syncthreads();
    int n = 0;
    for (long c = threadIdx.x; c < Cc; c += NT) n += p.col_valid[c] != 0;
    n = (int)warp_sum((float)n);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(&ncols_s, n);
    __syncthreads();
  }
  if (threadIdx.x < BWD_ROWS) {
    float a = 0.f, b = 0.f, l = 0.f;
    if (threadIdx.x < nrows) {
      const long r = r0 + threadIdx.x;
      const float w = p.wr[r];
      if (w != 0.f) {
        a = w / p.row_all[(size_t)s * R + r];
        const float rp = p.row_pos[(size_t)s * R + r];
        b = rp > 0.f ? w / rp : 0.f;
        if (p.own_row_mask && !(rp > 0.f)) l = w / (float)ncols_s;
      }
    }
    ra_s[threadIdx.x] = a * p.inv_temp; rp_s[threadIdx.x] = b * p.inv_temp; rl_s[threadIdx.x] = l * p.inv_temp;
  }
  __syncthreads();
  const long c0 = (long)blockIdx.y * CB + (long)threadIdx.x * CPT;
  if (c0 >= Cc) return;
  const long coff = (long)p.clip0 * p.N;
  const bool vec = (Cc % CPT) == 0;
  const int nvalid = (int)min((long)CPT, Cc - c0);
  int cb[CPT]; bool cv[CPT]; float ka[CPT], kp[CPT], kl[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    cb[j] = j < nvalid ? (int)((c0 + j) / p.N) - p.clip0 : -1 - p.clip0;
    cv[j] = j < nvalid && p.col_valid[c0 + j] != 0;
    const float w = j < nvalid ? p.wc[c0 + j] : 0.f;
    ka[j] = kp[j] = kl[j] = 0.f;
    if (cv[j] && w != 0.f) {
      ka[j] = p.inv_temp * w / p.col_all[(size_t)s * Cc + c0 + j];
      const float cpv = p.col_pos[(size_t)s * Cc + c0 + j];
      kp[j] = cpv > 0.f ? p.inv_temp * w / cpv : 0.f;
      if (p.own_row_mask && !(cpv > 0.f)) kl[j] = p.inv_temp * w / (float)((long)p.Bc * p.T);
    }
  }
  const bool fast = vec && p.dtype == EXG_BF16;
  for (int i0 = 0; i0 < nrows; i0 += LDB) {
    uint4 raw[LDB];
    if (fast) {
#pragma unroll
      for (int u = 0; u < LDB; ++u) {
        const long r = r0 + min(i0 + u, nrows - 1);
        raw[u] = __ldcs(reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.L) + base +
                                                       (size_t)r * Cc + c0));
      }
    }
#pragma unroll
    for (int u = 0; u < LDB; ++u) {
      const int i = i0 + u;
      if (i >= nrows) break;
      const long r = r0 + i;
      const int a = (int)(r / p.T), t = (int)(r % p.T);
      float v[8], g[8];
      if (fast) cvt8(raw[u], v);
      else load8_any(p.L, p.dtype, base + (size_t)r * Cc + c0, v, vec, nvalid);
      const float ra = ra_s[i], rp = rp_s[i];
      const bool own = cb[0] == a || cb[CPT - 1] == a;
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        float e = exp2f(fmaf(v[j], k1, -k1));
        const bool dropped = own && cb[j] == a && p.own_row_mask && p.own_row_mask[r];
        if (dropped) e = 0.f;
        float coef = ra + ka[j], leak = 0.f;
        if (own && cb[j] == a && p.tgt[((size_t)(c0 + j - coff)) * p.T + t]) {
          coef -= rp + kp[j];
          if (dropped) leak = rl_s[i] + kl[j];
        }
        g[j] = cv[j] ? e * coef - leak : 0.f;
      }
      store8_any(p.G, p.g_dtype, base + (size_t)r * Cc + c0, g, vec, nvalid);
    }
  }
}

__global__ void own_max_kernel(const void* __restrict__ L, int dtype, float* __restrict__ out, int stage,
                               int B, int T, int N, int Bc, int clip0) {
  const int b = blockIdx.x;
  const long R = (long)B * T, Cc = (long)Bc * N;
  for (int k = threadIdx.x; k < N; k += blockDim.x) {
    float m = -INFINITY;
    for (int t = 0; t < T; ++t)
      m = fmaxf(m, load1(L, dtype, (size_t)stage * R * Cc + ((size_t)b * T + t) * Cc + (size_t)(b + clip0) * N + k));
    out[(size_t)b * N + k] = m;
  }
}

}  // namespace
int exg_nce_tma_fwd(const void*, int, const uint32_t*, const uint8_t*, const uint8_t*, float*, float*, float*, float*,
                    int, int, int, int, int, int, float, cudaStream_t);
int exg_nce_tma_bwd(const void*, int, const uint32_t*, const uint8_t*, const uint8_t*, const float*, const float*,
                    const float*, const float*, const float*, const float*, void*, int, int, int, int, int, int, int,
                    float, cudaStream_t);
namespace {

int nce_check(const void* logits, int dtype, const uint8_t* tgt, const uint8_t* col_valid, int S, int B, int T,
              int N, int Bc, int clip0) {
  EXG_CHECK_ARG(logits && tgt && col_valid, "nce: null pointer");
  EXG_CHECK_ARG(exg_valid_dtype(dtype), "nce: bad dtype");
  EXG_CHECK_ARG(S > 0 && B > 0 && T > 0 && N > 0 && S <= 65535, "nce: bad sizes");
  EXG_CHECK_ARG(clip0 >= 0 && clip0 + B <= Bc, "nce: own column block [clip0, clip0 + B) outside the %d column clips", Bc);
  return 0;
}

}  // namespace

extern "C" int exg_nce_fwd(const void* logits, int dtype, const uint8_t* tgt, const uint32_t* tgt_bits,
                           const uint8_t* col_valid, const uint8_t* own_row_mask, float* row_all, float* row_pos, float* col_all, float* col_pos, int S, int B, int T,
                           int N, int Bc, int clip0, float inv_temp, void* stream) {
  int rc = nce_check(logits, dtype, tgt, col_valid, S, B, T, N, Bc, clip0);
  if (rc) return rc;
  EXG_CHECK_ARG(row_all && row_pos && col_all && col_pos, "nce_fwd: null output");
  NceArgs a = {};
  a.L = logits; a.dtype = dtype; a.tgt = tgt; a.col_valid = col_valid; a.own_row_mask = own_row_mask;
  a.row_all = row_all; a.row_pos = row_pos;
  a.col_all = col_all; a.col_pos = col_pos; a.S = S; a.B = B; a.T = T; a.N = N; a.inv_temp = inv_temp; a.Bc = Bc; a.clip0 = clip0;
  const long R = (long)B * T, Cc = (long)Bc * N;
  cudaStream_t st = (cudaStream_t)stream;
  // the kernel accumulates with atomics: zero the four outputs here (captured as memset nodes under a graph)
  EXG_CUDA(cudaMemsetAsync(row_all, 0, (size_t)S * R * 4, st));
  EXG_CUDA(cudaMemsetAsync(row_pos, 0, (size_t)S * R * 4, st));
  EXG_CUDA(cudaMemsetAsync(col_all, 0, (size_t)S * Cc * 4, st));
  EXG_CUDA(cudaMemsetAsync(col_pos, 0, (size_t)S * Cc * 4, st));
  dim3 grid((unsigned)((R + FWD_ROWS - 1) / FWD_ROWS), (unsigned)((Cc + CB - 1) / CB), (unsigned)S);
  ExgProfScope prof(EXG_PC_NCE_FWD, (double)S * R * Cc * exg_dtype_size(dtype), stream);
  rc = exg_nce_tma_fwd(logits, dtype, tgt_bits, col_valid, own_row_mask, row_all, row_pos, col_all, col_pos, S, B, T, N, Bc, clip0, inv_temp, st);
  if (rc) return rc == 1 ? 0 : rc;
  nce_fwd_kernel<<<grid, NT, 0, st>>>(a);
  EXG_LAUNCH_CHECK("nce_fwd");
  return 0;
}

extern "C" int exg_nce_bwd(const void* logits, int dtype, const uint8_t* tgt, const uint32_t* tgt_bits,
                           const uint8_t* col_valid, const uint8_t* own_row_mask, const float* row_all, const float* row_pos, const float* col_all,
                           const float* col_pos, const float* wr, const float* wc, void* G, int g_dtype, int S,
                           int B, int T, int N, int Bc, int clip0, float inv_temp, void* stream) {
  int rc = nce_check(logits, dtype, tgt, col_valid, S, B, T, N, Bc, clip0);
  if (rc) return rc;
  EXG_CHECK_ARG(row_all && row_pos && col_all && col_pos && wr && wc && G && exg_valid_dtype(g_dtype),
                "nce_bwd: bad argument");
  NceArgs a = {};
  a.L = logits; a.dtype = dtype; a.tgt = tgt; a.col_valid = col_valid; a.own_row_mask = own_row_mask;
  a.row_all = const_cast<float*>(row_all); a.row_pos = const_cast<float*>(row_pos);
  a.col_all = const_cast<float*>(col_all); a.col_pos = const_cast<float*>(col_pos);
  a.wr = wr; a.wc = wc; a.G = G; a.g_dtype = g_dtype; a.S = S; a.B = B; a.T = T; a.N = N; a.inv_temp = inv_temp; a.Bc = Bc; a.clip0 = clip0;
  const long R = (long)B * T, Cc = (long)Bc * N;
  dim3 grid((unsigned)((R + BWD_ROWS - 1) / BWD_ROWS), (unsigned)((Cc + CB - 1) / CB), (unsigned)S);
  ExgProfScope prof(EXG_PC_NCE_BWD, (double)S * R * Cc * (exg_dtype_size(dtype) + exg_dtype_size(g_dtype)), stream);
  rc = exg_nce_tma_bwd(logits, dtype, tgt_bits, col_valid, own_row_mask, row_all, row_pos, col_all, col_pos, wr, wc, G, g_dtype, S, B, T,
                       N, Bc, clip0, inv_temp, (cudaStream_t)stream);
  if (rc) return rc == 1 ? 0 : rc;
  nce_bwd_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("nce_bwd");
  return 0;
}

extern "C" int exg_own_max(const void* logits, int dtype, float* own_max, int stage, int B, int T, int N,
                           int Bc, int clip0, void* stream) {
  EXG_CHECK_ARG(logits && own_max && exg_valid_dtype(dtype) && B > 0 && T > 0 && N > 0 && stage >= 0 &&
                clip0 >= 0 && clip0 + B <= Bc, "own_max: bad argument");
  ExgProfScope prof(EXG_PC_ROWWISE, (double)B * T * N * exg_dtype_size(dtype), stream);
  own_max_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(logits, dtype, own_max, stage, B, T, N, Bc, clip0);
  EXG_LAUNCH_CHECK("own_max");
  return 0;
}
