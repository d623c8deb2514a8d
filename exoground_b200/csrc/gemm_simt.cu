// gemm_simt.cu — fp32-accumulate SIMT GEMM with the full epilogue set of exg_gemm.
// This is the fp32 (parity / bit-exact-argmax) path and the fallback for shapes the tcgen05 kernel
// does not take (unaligned strides, tiny N such as the 512->1 / 512->2 heads).  Any M,N,K, any
// transpose combination, fp32 or bf16 operands, fp32 FMA accumulation in k order.
#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 16, TM = 8, TN = 8, THREADS = 256;

struct SimtArgs {
  const void* A; const void* B; void* C;
  int M, N, K;
  long lda, ldb, ldc;
  int transA, transB;
  long sA, sB, sC;
  int in_dtype, out_dtype, epilogue;
  const float* bias;
  const void* aux; long ldaux, sAux;
  void* aux_out;
  float alpha;
  int splits, k_per_split;   // split-K (fp32 atomics) for short M x N, long K (head weight gradients)
};

__device__ __forceinline__ float epilogue_apply(const SimtArgs& p, float v, long m, long n, long z,
                                                float* pre_out) {
  *pre_out = v;
  switch (p.epilogue) {
    case EXG_EPI_QGELU: return quick_gelu(v);
    case EXG_EPI_RESID:
      return reinterpret_cast<const float*>(p.aux)[z * p.sAux + m * p.ldaux + n] + v;
    case EXG_EPI_QGELU_BWD:
      return v * quick_gelu_grad(load1(p.aux, p.in_dtype, z * p.sAux + m * p.ldaux + n));
    case EXG_EPI_RELU: return fmaxf(v, 0.f);
    case EXG_EPI_RELU_BWD:
      return load1(p.aux, p.in_dtype, z * p.sAux + m * p.ldaux + n) > 0.f ? v : 0.f;
    default: return v;
  }
}

__global__ void __launch_bounds__(THREADS) gemm_simt_kernel(SimtArgs p) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const long z = blockIdx.z / p.splits;
  const int split = blockIdx.z % p.splits;
  const int kbeg = split * p.k_per_split, kend = min(p.K, kbeg + p.k_per_split);
  const long m0 = (long)blockIdx.y * BM, n0 = (long)blockIdx.x * BN;
  const char* Ab = reinterpret_cast<const char*>(p.A);
  const char* Bb = reinterpret_cast<const char*>(p.B);
  const size_t es = p.in_dtype == EXG_F32 ? 4 : 2;
  const void* A = Ab + (size_t)z * p.sA * es;
  const void* B = Bb + (size_t)z * p.sB * es;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  const int tx = tid % 16, ty = tid / 16;   // micro-tile: rows ty*8.., cols tx*8..

  for (int k0 = kbeg; k0 < kend; k0 += BK) {
    // ---- stage A tile (BM x BK) into As[k][m]
    if (!p.transA) {
      const int r = tid / 2, kk = (tid % 2) * 8;
      const long gm = m0 + r;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int gk = k0 + kk + i;
        As[kk + i][r] = (gm < p.M && gk < kend) ? load1(A, p.in_dtype, gm * p.lda + gk) : 0.f;
      }
    } else {
      const int kk = tid / 16, mm = (tid % 16) * 8;
      const int gk = k0 + kk;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long gm = m0 + mm + i;
        As[kk][mm + i] = (gm < p.M && gk < kend) ? load1(A, p.in_dtype, (long)gk * p.lda + gm) : 0.f;
      }
    }
    // ---- stage B tile (BK x BN) into Bs[k][n]
    if (!p.transB) {   // stored [N,K]
      const int c = tid / 2, kk = (tid % 2) * 8;
      const long gn = n0 + c;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int gk = k0 + kk + i;
        Bs[kk + i][c] = (gn < p.N && gk < kend) ? load1(B, p.in_dtype, gn * p.ldb + gk) : 0.f;
      }
    } else {           // stored [K,N]
      const int kk = tid / 16, nn = (tid % 16) * 8;
      const int gk = k0 + kk;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long gn = n0 + nn + i;
        Bs[kk][nn + i] = (gn < p.N && gk < kend) ? load1(B, p.in_dtype, (long)gk * p.ldb + gn) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], b[TN];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * 8 + 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][tx * 8]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[k][tx * 8 + 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const long m = m0 + ty * 8 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const long n = n0 + tx * 8 + j;
      if (n >= p.N) continue;
      float v = acc[i][j] * p.alpha;
      if (p.bias) v += p.bias[n];
      float pre;
      const float o = epilogue_apply(p, v, m, n, z, &pre);
      const long ci = z * p.sC + m * p.ldc + n;
      if (p.splits > 1) { atomicAdd(reinterpret_cast<float*>(p.C) + ci, o); continue; }
      store1(p.C, p.out_dtype, ci, o);
      if (p.epilogue == EXG_EPI_QGELU && p.aux_out) store1(p.aux_out, p.out_dtype, ci, pre);
    }
  }
}

}  // namespace

int exg_gemm_simt(const exg_gemm_params* q, void* stream) {
  SimtArgs a;
  a.A = q->A; a.B = q->B; a.C = q->C; a.M = q->M; a.N = q->N; a.K = q->K;
  a.lda = q->lda; a.ldb = q->ldb; a.ldc = q->ldc; a.transA = q->transA; a.transB = q->transB;
  a.sA = q->strideA; a.sB = q->strideB; a.sC = q->strideC;
  a.in_dtype = q->in_dtype; a.out_dtype = q->out_dtype; a.epilogue = q->epilogue;
  a.bias = q->bias; a.aux = q->aux; a.ldaux = q->ldaux; a.sAux = q->strideAux; a.aux_out = q->aux_out;
  a.alpha = q->alpha;
  a.splits = 1; a.k_per_split = q->K;
  const long tiles = (long)((q->N + BN - 1) / BN) * ((q->M + BM - 1) / BM) * q->batch;
  if (q->epilogue == EXG_EPI_NONE && !q->bias && q->out_dtype == EXG_F32 && q->batch == 1 &&
      tiles * 2 <= exg_num_sms() && q->K >= 512) {
    int want = (int)(2L * exg_num_sms() / tiles);
    int maxs = q->K / 128;
    a.splits = want < maxs ? want : maxs;
    if (a.splits < 1) a.splits = 1;
    a.k_per_split = ((q->K + a.splits - 1) / a.splits + BK - 1) / BK * BK;
    a.splits = (q->K + a.k_per_split - 1) / a.k_per_split;
    if (a.splits > 1)
      EXG_CUDA(cudaMemset2DAsync(q->C, (size_t)q->ldc * 4, 0, (size_t)q->N * 4, (size_t)q->M, (cudaStream_t)stream));
  }
  dim3 grid((unsigned)((q->N + BN - 1) / BN), (unsigned)((q->M + BM - 1) / BM), (unsigned)(q->batch * a.splits));
  gemm_simt_kernel<<<grid, THREADS, 0, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("gemm_simt");
  return 0;
}
