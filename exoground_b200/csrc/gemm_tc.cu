// gemm_tc.cu — bf16 GEMM on 5th-gen tensor cores (sm_100a): TMA -> 128B-swizzled smem ring ->
// tcgen05.mma (single issuing thread, fp32 accumulator in TMEM) -> tcgen05.ld epilogue.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..5 = epilogue (each owns the 32 TMEM lanes its warp-id%4 may touch).
// Tile 128 x BN x 64; one output tile per CTA; optional split-K with fp32 atomics for the
// weight-gradient shapes (tiny M x N, long K).  Operands may be K-major or MN-major in global memory;
// MN-major tiles are fetched as 64-wide column blocks and described to the MMA with the
// transposed-operand bits of the instruction descriptor, so backward GEMMs need no transposes.
#include "common.cuh"
#include "sm100.cuh"

using namespace sm100;

namespace {

constexpr int BM = 128, BK = 64, UK = 16;
constexpr int kThreads = 192;
constexpr uint32_t kABytes = BM * BK * 2;   // 16 KB

struct TcArgs {
  void* C; long ldc, sC;
  int M, N, K;
  int a_mn, b_mn;
  int out_dtype, epilogue, aux_dtype;
  const float* bias;
  const void* aux; long ldaux, sAux;
  void* aux_out;
  float alpha;
  int splits, kblocks_per_split, atomic;
  int a_bcast, b_bcast;   // operand shared by all batches (batch stride 0): TMA z coordinate pinned to 0
};

template <int BN> struct Cfg {
  static constexpr uint32_t kBBytes = BN * BK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr uint32_t kSmem = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
};

template <int BN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, TcArgs p) {
  using C_ = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C_::kStages * C_::kStageBytes);
  uint64_t* empty = full + C_::kStages;
  uint64_t* tmem_full = empty + C_::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int z = blockIdx.z / p.splits, split = blockIdx.z % p.splits;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int kblocks_total = (p.K + BK - 1) / BK;
  const int kb0 = split * p.kblocks_per_split;
  const int kb1 = min(kblocks_total, kb0 + p.kblocks_per_split);
  const int nkb = kb1 - kb0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C_::kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
  if (warp == 1) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % C_::kStages;
        const uint32_t ph = (i / C_::kStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        uint8_t* sa = smem + (size_t)s * C_::kStageBytes;
        uint8_t* sb = sa + kABytes;
        const int k0 = (kb0 + i) * BK;
        mbar_arrive_expect_tx(&full[s], C_::kStageBytes);
        const int za = p.a_bcast ? 0 : z, zb = p.b_bcast ? 0 : z;
        if (!p.a_mn) {
          tma_load_3d(sa, &tmA, &full[s], k0, m0, za);
        } else {
#pragma unroll
          for (int j = 0; j < BM / 64; ++j) tma_load_3d(sa + j * 8192, &tmA, &full[s], m0 + j * 64, k0, za);
        }
        if (!p.b_mn) {
          tma_load_3d(sb, &tmB, &full[s], k0, n0, zb);
        } else {
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) tma_load_3d(sb + j * 8192, &tmB, &full[s], n0 + j * 64, k0, zb);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(BM, BN, p.a_mn, p.b_mn);
      const uint32_t a_lbo = p.a_mn ? 8192u : 16u, b_lbo = p.b_mn ? 8192u : 16u;
      const uint32_t a_kstep = p.a_mn ? (UK * 128u) : (UK * 2u);
      const uint32_t b_kstep = p.b_mn ? (UK * 128u) : (UK * 2u);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % C_::kStages;
        const uint32_t ph = (i / C_::kStages) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + (size_t)s * C_::kStageBytes);
        const uint32_t sb = sa + kABytes;
#pragma unroll
        for (int k = 0; k < BK / UK; ++k) {
          const uint64_t ad = umma_smem_desc(sa + k * a_kstep, a_lbo, 1024u);
          const uint64_t bd = umma_smem_desc(sb + k * b_kstep, b_lbo, 1024u);
          umma_bf16(tmem_base, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty[s]);     // frees the smem slot once these MMAs retire
      }
      umma_commit(tmem_full);       // accumulator complete
    }
  } else {
    // ===================== epilogue =====================
    const int lg = warp & 3;                       // TMEM lane group this warp may access
    const long row = (long)m0 + lg * 32 + lane;
    if (nkb > 0) {
      mbar_wait(tmem_full, 0);
      tc_fence_after();
    }
    const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16);
    const bool row_ok = row < p.M;
    const long crow = (long)z * p.sC + row * p.ldc;
    const long arow = (long)z * p.sAux + row * p.ldaux;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      if (n0 + c0 >= p.N) break;
      uint32_t r[32];
      if (nkb > 0) {
        tmem_ld_32x32(taddr + c0, r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0u;
      }
      if (!row_ok) continue;
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const int n = n0 + c0 + j;
        if (n >= p.N) break;
        float4 v = make_float4(__uint_as_float(r[j]) * p.alpha, __uint_as_float(r[j + 1]) * p.alpha,
                               __uint_as_float(r[j + 2]) * p.alpha, __uint_as_float(r[j + 3]) * p.alpha);
        if (p.bias) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n));
          v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
        }
        float4 o = v;
        switch (p.epilogue) {
          case EXG_EPI_QGELU:
            o = make_float4(quick_gelu(v.x), quick_gelu(v.y), quick_gelu(v.z), quick_gelu(v.w));
            if (p.aux_out) store4(p.aux_out, p.out_dtype, crow + n, v);
            break;
          case EXG_EPI_RESID: {
            const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux) + arow + n);
            o = make_float4(a.x + v.x, a.y + v.y, a.z + v.z, a.w + v.w);
          } break;
          case EXG_EPI_QGELU_BWD: {
            const float4 a = load4(p.aux, p.aux_dtype, arow + n);
            o = make_float4(v.x * quick_gelu_grad(a.x), v.y * quick_gelu_grad(a.y),
                            v.z * quick_gelu_grad(a.z), v.w * quick_gelu_grad(a.w));
          } break;
          case EXG_EPI_RELU:
            o = make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
            break;
          case EXG_EPI_RELU_BWD: {
            const float4 a = load4(p.aux, p.aux_dtype, arow + n);
            o = make_float4(a.x > 0.f ? v.x : 0.f, a.y > 0.f ? v.y : 0.f, a.z > 0.f ? v.z : 0.f,
                            a.w > 0.f ? v.w : 0.f);
          } break;
          default: break;
        }
        if (p.atomic) {
          float* c = reinterpret_cast<float*>(p.C) + crow + n;
          atomicAdd(c, o.x); atomicAdd(c + 1, o.y); atomicAdd(c + 2, o.z); atomicAdd(c + 3, o.w);
        } else {
          store4(p.C, p.out_dtype, crow + n, o);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BN);
  }
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 3-D bf16 tensor map over [batch][rows][inner] with a {64, box_rows, 1} box, 128B swizzle.
int make_tmap(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows, uint64_t batch,
              uint64_t row_stride_elems, uint64_t batch_stride_elems, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { exg_set_error("gemm_tc: cuTensorMapEncodeTiled unavailable"); return EXG_EUNSUPPORTED; }
  cuuint64_t dims[3] = {inner, rows, batch};
  cuuint64_t strides[2] = {row_stride_elems * 2, (batch > 1 ? batch_stride_elems : rows * row_stride_elems) * 2};
  cuuint32_t box[3] = {64, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    exg_set_error("gemm_tc: cuTensorMapEncodeTiled failed (%d) inner=%llu rows=%llu batch=%llu ld=%llu",
                  (int)r, (unsigned long long)inner, (unsigned long long)rows, (unsigned long long)batch,
                  (unsigned long long)row_stride_elems);
    return EXG_EINVAL;
  }
  return 0;
}

template <int BN>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const TcArgs& a, int batch, cudaStream_t s) {
  static bool attr_set = false;
  auto kfn = gemm_tc_kernel<BN>;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg<BN>::kSmem);
    if (e != cudaSuccess) { exg_set_error("gemm_tc: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr_set = true;
  }
  dim3 grid((unsigned)((a.N + BN - 1) / BN), (unsigned)((a.M + BM - 1) / BM), (unsigned)(batch * a.splits));
  kfn<<<grid, kThreads, Cfg<BN>::kSmem, s>>>(ta, tb, a);
  EXG_LAUNCH_CHECK("gemm_tc");
  return 0;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

extern "C" int exg_gemm_tc_supported(const exg_gemm_params* p) {
  if (!p || p->in_dtype != EXG_BF16) return 0;
  if (p->M <= 0 || p->N <= 0 || p->K <= 0 || p->batch <= 0) return 0;
  if (p->N % 8 != 0) return 0;
  if (!aligned16(p->A) || !aligned16(p->B) || !aligned16(p->C)) return 0;
  if (p->lda % 8 || p->ldb % 8 || p->ldc % 4) return 0;
  if (p->batch > 1 && (p->strideA % 8 || p->strideB % 8 || p->strideC % 4)) return 0;
  if (p->aux && (p->ldaux % 4 || !aligned16(p->aux) || (p->batch > 1 && p->strideAux % 4))) return 0;
  if (p->aux_out && !aligned16(p->aux_out)) return 0;
  if (p->bias && !aligned16(p->bias)) return 0;
  // the contiguous (inner) extent must cover at least one 16-byte unit
  if ((p->transA ? p->M : p->K) < 8 || (p->transB ? p->N : p->K) < 8) return 0;
  return 1;
}

int exg_gemm_tc(const exg_gemm_params* q, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  TcArgs a;
  a.C = q->C; a.ldc = q->ldc; a.sC = q->strideC;
  a.M = q->M; a.N = q->N; a.K = q->K;
  a.a_mn = q->transA ? 1 : 0; a.b_mn = q->transB ? 1 : 0;
  a.out_dtype = q->out_dtype; a.epilogue = q->epilogue; a.aux_dtype = q->in_dtype;
  a.bias = q->bias; a.aux = q->aux; a.ldaux = q->ldaux; a.sAux = q->strideAux; a.aux_out = q->aux_out;
  a.alpha = q->alpha;

  const int kblocks = (q->K + BK - 1) / BK;
  const long tiles256 = (long)((q->M + BM - 1) / BM) * ((q->N + 255) / 256) * q->batch;
  const int bn = (q->N >= 256 && tiles256 >= exg_num_sms()) ? 256 : 128;
  const long tiles = (long)((q->M + BM - 1) / BM) * ((q->N + bn - 1) / bn) * q->batch;
  a.splits = 1; a.atomic = 0;
  if (q->epilogue == EXG_EPI_NONE && !q->bias && q->out_dtype == EXG_F32 && q->batch == 1 &&
      tiles * 2 <= exg_num_sms() && kblocks >= 8) {
    int want = (int)((exg_num_sms() + tiles - 1) / tiles);
    int maxs = kblocks / 4;
    a.splits = want < maxs ? want : maxs;
    if (a.splits < 1) a.splits = 1;
  }
  a.kblocks_per_split = (kblocks + a.splits - 1) / a.splits;
  a.splits = (kblocks + a.kblocks_per_split - 1) / a.kblocks_per_split;
  if (a.splits > 1) {
    a.atomic = 1;
    EXG_CUDA(cudaMemset2DAsync(q->C, (size_t)q->ldc * 4, 0, (size_t)q->N * 4, (size_t)q->M, s));
  }

  CUtensorMap ta, tb;
  int rc;
  a.a_bcast = (q->batch > 1 && q->strideA == 0) ? 1 : 0;
  a.b_bcast = (q->batch > 1 && q->strideB == 0) ? 1 : 0;
  const int ba = a.a_bcast ? 1 : q->batch, bb = a.b_bcast ? 1 : q->batch;
  if (!q->transA) rc = make_tmap(&ta, q->A, q->K, q->M, ba, q->lda, q->strideA, BM);
  else            rc = make_tmap(&ta, q->A, q->M, q->K, ba, q->lda, q->strideA, BK);
  if (rc) return rc;
  if (!q->transB) rc = make_tmap(&tb, q->B, q->K, q->N, bb, q->ldb, q->strideB, bn);
  else            rc = make_tmap(&tb, q->B, q->N, q->K, bb, q->ldb, q->strideB, BK);
  if (rc) return rc;
  return bn == 256 ? launch<256>(ta, tb, a, q->batch, s) : launch<128>(ta, tb, a, q->batch, s);
}
