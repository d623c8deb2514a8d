// gemm_ln.cu — out-projection GEMM with the residual add and the LayerNorm that follow it fused into the epilogue.
//
//   s = resid + A @ W^T + bias          (fp32 residual stream, written once)
//   y = LayerNorm(s) * gamma + beta      (bf16: the operand of the next GEMM / the stage buffer row)
//
// replaces the pair {GEMM (N = 512), exg_add_layernorm_fwd} of every attention out-projection and every FFN c_proj of
// the encoder / decoder blocks (model/tfm_model.py:22,28,34-38: `x = x + attn(ln_1(x)); x = x + mlp(ln_2(x))` followed
// by the next block's ln_1 / ln_2 / the post-encoder LayerNorm).  The model width C = 512 fp32 columns is exactly one
// 128-lane x 512-column TMEM accumulator, so a CTA owns whole rows: thread = row, the row statistics never leave the
// thread (two warps per lane quarter split the columns and exchange two floats per row through shared memory), and the
// [M, 512] delta tensor and one full read + write pass over the residual stream disappear (~20 us per call at
// M = 12288, profiles/launches_r01f_summary.txt: ln_fwd 120 launches, 1.6 ms per step).
//
//   warp 0      : TMA producer  (A tile 128 x 64 and the whole 512 x 64 weight slab per k-block, 128B swizzle, 2 stages)
//   warp 1      : MMA issuer    (tcgen05.mma cta_group::1, two 128 x 256 x 16 instructions per k-step: columns 0-255 / 256-511)
//   warps 2..9  : epilogue      pass 1: tcgen05.ld 32 columns -> + bias + residual (16-byte global loads issued a chunk
//                                       ahead) -> s to global (16-byte stores) and back into TMEM (tcgen05.st), row sum / sum
//                                       of squares in registers;  pass 2: tcgen05.ld s -> normalise -> bf16 -> global.
// The accumulator fills TMEM, so MMA and epilogue of a tile do not overlap (the producer does run ahead into the next
// tile); per tile ~ K/16 * 256 + ~7000 cycles.  Tensor-bound for K = 2048, epilogue-bound for K = 512 - either way one
// launch instead of two and 4 KB less HBM traffic per row.
#include "common.cuh"
#include "sm100.cuh"

using namespace sm100;

namespace {

constexpr int BM = 128, BN = 512, BK = 64, UK = 16;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 32 * (2 + kEpiWarps);
constexpr int kStages = 2;
constexpr uint32_t kABytes = BM * BK * 2;              // 16 KB
constexpr uint32_t kBBytes = BN * BK * 2;              // 64 KB (two 256-row boxes)
constexpr uint32_t kStageBytes = kABytes + kBBytes;    // 80 KB
constexpr uint32_t kParamBytes = 3 * BN * 4;           // bias, gamma, beta
constexpr uint32_t kStatBytes = BM * 2 * 2 * 4;        // [row][half][sum, sumsq]
constexpr uint32_t kSmem = kStages * kStageBytes + kParamBytes + kStatBytes + 1024 /*align*/ + 256 /*barriers*/;

struct LnArgs {
  int M, K;
  const float* bias; const float* resid; const float* gamma; const float* beta;
  float* s_out; uint16_t* y_out;
  float eps;
  int tiles;
};

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
      "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
      "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint4 ldg_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_v4(void* p, uint4 v) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void bar_pair(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

__global__ void __launch_bounds__(kThreads, 1)
gemm_ln_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, LnArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* prm = reinterpret_cast<float*>(smem + kStages * kStageBytes);          // bias | gamma | beta
  float* stat = prm + 3 * BN;                                                    // [128][2][2]
  uint64_t* full = reinterpret_cast<uint64_t*>(stat + BM * 4);
  uint64_t* empty = full + kStages;
  uint64_t* tmem_full = empty + kStages;
  uint64_t* tmem_empty = tmem_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = (p.K + BK - 1) / BK;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(tmem_full, 1); mbar_init(tmem_empty, kEpiWarps);
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
  if (warp == 1) tmem_alloc(tmem_slot, BN);
  for (int i = threadIdx.x; i < BN; i += kThreads) {          // the whole row's parameters, once per CTA
    prm[i] = p.bias ? p.bias[i] : 0.f;
    prm[BN + i] = p.gamma[i];
    prm[2 * BN + i] = p.beta[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        const int m0 = tile * BM;
        for (int i = 0; i < nkb; ++i) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* sa = smem + (size_t)s * kStageBytes;
          uint8_t* sb = sa + kABytes;
          mbar_arrive_expect_tx(&full[s], kStageBytes);
          tma_load_3d(sa, &tmA, &full[s], i * BK, m0, 0);
          tma_load_3d(sb, &tmB, &full[s], i * BK, 0, 0);
          tma_load_3d(sb + kBBytes / 2, &tmB, &full[s], i * BK, 256, 0);
          if (++s == kStages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(BM, 256, 0, 0);
      int s = 0; uint32_t ph = 0, ti = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++ti) {
        mbar_wait(tmem_empty, (ti & 1) ^ 1);           // the epilogue has drained the previous tile
        tc_fence_after();
        for (int i = 0; i < nkb; ++i) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)s * kStageBytes);
          const uint32_t sb = sa + kABytes;
#pragma unroll
          for (int k = 0; k < BK / UK; ++k) {
            const uint64_t ad = umma_smem_desc(sa + k * UK * 2, 16u, 1024u);
            const uint64_t b0 = umma_smem_desc(sb + k * UK * 2, 16u, 1024u);
            const uint64_t b1 = umma_smem_desc(sb + kBBytes / 2 + k * UK * 2, 16u, 1024u);
            const uint32_t acc = (i > 0 || k > 0) ? 1u : 0u;
            umma_bf16(tmem_base, ad, b0, idesc, acc);
            umma_bf16(tmem_base + 256, ad, b1, idesc, acc);
          }
          umma_commit(&empty[s]);
          if (++s == kStages) { s = 0; ph ^= 1; }
        }
        umma_commit(tmem_full);
      }
    }
  } else {
    // ===================== epilogue: thread = row =====================
    const int ew = warp - 2;
    const int lg = warp & 3;                  // TMEM lane quarter
    const int half = ew >> 2;                 // columns [256 * half, 256 * half + 256)
    const int col0 = half * 256;
    const uint32_t lane_addr = (uint32_t)(lg * 32) << 16;
    const int rloc = lg * 32 + lane;          // row within the tile
    float* my_stat = stat + (rloc * 2 + half) * 2;
    const float* peer_stat = stat + (rloc * 2 + (half ^ 1)) * 2;
    const float* sb_ = prm, *sg = prm + BN, *sbeta = prm + 2 * BN;
    uint32_t ti = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++ti) {
      const long row = (long)tile * BM + rloc;
      const bool row_ok = row < p.M;
      const float* rres = p.resid + row * BN + col0;
      float* rs = p.s_out + row * BN + col0;
      uint16_t* ry = p.y_out + row * BN + col0;
      // the residual row chunk is requested before the wait for the accumulator and then one chunk ahead
      uint4 nx[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) nx[u] = row_ok ? ldg_v4(rres + 4 * u) : make_uint4(0u, 0u, 0u, 0u);
      mbar_wait(tmem_full, ti & 1);
      tc_fence_after();
      float sum = 0.f, sq = 0.f;
      // ---- pass 1: s = acc + bias + resid; s -> global and back into TMEM; row moments
#pragma unroll 1
      for (int c = 0; c < 256; c += 32) {
        uint4 cur[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) cur[u] = nx[u];
        if (c + 32 < 256) {
#pragma unroll
          for (int u = 0; u < 8; ++u) nx[u] = row_ok ? ldg_v4(rres + c + 32 + 4 * u) : make_uint4(0u, 0u, 0u, 0u);
        }
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + lane_addr + col0 + c, r);
        tmem_ld_wait();
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float4 b = *reinterpret_cast<const float4*>(sb_ + col0 + c + 4 * u);      // broadcast
          const float x0 = __uint_as_float(r[4 * u]) + b.x + __uint_as_float(cur[u].x);
          const float x1 = __uint_as_float(r[4 * u + 1]) + b.y + __uint_as_float(cur[u].y);
          const float x2 = __uint_as_float(r[4 * u + 2]) + b.z + __uint_as_float(cur[u].z);
          const float x3 = __uint_as_float(r[4 * u + 3]) + b.w + __uint_as_float(cur[u].w);
          sum += (x0 + x1) + (x2 + x3);
          sq = fmaf(x0, x0, fmaf(x1, x1, fmaf(x2, x2, fmaf(x3, x3, sq))));
          r[4 * u] = __float_as_uint(x0); r[4 * u + 1] = __float_as_uint(x1);
          r[4 * u + 2] = __float_as_uint(x2); r[4 * u + 3] = __float_as_uint(x3);
          if (row_ok) stg_v4(rs + c + 4 * u, make_uint4(r[4 * u], r[4 * u + 1], r[4 * u + 2], r[4 * u + 3]));
        }
        tmem_st_32x32(tmem_base + lane_addr + col0 + c, r);
      }
      tmem_st_wait();
      // ---- exchange the two halves' moments (the peer warp of this lane quarter owns the other 256 columns)
      my_stat[0] = sum; my_stat[1] = sq;
      bar_pair(1 + lg);
      const float tsum = sum + peer_stat[0], tsq = sq + peer_stat[1];
      const float mean = tsum * (1.0f / BN);
      const float var = fmaxf(tsq * (1.0f / BN) - mean * mean, 0.f);
      const float rstd = rsqrtf(var + p.eps);
      // ---- pass 2: y = (s - mean) * rstd * gamma + beta -> bf16 -> global
#pragma unroll 1
      for (int c = 0; c < 256; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + lane_addr + col0 + c, r);
        tmem_ld_wait();
        if (c + 32 >= 256) {
          // all TMEM reads of this tile are done: the MMA warp may overwrite the accumulator
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tmem_empty);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float y[8];
#pragma unroll
          for (int j = 0; j < 8; j += 4) {
            const float4 g = *reinterpret_cast<const float4*>(sg + col0 + c + 8 * u + j);
            const float4 be = *reinterpret_cast<const float4*>(sbeta + col0 + c + 8 * u + j);
            y[j] = fmaf((__uint_as_float(r[8 * u + j]) - mean) * rstd, g.x, be.x);
            y[j + 1] = fmaf((__uint_as_float(r[8 * u + j + 1]) - mean) * rstd, g.y, be.y);
            y[j + 2] = fmaf((__uint_as_float(r[8 * u + j + 2]) - mean) * rstd, g.z, be.z);
            y[j + 3] = fmaf((__uint_as_float(r[8 * u + j + 3]) - mean) * rstd, g.w, be.w);
          }
          uint4 q;
          q.x = pack_bf16x2(y[0], y[1]); q.y = pack_bf16x2(y[2], y[3]);
          q.z = pack_bf16x2(y[4], y[5]); q.w = pack_bf16x2(y[6], y[7]);
          if (row_ok) stg_v4(ry + c + 8 * u, q);
        }
      }
      bar_pair(1 + lg);          // the peer has read this warp's moments before the next tile overwrites them
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BN);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
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

int make_map(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows, uint64_t ld, uint32_t box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { exg_set_error("gemm_ln: cuTensorMapEncodeTiled unavailable"); return EXG_EUNSUPPORTED; }
  cuuint64_t dims[3] = {inner, rows, 1};
  cuuint64_t strides[2] = {ld * 2, rows * ld * 2};
  cuuint32_t box[3] = {64, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { exg_set_error("gemm_ln: cuTensorMapEncodeTiled failed (%d)", (int)r); return EXG_EINVAL; }
  return 0;
}

}  // namespace

// 1 if exg_gemm_add_ln takes these sizes / alignments (bf16 operands, N = 512).
extern "C" int exg_gemm_add_ln_supported(const void* A, const void* W, int M, int K, long lda, long ldb) {
  if (!A || !W || M <= 0 || K < 64 || K % 8) return 0;
  if (((uintptr_t)A & 15) || ((uintptr_t)W & 15) || lda % 8 || ldb % 8) return 0;
  return 1;
}

extern "C" int exg_gemm_add_ln(const void* A, const void* W, const float* bias, const float* resid, const float* gamma,
                               const float* beta, float eps, float* s_out, void* y_out, int M, int K, long lda,
                               long ldb, void* stream) {
  EXG_CHECK_ARG(exg_gemm_add_ln_supported(A, W, M, K, lda, ldb), "gemm_add_ln: unsupported operands (bf16, 16-byte rows)");
  EXG_CHECK_ARG(resid && gamma && beta && s_out && y_out, "gemm_add_ln: null pointer");
  EXG_CHECK_ARG(!((uintptr_t)resid & 15) && !((uintptr_t)s_out & 15) && !((uintptr_t)y_out & 15),
                "gemm_add_ln: residual / outputs must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  CUtensorMap ta, tb;
  int rc = make_map(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BM);
  if (rc) return rc;
  rc = make_map(&tb, W, (uint64_t)K, (uint64_t)BN, (uint64_t)ldb, 256);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    EXG_CUDA(cudaFuncSetAttribute(gemm_ln_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    attr_set = true;
  }
  LnArgs a;
  a.M = M; a.K = K; a.bias = bias; a.resid = resid; a.gamma = gamma; a.beta = beta; a.s_out = s_out;
  a.y_out = reinterpret_cast<uint16_t*>(y_out); a.eps = eps;
  a.tiles = (M + BM - 1) / BM;
  const int sms = exg_persistent_sms();
  const int grid = a.tiles < sms ? a.tiles : sms;
  ExgProfScope prof(EXG_PC_GEMM_TC, 2.0 * M * (double)BN * K, stream);
  gemm_ln_kernel<<<grid, kThreads, kSmem, s>>>(ta, tb, a);
  EXG_LAUNCH_CHECK("gemm_add_ln");
  return 0;
}
