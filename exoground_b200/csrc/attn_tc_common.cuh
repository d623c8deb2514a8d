// attn_tc_common.cuh -- shared by attention_tc.cu (<= 128 keys, single tile) and attention_flash.cu (any length):
// 128B-swizzled tile addressing, the Q K^T style MMA, tensor maps of the [B][S][ld] q/k/v/o views.
#pragma once
#include "common.cuh"
#include "sm100.cuh"
#include <math.h>

using namespace sm100;

namespace {

constexpr int TQ = 128;            // query rows per CTA (UMMA M)
constexpr int TK = 128;            // max keys per CTA  (UMMA N of the score MMAs)
constexpr int HD = 64;             // head dim
constexpr uint32_t kTile = TQ * HD * 2;     // 16 KB: [128 rows][64 bf16], 128B-swizzled
constexpr float kLog2e = 1.4426950408889634f;

struct AtcArgs {
  int B, H, Sq, Sk;
  float scale;
  const uint8_t* key_pad;
  float* lse;
  // attention_tc.cu only: rows of the smem query / key tiles (multiples of 16, <= 128) - short sequences get short
  // tiles, which is what lets several CTAs share an SM
  int tq_rows, tk_rows;
};

// ---- smem addressing of a 128B-swizzled [rows][64 bf16] tile: 16-byte unit u of row r
__device__ __forceinline__ uint32_t swz(int r, int u) { return (uint32_t)(r * 128 + ((u ^ (r & 7)) << 4)); }

__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 q;
  q.x = pack_bf16x2(f[0], f[1]); q.y = pack_bf16x2(f[2], f[3]);
  q.z = pack_bf16x2(f[4], f[5]); q.w = pack_bf16x2(f[6], f[7]);
  return q;
}

// S-type MMA: D[128 x N] (+)= A[128 x 64] * B[N x 64]^T, both K-major tiles (4 k-steps of 16)
__device__ __forceinline__ void mma_kmajor_k64(uint32_t d_tmem, uint32_t sa, uint32_t sb, int N) {
  const uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
#pragma unroll
  for (int k = 0; k < 4; ++k)
    umma_bf16(d_tmem, umma_smem_desc(sa + k * 32, 16, 1024), umma_smem_desc(sb + k * 32, 16, 1024), idesc, k > 0);
}

// key-validity bitmask for the 128 keys of this clip: bit (k & 31) of word k >> 5
__device__ __forceinline__ void build_mask(uint32_t* maskw, const AtcArgs& p, int b, int tid) {
  const bool ok = tid < p.Sk && !(p.key_pad && p.key_pad[(long)b * p.Sk + tid]);
  const uint32_t bits = __ballot_sync(0xffffffffu, ok);
  if ((tid & 31) == 0) maskw[tid >> 5] = bits;
}

// ------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
EncodeTiledFn enc_fn() {
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

// [B][S][ld] bf16 view of one of q / k / v / o (head h at column h*64): box = 64 columns x box_rows rows x 1 clip
int seq_map(CUtensorMap* out, const void* base, int B, int S, int cols, long ld, int box_rows = 128) {
  EncodeTiledFn fn = enc_fn();
  if (!fn) { exg_set_error("attn_tc: cuTensorMapEncodeTiled unavailable"); return EXG_EUNSUPPORTED; }
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)S, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)S * ld * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    exg_set_error("attn_tc: cuTensorMapEncodeTiled failed (%d) B=%d S=%d ld=%ld", (int)r, B, S, ld);
    return EXG_EINVAL;
  }
  return 0;
}

bool a16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace
