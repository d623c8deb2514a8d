// attention_tc.cu — fused attention on 5th-gen tensor cores for the short sequences of the TAN / ExoGround
// encoders (S = T or T+N <= 128 keys per clip; head_dim 64): one CTA per (batch, head, 128-query tile).
//
//   forward : TMA(Q,K,V) -> S = Q K^T (tcgen05, fp32 in TMEM) -> masked softmax in registers (thread = query
//             row, tcgen05.ld) -> P (bf16) to 128B-swizzled smem -> O = P V (tcgen05) -> TMA store, lse to global
//   backward: TMA(Q,K,V,dO) -> S = Q K^T, dP = dO V^T -> P = exp(S*scale - lse), dS = P (dP - delta) scale in
//             registers -> P, dS (bf16) to smem -> dV = P^T dO, dK = dS^T Q, dQ = dS K (tcgen05) -> TMA stores
//
// The score matrix never leaves the SM.  Sequences shorter than the 128-row MMA tile are handled by 3-D tensor
// maps [batch][seq][row stride]: rows past the sequence end are zero-filled on load and clipped on store, so a
// tile never touches the neighbouring clip.  The same P / dS smem tile serves as a K-major A operand (dQ = dS K,
// O = P V) and as an MN-major A operand (dV = P^T dO, dK = dS^T Q) - no transposes anywhere.
// Longer key sequences use the SIMT flash kernel in attention.cu.
// Replaces nn.MultiheadAttention's scaled-dot-product core (model/tfm_model.py:32,75,79) and its backward.
#include "attn_tc_common.cuh"

namespace {

// ------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(128)
attn_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, AtcArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* Qs = smem; uint8_t* Ks = smem + kTile; uint8_t* Vs = smem + 2 * kTile; uint8_t* Ps = smem + 3 * kTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 5 * kTile);      // [0] loads, [1] MMAs
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  uint32_t* maskw = tmem_slot + 1;                                     // [4]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * TQ;
  if (tid == 0) {
    mbar_init(&bars[0], 1); mbar_init(&bars[1], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
  }
  if (warp == 0) tmem_alloc(tmem_slot, 256);
  build_mask(maskw, p, b, tid);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + TK;

  if (tid == 0) {
    mbar_arrive_expect_tx(&bars[0], 3 * kTile);
    tma_load_3d(Qs, &tmQ, &bars[0], h * HD, q0, b);
    tma_load_3d(Ks, &tmK, &bars[0], h * HD, 0, b);
    tma_load_3d(Vs, &tmV, &bars[0], h * HD, 0, b);
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    mma_kmajor_k64(tS, smem_u32(Qs), smem_u32(Ks), TK);
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 0);
  tc_fence_after();

  // ---- masked softmax over the keys; thread = query row
  const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
  const uint32_t m0 = maskw[0], m1 = maskw[1], m2 = maskw[2], m3 = maskw[3];
  const int nchunk = (p.Sk + 31) / 32;
  const float sl2 = p.scale * kLog2e;
  float mx = -INFINITY;
  for (int c = 0; c < nchunk; ++c) {
    uint32_t r[32];
    tmem_ld_32x32(tS + lane_addr + c * 32, r);
    tmem_ld_wait();
    const uint32_t mw = c == 0 ? m0 : (c == 1 ? m1 : (c == 2 ? m2 : m3));
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if ((mw >> j) & 1) mx = fmaxf(mx, __uint_as_float(r[j]));
  }
  const float msafe = mx == -INFINITY ? 0.f : mx;
  float sum = 0.f;
  for (int c = 0; c < 4; ++c) {
    float pv[32];
    if (c < nchunk) {
      uint32_t r[32];
      tmem_ld_32x32(tS + lane_addr + c * 32, r);
      tmem_ld_wait();
      const uint32_t mw = c == 0 ? m0 : (c == 1 ? m1 : (c == 2 ? m2 : m3));
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float e = ((mw >> j) & 1) ? exp2f((__uint_as_float(r[j]) - msafe) * sl2) : 0.f;
        pv[j] = e; sum += e;
      }
    } else {
      if (c * 32 >= ((p.Sk + 15) / 16) * 16) break;      // columns no P V k-step will read
#pragma unroll
      for (int j = 0; j < 32; ++j) pv[j] = 0.f;
    }
    uint8_t* blk = Ps + (c >> 1) * kTile;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      *reinterpret_cast<uint4*>(blk + swz(tid, (c & 1) * 4 + u)) = pack8(&pv[u * 8]);
  }
  if (q0 + tid < p.Sq && p.lse)
    p.lse[((long)b * p.H + h) * p.Sq + q0 + tid] = mx * p.scale + __logf(sum);

  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    // O[128 x 64] = P[128 x Sk] V[Sk x 64]: A = P K-major (two 64-key blocks), B = V MN-major (keys x d)
    const uint32_t idesc = umma_idesc_bf16(128, HD, 0, 1);
    const int nk = (p.Sk + 15) / 16;
    for (int k = 0; k < nk; ++k) {
      const uint64_t ad = umma_smem_desc(smem_u32(Ps) + (k >> 2) * kTile + (k & 3) * 32, 16, 1024);
      const uint64_t bd = umma_smem_desc(smem_u32(Vs) + k * 2048, 8192, 1024);
      umma_bf16(tO, ad, bd, idesc, k > 0);
    }
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 1);
  tc_fence_after();

  // ---- O / l -> bf16 -> swizzled staging (the Q tile is free now) -> TMA store (rows >= Sq are clipped)
  {
    const float inv = 1.0f / sum;
    uint32_t r[64];
    tmem_ld_32x32(tO + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
    tmem_ld_32x32(tO + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
    tmem_ld_wait();
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[u * 8 + j]) * inv;
      *reinterpret_cast<uint4*>(Qs + swz(tid, u)) = pack8(f);
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tma_store_3d(&tmO, Qs, h * HD, q0, b);
    bulk_commit();
    bulk_wait_all();
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------ backward
__global__ void __launch_bounds__(128)
attn_tc_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                   const __grid_constant__ CUtensorMap tmdQ, const __grid_constant__ CUtensorMap tmdK,
                   const __grid_constant__ CUtensorMap tmdV, const void* __restrict__ o, long ldo, AtcArgs p,
                   int ps_tiles) {
  // smem: [P][dS][Q][K][V][dO].  P and dS are 128 x 128 (two 64-key blocks) in general; with <= 64 keys only
  // block 0 is ever written and ps_tiles = 1: the MN-major descriptors of the dV / dK MMAs then read the NEXT tile
  // as "keys 64..127" (dS for P, Q for dS - valid memory, finite or not), which only produces accumulator rows
  // >= 64 that the clipped TMA stores never write.  96 KB instead of 128 KB: two CTAs per SM.
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* Ps = smem; uint8_t* dSs = smem + ps_tiles * kTile;
  uint8_t* Qs = smem + 2 * ps_tiles * kTile; uint8_t* Ks = Qs + kTile; uint8_t* Vs = Qs + 2 * kTile;
  uint8_t* dOs = Qs + 3 * kTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(Qs + 4 * kTile);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  uint32_t* maskw = tmem_slot + 1;

  const int tid = threadIdx.x, warp = tid >> 5;
  const int b = blockIdx.z, h = blockIdx.y;
  if (tid == 0) {
    mbar_init(&bars[0], 1); mbar_init(&bars[1], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmdO);
  }
  if (warp == 0) tmem_alloc(tmem_slot, 256);
  build_mask(maskw, p, b, tid);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + TK;                       // phase 1
  const uint32_t tdV = tmem, tdK = tmem + HD, tdQ = tmem + 2 * HD; // phase 2 (S / dP are dead by then)

  if (tid == 0) {
    mbar_arrive_expect_tx(&bars[0], 4 * kTile);
    tma_load_3d(Qs, &tmQ, &bars[0], h * HD, 0, b);
    tma_load_3d(Ks, &tmK, &bars[0], h * HD, 0, b);
    tma_load_3d(Vs, &tmV, &bars[0], h * HD, 0, b);
    tma_load_3d(dOs, &tmdO, &bars[0], h * HD, 0, b);
  }
  // per-row softmax statistics while the tiles are in flight
  const bool row_ok = tid < p.Sq;
  const float lse_l2 = row_ok ? p.lse[((long)b * p.H + h) * p.Sq + tid] * kLog2e : 0.f;
  mbar_wait(&bars[0], 0);
  if (tid == 0) {
    tc_fence_after();
    mma_kmajor_k64(tS, smem_u32(Qs), smem_u32(Ks), TK);      // S  = Q  K^T
    mma_kmajor_k64(tdP, smem_u32(dOs), smem_u32(Vs), TK);    // dP = dO V^T
    umma_commit(&bars[1]);
  }
  // delta_r = <dO_r, O_r>  (dO row from the swizzled smem tile, O row straight from global: 128 contiguous bytes)
  float delta = 0.f;
  if (row_ok) {
    const uint4* orow = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(o) +
                                                       ((long)b * p.Sq + tid) * ldo + h * HD);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const uint4 a = *reinterpret_cast<const uint4*>(dOs + swz(tid, u));
      const uint4 c = __ldg(orow + u);
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, cw[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        delta += __uint_as_float(aw[j] << 16) * __uint_as_float(cw[j] << 16);
        delta += __uint_as_float(aw[j] & 0xffff0000u) * __uint_as_float(cw[j] & 0xffff0000u);
      }
    }
  }
  mbar_wait(&bars[1], 0);
  tc_fence_after();

  const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
  const uint32_t m0 = maskw[0], m1 = maskw[1], m2 = maskw[2], m3 = maskw[3];
  const int nchunk = (p.Sk + 31) / 32;
  const float sl2 = p.scale * kLog2e;
  for (int c = 0; c < nchunk; ++c) {
    uint32_t rs[32], rd[32];
    tmem_ld_32x32(tS + lane_addr + c * 32, rs);
    tmem_ld_32x32(tdP + lane_addr + c * 32, rd);
    tmem_ld_wait();
    const uint32_t mw = row_ok ? (c == 0 ? m0 : (c == 1 ? m1 : (c == 2 ? m2 : m3))) : 0u;
    float pv[32], ds[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float pr = ((mw >> j) & 1) ? exp2f(__uint_as_float(rs[j]) * sl2 - lse_l2) : 0.f;
      pv[j] = pr;
      ds[j] = pr * (__uint_as_float(rd[j]) - delta) * p.scale;
    }
    uint8_t* pb = Ps + (c >> 1) * kTile; uint8_t* db = dSs + (c >> 1) * kTile;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      *reinterpret_cast<uint4*>(pb + swz(tid, (c & 1) * 4 + u)) = pack8(&pv[u * 8]);
      *reinterpret_cast<uint4*>(db + swz(tid, (c & 1) * 4 + u)) = pack8(&ds[u * 8]);
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const int nkq = (min(p.Sq, TQ) + 15) / 16, nkk = (p.Sk + 15) / 16;
    // dV[key x d] = P^T dO ; dK = dS^T Q : A = P / dS read MN-major (keys contiguous, 2 blocks of 64), K-dim = queries
    const uint32_t id_t = umma_idesc_bf16(128, HD, 1, 1);
    for (int k = 0; k < nkq; ++k)
      umma_bf16(tdV, umma_smem_desc(smem_u32(Ps) + k * 2048, kTile, 1024),
                umma_smem_desc(smem_u32(dOs) + k * 2048, 8192, 1024), id_t, k > 0);
    for (int k = 0; k < nkq; ++k)
      umma_bf16(tdK, umma_smem_desc(smem_u32(dSs) + k * 2048, kTile, 1024),
                umma_smem_desc(smem_u32(Qs) + k * 2048, 8192, 1024), id_t, k > 0);
    // dQ[query x d] = dS K : A = dS K-major, B = K MN-major (keys x d)
    const uint32_t id_q = umma_idesc_bf16(128, HD, 0, 1);
    for (int k = 0; k < nkk; ++k)
      umma_bf16(tdQ, umma_smem_desc(smem_u32(dSs) + (k >> 2) * kTile + (k & 3) * 32, 16, 1024),
                umma_smem_desc(smem_u32(Ks) + k * 2048, 8192, 1024), id_q, k > 0);
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 1);
  tc_fence_after();
  // ---- gradients -> bf16 -> staging (input tiles are dead) -> TMA stores (rows past the sequence end are clipped)
  {
    const uint32_t src[3] = {tdV, tdK, tdQ};
    uint8_t* dst[3] = {Vs, Ks, Qs};
#pragma unroll
    for (int g = 0; g < 3; ++g) {
      uint32_t r[64];
      tmem_ld_32x32(src[g] + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
      tmem_ld_32x32(src[g] + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
      tmem_ld_wait();
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float f[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[u * 8 + j]);
        *reinterpret_cast<uint4*>(dst[g] + swz(tid, u)) = pack8(f);
      }
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tma_store_3d(&tmdV, Vs, h * HD, 0, b);
    tma_store_3d(&tmdK, Ks, h * HD, 0, b);
    tma_store_3d(&tmdQ, Qs, h * HD, 0, b);
    bulk_commit();
    bulk_wait_all();
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------ host side

constexpr size_t kFwdSmem = 5 * kTile + 1024 + 64;
constexpr size_t bwd_smem(int ps_tiles) { return (size_t)(4 + 2 * ps_tiles) * kTile + 1024 + 64; }
constexpr size_t kBwdSmem = bwd_smem(2);

}  // namespace

// 1 if the tensor-core attention kernels take this problem (bf16, <= 128 keys, TMA-compatible strides).
int exg_attn_tc_ok(int dtype, long ldq, long ldk, long ldv, long ldo, int Sq, int Sk, int d, bool bwd,
                   const void* q, const void* k, const void* v, const void* o) {
  static const int off = getenv("EXG_NO_ATTN_TC") ? atoi(getenv("EXG_NO_ATTN_TC")) : 0;
  if (off || dtype != EXG_BF16 || d != 64 || Sk > TK || Sk < 1) return 0;
  if (bwd && Sq > TQ) return 0;
  if (ldq % 8 || ldk % 8 || ldv % 8 || ldo % 8) return 0;
  if (!a16(q) || !a16(k) || !a16(v) || !a16(o)) return 0;
  return 1;
}

int exg_attn_tc_fwd(const void* q, const void* k, const void* v, long ldq, long ldk, long ldv,
                    const uint8_t* key_pad, void* o, long ldo, float* lse, int B, int H, int Sq, int Sk,
                    float scale, void* stream) {
  CUtensorMap mq, mk, mv, mo;
  int rc;
  if ((rc = seq_map(&mq, q, B, Sq, H * HD, ldq))) return rc;
  if ((rc = seq_map(&mk, k, B, Sk, H * HD, ldk))) return rc;
  if ((rc = seq_map(&mv, v, B, Sk, H * HD, ldv))) return rc;
  if ((rc = seq_map(&mo, o, B, Sq, H * HD, ldo))) return rc;
  static bool once = false;
  if (!once) {
    cudaError_t e = cudaFuncSetAttribute(attn_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kFwdSmem);
    if (e != cudaSuccess) { exg_set_error("attn_tc_fwd: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    once = true;
  }
  AtcArgs a = {B, H, Sq, Sk, scale, key_pad, lse};
  dim3 grid((Sq + TQ - 1) / TQ, H, B);
  attn_tc_fwd_kernel<<<grid, 128, kFwdSmem, (cudaStream_t)stream>>>(mq, mk, mv, mo, a);
  EXG_LAUNCH_CHECK("attn_tc_fwd");
  return 0;
}

int exg_attn_tc_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, long ldq,
                    long ldk, long ldv, long ldo, const uint8_t* key_pad, const float* lse, void* dq, void* dk,
                    void* dv, int B, int H, int Sq, int Sk, float scale, void* stream) {
  CUtensorMap mq, mk, mv, mdo, mdq, mdk, mdv;
  int rc;
  if ((rc = seq_map(&mq, q, B, Sq, H * HD, ldq))) return rc;
  if ((rc = seq_map(&mk, k, B, Sk, H * HD, ldk))) return rc;
  if ((rc = seq_map(&mv, v, B, Sk, H * HD, ldv))) return rc;
  if ((rc = seq_map(&mdo, dout, B, Sq, H * HD, ldo))) return rc;
  if ((rc = seq_map(&mdq, dq, B, Sq, H * HD, ldq))) return rc;
  if ((rc = seq_map(&mdk, dk, B, Sk, H * HD, ldk))) return rc;
  if ((rc = seq_map(&mdv, dv, B, Sk, H * HD, ldv))) return rc;
  static bool once = false;
  if (!once) {
    cudaError_t e = cudaFuncSetAttribute(attn_tc_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kBwdSmem);
    if (e != cudaSuccess) { exg_set_error("attn_tc_bwd: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    once = true;
  }
  AtcArgs a = {B, H, Sq, Sk, scale, key_pad, const_cast<float*>(lse)};
  dim3 grid(1, H, B);
  const int ps_tiles = Sk <= 64 ? 1 : 2;
  attn_tc_bwd_kernel<<<grid, 128, bwd_smem(ps_tiles), (cudaStream_t)stream>>>(mq, mk, mv, mdo, mdq, mdk, mdv, o, ldo, a,
                                                                            ps_tiles);
  EXG_LAUNCH_CHECK("attn_tc_bwd");
  return 0;
}
