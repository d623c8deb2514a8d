// attention_flash.cu — fused attention on tcgen05 for key sequences LONGER than one 128-key tile (ExoGround
// T = 1024 video tokens, the T = 64 -> 4096 inference sweep of eval/eval_zeroshot_align.py): flash-style online
// softmax, scores never leave the SM.  head_dim 64, bf16 in / out, fp32 accumulation and statistics.
//
//   forward : CTA = (clip, head, 128-query tile), 4 softmax warps (thread = query row) + 1 producer warp.
//             producer: TMA(Q), then per 128-key tile j: TMA(K_j) double-buffered, TMA(V_j), S_j = Q K_j^T
//             (tcgen05 -> TMEM), O_j = P_j V_j (tcgen05 -> TMEM).  softmax warps: running max / sum in registers,
//             P_j (bf16) -> 128B-swizzled smem, O accumulated in REGISTERS (O = O * alpha + O_j, tcgen05.ld of a
//             64-column tile) so no TMEM accumulator ever needs rescaling.  192 TMEM columns, 96 KB smem:
//             two CTAs per SM, one CTA's MMAs run under the other's exponentials.
//   backward: CTA = (clip, head, 128-key tile); K_j, V_j resident, loop over the query tiles i: S = Q_i K_j^T,
//             dP = dO_i V_j^T, P = exp(S*scale - lse), dS = P (dP - delta) scale -> smem, then dV_j += P^T dO_i,
//             dK_j += dS^T Q_i (TMEM accumulators across the whole loop) and dQ_i = dS K_j, which is added into
//             an fp32 workspace with a TMA reduce-add (the only cross-CTA accumulation), converted to bf16 by a
//             small kernel at the end.
// Replaces nn.MultiheadAttention's scaled-dot-product core (model/tfm_model.py:32,75,79) for S > 128.
#include "attn_tc_common.cuh"

namespace {

constexpr int FT = 160;                       // 4 softmax warps + 1 producer warp
constexpr int kMaxMaskWords = 1024;           // up to 32768 keys

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void softmax_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
constexpr float kLazyLog2 = 20.f;             // forward: largest 2^x a score may exceed the stale running maximum by
// CTA-wide OR over the 128 softmax threads (named barrier 2, so it cannot pair with a bar.sync 1 of another warp)
__device__ __forceinline__ bool softmax_bar_or(bool v) {
  uint32_t r;
  asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %1, 0;\n\tbar.red.or.pred q, 2, 128, p;\n\tselp.u32 %0, 1, 0, q;\n\t}"
               : "=r"(r) : "r"((uint32_t)v) : "memory");
  return r != 0;
}
constexpr int kBwdSoftmaxWarps = 8;           // backward: two warps per TMEM lane quarter
constexpr int FTB = 32 * (kBwdSoftmaxWarps + 1);
__device__ __forceinline__ void bwd_softmax_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// key-validity bit words for ALL keys of clip b (bit k & 31 of word k >> 5), rounded up to whole 128-key tiles
__device__ __forceinline__ void build_mask_all(uint32_t* maskw, const AtcArgs& p, int b, int nwords, int warp, int lane,
                                               int nwarps) {
  for (int w = warp; w < nwords; w += nwarps) {
    const int key = w * 32 + lane;
    const bool ok = key < p.Sk && !(p.key_pad && p.key_pad[(long)b * p.Sk + key]);
    const uint32_t bits = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) maskw[w] = bits;
  }
}

// O[128 x 64] = P[128 x 128 keys] V[128 keys x 64]: A = P K-major (two 64-key blocks), B = V MN-major (keys x d)
__device__ __forceinline__ void mma_pv(uint32_t tO, uint32_t sP, uint32_t sV, uint32_t acc) {
  const uint32_t idesc = umma_idesc_bf16(128, HD, 0, 1);
#pragma unroll
  for (int k = 0; k < 8; ++k)
    umma_bf16(tO, umma_smem_desc(sP + (k >> 2) * kTile + (k & 3) * 32, 16, 1024),
              umma_smem_desc(sV + k * 2048, 8192, 1024), idesc, acc | (uint32_t)(k > 0));
}

// ------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(FT, 2)
attn_flash_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, AtcArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* Qs = smem; uint8_t* Ks = smem + kTile; uint8_t* Vs = smem + 3 * kTile; uint8_t* Ps = smem + 4 * kTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * kTile);
  uint64_t *bar_q = bars, *bar_k = bars + 1 /* [2] */, *bar_v = bars + 3, *bar_s = bars + 4, *bar_p = bars + 5,
           *bar_o = bars + 6;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  uint32_t* pv_acc = tmem_slot + 1;               // 1: this tile's P V accumulates onto O in TMEM, 0: it restarts O
  uint32_t* maskw = tmem_slot + 4;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * TQ;
  const int nt = (p.Sk + TK - 1) / TK;
  if (tid == 0) {
    for (int i = 0; i < 7; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
  }
  if (warp == 4) tmem_alloc(tmem_slot, 256);
  build_mask_all(maskw, p, b, nt * 4, warp, lane, FT / 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + TK;

  if (warp == 4) {
    if (lane == 0) {
      mbar_arrive_expect_tx(bar_q, kTile);
      tma_load_3d(Qs, &tmQ, bar_q, h * HD, q0, b);
      mbar_arrive_expect_tx(&bar_k[0], kTile);
      tma_load_3d(Ks, &tmK, &bar_k[0], h * HD, 0, b);
      mbar_arrive_expect_tx(bar_v, kTile);
      tma_load_3d(Vs, &tmV, bar_v, h * HD, 0, b);
      if (nt > 1) {
        mbar_arrive_expect_tx(&bar_k[1], kTile);
        tma_load_3d(Ks + kTile, &tmK, &bar_k[1], h * HD, TK, b);
      }
      mbar_wait(bar_q, 0);
      mbar_wait(&bar_k[0], 0);
      tc_fence_after();
      mma_kmajor_k64(tS, smem_u32(Qs), smem_u32(Ks), TK);
      umma_commit(bar_s);
      for (int j = 0; j < nt; ++j) {
        mbar_wait(bar_p, j & 1);                 // P_j in smem; S_j consumed
        const uint32_t acc = *reinterpret_cast<volatile uint32_t*>(pv_acc);     // 0: exact tile, O restarts
        mbar_wait(bar_v, j & 1);
        tc_fence_after();
        mma_pv(tO, smem_u32(Ps), smem_u32(Vs), acc);
        umma_commit(bar_o);
        if (j + 1 < nt) {
          mbar_wait(&bar_k[(j + 1) & 1], ((j + 1) >> 1) & 1);
          tc_fence_after();
          mma_kmajor_k64(tS, smem_u32(Qs), smem_u32(Ks + ((j + 1) & 1) * kTile), TK);
          umma_commit(bar_s);
        }
        if (j + 2 < nt) {                        // K buffer j & 1 is free: S_j completed before P_j was written
          mbar_arrive_expect_tx(&bar_k[j & 1], kTile);
          tma_load_3d(Ks + (j & 1) * kTile, &tmK, &bar_k[j & 1], h * HD, (j + 2) * TK, b);
        }
        if (j + 1 < nt) {                        // the single V buffer is free once P_j V_j has completed
          mbar_wait(bar_o, j & 1);
          mbar_arrive_expect_tx(bar_v, kTile);
          tma_load_3d(Vs, &tmV, bar_v, h * HD, (j + 1) * TK, b);
        }
      }
    }
    __syncwarp();
  } else {
    const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
    const float sl2 = p.scale * kLog2e;
    float m = -INFINITY, l = 0.f;
    float o[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) o[i] = 0.f;
    // TMEM -> register reads are what bounds this kernel (64 B / clk / SM: one 128 x 128 fp32 score tile = 1024
    // cycles, twice the time of its two MMAs), so the common tile reads S exactly ONCE and never reads O:
    //   * the exponentials are taken relative to the running maximum of the PREVIOUS tiles (m); the tile maximum is
    //     tracked alongside.  Softmax is shift-invariant, so a stale reference only matters for range: as long as no
    //     row's tile maximum exceeds m by more than 2^kLazyLog2 the tile is done - P V accumulates onto O in TMEM.
    //   * otherwise (the first tile, and whenever some row's maximum jumps) the whole CTA takes the exact path for
    //     this tile: new maximum, O so far folded into registers with the rescale factor, second read of S, and the
    //     TMEM accumulation restarts (pv_acc = 0).  O = (registers + TMEM) / l at the end either way.
    for (int j = 0; j < nt; ++j) {
      mbar_wait(bar_s, j & 1);
      tc_fence_after();
      const uint32_t* mw = maskw + j * 4;
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      float sm4[4] = {0.f, 0.f, 0.f, 0.f};
      bool exact = j == 0;
      if (!exact) {
        const float nb = -(m == -INFINITY ? 0.f : m) * sl2;
        uint32_t ra[32], rb[32];
        tmem_ld_32x32(tS + lane_addr, ra);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t (&cur)[32] = (c & 1) ? rb : ra;
          uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
          tmem_ld_wait();
          if (c + 1 < 4) tmem_ld_32x32(tS + lane_addr + (c + 1) * 32, nxt);
          const uint32_t w = mw[c];
          float pv[32];
          if (w == 0xffffffffu) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float sv = __uint_as_float(cur[i]);
              mx4[i & 3] = fmaxf(mx4[i & 3], sv);
              pv[i] = ex2_approx(fmaf(sv, sl2, nb));
              sm4[i & 3] += pv[i];
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float sv = __uint_as_float(cur[i]);
              const bool on = (w >> i) & 1;
              if (on) mx4[i & 3] = fmaxf(mx4[i & 3], sv);
              pv[i] = on ? ex2_approx(fmaf(sv, sl2, nb)) : 0.f;
              sm4[i & 3] += pv[i];
            }
          }
          uint8_t* blk = Ps + (c >> 1) * kTile;
#pragma unroll
          for (int u = 0; u < 4; ++u)
            *reinterpret_cast<uint4*>(blk + swz(tid, (c & 1) * 4 + u)) = pack8(&pv[u * 8]);
        }
        const float mt = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        // (m = -inf with a finite tile maximum: the difference is +inf, i.e. exact path)
        const bool jump = mt != -INFINITY && (mt - m) * sl2 > kLazyLog2;
        exact = softmax_bar_or(jump);                // CTA-uniform
        if (!exact) l += (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]);
      }
      if (exact) {
        if (j == 0) {
          // tile maximum (later tiles know it from the single pass).  TMEM loads double-buffered; 4 independent
          // running maxima instead of one 32-deep dependency chain; unmasked fast path
          uint32_t ra[32], rb[32];
          tmem_ld_32x32(tS + lane_addr, ra);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t (&cur)[32] = (c & 1) ? rb : ra;
            uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
            tmem_ld_wait();
            if (c + 1 < 4) tmem_ld_32x32(tS + lane_addr + (c + 1) * 32, nxt);
            const uint32_t w = mw[c];
            if (w == 0xffffffffu) {
#pragma unroll
              for (int i = 0; i < 32; ++i) mx4[i & 3] = fmaxf(mx4[i & 3], __uint_as_float(cur[i]));
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if ((w >> i) & 1) mx4[i & 3] = fmaxf(mx4[i & 3], __uint_as_float(cur[i]));
            }
          }
        }
        const float mt = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const float m_new = fmaxf(m, mt);
        const float msafe = m_new == -INFINITY ? 0.f : m_new;
        const float alpha = ex2_approx((m - msafe) * sl2);        // m = -inf -> 0 (o and l are 0 then anyway)
        m = m_new;
        if (j > 0) {                                              // fold the O accumulated in TMEM since the last exact tile
          mbar_wait(bar_o, (j - 1) & 1);
          tc_fence_after();
          uint32_t r[64];
          tmem_ld_32x32(tO + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
          tmem_ld_32x32(tO + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < HD; ++i) o[i] = (o[i] + __uint_as_float(r[i])) * alpha;
        }
        const float nb = -msafe * sl2;
        sm4[0] = sm4[1] = sm4[2] = sm4[3] = 0.f;
        {
          uint32_t ra[32], rb[32];
          tmem_ld_32x32(tS + lane_addr, ra);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t (&cur)[32] = (c & 1) ? rb : ra;
            uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
            tmem_ld_wait();
            if (c + 1 < 4) tmem_ld_32x32(tS + lane_addr + (c + 1) * 32, nxt);
            const uint32_t w = mw[c];
            float pv[32];
            if (w == 0xffffffffu) {
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                pv[i] = ex2_approx(fmaf(__uint_as_float(cur[i]), sl2, nb));
                sm4[i & 3] += pv[i];
              }
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                pv[i] = ((w >> i) & 1) ? ex2_approx(fmaf(__uint_as_float(cur[i]), sl2, nb)) : 0.f;
                sm4[i & 3] += pv[i];
              }
            }
            uint8_t* blk = Ps + (c >> 1) * kTile;
#pragma unroll
            for (int u = 0; u < 4; ++u)
              *reinterpret_cast<uint4*>(blk + swz(tid, (c & 1) * 4 + u)) = pack8(&pv[u * 8]);
          }
        }
        l = l * alpha + (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]);
      }
      fence_proxy_async();
      tc_fence_before();
      softmax_bar_sync();
      if (tid == 0) {
        *pv_acc = exact ? 0u : 1u;                   // read by the producer after its acquire of bar_p
        mbar_arrive(bar_p);
      }
    }
    // last tile's O, normalise, stage (the Q tile is dead: every S MMA has completed), store
    mbar_wait(bar_o, (nt - 1) & 1);
    tc_fence_after();
    {
      uint32_t r[64];
      tmem_ld_32x32(tO + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
      tmem_ld_32x32(tO + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
      tmem_ld_wait();
      const float inv = 1.0f / l;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = (o[u * 8 + i] + __uint_as_float(r[u * 8 + i])) * inv;
        *reinterpret_cast<uint4*>(Qs + swz(tid, u)) = pack8(f);
      }
    }
    if (q0 + tid < p.Sq && p.lse) p.lse[((long)b * p.H + h) * p.Sq + q0 + tid] = m * p.scale + __logf(l);
    fence_proxy_async();
    tc_fence_before();
    softmax_bar_sync();
    if (tid == 0) {
      tma_store_3d(&tmO, Qs, h * HD, q0, b);
      bulk_commit();
      bulk_wait_all();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------ forward, version 2
// The version above serialises, per 128-key tile and CTA, softmax -> barrier -> P.V + the NEXT Q.K^T -> barrier: a
// quarter of all stall samples sit in the softmax warps' wait for S (profiles/ncu_full_r02o_flash_summary.txt).  Here
// the key tile is 64 keys, S and P are double-buffered (TMEM: S[2] x 64 columns + O 64 = 192; smem: Q 16 KB, K and V
// 3 x 8 KB each, P 2 x 16 KB = 96 KB, still two CTAs per SM), and the producer issues S_{j+2} = Q K_{j+2}^T as soon as
// the softmax warps have consumed S_j - i.e. while they work on S_{j+1} - so a score tile is always waiting for them.
// Softmax: as above (one read of S per tile, exponentials relative to the running maximum of the previous tiles, O
// accumulated in TMEM, exact re-base when some row's maximum jumps by more than 2^kLazyLog2).
constexpr int KT2 = 64;                        // keys per tile
constexpr uint32_t kKvTile = KT2 * HD * 2;     // 8 KB
constexpr int kKvStages = 3;

__global__ void __launch_bounds__(FT, 2)
attn_flash_fwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, AtcArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* Qs = smem; uint8_t* Ks = smem + kTile; uint8_t* Vs = Ks + kKvStages * kKvTile;
  uint8_t* Ps = Vs + kKvStages * kKvTile;        // [2][128 rows][64 keys]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * kTile);
  uint64_t *bar_q = bars, *bar_k = bars + 1 /* [3] */, *bar_v = bars + 4 /* [3] */, *bar_s = bars + 7 /* [2] */,
           *bar_p = bars + 9 /* [2] */, *bar_pv = bars + 11 /* [2] */;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);
  uint32_t* pv_acc = tmem_slot + 1;              // [2]: 1 = P_j V_j accumulates onto O in TMEM, 0 = it restarts O
  uint32_t* maskw = tmem_slot + 4;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * TQ;
  const int nt = (p.Sk + KT2 - 1) / KT2;
  if (tid == 0) {
    for (int i = 0; i < 13; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
  }
  if (warp == 4) tmem_alloc(tmem_slot, 256);
  build_mask_all(maskw, p, b, nt * 2, warp, lane, FT / 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem /* + 64 * buffer */, tO = tmem + 2 * KT2;

  if (warp == 4) {
    if (lane == 0) {
      mbar_arrive_expect_tx(bar_q, kTile);
      tma_load_3d(Qs, &tmQ, bar_q, h * HD, q0, b);
      for (int j = 0; j < kKvStages && j < nt; ++j) {
        mbar_arrive_expect_tx(&bar_k[j], kKvTile);
        tma_load_3d(Ks + j * kKvTile, &tmK, &bar_k[j], h * HD, j * KT2, b);
        mbar_arrive_expect_tx(&bar_v[j], kKvTile);
        tma_load_3d(Vs + j * kKvTile, &tmV, &bar_v[j], h * HD, j * KT2, b);
      }
      mbar_wait(bar_q, 0);
      for (int j = 0; j < 2 && j < nt; ++j) {    // S_0, S_1: both score buffers start free
        mbar_wait(&bar_k[j], 0);
        tc_fence_after();
        mma_kmajor_k64(tS + j * KT2, smem_u32(Qs), smem_u32(Ks + j * kKvTile), KT2);
        umma_commit(&bar_s[j]);
      }
      const uint32_t idesc_pv = umma_idesc_bf16(128, HD, 0, 1);
      for (int j = 0; j < nt; ++j) {
        const int bf = j & 1, st = j % kKvStages;
        mbar_wait(&bar_p[bf], (j >> 1) & 1);     // P_j in smem, S_j consumed
        const uint32_t acc = *reinterpret_cast<volatile uint32_t*>(&pv_acc[bf]);
        mbar_wait(&bar_v[st], (j / kKvStages) & 1);
        tc_fence_after();
        // O (+)= P_j[128 x 64 keys] V_j[64 keys x 64]: A K-major, B MN-major
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tO, umma_smem_desc(smem_u32(Ps) + bf * kTile + k * 32, 16, 1024),
                    umma_smem_desc(smem_u32(Vs) + st * kKvTile + k * 2048, 8192, 1024), idesc_pv, acc | (uint32_t)(k > 0));
        umma_commit(&bar_pv[bf]);
        if (j + 2 < nt) {                        // the score buffer S_j lived in takes S_{j+2}
          const int s2 = (j + 2) % kKvStages;
          mbar_wait(&bar_k[s2], ((j + 2) / kKvStages) & 1);
          tc_fence_after();
          mma_kmajor_k64(tS + bf * KT2, smem_u32(Qs), smem_u32(Ks + s2 * kKvTile), KT2);
          umma_commit(&bar_s[bf]);
        }
        if (j + kKvStages < nt) {                // K stage of tile j: S_j completed before the softmax warps read it
          mbar_arrive_expect_tx(&bar_k[st], kKvTile);
          tma_load_3d(Ks + st * kKvTile, &tmK, &bar_k[st], h * HD, (j + kKvStages) * KT2, b);
        }
        if (j >= 1 && j + 2 < nt) {              // V stage of tile j-1 is free once P_{j-1} V_{j-1} has completed
          mbar_wait(&bar_pv[(j - 1) & 1], ((j - 1) >> 1) & 1);
          const int sv = (j + 2) % kKvStages;    // == (j - 1) % 3
          mbar_arrive_expect_tx(&bar_v[sv], kKvTile);
          tma_load_3d(Vs + sv * kKvTile, &tmV, &bar_v[sv], h * HD, (j + 2) * KT2, b);
        }
      }
    }
    __syncwarp();
  } else {
    const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
    const float sl2 = p.scale * kLog2e;
    float m = -INFINITY, l = 0.f;
    float o[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) o[i] = 0.f;
    for (int j = 0; j < nt; ++j) {
      const int bf = j & 1;
      const uint32_t tSj = tS + bf * KT2;
      uint8_t* Pj = Ps + bf * kTile;
      mbar_wait(&bar_s[bf], (j >> 1) & 1);
      if (j >= 2) mbar_wait(&bar_pv[bf], ((j - 2) >> 1) & 1);      // P buffer: P_{j-2} V_{j-2} has read it
      tc_fence_after();
      const uint32_t mw0 = maskw[j * 2], mw1 = maskw[j * 2 + 1];
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      float sm4[4] = {0.f, 0.f, 0.f, 0.f};
      bool exact = j == 0;
      uint32_t ra[32], rb[32];
      if (!exact) {
        const float nb = -(m == -INFINITY ? 0.f : m) * sl2;
        tmem_ld_32x32(tSj + lane_addr, ra);
        tmem_ld_32x32(tSj + lane_addr + 32, rb);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t (&cur)[32] = c ? rb : ra;
          if (c == 0) tmem_ld_wait();            // (both loads were issued together; one wait covers them)
          const uint32_t w = c ? mw1 : mw0;
          float pv[32];
          if (w == 0xffffffffu) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float sv = __uint_as_float(cur[i]);
              mx4[i & 3] = fmaxf(mx4[i & 3], sv);
              pv[i] = ex2_approx(fmaf(sv, sl2, nb));
              sm4[i & 3] += pv[i];
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float sv = __uint_as_float(cur[i]);
              const bool on = (w >> i) & 1;
              if (on) mx4[i & 3] = fmaxf(mx4[i & 3], sv);
              pv[i] = on ? ex2_approx(fmaf(sv, sl2, nb)) : 0.f;
              sm4[i & 3] += pv[i];
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u)
            *reinterpret_cast<uint4*>(Pj + swz(tid, c * 4 + u)) = pack8(&pv[u * 8]);
        }
        const float mt = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const bool jump = mt != -INFINITY && (mt - m) * sl2 > kLazyLog2;
        exact = softmax_bar_or(jump);              // CTA-uniform
        if (!exact) l += (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]);
      }
      if (exact) {
        if (j == 0) {                              // tile maximum (later tiles know it from the single pass)
          tmem_ld_32x32(tSj + lane_addr, ra);
          tmem_ld_32x32(tSj + lane_addr + 32, rb);
          tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t (&cur)[32] = c ? rb : ra;
            const uint32_t w = c ? mw1 : mw0;
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if ((w >> i) & 1) mx4[i & 3] = fmaxf(mx4[i & 3], __uint_as_float(cur[i]));
          }
        }
        const float mt = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const float m_new = fmaxf(m, mt);
        const float msafe = m_new == -INFINITY ? 0.f : m_new;
        const float alpha = ex2_approx((m - msafe) * sl2);        // m = -inf -> 0 (o and l are 0 then anyway)
        m = m_new;
        if (j > 0) {                                              // fold the O accumulated in TMEM since the last exact tile
          mbar_wait(&bar_pv[(j - 1) & 1], ((j - 1) >> 1) & 1);
          tc_fence_after();
          uint32_t r[64];
          tmem_ld_32x32(tO + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
          tmem_ld_32x32(tO + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < HD; ++i) o[i] = (o[i] + __uint_as_float(r[i])) * alpha;
        }
        const float nb = -msafe * sl2;
        sm4[0] = sm4[1] = sm4[2] = sm4[3] = 0.f;
        if (j > 0) {                               // (j == 0: ra / rb still hold the tile)
          tmem_ld_32x32(tSj + lane_addr, ra);
          tmem_ld_32x32(tSj + lane_addr + 32, rb);
          tmem_ld_wait();
        }
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t (&cur)[32] = c ? rb : ra;
          const uint32_t w = c ? mw1 : mw0;
          float pv[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            pv[i] = ((w >> i) & 1) ? ex2_approx(fmaf(__uint_as_float(cur[i]), sl2, nb)) : 0.f;
            sm4[i & 3] += pv[i];
          }
#pragma unroll
          for (int u = 0; u < 4; ++u)
            *reinterpret_cast<uint4*>(Pj + swz(tid, c * 4 + u)) = pack8(&pv[u * 8]);
        }
        l = l * alpha + (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]);
      }
      fence_proxy_async();
      tc_fence_before();
      softmax_bar_sync();
      if (tid == 0) {
        pv_acc[bf] = exact ? 0u : 1u;              // read by the producer after its acquire of bar_p[bf]
        mbar_arrive(&bar_p[bf]);
      }
    }
    // last tile's O, normalise, stage (the Q tile is dead: every S MMA has completed), store
    mbar_wait(&bar_pv[(nt - 1) & 1], ((nt - 1) >> 1) & 1);
    tc_fence_after();
    {
      uint32_t r[64];
      tmem_ld_32x32(tO + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
      tmem_ld_32x32(tO + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
      tmem_ld_wait();
      const float inv = 1.0f / l;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = (o[u * 8 + i] + __uint_as_float(r[u * 8 + i])) * inv;
        *reinterpret_cast<uint4*>(Qs + swz(tid, u)) = pack8(f);
      }
    }
    if (q0 + tid < p.Sq && p.lse) p.lse[((long)b * p.H + h) * p.Sq + q0 + tid] = m * p.scale + __logf(l);
    fence_proxy_async();
    tc_fence_before();
    softmax_bar_sync();
    if (tid == 0) {
      tma_store_3d(&tmO, Qs, h * HD, q0, b);
      bulk_commit();
      bulk_wait_all();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------ forward, version 3
// Version 2 with 8 softmax warps (two per TMEM lane quarter): with 4 warps per CTA and two CTAs per SM every scheduler
// had two softmax warps to pick from, 31 % of the issue slots were used and the XU pipe (ex2) sat at 38 %.
constexpr int kFwd3Warps = 8;
constexpr int FT3 = 32 * (kFwd3Warps + 1);
__device__ __forceinline__ void fwd3_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ bool fwd3_bar_or(bool v) {
  uint32_t r;
  asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %1, 0;\n\tbar.red.or.pred q, 2, 256, p;\n\tselp.u32 %0, 1, 0, q;\n\t}"
               : "=r"(r) : "r"((uint32_t)v) : "memory");
  return r != 0;
}

__global__ void __launch_bounds__(FT3, 2)
attn_flash_fwd3_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, AtcArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* Qs = smem; uint8_t* Ks = smem + kTile; uint8_t* Vs = Ks + kKvStages * kKvTile;
  uint8_t* Ps = Vs + kKvStages * kKvTile;        // [2][128 rows][64 keys]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * kTile);
  uint64_t *bar_q = bars, *bar_k = bars + 1 /* [3] */, *bar_v = bars + 4 /* [3] */, *bar_s = bars + 7 /* [2] */,
           *bar_p = bars + 9 /* [2] */, *bar_pv = bars + 11 /* [2] */;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);
  uint32_t* pv_acc = tmem_slot + 1;              // [2]: 1 = P_j V_j accumulates onto O in TMEM, 0 = it restarts O
  float* xch = reinterpret_cast<float*>(tmem_slot + 4);      // [2][128]: row maxima / sums exchanged between the halves
  uint32_t* maskw = tmem_slot + 4 + 256;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * TQ;
  const int nt = (p.Sk + KT2 - 1) / KT2;
  if (tid == 0) {
    for (int i = 0; i < 13; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
  }
  if (warp == kFwd3Warps) tmem_alloc(tmem_slot, 256);
  build_mask_all(maskw, p, b, nt * 2, warp, lane, FT3 / 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem /* + 64 * buffer */, tO = tmem + 2 * KT2;

  if (warp == kFwd3Warps) {
    if (lane == 0) {
      mbar_arrive_expect_tx(bar_q, kTile);
      tma_load_3d(Qs, &tmQ, bar_q, h * HD, q0, b);
      for (int j = 0; j < kKvStages && j < nt; ++j) {
        mbar_arrive_expect_tx(&bar_k[j], kKvTile);
        tma_load_3d(Ks + j * kKvTile, &tmK, &bar_k[j], h * HD, j * KT2, b);
        mbar_arrive_expect_tx(&bar_v[j], kKvTile);
        tma_load_3d(Vs + j * kKvTile, &tmV, &bar_v[j], h * HD, j * KT2, b);
      }
      mbar_wait(bar_q, 0);
      for (int j = 0; j < 2 && j < nt; ++j) {    // S_0, S_1: both score buffers start free
        mbar_wait(&bar_k[j], 0);
        tc_fence_after();
        mma_kmajor_k64(tS + j * KT2, smem_u32(Qs), smem_u32(Ks + j * kKvTile), KT2);
        umma_commit(&bar_s[j]);
      }
      const uint32_t idesc_pv = umma_idesc_bf16(128, HD, 0, 1);
      for (int j = 0; j < nt; ++j) {
        const int bf = j & 1, st = j % kKvStages;
        mbar_wait(&bar_p[bf], (j >> 1) & 1);     // P_j in smem, S_j consumed
        const uint32_t acc = *reinterpret_cast<volatile uint32_t*>(&pv_acc[bf]);
        mbar_wait(&bar_v[st], (j / kKvStages) & 1);
        tc_fence_after();
        // O (+)= P_j[128 x 64 keys] V_j[64 keys x 64]: A K-major, B MN-major
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tO, umma_smem_desc(smem_u32(Ps) + bf * kTile + k * 32, 16, 1024),
                    umma_smem_desc(smem_u32(Vs) + st * kKvTile + k * 2048, 8192, 1024), idesc_pv, acc | (uint32_t)(k > 0));
        umma_commit(&bar_pv[bf]);
        if (j + 2 < nt) {                        // the score buffer S_j lived in takes S_{j+2}
          const int s2 = (j + 2) % kKvStages;
          mbar_wait(&bar_k[s2], ((j + 2) / kKvStages) & 1);
          tc_fence_after();
          mma_kmajor_k64(tS + bf * KT2, smem_u32(Qs), smem_u32(Ks + s2 * kKvTile), KT2);
          umma_commit(&bar_s[bf]);
        }
        if (j + kKvStages < nt) {                // K stage of tile j: S_j completed before the softmax warps read it
          mbar_arrive_expect_tx(&bar_k[st], kKvTile);
          tma_load_3d(Ks + st * kKvTile, &tmK, &bar_k[st], h * HD, (j + kKvStages) * KT2, b);
        }
        if (j >= 1 && j + 2 < nt) {              // V stage of tile j-1 is free once P_{j-1} V_{j-1} has completed
          mbar_wait(&bar_pv[(j - 1) & 1], ((j - 1) >> 1) & 1);
          const int sv = (j + 2) % kKvStages;    // == (j - 1) % 3
          mbar_arrive_expect_tx(&bar_v[sv], kKvTile);
          tma_load_3d(Vs + sv * kKvTile, &tmV, &bar_v[sv], h * HD, (j + 2) * KT2, b);
        }
      }
    }
    __syncwarp();
  } else {
    // 8 softmax warps: warps w and w + 4 share a TMEM lane quarter (query rows) and split every 64-key tile into two
    // 32-key halves, and the 64 output columns into two 32-column halves.  The lazy reference maximum is what makes
    // the halves independent: on an ordinary tile nothing crosses between them (partial row sums are added at the
    // end); only an exact tile exchanges the half maxima through shared memory so that both keep the same m.
    const int quarter = warp & 3, half = warp >> 2, row = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const float sl2 = p.scale * kLog2e;
    float m = -INFINITY, l = 0.f;                  // l: this half's share of the row sum
    float o[32];                                   // output columns [32 half, 32 half + 32)
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = 0.f;
    for (int j = 0; j < nt; ++j) {
      const int bf = j & 1;
      const uint32_t tSj = tS + bf * KT2 + half * 32;
      uint8_t* Pj = Ps + bf * kTile;
      mbar_wait(&bar_s[bf], (j >> 1) & 1);
      if (j >= 2) mbar_wait(&bar_pv[bf], ((j - 2) >> 1) & 1);      // P buffer: P_{j-2} V_{j-2} has read it
      tc_fence_after();
      const uint32_t w = maskw[j * 2 + half];
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      float sm4[4] = {0.f, 0.f, 0.f, 0.f};
      bool exact = j == 0;
      uint32_t ra[32];
      tmem_ld_32x32(tSj + lane_addr, ra);
      tmem_ld_wait();
      if (!exact) {
        const float nb = -(m == -INFINITY ? 0.f : m) * sl2;
        float pv[32];
        if (w == 0xffffffffu) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float sv = __uint_as_float(ra[i]);
            mx4[i & 3] = fmaxf(mx4[i & 3], sv);
            pv[i] = ex2_approx(fmaf(sv, sl2, nb));
            sm4[i & 3] += pv[i];
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float sv = __uint_as_float(ra[i]);
            const bool on = (w >> i) & 1;
            if (on) mx4[i & 3] = fmaxf(mx4[i & 3], sv);
            pv[i] = on ? ex2_approx(fmaf(sv, sl2, nb)) : 0.f;
            sm4[i & 3] += pv[i];
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          *reinterpret_cast<uint4*>(Pj + swz(row, half * 4 + u)) = pack8(&pv[u * 8]);
        const float mt = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const bool jump = mt != -INFINITY && (mt - m) * sl2 > kLazyLog2;
        exact = fwd3_bar_or(jump);                 // CTA-uniform over the 256 softmax threads
        if (!exact) l += (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if ((w >> i) & 1) mx4[i & 3] = fmaxf(mx4[i & 3], __uint_as_float(ra[i]));
      }
      if (exact) {
        // row maximum over both halves (the two warps of a row must keep the same reference)
        xch[half * 128 + row] = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        fwd3_bar_sync();
        const float mt = fmaxf(xch[row], xch[128 + row]);
        fwd3_bar_sync();                           // both halves have read the pair before anyone overwrites it
        const float m_new = fmaxf(m, mt);
        const float msafe = m_new == -INFINITY ? 0.f : m_new;
        const float alpha = ex2_approx((m - msafe) * sl2);        // m = -inf -> 0 (o and l are 0 then anyway)
        m = m_new;
        if (j > 0) {                                              // fold the O accumulated in TMEM since the last exact tile
          mbar_wait(&bar_pv[(j - 1) & 1], ((j - 1) >> 1) & 1);
          tc_fence_after();
          uint32_t r[32];
          tmem_ld_32x32(tO + lane_addr + half * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = (o[i] + __uint_as_float(r[i])) * alpha;
        }
        const float nb = -msafe * sl2;
        sm4[0] = sm4[1] = sm4[2] = sm4[3] = 0.f;
        float pv[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          pv[i] = ((w >> i) & 1) ? ex2_approx(fmaf(__uint_as_float(ra[i]), sl2, nb)) : 0.f;
          sm4[i & 3] += pv[i];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          *reinterpret_cast<uint4*>(Pj + swz(row, half * 4 + u)) = pack8(&pv[u * 8]);
        l = l * alpha + (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]);
      }
      fence_proxy_async();
      tc_fence_before();
      fwd3_bar_sync();
      if (tid == 0) {
        pv_acc[bf] = exact ? 0u : 1u;              // read by the producer after its acquire of bar_p[bf]
        mbar_arrive(&bar_p[bf]);
      }
    }
    // last tile's O, row sum over both halves, normalise, stage (the Q tile is dead), store
    xch[half * 128 + row] = l;
    mbar_wait(&bar_pv[(nt - 1) & 1], ((nt - 1) >> 1) & 1);
    tc_fence_after();
    fwd3_bar_sync();
    const float lsum = xch[row] + xch[128 + row];
    {
      uint32_t r[32];
      tmem_ld_32x32(tO + lane_addr + half * 32, r);
      tmem_ld_wait();
      const float inv = 1.0f / lsum;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = (o[u * 8 + i] + __uint_as_float(r[u * 8 + i])) * inv;
        *reinterpret_cast<uint4*>(Qs + swz(row, half * 4 + u)) = pack8(f);
      }
    }
    if (half == 0 && q0 + row < p.Sq && p.lse)
      p.lse[((long)b * p.H + h) * p.Sq + q0 + row] = m * p.scale + __logf(lsum);
    fence_proxy_async();
    tc_fence_before();
    fwd3_bar_sync();
    if (tid == 0) {
      tma_store_3d(&tmO, Qs, h * HD, q0, b);
      bulk_commit();
      bulk_wait_all();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kFwd3Warps) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------ backward
// delta[b, h, q] = <dO[b, q, h, :], O[b, q, h, :]>; one warp per (clip, query) row, 4 lanes per head.
__global__ void attn_delta_kernel(const uint16_t* __restrict__ o, const uint16_t* __restrict__ dout, long ldo,
                                  float* __restrict__ delta, int B, int H, int Sq) {
  const long row = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= (long)B * Sq) return;
  const long b = row / Sq, q = row % Sq;
  for (int c0 = 0; c0 < H * HD; c0 += 512) {
    const int col = c0 + lane * 16;
    float acc = 0.f;
    if (col < H * HD) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const uint4 a = *reinterpret_cast<const uint4*>(dout + row * ldo + col + u * 8);
        const uint4 c = *reinterpret_cast<const uint4*>(o + row * ldo + col + u * 8);
        const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, cw[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          acc += __uint_as_float(aw[j] << 16) * __uint_as_float(cw[j] << 16);
          acc += __uint_as_float(aw[j] & 0xffff0000u) * __uint_as_float(cw[j] & 0xffff0000u);
        }
      }
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if ((lane & 3) == 0 && col < H * HD) delta[(b * H + col / HD) * Sq + q] = acc;
  }
}

// dq (bf16, row stride ldq) = dq_acc (fp32, dense [B*Sq, H*64])
__global__ void attn_dq_convert_kernel(const float* __restrict__ acc, uint16_t* __restrict__ dq, long ldq, long rows,
                                       int cols) {
  const long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i >= rows * cols) return;
  const long r = i / cols; const int c = (int)(i % cols);
  const float4 a = *reinterpret_cast<const float4*>(acc + i), b2 = *reinterpret_cast<const float4*>(acc + i + 4);
  uint4 u;
  u.x = pack_bf16x2(a.x, a.y); u.y = pack_bf16x2(a.z, a.w); u.z = pack_bf16x2(b2.x, b2.y); u.w = pack_bf16x2(b2.z, b2.w);
  *reinterpret_cast<uint4*>(dq + r * ldq + c) = u;
}

struct FlashBwdArgs {
  AtcArgs a;
  const float* delta;
};

__global__ void __launch_bounds__(FTB, 1)
attn_flash_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                      const __grid_constant__ CUtensorMap tmdQacc, const __grid_constant__ CUtensorMap tmdK,
                      const __grid_constant__ CUtensorMap tmdV, FlashBwdArgs fa) {
  const AtcArgs& p = fa.a;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* Ks = smem; uint8_t* Vs = smem + kTile; uint8_t* Qs = smem + 2 * kTile /* [2] */;
  uint8_t* dOs = smem + 4 * kTile /* [2] */; uint8_t* Ps = smem + 6 * kTile /* 2 blocks */;
  uint8_t* dSs = smem + 8 * kTile /* 2 blocks */; float* stage = reinterpret_cast<float*>(smem + 10 * kTile);  // 32 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 12 * kTile);
  uint64_t *bar_kv = bars, *bar_qd = bars + 1 /* [2] */, *bar_s = bars + 3, *bar_p = bars + 4, *bar_g = bars + 5;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  uint32_t* maskw = tmem_slot + 4;                 // [4]: this key tile

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.z, h = blockIdx.y, k0 = blockIdx.x * TK;
  const int nq = (p.Sq + TQ - 1) / TQ;
  if (tid == 0) {
    for (int i = 0; i < 6; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmdO);
    tma_prefetch_desc(&tmdQacc); tma_prefetch_desc(&tmdK); tma_prefetch_desc(&tmdV);
  }
  if (warp == kBwdSoftmaxWarps) tmem_alloc(tmem_slot, 512);
  if (warp < 4) {
    const int key = k0 + tid;
    const bool ok = key < p.Sk && !(p.key_pad && p.key_pad[(long)b * p.Sk + key]);
    const uint32_t bits = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) maskw[warp] = bits;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + TK, tdV = tmem + 2 * TK, tdK = tdV + HD, tdQ = tdK + HD;

  if (warp == kBwdSoftmaxWarps) {
    if (lane == 0) {
      mbar_arrive_expect_tx(bar_kv, 2 * kTile);
      tma_load_3d(Ks, &tmK, bar_kv, h * HD, k0, b);
      tma_load_3d(Vs, &tmV, bar_kv, h * HD, k0, b);
      for (int i = 0; i < 2 && i < nq; ++i) {
        mbar_arrive_expect_tx(&bar_qd[i], 2 * kTile);
        tma_load_3d(Qs + i * kTile, &tmQ, &bar_qd[i], h * HD, i * TQ, b);
        tma_load_3d(dOs + i * kTile, &tmdO, &bar_qd[i], h * HD, i * TQ, b);
      }
      mbar_wait(bar_kv, 0);
      mbar_wait(&bar_qd[0], 0);
      tc_fence_after();
      mma_kmajor_k64(tS, smem_u32(Qs), smem_u32(Ks), TK);          // S  = Q_0  K^T
      mma_kmajor_k64(tdP, smem_u32(dOs), smem_u32(Vs), TK);        // dP = dO_0 V^T
      umma_commit(bar_s);
      const uint32_t id_t = umma_idesc_bf16(128, HD, 1, 1), id_q = umma_idesc_bf16(128, HD, 0, 1);
      for (int i = 0; i < nq; ++i) {
        const uint32_t sQ = smem_u32(Qs + (i & 1) * kTile), sdO = smem_u32(dOs + (i & 1) * kTile);
        mbar_wait(bar_p, i & 1);                 // P_i, dS_i in smem; S / dP / dQ_{i-1} consumed
        tc_fence_after();
        // dV += P^T dO_i ; dK += dS^T Q_i : A = P / dS MN-major (keys contiguous, 2 blocks of 64), K-dim = queries
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_bf16(tdV, umma_smem_desc(smem_u32(Ps) + k * 2048, kTile, 1024), umma_smem_desc(sdO + k * 2048, 8192, 1024),
                    id_t, i > 0 || k > 0);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_bf16(tdK, umma_smem_desc(smem_u32(dSs) + k * 2048, kTile, 1024), umma_smem_desc(sQ + k * 2048, 8192, 1024),
                    id_t, i > 0 || k > 0);
        // dQ_i = dS K : A = dS K-major, B = K MN-major (keys x d)
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_bf16(tdQ, umma_smem_desc(smem_u32(dSs) + (k >> 2) * kTile + (k & 3) * 32, 16, 1024),
                    umma_smem_desc(smem_u32(Ks) + k * 2048, 8192, 1024), id_q, k > 0);
        umma_commit(bar_g);
        if (i + 1 < nq) {                        // next tile's scores run under this tile's dQ drain
          mbar_wait(&bar_qd[(i + 1) & 1], ((i + 1) >> 1) & 1);
          tc_fence_after();
          mma_kmajor_k64(tS, smem_u32(Qs + ((i + 1) & 1) * kTile), smem_u32(Ks), TK);
          mma_kmajor_k64(tdP, smem_u32(dOs + ((i + 1) & 1) * kTile), smem_u32(Vs), TK);
          umma_commit(bar_s);
        }
        if (i + 2 < nq) {                        // Q / dO buffer i & 1 is free once this tile's MMAs are done
          mbar_wait(bar_g, i & 1);
          mbar_arrive_expect_tx(&bar_qd[i & 1], 2 * kTile);
          tma_load_3d(Qs + (i & 1) * kTile, &tmQ, &bar_qd[i & 1], h * HD, (i + 2) * TQ, b);
          tma_load_3d(dOs + (i & 1) * kTile, &tmdO, &bar_qd[i & 1], h * HD, (i + 2) * TQ, b);
        }
      }
    }
    __syncwarp();
  } else {
    // 8 softmax warps: warps w and w + 4 share a TMEM lane quarter (query rows 32 (w & 3) ..) and take one 64-key half
    // each - P = exp(S - lse) needs no row reduction, so the halves are independent.  Two warps per SM sub-partition:
    // one waits on TMEM / MUFU / shared memory while the other issues.
    const int quarter = warp & 3, half = warp >> 2, row = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const float sl2 = p.scale * kLog2e;
    const uint32_t mw[4] = {maskw[0], maskw[1], maskw[2], maskw[3]};
    for (int i = 0; i < nq; ++i) {
      const int q = i * TQ + row;
      const bool row_ok = q < p.Sq;
      const long sidx = ((long)b * p.H + h) * p.Sq + q;
      const float lse_l2 = row_ok ? p.lse[sidx] * kLog2e : 0.f;
      const float delta = row_ok ? fa.delta[sidx] : 0.f;
      mbar_wait(bar_s, i & 1);
      tc_fence_after();
      {
        uint32_t sa[32], da[32], sb2[32], db2[32];
        tmem_ld_32x32(tS + lane_addr + half * 64, sa);
        tmem_ld_32x32(tdP + lane_addr + half * 64, da);
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c = half * 2 + cc;                    // 32-key chunk of the tile
          uint32_t (&rs)[32] = cc ? sb2 : sa;
          uint32_t (&rd)[32] = cc ? db2 : da;
          tmem_ld_wait();
          if (cc == 0) {
            tmem_ld_32x32(tS + lane_addr + (c + 1) * 32, sb2);
            tmem_ld_32x32(tdP + lane_addr + (c + 1) * 32, db2);
          }
          const uint32_t w = row_ok ? (half ? (cc ? mw[3] : mw[2]) : (cc ? mw[1] : mw[0])) : 0u;
          float pv[32], ds[32];
          if (w == 0xffffffffu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              pv[j] = ex2_approx(fmaf(__uint_as_float(rs[j]), sl2, -lse_l2));
              ds[j] = pv[j] * ((__uint_as_float(rd[j]) - delta) * p.scale);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              pv[j] = ((w >> j) & 1) ? ex2_approx(fmaf(__uint_as_float(rs[j]), sl2, -lse_l2)) : 0.f;
              ds[j] = pv[j] * ((__uint_as_float(rd[j]) - delta) * p.scale);
            }
          }
          uint8_t* pb = Ps + half * kTile; uint8_t* db = dSs + half * kTile;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            *reinterpret_cast<uint4*>(pb + swz(row, cc * 4 + u)) = pack8(&pv[u * 8]);
            *reinterpret_cast<uint4*>(db + swz(row, cc * 4 + u)) = pack8(&ds[u * 8]);
          }
        }
      }
      fence_proxy_async();
      tc_fence_before();
      bwd_softmax_bar_sync();
      if (tid == 0) mbar_arrive(bar_p);
      // ---- drain dQ_i: TMEM -> fp32 staging -> TMA reduce-add into the fp32 workspace
      mbar_wait(bar_g, i & 1);
      tc_fence_after();
      if (tid == 0) bulk_wait_read<0>();         // the previous tile's reduce has finished reading the staging tile
      bwd_softmax_bar_sync();
      {
        uint32_t r[32];
        tmem_ld_32x32(tdQ + lane_addr + half * 32, r);
        tmem_ld_wait();
        // two [128 rows][32 fp32] halves (one per warp of a quarter), 128B-swizzled: conflict-free 16-byte stores
        uint8_t* sb = reinterpret_cast<uint8_t*>(stage) + half * kTile;
#pragma unroll
        for (int u = 0; u < 8; ++u)
          *reinterpret_cast<uint4*>(sb + swz(row, u)) = make_uint4(r[u * 4], r[u * 4 + 1], r[u * 4 + 2], r[u * 4 + 3]);
      }
      fence_proxy_async();
      tc_fence_before();
      bwd_softmax_bar_sync();
      if (tid == 0) {
        tma_reduce_add_3d(&tmdQacc, stage, h * HD, i * TQ, b);
        tma_reduce_add_3d(&tmdQacc, reinterpret_cast<uint8_t*>(stage) + kTile, h * HD + 32, i * TQ, b);
        bulk_commit();
      }
    }
    // ---- dV_j, dK_j -> bf16 -> staging (Q / dO buffers are dead) -> TMA stores (rows past Sk are clipped)
    {
      const uint32_t src[2] = {tdV, tdK};
      uint8_t* dst[2] = {Qs, dOs};
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        uint32_t r[32];
        tmem_ld_32x32(src[g] + lane_addr + half * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float f[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[u * 8 + j]);
          *reinterpret_cast<uint4*>(dst[g] + swz(row, half * 4 + u)) = pack8(f);
        }
      }
    }
    fence_proxy_async();
    tc_fence_before();
    bwd_softmax_bar_sync();
    if (tid == 0) {
      tma_store_3d(&tmdV, Qs, h * HD, k0, b);
      tma_store_3d(&tmdK, dOs, h * HD, k0, b);
      bulk_commit();
      bulk_wait_all();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kBwdSoftmaxWarps) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

constexpr size_t fwd_smem(int Sk) { return 6 * (size_t)kTile + 1024 + 128 + (size_t)((Sk + TK - 1) / TK) * 16; }
constexpr size_t kFlashBwdSmem = 12 * (size_t)kTile + 1024 + 256;

// fp32 [B][S][cols] dense accumulator: box = 32 columns (128 B) x 128 rows x 1 clip, 128B swizzle
int acc_map(CUtensorMap* out, const float* base, int B, int S, int cols) {
  EncodeTiledFn fn = enc_fn();
  if (!fn) { exg_set_error("attn_flash: cuTensorMapEncodeTiled unavailable"); return EXG_EUNSUPPORTED; }
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)S, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)cols * 4, (cuuint64_t)S * cols * 4};
  cuuint32_t box[3] = {32, 128, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { exg_set_error("attn_flash: fp32 tensor map failed (%d)", (int)r); return EXG_EINVAL; }
  return 0;
}

}  // namespace

// 1 if the multi-tile tensor-core kernels take this problem (bf16, head_dim 64, TMA-compatible strides).
int exg_attn_flash_ok(int dtype, long ldq, long ldk, long ldv, long ldo, int Sq, int Sk, int d, const void* q,
                      const void* k, const void* v, const void* o) {
  static const int off = getenv("EXG_NO_ATTN_FLASH") ? atoi(getenv("EXG_NO_ATTN_FLASH")) : 0;
  if (off || dtype != EXG_BF16 || d != 64 || Sk < 1 || Sk > kMaxMaskWords * 32) return 0;
  if (ldq % 8 || ldk % 8 || ldv % 8 || ldo % 8) return 0;
  if (!a16(q) || !a16(k) || !a16(v) || !a16(o)) return 0;
  return 1;
}

int exg_attn_flash_fwd(const void* q, const void* k, const void* v, long ldq, long ldk, long ldv,
                       const uint8_t* key_pad, void* o, long ldo, float* lse, int B, int H, int Sq, int Sk,
                       float scale, void* stream) {
  CUtensorMap mq, mk, mv, mo;
  int rc;
  // version 2 (64-key tiles, double-buffered scores) unless EXG_FLASH_FWD_V1=1
  static const int v1 = getenv("EXG_FLASH_FWD_V1") ? atoi(getenv("EXG_FLASH_FWD_V1")) : 0;
  if ((rc = seq_map(&mq, q, B, Sq, H * HD, ldq))) return rc;
  if ((rc = seq_map(&mk, k, B, Sk, H * HD, ldk, v1 ? 128 : KT2))) return rc;
  if ((rc = seq_map(&mv, v, B, Sk, H * HD, ldv, v1 ? 128 : KT2))) return rc;
  if ((rc = seq_map(&mo, o, B, Sq, H * HD, ldo))) return rc;
  if (!v1) {
    // version 3 (8 softmax warps) unless EXG_FLASH_FWD_V2=1; + 1 KB for the half-row exchange buffer
    static const int v2 = getenv("EXG_FLASH_FWD_V2") ? atoi(getenv("EXG_FLASH_FWD_V2")) : 0;
    static size_t cur2 = 0, cur3 = 0;
    const size_t need2 = fwd_smem(Sk) + (v2 ? 0 : 1024);
    size_t& cur = v2 ? cur2 : cur3;
    if (need2 > cur) {
      cudaError_t e = v2 ? cudaFuncSetAttribute(attn_flash_fwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need2)
                         : cudaFuncSetAttribute(attn_flash_fwd3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need2);
      if (e != cudaSuccess) { exg_set_error("attn_flash_fwd: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
      cur = need2;
    }
    AtcArgs a2 = {B, H, Sq, Sk, scale, key_pad, lse};
    dim3 grid2((Sq + TQ - 1) / TQ, H, B);
    if (v2) attn_flash_fwd2_kernel<<<grid2, FT, need2, (cudaStream_t)stream>>>(mq, mk, mv, mo, a2);
    else    attn_flash_fwd3_kernel<<<grid2, FT3, need2, (cudaStream_t)stream>>>(mq, mk, mv, mo, a2);
    EXG_LAUNCH_CHECK("attn_flash_fwd");
    return 0;
  }
  static size_t cur = 0;
  const size_t need = fwd_smem(Sk);
  if (need > cur) {
    cudaError_t e = cudaFuncSetAttribute(attn_flash_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need);
    if (e != cudaSuccess) { exg_set_error("attn_flash_fwd: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    cur = need;
  }
  AtcArgs a = {B, H, Sq, Sk, scale, key_pad, lse};
  dim3 grid((Sq + TQ - 1) / TQ, H, B);
  attn_flash_fwd_kernel<<<grid, FT, need, (cudaStream_t)stream>>>(mq, mk, mv, mo, a);
  EXG_LAUNCH_CHECK("attn_flash_fwd");
  return 0;
}

// workspace: fp32 [B*H*Sq] delta followed by fp32 [B*Sq*H*64] dQ accumulator (exg_attn_bwd_workspace)
int exg_attn_flash_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, long ldq,
                       long ldk, long ldv, long ldo, const uint8_t* key_pad, const float* lse, float* workspace,
                       void* dq, void* dk, void* dv, int B, int H, int Sq, int Sk, float scale, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  float* delta = workspace;
  float* acc = workspace + (((size_t)B * H * Sq + 31) & ~(size_t)31);       // keep the accumulator 128-byte aligned
  CUtensorMap mq, mk, mv, mdo, macc, mdk, mdv;
  int rc;
  if ((rc = seq_map(&mq, q, B, Sq, H * HD, ldq))) return rc;
  if ((rc = seq_map(&mk, k, B, Sk, H * HD, ldk))) return rc;
  if ((rc = seq_map(&mv, v, B, Sk, H * HD, ldv))) return rc;
  if ((rc = seq_map(&mdo, dout, B, Sq, H * HD, ldo))) return rc;
  if ((rc = acc_map(&macc, acc, B, Sq, H * HD))) return rc;
  if ((rc = seq_map(&mdk, dk, B, Sk, H * HD, ldk))) return rc;
  if ((rc = seq_map(&mdv, dv, B, Sk, H * HD, ldv))) return rc;
  static bool once = false;
  if (!once) {
    cudaError_t e = cudaFuncSetAttribute(attn_flash_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kFlashBwdSmem);
    if (e != cudaSuccess) { exg_set_error("attn_flash_bwd: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    once = true;
  }
  const long rows = (long)B * Sq;
  EXG_CUDA(cudaMemsetAsync(acc, 0, (size_t)rows * H * HD * 4, st));
  attn_delta_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(reinterpret_cast<const uint16_t*>(o),
                                                                reinterpret_cast<const uint16_t*>(dout), ldo, delta, B, H, Sq);
  EXG_LAUNCH_CHECK("attn_delta");
  FlashBwdArgs fa = {{B, H, Sq, Sk, scale, key_pad, const_cast<float*>(lse)}, delta};
  dim3 grid((Sk + TK - 1) / TK, H, B);
  attn_flash_bwd_kernel<<<grid, FTB, kFlashBwdSmem, st>>>(mq, mk, mv, mdo, macc, mdk, mdv, fa);
  EXG_LAUNCH_CHECK("attn_flash_bwd");
  const long n8 = rows * H * HD / 8;
  attn_dq_convert_kernel<<<(unsigned)((n8 + 255) / 256), 256, 0, st>>>(acc, reinterpret_cast<uint16_t*>(dq), ldq, rows,
                                                                      H * HD);
  EXG_LAUNCH_CHECK("attn_dq_convert");
  return 0;
}
