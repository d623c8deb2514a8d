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
//
// Occupancy is what these kernels live on: a (clip, head) problem is 24-32 KB of operands and a serial chain of
// TMA -> MMA -> TMEM read -> softmax -> MMA -> store (~5 us), i.e. latency-bound, and the only way to hide the
// chain is more CTAs per SM.  So every smem tile has ceil(S/16)*16 rows instead of 128 (the TMA box is that tall;
// an M = 128 MMA that reads an A operand past its tile reads the FOLLOWING tile - finite or not, that only produces
// accumulator rows >= S, which no store writes and no other row depends on), the score MMAs use N = ceil(Sk/16)*16
// and TMEM is allocated for what is live at one time:
//   forward  S = 64: 32 KB smem,  64 TMEM columns (O reuses the columns of S)  -> 6 CTAs / SM   (was 80 KB, 256: 2)
//            S = 96: 60 KB,      128 columns                                   -> 3 CTAs / SM   (was 2)
//   backward S = 64: 48 KB,      128 columns (dQ reuses dV's after its drain)  -> 4 CTAs / SM   (was 96 KB, 256: 2)
//            S = 96: 96 KB,      256 columns                                   -> 2 CTAs / SM   (was 128 KB: 1)
// Longer key sequences use the flash kernels in attention_flash.cu.
// Replaces nn.MultiheadAttention's scaled-dot-product core (model/tfm_model.py:32,75,79) and its backward.
#include "attn_tc_common.cuh"

namespace {

// bytes of an n-row tile; tiles start on 1024-byte boundaries (n is a multiple of 16)
__device__ __host__ __forceinline__ uint32_t tile_bytes(int rows) { return (uint32_t)rows * 128u; }

// ------------------------------------------------------------------------------------------ forward
// smem: [Q tq][P nblk x tq][K tk][V tk] (+ pad so that P's last block + 16 KB stays inside) [barriers, mask words]
__global__ void __launch_bounds__(128)
attn_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, AtcArgs p,
                   uint32_t tiles_bytes, uint32_t tmem_cols) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t tqb = tile_bytes(p.tq_rows), tkb = tile_bytes(p.tk_rows);
  const int nblk = (p.Sk + 63) >> 6;
  uint8_t* Qs = smem; uint8_t* Ps = Qs + tqb; uint8_t* Ks = Ps + nblk * tqb; uint8_t* Vs = Ks + tkb;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + tiles_bytes);    // [0] loads, [1] MMAs
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  uint32_t* maskw = tmem_slot + 1;                                     // [4]

  const int tid = threadIdx.x, warp = tid >> 5;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * TQ;
  if (tid == 0) {
    mbar_init(&bars[0], 1); mbar_init(&bars[1], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
  }
  if (warp == 0) tmem_alloc(tmem_slot, tmem_cols);
  build_mask(maskw, p, b, tid);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem;          // O accumulates over S's columns: every thread has read S by then

  if (tid == 0) {
    mbar_arrive_expect_tx(&bars[0], tqb + 2 * tkb);
    tma_load_3d(Qs, &tmQ, &bars[0], h * HD, q0, b);
    tma_load_3d(Ks, &tmK, &bars[0], h * HD, 0, b);
    tma_load_3d(Vs, &tmV, &bars[0], h * HD, 0, b);
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    mma_kmajor_k64(tS, smem_u32(Qs), smem_u32(Ks), p.tk_rows);
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 0);
  tc_fence_after();

  // ---- masked softmax over the keys; thread = query row (warps whose 32 rows lie past the tile sit this out)
  const bool warp_on = warp * 32 < p.tq_rows;
  const bool row_in = tid < p.tq_rows;                 // this thread's row exists in the smem tiles
  const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
  float sum = 1.f, mx = 0.f;
  if (warp_on) {
    const uint32_t m0 = maskw[0], m1 = maskw[1], m2 = maskw[2], m3 = maskw[3];
    const int nchunk = (p.Sk + 31) / 32;
    const float sl2 = p.scale * kLog2e;
    mx = -INFINITY;
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
    sum = 0.f;
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
        if (c * 32 >= p.tk_rows) break;                    // columns no P V k-step will read
#pragma unroll
        for (int j = 0; j < 32; ++j) pv[j] = 0.f;
      }
      if (row_in) {
        uint8_t* blk = Ps + (c >> 1) * tqb;
#pragma unroll
        for (int u = 0; u < 4; ++u)
          *reinterpret_cast<uint4*>(blk + swz(tid, (c & 1) * 4 + u)) = pack8(&pv[u * 8]);
      }
    }
    if (q0 + tid < p.Sq && p.lse)
      p.lse[((long)b * p.H + h) * p.Sq + q0 + tid] = mx * p.scale + __logf(sum);
  }

  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    // O[128 x 64] = P[128 x Sk] V[Sk x 64]: A = P K-major (64-key blocks of tq rows), B = V MN-major (keys x d)
    const uint32_t idesc = umma_idesc_bf16(128, HD, 0, 1);
    const int nk = p.tk_rows >> 4;
    for (int k = 0; k < nk; ++k) {
      const uint64_t ad = umma_smem_desc(smem_u32(Ps) + (k >> 2) * tqb + (k & 3) * 32, 16, 1024);
      const uint64_t bd = umma_smem_desc(smem_u32(Vs) + k * 2048, 8192, 1024);
      umma_bf16(tO, ad, bd, idesc, k > 0);
    }
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 1);
  tc_fence_after();

  // ---- O / l -> bf16 -> swizzled staging (the Q tile is free now) -> TMA store (rows >= Sq are clipped)
  if (warp_on) {
    const float inv = 1.0f / sum;
    uint32_t r[64];
    tmem_ld_32x32(tO + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
    tmem_ld_32x32(tO + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
    tmem_ld_wait();
    if (row_in) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float f[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[u * 8 + j]) * inv;
        *reinterpret_cast<uint4*>(Qs + swz(tid, u)) = pack8(f);
      }
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
    tmem_dealloc(tmem, tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------ backward
// smem: [P nblk x tq][dS nblk x tq][Q tq][dO tq][K tk][V tk] (+ pad so that dO + 16 KB stays inside) [barriers, mask]
// P and dS hold [tq query rows][64 keys] per 64-key block.  The MN-major descriptors of the dV / dK MMAs (M = 128
// keys) read "keys 64..127" from the block that FOLLOWS (block stride tq rows): with <= 64 keys that is the next
// tile (dS for P, Q for dS - valid memory, finite or not), which only produces accumulator rows >= 64 that the
// clipped TMA stores never write.
// TMEM: S | dP (N columns each), then dV | dK | dQ (64 each) over the same columns.  With N <= 64 the allocation
// is 128 columns and dQ = dS K is issued after dV has been drained (into dV's columns); otherwise 256 and all
// three products are issued together.
__global__ void __launch_bounds__(128)
attn_tc_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                   const __grid_constant__ CUtensorMap tmdQ, const __grid_constant__ CUtensorMap tmdK,
                   const __grid_constant__ CUtensorMap tmdV, const void* __restrict__ o, long ldo, AtcArgs p,
                   uint32_t tiles_bytes, uint32_t tmem_cols) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t tqb = tile_bytes(p.tq_rows), tkb = tile_bytes(p.tk_rows);
  const int nblk = (p.Sk + 63) >> 6;
  uint8_t* Ps = smem; uint8_t* dSs = Ps + nblk * tqb;
  uint8_t* Qs = dSs + nblk * tqb; uint8_t* dOs = Qs + tqb; uint8_t* Ks = dOs + tqb; uint8_t* Vs = Ks + tkb;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + tiles_bytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  uint32_t* maskw = tmem_slot + 1;

  const int tid = threadIdx.x, warp = tid >> 5;
  const int b = blockIdx.z, h = blockIdx.y;
  if (tid == 0) {
    mbar_init(&bars[0], 1); mbar_init(&bars[1], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmdO);
  }
  if (warp == 0) tmem_alloc(tmem_slot, tmem_cols);
  build_mask(maskw, p, b, tid);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int N = p.tk_rows;
  const bool split = tmem_cols < 3 * HD;                           // dQ after dV's drain, in dV's columns
  const uint32_t tS = tmem, tdP = tmem + N;                        // phase 1
  const uint32_t tdV = tmem, tdK = tmem + HD, tdQ = split ? tmem : tmem + 2 * HD;   // phase 2 (S / dP are dead)

  if (tid == 0) {
    mbar_arrive_expect_tx(&bars[0], 2 * tqb + 2 * tkb);
    tma_load_3d(Qs, &tmQ, &bars[0], h * HD, 0, b);
    tma_load_3d(Ks, &tmK, &bars[0], h * HD, 0, b);
    tma_load_3d(Vs, &tmV, &bars[0], h * HD, 0, b);
    tma_load_3d(dOs, &tmdO, &bars[0], h * HD, 0, b);
  }
  // per-row softmax statistics while the tiles are in flight
  const bool warp_on = warp * 32 < p.tq_rows;
  const bool row_in = tid < p.tq_rows;
  const bool row_ok = tid < p.Sq;
  const float lse_l2 = row_ok ? p.lse[((long)b * p.H + h) * p.Sq + tid] * kLog2e : 0.f;
  mbar_wait(&bars[0], 0);
  if (tid == 0) {
    tc_fence_after();
    mma_kmajor_k64(tS, smem_u32(Qs), smem_u32(Ks), N);       // S  = Q  K^T
    mma_kmajor_k64(tdP, smem_u32(dOs), smem_u32(Vs), N);     // dP = dO V^T
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
  if (warp_on) {
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
        ds[j] = ((mw >> j) & 1) ? pr * (__uint_as_float(rd[j]) - delta) * p.scale : 0.f;
      }
      if (row_in) {
        uint8_t* pb = Ps + (c >> 1) * tqb; uint8_t* db = dSs + (c >> 1) * tqb;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          *reinterpret_cast<uint4*>(pb + swz(tid, (c & 1) * 4 + u)) = pack8(&pv[u * 8]);
          *reinterpret_cast<uint4*>(db + swz(tid, (c & 1) * 4 + u)) = pack8(&ds[u * 8]);
        }
      }
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  const int nkq = p.tq_rows >> 4, nkk = N >> 4;
  const uint32_t id_q = umma_idesc_bf16(128, HD, 0, 1);
  if (tid == 0) {
    tc_fence_after();
    // dV[key x d] = P^T dO ; dK = dS^T Q : A = P / dS read MN-major (keys contiguous, 64-key blocks tq rows apart),
    // K-dim = queries
    const uint32_t id_t = umma_idesc_bf16(128, HD, 1, 1);
    for (int k = 0; k < nkq; ++k)
      umma_bf16(tdV, umma_smem_desc(smem_u32(Ps) + k * 2048, tqb, 1024),
                umma_smem_desc(smem_u32(dOs) + k * 2048, 8192, 1024), id_t, k > 0);
    for (int k = 0; k < nkq; ++k)
      umma_bf16(tdK, umma_smem_desc(smem_u32(dSs) + k * 2048, tqb, 1024),
                umma_smem_desc(smem_u32(Qs) + k * 2048, 8192, 1024), id_t, k > 0);
    if (!split) {
      // dQ[query x d] = dS K : A = dS K-major, B = K MN-major (keys x d)
      for (int k = 0; k < nkk; ++k)
        umma_bf16(tdQ, umma_smem_desc(smem_u32(dSs) + (k >> 2) * tqb + (k & 3) * 32, 16, 1024),
                  umma_smem_desc(smem_u32(Ks) + k * 2048, 8192, 1024), id_q, k > 0);
    }
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 1);
  tc_fence_after();

  // ---- gradients -> bf16 -> staging (dead input tiles) -> TMA stores (rows past the sequence end are clipped).
  // rows of dV / dK are keys (tk rows), rows of dQ are queries (tq rows)
  auto drain = [&](uint32_t src, uint8_t* dst, int rows) {
    if (warp * 32 < rows) {
      uint32_t r[64];
      tmem_ld_32x32(src + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
      tmem_ld_32x32(src + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
      tmem_ld_wait();
      if (tid < rows) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          float f[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[u * 8 + j]);
          *reinterpret_cast<uint4*>(dst + swz(tid, u)) = pack8(f);
        }
      }
    }
  };
  drain(tdV, Vs, p.tk_rows);                 // V has been dead since dP
  if (split) {
    tc_fence_before();
    __syncthreads();                         // every thread has read dV: its columns take dQ
    if (tid == 0) {
      tc_fence_after();
      for (int k = 0; k < nkk; ++k)
        umma_bf16(tdQ, umma_smem_desc(smem_u32(dSs) + (k >> 2) * tqb + (k & 3) * 32, 16, 1024),
                  umma_smem_desc(smem_u32(Ks) + k * 2048, 8192, 1024), id_q, k > 0);
      umma_commit(&bars[1]);
    }
    mbar_wait(&bars[1], 0);                  // third completion of this barrier
    tc_fence_after();
  }
  drain(tdK, Ks, p.tk_rows);                 // K is dead once dQ = dS K has completed
  drain(tdQ, Qs, p.tq_rows);
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
    tmem_dealloc(tmem, tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------ host side
constexpr uint32_t kTailBytes = 64;                      // barriers, TMEM slot, mask words
constexpr size_t kMaxSmem = 8 * kTile + 1024 + kTailBytes;

inline int round16(int n) { return (n + 15) / 16 * 16; }
inline uint32_t umax(uint32_t a, uint32_t b) { return a > b ? a : b; }

struct AtcPlan { int tq_rows, tk_rows; uint32_t tiles_bytes, tmem_cols; size_t smem; };

// forward: [Q][P x nblk][K][V]; the furthest M = 128 read of an A operand starts at P's last block
AtcPlan plan_fwd(int Sq, int Sk) {
  AtcPlan pl;
  pl.tq_rows = Sq >= TQ ? TQ : round16(Sq);
  pl.tk_rows = round16(Sk);
  const uint32_t tqb = pl.tq_rows * 128u, tkb = pl.tk_rows * 128u, nblk = (Sk + 63) / 64;
  const uint32_t used = tqb + nblk * tqb + 2 * tkb;
  const uint32_t reach = tqb + (nblk - 1) * tqb + kTile;
  pl.tiles_bytes = umax(used, reach);
  pl.tmem_cols = pl.tk_rows <= 64 ? 64 : 128;             // S (N columns), then O (64) over the same columns
  pl.smem = pl.tiles_bytes + 1024 + kTailBytes;
  return pl;
}

// backward: [P x nblk][dS x nblk][Q][dO][K][V]; the furthest M = 128 read starts at dO
AtcPlan plan_bwd(int Sq, int Sk) {
  AtcPlan pl;
  pl.tq_rows = round16(Sq);
  pl.tk_rows = round16(Sk);
  const uint32_t tqb = pl.tq_rows * 128u, tkb = pl.tk_rows * 128u, nblk = (Sk + 63) / 64;
  const uint32_t used = 2 * nblk * tqb + 2 * tqb + 2 * tkb;
  const uint32_t reach = 2 * nblk * tqb + tqb + kTile;
  pl.tiles_bytes = umax(used, reach);
  pl.tmem_cols = pl.tk_rows <= 64 ? 128 : 256;            // S | dP (2N), then dV | dK (| dQ)
  pl.smem = pl.tiles_bytes + 1024 + kTailBytes;
  return pl;
}

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
  const AtcPlan pl = plan_fwd(Sq, Sk);
  CUtensorMap mq, mk, mv, mo;
  int rc;
  if ((rc = seq_map(&mq, q, B, Sq, H * HD, ldq, pl.tq_rows))) return rc;
  if ((rc = seq_map(&mk, k, B, Sk, H * HD, ldk, pl.tk_rows))) return rc;
  if ((rc = seq_map(&mv, v, B, Sk, H * HD, ldv, pl.tk_rows))) return rc;
  if ((rc = seq_map(&mo, o, B, Sq, H * HD, ldo, pl.tq_rows))) return rc;
  static bool once = false;
  if (!once) {
    cudaError_t e = cudaFuncSetAttribute(attn_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kMaxSmem);
    if (e != cudaSuccess) { exg_set_error("attn_tc_fwd: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    once = true;
  }
  AtcArgs a = {B, H, Sq, Sk, scale, key_pad, lse, pl.tq_rows, pl.tk_rows};
  dim3 grid((Sq + TQ - 1) / TQ, H, B);
  attn_tc_fwd_kernel<<<grid, 128, pl.smem, (cudaStream_t)stream>>>(mq, mk, mv, mo, a, pl.tiles_bytes, pl.tmem_cols);
  EXG_LAUNCH_CHECK("attn_tc_fwd");
  return 0;
}

int exg_attn_tc_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, long ldq,
                    long ldk, long ldv, long ldo, const uint8_t* key_pad, const float* lse, void* dq, void* dk,
                    void* dv, int B, int H, int Sq, int Sk, float scale, void* stream) {
  const AtcPlan pl = plan_bwd(Sq, Sk);
  CUtensorMap mq, mk, mv, mdo, mdq, mdk, mdv;
  int rc;
  if ((rc = seq_map(&mq, q, B, Sq, H * HD, ldq, pl.tq_rows))) return rc;
  if ((rc = seq_map(&mk, k, B, Sk, H * HD, ldk, pl.tk_rows))) return rc;
  if ((rc = seq_map(&mv, v, B, Sk, H * HD, ldv, pl.tk_rows))) return rc;
  if ((rc = seq_map(&mdo, dout, B, Sq, H * HD, ldo, pl.tq_rows))) return rc;
  if ((rc = seq_map(&mdq, dq, B, Sq, H * HD, ldq, pl.tq_rows))) return rc;
  if ((rc = seq_map(&mdk, dk, B, Sk, H * HD, ldk, pl.tk_rows))) return rc;
  if ((rc = seq_map(&mdv, dv, B, Sk, H * HD, ldv, pl.tk_rows))) return rc;
  static bool once = false;
  if (!once) {
    cudaError_t e = cudaFuncSetAttribute(attn_tc_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kMaxSmem);
    if (e != cudaSuccess) { exg_set_error("attn_tc_bwd: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    once = true;
  }
  AtcArgs a = {B, H, Sq, Sk, scale, key_pad, const_cast<float*>(lse), pl.tq_rows, pl.tk_rows};
  dim3 grid(1, H, B);
  attn_tc_bwd_kernel<<<grid, 128, pl.smem, (cudaStream_t)stream>>>(mq, mk, mv, mdo, mdq, mdk, mdv, o, ldo, a,
                                                                 pl.tiles_bytes, pl.tmem_cols);
  EXG_LAUNCH_CHECK("attn_tc_bwd");
  return 0;
}
