// nce_tma.cu — the bf16 fast path of the masked bidirectional MIL-NCE (train/loss.py:234-278) over materialised
// stage-major logits L[S, R=B*T, Cc=B*N]: TMA-fed, warp-autonomous streaming kernels.
//
// One persistent CTA per SM, 16 warps.  Warp w owns one 256-column chunk of the CTA's column block and walks a
// contiguous range of 8-row units; its 8 x 256 bf16 tiles (4 KB) arrive through a private 3-deep TMA ring
// (cp.async.bulk.tensor + mbarrier), so 8 KB per warp / 128 KB per SM are in flight with no CTA-wide barrier
// anywhere.  A lane owns 8 consecutive columns: column partial sums stay in registers for the warp's whole row
// range (one atomic per column at the end), row partial sums are folded across the warp with a 9-shuffle
// transposing reduction per 8 rows.  Column masks (padded sentences, TMA-zero-filled columns past Cc) are folded
// into the exponent bias (-inf), so the inner loop is convert, fma, ex2, two adds per logit.
// The backward kernel rewrites each tile in place in shared memory and sends it out with a TMA store.
// HBM-bound: forward reads L once (2 B / logit); backward reads L once and writes G once (4 B / logit).
#include "common.cuh"
#include "sm100.cuh"
#include <cuda.h>
#include <math.h>
#include <stdlib.h>

using namespace sm100;

namespace {

constexpr int WARPS = 16;
constexpr int NTHREADS = WARPS * 32;
constexpr int WCOLS = 256;                 // columns per warp chunk (TMA box inner dim, 512 B)
constexpr int UROWS = 8;                   // rows per unit (TMA box rows)
constexpr int STAGES = 3;
constexpr int STAGE_BYTES = UROWS * WCOLS * 2;          // 4096
constexpr size_t SMEM_BYTES = (size_t)WARPS * STAGES * STAGE_BYTES + WARPS * STAGES * 8 + 1024;

struct NceT {
  const uint32_t* bits; const uint8_t* col_valid; const uint8_t* own_row_mask;   // bits: packed positives [R][W] (exg_nce_pack_targets)
  float *row_all, *row_pos, *col_all, *col_pos;
  const float *wr, *wc;
  int S, B, T, N, W;
  int Bc, clip0;        // column clips and the column clip owned by row clip 0 (see nce.cu NceArgs)
  int k, g, ncb;        // chunks per column block, row groups per CTA, column blocks
  float inv_temp;
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void cvt8(const uint4& u, float (&v)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[2 * j] = __uint_as_float(w[j] << 16);
    v[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// two bf16 of one 32-bit word as a float2 (element order preserved)
__device__ __forceinline__ float2 cvt2(uint32_t w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
__device__ __forceinline__ float2 ex2f2(float2 x) { return make_float2(ex2f(x.x), ex2f(x.y)); }
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// (column block, stage, unit) cursor.  Units are dealt round-robin: virtual group vg (one per CTA row group) takes
// units vg, vg + VG, ... of every (column block, stage), so at any moment the whole grid works inside a window of
// a few MB of L (few live TLB entries / DRAM pages) instead of 148 streams megabytes apart.
struct Cursor {
  int cb, s, u, u0, du;
  __device__ __forceinline__ void first(int vg, int VG) { cb = 0; s = 0; u = u0 = vg; du = VG; }
  __device__ __forceinline__ void next(int S, int U) {
    u += du;
    if (u >= U) { u = u0; if (++s == S) { s = 0; ++cb; } }
  }
};

// Work split shared by both kernels: this warp's chunk, virtual group and item count (0 = nothing to do).
__device__ __forceinline__ long warp_items(const NceT& p, int warp, int U, long Cc, int& chunk, int& vg, int& VG) {
  if (warp >= p.k * p.g) return 0;
  chunk = warp % p.k;
  VG = (int)gridDim.x * p.g;
  vg = (int)blockIdx.x * p.g + warp / p.k;
  if (vg >= U) return 0;
  const long cnt = (U - vg + VG - 1) / VG;
  // this warp's chunk does not exist in the last (partial) column block
  const int ncb = (((long)(p.ncb - 1) * p.k + chunk) * WCOLS >= Cc) ? p.ncb - 1 : p.ncb;
  return cnt * p.S * ncb;
}

// Per-(column block, stage) lane state: exponent bias per column (-k1, or -inf when masked) and each column's clip
struct ColState {
  float kb[8];
  long c0;            // first column of this lane
  int wb_lo, wb_hi;   // clip range covered by the warp's chunk
};
__device__ __forceinline__ void col_setup(ColState& c, const NceT& p, int cb, int chunk, int lane, long Cc, float k1) {
  const long w0 = ((long)cb * p.k + chunk) * WCOLS;
  c.c0 = w0 + lane * 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const long col = c.c0 + j;
    const bool ok = col < Cc && p.col_valid[col] != 0;
    c.kb[j] = ok ? -k1 : -INFINITY;
  }
  c.wb_lo = (int)(w0 / p.N) - p.clip0;               // in row-clip numbering
  c.wb_hi = (int)(min(w0 + WCOLS - 1, Cc - 1) / p.N) - p.clip0;
}

// bit j = 1 iff column c0 + j is a positive of (clip a, row r): the lane's 8 columns meet the clip's N columns in
// one contiguous run, i.e. at most two words of the packed row.
__device__ __forceinline__ uint32_t pos_mask(const NceT& p, long c0, int a, long r) {
  const long base = (long)(a + p.clip0) * p.N;
  const long lo = max(c0, base), hi = min(c0 + 8, base + p.N);
  if (lo >= hi) return 0u;
  const int n_lo = (int)(lo - base), len = (int)(hi - lo), sh = n_lo & 31;
  const uint32_t* wp = p.bits + (size_t)r * p.W + (n_lo >> 5);
  uint32_t run = __ldg(wp) >> sh;
  if (sh + len > 32) run |= __ldg(wp + 1) << (32 - sh);
  return (run & ((1u << len) - 1u)) << (int)(lo - c0);
}

// bit j = 1 iff column c0 + j belongs to clip a
__device__ __forceinline__ uint32_t own_cols(const NceT& p, long c0, int a) {
  const long base = (long)(a + p.clip0) * p.N;
  const long lo = max(c0, base), hi = min(c0 + 8, base + p.N);
  if (lo >= hi) return 0u;
  return ((1u << (int)(hi - lo)) - 1u) << (int)(lo - c0);
}

// Positive masks of the unit's 8 rows (0 when the warp's chunk holds no column of the row's clip); returns whether
// any row of the unit can have positives in this chunk (warp-uniform).
__device__ __forceinline__ bool unit_masks(const NceT& p, const ColState& c, int a, int t, int r0, int nrows,
                                           uint32_t (&m)[8]) {
  int a_last = a, tt = t + UROWS - 1;
  while (tt >= p.T) { tt -= p.T; ++a_last; }
  if (a > c.wb_hi || a_last < c.wb_lo) return false;
  int ai = a, ti = t;
#pragma unroll
  for (int i = 0; i < UROWS; ++i) {
    m[i] = (i < nrows && ai >= c.wb_lo && ai <= c.wb_hi) ? pos_mask(p, c.c0, ai, (long)r0 + i) : 0u;
    // bits 8..15: own-clip columns whose entries are dropped on this row (reference's in-place -6e4 fill)
    if (p.own_row_mask && i < nrows && ai >= c.wb_lo && ai <= c.wb_hi && p.own_row_mask[r0 + i])
      m[i] |= own_cols(p, c.c0, ai) << 8;
    if (++ti == p.T) { ti = 0; ++ai; }
  }
  return true;
}

// Sum 8 per-lane values over the warp; afterwards lane l with (l & 3) == 0 holds the total of index
// ((l >> 4) & 1) * 4 + ((l >> 3) & 1) * 2 + ((l >> 2) & 1).  9 shuffles instead of 40.
__device__ __forceinline__ float fold8(const float (&x)[8], int lane) {
  const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4;
  float a[4], b[2], c;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float send = h4 ? x[i] : x[i + 4];
    const float keep = h4 ? x[i + 4] : x[i];
    a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float send = h3 ? a[i] : a[i + 2];
    const float keep = h3 ? a[i + 2] : a[i];
    b[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  {
    const float send = h2 ? b[0] : b[1];
    const float keep = h2 ? b[1] : b[0];
    c = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  c += __shfl_xor_sync(0xffffffffu, c, 2);
  c += __shfl_xor_sync(0xffffffffu, c, 1);
  return c;
}

// ------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(NTHREADS, 1)
nce_fwd_tma_kernel(const __grid_constant__ CUtensorMap tmL, NceT p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int R = p.B * p.T, U = (R + UROWS - 1) / UROWS;
  const long Cc = (long)p.Bc * p.N;
  int chunk, vg, VG;
  const long lo = 0, hi = warp_items(p, warp, U, Cc, chunk, vg, VG);
  if (hi == 0) return;
  uint8_t* buf = smem + (size_t)warp * STAGES * STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)WARPS * STAGES * STAGE_BYTES) + warp * STAGES;
  if (lane == 0) {
#pragma unroll
    for (int st = 0; st < STAGES; ++st) mbar_init(full + st, 1);
    fence_barrier_init();
  }
  __syncwarp();
  const float k1 = p.inv_temp * 1.4426950408889634f;

  Cursor cur, pre;                       // item being consumed / next item to prefetch
  cur.first(vg, VG);
  pre = cur;
  long pre_item = lo;
  if (lane == 0) {
#pragma unroll
    for (int st = 0; st < STAGES; ++st) {
      if (pre_item < hi) {
        mbar_arrive_expect_tx(full + st, STAGE_BYTES);
        tma_load_3d(buf + st * STAGE_BYTES, &tmL, full + st, (pre.cb * p.k + chunk) * WCOLS, pre.u * UROWS, pre.s);
        pre.next(p.S, U);
        ++pre_item;
      }
    }
  }
  pre_item = __shfl_sync(0xffffffffu, pre_item, 0);      // only lane 0's cursor is used afterwards

  ColState cs;
  float ca[8], cp[8];
  int cs_cb = -1, cs_s = -1;
  int st = 0; uint32_t ph = 0;
  int a = 0, t = 0;                       // clip / second of the unit's first row
  for (long item = lo; item < hi; ++item) {
    if (cur.cb != cs_cb || cur.s != cs_s) {
      if (cs_cb >= 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (ca[j] != 0.f) atomicAdd(p.col_all + (size_t)cs_s * Cc + cs.c0 + j, ca[j]);
          if (cp[j] != 0.f) atomicAdd(p.col_pos + (size_t)cs_s * Cc + cs.c0 + j, cp[j]);
        }
      }
      cs_cb = cur.cb; cs_s = cur.s;
      col_setup(cs, p, cur.cb, chunk, lane, Cc, k1);
#pragma unroll
      for (int j = 0; j < 8; ++j) ca[j] = cp[j] = 0.f;
    }
    const int r0 = cur.u * UROWS;
    a = (int)((unsigned)r0 / (unsigned)p.T);
    t = r0 - a * p.T;
    const int nrows = min(UROWS, R - r0);
    uint32_t m[8];
    const bool unit_own = unit_masks(p, cs, a, t, r0, nrows, m);    // issued before the wait: overlaps the TMA latency
    mbar_wait(full + st, ph);
    const uint32_t sp = smem_u32(buf + st * STAGE_BYTES) + lane * 16;
    float ra[8];
    if (nrows == UROWS && !unit_own) {
      // common case: a full unit with no positives in this chunk -- packed fp32x2 math, no per-row branches
      const float2 k2 = make_float2(k1, k1);
      float2 kb2[4], c2[4];
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        kb2[jj] = make_float2(cs.kb[2 * jj], cs.kb[2 * jj + 1]);
        c2[jj] = make_float2(ca[2 * jj], ca[2 * jj + 1]);
      }
#pragma unroll
      for (int i = 0; i < UROWS; ++i) {
        const uint4 q = lds128(sp + i * (WCOLS * 2));
        const float2 e0 = ex2f2(__ffma2_rn(cvt2(q.x), k2, kb2[0]));
        const float2 e1 = ex2f2(__ffma2_rn(cvt2(q.y), k2, kb2[1]));
        const float2 e2 = ex2f2(__ffma2_rn(cvt2(q.z), k2, kb2[2]));
        const float2 e3 = ex2f2(__ffma2_rn(cvt2(q.w), k2, kb2[3]));
        c2[0] = __fadd2_rn(c2[0], e0); c2[1] = __fadd2_rn(c2[1], e1);
        c2[2] = __fadd2_rn(c2[2], e2); c2[3] = __fadd2_rn(c2[3], e3);
        const float2 r2 = __fadd2_rn(__fadd2_rn(e0, e1), __fadd2_rn(e2, e3));
        ra[i] = r2.x + r2.y;
      }
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) { ca[2 * jj] = c2[jj].x; ca[2 * jj + 1] = c2[jj].y; }
    } else {
      int ai = a, ti = t;
#pragma unroll
      for (int i = 0; i < UROWS; ++i) {
        const uint4 q = lds128(sp + i * (WCOLS * 2));
        float v[8], e[8];
        cvt8(q, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) e[j] = ex2f(fmaf(v[j], k1, cs.kb[j]));
        if (unit_own && (m[i] >> 8)) {
#pragma unroll
          for (int j = 0; j < 8; ++j) if ((m[i] >> (8 + j)) & 1u) e[j] = 0.f;
        }
        if (i < nrows) {
          ra[i] = ((e[0] + e[1]) + (e[2] + e[3])) + ((e[4] + e[5]) + (e[6] + e[7]));
#pragma unroll
          for (int j = 0; j < 8; ++j) ca[j] += e[j];
          if (unit_own && ai >= cs.wb_lo && ai <= cs.wb_hi) {   // warp-uniform
            float rp = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if ((m[i] >> j) & 1u) { rp += e[j]; cp[j] += e[j]; }
            }
            rp = warp_sum(rp);
            if (lane == 0 && rp != 0.f) atomicAdd(p.row_pos + (size_t)cur.s * R + r0 + i, rp);
          }
        } else {
          ra[i] = 0.f;
        }
        if (++ti == p.T) { ti = 0; ++ai; }
      }
    }
    const float tot = fold8(ra, lane);
    const int idx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    if ((lane & 3) == 0 && idx < nrows && tot != 0.f) atomicAdd(p.row_all + (size_t)cur.s * R + r0 + idx, tot);
    __syncwarp();
    if (lane == 0 && pre_item < hi) {                     // refill the stage just consumed
      mbar_arrive_expect_tx(full + st, STAGE_BYTES);
      tma_load_3d(buf + st * STAGE_BYTES, &tmL, full + st, (pre.cb * p.k + chunk) * WCOLS, pre.u * UROWS, pre.s);
      pre.next(p.S, U);
      ++pre_item;
    }
    if (++st == STAGES) { st = 0; ph ^= 1; }
    cur.next(p.S, U);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (ca[j] != 0.f) atomicAdd(p.col_all + (size_t)cs_s * Cc + cs.c0 + j, ca[j]);
    if (cp[j] != 0.f) atomicAdd(p.col_pos + (size_t)cs_s * Cc + cs.c0 + j, cp[j]);
  }
}

// ------------------------------------------------------------------------------------------ backward
// G = inv_temp * e * (wr/row_all + wc/col_all - pos * (wr/row_pos + wc/col_pos)), written in place over the staged
// tile and sent out with a TMA store (rows past R / columns past Cc are clipped by the tensor map).
__global__ void __launch_bounds__(NTHREADS, 1)
nce_bwd_tma_kernel(const __grid_constant__ CUtensorMap tmL, const __grid_constant__ CUtensorMap tmG, NceT p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int R = p.B * p.T, U = (R + UROWS - 1) / UROWS;
  const long Cc = (long)p.Bc * p.N;
  int chunk, vg, VG;
  const long lo = 0, hi = warp_items(p, warp, U, Cc, chunk, vg, VG);
  if (hi == 0) return;
  uint8_t* buf = smem + (size_t)warp * STAGES * STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)WARPS * STAGES * STAGE_BYTES) + warp * STAGES;
  if (lane == 0) {
#pragma unroll
    for (int st = 0; st < STAGES; ++st) mbar_init(full + st, 1);
    fence_barrier_init();
  }
  __syncwarp();
  const float k1 = p.inv_temp * 1.4426950408889634f;
  // leak terms of the own-row mask (see nce.cu nce_bwd_kernel): 1/#valid columns per row, 1/#rows per column
  float leak_r = 0.f, leak_c = 0.f;
  if (p.own_row_mask) {
    int n = 0;
    for (long c = lane; c < Cc; c += 32) n += p.col_valid[c] != 0;
#pragma unroll
    for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    leak_r = p.inv_temp / (float)max(n, 1);
    leak_c = p.inv_temp / (float)((long)p.Bc * p.T);
  }

  Cursor cur, pre;
  cur.first(vg, VG);
  pre = cur;
  long pre_item = lo;
  if (lane == 0) {                      // two tiles ahead; the third stage is the one being stored
#pragma unroll
    for (int st = 0; st < STAGES - 1; ++st) {
      if (pre_item < hi) {
        mbar_arrive_expect_tx(full + st, STAGE_BYTES);
        tma_load_3d(buf + st * STAGE_BYTES, &tmL, full + st, (pre.cb * p.k + chunk) * WCOLS, pre.u * UROWS, pre.s);
        pre.next(p.S, U);
        ++pre_item;
      }
    }
  }
  pre_item = __shfl_sync(0xffffffffu, pre_item, 0);

  ColState cs;
  float ka[8], kp[8], kl[8];
  int cs_cb = -1, cs_s = -1;
  int st = 0; uint32_t ph = 0;
  int a = 0, t = 0;
  for (long item = lo; item < hi; ++item) {
    if (cur.cb != cs_cb || cur.s != cs_s) {
      cs_cb = cur.cb; cs_s = cur.s;
      col_setup(cs, p, cur.cb, chunk, lane, Cc, k1);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        ka[j] = kp[j] = kl[j] = 0.f;
        const long col = cs.c0 + j;
        if (col < Cc) {
          const float w = p.wc[col];
          if (w != 0.f && cs.kb[j] != -INFINITY) {
            const float qa = p.col_all[(size_t)cur.s * Cc + col];
            ka[j] = qa > 0.f ? p.inv_temp * w / qa : 0.f;
            const float q = p.col_pos[(size_t)cur.s * Cc + col];
            kp[j] = q > 0.f ? p.inv_temp * w / q : 0.f;
            kl[j] = q > 0.f ? 0.f : leak_c * w;
          }
        }
      }
    }
    const int r0 = cur.u * UROWS;
    a = (int)((unsigned)r0 / (unsigned)p.T);
    t = r0 - a * p.T;
    const int nrows = min(UROWS, R - r0);
    // row coefficients of the unit's 8 rows: lane i computes row i, broadcast below
    float rowA = 0.f, rowP = 0.f, rowL = 0.f;
    if (lane < nrows) {
      const float w = p.wr[r0 + lane];
      if (w != 0.f) {
        const float qa = p.row_all[(size_t)cur.s * R + r0 + lane];
        rowA = qa > 0.f ? p.inv_temp * w / qa : 0.f;
        const float q = p.row_pos[(size_t)cur.s * R + r0 + lane];
        rowP = q > 0.f ? p.inv_temp * w / q : 0.f;
        rowL = q > 0.f ? 0.f : leak_r * w;
      }
    }
    uint32_t m[8];
    const bool unit_own = unit_masks(p, cs, a, t, r0, nrows, m);
    mbar_wait(full + st, ph);
    const uint32_t sp = smem_u32(buf + st * STAGE_BYTES) + lane * 16;
    if (!unit_own) {
      // common case: no positives in this chunk's rows -- packed fp32x2 math, no per-row branches (rows past R
      // carry a zero row coefficient and are clipped by the store)
      const float2 k2 = make_float2(k1, k1);
      float2 kb2[4], ka2[4];
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        kb2[jj] = make_float2(cs.kb[2 * jj], cs.kb[2 * jj + 1]);
        ka2[jj] = make_float2(ka[2 * jj], ka[2 * jj + 1]);
      }
#pragma unroll
      for (int i = 0; i < UROWS; ++i) {
        const float rA = __shfl_sync(0xffffffffu, rowA, i);
        const float2 rA2 = make_float2(rA, rA);
        const uint4 q = lds128(sp + i * (WCOLS * 2));
        const float2 g0 = __fmul2_rn(ex2f2(__ffma2_rn(cvt2(q.x), k2, kb2[0])), __fadd2_rn(rA2, ka2[0]));
        const float2 g1 = __fmul2_rn(ex2f2(__ffma2_rn(cvt2(q.y), k2, kb2[1])), __fadd2_rn(rA2, ka2[1]));
        const float2 g2 = __fmul2_rn(ex2f2(__ffma2_rn(cvt2(q.z), k2, kb2[2])), __fadd2_rn(rA2, ka2[2]));
        const float2 g3 = __fmul2_rn(ex2f2(__ffma2_rn(cvt2(q.w), k2, kb2[3])), __fadd2_rn(rA2, ka2[3]));
        uint4 o;
        o.x = pack2(g0.x, g0.y); o.y = pack2(g1.x, g1.y); o.z = pack2(g2.x, g2.y); o.w = pack2(g3.x, g3.y);
        sts128(sp + i * (WCOLS * 2), o);
      }
    } else {
      int ai = a, ti = t;
#pragma unroll
      for (int i = 0; i < UROWS; ++i) {
        const float rA = __shfl_sync(0xffffffffu, rowA, i);
        const uint4 q = lds128(sp + i * (WCOLS * 2));
        float v[8], g[8];
        cvt8(q, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = ex2f(fmaf(v[j], k1, cs.kb[j])) * (rA + ka[j]);
        if (i < nrows && ai >= cs.wb_lo && ai <= cs.wb_hi) {
          const float rP = __shfl_sync(0xffffffffu, rowP, i);
          const float rL = __shfl_sync(0xffffffffu, rowL, i);
          if (m[i]) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if ((m[i] >> j) & 1u) g[j] = ex2f(fmaf(v[j], k1, cs.kb[j])) * (rA + ka[j] - rP - kp[j]);
              // dropped entry: zero, except a dropped positive of a row / column left without any positive
              if ((m[i] >> (8 + j)) & 1u) g[j] = ((m[i] >> j) & 1u) && cs.kb[j] != -INFINITY ? -(rL + kl[j]) : 0.f;
            }
          }
        }
        uint4 o;
        o.x = pack2(g[0], g[1]); o.y = pack2(g[2], g[3]); o.z = pack2(g[4], g[5]); o.w = pack2(g[6], g[7]);
        sts128(sp + i * (WCOLS * 2), o);
        if (++ti == p.T) { ti = 0; ++ai; }
      }
    }
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) {
      tma_store_3d(&tmG, buf + st * STAGE_BYTES, (cur.cb * p.k + chunk) * WCOLS, r0, cur.s);
      bulk_commit();
      if (pre_item < hi) {
        // stage (st + 2) % 3 was stored one iteration ago: wait until that store has read it, then refill it
        bulk_wait_read<1>();
        const int st2 = st >= 1 ? st - 1 : STAGES - 1;
        mbar_arrive_expect_tx(full + st2, STAGE_BYTES);
        tma_load_3d(buf + st2 * STAGE_BYTES, &tmL, full + st2, (pre.cb * p.k + chunk) * WCOLS, pre.u * UROWS, pre.s);
        pre.next(p.S, U);
        ++pre_item;
      }
    }
    if (++st == STAGES) { st = 0; ph ^= 1; }
    cur.next(p.S, U);
  }
  if (lane == 0) bulk_wait_all();
}

// bits[r, w] bit n%32 = tgt[(clip(r) * N + 32 w + n%32) * T + second(r)] != 0   (tgt uint8 [B, N, T])
__global__ void pack_targets_kernel(const uint8_t* __restrict__ tgt, uint32_t* __restrict__ bits, int B, int T, int N,
                                    int W) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)B * T * W) return;
  const int w = (int)(i % W);
  const long r = i / W;
  const int a = (int)(r / T), t = (int)(r % T);
  uint32_t v = 0;
  const int n_end = min(N, 32 * w + 32);
  for (int n = 32 * w; n < n_end; ++n) v |= (uint32_t)(tgt[((size_t)a * N + n) * T + t] != 0) << (n & 31);
  bits[i] = v;
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

// [S][R][Cc] bf16, box = 256 columns x 8 rows x 1 stage, no swizzle (a lane reads 16 B at lane * 16: conflict-free)
int logits_map(CUtensorMap* out, const void* base, int S, long R, long Cc) {
  EncodeTiledFn fn = enc_fn();
  if (!fn) return EXG_EUNSUPPORTED;
  cuuint64_t dims[3] = {(cuuint64_t)Cc, (cuuint64_t)R, (cuuint64_t)S};
  cuuint64_t strides[2] = {(cuuint64_t)Cc * 2, (cuuint64_t)R * Cc * 2};
  cuuint32_t box[3] = {WCOLS, UROWS, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : EXG_EINVAL;
}

int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

bool plan(NceT& t, int S, int B, int T, int N, int Bc, unsigned& grid) {
  const long R = (long)B * T, Cc = (long)Bc * N;
  if (R >= (1L << 31) - 16 || Cc >= (1L << 31) - 512) return false;
  const int chunks = (int)((Cc + WCOLS - 1) / WCOLS);
  t.ncb = (chunks + WARPS - 1) / WARPS;
  t.k = (chunks + t.ncb - 1) / t.ncb;            // balanced chunks per column block (<= 16)
  t.g = WARPS / t.k;
  const long U = (R + UROWS - 1) / UROWS, tot = U * S * t.ncb;
  const long want = (tot + (long)t.g * 4 - 1) / ((long)t.g * 4);      // >= ~4 units per warp on small problems
  grid = (unsigned)(want < 1 ? 1 : (want > num_sms() ? num_sms() : want));
  return true;
}

bool a16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

bool enabled() {
  static const int off = getenv("EXG_NO_NCE_TMA") ? atoi(getenv("EXG_NO_NCE_TMA")) : 0;
  return !off;
}

}  // namespace

// Both return 1 when the TMA path took the call, 0 when the caller must use the generic kernels, < 0 / a CUDA error
// code on failure.  Outputs of the forward must have been zeroed by the caller.
int exg_nce_tma_fwd(const void* logits, int dtype, const uint32_t* bits, const uint8_t* col_valid,
                    const uint8_t* own_row_mask, float* row_all,
                    float* row_pos, float* col_all, float* col_pos, int S, int B, int T, int N, int Bc, int clip0,
                    float inv_temp, cudaStream_t st) {
  const long R = (long)B * T, Cc = (long)Bc * N;
  if (!enabled() || !bits || dtype != EXG_BF16 || Cc % 8 || !a16(logits)) return 0;
  NceT t = {};
  unsigned grid;
  if (!plan(t, S, B, T, N, Bc, grid)) return 0;
  CUtensorMap mL;
  if (logits_map(&mL, logits, S, R, Cc)) return 0;
  static bool attr = false;
  if (!attr) {
    EXG_CUDA(cudaFuncSetAttribute(nce_fwd_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    EXG_CUDA(cudaFuncSetAttribute(nce_bwd_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    attr = true;
  }
  t.bits = bits; t.W = (N + 31) / 32; t.col_valid = col_valid; t.own_row_mask = own_row_mask; t.row_all = row_all; t.row_pos = row_pos; t.col_all = col_all;
  t.col_pos = col_pos; t.S = S; t.B = B; t.T = T; t.N = N; t.inv_temp = inv_temp; t.Bc = Bc; t.clip0 = clip0;
  nce_fwd_tma_kernel<<<grid, NTHREADS, SMEM_BYTES, st>>>(mL, t);
  EXG_LAUNCH_CHECK("nce_fwd_tma");
  return 1;
}

int exg_nce_tma_bwd(const void* logits, int dtype, const uint32_t* bits, const uint8_t* col_valid,
                    const uint8_t* own_row_mask, const float* row_all, const float* row_pos, const float* col_all, const float* col_pos,
                    const float* wr, const float* wc, void* G, int g_dtype, int S, int B, int T, int N,
                    int Bc, int clip0, float inv_temp, cudaStream_t st) {
  const long R = (long)B * T, Cc = (long)Bc * N;
  if (!enabled() || !bits || dtype != EXG_BF16 || g_dtype != EXG_BF16 || Cc % 8 || !a16(logits) || !a16(G)) return 0;
  NceT t = {};
  unsigned grid;
  if (!plan(t, S, B, T, N, Bc, grid)) return 0;
  CUtensorMap mL, mG;
  if (logits_map(&mL, logits, S, R, Cc) || logits_map(&mG, G, S, R, Cc)) return 0;
  static bool attr = false;
  if (!attr) {
    EXG_CUDA(cudaFuncSetAttribute(nce_fwd_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    EXG_CUDA(cudaFuncSetAttribute(nce_bwd_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    attr = true;
  }
  t.bits = bits; t.W = (N + 31) / 32; t.col_valid = col_valid; t.own_row_mask = own_row_mask;
  t.row_all = const_cast<float*>(row_all); t.row_pos = const_cast<float*>(row_pos);
  t.col_all = const_cast<float*>(col_all); t.col_pos = const_cast<float*>(col_pos);
  t.wr = wr; t.wc = wc; t.S = S; t.B = B; t.T = T; t.N = N; t.inv_temp = inv_temp; t.Bc = Bc; t.clip0 = clip0;
  nce_bwd_tma_kernel<<<grid, NTHREADS, SMEM_BYTES, st>>>(mL, mG, t);
  EXG_LAUNCH_CHECK("nce_bwd_tma");
  return 1;
}

// Packs the [B, N, T] uint8 target mask into per-(clip, second) bit rows [B*T, ceil(N/32)] uint32, the layout the
// streaming kernels read one word per own row from (the byte layout costs one strided load per column).
extern "C" int exg_nce_pack_targets(const uint8_t* tgt, uint32_t* bits, int B, int T, int N, void* stream) {
  EXG_CHECK_ARG(tgt && bits && B > 0 && T > 0 && N > 0, "nce_pack_targets: bad argument");
  const int W = (N + 31) / 32;
  const long n = (long)B * T * W;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)B * T * N + (double)n * 4, stream);
  pack_targets_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(tgt, bits, B, T, N, W);
  EXG_LAUNCH_CHECK("nce_pack_targets");
  return 0;
}
