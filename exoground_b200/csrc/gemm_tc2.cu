// gemm_tc2.cu — persistent, warp-specialised bf16 GEMM for sm_100a.
//
//   warp 0      : TMA producer   (global -> 128B-swizzled smem ring, mbarrier complete_tx)
//   warp 1      : MMA issuer     (one elected lane, tcgen05.mma cta_group::1, 128 x BN x 16 per instruction)
//   warps 2..9  : epilogue       (tcgen05.ld TMEM -> registers -> bias / QuickGELU / ReLU -> 128B-swizzled smem
//                                 staging -> TMA store, or TMA reduce-add for split-K partial sums).  Two warps per
//                                 TMEM lane quarter (= per SM sub-partition) take alternate 64-column chunks, so each
//                                 scheduler has a second warp to issue from while one waits on TMEM / MUFU / TMA.
//
// With NCTA = 2 the kernel runs as clusters of two CTAs on the two SMs of a TPC (cta_group::2): the pair computes a
// 256 x BN tile, each CTA stages its own 128 rows of A and HALF of the B tile, the leader CTA issues
// tcgen05.mma.cta_group::2 for both, and each CTA's epilogue drains its own 128 accumulator lanes.  Per SM that is
// 2/3 of the L2 -> SM operand traffic of two independent 128 x BN tiles, which is what bounds these GEMMs.
// One CTA per SM loops over output tiles (static round-robin).  The fp32 accumulator is double-buffered in
// TMEM (2 x BN columns), so the epilogue of tile i overlaps the MMAs of tile i+1, and every global access is a
// TMA transfer (no per-thread global loads or stores in the kernel).  Operands may be K-major or MN-major in
// global memory (MN-major tiles are fetched as 64-wide column blocks and flagged in the instruction
// descriptor), so the backward GEMMs need no transposes; weight gradients (short M x N, long K) are split
// along K and accumulated in fp32 with TMA reduce-add.
#include "common.cuh"
#include "sm100.cuh"

using namespace sm100;

namespace {

constexpr int BM = 128, BK = 64, UK = 16;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 32 * (2 + kEpiWarps);
constexpr uint32_t kABytes = BM * BK * 2;          // 16 KB
constexpr int kOutBufs = 2;                        // output staging buffers per epilogue warp (plain epilogues): chunks are
                                                   // staged into them and stored by ONE fence + TMA-store group
constexpr int kAuxBufs = 3;                        // aux-in epilogues: ring of 3 buffers per warp; a buffer first receives an
                                                   // aux-in chunk (TMA, issued a whole chunk ahead), then stages the output
                                                   // chunk computed from it
constexpr uint32_t kStageBufBytes = 32 * 128;      // 32 rows x 128 B
constexpr uint32_t kBiasStripBytes = 128 * 4;      // per epilogue warp: the bias values of the columns it owns in a tile

struct Tc2Args {
  int M, N, K;
  int a_mn, b_mn;
  int out_f32, epilogue;
  const float* bias;
  float alpha;
  int tiles_m, tiles_n, splits, kblocks_per_split, batch;
  int a_bcast, b_bcast;
  int stages;        // depth of the operand ring (what fits beside the epilogue staging buffers)
  int has_aux_out;
  int has_aux_in;    // QGELU_BWD / RELU_BWD: bf16 aux tile streamed in by TMA, chunk by chunk, by the epilogue warps
  unsigned long long* trace;   // optional per-role event log of ONE CTA (exg_gemm_trace; scripts/gemm_trace.py)
  int trace_cap, trace_cta;
  int epi_bytes;               // shared memory of the epilogue region (staging buffers + bias strips)
  // 0: every staged chunk leaves with its own proxy fence + TMA store (the buffer it reuses was stored two chunks
  //    ago: a whole tile of slack for the store's shared-memory read, ~2000 cycles measured);
  // 1: the chunks of a tile leave as one fence + store group (fewer fences, but the next tile's first chunk then waits
  //    for stores issued a moment ago) - measured slower inside the train step (profiles/gemm_epilogue_r02.txt)
  int flush_mode;
  // development: 1 = the epilogue only waits for / releases the accumulator (no TMEM reads, math or stores): time of
  // the operand feed + MMA pipeline alone; 2 = no TMA loads and no MMAs (accumulator stages are handed over at once):
  // time of the epilogue + stores alone.  Results are garbage in both modes (scripts/bench_gemm.py ABLATE=...).
  int ablate;
};

// Event log for pipeline analysis without a profiler: role r (0 producer, 1 MMA issuer, 2..9 epilogue warps, 10 CTA)
// appends clock64() * 256 + tag to its own row of the buffer.  Costs one uniform predicate per event when off.
struct Tracer {
  unsigned long long* row; int cap, n;
  __device__ __forceinline__ Tracer(const Tc2Args& p, int role) {
    row = (p.trace && (int)blockIdx.x == p.trace_cta) ? p.trace + (size_t)role * p.trace_cap : nullptr;
    cap = p.trace_cap; n = 0;
  }
  __device__ __forceinline__ void ev(int tag) {
    if (row && n < cap) row[n++] = ((unsigned long long)clock64() << 8) | (unsigned)tag;
  }
};

constexpr int kMaxStages = 8;
constexpr uint32_t kMaxSmem = 232448;

template <int BN, int NCTA> struct Cfg2 {
  static constexpr uint32_t kBBytes = (BN / NCTA) * BK * 2;       // a CTA of a pair stages half of the B tile
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  __host__ __device__ static constexpr uint32_t epi_bytes(int aux_in) {
    return kEpiWarps * ((aux_in ? kAuxBufs : kOutBufs) * kStageBufBytes + kBiasStripBytes);
  }
  __host__ __device__ static constexpr int stages_for(uint32_t epi) {
    const int n = (int)((kMaxSmem - 1024 - 512 - epi) / kStageBytes);
    return n > kMaxStages ? kMaxStages : n;
  }
  __host__ __device__ static constexpr uint32_t smem_for(uint32_t epi) {
    return stages_for(epi) * kStageBytes + epi + 1024 /*align*/ + 512 /*barriers*/;
  }
};

struct TileCoord { int m0, n0, z, kb0, nkb; };

// n / d and n % d for 0 <= n < 2^24, 0 < d < 2^24 without the ~40-instruction integer division: every role of the CTA
// decodes a tile index once per tile, and on the single MMA-issuing thread that was a ~600-cycle bubble between tiles
__device__ __forceinline__ int fast_divmod(int n, int d, int& rem) {
  int q = __float2int_rz(__int2float_rz(n) * __frcp_rz(__int2float_rz(d)));
  int r = n - q * d;
  if (r >= d) { ++q; r -= d; }
  if (r < 0) { --q; r += d; }
  rem = r;
  return q;
}

__device__ __forceinline__ TileCoord tile_coord(const Tc2Args& p, int tile, int BN, int bm_tile, int cta_rank) {
  // order: n fastest, then m, then split, then batch -> CTAs running concurrently share A row blocks in L2
  TileCoord t;
  int tn, tm, sp;
  tile = fast_divmod(tile, p.tiles_n, tn);
  tile = fast_divmod(tile, p.tiles_m, tm);
  if (p.splits == 1) { sp = 0; t.z = tile; } else { t.z = fast_divmod(tile, p.splits, sp); }
  t.m0 = tm * bm_tile + cta_rank * BM; t.n0 = tn * BN;
  const int kbt = (p.K + BK - 1) >> 6;
  t.kb0 = sp * p.kblocks_per_split;
  const int kb1 = min(kbt, t.kb0 + p.kblocks_per_split);
  t.nkb = kb1 - t.kb0;
  return t;
}

template <int BN, int NCTA>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmAux,
                const __grid_constant__ CUtensorMap tmAuxIn, Tc2Args p) {
  using C_ = Cfg2<BN, NCTA>;
  const int cta_rank = NCTA == 2 ? (int)cluster_ctarank() : 0;
  const int unit = (int)blockIdx.x / NCTA, n_units = (int)gridDim.x / NCTA;     // CTA (pair) index / count
  const int n_stages = p.stages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* epi_smem = smem + (size_t)n_stages * C_::kStageBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(epi_smem + p.epi_bytes);
  uint64_t* empty = full + kMaxStages;
  uint64_t* tmem_full = empty + kMaxStages;      // [2]
  uint64_t* tmem_empty = tmem_full + 2;          // [2]
  uint64_t* aux_bar = tmem_empty + 2;            // [kEpiWarps][kAuxBufs]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux_bar + kAuxBufs * kEpiWarps);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.tiles_m * p.tiles_n * p.splits * p.batch;
  Tracer tr_cta(p, 10);
  if (threadIdx.x == 64) tr_cta.ev(11);

  if (threadIdx.x == 0) {
    for (int s = 0; s < n_stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tmem_full[s], 1); mbar_init(&tmem_empty[s], kEpiWarps * NCTA); }
    for (int s = 0; s < kAuxBufs * kEpiWarps; ++s) mbar_init(&aux_bar[s], 1);
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); tma_prefetch_desc(&tmC);
    if (p.has_aux_out) tma_prefetch_desc(&tmAux);
    if (p.has_aux_in) tma_prefetch_desc(&tmAuxIn);
  }
  if (warp == 1) { if (NCTA == 2) tmem_alloc_2sm(tmem_slot, 2 * BN); else tmem_alloc(tmem_slot, 2 * BN); }
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();      // the peer's barriers are initialised before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 64) tr_cta.ev(12);

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      Tracer tr(p, 0);
      int s = 0; uint32_t ph = 0;
      // a pair's loads all signal the LEADER's full barrier, which expects both CTAs' bytes
      uint32_t full_addr[kMaxStages];
#pragma unroll
      for (int i = 0; i < kMaxStages; ++i)
        full_addr[i] = NCTA == 2 ? mapa_u32(smem_u32(&full[i]), 0) : smem_u32(&full[i]);
      constexpr int BNL = BN / NCTA;              // B rows staged by this CTA
      for (int tile = unit; tile < ((p.ablate & 2) ? 0 : total_tiles); tile += n_units) {
        const TileCoord t = tile_coord(p, tile, BN, BM * NCTA, cta_rank);
        const int za = p.a_bcast ? 0 : t.z, zb = p.b_bcast ? 0 : t.z;
        const int nb0 = t.n0 + cta_rank * BNL;
        for (int i = 0; i < t.nkb; ++i) {
          mbar_wait(&empty[s], ph ^ 1);
          tr.ev(1);
          uint8_t* sa = smem + (size_t)s * C_::kStageBytes;
          uint8_t* sb = sa + kABytes;
          const int k0 = (t.kb0 + i) * BK;
          if (cta_rank == 0) mbar_arrive_expect_tx(&full[s], C_::kStageBytes * NCTA);
          auto load = [&](void* dst, const CUtensorMap* m, int c0, int c1, int c2) {
            if (NCTA == 2) tma_load_3d_2sm(dst, m, full_addr[s], c0, c1, c2);
            else tma_load_3d(dst, m, &full[s], c0, c1, c2);
          };
          if (!p.a_mn) {
            load(sa, &tmA, k0, t.m0, za);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) load(sa + j * 8192, &tmA, t.m0 + j * 64, k0, za);
          }
          if (!p.b_mn) {
            load(sb, &tmB, k0, nb0, zb);
          } else {
#pragma unroll
            for (int j = 0; j < BNL / 64; ++j) load(sb + j * 8192, &tmB, nb0 + j * 64, k0, zb);
          }
          if (++s == n_stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (the leader CTA of a pair issues for both) =====================
    if (lane == 0 && cta_rank == 0) {
      const uint32_t idesc = umma_idesc_bf16(BM * NCTA, BN, p.a_mn, p.b_mn);
      const uint32_t a_lbo = p.a_mn ? 8192u : 16u, b_lbo = p.b_mn ? 8192u : 16u;
      const uint32_t a_kstep = p.a_mn ? (UK * 128u) : (UK * 2u);
      const uint32_t b_kstep = p.b_mn ? (UK * 128u) : (UK * 2u);
      int s = 0; uint32_t ph = 0, ti = 0;
      Tracer tr(p, 1);
      for (int tile = unit; tile < total_tiles; tile += n_units, ++ti) {
        const TileCoord t = tile_coord(p, tile, BN, BM * NCTA, cta_rank);
        const uint32_t as = ti & 1, aph = (ti >> 1) & 1;
        mbar_wait(&tmem_empty[as], aph ^ 1);          // every epilogue warp (of both CTAs) has drained this stage
        tr.ev(2);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int i = 0; i < ((p.ablate & 2) ? 0 : t.nkb); ++i) {
          mbar_wait(&full[s], ph);
          tr.ev(3);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)s * C_::kStageBytes);
          const uint32_t sb = sa + kABytes;
#pragma unroll
          for (int k = 0; k < BK / UK; ++k) {
            const uint64_t ad = umma_smem_desc(sa + k * a_kstep, a_lbo, 1024u);
            const uint64_t bd = umma_smem_desc(sb + k * b_kstep, b_lbo, 1024u);
            if (NCTA == 2) umma_bf16_2sm(d_tmem, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
            else umma_bf16(d_tmem, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
          if (NCTA == 2) umma_commit_2sm(&empty[s], 3); else umma_commit(&empty[s]);
          if (++s == n_stages) { s = 0; ph ^= 1; }
        }
        if (NCTA == 2) umma_commit_2sm(&tmem_full[as], 3); else umma_commit(&tmem_full[as]);
        tr.ev(4);
      }
    }
  } else {
    // ===================== epilogue =====================
    const int ew = warp - 2;                      // staging slot
    Tracer tr(p, 2 + ew);
    if (lane != 0) tr.row = nullptr;
    const int lg = warp & 3;                      // TMEM lane group this warp may access
    const int half = ew >> 2;                     // which of the two interleaved chunk sequences this warp takes
    const uint32_t n_bufs = p.has_aux_in ? kAuxBufs : kOutBufs;
    uint8_t* stage = epi_smem + (size_t)ew * n_bufs * kStageBufBytes;
    float* bias_s = reinterpret_cast<float*>(epi_smem + p.epi_bytes - kEpiWarps * kBiasStripBytes) + ew * 128;
    uint32_t ti = 0;
    const int cols_per_chunk = p.out_f32 ? 32 : 64;
    const int c_first = half * cols_per_chunk, c_step = 2 * cols_per_chunk;
    uint64_t* my_aux_bar = aux_bar + kAuxBufs * ew;
    uint32_t ci = 0;                              // running chunk index of this warp (aux ring position)
    const uint32_t tmem_empty_addr[2] = {NCTA == 2 ? mapa_u32(smem_u32(&tmem_empty[0]), 0) : 0u,
                                         NCTA == 2 ? mapa_u32(smem_u32(&tmem_empty[1]), 0) : 0u};
    const bool use_alpha = p.alpha != 1.0f;
    // this lane's float4 of the warp's bias strip: strip element e = 4 * lane covers chunk e / cols_per_chunk
    const int b_chunk = (lane * 4) / cols_per_chunk, b_within = (lane * 4) % cols_per_chunk;
    const int b_col = c_first + b_chunk * c_step + b_within;
    // plain epilogues: chunks are staged into the warp's buffers and leave as ONE fence + store group per
    // kOutBufs buffers (one per tile for bf16 outputs), instead of a proxy fence per chunk
    uint32_t pend = 0, sbuf = 0;
    static_assert(kOutBufs == 2, "the pending-store bookkeeping below is written for two buffers");
    int pend_x0 = 0, pend_x1 = 0, pend_k0 = 0, pend_k1 = 0;    // kind: 0 store C, 1 reduce-add C, 2 store aux-out
    auto issue = [&](const uint8_t* b, int kind, int x, int row0, int z) {
      if (kind == 2) tma_store_3d(&tmAux, b, x, row0, z);
      else if (kind == 1) tma_reduce_add_3d(&tmC, b, x, row0, z);
      else tma_store_3d(&tmC, b, x, row0, z);
    };
    auto flush = [&](int row0, int z) {
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        issue(stage, pend_k0, pend_x0, row0, z);
        if (pend > 1) issue(stage + kStageBufBytes, pend_k1, pend_x1, row0, z);
        bulk_commit();
      }
      pend = 0;
    };
    auto note = [&](int x, int kind) {
      if (pend == 0) { pend_x0 = x; pend_k0 = kind; } else { pend_x1 = x; pend_k1 = kind; }
      ++pend;
    };
    auto stage_bf16 = [&](uint8_t* buf, const uint32_t (&r)[64]) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        uint4 q;
        q.x = pack_bf16x2(__uint_as_float(r[u * 8 + 0]), __uint_as_float(r[u * 8 + 1]));
        q.y = pack_bf16x2(__uint_as_float(r[u * 8 + 2]), __uint_as_float(r[u * 8 + 3]));
        q.z = pack_bf16x2(__uint_as_float(r[u * 8 + 4]), __uint_as_float(r[u * 8 + 5]));
        q.w = pack_bf16x2(__uint_as_float(r[u * 8 + 6]), __uint_as_float(r[u * 8 + 7]));
        *reinterpret_cast<uint4*>(buf + lane * 128 + ((u ^ (lane & 7)) << 4)) = q;
      }
    };
    if (p.has_aux_in && lane == 0 && unit < total_tiles) {
      const TileCoord t0 = tile_coord(p, unit, BN, BM * NCTA, cta_rank);
      mbar_arrive_expect_tx(&my_aux_bar[0], kStageBufBytes);
      tma_load_3d(stage, &tmAuxIn, &my_aux_bar[0], t0.n0 + c_first, t0.m0 + lg * 32, t0.z);
    }
    for (int tile = unit; tile < total_tiles; tile += n_units, ++ti) {
      const TileCoord t = tile_coord(p, tile, BN, BM * NCTA, cta_rank);
      const uint32_t as = ti & 1, aph = (ti >> 1) & 1;
      // the tile's bias values are requested BEFORE the wait for the accumulator (their L2 latency hides behind it) and
      // parked in shared memory; the math below reads them as warp-wide broadcasts
      float4 bpre = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p.bias && b_col < BN && t.n0 + b_col < p.N) bpre = __ldg(reinterpret_cast<const float4*>(p.bias + t.n0 + b_col));
      mbar_wait(&tmem_full[as], aph);
      tr.ev(5);
      tc_fence_after();
      if (p.bias) {
        __syncwarp();                                     // the previous tile's reads of the strip are done
        *reinterpret_cast<float4*>(bias_s + lane * 4) = bpre;
        __syncwarp();
      }
      const uint32_t taddr = tmem_base + as * BN + ((uint32_t)(lg * 32) << 16);
      const int row0 = t.m0 + lg * 32;
      const bool rows_live = row0 < p.M;
      const int n_live = min(BN, p.N - t.n0);
      int chunk_k = 0;                                    // index of the chunk within this warp's strip
      if (p.ablate & 1) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (NCTA == 2) mbar_arrive_remote(tmem_empty_addr[as]); else mbar_arrive(&tmem_empty[as]); }
        continue;
      }
#pragma unroll 1
      for (int c0 = c_first; c0 < BN; c0 += c_step, ++chunk_k) {
        const bool live = rows_live && c0 < n_live;      // warp-uniform
        const bool last_chunk = c0 + c_step >= BN;
        uint32_t r[64];
        uint4 ax[8];
        if (p.has_aux_in) {
          // prefetch the aux chunk after this one (same tile, or first chunk of this CTA's next tile) into the ring
          // buffer after this one: its last user is the store issued two chunks ago
          if (lane == 0) {
            int nt = tile, nc = c0 + c_step;
            if (nc >= BN) { nt = tile + n_units; nc = c_first; }
            if (nt < total_tiles) {
              const TileCoord tn = tile_coord(p, nt, BN, BM * NCTA, cta_rank);
              const uint32_t nb = (ci + 1) % kAuxBufs;
              bulk_wait_read<1>();       // only the previous chunk's store may still be reading (another buffer)
              mbar_arrive_expect_tx(&my_aux_bar[nb], kStageBufBytes);
              tma_load_3d(stage + nb * kStageBufBytes, &tmAuxIn, &my_aux_bar[nb], tn.n0 + nc,
                          tn.m0 + lg * 32, tn.z);
            }
          }
          // ... and take this chunk's aux rows (thread = row, 128B-swizzled) into registers
          const uint32_t cb = ci % kAuxBufs;
          mbar_wait(&my_aux_bar[cb], (ci / kAuxBufs) & 1);
          const uint8_t* abuf = stage + cb * kStageBufBytes;
#pragma unroll
          for (int u = 0; u < 8; ++u)
            ax[u] = *reinterpret_cast<const uint4*>(abuf + lane * 128 + ((u ^ (lane & 7)) << 4));
          ++ci;
        }
        if (live) {
          tmem_ld_32x32(taddr + c0, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
          if (!p.out_f32) tmem_ld_32x32(taddr + c0 + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
          tmem_ld_wait();
        }
        tr.ev(6);
        if (last_chunk) {
          // all TMEM reads of this tile are done: hand the accumulator stage back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) { if (NCTA == 2) mbar_arrive_remote(tmem_empty_addr[as]); else mbar_arrive(&tmem_empty[as]); }
        }
        if (!live) {
          if (last_chunk && pend) flush(row0, t.z);
          continue;
        }
        // ---- math in the thread = row layout
        const int ncols = cols_per_chunk;
        const int nbase = t.n0 + c0;
        if (use_alpha) {
#pragma unroll
          for (int j = 0; j < 64; ++j) if (j < ncols) r[j] = __float_as_uint(__uint_as_float(r[j]) * p.alpha);
        }
        if (p.bias) {
          const float* bs = bias_s + chunk_k * cols_per_chunk;
#pragma unroll
          for (int j = 0; j < 64; j += 4) {
            if (j >= ncols) break;
            const float4 b = *reinterpret_cast<const float4*>(bs + j);         // same address in every lane: broadcast
            r[j] = __float_as_uint(__uint_as_float(r[j]) + b.x);
            r[j + 1] = __float_as_uint(__uint_as_float(r[j + 1]) + b.y);
            r[j + 2] = __float_as_uint(__uint_as_float(r[j + 2]) + b.z);
            r[j + 3] = __float_as_uint(__uint_as_float(r[j + 3]) + b.w);
          }
        }
        // ---- optional second output: the pre-activation (bf16), staged first
        if (p.has_aux_out) {
          if (p.flush_mode == 0) {
            if (lane == 0) bulk_wait_read<kOutBufs - 1>();      // the store that last used this buffer is done reading
            __syncwarp();
            stage_bf16(stage + (size_t)(sbuf & 1) * kStageBufBytes, r);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) { tma_store_3d(&tmAux, stage + (size_t)(sbuf & 1) * kStageBufBytes, nbase, row0, t.z); bulk_commit(); }
            ++sbuf;
          } else {
            if (pend == 0) { if (lane == 0) bulk_wait_read<0>(); __syncwarp(); }
            stage_bf16(stage + (size_t)pend * kStageBufBytes, r);
            note(nbase, 2);
            if (pend == kOutBufs) flush(row0, t.z);
          }
        }
        if (p.epilogue == EXG_EPI_QGELU) {
#pragma unroll
          for (int j = 0; j < 64; ++j) if (j < ncols) r[j] = __float_as_uint(quick_gelu_fast(__uint_as_float(r[j])));
        } else if (p.epilogue == EXG_EPI_RELU) {
#pragma unroll
          for (int j = 0; j < 64; ++j) if (j < ncols) r[j] = __float_as_uint(fmaxf(__uint_as_float(r[j]), 0.f));
        } else if (p.epilogue == EXG_EPI_QGELU_BWD || p.epilogue == EXG_EPI_RELU_BWD) {
          const bool gelu = p.epilogue == EXG_EPI_QGELU_BWD;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t w[4] = {ax[u].x, ax[u].y, ax[u].z, ax[u].w};
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              const float a0 = __uint_as_float(w[h] << 16), a1 = __uint_as_float(w[h] & 0xffff0000u);
              const int j = u * 8 + h * 2;
              const float v0 = __uint_as_float(r[j]), v1 = __uint_as_float(r[j + 1]);
              r[j] = __float_as_uint(gelu ? v0 * quick_gelu_grad_fast(a0) : (a0 > 0.f ? v0 : 0.f));
              r[j + 1] = __float_as_uint(gelu ? v1 * quick_gelu_grad_fast(a1) : (a1 > 0.f ? v1 : 0.f));
            }
          }
        }
        tr.ev(7);
        if (p.has_aux_in) {
          // stage into the buffer this chunk's aux rows came in (they are in registers) and store it right away
          uint8_t* buf = stage + (size_t)((ci - 1) % kAuxBufs) * kStageBufBytes;
          __syncwarp();                   // every lane has taken its aux row out of the buffer
          tr.ev(8);
          stage_bf16(buf, r);
          fence_proxy_async();            // also orders the generic reads of the aux rows before later TMA writes
          __syncwarp();
          if (lane == 0) { tma_store_3d(&tmC, buf, nbase, row0, t.z); bulk_commit(); }
          tr.ev(9);
          continue;
        }
        // ---- stage (128B-swizzled rows) and store
        if (p.ablate & 8) { tr.ev(9); continue; }
        if (p.flush_mode == 0) {
          if (lane == 0) bulk_wait_read<kOutBufs - 1>();
          __syncwarp();
        } else if (pend == 0) {
          if (lane == 0) bulk_wait_read<0>();
          __syncwarp();
        }
        tr.ev(8);
        uint8_t* buf = stage + (size_t)(p.flush_mode == 0 ? (sbuf & 1) : pend) * kStageBufBytes;
        if (p.out_f32) {
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint4 q = make_uint4(r[u * 4], r[u * 4 + 1], r[u * 4 + 2], r[u * 4 + 3]);
            *reinterpret_cast<uint4*>(buf + lane * 128 + ((u ^ (lane & 7)) << 4)) = q;
          }
        } else {
          stage_bf16(buf, r);
        }
        if (p.ablate & 4) { ++sbuf; tr.ev(9); continue; }
        if (p.flush_mode == 0) {
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) { issue(buf, p.splits > 1 ? 1 : 0, nbase, row0, t.z); bulk_commit(); }
          ++sbuf;
        } else {
          note(nbase, p.splits > 1 ? 1 : 0);
          if (pend == kOutBufs || last_chunk) flush(row0, t.z);
        }
        tr.ev(9);
      }
      tr.ev(10);
    }
    if (lane == 0) bulk_wait_all();
    tr.ev(15);
  }
  if (threadIdx.x == 64) tr_cta.ev(13);
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();      // the peer may still be reading this CTA's smem / signalling its barriers
  if (warp == 1) {
    tc_fence_after();
    if (NCTA == 2) tmem_dealloc_2sm(tmem_base, 2 * BN); else tmem_dealloc(tmem_base, 2 * BN);
  }
  if (threadIdx.x == 64) tr_cta.ev(14);
}

// ---------------------------------------------------------------- host side
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

// 3-D tensor map over [batch][rows][inner] with a {box_inner, box_rows, 1} box, 128B swizzle.
int make_map(CUtensorMap* out, const void* base, int f32, uint64_t inner, uint64_t rows, uint64_t batch,
             uint64_t row_stride_elems, uint64_t batch_stride_elems, uint32_t box_inner, uint32_t box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { exg_set_error("gemm_tc2: cuTensorMapEncodeTiled unavailable"); return EXG_EUNSUPPORTED; }
  const uint64_t es = f32 ? 4 : 2;
  cuuint64_t dims[3] = {inner, rows, batch};
  cuuint64_t strides[2] = {row_stride_elems * es, (batch > 1 ? batch_stride_elems : rows * row_stride_elems) * es};
  cuuint32_t box[3] = {box_inner, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    exg_set_error("gemm_tc2: cuTensorMapEncodeTiled failed (%d) inner=%llu rows=%llu batch=%llu ld=%llu",
                  (int)r, (unsigned long long)inner, (unsigned long long)rows, (unsigned long long)batch,
                  (unsigned long long)row_stride_elems);
    return EXG_EINVAL;
  }
  return 0;
}

template <int BN, int NCTA>
int launch2(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tx,
            const CUtensorMap& ti, Tc2Args a, cudaStream_t s) {
  using C_ = Cfg2<BN, NCTA>;
  static bool attr_set = false;
  auto kfn = gemm_tc2_kernel<BN, NCTA>;
  a.epi_bytes = (int)C_::epi_bytes(a.has_aux_in);
  a.stages = C_::stages_for((uint32_t)a.epi_bytes);
  const uint32_t smem = C_::smem_for((uint32_t)a.epi_bytes);
  if (a.stages < 2 || smem > kMaxSmem) {
    exg_set_error("gemm_tc2: %u bytes of shared memory (BN=%d NCTA=%d aux_in=%d)", smem, BN, NCTA, a.has_aux_in);
    return EXG_EINVAL;
  }
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmem);
    if (e != cudaSuccess) { exg_set_error("gemm_tc2: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr_set = true;
  }
  const long total = (long)a.tiles_m * a.tiles_n * a.splits * a.batch;
  const long units = exg_persistent_sms() / NCTA;
  const int grid = (int)(total < units ? total : units) * NCTA;
  if (NCTA == 1) {
    kfn<<<grid, kThreads, smem, s>>>(ta, tb, tc, tx, ti, a);
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = NCTA; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kfn, ta, tb, tc, tx, ti, a);
    if (e != cudaSuccess) { exg_set_error("gemm_tc2: cluster launch failed: %s", cudaGetErrorString(e)); return (int)e; }
  }
  EXG_LAUNCH_CHECK("gemm_tc2");
  return 0;
}

bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

unsigned long long* g_trace = nullptr;
int g_trace_cap = 0, g_trace_cta = 0;

double wave_eff(long tiles) {
  const long sms = exg_num_sms();
  const long waves = (tiles + sms - 1) / sms;
  return (double)tiles / (double)(waves * sms);
}

}  // namespace

// Pipeline event log of the persistent GEMM (development aid, scripts/gemm_trace.py): while `buf` (device, 11 rows of
// `cap` uint64) is installed, CTA `cta` of every gemm_tc2 launch appends clock64() * 256 + tag per role.  buf = null
// switches it off.  Tags: 1 producer slot free, 2 MMA accumulator free, 3 MMA operands landed, 4 tile committed,
// 5 epilogue accumulator ready, 6 TMEM read done, 7 math done, 8 staging buffer free, 9 store issued, 10 tile done,
// 11 kernel entry, 12 setup done, 13 roles done, 14 teardown done, 15 stores drained.
extern "C" int exg_gemm_trace(void* buf, int cap, int cta) {
  g_trace = reinterpret_cast<unsigned long long*>(buf);
  g_trace_cap = cap; g_trace_cta = cta;
  return 0;
}

// 1 if the persistent TMA-epilogue kernel takes these parameters.
int exg_gemm_tc2_supported(const exg_gemm_params* p) {
  if (!p || p->in_dtype != EXG_BF16) return 0;
  if (p->M <= 0 || p->N <= 0 || p->K <= 0 || p->batch <= 0) return 0;
  const bool aux_in = p->epilogue == EXG_EPI_QGELU_BWD || p->epilogue == EXG_EPI_RELU_BWD;
  if (p->epilogue != EXG_EPI_NONE && p->epilogue != EXG_EPI_QGELU && p->epilogue != EXG_EPI_RELU && !aux_in) return 0;
  if (aux_in) {   // bf16 aux tile arrives by TMA: bf16 output, 16-byte aligned aux rows
    if (!p->aux || p->out_dtype != EXG_BF16 || !al16(p->aux) || p->ldaux % 8 ||
        (p->batch > 1 && p->strideAux % 8)) return 0;
  } else if (p->aux) return 0;
  if (p->aux_out && (p->epilogue != EXG_EPI_QGELU || p->out_dtype != EXG_BF16)) return 0;
  const int es_out = p->out_dtype == EXG_F32 ? 4 : 2;
  if (p->N < 128 / es_out) return 0;                       // the store box is 128 bytes wide
  if (p->N % 8 != 0) return 0;
  if (!al16(p->A) || !al16(p->B) || !al16(p->C) || (p->aux_out && !al16(p->aux_out))) return 0;
  if (p->lda % 8 || p->ldb % 8 || (p->ldc * es_out) % 16) return 0;
  if (p->batch > 1 && (p->strideA % 8 || p->strideB % 8 || (p->strideC * es_out) % 16)) return 0;
  if (p->bias && !al16(p->bias)) return 0;
  if ((p->transA ? p->M : p->K) < 8 || (p->transB ? p->N : p->K) < 8) return 0;
  return 1;
}

int exg_gemm_tc2(const exg_gemm_params* q, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  Tc2Args a;
  a.M = q->M; a.N = q->N; a.K = q->K;
  a.a_mn = q->transA ? 1 : 0; a.b_mn = q->transB ? 1 : 0;
  a.out_f32 = q->out_dtype == EXG_F32; a.epilogue = q->epilogue;
  a.bias = q->bias; a.alpha = q->alpha; a.batch = q->batch;
  a.has_aux_out = q->aux_out ? 1 : 0;
  a.has_aux_in = (q->epilogue == EXG_EPI_QGELU_BWD || q->epilogue == EXG_EPI_RELU_BWD) ? 1 : 0;
  a.trace = g_trace; a.trace_cap = g_trace_cap; a.trace_cta = g_trace_cta;
  static const int flush_env = getenv("EXG_GEMM_FLUSH") ? atoi(getenv("EXG_GEMM_FLUSH")) : 0;
  a.flush_mode = flush_env;
  static const int ablate_env = getenv("EXG_GEMM_ABLATE") ? atoi(getenv("EXG_GEMM_ABLATE")) : 0;
  a.ablate = ablate_env;

  const int kblocks = (q->K + BK - 1) / BK;
  const long tm1 = (q->M + BM - 1) / BM;
  // tile width: 256 everywhere it exists -- measured (scripts/bench_gemm.py, EXG_GEMM_BN / EXG_GEMM_NCTA sweeps) the
  // 128-wide tile loses on every shape of the step even where it fills the last wave better: these GEMMs are bound
  // by L2 -> SM operand bytes (48 KB per k-block for 128 x 256 vs 2 x 32 KB for two 128 x 128 tiles)
  int bn = q->N <= 128 ? 128 : 256;
  // CTA pairs (256 x BN tiles over two SMs): two thirds of the operand bytes per flop.  Round 1 kept pairs away from
  // epilogues with bias / activation / aux traffic and from short split-K items (the lock-step of the two epilogues
  // cost more than the operand bytes saved); with the CTA-scope peer arrive, the smem bias strip and the 3-deep aux
  // ring that is no longer so: in the train step pairs EVERYWHERE (M > 128) beat both the selective policy and no
  // pairs - 7.106 vs 7.124 vs 7.359 ms per step, GEMM class 5.48 / 5.55 / 6.03 ms (run r02p).
  static const int no_pair = getenv("EXG_NO_GEMM_PAIR") ? atoi(getenv("EXG_NO_GEMM_PAIR")) : 0;
  const bool splittable = q->epilogue == EXG_EPI_NONE && !q->bias && q->out_dtype == EXG_F32 && kblocks >= 8;
  int ncta = (!no_pair && q->M > BM) ? 2 : 1;
  {  // tuning overrides (scripts/bench_gemm.py)
    static const int f_bn = getenv("EXG_GEMM_BN") ? atoi(getenv("EXG_GEMM_BN")) : 0;
    static const int f_ncta = getenv("EXG_GEMM_NCTA") ? atoi(getenv("EXG_GEMM_NCTA")) : 0;
    if (f_ncta == 1 || (f_ncta == 2 && q->M > BM)) ncta = f_ncta;
    if ((f_bn == 128 || f_bn == 256) && q->N > 128) bn = f_bn;
  }
  const long tm = (q->M + BM * ncta - 1) / (BM * ncta);
  a.tiles_m = (int)tm; a.tiles_n = (q->N + bn - 1) / bn;
  const long tiles = tm * a.tiles_n * q->batch;
  const long units = exg_num_sms() / ncta;
  a.splits = 1;
  if (splittable && tiles * 2 <= units) {
    // pick the split count with the shortest critical path: rounds x (k-blocks per item + fixed per-item cost)
    const int maxs = kblocks / 4 < 64 ? kblocks / 4 : 64;
    double best = 1e30;
    for (int sp = 1; sp <= maxs; ++sp) {
      const long kb = (kblocks + sp - 1) / sp;
      const long items = tiles * ((kblocks + kb - 1) / kb);
      const long rounds = (items + units - 1) / units;
      // + a per-split term: every extra split adds one more fp32 reduce-add pass over the output tile in L2
      static const double split_pen = getenv("EXG_GEMM_SPLIT_PEN") ? atof(getenv("EXG_GEMM_SPLIT_PEN")) : 0.0;
      const double cost = (double)rounds * ((double)kb + 6.0) + split_pen * (double)sp;
      if (cost < best - 1e-9) { best = cost; a.splits = sp; }
    }
  }
  a.kblocks_per_split = (kblocks + a.splits - 1) / a.splits;
  a.splits = (kblocks + a.kblocks_per_split - 1) / a.kblocks_per_split;
  if (a.splits > 1) {
    for (int z = 0; z < q->batch; ++z)
      EXG_CUDA(cudaMemset2DAsync(reinterpret_cast<float*>(q->C) + (size_t)z * q->strideC, (size_t)q->ldc * 4, 0,
                                 (size_t)q->N * 4, (size_t)q->M, s));
  }

  CUtensorMap ta, tb, tc, tx, ti;
  int rc;
  a.a_bcast = (q->batch > 1 && q->strideA == 0) ? 1 : 0;
  a.b_bcast = (q->batch > 1 && q->strideB == 0) ? 1 : 0;
  const int ba = a.a_bcast ? 1 : q->batch, bb = a.b_bcast ? 1 : q->batch;
  if (!q->transA) rc = make_map(&ta, q->A, 0, q->K, q->M, ba, q->lda, q->strideA, 64, BM);
  else            rc = make_map(&ta, q->A, 0, q->M, q->K, ba, q->lda, q->strideA, 64, BK);
  if (rc) return rc;
  if (!q->transB) rc = make_map(&tb, q->B, 0, q->K, q->N, bb, q->ldb, q->strideB, 64, bn / ncta);
  else            rc = make_map(&tb, q->B, 0, q->N, q->K, bb, q->ldb, q->strideB, 64, BK);
  if (rc) return rc;
  rc = make_map(&tc, q->C, a.out_f32, q->N, q->M, q->batch, q->ldc, q->strideC, a.out_f32 ? 32 : 64, 32);
  if (rc) return rc;
  tx = tc;
  if (q->aux_out) {
    rc = make_map(&tx, q->aux_out, 0, q->N, q->M, q->batch, q->ldc, q->strideC, 64, 32);
    if (rc) return rc;
  }
  ti = tc;
  if (a.has_aux_in) {
    rc = make_map(&ti, q->aux, 0, q->N, q->M, q->batch, q->ldaux, q->strideAux, 64, 32);
    if (rc) return rc;
  }
  if (ncta == 2) return bn == 256 ? launch2<256, 2>(ta, tb, tc, tx, ti, a, s) : launch2<128, 2>(ta, tb, tc, tx, ti, a, s);
  return bn == 256 ? launch2<256, 1>(ta, tb, tc, tx, ti, a, s) : launch2<128, 1>(ta, tb, tc, tx, ti, a, s);
}
