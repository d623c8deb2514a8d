// attention.cu — flash-style fused attention with key-padding mask, forward and backward, fp32 math.
// softmax(q k^T * scale + mask) v per (batch, head); never materialises the Sq x Sk score matrix in
// global memory.  head_dim is fixed at 64 (width 512 / 8 heads: model/tfm_model.py:21,62,64).
// This SIMT kernel is the fp32 path and the reference for the tensor-core variant (attention_tc.cu).
//
// All 64x64 products share one helper: acc[i][j] += sum_c A[c][ty*8+i] * B[c][tx*4+j] with both
// operands stored contraction-major in shared memory (pitch 68 floats -> aligned float4 reads).
#include "common.cuh"
#include <math.h>

namespace {

constexpr int BT = 64;          // tile edge (queries, keys) and head_dim
constexpr int PITCH = 68;
constexpr int TILE = BT * PITCH;   // floats per smem matrix
constexpr int NTHREADS = 128;

__device__ __forceinline__ void mm64(float (&acc)[8][4], const float* __restrict__ A,
                                     const float* __restrict__ B, int ty, int tx) {
#pragma unroll 8
  for (int c = 0; c < BT; ++c) {
    const float4 a0 = *reinterpret_cast<const float4*>(A + c * PITCH + ty * 8);
    const float4 a1 = *reinterpret_cast<const float4*>(A + c * PITCH + ty * 8 + 4);
    const float4 b = *reinterpret_cast<const float4*>(B + c * PITCH + tx * 4);
    const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
  }
}

// Load a 64x64 tile (rows row0.., head columns) into natural [r][c] and/or transposed [c][r] smem.
__device__ __forceinline__ void load_tile(float* __restrict__ nat, float* __restrict__ trn,
                                          const void* __restrict__ src, int dtype, long ld, long base_row,
                                          int row0, int rows_total, int col0) {
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int idx = it * NTHREADS + threadIdx.x;
    const int r = idx >> 4, c4 = (idx & 15) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < rows_total) v = load4(src, dtype, (base_row + row0 + r) * ld + col0 + c4);
    if (nat) *reinterpret_cast<float4*>(nat + r * PITCH + c4) = v;
    if (trn) {
      trn[(c4 + 0) * PITCH + r] = v.x; trn[(c4 + 1) * PITCH + r] = v.y;
      trn[(c4 + 2) * PITCH + r] = v.z; trn[(c4 + 3) * PITCH + r] = v.w;
    }
  }
}

__device__ __forceinline__ float half_warp_max(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float half_warp_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct AttnArgs {
  const void *q, *k, *v, *o, *dout;
  void *out, *dq, *dk, *dv;
  int dtype;
  long ldq, ldk, ldv, ldo;
  const uint8_t* key_pad;
  float* lse; float* delta;
  int B, H, Sq, Sk;
  float scale;
};

// ------------------------------------------------------------------ forward
__global__ void __launch_bounds__(NTHREADS) attn_fwd_kernel(AttnArgs p) {
  extern __shared__ float sm[];
  float* Qt = sm; float* Kt = sm + TILE; float* Vn = sm + 2 * TILE; float* Pt = sm + 3 * TILE;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * BT;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  load_tile(nullptr, Qt, p.q, p.dtype, p.ldq, (long)b * p.Sq, q0, p.Sq, h * BT);
  float m[8], l[8], acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    m[i] = -INFINITY; l[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  }
  for (int k0 = 0; k0 < p.Sk; k0 += BT) {
    __syncthreads();   // previous tile fully consumed (also orders the Q stores on the first pass)
    load_tile(nullptr, Kt, p.k, p.dtype, p.ldk, (long)b * p.Sk, k0, p.Sk, h * BT);
    load_tile(Vn, nullptr, p.v, p.dtype, p.ldv, (long)b * p.Sk, k0, p.Sk, h * BT);
    __syncthreads();
    float s[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) s[i][j] = 0.f;
    mm64(s, Qt, Kt, ty, tx);
    bool ok[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int kk = k0 + tx * 4 + j;
      ok[j] = kk < p.Sk && !(p.key_pad && p.key_pad[(long)b * p.Sk + kk]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s[i][j] = ok[j] ? s[i][j] * p.scale : -INFINITY;
        mx = fmaxf(mx, s[i][j]);
      }
      mx = half_warp_max(mx);
      const float mnew = fmaxf(m[i], mx);
      const float msafe = mnew == -INFINITY ? 0.f : mnew;
      const float corr = __expf(m[i] - msafe);       // m = -inf -> 0
      float rs = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s[i][j] = __expf(s[i][j] - msafe);
        rs += s[i][j];
      }
      rs = half_warp_sum(rs);
      l[i] = l[i] * corr + rs;
      m[i] = mnew;
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] *= corr;
    }
    // P^T to smem: Pt[key j][query i]
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float* dst = Pt + (tx * 4 + j) * PITCH + ty * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(s[0][j], s[1][j], s[2][j], s[3][j]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(s[4][j], s[5][j], s[6][j], s[7][j]);
    }
    __syncthreads();
    mm64(acc, Pt, Vn, ty, tx);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int qi = q0 + ty * 8 + i;
    if (qi >= p.Sq) continue;
    const float inv = 1.0f / l[i];
    const float4 o = make_float4(acc[i][0] * inv, acc[i][1] * inv, acc[i][2] * inv, acc[i][3] * inv);
    store4(p.out, p.dtype, ((long)b * p.Sq + qi) * p.ldo + h * BT + tx * 4, o);
    if (tx == 0 && p.lse) p.lse[((long)b * p.H + h) * p.Sq + qi] = m[i] + __logf(l[i]);
  }
}

// ------------------------------------------------------------------ backward: dQ (+ delta)
__global__ void __launch_bounds__(NTHREADS) attn_bwd_dq_kernel(AttnArgs p) {
  extern __shared__ float sm[];
  float* Qt = sm; float* dOt = sm + TILE; float* Kt = sm + 2 * TILE; float* Kn = sm + 3 * TILE;
  float* Vt = sm + 4 * TILE; float* dSt = sm + 5 * TILE;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * BT;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  load_tile(nullptr, Qt, p.q, p.dtype, p.ldq, (long)b * p.Sq, q0, p.Sq, h * BT);
  load_tile(nullptr, dOt, p.dout, p.dtype, p.ldo, (long)b * p.Sq, q0, p.Sq, h * BT);
  // delta_i = <dO_i, O_i>: 16 lanes share a row, each owns 4 of the 64 columns
  float lse[8], dl[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int qi = q0 + ty * 8 + i;
    float d = 0.f;
    lse[i] = 0.f;
    if (qi < p.Sq) {
      const long off = ((long)b * p.Sq + qi) * p.ldo + h * BT + tx * 4;
      const float4 a = load4(p.dout, p.dtype, off), o = load4(p.o, p.dtype, off);
      d = a.x * o.x + a.y * o.y + a.z * o.z + a.w * o.w;
      lse[i] = p.lse[((long)b * p.H + h) * p.Sq + qi];
    }
    dl[i] = half_warp_sum(d);
    if (tx == 0 && qi < p.Sq) p.delta[((long)b * p.H + h) * p.Sq + qi] = dl[i];
  }
  float dq[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) dq[i][j] = 0.f;
  for (int k0 = 0; k0 < p.Sk; k0 += BT) {
    __syncthreads();
    load_tile(Kn, Kt, p.k, p.dtype, p.ldk, (long)b * p.Sk, k0, p.Sk, h * BT);
    load_tile(nullptr, Vt, p.v, p.dtype, p.ldv, (long)b * p.Sk, k0, p.Sk, h * BT);
    __syncthreads();
    float s[8][4], dp[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) { s[i][j] = 0.f; dp[i][j] = 0.f; }
    mm64(s, Qt, Kt, ty, tx);
    mm64(dp, dOt, Vt, ty, tx);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int kk = k0 + tx * 4 + j;
      const bool ok = kk < p.Sk && !(p.key_pad && p.key_pad[(long)b * p.Sk + kk]);
      float ds[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float pr = ok ? __expf(s[i][j] * p.scale - lse[i]) : 0.f;
        ds[i] = pr * (dp[i][j] - dl[i]) * p.scale;
      }
      float* dst = dSt + (tx * 4 + j) * PITCH + ty * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(ds[0], ds[1], ds[2], ds[3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(ds[4], ds[5], ds[6], ds[7]);
    }
    __syncthreads();
    mm64(dq, dSt, Kn, ty, tx);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int qi = q0 + ty * 8 + i;
    if (qi >= p.Sq) continue;
    store4(p.dq, p.dtype, ((long)b * p.Sq + qi) * p.ldq + h * BT + tx * 4,
           make_float4(dq[i][0], dq[i][1], dq[i][2], dq[i][3]));
  }
}

// ------------------------------------------------------------------ backward: dK, dV
__global__ void __launch_bounds__(NTHREADS) attn_bwd_dkv_kernel(AttnArgs p) {
  extern __shared__ float sm[];
  float* Kt = sm; float* Vt = sm + TILE; float* Qt = sm + 2 * TILE; float* Qn = sm + 3 * TILE;
  float* dOt = sm + 4 * TILE; float* dOn = sm + 5 * TILE; float* Pn = sm + 6 * TILE; float* dSn = sm + 7 * TILE;
  const int b = blockIdx.z, h = blockIdx.y, k0 = blockIdx.x * BT;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  load_tile(nullptr, Kt, p.k, p.dtype, p.ldk, (long)b * p.Sk, k0, p.Sk, h * BT);
  load_tile(nullptr, Vt, p.v, p.dtype, p.ldv, (long)b * p.Sk, k0, p.Sk, h * BT);
  bool ok[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int kk = k0 + tx * 4 + j;
    ok[j] = kk < p.Sk && !(p.key_pad && p.key_pad[(long)b * p.Sk + kk]);
  }
  float dk[8][4], dv[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) { dk[i][j] = 0.f; dv[i][j] = 0.f; }
  for (int q0 = 0; q0 < p.Sq; q0 += BT) {
    __syncthreads();
    load_tile(Qn, Qt, p.q, p.dtype, p.ldq, (long)b * p.Sq, q0, p.Sq, h * BT);
    load_tile(dOn, dOt, p.dout, p.dtype, p.ldo, (long)b * p.Sq, q0, p.Sq, h * BT);
    __syncthreads();
    float s[8][4], dp[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) { s[i][j] = 0.f; dp[i][j] = 0.f; }
    mm64(s, Qt, Kt, ty, tx);       // rows = queries (ty), cols = keys (tx)
    mm64(dp, dOt, Vt, ty, tx);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int qi = q0 + ty * 8 + i;
      float lse = 0.f, dl = 0.f;
      const bool qok = qi < p.Sq;
      if (qok) {
        lse = p.lse[((long)b * p.H + h) * p.Sq + qi];
        dl = p.delta[((long)b * p.H + h) * p.Sq + qi];
      }
      float pr[4], ds[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        pr[j] = (qok && ok[j]) ? __expf(s[i][j] * p.scale - lse) : 0.f;
        ds[j] = pr[j] * (dp[i][j] - dl) * p.scale;
      }
      *reinterpret_cast<float4*>(Pn + (ty * 8 + i) * PITCH + tx * 4) = make_float4(pr[0], pr[1], pr[2], pr[3]);
      *reinterpret_cast<float4*>(dSn + (ty * 8 + i) * PITCH + tx * 4) = make_float4(ds[0], ds[1], ds[2], ds[3]);
    }
    __syncthreads();
    mm64(dv, Pn, dOn, ty, tx);     // rows = keys (ty), cols = head dim (tx)
    mm64(dk, dSn, Qn, ty, tx);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int kk = k0 + ty * 8 + i;
    if (kk >= p.Sk) continue;
    store4(p.dk, p.dtype, ((long)b * p.Sk + kk) * p.ldk + h * BT + tx * 4,
           make_float4(dk[i][0], dk[i][1], dk[i][2], dk[i][3]));
    store4(p.dv, p.dtype, ((long)b * p.Sk + kk) * p.ldv + h * BT + tx * 4,
           make_float4(dv[i][0], dv[i][1], dv[i][2], dv[i][3]));
  }
}

template <typename K>
int set_smem(K kfn, size_t bytes, const char* name) {
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) { exg_set_error("%s: smem attribute: %s", name, cudaGetErrorString(e)); return (int)e; }
  return 0;
}

int check_common(int dtype, long ldq, long ldk, long ldv, long ldo, int B, int H, int Sq, int Sk, int d) {
  EXG_CHECK_ARG(exg_valid_dtype(dtype), "attn: bad dtype");
  EXG_CHECK_ARG(d == 64, "attn: head_dim must be 64 (got %d)", d);
  EXG_CHECK_ARG(B > 0 && H > 0 && Sq > 0 && Sk > 0, "attn: bad sizes");
  EXG_CHECK_ARG(ldq % 4 == 0 && ldk % 4 == 0 && ldv % 4 == 0 && ldo % 4 == 0, "attn: strides must be multiples of 4");
  EXG_CHECK_ARG(H <= 65535 && B <= 65535, "attn: B/H too large for the grid");
  return 0;
}

}  // namespace

int exg_attn_tc_ok(int dtype, long ldq, long ldk, long ldv, long ldo, int Sq, int Sk, int d, bool bwd,
                   const void* q, const void* k, const void* v, const void* o);
int exg_attn_tc_fwd(const void* q, const void* k, const void* v, long ldq, long ldk, long ldv,
                    const uint8_t* key_pad, void* o, long ldo, float* lse, int B, int H, int Sq, int Sk,
                    float scale, void* stream);
int exg_attn_tc_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, long ldq,
                    long ldk, long ldv, long ldo, const uint8_t* key_pad, const float* lse, void* dq, void* dk,
                    void* dv, int B, int H, int Sq, int Sk, float scale, void* stream);

int exg_attn_flash_ok(int dtype, long ldq, long ldk, long ldv, long ldo, int Sq, int Sk, int d, const void* q,
                      const void* k, const void* v, const void* o);
int exg_attn_flash_fwd(const void* q, const void* k, const void* v, long ldq, long ldk, long ldv,
                       const uint8_t* key_pad, void* o, long ldo, float* lse, int B, int H, int Sq, int Sk,
                       float scale, void* stream);
int exg_attn_flash_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, long ldq,
                       long ldk, long ldv, long ldo, const uint8_t* key_pad, const float* lse, float* workspace,
                       void* dq, void* dk, void* dv, int B, int H, int Sq, int Sk, float scale, void* stream);

// fp32 elements of the `delta` workspace of exg_attn_bwd: [B,H,Sq] row statistics, plus (sequences longer than one
// 128-row tile) the fp32 dQ accumulator [B*Sq, H*d] of the multi-tile tensor-core backward
extern "C" long exg_attn_bwd_workspace(int B, int H, int Sq, int Sk, int d) {
  long n = (((long)B * H * Sq + 31) / 32) * 32;
  if (Sq > 128 || Sk > 128) n += (long)B * Sq * H * d;
  return n;
}

extern "C" int exg_attn_fwd(const void* q, const void* k, const void* v, int dtype, long ldq, long ldk,
                            long ldv, const uint8_t* key_pad, void* o, long ldo, float* lse, int B, int H,
                            int Sq, int Sk, int d, float scale, void* stream) {
  EXG_CHECK_ARG(q && k && v && o, "attn_fwd: null pointer");
  int rc = check_common(dtype, ldq, ldk, ldv, ldo, B, H, Sq, Sk, d);
  if (rc) return rc;
  if (lse && exg_attn_tc_ok(dtype, ldq, ldk, ldv, ldo, Sq, Sk, d, false, q, k, v, o)) {
    ExgProfScope prof(EXG_PC_ATTN_FWD, 4.0 * B * H * Sq * Sk * d, stream);
    return exg_attn_tc_fwd(q, k, v, ldq, ldk, ldv, key_pad, o, ldo, lse, B, H, Sq, Sk, scale, stream);
  }
  if (exg_attn_flash_ok(dtype, ldq, ldk, ldv, ldo, Sq, Sk, d, q, k, v, o)) {
    ExgProfScope prof(EXG_PC_ATTN_FWD, 4.0 * B * H * Sq * Sk * d, stream);
    return exg_attn_flash_fwd(q, k, v, ldq, ldk, ldv, key_pad, o, ldo, lse, B, H, Sq, Sk, scale, stream);
  }
  AttnArgs a = {};
  a.q = q; a.k = k; a.v = v; a.out = o; a.dtype = dtype; a.ldq = ldq; a.ldk = ldk; a.ldv = ldv; a.ldo = ldo;
  a.key_pad = key_pad; a.lse = lse; a.B = B; a.H = H; a.Sq = Sq; a.Sk = Sk; a.scale = scale;
  const size_t smem = 4 * TILE * sizeof(float);
  static bool once = false;
  if (!once) { rc = set_smem(attn_fwd_kernel, smem, "attn_fwd"); if (rc) return rc; once = true; }
  dim3 grid((Sq + BT - 1) / BT, H, B);
  ExgProfScope prof(EXG_PC_ATTN_FWD, 4.0 * B * H * Sq * Sk * d, stream);
  attn_fwd_kernel<<<grid, NTHREADS, smem, (cudaStream_t)stream>>>(a);
  EXG_LAUNCH_CHECK("attn_fwd");
  return 0;
}

extern "C" int exg_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                            int dtype, long ldq, long ldk, long ldv, long ldo, const uint8_t* key_pad,
                            const float* lse, float* delta, void* dq, void* dk, void* dv, int B, int H,
                            int Sq, int Sk, int d, float scale, void* stream) {
  EXG_CHECK_ARG(q && k && v && o && dout && lse && delta && dq && dk && dv, "attn_bwd: null pointer");
  int rc = check_common(dtype, ldq, ldk, ldv, ldo, B, H, Sq, Sk, d);
  if (rc) return rc;
  if (exg_attn_tc_ok(dtype, ldq, ldk, ldv, ldo, Sq, Sk, d, true, q, k, v, o) &&
      exg_attn_tc_ok(dtype, ldq, ldk, ldv, ldo, Sq, Sk, d, true, dq, dk, dv, dout)) {
    ExgProfScope prof(EXG_PC_ATTN_BWD, 10.0 * B * H * Sq * Sk * d, stream);
    return exg_attn_tc_bwd(q, k, v, o, dout, ldq, ldk, ldv, ldo, key_pad, lse, dq, dk, dv, B, H, Sq, Sk, scale,
                           stream);
  }
  if ((Sq > 128 || Sk > 128) && exg_attn_flash_ok(dtype, ldq, ldk, ldv, ldo, Sq, Sk, d, q, k, v, o) &&
      exg_attn_flash_ok(dtype, ldq, ldk, ldv, ldo, Sq, Sk, d, dq, dk, dv, dout)) {
    ExgProfScope prof(EXG_PC_ATTN_BWD, 10.0 * B * H * Sq * Sk * d, stream);
    return exg_attn_flash_bwd(q, k, v, o, dout, ldq, ldk, ldv, ldo, key_pad, lse, delta, dq, dk, dv, B, H, Sq, Sk,
                              scale, stream);
  }
  AttnArgs a = {};
  a.q = q; a.k = k; a.v = v; a.o = o; a.dout = dout; a.dq = dq; a.dk = dk; a.dv = dv; a.dtype = dtype;
  a.ldq = ldq; a.ldk = ldk; a.ldv = ldv; a.ldo = ldo; a.key_pad = key_pad;
  a.lse = const_cast<float*>(lse); a.delta = delta; a.B = B; a.H = H; a.Sq = Sq; a.Sk = Sk; a.scale = scale;
  const size_t smem_dq = 6 * TILE * sizeof(float), smem_dkv = 8 * TILE * sizeof(float);
  static bool once = false;
  if (!once) {
    rc = set_smem(attn_bwd_dq_kernel, smem_dq, "attn_bwd_dq"); if (rc) return rc;
    rc = set_smem(attn_bwd_dkv_kernel, smem_dkv, "attn_bwd_dkv"); if (rc) return rc;
    once = true;
  }
  cudaStream_t s = (cudaStream_t)stream;
  ExgProfScope prof(EXG_PC_ATTN_BWD, 10.0 * B * H * Sq * Sk * d, stream);
  attn_bwd_dq_kernel<<<dim3((Sq + BT - 1) / BT, H, B), NTHREADS, smem_dq, s>>>(a);
  EXG_LAUNCH_CHECK("attn_bwd_dq");
  attn_bwd_dkv_kernel<<<dim3((Sk + BT - 1) / BT, H, B), NTHREADS, smem_dkv, s>>>(a);
  EXG_LAUNCH_CHECK("attn_bwd_dkv");
  return 0;
}
