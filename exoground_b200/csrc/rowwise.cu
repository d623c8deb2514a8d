// rowwise.cu — HBM-bound row-wise kernels: LayerNorm fwd/bwd, broadcast add, L2 normalise fwd/bwd,
// column sums, casts.  One warp owns one row; a lane owns float4 chunks (j*32+lane)*4 so every
// warp-level load/store is a contiguous 512-byte (fp32) or 256-byte (bf16) segment.
// Reference ops replaced: nn.LayerNorm (model/tfm_model.py:22,28,63,70,71; model/tan_model.py:44-48),
// pos-embed add (tan_model.py:161,193), x / x.norm() (tan_model.py:110-111,130-131).
#include "common.cuh"

namespace {

constexpr int kLnThreads = 128;  // 4 warps = 4 rows per CTA

template <int NCH>
__global__ void __launch_bounds__(kLnThreads)
ln_fwd_kernel(const void* __restrict__ x, int xd, const float* __restrict__ gamma,
              const float* __restrict__ beta, void* __restrict__ y, int yd, exg_rowmap map_y,
              void* __restrict__ y2, int y2d, const float* __restrict__ add, long add_period,
              float* __restrict__ mean_out, float* __restrict__ rstd_out, long rows, int C, float eps,
              const float* __restrict__ resid, float* __restrict__ sum_out) {
  const int lane = threadIdx.x & 31;
  const long row = (long)blockIdx.x * (kLnThreads / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;
  float4 v[NCH];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    int col = (j * 32 + lane) * 4;
    if (col < C) {
      v[j] = load4(x, xd, row * C + col);
      if (resid) {      // residual stream update fused in: s = resid + x, written back as the new stream
        const float4 r = *reinterpret_cast<const float4*>(resid + row * C + col);
        v[j].x += r.x; v[j].y += r.y; v[j].z += r.z; v[j].w += r.w;
        if (sum_out) *reinterpret_cast<float4*>(sum_out + row * C + col) = v[j];
      }
      s += v[j].x + v[j].y + v[j].z + v[j].w;
    } else {
      v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  const float mean = warp_sum(s) / (float)C;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    int col = (j * 32 + lane) * 4;
    if (col < C) {
      float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
      q += a * a + b * b + c * c + d * d;
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)C + eps);
  if (lane == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
  const long orow = exg_map_row(map_y, row);
  const long arow = add ? (long)((unsigned)row % (unsigned)add_period) : 0;
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    int col = (j * 32 + lane) * 4;
    if (col < C) {
      float4 g = *reinterpret_cast<const float4*>(gamma + col);
      float4 b = *reinterpret_cast<const float4*>(beta + col);
      float4 o;
      o.x = (v[j].x - mean) * rstd * g.x + b.x;
      o.y = (v[j].y - mean) * rstd * g.y + b.y;
      o.z = (v[j].z - mean) * rstd * g.z + b.z;
      o.w = (v[j].w - mean) * rstd * g.w + b.w;
      if (add) {
        float4 a = *reinterpret_cast<const float4*>(add + arow * C + col);
        o.x += a.x; o.y += a.y; o.z += a.z; o.w += a.w;
      }
      if (y) store4(y, yd, orow * C + col, o);
      if (y2) store4(y2, y2d, row * C + col, o);
    }
  }
}

constexpr int kLnBwdThreads = 256;  // 8 warps

__device__ __forceinline__ float4 bf16x4_to_float4(uint2 u) {
  float4 r;
  r.x = __uint_as_float(u.x << 16);
  r.y = __uint_as_float(u.x & 0xffff0000u);
  r.z = __uint_as_float(u.y << 16);
  r.w = __uint_as_float(u.y & 0xffff0000u);
  return r;
}

// FAST = the training configuration (x fp32, dy1 / dy2 bf16): the prefetched dy chunks stay RAW bf16 bits in
// registers and are converted / added when the row is consumed.  Converting inside fetch() (the generic path) makes
// the warp wait for its own prefetch at issue time, which serialises load latency and reductions (ncu: 80 % of the
// stall samples were long-scoreboard waits on the bf16 -> fp32 shifts, 19 % of DRAM peak).
template <int NCH, bool FAST>
__global__ void __launch_bounds__(kLnBwdThreads, 2)
ln_bwd_kernel(const void* __restrict__ dy1, int d1, exg_rowmap m1, const void* __restrict__ dy2, int d2,
              exg_rowmap m2, const void* __restrict__ x, int xd, const float* __restrict__ gamma,
              const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
              const float* __restrict__ dres, float* __restrict__ dx, void* __restrict__ dxT, int dxTd,
              float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dxsum, long rows, int C,
              float eps) {
  // One warp per row, software-pipelined: the loads of the warp's next row (x, dy, residual gradient) are issued
  // before the reductions of the current one, so every warp keeps two rows of HBM traffic in flight.  The
  // dgamma / dbeta column partials live in this warp's slice of shared memory (conflict-free float4 RMW), which
  // is what leaves the registers for the prefetch.
  // dxsum (optional): column sums of the OUTPUT rows = the bias gradient of the Linear whose output this
  // residual stream receives (out_proj / c_proj bias: tfm_model.py:36-37), saving a separate pass over dx.
  extern __shared__ float red[];  // [warps][NS*C], NS = 2 (+1 with dxsum)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = blockDim.x / 32;
  const bool want_cols = dgamma || dbeta;
  const int NS = dxsum ? 3 : 2;
  float* rg = red + (size_t)warp * NS * C;
  if (want_cols || dxsum) {
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
      const int col = (j * 32 + lane) * 4;
      if (col < C) {
        *reinterpret_cast<float4*>(rg + col) = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(rg + C + col) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (dxsum) *reinterpret_cast<float4*>(rg + 2 * C + col) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  const long stride = (long)gridDim.x * nwarps;
  long row = (long)blockIdx.x * nwarps + warp;
  float4 nx[NCH], nd[NCH], nr[NCH];
  uint2 q1[FAST ? NCH : 1], q2[FAST ? NCH : 1];      // FAST: raw bf16 bits of dy1 / dy2
  auto fetch = [&](long r) {
    const long r1 = exg_map_row(m1, r);
    const long r2 = dy2 ? exg_map_row(m2, r) : 0;
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
      const int col = (j * 32 + lane) * 4;
      if (col < C) {
        if constexpr (FAST) {
          nx[j] = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(x) + r * C + col);
          q1[j] = *reinterpret_cast<const uint2*>(reinterpret_cast<const uint16_t*>(dy1) + r1 * C + col);
          q2[j] = dy2 ? *reinterpret_cast<const uint2*>(reinterpret_cast<const uint16_t*>(dy2) + r2 * C + col)
                      : make_uint2(0u, 0u);
        } else {
          nx[j] = load4(x, xd, r * C + col);
          nd[j] = load4(dy1, d1, r1 * C + col);
          if (dy2) {
            const float4 t = load4(dy2, d2, r2 * C + col);
            nd[j].x += t.x; nd[j].y += t.y; nd[j].z += t.z; nd[j].w += t.w;
          }
        }
        nr[j] = dres ? *reinterpret_cast<const float4*>(dres + r * C + col) : make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        nx[j] = nd[j] = nr[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if constexpr (FAST) q1[j] = q2[j] = make_uint2(0u, 0u);
      }
    }
  };
  if (row < rows) fetch(row);
  for (; row < rows; row += stride) {
    float4 xv[NCH], dv[NCH], rv[NCH];
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
      xv[j] = nx[j]; rv[j] = nr[j];
      if constexpr (FAST) {
        const float4 a = bf16x4_to_float4(q1[j]), b = bf16x4_to_float4(q2[j]);
        dv[j] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
      } else {
        dv[j] = nd[j];
      }
    }
    if (row + stride < rows) fetch(row + stride);
    float mean, rstd;
    if (mean_in && rstd_in) {
      mean = mean_in[row];
      rstd = rstd_in[row];
    } else {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < NCH; ++j) s += xv[j].x + xv[j].y + xv[j].z + xv[j].w;
      mean = warp_sum(s) / (float)C;
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < NCH; ++j) {
        int col = (j * 32 + lane) * 4;
        if (col < C) {
          float a = xv[j].x - mean, b = xv[j].y - mean, c = xv[j].z - mean, d = xv[j].w - mean;
          q += a * a + b * b + c * c + d * d;
        }
      }
      rstd = rsqrtf(warp_sum(q) / (float)C + eps);
    }
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
      // xv <- xhat ; accumulate dgamma/dbeta ; dv <- dy*gamma
      xv[j].x = (xv[j].x - mean) * rstd; xv[j].y = (xv[j].y - mean) * rstd;
      xv[j].z = (xv[j].z - mean) * rstd; xv[j].w = (xv[j].w - mean) * rstd;
      const int col = (j * 32 + lane) * 4;
      if (col >= C) { xv[j] = make_float4(0.f, 0.f, 0.f, 0.f); continue; }
      if (want_cols) {
        float4 a = *reinterpret_cast<float4*>(rg + col), b = *reinterpret_cast<float4*>(rg + C + col);
        a.x += dv[j].x * xv[j].x; a.y += dv[j].y * xv[j].y; a.z += dv[j].z * xv[j].z; a.w += dv[j].w * xv[j].w;
        b.x += dv[j].x; b.y += dv[j].y; b.z += dv[j].z; b.w += dv[j].w;
        *reinterpret_cast<float4*>(rg + col) = a;
        *reinterpret_cast<float4*>(rg + C + col) = b;
      }
      const float4 g = *reinterpret_cast<const float4*>(gamma + col);
      dv[j].x *= g.x; dv[j].y *= g.y; dv[j].z *= g.z; dv[j].w *= g.w;
      c1 += dv[j].x + dv[j].y + dv[j].z + dv[j].w;
      c2 += dv[j].x * xv[j].x + dv[j].y * xv[j].y + dv[j].z * xv[j].z + dv[j].w * xv[j].w;
    }
    c1 = warp_sum(c1) / (float)C;
    c2 = warp_sum(c2) / (float)C;
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
      int col = (j * 32 + lane) * 4;
      if (col < C) {
        float4 o;
        o.x = (dv[j].x - c1 - xv[j].x * c2) * rstd + rv[j].x;
        o.y = (dv[j].y - c1 - xv[j].y * c2) * rstd + rv[j].y;
        o.z = (dv[j].z - c1 - xv[j].z * c2) * rstd + rv[j].z;
        o.w = (dv[j].w - c1 - xv[j].w * c2) * rstd + rv[j].w;
        if (dx) *reinterpret_cast<float4*>(dx + row * C + col) = o;
        if (dxT) store4(dxT, dxTd, row * C + col, o);
        if (dxsum) {
          float4 a = *reinterpret_cast<float4*>(rg + 2 * C + col);
          a.x += o.x; a.y += o.y; a.z += o.z; a.w += o.w;
          *reinterpret_cast<float4*>(rg + 2 * C + col) = a;
        }
      }
    }
  }
  if (!want_cols && !dxsum) return;
  // cross-warp reduction of the per-warp column partials, then one atomic per column per CTA
  __syncthreads();
  for (int i = threadIdx.x; i < NS * C; i += blockDim.x) {
    float t = 0.f;
    for (int w = 0; w < nwarps; ++w) t += red[(size_t)w * NS * C + i];
    if (i < C) { if (dgamma) atomicAdd(dgamma + i, t); }
    else if (i < 2 * C) { if (dbeta) atomicAdd(dbeta + (i - C), t); }
    else atomicAdd(dxsum + (i - 2 * C), t);
  }
}

__global__ void add_rows_kernel(const float* __restrict__ x, exg_rowmap mx, const float* __restrict__ x2,
                                exg_rowmap mx2, const float* __restrict__ add, long add_period,
                                void* __restrict__ out, int od, exg_rowmap mo, void* __restrict__ out2,
                                int o2d, long rows, int C) {
  const int c4 = C / 4;
  const long total = rows * c4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long r = i / c4;
    const int col = (int)(i % c4) * 4;
    float4 v = *reinterpret_cast<const float4*>(x + exg_map_row(mx, r) * C + col);
    if (x2) {
      float4 t = *reinterpret_cast<const float4*>(x2 + exg_map_row(mx2, r) * C + col);
      v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
    }
    if (add) {
      float4 t = *reinterpret_cast<const float4*>(add + (r % add_period) * C + col);
      v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
    }
    if (out) store4(out, od, exg_map_row(mo, r) * C + col, v);
    if (out2) store4(out2, o2d, r * C + col, v);
  }
}

template <int NCH>
__global__ void __launch_bounds__(kLnThreads)
l2n_fwd_kernel(const void* __restrict__ x, int xd, exg_rowmap mx, void* __restrict__ y, int yd,
               float* __restrict__ inv_norm, long rows, int C) {
  const int lane = threadIdx.x & 31;
  const long row = (long)blockIdx.x * (kLnThreads / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const long ir = exg_map_row(mx, row);
  float4 v[NCH];
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    int col = (j * 32 + lane) * 4;
    if (col < C) {
      v[j] = load4(x, xd, ir * C + col);
      ss += v[j].x * v[j].x + v[j].y * v[j].y + v[j].z * v[j].z + v[j].w * v[j].w;
    }
  }
  const float inv = 1.0f / sqrtf(warp_sum(ss));   // no eps (tan_model.py:110): a zero row gives NaN
  if (lane == 0 && inv_norm) inv_norm[row] = inv;
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    int col = (j * 32 + lane) * 4;
    if (col < C) {
      float4 o = make_float4(v[j].x * inv, v[j].y * inv, v[j].z * inv, v[j].w * inv);
      store4(y, yd, row * C + col, o);
    }
  }
}

template <int NCH>
__global__ void __launch_bounds__(kLnThreads)
l2n_bwd_kernel(const void* __restrict__ dy, int dyd, const void* __restrict__ y, int yd,
               const float* __restrict__ inv_norm, void* __restrict__ dx, int dxd, exg_rowmap mdx,
               int accumulate, long rows, int C) {
  const int lane = threadIdx.x & 31;
  const long row = (long)blockIdx.x * (kLnThreads / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;
  float4 g[NCH], yy[NCH];
  float dot = 0.f;
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    int col = (j * 32 + lane) * 4;
    if (col < C) {
      g[j] = load4(dy, dyd, row * C + col);
      yy[j] = load4(y, yd, row * C + col);
      dot += g[j].x * yy[j].x + g[j].y * yy[j].y + g[j].z * yy[j].z + g[j].w * yy[j].w;
    }
  }
  dot = warp_sum(dot);
  const float inv = inv_norm[row];
  const long orow = exg_map_row(mdx, row);
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    int col = (j * 32 + lane) * 4;
    if (col < C) {
      float4 o;
      o.x = (g[j].x - yy[j].x * dot) * inv; o.y = (g[j].y - yy[j].y * dot) * inv;
      o.z = (g[j].z - yy[j].z * dot) * inv; o.w = (g[j].w - yy[j].w * dot) * inv;
      if (accumulate) {
        float4 p = load4(dx, dxd, orow * C + col);
        o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
      }
      store4(dx, dxd, orow * C + col, o);
    }
  }
}

// out[n] += sum over a slab of rows (optionally weighted per row).  blockDim = (32, 8): x = 4-column group,
// y = row lane; 4 independent row loads in flight per thread.
__global__ void colsum_kernel(const void* __restrict__ x, int xd, const float* __restrict__ roww, long M, long N,
                              long ldx, float* __restrict__ out, long rows_per_block) {
  __shared__ float4 sm[8][32];
  const long col = ((long)blockIdx.x * 32 + threadIdx.x) * 4;
  const long r0 = (long)blockIdx.y * rows_per_block;
  const long r1 = min(M, r0 + rows_per_block);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (col < N) {
    long r = r0 + threadIdx.y;
    for (; r + 24 < r1; r += 32) {
      float4 v[4];
      float w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        v[u] = load4(x, xd, (r + 8 * u) * ldx + col);
        w[u] = roww ? roww[r + 8 * u] : 1.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        acc.x += v[u].x * w[u]; acc.y += v[u].y * w[u]; acc.z += v[u].z * w[u]; acc.w += v[u].w * w[u];
      }
    }
    for (; r < r1; r += 8) {
      const float4 v = load4(x, xd, r * ldx + col);
      const float w = roww ? roww[r] : 1.f;
      acc.x += v.x * w; acc.y += v.y * w; acc.z += v.z * w; acc.w += v.w * w;
    }
  }
  sm[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && col < N) {
#pragma unroll
    for (int k = 1; k < 8; ++k) {
      float4 v = sm[k][threadIdx.x];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    atomicAdd(out + col, acc.x); atomicAdd(out + col + 1, acc.y);
    atomicAdd(out + col + 2, acc.z); atomicAdd(out + col + 3, acc.w);
  }
}

// bf16 fast path of the above: 16-byte loads (8 columns per thread, a warp covers 512 contiguous bytes of a row) and 8
// rows in flight per thread (32 KB per block) - the 8-byte / 4-deep version above sustains ~1.4 TB/s on the
// [16 K x 1.5-2 K] gradient tensors of the ExoGround step, a third of what the same bytes stream at.
__global__ void __launch_bounds__(256) colsum_bf16x8_kernel(const uint16_t* __restrict__ x, const float* __restrict__ roww,
                                                            long M, long N, long ldx, float* __restrict__ out,
                                                            long rows_per_block) {
  __shared__ float sm[8][32][9];               // (+1: the cross-row reduction reads a column of this array)
  const long col = ((long)blockIdx.x * 32 + threadIdx.x) * 8;
  const long r0 = (long)blockIdx.y * rows_per_block;
  const long r1 = min(M, r0 + rows_per_block);
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  if (col < N) {
    const uint16_t* base = x + col;
    long r = r0 + threadIdx.y;
    for (; r + 56 < r1; r += 64) {
      uint4 v[8];
      float w[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        v[u] = *reinterpret_cast<const uint4*>(base + (r + 8 * u) * ldx);
        w[u] = roww ? roww[r + 8 * u] : 1.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t q[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          acc[2 * k] = fmaf(__uint_as_float(q[k] << 16), w[u], acc[2 * k]);
          acc[2 * k + 1] = fmaf(__uint_as_float(q[k] & 0xffff0000u), w[u], acc[2 * k + 1]);
        }
      }
    }
    for (; r < r1; r += 8) {
      const uint4 v = *reinterpret_cast<const uint4*>(base + r * ldx);
      const float w = roww ? roww[r] : 1.f;
      const uint32_t q[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        acc[2 * k] = fmaf(__uint_as_float(q[k] << 16), w, acc[2 * k]);
        acc[2 * k + 1] = fmaf(__uint_as_float(q[k] & 0xffff0000u), w, acc[2 * k + 1]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) sm[threadIdx.y][threadIdx.x][i] = acc[i];
  __syncthreads();
  // 256 threads, 256 columns of the strip: thread t sums column t over the 8 row lanes
  const int t = threadIdx.y * 32 + threadIdx.x;
  const long c = (long)blockIdx.x * 256 + t;
  if (c < N) {
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) sum += sm[k][t >> 3][t & 7];
    atomicAdd(out + c, sum);
  }
}

// y[r] = <x[r, :], w> (+ bias[0]): one warp per row (Linear with one output: the alignability head, tan_model.py:64-68)
__global__ void rowdot_kernel(const void* __restrict__ x, int xd, const float* __restrict__ w,
                              const float* __restrict__ bias, float* __restrict__ y, long rows, int C) {
  const int lane = threadIdx.x & 31;
  const long row = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  float a = 0.f;
  for (int c = lane * 4; c < C; c += 128) {
    const float4 v = load4(x, xd, row * C + c);
    const float4 ww = *reinterpret_cast<const float4*>(w + c);
    a += v.x * ww.x + v.y * ww.y + v.z * ww.z + v.w * ww.w;
  }
  a = warp_sum(a);
  if (lane == 0) y[row] = a + (bias ? bias[0] : 0.f);
}

// out[r, :] = g[r] * w   (gradient of rowdot w.r.t. x)
__global__ void outer_rows_kernel(const float* __restrict__ g, const float* __restrict__ w, void* __restrict__ out,
                                  int od, long rows, int C) {
  const long n4 = rows * (C / 4);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    const long r = i / (C / 4);
    const int c = (int)(i % (C / 4)) * 4;
    const float s = g[r];
    const float4 ww = *reinterpret_cast<const float4*>(w + c);
    store4(out, od, r * C + c, make_float4(s * ww.x, s * ww.y, s * ww.z, s * ww.w));
  }
}

// generic fallback (any N / ldx): one CTA per column
__global__ void colsum_scalar_kernel(const void* __restrict__ x, int xd, long M, long N, long ldx,
                                     float* __restrict__ out, int accumulate) {
  __shared__ float sm[8];
  const long n = blockIdx.x;
  float a = 0.f;
  for (long r = threadIdx.x; r < M; r += blockDim.x) a += load1(x, xd, r * ldx + n);
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sm[w];
    out[n] = accumulate ? out[n] + t : t;
  }
}

__global__ void cast_kernel(const void* __restrict__ src, int sd, void* __restrict__ dst, int dd, long n) {
  const long n4 = n / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x)
    store4(dst, dd, i * 4, load4(src, sd, i * 4));
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    long i = n4 * 4 + threadIdx.x;
    store1(dst, dd, i, load1(src, sd, i));
  }
}

__global__ void stage_mean_kernel(const void* __restrict__ x, int xd, long stride, int S,
                                  float* __restrict__ out, void* __restrict__ outT, int otd, long n) {
  const long n4 = n / 4;
  const float inv = 1.0f / (float)S;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s = 0; s < S; ++s) {
      float4 v = load4(x, xd, (long)s * stride + i * 4);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    a.x *= inv; a.y *= inv; a.z *= inv; a.w *= inv;
    if (out) *reinterpret_cast<float4*>(out + i * 4) = a;
    if (outT) store4(outT, otd, i * 4, a);
  }
}

__global__ void bcast_rows_kernel(const float* __restrict__ g, float alpha, void* __restrict__ out, int od,
                                  long stride, int S, long n) {
  const long n4 = n / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    float4 v = *reinterpret_cast<const float4*>(g + i * 4);
    v.x *= alpha; v.y *= alpha; v.z *= alpha; v.w *= alpha;
    for (int s = 0; s < S; ++s) store4(out, od, (long)s * stride + i * 4, v);
  }
}

inline int grid_for(long work_items, int threads, int max_blocks_per_sm = 8) {
  long b = (work_items + threads - 1) / threads;
  long cap = (long)exg_num_sms() * max_blocks_per_sm;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

bool rowmap_ok(const exg_rowmap& m) { return m.group > 0; }

}  // namespace

#define DISPATCH_NCH(C, ...)                                     \
  do {                                                           \
    if ((C) <= 512) { constexpr int NCH = 4; __VA_ARGS__; }      \
    else if ((C) <= 1024) { constexpr int NCH = 8; __VA_ARGS__; } \
    else if ((C) <= 2048) { constexpr int NCH = 16; __VA_ARGS__; } \
    else { constexpr int NCH = 32; __VA_ARGS__; }                \
  } while (0)

extern "C" int exg_layernorm_fwd(const void* x, int x_dtype, const float* gamma, const float* beta,
                                 void* y, int y_dtype, exg_rowmap map_y, void* y2, int y2_dtype,
                                 const float* add, long add_period, float* mean, float* rstd,
                                 long rows, int C, float eps, void* stream) {
  EXG_CHECK_ARG(x && gamma && beta && (y || y2), "layernorm_fwd: null pointer");
  EXG_CHECK_ARG(exg_valid_dtype(x_dtype) && exg_valid_dtype(y_dtype) && exg_valid_dtype(y2_dtype),
                "layernorm_fwd: bad dtype");
  EXG_CHECK_ARG(C > 0 && C % 4 == 0 && C <= 4096, "layernorm_fwd: C=%d must be a multiple of 4, <= 4096", C);
  EXG_CHECK_ARG(rowmap_ok(map_y) && (!add || add_period > 0), "layernorm_fwd: bad row map");
  if (rows <= 0) return 0;
  EXG_CHECK_ARG(rows <= EXG_MAX_ROWS, "row count exceeds 2^31 - 1");
  cudaStream_t s = (cudaStream_t)stream;
  int blocks = (int)((rows + 3) / 4);
  ExgProfScope prof(EXG_PC_LN_FWD, (double)rows * C * (exg_dtype_size(x_dtype) + (y ? exg_dtype_size(y_dtype) : 0) + (y2 ? exg_dtype_size(y2_dtype) : 0)), stream);
  DISPATCH_NCH(C, ln_fwd_kernel<NCH><<<blocks, kLnThreads, 0, s>>>(
                      x, x_dtype, gamma, beta, y, y_dtype, map_y, y2, y2_dtype, add, add_period, mean,
                      rstd, rows, C, eps, nullptr, nullptr));
  EXG_LAUNCH_CHECK("layernorm_fwd");
  return 0;
}

extern "C" int exg_add_layernorm_fwd(const void* x, int x_dtype, const float* resid, float* sum_out,
                                     const float* gamma, const float* beta, void* y, int y_dtype,
                                     exg_rowmap map_y, long rows, int C, float eps, void* stream) {
  EXG_CHECK_ARG(x && resid && sum_out && gamma && beta && y, "add_layernorm_fwd: null pointer");
  EXG_CHECK_ARG(exg_valid_dtype(x_dtype) && exg_valid_dtype(y_dtype), "add_layernorm_fwd: bad dtype");
  EXG_CHECK_ARG(C > 0 && C % 4 == 0 && C <= 4096, "add_layernorm_fwd: C=%d must be a multiple of 4, <= 4096", C);
  EXG_CHECK_ARG(rowmap_ok(map_y), "add_layernorm_fwd: bad row map");
  if (rows <= 0) return 0;
  EXG_CHECK_ARG(rows <= EXG_MAX_ROWS, "row count exceeds 2^31 - 1");
  cudaStream_t s = (cudaStream_t)stream;
  int blocks = (int)((rows + 3) / 4);
  ExgProfScope prof(EXG_PC_LN_FWD, (double)rows * C * (exg_dtype_size(x_dtype) + 8 + exg_dtype_size(y_dtype)), stream);
  DISPATCH_NCH(C, ln_fwd_kernel<NCH><<<blocks, kLnThreads, 0, s>>>(
                      x, x_dtype, gamma, beta, y, y_dtype, map_y, nullptr, 0, nullptr, 0, nullptr, nullptr,
                      rows, C, eps, resid, sum_out));
  EXG_LAUNCH_CHECK("add_layernorm_fwd");
  return 0;
}

extern "C" int exg_layernorm_bwd(const void* dy1, int dy1_dtype, exg_rowmap map1, const void* dy2,
                                 int dy2_dtype, exg_rowmap map2, const void* x, int x_dtype,
                                 const float* gamma, const float* mean, const float* rstd,
                                 const float* dres, float* dx, void* dxT, int dxT_dtype, float* dgamma,
                                 float* dbeta, float* dxsum, long rows, int C, void* stream) {
  EXG_CHECK_ARG(dy1 && x && gamma, "layernorm_bwd: null pointer");
  EXG_CHECK_ARG(exg_valid_dtype(dy1_dtype) && exg_valid_dtype(dy2_dtype) && exg_valid_dtype(x_dtype) &&
                    exg_valid_dtype(dxT_dtype), "layernorm_bwd: bad dtype");
  EXG_CHECK_ARG(C > 0 && C % 4 == 0 && C <= 4096, "layernorm_bwd: C=%d unsupported", C);
  EXG_CHECK_ARG(rowmap_ok(map1) && (!dy2 || rowmap_ok(map2)), "layernorm_bwd: bad row map");
  if (rows <= 0) return 0;
  EXG_CHECK_ARG(rows <= EXG_MAX_ROWS, "row count exceeds 2^31 - 1");
  cudaStream_t s = (cudaStream_t)stream;
  const int threads = C > 1024 ? 128 : kLnBwdThreads;   // keeps the [warps][3C] reduction buffer under 227 KB
  EXG_CHECK_ARG(!dxsum || C <= 2048, "layernorm_bwd: dxsum needs C <= 2048");
  const int nwarps = threads / 32;
  long want = (rows + nwarps - 1) / nwarps;
  int blocks = (int)(want < 2L * exg_num_sms() ? want : 2L * exg_num_sms());
  size_t smem = (size_t)nwarps * (dxsum ? 3 : 2) * C * sizeof(float);
  const float eps = 1e-5f;
  const bool fast = x_dtype == EXG_F32 && dy1_dtype == EXG_BF16 && (!dy2 || dy2_dtype == EXG_BF16);
  ExgProfScope prof(EXG_PC_LN_BWD, (double)rows * C * (exg_dtype_size(x_dtype) + exg_dtype_size(dy1_dtype) + (dy2 ? exg_dtype_size(dy2_dtype) : 0) + (dres ? 4 : 0) + (dx ? 4 : 0) + (dxT ? exg_dtype_size(dxT_dtype) : 0)), stream);
#define LN_BWD_CALL                                                                               \
  do {                                                                                            \
    auto kfn = fast ? ln_bwd_kernel<NCH, true> : ln_bwd_kernel<NCH, false>;                       \
    if (smem > 48 * 1024) cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    kfn<<<blocks, threads, smem, s>>>(dy1, dy1_dtype, map1, dy2, dy2_dtype, map2, x, x_dtype, \
                                            gamma, mean, rstd, dres, dx, dxT, dxT_dtype, dgamma,  \
                                            dbeta, dxsum, rows, C, eps);                          \
  } while (0)
  DISPATCH_NCH(C, LN_BWD_CALL);
#undef LN_BWD_CALL
  EXG_LAUNCH_CHECK("layernorm_bwd");
  return 0;
}

extern "C" int exg_add_rows(const float* x, exg_rowmap map_x, const float* x2, exg_rowmap map_x2,
                            const float* add, long add_period, void* out, int out_dtype,
                            exg_rowmap map_out, void* out2, int out2_dtype, long rows, int C,
                            void* stream) {
  EXG_CHECK_ARG(x && (out || out2), "add_rows: null pointer");
  EXG_CHECK_ARG(exg_valid_dtype(out_dtype) && exg_valid_dtype(out2_dtype), "add_rows: bad dtype");
  EXG_CHECK_ARG(C > 0 && C % 4 == 0, "add_rows: C must be a multiple of 4");
  EXG_CHECK_ARG(rowmap_ok(map_x) && rowmap_ok(map_out) && (!x2 || rowmap_ok(map_x2)) &&
                    (!add || add_period > 0), "add_rows: bad row map");
  if (rows <= 0) return 0;
  EXG_CHECK_ARG(rows <= EXG_MAX_ROWS, "row count exceeds 2^31 - 1");
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * C * (4.0 + (x2 ? 4 : 0) + (out ? exg_dtype_size(out_dtype) : 0) + (out2 ? exg_dtype_size(out2_dtype) : 0)), stream);
  add_rows_kernel<<<grid_for(rows * (C / 4), 256), 256, 0, (cudaStream_t)stream>>>(
      x, map_x, x2, map_x2, add, add_period, out, out_dtype, map_out, out2, out2_dtype, rows, C);
  EXG_LAUNCH_CHECK("add_rows");
  return 0;
}

extern "C" int exg_l2norm_fwd(const void* x, int x_dtype, exg_rowmap map_x, void* y, int y_dtype,
                              float* inv_norm, long rows, int C, void* stream) {
  EXG_CHECK_ARG(x && y, "l2norm_fwd: null pointer");
  EXG_CHECK_ARG(exg_valid_dtype(x_dtype) && exg_valid_dtype(y_dtype), "l2norm_fwd: bad dtype");
  EXG_CHECK_ARG(C > 0 && C % 4 == 0 && C <= 4096 && rowmap_ok(map_x), "l2norm_fwd: bad C / map");
  if (rows <= 0) return 0;
  EXG_CHECK_ARG(rows <= EXG_MAX_ROWS, "row count exceeds 2^31 - 1");
  int blocks = (int)((rows + 3) / 4);
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * C * (exg_dtype_size(x_dtype) + exg_dtype_size(y_dtype)), stream);
  DISPATCH_NCH(C, l2n_fwd_kernel<NCH><<<blocks, kLnThreads, 0, (cudaStream_t)stream>>>(
                      x, x_dtype, map_x, y, y_dtype, inv_norm, rows, C));
  EXG_LAUNCH_CHECK("l2norm_fwd");
  return 0;
}

extern "C" int exg_l2norm_bwd(const void* dy, int dy_dtype, const void* y, int y_dtype,
                              const float* inv_norm, void* dx, int dx_dtype, exg_rowmap map_dx,
                              int accumulate, long rows, int C, void* stream) {
  EXG_CHECK_ARG(dy && y && inv_norm && dx, "l2norm_bwd: null pointer");
  EXG_CHECK_ARG(exg_valid_dtype(dy_dtype) && exg_valid_dtype(y_dtype) && exg_valid_dtype(dx_dtype),
                "l2norm_bwd: bad dtype");
  EXG_CHECK_ARG(C > 0 && C % 4 == 0 && C <= 4096 && rowmap_ok(map_dx), "l2norm_bwd: bad C / map");
  if (rows <= 0) return 0;
  EXG_CHECK_ARG(rows <= EXG_MAX_ROWS, "row count exceeds 2^31 - 1");
  int blocks = (int)((rows + 3) / 4);
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * C * (exg_dtype_size(dy_dtype) + exg_dtype_size(y_dtype) + exg_dtype_size(dx_dtype) * (accumulate ? 2 : 1)), stream);
  DISPATCH_NCH(C, l2n_bwd_kernel<NCH><<<blocks, kLnThreads, 0, (cudaStream_t)stream>>>(
                      dy, dy_dtype, y, y_dtype, inv_norm, dx, dx_dtype, map_dx, accumulate, rows, C));
  EXG_LAUNCH_CHECK("l2norm_bwd");
  return 0;
}

static int colsum_impl(const void* x, int x_dtype, const float* roww, long M, long N, long ldx, float* out,
                       int accumulate, void* stream) {
  EXG_CHECK_ARG(x && out && exg_valid_dtype(x_dtype), "colsum: bad argument");
  EXG_CHECK_ARG(N > 0, "colsum: N must be positive");
  cudaStream_t s = (cudaStream_t)stream;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)M * N * exg_dtype_size(x_dtype), stream);
  if (N % 4 != 0 || ldx % 4 != 0) {
    EXG_CHECK_ARG(!roww, "wcolsum: N and ldx must be multiples of 4");
    EXG_CHECK_ARG(N <= 65535, "colsum: unaligned N too large");
    colsum_scalar_kernel<<<(unsigned)N, 256, 0, s>>>(x, x_dtype, M, N, ldx, out, accumulate);
    EXG_LAUNCH_CHECK("colsum_scalar");
    return 0;
  }
  if (!accumulate) EXG_CUDA(cudaMemsetAsync(out, 0, (size_t)N * sizeof(float), s));
  if (M <= 0) return 0;
  if (x_dtype == EXG_BF16 && N % 8 == 0 && ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    const long strips8 = (N + 255) / 256;
    const long target8 = 4L * exg_num_sms();
    long splits8 = (target8 + strips8 - 1) / strips8;
    long rpb = (M + splits8 - 1) / splits8;
    if (rpb < 64) rpb = 64;
    splits8 = (M + rpb - 1) / rpb;
    dim3 grid8((unsigned)strips8, (unsigned)splits8), block8(32, 8);
    colsum_bf16x8_kernel<<<grid8, block8, 0, s>>>(reinterpret_cast<const uint16_t*>(x), roww, M, N, ldx, out, rpb);
    EXG_LAUNCH_CHECK("colsum_bf16x8");
    return 0;
  }
  long strips = (N + 127) / 128;
  long target = 4L * exg_num_sms();
  long splits = (target + strips - 1) / strips;
  long rows_per_block = (M + splits - 1) / splits;
  if (rows_per_block < 64) rows_per_block = 64;
  splits = (M + rows_per_block - 1) / rows_per_block;
  dim3 grid((unsigned)strips, (unsigned)splits), block(32, 8);
  colsum_kernel<<<grid, block, 0, s>>>(x, x_dtype, roww, M, N, ldx, out, rows_per_block);
  EXG_LAUNCH_CHECK("colsum");
  return 0;
}

extern "C" int exg_colsum(const void* x, int x_dtype, long M, long N, long ldx, float* out,
                          int accumulate, void* stream) {
  return colsum_impl(x, x_dtype, nullptr, M, N, ldx, out, accumulate, stream);
}

extern "C" int exg_wcolsum(const void* x, int x_dtype, const float* row_w, long M, long N, long ldx, float* out,
                           int accumulate, void* stream) {
  EXG_CHECK_ARG(row_w, "wcolsum: null row weights");
  return colsum_impl(x, x_dtype, row_w, M, N, ldx, out, accumulate, stream);
}

extern "C" int exg_rowdot(const void* x, int x_dtype, const float* w, const float* bias, float* y, long rows, int C,
                          void* stream) {
  EXG_CHECK_ARG(x && w && y && exg_valid_dtype(x_dtype) && C > 0 && C % 4 == 0, "rowdot: bad argument");
  if (rows <= 0) return 0;
  EXG_CHECK_ARG(rows <= EXG_MAX_ROWS, "row count exceeds 2^31 - 1");
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * C * exg_dtype_size(x_dtype), stream);
  rowdot_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, x_dtype, w, bias, y, rows, C);
  EXG_LAUNCH_CHECK("rowdot");
  return 0;
}

extern "C" int exg_outer_rows(const float* g, const float* w, void* out, int out_dtype, long rows, int C,
                              void* stream) {
  EXG_CHECK_ARG(g && w && out && exg_valid_dtype(out_dtype) && C > 0 && C % 4 == 0, "outer_rows: bad argument");
  if (rows <= 0) return 0;
  EXG_CHECK_ARG(rows <= EXG_MAX_ROWS, "row count exceeds 2^31 - 1");
  ExgProfScope prof(EXG_PC_ROWWISE, (double)rows * C * exg_dtype_size(out_dtype), stream);
  outer_rows_kernel<<<grid_for(rows * (C / 4), 256), 256, 0, (cudaStream_t)stream>>>(g, w, out, out_dtype, rows, C);
  EXG_LAUNCH_CHECK("outer_rows");
  return 0;
}

extern "C" int exg_cast(const void* src, int src_dtype, void* dst, int dst_dtype, long n, void* stream) {
  EXG_CHECK_ARG(src && dst && exg_valid_dtype(src_dtype) && exg_valid_dtype(dst_dtype), "cast: bad argument");
  if (n <= 0) return 0;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)n * (exg_dtype_size(src_dtype) + exg_dtype_size(dst_dtype)), stream);
  cast_kernel<<<grid_for(n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(src, src_dtype, dst, dst_dtype, n);
  EXG_LAUNCH_CHECK("cast");
  return 0;
}

extern "C" int exg_stage_mean(const void* x, int x_dtype, long stage_stride, int S, float* out,
                              void* outT, int outT_dtype, long n, void* stream) {
  EXG_CHECK_ARG(x && (out || outT) && exg_valid_dtype(x_dtype) && exg_valid_dtype(outT_dtype) && S > 0,
                "stage_mean: bad argument");
  EXG_CHECK_ARG(n % 4 == 0 && stage_stride % 4 == 0, "stage_mean: n must be a multiple of 4");
  if (n <= 0) return 0;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)n * (S * exg_dtype_size(x_dtype) + (out ? 4 : 0) + (outT ? exg_dtype_size(outT_dtype) : 0)), stream);
  stage_mean_kernel<<<grid_for(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(x, x_dtype, stage_stride, S,
                                                                          out, outT, outT_dtype, n);
  EXG_LAUNCH_CHECK("stage_mean");
  return 0;
}

extern "C" int exg_bcast_rows(const float* g, float alpha, void* out, int out_dtype, long stage_stride, int S,
                              long n, void* stream) {
  EXG_CHECK_ARG(g && out && exg_valid_dtype(out_dtype) && S > 0, "bcast_rows: bad argument");
  EXG_CHECK_ARG(n % 4 == 0 && stage_stride % 4 == 0, "bcast_rows: n must be a multiple of 4");
  if (n <= 0) return 0;
  ExgProfScope prof(EXG_PC_ROWWISE, (double)n * (4.0 + S * exg_dtype_size(out_dtype)), stream);
  bcast_rows_kernel<<<grid_for(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(g, alpha, out, out_dtype, stage_stride,
                                                                          S, n);
  EXG_LAUNCH_CHECK("bcast_rows");
  return 0;
}
