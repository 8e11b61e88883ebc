// K2-K5 stand-alone interaction layers (CUDA-core fp32): Wide, FM (dense / one-hot),
// the reference's FFM, PNN inner / outer product, DCN cross.  One warp per sample unless noted.
// All are HBM/L2-bound integer+fp32 work; no tensor cores (interaction outputs must stay
// within 1e-5 relative of the fp32 reference, which TF32 cannot give).
#include <algorithm>
#include "common.cuh"

namespace ctrhot {

#define WARP_IDS()                                                                  \
  const int lane = threadIdx.x & 31;                                                \
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); \
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);

// ---------------------------------------------------------------- R2 WideLayer
__global__ void __launch_bounds__(256)
wide_dense_kernel(const float* __restrict__ x, int64_t x_ld, int64_t n, int64_t B,
                  const float* __restrict__ w0, const float* __restrict__ w, float* __restrict__ out) {
  WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    const float* row = x + b * x_ld;
    float acc = 0.f;
    for (int64_t j = lane; j < n; j += 32) {
      const float t = ldg_nc_f1(row + j);
      if (t != 0.f) acc = fmaf(t, __ldg(w + j), acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) out[b] = acc + __ldg(w0);
  }
}

template <int IDKIND>
__global__ void __launch_bounds__(256)
wide_onehot_kernel(const float* __restrict__ dense, int64_t dense_ld, int D,
                   const int32_t* __restrict__ dense_col, const void* __restrict__ ids,
                   int64_t ids_ld, int F, const int64_t* __restrict__ col_off, int64_t B,
                   const float* __restrict__ w0, const float* __restrict__ w, float* __restrict__ out) {
  WARP_IDS();
  // S samples per warp pass: all S * ceil(F/32) id loads are issued before the first dependent weight load, so a warp pays
  // two DRAM round trips per S samples instead of per sample (the kernel is latency-bound: 4 bytes per (sample, field))
  constexpr int S = 4;
  if (F <= 32) {
    const bool actf = lane < F;
    const int64_t c0 = actf ? col_off[lane] : 0, nc = actf ? col_off[lane + 1] - c0 : 0;
    const float wd = (lane < D) ? __ldg(w + __ldg(dense_col + lane)) : 0.f;
    const float bias = __ldg(w0);
    for (int64_t b0 = warp * S; b0 < B; b0 += nwarps * S) {
      int64_t id[S];
      float xd[S], acc[S];
#pragma unroll
      for (int s = 0; s < S; ++s) {
        const int64_t b = b0 + s;
        id[s] = (actf && b < B) ? IdLoad<IDKIND>::at(ids, b * ids_ld + lane) : -1;
        xd[s] = (lane < D && b < B) ? __ldg(dense + b * dense_ld + lane) : 0.f;
      }
#pragma unroll
      for (int s = 0; s < S; ++s) {
        acc[s] = xd[s] * wd;
        if (id[s] >= 0 && id[s] < nc) acc[s] += ldg_nc_f1(w + c0 + id[s]);
      }
      for (int j = 32 + lane; j < D; j += 32) {                  // more than 32 dense columns: rare
#pragma unroll
        for (int s = 0; s < S; ++s)
          if (b0 + s < B) acc[s] = fmaf(__ldg(dense + (b0 + s) * dense_ld + j), __ldg(w + __ldg(dense_col + j)), acc[s]);
      }
#pragma unroll
      for (int s = 0; s < S; ++s) acc[s] = warp_sum(acc[s]);
      if (lane < S && b0 + lane < B) {
        float v = acc[0];
#pragma unroll
        for (int s = 1; s < S; ++s) if (lane == s) v = acc[s];
        out[b0 + lane] = v + bias;
      }
    }
    return;
  }
  for (int64_t b = warp; b < B; b += nwarps) {
    float acc = 0.f;
    for (int j = lane; j < D; j += 32)
      acc = fmaf(__ldg(dense + b * dense_ld + j), __ldg(w + __ldg(dense_col + j)), acc);
    for (int f = lane; f < F; f += 32) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      const int64_t c0 = col_off[f], c1 = col_off[f + 1];
      if (id >= 0 && id < c1 - c0) acc += ldg_nc_f1(w + c0 + id);
    }
    acc = warp_sum(acc);
    if (lane == 0) out[b] = acc + __ldg(w0);
  }
}

// ---------------------------------------------------------------- R3 FMLayer, dense x
// lane c < kfm owns factor c: s_c = sum_j x_j v[j,c], t_c = sum_j x_j^2 v[j,c]^2; lane kfm owns the
// linear term.  Non-zero columns are found with a ballot so a one-hot row costs ~nnz v-row reads.
__global__ void __launch_bounds__(256)
fm_dense_kernel(const float* __restrict__ x, int64_t x_ld, int64_t n, int64_t B,
                const float* __restrict__ w0, const float* __restrict__ w1,
                const float* __restrict__ v, int kfm, float* __restrict__ out, int apply_sigmoid) {
  WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    const float* row = x + b * x_ld;
    float s = 0.f, t = 0.f;
    for (int64_t j0 = 0; j0 < n; j0 += 32) {
      const int64_t j = j0 + lane;
      const float xj = (j < n) ? ldg_nc_f1(row + j) : 0.f;
      unsigned m = __ballot_sync(0xffffffffu, xj != 0.f);
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const float xs = __shfl_sync(0xffffffffu, xj, src);
        const int64_t js = j0 + src;
        if (lane < kfm) {
          const float vv = __ldg(v + js * kfm + lane);
          s = fmaf(xs, vv, s);
          t = fmaf(xs * xs, vv * vv, t);
        } else if (lane == kfm) {
          s = fmaf(xs, __ldg(w1 + js), s);
        }
      }
    }
    float r = (lane < kfm) ? 0.5f * (s * s - t) : (lane == kfm ? s : 0.f);
    r = warp_sum(r);
    if (lane == 0) {
      r += __ldg(w0);
      out[b] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
    }
  }
}

// Same layer for WIDE inputs (config 5: x [B, 3341] dense; config 1: x [B, 43604] one-hot).  A CTA owns 32 samples, a warp
// FOUR of them, so every v row is used 32 times per staging and 4 times per register load.  v / w1 are staged per tile of
// 256 columns into shared memory TRANSPOSED ([factor][column]: lanes read consecutive words, no bank conflicts -- reading
// v[j, c] straight from global costs one L1 wavefront per lane because rows are kfm*4 bytes apart).  Per 32 columns:
//   all four rows zero  -> skip (one-hot inputs are mostly this);
//   <= 4 live columns   -> the sparse walk of fm_dense_kernel (lane c = factor c), for the four samples at once;
//   otherwise           -> dense: lane l owns column j0 + l, accumulates s[i][c] += x_i v[j,c] in registers (reduced
//                          across lanes once at the end) and t_i += x_i^2 * sum_c v[j,c]^2.
// HBM: x is read once (coalesced 128-B rows).
constexpr int FMW_TJ = 256;
template <int KF>
__global__ void __launch_bounds__(256)
fm_dense_wide_kernel(const float* __restrict__ x, int64_t x_ld, int64_t n, int64_t B,
                     const float* __restrict__ w0, const float* __restrict__ w1,
                     const float* __restrict__ v, int kfm, float* __restrict__ out, int apply_sigmoid) {
  __shared__ float vs[(KF + 1) * FMW_TJ];                     // planes 0..kfm-1 = v[:, c], plane kfm = w1
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const int64_t ngroups = (B + 3) >> 2;
  const int64_t nblocks = (ngroups + 7) >> 3;
  for (int64_t gb = blockIdx.x; gb < nblocks; gb += gridDim.x) {
    const int64_t g = gb * 8 + wic;
    const bool live_w = g < ngroups;
    const int64_t b0 = g << 2;
    const float* row[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) row[i] = x + ((live_w && b0 + i < B) ? b0 + i : B - 1) * x_ld;
    float s[4][KF], lin[4], t[4], sps[4], spt[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      lin[i] = t[i] = sps[i] = spt[i] = 0.f;
#pragma unroll
      for (int c = 0; c < KF; ++c) s[i][c] = 0.f;
    }
    for (int64_t jt = 0; jt < n; jt += FMW_TJ) {
      const int tj = (int)((n - jt < FMW_TJ) ? n - jt : FMW_TJ);
      __syncthreads();                                         // the previous tile is no longer being read
      for (int idx = threadIdx.x; idx < tj * kfm; idx += 256) {
        const int jl = idx / kfm, c = idx - jl * kfm;
        vs[c * FMW_TJ + jl] = __ldg(v + jt * kfm + idx);
      }
      for (int jl = threadIdx.x; jl < tj; jl += 256) vs[kfm * FMW_TJ + jl] = __ldg(w1 + jt + jl);
      __syncthreads();
      if (!live_w) continue;
      for (int jq = 0; jq < tj; jq += 128) {
        // 16 row loads in flight per lane (4 groups of 32 columns x 4 samples) before any of them is consumed: the
        // kernel is a stream over x, and 2 CTAs x 256 threads x 64 B is what covers the HBM latency
        float xq[4][4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int jl = jq + 32 * u + lane;
#pragma unroll
          for (int i = 0; i < 4; ++i) xq[u][i] = (jl < tj) ? ldg_nc_f1(row[i] + jt + jl) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int jg = jq + 32 * u;
          const int jl = jg + lane;
          const float (&xv)[4] = xq[u];
          unsigned m = __ballot_sync(0xffffffffu, (xv[0] != 0.f) | (xv[1] != 0.f) | (xv[2] != 0.f) | (xv[3] != 0.f));
          if (m == 0) {
          } else if (__popc(m) <= 4) {
            while (m) {
              const int src = __ffs(m) - 1;
              m &= m - 1;
              const float wv = (lane <= kfm) ? vs[lane * FMW_TJ + jg + src] : 0.f;
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float xs = __shfl_sync(0xffffffffu, xv[i], src);
                sps[i] = fmaf(xs, wv, sps[i]);
                spt[i] = fmaf(xs * xs, wv * wv, spt[i]);      // only lanes < kfm use it
              }
            }
          } else if (jl < tj) {
            const float wl = vs[kfm * FMW_TJ + jl];
            float q = 0.f;
#pragma unroll
            for (int c = 0; c < KF; ++c) {
              if (c < kfm) {
                const float vv = vs[c * FMW_TJ + jl];
                q = fmaf(vv, vv, q);
#pragma unroll
                for (int i = 0; i < 4; ++i) s[i][c] = fmaf(xv[i], vv, s[i][c]);
              }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              lin[i] = fmaf(xv[i], wl, lin[i]);
              t[i] = fmaf(xv[i] * xv[i], q, t[i]);
            }
          }
        }
      }
    }
    if (!live_w) continue;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float inter = 0.f;
#pragma unroll
      for (int c = 0; c < KF; ++c) {
        if (c < kfm) {
          const float tot = warp_sum(s[i][c]) + __shfl_sync(0xffffffffu, sps[i], c);
          inter = fmaf(tot, tot, inter);
        }
      }
      const float tt = warp_sum(t[i] + (lane < kfm ? spt[i] : 0.f));
      const float ll = warp_sum(lin[i] + (lane == kfm ? sps[i] : 0.f));
      if (lane == 0 && b0 + i < B) {
        const float r = __ldg(w0) + ll + 0.5f * (inter - tt);
        out[b0 + i] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
      }
    }
  }
}

// The same kernel for 16-byte aligned rows (x_ld % 4 == 0): lane l owns the FOUR consecutive columns 4l..4l+3 of a block
// of 128, so one warp instruction reads 512 contiguous bytes of a row (four full DRAM sectors runs instead of 128-byte
// pieces of four different DRAM pages -- measured 2.7 ms -> see DESIGN.md for x [262144, 3341]) and one LDS.128 fetches
// v[4l..4l+3, c].  skip / sparse / dense is decided per block of 128 columns.
template <int KF>
__global__ void __launch_bounds__(256)
fm_dense_wide_v4_kernel(const float* __restrict__ x, int64_t x_ld, int64_t n, int64_t B,
                        const float* __restrict__ w0, const float* __restrict__ w1,
                        const float* __restrict__ v, int kfm, float* __restrict__ out, int apply_sigmoid,
                        float* __restrict__ part = nullptr, int64_t cols_per_split = 0) {
  // part != NULL: few samples, many columns (the reference's own FM case: 2 000 rows x 43 604 one-hot columns would fill 63 CTAs):
  // blockIdx.y takes the column range [y, y + 1) * cols_per_split and leaves its KF + 2 partial sums per sample in
  // part[y][b][0 .. KF + 2) = {sum_j x_j v_jc (c < KF), sum_j x_j^2 sum_c v_jc^2, x . w1}; fm_dense_finish_kernel adds them in order
  const int64_t j_begin = part ? (int64_t)blockIdx.y * cols_per_split : 0;
  const int64_t j_end = part ? (j_begin + cols_per_split < n ? j_begin + cols_per_split : n) : n;
  __shared__ __align__(16) float vs[(KF + 1) * FMW_TJ];       // planes 0..kfm-1 = v[:, c], plane kfm = w1; zero past n
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const int64_t ngroups = (B + 3) >> 2;
  const int64_t nblocks = (ngroups + 7) >> 3;
  for (int64_t gb = blockIdx.x; gb < nblocks; gb += gridDim.x) {
    const int64_t g = gb * 8 + wic;
    const bool live_w = g < ngroups;
    const int64_t b0 = g << 2;
    const float* row[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) row[i] = x + ((live_w && b0 + i < B) ? b0 + i : B - 1) * x_ld;
    float s[4][KF], lin[4], t[4], sps[4], spt[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      lin[i] = t[i] = sps[i] = spt[i] = 0.f;
#pragma unroll
      for (int c = 0; c < KF; ++c) s[i][c] = 0.f;
    }
    for (int64_t jt = j_begin; jt < j_end; jt += FMW_TJ) {
      const int tj = (int)((j_end - jt < FMW_TJ) ? j_end - jt : FMW_TJ);
      __syncthreads();                                         // the previous tile is no longer being read
      for (int idx = threadIdx.x; idx < FMW_TJ * kfm; idx += 256) {
        const int jl = idx / kfm, c = idx - jl * kfm;
        vs[c * FMW_TJ + jl] = (jl < tj) ? __ldg(v + jt * kfm + idx) : 0.f;
      }
      vs[kfm * FMW_TJ + threadIdx.x] = ((int)threadIdx.x < tj) ? __ldg(w1 + jt + threadIdx.x) : 0.f;
      __syncthreads();
      if (!live_w) continue;
      float4 xq[2][4];                                          // both 128-column blocks of the tile are fetched up front
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int jl = 128 * h + 4 * lane;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          // x_ld % 4 == 0 and the columns [n, x_ld) of a row are readable, so a float4 straddling n stays inside the row;
          // its columns >= n meet zeros in vs
          xq[h][i] = (jl < tj) ? ldg_nc_f4(row[i] + jt + jl) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int jl = 128 * h + 4 * lane;
        float xa[4][4];                                         // [sample][column of mine]
#pragma unroll
        for (int i = 0; i < 4; ++i) { xa[i][0] = xq[h][i].x; xa[i][1] = xq[h][i].y; xa[i][2] = xq[h][i].z; xa[i][3] = xq[h][i].w; }
        if (jl < tj && jt + jl + 4 > n) {                       // the straddling float4: drop what lies past n
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (jt + jl + e >= n) { xa[0][e] = xa[1][e] = xa[2][e] = xa[3][e] = 0.f; }
        }
        unsigned nz = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e)
          nz |= ((xa[0][e] != 0.f) | (xa[1][e] != 0.f) | (xa[2][e] != 0.f) | (xa[3][e] != 0.f)) ? (1u << e) : 0u;
        unsigned m = __ballot_sync(0xffffffffu, nz != 0u);
        if (m == 0) {
        } else if (__popc(m) <= 4) {
          while (m) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const unsigned snz = __shfl_sync(0xffffffffu, nz, src);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              if (!(snz & (1u << e))) continue;                 // warp-uniform
              const float wv = (lane <= kfm) ? vs[lane * FMW_TJ + 128 * h + 4 * src + e] : 0.f;
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float xs = __shfl_sync(0xffffffffu, xa[i][e], src);
                sps[i] = fmaf(xs, wv, sps[i]);
                spt[i] = fmaf(xs * xs, wv * wv, spt[i]);      // only lanes < kfm use it
              }
            }
          }
        } else if (jl < tj) {
          const float4 wl = *reinterpret_cast<const float4*>(&vs[kfm * FMW_TJ + jl]);
          float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int c = 0; c < KF; ++c) {
            if (c < kfm) {
              const float4 vv = *reinterpret_cast<const float4*>(&vs[c * FMW_TJ + jl]);
              q.x = fmaf(vv.x, vv.x, q.x); q.y = fmaf(vv.y, vv.y, q.y); q.z = fmaf(vv.z, vv.z, q.z); q.w = fmaf(vv.w, vv.w, q.w);
#pragma unroll
              for (int i = 0; i < 4; ++i)
                s[i][c] = fmaf(xa[i][3], vv.w, fmaf(xa[i][2], vv.z, fmaf(xa[i][1], vv.y, fmaf(xa[i][0], vv.x, s[i][c]))));
            }
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            lin[i] = fmaf(xa[i][3], wl.w, fmaf(xa[i][2], wl.z, fmaf(xa[i][1], wl.y, fmaf(xa[i][0], wl.x, lin[i]))));
            t[i] = fmaf(xa[i][3] * xa[i][3], q.w, fmaf(xa[i][2] * xa[i][2], q.z,
                   fmaf(xa[i][1] * xa[i][1], q.y, fmaf(xa[i][0] * xa[i][0], q.x, t[i]))));
          }
        }
      }
    }
    if (!live_w) continue;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float inter = 0.f;
      float* prow = part ? part + (((int64_t)blockIdx.y * B) + (b0 + i < B ? b0 + i : 0)) * (KF + 2) : nullptr;
#pragma unroll
      for (int c = 0; c < KF; ++c) {
        if (c < kfm) {
          const float tot = warp_sum(s[i][c]) + __shfl_sync(0xffffffffu, sps[i], c);
          inter = fmaf(tot, tot, inter);
          if (prow && lane == 0 && b0 + i < B) prow[c] = tot;
        } else if (prow && lane == 0 && b0 + i < B) prow[c] = 0.f;
      }
      const float tt = warp_sum(t[i] + (lane < kfm ? spt[i] : 0.f));
      const float ll = warp_sum(lin[i] + (lane == kfm ? sps[i] : 0.f));
      if (lane == 0 && b0 + i < B) {
        if (prow) { prow[KF] = tt; prow[KF + 1] = ll; continue; }
        const float r = __ldg(w0) + ll + 0.5f * (inter - tt);
        out[b0 + i] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
      }
    }
  }
}

// out[b] = w0 + sum_y ll + 1/2 (sum_c (sum_y s_c)^2 - sum_y tt), the column splits added in order (deterministic)
__global__ void fm_dense_finish_kernel(const float* __restrict__ part, int nsplit, int64_t B, int KF2, const float* __restrict__ w0,
                                       float* __restrict__ out, int apply_sigmoid) {
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < B; b += (int64_t)gridDim.x * blockDim.x) {
    float inter = 0.f, tt = 0.f, ll = 0.f;
    for (int c = 0; c < KF2 - 2; ++c) {
      float tot = 0.f;
      for (int y = 0; y < nsplit; ++y) tot += part[((int64_t)y * B + b) * KF2 + c];
      inter = fmaf(tot, tot, inter);
    }
    for (int y = 0; y < nsplit; ++y) { tt += part[((int64_t)y * B + b) * KF2 + KF2 - 2]; ll += part[((int64_t)y * B + b) * KF2 + KF2 - 1]; }
    const float r = __ldg(w0) + ll + 0.5f * (inter - tt);
    out[b] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
  }
}

// ---------------------------------------------------------------- R3 FMLayer, one-hot as gather
template <int IDKIND>
__global__ void __launch_bounds__(256)
fm_onehot_kernel(const float* __restrict__ dense, int64_t dense_ld, int D,
                 const int32_t* __restrict__ dense_col, const void* __restrict__ ids,
                 int64_t ids_ld, int F, const int64_t* __restrict__ col_off, int64_t B,
                 const float* __restrict__ w0, const float* __restrict__ w1,
                 const float* __restrict__ v, int kfm, float* __restrict__ out, int apply_sigmoid) {
  WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    float s = 0.f, t = 0.f;
    for (int j = 0; j < D; ++j) {
      const float xs = __ldg(dense + b * dense_ld + j);
      if (xs == 0.f) continue;
      const int64_t js = __ldg(dense_col + j);
      if (lane < kfm) {
        const float vv = __ldg(v + js * kfm + lane);
        s = fmaf(xs, vv, s);
        t = fmaf(xs * xs, vv * vv, t);
      } else if (lane == kfm) {
        s = fmaf(xs, __ldg(w1 + js), s);
      }
    }
    for (int f = 0; f < F; ++f) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      const int64_t c0 = col_off[f], c1 = col_off[f + 1];
      if (id < 0 || id >= c1 - c0) continue;
      const int64_t js = c0 + id;
      if (lane < kfm) {
        const float vv = ldg_nc_f1(v + js * kfm + lane);
        s += vv;
        t = fmaf(vv, vv, t);
      } else if (lane == kfm) {
        s += ldg_nc_f1(w1 + js);
      }
    }
    float r = (lane < kfm) ? 0.5f * (s * s - t) : (lane == kfm ? s : 0.f);
    r = warp_sum(r);
    if (lane == 0) {
      r += __ldg(w0);
      out[b] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
    }
  }
}

// ---------------------------------------------------------------- R5 FFMLayer (reference formula, N4)
// field_f[i,:] = sum_j x_j v[j,i,:]  (a [field_num*k] vector: one gathered v-row per non-zero feature),
// inter = sum_{i<j} <field_f_i, field_f_j> = 0.5 (|sum_i field_f_i|^2 - sum_i |field_f_i|^2).
// Element e = i*k + c lives in lane e%32, slot e/32; k must divide 32 or be a multiple of 32 so
// that a lane's elements all share c = e % k (c group = lane % k for k<=32).
template <int IDKIND, int NS>
__global__ void __launch_bounds__(128)
ffm_kernel(const float* __restrict__ dense, int64_t dense_ld, int D, const void* __restrict__ ids,
           int64_t ids_ld, int F, const int64_t* __restrict__ col_off, int64_t B,
           const float* __restrict__ w0, const float* __restrict__ w, const float* __restrict__ v,
           int field_num, int k, float* __restrict__ out, int apply_sigmoid) {
  WARP_IDS();
  const int len = field_num * k;
  for (int64_t b = warp; b < B; b += nwarps) {
    float acc[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) acc[s] = 0.f;
    float lin = 0.f;
    for (int j = 0; j < D; ++j) {
      const float xs = __ldg(dense + b * dense_ld + j);
      if (xs == 0.f) continue;
      const float* vr = v + (size_t)j * len;
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const int e = lane + 32 * s;
        if (e < len) acc[s] = fmaf(xs, __ldg(vr + e), acc[s]);
      }
      if (lane == 0) lin = fmaf(xs, __ldg(w + j), lin);
    }
    for (int f = 0; f < F; ++f) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      const int64_t c0 = col_off[f], c1 = col_off[f + 1];
      if (id < 0 || id >= c1 - c0) continue;          // tf.one_hot: out-of-range -> zero row
      const float* vr = v + (size_t)(c0 + id) * len;
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const int e = lane + 32 * s;
        if (e < len) acc[s] += ldg_nc_f1(vr + e);
      }
      if (lane == 0) lin += ldg_nc_f1(w + c0 + id);
    }
    float sq = 0.f, tot = 0.f;
#pragma unroll
    for (int s = 0; s < NS; ++s) { sq = fmaf(acc[s], acc[s], sq); tot += acc[s]; }
    sq = warp_sum(sq);
    // tot currently sums this lane's elements (all share c = lane % k when k <= 32)
    float tsq;
    if (k <= 32) {
      for (int o = 16; o >= k; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
      tsq = (lane < k) ? tot * tot : 0.f;
      tsq = warp_sum(tsq);
    } else {
      // k multiple of 32: element e -> c = e % k = lane + 32*(s % (k/32)); regroup slots
      const int g = k >> 5;
      tsq = 0.f;
      for (int c = 0; c < g; ++c) {
        float tc = 0.f;
#pragma unroll
        for (int s = 0; s < NS; ++s) if ((s % g) == c) tc += acc[s];
        tsq = fmaf(tc, tc, tsq);
      }
      tsq = warp_sum(tsq);
    }
    if (lane == 0) {
      const float r = lin + __ldg(w0) + 0.5f * (tsq - sq);
      out[b] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
    }
  }
}

// ---------------------------------------------------------------- R3 FMLayer, dense x of up to 1024 columns in registers
// One warp per S samples: lane l holds the 16-byte pieces l, l + 32, ... of each row.  v is staged transposed in smem as
// sv[c][piece] = (v[4q][c] .. v[4q+3][c]) so that consecutive lanes read consecutive 16-byte words; two more planes carry w1
// and sum_c v[j][c]^2.  Per element: KF + 2 FMA; the KF + 2 partial sums of a sample are reduced by shuffles.
// out = w0 + x.w1 + 0.5 * (sum_c (x.v_c)^2 - sum_j x_j^2 sum_c v_jc^2)         (FMLayer.call, interaction.py:106-114)
template <int KF, int NQ, int S>
__global__ void __launch_bounds__(256)
fm_dense_reg_kernel(const float* __restrict__ x, int64_t x_ld, int n, int64_t B, const float* __restrict__ w0,
                    const float* __restrict__ w1, const float* __restrict__ v, int kfm, float* __restrict__ out,
                    int apply_sigmoid) {
  extern __shared__ float sm[];
  WARP_IDS();
  const int n4 = (n + 3) / 4;
  float4* sv = reinterpret_cast<float4*>(sm);                 // [KF + 2][n4]
  for (int i = threadIdx.x; i < (KF + 2) * n4; i += blockDim.x) {
    const int c = i / n4, q = i - c * n4;
    float r[4];
#pragma unroll
    for (int e2 = 0; e2 < 4; ++e2) {
      const int j = 4 * q + e2;
      float t = 0.f;
      if (j < n) {
        if (c < kfm) t = __ldg(v + (size_t)j * kfm + c);
        else if (c == KF) t = __ldg(w1 + j);
        else if (c == KF + 1) { for (int cc = 0; cc < kfm; ++cc) { const float u = __ldg(v + (size_t)j * kfm + cc); t = fmaf(u, u, t); } }
      }
      r[e2] = t;
    }
    sv[i] = make_float4(r[0], r[1], r[2], r[3]);
  }
  __syncthreads();
  const float bias = __ldg(w0);
  // the rows of the next S samples are requested as soon as the FMAs over the current ones are done (xv is dead from there on),
  // so their latency runs under the shuffle reductions; the first version's capture had 57 % of its stall samples on the first use
  // of a freshly loaded row at 16 resident warps per SM
  float4 xv[S][NQ];
  auto fetch = [&](int64_t b0) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        xv[s][i] = (q < n4 && b0 + s < B) ? ldg_nc_f4(x + (b0 + s) * x_ld + 4 * q) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  };
  fetch(warp * S);
  for (int64_t b0 = warp * S; b0 < B; b0 += nwarps * S) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        if (q < n4 && 4 * q + 3 >= n) {                  // last piece: columns past n are not part of x
          if (4 * q + 1 >= n) xv[s][i].y = 0.f;
          if (4 * q + 2 >= n) xv[s][i].z = 0.f;
          xv[s][i].w = 0.f;
        }
      }
    }
    float acc[S][KF + 2];
#pragma unroll
    for (int s = 0; s < S; ++s)
#pragma unroll
      for (int c = 0; c < KF + 2; ++c) acc[s][c] = 0.f;
#pragma unroll
    for (int i = 0; i < NQ; ++i) {
      const int q = lane + 32 * i;
      if (q < n4) {
#pragma unroll
        for (int c = 0; c < KF + 1; ++c) {
          const float4 w = sv[c * n4 + q];
#pragma unroll
          for (int s = 0; s < S; ++s)
            acc[s][c] = fmaf(xv[s][i].x, w.x, fmaf(xv[s][i].y, w.y, fmaf(xv[s][i].z, w.z, fmaf(xv[s][i].w, w.w, acc[s][c]))));
        }
        const float4 w = sv[(KF + 1) * n4 + q];
#pragma unroll
        for (int s = 0; s < S; ++s)
          acc[s][KF + 1] = fmaf(xv[s][i].x * xv[s][i].x, w.x, fmaf(xv[s][i].y * xv[s][i].y, w.y,
                           fmaf(xv[s][i].z * xv[s][i].z, w.z, fmaf(xv[s][i].w * xv[s][i].w, w.w, acc[s][KF + 1]))));
      }
    }
    fetch(b0 + nwarps * S);
#pragma unroll
    for (int s = 0; s < S; ++s) {
      // reduce-scatter of the KF + 2 partial sums (padded to 16): 31 shuffles instead of 5 per value, then one more
      // warp sum of the per-factor terms (shuffles, not FMAs, are what this kernel issues most of)
      float r16[16];
#pragma unroll
      for (int c = 0; c < 16; ++c) r16[c] = c < KF + 2 ? acc[s][c] : 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c) r16[c] += __shfl_xor_sync(0xffffffffu, r16[c], 16);
#pragma unroll
      for (int o = 8; o >= 1; o >>= 1) {
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int c = 0; c < o; ++c) {
          const float send = up ? r16[c] : r16[c + o];
          const float keep = up ? r16[c + o] : r16[c];
          r16[c] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
      }
      const float t = r16[0];                              // lane l: total of value l % 16
      float term = 0.f;
      if (lane < KF) term = 0.5f * t * t;
      else if (lane == KF) term = t;
      else if (lane == KF + 1) term = -0.5f * t;
      term = warp_sum(term);
      if (lane == 0 && b0 + s < B) {
        const float r = bias + term;
        out[b0 + s] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
      }
    }
  }
}

// ---------------------------------------------------------------- R6 InnerProductLayer
// One CTA per TS samples: E tile [TS][F*k] staged in smem, thread p computes <e_i, e_j>.
__global__ void __launch_bounds__(256)
pnn_inner_kernel(const float* __restrict__ e, int64_t e_ld, int F, int k, int64_t B,
                 float* __restrict__ out, int64_t out_ld) {
  extern __shared__ float sm[];
  const int P = F * (F - 1) / 2;
  const int fk = F * k;
  const int fkp = fk + 1;
  constexpr int TS = 8;
  int* prow = reinterpret_cast<int*>(sm + TS * fkp);
  int* pcol = prow + P;
  for (int i = threadIdx.x; i < F - 1; i += blockDim.x) {
    int base = i * (2 * F - i - 1) / 2;
    for (int j = i + 1; j < F; ++j) { prow[base + j - i - 1] = i; pcol[base + j - i - 1] = j; }
  }
  for (int64_t b0 = (int64_t)blockIdx.x * TS; b0 < B; b0 += (int64_t)gridDim.x * TS) {
    __syncthreads();
    for (int t = threadIdx.x; t < TS * fk; t += blockDim.x) {
      const int s = t / fk, c = t - s * fk;
      sm[s * fkp + c] = (b0 + s < B) ? ldg_nc_f1(e + (b0 + s) * e_ld + c) : 0.f;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < TS * P; t += blockDim.x) {
      const int s = t / P, p = t - s * P;
      if (b0 + s >= B) continue;
      const float* a = sm + s * fkp + prow[p] * k;
      const float* c = sm + s * fkp + pcol[p] * k;
      float acc = 0.f;
      for (int d = 0; d < k; ++d) acc = fmaf(a[d], c[d], acc);
      out[(b0 + s) * out_ld + p] = acc;
    }
  }
}

// ---------------------------------------------------------------- R6 InnerProductLayer, warp = sample, lane = field
// Lane i < F holds row e_i in registers (one coalesced 128-bit-per-piece read of the sample).  Round d = 1 .. F/2: every lane
// fetches row (i + d) mod F from its owner by warp shuffles and forms one dot product, so each unordered pair is computed
// once and the work is balanced ((F-1)/2 rounds for all lanes, plus half a round when F is even).  The 325 results are put
// in pair order in a per-warp smem row and written as consecutive floats.  No smem traffic for the operands at all.
// (Measured before: thread per (sample, pair) with scalar smem reads = 0.07 of HBM; a lane-per-sample tile variant was
// latency-bound at one warp per scheduler and slower still.)
template <int K>
__global__ void __launch_bounds__(256)
pnn_inner_rot_kernel(const float* __restrict__ e, int64_t e_ld, int F, int64_t B, float* __restrict__ out, int64_t out_ld) {
  extern __shared__ float sm[];
  WARP_IDS();
  constexpr int K4 = K / 4;
  const int P = F * (F - 1) / 2;
  float* st = sm + (size_t)(threadIdx.x >> 5) * P;
  const int full = (F - 1) / 2;                       // rounds in which every lane owns a pair
  const bool half = (F & 1) == 0;                     // F even: round F/2 pairs lane i < F/2 with i + F/2
  const bool act = lane < F;
  for (int64_t b = warp; b < B; b += nwarps) {
    float a[K];
#pragma unroll
    for (int c = 0; c < K4; ++c) {
      const float4 v = act ? ldg_nc_f4(e + b * e_ld + (lane * K4 + c) * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      a[4 * c] = v.x; a[4 * c + 1] = v.y; a[4 * c + 2] = v.z; a[4 * c + 3] = v.w;
    }
    __syncwarp();                                     // the previous sample's results have left st
    for (int d = 1; d <= full + (half ? 1 : 0); ++d) {
      int j = lane + d;
      if (j >= F) j -= F;
      const int src = act ? j : lane;
      float dot = 0.f;
#pragma unroll
      for (int c = 0; c < K; ++c) dot = fmaf(a[c], __shfl_sync(0xffffffffu, a[c], src), dot);
      const bool mine = act && (d <= full || lane < F / 2);
      if (mine) {
        const int lo = lane < j ? lane : j, hi = lane < j ? j : lane;
        st[lo * (2 * F - lo - 1) / 2 + hi - lo - 1] = dot;
      }
    }
    __syncwarp();
    for (int p = lane; p < P; p += 32) out[b * out_ld + p] = st[p];
  }
}

// ---------------------------------------------------------------- R7 OuterProductLayer
// out[b,p] = sum_a q[a] * (sum_c W[a,p,c] p_[c]),  p_=e_i, q=e_j.  CTA = (pair chunk, sample tile):
// W[:,p,:] for PC pairs staged in smem once and reused by TS samples.
__global__ void __launch_bounds__(256)
pnn_outer_kernel(const float* __restrict__ e, int64_t e_ld, int F, int k, int64_t B,
                 const float* __restrict__ W, float* __restrict__ out, int64_t out_ld, int PC, int TS) {
  extern __shared__ float sm[];
  const int P = F * (F - 1) / 2;
  const int fk = F * k;
  const int kk = k * k;
  float* sW = sm;                                  // [PC][k*k + 1]  (a-major: W[a,p,c] -> sW[pp][a*k+c])
  float* sE = sW + (size_t)PC * (kk + 1);          // [TS][fk + 1]
  const int p0 = blockIdx.x * PC;
  const int np = min(PC, P - p0);
  for (int t = threadIdx.x; t < np * kk; t += blockDim.x) {
    const int pp = t / kk, r = t - pp * kk;
    const int a = r / k, c = r - a * k;
    sW[(size_t)pp * (kk + 1) + r] = __ldg(W + ((size_t)a * P + (p0 + pp)) * k + c);
  }
  for (int64_t b0 = (int64_t)blockIdx.y * TS; b0 < B; b0 += (int64_t)gridDim.y * TS) {
    __syncthreads();
    for (int t = threadIdx.x; t < TS * fk; t += blockDim.x) {
      const int s = t / fk, c = t - s * fk;
      sE[s * (fk + 1) + c] = (b0 + s < B) ? ldg_nc_f1(e + (b0 + s) * e_ld + c) : 0.f;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < TS * np; t += blockDim.x) {
      const int s = t / np, pp = t - s * np;
      if (b0 + s >= B) continue;
      const int p = p0 + pp;
      // invert row-major pair index p -> (i,j)
      int i = 0, rem = p;
      while (rem >= F - 1 - i) { rem -= F - 1 - i; ++i; }
      const int j = i + 1 + rem;
      const float* pi = sE + s * (fk + 1) + i * k;
      const float* qj = sE + s * (fk + 1) + j * k;
      const float* w = sW + (size_t)pp * (kk + 1);
      float acc = 0.f;
      for (int a = 0; a < k; ++a) {
        float t2 = 0.f;
        for (int c = 0; c < k; ++c) t2 = fmaf(w[a * k + c], pi[c], t2);
        acc = fmaf(t2, qj[a], acc);
      }
      out[(b0 + s) * out_ld + p] = acc;
    }
  }
}

// ---------------------------------------------------------------- R7 OuterProductLayer, lane = sample
// out[b,p] = sum_a q[a] * (sum_c W[a,p,c] p_[c]) is 272 FMA per (sample, pair): compute-bound (88 K FMA per sample at F = 26,
// k = 16; 156 us at the fp32 peak for B = 65 536).  A warp owns 32 samples: their rows sit in a smem tile whose row stride is
// an odd number of 16-byte chunks (lane s reads ITS sample conflict-free), p_ = e_i and q = e_j are held in registers, and
// W[:, p, :] of a chunk of PC pairs is staged in smem once per CTA pass and read as warp-wide BROADCASTS, one LDS.128 per 4
// FMA, 16 independent accumulation chains per pair.  (The old kernel read W and both rows by scalar LDS per FMA: 4.2 ms.)
template <int K>
__global__ void __launch_bounds__(128)
pnn_outer_ls_kernel(const float* __restrict__ e, int64_t e_ld, int F, int64_t B, const float* __restrict__ W,
                    float* __restrict__ out, int64_t out_ld, int ES, int warps, int PC) {
  extern __shared__ float sm[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  constexpr int K4 = K / 4, KK = K * K;
  const int P = F * (F - 1) / 2;
  const int fk4 = F * K4;
  float* sW = sm;                                            // [PC][K][K]
  float* tile = sm + (size_t)PC * KK + (size_t)wid * 32 * ES;
  const float* my = tile + (size_t)lane * ES;
  const int64_t n_tiles = (B + 31) / 32;
  const int64_t n_groups = (n_tiles + warps - 1) / warps;
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t t = g * warps + wid;
    const bool live = (wid < warps) && (t < n_tiles);
    const int64_t b0 = t * 32;
    // W chunk of PC = 2048 / K^2 pairs = 16 floats per thread, fetched into registers one chunk AHEAD of its use: the L2
    // round trip of chunk n+1 runs under the FMAs of chunk n (it was 38 % of the stall samples when loaded in place)
    float wr[16];
    auto wload = [&](int p0) {
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const int idx = (int)threadIdx.x + 128 * u;
        const int pp = idx / KK, r = idx - pp * KK;
        const int a = r / K, c = r - a * K;
        wr[u] = (p0 + pp < P) ? __ldg(W + ((size_t)a * P + (p0 + pp)) * K + c) : 0.f;
      }
    };
    wload(0);
    if (live) {
      for (int s2 = 0; s2 < 32; s2 += 2) {                   // two samples' loads in flight per pass
        const int64_t ba = b0 + s2, bb = ba + 1;
        for (int c = lane; c < fk4; c += 32) {
          const float4 va = (ba < B) ? ldg_nc_f4(e + ba * e_ld + 4 * c) : make_float4(0.f, 0.f, 0.f, 0.f);
          const float4 vb = (bb < B) ? ldg_nc_f4(e + bb * e_ld + 4 * c) : make_float4(0.f, 0.f, 0.f, 0.f);
          *reinterpret_cast<float4*>(tile + (size_t)s2 * ES + 4 * c) = va;
          *reinterpret_cast<float4*>(tile + (size_t)(s2 + 1) * ES + 4 * c) = vb;
        }
      }
    }
    const int64_t b = b0 + lane;
    int i = 0, j = 1;                                        // pair p = (i, j), row-major over i < j
    for (int p0 = 0; p0 < P; p0 += PC) {
      const int np = min(PC, P - p0);
      __syncthreads();                                       // the previous chunk of W has been used by every warp
#pragma unroll
      for (int u = 0; u < 16; ++u) sW[(int)threadIdx.x + 128 * u] = wr[u];
      __syncthreads();
      if (p0 + PC < P) wload(p0 + PC);
      if (!live) {                                           // keep (i, j) in step; nothing to compute
        continue;
      }
      float pi[K];
      int i_loaded = -1;
      for (int pp = 0; pp < np; ++pp) {
        if (i != i_loaded) {
#pragma unroll
          for (int c = 0; c < K4; ++c) {
            const float4 v = *reinterpret_cast<const float4*>(my + (i * K4 + c) * 4);
            pi[4 * c] = v.x; pi[4 * c + 1] = v.y; pi[4 * c + 2] = v.z; pi[4 * c + 3] = v.w;
          }
          i_loaded = i;
        }
        float qj[K];
#pragma unroll
        for (int c = 0; c < K4; ++c) {
          const float4 v = *reinterpret_cast<const float4*>(my + (j * K4 + c) * 4);
          qj[4 * c] = v.x; qj[4 * c + 1] = v.y; qj[4 * c + 2] = v.z; qj[4 * c + 3] = v.w;
        }
        const float4* w = reinterpret_cast<const float4*>(sW + (size_t)pp * KK);
        // K independent chains t2[a] (c outermost): the dependent-FMA latency is hidden inside one warp, which matters at one
        // warp per scheduler (a 54 KB tile per warp); summation order per chain is still c = 0 .. K-1
        float t2[K];
#pragma unroll
        for (int a = 0; a < K; ++a) t2[a] = 0.f;
#pragma unroll
        for (int c = 0; c < K4; ++c) {
#pragma unroll
          for (int a = 0; a < K; ++a) {
            const float4 ww = w[a * K4 + c];
            t2[a] = fmaf(ww.x, pi[4 * c], t2[a]); t2[a] = fmaf(ww.y, pi[4 * c + 1], t2[a]);
            t2[a] = fmaf(ww.z, pi[4 * c + 2], t2[a]); t2[a] = fmaf(ww.w, pi[4 * c + 3], t2[a]);
          }
        }
        float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
        for (int a = 0; a < K; a += 2) { acc0 = fmaf(t2[a], qj[a], acc0); acc1 = fmaf(t2[a + 1], qj[a + 1], acc1); }
        const float acc = acc0 + acc1;
        if (b < B) out[b * out_ld + p0 + pp] = acc;
        if (++j == F) { ++i; j = i + 1; }
      }
    }
    __syncthreads();                                         // tiles are rewritten by the next group
  }
}

// ---------------------------------------------------------------- R8 CrossLayer, rows of up to 32 * 4 * NQ floats in registers
// One warp per sample, S samples per pass: lane l holds the 16-byte pieces l, l + 32, ... of x_0 and x_l; the L weight and
// bias vectors sit in smem (zero-padded to a multiple of 4, so the padding of the last piece contributes nothing).
// HBM-bound: one read of x, one write of the output (or 4 bytes when only the dot with final_w is wanted).
template <int NQ, int S>
__global__ void __launch_bounds__(256)
cross_reg_kernel(const float* __restrict__ x, int64_t x_ld, int d, int64_t B, int L,
                 const float* __restrict__ w, const float* __restrict__ bia, float* __restrict__ out,
                 int64_t out_ld, const float* __restrict__ final_w, float* __restrict__ dot_out) {
  extern __shared__ float sm[];
  WARP_IDS();
  const int d4 = (d + 3) / 4, dp = d4 * 4;
  float* s_w = sm;                       // [L][dp]
  float* s_b = sm + (size_t)L * dp;      // [L][dp]
  float* s_f = s_b + (size_t)L * dp;     // [dp]
  for (int i = threadIdx.x; i < L * dp; i += blockDim.x) {
    const int l = i / dp, j = i - l * dp;
    s_w[i] = j < d ? __ldg(w + (size_t)l * d + j) : 0.f;
    s_b[i] = j < d ? __ldg(bia + (size_t)l * d + j) : 0.f;
  }
  for (int j = threadIdx.x; j < dp; j += blockDim.x) s_f[j] = (final_w && j < d) ? __ldg(final_w + j) : 0.f;
  __syncthreads();
  const bool out_vec = out && ((out_ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
  const bool out_full = out_vec && out_ld >= dp;       // rows padded to whole 16-byte pieces: the last piece is stored whole (its
                                                       // columns past d are zeros: x, w and b are zero there)
  // the rows of the NEXT S samples are fetched while the current ones are worked on (the capture of the first version showed
  // 46 % of the warp slots filled and a third of all stall samples on the first use of the loaded row: latency, not bandwidth)
  float4 xn[S][NQ];
  auto fetch = [&](int64_t b0) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        xn[s][i] = (q < d4 && b0 + s < B) ? ldg_nc_f4(x + (b0 + s) * x_ld + 4 * q) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  };
  fetch(warp * S);
  for (int64_t b0 = warp * S; b0 < B; b0 += nwarps * S) {
    float4 x0[S][NQ], xl[S][NQ];
#pragma unroll
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        float4 v = xn[s][i];
        if (q < d4 && 4 * q + 3 >= d) {                // last piece: whatever lies behind column d is not part of x
          if (4 * q + 1 >= d) v.y = 0.f;
          if (4 * q + 2 >= d) v.z = 0.f;
          v.w = 0.f;
        }
        x0[s][i] = v; xl[s][i] = v;
      }
    }
    fetch(b0 + nwarps * S);
    for (int l = 0; l < L; ++l) {
      const float4* wl = reinterpret_cast<const float4*>(s_w + (size_t)l * dp);
      const float4* bl = reinterpret_cast<const float4*>(s_b + (size_t)l * dp);
      float dot[S];
#pragma unroll
      for (int s = 0; s < S; ++s) dot[s] = 0.f;
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        if (q < d4) {
          const float4 ww = wl[q];
#pragma unroll
          for (int s = 0; s < S; ++s)
            dot[s] = fmaf(xl[s][i].x, ww.x, fmaf(xl[s][i].y, ww.y, fmaf(xl[s][i].z, ww.z, fmaf(xl[s][i].w, ww.w, dot[s]))));
        }
      }
#pragma unroll
      for (int s = 0; s < S; ++s) dot[s] = warp_sum(dot[s]);
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        if (q < d4) {
          const float4 bb = bl[q];
#pragma unroll
          for (int s = 0; s < S; ++s) {
            xl[s][i].x = fmaf(x0[s][i].x, dot[s], bb.x) + xl[s][i].x;
            xl[s][i].y = fmaf(x0[s][i].y, dot[s], bb.y) + xl[s][i].y;
            xl[s][i].z = fmaf(x0[s][i].z, dot[s], bb.z) + xl[s][i].z;
            xl[s][i].w = fmaf(x0[s][i].w, dot[s], bb.w) + xl[s][i].w;
          }
        }
      }
    }
    if (out) {
#pragma unroll
      for (int s = 0; s < S; ++s) {
        if (b0 + s >= B) continue;
        float* row = out + (b0 + s) * out_ld;
#pragma unroll
        for (int i = 0; i < NQ; ++i) {
          const int q = lane + 32 * i;
          if (q >= d4) continue;
          if (out_vec && (4 * q + 3 < d || out_full)) stg_na_f4(row + 4 * q, xl[s][i]);
        }
        if (!out_vec || ((d & 3) && !out_full)) {
          // rows that are not 16-byte aligned (out_ld = d = 429): the row goes through a per-warp smem slab so that every
          // store instruction writes 32 CONSECUTIVE floats (4 sectors per instruction instead of 16; shuffles were slower)
          float* slab = s_f + dp + (size_t)(threadIdx.x >> 5) * dp;
          __syncwarp();
#pragma unroll
          for (int i = 0; i < NQ; ++i) {
            const int q = lane + 32 * i;
            if (q < d4) *reinterpret_cast<float4*>(slab + 4 * q) = xl[s][i];
          }
          __syncwarp();
          // with aligned rows only the tail piece (4 q + 3 >= d) is still to be written
          for (int col = out_vec ? (d & ~3) + lane : lane; col < d; col += 32) row[col] = slab[col];
        }
      }
    }
    if (final_w) {
      const float4* fw = reinterpret_cast<const float4*>(s_f);
      float dot[S];
#pragma unroll
      for (int s = 0; s < S; ++s) dot[s] = 0.f;
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        if (q < d4) {
          const float4 ww = fw[q];
#pragma unroll
          for (int s = 0; s < S; ++s)
            dot[s] = fmaf(xl[s][i].x, ww.x, fmaf(xl[s][i].y, ww.y, fmaf(xl[s][i].z, ww.z, fmaf(xl[s][i].w, ww.w, dot[s]))));
        }
      }
#pragma unroll
      for (int s = 0; s < S; ++s) {
        dot[s] = warp_sum(dot[s]);
        if (lane == 0 && b0 + s < B) dot_out[b0 + s] = dot[s];
      }
    }
  }
}

// ---------------------------------------------------------------- R8 CrossLayer (x already in HBM)
// One warp per sample, x0 / x_l held in a per-warp smem slab (any d); all L layers in one launch.
__global__ void __launch_bounds__(128)
cross_kernel(const float* __restrict__ x, int64_t x_ld, int d, int64_t B, int L,
             const float* __restrict__ w, const float* __restrict__ bia, float* __restrict__ out,
             int64_t out_ld, const float* __restrict__ final_w, float* __restrict__ dot_out) {
  extern __shared__ float sm[];
  WARP_IDS();
  float* x0 = sm + (size_t)(threadIdx.x >> 5) * 2 * d;
  float* xl = x0 + d;
  for (int64_t b = warp; b < B; b += nwarps) {
    for (int j = lane; j < d; j += 32) { const float t = ldg_nc_f1(x + b * x_ld + j); x0[j] = t; xl[j] = t; }
    __syncwarp();
    for (int l = 0; l < L; ++l) {
      const float* wl = w + (size_t)l * d;
      const float* bl = bia + (size_t)l * d;
      float s = 0.f;
      for (int j = lane; j < d; j += 32) s = fmaf(xl[j], __ldg(wl + j), s);
      s = warp_sum(s);
      for (int j = lane; j < d; j += 32) xl[j] = fmaf(x0[j], s, __ldg(bl + j)) + xl[j];
      __syncwarp();
    }
    if (out) for (int j = lane; j < d; j += 32) out[b * out_ld + j] = xl[j];
    if (final_w) {
      float s = 0.f;
      for (int j = lane; j < d; j += 32) s = fmaf(xl[j], __ldg(final_w + j), s);
      s = warp_sum(s);
      if (lane == 0) dot_out[b] = s;
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------- R10 Dice (elementwise), 128-bit path
// Per-column constants (alpha, mean, 1/sqrt(var + eps)) staged in smem once per CTA; the column of a thread's float4 is
// advanced incrementally (no 64-bit modulo per element, which was most of the old kernel's instruction count).
__global__ void __launch_bounds__(256)
dice_vec_kernel(const float* __restrict__ x, int64_t total4, int h4, const float* __restrict__ alphas,
                const float* __restrict__ mean, const float* __restrict__ var, float eps, float* __restrict__ out) {
  extern __shared__ float sm[];
  const int h = h4 * 4;
  float* s_a = sm; float* s_m = sm + h; float* s_r = sm + 2 * h;
  for (int c = threadIdx.x; c < h; c += blockDim.x) {
    s_a[c] = __ldg(alphas + c); s_m[c] = __ldg(mean + c); s_r[c] = 1.0f / sqrtf(__ldg(var + c) + eps);
  }
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c4 = (int)(i % h4);
  const int step = (int)(stride % h4);
  auto one = [&](float t, int c) {
    const float p = 1.0f / (1.0f + expf(-(t - s_m[c]) * s_r[c]));
    return s_a[c] * (1.0f - p) * t + p * t;
  };
  // four loads in flight per thread (one was 32 KB per SM at full occupancy: short of what the HBM latency needs)
  for (; i + 3 * stride < total4; i += 4 * stride) {
    float4 v[4];
    int c[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      v[u] = ldg_nc_f4(x + (i + u * stride) * 4);
      c[u] = c4 * 4;
      c4 += step;
      if (c4 >= h4) c4 -= h4;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      stg_na_f4(out + (i + u * stride) * 4, make_float4(one(v[u].x, c[u]), one(v[u].y, c[u] + 1), one(v[u].z, c[u] + 2), one(v[u].w, c[u] + 3)));
  }
  for (; i < total4; i += stride) {
    const float4 v = ldg_nc_f4(x + i * 4);
    const int c = c4 * 4;
    stg_na_f4(out + i * 4, make_float4(one(v.x, c), one(v.y, c + 1), one(v.z, c + 2), one(v.w, c + 3)));
    c4 += step;
    if (c4 >= h4) c4 -= h4;
  }
}

// ---------------------------------------------------------------- R10 Dice (elementwise)
__global__ void dice_kernel(const float* __restrict__ x, int64_t total, int h,
                            const float* __restrict__ alphas, const float* __restrict__ mean,
                            const float* __restrict__ var, float eps, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % h);
    const float t = x[i];
    const float p = 1.0f / (1.0f + expf(-(t - __ldg(mean + c)) / sqrtf(__ldg(var + c) + eps)));
    out[i] = __ldg(alphas + c) * (1.0f - p) * t + p * t;
  }
}

static inline int warp_grid(int64_t B, int warps_per_cta) {
  int64_t g = (B + warps_per_cta - 1) / warps_per_cta;
  const int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace ctrhot

using namespace ctrhot;
#define ST(s) static_cast<cudaStream_t>(s)
#define IDS_OK(kind) ((kind) == CTRHOT_IDS_I32 || (kind) == CTRHOT_IDS_F32)

extern "C" int ctrhot_wide_dense_f32(const float* x, int64_t x_ld, int64_t n, int64_t B,
                                     const float* w0, const float* w, float* out, void* stream) {
  CTRHOT_RANGE("ctrhot_wide_dense_f32");
  if (!x || !w0 || !w || !out || n <= 0 || B < 0 || x_ld < n) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  wide_dense_kernel<<<warp_grid(B, 8), 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w, out);
  CTRHOT_LAUNCH_OK("ctrhot_wide_dense_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_wide_onehot_f32(const float* dense, int64_t dense_ld, int D,
                                      const int32_t* dense_col, const void* ids, int ids_kind,
                                      int64_t ids_ld, int F, const int64_t* col_off, int64_t B,
                                      const float* w0, const float* w, float* out, void* stream) {
  CTRHOT_RANGE("ctrhot_wide_onehot_f32");
  if (!w0 || !w || !out || B < 0 || D < 0 || F < 0 || !IDS_OK(ids_kind)) return CTRHOT_EARG;
  if ((D > 0 && (!dense || !dense_col)) || (F > 0 && (!ids || !col_off))) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  if (ids_kind == CTRHOT_IDS_I32)
    wide_onehot_kernel<CTRHOT_IDS_I32><<<warp_grid(B, 8), 256, 0, ST(stream)>>>(
        dense, dense_ld, D, dense_col, ids, ids_ld, F, col_off, B, w0, w, out);
  else
    wide_onehot_kernel<CTRHOT_IDS_F32><<<warp_grid(B, 8), 256, 0, ST(stream)>>>(
        dense, dense_ld, D, dense_col, ids, ids_ld, F, col_off, B, w0, w, out);
  CTRHOT_LAUNCH_OK("ctrhot_wide_onehot_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_fm_dense_f32(const float* x, int64_t x_ld, int64_t n, int64_t B,
                                   const float* w0, const float* w1, const float* v, int kfm,
                                   float* out, int apply_sigmoid, void* stream) {
  CTRHOT_RANGE("ctrhot_fm_dense_f32");
  if (!x || !w0 || !w1 || !out || n <= 0 || B < 0 || x_ld < n || kfm < 0) return CTRHOT_EARG;
  if (kfm > 0 && !v) return CTRHOT_EARG;
  if (kfm > 31) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  if (n <= 512 && kfm >= 1 && kfm <= 12 && x_ld % 4 == 0 && aligned16(x)) {
    // rows fit the registers of a warp (DeepFM's x = [dense | embed], n = 429): register-resident kernel
    const int n4 = (int)((n + 3) / 4), nq = (n4 + 31) / 32;
    const size_t sm2 = (size_t)(12 + 2) * n4 * sizeof(float4);
#define FM_LAUNCH(KF, NQ, S)                                                                                          \
    fm_dense_reg_kernel<KF, NQ, S><<<fm_grid, 256, (size_t)(KF + 2) * n4 * sizeof(float4), ST(stream)>>>(             \
        x, x_ld, (int)n, B, w0, w1, v, kfm, out, apply_sigmoid)
    (void)sm2;
    // persistent: the weight planes are staged once per CTA, so no more CTAs than stay resident (2 per SM at 91 registers)
    const int fm_grid = (int)std::min<int64_t>((B + 15) / 16, (int64_t)kNumSMs * 2);
    if (kfm <= 4) { if (nq <= 1) FM_LAUNCH(4, 1, 4); else if (nq <= 2) FM_LAUNCH(4, 2, 2); else FM_LAUNCH(4, 4, 2); }
    else if (kfm <= 8) { if (nq <= 1) FM_LAUNCH(8, 1, 4); else if (nq <= 2) FM_LAUNCH(8, 2, 2); else FM_LAUNCH(8, 4, 2); }
    else { if (nq <= 1) FM_LAUNCH(12, 1, 4); else if (nq <= 2) FM_LAUNCH(12, 2, 2); else FM_LAUNCH(12, 4, 2); }
#undef FM_LAUNCH
    CTRHOT_LAUNCH_OK("ctrhot_fm_dense_f32");
    return CTRHOT_OK;
  }
  if (n >= 128 && kfm >= 1 && kfm <= 16) {                    // wide rows: four samples per warp, dense / sparse per 32 columns
    int64_t nb = ((B + 3) / 4 + 7) / 8;
    const int grid = (int)(nb < (int64_t)kNumSMs * 8 ? nb : (int64_t)kNumSMs * 8);
    const bool v4 = (x_ld % 4 == 0) && aligned16(x);
    if (v4 && kfm <= 8) fm_dense_wide_v4_kernel<8><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid);
    else if (v4 && kfm <= 12) fm_dense_wide_v4_kernel<12><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid);
    else if (v4) fm_dense_wide_v4_kernel<16><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid);
    else if (kfm <= 4) fm_dense_wide_kernel<4><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid);
    else if (kfm <= 8) fm_dense_wide_kernel<8><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid);
    else if (kfm <= 12) fm_dense_wide_kernel<12><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid);
    else fm_dense_wide_kernel<16><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid);
  } else {
    fm_dense_kernel<<<warp_grid(B, 8), 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid);
  }
  CTRHOT_LAUNCH_OK("ctrhot_fm_dense_f32");
  return CTRHOT_OK;
}

// ctrhot_fm_dense_f32 for FEW rows of MANY columns (FM on the reference's dense one-hot matrix, DL/model/fm.py:26-45): the columns
// are split over blockIdx.y so that the launch fills the GPU, partial sums in the caller's workspace, a finishing pass.
static int fm_dense_split_plan(int64_t n, int64_t B, int kfm, int* nsplit, int64_t* cols, int* KF) {
  *nsplit = 0;
  if (n < 4096 || kfm < 1 || kfm > 16 || B < 1) return 0;
  const int64_t nb = ((B + 3) / 4 + 7) / 8;
  if (nb >= (int64_t)kNumSMs * 2) return 0;                                 // enough row blocks already
  const int64_t tiles = (n + FMW_TJ - 1) / FMW_TJ;
  int64_t want = ((int64_t)kNumSMs * 4 + nb - 1) / nb;
  if (want > tiles / 4) want = tiles / 4;                                   // at least four 256-column tiles per split
  if (want < 2) return 0;
  const int64_t tps = (tiles + want - 1) / want;
  *cols = tps * FMW_TJ;
  *nsplit = (int)((tiles + tps - 1) / tps);
  *KF = kfm <= 8 ? 8 : (kfm <= 12 ? 12 : 16);
  return 1;
}
extern "C" int64_t ctrhot_fm_dense_ws_bytes(int64_t n, int64_t B, int kfm) {
  int ns, KF; int64_t cols;
  if (!fm_dense_split_plan(n, B, kfm, &ns, &cols, &KF)) return 0;
  return (int64_t)ns * B * (KF + 2) * (int64_t)sizeof(float);
}
extern "C" int ctrhot_fm_dense_ws_f32(const float* x, int64_t x_ld, int64_t n, int64_t B, const float* w0, const float* w1,
                                      const float* v, int kfm, float* out, int apply_sigmoid, void* workspace,
                                      int64_t workspace_bytes, void* stream) {
  CTRHOT_RANGE("ctrhot_fm_dense_ws_f32");
  int ns, KF; int64_t cols;
  if (!x || !w0 || !w1 || !out || n <= 0 || B < 0 || x_ld < n || kfm < 0 || (kfm > 0 && !v)) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  if (!workspace || !(x_ld % 4 == 0 && aligned16(x)) || !fm_dense_split_plan(n, B, kfm, &ns, &cols, &KF) ||
      workspace_bytes < (int64_t)ns * B * (KF + 2) * (int64_t)sizeof(float))
    return ctrhot_fm_dense_f32(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid, stream);
  float* part = static_cast<float*>(workspace);
  const int64_t nb = ((B + 3) / 4 + 7) / 8;
  dim3 grid((unsigned)nb, (unsigned)ns);
  if (KF == 8) fm_dense_wide_v4_kernel<8><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid, part, cols);
  else if (KF == 12) fm_dense_wide_v4_kernel<12><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid, part, cols);
  else fm_dense_wide_v4_kernel<16><<<grid, 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out, apply_sigmoid, part, cols);
  fm_dense_finish_kernel<<<(int)((B + 255) / 256), 256, 0, ST(stream)>>>(part, ns, B, KF + 2, w0, out, apply_sigmoid);
  CTRHOT_LAUNCH_OK("ctrhot_fm_dense_ws_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_fm_onehot_f32(const float* dense, int64_t dense_ld, int D,
                                    const int32_t* dense_col, const void* ids, int ids_kind,
                                    int64_t ids_ld, int F, const int64_t* col_off, int64_t B,
                                    const float* w0, const float* w1, const float* v, int kfm,
                                    float* out, int apply_sigmoid, void* stream) {
  CTRHOT_RANGE("ctrhot_fm_onehot_f32");
  if (!w0 || !w1 || !out || B < 0 || D < 0 || F < 0 || kfm < 0 || !IDS_OK(ids_kind)) return CTRHOT_EARG;
  if ((D > 0 && (!dense || !dense_col)) || (F > 0 && (!ids || !col_off)) || (kfm > 0 && !v)) return CTRHOT_EARG;
  if (kfm > 31) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  if (ids_kind == CTRHOT_IDS_I32)
    fm_onehot_kernel<CTRHOT_IDS_I32><<<warp_grid(B, 8), 256, 0, ST(stream)>>>(
        dense, dense_ld, D, dense_col, ids, ids_ld, F, col_off, B, w0, w1, v, kfm, out, apply_sigmoid);
  else
    fm_onehot_kernel<CTRHOT_IDS_F32><<<warp_grid(B, 8), 256, 0, ST(stream)>>>(
        dense, dense_ld, D, dense_col, ids, ids_ld, F, col_off, B, w0, w1, v, kfm, out, apply_sigmoid);
  CTRHOT_LAUNCH_OK("ctrhot_fm_onehot_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_ffm_onehot_f32(const float* dense, int64_t dense_ld, int D, const void* ids,
                                     int ids_kind, int64_t ids_ld, int F, const int64_t* col_off,
                                     int64_t B, const float* w0, const float* w, const float* v,
                                     int field_num, int k, float* out, int apply_sigmoid, void* stream) {
  CTRHOT_RANGE("ctrhot_ffm_onehot_f32");
  if (!w0 || !w || !v || !out || B < 0 || D < 0 || F < 0 || field_num <= 0 || k <= 0 || !IDS_OK(ids_kind))
    return CTRHOT_EARG;
  if ((D > 0 && !dense) || (F > 0 && (!ids || !col_off))) return CTRHOT_EARG;
  const bool kok = (k <= 32) ? (32 % k == 0) : (k % 32 == 0);
  const int len = field_num * k;
  if (!kok || len > 64 * 32) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  const int ns = (len + 31) / 32;
  const int grid = warp_grid(B, 4);
#define L2(KIND, NS) ffm_kernel<KIND, NS><<<grid, 128, 0, ST(stream)>>>(dense, dense_ld, D, ids, ids_ld, F, \
      col_off, B, w0, w, v, field_num, k, out, apply_sigmoid)
#define L1(KIND) if (ns <= 4) L2(KIND, 4); else if (ns <= 8) L2(KIND, 8); else if (ns <= 16) L2(KIND, 16); \
      else if (ns <= 32) L2(KIND, 32); else L2(KIND, 64);
  if (ids_kind == CTRHOT_IDS_I32) { L1(CTRHOT_IDS_I32) } else { L1(CTRHOT_IDS_F32) }
#undef L1
#undef L2
  CTRHOT_LAUNCH_OK("ctrhot_ffm_onehot_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_pnn_inner_f32(const float* e, int64_t e_ld, int F, int k, int64_t B,
                                    float* out, int64_t out_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_pnn_inner_f32");
  const int P = F * (F - 1) / 2;
  if (!e || !out || F < 2 || k <= 0 || B < 0 || e_ld < (int64_t)F * k || out_ld < P) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  if ((k == 16 || k == 8) && F <= 32 && e_ld % 4 == 0 && aligned16(e)) {
    const size_t sm2 = (size_t)8 * P * sizeof(float);
    if (k == 16) pnn_inner_rot_kernel<16><<<warp_grid(B, 8), 256, sm2, ST(stream)>>>(e, e_ld, F, B, out, out_ld);
    else pnn_inner_rot_kernel<8><<<warp_grid(B, 8), 256, sm2, ST(stream)>>>(e, e_ld, F, B, out, out_ld);
    CTRHOT_LAUNCH_OK("ctrhot_pnn_inner_f32");
    return CTRHOT_OK;
  }
  const size_t smem = (size_t)8 * (F * k + 1) * sizeof(float) + (size_t)2 * P * sizeof(int);
  if (smem > 200 * 1024) return CTRHOT_ESHAPE;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_inner_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem), "pnn_inner attr");
  int64_t g = (B + 7) / 8;
  if (g > kNumSMs * 8) g = kNumSMs * 8;
  pnn_inner_kernel<<<(int)g, 256, smem, ST(stream)>>>(e, e_ld, F, k, B, out, out_ld);
  CTRHOT_LAUNCH_OK("ctrhot_pnn_inner_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_pnn_outer_f32(const float* e, int64_t e_ld, int F, int k, int64_t B,
                                    const float* W, float* out, int64_t out_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_pnn_outer_f32");
  const int P = F * (F - 1) / 2;
  if (!e || !W || !out || F < 2 || k <= 0 || B < 0 || e_ld < (int64_t)F * k || out_ld < P) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  if ((k == 16 || k == 8) && e_ld % 4 == 0 && aligned16(e)) {
    const int ES = ((F * k / 4) | 1) * 4;                                  // odd number of 16-byte chunks per tile row
    const size_t tile_bytes = (size_t)32 * ES * sizeof(float);
    int warps = 0, pc = 0;
    for (int cand : {2048 / (k * k)}) {                                    // pairs of W per smem chunk (16 floats per thread)
      const size_t wb = (size_t)cand * k * k * sizeof(float);
      if (wb + tile_bytes > 220 * 1024) continue;
      int wq = (int)((220 * 1024 - wb) / tile_bytes);
      if (wq > 4) wq = 4;
      if (wq > warps) { warps = wq; pc = cand; }
    }
    if (warps >= 1) {
      const size_t sm2 = (size_t)pc * k * k * sizeof(float) + tile_bytes * warps;
      const int64_t n_groups = ((B + 31) / 32 + warps - 1) / warps;
      const int g2 = (int)(n_groups < kNumSMs ? n_groups : kNumSMs);
      if (k == 16) {
        CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_outer_ls_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2),
                       "pnn_outer attr");
        pnn_outer_ls_kernel<16><<<g2, 128, sm2, ST(stream)>>>(e, e_ld, F, B, W, out, out_ld, ES, warps, pc);
      } else {
        CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_outer_ls_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2),
                       "pnn_outer attr");
        pnn_outer_ls_kernel<8><<<g2, 128, sm2, ST(stream)>>>(e, e_ld, F, B, W, out, out_ld, ES, warps, pc);
      }
      CTRHOT_LAUNCH_OK("ctrhot_pnn_outer_f32");
      return CTRHOT_OK;
    }
  }
  const int TS = 32;
  int PC = (int)((160 * 1024 / sizeof(float) - (size_t)TS * (F * k + 1)) / (k * k + 1));
  if (PC < 1) return CTRHOT_ESHAPE;
  if (PC > 32) PC = 32;
  if (PC > P) PC = P;
  const size_t smem = ((size_t)PC * (k * k + 1) + (size_t)TS * (F * k + 1)) * sizeof(float);
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_outer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem), "pnn_outer attr");
  const int gx = (P + PC - 1) / PC;
  int64_t gy = (B + TS - 1) / TS;
  const int64_t cap = (kNumSMs * 4 + gx - 1) / gx;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  pnn_outer_kernel<<<dim3(gx, (unsigned)gy), 256, smem, ST(stream)>>>(e, e_ld, F, k, B, W, out, out_ld, PC, TS);
  CTRHOT_LAUNCH_OK("ctrhot_pnn_outer_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_cross_f32(const float* x, int64_t x_ld, int d, int64_t B, int L,
                                const float* w, const float* b, float* out, int64_t out_ld,
                                const float* final_w, float* dot, void* stream) {
  CTRHOT_RANGE("ctrhot_cross_f32");
  if (!x || d <= 0 || B < 0 || L < 0 || x_ld < d || (L > 0 && (!w || !b))) return CTRHOT_EARG;
  if (!out && !final_w) return CTRHOT_EARG;
  if ((out && out_ld < d) || (final_w && !dot)) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  if (x_ld % 4 == 0 && aligned16(x) && d <= 1024 && L <= 16) {
    // rows fit the registers of a warp: register-resident kernel, 128-bit accesses
    const int d4 = (d + 3) / 4;
    const size_t sm2 = (size_t)(2 * L + 1 + 8) * d4 * 4 * sizeof(float);      // weights, biases, final_w, 8 per-warp slabs
    const int nq = (d4 + 31) / 32;
#define CROSS_LAUNCH(NQ, S)                                                                                         \
    do {                                                                                                              \
      CTRHOT_CUDA_OK(cudaFuncSetAttribute(cross_reg_kernel<NQ, S>, cudaFuncAttributeMaxDynamicSharedMemorySize,       \
                                          (int)sm2), "cross attr");                                                  \
      int occ = 0;                                                                                                    \
      CTRHOT_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, cross_reg_kernel<NQ, S>, 256, sm2), "cross occ"); \
      /* persistent: one wave of resident CTAs, every warp walks its share (a capped grid of several waves left 13 % idle) */ \
      int64_t g_ = (B + 8 * S - 1) / (8 * S);                                                                         \
      if (occ > 0 && g_ > (int64_t)kNumSMs * occ) g_ = (int64_t)kNumSMs * occ;                                        \
      cross_reg_kernel<NQ, S><<<(int)g_, 256, sm2, ST(stream)>>>(x, x_ld, d, B, L, w, b, out, out_ld, final_w, dot);  \
    } while (0)
    if (nq <= 1) CROSS_LAUNCH(1, 4);
    else if (nq <= 2) CROSS_LAUNCH(2, 4);
    else if (nq <= 4) CROSS_LAUNCH(4, 1);
    else CROSS_LAUNCH(8, 1);
#undef CROSS_LAUNCH
    CTRHOT_LAUNCH_OK("ctrhot_cross_f32");
    return CTRHOT_OK;
  }
  const size_t smem = (size_t)4 * 2 * d * sizeof(float);
  if (smem > 220 * 1024) return CTRHOT_ESHAPE;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(cross_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem), "cross attr");
  cross_kernel<<<warp_grid(B, 4), 128, smem, ST(stream)>>>(x, x_ld, d, B, L, w, b, out, out_ld, final_w, dot);
  CTRHOT_LAUNCH_OK("ctrhot_cross_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_dice_f32(const float* x, int64_t rows, int h, const float* alphas,
                               const float* mean, const float* var, float eps, float* out, void* stream) {
  CTRHOT_RANGE("ctrhot_dice_f32");
  if (!x || !alphas || !mean || !var || !out || rows < 0 || h <= 0) return CTRHOT_EARG;
  const int64_t total = rows * h;
  if (total == 0) return CTRHOT_OK;
  if (h % 4 == 0 && h <= 4096 && aligned16(x) && aligned16(out)) {
    const int64_t total4 = total / 4;
    int64_t g4 = (total4 + 255) / 256;
    if (g4 > kNumSMs * 8) g4 = kNumSMs * 8;
    dice_vec_kernel<<<(int)g4, 256, (size_t)3 * h * sizeof(float), ST(stream)>>>(x, total4, h / 4, alphas, mean, var, eps, out);
    CTRHOT_LAUNCH_OK("ctrhot_dice_f32");
    return CTRHOT_OK;
  }
  int64_t g = (total + 255) / 256;
  if (g > kNumSMs * 16) g = kNumSMs * 16;
  dice_kernel<<<(int)g, 256, 0, ST(stream)>>>(x, total, h, alphas, mean, var, eps, out);
  CTRHOT_LAUNCH_OK("ctrhot_dice_f32");
  return CTRHOT_OK;
}
