// K1: embedding-row gather (EmbedLayer.call, DL/layer/core.py:273-280) and the fused model
// front ends built on it (DeepFM: gather -> FM; DCN: gather -> L cross layers).
//
// Layout: one warp owns S samples at a time.  The F*k floats of a sample are F*k/4 128-bit
// chunks; lane l owns chunks q = l, l+32, ... so k/4 adjacent lanes read one table row with
// ld.global.nc.L1::no_allocate.v4 (whole 32-B sectors, no L1 pollution) and the warp's stores
// to x are fully coalesced.  All NQ*S loads of a pass are issued before any is consumed.
// HBM-bound: algorithmic bytes/sample = ids (4F) + dense (4D) + rows (4Fk) (+ x write when the
// consumer is not fused).
#include "common.cuh"

namespace ctrhot {

// ------------------------------------------------------------------ plain gather (any k)
template <int IDKIND, bool VEC>
__global__ void __launch_bounds__(256)
gather_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
              const void* __restrict__ ids, int64_t ids_ld, int64_t B,
              const float* __restrict__ dense, int64_t dense_ld, int D,
              float* __restrict__ out, int64_t out_ld, int* __restrict__ err_flag) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  constexpr int U = 4;
  if (VEC) {
    // Work items = (sample, pass of 32*U chunks).  The ids of item i+1 are loaded BEFORE the row loads of item i are
    // issued, so a warp pays one DRAM round trip per item (rows), not two (ids, then rows).
    const int k4 = k >> 2;
    const int total = F * k4;
    const int span = 32 * U;
    auto load_ids = [&](int64_t b, int q0, int64_t (&id)[U]) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int q = q0 + u * 32 + lane;
        id[u] = -1;
        if (q < total && b < B) id[u] = IdLoad<IDKIND>::at(ids, b * ids_ld + q / k4);
      }
    };
    int64_t b = warp;
    int q0 = 0;
    int64_t idc[U], idn[U];
    load_ids(b, q0, idc);
    while (b < B) {
      int64_t bn = b;
      int qn = q0 + span;
      if (qn >= total) { qn = 0; bn = b + nwarps; }
      load_ids(bn, qn, idn);
      float4 buf[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int q = q0 + u * 32 + lane;
        buf[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q < total) {
          const int f = q / k4, c = q - f * k4;
          const int64_t r0 = __ldg(row_off + f), r1 = __ldg(row_off + f + 1);
          if (idc[u] >= 0 && idc[u] < r1 - r0) buf[u] = ldg_nc_f4(arena + (r0 + idc[u]) * k + c * 4);
          else if (err_flag) *err_flag = 1;
        }
      }
      if (q0 == 0)
        for (int j = lane; j < D; j += 32) out[b * out_ld + j] = __ldg(dense + b * dense_ld + j);
      float* orow = out + b * out_ld + D;
      const bool al = (reinterpret_cast<uintptr_t>(orow) & 15) == 0;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int q = q0 + u * 32 + lane;
        if (q < total) {
          if (al) {
            stg_na_f4(orow + q * 4, buf[u]);
          } else {
            orow[q * 4 + 0] = buf[u].x; orow[q * 4 + 1] = buf[u].y;
            orow[q * 4 + 2] = buf[u].z; orow[q * 4 + 3] = buf[u].w;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) idc[u] = idn[u];
      b = bn; q0 = qn;
    }
  } else {
    const int total = F * k;
    for (int64_t b = warp; b < B; b += nwarps) {
      for (int j = lane; j < D; j += 32) out[b * out_ld + j] = __ldg(dense + b * dense_ld + j);
      for (int e = lane; e < total; e += 32) {
        const int f = e / k, c = e - f * k;
        const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
        const int64_t r0 = row_off[f], r1 = row_off[f + 1];
        float v = 0.f;
        if (id >= 0 && id < r1 - r0) v = ldg_nc_f1(arena + (r0 + id) * k + c);
        else if (err_flag) *err_flag = 1;
        out[b * out_ld + D + e] = v;
      }
    }
  }
}

// ------------------------------------------------------------------ gather, F*k/4 <= 128 chunks per sample (k = 16: 104)
// The ncu capture of gather_kernel at k = 16 (profiles/r01e_ncu_summary.md) showed DRAM at 29 % and the issue slots at 54 %:
// the per-chunk index arithmetic (q / k4, 64-bit offsets, two row_off loads, range test) is what the warps spend their
// time on.  With at most 4 chunks per lane all of that is a per-LANE constant: it is hoisted out of the sample loop, which
// is left with 4 id loads (one sample ahead), 4 row loads and 4 stores per lane.
#ifndef FU_GSMINB_
#define FU_GSMINB_ 4          // resident CTAs per SM the register budget is cut for (tuning builds: profiles/build_variants.py)
#endif
template <int IDKIND>
__global__ void __launch_bounds__(256, FU_GSMINB_)
gather_small_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                    const void* __restrict__ ids, int64_t ids_ld, int64_t B,
                    const float* __restrict__ dense, int64_t dense_ld, int D,
                    float* __restrict__ out, int64_t out_ld, int* __restrict__ err_flag) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  constexpr int U = 4;
  const int k4 = k >> 2, total = F * k4;
  const float* base[U];
  uint32_t nrow[U];
  int fld[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int q = u * 32 + lane;
    const bool act = q < total;
    const int f = act ? q / k4 : 0, c = q - f * k4;
    const int64_t r0 = row_off[f], nr = row_off[f + 1] - r0;
    base[u] = arena + r0 * k + c * 4;
    nrow[u] = !act ? 0u : (nr > 0x7FFFFFFF ? 0x7FFFFFFFu : (nr < 0 ? 0u : (uint32_t)nr));   // ids are int32
    fld[u] = f;
  }
  const bool al = ((reinterpret_cast<uintptr_t>(out + D) & 15) == 0) && ((out_ld & 3) == 0);
  auto load_ids = [&](int64_t b, uint32_t (&id)[U]) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      id[u] = 0xFFFFFFFFu;
      if (b < B && u * 32 + lane < total) {
        if (IDKIND == CTRHOT_IDS_I32) id[u] = (uint32_t)__ldg(static_cast<const int32_t*>(ids) + b * ids_ld + fld[u]);
        else id[u] = (uint32_t)__float2int_rz(__ldg(static_cast<const float*>(ids) + b * ids_ld + fld[u]));
      }
    }
  };
  uint32_t idc[U], idn[U];
  int64_t b = warp;
  load_ids(b, idc);
  bool bad = false;
  while (b < B) {
    const int64_t bn = b + nwarps;
    load_ids(bn, idn);
    float4 buf[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      buf[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (idc[u] < nrow[u]) buf[u] = ldg_nc_f4(base[u] + (uint64_t)idc[u] * (uint32_t)k);
      else if (u * 32 + lane < total) bad = true;
    }
    for (int j = lane; j < D; j += 32) out[b * out_ld + j] = __ldg(dense + b * dense_ld + j);
    float* orow = out + b * out_ld + D;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int q = u * 32 + lane;
      if (q < total) {
        if (al) {
          stg_na_f4(orow + q * 4, buf[u]);
        } else {
          orow[q * 4 + 0] = buf[u].x; orow[q * 4 + 1] = buf[u].y;
          orow[q * 4 + 2] = buf[u].z; orow[q * 4 + 3] = buf[u].w;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) idc[u] = idn[u];
    b = bn;
  }
  if (bad && err_flag) *err_flag = 1;
}

// ------------------------------------------------------------------ NFM bi-interaction
// NFM.call, DL/model/nfm.py:24-30:  bi = 0.5 * ((sum_f e_f)^2 - sum_f e_f^2)  over the F gathered rows, x = [dense | bi].
// One warp per sample; lane l takes the 16-byte pieces l, l + 32, ... of the sample's F*k/4 pieces.  k/4 divides 32, so all
// pieces of a lane belong to the same 4 columns: the lane keeps sum and sum of squares of those columns in registers, lanes
// with equal l mod k/4 are combined by xor-shuffles, and lanes 0 .. k/4-1 write the k results.  The rows never reach HBM.
template <int IDKIND>
__global__ void __launch_bounds__(256)
nfm_bi_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
              const void* __restrict__ ids, int64_t ids_ld, int64_t B,
              const float* __restrict__ dense, int64_t dense_ld, int D,
              float* __restrict__ out, int64_t out_ld, int* __restrict__ err_flag) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int k4 = k >> 2, total = F * k4;
  const int sh = 31 - __clz(k4);                       // k4 = 1 << sh
  const int c = lane & (k4 - 1);
  constexpr int U = 4;
  bool bad = false;
  for (int64_t b = warp; b < B; b += nwarps) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f), ss = s;
    for (int q0 = 0; q0 < total; q0 += 32 * U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int q = q0 + 32 * u + lane;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q < total) {
          const int f = q >> sh;
          const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
          const int64_t r0 = __ldg(row_off + f), r1 = __ldg(row_off + f + 1);
          if (id >= 0 && id < r1 - r0) v[u] = ldg_nc_f4(arena + (r0 + id) * k + c * 4);
          else bad = true;
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {                    // fixed order: the sum over fields is deterministic
        s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w;
        ss.x = fmaf(v[u].x, v[u].x, ss.x); ss.y = fmaf(v[u].y, v[u].y, ss.y);
        ss.z = fmaf(v[u].z, v[u].z, ss.z); ss.w = fmaf(v[u].w, v[u].w, ss.w);
      }
    }
    for (int o = 16; o >= k4; o >>= 1) {
      s.x += __shfl_xor_sync(0xffffffffu, s.x, o); s.y += __shfl_xor_sync(0xffffffffu, s.y, o);
      s.z += __shfl_xor_sync(0xffffffffu, s.z, o); s.w += __shfl_xor_sync(0xffffffffu, s.w, o);
      ss.x += __shfl_xor_sync(0xffffffffu, ss.x, o); ss.y += __shfl_xor_sync(0xffffffffu, ss.y, o);
      ss.z += __shfl_xor_sync(0xffffffffu, ss.z, o); ss.w += __shfl_xor_sync(0xffffffffu, ss.w, o);
    }
    float* row = out + b * out_ld;
    for (int j = lane; j < D; j += 32) row[j] = __ldg(dense + b * dense_ld + j);
    if (lane < k4) {
      row[D + 4 * lane + 0] = 0.5f * (s.x * s.x - ss.x); row[D + 4 * lane + 1] = 0.5f * (s.y * s.y - ss.y);
      row[D + 4 * lane + 2] = 0.5f * (s.z * s.z - ss.z); row[D + 4 * lane + 3] = 0.5f * (s.w * s.w - ss.w);
    }
    for (int j = D + k + lane; j < out_ld; j += 32) row[j] = 0.f;         // padding of the row (the tower reads 16-byte pieces)
  }
  if (bad && err_flag) *err_flag = 1;
}

// ------------------------------------------------------------------ AFM pooling over the pairwise products
// AFMLayer.call, DL/layer/interaction.py:337-351: embed -> InteractionLayer (all e_i * e_j, i < j: [B, P, k]) -> pooling over the
// P pairs -> Dense(1) -> sigmoid; AFM.call (DL/model/afm.py) applies a second sigmoid.  None of the [B, P, k] tensor is built:
//   avg  mean over pairs                = (0.5 ((sum e)^2 - sum e^2)) / P
//   att  AttentionLayer (:300-319) takes softmax over the LAST axis of a [B, P, 1] tensor, i.e. every weight is exactly 1:
//        its output is the plain sum over pairs = 0.5 ((sum e)^2 - sum e^2), whatever its two Dense layers hold
//   max  max over pairs of x_i x_j per column = max(product of the two largest, product of the two smallest values)
// Same warp-per-sample gather as nfm_bi_kernel; the k pooled values meet out_w in registers, 4 bytes per sample leave.
template <int IDKIND>
__global__ void __launch_bounds__(256)
afm_pool_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                const void* __restrict__ ids, int64_t ids_ld, int64_t B, int mode,
                const float* __restrict__ out_w, const float* __restrict__ out_b, int n_sigmoid,
                float* __restrict__ out, int* __restrict__ err_flag) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int k4 = k >> 2, total = F * k4;
  const int sh = 31 - __clz(k4);
  const int c = lane & (k4 - 1);
  const float inv_pairs = 2.0f / ((float)F * (float)(F - 1));
  const float4 w = (lane < k4) ? __ldg(reinterpret_cast<const float4*>(out_w) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
  const float bias = __ldg(out_b);
  constexpr int U = 4;
  const float INF = __int_as_float(0x7f800000);
  bool bad = false;
  for (int64_t b = warp; b < B; b += nwarps) {
    float s[4] = {0.f, 0.f, 0.f, 0.f}, ss[4] = {0.f, 0.f, 0.f, 0.f};
    float a1[4], a2[4], b1[4], b2[4];                  // two largest / two smallest values per column (mode max)
#pragma unroll
    for (int e = 0; e < 4; ++e) { a1[e] = -INF; a2[e] = -INF; b1[e] = INF; b2[e] = INF; }
    for (int q0 = 0; q0 < total; q0 += 32 * U) {
      float4 v[U];
      bool have[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int q = q0 + 32 * u + lane;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        have[u] = q < total;
        if (have[u]) {
          const int f = q >> sh;
          const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
          const int64_t r0 = __ldg(row_off + f), r1 = __ldg(row_off + f + 1);
          if (id >= 0 && id < r1 - r0) v[u] = ldg_nc_f4(arena + (r0 + id) * k + c * 4);
          else bad = true;                              // the row counts as zeros, like the gather's
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!have[u]) continue;
        const float x[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          s[e] += x[e];
          ss[e] = fmaf(x[e], x[e], ss[e]);
          a2[e] = fmaxf(a2[e], fminf(a1[e], x[e])); a1[e] = fmaxf(a1[e], x[e]);
          b2[e] = fminf(b2[e], fmaxf(b1[e], x[e])); b1[e] = fminf(b1[e], x[e]);
        }
      }
    }
    for (int o = 16; o >= k4; o >>= 1) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        s[e] += __shfl_xor_sync(0xffffffffu, s[e], o);
        ss[e] += __shfl_xor_sync(0xffffffffu, ss[e], o);
        const float oa1 = __shfl_xor_sync(0xffffffffu, a1[e], o), oa2 = __shfl_xor_sync(0xffffffffu, a2[e], o);
        const float ob1 = __shfl_xor_sync(0xffffffffu, b1[e], o), ob2 = __shfl_xor_sync(0xffffffffu, b2[e], o);
        a2[e] = fmaxf(fminf(a1[e], oa1), fmaxf(a2[e], oa2)); a1[e] = fmaxf(a1[e], oa1);
        b2[e] = fminf(fmaxf(b1[e], ob1), fminf(b2[e], ob2)); b1[e] = fminf(b1[e], ob1);
      }
    }
    float dot = 0.f;
    if (lane < k4) {
      const float ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float pooled;
        if (mode == 2) pooled = fmaxf(a1[e] * a2[e], b1[e] * b2[e]);
        else pooled = 0.5f * (s[e] * s[e] - ss[e]) * (mode == 0 ? inv_pairs : 1.0f);
        dot = fmaf(pooled, ww[e], dot);
      }
    }
    dot = warp_sum(dot);
    if (lane == 0) {
      float r = dot + bias;
      for (int i = 0; i < n_sigmoid; ++i) r = 1.0f / (1.0f + expf(-r));
      out[b] = r;
    }
  }
  if (bad && err_flag) *err_flag = 1;
}

// ------------------------------------------------------------------ shared front-end loader
// Loads S samples' rows into registers: xv[s][i] = chunk q = lane + 32 i, xd[s] = dense[lane].
template <int IDKIND, int NQ, int S>
__device__ __forceinline__ void load_samples(
    const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k4,
    const void* __restrict__ ids, int64_t ids_ld, const float* __restrict__ dense,
    int64_t dense_ld, int D, int64_t b0, int64_t B, int lane, int* __restrict__ err_flag,
    float4 (&xv)[S][NQ], float (&xd)[S]) {
  const int total = F * k4;
  const int k = k4 * 4;
#pragma unroll
  for (int i = 0; i < NQ; ++i) {
    const int q = lane + 32 * i;
    const bool act = q < total;
    const int f = act ? q / k4 : 0;
    const int c = q - f * k4;
    const int64_t r0 = act ? row_off[f] : 0, r1 = act ? row_off[f + 1] : 0;
#pragma unroll
    for (int s = 0; s < S; ++s) {
      const int64_t b = b0 + s;
      xv[s][i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (act && b < B) {
        const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
        if (id >= 0 && id < r1 - r0) xv[s][i] = ldg_nc_f4(arena + (r0 + id) * k + c * 4);
        else if (err_flag) *err_flag = 1;
      }
    }
  }
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const int64_t b = b0 + s;
    xd[s] = (lane < D && b < B) ? __ldg(dense + b * dense_ld + lane) : 0.f;
  }
}

// x row layout written for the tower: [embeddings F*k | dense D | zero pad to x_ld]
// ("internal" column order; the tower/FM/cross weights are permuted to match at pack time).
template <int NQ, int S>
__device__ __forceinline__ void store_x(float* __restrict__ x, int64_t x_ld, int total, int D,
                                        int64_t b0, int64_t B, int lane,
                                        const float4 (&xv)[S][NQ], const float (&xd)[S]) {
  if (x == nullptr) return;
  const int emb = total * 4;
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const int64_t b = b0 + s;
    if (b >= B) continue;
    float* row = x + b * x_ld;
#pragma unroll
    for (int i = 0; i < NQ; ++i) {
      const int q = lane + 32 * i;
      if (q < total) stg_na_f4(row + q * 4, xv[s][i]);
    }
    // dense tail + zero padding (x_ld - emb - D < 32 + D is required by the host wrapper)
    // lane l < D holds dense[l] (xd is 0 for l >= D); everything past emb + 32 is padding
    for (int j = emb + lane; j < x_ld; j += 32) row[j] = (j - emb == lane) ? xd[s] : 0.f;
  }
}

// reduce-scatter over a warp: on return lane l holds the warp-wide sum of v[l] (N == 32),
// or of v[l % N] (N < 32, every lane a full sum).
template <int N>
__device__ __forceinline__ float warp_reduce_scatter(float (&v)[N], int lane) {
  static_assert(N == 32 || N == 16, "N");
  if (N == 16) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], 16);
  }
#pragma unroll
  for (int o = (N == 32 ? 16 : 8); o >= 1; o >>= 1) {
    const bool up = (lane & o) != 0;
#pragma unroll
    for (int i = 0; i < o; ++i) {
      const float send = up ? v[i] : v[i + o];
      const float keep = up ? v[i + o] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
    }
  }
  return v[0];
}

// ------------------------------------------------------------------ DeepFM front: gather -> FM (+ x)
// planes: [KFM+2][n_pad] in internal column order: plane c<KFM = v[:,c]; plane KFM = w1;
// plane KFM+1 = s_j = sum_c v[j,c]^2   (so  sum_c (x^2 @ v^2)_c = x^2 . s, DL/layer/interaction.py:110-111).
template <int IDKIND, int NQ, int S, int KFM>
__global__ void __launch_bounds__(256, (NQ * S <= 16) ? 2 : 1)
deepfm_front_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k4,
                    const void* __restrict__ ids, int64_t ids_ld, const float* __restrict__ dense,
                    int64_t dense_ld, int D, int64_t B, const float* __restrict__ planes, int n_pad,
                    const float* __restrict__ w0, float* __restrict__ x, int64_t x_ld,
                    float* __restrict__ fm_out, int* __restrict__ err_flag) {
  // The KFM+2 planes are processed in NG groups of GP <= 8 so that only S*GP accumulators are live
  // (register budget <= 128 -> two CTAs per SM); each group ends in one 32-value warp reduce-scatter
  // (value index = s*8 + c).
  constexpr int NP = KFM + 2;
  constexpr int NG = (NP + 7) / 8;
  constexpr int GP = (NP + NG - 1) / NG;
  static_assert(S * 8 <= 32, "S");
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int total = F * k4;
  const int emb = total * 4;
  const float bias0 = __ldg(w0);
  for (int64_t b0 = warp * S; b0 < B; b0 += nwarps * S) {
    float4 xv[S][NQ];
    float xd[S];
    load_samples<IDKIND, NQ, S>(arena, row_off, F, k4, ids, ids_ld, dense, dense_ld, D, b0, B, lane,
                                err_flag, xv, xd);
    store_x<NQ, S>(x, x_ld, total, D, b0, B, lane, xv, xd);
    float term = 0.f;
#pragma unroll
    for (int g = 0; g < NG; ++g) {
      float acc[S][GP];
#pragma unroll
      for (int s = 0; s < S; ++s)
#pragma unroll
        for (int c = 0; c < GP; ++c) acc[s][c] = 0.f;
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        if (q < total) {
#pragma unroll
          for (int c = 0; c < GP; ++c) {
            const int pl = g * GP + c;
            if (pl < NP) {
              const float4 w = __ldg(reinterpret_cast<const float4*>(planes + (size_t)pl * n_pad) + q);
#pragma unroll
              for (int s = 0; s < S; ++s) {
                const float4 t = xv[s][i];
                if (pl == NP - 1)
                  acc[s][c] = fmaf(t.x * t.x, w.x, fmaf(t.y * t.y, w.y, fmaf(t.z * t.z, w.z, fmaf(t.w * t.w, w.w, acc[s][c]))));
                else
                  acc[s][c] = fmaf(t.x, w.x, fmaf(t.y, w.y, fmaf(t.z, w.z, fmaf(t.w, w.w, acc[s][c]))));
              }
            }
          }
        }
      }
      if (lane < D) {
#pragma unroll
        for (int c = 0; c < GP; ++c) {
          const int pl = g * GP + c;
          if (pl < NP) {
            const float w = __ldg(planes + (size_t)pl * n_pad + emb + lane);
#pragma unroll
            for (int s = 0; s < S; ++s) acc[s][c] = fmaf(pl == NP - 1 ? xd[s] * xd[s] : xd[s], w, acc[s][c]);
          }
        }
      }
      float vals[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int s = j >> 3, c = j & 7;
        vals[j] = (s < S && c < GP && g * GP + c < NP) ? acc[s < S ? s : 0][c < GP ? c : 0] : 0.f;
      }
      const float tot = warp_reduce_scatter<32>(vals, lane);
      const int pl = g * GP + (lane & 7);
      if ((lane & 7) < GP) term += (pl < KFM) ? 0.5f * tot * tot : (pl == KFM ? tot : (pl == KFM + 1 ? -0.5f * tot : 0.f));
    }
    term += __shfl_xor_sync(0xffffffffu, term, 4);
    term += __shfl_xor_sync(0xffffffffu, term, 2);
    term += __shfl_xor_sync(0xffffffffu, term, 1);
    const int s = lane >> 3;
    if ((lane & 7) == 0 && s < S && b0 + s < B) fm_out[b0 + s] = term + bias0;
  }
}

// ------------------------------------------------------------------ DCN front: gather -> L cross layers (+ x)
// cw/cb: [L][n_pad] internal order; final_w: [n_pad] (cross half of DCN's last Dense) or NULL.
template <int IDKIND, int NQ, int S>
__global__ void __launch_bounds__(256)
dcn_front_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k4,
                 const void* __restrict__ ids, int64_t ids_ld, const float* __restrict__ dense,
                 int64_t dense_ld, int D, int64_t B, int L, const float* __restrict__ cw,
                 const float* __restrict__ cb, const float* __restrict__ final_w, int n_pad,
                 float* __restrict__ x, int64_t x_ld, float* __restrict__ cross_out, int64_t cross_ld,
                 float* __restrict__ cross_dot, int* __restrict__ err_flag) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int total = F * k4;
  const int emb = total * 4;
  for (int64_t b0 = warp * S; b0 < B; b0 += nwarps * S) {
    float4 x0[S][NQ];
    float d0[S];
    load_samples<IDKIND, NQ, S>(arena, row_off, F, k4, ids, ids_ld, dense, dense_ld, D, b0, B, lane,
                                err_flag, x0, d0);
    store_x<NQ, S>(x, x_ld, total, D, b0, B, lane, x0, d0);
    float4 xl[S][NQ];
    float dl[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
      dl[s] = d0[s];
#pragma unroll
      for (int i = 0; i < NQ; ++i) xl[s][i] = x0[s][i];
    }
    for (int l = 0; l < L; ++l) {
      const float* wl = cw + (size_t)l * n_pad;
      const float* bl = cb + (size_t)l * n_pad;
      float dot[S];
#pragma unroll
      for (int s = 0; s < S; ++s) dot[s] = 0.f;
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        if (q < total) {
          const float4 w = __ldg(reinterpret_cast<const float4*>(wl) + q);
#pragma unroll
          for (int s = 0; s < S; ++s)
            dot[s] = fmaf(xl[s][i].x, w.x, fmaf(xl[s][i].y, w.y, fmaf(xl[s][i].z, w.z, fmaf(xl[s][i].w, w.w, dot[s]))));
        }
      }
      if (lane < D) {
        const float w = __ldg(wl + emb + lane);
#pragma unroll
        for (int s = 0; s < S; ++s) dot[s] = fmaf(dl[s], w, dot[s]);
      }
#pragma unroll
      for (int s = 0; s < S; ++s) dot[s] = warp_sum(dot[s]);
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        if (q < total) {
          const float4 bb = __ldg(reinterpret_cast<const float4*>(bl) + q);
#pragma unroll
          for (int s = 0; s < S; ++s) {
            xl[s][i].x = fmaf(x0[s][i].x, dot[s], bb.x) + xl[s][i].x;
            xl[s][i].y = fmaf(x0[s][i].y, dot[s], bb.y) + xl[s][i].y;
            xl[s][i].z = fmaf(x0[s][i].z, dot[s], bb.z) + xl[s][i].z;
            xl[s][i].w = fmaf(x0[s][i].w, dot[s], bb.w) + xl[s][i].w;
          }
        }
      }
      if (lane < D) {
        const float bb = __ldg(bl + emb + lane);
#pragma unroll
        for (int s = 0; s < S; ++s) dl[s] = fmaf(d0[s], dot[s], bb) + dl[s];
      }
    }
    if (cross_out != nullptr) store_x<NQ, S>(cross_out, cross_ld, total, D, b0, B, lane, xl, dl);
    if (final_w != nullptr) {
      float dot[S];
#pragma unroll
      for (int s = 0; s < S; ++s) dot[s] = 0.f;
#pragma unroll
      for (int i = 0; i < NQ; ++i) {
        const int q = lane + 32 * i;
        if (q < total) {
          const float4 w = __ldg(reinterpret_cast<const float4*>(final_w) + q);
#pragma unroll
          for (int s = 0; s < S; ++s)
            dot[s] = fmaf(xl[s][i].x, w.x, fmaf(xl[s][i].y, w.y, fmaf(xl[s][i].z, w.z, fmaf(xl[s][i].w, w.w, dot[s]))));
        }
      }
      if (lane < D) {
        const float w = __ldg(final_w + emb + lane);
#pragma unroll
        for (int s = 0; s < S; ++s) dot[s] = fmaf(dl[s], w, dot[s]);
      }
#pragma unroll
      for (int s = 0; s < S; ++s) {
        dot[s] = warp_sum(dot[s]);
        if (lane == 0 && b0 + s < B) cross_dot[b0 + s] = dot[s];
      }
    }
  }
}

// ------------------------------------------------------------------ pack helpers (weights -> internal order)
// internal column j_int: [0, emb) <- reference column D + j_int ; [emb, emb+D) <- reference column j_int - emb
__global__ void fm_pack_planes_kernel(const float* __restrict__ w1, const float* __restrict__ v, int n,
                                      int kfm, int D, int n_pad, float* __restrict__ planes) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_pad) return;
  const int emb = n - D;
  const int jr = (j < emb) ? (D + j) : (j - emb);
  float s = 0.f;
  for (int c = 0; c < kfm; ++c) {
    const float t = (j < n) ? v[(size_t)jr * kfm + c] : 0.f;
    planes[(size_t)c * n_pad + j] = t;
    s = fmaf(t, t, s);
  }
  planes[(size_t)kfm * n_pad + j] = (j < n) ? w1[jr] : 0.f;
  planes[(size_t)(kfm + 1) * n_pad + j] = s;
}

__global__ void permute_cols_kernel(const float* __restrict__ src, int rows, int n, int D, int n_pad,
                                    float* __restrict__ dst) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (j >= n_pad || r >= rows) return;
  const int emb = n - D;
  const int jr = (j < emb) ? (D + j) : (j - emb);
  dst[(size_t)r * n_pad + j] = (j < n) ? src[(size_t)r * n + jr] : 0.f;
}

__global__ void combine_sigmoid_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                       const float* __restrict__ b_scale, const float* __restrict__ bias,
                                       float scale, int64_t B, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  float t = a[i];
  if (b) t = fmaf(b[i], b_scale ? __ldg(b_scale) : 1.0f, t);
  if (bias) t += __ldg(bias);
  out[i] = 1.0f / (1.0f + expf(-scale * t));
}

static inline int grid_for(int64_t B, int samples_per_cta) {
  int64_t g = (B + samples_per_cta - 1) / samples_per_cta;
  const int64_t cap = (int64_t)kNumSMs * 8;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int ctrhot_gather_f32(const float* arena, const int64_t* row_off, int F, int k,
                                 const void* ids, int ids_kind, int64_t ids_ld, int64_t B,
                                 const float* dense, int64_t dense_ld, int D,
                                 float* out, int64_t out_ld, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_gather_f32");
  if (!arena || !row_off || !ids || !out || F <= 0 || k <= 0 || B < 0 || D < 0) return CTRHOT_EARG;
  if (D > 0 && !dense) return CTRHOT_EARG;
  if (out_ld < (int64_t)D + (int64_t)F * k) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool vec = (k % 4 == 0) && aligned16(arena);
  const int grid = grid_for(B, 8);
#define LAUNCH(KIND, VEC)                                                                         \
  gather_kernel<KIND, VEC><<<grid, 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, dense,     \
                                                 dense_ld, D, out, out_ld, err_flag)
  if (vec && F * (k / 4) <= 128) {
    if (ids_kind == CTRHOT_IDS_I32)
      gather_small_kernel<CTRHOT_IDS_I32><<<grid, 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, dense, dense_ld, D, out,
                                                                out_ld, err_flag);
    else
      gather_small_kernel<CTRHOT_IDS_F32><<<grid, 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, dense, dense_ld, D, out,
                                                                out_ld, err_flag);
    CTRHOT_LAUNCH_OK("ctrhot_gather_f32");
    return CTRHOT_OK;
  }
  if (ids_kind == CTRHOT_IDS_I32) { if (vec) LAUNCH(CTRHOT_IDS_I32, true); else LAUNCH(CTRHOT_IDS_I32, false); }
  else                            { if (vec) LAUNCH(CTRHOT_IDS_F32, true); else LAUNCH(CTRHOT_IDS_F32, false); }
#undef LAUNCH
  CTRHOT_LAUNCH_OK("ctrhot_gather_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_nfm_bi_f32(const float* arena, const int64_t* row_off, int F, int k,
                                 const void* ids, int ids_kind, int64_t ids_ld, int64_t B,
                                 const float* dense, int64_t dense_ld, int D,
                                 float* out, int64_t out_ld, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_nfm_bi_f32");
  if (!arena || !row_off || !ids || !out || F <= 0 || k <= 0 || B < 0 || D < 0) return CTRHOT_EARG;
  if (D > 0 && !dense) return CTRHOT_EARG;
  if (out_ld < (int64_t)D + k) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (k % 4 != 0 || ((k / 4) & (k / 4 - 1)) != 0 || k > 128 || !aligned16(arena)) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(B, 8);
  if (ids_kind == CTRHOT_IDS_I32)
    nfm_bi_kernel<CTRHOT_IDS_I32><<<grid, 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, dense, dense_ld, D, out, out_ld,
                                                        err_flag);
  else
    nfm_bi_kernel<CTRHOT_IDS_F32><<<grid, 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, dense, dense_ld, D, out, out_ld,
                                                        err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_nfm_bi_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_afm_pool_f32(const float* arena, const int64_t* row_off, int F, int k,
                                   const void* ids, int ids_kind, int64_t ids_ld, int64_t B, int mode,
                                   const float* out_w, const float* out_b, int n_sigmoid, float* out, int* err_flag,
                                   void* stream) {
  CTRHOT_RANGE("ctrhot_afm_pool_f32");
  if (!arena || !row_off || !ids || !out || !out_w || !out_b || F < 2 || k <= 0 || B < 0) return CTRHOT_EARG;
  if (mode < 0 || mode > 2 || n_sigmoid < 0 || n_sigmoid > 2) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (k % 4 != 0 || ((k / 4) & (k / 4 - 1)) != 0 || k > 128 || !aligned16(arena) || !aligned16(out_w)) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(B, 8);
  if (ids_kind == CTRHOT_IDS_I32)
    afm_pool_kernel<CTRHOT_IDS_I32><<<grid, 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, mode, out_w, out_b, n_sigmoid, out,
                                                          err_flag);
  else
    afm_pool_kernel<CTRHOT_IDS_F32><<<grid, 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, mode, out_w, out_b, n_sigmoid, out,
                                                          err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_afm_pool_f32");
  return CTRHOT_OK;
}

extern "C" int64_t ctrhot_fm_planes_bytes(int n, int kfm) {
  if (n <= 0 || kfm < 0) return CTRHOT_EARG;
  const int n_pad = (n + 31) / 32 * 32;
  return (int64_t)(kfm + 2) * n_pad * sizeof(float);
}

extern "C" int ctrhot_fm_pack_planes(const float* w1, const float* v, int n, int kfm, int D,
                                     float* planes, void* stream) {
  CTRHOT_RANGE("ctrhot_fm_pack_planes");
  if (!w1 || (!v && kfm > 0) || !planes || n <= 0 || kfm < 0 || D < 0 || D > n) return CTRHOT_EARG;
  const int n_pad = (n + 31) / 32 * 32;
  fm_pack_planes_kernel<<<(n_pad + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      w1, v, n, kfm, D, n_pad, planes);
  CTRHOT_LAUNCH_OK("ctrhot_fm_pack_planes");
  return CTRHOT_OK;
}

extern "C" int ctrhot_permute_cols(const float* src, int rows, int n, int D, float* dst, void* stream) {
  CTRHOT_RANGE("ctrhot_permute_cols");
  if (!src || !dst || rows <= 0 || n <= 0 || D < 0 || D > n) return CTRHOT_EARG;
  const int n_pad = (n + 31) / 32 * 32;
  dim3 grid((n_pad + 127) / 128, rows);
  permute_cols_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(src, rows, n, D, n_pad, dst);
  CTRHOT_LAUNCH_OK("ctrhot_permute_cols");
  return CTRHOT_OK;
}

namespace {
struct FrontShape { int nq; int s; };
// NQ buckets (chunks per lane) -> samples per warp pass
inline bool pick_front(int total_q, int& nq) {
  const int need = (total_q + 31) / 32;
  const int buckets[] = {1, 2, 4, 7, 13, 26};
  for (int b : buckets) if (need <= b) { nq = b; return true; }
  return false;
}
}  // namespace

#define FRONT_COMMON_CHECKS()                                                                     \
  if (!arena || !row_off || !ids || F <= 0 || k <= 0 || B < 0 || D < 0 || D > 32) return CTRHOT_EARG; \
  if (D > 0 && !dense) return CTRHOT_EARG;                                                        \
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;               \
  if (k % 4 != 0 || !aligned16(arena)) return CTRHOT_ESHAPE;                                      \
  const int n = D + F * k;                                                                        \
  const int n_pad = (n + 31) / 32 * 32;                                                           \
  if (x && (x_ld < n || x_ld % 4 != 0 || x_ld > n_pad + 32 || !aligned16(x))) return CTRHOT_EARG; \
  int nq = 0;                                                                                     \
  if (!pick_front(F * (k / 4), nq)) return CTRHOT_ESHAPE;                                         \
  if (B == 0) return CTRHOT_OK;                                                                   \
  cudaStream_t st = static_cast<cudaStream_t>(stream);

extern "C" int ctrhot_deepfm_front_f32(const float* arena, const int64_t* row_off, int F, int k,
                                       const void* ids, int ids_kind, int64_t ids_ld,
                                       const float* dense, int64_t dense_ld, int D, int64_t B,
                                       const float* w0, const float* planes, int kfm,
                                       float* x, int64_t x_ld, float* fm_out, int* err_flag,
                                       void* stream) {
  CTRHOT_RANGE("ctrhot_deepfm_front_f32");
  if (!w0 || !planes || !fm_out) return CTRHOT_EARG;
  FRONT_COMMON_CHECKS();
  const int k4 = k / 4;
#define LAUNCH3(KIND, NQ, S, KFM)                                                                  \
  deepfm_front_kernel<KIND, NQ, S, KFM><<<grid_for(B, 8 * S), 256, 0, st>>>(                       \
      arena, row_off, F, k4, ids, ids_ld, dense, dense_ld, D, B, planes, n_pad, w0, x, x_ld,       \
      fm_out, err_flag)
#define LAUNCH2(KIND, KFM)                                                                         \
  switch (nq) {                                                                                    \
    case 1: LAUNCH3(KIND, 1, 4, KFM); break;                                                       \
    case 2: LAUNCH3(KIND, 2, 4, KFM); break;                                                       \
    case 4: LAUNCH3(KIND, 4, 4, KFM); break;                                                       \
    case 7: LAUNCH3(KIND, 7, 2, KFM); break;                                                       \
    case 13: LAUNCH3(KIND, 13, 2, KFM); break;                                                     \
    default: LAUNCH3(KIND, 26, 1, KFM); break;                                                     \
  }
#define LAUNCH1(KFM)                                                                               \
  if (ids_kind == CTRHOT_IDS_I32) { LAUNCH2(CTRHOT_IDS_I32, KFM) } else { LAUNCH2(CTRHOT_IDS_F32, KFM) }
  switch (kfm) {
    case 4: LAUNCH1(4); break;
    case 8: LAUNCH1(8); break;
    case 10: LAUNCH1(10); break;
    case 16: LAUNCH1(16); break;
    default: return CTRHOT_ESHAPE;
  }
#undef LAUNCH1
#undef LAUNCH2
#undef LAUNCH3
  CTRHOT_LAUNCH_OK("ctrhot_deepfm_front_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_dcn_front_f32(const float* arena, const int64_t* row_off, int F, int k,
                                    const void* ids, int ids_kind, int64_t ids_ld,
                                    const float* dense, int64_t dense_ld, int D, int64_t B,
                                    int L, const float* cw, const float* cb, const float* final_w,
                                    float* x, int64_t x_ld, float* cross_out, int64_t cross_ld,
                                    float* cross_dot, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_dcn_front_f32");
  if (L < 0 || (L > 0 && (!cw || !cb)) || (final_w && !cross_dot)) return CTRHOT_EARG;
  FRONT_COMMON_CHECKS();
  if (cross_out && (cross_ld < n || cross_ld % 4 != 0 || cross_ld > n_pad + 32 || !aligned16(cross_out)))
    return CTRHOT_EARG;
  const int k4 = k / 4;
#define LAUNCH3(KIND, NQ, S)                                                                       \
  dcn_front_kernel<KIND, NQ, S><<<grid_for(B, 8 * S), 256, 0, st>>>(                               \
      arena, row_off, F, k4, ids, ids_ld, dense, dense_ld, D, B, L, cw, cb, final_w, n_pad, x,     \
      x_ld, cross_out, cross_ld, cross_dot, err_flag)
#define LAUNCH2(KIND)                                                                              \
  switch (nq) {                                                                                    \
    case 1: LAUNCH3(KIND, 1, 4); break;                                                            \
    case 2: LAUNCH3(KIND, 2, 4); break;                                                            \
    case 4: LAUNCH3(KIND, 4, 2); break;                                                            \
    case 7: LAUNCH3(KIND, 7, 2); break;                                                            \
    case 13: LAUNCH3(KIND, 13, 1); break;                                                          \
    default: LAUNCH3(KIND, 26, 1); break;                                                          \
  }
  if (ids_kind == CTRHOT_IDS_I32) { LAUNCH2(CTRHOT_IDS_I32) } else { LAUNCH2(CTRHOT_IDS_F32) }
#undef LAUNCH2
#undef LAUNCH3
  CTRHOT_LAUNCH_OK("ctrhot_dcn_front_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_combine_sigmoid_f32(const float* a, const float* b_, const float* b_scale,
                                          const float* bias, float scale, int64_t B, float* out,
                                          void* stream) {
  CTRHOT_RANGE("ctrhot_combine_sigmoid_f32");
  if (!a || !out || B < 0) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  combine_sigmoid_kernel<<<(unsigned)((B + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      a, b_, b_scale, bias, scale, B, out);
  CTRHOT_LAUNCH_OK("ctrhot_combine_sigmoid_f32");
  return CTRHOT_OK;
}
