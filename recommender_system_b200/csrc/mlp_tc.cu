// K7: Dense tower on the 5th-gen tensor cores.  y = act(x @ W + b), TF32 multiply / fp32 accumulate.
//
// A layer is a persistent kernel (one CTA per SM over 128-row tiles, column tiles of BN <= 256 accumulated in two
// alternating 256-column TMEM buffers); see dense_tc_kernel below.  The chain of layers goes through HBM (fp32), a
// trailing Dense(N -> 1) is folded into the previous layer's epilogue.  Used for every tower shape the single-kernel
// towers (tower_fused.cu, fused.cu) do not take, e.g. DCN's 400-wide tower (cfg 3).
// Roofline per layer: tensor 2*M*K*N flop at the TF32 rate, HBM 4*M*(K + N) bytes.
#include "common.cuh"
#include "epilogue.cuh"
#include "mlp.cuh"
#include "tc05.cuh"

namespace ctrhot {

using namespace tc;

constexpr int TC_BM = 128;
constexpr int TC_BK = 32;                    // fp32 elements per 128-byte swizzle row
constexpr int TC_A_BYTES = TC_BM * 128;      // 16 KB

struct TcParams {
  const float* X; int64_t x_ld; int64_t M; int K; int N;
  const float* Wp; const float* bias; Epilogue ep;
  float* Y; int64_t y_ld;
  const float* dot_w; const float* dot_b; float* y_dot;   // fused Dense(N -> 1) (N fits one tile)
  int KC, BN, stages, tmem_cols, n_tiles;
  unsigned long long* trace;     // tuning only: [grid][4] cycles the MMA thread waited for {x/W stage, free accumulator}, total, steps
};

TcGeom tc_geom(int K, int N) {
  TcGeom g;
  g.KC = (K + TC_BK - 1) / TC_BK;
  g.n_tiles = (N + 255) / 256;
  const int per = (N + g.n_tiles - 1) / g.n_tiles;
  g.BN = (per + 15) / 16 * 16;
  g.stages = (g.BN > 128) ? 2 : 3;
  g.tmem_cols = 32;
  while (g.tmem_cols < g.BN) g.tmem_cols <<= 1;
  return g;
}

__global__ void mlp_pack_kernel(const float* __restrict__ W, int K, int N, int KC, int n_tiles, int BN,
                                float* __restrict__ Wp) {
  const int64_t total = (int64_t)n_tiles * KC * BN * TC_BK;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % TC_BK);
    const int nl = (int)((i / TC_BK) % BN);
    const int kc = (int)((i / ((int64_t)TC_BK * BN)) % KC);
    const int nt = (int)(i / ((int64_t)TC_BK * BN * KC));
    const int k = kc * TC_BK + kk, n = nt * BN + nl;
    const float v = (k < K && n < N) ? to_tf32_rna(W[(size_t)k * N + n]) : 0.f;
    const size_t tile = ((size_t)nt * KC + kc) * BN * TC_BK;        // floats
    Wp[tile + (sw128_offset(nl, kk >> 2) >> 2) + (kk & 3)] = v;
  }
}

// ---------------------------------------------------------------- persistent layer kernel
// One CTA per SM walks 128-row tiles; per row tile it runs the layer's n_tiles column tiles back to back (the second
// pass over the x rows hits L2), each into one of TWO TMEM accumulator buffers so that the epilogue of one step runs
// under the MMAs of the next.  14 warps:
//   warps 0-3   epilogue: tcgen05.ld of their own 32 TMEM lanes -> bias + activation -> 128 B per row and chunk to HBM,
//               or the fused trailing Dense(N -> 1) dot (accumulated in registers across the column tiles).
//   warps 4-11  A producers: coalesced 128-bit loads of x (4 rows per warp instruction) held in registers TC_PF chunks
//               ahead of their use -> round-to-nearest TF32 -> st.shared in the 128-byte-swizzled K-major layout.
//   warp 12     lane 0 issues tcgen05.mma.kind::tf32 (M=128, N=BN, K=8), four per 32-column chunk.
//   warp 13     lane 0 streams the pre-swizzled W chunk (BN x 128 B) with one cp.async.bulk per stage.
constexpr int TC_THREADS = 14 * 32;
constexpr int TC_PF = 4;                     // chunks of x in registers ahead of the smem ring
constexpr int TC_ACC_COLS = 256;             // columns per accumulator buffer

constexpr int TC_MAXN = 1024;                // widest layer the persistent kernel stages vectors for (n_tiles * BN)

// bias + activation (+ the trailing-dot partial sum) on W columns held by one thread; padding columns come out as 0
template <int W>
__device__ __forceinline__ void tc_epi_chunk(float (&acc)[W], const float* sb, const float* sa, const float* sd, float& dot) {
  float d0 = 0.f, d1 = 0.f;
#pragma unroll
  for (int j = 0; j < W; j += 4) {
    const float4 b = *reinterpret_cast<const float4*>(sb + j);
    const float4 a = *reinterpret_cast<const float4*>(sa + j);
    const float4 w = *reinterpret_cast<const float4*>(sd + j);
    float t;
    t = acc[j + 0] + b.x; t = fmaxf(t, 0.f) + a.x * fminf(t, 0.f); acc[j + 0] = t; d0 = fmaf(t, w.x, d0);
    t = acc[j + 1] + b.y; t = fmaxf(t, 0.f) + a.y * fminf(t, 0.f); acc[j + 1] = t; d1 = fmaf(t, w.y, d1);
    t = acc[j + 2] + b.z; t = fmaxf(t, 0.f) + a.z * fminf(t, 0.f); acc[j + 2] = t; d0 = fmaf(t, w.z, d0);
    t = acc[j + 3] + b.w; t = fmaxf(t, 0.f) + a.w * fminf(t, 0.f); acc[j + 3] = t; d1 = fmaf(t, w.w, d1);
  }
  dot += d0 + d1;
}

// W consecutive columns [n0, n0 + W) of one output row
template <int W>
__device__ __forceinline__ void tc_store_chunk(const float (&acc)[W], float* row, int n0, int N, int64_t y_ld, bool vec) {
  float* dst = row + n0;
  if (vec && n0 + W <= y_ld) {
#pragma unroll
    for (int j = 0; j < W; j += 4)
      stg_na_f4(dst + j, make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]));   // asm: no alignment inference
  } else {
#pragma unroll
    for (int j = 0; j < W; ++j) if (n0 + j < N) dst[j] = acc[j];
  }
}

struct TcCursor {
  int lt, nt, kc;
  __device__ __forceinline__ void advance(int KC, int n_tiles) {
    if (++kc == KC) { kc = 0; if (++nt == n_tiles) { nt = 0; ++lt; } }
  }
};

template <bool VEC>
__global__ void __launch_bounds__(TC_THREADS, 1)
dense_tc_kernel(const TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw_addr);
  const uint32_t b_bytes = (uint32_t)p.BN * 128u;
  const uint32_t stage_bytes = TC_A_BYTES + b_bytes;
  const uint32_t bars = base + p.stages * stage_bytes;          // full[S], empty[S], acc_full[2], acc_free[2], tmem slot
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (p.stages + s); };
  auto acc_full = [&](int b) { return bars + 8u * (2 * p.stages + b); };
  auto acc_free = [&](int b) { return bars + 8u * (2 * p.stages + 2 + b); };
  const uint32_t slot_addr = bars + 8u * (2 * p.stages + 4);
  volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + (slot_addr - base));
  // per-column vectors of the layer, zero-padded to n_tiles * BN: bias | slope of the negative side | trailing-dot weights
  float* s_vec = reinterpret_cast<float*>(base_ptr + (bars + 256u - base));
  const int NP = p.n_tiles * p.BN;
  float* s_bias = s_vec;
  float* s_alpha = s_vec + TC_MAXN;
  float* s_dot = s_vec + 2 * TC_MAXN;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // none / relu / PReLU with one alpha row: act(t) = max(t, 0) + a * min(t, 0) with a = 1 / 0 / alpha[n]
  const bool fast_act = p.ep.act == CTRHOT_ACT_NONE || p.ep.act == CTRHOT_ACT_RELU ||
                        (p.ep.act == CTRHOT_ACT_PRELU && p.ep.alpha_rows == 1);
  for (int i = tid; i < NP; i += TC_THREADS) {
    s_bias[i] = (i < p.N && p.bias) ? __ldg(p.bias + i) : 0.f;
    float a = 0.f;
    if (i < p.N) a = (p.ep.act == CTRHOT_ACT_NONE) ? 1.f : (p.ep.act == CTRHOT_ACT_PRELU && fast_act ? __ldg(p.ep.p + i) : 0.f);
    s_alpha[i] = a;
    s_dot[i] = (i < p.N && p.dot_w) ? __ldg(p.dot_w + i) : 0.f;
  }
  if (tid == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar(s), 8 + 1); mbar_init(empty_bar(s), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(acc_full(b), 1); mbar_init(acc_free(b), 4); }
    mbar_fence_init();
  }
  if (warp == 12) { tmem_alloc(slot_addr, 512u); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot_ptr;

  const int n_row_tiles = (int)((p.M + TC_BM - 1) / TC_BM);
  const int n_my = ((int)blockIdx.x < n_row_tiles) ? (n_row_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int KC = p.KC, NT = p.n_tiles;
  const int total_chunks = n_my * NT * KC;

  if (warp >= 4 && warp < 12) {
    // ------------------------------------------------ A producers (256 threads: piece c of rows r0 + 32 i)
    const int t = tid - 128;
    const int c = t & 7;
    const int r0 = t >> 3;                      // 0..31
    TcCursor cf{0, 0, 0}, cp{0, 0, 0};
    int n_fetched = 0;
    auto fetch = [&](float4 (&v)[4]) {
      if (n_fetched >= total_chunks) return;
      ++n_fetched;
      const int64_t m0 = ((int64_t)blockIdx.x + (int64_t)cf.lt * gridDim.x) * TC_BM;
      const int col = cf.kc * TC_BK + c * 4;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t gm = m0 + r0 + 32 * i;
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gm < p.M) {
          const float* src = p.X + gm * p.x_ld + col;
          if (VEC && col + 3 < p.K) {
            v[i] = ldg_nc_f4(src);
          } else {
            if (col + 0 < p.K) v[i].x = ldg_nc_f1(src + 0);
            if (col + 1 < p.K) v[i].y = ldg_nc_f1(src + 1);
            if (col + 2 < p.K) v[i].z = ldg_nc_f1(src + 2);
            if (col + 3 < p.K) v[i].w = ldg_nc_f1(src + 3);
          }
        }
      }
      cf.advance(KC, NT);
    };
    auto put = [&](int g, const float4 (&v)[4]) {
      const int s = g % p.stages;
      mbar_wait(empty_bar(s), (((uint32_t)(g / p.stages)) & 1u) ^ 1u);
      uint8_t* a_tile = base_ptr + (size_t)s * stage_bytes;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float4 q = v[i];
        q.x = to_tf32_rna(q.x); q.y = to_tf32_rna(q.y); q.z = to_tf32_rna(q.z); q.w = to_tf32_rna(q.w);
        *reinterpret_cast<float4*>(a_tile + sw128_offset((uint32_t)(r0 + 32 * i), (uint32_t)c)) = q;
      }
      // No proxy fence here: a thread with loads in flight would sit in it until they land (the fence drains the
      // thread's outstanding memory operations), serialising the prefetch.  The MMA thread fences after its wait on
      // full_bar instead -- the fence only has to lie on the causality path between these stores and the MMA's reads.
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar(s));
    };
    static_assert(TC_PF == 4, "the loop below is unrolled for four register sets");
    float4 va[4], vb[4], vc[4], vd[4];
    fetch(va); fetch(vb); fetch(vc); fetch(vd);
    for (int g = 0; g < total_chunks; g += TC_PF) {
      put(g, va); fetch(va);
      if (g + 1 < total_chunks) { put(g + 1, vb); fetch(vb); }
      if (g + 2 < total_chunks) { put(g + 2, vc); fetch(vc); }
      if (g + 3 < total_chunks) { put(g + 3, vd); fetch(vd); }
    }
    (void)cp;
  } else if (warp == 12) {
    // ------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(TC_BM, (uint32_t)p.BN);
      int g = 0;
      uint32_t step = 0;
      long long tw_full = 0, tw_acc = 0;
      const long long t_begin = p.trace ? clock64() : 0;
      for (int lt = 0; lt < n_my; ++lt) {
        for (int nt = 0; nt < NT; ++nt, ++step) {
          const uint32_t buf = step & 1u;
          if (p.trace) {
            const long long t0 = clock64();
            mbar_wait(acc_free(buf), ((step >> 1) & 1u) ^ 1u);
            tw_acc += clock64() - t0;
          }
          mbar_wait(acc_free(buf), ((step >> 1) & 1u) ^ 1u);          // the epilogue has drained this buffer
          tc_fence_after();
          const uint32_t d_t = tmem_base + buf * TC_ACC_COLS;
          for (int kc = 0; kc < KC; ++kc, ++g) {
            const int s = g % p.stages;
            if (p.trace) {
              const long long t0 = clock64();
              mbar_wait(full_bar(s), ((uint32_t)(g / p.stages)) & 1u);
              tw_full += clock64() - t0;
            }
            mbar_wait(full_bar(s), ((uint32_t)(g / p.stages)) & 1u);
            fence_proxy_async_smem();             // generic-proxy stores of the A producers -> async-proxy reads of the MMA
            tc_fence_after();
            const uint32_t a_addr = base + s * stage_bytes;
            const uint64_t da = make_desc_sw128(a_addr);
            const uint64_t db = make_desc_sw128(a_addr + TC_A_BYTES);
#pragma unroll
            for (int k = 0; k < 4; ++k)       // 4 x (K=8 tf32 = 32 bytes) per 128-byte swizzle row
              mma_tf32_ss(d_t, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kc | k) != 0));
            mma_commit(empty_bar(s));
          }
          mma_commit(acc_full(buf));
        }
      }
      if (p.trace) {
        unsigned long long* t = p.trace + 4 * blockIdx.x;
        t[0] = (unsigned long long)tw_full; t[1] = (unsigned long long)tw_acc;
        t[2] = (unsigned long long)(clock64() - t_begin); t[3] = step;
      }
    }
    __syncwarp();
  } else if (warp == 13) {
    // ------------------------------------------------ W producer (bulk async copy)
    if (lane == 0) {
      int g = 0;
      for (int lt = 0; lt < n_my; ++lt) {
        for (int nt = 0; nt < NT; ++nt) {
          const float* wsrc = p.Wp + (size_t)nt * KC * p.BN * TC_BK;
          for (int kc = 0; kc < KC; ++kc, ++g) {
            const int s = g % p.stages;
            mbar_wait(empty_bar(s), (((uint32_t)(g / p.stages)) & 1u) ^ 1u);
            mbar_arrive_expect_tx(full_bar(s), b_bytes);
            bulk_g2s(base + s * stage_bytes + TC_A_BYTES, wsrc + (size_t)kc * p.BN * TC_BK, b_bytes, full_bar(s));
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------ epilogue (warps 0-3 = TMEM lane quarters 0-3)
    const int row = warp * 32 + lane;
    const bool y_vec = p.Y && ((p.y_ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.Y) & 15) == 0);
    uint32_t step = 0;
    for (int lt = 0; lt < n_my; ++lt) {
      const int64_t gm = ((int64_t)blockIdx.x + (int64_t)lt * gridDim.x) * TC_BM + row;
      float dot = 0.f;
      for (int nt = 0; nt < NT; ++nt, ++step) {
        const uint32_t buf = step & 1u;
        mbar_wait(acc_full(buf), (step >> 1) & 1u);
        tc_fence_after();
        const uint32_t t_lane = tmem_base + ((uint32_t)(warp * 32) << 16) + buf * TC_ACC_COLS;
        if (fast_act) {
          for (int c0 = 0; c0 < p.BN; c0 += 32) {
            const int n0 = nt * p.BN + c0;
            if (c0 + 32 <= p.BN) {
              float acc[32];
              tmem_ld32(t_lane + (uint32_t)c0, acc);
              tc_epi_chunk<32>(acc, s_bias + n0, s_alpha + n0, s_dot + n0, dot);
              if (gm < p.M && p.Y) tc_store_chunk<32>(acc, p.Y + gm * p.y_ld, n0, p.N, p.y_ld, y_vec);
            } else {
              float acc[16];
              tmem_ld16(t_lane + (uint32_t)c0, acc);
              tc_epi_chunk<16>(acc, s_bias + n0, s_alpha + n0, s_dot + n0, dot);
              if (gm < p.M && p.Y) tc_store_chunk<16>(acc, p.Y + gm * p.y_ld, n0, p.N, p.y_ld, y_vec);
            }
          }
        } else {
          for (int c0 = 0; c0 < p.BN; c0 += 16) {
            float acc[16];
            tmem_ld16(t_lane + (uint32_t)c0, acc);
            const int n0 = nt * p.BN + c0;
            if (gm < p.M) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const int n = n0 + j;
                float v = 0.f;
                if (n < p.N) {
                  v = p.ep.apply(acc[j] + s_bias[n], gm, n, p.N);
                  dot = fmaf(v, s_dot[n], dot);
                }
                acc[j] = v;
              }
              if (p.Y) tc_store_chunk<16>(acc, p.Y + gm * p.y_ld, n0, p.N, p.y_ld, y_vec);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_free(buf));
      }
      if (p.dot_w && gm < p.M) p.y_dot[gm] = dot + (p.dot_b ? __ldg(p.dot_b) : 0.f);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 12) { tc_fence_after(); tmem_dealloc(tmem_base, 512u); }
}

static unsigned long long* g_tc_trace = nullptr;   // tuning only (ctrhot_debug_tc_trace)

int launch_dense_tc(const float* X, int64_t x_ld, int64_t M, const ctrhot_dense_desc& d, float* Y, int64_t y_ld,
                    const float* dot_w, const float* dot_b, float* y_dot, cudaStream_t st) {
  const TcGeom g = tc_geom(d.in_dim, d.out_dim);
  TcParams p;
  p.X = X; p.x_ld = x_ld; p.M = M; p.K = d.in_dim; p.N = d.out_dim;
  p.Wp = static_cast<const float*>(d.packed); p.bias = d.bias; p.ep = Epilogue::from(d);
  p.Y = Y; p.y_ld = y_ld; p.dot_w = dot_w; p.dot_b = dot_b; p.y_dot = y_dot;
  p.KC = g.KC; p.BN = g.BN; p.n_tiles = g.n_tiles; p.tmem_cols = 512; p.trace = g_tc_trace;
  const size_t stage = (size_t)TC_A_BYTES + (size_t)g.BN * 128;
  if (g.n_tiles * g.BN > TC_MAXN) return CTRHOT_ESHAPE;
  const size_t tail = 1024 + 256 + 3 * TC_MAXN * sizeof(float);      // alignment slack, barriers, per-column vectors
  int stages = (int)((227 * 1024 - tail) / stage);
  if (stages > 6) stages = 6;
  if (stages < 2) return CTRHOT_ESHAPE;
  p.stages = stages;
  const size_t smem = (size_t)stages * stage + tail;
  const bool vec = (x_ld % 4 == 0) && aligned16(X);
  auto kern = vec ? dense_tc_kernel<true> : dense_tc_kernel<false>;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "mlp_tc attr");
  const int64_t row_tiles = (M + TC_BM - 1) / TC_BM;
  const int grid = (int)(row_tiles < kNumSMs ? row_tiles : kNumSMs);
  kern<<<grid, TC_THREADS, smem, st>>>(p);
  CTRHOT_LAUNCH_OK("ctrhot_mlp_tf32");
  return CTRHOT_OK;
}

}  // namespace ctrhot

using namespace ctrhot;

// tuning only: device buffer of kNumSMs * 4 counters filled by the next per-layer launches (NULL switches it off)
extern "C" int ctrhot_debug_tc_trace(uint64_t* d_buf) {
  CTRHOT_RANGE("ctrhot_debug_tc_trace");
  ctrhot::g_tc_trace = reinterpret_cast<unsigned long long*>(d_buf);
  return CTRHOT_OK;
}

extern "C" int64_t ctrhot_mlp_pack_bytes(int in_dim, int out_dim) {
  if (in_dim <= 0 || out_dim <= 0) return CTRHOT_EARG;
  const TcGeom g = tc_geom(in_dim, out_dim);
  return (int64_t)g.n_tiles * g.KC * g.BN * TC_BK * sizeof(float);
}

extern "C" int ctrhot_mlp_pack(const float* weight, int in_dim, int out_dim, void* packed, void* stream) {
  CTRHOT_RANGE("ctrhot_mlp_pack");
  if (!weight || !packed || in_dim <= 0 || out_dim <= 0) return CTRHOT_EARG;
  if (!aligned16(packed)) return CTRHOT_EARG;
  const TcGeom g = tc_geom(in_dim, out_dim);
  const int64_t total = (int64_t)g.n_tiles * g.KC * g.BN * TC_BK;
  int64_t grid = (total + 255) / 256;
  if (grid > kNumSMs * 16) grid = kNumSMs * 16;
  mlp_pack_kernel<<<(int)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(weight, in_dim, out_dim, g.KC, g.n_tiles,
                                                                           g.BN, static_cast<float*>(packed));
  CTRHOT_LAUNCH_OK("ctrhot_mlp_pack");
  return CTRHOT_OK;
}

extern "C" int ctrhot_mlp_tf32(const float* x, int64_t x_ld, int64_t M, const ctrhot_dense_desc* h_layers,
                               int n_layers, float* out, int64_t out_ld, void* workspace,
                               int64_t workspace_bytes, void* stream) {
  CTRHOT_RANGE("ctrhot_mlp_tf32");
  if (!x || !h_layers || n_layers <= 0 || !out || M < 0) return CTRHOT_EARG;
  if (x_ld < h_layers[0].in_dim || out_ld < h_layers[n_layers - 1].out_dim) return CTRHOT_EARG;
  for (int l = 0; l < n_layers; ++l) {
    const ctrhot_dense_desc& d = h_layers[l];
    if (!d.weight || d.in_dim <= 0 || d.out_dim <= 0) return CTRHOT_EARG;
    if (l > 0 && d.in_dim != h_layers[l - 1].out_dim) return CTRHOT_EARG;
    if ((d.act == CTRHOT_ACT_PRELU || d.act == CTRHOT_ACT_DICE) && !d.act_param) return CTRHOT_EARG;
    if (d.act < CTRHOT_ACT_NONE || d.act > CTRHOT_ACT_TANH) return CTRHOT_EARG;
    if (d.out_dim >= 8 && (!d.packed || !aligned16(d.packed))) return CTRHOT_EARG;   // tensor path needs ctrhot_mlp_pack
  }
  if (M == 0) return CTRHOT_OK;
  int64_t mx = 0;
  for (int l = 0; l + 1 < n_layers; ++l) {
    int64_t w = (h_layers[l].out_dim + 31) / 32 * 32;
    if (w > mx) mx = w;
  }
  if (n_layers > 1 && (!workspace || workspace_bytes < 2 * M * mx * (int64_t)sizeof(float))) return CTRHOT_EARG;
  float* buf[2] = {static_cast<float*>(workspace), static_cast<float*>(workspace) + M * mx};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float* cur = x;
  int64_t cur_ld = x_ld;
  for (int l = 0; l < n_layers; ++l) {
    const ctrhot_dense_desc& d = h_layers[l];
    const bool last = (l == n_layers - 1);
    // fuse a trailing Dense(h -> 1, no activation) into this layer's epilogue
    const bool fuse_next = (l == n_layers - 2) && h_layers[l + 1].out_dim == 1 &&
                           h_layers[l + 1].act == CTRHOT_ACT_NONE && d.out_dim >= 8 && out_ld == 1;
    float* y = last ? out : buf[l & 1];
    const int64_t y_ld = last ? out_ld : mx;
    int rc;
    if (d.out_dim < 8) {
      rc = launch_dense_f32(cur, cur_ld, M, d, y, y_ld, st);      // N too narrow for an MMA tile: CUDA-core dot
    } else if (fuse_next) {
      rc = launch_dense_tc(cur, cur_ld, M, d, nullptr, 0, h_layers[l + 1].weight, h_layers[l + 1].bias, out, st);
      if (rc != CTRHOT_OK) return rc;
      return CTRHOT_OK;
    } else {
      rc = launch_dense_tc(cur, cur_ld, M, d, y, y_ld, nullptr, nullptr, nullptr, st);
    }
    if (rc != CTRHOT_OK) return rc;
    cur = y;
    cur_ld = y_ld;
  }
  return CTRHOT_OK;
}
