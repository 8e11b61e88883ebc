// K7: Dense tower on the 5th-gen tensor cores.  y = act(x @ W + b), TF32 multiply / fp32 accumulate.
//
// One CTA = one 128-row x BN-column output tile (BN <= 256, fp32 accumulator = BN TMEM columns).
//   warps 0-3  A producers: 128-bit global loads of the x tile -> round-to-nearest TF32 -> st.shared in the
//              128-byte-swizzled K-major layout the UMMA descriptor expects; afterwards the epilogue
//              (tcgen05.ld of their own 32 TMEM lanes -> bias + activation -> store, or the fused
//              final Dense(h->1) dot).
//   warp 4     owns TMEM (tcgen05.alloc/dealloc); lane 0 issues tcgen05.mma.kind::tf32 (M=128,N=BN,K=8)
//              four per 32-column K chunk and tcgen05.commit's the stage back to the producers.
//   warp 5     lane 0 streams the pre-packed, pre-swizzled W chunk (BN x 128 B, one contiguous
//              block) with a single cp.async.bulk per stage (TMA engine, mbarrier complete_tx).
// Two CTAs are resident per SM (<= 100 KB smem, <= 256 TMEM columns each) so one CTA's epilogue
// overlaps the other's main loop.  Tensor-bound roofline: 2*M*K*N flop per layer.
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
  int KC, BN, stages, tmem_cols;
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

template <bool VEC>
__global__ void __launch_bounds__(192, 2)
dense_tc_kernel(const TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw_addr);
  const uint32_t b_bytes = (uint32_t)p.BN * 128u;
  const uint32_t stage_bytes = TC_A_BYTES + b_bytes;
  const uint32_t bars = base + p.stages * stage_bytes;          // full[S], empty[S], accum, tmem slot
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (p.stages + s); };
  const uint32_t accum_bar = bars + 8u * (2 * p.stages);
  const uint32_t slot_addr = accum_bar + 8u;
  volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + (slot_addr - base));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar(s), 4 + 1); mbar_init(empty_bar(s), 1); }
    mbar_init(accum_bar, 1);
    mbar_fence_init();
  }
  if (warp == 4) { tmem_alloc(slot_addr, (uint32_t)p.tmem_cols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot_ptr;

  const int64_t m0 = (int64_t)blockIdx.y * TC_BM;
  const int nt = blockIdx.x;

  if (warp < 4) {
    // ------------------------------------------------ A producer
    const int c = tid & 7;                       // 16-byte chunk within the 128-byte row
    const int r0 = tid >> 3;                     // 0..15
    for (int kc = 0; kc < p.KC; ++kc) {
      const int s = kc % p.stages;
      const uint32_t ph = (uint32_t)(kc / p.stages) & 1u;
      float4 v[8];
      const int col = kc * TC_BK + c * 4;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int64_t gm = m0 + r0 + 16 * i;
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
      mbar_wait(empty_bar(s), ph ^ 1u);
      uint8_t* a_tile = base_ptr + (size_t)s * stage_bytes;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint32_t row = (uint32_t)(r0 + 16 * i);
        float4 t = v[i];
        t.x = to_tf32_rna(t.x); t.y = to_tf32_rna(t.y); t.z = to_tf32_rna(t.z); t.w = to_tf32_rna(t.w);
        *reinterpret_cast<float4*>(a_tile + sw128_offset(row, (uint32_t)c)) = t;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar(s));
    }
    // ------------------------------------------------ epilogue
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;
    const int64_t gm = m0 + row;
    const uint32_t t_lane = tmem_base + ((uint32_t)(warp * 32) << 16);
    float dot = 0.f;
    for (int c0 = 0; c0 < p.BN; c0 += 16) {
      float acc[16];
      tmem_ld16(t_lane + (uint32_t)c0, acc);
      const int n0 = nt * p.BN + c0;
      if (gm < p.M) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int n = n0 + j;
          float t = 0.f;
          if (n < p.N) {
            t = acc[j] + (p.bias ? __ldg(p.bias + n) : 0.f);
            t = p.ep.apply(t, gm, n, p.N);
            if (p.dot_w) dot = fmaf(t, __ldg(p.dot_w + n), dot);
          }
          acc[j] = t;
        }
        if (p.Y) {
          float* dst = p.Y + gm * p.y_ld + n0;
          if (((p.y_ld & 3) == 0) && n0 + 16 <= p.y_ld && ((reinterpret_cast<uintptr_t>(p.Y) & 15) == 0)) {
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              if (n0 + j < p.y_ld) stg_na_f4(dst + j, make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]));   // asm: no alignment inference
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) if (n0 + j < p.N) dst[j] = acc[j];
          }
        }
      }
    }
    if (p.dot_w && gm < p.M) p.y_dot[gm] = dot + (p.dot_b ? __ldg(p.dot_b) : 0.f);
    tc_fence_before();
  } else if (warp == 4) {
    // ------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(TC_BM, (uint32_t)p.BN);
      for (int kc = 0; kc < p.KC; ++kc) {
        const int s = kc % p.stages;
        const uint32_t ph = (uint32_t)(kc / p.stages) & 1u;
        mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t a_addr = base + s * stage_bytes;
        const uint64_t da = make_desc_sw128(a_addr);
        const uint64_t db = make_desc_sw128(a_addr + TC_A_BYTES);
#pragma unroll
        for (int k = 0; k < 4; ++k)       // 4 x (K=8 tf32 = 32 bytes) per 128-byte swizzle row
          mma_tf32_ss(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kc | k) != 0));
        mma_commit(empty_bar(s));
      }
      mma_commit(accum_bar);
    }
    __syncwarp();
  } else {
    // ------------------------------------------------ W producer (bulk async copy)
    if (lane == 0) {
      const float* wsrc = p.Wp + (size_t)nt * p.KC * p.BN * TC_BK;
      for (int kc = 0; kc < p.KC; ++kc) {
        const int s = kc % p.stages;
        const uint32_t ph = (uint32_t)(kc / p.stages) & 1u;
        mbar_wait(empty_bar(s), ph ^ 1u);
        mbar_arrive_expect_tx(full_bar(s), b_bytes);
        bulk_g2s(base + s * stage_bytes + TC_A_BYTES, wsrc + (size_t)kc * p.BN * TC_BK, b_bytes, full_bar(s));
      }
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == 4) { tc_fence_after(); tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols); }
}

int launch_dense_tc(const float* X, int64_t x_ld, int64_t M, const ctrhot_dense_desc& d, float* Y, int64_t y_ld,
                    const float* dot_w, const float* dot_b, float* y_dot, cudaStream_t st) {
  const TcGeom g = tc_geom(d.in_dim, d.out_dim);
  TcParams p;
  p.X = X; p.x_ld = x_ld; p.M = M; p.K = d.in_dim; p.N = d.out_dim;
  p.Wp = static_cast<const float*>(d.packed); p.bias = d.bias; p.ep = Epilogue::from(d);
  p.Y = Y; p.y_ld = y_ld; p.dot_w = dot_w; p.dot_b = dot_b; p.y_dot = y_dot;
  p.KC = g.KC; p.BN = g.BN; p.stages = g.stages; p.tmem_cols = g.tmem_cols;
  const size_t smem = (size_t)g.stages * (TC_A_BYTES + g.BN * 128) + 1024 + 128;
  const bool vec = (x_ld % 4 == 0) && aligned16(X);
  auto kern = vec ? dense_tc_kernel<true> : dense_tc_kernel<false>;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "mlp_tc attr");
  dim3 grid(g.n_tiles, (unsigned)((M + TC_BM - 1) / TC_BM));
  kern<<<grid, 192, smem, st>>>(p);
  CTRHOT_LAUNCH_OK("ctrhot_mlp_tf32");
  return CTRHOT_OK;
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int64_t ctrhot_mlp_pack_bytes(int in_dim, int out_dim) {
  if (in_dim <= 0 || out_dim <= 0) return CTRHOT_EARG;
  const TcGeom g = tc_geom(in_dim, out_dim);
  return (int64_t)g.n_tiles * g.KC * g.BN * TC_BK * sizeof(float);
}

extern "C" int ctrhot_mlp_pack(const float* weight, int in_dim, int out_dim, void* packed, void* stream) {
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
                           h_layers[l + 1].act == CTRHOT_ACT_NONE && d.out_dim >= 8 &&
                           tc_geom(d.in_dim, d.out_dim).n_tiles == 1 && out_ld == 1;
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
