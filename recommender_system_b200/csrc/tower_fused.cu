// K7f: the whole Dense tower in ONE persistent kernel, optionally fed straight from the embedding
// tables (the gathered rows never reach HBM as x).   tcgen05 / TMEM / TF32.
//
//   out[m] = post( w_out . act_L(... act_1(x_m W_1 + b_1) ...) + b_out )            (DNNLayer, interaction.py:40-46)
//
// One CTA per SM, looping over 128-row tiles.  Warp roles (320 threads):
//   warps 0-3  A producers.  Row source is either x[M, x_ld] or (ids, tables, dense): every 16-byte piece of
//              a tile row is moved global -> shared with cp.async.cg (LDGSTS, no register staging) directly
//              into the 128B-swizzled K-major stage the UMMA descriptor reads; three stages stay in flight per
//              thread (commit_group / wait_group), then fence.proxy.async + mbarrier arrive.
//   warps 4-7  epilogue: tcgen05.ld the fp32 accumulator (their own 32 TMEM lanes) -> bias + activation ->
//              round to TF32 -> tcgen05.st back into TMEM as the NEXT layer's A operand.  Hidden activations
//              never touch shared or global memory.  The last hidden layer's epilogue does the final
//              Dense(h -> 1) dot, the optional "+ fm" / sigmoid, and writes 4 bytes per row.
//   warp 8     TMEM owner + MMA issuer (one lane): layer 1 = SS (A,B from smem), layers >= 2 = TS (A from TMEM).
//   warp 9     W producer: one cp.async.bulk (TMA engine) per 32-column K chunk of the pre-swizzled weights.
// TMEM (512 columns): D (accumulator) at column 0, H (next A operand) at column 256.
#include "common.cuh"
#include "epilogue.cuh"
#include "mlp.cuh"
#include "tc05.cuh"

namespace ctrhot {
using namespace tc;

constexpr int FT_BM = 128;
constexpr int FT_NA = 4;                 // A ring stages (16 KB each)
constexpr int FT_NB = 3;                 // W ring stages (<= 32 KB each)
constexpr int FT_PA = 3;                 // cp.async groups in flight per producer thread
constexpr int FT_A_BYTES = FT_BM * 128;
constexpr int FT_MAXL = 3;
constexpr int FT_H_COL = 256;
constexpr int FT_THREADS = 320;

struct FtLayer {
  const float* Wp; const float* bias; Epilogue ep;
  int K, N, KC, BN;                      // BN = N rounded up to 16 (MMA N), KC = K chunks of 32
};

struct FtParams {
  // row source
  const float* x; int64_t x_ld;
  const float* arena; const int64_t* row_off; int F; int k4; const void* ids; int64_t ids_ld;
  const float* dense; int64_t dense_ld; int D; int emb;          // emb = F*k (columns before the dense tail)
  int64_t M; int n_tiles;
  FtLayer L[FT_MAXL]; int nl;
  const float* w_out; const float* b_out;                        // final Dense(N_last -> 1)
  const float* y_scale; const float* add_vec; const float* add_bias; float scale; int apply_sigmoid;
  float* out; int* err_flag;
  uint32_t b_stage_bytes;
};

template <int IDKIND, bool GATHER>
__global__ void __launch_bounds__(FT_THREADS, 1)
tower_fused_kernel(const FtParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw_addr);
  const uint32_t a_base = base;
  const uint32_t b_base = a_base + FT_NA * FT_A_BYTES;
  const uint32_t ids_base = b_base + FT_NB * p.b_stage_bytes;            // 2 x [128][F] int32 (gather mode)
  const uint32_t ids_bytes = GATHER ? (uint32_t)(FT_BM * p.F * 4 + 15) & ~15u : 0u;
  const uint32_t bars = ids_base + 2 * ids_bytes;
  auto full_a = [&](int s) { return bars + 8u * s; };
  auto empty_a = [&](int s) { return bars + 8u * (FT_NA + s); };
  auto full_b = [&](int s) { return bars + 8u * (2 * FT_NA + s); };
  auto empty_b = [&](int s) { return bars + 8u * (2 * FT_NA + FT_NB + s); };
  const uint32_t d_full = bars + 8u * (2 * FT_NA + 2 * FT_NB);
  const uint32_t h_ready = d_full + 8u;
  const uint32_t d_free = d_full + 16u;
  const uint32_t slot_addr = d_full + 24u;
  volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + (slot_addr - base));
  int32_t* s_ids = reinterpret_cast<int32_t*>(base_ptr + (ids_base - base));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < FT_NA; ++s) { mbar_init(full_a(s), 128); mbar_init(empty_a(s), 1); }
    for (int s = 0; s < FT_NB; ++s) { mbar_init(full_b(s), 1); mbar_init(empty_b(s), 1); }
    mbar_init(d_full, 1);
    mbar_init(h_ready, 128);
    mbar_init(d_free, 128);
    mbar_fence_init();
  }
  if (warp == 8) { tmem_alloc(slot_addr, 512u); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot_ptr;
  const int KC1 = p.L[0].KC;

  if (warp < 4) {
    // ================================================================ A producer
    const int F = p.F;
    auto load_ids = [&](int tile, int32_t (&reg)[32]) {
      // element e = i*128 + tid of the tile's [128][F] id block
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        reg[i] = -1;
        if (i < F) {
          const int e = i * 128 + tid;
          const int smp = e / F, f = e - smp * F;
          const int64_t gm = (int64_t)tile * FT_BM + smp;
          if (gm < p.M) reg[i] = (int32_t)IdLoad<IDKIND>::at(p.ids, gm * p.ids_ld + f);
        }
      }
    };
    auto store_ids = [&](int buf, const int32_t (&reg)[32]) {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < F) s_ids[buf * (ids_bytes >> 2) + i * 128 + tid] = reg[i];
    };
    int32_t idreg[32];
    int local_tile = 0;
    if (GATHER) {
      if ((int)blockIdx.x < p.n_tiles) { load_ids(blockIdx.x, idreg); store_ids(0, idreg); }
      named_bar_sync(1, 128);
    }
    uint32_t it = 0;                 // global A-chunk counter (ring position)
    uint32_t signalled = 0;          // chunks whose full barrier has been arrived on
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++local_tile) {
      const int buf = local_tile & 1;
      const int next_tile = tile + gridDim.x;
      if (GATHER && next_tile < p.n_tiles) load_ids(next_tile, idreg);     // latency hidden behind this tile's stages
      const int64_t row0 = (int64_t)tile * FT_BM;
      for (int kc = 0; kc < KC1; ++kc, ++it) {
        const int s = it % FT_NA;
        mbar_wait(empty_a(s), ((it / FT_NA) & 1u) ^ 1u);
        const uint32_t stage = a_base + s * FT_A_BYTES;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int g = i * 128 + tid;
          const int smp = g >> 3, c = g & 7;
          const int64_t gm = row0 + smp;
          const uint32_t dst = stage + sw128_offset((uint32_t)smp, (uint32_t)c);
          const int col = kc * 32 + c * 4;
          if (!GATHER) {
            const bool ok = gm < p.M && col + 3 < p.L[0].K;
            if (ok || gm >= p.M || col >= p.L[0].K) {
              cp_async16(dst, ok ? (const void*)(p.x + gm * p.x_ld + col) : (const void*)p.x, ok ? 16u : 0u);
            } else {                                   // ragged K tail: element-wise, zero padded
              float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
              const float* src = p.x + gm * p.x_ld + col;
              if (col + 0 < p.L[0].K) t.x = __ldg(src + 0);
              if (col + 1 < p.L[0].K) t.y = __ldg(src + 1);
              if (col + 2 < p.L[0].K) t.z = __ldg(src + 2);
              *reinterpret_cast<float4*>(base_ptr + (dst - base)) = t;
            }
          } else if (col < p.emb) {
            const int q = col >> 2;
            const int f = q / p.k4, part = q - f * p.k4;
            const int32_t id = s_ids[buf * (ids_bytes >> 2) + smp * F + f];
            const int64_t r0 = p.row_off[f], r1 = p.row_off[f + 1];
            const bool ok = gm < p.M && id >= 0 && id < r1 - r0;
            if (!ok && gm < p.M && p.err_flag) *p.err_flag = 1;
            cp_async16(dst, ok ? (const void*)(p.arena + ((r0 + id) * p.k4 + part) * 4) : (const void*)p.arena,
                       ok ? 16u : 0u);
          } else {                                     // dense tail [emb, emb+D) then zero padding
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            const int j = col - p.emb;
            if (gm < p.M) {
              const float* src = p.dense + gm * p.dense_ld;
              if (j + 0 < p.D) t.x = __ldg(src + j + 0);
              if (j + 1 < p.D) t.y = __ldg(src + j + 1);
              if (j + 2 < p.D) t.z = __ldg(src + j + 2);
              if (j + 3 < p.D) t.w = __ldg(src + j + 3);
            }
            *reinterpret_cast<float4*>(base_ptr + (dst - base)) = t;
          }
        }
        cp_async_commit();
        if (it + 1 - signalled >= (uint32_t)FT_PA) {
          cp_async_wait<FT_PA - 1>();
          fence_proxy_async_smem();
          mbar_arrive(full_a(signalled % FT_NA));
          ++signalled;
        }
      }
      if (GATHER) {
        if (next_tile < p.n_tiles) store_ids(buf ^ 1, idreg);
        named_bar_sync(1, 128);
      }
    }
    // drain the cp.async pipeline
    cp_async_wait<0>();
    fence_proxy_async_smem();
    while (signalled < it) { mbar_arrive(full_a(signalled % FT_NA)); ++signalled; }
  } else if (warp < 8) {
    // ================================================================ epilogue
    const int ew = warp - 4;
    const uint32_t t_lane = tmem_base + ((uint32_t)(ew * 32) << 16);
    uint32_t dcount = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t gm = (int64_t)tile * FT_BM + ew * 32 + lane;
      for (int l = 0; l < p.nl; ++l) {
        const FtLayer& L = p.L[l];
        const bool last = (l == p.nl - 1);
        mbar_wait(d_full, dcount & 1u);
        ++dcount;
        tc_fence_after();
        float dot = 0.f;
        const int ncols = last ? L.BN : ((L.N + 31) / 32 * 32);
        for (int c0 = 0; c0 < ncols; c0 += 16) {
          float acc[16];
          if (c0 < L.BN) {
            tmem_ld16(t_lane + (uint32_t)c0, acc);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int n = c0 + j;
              float t = 0.f;
              if (n < L.N) {
                t = acc[j] + (L.bias ? __ldg(L.bias + n) : 0.f);
                t = L.ep.apply(t, gm, n, L.N);
                if (last) dot = fmaf(t, __ldg(p.w_out + n), dot);
                else t = to_tf32_rna(t);
              }
              acc[j] = t;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = 0.f;
          }
          if (!last) tmem_st16(t_lane + (uint32_t)(FT_H_COL + c0), acc);
        }
        if (!last) {
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(h_ready);
        } else {
          tc_fence_before();
          mbar_arrive(d_free);
          if (gm < p.M) {
            float y = dot + (p.b_out ? __ldg(p.b_out) : 0.f);
            if (p.y_scale) y *= __ldg(p.y_scale);
            if (p.add_vec) y += __ldg(p.add_vec + gm);
            if (p.add_bias) y += __ldg(p.add_bias);
            y *= p.scale;
            p.out[gm] = p.apply_sigmoid ? 1.0f / (1.0f + expf(-y)) : y;
          }
        }
      }
    }
  } else if (warp == 8) {
    // ================================================================ MMA issuer
    if (lane == 0) {
      uint32_t ia = 0, ib = 0, hcount = 0, fcount = 0;
      int local_tile = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++local_tile) {
        if (local_tile > 0) { mbar_wait(d_free, fcount & 1u); ++fcount; tc_fence_after(); }
        for (int l = 0; l < p.nl; ++l) {
          const FtLayer& L = p.L[l];
          const uint32_t idesc = make_idesc_tf32(FT_BM, (uint32_t)L.BN);
          if (l > 0) { mbar_wait(h_ready, hcount & 1u); ++hcount; tc_fence_after(); }
          for (int kc = 0; kc < L.KC; ++kc, ++ib) {
            const int sb = ib % FT_NB;
            const uint64_t db = make_desc_sw128(b_base + sb * p.b_stage_bytes);
            if (l == 0) {
              const int sa = ia % FT_NA;
              mbar_wait(full_a(sa), (ia / FT_NA) & 1u);
              mbar_wait(full_b(sb), (ib / FT_NB) & 1u);
              tc_fence_after();
              const uint64_t da = make_desc_sw128(a_base + sa * FT_A_BYTES);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                mma_tf32_ss(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kc | k) != 0));
              mma_commit(empty_a(sa));
              ++ia;
            } else {
              mbar_wait(full_b(sb), (ib / FT_NB) & 1u);
              tc_fence_after();
#pragma unroll
              for (int k = 0; k < 4; ++k)
                mma_tf32_ts(tmem_base, tmem_base + (uint32_t)(FT_H_COL + kc * 32 + k * 8), db + (uint64_t)(2 * k), idesc,
                            (uint32_t)((kc | k) != 0));
            }
            mma_commit(empty_b(sb));
          }
          mma_commit(d_full);
        }
      }
    }
    __syncwarp();
  } else {
    // ================================================================ W producer
    if (lane == 0) {
      uint32_t ib = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        for (int l = 0; l < p.nl; ++l) {
          const FtLayer& L = p.L[l];
          const uint32_t bytes = (uint32_t)L.BN * 128u;
          for (int kc = 0; kc < L.KC; ++kc, ++ib) {
            const int sb = ib % FT_NB;
            mbar_wait(empty_b(sb), ((ib / FT_NB) & 1u) ^ 1u);
            mbar_arrive_expect_tx(full_b(sb), bytes);
            bulk_g2s(b_base + sb * p.b_stage_bytes, L.Wp + (size_t)kc * L.BN * 32, bytes, full_b(sb));
          }
        }
      }
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == 8) { tc_fence_after(); tmem_dealloc(tmem_base, 512u); }
}

}  // namespace ctrhot

using namespace ctrhot;

static int ft_setup(FtParams& p, const ctrhot_dense_desc* h, int n_layers, size_t& smem, bool gather) {
  if (!h || n_layers < 2 || n_layers > FT_MAXL + 1) return CTRHOT_ESHAPE;
  const ctrhot_dense_desc& o = h[n_layers - 1];
  if (o.out_dim != 1 || o.act != CTRHOT_ACT_NONE || !o.weight) return CTRHOT_ESHAPE;
  p.nl = n_layers - 1;
  int bn_max = 16;
  for (int l = 0; l < p.nl; ++l) {
    const ctrhot_dense_desc& d = h[l];
    if (!d.weight || !d.packed || d.in_dim <= 0 || d.out_dim <= 0) return CTRHOT_EARG;
    if (d.out_dim > 256 || d.out_dim < 8) return CTRHOT_ESHAPE;
    if (l > 0 && d.in_dim != h[l - 1].out_dim) return CTRHOT_EARG;
    if ((d.act == CTRHOT_ACT_PRELU || d.act == CTRHOT_ACT_DICE) && !d.act_param) return CTRHOT_EARG;
    const TcGeom g = tc_geom(d.in_dim, d.out_dim);
    FtLayer& L = p.L[l];
    L.Wp = static_cast<const float*>(d.packed); L.bias = d.bias; L.ep = Epilogue::from(d);
    L.K = d.in_dim; L.N = d.out_dim; L.KC = g.KC; L.BN = g.BN;
    if (g.BN > bn_max) bn_max = g.BN;
  }
  if (o.in_dim != h[n_layers - 2].out_dim) return CTRHOT_EARG;
  p.w_out = o.weight; p.b_out = o.bias;
  p.b_stage_bytes = (uint32_t)bn_max * 128u;
  const size_t ids_bytes = gather ? (((size_t)FT_BM * p.F * 4 + 15) & ~(size_t)15) : 0;
  smem = (size_t)FT_NA * FT_A_BYTES + (size_t)FT_NB * p.b_stage_bytes + 2 * ids_bytes + 256 + 1024;
  if (smem > 227 * 1024) return CTRHOT_ESHAPE;
  return CTRHOT_OK;
}

template <int IDKIND, bool GATHER>
static int ft_launch(const FtParams& p, size_t smem, cudaStream_t st) {
  auto kern = tower_fused_kernel<IDKIND, GATHER>;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "tower_fused attr");
  const int grid = p.n_tiles < kNumSMs ? p.n_tiles : kNumSMs;
  kern<<<grid, FT_THREADS, smem, st>>>(p);
  CTRHOT_LAUNCH_OK("ctrhot_tower_fused_tf32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_tower_fused_tf32(const float* x, int64_t x_ld, int64_t M,
                                       const ctrhot_dense_desc* h_layers, int n_layers,
                                       const float* y_scale, const float* add_vec, const float* add_bias,
                                       float scale, int apply_sigmoid, float* out, void* stream) {
  CTRHOT_RANGE("ctrhot_tower_fused_tf32");
  if (!x || !out || M < 0) return CTRHOT_EARG;
  if (x_ld % 4 != 0 || !aligned16(x)) return CTRHOT_ESHAPE;
  FtParams p{};
  p.x = x; p.x_ld = x_ld; p.M = M; p.n_tiles = (int)((M + FT_BM - 1) / FT_BM);
  p.y_scale = y_scale; p.add_vec = add_vec; p.add_bias = add_bias; p.scale = scale; p.apply_sigmoid = apply_sigmoid;
  p.out = out;
  size_t smem = 0;
  const int rc = ft_setup(p, h_layers, n_layers, smem, false);
  if (rc != CTRHOT_OK) return rc;
  if (x_ld < p.L[0].K) return CTRHOT_EARG;
  if (M == 0) return CTRHOT_OK;
  return ft_launch<CTRHOT_IDS_I32, false>(p, smem, static_cast<cudaStream_t>(stream));
}

extern "C" int ctrhot_tower_gather_tf32(const float* arena, const int64_t* row_off, int F, int k,
                                        const void* ids, int ids_kind, int64_t ids_ld,
                                        const float* dense, int64_t dense_ld, int D, int64_t M,
                                        const ctrhot_dense_desc* h_layers, int n_layers,
                                        const float* y_scale, const float* add_vec, const float* add_bias,
                                        float scale, int apply_sigmoid, float* out, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_tower_gather_tf32");
  if (!arena || !row_off || !ids || !out || F <= 0 || k <= 0 || M < 0 || D < 0) return CTRHOT_EARG;
  if (D > 0 && !dense) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (k % 4 != 0 || !aligned16(arena) || F > 32) return CTRHOT_ESHAPE;
  FtParams p{};
  p.arena = arena; p.row_off = row_off; p.F = F; p.k4 = k / 4; p.ids = ids; p.ids_ld = ids_ld;
  p.dense = dense; p.dense_ld = dense_ld; p.D = D; p.emb = F * k;
  p.M = M; p.n_tiles = (int)((M + FT_BM - 1) / FT_BM);
  p.y_scale = y_scale; p.add_vec = add_vec; p.add_bias = add_bias; p.scale = scale; p.apply_sigmoid = apply_sigmoid;
  p.out = out; p.err_flag = err_flag;
  size_t smem = 0;
  const int rc = ft_setup(p, h_layers, n_layers, smem, true);
  if (rc != CTRHOT_OK) return rc;
  if (p.L[0].K != D + F * k) return CTRHOT_EARG;
  if (M == 0) return CTRHOT_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return ids_kind == CTRHOT_IDS_I32 ? ft_launch<CTRHOT_IDS_I32, true>(p, smem, st)
                                    : ft_launch<CTRHOT_IDS_F32, true>(p, smem, st);
}
