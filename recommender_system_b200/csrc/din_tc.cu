// K6: DIN local-activation attention on the tensor pipe (Attention.call, DL/layer/interaction.py:369-406, fed by the
// history / candidate embedding lookups of DIN.call, DL/model/din.py:76-81).
//
//   a_t   = [q, k_t, q - k_t, q * k_t]                     (interaction.py:381-391; never materialised)
//   h1_t  = PReLU_{alpha1[t,:]}(a_t W1 + b1)               W1:[4K,h1]  (interaction.py:366,393-394)
//   h2_t  = PReLU_{alpha2[t,:]}(h1_t W2 + b2)              W2:[h1,h2]
//   s_t   = h2_t . w3 + b3;  s_t = -2^32+1 where hist id == 0;  p = softmax_T(s);  out = sum_t p_t k_t   (interaction.py:396-405)
//
// The first layer is split by linearity:  a_t W1 = q (W1a + W1c) + k_t (W1b - W1c) + (q * k_t) W1d.  The q-term is one
// [K] x [K,h1] product per SAMPLE (done by din_prep_kernel with the candidate gather, fp32), so the per-(sample, t) GEMM
// has K-extent 2K instead of 4K.  This changes rounding only (SURVEY 8(d) cfg 4 allows it).
//
// One persistent CTA per SM, one 128-row MMA tile per SAMPLE (rows t >= T are zero), 17 warps:
//   warps 4-7   gather: 16-byte cp.async of the T history rows (+ q and u = q Wq + b1 of this sample) into a 3-stage smem
//               ring; history ids prefetched one sample ahead; completion by mbarrier.
//   warps 8-11  stager: thread t reads key row t from smem and writes [k_t | q * k_t] into the TMEM A slot (tcgen05.st).
//   warp 12     MMA issuer: D1 = A W' (TS form, B resident in smem), D2 = H1 W2; two TMEM slots so that the epilogue of
//               sample s overlaps the MMAs of s+1:   L1(s), L2(s-1), L1(s+1), L2(s), ...
//   warps 0-3   epilogue 1: thread t owns TMEM lane t.  D1 + u -> PReLU -> TF32 back in place (A of layer 2).
//   warps 13-16 epilogue 2: D2 + b2 -> PReLU -> . w3 -> mask -> softmax over the CTA's 128 lanes -> out = p @ keys (keys
//               re-read from the smem stage) -> 4K bytes per sample to HBM.  Two warp groups so that the two epilogues
//               of consecutive samples run concurrently.
// All weights (W', W2, alpha1, alpha2, ...) stay resident in shared memory for the whole launch.
// HBM traffic per sample = (T + 1) ids + (T + 1) rows + u + out; tensor work = 2 T (2K h1 + h1 h2) flop.
#include "common.cuh"
#include "tc05.cuh"

namespace ctrhot {
using namespace tc;

constexpr int DT_BM = 128;
constexpr int DT_NRAW = 4;
constexpr int DT_THREADS = 17 * 32;
constexpr int DT_HDR_BYTES = 1536;            // per stage: ids[128] int32 | q[<=64] | u[<=96]  (1024 + 96 * 4 <= 1536)
constexpr int DT_HDR_Q = 512, DT_HDR_U = 1024;

struct DinShape {
  int T, K, h1, h2;
  int N1p, N2p;            // out dims rounded up to 16 (MMA N)
  int AK;                  // 2K: K-extent of layer 1
  int nchK, nchA, nch2;    // 32-column chunks: key tile, W', W2
  int SR;                  // key rows a stage holds: T rounded up to 8 (rows T..127 of the MMA tile are zeros made by the stager)
  int S1, S2;              // padded alpha row strides (floats): conflict-free one-row-per-thread reads
  // packed blob offsets (floats)
  int o_wq, o_b1, o_wa, o_w2, o_al1, o_al2, o_b2, o_w3, o_b3, total;
  // TMEM columns inside a slot
  int c_a, c_d1, c_d2, slot_cols;
};

__host__ __device__ inline int dt_round(int v, int m) { return (v + m - 1) / m * m; }

__host__ __device__ inline bool din_shape(int T, int K, int h1, int h2, DinShape& s) {
  if (T < 1 || T > DT_BM) return false;
  if (!(K == 8 || K == 16 || K == 32 || K == 64)) return false;
  if (h1 < 8 || h1 > 96 || h2 < 1 || h2 > 64) return false;
  s.T = T; s.K = K; s.h1 = h1; s.h2 = h2;
  s.SR = dt_round(T, 8);
  s.N1p = dt_round(h1, 16); s.N2p = dt_round(h2, 16);
  s.AK = 2 * K;
  s.nchK = (K + 31) / 32; s.nchA = (s.AK + 31) / 32; s.nch2 = (s.N1p + 31) / 32;
  s.S1 = s.N1p + 4; s.S2 = s.N2p + 4;
  int o = 0;
  s.o_wq = o; o += K * s.N1p;
  s.o_b1 = o; o += s.N1p;
  s.o_wa = o; o += s.nchA * s.N1p * 32;
  s.o_w2 = o; o += s.nch2 * s.N2p * 32;
  s.o_al1 = o; o += T * s.S1;
  s.o_al2 = o; o += T * s.S2;
  s.o_b2 = o; o += s.N2p;
  s.o_w3 = o; o += s.N2p;
  s.o_b3 = o; o += 4;
  s.total = o;
  s.c_a = 0; s.c_d1 = s.AK; s.c_d2 = s.AK + s.N1p; s.slot_cols = s.AK + s.N1p + s.N2p;
  return s.slot_cols <= 256;
}

// ---------------------------------------------------------------- pack (weights -> one device blob)
__global__ void din_pack_kernel(const float* __restrict__ W1, const float* __restrict__ b1, const float* __restrict__ al1,
                                const float* __restrict__ W2, const float* __restrict__ b2, const float* __restrict__ al2,
                                const float* __restrict__ w3, const float* __restrict__ b3, int ar1, int ar2, DinShape s,
                                float* __restrict__ P) {
  const int K = s.K, h1 = s.h1, h2 = s.h2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < s.total; i += gridDim.x * blockDim.x) {
    float v = 0.f;
    if (i < s.o_b1) {                                          // Wq[c][n] = W1a + W1c
      const int c = i / s.N1p, n = i - c * s.N1p;
      if (n < h1) v = W1[(size_t)c * h1 + n] + W1[(size_t)(2 * K + c) * h1 + n];
    } else if (i < s.o_wa) {
      const int n = i - s.o_b1;
      if (n < h1 && b1) v = b1[n];
    } else if (i < s.o_w2) {                                   // W' chunks, K-major, 128B swizzle
      const int e = i - s.o_wa;
      const int ch = e / (s.N1p * 32), r = (e / 32) % s.N1p, w = e % 32;
      const int piece = (w >> 2) ^ (r & 7), kin = ch * 32 + piece * 4 + (w & 3);     // inverse of the swizzle
      if (r < h1 && kin < s.AK) {
        const float t = (kin < K) ? (W1[(size_t)(K + kin) * h1 + r] - W1[(size_t)(2 * K + kin) * h1 + r])
                                  : W1[(size_t)(3 * K + (kin - K)) * h1 + r];
        v = to_tf32_rna(t);
      }
    } else if (i < s.o_al1) {
      const int e = i - s.o_w2;
      const int ch = e / (s.N2p * 32), r = (e / 32) % s.N2p, w = e % 32;
      const int piece = (w >> 2) ^ (r & 7), kin = ch * 32 + piece * 4 + (w & 3);
      if (r < h2 && kin < h1) v = to_tf32_rna(W2[(size_t)kin * h2 + r]);
    } else if (i < s.o_al2) {
      const int e = i - s.o_al1;
      const int t = e / s.S1, n = e - t * s.S1;
      if (n < h1 && al1) v = al1[(size_t)(t % ar1) * h1 + n];
    } else if (i < s.o_b2) {
      const int e = i - s.o_al2;
      const int t = e / s.S2, n = e - t * s.S2;
      if (n < h2 && al2) v = al2[(size_t)(t % ar2) * h2 + n];
    } else if (i < s.o_w3) {
      const int n = i - s.o_b2;
      if (n < h2 && b2) v = b2[n];
    } else if (i < s.o_b3) {
      const int n = i - s.o_w3;
      if (n < h2) v = w3[n];
    } else if (i == s.o_b3) {
      v = b3 ? b3[0] : 0.f;
    }
    P[i] = v;
  }
}

// ---------------------------------------------------------------- prep: candidate gather + u = q Wq + b1
// One CTA = DP_S samples: the candidate rows are gathered (coalesced, and written to item_out), kept in shared memory,
// then thread (column n, sample group) forms u[s][n] for its samples with one Wq load per (c, n) shared by all of them.
constexpr int DP_S = 32;          // samples per CTA
constexpr int DP_T = 256;
template <int IDKIND>
__global__ void __launch_bounds__(DP_T)
din_prep_kernel(const float* __restrict__ table, int64_t V, int K, const void* __restrict__ cand_ids, int64_t cand_ld,
                int64_t B, const float* __restrict__ Wq, const float* __restrict__ b1, int N1p,
                float* __restrict__ item_out, int64_t item_ld, float* __restrict__ U, int* __restrict__ err_flag) {
  __shared__ float sq[DP_S][64 + 1];
  const int tid = threadIdx.x;
  const int64_t b0 = (int64_t)blockIdx.x * DP_S;
  // two batches of independent loads (all ids of this thread's elements, then all rows) instead of an id -> row round trip
  // per element: the loop was 2 x DRAM latency x 8 iterations long
  constexpr int PE = 8;                               // elements per thread and pass
  for (int e0 = 0; e0 < DP_S * K; e0 += DP_T * PE) {
    int64_t id[PE];
    float v[PE];
#pragma unroll
    for (int i = 0; i < PE; ++i) {
      const int e = e0 + tid + DP_T * i;
      const int s = e / K;
      id[i] = -1;
      if (e < DP_S * K && b0 + s < B) id[i] = IdLoad<IDKIND>::at(cand_ids, (b0 + s) * cand_ld);
    }
#pragma unroll
    for (int i = 0; i < PE; ++i) {
      const int e = e0 + tid + DP_T * i;
      const int s = e / K, c = e - s * K;
      v[i] = 0.f;
      if (e < DP_S * K && b0 + s < B) {
        const bool ok = id[i] >= 0 && id[i] < V;
        if (!ok && err_flag && c == 0) *err_flag = 1;
        if (ok) v[i] = ldg_nc_f1(table + id[i] * K + c);
      }
    }
#pragma unroll
    for (int i = 0; i < PE; ++i) {
      const int e = e0 + tid + DP_T * i;
      const int s = e / K, c = e - s * K;
      if (e < DP_S * K) {
        if (b0 + s < B) item_out[(b0 + s) * item_ld + c] = v[i];
        sq[s][c] = v[i];
      }
    }
  }
  __syncthreads();
  const int groups = DP_T / N1p;                      // column n = tid % N1p, sample group = tid / N1p
  const int n = tid % N1p, g = tid / N1p;
  if (g >= groups) return;
  constexpr int MAXS = 16;
  const int per = (DP_S + groups - 1) / groups;       // samples per group (<= MAXS for N1p <= 96 -> groups >= 2)
  float acc[MAXS];
  const float bias = __ldg(b1 + n);
#pragma unroll
  for (int i = 0; i < MAXS; ++i) acc[i] = bias;
  for (int c0 = 0; c0 < K; c0 += 4) {                  // K is 8, 16, 32 or 64: four weight loads in flight
    float w[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) w[u] = __ldg(Wq + (size_t)(c0 + u) * N1p + n);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int i = 0; i < MAXS; ++i)
        if (i < per) acc[i] = fmaf(sq[(g * per + i) % DP_S][c0 + u], w[u], acc[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < MAXS; ++i) {
    const int s = g * per + i;
    if (i < per && s < DP_S && b0 + s < B) U[(b0 + s) * N1p + n] = acc[i];
  }
}

// ---------------------------------------------------------------- tcgen05 helpers local to this kernel
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
                  "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
                  "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])) : "memory");
}
__device__ __forceinline__ float4 dt_lds128(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void dt_ld16_nowait(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void dt_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float dt_prelu(float v, float a) { return fmaxf(v, 0.f) + a * fminf(v, 0.f); }
// round-to-nearest (ties away) TF32 of a finite value: two integer ops instead of cvt.rna's NaN-preserving sequence
__device__ __forceinline__ float dt_tf32_fast(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u); }

struct DinParams {
  const float* table; int64_t V;
  const void* hist_ids; int64_t hist_ld; int ids_kind;
  const float* item; int64_t item_ld;        // q rows (written by din_prep_kernel)
  const float* U;                            // [B, N1p]
  const float* packed;
  float* out; int64_t out_ld;
  int64_t B;
  int* err_flag;
  int nraw;
  DinShape s;
};

// shared-memory map (bytes from the 1024-aligned base); the same function sizes the launch on the host
struct DinSmem { uint32_t wa, w2, stage0, stage_bytes, al1, al2, small, bars, total; };
__host__ __device__ inline DinSmem din_smem(const DinShape& s, int nraw) {
  DinSmem m;
  uint32_t o = 0;
  m.wa = o; o += (uint32_t)s.nchA * s.N1p * 128u;
  m.w2 = o; o += (uint32_t)s.nch2 * s.N2p * 128u;
  o = (o + 1023u) & ~1023u;
  m.stage_bytes = (uint32_t)s.nchK * (uint32_t)s.SR * 128u + DT_HDR_BYTES;
  m.stage0 = o; o += (uint32_t)nraw * m.stage_bytes;
  m.al1 = o; o += (uint32_t)s.T * s.S1 * 4u;
  m.al2 = o; o += (uint32_t)s.T * s.S2 * 4u;
  m.small = o; o += (uint32_t)(2 * s.N2p + 4 + DT_BM + 16 + 4 * DT_BM + 64) * 4u; // b2 | w3 | b3 | P[128] | red[16] | part[512] | row0[64]
  o = (o + 15u) & ~15u;
  m.bars = o; o += 8u * (2 * DT_NRAW + 12) + 16u;
  m.total = o;
  return m;
}

#ifdef DT_TRACE
// tuning builds only: clock64 of pipeline events of CTA 0, [event][sample]
__device__ long long dt_trace[8][128];
#define DT_MARK(ev, i) do { if (blockIdx.x == 0 && (i) < 128 && (threadIdx.x & 31) == 0) dt_trace[ev][i] = clock64(); } while (0)
#else
#define DT_MARK(ev, i) do { } while (0)
#endif

template <int IDKIND>
__global__ void __launch_bounds__(DT_THREADS, 1)
din_tc_kernel(const __grid_constant__ DinParams p) {
  extern __shared__ uint8_t smem_raw[];
  const DinShape& s = p.s;
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* bp = smem_raw + (base - raw_addr);
  const DinSmem mp = din_smem(s, p.nraw);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int nraw = p.nraw;
  const uint32_t bars = base + mp.bars;
  auto full_raw = [&](int i) { return bars + 8u * i; };
  auto raw_free = [&](int i) { return bars + 8u * (DT_NRAW + i); };
  const uint32_t mb = bars + 8u * (2 * DT_NRAW);
  auto a_full = [&](int i) { return mb + 8u * i; };
  auto a_free = [&](int i) { return mb + 8u * (2 + i); };
  auto d1_full = [&](int i) { return mb + 8u * (4 + i); };
  auto h1_ready = [&](int i) { return mb + 8u * (6 + i); };
  auto d2_full = [&](int i) { return mb + 8u * (8 + i); };
  auto d2_free = [&](int i) { return mb + 8u * (10 + i); };
  const uint32_t slot_addr = mb + 8u * 12;
  volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(bp + (slot_addr - base));
  float* f_small = reinterpret_cast<float*>(bp + mp.small);
  float* s_b2 = f_small; float* s_w3 = s_b2 + s.N2p; float* s_b3 = s_w3 + s.N2p;
  float* s_P = s_b3 + 4; float* s_red = s_P + DT_BM; float* s_part = s_red + 16;
  float* s_row0 = s_part + 4 * DT_BM;            // table row 0: the key of every padding position (history id 0)

  if (tid == 0) {
    for (int i = 0; i < DT_NRAW; ++i) { mbar_init(full_raw(i), 256); mbar_init(raw_free(i), 128); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(a_full(i), 128); mbar_init(a_free(i), 1); mbar_init(d1_full(i), 1);
      mbar_init(h1_ready(i), 128); mbar_init(d2_full(i), 1); mbar_init(d2_free(i), 128);
    }
    mbar_fence_init();
  }
  // resident weights: the packed blob is already in its shared-memory layout
  {
    const float4* src = reinterpret_cast<const float4*>(p.packed + s.o_wa);
    float4* dst = reinterpret_cast<float4*>(bp + mp.wa);
    const int n4 = (s.nchA * s.N1p + s.nch2 * s.N2p) * 8;
    for (int i = tid; i < n4; i += DT_THREADS) dst[i] = __ldg(src + i);
    float* a1 = reinterpret_cast<float*>(bp + mp.al1);
    for (int i = tid; i < s.T * s.S1; i += DT_THREADS) a1[i] = __ldg(p.packed + s.o_al1 + i);
    float* a2 = reinterpret_cast<float*>(bp + mp.al2);
    for (int i = tid; i < s.T * s.S2; i += DT_THREADS) a2[i] = __ldg(p.packed + s.o_al2 + i);
    for (int i = tid; i < 2 * s.N2p + 4; i += DT_THREADS) f_small[i] = __ldg(p.packed + s.o_b2 + i);
    for (int i = tid; i < s.K; i += DT_THREADS) s_row0[i] = (p.V > 0) ? __ldg(p.table + i) : 0.f;
  }
  fence_proxy_async_smem();
  if (warp == 12) { tmem_alloc(slot_addr, 512u); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot_ptr;
  const int64_t n_my = (p.B > blockIdx.x) ? (p.B - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int T = s.T, K = s.K;

  if (warp >= 4 && warp < 8) {
    // ================================================================ gather
    // Thread tg fetches the history id of row tg (two samples ahead), the 128 ids go through the stage header (the mask
    // needs them there anyway); then PPR adjacent threads copy one key row together (coalesced 16-byte pieces).
    const int tg = tid - 128;
    const int PPR = K / 4;                                   // 16-byte pieces per key row (2, 4, 8 or 16)
    const int c = tg % PPR, rsub = tg / PPR, rpp = DT_BM / PPR;  // this thread: piece c of rows rsub + rpp * j, j < PPR
    const uint32_t rowbytes = (uint32_t)K * 4u;
    const uint32_t chunk = (uint32_t)(c >> 3) * ((uint32_t)s.SR * 128u), pc = (uint32_t)(c & 7);
    auto fetch = [&](int64_t i) -> int32_t {
      if (i >= n_my || tg >= T) return 0;
      const int64_t b = blockIdx.x + i * gridDim.x;
      return __ldg(static_cast<const int32_t*>(p.hist_ids) + b * p.hist_ld + tg);
    };
    auto issue = [&](int64_t i, int32_t raw_id) {
      const int st = (int)(i % nraw);
      const int64_t b = blockIdx.x + i * gridDim.x;
      mbar_wait(raw_free(st), (uint32_t)((i / nraw) & 1) ^ 1u);
      const uint32_t stage = base + mp.stage0 + (uint32_t)st * mp.stage_bytes;
      const uint32_t hdr = stage + (uint32_t)s.nchK * (uint32_t)s.SR * 128u;
      int32_t* ids_s = reinterpret_cast<int32_t*>(bp + (hdr - base));
      {
        const int32_t id = (IDKIND == CTRHOT_IDS_I32) ? raw_id : __float2int_rz(__int_as_float(raw_id));
        const bool ok = tg >= T || (id >= 0 && (int64_t)id < p.V);
        if (!ok && p.err_flag) *p.err_flag = 1;
        ids_s[tg] = (tg < T) ? (ok ? id : -1) : 0;           // -1: out of range -> zero row, not masked
      }
      named_bar_sync(2, 128);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int r = rsub + rpp * j;
        if (j < PPR && r < s.SR) {                       // rows >= SR are not stored: the stager writes zeros for them
          const int32_t id = ids_s[r];
          // id 0 = padding (masked: softmax weight exactly 0 unless the WHOLE history is padding): its key is table row 0,
          // which epilogue 2 adds from smem in that one case -- not fetching it halves the gather of a half-padded batch and
          // removes 16384 x ~50 reads of one 256-byte line (all served by a single L2 slice)
          const bool ok = r < T && id > 0;
          cp_async16(stage + chunk + (uint32_t)r * 128u + ((pc ^ (uint32_t)(r & 7)) << 4),
                     reinterpret_cast<const char*>(p.table) + (ok ? (uint64_t)(uint32_t)id * rowbytes : 0ull) + (uint32_t)c * 16u,
                     ok ? 16u : 0u);
        }
      }
      if (tg < PPR) cp_async16(hdr + DT_HDR_Q + tg * 16u, p.item + b * p.item_ld + tg * 4, 16u);
      else if (tg < PPR + s.N1p / 4) cp_async16(hdr + DT_HDR_U + (tg - PPR) * 16u, p.U + b * s.N1p + (tg - PPR) * 4, 16u);
      cp_async_mbar_arrive_noinc(full_raw(st));
      mbar_arrive(full_raw(st));                             // releases the ids written with plain stores
      if (tg == 0) DT_MARK(0, i);
    };
    int32_t ra = fetch(0), rb = fetch(1), rc;
    for (int64_t i = 0; i < n_my; i += 3) {
      rc = fetch(i + 2); issue(i, ra);
      if (i + 1 < n_my) { ra = fetch(i + 3); issue(i + 1, rb); }
      if (i + 2 < n_my) { rb = fetch(i + 4); issue(i + 2, rc); }
    }
  } else if (warp >= 8 && warp < 12) {
    // ================================================================ stager: keys -> TMEM A slot [k | q * k]
    const int t = tid - 256;
    const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t rx = (uint32_t)(t & 7);
    for (int64_t i = 0; i < n_my; ++i) {
      const int st = (int)(i % nraw), slot = (int)(i & 1);
      const uint32_t stage = base + mp.stage0 + (uint32_t)st * mp.stage_bytes;
      const uint32_t hq = stage + (uint32_t)s.nchK * (uint32_t)s.SR * 128u + DT_HDR_Q;
      mbar_wait(full_raw(st), (uint32_t)((i / nraw) & 1));
      if (t == 0) DT_MARK(1, i);
      mbar_wait(a_free(slot), (uint32_t)((i >> 1) & 1) ^ 1u);
      tc_fence_after();
      const uint32_t a_t = t_lane + (uint32_t)(slot * s.slot_cols + s.c_a);
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        if (g >= K / 8) break;
        const uint32_t row = stage + (uint32_t)(g >> 2) * ((uint32_t)s.SR * 128u) + (uint32_t)(t < s.SR ? t : 0) * 128u;
        const uint32_t p0 = (uint32_t)((2 * g) & 7), p1 = p0 + 1;
        float4 k0 = dt_lds128(row + ((p0 ^ rx) << 4)), k1 = dt_lds128(row + ((p1 ^ rx) << 4));
        if (t >= s.SR) { k0 = make_float4(0.f, 0.f, 0.f, 0.f); k1 = k0; }
        const float4 q0 = dt_lds128(hq + (uint32_t)g * 32u), q1 = dt_lds128(hq + (uint32_t)g * 32u + 16u);
        const float kv[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
        const float qk[8] = {k0.x * q0.x, k0.y * q0.y, k0.z * q0.z, k0.w * q0.w, k1.x * q1.x, k1.y * q1.y, k1.z * q1.z, k1.w * q1.w};
        tmem_st8(a_t + (uint32_t)(8 * g), kv);
        tmem_st8(a_t + (uint32_t)(K + 8 * g), qk);
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(a_full(slot));
      if (t == 0) DT_MARK(2, i);
    }
  } else if (warp == 12) {
    // ================================================================ MMA issuer: L1(i), L2(i-1), L1(i+1), ...
    if (lane == 0) {
      const uint32_t idesc1 = make_idesc_tf32(DT_BM, (uint32_t)s.N1p), idesc2 = make_idesc_tf32(DT_BM, (uint32_t)s.N2p);
      const uint32_t wa = base + mp.wa, w2 = base + mp.w2;
      auto layer2 = [&](int64_t i) {
        const int slot = (int)(i & 1);
        tc_fence_after();
        const uint32_t a_t = tmem_base + (uint32_t)(slot * s.slot_cols + s.c_d1);
        const uint32_t d_t = tmem_base + (uint32_t)(slot * s.slot_cols + s.c_d2);
        for (int kk = 0; kk < s.N1p / 8; ++kk) {
          const uint64_t db = make_desc_sw128(w2 + (uint32_t)(kk >> 2) * (uint32_t)s.N2p * 128u) + (uint64_t)(2 * (kk & 3));
          mma_tf32_ts(d_t, a_t + (uint32_t)(8 * kk), db, idesc2, (uint32_t)(kk != 0));
        }
        mma_commit(d2_full(slot));
      };
      auto layer1 = [&](int64_t i) {
        const int slot = (int)(i & 1);
        tc_fence_after();
        const uint32_t a_t = tmem_base + (uint32_t)(slot * s.slot_cols + s.c_a);
        const uint32_t d_t = tmem_base + (uint32_t)(slot * s.slot_cols + s.c_d1);
        for (int kk = 0; kk < s.AK / 8; ++kk) {
          const uint64_t db = make_desc_sw128(wa + (uint32_t)(kk >> 2) * (uint32_t)s.N1p * 128u) + (uint64_t)(2 * (kk & 3));
          mma_tf32_ts(d_t, a_t + (uint32_t)(8 * kk), db, idesc1, (uint32_t)(kk != 0));
        }
        mma_commit(a_free(slot));
        mma_commit(d1_full(slot));
      };
      // Whatever is ready goes first, finishing samples before starting new ones: L2(i2) needs H1(i2) and a free D2 slot;
      // L1(i1) needs A(i1) and must not overwrite H1(i1 - 2) before L2(i1 - 2) was issued (two TMEM slots).
      int64_t i1 = 0, i2 = 0;
      while (i2 < n_my) {
        if (i2 < i1) {
          const int slot = (int)(i2 & 1);
          const uint32_t ph = (uint32_t)((i2 >> 1) & 1);
          if (mbar_try_wait(h1_ready(slot), ph) && mbar_try_wait(d2_free(slot), ph ^ 1u)) { layer2(i2); ++i2; continue; }
        }
        if (i1 < n_my && i1 - i2 < 2) {
          if (mbar_try_wait(a_full((int)(i1 & 1)), (uint32_t)((i1 >> 1) & 1))) { layer1(i1); ++i1; }
        }
      }
    }
    __syncwarp();
  } else if (warp < 4) {
    // ================================================================ epilogue 1 (warps 0-3): thread t = TMEM lane t
    const int t = tid;
    const uint32_t t_lane = tmem_base + ((uint32_t)(warp * 32) << 16);
    const float* al1 = reinterpret_cast<const float*>(bp + mp.al1) + (size_t)(t < T ? t : 0) * s.S1;
    const bool exact = (s.h1 == s.N1p);
    auto chunk = [&](uint32_t (&r)[16], int n0, const float* u, uint32_t d_t) {
      float acc[16];
#pragma unroll
      for (int j = 0; j < 16; j += 4) {
        const float4 uu = *reinterpret_cast<const float4*>(u + n0 + j);
        const float4 aa = *reinterpret_cast<const float4*>(al1 + n0 + j);
        acc[j + 0] = dt_tf32_fast(dt_prelu(__uint_as_float(r[j + 0]) + uu.x, aa.x));
        acc[j + 1] = dt_tf32_fast(dt_prelu(__uint_as_float(r[j + 1]) + uu.y, aa.y));
        acc[j + 2] = dt_tf32_fast(dt_prelu(__uint_as_float(r[j + 2]) + uu.z, aa.z));
        acc[j + 3] = dt_tf32_fast(dt_prelu(__uint_as_float(r[j + 3]) + uu.w, aa.w));
      }
      if (!exact) {
#pragma unroll
        for (int j = 0; j < 16; ++j) if (n0 + j >= s.h1) acc[j] = 0.f;
      }
      tmem_st16(d_t + (uint32_t)n0, acc);
    };
    for (int64_t i = 0; i < n_my; ++i) {
      const int st = (int)(i % nraw), slot = (int)(i & 1);
      const float* u = reinterpret_cast<const float*>(bp + mp.stage0 + (uint32_t)st * mp.stage_bytes +
                                                      (uint32_t)s.nchK * (uint32_t)s.SR * 128u + DT_HDR_U);
      mbar_wait(d1_full(slot), (uint32_t)((i >> 1) & 1));
      if (t == 0) DT_MARK(3, i);
      tc_fence_after();
      const uint32_t d_t = t_lane + (uint32_t)(slot * s.slot_cols + s.c_d1);
      uint32_t r0[16], r1[16];                              // two chunks in flight: load n+1 while n is processed
      dt_ld16_nowait(d_t, r0);
      dt_ld_wait();
      for (int n0 = 0; n0 < s.N1p; n0 += 32) {
        if (n0 + 16 < s.N1p) dt_ld16_nowait(d_t + (uint32_t)(n0 + 16), r1);
        chunk(r0, n0, u, d_t);
        dt_ld_wait();
        if (n0 + 32 < s.N1p) dt_ld16_nowait(d_t + (uint32_t)(n0 + 32), r0);
        if (n0 + 16 < s.N1p) chunk(r1, n0 + 16, u, d_t);
        dt_ld_wait();
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(h1_ready(slot));
      if (t == 0) DT_MARK(4, i);
    }
  } else {
    // ================================================================ epilogue 2 (warps 13-16): scores, softmax, weighted sum
    const int q4 = warp & 3;
    const int t = q4 * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
    const float* al2 = reinterpret_cast<const float*>(bp + mp.al2) + (size_t)(t < T ? t : 0) * s.S2;
    const int NQ = K / 4, G = DT_BM / NQ, cq = t % NQ, g = t / NQ;   // column quads, row groups
    for (int64_t i = 0; i < n_my; ++i) {
      const int st = (int)(i % nraw), slot = (int)(i & 1);
      const int64_t b = blockIdx.x + i * gridDim.x;
      const uint32_t stage_off = mp.stage0 + (uint32_t)st * mp.stage_bytes;
      const int32_t* ids_s = reinterpret_cast<const int32_t*>(bp + stage_off + (uint32_t)s.nchK * (uint32_t)s.SR * 128u);
      mbar_wait(d2_full(slot), (uint32_t)((i >> 1) & 1));
      if (lane == 0 && q4 == 0) DT_MARK(5, i);
      tc_fence_after();
      const uint32_t d_t = t_lane + (uint32_t)(slot * s.slot_cols + s.c_d2);
      float sc = s_b3[0];
      for (int n0 = 0; n0 < s.N2p; n0 += 16) {
        float acc[16];
        tmem_ld16(d_t + (uint32_t)n0, acc);
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          const float4 bb = *reinterpret_cast<const float4*>(s_b2 + n0 + j);
          const float4 aa = *reinterpret_cast<const float4*>(al2 + n0 + j);
          const float4 ww = *reinterpret_cast<const float4*>(s_w3 + n0 + j);
          sc = fmaf(dt_prelu(acc[j + 0] + bb.x, aa.x), ww.x, sc);
          sc = fmaf(dt_prelu(acc[j + 1] + bb.y, aa.y), ww.y, sc);
          sc = fmaf(dt_prelu(acc[j + 2] + bb.z, aa.z), ww.z, sc);
          sc = fmaf(dt_prelu(acc[j + 3] + bb.w, aa.w), ww.w, sc);
        }
      }
      tc_fence_before();
      mbar_arrive(d2_free(slot));
      // mask (interaction.py:399-401: where(mask == 0, -2^32 + 1, s)), rows past T do not exist
      const bool live = t < T;
      if (live && ids_s[t] == 0) sc = -4294967295.0f;
      if (!live) sc = -INFINITY;
      float m = sc;
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      if (lane == 0) s_red[q4] = m;
      named_bar_sync(1, 128);
      m = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
      const float e = live ? __expf(sc - m) : 0.f;
      float sum = e;
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      if (lane == 0) s_red[4 + q4] = sum;
      named_bar_sync(1, 128);
      sum = (s_red[4] + s_red[5]) + (s_red[6] + s_red[7]);
      s_P[t] = e / sum;
      named_bar_sync(1, 128);
      // out[c] = sum_t p_t k[t][c]: thread = (four columns 4 cq .., group g of rows t = g, g + G, ...)
      {
        const uint8_t* kt = bp + stage_off + (uint32_t)(cq >> 3) * ((uint32_t)s.SR * 128u);
        const uint32_t pq = (uint32_t)(cq & 7);
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
        int r = g;
        for (; r + G < T; r += 2 * G) {
          const float4 k0 = *reinterpret_cast<const float4*>(kt + (uint32_t)r * 128u + ((pq ^ (uint32_t)(r & 7)) << 4));
          const float4 k1 = *reinterpret_cast<const float4*>(kt + (uint32_t)(r + G) * 128u + ((pq ^ (uint32_t)((r + G) & 7)) << 4));
          const float p0 = s_P[r], p1 = s_P[r + G];
          a0.x = fmaf(p0, k0.x, a0.x); a0.y = fmaf(p0, k0.y, a0.y); a0.z = fmaf(p0, k0.z, a0.z); a0.w = fmaf(p0, k0.w, a0.w);
          a1.x = fmaf(p1, k1.x, a1.x); a1.y = fmaf(p1, k1.y, a1.y); a1.z = fmaf(p1, k1.z, a1.z); a1.w = fmaf(p1, k1.w, a1.w);
        }
        if (r < T) {
          const float4 k0 = *reinterpret_cast<const float4*>(kt + (uint32_t)r * 128u + ((pq ^ (uint32_t)(r & 7)) << 4));
          const float p0 = s_P[r];
          a0.x = fmaf(p0, k0.x, a0.x); a0.y = fmaf(p0, k0.y, a0.y); a0.z = fmaf(p0, k0.z, a0.z); a0.w = fmaf(p0, k0.w, a0.w);
        }
        *reinterpret_cast<float4*>(s_part + (size_t)g * K + 4 * cq) = make_float4(a0.x + a1.x, a0.y + a1.y, a0.z + a1.z, a0.w + a1.w);
      }
      named_bar_sync(1, 128);
      mbar_arrive(raw_free(st));                             // keys no longer needed: the gather may refill the stage
      if (t < K) {
        float o = 0.f;
        for (int gg = 0; gg < G; ++gg) o += s_part[gg * K + t];
        // every position masked: the softmax is uniform over T copies of table row 0 (whose rows were not fetched)
        if (m == -4294967295.0f) o += s_row0[t];
        p.out[b * p.out_ld + t] = o;
      }
      if (lane == 0 && q4 == 0) DT_MARK(6, i);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 12) { tc_fence_after(); tmem_dealloc(tmem_base, 512u); }
}

}  // namespace ctrhot

using namespace ctrhot;

static bool din_layers_ok(const ctrhot_dense_desc* L, int n_layers, int K) {
  if (!L || n_layers != 2) return false;
  if (!L[0].weight || !L[1].weight) return false;
  if (L[0].in_dim != 4 * K || L[1].in_dim != L[0].out_dim) return false;
  for (int i = 0; i < 2; ++i) {
    if (L[i].act == CTRHOT_ACT_PRELU) { if (!L[i].act_param || L[i].alpha_rows < 1) return false; }
    else if (L[i].act != CTRHOT_ACT_RELU && L[i].act != CTRHOT_ACT_NONE) return false;
  }
  return true;
}

extern "C" int64_t ctrhot_din_pack_bytes(int T, int K, int h1, int h2) {
  DinShape s;
  if (!din_shape(T, K, h1, h2, s)) return CTRHOT_ESHAPE;
  return (int64_t)s.total * sizeof(float);
}

extern "C" int ctrhot_din_pack(const ctrhot_dense_desc* h_layers, int n_layers, const float* out_w, const float* out_b,
                               int T, int K, void* packed, void* stream) {
  CTRHOT_RANGE("ctrhot_din_pack");
  if (!h_layers || !out_w || !packed || !aligned16(packed)) return CTRHOT_EARG;
  if (!din_layers_ok(h_layers, n_layers, K)) return CTRHOT_ESHAPE;
  DinShape s;
  if (!din_shape(T, K, h_layers[0].out_dim, h_layers[1].out_dim, s)) return CTRHOT_ESHAPE;
  // ReLU = PReLU with alpha 0 (the blob is zero where no alpha is given); "none" is not representable here
  if (h_layers[0].act == CTRHOT_ACT_NONE || h_layers[1].act == CTRHOT_ACT_NONE) return CTRHOT_ESHAPE;
  const float* a1 = h_layers[0].act == CTRHOT_ACT_PRELU ? h_layers[0].act_param : nullptr;
  const float* a2 = h_layers[1].act == CTRHOT_ACT_PRELU ? h_layers[1].act_param : nullptr;
  din_pack_kernel<<<64, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      h_layers[0].weight, h_layers[0].bias, a1, h_layers[1].weight, h_layers[1].bias, a2, out_w, out_b,
      a1 ? h_layers[0].alpha_rows : 1, a2 ? h_layers[1].alpha_rows : 1, s, static_cast<float*>(packed));
  CTRHOT_LAUNCH_OK("ctrhot_din_pack");
  return CTRHOT_OK;
}

extern "C" int64_t ctrhot_din_tc_workspace_bytes(int64_t B, int h1) {
  if (B < 0 || h1 <= 0) return CTRHOT_EARG;
  return B * (int64_t)dt_round(h1, 16) * (int64_t)sizeof(float);
}

extern "C" int ctrhot_din_attention_gather_tf32(const float* table, int64_t V, int K,
                                                const void* hist_ids, int ids_kind, int64_t hist_ld,
                                                const void* cand_ids, int64_t cand_ld, int64_t B, int T,
                                                const void* packed, int h1, int h2,
                                                float* att_out, int64_t att_ld, float* item_out, int64_t item_ld,
                                                void* workspace, int64_t workspace_bytes, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_din_attention_gather_tf32");
  if (!table || !hist_ids || !cand_ids || !packed || !att_out || !item_out || !workspace || B < 0 || V <= 0) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  DinShape s;
  if (!din_shape(T, K, h1, h2, s)) return CTRHOT_ESHAPE;
  if (att_ld < K || item_ld < K || hist_ld < T) return CTRHOT_EARG;
  if (!aligned16(table) || !aligned16(packed) || !aligned16(item_out) || !aligned16(workspace) || item_ld % 4 != 0) return CTRHOT_ESHAPE;
  if (workspace_bytes < ctrhot_din_tc_workspace_bytes(B, h1)) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float* P = static_cast<const float*>(packed);
  float* U = static_cast<float*>(workspace);
  {
    const int64_t grid = (B + DP_S - 1) / DP_S;
    if (ids_kind == CTRHOT_IDS_I32)
      din_prep_kernel<CTRHOT_IDS_I32><<<(int)grid, 256, 0, st>>>(table, V, K, cand_ids, cand_ld, B, P + s.o_wq, P + s.o_b1, s.N1p,
                                                              item_out, item_ld, U, err_flag);
    else
      din_prep_kernel<CTRHOT_IDS_F32><<<(int)grid, 256, 0, st>>>(table, V, K, cand_ids, cand_ld, B, P + s.o_wq, P + s.o_b1, s.N1p,
                                                              item_out, item_ld, U, err_flag);
    CTRHOT_LAUNCH_OK("din_prep");
  }
  DinParams p{};
  p.table = table; p.V = V; p.hist_ids = hist_ids; p.hist_ld = hist_ld; p.ids_kind = ids_kind;
  p.item = item_out; p.item_ld = item_ld; p.U = U; p.packed = P; p.out = att_out; p.out_ld = att_ld; p.B = B;
  p.err_flag = err_flag; p.s = s;
  DinSmem mp{};
  for (p.nraw = DT_NRAW; p.nraw >= 2; --p.nraw) {
    mp = din_smem(s, p.nraw);
    if (mp.total + 1024u <= 227u * 1024u) break;
  }
  if (p.nraw < 2) return CTRHOT_ESHAPE;
  const size_t smem = mp.total + 1024u;
  const int grid = (int)(B < kNumSMs ? B : kNumSMs);
  if (ids_kind == CTRHOT_IDS_I32) {
    CTRHOT_CUDA_OK(cudaFuncSetAttribute(din_tc_kernel<CTRHOT_IDS_I32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "din_tc attr");
    din_tc_kernel<CTRHOT_IDS_I32><<<grid, DT_THREADS, smem, st>>>(p);
  } else {
    CTRHOT_CUDA_OK(cudaFuncSetAttribute(din_tc_kernel<CTRHOT_IDS_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "din_tc attr");
    din_tc_kernel<CTRHOT_IDS_F32><<<grid, DT_THREADS, smem, st>>>(p);
  }
  CTRHOT_LAUNCH_OK("ctrhot_din_attention_gather_tf32");
  return CTRHOT_OK;
}

#ifdef DT_TRACE
extern "C" int ctrhot_debug_din_trace(long long* h_out) {
  CTRHOT_RANGE("ctrhot_debug_din_trace");
  return cudaMemcpyFromSymbol(h_out, ctrhot::dt_trace, sizeof(long long) * 8 * 128) == cudaSuccess ? 0 : -3;
}
#endif
