// K7: the whole front end of DIN.call (DL/model/din.py:56-90) in ONE launch -- candidate lookup, history lookup, the
// local-activation attention unit (Attention.call, DL/layer/interaction.py:369-406), the other sparse lookups and the dense
// columns, written as the tower's input matrix  emb = [att | item | other ... | dense]  (din.py:84-90's concat).
//
//   a_t   = [q, k_t, q - k_t, q * k_t]                     (interaction.py:381-391; never materialised)
//   h1_t  = PReLU_{alpha1[t,:]}(a_t W1 + b1)               W1:[4K,h1]  (interaction.py:366,393-394)
//   h2_t  = PReLU_{alpha2[t,:]}(h1_t W2 + b2)              W2:[h1,h2]
//   s_t   = h2_t . w3 + b3;  s_t = -2^32+1 where hist id == 0;  p = softmax_T(s);  att = sum_t p_t k_t   (interaction.py:396-405)
//
// The first layer is rearranged by linearity so that its per-(sample, t) GEMM has K-extent K instead of 4K:
//   a_t W1 = q (W1a + W1c)  +  k_t (W1b - W1c)  +  (q * k_t) W1d
//          = u_s            +  k_t Ws,            Ws = (W1b - W1c) + diag(q) W1d        -- a [K, h1] matrix per SAMPLE.
// u_s (fp32, CUDA cores) is made in a prologue together with the candidate / other-feature lookups; Ws is rebuilt for every
// sample by 8 "builder" warps that hold W1b - W1c and W1d in registers (one FMA + one TF32 rounding + a quarter of a 16-byte
// shared-memory store per element); the history rows land in shared memory in exactly the K-major 128-byte-swizzled layout
// tcgen05.mma reads, so layer 1 is an SS-form MMA straight from the gather's stage (no staging pass, no A operand in TMEM):
// 8 + 10 MMA instructions per sample at cfg 4 instead of 16 + 10, and a TMEM slot is 128 columns, so FOUR samples are in
// flight where din_tc_kernel (K6) had two.  This changes rounding only (SURVEY 8(d) cfg 4 allows it).
//
// One persistent CTA per SM, one 128-row MMA tile per sample (rows t >= T: whatever follows the stage, never used), 21 warps:
//   warps 0-3    epilogue 1: thread t = TMEM lane t.  D1 + u -> PReLU -> TF32, back in place (A of layer 2, TS form).
//   warps 4-7    gather: 16-byte cp.async of the T history rows (+ q, u of this sample) into a 4-stage ring, ids two ahead.
//   warps 8-15   builders: Ws(i) into one of two weight buffers.
//   warp  16     MMA issuer (converged warp, elect.sync): L1(i) = keys(i) x Ws(i), L2(i) = H1(i) x W2, ready-first order.
//   warps 17-20  epilogue 2: D2 + b2 -> PReLU -> . w3 -> mask -> softmax over the CTA's lanes -> att = p @ keys (keys re-read
//                from the stage) -> emb[:, 0:K].
// HBM traffic per sample = (T + 1) ids + (live rows + 1) rows of 4K bytes + u + the emb row; tensor work 2 T (K h1 + h1 h2).
#include "common.cuh"
#include "tc05.cuh"

namespace ctrhot {
using namespace tc;

constexpr int DF_BM = 128;
constexpr int DF_THREADS = 24 * 32;          // (ptxas budgets registers for a multiple of 4 warps: 21 warps cost what 24 do)
constexpr int DF_NSUM = 96;                  // threads of the weighted-sum group (warps 21-23)
constexpr int DF_NRAW = 4;                   // key stages (fewer if shared memory is short)
constexpr int DF_NSLOT = 4;                  // TMEM slots [D1/H1 | D2]
constexpr int DF_NBLD = 256;                 // builder threads
constexpr int DF_HDR_BYTES = 2048;           // per stage: ids[128] int32 | q[<=64] at 512 | q of the NEXT sample at 768 | u[<=96] at 1024
constexpr int DF_HDR_Q = 512, DF_HDR_QN = 768, DF_HDR_U = 1024;   // (2048 keeps the stages 1024-aligned)
constexpr int DF_MAX_EXTRA = 8;
constexpr int DF_MAXPASS = 6;                // builder passes: ceil(N1p / (256 / (K/4))) <= 6  (K = 64, N1p = 96)

struct DfShape {
  int T, K, h1, h2;
  int N1p, N2p;            // out dims rounded up to 16 (MMA N)
  int nchK, nch2;          // 32-column chunks: keys / Ws, W2
  int SR;                  // key rows a stage holds: T rounded up to 8
  int npass;               // builder passes
  int o_wq, o_b1, o_wb, o_wd, o_w2, o_al1, o_al2, o_b2, o_w3, o_b3, total;   // packed blob offsets (floats)
  int slot_cols, nslot;
};

__host__ __device__ inline int df_round(int v, int m) { return (v + m - 1) / m * m; }

__host__ __device__ inline bool df_shape(int T, int K, int h1, int h2, DfShape& s) {
  if (T < 1 || T > DF_BM) return false;
  if (!(K == 32 || K == 64)) return false;             // one or two 128-byte swizzle chunks per key row (K6 takes K = 8, 16)
  if (h1 < 8 || h1 > 96 || h2 < 1 || h2 > 64) return false;
  s.T = T; s.K = K; s.h1 = h1; s.h2 = h2;
  s.SR = df_round(T, 8);
  s.N1p = df_round(h1, 16); s.N2p = df_round(h2, 16);
  s.nchK = K / 32; s.nch2 = (s.N1p + 31) / 32;
  const int rows_per_pass = DF_NBLD / (K / 4);
  s.npass = (s.N1p + rows_per_pass - 1) / rows_per_pass;
  if (s.npass > DF_MAXPASS) return false;
  int o = 0;
  s.o_wq = o; o += K * s.N1p;
  s.o_b1 = o; o += s.N1p;
  s.o_wb = o; o += s.npass * DF_NBLD * 4;
  s.o_wd = o; o += s.npass * DF_NBLD * 4;
  s.o_w2 = o; o += s.nch2 * s.N2p * 32;
  s.o_al1 = o; o += T * s.N1p;
  s.o_al2 = o; o += T * s.N2p;
  s.o_b2 = o; o += s.N2p;
  s.o_w3 = o; o += s.N2p;
  s.o_b3 = o; o += 4;
  s.total = o;
  s.slot_cols = s.N1p + s.N2p;
  s.nslot = 512 / s.slot_cols; if (s.nslot > DF_NSLOT) s.nslot = DF_NSLOT;
  return s.nslot >= 2;
}

// alpha rows are read one row per thread with 16-byte loads: instead of padding the row stride, float4 j of row t is stored at
// j ^ sw(t) so that 8 consecutive rows hit 8 different 16-byte bank groups (N / 4 is 4 or 0 mod 8 for N a multiple of 16)
__host__ __device__ inline int df_alpha_sw(int N, int t) { return (((N >> 2) & 7) == 4) ? ((t >> 1) & 3) : (t & 7); }

// ---------------------------------------------------------------- pack (weights -> one device blob)
__global__ void din_fwd_pack_kernel(const float* __restrict__ W1, const float* __restrict__ b1, const float* __restrict__ al1,
                                    const float* __restrict__ W2, const float* __restrict__ b2, const float* __restrict__ al2,
                                    const float* __restrict__ w3, const float* __restrict__ b3, int ar1, int ar2, DfShape s,
                                    float* __restrict__ P) {
  const int K = s.K, h1 = s.h1, h2 = s.h2;
  const int NQ = K / 4, rpp = DF_NBLD / NQ;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < s.total; i += gridDim.x * blockDim.x) {
    float v = 0.f;
    if (i < s.o_b1) {                                          // Wq[c][n] = W1a + W1c
      const int c = i / s.N1p, n = i - c * s.N1p;
      if (n < h1) v = W1[(size_t)c * h1 + n] + W1[(size_t)(2 * K + c) * h1 + n];
    } else if (i < s.o_wb) {
      const int n = i - s.o_b1;
      if (n < h1 && b1) v = b1[n];
    } else if (i < s.o_w2) {                                   // builder constants, [pass][thread][4]: (W1b - W1c) then W1d
      const bool second = i >= s.o_wd;
      const int e = i - (second ? s.o_wd : s.o_wb);
      const int j = e / (DF_NBLD * 4), tb = (e >> 2) % DF_NBLD, q = e & 3;
      const int c = 4 * (tb % NQ) + q, n = tb / NQ + rpp * j;
      if (n < h1) v = second ? W1[(size_t)(3 * K + c) * h1 + n]
                             : W1[(size_t)(K + c) * h1 + n] - W1[(size_t)(2 * K + c) * h1 + n];
    } else if (i < s.o_al1) {                                  // W2 chunks, K-major, 128B swizzle, TF32
      const int e = i - s.o_w2;
      const int ch = e / (s.N2p * 32), r = (e / 32) % s.N2p, w = e % 32;
      const int piece = (w >> 2) ^ (r & 7), kin = ch * 32 + piece * 4 + (w & 3);
      if (r < h2 && kin < h1) v = to_tf32_rna(W2[(size_t)kin * h2 + r]);
    } else if (i < s.o_al2) {
      const int e = i - s.o_al1;
      const int t = e / s.N1p, w = e - t * s.N1p;
      const int n = (((w >> 2) ^ df_alpha_sw(s.N1p, t)) << 2) | (w & 3);       // the column stored at position w of row t
      if (n < h1 && al1) v = al1[(size_t)(t % ar1) * h1 + n];
    } else if (i < s.o_b2) {
      const int e = i - s.o_al2;
      const int t = e / s.N2p, w = e - t * s.N2p;
      const int n = (((w >> 2) ^ df_alpha_sw(s.N2p, t)) << 2) | (w & 3);
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

// ---------------------------------------------------------------- device helpers
__device__ __forceinline__ float4 df_lds128(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void df_sts128(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" :: "r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void df_ld16_nowait(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void df_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
         "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void df_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float df_prelu(float v, float a) { return fmaxf(v, 0.f) + a * fminf(v, 0.f); }
// round-to-nearest (ties away) TF32 of a finite value
__device__ __forceinline__ float df_tf32(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u); }
__device__ __forceinline__ int64_t df_id(const void* p, int kind, int64_t i) {
  return kind == CTRHOT_IDS_I32 ? (int64_t)__ldg(static_cast<const int32_t*>(p) + i)
                                : (int64_t)__float2int_rz(__ldg(static_cast<const float*>(p) + i));
}

struct DfExtra { const float* table; int64_t rows; const void* ids; int64_t ids_ld; int dim; int col; };

struct DfParams {
  const float* table; int64_t V;
  const void* hist_ids; int64_t hist_ld;
  const void* cand_ids; int64_t cand_ld;
  int ids_kind;
  float* U;                                  // workspace [B, N1p]
  const float* packed;
  float* emb; int64_t emb_ld;                // [att | item | extras | dense]
  int64_t B;
  int* err_flag;
  int nraw;
  int n_extra;
  DfExtra extra[DF_MAX_EXTRA];
  const float* dense; int64_t dense_ld; int n_dense; int dense_col;
  DfShape s;
};

// shared-memory map (bytes from the 1024-aligned base); the same function sizes the launch on the host
struct DfSmem { uint32_t ws, ws_bytes, w2, stage0, stage_bytes, al1, al2, small, bars, total; };
__host__ __device__ inline DfSmem df_smem(const DfShape& s, int nraw) {
  DfSmem m;
  uint32_t o = 0;
  m.ws_bytes = (uint32_t)s.nchK * s.N1p * 128u;
  m.ws = o; o += 2u * m.ws_bytes;
  m.w2 = o; o += (uint32_t)s.nch2 * s.N2p * 128u;
  o = (o + 1023u) & ~1023u;
  m.stage_bytes = (uint32_t)s.nchK * (uint32_t)s.SR * 128u + DF_HDR_BYTES;
  m.stage0 = o; o += (uint32_t)nraw * m.stage_bytes;
  m.al1 = o; o += (uint32_t)s.T * s.N1p * 4u;
  m.al2 = o; o += (uint32_t)s.T * s.N2p * 4u;
  m.small = o; o += (uint32_t)(2 * s.N2p + 4 + 2 * DF_BM + 16 + 4 * DF_BM + 64 + 4) * 4u;   // b2 | w3 | b3 | P[2][128] | red[16] | part[512] | row0[64] | aux[4]
  o = (o + 15u) & ~15u;
  m.bars = o; o += 8u * (2 * DF_NRAW + 4 + 4 + 4 * DF_NSLOT) + 16u;
  // the layer-1 MMA reads 128 rows from every key chunk: the rows past SR of the LAST stage's last chunk must still be inside
  // the allocation (their contents are never used)
  const uint32_t reach = m.stage0 + (uint32_t)(nraw - 1) * m.stage_bytes + (uint32_t)(s.nchK - 1) * (uint32_t)s.SR * 128u + DF_BM * 128u;
  m.total = o > reach ? o : reach;
  return m;
}

#ifndef DF_ROUND_H1
#define DF_ROUND_H1 1      // 1: H1 is rounded to nearest TF32 before layer 2 reads it; 0: left to the MMA's truncation
#endif
#ifndef DF_SKIP_PAD
#define DF_SKIP_PAD 0      // 1: padding positions (history id 0) cost no row copy, no epilogue arithmetic, no weighted-sum read
#endif
#ifndef DF_SLEEP_
#define DF_SLEEP_ 0        // ns of back-off between two polls of an mbarrier (0: spin on try_wait)
#endif
__device__ __forceinline__ void df_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) { if (DF_SLEEP_ > 0) __nanosleep(DF_SLEEP_); }
}
#ifdef DF_TRACE
// tuning builds only: clock64 of pipeline events of CTA 0, [event][sample]
__device__ long long df_trace[14][128];
#ifndef DF_TRACE_CTA
#define DF_TRACE_CTA 0
#endif
#define DF_MARK(ev, i) do { if (blockIdx.x == DF_TRACE_CTA && (i) < 128 && (threadIdx.x & 31) == 0) df_trace[ev][i] = clock64(); } while (0)
__device__ unsigned long long df_cta[148][4];          // per CTA: globaltimer (ns) at kernel start, prologue end, last att, exit
__device__ __forceinline__ unsigned long long df_gt() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define DF_CTA(k) do { if (threadIdx.x == 0 && blockIdx.x < 148) df_cta[blockIdx.x][k] = df_gt(); } while (0)
#else
#define DF_CTA(k) do { } while (0)
#define DF_MARK(ev, i) do { } while (0)
#endif

template <int IDKIND>
__global__ void __launch_bounds__(DF_THREADS, 1)
din_fwd_kernel(const __grid_constant__ DfParams p) {
  extern __shared__ uint8_t smem_raw[];
  const DfShape& s = p.s;
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* bp = smem_raw + (base - raw_addr);
  const int nraw = p.nraw, NS = s.nslot;
  const DfSmem mp = df_smem(s, nraw);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t bars = base + mp.bars;
  auto full_raw = [&](int i) { return bars + 8u * i; };
  auto p_full = [&](int i) { return bars + 8u * (DF_NRAW + i); };          // softmax weights of sample i in s_P[i & 1]
  auto p_free = [&](int i) { return bars + 8u * (DF_NRAW + 2 + i); };
  auto raw_free = [&](int i) { return bars + 8u * (DF_NRAW + 4 + i); };    // att of the stage's sample is formed: refill it
  const uint32_t mb = bars + 8u * (2 * DF_NRAW + 4);
  auto ws_full = [&](int i) { return mb + 8u * i; };
  auto ws_free = [&](int i) { return mb + 8u * (2 + i); };
  auto d1_full = [&](int i) { return mb + 8u * (4 + i); };
  auto h1_ready = [&](int i) { return mb + 8u * (4 + DF_NSLOT + i); };
  auto d2_full = [&](int i) { return mb + 8u * (4 + 2 * DF_NSLOT + i); };
  auto d2_free = [&](int i) { return mb + 8u * (4 + 3 * DF_NSLOT + i); };
  const uint32_t slot_addr = mb + 8u * (4 + 4 * DF_NSLOT);
  volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(bp + (slot_addr - base));
  float* f_small = reinterpret_cast<float*>(bp + mp.small);
  float* s_b2 = f_small; float* s_w3 = s_b2 + s.N2p; float* s_b3 = s_w3 + s.N2p;
  float* s_P = s_b3 + 4; float* s_red = s_P + 2 * DF_BM; float* s_part = s_red + 16;   // s_P: two samples
  float* s_row0 = s_part + 4 * DF_BM;
  float* s_aux = s_row0 + 64;                   // [2][2]: 1 / sum, all-masked flag            // table row 0: the key of every padding position (history id 0)
  const int T = s.T, K = s.K;
  const uint32_t keys_bytes = (uint32_t)s.nchK * (uint32_t)s.SR * 128u;   // header offset inside a stage

  if (tid == 0) {
    for (int i = 0; i < DF_NRAW; ++i) mbar_init(full_raw(i), 256);
    for (int i = 0; i < 2; ++i) { mbar_init(p_full(i), 128); mbar_init(p_free(i), DF_NSUM); }
    for (int i = 0; i < DF_NRAW; ++i) mbar_init(raw_free(i), DF_NSUM);
    for (int i = 0; i < 2; ++i) { mbar_init(ws_full(i), DF_NBLD); mbar_init(ws_free(i), 1); }
    for (int i = 0; i < DF_NSLOT; ++i) {
      mbar_init(d1_full(i), 1); mbar_init(h1_ready(i), 128); mbar_init(d2_full(i), 1); mbar_init(d2_free(i), 128);
    }
    mbar_fence_init();
  }
  // resident weights: W2 is already in its shared-memory layout, the alpha rows in their swizzled order
  {
    const float4* src = reinterpret_cast<const float4*>(p.packed + s.o_w2);
    float4* dst = reinterpret_cast<float4*>(bp + mp.w2);
    const int n4 = s.nch2 * s.N2p * 8;
    for (int i = tid; i < n4; i += DF_THREADS) dst[i] = __ldg(src + i);
    float4* a1 = reinterpret_cast<float4*>(bp + mp.al1);
    const float4* g1 = reinterpret_cast<const float4*>(p.packed + s.o_al1);
    for (int i = tid; i < s.T * s.N1p / 4; i += DF_THREADS) a1[i] = __ldg(g1 + i);
    float4* a2 = reinterpret_cast<float4*>(bp + mp.al2);
    const float4* g2 = reinterpret_cast<const float4*>(p.packed + s.o_al2);
    for (int i = tid; i < s.T * s.N2p / 4; i += DF_THREADS) a2[i] = __ldg(g2 + i);
    for (int i = tid; i < 2 * s.N2p + 4; i += DF_THREADS) f_small[i] = __ldg(p.packed + s.o_b2 + i);
    for (int i = tid; i < K; i += DF_THREADS) s_row0[i] = (p.V > 0) ? __ldg(p.table + i) : 0.f;
  }
  fence_proxy_async_smem();
  if (warp == 16) { tmem_alloc(slot_addr, 512u); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot_ptr;
  const int64_t n_my = (p.B > blockIdx.x) ? (p.B - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (tid == 0) DF_MARK(7, 0);
  DF_CTA(0);
  // ================================================================ prologue (all warps): per-sample lookups and u
  // For this CTA's samples, PB at a time: ids -> candidate rows (to emb[:, K:2K] and to shared memory), other-feature rows and
  // dense columns (to emb) -> u = q (W1a + W1c) + b1 in fp32 (to the workspace; the gather warps fetch it back per sample).
  // The key stages and the Ws buffers are not in use yet: they hold the candidate rows, the ids and Wq.
  // (Everything in this kernel is written as ROLLED loops on purpose: the first version, unrolled the usual way, was 132 KB of
  // SASS with a hot set of ~64 KB, and ran at 77 % instruction-cache hit rate with the GPC-level instruction cache at 81 % of
  // its request peak -- instruction fetch, not any pipe, bounded it.  profiles/r02_din_fwd.md.)
  {
    float* sq = reinterpret_cast<float*>(bp + mp.stage0);
    const int QS = K + 4, nf = 1 + p.n_extra;
    int PB = (int)(((uint32_t)nraw * mp.stage_bytes) / (uint32_t)(QS * 4 + nf * 4));
    if (PB > DF_BM) PB = DF_BM;
    if (PB >= 4) PB &= ~3;                                  // the u tiles take four samples at a time
    int32_t* pid = reinterpret_cast<int32_t*>(sq + (size_t)PB * QS);
    const int N1p = s.N1p;
    float* sWq = reinterpret_cast<float*>(bp + mp.ws);
    for (int i = tid; i < K * N1p / 4; i += DF_THREADS)
      reinterpret_cast<float4*>(sWq)[i] = __ldg(reinterpret_cast<const float4*>(p.packed + s.o_wq) + i);
#pragma unroll 1
    for (int64_t i0 = 0; i0 < n_my; i0 += PB) {
      const int ns = (int)((n_my - i0 < PB) ? (n_my - i0) : PB);
#pragma unroll 1
      for (int e = tid; e < ns * nf; e += DF_THREADS) {
        const int f = e / ns, si = e - f * ns;
        const int64_t b = blockIdx.x + (i0 + si) * (int64_t)gridDim.x;
        int64_t id, rows;
        if (f == 0) { id = IdLoad<IDKIND>::at(p.cand_ids, b * p.cand_ld); rows = p.V; }
        else { id = IdLoad<IDKIND>::at(p.extra[f - 1].ids, b * p.extra[f - 1].ids_ld); rows = p.extra[f - 1].rows; }
        const bool ok = id >= 0 && id < rows;
        if (!ok && p.err_flag) *p.err_flag = 1;
        pid[f * PB + si] = ok ? (int32_t)id : -1;
      }
      __syncthreads();
      {                                                       // candidate rows, 16 bytes per element, four loads in flight
        const int q4n = K / 4, n_el = ns * q4n;
#pragma unroll 1
        for (int e0 = 0; e0 < n_el; e0 += DF_THREADS * 4) {
          float4 v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int e = e0 + tid + DF_THREADS * u;
            v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (e < n_el) {
              const int32_t id = pid[e / q4n];
              if (id >= 0) v[u] = __ldg(reinterpret_cast<const float4*>(p.table + (int64_t)id * K) + (e % q4n));
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int e = e0 + tid + DF_THREADS * u;
            if (e < n_el) {
              const int si = e / q4n, c4 = e - si * q4n;
              const int64_t b = blockIdx.x + (i0 + si) * (int64_t)gridDim.x;
              *reinterpret_cast<float4*>(sq + (size_t)si * QS + 4 * c4) = v[u];
              *reinterpret_cast<float4*>(p.emb + b * p.emb_ld + K + 4 * c4) = v[u];
            }
          }
        }
      }
#pragma unroll 1
      for (int f = 0; f < p.n_extra; ++f) {                   // other sparse features: Embedding rows straight into emb
        const float* tab = p.extra[f].table;
        const int dim = p.extra[f].dim, col = p.extra[f].col, n_el = ns * dim;
        const int32_t* idf = pid + (f + 1) * PB;
#pragma unroll 1
        for (int e0 = 0; e0 < n_el; e0 += DF_THREADS * 4) {
          float v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int e = e0 + tid + DF_THREADS * u;
            v[u] = 0.f;
            if (e < n_el) {
              const int si = e / dim;
              const int32_t id = idf[si];
              if (id >= 0) v[u] = ldg_nc_f1(tab + (int64_t)id * dim + (e - si * dim));
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int e = e0 + tid + DF_THREADS * u;
            if (e < n_el) {
              const int si = e / dim;
              const int64_t b = blockIdx.x + (i0 + si) * (int64_t)gridDim.x;
              p.emb[b * p.emb_ld + col + (e - si * dim)] = v[u];
            }
          }
        }
      }
#pragma unroll 1
      for (int e = tid; e < ns * p.n_dense; e += DF_THREADS) { // dense columns
        const int si = e / p.n_dense, c = e - si * p.n_dense;
        const int64_t b = blockIdx.x + (i0 + si) * (int64_t)gridDim.x;
        p.emb[b * p.emb_ld + p.dense_col + c] = __ldg(p.dense + b * p.dense_ld + c);
      }
      __syncthreads();
      // u[s][n] = b1[n] + sum_c q_s[c] Wq[c][n]: a 4-sample x 4-column tile per thread (8 shared-memory loads per 64 FMA; one
      // (sample, column) per thread made this loop shared-memory-bound: 22 K wavefronts per batch)
      {
        const int n4n = N1p / 4, n_tiles = ((ns + 3) / 4) * n4n;
#pragma unroll 1
        for (int e = tid; e < n_tiles; e += DF_THREADS) {
          const int s4 = e / n4n, n0 = 4 * (e - s4 * n4n);
          const float4 bias = __ldg(reinterpret_cast<const float4*>(p.packed + s.o_b1 + n0));
          float4 acc[4];
#pragma unroll
          for (int r = 0; r < 4; ++r) acc[r] = bias;
          const float* q0 = sq + (size_t)(4 * s4) * QS;      // rows past ns hold stale (finite or not) data: never stored
#pragma unroll 1
          for (int c0 = 0; c0 < K; c0 += 4) {
            float4 w[4], q[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) w[u] = *reinterpret_cast<const float4*>(sWq + (c0 + u) * N1p + n0);
#pragma unroll
            for (int r = 0; r < 4; ++r) q[r] = *reinterpret_cast<const float4*>(q0 + (size_t)r * QS + c0);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              acc[r].x = fmaf(q[r].x, w[0].x, acc[r].x); acc[r].y = fmaf(q[r].x, w[0].y, acc[r].y);
              acc[r].z = fmaf(q[r].x, w[0].z, acc[r].z); acc[r].w = fmaf(q[r].x, w[0].w, acc[r].w);
              acc[r].x = fmaf(q[r].y, w[1].x, acc[r].x); acc[r].y = fmaf(q[r].y, w[1].y, acc[r].y);
              acc[r].z = fmaf(q[r].y, w[1].z, acc[r].z); acc[r].w = fmaf(q[r].y, w[1].w, acc[r].w);
              acc[r].x = fmaf(q[r].z, w[2].x, acc[r].x); acc[r].y = fmaf(q[r].z, w[2].y, acc[r].y);
              acc[r].z = fmaf(q[r].z, w[2].z, acc[r].z); acc[r].w = fmaf(q[r].z, w[2].w, acc[r].w);
              acc[r].x = fmaf(q[r].w, w[3].x, acc[r].x); acc[r].y = fmaf(q[r].w, w[3].y, acc[r].y);
              acc[r].z = fmaf(q[r].w, w[3].z, acc[r].z); acc[r].w = fmaf(q[r].w, w[3].w, acc[r].w);
            }
          }
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const int si = 4 * s4 + r;
            if (si < ns) {
              const int64_t b = blockIdx.x + (i0 + si) * (int64_t)gridDim.x;
              *reinterpret_cast<float4*>(p.U + b * N1p + n0) = acc[r];
            }
          }
        }
      }
      __syncthreads();
    }
  }
  // Rows of padding positions (history id 0) are never fetched and never used (softmax weight exactly 0; an all-padding history
  // is answered with table row 0 directly): what the stage holds there must only be finite -- zeros now, stale keys later.
  for (uint32_t i = tid; i < (uint32_t)nraw * mp.stage_bytes / 16u; i += DF_THREADS)
    df_sts128(base + mp.stage0 + 16u * i, 0.f, 0.f, 0.f, 0.f);
  fence_proxy_async_smem();
  __syncthreads();
  if (tid == 0) DF_MARK(8, 0);
  DF_CTA(1);

  if (warp >= 4 && warp < 8) {
    // ================================================================ gather
    // Thread tg fetches the history id of row tg (two samples ahead); the 128 ids go through the stage header (the mask needs
    // them there anyway); then PPR adjacent threads copy one key row together (coalesced 16-byte pieces) into the swizzled
    // K-major layout the layer-1 MMA reads.  The stage's LAST reader is this warp group itself: before it refills a stage it
    // forms att(j) = (1 / sum) sum_t P_t k_t of the sample j that holds it (no "stage free" barrier).
    const int tg = tid - 128;
    const int PPR = K / 4;                                   // 16-byte pieces per key row (8 or 16)
    const int c = tg % PPR, rsub = tg / PPR, rpp = DF_BM / PPR;  // this thread: piece c of rows rsub + rpp * j, j < PPR
    const uint32_t rowbytes = (uint32_t)K * 4u;
    const uint32_t chunk = (uint32_t)(c >> 3) * ((uint32_t)s.SR * 128u), pc = (uint32_t)(c & 7);
    auto fetch = [&](int64_t i) -> int32_t {
      if (i >= n_my || tg >= T) return 0;
      const int64_t b = blockIdx.x + i * gridDim.x;
      return __ldg(static_cast<const int32_t*>(p.hist_ids) + b * p.hist_ld + tg);
    };
    int st = 0; uint32_t sph = 0;
    int32_t ra = fetch(0), rb = fetch(1);
#pragma unroll 1
    for (int64_t i = 0; i < n_my; ++i) {
      const int32_t rc = fetch(i + 2);
      const int64_t b = blockIdx.x + i * gridDim.x;
      df_wait(raw_free(st), sph ^ 1u);                       // warps 21-23 have formed att of the sample that held this stage
      const uint32_t stage = base + mp.stage0 + (uint32_t)st * mp.stage_bytes;
      const uint32_t hdr = stage + keys_bytes;
      int32_t* ids_s = reinterpret_cast<int32_t*>(bp + (hdr - base));
      {
        const int32_t id = (IDKIND == CTRHOT_IDS_I32) ? ra : __float2int_rz(__int_as_float(ra));
        const bool ok = tg >= T || (id >= 0 && (int64_t)id < p.V);
        if (!ok && p.err_flag) *p.err_flag = 1;
        ids_s[tg] = (tg < T) ? (ok ? id : -1) : 0;           // -1: out of range -> zero row, not masked
      }
      named_bar_sync(2, 128);
      const char* src0 = reinterpret_cast<const char*>(p.table) + (uint32_t)c * 16u;
      // (shared-memory loads take ~250 cycles in this kernel: all of a thread's ids are fetched before the first copy is issued)
#pragma unroll 8
      for (int r = rsub; r < s.SR; r += rpp) {
        const int32_t id = ids_s[r];
        // id 0 = padding (masked: softmax weight exactly 0 unless the WHOLE history is padding): its key is table row 0,
        // which is added from smem in that one case; the stage row is zero-filled
        const bool ok = r < T && id > 0;
        // (skipped rows keep what they held: zeros from the start of the launch or an earlier sample's keys -- finite, and
        // multiplied by a softmax weight of exactly 0)
        if (!DF_SKIP_PAD || id != 0)
          cp_async16(stage + chunk + (uint32_t)r * 128u + ((pc ^ (uint32_t)(r & 7)) << 4),
                     src0 + (ok ? (uint64_t)(uint32_t)id * rowbytes : 0ull), ok ? 16u : 0u);
      }
      // q of the NEXT sample travels with this one: its layer-1 weights are built while its own rows are still in flight
      if (tg < PPR) {
        if (i == 0) cp_async16(hdr + DF_HDR_Q + tg * 16u, p.emb + b * p.emb_ld + K + tg * 4, 16u);
        if (i + 1 < n_my) cp_async16(hdr + DF_HDR_QN + tg * 16u, p.emb + (b + gridDim.x) * p.emb_ld + K + tg * 4, 16u);
      } else if (tg < PPR + s.N1p / 4) cp_async16(hdr + DF_HDR_U + (tg - PPR) * 16u, p.U + b * s.N1p + (tg - PPR) * 4, 16u);
      cp_async_mbar_arrive_noinc(full_raw(st));
      mbar_arrive(full_raw(st));                             // releases the ids written with plain stores
      if (tg == 0) DF_MARK(0, i);
      if (++st == nraw) { st = 0; sph ^= 1u; }
      ra = rb; rb = rc;
    }
  } else if (warp >= 8 && warp < 16) {
    // ================================================================ builders: Ws(i) = tf32((W1b - W1c) + diag(q_i) W1d)
    // thread = (k quad cq, rows n = nrow0 + rpp * j): its 4 * npass constants of each matrix stay in registers for the launch
    const int tb = tid - 256;
    const int NQ = K / 4, cq = tb % NQ, nrow0 = tb / NQ, rpp = DF_NBLD / NQ;
    float4 wb[DF_MAXPASS], wd[DF_MAXPASS];
#pragma unroll
    for (int j = 0; j < DF_MAXPASS; ++j) {
      wb[j] = make_float4(0.f, 0.f, 0.f, 0.f); wd[j] = wb[j];
      if (j < s.npass) {
        wb[j] = __ldg(reinterpret_cast<const float4*>(p.packed + s.o_wb) + j * DF_NBLD + tb);
        wd[j] = __ldg(reinterpret_cast<const float4*>(p.packed + s.o_wd) + j * DF_NBLD + tb);
      }
    }
    const uint32_t dst0 = base + mp.ws + (uint32_t)(cq >> 3) * ((uint32_t)s.N1p * 128u);
    const uint32_t piece = (uint32_t)(cq & 7);
    // Ws(i + 1) is built from the copy of q(i + 1) that arrived with sample i: one sample ahead of the rows it multiplies
    auto build = [&](int64_t i, float4 q) {
      const int wbuf = (int)(i & 1);
      df_wait(ws_free(wbuf), (uint32_t)((i >> 1) & 1) ^ 1u);
      const uint32_t dst = dst0 + (uint32_t)wbuf * mp.ws_bytes;
#pragma unroll
      for (int j = 0; j < DF_MAXPASS; ++j) {
        const int n = nrow0 + rpp * j;
        if (j < s.npass && n < s.N1p)
          df_sts128(dst + (uint32_t)n * 128u + ((piece ^ (uint32_t)(n & 7)) << 4),
                    df_tf32(fmaf(q.x, wd[j].x, wb[j].x)), df_tf32(fmaf(q.y, wd[j].y, wb[j].y)),
                    df_tf32(fmaf(q.z, wd[j].z, wb[j].z)), df_tf32(fmaf(q.w, wd[j].w, wb[j].w)));
      }
      fence_proxy_async_smem();                              // generic-proxy writes -> the MMA's (async) proxy
      mbar_arrive(ws_full(wbuf));
      if (tb == 0) DF_MARK(2, i);
    };
    int st = 0; uint32_t sph = 0;
#pragma unroll 1
    for (int64_t i = 0; i < n_my; ++i) {
      const uint32_t hdr = base + mp.stage0 + (uint32_t)st * mp.stage_bytes + keys_bytes;
      df_wait(full_raw(st), sph);
      if (tb == 0) DF_MARK(1, i);
      if (i == 0) build(0, df_lds128(hdr + DF_HDR_Q + (uint32_t)cq * 16u));
      if (i + 1 < n_my) build(i + 1, df_lds128(hdr + DF_HDR_QN + (uint32_t)cq * 16u));
      if (++st == nraw) { st = 0; sph ^= 1u; }
    }
  } else if (warp == 16) {
    // ================================================================ MMA issuer: a converged warp, one elected lane issues
    const uint32_t idesc1 = make_idesc_tf32(DF_BM, (uint32_t)s.N1p), idesc2 = make_idesc_tf32(DF_BM, (uint32_t)s.N2p);
    const uint32_t w2 = base + mp.w2;
    // Whatever is ready goes first, finishing samples before starting new ones: L2(i2) needs H1(i2) and a free D2 slot;
    // L1(i1) needs Ws(i1) and keys(i1), and must not overwrite H1(i1 - NS) before L2(i1 - NS) was issued.
    int64_t i1 = 0, i2 = 0;
    int slot1 = 0, slot2 = 0, st1 = 0; uint32_t ph2 = 0, sph1 = 0;
#pragma unroll 1
    while (i2 < n_my) {
      bool did = false;
      if (i2 < i1) {
        bool r = mbar_test_wait(h1_ready(slot2), ph2) && mbar_test_wait(d2_free(slot2), ph2 ^ 1u);
        r = __all_sync(0xffffffffu, r);
        if (r) {
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_t = tmem_base + (uint32_t)(slot2 * s.slot_cols);
            const uint32_t d_t = a_t + (uint32_t)s.N1p;
#pragma unroll 1
            for (int kk = 0; kk < s.N1p / 8; ++kk) {
              const uint64_t db = make_desc_sw128(w2 + (uint32_t)(kk >> 2) * (uint32_t)s.N2p * 128u) + (uint64_t)(2 * (kk & 3));
              mma_tf32_ts(d_t, a_t + (uint32_t)(8 * kk), db, idesc2, (uint32_t)(kk != 0));
            }
            mma_commit(d2_full(slot2));
          }
          __syncwarp();
          ++i2; did = true;
          if (++slot2 == NS) { slot2 = 0; ph2 ^= 1u; }
        }
      }
      if (!did && i1 < n_my && i1 - i2 < NS) {
        const int wbuf = (int)(i1 & 1);
        bool r = mbar_test_wait(ws_full(wbuf), (uint32_t)((i1 >> 1) & 1)) && mbar_test_wait(full_raw(st1), sph1);
        r = __all_sync(0xffffffffu, r);
        if (r) {
          fence_proxy_async_smem();
          tc_fence_after();
          if (elect_one()) {
            const uint32_t keys = base + mp.stage0 + (uint32_t)st1 * mp.stage_bytes;
            const uint32_t ws = base + mp.ws + (uint32_t)wbuf * mp.ws_bytes;
            const uint32_t d_t = tmem_base + (uint32_t)(slot1 * s.slot_cols);
#pragma unroll 1
            for (int kk = 0; kk < K / 8; ++kk) {
              const uint64_t da = make_desc_sw128(keys + (uint32_t)(kk >> 2) * (uint32_t)s.SR * 128u) + (uint64_t)(2 * (kk & 3));
              const uint64_t db = make_desc_sw128(ws + (uint32_t)(kk >> 2) * (uint32_t)s.N1p * 128u) + (uint64_t)(2 * (kk & 3));
              mma_tf32_ss(d_t, da, db, idesc1, (uint32_t)(kk != 0));
            }
            mma_commit(ws_free(wbuf));
            mma_commit(d1_full(slot1));
          }
          __syncwarp();
          ++i1;
          if (++slot1 == NS) slot1 = 0;
          if (++st1 == nraw) { st1 = 0; sph1 ^= 1u; }
        }
      }
    }
  } else if (warp < 4) {
    // ================================================================ epilogue 1 (warps 0-3): thread t = TMEM lane t
    // 16 columns per step of a rolled loop; the next step's TMEM load is in flight while this one is processed in place.
    const int t = tid;
    const uint32_t t_lane = tmem_base + ((uint32_t)(warp * 32) << 16);
    const int tr = t < T ? t : 0;
    const float* al1 = reinterpret_cast<const float*>(bp + mp.al1) + (size_t)tr * s.N1p;
    const int x1 = df_alpha_sw(s.N1p, tr);
    const int N1p = s.N1p, h1 = s.h1;
    int st = 0, slot = 0; uint32_t ph = 0;
#pragma unroll 1
    for (int64_t i = 0; i < n_my; ++i) {
      const float* u = reinterpret_cast<const float*>(bp + mp.stage0 + (uint32_t)st * mp.stage_bytes + keys_bytes + DF_HDR_U);
      df_wait(d1_full(slot), ph);
      if (t == 0) DF_MARK(3, i);
      // a padding position's row of H1 feeds a score that the mask replaces: leave whatever layer 1 produced there (finite)
      const bool dead = DF_SKIP_PAD && (t >= T || reinterpret_cast<const int32_t*>(u - (DF_HDR_U / 4))[t] == 0);
      tc_fence_after();
      const uint32_t d_t = t_lane + (uint32_t)(slot * s.slot_cols);
      uint32_t a[16], nx[16];
      df_ld16_nowait(d_t, nx);
#pragma unroll 1
      for (int n0 = 0; n0 < N1p; n0 += 16) {
        df_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) a[j] = nx[j];
        if (n0 + 16 < N1p) df_ld16_nowait(d_t + (uint32_t)(n0 + 16), nx);
        if (!dead) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          const float4 uu = *reinterpret_cast<const float4*>(u + n0 + j);
          const float4 aa = *reinterpret_cast<const float4*>(al1 + ((((n0 + j) >> 2) ^ x1) << 2));
          a[j + 0] = __float_as_uint(DF_ROUND_H1 ? df_tf32(df_prelu(__uint_as_float(a[j + 0]) + uu.x, aa.x)) : df_prelu(__uint_as_float(a[j + 0]) + uu.x, aa.x));
          a[j + 1] = __float_as_uint(DF_ROUND_H1 ? df_tf32(df_prelu(__uint_as_float(a[j + 1]) + uu.y, aa.y)) : df_prelu(__uint_as_float(a[j + 1]) + uu.y, aa.y));
          a[j + 2] = __float_as_uint(DF_ROUND_H1 ? df_tf32(df_prelu(__uint_as_float(a[j + 2]) + uu.z, aa.z)) : df_prelu(__uint_as_float(a[j + 2]) + uu.z, aa.z));
          a[j + 3] = __float_as_uint(DF_ROUND_H1 ? df_tf32(df_prelu(__uint_as_float(a[j + 3]) + uu.w, aa.w)) : df_prelu(__uint_as_float(a[j + 3]) + uu.w, aa.w));
        }
        }
        if (n0 + 16 > h1) {                                   // zero the padding columns (they are layer 2's K padding)
#pragma unroll
          for (int j = 0; j < 16; ++j) if (n0 + j >= h1) a[j] = 0u;
        }
        df_st16(d_t + (uint32_t)n0, a);
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(h1_ready(slot));
      if (t == 0) DF_MARK(4, i);
      if (++st == nraw) st = 0;
      if (++slot == NS) { slot = 0; ph ^= 1u; }
    }
  } else if (warp < 21) {
    // ================================================================ epilogue 2 (warps 17-20): scores and softmax
    // -> s_P[i & 1][t] = exp(s_t - max) (not normalised), s_aux[i & 1] = {1 / sum, "every position is masked"}; the weighted sum
    // over the keys is the gather warps' last act on the stage before they refill it.
    const int q4 = warp & 3;
    const int t = q4 * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
    const int tr = t < T ? t : 0;
    const float* al2 = reinterpret_cast<const float*>(bp + mp.al2) + (size_t)tr * s.N2p;
    const int x2 = df_alpha_sw(s.N2p, tr);
    const int N2p = s.N2p;
    const bool live = t < T;
    int st = 0, slot = 0; uint32_t ph = 0;
#pragma unroll 1
    for (int64_t i = 0; i < n_my; ++i) {
      const int pb = (int)(i & 1);
      const int32_t* ids_s = reinterpret_cast<const int32_t*>(bp + mp.stage0 + (uint32_t)st * mp.stage_bytes + keys_bytes);
      df_wait(d2_full(slot), ph);
      if (lane == 0 && q4 == 0) DF_MARK(5, i);
      tc_fence_after();
      const uint32_t d_t = t_lane + (uint32_t)(slot * s.slot_cols + s.N1p);
      float sc = s_b3[0];
      const bool dead = DF_SKIP_PAD && (!live || ids_s[t] == 0);
      uint32_t v[16], nx[16];
      df_ld16_nowait(d_t, nx);
#pragma unroll 1
      for (int n0 = 0; n0 < N2p; n0 += 16) {
        df_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = nx[j];
        if (n0 + 16 < N2p) df_ld16_nowait(d_t + (uint32_t)(n0 + 16), nx);
        if (!dead)
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          const int n = n0 + j;
          const float4 bb = *reinterpret_cast<const float4*>(s_b2 + n);
          const float4 aa = *reinterpret_cast<const float4*>(al2 + (((n >> 2) ^ x2) << 2));
          const float4 ww = *reinterpret_cast<const float4*>(s_w3 + n);
          sc = fmaf(df_prelu(__uint_as_float(v[j + 0]) + bb.x, aa.x), ww.x, sc);
          sc = fmaf(df_prelu(__uint_as_float(v[j + 1]) + bb.y, aa.y), ww.y, sc);
          sc = fmaf(df_prelu(__uint_as_float(v[j + 2]) + bb.z, aa.z), ww.z, sc);
          sc = fmaf(df_prelu(__uint_as_float(v[j + 3]) + bb.w, aa.w), ww.w, sc);
        }
      }
      tc_fence_before();
      mbar_arrive(d2_free(slot));
      // mask (interaction.py:399-401: where(mask == 0, -2^32 + 1, s)), rows past T do not exist
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
      df_wait(p_free(pb), (uint32_t)((i >> 1) & 1) ^ 1u);    // the gather warps are done with P(i - 2)
      if (lane == 0) s_red[4 + q4] = sum;
      s_P[pb * DF_BM + t] = e;
      named_bar_sync(1, 128);
      if (t == 0) {
        s_aux[2 * pb] = 1.f / ((s_red[4] + s_red[5]) + (s_red[6] + s_red[7]));
        s_aux[2 * pb + 1] = (m == -4294967295.0f) ? 1.f : 0.f;   // every position masked: uniform over T copies of table row 0
      }
      // s_red[0..3] is next written after this barrier, s_red[4..7] after the next one, and thread 0 takes part in both:
      // nobody overwrites what it is reading here
      mbar_arrive(p_full(pb));
      if (lane == 0 && q4 == 0) DF_MARK(6, i);
      if (++st == nraw) st = 0;
      if (++slot == NS) { slot = 0; ph ^= 1u; }
    }
  } else {
    // ================================================================ weighted sum (warps 21-23): att(j) = (1 / sum) sum_t P_t k_t
    // The stage's last reader: thread = (four columns 4 c .., rows rsub, rsub + G, ...); when it is done the gather may refill.
    const int tq = tid - 21 * 32;
    const int NQ = K / 4, c = tq % NQ, rsub = tq / NQ, G = DF_NSUM / NQ;
    const uint32_t pc = (uint32_t)(c & 7);
    const uint8_t* ksub = bp + mp.stage0 + (uint32_t)(c >> 3) * ((uint32_t)s.SR * 128u);
    int stj = 0;
#pragma unroll 1
    for (int64_t j = 0; j < n_my; ++j) {
      const int pb = (int)(j & 1);
      const int64_t b = blockIdx.x + j * gridDim.x;
      df_wait(p_full(pb), (uint32_t)((j >> 1) & 1));
      const float* P = s_P + pb * DF_BM;
      const uint8_t* kt = ksub + (uint32_t)stj * mp.stage_bytes;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 6
      for (int r = rsub; r < T; r += G) {
        const float p0 = P[r];
        if (!DF_SKIP_PAD || p0 != 0.f) {
          const float4 k0 = *reinterpret_cast<const float4*>(kt + (uint32_t)r * 128u + ((pc ^ (uint32_t)(r & 7)) << 4));
          a.x = fmaf(p0, k0.x, a.x); a.y = fmaf(p0, k0.y, a.y); a.z = fmaf(p0, k0.z, a.z); a.w = fmaf(p0, k0.w, a.w);
        }
      }
      const float inv = s_aux[2 * pb], all_masked = s_aux[2 * pb + 1];
      named_bar_sync(3, DF_NSUM);                            // the previous sample's partial sums have been read
      *reinterpret_cast<float4*>(s_part + (size_t)rsub * K + 4 * c) = a;
      named_bar_sync(3, DF_NSUM);                            // partial sums complete; every thread is done with the stage and P
      mbar_arrive(p_free(pb));
      mbar_arrive(raw_free(stj));
      if (tq < K) {
        float o = 0.f;
#pragma unroll 2
        for (int gg = 0; gg < G; ++gg) o += s_part[gg * K + tq];
        // every position masked: the softmax is uniform over T copies of table row 0 (interaction.py:399-404)
        p.emb[b * p.emb_ld + tq] = (all_masked != 0.f) ? s_row0[tq] : o * inv;
      }
      if (tq == 0) DF_MARK(9, j);
      if (++stj == nraw) stj = 0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) DF_MARK(13, 0);
  DF_CTA(3);
  if (warp == 16) { tc_fence_after(); tmem_dealloc(tmem_base, 512u); }
}

}  // namespace ctrhot

using namespace ctrhot;

static bool df_layers_ok(const ctrhot_dense_desc* L, int n_layers, int K) {
  if (!L || n_layers != 2) return false;
  if (!L[0].weight || !L[1].weight) return false;
  if (L[0].in_dim != 4 * K || L[1].in_dim != L[0].out_dim) return false;
  for (int i = 0; i < 2; ++i) {
    if (L[i].act == CTRHOT_ACT_PRELU) { if (!L[i].act_param || L[i].alpha_rows < 1) return false; }
    else if (L[i].act != CTRHOT_ACT_RELU) return false;       // ReLU = PReLU with alpha 0; "none" is not representable here
  }
  return true;
}

extern "C" int64_t ctrhot_din_fwd_pack_bytes(int T, int K, int h1, int h2) {
  DfShape s;
  if (!df_shape(T, K, h1, h2, s)) return CTRHOT_ESHAPE;
  for (int nraw = DF_NRAW; nraw >= 2; --nraw)
    if (df_smem(s, nraw).total + 1024u <= 227u * 1024u) return (int64_t)s.total * sizeof(float);
  return CTRHOT_ESHAPE;
}

extern "C" int ctrhot_din_fwd_pack(const ctrhot_dense_desc* h_layers, int n_layers, const float* out_w, const float* out_b,
                                   int T, int K, void* packed, void* stream) {
  CTRHOT_RANGE("ctrhot_din_fwd_pack");
  if (!h_layers || !out_w || !packed || !aligned16(packed)) return CTRHOT_EARG;
  if (!df_layers_ok(h_layers, n_layers, K)) return CTRHOT_ESHAPE;
  DfShape s;
  if (!df_shape(T, K, h_layers[0].out_dim, h_layers[1].out_dim, s)) return CTRHOT_ESHAPE;
  const float* a1 = h_layers[0].act == CTRHOT_ACT_PRELU ? h_layers[0].act_param : nullptr;
  const float* a2 = h_layers[1].act == CTRHOT_ACT_PRELU ? h_layers[1].act_param : nullptr;
  din_fwd_pack_kernel<<<64, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      h_layers[0].weight, h_layers[0].bias, a1, h_layers[1].weight, h_layers[1].bias, a2, out_w, out_b,
      a1 ? h_layers[0].alpha_rows : 1, a2 ? h_layers[1].alpha_rows : 1, s, static_cast<float*>(packed));
  CTRHOT_LAUNCH_OK("ctrhot_din_fwd_pack");
  return CTRHOT_OK;
}

extern "C" int64_t ctrhot_din_fwd_workspace_bytes(int64_t B, int h1) {
  if (B < 0 || h1 <= 0) return CTRHOT_EARG;
  return B * (int64_t)df_round(h1, 16) * (int64_t)sizeof(float);
}

// h_extras: n_extras x 6 int64 on the HOST: {table (device pointer), rows, dim, ids (device pointer, same ids_kind as the
// history, one id per sample), ids_ld, first column in emb}
extern "C" int ctrhot_din_fwd_tf32(const float* table, int64_t V, int K, const void* hist_ids, int ids_kind, int64_t hist_ld,
                                   const void* cand_ids, int64_t cand_ld, int64_t B, int T, const void* packed, int h1, int h2,
                                   float* emb, int64_t emb_ld, const int64_t* h_extras, int n_extras,
                                   const float* dense, int64_t dense_ld, int n_dense, int dense_col,
                                   void* workspace, int64_t workspace_bytes, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_din_fwd_tf32");
  if (!table || !hist_ids || !cand_ids || !packed || !emb || !workspace || B < 0 || V <= 0) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (n_extras < 0 || n_extras > DF_MAX_EXTRA || (n_extras > 0 && !h_extras) || n_dense < 0 || (n_dense > 0 && !dense)) return CTRHOT_EARG;
  DfShape s;
  if (!df_shape(T, K, h1, h2, s)) return CTRHOT_ESHAPE;
  if (emb_ld < 2 * K || hist_ld < T || cand_ld < 1) return CTRHOT_EARG;
  if (!aligned16(table) || !aligned16(packed) || !aligned16(emb) || !aligned16(workspace) || emb_ld % 4 != 0) return CTRHOT_ESHAPE;
  if (workspace_bytes < ctrhot_din_fwd_workspace_bytes(B, h1)) return CTRHOT_EARG;
  DfParams p{};
  p.table = table; p.V = V; p.hist_ids = hist_ids; p.hist_ld = hist_ld; p.cand_ids = cand_ids; p.cand_ld = cand_ld;
  p.ids_kind = ids_kind; p.U = static_cast<float*>(workspace); p.packed = static_cast<const float*>(packed);
  p.emb = emb; p.emb_ld = emb_ld; p.B = B; p.err_flag = err_flag; p.s = s;
  p.n_extra = n_extras;
  for (int i = 0; i < n_extras; ++i) {
    const int64_t* e = h_extras + 6 * i;
    DfExtra& x = p.extra[i];
    x.table = reinterpret_cast<const float*>(static_cast<uintptr_t>(e[0])); x.rows = e[1]; x.dim = (int)e[2];
    x.ids = reinterpret_cast<const void*>(static_cast<uintptr_t>(e[3])); x.ids_ld = e[4]; x.col = (int)e[5];
    if (!x.table || !x.ids || x.rows <= 0 || x.dim <= 0 || x.ids_ld < 1 || x.col < 2 * K || x.col + x.dim > emb_ld) return CTRHOT_EARG;
  }
  p.dense = dense; p.dense_ld = dense_ld; p.n_dense = n_dense; p.dense_col = dense_col;
  if (n_dense > 0 && (dense_ld < n_dense || dense_col < 2 * K || dense_col + n_dense > emb_ld)) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  DfSmem mp{};
  for (p.nraw = DF_NRAW; p.nraw >= 2; --p.nraw) {
    mp = df_smem(s, p.nraw);
    if (mp.total + 1024u <= 227u * 1024u) break;
  }
  if (p.nraw < 2) return CTRHOT_ESHAPE;
  const size_t smem = mp.total + 1024u;
  const int grid = (int)(B < kNumSMs ? B : kNumSMs);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (ids_kind == CTRHOT_IDS_I32) {
    CTRHOT_CUDA_OK(cudaFuncSetAttribute(din_fwd_kernel<CTRHOT_IDS_I32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "din_fwd attr");
    din_fwd_kernel<CTRHOT_IDS_I32><<<grid, DF_THREADS, smem, st>>>(p);
  } else {
    CTRHOT_CUDA_OK(cudaFuncSetAttribute(din_fwd_kernel<CTRHOT_IDS_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "din_fwd attr");
    din_fwd_kernel<CTRHOT_IDS_F32><<<grid, DF_THREADS, smem, st>>>(p);
  }
  CTRHOT_LAUNCH_OK("ctrhot_din_fwd_tf32");
  return CTRHOT_OK;
}

#ifdef DF_TRACE
extern "C" int ctrhot_debug_din_fwd_cta(unsigned long long* h_out) {
  CTRHOT_RANGE("ctrhot_debug_din_fwd_cta");
  return cudaMemcpyFromSymbol(h_out, ctrhot::df_cta, sizeof(unsigned long long) * 148 * 4) == cudaSuccess ? 0 : -3;
}
extern "C" int ctrhot_debug_din_fwd_trace(long long* h_out) {
  CTRHOT_RANGE("ctrhot_debug_din_fwd_trace");
  return cudaMemcpyFromSymbol(h_out, ctrhot::df_trace, sizeof(long long) * 14 * 128) == cudaSuccess ? 0 : -3;
}
#endif
