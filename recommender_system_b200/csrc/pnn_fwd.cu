// R15 PNN front: EmbedLayer -> [ flat | InnerProductLayer | OuterProductLayer ] in ONE launch (DL/model/pnn.py:28-49,
// DL/layer/interaction.py:166-215).  The gathered rows of 128 samples live in shared memory only: they are the A operand of
// the outer product's contraction, the q / p vectors of the epilogue and the source of the flat columns; what reaches HBM is
// the product vector the tower reads.
//
// The outer product  out[b, p] = sum_a e_j[a] * (sum_c W[a, p, c] e_i[c])  is a contraction with per-pair weights
// (k^2 FMA per pair: 20.8 K per sample at F = 26, k = 8), so its inner sum runs on the tensor pipe:
//     D[b, (slot, a)] = sum_c E_i[b, c] * Wb[(slot, a), c]        M = 128 samples, N = 8 x slots of run i, K = 8 (one TF32 k-step)
// with fp32 accuracy from a three-term split (A = At + Al, W = Wh + Wl:  At Wh + Al Wh + At Wl; At is what the MMA reads from
// the fp32 tile itself -- it truncates --, Al = tf32(x - At) is built per 4-field panel, Wh / Wl are packed once), and the
// outer sum  sum_a D[b, (slot, a)] e_j[b, a]  plus the inner product <e_i, e_j> run in the epilogue warps on TMEM lanes = rows.
//
// shared memory (k = 8, F <= 28):  E tile  ceil(F/4) panels of [128 rows x 128 B] (4 fields per 128-byte row, 128-byte
// swizzle: field f is k-slice f % 4 of panel f / 4), two Al panels, a 2-deep ring of 16 KB W tiles (one 256-column chunk:
// k-slices {hi, lo} x {columns 0..127, 128..255}), per-warp transpose buffers for coalesced stores.
// TMEM: two 256-column accumulators (MMA of chunk c + 1 under the epilogue of chunk c).
//
// Measured (B = 65 536, F = 26, one B200): 0.157 ms inner + outer, 0.130 ms outer only -- against 0.665 / 0.610 ms for gather +
// InnerProductLayer + OuterProductLayer as separate launches.  What bounds it (profiles/r02_pnn_front.md): the epilogue warps
// (two per scheduler) run ~1 000 dependent instructions per 16 pair slots -- table word -> row of field j from shared memory ->
// 16 FMA -> transpose -> store -- at ~0.3 IPC; removing the global stores, the TMEM loads or the transposes each moves the
// time by < 20 %, the tensor pipe is 5 % busy and HBM traffic (256 MB) would take 40 us.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "common.cuh"
#include "tc05.cuh"

namespace ctrhot {
using namespace tc;
namespace {

constexpr int PF_K = 8;
constexpr int PF_BM = 128;
constexpr int PF_MAXF = 28;
constexpr int PF_MAXSLOT = 416;                 // 13 chunks of 32 pair slots
constexpr int PF_MAXOPS = 128;
constexpr int PF_NEPI = 8, PF_NGAT = 4;         // epilogue warps 0-7, gather/builder warps 8-11, MMA 12, W producer 13
constexpr int PF_THREADS = (PF_NEPI + PF_NGAT + 2) * 32;
constexpr int PF_WRING = 2;
constexpr uint32_t PF_PANEL = 16384;
// packed buffer: header, slot tables, MMA op list, then one 16 KB W tile per chunk
constexpr int PF_O_IJ = 64, PF_O_P = PF_O_IJ + 4 * PF_MAXSLOT, PF_O_OPS = PF_O_P + 2 * PF_MAXSLOT;
constexpr int PF_TAB_BYTES = PF_O_OPS + 4 * PF_MAXOPS;   // 3072
constexpr int PF_O_W = 4096;

struct PfPlan { int F, P, nslots, nch, nops, npa; };
__host__ __device__ inline PfPlan pf_plan(int F) {
  PfPlan q;
  q.F = F; q.P = F * (F - 1) / 2;
  int ns = 0, nops = 0;
  for (int i = 0; i < F - 1; ++i) {
    const int len = F - 1 - i, plen = len + (len & 1);
    int s = ns;
    while (s < ns + plen) { int e = (s / 16 + 1) * 16; if (e > ns + plen) e = ns + plen; ++nops; s = e; }
    ns += plen;
  }
  q.nslots = ns; q.nch = (ns + 31) / 32; q.nops = nops; q.npa = (F - 2) / 4 + 1;
  return q;
}

// one block: thread 0 lays out the pair slots (run i = pairs (i, i+1..F-1), padded to an even count so that every MMA has
// N % 16 == 0) and the op list; then all threads split W[a, p, c] into tf32 hi / lo in the tiles' swizzled layout
__global__ void pnn_pack_kernel(const float* __restrict__ W, int F, uint8_t* __restrict__ packed) {
  const PfPlan q = pf_plan(F);
  int32_t* hdr = reinterpret_cast<int32_t*>(packed);
  uint32_t* ij = reinterpret_cast<uint32_t*>(packed + PF_O_IJ);   // i << 20 | byte offset of field j in the E tile
  int16_t* sp = reinterpret_cast<int16_t*>(packed + PF_O_P);
  uint32_t* ops = reinterpret_cast<uint32_t*>(packed + PF_O_OPS);
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    hdr[0] = F; hdr[1] = q.P; hdr[2] = q.nslots; hdr[3] = q.nch; hdr[4] = q.nops; hdr[5] = q.npa;
    // bytes 32..47: per chunk, the E panels no later chunk reads (panel m: fields 4m..4m+3 are last used by run 4m+3, the
    // panel of field F-1 by the last run) -- the epilogue frees them for the next tile's gather as it goes
    uint8_t* cf = packed + 32;
    for (int c = 0; c < 16; ++c) cf[c] = 0;
    const int npan = (F + 3) / 4;
    int ns = 0, no = 0, p = 0;
    for (int i = 0; i < F - 1; ++i) {
      const int len = F - 1 - i, plen = len + (len & 1);
      for (int m = 0; m < npan; ++m)
        if (i == ((m == npan - 1) ? F - 2 : 4 * m + 3)) cf[(ns + plen - 1) / 32] |= (uint8_t)(1u << m);
      for (int t = 0; t < plen; ++t) {
        const int j = (t < len) ? i + 1 + t : i;                 // a padding slot reads field i and stores nothing
        ij[ns + t] = ((uint32_t)i << 20) | ((uint32_t)(j >> 2) * PF_PANEL + (uint32_t)(j & 3) * 32u);
        sp[ns + t] = (t < len) ? (int16_t)p++ : (int16_t)-1;
      }
      int s = ns;
      while (s < ns + plen) {
        int e = (s / 16 + 1) * 16; if (e > ns + plen) e = ns + plen;
        ops[no++] = ((uint32_t)(s / 32) << 24) | ((uint32_t)(s % 32) << 16) | ((uint32_t)(e - s) << 8) | (uint32_t)i;
        s = e;
      }
      ns += plen;
    }
    for (int t = ns; t < PF_MAXSLOT; ++t) { ij[t] = 0u; sp[t] = -1; }
    for (int t = no; t < PF_MAXOPS; ++t) ops[t] = 0xFFFFFFFFu;
  }
  __syncthreads();
  // tile of chunk c: row n' (0..127), k-slice s: 0 hi / 1 lo of column n', 2 hi / 3 lo of column 128 + n'; column n = slot * 8 + a
  const int total = q.nch * 256 * PF_K;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int c = t & 7, n = (t >> 3) & 255, ch = t >> 11;
    const int slot = ch * 32 + (n >> 3), a = n & 7;
    const int p = (slot < PF_MAXSLOT) ? sp[slot] : -1;
    const float w = (p >= 0) ? W[((size_t)a * q.P + p) * PF_K + c] : 0.f;
    const float hi = to_tf32_rna(w), lo = to_tf32_rna(w - hi);
    const uint32_t row = n & 127, half = n >> 7;
    uint8_t* tile = packed + PF_O_W + (size_t)ch * PF_PANEL;
    *reinterpret_cast<float*>(tile + sw128_offset(row, half * 4 + (c >> 2)) + (c & 3) * 4) = hi;
    *reinterpret_cast<float*>(tile + sw128_offset(row, half * 4 + 2 + (c >> 2)) + (c & 3) * 4) = lo;
  }
}

struct PfParams {
  const float* arena; const int64_t* row_off; const void* ids; int64_t ids_ld; int64_t B;
  const uint8_t* packed; float* out; int64_t out_ld; int* err_flag;
  int F, do_inner, col_inner, col_outer, width;
  int32_t* signal_flags; int epoch, overlap;
  int write_flat;          // 0: products only (ids == NULL: `arena` is a gathered [B, F, 8] matrix with row stride ids_ld)   // chained launches: tile t of `out` is published by storing epoch into signal_flags[t]
};

struct PfSmem { uint32_t e, lo, wr, stg, tab, bars, total; };
__host__ __device__ inline PfSmem pf_smem(int F) {
  PfSmem m;
  const uint32_t npan = (uint32_t)(F + 3) / 4;
  m.e = 0; m.lo = npan * PF_PANEL; m.wr = m.lo + 2 * PF_PANEL; m.stg = m.wr + PF_WRING * PF_PANEL;
  m.tab = m.stg + PF_NEPI * 2 * 32 * 17 * 4; m.bars = m.tab + ((PF_TAB_BYTES + 15) / 16) * 16;
  m.total = m.bars + 192 + 8 * (PF_MAXF + 4);
  return m;
}

#ifdef PF_TRACE
#define PF_T0() long long _t = clock64()
#define PF_ACC(k) do { long long _n = clock64(); tr[k] += _n - _t; _t = _n; } while (0)
#else
#define PF_T0() do { } while (0)
#define PF_ACC(k) do { } while (0)
#endif

template <int IDKIND, bool INNER>
__global__ void __launch_bounds__(PF_THREADS, 1)
pnn_front_kernel(const __grid_constant__ PfParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* bp = smem_raw + (base - raw_addr);
  const int F = p.F;
  const PfSmem mp = pf_smem(F);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t bars = base + mp.bars;
  const uint32_t e_full = bars;
  auto p_free = [&](int i) { return bars + 136u + 8u * i; };      // E panel i is dead for this tile (8 epilogue warps)
  auto lo_full = [&](int i) { return bars + 16u + 8u * i; };
  auto lo_free = [&](int i) { return bars + 32u + 8u * i; };
  auto w_full = [&](int i) { return bars + 48u + 8u * i; };
  auto w_free = [&](int i) { return bars + 72u + 8u * i; };
  auto d_full = [&](int i) { return bars + 96u + 8u * i; };
  auto d_free = [&](int i) { return bars + 112u + 8u * i; };
  const uint32_t slot_addr = bars + 128u;
  volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(bp + mp.bars + 128);
  const int32_t* hdr = reinterpret_cast<const int32_t*>(bp + mp.tab);
  const uint8_t* t_cf = bp + mp.tab + 32;
  const uint32_t* t_sl = reinterpret_cast<const uint32_t*>(bp + mp.tab + PF_O_IJ);
  const int16_t* t_p = reinterpret_cast<const int16_t*>(bp + mp.tab + PF_O_P);
  const uint32_t* t_ops = reinterpret_cast<const uint32_t*>(bp + mp.tab + PF_O_OPS);

  if (tid == 0) {
    mbar_init(e_full, PF_NGAT * 32);
    for (int i = 0; i < 7; ++i) mbar_init(p_free(i), PF_NEPI);
    for (int i = 0; i < 2; ++i) { mbar_init(lo_full(i), PF_NGAT * 32); mbar_init(lo_free(i), 1); }
    for (int i = 0; i < PF_WRING; ++i) { mbar_init(w_full(i), 1); mbar_init(w_free(i), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(d_full(i), 1); mbar_init(d_free(i), PF_NEPI); }
    mbar_fence_init();
  }
  for (int i = tid; i < PF_TAB_BYTES / 4; i += PF_THREADS)
    reinterpret_cast<uint32_t*>(bp + mp.tab)[i] = __ldg(reinterpret_cast<const uint32_t*>(p.packed) + i);
  int64_t* s_off = reinterpret_cast<int64_t*>(bp + mp.bars + 192);
  for (int i = tid; i <= PF_MAXF; i += PF_THREADS) s_off[i] = (i <= F && p.row_off) ? __ldg(p.row_off + i) : 0;
  if (warp == PF_NEPI + PF_NGAT) { tmem_alloc(slot_addr, 512u); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot_ptr;
  // the tower launched behind this kernel may start on the SMs it frees (it waits for signal_flags, not for this grid)
  if (p.overlap && threadIdx.x == 0) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int nch = hdr[3], nops = hdr[4], npa = hdr[5];
  const int64_t n_tiles = (p.B + PF_BM - 1) / PF_BM;
  const int64_t my_tiles = (n_tiles > blockIdx.x) ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

#ifdef PF_TRACE
  long long tr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long t_start = clock64();
#endif
  if (warp < PF_NEPI) {
    // ================================================================ epilogue: lanes = rows, warp = (lane quarter, column half)
    const int q = warp & 3, h = warp >> 2;
    const uint32_t row = (uint32_t)(q * 32 + lane);
    float* so = reinterpret_cast<float*>(bp + mp.stg) + warp * 2 * 32 * 17;   // outer values [row][17], then the inner ones
    float* si = so + 32 * 17;
    const uint8_t* erow = bp + mp.e + row * 128u;                  // this row in panel 0; a field's bytes are at (offset ^ r7)
    const uint32_t r7 = (row & 7u) << 4;
    const uint32_t erow_a = base + mp.e + row * 128u;
    const int rd = (lane >> 4) * 17 + (lane & 15);                 // transposed read: rows 2 r2 + (lane >> 4), column lane & 15
    uint32_t dc = 0;
#pragma unroll 1
    for (int64_t it = 0; it < my_tiles; ++it) {
      const int64_t b0 = (blockIdx.x + it * gridDim.x) * PF_BM;
      const int64_t rem = p.B - (b0 + q * 32);
      const int nrow = rem > 32 ? 32 : (int)rem;                   // rows of this warp that exist (may be <= 0)
      float* orow = p.out + (b0 + q * 32 + (lane >> 4)) * p.out_ld;
      PF_T0();
      mbar_wait(e_full, (uint32_t)(it & 1));
      PF_ACC(0);
      uint32_t cur_i = 0xFFFFFFFFu;
      float ei[PF_K];
#pragma unroll
      for (int a = 0; a < PF_K; ++a) ei[a] = 0.f;
#pragma unroll 1
      for (int c = 0; c < nch; ++c, ++dc) {
        const uint32_t buf = dc & 1u;
        mbar_wait(d_full(buf), (dc >> 1) & 1u);
        tc_fence_after();
        PF_ACC(1);
        // 16 columns (two pair slots) per TMEM load, the next load in flight under the FMAs of the current one
        const uint32_t tcol = tmem_base + ((uint32_t)(q * 32) << 16) + buf * 256u + (uint32_t)h * 128u;
        uint32_t dA[16], dB[16];
        tmem_ld16_issue(tcol, dA);
        tmem_ld16_wait(dA);
        // the 16 slot words of this half chunk (warp-uniform) and the rows of field j, one slot AHEAD of its FMAs;
        // straight-line code on purpose: the loads of later slots are scheduled under the FMAs of earlier ones
        uint32_t swv[16];
        {
          const uint4* t4 = reinterpret_cast<const uint4*>(t_sl + c * 32 + h * 16);
#pragma unroll
          for (int u = 0; u < 4; ++u) { const uint4 t = t4[u]; swv[4 * u] = t.x; swv[4 * u + 1] = t.y; swv[4 * u + 2] = t.z; swv[4 * u + 3] = t.w; }
        }
        float4 yA0, yA1, yB0, yB1;
        {
          const uint32_t oj = (swv[0] & 0xFFFFFu) ^ r7;
          yA0 = *reinterpret_cast<const float4*>(erow + oj); yA1 = *reinterpret_cast<const float4*>(erow + (oj ^ 16u));
        }
        auto one_slot = [&](const uint32_t* d, int sl, const float4& y0, const float4& y1, float4& n0, float4& n1) {
          if (sl + 1 < 16) {
            const uint32_t oj = (swv[sl + 1] & 0xFFFFFu) ^ r7;
            n0 = *reinterpret_cast<const float4*>(erow + oj); n1 = *reinterpret_cast<const float4*>(erow + (oj ^ 16u));
          }
          if (INNER) {                                                        // a new run: its field i (predicated, no branch)
            const uint32_t ni = swv[sl] >> 20;
            const uint32_t oi = (((ni >> 2) * PF_PANEL) | ((ni & 3u) * 32u)) ^ r7;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %8, %9;\n\t"
                         "@p ld.shared.v4.f32 {%0,%1,%2,%3}, [%10];\n\t@p ld.shared.v4.f32 {%4,%5,%6,%7}, [%11];\n\t}"
                         : "+f"(ei[0]), "+f"(ei[1]), "+f"(ei[2]), "+f"(ei[3]), "+f"(ei[4]), "+f"(ei[5]), "+f"(ei[6]), "+f"(ei[7])
                         : "r"(ni), "r"(cur_i), "r"(erow_a + oi), "r"(erow_a + (oi ^ 16u)));
            cur_i = ni;
          }
          const float ej[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
          float o = 0.f, n = 0.f;
#pragma unroll
          for (int a = 0; a < PF_K; ++a) {
            o = fmaf(__uint_as_float(d[a]), ej[a], o);
            if (INNER) n = fmaf(ei[a], ej[a], n);
          }
          so[lane * 17 + sl] = o;
          if (INNER) si[lane * 17 + sl] = n;
        };
        auto two_slots = [&](const uint32_t (&d)[16], int g) {
          one_slot(d, 2 * g, yA0, yA1, yB0, yB1);
          one_slot(d + 8, 2 * g + 1, yB0, yB1, yA0, yA1);
        };
#pragma unroll
        for (int g = 0; g < 8; g += 2) {
          tmem_ld16_issue(tcol + (uint32_t)(g + 1) * 16u, dB);
          two_slots(dA, g);
          tmem_ld16_wait(dB);
          if (g + 2 < 8) tmem_ld16_issue(tcol + (uint32_t)(g + 2) * 16u, dA);
          two_slots(dB, g + 1);
          if (g + 2 < 8) tmem_ld16_wait(dA);
        }
        tc_fence_before();
        __syncwarp();
        PF_ACC(2);
        if (lane == 0) mbar_arrive(d_free(buf));
        // transposed write-out: one store = 2 rows x 16 consecutive pair columns (padding slots have no column)
        const int pcol = t_p[c * 32 + h * 16 + (lane & 15)];
        const int lim = (pcol >= 0) ? nrow - (lane >> 4) : 0;                 // row 2 r2 + (lane >> 4) exists iff 2 r2 < lim
        const int64_t ld2 = 2 * p.out_ld;
#pragma unroll
        for (int pass = INNER ? 0 : 1; pass < 2; ++pass) {
          const float* src = (pass ? so : si) + rd;
          float* op = orow + (pass ? p.col_outer : p.col_inner) + pcol;
#pragma unroll
          for (int r0 = 0; r0 < 16; r0 += 8) {
            float v[8];
#pragma unroll
            for (int r2 = 0; r2 < 8; ++r2) v[r2] = src[(r0 + r2) * 34];
#pragma unroll
            for (int r2 = 0; r2 < 8; ++r2) {
              if (2 * (r0 + r2) < lim) asm volatile("st.global.f32 [%0], %1;" :: "l"(op), "f"(v[r2]) : "memory");
              op += ld2;
            }
          }
        }
        __syncwarp();
        PF_ACC(3);
        if (lane == 0) {
          for (uint32_t fm = t_cf[c]; fm; fm &= fm - 1) mbar_arrive(p_free(__ffs((int)fm) - 1));
        }
      }
    }
  } else if (warp < PF_NEPI + PF_NGAT) {
    // ================================================================ gather (thread = row) + flat columns + Al panels
    const int gt = tid - PF_NEPI * 32, gw = warp - PF_NEPI;
    const int npan = (F + 3) / 4;
    uint32_t lp = 0;                                               // Al panel sequence number (buffer lp & 1)
    // this row's F ids, fetched one tile AHEAD (all loads in flight at once): -1 = out of range / no such row
    int32_t rid[PF_MAXF];
    auto load_ids = [&](int64_t it) {
      const int64_t b = (blockIdx.x + it * gridDim.x) * PF_BM + gt;
      const bool live = it < my_tiles && b < p.B;
#pragma unroll
      for (int f = 0; f < PF_MAXF; ++f) {
        int64_t v = (live && f < F && p.ids) ? IdLoad<IDKIND>::at(p.ids, b * p.ids_ld + f) : 0;
        rid[f] = (v < 0 || v > 0x7FFFFFFF) ? -1 : (int32_t)v;
      }
    };
    load_ids(0);
#pragma unroll 1
    for (int64_t it = 0; it < my_tiles; ++it) {
      const int64_t b0 = (blockIdx.x + it * gridDim.x) * PF_BM;
      PF_T0();
      const bool live = b0 + gt < p.B;
#pragma unroll
      for (int f = 0; f < PF_MAXF; ++f) {
        if (f < 4 * npan) {
          if ((f & 3) == 0 && it > 0) mbar_wait(p_free(f >> 2), (uint32_t)((it - 1) & 1));
          const float* src = p.arena;
          uint32_t nb = 0;
          if (live && f < F) {
            if (!p.ids) { src = p.arena + (b0 + gt) * p.ids_ld + f * PF_K; nb = 16; }      // rows of an [B, F, 8] matrix
            else {
              const int64_t r0 = s_off[f], r1 = s_off[f + 1];
              if (rid[f] >= 0 && rid[f] < r1 - r0) { src = p.arena + (r0 + rid[f]) * PF_K; nb = 16; }
              else if (p.err_flag) *p.err_flag = 1;
            }
          }
          const uint32_t dst = base + mp.e + (uint32_t)(f >> 2) * PF_PANEL;
          cp_async16(dst + sw128_offset((uint32_t)gt, (uint32_t)(f & 3) * 2), src, nb);
          cp_async16(dst + sw128_offset((uint32_t)gt, (uint32_t)(f & 3) * 2 + 1), src + 4, nb);
        }
      }
      PF_ACC(0);
      load_ids(it + 1);
      cp_async_commit();
      cp_async_wait<0>();
      PF_ACC(1);
      fence_proxy_async_smem();
      mbar_arrive(e_full);
      named_bar_sync(1, PF_NGAT * 32);
      // tile it - 1 is complete in HBM: its last E panel was released (every epilogue warp is past its last store) and every
      // gather warp is past its flat columns (the barrier above)
      if (p.signal_flags && it > 0 && gt == 0) {
        __threadfence();
        asm volatile("st.release.gpu.global.s32 [%0], %1;" :: "l"(p.signal_flags + (blockIdx.x + (it - 1) * gridDim.x)), "r"(p.epoch) : "memory");
      }
      // Al = tf32(x - trunc_tf32(x)) for the panels that hold a run's field i (at most two ahead of the MMAs); the first two
      // are on the critical path of the tile and go before the flat columns
      auto build_lo = [&](int m) {
        const uint32_t bf = lp & 1u;
        mbar_wait(lo_free(bf), ((lp >> 1) & 1u) ^ 1u);
        const float4* s4 = reinterpret_cast<const float4*>(bp + mp.e + (uint32_t)m * PF_PANEL);
        float4* d4 = reinterpret_cast<float4*>(bp + mp.lo + bf * PF_PANEL);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          float4 x = s4[gt + 128 * u];
          x.x = to_tf32_rna(x.x - __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u));
          x.y = to_tf32_rna(x.y - __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u));
          x.z = to_tf32_rna(x.z - __uint_as_float(__float_as_uint(x.z) & 0xFFFFE000u));
          x.w = to_tf32_rna(x.w - __uint_as_float(__float_as_uint(x.w) & 0xFFFFE000u));
          d4[gt + 128 * u] = x;
        }
        fence_proxy_async_smem();
        mbar_arrive(lo_full(bf));
        ++lp;
      };
      const int n_first = npa < 2 ? npa : 2;
#pragma unroll 1
      PF_ACC(2);
      for (int m = 0; m < n_first; ++m) build_lo(m);
      PF_ACC(3);
      // flat columns [0, F k) and the zero padding [width, out_ld): coalesced 16-byte stores, rows of this warp
      {
        const int64_t remr = p.B - (b0 + gw * 32);
        const int nr = remr > 32 ? 32 : (int)remr;
        const int l1 = lane + 32;                                          // second 16-byte piece of the row, if 2 F > 32
        const uint32_t o0 = (uint32_t)(lane >> 3) * PF_PANEL, c0 = (uint32_t)(lane & 7);
        const uint32_t o1 = (uint32_t)(l1 >> 3) * PF_PANEL, c1 = (uint32_t)(l1 & 7);
        const bool h0 = lane < 2 * F, h1 = l1 < 2 * F;
#pragma unroll 4
        for (int r = 0; r < nr; ++r) {
          const uint32_t rr = (uint32_t)(gw * 32 + r);
          float* orow = p.out + (b0 + rr) * p.out_ld;
          const uint8_t* er = bp + mp.e + rr * 128u;
          float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
          if (h0) v0 = *reinterpret_cast<const float4*>(er + o0 + ((c0 ^ (rr & 7u)) << 4));
          if (h1) v1 = *reinterpret_cast<const float4*>(er + o1 + ((c1 ^ (rr & 7u)) << 4));
          if (h0 && p.write_flat) *reinterpret_cast<float4*>(orow + 4 * lane) = v0;
          if (h1 && p.write_flat) *reinterpret_cast<float4*>(orow + 4 * l1) = v1;
          for (int l = p.width + lane; l < p.out_ld; l += 32) orow[l] = 0.f;
        }
      }
      PF_ACC(4);
      for (int m = n_first; m < npa; ++m) build_lo(m);
      PF_ACC(5);
    }
    if (p.signal_flags && my_tiles > 0) {                            // the CTA's last tile
      for (int m = 0; m < npan; ++m) mbar_wait(p_free(m), (uint32_t)((my_tiles - 1) & 1));
      named_bar_sync(1, PF_NGAT * 32);
      if (gt == 0) {
        __threadfence();
        asm volatile("st.release.gpu.global.s32 [%0], %1;" :: "l"(p.signal_flags + (blockIdx.x + (my_tiles - 1) * gridDim.x)), "r"(p.epoch) : "memory");
      }
    }
  } else if (warp == PF_NEPI + PF_NGAT) {
    // ================================================================ MMA issuer (converged warp, one elected lane issues)
    uint32_t wc = 0, dc = 0, lp = 0;
#pragma unroll 1
    for (int64_t it = 0; it < my_tiles; ++it) {
      PF_T0();
      mbar_wait(e_full, (uint32_t)(it & 1));
      PF_ACC(0);
      int cur_pan = -1, oi = 0;
#pragma unroll 1
      for (int c = 0; c < nch; ++c, ++wc, ++dc) {
        const uint32_t st = wc % PF_WRING, buf = dc & 1u;
        mbar_wait(w_full(st), (wc / PF_WRING) & 1u);
        PF_ACC(1);
        mbar_wait(d_free(buf), ((dc >> 1) & 1u) ^ 1u);
        tc_fence_after();
        PF_ACC(2);
        const uint32_t wt = base + mp.wr + st * PF_PANEL;
#pragma unroll 1
        while (oi < nops && (int)(t_ops[oi] >> 24) == c) {
          const uint32_t op = t_ops[oi++];
          const uint32_t s0 = (op >> 16) & 255u, ns = (op >> 8) & 255u, i = op & 255u;
          const int pan = (int)(i >> 2);
          if (pan != cur_pan) {
            if (cur_pan >= 0) { if (elect_one()) mma_commit(lo_free((lp - 1) & 1u)); __syncwarp(); }
            PF_ACC(3);
            mbar_wait(lo_full(lp & 1u), (lp >> 1) & 1u);
            tc_fence_after();
            PF_ACC(4);
            ++lp; cur_pan = pan;
          }
          const uint32_t half = s0 >> 4, sh = s0 & 15u;
          const uint64_t a_hi = make_desc_sw128(base + mp.e + (uint32_t)pan * PF_PANEL + (i & 3u) * 32u);
          const uint64_t a_lo = make_desc_sw128(base + mp.lo + ((lp - 1) & 1u) * PF_PANEL + (i & 3u) * 32u);
          const uint64_t b_hi = make_desc_sw128(wt + sh * 1024u + half * 64u);
          const uint64_t b_lo = make_desc_sw128(wt + sh * 1024u + half * 64u + 32u);
          const uint32_t idesc = make_idesc_tf32(PF_BM, ns * 8u);
          const uint32_t dcol = tmem_base + buf * 256u + s0 * 8u;
          if (elect_one()) {
            mma_tf32_ss(dcol, a_hi, b_hi, idesc, 0u);
            mma_tf32_ss(dcol, a_lo, b_hi, idesc, 1u);
            mma_tf32_ss(dcol, a_hi, b_lo, idesc, 1u);
          }
          __syncwarp();
        }
        if (elect_one()) { mma_commit(d_full(buf)); mma_commit(w_free(st)); }
        __syncwarp();
        PF_ACC(3);
      }
      if (elect_one()) mma_commit(lo_free((lp - 1) & 1u));
      __syncwarp();
    }
  } else {
    // ================================================================ W producer: one 16 KB tile per chunk through the ring
    uint32_t wc = 0;
#pragma unroll 1
    for (int64_t it = 0; it < my_tiles; ++it) {
#pragma unroll 1
      for (int c = 0; c < nch; ++c, ++wc) {
        const uint32_t st = wc % PF_WRING;
        mbar_wait(w_free(st), ((wc / PF_WRING) & 1u) ^ 1u);
        if (lane == 0) {
          mbar_arrive_expect_tx(w_full(st), PF_PANEL);
          bulk_g2s(base + mp.wr + st * PF_PANEL, p.packed + PF_O_W + (size_t)c * PF_PANEL, PF_PANEL, w_full(st));
        }
        __syncwarp();
      }
    }
  }
#ifdef PF_TRACE
  if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 5 || warp == 8 || warp == 12))
    printf("w%d tiles %d total %lld : %lld %lld %lld %lld %lld %lld\n", warp, (int)my_tiles, clock64() - t_start, tr[0], tr[1], tr[2], tr[3], tr[4], tr[5]);
#endif
  tc_fence_before();
  __syncthreads();
  if (warp == PF_NEPI + PF_NGAT) { tc_fence_after(); tmem_dealloc(tmem_base, 512u); }
}

}  // namespace
}  // namespace ctrhot

using namespace ctrhot;

extern "C" int64_t ctrhot_pnn_pack_bytes(int F, int k) {
  if (k != PF_K || F < 2 || F > PF_MAXF) return CTRHOT_ESHAPE;
  return (int64_t)PF_O_W + (int64_t)pf_plan(F).nch * PF_PANEL;
}

extern "C" int ctrhot_pnn_pack(const float* W, int F, int k, void* packed, void* stream) {
  CTRHOT_RANGE("ctrhot_pnn_pack");
  if (!W || !packed) return CTRHOT_EARG;
  if (k != PF_K || F < 2 || F > PF_MAXF || !aligned16(packed)) return CTRHOT_ESHAPE;
  pnn_pack_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(W, F, static_cast<uint8_t*>(packed));
  CTRHOT_LAUNCH_OK("ctrhot_pnn_pack");
  return CTRHOT_OK;
}

extern "C" int ctrhot_pnn_front_f32(const float* arena, const int64_t* row_off, int F, int k,
                                    const void* ids, int ids_kind, int64_t ids_ld, int64_t B,
                                    const void* packed, int do_inner, float* out, int64_t out_ld,
                                    int* err_flag, int32_t* signal_flags, int epoch, int launch, void* stream) {
  CTRHOT_RANGE("ctrhot_pnn_front_f32");
  if (!arena || !row_off || !ids || !packed || !out || B < 0 || ids_ld < F) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (k != PF_K || F < 2 || F > PF_MAXF) return CTRHOT_ESHAPE;
  const PfPlan q = pf_plan(F);
  const int width = F * k + q.P * (do_inner ? 2 : 1);
  if (out_ld < width || out_ld % 4 != 0 || !aligned16(out) || !aligned16(arena) || !aligned16(packed)) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  PfParams p;
  p.arena = arena; p.row_off = row_off; p.ids = ids; p.ids_ld = ids_ld; p.B = B;
  p.packed = static_cast<const uint8_t*>(packed); p.out = out; p.out_ld = out_ld; p.err_flag = err_flag;
  p.write_flat = 1;
  p.signal_flags = signal_flags; p.epoch = epoch; p.overlap = (launch & CTRHOT_LAUNCH_OVERLAP) && signal_flags ? 1 : 0;
  p.F = F; p.do_inner = do_inner ? 1 : 0; p.col_inner = F * k; p.col_outer = F * k + (do_inner ? q.P : 0); p.width = width;
  const size_t smem = pf_smem(F).total + 1024;
  const int64_t n_tiles = (B + PF_BM - 1) / PF_BM;
  const int grid = (int)(n_tiles < kNumSMs ? n_tiles : kNumSMs);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define PF_LAUNCH(KIND, INN)                                                                                              \
  do {                                                                                                                    \
    CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_front_kernel<KIND, INN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), \
                   "pnn_front attr");                                                                                     \
    pnn_front_kernel<KIND, INN><<<grid, PF_THREADS, smem, st>>>(p);                                                       \
  } while (0)
  if (ids_kind == CTRHOT_IDS_I32) { if (do_inner) PF_LAUNCH(CTRHOT_IDS_I32, true); else PF_LAUNCH(CTRHOT_IDS_I32, false); }
  else { if (do_inner) PF_LAUNCH(CTRHOT_IDS_F32, true); else PF_LAUNCH(CTRHOT_IDS_F32, false); }
#undef PF_LAUNCH
  CTRHOT_LAUNCH_OK("ctrhot_pnn_front_f32");
  return CTRHOT_OK;
}

// InnerProductLayer / OuterProductLayer on an already gathered e [B, F, 8] (row stride e_ld) by the same kernel:
// out[b, 0:P) = <e_i, e_j> (if do_inner), then out[b, .. + P) = e_j^T W_p e_i; columns beyond are not touched.
extern "C" int ctrhot_pnn_products_f32(const float* e, int64_t e_ld, int F, int k, int64_t B, const void* packed, int do_inner,
                                       float* out, int64_t out_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_pnn_products_f32");
  if (!e || !packed || !out || B < 0 || e_ld < (int64_t)F * k) return CTRHOT_EARG;
  if (k != PF_K || F < 2 || F > PF_MAXF) return CTRHOT_ESHAPE;
  const PfPlan q = pf_plan(F);
  const int width = q.P * (do_inner ? 2 : 1);
  if (out_ld < width || out_ld > 0x7FFFFFFF || e_ld % 4 != 0 || !aligned16(e) || !aligned16(packed)) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  PfParams p;
  p.arena = e; p.row_off = nullptr; p.ids = nullptr; p.ids_ld = e_ld; p.B = B;
  p.packed = static_cast<const uint8_t*>(packed); p.out = out; p.out_ld = out_ld; p.err_flag = nullptr;
  p.write_flat = 0; p.signal_flags = nullptr; p.epoch = 0; p.overlap = 0;
  p.F = F; p.do_inner = do_inner ? 1 : 0; p.col_inner = 0; p.col_outer = do_inner ? q.P : 0;
  p.width = (int)out_ld;                                   // nothing outside [0, width) of a row is written (out may be a view)
  const size_t smem = pf_smem(F).total + 1024;
  const int64_t n_tiles = (B + PF_BM - 1) / PF_BM;
  const int grid = (int)(n_tiles < kNumSMs ? n_tiles : kNumSMs);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (do_inner) {
    CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_front_kernel<CTRHOT_IDS_I32, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "pnn_products attr");
    pnn_front_kernel<CTRHOT_IDS_I32, true><<<grid, PF_THREADS, smem, st>>>(p);
  } else {
    CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_front_kernel<CTRHOT_IDS_I32, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "pnn_products attr");
    pnn_front_kernel<CTRHOT_IDS_I32, false><<<grid, PF_THREADS, smem, st>>>(p);
  }
  CTRHOT_LAUNCH_OK("ctrhot_pnn_products_f32");
  return CTRHOT_OK;
}
