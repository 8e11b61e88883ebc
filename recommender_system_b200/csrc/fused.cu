// K-fused: the whole CTR forward for embedding + dense-interaction + tower models in ONE persistent kernel
// (DeepFM, DL/model/deepFM.py:23-31; DCN, DL/model/dcn.py:24-34; plain EmbedLayer -> DNNLayer towers).
//
//   rows      x_m = [ Embed(ids_m) | dense_m ]            gathered ONCE from the tables, never written to HBM
//   tower     y_m = w_out . act_L( ... act_1(x_m W_1 + b_1) ... ) + b_out        tcgen05 kind::tf32, fp32 accumulate
//   side      S_m = x_m @ side[:, 0:16]                    fp32-accurate on the tensor pipe by a split-TF32 product:
//                   x = x_hi + x_lo,  side = s_hi + s_lo,  S = x_hi s_hi + x_hi s_lo + x_lo s_hi   (error ~2^-21 |x||s|)
//             q_m = sum_j x_mj^2 sq_w[j]                   CUDA cores (fp32)
//   combine   DeepFM: sigmoid(0.5 * (w0 + S[kfm] + 0.5 * (sum_c S[c]^2 - q) + y))     (FMLayer, interaction.py:106-114)
//             DCN   : the L cross layers collapse to scalars: x_l = a_l x_0 + c_l with a_0 = 1, c_l = sum_{i<l} b_i,
//                     d_l = a_l (x_0 . w_l) + c_l . w_l,  a_{l+1} = a_l + d_l     (CrossLayer, interaction.py:75-83)
//                     so side = [w_0 .. w_{L-1}, final_w] and  out = sigmoid(a_L S[L] + c_L.final_w + wy * y + bias)
//
// One CTA per SM (persistent over 128-row tiles), 18 warps:
//   warps 0-7   epilogue: tcgen05.ld the accumulator -> bias + activation -> TF32 -> tcgen05.st back IN PLACE as the next
//               layer's A operand (hidden activations never leave TMEM); the last layer's epilogue does the final
//               Dense(h -> 1) dot, the combine and writes 4 bytes per row.  Warp w owns TMEM lanes 32*(w%4).., column half w/4.
//   warps 8-11  gather: 16-byte cp.async (LDGSTS) straight from the embedding rows into a smem ring of [128 x 32] chunks;
//               completion is tracked by mbarrier (cp.async.mbarrier.arrive.noinc), ids are fetched 3 chunks ahead.
//   warps 12-15 stager: one thread per tile row = per TMEM lane.  Reads its 32 floats of the chunk from smem and writes
//               x and x_lo = x - trunc_tf32(x) into a TMEM ring with tcgen05.st; accumulates q_m.  The first layer's A
//               operand therefore comes from TMEM: shared memory carries each gathered byte once in and once out, which
//               matters because a TF32 SS-mode MMA at M=128 would spend the whole smem bandwidth (128 B/clk) on operands.
//   warp 16     TMEM owner + MMA issuer (one lane).  Per 32-column K chunk: D1 += x W1 (N = N1), DS += x [s_hi | s_lo]
//               (N = 32), DS[0:16] += x_lo s_hi (N = 16) -- the tensor pipe drops the low 13 mantissa bits of x itself, i.e.
//               it multiplies by trunc_tf32(x); every layer reads A from TMEM (TS form).
//   warp 17     W producer: one cp.async.bulk (TMA engine) per K chunk of the pre-swizzled weights, multicast to the
//               CLUSTER so one L2 read feeds C SMs.
// TMEM columns: D1/H1 at 0, DS (32) next, then a region shared in time by the A ring (2..4 slots of x|x_lo, live during
// layer 1) and D2/H2, D3 (live afterwards).  HBM traffic = ids + dense + each needed row once + 4 B out.
#include <new>
#include "common.cuh"
#include "tc05.cuh"
#include "mlp.cuh"

namespace ctrhot {
using namespace tc;

constexpr int FU_BM = 128;
#ifndef FU_NB_
#define FU_NB_ 4
#endif
constexpr int FU_NB = FU_NB_;               // W ring, at most (the host takes one stage off when shared memory is short)
constexpr int FU_NAS = 4;                   // A slots in TMEM, at most (64 columns each: x | x_lo)
constexpr int FU_TILE_BYTES = FU_BM * 128;  // 16 KB
constexpr int FU_MAXL = 3;
constexpr int FU_THREADS = 18 * 32;
constexpr int FU_SIDE = 16;                 // side columns
constexpr int FU_NCB = 24;                  // combine constants
constexpr int FU_MAXN1 = 416;               // widest single layer (two MMA column halves, 416 TMEM columns + a 96-column A ring)

enum { FU_MODE_TOWER = 0, FU_MODE_DEEPFM = 1, FU_MODE_DCN = 2 };

struct FuLayer {
  const float* Wp; const float* bias; const float* alpha;   // alpha: PReLU [N] or NULL
  int N, KC, BN, KP, col, act;                              // KP = N rounded up to 32 (= K extent of the next layer)
  int BNa;                                                  // columns of the first MMA per K step (= BN unless BN > 256:
                                                            // a wide single layer is issued as two column halves)
};

constexpr int FU_MAXG = 16;
struct FuParams {
  const float* arena; const int64_t* row_off; int F; int k4; const void* ids; int ids_kind; int64_t ids_ld;
  int row_stride4;                                         // 16-byte units between consecutive table rows (= k4 for a table;
                                                           // = x_ld / 4 when the 'table' is an activation matrix, rows mode)
  int G; const float* peer[FU_MAXG];                       // row-sharded tables (cfg 5): owner = id mod G, arenas mapped over NVLink;
                                                           // row_off is then [G, F+1].  G = 1: peer[0] = arena
  const int32_t* pack_desc;                                // CTRHOT_IDS_PACKED: per field (byte offset in the record) | (width in bytes) << 16;
                                                           // ids = first record, ids_ld = record bytes (a multiple of 4)
  const float* dense; int64_t dense_ld; int D; int emb;
  int64_t M; int n_tiles; int n_super;
  FuLayer L[FU_MAXL];
  const float* w_out; const float* b_out;
  const float* sq_w;                                       // [KC1*32] internal order or NULL
  int mode; int n_side; float cb[FU_NCB];
  int apply_sigmoid;
  int overlap;                                             // CTRHOT_LAUNCH_OVERLAP: let the next launch in the stream start early
  int ds_col;                                              // TMEM column of the side accumulator
  int a_col, n_as, a_w;                                    // TMEM A ring: first column, slots, columns per slot (x | x_lo = 64
                                                           // with side products, x = 32 without)
  int nr;                                                  // raw-row ring stages in use (<= FU_NR)
  int side_rows;                                           // rows of [s_hi | s_lo] behind W1 in every first-layer chunk (32 or 0)
  float* h_out; int64_t h_ld;                              // single hidden layer: act(x W + b), TF32-rounded, to HBM (no final dot)
  // chained launches (rows of one launch are the input of the next): tile t of the next launch may be read once
  // signal_flags[t] == epoch; a launch in rows mode waits on wait_flags[t] before its first copy of tile t.  Lets launch l + 1
  // start (programmatic dependent launch) while launch l is still finishing its last tiles.
  const int* wait_flags; int* signal_flags; int epoch;
  int nb;                                                  // W ring stages in use (<= FU_NB)
  const float* add_vec; float out_scale, add_scale, add_const;   // tower mode: out = out_scale * y + add_scale * add_vec[m] + add_const
  const float* cc_w; int cc_n;                             // DCN with few cross layers: side products on the CUDA cores (stager),
                                                           // cc_w [KC1*32][4] internal column order, cc_n = n_side + 1 <= 4
  float* side_out;                                         // store mode: the cross term of every row (fed to the last launch)
  // Wide&Deep: the wide logit w0 + sum_j dense_j wd_j + sum_f w[wide_off[f] + id_f] formed inside this kernel: the gather warps
  // fetch the 4-byte weights next to the rows (a 2 KB tail of every raw stage), the stager threads add them up per row
  const float* wide_w; const int64_t* wide_off; const float* wide_dense; int64_t wide_dense_ld; int wide_D; float wide_dw[16];
  uint32_t raw_stride;                                     // bytes per raw stage (16 KB, + 2 KB with the wide tail)
  float* out; int* err_flag;
  uint32_t b_stage_bytes;
};

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      :: "r"(taddr),
         "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
         "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
         "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
         "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15])),
         "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])), "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])),
         "r"(__float_as_uint(v[20])), "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])), "r"(__float_as_uint(v[23])),
         "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])), "r"(__float_as_uint(v[27])),
         "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])), "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
      : "memory");
}

// arrive on the same mbarrier offset in every CTA of `mask` when the issuing thread's prior MMAs complete
__device__ __forceinline__ void mma_commit_mc(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               :: "r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void bulk_g2s_mc(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar, uint16_t mask) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
               :: "r"(dst_smem), "l"(src), "r"(bytes), "r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
  uint32_t r; asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r)); return r;
}
__device__ __forceinline__ uint32_t cluster_nctaid_x() {
  uint32_t r; asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r)); return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void sts128(uint32_t addr, const float4& v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" :: "r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

#ifndef FU_NR_
#define FU_NR_ 4
#endif
constexpr int FU_NR = FU_NR_;               // raw-row ring (gather depth)
constexpr int FU_PF = 3;                    // chunks by which the id loads run ahead of the row copies

// float-region layout (shared by host sizing and device carving); all offsets in floats, 16-byte aligned
struct FuSmemMap {
  int rowoff, sq, bias[FU_MAXL], alpha[FU_MAXL], wout, q, part, ccw, wb, pk, total;
};
__host__ __device__ inline FuSmemMap fu_smem_map(int F /* (owner, field) entries */, int KC1, const int* KP, int nl,
                                                 bool cc = false, bool wide = false, bool packed = false) {
  FuSmemMap m;
  int o = 0;
  m.rowoff = o; o += (F * 3 + 3) / 4 * 4;          // per field: table base address (8 B) x F, then row count (4 B) x F
  m.sq = o; o += KC1 * 32;
  for (int l = 0; l < FU_MAXL; ++l) {
    m.bias[l] = o; if (l < nl) o += KP[l];
    m.alpha[l] = o; if (l < nl) o += KP[l];
  }
  m.wout = o; o += KP[nl - 1];
  m.q = o; o += 2 * FU_BM;
  m.part = o; o += 2 * FU_BM;
  m.ccw = o; if (cc) o += KC1 * 32 * 4;
  m.wb = o; if (wide) o += (3 * F + 3) / 4 * 4;     // per field: address of its block of the wide weight vector (8 B) x F,
                                                    // then its one-hot width (4 B) x F
  m.pk = o; if (packed) o += (F + 3) / 4 * 4;       // per field: packed-id descriptor
  m.total = o;
  return m;
}

__device__ __forceinline__ float fu_act(float t, float a) { return fmaxf(t, 0.f) + a * fminf(t, 0.f); }

// round-to-nearest (ties away) TF32 of a finite value by integer arithmetic on its bit pattern: two full-rate ALU ops
// instead of the NaN-preserving sequence cvt.rna.tf32.f32 compiles to (the epilogues below are issue-bound)
__device__ __forceinline__ float tf32_rna_fast(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}

// one 32-column chunk of a hidden layer's epilogue: acc -> act(acc + b) -> tf32 -> back to TMEM in place
#ifndef FU_CHUNK_INLINE
#define FU_CHUNK_INLINE __noinline__
#endif
__device__ FU_CHUNK_INLINE void fu_hidden_chunk(uint32_t taddr, const float* sb, const float* sa, int c0, int N, bool relu) {
  float acc[32];
  tmem_ld32(taddr, acc);
  if (relu && c0 + 32 <= N) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 b = *reinterpret_cast<const float4*>(sb + c0 + j);
      acc[j + 0] = tf32_rna_fast(fmaxf(acc[j + 0] + b.x, 0.f));
      acc[j + 1] = tf32_rna_fast(fmaxf(acc[j + 1] + b.y, 0.f));
      acc[j + 2] = tf32_rna_fast(fmaxf(acc[j + 2] + b.z, 0.f));
      acc[j + 3] = tf32_rna_fast(fmaxf(acc[j + 3] + b.w, 0.f));
    }
  } else if (c0 + 32 <= N) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 b = *reinterpret_cast<const float4*>(sb + c0 + j);
      const float4 a = *reinterpret_cast<const float4*>(sa + c0 + j);
      acc[j + 0] = tf32_rna_fast(fu_act(acc[j + 0] + b.x, a.x));
      acc[j + 1] = tf32_rna_fast(fu_act(acc[j + 1] + b.y, a.y));
      acc[j + 2] = tf32_rna_fast(fu_act(acc[j + 2] + b.z, a.z));
      acc[j + 3] = tf32_rna_fast(fu_act(acc[j + 3] + b.w, a.w));
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      acc[j] = (c0 + j < N) ? tf32_rna_fast(fu_act(acc[j] + sb[c0 + j], sa[c0 + j])) : 0.f;
  }
  tmem_st32(taddr, acc);
}

// one 32-column chunk of the LAST hidden layer: returns sum_n act(acc + b)[n] * w_out[n]
__device__ FU_CHUNK_INLINE float fu_last_chunk(uint32_t taddr, const float* sb, const float* sa, const float* sw,
                                               int c0, int N, bool relu) {
  float acc[32];
  tmem_ld32(taddr, acc);
  float d0 = 0.f, d1 = 0.f;
  if (relu && c0 + 32 <= N) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 b = *reinterpret_cast<const float4*>(sb + c0 + j);
      const float4 w = *reinterpret_cast<const float4*>(sw + c0 + j);
      d0 = fmaf(fmaxf(acc[j + 0] + b.x, 0.f), w.x, d0);
      d1 = fmaf(fmaxf(acc[j + 1] + b.y, 0.f), w.y, d1);
      d0 = fmaf(fmaxf(acc[j + 2] + b.z, 0.f), w.z, d0);
      d1 = fmaf(fmaxf(acc[j + 3] + b.w, 0.f), w.w, d1);
    }
  } else if (c0 + 32 <= N) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 b = *reinterpret_cast<const float4*>(sb + c0 + j);
      const float4 a = *reinterpret_cast<const float4*>(sa + c0 + j);
      const float4 w = *reinterpret_cast<const float4*>(sw + c0 + j);
      d0 = fmaf(fu_act(acc[j + 0] + b.x, a.x), w.x, d0);
      d1 = fmaf(fu_act(acc[j + 1] + b.y, a.y), w.y, d1);
      d0 = fmaf(fu_act(acc[j + 2] + b.z, a.z), w.z, d0);
      d1 = fmaf(fu_act(acc[j + 3] + b.w, a.w), w.w, d1);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (c0 + j < N) d0 = fmaf(fu_act(acc[j] + sb[c0 + j], sa[c0 + j]), sw[c0 + j], d0);
  }
  return d0 + d1;
}

#ifndef FU_EXP_
#define FU_EXP_ 0         // tuning builds only (results are then wrong): bit 0 drops the N=32 side MMA, bit 1 the N=16 one,
                          // bit 2: stager skips the row LDS, bit 3: skips q, bit 4: skips x_lo,
                          // bit 6: ids from a hash instead of loads, bit 7: rows of tables of <= 64 rows are not fetched
#endif
#ifdef FU_TRACE
// tuning builds only (profiles/build_variants.py TRACE=1): cycles the MMA thread of every CTA spent waiting, per cause
__device__ unsigned long long fu_trace[256][8];
__device__ unsigned long long fu_times[256][4];     // globaltimer ns: CTA entry, MMA loop begin, MMA loop end, CTA exit
// cycles other roles waited: [0] stager on full_raw, [1] stager on a_free, [2] stager on last_free, [3] stager total,
// [4] gather on empty_raw, [5] gather total, [6] epilogue (warp 0) on d_full, [7] epilogue total
__device__ unsigned long long fu_trace2[256][8];
// [0] stager: cycles from a chunk's rows being ready to their last use in registers (LDS + q + x_lo + side sums),
// [1] stager: cycles from the A slot being free to the a_full arrive (tcgen05.st x2 + wait::st + fences), [2] chunks
__device__ unsigned long long fu_trace3[256][8];
#define FU_T2(var, stmt) do { const long long t0_ = clock64(); stmt; var += clock64() - t0_; } while (0)
__device__ __forceinline__ unsigned long long fu_now() {
  unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t;
}
#define FU_TWAIT(slot, stmt) do { const long long t0_ = clock64(); stmt; tw[slot] += clock64() - t0_; } while (0)
#else
#define FU_TWAIT(slot, stmt) do { stmt; } while (0)
#define FU_T2(var, stmt) do { stmt; } while (0)
#endif

template <int NL, int C, bool WIDE>
__global__ void __launch_bounds__(FU_THREADS, 1)
ctr_fused_kernel(const __grid_constant__ FuParams p) {
  extern __shared__ uint8_t smem_raw[];
  constexpr uint32_t RAW_STRIDE = FU_TILE_BYTES + (WIDE ? 2048u : 0u);   // raw stage: 128 x 128 B of rows (+ the wide tail)
#ifdef FU_TRACE
  if (threadIdx.x == 0 && blockIdx.x < 256) fu_times[blockIdx.x][0] = fu_now();
#endif
  // Programmatic dependent launch, opt-in: a following launch that was ALLOWED to (same flag) may start its CTAs on the SMs
  // this launch's CTAs leave -- a batch of 512 tiles on 148 SMs ends with 80 SMs idle for a quarter of the kernel.  Nothing
  // here waits for the previous launch (no griddepcontrol.wait): the caller promises that consecutive launches are independent.
  if (p.overlap && threadIdx.x == 0) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw_addr);
  const uint32_t r_base = base;                                         // FU_NR raw chunks
  const uint32_t b_base = r_base + (uint32_t)p.nr * RAW_STRIDE;       // p.nb weight stages
  const uint32_t f_base = b_base + (uint32_t)p.nb * p.b_stage_bytes;    // float region
  int KP[FU_MAXL];
#pragma unroll
  for (int l = 0; l < FU_MAXL; ++l) KP[l] = (l < NL) ? p.L[l].KP : 0;
  const int KC1 = p.L[0].KC;
  const FuSmemMap mp = fu_smem_map(p.F * p.G, KC1, KP, NL, p.cc_n > 0, WIDE, p.pack_desc != nullptr);
  float* fptr = reinterpret_cast<float*>(base_ptr + (f_base - base));
  const uint32_t bars = f_base + (uint32_t)mp.total * 4u;
  // barrier map
  auto full_raw = [&](int s) { return bars + 8u * s; };
  auto empty_raw = [&](int s) { return bars + 8u * (FU_NR + s); };
  auto a_full = [&](int s) { return bars + 8u * (2 * FU_NR + s); };
  auto a_free = [&](int s) { return bars + 8u * (2 * FU_NR + FU_NAS + s); };
  auto full_b = [&](int s) { return bars + 8u * (2 * FU_NR + 2 * FU_NAS + s); };
  auto empty_b = [&](int s) { return bars + 8u * (2 * FU_NR + 2 * FU_NAS + FU_NB + s); };
  const uint32_t misc = bars + 8u * (2 * FU_NR + 2 * FU_NAS + 2 * FU_NB);
  auto d_full = [&](int l) { return misc + 8u * l; };
  auto h_ready = [&](int l) { return misc + 8u * (FU_MAXL + l); };
  const uint32_t last_free = misc + 8u * (2 * FU_MAXL);
  const uint32_t slot_addr = last_free + 8u;
  // hidden activations are handed to the next layer per 32-column chunk: its MMAs over chunk kc start when the four epilogue
  // warps owning that chunk are done, while the other chunks of the same epilogue are still being converted
  auto h_chunk = [&](int l, int c) { return slot_addr + 8u + 8u * (uint32_t)(l * 8 + c); };
  volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + (slot_addr - base));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t crank = (C > 1) ? cluster_ctarank() : 0u;
  const uint32_t cid = (C > 1) ? cluster_id_x() : blockIdx.x;
  const uint32_t ncl = (C > 1) ? cluster_nctaid_x() : gridDim.x;
  const uint16_t cmask = (uint16_t)((1u << C) - 1u);

  if (tid == 0) {
    for (int s = 0; s < FU_NR; ++s) { mbar_init(full_raw(s), 128); mbar_init(empty_raw(s), 128); }
    for (int s = 0; s < FU_NAS; ++s) { mbar_init(a_full(s), 128); mbar_init(a_free(s), 1); }
    for (int s = 0; s < FU_NB; ++s) { mbar_init(full_b(s), 1); mbar_init(empty_b(s), C); }
    for (int l = 0; l < FU_MAXL; ++l) { mbar_init(d_full(l), 1); mbar_init(h_ready(l), 256); }
    for (int l = 0; l < FU_MAXL; ++l) for (int c = 0; c < 8; ++c) mbar_init(h_chunk(l, c), 128);
    mbar_init(last_free, 256);
    mbar_fence_init();
  }
  // stage the small per-column vectors
  {
    const float** s_fb = reinterpret_cast<const float**>(fptr + mp.rowoff);
    uint32_t* s_fn = reinterpret_cast<uint32_t*>(fptr + mp.rowoff + 2 * p.F * p.G);
    for (int i = tid; i < p.F * p.G; i += FU_THREADS) {       // entry (owner o, field f) at o * F + f
      const int o = i / p.F, f = i - o * p.F;
      const int64_t r0 = p.row_off ? p.row_off[o * (p.F + 1) + f] : 0;      // rows mode: one 'field', M rows, no offsets
      const int64_t nr = p.row_off ? p.row_off[o * (p.F + 1) + f + 1] - r0 : p.M;
      s_fb[i] = p.peer[o] + r0 * (int64_t)(p.row_stride4 * 4);
      s_fn[i] = nr < 0 ? 0u : (nr > 0x7FFFFFFF ? 0x7FFFFFFFu : (uint32_t)nr);   // ids are int32: larger tables cannot be addressed anyway
    }
    for (int i = tid; i < KC1 * 32; i += FU_THREADS) fptr[mp.sq + i] = p.sq_w ? __ldg(p.sq_w + i) : 0.f;
    if (p.cc_n > 0) for (int i = tid; i < KC1 * 32 * 4; i += FU_THREADS) fptr[mp.ccw + i] = __ldg(p.cc_w + i);
    if (p.pack_desc) {
      int32_t* s_pk = reinterpret_cast<int32_t*>(fptr + mp.pk);
      for (int i = tid; i < p.F; i += FU_THREADS) s_pk[i] = __ldg(p.pack_desc + i);
    }
    if (WIDE) {
      const float** s_wb = reinterpret_cast<const float**>(fptr + mp.wb);
      uint32_t* s_ww = reinterpret_cast<uint32_t*>(fptr + mp.wb + 2 * p.F);
      for (int i = tid; i < p.F; i += FU_THREADS) {
        s_wb[i] = p.wide_w + p.wide_off[i];
        // the one-hot block of a field may be narrower than its embedding table (get_dummies keeps only the labels it saw,
        // dataset.py:53-58): an id in [width, vocab) has no column, i.e. contributes 0 -- as wide_onehot_kernel does
        const int64_t wd = p.wide_off[i + 1] - p.wide_off[i];
        s_ww[i] = wd < 0 ? 0u : (wd > 0x7FFFFFFF ? 0x7FFFFFFFu : (uint32_t)wd);
      }
    }
#pragma unroll
    for (int l = 0; l < NL; ++l) {
      const FuLayer& L = p.L[l];
      for (int i = tid; i < L.KP; i += FU_THREADS) {
        fptr[mp.bias[l] + i] = (i < L.N && L.bias) ? __ldg(L.bias + i) : 0.f;
        float a = 0.f;
        if (i < L.N) a = (L.act == CTRHOT_ACT_NONE) ? 1.f : (L.act == CTRHOT_ACT_PRELU ? __ldg(L.alpha + i) : 0.f);
        fptr[mp.alpha[l] + i] = a;
      }
    }
    const FuLayer& LL = p.L[NL - 1];
    for (int i = tid; i < LL.KP; i += FU_THREADS) fptr[mp.wout + i] = (i < LL.N && p.w_out) ? __ldg(p.w_out + i) : 0.f;
  }
  if (warp == 16) { tmem_alloc(slot_addr, 512u); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  if (C > 1) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *slot_ptr;
  const int n_my = ((int)cid < p.n_super) ? (p.n_super - (int)cid + (int)ncl - 1) / (int)ncl : 0;

  if (warp >= 8 && warp < 12) {
    // ================================================================ gather (cp.async, completion tracked by mbarrier)
    // Thread (c, rsub) owns 16-byte piece c of stage rows rsub + 16 i.  The ids of a chunk are fetched FU_PF chunks ahead
    // of the copies that need them (their latency is a DRAM round trip); the copies themselves are never waited for by
    // this thread: cp.async.mbarrier.arrive.noinc hands their completion to full_raw, so FU_NR chunks stay in flight.
    // Two cursors (fetch, issue) walk the same (tile, chunk) sequence; (f, part) advance incrementally -- no divisions.
    const int tg = tid - 256;
    const int c = tg & 7;
    const int rsub = tg >> 3;
    const float* const* s_fb = reinterpret_cast<const float* const*>(fptr + mp.rowoff);
    const uint32_t* s_fn = reinterpret_cast<const uint32_t*>(fptr + mp.rowoff + 2 * p.F * p.G);
    const uint32_t G = (uint32_t)p.G;
    const int total_chunks = n_my * KC1;
    const int k4 = p.k4;
    const int f0 = c / k4, part0 = c - f0 * k4;
    const uint32_t rowbytes = (uint32_t)p.row_stride4 * 16u;
    const uint32_t dst0 = r_base + (uint32_t)rsub * 128u + (uint32_t)((c ^ (rsub & 7)) << 4);   // rows rsub + 16 i: + 2048 i
    const int64_t tile_stride = (int64_t)ncl * C * FU_BM;
    const int64_t first_row0 = ((int64_t)cid * C + crank) * FU_BM;
    struct Cur { int kc, f, part; int64_t row0; };
    auto advance = [&](Cur& u) {
      ++u.kc; u.part += 8;
      while (u.part >= k4) { u.part -= k4; ++u.f; }
      if (u.kc == KC1) { u.kc = 0; u.f = f0; u.part = part0; u.row0 += tile_stride; }
    };
    Cur cf{0, f0, part0, first_row0}, ci = cf;
    int n_fetched = 0;
    long long t2_fetch = 0, t2_issue = 0; (void)t2_fetch; (void)t2_issue;
    auto fetch = [&](int32_t (&reg)[8]) {
      if (n_fetched >= total_chunks) return;
#ifdef FU_TRACE
      const long long tf0 = clock64();
#endif
      ++n_fetched;
      if (cf.f < p.F) {
        if (p.ids == nullptr) {                            // rows mode: row m of the tile IS row m of the matrix
#pragma unroll
          for (int i = 0; i < 8; ++i) reg[i] = (cf.row0 + rsub + 16 * i < p.M) ? (int32_t)(cf.row0 + rsub + 16 * i) : 0;
        } else if (p.ids_kind == CTRHOT_IDS_PACKED) {
          // compact transport records: field f is `w` little-endian bytes at byte `off` of each record (any alignment).
          // Two aligned 32-bit loads + a funnel shift; neighbouring fields share the words, so these hit L1.
          uint32_t bit_off, w8;                                          // byte- or bit-granular field (common.cuh)
          rows_field((uint32_t)reinterpret_cast<const int32_t*>(fptr + mp.pk)[cf.f], bit_off, w8);
          const uint32_t mask = w8 >= 32u ? 0xFFFFFFFFu : ((1u << w8) - 1u);
          const uint8_t* rec = static_cast<const uint8_t*>(p.ids) + (cf.row0 + rsub) * p.ids_ld + (bit_off >> 3);
          const uint32_t bsh = bit_off & 7u;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            reg[i] = 0;
            if (cf.row0 + rsub + 16 * i < p.M) {
              const uintptr_t a = reinterpret_cast<uintptr_t>(rec + (int64_t)(16 * i) * p.ids_ld);
              const uint32_t* al = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
              const uint32_t sh = (uint32_t)(a & 3u) * 8u + bsh;             // <= 31
              const uint32_t lo = __ldg(al);
              const uint32_t hi = (sh + w8 > 32u) ? __ldg(al + 1) : 0u;    // never beyond the record: it ends on a word
              reg[i] = (int32_t)(__funnelshift_r(lo, hi, sh) & mask);
            }
          }
        } else if (FU_EXP_ & 64) {                        // tuning: ids from a hash, no id loads at all (timing only)
#pragma unroll
          for (int i = 0; i < 8; ++i)
            reg[i] = (int32_t)((((uint32_t)(cf.row0 + rsub + 16 * i) * 2654435761u + (uint32_t)cf.f * 40503u) >> 4) % s_fn[cf.f]);
        } else {
          const int32_t* src = static_cast<const int32_t*>(p.ids) + (cf.row0 + rsub) * p.ids_ld + cf.f;   // int32 or float bits
#pragma unroll
          for (int i = 0; i < 8; ++i)                      // rows past M read id 0: a valid row, never consumed
            reg[i] = (cf.row0 + rsub + 16 * i < p.M) ? __ldg(src + (int64_t)(16 * i) * p.ids_ld) : 0;
        }
      }
      advance(cf);
#ifdef FU_TRACE
      t2_fetch += clock64() - tf0;
#endif
    };
    int rs = 0; uint32_t rph = 0;                        // raw stage of the next issue and its phase
    constexpr bool wide_on = WIDE;
    long long t2_wait = 0; (void)t2_wait;
    auto issue = [&](int g, const int32_t (&reg)[8]) {
      const int s = rs;
      FU_T2(t2_wait, mbar_wait(empty_raw(s), rph ^ 1u));
#ifdef FU_TRACE
      const long long ti0 = clock64();
#endif
      if (++rs == p.nr) { rs = 0; rph ^= 1u; }
      if (p.wait_flags != nullptr && ci.kc == 0 && ci.row0 < p.M) {
        // first chunk of a tile whose rows the previous launch writes: wait until that launch has published the tile
        const int* fl = p.wait_flags + ci.row0 / FU_BM;
        int v;
        do { asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(fl) : "memory"); } while (v != p.epoch);
      }
      const uint32_t dst = dst0 + (uint32_t)s * RAW_STRIDE;
      // wide tail of the stage: [128 rows][4 slots] floats; this thread owns a slot when its piece is the first of a field
      bool wide_here = false;
      uint32_t wdst = 0;
      const float* wbase = nullptr;
      uint32_t wwidth = 0;
      if (wide_on && ci.f < p.F && ci.part == 0) {
        wide_here = true;
        wwidth = reinterpret_cast<const uint32_t*>(fptr + mp.wb + 2 * p.F)[ci.f];
        wdst = r_base + (uint32_t)s * RAW_STRIDE + FU_TILE_BYTES + (uint32_t)rsub * 16u + (k4 < 8 ? (uint32_t)(c / k4) * 4u : 0u);
        wbase = reinterpret_cast<const float* const*>(fptr + mp.wb)[ci.f];
      }
      if ((FU_EXP_ & 128) && ci.f < p.F && s_fn[ci.f] <= 64u) {
        // tuning: tiny tables cost nothing (as if their rows came from a shared-memory copy)
      } else if (ci.f < p.F) {
        const char* fb = reinterpret_cast<const char*>(s_fb[ci.f]) + ci.part * 16;
        const uint32_t nrow = s_fn[ci.f];
        bool bad = false;
        if (G > 1) {
          // row-sharded: the row lives in arena (id mod G) at local row (id div G); that arena may be a peer's HBM mapped
          // over NVLink -- the same 16-byte LDGSTS, only the latency differs (the raw ring keeps FU_NR chunks in flight)
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t id = (p.ids_kind != CTRHOT_IDS_F32) ? (uint32_t)reg[i]
                                                               : (uint32_t)__float2int_rz(__int_as_float(reg[i]));
            const uint32_t lr = id / G, o = id - lr * G;
            const int e = (int)o * p.F + ci.f;
            const bool ok = (int32_t)id >= 0 && lr < s_fn[e];
            bad |= !ok;
            const char* src = reinterpret_cast<const char*>(s_fb[e]) + ci.part * 16 + (uint64_t)lr * rowbytes;
            cp_async16(dst + 2048u * i, ok ? (const void*)src : (const void*)p.arena, ok ? 16u : 0u);
          }
        } else if (p.ids_kind != CTRHOT_IDS_F32) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t id = (uint32_t)reg[i];
            const bool ok = id < nrow;
            bad |= !ok;
            cp_async16(dst + 2048u * i, ok ? (const void*)(fb + (uint64_t)id * rowbytes) : (const void*)p.arena, ok ? 16u : 0u);
            if (wide_here) {
              const bool wok = ok && id < wwidth;
              cp_async4(wdst + 256u * i, wok ? (const void*)(wbase + id) : (const void*)p.arena, wok ? 4u : 0u);
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t id = (uint32_t)__float2int_rz(__int_as_float(reg[i]));
            const bool ok = id < nrow;
            bad |= !ok;
            cp_async16(dst + 2048u * i, ok ? (const void*)(fb + (uint64_t)id * rowbytes) : (const void*)p.arena, ok ? 16u : 0u);
            if (wide_here) {
              const bool wok = ok && id < wwidth;
              cp_async4(wdst + 256u * i, wok ? (const void*)(wbase + id) : (const void*)p.arena, wok ? 4u : 0u);
            }
          }
        }
        if (bad && p.err_flag) *p.err_flag = 1;
      } else {                                         // dense tail [emb, emb + D) then zero padding: 4-byte copies
        const int j = ((ci.f - p.F) * k4 + ci.part) * 4;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int64_t gm = ci.row0 + rsub + 16 * i;
          const bool live = gm < p.M;
          const float* src = p.dense + (live ? gm : 0) * p.dense_ld + j;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const bool ok = live && (j + e < p.D);
            cp_async4(dst + 2048u * i + 4u * e, ok ? (const void*)(src + e) : (const void*)p.arena, ok ? 4u : 0u);
          }
        }
      }
      cp_async_mbar_arrive_noinc(full_raw(s));
      advance(ci);
#ifdef FU_TRACE
      t2_issue += clock64() - ti0;
#endif
    };
    static_assert(FU_PF == 3, "the loop below is unrolled for three id register sets");
#ifdef FU_TRACE
    const long long t2_begin = clock64();
#endif
    int32_t ra[8], rb[8], rc[8];
    fetch(ra); fetch(rb); fetch(rc);
    for (int g = 0; g < total_chunks; g += FU_PF) {
      issue(g, ra); fetch(ra);
      if (g + 1 < total_chunks) { issue(g + 1, rb); fetch(rb); }
      if (g + 2 < total_chunks) { issue(g + 2, rc); fetch(rc); }
    }
#ifdef FU_TRACE
    if (tg == 0 && blockIdx.x < 256) {
      fu_trace2[blockIdx.x][4] = t2_wait; fu_trace2[blockIdx.x][5] = clock64() - t2_begin;
      fu_trace3[blockIdx.x][3] = t2_fetch; fu_trace3[blockIdx.x][4] = t2_issue;
    }
#endif
  } else if (warp >= 12 && warp < 16) {
    // ================================================================ stager: smem chunk -> TMEM A ring (x | x_lo), q = sum x^2 sq_w
    const int row = tid - 384;                          // = TMEM lane; warps 12..15 own lane quarters 0..3
    const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)p.a_col;
    const uint32_t src0 = r_base + (uint32_t)row * 128u;
    const uint32_t rx = (uint32_t)(row & 7);
    const int n_as = p.n_as;
    uint32_t sa = 0, pa = 0;                            // A slot and its phase
    int sr = 0; uint32_t pr = 0;                        // raw stage and its phase
    const bool with_lo = (p.a_w == 64);
    long long s2_raw = 0, s2_af = 0, s2_lf = 0, s2_math = 0, s2_store = 0;
    (void)s2_raw; (void)s2_af; (void)s2_lf; (void)s2_math; (void)s2_store;
#ifdef FU_TRACE
    const long long s2_begin = clock64();
#endif
    // (Loading the rows of chunk g + 1 while chunk g is processed was built and measured SLOWER, 76.5 -> 81.8 us per step:
    //  the stager is not the pace-setter -- the gather is, at ~1 700 cycles per chunk against 936 of MMA work and ~690 here;
    //  profiles/r02_fused_notes.md.)
    for (int lt = 0; lt < n_my; ++lt) {
      float q = 0.f;
      float S0 = 0.f, S1 = 0.f, S2 = 0.f, S3 = 0.f;     // CUDA-core side products (DCN: x . w_l and x . final_w)
      float wsum = 0.f;                                 // Wide&Deep: wide logit of this row (without w0)
      if (WIDE && p.wide_D > 0) {
        const int64_t gm = (((int64_t)cid + (int64_t)lt * ncl) * C + crank) * FU_BM + row;
        if (gm < p.M) {
          const float* dr = p.wide_dense + gm * p.wide_dense_ld;
#pragma unroll
          for (int j = 0; j < 16; ++j) if (j < p.wide_D) wsum = fmaf(__ldg(dr + j), p.wide_dw[j], wsum);
        }
      }
      for (int kc = 0; kc < KC1; ++kc) {
        FU_T2(s2_raw, mbar_wait(full_raw(sr), pr));
#ifdef FU_TRACE
        const long long s2_t0 = clock64();
#endif
        const uint32_t raw = src0 + sr * RAW_STRIDE;
        float x[32];
#pragma unroll
        for (int cc = 0; cc < 8; ++cc) {
          const float4 v = (FU_EXP_ & 4) ? make_float4(0.25f, 0.5f, 0.125f, 1.f) : lds128(raw + (((uint32_t)cc ^ rx) << 4));
          x[4 * cc + 0] = v.x; x[4 * cc + 1] = v.y; x[4 * cc + 2] = v.z; x[4 * cc + 3] = v.w;
        }
        const float* sq = fptr + mp.sq + kc * 32;
        if (p.sq_w && !(FU_EXP_ & 8)) {                 // FM's sum x^2 (sum_c v^2) term
#pragma unroll
          for (int cc = 0; cc < 8; ++cc) {
            const float4 w = *reinterpret_cast<const float4*>(sq + cc * 4);
            q = fmaf(x[4 * cc + 0] * x[4 * cc + 0], w.x, q); q = fmaf(x[4 * cc + 1] * x[4 * cc + 1], w.y, q);
            q = fmaf(x[4 * cc + 2] * x[4 * cc + 2], w.z, q); q = fmaf(x[4 * cc + 3] * x[4 * cc + 3], w.w, q);
          }
        }
        if (WIDE) {
          // the fields whose first piece lies in this chunk left their wide weight in the stage tail, in field order
          const float4 wv = lds128(r_base + sr * RAW_STRIDE + FU_TILE_BYTES + (uint32_t)row * 16u);
          const int k4 = p.k4;
          const int piece0 = kc * 8;
          if (k4 >= 8) {
            if ((piece0 & (k4 - 1)) == 0 && piece0 < p.F * k4) wsum += wv.x;
          } else {
            const int nf = 8 / k4, f0 = piece0 / k4;
            if (f0 < p.F) wsum += wv.x;
            if (nf > 1 && f0 + 1 < p.F) wsum += wv.y;
            if (nf > 2 && f0 + 2 < p.F) wsum += wv.z;
            if (nf > 2 && f0 + 3 < p.F) wsum += wv.w;
          }
        }
        if (p.cc_n > 0) {                               // exact fp32: the weights are read as warp-wide broadcasts
          const float4* cw = reinterpret_cast<const float4*>(fptr + mp.ccw) + kc * 32;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float4 w = cw[j];
            S0 = fmaf(x[j], w.x, S0); S1 = fmaf(x[j], w.y, S1); S2 = fmaf(x[j], w.z, S2); S3 = fmaf(x[j], w.w, S3);
          }
        }
        if (kc == KC1 - 1) {
          // per-row scalar for the epilogue, published before the arrive that lets the last MMAs of the tile go
          float v = WIDE ? wsum : q;
          if (p.cc_n > 0) {
            const float S[4] = {S0, S1, S2, S3};
            float a = 1.f, fin = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (j < p.n_side) a += fmaf(a, S[j], p.cb[j]);
              else if (j == p.n_side) fin = S[j];
            }
            v = fmaf(a, fin, p.cb[p.n_side]);
          }
          fptr[mp.q + (lt & 1) * FU_BM + row] = v;
        }
#ifdef FU_TRACE
        s2_math += clock64() - s2_t0;
#endif
        mbar_arrive(empty_raw(sr));                     // the chunk is in registers: the gather may refill the stage
        if (++sr == p.nr) { sr = 0; pr ^= 1u; }
        // the A ring shares its TMEM columns with the later layers' accumulators: wait until the previous tile is done with them
        if (NL > 1 && kc == 0 && lt > 0) FU_T2(s2_lf, mbar_wait(last_free, (uint32_t)(lt - 1) & 1u));
        FU_T2(s2_af, mbar_wait(a_free(sa), pa ^ 1u));
#ifdef FU_TRACE
        const long long s2_t1 = clock64();
#endif
        tc_fence_after();
        tmem_st32(t_lane + sa * (uint32_t)p.a_w, x);
        if (with_lo) {
          // x_lo = x - trunc_tf32(x): what the tensor pipe drops of x (it truncates its operands to TF32)
#pragma unroll
          if (!(FU_EXP_ & 16)) for (int j = 0; j < 32; ++j) x[j] = x[j] - trunc_tf32(x[j]);
          tmem_st32(t_lane + sa * 64u + 32u, x);
        }
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(a_full(sa));
#ifdef FU_TRACE
        s2_store += clock64() - s2_t1;
#endif
        if (++sa == (uint32_t)n_as) { sa = 0; pa ^= 1u; }
      }
    }
#ifdef FU_TRACE
    if (row == 0 && blockIdx.x < 256) {
      fu_trace2[blockIdx.x][0] = s2_raw; fu_trace2[blockIdx.x][1] = s2_af; fu_trace2[blockIdx.x][2] = s2_lf;
      fu_trace2[blockIdx.x][3] = clock64() - s2_begin;
      fu_trace3[blockIdx.x][0] = s2_math; fu_trace3[blockIdx.x][1] = s2_store; fu_trace3[blockIdx.x][2] = (long long)n_my * KC1;
    }
#endif
  } else if (warp == 16) {
    // ================================================================ MMA issuer: the warp stays converged, one elected lane issues
    {
#ifdef FU_TRACE
      long long tw[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      const long long t_begin = clock64();
      if (lane == 0 && blockIdx.x < 256) fu_times[blockIdx.x][1] = fu_now();
#endif
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0;     // A slot / W stage and their phases
      uint32_t hcount[FU_MAXL] = {0, 0, 0};
      const uint32_t ds_t = tmem_base + (uint32_t)p.ds_col;
      const uint32_t a_t0 = tmem_base + (uint32_t)p.a_col;
      for (int lt = 0; lt < n_my; ++lt) {
        {
          const FuLayer& L = p.L[0];
          const uint32_t idesc = make_idesc_tf32(FU_BM, (uint32_t)L.BNa);
          const uint32_t idesc_b = make_idesc_tf32(FU_BM, (uint32_t)(L.BN - L.BNa));      // second column half (wide layer)
          const bool two = L.BN > L.BNa;
          const uint32_t idesc32 = make_idesc_tf32(FU_BM, 32u), idesc16 = make_idesc_tf32(FU_BM, 16u);
          const uint32_t d_t = tmem_base + (uint32_t)L.col;
          const bool side_on = (p.side_rows > 0);              // tensor-pipe side products (none for a plain tower / CUDA-core sides)
          if (NL == 1 && lt > 0) { FU_TWAIT(4, mbar_wait(last_free, (uint32_t)(lt - 1) & 1u)); tc_fence_after(); }
          for (int kc = 0; kc < L.KC; ++kc) {
            FU_TWAIT(0, mbar_wait(full_b(sb), pb));
            FU_TWAIT(1, mbar_wait(a_full(sa), pa));
            tc_fence_after();
            const uint32_t a_hi = a_t0 + sa * (uint32_t)p.a_w, a_lo = a_hi + 32u;
            const uint64_t db = make_desc_sw128(b_base + sb * p.b_stage_bytes);
            const uint64_t db2 = make_desc_sw128(b_base + sb * p.b_stage_bytes + (uint32_t)L.BNa * 128u);
            const uint64_t ds = make_desc_sw128(b_base + sb * p.b_stage_bytes + (uint32_t)L.BN * 128u);
            if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              mma_tf32_ts(d_t, a_hi + (uint32_t)(8 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kc | k) != 0));
              if (two) mma_tf32_ts(d_t + (uint32_t)L.BNa, a_hi + (uint32_t)(8 * k), db2 + (uint64_t)(2 * k), idesc_b,
                                   (uint32_t)((kc | k) != 0));
              // (folding the side columns into the main MMA as two N = 144 halves was measured SLOWER, 88.9 -> 92.5 us:
              //  an MMA wider than 128 columns costs the full 138 cycles; kept in git history only)
              if (side_on && !(FU_EXP_ & 1))
                mma_tf32_ts(ds_t, a_hi + (uint32_t)(8 * k), ds + (uint64_t)(2 * k), idesc32, (uint32_t)((kc | k) != 0));
            }
            if (side_on && !(FU_EXP_ & 2)) {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                mma_tf32_ts(ds_t, a_lo + (uint32_t)(8 * k), ds + (uint64_t)(2 * k), idesc16, 1u);
            }
            mma_commit(a_free(sa));
            if (C > 1) mma_commit_mc(empty_b(sb), cmask); else mma_commit(empty_b(sb));
            }
            __syncwarp();
            if (++sa == (uint32_t)p.n_as) { sa = 0; pa ^= 1u; }
            if (++sb == (uint32_t)p.nb) { sb = 0; pb ^= 1u; }
          }
          if (elect_one()) mma_commit(d_full(0));
          __syncwarp();
        }
#pragma unroll
        for (int l = 1; l < NL; ++l) {
          const FuLayer& L = p.L[l];
          const uint32_t idesc = make_idesc_tf32(FU_BM, (uint32_t)L.BN);
          const uint32_t d_t = tmem_base + (uint32_t)L.col;
          const uint32_t a_t = tmem_base + (uint32_t)p.L[l - 1].col;
          const uint32_t hph = hcount[l - 1] & 1u; ++hcount[l - 1];
          if (l == NL - 1 && lt > 0) FU_TWAIT(4, mbar_wait(last_free, (uint32_t)(lt - 1) & 1u));
          for (int kc = 0; kc < L.KC; ++kc) {
            FU_TWAIT(3, mbar_wait(h_chunk(l - 1, kc), hph));       // columns [32 kc, 32 kc + 32) of the previous layer are in TMEM
            FU_TWAIT(5, mbar_wait(full_b(sb), pb));
            tc_fence_after();
            const uint64_t db = make_desc_sw128(b_base + sb * p.b_stage_bytes);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                mma_tf32_ts(d_t, a_t + (uint32_t)(kc * 32 + k * 8), db + (uint64_t)(2 * k), idesc, (uint32_t)((kc | k) != 0));
              if (C > 1) mma_commit_mc(empty_b(sb), cmask); else mma_commit(empty_b(sb));
            }
            __syncwarp();
            if (++sb == (uint32_t)p.nb) { sb = 0; pb ^= 1u; }
          }
          if (elect_one()) mma_commit(d_full(l));
          __syncwarp();
        }
      }
#ifdef FU_TRACE
      tw[7] = clock64() - t_begin;
      tw[6] = n_my;
      if (lane == 0 && blockIdx.x < 256) fu_times[blockIdx.x][2] = fu_now();
      if (lane == 0 && blockIdx.x < 256) for (int i = 0; i < 8; ++i) fu_trace[blockIdx.x][i] = (unsigned long long)tw[i];
#endif
    }
    __syncwarp();
  } else if (warp == 17) {
    // ================================================================ W producer (bulk async copies, cluster multicast)
    // (converged warp, one elected lane issues: a bulk copy under a divergent `lane == 0` branch compiles to a waterfall loop)
    {
      uint32_t sb = 0, pb = 0;
      for (int lt = 0; lt < n_my; ++lt) {
#pragma unroll
        for (int l = 0; l < NL; ++l) {
          const FuLayer& L = p.L[l];
          const uint32_t bytes = (uint32_t)(L.BN + (l == 0 ? p.side_rows : 0)) * 128u;
          const size_t stride = (size_t)(L.BN + (l == 0 ? 2 * FU_SIDE : 0)) * 128u;      // the pack always carries the side rows
          const uint32_t piece = bytes / C;
          for (int kc = 0; kc < L.KC; ++kc) {
            mbar_wait(empty_b(sb), pb ^ 1u);
            const uint8_t* src = reinterpret_cast<const uint8_t*>(L.Wp) + (size_t)kc * stride + (size_t)crank * piece;
            const uint32_t dst = b_base + sb * p.b_stage_bytes + crank * piece;
            if (elect_one()) {
              mbar_arrive_expect_tx(full_b(sb), bytes);
              if (C > 1) bulk_g2s_mc(dst, src, piece, full_b(sb), cmask);
              else bulk_g2s(dst, src, bytes, full_b(sb));
            }
            __syncwarp();
            if (++sb == (uint32_t)p.nb) { sb = 0; pb ^= 1u; }
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ================================================================ epilogue (warps 0-7)
    const int q4 = warp & 3, half = warp >> 2;
    const int row = q4 * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
    uint32_t dcount[FU_MAXL] = {0, 0, 0};
    long long e2_wait = 0; (void)e2_wait;
#ifdef FU_TRACE
    const long long e2_begin = clock64();
#endif
    for (int lt = 0; lt < n_my; ++lt) {
      const int64_t gm = (((int64_t)cid + (int64_t)lt * ncl) * C + crank) * FU_BM + row;
      float side = 0.f;
#pragma unroll
      for (int l = 0; l < NL; ++l) {
        const FuLayer& L = p.L[l];
        FU_T2(e2_wait, mbar_wait(d_full(l), dcount[l] & 1u)); ++dcount[l];
        tc_fence_after();
        if (l == 0 && half == 0 && p.cc_n > 0) {
          side = fptr[mp.q + (lt & 1) * FU_BM + row];
          if (p.side_out && gm < p.M) p.side_out[gm] = side;
        } else if (l == 0 && half == 0 && p.mode != FU_MODE_TOWER) {
          float S[32];
          tmem_ld32(t_lane + (uint32_t)p.ds_col, S);
#pragma unroll
          for (int j = 0; j < FU_SIDE; ++j) S[j] += S[FU_SIDE + j];
          if (p.mode == FU_MODE_DEEPFM) {
            float sq = 0.f, lin = 0.f;
#pragma unroll
            for (int j = 0; j < FU_SIDE; ++j) {
              if (j < p.n_side) sq = fmaf(S[j], S[j], sq);
              else if (j == p.n_side) lin = S[j];
            }
            side = p.cb[0] + lin + 0.5f * (sq - fptr[mp.q + (lt & 1) * FU_BM + row]);
          } else {
            float a = 1.f, fin = 0.f;
#pragma unroll
            for (int j = 0; j < FU_SIDE; ++j) {
              if (j < p.n_side) a += fmaf(a, S[j], p.cb[j]);
              else if (j == p.n_side) fin = S[j];
            }
            side = fmaf(a, fin, p.cb[p.n_side]);
          }
        }
        const float* sb = fptr + mp.bias[l];
        const float* sa = fptr + mp.alpha[l];
        if (l < NL - 1) {
          for (int c0 = half * 32; c0 < L.KP; c0 += 64) {
            fu_hidden_chunk(t_lane + (uint32_t)(L.col + c0), sb, sa, c0, L.N, L.act == CTRHOT_ACT_RELU);
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(h_chunk(l, c0 >> 5));
          }
        } else if (p.h_out) {
          const bool h_vec8 = ((reinterpret_cast<uintptr_t>(p.h_out) & 31) == 0) && ((p.h_ld & 7) == 0);
          // single wide layer: the activations go to HBM (TF32-rounded, so the next layer's MMA reads them exactly)
          for (int c0 = half * 32; c0 < L.KP; c0 += 64) {
            float acc[32];
            tmem_ld32(t_lane + (uint32_t)(L.col + c0), acc);
            if (gm < p.M && h_vec8) {
              // 32 bytes per store: every instruction fills whole sectors (half as many LSU transactions as 16-byte stores;
              // this epilogue holds up the next tile's MMAs, the accumulator is single-buffered at this width)
              float* dst = p.h_out + gm * p.h_ld + c0;
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                float o[8];
                const float4 b0 = *reinterpret_cast<const float4*>(sb + c0 + j), b1 = *reinterpret_cast<const float4*>(sb + c0 + j + 4);
                if (L.act == CTRHOT_ACT_RELU) {
                  o[0] = fmaxf(acc[j + 0] + b0.x, 0.f); o[1] = fmaxf(acc[j + 1] + b0.y, 0.f);
                  o[2] = fmaxf(acc[j + 2] + b0.z, 0.f); o[3] = fmaxf(acc[j + 3] + b0.w, 0.f);
                  o[4] = fmaxf(acc[j + 4] + b1.x, 0.f); o[5] = fmaxf(acc[j + 5] + b1.y, 0.f);
                  o[6] = fmaxf(acc[j + 6] + b1.z, 0.f); o[7] = fmaxf(acc[j + 7] + b1.w, 0.f);
                } else {
                  const float4 a0 = *reinterpret_cast<const float4*>(sa + c0 + j), a1 = *reinterpret_cast<const float4*>(sa + c0 + j + 4);
                  o[0] = fu_act(acc[j + 0] + b0.x, a0.x); o[1] = fu_act(acc[j + 1] + b0.y, a0.y);
                  o[2] = fu_act(acc[j + 2] + b0.z, a0.z); o[3] = fu_act(acc[j + 3] + b0.w, a0.w);
                  o[4] = fu_act(acc[j + 4] + b1.x, a1.x); o[5] = fu_act(acc[j + 5] + b1.y, a1.y);
                  o[6] = fu_act(acc[j + 6] + b1.z, a1.z); o[7] = fu_act(acc[j + 7] + b1.w, a1.w);
                }
#pragma unroll
                for (int e = 0; e < 8; ++e) o[e] = tf32_rna_fast(o[e]);
                if (c0 + j < p.h_ld)
                  asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                               :: "l"(dst + j), "f"(o[0]), "f"(o[1]), "f"(o[2]), "f"(o[3]), "f"(o[4]), "f"(o[5]), "f"(o[6]),
                                  "f"(o[7]) : "memory");
              }
            } else if (gm < p.M) {
              float* dst = p.h_out + gm * p.h_ld + c0;
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                const float4 b = *reinterpret_cast<const float4*>(sb + c0 + j);
                const float4 a = *reinterpret_cast<const float4*>(sa + c0 + j);
                float4 o;
                o.x = tf32_rna_fast(fu_act(acc[j + 0] + b.x, a.x)); o.y = tf32_rna_fast(fu_act(acc[j + 1] + b.y, a.y));
                o.z = tf32_rna_fast(fu_act(acc[j + 2] + b.z, a.z)); o.w = tf32_rna_fast(fu_act(acc[j + 3] + b.w, a.w));
                if (c0 + j < p.h_ld) stg_na_f4(dst + j, o);       // columns in [N, h_ld) are written as zeros
              }
            }
          }
          if (p.signal_flags != nullptr) {
            // publish the tile: every epilogue thread's stores, then one release store of the epoch
            __threadfence();
            named_bar_sync(7, 256);
            if (tid == 0) {
              int* fl = p.signal_flags + (((int64_t)cid + (int64_t)lt * ncl) * C + crank);
              asm volatile("st.release.gpu.global.s32 [%0], %1;" :: "l"(fl), "r"(p.epoch) : "memory");
            }
          }
          tc_fence_before();
          mbar_arrive(last_free);
        } else {
          float dot = 0.f;
          for (int c0 = half * 32; c0 < L.KP; c0 += 64)
            dot += fu_last_chunk(t_lane + (uint32_t)(L.col + c0), sb, sa, fptr + mp.wout, c0, L.N, L.act == CTRHOT_ACT_RELU);
          tc_fence_before();
          mbar_arrive(last_free);
          float* part = fptr + mp.part + (lt & 1) * FU_BM;
          if (half == 1) part[row] = dot;
          named_bar_sync(1 + q4, 64);
          if (half == 0 && gm < p.M) {
            const float y = dot + part[row] + (p.b_out ? __ldg(p.b_out) : 0.f);
            float r;
            if (p.mode == FU_MODE_DEEPFM) r = 0.5f * (side + y);
            else if (p.mode == FU_MODE_DCN) r = side + fmaf(y, p.cb[p.n_side + 1], p.cb[p.n_side + 2]);
            else if (WIDE) r = fmaf(p.out_scale, y, fmaf(p.add_scale, fptr[mp.q + (lt & 1) * FU_BM + row], p.add_const));
            else r = fmaf(p.out_scale, y, fmaf(p.add_scale, p.add_vec ? __ldg(p.add_vec + gm) : 0.f, p.add_const));
            p.out[gm] = p.apply_sigmoid ? 1.0f / (1.0f + __expf(-r)) : r;
          }
        }
      }
    }
#ifdef FU_TRACE
    if (tid == 0 && blockIdx.x < 256) { fu_trace2[blockIdx.x][6] = e2_wait; fu_trace2[blockIdx.x][7] = clock64() - e2_begin; }
#endif
  }
  tc_fence_before();
  __syncthreads();
  if (C > 1) cluster_sync_all();
  if (warp == 16) { tc_fence_after(); tmem_dealloc(tmem_base, 512u); }
#ifdef FU_TRACE
  if (threadIdx.x == 0 && blockIdx.x < 256) fu_times[blockIdx.x][3] = fu_now();
#endif
}

// ---------------------------------------------------------------- first-layer pack: [KC1][(BN1 + 32) rows][32 cols]
// rows [0, BN1)        = W1[:, n]          (tf32, round to nearest)
// rows [BN1, BN1+16)   = trunc_tf32(side[:, c])             "s_hi"
// rows [BN1+16, BN1+32)= side[:, c] - s_hi                  "s_lo"   (the MMA keeps its leading 11 bits)
// Columns are the INTERNAL order [embeddings | dense D | zero pad]; W1 / side arrive in the reference order
// [dense D | embeddings] (deepFM.py:26, dcn.py:27).  Each 32-column chunk is stored 128B-swizzled, K-major.
__global__ void fused_pack_kernel(const float* __restrict__ W1, const float* __restrict__ side, int n, int N1, int D,
                                  int KC1, int BN1, float* __restrict__ Wp) {
  const int rows = BN1 + 2 * FU_SIDE;
  const int64_t total = (int64_t)KC1 * rows * 32;
  const int emb = n - D;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % 32);
    const int r = (int)((i / 32) % rows);
    const int kc = (int)(i / ((int64_t)32 * rows));
    const int j = kc * 32 + kk;                                  // internal column
    float v = 0.f;
    if (j < n) {
      const int jr = (j < emb) ? (D + j) : (j - emb);            // reference column
      if (r < BN1) {
        if (r < N1) v = to_tf32_rna(W1[(size_t)jr * N1 + r]);
      } else if (side != nullptr) {
        const int c = (r - BN1) % FU_SIDE;
        const float s = side[(size_t)jr * FU_SIDE + c];
        const float hi = __uint_as_float(__float_as_uint(s) & 0xFFFFE000u);
        v = (r - BN1 < FU_SIDE) ? hi : (s - hi);
      }
    }
    Wp[(size_t)kc * rows * 32 + (sw128_offset((uint32_t)r, (uint32_t)(kk >> 2)) >> 2) + (kk & 3)] = v;
  }
}

}  // namespace ctrhot

using namespace ctrhot;

static inline int fu_round(int v, int m) { return (v + m - 1) / m * m; }

extern "C" int64_t ctrhot_fused_pack_bytes(int n, int n1) {
  if (n <= 0 || n1 <= 0 || n1 > FU_MAXN1) return CTRHOT_EARG;
  const int KC1 = (n + 31) / 32, BN1 = fu_round(n1, 16);
  return (int64_t)KC1 * (BN1 + 2 * FU_SIDE) * 32 * sizeof(float);
}

extern "C" int ctrhot_fused_pack(const float* W1, const float* side, int n, int n1, int D, void* packed, void* stream) {
  CTRHOT_RANGE("ctrhot_fused_pack");
  if (!W1 || !packed || n <= 0 || n1 <= 0 || n1 > FU_MAXN1 || D < 0 || D > n || !aligned16(packed)) return CTRHOT_EARG;
  const int KC1 = (n + 31) / 32, BN1 = fu_round(n1, 16);
  const int64_t total = (int64_t)KC1 * (BN1 + 2 * FU_SIDE) * 32;
  int64_t grid = (total + 255) / 256;
  if (grid > kNumSMs * 16) grid = kNumSMs * 16;
  fused_pack_kernel<<<(int)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(W1, side, n, n1, D, KC1, BN1,
                                                                             static_cast<float*>(packed));
  CTRHOT_LAUNCH_OK("ctrhot_fused_pack");
  return CTRHOT_OK;
}

#ifdef FU_TRACE
extern "C" int ctrhot_debug_fused_trace(unsigned long long* h_out) {
  CTRHOT_RANGE("ctrhot_debug_fused_trace");
  return cudaMemcpyFromSymbol(h_out, ctrhot::fu_trace, sizeof(unsigned long long) * 256 * 8) == cudaSuccess ? 0 : -3;
}
extern "C" int ctrhot_debug_fused_trace2(unsigned long long* h_out) {
  CTRHOT_RANGE("ctrhot_debug_fused_trace2");
  return cudaMemcpyFromSymbol(h_out, ctrhot::fu_trace2, sizeof(unsigned long long) * 256 * 8) == cudaSuccess ? 0 : -3;
}
extern "C" int ctrhot_debug_fused_trace3(unsigned long long* h_out) {
  CTRHOT_RANGE("ctrhot_debug_fused_trace3");
  return cudaMemcpyFromSymbol(h_out, ctrhot::fu_trace3, sizeof(unsigned long long) * 256 * 8) == cudaSuccess ? 0 : -3;
}
extern "C" int ctrhot_debug_fused_times(unsigned long long* h_out) {
  CTRHOT_RANGE("ctrhot_debug_fused_times");
  return cudaMemcpyFromSymbol(h_out, ctrhot::fu_times, sizeof(unsigned long long) * 256 * 4) == cudaSuccess ? 0 : -3;
}
#endif

namespace {
// Resident clusters of C CTAs for this kernel variant (a cluster must fit inside one GPC: 33 clusters of 4 = 132 CTAs,
// 74 of 2 = 148 on B200).  The occupancy query costs host time, so it is made once per (variant, smem size).
template <int NL, int C, bool WIDE>
int fu_max_clusters(size_t smem) {
  static size_t cached_smem = ~(size_t)0;
  static int cached_n = 0;
  if (cached_smem == smem && cached_n > 0) return cached_n;
  auto kern = ctr_fused_kernel<NL, C, WIDE>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    (void)cudaGetLastError();
    return kNumSMs / C;
  }
  int n = kNumSMs / C;
  if (C > 1) {
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(FU_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.gridDim = dim3(kNumSMs / C * C);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int q = 0;
    if (cudaOccupancyMaxActiveClusters(&q, kern, &cfg) == cudaSuccess && q > 0) n = q;
    else (void)cudaGetLastError();
  }
  cached_smem = smem;
  cached_n = n;
  return n;
}

template <int NL, int C, bool WIDE>
int fu_launch(const FuParams& p0, size_t smem, cudaStream_t st) {
  FuParams p = p0;
  auto kern = ctr_fused_kernel<NL, C, WIDE>;
  const int max_clusters = fu_max_clusters<NL, C, WIDE>(smem);      // also sets the dynamic shared memory attribute
  p.n_super = (p.n_tiles + C - 1) / C;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(FU_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (p.overlap) {
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.numAttrs = 2;
  }
  const int n_clusters = p.n_super < max_clusters ? p.n_super : max_clusters;
  cfg.gridDim = dim3(n_clusters * C);
  CTRHOT_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, p), "ctrhot_ctr_fused_tf32 launch");
  return CTRHOT_OK;
}

template <int C>
int fu_max_clusters_nl(int nl, size_t smem, bool wide) {
  switch (nl) {
    case 1: return wide ? fu_max_clusters<1, C, true>(smem) : fu_max_clusters<1, C, false>(smem);
    case 2: return wide ? fu_max_clusters<2, C, true>(smem) : fu_max_clusters<2, C, false>(smem);
    default: return wide ? fu_max_clusters<3, C, true>(smem) : fu_max_clusters<3, C, false>(smem);
  }
}

template <int C>
int fu_launch_nl(const FuParams& p, int nl, size_t smem, cudaStream_t st, bool wide) {
  switch (nl) {
    case 1: return wide ? fu_launch<1, C, true>(p, smem, st) : fu_launch<1, C, false>(p, smem, st);
    case 2: return wide ? fu_launch<2, C, true>(p, smem, st) : fu_launch<2, C, false>(p, smem, st);
    default: return wide ? fu_launch<3, C, true>(p, smem, st) : fu_launch<3, C, false>(p, smem, st);
  }
}

// A prepared whole-model launch: everything but the batch (ids / records, dense, row count, out) frozen.
struct FuPlan {
  FuParams p; int nl; size_t smem; bool wide; int cluster; int dense_off; bool packed;
};

struct FuExtra {
  const float* add_vec = nullptr; float out_scale = 1.f, add_scale = 1.f, add_const = 0.f;   // tower-mode post
  float* h_out = nullptr; int64_t h_ld = 0;                                                  // store mode
  const float* cc_w = nullptr; float* side_out = nullptr;                                    // CUDA-core side products (DCN)
  const float* wide_w = nullptr; const int64_t* wide_off = nullptr; const float* wide_dense = nullptr;   // Wide&Deep wide term
  int64_t wide_dense_ld = 0; int wide_D = 0; const float* wide_dw = nullptr;
  const int32_t* pack_desc = nullptr;                                                        // CTRHOT_IDS_PACKED
  const int* wait_flags = nullptr; int* signal_flags = nullptr; int epoch = 0;               // chained launches (FuParams)
  struct FuPlan* capture = nullptr;      // plan creation: keep the prepared launch instead of making it
};

int fu_dispatch(const FuParams& p, int nl, size_t smem, bool wide, int cluster, cudaStream_t st);

// peers[0..G): arena of owner g (this rank's own at its index; the others mapped through ctrhot_peer_open); G = 1 is the
// unsharded case.  row_off is [G, F+1].
int fu_entry(const float* const* peers, int G, int rank, const int64_t* row_off, int row_stride, int F, int k,
             const void* ids, int ids_kind, int64_t ids_ld,
             const float* dense, int64_t dense_ld, int D, int64_t M,
             const ctrhot_dense_desc* h_layers, int n_layers,
             const float* sq_w, int mode, int n_side, const float* h_combine, int n_combine,
             int apply_sigmoid, int cluster, float* out, int* err_flag, void* stream, const FuExtra& ex = FuExtra()) {
  float* const h_out = ex.h_out;
  const int64_t h_ld = ex.h_ld;
  if (!peers || G < 1 || G > FU_MAXG || rank < 0 || rank >= G) return CTRHOT_EARG;
  for (int g = 0; g < G; ++g) if (!peers[g] || !aligned16(peers[g])) return CTRHOT_EARG;
  if ((int64_t)F * G > 4096) return CTRHOT_ESHAPE;
  const float* arena = peers[rank];
  const bool rows_mode = (row_off == nullptr && ids == nullptr);       // ctrhot_tower_rows_tf32: the table is x itself
  if (!arena || (!out && !h_out) || !h_layers || F <= 0 || k <= 0 || M < 0 || D < 0) return CTRHOT_EARG;
  if (!rows_mode && (!row_off || !ids)) return CTRHOT_EARG;
  if (rows_mode && (F != 1 || G != 1 || M > 0x7FFFFFFF || row_stride < k || row_stride % 4 != 0)) return CTRHOT_EARG;
  if (D > 0 && !dense) return CTRHOT_EARG;
  if (ids_kind == CTRHOT_IDS_PACKED) {
    if (!ex.pack_desc || rows_mode || ids_ld <= 0 || ids_ld % 4 != 0 || (reinterpret_cast<uintptr_t>(ids) & 3u)) return CTRHOT_EARG;
  } else if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) {
    return CTRHOT_EARG;
  }
  if (mode < FU_MODE_TOWER || mode > FU_MODE_DCN || n_side < 0 || n_side >= FU_SIDE) return CTRHOT_EARG;
  if (n_combine < 0 || n_combine > FU_NCB || (n_combine > 0 && !h_combine)) return CTRHOT_EARG;
  if (mode == FU_MODE_DEEPFM && n_combine < 1) return CTRHOT_EARG;
  if (mode == FU_MODE_DCN && n_combine < n_side + 3) return CTRHOT_EARG;
  const int overlap = (cluster & CTRHOT_LAUNCH_OVERLAP) ? 1 : 0;
  cluster &= ~CTRHOT_LAUNCH_OVERLAP;
  if (cluster != 0 && cluster != 1 && cluster != 2 && cluster != 4) return CTRHOT_EARG;
  // chained launches read what the previous one wrote: they may overlap only under the per-tile flags
  if (overlap && ((ex.h_out && !ex.signal_flags) || (rows_mode && !ex.wait_flags))) return CTRHOT_EARG;
  if ((ex.wait_flags && !rows_mode) || (ex.signal_flags && !ex.h_out)) return CTRHOT_EARG;
  if (k % 4 != 0 || !aligned16(arena) || F > 63) return CTRHOT_ESHAPE;
  // h_out: ONE hidden layer whose activations go to HBM (no final Dense); otherwise hidden layers + the final Dense(h -> 1)
  if (h_out ? (n_layers != 1 || mode == FU_MODE_DEEPFM) : (n_layers < 2 || n_layers > FU_MAXL + 1)) return CTRHOT_ESHAPE;
  if (ex.cc_w && (mode != FU_MODE_DCN || n_side + 1 > 4)) return CTRHOT_ESHAPE;
  if (h_out && mode == FU_MODE_DCN && !ex.cc_w) return CTRHOT_ESHAPE;
  const ctrhot_dense_desc& o = h_layers[n_layers - 1];
  if (!h_out && (o.out_dim != 1 || o.act != CTRHOT_ACT_NONE || !o.weight)) return CTRHOT_ESHAPE;
  FuParams p{};
  const int nl = h_out ? 1 : n_layers - 1;
  const bool side_on = (mode != FU_MODE_TOWER) && !ex.cc_w;      // side products on the tensor pipe
  const int n = D + F * k;
  int KP[FU_MAXL] = {0, 0, 0};
  int col = 0;
  uint32_t bmax = 0;
  for (int l = 0; l < nl; ++l) {
    const ctrhot_dense_desc& d = h_layers[l];
    if (!d.weight || !d.packed || !aligned16(d.packed) || d.in_dim <= 0 || d.out_dim <= 0) return CTRHOT_EARG;
    if (d.out_dim > (nl == 1 ? FU_MAXN1 : 256) || d.out_dim < 8) return CTRHOT_ESHAPE;
    if (l == 0 ? d.in_dim != n : d.in_dim != h_layers[l - 1].out_dim) return CTRHOT_EARG;
    if (d.act == CTRHOT_ACT_PRELU) {
      if (!d.act_param) return CTRHOT_EARG;
      if (d.alpha_rows > 1) return CTRHOT_ESHAPE;
    } else if (d.act != CTRHOT_ACT_NONE && d.act != CTRHOT_ACT_RELU) {
      return CTRHOT_ESHAPE;
    }
    FuLayer& L = p.L[l];
    L.Wp = static_cast<const float*>(d.packed); L.bias = d.bias; L.alpha = d.act_param; L.act = d.act;
    L.N = d.out_dim; L.KC = (d.in_dim + 31) / 32; L.BN = fu_round(d.out_dim, 16); L.KP = fu_round(d.out_dim, 32);
    L.BNa = L.BN <= 256 ? L.BN : fu_round(L.BN / 2, 16);
    L.col = col;
    col += L.KP;
    if (l == 0) {
      if (side_on) { p.ds_col = col; col += 2 * FU_SIDE; }
      p.a_col = col;
    }
    KP[l] = L.KP;
    const uint32_t b = (uint32_t)(L.BN + (l == 0 && side_on ? 2 * FU_SIDE : 0)) * 128u;
    if (b > bmax) bmax = b;
  }
  if (col > 512) return CTRHOT_ESHAPE;
  p.a_w = side_on ? 64 : 32;
  p.side_rows = side_on ? 2 * FU_SIDE : 0;
  p.n_as = (512 - p.a_col) / p.a_w < FU_NAS ? (512 - p.a_col) / p.a_w : FU_NAS;
  if (p.n_as < 2) return CTRHOT_ESHAPE;
  if (!h_out && o.in_dim != h_layers[nl - 1].out_dim) return CTRHOT_EARG;
  if (h_out && (h_ld < h_layers[0].out_dim || h_ld % 4 != 0 || !aligned16(h_out))) return CTRHOT_EARG;
  if (M == 0) return CTRHOT_OK;
  p.G = G;
  for (int g = 0; g < FU_MAXG; ++g) p.peer[g] = g < G ? peers[g] : nullptr;
  p.row_stride4 = rows_mode ? row_stride / 4 : k / 4;
  p.arena = arena; p.row_off = row_off; p.F = F; p.k4 = k / 4; p.ids = ids; p.ids_kind = ids_kind; p.ids_ld = ids_ld;
  p.dense = dense; p.dense_ld = dense_ld; p.D = D; p.emb = F * k;
  p.M = M; p.n_tiles = (int)((M + FU_BM - 1) / FU_BM);
  p.w_out = h_out ? nullptr : o.weight; p.b_out = h_out ? nullptr : o.bias; p.sq_w = sq_w;
  p.h_out = h_out; p.h_ld = h_ld;
  p.wait_flags = ex.wait_flags; p.signal_flags = ex.signal_flags; p.epoch = ex.epoch;
  p.mode = mode; p.n_side = n_side; p.apply_sigmoid = apply_sigmoid; p.overlap = overlap;
  for (int i = 0; i < n_combine; ++i) p.cb[i] = h_combine[i];
  p.out = out; p.err_flag = err_flag; p.b_stage_bytes = bmax;
  p.add_vec = ex.add_vec; p.out_scale = ex.out_scale; p.add_scale = ex.add_scale; p.add_const = ex.add_const;
  p.cc_w = ex.cc_w; p.cc_n = ex.cc_w ? n_side + 1 : 0; p.side_out = ex.side_out;
  p.pack_desc = (ids_kind == CTRHOT_IDS_PACKED) ? ex.pack_desc : nullptr;
  if (ex.wide_w) {
    // the stage tail holds at most four fields per 32-column chunk and assumes fields start on piece boundaries
    if (mode != FU_MODE_TOWER || G != 1 || rows_mode || !ex.wide_off || ex.wide_D < 0 || ex.wide_D > 16 ||
        (ex.wide_D > 0 && (!ex.wide_dense || !ex.wide_dw)) || (k & (k - 1)) != 0 || k < 8) return CTRHOT_ESHAPE;
    p.wide_w = ex.wide_w; p.wide_off = ex.wide_off; p.wide_dense = ex.wide_dense; p.wide_dense_ld = ex.wide_dense_ld;
    p.wide_D = ex.wide_D;
    for (int j = 0; j < 16; ++j) p.wide_dw[j] = j < ex.wide_D ? ex.wide_dw[j] : 0.f;
  }
  const bool wide = (ex.wide_w != nullptr);
  p.raw_stride = FU_TILE_BYTES + (wide ? 2048u : 0u);
  const FuSmemMap mp = fu_smem_map(F * G, p.L[0].KC, KP, nl, p.cc_n > 0, p.wide_w != nullptr, p.pack_desc != nullptr);
  const size_t n_bars = 2 * FU_NR + 2 * FU_NAS + 2 * FU_NB + 2 * FU_MAXL + 1 + 1 + 8 * FU_MAXL;   // + TMEM slot word, h_chunk
  size_t smem = 0;
  bool fits = false;
  // deepest rings that fit: weight stages first (each is a round trip to L2), then raw-row stages
  for (p.nb = FU_NB; p.nb >= 2 && !fits; --p.nb) {
    for (p.nr = FU_NR; p.nr >= (p.nb >= 3 ? 3 : 2) && !fits; --p.nr) {
      smem = 1024 + (size_t)p.nr * p.raw_stride + (size_t)p.nb * bmax + (size_t)mp.total * 4 + 8 * n_bars + 16;
      fits = smem <= 227 * 1024;
      if (fits) break;
    }
    if (fits) break;
  }
  if (!fits) return CTRHOT_ESHAPE;
  if (ex.capture) {
    ex.capture->p = p; ex.capture->nl = nl; ex.capture->smem = smem; ex.capture->wide = wide; ex.capture->cluster = cluster;
    return CTRHOT_OK;
  }
  return fu_dispatch(p, nl, smem, wide, cluster, static_cast<cudaStream_t>(stream));
}

int fu_dispatch(const FuParams& p, int nl, size_t smem, bool wide, int cluster, cudaStream_t st) {
  int C = cluster;
  // Overlapped launches (CTRHOT_LAUNCH_OVERLAP): the next launch's CTAs take the SMs this one leaves idle at its tail, which
  // only works SM by SM -- a cluster needs 2 or 4 free SMs of one GPC at once, and 4-clusters occupy 132 of the 148 SMs to begin
  // with.  Measured at cfg 2a (profiles/cluster_overlap_sweep.py): 58.4 us per step with single CTAs, 60.9 with pairs, 67.0 with
  // clusters of 4 (without overlap the three are within 1 %: 75-77 us).
  if (C == 0 && p.overlap) C = 1;
  if (C == 0) {
    // larger clusters share each weight chunk among more CTAs (less L2 traffic) but leave SMs idle when a GPC's SM count
    // is not a multiple of the cluster size: take 4 unless 2 saves a whole wave of tiles
    if (p.n_tiles < 2) C = 1;
    else if (p.n_tiles < 4 * 32) C = 2;
    else {
      const int n4 = fu_max_clusters_nl<4>(nl, smem, wide), n2 = fu_max_clusters_nl<2>(nl, smem, wide);
      const int w4 = ((p.n_tiles + 3) / 4 + n4 - 1) / n4, w2 = ((p.n_tiles + 1) / 2 + n2 - 1) / n2;
      C = (w4 <= w2) ? 4 : 2;
    }
  }
  // every multicast piece must stay 16-byte granular (it is: rows * 128 / C)
  if (C == 4) return fu_launch_nl<4>(p, nl, smem, st, wide);
  if (C == 2) return fu_launch_nl<2>(p, nl, smem, st, wide);
  return fu_launch_nl<1>(p, nl, smem, st, wide);
}
}  // namespace

extern "C" int ctrhot_ctr_fused_tf32(const float* arena, const int64_t* row_off, int F, int k,
                                     const void* ids, int ids_kind, int64_t ids_ld,
                                     const float* dense, int64_t dense_ld, int D, int64_t M,
                                     const ctrhot_dense_desc* h_layers, int n_layers,
                                     const float* sq_w, int mode, int n_side, const float* h_combine, int n_combine,
                                     int apply_sigmoid, int cluster, float* out, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_ctr_fused_tf32");
  if (!arena) return CTRHOT_EARG;
  return fu_entry(&arena, 1, 0, row_off, 0, F, k, ids, ids_kind, ids_ld, dense, dense_ld, D, M, h_layers, n_layers, sq_w, mode,
                  n_side, h_combine, n_combine, apply_sigmoid, cluster, out, err_flag, stream);
}

extern "C" int ctrhot_ctr_fused_sharded_tf32(const void* const* arenas, const int64_t* row_off, int G, int rank, int F, int k,
                                             const void* ids, int ids_kind, int64_t ids_ld,
                                             const float* dense, int64_t dense_ld, int D, int64_t M,
                                             const ctrhot_dense_desc* h_layers, int n_layers,
                                             const float* sq_w, int mode, int n_side, const float* h_combine,
                                             int n_combine, int apply_sigmoid, int cluster, float* out, int* err_flag,
                                             void* stream) {
  CTRHOT_RANGE("ctrhot_ctr_fused_sharded_tf32");
  if (!arenas || G < 1 || G > FU_MAXG) return CTRHOT_EARG;
  const float* peers[FU_MAXG];
  for (int g = 0; g < G; ++g) peers[g] = static_cast<const float*>(arenas[g]);
  return fu_entry(peers, G, rank, row_off, 0, F, k, ids, ids_kind, ids_ld, dense, dense_ld, D, M, h_layers, n_layers, sq_w,
                  mode, n_side, h_combine, n_combine, apply_sigmoid, cluster, out, err_flag, stream);
}

// ctrhot_ctr_fused_tf32 fed from compact transport records (csrc/transport.cu): record r = [ ... dense D floats at byte
// dense_off ... | ids at their per-field byte widths ... ], rec_bytes a multiple of 4.  The gather warps decode the ids
// themselves, the dense columns are read in place: one host-to-device copy per batch, no unpack pass, no int32 id matrix.
extern "C" int ctrhot_ctr_fused_packed_tf32(const float* arena, const int64_t* row_off, int F, int k,
                                            const void* records, int64_t rec_bytes, const int32_t* pack_desc,
                                            int dense_off, int D, int64_t M,
                                            const ctrhot_dense_desc* h_layers, int n_layers,
                                            const float* sq_w, int mode, int n_side, const float* h_combine, int n_combine,
                                            int apply_sigmoid, int cluster, float* out, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_ctr_fused_packed_tf32");
  if (!arena || !records || !pack_desc || rec_bytes <= 0 || rec_bytes % 4 != 0 || dense_off < 0 || dense_off % 4 != 0 ||
      (int64_t)dense_off + 4 * (int64_t)D > rec_bytes || !aligned16(records)) return CTRHOT_EARG;
  FuExtra ex;
  ex.pack_desc = pack_desc;
  const float* dense = reinterpret_cast<const float*>(static_cast<const char*>(records) + dense_off);
  return fu_entry(&arena, 1, 0, row_off, 0, F, k, records, CTRHOT_IDS_PACKED, rec_bytes, D > 0 ? dense : nullptr, rec_bytes / 4,
                  D, M, h_layers, n_layers, sq_w, mode, n_side, h_combine, n_combine, apply_sigmoid, cluster, out, err_flag,
                  stream, ex);
}

// ---------------------------------------------------------------- plans: one cheap call per batch
// ctrhot_ctr_fused_tf32 / _packed_tf32 validate their ~25 arguments, size shared memory and pick a cluster shape on every
// call; a serving loop calls them every ~70 us with the same model.  A plan does that once: create freezes everything but the
// batch, run takes the batch.  Host memory only (a few hundred bytes, freed by destroy); the device pointers inside (tables,
// packed weights, sq_w, pack_desc, err_flag) must stay valid and unchanged for the plan's lifetime -- re-create it when the
// weights are re-packed.  ids_kind I32 / F32: `batch` = ids [M, ids_ld], `dense` [M, dense_ld]; PACKED: `batch` = records
// (ids_ld = record bytes), dense is taken from the records at dense_off and the `dense` argument of run is ignored.
extern "C" int ctrhot_fused_plan_create(const float* arena, const int64_t* row_off, int F, int k,
                                        int ids_kind, int64_t ids_ld, const int32_t* pack_desc, int dense_off,
                                        int64_t dense_ld, int D,
                                        const ctrhot_dense_desc* h_layers, int n_layers,
                                        const float* sq_w, int mode, int n_side, const float* h_combine, int n_combine,
                                        int apply_sigmoid, int cluster, int* err_flag, void** plan) {
  CTRHOT_RANGE("ctrhot_fused_plan_create");
  if (!arena || !plan) return CTRHOT_EARG;
  FuPlan* pl = new (std::nothrow) FuPlan();
  if (!pl) return CTRHOT_EARG;
  FuExtra ex;
  ex.capture = pl;
  pl->packed = (ids_kind == CTRHOT_IDS_PACKED);
  pl->dense_off = dense_off;
  // representative batch pointers for the argument checks: any aligned non-null address (nothing is launched)
  const void* fake = arena;
  float* fake_out = const_cast<float*>(arena);
  int rc;
  if (pl->packed) {
    if (!pack_desc || ids_ld <= 0 || ids_ld % 4 != 0 || dense_off < 0 || dense_off % 4 != 0 ||
        (int64_t)dense_off + 4 * (int64_t)D > ids_ld) { delete pl; return CTRHOT_EARG; }
    ex.pack_desc = pack_desc;
    rc = fu_entry(&arena, 1, 0, row_off, 0, F, k, fake, CTRHOT_IDS_PACKED, ids_ld, D > 0 ? arena : nullptr, ids_ld / 4, D, 128,
                  h_layers, n_layers, sq_w, mode, n_side, h_combine, n_combine, apply_sigmoid, cluster, fake_out, err_flag,
                  nullptr, ex);
  } else {
    rc = fu_entry(&arena, 1, 0, row_off, 0, F, k, fake, ids_kind, ids_ld, D > 0 ? arena : nullptr, dense_ld, D, 128, h_layers,
                  n_layers, sq_w, mode, n_side, h_combine, n_combine, apply_sigmoid, cluster, fake_out, err_flag, nullptr, ex);
  }
  if (rc != CTRHOT_OK) { delete pl; return rc; }
  *plan = pl;
  return CTRHOT_OK;
}

extern "C" int ctrhot_fused_plan_run(const void* plan, const void* batch, const float* dense, int64_t M, float* out,
                                     void* stream) {
  CTRHOT_RANGE("ctrhot_fused_plan_run");
  const FuPlan* pl = static_cast<const FuPlan*>(plan);
  if (!pl || !batch || !out || M < 0) return CTRHOT_EARG;
  if (M == 0) return CTRHOT_OK;
  FuParams p = pl->p;
  if (pl->packed) {
    if (!aligned16(batch)) return CTRHOT_EARG;
    p.ids = batch;
    p.dense = p.D > 0 ? reinterpret_cast<const float*>(static_cast<const char*>(batch) + pl->dense_off) : nullptr;
  } else {
    if (p.D > 0 && !dense) return CTRHOT_EARG;
    p.ids = batch;
    p.dense = dense;
  }
  p.M = M;
  p.n_tiles = (int)((M + FU_BM - 1) / FU_BM);
  p.out = out;
  return fu_dispatch(p, pl->nl, pl->smem, pl->wide, pl->cluster, static_cast<cudaStream_t>(stream));
}

extern "C" int ctrhot_fused_plan_destroy(void* plan) {
  CTRHOT_RANGE("ctrhot_fused_plan_destroy");
  delete static_cast<FuPlan*>(plan);
  return CTRHOT_OK;
}

// DNNLayer.call (DL/layer/interaction.py:40-46) on an activation matrix that already lives in HBM: the same persistent
// kernel with the matrix as its one 'table' (row m of a tile = row m of x, no ids, no offsets).  The first K - K % 4
// columns are fetched as 16-byte pieces, the K % 4 tail through the dense path (4-byte copies, zero-filled), so neither
// the padding of a row nor whatever lies behind column K is ever multiplied.
extern "C" int ctrhot_tower_rows_tf32(const float* x, int64_t x_ld, int K, int64_t M,
                                      const ctrhot_dense_desc* h_layers, int n_layers, int apply_sigmoid, int cluster,
                                      float* out, void* stream) {
  CTRHOT_RANGE("ctrhot_tower_rows_tf32");
  if (!x || K < 4 || x_ld < K || x_ld % 4 != 0 || x_ld > 0x7FFFFFFF || !aligned16(x)) return CTRHOT_EARG;
  const int r = K % 4, k = K - r;
  return fu_entry(&x, 1, 0, nullptr, (int)x_ld, 1, k, nullptr, CTRHOT_IDS_I32, 0, r ? x + k : nullptr, x_ld, r, M, h_layers,
                  n_layers, nullptr, FU_MODE_TOWER, 0, nullptr, 0, apply_sigmoid, cluster, out, nullptr, stream);
}

// One Dense layer of up to 416 units over a matrix in HBM, activations written back to HBM: the building block of towers
// too wide to be chained through TMEM (DCN cfg 3: 429 -> 400 -> 400 -> 400 -> 1 runs as store, store, then
// ctrhot_tower_rows_tf32 with the last hidden layer + the final Dense).  h[m, 0:N] = tf32(act(x_m W + b)), h[m, N:h_ld) = 0.
extern "C" int ctrhot_dense_rows_tf32(const float* x, int64_t x_ld, int K, int64_t M, const ctrhot_dense_desc* layer,
                                      float* h, int64_t h_ld, int cluster, const int32_t* wait_flags, int32_t* signal_flags,
                                      int epoch, void* stream) {
  CTRHOT_RANGE("ctrhot_dense_rows_tf32");
  if (!x || !layer || !h || K < 4 || x_ld < K || x_ld % 4 != 0 || x_ld > 0x7FFFFFFF || !aligned16(x)) return CTRHOT_EARG;
  const int r = K % 4, k = K - r;
  FuExtra ex;
  ex.h_out = h; ex.h_ld = h_ld;
  ex.wait_flags = wait_flags; ex.signal_flags = signal_flags; ex.epoch = epoch;
  return fu_entry(&x, 1, 0, nullptr, (int)x_ld, 1, k, nullptr, CTRHOT_IDS_I32, 0, r ? x + k : nullptr, x_ld, r, M, layer, 1,
                  nullptr, FU_MODE_TOWER, 0, nullptr, 0, 0, cluster, nullptr, nullptr, stream, ex);
}

// WideDeep.call (DL/model/wideDeep.py:24-34): out = sigmoid(scale * (add_vec[m] + DNN(Embed(ids))[m])) with the deep tower
// fed straight from the embedding arena by the persistent kernel (no side products: main MMAs only).
extern "C" int ctrhot_ctr_tower_tf32(const float* arena, const int64_t* row_off, int F, int k,
                                     const void* ids, int ids_kind, int64_t ids_ld,
                                     const float* dense, int64_t dense_ld, int D, int64_t M,
                                     const ctrhot_dense_desc* h_layers, int n_layers,
                                     const float* add_vec, float scale, int apply_sigmoid, int cluster,
                                     float* out, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_ctr_tower_tf32");
  if (!arena) return CTRHOT_EARG;
  FuExtra ex;
  ex.add_vec = add_vec; ex.out_scale = scale; ex.add_scale = scale;
  return fu_entry(&arena, 1, 0, row_off, 0, F, k, ids, ids_kind, ids_ld, dense, dense_ld, D, M, h_layers, n_layers, nullptr,
                  FU_MODE_TOWER, 0, nullptr, 0, apply_sigmoid, cluster, out, err_flag, stream, ex);
}

// WideDeep.call (DL/model/wideDeep.py:22-34) in ONE launch: the deep tower fed from the embedding arena as in
// ctrhot_ctr_tower_tf32, and the wide logit formed inside the same kernel -- the gather warps fetch w[wide_off[f] + id_f] next to
// the embedding rows (the ids are in their registers anyway), the stager threads add them and the dense term per row:
//   out[m] = (sigmoid of) scale * (tower(m) + w0 + sum_j dense[m, j] dw[j] + sum_f w[wide_off[f] + ids[m, f]])
// wide_off: DEVICE int64 [F + 1] (field f's one-hot block is [wide_off[f], wide_off[f+1]); an id at or beyond its width adds 0,
// like a get_dummies column that does not exist); dw: HOST float [wide_D <= 16] (the reference's wide input carries the dense block twice: pass the sum
// of the two weight blocks).  k must be a power of two >= 8.
extern "C" int ctrhot_widedeep_fused_tf32(const float* arena, const int64_t* row_off, int F, int k,
                                          const void* ids, int ids_kind, int64_t ids_ld, int64_t M,
                                          const ctrhot_dense_desc* h_layers, int n_layers,
                                          const float* wide_w, const int64_t* wide_off, const float* wide_dense,
                                          int64_t wide_dense_ld, int wide_D, const float* h_dw, float w0, float scale,
                                          int apply_sigmoid, int cluster, float* out, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_widedeep_fused_tf32");
  if (!arena || !wide_w || !wide_off) return CTRHOT_EARG;
  FuExtra ex;
  ex.out_scale = scale; ex.add_scale = scale; ex.add_const = scale * w0;
  ex.wide_w = wide_w; ex.wide_off = wide_off; ex.wide_dense = wide_dense; ex.wide_dense_ld = wide_dense_ld;
  ex.wide_D = wide_D; ex.wide_dw = h_dw;
  return fu_entry(&arena, 1, 0, row_off, 0, F, k, ids, ids_kind, ids_ld, nullptr, 0, 0, M, h_layers, n_layers, nullptr,
                  FU_MODE_TOWER, 0, nullptr, 0, apply_sigmoid, cluster, out, err_flag, stream, ex);
}

// DCN whose tower is too wide for TMEM chaining (cfg 3: 400-400-400), first launch: gather -> { L cross layers as scalars on
// the CUDA cores (exact fp32, in the stager threads) | first Dense layer on the tensor pipe }; h1 and the per-row cross term
// side[m] = a_L (x . final_w) + c_L . final_w go to HBM, x = [dense | embed] does not.  n_side = L <= 3.
extern "C" int ctrhot_dcn_first_layer_tf32(const float* arena, const int64_t* row_off, int F, int k,
                                           const void* ids, int ids_kind, int64_t ids_ld,
                                           const float* dense, int64_t dense_ld, int D, int64_t M,
                                           const ctrhot_dense_desc* layer, const float* cc_w, int n_side,
                                           const float* h_combine, int n_combine, float* h, int64_t h_ld, float* side_out,
                                           int cluster, int* err_flag, int32_t* signal_flags, int epoch, void* stream) {
  CTRHOT_RANGE("ctrhot_dcn_first_layer_tf32");
  if (!arena || !cc_w || !side_out || !h) return CTRHOT_EARG;
  FuExtra ex;
  ex.h_out = h; ex.h_ld = h_ld; ex.cc_w = cc_w; ex.side_out = side_out;
  ex.signal_flags = signal_flags; ex.epoch = epoch;
  return fu_entry(&arena, 1, 0, row_off, 0, F, k, ids, ids_kind, ids_ld, dense, dense_ld, D, M, layer, 1, nullptr,
                  FU_MODE_DCN, n_side, h_combine, n_combine, 0, cluster, nullptr, err_flag, stream, ex);
}

// ctrhot_tower_rows_tf32 with an affine post: out[m] = (sigmoid of) y_scale * tower(x_m) + add_scale * add_vec[m] + add_const
extern "C" int ctrhot_tower_rows_post_tf32(const float* x, int64_t x_ld, int K, int64_t M,
                                           const ctrhot_dense_desc* h_layers, int n_layers,
                                           const float* add_vec, float y_scale, float add_scale, float add_const,
                                           int apply_sigmoid, int cluster, float* out, const int32_t* wait_flags, int epoch,
                                           void* stream) {
  CTRHOT_RANGE("ctrhot_tower_rows_post_tf32");
  if (!x || K < 4 || x_ld < K || x_ld % 4 != 0 || x_ld > 0x7FFFFFFF || !aligned16(x)) return CTRHOT_EARG;
  const int r = K % 4, k = K - r;
  FuExtra ex;
  ex.add_vec = add_vec; ex.out_scale = y_scale; ex.add_scale = add_scale; ex.add_const = add_const;
  ex.wait_flags = wait_flags; ex.epoch = epoch;
  return fu_entry(&x, 1, 0, nullptr, (int)x_ld, 1, k, nullptr, CTRHOT_IDS_I32, 0, r ? x + k : nullptr, x_ld, r, M, h_layers,
                  n_layers, nullptr, FU_MODE_TOWER, 0, nullptr, 0, apply_sigmoid, cluster, out, nullptr, stream, ex);
}
