// Compact transport records for Criteo-shaped batches: the step on the LEFT of the hot path, reduced to
// what moves over PCIe.  The reference hands TensorFlow one float64 [B, 39] matrix (312 B per sample, DL/utils/dataset.py:61,
// DL/model/deepFM.py:24); the int32 fast path of this library takes 156 B.  A record keeps the 13 dense floats as they are
// (bit-exact) and stores every id at the byte width its vocabulary needs:
//
//     record = [ dense: D x float32 | id_0 .. id_{F-1}, field f in width[f] little-endian bytes | zero pad to 4 bytes ]
//     width[f] = 1 / 2 / 3 / 4 bytes for feat_onehot_dim <= 2^8 / 2^16 / 2^24 / 2^31
//
// 104 B per sample on the Criteo-Kaggle cardinalities; with ctrhot_rows_layout_bits every id takes ceil(log2(vocab)) BITS instead
// (96 B; fields then share bytes).  ids beyond a field's width cannot be represented: the packer reports them
// (CTRHOT_EID) instead of truncating.  The whole-model kernel decodes records in its gather warps (CTRHOT_IDS_PACKED,
// csrc/fused.cu); ctrhot_rows_unpack turns them back into (dense, ids) on the device for every other entry point.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <thread>
#include <vector>
#include "common.cuh"

namespace ctrhot {

__global__ void rows_unpack_kernel(const uint8_t* __restrict__ rec, int64_t rec_bytes, const int32_t* __restrict__ desc, int F,
                                   int dense_off, int D, int64_t M, float* __restrict__ dense, int64_t dense_ld,
                                   int32_t* __restrict__ ids, int64_t ids_ld) {
  // one thread per (sample, column): columns [0, D) are the dense floats, [D, D + F) the ids
  const int C = D + F;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M * C; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = i / C;
    const int c = (int)(i - m * C);
    const uint8_t* r = rec + m * rec_bytes;
    if (c < D) {
      dense[m * dense_ld + c] = *reinterpret_cast<const float*>(r + dense_off + 4 * c);
    } else {
      uint32_t bit_off, bits;
      rows_field((uint32_t)__ldg(desc + (c - D)), bit_off, bits);
      const uint8_t* a = r + (bit_off >> 3);
      const uint32_t sh = bit_off & 7u;
      const int nbytes = (int)((sh + bits + 7u) >> 3);                 // <= 5
      uint64_t v = 0;
      for (int b = 0; b < nbytes; ++b) v |= (uint64_t)a[b] << (8 * b);
      const uint64_t mask = bits >= 32u ? 0xFFFFFFFFull : ((1ull << bits) - 1ull);
      ids[m * ids_ld + (c - D)] = (int32_t)((v >> sh) & mask);
    }
  }
}

}  // namespace ctrhot

using namespace ctrhot;

// h_vocab [F] -> h_desc [F] = (byte offset of field f in the record) | (width in bytes) << 16; returns the record size in bytes
// (a multiple of 4), or a negative status.  The dense block sits at byte 0.
extern "C" int ctrhot_rows_layout(const int64_t* h_vocab, int F, int D, int32_t* h_desc) {
  CTRHOT_RANGE("ctrhot_rows_layout");
  if (!h_vocab || !h_desc || F <= 0 || D < 0) return CTRHOT_EARG;
  int64_t off = 4 * (int64_t)D;
  for (int f = 0; f < F; ++f) {
    const int64_t v = h_vocab[f];
    if (v <= 0 || v > 0x7FFFFFFFLL) return CTRHOT_ESHAPE;
    const int w = v <= 0x100 ? 1 : v <= 0x10000 ? 2 : v <= 0x1000000 ? 3 : 4;
    if (off > 0xFFFF) return CTRHOT_ESHAPE;
    h_desc[f] = (int32_t)((uint32_t)off | ((uint32_t)w << 16));
    off += w;
  }
  return (int)((off + 3) / 4 * 4);
}

// The same with every id at the BIT width its vocabulary needs (ceil(log2(vocab)), at least 1): h_desc[f] = 0x80000000 |
// (bits << 24) | (bit offset in the record).  96 B per sample on the Criteo-Kaggle cardinalities (52 of dense floats + 321 bits).
extern "C" int ctrhot_rows_layout_bits(const int64_t* h_vocab, int F, int D, int32_t* h_desc) {
  CTRHOT_RANGE("ctrhot_rows_layout_bits");
  if (!h_vocab || !h_desc || F <= 0 || D < 0) return CTRHOT_EARG;
  int64_t off = 32 * (int64_t)D;
  for (int f = 0; f < F; ++f) {
    const int64_t v = h_vocab[f];
    if (v <= 0 || v > 0x7FFFFFFFLL) return CTRHOT_ESHAPE;
    int bits = 1;
    while (bits < 31 && ((int64_t)1 << bits) < v) ++bits;
    if (off > 0xFFFFFF) return CTRHOT_ESHAPE;
    h_desc[f] = (int32_t)(0x80000000u | ((uint32_t)bits << 24) | (uint32_t)off);
    off += bits;
  }
  return (int)((off + 31) / 32 * 4);
}

// HOST function (no GPU work): (dense float32 [B, D], ids int32 [B, F]) -> records, `n_threads` host threads (<= 0: all).
// Returns CTRHOT_EID if some id does not fit its field's width (nothing is truncated silently: the record then holds 0).
extern "C" int ctrhot_rows_pack_host(const float* h_dense, int64_t dense_ld, const int32_t* h_ids, int64_t ids_ld, int64_t B,
                                     const int32_t* h_desc, int F, int D, int64_t rec_bytes, void* h_records, int n_threads) {
  CTRHOT_RANGE("ctrhot_rows_pack_host");
  if (!h_ids || !h_desc || !h_records || B < 0 || F <= 0 || D < 0 || (D > 0 && !h_dense) || rec_bytes <= 0) return CTRHOT_EARG;
  if (ids_ld < F || (D > 0 && dense_ld < D)) return CTRHOT_EARG;
  int64_t end = 4 * (int64_t)D;
  bool any_bits = false;
  for (int f = 0; f < F; ++f) {
    uint32_t bit_off, bits;
    rows_field((uint32_t)h_desc[f], bit_off, bits);
    any_bits = any_bits || (((uint32_t)h_desc[f]) >> 31);
    if (bits < 1 || bits > 32 || (int64_t)bit_off < 32 * (int64_t)D) return CTRHOT_EARG;
    end = std::max<int64_t>(end, ((int64_t)bit_off + bits + 7) / 8);
  }
  if (end > rec_bytes) return CTRHOT_EARG;
  int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
  nt = (int)std::max<int64_t>(1, std::min<int64_t>(nt, (B + 4095) / 4096));
  std::atomic<int> bad(0);
  auto work = [&](int64_t b0, int64_t b1) {
    uint8_t* out = static_cast<uint8_t*>(h_records) + b0 * rec_bytes;
    int local_bad = 0;
    for (int64_t b = b0; b < b1; ++b, out += rec_bytes) {
      if (D > 0) std::memcpy(out, h_dense + b * dense_ld, 4 * (size_t)D);
      const int32_t* id = h_ids + b * ids_ld;
      if (any_bits) std::memset(out + 4 * (size_t)D, 0, (size_t)(rec_bytes - 4 * (int64_t)D));   // fields share bytes: OR into zeros
      for (int f = 0; f < F; ++f) {
        uint32_t bit_off, bits;
        rows_field((uint32_t)h_desc[f], bit_off, bits);
        uint32_t v = (uint32_t)id[f];
        if (bits < 32 && (v >> bits) != 0) { local_bad = 1; v = 0; }
        uint8_t* a = out + (bit_off >> 3);
        const uint32_t sh = bit_off & 7u;
        const uint64_t val = (uint64_t)v << sh;
        const int nbytes = (int)((sh + bits + 7u) >> 3);
        if (any_bits) { for (int k = 0; k < nbytes; ++k) a[k] |= (uint8_t)(val >> (8 * k)); }
        else { for (int k = 0; k < nbytes; ++k) a[k] = (uint8_t)(val >> (8 * k)); }
      }
      if (!any_bits) for (int64_t z = end; z < rec_bytes; ++z) out[z] = 0;
    }
    if (local_bad) bad.store(1);
  };
  if (nt == 1) {
    work(0, B);
  } else {
    std::vector<std::thread> th;
    const int64_t per = (B + nt - 1) / nt;
    for (int t = 0; t < nt; ++t) {
      const int64_t b0 = t * per, b1 = std::min<int64_t>(B, b0 + per);
      if (b0 < b1) th.emplace_back(work, b0, b1);
    }
    for (auto& t : th) t.join();
  }
  return bad.load() ? CTRHOT_EID : CTRHOT_OK;
}

// records (DEVICE) -> dense float32 [M, dense_ld] and ids int32 [M, ids_ld] (DEVICE): the generic way into every entry point
// that takes (dense, ids).  desc: DEVICE int32 [F].
extern "C" int ctrhot_rows_unpack(const void* records, int64_t rec_bytes, const int32_t* desc, int F, int dense_off, int D,
                                  int64_t M, float* dense, int64_t dense_ld, int32_t* ids, int64_t ids_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_rows_unpack");
  if (!records || !desc || !ids || F <= 0 || D < 0 || M < 0 || rec_bytes <= 0 || rec_bytes % 4 != 0 || dense_off < 0 ||
      dense_off % 4 != 0 || (D > 0 && !dense) || ids_ld < F || (D > 0 && dense_ld < D)) return CTRHOT_EARG;
  if ((reinterpret_cast<uintptr_t>(records) & 3u) != 0) return CTRHOT_EARG;
  if (M == 0) return CTRHOT_OK;
  const int64_t n = M * (D + F);
  int64_t grid = (n + 255) / 256;
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  rows_unpack_kernel<<<(int)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(records), rec_bytes, desc, F, dense_off, D, M, dense, dense_ld, ids, ids_ld);
  CTRHOT_LAUNCH_OK("ctrhot_rows_unpack");
  return CTRHOT_OK;
}
