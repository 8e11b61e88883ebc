// The input step on the left of the hot path: Criteo TSV text -> MinMax-scaled dense features and
// label-encoded ids, on the device.  Reference: create_criteo_dataset, DL/utils/dataset.py:36-66 -- pandas read_csv (:38),
// fillna 0 / '-1' (:42-43), sklearn MinMaxScaler on the 13 counts (:46), sklearn LabelEncoder per categorical column (:51-52).
//
//   parse : one thread per text line: label, 13 integer counts (empty -> missing), 26 tokens of 8 lowercase hex digits (empty ->
//           missing).  A token is kept as the 32-bit value of its digits: for fixed-width lowercase hex the string order
//           LabelEncoder sorts by IS the numeric order, and '-1' (the fill value) sorts before every digit string.
//   encode: dense = fp32( x * scale + min_ ) evaluated in float64 exactly as MinMaxScaler.transform does (scale = 1 / (max - min),
//           1 where the column is constant; min_ = -min * scale); id = rank of the token in the column's sorted distinct values
//           (missing = the fill value '-1' = rank 0 when the column has any), by binary search.  A token that is not in the
//           vocabulary sets *err_flag and gets id -1 (LabelEncoder.transform raises on unseen labels).
// The vocabularies and min / max are FIT once per data set (host side of this file's callers: utils/dataset.py); encode is
// what runs per batch.
#include "common.cuh"

namespace ctrhot {

constexpr int ENC_D = 13, ENC_F = 26;

__global__ void criteo_parse_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ line_start, int64_t N,
                                    int has_label, int8_t* __restrict__ label, double* __restrict__ dense_raw,
                                    int64_t* __restrict__ tokens, int* __restrict__ err_flag) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < N; r += (int64_t)gridDim.x * blockDim.x) {
    const uint8_t* p = text + line_start[r];
    const uint8_t* end = text + line_start[r + 1];             // one past the line's '\n' (or the end of the buffer)
    while (end > p && (end[-1] == '\n' || end[-1] == '\r')) --end;
    bool bad = false;
    const int n_fields = (has_label ? 1 : 0) + ENC_D + ENC_F;
    for (int c = 0; c < n_fields; ++c) {
      const uint8_t* q = p;
      while (q < end && *q != '\t') ++q;                        // field = [p, q)
      const int col = c - (has_label ? 1 : 0);
      if (col < 0) {
        label[r] = (q > p && *p == '1') ? 1 : 0;
      } else if (col < ENC_D) {
        if (q == p) {
          dense_raw[r * ENC_D + col] = __longlong_as_double(0x7FF8000000000000LL);     // NaN = missing (pandas' NaN)
        } else {
          const uint8_t* s = p;
          const bool neg = (*s == '-');
          if (neg) ++s;
          long long v = 0;
          if (s == q) bad = true;
          for (; s < q; ++s) {
            if (*s < '0' || *s > '9') { bad = true; break; }
            v = v * 10 + (*s - '0');
          }
          dense_raw[r * ENC_D + col] = (double)(neg ? -v : v);
        }
      } else {
        const int f = col - ENC_D;
        if (q == p) {
          tokens[r * ENC_F + f] = -1;                           // missing -> the fill value '-1', which sorts first
        } else {
          long long v = 0;
          if (q - p != 8) bad = true;
          for (const uint8_t* s = p; s < q; ++s) {
            const int ch = *s;
            const int d = (ch >= '0' && ch <= '9') ? ch - '0' : (ch >= 'a' && ch <= 'f') ? ch - 'a' + 10 : -1;
            if (d < 0) { bad = true; break; }
            v = (v << 4) | d;
          }
          tokens[r * ENC_F + f] = v;
        }
      }
      p = (q < end) ? q + 1 : end;
      if (q >= end && c + 1 < n_fields) {
        // a short line: the remaining fields are missing (pandas fills trailing columns with NaN)
        for (int c2 = c + 1; c2 < n_fields; ++c2) {
          const int col2 = c2 - (has_label ? 1 : 0);
          if (col2 < ENC_D) dense_raw[r * ENC_D + col2] = __longlong_as_double(0x7FF8000000000000LL);
          else tokens[r * ENC_F + (col2 - ENC_D)] = -1;
        }
        break;
      }
    }
    if (bad && err_flag) *err_flag = 1;
  }
}

__global__ void criteo_encode_kernel(const double* __restrict__ dense_raw, const int64_t* __restrict__ tokens, int64_t N,
                                     const double* __restrict__ d_scale, const double* __restrict__ d_min,
                                     const int64_t* __restrict__ vocab, const int64_t* __restrict__ vocab_off,
                                     float* __restrict__ dense, int64_t dense_ld, int32_t* __restrict__ ids, int64_t ids_ld,
                                     int* __restrict__ err_flag) {
  const int C = ENC_D + ENC_F;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N * C; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / C;
    const int c = (int)(i - r * C);
    if (c < ENC_D) {
      double x = dense_raw[r * ENC_D + c];
      if (x != x) x = 0.0;                                                 // fillna(0), dataset.py:42
      // MinMaxScaler.transform: X *= scale_; X += min_  -- two rounded float64 operations, not an fma
      const double y = __dadd_rn(__dmul_rn(x, d_scale[c]), d_min[c]);
      dense[r * dense_ld + c] = (float)y;
    } else {
      const int f = c - ENC_D;
      const int64_t t = tokens[r * ENC_F + f];
      const int64_t* v = vocab + vocab_off[f];
      int64_t lo = 0, hi = vocab_off[f + 1] - vocab_off[f];                // first index with v[idx] >= t
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(v + mid) < t) lo = mid + 1; else hi = mid;
      }
      const bool found = lo < vocab_off[f + 1] - vocab_off[f] && __ldg(v + lo) == t;
      if (!found && err_flag) *err_flag = 1;
      ids[r * ids_ld + f] = found ? (int32_t)lo : -1;
    }
  }
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int ctrhot_criteo_parse_tsv(const void* text, const int64_t* line_start, int64_t N, int has_label, int8_t* label,
                                       double* dense_raw, int64_t* tokens, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_criteo_parse_tsv");
  if (!text || !line_start || N < 0 || !dense_raw || !tokens || (has_label && !label)) return CTRHOT_EARG;
  if (N == 0) return CTRHOT_OK;
  int64_t grid = (N + 127) / 128;
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  criteo_parse_kernel<<<(int)grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(text), line_start, N, has_label, label, dense_raw, tokens, err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_criteo_parse_tsv");
  return CTRHOT_OK;
}

extern "C" int ctrhot_criteo_encode(const double* dense_raw, const int64_t* tokens, int64_t N, const double* d_scale,
                                    const double* d_min, const int64_t* vocab, const int64_t* vocab_off, float* dense,
                                    int64_t dense_ld, int32_t* ids, int64_t ids_ld, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_criteo_encode");
  if (!dense_raw || !tokens || N < 0 || !d_scale || !d_min || !vocab || !vocab_off || !dense || !ids || dense_ld < ENC_D ||
      ids_ld < ENC_F) return CTRHOT_EARG;
  if (N == 0) return CTRHOT_OK;
  const int64_t n = N * (ENC_D + ENC_F);
  int64_t grid = (n + 255) / 256;
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  criteo_encode_kernel<<<(int)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(dense_raw, tokens, N, d_scale, d_min, vocab,
                                                                                 vocab_off, dense, dense_ld, ids, ids_ld, err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_criteo_encode");
  return CTRHOT_OK;
}
