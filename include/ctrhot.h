/* libctrhot -- C-ABI of the B200-native CTR forward hot path.
 *
 * The reference (Hcyand/recommender_system) is pure Python on TensorFlow and has
 * no FFI of its own; its only seam is the Keras object protocol
 * (Layer.build/call, Model.call).  Each entry point below replaces the
 * arithmetic of one such call() body; the citation names it.  Citations are
 * relative to /root/reference/algorithm/deep_learning/ ("DL/").
 *
 * Conventions
 *  - extern "C", plain pointers and sizes; every function returns an int:
 *    0 ok, <0 a ctrhot_status.  Nothing throws, nothing allocates
 *    caller-visible memory: the caller owns every input/output/workspace buffer.
 *  - All data pointers are DEVICE pointers unless the name starts with h_
 *    (host).  Matrices are row-major with an explicit leading dimension (ld, in
 *    elements).  Vectorised paths need 16-byte aligned bases and ld % 4 == 0;
 *    otherwise a scalar path is taken (never an error).
 *  - Launches are asynchronous on `stream` (a cudaStream_t passed as void*);
 *    entry points are re-entrant and keep no mutable global state.
 *  - Embedding tables of one layer live in one arena [sum(V_f), k] fp32;
 *    row_off[f] is the first arena row of field f, row_off[F] the total
 *    (int64, device).  Out-of-range ids write zeros and set *err_flag (device
 *    int, may be NULL); ctrhot_check_flag() turns it into CTRHOT_EID.
 *  - ids come either as int32 [B, ids_ld] (CTRHOT_IDS_I32) or as the
 *    reference's float columns (CTRHOT_IDS_F32, truncated like keras Embedding's
 *    int32 cast).
 */
#ifndef CTRHOT_H_
#define CTRHOT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CTRHOT_VERSION 100

typedef enum {
  CTRHOT_OK = 0,
  CTRHOT_EARG = -1,        /* bad argument (null pointer, negative size, ...) */
  CTRHOT_ESHAPE = -2,      /* shape not supported by any kernel variant */
  CTRHOT_ECUDA = -3,       /* CUDA error; text via ctrhot_last_cuda_error() */
  CTRHOT_EID = -4          /* an embedding id was out of range */
} ctrhot_status;

enum { CTRHOT_LAUNCH_OVERLAP = 0x100 };   /* OR into `cluster` of the whole-model entry points, see ctrhot_ctr_fused_tf32 */
enum { CTRHOT_IDS_I32 = 0, CTRHOT_IDS_F32 = 1, CTRHOT_IDS_PACKED = 2 /* transport records, see ctrhot_rows_* */ };
enum { CTRHOT_ACT_NONE = 0, CTRHOT_ACT_RELU = 1, CTRHOT_ACT_PRELU = 2, CTRHOT_ACT_DICE = 3,
       CTRHOT_ACT_SIGMOID = 4, CTRHOT_ACT_TANH = 5 };

int ctrhot_version(void);
const char* ctrhot_status_str(int status);
/* thread-local text of the last CUDA error seen by this host thread */
const char* ctrhot_last_cuda_error(void);
/* synchronises `stream`, reads *err_flag, clears it; CTRHOT_EID if it was set */
int ctrhot_check_flag(int* err_flag, void* stream);

/* ---- R1  EmbedLayer.call, DL/layer/core.py:273-280 (+ keras Embedding) --------
 * out[b, D + f*k + c] = arena[row_off[f] + ids[b,f], c]; out[b, 0:D] = dense[b,:]
 * (D = 0 / dense = NULL gives the plain EmbedLayer output [B, F*k]; D = 13
 * gives the x = concat([dense, embed]) of DL/model/deepFM.py:26, dcn.py:27). */
int ctrhot_gather_f32(const float* arena, const int64_t* row_off, int F, int k,
                      const void* ids, int ids_kind, int64_t ids_ld, int64_t B,
                      const float* dense, int64_t dense_ld, int D,
                      float* out, int64_t out_ld, int* err_flag, void* stream);

/* ---- 8(f) next: NFM bi-interaction, NFM.call DL/model/nfm.py:24-30 (with the [B, F, k] view of the embeddings the model
 * means: SURVEY N2) ------------------------------------------------------------------------------------------------------
 * out[b, 0:D] = dense[b, :],  out[b, D:D+k] = 0.5 * ((sum_f e_f)^2 - sum_f e_f^2),  out[b, D+k:out_ld) = 0  -- gather and
 * pooling in registers, the F rows never reach HBM.  k a multiple of 4 with k/4 a power of two, k <= 128. */
int ctrhot_nfm_bi_f32(const float* arena, const int64_t* row_off, int F, int k,
                      const void* ids, int ids_kind, int64_t ids_ld, int64_t B,
                      const float* dense, int64_t dense_ld, int D,
                      float* out, int64_t out_ld, int* err_flag, void* stream);
/* AFMLayer.call, DL/layer/interaction.py:322-351 (+ InteractionLayer :281-297, AttentionLayer :300-319), and AFM.call,
 * DL/model/afm.py: out[b] = sigmoid^n_sigmoid(pool_p(e_i * e_j) . out_w + out_b) without building the [B, P, k] pair tensor.
 * mode 0 'avg' (mean over the P pairs), 1 'att' (the reference's softmax runs over a size-1 axis: every weight is 1, the
 * pooling is the SUM over pairs), 2 'max'.  out_w: device [k] (16-byte aligned), out_b: device [1]; n_sigmoid 1 for the layer,
 * 2 for the model (it applies sigmoid again).  k as for ctrhot_nfm_bi_f32. */
int ctrhot_afm_pool_f32(const float* arena, const int64_t* row_off, int F, int k,
                        const void* ids, int ids_kind, int64_t ids_ld, int64_t B, int mode,
                        const float* out_w, const float* out_b, int n_sigmoid, float* out, int* err_flag,
                        void* stream);

/* ---- R2  WideLayer.call, DL/layer/interaction.py:24-26 -----------------------
 * one-hot form: out[b] = w0 + sum_j dense[b,j]*w[dense_col[j]] + sum_f w[col_off[f]+ids[b,f]]
 * (dense_col int32[D], col_off int64[F+1]; the reference's wide input repeats
 * the dense columns, DL/model/wideDeep.py:23-27 -> pass D=26 dense entries). */
int ctrhot_wide_onehot_f32(const float* dense, int64_t dense_ld, int D, const int32_t* dense_col,
                           const void* ids, int ids_kind, int64_t ids_ld, int F,
                           const int64_t* col_off, int64_t B,
                           const float* w0, const float* w, float* out, void* stream);
/* dense form: out[b] = w0 + x[b,:] @ w */
int ctrhot_wide_dense_f32(const float* x, int64_t x_ld, int64_t n, int64_t B,
                          const float* w0, const float* w, float* out, void* stream);

/* ---- R3  FMLayer.call, DL/layer/interaction.py:106-114 -----------------------
 * dense form over x[B,n] (DeepFM's FM term, DL/model/deepFM.py:28; also the
 * reference FM model's dense one-hot matrix -- zeros are skipped):
 *   out[b] = w0 + x@w1 + 0.5*sum_k[(x@v)_k^2 - (x^2 @ v^2)_k]          v:[n,kfm]
 * apply_sigmoid != 0 adds FM.call's sigmoid (DL/model/fm.py:22). */
int ctrhot_fm_dense_f32(const float* x, int64_t x_ld, int64_t n, int64_t B,
                        const float* w0, const float* w1, const float* v, int kfm,
                        float* out, int apply_sigmoid, void* stream);
/* ctrhot_fm_dense_f32 for few rows of many columns (the reference's FM on its dense one-hot matrix, DL/model/fm.py:26-45:
 * 2 000 x 43 604): the columns are split across CTAs so that the launch fills the GPU; partial sums in `workspace`
 * (ctrhot_fm_dense_ws_bytes; 0 = the shape does not need it), added in a fixed order by a finishing pass.  Falls through to
 * ctrhot_fm_dense_f32 when the workspace is absent / too small. */
int64_t ctrhot_fm_dense_ws_bytes(int64_t n, int64_t B, int kfm);
int ctrhot_fm_dense_ws_f32(const float* x, int64_t x_ld, int64_t n, int64_t B, const float* w0, const float* w1,
                           const float* v, int kfm, float* out, int apply_sigmoid, void* workspace,
                           int64_t workspace_bytes, void* stream);
/* one-hot form = north_star's (sum v)^2 - sum v^2 over gathered rows */
int ctrhot_fm_onehot_f32(const float* dense, int64_t dense_ld, int D, const int32_t* dense_col,
                         const void* ids, int ids_kind, int64_t ids_ld, int F,
                         const int64_t* col_off, int64_t B,
                         const float* w0, const float* w1, const float* v, int kfm,
                         float* out, int apply_sigmoid, void* stream);

/* ---- R5  FFMLayer.call, DL/layer/interaction.py:141-163 ----------------------
 * w:[n_feat], v:[n_feat, field_num, k]; features = D dense columns then the F
 * one-hot blocks (col_off[f] = D + sum_{g<f} V_g, col_off[F] = n_feat).  Ids
 * outside [0,V_f) contribute nothing (tf.one_hot semantics). */
int ctrhot_ffm_onehot_f32(const float* dense, int64_t dense_ld, int D,
                          const void* ids, int ids_kind, int64_t ids_ld, int F,
                          const int64_t* col_off, int64_t B,
                          const float* w0, const float* w, const float* v, int field_num, int k,
                          float* out, int apply_sigmoid, void* stream);

/* ---- R6/R7  InnerProductLayer / OuterProductLayer, interaction.py:166-215 ----
 * e:[B,F,k] (ld = elements per sample), out[b,p] over row-major pairs i<j.
 * outer: out[b,p] = sum_{a,c} e[b,j,a] * W[a,p,c] * e[b,i,c],  W:[k,P,k]. */
int ctrhot_pnn_inner_f32(const float* e, int64_t e_ld, int F, int k, int64_t B,
                         float* out, int64_t out_ld, void* stream);
int ctrhot_pnn_outer_f32(const float* e, int64_t e_ld, int F, int k, int64_t B, const float* W,
                         float* out, int64_t out_ld, void* stream);

/* ---- R8  CrossLayer.call, DL/layer/interaction.py:75-83 ----------------------
 * all L layers in one launch; w,b:[L,d].  out may alias x.
 * If final_w != NULL also writes dot[b] = sum_j out[b,j]*final_w[j] (the cross
 * half of DCN's last Dense, DL/model/dcn.py:32-33); out may then be NULL. */
int ctrhot_cross_f32(const float* x, int64_t x_ld, int d, int64_t B, int L,
                     const float* w, const float* b, float* out, int64_t out_ld,
                     const float* final_w, float* dot, void* stream);

/* ---- R4/R10  DNNLayer / Dense(+PReLU/Dice) towers, interaction.py:30-46 ------
 * One descriptor per Dense layer.  y = act(x @ weight + bias).
 *  weight  [in_dim, out_dim] row-major = the Keras `kernel`
 *  act_param PRELU: alpha[alpha_rows, out_dim]; row = (global row % alpha_rows)
 *            (alpha_rows = T for DIN's attention Dense, interaction.py:366; 1 otherwise)
 *            DICE : [3, out_dim] = alphas, moving_mean, moving_var (eps in dice_eps)
 * packed  NULL, or the buffer filled by ctrhot_mlp_pack() for the tensor-core path. */
typedef struct {
  const float* weight;
  const float* bias;
  const float* act_param;
  const void* packed;
  int32_t in_dim, out_dim, act, alpha_rows;
  float dice_eps;
  int32_t reserved;
} ctrhot_dense_desc;

/* exact fp32 CUDA-core path (FFMA, fp32 accumulate) */
int ctrhot_mlp_f32(const float* x, int64_t x_ld, int64_t M, const ctrhot_dense_desc* h_layers,
                   int n_layers, float* out, int64_t out_ld, void* workspace,
                   int64_t workspace_bytes, void* stream);
/* ctrhot_mlp_f32 without any allocation, by the caller or by the library.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arithmetic, same kernels and same results as ctrhot_mlp_f32; the hidden
 * activations ping-pong through a file-scope device array of CTRHOT_MLP_STATIC_SCRATCH_BYTES (16 MiB) that every GPU which
 * loads the library holds until exit.  Needs 2 * M * round32(widest hidden layer) * 4 bytes of it: more returns
 * CTRHOT_ESHAPE with the sizes in ctrhot_last_cuda_error() (e.g. 8 192 rows at 256 hidden units fit exactly).  Calls no
 * cudaMalloc / cudaFree / allocating library routine; the first call on each device resolves the array's device address
 * (cudaGetSymbolAddress) -- make that call outside a graph capture.  ALL CALLS SHARE THE ONE ARRAY PER DEVICE: calls that
 * can overlap on the GPU (different streams, no ordering between them) race on it -- serialise them.  Nothing in the array
 * needs clearing between calls: every element is written by layer l before layer l + 1 reads it. */
#define CTRHOT_MLP_STATIC_SCRATCH_BYTES (16 << 20)
int ctrhot_mlp_f32_static(const float* x, int64_t x_ld, int64_t M, const ctrhot_dense_desc* h_layers,
                          int n_layers, float* out, int64_t out_ld, void* stream);
/* tcgen05/TMEM TF32 path (fp32 in HBM, TF32 multiply, fp32 accumulate) */
int64_t ctrhot_mlp_pack_bytes(int in_dim, int out_dim);
int ctrhot_mlp_pack(const float* weight, int in_dim, int out_dim, void* packed, void* stream);
int ctrhot_mlp_tf32(const float* x, int64_t x_ld, int64_t M, const ctrhot_dense_desc* h_layers,
                    int n_layers, float* out, int64_t out_ld, void* workspace,
                    int64_t workspace_bytes, void* stream);
int64_t ctrhot_mlp_workspace_bytes(int64_t M, const ctrhot_dense_desc* h_layers, int n_layers);
/* tuning hook: d_counters = device buffer of 148 * 4 uint64 that the following ctrhot_mlp_tf32 launches fill per CTA with
 * {cycles the MMA thread waited for an operand stage, for a free accumulator, its total cycles, steps}; NULL = off */
int ctrhot_debug_tc_trace(uint64_t* d_counters);

/* Whole tower in ONE persistent tcgen05 kernel: hidden activations stay in TMEM (layer l's epilogue
 * writes layer l+1's A operand back with tcgen05.st), weights stream through the TMA engine.
 * h_layers = 1..3 hidden Dense (out_dim 8..256, packed) + a final Dense(h -> 1, no activation).
 *   out[m] = post(tower(x_m)),  post(y) = sigmoid?( scale * ( y * (y_scale?*y_scale:1) + (add_vec?add_vec[m]:0)
 *                                                      + (add_bias?*add_bias:0) ) )
 * (DeepFM: add_vec = FM term, scale 0.5, sigmoid -- deepFM.py:30; DCN: y_scale/add_vec/add_bias = the
 * final Dense over [cross, dnn] -- dcn.py:32-33).  _gather: rows come straight from the embedding
 * tables (x = [embeddings | dense] internal order never exists in HBM); first layer packed from the
 * permuted kernel.  CTRHOT_ESHAPE when the shape does not fit (use ctrhot_mlp_tf32 then). */
int ctrhot_tower_fused_tf32(const float* x, int64_t x_ld, int64_t M,
                            const ctrhot_dense_desc* h_layers, int n_layers,
                            const float* y_scale, const float* add_vec, const float* add_bias,
                            float scale, int apply_sigmoid, float* out, void* stream);
int ctrhot_tower_gather_tf32(const float* arena, const int64_t* row_off, int F, int k,
                             const void* ids, int ids_kind, int64_t ids_ld,
                             const float* dense, int64_t dense_ld, int D, int64_t M,
                             const ctrhot_dense_desc* h_layers, int n_layers,
                             const float* y_scale, const float* add_vec, const float* add_bias,
                             float scale, int apply_sigmoid, float* out, int* err_flag, void* stream);

/* ---- R12/R14 whole-model kernel: gather -> {FM | cross} -> tower -> sigmoid in ONE persistent launch ------
 * (DeepFM.call, DL/model/deepFM.py:23-31; DCN.call, DL/model/dcn.py:24-34; EmbedLayer -> DNNLayer towers).
 * Each embedding row is read from HBM once; x = [dense, embed] and the hidden activations never reach HBM.
 * h_layers = 1..3 hidden Dense (out_dim 8..256, act none/relu/PReLU[out_dim]) + the final Dense(h -> 1).
 *   h_layers[0].packed = ctrhot_fused_pack(W1 [n, n1], side [n, 16] or NULL): W1 (TF32) plus 16 "side" columns whose
 *                        products x @ side are formed fp32-accurately on the tensor pipe (split-TF32, three MMAs);
 *   h_layers[l>0].packed = ctrhot_mlp_pack.
 * sq_w  [round32(n)] internal column order [embeddings | dense | 0] or NULL:  q = sum_j x_j^2 sq_w[j]   (fp32).
 * mode 0 (tower):  out = y
 * mode 1 (DeepFM): side[:, c < n_side] = v[:, c], side[:, n_side] = w1, sq_w[j] = sum_c v[j,c]^2, h_combine = {w0}:
 *                  out = 0.5 * (w0 + S[n_side] + 0.5 * (sum_c S[c]^2 - q) + y)              (interaction.py:106-114)
 * mode 2 (DCN):    side[:, l < n_side] = cross w_l, side[:, n_side] = final Dense kernel over the cross half,
 *                  h_combine = {c_l . w_l (l < n_side), c_L . final_w, wy, bias} with c_l = sum_{i<l} b_i:
 *                  a = 1; a += a * S[l] + h_combine[l];  out = a * S[L] + h_combine[L] + wy * y + bias
 *                  (the L cross layers of interaction.py:75-83 collapse to these scalars because x_l = a_l x_0 + c_l)
 * apply_sigmoid != 0 applies the model's final sigmoid.  h_combine is a HOST array (<= 24 floats).
 * cluster: 0 = auto, else 1 | 2 | 4 CTAs share each weight chunk through TMA multicast.  OR-ing CTRHOT_LAUNCH_OVERLAP into it
 *          allows this launch to begin while the PREVIOUS launch of this entry point in the same stream (made with the same
 *          flag) is still finishing its last tiles (programmatic dependent launch): 65 536 samples are 512 tiles on 148 SMs,
 *          so 80 SMs would otherwise idle through the last quarter of every launch.  The caller promises that the two
 *          launches are independent: no input of this one (ids, dense, records, tables, weights) is written by that one, and
 *          `out` buffers differ.  Launches made without the flag are ordered as usual.
 * CTRHOT_ESHAPE when the tower does not fit (TMEM columns / shared memory / activation kind): use the layered path. */
int64_t ctrhot_fused_pack_bytes(int n, int n1);
int ctrhot_fused_pack(const float* W1, const float* side, int n, int n1, int D, void* packed, void* stream);
int ctrhot_ctr_fused_tf32(const float* arena, const int64_t* row_off, int F, int k,
                          const void* ids, int ids_kind, int64_t ids_ld,
                          const float* dense, int64_t dense_ld, int D, int64_t M,
                          const ctrhot_dense_desc* h_layers, int n_layers,
                          const float* sq_w, int mode, int n_side, const float* h_combine, int n_combine,
                          int apply_sigmoid, int cluster, float* out, int* err_flag, void* stream);
/* ---- 8(f) rank 4, the remaining towers (each a Dense stack -- ctrhot_mlp_* / ctrhot_tower_* -- plus one pass below) --------
 * ctrhot_residual_relu_f32: out = relu(x + y)                               ResLayer.call, DL/layer/interaction.py:271-278
 * ctrhot_mmoe_combine_f32 : out[t][b, h] = sum_e softmax(gate[b, t*E : (t+1)*E])[e] * expert[b, h*E + e]
 *                           mmoe_layer.call, DL/layer/interaction.py:481-509.  expert = relu(x @ expert_matrix.view(in, H*E) +
 *                           expert_bias.view(H*E)) -- the reference's [in, hidden, E] weight IS that [in, H*E] matrix in memory;
 *                           gate = x @ [gate_0 | .. | gate_{T-1}] + biases.  out: T matrices [B, out_ld] out_task_stride apart.
 * ctrhot_lau_concat_f32   : out[b*T + t, :] = [q_b, k_bt, q_b - k_bt, q_b * k_bt]   LocalActivationUnit.call, DL/layer/core.py:96-101 */
int ctrhot_residual_relu_f32(const float* x, int64_t x_ld, const float* y, int64_t y_ld, int64_t M, int N,
                             float* out, int64_t out_ld, void* stream);
int ctrhot_mmoe_combine_f32(const float* expert, int64_t e_ld, const float* gate, int64_t g_ld, int64_t B, int H,
                            int E, int T, float* out, int64_t out_task_stride, int64_t out_ld, void* stream);
int ctrhot_lau_concat_f32(const float* query, int64_t q_ld, const float* keys, int64_t k_ld, int64_t B, int T, int K,
                          float* out, int64_t out_ld, void* stream);
/* ---- 8(f) rank 2, Hash.call without a vocabulary file, DL/layer/utils.py:35-56 -------------------------------------------
 * out[i] = FarmHash Fingerprint64(str(x[i])) % nb   with nb = num_buckets - (mask_zero ? 1 : 0)   (tf.strings.to_hash_bucket_fast),
 * shifted up by one with the value "0" -> 0 when mask_zero (utils.py:52-54).  _int: x = DEVICE int32 / int64 [N], hashed as its
 * decimal string (tf.as_string, :37-38).  _bytes: strings of len[i] <= 64 bytes at s + i * stride (longer: out = -1, *err_flag).
 * Pinned on the example in TensorFlow's documentation of the op (["Hello", "TensorFlow", "2.x"], 3 -> [0, 2, 2]): lengths 1-16,
 * i.e. every integer of up to 16 digits; the 17-64 byte branches follow the published FarmHash source unpinned. */
int ctrhot_hash_bucket_int(const void* x, int is_i64, int64_t N, int64_t num_buckets, int mask_zero, int64_t* out, void* stream);
int ctrhot_hash_bucket_bytes(const uint8_t* s, int64_t stride, const int32_t* len, int64_t N, int64_t num_buckets,
                             int mask_zero, int64_t* out, int* err_flag, void* stream);
/* ---- 8(f) rank 2, the input step: create_criteo_dataset, DL/utils/dataset.py:36-66 (read_csv :38, fillna :42-43,
 * MinMaxScaler :46, LabelEncoder per column :51-52) -----------------------------------------------------------------------
 * ctrhot_criteo_parse_tsv: text = the file's bytes on the DEVICE; line_start = device int64 [N + 1] (byte offset of every
 *   line, then the end of the last one).  Per line: label (int8, when has_label), 13 counts -> dense_raw float64 [N, 13] (NaN
 *   = empty field, as pandas reads it), 26 tokens -> int64 [N, 26]: the 32-bit value of the 8 lowercase hex digits, -1 for an
 *   empty field (for fixed-width lowercase hex the string order LabelEncoder sorts by is the numeric order, and the fill value
 *   '-1' sorts first).  Malformed fields set *err_flag.
 * ctrhot_criteo_encode: dense[r, c] = fp32(x * d_scale[c] + d_min[c]) with x = 0 for NaN, evaluated in float64 as
 *   MinMaxScaler.transform does (two rounded operations); ids[r, f] = index of tokens[r, f] in vocab[vocab_off[f] ..
 *   vocab_off[f + 1]) (sorted distinct values of the column = LabelEncoder.classes_), -1 and *err_flag when absent
 *   (LabelEncoder raises on unseen labels).  All pointers DEVICE. */
int ctrhot_criteo_parse_tsv(const void* text, const int64_t* line_start, int64_t N, int has_label, int8_t* label,
                            double* dense_raw, int64_t* tokens, int* err_flag, void* stream);
int ctrhot_criteo_encode(const double* dense_raw, const int64_t* tokens, int64_t N, const double* d_scale,
                         const double* d_min, const int64_t* vocab, const int64_t* vocab_off, float* dense,
                         int64_t dense_ld, int32_t* ids, int64_t ids_ld, int* err_flag, void* stream);
/* ---- transport records: the copy on the left of the path (the reference passes one float64 [B, 39]
 * matrix, DL/utils/dataset.py:61, DL/model/deepFM.py:24 -- 312 B per sample; the int32 fast path above takes 156 B) --------
 *   record = [ dense: D x float32 (bit-exact) | id_f in width[f] little-endian bytes, f = 0..F-1 | zero pad to 4 bytes ]
 *   width[f] = 1 / 2 / 3 / 4 bytes for feat_onehot_dim <= 2^8 / 2^16 / 2^24 / 2^31      (104 B on Criteo-Kaggle cardinalities)
 * ctrhot_rows_layout   : HOST.  h_vocab [F] -> h_desc [F] = (byte offset in the record) | width << 16; returns the record
 *                        size in bytes (> 0, a multiple of 4) or a negative status.  Dense block at byte 0.
 * ctrhot_rows_pack_host: HOST, no GPU work.  (dense [B, D], ids int32 [B, F]) -> records, n_threads host threads (<= 0: all).
 *                        CTRHOT_EID if an id does not fit its width (that id is stored as 0, never truncated silently).
 * ctrhot_rows_unpack   : DEVICE records -> dense float32 [M, dense_ld], ids int32 [M, ids_ld]; desc = DEVICE int32 [F].
 *                        The generic way into every (dense, ids) entry point: one extra launch.
 * ctrhot_ctr_fused_packed_tf32: ctrhot_ctr_fused_tf32 reading the records directly -- its gather warps decode the ids,
 *                        the dense columns are read in place: ONE host-to-device copy per batch, no unpack pass.
 *                        records 16-byte aligned, rec_bytes % 4 == 0, pack_desc = DEVICE int32 [F]. */
int ctrhot_rows_layout(const int64_t* h_vocab, int F, int D, int32_t* h_desc);
/* the same at BIT granularity (width = ceil(log2(vocab)) bits): h_desc[f] = 0x80000000 | bits << 24 | bit offset; every consumer of
 * a descriptor (pack, unpack, the whole-model kernels) takes either form.  96 B per sample on the Criteo-Kaggle cardinalities. */
int ctrhot_rows_layout_bits(const int64_t* h_vocab, int F, int D, int32_t* h_desc);
int ctrhot_rows_pack_host(const float* h_dense, int64_t dense_ld, const int32_t* h_ids, int64_t ids_ld, int64_t B,
                          const int32_t* h_desc, int F, int D, int64_t rec_bytes, void* h_records, int n_threads);
int ctrhot_rows_unpack(const void* records, int64_t rec_bytes, const int32_t* desc, int F, int dense_off, int D,
                       int64_t M, float* dense, int64_t dense_ld, int32_t* ids, int64_t ids_ld, void* stream);
int ctrhot_ctr_fused_packed_tf32(const float* arena, const int64_t* row_off, int F, int k,
                                 const void* records, int64_t rec_bytes, const int32_t* pack_desc,
                                 int dense_off, int D, int64_t M,
                                 const ctrhot_dense_desc* h_layers, int n_layers,
                                 const float* sq_w, int mode, int n_side, const float* h_combine, int n_combine,
                                 int apply_sigmoid, int cluster, float* out, int* err_flag, void* stream);
/* Plans: ctrhot_ctr_fused_tf32 / _packed_tf32 with everything but the batch frozen, for loops that call them every ~70 us
 * (the argument checks, shared-memory sizing and cluster choice are done once).  create: arguments as for the two entry points
 * above -- ids_kind I32 / F32 with ids_ld, dense_ld; or PACKED with ids_ld = record bytes, pack_desc, dense_off.  The plan is
 * a few hundred bytes of HOST memory (destroy frees it); the device pointers inside (tables, packed weights, sq_w, pack_desc,
 * err_flag) must stay valid and unchanged -- re-create the plan when weights are re-packed.
 * run: `batch` = ids [M, ids_ld] (with `dense` [M, dense_ld]) or the records (then `dense` is ignored). */
int ctrhot_fused_plan_create(const float* arena, const int64_t* row_off, int F, int k,
                             int ids_kind, int64_t ids_ld, const int32_t* pack_desc, int dense_off,
                             int64_t dense_ld, int D,
                             const ctrhot_dense_desc* h_layers, int n_layers,
                             const float* sq_w, int mode, int n_side, const float* h_combine, int n_combine,
                             int apply_sigmoid, int cluster, int* err_flag, void** plan);
int ctrhot_fused_plan_run(const void* plan, const void* batch, const float* dense, int64_t M, float* out, void* stream);
int ctrhot_fused_plan_destroy(void* plan);
/* The same kernel over ROW-SHARDED tables (8(e), config 5): arenas = HOST array of G device pointers (arena of owner g;
 * this rank's own at index `rank`, the peers' mapped by ctrhot_peer_open), row_off = device int64 [G, F+1], ids = GLOBAL
 * row numbers: row id of field f is read at arenas[id mod G] + (row_off[id mod G][f] + id div G) * k -- the gather warps
 * pull remote rows over NVLink/NVSwitch with the same 16-byte async copies, under the tensor work of earlier tiles.
 * No all-to-all, no x in HBM.  G <= 16. */
int ctrhot_ctr_fused_sharded_tf32(const void* const* arenas, const int64_t* row_off, int G, int rank, int F, int k,
                                  const void* ids, int ids_kind, int64_t ids_ld,
                                  const float* dense, int64_t dense_ld, int D, int64_t M,
                                  const ctrhot_dense_desc* h_layers, int n_layers,
                                  const float* sq_w, int mode, int n_side, const float* h_combine, int n_combine,
                                  int apply_sigmoid, int cluster, float* out, int* err_flag, void* stream);

/* WideDeep.call (DL/model/wideDeep.py:24-34): out[m] = (sigmoid of) scale * (add_vec[m] + DNN([Embed(ids), dense])[m]) --
 * the deep tower fed straight from the embedding arena by the persistent kernel; add_vec (device [M], may be NULL) is the
 * wide logit from ctrhot_wide_onehot_f32.  h_layers[0].packed = ctrhot_fused_pack(W1, NULL, n, n1, D). */
int ctrhot_ctr_tower_tf32(const float* arena, const int64_t* row_off, int F, int k,
                          const void* ids, int ids_kind, int64_t ids_ld,
                          const float* dense, int64_t dense_ld, int D, int64_t M,
                          const ctrhot_dense_desc* h_layers, int n_layers,
                          const float* add_vec, float scale, int apply_sigmoid, int cluster,
                          float* out, int* err_flag, void* stream);

/* WideDeep.call in ONE launch: as ctrhot_ctr_tower_tf32 (no dense columns in the deep input), with the wide logit
 * w0 + sum_j dense[m, j] dw[j] + sum_f w[wide_off[f] + ids[m, f]] formed inside the kernel: the gather warps fetch the 4-byte
 * weights next to the embedding rows, the stager threads add them up per row; out = (sigmoid of) scale * (tower + wide).
 * wide_off: device int64 [F + 1] -- field f's one-hot block is [wide_off[f], wide_off[f+1]); an id at or beyond that width
 * contributes 0 (a get_dummies column that does not exist, dataset.py:53-58), exactly as ctrhot_wide_onehot_f32 does;
 * h_dw: HOST float [wide_D <= 16]; k a power of two >= 8. */
int ctrhot_widedeep_fused_tf32(const float* arena, const int64_t* row_off, int F, int k,
                               const void* ids, int ids_kind, int64_t ids_ld, int64_t M,
                               const ctrhot_dense_desc* h_layers, int n_layers,
                               const float* wide_w, const int64_t* wide_off, const float* wide_dense,
                               int64_t wide_dense_ld, int wide_D, const float* h_dw, float w0, float scale,
                               int apply_sigmoid, int cluster, float* out, int* err_flag, void* stream);

/* DNNLayer.call (DL/layer/interaction.py:40-46) on a matrix x [M, K] (row stride x_ld, x_ld % 4 == 0, x 16-byte aligned)
 * that already lives in HBM, by the same persistent kernel: out[m] = (sigmoid of) Dense(h->1)(tower(x[m, :])).
 * h_layers as for ctrhot_ctr_fused_tf32 with h_layers[0].packed = ctrhot_fused_pack(W1r, NULL, K, n1, K % 4) where
 * W1r = rows [K - K % 4, K) of W1 followed by rows [0, K - K % 4)  (the kernel reads the 16-byte pieces first). */
int ctrhot_tower_rows_tf32(const float* x, int64_t x_ld, int K, int64_t M,
                           const ctrhot_dense_desc* h_layers, int n_layers, int apply_sigmoid, int cluster,
                           float* out, void* stream);
/* One Dense layer (8 <= out_dim <= 416, act none/relu/PReLU[out_dim], packed by ctrhot_fused_pack(W, NULL, K, out_dim,
 * K % 4, ..) from W with its last K % 4 rows moved to the front) over x [M, x_ld >= K] in HBM through the same persistent
 * kernel (two MMA column halves when out_dim > 256): h[m, 0:out_dim] = tf32(act(x_m W + b)), h[m, out_dim:h_ld) = 0.
 * Chains layers too wide for TMEM through HBM (DCN cfg 3's 400-wide tower, DL/model/dcn.py:45); the last hidden layer
 * and the final Dense(h -> 1) then go through ctrhot_tower_rows_tf32 (one hidden layer may be up to 416 wide there). */
/* CHAINED LAUNCHES (wait_flags / signal_flags / epoch; all NULL / 0 = plain stream order): each launch publishes tile t (rows
 * [128 t, 128 t + 128)) of its output by storing `epoch` into signal_flags[t] (int32 [ceil(M / 128)], device), and a launch whose
 * input is another launch's output waits for wait_flags[t] == epoch before its first read of tile t.  With the flags in place
 * `cluster` may carry CTRHOT_LAUNCH_OVERLAP: launch l + 1 then starts on the SMs launch l frees while it finishes its last
 * tiles (512 tiles on 148 SMs leave 80 SMs idle for the last quarter of every launch otherwise).  Use a new epoch per
 * forward pass and keep x / h in buffers that the overlapping neighbour launches do not write (the caller's contract:
 * layer/core.py alternates two workspaces). */
int ctrhot_dense_rows_tf32(const float* x, int64_t x_ld, int K, int64_t M, const ctrhot_dense_desc* layer,
                           float* h, int64_t h_ld, int cluster, const int32_t* wait_flags, int32_t* signal_flags,
                           int epoch, void* stream);
/* ctrhot_tower_rows_tf32 with an affine post:
 *   out[m] = (sigmoid of) y_scale * Dense(h->1)(tower(x[m, :])) + add_scale * add_vec[m] + add_const     (add_vec may be NULL) */
int ctrhot_tower_rows_post_tf32(const float* x, int64_t x_ld, int K, int64_t M,
                                const ctrhot_dense_desc* h_layers, int n_layers,
                                const float* add_vec, float y_scale, float add_scale, float add_const,
                                int apply_sigmoid, int cluster, float* out, const int32_t* wait_flags, int epoch,
                                void* stream);
/* DCN.call (DL/model/dcn.py:24-34) with a tower too wide to chain through TMEM (cfg 3: 400-400-400), FIRST launch:
 * gather -> { the L <= 3 cross layers collapsed to scalars (see mode 2 of ctrhot_ctr_fused_tf32), exact fp32 on the CUDA
 * cores | first Dense layer (out_dim <= 416) on the tensor pipe }.  h[m, :] = tf32(act(x_m W1 + b1)) and
 * side[m] = a_L (x_m . final_w) + c_L . final_w reach HBM; x = [dense | embed] never does.
 * layer->packed = ctrhot_fused_pack(W1, NULL, n, out_dim, D); cc_w = device [round32(n), 4], INTERNAL column order
 * [embeddings | dense | 0], columns {w_0 .. w_{L-1}, final_w}; h_combine as for mode 2.  The remaining layers run as
 * ctrhot_dense_rows_tf32 and ctrhot_tower_rows_post_tf32(add_vec = side, y_scale = wy, add_scale = 1, add_const = bias). */
int ctrhot_dcn_first_layer_tf32(const float* arena, const int64_t* row_off, int F, int k,
                                const void* ids, int ids_kind, int64_t ids_ld,
                                const float* dense, int64_t dense_ld, int D, int64_t M,
                                const ctrhot_dense_desc* layer, const float* cc_w, int n_side,
                                const float* h_combine, int n_combine, float* h, int64_t h_ld, float* side_out,
                                int cluster, int* err_flag, int32_t* signal_flags, int epoch, void* stream);

/* ---- 8(f) rank 3, DCN under compile_fit (DL/model/dcn.py:15-34, CrossLayer DL/layer/interaction.py:49-83) ---------------------
 * One step of keras SGD(lr) on mean clipped BCE + CrossLayer's l2 regularisers (reg_w sum w_l^2 + reg_b sum b_l^2 per layer),
 * updating in place: the embedding rows that were looked up, the L cross weights / biases (h_cross_w / h_cross_b: HOST arrays of
 * L device pointers to the [n, 1] vectors), the tower (h_layers as for ctrhot_deepfm_train_step_f32: ReLU hidden layers with
 * Dropout, then Dense(h -> 1)) and the final Dense(n + 1 -> 1) (final_k [n + 1], final_b [1]).  Cross backward per layer, with
 * g = dL/dx_{l+1} and c = g . x_0:  dw_l = sum_b c x_l,  db_l = sum_b g,  dL/dx_l = g + c w_l,  dL/dx_0 += g s_l.
 * n = D + F k <= 1024, L <= 8.  loss_out = {mean BCE before the update, correct predictions}.  fp32 CUDA cores. */
int64_t ctrhot_dcn_train_workspace_bytes(int64_t B, int n, int L, const ctrhot_dense_desc* h_layers, int n_layers);
int ctrhot_dcn_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                              int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y, int64_t B,
                              const int64_t* h_cross_w, const int64_t* h_cross_b, int L, float reg_w, float reg_b,
                              const ctrhot_dense_desc* h_layers, int n_layers, float* final_k, float* final_b,
                              float lr, float dropout, uint64_t seed, float* loss_out, void* workspace,
                              int64_t workspace_bytes, int* err_flag, void* stream);

/* ---- 8(f) rank 3, PNN's own training loop, DL/model/pnn.py:74-85 (batch 32, SGD(0.01), loss = mean binary_crossentropy(y, model(X))
 * with PNN.call's RAW output: keras clips it to [1e-7, 1 - 1e-7], so d loss / dz = (z - y) / (z (1 - z)) / B inside the clip and 0
 * outside; tape.gradient(loss) leaves the layers' regularisers out).  mode: 1 inner, 2 outer, 3 both (tower input =
 * [rows | inner | outer], no FGCNN branch).  Product layers backward, pair p = (i, j), i < j, upstream gi_p / go_p:
 * de_i += gi e_j + go W_p^T e_j,  de_j += gi e_i + go W_p e_i,  dW[a, p, c] = sum_b go e_j[a] e_i[c].  Updates in place: the
 * looked-up arena rows, outer_w [k, P, k], the tower.  loss_out = {loss before the update, predictions with z > 0.5 == y}. */
int64_t ctrhot_pnn_train_workspace_bytes(int64_t B, int F, int k, int mode, const ctrhot_dense_desc* h_layers, int n_layers);
int ctrhot_pnn_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                              int64_t ids_ld, const float* y, int64_t B, float* outer_w, int mode,
                              const ctrhot_dense_desc* h_layers, int n_layers, float lr, float dropout,
                              uint64_t seed, float* loss_out, void* workspace, int64_t workspace_bytes,
                              int* err_flag, void* stream);

/* ---- 8(f) rank 3, WideDeep under compile_fit (DL/model/wideDeep.py:14-34; WideLayer DL/layer/interaction.py:11-26 with its
 * l2(1e-4) on w) -- p = sigmoid(0.5 (wide + deep)), the wide input [dense | dense again | one 0/1 block per field]
 * (DL/utils/dataset.py:53-58) never materialised: col_off = DEVICE int64 [F + 1], first wide column of every field (2 D + ...),
 * ids at or beyond a field's width have no column.  Updates in place: w0, w [n_wide] (decay on all of it, the dense columns
 * twice, the looked-up one-hot columns by atomic adds), the arena rows, the tower (which sees the embeddings only). */
int64_t ctrhot_widedeep_train_workspace_bytes(int64_t B, int n, const ctrhot_dense_desc* h_layers, int n_layers);
int ctrhot_widedeep_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                   int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y,
                                   int64_t B, float* wide_w0, float* wide_w, const int64_t* col_off, int64_t n_wide,
                                   float reg_w, const ctrhot_dense_desc* h_layers, int n_layers, float lr,
                                   float dropout, uint64_t seed, float* loss_out, void* workspace,
                                   int64_t workspace_bytes, int* err_flag, void* stream);

/* ---- 8(f) rank 3, DeepCrossing under compile_fit (DL/model/deepCrossing.py:14-32; ResLayer DL/layer/interaction.py:261-278) ---
 * x_0 = [dense | rows]; unit r: x_{r+1} = relu(x_r + Dense_n(relu-Dense chain(x_r))); p = sigmoid(Dense_1(x_R)); mean clipped BCE,
 * no regularisers, no dropout.  h_layers = n_res groups of per_res descriptors (per_res - 1 ReLU layers, then the unit's Dense back
 * to width n = D + F k, act none; every layer with a bias).  Updates in place: arena rows, all units, final_k [n], final_b [1]. */
int64_t ctrhot_deepcrossing_train_workspace_bytes(int64_t B, int n, const ctrhot_dense_desc* h_layers, int n_res, int per_res);
int ctrhot_deepcrossing_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                       int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y,
                                       int64_t B, const ctrhot_dense_desc* h_layers, int n_res, int per_res,
                                       float* final_k, float* final_b, float lr, float* loss_out, void* workspace,
                                       int64_t workspace_bytes, int* err_flag, void* stream);

/* ---- 8(f) ranks 1 + 3, NFM under compile_fit (DL/model/nfm.py:15-34 in training mode) -----------------------------------------
 * x = [dense | 0.5 ((sum_f e_f)^2 - sum_f e_f^2)]; keras BatchNormalization on the BATCH's mean and biased variance (eps 1e-3), the
 * moving statistics moved by m <- momentum m + (1 - momentum) stat (0.99); DNNLayer (ReLU + Dropout hidden layers, then
 * Dense(output_dim), act none = the last descriptor); final Dense(output_dim -> 1); sigmoid; mean clipped BCE, no regularisers.
 * Backward through all of it, the batch statistics included: gx = gamma rstd (g - mean_b g - xhat mean_b (g xhat)),
 * d bi_c / d e_fc = S_c - e_fc.  Updates in place: arena rows, gamma / beta / moving stats, tower, final_k [output_dim], final_b. */
int64_t ctrhot_nfm_train_workspace_bytes(int64_t B, int D, int F, int k, const ctrhot_dense_desc* h_layers, int n_layers);
int ctrhot_nfm_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                              int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y, int64_t B,
                              float* bn_gamma, float* bn_beta, float* bn_mean, float* bn_var, float momentum, float eps,
                              const ctrhot_dense_desc* h_layers, int n_layers, float* final_k, float* final_b,
                              float lr, float dropout, uint64_t seed, float* loss_out, void* workspace,
                              int64_t workspace_bytes, int* err_flag, void* stream);

/* ---- 8(f) ranks 1 + 3, AFM under compile_fit (DL/model/afm.py; AFMLayer DL/layer/interaction.py:322-351) --------------------------
 * x = pool over pairs of e_i * e_j (mode 0 mean, 1 'att' = sum: the softmax over a size-1 axis, :316, makes every weight 1 and leaves
 * AttentionLayer's Dense layers without gradient; 2 max), q = sigmoid(x w + b), p = sigmoid(q) (n_sigmoid 2: AFM.call) or q
 * (1: AFMLayer alone); mean clipped BCE.  Updates in place: arena rows (sum / mean: d / d e_fc = S_c - e_fc; max: the arg-max
 * pair only), out_k [k], out_b [1]. */
int64_t ctrhot_afm_train_workspace_bytes(int64_t B, int F, int k);
int ctrhot_afm_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                              int64_t ids_ld, const float* y, int64_t B, int mode, int n_sigmoid, float* out_k,
                              float* out_b, float lr, float* loss_out, void* workspace, int64_t workspace_bytes,
                              int* err_flag, void* stream);

/* ---- 8(f) rank 3, FFM under compile_fit (DL/model/ffm.py:14-23; FFMLayer DL/layer/interaction.py:117-163) ------------------------
 * p = sigmoid(w0 + x.w + sum_{f<g} <ff_f, ff_g>), ff[f, :] = sum_i x_i v[i, f, :] over the one-hot expansion x of [dense | ids]
 * (never materialised: D counts and one row per sparse field; ids outside their column block give tf.one_hot's zero row).
 * Mean clipped BCE + l2(reg_w)(w) + l2(reg_v)(v); one step of SGD(lr), in place: w0 [1], w [n_features], v [n_features,
 * field_num, k] (all of w and v decay by the regularisers; the batch's rows receive x_i (S - ff_f) through atomics). */
int64_t ctrhot_ffm_train_workspace_bytes(int64_t B, int field_num, int k);
int ctrhot_ffm_train_step_f32(const float* dense, int64_t dense_ld, int D, const void* ids, int ids_kind, int64_t ids_ld,
                              int F, const int64_t* col_off, const float* y, int64_t B, float* w0, float* w, float* v,
                              int64_t n_features, int field_num, int k, float reg_w, float reg_v, float lr,
                              float* loss_out, void* workspace, int64_t workspace_bytes, void* stream);

/* ---- 8(f) rank 3, DIN under compile_fit (DL/model/din.py:56-95; Attention DL/layer/interaction.py:369-406) ----------------------
 * One behaviour feature: seq_table [seq_rows, K] serves the history hist [B, T] (int32, 0 = padding) and the candidate cand [B].
 * Forward in TRAINING mode: A = [q, k, q - k, q * k] -> att_layers (Dense + PReLU, act_param = alpha [T, h], alpha_rows = T) ->
 * att_out -> scores, -2^32 + 1 where hist == 0, softmax over T, att = sum_t pw_t k_t; emb = [att | q | other tables' rows |
 * dense] (n columns) -> BatchNormalization on the batch's own mean / biased variance (moving statistics move by `momentum`) ->
 * tower_layers (Dense + PReLU, alpha [h], alpha_rows = 1) -> Dense(1) -> sigmoid; mean clipped BCE; dnn_dropout = 0 (din.py:17).
 * One step of SGD(lr) on EVERY variable, in place: seq_table (history rows and candidate row, atomics), other_tables[i]
 * [other_rows[i], other_dims[i]] (HOST arrays of n_other entries; other_ids [B, n_other] int32 on the device), the layers'
 * weight / bias / act_param (written through the const pointers of the descriptors), att_out_k [h, 1], att_out_b [1], bn_*,
 * out_k [h, 1], out_b [1].  loss_out [2] = mean BCE before the update, number of correct predictions. */
int64_t ctrhot_din_train_workspace_bytes(int64_t B, int T, int K, int n, const ctrhot_dense_desc* att_layers, int n_att,
                                         const ctrhot_dense_desc* tower_layers, int n_tower);
int ctrhot_din_train_step_f32(float* seq_table, int64_t seq_rows, int K, const int32_t* hist, int64_t hist_ld, int T,
                              const int32_t* cand, float* const* other_tables, const int64_t* other_rows,
                              const int32_t* other_dims, int n_other, const int32_t* other_ids, int64_t other_ids_ld,
                              const float* dense, int64_t dense_ld, int D, const float* y, int64_t B,
                              const ctrhot_dense_desc* att_layers, int n_att, float* att_out_k, float* att_out_b,
                              float* bn_gamma, float* bn_beta, float* bn_mean, float* bn_var, float momentum, float eps,
                              const ctrhot_dense_desc* tower_layers, int n_tower, float* out_k, float* out_b, float lr,
                              float* loss_out, void* workspace, int64_t workspace_bytes, int* err_flag, void* stream);

/* ---- R15 PNN front: EmbedLayer -> [ flat | InnerProductLayer | OuterProductLayer ] in ONE launch ---------------------------
 * PNN.call DL/model/pnn.py:28-49 up to the concat the tower reads; InnerProductLayer / OuterProductLayer.call
 * DL/layer/interaction.py:170-183 / 201-215 (SURVEY 8(b) `ctrhot_pnn_fwd`: the gathered rows stay in shared memory).
 *   out[b, 0 : F k)              = arena[row_off[f] + ids[b, f], :]                        (bit-exact)
 *   out[b, F k : F k + P)        = <e_i, e_j>, pairs row-major over i < j                  (if do_inner)
 *   out[b, .. : .. + P)          = sum_a e_j[a] sum_c W[a, p, c] e_i[c]                    W = [k, P, k]
 *   out[b, width : out_ld)       = 0
 * The inner sum of the outer product runs on the tensor pipe as a three-term TF32 split (fp32-accurate: <= 1e-5 relative),
 * the rest on the CUDA cores from the same shared-memory tile.  k == 8 (the reference's embed_dim), 2 <= F <= 28;
 * out 16-byte aligned, out_ld % 4 == 0.  packed = ctrhot_pnn_pack(W) (ctrhot_pnn_pack_bytes bytes, re-pack when W changes).
 * signal_flags / epoch / launch: CHAINED LAUNCHES as for ctrhot_dense_rows_tf32 -- tile t (rows [128 t, 128 t + 128)) of `out` is
 * published by storing epoch into signal_flags[t]; with launch = CTRHOT_LAUNCH_OVERLAP the tower launched behind it
 * (ctrhot_tower_rows_post_tf32 with wait_flags) may start on the SMs this kernel frees.  NULL / 0 / 0 = plain stream order. */
int64_t ctrhot_pnn_pack_bytes(int F, int k);
int ctrhot_pnn_pack(const float* W, int F, int k, void* packed, void* stream);
int ctrhot_pnn_front_f32(const float* arena, const int64_t* row_off, int F, int k,
                         const void* ids, int ids_kind, int64_t ids_ld, int64_t B,
                         const void* packed, int do_inner, float* out, int64_t out_ld,
                         int* err_flag, int32_t* signal_flags, int epoch, int launch, void* stream);
/* The two product layers on an already gathered e [B, F, 8] (row stride e_ld, 16-byte aligned, e_ld % 4 == 0) by the same kernel:
 * out[b, 0:P) = <e_i, e_j> (if do_inner), out[b, .. + P) = e_j^T W_p e_i (columns beyond are not touched).  InnerProductLayer.call /
 * OuterProductLayer.call, DL/layer/interaction.py:170-183 / 201-215, for the reference's embed_dim (k == 8, F <= 28). */
int ctrhot_pnn_products_f32(const float* e, int64_t e_ld, int F, int k, int64_t B, const void* packed, int do_inner,
                            float* out, int64_t out_ld, void* stream);

/* Dice.call, interaction.py:422-425, elementwise over x[rows, h] */
int ctrhot_dice_f32(const float* x, int64_t rows, int h, const float* alphas, const float* mean,
                    const float* var, float eps, float* out, void* stream);

/* ---- R9  DIN Attention.call, DL/layer/interaction.py:369-406 -----------------
 * query [B,K]; key = value [B,T,K] (key_ld_b = elements per sample); mask [B,T] float (0 = padding);
 * a layer with weight == NULL is an element-wise activation only (the 'dice' branch);
 * 'prelu' branch: layers = Dense(4K->h1,PReLU[T,h1]), Dense(h1->h2,PReLU[T,h2]), ...
 * then Dense(h_last->1); masked softmax with fill -2^32+1; out[B,K] = p @ value. */
int ctrhot_din_attention_f32(const float* query, int64_t q_ld, const float* key, int64_t key_ld_b,
                             const float* mask, int64_t B, int T, int K,
                             const ctrhot_dense_desc* h_layers, int n_layers,
                             const float* out_w, const float* out_b,
                             float* out, int64_t out_ld, void* stream);
/* fused: history/candidate gather + attention; keys never reach HBM.
 * hist_ids [B,T] and cand_ids [B] index `table` [V,K]; mask = (hist id != 0)
 * (DL/model/din.py:76-81).  item_out [B,K] receives the candidate embedding. */
int ctrhot_din_attention_gather_f32(const float* table, int64_t V, int K,
                                    const void* hist_ids, int ids_kind, int64_t hist_ld,
                                    const void* cand_ids, int64_t cand_ld, int64_t B, int T,
                                    const ctrhot_dense_desc* h_layers, int n_layers,
                                    const float* out_w, const float* out_b,
                                    float* att_out, int64_t att_ld, float* item_out, int64_t item_ld,
                                    int* err_flag, void* stream);

/* tcgen05 version of the fused gather + attention unit ('prelu' / relu branch with exactly two hidden Dense layers,
 * K in {8,16,32,64}, T <= 128, h1 <= 96, h2 <= 64; CTRHOT_ESHAPE otherwise -> use the _f32 entry point).
 * TF32 multiply, fp32 accumulate; the first layer is evaluated as q (W1a + W1c) + k (W1b - W1c) + (q*k) W1d.
 * packed = ctrhot_din_pack(h_layers[2], out_w, out_b): weights in the kernel's shared-memory layout (re-pack when they
 * change).  workspace: ctrhot_din_tc_workspace_bytes(B, h1) bytes, 16-byte aligned.  item_out is REQUIRED (it carries the
 * candidate rows between the two launches) and item_ld % 4 == 0. */
int64_t ctrhot_din_pack_bytes(int T, int K, int h1, int h2);
int ctrhot_din_pack(const ctrhot_dense_desc* h_layers, int n_layers, const float* out_w, const float* out_b,
                    int T, int K, void* packed, void* stream);
int64_t ctrhot_din_tc_workspace_bytes(int64_t B, int h1);
int ctrhot_din_attention_gather_tf32(const float* table, int64_t V, int K,
                                     const void* hist_ids, int ids_kind, int64_t hist_ld,
                                     const void* cand_ids, int64_t cand_ld, int64_t B, int T,
                                     const void* packed, int h1, int h2,
                                     float* att_out, int64_t att_ld, float* item_out, int64_t item_ld,
                                     void* workspace, int64_t workspace_bytes, int* err_flag, void* stream);

/* ---- R16 DIN.call's front end in ONE launch (DL/model/din.py:56-90) ------------------------------------------
 * candidate lookup (din.py:78-79), history lookup (:76-77), the attention unit (:81, interaction.py:369-406), the other
 * sparse lookups (:65-69) and the dense columns (:60-61), written as the tower's input matrix (the concat of din.py:84-90):
 *     emb[b] = [ att (K) | item (K) | extras at their columns | dense at dense_col ]      emb_ld % 4 == 0, 16-byte aligned
 * The mask is hist id != 0 (din.py:80).  K in {32, 64}, T <= 128, exactly two hidden Dense (PReLU[T,h] or ReLU), h1 <= 96,
 * h2 <= 64; CTRHOT_ESHAPE otherwise (use ctrhot_din_attention_gather_* + ctrhot_gather_f32 then).  TF32 multiply, fp32
 * accumulate; the first layer is evaluated as q (W1a + W1c) + k ((W1b - W1c) + diag(q) W1d): a per-sample [K, h1] weight
 * matrix built in shared memory, so that layer 1 reads the gathered rows in place.
 * packed = ctrhot_din_fwd_pack(...) (re-pack when the weights change); workspace = ctrhot_din_fwd_workspace_bytes(B, h1).
 * h_extras: n_extras (<= 8) x 6 int64 on the HOST: {table (device pointer) [rows, dim], rows, dim, ids (device pointer, one id
 * per sample, same ids_kind as the history), ids_ld, first column in emb}.  Bad ids: zero row + *err_flag = 1. */
int64_t ctrhot_din_fwd_pack_bytes(int T, int K, int h1, int h2);
int ctrhot_din_fwd_pack(const ctrhot_dense_desc* h_layers, int n_layers, const float* out_w, const float* out_b,
                        int T, int K, void* packed, void* stream);
int64_t ctrhot_din_fwd_workspace_bytes(int64_t B, int h1);
int ctrhot_din_fwd_tf32(const float* table, int64_t V, int K, const void* hist_ids, int ids_kind, int64_t hist_ld,
                        const void* cand_ids, int64_t cand_ld, int64_t B, int T, const void* packed, int h1, int h2,
                        float* emb, int64_t emb_ld, const int64_t* h_extras, int n_extras,
                        const float* dense, int64_t dense_ld, int n_dense, int dense_col,
                        void* workspace, int64_t workspace_bytes, int* err_flag, void* stream);

/* ---- 8(f) rank 4: AttentionSequencePoolingLayer.call, DL/layer/sequence.py:242-276, after its LocalActivationUnit ------------
 * s_t = valid(t) ? scores[b, t] : (weight_normalization ? -2^32 + 1 : 0);  p = weight_normalization ? softmax over T : s;
 * return_score ? out[b, 0:T] = p : out[b, 0:E] = sum_t p_t keys[b, t, :].   valid(t) = mask[b, t] != 0 (uint8, the keras mask
 * of supports_masking=True) or, with mask NULL, t < lengths[b] (tf.sequence_mask(keys_length, T)).  keys [B, T, E], the
 * sample stride key_ld_b >= T * E. */
int ctrhot_seq_attn_pool_f32(const float* scores, int64_t score_ld, const float* keys, int64_t key_ld_b,
                             const uint8_t* mask, int64_t mask_ld, const int32_t* lengths, int64_t B, int T, int E,
                             int weight_normalization, int return_score, float* out, int64_t out_ld, void* stream);

/* ---- 8(f) rank 4: AUGRU (DIEN), AUGRUCell.call DL/layer/activation.py:113-142 unrolled over time as rnn_augru does without a
 * mask (:237-244; AUGRU.call, DL/layer/sequence.py:293-395) --------------------------------------------------------------------
 *   z = ra(x_t Wz + h Uz) * att[b, t];  r = ra(x_t Wr + h Ur);  hh = act(x_t Wh + (r * h) Uh);  h <- z h + (1 - z) hh   (no biases)
 * x [B, T, E] (sample stride x_ld_b), att [B, T]; kernel [E, 3H], recurrent_kernel [H, 3H] hold the z | r | h blocks; h0 [B, H] or
 * NULL (zeros, keras' initial state); seq_out [B, T, H] and / or last_out [B, H].  recurrent_hard_sigmoid: 0 = sigmoid (what the
 * reference's cell runs with under TF 2: keras' GRUCell.__init__ overwrites the 'hard_sigmoid' AUGRUCell sets, activation.py:97-100),
 * 1 = clip(0.2 v + 0.5, 0, 1); linear_activation: 0 = tanh.  fp32 CUDA cores; 3H <= 1024 and the weights must fit shared memory. */
int ctrhot_augru_f32(const float* x, int64_t x_ld_b, const float* att, int64_t att_ld, const float* kernel,
                     const float* recurrent_kernel, const float* h0, int64_t B, int T, int E, int H,
                     int recurrent_hard_sigmoid, int linear_activation, float* seq_out, float* last_out,
                     void* stream);

/* ---- one SGD step of DeepFM, the step on the right of the path ------------------------------
 * What compile_fit (DL/utils/compile_fit.py:9-15: SGD(lr), loss 'binary_crossentropy', model.fit) does to DeepFM
 * (DL/model/deepFM.py:15-31) for one batch: forward, mean clipped BCE + the l2 regularisers of FMLayer.build
 * (DL/layer/interaction.py:97-104: reg_w sum w1^2 + reg_v sum v^2), exact fp32 gradients, w <- w - lr * grad.
 * UPDATED IN PLACE: the embedding arena (only the rows looked up; a row used by several samples gets the sum of their
 * gradients, as keras' sparse SGD does), fm_w0 [1], fm_w1 [n,1], fm_v [n,k_fm] (n = D + F k), and weight / bias of every
 * h_layers entry: n_layers - 1 hidden Dense with ReLU + the final Dense(h -> 1) (act NONE).  dropout in [0,1) is applied after
 * every hidden layer (DNNLayer.call, interaction.py:40-46) with a stateless generator keyed by `seed` (pass a new seed per
 * step): same distribution as keras' Dropout, not TensorFlow's random stream.  loss_out (DEVICE, 2 floats) = mean BCE of the
 * batch BEFORE the update (without the regularisers, which is what keras prints minus its regularisation term) and the
 * number of samples on the right side of 0.5.  ids [B, F] local row numbers; dense [B, D]; y [B] in {0, 1}.
 * fp32 CUDA cores throughout (the reference trains at batch 32).  Bad ids: the row contributes zeros, *err_flag = 1.
 * Summation order (all ctrhot_*_train_step_f32): embedding rows take their samples' gradients through atomic adds; from 512
 * rows per batch the weight gradients, and from 2 048 the bias gradients, are summed over up to 32 row slices the same way --
 * results repeat to fp32 round-off (~1e-7 of a weight), not bit for bit. */
int64_t ctrhot_deepfm_train_workspace_bytes(int64_t B, int n, int k_fm, const ctrhot_dense_desc* h_layers,
                                            int n_layers);
int ctrhot_deepfm_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                 int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y,
                                 int64_t B, float* fm_w0, float* fm_w1, float* fm_v, int k_fm, float reg_w,
                                 float reg_v, const ctrhot_dense_desc* h_layers, int n_layers, float lr,
                                 float dropout, uint64_t seed, float* loss_out, void* workspace,
                                 int64_t workspace_bytes, int* err_flag, void* stream);
/* The same step driven by a 32-byte DEVICE control block {uint64 step; int64 row0; float acc_loss; float acc_correct; int64 pad}
 * (caller-owned, 8-byte aligned, zero it before an epoch): the batch is rows [row0, row0 + B) of ids / dense / y, `step` is added
 * to the dropout seed, the batch's summed BCE and correct count are added to acc_*, and the LAST kernel of the step does
 * step += 1, row0 += B.  Nothing about a batch is in the launch parameters any more, so one captured CUDA graph of this call
 * (no allocation, no synchronisation inside) is replayed once per batch: at the reference's batch size of 32 the step is ~25
 * small launches and the graph removes their host cost. */
int ctrhot_deepfm_train_step_ctl_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids,
                                     int ids_kind, int64_t ids_ld, const float* dense, int64_t dense_ld, int D,
                                     const float* y, int64_t B, float* fm_w0, float* fm_w1, float* fm_v, int k_fm,
                                     float reg_w, float reg_v, const ctrhot_dense_desc* h_layers, int n_layers,
                                     float lr, float dropout, uint64_t seed, float* loss_out, void* workspace,
                                     int64_t workspace_bytes, int* err_flag, void* d_ctl, void* stream);

/* FM under compile_fit (DL/model/fm.py:14-23: p = sigmoid(FMLayer(x)) on the dense one-hot matrix x [B, n], x_ld == n): one SGD
 * step on mean clipped BCE + reg_w sum w1^2 + reg_v sum v^2, weights updated in place; loss_out as above. */
int64_t ctrhot_fm_train_workspace_bytes(int64_t B, int n, int k_fm);
int ctrhot_fm_train_step_f32(const float* x, int64_t x_ld, int n, const float* y, int64_t B, float* fm_w0,
                             float* fm_w1, float* fm_v, int k_fm, float reg_w, float reg_v, float lr,
                             float* loss_out, void* workspace, int64_t workspace_bytes, void* stream);

/* ---- R12/R14 fused model front ends ------------------------------------------
 * gather -> x=[dense, embed] -> FM (DeepFM, DL/model/deepFM.py:24-28) or L cross
 * layers (DCN, DL/model/dcn.py:25-29) in ONE kernel; the gathered rows are
 * consumed from registers, x is written once for the tower (or not at all when
 * x == NULL).  x uses the INTERNAL column order [embeddings F*k | dense D | zero
 * pad to x_ld] so all stores are 128-bit; weights are permuted to that order by
 * the pack helpers below (n = D + F*k reference columns, n_pad = n rounded up
 * to 32):
 *   ctrhot_fm_pack_planes : planes[kfm+2][n_pad] = v[:,c] | w1 | s_j=sum_c v[j,c]^2
 *   ctrhot_permute_cols   : dst[rows][n_pad] from src[rows][n] (cross w/b, final w)
 * fm_out / cross_dot are [B]; cross_out (optional) is [B, cross_ld] internal order. */
int64_t ctrhot_fm_planes_bytes(int n, int kfm);
int ctrhot_fm_pack_planes(const float* w1, const float* v, int n, int kfm, int D,
                          float* planes, void* stream);
int ctrhot_permute_cols(const float* src, int rows, int n, int D, float* dst, void* stream);
int ctrhot_deepfm_front_f32(const float* arena, const int64_t* row_off, int F, int k,
                            const void* ids, int ids_kind, int64_t ids_ld,
                            const float* dense, int64_t dense_ld, int D, int64_t B,
                            const float* w0, const float* planes, int kfm,
                            float* x, int64_t x_ld, float* fm_out, int* err_flag, void* stream);
int ctrhot_dcn_front_f32(const float* arena, const int64_t* row_off, int F, int k,
                         const void* ids, int ids_kind, int64_t ids_ld,
                         const float* dense, int64_t dense_ld, int D, int64_t B,
                         int L, const float* cw, const float* cb, const float* final_w,
                         float* x, int64_t x_ld, float* cross_out, int64_t cross_ld,
                         float* cross_dot, int* err_flag, void* stream);
/* out[b] = sigmoid(scale * (a[b] + (b_ ? b_[b] * (b_scale ? *b_scale : 1) : 0) + (bias ? *bias : 0)))
 * -- DeepFM/WideDeep: scale 0.5 (deepFM.py:30, wideDeep.py:33); DCN: a = cross dot, b_ = tower
 * output, b_scale = last weight of the final Dense, bias = its bias (dcn.py:32-33). */
int ctrhot_combine_sigmoid_f32(const float* a, const float* b_, const float* b_scale, const float* bias,
                               float scale, int64_t B, float* out, void* stream);

/* ---- 8(e) config 5: row-sharded tables, owner = id mod G, local row = id div G -
 * route: bucket this rank's [B,F] ids by owner; send_ids/send_slot are laid out
 * owner-major using counts/offsets computed by the same call (device int32[G]). */
int ctrhot_shard_route_ids(const int32_t* ids, int64_t B, int F, int G,
                           int32_t* send_local_row, int32_t* send_field, int32_t* send_slot,
                           int32_t* counts, void* workspace, int64_t workspace_bytes, void* stream);
int64_t ctrhot_shard_route_workspace_bytes(int64_t B, int F, int G);
/* owner side: rows[n, k] = arena[row_off[field[n]] + local_row[n]] */
int ctrhot_shard_gather_pack(const float* arena, const int64_t* row_off, int F, int k,
                             const int32_t* local_row, const int32_t* field, int64_t n,
                             float* rows, int* err_flag, void* stream);
/* requester side: x[slot/F, D + (slot%F)*k + c] = rows[n, c] */
int ctrhot_shard_unpack(const float* rows, const int32_t* slot, int64_t n, int F, int k, int D,
                        float* x, int64_t x_ld, void* stream);

/* synthetic config-5 tables: arena[row_off[f] + lr, c] = hash(seed, f, lr * G + rank, c) in U(-0.05, 0.05); every rank
 * can regenerate any global row for checking (SURVEY 8(d) cfg 5).  total_rows = row_off[F] (host copy of it). */
int ctrhot_shard_init_hash_f32(float* arena, const int64_t* row_off, int64_t total_rows, int F, int k,
                               int G, int rank, uint64_t seed, void* stream);

/* ---- 8(e) config 5 over NVSwitch peer memory: the requester gathers straight out of the owners' HBM (P2P loads over
 * NVLink), fusing route + all-to-all(ids) + gather_pack + all-to-all(rows) + unpack into ONE kernel (csrc/peer.cu).
 * ctrhot_peer_alloc cudaMalloc's an arena whose base a CUDA IPC handle names; every rank exports its handle
 * (ctrhot_peer_handle_bytes() opaque bytes, exchanged once by the host over torch.distributed at load time -- not on the
 * data path) and opens its peers'.  arenas = HOST array of G device pointers (own arena at index rank), row_off = device
 * int64 [G, F+1] (rank g's per-field first rows), ids [B, F] int32 global row numbers.  Writes
 * x[b, D + f*k + c] = arena_{id mod G}[row_off[id mod G][f] + id div G, c]; with dense != NULL also x[b, 0:D] = dense[b, :]
 * and zeroes columns [D + F*k, x_ld).  Ids outside their table raise *err_flag (row reads as zeros). */
int ctrhot_peer_handle_bytes(void);
int ctrhot_peer_alloc(int64_t bytes, void** ptr);
int ctrhot_peer_free(void* ptr);
int ctrhot_peer_export(void* ptr, void* handle_out);
int ctrhot_peer_open(const void* handle, void** ptr);
int ctrhot_peer_close(void* ptr);
int ctrhot_peer_enable(int peer_device);   /* single process, several GPUs: cudaDeviceEnablePeerAccess for the current device */
int ctrhot_shard_lookup_peer_f32(const void* const* arenas, const int64_t* row_off, int F, int k, int G,
                                 const int32_t* ids, int64_t B, int D, const float* dense, int64_t dense_ld,
                                 float* x, int64_t x_ld, int* err_flag, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CTRHOT_H_ */
