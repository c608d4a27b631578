/*
 * hop_sm100.h -- C-ABI of the B200-native (sm_100a) Hopfield retrieval library.
 *
 * This is the drop-in boundary for ONE path of joejenb/Hop-VAE: the
 * `hflayers.HopfieldLayer` association that HopVAE.py places between encoder
 * and decoder.  The reference has no FFI of its own for this path (it is a
 * Python nn.Module API, /root/reference/HopVAE.py:7 `from hflayers import
 * HopfieldLayer`); every entry point below therefore cites the reference /
 * upstream interface it stands in for, and `INTEGRATION.md` shows the ctypes
 * stub that binds it behind the unchanged module API.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types;
 *   - every pointer is a DEVICE pointer unless the name ends in `_host`;
 *   - all tensors are fp32, row-major, contiguous, 16-byte aligned;
 *   - the library never allocates or frees device memory and never
 *     synchronises the host: the caller passes outputs and a workspace whose
 *     size it queried first, and a `cudaStream_t` (as `void*`) on which all
 *     work is enqueued;
 *   - every function returns a `hop_status`; `hop_sm100_last_error()` gives
 *     the thread-local message of the last non-zero status;
 *   - there is NO CPU fallback and no non-sm_100 code path.
 *
 * Shapes:  N query rows ("latents"), P stored patterns, Din raw width
 * (`input_size`), H association width, Dv pattern-projection width, Dout
 * output width, T extra retrieve updates (`update_steps_max`).
 */
#ifndef HOP_SM100_H_
#define HOP_SM100_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HOP_SM100_ABI_VERSION 3u

typedef enum hop_status {
  HOP_OK = 0,
  HOP_ERR_BAD_ARG = 1,      /* null pointer, bad abi_version, bad enum            */
  HOP_ERR_BAD_SHAPE = 2,    /* a dimension is <= 0 or outside what is built       */
  HOP_ERR_BAD_ALIGN = 3,    /* a pointer is not 16-byte aligned                   */
  HOP_ERR_BAD_ARCH = 4,     /* current device is not compute capability 10.x      */
  HOP_ERR_CUDA = 5,         /* a CUDA call failed (message has the CUDA error)    */
  HOP_ERR_WORKSPACE = 6,    /* workspace too small                                */
  HOP_ERR_UNSUPPORTED = 7   /* valid request that this build does not implement   */
} hop_status;

/* hop_desc.flags -- which optional stages of upstream `Hopfield._associate` /
 * `HopfieldCore` are present.  Hop-VAE's three layers (HopVAE.py:149-170,
 * stored/state patterns static) use NORM_* | PROJ_PATTERN | OUT_PROJ.         */
enum {
  HOP_F_NORM_STATE   = 1u << 0,  /* LayerNorm on queries   (normalize_state_pattern)      */
  HOP_F_NORM_STORED  = 1u << 1,  /* LayerNorm on patterns  (normalize_stored_pattern)     */
  HOP_F_NORM_PROJ    = 1u << 2,  /* LayerNorm on patterns  (normalize_pattern_projection) */
  HOP_F_PROJ_STATE   = 1u << 3,  /* Q projection present   (state_pattern_as_static=False)  */
  HOP_F_PROJ_STORED  = 1u << 4,  /* K projection present   (stored_pattern_as_static=False) */
  HOP_F_PROJ_PATTERN = 1u << 5,  /* V projection present   (pattern_projection_as_static=False) */
  HOP_F_OUT_PROJ     = 1u << 6,  /* out_proj present       (disable_out_projection=False)  */
  /* behaviour switches of this library (no upstream counterpart) */
  HOP_F_EXACT_ARGMAX = 1u << 7,  /* tensor-core paths: rows whose nearest pattern the 16-bit score arithmetic cannot
                                    decide (on-chip top-1 / top-2 bound, see hop_sm100_forward) are re-scored in fp32 */
  HOP_F_TRAIN        = 1u << 8   /* prepare_patterns also writes the packs only the tensor-core backward reads
                                    (kt16, v16) */
};

/* hop_desc.tune_flags -- scheduling overrides for tests and experiments; 0 = the library's own heuristics.
 * They travel in the descriptor (not in environment variables): the library keeps no process-wide state. */
enum {
  HOP_TUNE_STREAMK_ON   = 1u << 0,  /* force / forbid the stream-K schedule of the T = 0 forward             */
  HOP_TUNE_STREAMK_OFF  = 1u << 1,
  HOP_TUNE_ROWS128      = 1u << 2,  /* D >= 512 on the 128-row kernel (score tile recomputed per column chunk) */
  HOP_TUNE_MULTI_ON     = 1u << 3,  /* retrieve updates as one launch per pass / on chip in one launch         */
  HOP_TUNE_MULTI_OFF    = 1u << 4,
  HOP_TUNE_PAIR_ON      = 1u << 5,  /* tcgen05 CTA pairs (cta_group::2) forced on / off                        */
  HOP_TUNE_PAIR_OFF     = 1u << 6,
  HOP_TUNE_PLAIN_UPDATES = 1u << 7, /* retrieve updates with plain 16-bit operands even when T >= 2 (default: the
                                       update passes of T >= 2 use hi + lo operand pairs)   */
  HOP_TUNE_BWD_SIMT     = 1u << 8,  /* backward on the fp32 CUDA-core kernels even where the tcgen05 one applies */
  HOP_TUNE_PREP_SIMT    = 1u << 9,  /* pattern preparation on the fp32 CUDA-core kernels even where the tcgen05
                                       GEMM applies (hop_patterns.scratch given, widths multiples of 64)      */
  HOP_TUNE_BWD_TAILS_SIMT = 1u << 10 /* projection / weight-gradient products of a tensor-core backward on the fp32
                                       CUDA-core kernels even where the tcgen05 GEMM applies (widths % 128)   */
};

/* hop_desc.path -- which kernel family evaluates the association.            */
enum {
  HOP_PATH_AUTO      = 0,  /* library picks (see hop_sm100_select_path)                         */
  HOP_PATH_FP32_SIMT = 1,  /* exact fp32 CUDA-core kernels (small / argmax-tight problems, D<16) */
  HOP_PATH_TC_F16    = 2,  /* tcgen05 kind::f16 MMAs, fp16-encoded operands, fp32 accumulate;
                              H == Dv in {64, 128, 256, 512, 1024}, else HOP_ERR_UNSUPPORTED    */
  HOP_PATH_TC_BF16   = 3   /* same kernels, bf16-encoded operands (8 significand bits: fails the 1e-3
                              parity bar on tight problems; kept for comparison, never picked by AUTO) */
};

typedef struct hop_desc {
  uint32_t abi_version;     /* HOP_SM100_ABI_VERSION                                   */
  uint32_t flags;           /* HOP_F_*                                                 */
  int32_t  path;            /* HOP_PATH_*                                              */
  int32_t  n_rows;          /* N                                                       */
  int32_t  n_patterns;      /* P  (`quantity`; this shard's count when sharded)        */
  int32_t  d_in;            /* Din (`input_size`)                                      */
  int32_t  d_assoc;         /* H  (`hidden_size`, = Din when Q/K are static)           */
  int32_t  d_v;             /* Dv (`pattern_size`, = H by default)                     */
  int32_t  d_out;           /* Dout (`output_size`), ignored without HOP_F_OUT_PROJ    */
  int32_t  update_steps;    /* T  (`update_steps_max`): at most T retrieve updates; < 0 = until the stop rule
                               fires (fp32 path, at most 64 updates)                     */
  int32_t  pattern_offset;  /* global index of this shard's first pattern (argmax)     */
  float    beta;            /* `scaling` (1/sqrt(H) when upstream's is None)           */
  float    ln_eps;          /* LayerNorm eps (1e-5)                                    */
  float    update_eps;      /* `update_steps_eps`: upstream's stop rule -- after pass t >= 1 the updates end
                               when max over the batch of ||A_t - A_(t-1)||_F <= update_eps; 0 = exactly T
                               updates.  Honoured by the fp32 path; the 16-bit tensor-core paths cannot resolve
                               1e-4 changes of A and always run exactly T updates                              */
  int32_t  rows_per_item;   /* query rows per batch item (`L` of a (B, L, D) input) for the stop rule's
                               per-item Frobenius norm; 0 = all N rows are one item                           */
  int32_t  tune_splits;     /* pattern splits per row tile of the tensor-core forward, 0 = heuristic          */
  uint32_t tune_flags;      /* HOP_TUNE_*, 0 = heuristics                                                     */
} hop_desc;

/* Module parameters, exactly the tensors of the reference's state-dict
 * (checkpoints/MNIST-28.ckpt keys).  NULL where a stage is
 * absent; LayerNorm weight/bias NULL means "no affine".                      */
typedef struct hop_params {
  const float* lookup;         /* lookup_weights[0]                         (P, Din)  */
  const float* lookup_targets; /* lookup_targets[0] or NULL (= lookup)      (P, Din)  */
  const float* ln_state_w;     /* hopfield.norm_state_pattern.weight        (Din)     */
  const float* ln_state_b;     /* hopfield.norm_state_pattern.bias          (Din)     */
  const float* ln_stored_w;    /* hopfield.norm_stored_pattern.weight       (Din)     */
  const float* ln_stored_b;    /* hopfield.norm_stored_pattern.bias         (Din)     */
  const float* ln_proj_w;      /* hopfield.norm_pattern_projection.weight   (Din)     */
  const float* ln_proj_b;      /* hopfield.norm_pattern_projection.bias     (Din)     */
  const float* wq; const float* bq;  /* in_proj_weight/bias q-slice  (H, Din), (H)    */
  const float* wk; const float* bk;  /* in_proj_weight/bias k-slice  (H, Din), (H)    */
  const float* wv; const float* bv;  /* in_proj_weight/bias v-slice  (Dv, Din), (Dv)  */
  const float* wo; const float* bo;  /* out_proj.weight/bias         (Dout, Dv), (Dout) */
} hop_params;

/* Prepared pattern-side operands: written by hop_sm100_prepare_patterns, read
 * by forward/backward.  Sizes from hop_sm100_pattern_bytes.  The 16-bit packs
 * are only needed (and only written) for the tensor-core paths.              */
typedef struct hop_patterns {
  float* khat;   /* K^ = [LN_k](W)[Wk^T + bk]            (P, H)   fp32                 */
  float* vhat;   /* V^ = [LN_v](Wt)[Wv^T + bv]           (P, Dv)  fp32                 */
  void*  k16;    /* K^ 16-bit, K-major for the score MMA (Ppad, H)                     */
  void*  kt16;   /* K^ transposed, for retrieve updates  (H, Ppad)  (T > 0 only)       */
  void*  vt16;   /* V^ transposed, for the read-out MMA  (Dv, Ppad)                    */
  void*  vot16;  /* (V^ . Wo^T) transposed (Dout, Ppad): upstream `out_proj` folded into
                    the read-out operand, A.(V^ Wo^T) + bo == (A.V^) Wo^T + bo; written and
                    used on the tensor-core paths when HOP_F_OUT_PROJ and Dout == H == Dv,
                    so that inference needs no separate out-projection launch            */
  void*  k16lo;  /* K^ - fp16(K^) as fp16 (Ppad, H): low halves of the score operand, used by the
                    update passes of T >= 2 (hi + lo pairs, 3 MMAs per score tile)          */
  void*  kt16lo; /* same, transposed (H, Ppad): low halves of the retrieve-update operand    */
  void*  v16;    /* V^ 16-bit, row-major (Ppad, Dv): tensor-core backward (HOP_F_TRAIN)      */
  float* kstats; /* 4 floats: [0] = max_j ||K^_j||_2 (bounds the 16-bit score noise for the
                    HOP_F_EXACT_ARGMAX certificate); tensor-core paths                       */
  void*  scratch;/* temporary of hop_sm100_prepare_patterns on the tensor-core paths (operand packs of the
                    V-projection / out_proj fold, which then run on tcgen05); only that call touches it --
                    free it afterwards.  NULL: the projection runs on the fp32 CUDA-core kernel instead */
} hop_patterns;

/* Byte sizes of the hop_patterns buffers (0 where a buffer is not needed).   */
typedef struct hop_pattern_sizes {
  size_t khat, vhat, k16, kt16, vt16, vot16, k16lo, kt16lo, v16, kstats, scratch;
} hop_pattern_sizes;

/* Gradients written (overwritten, not accumulated) by hop_sm100_backward.
 * Same shapes as the corresponding hop_params members; NULL = not wanted.    */
typedef struct hop_grads {
  float* d_x;            /* (N, Din)  gradient w.r.t. the queries */
  float* d_lookup;       /* (P, Din)  via K^ and via V^ (summed; lookup_targets == NULL) */
  float* d_lookup_targets;
  float* d_ln_state_w;  float* d_ln_state_b;
  float* d_ln_stored_w; float* d_ln_stored_b;
  float* d_ln_proj_w;   float* d_ln_proj_b;
  float* d_wq; float* d_bq;
  float* d_wk; float* d_bk;
  float* d_wv; float* d_bv;
  float* d_wo; float* d_bo;
} hop_grads;

/* ABI / build identification.  Returns HOP_SM100_ABI_VERSION.                */
uint32_t hop_sm100_version(void);

/* Thread-local message of the last failing call on this thread ("" if none). */
const char* hop_sm100_last_error(void);

/* Path the library would run for `desc` (resolves HOP_PATH_AUTO): the exact
 * fp32 kernels for small problems or H < 16, the tcgen05 kernels otherwise.  */
int hop_sm100_select_path(const hop_desc* desc, int32_t* path_out);

/* Path a TRAINING step (hop_sm100_forward + hop_sm100_backward) of `desc` runs on: the tcgen05 path where the
 * tensor-core backward applies (T == 0, H == Dv in {128, 256, 512}, large N x P, LayerNorm'd operands), else the
 * exact fp32 kernels (H, Dv <= 256); HOP_ERR_UNSUPPORTED where neither does.  The caller passes the result as
 * desc.path (with HOP_F_TRAIN on a tensor-core path) to prepare_patterns, forward and backward of that step.   */
int hop_sm100_select_train_path(const hop_desc* desc, int32_t* path_out);

/* Byte sizes of the hop_patterns buffers for `desc` (0 where not needed).     */
int hop_sm100_pattern_bytes(const hop_desc* desc, hop_pattern_sizes* sizes);

/* Workspace bytes needed by forward (for_backward == 0) or backward (!= 0).  */
int hop_sm100_workspace_bytes(const hop_desc* desc, int for_backward, size_t* bytes);

/* Pattern-side preparation, once per parameter update (replaces the B-times
 * redundant LayerNorm / projection of the batch-expanded lookup matrix that
 * upstream `HopfieldLayer._prepare_input` + `Hopfield._associate` +
 * `HopfieldCore` in-projection perform; called from HopVAE.py:193,215,216,
 * 234,255,257,271 through `HopfieldLayer.forward`).                          */
int hop_sm100_prepare_patterns(const hop_desc* desc, const hop_params* params,
                               const hop_patterns* patterns, void* stream);

/* The association itself: upstream `hopfield_core_forward` as reached from
 * `HopfieldLayer.forward` (call sites HopVAE.py:193,215,216,234,255,257,271):
 * LN_q, optional Q projection, beta scaling, score contraction, row softmax,
 * T retrieve updates, read-out, out-projection.
 *   x          (N, Din)   queries
 *   out        (N, Dout)  layer output ((N, Dv) without HOP_F_OUT_PROJ)
 *   retrieved  (N, Dv)    A.V^ before out_proj            (optional, NULL ok)
 *   qhat       (N, H)     beta * Q^ of the LAST pass      (optional; needed for backward)
 *   lse        (N)        log-sum-exp of the last pass' scores (natural log)
 *   argmax     (N)        index of the nearest (highest-score) pattern, + pattern_offset
 * When sharded, `out` of each shard is the shard-local normalised result and
 * (`retrieved`, `lse`, `argmax`, `rowmax`) feed hop_sm100_merge_shards.
 *   rowmax     (N)        max score of the last pass (optional, NULL ok)
 * On the tensor-core paths, when `retrieved` is NULL and patterns->vot16 is
 * given, the folded read-out operand is used and `out` comes straight from the
 * fused kernel (no out-projection launch).                                   */
int hop_sm100_forward(const hop_desc* desc, const hop_params* params,
                      const hop_patterns* patterns, const float* x,
                      float* out, float* retrieved, float* qhat, float* lse,
                      int32_t* argmax, float* rowmax,
                      void* workspace, size_t workspace_bytes, void* stream);

/* Backward of hop_sm100_forward for T == 0 (what every reference config
 * trains with: update_steps_max defaults to 0, main.py:35 `loss.backward()`).
 * Flash-style: the N x P matrices are recomputed from (qhat, lse), never
 * stored.  `retrieved` is the forward's A.V^ (N, Dv).  desc.path = the path
 * the forward of this step ran on (hop_sm100_select_train_path): on a
 * tensor-core path the association core (dq, dK^, dV^: the five dense
 * contractions) runs on tcgen05 from the kt16 / v16 packs (HOP_F_TRAIN);
 * LayerNorm / projection backwards are fp32 CUDA-core kernels either way.    */
int hop_sm100_backward(const hop_desc* desc, const hop_params* params,
                       const hop_patterns* patterns, const float* x,
                       const float* qhat, const float* retrieved, const float* lse,
                       const int32_t* argmax,   /* the forward's nearest-pattern indices (pattern_offset 0);
                                                   needed on a tensor-core path, NULL ok on the fp32 path */
                       const float* grad_out, const hop_grads* grads,
                       void* workspace, size_t workspace_bytes, void* stream);

/* Hop-VAE's bottleneck as one launch (replaces the caller's
 * glue around three HopfieldLayer.forward calls, HopVAE.py:255-272 and
 * utils/__init__.py:51-55):
 *   emb       = hopfield(z)                               (N, D)
 *   logits    = embedding_to_index(emb)                   (N, I)
 *   idx_q     = round(sigmoid(logits) * (num_levels - 1)) (N, I)  quantised levels
 *   emb_recon = index_to_embedding(idx_q / (num_levels - 1))      (N, D)
 *   mse       = mean((emb_recon - emb)^2)                 (1)     device scalar
 * descs / params / patterns: arrays of 3 (hopfield, embedding_to_index,
 * index_to_embedding), each prepared with hop_sm100_prepare_patterns on the
 * fp32 path.  Forward only (evaluation; training keeps the three layer calls
 * and their autograd).  Every output except emb and mse may be NULL.
 * HOP_ERR_UNSUPPORTED unless all three layers have the Hop-VAE shape (static
 * state / stored patterns, widths <= 64, I <= 4, update_steps == 0).          */
int hop_sm100_bottleneck_workspace_bytes(int32_t n_rows, size_t* bytes);
int hop_sm100_bottleneck_forward(const hop_desc* descs, const hop_params* params,
                                 const hop_patterns* patterns, int32_t num_levels,
                                 const float* z, float* emb, float* logits, float* idx_q,
                                 float* emb_recon, float* mse,
                                 void* workspace, size_t workspace_bytes, void* stream);

/* hop_sm100_bottleneck_forward without ANY allocation and without a caller
 * workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  The kernel's
 * scratch (completion counter + per-tile partial sums, 65,792 bytes) lives in a
 * file-scope __device__ array of the library, held by every GPU that loads it
 * until exit.  Limits: n_rows <= hop_sm100_bottleneck_static_max_rows()
 * (262,144; HOP_ERR_BAD_SHAPE beyond -- use the workspace form); all calls on
 * one GPU share the array, so calls that OVERLAP on that GPU (two streams)
 * race -- serialise them.  Same arguments and results as
 * hop_sm100_bottleneck_forward otherwise; calls no cudaMalloc / cudaFree.      */
int hop_sm100_bottleneck_forward_static(const hop_desc* descs, const hop_params* params,
                                        const hop_patterns* patterns, int32_t num_levels,
                                        const float* z, float* emb, float* logits, float* idx_q,
                                        float* emb_recon, float* mse, void* stream);
int hop_sm100_bottleneck_static_max_rows(void);

/* Diagnostic of HOP_F_EXACT_ARGMAX: of the last forward run with this descriptor
 * (argmax != NULL) into this workspace, how many rows took the full fp32
 * re-scoring pass (rows_out) and how many had failed the on-chip top-1 / top-2
 * certificate before the 64-pattern block check settled most of them
 * (flagged_out, NULL ok).  Waits for `stream`.  0 on the fp32 path (whose index
 * is exact by construction).                                                   */
int hop_sm100_repair_rows(const hop_desc* desc, const void* workspace, void* stream, int32_t* rows_out,
                          int32_t* flagged_out);

/* Pattern-sharded mode (new; the reference is single-device): combine G
 * shard-local results with the log-sum-exp monoid.
 *   lse_all     (G, N)      every shard's lse   (after an all-gather)
 *   rowmax_all  (G, N)      every shard's row max, argmax_all (G, N) its index
 *   retrieved   (N, Dv)     in: this shard's normalised A_g.V^_g ; out: scaled by
 *                            w_g = exp(lse_g - lse) so that a SUM all-reduce over
 *                            shards yields the global read-out
 *   lse_out     (N)         global lse;  argmax_out (N) global nearest pattern
 *                            (ties -> lowest index)                              */
int hop_sm100_merge_shards(int32_t n_rows, int32_t d_v, int32_t n_shards, int32_t shard,
                           const float* lse_all, const float* rowmax_all,
                           const int32_t* argmax_all, float* retrieved,
                           float* lse_out, int32_t* argmax_out, void* stream);

/* Pattern-sharded mode, exchange fused over NVLink / NVSwitch peer memory (new;
 * replaces all_gather + merge + all_reduce by ONE kernel per rank):
 * every rank g keeps its shard-local (O_g, lse_g, rowmax_g, argmax_g) in a
 * peer-mapped buffer; rank r finishes rows [row_begin, row_begin+row_count):
 * it waits (flag barrier in peer memory) until every rank's partials of this
 * `epoch` are complete, pulls the G partial rows with P2P loads, merges them
 * with the log-sum-exp monoid and stores the result into its own `out` buffer
 * (push_all = 0, reduce-scatter semantics) or into every rank's (push_all = 1,
 * all-reduce semantics; P2P stores).  A closing flag barrier makes the kernel's
 * completion on rank r imply that every peer's rows have landed in r's buffers
 * and that nobody still reads r's partials.
 *   o[g] (N,Dv), lse[g] (N), rowmax[g] (N), argmax[g] (N): rank g's partials
 *   out[g] (N,Dv), lse_out[g] (N), argmax_out[g] (N): rank g's result buffers
 *   flags[g]: rank g's flag block, HOP_PEER_FLAG_WORDS uint32, zeroed once at
 *             set-up and owned by the library afterwards
 *   epoch: the same strictly increasing non-zero number on every rank per call
 * All pointers must be dereferenceable from the calling rank's device
 * (cudaIpcOpenMemHandle / cuMem mappings + hop_sm100_enable_peer_access).      */
#define HOP_MAX_PEERS 16
#define HOP_PEER_FLAG_WORDS 64
typedef struct hop_peer_merge {
  int32_t n_rows, d_v, n_ranks, rank;
  int32_t row_begin, row_count;
  int32_t push_all;
  uint32_t epoch;
  const float* o[HOP_MAX_PEERS];
  const float* lse[HOP_MAX_PEERS];
  const float* rowmax[HOP_MAX_PEERS];
  const int32_t* argmax[HOP_MAX_PEERS];
  float* out[HOP_MAX_PEERS];
  float* lse_out[HOP_MAX_PEERS];
  int32_t* argmax_out[HOP_MAX_PEERS];
  uint32_t* flags[HOP_MAX_PEERS];
} hop_peer_merge;
int hop_sm100_peer_merge(const hop_peer_merge* m, void* stream);

/* cudaDeviceEnablePeerAccess(peer_device) from the current device (already
 * enabled is not an error).                                                   */
int hop_sm100_enable_peer_access(int32_t peer_device);

/* Sharing the exchange block between the ranks of one node (one process per
 * GPU).  export: CUDA IPC handle of the cudaMalloc allocation that contains
 * `ptr` plus ptr's offset inside it.  import: maps that allocation into the
 * calling process on the CURRENT device (cudaIpcMemLazyEnablePeerAccess, so the
 * calling rank's kernels may load/store it over NVLink) and returns the
 * address that corresponds to the exporter's `ptr`; the mapping lives until
 * hop_sm100_ipc_close(ptr, offset) or process exit.  The library allocates
 * nothing: the block itself belongs to the caller (a torch tensor).           */
#define HOP_IPC_HANDLE_BYTES 64
int hop_sm100_ipc_export(const void* ptr, unsigned char handle[HOP_IPC_HANDLE_BYTES], uint64_t* offset);
int hop_sm100_ipc_import(const unsigned char handle[HOP_IPC_HANDLE_BYTES], uint64_t offset, void** ptr);
int hop_sm100_ipc_close(void* ptr, uint64_t offset);

/* out = retrieved . Wo^T + bo on (N, Dv) -> (N, Dout); used after a shard
 * merge (upstream `out_proj`, ckpt keys ...association_core.out_proj.*).     */
int hop_sm100_out_projection(int32_t n_rows, int32_t d_v, int32_t d_out,
                             const float* retrieved, const float* wo, const float* bo,
                             float* out, void* stream);

/* Number of kernel launches this library has enqueued since load (for the
 * bench's `gpu_launches` claim); thread-safe, monotonically increasing.      */
uint64_t hop_sm100_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* HOP_SM100_H_ */
