// Shared host/device helpers of the sm_100a Hopfield retrieval library.
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/hop_sm100.h"

namespace hop {

// ---------------------------------------------------------------- host side
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define HOP_FAIL(code, ...)            \
  do {                                 \
    ::hop::set_error(__VA_ARGS__);     \
    return (code);                     \
  } while (0)

#define HOP_CUDA(expr)                                                              \
  do {                                                                              \
    cudaError_t e__ = (expr);                                                       \
    if (e__ != cudaSuccess)                                                         \
      HOP_FAIL(HOP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
               __FILE__, __LINE__);                                                 \
  } while (0)

#define HOP_LAUNCH_CHECK(name)                                                     \
  do {                                                                             \
    cudaError_t e__ = cudaGetLastError();                                          \
    if (e__ != cudaSuccess)                                                        \
      HOP_FAIL(HOP_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(e__)); \
    ::hop::count_launch();                                                         \
  } while (0)

#define HOP_TRY(expr)          \
  do {                         \
    int s__ = (expr);          \
    if (s__ != HOP_OK) return s__; \
  } while (0)

static inline int round_up(int x, int m) { return (x + m - 1) / m * m; }
static inline size_t round_up_sz(size_t x, size_t m) { return (x + m - 1) / m * m; }
// scheduling overrides of the tensor-core forward (hop_desc.tune_splits / tune_flags)
struct TcTune { int splits; uint32_t flags; };
static inline TcTune tune_of(const hop_desc& d) { return TcTune{d.tune_splits, d.tune_flags}; }
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Pattern count padded for the 16-bit transposed packs (TMA row pitch must be a
// multiple of 16 bytes; we pad to a full 128-pattern tile so tiles never straddle).
static inline int padded_patterns(int p) { return round_up(p, 128); }

int validate_desc(const hop_desc* d);   // common checks, sets error
int resolve_path(const hop_desc* d);    // HOP_PATH_* actually run (never AUTO), <0 on error
int num_sms();
bool out_proj_foldable(const hop_desc& d);  // HOP_F_OUT_PROJ can be folded into the TC read-out operand

// Entry points implemented per translation unit ------------------------------
// hop_prep.cu
int launch_prepare_patterns(const hop_desc& d, const hop_params& p, const hop_patterns& pt, int path,
                            cudaStream_t st);
int launch_prepare_queries(const hop_desc& d, const hop_params& p, const float* x, float* qhat,
                           cudaStream_t st);  // qhat = beta * ([LN](x)[Wq^T+bq])  (N,H)
int launch_linear(int n, int din, int dout, const float* x, const float* w, const float* b, float* y,
                  cudaStream_t st);           // y = x w^T + b
// hop_fwd_simt.cu
// One association pass on prepared operands.  If `x_raw` != nullptr the kernel
// fuses LN_q (+beta) itself, else it reads prepared `qs` (N,H).
// `bmat` is the read-out matrix (P,Dr) (V^ for the last pass, K^ for an update
// pass).  Writes res (N,Dr) scaled by out_scale, optional lse/argmax/rowmax/qsave,
// and the fused out-projection when wo != nullptr.
struct SimtPassArgs {
  const float* x_raw; const float* lnq_w; const float* lnq_b; float ln_eps; int has_ln; float beta;
  const float* qs;
  const float* khat; const float* bmat;
  int n, p, h, dr;
  float out_scale;
  float* res;       // (N, Dr) or nullptr
  float* qsave;     // (N, H)  or nullptr
  float* lse; int32_t* argmax; float* rowmax; int pattern_offset;
  const float* wo; const float* bo; int dout; float* out;   // fused out-proj (optional)
  const uint32_t* stopped;   // device flag of the stop rule or nullptr: when set, an update pass copies qs -> res
};
int launch_simt_pass(const SimtPassArgs& a, cudaStream_t st);
// hop_bwd_simt.cu
int launch_simt_backward(const hop_desc& d, const hop_params& p, const hop_patterns& pt, const float* x,
                         const float* qhat, const float* retrieved, const float* lse, const int32_t* argmax,
                         const float* grad_out, const hop_grads& g, void* ws, size_t ws_bytes,
                         cudaStream_t st);
size_t simt_backward_workspace(const hop_desc& d);
size_t gemm_tail_workspace(const hop_desc& d);   // packs of the GEMM-shaped tails (behind the tensor-core core's workspace)
// hop_repair.cu -- exact nearest-pattern index behind the 16-bit tensor-core scores (HOP_F_EXACT_ARGMAX) and the
// fp32 path's stop rule of the retrieve updates (update_eps)
struct RepairArgs {
  // query rows in the arithmetic of the last pass: raw x + LN_q + beta (x_raw) or a prepared fp32 state (qs)
  const float* x_raw; const float* lnq_w; const float* lnq_b; float ln_eps; int has_ln; float beta;
  const float* qs;
  const float* khat;          // (P, H) fp32
  const float* kstats;        // [0] = max_j ||K^_j||
  int n, p, h, pattern_offset;
  const float* lse; float* rowmax; int32_t* argmax;   // results of the 16-bit pass; argmax / rowmax are corrected in place
  float* rowsec;              // CertArgs::sec: exact maximum of the 16-bit scores outside each row's top 64-column tile
  const float* rowtop;        // CertArgs::top: bound on the other columns of that tile (read when !precertified)
  void* ws; size_t ws_bytes;  // repair_workspace_bytes(n, h); its first 16 bytes are the control words, which the
                              // forward kernel of the same call clears (TcPassArgs::zero_words)
  int precertified;           // the forward's merge kernel already evaluated the certificate and built the list
};
// Where the merge kernels of the tensor-core forward evaluate the certificate themselves (they hold the row maximum and
// the runner-up bound of every row in registers): the repair list of hop_repair.cu, the query norms the forward's
// prologue wrote, and -- for a forward that ends without a merge kernel -- where it leaves the runner-up bound.
struct CertArgs {
  uint32_t* ctrl; int32_t* list; unsigned long long* best;   // repair list (ctrl == nullptr: no certificate)
  const float* kstats; float* qn;                             // max_j ||K^_j||, per-row ||q|| (N floats, written by the forward)
  float* sec;       // per row: EXACT maximum of the 16-bit scores outside the top 64-column thread tile (N floats; written by
                    // the forward without a merge kernel, else by the merge kernel for the rows it flags)
  float* top;       // per row: bound on the other columns of the top tile (N floats; forward without a merge kernel only)
};
CertArgs repair_cert_args(void* ws, int n, int h, const float* kstats, float* qn, float* sec, float* top);
size_t repair_workspace_bytes(int n, int h);
int launch_argmax_repair(const RepairArgs& a, cudaStream_t st);
// ||A_t - A_(t-1)||_F per batch item, max over items, compared with eps: sets *stopped (device) to 1 when the
// retrieve updates have converged (upstream's rule).  fp32, CUDA cores.
int launch_update_delta(const float* q_new, const float* q_old, const float* khat, int n, int p, int h,
                        int rows_per_item, float eps, float* row_scratch, uint32_t* stopped, cudaStream_t st);
// hop_fwd_tc.cu
struct TcPassArgs {
  const float* x_raw; const float* lnq_w; const float* lnq_b; float ln_eps; int has_ln; float beta;
  const float* qs;          // prepared beta*Q^ (N,H) fp32 if x_raw == nullptr
  const void* k16; const void* kt16; const void* vt16;
  const void* k16lo; const void* kt16lo;   // low halves (update passes of T >= 2 with hi + lo operand pairs) or nullptr
  const float* bias;        // (Dv) added to the read-out (folded out_proj bias) or nullptr
  float out_scale;          // scale of the normalised read-out (1 except for an update pass run as its own launch)
  float* qtmp[2];           // two (N,H) fp32 state buffers, needed when T > 0 and H > 256 (one launch per pass)
  int n, p, h, dv, update_steps; int bf16;
  float* retrieved;         // (N,Dv)
  float* qsave;             // (N,H) or nullptr
  float* lse; int32_t* argmax; float* rowmax; int pattern_offset;
  void* ws; size_t ws_bytes;   // partial results of the pattern splits (tc_forward_workspace bytes)
  TcTune tune;
  uint32_t* zero_words;        // 4 control words the first launch clears (repair counters) or nullptr
  CertArgs cert;               // certificate evaluated by the merge kernel of the launch that produces argmax
  bool* cert_done;             // set when that happened (else the caller runs certify_rows_kernel)
};
int tc_choose_splits(int n, int p, int d, int update_steps, const TcTune& tune);
bool tc_use_pair(int n, int d, const TcTune& tune);
bool tc_precise_updates(int d, int update_steps, const TcTune& tune);
size_t tc_forward_workspace(int n, int p, int d, int update_steps, const TcTune& tune);
int launch_tc_forward(const TcPassArgs& a, cudaStream_t st);
bool tc_supported(int h, int dv);
// hop_chain.cu -- Hop-VAE's bottleneck (three retrievals + sigmoid / round / mse glue) as one launch, fp32
size_t chain_workspace_bytes(int n);
int chain_supported(const hop_desc* d3);
int launch_chain(const hop_desc* d3, const hop_params* p3, const hop_patterns* pt3, float levels, const float* z,
                 float* emb, float* logits, float* idx_q, float* emb_recon, float* mse, void* ws, size_t ws_bytes,
                 cudaStream_t st);
// (allocation-free variant: the workspace is a file-scope device array of hop_chain.cu, rows <= chain_static_max_rows())
int chain_static_max_rows();
int launch_chain_static(const hop_desc* d3, const hop_params* p3, const hop_patterns* pt3, float levels, const float* z,
                        float* emb, float* logits, float* idx_q, float* emb_recon, float* mse, cudaStream_t st);
// hop_gemm_tc.cu -- C[M x N] = A[M x K] . B[N x K]^T (+ bias) on tcgen05, fp16 K-major operand packs, optional hi + lo pairs
constexpr float kGemmLoScale = 2048.f;   // low halves of the GEMM's operand pairs are stored times 2^11
struct TcGemmArgs {
  const void* a16; const void* a16lo;   // (m_pad, k) fp16, rows >= m zero
  const void* b16; const void* b16lo;   // (n, k) fp16
  int m, m_pad, n, k, ldc;
  const float* bias;                    // (n) or nullptr
  float* c;                             // (m, ldc) fp32
  int split;                            // operands are hi + lo pairs: three MMAs per product, ~2^-21 relative
  int k_splits;                         // > 1: the contraction is cut into that many ranges (grid.z), every CTA adds its
                                        // partial tile into a ZEROED c with fp32 REDs (short-and-wide products: weight gradients)
  const float* out_scale_dev;           // device scalar multiplied into the result (undoes an operand's range scaling) or nullptr
};
bool tc_gemm_supported(int n, int k);
int launch_tc_gemm(const TcGemmArgs& a, cudaStream_t st);
// hop_prep.cu: one HBM pass x [-> LayerNorm] [/ *range_dev] -> fp16 hi + lo operand packs of hop_gemm_tc.cu (low halves
// times kGemmLoScale), row-major (rows_pad, width) and / or transposed (width, rows_pad); rows >= `rows` are zeros
struct Pack16Args {
  const float* x; int rows, rows_pad, width;
  int has_ln; const float* ln_w; const float* ln_b; float ln_eps;
  const float* range_dev;
  void* y16; void* y16lo; void* yt16; void* yt16lo;
};
int launch_pack16(const Pack16Args& p, cudaStream_t st);
size_t prepare_scratch_bytes(const hop_desc& d);   // hop_prep.cu: scratch of the tensor-core pattern preparation (0: FFMA form)
// hop_bwd_tc.cu -- association core of the backward on the tensor pipe (dq, dK^, dV^)
struct TcBwdArgs {
  const float* qs; const float* lse; const float* dr; const float* r;       // (N,H), (N), dR (N,Dv), R = A V^ (N,Dv)
  const void* k16; const void* v16; const void* kt16;                        // pattern packs
  const float* khat; const int32_t* argmax;                                  // fp32 K^ (P,H), the forward's nearest pattern (N)
  float* dqs; float* dkhat; float* dvhat;                                    // (N,H), (P,H), (P,Dv)
  int n, p, h, dv;
  void* ws; size_t ws_bytes;
};
bool tc_backward_supported(int n, int p, int h, int dv);
size_t tc_backward_workspace(int n, int p, int h, int dv);
int launch_tc_backward(const TcBwdArgs& a, cudaStream_t st);
// hop_merge.cu
int launch_merge(int n, int dv, int g, int shard, const float* lse_all, const float* rowmax_all,
                 const int32_t* argmax_all, float* retrieved, float* lse_out, int32_t* argmax_out,
                 cudaStream_t st);

int launch_peer_merge(const hop_peer_merge& m, cudaStream_t st);

// -------------------------------------------------------------- device side
#ifdef __CUDACC__

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// LayerNorm of one row held by one warp: lane owns elements lane, lane+32, ...
// Two-pass (mean, then centred variance) in registers -- matches nn.LayerNorm's
// biased variance.  Returns (mean, rstd).
__device__ __forceinline__ void warp_row_stats(const float* __restrict__ row, int d, float eps,
                                               float& mean, float& rstd) {
  const int lane = threadIdx.x & 31;
  float s = 0.f;
  for (int i = lane; i < d; i += 32) s += row[i];
  mean = warp_sum(s) / (float)d;
  float v = 0.f;
  for (int i = lane; i < d; i += 32) {
    const float c = row[i] - mean;
    v += c * c;
  }
  rstd = rsqrtf(warp_sum(v) / (float)d + eps);
}

#endif  // __CUDACC__

}  // namespace hop
