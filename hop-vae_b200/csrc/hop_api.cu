// C-ABI entry points (include/hop_sm100.h): argument checking, path selection
// and orchestration of the kernels.  No device allocation, no host sync.
#include <cuda.h>

#include <atomic>
#include <cstdarg>
#include <cstring>
#include <mutex>

#include "hop_common.cuh"

namespace hop {

namespace {
thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};
std::mutex g_dev_mu;
int g_dev_checked[64];   // 0 unknown, 1 ok (cc 10.x), 2 wrong arch
int g_dev_sms[64];
}  // namespace

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

static int device_info(int* sms_out) {
  int dev = 0;
  HOP_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) HOP_FAIL(HOP_ERR_BAD_ARG, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_dev_mu);
  if (g_dev_checked[dev] == 0) {
    int major = 0, sms = 0;
    HOP_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    HOP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    g_dev_sms[dev] = sms;
    g_dev_checked[dev] = (major == 10) ? 1 : 2;
  }
  if (g_dev_checked[dev] != 1)
    HOP_FAIL(HOP_ERR_BAD_ARCH, "device %d is not compute capability 10.x (this library is sm_100a only)", dev);
  if (sms_out) *sms_out = g_dev_sms[dev];
  return HOP_OK;
}

int num_sms() {
  int sms = 148;
  if (device_info(&sms) != HOP_OK) return 148;
  return sms;
}

int validate_desc(const hop_desc* d) {
  if (!d) HOP_FAIL(HOP_ERR_BAD_ARG, "desc is NULL");
  if (d->abi_version != HOP_SM100_ABI_VERSION)
    HOP_FAIL(HOP_ERR_BAD_ARG, "desc.abi_version %u != library %u", d->abi_version, HOP_SM100_ABI_VERSION);
  if (d->n_rows <= 0 || d->n_patterns <= 0 || d->d_in <= 0 || d->d_assoc <= 0 || d->d_v <= 0)
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "non-positive dimension (N=%d P=%d Din=%d H=%d Dv=%d)", d->n_rows, d->n_patterns,
             d->d_in, d->d_assoc, d->d_v);
  if ((d->flags & HOP_F_OUT_PROJ) && d->d_out <= 0) HOP_FAIL(HOP_ERR_BAD_SHAPE, "HOP_F_OUT_PROJ with d_out=%d", d->d_out);
  if (!(d->update_eps >= 0.f)) HOP_FAIL(HOP_ERR_BAD_ARG, "update_eps must be >= 0");
  if (d->update_steps < 0 && !(d->update_eps > 0.f))
    HOP_FAIL(HOP_ERR_BAD_ARG, "update_steps < 0 (iterate until convergence) needs update_eps > 0");
  if (d->rows_per_item < 0 || (d->rows_per_item > 0 && d->n_rows % d->rows_per_item != 0))
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "rows_per_item (%d) must divide n_rows (%d)", d->rows_per_item, d->n_rows);
  if (d->tune_splits < 0) HOP_FAIL(HOP_ERR_BAD_ARG, "tune_splits must be >= 0");
  if (!(d->flags & HOP_F_PROJ_STATE) && d->d_assoc != d->d_in)
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "static state pattern needs d_assoc == d_in (%d vs %d)", d->d_assoc, d->d_in);
  if (!(d->flags & HOP_F_PROJ_STORED) && d->d_assoc != d->d_in)
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "static stored pattern needs d_assoc == d_in (%d vs %d)", d->d_assoc, d->d_in);
  if (!(d->flags & HOP_F_PROJ_PATTERN) && d->d_v != d->d_in)
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "static pattern projection needs d_v == d_in (%d vs %d)", d->d_v, d->d_in);
  if (d->path < HOP_PATH_AUTO || d->path > HOP_PATH_TC_BF16) HOP_FAIL(HOP_ERR_BAD_ARG, "unknown path %d", d->path);
  if (!(d->beta > 0.f) || !(d->ln_eps >= 0.f)) HOP_FAIL(HOP_ERR_BAD_ARG, "beta must be > 0 and ln_eps >= 0");
  return HOP_OK;
}

// out_proj can be folded into the read-out operand when the fused kernel's single width serves it
bool out_proj_foldable(const hop_desc& d) {
  return (d.flags & HOP_F_OUT_PROJ) && d.d_out == d.d_assoc && tc_supported(d.d_assoc, d.d_v);
}

int resolve_path(const hop_desc* d) {
  if (d->path == HOP_PATH_FP32_SIMT) return HOP_PATH_FP32_SIMT;
  const bool tc_ok = tc_supported(d->d_assoc, d->d_v);
  if (d->path == HOP_PATH_TC_F16 || d->path == HOP_PATH_TC_BF16) {
    if (!tc_ok) {
      set_error("tensor-core path is not built for H=%d Dv=%d (needs H == Dv in {64,128,256,512,1024})", d->d_assoc, d->d_v);
      return -HOP_ERR_UNSUPPORTED;
    }
    return d->path;
  }
  // AUTO: exact fp32 kernels for the reference's own (launch-bound, argmax-tight) sizes, for widths the MMA cannot
  // take, for update_steps < 0 and for layers whose score operands are not LayerNorm'd (fp16 encodes O(1) operands
  // with 11 bits but saturates at 65504: without normalize_state_pattern / normalize_stored_pattern nothing bounds
  // them, so AUTO keeps such layers in fp32 while the width allows); tcgen05 once the N x P score block is large.
  const long long np = (long long)d->n_rows * d->n_patterns;
  const bool normed = (d->flags & HOP_F_NORM_STATE) && (d->flags & HOP_F_NORM_STORED);
  if (d->d_assoc > 256 && tc_ok) return HOP_PATH_TC_F16;          // the fp32 kernels stop at width 256
  if (d->update_steps < 0 || !normed) return HOP_PATH_FP32_SIMT;
  if (tc_ok && np >= (1ll << 25)) return HOP_PATH_TC_F16;
  return HOP_PATH_FP32_SIMT;
}

// maximum number of retrieve updates the fp32 path enqueues for update_steps < 0 ("until convergence"): upstream
// loops on a host-side `.any()`; here the stop flag lives on the device and the passes behind it are identity passes
constexpr int kMaxUpdatesUntilConverged = 64;
static inline int simt_update_count(const hop_desc& d) { return d.update_steps < 0 ? kMaxUpdatesUntilConverged : d.update_steps; }
static inline bool simt_stop_rule(const hop_desc& d) {
  return d.update_eps > 0.f && (d.update_steps < 0 || d.update_steps >= 2);
}
static inline bool tc_exact_argmax(const hop_desc& d, const int32_t* argmax) {
  return (d.flags & HOP_F_EXACT_ARGMAX) && argmax != nullptr;
}

}  // namespace hop

using namespace hop;

extern "C" {

uint32_t hop_sm100_version(void) { return HOP_SM100_ABI_VERSION; }

const char* hop_sm100_last_error(void) { return g_err; }

uint64_t hop_sm100_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int hop_sm100_select_path(const hop_desc* desc, int32_t* path_out) {
  HOP_TRY(validate_desc(desc));
  if (!path_out) HOP_FAIL(HOP_ERR_BAD_ARG, "path_out is NULL");
  const int p = resolve_path(desc);
  if (p < 0) return -p;
  *path_out = p;
  return HOP_OK;
}

int hop_sm100_select_train_path(const hop_desc* desc, int32_t* path_out) {
  HOP_TRY(validate_desc(desc));
  if (!path_out) HOP_FAIL(HOP_ERR_BAD_ARG, "path_out is NULL");
  if (desc->update_steps != 0)
    HOP_FAIL(HOP_ERR_UNSUPPORTED, "backward is implemented for update_steps == 0 (what the reference trains with)");
  const int fwd = resolve_path(desc);
  if (fwd < 0) return -fwd;
  const bool tc_bwd = tc_backward_supported(desc->n_rows, desc->n_patterns, desc->d_assoc, desc->d_v) &&
                      !(desc->tune_flags & HOP_TUNE_BWD_SIMT);
  if (fwd != HOP_PATH_FP32_SIMT && tc_bwd) { *path_out = fwd; return HOP_OK; }
  if (desc->d_assoc > 256 || desc->d_v > 256)
    HOP_FAIL(HOP_ERR_UNSUPPORTED, "training at H=%d Dv=%d: the fp32 kernels stop at width 256 and the tensor-core "
             "backward is built for H == Dv in {128, 256, 512}", desc->d_assoc, desc->d_v);
  *path_out = HOP_PATH_FP32_SIMT;
  return HOP_OK;
}

int hop_sm100_pattern_bytes(const hop_desc* desc, hop_pattern_sizes* sz) {
  HOP_TRY(validate_desc(desc));
  if (!sz) HOP_FAIL(HOP_ERR_BAD_ARG, "sizes is NULL");
  const int path = resolve_path(desc);
  if (path < 0) return -path;
  const bool tc = (path != HOP_PATH_FP32_SIMT);
  const size_t P = (size_t)desc->n_patterns, Pp = (size_t)padded_patterns(desc->n_patterns);
  const bool train = tc && (desc->flags & HOP_F_TRAIN);
  const bool precise = tc && path == HOP_PATH_TC_F16 && tc_precise_updates(desc->d_assoc, desc->update_steps, tune_of(*desc));
  *sz = hop_pattern_sizes{};
  sz->khat = P * desc->d_assoc * sizeof(float);
  sz->vhat = P * desc->d_v * sizeof(float);
  sz->k16 = tc ? Pp * desc->d_assoc * 2 : 0;
  sz->kt16 = (tc && (desc->update_steps > 0 || train)) ? Pp * desc->d_assoc * 2 : 0;
  sz->vt16 = tc ? Pp * desc->d_v * 2 : 0;
  sz->vot16 = (tc && out_proj_foldable(*desc)) ? Pp * desc->d_out * 2 : 0;
  sz->k16lo = precise ? Pp * desc->d_assoc * 2 : 0;
  sz->kt16lo = precise ? Pp * desc->d_assoc * 2 : 0;
  sz->v16 = train ? Pp * desc->d_v * 2 : 0;
  sz->kstats = tc ? 4 * sizeof(float) : 0;
  sz->scratch = tc ? prepare_scratch_bytes(*desc) : 0;
  return HOP_OK;
}

namespace {
// forward workspace: [q state 0][q state 1][read-out][tensor-core split partials][repair][final state][lse, rowmax]
// (fp32 path with the stop rule: [q state 0][q state 1][read-out][row scratch (N)][flag])
struct FwdWs {
  size_t q, r, tc, repair, qfinal, stat, total;   // sizes of the regions (q: one state buffer)
};
FwdWs fwd_ws(const hop_desc& d, int path, bool exact) {
  FwdWs w{};
  const size_t N = (size_t)d.n_rows;
  w.q = round_up_sz(N * d.d_assoc * sizeof(float), 256);
  w.r = round_up_sz(N * d.d_v * sizeof(float), 256);
  if (path != HOP_PATH_FP32_SIMT) {
    w.tc = tc_forward_workspace(d.n_rows, d.n_patterns, d.d_assoc, d.update_steps, tune_of(d));
    if (exact) {
      w.repair = repair_workspace_bytes(d.n_rows, d.d_assoc);
      w.qfinal = d.update_steps > 0 ? w.q : 0;
      w.stat = round_up_sz(N * sizeof(float), 256);
    }
  } else if (simt_stop_rule(d)) {
    w.stat = round_up_sz(N * sizeof(float), 256);
  }
  w.total = 2 * w.q + w.r + w.tc + w.repair + w.qfinal + 5 * w.stat + (w.stat ? 256 : 0);   // stat: lse, rowmax, ||q||, runner-up (others, top tile)
  return w;
}
}  // namespace

int hop_sm100_workspace_bytes(const hop_desc* desc, int for_backward, size_t* bytes) {
  HOP_TRY(validate_desc(desc));
  if (!bytes) HOP_FAIL(HOP_ERR_BAD_ARG, "bytes is NULL");
  if (for_backward) {
    *bytes = simt_backward_workspace(*desc);      // (its total is a multiple of 256: the tensor-core part follows it)
    if (desc->path != HOP_PATH_FP32_SIMT && desc->path != HOP_PATH_AUTO && !(desc->tune_flags & HOP_TUNE_BWD_SIMT) &&
        tc_backward_supported(desc->n_rows, desc->n_patterns, desc->d_assoc, desc->d_v))
      *bytes += tc_backward_workspace(desc->n_rows, desc->n_patterns, desc->d_assoc, desc->d_v) + gemm_tail_workspace(*desc);
  } else {
    const int path = resolve_path(desc);
    if (path < 0) return -path;
    // (sized as if the nearest-pattern index were requested: the caller asks before it knows)
    *bytes = fwd_ws(*desc, path, (desc->flags & HOP_F_EXACT_ARGMAX) != 0).total;
  }
  return HOP_OK;
}

int hop_sm100_prepare_patterns(const hop_desc* desc, const hop_params* params, const hop_patterns* patterns,
                               void* stream) {
  HOP_TRY(validate_desc(desc));
  HOP_TRY(device_info(nullptr));
  if (!params || !patterns) HOP_FAIL(HOP_ERR_BAD_ARG, "params / patterns is NULL");
  if (!params->lookup || !patterns->khat || !patterns->vhat) HOP_FAIL(HOP_ERR_BAD_ARG, "lookup / khat / vhat is NULL");
  const int path = resolve_path(desc);
  if (path < 0) return -path;
  if (path != HOP_PATH_FP32_SIMT) {
    if (!patterns->k16 || !patterns->vt16 || (desc->update_steps > 0 && !patterns->kt16))
      HOP_FAIL(HOP_ERR_BAD_ARG, "tensor-core path needs the k16 / vt16 (/ kt16 when T > 0) buffers");
    if (path == HOP_PATH_TC_F16 && tc_precise_updates(desc->d_assoc, desc->update_steps, tune_of(*desc)) &&
        (!patterns->k16lo || !patterns->kt16lo))
      HOP_FAIL(HOP_ERR_BAD_ARG, "T >= 2 at D <= 256 runs its update passes with hi + lo operand pairs: k16lo / kt16lo are NULL");
    if ((desc->flags & HOP_F_TRAIN) && (!patterns->kt16 || !patterns->v16))
      HOP_FAIL(HOP_ERR_BAD_ARG, "HOP_F_TRAIN: the tensor-core backward needs the kt16 / v16 buffers");
    if (!aligned16(patterns->k16) || !aligned16(patterns->vt16) || !aligned16(patterns->kt16) ||
        !aligned16(patterns->k16lo) || !aligned16(patterns->kt16lo) || !aligned16(patterns->v16) ||
        !aligned16(patterns->kstats) || !aligned16(patterns->scratch))
      HOP_FAIL(HOP_ERR_BAD_ALIGN, "16-bit pattern packs must be 16-byte aligned");
    if (patterns->vot16 && (!out_proj_foldable(*desc) || !params->wo || !aligned16(patterns->vot16)))
      HOP_FAIL(HOP_ERR_BAD_ARG, "vot16 given but out_proj cannot be folded (needs HOP_F_OUT_PROJ, Dout == H == Dv, wo, alignment)");
  }
  if (!aligned16(params->lookup) || !aligned16(patterns->khat) || !aligned16(patterns->vhat))
    HOP_FAIL(HOP_ERR_BAD_ALIGN, "lookup / khat / vhat must be 16-byte aligned");
  return launch_prepare_patterns(*desc, *params, *patterns, path, (cudaStream_t)stream);
}

int hop_sm100_forward(const hop_desc* desc, const hop_params* params, const hop_patterns* patterns, const float* x,
                      float* out, float* retrieved, float* qhat, float* lse, int32_t* argmax, float* rowmax,
                      void* workspace, size_t workspace_bytes, void* stream) {
  HOP_TRY(validate_desc(desc));
  HOP_TRY(device_info(nullptr));
  if (!params || !patterns || !x || !out) HOP_FAIL(HOP_ERR_BAD_ARG, "params / patterns / x / out is NULL");
  if (!aligned16(x) || !aligned16(out) || !aligned16(retrieved) || !aligned16(qhat) || !aligned16(workspace))
    HOP_FAIL(HOP_ERR_BAD_ALIGN, "x / out / retrieved / qhat / workspace must be 16-byte aligned");
  const hop_desc& d = *desc;
  const hop_params& p = *params;
  const int path = resolve_path(desc);
  if (path < 0) return -path;
  const bool tc = path != HOP_PATH_FP32_SIMT;
  if (tc && d.update_steps < 0)
    HOP_FAIL(HOP_ERR_UNSUPPORTED, "update_steps < 0 (iterate until convergence) runs on the fp32 path only: the 16-bit "
                                  "score arithmetic cannot resolve update_eps-sized changes of the association");
  const bool exact = tc && tc_exact_argmax(d, argmax);
  const bool stop_rule = !tc && simt_stop_rule(d);
  const FwdWs W = fwd_ws(d, path, exact);
  const bool out_proj = (d.flags & HOP_F_OUT_PROJ) != 0;
  const bool q_proj = (d.flags & HOP_F_PROJ_STATE) != 0;
  const bool need_ws = q_proj || d.update_steps != 0 || W.tc > 0 || exact ||
                       (tc && out_proj && !retrieved && !(patterns->vot16 && out_proj_foldable(d)));
  if (need_ws && (!workspace || workspace_bytes < W.total))
    HOP_FAIL(HOP_ERR_WORKSPACE, "workspace of %zu bytes needed, %zu given", W.total, workspace_bytes);
  if (out_proj && (!p.wo)) HOP_FAIL(HOP_ERR_BAD_ARG, "HOP_F_OUT_PROJ set but params.wo is NULL");
  cudaStream_t st = (cudaStream_t)stream;

  char* wsb = static_cast<char*>(workspace);
  auto region = [&](size_t off) -> char* { return wsb ? wsb + off : nullptr; };
  float* qbuf[2] = {reinterpret_cast<float*>(region(0)), reinterpret_cast<float*>(region(W.q))};
  float* rbuf = reinterpret_cast<float*>(region(2 * W.q));
  const size_t off_tc = 2 * W.q + W.r, off_repair = off_tc + W.tc, off_qfinal = off_repair + W.repair;
  const size_t off_stat = off_qfinal + W.qfinal;

  const float* qs = nullptr;  // prepared beta*Q^ when not fused
  if (q_proj) {
    HOP_TRY(launch_prepare_queries(d, p, x, qbuf[0], st));
    qs = qbuf[0];
  }

  if (!tc) {
    // fp32 path: one launch per pass.  With the stop rule (update_eps > 0 and at least two possible updates) the
    // passes behind a converged iteration become identity passes -- decided on the device, no host synchronisation.
    const int T = simt_update_count(d);
    float* row_scratch = stop_rule ? reinterpret_cast<float*>(region(off_stat)) : nullptr;
    uint32_t* stopped = stop_rule ? reinterpret_cast<uint32_t*>(region(off_stat + 3 * W.stat)) : nullptr;
    if (stop_rule) HOP_CUDA(cudaMemsetAsync(stopped, 0, sizeof(uint32_t), st));
    float* cur = const_cast<float*>(qs);      // buffer holding the current state (nullptr: still the raw x, fused)
    float* prev = nullptr;
    for (int t = 0; t < T; ++t) {
      if (stop_rule && t >= 1)
        HOP_TRY(launch_update_delta(cur, prev, patterns->khat, d.n_rows, d.n_patterns, d.d_assoc, d.rows_per_item,
                                    d.update_eps, row_scratch, stopped, st));
      SimtPassArgs a{};
      a.x_raw = cur ? nullptr : x; a.lnq_w = p.ln_state_w; a.lnq_b = p.ln_state_b; a.ln_eps = d.ln_eps;
      a.has_ln = (d.flags & HOP_F_NORM_STATE) ? 1 : 0; a.beta = d.beta; a.qs = cur;
      a.khat = patterns->khat; a.bmat = patterns->khat;
      a.n = d.n_rows; a.p = d.n_patterns; a.h = d.d_assoc; a.dr = d.d_assoc; a.out_scale = d.beta;
      float* nxt = (cur == qbuf[0]) ? qbuf[1] : qbuf[0];
      a.res = nxt;
      if (stop_rule && !cur) a.qsave = qbuf[1];          // the first state, needed by the check behind update 1
      a.stopped = (stop_rule && t >= 1) ? stopped : nullptr;
      HOP_TRY(launch_simt_pass(a, st));
      prev = cur ? cur : qbuf[1];
      cur = nxt;
    }
    SimtPassArgs a{};
    a.x_raw = cur ? nullptr : x; a.lnq_w = p.ln_state_w; a.lnq_b = p.ln_state_b; a.ln_eps = d.ln_eps;
    a.has_ln = (d.flags & HOP_F_NORM_STATE) ? 1 : 0; a.beta = d.beta; a.qs = cur;
    a.khat = patterns->khat; a.bmat = patterns->vhat;
    a.n = d.n_rows; a.p = d.n_patterns; a.h = d.d_assoc; a.dr = d.d_v; a.out_scale = 1.f;
    a.qsave = qhat; a.lse = lse; a.argmax = argmax; a.rowmax = rowmax; a.pattern_offset = d.pattern_offset;
    if (out_proj) {
      a.res = retrieved; a.wo = p.wo; a.bo = p.bo; a.dout = d.d_out; a.out = out;
    } else {
      a.res = out;
    }
    HOP_TRY(launch_simt_pass(a, st));
    if (!out_proj && retrieved && retrieved != out)
      HOP_CUDA(cudaMemcpyAsync(retrieved, out, (size_t)d.n_rows * d.d_v * sizeof(float), cudaMemcpyDeviceToDevice, st));
    return HOP_OK;
  }

  // tensor-core path: one persistent kernel runs all T+1 passes (or one launch per pass, see launch_tc_forward).  With
  // the folded operand (V^ Wo^T)^T the kernel's read-out IS the layer output; otherwise out_proj is a second launch.
  const bool folded = out_proj && !retrieved && patterns->vot16 && out_proj_foldable(d);
  if (exact && (!patterns->kstats || !patterns->khat))
    HOP_FAIL(HOP_ERR_BAD_ARG, "HOP_F_EXACT_ARGMAX needs patterns.kstats and the fp32 K^");
  float* lse_w = lse, *rowmax_w = rowmax;
  if (exact) {      // the certificate needs both statistics even when the caller does not ask for them
    if (!lse_w) lse_w = reinterpret_cast<float*>(region(off_stat));
    if (!rowmax_w) rowmax_w = reinterpret_cast<float*>(region(off_stat + W.stat));
  }
  float* qfinal = (exact && d.update_steps > 0) ? reinterpret_cast<float*>(region(off_qfinal)) : nullptr;
  TcPassArgs a{};
  a.x_raw = qs ? nullptr : x; a.lnq_w = p.ln_state_w; a.lnq_b = p.ln_state_b; a.ln_eps = d.ln_eps;
  a.has_ln = (d.flags & HOP_F_NORM_STATE) ? 1 : 0; a.beta = d.beta; a.qs = qs;
  a.k16 = patterns->k16; a.kt16 = patterns->kt16; a.vt16 = folded ? patterns->vot16 : patterns->vt16;
  a.k16lo = patterns->k16lo; a.kt16lo = patterns->kt16lo;
  a.bias = folded ? p.bo : nullptr; a.out_scale = 1.f; a.qtmp[0] = qbuf[0]; a.qtmp[1] = qbuf[1];
  a.n = d.n_rows; a.p = d.n_patterns; a.h = d.d_assoc; a.dv = d.d_v; a.update_steps = d.update_steps;
  a.bf16 = (path == HOP_PATH_TC_BF16);
  a.retrieved = folded ? out : (out_proj ? (retrieved ? retrieved : rbuf) : out);
  a.qsave = qhat ? qhat : qfinal; a.lse = lse_w; a.argmax = argmax; a.rowmax = rowmax_w; a.pattern_offset = d.pattern_offset;
  a.ws = region(off_tc);
  a.ws_bytes = W.tc;
  a.tune = tune_of(d);
  a.zero_words = exact ? reinterpret_cast<uint32_t*>(region(off_repair)) : nullptr;
  bool cert_done = false;
  if (exact) {
    a.cert = repair_cert_args(region(off_repair), d.n_rows, d.d_assoc, patterns->kstats,
                              reinterpret_cast<float*>(region(off_stat + 2 * W.stat)),
                              reinterpret_cast<float*>(region(off_stat + 3 * W.stat)),
                              reinterpret_cast<float*>(region(off_stat + 4 * W.stat)));
    a.cert_done = &cert_done;
  }
  HOP_TRY(launch_tc_forward(a, st));
  if (exact) {
    RepairArgs r{};
    if (d.update_steps > 0) { r.qs = a.qsave; }
    else if (qs) { r.qs = qs; }
    else { r.x_raw = x; r.lnq_w = p.ln_state_w; r.lnq_b = p.ln_state_b; r.ln_eps = d.ln_eps; r.has_ln = a.has_ln; }
    r.beta = d.beta;
    r.khat = patterns->khat; r.kstats = patterns->kstats;
    r.n = d.n_rows; r.p = d.n_patterns; r.h = d.d_assoc; r.pattern_offset = d.pattern_offset;
    r.lse = lse_w; r.rowmax = rowmax_w; r.argmax = argmax; r.rowsec = a.cert.sec; r.rowtop = a.cert.top;
    r.ws = region(off_repair); r.ws_bytes = W.repair;
    r.precertified = cert_done ? 1 : 0;
    HOP_TRY(launch_argmax_repair(r, st));
  }
  if (folded) return HOP_OK;
  if (out_proj) HOP_TRY(launch_linear(d.n_rows, d.d_v, d.d_out, a.retrieved, p.wo, p.bo, out, st));
  else if (retrieved && retrieved != out)
    HOP_CUDA(cudaMemcpyAsync(retrieved, out, (size_t)d.n_rows * d.d_v * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return HOP_OK;
}

int hop_sm100_bottleneck_workspace_bytes(int32_t n_rows, size_t* bytes) {
  if (!bytes || n_rows <= 0) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck workspace: bytes is NULL or n_rows <= 0");
  *bytes = chain_workspace_bytes(n_rows);
  return HOP_OK;
}

int hop_sm100_bottleneck_forward(const hop_desc* descs, const hop_params* params, const hop_patterns* patterns,
                                 int32_t num_levels, const float* z, float* emb, float* logits, float* idx_q,
                                 float* emb_recon, float* mse, void* workspace, size_t workspace_bytes, void* stream) {
  if (!descs || !params || !patterns) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck: descs / params / patterns is NULL");
  for (int l = 0; l < 3; ++l) HOP_TRY(validate_desc(descs + l));
  HOP_TRY(device_info(nullptr));
  if (num_levels < 2) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck: num_levels must be >= 2");
  if (!emb) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck: emb is NULL");
  if (!aligned16(z) || !aligned16(workspace)) HOP_FAIL(HOP_ERR_BAD_ALIGN, "bottleneck: z / workspace must be 16-byte aligned");
  return launch_chain(descs, params, patterns, (float)(num_levels - 1), z, emb, logits, idx_q, emb_recon, mse, workspace,
                      workspace_bytes, (cudaStream_t)stream);
}

int hop_sm100_bottleneck_forward_static(const hop_desc* descs, const hop_params* params, const hop_patterns* patterns,
                                        int32_t num_levels, const float* z, float* emb, float* logits, float* idx_q,
                                        float* emb_recon, float* mse, void* stream) {
  if (!descs || !params || !patterns) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck: descs / params / patterns is NULL");
  for (int l = 0; l < 3; ++l) HOP_TRY(validate_desc(descs + l));
  HOP_TRY(device_info(nullptr));
  if (num_levels < 2) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck: num_levels must be >= 2");
  if (!emb) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck: emb is NULL");
  if (!aligned16(z)) HOP_FAIL(HOP_ERR_BAD_ALIGN, "bottleneck: z must be 16-byte aligned");
  return launch_chain_static(descs, params, patterns, (float)(num_levels - 1), z, emb, logits, idx_q, emb_recon, mse,
                             (cudaStream_t)stream);
}

int hop_sm100_bottleneck_static_max_rows(void) { return chain_static_max_rows(); }

int hop_sm100_repair_rows(const hop_desc* desc, const void* workspace, void* stream, int32_t* rows_out,
                          int32_t* flagged_out) {
  HOP_TRY(validate_desc(desc));
  if (!rows_out) HOP_FAIL(HOP_ERR_BAD_ARG, "rows_out is NULL");
  *rows_out = 0;
  const int path = resolve_path(desc);
  if (path < 0) return -path;
  if (path == HOP_PATH_FP32_SIMT || !(desc->flags & HOP_F_EXACT_ARGMAX)) return HOP_OK;
  if (!workspace) HOP_FAIL(HOP_ERR_BAD_ARG, "workspace is NULL");
  const FwdWs W = fwd_ws(*desc, path, true);
  uint32_t ctrl[4] = {0, 0, 0, 0};   // control words of the repair region (hop_repair.cu): [0] rows that failed the
                                     // certificate, [2] rows the block check could not settle -> full fp32 pass
  HOP_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  HOP_CUDA(cudaMemcpy(ctrl, static_cast<const char*>(workspace) + 2 * W.q + W.r + W.tc, sizeof(ctrl), cudaMemcpyDeviceToHost));
  *rows_out = (int32_t)ctrl[2];
  if (flagged_out) *flagged_out = (int32_t)ctrl[0];
  return HOP_OK;
}

int hop_sm100_backward(const hop_desc* desc, const hop_params* params, const hop_patterns* patterns, const float* x,
                       const float* qhat, const float* retrieved, const float* lse, const int32_t* argmax,
                       const float* grad_out, const hop_grads* grads, void* workspace, size_t workspace_bytes,
                       void* stream) {
  HOP_TRY(validate_desc(desc));
  HOP_TRY(device_info(nullptr));
  if (!params || !patterns || !x || !qhat || !retrieved || !lse || !grad_out || !grads)
    HOP_FAIL(HOP_ERR_BAD_ARG, "a required backward argument is NULL");
  if (desc->update_steps != 0)
    HOP_FAIL(HOP_ERR_UNSUPPORTED, "backward is implemented for update_steps == 0 (what the reference trains with)");
  size_t need = 0;
  HOP_TRY(hop_sm100_workspace_bytes(desc, 1, &need));
  if (!workspace || workspace_bytes < need)
    HOP_FAIL(HOP_ERR_WORKSPACE, "backward workspace of %zu bytes needed, %zu given", need, workspace_bytes);
  if (!aligned16(workspace) || !aligned16(grad_out) || !aligned16(qhat) || !aligned16(retrieved))
    HOP_FAIL(HOP_ERR_BAD_ALIGN, "backward buffers must be 16-byte aligned");
  return launch_simt_backward(*desc, *params, *patterns, x, qhat, retrieved, lse, argmax, grad_out, *grads, workspace,
                              workspace_bytes, (cudaStream_t)stream);
}

int hop_sm100_merge_shards(int32_t n_rows, int32_t d_v, int32_t n_shards, int32_t shard, const float* lse_all,
                           const float* rowmax_all, const int32_t* argmax_all, float* retrieved, float* lse_out,
                           int32_t* argmax_out, void* stream) {
  HOP_TRY(device_info(nullptr));
  if (n_rows <= 0 || d_v <= 0 || n_shards <= 0 || shard < 0 || shard >= n_shards)
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "merge: bad sizes (N=%d Dv=%d G=%d g=%d)", n_rows, d_v, n_shards, shard);
  if (!lse_all || !retrieved) HOP_FAIL(HOP_ERR_BAD_ARG, "merge: lse_all / retrieved is NULL");
  if (!aligned16(retrieved)) HOP_FAIL(HOP_ERR_BAD_ALIGN, "merge: retrieved must be 16-byte aligned");
  return launch_merge(n_rows, d_v, n_shards, shard, lse_all, rowmax_all, argmax_all, retrieved, lse_out, argmax_out,
                      (cudaStream_t)stream);
}

int hop_sm100_out_projection(int32_t n_rows, int32_t d_v, int32_t d_out, const float* retrieved, const float* wo,
                             const float* bo, float* out, void* stream) {
  HOP_TRY(device_info(nullptr));
  if (n_rows <= 0 || d_v <= 0 || d_out <= 0) HOP_FAIL(HOP_ERR_BAD_SHAPE, "out_projection: bad sizes");
  if (!retrieved || !wo || !out) HOP_FAIL(HOP_ERR_BAD_ARG, "out_projection: NULL argument");
  return launch_linear(n_rows, d_v, d_out, retrieved, wo, bo, out, (cudaStream_t)stream);
}

int hop_sm100_peer_merge(const hop_peer_merge* m, void* stream) {
  HOP_TRY(device_info(nullptr));
  if (!m) HOP_FAIL(HOP_ERR_BAD_ARG, "peer_merge: descriptor is NULL");
  if (m->n_ranks < 1 || m->n_ranks > HOP_MAX_PEERS || m->rank < 0 || m->rank >= m->n_ranks)
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "peer_merge: rank %d of %d (at most %d peers)", m->rank, m->n_ranks, HOP_MAX_PEERS);
  if (m->n_rows <= 0 || m->d_v <= 0 || (m->d_v & 3) || m->row_begin < 0 || m->row_count < 0 ||
      m->row_begin + m->row_count > m->n_rows)
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "peer_merge: bad sizes (N=%d Dv=%d rows [%d, +%d))", m->n_rows, m->d_v, m->row_begin,
             m->row_count);
  if (m->epoch == 0) HOP_FAIL(HOP_ERR_BAD_ARG, "peer_merge: epoch must be non-zero");
  for (int g = 0; g < m->n_ranks; ++g) {
    if (!m->o[g] || !m->lse[g] || !m->rowmax[g] || !m->argmax[g] || !m->out[g] || !m->flags[g])
      HOP_FAIL(HOP_ERR_BAD_ARG, "peer_merge: a buffer of rank %d is NULL", g);
    if (!aligned16(m->o[g]) || !aligned16(m->out[g])) HOP_FAIL(HOP_ERR_BAD_ALIGN, "peer_merge: o / out must be 16-byte aligned");
  }
  return launch_peer_merge(*m, (cudaStream_t)stream);
}

int hop_sm100_enable_peer_access(int32_t peer_device) {
  int dev = 0;
  HOP_CUDA(cudaGetDevice(&dev));
  if (peer_device == dev) return HOP_OK;
  int can = 0;
  HOP_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  if (!can) HOP_FAIL(HOP_ERR_UNSUPPORTED, "device %d cannot access device %d's memory (no P2P path)", dev, peer_device);
  const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); return HOP_OK; }
  HOP_CUDA(e);
  return HOP_OK;
}

int hop_sm100_ipc_export(const void* ptr, unsigned char handle[HOP_IPC_HANDLE_BYTES], uint64_t* offset) {
  static_assert(sizeof(cudaIpcMemHandle_t) == HOP_IPC_HANDLE_BYTES, "CUDA IPC handle size");
  if (!ptr || !handle || !offset) HOP_FAIL(HOP_ERR_BAD_ARG, "ipc_export: NULL argument");
  using RangeFn = CUresult (*)(CUdeviceptr*, size_t*, CUdeviceptr);
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fp, cudaEnableDefault, &q) != cudaSuccess || !fp)
    HOP_FAIL(HOP_ERR_CUDA, "cuMemGetAddressRange is not available from the driver");
  CUdeviceptr base = 0;
  size_t size = 0;
  const CUresult r = reinterpret_cast<RangeFn>(fp)(&base, &size, (CUdeviceptr)(uintptr_t)ptr);
  if (r != CUDA_SUCCESS) HOP_FAIL(HOP_ERR_CUDA, "cuMemGetAddressRange failed (%d)", (int)r);
  cudaIpcMemHandle_t h;
  HOP_CUDA(cudaIpcGetMemHandle(&h, reinterpret_cast<void*>((uintptr_t)base)));
  memcpy(handle, &h, sizeof(h));
  *offset = (uint64_t)((uintptr_t)ptr - (uintptr_t)base);
  return HOP_OK;
}

int hop_sm100_ipc_import(const unsigned char handle[HOP_IPC_HANDLE_BYTES], uint64_t offset, void** ptr) {
  if (!handle || !ptr) HOP_FAIL(HOP_ERR_BAD_ARG, "ipc_import: NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* base = nullptr;
  HOP_CUDA(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr = static_cast<char*>(base) + offset;
  return HOP_OK;
}

int hop_sm100_ipc_close(void* ptr, uint64_t offset) {
  if (!ptr) return HOP_OK;
  HOP_CUDA(cudaIpcCloseMemHandle(static_cast<char*>(ptr) - offset));
  return HOP_OK;
}

}  // extern "C"
