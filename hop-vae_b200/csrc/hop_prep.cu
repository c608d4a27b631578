// Row-wise [LayerNorm] -> [Linear] -> scale, with optional 16-bit row-major and
// transposed packs for the tensor-core path.
//
// One kernel serves:
//   K^ = [LN_k](W)[Wk^T + bk]       -> fp32 (P,H), 16-bit (Ppad,H), 16-bit^T (H,Ppad)
//   V^ = [LN_v](Wt)[Wv^T + bv]      -> fp32 (P,Dv), 16-bit^T (Dv,Ppad)
//   Q^ = beta * [LN_q](x)[Wq^T+bq]  -> fp32 (N,H)         (only when Q is projected)
//   out = retrieved . Wo^T + bo     -> fp32 (N,Dout)      (after a shard merge / TC path)
// Upstream does the K^/V^ work on the batch-expanded lookup matrix, i.e. B
// times per forward; here it runs once per parameter update.
#include "hop_common.cuh"

namespace hop {

namespace {

constexpr int NT = 256;
constexpr int TN = 64;        // output-column tile
constexpr int KC = 64;        // contraction chunk
constexpr int XS = KC + 4;    // smem stride of X / W chunks
constexpr int YS = TN + 1;    // smem stride of the staged output tile

struct LnLinearArgs {
  const float* x; int rows; int rows_pad; int din;
  int has_ln; const float* ln_w; const float* ln_b; float ln_eps;
  const float* w; const float* b; int dout;   // w == nullptr -> identity (dout == din)
  float scale;
  float* y;            // (rows, dout) fp32 or nullptr
  void* y16;           // (rows_pad, dout) 16-bit or nullptr
  void* yt16;          // (dout, rows_pad) 16-bit or nullptr
  int bf16;            // 16-bit encoding: 0 = fp16 (saturating), 1 = bf16
  void* y16lo;         // (rows_pad, dout) 16-bit: y - decode(y16), the low half of a hi + lo operand pair, or nullptr
  void* yt16lo;        // (dout, rows_pad) same, transposed, or nullptr
  float lo_scale;      // the low halves hold lo_scale * (y - decode(y16)) (0 = 1; hop_gemm_tc.cu takes 2^11, which keeps the
                       // residual of small-magnitude operands out of the fp16 subnormals)
  float* norm_max;     // [1]: max over rows of ||y_row||_2 (atomic max on the float's bits; zeroed by the caller) or nullptr
  const float* range_dev;   // device scalar r > 0: the output is divided by it (gradients of any magnitude -> fp16's range) or nullptr
};

__device__ __forceinline__ uint16_t to16(float v, int bf16) {
  if (bf16) return __bfloat16_as_ushort(__float2bfloat16_rn(v));
  // fp16: clamp to the finite range so an outlier cannot turn into inf inside the MMA
  v = fminf(fmaxf(v, -65504.f), 65504.f);
  return __half_as_ushort(__float2half_rn(v));
}

__device__ __forceinline__ float from16(uint16_t w, int bf16) {
  return bf16 ? __bfloat162float(__ushort_as_bfloat16(w)) : __half2float(__ushort_as_half(w));
}
// low half of a hi + lo operand pair: what the 16-bit encoding of v lost, encoded again
__device__ __forceinline__ uint16_t to16_lo(float v, int bf16, float scale) {
  return to16((v - from16(to16(v, bf16), bf16)) * scale, bf16);
}

template <int RM>
__device__ __forceinline__ void ln_linear_body(const LnLinearArgs& a) {
  constexpr int BR = 16 * RM;
  __shared__ __align__(16) float Xs[BR * XS];
  __shared__ __align__(16) float Ws[TN * XS];
  __shared__ float s_mean[BR], s_rstd[BR], s_sq[BR];
  float* Ys = Xs;  // the staged output tile aliases the input chunk (BR*YS <= BR*XS), see the barriers below
  static_assert(YS <= XS, "output staging tile must fit in the input chunk");

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * BR;
  float out_scale = a.scale;
  if (a.range_dev) { const float r = __ldg(a.range_dev); out_scale = r > 0.f ? a.scale / r : a.scale; }

  // per-row LayerNorm statistics
  for (int r = warp; r < BR; r += NT / 32) {
    float mean = 0.f, rstd = 1.f;
    if (a.has_ln && row0 + r < a.rows) warp_row_stats(a.x + (size_t)(row0 + r) * a.din, a.din, a.ln_eps, mean, rstd);
    if (lane == 0) {
      s_mean[r] = mean;
      s_rstd[r] = rstd;
      s_sq[r] = 0.f;
    }
  }
  __syncthreads();

  for (int c0 = 0; c0 < a.dout; c0 += TN) {
    float acc[RM][4];
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int bb = 0; bb < 4; ++bb) acc[i][bb] = 0.f;

    const int k_begin = a.w ? 0 : c0;
    const int k_end = a.w ? a.din : min(c0 + KC, a.din);
    for (int k0 = k_begin; k0 < k_end; k0 += KC) {
      __syncthreads();
      // normalised input chunk
      for (int idx = tid; idx < BR * KC; idx += NT) {
        const int r = idx / KC, k = idx - r * KC;
        float v = 0.f;
        if (row0 + r < a.rows && k0 + k < a.din) {
          v = __ldg(a.x + (size_t)(row0 + r) * a.din + k0 + k);
          if (a.has_ln) {
            v = (v - s_mean[r]) * s_rstd[r];
            if (a.ln_w) v *= __ldg(a.ln_w + k0 + k);
            if (a.ln_b) v += __ldg(a.ln_b + k0 + k);
          }
        }
        Xs[r * XS + k] = v;
      }
      if (a.w) {
        for (int idx = tid; idx < TN * KC; idx += NT) {
          const int c = idx / KC, k = idx - c * KC;
          float v = 0.f;
          if (c0 + c < a.dout && k0 + k < a.din) v = __ldg(a.w + (size_t)(c0 + c) * a.din + k0 + k);
          Ws[c * XS + k] = v;
        }
      }
      __syncthreads();
      if (a.w) {
#pragma unroll 4
        for (int k = 0; k < KC; k += 4) {
          float4 xv[RM];
#pragma unroll
          for (int i = 0; i < RM; ++i) xv[i] = *reinterpret_cast<const float4*>(Xs + (ty * RM + i) * XS + k);
#pragma unroll
          for (int bb = 0; bb < 4; ++bb) {
            const float4 wv = *reinterpret_cast<const float4*>(Ws + (tx + 16 * bb) * XS + k);
#pragma unroll
            for (int i = 0; i < RM; ++i) {
              acc[i][bb] = fmaf(xv[i].x, wv.x, acc[i][bb]);
              acc[i][bb] = fmaf(xv[i].y, wv.y, acc[i][bb]);
              acc[i][bb] = fmaf(xv[i].z, wv.z, acc[i][bb]);
              acc[i][bb] = fmaf(xv[i].w, wv.w, acc[i][bb]);
            }
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < RM; ++i)
#pragma unroll
          for (int bb = 0; bb < 4; ++bb) acc[i][bb] = Xs[(ty * RM + i) * XS + tx + 16 * bb];
      }
    }

    // stage the finished (BR x TN) tile, then write every requested layout coalesced
    __syncthreads();
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int bb = 0; bb < 4; ++bb) {
        const int c = tx + 16 * bb;
        float v = acc[i][bb];
        if (a.w && a.b && c0 + c < a.dout) v += __ldg(a.b + c0 + c);
        Ys[(ty * RM + i) * YS + c] = v * out_scale;
      }
    __syncthreads();
    for (int idx = tid; idx < BR * TN; idx += NT) {
      const int r = idx / TN, c = idx - r * TN;
      const int gr = row0 + r, gc = c0 + c;
      if (gc >= a.dout) continue;
      const float v = (gr < a.rows) ? Ys[r * YS + c] : 0.f;
      if (a.y && gr < a.rows) a.y[(size_t)gr * a.dout + gc] = v;
      if (a.y16 && gr < a.rows_pad) reinterpret_cast<uint16_t*>(a.y16)[(size_t)gr * a.dout + gc] = to16(v, a.bf16);
      if (a.y16lo && gr < a.rows_pad) reinterpret_cast<uint16_t*>(a.y16lo)[(size_t)gr * a.dout + gc] = to16_lo(v, a.bf16, a.lo_scale > 0.f ? a.lo_scale : 1.f);
      if (a.norm_max) atomicAdd(&s_sq[r], v * v);
    }
    if (a.yt16) {
      for (int idx = tid; idx < BR * TN; idx += NT) {
        const int c = idx / BR, r = idx - c * BR;
        const int gr = row0 + r, gc = c0 + c;
        if (gc >= a.dout || gr >= a.rows_pad) continue;
        const float v = (gr < a.rows) ? Ys[r * YS + c] : 0.f;
        reinterpret_cast<uint16_t*>(a.yt16)[(size_t)gc * a.rows_pad + gr] = to16(v, a.bf16);
        if (a.yt16lo) reinterpret_cast<uint16_t*>(a.yt16lo)[(size_t)gc * a.rows_pad + gr] = to16_lo(v, a.bf16, a.lo_scale > 0.f ? a.lo_scale : 1.f);
      }
    }
  }
  if (a.norm_max) {
    __syncthreads();
    if (tid < BR) atomicMax(reinterpret_cast<int*>(a.norm_max), __float_as_int(sqrtf(s_sq[tid])));   // (>= 0: bit order = value order)
  }
}

template <int RM>
__global__ void __launch_bounds__(NT) ln_linear_kernel(LnLinearArgs a) { ln_linear_body<RM>(a); }

// Two independent jobs over the same number of rows in one launch (blockIdx.y picks the job): the K^ and V^ packs of
// prepare_patterns -- at the reference's sizes each is a ~5 us launch-latency-bound kernel, once per training step.
template <int RM>
__global__ void __launch_bounds__(NT) ln_linear_pair_kernel(LnLinearArgs a0, LnLinearArgs a1) {
  ln_linear_body<RM>(blockIdx.y == 0 ? a0 : a1);
}

int launch_ln_linear_pair(const LnLinearArgs& a0, const LnLinearArgs& a1, cudaStream_t st) {
  const int total_rows = (a0.y16 || a0.yt16 || a0.y16lo || a1.y16 || a1.yt16) ? a0.rows_pad : a0.rows;
  if (total_rows <= 0) return HOP_OK;
  const int sms = num_sms();
  if (total_rows >= 64 * sms) {
    ln_linear_pair_kernel<4><<<dim3((total_rows + 63) / 64, 2), NT, 0, st>>>(a0, a1);
  } else if (total_rows >= 32 * sms) {
    ln_linear_pair_kernel<2><<<dim3((total_rows + 31) / 32, 2), NT, 0, st>>>(a0, a1);
  } else {
    ln_linear_pair_kernel<1><<<dim3((total_rows + 15) / 16, 2), NT, 0, st>>>(a0, a1);
  }
  HOP_LAUNCH_CHECK("ln_linear_pair_kernel");
  return HOP_OK;
}

int launch_ln_linear(const LnLinearArgs& a, cudaStream_t st) {
  const int total_rows = (a.y16 || a.yt16 || a.y16lo) ? a.rows_pad : a.rows;
  if (total_rows <= 0) return HOP_OK;
  const int sms = num_sms();
  if (total_rows >= 64 * sms) {
    ln_linear_kernel<4><<<(total_rows + 63) / 64, NT, 0, st>>>(a);
  } else if (total_rows >= 32 * sms) {
    ln_linear_kernel<2><<<(total_rows + 31) / 32, NT, 0, st>>>(a);
  } else {
    ln_linear_kernel<1><<<(total_rows + 15) / 16, NT, 0, st>>>(a);
  }
  HOP_LAUNCH_CHECK("ln_linear_kernel");
  return HOP_OK;
}

}  // namespace

// ---- tensor-core form of the V path (V-projection and out_proj fold as tcgen05 GEMMs, hop_gemm_tc.cu) -------------
// scratch: [A hi | A lo] (Ppad x max(Din, Dv) fp16 each: LN_v(W), later V^), [B hi | B lo] (max(Dv x Din, Dout x Dv)),
//          V^ Wo^T fp32 (P x Dout)
namespace {
struct PrepScratch { char *a_hi, *a_lo, *b_hi, *b_lo; float* vo; size_t total; };
PrepScratch carve_prep(void* base, const hop_desc& d) {
  PrepScratch s{};
  size_t off = 0;
  auto take = [&](size_t bytes) {
    char* p = base ? static_cast<char*>(base) + off : nullptr;
    off += round_up_sz(bytes, 1024);
    return p;
  };
  const size_t ppad = (size_t)padded_patterns(d.n_patterns);
  const size_t wa = (size_t)(d.d_in > d.d_v ? d.d_in : d.d_v);
  size_t wb = (size_t)d.d_v * d.d_in;
  const bool fold = out_proj_foldable(d);
  if (fold && (size_t)d.d_out * d.d_v > wb) wb = (size_t)d.d_out * d.d_v;
  s.a_hi = take(ppad * wa * 2); s.a_lo = take(ppad * wa * 2);
  s.b_hi = take(wb * 2); s.b_lo = take(wb * 2);
  s.vo = reinterpret_cast<float*>(take(fold ? (size_t)d.n_patterns * d.d_out * sizeof(float) : 0));
  s.total = off;
  return s;
}
bool prep_uses_gemm(const hop_desc& d) {
  return (d.flags & HOP_F_PROJ_PATTERN) && !(d.tune_flags & HOP_TUNE_PREP_SIMT) && tc_gemm_supported(d.d_v, d.d_in) &&
         d.n_patterns >= 1024 &&
         (!out_proj_foldable(d) || tc_gemm_supported(d.d_out, d.d_v));
}
}  // namespace

size_t prepare_scratch_bytes(const hop_desc& d) { return prep_uses_gemm(d) ? carve_prep(nullptr, d).total : 0; }

int launch_ln_pack(const LnLinearArgs& a, cudaStream_t st);   // below: the pack pass where the job is a pure LayerNorm

int launch_prepare_patterns(const hop_desc& d, const hop_params& p, const hop_patterns& pt, int path,
                            cudaStream_t st) {
  const bool tc = (path == HOP_PATH_TC_F16 || path == HOP_PATH_TC_BF16);
  const int ppad = padded_patterns(d.n_patterns);
  LnLinearArgs k{};
  k.x = p.lookup; k.rows = d.n_patterns; k.rows_pad = ppad; k.din = d.d_in;
  k.has_ln = (d.flags & HOP_F_NORM_STORED) ? 1 : 0; k.ln_w = p.ln_stored_w; k.ln_b = p.ln_stored_b; k.ln_eps = d.ln_eps;
  k.w = (d.flags & HOP_F_PROJ_STORED) ? p.wk : nullptr; k.b = (d.flags & HOP_F_PROJ_STORED) ? p.bk : nullptr;
  k.dout = d.d_assoc; k.scale = 1.f; k.y = pt.khat;
  k.y16 = tc ? pt.k16 : nullptr; k.yt16 = tc ? pt.kt16 : nullptr;   // (kt16: only when T > 0 or HOP_F_TRAIN, else NULL)
  k.y16lo = tc ? pt.k16lo : nullptr; k.yt16lo = (tc && pt.kt16) ? pt.kt16lo : nullptr;
  k.bf16 = (path == HOP_PATH_TC_BF16);
  if (tc && pt.kstats) {
    HOP_CUDA(cudaMemsetAsync(pt.kstats, 0, 4 * sizeof(float), st));
    k.norm_max = pt.kstats;
  }
  if ((d.flags & HOP_F_PROJ_STORED) && !p.wk) HOP_FAIL(HOP_ERR_BAD_ARG, "HOP_F_PROJ_STORED set but params.wk is NULL");
  if ((d.flags & HOP_F_PROJ_PATTERN) && !p.wv) HOP_FAIL(HOP_ERR_BAD_ARG, "HOP_F_PROJ_PATTERN set but params.wv is NULL");
  const float* Wt = p.lookup_targets ? p.lookup_targets : p.lookup;
  const int bf16 = (path == HOP_PATH_TC_BF16);

  if (tc && !bf16 && pt.scratch && prep_uses_gemm(d)) {
    // V^ = LN_v(Wt) Wv^T + bv and (V^ Wo^T) on the tensor pipe; every pack kernel around them is a pass over HBM
    const PrepScratch s = carve_prep(pt.scratch, d);
    HOP_TRY(launch_ln_pack(k, st));                                     // K^ path (no contraction when K is static)
    auto pack = [&](const float* x, int rows, int rows_pad, int w, int has_ln, const float* lw, const float* lb,
                    float* y, void* y16, void* y16lo, void* yt16) {
      (void)y;                            // (no fp32 copy is needed anywhere on this path)
      Pack16Args a{x, rows, rows_pad, w, has_ln, lw, lb, d.ln_eps, nullptr, y16, y16lo, yt16, nullptr};
      return launch_pack16(a, st);
    };
    HOP_TRY(pack(Wt, d.n_patterns, ppad, d.d_in, (d.flags & HOP_F_NORM_PROJ) ? 1 : 0, p.ln_proj_w, p.ln_proj_b, nullptr,
                 s.a_hi, s.a_lo, nullptr));                             // LN_v(Wt) as a hi + lo pair
    HOP_TRY(pack(p.wv, d.d_v, d.d_v, d.d_in, 0, nullptr, nullptr, nullptr, s.b_hi, s.b_lo, nullptr));
    TcGemmArgs g{};
    g.a16 = s.a_hi; g.a16lo = s.a_lo; g.b16 = s.b_hi; g.b16lo = s.b_lo;
    g.m = d.n_patterns; g.m_pad = ppad; g.n = d.d_v; g.k = d.d_in; g.ldc = d.d_v; g.bias = p.bv; g.c = pt.vhat; g.split = 1;
    HOP_TRY(launch_tc_gemm(g, st));
    const bool fold = pt.vot16 != nullptr;
    // V^ -> vt16 [, v16]; for the fold also the hi + lo pair of V^ (the A operand of the second product)
    void* vhi = pt.v16 ? pt.v16 : (fold ? (void*)s.a_hi : nullptr);
    HOP_TRY(pack(pt.vhat, d.n_patterns, ppad, d.d_v, 0, nullptr, nullptr, nullptr, vhi, fold ? (void*)s.a_lo : nullptr, pt.vt16));
    if (fold) {
      HOP_TRY(pack(p.wo, d.d_out, d.d_out, d.d_v, 0, nullptr, nullptr, nullptr, s.b_hi, s.b_lo, nullptr));
      TcGemmArgs f{};
      f.a16 = vhi; f.a16lo = s.a_lo; f.b16 = s.b_hi; f.b16lo = s.b_lo;
      f.m = d.n_patterns; f.m_pad = ppad; f.n = d.d_out; f.k = d.d_v; f.ldc = d.d_out; f.bias = nullptr; f.c = s.vo; f.split = 1;
      HOP_TRY(launch_tc_gemm(f, st));
      HOP_TRY(pack(s.vo, d.n_patterns, ppad, d.d_out, 0, nullptr, nullptr, nullptr, nullptr, nullptr, pt.vot16));
    }
    return HOP_OK;
  }

  LnLinearArgs v{};
  v.x = Wt; v.rows = d.n_patterns; v.rows_pad = ppad; v.din = d.d_in;
  v.has_ln = (d.flags & HOP_F_NORM_PROJ) ? 1 : 0; v.ln_w = p.ln_proj_w; v.ln_b = p.ln_proj_b; v.ln_eps = d.ln_eps;
  v.w = (d.flags & HOP_F_PROJ_PATTERN) ? p.wv : nullptr; v.b = (d.flags & HOP_F_PROJ_PATTERN) ? p.bv : nullptr;
  v.dout = d.d_v; v.scale = 1.f; v.y = pt.vhat;
  v.y16 = tc ? pt.v16 : nullptr; v.yt16 = tc ? pt.vt16 : nullptr;
  v.bf16 = bf16;
  HOP_TRY(launch_ln_linear_pair(k, v, st));     // K^ and V^ packs: same rows, one launch

  if (tc && pt.vot16) {
    // out_proj folded into the read-out operand: (V^ Wo^T)^T, 16-bit, from the fp32 V^ just written
    LnLinearArgs o{};
    o.x = pt.vhat; o.rows = d.n_patterns; o.rows_pad = ppad; o.din = d.d_v; o.has_ln = 0; o.ln_eps = 0.f;
    o.w = p.wo; o.b = nullptr; o.dout = d.d_out; o.scale = 1.f; o.y = nullptr; o.y16 = nullptr; o.yt16 = pt.vot16;
    o.bf16 = bf16;
    HOP_TRY(launch_ln_linear(o, st));
  }
  return HOP_OK;
}

// ------------------------------------------------------------------ dedicated pack pass
// x [-> LayerNorm] [/ *range] -> fp16 hi [+ lo] packs, row-major and / or transposed, as ONE streaming pass: 64 rows per
// CTA, 64-column chunks through a shared-memory tile for the transposed layout (each column's 64 rows leave as one
// 128-byte segment).  ln_linear_kernel does the same through its GEMM tiling at ~1.3 TB/s; this is the copy it should be.
struct PackArgs {
  const float* x; int rows, rows_pad, width;
  int has_ln; const float* ln_w; const float* ln_b; float ln_eps;
  const float* range_dev; float lo_scale;
  uint16_t* y16; uint16_t* y16lo; uint16_t* yt16; uint16_t* yt16lo;
  float* y;            // fp32 copy (rows, width) or nullptr
  float* norm_max;     // [1]: max over rows of ||y_row||_2 (atomic max on the float's bits; zeroed by the caller) or nullptr
};
__global__ void __launch_bounds__(NT) pack16_kernel(const PackArgs a) {
  __shared__ float tile[64][65];
  __shared__ float s_mean[64], s_rstd[64];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * 64;
  float sc = 1.f;
  if (a.range_dev) { const float r = __ldg(a.range_dev); sc = r > 0.f ? 1.f / r : 1.f; }
  if (a.has_ln) {
    for (int r = warp; r < 64; r += NT / 32) {
      float mean = 0.f, rstd = 1.f;
      if (row0 + r < a.rows) warp_row_stats(a.x + (size_t)(row0 + r) * a.width, a.width, a.ln_eps, mean, rstd);
      if (lane == 0) { s_mean[r] = mean; s_rstd[r] = rstd; }
    }
    __syncthreads();
  }
  const bool tr = a.yt16 != nullptr;
  const int c4 = (tid & 15) * 4, rq = tid >> 4;            // this thread's 4 columns of the chunk, its row within 16
  float sq[4] = {0.f, 0.f, 0.f, 0.f};                      // partial ||row||^2 of this thread's four rows
  for (int c0 = 0; c0 < a.width; c0 += 64) {
    const int c = c0 + c4;
    float4 w4 = make_float4(1.f, 1.f, 1.f, 1.f), b4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.has_ln && c < a.width) {
      if (a.ln_w) w4 = __ldg(reinterpret_cast<const float4*>(a.ln_w + c));
      if (a.ln_b) b4 = __ldg(reinterpret_cast<const float4*>(a.ln_b + c));
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = rq + 16 * i, gr = row0 + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gr < a.rows && c < a.width) {
        v = __ldg(reinterpret_cast<const float4*>(a.x + (size_t)gr * a.width + c));
        if (a.has_ln) {
          const float m = s_mean[r], rs = s_rstd[r];
          v.x = (v.x - m) * rs * w4.x + b4.x; v.y = (v.y - m) * rs * w4.y + b4.y;
          v.z = (v.z - m) * rs * w4.z + b4.z; v.w = (v.w - m) * rs * w4.w + b4.w;
        }
        v.x *= sc; v.y *= sc; v.z *= sc; v.w *= sc;
        if (a.y) *reinterpret_cast<float4*>(a.y + (size_t)gr * a.width + c) = v;
        sq[i] += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
      }
      if (gr < a.rows_pad && c < a.width) {
        const uint16_t h0 = to16(v.x, 0), h1 = to16(v.y, 0), h2 = to16(v.z, 0), h3 = to16(v.w, 0);
        if (a.y16)
          *reinterpret_cast<uint2*>(a.y16 + (size_t)gr * a.width + c) = make_uint2(h0 | ((uint32_t)h1 << 16), h2 | ((uint32_t)h3 << 16));
        if (a.y16lo) {
          const uint16_t l0 = to16((v.x - from16(h0, 0)) * a.lo_scale, 0), l1 = to16((v.y - from16(h1, 0)) * a.lo_scale, 0);
          const uint16_t l2 = to16((v.z - from16(h2, 0)) * a.lo_scale, 0), l3 = to16((v.w - from16(h3, 0)) * a.lo_scale, 0);
          *reinterpret_cast<uint2*>(a.y16lo + (size_t)gr * a.width + c) = make_uint2(l0 | ((uint32_t)l1 << 16), l2 | ((uint32_t)l3 << 16));
        }
      }
      if (tr) { tile[r][c4] = v.x; tile[r][c4 + 1] = v.y; tile[r][c4 + 2] = v.z; tile[r][c4 + 3] = v.w; }
    }
    if (tr) {
      __syncthreads();
      // column (tid >> 2) of the chunk, rows 16 (tid & 3) .. +15: 32 contiguous bytes per thread, 128 per column
      const int col = tid >> 2, rs0 = (tid & 3) * 16, gc = c0 + col;
      if (gc < a.width && row0 + rs0 < a.rows_pad) {
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float v0 = tile[rs0 + 2 * e][col], v1 = tile[rs0 + 2 * e + 1][col];
          const uint16_t h0 = to16(v0, 0), h1 = to16(v1, 0);
          hi[e] = h0 | ((uint32_t)h1 << 16);
          lo[e] = to16((v0 - from16(h0, 0)) * a.lo_scale, 0) | ((uint32_t)to16((v1 - from16(h1, 0)) * a.lo_scale, 0) << 16);
        }
        uint4* dst = reinterpret_cast<uint4*>(a.yt16 + (size_t)gc * a.rows_pad + row0 + rs0);
        dst[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]); dst[1] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
        if (a.yt16lo) {
          uint4* dl = reinterpret_cast<uint4*>(a.yt16lo + (size_t)gc * a.rows_pad + row0 + rs0);
          dl[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]); dl[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
        }
      }
      __syncthreads();
    }
  }
  if (a.norm_max) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float v = sq[i];
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);     // the 16 threads of a row
      if ((tid & 15) == 0 && v > 0.f) atomicMax(reinterpret_cast<int*>(a.norm_max), __float_as_int(sqrtf(v)));
    }
  }
}
// applies when the layouts allow the vector accesses above; else the caller takes ln_linear_kernel
static bool pack16_fast_ok(const Pack16Args& p) {
  auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  return p.width % 4 == 0 && p.rows_pad % 64 == 0 && p.rows_pad > 0 && al(p.x) && al(p.ln_w) && al(p.ln_b) && al(p.y16) &&
         al(p.y16lo) && al(p.yt16) && al(p.yt16lo);
}
static int launch_pack16_fast(const PackArgs& a, cudaStream_t st) {
  pack16_kernel<<<a.rows_pad / 64, NT, 0, st>>>(a);
  HOP_LAUNCH_CHECK("pack16_kernel");
  return HOP_OK;
}

int launch_ln_pack(const LnLinearArgs& a, cudaStream_t st) {
  auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  const bool fast = a.w == nullptr && !a.bf16 && a.scale == 1.f && a.dout == a.din && a.din % 4 == 0 && a.rows_pad % 64 == 0 &&
                    a.rows_pad >= a.rows && (a.y16 || a.yt16 || a.y16lo) && al(a.x) && al(a.ln_w) && al(a.ln_b) && al(a.y) &&
                    al(a.y16) && al(a.y16lo) && al(a.yt16) && al(a.yt16lo) && (a.yt16 || !a.yt16lo);
  if (!fast) return launch_ln_linear(a, st);
  PackArgs f{a.x, a.rows, a.rows_pad, a.din, a.has_ln, a.ln_w, a.ln_b, a.ln_eps, a.range_dev, a.lo_scale > 0.f ? a.lo_scale : 1.f,
             static_cast<uint16_t*>(a.y16), static_cast<uint16_t*>(a.y16lo), static_cast<uint16_t*>(a.yt16),
             static_cast<uint16_t*>(a.yt16lo), a.y, a.norm_max};
  return launch_pack16_fast(f, st);
}

int launch_pack16(const Pack16Args& p, cudaStream_t st) {
  if (pack16_fast_ok(p)) {
    PackArgs f{p.x, p.rows, p.rows_pad, p.width, p.has_ln, p.ln_w, p.ln_b, p.ln_eps, p.range_dev, kGemmLoScale,
               static_cast<uint16_t*>(p.y16), static_cast<uint16_t*>(p.y16lo), static_cast<uint16_t*>(p.yt16),
               static_cast<uint16_t*>(p.yt16lo), nullptr, nullptr};
    return launch_pack16_fast(f, st);
  }
  LnLinearArgs a{};
  a.x = p.x; a.rows = p.rows; a.rows_pad = p.rows_pad; a.din = p.width; a.has_ln = p.has_ln; a.ln_w = p.ln_w; a.ln_b = p.ln_b;
  a.ln_eps = p.ln_eps; a.w = nullptr; a.b = nullptr; a.dout = p.width; a.scale = 1.f; a.y = nullptr;
  a.y16 = p.y16; a.y16lo = p.y16lo; a.yt16 = p.yt16; a.yt16lo = p.yt16lo; a.bf16 = 0; a.lo_scale = kGemmLoScale;
  a.range_dev = p.range_dev;
  return launch_ln_linear(a, st);
}

int launch_prepare_queries(const hop_desc& d, const hop_params& p, const float* x, float* qhat, cudaStream_t st) {
  LnLinearArgs q{};
  q.x = x; q.rows = d.n_rows; q.rows_pad = d.n_rows; q.din = d.d_in;
  q.has_ln = (d.flags & HOP_F_NORM_STATE) ? 1 : 0; q.ln_w = p.ln_state_w; q.ln_b = p.ln_state_b; q.ln_eps = d.ln_eps;
  q.w = (d.flags & HOP_F_PROJ_STATE) ? p.wq : nullptr; q.b = (d.flags & HOP_F_PROJ_STATE) ? p.bq : nullptr;
  q.dout = d.d_assoc; q.scale = d.beta; q.y = qhat;
  if ((d.flags & HOP_F_PROJ_STATE) && !p.wq) HOP_FAIL(HOP_ERR_BAD_ARG, "HOP_F_PROJ_STATE set but params.wq is NULL");
  return launch_ln_linear(q, st);
}

int launch_linear(int n, int din, int dout, const float* x, const float* w, const float* b, float* y,
                  cudaStream_t st) {
  LnLinearArgs l{};
  l.x = x; l.rows = n; l.rows_pad = n; l.din = din; l.has_ln = 0; l.ln_eps = 0.f;
  l.w = w; l.b = b; l.dout = dout; l.scale = 1.f; l.y = y;
  return launch_ln_linear(l, st);
}

}  // namespace hop
