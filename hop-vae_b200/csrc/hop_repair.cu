// Exact nearest-pattern index behind the 16-bit tensor-core scores, and the stop rule of the retrieve updates.
//
// (1) HOP_F_EXACT_ARGMAX.  The fused tcgen05 forward (hop_fwd_tc.cu) finds each row's nearest pattern on scores whose
// operands were rounded to 16 bits: |s16 - s| <= |q| |k| 2^-10 (two relative roundings of 2^-11 per product, summed with
// Cauchy-Schwarz).  Its index equals the fp32 reference's ("nearest-pattern argmax indices identical", BASELINE.json)
// whenever the top-1 / top-2 gap exceeds tau = 2 |q| max_j|k_j| 2^-10.  The forward carries a certificate for that: next
// to each row's maximum its softmax threads keep an upper bound on the runner-up score (`second_bound` in
// hop_fwd_tc.cu: exact maxima of every tile and max chain that does not hold the running maximum, and for the 8 .. 16
// columns sharing a chain with it the log of their summed exponentials -- a bound that does NOT grow with the pattern
// count, unlike lse - rowmax, whose tail of P - 1 exponentials reaches e^-tau long before P = 10^6);  rowmax - bound > tau
// proves that EVERY other score lies more than tau below the maximum.  `certify_rows_kernel` evaluates it per row (|q_r|
// recomputed in fp32 from the raw row, mirroring the forward's LN_q prologue; where the forward ends in a merge kernel --
// every large problem -- that kernel evaluates the certificate itself from the norms the forward's prologue wrote, and this
// launch is skipped); rows that fail are appended to a compact list.  `block_check_kernel` then settles every row whose
// bound was merely loose by re-scoring the 64 patterns around the 16-bit winner in fp32 (see its comment); what is left --
// rows whose true gap is within the 16-bit noise -- `argmax_repair_kernel` re-scores against all of the fp32 K^ on CUDA
// cores (64 x 64 register-tiled FFMA, query rows rebuilt from the raw input), merging (score, index) with a packed
// 64-bit atomicMax; the last CTA to finish writes the corrected index / row maximum back.  With well-separated
// retrievals (BASELINE config 5's recipe: gap ~ 10 nats) no row is flagged and the launches cost ~5 us; with diffuse
// scores every row is flagged, the block check settles all but a few per cent, and the full pass runs on those.
//
// (2) update_eps (upstream hflayers: `update_steps_eps`).  After update t >= 1 the iteration stops when
// max over batch items of ||A_t - A_(t-1)||_F <= eps.  `delta_rows_kernel` recomputes both score rows in fp32 (two
// sweeps: log-sum-exp of both, then the squared difference of the normalised weights, element by element -- the
// difference of three O(1) sums would cancel catastrophically at eps = 1e-4), `delta_decide_kernel` reduces per item and
// sets a device flag that turns the remaining update passes into identity passes: no host synchronisation.
#include "hop_common.cuh"

namespace hop {

namespace {

constexpr int NT = 256;
constexpr int kCtrlBytes = 256;       // ctrl[0] = rows that failed the certificate, ctrl[1] = CTAs done, ctrl[2] = rows left for
                                      // the full pass (all cleared by the forward kernel)
constexpr int BP = 64;                // patterns per tile
constexpr int KC = 64;                // contraction chunk
constexpr int TS = KC + 4;            // smem stride

struct RepairWs {
  uint32_t* ctrl; int32_t* list; unsigned long long* best; int32_t* list2; size_t total;
};

RepairWs carve_repair(void* ws, int n, int h) {
  RepairWs w{};
  size_t off = 0;
  auto take = [&](size_t bytes) {
    char* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += round_up_sz(bytes, 256);
    return p;
  };
  w.ctrl = reinterpret_cast<uint32_t*>(take(kCtrlBytes));
  w.list = reinterpret_cast<int32_t*>(take((size_t)n * sizeof(int32_t)));
  w.best = reinterpret_cast<unsigned long long*>(take((size_t)n * sizeof(unsigned long long)));
  w.list2 = reinterpret_cast<int32_t*>(take((size_t)n * sizeof(int32_t)));
  (void)h;
  w.total = off;
  return w;
}

__device__ __forceinline__ uint32_t ordered_bits(float s) {
  const uint32_t u = __float_as_uint(s);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float from_ordered(uint32_t o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

// element i of the fp32 query row in the arithmetic of the forward's prologue
__device__ __forceinline__ float query_elem(const RepairArgs& a, const float* src, int i, bool ln, float mean, float rstd) {
  float v = __ldg(src + i);
  if (a.x_raw) {
    if (ln) {
      v = (v - mean) * rstd;
      if (a.lnq_w) v *= __ldg(a.lnq_w + i);
      if (a.lnq_b) v += __ldg(a.lnq_b + i);
    }
    v *= a.beta;
  }
  return v;
}

// One warp per query row: fp32 query in the forward's arithmetic, its norm, the certificate.
__global__ void __launch_bounds__(NT) certify_rows_kernel(RepairArgs a, RepairWs w) {
  const int warp = (blockIdx.x * NT + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= a.n) return;
  const long long r = warp;
  const float gap = a.rowmax[r] - fmaxf(a.rowsec[r], a.rowtop[r]);   // lower bound on the top-1 / top-2 gap of the 16-bit scores
  const float* src = (a.x_raw ? a.x_raw : a.qs) + r * a.h;
  float mean = 0.f, rstd = 1.f;
  const bool ln = a.x_raw && a.has_ln;
  if (ln) warp_row_stats(src, a.h, a.ln_eps, mean, rstd);
  float qn2 = 0.f;
  for (int i = lane; i < a.h; i += 32) {
    const float v = query_elem(a, src, i, ln, mean, rstd);
    qn2 = fmaf(v, v, qn2);
  }
  qn2 = warp_sum(qn2);
  const float tau = 2.f * sqrtf(qn2) * __ldg(a.kstats) * (1.f / 1024.f) * 1.02f + 1e-5f;
  const bool certified = gap > tau;
  if (certified) return;
  if (lane == 0) {
    const uint32_t slot = atomicAdd(w.ctrl, 1u);
    w.list[slot] = (int32_t)r;
    w.best[slot] = 0ull;
  }
}

// Block check.  A row fails the certificate when the BOUND on the other columns of its top tile is too loose (a diffuse
// retrieval: the 63 neighbours of the top column sum to more than the top term), which says nothing about its true gap.
// What the forward knows exactly is `others`, the maximum 16-bit score outside the 64-aligned pattern block B that holds the
// 16-bit winner; with |s16 - s| <= eps = |q| max|k| 2^-10 every fp32 score outside B is <= others + eps.  So one warp
// per flagged row re-scores just the 64 patterns of B in fp32 (64 x H FMAs): if the block's fp32 maximum b1 exceeds both
// the block's second value and others + eps, the fp32 nearest pattern is that column -- settled.  Only rows whose true gap
// is within the 16-bit noise go on to the full pass (list2).  On random queries against 64k patterns that is a few per
// cent of the rows instead of all of them.
__global__ void __launch_bounds__(NT) block_check_kernel(RepairArgs a, RepairWs w) {
  const int count = (int)*reinterpret_cast<volatile uint32_t*>(w.ctrl);
  const int lane = threadIdx.x & 31;
  for (int slot = (blockIdx.x * NT + threadIdx.x) >> 5; slot < count; slot += (gridDim.x * NT) >> 5) {
    const int row = w.list[slot];
    const float* src = (a.x_raw ? a.x_raw : a.qs) + (size_t)row * a.h;
    float mean = 0.f, rstd = 1.f;
    const bool ln = a.x_raw && a.has_ln;
    if (ln) warp_row_stats(src, a.h, a.ln_eps, mean, rstd);
    const int p0 = (a.argmax[row] - a.pattern_offset) & ~63;
    const int j0 = p0 + lane, j1 = p0 + lane + 32;               // this lane's two patterns of the block
    const float* k0 = a.khat + (size_t)min(j0, a.p - 1) * a.h;
    const float* k1 = a.khat + (size_t)min(j1, a.p - 1) * a.h;
    float s0 = 0.f, s1 = 0.f, qn2 = 0.f;
    for (int i = 0; i < a.h; i += 4) {
      float q[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) q[e] = query_elem(a, src, i + e, ln, mean, rstd);
      const float4 a0 = __ldg(reinterpret_cast<const float4*>(k0 + i)), a1 = __ldg(reinterpret_cast<const float4*>(k1 + i));
      s0 = fmaf(q[0], a0.x, fmaf(q[1], a0.y, fmaf(q[2], a0.z, fmaf(q[3], a0.w, s0))));
      s1 = fmaf(q[0], a1.x, fmaf(q[1], a1.y, fmaf(q[2], a1.z, fmaf(q[3], a1.w, s1))));
      qn2 = fmaf(q[0], q[0], fmaf(q[1], q[1], fmaf(q[2], q[2], fmaf(q[3], q[3], qn2))));
    }
    if (j0 >= a.p) s0 = -INFINITY;
    if (j1 >= a.p) s1 = -INFINITY;
    // block maximum (lowest index on ties), then the maximum of everything else in the block
    float b1 = s0 >= s1 ? s0 : s1;
    int i1 = s0 >= s1 ? j0 : j1;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, b1, off);
      const int oi = __shfl_xor_sync(0xffffffffu, i1, off);
      if (om > b1 || (om == b1 && oi < i1)) { b1 = om; i1 = oi; }
    }
    float b2 = fmaxf(j0 == i1 ? -INFINITY : s0, j1 == i1 ? -INFINITY : s1);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) b2 = fmaxf(b2, __shfl_xor_sync(0xffffffffu, b2, off));
    const float eps = sqrtf(qn2) * __ldg(a.kstats) * (1.f / 1024.f) * 1.02f + 5e-6f;
    const bool settled = b1 > b2 && b1 > a.rowsec[row] + eps;
    if (lane == 0) {
      if (settled) {
        a.argmax[row] = i1 + a.pattern_offset;
        a.rowmax[row] = b1;
      } else {
        const uint32_t s2 = atomicAdd(w.ctrl + 2, 1u);
        w.list2[s2] = row;
      }
    }
  }
}

// (flagged-row tile) x (pattern chunk) work items, 16*RM rows per tile, fp32 FFMA scores, running (max, index)
template <int RM>
__device__ __forceinline__ void repair_items(const RepairArgs& a, const RepairWs& w, int count, float* Qs, float* Ks,
                                             float* s_mean, float* s_rstd, int* s_row) {
  constexpr int BM = 16 * RM;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int warp = tid >> 5;
  const float* qsrc = a.x_raw ? a.x_raw : a.qs;
  const bool ln = a.x_raw && a.has_ln;
  const int chunk_tiles = 4;                                   // 256 patterns per item
  const int n_tiles = (a.p + BP - 1) / BP;
  const int n_pc = (n_tiles + chunk_tiles - 1) / chunk_tiles;
  const int n_rt = (count + BM - 1) / BM;
  const long long items = (long long)n_rt * n_pc;
  for (long long item = blockIdx.x; item < items; item += gridDim.x) {
    const int rt = (int)(item / n_pc), pc = (int)(item - (long long)rt * n_pc);
    const int slot0 = rt * BM;
    // the flagged rows of this tile and their LayerNorm statistics (the query rows are rebuilt from the raw input in
    // the forward's arithmetic: nothing but the row list travels between the kernels)
    __syncthreads();
    for (int r = warp; r < BM; r += NT / 32) {
      const int row = slot0 + r < count ? w.list2[slot0 + r] : -1;
      float mean = 0.f, rstd = 1.f;
      if (row >= 0 && ln) warp_row_stats(qsrc + (size_t)row * a.h, a.h, a.ln_eps, mean, rstd);
      if ((tid & 31) == 0) { s_row[r] = row; s_mean[r] = mean; s_rstd[r] = rstd; }
    }
    __syncthreads();
    float best[RM];
    int barg[RM];
#pragma unroll
    for (int i = 0; i < RM; ++i) { best[i] = -INFINITY; barg[i] = 0x7fffffff; }
    const int t_end = min(n_tiles, (pc + 1) * chunk_tiles);
    for (int t = pc * chunk_tiles; t < t_end; ++t) {
      const int p0 = t * BP;
      float s[RM][4];
#pragma unroll
      for (int i = 0; i < RM; ++i)
#pragma unroll
        for (int b = 0; b < 4; ++b) s[i][b] = 0.f;
      for (int k0 = 0; k0 < a.h; k0 += KC) {
        __syncthreads();
        for (int idx = tid; idx < BM * (KC / 4); idx += NT) {
          const int r = idx / (KC / 4), c = (idx - r * (KC / 4)) * 4;
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          const int row = s_row[r];
          if (row >= 0 && k0 + c < a.h) {
            const float* src = qsrc + (size_t)row * a.h;
            v.x = query_elem(a, src, k0 + c, ln, s_mean[r], s_rstd[r]);
            v.y = query_elem(a, src, k0 + c + 1, ln, s_mean[r], s_rstd[r]);
            v.z = query_elem(a, src, k0 + c + 2, ln, s_mean[r], s_rstd[r]);
            v.w = query_elem(a, src, k0 + c + 3, ln, s_mean[r], s_rstd[r]);
          }
          *reinterpret_cast<float4*>(Qs + r * TS + c) = v;
        }
        for (int idx = tid; idx < BP * (KC / 4); idx += NT) {
          const int r = idx / (KC / 4), c = (idx - r * (KC / 4)) * 4;
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (p0 + r < a.p && k0 + c < a.h) v = __ldg(reinterpret_cast<const float4*>(a.khat + (size_t)(p0 + r) * a.h + k0 + c));
          *reinterpret_cast<float4*>(Ks + r * TS + c) = v;
        }
        __syncthreads();
#pragma unroll 4
        for (int k = 0; k < KC; k += 4) {
          float4 q[RM];
#pragma unroll
          for (int i = 0; i < RM; ++i) q[i] = *reinterpret_cast<const float4*>(Qs + (ty * RM + i) * TS + k);
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const float4 kk = *reinterpret_cast<const float4*>(Ks + (tx + 16 * b) * TS + k);
#pragma unroll
            for (int i = 0; i < RM; ++i) {
              s[i][b] = fmaf(q[i].x, kk.x, s[i][b]);
              s[i][b] = fmaf(q[i].y, kk.y, s[i][b]);
              s[i][b] = fmaf(q[i].z, kk.z, s[i][b]);
              s[i][b] = fmaf(q[i].w, kk.w, s[i][b]);
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < RM; ++i)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int col = p0 + tx + 16 * b;
          if (col < a.p && s[i][b] > best[i]) { best[i] = s[i][b]; barg[i] = col; }   // strict: lowest column keeps ties
        }
    }
#pragma unroll
    for (int i = 0; i < RM; ++i) {
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, best[i], off);
        const int oa = __shfl_xor_sync(0xffffffffu, barg[i], off);
        if (om > best[i] || (om == best[i] && oa < barg[i])) { best[i] = om; barg[i] = oa; }
      }
      const int slot = slot0 + ty * RM + i;
      if (tx == 0 && slot < count && barg[i] != 0x7fffffff) {
        const unsigned long long key = ((unsigned long long)ordered_bits(best[i]) << 32) |
                                       (unsigned long long)(0xffffffffu - (uint32_t)barg[i]);
        atomicMax(w.best + slot, key);
      }
    }
  }
}

__global__ void __launch_bounds__(NT) argmax_repair_kernel(RepairArgs a, RepairWs w) {
  __shared__ __align__(16) float Qs[64 * TS];
  __shared__ __align__(16) float Ks[BP * TS];
  __shared__ float s_mean[64], s_rstd[64];
  __shared__ int s_row[64];
  __shared__ int s_last;
  const int count = (int)*reinterpret_cast<volatile uint32_t*>(w.ctrl + 2);
  if (count == 0) return;
  if (count <= 16) repair_items<1>(a, w, count, Qs, Ks, s_mean, s_rstd, s_row);
  else repair_items<4>(a, w, count, Qs, Ks, s_mean, s_rstd, s_row);
  // the last CTA to finish writes the corrected (index, row maximum) of every flagged row
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = (atomicAdd(w.ctrl + 1, 1u) == gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int slot = threadIdx.x; slot < count; slot += NT) {
    const unsigned long long key = __ldcg(w.best + slot);
    if (key == 0ull) continue;
    const int row = w.list2[slot];
    a.argmax[row] = (int32_t)(0xffffffffu - (uint32_t)(key & 0xffffffffull)) + a.pattern_offset;
    if (a.rowmax) a.rowmax[row] = from_ordered((uint32_t)(key >> 32));
  }
}

// ------------------------------------------------------------------ stop rule of the retrieve updates
// One warp per query row; lanes split the contraction, so every lane ends up with the full score.
__global__ void __launch_bounds__(NT) delta_rows_kernel(const float* __restrict__ q_new, const float* __restrict__ q_old,
                                                        const float* __restrict__ khat, int n, int p, int h,
                                                        float* __restrict__ row_out, const uint32_t* stopped) {
  if (*stopped) return;
  const int warp = (blockIdx.x * NT + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  constexpr int kMaxPerLane = 8;                     // H <= 256 (the fp32 path's limit)
  float qa[kMaxPerLane], qb[kMaxPerLane];
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) {
    const int c = lane + 32 * i;
    qa[i] = c < h ? __ldg(q_new + (size_t)warp * h + c) : 0.f;
    qb[i] = c < h ? __ldg(q_old + (size_t)warp * h + c) : 0.f;
  }
  auto scores = [&](int j, float& sa, float& sb) {
    const float* k = khat + (size_t)j * h;
    float pa = 0.f, pb = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxPerLane; ++i) {
      const int c = lane + 32 * i;
      const float kv = c < h ? __ldg(k + c) : 0.f;
      pa = fmaf(qa[i], kv, pa);
      pb = fmaf(qb[i], kv, pb);
    }
    sa = warp_sum(pa);
    sb = warp_sum(pb);
  };
  float ma = -INFINITY, mb = -INFINITY, la = 0.f, lb = 0.f;
  for (int j = 0; j < p; ++j) {
    float sa, sb;
    scores(j, sa, sb);
    const float na = fmaxf(ma, sa), nb = fmaxf(mb, sb);
    la = la * __expf(ma - na) + __expf(sa - na);
    lb = lb * __expf(mb - nb) + __expf(sb - nb);
    ma = na; mb = nb;
  }
  const float lse_a = ma + __logf(la), lse_b = mb + __logf(lb);
  float d2 = 0.f;
  for (int j = 0; j < p; ++j) {
    float sa, sb;
    scores(j, sa, sb);
    const float d = __expf(sa - lse_a) - __expf(sb - lse_b);
    d2 = fmaf(d, d, d2);
  }
  if (lane == 0) row_out[warp] = d2;
}

__global__ void __launch_bounds__(NT) delta_decide_kernel(const float* __restrict__ row_d2, int n, int rows_per_item,
                                                          float eps, uint32_t* stopped) {
  if (*stopped) return;
  __shared__ float s_max[NT / 32];
  const int items = rows_per_item > 0 ? (n + rows_per_item - 1) / rows_per_item : 1;
  const int per = rows_per_item > 0 ? rows_per_item : n;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float worst = 0.f;
  for (int it = warp; it < items; it += NT / 32) {          // one warp per batch item
    float sum = 0.f;
    const int r1 = min(n, (it + 1) * per);
    for (int r = it * per + lane; r < r1; r += 32) sum += row_d2[r];
    sum = warp_sum(sum);
    worst = fmaxf(worst, sqrtf(sum));
  }
  if (lane == 0) s_max[warp] = worst;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = 0.f;
    for (int i = 0; i < NT / 32; ++i) m = fmaxf(m, s_max[i]);
    if (!(m > eps)) *stopped = 1u;
  }
}

}  // namespace

size_t repair_workspace_bytes(int n, int h) { return carve_repair(nullptr, n, h).total; }

CertArgs repair_cert_args(void* ws, int n, int h, const float* kstats, float* qn, float* sec, float* top) {
  const RepairWs w = carve_repair(ws, n, h);
  return CertArgs{w.ctrl, w.list, w.best, kstats, qn, sec, top};
}

int launch_argmax_repair(const RepairArgs& a, cudaStream_t st) {
  if (!a.argmax || !a.rowmax || !a.khat || !a.kstats || !a.rowsec || (!a.precertified && !a.rowtop))
    HOP_FAIL(HOP_ERR_BAD_ARG, "exact argmax needs argmax / rowmax outputs, the runner-up bound, the fp32 K^ and kstats");
  const size_t need = repair_workspace_bytes(a.n, a.h);
  if (!a.ws || a.ws_bytes < need) HOP_FAIL(HOP_ERR_WORKSPACE, "argmax repair needs %zu workspace bytes, %zu given", need, a.ws_bytes);
  if ((a.h & 3) != 0) HOP_FAIL(HOP_ERR_BAD_SHAPE, "argmax repair needs H %% 4 == 0");
  const RepairWs w = carve_repair(a.ws, a.n, a.h);
  const int rows_per_block = NT / 32;
  if (!a.precertified) {
    certify_rows_kernel<<<(a.n + rows_per_block - 1) / rows_per_block, NT, 0, st>>>(a, w);
    HOP_LAUNCH_CHECK("certify_rows_kernel");
  }
  block_check_kernel<<<2 * num_sms(), NT, 0, st>>>(a, w);
  HOP_LAUNCH_CHECK("block_check_kernel");
  argmax_repair_kernel<<<2 * num_sms(), NT, 0, st>>>(a, w);
  HOP_LAUNCH_CHECK("argmax_repair_kernel");
  return HOP_OK;
}

int launch_update_delta(const float* q_new, const float* q_old, const float* khat, int n, int p, int h,
                        int rows_per_item, float eps, float* row_scratch, uint32_t* stopped, cudaStream_t st) {
  if (h > 256) HOP_FAIL(HOP_ERR_UNSUPPORTED, "stop rule: the fp32 path supports association width <= 256");
  const int rows_per_block = NT / 32;
  delta_rows_kernel<<<(n + rows_per_block - 1) / rows_per_block, NT, 0, st>>>(q_new, q_old, khat, n, p, h, row_scratch, stopped);
  HOP_LAUNCH_CHECK("delta_rows_kernel");
  delta_decide_kernel<<<1, NT, 0, st>>>(row_scratch, n, rows_per_item, eps, stopped);
  HOP_LAUNCH_CHECK("delta_decide_kernel");
  return HOP_OK;
}

}  // namespace hop
