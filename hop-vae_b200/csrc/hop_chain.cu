// Hop-VAE's bottleneck as ONE launch: the three Hopfield retrievals of HopVAE.forward and the glue
// between them,
//
//     emb    = hopfield(z)                         HopVAE.py:255   (64 -> 64, P stored patterns)
//     logits = embedding_to_index(emb)             HopVAE.py:257   (64 -> 3)
//     idx    = round(sigmoid(logits) * L) / L      HopVAE.py:264-267, utils/__init__.py:51-55 (forward of the STE round)
//     recon  = index_to_embedding(idx)             HopVAE.py:271   (3 -> 64)
//     loss   = mse_loss(recon, emb)                HopVAE.py:272
//
// for the reference's sizes (P = 512 patterns per layer, 2k .. 9k latent rows per batch), where the three layers are
// launch- and latency-bound as separate kernels (88 / 105 / 152 us per bottleneck under a CUDA graph at configs 1 / 2 / 3,
// profiles/r01_model_configs.jsonl).  A cluster of S CTAs (256 threads each) owns BM = 64 latent rows and walks them
// through all three layers: LN_q + beta in place, the flash-style association against 64-pattern tiles of that layer's
// K^ / V^ (cp.async double buffered, exact fp32 FFMA, the arithmetic of hop_fwd_simt.cu), the out-projection, the glue --
// the row tile never leaves shared memory between layers.  2,048 rows are only 32 such tiles, so inside a layer the S
// CTAs of a cluster split the PATTERN axis: each runs its share of the tiles on the same rows, leaves its un-normalised
// partial (O, max, sum) in its own shared memory, and after a cluster barrier every CTA merges all partials out of its
// peers' shared memory (DSMEM) in rank order -- so all S CTAs hold the identical layer output and go on to the next
// layer without exchanging anything else.  (A 16-row tile per CTA instead -- first version -- is shared-memory bound:
// one K^ load per four FMAs, 100 us at 2,048 rows; 4 x 4 register tiles need 64 rows.)  The squared error is reduced per
// cluster by its rank-0 CTA and the last one to finish adds the partials in tile order (deterministic) and writes the mean.
//
// Inference / evaluation only (no autograd state is saved); every layer is the Hop-VAE shape: static state and stored
// patterns (H = D_in), pattern projection D_in -> D_in, out-projection.  Widths up to 64.
#include <cooperative_groups.h>

#include <mutex>

#include "hop_common.cuh"

namespace cg = cooperative_groups;

namespace hop {

namespace {

constexpr int NT = 256;
constexpr int BP = 64;           // patterns per tile
constexpr int HS = 68;           // row stride of the 64-wide tiles (floats)
constexpr int DS = 64;           // row stride of the read-out tile

struct ChainLayer {
  const float* khat; const float* vhat;          // (P, h), (P, dv) fp32 (hop_sm100_prepare_patterns)
  const float* lnq_w; const float* lnq_b; int has_ln; float ln_eps; float beta;
  const float* wo; const float* bo;              // (dout, dv), (dout)
  int p, h, dv, dout;
  float* out;                                    // (N, dout) or nullptr
};
struct ChainArgs {
  ChainLayer layer[3];
  const float* z; int n;
  float levels;                                  // num_levels - 1
  float* idx_q;                                  // (N, index_dim) quantised levels (what the prior is fitted on) or nullptr
  float* partial; uint32_t* done; float* mse;    // per-CTA squared-error sums, completion counter, mean squared error
};

__device__ __forceinline__ void cp16(float* dst_smem, const float* src, bool valid) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst_smem);
  const int bytes = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// rows [r0, r0 + rows) x [0, w) of a (total, w) fp32 matrix -> smem with row stride `stride`, zero-filled up to wpad
__device__ __forceinline__ void tile_in(float* dst, const float* src, int r0, int rows, int total, int w, int wpad,
                                        int stride) {
  if ((w & 3) == 0) {
    const int w4 = wpad >> 2;
    for (int idx = threadIdx.x; idx < rows * w4; idx += NT) {
      const int r = idx / w4, c = (idx - r * w4) << 2;
      const bool valid = r0 + r < total && c < w;
      cp16(dst + r * stride + c, valid ? src + (size_t)(r0 + r) * w + c : src, valid);
    }
  } else {
    for (int idx = threadIdx.x; idx < rows * wpad; idx += NT) {
      const int r = idx / wpad, c = idx - r * wpad;
      dst[r * stride + c] = (r0 + r < total && c < w) ? __ldg(src + (size_t)(r0 + r) * w + c) : 0.f;
    }
  }
}

// One retrieval of the row tile in X (BM x HS, this layer's raw input in columns [0, h)); leaves the layer output in
// X (columns [0, dout), zeros up to 64) and, when asked, in global memory.  HF = association width padded to 4.
template <int RM, int HF>
__device__ __forceinline__ void chain_layer(const ChainLayer& L, float* X, float* Os, float* Ps, float* St, float* K0,
                                            float* B0, int row0, int n, int rank, int splits, bool writer) {
  constexpr int BM = 16 * RM;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4, warp = tid >> 5, lane = tid & 31;
  const int dv4 = (L.dv + 3) & ~3;
  const int n_tiles_all = (L.p + BP - 1) / BP;
  const int per = (n_tiles_all + splits - 1) / splits;
  const int t_begin = min(rank * per, n_tiles_all), t_end = min(t_begin + per, n_tiles_all);
  const int n_tiles = t_end - t_begin;        // this CTA's share of the pattern tiles (local index 0 .. n_tiles)
  // one K^ tile buffer (tile t+1 is fetched behind the scores of tile t, under its read-out), two read-out buffers
  auto issue_k = [&](int t) { tile_in(K0, L.khat, (t_begin + t) * BP, BP, L.p, L.h, HF, HS); cp_commit(); };
  auto issue_b = [&](int t) {
    tile_in(B0 + (t & 1) * BP * DS, L.vhat, (t_begin + t) * BP, BP, L.p, L.dv, DS, DS);
    cp_commit();
  };
  if (n_tiles > 0) { issue_k(0); issue_b(0); }   // overlap the LayerNorm below
  // LN_q + beta in place (a warp per row)
  for (int r = warp; r < BM; r += NT / 32) {
    float* qr = X + r * HS;
    float mean = 0.f, rstd = 1.f;
    if (L.has_ln) warp_row_stats(qr, L.h, L.ln_eps, mean, rstd);
    for (int i = lane; i < HF; i += 32) {
      float v = 0.f;
      if (i < L.h) {
        v = qr[i];
        if (L.has_ln) {
          v = (v - mean) * rstd;
          if (L.lnq_w) v *= __ldg(L.lnq_w + i);
          if (L.lnq_b) v += __ldg(L.lnq_b + i);
        }
        v *= L.beta;
      }
      qr[i] = v;
    }
  }

  float m_run[RM], l_run[RM], o[RM][4];
#pragma unroll
  for (int i = 0; i < RM; ++i) {
    m_run[i] = -INFINITY; l_run[i] = 0.f;
#pragma unroll
    for (int e = 0; e < 4; ++e) o[i][e] = 0.f;
  }
  for (int t = 0; t < n_tiles; ++t) {
    __syncthreads();                          // the other read-out buffer's readers are done; X is normalised
    if (t + 1 < n_tiles) { issue_b(t + 1); cp_wait<1>(); } else { cp_wait<0>(); }   // K(t), B(t) have landed
    const float* Ks = K0;
    const float* Bs = B0 + (t & 1) * BP * DS;
    __syncthreads();
    float s[RM][4];
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int b = 0; b < 4; ++b) s[i][b] = 0.f;
#pragma unroll(HF >= 16 ? 4 : 1)
    for (int k = 0; k < HF; k += 4) {
      float4 q[RM];
#pragma unroll
      for (int i = 0; i < RM; ++i) q[i] = *reinterpret_cast<const float4*>(X + (ty * RM + i) * HS + k);
      float4 kk[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) kk[b] = *reinterpret_cast<const float4*>(Ks + (tx + 16 * b) * HS + k);
      // sixteen independent accumulators between two FMAs on the same one
#pragma unroll
      for (int b = 0; b < 4; ++b)
#pragma unroll
        for (int i = 0; i < RM; ++i) s[i][b] = fmaf(q[i].x, kk[b].x, s[i][b]);
#pragma unroll
      for (int b = 0; b < 4; ++b)
#pragma unroll
        for (int i = 0; i < RM; ++i) s[i][b] = fmaf(q[i].y, kk[b].y, s[i][b]);
#pragma unroll
      for (int b = 0; b < 4; ++b)
#pragma unroll
        for (int i = 0; i < RM; ++i) s[i][b] = fmaf(q[i].z, kk[b].z, s[i][b]);
#pragma unroll
      for (int b = 0; b < 4; ++b)
#pragma unroll
        for (int i = 0; i < RM; ++i) s[i][b] = fmaf(q[i].w, kk[b].w, s[i][b]);
    }
    const int p0 = (t_begin + t) * BP;
#pragma unroll
    for (int i = 0; i < RM; ++i) {
      float tmax = -INFINITY;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        if (p0 + tx + 16 * b >= L.p) s[i][b] = -INFINITY;
        tmax = fmaxf(tmax, s[i][b]);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, off));
      const float m_new = fmaxf(m_run[i], tmax);
      const float corr = __expf(m_run[i] - m_new);
      float psum = 0.f;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const float pr = __expf(s[i][b] - m_new);
        psum += pr;
        Ps[(ty * RM + i) * HS + tx + 16 * b] = pr;
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) psum += __shfl_xor_sync(0xffffffffu, psum, off);
      l_run[i] = l_run[i] * corr + psum;
      m_run[i] = m_new;
#pragma unroll
      for (int e = 0; e < 4; ++e) o[i][e] *= corr;
    }
    __syncthreads();                          // scores of tile t are done: the K buffer is free
    if (t + 1 < n_tiles) issue_k(t + 1);
    if (HF == 4 && dv4 == 4) {
      // index_to_embedding reads out <= 4 columns: the 16 threads of a row group split the 64 PATTERNS of the tile
      // (4 each) instead of the columns and keep private partial sums, reduced across the group once per layer
      float4 pv[RM];
#pragma unroll
      for (int i = 0; i < RM; ++i) pv[i] = *reinterpret_cast<const float4*>(Ps + (ty * RM + i) * HS + 4 * tx);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const float4 v = *reinterpret_cast<const float4*>(Bs + (4 * tx + jj) * DS);
#pragma unroll
        for (int i = 0; i < RM; ++i) {
          const float pr = jj == 0 ? pv[i].x : jj == 1 ? pv[i].y : jj == 2 ? pv[i].z : pv[i].w;
          o[i][0] = fmaf(pr, v.x, o[i][0]);
          o[i][1] = fmaf(pr, v.y, o[i][1]);
          o[i][2] = fmaf(pr, v.z, o[i][2]);
          o[i][3] = fmaf(pr, v.w, o[i][3]);
        }
      }
    } else if (4 * tx < dv4) {
      for (int j = 0; j < BP; j += 4) {
        float4 pv[RM];
#pragma unroll
        for (int i = 0; i < RM; ++i) pv[i] = *reinterpret_cast<const float4*>(Ps + (ty * RM + i) * HS + j);
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const float4 v = *reinterpret_cast<const float4*>(Bs + (j + jj) * DS + 4 * tx);
#pragma unroll
          for (int i = 0; i < RM; ++i) {
            const float pr = jj == 0 ? pv[i].x : jj == 1 ? pv[i].y : jj == 2 ? pv[i].z : pv[i].w;
            o[i][0] = fmaf(pr, v.x, o[i][0]);
            o[i][1] = fmaf(pr, v.y, o[i][1]);
            o[i][2] = fmaf(pr, v.z, o[i][2]);
            o[i][3] = fmaf(pr, v.w, o[i][3]);
          }
        }
      }
    }
  }
  if (HF == 4 && dv4 == 4) {                  // pattern-split read-out: sum the group's partials, keep them in tx == 0
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = o[i][e];
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
        o[i][e] = tx == 0 ? v : 0.f;
      }
  }
  // un-normalised partial (O, max, sum) of this CTA's pattern range -> own shared memory
  __syncthreads();                            // (an empty range never entered the loop: X readers / LN writers meet here)
#pragma unroll
  for (int i = 0; i < RM; ++i) {
    const int r = ty * RM + i;
    *reinterpret_cast<float4*>(Os + r * HS + 4 * tx) = make_float4(o[i][0], o[i][1], o[i][2], o[i][3]);
    if (tx == 0) { St[r] = m_run[i]; St[BM + r] = l_run[i]; }
  }
  cg::cluster_group cluster = cg::this_cluster();
  cluster.sync();
  // Wo -> the first K buffer (both K buffers are free: the last tile's scores are done); merge of all partials in rank
  // order -> Ps (free as well), the same in every CTA of the cluster
  for (int idx = tid; idx < BP * DS; idx += NT) {
    const int oc = idx >> 6, j = idx & 63;
    K0[oc * HS + j] = (oc < L.dout && j < L.dv) ? __ldg(L.wo + (size_t)oc * L.dv + j) : 0.f;
  }
  // (every DSMEM load of a phase is issued before the first use: a chain of dependent remote loads costs ~1 us each)
  {
    constexpr int kMaxSplits = 8;
    const float* pst[kMaxSplits];
    const float* pos[kMaxSplits];
#pragma unroll
    for (int sp = 0; sp < kMaxSplits; ++sp) {
      const int src = sp < splits ? sp : rank;                   // (unused slots alias this CTA: weight 0 below)
      pst[sp] = cluster.map_shared_rank(St, src);
      pos[sp] = cluster.map_shared_rank(Os, src);
    }
    float ms[RM][kMaxSplits], ls[RM][kMaxSplits];
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int sp = 0; sp < kMaxSplits; ++sp) {
        const int r = ty * RM + i;
        ms[i][sp] = sp < splits ? pst[sp][r] : -INFINITY;
        ls[i][sp] = sp < splits ? pst[sp][BM + r] : 0.f;
      }
#pragma unroll
    for (int i = 0; i < RM; ++i) {
      const int r = ty * RM + i;
      float M = -INFINITY;
#pragma unroll
      for (int sp = 0; sp < kMaxSplits; ++sp) M = fmaxf(M, ms[i][sp]);
      float4 v[kMaxSplits];
#pragma unroll
      for (int sp = 0; sp < kMaxSplits; ++sp)
        v[sp] = sp < splits ? *reinterpret_cast<const float4*>(pos[sp] + r * HS + 4 * tx) : make_float4(0.f, 0.f, 0.f, 0.f);
      float Lsum = 0.f;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int sp = 0; sp < kMaxSplits; ++sp) {                  // rank order: the same sum in every CTA
        const float w = __expf(ms[i][sp] - M);                   // exp(-inf) = 0 for an empty range / unused slot
        Lsum = fmaf(w, ls[i][sp], Lsum);
        acc.x = fmaf(w, v[sp].x, acc.x); acc.y = fmaf(w, v[sp].y, acc.y);
        acc.z = fmaf(w, v[sp].z, acc.z); acc.w = fmaf(w, v[sp].w, acc.w);
      }
      const float inv = 1.f / Lsum;
      *reinterpret_cast<float4*>(Ps + r * HS + 4 * tx) = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
    }
  }
  cluster.sync();                             // peers are done with this CTA's partials; Ps and Wo are complete
  float acc[RM][4];
#pragma unroll
  for (int i = 0; i < RM; ++i)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[i][b] = 0.f;
  for (int k = 0; k < dv4; k += 4) {
    float4 ov[RM];
#pragma unroll
    for (int i = 0; i < RM; ++i) ov[i] = *reinterpret_cast<const float4*>(Ps + (ty * RM + i) * HS + k);
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const float4 ww = *reinterpret_cast<const float4*>(K0 + (tx + 16 * b) * HS + k);
#pragma unroll
      for (int i = 0; i < RM; ++i) {
        acc[i][b] = fmaf(ov[i].x, ww.x, acc[i][b]);
        acc[i][b] = fmaf(ov[i].y, ww.y, acc[i][b]);
        acc[i][b] = fmaf(ov[i].z, ww.z, acc[i][b]);
        acc[i][b] = fmaf(ov[i].w, ww.w, acc[i][b]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < RM; ++i)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int oc = tx + 16 * b;
      X[(ty * RM + i) * HS + oc] = oc < L.dout ? acc[i][b] + (L.bo ? __ldg(L.bo + oc) : 0.f) : 0.f;
    }
  __syncthreads();
  if (L.out != nullptr && writer) {
    for (int idx = tid; idx < BM * L.dout; idx += NT) {
      const int r = idx / L.dout, c = idx - r * L.dout;
      if (row0 + r < n) L.out[(size_t)(row0 + r) * L.dout + c] = X[r * HS + c];
    }
  }
  __syncthreads();                            // the caller rewrites X next
}

template <int RM>
__global__ void __launch_bounds__(NT, 2) hop_chain_kernel(const __grid_constant__ ChainArgs a, int splits) {
  extern __shared__ __align__(16) float smem[];
  constexpr int BM = 16 * RM;
  float* X = smem;                    // the row tile: input / output of the current layer
  float* E = X + BM * HS;             // emb, kept for the squared error
  float* Ps = E + BM * HS;            // probability tile; merged read-out behind the tile loop
  float* St = Ps + BM * HS;           // this CTA's partial (max, sum) per row
  float* K0 = St + 2 * BM;            // one K^ tile
  float* B0 = K0 + BP * HS;           // two read-out tiles
  float* Os = B0;                     // this CTA's partial read-out (read by the whole cluster), behind the tile loop
  static_assert(BM * HS <= 2 * BP * DS, "the partial read-out must fit the read-out tile buffers");
  __shared__ float red[NT / 32];
  __shared__ bool is_last;
  const int tid = threadIdx.x;
  const int rank = (int)cg::this_cluster().block_rank();
  const int tile = blockIdx.x / splits, n_row_tiles = gridDim.x / splits;
  const int row0 = tile * BM;
  const bool writer = rank == 0;      // every CTA of a cluster holds the same results; one of them stores them
  const int d0 = a.layer[0].h;

  tile_in(X, a.z, row0, BM, a.n, d0, 64, HS);
  cp_commit();
  cp_wait<0>();
  __syncthreads();
#pragma unroll 1
  for (int l = 0; l < 2; ++l) {       // hopfield, embedding_to_index: one copy of the 64-wide layer code
    chain_layer<RM, 64>(a.layer[l], X, Os, Ps, St, K0, B0, row0, a.n, rank, splits, writer);
    if (l == 0) {
      for (int idx = tid; idx < BM * 64; idx += NT) E[(idx >> 6) * HS + (idx & 63)] = X[(idx >> 6) * HS + (idx & 63)];
      __syncthreads();
    }
  }
  // sigmoid -> round to the level grid (forward of utils.straight_through_round) -> back to [0, 1]
  const int di = a.layer[1].dout;
  for (int idx = tid; idx < BM * di; idx += NT) {
    const int r = idx / di, c = idx - r * di;
    const float sg = 1.f / (1.f + expf(-X[r * HS + c]));
    const float q = rintf(sg * a.levels);
    if (a.idx_q != nullptr && writer && row0 + r < a.n) a.idx_q[(size_t)(row0 + r) * di + c] = q;
    X[r * HS + c] = q / a.levels;
  }
  __syncthreads();
  chain_layer<RM, 4>(a.layer[2], X, Os, Ps, St, K0, B0, row0, a.n, rank, splits, writer);       // index_to_embedding
  if (!writer) return;
  // squared error of this tile's valid rows
  const int de = a.layer[2].dout;
  float sq = 0.f;
  for (int idx = tid; idx < BM * de; idx += NT) {
    const int r = idx / de, c = idx - r * de;
    if (row0 + r < a.n) {
      const float df = X[r * HS + c] - E[r * HS + c];
      sq = fmaf(df, df, sq);
    }
  }
  sq = warp_sum(sq);
  if ((tid & 31) == 0) red[tid >> 5] = sq;
  __syncthreads();
  if (tid == 0) {
    float tot = 0.f;
    for (int w = 0; w < NT / 32; ++w) tot += red[w];
    a.partial[tile] = tot;
    __threadfence();
    is_last = atomicAdd(a.done, 1u) == (unsigned)n_row_tiles - 1;
  }
  __syncthreads();
  if (is_last && tid < 32) {
    __threadfence();
    float tot = 0.f;                  // tile order, fixed lane assignment: the same sum on every run
    for (int i = tid; i < n_row_tiles; i += 32) tot += __ldcg(a.partial + i);
    tot = warp_sum(tot);
    if (tid == 0) {
      *a.mse = tot / ((float)a.n * (float)de);
      *a.done = 0u;                   // ready for the next launch on this workspace
    }
  }
}

constexpr int kChainRM = 4;           // 64 rows per cluster: 4 x 4 register tiles

// Pattern splits per row tile.  Measured (B200, graph replays, `scripts/probes/chain_time.py`): 2,048 rows (32 tiles)
// 166 / 111 / 86 / 119 us with 1 / 2 / 4 / 8 splits, 2,592 rows (41 tiles) 167 / 111 / 112 / 218, 9,248 rows (145 tiles)
// 170 / 158 / 291 / 536: fill the SMs once -- a second co-resident CTA (103 KB of shared memory, 128 registers x 256
// threads each) helps a little, a second wave or more merge partners than that do not.
int chain_splits(int n) {
  const int tiles = (n + 16 * kChainRM - 1) / (16 * kChainRM);
  const int sms = num_sms();
  if (2 * tiles >= sms) return tiles * 2 <= 2 * sms ? 2 : 1;
  int s = 1;
  while (s < 8 && tiles * s * 2 <= sms + sms / 8) s *= 2;
  return s;
}

int launch_chain_inst(const ChainArgs& a, int tune_splits, cudaStream_t st) {
  constexpr int BM = 16 * kChainRM;
  const size_t smem = (size_t)(3 * BM * HS + 2 * BM + BP * HS + 2 * BP * DS) * sizeof(float);
  auto* kern = hop_chain_kernel<kChainRM>;
  HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // two CTAs per SM need 2 x 103 KB: ask for the full shared-memory carve-out (the default keeps more L1 and fits one)
  HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  const int splits = (tune_splits == 1 || tune_splits == 2 || tune_splits == 4 || tune_splits == 8) ? tune_splits
                                                                                                  : chain_splits(a.n);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)((a.n + BM - 1) / BM * splits));
  cfg.blockDim = dim3(NT);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)splits; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  HOP_CUDA(cudaLaunchKernelEx(&cfg, kern, a, splits));
  HOP_LAUNCH_CHECK("hop_chain_kernel");
  return HOP_OK;
}

}  // namespace

size_t chain_workspace_bytes(int n) { return 256 + round_up_sz((size_t)((n + 15) / 16) * sizeof(float), 256); }

int chain_supported(const hop_desc* d) {
  // the Hop-VAE bottleneck: width w -> w -> small -> w, static q / k, projected v, out-projection, no retrieve updates
  const unsigned need = HOP_F_PROJ_PATTERN | HOP_F_OUT_PROJ;
  for (int l = 0; l < 3; ++l) {
    if ((d[l].flags & need) != need || (d[l].flags & (HOP_F_PROJ_STATE | HOP_F_PROJ_STORED))) return 0;
    if (d[l].update_steps != 0 || d[l].n_rows != d[0].n_rows || d[l].pattern_offset != 0) return 0;
    if (d[l].d_assoc != d[l].d_in || d[l].d_v > 64 || d[l].d_out > 64 || d[l].d_in > 64) return 0;
  }
  if (d[0].d_in > 64 || d[1].d_in != d[0].d_out || d[2].d_in != d[1].d_out || d[2].d_out != d[0].d_out) return 0;
  if (d[2].d_in > 4) return 0;
  return 1;
}

int launch_chain(const hop_desc* d, const hop_params* p, const hop_patterns* pt, float levels, const float* z,
                 float* emb, float* logits, float* idx_q, float* emb_recon, float* mse, void* ws, size_t ws_bytes,
                 cudaStream_t st) {
  if (!chain_supported(d)) HOP_FAIL(HOP_ERR_UNSUPPORTED, "bottleneck chain: the three layers are not the Hop-VAE shape "
                                    "(static q / k, widths <= 64, index width <= 4, no retrieve updates)");
  const int n = d[0].n_rows;
  const size_t need = chain_workspace_bytes(n);
  if (!ws || ws_bytes < need) HOP_FAIL(HOP_ERR_WORKSPACE, "bottleneck chain needs %zu workspace bytes, %zu given", need, ws_bytes);
  if (!z || !mse) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck chain: z / mse is NULL");
  ChainArgs a{};
  float* outs[3] = {emb, logits, emb_recon};
  for (int l = 0; l < 3; ++l) {
    if (!pt[l].khat || !pt[l].vhat || !p[l].wo) HOP_FAIL(HOP_ERR_BAD_ARG, "bottleneck chain: layer %d lacks K^ / V^ / Wo", l);
    ChainLayer& L = a.layer[l];
    L.khat = pt[l].khat; L.vhat = pt[l].vhat;
    L.has_ln = (d[l].flags & HOP_F_NORM_STATE) ? 1 : 0; L.lnq_w = p[l].ln_state_w; L.lnq_b = p[l].ln_state_b;
    L.ln_eps = d[l].ln_eps; L.beta = d[l].beta; L.wo = p[l].wo; L.bo = p[l].bo;
    L.p = d[l].n_patterns; L.h = d[l].d_assoc; L.dv = d[l].d_v; L.dout = d[l].d_out; L.out = outs[l];
  }
  a.z = z; a.n = n; a.levels = levels; a.idx_q = idx_q;
  a.done = static_cast<uint32_t*>(ws);
  a.partial = reinterpret_cast<float*>(static_cast<char*>(ws) + 256);
  a.mse = mse;
  HOP_CUDA(cudaMemsetAsync(a.done, 0, sizeof(uint32_t), st));
  return launch_chain_inst(a, d[0].tune_splits, st);      // (hop_desc.tune_splits of the first layer: experiments)
}

// ------------------------------------------------------------------ allocation-free variant
// hop_sm100_bottleneck_forward_static: the same launch with its workspace (completion counter + per-tile squared-error
// partials) in a file-scope device array instead of caller memory -- for callers that cannot allocate (CUDA graph
// capture, tight serving loops).  Every GPU that loads the library holds the array until exit; every call on a device
// shares it, so calls that overlap on that GPU (two streams) race on the counter and the partials.  The launch clears
// the counter itself (launch_chain's 4-byte memset) and every partial slot it reads is written first, so nothing else
// has to start at zero.
constexpr int kStaticMaxRows = 262144;                       // 28 x config 3's 9,248 latent rows
constexpr size_t kStaticScratchBytes = 256 + (size_t)(kStaticMaxRows / 16) * sizeof(float);     // = chain_workspace_bytes(max)
static __device__ __align__(256) unsigned char g_chain_scratch[kStaticScratchBytes];

int chain_static_max_rows() { return kStaticMaxRows; }

int launch_chain_static(const hop_desc* d, const hop_params* p, const hop_patterns* pt, float levels, const float* z,
                        float* emb, float* logits, float* idx_q, float* emb_recon, float* mse, cudaStream_t st) {
  if (d[0].n_rows > kStaticMaxRows)
    HOP_FAIL(HOP_ERR_BAD_SHAPE, "bottleneck (static scratch): %d rows exceed the built-in maximum of %d; use "
             "hop_sm100_bottleneck_forward with a caller workspace", d[0].n_rows, kStaticMaxRows);
  // In host code `g_chain_scratch` is only a stand-in for the symbol: its device address is looked up once per device,
  // with that device current, and that pointer is what the kernel gets.
  constexpr int kMaxDevices = 64;
  static void* addr[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  HOP_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) HOP_FAIL(HOP_ERR_CUDA, "bottleneck (static scratch): device ordinal %d out of range", dev);
  void* ptr = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (addr[dev] == nullptr) {
      void* q = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&q, g_chain_scratch);
      if (e != cudaSuccess || q == nullptr)
        HOP_FAIL(HOP_ERR_CUDA, "bottleneck (static scratch): cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      addr[dev] = q;
    }
    ptr = addr[dev];
  }
  return launch_chain(d, p, pt, levels, z, emb, logits, idx_q, emb_recon, mse, ptr, kStaticScratchBytes, st);
}

}  // namespace hop
