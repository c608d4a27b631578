// Backward of the association for update_steps == 0 (SURVEY.md section 8a row
// a11: the autograd mirrors of LayerNorm x3, V-projection, score bmm, softmax,
// read-out bmm and out_proj that `loss.backward()` at main.py:35 runs).
//
// Flash-style: A = exp(S - lse) is recomputed tile by tile from the saved
// beta*Q^ and lse; no N x P tensor is ever stored.
//
//   dR   = G Wo                      (N,Dv)      [linear, transposed weight]
//   dWo  = G^T R, dbo = sum G                     [wgrad]
//   delta= rowsum(dR * R)
//   dS   = A * (dR V^T - delta)      (tile)
//   dqs  = dS K^            (N,H)                 [core_q]
//   dK^  = dS^T qs          (P,H),  dV^ = A^T dR  (P,Dv)   [core_kv, split over N, fp32 RED]
//   then LayerNorm / projection backward on the query side and (twice) on the
//   pattern side, both paths summing into d_lookup.
#include <cooperative_groups.h>

#include "hop_common.cuh"

namespace cg = cooperative_groups;

namespace hop {

namespace {

constexpr int NT = 256;
constexpr int BP = 64;
constexpr int PS = BP + 4;

__device__ __forceinline__ void load_tile_pad(float* __restrict__ dst, const float* __restrict__ src, int r0,
                                              int rows, int nrows_total, int w, int wpad, int stride) {
  if ((w & 3) == 0) {
    const int w4 = wpad >> 2;
    for (int idx = threadIdx.x; idx < rows * w4; idx += NT) {
      const int r = idx / w4, c = (idx - r * w4) << 2;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r0 + r < nrows_total && c < w) v = __ldg(reinterpret_cast<const float4*>(src + (size_t)(r0 + r) * w + c));
      *reinterpret_cast<float4*>(dst + r * stride + c) = v;
    }
  } else {
    for (int idx = threadIdx.x; idx < rows * wpad; idx += NT) {
      const int r = idx / wpad, c = idx - r * wpad;
      float v = 0.f;
      if (r0 + r < nrows_total && c < w) v = __ldg(src + (size_t)(r0 + r) * w + c);
      dst[r * stride + c] = v;
    }
  }
}

// acc[i][b] += sum_k A[(ty*RM+i)][k] * B[(tx+16b)][k]   (both k-contiguous, stride given)
template <int RM>
__device__ __forceinline__ void mma_nt(float (&acc)[RM][4], const float* __restrict__ A, int as,
                                       const float* __restrict__ B, int bs, int k4, int ty, int tx) {
  for (int k = 0; k < k4; k += 4) {
    float4 a[RM];
#pragma unroll
    for (int i = 0; i < RM; ++i) a[i] = *reinterpret_cast<const float4*>(A + (ty * RM + i) * as + k);
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const float4 bb = *reinterpret_cast<const float4*>(B + (tx + 16 * b) * bs + k);
#pragma unroll
      for (int i = 0; i < RM; ++i) {
        acc[i][b] = fmaf(a[i].x, bb.x, acc[i][b]);
        acc[i][b] = fmaf(a[i].y, bb.y, acc[i][b]);
        acc[i][b] = fmaf(a[i].z, bb.z, acc[i][b]);
        acc[i][b] = fmaf(a[i].w, bb.w, acc[i][b]);
      }
    }
  }
}

struct CoreArgs {
  const float* qs; const float* khat; const float* vhat; const float* lse;
  const float* dr; const float* r;   // dR (N,Dv), R (N,Dv)
  float* delta;                      // (N)  written by core_q, read by core_kv
  float* dqs;                        // (N,H)
  float* dkhat; float* dvhat;        // (P,H), (P,Dv) accumulated with RED (pre-zeroed)
  int n, p, h, dv;
  int rows_per_split;                // core_kv: query rows handled by one blockIdx.y
};

// ------------------------------------------------------------------ core_q
// Like the forward (hop_fwd_simt.cu): 16*RM query rows per CTA; when the row tiles alone do not fill the SMs the
// pattern axis is cut into `splits` ranges across a thread-block cluster (1 x splits x 1) and the partial dq tiles
// are summed through distributed shared memory -- no atomics, no partials in HBM.
template <int RM, int HC, int DVC>
__global__ void __launch_bounds__(NT) bwd_core_q_kernel(CoreArgs a, int splits) {
  extern __shared__ __align__(16) float smem[];
  constexpr int BM = 16 * RM;
  constexpr int HS = 64 * HC + 4, VS = 64 * DVC + 4;
  float* Qs = smem;
  float* Gs = Qs + BM * HS;
  float* Ks = Gs + BM * VS;
  float* Vs = Ks + BP * HS;
  float* Ps = Vs + BP * VS;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int row0 = blockIdx.x * BM;
  const int h4 = (a.h + 3) & ~3, dv4 = (a.dv + 3) & ~3;

  load_tile_pad(Qs, a.qs, row0, BM, a.n, a.h, 64 * HC, HS);
  load_tile_pad(Gs, a.dr, row0, BM, a.n, a.dv, 64 * DVC, VS);

  float lse_r[RM], delta_r[RM];
#pragma unroll
  for (int i = 0; i < RM; ++i) {
    const int gr = row0 + ty * RM + i;
    float part = 0.f;
    if (gr < a.n)
      for (int c = tx; c < a.dv; c += 16) part = fmaf(__ldg(a.dr + (size_t)gr * a.dv + c), __ldg(a.r + (size_t)gr * a.dv + c), part);
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
    delta_r[i] = part;
    lse_r[i] = gr < a.n ? __ldg(a.lse + gr) : 0.f;
    if (tx == 0 && gr < a.n && blockIdx.y == 0) a.delta[gr] = part;
  }

  float dq[RM][HC][4];
#pragma unroll
  for (int i = 0; i < RM; ++i)
#pragma unroll
    for (int c = 0; c < HC; ++c)
#pragma unroll
      for (int e = 0; e < 4; ++e) dq[i][c][e] = 0.f;

  const int n_tiles = (a.p + BP - 1) / BP;
  const int tiles_per_split = (n_tiles + splits - 1) / splits;
  const int t_begin = (int)blockIdx.y * tiles_per_split;
  const int t_end = min(n_tiles, t_begin + tiles_per_split);
  for (int t = t_begin; t < t_end; ++t) {
    const int p0 = t * BP;
    __syncthreads();
    load_tile_pad(Ks, a.khat, p0, BP, a.p, a.h, 64 * HC, HS);
    load_tile_pad(Vs, a.vhat, p0, BP, a.p, a.dv, 64 * DVC, VS);
    __syncthreads();
    float s[RM][4], da[RM][4];
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int b = 0; b < 4; ++b) s[i][b] = da[i][b] = 0.f;
    mma_nt<RM>(s, Qs, HS, Ks, HS, h4, ty, tx);
    mma_nt<RM>(da, Gs, VS, Vs, VS, dv4, ty, tx);
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int col = p0 + tx + 16 * b;
        const float prob = (col < a.p) ? __expf(s[i][b] - lse_r[i]) : 0.f;
        Ps[(ty * RM + i) * PS + tx + 16 * b] = prob * (da[i][b] - delta_r[i]);
      }
    __syncthreads();
    for (int j = 0; j < BP; j += 4) {
      float4 pv[RM];
#pragma unroll
      for (int i = 0; i < RM; ++i) pv[i] = *reinterpret_cast<const float4*>(Ps + (ty * RM + i) * PS + j);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj)
#pragma unroll
        for (int c = 0; c < HC; ++c) {
          const float4 kk = *reinterpret_cast<const float4*>(Ks + (j + jj) * HS + 4 * tx + 64 * c);
#pragma unroll
          for (int i = 0; i < RM; ++i) {
            const float w = jj == 0 ? pv[i].x : jj == 1 ? pv[i].y : jj == 2 ? pv[i].z : pv[i].w;
            dq[i][c][0] = fmaf(w, kk.x, dq[i][c][0]);
            dq[i][c][1] = fmaf(w, kk.y, dq[i][c][1]);
            dq[i][c][2] = fmaf(w, kk.z, dq[i][c][2]);
            dq[i][c][3] = fmaf(w, kk.w, dq[i][c][3]);
          }
        }
    }
  }
  if (splits > 1) {
    // partial dq tile -> own shared memory (the Q tile is no longer needed), summed across the cluster
    __syncthreads();
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int c = 0; c < HC; ++c)
        *reinterpret_cast<float4*>(Qs + (ty * RM + i) * HS + 4 * tx + 64 * c) =
            make_float4(dq[i][c][0], dq[i][c][1], dq[i][c][2], dq[i][c][3]);
    cg::cluster_group cluster = cg::this_cluster();
    cluster.sync();
    const int rows_n = BM / splits, rows_lo = (int)blockIdx.y * rows_n;
    constexpr int W4 = 16 * HC;    // float4 per row
    for (int idx = tid; idx < rows_n * W4; idx += NT) {
      const int r = rows_lo + idx / W4, c4 = (idx % W4) * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int sp = 0; sp < splits; ++sp) {
        const float* peer = cluster.map_shared_rank(Qs, sp);
        const float4 v = *reinterpret_cast<const float4*>(peer + r * HS + c4);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      const int gr = row0 + r;
      if (gr < a.n) {
        const float vv[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (c4 + e < a.h) a.dqs[(size_t)gr * a.h + c4 + e] = vv[e];
      }
    }
    cluster.sync();   // peers are done reading this CTA's partial
    return;
  }
#pragma unroll
  for (int i = 0; i < RM; ++i) {
    const int gr = row0 + ty * RM + i;
    if (gr >= a.n) continue;
#pragma unroll
    for (int c = 0; c < HC; ++c)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int col = 4 * tx + 64 * c + e;
        if (col < a.h) a.dqs[(size_t)gr * a.h + col] = dq[i][c][e];
      }
  }
}

// v[0..3] += into row[col..col+3]: one 16-byte RED (sm_90+ float4 atomicAdd) when the four columns exist and the
// address is 16-byte aligned (width % 4 == 0), scalar REDs at a ragged edge.
__device__ __forceinline__ void red_add4(float* row, int col, int width, const float (&v)[4]) {
  if (col + 3 < width && (width & 3) == 0) {
    atomicAdd(reinterpret_cast<float4*>(row + col), make_float4(v[0], v[1], v[2], v[3]));
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (col + e < width) atomicAdd(row + col + e, v[e]);
  }
}

// ----------------------------------------------------------------- core_kv
template <int HC, int DVC, int RQ>
__global__ void __launch_bounds__(NT) bwd_core_kv_kernel(CoreArgs a) {
  extern __shared__ __align__(16) float smem[];
  constexpr int BQ = 16 * RQ;      // query rows per step (4 x 4 register tile for the score recompute when RQ = 4)
  constexpr int HS = 64 * HC + 4, VS = 64 * DVC + 4;
  float* Ks = smem;
  float* Vs = Ks + BP * HS;
  float* Qs = Vs + BP * VS;
  float* Gs = Qs + BQ * HS;
  float* As = Gs + BQ * VS;
  float* Ds = As + BQ * PS;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int p0 = blockIdx.x * BP;
  const int h4 = (a.h + 3) & ~3, dv4 = (a.dv + 3) & ~3;
  const int q_begin = blockIdx.y * a.rows_per_split;
  const int q_end = min(a.n, q_begin + a.rows_per_split);

  load_tile_pad(Ks, a.khat, p0, BP, a.p, a.h, 64 * HC, HS);
  load_tile_pad(Vs, a.vhat, p0, BP, a.p, a.dv, 64 * DVC, VS);

  float dk[4][HC][4], dvv[4][DVC][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int c = 0; c < HC; ++c)
#pragma unroll
      for (int e = 0; e < 4; ++e) dk[i][c][e] = 0.f;
#pragma unroll
    for (int c = 0; c < DVC; ++c)
#pragma unroll
      for (int e = 0; e < 4; ++e) dvv[i][c][e] = 0.f;
  }

  for (int q0 = q_begin; q0 < q_end; q0 += BQ) {
    __syncthreads();
    load_tile_pad(Qs, a.qs, q0, BQ, q_end, a.h, 64 * HC, HS);
    load_tile_pad(Gs, a.dr, q0, BQ, q_end, a.dv, 64 * DVC, VS);
    __syncthreads();
    float s[RQ][4], da[RQ][4];
#pragma unroll
    for (int i = 0; i < RQ; ++i)
#pragma unroll
      for (int b = 0; b < 4; ++b) s[i][b] = da[i][b] = 0.f;
    mma_nt<RQ>(s, Qs, HS, Ks, HS, h4, ty, tx);
    mma_nt<RQ>(da, Gs, VS, Vs, VS, dv4, ty, tx);
#pragma unroll
    for (int i = 0; i < RQ; ++i) {
      const int gr = q0 + ty * RQ + i;
      const bool rv = gr < q_end;
      const float lse_r = rv ? __ldg(a.lse + gr) : 0.f;
      const float delta_r = rv ? __ldg(a.delta + gr) : 0.f;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int col = p0 + tx + 16 * b;
        const float prob = (rv && col < a.p) ? __expf(s[i][b] - lse_r) : 0.f;
        As[(ty * RQ + i) * PS + tx + 16 * b] = prob;
        Ds[(ty * RQ + i) * PS + tx + 16 * b] = prob * (da[i][b] - delta_r);
      }
    }
    __syncthreads();
    for (int r = 0; r < BQ; ++r) {
      const float4 a4 = *reinterpret_cast<const float4*>(As + r * PS + ty * 4);
      const float4 d4 = *reinterpret_cast<const float4*>(Ds + r * PS + ty * 4);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w};
      const float dvl[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
      for (int c = 0; c < DVC; ++c) {
        const float4 g = *reinterpret_cast<const float4*>(Gs + r * VS + 4 * tx + 64 * c);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          dvv[i][c][0] = fmaf(av[i], g.x, dvv[i][c][0]);
          dvv[i][c][1] = fmaf(av[i], g.y, dvv[i][c][1]);
          dvv[i][c][2] = fmaf(av[i], g.z, dvv[i][c][2]);
          dvv[i][c][3] = fmaf(av[i], g.w, dvv[i][c][3]);
        }
      }
#pragma unroll
      for (int c = 0; c < HC; ++c) {
        const float4 q = *reinterpret_cast<const float4*>(Qs + r * HS + 4 * tx + 64 * c);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          dk[i][c][0] = fmaf(dvl[i], q.x, dk[i][c][0]);
          dk[i][c][1] = fmaf(dvl[i], q.y, dk[i][c][1]);
          dk[i][c][2] = fmaf(dvl[i], q.z, dk[i][c][2]);
          dk[i][c][3] = fmaf(dvl[i], q.w, dk[i][c][3]);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gp = p0 + ty * 4 + i;
    if (gp >= a.p) continue;
#pragma unroll
    for (int c = 0; c < HC; ++c) red_add4(a.dkhat + (size_t)gp * a.h, 4 * tx + 64 * c, a.h, dk[i][c]);
#pragma unroll
    for (int c = 0; c < DVC; ++c) red_add4(a.dvhat + (size_t)gp * a.dv, 4 * tx + 64 * c, a.dv, dvv[i][c]);
  }
}

// ------------------------------------------------------- tiny-width core (H, Dv <= 4)
// Hop-VAE's `index_to_embedding` layer has 3-wide patterns (HopVAE.py:164-170): the 64-column tiles above spend
// 95 % of their arithmetic on padding there (core_q + core_kv: 42 us at N = 2,048, as much as a 64-wide layer).
// Here all P patterns sit in shared memory as float4 and one launch does both halves of the core backward:
//   phase A (row-major)      8 lanes share a query row, lane s scans patterns s, s+8, ...: dq = sum_j dS_j k_j,
//                            combined with shuffles;
//   phase B (pattern-major)  thread t owns patterns t, t+128, ... of a 512-pattern chunk and walks the CTA's rows
//                            (broadcast reads): dK_j += dS_rj q_r, dV_j += A_rj dR_r in registers, one RED per
//                            element at the end.
// A score costs four FMAs, so computing it once per phase is cheaper than exchanging it.
constexpr int kTinyBwdThreads = 128;
constexpr int kTinyBwdLanes = 8;
constexpr int kTinyBwdRows = kTinyBwdThreads / kTinyBwdLanes;   // 16 query rows per step
constexpr int kTinyBwdPat = 4;                                   // patterns per thread and chunk

__device__ __forceinline__ float dot4(const float4& a, const float4& b) {
  return fmaf(a.w, b.w, fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)));
}

__global__ void __launch_bounds__(kTinyBwdThreads) bwd_core_tiny_kernel(CoreArgs a) {
  extern __shared__ __align__(16) float smem[];
  float4* Ks = reinterpret_cast<float4*>(smem);   // [P]
  float4* Vs = Ks + a.p;                          // [P]
  float4* Qr = Vs + a.p;                          // [rows] query rows of the current step
  float4* Gr = Qr + kTinyBwdRows;                 // [rows] dR rows
  float* Lr = reinterpret_cast<float*>(Gr + kTinyBwdRows);   // [rows] lse
  float* Dr = Lr + kTinyBwdRows;                  // [rows] delta
  const int tid = threadIdx.x;
  for (int j = tid; j < a.p; j += kTinyBwdThreads) {
    float kv[4] = {0.f, 0.f, 0.f, 0.f}, vv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 4; ++c) {                 // (constant indices: the arrays stay in registers)
      if (c < a.h) kv[c] = __ldg(a.khat + (size_t)j * a.h + c);
      if (c < a.dv) vv[c] = __ldg(a.vhat + (size_t)j * a.dv + c);
    }
    Ks[j] = make_float4(kv[0], kv[1], kv[2], kv[3]);
    Vs[j] = make_float4(vv[0], vv[1], vv[2], vv[3]);
  }
  const int sub = tid & (kTinyBwdLanes - 1), rloc = tid / kTinyBwdLanes;
  const int chunk = kTinyBwdThreads * kTinyBwdPat;
  const int n_chunks = (a.p + chunk - 1) / chunk;
  const int n_groups = (a.n + kTinyBwdRows - 1) / kTinyBwdRows;
  float dk[kTinyBwdPat][4], dv[kTinyBwdPat][4];
#pragma unroll
  for (int m = 0; m < kTinyBwdPat; ++m)
#pragma unroll
    for (int e = 0; e < 4; ++e) dk[m][e] = dv[m][e] = 0.f;

  auto flush = [&](int c0) {
#pragma unroll
    for (int m = 0; m < kTinyBwdPat; ++m) {
      const int j = c0 + tid + kTinyBwdThreads * m;
      if (j < a.p) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (c < a.h) atomicAdd(a.dkhat + (size_t)j * a.h + c, dk[m][c]);
          if (c < a.dv) atomicAdd(a.dvhat + (size_t)j * a.dv + c, dv[m][c]);
        }
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) dk[m][e] = dv[m][e] = 0.f;
    }
  };

  for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {
    const long long gr = (long long)grp * kTinyBwdRows + rloc;
    const bool ok = gr < a.n;
    // ---- the row: q, dR, lse, delta = <dR, R>
    float qv[4] = {0.f, 0.f, 0.f, 0.f}, gv[4] = {0.f, 0.f, 0.f, 0.f};
    float lse = INFINITY, delta = 0.f;            // exp(s - inf) = 0: rows past N contribute nothing
    if (ok) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (c < a.h) qv[c] = __ldg(a.qs + gr * a.h + c);
        if (c < a.dv) {
          gv[c] = __ldg(a.dr + gr * a.dv + c);
          delta = fmaf(gv[c], __ldg(a.r + gr * a.dv + c), delta);
        }
      }
      lse = __ldg(a.lse + gr);
    }
    const float4 q = make_float4(qv[0], qv[1], qv[2], qv[3]);
    const float4 g = make_float4(gv[0], gv[1], gv[2], gv[3]);
    __syncthreads();                               // phase B of the previous step is done with Qr.. (and Ks/Vs are staged)
    if (sub == 0) {
      Qr[rloc] = q; Gr[rloc] = g; Lr[rloc] = lse; Dr[rloc] = delta;
      if (ok) a.delta[gr] = delta;
    }
    // ---- phase A: dq of this row
    float dq[4] = {0.f, 0.f, 0.f, 0.f};
    for (int j = sub; j < a.p; j += kTinyBwdLanes) {
      const float4 k = Ks[j];
      const float prob = __expf(dot4(q, k) - lse);
      const float ds = prob * (dot4(g, Vs[j]) - delta);
      dq[0] = fmaf(ds, k.x, dq[0]); dq[1] = fmaf(ds, k.y, dq[1]);
      dq[2] = fmaf(ds, k.z, dq[2]); dq[3] = fmaf(ds, k.w, dq[3]);
    }
#pragma unroll
    for (int off = 1; off < kTinyBwdLanes; off <<= 1)
#pragma unroll
      for (int c = 0; c < 4; ++c) dq[c] += __shfl_xor_sync(0xffffffffu, dq[c], off);
    if (ok && sub == 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (c < a.h) a.dqs[gr * a.h + c] = dq[c];
    }
    __syncthreads();
    // ---- phase B: dK^, dV^ of this thread's patterns over the step's rows
    for (int ch = 0; ch < n_chunks; ++ch) {
      float4 k[kTinyBwdPat], v[kTinyBwdPat];
#pragma unroll
      for (int m = 0; m < kTinyBwdPat; ++m) {
        const int j = ch * chunk + tid + kTinyBwdThreads * m;
        k[m] = j < a.p ? Ks[j] : make_float4(0.f, 0.f, 0.f, 0.f);
        v[m] = j < a.p ? Vs[j] : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll 4
      for (int r = 0; r < kTinyBwdRows; ++r) {
        const float4 qr = Qr[r], gr4 = Gr[r];
        const float lr = Lr[r], dr = Dr[r];
#pragma unroll
        for (int m = 0; m < kTinyBwdPat; ++m) {
          const float prob = __expf(dot4(qr, k[m]) - lr);
          const float ds = prob * (dot4(gr4, v[m]) - dr);
          dk[m][0] = fmaf(ds, qr.x, dk[m][0]); dk[m][1] = fmaf(ds, qr.y, dk[m][1]);
          dk[m][2] = fmaf(ds, qr.z, dk[m][2]); dk[m][3] = fmaf(ds, qr.w, dk[m][3]);
          dv[m][0] = fmaf(prob, gr4.x, dv[m][0]); dv[m][1] = fmaf(prob, gr4.y, dv[m][1]);
          dv[m][2] = fmaf(prob, gr4.z, dv[m][2]); dv[m][3] = fmaf(prob, gr4.w, dv[m][3]);
        }
      }
      if (n_chunks > 1) flush(ch * chunk);         // several chunks share the registers: hand each over at once
    }
  }
  if (n_chunks == 1) flush(0);                     // one chunk: the sums ran across all of this CTA's steps
}

int launch_core_tiny(const CoreArgs& a, cudaStream_t st) {
  const size_t smem = sizeof(float4) * (2 * (size_t)a.p + 2 * kTinyBwdRows) + sizeof(float) * 2 * kTinyBwdRows;
  if (smem > 48 * 1024)
    HOP_CUDA(cudaFuncSetAttribute(bwd_core_tiny_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int n_groups = (a.n + kTinyBwdRows - 1) / kTinyBwdRows;
  int grid = 2 * num_sms();
  if (grid > n_groups) grid = n_groups;
  bwd_core_tiny_kernel<<<grid < 1 ? 1 : grid, kTinyBwdThreads, smem, st>>>(a);
  HOP_LAUNCH_CHECK("bwd_core_tiny_kernel");
  return HOP_OK;
}

// ------------------------------------------------------- LayerNorm backward
// For rows of y = LN(x) (or y = x when !has_ln):  dx (+)= d/dx <scale*g, y>, and the
// affine grads dgamma += sum_rows scale*g*xhat, dbeta += sum_rows scale*g (fp32 RED).
struct LnBwdArgs {
  const float* x; const float* g; int rows; int d;
  int has_ln; const float* gamma; float eps; float scale;
  float* dx; int accumulate;
  float* dgamma; float* dbeta;   // pre-zeroed; nullptr when no affine / not wanted
};

__device__ __forceinline__ void ln_bwd_body(const LnBwdArgs& a, float* sh, int vbx, int vgx) {
  float* sg = sh;                // [2*d] column partials
  float* sb = sh + a.d;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool want_affine = a.has_ln && (a.dgamma || a.dbeta);
  constexpr int kRegCols = 8;                       // columns per lane held in registers: widths up to 256
  const bool reg_row = a.has_ln && a.d <= 32 * kRegCols;       // the row fits the lanes' registers
  const bool reg_affine = want_affine && a.d <= 32 * kRegCols;
  float ag[kRegCols], ab[kRegCols], gam[kRegCols];
#pragma unroll
  for (int k = 0; k < kRegCols; ++k) {
    ag[k] = 0.f; ab[k] = 0.f;
    gam[k] = (reg_row && a.gamma && lane + 32 * k < a.d) ? __ldg(a.gamma + lane + 32 * k) : 1.f;
  }
  if (want_affine) {
    for (int i = threadIdx.x; i < 2 * a.d; i += NT) sh[i] = 0.f;
    __syncthreads();
  }
  for (int r = vbx * (NT / 32) + warp; r < a.rows; r += vgx * (NT / 32)) {
    const float* xr = a.x + (size_t)r * a.d;
    const float* gr = a.g + (size_t)r * a.d;
    float* dxr = a.dx + (size_t)r * a.d;
    if (!a.has_ln) {
      for (int i = lane; i < a.d; i += 32) {
        const float v = a.scale * gr[i];
        dxr[i] = a.accumulate ? dxr[i] + v : v;
      }
      continue;
    }
    if (reg_row) {
      // the whole row lives in registers: x and g are read ONCE (16 independent loads per lane in flight), the statistics
      // are warp reductions over registers in warp_row_stats' order, the column sums stay in registers across rows.
      // (Before: five passes over x / g through L1 with dependent reductions between them -- 234 us per P x D matrix at
      // P = 64k, D = 256, against ~35 us for the bytes.)
      float xv[kRegCols], gv[kRegCols];
#pragma unroll
      for (int k = 0; k < kRegCols; ++k) {
        const int i = lane + 32 * k;
        xv[k] = i < a.d ? __ldg(xr + i) : 0.f;
        gv[k] = i < a.d ? a.scale * __ldg(gr + i) : 0.f;
      }
      float sm = 0.f;
#pragma unroll
      for (int k = 0; k < kRegCols; ++k) sm += xv[k];
      const float mean = warp_sum(sm) / (float)a.d;
      float sv = 0.f;
#pragma unroll
      for (int k = 0; k < kRegCols; ++k) { const float c = xv[k] - mean; sv += (lane + 32 * k < a.d) ? c * c : 0.f; }
      const float rstd = rsqrtf(warp_sum(sv) / (float)a.d + a.eps);
      float c1 = 0.f, c2 = 0.f;
#pragma unroll
      for (int k = 0; k < kRegCols; ++k) {
        xv[k] = (xv[k] - mean) * rstd;                          // x^
        const float gy = gv[k] * gam[k];
        c1 += (lane + 32 * k < a.d) ? gy : 0.f;
        c2 += (lane + 32 * k < a.d) ? gy * xv[k] : 0.f;
      }
      c1 = warp_sum(c1) / (float)a.d;
      c2 = warp_sum(c2) / (float)a.d;
#pragma unroll
      for (int k = 0; k < kRegCols; ++k) {
        const int i = lane + 32 * k;
        if (i < a.d) {
          const float v = rstd * (gv[k] * gam[k] - c1 - xv[k] * c2);
          dxr[i] = a.accumulate ? dxr[i] + v : v;
          ag[k] = fmaf(gv[k], xv[k], ag[k]);
          ab[k] += gv[k];
        }
      }
      continue;
    }
    float mean, rstd;
    warp_row_stats(xr, a.d, a.eps, mean, rstd);
    float c1 = 0.f, c2 = 0.f;
    for (int i = lane; i < a.d; i += 32) {
      const float xh = (xr[i] - mean) * rstd;
      const float gy = a.scale * gr[i] * (a.gamma ? __ldg(a.gamma + i) : 1.f);
      c1 += gy;
      c2 += gy * xh;
    }
    c1 = warp_sum(c1) / (float)a.d;
    c2 = warp_sum(c2) / (float)a.d;
    if (reg_affine) {
      // column sums of this lane's columns stay in registers across its rows (two shared-memory atomics per ELEMENT
      // made this pass 291 us per P x D matrix at P = 64k, D = 256; one per column per warp at the end instead)
#pragma unroll
      for (int k = 0; k < kRegCols; ++k) {
        const int i = lane + 32 * k;
        if (i < a.d) {
          const float xh = (xr[i] - mean) * rstd;
          const float gs = a.scale * gr[i];
          const float gy = gs * (a.gamma ? __ldg(a.gamma + i) : 1.f);
          const float v = rstd * (gy - c1 - xh * c2);
          dxr[i] = a.accumulate ? dxr[i] + v : v;
          ag[k] = fmaf(gs, xh, ag[k]);
          ab[k] += gs;
        }
      }
      continue;
    }
    for (int i = lane; i < a.d; i += 32) {
      const float xh = (xr[i] - mean) * rstd;
      const float gs = a.scale * gr[i];
      const float gy = gs * (a.gamma ? __ldg(a.gamma + i) : 1.f);
      const float v = rstd * (gy - c1 - xh * c2);
      dxr[i] = a.accumulate ? dxr[i] + v : v;
      if (want_affine) {
        atomicAdd(sg + i, gs * xh);
        atomicAdd(sb + i, gs);
      }
    }
  }
  if (reg_affine) {            // (reg_row implies the sums are in ag / ab; without affine outputs they are dropped)
#pragma unroll
    for (int k = 0; k < kRegCols; ++k) {
      const int i = lane + 32 * k;
      if (i < a.d) { atomicAdd(sg + i, ag[k]); atomicAdd(sb + i, ab[k]); }
    }
  }
  if (want_affine) {
    __syncthreads();
    for (int i = threadIdx.x; i < a.d; i += NT) {
      if (a.dgamma) atomicAdd(a.dgamma + i, sg[i]);
      if (a.dbeta) atomicAdd(a.dbeta + i, sb[i]);
    }
  }
}

int ln_bwd_grid(const LnBwdArgs& a) {
  int grid = (a.rows + 7) / 8;
  if (grid > 4 * num_sms()) grid = 4 * num_sms();
  return grid < 1 ? 1 : grid;
}

// ------------------------------------------------------------ weight grads
// dW[o][i] += sum_n dY[n][o] * Xn[n][i],  db[o] += sum_n dY[n][o]   (fp32 RED; pre-zeroed)
// Xn = LN(x) with affine when has_ln (recomputed), else x.
struct WgradArgs {
  const float* dy; const float* x; int rows; int dout; int din;
  int has_ln; const float* gamma; const float* beta_ln; float eps;
  float* dw; float* db; int rows_per_cta; float scale;
};

__device__ __forceinline__ void wgrad_body(const WgradArgs& a, int vbx, int vby, int vbz) {
  constexpr int BR = 32, ST = 68;
  __shared__ __align__(16) float Ys[BR * ST];
  __shared__ __align__(16) float Xs[BR * ST];
  __shared__ float s_mean[BR], s_rstd[BR];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4, warp = tid >> 5, lane = tid & 31;
  const int o0 = vbx * 64, i0 = vby * 64;
  const int r_begin = vbz * a.rows_per_cta, r_end = min(a.rows, r_begin + a.rows_per_cta);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float bsum[4] = {0.f, 0.f, 0.f, 0.f};
  for (int r0 = r_begin; r0 < r_end; r0 += BR) {
    __syncthreads();
    if (a.has_ln) {
      for (int r = warp; r < BR; r += NT / 32) {
        float mean = 0.f, rstd = 1.f;
        if (r0 + r < r_end) warp_row_stats(a.x + (size_t)(r0 + r) * a.din, a.din, a.eps, mean, rstd);
        if (lane == 0) {
          s_mean[r] = mean;
          s_rstd[r] = rstd;
        }
      }
      __syncthreads();
    }
    for (int idx = tid; idx < BR * 64; idx += NT) {
      const int r = idx >> 6, c = idx & 63;
      float yv = 0.f, xv = 0.f;
      if (r0 + r < r_end) {
        if (o0 + c < a.dout) yv = __ldg(a.dy + (size_t)(r0 + r) * a.dout + o0 + c);
        if (i0 + c < a.din) {
          xv = __ldg(a.x + (size_t)(r0 + r) * a.din + i0 + c);
          if (a.has_ln) {
            xv = (xv - s_mean[r]) * s_rstd[r];
            if (a.gamma) xv *= __ldg(a.gamma + i0 + c);
            if (a.beta_ln) xv += __ldg(a.beta_ln + i0 + c);
          }
        }
      }
      Ys[r * ST + c] = yv;
      Xs[r * ST + c] = xv;
    }
    __syncthreads();
#pragma unroll 4
    for (int r = 0; r < BR; ++r) {
      const float4 y4 = *reinterpret_cast<const float4*>(Ys + r * ST + ty * 4);
      const float4 x4 = *reinterpret_cast<const float4*>(Xs + r * ST + tx * 4);
      const float yv[4] = {y4.x, y4.y, y4.z, y4.w};
      const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        bsum[i] += yv[i];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(yv[i], xv[j], acc[i][j]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = o0 + ty * 4 + i;
    if (o >= a.dout) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = i0 + tx * 4 + j;
      if (c < a.din) atomicAdd(a.dw + (size_t)o * a.din + c, acc[i][j] * a.scale);
    }
    if (a.db && vby == 0 && tx == 0) atomicAdd(a.db + o, bsum[i] * a.scale);
  }
}

// grid (dout/64, din/64, row splits); fills a.rows_per_cta
dim3 wgrad_grid(WgradArgs& a) {
  const int sms = num_sms();
  const int tiles = ((a.dout + 63) / 64) * ((a.din + 63) / 64);
  int splits = (2 * sms + tiles - 1) / tiles;
  const int max_splits = (a.rows + 31) / 32;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  a.rows_per_cta = round_up((a.rows + splits - 1) / splits, 32);
  splits = (a.rows + a.rows_per_cta - 1) / a.rows_per_cta;
  return dim3((a.dout + 63) / 64, (a.din + 63) / 64, splits);
}

// y[n][j] = sum_o g[n][o] * w[o][j]   (w row-major (dout_w, din_w); i.e. g . W, no transpose)
struct LinearNnArgs { const float* g; const float* w; float* y; int n, k, m; };

__device__ __forceinline__ void linear_nn_body(const LinearNnArgs& a, int vbx, int vby) {
  const float* __restrict__ g = a.g;
  const float* __restrict__ w = a.w;
  float* __restrict__ y = a.y;
  const int n = a.n, k = a.k, m = a.m;
  // CTA: 32 rows x 64 output columns; contraction streamed in chunks of 32
  constexpr int BR = 32, KC = 32;
  __shared__ __align__(16) float Gs[BR * (KC + 4)];
  __shared__ __align__(16) float Ws[KC * 68];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int row0 = vbx * BR, c0 = vby * 64;
  float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  for (int k0 = 0; k0 < k; k0 += KC) {
    __syncthreads();
    for (int idx = tid; idx < BR * KC; idx += NT) {
      const int r = idx / KC, kk = idx - r * KC;
      Gs[r * (KC + 4) + kk] = (row0 + r < n && k0 + kk < k) ? __ldg(g + (size_t)(row0 + r) * k + k0 + kk) : 0.f;
    }
    for (int idx = tid; idx < KC * 64; idx += NT) {
      const int kk = idx >> 6, c = idx & 63;
      Ws[kk * 68 + c] = (k0 + kk < k && c0 + c < m) ? __ldg(w + (size_t)(k0 + kk) * m + c0 + c) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < KC; ++kk) {
      const float4 wv = *reinterpret_cast<const float4*>(Ws + kk * 68 + 4 * tx);
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const float gv = Gs[(ty * 2 + i) * (KC + 4) + kk];
        acc[i][0] = fmaf(gv, wv.x, acc[i][0]);
        acc[i][1] = fmaf(gv, wv.y, acc[i][1]);
        acc[i][2] = fmaf(gv, wv.z, acc[i][2]);
        acc[i][3] = fmaf(gv, wv.w, acc[i][3]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int gr = row0 + ty * 2 + i;
    if (gr >= n) continue;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int c = c0 + 4 * tx + e;
      if (c < m) y[(size_t)gr * m + c] = acc[i][e];
    }
  }
}

// ------------------------------------------------------------ batched small launches
// The projections, weight gradients and LayerNorm backwards either side of the core are 3-10 us launch-latency-bound
// kernels at the reference's sizes, ten per layer.  Jobs that are ready at the same time run as ONE launch: the
// grid is the concatenation of the jobs' own grids, a block finds its job from the prefix sums.
enum BwdJobKind { kJobLnBwd = 0, kJobWgrad = 1, kJobLinearNn = 2 };
struct BwdJob {
  int kind, block0, gx, gy;       // first block of the job in the merged grid; its own grid is (gx, gy, *)
  LnBwdArgs ln; WgradArgs wg; LinearNnArgs lin;
};
struct BwdJobBatch {
  static constexpr int kMax = 8;
  BwdJob job[kMax];
  int count, blocks;
  size_t smem;
};

__global__ void __launch_bounds__(NT) bwd_job_batch_kernel(const __grid_constant__ BwdJobBatch b) {
  extern __shared__ float sh[];
  int j = 0;
  while (j + 1 < b.count && (int)blockIdx.x >= b.job[j + 1].block0) ++j;
  const BwdJob& job = b.job[j];
  const int v = (int)blockIdx.x - job.block0;
  if (job.kind == kJobLnBwd) {
    ln_bwd_body(job.ln, sh, v, job.gx);
  } else if (job.kind == kJobWgrad) {
    const int per_z = job.gx * job.gy;
    const int vz = v / per_z, rem = v - vz * per_z;
    wgrad_body(job.wg, rem % job.gx, rem / job.gx, vz);
  } else {
    linear_nn_body(job.lin, v % job.gx, v / job.gx);
  }
}

int batch_add_ln_bwd(BwdJobBatch& b, const LnBwdArgs& a) {
  if (b.count >= BwdJobBatch::kMax) HOP_FAIL(HOP_ERR_BAD_ARG, "backward: job batch overflow");
  BwdJob& j = b.job[b.count++];
  j.kind = kJobLnBwd; j.block0 = b.blocks; j.gx = ln_bwd_grid(a); j.gy = 1; j.ln = a;
  b.blocks += j.gx;
  const size_t smem = 2 * (size_t)a.d * sizeof(float);
  if (smem > b.smem) b.smem = smem;
  return HOP_OK;
}
int batch_add_wgrad(BwdJobBatch& b, WgradArgs a) {
  if (b.count >= BwdJobBatch::kMax) HOP_FAIL(HOP_ERR_BAD_ARG, "backward: job batch overflow");
  const dim3 g = wgrad_grid(a);
  BwdJob& j = b.job[b.count++];
  j.kind = kJobWgrad; j.block0 = b.blocks; j.gx = (int)g.x; j.gy = (int)g.y; j.wg = a;
  b.blocks += (int)(g.x * g.y * g.z);
  return HOP_OK;
}
int batch_add_linear_nn(BwdJobBatch& b, const float* g, const float* w, float* y, int n, int k, int m) {
  if (b.count >= BwdJobBatch::kMax) HOP_FAIL(HOP_ERR_BAD_ARG, "backward: job batch overflow");
  BwdJob& j = b.job[b.count++];
  j.kind = kJobLinearNn; j.block0 = b.blocks; j.gx = (n + 31) / 32; j.gy = (m + 63) / 64;
  j.lin = LinearNnArgs{g, w, y, n, k, m};
  b.blocks += j.gx * j.gy;
  return HOP_OK;
}
int launch_job_batch(BwdJobBatch& b, cudaStream_t st) {
  if (b.count == 0 || b.blocks == 0) return HOP_OK;
  bwd_job_batch_kernel<<<b.blocks, NT, b.smem, st>>>(b);
  HOP_LAUNCH_CHECK("bwd_job_batch_kernel");
  b = BwdJobBatch{};
  return HOP_OK;
}

template <int RM, int HC, int DVC>
int launch_core_q(const CoreArgs& a, cudaStream_t st) {
  constexpr int BM = 16 * RM, HS = 64 * HC + 4, VS = 64 * DVC + 4;
  const size_t smem = sizeof(float) * (size_t)(BM * HS + BM * VS + BP * HS + BP * VS + BM * PS);
  auto kern = bwd_core_q_kernel<RM, HC, DVC>;
  if (smem > 48 * 1024) HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // pattern splits across a cluster while everything stays in one wave and every CTA keeps >= 2 tiles
  const int sms = num_sms();
  int per_sm = (int)((227u * 1024u) / smem);
  per_sm = per_sm > 2 ? 2 : (per_sm < 1 ? 1 : per_sm);
  const int row_tiles = (a.n + BM - 1) / BM, n_tiles = (a.p + BP - 1) / BP;
  int splits = 1;
  while (splits < 8 && splits * 2 <= BM && n_tiles >= 4 * splits && row_tiles * splits * 2 <= sms * per_sm) splits <<= 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(row_tiles, splits, 1);
  cfg.blockDim = dim3(NT);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = splits; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = splits > 1 ? 1 : 0;
  HOP_CUDA(cudaLaunchKernelEx(&cfg, kern, a, splits));
  HOP_LAUNCH_CHECK("bwd_core_q_kernel");
  return HOP_OK;
}

template <int HC, int DVC>
int launch_core_kv(CoreArgs a, cudaStream_t st) {
  constexpr int RQ = (HC == 1 && DVC == 1) ? 4 : 2;
  constexpr int BQ = 16 * RQ, HS = 64 * HC + 4, VS = 64 * DVC + 4;
  const size_t smem = sizeof(float) * (size_t)(BP * HS + BP * VS + BQ * HS + BQ * VS + 2 * BQ * PS);
  auto kern = bwd_core_kv_kernel<HC, DVC, RQ>;
  if (smem > 48 * 1024) HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int sms = num_sms();
  const int ptiles = (a.p + BP - 1) / BP;
  // one full wave of resident CTAs (2 per SM at 104 KB of shared memory): with 3 x SMs CTAs the last half wave ran
  // alone -- FFHQ-64 (N = 9,248): 895 -> 838 us per bottleneck training step
  int per_sm = (int)((227u * 1024u) / smem);
  per_sm = per_sm > 2 ? 2 : (per_sm < 1 ? 1 : per_sm);
  const int waves = per_sm;
  int splits = (waves * sms + ptiles - 1) / ptiles;
  const int max_splits = (a.n + BQ - 1) / BQ;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  a.rows_per_split = round_up((a.n + splits - 1) / splits, BQ);
  splits = (a.n + a.rows_per_split - 1) / a.rows_per_split;
  kern<<<dim3(ptiles, splits), NT, smem, st>>>(a);
  HOP_LAUNCH_CHECK("bwd_core_kv_kernel");
  return HOP_OK;
}

template <int HC, int DVC>
int launch_core(const CoreArgs& a, cudaStream_t st) {
  // 64-row tiles (4 x 4 register tile) where the shared-memory budget allows, else 32
  if constexpr (HC == 1 && DVC == 1) HOP_TRY((launch_core_q<4, HC, DVC>(a, st)));
  else HOP_TRY((launch_core_q<2, HC, DVC>(a, st)));
  return launch_core_kv<HC, DVC>(a, st);
}

// One launch zeroes every accumulation target of a backward call (the RED / atomicAdd destinations): at the
// reference's sizes a dozen cudaMemsetAsync calls cost more host time than the arithmetic.
struct ZeroBatch {
  static constexpr int kMax = 20;
  float* ptr[kMax]; unsigned long long elems[kMax]; int count;
};

__global__ void __launch_bounds__(NT) multi_zero_kernel(ZeroBatch z) {
  float* p = z.ptr[blockIdx.y];
  const unsigned long long n = z.elems[blockIdx.y];
  for (unsigned long long i = (unsigned long long)blockIdx.x * NT + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * NT)
    p[i] = 0.f;
}

// ------------------------------------------------------------ GEMM-shaped tails on the tensor pipe
// Around the tensor-core association core (hop_bwd_tc.cu) the out-projection and V-projection backwards are four dense
// products -- dR = G Wo, dWo = G^T R, dVn = dV^ Wv, dWv = dV^^T LN_v(W) -- that the FFMA job kernels above run at ~10
// TFLOP/s (2.06 ms of a 5.5 ms backward at N = 16k, P = 64k, D = 256; profiles/r02_bwd_launches.csv).  Where the widths
// are multiples of 128 they go to hop_gemm_tc.cu instead: every operand is packed once as an fp16 hi + lo pair (one HBM
// pass, row-major and transposed as the product needs it), the gradient operands divided by their own max|.| first (a
// device scalar from absmax_kernel, multiplied back in the GEMM's epilogue) so that gradients of any magnitude sit in
// fp16's range; the weight gradients contract over N or P rows and run split over that axis with fp32 REDs.  The bias
// gradients are column sums.  LayerNorm backwards stay on the job kernels (they are HBM passes).
__global__ void __launch_bounds__(NT) absmax_kernel(const float* __restrict__ x, size_t n, float* __restrict__ out) {
  float m = 0.f;
  const size_t n4 = n >> 2;
  for (size_t i = (size_t)blockIdx.x * NT + threadIdx.x; i < n4; i += (size_t)gridDim.x * NT) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
  }
  for (size_t i = (n4 << 2) + (size_t)blockIdx.x * NT + threadIdx.x; i < n; i += (size_t)gridDim.x * NT) m = fmaxf(m, fabsf(x[i]));
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(reinterpret_cast<int*>(out), __float_as_int(m));   // (>= 0: bit order = value order)
}

// out[c] += scale * sum_r x[r][c]   (out zeroed by the caller); 64 columns x a row range per CTA
__global__ void __launch_bounds__(NT) colsum_kernel(const float* __restrict__ x, int rows, int cols, int rows_per_cta,
                                                    float* __restrict__ out) {
  __shared__ float red[4][64];
  const int c = blockIdx.x * 64 + (threadIdx.x & 63), part = threadIdx.x >> 6;
  const int r0 = blockIdx.y * rows_per_cta, r1 = min(rows, r0 + rows_per_cta);
  float acc = 0.f;
  if (c < cols)
    for (int r = r0 + part; r < r1; r += 4) acc += __ldg(x + (size_t)r * cols + c);
  red[part][threadIdx.x & 63] = acc;
  __syncthreads();
  if (part == 0 && c < cols) atomicAdd(out + c, (red[0][threadIdx.x] + red[1][threadIdx.x]) + (red[2][threadIdx.x] + red[3][threadIdx.x]));
}

int launch_absmax(const float* x, size_t n, float* out, cudaStream_t st) {
  size_t blocks = (n / 4 + NT - 1) / NT;
  const size_t cap = (size_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  absmax_kernel<<<(int)blocks, NT, 0, st>>>(x, n, out);
  HOP_LAUNCH_CHECK("absmax_kernel");
  return HOP_OK;
}
int launch_colsum(const float* x, int rows, int cols, float* out, cudaStream_t st) {
  const int cgrid = (cols + 63) / 64;
  int splits = (2 * num_sms() + cgrid - 1) / cgrid;
  if (splits > (rows + 63) / 64) splits = (rows + 63) / 64;
  if (splits < 1) splits = 1;
  const int per = (rows + splits - 1) / splits;
  colsum_kernel<<<dim3(cgrid, (rows + per - 1) / per), NT, 0, st>>>(x, rows, cols, per, out);
  HOP_LAUNCH_CHECK("colsum_kernel");
  return HOP_OK;
}

struct TailWs {
  float* range;                                   // [0] = max|G|, [1] = max|dV^| (zeroed with the accumulation targets)
  void *g_hi, *g_lo, *gt_hi, *gt_lo;              // G (Npad, Dout) and G^T (Dout, Npad)
  void *rt_hi, *rt_lo;                            // R^T (Dv, Npad)
  void *wot_hi, *wot_lo;                          // Wo^T (Dv, Dout)
  void *dv_hi, *dv_lo, *dvt_hi, *dvt_lo;          // dV^ (Ppad, Dv) and dV^^T (Dv, Ppad)
  void *xt_hi, *xt_lo;                            // LN_v(Wt)^T (Din, Ppad)
  void *wvt_hi, *wvt_lo;                          // Wv^T (Din, Dv)
  size_t total;
};
// which of the two tails run as tensor-core GEMMs for this layer (needs the tensor-core core)
static inline bool tail_out_gemm(const hop_desc& d) {
  return (d.flags & HOP_F_OUT_PROJ) && d.d_out % 128 == 0 && d.d_v % 128 == 0 && d.d_out <= 1024 && d.d_v <= 1024;
}
static inline bool tail_v_gemm(const hop_desc& d) {
  return (d.flags & HOP_F_PROJ_PATTERN) && d.d_v % 128 == 0 && d.d_in % 128 == 0 && d.d_in <= 1024;
}
TailWs carve_tail(const hop_desc& d, void* ws) {
  TailWs w{};
  size_t off = 0;
  auto take = [&](size_t bytes) -> void* {
    void* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += round_up_sz(bytes, 256);
    return p;
  };
  const size_t npad = round_up(d.n_rows, 128), ppad = padded_patterns(d.n_patterns);
  w.range = static_cast<float*>(take(256));
  if (tail_out_gemm(d)) {
    w.g_hi = take(npad * d.d_out * 2); w.g_lo = take(npad * d.d_out * 2);
    w.gt_hi = take(npad * d.d_out * 2); w.gt_lo = take(npad * d.d_out * 2);
    w.rt_hi = take(npad * d.d_v * 2); w.rt_lo = take(npad * d.d_v * 2);
    w.wot_hi = take((size_t)d.d_out * d.d_v * 2); w.wot_lo = take((size_t)d.d_out * d.d_v * 2);
  }
  if (tail_v_gemm(d)) {
    w.dv_hi = take(ppad * d.d_v * 2); w.dv_lo = take(ppad * d.d_v * 2);
    w.dvt_hi = take(ppad * d.d_v * 2); w.dvt_lo = take(ppad * d.d_v * 2);
    w.xt_hi = take(ppad * d.d_in * 2); w.xt_lo = take(ppad * d.d_in * 2);
    w.wvt_hi = take((size_t)d.d_in * d.d_v * 2); w.wvt_lo = take((size_t)d.d_in * d.d_v * 2);
  }
  w.total = off;
  return w;
}
int weight_grad_splits(int m, int n, int kchunks) {
  const int tiles = ((m + 127) / 128) * ((n + 255) / 256);
  int s = (num_sms() + tiles - 1) / tiles;
  if (s > kchunks) s = kchunks;
  return s < 1 ? 1 : s;
}

struct BwdWs {
  float *dr, *dqs, *delta, *dkhat, *dvhat, *dvn, *dkn, *dqn;
  size_t total;
};

BwdWs carve(const hop_desc& d, void* ws) {
  BwdWs w{};
  size_t off = 0;
  auto take = [&](size_t elems) {
    float* p = ws ? reinterpret_cast<float*>(static_cast<char*>(ws) + off) : nullptr;
    off += round_up_sz(elems * sizeof(float), 256);
    return p;
  };
  const size_t N = d.n_rows, P = d.n_patterns;
  w.dr = take(N * d.d_v);
  w.dqs = take(N * d.d_assoc);
  w.delta = take(N);
  w.dkhat = take(P * d.d_assoc);
  w.dvhat = take(P * d.d_v);
  w.dvn = take(P * d.d_in);
  w.dkn = take(P * d.d_in);
  w.dqn = take(N * d.d_in);
  w.total = off;
  return w;
}

}  // namespace

size_t simt_backward_workspace(const hop_desc& d) { return carve(d, nullptr).total; }
size_t gemm_tail_workspace(const hop_desc& d) {
  if (d.tune_flags & HOP_TUNE_BWD_TAILS_SIMT) return 0;
  return (tail_out_gemm(d) || tail_v_gemm(d)) ? carve_tail(d, nullptr).total : 0;
}

int launch_simt_backward(const hop_desc& d, const hop_params& p, const hop_patterns& pt, const float* x,
                         const float* qhat, const float* retrieved, const float* lse, const int32_t* argmax,
                         const float* grad_out, const hop_grads& g, void* ws, size_t ws_bytes, cudaStream_t st) {
  // association core: tcgen05 (hop_bwd_tc.cu) when the step ran on a tensor-core path, else the fp32 kernels below
  const bool tc_core = d.path != HOP_PATH_FP32_SIMT && d.path != HOP_PATH_AUTO && !(d.tune_flags & HOP_TUNE_BWD_SIMT) &&
                       tc_backward_supported(d.n_rows, d.n_patterns, d.d_assoc, d.d_v);
  if (!tc_core && (d.d_assoc > 128 || d.d_v > 128))
    HOP_FAIL(HOP_ERR_UNSUPPORTED, "fp32 backward core is built for H, Dv <= 128 (got %d, %d); wider layers train on the "
             "tensor-core path (hop_sm100_select_train_path)", d.d_assoc, d.d_v);
  const BwdWs w = carve(d, ws);
  const int N = d.n_rows, P = d.n_patterns, H = d.d_assoc, Dv = d.d_v, Din = d.d_in;
  // GEMM-shaped tails on the tensor pipe (see carve_tail); their packs sit behind the tensor-core core's workspace
  const bool tails_tc = tc_core && !(d.tune_flags & HOP_TUNE_BWD_TAILS_SIMT);
  const bool out_gemm = tails_tc && tail_out_gemm(d), v_gemm = tails_tc && tail_v_gemm(d);
  const size_t tc_ws = tc_core ? tc_backward_workspace(N, P, H, Dv) : 0;
  TailWs tw{};
  if (out_gemm || v_gemm) {
    tw = carve_tail(d, static_cast<char*>(ws) + w.total + tc_ws);
    if (ws_bytes < w.total + tc_ws + tw.total)
      HOP_FAIL(HOP_ERR_WORKSPACE, "backward: %zu workspace bytes needed, %zu given", w.total + tc_ws + tw.total, ws_bytes);
  }
  const int npad = round_up(N, 128), ppad = padded_patterns(P);
#ifdef HOP_TAIL_NO_RANGE        // (experiment: operands packed unscaled)
  float* const range_g = nullptr; float* const range_v = nullptr;
#else
  float* const range_g = tw.range; float* const range_v = tw.range ? tw.range + 1 : nullptr;
#endif
  const bool out_proj = d.flags & HOP_F_OUT_PROJ, q_proj = d.flags & HOP_F_PROJ_STATE;
  const bool k_proj = d.flags & HOP_F_PROJ_STORED, v_proj = d.flags & HOP_F_PROJ_PATTERN;
  const bool sep_targets = p.lookup_targets != nullptr;

  // every accumulation target is known up front: collect them, zero them with one launch
  ZeroBatch zb{};
  size_t zmax = 0;
  auto zero = [&](float* ptr, size_t elems) -> int {
    if (ptr && elems) {
      if (zb.count >= ZeroBatch::kMax) HOP_FAIL(HOP_ERR_BAD_ARG, "backward: too many accumulation targets");
      zb.ptr[zb.count] = ptr; zb.elems[zb.count] = elems; ++zb.count;
      if (elems > zmax) zmax = elems;
    }
    return HOP_OK;
  };
  {
    const bool has_wt = (sep_targets ? g.d_lookup_targets : g.d_lookup) != nullptr;
    if (out_proj && g.d_wo) { HOP_TRY(zero(g.d_wo, (size_t)d.d_out * Dv)); HOP_TRY(zero(g.d_bo, d.d_out)); }
    if (out_gemm || v_gemm) HOP_TRY(zero(tw.range, 4));
    HOP_TRY(zero(w.dkhat, (size_t)P * H));
    HOP_TRY(zero(w.dvhat, (size_t)P * Dv));
    if (tc_core) HOP_TRY(zero(w.dqs, (size_t)N * H));      // the tensor-core core accumulates dq with REDs too
    if (q_proj && g.d_wq) { HOP_TRY(zero(g.d_wq, (size_t)H * Din)); HOP_TRY(zero(g.d_bq, H)); }
    if (g.d_x) { HOP_TRY(zero(g.d_ln_state_w, Din)); HOP_TRY(zero(g.d_ln_state_b, Din)); }
    if (v_proj && g.d_wv) { HOP_TRY(zero(g.d_wv, (size_t)Dv * Din)); HOP_TRY(zero(g.d_bv, Dv)); }
    if (has_wt) { HOP_TRY(zero(g.d_ln_proj_w, Din)); HOP_TRY(zero(g.d_ln_proj_b, Din)); }
    if (k_proj && g.d_wk) { HOP_TRY(zero(g.d_wk, (size_t)H * Din)); HOP_TRY(zero(g.d_bk, H)); }
    if (g.d_lookup) { HOP_TRY(zero(g.d_ln_stored_w, Din)); HOP_TRY(zero(g.d_ln_stored_b, Din)); }
    if (zb.count) {
      int bx = (int)((zmax + NT * 4 - 1) / (NT * 4));
      if (bx > 64) bx = 64;
      multi_zero_kernel<<<dim3(bx < 1 ? 1 : bx, zb.count), NT, 0, st>>>(zb);
      HOP_LAUNCH_CHECK("multi_zero_kernel");
    }
  }

  // 1. out-projection backward: dR = G . Wo and dWo = G^T . R are independent -> one launch
  BwdJobBatch stage[3] = {};
  const float* dR = grad_out;
  if (out_proj && out_gemm) {
    HOP_TRY(launch_absmax(grad_out, (size_t)N * d.d_out, tw.range, st));
    Pack16Args pg{grad_out, N, npad, d.d_out, 0, nullptr, nullptr, 0.f, range_g, tw.g_hi, tw.g_lo,
                  g.d_wo ? tw.gt_hi : nullptr, g.d_wo ? tw.gt_lo : nullptr};
    HOP_TRY(launch_pack16(pg, st));
    Pack16Args pw{p.wo, d.d_out, d.d_out, Dv, 0, nullptr, nullptr, 0.f, nullptr, nullptr, nullptr, tw.wot_hi, tw.wot_lo};
    HOP_TRY(launch_pack16(pw, st));
    TcGemmArgs m1{};                      // dR (N, Dv) = G (N, Dout) . Wo
    m1.a16 = tw.g_hi; m1.a16lo = tw.g_lo; m1.b16 = tw.wot_hi; m1.b16lo = tw.wot_lo;
    m1.m = N; m1.m_pad = npad; m1.n = Dv; m1.k = d.d_out; m1.ldc = Dv; m1.c = w.dr; m1.split = 1; m1.out_scale_dev = range_g;
    HOP_TRY(launch_tc_gemm(m1, st));
    dR = w.dr;
    if (g.d_wo) {
      Pack16Args pr{retrieved, N, npad, Dv, 0, nullptr, nullptr, 0.f, nullptr, nullptr, nullptr, tw.rt_hi, tw.rt_lo};
      HOP_TRY(launch_pack16(pr, st));
      TcGemmArgs m2{};                    // dWo (Dout, Dv) += G^T . R, contraction over the N rows split across CTAs
      m2.a16 = tw.gt_hi; m2.a16lo = tw.gt_lo; m2.b16 = tw.rt_hi; m2.b16lo = tw.rt_lo;
      m2.m = d.d_out; m2.m_pad = d.d_out; m2.n = Dv; m2.k = npad; m2.ldc = Dv; m2.c = g.d_wo; m2.split = 1;
      m2.k_splits = weight_grad_splits(d.d_out, Dv, npad / 64); m2.out_scale_dev = range_g;
      if (m2.k_splits < 2) m2.k_splits = 2;         // (always the accumulating form: d_wo is zeroed above)
      HOP_TRY(launch_tc_gemm(m2, st));
      if (g.d_bo) HOP_TRY(launch_colsum(grad_out, N, d.d_out, g.d_bo, st));
    }
  } else if (out_proj) {
    HOP_TRY(batch_add_linear_nn(stage[0], grad_out, p.wo, w.dr, N, d.d_out, Dv));
    dR = w.dr;
    if (g.d_wo) {
      WgradArgs wa{grad_out, retrieved, N, d.d_out, Dv, 0, nullptr, nullptr, 0.f, g.d_wo, g.d_bo, 0, 1.f};
      HOP_TRY(batch_add_wgrad(stage[0], wa));
    }
  }
  HOP_TRY(launch_job_batch(stage[0], st));

  // 2. association core
  CoreArgs ca{qhat, pt.khat, pt.vhat, lse, dR, retrieved, w.delta, w.dqs, w.dkhat, w.dvhat, N, P, H, Dv, 0};
  const int hc = (H + 63) / 64, dvc = (Dv + 63) / 64;
  if (tc_core) {
    TcBwdArgs ta{};
    ta.qs = qhat; ta.lse = lse; ta.dr = dR; ta.r = retrieved;
    ta.k16 = pt.k16; ta.v16 = pt.v16; ta.kt16 = pt.kt16; ta.khat = pt.khat; ta.argmax = argmax;
    ta.dqs = w.dqs; ta.dkhat = w.dkhat; ta.dvhat = w.dvhat;
    ta.n = N; ta.p = P; ta.h = H; ta.dv = Dv;
    ta.ws = static_cast<char*>(ws) + w.total;
    ta.ws_bytes = ws_bytes > w.total ? ws_bytes - w.total : 0;
    HOP_TRY(launch_tc_backward(ta, st));
  } else if (H <= 4 && Dv <= 4 && P <= 4096) HOP_TRY(launch_core_tiny(ca, st));
  else if (hc == 1 && dvc == 1) HOP_TRY((launch_core<1, 1>(ca, st)));
  else if (hc == 1 && dvc == 2) HOP_TRY((launch_core<1, 2>(ca, st)));
  else if (hc == 2 && dvc == 1) HOP_TRY((launch_core<2, 1>(ca, st)));
  else HOP_TRY((launch_core<2, 2>(ca, st)));

  // 3-5. the three tails (query side, K^ path, V^ path) in stages: stage 0 = everything that needs only the core's
  // outputs (weight gradients, the Linear backwards, LayerNorm backwards without a projection in front), stage 1 =
  // LayerNorm backwards behind a projection, stage 2 = a second writer of d_lookup (the K^ and V^ paths share it
  // unless lookup_targets exists; the later one accumulates, so they must not run in the same launch).
  // query side: dqs -> [Wq] -> LN_q -> dx
  {
    const float* gq = w.dqs;     // gradient w.r.t. beta*Qproj
    if (q_proj) {
      // dWq = (beta dqs)^T LN(x), dbq = sum beta dqs ; dqn = beta dqs . Wq
      if (g.d_wq) {
        WgradArgs wa{w.dqs, x, N, H, Din, (d.flags & HOP_F_NORM_STATE) ? 1 : 0, p.ln_state_w, p.ln_state_b, d.ln_eps,
                     g.d_wq, g.d_bq, 0, d.beta};
        HOP_TRY(batch_add_wgrad(stage[0], wa));
      }
      HOP_TRY(batch_add_linear_nn(stage[0], w.dqs, p.wq, w.dqn, N, H, Din));
      gq = w.dqn;
    }
    if (g.d_x) {
      LnBwdArgs la{x, gq, N, Din, (d.flags & HOP_F_NORM_STATE) ? 1 : 0, p.ln_state_w, d.ln_eps, d.beta,
                   g.d_x, 0, g.d_ln_state_w, g.d_ln_state_b};
      HOP_TRY(batch_add_ln_bwd(stage[q_proj ? 1 : 0], la));
    }
  }
  // pattern side, K^ path: dK^ -> [Wk] -> LN_k -> d_lookup (first writer)
  int k_stage = -1;
  {
    const float* gk = w.dkhat;
    if (k_proj) {
      if (g.d_wk) {
        WgradArgs wa{w.dkhat, p.lookup, P, H, Din, (d.flags & HOP_F_NORM_STORED) ? 1 : 0, p.ln_stored_w, p.ln_stored_b,
                     d.ln_eps, g.d_wk, g.d_bk, 0, 1.f};
        HOP_TRY(batch_add_wgrad(stage[0], wa));
      }
      HOP_TRY(batch_add_linear_nn(stage[0], w.dkhat, p.wk, w.dkn, P, H, Din));
      gk = w.dkn;
    }
    if (g.d_lookup) {
      k_stage = k_proj ? 1 : 0;
      LnBwdArgs la{p.lookup, gk, P, Din, (d.flags & HOP_F_NORM_STORED) ? 1 : 0, p.ln_stored_w, d.ln_eps, 1.f,
                   g.d_lookup, 0, g.d_ln_stored_w, g.d_ln_stored_b};
      HOP_TRY(batch_add_ln_bwd(stage[k_stage], la));
    }
  }
  // pattern side, V^ path: dV^ -> [Wv] -> LN_v -> d_lookup(_targets) (accumulates behind the K^ path when shared)
  {
    const float* Wt = sep_targets ? p.lookup_targets : p.lookup;
    float* d_wt = sep_targets ? g.d_lookup_targets : g.d_lookup;
    const float* gv = w.dvhat;
    if (v_proj && v_gemm) {
      HOP_TRY(launch_absmax(w.dvhat, (size_t)P * Dv, tw.range + 1, st));
      Pack16Args pd{w.dvhat, P, ppad, Dv, 0, nullptr, nullptr, 0.f, range_v, tw.dv_hi, tw.dv_lo,
                    g.d_wv ? tw.dvt_hi : nullptr, g.d_wv ? tw.dvt_lo : nullptr};
      HOP_TRY(launch_pack16(pd, st));
      Pack16Args pw{p.wv, Dv, Dv, Din, 0, nullptr, nullptr, 0.f, nullptr, nullptr, nullptr, tw.wvt_hi, tw.wvt_lo};
      HOP_TRY(launch_pack16(pw, st));
      TcGemmArgs m3{};                    // dVn (P, Din) = dV^ (P, Dv) . Wv
      m3.a16 = tw.dv_hi; m3.a16lo = tw.dv_lo; m3.b16 = tw.wvt_hi; m3.b16lo = tw.wvt_lo;
      m3.m = P; m3.m_pad = ppad; m3.n = Din; m3.k = Dv; m3.ldc = Din; m3.c = w.dvn; m3.split = 1; m3.out_scale_dev = range_v;
      HOP_TRY(launch_tc_gemm(m3, st));
      if (g.d_wv) {
        Pack16Args px{Wt, P, ppad, Din, (d.flags & HOP_F_NORM_PROJ) ? 1 : 0, p.ln_proj_w, p.ln_proj_b, d.ln_eps, nullptr,
                      nullptr, nullptr, tw.xt_hi, tw.xt_lo};
        HOP_TRY(launch_pack16(px, st));
        TcGemmArgs m4{};                  // dWv (Dv, Din) += dV^^T . LN_v(Wt), contraction over the P patterns split
        m4.a16 = tw.dvt_hi; m4.a16lo = tw.dvt_lo; m4.b16 = tw.xt_hi; m4.b16lo = tw.xt_lo;
        m4.m = Dv; m4.m_pad = Dv; m4.n = Din; m4.k = ppad; m4.ldc = Din; m4.c = g.d_wv; m4.split = 1;
        m4.k_splits = weight_grad_splits(Dv, Din, ppad / 64); m4.out_scale_dev = range_v;
        if (m4.k_splits < 2) m4.k_splits = 2;
        HOP_TRY(launch_tc_gemm(m4, st));
        if (g.d_bv) HOP_TRY(launch_colsum(w.dvhat, P, Dv, g.d_bv, st));
      }
      gv = w.dvn;
    } else if (v_proj) {
      if (g.d_wv) {
        WgradArgs wa{w.dvhat, Wt, P, Dv, Din, (d.flags & HOP_F_NORM_PROJ) ? 1 : 0, p.ln_proj_w, p.ln_proj_b, d.ln_eps,
                     g.d_wv, g.d_bv, 0, 1.f};
        HOP_TRY(batch_add_wgrad(stage[0], wa));
      }
      HOP_TRY(batch_add_linear_nn(stage[0], w.dvhat, p.wv, w.dvn, P, Dv, Din));
      gv = w.dvn;
    }
    if (d_wt) {
      const int accumulate = (!sep_targets && k_stage >= 0) ? 1 : 0;
      int v_stage = v_proj ? 1 : 0;
      if (accumulate && v_stage <= k_stage) v_stage = k_stage + 1;
      LnBwdArgs la{Wt, gv, P, Din, (d.flags & HOP_F_NORM_PROJ) ? 1 : 0, p.ln_proj_w, d.ln_eps, 1.f,
                   d_wt, accumulate, g.d_ln_proj_w, g.d_ln_proj_b};
      HOP_TRY(batch_add_ln_bwd(stage[v_stage], la));
    }
  }
  for (int i = 0; i < 3; ++i) HOP_TRY(launch_job_batch(stage[i], st));
  return HOP_OK;
}

}  // namespace hop
