// Exact fp32 (CUDA-core) association pass: flash-style, the N x P score matrix
// lives only in registers / shared memory.
//
// Replaces, for small or argmax-tight problems (the reference's own sizes:
// P=512, D in {64, 3}):
//   q*scaling, torch.bmm(q,k^T), softmax(dim=-1), torch.bmm(A,v), out_proj
// of upstream hopfield_core_forward, reached from HopVAE.py:193,215,216,234,
// 255,257,271.  LN_q (row a3) is fused into the prologue, out_proj (a10) into
// the epilogue.
//
// Tiling: a CTA of 256 threads owns BM = 16*RM query rows and sweeps the
// patterns in tiles of 64.  Threads form a 16 x 16 grid: ty owns RM rows, tx
// owns 4 score columns (tx + 16 b) and 4*DVC read-out columns (4 tx + 64 c).
// The 16 threads of a row group keep the running (max, sum, argmax) replicated
// via half-warp butterfly shuffles.
//
// Small problems (the reference's N = 2k..9k rows against P = 512 patterns) do not fill 148 SMs with row tiles
// alone: the pattern axis is then cut into `splits` ranges, one CTA each, and the CTAs of one row tile form a
// thread-block cluster (1 x splits x 1).  Every CTA leaves its un-normalised partial (O, max, sum, argmax) in
// its own shared memory; after a cluster barrier CTA r merges rows [r BM/splits, (r+1) BM/splits) straight out
// of its peers' shared memory (DSMEM), runs the fused out-projection on them and writes the result -- still
// one launch, and no partials in HBM.
#include <cooperative_groups.h>

#include "hop_common.cuh"

namespace cg = cooperative_groups;

namespace hop {

namespace {

constexpr int NT = 256;
constexpr int BP = 64;
constexpr int PS = BP + 4;  // stride of the probability tile

struct SimtSmemLayout {
  int hs;      // stride of Q / K tiles (floats)
  int ds;      // stride of read-out tile (floats) = 64*DVC
  int q_off, k_off, b_off, p_off, st_off, total;  // float offsets
  int k_elems, b_elems, nbuf;                     // per-buffer sizes of the K / read-out tiles, buffers (1 | 2)
};

__host__ __device__ inline SimtSmemLayout simt_layout_n(int bm, int h, int dvc, int nbuf) {
  SimtSmemLayout L;
  const int h4 = (h + 3) / 4 * 4;
  L.hs = h4 + 4;
  L.ds = 64 * dvc;
  L.nbuf = nbuf;
  L.q_off = 0;
  L.k_off = L.q_off + bm * L.hs;
  L.k_elems = BP * L.hs;
  L.b_off = L.k_off + nbuf * L.k_elems;
  L.b_elems = (BP > bm ? BP : bm) * L.ds;  // buffer 0 is reused as the O tile in the epilogue
  L.p_off = L.b_off + nbuf * L.b_elems;
  L.st_off = L.p_off + bm * PS;      // per-row (max, sum, argmax) of this CTA's pattern range
  L.total = L.st_off + 3 * bm;
  return L;
}

// Two K^ / read-out tile buffers (cp.async prefetch of tile t+1 under the arithmetic of tile t) while two CTAs
// still fit an SM, else one.
__host__ __device__ inline SimtSmemLayout simt_layout(int bm, int h, int dvc) {
  const SimtSmemLayout L2 = simt_layout_n(bm, h, dvc, 2);
  if ((size_t)L2.total * sizeof(float) <= 110u * 1024u) return L2;
  return simt_layout_n(bm, h, dvc, 1);
}

__device__ __forceinline__ void cp_async16(float* dst_smem, const float* src, bool valid) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst_smem);
  const int bytes = valid ? 16 : 0;            // src-size 0: the 16 destination bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Load rows [r0, r0+rows) x [0, w) of a (nrows_total, w) fp32 matrix into smem
// with stride `stride`, zero-filling rows >= nrows_total and columns in [w, wpad).
__device__ __forceinline__ void load_tile(float* __restrict__ dst, const float* __restrict__ src, int r0,
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

// Same tile as load_tile, issued as 16-byte cp.async copies when the row width allows (multiple of 4 floats);
// other widths fall back to the synchronous loader.  The caller commits / waits.
__device__ __forceinline__ void load_tile_async(float* __restrict__ dst, const float* __restrict__ src, int r0,
                                                int rows, int nrows_total, int w, int wpad, int stride) {
  if ((w & 3) != 0) {
    load_tile(dst, src, r0, rows, nrows_total, w, wpad, stride);
    return;
  }
  const int w4 = wpad >> 2;
  for (int idx = threadIdx.x; idx < rows * w4; idx += NT) {
    const int r = idx / w4, c = (idx - r * w4) << 2;
    const bool valid = r0 + r < nrows_total && c < w;
    cp_async16(dst + r * stride + c, valid ? src + (size_t)(r0 + r) * w + c : src, valid);
  }
}

// HF: association width known at compile time (64 -- Hop-VAE's `hopfield` and `embedding_to_index` layers) or 0.
// With a runtime width 30 % of the issued instructions were address arithmetic and loop control of the two
// contraction loops (IMAD / ISETP / BRA / LEA); the fixed width folds the strides and unrolls them.
template <int RM, int DVC, int HF>
__global__ void __launch_bounds__(NT) hop_fwd_simt_kernel(SimtPassArgs a, int splits) {
  extern __shared__ __align__(16) float smem[];
  constexpr int BM = 16 * RM;
  const SimtSmemLayout L = simt_layout(BM, HF ? HF : a.h, DVC);
  const int hs = HF ? HF + 4 : L.hs;            // stride of the Q / K tiles
  float* Qs = smem + L.q_off;
  float* Ks = smem + L.k_off;
  float* Bs = smem + L.b_off;
  float* Ps = smem + L.p_off;

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * BM;
  const int h4 = HF ? HF : (a.h + 3) / 4 * 4;

  if (a.stopped != nullptr && *a.stopped != 0u) {
    // the stop rule has fired (hop_repair.cu): this update pass carries the state through unchanged.  Every CTA of a
    // cluster reads the same flag, so all of them leave before the first cluster barrier.
    if (blockIdx.y == 0 && a.res != nullptr && a.qs != nullptr)
      for (int idx = tid; idx < BM * a.h; idx += NT) {
        const int r = idx / a.h, c = idx - r * a.h;
        if (row0 + r < a.n) a.res[(size_t)(row0 + r) * a.dr + c] = a.qs[(size_t)(row0 + r) * a.h + c];
      }
    return;
  }

  const int n_tiles = (a.p + BP - 1) / BP;
  const int tiles_per_split = (n_tiles + splits - 1) / splits;
  const int t_begin = (int)blockIdx.y * tiles_per_split;
  const int t_end = min(n_tiles, t_begin + tiles_per_split);
  float* const K0 = Ks;
  float* const B0 = Bs;
  auto issue = [&](int t) {            // tile t -> buffer (t - t_begin) & 1 (or the only buffer)
    const int buf = L.nbuf == 2 ? ((t - t_begin) & 1) : 0;
    load_tile_async(K0 + buf * L.k_elems, a.khat, t * BP, BP, a.p, a.h, h4, L.hs);
    load_tile_async(B0 + buf * L.b_elems, a.bmat, t * BP, BP, a.p, a.dr, L.ds, L.ds);
    cp_async_commit();
  };
  // ---- prologue: queries -> smem (fused LayerNorm + beta when x_raw given) ----
  // The raw rows are staged with one batch of copies (all in flight at once) and normalised in place: a warp that
  // reads its rows from global memory one after another exposes the full memory latency once per row (ncu: 29 %
  // of this kernel's stall samples at N = 2,048).
  if (a.x_raw != nullptr) {
    load_tile_async(Qs, a.x_raw, row0, BM, a.n, a.h, h4, L.hs);
    cp_async_commit();
  }
  const bool prefetch0 = L.nbuf == 2 && t_begin < t_end;
  if (prefetch0) issue(t_begin);                            // first pattern tile: overlaps the LayerNorm below
  if (a.x_raw != nullptr) {
    if (prefetch0) cp_async_wait<1>(); else cp_async_wait<0>();
    __syncthreads();
    for (int r = warp; r < BM; r += NT / 32) {
      const int gr = row0 + r;
      float* qr = Qs + r * L.hs;
      if (gr < a.n) {
        float mean = 0.f, rstd = 1.f;
        if (a.has_ln) warp_row_stats(qr, a.h, a.ln_eps, mean, rstd);
        for (int i = lane; i < a.h; i += 32) {
          float v = qr[i];
          if (a.has_ln) {
            v = (v - mean) * rstd;
            if (a.lnq_w) v *= __ldg(a.lnq_w + i);
            if (a.lnq_b) v += __ldg(a.lnq_b + i);
          }
          v *= a.beta;
          if (a.qsave && blockIdx.y == 0) a.qsave[(size_t)gr * a.h + i] = v;
          qr[i] = v;
        }
      }
      for (int i = h4 + lane; i < L.hs; i += 32) qr[i] = 0.f;    // stride padding
    }
  } else {
    load_tile(Qs, a.qs, row0, BM, a.n, a.h, h4, L.hs);
    if (a.qsave && blockIdx.y == 0) {
      for (int idx = tid; idx < BM * a.h; idx += NT) {
        const int r = idx / a.h, c = idx - r * a.h;
        if (row0 + r < a.n) a.qsave[(size_t)(row0 + r) * a.h + c] = a.qs[(size_t)(row0 + r) * a.h + c];
      }
    }
  }

  float m_run[RM], l_run[RM];
  int arg_run[RM];
  float o[RM][DVC][4];
#pragma unroll
  for (int i = 0; i < RM; ++i) {
    m_run[i] = -INFINITY;
    l_run[i] = 0.f;
    arg_run[i] = 0;
#pragma unroll
    for (int c = 0; c < DVC; ++c)
#pragma unroll
      for (int e = 0; e < 4; ++e) o[i][c][e] = 0.f;
  }

  for (int t = t_begin; t < t_end; ++t) {
    const int p0 = t * BP;
    __syncthreads();  // previous tile's consumers are done with their K / read-out / P tiles (and Qs is written)
    if (L.nbuf == 2) {
      if (t + 1 < t_end) { issue(t + 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
      Ks = K0 + ((t - t_begin) & 1) * L.k_elems;
      Bs = B0 + ((t - t_begin) & 1) * L.b_elems;
    } else {
      issue(t);
      cp_async_wait<0>();
    }
    __syncthreads();

    // ---- phase 1: scores  s[i][b] = <q_row, k_col> ----
    float s[RM][4];
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
      for (int b = 0; b < 4; ++b) s[i][b] = 0.f;
#pragma unroll(HF ? 4 : 1)
    for (int k = 0; k < h4; k += 4) {
      float4 q[RM];
#pragma unroll
      for (int i = 0; i < RM; ++i) q[i] = *reinterpret_cast<const float4*>(Qs + (ty * RM + i) * hs + k);
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const float4 kk = *reinterpret_cast<const float4*>(Ks + (tx + 16 * b) * hs + k);
#pragma unroll
        for (int i = 0; i < RM; ++i) {
          s[i][b] = fmaf(q[i].x, kk.x, s[i][b]);
          s[i][b] = fmaf(q[i].y, kk.y, s[i][b]);
          s[i][b] = fmaf(q[i].z, kk.z, s[i][b]);
          s[i][b] = fmaf(q[i].w, kk.w, s[i][b]);
        }
      }
    }

    // ---- phase 2: online softmax over this tile ----
#pragma unroll
    for (int i = 0; i < RM; ++i) {
      float tmax = -INFINITY;
      int targ = 0x7fffffff;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int col = p0 + tx + 16 * b;
        if (col >= a.p) s[i][b] = -INFINITY;
        if (s[i][b] > tmax) {  // strict: lowest column wins ties inside the thread
          tmax = s[i][b];
          targ = col;
        }
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, tmax, off);
        const int oa = __shfl_xor_sync(0xffffffffu, targ, off);
        if (om > tmax || (om == tmax && oa < targ)) {
          tmax = om;
          targ = oa;
        }
      }
      if (tmax > m_run[i]) arg_run[i] = targ;  // strict: earlier tile keeps ties
      const float m_new = fmaxf(m_run[i], tmax);
      const float corr = __expf(m_run[i] - m_new);  // exp(-inf) = 0 on the first tile
      float psum = 0.f;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const float p = __expf(s[i][b] - m_new);
        psum += p;
        Ps[(ty * RM + i) * PS + tx + 16 * b] = p;
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) psum += __shfl_xor_sync(0xffffffffu, psum, off);
      l_run[i] = l_run[i] * corr + psum;
      m_run[i] = m_new;
#pragma unroll
      for (int c = 0; c < DVC; ++c)
#pragma unroll
        for (int e = 0; e < 4; ++e) o[i][c][e] *= corr;
    }
    __syncthreads();

    // ---- phase 3: read-out  o += P . B ----
    for (int j = 0; j < BP; j += 4) {
      float4 pv[RM];
#pragma unroll
      for (int i = 0; i < RM; ++i) pv[i] = *reinterpret_cast<const float4*>(Ps + (ty * RM + i) * PS + j);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
#pragma unroll
        for (int c = 0; c < DVC; ++c) {
          const float4 v = *reinterpret_cast<const float4*>(Bs + (j + jj) * L.ds + 4 * tx + 64 * c);
#pragma unroll
          for (int i = 0; i < RM; ++i) {
            const float p = jj == 0 ? pv[i].x : jj == 1 ? pv[i].y : jj == 2 ? pv[i].z : pv[i].w;
            o[i][c][0] = fmaf(p, v.x, o[i][c][0]);
            o[i][c][1] = fmaf(p, v.y, o[i][c][1]);
            o[i][c][2] = fmaf(p, v.z, o[i][c][2]);
            o[i][c][3] = fmaf(p, v.w, o[i][c][3]);
          }
        }
      }
    }
  }

  // ---- epilogue ----
  __syncthreads();
  Ks = K0;
  Bs = B0;
  // out-projection operand: when one K-tile row holds a whole Wo row, the first 64 rows of Wo start loading now
  // (the K tiles are free) and arrive under the merge below
  const int kc = h4 < L.ds ? h4 : L.ds;          // contraction chunk that fits a K-tile row
  const bool wo_async = a.wo != nullptr && kc == a.dr && (a.dr & 3) == 0;
  if (wo_async) {
    load_tile_async(Ks, a.wo, 0, BP, a.dout, a.dr, kc, L.hs);
    cp_async_commit();
  }
  float* Os = Bs;  // [BM][ds]: read-out rows for the merge / the fused out-projection
  int rows_lo = 0, rows_n = BM;   // rows of the tile this CTA finishes
  if (splits > 1) {
    // un-normalised partials -> own shared memory, then merge this CTA's share of the rows across the cluster
    float* St = smem + L.st_off;
#pragma unroll
    for (int i = 0; i < RM; ++i) {
      const int r = ty * RM + i;
#pragma unroll
      for (int c = 0; c < DVC; ++c)
        *reinterpret_cast<float4*>(Os + r * L.ds + 4 * tx + 64 * c) =
            make_float4(o[i][c][0], o[i][c][1], o[i][c][2], o[i][c][3]);
      if (tx == 0) {
        St[r] = m_run[i];
        St[BM + r] = l_run[i];
        St[2 * BM + r] = __int_as_float(arg_run[i]);
      }
    }
    cg::cluster_group cluster = cg::this_cluster();
    cluster.sync();
    rows_n = BM / splits;
    rows_lo = (int)blockIdx.y * rows_n;
    float* Ms = Ps;   // merged rows [rows_n][ds] (the probability tile is free now)
    float* Mst = Ps + rows_n * L.ds;   // merged (scale per split ..., lse, max, arg) scratch: [rows_n][splits + 3]
    const int sw = splits + 3;
    for (int r = tid; r < rows_n; r += NT) {
      float M = -INFINITY;
      int arg = 0;
      for (int sp = 0; sp < splits; ++sp) {
        const float* pst = cluster.map_shared_rank(St, sp);
        const float ms = pst[rows_lo + r];
        if (ms > M) { M = ms; arg = __float_as_int(pst[2 * BM + rows_lo + r]); }   // strict: lower range keeps ties
      }
      float Lsum = 0.f;
      for (int sp = 0; sp < splits; ++sp) {
        const float* pst = cluster.map_shared_rank(St, sp);
        const float w = __expf(pst[rows_lo + r] - M);     // exp(-inf) = 0 for an empty range
        Mst[r * sw + sp] = w;
        Lsum += w * pst[BM + rows_lo + r];
      }
      const float inv = 1.f / Lsum;
      for (int sp = 0; sp < splits; ++sp) Mst[r * sw + sp] *= inv;
      Mst[r * sw + splits] = M + __logf(Lsum);
      Mst[r * sw + splits + 1] = M;
      Mst[r * sw + splits + 2] = __int_as_float(arg);
    }
    __syncthreads();
    for (int idx = tid; idx < rows_n * (L.ds / 4); idx += NT) {
      const int r = idx / (L.ds / 4), c4 = (idx - r * (L.ds / 4)) * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int sp = 0; sp < splits; ++sp) {
        const float* pos = cluster.map_shared_rank(Os, sp);
        const float4 v = *reinterpret_cast<const float4*>(pos + (rows_lo + r) * L.ds + c4);
        const float w = Mst[r * sw + sp];
        acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y); acc.z = fmaf(w, v.z, acc.z); acc.w = fmaf(w, v.w, acc.w);
      }
      *reinterpret_cast<float4*>(Ms + r * L.ds + c4) = acc;
      const int gr = row0 + rows_lo + r;
      if (a.res != nullptr && gr < a.n) {
        const float vv[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (c4 + e < a.dr) a.res[(size_t)gr * a.dr + c4 + e] = vv[e] * a.out_scale;
      }
    }
    for (int r = tid; r < rows_n; r += NT) {
      const int gr = row0 + rows_lo + r;
      if (gr < a.n) {
        if (a.lse) a.lse[gr] = Mst[r * sw + splits];
        if (a.rowmax) a.rowmax[gr] = Mst[r * sw + splits + 1];
        if (a.argmax) a.argmax[gr] = __float_as_int(Mst[r * sw + splits + 2]) + a.pattern_offset;
      }
    }
    cluster.sync();   // peers are done reading this CTA's partials (and Ms is complete for the out-projection)
    Os = Ms;
  } else {
#pragma unroll
    for (int i = 0; i < RM; ++i) {
      const int r = ty * RM + i, gr = row0 + r;
      const float inv = 1.f / l_run[i];
#pragma unroll
      for (int c = 0; c < DVC; ++c) {
        float4 v = make_float4(o[i][c][0] * inv, o[i][c][1] * inv, o[i][c][2] * inv, o[i][c][3] * inv);
        *reinterpret_cast<float4*>(Os + r * L.ds + 4 * tx + 64 * c) = v;
        if (a.res != nullptr && gr < a.n) {
          const int col = 4 * tx + 64 * c;
          const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (col + e < a.dr) a.res[(size_t)gr * a.dr + col + e] = vv[e] * a.out_scale;
        }
      }
      if (tx == 0 && gr < a.n) {
        if (a.lse) a.lse[gr] = m_run[i] + __logf(l_run[i]);
        if (a.argmax) a.argmax[gr] = arg_run[i] + a.pattern_offset;
        if (a.rowmax) a.rowmax[gr] = m_run[i];
      }
    }
    __syncthreads();
  }
  if (a.wo != nullptr) {
    // fused out-projection of the rows this CTA finished (Os row r <-> tile row rows_lo + r):
    // out = Os . Wo^T is the score contraction again, with 64-row chunks of Wo staged in the K tile
    for (int oc0 = 0; oc0 < a.dout; oc0 += BP) {
      float acc[RM][4];
#pragma unroll
      for (int i = 0; i < RM; ++i)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[i][b] = 0.f;
      for (int j0 = 0; j0 < a.dr; j0 += kc) {
        __syncthreads();
        if (wo_async) {
          if (oc0 > 0) {
            load_tile_async(Ks, a.wo, oc0, BP, a.dout, a.dr, kc, L.hs);
            cp_async_commit();
          }
          cp_async_wait<0>();
        } else {
          for (int idx = tid; idx < BP * kc; idx += NT) {
            const int oc = idx / kc, j = idx - oc * kc;
            float v = 0.f;
            if (oc0 + oc < a.dout && j0 + j < a.dr) v = __ldg(a.wo + (size_t)(oc0 + oc) * a.dr + j0 + j);
            Ks[oc * L.hs + j] = v;
          }
        }
        __syncthreads();
        if (ty * RM < rows_n) {
#pragma unroll(HF ? 4 : 1)
          for (int k = 0; k < kc; k += 4) {
            float4 ov[RM];
#pragma unroll
            for (int i = 0; i < RM; ++i) ov[i] = *reinterpret_cast<const float4*>(Os + (ty * RM + i) * L.ds + j0 + k);
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              const float4 ww = *reinterpret_cast<const float4*>(Ks + (tx + 16 * b) * hs + k);
#pragma unroll
              for (int i = 0; i < RM; ++i) {
                acc[i][b] = fmaf(ov[i].x, ww.x, acc[i][b]);
                acc[i][b] = fmaf(ov[i].y, ww.y, acc[i][b]);
                acc[i][b] = fmaf(ov[i].z, ww.z, acc[i][b]);
                acc[i][b] = fmaf(ov[i].w, ww.w, acc[i][b]);
              }
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < RM; ++i) {
        const int r = ty * RM + i, gr = row0 + rows_lo + r;
        if (r >= rows_n || gr >= a.n) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int oc = oc0 + tx + 16 * b;
          if (oc < a.dout) a.out[(size_t)gr * a.dout + oc] = acc[i][b] + (a.bo ? __ldg(a.bo + oc) : 0.f);
        }
      }
    }
  }
}

// ------------------------------------------------------------------ tiny-width path (H, Dr <= 4)
// Hop-VAE's `index_to_embedding` layer retrieves over 3-dimensional patterns (HopVAE.py:164-170; all of BASELINE
// config 4 is this layer): a 64 x 64 tile wastes 15/16 of its read-out lanes on padding.  Here all P patterns
// (K^ and the read-out matrix, padded to float4) sit in shared memory, four lanes share a query row -- lane q
// scans patterns q, q+4, ... -- and combine with quad shuffles.  Two passes over the scores (max, then exp /
// sum / read-out) instead of an online rescale: with H <= 4 recomputing a score costs less than the dependent
// rescale chain.  LN_q + beta in the prologue, out-projection in the epilogue, like the tiled kernel.
//
// LPR lanes share a query row (4 | 8 | 16 | 32): at N = 2,048 a quad per row gives 64 CTAs of four warps, each
// lane walking 128 patterns twice -- one warp per scheduler on 64 of 148 SMs, pure latency.  More lanes per row
// shorten the serial scan and multiply the CTA count; the launcher picks the smallest LPR that fills the GPU.
constexpr int kTinyThreads = 128;

template <int LPR>
__global__ void __launch_bounds__(kTinyThreads) hop_fwd_tiny_kernel(SimtPassArgs a) {
  constexpr int kTinyRows = kTinyThreads / LPR;
  extern __shared__ __align__(16) float smem[];
  float4* Ks = reinterpret_cast<float4*>(smem);                 // [P]
  float4* Bs = Ks + a.p;                                        // [P]
  float4* Ws = Bs + a.p;                                        // [dout] rows of Wo (dr <= 4 columns), if any
  float* bs = reinterpret_cast<float*>(Ws + (a.wo ? a.dout : 0));   // [dout] bias
  const int tid = threadIdx.x;
  if (a.stopped != nullptr && *a.stopped != 0u) {      // stop rule fired: identity pass (see hop_fwd_simt_kernel)
    const long long r = (long long)blockIdx.x * kTinyRows + tid / LPR;
    if (r < a.n && (tid & (LPR - 1)) == 0 && a.res != nullptr && a.qs != nullptr)
      for (int c = 0; c < a.h; ++c) a.res[r * a.dr + c] = a.qs[r * a.h + c];
    return;
  }
  for (int j = tid; j < a.p; j += kTinyThreads) {
    float kv[4] = {0.f, 0.f, 0.f, 0.f}, bv[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < a.h; ++c) kv[c] = __ldg(a.khat + (size_t)j * a.h + c);
    for (int c = 0; c < a.dr; ++c) bv[c] = __ldg(a.bmat + (size_t)j * a.dr + c);
    Ks[j] = make_float4(kv[0], kv[1], kv[2], kv[3]);
    Bs[j] = make_float4(bv[0], bv[1], bv[2], bv[3]);
  }
  if (a.wo) {
    for (int oc = tid; oc < a.dout; oc += kTinyThreads) {
      float wv[4] = {0.f, 0.f, 0.f, 0.f};
      for (int c = 0; c < a.dr; ++c) wv[c] = __ldg(a.wo + (size_t)oc * a.dr + c);
      Ws[oc] = make_float4(wv[0], wv[1], wv[2], wv[3]);
      bs[oc] = a.bo ? __ldg(a.bo + oc) : 0.f;
    }
  }
  __syncthreads();

  const int quad = tid & (LPR - 1);           // lane of the row group
  const long long gr = (long long)blockIdx.x * kTinyRows + tid / LPR;
  const bool ok = gr < a.n;
  // ---- query row: LN_q + beta (every lane of the row group redundantly; H <= 4)
  float q[4] = {0.f, 0.f, 0.f, 0.f};
  if (ok) {
    const float* src = (a.x_raw ? a.x_raw : a.qs) + gr * a.h;
    for (int c = 0; c < a.h; ++c) q[c] = __ldg(src + c);
    if (a.x_raw) {
      if (a.has_ln) {
        float mean = 0.f;
        for (int c = 0; c < a.h; ++c) mean += q[c];
        mean /= (float)a.h;
        float var = 0.f;
        for (int c = 0; c < a.h; ++c) { const float d0 = q[c] - mean; var += d0 * d0; }
        const float rstd = rsqrtf(var / (float)a.h + a.ln_eps);
        for (int c = 0; c < a.h; ++c) {
          q[c] = (q[c] - mean) * rstd;
          if (a.lnq_w) q[c] *= __ldg(a.lnq_w + c);
          if (a.lnq_b) q[c] += __ldg(a.lnq_b + c);
        }
      }
      for (int c = 0; c < a.h; ++c) q[c] *= a.beta;
    }
    if (a.qsave && quad == 0)
      for (int c = 0; c < a.h; ++c) a.qsave[gr * a.h + c] = q[c];
  }
  // ---- pass 1: row maximum and nearest pattern (strict >: the lowest index wins ties inside a lane)
  float m = -INFINITY;
  int arg = 0x7fffffff;
  for (int j = quad; j < a.p; j += LPR) {
    const float4 k = Ks[j];
    const float s = fmaf(q[3], k.w, fmaf(q[2], k.z, fmaf(q[1], k.y, q[0] * k.x)));
    if (s > m) { m = s; arg = j; }
  }
#pragma unroll
  for (int off = 1; off < LPR; off <<= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, off);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, off);
    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
  }
  // ---- pass 2: exp, sum, read-out
  float l = 0.f, acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int j = quad; j < a.p; j += LPR) {
    const float4 k = Ks[j];
    const float s = fmaf(q[3], k.w, fmaf(q[2], k.z, fmaf(q[1], k.y, q[0] * k.x)));
    const float pexp = __expf(s - m);
    const float4 b = Bs[j];
    l += pexp;
    acc[0] = fmaf(pexp, b.x, acc[0]); acc[1] = fmaf(pexp, b.y, acc[1]);
    acc[2] = fmaf(pexp, b.z, acc[2]); acc[3] = fmaf(pexp, b.w, acc[3]);
  }
#pragma unroll
  for (int off = 1; off < LPR; off <<= 1) {
    l += __shfl_xor_sync(0xffffffffu, l, off);
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], off);
  }
  if (!ok) return;
  const float inv = 1.f / l;
  float r[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) r[c] = acc[c] * inv;
  if (quad == 0) {
    if (a.res)
      for (int c = 0; c < a.dr; ++c) a.res[gr * a.dr + c] = r[c] * a.out_scale;
    if (a.lse) a.lse[gr] = m + __logf(l);
    if (a.argmax) a.argmax[gr] = arg + a.pattern_offset;
    if (a.rowmax) a.rowmax[gr] = m;
  }
  if (a.wo) {
    // out-projection: lane `quad` writes output columns quad, quad + LPR, ... (a row group covers 4 LPR consecutive bytes)
    float* dst = a.out + gr * a.dout;
    for (int oc = quad; oc < a.dout; oc += LPR) {
      const float4 w = Ws[oc];
      dst[oc] = fmaf(r[3], w.w, fmaf(r[2], w.z, fmaf(r[1], w.y, fmaf(r[0], w.x, bs[oc]))));
    }
  }
}

template <int LPR>
int launch_tiny_lpr(const SimtPassArgs& a, size_t smem, cudaStream_t st) {
  constexpr int rows = kTinyThreads / LPR;
  if (smem > 48 * 1024)
    HOP_CUDA(cudaFuncSetAttribute(hop_fwd_tiny_kernel<LPR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  hop_fwd_tiny_kernel<LPR><<<(a.n + rows - 1) / rows, kTinyThreads, smem, st>>>(a);
  HOP_LAUNCH_CHECK("hop_fwd_tiny_kernel");
  return HOP_OK;
}

int launch_tiny(const SimtPassArgs& a, cudaStream_t st) {
  const size_t smem = sizeof(float4) * (2 * (size_t)a.p + (a.wo ? a.dout : 0)) + sizeof(float) * (a.wo ? a.dout : 0);
  // smallest row group that still gives every SM two CTAs (each CTA stages all patterns: keep the groups small
  // when there are rows enough).  Measured (B200, P = 512): N = 64: 16.5 / 10.4 / 8.3 us with 4 / 8 / 32 lanes per
  // row; N = 65,536: 81 / 66 / 87 / 135 us with 4 / 8 / 16 / 32.
  int lpr = 8;
  while (lpr < 32 && (long long)a.n * lpr < 2LL * num_sms() * kTinyThreads) lpr <<= 1;
  switch (lpr) {
    case 4: return launch_tiny_lpr<4>(a, smem, st);
    case 8: return launch_tiny_lpr<8>(a, smem, st);
    case 16: return launch_tiny_lpr<16>(a, smem, st);
    default: return launch_tiny_lpr<32>(a, smem, st);
  }
}

template <int RM, int DVC, int HF>
int launch_inst_hf(const SimtPassArgs& a, int splits, cudaStream_t st) {
  constexpr int BM = 16 * RM;
  const SimtSmemLayout L = simt_layout(BM, a.h, DVC);
  const size_t smem = (size_t)L.total * sizeof(float);
  if (smem > 227 * 1024) HOP_FAIL(HOP_ERR_BAD_SHAPE, "fp32 path: tile for H=%d Dr=%d needs %zu B of shared memory", a.h, a.dr, smem);
  // the merged rows + their scratch live in the probability tile: (BM/splits) * (ds + splits + 3) <= BM * PS
  while (splits > 1 && (BM % splits != 0 || (BM / splits) * (L.ds + splits + 3) > BM * PS)) splits >>= 1;
  auto kern = hop_fwd_simt_kernel<RM, DVC, HF>;
  if (smem > 48 * 1024) HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = (a.n + BM - 1) / BM;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid, splits, 1);
  cfg.blockDim = dim3(NT);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = splits; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = splits > 1 ? 1 : 0;
  HOP_CUDA(cudaLaunchKernelEx(&cfg, kern, a, splits));
  HOP_LAUNCH_CHECK("hop_fwd_simt_kernel");
  return HOP_OK;
}

template <int RM, int DVC>
int launch_inst(const SimtPassArgs& a, int splits, cudaStream_t st) {
  if constexpr (RM == 4 && DVC == 1) {
    if (a.h == 64) return launch_inst_hf<RM, DVC, 64>(a, splits, st);
  }
  return launch_inst_hf<RM, DVC, 0>(a, splits, st);
}

template <int RM>
int launch_rm(const SimtPassArgs& a, int splits, cudaStream_t st) {
  const int dvc = (a.dr + 63) / 64;
  switch (dvc) {
    case 1: return launch_inst<RM, 1>(a, splits, st);
    case 2: return launch_inst<RM, 2>(a, splits, st);
    case 3:
    case 4: return launch_inst<RM, 4>(a, splits, st);
    default: HOP_FAIL(HOP_ERR_UNSUPPORTED, "fp32 path supports read-out width <= 256 (got %d)", a.dr);
  }
}

}  // namespace

int launch_simt_pass(const SimtPassArgs& a, cudaStream_t st) {
  if (a.h > 256) HOP_FAIL(HOP_ERR_UNSUPPORTED, "fp32 path supports association width <= 256 (got %d)", a.h);
  if (a.h <= 4 && a.dr <= 4 && a.p <= 4096 && (!a.wo || a.dout <= 1024))
    return launch_tiny(a, st);
  // Rows per CTA: 64-row tiles keep the inner loops FMA-bound (4 x 4 register tile per thread); when the row
  // tiles alone leave SMs idle the pattern axis is split across a cluster (see the header) while everything
  // still fits one wave of resident CTAs and every CTA keeps at least two 64-pattern tiles.
  const int sms = num_sms();
  int rm = 4;
  if (a.h > 128 || a.dr > 128) rm = 2;  // keep the tile under the shared-memory limit
  const int row_tiles = (a.n + 16 * rm - 1) / (16 * rm);
  const int n_tiles = (a.p + BP - 1) / BP;
  // one wave: at most ~3 CTAs of this kernel are resident per SM (registers / shared memory)
  const size_t tile_bytes = (size_t)simt_layout(16 * rm, a.h, (a.dr + 63) / 64).total * sizeof(float);
  int per_sm = (int)((227u * 1024u) / (tile_bytes ? tile_bytes : 1));
  per_sm = per_sm > 3 ? 3 : (per_sm < 1 ? 1 : per_sm);
  const int slots = sms * per_sm;
  int splits = 1;
  while (splits < 8 && n_tiles >= 4 * splits && row_tiles * splits * 2 <= slots) splits <<= 1;
  switch (rm) {
    case 4: return launch_rm<4>(a, splits, st);
    case 2: return launch_rm<2>(a, splits, st);
    default: return launch_rm<1>(a, splits, st);
  }
}

}  // namespace hop
