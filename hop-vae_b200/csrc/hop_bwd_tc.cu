// Association core of the backward on the sm_100a tensor pipe (tcgen05 + TMEM + TMA).
//
// What `loss.backward()` at /root/reference/main.py:35 walks through one Hopfield layer (upstream hflayers
// `hopfield_core_forward`: torch.bmm, softmax, torch.bmm) is five dense contractions over the N x P score block
// (F_bwd = 2 N P (3 D + 2 Dv)); flash-style, nothing N x P is stored:
//     S  = q K^T          A = exp(S - lse)          (recomputed from the saved beta*Q^ and lse)
//     dA = dR V^T         dS = A o (dA - delta)      delta = rowsum(dR o R)
//     dq = dS K^          dK^ = dS^T q               dV^ = A^T dR
// Three launches of ONE kernel template, each a "two score tiles -> element-wise -> accumulate" pipeline in the
// 64-rows-per-CTA pair form of the forward (hop_fwd_tc64_kernel: cta_group::2 with M = 128, an N-wide accumulator
// costs N/2 TMEM columns, so two score tiles, their double buffers AND a 512-wide accumulator fit the 512 columns):
//     MODE_DQ   stationary rows = queries   (q16, dr16)      streamed = patterns (k16, v16)   acc += dS  . K^   (kt16)
//     MODE_DK   stationary rows = patterns  (k16, v16)       streamed = queries  (q16, dr16)  acc += dS^T. q    (qt16)
//     MODE_DV   stationary rows = patterns  (k16)            streamed = queries  (q16)        acc += A^T . dR   (drt16)
// In the pattern-stationary modes the tile is computed transposed (S^T = K^ q^T: TMEM lanes = patterns, columns =
// queries), so the 16-bit weights go back into TMEM as the A operand of the accumulate MMA (TS mode) exactly like P~ in
// the forward; the softmax statistics then belong to the COLUMNS and are staged per tile in shared memory.
// Executed FLOPs: 3 + 3 + 2 = 8 contraction units against 5 algorithmic (S and dA are recomputed per output).
//
// Operands are fp16-encoded (fp32 accumulate): q16 carries beta*Q^ * log2(e) (the score tile is then in log2 units and
// A = 2^(S - lse*log2 e)), dR is scaled by 1/g with g = max|dR| (one device scalar) so that gradients of any magnitude
// sit in fp16's range; the epilogue multiplies g back.  Outputs are accumulated with 16-byte fp32 REDs into zeroed
// buffers (the streamed axis may be split across CTAs to fill the SMs).
#include <cuda.h>

#include <mutex>

#include "hop_common.cuh"
#include "hop_tc_ptx.cuh"

namespace hop {

namespace {

constexpr int kRows = 64;            // stationary rows per CTA (M = 128 over the pair)
constexpr int kTP = 128;             // streamed rows per tile
constexpr int kChunkBytes = 16384;   // one ring stage
constexpr int kSoftWarps = 8;
constexpr int kSoftThreads = kSoftWarps * 32;
constexpr int kThreads = 64 + kSoftThreads;
constexpr uint32_t kTmemCols = 512;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kPadLse = 1.0e30f;   // lse of a padded query row: its weights are 2^(s - 1e30) = 0

enum { MODE_DQ = 0, MODE_DK = 1, MODE_DV = 2 };

template <int H, int MODE>
struct BwdCfg {
  static constexpr bool kTwo = MODE != MODE_DV;            // second score tile (dA) present
  static constexpr bool kColStats = MODE != MODE_DQ;       // softmax statistics belong to the streamed rows (columns)
  static constexpr int kKChunks = H / 64;                  // 64-wide K chunks of the score contractions
  static constexpr int kXChunkBytes = kRows * 128;         // this CTA's 64 rows of one chunk
  static constexpr int kXTileBytes = kKChunks * kXChunkBytes;
  static constexpr int kXBytes = (kTwo ? 2 : 1) * kXTileBytes;
  static constexpr int kSCols = kTP / 2;                   // TMEM columns of one N = 128 score tile
  static constexpr int kBufCols = (kTwo ? 2 : 1) * kSCols; // one buffer: S [, dA]
  // Score buffers in TMEM.  S(j+1) overwrites the buffer whose W~ the accumulate MMAs of an earlier tile read as their A
  // operand, and the tensor pipe does not interlock that read against a later instruction's D write: those MMAs must
  // have COMPLETED before S(j+1) is issued.  With three buffers that is tile j-2 -- long done, no bubble; H = 512 has
  // TMEM room for two, where the issuer waits for tile j-1 (a short bubble per tile).
#ifdef HOP_BWD_BUFS             // (experiment builds)
  static constexpr int kBufs = HOP_BWD_BUFS;
#else
  static constexpr int kBufs = (3 * kBufCols + H / 2 <= 512) ? 3 : 2;
#endif
  static constexpr uint32_t kTmemAcc = kBufs * kBufCols;
  static constexpr int kAccN = H >= 256 ? 256 : H;         // N of one accumulate instruction
  static constexpr int kNH = H / kAccN;                    // instructions per K step
  static constexpr int kAccStageBytes = (kAccN / 2) * 128; // this CTA's half of one (64-k, N) operand slice
  static constexpr int kPC = kTP / 64;                     // 64-row K steps of the accumulate per tile
  static constexpr int kC = kTP / 4;                       // score columns per element-wise thread (32)
  static constexpr int kW = kC / 2;                        // packed words per thread (16)
  static constexpr int kBarBytes = 512;
  static constexpr int kStatBytes = kColStats ? 2 * 2 * kTP * 4 : 0;   // [slot][lse | delta][128]
  static constexpr int kXpBytes = (kW / 4) * 2 * 128 * 16; // [16-byte chunk][sub][lane]
  static constexpr int kRoom = (232448 - kXBytes - kBarBytes - kStatBytes - kXpBytes) / kChunkBytes;
  static constexpr int kStages = kRoom > 10 ? 10 : kRoom;
  static constexpr int kSmemBytes = kXBytes + kStages * kChunkBytes + kBarBytes + kStatBytes + kXpBytes;
  static_assert(H == 128 || H == 256 || H == 512, "widths of the tensor-core backward");
  static_assert(kTmemAcc + kNH * (kAccN / 2) <= 512 && kStages >= 4, "TMEM / shared-memory budget");
};

struct BwdKernelArgs {
  int n_stat, n_stream;          // valid stationary / streamed rows (outputs beyond n_stat are not written)
  int n_stream_valid;            // MODE_DQ: streamed rows (patterns) >= this are padding and get zero weight
  int tiles_per_split;           // streamed tiles (of 128) per CTA pair
  int w;                         // output row pitch (= H)
  const float* row_lse2;         // MODE_DQ: per stationary row  lse * log2(e)   (padded rows: kPadLse)
  const float* row_delta;        // MODE_DQ: per stationary row  delta / g
  const float* col_lse2;         // MODE_DK / DV: per streamed row (padded to the tile)
  const float* col_delta;        // MODE_DK
  const float* gscale;           // [0] = g = max|dR| (device scalar), multiplied back in the epilogue
  float out_scale;               // constant factor of the output (1/log2 e where the operand carried log2 e)
  int use_g;                     // multiply by *gscale
  float* out;                    // (n_stat, w) fp32, zeroed; accumulated with REDs
  float* csum;                   // MODE_DQ: per stationary row, sum_j dS_j (zeroed; RED) -- see bwd_fix_dq_kernel
};

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void ld_shared_v4(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ld_shared_f4(uint32_t addr, float (&v)[4]) {
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "r"(addr) : "memory");
}

// One CTA pair: 2 x 64 stationary rows, all streamed tiles of its split.
//   warp 0     TMA producer: the stationary tiles once, then per streamed tile the score operand(s) (8 KiB chunks, two
//              per ring stage), the column statistics (bulk copies, MODE_DK / DV) and the accumulate operand slices
//   warp 1     MMA issuer (leader CTA, one thread): S [and dA] of tile j+1 into the other TMEM buffer, then
//              acc += W~(j) . Z(j) with W~ read from TMEM
//   warps 2-9  element-wise: TMEM lane L = 32 (warp % 4) + lane holds stationary row L & 63 and tile-column half L >> 6;
//              two threads per lane (sub), 32 columns each.  Every lane of a row needs the row's whole W~ (measured in
//              round 1: the TS-mode MMA reads its A rows lane by lane), so the packed words are swapped with the
//              partner lane L ^ 64 through shared memory, as in hop_fwd_tc64_kernel.
template <int H, int MODE>
__global__ void __launch_bounds__(kThreads, 1)
hop_bwd_tc_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_x2,
                  const __grid_constant__ CUtensorMap tm_y, const __grid_constant__ CUtensorMap tm_y2,
                  const __grid_constant__ CUtensorMap tm_z, const BwdKernelArgs a) {
  using Cfg = BwdCfg<H, MODE>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = ptx::smem_u32(smem_raw);
  if ((smem_base & 1023u) != 0) __trap();
  const uint32_t x_smem = smem_base;
  const uint32_t ring = x_smem + Cfg::kXBytes;
  const uint32_t bars = ring + Cfg::kStages * kChunkBytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (Cfg::kStages + s); };
  const uint32_t misc = bars + 16u * Cfg::kStages;
  auto s_full = [&](int b) { return misc + 8u * b; };                 // (three slots each: Cfg::kBufs <= 3)
  auto p_full = [&](int b) { return misc + 24u + 8u * b; };
  auto p_empty = [&](int b) { return misc + 48u + 8u * b; };
  const uint32_t x_ready = misc + 72u;
  auto st_full = [&](int s) { return misc + 80u + 8u * s; };
  auto st_empty = [&](int s) { return misc + 96u + 8u * s; };
  const uint32_t tmem_slot = misc + 112u;
  constexpr int kBufs = Cfg::kBufs;
  const uint32_t stat = bars + Cfg::kBarBytes;
  auto stat_at = [&](int slot, int which, int col) { return stat + (uint32_t)(((slot * 2 + which) * kTP + col) * 4); };
  const uint32_t xp = stat + Cfg::kStatBytes;
  auto xp_at = [&](int c, int sub, int L) { return xp + (uint32_t)((((c * 2 + sub) * 128) + L) * 16); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();
  constexpr uint32_t kArrivers = kSoftThreads * 2;
  const int cid = (int)blockIdx.x / 2;
  const int nt_total = (a.n_stream + kTP - 1) / kTP;
  const int t0 = (int)blockIdx.y * a.tiles_per_split;
  const int nt = min(a.tiles_per_split, nt_total - t0);
  constexpr int kKStages = Cfg::kKChunks / 2;          // two 8 KiB score-operand chunks per ring stage

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    for (int b = 0; b < kBufs; ++b) {
      ptx::mbar_init(s_full(b), 1);
      ptx::mbar_init(p_full(b), kArrivers); ptx::mbar_init(p_empty(b), 1);
    }
    for (int b = 0; b < 2; ++b) { ptx::mbar_init(st_full(b), 1); ptx::mbar_init(st_empty(b), kSoftThreads); }
    ptx::mbar_init(x_ready, 1);
    ptx::fence_barrier_init();
    ptx::prefetch_tmap(&tm_x); ptx::prefetch_tmap(&tm_y); ptx::prefetch_tmap(&tm_z);
    if constexpr (Cfg::kTwo) { ptx::prefetch_tmap(&tm_x2); ptx::prefetch_tmap(&tm_y2); }
  }
  if (warp == 0) {
    __syncwarp();
    ptx::tmem_alloc_pair(tmem_slot, kTmemCols); ptx::tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (each CTA: its half of every tile)
    if (ptx::elect_one()) {
      // stationary tiles: this CTA's 64 rows, all chunks; the bytes of both CTAs are counted on the leader's barrier
      {
        const uint32_t xb = ptx::map_to_cta(x_ready, 0);
        if (crank == 0) ptx::mbar_arrive_expect_tx(x_ready, 2u * Cfg::kXBytes);
        const int row0 = (cid * 2 + (int)crank) * kRows;
        for (int kc = 0; kc < Cfg::kKChunks; ++kc)
          ptx::tma_load_2d_pair(x_smem + kc * Cfg::kXChunkBytes, &tm_x, xb, kc * 64, row0);
        if constexpr (Cfg::kTwo)
          for (int kc = 0; kc < Cfg::kKChunks; ++kc)
            ptx::tma_load_2d_pair(x_smem + Cfg::kXTileBytes + kc * Cfg::kXChunkBytes, &tm_x2, xb, kc * 64, row0);
      }
      uint32_t it = 0;
      auto acquire = [&](uint32_t total_bytes) {
        const int s = it % Cfg::kStages;
        ptx::mbar_wait(empty_bar(s), ((it / Cfg::kStages) & 1) ^ 1);
        if (crank == 0) ptx::mbar_arrive_expect_tx(full_bar(s), total_bytes);
        return s;
      };
      auto load_scores = [&](int tile) {
        for (int two = 0; two < (Cfg::kTwo ? 2 : 1); ++two) {
          const CUtensorMap* m = two ? &tm_y2 : &tm_y;
          for (int ks = 0; ks < kKStages; ++ks) {
            const int s = acquire(2u * 2u * 8192u);
            const uint32_t fb = ptx::map_to_cta(full_bar(s), 0);
            for (int c = 0; c < 2; ++c)
              ptx::tma_load_2d_pair(ring + s * kChunkBytes + c * 8192, m, fb, (ks * 2 + c) * 64,
                                    tile * kTP + (int)crank * (kTP / 2));
            ++it;
          }
        }
      };
      auto load_stats = [&](int j) {
        if constexpr (Cfg::kColStats) {
          const int slot = j & 1;
          ptx::mbar_wait(st_empty(slot), ((j >> 1) & 1) ^ 1);
          ptx::mbar_arrive_expect_tx(st_full(slot), (MODE == MODE_DK ? 2u : 1u) * kTP * 4u);
          ptx::bulk_load(stat_at(slot, 0, 0), a.col_lse2 + (size_t)(t0 + j) * kTP, kTP * 4, st_full(slot));
          if constexpr (MODE == MODE_DK)
            ptx::bulk_load(stat_at(slot, 1, 0), a.col_delta + (size_t)(t0 + j) * kTP, kTP * 4, st_full(slot));
        }
      };
      auto load_acc = [&](int tile) {
        for (int kc = 0; kc < Cfg::kPC; ++kc)
          for (int nh = 0; nh < Cfg::kNH; ++nh) {
            const int s = acquire(2u * Cfg::kAccStageBytes);
            ptx::tma_load_2d_pair(ring + s * kChunkBytes, &tm_z, ptx::map_to_cta(full_bar(s), 0),
                                  tile * kTP + kc * 64, nh * Cfg::kAccN + (int)crank * (Cfg::kAccN / 2));
            ++it;
          }
      };
      if (nt > 0) {
        load_scores(t0);
        load_stats(0);
        for (int j = 0; j < nt; ++j) {
          if (j + 1 < nt) { load_scores(t0 + j + 1); load_stats(j + 1); }
          load_acc(t0 + j);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer (leader CTA, one thread)
    if (crank == 0 && nt > 0 && ptx::elect_one()) {
      constexpr uint32_t idesc_s = ptx::umma_idesc_f16(128, kTP, 0);
      constexpr uint32_t idesc_acc = ptx::umma_idesc_f16(128, Cfg::kAccN, 0);
      uint32_t it = 0;
      ptx::mbar_wait_cluster(x_ready, 0);
      ptx::tc_fence_after();
      auto issue_scores = [&](int j) {
        const int b = j % kBufs;
        for (int two = 0; two < (Cfg::kTwo ? 2 : 1); ++two)
          for (int ks = 0; ks < kKStages; ++ks) {
            const int s = it % Cfg::kStages;
            ptx::mbar_wait_cluster(full_bar(s), (it / Cfg::kStages) & 1);
            ptx::tc_fence_after();
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              const uint64_t ad = ptx::umma_desc_sw128(x_smem + two * Cfg::kXTileBytes + (ks * 2 + c) * Cfg::kXChunkBytes);
              const uint64_t bd = ptx::umma_desc_sw128(ring + s * kChunkBytes + c * 8192);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                ptx::umma2_f16_ss(tmem_base + b * Cfg::kBufCols + two * Cfg::kSCols, ad + 2 * k, bd + 2 * k, idesc_s,
                                  (ks | c | k) != 0);
            }
            ptx::umma2_commit_both(empty_bar(s));
            ++it;
          }
        ptx::umma2_commit_both(s_full(b));
      };
      issue_scores(0);
      for (int j = 0; j < nt; ++j) {
        if (j + 1 < nt) {
          // the accumulate MMAs that read W~ out of S(j+1)'s buffer (tile j + 1 - kBufs) must have completed -- measured
          // without this: 16-row blocks of dV^ off by 1-4 % from run to run wherever the score operands were already
          // waiting in the ring (see BwdCfg::kBufs)
          if (j + 1 >= kBufs) {
            const int jp = j + 1 - kBufs;
            ptx::mbar_wait_cluster(p_empty(jp % kBufs), (jp / kBufs) & 1);
            ptx::tc_fence_after();
          }
          issue_scores(j + 1);
        }
        const int b = j % kBufs;
        ptx::mbar_wait_cluster(p_full(b), (j / kBufs) & 1);
        ptx::tc_fence_after();
        for (int kc = 0; kc < Cfg::kPC; ++kc)
          for (int nh = 0; nh < Cfg::kNH; ++nh) {
            const int s = it % Cfg::kStages;
            ptx::mbar_wait_cluster(full_bar(s), (it / Cfg::kStages) & 1);
            ptx::tc_fence_after();
            const uint32_t at = tmem_base + b * Cfg::kBufCols + kc * 32;
            const uint64_t bd = ptx::umma_desc_sw128(ring + s * kChunkBytes);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma2_f16_ts(tmem_base + Cfg::kTmemAcc + nh * (Cfg::kAccN / 2), at + 8 * k, bd + 2 * k, idesc_acc,
                                (j | kc | k) != 0);
            ptx::umma2_commit_both(empty_bar(s));
            ++it;
          }
        ptx::umma2_commit_both(p_empty(b));
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------ element-wise + epilogue
    const int wq = warp & 3;                      // TMEM lane quarter
    const int sub = (warp - 2) >> 2;              // which 32 of the lane's 64 score columns
    const int L = wq * 32 + lane;                 // TMEM lane
    const int h = L >> 6, row = L & 63;           // tile-column half, stationary row of this CTA
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    const uint32_t grp_bar = 1 + (wq & 1);        // the four warps that share 32 rows (128 threads)
    const long long grow = ((long long)cid * 2 + (long long)crank) * kRows + row;
    const bool row_ok = grow < a.n_stat;
    const int cbase = h * (kTP / 2) + sub * Cfg::kC;          // first tile column of this thread
    auto arrive_issuer = [&](uint32_t bar) { ptx::mbar_arrive_cluster(ptx::map_to_cta(bar, 0)); };
    float lse_r = kPadLse, del_r = 0.f, cacc = 0.f;
    if constexpr (MODE == MODE_DQ) {
      if (row_ok) { lse_r = __ldg(a.row_lse2 + grow); del_r = __ldg(a.row_delta + grow); }
    }

    for (int j = 0; j < nt; ++j) {
      const int b = j % kBufs;
      ptx::mbar_wait(s_full(b), (j / kBufs) & 1);
      ptx::tc_fence_after();
      uint32_t v[Cfg::kC], u[Cfg::kTwo ? Cfg::kC : 1];
      ptx::tmem_ld_x32p(lane_base + b * Cfg::kBufCols + sub * Cfg::kC, v);
      if constexpr (Cfg::kTwo) ptx::tmem_ld_x32p(lane_base + b * Cfg::kBufCols + Cfg::kSCols + sub * Cfg::kC, u);
      ptx::tmem_wait_ld();
      const int slot = j & 1;
      if constexpr (Cfg::kColStats) ptx::mbar_wait(st_full(slot), (j >> 1) & 1);
      // MODE_DQ: padded patterns have zero K^ rows, i.e. score 0 and weight 2^-lse != 0: they must not enter the row sum
      // that defines the consistent delta (padded QUERIES carry lse = 1e30 and vanish by themselves)
      const int valid = MODE == MODE_DQ ? a.n_stream_valid - ((t0 + j) * kTP + cbase) : Cfg::kC;
      uint32_t pk[Cfg::kW];
#pragma unroll
      for (int i4 = 0; i4 < Cfg::kC / 4; ++i4) {
        float l4[4], d4[4];
        if constexpr (Cfg::kColStats) {
          ld_shared_f4(stat_at(slot, 0, cbase + 4 * i4), l4);
          if constexpr (MODE == MODE_DK) ld_shared_f4(stat_at(slot, 1, cbase + 4 * i4), d4);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) { l4[e] = lse_r; d4[e] = del_r; }
        }
        float w4[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float p = ptx::ex2(__uint_as_float(v[4 * i4 + e]) - l4[e]);
          if constexpr (Cfg::kTwo) w4[e] = p * (__uint_as_float(u[4 * i4 + e]) - d4[e]);
          else w4[e] = p;
          if (MODE == MODE_DQ && 4 * i4 + e >= valid) w4[e] = 0.f;
        }
        if constexpr (MODE == MODE_DQ) cacc += (w4[0] + w4[1]) + (w4[2] + w4[3]);
        pk[2 * i4] = pack_h2(w4[0], w4[1]);
        pk[2 * i4 + 1] = pack_h2(w4[2], w4[3]);
      }
      if constexpr (Cfg::kColStats) ptx::mbar_arrive(st_empty(slot));
      ptx::named_bar_sync(grp_bar, 128);           // the partner lanes have read the previous tile's words
#pragma unroll
      for (int c = 0; c < Cfg::kW / 4; ++c) st_shared_v4(xp_at(c, sub, L), pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      ptx::named_bar_sync(grp_bar, 128);           // ... and every S / dA column of these rows is in registers
      uint32_t pw[Cfg::kW];
#pragma unroll
      for (int c = 0; c < Cfg::kW / 4; ++c) ld_shared_v4(xp_at(c, sub, L ^ 64), pw[4 * c], pw[4 * c + 1], pw[4 * c + 2], pw[4 * c + 3]);
      // W~ row of this lane: tile column (TP/2) h' + 32 sub' + i -> TMEM column (TP/4) h' + 16 sub' + i / 2, over the S tile
      const uint32_t p_own = lane_base + b * Cfg::kBufCols + h * (kTP / 4) + sub * Cfg::kW;
      const uint32_t p_oth = lane_base + b * Cfg::kBufCols + (h ^ 1) * (kTP / 4) + sub * Cfg::kW;
      ptx::tmem_st_x16(p_own, pk);
      ptx::tmem_st_x16(p_oth, pw);
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      arrive_issuer(p_full(b));
    }

    if constexpr (MODE == MODE_DQ) {
      if (row_ok && nt > 0 && a.csum) atomicAdd(a.csum + grow, cacc);
    }
    if (nt > 0) {
      const uint32_t last = nt - 1;
      ptx::mbar_wait(p_empty(last % kBufs), (last / kBufs) & 1);
      ptx::tc_fence_after();
      float osc = a.out_scale;
      if (a.use_g) osc *= __ldg(a.gscale);
      constexpr int kOGroups = H / 4 / 32;          // 32-column TMEM groups of the accumulator per thread
      const int og0 = sub * kOGroups;
#pragma unroll 1
      for (int c = og0; c < og0 + kOGroups; ++c) {
        uint32_t o[32];
        ptx::tmem_ld_x32(lane_base + Cfg::kTmemAcc + c * 32, o);
        ptx::tmem_wait_ld();
        if (row_ok) {
          const int tcol = c * 32;                  // TMEM column: instruction nh, column `within` of its lane half
          const int nh = tcol / (Cfg::kAccN / 2), within = tcol - nh * (Cfg::kAccN / 2);
          float* dst = a.out + grow * a.w + nh * Cfg::kAccN + h * (Cfg::kAccN / 2) + within;
#pragma unroll
          for (int q4 = 0; q4 < 8; ++q4)
            atomicAdd(reinterpret_cast<float4*>(dst) + q4,
                      make_float4(__uint_as_float(o[4 * q4]) * osc, __uint_as_float(o[4 * q4 + 1]) * osc,
                                  __uint_as_float(o[4 * q4 + 2]) * osc, __uint_as_float(o[4 * q4 + 3]) * osc));
        }
      }
      // The REDs above must be PERFORMED before this CTA retires: without the fence the kernels that read the result right
      // behind this launch (absmax / pack / column sums of dV^, the dq fix-up) saw run-to-run differences of 1e-4 .. 1e-3
      // wherever the grid ran in more than one wave or several CTAs added into the same rows (measured with
      // scripts/probes/bwd_determinism.py; 1e-7 with the fence)
#ifndef HOP_BWD_NO_FENCE        // (experiment builds)
      __threadfence();
#endif
      ptx::tc_fence_before();
    }
  }

  __syncthreads();
  ptx::cluster_sync();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------ operand packs of the row side
// delta = rowsum(dR o R), lse2 = lse log2(e) (padded rows: kPadLse), g = max |dR| (atomic max on the float's bits)
__global__ void __launch_bounds__(256) bwd_stats_kernel(const float* __restrict__ dr, const float* __restrict__ r,
                                                        const float* __restrict__ lse, int n, int npad, int dv,
                                                        float* __restrict__ delta, float* __restrict__ lse2,
                                                        float* __restrict__ gmax) {
  const int row = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= npad) return;
  if (row >= n) {
    if (lane == 0) { delta[row] = 0.f; lse2[row] = kPadLse; }
    return;
  }
  float s = 0.f, m = 0.f;
  for (int c = lane * 4; c < dv; c += 128) {
    const float4 g4 = __ldg(reinterpret_cast<const float4*>(dr + (size_t)row * dv + c));
    const float4 r4 = __ldg(reinterpret_cast<const float4*>(r + (size_t)row * dv + c));
    s = fmaf(g4.x, r4.x, fmaf(g4.y, r4.y, fmaf(g4.z, r4.z, fmaf(g4.w, r4.w, s))));
    m = fmaxf(m, fmaxf(fmaxf(fabsf(g4.x), fabsf(g4.y)), fmaxf(fabsf(g4.z), fabsf(g4.w))));
  }
  s = warp_sum(s);
  m = warp_max(m);
  if (lane == 0) {
    delta[row] = s;
    lse2[row] = lse[row] * kLog2e;
    if (m > 0.f) atomicMax(reinterpret_cast<int*>(gmax), __float_as_int(m));
  }
}

// (rows, w) fp32 -> 16-bit row-major (rows_pad, w) and transposed (w, rows_pad), each with its own scale; rows >= rows
// are written as zeros.  32 x 32 tiles through shared memory, both layouts written coalesced.
struct PackJob {
  const float* src; int rows, rows_pad, w;
  __half* rm; float rm_scale;            // row-major pack and its constant scale
  __half* tr; float tr_scale;            // transposed pack
  int by_g;                              // additionally divide both by *g
};
__global__ void __launch_bounds__(256) bwd_pack_kernel(PackJob j0, PackJob j1, const float* __restrict__ g,
                                                       const float* __restrict__ delta, float* __restrict__ deltag,
                                                       int npad) {
  __shared__ float tile[32][33];
  const PackJob& j = blockIdx.z == 0 ? j0 : j1;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  if (c0 >= j.w) return;
  const float gv = __ldg(g);
  const float inv_g = gv > 0.f ? 1.f / gv : 0.f;
  const float extra = j.by_g ? inv_g : 1.f;
  if (blockIdx.z == 1 && blockIdx.x == 0 && ty == 0 && r0 + tx < npad) deltag[r0 + tx] = delta[r0 + tx] * inv_g;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = r0 + ty + 8 * i, c = c0 + tx;
    float v = 0.f;
    if (r < j.rows && c < j.w) v = __ldg(j.src + (size_t)r * j.w + c);
    tile[ty + 8 * i][tx] = v * extra;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = r0 + ty + 8 * i, c = c0 + tx;
    if (r < j.rows_pad && c < j.w) j.rm[(size_t)r * j.w + c] = __float2half_rn(tile[ty + 8 * i][tx] * j.rm_scale);
    const int ct = c0 + ty + 8 * i, rt = r0 + tx;                  // transposed: rows along threadIdx.x
    if (ct < j.w && rt < j.rows_pad) j.tr[(size_t)ct * j.rows_pad + rt] = __float2half_rn(tile[tx][ty + 8 * i] * j.tr_scale);
  }
}

// Softmax-Jacobian consistency.  dS_j = A_j (dA_j - delta) is a difference of nearly equal numbers on a saturated row
// (A ~ one-hot: dA_top ~ delta), and delta = rowsum(dR o R) is exact fp32 while dA comes from fp16-rounded operands: the
// mismatch (2^-11 |dA|) becomes a common-mode error c_r = sum_j dS_j != 0 that swamps the (tiny) true gradient --
// measured: d_x off by 3e-2 at w = 0.999, by 10x at D = 512.  The DQ launch accumulates c_r (one fp32 sum per thread, one
// RED per row); the delta that is consistent with the rounded dA is delta' = delta + c_r, and the consistent gradient
//     dq* = dq - c_r sum_j A_j k_j  ~=  dq - c_r K^[nearest pattern]
// -- the read-out A.K^ only multiplies the rounding-sized c_r, so its leading term (the forward's nearest pattern) is
// enough: the residual is c_r (1 - w) |k|, second order.  dK^ then runs with delta'.  One warp per query row.
__global__ void __launch_bounds__(256) bwd_fix_dq_kernel(float* __restrict__ dqs, const float* __restrict__ khat,
                                                         const int32_t* __restrict__ argmax, const float* __restrict__ csum,
                                                         const float* __restrict__ gmax, float* __restrict__ deltag,
                                                         int n, int h) {
  const int row = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= n) return;
  const float c = csum[row];
  const float cg = c * __ldg(gmax);
  const float* k = khat + (size_t)argmax[row] * h;
  float* q = dqs + (size_t)row * h;
  for (int i = lane * 4; i < h; i += 128) {
    const float4 kv = __ldg(reinterpret_cast<const float4*>(k + i));
    float4 v = *reinterpret_cast<float4*>(q + i);
    v.x = fmaf(-cg, kv.x, v.x); v.y = fmaf(-cg, kv.y, v.y); v.z = fmaf(-cg, kv.z, v.z); v.w = fmaf(-cg, kv.w, v.w);
    *reinterpret_cast<float4*>(q + i) = v;
  }
  if (lane == 0) deltag[row] += c;
}

// ------------------------------------------------------------------ host side
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn bwd_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}
// 2-D row-major fp16 matrix (rows, cols) -> boxes of (box_rows x 64 columns), 128-byte swizzle
int bwd_map(CUtensorMap* m, const void* base, int rows, int cols, int box_rows) {
  EncodeTiledFn fn = bwd_encode_fn();
  if (!fn) HOP_FAIL(HOP_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) HOP_FAIL(HOP_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for a %d x %d matrix", (int)r, rows, cols);
  return HOP_OK;
}

struct TcBwdWs {
  __half *q16, *qt16, *dr16, *drt16;
  float *lse2, *delta, *deltag, *csum, *gmax;
  size_t total;
};
TcBwdWs carve_bwd(void* ws, int n, int h, int dv) {
  TcBwdWs w{};
  size_t off = 0;
  auto take = [&](size_t bytes) {
    char* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += round_up_sz(bytes, 256);
    return p;
  };
  const size_t npad = (size_t)round_up(n, kTP);
  w.q16 = reinterpret_cast<__half*>(take(npad * h * 2));
  w.qt16 = reinterpret_cast<__half*>(take(npad * h * 2));
  w.dr16 = reinterpret_cast<__half*>(take(npad * dv * 2));
  w.drt16 = reinterpret_cast<__half*>(take(npad * dv * 2));
  w.lse2 = reinterpret_cast<float*>(take(npad * 4));
  w.delta = reinterpret_cast<float*>(take(npad * 4));
  w.deltag = reinterpret_cast<float*>(take(npad * 4));
  w.csum = reinterpret_cast<float*>(take(npad * 4));
  w.gmax = reinterpret_cast<float*>(take(256));
  w.total = off;
  return w;
}

// streamed-axis splits: the smallest count that fills >= 90 % of the last wave of CTA pairs, >= 8 tiles each
int bwd_splits(int pairs, int tiles) {
#ifdef HOP_BWD_SPLITS           // (experiment builds)
  return HOP_BWD_SPLITS < tiles ? HOP_BWD_SPLITS : tiles;
#endif
  const int slots = num_sms() / 2;
  int best = 1;
  double best_eff = 0.0;
  for (int s = 1; s <= 16; ++s) {
    if (s > 1 && tiles / s < 8) break;
    const long long items = (long long)pairs * s;
    const long long waves = (items + slots - 1) / slots;
    const double eff = (double)items / (double)(waves * slots);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
    if (eff >= 0.9) break;
  }
  return best;
}

template <int H, int MODE>
int launch_mode(const CUtensorMap& tx, const CUtensorMap& tx2, const CUtensorMap& ty, const CUtensorMap& ty2,
                const CUtensorMap& tz, BwdKernelArgs k, int stat_rows_pad, cudaStream_t st) {
  using Cfg = BwdCfg<H, MODE>;
  auto* kern = hop_bwd_tc_kernel<H, MODE>;
  HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  const int pairs = (stat_rows_pad / kRows + 1) / 2;
  const int tiles = (k.n_stream + kTP - 1) / kTP;
  const int splits = bwd_splits(pairs, tiles);
  k.tiles_per_split = (tiles + splits - 1) / splits;
  const int gy = (tiles + k.tiles_per_split - 1) / k.tiles_per_split;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(pairs * 2, gy, 1);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  HOP_CUDA(cudaLaunchKernelEx(&cfg, kern, tx, tx2, ty, ty2, tz, k));
  HOP_LAUNCH_CHECK("hop_bwd_tc_kernel");
  return HOP_OK;
}

template <int H>
int launch_core_tc(const TcBwdArgs& a, const TcBwdWs& w, cudaStream_t st) {
  using CfgQ = BwdCfg<H, MODE_DQ>;
  const int npad = round_up(a.n, kTP), ppad = padded_patterns(a.p);
  CUtensorMap mq, mdr, mk, mv, mkt, mqt, mdrt;
  HOP_TRY(bwd_map(&mq, w.q16, npad, H, 64));
  HOP_TRY(bwd_map(&mdr, w.dr16, npad, H, 64));
  HOP_TRY(bwd_map(&mk, a.k16, ppad, H, 64));
  HOP_TRY(bwd_map(&mv, a.v16, ppad, H, 64));
  HOP_TRY(bwd_map(&mkt, a.kt16, H, ppad, CfgQ::kAccN / 2));
  HOP_TRY(bwd_map(&mqt, w.qt16, H, npad, CfgQ::kAccN / 2));
  HOP_TRY(bwd_map(&mdrt, w.drt16, H, npad, CfgQ::kAccN / 2));
  BwdKernelArgs k{};
  k.w = H; k.gscale = w.gmax; k.out_scale = 1.f; k.use_g = 1;
  // dq: queries stationary, patterns streamed
  k.n_stat = a.n; k.n_stream = ppad; k.n_stream_valid = a.p; k.row_lse2 = w.lse2; k.row_delta = w.deltag; k.out = a.dqs; k.csum = w.csum;
  HOP_TRY((launch_mode<H, MODE_DQ>(mq, mdr, mk, mv, mkt, k, npad, st)));
  k.csum = nullptr;
  // consistent delta for the dK^ launch, common-mode correction of dq (pattern_offset-free indices of this call)
  bwd_fix_dq_kernel<<<(a.n + 7) / 8, 256, 0, st>>>(a.dqs, a.khat, a.argmax, w.csum, w.gmax, w.deltag, a.n, H);
  HOP_LAUNCH_CHECK("bwd_fix_dq_kernel");
  // dK^ and dV^: patterns stationary, queries streamed
  k.n_stat = a.p; k.n_stream = npad; k.row_lse2 = nullptr; k.row_delta = nullptr;
  k.col_lse2 = w.lse2; k.col_delta = w.deltag; k.out = a.dkhat;
  HOP_TRY((launch_mode<H, MODE_DK>(mk, mv, mq, mdr, mqt, k, ppad, st)));
  k.out = a.dvhat; k.col_delta = nullptr;
  HOP_TRY((launch_mode<H, MODE_DV>(mk, mk, mq, mq, mdrt, k, ppad, st)));
  return HOP_OK;
}

}  // namespace

// The tensor-core core applies where the fused forward does: equal widths the pair form is built for and an N x P
// block large enough that five dense contractions outweigh the pack launches.
bool tc_backward_supported(int n, int p, int h, int dv) {
  return h == dv && (h == 128 || h == 256 || h == 512) && (long long)n * p >= (1ll << 25);
}

size_t tc_backward_workspace(int n, int p, int h, int dv) {
  (void)p;
  return carve_bwd(nullptr, n, h, dv).total;
}

// dqs / dkhat / dvhat must be zeroed by the caller (they are accumulated with REDs).
int launch_tc_backward(const TcBwdArgs& a, cudaStream_t st) {
  if (!tc_backward_supported(a.n, a.p, a.h, a.dv)) HOP_FAIL(HOP_ERR_UNSUPPORTED, "tensor-core backward: H == Dv in {128, 256, 512}");
  if (!a.k16 || !a.v16 || !a.kt16) HOP_FAIL(HOP_ERR_BAD_ARG, "tensor-core backward needs the k16 / v16 / kt16 packs (HOP_F_TRAIN)");
  if (!a.argmax || !a.khat) HOP_FAIL(HOP_ERR_BAD_ARG, "tensor-core backward needs the forward's nearest-pattern indices and the fp32 K^");
  const size_t need = tc_backward_workspace(a.n, a.p, a.h, a.dv);
  if (!a.ws || a.ws_bytes < need) HOP_FAIL(HOP_ERR_WORKSPACE, "tensor-core backward needs %zu workspace bytes, %zu given", need, a.ws_bytes);
  const TcBwdWs w = carve_bwd(a.ws, a.n, a.h, a.dv);
  const int npad = round_up(a.n, kTP);
  HOP_CUDA(cudaMemsetAsync(w.gmax, 0, 256, st));
  HOP_CUDA(cudaMemsetAsync(w.csum, 0, (size_t)npad * sizeof(float), st));
  bwd_stats_kernel<<<(npad + 7) / 8, 256, 0, st>>>(a.dr, a.r, a.lse, a.n, npad, a.dv, w.delta, w.lse2, w.gmax);
  HOP_LAUNCH_CHECK("bwd_stats_kernel");
  PackJob jq{a.qs, a.n, npad, a.h, w.q16, kLog2e, w.qt16, 1.f, 0};
  PackJob jd{a.dr, a.n, npad, a.dv, w.dr16, 1.f, w.drt16, 1.f, 1};
  const int wmax = a.h > a.dv ? a.h : a.dv;
  bwd_pack_kernel<<<dim3((wmax + 31) / 32, npad / 32, 2), 256, 0, st>>>(jq, jd, w.gmax, w.delta, w.deltag, npad);
  HOP_LAUNCH_CHECK("bwd_pack_kernel");
  switch (a.h) {
    case 128: return launch_core_tc<128>(a, w, st);
    case 256: return launch_core_tc<256>(a, w, st);
    default: return launch_core_tc<512>(a, w, st);
  }
}

}  // namespace hop
