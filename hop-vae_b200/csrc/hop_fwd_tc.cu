// Fused Hopfield retrieval forward on the sm_100a tensor pipe (tcgen05 + TMEM + TMA).
//
// Replaces, in one launch, upstream hflayers `hopfield_core_forward` as reached from
// `HopfieldLayer.forward` (call sites /root/reference/HopVAE.py:193,215,216,234,255,257,271):
// LN_q + beta, score contraction, row softmax, T retrieve updates, pattern read-out.  The N x P
// score matrix lives only in TMEM / registers / shared memory.
//
// One CTA owns 128 query rows for all T+1 passes and, optionally, one of `splits` contiguous
// pattern ranges (flash-decoding style; partials are combined by tc_combine_kernel).
//   warp 0      TMA producer: ring of 16 KiB operand chunks (K^ rows for the score MMA, then
//               V^T / K^T rows for the read-out MMA), 128-byte swizzle, mbarrier-tracked
//   warp 1      MMA issuer (one elected lane): S = Q K^T into a double-buffered TMEM tile,
//               O += P~ B into TMEM columns [256, 256+D) with one N = D instruction per K step
//   warps 2..9  softmax, two warps per TMEM lane quarter (= two per SM sub-partition, so each
//               hides the other's MUFU / TMEM latency): warp pair (q, h) owns rows 32q..32q+31 and
//               score columns 64h..64h+63 of every tile; the pair agrees on the row maximum through
//               shared memory + a 64-thread named barrier, then each half does ex2, the 16-bit P~
//               pack into swizzled shared memory and half of every O-column job (lazy rescale,
//               epilogue, next pass' Q tile).
//
// Two kernels: hop_fwd_tc_kernel (128 query rows per CTA; D <= 256, and D >= 512 under HOP_TC_HALFM=0) and
// hop_fwd_tc64_kernel (64 rows per CTA under M = 128 pair MMAs, whose accumulator layout fits a 512-column
// read-out next to the score tiles; D = 512 and 1024).
#include <cuda.h>

#include <mutex>

#include "hop_common.cuh"
#include "hop_tc_ptx.cuh"

namespace hop {

namespace {

constexpr int kTileM = 128;          // query rows per CTA (TMEM lanes)
constexpr int kTileP = 128;          // patterns per score tile
constexpr int kHalfP = 64;           // score columns per softmax thread
constexpr int kChunkBytes = 16384;   // 128 rows x 64 16-bit elements, 128B-swizzled
constexpr int kSoftmaxWarps = 8;
constexpr int kSoftmaxThreads = kSoftmaxWarps * 32;
constexpr int kThreads = 64 + kSoftmaxThreads;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemO = 256;     // O accumulator starts after the two S buffers
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kRescaleThreshold = 15.f;  // log2 units: P~ <= 2^15 = 32768 < 65504, the fp16 maximum

#ifdef HOP_TC_TRACE
// Debug build only (libhop_sm100_trace.so): per-tile clock64() timestamps of CTA (0,0) for timeline analysis.
constexpr int kTraceTiles = 64, kTraceEvents = 8, kTraceRoles = 9;
__device__ long long g_trace[kTraceRoles * kTraceTiles * kTraceEvents];
#define HOP_TRACE(role, tile, ev)                                                                  \
  do {                                                                                             \
    if (blockIdx.x == 0 && blockIdx.y == 0 && (tile) < kTraceTiles)                                \
      g_trace[((role) * kTraceTiles + (tile)) * kTraceEvents + (ev)] = clock64();                  \
  } while (0)
#else
#define HOP_TRACE(role, tile, ev) do {} while (0)
#endif

// Every kPolyPeriod-th score column takes ex2_poly instead of MUFU (0 = never).  Tunable at build time.
#ifndef HOP_POLY_D256
#define HOP_POLY_D256 4
#endif
#ifndef HOP_POLY_D128
#define HOP_POLY_D128 4
#endif
#ifndef HOP_TC64_STAGES_512
#define HOP_TC64_STAGES_512 8   // ring stages of the 64-row pair kernel at H = 512 (9 still fits)
#endif
#ifndef HOP_POLY_D64
#define HOP_POLY_D64 4
#endif

// H  = association width (contraction length of the score MMA, Q tile width);
// DV = read-out columns accumulated by ONE CTA.  The O accumulator owns TMEM columns [256, 256+DV), so DV <= 256:
// a wider read-out (Dv = 512) is cut into Dv/DV column chunks along gridDim.z, each CTA recomputing the score
// tile of its rows (the redundant score MMA is the price of the 512-column TMEM).
// kPrecise: hi + lo operand pairs (fp16(v) and fp16(v - fp16(v))) for the state, K^ and the read-out operand: the
// score tile is three MMAs (Qh Kh + Ql Kh + Qh Kl, the lo x lo term is below 2^-22) and the read-out two (P~ Bh +
// P~ Bl), ~2^-21 relative operand error instead of 2^-11.  Used for the update passes of T >= 2, where the retrieve
// iteration amplifies operand rounding; the Q tile holds both halves, so the ring is shorter.
template <int H, int DV, bool kPrecise = false>
struct TcCfg {
  // A hi + lo pass computes every exponential on the MUFU: the degree-3 polynomial's 7.5e-5 relative error is systematic,
  // and the retrieve iteration amplifies it like any other rounding (CPU emulation of the kernel's arithmetic at the
  // bifurcation point P = 256k, D = 256, T = 3: 2.2e-4 without, 1.1e-3 with the polynomial in the update passes; measured
  // with it: 0.97e-3).  These passes issue 2.5x the MMA work per exponential, so the MUFU is not their limit.
  static constexpr int kPolyPeriod = kPrecise ? 0 : (H >= 256 ? HOP_POLY_D256 : (H >= 128 ? HOP_POLY_D128 : HOP_POLY_D64));
  static constexpr int kKChunks = H / 64;                 // 64-wide K chunks of the Q / K^ tiles
  static constexpr int kBStages = DV > 128 ? DV / 128 : 1;  // ring stages of one read-out chunk (DV rows x 64 patterns)
  static constexpr int kBRows = DV / kBStages;            // rows per TMA box of the read-out operand
  // Query rows a CTA really owns.  H = 1024: a 128-row fp16 Q tile (256 KiB) does not fit shared memory, so the
  // tile keeps 64 rows (8 KiB per 64-column chunk).  The MMAs stay M = 128: the A descriptor of chunk kc reads
  // 16 KiB from q_smem + kc * 8 KiB, i.e. TMEM lanes 64..127 accumulate rows that belong to nobody (their operand
  // bytes are the next chunk's / the ring's) and are never read back; their softmax warps only keep the barrier
  // counts.  Half of the issued tensor work is thrown away -- the functional first form of this width.
  static constexpr int kRows = H > 512 ? 64 : kTileM;
  static constexpr int kQChunkBytes = kRows * 128;        // one 64-column chunk of the Q tile
  static constexpr int kQTileBytes = kKChunks * kQChunkBytes;             // one 16-bit copy of the Q tile
  static constexpr int kQBytes = (kPrecise ? 2 : 1) * kQTileBytes;        // hi [+ lo]
  static constexpr int kBarBytes = 512;                   // mbarriers + TMEM slot
  static constexpr int kXchBytes = 2 * 2 * kTileM * 4;    // [parity][half][row] floats exchanged by a warp pair
  static constexpr int kRoom = (232448 - kBarBytes - kXchBytes - kQBytes) / kChunkBytes / 2 * 2;
  // even: a 2-stage read-out chunk never wraps
  static constexpr int kStages = !kPrecise ? ((H >= 512) ? 6 : (H >= 256) ? 10 : 12) : (kRoom > 12 ? 12 : kRoom);
  static constexpr int kSmemBytes = kQBytes + kStages * kChunkBytes + kBarBytes + kXchBytes;
  static_assert(kStages % 2 == 0 && (kKChunks % 2 == 0 || kBStages == 1), "2-stage chunks must stay pair-aligned");
  static_assert(DV <= 256 && DV <= H && H % DV == 0 && kSmemBytes <= 232448, "TMEM / shared-memory budget");
  static_assert(!kPrecise || (H == DV && H <= 256 && kStages >= 6), "hi + lo operand pairs: D <= 256");
};

size_t ws_for_splits_fwd(int splits, int n, int d);   // defined with the host-side heuristics below
bool streamk_wanted(int n, int p, int d, int update_steps, const TcTune& tune);

struct TcKernelArgs {
  const float* x_raw; const float* lnq_w; const float* lnq_b; float ln_eps; int has_ln; float beta;
  const float* qs; const float* bias;
  int n, p, update_steps;
  int dv_total;                         // row pitch of the read-out (gridDim.z chunks of DV columns)
  float out_scale;                      // applied to the normalised read-out (beta when it is the next state)
  int splits, tiles_per_split;          // pattern tiles (of 128) handled by one CTA
  // stream-K form (T = 0): the (row group x pattern tile) units are cut into one contiguous share per resident
  // cluster; a cluster walks its share segment by segment (a segment = one row group, a contiguous tile range)
  int streamk, sk_base, sk_rem, sk_max_seg;
  float* retrieved; float* qsave; float* lse; int32_t* argmax; float* rowmax; int pattern_offset;
  float* rowsec;                        // per-row EXACT maximum of the 16-bit scores outside the top 64-column thread tile
  float* rowtop;                        // per-row bound on the other columns of that tile (both: certificate) or nullptr
  uint32_t* zero_words;                 // 4 control words of the argmax repair, cleared by CTA (0,0,0) (or nullptr)
  float* qn;                            // per-row ||q|| of the pass (exact-argmax certificate) or nullptr
  // partial outputs when splits > 1: (splits, N, D), (splits, N) x3
  float* part_o; float* part_lse; float* part_max; int32_t* part_arg; float* part_sec; float* part_top;
};

// Runner-up bound of one softmax thread (exact-index certificate, hop_repair.cu).  Next to its running maximum `tmax` a
// thread keeps `tsec`, the exact maximum of everything that is known to lie below it -- the maxima of all other tiles and
// every previous value of tmax -- at two instructions per tile.  What is left are the other 31 / 63 columns of the one
// tile that holds tmax: for those it remembers that tile's sum of exponentials and its reference (two predicated moves
// per tile); no single one of them can exceed the sum of all of them, so at the end of the pass
//     second <= ref + log2(sum - 2^(tmax - ref))
// closes the bound with two MUFU operations per pass instead of a second max-scan per tile.  The bound is >= the true
// runner-up (a duplicate of tmax in the same tile makes the rest >= the top term, i.e. the bound >= tmax): rows
// certified with it are safe, a loose bound only sends a row to the fp32 repair.  `kSecondSlack` covers the rounding
// of the sum and the 7.5e-5 of the polynomial exponential on the top term.
constexpr float kSecondSlack = 4e-4f;
__device__ __forceinline__ float second_bound(float tmax, float top_sum, float top_ref) {
  const float pt = ptx::ex2(tmax - top_ref);
  const float rest = fmaxf(top_sum - pt, 0.f) + kSecondSlack * pt;
  return top_ref + ptx::lg2(rest);
}

template <bool kBf16>
__device__ __forceinline__ uint32_t pack16(float a, float b) {
  if constexpr (kBf16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}

template <bool kBf16>
__device__ __forceinline__ float2 unpack16(uint32_t w) {
  if constexpr (kBf16) return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w));
  else return __half22float2(*reinterpret_cast<__half2*>(&w));
}
// low half of a hi + lo operand pair: the 16-bit encoding of what the hi word lost
template <bool kBf16>
__device__ __forceinline__ uint32_t pack16_lo(float a, float b, uint32_t hi) {
  const float2 h = unpack16<kBf16>(hi);
  return pack16<kBf16>(a - h.x, b - h.y);
}

// Byte offset of 16-byte chunk `c16` (0..7) of row `r` inside a 128B-swizzled [rows x 64] tile.
__device__ __forceinline__ uint32_t sw128_off(int r, int c16) {
  return (uint32_t)(r * 128 + ((c16 ^ (r & 7)) << 4));
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st_shared_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float ld_shared_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

// Next segment of this cluster's unit range [u, end): row group `rg`, first tile `t0`, `nt` tiles.
__device__ __forceinline__ bool seg_next(int& u, int end, int nt_total, int& rg, int& t0, int& nt) {
  if (u >= end) return false;
  rg = u / nt_total;
  t0 = u - rg * nt_total;
  nt = min(end - u, nt_total - t0);
  u += nt;
  return true;
}

// kCluster CTAs (consecutive row tiles, same pattern split) form a cluster and share every K^ / V^T stage:
// each loads 1/kCluster of the rows of a stage and multicasts it into all of them, so the L2 -> SM
// traffic of the streamed operands drops by kCluster (the kernel is L2-bandwidth-bound otherwise).
//
// kPair: the two CTAs of a cluster form a tcgen05 CTA pair (cta_group::2, M = 256).  Each keeps its own 128 query
// rows (Q tile, S / P~ / O in its own TMEM, its own softmax warps) but holds only HALF of every streamed K^ / V^T
// tile in shared memory; one thread of the leader CTA issues the MMAs for both.  Per SM this halves the TMA
// writes and the B-operand reads of shared memory -- the SS-mode score MMA alone needs 128 B/clk of operand
// fetch in the single-CTA form, i.e. the whole shared-memory bandwidth (measured: 1400 instead of 1024 cycles).
template <int H, int DV, bool kBf16, int kCluster, bool kPair, bool kPrecise>
__global__ void __launch_bounds__(kThreads, 1)
hop_fwd_tc_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_kt,
                  const __grid_constant__ CUtensorMap tm_vt, const __grid_constant__ CUtensorMap tm_klo,
                  const __grid_constant__ CUtensorMap tm_blo, const TcKernelArgs a) {
  using Cfg = TcCfg<H, DV, kPrecise>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = ptx::smem_u32(smem_raw);
  if ((smem_base & 1023u) != 0) __trap();       // the 128B-swizzle atoms need a 1 KiB-aligned base
  const uint32_t q_smem = smem_base;
  const uint32_t ring = q_smem + Cfg::kQBytes;
  const uint32_t bars = ring + Cfg::kStages * kChunkBytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (Cfg::kStages + s); };
  const uint32_t misc = bars + 16u * Cfg::kStages;
  auto s_full = [&](int b) { return misc + 8u * b; };
  auto p_full = [&](int b) { return misc + 32u + 8u * b; };
  auto p_empty = [&](int b) { return misc + 48u + 8u * b; };
  const uint32_t q_ready = misc + 64u;
  const uint32_t tmem_slot = misc + 72u;
  const uint32_t xch = bars + Cfg::kBarBytes;   // [parity][half][row]
  auto xch_at = [&](int parity, int half, int r) { return xch + (uint32_t)(((parity * 2 + half) * kTileM + r) * 4); };

  static_assert(!kPair || kCluster == 2, "a CTA pair is a cluster of two");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t crank = kCluster > 1 ? ptx::cluster_ctarank() : 0u;
  constexpr uint32_t kArrivers = kSoftmaxThreads * (kPair ? 2 : 1);   // softmax threads feeding the MMA issuer
  const int split = blockIdx.y;
  const int zoff = (int)blockIdx.z * DV;       // first read-out column of this CTA
  const int nt_total = (a.p + kTileP - 1) / kTileP;
  const int cid = (int)blockIdx.x / kCluster;  // cluster index along x (row group in the plain form)
  // unit range of this cluster: one segment in the plain form, a contiguous share in the stream-K form
  int u_start, u_end;
  if (a.streamk) {
    u_start = cid * a.sk_base + min(cid, a.sk_rem);
    u_end = u_start + a.sk_base + (cid < a.sk_rem ? 1 : 0);
  } else {
    const int ts = split * a.tiles_per_split;
    u_start = cid * nt_total + ts;
    u_end = u_start + min(a.tiles_per_split, nt_total - ts);
  }
  const int passes = a.update_steps + 1;

  if (threadIdx.x == 0) {
    if (a.zero_words && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0)
      *reinterpret_cast<uint4*>(a.zero_words) = make_uint4(0u, 0u, 0u, 0u);
    for (int s = 0; s < Cfg::kStages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), kPair ? 1 : kCluster);   // pair: one multicast commit of the leader's MMA thread
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(s_full(b), 1);
      ptx::mbar_init(p_full(b), kArrivers); ptx::mbar_init(p_empty(b), 1);
    }
    ptx::mbar_init(q_ready, kArrivers);
    ptx::fence_barrier_init();
    ptx::prefetch_tmap(&tm_k); ptx::prefetch_tmap(&tm_kt); ptx::prefetch_tmap(&tm_vt);
    if constexpr (kPrecise) { ptx::prefetch_tmap(&tm_klo); ptx::prefetch_tmap(&tm_blo); }
  }
  if (warp == 0) {
    __syncwarp();
    if constexpr (kPair) { ptx::tmem_alloc_pair(tmem_slot, kTmemCols); ptx::tmem_relinquish_pair(); }
    else { ptx::tmem_alloc(tmem_slot, kTmemCols); ptx::tmem_relinquish(); }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if constexpr (kCluster > 1) ptx::cluster_sync();   // peers' barriers are initialised before any remote arrive
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if constexpr (kPair) {
      if (ptx::elect_one()) {
        // each CTA streams ITS half of every tile into its own ring; bytes are counted on the leader's barrier
        uint32_t it = 0;
        constexpr int kKPerStage = Cfg::kKChunks >= 2 ? 2 : 1;          // 8 KiB K^ chunks (64 patterns x 64) per stage
        constexpr int kKStages = Cfg::kKChunks / kKPerStage;
        auto acquire = [&](uint32_t total_bytes) {
          const int s = it % Cfg::kStages;
          ptx::mbar_wait(empty_bar(s), ((it / Cfg::kStages) & 1) ^ 1);
          if (crank == 0) ptx::mbar_arrive_expect_tx(full_bar(s), total_bytes);
          return s;
        };
        int u = u_start, rg, t0, nt;
        while (seg_next(u, u_end, nt_total, rg, t0, nt))
        for (int pass = 0; pass < passes; ++pass) {
          const CUtensorMap* bm = (pass < a.update_steps) ? &tm_kt : &tm_vt;
          auto load_k = [&](int tile) {
            for (int half = 0; half < (kPrecise ? 2 : 1); ++half) {       // hi chunks, then the lo chunks
              const CUtensorMap* km = half ? &tm_klo : &tm_k;
              for (int ks = 0; ks < kKStages; ++ks) {
                const int s = acquire(2u * kKPerStage * 8192u);
                const uint32_t fb = ptx::map_to_cta(full_bar(s), 0);
                for (int c = 0; c < kKPerStage; ++c)
                  ptx::tma_load_2d_pair(ring + s * kChunkBytes + c * 8192, km, fb, (ks * kKPerStage + c) * 64,
                                        tile * kTileP + (int)crank * 64);
                ++it;
              }
            }
          };
          auto load_b = [&](int tile) {
            for (int half = 0; half < (kPrecise ? 2 : 1); ++half) {
              const CUtensorMap* m = half ? &tm_blo : bm;
              for (int kc = 0; kc < 2; ++kc) {
                const int s = acquire(2u * (DV / 2) * 128u);
                ptx::tma_load_2d_pair(ring + s * kChunkBytes, m, ptx::map_to_cta(full_bar(s), 0),
                                      tile * kTileP + kc * 64, zoff + (int)crank * (DV / 2));
                ++it;
              }
            }
          };
          load_k(t0);
          for (int j = 0; j < nt; ++j) {
            if (j + 1 < nt) load_k(t0 + j + 1);
            load_b(t0 + j);
          }
        }
      }
    } else
    if (ptx::elect_one()) {
      uint32_t it = 0;
      constexpr uint16_t kMask = (uint16_t)((1u << kCluster) - 1u);
      // `rows` = rows of the whole stage box (coordinate c1 counts rows); this CTA loads rows/kCluster of them
      auto produce = [&](const CUtensorMap* m, int c0, int c1, int rows) {
        const int s = it % Cfg::kStages;
        ptx::mbar_wait(empty_bar(s), ((it / Cfg::kStages) & 1) ^ 1);    // all kCluster consumers released it
        ptx::mbar_arrive_expect_tx(full_bar(s), (uint32_t)rows * 128u);
        if constexpr (kCluster > 1) {
          const int part = rows / kCluster;
          ptx::tma_load_2d_mc(ring + s * kChunkBytes + (int)crank * part * 128, m, full_bar(s), c0,
                              c1 + (int)crank * part, kMask);
        } else {
          ptx::tma_load_2d(ring + s * kChunkBytes, m, full_bar(s), c0, c1);
        }
        ++it;
      };
      int u = u_start, rg, t0, nt;
      while (seg_next(u, u_end, nt_total, rg, t0, nt))
      for (int pass = 0; pass < passes; ++pass) {
        const CUtensorMap* bm = (pass < a.update_steps) ? &tm_kt : &tm_vt;
        auto load_k = [&](int tile) {
          for (int kc = 0; kc < Cfg::kKChunks; ++kc) produce(&tm_k, kc * 64, tile * kTileP, kTileP);
          if constexpr (kPrecise)
            for (int kc = 0; kc < Cfg::kKChunks; ++kc) produce(&tm_klo, kc * 64, tile * kTileP, kTileP);
        };
        auto load_b = [&](int tile) {
          for (int half = 0; half < (kPrecise ? 2 : 1); ++half)
            for (int kc = 0; kc < 2; ++kc)
              for (int bs = 0; bs < Cfg::kBStages; ++bs)
                produce(half ? &tm_blo : bm, tile * kTileP + kc * 64, zoff + bs * Cfg::kBRows, Cfg::kBRows);
        };
        load_k(t0);
        for (int j = 0; j < nt; ++j) {
          if (j + 1 < nt) load_k(t0 + j + 1);
          load_b(t0 + j);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    if constexpr (kPair) {
      if (crank == 0 && ptx::elect_one()) {
        constexpr int fmt = kBf16 ? 1 : 0;
        constexpr uint32_t idesc_qk = ptx::umma_idesc_f16(2 * kTileM, kTileP, fmt);
        constexpr uint32_t idesc_pv = ptx::umma_idesc_f16(2 * kTileM, DV, fmt);
        constexpr int kKPerStage = Cfg::kKChunks >= 2 ? 2 : 1;
        constexpr int kKStages = Cfg::kKChunks / kKPerStage;
        uint32_t it = 0, g = 0, qphase = 0;
        int u = u_start, rg, t0, nt;
        while (seg_next(u, u_end, nt_total, rg, t0, nt))
        for (int pass = 0; pass < passes; ++pass) {
          ptx::mbar_wait_cluster(q_ready, (qphase++) & 1);     // both CTAs' Q tiles are in their shared memory
          ptx::tc_fence_after();
          auto issue_qk = [&](uint32_t gg) {
            const int b = gg & 1;
            for (int half = 0; half < (kPrecise ? 2 : 1); ++half)      // K^ hi stages (x Qh [, x Ql]), then K^ lo (x Qh)
            for (int ks = 0; ks < kKStages; ++ks) {
              const int s = it % Cfg::kStages;
              ptx::mbar_wait_cluster(full_bar(s), (it / Cfg::kStages) & 1);
              ptx::tc_fence_after();
#pragma unroll
              for (int c = 0; c < kKPerStage; ++c) {
                const uint64_t ad = ptx::umma_desc_sw128(q_smem + (ks * kKPerStage + c) * Cfg::kQChunkBytes);
                const uint64_t bd = ptx::umma_desc_sw128(ring + s * kChunkBytes + c * 8192);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  ptx::umma2_f16_ss(tmem_base + b * kTileP, ad + 2 * k, bd + 2 * k, idesc_qk, (half | ks | c | k) != 0);
                if (kPrecise && half == 0) {
                  const uint64_t al = ptx::umma_desc_sw128(q_smem + Cfg::kQTileBytes + (ks * kKPerStage + c) * Cfg::kQChunkBytes);
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    ptx::umma2_f16_ss(tmem_base + b * kTileP, al + 2 * k, bd + 2 * k, idesc_qk, 1u);
                }
              }
              ptx::umma2_commit_both(empty_bar(s));
              ++it;
            }
            ptx::umma2_commit_both(s_full(b));
          };
          issue_qk(g);
          for (int j = 0; j < nt; ++j) {
            HOP_TRACE(0, j, 0);
            if (j + 1 < nt) issue_qk(g + j + 1);
            HOP_TRACE(0, j, 1);
            const uint32_t gg = g + j;
            const int b = gg & 1;
            ptx::mbar_wait_cluster(p_full(b), (gg >> 1) & 1);  // P~ of both CTAs is in their TMEM
            ptx::tc_fence_after();
            HOP_TRACE(0, j, 2);
            for (int half = 0; half < (kPrecise ? 2 : 1); ++half)      // read-out operand hi, then lo
            for (int kc = 0; kc < 2; ++kc) {
              const int s = it % Cfg::kStages;
              ptx::mbar_wait_cluster(full_bar(s), (it / Cfg::kStages) & 1);
              ptx::tc_fence_after();
              const uint32_t at = tmem_base + b * kTileP + kc * 32;
              const uint64_t bd = ptx::umma_desc_sw128(ring + s * kChunkBytes);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                ptx::umma2_f16_ts(tmem_base + kTmemO, at + 8 * k, bd + 2 * k, idesc_pv, (j | half | kc | k) != 0);
              ptx::umma2_commit_both(empty_bar(s));
              ++it;
            }
            ptx::umma2_commit_both(p_empty(b));
            HOP_TRACE(0, j, 3);
          }
          g += nt;
        }
      }
    } else
    if (ptx::elect_one()) {
      constexpr int fmt = kBf16 ? 1 : 0;
      constexpr uint32_t idesc_qk = ptx::umma_idesc_f16(kTileM, kTileP, fmt);
      constexpr uint32_t idesc_pv = ptx::umma_idesc_f16(kTileM, DV, fmt);
      uint32_t it = 0, g = 0;
      // a ring stage is free again once the MMAs of EVERY CTA of the cluster that read it have completed
      auto release = [&](uint32_t bar) {
        if constexpr (kCluster > 1) ptx::umma_commit_mc(bar, (uint16_t)((1u << kCluster) - 1u));
        else ptx::umma_commit(bar);
      };
      uint32_t qphase = 0;
      int u = u_start, rg, t0, nt;
      while (seg_next(u, u_end, nt_total, rg, t0, nt))
      for (int pass = 0; pass < passes; ++pass) {
        ptx::mbar_wait(q_ready, (qphase++) & 1);
        ptx::tc_fence_after();
        auto issue_qk = [&](uint32_t gg) {
          // S buffer b also holds P~ of tile gg-2: its read-out MMA was issued before this point and
          // tcgen05.mma instructions of one thread execute in issue order, so no barrier is needed
          const int b = gg & 1;
          for (int half = 0; half < (kPrecise ? 2 : 1); ++half)      // K^ hi chunks (x Qh [, x Ql]), then K^ lo (x Qh)
          for (int kc = 0; kc < Cfg::kKChunks; ++kc) {
            const int s = it % Cfg::kStages;
            ptx::mbar_wait(full_bar(s), (it / Cfg::kStages) & 1);
            ptx::tc_fence_after();
            const uint64_t ad = ptx::umma_desc_sw128(q_smem + kc * Cfg::kQChunkBytes);
            const uint64_t bd = ptx::umma_desc_sw128(ring + s * kChunkBytes);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_f16_ss(tmem_base + b * kTileP, ad + 2 * k, bd + 2 * k, idesc_qk, (half | kc | k) != 0);
            if (kPrecise && half == 0) {
              const uint64_t al = ptx::umma_desc_sw128(q_smem + Cfg::kQTileBytes + kc * Cfg::kQChunkBytes);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                ptx::umma_f16_ss(tmem_base + b * kTileP, al + 2 * k, bd + 2 * k, idesc_qk, 1u);
            }
            release(empty_bar(s));
            ++it;
          }
          ptx::umma_commit(s_full(b));
        };
        issue_qk(g);
        for (int j = 0; j < nt; ++j) {
          HOP_TRACE(0, j, 0);
          if (j + 1 < nt) issue_qk(g + j + 1);
          HOP_TRACE(0, j, 1);
          const uint32_t gg = g + j;
          const int b = gg & 1;
          ptx::mbar_wait(p_full(b), (gg >> 1) & 1);
          ptx::tc_fence_after();
          HOP_TRACE(0, j, 2);
          for (int half = 0; half < (kPrecise ? 2 : 1); ++half)      // read-out operand hi, then lo
          for (int kc = 0; kc < 2; ++kc) {
            // one read-out chunk = D rows x 64 patterns = kBStages consecutive (pair-aligned) ring stages
            const int s = it % Cfg::kStages;
            for (int bs = 0; bs < Cfg::kBStages; ++bs) {
              const uint32_t i2 = it + bs;
              ptx::mbar_wait(full_bar(i2 % Cfg::kStages), (i2 / Cfg::kStages) & 1);
            }
            ptx::tc_fence_after();
            const uint32_t at = tmem_base + b * kTileP + kc * 32;   // P~ (16-bit pairs) aliases the S buffer
            const uint64_t bd = ptx::umma_desc_sw128(ring + s * kChunkBytes);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_f16_ts(tmem_base + kTmemO, at + 8 * k, bd + 2 * k, idesc_pv, (j | half | kc | k) != 0);
            for (int bs = 0; bs < Cfg::kBStages; ++bs) {
              release(empty_bar(it % Cfg::kStages));
              ++it;
            }
          }
          ptx::umma_commit(p_empty(b));
          HOP_TRACE(0, j, 3);
        }
        g += nt;
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------ softmax / epilogue
    const int wq = warp & 3;                      // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;             // which 64 score columns / which half of the O columns
    const int row = wq * 32 + lane;               // row of the tile owned by this thread (shared with the partner warp)
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    const float qscale = a.beta;                  // natural-units scale of the state pattern
    const uint32_t pair_bar = 1 + wq;             // named barrier of the warp pair (64 threads)
    constexpr int kOGroups = DV / 64;             // 32-column O groups per half
    const int og0 = half * kOGroups;
    // q_ready / p_full are waited on by the MMA issuer: in a CTA pair that is the leader's thread
    auto arrive_issuer = [&](uint32_t bar) {
      if constexpr (kPair) ptx::mbar_arrive_cluster(ptx::map_to_cta(bar, 0));
      else ptx::mbar_arrive(bar);
    };

    uint32_t g = 0;
    int u = u_start, rg, t0, nt, seg_local = 0;
    while (seg_next(u, u_end, nt_total, rg, t0, nt)) {
    const long long row_tile = (long long)rg * kCluster + (long long)crank;   // (plain form: == blockIdx.x)
    const long long grow = row_tile * Cfg::kRows + row;
    const bool live = row < Cfg::kRows;            // (H = 1024: TMEM lanes 64..127 hold no query row)
    const bool row_ok = live && grow < a.n;
    // row of this thread inside the partial-result arrays
    const long long part_row = a.streamk
        ? ((long long)cid * a.sk_max_seg + seg_local) * (kCluster * kTileM) + (long long)crank * kTileM + row
        : (long long)split * a.n + grow;
    ++seg_local;
    // ---- pass-0 Q tile: one warp per row, 4 rows in flight; LN_q + beta, fp32 qsave, 16-bit swizzled smem
    // (a later segment may overwrite the Q tile: its predecessor ended by waiting for its last MMA)
    {
      constexpr int kVec = (H + 127) / 128;       // float4 per lane
      constexpr int kRowsInFlight = H > 256 ? 2 : 4;
      constexpr int kRowsPerWarp = Cfg::kRows / kSoftmaxWarps;
      const int sw = warp - 2;
      float4 w4[kVec], b4[kVec];
#pragma unroll
      for (int v = 0; v < kVec; ++v) {
        const int e = lane * 4 + v * 128;
        w4[v] = make_float4(1.f, 1.f, 1.f, 1.f);
        b4[v] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (e < H && a.x_raw && a.has_ln) {
          if (a.lnq_w) w4[v] = *reinterpret_cast<const float4*>(a.lnq_w + e);
          if (a.lnq_b) b4[v] = *reinterpret_cast<const float4*>(a.lnq_b + e);
        }
      }
      const float* src = a.x_raw ? a.x_raw : a.qs;
      for (int rr = 0; rr < kRowsPerWarp; rr += kRowsInFlight) {
        float4 xv[kRowsInFlight][kVec];
#pragma unroll
        for (int u = 0; u < kRowsInFlight; ++u) {
          const long long gr = row_tile * Cfg::kRows + sw * kRowsPerWarp + rr + u;
#pragma unroll
          for (int v = 0; v < kVec; ++v) {
            const int e = lane * 4 + v * 128;
            xv[u][v] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (e < H && gr < a.n) xv[u][v] = __ldg(reinterpret_cast<const float4*>(src + gr * H + e));
          }
        }
#pragma unroll
        for (int u = 0; u < kRowsInFlight; ++u) {
          const int r = sw * kRowsPerWarp + rr + u;
          const long long gr = row_tile * Cfg::kRows + r;
          if (a.x_raw) {
            if (a.has_ln) {
              float s = 0.f;
#pragma unroll
              for (int v = 0; v < kVec; ++v) s += (xv[u][v].x + xv[u][v].y) + (xv[u][v].z + xv[u][v].w);
              const float mean = warp_sum(s) * (1.f / H);
              float q = 0.f;
#pragma unroll
              for (int v = 0; v < kVec; ++v) {
                const int e = lane * 4 + v * 128;
                if (e < H) {
                  xv[u][v].x -= mean; xv[u][v].y -= mean; xv[u][v].z -= mean; xv[u][v].w -= mean;
                  q += xv[u][v].x * xv[u][v].x + xv[u][v].y * xv[u][v].y + xv[u][v].z * xv[u][v].z +
                       xv[u][v].w * xv[u][v].w;
                }
              }
              const float rstd = rsqrtf(warp_sum(q) * (1.f / H) + a.ln_eps);
#pragma unroll
              for (int v = 0; v < kVec; ++v) {
                xv[u][v].x = xv[u][v].x * rstd * w4[v].x + b4[v].x; xv[u][v].y = xv[u][v].y * rstd * w4[v].y + b4[v].y;
                xv[u][v].z = xv[u][v].z * rstd * w4[v].z + b4[v].z; xv[u][v].w = xv[u][v].w * rstd * w4[v].w + b4[v].w;
              }
            }
#pragma unroll
            for (int v = 0; v < kVec; ++v) {
              xv[u][v].x *= qscale; xv[u][v].y *= qscale; xv[u][v].z *= qscale; xv[u][v].w *= qscale;
            }
          }
          if (a.qn != nullptr && t0 == 0 && blockIdx.z == 0) {       // ||q|| of the row: bounds the 16-bit score noise
            float s2 = 0.f;
#pragma unroll
            for (int v = 0; v < kVec; ++v)
              s2 += xv[u][v].x * xv[u][v].x + xv[u][v].y * xv[u][v].y + xv[u][v].z * xv[u][v].z + xv[u][v].w * xv[u][v].w;
            s2 = warp_sum(s2);
            if (lane == 0 && gr < a.n) a.qn[gr] = sqrtf(s2);
          }
#pragma unroll
          for (int v = 0; v < kVec; ++v) {
            const int e = lane * 4 + v * 128;
            if (e < H) {
              if (a.qsave && a.update_steps == 0 && gr < a.n && t0 == 0 && blockIdx.z == 0)
                *reinterpret_cast<float4*>(a.qsave + gr * H + e) = xv[u][v];
              const uint32_t lo = pack16<kBf16>(xv[u][v].x * kLog2e, xv[u][v].y * kLog2e);
              const uint32_t hi = pack16<kBf16>(xv[u][v].z * kLog2e, xv[u][v].w * kLog2e);
              const uint32_t addr = q_smem + (e >> 6) * Cfg::kQChunkBytes + sw128_off(r, (e & 63) >> 3) + ((e & 4) << 1);
              asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(lo), "r"(hi) : "memory");
              if constexpr (kPrecise) {         // second copy of the tile: what the 16-bit words above lost
                const uint32_t rlo = pack16_lo<kBf16>(xv[u][v].x * kLog2e, xv[u][v].y * kLog2e, lo);
                const uint32_t rhi = pack16_lo<kBf16>(xv[u][v].z * kLog2e, xv[u][v].w * kLog2e, hi);
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr + Cfg::kQTileBytes), "r"(rlo), "r"(rhi) : "memory");
              }
            }
          }
        }
      }
    }

    for (int pass = 0; pass < passes; ++pass) {
      ptx::fence_proxy_async_smem();
      arrive_issuer(q_ready);

      // m_run is shared by the pair (identical in both threads of a row); l, tmax, targ are per half
      float m_run = -INFINITY, l = 0.f, tmax = -INFINITY, tsec = -INFINITY, top_sum = 0.f, top_ref = 0.f;
      int targ = 0;
      for (int j = 0; j < nt; ++j) {
        const uint32_t gg = g + j;
        const int b = gg & 1;
        const int col0 = (t0 + j) * kTileP + half * kHalfP;   // global pattern index of this thread's column 0
        const int valid = a.p - col0;                          // columns >= valid are padding
#ifdef HOP_TC_TRACE
        const int trole = (lane == 0) ? warp - 1 : -1;   // roles 1..8 = softmax warps 2..9
#define HOP_STRACE(ev) do { if (trole > 0) HOP_TRACE(trole, j, ev); } while (0)
#else
#define HOP_STRACE(ev) do {} while (0)
#endif
        HOP_STRACE(0);
        ptx::mbar_wait(s_full(b), (gg >> 1) & 1);
        ptx::tc_fence_after();
        HOP_STRACE(1);
        if (!live) {                         // whole warps (lane quarters 2, 3): stay in step, feed the arrival count
          ptx::tc_fence_before();
          arrive_issuer(p_full(b));
          continue;
        }
        const uint32_t s_addr = lane_base + b * kTileP + half * kHalfP;

        // this thread's 64 scores into registers (one TMEM round trip), then free the S buffer at once
        uint32_t v[kHalfP];
        ptx::tmem_ld_x32p(s_addr, v);
        ptx::tmem_ld_x32p(s_addr + 32, v + 32);
        ptx::tmem_wait_ld();
        HOP_STRACE(2);
        if (valid < kHalfP) {
#pragma unroll
          for (int i = 0; i < kHalfP; ++i)
            if (i >= valid) v[i] = __float_as_uint(-INFINITY);
        }

        // max of the 64 columns on four independent chains (FMNMX3); chain u holds elements 8g + 2u + {0,1}
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < kHalfP; i += 8) {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            mx[u] = fmaxf(mx[u], fmaxf(__uint_as_float(v[i + 2 * u]), __uint_as_float(v[i + 2 * u + 1])));
        }
        const float hmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
        // publish the half maximum; the row maximum decides the (rare) rescale identically in both halves
        st_shared_f32(xch_at(b, half, row), hmax);
        // nearest-pattern index, only when the half maximum moved: scan just the chains that hold it.  The runner-up
        // `tsec` takes the old maximum then (see second_bound), else the tile maximum
        const bool moved = hmax > tmax;
#ifndef HOP_NO_TOP2          // (timing experiments only: without the tracking every row fails the certificate)
        tsec = fmaxf(tsec, moved ? tmax : hmax);
#endif
        if (moved) {
          int idx = kHalfP - 1;
#pragma unroll
          for (int u = 3; u >= 0; --u) {
            if (mx[u] == hmax) {
              int iu = kHalfP;
#pragma unroll
              for (int gq = 7; gq >= 0; --gq) {
                iu = (__uint_as_float(v[8 * gq + 2 * u + 1]) == hmax) ? 8 * gq + 2 * u + 1 : iu;
                iu = (__uint_as_float(v[8 * gq + 2 * u]) == hmax) ? 8 * gq + 2 * u : iu;
              }
              idx = min(idx, iu);
            }
          }
          tmax = hmax;
          targ = col0 + idx;
        }
        ptx::named_bar_sync(pair_bar, 64);
        const float cmax = fmaxf(hmax, ld_shared_f32(xch_at(b, half ^ 1, row)));
        HOP_STRACE(3);

        // lazy rescale of (l, O): only when some row of the warp outgrew its reference max by 2^15.  The new
        // reference is fixed here, but O is touched only AFTER this tile's exponentials: by then the previous
        // tile's read-out MMA (which must finish before O may be rescaled) is done instead of being waited for.
        const bool need = cmax > m_run + kRescaleThreshold;
        const bool any_need = __any_sync(0xffffffffu, need);
        const float m_old = m_run;
        if (any_need) m_run = fmaxf(m_run, cmax);

        HOP_STRACE(4);
        // P~ = 2^(s - m_run) as 16-bit pairs into TMEM, over the first 64 columns of the S buffer just read
        // (this half's 64 patterns -> columns 32h..32h+31; the partner finished its tcgen05.ld of those
        // columns before the pair barrier above); the A operand of the read-out MMA
        float ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < kHalfP; ++i) {
          const float xs = __uint_as_float(v[i]) - m_run;
          const bool poly = Cfg::kPolyPeriod > 0 && (i % (Cfg::kPolyPeriod > 0 ? Cfg::kPolyPeriod : 1)) == 0;
          v[i] = __float_as_uint(poly ? ptx::ex2_poly(xs) : ptx::ex2(xs));
        }
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float p0 = __uint_as_float(v[2 * i]), p1 = __uint_as_float(v[2 * i + 1]);
          ls[i & 3] += p0 + p1;
          pk[i] = pack16<kBf16>(p0, p1);
        }
        if (any_need && j > 0) {
          const uint32_t pg = gg - 1;
          ptx::mbar_wait(p_empty(pg & 1), (pg >> 1) & 1);   // read-out MMA of the previous tile done
          ptx::tc_fence_after();
          const float f = ptx::ex2(m_old - m_run);
          l *= f;
#pragma unroll 1
          for (int c = og0; c < og0 + kOGroups; ++c) {
            uint32_t o[32];
            ptx::tmem_ld_x32(lane_base + kTmemO + c * 32, o);
            ptx::tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * f);
            ptx::tmem_st_x32(lane_base + kTmemO + c * 32, o);
          }
        }
        HOP_STRACE(5);
        ptx::tmem_st_x32(lane_base + b * kTileP + half * 32, pk);
        const float tsum = (ls[0] + ls[1]) + (ls[2] + ls[3]);
#ifndef HOP_NO_TOP2
        top_sum = moved ? tsum : top_sum;
        top_ref = moved ? m_run : top_ref;
#endif
        l += tsum;
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        arrive_issuer(p_full(b));
        HOP_STRACE(6);
      }

      if (!live) {                           // the Q tile may be rewritten only after the last MMA that reads it
        const uint32_t lastd = g + nt - 1;
        ptx::mbar_wait(p_empty(lastd & 1), (lastd >> 1) & 1);
        ptx::tc_fence_after();
        ptx::tc_fence_before();
        g += nt;
        continue;
      }
      // ---- end of the pass: the pair combines (l, tmax, targ); slots of the other parity are idle
      const int pl = (g + nt - 1) & 1;
      float ttop = second_bound(tmax, top_sum, top_ref);   // (tsec stays the exact maximum outside the top tile)
      st_shared_f32(xch_at(pl ^ 1, half, row), l);
      ptx::named_bar_sync(pair_bar, 64);
      const float l_tot = l + ld_shared_f32(xch_at(pl ^ 1, half ^ 1, row));
      if (half == 1) {
        st_shared_f32(xch_at(pl, 1, row), tmax);
        st_shared_f32(xch_at(pl, 0, row), __int_as_float(targ));
        st_shared_f32(xch_at(pl ^ 1, 0, row), tsec);   // half 0's l slot: this thread was its only reader
      }
      ptx::named_bar_sync(pair_bar, 64);
      float om = -INFINITY, osec = -INFINITY;
      int oa = 0;
      if (half == 0) {
        om = ld_shared_f32(xch_at(pl, 1, row));
        oa = __float_as_int(ld_shared_f32(xch_at(pl, 0, row)));
        osec = ld_shared_f32(xch_at(pl ^ 1, 0, row));
      }
      ptx::named_bar_sync(pair_bar, 64);               // (a fourth value needs a slot that half 0 has finished reading)
      if (half == 1) st_shared_f32(xch_at(pl, 1, row), ttop);
      ptx::named_bar_sync(pair_bar, 64);
      if (half == 0) {
        // everything the losing half saw lies at or below ITS maximum; the winner keeps its own tile's bound
        if (om > tmax) { tsec = fmaxf(osec, tmax); ttop = ld_shared_f32(xch_at(pl, 1, row)); tmax = om; targ = oa; }
        else tsec = fmaxf(tsec, om);                   // ties keep the lower (half-0) index
      }
      // wait for the last read-out MMA, normalise
      const uint32_t last = g + nt - 1;
      ptx::mbar_wait(p_empty(last & 1), (last >> 1) & 1);
      ptx::tc_fence_after();
      const float inv = 1.f / l_tot;
      if (H == DV && pass < a.update_steps) {   // (a chunked read-out runs its updates as separate launches)
        // retrieve update: q <- beta * (A K^) becomes the next pass' Q tile (never leaves the SM)
        const bool save = (pass == a.update_steps - 1) && a.qsave && row_ok;
        const float sc = inv * qscale;
#pragma unroll 1
        for (int c = og0; c < og0 + kOGroups; ++c) {
          uint32_t o[32];
          ptx::tmem_ld_x32(lane_base + kTmemO + c * 32, o);
          ptx::tmem_wait_ld();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float q0 = __uint_as_float(o[2 * i]) * sc, q1 = __uint_as_float(o[2 * i + 1]) * sc;
            o[2 * i] = __float_as_uint(q0); o[2 * i + 1] = __float_as_uint(q1);
            pk[i] = pack16<kBf16>(q0 * kLog2e, q1 * kLog2e);
          }
          if (save) {
            float4* dst = reinterpret_cast<float4*>(a.qsave + grow * H + c * 32);
#pragma unroll
            for (int q4 = 0; q4 < 8; ++q4)
              dst[q4] = make_float4(__uint_as_float(o[4 * q4]), __uint_as_float(o[4 * q4 + 1]),
                                    __uint_as_float(o[4 * q4 + 2]), __uint_as_float(o[4 * q4 + 3]));
          }
          const uint32_t base = q_smem + (c >> 1) * Cfg::kQChunkBytes;
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4)
            st_shared_v4(base + sw128_off(row, (c & 1) * 4 + q4), pk[4 * q4], pk[4 * q4 + 1], pk[4 * q4 + 2],
                         pk[4 * q4 + 3]);
        }
        ptx::tc_fence_before();
      } else {
        float* o_dst = ((a.splits > 1) ? a.part_o + part_row * a.dv_total
                                       : a.retrieved + grow * a.dv_total) + zoff;
        const float osc = (a.splits > 1) ? inv : inv * a.out_scale;
#pragma unroll 1
        for (int c = og0; c < og0 + kOGroups; ++c) {
          uint32_t o[32];
          ptx::tmem_ld_x32(lane_base + kTmemO + c * 32, o);
          ptx::tmem_wait_ld();
          if (row_ok) {
            float4* dst = reinterpret_cast<float4*>(o_dst + c * 32);
            const float4* bia = (a.bias && a.splits == 1) ? reinterpret_cast<const float4*>(a.bias + zoff + c * 32) : nullptr;
#pragma unroll
            for (int q4 = 0; q4 < 8; ++q4) {
              float4 r = make_float4(__uint_as_float(o[4 * q4]) * osc, __uint_as_float(o[4 * q4 + 1]) * osc,
                                     __uint_as_float(o[4 * q4 + 2]) * osc, __uint_as_float(o[4 * q4 + 3]) * osc);
              if (bia) { const float4 bb = __ldg(bia + q4); r.x += bb.x; r.y += bb.y; r.z += bb.z; r.w += bb.w; }
              dst[q4] = r;
            }
          }
        }
        if (row_ok && half == 0 && blockIdx.z == 0) {
          const float lse = kLn2 * (m_run + ptx::lg2(l_tot));
          if (a.splits > 1) {
            const long long o = part_row;
            a.part_lse[o] = lse; a.part_max[o] = tmax * kLn2; a.part_arg[o] = targ + a.pattern_offset;
            a.part_sec[o] = tsec * kLn2; a.part_top[o] = ttop * kLn2;
          } else {
            if (a.lse) a.lse[grow] = lse;
            if (a.rowmax) a.rowmax[grow] = tmax * kLn2;
            if (a.rowsec) { a.rowsec[grow] = tsec * kLn2; a.rowtop[grow] = ttop * kLn2; }
            if (a.argmax) a.argmax[grow] = targ + a.pattern_offset;
          }
        }
        ptx::tc_fence_before();
      }
      g += nt;
    }
    }  // segments of this cluster
  }

  __syncthreads();
  if constexpr (kCluster > 1) ptx::cluster_sync();   // no CTA leaves while a peer may still signal its barriers
  if (warp == 0) {
    ptx::tc_fence_after();
    if constexpr (kPair) ptx::tmem_dealloc_pair(tmem_base, kTmemCols);
    else ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------------------
// 64-rows-per-CTA form for the wide widths (H = 512, 1024).  A tcgen05 CTA pair with M = 128: each CTA supplies 64
// query rows and half of every streamed operand tile.  Measured layout (scripts/probes/umma_layout_probe.cu): the
// accumulator row r of a CTA sits in TMEM lane r for columns n < N/2 and in lane r + 64 for n >= N/2, both in TMEM
// columns [0, N/2) -- an N-wide accumulator costs N/2 TMEM columns.  So one CTA holds two TP-pattern score tiles
// (TP/2 columns each) AND a 512-column read-out (256 columns, two N = 256 instructions): no score recompute at
// D = 512, two column chunks at D = 1024, and a Q tile of 64 KiB / 128 KiB.
//   TMEM   S0 [0,TP/2)  S1 [TP/2,TP)  O [TP, TP + DV/2);  P~ (16-bit pairs, TP/2 columns) over the score tile
//   lane L = 32 (warp % 4) + lane: row L & 63, pattern half h = L >> 6; two threads per lane (sub = 0 | 1), TP/4
//   score columns each.  The four threads of a row agree on the row maximum through shared memory; the read-out
//   MMA takes P~ from TMEM and every lane reads ITS OWN lane's data (measured), so the TP/8 packed words of a
//   thread are swapped with the partner lane (L ^ 64) through shared memory and stored into both lanes.
// TP = patterns per score tile: 256 at H = 512 (the M = 128 pair instruction issues at 3,640 FLOP/clk/SM with N = 128
// but 5,040 with N = 256, profiles/r01_umma_layout_probe.txt), 128 at H = 1024 (the 32 KiB exchange buffer of the
// wide tile leaves room for four ring stages only next to the 128 KiB Q tile: measured slower, HOP_TC64_TILE=256).
template <int H, int DV, int TP>
struct Tc64Cfg {
  static constexpr int kPolyPeriod = HOP_POLY_D256;
  static constexpr int kRows = 64;
  static constexpr int kTP = TP;
  static constexpr int kSCols = TP / 2;                     // TMEM columns of one score tile (and of its P~)
  static constexpr int kC = TP / 4;                         // score columns per softmax thread
  static constexpr int kW = kC / 2;                         // packed P~ words per softmax thread
  static constexpr int kPC = TP / 64;                       // 64-pattern K steps of the read-out per tile
  static constexpr int kKChunks = H / 64;
  static constexpr int kKPerStage = TP == 128 ? 2 : 1;      // K^ chunks (TP/2 patterns x 64) per 16 KiB ring stage
  static constexpr int kKChunkBytes = (TP / 2) * 128;       // this CTA's half of one K^ chunk
  static constexpr int kKStages = kKChunks / kKPerStage;    // ring stages of one K^ tile
  static constexpr int kQChunkBytes = kRows * 128;
  static constexpr int kQBytes = kKChunks * kQChunkBytes;
  static constexpr int kNH = DV / 256;                      // N = 256 read-out instructions per K step
  static constexpr int kStages = H >= 1024 ? (TP == 256 ? 4 : 5) : (TP == 256 ? 8 : HOP_TC64_STAGES_512);
  static constexpr int kBarBytes = 512;
  static constexpr int kXmaxBytes = 2 * 2 * 128 * 4;        // [parity][sub][lane]
  static constexpr int kXpBytes = (kW / 4) * 2 * 128 * 16;  // [16-byte chunk][sub][lane]
  static constexpr int kSmemBytes = kQBytes + kStages * kChunkBytes + kBarBytes + kXmaxBytes + kXpBytes;
  static constexpr uint32_t kTmemO = 2 * kSCols;
  static_assert(TP == 128 || TP == 256, "score tile");
  static_assert(DV % 256 == 0 && DV <= 512 && H % DV == 0 && kSmemBytes <= 232448 && 2 * kSCols + DV / 2 <= 512,
                "TMEM / shared-memory budget");
};

__device__ __forceinline__ void ld_shared_v4(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr) : "memory");
}

template <int H, int DV, int TP, bool kBf16>
__global__ void __launch_bounds__(kThreads, 1)
hop_fwd_tc64_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_vt,
                    const TcKernelArgs a) {
  using Cfg = Tc64Cfg<H, DV, TP>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = ptx::smem_u32(smem_raw);
  if ((smem_base & 1023u) != 0) __trap();
  const uint32_t q_smem = smem_base;
  const uint32_t ring = q_smem + Cfg::kQBytes;
  const uint32_t bars = ring + Cfg::kStages * kChunkBytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (Cfg::kStages + s); };
  const uint32_t misc = bars + 16u * Cfg::kStages;
  auto s_full = [&](int b) { return misc + 8u * b; };
  auto p_full = [&](int b) { return misc + 32u + 8u * b; };
  auto p_empty = [&](int b) { return misc + 48u + 8u * b; };
  const uint32_t q_ready = misc + 64u;
  const uint32_t tmem_slot = misc + 72u;
  const uint32_t xmax = bars + Cfg::kBarBytes;
  const uint32_t xp = xmax + Cfg::kXmaxBytes;
  auto xmax_at = [&](int parity, int sub, int L) { return xmax + (uint32_t)((((parity * 2 + sub) * 128) + L) * 4); };
  auto xp_at = [&](int c, int sub, int L) { return xp + (uint32_t)((((c * 2 + sub) * 128) + L) * 16); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();
  constexpr uint32_t kArrivers = kSoftmaxThreads * 2;
  const int split = blockIdx.y;
  const int zoff = (int)blockIdx.z * DV;
  const int nt_total = (a.p + TP - 1) / TP;
  const int cid = (int)blockIdx.x / 2;
  const int t0 = split * a.tiles_per_split;
  const int nt = min(a.tiles_per_split, nt_total - t0);

  if (threadIdx.x == 0) {
    if (a.zero_words && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0)
      *reinterpret_cast<uint4*>(a.zero_words) = make_uint4(0u, 0u, 0u, 0u);
    for (int s = 0; s < Cfg::kStages; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(s_full(b), 1);
      ptx::mbar_init(p_full(b), kArrivers); ptx::mbar_init(p_empty(b), 1);
    }
    ptx::mbar_init(q_ready, kArrivers);
    ptx::fence_barrier_init();
    ptx::prefetch_tmap(&tm_k); ptx::prefetch_tmap(&tm_vt);
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
      uint32_t it = 0;
      auto acquire = [&](uint32_t total_bytes) {
        const int s = it % Cfg::kStages;
        ptx::mbar_wait(empty_bar(s), ((it / Cfg::kStages) & 1) ^ 1);
        if (crank == 0) ptx::mbar_arrive_expect_tx(full_bar(s), total_bytes);
        return s;
      };
      auto load_k = [&](int tile) {
        for (int ks = 0; ks < Cfg::kKStages; ++ks) {
          const int s = acquire(2u * Cfg::kKPerStage * Cfg::kKChunkBytes);
          const uint32_t fb = ptx::map_to_cta(full_bar(s), 0);
          for (int c = 0; c < Cfg::kKPerStage; ++c)
            ptx::tma_load_2d_pair(ring + s * kChunkBytes + c * Cfg::kKChunkBytes, &tm_k, fb,
                                  (ks * Cfg::kKPerStage + c) * 64, tile * TP + (int)crank * (TP / 2));
          ++it;
        }
      };
      auto load_b = [&](int tile) {
        for (int kc = 0; kc < Cfg::kPC; ++kc)
          for (int nh = 0; nh < Cfg::kNH; ++nh) {
            const int s = acquire(2u * 128u * 128u);
            ptx::tma_load_2d_pair(ring + s * kChunkBytes, &tm_vt, ptx::map_to_cta(full_bar(s), 0),
                                  tile * TP + kc * 64, zoff + nh * 256 + (int)crank * 128);
            ++it;
          }
      };
      load_k(t0);
      for (int j = 0; j < nt; ++j) {
        if (j + 1 < nt) load_k(t0 + j + 1);
        load_b(t0 + j);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer (leader CTA, one thread)
    if (crank == 0 && ptx::elect_one()) {
      constexpr int fmt = kBf16 ? 1 : 0;
      constexpr uint32_t idesc_qk = ptx::umma_idesc_f16(128, TP, fmt);
      constexpr uint32_t idesc_pv = ptx::umma_idesc_f16(128, 256, fmt);
      uint32_t it = 0;
      ptx::mbar_wait_cluster(q_ready, 0);
      ptx::tc_fence_after();
      auto issue_qk = [&](uint32_t gg) {
        const int b = gg & 1;
        for (int ks = 0; ks < Cfg::kKStages; ++ks) {
          const int s = it % Cfg::kStages;
          ptx::mbar_wait_cluster(full_bar(s), (it / Cfg::kStages) & 1);
          ptx::tc_fence_after();
#pragma unroll
          for (int c = 0; c < Cfg::kKPerStage; ++c) {
            const uint64_t ad = ptx::umma_desc_sw128(q_smem + (ks * Cfg::kKPerStage + c) * Cfg::kQChunkBytes);
            const uint64_t bd = ptx::umma_desc_sw128(ring + s * kChunkBytes + c * Cfg::kKChunkBytes);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma2_f16_ss(tmem_base + b * Cfg::kSCols, ad + 2 * k, bd + 2 * k, idesc_qk, (ks | c | k) != 0);
          }
          ptx::umma2_commit_both(empty_bar(s));
          ++it;
        }
        ptx::umma2_commit_both(s_full(b));
      };
      issue_qk(0);
      for (int j = 0; j < nt; ++j) {
        if (j + 1 < nt) issue_qk(j + 1);
        const int b = j & 1;
        ptx::mbar_wait_cluster(p_full(b), (j >> 1) & 1);
        ptx::tc_fence_after();
        for (int kc = 0; kc < Cfg::kPC; ++kc)
          for (int nh = 0; nh < Cfg::kNH; ++nh) {
            const int s = it % Cfg::kStages;
            ptx::mbar_wait_cluster(full_bar(s), (it / Cfg::kStages) & 1);
            ptx::tc_fence_after();
            const uint32_t at = tmem_base + b * Cfg::kSCols + kc * 32;
            const uint64_t bd = ptx::umma_desc_sw128(ring + s * kChunkBytes);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma2_f16_ts(tmem_base + Cfg::kTmemO + nh * 128, at + 8 * k, bd + 2 * k, idesc_pv, (j | kc | k) != 0);
            ptx::umma2_commit_both(empty_bar(s));
            ++it;
          }
        ptx::umma2_commit_both(p_empty(b));
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------ softmax / epilogue
    const int wq = warp & 3;                      // TMEM lane quarter
    const int sub = (warp - 2) >> 2;              // which half of the lane's TP/2 score columns
    const int L = wq * 32 + lane;                 // TMEM lane
    const int h = L >> 6, row = L & 63;           // pattern half of the tile, query row of the CTA
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    const uint32_t grp_bar = 1 + (wq & 1);        // the four warps that share 32 rows (128 threads)
    const float qscale = a.beta;
    const long long row_tile = (long long)cid * 2 + (long long)crank;
    const long long grow = row_tile * Cfg::kRows + row;
    const bool row_ok = grow < a.n;
    const long long part_row = (long long)split * a.n + grow;
    auto arrive_issuer = [&](uint32_t bar) { ptx::mbar_arrive_cluster(ptx::map_to_cta(bar, 0)); };

    // ---- Q tile: one warp per row, 2 rows in flight; LN_q + beta, fp32 qsave, 16-bit swizzled smem
    {
      constexpr int kVec = (H + 127) / 128;
      constexpr int kRowsInFlight = 2;
      constexpr int kRowsPerWarp = Cfg::kRows / kSoftmaxWarps;
      const int sw = warp - 2;
      float4 w4[kVec], b4[kVec];
#pragma unroll
      for (int v = 0; v < kVec; ++v) {
        const int e = lane * 4 + v * 128;
        w4[v] = make_float4(1.f, 1.f, 1.f, 1.f);
        b4[v] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.x_raw && a.has_ln) {
          if (a.lnq_w) w4[v] = *reinterpret_cast<const float4*>(a.lnq_w + e);
          if (a.lnq_b) b4[v] = *reinterpret_cast<const float4*>(a.lnq_b + e);
        }
      }
      const float* src = a.x_raw ? a.x_raw : a.qs;
      for (int rr = 0; rr < kRowsPerWarp; rr += kRowsInFlight) {
        float4 xv[kRowsInFlight][kVec];
#pragma unroll
        for (int u = 0; u < kRowsInFlight; ++u) {
          const long long gr = row_tile * Cfg::kRows + sw * kRowsPerWarp + rr + u;
#pragma unroll
          for (int v = 0; v < kVec; ++v) {
            const int e = lane * 4 + v * 128;
            xv[u][v] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gr < a.n) xv[u][v] = __ldg(reinterpret_cast<const float4*>(src + gr * H + e));
          }
        }
#pragma unroll
        for (int u = 0; u < kRowsInFlight; ++u) {
          const int r = sw * kRowsPerWarp + rr + u;
          const long long gr = row_tile * Cfg::kRows + r;
          if (a.x_raw) {
            if (a.has_ln) {
              float s = 0.f;
#pragma unroll
              for (int v = 0; v < kVec; ++v) s += (xv[u][v].x + xv[u][v].y) + (xv[u][v].z + xv[u][v].w);
              const float mean = warp_sum(s) * (1.f / H);
              float q = 0.f;
#pragma unroll
              for (int v = 0; v < kVec; ++v) {
                xv[u][v].x -= mean; xv[u][v].y -= mean; xv[u][v].z -= mean; xv[u][v].w -= mean;
                q += xv[u][v].x * xv[u][v].x + xv[u][v].y * xv[u][v].y + xv[u][v].z * xv[u][v].z +
                     xv[u][v].w * xv[u][v].w;
              }
              const float rstd = rsqrtf(warp_sum(q) * (1.f / H) + a.ln_eps);
#pragma unroll
              for (int v = 0; v < kVec; ++v) {
                xv[u][v].x = xv[u][v].x * rstd * w4[v].x + b4[v].x; xv[u][v].y = xv[u][v].y * rstd * w4[v].y + b4[v].y;
                xv[u][v].z = xv[u][v].z * rstd * w4[v].z + b4[v].z; xv[u][v].w = xv[u][v].w * rstd * w4[v].w + b4[v].w;
              }
            }
#pragma unroll
            for (int v = 0; v < kVec; ++v) {
              xv[u][v].x *= qscale; xv[u][v].y *= qscale; xv[u][v].z *= qscale; xv[u][v].w *= qscale;
            }
          }
          if (a.qn != nullptr && t0 == 0 && blockIdx.z == 0) {       // ||q|| of the row: bounds the 16-bit score noise
            float s2 = 0.f;
#pragma unroll
            for (int v = 0; v < kVec; ++v)
              s2 += xv[u][v].x * xv[u][v].x + xv[u][v].y * xv[u][v].y + xv[u][v].z * xv[u][v].z + xv[u][v].w * xv[u][v].w;
            s2 = warp_sum(s2);
            if (lane == 0 && gr < a.n) a.qn[gr] = sqrtf(s2);
          }
#pragma unroll
          for (int v = 0; v < kVec; ++v) {
            const int e = lane * 4 + v * 128;
            if (a.qsave && gr < a.n && t0 == 0 && blockIdx.z == 0)
              *reinterpret_cast<float4*>(a.qsave + gr * H + e) = xv[u][v];
            const uint32_t lo = pack16<kBf16>(xv[u][v].x * kLog2e, xv[u][v].y * kLog2e);
            const uint32_t hi = pack16<kBf16>(xv[u][v].z * kLog2e, xv[u][v].w * kLog2e);
            const uint32_t addr = q_smem + (e >> 6) * Cfg::kQChunkBytes + sw128_off(r, (e & 63) >> 3) + ((e & 4) << 1);
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(lo), "r"(hi) : "memory");
          }
        }
      }
    }
    ptx::fence_proxy_async_smem();
    arrive_issuer(q_ready);

    float m_run = -INFINITY, l = 0.f, tmax = -INFINITY, tsec = -INFINITY, top_sum = 0.f, top_ref = 0.f;
    int targ = 0;
    constexpr int kOGroups = DV / 4 / 32;          // 32-column TMEM groups of O handled by this thread
    const int og0 = sub * kOGroups;
    for (int j = 0; j < nt; ++j) {
      const int b = j & 1;
      constexpr int kC = Cfg::kC, kW = Cfg::kW;
      const int col0 = (t0 + j) * TP + h * (TP / 2) + sub * kC;  // global pattern index of this thread's column 0
      const int valid = a.p - col0;
      ptx::mbar_wait(s_full(b), (j >> 1) & 1);
      ptx::tc_fence_after();
      uint32_t v[kC];
#pragma unroll
      for (int c = 0; c < kC; c += 32) ptx::tmem_ld_x32p(lane_base + b * Cfg::kSCols + sub * kC + c, v + c);
      ptx::tmem_wait_ld();
      if (valid < kC) {
#pragma unroll
        for (int i = 0; i < kC; ++i)
          if (i >= valid) v[i] = __float_as_uint(-INFINITY);
      }
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < kC; i += 8) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
          mx[u] = fmaxf(mx[u], fmaxf(__uint_as_float(v[i + 2 * u]), __uint_as_float(v[i + 2 * u + 1])));
      }
      const float hmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      st_shared_f32(xmax_at(b, sub, L), hmax);
      const bool moved = hmax > tmax;                // runner-up tracking as in hop_fwd_tc_kernel
      tsec = fmaxf(tsec, moved ? tmax : hmax);
      if (moved) {
        int idx = kC - 1;
#pragma unroll
        for (int i = kC - 1; i >= 0; --i) idx = (__uint_as_float(v[i]) == hmax) ? i : idx;
        tmax = hmax;
        targ = col0 + idx;
      }
      ptx::named_bar_sync(grp_bar, 128);
      const float cmax = fmaxf(fmaxf(ld_shared_f32(xmax_at(b, 0, row)), ld_shared_f32(xmax_at(b, 1, row))),
                               fmaxf(ld_shared_f32(xmax_at(b, 0, row + 64)), ld_shared_f32(xmax_at(b, 1, row + 64))));
      const bool need = cmax > m_run + kRescaleThreshold;
      const bool any_need = __any_sync(0xffffffffu, need);
      const float m_old = m_run;
      if (any_need) m_run = fmaxf(m_run, cmax);

      float ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < kC; ++i) {
        const float xs = __uint_as_float(v[i]) - m_run;
        const bool poly = Cfg::kPolyPeriod > 0 && (i % (Cfg::kPolyPeriod > 0 ? Cfg::kPolyPeriod : 1)) == 0;
        v[i] = __float_as_uint(poly ? ptx::ex2_poly(xs) : ptx::ex2(xs));
      }
      uint32_t pk[kW];
#pragma unroll
      for (int i = 0; i < kW; ++i) {
        const float p0 = __uint_as_float(v[2 * i]), p1 = __uint_as_float(v[2 * i + 1]);
        ls[i & 3] += p0 + p1;
        pk[i] = pack16<kBf16>(p0, p1);
      }
#pragma unroll
      for (int c = 0; c < kW / 4; ++c) st_shared_v4(xp_at(c, sub, L), pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      if (any_need && j > 0) {
        const uint32_t pg = j - 1;
        ptx::mbar_wait(p_empty(pg & 1), (pg >> 1) & 1);   // read-out MMA of the previous tile done
        ptx::tc_fence_after();
        const float f = ptx::ex2(m_old - m_run);
        l *= f;
#pragma unroll 1
        for (int c = og0; c < og0 + kOGroups; ++c) {
          uint32_t o[32];
          ptx::tmem_ld_x32(lane_base + Cfg::kTmemO + c * 32, o);
          ptx::tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * f);
          ptx::tmem_st_x32(lane_base + Cfg::kTmemO + c * 32, o);
        }
      }
      ptx::named_bar_sync(grp_bar, 128);
      uint32_t pw[kW];
#pragma unroll
      for (int c = 0; c < kW / 4; ++c) ld_shared_v4(xp_at(c, sub, L ^ 64), pw[4 * c], pw[4 * c + 1], pw[4 * c + 2], pw[4 * c + 3]);
      // P~ row of this lane: patterns (TP/2) h' + kC sub' + i -> column (TP/4) h' + kW sub' + i / 2
      const uint32_t p_own = lane_base + b * Cfg::kSCols + h * (TP / 4) + sub * kW;
      const uint32_t p_oth = lane_base + b * Cfg::kSCols + (h ^ 1) * (TP / 4) + sub * kW;
      if constexpr (kW == 16) { ptx::tmem_st_x16(p_own, pk); ptx::tmem_st_x16(p_oth, pw); }
      else { ptx::tmem_st_x32(p_own, pk); ptx::tmem_st_x32(p_oth, pw); }
      const float tsum = (ls[0] + ls[1]) + (ls[2] + ls[3]);
      top_sum = moved ? tsum : top_sum;
      top_ref = moved ? m_run : top_ref;
      l += tsum;
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      arrive_issuer(p_full(b));
    }

    // ---- end of the pass: the four threads of a row combine (l, tmax, targ) in pattern order
    ptx::named_bar_sync(grp_bar, 128);             // every P~ word of the last tile has been read
    const float ttop = second_bound(tmax, top_sum, top_ref);    // (tsec stays the exact maximum outside the top tile)
    st_shared_v4(xp_at(0, sub, L), __float_as_uint(l), __float_as_uint(tmax), (uint32_t)targ, __float_as_uint(tsec));
    st_shared_v4(xp_at(1, sub, L), __float_as_uint(ttop), 0u, 0u, 0u);
    ptx::named_bar_sync(grp_bar, 128);
    float l_tot = 0.f, best = -INFINITY, sec = -INFINITY, top = -INFINITY;
    int best_arg = 0;
#pragma unroll
    for (int hh = 0; hh < 2; ++hh)
#pragma unroll
      for (int ss = 0; ss < 2; ++ss) {
        uint32_t w0, w1, w2, w3;
        ld_shared_v4(xp_at(0, ss, row + 64 * hh), w0, w1, w2, w3);
        l_tot += __uint_as_float(w0);
        const float om = __uint_as_float(w1);
        if (om > best) {                                     // pattern order: ties keep the lowest index
          uint32_t t0w, t1w, t2w, t3w;
          ld_shared_v4(xp_at(1, ss, row + 64 * hh), t0w, t1w, t2w, t3w);
          sec = fmaxf(__uint_as_float(w3), best); top = __uint_as_float(t0w);
          best = om; best_arg = (int)w2;
        } else {
          sec = fmaxf(sec, om);
        }
      }
    const uint32_t last = nt - 1;
    ptx::mbar_wait(p_empty(last & 1), (last >> 1) & 1);
    ptx::tc_fence_after();
    const float inv = 1.f / l_tot;
    float* o_row = (a.splits > 1) ? a.part_o + part_row * a.dv_total : a.retrieved + grow * a.dv_total;
    const float osc = (a.splits > 1) ? inv : inv * a.out_scale;
#pragma unroll 1
    for (int c = og0; c < og0 + kOGroups; ++c) {
      uint32_t o[32];
      ptx::tmem_ld_x32(lane_base + Cfg::kTmemO + c * 32, o);
      ptx::tmem_wait_ld();
      if (row_ok) {
        const int tcol = c * 32;                               // TMEM column of O: instruction nh, column jj of its half
        const int ocol = zoff + (tcol >> 7) * 256 + h * 128 + (tcol & 127);
        float4* dst = reinterpret_cast<float4*>(o_row + ocol);
        const float4* bia = (a.bias && a.splits == 1) ? reinterpret_cast<const float4*>(a.bias + ocol) : nullptr;
#pragma unroll
        for (int q4 = 0; q4 < 8; ++q4) {
          float4 r = make_float4(__uint_as_float(o[4 * q4]) * osc, __uint_as_float(o[4 * q4 + 1]) * osc,
                                 __uint_as_float(o[4 * q4 + 2]) * osc, __uint_as_float(o[4 * q4 + 3]) * osc);
          if (bia) { const float4 bb = __ldg(bia + q4); r.x += bb.x; r.y += bb.y; r.z += bb.z; r.w += bb.w; }
          dst[q4] = r;
        }
      }
    }
    if (row_ok && h == 0 && sub == 0 && blockIdx.z == 0) {
      const float lse = kLn2 * (m_run + ptx::lg2(l_tot));
      if (a.splits > 1) {
        a.part_lse[part_row] = lse; a.part_max[part_row] = best * kLn2; a.part_arg[part_row] = best_arg + a.pattern_offset;
        a.part_sec[part_row] = sec * kLn2; a.part_top[part_row] = top * kLn2;
      } else {
        if (a.lse) a.lse[grow] = lse;
        if (a.rowmax) a.rowmax[grow] = best * kLn2;
        if (a.rowsec) { a.rowsec[grow] = sec * kLn2; a.rowtop[grow] = top * kLn2; }
        if (a.argmax) a.argmax[grow] = best_arg + a.pattern_offset;
      }
    }
    ptx::tc_fence_before();
  }

  __syncthreads();
  ptx::cluster_sync();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair(tmem_base, kTmemCols);
  }
}

// The exact-argmax certificate of hop_repair.cu, evaluated where the row maximum and the runner-up bound are already
// in registers.
__device__ __forceinline__ void certify_row(const CertArgs& c, long long r, float rowmax, float others, float top) {
  if (c.ctrl == nullptr) return;
  const float tau = 2.f * c.qn[r] * __ldg(c.kstats) * (1.f / 1024.f) * 1.02f + 1e-5f;
  if (rowmax - fmaxf(others, top) > tau) return;                       // every other score is more than tau below
  c.sec[r] = others;                                                   // for the block check of hop_repair.cu
  const uint32_t slot = atomicAdd(c.ctrl, 1u);
  c.list[slot] = (int32_t)r;
  c.best[slot] = 0ull;
}

// Combine `splits` partial results of one row (flash-decoding merge): weights exp(lse_s - lse).
template <int D>
__global__ void tc_combine_kernel(int n, int splits, const float* __restrict__ part_o,
                                  const float* __restrict__ part_lse, const float* __restrict__ part_max,
                                  const int32_t* __restrict__ part_arg, const float* __restrict__ part_sec, const float* __restrict__ part_top,
                                  const float* __restrict__ bias, float out_scale, float* __restrict__ retrieved, float* __restrict__ lse,
                                  int32_t* __restrict__ argmax, float* __restrict__ rowmax, const CertArgs cert) {
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp_global >= n) return;
  const long long r = warp_global;
  float m = -INFINITY;
  for (int s = 0; s < splits; ++s) m = fmaxf(m, part_lse[(long long)s * n + r]);
  float z = 0.f;
  for (int s = 0; s < splits; ++s) z += __expf(part_lse[(long long)s * n + r] - m);
  const float inv = 1.f / z;
  constexpr int kVec = (D + 127) / 128;
  float4 acc[kVec];
#pragma unroll
  for (int v = 0; v < kVec; ++v) {
    const int e = lane * 4 + v * 128;
    acc[v] = (bias && e < D) ? __ldg(reinterpret_cast<const float4*>(bias + e)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float best = -INFINITY, sec = -INFINITY, top = -INFINITY;
  int arg = 0;
  for (int s = 0; s < splits; ++s) {
    const long long o = (long long)s * n + r;
    const float w = __expf(part_lse[o] - m) * inv * out_scale;
    const float pm = part_max[o];
    if (pm > best) { sec = fmaxf(part_sec[o], best); top = part_top[o]; best = pm; arg = part_arg[o]; }
    else sec = fmaxf(sec, pm);                         // splits are in pattern order: ties keep the lowest index
#pragma unroll
    for (int v = 0; v < kVec; ++v) {
      const int e = lane * 4 + v * 128;
      if (e < D) {
        const float4 t = *reinterpret_cast<const float4*>(part_o + o * D + e);
        acc[v].x += w * t.x; acc[v].y += w * t.y; acc[v].z += w * t.z; acc[v].w += w * t.w;
      }
    }
  }
#pragma unroll
  for (int v = 0; v < kVec; ++v) {
    const int e = lane * 4 + v * 128;
    if (e < D) *reinterpret_cast<float4*>(retrieved + r * D + e) = acc[v];
  }
  if (lane == 0) {
    if (lse) lse[r] = m + __logf(z);
    if (rowmax) rowmax[r] = best;
    if (argmax) argmax[r] = arg;
    certify_row(cert, r, best, sec, top);
  }
}

// Stream-K form of the merge: the partials of row group rg live in the slots of the clusters whose unit shares
// intersect [rg * nt_total, (rg + 1) * nt_total) -- one segment each, in pattern order.
struct StreamKPlan { int nt_total, base, rem, max_seg, cluster; };

__device__ __forceinline__ int sk_cluster_of(const StreamKPlan& s, int u) {
  const int big = s.rem * (s.base + 1);
  return u < big ? u / (s.base + 1) : s.rem + (u - big) / s.base;
}

template <int D>
__global__ void tc_combine_streamk_kernel(int n, StreamKPlan sk, const float* __restrict__ part_o,
                                          const float* __restrict__ part_lse, const float* __restrict__ part_max,
                                          const int32_t* __restrict__ part_arg, const float* __restrict__ part_sec, const float* __restrict__ part_top,
                                          const float* __restrict__ bias, float out_scale, float* __restrict__ retrieved, float* __restrict__ lse,
                                          int32_t* __restrict__ argmax, float* __restrict__ rowmax, const CertArgs cert) {
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp_global >= n) return;
  const long long r = warp_global;
  const int row_tile = warp_global / kTileM;
  const int rg = row_tile / sk.cluster;
  const int in_rg = (row_tile - rg * sk.cluster) * kTileM + (warp_global - row_tile * kTileM);
  const int c0 = sk_cluster_of(sk, rg * sk.nt_total), c1 = sk_cluster_of(sk, (rg + 1) * sk.nt_total - 1);
  auto slot_row = [&](int c) -> long long {
    const int start = c * sk.base + min(c, sk.rem);
    const int local = rg - start / sk.nt_total;
    return ((long long)c * sk.max_seg + local) * (sk.cluster * kTileM) + in_rg;
  };
  float m = -INFINITY;
  for (int c = c0; c <= c1; ++c) m = fmaxf(m, part_lse[slot_row(c)]);
  float z = 0.f;
  for (int c = c0; c <= c1; ++c) z += __expf(part_lse[slot_row(c)] - m);
  const float inv = 1.f / z;
  constexpr int kVec = (D + 127) / 128;
  float4 acc[kVec];
#pragma unroll
  for (int v = 0; v < kVec; ++v) {
    const int e = lane * 4 + v * 128;
    acc[v] = (bias && e < D) ? __ldg(reinterpret_cast<const float4*>(bias + e)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float best = -INFINITY, sec = -INFINITY, top = -INFINITY;
  int arg = 0;
  for (int c = c0; c <= c1; ++c) {
    const long long o = slot_row(c);
    const float w = __expf(part_lse[o] - m) * inv * out_scale;
    const float pm = part_max[o];
    if (pm > best) { sec = fmaxf(part_sec[o], best); top = part_top[o]; best = pm; arg = part_arg[o]; }
    else sec = fmaxf(sec, pm);                         // segments are in pattern order: ties keep the lowest index
#pragma unroll
    for (int v = 0; v < kVec; ++v) {
      const int e = lane * 4 + v * 128;
      if (e < D) {
        const float4 t = *reinterpret_cast<const float4*>(part_o + o * D + e);
        acc[v].x += w * t.x; acc[v].y += w * t.y; acc[v].z += w * t.z; acc[v].w += w * t.w;
      }
    }
  }
#pragma unroll
  for (int v = 0; v < kVec; ++v) {
    const int e = lane * 4 + v * 128;
    if (e < D) *reinterpret_cast<float4*>(retrieved + r * D + e) = acc[v];
  }
  if (lane == 0) {
    if (lse) lse[r] = m + __logf(z);
    if (rowmax) rowmax[r] = best;
    if (argmax) argmax[r] = arg;
    certify_row(cert, r, best, sec, top);
  }
}

// ------------------------------------------------------------------ host side
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
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

// 2-D row-major 16-bit matrix (rows, cols) -> boxes of (box_rows x 64 columns), 128-byte swizzle.
int make_map(CUtensorMap* m, const void* base, int rows, int cols, int box_rows, int bf16) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) HOP_FAIL(HOP_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = fn(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                        const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) HOP_FAIL(HOP_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for a %d x %d matrix", (int)r, rows, cols);
  return HOP_OK;
}

// internal per-launch switches of launch_tc_one (not part of TcPassArgs: they belong to one launch of a call)
struct TcLaunchOpt {
  bool precise;       // hi + lo operand pairs (update passes of T >= 2)
  bool allow_streamk; // false for every launch of the one-launch-per-pass form (its workspace is sized for the plain grid
                      // and its accumulation chains are capped at kMaxChainTilesUpdate tiles)
  uint32_t* zero_words;
};

template <int H, int DV, bool kBf16, int kCluster, bool kPair, bool kPrecise>
int launch_d(const TcPassArgs& a, int splits, const TcLaunchOpt& opt, cudaStream_t st) {
  using Cfg = TcCfg<H, DV, kPrecise>;
  const int ppad = padded_patterns(a.p);
  if (a.update_steps > 0 && H != DV) HOP_FAIL(HOP_ERR_UNSUPPORTED, "on-chip retrieve updates need H == DV (internal)");
  if (kPrecise && (a.update_steps > 0 || !a.k16lo || !a.kt16lo))
    HOP_FAIL(HOP_ERR_BAD_ARG, "hi + lo update passes are separate launches and need the k16lo / kt16lo packs");
  CUtensorMap tk, tkt, tvt, tklo, tblo;
  // TMA boxes: pair = this CTA's half of a tile (64 patterns of K^, DV/2 rows of V^T); else 1/kCluster of a stage
  const int k_rows = kPair ? kTileP / 2 : kTileP / kCluster;
  const int b_rows = kPair ? DV / 2 : Cfg::kBRows / kCluster;
  HOP_TRY(make_map(&tk, a.k16, ppad, H, k_rows, a.bf16));
  HOP_TRY(make_map(&tvt, a.vt16, a.dv, ppad, b_rows, a.bf16));
  if (a.update_steps > 0) HOP_TRY(make_map(&tkt, a.kt16, H, ppad, b_rows, a.bf16));
  else tkt = tvt;
  if (kPrecise) {
    HOP_TRY(make_map(&tklo, a.k16lo, ppad, H, k_rows, a.bf16));
    HOP_TRY(make_map(&tblo, a.kt16lo, a.dv, ppad, b_rows, a.bf16));   // (a precise launch reads out over K^: vt16 = kt16)
  } else {
    tklo = tk; tblo = tvt;
  }

  TcKernelArgs k{};
  k.x_raw = a.x_raw; k.lnq_w = a.lnq_w; k.lnq_b = a.lnq_b; k.ln_eps = a.ln_eps; k.has_ln = a.has_ln; k.beta = a.beta;
  k.qs = a.qs; k.bias = a.bias; k.n = a.n; k.p = a.p; k.update_steps = a.update_steps;
  k.dv_total = a.dv; k.out_scale = a.out_scale;
  const int nt_total = ppad / kTileP;
  k.tiles_per_split = (nt_total + splits - 1) / splits;
  k.splits = (nt_total + k.tiles_per_split - 1) / k.tiles_per_split;   // every split owns >= 1 tile
  k.retrieved = a.retrieved; k.qsave = a.qsave; k.lse = a.lse; k.argmax = a.argmax; k.rowmax = a.rowmax; k.rowsec = a.argmax ? a.cert.sec : nullptr; k.rowtop = a.argmax ? a.cert.top : nullptr;
  k.pattern_offset = a.pattern_offset;
  k.zero_words = opt.zero_words;
  const CertArgs cert = (a.argmax && a.lse && a.rowmax) ? a.cert : CertArgs{};
  k.qn = cert.ctrl ? cert.qn : nullptr;
  auto* kern = hop_fwd_tc_kernel<H, DV, kBf16, kCluster, kPair, kPrecise>;
  const int row_tiles = (a.n + Cfg::kRows - 1) / Cfg::kRows;
  const int row_groups = (row_tiles + kCluster - 1) / kCluster;

  // ---- stream-K form: one contiguous share of the (row group x tile) units per resident cluster
  StreamKPlan sk{};
  constexpr int kZ = H / DV;           // read-out column chunks: every z-plane walks the same partition
  bool streamk = !kPrecise && opt.allow_streamk && a.update_steps == 0 &&
                 streamk_wanted(a.n, a.p, a.h, a.update_steps, a.tune);
  if (streamk) {
    HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    cudaLaunchConfig_t oc{};
    oc.gridDim = dim3(num_sms() / kCluster * kCluster, 1, 1);
    oc.blockDim = dim3(kThreads);
    oc.dynamicSmemBytes = Cfg::kSmemBytes;
    cudaLaunchAttribute oattr[1];
    oattr[0].id = cudaLaunchAttributeClusterDimension;
    oattr[0].val.clusterDim.x = kCluster; oattr[0].val.clusterDim.y = 1; oattr[0].val.clusterDim.z = 1;
    oc.attrs = oattr; oc.numAttrs = 1;
    // resident clusters of this kernel on this device: queried once (the query costs tens of microseconds)
    static int resident_cache[64];
    int dev = 0;
    HOP_CUDA(cudaGetDevice(&dev));
    int resident = (dev >= 0 && dev < 64) ? resident_cache[dev] : 0;
    if (resident == 0) {
      if (cudaOccupancyMaxActiveClusters(&resident, kern, &oc) != cudaSuccess) { (void)cudaGetLastError(); resident = 0; }
      if (dev >= 0 && dev < 64) resident_cache[dev] = resident;
    }
    const int cap = num_sms() / kCluster;
    int n_clusters = (resident < cap ? resident : cap) / kZ;
    const long long units = (long long)row_groups * nt_total;
    if (n_clusters > units) n_clusters = (int)units;
    if (n_clusters < 1) streamk = false;
    if (streamk) {
      sk.nt_total = nt_total; sk.cluster = kCluster;
      sk.base = (int)(units / n_clusters); sk.rem = (int)(units % n_clusters);
      sk.max_seg = (sk.base + 1 + nt_total - 1) / nt_total + 1;
      const size_t rows = (size_t)n_clusters * sk.max_seg * kCluster * kTileM;
      const size_t need = round_up_sz(rows * a.dv * sizeof(float), 256) + 5 * round_up_sz(rows * sizeof(float), 256);
      if (!a.ws || a.ws_bytes < need) HOP_FAIL(HOP_ERR_WORKSPACE, "stream-K forward needs %zu workspace bytes, %zu given", need, a.ws_bytes);
      char* w = static_cast<char*>(a.ws);
      k.part_o = reinterpret_cast<float*>(w); w += round_up_sz(rows * a.dv * sizeof(float), 256);
      k.part_lse = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
      k.part_max = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
      k.part_arg = reinterpret_cast<int32_t*>(w); w += round_up_sz(rows * sizeof(int32_t), 256);
    k.part_sec = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
    k.part_top = reinterpret_cast<float*>(w);
      k.streamk = 1; k.sk_base = sk.base; k.sk_rem = sk.rem; k.sk_max_seg = sk.max_seg;
      k.splits = 2;                       // (> 1: every segment leaves a partial for the merge)
      cudaLaunchConfig_t cfg = oc;
      cfg.gridDim = dim3(n_clusters * kCluster, 1, kZ);
      cfg.stream = st;
      HOP_CUDA(cudaLaunchKernelEx(&cfg, kern, tk, tkt, tvt, tklo, tblo, k));
      HOP_LAUNCH_CHECK("hop_fwd_tc_kernel");
      const int threads = 256, rows_per_block = threads / 32;
      tc_combine_streamk_kernel<H><<<(a.n + rows_per_block - 1) / rows_per_block, threads, 0, st>>>(
          a.n, sk, k.part_o, k.part_lse, k.part_max, k.part_arg, k.part_sec, k.part_top, a.bias, a.out_scale, a.retrieved, a.lse, a.argmax,
          a.rowmax, cert);
      HOP_LAUNCH_CHECK("tc_combine_streamk_kernel");
      if (cert.ctrl && a.cert_done) *a.cert_done = true;
      return HOP_OK;
    }
  }
  if (k.splits > 1) {
    const size_t need = ws_for_splits_fwd(splits, a.n, a.dv);
    if (!a.ws || a.ws_bytes < need)
      HOP_FAIL(HOP_ERR_WORKSPACE, "tensor-core forward needs %zu workspace bytes for %d pattern splits, %zu given", need,
               k.splits, a.ws_bytes);
    char* w = static_cast<char*>(a.ws);
    const size_t rows = (size_t)splits * a.n;
    k.part_o = reinterpret_cast<float*>(w); w += round_up_sz(rows * a.dv * sizeof(float), 256);
    k.part_lse = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
    k.part_max = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
    k.part_arg = reinterpret_cast<int32_t*>(w); w += round_up_sz(rows * sizeof(int32_t), 256);
    k.part_sec = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
    k.part_top = reinterpret_cast<float*>(w);
  }

  // the attribute is per device and cheap to set: do it on every launch so that any current device works
  HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  cudaLaunchConfig_t cfg{};
  // whole clusters (extra CTAs are masked) x pattern splits x read-out column chunks
  cfg.gridDim = dim3((row_tiles + kCluster - 1) / kCluster * kCluster, k.splits, a.dv / DV);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  HOP_CUDA(cudaLaunchKernelEx(&cfg, kern, tk, tkt, tvt, tklo, tblo, k));
  HOP_LAUNCH_CHECK("hop_fwd_tc_kernel");
  if (k.splits > 1) {
    const int threads = 256, rows_per_block = threads / 32;
    tc_combine_kernel<H><<<(a.n + rows_per_block - 1) / rows_per_block, threads, 0, st>>>(
        a.n, k.splits, k.part_o, k.part_lse, k.part_max, k.part_arg, k.part_sec, k.part_top, a.bias, a.out_scale, a.retrieved, a.lse, a.argmax,
        a.rowmax, cert);
    HOP_LAUNCH_CHECK("tc_combine_kernel");
    if (cert.ctrl && a.cert_done) *a.cert_done = true;
  }
  return HOP_OK;
}

// 64-rows-per-CTA pair form (hop_fwd_tc64_kernel): plain grid (row pairs x pattern splits x column chunks), T = 0
template <int H, int DV, int TP, bool kBf16>
int launch_h64(const TcPassArgs& a, int splits, const TcLaunchOpt& opt, cudaStream_t st) {
  using Cfg = Tc64Cfg<H, DV, TP>;
  const int ppad = padded_patterns(a.p);
  if (a.update_steps > 0) HOP_FAIL(HOP_ERR_UNSUPPORTED, "the 64-row form runs retrieve updates as separate launches (internal)");
  CUtensorMap tk, tvt;
  HOP_TRY(make_map(&tk, a.k16, ppad, H, TP / 2, a.bf16));         // this CTA's TP/2 patterns of a K^ chunk
  HOP_TRY(make_map(&tvt, a.vt16, a.dv, ppad, 128, a.bf16));       // this CTA's 128 rows of an N = 256 read-out chunk
  TcKernelArgs k{};
  k.x_raw = a.x_raw; k.lnq_w = a.lnq_w; k.lnq_b = a.lnq_b; k.ln_eps = a.ln_eps; k.has_ln = a.has_ln; k.beta = a.beta;
  k.qs = a.qs; k.bias = a.bias; k.n = a.n; k.p = a.p; k.update_steps = 0;
  k.dv_total = a.dv; k.out_scale = a.out_scale;
  const int nt_total = (ppad + TP - 1) / TP;          // (rows past ppad read as zeros and are masked)
  k.tiles_per_split = (nt_total + splits - 1) / splits;
  k.splits = (nt_total + k.tiles_per_split - 1) / k.tiles_per_split;
  k.retrieved = a.retrieved; k.qsave = a.qsave; k.lse = a.lse; k.argmax = a.argmax; k.rowmax = a.rowmax; k.rowsec = a.argmax ? a.cert.sec : nullptr; k.rowtop = a.argmax ? a.cert.top : nullptr;
  k.pattern_offset = a.pattern_offset;
  k.zero_words = opt.zero_words;
  const CertArgs cert = (a.argmax && a.lse && a.rowmax) ? a.cert : CertArgs{};
  k.qn = cert.ctrl ? cert.qn : nullptr;
  if (k.splits > 1) {
    const size_t need = ws_for_splits_fwd(splits, a.n, a.dv);
    if (!a.ws || a.ws_bytes < need)
      HOP_FAIL(HOP_ERR_WORKSPACE, "tensor-core forward needs %zu workspace bytes for %d pattern splits, %zu given", need,
               k.splits, a.ws_bytes);
    char* w = static_cast<char*>(a.ws);
    const size_t rows = (size_t)splits * a.n;
    k.part_o = reinterpret_cast<float*>(w); w += round_up_sz(rows * a.dv * sizeof(float), 256);
    k.part_lse = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
    k.part_max = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
    k.part_arg = reinterpret_cast<int32_t*>(w); w += round_up_sz(rows * sizeof(int32_t), 256);
    k.part_sec = reinterpret_cast<float*>(w); w += round_up_sz(rows * sizeof(float), 256);
    k.part_top = reinterpret_cast<float*>(w);
  }
  auto* kern = hop_fwd_tc64_kernel<H, DV, TP, kBf16>;
  HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  const int row_tiles = (a.n + Cfg::kRows - 1) / Cfg::kRows;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((row_tiles + 1) / 2 * 2, k.splits, a.dv / DV);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  HOP_CUDA(cudaLaunchKernelEx(&cfg, kern, tk, tvt, k));
  HOP_LAUNCH_CHECK("hop_fwd_tc64_kernel");
  if (k.splits > 1) {
    const int threads = 256, rows_per_block = threads / 32;
    tc_combine_kernel<H><<<(a.n + rows_per_block - 1) / rows_per_block, threads, 0, st>>>(
        a.n, k.splits, k.part_o, k.part_lse, k.part_max, k.part_arg, k.part_sec, k.part_top, a.bias, a.out_scale, a.retrieved, a.lse, a.argmax,
        a.rowmax, cert);
    HOP_LAUNCH_CHECK("tc_combine_kernel");
    if (cert.ctrl && a.cert_done) *a.cert_done = true;
  }
  return HOP_OK;
}

}  // namespace

#ifdef HOP_TC_TRACE
extern "C" int hop_sm100_debug_trace(long long* dst_host, int n) {
  const int total = kTraceRoles * kTraceTiles * kTraceEvents;
  return (int)cudaMemcpyFromSymbol(dst_host, g_trace, sizeof(long long) * (n < total ? n : total));
}
#endif

bool tc_supported(int h, int dv) { return h == dv && (h == 64 || h == 128 || h == 256 || h == 512 || h == 1024); }

// D >= 512 runs the 64-rows-per-CTA pair form (hop_fwd_tc64_kernel); HOP_TUNE_ROWS128 keeps the 128-row kernel with its
// recomputed score tiles (D = 512) / discarded half-MMAs (D = 1024) for comparison.
static inline bool tc_halfm(int d, const TcTune& t) { return d >= 512 && !(t.flags & HOP_TUNE_ROWS128); }

// query rows per CTA: 64 in the pair form above and under the 1024-wide Q tile of the 128-row kernel (TcCfg::kRows)
static inline int tc_rows(int d, const TcTune& t) { return (tc_halfm(d, t) || d > 512) ? 64 : kTileM; }

// read-out column chunks (gridDim.z): the O accumulator of one CTA holds at most 256 fp32 TMEM columns -- 256
// read-out columns in the 128-row kernel, 512 in the 64-row pair form
static inline int tc_dv_chunks(int d, const TcTune& t) { return tc_halfm(d, t) ? d / 512 : (d > 256 ? d / 256 : 1); }

// Update passes with hi + lo operand pairs (TcCfg<.., kPrecise>): default for T >= 2 at D <= 256.  The retrieve
// iteration amplifies operand rounding by its gain beta w (1 - w) |k|^2 per pass (w = softmax weight of the nearest
// pattern): at the bifurcation of BASELINE config 5's recipe (P = 256k, D = 256: w ~ 0.79, gain ~ 2.7) plain 16-bit
// operands end at 3.3e-3 after three updates, the pairs at 2e-4 (CPU emulation of both: DESIGN.md 4.1).  D >= 512 is
// contractive in that recipe (w ~ 0.997) and keeps plain operands.
bool tc_precise_updates(int d, int update_steps, const TcTune& t) {
  return update_steps >= 2 && d <= 256 && !(t.flags & HOP_TUNE_PLAIN_UPDATES);
}

// Retrieve updates run as one launch per pass (state through the fp32 workspace buffers, every pass free to
// split the pattern axis) instead of on-chip in one persistent launch.  Always at D >= 512 (a CTA holds part of the
// state columns) and for the hi + lo update passes.  HOP_TUNE_MULTI_ON / _OFF force it for experiments.
//
// Also for D <= 256 once one CTA would sweep more than 512 pattern tiles (P > 64k): the tensor core adds into its
// fp32 accumulator with truncation, a bias that grows with the length of the accumulation chain (measured: the
// state shrinks by ~1e-3 over 2,048 tiles) and that the retrieve iteration amplifies near its bifurcation
// (P = 256k, D = 256: rel-L2 2.1e-3 / 9.3e-3 at T = 1 / 3 on-chip).  Short per-split chains, merged in fp32 by
// the combine kernel, keep it at the operand-rounding level.
constexpr int kMaxTilesOnChipUpdates = 512;
constexpr int kMaxChainTilesUpdate = 256;    // pattern tiles one CTA accumulates when T > 0 (multi-launch form)
constexpr int kMaxChainTilesPrecise = 32;    // ... in a hi + lo update pass: with the operand rounding gone, the truncation
                                             // bias of the chain (~2^-25 per accumulate instruction, 16 per tile) is what the
                                             // iteration amplifies -- measured at P = 256k, D = 256, T = 3: 1.27e-3 with
                                             // 256-tile chains = 1.2e-4 per pass x (1 + 2.7 + 2.7^2)
constexpr int kMaxSplitsPrecise = 64;        // (bounds the partial-result traffic: P = 1M keeps 128-tile chains)
constexpr int kMaxChainTilesReadout = 512;   // ... when T = 0 (errors are not amplified)
static inline bool tc_multi_launch(int p, int d, int update_steps, const TcTune& t) {
  if (update_steps <= 0) return false;
  if (d >= 512 || tc_precise_updates(d, update_steps, t)) return true;
  if (t.flags & HOP_TUNE_MULTI_ON) return true;
  if (t.flags & HOP_TUNE_MULTI_OFF) return false;
  return padded_patterns(p) / kTileP > kMaxTilesOnChipUpdates;
}

// tcgen05 CTA pairs (cta_group::2): on for D = 256 when there are at least two row tiles (HOP_TUNE_PAIR_ON / _OFF
// force it for every D <= 256).  Measured (r01, P = 64k / 256k): D = 256 1174 -> 1220 / 1226 -> 1367 TFLOP/s; D = 128
// and 64 are MUFU-bound and do not gain (824 -> 798, 440 -> 438).
bool tc_use_pair(int n, int d, const TcTune& t) {
  if (tc_halfm(d, t)) return true;
  if (t.flags & HOP_TUNE_PAIR_ON) return n > tc_rows(d, t);
  if (t.flags & HOP_TUNE_PAIR_OFF) return false;
  return n > tc_rows(d, t) && d >= 256;
}
static inline int tc_cluster(int n, int d, const TcTune& t) { return tc_use_pair(n, d, t) ? 2 : 1; }

// Pattern splits per 128-row tile: the smallest count that fills >= 95 % of the last wave of SMs
// (on-chip retrieve updates need the full softmax per pass, so T > 0 never splits there; the one-launch-per-pass
// form may split every launch).
int tc_choose_splits(int n, int p, int d, int update_steps, const TcTune& t) {
  if (update_steps > 0 && !tc_multi_launch(p, d, update_steps, t)) return 1;
  const int nt_all = padded_patterns(p) / kTileP;
  if (t.splits >= 1) return t.splits < nt_all ? t.splits : nt_all;      // forced (tests)
  const int sms = num_sms();
  const int cl = tc_cluster(n, d, t);
  const int row_tiles = ((n + tc_rows(d, t) - 1) / tc_rows(d, t) + cl - 1) / cl * cl * tc_dv_chunks(d, t);
  const int nt_total = padded_patterns(p) / kTileP;
  int chain = update_steps > 0 ? kMaxChainTilesUpdate : kMaxChainTilesReadout;
  if (tc_precise_updates(d, update_steps, t)) {
    chain = kMaxChainTilesPrecise;
    if ((nt_total + chain - 1) / chain > kMaxSplitsPrecise) chain = (nt_total + kMaxSplitsPrecise - 1) / kMaxSplitsPrecise;
  }
  const int min_splits = (nt_total + chain - 1) / chain;            // accuracy: bound the accumulation chain
  int max_splits = nt_total < 32 ? nt_total : 32;
  if (max_splits < min_splits + 8) max_splits = min_splits + 8 < nt_total ? min_splits + 8 : nt_total;
  int best = min_splits;
  double best_eff = 0.0;
  for (int s = min_splits; s <= max_splits; ++s) {
    if (nt_total / s < 4 && s > min_splits) break;        // keep >= 4 tiles per CTA
    const long long items = (long long)row_tiles * s;
    const long long waves = (items + sms - 1) / sms;
    const double eff = (double)items / (double)(waves * sms);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
    if (eff >= 0.95) break;
  }
  return best;
}

static size_t ws_for_splits(int splits, int n, int d);
namespace { size_t ws_for_splits_fwd(int splits, int n, int d) { return ws_for_splits(splits, n, d); } }
static size_t ws_for_splits(int splits, int n, int d) {
  if (splits <= 1) return 0;
  const size_t rows = (size_t)splits * n;
  return round_up_sz(rows * d * sizeof(float), 256) + 5 * round_up_sz(rows * sizeof(float), 256);
}

// Stream-K scheduling of the T = 0 forward (HOP_TUNE_STREAMK_ON / _OFF force it): used when the plain grid would give
// every CTA a short pattern range (fixed prologue / epilogue cost per CTA, P = 64k: 64 tiles each) and the share
// of one cluster stays within the accumulation-chain cap.  Never for a launch of the one-launch-per-pass form.
static bool streamk_ok(int n, int p, int d, int update_steps, const TcTune& t) {
  if (update_steps != 0 || t.splits >= 1 || (t.flags & HOP_TUNE_STREAMK_OFF)) return false;
  if (d > 512 || tc_halfm(d, t)) return false;  // (the partial-slot arithmetic assumes 128-row tiles)
  const int cl = tc_cluster(n, d, t);
  const int sms = num_sms();
  const long long nt_total = padded_patterns(p) / kTileP;
  const long long row_groups = ((n + kTileM - 1) / kTileM + cl - 1) / cl;
  const int planes = sms / cl / tc_dv_chunks(d, t);     // clusters per read-out column chunk
  const long long share = row_groups * nt_total / (planes > 0 ? planes : 1);
  if (t.flags & HOP_TUNE_STREAMK_ON) return true;
  // The clusters of a static partition sit at different pattern positions, so together they stream ALL of K^ / V^
  // at once: that only works out of the L2 (126 MB).  Measured at D = 512, P = 64k (128 MB of packs): 0.61 -> 0.55.
  const long long pack_bytes = (long long)padded_patterns(p) * 2 * d * 2;
  if (pack_bytes > (96ll << 20)) return false;
  return share >= 200 && share <= 2 * kMaxChainTilesReadout;   // (measured: below ~200 tiles per cluster the plain grid wins)
}
namespace {
bool streamk_wanted(int n, int p, int d, int update_steps, const TcTune& t) { return streamk_ok(n, p, d, update_steps, t); }
}

// upper bound of the stream-K partial slots: sum of segments <= row groups + 3 x clusters
static size_t streamk_ws(int n, int d, const TcTune& t) {
  const int cl = tc_cluster(n, d, t);
  const size_t row_groups = ((size_t)(n + kTileM - 1) / kTileM + cl - 1) / cl;
  const size_t rows = (row_groups + 3 * (size_t)(num_sms() / cl)) * cl * kTileM;
  return round_up_sz(rows * d * sizeof(float), 256) + 5 * round_up_sz(rows * sizeof(float), 256);
}

// Covers every launch form launch_tc_forward can choose for (n, p, d, T): the plain grid's split partials (all
// launches of one call share one split count) and, for the single T = 0 launch only, the stream-K slots.
size_t tc_forward_workspace(int n, int p, int d, int update_steps, const TcTune& t) {
  const size_t plain = ws_for_splits(tc_choose_splits(n, p, d, update_steps, t), n, d);
  if (!streamk_ok(n, p, d, update_steps, t)) return plain;
  const size_t sk = streamk_ws(n, d, t);
  return sk > plain ? sk : plain;
}

static int launch_tc_one(const TcPassArgs& a, int splits, const TcLaunchOpt& opt, cudaStream_t st) {
  if (tc_halfm(a.h, a.tune)) {
    // D = 512: 256-pattern score tiles, one 512-column read-out; D = 1024: 128-pattern tiles, two read-out chunks
    if (a.h == 512)
      return a.bf16 ? launch_h64<512, 512, 256, true>(a, splits, opt, st) : launch_h64<512, 512, 256, false>(a, splits, opt, st);
    return a.bf16 ? launch_h64<1024, 512, 128, true>(a, splits, opt, st) : launch_h64<1024, 512, 128, false>(a, splits, opt, st);
  }
  const bool pair = tc_use_pair(a.n, a.h, a.tune);
#define HOP_TC_DISPATCH(HH, DVC, BF)                                                        \
  do {                                                                                      \
    if (pair) return launch_d<HH, DVC, BF, 2, true, false>(a, splits, opt, st);             \
    return launch_d<HH, DVC, BF, 1, false, false>(a, splits, opt, st);                      \
  } while (0)
#define HOP_TC_DISPATCH_PRECISE(HH)                                                         \
  do {                                                                                      \
    if (pair) return launch_d<HH, HH, false, 2, true, true>(a, splits, opt, st);            \
    return launch_d<HH, HH, false, 1, false, true>(a, splits, opt, st);                     \
  } while (0)
  if (opt.precise) {
    if (a.bf16) HOP_FAIL(HOP_ERR_UNSUPPORTED, "hi + lo update passes are built for fp16-encoded operands");
    switch (a.h) {
      case 64: HOP_TC_DISPATCH_PRECISE(64);
      case 128: HOP_TC_DISPATCH_PRECISE(128);
      case 256: HOP_TC_DISPATCH_PRECISE(256);
      default: HOP_FAIL(HOP_ERR_UNSUPPORTED, "hi + lo update passes are built for D <= 256");
    }
  }
  if (a.bf16) {
    switch (a.h) {
      case 64: HOP_TC_DISPATCH(64, 64, true);
      case 128: HOP_TC_DISPATCH(128, 128, true);
      case 256: HOP_TC_DISPATCH(256, 256, true);
      case 512: HOP_TC_DISPATCH(512, 256, true);
      default: HOP_TC_DISPATCH(1024, 256, true);
    }
  }
  switch (a.h) {
    case 64: HOP_TC_DISPATCH(64, 64, false);
    case 128: HOP_TC_DISPATCH(128, 128, false);
    case 256: HOP_TC_DISPATCH(256, 256, false);
    case 512: HOP_TC_DISPATCH(512, 256, false);
    default: HOP_TC_DISPATCH(1024, 256, false);
  }
#undef HOP_TC_DISPATCH
#undef HOP_TC_DISPATCH_PRECISE
}

int launch_tc_forward(const TcPassArgs& a, cudaStream_t st) {
  if (!tc_supported(a.h, a.dv)) HOP_FAIL(HOP_ERR_UNSUPPORTED, "tensor-core path needs H == Dv in {64,128,256,512,1024}");
  if (!a.retrieved) HOP_FAIL(HOP_ERR_BAD_ARG, "tensor-core forward: retrieved is NULL");
  const int splits = tc_choose_splits(a.n, a.p, a.h, a.update_steps, a.tune);   // one count for every pass of this call
  if (!tc_multi_launch(a.p, a.h, a.update_steps, a.tune))
    return launch_tc_one(a, splits, TcLaunchOpt{false, true, a.zero_words}, st);
  // one launch per pass: read-out operand = K^T, result beta * A K^ = the next launch's prepared state (fp32).
  // No stream-K here: the workspace is sized for `splits` partials and the chains stay within kMaxChainTilesUpdate.
  if (!a.kt16 || !a.qtmp[0] || !a.qtmp[1])
    HOP_FAIL(HOP_ERR_WORKSPACE, "tensor-core forward with one launch per retrieve update needs kt16 and the two state buffers");
  const bool precise = tc_precise_updates(a.h, a.update_steps, a.tune) && !a.bf16;
  if (precise && (!a.k16lo || !a.kt16lo))
    HOP_FAIL(HOP_ERR_BAD_ARG, "T >= 2 at D <= 256 runs its update passes with hi + lo operand pairs: k16lo / kt16lo are NULL");
  const float* qs = a.qs;
  const float* x_raw = a.x_raw;
  for (int t = 0; t < a.update_steps; ++t) {
    TcPassArgs b = a;
    b.x_raw = x_raw; b.qs = qs; b.update_steps = 0; b.vt16 = a.kt16; b.bias = nullptr; b.out_scale = a.beta;
    b.retrieved = (qs == a.qtmp[0]) ? a.qtmp[1] : a.qtmp[0];
    b.qsave = nullptr; b.lse = nullptr; b.argmax = nullptr; b.rowmax = nullptr;
    HOP_TRY(launch_tc_one(b, splits, TcLaunchOpt{precise, false, t == 0 ? a.zero_words : nullptr}, st));
    qs = b.retrieved; x_raw = nullptr;
  }
  TcPassArgs b = a;
  b.x_raw = nullptr; b.qs = qs; b.update_steps = 0;
  return launch_tc_one(b, splits, TcLaunchOpt{false, false, nullptr}, st);
}

}  // namespace hop
