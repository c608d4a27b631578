// Small-N / small-K dense products on the sm_100a tensor pipe: the GEMM-shaped pieces AROUND the fused association --
// the V-projection and the out_proj fold of the pattern-side preparation (upstream HopfieldCore's in-projection,
// V^ = LN_v(W) Wv^T + bv over P rows, P up to 1M) -- which round 1 ran as an FFMA GEMM at
// ~0.12 of the HBM roofline.
//
//     C[M x N] (fp32, row-major)  =  A[M x K] . B[N x K]^T  (+ bias[N])
//
// A and B are K-major fp16 packs streamed by TMA (128-byte swizzle).  With `split` each operand is a hi + lo pair
// (fp16(v), fp16(2^11 (v - fp16(v)))) and a product is three MMAs, Ah Bh into one TMEM accumulator and Al Bh + Ah Bl into
// a second one that the epilogue adds back times 2^-11: ~2^-21 relative, i.e. the fp32 V^ the other kernels read stays an
// fp32-grade quantity although it is produced by tcgen05.mma.kind::f16.  (The residual is stored scaled because the
// unscaled one of a weight of magnitude 0.05 is ~2^-16, an fp16 subnormal with 2^-25 absolute spacing -- measured 1.1e-6
// rel-L2 on V^ against 3e-7 for the fp32 kernels; scaled, the residual has the hi word's own exponent range.)
//
// One CTA = one 128-row tile of C, up to 256 columns (grid.y walks wider N): warp 0 TMA producer, warp 1 MMA issuer
// (cta_group::1, M = 128, N = the tile width in one instruction), warps 2-5 epilogue (TMEM lane = row: 128-byte row
// segments per thread, bias added, plain stores).  K is short here (64..1024), so a tile is a few microseconds: the grid
// is the row tiles, CTAs of neighbouring tiles overlap one tile's epilogue with another's loads where shared memory lets two reside
// (TMEM: TN columns per CTA, 2 TN with operand pairs).
#include <cuda.h>

#include <mutex>

#include "hop_common.cuh"
#include "hop_tc_ptx.cuh"

namespace hop {

namespace {

constexpr int kGM = 128;                  // rows of C per CTA
constexpr int kGThreads = 64 + 128;       // producer, issuer, 4 epilogue warps

struct GemmArgs {
  int m, n, k;                 // valid extents (k a multiple of 64)
  int ldc;
  const float* bias;           // (n) or nullptr
  float* c;
  int chunks_per_split;        // 64-wide K chunks per grid.z slice
  int atomic;                  // add into c with REDs (split-K) instead of storing
  const float* out_scale_dev;  // device scalar multiplied into the result or nullptr
};

// TN = columns of C per CTA (64 | 128 | 256), kSplit = hi + lo operand pairs
template <int TN, bool kSplit>
struct GemmCfg {
  static constexpr int kABytes = kGM * 128;                       // one 64-wide K chunk of A (hi or lo)
  static constexpr int kBBytes = TN * 128;
  static constexpr int kStageBytes = (kSplit ? 2 : 1) * (kABytes + kBBytes);
  static constexpr int kStages = kSplit ? 2 : 3;
  static constexpr int kSmemBytes = kStages * kStageBytes + 256 + 1024;   // + barriers, + alignment slack
  static constexpr uint32_t kTmemCols = (kSplit ? 2 : 1) * TN;     // 64 .. 512: a power of two
};

template <int TN, bool kSplit>
__global__ void __launch_bounds__(kGThreads)
hop_gemm_tc_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_alo,
                   const __grid_constant__ CUtensorMap tm_b, const __grid_constant__ CUtensorMap tm_blo,
                   const GemmArgs g) {
  using Cfg = GemmCfg<TN, kSplit>;
  extern __shared__ uint8_t smem_dyn[];
  const uint32_t smem_base = (ptx::smem_u32(smem_dyn) + 1023u) & ~1023u;      // swizzle atoms need 1 KiB alignment
  const uint32_t ring = smem_base;
  const uint32_t bars = ring + Cfg::kStages * Cfg::kStageBytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (Cfg::kStages + s); };
  const uint32_t acc_full = bars + 16u * Cfg::kStages;
  const uint32_t tmem_slot = acc_full + 8u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = blockIdx.x * kGM, col0 = blockIdx.y * TN;
  const int kc_begin = blockIdx.z * g.chunks_per_split;
  const int kchunks = min(g.k / 64 - kc_begin, g.chunks_per_split);      // (the host leaves no slice empty)

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    ptx::mbar_init(acc_full, 1);
    ptx::fence_barrier_init();
    ptx::prefetch_tmap(&tm_a); ptx::prefetch_tmap(&tm_b);
  }
  if (warp == 0) {
    __syncwarp();
    ptx::tmem_alloc(tmem_slot, Cfg::kTmemCols); ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    if (ptx::elect_one()) {
      for (int kc = 0; kc < kchunks; ++kc) {
        const int s = kc % Cfg::kStages;
        ptx::mbar_wait(empty_bar(s), ((kc / Cfg::kStages) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(full_bar(s), Cfg::kStageBytes);
        const uint32_t base = ring + s * Cfg::kStageBytes;
        const int kx = (kc_begin + kc) * 64;
        ptx::tma_load_2d(base, &tm_a, full_bar(s), kx, row0);
        ptx::tma_load_2d(base + Cfg::kABytes, &tm_b, full_bar(s), kx, col0);
        if constexpr (kSplit) {
          ptx::tma_load_2d(base + Cfg::kABytes + Cfg::kBBytes, &tm_alo, full_bar(s), kx, row0);
          ptx::tma_load_2d(base + 2 * Cfg::kABytes + Cfg::kBBytes, &tm_blo, full_bar(s), kx, col0);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (ptx::elect_one()) {
      constexpr uint32_t idesc = ptx::umma_idesc_f16(kGM, TN, 0);
      for (int kc = 0; kc < kchunks; ++kc) {
        const int s = kc % Cfg::kStages;
        ptx::mbar_wait(full_bar(s), (kc / Cfg::kStages) & 1);
        ptx::tc_fence_after();
        const uint32_t base = ring + s * Cfg::kStageBytes;
        const uint64_t ah = ptx::umma_desc_sw128(base), bh = ptx::umma_desc_sw128(base + Cfg::kABytes);
#pragma unroll
        for (int k = 0; k < 4; ++k) ptx::umma_f16_ss(tmem_base, ah + 2 * k, bh + 2 * k, idesc, (kc | k) != 0);
        if constexpr (kSplit) {
          const uint64_t al = ptx::umma_desc_sw128(base + Cfg::kABytes + Cfg::kBBytes);
          const uint64_t bl = ptx::umma_desc_sw128(base + 2 * Cfg::kABytes + Cfg::kBBytes);
#pragma unroll
          for (int k = 0; k < 4; ++k) ptx::umma_f16_ss(tmem_base + TN, al + 2 * k, bh + 2 * k, idesc, (kc | k) != 0);
#pragma unroll
          for (int k = 0; k < 4; ++k) ptx::umma_f16_ss(tmem_base + TN, ah + 2 * k, bl + 2 * k, idesc, 1u);
        }
        ptx::umma_commit(empty_bar(s));
      }
      ptx::umma_commit(acc_full);
    }
    __syncwarp();
  } else {
    // epilogue: TMEM lane quarter of this warp = (warp % 4); row = lane of the accumulator
    const int wq = warp & 3;
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    const long long grow = (long long)row0 + wq * 32 + lane;
    ptx::mbar_wait(acc_full, 0);
    ptx::tc_fence_after();
    const float osc = g.out_scale_dev ? __ldg(g.out_scale_dev) : 1.f;
#pragma unroll 1
    for (int c = 0; c < TN / 32; ++c) {
      uint32_t o[32];
      ptx::tmem_ld_x32(lane_base + c * 32, o);
      if constexpr (kSplit) {
        uint32_t ol[32];
        ptx::tmem_ld_x32(lane_base + TN + c * 32, ol);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i)
          o[i] = __float_as_uint(fmaf(__uint_as_float(ol[i]), 1.f / kGemmLoScale, __uint_as_float(o[i])));
      } else {
        ptx::tmem_wait_ld();
      }
      const int gc = col0 + c * 32;
      if (grow < g.m && gc < g.n) {
        float* dst = g.c + grow * g.ldc + gc;
#pragma unroll
        for (int q4 = 0; q4 < 8; ++q4) {
          float4 r = make_float4(__uint_as_float(o[4 * q4]) * osc, __uint_as_float(o[4 * q4 + 1]) * osc,
                                 __uint_as_float(o[4 * q4 + 2]) * osc, __uint_as_float(o[4 * q4 + 3]) * osc);
          if (g.atomic) {
            if (gc + 4 * q4 + 3 < g.n) atomicAdd(reinterpret_cast<float4*>(dst) + q4, r);
            else {
              const float rv[4] = {r.x, r.y, r.z, r.w};
              for (int e = 0; e < 4; ++e) if (gc + 4 * q4 + e < g.n) atomicAdd(dst + 4 * q4 + e, rv[e]);
            }
            continue;
          }
          if (g.bias) {
            const float4 bb = __ldg(reinterpret_cast<const float4*>(g.bias + gc) + q4);
            r.x += bb.x; r.y += bb.y; r.z += bb.z; r.w += bb.w;
          }
          if (gc + 4 * q4 + 3 < g.n) reinterpret_cast<float4*>(dst)[q4] = r;
          else {
            const float rv[4] = {r.x, r.y, r.z, r.w};
            for (int e = 0; e < 4; ++e) if (gc + 4 * q4 + e < g.n) dst[4 * q4 + e] = rv[e];
          }
        }
      }
    }
    if (g.atomic) __threadfence();      // REDs performed before the CTA retires (see hop_bwd_tc.cu's epilogue)
    ptx::tc_fence_before();
  }
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn gemm_encode_fn() {
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
// (rows, cols) fp16 row-major with row pitch `ld` elements -> boxes of box_rows x 64 columns, 128-byte swizzle;
// rows beyond `rows` read as zeros (TMA out-of-bounds fill)
int gemm_map(CUtensorMap* m, const void* base, int rows, int cols, int ld, int box_rows) {
  EncodeTiledFn fn = gemm_encode_fn();
  if (!fn) HOP_FAIL(HOP_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) HOP_FAIL(HOP_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for a %d x %d matrix", (int)r, rows, cols);
  return HOP_OK;
}

template <int TN, bool kSplit>
int launch_gemm_inst(const TcGemmArgs& a, cudaStream_t st) {
  using Cfg = GemmCfg<TN, kSplit>;
  CUtensorMap ta, tal, tb, tbl;
  HOP_TRY(gemm_map(&ta, a.a16, a.m_pad, a.k, a.k, kGM));
  HOP_TRY(gemm_map(&tb, a.b16, a.n, a.k, a.k, TN));
  if (kSplit) {
    HOP_TRY(gemm_map(&tal, a.a16lo, a.m_pad, a.k, a.k, kGM));
    HOP_TRY(gemm_map(&tbl, a.b16lo, a.n, a.k, a.k, TN));
  } else { tal = ta; tbl = tb; }
  const int kchunks = a.k / 64;
  const int want = a.k_splits > 1 ? (a.k_splits < kchunks ? a.k_splits : kchunks) : 1;
  const int per = (kchunks + want - 1) / want;
  const int zs = (kchunks + per - 1) / per;                    // no empty slice
  GemmArgs g{a.m, a.n, a.k, a.ldc, a.bias, a.c, per, zs > 1 || a.k_splits > 1 ? 1 : 0, a.out_scale_dev};
  auto* kern = hop_gemm_tc_kernel<TN, kSplit>;
  HOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  kern<<<dim3((a.m + kGM - 1) / kGM, (a.n + TN - 1) / TN, zs), kGThreads, Cfg::kSmemBytes, st>>>(ta, tal, tb, tbl, g);
  HOP_LAUNCH_CHECK("hop_gemm_tc_kernel");
  return HOP_OK;
}

}  // namespace

bool tc_gemm_supported(int n, int k) { return k % 64 == 0 && k >= 64 && n % 64 == 0 && n >= 64; }

int launch_tc_gemm(const TcGemmArgs& a, cudaStream_t st) {
  if (!tc_gemm_supported(a.n, a.k)) HOP_FAIL(HOP_ERR_UNSUPPORTED, "tensor-core GEMM: N and K must be multiples of 64");
  if (!a.a16 || !a.b16 || !a.c || (a.split && (!a.a16lo || !a.b16lo))) HOP_FAIL(HOP_ERR_BAD_ARG, "tensor-core GEMM: NULL operand");
  if ((a.ldc & 3) != 0) HOP_FAIL(HOP_ERR_BAD_ALIGN, "tensor-core GEMM: ldc must be a multiple of 4");
  if (a.k_splits > 1 && a.bias) HOP_FAIL(HOP_ERR_BAD_ARG, "tensor-core GEMM: a split contraction takes no bias");
  const int tn = a.n % 256 == 0 ? 256 : (a.n % 128 == 0 ? 128 : 64);
  if (a.split) {
    if (tn == 256) return launch_gemm_inst<256, true>(a, st);
    if (tn == 128) return launch_gemm_inst<128, true>(a, st);
    return launch_gemm_inst<64, true>(a, st);
  }
  if (tn == 256) return launch_gemm_inst<256, false>(a, st);
  if (tn == 128) return launch_gemm_inst<128, false>(a, st);
  return launch_gemm_inst<64, false>(a, st);
}

}  // namespace hop
