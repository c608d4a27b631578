// Probe (diagnostic, not product): where does tcgen05.mma put the accumulator rows / columns in TMEM for the
// shapes the guides do not document, and how fast does each shape issue?
//
//   mode 0  cta_group::2, M = 256, N = 128   (known: lane i = row i of each CTA, column j = n j) -- control
//   mode 1  cta_group::2, M = 128, N = 128   (each CTA contributes 64 rows of A, 64 rows of B)
//   mode 2  cta_group::2, M = 128, N = 256
//   mode 3  cta_group::1, M = 64,  N = 128
//   mode 4  cta_group::1, M = 128, N = 128   -- control
//   mode 5  cta_group::2, M = 128, N = 128, A from TMEM (TS form): which lanes' A data feed which accumulator lanes?
//           (A is written to TMEM columns [256, 264) of EVERY lane L as A[L, 0] = L + 1 (+ 128 in CTA 1), so the
//           "row" decoded from lane l's accumulator names the lane whose A data it used)
//   mode 6  cta_group::2, M = 256, N = 128, A from TMEM -- control
//
// Operands are chosen so that D[row, n] = 256 (row + 1) + (n + 1) exactly (fp16 inputs, fp32 accumulate):
// A[row, 0] = row + 1, A[row, 1] = 1, B[n, 0] = 256, B[n, 1] = n + 1, everything else 0; TMEM is pre-filled with a
// sentinel.  Every CTA then dumps its 128 lanes x 256 columns; the host prints, per CTA, the (row, n) rectangle
// each (lane range, column range) holds.  A second kernel times back-to-back MMAs (N = 256, K = 16) per shape.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I hop-vae_b200/csrc \
//        scripts/probes/umma_layout_probe.cu -o gpurun_out/umma_probe && gpurun_out/umma_probe
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "hop_tc_ptx.cuh"

using namespace hop;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t sw128_off(int r, int c16) { return (uint32_t)(r * 128 + ((c16 ^ (r & 7)) << 4)); }

struct Mode { int pair, m, n, ts; };
__constant__ Mode kModes[7] = {{1, 256, 128, 0}, {1, 128, 128, 0}, {1, 128, 256, 0}, {0, 64, 128, 0}, {0, 128, 128, 0},
                               {1, 128, 128, 1}, {1, 256, 128, 1}};
constexpr uint32_t kSentinel = 0x7fc0dead;   // a NaN pattern

__device__ __forceinline__ void put16(uint8_t* tile, int r, int k, float v) {
  *reinterpret_cast<__half*>(tile + sw128_off(r, k >> 3) + (k & 7) * 2) = __float2half_rn(v);
}

// one cluster of two CTAs, 128 threads each
__global__ void __launch_bounds__(128, 1) probe_layout(int mode, uint32_t* out /* [2][128][256] */) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* a_tile = smem;                 // up to 128 rows x 64 halves (16 KiB)
  uint8_t* b_tile = smem + 16384;         // up to 256 rows x 64 halves (32 KiB)
  const uint32_t bar = ptx::smem_u32(smem + 49152), slot = bar + 16;
  const Mode md = kModes[mode];
  const uint32_t crank = ptx::cluster_ctarank();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int a_rows = md.pair ? md.m / 2 : md.m;         // rows of A this CTA holds
  const int b_rows = md.pair ? md.n / 2 : md.n;
  for (int i = threadIdx.x; i < 49152 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  __syncthreads();
  for (int r = threadIdx.x; r < a_rows; r += blockDim.x) {
    const int grow = md.pair ? (int)crank * a_rows + r : r;
    put16(a_tile, r, 0, (float)(grow + 1));
    put16(a_tile, r, 1, 1.f);
  }
  for (int r = threadIdx.x; r < b_rows; r += blockDim.x) {
    const int gn = md.pair ? (int)crank * b_rows + r : r;
    put16(b_tile, r, 0, 256.f);
    put16(b_tile, r, 1, (float)(gn + 1));
  }
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  ptx::fence_proxy_async_smem();
  if (warp == 0) {
    __syncwarp();
    if (md.pair) { ptx::tmem_alloc_pair(slot, 512); ptx::tmem_relinquish_pair(); }
    else { ptx::tmem_alloc(slot, 512); ptx::tmem_relinquish(); }
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(slot));
  const uint32_t lane_base = tmem_base + ((uint32_t)(warp * 32) << 16);
  {
    uint32_t s[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) s[i] = kSentinel;
    for (int c = 0; c < 256; c += 32) ptx::tmem_st_x32(lane_base + c, s);
    if (md.ts) {   // A operand in TMEM: 16-bit pairs, K = 16 -> 8 columns; the rest of the 32-column store is zero
      const int L = warp * 32 + lane + (int)crank * 128;
      __half2 h2 = __floats2half2_rn((float)(L + 1), 1.f);
      for (int i = 0; i < 32; ++i) s[i] = 0;
      s[0] = *reinterpret_cast<uint32_t*>(&h2);
      ptx::tmem_st_x32(lane_base + 256, s);
    }
    ptx::tmem_wait_st();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  if (warp == 1 && ptx::elect_one()) {
    const uint32_t idesc = ptx::umma_idesc_f16(md.m, md.n, 0);
    const uint64_t ad = ptx::umma_desc_sw128(ptx::smem_u32(a_tile)), bd = ptx::umma_desc_sw128(ptx::smem_u32(b_tile));
    if (md.pair) {
      if (crank == 0) {
        if (md.ts) ptx::umma2_f16_ts(tmem_base, tmem_base + 256, bd, idesc, 0);
        else ptx::umma2_f16_ss(tmem_base, ad, bd, idesc, 0);
        ptx::umma2_commit_both(bar);
      }
    } else {
      ptx::umma_f16_ss(tmem_base, ad, bd, idesc, 0); ptx::umma_commit(bar);
    }
  }
  __syncwarp();
  ptx::mbar_wait_cluster(bar, 0);
  ptx::tc_fence_after();
  for (int c = 0; c < 256; c += 32) {
    uint32_t v[32];
    ptx::tmem_ld_x32(lane_base + c, v);
    ptx::tmem_wait_ld();
    for (int i = 0; i < 32; ++i) out[((size_t)crank * 128 + warp * 32 + lane) * 256 + c + i] = v[i];
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 0) {
    ptx::tc_fence_after();
    if (md.pair) ptx::tmem_dealloc_pair(tmem_base, 512); else ptx::tmem_dealloc(tmem_base, 512);
  }
}

// issue rate: `reps` x 4 MMAs (N = 256, K = 16 each) on resident operands, clock64 around them
__global__ void __launch_bounds__(128, 1) probe_rate(int pair, int m, int n, int reps, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t bar = ptx::smem_u32(smem + 49152), slot = bar + 16;
  const uint32_t crank = ptx::cluster_ctarank();
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 49152 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  ptx::fence_proxy_async_smem();
  if (warp == 0) {
    __syncwarp();
    if (pair) { ptx::tmem_alloc_pair(slot, 512); ptx::tmem_relinquish_pair(); }
    else { ptx::tmem_alloc(slot, 512); ptx::tmem_relinquish(); }
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(slot));
  if (warp == 1 && ptx::elect_one() && (!pair || crank == 0)) {
    const uint32_t idesc = ptx::umma_idesc_f16(m, n, 0);
    const uint64_t ad = ptx::umma_desc_sw128(ptx::smem_u32(smem)), bd = ptx::umma_desc_sw128(ptx::smem_u32(smem + 16384));
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (pair) ptx::umma2_f16_ss(tmem_base + (r & 1) * 256, ad + 2 * k, bd + 2 * k, idesc, 1);
        else ptx::umma_f16_ss(tmem_base + (r & 1) * 256, ad + 2 * k, bd + 2 * k, idesc, 1);
      }
    }
    if (pair) ptx::umma2_commit_both(bar); else ptx::umma_commit(bar);
    ptx::mbar_wait_cluster(bar, 0);
    const long long t1 = clock64();
    cycles[blockIdx.x] = t1 - t0;
  }
  __syncwarp();
  if (!(warp == 1) ) { /* others wait for the barrier too so that nobody leaves early */ }
  ptx::mbar_wait_cluster(bar, 0);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 0) {
    ptx::tc_fence_after();
    if (pair) ptx::tmem_dealloc_pair(tmem_base, 512); else ptx::tmem_dealloc(tmem_base, 512);
  }
}

static int launch(void* fn, int grid, int smem, void** args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = 0;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return (int)cudaLaunchKernelExC(&cfg, fn, args);
}

int main() {
  const int smem = 49152 + 64;
  CK(cudaFuncSetAttribute(probe_layout, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(probe_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  uint32_t* d_out; long long* d_cyc;
  CK(cudaMalloc(&d_out, 2 * 128 * 256 * 4));
  CK(cudaMalloc(&d_cyc, 2 * sizeof(long long)));
  std::vector<uint32_t> h(2 * 128 * 256);
  const char* names[7] = {"pair M=256 N=128", "pair M=128 N=128", "pair M=128 N=256", "single M=64 N=128", "single M=128 N=128",
                          "pair M=128 N=128 A-from-TMEM (row = source lane, +128 in cta1)", "pair M=256 N=128 A-from-TMEM"};
  for (int mode = 0; mode < 7; ++mode) {
    CK(cudaMemset(d_out, 0, 2 * 128 * 256 * 4));
    void* args[] = {&mode, &d_out};
    int rc = launch((void*)probe_layout, 2, smem, args);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc != 0 || e != cudaSuccess) { printf("mode %d (%s): launch rc=%d sync=%s\n", mode, names[mode], rc, cudaGetErrorString(e)); return 1; }
    CK(cudaMemcpy(h.data(), d_out, h.size() * 4, cudaMemcpyDeviceToHost));
    printf("== mode %d: %s\n", mode, names[mode]);
    for (int cta = 0; cta < 2; ++cta) {
      // run-length summary: consecutive lanes whose (row - lane) offset and column->n map agree are merged
      int start = 0;
      auto sig = [&](int ln, char* buf) {
        // describe a lane: its row (or -1), and the column segments -> n
        const uint32_t* v = &h[((size_t)cta * 128 + ln) * 256];
        int pos = 0, c = 0, row_seen = -2;
        while (c < 256) {
          if (v[c] == kSentinel) { int c0 = c; while (c < 256 && v[c] == kSentinel) ++c; pos += sprintf(buf + pos, "[c%d-%d:untouched]", c0, c - 1); continue; }
          const float f = *reinterpret_cast<const float*>(&v[c]);
          const int iv = (int)f, row = iv / 256 - 1, n = iv % 256 - 1;
          int c0 = c, n0 = n;
          while (c < 256 && v[c] != kSentinel) {
            const int jv = (int)*reinterpret_cast<const float*>(&v[c]);
            if (jv / 256 - 1 != row || jv % 256 - 1 != n0 + (c - c0)) break;
            ++c;
          }
          if (c == c0) { pos += sprintf(buf + pos, "[c%d:?%08x]", c, v[c]); ++c; continue; }
          if (row_seen == -2) row_seen = row;
          pos += sprintf(buf + pos, "[c%d-%d:row%+d n%d-%d]", c0, c - 1, row - ln, n0, n0 + (c - c0) - 1);
          if (pos > 900) break;
        }
        (void)row_seen;
      };
      char cur[1024], nxt[1024];
      sig(0, cur);
      for (int ln = 1; ln <= 128; ++ln) {
        if (ln < 128) sig(ln, nxt);
        if (ln == 128 || strcmp(cur, nxt) != 0) {
          printf("  cta%d lanes %3d-%3d: %s   (rowΔ = accumulator row minus lane; rows are GLOBAL over the pair)\n", cta, start, ln - 1, cur);
          start = ln;
          if (ln < 128) memcpy(cur, nxt, sizeof(cur));
        }
      }
    }
  }
  // rates
  struct R { int pair, m; const char* name; } rs[4] = {{1, 256, "pair M=256"}, {1, 128, "pair M=128"}, {0, 128, "single M=128"}, {0, 64, "single M=64"}};
  for (int n : {256, 128})
  for (auto& r : rs) {
    int reps = 4000;
    void* args[] = {&r.pair, &r.m, &n, &reps, &d_cyc};
    CK(cudaMemset(d_cyc, 0, 2 * sizeof(long long)));
    int rc = launch((void*)probe_rate, 2, smem, args);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc != 0 || e != cudaSuccess) { printf("rate %s: launch rc=%d sync=%s\n", r.name, rc, cudaGetErrorString(e)); return 1; }
    long long c[2];
    CK(cudaMemcpy(c, d_cyc, sizeof(c), cudaMemcpyDeviceToHost));
    const double per_mma = (double)c[0] / (reps * 4.0);
    const double flops_per_sm = 2.0 * r.m * n * 16 / (r.pair ? 2.0 : 1.0);   // per MMA, per SM involved
    printf("rate %-13s N=%d K=16: %.1f cycles / MMA, %.0f FLOP / cycle / SM\n", r.name, n, per_mma, flops_per_sm / per_mma);
  }
  return 0;
}
