// Pattern-sharded mode: log-sum-exp merge of G shard-local retrievals
// (new design -- the reference is single-device).
//
// Each shard g holds for every query row: the normalised read-out O_g (N,Dv),
// lse_g = log sum_j exp(s_j) over ITS patterns, the row max and its index.
// The global read-out is sum_g w_g O_g with w_g = exp(lse_g - lse),
// lse = log sum_g exp(lse_g).  This kernel rescales the local O_g by w_g in
// place (so one SUM all-reduce over NVLink finishes the job) and resolves the
// global lse / nearest pattern.  HBM-bound: reads and writes N*Dv*4 bytes.
#include "hop_common.cuh"

namespace hop {

namespace {

__global__ void __launch_bounds__(256) merge_kernel(int n, int dv, int g, int shard,
                                                    const float* __restrict__ lse_all,
                                                    const float* __restrict__ rowmax_all,
                                                    const int32_t* __restrict__ argmax_all,
                                                    float* __restrict__ retrieved, float* __restrict__ lse_out,
                                                    int32_t* __restrict__ argmax_out) {
  const int dv4 = dv >> 2;  // host guarantees dv % 4 == 0
  const long long total = (long long)n * dv4;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int row = (int)(idx / dv4);
    const int c4 = (int)(idx - (long long)row * dv4);
    float m = -INFINITY;
    for (int s = 0; s < g; ++s) m = fmaxf(m, lse_all[(size_t)s * n + row]);
    float sum = 0.f;
    for (int s = 0; s < g; ++s) sum += __expf(lse_all[(size_t)s * n + row] - m);
    const float lse = m + __logf(sum);
    const float w = __expf(lse_all[(size_t)shard * n + row] - lse);
    float4* p = reinterpret_cast<float4*>(retrieved + (size_t)row * dv) + c4;
    float4 v = *p;
    v.x *= w; v.y *= w; v.z *= w; v.w *= w;
    *p = v;
    if (c4 == 0) {
      if (lse_out) lse_out[row] = lse;
      if (argmax_out && rowmax_all && argmax_all) {
        float best = -INFINITY;
        int32_t arg = 0x7fffffff;
        for (int s = 0; s < g; ++s) {
          const float rm = rowmax_all[(size_t)s * n + row];
          const int32_t am = argmax_all[(size_t)s * n + row];
          if (rm > best || (rm == best && am < arg)) {
            best = rm;
            arg = am;
          }
        }
        argmax_out[row] = arg;
      }
    }
  }
}

// ------------------------------------------------------------------ peer-memory merge (hop_sm100_peer_merge)
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Bounded flag wait: a peer that never arrives (dead rank, mismatched epoch) becomes a trap ("unspecified launch
// failure" on this rank) after kPeerWaitNs of wall time instead of a hung GPU.  The bound is wall time, not a spin
// count: ranks legitimately start seconds apart (host-side skew of a first call).
constexpr unsigned long long kPeerWaitNs = 30ull * 1000ull * 1000ull * 1000ull;
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
template <bool kSys>
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t epoch) {
  unsigned long long t0 = 0;
  uint32_t spins = 0;
  while ((int32_t)((kSys ? ld_acquire_sys(p) : ld_acquire_gpu(p)) - epoch) < 0) {
    __nanosleep(32);
    if ((++spins & 0x3fffu) == 0) {
      const unsigned long long now = global_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > kPeerWaitNs) __trap();
    }
  }
}
// flag block of one rank: [0,16) ready[src], [16,32) done[src], [32] block counter, [33] local "go" flag
constexpr int kFlagDone = HOP_MAX_PEERS, kFlagCounter = 2 * HOP_MAX_PEERS, kFlagGo = 2 * HOP_MAX_PEERS + 1;

// One work item = (row, 32-lane column chunk): a lane pulls one float4 from each of the G ranks.  R items are in
// flight per warp so that enough bytes are outstanding to cover the NVLink round trip.
template <int GMAX, int R>
__global__ void __launch_bounds__(256) peer_merge_kernel(const hop_peer_merge m) {
  const int G = m.n_ranks, me = m.rank;
  // ---- opening barrier.  Block 0 talks to the peers (system scope): "my partials are complete" (stream order),
  // then waits for the same from everyone; the other blocks (all co-resident: a cooperative launch whose grid is
  // capped by the occupancy query, see launch_peer_merge) wait on a local device-scope flag -- hundreds of blocks polling system-scope flags cost more than the merge itself.
  if (blockIdx.x == 0) {
    if (threadIdx.x < G) {
      __threadfence_system();
      st_release_sys(m.flags[threadIdx.x] + me, m.epoch);
      wait_flag<true>(m.flags[me] + threadIdx.x, m.epoch);
    }
    __syncthreads();
    if (threadIdx.x == 0) st_release_gpu(m.flags[me] + kFlagGo, m.epoch);
  } else {
    if (threadIdx.x == 0) wait_flag<false>(m.flags[me] + kFlagGo, m.epoch);
    __syncthreads();
  }

  const int dv4 = m.d_v >> 2;
  const int chunks = (dv4 + 31) >> 5;
  const long long items = (long long)m.row_count * chunks;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  for (long long base = ((long long)blockIdx.x * wpb + warp) * R; base < items; base += (long long)gridDim.x * wpb * R) {
    float4 v[R][GMAX];
    float l[R][GMAX], rm[R][GMAX];
    int32_t am[R][GMAX];
    size_t row[R];
    int c4[R];
    bool valid[R], first[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const long long it = base + r;
      valid[r] = it < items;
      const long long ri = valid[r] ? it / chunks : 0;
      row[r] = (size_t)m.row_begin + (size_t)ri;
      const int ch = valid[r] ? (int)(it - ri * chunks) : 0;
      first[r] = ch == 0;
      c4[r] = ch * 32 + lane;
#pragma unroll
      for (int g = 0; g < GMAX; ++g) {
        l[r][g] = -INFINITY; rm[r][g] = -INFINITY; am[r][g] = 0x7fffffff;
        v[r][g] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g < G && valid[r]) {
          l[r][g] = __ldcg(m.lse[g] + row[r]);
          if (first[r]) { rm[r][g] = __ldcg(m.rowmax[g] + row[r]); am[r][g] = __ldcg(m.argmax[g] + row[r]); }
          if (c4[r] < dv4) v[r][g] = __ldcg(reinterpret_cast<const float4*>(m.o[g] + row[r] * m.d_v) + c4[r]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (!valid[r]) continue;
      float mx = -INFINITY;
#pragma unroll
      for (int g = 0; g < GMAX; ++g) mx = fmaxf(mx, l[r][g]);
      float sum = 0.f;
#pragma unroll
      for (int g = 0; g < GMAX; ++g) { l[r][g] = (g < G) ? __expf(l[r][g] - mx) : 0.f; sum += l[r][g]; }
      const float inv = 1.f / sum;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int g = 0; g < GMAX; ++g) {
        const float w = l[r][g] * inv;
        acc.x = fmaf(w, v[r][g].x, acc.x); acc.y = fmaf(w, v[r][g].y, acc.y);
        acc.z = fmaf(w, v[r][g].z, acc.z); acc.w = fmaf(w, v[r][g].w, acc.w);
      }
      if (c4[r] < dv4) {
        if (m.push_all) {
#pragma unroll
          for (int g = 0; g < GMAX; ++g)
            if (g < G) reinterpret_cast<float4*>(m.out[g] + row[r] * m.d_v)[c4[r]] = acc;
        } else {
          reinterpret_cast<float4*>(m.out[me] + row[r] * m.d_v)[c4[r]] = acc;
        }
      }
      if (first[r] && lane == 0) {
        float best = -INFINITY;
        int32_t arg = 0x7fffffff;
#pragma unroll
        for (int g = 0; g < GMAX; ++g)
          if (rm[r][g] > best || (rm[r][g] == best && am[r][g] < arg)) { best = rm[r][g]; arg = am[r][g]; }
        const float lse = mx + __logf(sum);
        if (m.push_all) {
          for (int g = 0; g < G; ++g) {
            if (m.lse_out[g]) m.lse_out[g][row[r]] = lse;
            if (m.argmax_out[g]) m.argmax_out[g][row[r]] = arg;
          }
        } else {
          if (m.lse_out[me]) m.lse_out[me][row[r]] = lse;
          if (m.argmax_out[me]) m.argmax_out[me][row[r]] = arg;
        }
      }
    }
  }

  // ---- closing barrier: the last block of this rank tells every peer "my reads of your partials and my stores
  // into your buffers are done", then waits for the same from all of them before the kernel may complete.
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) {
    __threadfence_system();     // cumulative over the CTA's stores ordered before it by the barrier
    uint32_t* counter = m.flags[me] + kFlagCounter;
    s_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1 : 0;
    if (s_last) *counter = 0u;
  }
  __syncthreads();
  if (s_last && threadIdx.x < G) {
    __threadfence_system();
    st_release_sys(m.flags[threadIdx.x] + kFlagDone + me, m.epoch);
    wait_flag<true>(m.flags[me] + kFlagDone + threadIdx.x, m.epoch);
  }
}

}  // namespace

int launch_merge(int n, int dv, int g, int shard, const float* lse_all, const float* rowmax_all,
                 const int32_t* argmax_all, float* retrieved, float* lse_out, int32_t* argmax_out,
                 cudaStream_t st) {
  if (dv % 4 != 0) HOP_FAIL(HOP_ERR_BAD_SHAPE, "merge: d_v must be a multiple of 4 (got %d)", dv);
  const long long total = (long long)n * (dv / 4);
  const int sms = num_sms();
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)sms * 8) blocks = (long long)sms * 8;  // grid-stride, 8 CTAs per SM
  if (blocks < 1) blocks = 1;
  merge_kernel<<<(int)blocks, 256, 0, st>>>(n, dv, g, shard, lse_all, rowmax_all, argmax_all, retrieved,
                                            lse_out, argmax_out);
  HOP_LAUNCH_CHECK("merge_kernel");
  return HOP_OK;
}

template <int GMAX, int R>
static int launch_peer_merge_inst(const hop_peer_merge& m, long long want, cudaStream_t st) {
  auto* kern = peer_merge_kernel<GMAX, R>;
  // every block must be resident at once (blocks != 0 wait for block 0's flag): cap the grid by the occupancy of
  // THIS kernel on THIS device and launch cooperatively, so that the driver refuses a grid it cannot co-schedule
  // instead of letting it deadlock
  static int per_sm_cache[64];
  int dev = 0;
  HOP_CUDA(cudaGetDevice(&dev));
  int per_sm = (dev >= 0 && dev < 64) ? per_sm_cache[dev] : 0;
  if (per_sm == 0) {
    HOP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, 0));
    if (per_sm < 1) HOP_FAIL(HOP_ERR_CUDA, "peer_merge_kernel cannot be resident on this device");
    if (dev >= 0 && dev < 64) per_sm_cache[dev] = per_sm;
  }
  const long long cap = (long long)(per_sm > 2 ? 2 : per_sm) * num_sms();
  const int blocks = (int)(want < 1 ? 1 : (want > cap ? cap : want));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(blocks);
  cfg.blockDim = dim3(256);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  HOP_CUDA(cudaLaunchKernelEx(&cfg, kern, m));
  HOP_LAUNCH_CHECK("peer_merge_kernel");
  return HOP_OK;
}

int launch_peer_merge(const hop_peer_merge& m, cudaStream_t st) {
  const int chunks = ((m.d_v >> 2) + 31) >> 5;
  const long long items = (long long)m.row_count * chunks;
  const int r = m.n_ranks <= 2 ? 4 : (m.n_ranks <= 4 ? 2 : 1);
  const long long want = (items + 8 * r - 1) / (8 * r);      // 8 warps per block, r items per warp per trip
  if (m.n_ranks <= 2) return launch_peer_merge_inst<2, 4>(m, want, st);
  if (m.n_ranks <= 4) return launch_peer_merge_inst<4, 2>(m, want, st);
  if (m.n_ranks <= 8) return launch_peer_merge_inst<8, 1>(m, want, st);
  return launch_peer_merge_inst<16, 1>(m, want, st);
}

}  // namespace hop
