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

}  // namespace hop
