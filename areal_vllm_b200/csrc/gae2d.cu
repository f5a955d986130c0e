// Padded [B, H] GAE with episode boundaries inside a row ("2-D, overlapped" layout).
//   ref: csrc/cugae/gae.cu:62-131 (gae_kernel_2d_olp + gae_2d_olp<16,16>) and :133-206
//        (gae_2d_nolp, whose host function launches the olp kernel, gae.cu:199).
// No Python caller exists in the reference; built for completeness of the cugae binding.
// The reference gives every (row, segment) ONE thread; here a warp takes the segment: 32
// time steps are loaded coalesced, lane 0 runs the dependent chain in the reference's exact
// operation order (bit-identical), results are written back coalesced.
#include "common.cuh"

namespace b200rl {
namespace {

constexpr int kWarps2d = 8;

__global__ void __launch_bounds__(kWarps2d * 32)
gae_2d_olp_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                  const uint8_t* __restrict__ done, const int32_t* __restrict__ done_y_indices,
                  const int32_t* __restrict__ cu_num_dones, const uint8_t* __restrict__ is_truncate,
                  float* __restrict__ adv_out, float* __restrict__ ret_out, int batch, int horizon,
                  float gamma, float lmbda) {
  const int b = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = cu_num_dones[b];
  const int breadth = cu_num_dones[b + 1] - c0 + 1;
  const float* r = rewards + (int64_t)b * horizon;
  const float* v = values + (int64_t)b * (horizon + 1);
  const uint8_t* dn = done + (int64_t)b * (horizon + 1);
  float* adv = adv_out + (int64_t)b * horizon;
  float* ret = ret_out + (int64_t)b * horizon;
  const float c = gamma * lmbda;
  for (int seq = blockIdx.y * kWarps2d + warp; seq < breadth; seq += gridDim.y * kWarps2d) {
    const int t_start = seq == 0 ? 0 : done_y_indices[c0 + seq - 1];
    const int t_end = seq == breadth - 1 ? horizon : done_y_indices[c0 + seq];
    const bool abandon = seq != breadth - 1 && is_truncate[c0 + seq] != 0;  // gae.cu:77-78
    float last = 0.f;
    for (int hi = t_end; hi > t_start; hi -= 32) {  // chunks of 32 steps, from the end
      const int t = hi - 1 - lane;                   // lane 0 holds the latest step
      float rr = 0.f, cur = 0.f, nex = 0.f;
      if (t >= t_start) {
        rr = r[t];
        cur = v[t];
        nex = (t == t_end - 1 && dn[t + 1]) ? 0.f : v[t + 1];  // :81-83
      }
      const float delta = fmaf(gamma, nex, rr) - cur;          // :85
      float my_adv = 0.f;
      const int n = min(32, hi - t_start);
      for (int k = 0; k < n; k++) {  // the dependent chain, in the reference's order
        const float d = __shfl_sync(0xffffffffu, delta, k);
        last = fmaf(c, last, d);                                // :86
        if (hi - 1 - k == t_end - 1 && abandon) last = 0.f;     // :87
        if (lane == k) my_adv = last;
      }
      if (t >= t_start) {
        adv[t] = my_adv;
        ret[t] = my_adv + cur;                                  // :89
      }
    }
  }
}

}  // namespace
}  // namespace b200rl

extern "C" int b200rl_gae_2d_olp(const float* rewards, const float* values, const uint8_t* done,
                                 const int32_t* done_y_indices, const int32_t* cu_num_dones,
                                 const uint8_t* is_truncate, float* adv, float* ret, int32_t batch,
                                 int32_t horizon, int32_t max_num_dones, float gamma, float lam,
                                 void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(batch >= 0 && horizon >= 0 && max_num_dones >= 0, "negative size");
  if (batch == 0 || horizon == 0) return B200RL_OK;
  B200RL_REQUIRE(rewards && values && done && cu_num_dones && adv && ret, "null pointer");
  B200RL_REQUIRE(max_num_dones == 0 || (done_y_indices && is_truncate), "null pointer");
  B200RL_REQUIRE(batch <= 65535 * 1024, "batch too large");
  int gy = (max_num_dones + 1 + kWarps2d - 1) / kWarps2d;
  if (gy > 64) gy = 64;
  dim3 grid((unsigned)batch, (unsigned)gy);
  gae_2d_olp_kernel<<<grid, kWarps2d * 32, 0, (cudaStream_t)stream>>>(
      rewards, values, done, done_y_indices, cu_num_dones, is_truncate, adv, ret, batch, horizon,
      gamma, lam);
  return check_launch("gae_2d_olp_kernel");
}
