// Small [T]-sized companions of the GAE path in the legacy PPO interface:
//   * clipped value (critic) loss, forward + gradient in one launch
//       ref: realhf/impl/model/utils/ppo_functional.py:150-225 (_huber_loss, _mse_loss, critic_loss_fn)
//   * KL-shaped packed rewards
//       ref: realhf/impl/model/utils/ppo_functional.py:228-253 (get_packed_rewards)
// Latency-bound: the win is one or two launches and no host sync instead of ~15 eager kernels.
#include "common.cuh"

namespace b200rl {
namespace {

constexpr int kThreads = 256;
constexpr int kPerThread = 4;
constexpr int kChunk = kThreads * kPerThread;
constexpr int kMaxBlocks = 148 * 4;

__device__ __forceinline__ float tmaxf(float a, float b) {
  return (a != a || b != b) ? __int_as_float(0x7fc00000) : fmaxf(a, b);
}

// loss and its derivative w.r.t. x; type 0 = mse 0.5*(x-y)^2, type 1 = huber(delta)
__device__ __forceinline__ void value_loss(float x, float y, int type, float delta, float& l, float& dl) {
  const float e = x - y;
  if (type == 0) {
    l = 0.5f * (e * e);  // 0.5 * (x - y) ** 2
    dl = e;
  } else {
    const float a = fabsf(e);
    if (a < delta) {
      l = 0.5f * (a * a);
      dl = e;
    } else {
      l = delta * (a - 0.5f * delta);
      dl = e > 0.f ? delta : (e < 0.f ? -delta : 0.f);
    }
  }
}

struct CriticWs {
  unsigned int* counter;
  double* partial;
  int blocks;
  size_t bytes;
};
CriticWs critic_ws(void* ws, int64_t T) {
  CriticWs W;
  int64_t b = (T + kChunk - 1) / kChunk;
  if (b < 1) b = 1;
  if (b > kMaxBlocks) b = kMaxBlocks;
  W.blocks = (int)b;
  W.counter = (unsigned int*)ws;
  W.partial = (double*)((char*)ws + 256);
  W.bytes = ((256 + (size_t)W.blocks * 2 * sizeof(double)) + 255) & ~(size_t)255;
  return W;
}

__global__ void __launch_bounds__(kThreads)
critic_loss_fwd_kernel(const float* __restrict__ value, const float* __restrict__ old_value,
                       const float* __restrict__ target, const uint8_t* __restrict__ mask, int64_t T,
                       float eps, int type, float delta, float* __restrict__ s_loss,
                       uint8_t* __restrict__ s_clip, float* __restrict__ dvalue,
                       float* __restrict__ loss, float* __restrict__ denom,
                       unsigned int* __restrict__ counter, double* __restrict__ partial) {
  __shared__ double red[2][kThreads / 32];
  __shared__ bool is_last;
  const int tid = threadIdx.x;
  float tsum = 0.f;
  int tcnt = 0;
  for (int64_t base = (int64_t)blockIdx.x * kChunk; base < T; base += (int64_t)gridDim.x * kChunk) {
#pragma unroll
    for (int k = 0; k < kPerThread; k++) {
      const int64_t t = base + k * kThreads + tid;
      if (t >= T) break;
      const bool m = mask ? mask[t] != 0 : true;
      const float v = value[t], ov = old_value[t], y = target[t];
      float lo, go, lc, gc;
      value_loss(v, y, type, delta, lo, go);                        // :205
      const float dv = v - ov;
      const float dcl = fminf(fmaxf(dv, -eps), eps);                // :207-209 clamp(-eps, eps)
      const float vc = ov + (dv != dv ? dv : dcl);
      value_loss(vc, y, type, delta, lc, gc);                       // :211
      if (!(dv >= -eps && dv <= eps)) gc = 0.f;                     // clamp blocks the gradient outside
      const float vl = tmaxf(lo, lc);                               // :213
      const float g = (lo > lc) ? go : ((lo < lc) ? gc : 0.5f * (go + gc));
      const bool clip = (lc > lo) && m;                             // :216-218
      if (s_loss) s_loss[t] = vl;
      if (s_clip) s_clip[t] = clip;
      if (dvalue) dvalue[t] = m ? g : 0.f;
      if (m) {
        tsum += vl;
        tcnt += 1;
      }
    }
  }
  double s = warp_sum((double)tsum);
  double c = warp_sum((double)tcnt);
  if ((tid & 31) == 0) {
    red[0][tid >> 5] = s;
    red[1][tid >> 5] = c;
  }
  __syncthreads();
  if (tid == 0) {
    double bs = 0, bc = 0;
    for (int w = 0; w < kThreads / 32; w++) {
      bs += red[0][w];
      bc += red[1][w];
    }
    partial[2 * blockIdx.x] = bs;
    partial[2 * blockIdx.x + 1] = bc;
    __threadfence();
    is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last && tid < 32) {
    __threadfence();
    const volatile double* part = partial;
    double fs = 0, fc = 0;
    for (int b = tid; b < (int)gridDim.x; b += 32) {
      fs += part[2 * b];
      fc += part[2 * b + 1];
    }
    fs = warp_sum(fs);
    fc = warp_sum(fc);
    if (tid == 0) {
      // masked: sum / count_nonzero (no `or 1` in the reference, :222); unmasked: mean (:225)
      const float N = (float)fc;
      *denom = N;
      *loss = (float)fs / N;
    }
  }
}

__global__ void __launch_bounds__(256)
packed_rewards_kernel(const float* __restrict__ lp, const float* __restrict__ ref, int64_t T,
                      float kl_ctl, float* __restrict__ kl_rewards, float* __restrict__ tot_rewards) {
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < T; t += (int64_t)gridDim.x * 256) {
    const float r = -kl_ctl * (lp[t] - ref[t]);  // :241
    kl_rewards[t] = r;
    tot_rewards[t] = r;
  }
}

__global__ void __launch_bounds__(256)
packed_rewards_score_kernel(const float* __restrict__ score, const int32_t* __restrict__ cu,
                            const uint8_t* __restrict__ no_eos, int batch, int64_t T, float clip,
                            int mask_no_eos_with_zero, float* __restrict__ tot_rewards) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= batch) return;
  float s = fminf(fmaxf(score[i], -clip), clip);  // :243
  if (mask_no_eos_with_zero && no_eos[i]) s = 0.f;  // :245-248
  int64_t pos = (int64_t)cu[i + 1] - 1;
  if (pos < 0) pos += T;  // python negative index (an empty first sequence)
  if (pos >= 0 && pos < T) tot_rewards[pos] += s;  // one writer per position for non-empty sequences
}

}  // namespace
}  // namespace b200rl

extern "C" {

size_t b200rl_critic_loss_workspace_bytes(int64_t T) {
  if (T < 0) return 0;
  return b200rl::critic_ws(nullptr, T).bytes;
}

int b200rl_critic_loss_fwd(const float* value, const float* old_value, const float* target_value,
                           const uint8_t* mask, int64_t T, float value_eps_clip, int loss_type,
                           float huber_delta, float* stat_loss, uint8_t* stat_clip_mask,
                           float* dvalue, float* loss, float* denom, void* workspace,
                           size_t workspace_bytes, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(T >= 0, "negative T");
  B200RL_REQUIRE(loss && denom && workspace, "loss/denom/workspace must not be null");
  B200RL_REQUIRE(T == 0 || (value && old_value && target_value), "null input pointer");
  B200RL_REQUIRE(loss_type == 0 || loss_type == 1, "loss_type must be 0 (mse) or 1 (huber)");
  B200RL_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "workspace must be 256-B aligned");
  CriticWs W = critic_ws(workspace, T);
  if (workspace_bytes < W.bytes) {
    set_last_error("critic loss workspace too small: %zu < %zu", workspace_bytes, W.bytes);
    return B200RL_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200RL_CUDA(cudaMemsetAsync(workspace, 0, 256, st));
  critic_loss_fwd_kernel<<<W.blocks, kThreads, 0, st>>>(value, old_value, target_value, mask, T,
                                                       value_eps_clip, loss_type, huber_delta,
                                                       stat_loss, stat_clip_mask, dvalue, loss,
                                                       denom, W.counter, W.partial);
  return check_launch("critic_loss_fwd_kernel");
}

int b200rl_packed_rewards(const float* log_probs, const float* ref_log_probs,
                          const float* reward_score, const int32_t* short1cu_seqlens,
                          const uint8_t* seq_no_eos_mask, int32_t batch, int64_t T, float kl_ctl,
                          float clip_reward_value, int mask_no_eos_with_zero, float* kl_rewards,
                          float* tot_rewards, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(T >= 0 && batch >= 0, "negative size");
  if (T == 0) return B200RL_OK;
  B200RL_REQUIRE(log_probs && ref_log_probs && kl_rewards && tot_rewards, "null pointer");
  B200RL_REQUIRE(batch == 0 || (reward_score && short1cu_seqlens && seq_no_eos_mask), "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  int64_t blocks = (T + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  packed_rewards_kernel<<<(unsigned)blocks, 256, 0, st>>>(log_probs, ref_log_probs, T, kl_ctl,
                                                         kl_rewards, tot_rewards);
  if (batch > 0)
    packed_rewards_score_kernel<<<(batch + 255) / 256, 256, 0, st>>>(
        reward_score, short1cu_seqlens, seq_no_eos_mask, batch, T, clip_reward_value,
        mask_no_eos_with_zero, tot_rewards);
  return check_launch("packed_rewards kernels");
}

}  // extern "C"
