// H2 — clipped PPO / GRPO actor loss (decoupled PPO), forward + analytic gradient in ONE launch.
//
// Replaces ~25 eager aten kernels + a host sync (`loss_mask.count_nonzero() or 1`) in
// areal/utils/functional.py:124-188. Everything is [T]-sized (~48 B/token), so this kernel is
// latency/launch bound, not bandwidth bound: the win is one launch, no sync, and the gradient
// d(N*loss)/d logp emitted in the same pass (torch autograd tie semantics of max/min/clamp are
// reproduced, see the `gp` computation). Optional fused advantage normalisation (per-token group
// mean/denominator lookup in fp64, the arithmetic of masked_normalization).
//
// Reduction: per-CTA partial sums (fp64) -> workspace -> the last CTA to finish adds them in
// index order. Fixed order => deterministic loss for a given T.
#include "common.cuh"
#include "ppo_math.cuh"

namespace b200rl {
namespace {

constexpr int kThreads = 256;
constexpr int kPerThread = 4;
constexpr int kChunk = kThreads * kPerThread;
constexpr int kMaxBlocks = 148 * 4;

struct LossWs {
  unsigned int* counter;  // CTAs finished
  double* partial;        // [blocks][2] : sum, count
  int blocks;
  size_t bytes;
};
LossWs loss_ws(void* ws, int64_t T) {
  LossWs W;
  int64_t b = (T + kChunk - 1) / kChunk;
  if (b < 1) b = 1;
  if (b > kMaxBlocks) b = kMaxBlocks;
  W.blocks = (int)b;
  W.counter = (unsigned int*)ws;
  W.partial = (double*)((char*)ws + 256);
  W.bytes = 256 + (size_t)W.blocks * 2 * sizeof(double);
  W.bytes = (W.bytes + 255) & ~(size_t)255;
  return W;
}

struct Params : PpoLossArgs {
  unsigned int* counter;
  double* partial;
};

__global__ void __launch_bounds__(kThreads) ppo_loss_fwd_kernel(const Params p) {
  __shared__ double red[2][kThreads / 32];
  __shared__ bool is_last;
  const int tid = threadIdx.x;
  const bool norm = p.adv_mean != nullptr || p.adv_denom != nullptr;
  float tsum = 0.f;
  int tcnt = 0;
  for (int64_t base = (int64_t)blockIdx.x * kChunk; base < p.T; base += (int64_t)gridDim.x * kChunk) {
#pragma unroll
    for (int k = 0; k < kPerThread; k++) {
      const int64_t t = base + k * kThreads + tid;
      if (t >= p.T) break;
      const bool m = p.mask ? p.mask[t] != 0 : p.mask_f32[t] != 0.f;
      const float lp = p.logp[t], px = p.prox[t], ol = p.old[t];
      float A = p.adv[t];
      if (norm) {
        const int g = p.adv_group ? p.adv_group[t] : 0;
        const double mu = p.adv_mean ? p.adv_mean[g] : 0.0;
        const double dn = p.adv_denom ? p.adv_denom[g] : 1.0;
        A = (float)(((double)A * (m ? 1.0 : 0.0) - mu) / dn);
        if (p.adv_out) p.adv_out[t] = A;
      }
      const PpoToken tk = ppo_token(lp, px, ol, A, m, p.clip_lo, p.clip_hi, p.c_clip, p.cap);
      const float pgw = tk.pgw;
      if (m) {
        tsum += pgw;
        tcnt += 1;
      }
      if (p.s_loss) p.s_loss[t] = pgw;
      if (p.s_iw) p.s_iw[t] = tk.ratio;
      if (p.s_kl) p.s_kl[t] = tk.kl;
      if (p.s_clip) p.s_clip[t] = tk.clip;
      if (p.s_dual) p.s_dual[t] = tk.dual;
      if (p.s_bw) p.s_bw[t] = tk.w;
      if (p.s_bkl) p.s_bkl[t] = tk.bkl;
      if (p.s_bm) p.s_bm[t] = tk.bmask;
      if (p.dlogp) p.dlogp[t] = tk.dlogp;
      if (p.s_valid) p.s_valid[t] = m;
      if (p.s_clip_ratio) p.s_clip_ratio[t] = tk.clip ? 1.f : 0.f;
      if (p.s_dual_ratio) p.s_dual_ratio[t] = tk.dual ? 1.f : 0.f;
      if (p.s_clip_new) p.s_clip_new[t] = tk.clip ? lp : 0.f;
      if (p.s_clip_old) p.s_clip_old[t] = tk.clip ? ol : 0.f;
    }
  }
  // block reduce (fp64)
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
    p.partial[2 * blockIdx.x] = bs;
    p.partial[2 * blockIdx.x + 1] = bc;
    __threadfence();
    const unsigned int done = atomicAdd(p.counter, 1u);
    is_last = (done == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last && tid < 32) {
    __threadfence();
    double fs = 0, fc = 0;
    const volatile double* part = p.partial;
    // lane-strided then a fixed shuffle tree: order depends only on gridDim
    for (int b = tid; b < (int)gridDim.x; b += 32) {
      fs += part[2 * b];
      fc += part[2 * b + 1];
    }
    fs = warp_sum(fs);
    fc = warp_sum(fc);
    if (tid == 0) {
      const float N = fc > 0.0 ? (float)fc : 1.f;  // `count_nonzero() or 1`
      *p.denom = N;
      *p.loss = (float)fs / N;  // fp32 sum divided by the count (functional.py:175)
    }
  }
}

__global__ void __launch_bounds__(256)
ppo_loss_bwd_kernel(const float* __restrict__ dlogp, const float* __restrict__ grad_loss,
                    const float* __restrict__ denom, float* __restrict__ g, int64_t T) {
  const float s = grad_loss[0] / denom[0];
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < T; t += (int64_t)gridDim.x * 256)
    g[t] = dlogp[t] * s;
}

}  // namespace

int ppo_loss_fwd_launch(const PpoLossArgs& a, void* workspace, size_t workspace_bytes, void* stream) {
  B200RL_REQUIRE(a.T >= 0, "negative T");
  B200RL_REQUIRE(a.loss && a.denom && workspace, "loss/denom/workspace must not be null");
  B200RL_REQUIRE(a.T == 0 || (a.logp && a.prox && a.old && a.adv && (a.mask || a.mask_f32)),
                 "null input pointer");
  B200RL_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "workspace must be 256-B aligned");
  B200RL_REQUIRE(!(a.c_clip > 0.f) || a.c_clip > 1.f, "c_clip must be > 1 (functional.py:157), got %f",
                 a.c_clip);
  LossWs W = loss_ws(workspace, a.T);
  if (workspace_bytes < W.bytes) {
    set_last_error("ppo loss workspace too small: %zu < %zu", workspace_bytes, W.bytes);
    return B200RL_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200RL_CUDA(cudaMemsetAsync(workspace, 0, 256, st));
  Params p;
  static_cast<PpoLossArgs&>(p) = a;
  p.counter = W.counter;
  p.partial = W.partial;
  ppo_loss_fwd_kernel<<<W.blocks, kThreads, 0, st>>>(p);
  return check_launch("ppo_loss_fwd_kernel");
}

}  // namespace b200rl

extern "C" {

size_t b200rl_ppo_loss_workspace_bytes(int64_t T) {
  if (T < 0) return 0;
  return b200rl::loss_ws(nullptr, T).bytes;
}

int b200rl_ppo_loss_fwd(const float* logp, const float* prox_logp, const float* old_logp,
                        const float* adv, const uint8_t* mask, int64_t T, float clip_lo,
                        float clip_hi, float c_clip, float cap, const double* adv_mean,
                        const double* adv_denom, const int32_t* adv_group, float* adv_out,
                        float* stat_loss, float* stat_imp_weight, float* stat_approx_kl,
                        uint8_t* stat_clip_mask, uint8_t* stat_dual_clip_mask,
                        float* stat_behave_imp_weight, float* stat_behave_approx_kl,
                        uint8_t* stat_behave_mask, float* dlogp, float* loss, float* denom,
                        void* workspace, size_t workspace_bytes, void* stream) {
  b200rl::PpoLossArgs p{};
  p.logp = logp; p.prox = prox_logp; p.old = old_logp; p.adv = adv; p.mask = mask; p.T = T;
  p.clip_lo = clip_lo; p.clip_hi = clip_hi; p.c_clip = c_clip; p.cap = cap;
  p.adv_mean = adv_mean; p.adv_denom = adv_denom; p.adv_group = adv_group; p.adv_out = adv_out;
  p.s_loss = stat_loss; p.s_iw = stat_imp_weight; p.s_kl = stat_approx_kl;
  p.s_clip = stat_clip_mask; p.s_dual = stat_dual_clip_mask;
  p.s_bw = stat_behave_imp_weight; p.s_bkl = stat_behave_approx_kl; p.s_bm = stat_behave_mask;
  p.dlogp = dlogp; p.loss = loss; p.denom = denom;
  return b200rl::ppo_loss_fwd_launch(p, workspace, workspace_bytes, stream);
}

int b200rl_ppo_loss_bwd(const float* dlogp, const float* grad_loss, const float* denom,
                        float* g_logp, int64_t T, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(T >= 0, "negative T");
  if (T == 0) return B200RL_OK;
  B200RL_REQUIRE(dlogp && grad_loss && denom && g_logp, "null pointer");
  int64_t blocks = (T + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  ppo_loss_bwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(dlogp, grad_loss, denom,
                                                                         g_logp, T);
  return check_launch("ppo_loss_bwd_kernel");
}

}  // extern "C"
