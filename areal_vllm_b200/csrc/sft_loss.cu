// Packed SFT loss tail: masked mean of -logp, per-sequence mean log-prob / perplexity, and the
// gradient seed, in one launch.
//   ref: areal/engine/sft/lm_engine.py:45-61 (compute_packed_sft_loss: where/sum/count_nonzero and
//        the python loop over cu_seqlens that syncs the host once per sequence)
// [T]-sized and latency-bound; sums are fp64 and added in a fixed order (deterministic).
#include "common.cuh"

namespace b200rl {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxBlocks = 148 * 2;

struct SftWs {
  unsigned int* counter;
  double* partial;
  int blocks;
  size_t bytes;
};
SftWs sft_ws(void* ws, int64_t T, int64_t batch) {
  SftWs W;
  int64_t need = (T + kThreads * 8 - 1) / (kThreads * 8);
  const int64_t need_b = (batch + kWarps - 1) / kWarps;
  if (need_b > need) need = need_b;
  if (need < 1) need = 1;
  if (need > kMaxBlocks) need = kMaxBlocks;
  W.blocks = (int)need;
  W.counter = (unsigned int*)ws;
  W.partial = (double*)((char*)ws + 256);
  W.bytes = ((256 + (size_t)W.blocks * 2 * sizeof(double)) + 255) & ~(size_t)255;
  return W;
}

template <typename CuT>
__global__ void __launch_bounds__(kThreads)
sft_loss_kernel(const float* __restrict__ logp, const uint8_t* __restrict__ mask,
                const CuT* __restrict__ cu, int64_t T, int batch, float* __restrict__ masked_logp,
                float* __restrict__ dlogp, double* __restrict__ seqlogp, float* __restrict__ ppl,
                float* __restrict__ loss, float* __restrict__ denom,
                unsigned int* __restrict__ counter, double* __restrict__ partial) {
  __shared__ double red[2][kWarps];
  __shared__ bool is_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // (1) token pass: where(mask, logp, 0), gradient seed, total sum / count
  double s = 0.0;
  int c = 0;
  for (int64_t t = (int64_t)blockIdx.x * kThreads + tid; t < T; t += (int64_t)gridDim.x * kThreads) {
    const bool m = mask[t] != 0;
    const float v = m ? logp[t] : 0.f;
    if (masked_logp) masked_logp[t] = v;
    if (dlogp) dlogp[t] = m ? -1.f : 0.f;  // d(N * loss)/d logp
    s += (double)v;
    c += m;
  }
  s = warp_sum(s);
  double cd = warp_sum((double)c);
  if (lane == 0) {
    red[0][warp] = s;
    red[1][warp] = cd;
  }

  // (2) one warp per sequence: mean log-prob over its valid tokens (0/0 = nan, as the reference)
  for (int64_t i = (int64_t)blockIdx.x * kWarps + warp; i < batch; i += (int64_t)gridDim.x * kWarps) {
    int64_t b = (int64_t)cu[i], e = (int64_t)cu[i + 1];
    if (b < 0) b = 0;
    if (e > T) e = T;
    int cc = 0;
    double acc = 0.0;
    for (int64_t t = b + lane; t < e; t += 32) {
      const bool m = mask[t] != 0;
      acc += m ? (double)logp[t] : 0.0;
      cc += m;
    }
    acc = warp_sum(acc);
    const double n = warp_sum((double)cc);
    if (lane == 0) {
      const float mean = (float)acc / (float)n;  // fp32 quotient stored into the fp64 array
      if (seqlogp) seqlogp[i] = (double)mean;
      if (ppl) ppl[i] = (float)exp(-(double)mean);
    }
  }

  __syncthreads();
  if (tid == 0) {
    double bs = 0, bc = 0;
    for (int w = 0; w < kWarps; w++) {
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
      *denom = (float)fc;
      *loss = -(float)fs / (float)fc;  // lm_engine.py:53: no `or 1`
    }
  }
}

}  // namespace
}  // namespace b200rl

extern "C" {

size_t b200rl_sft_loss_workspace_bytes(int64_t T, int64_t batch) {
  if (T < 0 || batch < 0) return 0;
  return b200rl::sft_ws(nullptr, T, batch).bytes;
}

int b200rl_sft_loss_fwd(const float* logp, const uint8_t* mask, const void* cu_seqlens,
                        int cu_is_int64, int64_t T, int32_t batch, float* masked_logp, float* dlogp,
                        double* seqlogp, float* ppl, float* loss, float* denom, void* workspace,
                        size_t workspace_bytes, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(T >= 0 && batch >= 0, "negative size");
  B200RL_REQUIRE(loss && denom && workspace, "loss/denom/workspace must not be null");
  B200RL_REQUIRE(T == 0 || (logp && mask), "null input pointer");
  B200RL_REQUIRE(batch == 0 || cu_seqlens, "cu_seqlens must not be null");
  B200RL_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "workspace must be 256-B aligned");
  SftWs W = sft_ws(workspace, T, batch);
  if (workspace_bytes < W.bytes) {
    set_last_error("sft loss workspace too small: %zu < %zu", workspace_bytes, W.bytes);
    return B200RL_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200RL_CUDA(cudaMemsetAsync(workspace, 0, 256, st));
  if (cu_is_int64)
    sft_loss_kernel<int64_t><<<W.blocks, kThreads, 0, st>>>(
        logp, mask, (const int64_t*)cu_seqlens, T, batch, masked_logp, dlogp, seqlogp, ppl, loss,
        denom, W.counter, W.partial);
  else
    sft_loss_kernel<int32_t><<<W.blocks, kThreads, 0, st>>>(
        logp, mask, (const int32_t*)cu_seqlens, T, batch, masked_logp, dlogp, seqlogp, ppl, loss,
        denom, W.counter, W.partial);
  return check_launch("sft_loss_kernel");
}

}  // extern "C"
