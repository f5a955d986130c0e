// H2 — masked moments and normalisation for advantage normalisation (batch / group level).
//
// Replaces the python loops of tiny aten kernels in areal/utils/functional.py:83-121
// (masked_normalization: ~12 launches per call, 64 calls for 512 sequences in groups of 8) and
// areal/engine/ppo/actor.py:538-604 (AdvNorm._compute_mean/_compute_std). All moment groups are
// reduced in ONE launch (grid.y = group), in fp64 like the reference's `high_precision=True`;
// the three moments of all groups sit in one contiguous [G,3] fp64 buffer so that the
// data-parallel case needs exactly one all-reduce instead of the reference's three.
//
// Bound: HBM, 8 B per element per pass (x + mask) + 4 B written by normalize — tiny tensors
// ([512,1024] at the GRPO config), so in practice launch-latency bound.
#include "common.cuh"

namespace b200rl {
namespace {

constexpr int kThreads = 256;
constexpr int kSplitElems = 8192;  // elements per CTA
constexpr int kMaxSplit = 64;

struct MomWs {
  unsigned int* counters;  // [n_groups]
  double* partial;         // [n_groups][nsplit][3]
  int nsplit;
  int64_t n_groups;
  size_t bytes;
};
MomWs mom_ws(void* ws, int64_t n_elems, int64_t group_elems) {
  MomWs W;
  W.n_groups = group_elems > 0 ? n_elems / group_elems : 0;
  int64_t s = (group_elems + kSplitElems - 1) / kSplitElems;
  if (s < 1) s = 1;
  if (s > kMaxSplit) s = kMaxSplit;
  W.nsplit = (int)s;
  size_t cbytes = ((size_t)W.n_groups * sizeof(unsigned int) + 255) & ~(size_t)255;
  W.counters = (unsigned int*)ws;
  W.partial = (double*)((char*)ws + cbytes);
  W.bytes = cbytes + (size_t)W.n_groups * W.nsplit * 3 * sizeof(double);
  W.bytes = (W.bytes + 255) & ~(size_t)255;
  return W;
}

__global__ void __launch_bounds__(kThreads)
masked_moments_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                      int64_t group_elems, const double* __restrict__ center,
                      int64_t center_group_elems, double* __restrict__ moments,
                      unsigned int* __restrict__ counters, double* __restrict__ partial,
                      int nsplit) {
  __shared__ double red[3][kThreads / 32];
  __shared__ bool is_last;
  const int tid = threadIdx.x;
  const int64_t g = blockIdx.y;
  const int sp = blockIdx.x;
  // contiguous sub-slice of the group for this CTA
  const int64_t per = (group_elems + nsplit - 1) / nsplit;
  const int64_t lo = g * group_elems + (int64_t)sp * per;
  int64_t hi = lo + per;
  const int64_t gend = (g + 1) * group_elems;
  if (hi > gend) hi = gend;
  double m0 = 0, m1 = 0, m2 = 0;
  for (int64_t i = lo + tid; i < hi; i += kThreads) {
    const double w = mask ? (double)mask[i] : 1.0;
    const double xv = (double)x[i];
    const double xm = xv * w;  // functional.py:103 `x = x * mask`
    m0 += w;
    m1 += xm;
    if (center) {
      // actor.py:590 `x_centered = x_masked - mean * mask`
      const double cm = xm - center[i / center_group_elems] * w;
      m2 += cm * cm;
    } else {
      m2 += xm * xm;  // functional.py:106
    }
  }
  m0 = warp_sum(m0);
  m1 = warp_sum(m1);
  m2 = warp_sum(m2);
  if ((tid & 31) == 0) {
    red[0][tid >> 5] = m0;
    red[1][tid >> 5] = m1;
    red[2][tid >> 5] = m2;
  }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0, c = 0;
    for (int w = 0; w < kThreads / 32; w++) {
      a += red[0][w];
      b += red[1][w];
      c += red[2][w];
    }
    double* out = partial + ((size_t)g * nsplit + sp) * 3;
    out[0] = a;
    out[1] = b;
    out[2] = c;
    __threadfence();
    is_last = atomicAdd(&counters[g], 1u) == (unsigned)nsplit - 1;
  }
  __syncthreads();
  if (is_last && tid == 0) {
    __threadfence();
    const volatile double* in = partial + (size_t)g * nsplit * 3;
    double a = 0, b = 0, c = 0;
    for (int s = 0; s < nsplit; s++) {  // fixed order
      a += in[3 * s];
      b += in[3 * s + 1];
      c += in[3 * s + 2];
    }
    moments[3 * g] = a;
    moments[3 * g + 1] = b;
    moments[3 * g + 2] = c;
  }
}

__global__ void moments_finalize_kernel(const double* __restrict__ moments, int64_t n_groups,
                                        int centered, int unbiased, double eps,
                                        double* __restrict__ mean, double* __restrict__ denom) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_groups) return;
  const double n = moments[3 * g], s1 = moments[3 * g + 1], s2 = moments[3 * g + 2];
  const double mu = s1 / n;
  double var = centered ? s2 / n : s2 / n - mu * mu;  // functional.py:116-118 / actor.py:600
  if (unbiased) var *= n / (n - 1.0);
  if (mean) mean[g] = mu;
  if (denom) denom[g] = sqrt(var) + eps;
}

__global__ void __launch_bounds__(256)
masked_normalize_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                        int64_t n_elems, const double* __restrict__ mean,
                        int64_t mean_group_elems, const double* __restrict__ denom,
                        int64_t denom_group_elems, int mask_x, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n_elems;
       i += (int64_t)gridDim.x * 256) {
    double xv = (double)x[i];
    if (mask_x && mask) xv *= (double)mask[i];
    const double mu = mean ? mean[i / mean_group_elems] : 0.0;
    const double dn = denom ? denom[i / denom_group_elems] : 1.0;
    out[i] = (float)((xv - mu) / dn);  // functional.py:121
  }
}

}  // namespace
}  // namespace b200rl

extern "C" {

size_t b200rl_moments_workspace_bytes(int64_t n_elems, int64_t group_elems) {
  if (n_elems < 0 || group_elems <= 0) return 0;
  return b200rl::mom_ws(nullptr, n_elems, group_elems).bytes;
}

int b200rl_masked_moments(const float* x, const float* mask, int64_t n_elems,
                          int64_t group_elems, const double* center, int64_t center_group_elems,
                          double* moments, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(n_elems >= 0 && group_elems > 0, "bad sizes");
  B200RL_REQUIRE(n_elems % group_elems == 0, "n_elems %lld is not a multiple of group_elems %lld",
                 (long long)n_elems, (long long)group_elems);
  if (n_elems == 0) return B200RL_OK;
  B200RL_REQUIRE(x && moments && workspace, "null pointer");
  B200RL_REQUIRE(!center || center_group_elems > 0, "center_group_elems must be > 0");
  B200RL_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "workspace must be 256-B aligned");
  MomWs W = mom_ws(workspace, n_elems, group_elems);
  B200RL_REQUIRE(W.n_groups <= 65535, "at most 65535 moment groups per call");
  if (workspace_bytes < W.bytes) {
    set_last_error("moments workspace too small: %zu < %zu", workspace_bytes, W.bytes);
    return B200RL_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200RL_CUDA(cudaMemsetAsync(W.counters, 0, (size_t)W.n_groups * sizeof(unsigned int), st));
  dim3 grid((unsigned)W.nsplit, (unsigned)W.n_groups);
  masked_moments_kernel<<<grid, kThreads, 0, st>>>(x, mask, group_elems, center,
                                                  center ? center_group_elems : 1, moments,
                                                  W.counters, W.partial, W.nsplit);
  return check_launch("masked_moments_kernel");
}

int b200rl_moments_finalize(const double* moments, int64_t n_groups, int centered, int unbiased,
                            double eps, double* mean, double* denom, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(n_groups >= 0, "negative n_groups");
  if (n_groups == 0) return B200RL_OK;
  B200RL_REQUIRE(moments != nullptr, "null pointer");
  moments_finalize_kernel<<<(unsigned)((n_groups + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      moments, n_groups, centered, unbiased, eps, mean, denom);
  return check_launch("moments_finalize_kernel");
}

int b200rl_masked_normalize(const float* x, const float* mask, int64_t n_elems,
                            const double* mean, int64_t mean_group_elems, const double* denom,
                            int64_t denom_group_elems, int mask_x, float* out, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(n_elems >= 0, "negative n_elems");
  if (n_elems == 0) return B200RL_OK;
  B200RL_REQUIRE(x && out, "null pointer");
  B200RL_REQUIRE(!mean || mean_group_elems > 0, "mean_group_elems must be > 0");
  B200RL_REQUIRE(!denom || denom_group_elems > 0, "denom_group_elems must be > 0");
  int64_t blocks = (n_elems + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  masked_normalize_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      x, mask, n_elems, mean, mean ? mean_group_elems : 1, denom, denom ? denom_group_elems : 1,
      mask_x, out);
  return check_launch("masked_normalize_kernel");
}

int b200rl_masked_normalize_groups(const float* x, const float* mask, int64_t n_elems,
                                   int64_t group_elems, int unbiased, double eps, float* out,
                                   double* moments, double* mean, double* denom, void* workspace,
                                   size_t workspace_bytes, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(n_elems == 0 || (moments && mean && denom), "moments/mean/denom must not be null");
  int rc = b200rl_masked_moments(x, mask, n_elems, group_elems, nullptr, 1, moments, workspace,
                                 workspace_bytes, stream);
  if (rc || n_elems == 0) return rc;
  rc = b200rl_moments_finalize(moments, n_elems / group_elems, 0, unbiased, eps, mean, denom, stream);
  if (rc) return rc;
  return b200rl_masked_normalize(x, mask, n_elems, mean, group_elems, denom, group_elems, 1, out, stream);
}

}  // extern "C"
