// Shared device helpers for the b200rl kernels (sm_100a only).
// No torch headers anywhere under csrc/: the library is a plain C-ABI .so.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200rl.h"

namespace b200rl {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// ---- error plumbing -------------------------------------------------------
void set_last_error(const char* fmt, ...);

#define B200RL_REQUIRE(cond, ...)            \
  do {                                       \
    if (!(cond)) {                           \
      b200rl::set_last_error(__VA_ARGS__);   \
      return B200RL_EINVAL;                  \
    }                                        \
  } while (0)

#define B200RL_CUDA(call)                                                   \
  do {                                                                      \
    cudaError_t e__ = (call);                                               \
    if (e__ != cudaSuccess) {                                               \
      b200rl::set_last_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,   \
                             cudaGetErrorString(e__));                      \
      return (int)e__;                                                      \
    }                                                                       \
  } while (0)

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("launch of %s failed: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

int sm_count();  // SM count of the current device (cached per device)
// opt a kernel into `bytes` of dynamic shared memory, once per (kernel, current device); thread-safe
int ensure_dynamic_smem(const void* kernel, int bytes);

// ---- launchers shared between translation units (not part of the C ABI) ----------------------
// logprob.cu. `roll` != 0: `labels` is the packed input_ids, row t's label is labels[(t+1) % T].
// bwd: when g_num is non-null the per-row upstream gradient is g_logp[t] * (g_num[0] / g_den[0]).
int logprob_fwd_impl(const void* logits, int dtype, int64_t ld, const int64_t* labels, int64_t T,
                     int64_t V, float temperature, float* logp, float* entropy, float* lse,
                     float* rowmax, float* rowmin, float* partial, bool want_ent, bool want_mm,
                     int roll, void* stream);
int logprob_bwd_impl(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                     const float* lse, const float* entropy, const float* g_logp,
                     const float* g_ent, int64_t T, int64_t V, float temperature, void* dlogits,
                     int64_t ld_out, int roll, const float* g_num, const float* g_den,
                     void* stream);

// ppo_loss.cu: everything b200rl_ppo_loss_fwd takes, plus an fp32 mask alternative and the
// derived logging tensors of grpo_loss_fn (areal/engine/ppo/actor.py:339-365).
struct PpoLossArgs {
  const float *logp, *prox, *old, *adv;
  const uint8_t* mask;     // bytes (torch.bool) ...
  const float* mask_f32;   // ... or fp32 (non-zero = valid) when mask == nullptr
  int64_t T;
  float clip_lo, clip_hi, c_clip, cap;
  const double *adv_mean, *adv_denom;
  const int32_t* adv_group;
  float* adv_out;
  float *s_loss, *s_iw, *s_kl;
  uint8_t *s_clip, *s_dual;
  float *s_bw, *s_bkl;
  uint8_t* s_bm;
  // extras (all nullable): the mask as bytes, clip_mask / dual_clip_mask as fp32, and
  // where(clip_mask, logp, 0) / where(clip_mask, old_logp, 0)
  uint8_t* s_valid;
  float *s_clip_ratio, *s_dual_ratio, *s_clip_new, *s_clip_old;
  float *dlogp, *loss, *denom;
};
int ppo_loss_fwd_launch(const PpoLossArgs& a, void* workspace, size_t workspace_bytes, void* stream);

// ---- small PTX wrappers ---------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Packed fp32x2 arithmetic (sm_100: FFMA2 / FADD2): two IEEE-rounded fp32 operations per issued
// instruction, bit-identical to the scalar forms. The streaming kernels are issue-bound, not
// FMA-pipe-bound, so halving the instruction count of their per-element FMAs/ADDs is what pays.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long ra, rb, rc, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  unsigned long long ra, rb, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
}

// bf16 pair packed in one 32-bit register -> two fp32 (exact)
__device__ __forceinline__ float bf16lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

__device__ __forceinline__ uint32_t max_bf16x2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t min_bf16x2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("min.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
// two fp32 -> packed bf16x2 (round-to-nearest-even), lo in bits [0,16)
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// mbarrier (shared::cta) -----------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// thread-block cluster: distributed shared memory of the CTAs of ONE cluster (same GPU) ---------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cta address of this CTA -> shared::cluster address of the same offset in CTA `rank`
__device__ __forceinline__ uint32_t dsmem_addr(uint32_t cta_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
  return r;
}
// Asynchronous store into the shared memory of any CTA of the cluster; the bytes count as
// complete_tx on the mbarrier `mbar` of the SAME target CTA, and whoever then sees that barrier's
// phase complete sees the data (the mechanism TMA loads use) - no cluster-scope fence involved.
__device__ __forceinline__ void st_async_v4(uint32_t addr, float a, float b, float c, float d,
                                            uint32_t mbar) {
  asm volatile(
      "st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(addr),
      "r"(__float_as_uint(a)), "r"(__float_as_uint(b)), "r"(__float_as_uint(c)),
      "r"(__float_as_uint(d)), "r"(mbar)
      : "memory");
}
__device__ __forceinline__ void st_async_b32(uint32_t addr, float a, uint32_t mbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(addr),
               "r"(__float_as_uint(a)), "r"(mbar)
               : "memory");
}
// one arrival + `bytes` more expected on an mbarrier of any CTA of the cluster (orders nothing)
__device__ __forceinline__ void mbar_arrive_expect_tx_dsmem(uint32_t mbar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.relaxed.cluster.shared::cluster.b64 _, [%0], %1;" ::"r"(mbar),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// L2 eviction policy for data that is streamed exactly once.
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// TMA 1-D bulk copy global -> shared::cta, completion on an mbarrier
// (SASS: UBLKCP). dst/src 16-B aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes,
                                         uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}

// streaming (no L1 allocate) 16-B global accesses
__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
// (plain, .cs and .wt variants of this store were measured in the backward: within 1 % of each other)
__device__ __forceinline__ void stg_na_v4(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// 16-B streaming store with an explicit L2 eviction policy
__device__ __forceinline__ void stg_hint_v4(void* p, uint4 v, uint64_t policy) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p),
               "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(policy)
               : "memory");
}

// ---- warp reductions ------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace b200rl
