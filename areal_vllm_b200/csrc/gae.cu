// H3 — GAE over packed variable-length sequences.
//
// Replaces csrc/cugae/gae.cu:10-60 of the reference (one *thread* per sequence, serial and
// uncoalesced: 256 sequences occupy one CTA on one SM). Two kernels:
//
//  gae_serial_kernel  one warp per sequence. Loads/stores are warp-coalesced; the recurrence
//                     itself is evaluated by one lane in exactly the reference's order with
//                     the reference's fp32 contractions, so results are bit-identical to
//                     gae_kernel_1d_nolp_misalign. Used when gamma*lam ~ 1, where any
//                     re-association of the fp32 suffix sum moves results by more than the
//                     1e-5 parity bar (the sum has no forgetting, errors random-walk).
//                     gae_serial_relay_kernel (same bits, the chain relayed lane to lane in
//                     registers) takes over from 1024 sequences up, gae_serial_slots_kernel
//                     (8 ticket-fed chains per warp) for very many short sequences.
//
//  gae_scan_kernel    the packed token stream is cut into fixed tiles of 1024 tokens, one warp
//                     per tile, whatever the sequence lengths are (perfect balance: a 16k-token
//                     sequence is spread over 16 warps, 16 sequences of 64 tokens share one).
//                     Inside a tile: segmented reverse affine scan
//                         (a1,b1) o (a2,b2) = (a1*a2, b1 + a1*b2),  a_t = gamma*lam (0 at a
//                     sequence's last token), b_t = delta_t. Across tiles: single-pass decoupled
//                     look-back on 64-bit {flag,value} words (tile g needs adv at the first
//                     token of tile g+1; tiles take tickets right-to-left so a tile only ever
//                     waits on tiles that already run).
//
// Bound: HBM. Algorithmic bytes: 16 B per token (r, v read; adv, ret written) + 9 B/sequence.
#include "common.cuh"

#include <math.h>

namespace b200rl {
namespace {

// ============================ SERIAL =========================================
// The kernel is ISSUE-bound (ncu, round 1: 4.1 warp instructions per token, issue slots 64 % busy,
// DRAM 33 %): what it costs is instructions, not bytes. Round 2 therefore (i) anchors every
// 256-token block at its TOP, so the block maps 1:1 onto the shared-memory slots, the chain always
// starts on a 16-byte boundary and only the block that reaches the sequence's start is ragged;
// (ii) addresses the block through three per-lane base pointers with compile-time offsets and takes
// an unpredicated path for full blocks (the round-1 code spent ~1.4 instructions per token on 64-bit
// index arithmetic and predicates); (iii) runs the chain 32 tokens per iteration: 32 dependent FMAs
// + 8 LDS.128 + 8 STS.128, i.e. 1.5 instructions per token plus ~0.1 of loop overhead.
constexpr int kSerWarps = 4;
#ifndef B200RL_SER_PERLANE
#define B200RL_SER_PERLANE 8
#endif
constexpr int kSerPerLane = B200RL_SER_PERLANE;  // tokens per lane per block
constexpr int kSerBlk = 32 * kSerPerLane;    // 256 tokens per warp iteration: the 4-cycle/token
                                             // chain of one block (~1000 cycles) covers the DRAM
                                             // latency of the next block's prefetch

struct SerBlock {
  float r[kSerPerLane], v[kSerPerLane], vn[kSerPerLane];
};

// tokens [hi - 256, hi) of one sequence; slot j = 32*k + lane <-> token hi - 256 + j. Slots of
// tokens below 0 (only in the block that reaches the sequence's start) are not loaded.
__device__ __forceinline__ void ser_load(SerBlock& B, const float* __restrict__ rbase,
                                         const float* __restrict__ vbase, int hi, int lane) {
  const int lo = hi - kSerBlk;
  const float* rp = rbase + lo + lane;
  const float* vp = vbase + lo + lane;
  if (lo >= 0) {  // warp-uniform: the common, unpredicated path
#pragma unroll
    for (int k = 0; k < kSerPerLane; k++) {
      B.r[k] = __ldg(rp + 32 * k);
      B.v[k] = __ldg(vp + 32 * k);
      B.vn[k] = __ldg(vp + 32 * k + 1);
    }
  } else {
    const int first = -lo - lane;  // slot 32*k + lane holds a token iff 32*k >= first
#pragma unroll
    for (int k = 0; k < kSerPerLane; k++) {
      const bool ok = 32 * k >= first;
      B.r[k] = ok ? __ldg(rp + 32 * k) : 0.f;
      B.v[k] = ok ? __ldg(vp + 32 * k) : 0.f;
      B.vn[k] = ok ? __ldg(vp + 32 * k + 1) : 0.f;
    }
  }
}

__global__ void __launch_bounds__(kSerWarps * 32)
gae_serial_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                  const int32_t* __restrict__ cu, const uint8_t* __restrict__ boot,
                  float* __restrict__ adv, float* __restrict__ ret, int batch, float gamma,
                  float lam) {
  __shared__ __align__(16) float sm[kSerWarps][kSerBlk];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float* d = sm[w];
  const float c = __fmul_rn(gamma, lam);  // ref gae.cu:24 `gamma * lmbda * lastgae`
  const int nwarps = gridDim.x * kSerWarps;
  for (int i = blockIdx.x * kSerWarps + w; i < batch; i += nwarps) {
    const int rs = cu[i], re = cu[i + 1];
    const int n = re - rs;
    if (n <= 0) continue;
    const float* rbase = rewards + rs;
    const float* vbase = values + ((int64_t)rs + i);  // sequence i's values start at cu[i] + i
    const bool bs = boot[i] != 0;
    float X = 0.f;  // lastgae
    SerBlock B;
    int hi = n;
    ser_load(B, rbase, vbase, hi, lane);
    // the sequence's last step without bootstrap sees next value 0 (gae.cu:20-21): slot 255 of
    // the top block
    if (!bs && lane == 31) B.vn[kSerPerLane - 1] = 0.f;
    while (hi > 0) {
      const int lo = hi - kSerBlk;
      const int skip = lo < 0 ? -lo : 0;  // slots below `skip` hold no token
      // delta in the reference's contraction: fma(gamma, next_v, r) - v   (gae.cu:22-23)
      float vk[kSerPerLane];
#pragma unroll
      for (int k = 0; k < kSerPerLane; k++) {
        d[32 * k + lane] = __fsub_rn(__fmaf_rn(gamma, B.vn[k], B.r[k]), B.v[k]);
        vk[k] = B.v[k];
      }
      // prefetch the next block (to the left) while the chain runs
      if (lo > 0) ser_load(B, rbase, vbase, lo, lane);
      __syncwarp();
      if (lane == 0) {
        int q = kSerBlk - 1;
        // 32 tokens per iteration: eight 16-byte loads up front (the first FMA needs only the first
        // of them), 32 dependent FMAs in the reference's order, eight 16-byte stores
#pragma unroll 1
        for (; q - 31 >= skip; q -= 32) {
          float4 t[8];
#pragma unroll
          for (int g = 0; g < 8; g++) t[g] = *reinterpret_cast<const float4*>(d + q - 3 - 4 * g);
#pragma unroll
          for (int g = 0; g < 8; g++) {
            t[g].w = X = __fmaf_rn(c, X, t[g].w);
            t[g].z = X = __fmaf_rn(c, X, t[g].z);
            t[g].y = X = __fmaf_rn(c, X, t[g].y);
            t[g].x = X = __fmaf_rn(c, X, t[g].x);
          }
#pragma unroll
          for (int g = 0; g < 8; g++) *reinterpret_cast<float4*>(d + q - 3 - 4 * g) = t[g];
        }
#pragma unroll 1
        for (; q - 3 >= skip; q -= 4) {
          float4 t = *reinterpret_cast<const float4*>(d + q - 3);
          t.w = X = __fmaf_rn(c, X, t.w);
          t.z = X = __fmaf_rn(c, X, t.z);
          t.y = X = __fmaf_rn(c, X, t.y);
          t.x = X = __fmaf_rn(c, X, t.x);
          *reinterpret_cast<float4*>(d + q - 3) = t;
        }
#pragma unroll 1
        for (; q >= skip; q--) {
          X = __fmaf_rn(c, X, d[q]);
          d[q] = X;
        }
      }
      __syncwarp();
      X = __shfl_sync(0xffffffffu, X, 0);
      float* ap = adv + rs + lo + lane;
      float* tp = ret + rs + lo + lane;
      if (lo >= 0) {
#pragma unroll
        for (int k = 0; k < kSerPerLane; k++) {
          const float A = d[32 * k + lane];
          ap[32 * k] = A;
          tp[32 * k] = __fadd_rn(A, vk[k]);  // gae.cu:26
        }
      } else {
        const int first = skip - lane;
#pragma unroll
        for (int k = 0; k < kSerPerLane; k++) {
          if (32 * k >= first) {
            const float A = d[32 * k + lane];
            ap[32 * k] = A;
            tp[32 * k] = __fadd_rn(A, vk[k]);
          }
        }
      }
      __syncwarp();
      hi = lo;
    }
  }
}

// ============================ SERIAL, team ===================================
// Few sequences (fewer than ~7 per SM) make the kernel a LATENCY problem: the time is the longest
// sequence's dependent FMA chain, 4 cycles per token at best, and gae_serial_kernel spends ~10
// because one warp does everything in turn (deltas -> shared memory, chain, results -> global) and
// the chain's own 16-byte shared loads/stores sit between its FMAs. Here two warps share a
// sequence through a ring of 256-token blocks in shared memory: warp 0 loads (four blocks ahead),
// forms the deltas and writes finished blocks out; lane 0 of warp 1 does nothing but the chain,
// with the next 32 deltas already in registers while the current 32 FMAs run, so the FMA pipe of
// that one lane is the only thing on the critical path. Same operations in the same order as the
// reference (gae.cu:10-28): same bits. Hand-off: one mbarrier per ring slot and direction.
constexpr int kTeamSlots = 4;  // ring depth
constexpr int kTeamAhead = 4;  // blocks of global loads in flight (registers of warp 0)
constexpr int kTeamLag = 2;    // warp 0 writes block j out after it has produced block j + kTeamLag
static_assert(kTeamSlots > kTeamLag, "a slot is reused only after its results were written out");

__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w));
}
__device__ __forceinline__ void team_wait(uint64_t* bar, uint32_t parity) {
  // microseconds at most; the bound turns a lost arrival into an error instead of a hung GPU
  for (uint32_t polls = 0; !mbar_try_wait(bar, parity);)
    if (++polls > (1u << 26)) __trap();
}

__global__ void __launch_bounds__(64)
gae_serial_team_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                       const int32_t* __restrict__ cu, const uint8_t* __restrict__ boot,
                       float* __restrict__ adv, float* __restrict__ ret, int batch, float gamma,
                       float lam) {
  __shared__ __align__(16) float sd[kTeamSlots][kSerBlk];  // deltas in, advantages out
  __shared__ __align__(16) float sv[kTeamSlots][kSerBlk];  // values, for ret = adv + v
  __shared__ uint64_t full[kTeamSlots], done[kTeamSlots];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kTeamSlots; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&done[s], 1);
    }
    fence_barrier_init();
  }
  __syncthreads();
  const float c = __fmul_rn(gamma, lam);  // ref gae.cu:24 `gamma * lmbda * lastgae`
  uint32_t gb = 0;  // blocks this team has been through: block g uses slot g % S in phase (g / S) & 1
  for (int i = blockIdx.x; i < batch; i += gridDim.x) {
    const int rs = cu[i], re = cu[i + 1];
    const int n = re - rs;
    if (n <= 0) continue;
    const int nb = (n + kSerBlk - 1) / kSerBlk;  // block j = tokens [n - 256 (j+1), n - 256 j)
    if (w == 0) {
      // ---------------- producer / writer ----------------
      const float* rbase = rewards + rs;
      const float* vbase = values + ((int64_t)rs + i);  // sequence i's values start at cu[i] + i
      // four blocks of loads in flight: the chain needs ~0.5 us per block, DRAM ~1 us
      SerBlock Ba, Bb, Bc, Bd;
      ser_load(Ba, rbase, vbase, n, lane);
      // the sequence's last step without bootstrap sees next value 0 (gae.cu:20-21)
      if (boot[i] == 0 && lane == 31) Ba.vn[kSerPerLane - 1] = 0.f;
      if (nb > 1) ser_load(Bb, rbase, vbase, n - kSerBlk, lane);
      if (nb > 2) ser_load(Bc, rbase, vbase, n - 2 * kSerBlk, lane);
      if (nb > 3) ser_load(Bd, rbase, vbase, n - 3 * kSerBlk, lane);
      auto step = [&](SerBlock& B, int j) {
        if (j < nb) {
          const uint32_t g = gb + (uint32_t)j;
          float* d = sd[g % kTeamSlots];
          float* v = sv[g % kTeamSlots];
#pragma unroll
          for (int k = 0; k < kSerPerLane; k++) {
            // delta in the reference's contraction: fma(gamma, next_v, r) - v   (gae.cu:22-23)
            d[32 * k + lane] = __fsub_rn(__fmaf_rn(gamma, B.vn[k], B.r[k]), B.v[k]);
            v[32 * k + lane] = B.v[k];
          }
          if (j + kTeamAhead < nb) ser_load(B, rbase, vbase, n - (j + kTeamAhead) * kSerBlk, lane);
          __syncwarp();
          if (lane == 0) mbar_arrive(&full[g % kTeamSlots]);
        }
        const int jo = j - kTeamLag;
        if (jo >= 0 && jo < nb) {
          const uint32_t g = gb + (uint32_t)jo;
          team_wait(&done[g % kTeamSlots], (g / kTeamSlots) & 1u);
          const float* d = sd[g % kTeamSlots];
          const float* v = sv[g % kTeamSlots];
          const int lo = n - (jo + 1) * kSerBlk;
          float* ap = adv + rs + lo + lane;
          float* tp = ret + rs + lo + lane;
          if (lo >= 0) {
#pragma unroll
            for (int k = 0; k < kSerPerLane; k++) {
              const float A = d[32 * k + lane];
              ap[32 * k] = A;
              tp[32 * k] = __fadd_rn(A, v[32 * k + lane]);  // gae.cu:26
            }
          } else {
            const int first = -lo - lane;  // slot 32*k + lane holds a token iff 32*k >= first
#pragma unroll
            for (int k = 0; k < kSerPerLane; k++) {
              if (32 * k >= first) {
                const float A = d[32 * k + lane];
                ap[32 * k] = A;
                tp[32 * k] = __fadd_rn(A, v[32 * k + lane]);
              }
            }
          }
          __syncwarp();  // every lane has read the slot before this warp writes into it again
        }
      };
      for (int j = 0; j < nb + kTeamLag; j += kTeamAhead) {
        step(Ba, j);
        step(Bb, j + 1);
        step(Bc, j + 2);
        step(Bd, j + 3);
      }
    } else {
      // ---------------- the chain ----------------
      float X = 0.f;  // lastgae
      for (int j = 0; j < nb; j++) {
        const uint32_t g = gb + (uint32_t)j;
        team_wait(&full[g % kTeamSlots], (g / kTeamSlots) & 1u);
        if (lane == 0) {
          float* d = sd[g % kTeamSlots];
          const int lo = n - (j + 1) * kSerBlk;
          if (lo >= 0) {
            // a whole block: 8 groups of 32 tokens, the next group's 8 vectors are requested before
            // this group's 32 dependent FMAs. (ptxas interleaves one 16-byte load and one store
            // per four FMAs: 1025 cycles per block by its own stall counts, and ncu's sampling
            // agrees; the ~5.5 cycles per token measured end to end include what happens between
            // blocks)
            const uint32_t da = smem_u32(d);
            uint4 cur[8], nxt[8];
#pragma unroll
            for (int q4 = 0; q4 < 8; q4++) cur[q4] = lds128(da + 4u * (kSerBlk - 4 - 4 * q4));
#pragma unroll
            for (int grp = 0; grp < kSerBlk / 32; grp++) {
              const int top = kSerBlk - 32 * grp;  // this group = slots [top - 32, top)
              if (grp + 1 < kSerBlk / 32) {
#pragma unroll
                for (int q4 = 0; q4 < 8; q4++) nxt[q4] = lds128(da + 4u * (top - 32 - 4 - 4 * q4));
              }
              uint4 o[8];
#pragma unroll
              for (int q4 = 0; q4 < 8; q4++) {
                X = __fmaf_rn(c, X, __uint_as_float(cur[q4].w));
                o[q4].w = __float_as_uint(X);
                X = __fmaf_rn(c, X, __uint_as_float(cur[q4].z));
                o[q4].z = __float_as_uint(X);
                X = __fmaf_rn(c, X, __uint_as_float(cur[q4].y));
                o[q4].y = __float_as_uint(X);
                X = __fmaf_rn(c, X, __uint_as_float(cur[q4].x));
                o[q4].x = __float_as_uint(X);
              }
#pragma unroll
              for (int q4 = 0; q4 < 8; q4++) sts128(da + 4u * (top - 4 - 4 * q4), o[q4]);
              if (grp + 1 < kSerBlk / 32) {
#pragma unroll
                for (int q4 = 0; q4 < 8; q4++) cur[q4] = nxt[q4];
              }
            }
          } else {
            // the block that reaches the sequence's start: slots below `skip` hold no token
            const int skip = -lo;
            int q = kSerBlk - 1;
#pragma unroll 1
            for (; q - 3 >= skip; q -= 4) {
              float4 t = *reinterpret_cast<const float4*>(d + q - 3);
              t.w = X = __fmaf_rn(c, X, t.w);
              t.z = X = __fmaf_rn(c, X, t.z);
              t.y = X = __fmaf_rn(c, X, t.y);
              t.x = X = __fmaf_rn(c, X, t.x);
              *reinterpret_cast<float4*>(d + q - 3) = t;
            }
#pragma unroll 1
            for (; q >= skip; q--) {
              X = __fmaf_rn(c, X, d[q]);
              d[q] = X;
            }
          }
          mbar_arrive(&done[g % kTeamSlots]);
        }
        __syncwarp();
      }
    }
    gb += (uint32_t)nb;
  }
}

// ============================ SERIAL, relay ==================================
// Same warp-per-sequence walk as gae_serial_kernel, but the chain never touches shared memory:
// after the coalesced loads the 256 deltas are transposed ONCE (8 scalar stores + 2 16-byte loads
// per lane, all 32 lanes busy) so that lane L holds the 8 consecutive tokens 8L..8L+7 in
// registers; the recurrence then runs as a relay - lane 31 does its 8 dependent FMAs, hands X to
// lane 30 with one shuffle, and so on down to lane 0 - in exactly the reference's order, so the
// bits are the same. Per 256-token block that is 32 SHFL + 20 full-warp shared-memory
// instructions instead of 128 single-lane LDS.128/STS.128 + 16, which is what the warps of
// gae_serial_kernel queue for (mio_throttle / short_scoreboard).
constexpr int kRelayPad = kSerBlk + (kSerBlk / 32) * 4;  // 4 floats of padding per 32: both the strided
                                                          // scalar and the consecutive 16-byte accesses
                                                          // are bank-conflict free
__device__ __forceinline__ int relay_slot(int j) { return j + ((j >> 5) << 2); }

#ifndef B200RL_RELAY_MINB
#define B200RL_RELAY_MINB 7
#endif
__global__ void __launch_bounds__(kSerWarps * 32, B200RL_RELAY_MINB)
gae_serial_relay_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                        const int32_t* __restrict__ cu, const uint8_t* __restrict__ boot,
                        float* __restrict__ adv, float* __restrict__ ret, int batch, float gamma,
                        float lam) {
  static_assert(kSerPerLane == 8, "the relay owns 8 consecutive tokens per lane");
  __shared__ __align__(16) float sm[kSerWarps][kRelayPad];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float* d = sm[w];
  const float c = __fmul_rn(gamma, lam);  // ref gae.cu:24
  const int nwarps = gridDim.x * kSerWarps;
  for (int i = blockIdx.x * kSerWarps + w; i < batch; i += nwarps) {
    const int rs = cu[i], re = cu[i + 1];
    const int n = re - rs;
    if (n <= 0) continue;
    const float* rbase = rewards + rs;
    const float* vbase = values + ((int64_t)rs + i);
    const bool bs = boot[i] != 0;
    float X = 0.f;  // lastgae entering the block at its top (lane 31)
    SerBlock B;
    int hi = n;
    ser_load(B, rbase, vbase, hi, lane);
    if (!bs && lane == 31) B.vn[kSerPerLane - 1] = 0.f;  // gae.cu:20-21
    while (hi > 0) {
      const int lo = hi - kSerBlk;
      const int skip = lo < 0 ? -lo : 0;
      float vk[kSerPerLane];
#pragma unroll
      for (int k = 0; k < kSerPerLane; k++) {
        d[36 * k + lane] = __fsub_rn(__fmaf_rn(gamma, B.vn[k], B.r[k]), B.v[k]);  // slot 32k+lane, padded
        vk[k] = B.v[k];
      }
      if (lo > 0) ser_load(B, rbase, vbase, lo, lane);
      __syncwarp();
      float4 e0 = *reinterpret_cast<const float4*>(d + relay_slot(8 * lane));      // tokens 8L .. 8L+3
      float4 e1 = *reinterpret_cast<const float4*>(d + relay_slot(8 * lane + 4));  // tokens 8L+4 .. 8L+7
      // Relay from lane 31 (top of the block) down to the lane that holds the sequence's first
      // token, written without any predication: at EVERY step every lane re-runs its 8 FMAs on the
      // X its upper neighbour holds at that moment. Lane 31's input is the carry, so its result is
      // final after step 1 and never changes again; lane 30's is therefore final after step 2, and
      // so on: after k steps lanes 31..32-k hold exactly the serial chain's values (same operands,
      // same order, same roundings), whatever the lower lanes computed on the way. 10 instructions
      // per step (SHFL, SEL, 8 FFMA) instead of a divergent branch around them.
      const int steps = 32 - (skip >> 3);
      float4 a0 = e0, a1 = e1;
      float Xl = X;
#pragma unroll 4
      for (int s = 0; s < steps; ++s) {
        float x = __shfl_down_sync(0xffffffffu, Xl, 1);
        if (lane == 31) x = X;
        a1.w = x = __fmaf_rn(c, x, e1.w);
        a1.z = x = __fmaf_rn(c, x, e1.z);
        a1.y = x = __fmaf_rn(c, x, e1.y);
        a1.x = x = __fmaf_rn(c, x, e1.x);
        a0.w = x = __fmaf_rn(c, x, e0.w);
        a0.z = x = __fmaf_rn(c, x, e0.z);
        a0.y = x = __fmaf_rn(c, x, e0.y);
        a0.x = x = __fmaf_rn(c, x, e0.x);
        Xl = x;
      }
      X = __shfl_sync(0xffffffffu, Xl, 0);  // only meaningful while the sequence continues (32 steps)
      e0 = a0;
      e1 = a1;
      __syncwarp();
      *reinterpret_cast<float4*>(d + relay_slot(8 * lane)) = e0;
      *reinterpret_cast<float4*>(d + relay_slot(8 * lane + 4)) = e1;
      __syncwarp();
      float* ap = adv + rs + lo + lane;
      float* tp = ret + rs + lo + lane;
      if (lo >= 0) {
#pragma unroll
        for (int k = 0; k < kSerPerLane; k++) {
          const float A = d[36 * k + lane];
          ap[32 * k] = A;
          tp[32 * k] = __fadd_rn(A, vk[k]);  // gae.cu:26
        }
      } else {
        const int first = skip - lane;
#pragma unroll
        for (int k = 0; k < kSerPerLane; k++) {
          if (32 * k >= first) {
            const float A = d[36 * k + lane];
            ap[32 * k] = A;
            tp[32 * k] = __fadd_rn(A, vk[k]);
          }
        }
      }
      __syncwarp();
      hi = lo;
    }
  }
}

// ============================ SERIAL, many sequences =========================
// With very many SHORT sequences the one-warp-per-sequence kernel wastes most of each warp (a
// 20-token sequence uses 20 lanes once). Here a warp owns GS sequences and lane q runs the chain
// of sequence q, so one FMA instruction advances GS chains. (For long sequences the row-wise
// staging below costs more instructions per token than the warp kernel's: measured, not used.)
// Data moves through shared memory in CH-token chunks (from each sequence's end towards its
// start): cp.async copies every sequence's chunk of rewards and values row-wise (coalesced) into a
// ring of STAGES buffers, one or two chunks ahead of the arithmetic; the chains then run
// column-wise - delta and recurrence in the reference's order and contractions, so results are
// bit-identical - overwriting the rewards with adv and the (no longer needed) next-value slot with
// ret; finally the rows are written out coalesced. Row pitch CH+1 floats keeps the column accesses
// free of bank conflicts.
constexpr int kBatWarps = 4;

__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int GS, int CH, int STAGES>
__global__ void __launch_bounds__(kBatWarps * 32)
gae_serial_batched_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                          const int32_t* __restrict__ cu, const uint8_t* __restrict__ boot,
                          float* __restrict__ adv, float* __restrict__ ret, int batch, float gamma,
                          float lam) {
  static_assert(CH % 32 == 0 && GS <= 32 && STAGES >= 2, "shape");
  constexpr int LD = CH + 1;
  constexpr int PER = CH / 32;          // tokens per lane per row
  constexpr int STAGE = 2 * GS * LD;    // floats per stage: R rows then V rows
  extern __shared__ float bat_smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float* ring = bat_smem + (size_t)w * (STAGES * STAGE);
  const float c = __fmul_rn(gamma, lam);  // ref gae.cu:24
  const int ngroups = (batch + GS - 1) / GS;
  for (int grp = blockIdx.x * kBatWarps + w; grp < ngroups; grp += gridDim.x * kBatWarps) {
    const int i = grp * GS + lane;
    const bool valid = lane < GS && i < batch;
    const int rs = valid ? cu[i] : 0;
    int n = valid ? cu[i + 1] - rs : 0;
    if (n < 0) n = 0;
    const bool bs = valid ? boot[i] != 0 : false;
    int maxn = n;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) maxn = max(maxn, __shfl_xor_sync(0xffffffffu, maxn, o));
    const int nchunks = (maxn + CH - 1) / CH;

    auto issue = [&](int ck) {
      float* R = ring + (ck % STAGES) * STAGE;
      float* V = R + GS * LD;
      const int off = ck * CH;
#pragma unroll 4
      for (int q = 0; q < GS; q++) {
        const int nq = __shfl_sync(0xffffffffu, n, q);
        const int rsq = __shfl_sync(0xffffffffu, rs, q);
        const int hi = nq - off;
        if (hi <= 0) continue;  // warp-uniform: this sequence is finished
        const int lo = max(hi - CH, 0);
        const int cnt = hi - lo;
        const float* rp = rewards + rsq + lo;
        const float* vp = values + ((int64_t)rsq + grp * GS + q) + lo;  // values of sequence i start at cu[i] + i
#pragma unroll
        for (int k = 0; k < PER; k++) {
          const int jj = 32 * k + lane;
          if (jj < cnt) cp_async4(R + q * LD + jj, rp + jj);
          if (jj <= cnt) cp_async4(V + q * LD + jj, vp + jj);
        }
        if (lane == 0 && cnt == CH) cp_async4(V + q * LD + CH, vp + CH);  // the chunk's next value
      }
    };

#pragma unroll
    for (int s0 = 0; s0 < STAGES - 1; s0++) {
      if (s0 < nchunks) issue(s0);
      cp_async_commit();
    }
    float X = 0.f;  // lastgae of this lane's sequence
    for (int ck = 0; ck < nchunks; ck++) {
      if (ck + STAGES - 1 < nchunks) issue(ck + STAGES - 1);
      cp_async_commit();
      cp_async_wait<STAGES - 1>();
      __syncwarp();
      float* R = ring + (ck % STAGES) * STAGE;
      float* V = R + GS * LD;
      const int off = ck * CH;
      // ---- GS chains, one per lane, in the reference's order ------------------------------
      if (lane < GS) {
        const int hi = n - off;
        if (hi > 0) {
          const int cnt = min(hi, CH);
          float* rrow = R + lane * LD;
          float* vrow = V + lane * LD;
          int k = cnt - 1;
          float nv = vrow[k + 1];
          if (ck == 0 && !bs) nv = 0.f;  // gae.cu:20-21: the sequence's last step without bootstrap
#pragma unroll 4
          for (; k >= 0; k--) {
            const float cur = vrow[k];
            const float d = __fsub_rn(__fmaf_rn(gamma, nv, rrow[k]), cur);  // :22-23
            X = __fmaf_rn(c, X, d);                                       // :24
            rrow[k] = X;                                                  // adv
            vrow[k + 1] = __fadd_rn(X, cur);                              // ret (:26), slot k+1 is free now
            nv = cur;
          }
        }
      }
      __syncwarp();
      // ---- rows out ---------------------------------------------------------------------
#pragma unroll 4
      for (int q = 0; q < GS; q++) {
        const int nq = __shfl_sync(0xffffffffu, n, q);
        const int rsq = __shfl_sync(0xffffffffu, rs, q);
        const int hi = nq - off;
        if (hi <= 0) continue;
        const int lo = max(hi - CH, 0);
        const int cnt = hi - lo;
#pragma unroll
        for (int k = 0; k < PER; k++) {
          const int jj = 32 * k + lane;
          if (jj < cnt) {
            adv[rsq + lo + jj] = R[q * LD + jj];
            ret[rsq + lo + jj] = V[q * LD + jj + 1];
          }
        }
      }
      __syncwarp();  // the stage may be refilled by the next iteration's issue
    }
    cp_async_wait<0>();
    __syncwarp();
  }
}

template <int GS, int CH, int STAGES>
int launch_serial_batched(const float* rewards, const float* values, const int32_t* cu,
                          const uint8_t* boot, float* adv, float* ret, int batch, float gamma,
                          float lam, cudaStream_t st) {
  constexpr size_t smem = (size_t)kBatWarps * STAGES * 2 * GS * (CH + 1) * sizeof(float);
  auto k = gae_serial_batched_kernel<GS, CH, STAGES>;
  if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(k), (int)smem)) return rc;
  const int ngroups = (batch + GS - 1) / GS;
  int grid = (ngroups + kBatWarps - 1) / kBatWarps;
  const int cap = sm_count() * (int)(220 * 1024 / (smem + 1024));
  if (grid > cap) grid = cap;
  k<<<grid, kBatWarps * 32, smem, st>>>(rewards, values, cu, boot, adv, ret, batch, gamma, lam);
  return check_launch("gae_serial_batched_kernel");
}

// ============================ SERIAL, slots ==================================
// What limits gae_serial_kernel is not arithmetic or issue slots but the shared-memory pipe (ncu,
// round 2: mio_throttle 4.0 + short_scoreboard 4.4 stalled warps per issue, issue slots 45 % busy):
// its chain lane moves every token through shared memory with single-lane 16-byte accesses, and
// a one-lane LDS.128/STS.128 occupies the pipe like a full-warp one. Here a warp owns EIGHT
// sequence "slots" and lane s < 8 runs the chain of slot s, so one LDS.128/FFMA/STS.128 warp
// instruction serves eight chains (the pipe cost per token drops 8x). Unlike
// gae_serial_batched_kernel, slots are independent: each walks its own sequence from the top in
// 64-token blocks and takes the next unprocessed sequence (a global ticket counter) as soon as its
// own ends, so unequal lengths cost nothing (the batched kernel waits for the longest of its 8).
// Raw rewards/values of the NEXT block of all 8 slots are staged by 4-byte cp.async (any
// alignment, coalesced per slot) into the other half of a 2-stage ring while the chains run; the
// chain lane computes delta itself - fma(gamma, v_next, r) - v, then fma(c, X, delta), the
// reference's order and contractions, so the result is bit-identical - and overwrites r with adv
// and v with ret in place; all 32 lanes then write the rows out coalesced.
#ifndef B200RL_RELAY_MIN_BATCH
#define B200RL_RELAY_MIN_BATCH 1024
#endif
constexpr int kSlots = 8;            // sequences in flight per warp
#ifndef B200RL_SLOT_BLK
#define B200RL_SLOT_BLK 64
#endif
constexpr int kSlotBlk = B200RL_SLOT_BLK;  // tokens per slot per round (32 or 64)
constexpr int kSlotPitch = kSlotBlk + 4;   // floats per row: the tokens + the block's next value, padded so
                                     // that the 8 chain lanes' 16-byte accesses hit distinct banks
constexpr int kSlotWarps = 4;
constexpr int kSlotStage = 2 * kSlots * kSlotPitch;  // floats per stage: R rows then V rows
struct __align__(16) SlotDesc {
  int rpos, vpos;  // global index of row slot 0 in rewards (= adv/ret) and in values
  int skip;        // row slots < skip hold no token (kSlotBlk = the slot is idle)
  int flags;       // bit 0: the block is its sequence's top; bit 1: that sequence bootstraps
};
#ifndef B200RL_SLOT_STAGES
#define B200RL_SLOT_STAGES 2
#endif
constexpr int kSlotStages = B200RL_SLOT_STAGES;  // ring depth: blocks staged ahead of the chains
constexpr size_t kSlotWarpBytes = kSlotStages * (kSlotStage * sizeof(float) + kSlots * sizeof(SlotDesc));
constexpr size_t kSlotSmem = (size_t)kSlotWarps * kSlotWarpBytes;

__device__ __forceinline__ void cp_async4_s(uint32_t smem_dst, const float* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}

__global__ void __launch_bounds__(kSlotWarps * 32)
gae_serial_slots_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                        const int32_t* __restrict__ cu, const uint8_t* __restrict__ boot,
                        float* __restrict__ adv, float* __restrict__ ret, int batch, float gamma,
                        float lam, unsigned int* __restrict__ ticket) {
  extern __shared__ __align__(16) unsigned char slot_smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float* ring = reinterpret_cast<float*>(slot_smem + (size_t)w * kSlotWarpBytes);
  SlotDesc* dsc = reinterpret_cast<SlotDesc*>(ring + kSlotStages * kSlotStage);  // [kSlotStages][kSlots]
  const float c = __fmul_rn(gamma, lam);  // ref gae.cu:24
  const bool chain_lane = lane < kSlots;

  // ---- per-slot state, held by lane s < 8 ------------------------------------------------------
  int hi = 0, rs = 0, vs = 0, n = 0;      // current sequence: tokens [0, hi) still to do
  bool bflag = false;
  int nx_i = batch, nx_rs = 0, nx_re = 0;  // the sequence after it, fetched one sequence ahead
  bool nx_boot = false;
  // without a ticket counter: static round-robin over all slots of the grid
  int static_next = (blockIdx.x * kSlotWarps + w) * kSlots + lane;
  const int static_stride = gridDim.x * kSlotWarps * kSlots;
  auto fetch_next = [&]() {
    int i;
    if (ticket) {
      i = (int)atomicAdd(ticket, 1u);
    } else {
      i = static_next;
      static_next += static_stride;
    }
    if (i < 0 || i >= batch) i = batch;
    nx_i = i;
    if (i < batch) {
      nx_rs = cu[i];
      nx_re = cu[i + 1];
      nx_boot = boot[i] != 0;
    }
  };
  auto refill = [&]() {  // rare: once per sequence
    while (nx_i < batch && nx_re - nx_rs <= 0) fetch_next();  // empty sequences produce nothing
    if (nx_i >= batch) return;
    rs = nx_rs;
    n = nx_re - nx_rs;
    vs = nx_rs + nx_i;  // values of sequence i start at cu[i] + i
    bflag = nx_boot;
    hi = n;
    fetch_next();
  };
  auto next_desc = [&]() {
    SlotDesc D{0, 0, kSlotBlk, 0};
    if (!chain_lane) return D;
    if (hi <= 0) refill();
    if (hi <= 0) return D;
    const int lo = hi - kSlotBlk;
    D.skip = lo < 0 ? -lo : 0;
    D.rpos = rs + lo;
    D.vpos = vs + lo;
    D.flags = (hi == n ? 1 : 0) | (bflag ? 2 : 0);
    hi = lo > 0 ? lo : 0;
    return D;
  };
  // all 32 lanes: stage the blocks described by dsc[stage] into ring[stage]. Branch-free: the
  // eight descriptors are loaded first (one 16-byte broadcast load each, all in flight together),
  // every copy is predicated on its row slot holding a token (idle slots have skip = kSlotBlk).
  auto issue = [&](int stage) {
    const uint32_t sR = smem_u32(ring + stage * kSlotStage) + lane * 4;
    const uint32_t sV = sR + kSlots * kSlotPitch * 4;
    const SlotDesc* D = dsc + stage * kSlots;
    SlotDesc d[kSlots];
#pragma unroll
    for (int q = 0; q < kSlots; q++) d[q] = D[q];
#pragma unroll
    for (int q = 0; q < kSlots; q++) {
      const float* rp = rewards + d[q].rpos + lane;
      const float* vp = values + d[q].vpos + lane;
      const uint32_t r0 = sR + q * kSlotPitch * 4, v0 = sV + q * kSlotPitch * 4;
      if (lane >= d[q].skip) {
        cp_async4_s(r0, rp);
        cp_async4_s(v0, vp);
      }
      if (kSlotBlk > 32 && lane + 32 >= d[q].skip) {
        cp_async4_s(r0 + 128, rp + 32);
        cp_async4_s(v0 + 128, vp + 32);
      }
      // the value after the block's top token
      if (lane == 0 && d[q].skip < kSlotBlk) cp_async4_s(v0 + kSlotBlk * 4, vp + kSlotBlk);
    }
  };

  if (chain_lane) fetch_next();
  // prologue: kSlotStages - 1 blocks per slot in flight
#pragma unroll
  for (int p = 0; p < kSlotStages - 1; p++) {
    const SlotDesc d0 = next_desc();
    if (chain_lane) dsc[p * kSlots + lane] = d0;
    __syncwarp();
    issue(p);
    cp_async_commit();
  }
  float X = 0.f;  // lastgae of this lane's slot
  int stage = 0;
  for (;;) {
    const SlotDesc cur = dsc[stage * kSlots + (lane & (kSlots - 1))];
    if (!__any_sync(0xffffffffu, chain_lane && cur.skip < kSlotBlk)) break;  // every slot has run dry
    const int fill = stage == 0 ? kSlotStages - 1 : stage - 1;  // the stage freed by the previous round
    const SlotDesc nxt = next_desc();
    if (chain_lane) dsc[fill * kSlots + lane] = nxt;
    __syncwarp();
    issue(fill);
    cp_async_commit();
    cp_async_wait<kSlotStages - 1>();
    __syncwarp();
    float* R = ring + stage * kSlotStage;
    float* V = R + kSlots * kSlotPitch;
    // ---- 8 chains, one per lane, 4 tokens per step, in the reference's order --------------------
    if (chain_lane && cur.skip < kSlotBlk) {
      float* rrow = R + lane * kSlotPitch;
      const float* vrow = V + lane * kSlotPitch;
      if (cur.flags & 1) X = 0.f;
      float nv = vrow[kSlotBlk];
      if ((cur.flags & 3) == 1) nv = 0.f;  // gae.cu:20-21: the last step without bootstrap
      const int gmin = cur.skip >> 2;      // groups below hold no token (the lowest may be partial:
                                           // its empty slots are computed on stale data and never
                                           // written out, and the sequence ends there)
      // two register sets in ping-pong (no copies between them): while one group's four
      // dependent FMAs run, the other group's two 16-byte loads are in flight
      auto step4 = [&](const float4& r4, const float4& v4, float* at) {
        float4 a4;
        a4.w = X = __fmaf_rn(c, X, __fsub_rn(__fmaf_rn(gamma, nv, r4.w), v4.w));    // gae.cu:22-24
        a4.z = X = __fmaf_rn(c, X, __fsub_rn(__fmaf_rn(gamma, v4.w, r4.z), v4.z));
        a4.y = X = __fmaf_rn(c, X, __fsub_rn(__fmaf_rn(gamma, v4.z, r4.y), v4.y));
        a4.x = X = __fmaf_rn(c, X, __fsub_rn(__fmaf_rn(gamma, v4.y, r4.x), v4.x));
        nv = v4.x;
        *reinterpret_cast<float4*>(at) = a4;  // adv over the rewards; the values stay
      };
      int g = kSlotBlk / 4 - 1;
      float4 ra = *reinterpret_cast<const float4*>(rrow + 4 * g);
      float4 va = *reinterpret_cast<const float4*>(vrow + 4 * g);
      float4 rb = ra, vb = va;
#pragma unroll 1
      while (true) {
        const bool more1 = g - 1 >= gmin;
        if (more1) {
          rb = *reinterpret_cast<const float4*>(rrow + 4 * g - 4);
          vb = *reinterpret_cast<const float4*>(vrow + 4 * g - 4);
        }
        step4(ra, va, rrow + 4 * g);
        if (!more1) break;
        const bool more2 = g - 2 >= gmin;
        if (more2) {
          ra = *reinterpret_cast<const float4*>(rrow + 4 * g - 8);
          va = *reinterpret_cast<const float4*>(vrow + 4 * g - 8);
        }
        step4(rb, vb, rrow + 4 * g - 4);
        if (!more2) break;
        g -= 2;
      }
    }
    __syncwarp();
    // ---- rows out, coalesced; ret = adv + v (gae.cu:26) by all lanes; branch-free like issue() ----
    {
      const float* Rl = R + lane;
      const float* Vl = V + lane;
      const SlotDesc* D = dsc + stage * kSlots;
      SlotDesc d[kSlots];
#pragma unroll
      for (int q = 0; q < kSlots; q++) d[q] = D[q];
#pragma unroll
      for (int q = 0; q < kSlots; q++) {
        float* ap = adv + d[q].rpos + lane;
        float* tp = ret + d[q].rpos + lane;
        const float a0 = Rl[q * kSlotPitch];
        const float t0 = __fadd_rn(a0, Vl[q * kSlotPitch]);
        if (lane >= d[q].skip) {
          ap[0] = a0;
          tp[0] = t0;
        }
        if (kSlotBlk > 32) {
          const float a1 = Rl[q * kSlotPitch + 32];
          const float t1 = __fadd_rn(a1, Vl[q * kSlotPitch + 32]);
          if (lane + 32 >= d[q].skip) {
            ap[32] = a1;
            tp[32] = t1;
          }
        }
      }
    }
    __syncwarp();  // the stage (and its descriptors) are rewritten by the next round
    stage = stage + 1 == kSlotStages ? 0 : stage + 1;
  }
  cp_async_wait<0>();
}

// ============================ SCAN ===========================================
constexpr int kTile = 1024;   // tokens per warp tile
constexpr int kBlk = 128;     // tokens per scan step (32 lanes x 4)
constexpr int kScanWarps = 4;
constexpr uint32_t kFlagAgg = 1u, kFlagIncl = 2u;
constexpr int kMetaLast = 1 << 30;
constexpr int kMetaBoot = 1 << 29;
constexpr int kMetaCnt = (1 << 24) - 1;

struct ScanSmem {
  float d[kTile];    // delta, then adv (zero carry-in)
  int meta[kTile];   // start-count / relative sequence id, LAST/BOOT flags
};

__device__ __forceinline__ uint64_t ld_state(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_state(uint64_t* p, uint32_t flag, float val) {
  uint64_t v = ((uint64_t)flag << 32) | (uint64_t)__float_as_uint(val);
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// #{k in [1, batch] : cu[k] <= x}, warp-cooperative 32-ary search (all lanes return it)
__device__ __forceinline__ int count_ends_le(const int32_t* __restrict__ cu, int batch, int x,
                                             int lane) {
  int lo = 1, hi = batch + 1;  // predicate true for k < lo, false for k >= hi
  while (lo < hi) {
    const int width = hi - lo;
    const int step = (width + 31) >> 5;
    int k = lo + (lane + 1) * step - 1;
    if (k > hi - 1) k = hi - 1;
    const bool pred = cu[k] <= x;
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    const int nt = __popc(m);  // predicate is monotone: the first nt lanes are true
    if (nt == 32) {
      lo = hi;  // last probe is hi-1 (clamped) and true
    } else {
      int kf = lo + (nt + 1) * step - 1;  // first false probe
      if (kf > hi - 1) kf = hi - 1;
      const int new_lo = nt ? min(lo + nt * step, hi) : lo;
      hi = kf;
      lo = new_lo;
    }
  }
  return lo - 1;
}

__global__ void __launch_bounds__(kScanWarps * 32, 6)
gae_scan_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                const int32_t* __restrict__ cu, const uint8_t* __restrict__ boot,
                float* __restrict__ adv, float* __restrict__ ret, int batch, int total,
                float gamma, float lam, uint64_t* __restrict__ state,
                unsigned int* __restrict__ ticket_ctr, int num_tiles) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  ScanSmem& S = reinterpret_cast<ScanSmem*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const float c = gamma * lam;
  const float lc = log2f(c);        // -inf when c == 0: powers below become exact zeros
  const float c2 = c * c, c3 = c2 * c, c4 = c2 * c2;
  (void)c4;

  while (true) {
    // a ticket is taken only when the warp is ready to work on it: a tile that looks back at its
    // right neighbour then finds that neighbour already in progress, not queued behind other work
    unsigned int ticket = 0;
    if (lane == 0) ticket = atomicAdd(ticket_ctr, 1u);
    ticket = __shfl_sync(0xffffffffu, ticket, 0);
    if (ticket >= (unsigned)num_tiles) break;
    const int g = num_tiles - 1 - (int)ticket;  // right-most tile first
    const int tlo = g * kTile;
    const int thi = min(tlo + kTile, total);
    const int tlen = thi - tlo;

    // ---- prologue: sequence ids and flags of the tile's tokens --------------------------
    const int i_lo = count_ends_le(cu, batch, tlo, lane);  // sequence that owns token tlo
#pragma unroll
    for (int k = 0; k < kTile / 128; k++)
      *reinterpret_cast<int4*>(&S.meta[k * 128 + lane * 4]) = make_int4(0, 0, 0, 0);
    __syncwarp();
    for (int base = i_lo;; base += 32) {
      const int i = base + lane;
      bool more = false;
      if (i < batch) {
        const int s = cu[i], e = cu[i + 1];
        if (i > i_lo && s < thi) atomicAdd(&S.meta[s - tlo], 1);  // a sequence starts here
        if (e > s && e - 1 >= tlo && e - 1 < thi)
          atomicOr(&S.meta[e - 1 - tlo], kMetaLast | (boot[i] ? kMetaBoot : 0));
        more = e < thi;  // sequences after this one may still start inside the tile
      }
      if (__ballot_sync(0xffffffffu, more) != 0xffffffffu) break;
    }
    __syncwarp();
    // inclusive prefix sum of the start counts -> relative sequence id per token
    {
      int run = 0;
#pragma unroll 1
      for (int k = 0; k < kTile / 128; k++) {
        int4 m = *reinterpret_cast<int4*>(&S.meta[k * 128 + lane * 4]);
        int c0 = m.x & kMetaCnt, c1 = m.y & kMetaCnt, c2_ = m.z & kMetaCnt, c3_ = m.w & kMetaCnt;
        int s0 = c0, s1 = s0 + c1, s2 = s1 + c2_, s3 = s2 + c3_;
        int inc = s3;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          int t = __shfl_up_sync(0xffffffffu, inc, o);
          if (lane >= o) inc += t;
        }
        const int ex = run + inc - s3;
        m.x = (m.x & ~kMetaCnt) | (ex + s0);
        m.y = (m.y & ~kMetaCnt) | (ex + s1);
        m.z = (m.z & ~kMetaCnt) | (ex + s2);
        m.w = (m.w & ~kMetaCnt) | (ex + s3);
        *reinterpret_cast<int4*>(&S.meta[k * 128 + lane * 4]) = m;
        run += __shfl_sync(0xffffffffu, inc, 31);
      }
    }
    __syncwarp();

    // ---- load (coalesced, lane-strided) and delta ---------------------------------------
    // 8 lane-strided rows at a time: 24 independent global loads in flight per lane
#pragma unroll 1
    for (int kb = 0; kb < kTile / 32; kb += 8) {
      float rr[8], vv[8], nn[8];
      int mm[8];
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const int o = (kb + j) * 32 + lane;
        const bool ok = o < tlen;
        mm[j] = ok ? S.meta[o] : 0;
        const int t = tlo + o;
        const int64_t vi = (int64_t)t + i_lo + (mm[j] & kMetaCnt);
        rr[j] = ok ? rewards[t] : 0.f;
        vv[j] = ok ? values[vi] : 0.f;
        nn[j] = ok ? values[vi + 1] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const int o = (kb + j) * 32 + lane;
        float nv = nn[j];
        if ((mm[j] & kMetaLast) && !(mm[j] & kMetaBoot)) nv = 0.f;
        S.d[o] = (o < tlen) ? fmaf(gamma, nv, rr[j]) - vv[j] : 0.f;
        if (o >= tlen) S.meta[o] = 0;
      }
    }
    __syncwarp();

    // ---- segmented reverse scan with zero carry-in at the tile's right edge --------------
    int last_boundary = -1;  // tile-local index of the right-most LAST token
    float X = 0.f;           // adv of the first token of the block to the right (carry)
#pragma unroll 1
    for (int blk = kTile / kBlk - 1; blk >= 0; blk--) {
      const int o = blk * kBlk + lane * 4;
      float4 dl = *reinterpret_cast<float4*>(&S.d[o]);
      int4 m = *reinterpret_cast<int4*>(&S.meta[o]);
      const float a0 = (m.x & kMetaLast) ? 0.f : c, a1 = (m.y & kMetaLast) ? 0.f : c,
                  a2 = (m.z & kMetaLast) ? 0.f : c, a3 = (m.w & kMetaLast) ? 0.f : c;
      // lane-local: L_k = adv with zero carry, P_k = prod a_k..a_3
      const float L3 = dl.w, P3 = a3;
      const float L2 = fmaf(a2, L3, dl.z), P2 = a2 * P3;
      const float L1 = fmaf(a1, L2, dl.y), P1 = a1 * P2;
      const float L0 = fmaf(a0, L1, dl.x), P0 = a0 * P1;
      // suffix (towards higher lanes) inclusive scan of (P0, L0)
      float Q = P0, Sx = L0;
#pragma unroll
      for (int sft = 1; sft < 32; sft <<= 1) {
        const float q2 = __shfl_down_sync(0xffffffffu, Q, sft);
        const float s2 = __shfl_down_sync(0xffffffffu, Sx, sft);
        if (lane + sft < 32) {
          Sx = fmaf(Q, s2, Sx);
          Q = Q * q2;
        }
      }
      // composite of the lanes to my right, applied to the block carry X
      float Qr = __shfl_down_sync(0xffffffffu, Q, 1);
      float Sr = __shfl_down_sync(0xffffffffu, Sx, 1);
      if (lane == 31) {
        Qr = 1.f;
        Sr = 0.f;
      }
      const float carry = fmaf(Qr, X, Sr);
      float4 A;
      A.w = fmaf(P3, carry, L3);
      A.z = fmaf(P2, carry, L2);
      A.y = fmaf(P1, carry, L1);
      A.x = fmaf(P0, carry, L0);
      *reinterpret_cast<float4*>(&S.d[o]) = A;
      X = __shfl_sync(0xffffffffu, A.x, 0);
      if (last_boundary < 0) {
        const unsigned lm = ((m.x & kMetaLast) ? 1u : 0u) | ((m.y & kMetaLast) ? 2u : 0u) |
                            ((m.z & kMetaLast) ? 4u : 0u) | ((m.w & kMetaLast) ? 8u : 0u);
        const unsigned any = __ballot_sync(0xffffffffu, lm != 0);
        if (any) {
          const int hl = 31 - __clz(any);  // highest lane with a boundary
          const unsigned hm = __shfl_sync(0xffffffffu, lm, hl);
          last_boundary = blk * kBlk + hl * 4 + (31 - __clz(hm));
        }
      }
    }
    const float first_zero_carry = X;  // adv of token tlo assuming zero carry-in

    // ---- publish / look back -------------------------------------------------------------
    const bool need_carry = thi < total && last_boundary != tlen - 1;
    const bool closed = last_boundary >= 0;  // first token's adv does not depend on the carry
    if (lane == 0 && g > 0) {
      if (closed || !need_carry) st_state(&state[g], kFlagIncl, first_zero_carry);
      else st_state(&state[g], kFlagAgg, first_zero_carry);
    }
    float Xin = 0.f;
    if (need_carry) {
      if (lane == 0) {
        const float cT = exp2f(lc * (float)kTile);
        float mult = 1.f;
        int j = g + 1;
        while (true) {
          uint64_t st;
          while (((st = ld_state(&state[j])) >> 32) == 0) __nanosleep(32);
          Xin = fmaf(mult, __uint_as_float((uint32_t)st), Xin);
          if ((uint32_t)(st >> 32) == kFlagIncl) break;
          mult *= cT;
          j++;
        }
        if (!closed && g > 0)
          st_state(&state[g], kFlagIncl, fmaf(exp2f(lc * (float)tlen), Xin, first_zero_carry));
      }
      Xin = __shfl_sync(0xffffffffu, Xin, 0);
    }

    // ---- write back (coalesced); tokens right of the last boundary take the carry -----------
    // v is read again for ret = adv + v: the tile was streamed microseconds ago, so this is an
    // L2 hit, and not keeping it in shared memory buys 50% more resident warps.
#pragma unroll 1
    for (int kb = 0; kb < kTile / 32; kb += 8) {
      float vv[8];
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const int o = (kb + j) * 32 + lane;
        const int64_t vi = (int64_t)tlo + o + i_lo + (S.meta[o] & kMetaCnt);
        vv[j] = (o < tlen) ? values[vi] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const int o = (kb + j) * 32 + lane;
        if (o < tlen) {
          float A = S.d[o];
          if (need_carry && o > last_boundary) A = fmaf(ex2_approx(lc * (float)(tlen - o)), Xin, A);
          adv[tlo + o] = A;
          ret[tlo + o] = A + vv[j];
        }
      }
    }
    __syncwarp();
    (void)c2; (void)c3;
  }
}

// ============================ SCAN, piece-wise ================================
// Same tiling, tickets and look-back protocol as gae_scan_kernel, but a tile is processed one
// SEQUENCE PIECE at a time (the intersection of a sequence with the tile) instead of token by token
// with per-token sequence ids: inside a piece the value offset is a constant and the scan is a plain
// geometric one (a = gamma*lam everywhere), so a lane owns 4 consecutive tokens, rewards / adv / ret
// move as 16-byte vectors straight between global memory and registers, and whole blocks inside a
// piece take a branch-free path with constant multipliers. ~4x fewer instructions per token than the
// per-token kernel, which was issue-bound (ncu: profiles/r01_ncu_gae.txt). Requires 16-byte aligned
// rewards / adv / ret (else the per-token kernel is used).
constexpr int kVwPad = 64;  // value slots staged beyond the tile: one per sequence start in the tile
struct __align__(16) PieceSmem {
  float r[kTile];            // the tile's rewards, staged by one TMA bulk copy; once a block's rewards
                             // are in registers the same slots take the zero-carry advantages of the
                             // piece that waits for a look-back carry
  float vw[kTile + kVwPad];  // the tile's window of `values`, staged by one TMA bulk copy
  uint64_t bar;              // mbarrier of the two copies
  uint64_t pad_;
};

struct PieceConst {
  float c, c2, c3, c4;
  float cpow[5];  // c^(4*2^k), k = 0..4: Kogge-Stone multipliers for lane distance 1,2,4,8,16
  float qlane;    // c^(4*(31-lane))
  float gamma, lc;
};

// Powers of c = gamma*lam, rounded once from double on the host: repeated fp32 squaring would put
// ~4e-6 of relative error on c^64, which at |adv| ~ 10 is already the whole 1e-5 parity budget.
struct PiecePowers {
  float c, c2, c3, c4, cT;
  float cpow[5];
  float qtab[32];  // c^(4*(31-lane))
};

// The inputs of one 128-token block of a piece: 4 rewards and 5 values per lane (the 5th is the
// next-value of the lane's last token), read from the tile's TMA-staged shared-memory copies.
// The values of a piece are contiguous but shifted by the sequence index, so they are misaligned
// against the 4-token lane groups by m = seq & 3 — the same for every lane: two aligned 16-byte
// shared loads and a static select. Anything outside the staged windows (last partial tile, tiles
// with more than kVwPad sequence starts) falls back to global loads.
struct BlkData {
  float r[4], v[5];
};

struct TileStage {
  int tlo, rcount, vw0, vcount, total, vend;
};

__device__ __forceinline__ BlkData load_block(const PieceSmem& S, const TileStage& G,
                                              const float* __restrict__ rewards,
                                              const float* __restrict__ values, int seq, int tb,
                                              int ps, int pe, int lane) {
  BlkData D;
  const int t0 = tb + lane * 4;
  const int o = t0 - G.tlo;
  if (tb - G.tlo + 128 <= G.rcount) {
    const float4 r4 = *reinterpret_cast<const float4*>(&S.r[o]);
    D.r[0] = r4.x; D.r[1] = r4.y; D.r[2] = r4.z; D.r[3] = r4.w;
  } else {
#pragma unroll
    for (int k = 0; k < 4; k++)
      D.r[k] = (o + k < G.rcount) ? S.r[o + k] : ((t0 + k < G.total) ? __ldg(rewards + t0 + k) : 0.f);
  }
  const int bi = t0 + seq - G.vw0;  // window index of the lane's first value
  const int m = bi & 3;             // = (seq - vw0) & 3, warp-uniform
  const int last = tb + 124 + seq - G.vw0 - m + 8;  // one past the last slot lane 31 touches
  if (last <= G.vcount) {
    const float4 w0 = *reinterpret_cast<const float4*>(&S.vw[bi - m]);
    const float4 w1 = *reinterpret_cast<const float4*>(&S.vw[bi - m + 4]);
    switch (m) {
      case 0: D.v[0] = w0.x; D.v[1] = w0.y; D.v[2] = w0.z; D.v[3] = w0.w; D.v[4] = w1.x; break;
      case 1: D.v[0] = w0.y; D.v[1] = w0.z; D.v[2] = w0.w; D.v[3] = w1.x; D.v[4] = w1.y; break;
      case 2: D.v[0] = w0.z; D.v[1] = w0.w; D.v[2] = w1.x; D.v[3] = w1.y; D.v[4] = w1.z; break;
      default: D.v[0] = w0.w; D.v[1] = w1.x; D.v[2] = w1.y; D.v[3] = w1.z; D.v[4] = w1.w; break;
    }
  } else {
#pragma unroll
    for (int k = 0; k < 5; k++) {
      const int t = t0 + k;
      D.v[k] = (t >= ps && t <= pe && t + seq < G.vend) ? __ldg(values + t + seq) : 0.f;
    }
  }
  return D;
}

// One 128-token block of a piece. [ps, pe): the piece; tb: first token of the block (multiple of
// 4); X: adv at token tb+128 (carry from the right); returns adv of the block's first token in X.
template <bool STASH>
__device__ __forceinline__ void piece_block(const PieceConst& K, const BlkData& D,
                                            float* __restrict__ adv, float* __restrict__ ret, int tb,
                                            int ps, int pe, int last_tok, bool boot, float& X,
                                            PieceSmem& S, int tlo, int lane) {
  const int t0 = tb + lane * 4;
  const bool full = tb >= ps && tb + 128 <= pe;
  const float* r = D.r;
  const float* v = D.v;
  float d[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    float nv = v[k + 1];
    if (t0 + k == last_tok && !boot) nv = 0.f;  // sequence ended without truncation: no bootstrap
    d[k] = fmaf(K.gamma, nv, r[k]) - v[k];
  }
  float A[4];
  if (!full) {
    // Ragged block: tokens outside the piece contribute delta = 0 and keep the multiplier c. That is
    // exact: to the right of the piece's end the carry is 0 by construction (a piece that ends
    // inside the tile starts from X = 0, and the piece that continues ends at the tile edge, which
    // is a block edge), and what is computed left of the piece's start is never used.
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int t = t0 + k;
      if (t < ps || t >= pe) d[k] = 0.f;
    }
  }
  {
    // uniform multipliers: lane composite (c^4, L0), suffix scan over lanes with constants
    const float L3 = d[3];
    const float L2 = fmaf(K.c, L3, d[2]);
    const float L1 = fmaf(K.c, L2, d[1]);
    const float L0 = fmaf(K.c, L1, d[0]);
    float Sx = L0;
#pragma unroll
    for (int j = 0; j < 5; j++) {
      const float s2 = __shfl_down_sync(0xffffffffu, Sx, 1 << j);
      if (lane + (1 << j) < 32) Sx = fmaf(K.cpow[j], s2, Sx);
    }
    float Sr = __shfl_down_sync(0xffffffffu, Sx, 1);
    if (lane == 31) Sr = 0.f;
    const float carry = fmaf(K.qlane, X, Sr);
    A[3] = fmaf(K.c, carry, L3);
    A[2] = fmaf(K.c2, carry, L2);
    A[1] = fmaf(K.c3, carry, L1);
    A[0] = fmaf(K.c4, carry, L0);
  }
  if (tb >= ps) {
    X = __shfl_sync(0xffffffffu, A[0], 0);
  } else {
    // the piece starts inside this block: hand back adv at its first token
    const int rel = ps - tb;
    const float sel = (rel & 3) == 0 ? A[0] : ((rel & 3) == 1 ? A[1] : ((rel & 3) == 2 ? A[2] : A[3]));
    X = __shfl_sync(0xffffffffu, sel, rel >> 2);
  }
  if (STASH) {
    // the piece's carry is not known yet: keep the zero-carry adv for the final sweep (its values
    // stay in the staged window)
    const int o = t0 - tlo;
    *reinterpret_cast<float4*>(&S.r[o]) = make_float4(A[0], A[1], A[2], A[3]);
  } else if (full) {
    *reinterpret_cast<float4*>(adv + t0) = make_float4(A[0], A[1], A[2], A[3]);
    *reinterpret_cast<float4*>(ret + t0) = make_float4(A[0] + v[0], A[1] + v[1], A[2] + v[2], A[3] + v[3]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int t = t0 + k;
      if (t >= ps && t < pe) {
        adv[t] = A[k];
        ret[t] = A[k] + v[k];
      }
    }
  }
}

// Pre-pass of the piece-wise kernel: for every tile, the sequence that owns its first token and that
// sequence's bounds and bootstrap flag. One thread per tile; all the dependent binary-search loads of
// all tiles overlap here instead of sitting at the head of every tile's critical path.
__global__ void __launch_bounds__(256)
gae_tile_index_kernel(const int32_t* __restrict__ cu, const uint8_t* __restrict__ boot, int batch,
                      int num_tiles, int4* __restrict__ table) {
  const int g = blockIdx.x * 256 + threadIdx.x;
  if (g >= num_tiles) return;
  const int x = g * kTile;
  int lo = 1, hi = batch + 1;  // #{k in [1,batch] : cu[k] <= x}
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (cu[mid] <= x) lo = mid + 1; else hi = mid;
  }
  const int i = lo - 1;
  table[g] = make_int4(i, cu[i], cu[i + 1], (int)boot[i]);
}

__global__ void __launch_bounds__(kScanWarps * 32, 6)
gae_pieces_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                  const int32_t* __restrict__ cu, const uint8_t* __restrict__ boot,
                  float* __restrict__ adv, float* __restrict__ ret, int batch, int total,
                  float gamma, const PiecePowers PW, uint64_t* __restrict__ state,
                  unsigned int* __restrict__ ticket_ctr, const int4* __restrict__ table,
                  int num_tiles) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PieceSmem& S = reinterpret_cast<PieceSmem*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  PieceConst K;
  K.gamma = gamma;
  K.c = PW.c;
  K.c2 = PW.c2;
  K.c3 = PW.c3;
  K.c4 = PW.c4;
  K.lc = log2f(K.c);
#pragma unroll
  for (int j = 0; j < 5; j++) K.cpow[j] = PW.cpow[j];
  K.qlane = PW.qtab[lane];
  if (lane == 0) {
    mbar_init(&S.bar, 1);
    fence_barrier_init();
  }
  __syncwarp();
  uint32_t bar_phase = 0;
  const uint64_t pol = policy_evict_first();

  // The ticket and the index-table entry of the NEXT tile are fetched while the current one is
  // being processed (every warp does the same, so neighbouring tiles still start at about the same
  // time and the look-back finds its right neighbour in progress).
  // {sequence owning the tile's first token, its start, its end, its bootstrap flag}: from the
  // pre-pass table for large inputs, searched here for small ones (where one more launch would
  // cost more than the search)
  auto tile_entry = [&](unsigned int tk) {
    const int gg = num_tiles - 1 - (int)tk;
    if (table) return table[gg];
    const int i = count_ends_le(cu, batch, gg * kTile, lane);
    return make_int4(i, cu[i], cu[i + 1], (int)boot[i]);
  };
  unsigned int ticket = 0;
  if (lane == 0) ticket = atomicAdd(ticket_ctr, 1u);
  ticket = __shfl_sync(0xffffffffu, ticket, 0);
  int4 te = make_int4(0, 0, 0, 0);
  if (ticket < (unsigned)num_tiles) te = tile_entry(ticket);
  while (ticket < (unsigned)num_tiles) {
    const int g = num_tiles - 1 - (int)ticket;
    const int tlo = g * kTile;
    const int thi = min(tlo + kTile, total);
    unsigned int nticket = 0;
    if (lane == 0) nticket = atomicAdd(ticket_ctr, 1u);
    const int i_lo = te.x;
    // stage the tile: its rewards and its window of values, one TMA bulk copy each
    TileStage G;
    G.tlo = tlo;
    G.total = total;
    G.vend = total + batch;
    G.rcount = (thi - tlo) & ~3;
    G.vw0 = (tlo + i_lo) & ~3;
    G.vcount = min(kTile + kVwPad, (G.vend - G.vw0) & ~3);
    __syncwarp();  // every lane is done reading the previous tile's staged data
    if (lane == 0 && G.rcount + G.vcount > 0) {
      mbar_expect_tx(&S.bar, (uint32_t)(G.rcount + G.vcount) * 4u);
      if (G.rcount > 0) bulk_g2s(S.r, rewards + tlo, (uint32_t)G.rcount * 4u, &S.bar, pol);
      if (G.vcount > 0) bulk_g2s(S.vw, values + G.vw0, (uint32_t)G.vcount * 4u, &S.bar, pol);
    }

    bool first = true;          // the piece that contains tlo has not been published yet
    int carry_ps = -1;          // first token of the piece that waits for the look-back carry
    int carry_seq = 0;          // its sequence
    float carry_first = 0.f;    // its zero-carry adv at carry_ps
    bool carry_is_first = false;

    auto do_piece = [&](int seq, int s, int e, bool b) {
      const int ps = max(s, tlo), pe = min(e, thi);
      const bool ends_here = e <= thi;
      const bool waits = !ends_here;  // continues into the tile on the right: needs its adv
      const int last_tok = ends_here ? e - 1 : -1;  // value of token t of sequence seq: values[t + seq]
      float X = 0.f;
      const int b_hi = (pe - 1 - tlo) >> 7, b_lo = (ps - tlo) >> 7;
#pragma unroll 1
      for (int blk = b_hi; blk >= b_lo; blk--) {
        const BlkData cur = load_block(S, G, rewards, values, seq, tlo + blk * 128, ps, pe, lane);
        if (waits) piece_block<true>(K, cur, adv, ret, tlo + blk * 128, ps, pe, last_tok, b, X, S, tlo, lane);
        else piece_block<false>(K, cur, adv, ret, tlo + blk * 128, ps, pe, last_tok, b, X, S, tlo, lane);
      }
      if (waits) {
        carry_ps = ps;
        carry_seq = seq;
        carry_first = X;
        carry_is_first = first;
      }
      if (first) {
        first = false;
        if (lane == 0 && g > 0) st_state(&state[g], waits ? kFlagAgg : kFlagIncl, X);
      }
    };

    // the sequences after the first one: their bounds are fetched while the first piece is scanned
    int base = i_lo + 1;
    int idx = base + lane;
    int cs = idx <= batch ? cu[idx] : 0x7fffffff;
    int ce = idx < batch ? cu[idx + 1] : 0x7fffffff;
    int bt = idx < batch ? (int)boot[idx] : 0;
    if (G.rcount + G.vcount > 0) {
      while (!mbar_try_wait(&S.bar, bar_phase)) {
      }
      bar_phase ^= 1;
    }
    nticket = __shfl_sync(0xffffffffu, nticket, 0);
    int4 nte = make_int4(0, 0, 0, 0);
    if (nticket < (unsigned)num_tiles) nte = tile_entry(nticket);
    do_piece(i_lo, te.y, te.z, te.w != 0);
    bool done = te.z >= thi;
    while (!done) {
#pragma unroll 1
      for (int k = 0; k < 32; k++) {
        const int seq = base + k;
        const int s_ = __shfl_sync(0xffffffffu, cs, k);
        const int e_ = __shfl_sync(0xffffffffu, ce, k);
        const bool b_ = __shfl_sync(0xffffffffu, bt, k) != 0;
        if (seq >= batch || s_ >= thi) {
          done = true;
          break;
        }
        if (e_ <= s_) continue;  // empty sequence
        do_piece(seq, s_, e_, b_);
      }
      if (!done) {
        base += 32;
        idx = base + lane;
        cs = idx <= batch ? cu[idx] : 0x7fffffff;
        ce = idx < batch ? cu[idx + 1] : 0x7fffffff;
        bt = idx < batch ? (int)boot[idx] : 0;
      }
    }

    if (carry_ps >= 0) {
      // ---- look back for adv at token thi, then finish the waiting piece -------------------
      float Xin = 0.f;
      if (lane == 0) {
        const float cT = PW.cT;
        float mult = 1.f;
        int j = g + 1;
        while (true) {
          uint64_t stv;
          while (((stv = ld_state(&state[j])) >> 32) == 0) __nanosleep(32);
          Xin = fmaf(mult, __uint_as_float((uint32_t)stv), Xin);
          if ((uint32_t)(stv >> 32) == kFlagIncl) break;
          mult *= cT;
          j++;
        }
        if (carry_is_first && g > 0)
          st_state(&state[g], kFlagIncl, fmaf(exp2f(K.lc * (float)(thi - carry_ps)), Xin, carry_first));
      }
      Xin = __shfl_sync(0xffffffffu, Xin, 0);
      __syncwarp();
      const int b_hi = (thi - 1 - tlo) >> 7, b_lo = (carry_ps - tlo) >> 7;
#pragma unroll 1
      for (int blk = b_hi; blk >= b_lo; blk--) {
        const int o = blk * 128 + lane * 4;
        const int t0 = tlo + o;
        const float4 a4 = *reinterpret_cast<const float4*>(&S.r[o]);
        const BlkData D = load_block(S, G, rewards, values, carry_seq, tlo + blk * 128, carry_ps, thi, lane);
        const float4 v4 = make_float4(D.v[0], D.v[1], D.v[2], D.v[3]);
        // multiplier of the carry for token t: c^(thi - t)
        const float m3 = ex2_approx(K.lc * (float)(thi - (t0 + 3)));
        const float m2 = m3 * K.c, m1 = m2 * K.c, m0 = m1 * K.c;
        const float A0 = fmaf(m0, Xin, a4.x), A1 = fmaf(m1, Xin, a4.y), A2 = fmaf(m2, Xin, a4.z),
                    A3 = fmaf(m3, Xin, a4.w);
        if (t0 >= carry_ps && t0 + 4 <= thi) {
          *reinterpret_cast<float4*>(adv + t0) = make_float4(A0, A1, A2, A3);
          *reinterpret_cast<float4*>(ret + t0) = make_float4(A0 + v4.x, A1 + v4.y, A2 + v4.z, A3 + v4.w);
        } else {
          const float Ak[4] = {A0, A1, A2, A3};
          const float vk[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
          for (int k = 0; k < 4; k++) {
            const int t = t0 + k;
            if (t >= carry_ps && t < thi) {
              adv[t] = Ak[k];
              ret[t] = Ak[k] + vk[k];
            }
          }
        }
      }
      __syncwarp();
    }
    ticket = nticket;
    te = nte;
  }
}

struct ScanWs {
  unsigned int* ticket;
  uint64_t* state;
  int num_tiles;
  size_t bytes;
};
ScanWs scan_ws(void* ws, int64_t total) {
  ScanWs W;
  W.num_tiles = (int)((total + kTile - 1) / kTile);
  W.ticket = (unsigned int*)ws;
  W.state = (uint64_t*)((char*)ws + 256);
  W.bytes = 256 + (size_t)(W.num_tiles + 1) * sizeof(uint64_t);
  W.bytes = (W.bytes + 255) & ~(size_t)255;
  return W;
}

}  // namespace
}  // namespace b200rl

extern "C" {

size_t b200rl_gae_workspace_bytes(int32_t batch, int64_t total) {
  (void)batch;
  if (total < 0) return 0;
  const b200rl::ScanWs W = b200rl::scan_ws(nullptr, total);
  return W.bytes + (size_t)W.num_tiles * sizeof(int4);  // look-back state + the tile index table
}

int b200rl_gae_1d_nolp_misalign(const float* rewards, const float* values,
                                const int32_t* cu_seqlens, const uint8_t* bootstrap, float* adv,
                                float* ret, int32_t batch, int64_t total, float gamma, float lam,
                                int mode, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(batch >= 0 && total >= 0, "negative size");
  B200RL_REQUIRE(total < (int64_t)INT32_MAX - batch, "total tokens must fit int32 (cu_seqlens is int32)");
  if (batch == 0 || total == 0) return B200RL_OK;
  B200RL_REQUIRE(rewards && values && cu_seqlens && bootstrap && adv && ret, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == B200RL_GAE_AUTO) {
    // SCAN re-associates fp32 additions; its deviation from the serial order stays below the
    // 1e-5 bar only while gamma*lam forgets fast enough (measured: tests/test_gae.py).
    const float c = gamma * lam;
    mode = (c > 0.99f || !(c >= 0.f)) ? B200RL_GAE_SERIAL : B200RL_GAE_SCAN;
  }
  // SERIAL picks among kernels that produce the same bits (measured on B200,
  // scripts/time_gae_modes.py -> profiles/r02_gae_modes.txt):
  //   relay   one warp per sequence, chain relayed lane to lane in registers: >= 1024 sequences of
  //           >= 64 tokens on average (65,536 x [64,2048]: 253 us vs 386 us warp / 360 us slots;
  //           8,192: 46 vs 64 us; 4,096 x [1k,16k]: 188 vs 266 us; 1,024: 18 vs 20 us);
  //   slots   8 independent chains per warp with ticket refill: >= 32k SHORT sequences (500k x
  //           [1,64]: 84 us vs 218 us warp / 144 us for the 32-per-warp kernel);
  //   batch8  8 sequences per warp in lock step: 16k-32k short sequences (< 64 tokens on average);
  //   team    two warps per sequence (feeder/writer + chain lane): at most 4 sequences per SM of
  //           >= 512 tokens on average, where the longest chain is the run time (120 x [1k,16k]:
  //           55 us vs 86 us warp; 512 x [1k,16k]: 70 vs 88 us; 1000 x [1k,16k]: 113 vs 105 us);
  //   warp    one warp per sequence, chain in lane 0: everything else.
  const bool short_seqs = total < (int64_t)batch * 64;
  const bool use_slots = mode == B200RL_GAE_SERIAL_SLOTS ||
                         (mode == B200RL_GAE_SERIAL && short_seqs && batch >= 32768);
  if (mode == B200RL_GAE_SERIAL && !short_seqs && batch >= B200RL_RELAY_MIN_BATCH) mode = B200RL_GAE_SERIAL_RELAY;
  if (mode == B200RL_GAE_SERIAL_BATCH32)
    return launch_serial_batched<32, 32, 2>(rewards, values, cu_seqlens, bootstrap, adv, ret, batch, gamma, lam, st);
  if (mode == B200RL_GAE_SERIAL_BATCH8 || (mode == B200RL_GAE_SERIAL && !use_slots && short_seqs && batch >= 16384))
    return launch_serial_batched<8, 128, 2>(rewards, values, cu_seqlens, bootstrap, adv, ret, batch, gamma, lam, st);
  if (use_slots) {
    if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(gae_serial_slots_kernel), (int)kSlotSmem))
      return rc;
    unsigned int* ticket = nullptr;  // with a workspace: sequences are handed out by a ticket counter
    if (workspace && workspace_bytes >= sizeof(unsigned int) &&
        (reinterpret_cast<uintptr_t>(workspace) & 3) == 0) {
      ticket = static_cast<unsigned int*>(workspace);
      B200RL_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st));
    }
    int grid = (batch + kSlotWarps * kSlots - 1) / (kSlotWarps * kSlots);
    const int cap = sm_count() * (int)(220 * 1024 / (kSlotSmem + 1024));
    if (grid > cap) grid = cap;
    gae_serial_slots_kernel<<<grid, kSlotWarps * 32, kSlotSmem, st>>>(
        rewards, values, cu_seqlens, bootstrap, adv, ret, batch, gamma, lam, ticket);
    return check_launch("gae_serial_slots_kernel");
  }
  // team: few long sequences, where the longest chain IS the run time (120 x [1k,16k]: see
  // profiles/r02_gae_modes.txt)
  if (mode == B200RL_GAE_SERIAL && batch <= sm_count() * 4 && total >= (int64_t)batch * 2 * kSerBlk)
    mode = B200RL_GAE_SERIAL_TEAM;
  if (mode == B200RL_GAE_SERIAL_TEAM) {
    int grid = batch;
    const int cap = sm_count() * 8;
    if (grid > cap) grid = cap;
    gae_serial_team_kernel<<<grid, 64, 0, st>>>(rewards, values, cu_seqlens, bootstrap, adv, ret, batch,
                                               gamma, lam);
    return check_launch("gae_serial_team_kernel");
  }
  if (mode == B200RL_GAE_SERIAL_RELAY) {
    int grid = (batch + kSerWarps - 1) / kSerWarps;
    const int cap = sm_count() * 16;
    if (grid > cap) grid = cap;
    gae_serial_relay_kernel<<<grid, kSerWarps * 32, 0, st>>>(rewards, values, cu_seqlens, bootstrap, adv,
                                                            ret, batch, gamma, lam);
    return check_launch("gae_serial_relay_kernel");
  }
  if (mode == B200RL_GAE_SERIAL || mode == B200RL_GAE_SERIAL_WARP) {
    int grid = (batch + kSerWarps - 1) / kSerWarps;
    const int cap = sm_count() * 16;
    if (grid > cap) grid = cap;
    gae_serial_kernel<<<grid, kSerWarps * 32, 0, st>>>(rewards, values, cu_seqlens, bootstrap, adv,
                                                      ret, batch, gamma, lam);
    return check_launch("gae_serial_kernel");
  }
  if (mode == B200RL_GAE_SCAN) {
    B200RL_REQUIRE(workspace != nullptr, "SCAN mode needs a workspace");
    B200RL_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "workspace must be 256-B aligned");
    ScanWs W = scan_ws(workspace, total);
    if (workspace_bytes < W.bytes + (size_t)W.num_tiles * sizeof(int4)) {
      set_last_error("gae workspace too small: %zu < %zu", workspace_bytes,
                     W.bytes + (size_t)W.num_tiles * sizeof(int4));
      return B200RL_EWORKSPACE;
    }
    B200RL_CUDA(cudaMemsetAsync(workspace, 0, W.bytes, st));
    const size_t smem = sizeof(ScanSmem) * kScanWarps;
    const size_t smem_p = sizeof(PieceSmem) * kScanWarps;
    if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(gae_scan_kernel), (int)smem)) return rc;
    if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(gae_pieces_kernel), (int)smem_p)) return rc;
    int grid = (W.num_tiles + kScanWarps - 1) / kScanWarps;
    const int cap = sm_count() * 6;  // 6 CTAs x 4 warps x 8 KiB = 192 KiB smem per SM
    if (grid > cap) grid = cap;
    // The piece-wise kernel wants 16-byte aligned rewards/adv/ret and pays a fixed cost per
    // sequence piece, so batches of very short sequences (< 128 tokens on average) stay on the
    // per-token kernel (measured: 500k sequences of 1-64 tokens, 103 us vs 232 us).
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(rewards) | reinterpret_cast<uintptr_t>(adv) |
                          reinterpret_cast<uintptr_t>(ret) | reinterpret_cast<uintptr_t>(values)) & 15) == 0;
    if (vec_ok && total >= (int64_t)batch * 128) {
      PiecePowers PW;
      const double c = (double)gamma * (double)lam;  // the kernels use fl32(gamma*lam) as c
      const double cf = (double)(gamma * lam);
      (void)c;
      PW.c = (float)cf;
      PW.c2 = (float)(cf * cf);
      PW.c3 = (float)(cf * cf * cf);
      PW.c4 = (float)(cf * cf * cf * cf);
      PW.cT = (float)pow(cf, (double)kTile);
      for (int j = 0; j < 5; j++) PW.cpow[j] = (float)pow(cf, 4.0 * (double)(1 << j));
      for (int l = 0; l < 32; l++) PW.qtab[l] = (float)pow(cf, 4.0 * (double)(31 - l));
      int4* table = nullptr;
      if (W.num_tiles >= 4096) {  // below that the in-kernel search is cheaper than another launch
        table = reinterpret_cast<int4*>((char*)workspace + W.bytes);
        gae_tile_index_kernel<<<(W.num_tiles + 255) / 256, 256, 0, st>>>(cu_seqlens, bootstrap, batch,
                                                                        W.num_tiles, table);
      }
      int grid_p = (W.num_tiles + kScanWarps - 1) / kScanWarps;
      if (grid_p > sm_count() * 6) grid_p = sm_count() * 6;  // 6 CTAs x 4 warps x 8.3 KiB per SM
      gae_pieces_kernel<<<grid_p, kScanWarps * 32, smem_p, st>>>(rewards, values, cu_seqlens, bootstrap,
                                                          adv, ret, batch, (int)total, gamma, PW,
                                                          W.state, W.ticket, table, W.num_tiles);
      return check_launch("gae_pieces_kernel");
    }
    gae_scan_kernel<<<grid, kScanWarps * 32, smem, st>>>(rewards, values, cu_seqlens, bootstrap,
                                                        adv, ret, batch, (int)total, gamma, lam,
                                                        W.state, W.ticket, W.num_tiles);
    return check_launch("gae_scan_kernel");
  }
  set_last_error("unknown GAE mode %d", mode);
  return B200RL_EINVAL;
}

}  // extern "C"

// =============================================================================================
// Allocation-free variant of the tiled scan.
//
// Who needs it: a process that shares its GPU with an inference server which has already
// reserved nearly all of HBM for its KV cache (the usual AReaL colocated layout), a plain C / FFI
// host that has no device allocator of its own to hand us a workspace from, or any caller that
// runs at a point where device allocation is forbidden or may fail. For those callers the look-back
// state of gae_scan_kernel lives in the library image instead of in a caller buffer: every GPU that
// loads libb200rl.so carries this array (about 33 KB) until the process exits.
//
// The array below is named from HOST code only to ask the runtime where it lives
// (cudaGetSymbolAddress); the name itself is not a device pointer and must never be passed to a
// kernel. The lookup happens once per device, on that device, the first time the entry point is
// called there.
// =============================================================================================
namespace b200rl {
namespace {
constexpr int64_t kNoAllocMaxTokens = 4194304;  // 4 Mi packed tokens = 4096 tiles
constexpr size_t kNoAllocBytes = 256 + ((size_t)(kNoAllocMaxTokens / kTile) + 1) * sizeof(uint64_t);
constexpr size_t kNoAllocBytesPadded = (kNoAllocBytes + 255) & ~(size_t)255;
}  // namespace
}  // namespace b200rl

static __device__ __align__(256) unsigned char g_gae_noalloc_scratch[b200rl::kNoAllocBytesPadded];

namespace b200rl {
namespace {
// Device address of the scratch array on the CURRENT device (resolved and cached on first use).
int gae_noalloc_scratch(void** out) {
  static void* resolved[64] = {nullptr};
  int dev = 0;
  B200RL_CUDA(cudaGetDevice(&dev));
  B200RL_REQUIRE(dev >= 0 && dev < 64, "device ordinal %d out of range", dev);
  if (resolved[dev] == nullptr) {
    void* p = nullptr;
    B200RL_CUDA(cudaGetSymbolAddress((void**)&p, g_gae_noalloc_scratch));
    B200RL_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    B200RL_CUDA(cudaFuncSetAttribute(gae_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(sizeof(ScanSmem) * kScanWarps)));
    resolved[dev] = p;
  }
  *out = resolved[dev];
  return B200RL_OK;
}
}  // namespace
}  // namespace b200rl

extern "C" int b200rl_gae_1d_nolp_misalign_noalloc(const float* rewards, const float* values,
                                                   const int32_t* cu_seqlens,
                                                   const uint8_t* bootstrap, float* adv, float* ret,
                                                   int32_t batch, int64_t total, float gamma,
                                                   float lam, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(batch >= 0 && total >= 0, "negative size");
  if (total > kNoAllocMaxTokens) {
    set_last_error(
        "b200rl_gae_1d_nolp_misalign_noalloc: %lld tokens exceed the built-in scratch, which covers "
        "at most %lld; call b200rl_gae_1d_nolp_misalign with a caller-provided workspace instead",
        (long long)total, (long long)kNoAllocMaxTokens);
    return B200RL_EINVAL;
  }
  if (batch == 0 || total == 0) return B200RL_OK;
  B200RL_REQUIRE(rewards && values && cu_seqlens && bootstrap && adv && ret, "null pointer");
  void* scratch = nullptr;
  int rc = gae_noalloc_scratch(&scratch);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  ScanWs W = scan_ws(scratch, total);
  // the array keeps whatever the previous call left: ticket counter and tile states must be zero
  B200RL_CUDA(cudaMemsetAsync(scratch, 0, W.bytes, st));
  int grid = (W.num_tiles + kScanWarps - 1) / kScanWarps;
  const int cap = sm_count() * 6;
  if (grid > cap) grid = cap;
  gae_scan_kernel<<<grid, kScanWarps * 32, sizeof(ScanSmem) * kScanWarps, st>>>(
      rewards, values, cu_seqlens, bootstrap, adv, ret, batch, (int)total, gamma, lam, W.state,
      W.ticket, W.num_tiles);
  return check_launch("gae_scan_kernel (noalloc)");
}

