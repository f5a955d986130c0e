// Fused GRPO actor step: logits -> log-prob, entropy, row max/min, clipped decoupled-PPO loss with
// all its statistics, AND dlogits, in one kernel and one trip of the logits through HBM.
//
// The reference (and the drop-in path of this library) does forward and backward as separate
// passes: 2*V bytes read, then 2*V read again + 2*V written = 6*V bytes per token. But the PPO
// loss is separable per token once the number of valid tokens N is known, so the gradient of the
// loss w.r.t. a token's log-prob is available the moment that token's row statistics are. Here a
// CTA therefore streams a row from HBM for the forward statistics, finishes the token's loss
// math, and immediately streams the SAME row again for the backward — now out of the 126 MB L2
// (148 rows in flight are 45 MB) — writing dlogits. HBM traffic drops to 4*V bytes per token.
//
// What makes this legal is knowing the upstream gradient of the loss before backward() runs:
// AReaL's engine multiplies the loss by a `loss_scale` that does not depend on the loss
// (areal/engine/fsdp_engine.py:543-549), so it can be handed to this kernel up front; the
// autograd node's backward then only checks that the incoming gradient is 1 and rescales
// dlogits in place if it is not (b200rl_scale_inplace: no traffic when the factor is 1).
//
// Bound: HBM, 4*V bytes per token (bf16); L2 sees 6*V.
#include "logprob_common.cuh"
#include "ppo_math.cuh"

#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace b200rl {
namespace {

using namespace lp;

#ifndef B200RL_FUSED_WARPS
#define B200RL_FUSED_WARPS 24
#endif
#ifndef B200RL_FUSED_STAGES
#define B200RL_FUSED_STAGES 2
#endif
// Skewed order (backward of row r after the forward of row r+1) hides the wait for the row's
// finalisation but keeps two rows per SM live: 90 MB plus the dlogits stream no longer fits the
// L2 (measured: DRAM reads 1055 MB instead of 626 MB per 2048 rows, 6.8 M instead of 9.5 M tok/s).
#ifndef B200RL_FUSED_SKEW
#define B200RL_FUSED_SKEW 0
#endif
#ifndef B200RL_FUSED_PAIR_DEFAULT
#define B200RL_FUSED_PAIR_DEFAULT 1
#endif
#ifndef B200RL_FUSED_PAIR_DELAY_DEFAULT
#define B200RL_FUSED_PAIR_DELAY_DEFAULT 8000
#endif
constexpr int kFusedWarps = B200RL_FUSED_WARPS, kFusedStages = B200RL_FUSED_STAGES;
constexpr bool kFusedSkew = B200RL_FUSED_SKEW != 0;

// G = rows a CTA works on at once: 1 (all warps on one row) or 2 (pair mode, below)
template <int WARPS, int STAGES, int G = 1>
struct FusedSmem {
  static constexpr int kData = WARPS * STAGES * kChunkBytes;
  static constexpr int kBarOff = kData;
  static constexpr int kPartOff = kBarOff + WARPS * STAGES * 8;       // [G][2][WARPS][8] float (G=1: [2][5][WARPS])
  static constexpr int kTokOff = kPartOff + G * 2 * 8 * WARPS * 4;    // [G][2][8] float: xy,prox,old,adv,mask
  static constexpr int kBwdOff = kTokOff + G * 2 * 8 * 4;             // [G][2][4]: nlse2, A0, lab_add, label
  static constexpr int kSyncOff = kBwdOff + G * 2 * 4 * 4;            // [G] x (cnt[2], ready[2]) int
  static constexpr int kRdyOff = (kSyncOff + G * 4 * 4 + 7) & ~7;     // pair mode: [G][2] "constants ready"
  static constexpr int kArrOff = kRdyOff + G * 2 * 8;                 //            [G][2] "row statistics in"
  static constexpr int kSumOff = kArrOff + G * 2 * 8;                 // [G] x (double sum, double count)
  static constexpr int kBytes = kSumOff + G * 16;
  static_assert(kPartOff % 16 == 0 && kBwdOff % 16 == 0, "st.async.v4 targets");
};
static_assert(FusedSmem<kFusedWarps, kFusedStages, 2>::kBytes <= 227 * 1024, "fused ring exceeds shared memory");
static_assert(kFusedWarps % 2 == 0, "pair mode splits the warps in two groups");

struct FusedParams {
  const void* logits;
  int64_t ld;
  const int64_t* labels;
  int T, V;
  float scale, inv_temp;
  const float *prox, *old, *adv;
  const uint8_t* mask;     // bytes (torch.bool) ...
  const float* mask_f32;   // ... or fp32 (non-zero = valid) when mask == nullptr
  int roll;                // labels = input_ids, label of row t = labels[(t+1) % T]
  float clip_lo, clip_hi, c_clip, cap;
  const float* denom;     // N = max(count(mask), 1), device
  const float* upstream;  // loss scale known in advance, device, may be null (= 1)
  float *logp, *ent, *lse, *rowmax, *rowmin;
  float *s_loss, *s_iw, *s_kl;
  uint8_t *s_clip, *s_dual;
  float *s_bw, *s_bkl;
  uint8_t* s_bm;
  // the derived tensors grpo_loss_fn logs (actor.py:339-365), all nullable
  uint8_t* s_valid;
  float *s_clip_ratio, *s_dual_ratio, *s_clip_new, *s_clip_old;
  void* dlogits;
  int64_t ld_out;
  double* partial;        // [grid][2]
  unsigned int* counter;
  float* loss;            // sum(mask * pg * w) / N  (not multiplied by upstream)
  int pair_delay_ns;      // pair mode: head start of row stream 0
};

__global__ void __launch_bounds__(1024)
mask_count_kernel(const uint8_t* __restrict__ mask, const float* __restrict__ mask_f32, int64_t T,
                  float* __restrict__ denom, int or_one) {
  __shared__ int red[32];
  int c = 0;
  if (mask) {
    for (int64_t i = threadIdx.x; i < T; i += 1024) c += mask[i] != 0;
  } else {
    for (int64_t i = threadIdx.x; i < T; i += 1024) c += mask_f32[i] != 0.f;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 32; w++) t += red[w];
    // PPO: `count_nonzero() or 1` (functional.py:143). SFT: plain count, 0 gives nan (lm_engine.py:53)
    denom[0] = (t > 0 || !or_one) ? (float)t : 1.f;
  }
}

// PAIR: the kernel runs as clusters of two CTAs (two SMs) that share TWO row streams. Warps
// 0..WARPS/2-1 of both CTAs work on the rows of stream 0, the other halves on stream 1, so a row
// is still cut in WARPS slices and a cluster still has one row per SM in flight (the L2 footprint
// of the re-read is unchanged) - but while one stream sits in the gap between its two passes
// (the wait for the slowest slice plus the token's loss math, 11 % of the warp time in the
// one-row-per-CTA form, profiles/r02_ncu_fused.txt) the warps of the other stream keep the MUFU
// and issue pipes of BOTH SMs busy. All traffic between the CTAs is st.async into distributed
// shared memory completing on an mbarrier of the receiving CTA (no cluster-scope fence: the
// release/acquire form of the same protocol cost a MEMBAR per warp and row, which has to wait for
// the warp's dlogits stores): each warp sends its slice statistics to CTA 0, where a rotating
// "finisher" warp of the stream waits for all of them, does the token's loss and sends the four
// backward constants to both CTAs.
template <int DT, int WARPS, int STAGES, bool MINMAX, bool SKEW, bool PAIR>
__global__ void __launch_bounds__(WARPS * 32, 1) grpo_fused_kernel(const FusedParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int G = PAIR ? 2 : 1;
  using L = FusedSmem<WARPS, STAGES, G>;
  using E = Elem<DT>;
  constexpr uint32_t kNeutral = E::kNeutralWord;
  constexpr bool ENT = true;
  static_assert(!(PAIR && SKEW), "pair mode has its own overlap");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cr = PAIR ? cluster_ctarank() : 0u;         // CTA within the cluster
  const int g = PAIR ? warp / (WARPS / 2) : 0;                // row stream of this warp
  const int seg = PAIR ? (int)cr * (WARPS / 2) + warp % (WARPS / 2) : warp;  // slice of the row
  const int first_row = PAIR ? (int)(blockIdx.x - cr) + g : (int)blockIdx.x;
  float* part = reinterpret_cast<float*>(smem + L::kPartOff) + g * 2 * 8 * WARPS;
  float* tok = reinterpret_cast<float*>(smem + L::kTokOff) + g * 2 * 8;
  float* bwdc = reinterpret_cast<float*>(smem + L::kBwdOff) + g * 2 * 4;
  int* cnt = reinterpret_cast<int*>(smem + L::kSyncOff) + g * 4;
  volatile int* ready = cnt + 2;
  uint64_t* rdy = reinterpret_cast<uint64_t*>(smem + L::kRdyOff) + g * 2;
  uint64_t* arr = reinterpret_cast<uint64_t*>(smem + L::kArrOff) + g * 2;
  double* cta_acc = reinterpret_cast<double*>(smem + L::kSumOff) + g * 2;
  if (threadIdx.x < 4 * G) reinterpret_cast<int*>(smem + L::kSyncOff)[threadIdx.x] = 0;
  if (threadIdx.x == 0) {
    double* acc0 = reinterpret_cast<double*>(smem + L::kSumOff);
#pragma unroll
    for (int i = 0; i < 2 * G; i++) acc0[i] = 0.0;
    if (PAIR) {
      uint64_t* r0 = reinterpret_cast<uint64_t*>(smem + L::kRdyOff);
      uint64_t* a0 = reinterpret_cast<uint64_t*>(smem + L::kArrOff);
#pragma unroll
      for (int i = 0; i < 2 * G; i++) {
        mbar_init(&r0[i], 1);          // the finisher's arrive.expect_tx
        mbar_init(&a0[i], WARPS / 2);  // the stream's warps of CTA 0 (CTA 1's only send bytes)
      }
      fence_barrier_init();
    }
  }
  __syncthreads();
  if (PAIR) cluster_sync_all();  // nobody touches the other CTA's shared memory before it is set up

  const int T = p.T, V = p.V;
  const float scale = p.scale;
  const float up_over_n = (p.upstream ? p.upstream[0] : 1.f) / p.denom[0];
  const int nrows = first_row < T ? (T - first_row + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  WarpStream<WARPS, STAGES, 2, SKEW> ws;
  ws.init_seg(p.logits, p.ld * E::kBytes, V / E::kPerVec, nrows, smem + warp * STAGES * kChunkBytes,
              reinterpret_cast<uint64_t*>(smem + L::kBarOff) + warp * STAGES, seg, WARPS, first_row,
              (int)gridDim.x, lane);
  if (PAIR && g == 1 && p.pair_delay_ns > 0) {
    // start the second stream out of phase with the first one, so that their gaps do not coincide
    uint64_t t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do {
      __nanosleep(256);
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    } while (t1 - t0 < (uint64_t)p.pair_delay_ns);
  }

  // pass 1 of row-iteration `it` (forward statistics + the token's loss, row from HBM)
  auto pass1 = [&](int it) {
      const int row = first_row + it * gridDim.x;
      const int slot = it & 1, use = it >> 1;
      // one (rotating) warp fetches the token's scalars early
      // (pair mode: the row's finisher, a warp of CTA 0, which then keeps them in registers)
      const bool finisher = PAIR && cr == 0 && (it % (WARPS / 2)) == seg;
      const bool fetch = (PAIR ? finisher : (it % WARPS) == seg) && lane == 0;
      float xy = 0.f, t_prox = 0.f, t_old = 0.f, t_adv = 0.f, t_mask = 0.f;
      int t_lab = -1;
      if (fetch) {
        const int64_t lab = p.labels[label_row(row, T, p.roll)];
        const bool ok = lab >= 0 && lab < V;
        xy = ok ? E::load1(p.logits, (int64_t)row * p.ld + lab) : __int_as_float(0x7fc00000);
        t_lab = ok ? (int)lab : -1;
        t_prox = p.prox[row];
        t_old = p.old[row];
        t_adv = p.adv[row];
        t_mask = (p.mask ? p.mask[row] != 0 : p.mask_f32[row] != 0.f) ? 1.f : 0.f;
      }

      // ---------------- pass 1: forward statistics (row from HBM) ----------------
      LaneState<E> S{kSentinel, 0.f, 0.f, 0.f, 0.f, E::max_init(), E::min_init()};
      int left = ws.seg_len;
      while (left > 0) {
        const int nv = min(kChunkVec, left);
        left -= nv;
        const uint32_t base = ws.wait() + lane * 16;
        if (nv == kChunkVec) {
          uint4 v[8];
#pragma unroll
          for (int i = 0; i < 8; i++) v[i] = lds128(base + i * 512);
          ws.release(lane);
          typename E::MaxT cm = E::vmax(v[0]);
#pragma unroll
          for (int i = 1; i < 8; i++) cm = E::cmax(cm, E::vmax(v[i]));
          S.pmx = E::cmax(S.pmx, cm);
          if (MINMAX) {
#pragma unroll
            for (int i = 0; i < 8; i++) S.pmn = E::cmin(S.pmn, E::vmin(v[i]));
          }
          lazy_rescale<E, ENT>(S, cm, scale);
#pragma unroll
          for (int i = 0; i < 8; i++) accumulate_vec<E, ENT>(S, v[i], scale);
        } else {
          const int nr = (nv + 31) >> 5;
#pragma unroll 1
          for (int r = 0; r < nr; r++) {
            const bool act = r * 32 + lane < nv;
            uint4 v = make_uint4(kNeutral, kNeutral, kNeutral, kNeutral);
            if (act) v = lds128(base + r * 512);
            const typename E::MaxT cm = E::vmax(v);
            S.pmx = E::cmax(S.pmx, cm);
            if (MINMAX && act) S.pmn = E::cmin(S.pmn, E::vmin(v));
            lazy_rescale<E, ENT>(S, cm, scale);
            accumulate_vec<E, ENT>(S, v, scale);
          }
          ws.release(lane);
        }
      }
      {
        const float m = -S.negm;
        const float s = S.s0 + S.s1, t = S.t0 + S.t1;
        const float M = warp_max(m);
        const float dl = m - M;
        const float f = ex2_approx(dl);
        const float Ssum = warp_sum(s * f);
        const float Tt = warp_sum(f * fmaf(dl, s, t));
        float mx = 0.f, mn = 0.f;
        if (MINMAX) {
          mx = warp_max(E::max_f(S.pmx));
          mn = warp_min(E::min_f(S.pmn));
        }
        float* q = part + slot * (PAIR ? 8 : 5) * WARPS;
        int old = 0;
        bool finish;
        float Mw = -kSentinel, Sw = 0.f, Tw = 0.f, mxw = -INFINITY, mnw = INFINITY;
        if (PAIR) {
          // slice statistics -> CTA 0's part[slot][seg][0..4], counted on its arr[slot]
          if (lane == 0) {
            const uint32_t q_c = dsmem_addr(smem_u32(q + seg * 8), 0);
            const uint32_t arr_c = dsmem_addr(smem_u32(&arr[slot]), 0);
            st_async_v4(q_c, M, Ssum, Tt, mx, arr_c);
            st_async_b32(q_c + 16, mn, arr_c);
            if (finisher) mbar_expect_tx(&arr[slot], WARPS * 20);
            else if (cr == 0) mbar_arrive(&arr[slot]);
          }
          finish = finisher;
          if (finish) {
            for (uint32_t polls = 0; !mbar_try_wait(&arr[slot], (uint32_t)use & 1u);)
              if (++polls > (1u << 24)) __trap();
            if (lane < WARPS) {
              const volatile float* vq = q + lane * 8;
              Mw = vq[0], Sw = vq[1], Tw = vq[2], mxw = vq[3], mnw = vq[4];
            }
          }
        } else {
          if (lane == 0) {
            q[0 * WARPS + warp] = M;
            q[1 * WARPS + warp] = Ssum;
            q[2 * WARPS + warp] = Tt;
            q[3 * WARPS + warp] = mx;
            q[4 * WARPS + warp] = mn;
            if (fetch) {
              float* tk = tok + slot * 8;
              tk[0] = xy;
              tk[1] = t_prox;
              tk[2] = t_old;
              tk[3] = t_adv;
              tk[4] = t_mask;
              tk[5] = __int_as_float(t_lab);
            }
            __threadfence_block();
            old = atomicAdd(&cnt[slot], 1);
          }
          old = __shfl_sync(0xffffffffu, old, 0);
          finish = old == WARPS - 1;
          if (finish) {
            __threadfence_block();
            const volatile float* vq = q;
            if (lane < WARPS) {
              Mw = vq[0 * WARPS + lane];
              Sw = vq[1 * WARPS + lane];
              Tw = vq[2 * WARPS + lane];
              mxw = vq[3 * WARPS + lane];
              mnw = vq[4 * WARPS + lane];
            }
          }
        }
        if (finish) {
          // ---- last warp to arrive / the finisher: row statistics, the token's loss, the
          // backward constants ----
          const float Mr = warp_max(Mw);
          const float dr = Mw - Mr;
          const float fr = ex2_approx(dr);
          const float Sr = warp_sum(Sw * fr);
          const float Tr = warp_sum(fr * fmaf(dr, Sw, Tw));
          float mxr = 0.f, mnr = 0.f;
          if (MINMAX) {
            mxr = warp_max(mxw);
            mnr = warp_min(mnw);
          }
          if (lane == 0) {
            float xyr, px, ol, A, mk;
            int lab;
            if (PAIR) {
              xyr = xy, px = t_prox, ol = t_old, A = t_adv, mk = t_mask, lab = t_lab;
            } else {
              cnt[slot] = 0;
              const volatile float* tk = tok + slot * 8;
              xyr = tk[0], px = tk[1], ol = tk[2], A = tk[3], mk = tk[4];
              lab = __float_as_int(tk[5]);
            }
            const bool m = mk != 0.f;
            const float l2s = log2f(Sr);
            const float lse_nat = (Mr + l2s) * kLn2;
            const float lpv = (fmaf(xyr, scale, -Mr) - l2s) * kLn2;
            const PpoToken k = ppo_token(lpv, px, ol, A, m, p.clip_lo, p.clip_hi, p.c_clip, p.cap);
            // The other warps of the row are waiting for these four numbers: publish them first,
            // write the row's outputs to global memory afterwards (off their critical path).
            // same operation order as ppo_loss_bwd_kernel + logprob_bwd_kernel (bit-identical dlogits)
            const float glp = k.dlogp * up_over_n;
            float* bc = bwdc + slot * 4;
            if (PAIR) {
#pragma unroll
              for (uint32_t r = 0; r < 2; r++) {
                const uint32_t rdy_c = dsmem_addr(smem_u32(&rdy[slot]), r);
                mbar_arrive_expect_tx_dsmem(rdy_c, 16);
                st_async_v4(dsmem_addr(smem_u32(bc), r), -lse_nat * kLog2e, -glp * p.inv_temp,
                            glp * p.inv_temp, __int_as_float(lab), rdy_c);
              }
            } else {
              bc[0] = -lse_nat * kLog2e;
              bc[1] = -glp * p.inv_temp;
              bc[2] = glp * p.inv_temp;
              bc[3] = __int_as_float(lab);
              __threadfence_block();
              ready[slot] = use + 1;
            }
            p.lse[row] = lse_nat;
            p.logp[row] = lpv;
            if (p.ent) p.ent[row] = kLn2 * (l2s - Tr / Sr);
            if (MINMAX) {
              if (p.rowmax) p.rowmax[row] = mxr;
              if (p.rowmin) p.rowmin[row] = mnr;
            }
            if (p.s_loss) p.s_loss[row] = k.pgw;
            if (p.s_iw) p.s_iw[row] = k.ratio;
            if (p.s_kl) p.s_kl[row] = k.kl;
            if (p.s_clip) p.s_clip[row] = k.clip;
            if (p.s_dual) p.s_dual[row] = k.dual;
            if (p.s_bw) p.s_bw[row] = k.w;
            if (p.s_bkl) p.s_bkl[row] = k.bkl;
            if (p.s_bm) p.s_bm[row] = k.bmask;
            if (p.s_valid) p.s_valid[row] = m;
            if (p.s_clip_ratio) p.s_clip_ratio[row] = k.clip ? 1.f : 0.f;
            if (p.s_dual_ratio) p.s_dual_ratio[row] = k.dual ? 1.f : 0.f;
            if (p.s_clip_new) p.s_clip_new[row] = k.clip ? lpv : 0.f;
            if (p.s_clip_old) p.s_clip_old[row] = k.clip ? ol : 0.f;
            if (m) {  // rows of a stream are finalised strictly in order: a deterministic sum
              // (pair mode: the finisher is always a warp of CTA 0, and the next row's finisher
              // passes arr[] only after this warp's own release-arrive for that row)
              cta_acc[0] += (double)k.pgw;
              cta_acc[1] += 1.0;
            }
          }
        }
      }

  };
  // pass 2 of row-iteration `it` (backward, row from L2)
  auto pass2 = [&](int it) {
    const int row = first_row + it * gridDim.x;
    const int slot = it & 1, use = it >> 1;
      // ---------------- pass 2: backward (row again, from L2) ----------------
      if (PAIR) {
        // (a wait of microseconds; the bound turns a lost arrival into an error, not a hung GPU)
        for (uint32_t polls = 0; !mbar_try_wait(&rdy[slot], (uint32_t)use & 1u);)
          if (++polls > (1u << 24)) __trap();
      } else {
        if (lane == 0) {
          while (ready[slot] < use + 1) {
          }
        }
        __syncwarp();
        __threadfence_block();
      }
      const volatile float* bc = bwdc + slot * 4;
      const float nlse2 = bc[0], A0 = bc[1], lab_add = bc[2];
      const int lab = __float_as_int(bc[3]);
      float nlse2f;
      uint32_t sgn;
      bwd_row_consts<E>(nlse2, A0, nlse2f, sgn);
      const int lvec = lab >= 0 ? lab / E::kPerVec : -1;
      const int lsub = lab >= 0 ? lab % E::kPerVec : 0;
      char* out_row = reinterpret_cast<char*>(p.dlogits) + ((int64_t)row * p.ld_out) * E::kBytes;
      int left = ws.seg_len;
      int v0 = ws.sv + lane;
      while (left > 0) {
        const int nv = min(kChunkVec, left);
        left -= nv;
        const uint32_t base = ws.wait() + lane * 16;
        if (nv == kChunkVec) {
          uint4 v[8];
#pragma unroll
          for (int i = 0; i < 8; i++) v[i] = lds128(base + i * 512);
          ws.release(lane);
#pragma unroll
          for (int i = 0; i < 8; i++)
            bwd_vec_folded<E>(v[i], v0 + i * 32, true, scale, nlse2f, sgn, lab_add, lvec, lsub, out_row);
        } else {
          const int nr = (nv + 31) >> 5;
#pragma unroll 1
          for (int r = 0; r < nr; r++) {
            const bool act = r * 32 + lane < nv;
            uint4 v = make_uint4(kNeutral, kNeutral, kNeutral, kNeutral);
            if (act) v = lds128(base + r * 512);
            bwd_vec_folded<E>(v, v0 + r * 32, act, scale, nlse2f, sgn, lab_add, lvec, lsub, out_row);
          }
          ws.release(lane);
        }
        v0 += kChunkVec;
      }
  };
  if (SKEW) {
    // A0, A1, B0, A2, B1, ...: by the time a warp turns to the backward of row r it has streamed
    // row r+1, so the last-arriving warp of row r has long published the backward constants
    for (int it = 0; it < nrows; ++it) {
      pass1(it);
      if (it >= 1) pass2(it - 1);
    }
    if (nrows > 0) pass2(nrows - 1);
  } else {
    for (int it = 0; it < nrows; ++it) {
      pass1(it);
      pass2(it);
    }
  }

  // ---- loss: per-stream partial, the last CTA adds them in index order -----------------------
  // (partial b holds the rows b, b + grid, ... in both modes: the same sums in the same order)
  __syncthreads();
  if (PAIR) cluster_sync_all();  // every remote update of CTA 0's sums has landed; CTA 1 is done
  __shared__ bool is_last;
  if (PAIR && cr != 0) return;
  if (threadIdx.x == 0) {
    const double* acc0 = reinterpret_cast<const double*>(smem + L::kSumOff);
#pragma unroll
    for (int i = 0; i < 2 * G; i++) p.partial[2 * blockIdx.x + i] = acc0[i];
    __threadfence();
    is_last = atomicAdd(p.counter, (unsigned)G) == gridDim.x - G;
  }
  __syncthreads();
  if (is_last && threadIdx.x == 0) {
    __threadfence();
    const volatile double* part_g = p.partial;
    double s = 0.0;
    for (unsigned b = 0; b < gridDim.x; b++) s += part_g[2 * b];
    p.loss[0] = (float)s / p.denom[0];
  }
}

__global__ void __launch_bounds__(256)
scale_inplace_kernel(uint4* __restrict__ buf, int64_t nvec, int dtype, const float* __restrict__ g) {
  const float s = g[0];
  if (s == 1.f) return;  // the common case: nothing is read or written
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * 256) {
    uint4 v = buf[i];
    if (dtype == B200RL_FP32) {
      float x[4];
      Elem<B200RL_FP32>::unpack(v, x);
#pragma unroll
      for (int j = 0; j < 4; j++) x[j] *= s;
      buf[i] = Elem<B200RL_FP32>::pack(x);
    } else if (dtype == B200RL_BF16) {
      float x[8];
      Elem<B200RL_BF16>::unpack(v, x);
#pragma unroll
      for (int j = 0; j < 8; j++) x[j] *= s;
      buf[i] = Elem<B200RL_BF16>::pack(x);
    } else {
      float x[8];
      Elem<B200RL_FP16>::unpack(v, x);
#pragma unroll
      for (int j = 0; j < 8; j++) x[j] *= s;
      buf[i] = Elem<B200RL_FP16>::pack(x);
    }
  }
}

struct FusedWs {
  unsigned int* counter;
  double* partial;
  size_t bytes;
};
FusedWs fused_ws(void* ws) {
  FusedWs W;
  W.counter = (unsigned int*)ws;
  W.partial = (double*)((char*)ws + 256);
  W.bytes = 256 + 2 * sizeof(double) * 1024;  // up to 1024 CTAs
  return W;
}

template <typename K>
int set_smem_once(K kernel, int bytes) {
  return ensure_dynamic_smem(reinterpret_cast<const void*>(kernel), bytes);
}

// B200RL_FUSED_PAIR=0 selects the one-row-per-CTA form; B200RL_FUSED_PAIR_DELAY_NS the head
// start of the first row stream (both read once).
struct PairConfig {
  bool on;
  int delay_ns;
};
const PairConfig& pair_config() {
  static const PairConfig c = [] {
    PairConfig r{B200RL_FUSED_PAIR_DEFAULT != 0, B200RL_FUSED_PAIR_DELAY_DEFAULT};
    if (const char* e = getenv("B200RL_FUSED_PAIR")) r.on = atoi(e) != 0;
    if (const char* e = getenv("B200RL_FUSED_PAIR_DELAY_NS")) r.delay_ns = atoi(e);
    return r;
  }();
  return c;
}

// clusters of two CTAs that can be resident at once on this device (cached per device)
template <typename K>
int max_pair_clusters(K kernel, int smem) {
  static std::mutex mu;
  static int cached[64];
  static bool have[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  std::lock_guard<std::mutex> lk(mu);
  if (!have[dev]) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * sm_count());
    cfg.blockDim = dim3(kFusedWarps * 32);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess) {
      cudaGetLastError();
      n = 0;
    }
    cached[dev] = n;
    have[dev] = true;
    if (getenv("B200RL_FUSED_PAIR_VERBOSE")) fprintf(stderr, "b200rl: fused pair mode: %d clusters of 2 fit (device %d, %d SMs)\n", n, dev, sm_count());
  }
  return cached[dev];
}

template <int DT, bool MINMAX>
int launch_fused(FusedParams p, cudaStream_t st) {
  const PairConfig& pc = pair_config();
  if (pc.on && !kFusedSkew && p.T >= 2) {
    auto k = grpo_fused_kernel<DT, kFusedWarps, kFusedStages, MINMAX, false, true>;
    constexpr int smem = FusedSmem<kFusedWarps, kFusedStages, 2>::kBytes;
    int rc = set_smem_once(k, smem);
    if (rc) return rc;
    const int clusters = min(max_pair_clusters(k, smem), min((p.T + 1) / 2, 512));
    if (clusters > 0) {
      p.pair_delay_ns = pc.delay_ns;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(2 * clusters);
      cfg.blockDim = dim3(kFusedWarps * 32);
      cfg.dynamicSmemBytes = smem;
      cfg.stream = st;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2;
      at[0].val.clusterDim.y = 1;
      at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      if (cudaLaunchKernelEx(&cfg, k, p) == cudaSuccess) return check_launch("grpo_fused_kernel (pair)");
      // A cluster launch this context refuses is not an error of the call: the one-row-per-CTA
      // form below computes the same bits. (A sticky error from earlier work comes back from that
      // launch as well, so nothing is hidden.)
      cudaGetLastError();
    }
  }
  auto k = grpo_fused_kernel<DT, kFusedWarps, kFusedStages, MINMAX, kFusedSkew, false>;
  constexpr int smem = FusedSmem<kFusedWarps, kFusedStages, 1>::kBytes;
  int rc = set_smem_once(k, smem);
  if (rc) return rc;
  const int grid = min(p.T, min(sm_count(), 1024));
  p.pair_delay_ns = 0;
  k<<<grid, kFusedWarps * 32, smem, st>>>(p);
  return check_launch("grpo_fused_kernel");
}

}  // namespace
}  // namespace b200rl

extern "C" {

size_t b200rl_grpo_fused_workspace_bytes(int64_t T) {
  (void)T;
  return b200rl::fused_ws(nullptr).bytes;
}

int b200rl_grpo_fused_step(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                           int64_t T, int64_t V, float temperature, const float* prox_logp,
                           const float* old_logp, const float* adv, const uint8_t* mask,
                           float clip_lo, float clip_hi, float c_clip, float cap,
                           const float* upstream, float* logp, float* entropy, float* lse,
                           float* rowmax, float* rowmin, float* stat_loss, float* stat_imp_weight,
                           float* stat_approx_kl, uint8_t* stat_clip_mask,
                           uint8_t* stat_dual_clip_mask, float* stat_behave_imp_weight,
                           float* stat_behave_approx_kl, uint8_t* stat_behave_mask, void* dlogits,
                           int64_t ld_out, float* loss, float* denom, void* workspace,
                           size_t workspace_bytes, void* stream) {
  return b200rl_grpo_fused_step_ids(logits, dtype, ld, labels, 0, T, V, temperature, prox_logp, old_logp, adv,
                                    mask, 0, clip_lo, clip_hi, c_clip, cap, upstream, logp, entropy, lse,
                                    rowmax, rowmin, stat_loss, stat_imp_weight, stat_approx_kl, stat_clip_mask,
                                    stat_dual_clip_mask, stat_behave_imp_weight, stat_behave_approx_kl,
                                    stat_behave_mask, nullptr, nullptr, nullptr, nullptr, nullptr, dlogits,
                                    ld_out, loss, denom, workspace, workspace_bytes, stream);
}

int b200rl_grpo_fused_step_ids(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                               int label_roll, int64_t T, int64_t V, float temperature,
                               const float* prox_logp, const float* old_logp, const float* adv,
                               const void* mask_any, int mask_is_f32, float clip_lo, float clip_hi,
                               float c_clip, float cap, const float* upstream, float* logp,
                               float* entropy, float* lse, float* rowmax, float* rowmin,
                               float* stat_loss, float* stat_imp_weight, float* stat_approx_kl,
                               uint8_t* stat_clip_mask, uint8_t* stat_dual_clip_mask,
                               float* stat_behave_imp_weight, float* stat_behave_approx_kl,
                               uint8_t* stat_behave_mask, uint8_t* stat_valid, float* stat_clip_ratio,
                               float* stat_dual_clip_ratio, float* stat_clipped_new_logp,
                               float* stat_clipped_old_logp, void* dlogits, int64_t ld_out, float* loss,
                               float* denom, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace b200rl;
  const uint8_t* mask = mask_is_f32 ? nullptr : static_cast<const uint8_t*>(mask_any);
  const float* mask_f32 = mask_is_f32 ? static_cast<const float*>(mask_any) : nullptr;
  B200RL_REQUIRE(T >= 0 && V > 0, "bad shape T=%lld V=%lld", (long long)T, (long long)V);
  B200RL_REQUIRE(ld >= V && ld_out >= V, "row stride < V");
  B200RL_REQUIRE(temperature > 0.f, "temperature must be > 0, got %f", temperature);
  B200RL_REQUIRE(T < INT32_MAX && V < INT32_MAX, "T and V must fit int32");
  B200RL_REQUIRE(loss && denom && workspace, "loss/denom/workspace must not be null");
  B200RL_REQUIRE(!(c_clip > 0.f) || c_clip > 1.f, "c_clip must be > 1 (functional.py:157), got %f", c_clip);
  B200RL_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "workspace must be 256-B aligned");
  FusedWs W = fused_ws(workspace);
  if (workspace_bytes < W.bytes) {
    set_last_error("fused workspace too small: %zu < %zu", workspace_bytes, W.bytes);
    return B200RL_EWORKSPACE;
  }
  B200RL_REQUIRE(dtype == B200RL_BF16 || dtype == B200RL_FP16 || dtype == B200RL_FP32,
                 "unsupported logits dtype code %d", dtype);
  const int pv = dtype == B200RL_FP32 ? 4 : 8;
  const bool fast = (V % pv == 0) && (ld % pv == 0) && (ld_out % pv == 0) &&
                    (reinterpret_cast<uintptr_t>(logits) % 16 == 0) &&
                    (reinterpret_cast<uintptr_t>(dlogits) % 16 == 0);
  if (!fast) {
    set_last_error("fused step needs 16-byte aligned rows (V %% %d == 0); use the separate fwd/bwd entry points", pv);
    return B200RL_EUNSUPPORTED;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200RL_CUDA(cudaMemsetAsync(workspace, 0, 256, st));
  if (T == 0) {
    B200RL_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), st));
    return B200RL_OK;
  }
  B200RL_REQUIRE(logits && labels && prox_logp && old_logp && adv && mask_any && logp && lse && dlogits,
                 "null pointer");
  mask_count_kernel<<<1, 1024, 0, st>>>(mask, mask_f32, T, denom, cap < -2.5f ? 0 : 1);
  FusedParams p;
  p.logits = logits; p.ld = ld; p.labels = labels; p.T = (int)T; p.V = (int)V;
  p.scale = (float)(1.4426950408889634 / (double)temperature);
  p.inv_temp = (float)(1.0 / (double)temperature);
  p.prox = prox_logp; p.old = old_logp; p.adv = adv; p.mask = mask; p.mask_f32 = mask_f32;
  p.roll = label_roll != 0;
  p.s_valid = stat_valid; p.s_clip_ratio = stat_clip_ratio; p.s_dual_ratio = stat_dual_clip_ratio;
  p.s_clip_new = stat_clipped_new_logp; p.s_clip_old = stat_clipped_old_logp;
  p.clip_lo = clip_lo; p.clip_hi = clip_hi; p.c_clip = c_clip; p.cap = cap;
  p.denom = denom; p.upstream = upstream;
  p.logp = logp; p.ent = entropy; p.lse = lse; p.rowmax = rowmax; p.rowmin = rowmin;
  p.s_loss = stat_loss; p.s_iw = stat_imp_weight; p.s_kl = stat_approx_kl;
  p.s_clip = stat_clip_mask; p.s_dual = stat_dual_clip_mask;
  p.s_bw = stat_behave_imp_weight; p.s_bkl = stat_behave_approx_kl; p.s_bm = stat_behave_mask;
  p.dlogits = dlogits; p.ld_out = ld_out;
  p.partial = W.partial; p.counter = W.counter; p.loss = loss;
  const bool mm = rowmax != nullptr || rowmin != nullptr;
  switch (dtype) {
    case B200RL_BF16: return mm ? launch_fused<B200RL_BF16, true>(p, st) : launch_fused<B200RL_BF16, false>(p, st);
    case B200RL_FP16: return mm ? launch_fused<B200RL_FP16, true>(p, st) : launch_fused<B200RL_FP16, false>(p, st);
    default: return mm ? launch_fused<B200RL_FP32, true>(p, st) : launch_fused<B200RL_FP32, false>(p, st);
  }
}

int b200rl_scale_inplace(void* buf, int dtype, int64_t numel, const float* scale, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(numel >= 0, "negative numel");
  if (numel == 0) return B200RL_OK;
  B200RL_REQUIRE(buf && scale, "null pointer");
  B200RL_REQUIRE(dtype == B200RL_BF16 || dtype == B200RL_FP16 || dtype == B200RL_FP32,
                 "unsupported dtype code %d", dtype);
  const int pv = dtype == B200RL_FP32 ? 4 : 8;
  B200RL_REQUIRE(numel % pv == 0 && (reinterpret_cast<uintptr_t>(buf) & 15) == 0,
                 "buffer must be 16-byte aligned and a whole number of 16-byte vectors");
  const int64_t nvec = numel / pv;
  int64_t blocks = (nvec + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  scale_inplace_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((uint4*)buf, nvec, dtype, scale);
  return check_launch("scale_inplace_kernel");
}

}  // extern "C"
