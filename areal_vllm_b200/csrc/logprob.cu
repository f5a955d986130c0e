// H1 — streaming log-prob / entropy / gather over [T, V] logits, forward and fused backward.
//
// Replaces the torch.compile'd log_softmax + gather (+ exp*mul*sum for the entropy) of
// areal/utils/functional.py:10-80 and the autograd graph behind it, which saves an fp32
// [chunk, V] log-prob tensor for the backward. Here:
//
//  forward   one pass over each bf16 row (2*V bytes/token): every warp of a persistent CTA
//            streams its own slice of the row through a private ring of TMA bulk copies
//            (cp.async.bulk global->shared, mbarrier complete_tx) and keeps, per lane, an
//            online softmax state (m, s = sum 2^(y-m), t = sum 2^(y-m)*(y-m)) in the log2 domain,
//            y = x*log2(e)/temperature: exactly one MUFU.EX2 per element, the reference max
//            m only ever moves when an element exceeds it by more than 2^8 (lazy rescale), so
//            the common path has no dependent max chain. Per row: warp shuffles, one
//            __syncthreads, and warp 0 emits logp, entropy, lse (and the row's raw max/min).
//  backward  dlogits = (g_logp*(onehot - p) - g_ent*p*(log p + H)) / temperature, with p rebuilt
//            from the saved lse: one read of the row + one bf16 write (4*V bytes/token). The
//            softmax is never materialised. Shaped like a copy kernel (one CTA per 8 KiB tile,
//            register-staged 16-byte loads and stores): measured 13 % faster than streaming it
//            through the forward's TMA ring.
//
// Bound: HBM. Algorithmic bytes per token: fwd 2*V, bwd 4*V (bf16 logits).
// The same kernels are instantiated for fp16 and fp32 logits (fp32: 4 elements per 16-byte vector,
// 4*V / 8*V bytes per token). Vocabularies whose rows are not 16-byte aligned take a plain
// (non-TMA) kernel further down: correct, not tuned.
#include "logprob_common.cuh"

namespace b200rl {
namespace {

using namespace lp;

// floats per row of the vocabulary-shard partial record: (M, S, W, x_y, max, min, 0, 0)
constexpr int kPartialStride = 8;

// ------------------------------------------------------------------------------------------
// forward, rows 16-B aligned (bf16 / fp16 / fp32)
// ------------------------------------------------------------------------------------------
template <int DT, int WARPS, int STAGES, bool ENT, bool MINMAX>
__global__ void __launch_bounds__(WARPS * 32, 1)
logprob_fwd_kernel(const void* __restrict__ logits, int64_t ld,
                        const int64_t* __restrict__ labels, int T, int V, float scale,
                        float* __restrict__ logp, float* __restrict__ ent,
                        float* __restrict__ lse, float* __restrict__ rowmax,
                        float* __restrict__ rowmin, float* __restrict__ partial, int roll) {
  extern __shared__ __align__(128) unsigned char smem[];
  using L = FwdSmem<WARPS, STAGES>;
  using E = Elem<DT>;
  constexpr uint32_t kNeutral = E::kNeutralWord;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* part = reinterpret_cast<float*>(smem + L::kPartOff);   // [kSlots][6][WARPS]
  int* cnt = reinterpret_cast<int*>(smem + L::kCntOff);         // arrivals per slot
  volatile int* gen = cnt + kSlots;                             // completed uses per slot
  if (threadIdx.x < 2 * kSlots) cnt[threadIdx.x] = 0;
  __syncthreads();

  const int nrows = (T - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  WarpStream<WARPS, STAGES> ws;
  ws.init(logits, ld * E::kBytes, V / E::kPerVec, nrows, smem + warp * STAGES * kChunkBytes,
          reinterpret_cast<uint64_t*>(smem + L::kBarOff) + warp * STAGES, warp, lane);

  for (int it = 0; it < nrows; ++it) {
    const int row = blockIdx.x + it * gridDim.x;
    // the label logit is a single-element gather; one (rotating) warp fetches it early
    float xy = 0.f;
    const bool fetch_xy = (it % WARPS) == warp && lane == 0;
    if (fetch_xy) {
      const int64_t lab = labels[label_row(row, T, roll)];
      xy = (lab >= 0 && lab < V) ? E::load1(logits, (int64_t)row * ld + lab)
                                 : __int_as_float(0x7fc00000);
    }
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
        // tail of the slice: 1..7 warp-rows, only the vocabulary's last one can be ragged
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

    // ---- combine the 32 lanes of this warp -------------------------------------------------
    const float m = -S.negm;
    const float s = S.s0 + S.s1, t = S.t0 + S.t1;
    const float M = warp_max(m);
    const float dl = m - M;  // <= 0
    const float f = ex2_approx(dl);
    const float Ssum = warp_sum(s * f);
    float Tt = 0.f;
    if (ENT) Tt = warp_sum(f * fmaf(dl, s, t));
    float mx = 0.f, mn = 0.f;
    if (MINMAX) {
      mx = warp_max(E::max_f(S.pmx));
      mn = warp_min(E::min_f(S.pmn));
    }
    // ---- publish; the last warp to arrive finalises the row (no CTA-wide barrier) -----------
    const int slot = it & (kSlots - 1);
    const int use = it / kSlots;
    float* p = part + slot * 6 * WARPS;
    int old = 0;
    if (lane == 0) {
      while (gen[slot] < use) {  // slot still being read for row it - kSlots (rare)
      }
      p[0 * WARPS + warp] = M;
      p[1 * WARPS + warp] = Ssum;
      p[2 * WARPS + warp] = Tt;
      p[3 * WARPS + warp] = mx;
      p[4 * WARPS + warp] = mn;
      if (fetch_xy) p[5 * WARPS] = xy;
      __threadfence_block();
      old = atomicAdd(&cnt[slot], 1);
    }
    old = __shfl_sync(0xffffffffu, old, 0);
    if (old == WARPS - 1) {
      __threadfence_block();
      const volatile float* q = p;
      const bool has = lane < WARPS;
      const float Mw = has ? q[0 * WARPS + lane] : -kSentinel;
      const float Sw = has ? q[1 * WARPS + lane] : 0.f;
      const float Tw = has ? q[2 * WARPS + lane] : 0.f;
      const float mxw = has ? q[3 * WARPS + lane] : -INFINITY;
      const float mnw = has ? q[4 * WARPS + lane] : INFINITY;
      const float xyr = q[5 * WARPS];
      const float Mr = warp_max(Mw);
      const float dr = Mw - Mr;
      const float fr = ex2_approx(dr);
      const float Sr = warp_sum(Sw * fr);
      float Tr = 0.f;
      if (ENT) Tr = warp_sum(fr * fmaf(dr, Sw, Tw));
      float mxr = 0.f, mnr = 0.f;
      if (MINMAX) {
        mxr = warp_max(mxw);
        mnr = warp_min(mnw);
      }
      if (lane == 0) {
        cnt[slot] = 0;
        if (partial) {
          // vocabulary-shard mode: raw online-softmax state of this shard's slice of the row
          float4* pr = reinterpret_cast<float4*>(partial + (int64_t)row * kPartialStride);
          pr[0] = make_float4(Mr, Sr, Tr, xyr);
          pr[1] = make_float4(mxr, mnr, 0.f, 0.f);
        } else {
          const float l2s = log2f(Sr);
          lse[row] = (Mr + l2s) * kLn2;
          // (x_y*scale - M) first: exact cancellation when the label is the row maximum, however
          // large |M| is; only then the (small) log2 of the sum
          logp[row] = (fmaf(xyr, scale, -Mr) - l2s) * kLn2;
          if (ENT && ent) ent[row] = kLn2 * (l2s - Tr / Sr);
          if (MINMAX) {
            if (rowmax) rowmax[row] = mxr;
            if (rowmin) rowmin[row] = mnr;
          }
        }
        __threadfence_block();
        gen[slot] = use + 1;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// backward, rows 16-B aligned on input and output (bf16 / fp16 / fp32)
//
// No shared-memory ring here. Round 1 streamed the backward through the same per-warp TMA ring as
// the forward (persistent grid of one CTA per SM); that structure tops out at 6.0-6.2 TB/s for
// this read+write kernel - and does so even with the arithmetic compiled out - while torch's
// copy kernel reaches 6.5-6.6 TB/s on the same box and traffic mix. The shape that wins is the
// copy kernel's: one small CTA per 8 KiB tile of a row, every thread issues its 16-byte loads up
// front, folds the gradient in registers and stores; the CTAs sweep memory in address order.
// Measured (scripts/time_h1.py, T = 10,240, V = 151,936, bf16): 892-904 us = 6.9-7.0 TB/s against
// 1012 us = 6.15 TB/s for the ring version (tiles of 8 / 16 / 32 KiB: 892 / 904 / 903 us).
// ------------------------------------------------------------------------------------------
#ifndef B200RL_BWD_TILE_U
#define B200RL_BWD_TILE_U 2
#endif
constexpr int kBwdThreads = 256;
constexpr int kBwdU = B200RL_BWD_TILE_U;            // 16-byte vectors per thread
constexpr int kBwdTileVec = kBwdThreads * kBwdU;    // vectors per CTA (8 KiB)

template <int DT, bool HAS_GH>
__global__ void __launch_bounds__(kBwdThreads)
logprob_bwd_kernel(const void* __restrict__ logits, int64_t ld, const int64_t* __restrict__ labels,
                   const float* __restrict__ lse, const float* __restrict__ entropy,
                   const float* __restrict__ g_logp, const float* __restrict__ g_ent, int T, int V,
                   float scale, float inv_temp, void* __restrict__ dlogits, int64_t ld_out, int roll,
                   const float* __restrict__ g_num, const float* __restrict__ g_den, int tiles_per_row) {
  using E = Elem<DT>;
  const int row = blockIdx.x / tiles_per_row;
  const int tile = blockIdx.x - row * tiles_per_row;
  const int nvec = V / E::kPerVec;
  const char* in_row = reinterpret_cast<const char*>(logits) + ((int64_t)row * ld) * E::kBytes;
  char* out_row = reinterpret_cast<char*>(dlogits) + ((int64_t)row * ld_out) * E::kBytes;
  const int v0 = tile * kBwdTileVec + threadIdx.x;
  uint4 v[kBwdU];
#pragma unroll
  for (int u = 0; u < kBwdU; u++) {
    const int vi = v0 + u * kBwdThreads;
    v[u] = make_uint4(E::kNeutralWord, E::kNeutralWord, E::kNeutralWord, E::kNeutralWord);
    if (vi < nvec) v[u] = ldg_nc_v4(in_row + (int64_t)vi * 16);
  }
  // per-row constants (the same for every thread of the CTA: broadcast loads).
  // g_logp scaled by g_num[0] / g_den[0] when given: the scalar loss's backward
  // (ppo_loss_bwd_kernel's `dlogp[t] * (grad_loss / N)`, same operation order) folded in
  const float glp = g_num ? g_logp[row] * (g_num[0] / g_den[0]) : g_logp[row];
  const float nlse2 = -lse[row] * kLog2e;
  float gh = 0.f, H = 0.f;
  if (HAS_GH) {
    gh = g_ent[row];
    H = entropy[row];
  }
  const float A0 = -(glp + gh * H) * inv_temp;
  const float B0 = -gh * kLn2 * inv_temp;
  const float lab_add = glp * inv_temp;
  const int64_t lab = labels[label_row(row, T, roll)];
  const bool ok = lab >= 0 && lab < V;
  const int lvec = ok ? (int)(lab / E::kPerVec) : -1;
  const int lsub = (int)(lab % E::kPerVec);
  float nlse2f = 0.f;
  uint32_t sgn = 0u;
  if (!HAS_GH) bwd_row_consts<E>(nlse2, A0, nlse2f, sgn);
#pragma unroll
  for (int u = 0; u < kBwdU; u++) {
    const int vi = v0 + u * kBwdThreads;
    if (HAS_GH)
      bwd_vec<E, true>(v[u], vi, vi < nvec, scale, nlse2, A0, B0, lab_add, lvec, lsub, out_row);
    else
      bwd_vec_folded<E>(v[u], vi, vi < nvec, scale, nlse2f, sgn, lab_add, lvec, lsub, out_row);
  }
}

// ------------------------------------------------------------------------------------------
// plain kernels: any of bf16 / fp16 / fp32, any V, any stride. One CTA per row, two passes.
// ------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }

constexpr int kPlainThreads = 512;

__device__ __forceinline__ float block_reduce(float v, float* sh, int op /*0 sum,1 max,2 min*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = op == 0 ? warp_sum(v) : (op == 1 ? warp_max(v) : warp_min(v));
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float r = sh[0];
  for (int w = 1; w < kPlainThreads / 32; w++)
    r = op == 0 ? r + sh[w] : (op == 1 ? fmaxf(r, sh[w]) : fminf(r, sh[w]));
  return r;
}

template <typename T>
__global__ void __launch_bounds__(kPlainThreads)
logprob_fwd_plain_kernel(const T* __restrict__ logits, int64_t ld, const int64_t* __restrict__ labels,
                         int64_t V, float scale, float* __restrict__ logp, float* __restrict__ ent,
                         float* __restrict__ lse, float* __restrict__ rowmax,
                         float* __restrict__ rowmin, float* __restrict__ partial, int roll) {
  __shared__ float sh[kPlainThreads / 32];
  const int64_t row = blockIdx.x;
  const T* x = logits + row * ld;
  float mx = -INFINITY, mn = INFINITY;
  for (int64_t j = threadIdx.x; j < V; j += kPlainThreads) {
    const float f = to_f<T>(x[j]);
    mx = fmaxf(mx, f);
    mn = fminf(mn, f);
  }
  mx = block_reduce(mx, sh, 1);
  mn = block_reduce(mn, sh, 2);
  const float m2 = mx * scale;
  float s = 0.f, t = 0.f;
  for (int64_t j = threadIdx.x; j < V; j += kPlainThreads) {
    const float y = fmaf(to_f<T>(x[j]), scale, -m2);
    const float e = exp2f(y);
    s += e;
    t = fmaf(e, y, t);
  }
  s = block_reduce(s, sh, 0);
  t = block_reduce(t, sh, 0);
  if (threadIdx.x == 0) {
    const float l2s = log2f(s);
    const float lse2 = m2 + l2s;
    const int64_t lab = labels[label_row((int)row, (int)gridDim.x, roll)];
    const float xy = (lab >= 0 && lab < V) ? to_f<T>(x[lab]) : __int_as_float(0x7fc00000);
    if (partial) {
      float4* pr = reinterpret_cast<float4*>(partial + row * kPartialStride);
      pr[0] = make_float4(m2, s, t, xy);
      pr[1] = make_float4(mx, mn, 0.f, 0.f);
      return;
    }
    lse[row] = lse2 * kLn2;
    logp[row] = (fmaf(xy, scale, -m2) - l2s) * kLn2;
    if (ent) ent[row] = kLn2 * (l2s - t / s);
    if (rowmax) rowmax[row] = mx;
    if (rowmin) rowmin[row] = mn;
  }
}

template <typename T>
__global__ void __launch_bounds__(kPlainThreads)
logprob_bwd_plain_kernel(const T* __restrict__ logits, int64_t ld, const int64_t* __restrict__ labels,
                         const float* __restrict__ lse, const float* __restrict__ entropy,
                         const float* __restrict__ g_logp, const float* __restrict__ g_ent,
                         int64_t V, float scale, float inv_temp, T* __restrict__ dlogits,
                         int64_t ld_out, int roll, const float* __restrict__ g_num,
                         const float* __restrict__ g_den) {
  const int64_t row = blockIdx.x;
  const T* x = logits + row * ld;
  T* o = dlogits + row * ld_out;
  const float glp = g_num ? g_logp[row] * (g_num[0] / g_den[0]) : g_logp[row];
  const float gh = g_ent ? g_ent[row] : 0.f;
  const float H = g_ent ? entropy[row] : 0.f;
  const float nlse2 = -lse[row] * kLog2e;
  const float A0 = -(glp + gh * H) * inv_temp;
  const float B0 = -gh * kLn2 * inv_temp;
  const int64_t lab = labels[label_row((int)row, (int)gridDim.x, roll)];
  for (int64_t j = threadIdx.x; j < V; j += kPlainThreads) {
    const float y = fmaf(to_f<T>(x[j]), scale, nlse2);
    float d = exp2f(y) * fmaf(B0, y, A0);
    if (j == lab) d += glp * inv_temp;
    o[j] = from_f<T>(d);
  }
}

template <typename K>
int set_smem(K kernel, int bytes) {
  return ensure_dynamic_smem(reinterpret_cast<const void*>(kernel), bytes);
}

// Ring geometry (one CTA per SM). Measured on B200 (scripts/gpu_tune.sh): the forward, whose body
// is MUFU/issue heavy, gains from more warps to hide latency (24 warps x 2 stages: 6.84 TB/s vs
// 6.37 TB/s at 16 x 3); the backward, whose body is light and whose stores compete with the
// loads, prefers the deeper ring (16 x 3: 6.09 TB/s vs 5.93 TB/s at 24 x 2).
#ifndef B200RL_FWD_WARPS
#define B200RL_FWD_WARPS 24
#endif
#ifndef B200RL_FWD_STAGES
#define B200RL_FWD_STAGES 2
#endif
constexpr int kFwdWarps = B200RL_FWD_WARPS, kFwdStages = B200RL_FWD_STAGES;
static_assert(FwdSmem<kFwdWarps, kFwdStages>::kBytes <= 227 * 1024, "forward ring exceeds shared memory");
static_assert(kFwdWarps <= 32, "the finalising warp reads one partial per lane");

// ------------------------------------------------------------------------------------------
// vocabulary-parallel combine: one thread per row folds the n_shards partial records
// (after the tensor-parallel all-gather) with the same algebra the warps use inside a row.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
logprob_shard_combine_kernel(const float* __restrict__ partials, int n_shards, int64_t T, float scale,
                             float* __restrict__ logp, float* __restrict__ ent,
                             float* __restrict__ lse, float* __restrict__ rowmax,
                             float* __restrict__ rowmin) {
  const int64_t row = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (row >= T) return;
  combine_shard_row(reinterpret_cast<const float4*>(partials), n_shards, T, row, scale, logp, ent, lse, rowmax,
                    rowmin);
}

template <int DT, bool ENT, bool MINMAX>
int launch_fwd_fast(const void* logits, int64_t ld, const int64_t* labels, int T, int V,
                    float scale, float* logp, float* ent, float* lse, float* rowmax,
                    float* rowmin, float* partial, int roll, cudaStream_t st) {
  auto k = logprob_fwd_kernel<DT, kFwdWarps, kFwdStages, ENT, MINMAX>;
  constexpr int smem = FwdSmem<kFwdWarps, kFwdStages>::kBytes;
  int rc = set_smem(k, smem);
  if (rc) return rc;
  const int grid = min(T, sm_count());
  k<<<grid, kFwdWarps * 32, smem, st>>>(logits, ld, labels, T, V, scale, logp, ent, lse, rowmax, rowmin,
                                        partial, roll);
  return check_launch("logprob_fwd_kernel");
}

template <int DT>
int dispatch_fwd_fast(const void* logits, int64_t ld, const int64_t* labels, int T, int V,
                      float scale, float* logp, float* ent, float* lse, float* rowmax,
                      float* rowmin, float* partial, bool want_ent, bool want_mm, int roll,
                      cudaStream_t st) {
  // (Timing the instances back to back shows differences of up to 8 % that follow the ORDER of the
  // runs, not the instance: the forward is MUFU/issue-bound and tracks the SM clock, which sags under
  // sustained load. In the first seconds of a run all instances take 458-462 us per 10,240 rows.)
  if (want_ent && want_mm)
    return launch_fwd_fast<DT, true, true>(logits, ld, labels, T, V, scale, logp, ent, lse, rowmax, rowmin, partial, roll, st);
  if (want_ent)
    return launch_fwd_fast<DT, true, false>(logits, ld, labels, T, V, scale, logp, ent, lse, rowmax, rowmin, partial, roll, st);
  return want_mm ? launch_fwd_fast<DT, false, true>(logits, ld, labels, T, V, scale, logp, ent, lse, rowmax, rowmin, partial, roll, st)
                 : launch_fwd_fast<DT, false, false>(logits, ld, labels, T, V, scale, logp, ent, lse, rowmax, rowmin, partial, roll, st);
}

template <int DT, bool HAS_GH>
int launch_bwd_fast(const void* logits, int64_t ld, const int64_t* labels, const float* lse,
                    const float* entropy, const float* g_logp, const float* g_ent, int T, int V,
                    float scale, float inv_temp, void* dlogits, int64_t ld_out, int roll,
                    const float* g_num, const float* g_den, cudaStream_t st) {
  const int tiles = (V / Elem<DT>::kPerVec + kBwdTileVec - 1) / kBwdTileVec;
  const int64_t grid = (int64_t)T * tiles;
  if (grid > 0x7fffffffLL) {
    set_last_error("T * tiles = %lld exceeds the grid limit", (long long)grid);
    return B200RL_EINVAL;
  }
  logprob_bwd_kernel<DT, HAS_GH><<<(unsigned)grid, kBwdThreads, 0, st>>>(
      logits, ld, labels, lse, entropy, g_logp, g_ent, T, V, scale, inv_temp, dlogits, ld_out, roll, g_num,
      g_den, tiles);
  return check_launch("logprob_bwd_kernel");
}

template <int DT>
int dispatch_bwd_fast(const void* logits, int64_t ld, const int64_t* labels, const float* lse,
                      const float* entropy, const float* g_logp, const float* g_ent, int T, int V,
                      float scale, float inv_temp, void* dlogits, int64_t ld_out, int roll,
                      const float* g_num, const float* g_den, cudaStream_t st) {
  return g_ent ? launch_bwd_fast<DT, true>(logits, ld, labels, lse, entropy, g_logp, g_ent, T, V, scale,
                                           inv_temp, dlogits, ld_out, roll, g_num, g_den, st)
               : launch_bwd_fast<DT, false>(logits, ld, labels, lse, entropy, g_logp, g_ent, T, V, scale,
                                            inv_temp, dlogits, ld_out, roll, g_num, g_den, st);
}

inline int per_vec(int dtype) { return dtype == B200RL_FP32 ? 4 : 8; }

}  // namespace

// shared by the whole-vocabulary forward, the vocabulary-shard (partial) forward and the
// grpo_loss composite (grpo_loss.cu); declared in common.cuh
int logprob_fwd_impl(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                     int64_t T, int64_t V, float temperature, float* logp, float* entropy,
                     float* lse, float* rowmax, float* rowmin, float* partial, bool want_ent,
                     bool want_mm, int roll, void* stream) {
  B200RL_REQUIRE(T >= 0 && V > 0, "bad shape T=%lld V=%lld", (long long)T, (long long)V);
  B200RL_REQUIRE(ld >= V, "row stride %lld < V %lld", (long long)ld, (long long)V);
  B200RL_REQUIRE(temperature > 0.f, "temperature must be > 0, got %f", temperature);
  B200RL_REQUIRE(T < INT32_MAX && V < INT32_MAX, "T and V must fit int32");
  if (T == 0) return B200RL_OK;
  B200RL_REQUIRE(logits && labels && (partial || (logp && lse)), "null pointer");
  B200RL_REQUIRE(!partial || (reinterpret_cast<uintptr_t>(partial) % 16 == 0),
                 "partial buffer must be 16-B aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const float scale = (float)(1.4426950408889634 / (double)temperature);
  B200RL_REQUIRE(dtype == B200RL_BF16 || dtype == B200RL_FP16 || dtype == B200RL_FP32,
                 "unsupported logits dtype code %d", dtype);
  const int pv = per_vec(dtype);
  const bool fast = (V % pv == 0) && (ld % pv == 0) && (reinterpret_cast<uintptr_t>(logits) % 16 == 0);
  if (fast) {
    switch (dtype) {
      case B200RL_BF16:
        return dispatch_fwd_fast<B200RL_BF16>(logits, ld, labels, (int)T, (int)V, scale, logp, entropy, lse, rowmax, rowmin, partial, want_ent, want_mm, roll, st);
      case B200RL_FP16:
        return dispatch_fwd_fast<B200RL_FP16>(logits, ld, labels, (int)T, (int)V, scale, logp, entropy, lse, rowmax, rowmin, partial, want_ent, want_mm, roll, st);
      default:
        return dispatch_fwd_fast<B200RL_FP32>(logits, ld, labels, (int)T, (int)V, scale, logp, entropy, lse, rowmax, rowmin, partial, want_ent, want_mm, roll, st);
    }
  }
  const unsigned grid = (unsigned)T;
  switch (dtype) {
    case B200RL_BF16:
      logprob_fwd_plain_kernel<__nv_bfloat16><<<grid, kPlainThreads, 0, st>>>(
          (const __nv_bfloat16*)logits, ld, labels, V, scale, logp, entropy, lse, rowmax, rowmin, partial, roll);
      break;
    case B200RL_FP16:
      logprob_fwd_plain_kernel<__half><<<grid, kPlainThreads, 0, st>>>(
          (const __half*)logits, ld, labels, V, scale, logp, entropy, lse, rowmax, rowmin, partial, roll);
      break;
    case B200RL_FP32:
      logprob_fwd_plain_kernel<float><<<grid, kPlainThreads, 0, st>>>(
          (const float*)logits, ld, labels, V, scale, logp, entropy, lse, rowmax, rowmin, partial, roll);
      break;
    default:
      set_last_error("unsupported logits dtype code %d", dtype);
      return B200RL_EUNSUPPORTED;
  }
  return check_launch("logprob_fwd_plain_kernel");
}

}  // namespace b200rl

extern "C" {

int b200rl_logprob_entropy_fwd(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                               int64_t T, int64_t V, float temperature, float* logp,
                               float* entropy, float* lse, float* rowmax, float* rowmin,
                               void* stream) {
  return b200rl::logprob_fwd_impl(logits, dtype, ld, labels, T, V, temperature, logp, entropy, lse,
                                  rowmax, rowmin, nullptr, entropy != nullptr,
                                  rowmax != nullptr || rowmin != nullptr, 0, stream);
}

int b200rl_logprob_shard_fwd(const void* logits, int dtype, int64_t ld, const int64_t* local_labels,
                             int64_t T, int64_t V_shard, float temperature, int need_entropy,
                             int need_minmax, float* partial, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(partial || T == 0, "partial must not be null");
  return logprob_fwd_impl(logits, dtype, ld, local_labels, T, V_shard, temperature, nullptr, nullptr,
                          nullptr, nullptr, nullptr, partial, need_entropy != 0, need_minmax != 0, 0,
                          stream);
}

int b200rl_logprob_shard_combine(const float* partials, int32_t n_shards, int64_t T,
                                 float temperature, float* logp, float* entropy, float* lse,
                                 float* rowmax, float* rowmin, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(T >= 0 && n_shards >= 1, "bad shape T=%lld n_shards=%d", (long long)T, n_shards);
  B200RL_REQUIRE(temperature > 0.f, "temperature must be > 0, got %f", temperature);
  if (T == 0) return B200RL_OK;
  B200RL_REQUIRE(partials && logp && lse, "null pointer");
  B200RL_REQUIRE(reinterpret_cast<uintptr_t>(partials) % 16 == 0, "partials must be 16-B aligned");
  const float scale = (float)(1.4426950408889634 / (double)temperature);
  const unsigned grid = (unsigned)((T + 255) / 256);
  logprob_shard_combine_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      partials, n_shards, T, scale, logp, entropy, lse, rowmax, rowmin);
  return check_launch("logprob_shard_combine_kernel");
}

int b200rl_logprob_entropy_bwd(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                               const float* lse, const float* entropy, const float* g_logp,
                               const float* g_ent, int64_t T, int64_t V, float temperature,
                               void* dlogits, int64_t ld_out, void* stream) {
  return b200rl::logprob_bwd_impl(logits, dtype, ld, labels, lse, entropy, g_logp, g_ent, T, V,
                                  temperature, dlogits, ld_out, 0, nullptr, nullptr, stream);
}

}  // extern "C"

namespace b200rl {

int logprob_bwd_impl(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                     const float* lse, const float* entropy, const float* g_logp,
                     const float* g_ent, int64_t T, int64_t V, float temperature, void* dlogits,
                     int64_t ld_out, int roll, const float* g_num, const float* g_den,
                     void* stream) {
  B200RL_REQUIRE(T >= 0 && V > 0, "bad shape T=%lld V=%lld", (long long)T, (long long)V);
  B200RL_REQUIRE(ld >= V && ld_out >= V, "row stride < V");
  B200RL_REQUIRE(temperature > 0.f, "temperature must be > 0, got %f", temperature);
  B200RL_REQUIRE(T < INT32_MAX && V < INT32_MAX, "T and V must fit int32");
  if (T == 0) return B200RL_OK;
  B200RL_REQUIRE(logits && labels && lse && g_logp && dlogits, "null pointer");
  B200RL_REQUIRE(!g_ent || entropy, "g_ent given without the forward's entropy");
  cudaStream_t st = (cudaStream_t)stream;
  const float scale = (float)(1.4426950408889634 / (double)temperature);
  const float inv_temp = (float)(1.0 / (double)temperature);
  B200RL_REQUIRE(dtype == B200RL_BF16 || dtype == B200RL_FP16 || dtype == B200RL_FP32,
                 "unsupported logits dtype code %d", dtype);
  const int pv = per_vec(dtype);
  const bool fast = (V % pv == 0) && (ld % pv == 0) && (ld_out % pv == 0) &&
                    (reinterpret_cast<uintptr_t>(logits) % 16 == 0) &&
                    (reinterpret_cast<uintptr_t>(dlogits) % 16 == 0);
  if (fast) {
    switch (dtype) {
      case B200RL_BF16:
        return dispatch_bwd_fast<B200RL_BF16>(logits, ld, labels, lse, entropy, g_logp, g_ent, (int)T, (int)V, scale, inv_temp, dlogits, ld_out, roll, g_num, g_den, st);
      case B200RL_FP16:
        return dispatch_bwd_fast<B200RL_FP16>(logits, ld, labels, lse, entropy, g_logp, g_ent, (int)T, (int)V, scale, inv_temp, dlogits, ld_out, roll, g_num, g_den, st);
      default:
        return dispatch_bwd_fast<B200RL_FP32>(logits, ld, labels, lse, entropy, g_logp, g_ent, (int)T, (int)V, scale, inv_temp, dlogits, ld_out, roll, g_num, g_den, st);
    }
  }
  const unsigned grid = (unsigned)T;
  switch (dtype) {
    case B200RL_BF16:
      logprob_bwd_plain_kernel<__nv_bfloat16><<<grid, kPlainThreads, 0, st>>>(
          (const __nv_bfloat16*)logits, ld, labels, lse, entropy, g_logp, g_ent, V, scale,
          inv_temp, (__nv_bfloat16*)dlogits, ld_out, roll, g_num, g_den);
      break;
    case B200RL_FP16:
      logprob_bwd_plain_kernel<__half><<<grid, kPlainThreads, 0, st>>>(
          (const __half*)logits, ld, labels, lse, entropy, g_logp, g_ent, V, scale, inv_temp,
          (__half*)dlogits, ld_out, roll, g_num, g_den);
      break;
    case B200RL_FP32:
      logprob_bwd_plain_kernel<float><<<grid, kPlainThreads, 0, st>>>(
          (const float*)logits, ld, labels, lse, entropy, g_logp, g_ent, V, scale, inv_temp,
          (float*)dlogits, ld_out, roll, g_num, g_den);
      break;
    default:
      set_last_error("unsupported logits dtype code %d", dtype);
      return B200RL_EUNSUPPORTED;
  }
  return check_launch("logprob_bwd_plain_kernel");
}

}  // namespace b200rl
