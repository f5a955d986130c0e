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
//            softmax is never materialised.
//
// Bound: HBM. Algorithmic bytes per token: fwd 2*V, bwd 4*V (bf16 logits).
// fp16/fp32 logits and vocabularies whose rows are not 16-byte aligned take a plain
// (non-TMA) kernel further down: correct, not tuned.
#include "common.cuh"

namespace b200rl {
namespace {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kRescaleThr = 8.0f;       // log2 units a new element may exceed the lane's reference
constexpr float kSentinel = 1.0e30f;      // -reference of a lane that has seen nothing yet
constexpr uint32_t kNeutral = 0xF14AF14Au;  // bf16 pair ~ -1e30: exp -> 0, finite products
constexpr uint32_t kNegInf2 = 0xFF80FF80u;
constexpr uint32_t kPosInf2 = 0x7F807F80u;

constexpr int kChunkVec = 256;            // 16-B vectors per TMA chunk
constexpr int kChunkBytes = kChunkVec * 16;

template <int WARPS, int STAGES>
struct FwdSmem {
  static constexpr int kData = WARPS * STAGES * kChunkBytes;
  static constexpr int kBarOff = kData;
  static constexpr int kPartOff = kBarOff + WARPS * STAGES * 8;
  static constexpr int kBytes = kPartOff + 2 * 5 * WARPS * 4;
};

// Producer state of one warp: walks (row iteration, chunk) pairs of the warp's slice.
struct Producer {
  int it, ck;
  int64_t q, total;
  int stage;
};

// One TMA chunk (<= 256 16-B vectors) of the forward pass for one warp. FULL = all 8 vector
// slots of every lane are live (no predication, no neutral fill).
template <bool FULL, bool ENT, bool MINMAX, typename Refill>
__device__ __forceinline__ void fwd_chunk(uint32_t base, int nv, int lane, float scale,
                                          float& negm, float& s0, float& s1, float& t0,
                                          float& t1, uint32_t& pmx, uint32_t& pmn,
                                          Refill&& refill) {
  uint4 v[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    if (FULL || i * 32 + lane < nv) v[i] = lds128(base + i * 512);
    else v[i] = make_uint4(kNeutral, kNeutral, kNeutral, kNeutral);
  }
  refill();  // the slot is free again
  // chunk maximum (packed bf16x2), lazy rescale of the lane's reference
  uint32_t cm = max_bf16x2(max_bf16x2(v[0].x, v[0].y), max_bf16x2(v[0].z, v[0].w));
#pragma unroll
  for (int i = 1; i < 8; i++)
    cm = max_bf16x2(cm, max_bf16x2(max_bf16x2(v[i].x, v[i].y), max_bf16x2(v[i].z, v[i].w)));
  pmx = max_bf16x2(pmx, cm);
  if (MINMAX) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (FULL || i * 32 + lane < nv)
        pmn = min_bf16x2(pmn, min_bf16x2(min_bf16x2(v[i].x, v[i].y), min_bf16x2(v[i].z, v[i].w)));
    }
  }
  const float cmf = fmaxf(bf16lo(cm), bf16hi(cm)) * scale;
  if (cmf + negm > kRescaleThr) {
    const float d = cmf + negm;  // new reference - old reference > 0
    const float f = ex2_approx(-d);
    const float s = s0 + s1, t = t0 + t1;
    if (ENT) {
      t0 = f * (t - d * s);
      t1 = 0.f;
    }
    s0 = f * s;
    s1 = 0.f;
    negm = -cmf;
  }
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const float y0 = fmaf(bf16lo(w[j]), scale, negm);
      const float y1 = fmaf(bf16hi(w[j]), scale, negm);
      const float e0 = ex2_approx(y0);
      const float e1 = ex2_approx(y1);
      s0 += e0;
      s1 += e1;
      if (ENT) {
        t0 = fmaf(e0, y0, t0);
        t1 = fmaf(e1, y1, t1);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// forward, bf16, rows 16-B aligned
// ------------------------------------------------------------------------------------------
template <int WARPS, int STAGES, bool ENT, bool MINMAX>
__global__ void __launch_bounds__(WARPS * 32, 1)
logprob_fwd_bf16_kernel(const __nv_bfloat16* __restrict__ logits, int64_t ld,
                        const int64_t* __restrict__ labels, int T, int V, float scale,
                        float* __restrict__ logp, float* __restrict__ ent,
                        float* __restrict__ lse, float* __restrict__ rowmax,
                        float* __restrict__ rowmin) {
  extern __shared__ __align__(128) unsigned char smem[];
  using L = FwdSmem<WARPS, STAGES>;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* my_data = smem + warp * STAGES * kChunkBytes;
  uint64_t* my_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOff) + warp * STAGES;
  float* part = reinterpret_cast<float*>(smem + L::kPartOff);  // [2][5][WARPS]

  const int nvec = V >> 3;
  const int seg = (nvec + WARPS - 1) / WARPS;
  const int sv = warp * seg;
  const int seg_len = max(0, min(seg, nvec - sv));
  const int nck = (seg_len + kChunkVec - 1) / kChunkVec;
  const int nrows = (T - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; s++) mbar_init(&my_bar[s], 1);
    fence_barrier_init();
    fence_proxy_async();
  }
  __syncwarp();

  const uint64_t pol = policy_evict_first();
  Producer P{0, 0, 0, (int64_t)nrows * nck, 0};
  auto issue = [&]() {  // lane 0 only
    if (P.q < P.total) {
      const int row = blockIdx.x + P.it * gridDim.x;
      const int nv = min(kChunkVec, seg_len - P.ck * kChunkVec);
      const char* src = reinterpret_cast<const char*>(logits) + ((int64_t)row * ld) * 2 +
                        ((int64_t)(sv + P.ck * kChunkVec)) * 16;
      mbar_expect_tx(&my_bar[P.stage], nv * 16);
      bulk_g2s(my_data + P.stage * kChunkBytes, src, nv * 16, &my_bar[P.stage], pol);
      P.q++;
      if (++P.stage == STAGES) P.stage = 0;
      if (++P.ck == nck) {
        P.ck = 0;
        P.it++;
      }
    }
  };
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; s++) issue();
  }

  auto refill = [&]() {  // called by all lanes once the chunk sits in registers
    __syncwarp();
    if (lane == 0) {
      fence_proxy_async();
      issue();
    }
  };

  int c_stage = 0;
  uint32_t c_par = 0;
  for (int it = 0; it < nrows; ++it) {
    const int row = blockIdx.x + it * gridDim.x;
    // label logit: a single 2-byte gather, issued early by the finalising warp
    float xy = 0.f;
    if (warp == 0 && lane == 0) {
      const int64_t lab = labels[row];
      xy = (lab >= 0 && lab < V)
               ? __bfloat162float(logits[(int64_t)row * ld + lab])
               : __int_as_float(0x7fc00000);
    }
    float negm = kSentinel;  // minus the lane's reference exponent (log2 domain)
    float s0 = 0.f, s1 = 0.f, t0 = 0.f, t1 = 0.f;
    uint32_t pmx = kNegInf2, pmn = kPosInf2;

    for (int ck = 0; ck < nck; ++ck) {
      const int nv = min(kChunkVec, seg_len - ck * kChunkVec);
      mbar_wait(&my_bar[c_stage], c_par);
      const uint32_t base = smem_u32(my_data + c_stage * kChunkBytes) + lane * 16;
      if (nv == kChunkVec)
        fwd_chunk<true, ENT, MINMAX>(base, nv, lane, scale, negm, s0, s1, t0, t1, pmx, pmn, refill);
      else
        fwd_chunk<false, ENT, MINMAX>(base, nv, lane, scale, negm, s0, s1, t0, t1, pmx, pmn, refill);
      if (++c_stage == STAGES) {
        c_stage = 0;
        c_par ^= 1;
      }
    }

    // ---- combine the 32 lanes of this warp -------------------------------------------------
    {
      const float m = -negm;
      const float s = s0 + s1, t = t0 + t1;
      const float M = warp_max(m);
      const float dl = m - M;  // <= 0
      const float f = ex2_approx(dl);
      const float S = warp_sum(s * f);
      float Tt = 0.f;
      if (ENT) Tt = warp_sum(f * fmaf(dl, s, t));
      float mx = 0.f, mn = 0.f;
      if (MINMAX) {
        mx = warp_max(fmaxf(bf16lo(pmx), bf16hi(pmx)));
        mn = warp_min(fminf(bf16lo(pmn), bf16hi(pmn)));
      }
      if (lane == 0) {
        float* p = part + (it & 1) * 5 * WARPS;
        p[0 * WARPS + warp] = M;
        p[1 * WARPS + warp] = S;
        p[2 * WARPS + warp] = Tt;
        p[3 * WARPS + warp] = mx;
        p[4 * WARPS + warp] = mn;
      }
    }
    __syncthreads();
    if (warp == 0) {
      const float* p = part + (it & 1) * 5 * WARPS;
      const bool has = lane < WARPS;
      const float Mw = has ? p[0 * WARPS + lane] : -kSentinel;
      const float Sw = has ? p[1 * WARPS + lane] : 0.f;
      const float Tw = has ? p[2 * WARPS + lane] : 0.f;
      const float M = warp_max(Mw);
      const float dl = Mw - M;
      const float f = ex2_approx(dl);
      const float S = warp_sum(Sw * f);
      float Tt = 0.f;
      if (ENT) Tt = warp_sum(f * fmaf(dl, Sw, Tw));
      float mx = 0.f, mn = 0.f;
      if (MINMAX) {
        mx = warp_max(has ? p[3 * WARPS + lane] : -INFINITY);
        mn = warp_min(has ? p[4 * WARPS + lane] : INFINITY);
      }
      if (lane == 0) {
        const float l2s = log2f(S);
        const float lse2 = M + l2s;  // log2 domain
        lse[row] = lse2 * kLn2;
        // (x_y*scale - M) first: exact cancellation when the label is the row maximum, however
        // large |M| is; only then the (small) log2 of the sum
        logp[row] = (fmaf(xy, scale, -M) - l2s) * kLn2;
        if (ENT) ent[row] = kLn2 * (l2s - Tt / S);
        if (MINMAX) {
          if (rowmax) rowmax[row] = mx;
          if (rowmin) rowmin[row] = mn;
        }
      }
    }
  }
}

// One TMA chunk of the backward pass for one warp.
template <bool FULL, bool HAS_GH, typename Refill>
__device__ __forceinline__ void bwd_chunk(uint32_t base, int nv, int lane, int v0, float scale,
                                          float nlse2, float A0, float B0, float lab_add,
                                          int lvec, int lsub, char* out_row, Refill&& refill) {
  uint4 v[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    if (FULL || i * 32 + lane < nv) v[i] = lds128(base + i * 512);
    else v[i] = make_uint4(kNeutral, kNeutral, kNeutral, kNeutral);
  }
  refill();
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
    float d[8];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const float y0 = fmaf(bf16lo(w[j]), scale, nlse2);
      const float y1 = fmaf(bf16hi(w[j]), scale, nlse2);
      const float p0 = ex2_approx(y0);
      const float p1 = ex2_approx(y1);
      if (HAS_GH) {
        d[2 * j] = p0 * fmaf(B0, y0, A0);
        d[2 * j + 1] = p1 * fmaf(B0, y1, A0);
      } else {
        d[2 * j] = p0 * A0;
        d[2 * j + 1] = p1 * A0;
      }
    }
    const int vi = v0 + i * 32;
    if (vi == lvec) {  // the one-hot term of the label, one lane once per row
#pragma unroll
      for (int k = 0; k < 8; k++)
        if (k == lsub) d[k] += lab_add;
    }
    if (FULL || i * 32 + lane < nv) {
      uint4 o;
      o.x = pack_bf16x2(d[0], d[1]);
      o.y = pack_bf16x2(d[2], d[3]);
      o.z = pack_bf16x2(d[4], d[5]);
      o.w = pack_bf16x2(d[6], d[7]);
      stg_na_v4(out_row + (int64_t)vi * 16, o);
    }
  }
}

// ------------------------------------------------------------------------------------------
// backward, bf16, rows 16-B aligned (input and output)
// ------------------------------------------------------------------------------------------
template <int WARPS, int STAGES, bool HAS_GH>
__global__ void __launch_bounds__(WARPS * 32, 1)
logprob_bwd_bf16_kernel(const __nv_bfloat16* __restrict__ logits, int64_t ld,
                        const int64_t* __restrict__ labels, const float* __restrict__ lse,
                        const float* __restrict__ entropy, const float* __restrict__ g_logp,
                        const float* __restrict__ g_ent, int T, int V, float scale,
                        float inv_temp, __nv_bfloat16* __restrict__ dlogits, int64_t ld_out) {
  extern __shared__ __align__(128) unsigned char smem[];
  using L = FwdSmem<WARPS, STAGES>;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* my_data = smem + warp * STAGES * kChunkBytes;
  uint64_t* my_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOff) + warp * STAGES;

  const int nvec = V >> 3;
  const int seg = (nvec + WARPS - 1) / WARPS;
  const int sv = warp * seg;
  const int seg_len = max(0, min(seg, nvec - sv));
  const int nck = (seg_len + kChunkVec - 1) / kChunkVec;
  const int nrows = (T - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  if (nck == 0) return;  // this warp owns no part of the row (tiny vocabularies)

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; s++) mbar_init(&my_bar[s], 1);
    fence_barrier_init();
    fence_proxy_async();
  }
  __syncwarp();

  const uint64_t pol = policy_evict_first();
  Producer P{0, 0, 0, (int64_t)nrows * nck, 0};
  auto issue = [&]() {
    if (P.q < P.total) {
      const int row = blockIdx.x + P.it * gridDim.x;
      const int nv = min(kChunkVec, seg_len - P.ck * kChunkVec);
      const char* src = reinterpret_cast<const char*>(logits) + ((int64_t)row * ld) * 2 +
                        ((int64_t)(sv + P.ck * kChunkVec)) * 16;
      mbar_expect_tx(&my_bar[P.stage], nv * 16);
      bulk_g2s(my_data + P.stage * kChunkBytes, src, nv * 16, &my_bar[P.stage], pol);
      P.q++;
      if (++P.stage == STAGES) P.stage = 0;
      if (++P.ck == nck) {
        P.ck = 0;
        P.it++;
      }
    }
  };
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; s++) issue();
  }

  auto refill = [&]() {
    __syncwarp();
    if (lane == 0) {
      fence_proxy_async();
      issue();
    }
  };

  // per-row constants, fetched one row ahead
  auto load_row = [&](int row, float& nlse2, float& A0, float& B0, float& lab_add, int& lvec,
                      int& lsub) {
    const float glp = g_logp[row];
    nlse2 = -lse[row] * kLog2e;
    float gh = 0.f, H = 0.f;
    if (HAS_GH) {
      gh = g_ent[row];
      H = entropy[row];
    }
    A0 = -(glp + gh * H) * inv_temp;
    B0 = -gh * kLn2 * inv_temp;
    lab_add = glp * inv_temp;
    const int64_t lab = labels[row];
    const bool ok = lab >= 0 && lab < V;
    lvec = ok ? (int)(lab >> 3) : -1;
    lsub = (int)(lab & 7);
  };
  float n_nlse2 = 0.f, n_A0 = 0.f, n_B0 = 0.f, n_lab = 0.f;
  int n_lvec = -1, n_lsub = 0;
  if (nrows > 0) load_row(blockIdx.x, n_nlse2, n_A0, n_B0, n_lab, n_lvec, n_lsub);

  int c_stage = 0;
  uint32_t c_par = 0;
  for (int it = 0; it < nrows; ++it) {
    const int row = blockIdx.x + it * gridDim.x;
    const float nlse2 = n_nlse2, A0 = n_A0, B0 = n_B0, lab_add = n_lab;
    const int lvec = n_lvec, lsub = n_lsub;
    if (it + 1 < nrows) load_row(row + gridDim.x, n_nlse2, n_A0, n_B0, n_lab, n_lvec, n_lsub);
    char* out_row = reinterpret_cast<char*>(dlogits) + ((int64_t)row * ld_out) * 2;

    for (int ck = 0; ck < nck; ++ck) {
      const int nv = min(kChunkVec, seg_len - ck * kChunkVec);
      mbar_wait(&my_bar[c_stage], c_par);
      const uint32_t base = smem_u32(my_data + c_stage * kChunkBytes) + lane * 16;
      const int v0 = sv + ck * kChunkVec + lane;  // global 16-B vector index of slot 0
      if (nv == kChunkVec)
        bwd_chunk<true, HAS_GH>(base, nv, lane, v0, scale, nlse2, A0, B0, lab_add, lvec, lsub, out_row, refill);
      else
        bwd_chunk<false, HAS_GH>(base, nv, lane, v0, scale, nlse2, A0, B0, lab_add, lvec, lsub, out_row, refill);
      if (++c_stage == STAGES) {
        c_stage = 0;
        c_par ^= 1;
      }
    }
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
                         float* __restrict__ rowmin) {
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
    const int64_t lab = labels[row];
    const float xy = (lab >= 0 && lab < V) ? to_f<T>(x[lab]) : __int_as_float(0x7fc00000);
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
                         int64_t ld_out) {
  const int64_t row = blockIdx.x;
  const T* x = logits + row * ld;
  T* o = dlogits + row * ld_out;
  const float glp = g_logp[row];
  const float gh = g_ent ? g_ent[row] : 0.f;
  const float H = g_ent ? entropy[row] : 0.f;
  const float nlse2 = -lse[row] * kLog2e;
  const float A0 = -(glp + gh * H) * inv_temp;
  const float B0 = -gh * kLn2 * inv_temp;
  const int64_t lab = labels[row];
  for (int64_t j = threadIdx.x; j < V; j += kPlainThreads) {
    const float y = fmaf(to_f<T>(x[j]), scale, nlse2);
    float d = exp2f(y) * fmaf(B0, y, A0);
    if (j == lab) d += glp * inv_temp;
    o[j] = from_f<T>(d);
  }
}

template <typename K>
int set_smem(K kernel, int bytes) {
  // once per (kernel, device); a benign race at worst sets the attribute twice
  static const void* seen_fn[64];
  static int seen_dev[64];
  static int n_seen = 0;
  int dev = 0;
  cudaGetDevice(&dev);
  const void* fn = reinterpret_cast<const void*>(kernel);
  for (int i = 0; i < n_seen && i < 64; i++)
    if (seen_fn[i] == fn && seen_dev[i] == dev) return 0;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) {
    set_last_error("cudaFuncSetAttribute(smem=%d): %s", bytes, cudaGetErrorString(e));
    return (int)e;
  }
  if (n_seen < 64) {
    seen_fn[n_seen] = fn;
    seen_dev[n_seen] = dev;
    n_seen++;
  }
  return 0;
}

constexpr int kWarps = 16;
constexpr int kStages = 3;

template <bool ENT, bool MINMAX>
int launch_fwd_fast(const void* logits, int64_t ld, const int64_t* labels, int T, int V,
                    float scale, float* logp, float* ent, float* lse, float* rowmax,
                    float* rowmin, cudaStream_t st) {
  auto k = logprob_fwd_bf16_kernel<kWarps, kStages, ENT, MINMAX>;
  constexpr int smem = FwdSmem<kWarps, kStages>::kBytes;
  int rc = set_smem(k, smem);
  if (rc) return rc;
  const int grid = min(T, sm_count());
  k<<<grid, kWarps * 32, smem, st>>>((const __nv_bfloat16*)logits, ld, labels, T, V, scale, logp,
                                     ent, lse, rowmax, rowmin);
  return check_launch("logprob_fwd_bf16_kernel");
}

}  // namespace
}  // namespace b200rl

extern "C" {

int b200rl_logprob_entropy_fwd(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                               int64_t T, int64_t V, float temperature, float* logp,
                               float* entropy, float* lse, float* rowmax, float* rowmin,
                               void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(T >= 0 && V > 0, "bad shape T=%lld V=%lld", (long long)T, (long long)V);
  B200RL_REQUIRE(ld >= V, "row stride %lld < V %lld", (long long)ld, (long long)V);
  B200RL_REQUIRE(temperature > 0.f, "temperature must be > 0, got %f", temperature);
  B200RL_REQUIRE(T < INT32_MAX && V < INT32_MAX, "T and V must fit int32");
  if (T == 0) return B200RL_OK;
  B200RL_REQUIRE(logits && labels && logp && lse, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const float scale = (float)(1.4426950408889634 / (double)temperature);
  const bool fast = dtype == B200RL_BF16 && (V % 8 == 0) && (ld % 8 == 0) &&
                    (reinterpret_cast<uintptr_t>(logits) % 16 == 0);
  if (fast) {
    const bool mm = rowmax != nullptr || rowmin != nullptr;
    if (entropy) {
      return mm ? launch_fwd_fast<true, true>(logits, ld, labels, (int)T, (int)V, scale, logp,
                                              entropy, lse, rowmax, rowmin, st)
                : launch_fwd_fast<true, false>(logits, ld, labels, (int)T, (int)V, scale, logp,
                                               entropy, lse, rowmax, rowmin, st);
    }
    return mm ? launch_fwd_fast<false, true>(logits, ld, labels, (int)T, (int)V, scale, logp,
                                             entropy, lse, rowmax, rowmin, st)
              : launch_fwd_fast<false, false>(logits, ld, labels, (int)T, (int)V, scale, logp,
                                              entropy, lse, rowmax, rowmin, st);
  }
  const unsigned grid = (unsigned)T;
  switch (dtype) {
    case B200RL_BF16:
      logprob_fwd_plain_kernel<__nv_bfloat16><<<grid, kPlainThreads, 0, st>>>(
          (const __nv_bfloat16*)logits, ld, labels, V, scale, logp, entropy, lse, rowmax, rowmin);
      break;
    case B200RL_FP16:
      logprob_fwd_plain_kernel<__half><<<grid, kPlainThreads, 0, st>>>(
          (const __half*)logits, ld, labels, V, scale, logp, entropy, lse, rowmax, rowmin);
      break;
    case B200RL_FP32:
      logprob_fwd_plain_kernel<float><<<grid, kPlainThreads, 0, st>>>(
          (const float*)logits, ld, labels, V, scale, logp, entropy, lse, rowmax, rowmin);
      break;
    default:
      set_last_error("unsupported logits dtype code %d", dtype);
      return B200RL_EUNSUPPORTED;
  }
  return check_launch("logprob_fwd_plain_kernel");
}

int b200rl_logprob_entropy_bwd(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                               const float* lse, const float* entropy, const float* g_logp,
                               const float* g_ent, int64_t T, int64_t V, float temperature,
                               void* dlogits, int64_t ld_out, void* stream) {
  using namespace b200rl;
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
  const bool fast = dtype == B200RL_BF16 && (V % 8 == 0) && (ld % 8 == 0) && (ld_out % 8 == 0) &&
                    (reinterpret_cast<uintptr_t>(logits) % 16 == 0) &&
                    (reinterpret_cast<uintptr_t>(dlogits) % 16 == 0);
  if (fast) {
    constexpr int smem = FwdSmem<kWarps, kStages>::kBytes;
    const int grid = min((int)T, sm_count());
    if (g_ent) {
      auto k = logprob_bwd_bf16_kernel<kWarps, kStages, true>;
      int rc = set_smem(k, smem);
      if (rc) return rc;
      k<<<grid, kWarps * 32, smem, st>>>((const __nv_bfloat16*)logits, ld, labels, lse, entropy,
                                         g_logp, g_ent, (int)T, (int)V, scale, inv_temp,
                                         (__nv_bfloat16*)dlogits, ld_out);
    } else {
      auto k = logprob_bwd_bf16_kernel<kWarps, kStages, false>;
      int rc = set_smem(k, smem);
      if (rc) return rc;
      k<<<grid, kWarps * 32, smem, st>>>((const __nv_bfloat16*)logits, ld, labels, lse, entropy,
                                         g_logp, g_ent, (int)T, (int)V, scale, inv_temp,
                                         (__nv_bfloat16*)dlogits, ld_out);
    }
    return check_launch("logprob_bwd_bf16_kernel");
  }
  const unsigned grid = (unsigned)T;
  switch (dtype) {
    case B200RL_BF16:
      logprob_bwd_plain_kernel<__nv_bfloat16><<<grid, kPlainThreads, 0, st>>>(
          (const __nv_bfloat16*)logits, ld, labels, lse, entropy, g_logp, g_ent, V, scale,
          inv_temp, (__nv_bfloat16*)dlogits, ld_out);
      break;
    case B200RL_FP16:
      logprob_bwd_plain_kernel<__half><<<grid, kPlainThreads, 0, st>>>(
          (const __half*)logits, ld, labels, lse, entropy, g_logp, g_ent, V, scale, inv_temp,
          (__half*)dlogits, ld_out);
      break;
    case B200RL_FP32:
      logprob_bwd_plain_kernel<float><<<grid, kPlainThreads, 0, st>>>(
          (const float*)logits, ld, labels, lse, entropy, g_logp, g_ent, V, scale, inv_temp,
          (float*)dlogits, ld_out);
      break;
    default:
      set_last_error("unsupported logits dtype code %d", dtype);
      return B200RL_EUNSUPPORTED;
  }
  return check_launch("logprob_bwd_plain_kernel");
}

}  // extern "C"
