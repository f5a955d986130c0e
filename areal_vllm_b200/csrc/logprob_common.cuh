// Device code shared by the log-prob kernels (logprob.cu) and the fused GRPO step kernel
// (grpo_fused.cu): element traits, the per-warp TMA ring, the online-softmax lane state.
#pragma once
#include "common.cuh"

namespace b200rl {
namespace lp {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kRescaleThr = 8.0f;       // log2 units a new element may exceed the lane's reference
constexpr float kSentinel = 1.0e30f;      // -reference of a lane that has seen nothing yet
constexpr uint32_t kNegInf2 = 0xFF80FF80u;
constexpr uint32_t kPosInf2 = 0x7F807F80u;


// Element traits of the three logits dtypes. A 16-byte vector holds kPerVec elements on the way
// in (logits) and on the way out (dlogits), so vector indices coincide for every dtype.
template <int DT> struct Elem;

template <> struct Elem<B200RL_BF16> {
  static constexpr int kPerVec = 8, kBytes = 2;
  static constexpr uint32_t kNeutralWord = 0xF14AF14Au;  // pair of ~-1e30: exp -> 0, finite products
  static constexpr uint32_t kSignWord = 0x80008000u;     // sign bits of a packed pair
  using MaxT = uint32_t;  // packed bf16x2 running extremum
  static __device__ __forceinline__ MaxT max_init() { return kNegInf2; }
  static __device__ __forceinline__ MaxT min_init() { return kPosInf2; }
  static __device__ __forceinline__ MaxT cmax(MaxT a, MaxT b) { return max_bf16x2(a, b); }
  static __device__ __forceinline__ MaxT cmin(MaxT a, MaxT b) { return min_bf16x2(a, b); }
  static __device__ __forceinline__ MaxT vmax(const uint4& v) { return cmax(cmax(v.x, v.y), cmax(v.z, v.w)); }
  static __device__ __forceinline__ MaxT vmin(const uint4& v) { return cmin(cmin(v.x, v.y), cmin(v.z, v.w)); }
  static __device__ __forceinline__ float max_f(MaxT m) { return fmaxf(bf16lo(m), bf16hi(m)); }
  static __device__ __forceinline__ float min_f(MaxT m) { return fminf(bf16lo(m), bf16hi(m)); }
  static __device__ __forceinline__ void unpack(const uint4& v, float (&x)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
      x[2 * j] = bf16lo(w[j]);
      x[2 * j + 1] = bf16hi(w[j]);
    }
  }
  static __device__ __forceinline__ uint4 pack(const float (&d)[8]) {
    return make_uint4(pack_bf16x2(d[0], d[1]), pack_bf16x2(d[2], d[3]), pack_bf16x2(d[4], d[5]),
                      pack_bf16x2(d[6], d[7]));
  }
  static __device__ __forceinline__ float load1(const void* base, int64_t i) {
    return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[i]);
  }
};

template <> struct Elem<B200RL_FP16> {
  static constexpr int kPerVec = 8, kBytes = 2;
  static constexpr uint32_t kNeutralWord = 0xFBFFFBFFu;  // pair of -65504 (largest finite |fp16|)
  static constexpr uint32_t kSignWord = 0x80008000u;
  using MaxT = uint32_t;  // packed f16x2
  static __device__ __forceinline__ MaxT max_init() { return 0xFC00FC00u; }
  static __device__ __forceinline__ MaxT min_init() { return 0x7C007C00u; }
  static __device__ __forceinline__ MaxT cmax(MaxT a, MaxT b) {
    uint32_t r;
    asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
  }
  static __device__ __forceinline__ MaxT cmin(MaxT a, MaxT b) {
    uint32_t r;
    asm("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
  }
  static __device__ __forceinline__ MaxT vmax(const uint4& v) { return cmax(cmax(v.x, v.y), cmax(v.z, v.w)); }
  static __device__ __forceinline__ MaxT vmin(const uint4& v) { return cmin(cmin(v.x, v.y), cmin(v.z, v.w)); }
  static __device__ __forceinline__ float2 f2(uint32_t w) {
    return __half22float2(*reinterpret_cast<const __half2*>(&w));
  }
  static __device__ __forceinline__ float max_f(MaxT m) { float2 f = f2(m); return fmaxf(f.x, f.y); }
  static __device__ __forceinline__ float min_f(MaxT m) { float2 f = f2(m); return fminf(f.x, f.y); }
  static __device__ __forceinline__ void unpack(const uint4& v, float (&x)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const float2 f = f2(w[j]);
      x[2 * j] = f.x;
      x[2 * j + 1] = f.y;
    }
  }
  static __device__ __forceinline__ uint4 pack(const float (&d)[8]) {
    uint4 o;
    __half2 h;
    h = __floats2half2_rn(d[0], d[1]); o.x = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2half2_rn(d[2], d[3]); o.y = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2half2_rn(d[4], d[5]); o.z = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2half2_rn(d[6], d[7]); o.w = *reinterpret_cast<uint32_t*>(&h);
    return o;
  }
  static __device__ __forceinline__ float load1(const void* base, int64_t i) {
    return __half2float(reinterpret_cast<const __half*>(base)[i]);
  }
};

template <> struct Elem<B200RL_FP32> {
  static constexpr int kPerVec = 4, kBytes = 4;
  static constexpr uint32_t kNeutralWord = 0xF149F2CAu;  // -1e30f
  static constexpr uint32_t kSignWord = 0x80000000u;
  using MaxT = float;
  static __device__ __forceinline__ MaxT max_init() { return -INFINITY; }
  static __device__ __forceinline__ MaxT min_init() { return INFINITY; }
  static __device__ __forceinline__ MaxT cmax(MaxT a, MaxT b) { return fmaxf(a, b); }
  static __device__ __forceinline__ MaxT cmin(MaxT a, MaxT b) { return fminf(a, b); }
  static __device__ __forceinline__ MaxT vmax(const uint4& v) {
    return fmaxf(fmaxf(__uint_as_float(v.x), __uint_as_float(v.y)),
                 fmaxf(__uint_as_float(v.z), __uint_as_float(v.w)));
  }
  static __device__ __forceinline__ MaxT vmin(const uint4& v) {
    return fminf(fminf(__uint_as_float(v.x), __uint_as_float(v.y)),
                 fminf(__uint_as_float(v.z), __uint_as_float(v.w)));
  }
  static __device__ __forceinline__ float max_f(MaxT m) { return m; }
  static __device__ __forceinline__ float min_f(MaxT m) { return m; }
  static __device__ __forceinline__ void unpack(const uint4& v, float (&x)[4]) {
    x[0] = __uint_as_float(v.x);
    x[1] = __uint_as_float(v.y);
    x[2] = __uint_as_float(v.z);
    x[3] = __uint_as_float(v.w);
  }
  static __device__ __forceinline__ uint4 pack(const float (&d)[4]) {
    return make_uint4(__float_as_uint(d[0]), __float_as_uint(d[1]), __float_as_uint(d[2]),
                      __float_as_uint(d[3]));
  }
  static __device__ __forceinline__ float load1(const void* base, int64_t i) {
    return reinterpret_cast<const float*>(base)[i];
  }
};

constexpr int kChunkVec = 256;            // 16-B vectors per full TMA chunk (8 warp-rows of 32)
constexpr int kChunkBytes = kChunkVec * 16;
constexpr int kSlots = 4;                 // row-result slots in flight per CTA (forward)

template <int WARPS, int STAGES>
struct FwdSmem {
  static constexpr int kData = WARPS * STAGES * kChunkBytes;
  static constexpr int kBarOff = kData;
  static constexpr int kPartOff = kBarOff + WARPS * STAGES * 8;
  static constexpr int kCntOff = kPartOff + kSlots * 6 * WARPS * 4;  // [slot][M,S,T,mx,mn,xy][warp]
  static constexpr int kBytes = kCntOff + 2 * kSlots * 4;            // cnt[kSlots], gen[kSlots]
};

// A warp's slice of every row it visits, and the TMA ring that streams it.
// The vocabulary is cut into warp-rows of 32 16-byte vectors; warp w owns a contiguous run of
// warp-rows (balanced to within one), so only the very last warp-row of the vocabulary can be
// ragged. Rows of the [T, V] matrix are visited round-robin over CTAs.
//
// PASSES > 1 streams every row that many times back to back (the fused GRPO kernel reads a row for
// the forward statistics and immediately again, out of L2, for the backward): all passes but the
// last load with the normal L2 policy so the lines stay resident, the last with evict-first.
// SKEW (with PASSES == 2) orders the stream as A0, A1, B0, A2, B1, ... (A = first pass, B = second
// pass of a row): the second pass of a row comes one row later, by which time whatever had to be
// computed between the passes is long done.
template <int WARPS, int STAGES, int PASSES = 1, bool SKEW = false>
struct WarpStream {
  // static geometry
  int sv, seg_len;        // first vector / number of vectors of this warp's slice
  int64_t row_jump;       // bytes from the end of this slice in row r to its start in row r+grid
  // producer (lane 0 issues, every lane tracks the same state)
  const char* p_src;
  int p_left, p_rows_left, p_pass;
  const char *a_ptr, *b_ptr;   // SKEW: slice start of the next first-pass / second-pass row
  int a_left, a_done;          // SKEW: first-pass rows still to queue / already queued
  int64_t row_stride;
  uint32_t p_dst, p_bar, ring_dst0, ring_bar0;
  int p_stage;
  uint64_t pol, pol_keep;
  // consumer
  uint32_t c_dst, c_bar;
  int c_stage;
  uint32_t c_par;

  __device__ __forceinline__ void init(const void* base, int64_t ld_bytes, int nvec, int nrows,
                                       unsigned char* my_data, uint64_t* my_bar, int warp,
                                       int lane) {
    init_seg(base, ld_bytes, nvec, nrows, my_data, my_bar, warp, WARPS, (int)blockIdx.x,
             (int)gridDim.x, lane);
  }

  // General form: this warp streams slice `seg` of `nsegs` of the rows first_row,
  // first_row + row_step, ... (`nrows` of them).
  __device__ __forceinline__ void init_seg(const void* base, int64_t ld_bytes, int nvec, int nrows,
                                           unsigned char* my_data, uint64_t* my_bar, int seg,
                                           int nsegs, int first_row, int row_step, int lane) {
    const int R = (nvec + 31) >> 5;
    const int q = R / nsegs, rem = R % nsegs;
    const int my_rows = q + (seg < rem ? 1 : 0);
    sv = (seg * q + min(seg, rem)) * 32;
    seg_len = max(0, min(my_rows * 32, nvec - sv));
    row_jump = (int64_t)row_step * ld_bytes - (int64_t)seg_len * 16;
    p_src = reinterpret_cast<const char*>(base) + (int64_t)first_row * ld_bytes + (int64_t)sv * 16;
    p_left = seg_len;
    p_pass = 0;
    p_rows_left = seg_len > 0 ? nrows : 0;
    row_stride = (int64_t)row_step * ld_bytes;
    a_ptr = b_ptr = p_src;
    a_left = nrows;
    a_done = 0;
    ring_dst0 = smem_u32(my_data);
    ring_bar0 = smem_u32(my_bar);
    p_dst = c_dst = ring_dst0;
    p_bar = c_bar = ring_bar0;
    p_stage = c_stage = 0;
    c_par = 0;
    pol = policy_evict_first();
    pol_keep = pol;
    if (PASSES > 1) asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol_keep));
    if (lane == 0) {
#pragma unroll
      for (int s = 0; s < STAGES; s++) mbar_init(&my_bar[s], 1);
      fence_barrier_init();
    }
    __syncwarp();
#pragma unroll
    for (int s = 0; s < STAGES; s++) issue(lane);
  }

  // Queue the next chunk of the slice into the next ring slot (no-op once all rows are queued).
  __device__ __forceinline__ void issue(int lane) {
    if (p_rows_left > 0) {
      const int nv = min(kChunkVec, p_left);
      const uint32_t bytes = (uint32_t)nv * 16u;
      if (lane == 0) {
        const uint64_t policy = (PASSES > 1 && p_pass < PASSES - 1) ? pol_keep : pol;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(p_bar), "r"(bytes)
                     : "memory");
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
            "[%0], [%1], %2, [%3], %4;" ::"r"(p_dst), "l"(p_src), "r"(bytes), "r"(p_bar), "l"(policy)
            : "memory");
      }
      p_src += bytes;
      p_left -= nv;
      if (p_left == 0) {
        p_left = seg_len;
        if (SKEW) {
          if (p_pass == 0) {  // finished a first pass
            a_ptr += row_stride;
            a_left--;
            a_done++;
            p_pass = (a_done >= 2 || a_left == 0) ? 1 : 0;
          } else {            // finished a second pass: that row is complete
            b_ptr += row_stride;
            p_rows_left--;
            p_pass = a_left > 0 ? 0 : 1;
          }
          p_src = p_pass == 0 ? a_ptr : b_ptr;
        } else if (PASSES > 1 && ++p_pass < PASSES) {
          p_src -= (int64_t)seg_len * 16;  // same slice of the same row again
        } else {
          p_pass = 0;
          p_src += row_jump;
          p_rows_left--;
        }
      }
      p_dst += kChunkBytes;
      p_bar += 8;
      if (++p_stage == STAGES) {
        p_stage = 0;
        p_dst = ring_dst0;
        p_bar = ring_bar0;
      }
    }
  }

  // Wait until the current consumer slot holds its chunk; returns its shared address.
  __device__ __forceinline__ uint32_t wait() {
    uint32_t ok;
    do {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(ok)
          : "r"(c_bar), "r"(c_par)
          : "memory");
    } while (!ok);
    return c_dst;
  }

  // The warp has read the current slot into registers: hand it back to the producer.
  // (generic-proxy reads followed by an async-proxy overwrite need no proxy fence; the
  // __syncwarp orders every lane's reads before lane 0 issues the copy)
  __device__ __forceinline__ void release(int lane) {
    __syncwarp();
    issue(lane);
    c_dst += kChunkBytes;
    c_bar += 8;
    if (++c_stage == STAGES) {
      c_stage = 0;
      c_dst = ring_dst0;
      c_bar = ring_bar0;
      c_par ^= 1;
    }
  }
};

template <typename E>
struct LaneState {
  float negm, s0, s1, t0, t1;
  typename E::MaxT pmx, pmn;
};

template <typename E, bool ENT>
__device__ __forceinline__ void lazy_rescale(LaneState<E>& L, typename E::MaxT cm, float scale) {
  const float cmf = E::max_f(cm) * scale;
  if (cmf + L.negm > kRescaleThr) {
    const float d = cmf + L.negm;  // new reference - old reference > 0
    const float f = ex2_approx(-d);
    const float s = L.s0 + L.s1, t = L.t0 + L.t1;
    if (ENT) {
      L.t0 = f * (t - d * s);
      L.t1 = 0.f;
    }
    L.s0 = f * s;
    L.s1 = 0.f;
    L.negm = -cmf;
  }
}

template <typename E, bool ENT>
__device__ __forceinline__ void accumulate_vec(LaneState<E>& L, const uint4& v, float scale) {
  float x[E::kPerVec];
  E::unpack(v, x);
  // pairs of elements through FFMA2 / FADD2 (same roundings as the scalar forms): per pair one
  // FFMA2 (y), two MUFU.EX2, one FADD2 (s) and, with the entropy, one FFMA2 (t)
  const float2 sc2 = make_float2(scale, scale), nm2 = make_float2(L.negm, L.negm);
  float2 s = make_float2(L.s0, L.s1), t = make_float2(L.t0, L.t1);
#pragma unroll
  for (int j = 0; j < E::kPerVec; j += 2) {
    const float2 y = ffma2(make_float2(x[j], x[j + 1]), sc2, nm2);
    const float2 e = make_float2(ex2_approx(y.x), ex2_approx(y.y));
    s = fadd2(s, e);
    if (ENT) t = ffma2(e, y, t);
  }
  L.s0 = s.x;
  L.s1 = s.y;
  if (ENT) {
    L.t0 = t.x;
    L.t1 = t.y;
  }
}


// One 16-byte vector of the backward pass.
template <typename E, bool HAS_GH>
__device__ __forceinline__ void bwd_vec(const uint4& v, int vi, bool act, float scale, float nlse2,
                                        float A0, float B0, float lab_add, int lvec, int lsub,
                                        char* out_row) {
  float x[E::kPerVec], d[E::kPerVec];
  E::unpack(v, x);
#pragma unroll
  for (int j = 0; j < E::kPerVec; j++) {
    const float y = fmaf(x[j], scale, nlse2);
    const float p = ex2_approx(y);
    d[j] = HAS_GH ? p * fmaf(B0, y, A0) : p * A0;
  }
  if (vi == lvec) {  // the one-hot term of the label, one lane once per row
#pragma unroll
    for (int k = 0; k < E::kPerVec; k++)
      if (k == lsub) d[k] += lab_add;
  }
  // (an evict-first L2 policy on these stores was measured: no change, the write-back of the
  // dirty tail only moves between this kernel and the next one)
  if (act) stg_na_v4(out_row + (int64_t)vi * 16, E::pack(d));
}

// The same vector when the entropy carries no gradient (the GRPO call site): d = p * A0 with a
// per-row constant A0, so the multiplication is folded into the exponent,
//   p * A0 = sign(A0) * 2^(x*scale + nlse2 + log2|A0|),
// and the sign is one XOR per packed word: one FFMA + one MUFU.EX2 per element and half a LOP,
// instead of FFMA + MUFU + FMUL. `nlse2f` = nlse2 + log2|A0| (-inf when A0 == 0: masked rows give
// exact zeros), `sgn` = the dtype's packed sign mask when A0 < 0, else 0 (bwd_row_consts).
template <typename E>
__device__ __forceinline__ void bwd_vec_folded(const uint4& v, int vi, bool act, float scale,
                                               float nlse2f, uint32_t sgn, float lab_add, int lvec,
                                               int lsub, char* out_row) {
  float x[E::kPerVec], d[E::kPerVec];
  E::unpack(v, x);
  const float2 sc2 = make_float2(scale, scale), nl2 = make_float2(nlse2f, nlse2f);
#pragma unroll
  for (int j = 0; j < E::kPerVec; j += 2) {
    const float2 y = ffma2(make_float2(x[j], x[j + 1]), sc2, nl2);
    d[j] = ex2_approx(y.x);
    d[j + 1] = ex2_approx(y.y);
  }
  if (vi == lvec) {  // the one-hot term of the label, one lane once per row: sign first, in fp32
#pragma unroll
    for (int k = 0; k < E::kPerVec; k++) {
      if (sgn) d[k] = -d[k];
      if (k == lsub) d[k] += lab_add;
    }
    if (act) stg_na_v4(out_row + (int64_t)vi * 16, E::pack(d));
    return;
  }
  uint4 o = E::pack(d);
  o.x ^= sgn;
  o.y ^= sgn;
  o.z ^= sgn;
  o.w ^= sgn;
  if (act) stg_na_v4(out_row + (int64_t)vi * 16, o);
}

// roll != 0: `labels` holds the packed input_ids and the label of row t is the NEXT token,
// labels[(t + 1) % T] = torch.roll(input_ids, -1)[t] (ref areal/engine/ppo/actor.py:308)
__device__ __forceinline__ int label_row(int row, int T, int roll) {
  return roll ? (row + 1 == T ? 0 : row + 1) : row;
}

// Vocabulary-parallel combine of one row: folds the n_shards 8-float records
// (M, S, W, x_y, max, min, -, -) with the algebra the warps use inside a row. `base` is
// [n_shards][shard_stride_rows][2] float4; shard k's record of `row` is base[(k*stride + row)*2].
__device__ __forceinline__ void combine_shard_row(const float4* base, int n_shards,
                                                  int64_t shard_stride_rows, int64_t row, float scale,
                                                  float* __restrict__ logp, float* __restrict__ ent,
                                                  float* __restrict__ lse, float* __restrict__ rowmax,
                                                  float* __restrict__ rowmin) {
  float M = -INFINITY;
  // __ldcg: straight from L2 - the records may have been written by other GPUs moments ago
  for (int k = 0; k < n_shards; k++) M = fmaxf(M, __ldcg(base + ((int64_t)k * shard_stride_rows + row) * 2).x);
  float S = 0.f, W = 0.f, xy = __int_as_float(0x7fc00000), mx = -INFINITY, mn = INFINITY;
  for (int k = 0; k < n_shards; k++) {
    const float4 a = __ldcg(base + ((int64_t)k * shard_stride_rows + row) * 2);
    const float4 b = __ldcg(base + ((int64_t)k * shard_stride_rows + row) * 2 + 1);
    const float d = a.x - M;  // <= 0
    const float f = ex2_approx(d);
    S = fmaf(a.y, f, S);
    W = fmaf(f, fmaf(d, a.y, a.z), W);
    if (a.w == a.w) xy = a.w;  // the one shard that owns the label column
    mx = fmaxf(mx, b.x);
    mn = fminf(mn, b.y);
  }
  const float l2s = log2f(S);
  lse[row] = (M + l2s) * kLn2;
  logp[row] = (fmaf(xy, scale, -M) - l2s) * kLn2;
  if (ent) ent[row] = kLn2 * (l2s - W / S);
  if (rowmax) rowmax[row] = mx;
  if (rowmin) rowmin[row] = mn;
}

// per-row constants of the folded form
template <typename E>
__device__ __forceinline__ void bwd_row_consts(float nlse2, float A0, float& nlse2f, uint32_t& sgn) {
  nlse2f = nlse2 + lg2_approx(fabsf(A0));
  sgn = A0 < 0.f ? E::kSignWord : 0u;
}

}  // namespace lp
}  // namespace b200rl
