// Vocabulary-parallel exchange without a collective library call.
//
// ref: areal/experimental/utils/mcore/functional.py:8-52 and megatron_actor.py:60-66, 323-331 do
// three + three all-reduces over the tensor-parallel group per call. Round 1 here: one streaming
// pass per shard, ONE NCCL all_gather of the [T, 8] records, one combine launch; at tp = 8 and
// micro-batches of a few thousand tokens that step was bound by the all-gather's launch and the
// host work around it (0.53 ms per forward+backward of which the kernels need 0.15 ms).
// Here the 32-byte records travel as plain peer stores over NVLink into a buffer that every rank
// of the group has mapped (CUDA IPC), and the combine kernel itself waits for them:
//
//   tp_push_kernel      reads this rank's [T, 8] records and stores them into slot `epoch & 1` of
//                       EVERY rank's exchange buffer (its own included), then - after a
//                       system-scope fence - writes `epoch` into its per-CTA flag on every rank.
//   tp_combine_kernel   CTA b covers the same 256 rows as push CTA b of every rank, so it polls
//                       just those tp flags until they hold `epoch` (bounded: after ~2 s it sets
//                       *status and carries on, it can not hang the GPU), then folds the tp
//                       records of its rows exactly like logprob_shard_combine_kernel.
//
// Two slots are enough: a rank can be at most one call ahead of a peer (its own combine of call
// e needs the peer's push of call e, which the peer issues after its combine of call e-1).
// Buffer layout (b200rl_tp_exchange_bytes): flags [2][kMaxTp][kMaxPushCtas] u32, then records
// [2][tp][max_tokens][8] f32. The buffer must be zero-filled once before the first call.
#include "logprob_common.cuh"

namespace b200rl {
namespace {

using namespace lp;

constexpr int kMaxTp = 16;
constexpr int kPushThreads = 256;
constexpr int kMaxPushCtas = 1024;  // rows per call <= 262,144
constexpr size_t kFlagBytes = (size_t)2 * kMaxTp * kMaxPushCtas * sizeof(uint32_t);

struct Peers {
  unsigned char* base[kMaxTp];
};

__device__ __forceinline__ uint32_t* flag_ptr(unsigned char* buf, int slot, int producer, int cta) {
  return reinterpret_cast<uint32_t*>(buf) + ((size_t)slot * kMaxTp + producer) * kMaxPushCtas + cta;
}
__device__ __forceinline__ float4* record_ptr(unsigned char* buf, int slot, int tp, int64_t max_tokens,
                                              int producer, int64_t row) {
  return reinterpret_cast<float4*>(buf + kFlagBytes) +
         (((size_t)slot * tp + producer) * (size_t)max_tokens + (size_t)row) * 2;
}

__global__ void __launch_bounds__(kPushThreads)
tp_push_kernel(const float* __restrict__ partial, int64_t T, Peers peers, int tp, int rank,
               int64_t max_tokens, uint32_t epoch) {
  const int slot = epoch & 1;
  const int64_t row = (int64_t)blockIdx.x * kPushThreads + threadIdx.x;
  if (row < T) {
    const float4* src = reinterpret_cast<const float4*>(partial) + row * 2;
    const float4 a = src[0], b = src[1];
    for (int p = 0; p < tp; p++) {
      float4* dst = record_ptr(peers.base[p], slot, tp, max_tokens, rank, row);
      dst[0] = a;  // peer stores over NVLink (a local store for p == rank)
      dst[1] = b;
    }
  }
  __threadfence_system();  // the records are visible system-wide before the flag is
  __syncthreads();
  if (threadIdx.x < tp) {
    uint32_t* f = flag_ptr(peers.base[threadIdx.x], slot, rank, blockIdx.x);
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(epoch) : "memory");
  }
}

__global__ void __launch_bounds__(256)
tp_combine_kernel(unsigned char* mine, int tp, int64_t T, int64_t max_tokens, uint32_t epoch,
                  float scale, float* __restrict__ logp, float* __restrict__ ent,
                  float* __restrict__ lse, float* __restrict__ rowmax, float* __restrict__ rowmin,
                  int* __restrict__ status) {
  const int slot = epoch & 1;
  static_assert(kPushThreads == 256, "push CTA b and combine CTA b cover the same rows");
  // ---- wait for push CTA `blockIdx.x` of every rank (bounded) --------------------------------
  bool late = false;
  if (threadIdx.x < tp) {
    const long long t0 = clock64();
    const uint32_t* f = flag_ptr(mine, slot, threadIdx.x, blockIdx.x);
    uint32_t v;
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
      if (v == epoch) break;
      if (clock64() - t0 > 4000000000LL) {  // ~2 s at 1.9 GHz: a peer is gone; do not hang the GPU
        late = true;
        break;
      }
      __nanosleep(64);
    } while (true);
  }
  if (late) atomicExch(status, 1);
  __syncthreads();
  const int64_t row = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (row >= T) return;
  const float4* base = reinterpret_cast<const float4*>(mine + kFlagBytes) + (size_t)slot * tp * (size_t)max_tokens * 2;
  combine_shard_row(base, tp, max_tokens, row, scale, logp, ent, lse, rowmax, rowmin);
}

}  // namespace
}  // namespace b200rl

extern "C" {

size_t b200rl_tp_exchange_bytes(int32_t tp, int64_t max_tokens) {
  using namespace b200rl;
  if (tp < 1 || tp > kMaxTp || max_tokens < 1) return 0;
  return kFlagBytes + (size_t)2 * tp * (size_t)max_tokens * 8 * sizeof(float);
}

int b200rl_tp_push(const float* partial, int64_t T, void* const* peer_bufs, int32_t tp, int32_t rank,
                   int64_t max_tokens, uint32_t epoch, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(tp >= 1 && tp <= kMaxTp && rank >= 0 && rank < tp, "bad tp/rank %d/%d", tp, rank);
  B200RL_REQUIRE(T >= 0 && T <= max_tokens, "T %lld exceeds the exchange buffer's %lld rows", (long long)T,
                 (long long)max_tokens);
  B200RL_REQUIRE(epoch != 0, "epoch must start at 1 (the buffer is zero-filled)");
  if (T == 0) return B200RL_OK;
  B200RL_REQUIRE(partial && peer_bufs, "null pointer");
  const int64_t ctas = (T + kPushThreads - 1) / kPushThreads;
  B200RL_REQUIRE(ctas <= kMaxPushCtas, "at most %d rows per call", kMaxPushCtas * kPushThreads);
  Peers P;
  for (int p = 0; p < kMaxTp; p++) P.base[p] = p < tp ? static_cast<unsigned char*>(peer_bufs[p]) : nullptr;
  for (int p = 0; p < tp; p++) B200RL_REQUIRE(P.base[p] != nullptr, "peer buffer %d is null", p);
  tp_push_kernel<<<(unsigned)ctas, kPushThreads, 0, (cudaStream_t)stream>>>(partial, T, P, tp, rank, max_tokens, epoch);
  return check_launch("tp_push_kernel");
}

int b200rl_tp_combine(void* my_buf, int32_t tp, int64_t T, int64_t max_tokens, uint32_t epoch,
                      float temperature, float* logp, float* entropy, float* lse, float* rowmax,
                      float* rowmin, int32_t* status, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(tp >= 1 && tp <= kMaxTp, "bad tp %d", tp);
  B200RL_REQUIRE(T >= 0 && T <= max_tokens, "T %lld exceeds the exchange buffer's %lld rows", (long long)T,
                 (long long)max_tokens);
  B200RL_REQUIRE(temperature > 0.f, "temperature must be > 0, got %f", temperature);
  if (T == 0) return B200RL_OK;
  B200RL_REQUIRE(my_buf && logp && lse && status, "null pointer");
  const float scale = (float)(1.4426950408889634 / (double)temperature);
  B200RL_REQUIRE((T + 255) / 256 <= kMaxPushCtas, "at most %d rows per call", kMaxPushCtas * kPushThreads);
  tp_combine_kernel<<<(unsigned)((T + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      static_cast<unsigned char*>(my_buf), tp, T, max_tokens, epoch, scale, logp, entropy, lse, rowmax, rowmin,
      status);
  return check_launch("tp_combine_kernel");
}

}  // extern "C"
