// H4 — interval gather/scatter on a flat parameter buffer (weight resharding).
//
// Replaces csrc/interval_op/interval_op.cu:12-114 of the reference, whose copyDataKernel gives
// every *thread* a serial <=4 KiB memcpy of its own interval chunk (uncoalesced, grid sized by
// the largest interval). Here the PACKED side of the copy is cut into fixed 16 KiB tiles, one
// warp per tile, so the launch is balanced whatever the interval-size skew is; inside a tile the
// warp walks the intervals that overlap it and moves each piece with 16-byte vector accesses
// aligned on the store side (narrower vectors when source and destination are mutually
// misaligned). The per-interval packed offsets (exclusive prefix sum of the lengths, which the
// reference builds with three aten ops per call, interval_op.cu:49-51,84-86) and the
// tile -> first-interval map form a "plan" kept in caller-owned workspace and reusable across
// calls with the same interval list.
//
// Bound: HBM. Algorithmic bytes = 2 * elem_size per copied element (+16 B per interval).
#include "common.cuh"
#include <string.h>

namespace b200rl {
namespace {

#ifndef B200RL_IV_TILE
#define B200RL_IV_TILE 16384
#endif
#ifndef B200RL_IV_WARPS
#define B200RL_IV_WARPS 8
#endif
#ifndef B200RL_IV_OCC
#define B200RL_IV_OCC 4
#endif
#ifndef B200RL_IV_U
#define B200RL_IV_U 4
#endif
#ifndef B200RL_IV_SHIFT_MIN
#define B200RL_IV_SHIFT_MIN 512  // bytes: shorter misaligned pieces take the narrow-vector movers
                                 // (measured: 4096 -> 512 lifts lists of short odd intervals from 0.39 to 0.46 of peak)
#endif
#ifndef B200RL_IV_US
#define B200RL_IV_US 4
#endif
#ifndef B200RL_IV_TILE_MAX
#define B200RL_IV_TILE_MAX 32768
#endif
constexpr int kTileBytes = B200RL_IV_TILE;   // smallest packed bytes per warp tile (workspace sizing)
constexpr int kTileBytesMax = B200RL_IV_TILE_MAX;
constexpr int kCopyWarps = B200RL_IV_WARPS;  // warps per CTA in the copy kernel
constexpr int kCopyOcc = B200RL_IV_OCC;      // CTAs per SM the copy kernels are compiled for
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;         // intervals per thread in the local scan
constexpr int kScanBlock = kScanThreads * kScanItems;  // 2048 intervals per CTA

struct PlanLayout {
  int64_t* hdr;         // [0] total elems, [1] error flag, [2] n, [3] packed_numel
  int64_t* offsets;     // [n+1]
  int64_t* block_sums;  // [nb]
  int32_t* tile_first;  // [num_tiles+1]
  int64_t nb, num_tiles;
  size_t bytes;
};

// Tile size: 16 KiB keeps small copies parallel; large copies take bigger tiles so that the
// per-tile start-up (three dependent metadata loads) is amortised. A pure function of the packed
// size, so a cached plan and a later call agree.
__host__ int pick_tile_bytes(int64_t packed_bytes) {
  int tile = kTileBytes;
  while (tile < kTileBytesMax && packed_bytes / (2 * (int64_t)tile) >= (int64_t)kNumSMs * 64) tile *= 2;
  return tile;
}

__host__ PlanLayout plan_layout(void* ws, int64_t n, int64_t packed_numel, int elem_size,
                                int tile_bytes = kTileBytes) {
  PlanLayout L;
  L.nb = (n + kScanBlock - 1) / kScanBlock;
  if (L.nb < 1) L.nb = 1;
  int64_t packed_bytes = packed_numel * (int64_t)elem_size;
  L.num_tiles = (packed_bytes + tile_bytes - 1) / tile_bytes;
  char* p = (char*)ws;
  size_t off = 0;
  L.hdr = (int64_t*)(p + off);
  off += 4 * sizeof(int64_t);
  L.offsets = (int64_t*)(p + off);
  off += (size_t)(n + 1) * sizeof(int64_t);
  L.block_sums = (int64_t*)(p + off);
  off += (size_t)L.nb * sizeof(int64_t);
  L.tile_first = (int32_t*)(p + off);
  off += (size_t)(L.num_tiles + 1) * sizeof(int32_t);
  L.bytes = (off + 255) & ~(size_t)255;
  return L;
}

// ---- plan kernels ----------------------------------------------------------
// K1: lengths + validation + CTA-local exclusive scan.
__global__ void __launch_bounds__(kScanThreads)
interval_local_scan(const int64_t* __restrict__ intervals, int64_t n, int64_t big_numel,
                    int64_t* __restrict__ offsets, int64_t* __restrict__ block_sums,
                    int64_t* __restrict__ hdr) {
  __shared__ int64_t warp_tot[kScanThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t base = (int64_t)blockIdx.x * kScanBlock + (int64_t)tid * kScanItems;
  int64_t len[kScanItems];
  int64_t tsum = 0;
  bool bad = false;
#pragma unroll
  for (int k = 0; k < kScanItems; k++) {
    int64_t i = base + k;
    int64_t l = 0;
    if (i < n) {
      // one 16-B load per interval: {start, end}
      longlong2 se = *reinterpret_cast<const longlong2*>(intervals + 2 * i);
      l = se.y - se.x;
      if (se.x < 0 || l < 0 || se.y > big_numel) {
        bad = true;
        l = 0;  // malformed intervals copy nothing
      }
    }
    len[k] = l;
    tsum += l;
  }
  // warp inclusive scan of per-thread sums
  int64_t inc = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int64_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  int64_t wbase = 0;
#pragma unroll
  for (int w = 0; w < kScanThreads / 32; w++)
    if (w < warp) wbase += warp_tot[w];
  int64_t run = wbase + inc - tsum;
#pragma unroll
  for (int k = 0; k < kScanItems; k++) {
    int64_t i = base + k;
    if (i < n) offsets[i] = run;
    run += len[k];
  }
  if (tid == kScanThreads - 1) block_sums[blockIdx.x] = run;
  if (bad) hdr[1] = 1;
}

// K2: exclusive scan of the per-CTA sums (single CTA) + header.
__global__ void __launch_bounds__(1024)
interval_block_scan(int64_t* __restrict__ block_sums, int64_t nb, int64_t* __restrict__ hdr,
                    int64_t n, int64_t packed_numel) {
  __shared__ int64_t part[1024];
  const int tid = threadIdx.x;
  const int64_t per = (nb + 1023) / 1024;
  const int64_t lo = tid * per, hi = min(lo + per, nb);
  int64_t s = 0;
  for (int64_t i = lo; i < hi; i++) s += block_sums[i];
  part[tid] = s;
  __syncthreads();
  // Hillis-Steele over 1024 partials
  for (int o = 1; o < 1024; o <<= 1) {
    int64_t v = (tid >= o) ? part[tid - o] : 0;
    __syncthreads();
    part[tid] += v;
    __syncthreads();
  }
  int64_t run = part[tid] - s;
  for (int64_t i = lo; i < hi; i++) {
    int64_t v = block_sums[i];
    block_sums[i] = run;
    run += v;
  }
  if (tid == 1023) {
    hdr[0] = part[1023];
    hdr[2] = n;
    hdr[3] = packed_numel;
    if (part[1023] != packed_numel) hdr[1] |= 2;  // sum of lengths != packed size
  }
}

// K3: add the CTA base, write offsets[n] = total.
__global__ void __launch_bounds__(256)
interval_add_base(int64_t* __restrict__ offsets, const int64_t* __restrict__ block_sums,
                  const int64_t* __restrict__ hdr, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < n) offsets[i] += block_sums[i / kScanBlock];
  if (i == n) offsets[n] = hdr[0];
}

// K4: tile_first[t] = last interval whose packed offset is <= the tile's first element
// (upper_bound - 1, which skips empty intervals).
__global__ void __launch_bounds__(256)
interval_partition(const int64_t* __restrict__ offsets, int64_t n, int32_t* __restrict__ tile_first,
                   int64_t num_tiles, int elem_shift, int tile_bytes) {
  int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t > num_tiles) return;
  if (t == num_tiles) {
    tile_first[t] = (int32_t)n;
    return;
  }
  const int64_t pos = (t * (int64_t)tile_bytes) >> elem_shift;  // first packed element of tile
  int64_t lo = 0, hi = n + 1;  // upper_bound over offsets[0..n]
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (offsets[mid] <= pos) lo = mid + 1; else hi = mid;
  }
  int64_t idx = lo - 1;
  if (idx < 0) idx = 0;
  if (idx > n - 1) idx = n - 1;
  tile_first[t] = (int32_t)idx;
}

// ---- copy ------------------------------------------------------------------
template <int W> struct VecT;
template <> struct VecT<16> { using type = uint4; };
template <> struct VecT<8> { using type = uint2; };
template <> struct VecT<4> { using type = uint32_t; };
template <> struct VecT<2> { using type = uint16_t; };
template <> struct VecT<1> { using type = uint8_t; };

template <int W>
__device__ __forceinline__ typename VecT<W>::type ld_stream(const char* p) {
  if constexpr (W == 16) {
    return ldg_nc_v4(p);
  } else {
    return __ldg(reinterpret_cast<const typename VecT<W>::type*>(p));
  }
}
template <int W>
__device__ __forceinline__ void st_stream(char* p, typename VecT<W>::type v) {
  if constexpr (W == 16) {
    stg_na_v4(p, v);
  } else {
    *reinterpret_cast<typename VecT<W>::type*>(p) = v;
  }
}

// element-granular copy of `bytes` (< 32*ES*k small) by the warp
template <int ES>
__device__ __forceinline__ void copy_small(const char* s, char* d, int bytes, int lane) {
  for (int o = lane * ES; o < bytes; o += 32 * ES) st_stream<ES>(d + o, ld_stream<ES>(s + o));
}

// Warp-cooperative copy of `bytes` from s to d using W-byte accesses on the body.
// Requires (s - d) % W == 0 and ES | W; d, s are ES-aligned.
template <int W, int ES>
__device__ __forceinline__ void copy_piece_w(const char* s, char* d, int64_t bytes, int lane) {
  if constexpr (W > ES) {
    int head = (int)((W - (reinterpret_cast<uintptr_t>(d) & (W - 1))) & (W - 1));
    if (head > bytes) head = (int)bytes;
    copy_small<ES>(s, d, head, lane);
    s += head;
    d += head;
    bytes -= head;
  }
  const int64_t nv = bytes / W;
  constexpr int U = B200RL_IV_U;
  int64_t v = lane;
  for (; v + 32 * (U - 1) < nv; v += 32 * U) {
    typename VecT<W>::type r[U];
#pragma unroll
    for (int u = 0; u < U; u++) r[u] = ld_stream<W>(s + (v + 32 * u) * W);
#pragma unroll
    for (int u = 0; u < U; u++) st_stream<W>(d + (v + 32 * u) * W, r[u]);
  }
  for (; v < nv; v += 32) st_stream<W>(d + v * W, ld_stream<W>(s + v * W));
  if constexpr (W > ES) {
    int tail = (int)(bytes - nv * W);
    copy_small<ES>(s + nv * W, d + nv * W, tail, lane);
  }
}


// Mutually misaligned piece ((s - d) % 16 != 0) moved with 16-byte accesses on BOTH sides: stores
// are aligned on d; every lane loads the aligned source vector under its store, takes the next one
// from its neighbour lane by shuffle, and funnel-shifts the 32-byte window by the (warp-uniform)
// byte phase. Reads stay inside [floor16(s'), s' + body + 16), i.e. inside the interval except for
// < 16 bytes before it, which is still inside the (16-byte aligned) source allocation.
template <int K>
__device__ __forceinline__ uint4 shift_window(const uint4& lo, const uint4& hi, unsigned bits) {
  const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
  uint4 o;
  o.x = __funnelshift_r(w[K + 0], w[K + 1], bits);
  o.y = __funnelshift_r(w[K + 1], w[K + 2], bits);
  o.z = __funnelshift_r(w[K + 2], w[K + 3], bits);
  o.w = __funnelshift_r(w[K + 3], w[K + 4], bits);
  return o;
}

template <int ES>
__device__ __forceinline__ void copy_piece_shift(const char* s, char* d, int64_t bytes, int lane) {
  int head = (int)((16 - (reinterpret_cast<uintptr_t>(d) & 15)) & 15);
  if (head > bytes) head = (int)bytes;
  copy_small<ES>(s, d, head, lane);
  s += head;
  d += head;
  bytes -= head;
  const unsigned ph = (unsigned)(reinterpret_cast<uintptr_t>(s) & 15);  // 1..15, warp-uniform
  const char* a = s - ph;                                               // 16-B aligned
  const int64_t nv = bytes >= 32 ? (bytes - 16) / 16 : 0;               // keep the over-read inside
  const unsigned k = ph >> 2, bits = (ph & 3) * 8;
  constexpr int U = B200RL_IV_US;
  for (int64_t v0 = 0; v0 < nv; v0 += 32 * U) {
    uint4 lo[U], ex[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t v = v0 + 32 * u + lane;
      lo[u] = make_uint4(0, 0, 0, 0);
      ex[u] = make_uint4(0, 0, 0, 0);
      if (v <= nv) lo[u] = ldg_nc_v4(a + v * 16);  // vector nv is the last window's upper half
      if (lane == 31 && v < nv) ex[u] = ldg_nc_v4(a + (v + 1) * 16);  // neighbour of the last lane
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t v = v0 + 32 * u + lane;
      uint4 hi;
      hi.x = __shfl_down_sync(0xffffffffu, lo[u].x, 1);
      hi.y = __shfl_down_sync(0xffffffffu, lo[u].y, 1);
      hi.z = __shfl_down_sync(0xffffffffu, lo[u].z, 1);
      hi.w = __shfl_down_sync(0xffffffffu, lo[u].w, 1);
      if (lane == 31) hi = ex[u];
      uint4 o;
      switch (k) {
        case 0: o = shift_window<0>(lo[u], hi, bits); break;
        case 1: o = shift_window<1>(lo[u], hi, bits); break;
        case 2: o = shift_window<2>(lo[u], hi, bits); break;
        default: o = shift_window<3>(lo[u], hi, bits); break;
      }
      if (v < nv) stg_na_v4(d + v * 16, o);
    }
  }
  copy_small<ES>(s + nv * 16, d + nv * 16, (int)(bytes - nv * 16), lane);
}

template <int ES>
__device__ __forceinline__ void copy_piece(const char* s, char* d, int64_t bytes, int lane, bool src_base_aligned) {
  const unsigned rel = (unsigned)((reinterpret_cast<uintptr_t>(s) - reinterpret_cast<uintptr_t>(d)) & 15u);
  if (rel == 0) {
    copy_piece_w<16, ES>(s, d, bytes, lane);
  } else if (src_base_aligned && bytes >= B200RL_IV_SHIFT_MIN) {
    copy_piece_shift<ES>(s, d, bytes, lane);
  } else if constexpr (ES < 8) {
    if ((rel & 7u) == 0) {
      copy_piece_w<8, ES>(s, d, bytes, lane);
    } else if constexpr (ES < 4) {
      if ((rel & 3u) == 0) {
        copy_piece_w<4, ES>(s, d, bytes, lane);
      } else if constexpr (ES < 2) {
        if ((rel & 1u) == 0) copy_piece_w<2, ES>(s, d, bytes, lane);
        else copy_piece_w<1, ES>(s, d, bytes, lane);
      } else {
        copy_piece_w<2, ES>(s, d, bytes, lane);
      }
    } else {
      copy_piece_w<4, ES>(s, d, bytes, lane);
    }
  } else {
    copy_piece_w<8, ES>(s, d, bytes, lane);
  }
}

// MODE 0: slice (big -> packed).  MODE 1: set (packed -> big).
template <int MODE, int ES>
__global__ void __launch_bounds__(kCopyWarps * 32, kCopyOcc)
interval_copy_kernel(const char* __restrict__ src, char* __restrict__ dst,
                     const int64_t* __restrict__ intervals, const int64_t* __restrict__ offsets,
                     const int32_t* __restrict__ tile_first, const int64_t* __restrict__ hdr,
                     int64_t n, int64_t num_tiles, int tile_bytes) {
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * kCopyWarps + (threadIdx.x >> 5);
  if (tile >= num_tiles) return;
  const bool src_ok = (reinterpret_cast<uintptr_t>(src) & 15) == 0;  // allows aligned over-reads
  int64_t total_elems = hdr[0];
  if (total_elems > hdr[3]) total_elems = hdr[3];  // never write past the packed buffer
  const int64_t total_bytes = total_elems * ES;
  int64_t p0 = tile * (int64_t)tile_bytes;
  int64_t p1 = p0 + tile_bytes;
  if (p1 > total_bytes) p1 = total_bytes;
  int64_t i = tile_first[tile];
  // (requesting interval i+1's metadata before interval i's bytes are moved was measured on the
  // list of 200k short odd intervals: 183 us vs 178 us, no gain - those tiles are bound by the
  // per-piece head/body/tail sequence, not by the offsets -> intervals load chain)
  while (p0 < p1 && i < n) {
    const int64_t ob = offsets[i] * ES;
    const int64_t oe = offsets[i + 1] * ES;
    if (oe <= p0) {  // empty interval (or one that ends before this tile)
      i++;
      continue;
    }
    const int64_t start = intervals[2 * i] * ES;
    const int64_t b = oe < p1 ? oe : p1;
    const int64_t big_off = start + (p0 - ob);
    if (MODE == 0) copy_piece<ES>(src + big_off, dst + p0, b - p0, lane, src_ok);
    else           copy_piece<ES>(src + p0, dst + big_off, b - p0, lane, src_ok);
    p0 = b;
    if (b == oe) i++;
  }
}

// MOVE: big -> big without the packed staging buffer (the kernel the reference's TODO at
// real_llm_api.py:719-720 asks for). Two plans over the same virtual packed stream: the source
// intervals cut it one way, the destination intervals another; a tile walks both lists and moves
// the intersections. `src` may be peer memory (NVLink loads): nothing here waits on another GPU.
template <int ES>
__global__ void __launch_bounds__(kCopyWarps * 32, kCopyOcc)
interval_move_kernel(const char* __restrict__ src, char* __restrict__ dst,
                     const int64_t* __restrict__ s_iv, const int64_t* __restrict__ s_off,
                     const int32_t* __restrict__ s_first, const int64_t* __restrict__ s_hdr, int64_t s_n,
                     const int64_t* __restrict__ d_iv, const int64_t* __restrict__ d_off,
                     const int32_t* __restrict__ d_first, const int64_t* __restrict__ d_hdr, int64_t d_n,
                     int64_t num_tiles, int tile_bytes) {
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * kCopyWarps + (threadIdx.x >> 5);
  if (tile >= num_tiles) return;
  const bool src_ok = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
  int64_t total_elems = s_hdr[0] < d_hdr[0] ? s_hdr[0] : d_hdr[0];
  if (total_elems > s_hdr[3]) total_elems = s_hdr[3];
  const int64_t total_bytes = total_elems * ES;
  int64_t p0 = tile * (int64_t)tile_bytes;
  int64_t p1 = p0 + tile_bytes;
  if (p1 > total_bytes) p1 = total_bytes;
  int64_t i = s_first[tile], j = d_first[tile];
  int64_t sob = 0, soe = -1, sst = 0, dob = 0, doe = -1, dst_start = 0;
  while (p0 < p1) {
    if (soe <= p0) {  // (re)load the source interval covering p0, skipping empty ones
      if (i >= s_n) break;
      sob = s_off[i] * ES;
      soe = s_off[i + 1] * ES;
      sst = s_iv[2 * i] * ES;
      i++;
      continue;
    }
    if (doe <= p0) {
      if (j >= d_n) break;
      dob = d_off[j] * ES;
      doe = d_off[j + 1] * ES;
      dst_start = d_iv[2 * j] * ES;
      j++;
      continue;
    }
    int64_t b = soe < doe ? soe : doe;
    if (b > p1) b = p1;
    copy_piece<ES>(src + sst + (p0 - sob), dst + dst_start + (p0 - dob), b - p0, lane, src_ok);
    p0 = b;
  }
}

int elem_shift_of(int elem_size) {
  switch (elem_size) {
    case 1: return 0;
    case 2: return 1;
    case 4: return 2;
    case 8: return 3;
    default: return -1;
  }
}

int build_plan(const PlanLayout& L, const int64_t* intervals, int64_t n, int64_t big_numel,
               int64_t packed_numel, int elem_shift, int tile_bytes, cudaStream_t st) {
  B200RL_CUDA(cudaMemsetAsync(L.hdr, 0, 4 * sizeof(int64_t), st));
  interval_local_scan<<<(unsigned)L.nb, kScanThreads, 0, st>>>(intervals, n, big_numel, L.offsets,
                                                               L.block_sums, L.hdr);
  interval_block_scan<<<1, 1024, 0, st>>>(L.block_sums, L.nb, L.hdr, n, packed_numel);
  interval_add_base<<<(unsigned)((n + 1 + 255) / 256), 256, 0, st>>>(L.offsets, L.block_sums, L.hdr, n);
  interval_partition<<<(unsigned)((L.num_tiles + 1 + 255) / 256), 256, 0, st>>>(
      L.offsets, n, L.tile_first, L.num_tiles, elem_shift, tile_bytes);
  return check_launch("interval plan");
}

template <int MODE>
int run_copy(const void* src, void* dst, int64_t big_numel, int64_t packed_numel,
             const int64_t* intervals, int64_t n, int elem_size, void* ws, size_t ws_bytes,
             int reuse_plan, void* stream) {
  B200RL_REQUIRE(n >= 0 && packed_numel >= 0 && big_numel >= 0, "negative size");
  const int shift = elem_shift_of(elem_size);
  B200RL_REQUIRE(shift >= 0, "elem_size %d not in {1,2,4,8}", elem_size);
  if (n == 0 || packed_numel == 0) return B200RL_OK;
  B200RL_REQUIRE(src && dst && intervals && ws, "null pointer");
  B200RL_REQUIRE(n < (int64_t)INT32_MAX, "too many intervals");
  B200RL_REQUIRE((reinterpret_cast<uintptr_t>(intervals) & 15) == 0, "intervals must be 16-B aligned");
  B200RL_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0, "workspace must be 16-B aligned");
  B200RL_REQUIRE(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & (elem_size - 1)) == 0,
                 "src/dst must be element aligned");
  const int tile_bytes = pick_tile_bytes(packed_numel * (int64_t)elem_size);
  PlanLayout L = plan_layout(ws, n, packed_numel, elem_size, tile_bytes);
  if (ws_bytes < L.bytes) {
    set_last_error("interval workspace too small: %zu < %zu", ws_bytes, L.bytes);
    return B200RL_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (!reuse_plan) {
    int rc = build_plan(L, intervals, n, big_numel, packed_numel, shift, tile_bytes, st);
    if (rc) return rc;
  }
  const unsigned grid = (unsigned)((L.num_tiles + kCopyWarps - 1) / kCopyWarps);
  const char* s = (const char*)src;
  char* d = (char*)dst;
#define LAUNCH(ES)                                                                             \
  interval_copy_kernel<MODE, ES><<<grid, kCopyWarps * 32, 0, st>>>(s, d, intervals, L.offsets, \
                                                                  L.tile_first, L.hdr, n, L.num_tiles, tile_bytes)
  switch (elem_size) {
    case 1: LAUNCH(1); break;
    case 2: LAUNCH(2); break;
    case 4: LAUNCH(4); break;
    default: LAUNCH(8); break;
  }
#undef LAUNCH
  return check_launch("interval_copy_kernel");
}

int run_move(const void* src, int64_t src_numel, void* dst, int64_t dst_numel,
             const int64_t* s_iv, int64_t s_n, const int64_t* d_iv, int64_t d_n, int64_t total_numel,
             int elem_size, void* ws, size_t ws_bytes, int reuse_plan, void* stream) {
  B200RL_REQUIRE(s_n >= 0 && d_n >= 0 && total_numel >= 0 && src_numel >= 0 && dst_numel >= 0, "negative size");
  const int shift = elem_shift_of(elem_size);
  B200RL_REQUIRE(shift >= 0, "elem_size %d not in {1,2,4,8}", elem_size);
  if (total_numel == 0) return B200RL_OK;
  B200RL_REQUIRE(s_n > 0 && d_n > 0, "total_numel > 0 with an empty interval list");
  B200RL_REQUIRE(src && dst && s_iv && d_iv && ws, "null pointer");
  B200RL_REQUIRE(s_n < (int64_t)INT32_MAX && d_n < (int64_t)INT32_MAX, "too many intervals");
  B200RL_REQUIRE(((reinterpret_cast<uintptr_t>(s_iv) | reinterpret_cast<uintptr_t>(d_iv)) & 15) == 0,
                 "intervals must be 16-B aligned");
  B200RL_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "workspace must be 256-B aligned");
  B200RL_REQUIRE(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & (elem_size - 1)) == 0,
                 "src/dst must be element aligned");
  const int tile_bytes = pick_tile_bytes(total_numel * (int64_t)elem_size);
  PlanLayout Ls = plan_layout(ws, s_n, total_numel, elem_size, tile_bytes);
  PlanLayout Ld = plan_layout((char*)ws + Ls.bytes, d_n, total_numel, elem_size, tile_bytes);
  if (ws_bytes < Ls.bytes + Ld.bytes) {
    set_last_error("interval move workspace too small: %zu < %zu", ws_bytes, Ls.bytes + Ld.bytes);
    return B200RL_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (!reuse_plan) {
    int rc = build_plan(Ls, s_iv, s_n, src_numel, total_numel, shift, tile_bytes, st);
    if (rc) return rc;
    rc = build_plan(Ld, d_iv, d_n, dst_numel, total_numel, shift, tile_bytes, st);
    if (rc) return rc;
  }
  const unsigned grid = (unsigned)((Ls.num_tiles + kCopyWarps - 1) / kCopyWarps);
  const char* s = (const char*)src;
  char* d = (char*)dst;
#define LAUNCH(ES)                                                                                \
  interval_move_kernel<ES><<<grid, kCopyWarps * 32, 0, st>>>(s, d, s_iv, Ls.offsets, Ls.tile_first, \
                                                            Ls.hdr, s_n, d_iv, Ld.offsets,        \
                                                            Ld.tile_first, Ld.hdr, d_n, Ls.num_tiles, tile_bytes)
  switch (elem_size) {
    case 1: LAUNCH(1); break;
    case 2: LAUNCH(2); break;
    case 4: LAUNCH(4); break;
    default: LAUNCH(8); break;
  }
#undef LAUNCH
  return check_launch("interval_move_kernel");
}

}  // namespace
}  // namespace b200rl

extern "C" {

size_t b200rl_interval_workspace_bytes(int64_t n_intervals, int64_t packed_numel, int elem_size) {
  if (n_intervals < 0 || packed_numel < 0 || elem_size <= 0) return 0;
  return b200rl::plan_layout(nullptr, n_intervals, packed_numel, elem_size).bytes;
}

int b200rl_slice_intervals(const void* big, int64_t big_numel, void* packed, int64_t packed_numel,
                           const int64_t* intervals, int64_t n_intervals, int elem_size,
                           void* workspace, size_t workspace_bytes, int reuse_plan, void* stream) {
  return b200rl::run_copy<0>(big, packed, big_numel, packed_numel, intervals, n_intervals,
                             elem_size, workspace, workspace_bytes, reuse_plan, stream);
}

int b200rl_set_intervals(const void* packed, int64_t packed_numel, void* big, int64_t big_numel,
                         const int64_t* intervals, int64_t n_intervals, int elem_size,
                         void* workspace, size_t workspace_bytes, int reuse_plan, void* stream) {
  return b200rl::run_copy<1>(packed, big, big_numel, packed_numel, intervals, n_intervals,
                             elem_size, workspace, workspace_bytes, reuse_plan, stream);
}

size_t b200rl_move_workspace_bytes(int64_t n_src, int64_t n_dst, int64_t total_numel, int elem_size) {
  if (n_src < 0 || n_dst < 0 || total_numel < 0 || elem_size <= 0) return 0;
  return b200rl::plan_layout(nullptr, n_src, total_numel, elem_size).bytes +
         b200rl::plan_layout(nullptr, n_dst, total_numel, elem_size).bytes;
}

int b200rl_move_intervals(const void* src, int64_t src_numel, void* dst, int64_t dst_numel,
                          const int64_t* src_intervals, int64_t n_src, const int64_t* dst_intervals,
                          int64_t n_dst, int64_t total_numel, int elem_size, void* workspace,
                          size_t workspace_bytes, int reuse_plan, void* stream) {
  return b200rl::run_move(src, src_numel, dst, dst_numel, src_intervals, n_src, dst_intervals, n_dst,
                          total_numel, elem_size, workspace, workspace_bytes, reuse_plan, stream);
}

int b200rl_enable_peer_access(int device, int peer) {
  using namespace b200rl;
  if (device == peer) return B200RL_OK;
  int can = 0;
  B200RL_CUDA(cudaDeviceCanAccessPeer(&can, device, peer));
  if (!can) {
    set_last_error("device %d cannot access peer %d", device, peer);
    return B200RL_EUNSUPPORTED;
  }
  int prev = 0;
  B200RL_CUDA(cudaGetDevice(&prev));
  B200RL_CUDA(cudaSetDevice(device));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    (void)cudaGetLastError();
    e = cudaSuccess;
  }
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    set_last_error("cudaDeviceEnablePeerAccess(%d -> %d): %s", device, peer, cudaGetErrorString(e));
    return (int)e;
  }
  return B200RL_OK;
}

// ---- CUDA IPC: export a buffer of this process / map a peer process's buffer ----------------
int b200rl_ipc_export(const void* ptr, unsigned char* handle64, int64_t* offset_bytes) {
  using namespace b200rl;
  B200RL_REQUIRE(ptr && handle64 && offset_bytes, "null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  // base of the allocation that contains ptr (the handle always names the whole allocation)
  typedef int (*GetRangeFn)(unsigned long long*, size_t*, unsigned long long);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  B200RL_CUDA(cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q));
  if (!fn || q != cudaDriverEntryPointSuccess) {
    set_last_error("cuMemGetAddressRange is not available");
    return B200RL_EUNSUPPORTED;
  }
  unsigned long long base = 0;
  size_t size = 0;
  const int drc = ((GetRangeFn)fn)(&base, &size, (unsigned long long)(uintptr_t)ptr);
  if (drc != 0) {
    set_last_error("cuMemGetAddressRange failed with driver error %d", drc);
    return B200RL_EINVAL;
  }
  cudaIpcMemHandle_t h;
  B200RL_CUDA(cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base));
  memcpy(handle64, &h, 64);
  *offset_bytes = (int64_t)((uintptr_t)ptr - (uintptr_t)base);
  return B200RL_OK;
}

int b200rl_ipc_open(const unsigned char* handle64, int device, void** base_out) {
  using namespace b200rl;
  B200RL_REQUIRE(handle64 && base_out, "null pointer");
  int prev = 0;
  B200RL_CUDA(cudaGetDevice(&prev));
  B200RL_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_last_error("cudaIpcOpenMemHandle on device %d: %s", device, cudaGetErrorString(e));
    return (int)e;
  }
  *base_out = p;
  return B200RL_OK;
}

int b200rl_ipc_close(void* base, int device) {
  using namespace b200rl;
  if (!base) return B200RL_OK;
  int prev = 0;
  B200RL_CUDA(cudaGetDevice(&prev));
  B200RL_CUDA(cudaSetDevice(device));
  cudaError_t e = cudaIpcCloseMemHandle(base);
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_last_error("cudaIpcCloseMemHandle: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return B200RL_OK;
}

}  // extern "C"
