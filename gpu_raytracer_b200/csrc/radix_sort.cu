// radix_sort.cu — A3: onesweep least-significant-digit radix sort of
// (Morton key, primitive index) pairs (Adinets & Merrill
// 2022, decoupled look-back of Merrill & Garland 2016).
//
// One kernel launch per 8-bit digit.  The digit histograms of all passes were
// accumulated while the keys were produced (morton.cu), so each pass reads every
// pair once and writes it once: 16 B/element/pass for 32-bit keys, 24 B for
// 64-bit keys.  Per tile:
//   1. take a tile id from an atomic counter (ids follow start order, so every
//      predecessor a tile waits for is already running: forward progress);
//   2. warp-striped coalesced key loads;
//   3. per-warp ranking: lanes with equal digits find each other (match.any or
//      an 8-ballot loop), the lowest such lane bumps the warp's private
//      shared-memory counter — no shared atomics, stable by construction;
//   4. one thread per bin folds the warp counters, publishes the tile's count
//      (flag LOCAL), looks back over predecessors until it meets an INCLUSIVE
//      word, then publishes its own inclusive prefix.  Count and flag share one
//      32-bit word, so a single relaxed store/load is atomic;
//   5. pairs are first placed in tile-sorted order in shared memory, then
//      written out so consecutive threads write consecutive addresses of one bin.
//
// Bound: HBM (integer/byte work, no tensor cores).  Grid: one CTA per tile,
// 2-3 CTAs resident per SM (shared memory sized for the 227 KB of sm_100a).
#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

constexpr int kRadix = 256;
constexpr uint32_t kFlagLocal = 1u << 30;
constexpr uint32_t kFlagIncl = 2u << 30;
constexpr uint32_t kValueMask = (1u << 30) - 1;

__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <bool USE_MATCH>
__device__ __forceinline__ uint32_t peers_of(uint32_t digit) {
  if (USE_MATCH) {
    return __match_any_sync(0xFFFFFFFFu, digit);
  } else {
    uint32_t mask = 0xFFFFFFFFu;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      bool bit = (digit >> b) & 1u;
      uint32_t vote = __ballot_sync(0xFFFFFFFFu, bit);
      mask &= bit ? vote : ~vote;
    }
    return mask;
  }
}

// exclusive prefix sums of the per-pass digit histograms: base[p][b]
__global__ void __launch_bounds__(kRadix) scan_hist_kernel(const uint32_t* __restrict__ hist,
                                                           uint32_t* __restrict__ base) {
  __shared__ uint32_t s_warp[8];
  const int b = threadIdx.x, lane = b & 31, warp = b >> 5;
  uint32_t v = hist[blockIdx.x * kRadix + b];
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t off = 0;
  for (int w = 0; w < warp; ++w) off += s_warp[w];
  base[blockIdx.x * kRadix + b] = off + incl - v;
}

template <typename KeyT, int THREADS, int IPT>
struct SortSmem {
  static constexpr int kWarps = THREADS / 32;
  static constexpr int kTile = THREADS * IPT;
  uint32_t warp_hist[kWarps][kRadix];
  uint32_t local_start[kRadix];
  uint32_t bin_base[kRadix];
  uint32_t scan_warp[8];
  uint32_t tile;
  KeyT keys[kTile];
  uint32_t vals[kTile];
};

template <typename KeyT, int THREADS, int IPT, int MIN_BLOCKS, bool IOTA, bool USE_MATCH>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
onesweep_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                const uint32_t* __restrict__ digit_base, uint32_t* __restrict__ lookback,
                uint32_t* __restrict__ tile_counter, uint32_t n, int shift) {
  using Smem = SortSmem<KeyT, THREADS, IPT>;
  constexpr int kTile = Smem::kTile;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) sm.tile = atomicAdd(tile_counter, 1u);
  // each warp clears its private counters
#pragma unroll
  for (int b = lane; b < kRadix; b += 32) sm.warp_hist[warp][b] = 0;
  __syncthreads();
  const uint32_t tile = sm.tile;
  const uint32_t tile_base = tile * (uint32_t)kTile;
  const uint32_t tile_n = min((uint32_t)kTile, n - tile_base);
  const bool full = (tile_n == (uint32_t)kTile);
  const uint32_t warp_base = tile_base + (uint32_t)warp * 32u * IPT + lane;

  // ---- 2. keys: warp-striped, coalesced -----------------------------------
  KeyT key[IPT];
  if (full) {
#pragma unroll
    for (int k = 0; k < IPT; ++k) key[k] = keys_in[warp_base + k * 32];
  } else {
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      uint32_t idx = warp_base + k * 32;
      key[k] = (idx < n) ? keys_in[idx] : ~KeyT(0);
    }
  }

  // ---- 3. ranking inside the warp -------------------------------------------
  uint32_t rank[IPT];
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint32_t* wh = sm.warp_hist[warp];
#pragma unroll
  for (int k = 0; k < IPT; ++k) {
    const uint32_t d = (uint32_t)(key[k] >> shift) & (kRadix - 1);
    const uint32_t peers = peers_of<USE_MATCH>(d);
    const int leader = __ffs(peers) - 1;
    uint32_t old = 0;
    if (lane == leader) {
      old = wh[d];
      wh[d] = old + __popc(peers);
    }
    old = __shfl_sync(0xFFFFFFFFu, old, leader);
    rank[k] = old + __popc(peers & lt_mask);
    __syncwarp();
  }

  // payload: issued now so the loads fly during the look-back
  uint32_t val[IPT];
  if (IOTA) {
#pragma unroll
    for (int k = 0; k < IPT; ++k) val[k] = warp_base + k * 32;
  } else if (full) {
#pragma unroll
    for (int k = 0; k < IPT; ++k) val[k] = vals_in[warp_base + k * 32];
  } else {
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      uint32_t idx = warp_base + k * 32;
      val[k] = (idx < n) ? vals_in[idx] : 0u;
    }
  }
  __syncthreads();

  // ---- 4. per-bin fold, publish, look back -----------------------------------
  uint32_t total = 0, excl = 0;
  if (tid < kRadix) {
#pragma unroll
    for (int w = 0; w < Smem::kWarps; ++w) {
      uint32_t c = sm.warp_hist[w][tid];
      sm.warp_hist[w][tid] = total;
      total += c;
    }
    uint32_t* mine = lookback + (size_t)tile * kRadix + tid;
    if (tile == 0) {
      st_relaxed(mine, total | kFlagIncl);
    } else {
      st_relaxed(mine, total | kFlagLocal);
    }
    // exclusive scan of the tile's bin totals -> position of the bin in the tile
    uint32_t incl = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) sm.scan_warp[warp] = incl;
    asm volatile("bar.sync 1, 256;" ::: "memory");  // the 256 bin threads only
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) off += (w < warp) ? sm.scan_warp[w] : 0u;
    const uint32_t lstart = off + incl - total;
    sm.local_start[tid] = lstart;

    if (tile != 0) {
      int t = (int)tile - 1;
      while (true) {
        uint32_t v = ld_relaxed(lookback + (size_t)t * kRadix + tid);
        uint32_t flag = v & ~kValueMask;
        if (flag == 0) continue;  // predecessor has not published yet
        excl += v & kValueMask;
        if (flag == kFlagIncl) break;
        --t;
      }
      st_relaxed(mine, (excl + total) | kFlagIncl);
    }
    sm.bin_base[tid] = digit_base[tid] + excl - lstart;
  }
  __syncthreads();

  // ---- 5. tile-local placement, then coalesced write-out ---------------------
#pragma unroll
  for (int k = 0; k < IPT; ++k) {
    const uint32_t d = (uint32_t)(key[k] >> shift) & (kRadix - 1);
    const uint32_t pos = sm.local_start[d] + wh[d] + rank[k];
    sm.keys[pos] = key[k];
    sm.vals[pos] = val[k];
  }
  __syncthreads();
  if (full) {
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      const uint32_t p = tid + k * THREADS;
      const KeyT kk = sm.keys[p];
      const uint32_t d = (uint32_t)(kk >> shift) & (kRadix - 1);
      const uint32_t dst = sm.bin_base[d] + p;
      keys_out[dst] = kk;
      vals_out[dst] = sm.vals[p];
    }
  } else {
    for (uint32_t p = tid; p < tile_n; p += THREADS) {
      const KeyT kk = sm.keys[p];
      const uint32_t d = (uint32_t)(kk >> shift) & (kRadix - 1);
      const uint32_t dst = sm.bin_base[d] + p;
      keys_out[dst] = kk;
      vals_out[dst] = sm.vals[p];
    }
  }
}

// ---- variants ----------------------------------------------------------------
struct Variant {
  int threads, ipt;
};
// index = params().sort_variant; [key width][variant]
constexpr int kNumVariants = 4;
constexpr Variant kVariants32[kNumVariants] = {{512, 15}, {384, 18}, {256, 23}, {512, 15}};
constexpr Variant kVariants64[kNumVariants] = {{512, 10}, {384, 13}, {256, 16}, {512, 10}};
constexpr int kMinTile32 = 256 * 23, kMinTile64 = 256 * 16;

template <typename KeyT, int THREADS, int IPT, int MIN_BLOCKS, bool USE_MATCH>
int launch_pass(const KeyT* kin, KeyT* kout, const uint32_t* vin, uint32_t* vout,
                const uint32_t* base, uint32_t* lookback, uint32_t* counter, uint32_t n, int shift,
                bool iota, cudaStream_t s) {
  using Smem = SortSmem<KeyT, THREADS, IPT>;
  const uint32_t tiles = (n + Smem::kTile - 1) / Smem::kTile;
  const size_t smem = sizeof(Smem);
  if (iota) {
    auto kern = onesweep_kernel<KeyT, THREADS, IPT, MIN_BLOCKS, true, USE_MATCH>;
    RT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<tiles, THREADS, smem, s>>>(kin, kout, vin, vout, base, lookback, counter, n, shift);
  } else {
    auto kern = onesweep_kernel<KeyT, THREADS, IPT, MIN_BLOCKS, false, USE_MATCH>;
    RT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<tiles, THREADS, smem, s>>>(kin, kout, vin, vout, base, lookback, counter, n, shift);
  }
  RT_LAUNCHED();
  return 0;
}

template <typename KeyT>
int dispatch_pass(int variant, const KeyT* kin, KeyT* kout, const uint32_t* vin, uint32_t* vout,
                  const uint32_t* base, uint32_t* lookback, uint32_t* counter, uint32_t n,
                  int shift, bool iota, cudaStream_t s) {
  constexpr bool k64 = sizeof(KeyT) == 8;
  switch (variant) {
    default:
    case 0:
      return launch_pass<KeyT, 512, k64 ? 10 : 15, 2, false>(kin, kout, vin, vout, base, lookback,
                                                            counter, n, shift, iota, s);
    case 1:
      return launch_pass<KeyT, 384, k64 ? 13 : 18, 2, false>(kin, kout, vin, vout, base, lookback,
                                                            counter, n, shift, iota, s);
    case 2:
      return launch_pass<KeyT, 256, k64 ? 16 : 23, 3, false>(kin, kout, vin, vout, base, lookback,
                                                            counter, n, shift, iota, s);
    case 3:
      return launch_pass<KeyT, 512, k64 ? 10 : 15, 2, true>(kin, kout, vin, vout, base, lookback,
                                                             counter, n, shift, iota, s);
  }
}

}  // namespace

int sort_num_passes(uint32_t key_bits) { return key_bits == 30 ? 4 : 8; }

// [passes][256] tile counters (one word used per pass, padded) then
// [passes][tiles_max][256] look-back words
uint64_t sort_lookback_words(uint64_t n, uint32_t key_bits) {
  const uint64_t min_tile = key_bits == 30 ? kMinTile32 : kMinTile64;
  const uint64_t tiles = (n + min_tile - 1) / min_tile;
  const uint64_t passes = sort_num_passes(key_bits);
  return passes * kRadix + passes * tiles * kRadix;
}

int launch_sort(void* keys, uint64_t n, uint32_t key_bits, const SortWorkspace& ws,
                bool iota_payload, void** keys_out, uint32_t** vals_out, cudaStream_t s) {
  *keys_out = keys;
  *vals_out = ws.vals;
  if (n > ((1ull << 30) - 1)) return fail(-5, "sort: n = %llu beyond 2^30-1", (unsigned long long)n);
  const int passes = sort_num_passes(key_bits);
  if (n == 0) return 0;
  int variant = params().sort_variant;
  if (variant < 0 || variant >= kNumVariants) variant = 0;
  const uint64_t min_tile = key_bits == 30 ? kMinTile32 : kMinTile64;
  const uint64_t tiles_max = (n + min_tile - 1) / min_tile;
  RT_CUDA(cudaMemsetAsync(ws.lookback, 0, ws.lookback_words * sizeof(uint32_t), s));
  uint32_t* base = ws.hist + (size_t)passes * kRadix;
  scan_hist_kernel<<<passes, kRadix, 0, s>>>(ws.hist, base);
  RT_LAUNCHED();
  uint32_t* counters = ws.lookback;
  uint32_t* lb = ws.lookback + (size_t)passes * kRadix;
  for (int p = 0; p < passes; ++p) {
    const bool fwd = (p & 1) == 0;
    const void* kin = fwd ? keys : ws.keys_alt;
    void* kout = fwd ? ws.keys_alt : keys;
    const uint32_t* vin = fwd ? ws.vals : ws.vals_alt;
    uint32_t* vout = fwd ? ws.vals_alt : ws.vals;
    const bool iota = iota_payload && p == 0;
    uint32_t* lbp = lb + (size_t)p * tiles_max * kRadix;
    if (key_bits == 30) {
      RT_TRY(dispatch_pass<uint32_t>(variant, (const uint32_t*)kin, (uint32_t*)kout, vin, vout,
                                     base + p * kRadix, lbp, counters + p * kRadix, (uint32_t)n,
                                     8 * p, iota, s));
    } else {
      RT_TRY(dispatch_pass<uint64_t>(variant, (const uint64_t*)kin, (uint64_t*)kout, vin, vout,
                                     base + p * kRadix, lbp, counters + p * kRadix, (uint32_t)n,
                                     8 * p, iota, s));
    }
  }
  // passes is even (4 or 8): the result is back in the caller's buffers
  return 0;
}

}  // namespace rt
