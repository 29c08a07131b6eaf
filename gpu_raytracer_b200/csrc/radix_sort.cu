// radix_sort.cu — A3: least-significant-digit radix sort of (Morton key,
// primitive index) pairs.
//
// B200 shape of the algorithm.  A decoupled look-back onesweep (Adinets &
// Merrill 2022) was built and measured first (round 1; the kernel is in the history, git show
// 8291441:gpu_raytracer_b200/csrc/radix_sort.cu; its measurements were not kept): at ~0.05 us
// between tile starts and ~0.7 us per L2 round trip every tile had to walk
// 10-20 predecessors for each of its 256-1024 bins, and the look-back words
// cost more L2 traffic than the pairs themselves.  This version has no
// inter-CTA dependency at all:
//
//   chunk_hist_kernel    the input is cut into one contiguous chunk per resident
//                        CTA; each CTA histograms the digit of its chunk in
//                        shared memory and stores H[chunk][bin] (+ global bin
//                        totals).  Keys only: 4 or 8 B/element, mostly L2 hits
//                        because the previous pass has just written them.
//   chunk_offsets_kernel O[chunk][bin] = (start of the bin) + (elements of that
//                        bin in earlier chunks): one column scan per bin.
//   scatter_kernel       persistent, one CTA per chunk, tiles of the chunk in
//                        order.  Per tile: warp-striped key loads, per-warp
//                        stable ranking against private 16-bit counters,
//                        fold + scan of the tile's bin counts, placement in
//                        shared memory in tile-sorted order, coalesced write-out.
//                        The running output offset of every bin lives in the
//                        registers of the bin's owner thread; the next tile's
//                        keys are fetched while the current tile is written.
//
// Digits are wide (10 bits for 30-bit keys: 3 passes; 9 bits for 63-bit keys:
// 7 passes): every pass saved is 16 or 24 B/element that never move.
//
// Ranking: every lane reads its digit's counter, writes back count+1 tagged
// with its lane id and re-reads.  Lanes that find their own tag are alone with
// their digit.  Otherwise the tag read back names one lane per digit group (5
// bits), so five ballots on it give every lane its peer mask, instead of one
// ballot per digit bit.  No shared atomics, stable by construction.
//
// Bound: HBM by the byte count (integer/byte work, no tensor cores); the
// measured limiter is instruction issue (profiles/).
#include <type_traits>

#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

constexpr uint32_t kFull = 0xFFFFFFFFu;

// ---- per-chunk digit histogram ------------------------------------------------
constexpr int kHistThreads = 1024;

template <int BITS>
__device__ __forceinline__ void hist_add(uint32_t* sh, uint32_t d, bool valid) {
  // coherent input: a warp often holds one digit only; count it with one atomic
  const uint32_t d0 = __shfl_sync(kFull, d, 0);
  if (__all_sync(kFull, valid && d == d0)) {
    if ((threadIdx.x & 31) == 0) atomicAdd(&sh[d], 32u);
  } else if (valid) {
    atomicAdd(&sh[d], 1u);
  }
}

template <typename KeyT, int BITS>
__global__ void __launch_bounds__(kHistThreads)
chunk_hist_kernel(const KeyT* __restrict__ keys, ChunkGeom g, int shift,
                  uint32_t* __restrict__ chunk_hist, uint32_t* __restrict__ totals) {
  constexpr int kBins = 1 << BITS;
  __shared__ uint32_t sh[kBins];
  pdl_enter();
  for (int b = threadIdx.x; b < kBins; b += kHistThreads) sh[b] = 0u;
  __syncthreads();
  const uint32_t c = blockIdx.x, begin = g.begin(c), end = g.end(c);
  constexpr uint32_t kUnroll = 4;
  uint32_t base = begin;
  for (; base + kUnroll * kHistThreads <= end; base += kUnroll * kHistThreads) {
    KeyT k[kUnroll];
#pragma unroll
    for (uint32_t u = 0; u < kUnroll; ++u) k[u] = keys[base + u * kHistThreads + threadIdx.x];
#pragma unroll
    for (uint32_t u = 0; u < kUnroll; ++u)
      hist_add<BITS>(sh, (uint32_t)(k[u] >> shift) & (kBins - 1), true);
  }
  for (; base < end; base += kHistThreads) {
    const uint32_t i = base + threadIdx.x;
    const bool valid = i < end;
    const KeyT k = valid ? keys[i] : KeyT(0);
    hist_add<BITS>(sh, (uint32_t)(k >> shift) & (kBins - 1), valid);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < kBins; b += kHistThreads) {
    const uint32_t v = sh[b];
    chunk_hist[(size_t)c * kBins + b] = v;
    if (v) atomicAdd(&totals[b], v);
  }
}

// ---- chunk offsets ---------------------------------------------------------------
// in place: chunk_hist[c][b] <- start of bin b + elements of bin b in chunks < c.
// One warp per bin: lanes stride over the chunks, warp-shuffle scan with carry.
constexpr int kOffWarps = 8;
constexpr int kOffMaxRounds = 24;  // chunks <= 32 * 24

template <int BITS>
__global__ void __launch_bounds__(kOffWarps * 32)
chunk_offsets_kernel(uint32_t* __restrict__ chunk_hist, const uint32_t* __restrict__ totals,
                     uint32_t chunks) {
  constexpr int kBins = 1 << BITS;
  pdl_enter();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int bin = blockIdx.x * kOffWarps + warp;
  const int rounds = (int)((chunks + 31) / 32);
  uint32_t v[kOffMaxRounds];
#pragma unroll
  for (int j = 0; j < kOffMaxRounds; ++j) {
    const uint32_t c = (uint32_t)lane + 32u * j;
    v[j] = (j < rounds && c < chunks) ? chunk_hist[(size_t)c * kBins + bin] : 0u;
  }
  // start of the bin: totals of all lower bins, summed by the whole CTA at once
  __shared__ uint32_t s_part[kOffWarps];
  __shared__ uint32_t s_tot[kOffWarps];
  {
    constexpr int kPer = (kBins + kOffWarps * 32 - 1) / (kOffWarps * 32);
    const int first_bin = blockIdx.x * kOffWarps;
    uint32_t part = 0;
#pragma unroll
    for (int i = 0; i < kPer; ++i) {
      const int b = (int)threadIdx.x + i * kOffWarps * 32;
      if (b < first_bin) part += totals[b];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(kFull, part, o);
    if (lane == 0) {
      s_part[warp] = part;
      s_tot[warp] = totals[bin];
    }
  }
  __syncthreads();
  uint32_t run = 0;
#pragma unroll
  for (int w = 0; w < kOffWarps; ++w) run += s_part[w] + ((w < warp) ? s_tot[w] : 0u);
#pragma unroll
  for (int j = 0; j < kOffMaxRounds; ++j) {
    if (j < rounds) {  // warp-uniform
      uint32_t incl = v[j];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += t;
      }
      const uint32_t c = (uint32_t)lane + 32u * j;
      if (c < chunks) chunk_hist[(size_t)c * kBins + bin] = run + incl - v[j];
      run += __shfl_sync(kFull, incl, 31);
    }
  }
}

// ---- scatter -----------------------------------------------------------------------
template <typename KeyT, int BITS, int THREADS, int IPT>
struct SortCfg {
  static constexpr int kBins = 1 << BITS;
  static constexpr int kWarps = THREADS / 32;
  static constexpr int kTile = THREADS * IPT;
  // bin owners: each owns kBPT consecutive bins (an even number, handled as
  // kPPT packed pairs of 16-bit counters)
  static constexpr int kOwners = (THREADS < kBins / 2) ? THREADS : kBins / 2;
  static constexpr int kBPT = kBins / kOwners;
  static constexpr int kPPT = kBPT / 2;
  static_assert(kBPT % 2 == 0 && kBPT >= 2, "bin owners handle packed pairs");
  static_assert(IPT * 32 < 2048, "11-bit per-warp counters next to a 5-bit lane tag");
  static_assert(kTile < 65536, "16-bit tile-local positions");
};

// 32-bit keys travel between the passes as packed 64-bit records
// (key << 32 | payload): one load, one shared-memory slot and one scattered
// store per element instead of two.  64-bit keys keep separate arrays.
template <typename KeyT, int BITS, int THREADS, int IPT, bool TMA = false>
struct SortSmem {
  using C = SortCfg<KeyT, BITS, THREADS, IPT>;
  static constexpr bool kPacked = sizeof(KeyT) == 4;
  alignas(16) uint64_t in_tile[TMA ? C::kTile : 2];  // TMA: the next input tile, landed by cp.async.bulk
  alignas(8) uint64_t mbar;                            // TMA: its mbarrier
  uint64_t keys[C::kTile];                   // packed records, or the 64-bit keys
  uint32_t vals[kPacked ? 1 : C::kTile];     // payload of 64-bit keys
  uint32_t bin_base[C::kBins];
  uint16_t warp_hist[C::kWarps][C::kBins];  // (count << 5 | lane tag), later exclusive offsets
  uint16_t local_start[C::kBins];
  uint32_t scan_warp[32];
  uint32_t peer_tbl[C::kWarps][32];
};

enum : int { kInSplit = 0, kInIota = 1, kInRec = 2 };
enum : int { kOutSplit = 0, kOutRec = 1 };

// Rank of this lane's key among the keys of the warp seen so far with the same
// digit; bumps the warp's private counter.  wh[d] = (count << 5) | lane tag.
// The tag read back names one lane per group of equal digits; the lanes of a
// group OR their bit into that lane's slot of a 32-word table, which is then the
// group's peer mask: three shared-memory operations instead of one ballot per
// digit bit.  A warp holding one digit only (coherent input) skips the atomics.
__device__ __forceinline__ uint32_t warp_rank(volatile uint16_t* wh, volatile uint32_t* peer_tbl,
                                              uint32_t d, uint32_t lane, uint32_t lt_mask) {
  const uint32_t c = (uint32_t)wh[d] >> 5;
  wh[d] = (uint16_t)(((c + 1u) << 5) | lane);
  __syncwarp();
  const uint32_t winner = (uint32_t)wh[d] & 31u;  // one lane id per group of equal digits
  // 32 random digits out of 1024 are all different in 3 rounds of 5: every lane is alone in
  // its group, and the counter it wrote (count + 1, the tag bits are never counted) stands
  if (__ballot_sync(kFull, winner != lane) == 0u) {
    __syncwarp();
    return c;
  }
  uint32_t peers;
  if (__all_sync(kFull, winner == __shfl_sync(kFull, winner, 0))) {
    peers = kFull;
  } else {
    peer_tbl[lane] = 0u;
    __syncwarp();
    atomicOr(const_cast<uint32_t*>(peer_tbl) + winner, 1u << lane);
    __syncwarp();
    peers = peer_tbl[winner];
  }
  if ((peers & lt_mask) == 0u) wh[d] = (uint16_t)((c + __popc(peers)) << 5);
  __syncwarp();
  return c + __popc(peers & lt_mask);
}

// ---- TMA-staged input tiles (variant 3; BASELINE.json:5 "shared-memory/TMA-staged tiles") ----------
// One elected thread asks the copy engine for the whole tile (a 1-D cp.async.bulk, SASS
// UBLKCP) and the CTA waits on an mbarrier; the LSU then carries no global loads for the
// records.  Measured against the register prefetch: profiles/r02j_sort_tma_ab.md.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}

// IN: kInSplit (keys_in + vals_in), kInIota (keys_in, payload = position),
//     kInRec (recs_in, 32-bit keys only).   OUT: kOutSplit or kOutRec.
template <typename KeyT, int BITS, int THREADS, int IPT, int MIN_BLOCKS, int IN, int OUT, bool TMA = false>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
scatter_kernel(const KeyT* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
               const uint2* __restrict__ recs_in, KeyT* __restrict__ keys_out,
               uint32_t* __restrict__ vals_out, uint2* __restrict__ recs_out,
               const uint32_t* __restrict__ chunk_offsets, ChunkGeom g, int shift) {
  using C = SortCfg<KeyT, BITS, THREADS, IPT>;
  using Smem = SortSmem<KeyT, BITS, THREADS, IPT, TMA>;
  constexpr bool kPacked = Smem::kPacked;
  static_assert(kPacked || (IN != kInRec && OUT != kOutRec), "records carry 32-bit keys");
  static_assert(!TMA || IN == kInRec, "TMA stages packed records");
  constexpr int kTile = C::kTile, kBins = C::kBins, kBPT = C::kBPT, kPPT = C::kPPT;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  pdl_enter();

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t chunk = blockIdx.x;
  const uint32_t chunk_begin = g.begin(chunk), chunk_end = g.end(chunk);
  const uint32_t m = chunk_end - chunk_begin;
  // equal tiles: T = ceil(m / kTile) tiles of `steps` key-steps each
  const uint32_t n_tiles = (m + kTile - 1) / kTile;
  const uint32_t per_tile = (m + n_tiles - 1) / n_tiles;
  const int steps = (int)((per_tile + THREADS - 1) / THREADS);  // <= IPT
  const uint32_t tile_size = (uint32_t)steps * THREADS;
  const uint32_t seg = (uint32_t)steps * 32u;  // keys per warp per tile

  // running output position of the owned bins (registers, carried across tiles)
  const bool owner = tid < C::kOwners;
  const int bin0 = tid * kBPT;
  uint32_t run_base[kBPT];
#pragma unroll
  for (int i = 0; i < kBPT; ++i)
    run_base[i] = owner ? chunk_offsets[(size_t)chunk * kBins + bin0 + i] : 0u;

  KeyT key[IPT];
  uint32_t val[IPT];
  // keys (and, for records, payloads) of one tile; out-of-range slots hold all-ones
  auto load_keys = [&](uint32_t tile_begin) {
    const uint32_t wb = tile_begin + (uint32_t)warp * seg + lane;
    const bool whole = (steps == IPT) && (tile_begin + tile_size <= chunk_end);
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      const uint32_t idx = wb + k * 32;
      const bool ok = whole || (k < steps && idx < chunk_end);
      if (IN == kInRec) {
        const uint2 r = ok ? recs_in[idx] : make_uint2(0u, 0xFFFFFFFFu);
        key[k] = (KeyT)r.y;
        val[k] = r.x;
      } else {
        key[k] = ok ? keys_in[idx] : ~KeyT(0);
      }
    }
  };
  auto load_vals = [&](uint32_t tile_begin) {
    const uint32_t wb = tile_begin + (uint32_t)warp * seg + lane;
    const bool whole = (steps == IPT) && (tile_begin + tile_size <= chunk_end);
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      const uint32_t idx = wb + k * 32;
      if (IN == kInIota) val[k] = idx;
      else if (IN == kInSplit) val[k] = (whole || (k < steps && idx < chunk_end)) ? vals_in[idx] : 0u;
    }
  };
  // TMA: a tile whose byte count is a multiple of 16 comes through the copy engine, any
  // other (only the odd-sized last tile of the array) through plain loads
  auto tma_ok = [&](uint32_t tile_begin) {
    return TMA && ((min(tile_size, chunk_end - tile_begin) & 1u) == 0u);
  };
  auto tma_issue = [&](uint32_t tile_begin) {  // one thread; in_tile is free (everybody has read it)
    if (TMA && tid == 0)
      tma_load_1d(sm.in_tile, recs_in + tile_begin, min(tile_size, chunk_end - tile_begin) * 8u, &sm.mbar);
  };
  auto tma_consume = [&](uint32_t tile_begin, uint32_t parity) {
    mbar_wait(&sm.mbar, parity);
    const uint32_t off = (uint32_t)warp * seg + lane;
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      const uint32_t idx = tile_begin + off + k * 32;
      const bool ok = k < steps && idx < chunk_end;
      const uint64_t r = ok ? sm.in_tile[off + k * 32] : 0xFFFFFFFF00000000ull;
      key[k] = (KeyT)(uint32_t)(r >> 32);
      val[k] = (uint32_t)r;
    }
  };
  uint32_t tma_phase = 0;
  if (TMA) {
    if (tid == 0) mbar_init(&sm.mbar, 1);
    __syncthreads();
  }
  if (n_tiles) {
    if (tma_ok(chunk_begin)) tma_issue(chunk_begin); else load_keys(chunk_begin);
  }

  for (uint32_t t = 0; t < n_tiles; ++t) {
    const uint32_t tile_begin = chunk_begin + t * tile_size;
    const uint32_t tile_n = min(tile_size, chunk_end - tile_begin);
    if (tma_ok(tile_begin)) {
      tma_consume(tile_begin, tma_phase);
      tma_phase ^= 1u;
    }

    {  // each warp clears its private counters (32-bit words = two counters)
      uint32_t* w32 = reinterpret_cast<uint32_t*>(sm.warp_hist[warp]);
#pragma unroll
      for (int b = lane; b < kBins / 2; b += 32) w32[b] = 0u;
      __syncwarp();
    }

    // ---- ranking inside the warp ---------------------------------------------
    uint32_t rank[IPT];
    {
      volatile uint16_t* wh = sm.warp_hist[warp];
      volatile uint32_t* ptbl = sm.peer_tbl[warp];
#pragma unroll
      for (int k = 0; k < IPT; ++k) {
        if (k < steps) {  // warp-uniform
          const uint32_t d = (uint32_t)(key[k] >> shift) & (kBins - 1);
          rank[k] = warp_rank(wh, ptbl, d, (uint32_t)lane, lt_mask);
        } else {
          rank[k] = 0u;
        }
      }
    }
    // payload: issued now so the loads fly during the fold
    if (IN != kInRec) load_vals(tile_begin);
    __syncthreads();
    // TMA: everybody has taken its records out of in_tile: the next tile may land in it
    if (TMA && t + 1 < n_tiles && tma_ok(tile_begin + tile_size)) tma_issue(tile_begin + tile_size);

    // ---- fold the warp counters, scan the bin totals ---------------------------
    uint32_t total[kBPT];
    uint32_t thread_sum = 0;
    if (owner) {
      uint32_t run[kPPT];
#pragma unroll
      for (int q = 0; q < kPPT; ++q) run[q] = 0u;
#pragma unroll
      for (int w = 0; w < C::kWarps; ++w) {
        uint32_t* w32 = reinterpret_cast<uint32_t*>(sm.warp_hist[w]) + tid * kPPT;
#pragma unroll
        for (int q = 0; q < kPPT; ++q) {
          const uint32_t x = w32[q];
          w32[q] = run[q];
          run[q] += (x >> 5) & 0x07FF07FFu;
        }
      }
#pragma unroll
      for (int q = 0; q < kPPT; ++q) {
        total[2 * q] = run[q] & 0xFFFFu;
        total[2 * q + 1] = run[q] >> 16;
        thread_sum += total[2 * q] + total[2 * q + 1];
      }
    } else {
#pragma unroll
      for (int i = 0; i < kBPT; ++i) total[i] = 0u;
    }
    {  // block-wide exclusive scan of thread_sum -> tile-local start of every bin
      uint32_t incl = thread_sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) sm.scan_warp[warp] = incl;
      __syncthreads();
      uint32_t run = incl - thread_sum;
#pragma unroll
      for (int w = 0; w < C::kWarps; ++w) run += (w < warp) ? sm.scan_warp[w] : 0u;
      if (owner) {
#pragma unroll
        for (int i = 0; i < kBPT; ++i) {
          sm.local_start[bin0 + i] = (uint16_t)run;
          sm.bin_base[bin0 + i] = run_base[i] - run;
          run_base[i] += total[i];
          run += total[i];
        }
      }
    }
    __syncthreads();

    // ---- tile-local placement ------------------------------------------------------
    {
      const uint16_t* wh = sm.warp_hist[warp];
#pragma unroll
      for (int k = 0; k < IPT; ++k) {
        if (k < steps) {
          const uint32_t d = (uint32_t)(key[k] >> shift) & (kBins - 1);
          const uint32_t pos = (uint32_t)sm.local_start[d] + (uint32_t)wh[d] + rank[k];
          if (kPacked) {
            sm.keys[pos] = ((uint64_t)key[k] << 32) | val[k];
          } else {
            sm.keys[pos] = key[k];
            sm.vals[pos] = val[k];
          }
        }
      }
    }
    __syncthreads();

    // next tile's keys travel while this tile is written out
    if (t + 1 < n_tiles && !tma_ok(tile_begin + tile_size)) load_keys(tile_begin + tile_size);

    // ---- coalesced write-out -------------------------------------------------------
    auto emit = [&](uint32_t p) {
      const uint64_t kk = sm.keys[p];
      const uint32_t d = (uint32_t)(kk >> (shift + (kPacked ? 32 : 0))) & (kBins - 1);
      const uint32_t dst = sm.bin_base[d] + p;
      if (kPacked) {
        if (OUT == kOutRec) {
          recs_out[dst] = make_uint2((uint32_t)kk, (uint32_t)(kk >> 32));
        } else {
          keys_out[dst] = (KeyT)(kk >> 32);
          vals_out[dst] = (uint32_t)kk;
        }
      } else {
        keys_out[dst] = (KeyT)kk;
        vals_out[dst] = sm.vals[p];
      }
    };
    if (steps == IPT && tile_n == (uint32_t)kTile) {
#pragma unroll
      for (int k = 0; k < IPT; ++k) emit(tid + k * THREADS);
    } else {
      for (uint32_t p = tid; p < tile_n; p += THREADS) emit(p);
    }
    // no barrier here: the next tile's placement sits behind three more barriers
  }
}

// ---- launchers -----------------------------------------------------------------
struct Shape {
  int threads, ipt32, ipt64, ctas_per_sm;
};
// variant 0: 512 threads, 2 CTAs/SM; variant 1: 256 threads, 4 CTAs/SM; variants 2 and 3 (30-bit
// keys, 10-bit digits only; experiments): 512 threads x 9 keys, 2 CTAs/SM, without and with
// TMA-staged record tiles (two tile buffers of 36 KB fit next to the counters twice per SM)
constexpr int kNumVariants = 4;
constexpr Shape kShapes[kNumVariants] = {{512, 15, 10, 2}, {256, 16, 11, 4}, {512, 9, 10, 2}, {512, 9, 10, 2}};

int clamp_variant() {
  int v = params().sort_variant;
  return (v < 0 || v >= kNumVariants) ? 0 : v;
}

}  // namespace
ChunkGeom sort_geometry(uint64_t n, uint32_t key_bits) {
  int v = clamp_variant();
  if (v >= 2 && (key_bits != 30 || sort_plan(30).bits != 10)) v = 0;  // variants 2, 3: 30-bit keys, 10-bit digits
  const Shape& sh = kShapes[v];
  const uint64_t tile = (uint64_t)sh.threads * (key_bits == 30 ? sh.ipt32 : sh.ipt64);
  uint64_t max_ctas = (uint64_t)num_sms() * sh.ctas_per_sm;
  // chunk_offsets_kernel scans at most 32 * kOffMaxRounds chunks per bin
  if (max_ctas > 32ull * kOffMaxRounds) max_ctas = 32ull * kOffMaxRounds;
  ChunkGeom g;
  g.n = (uint32_t)n;
  const uint64_t tiles = (n + tile - 1) / tile;
  if (tiles <= max_ctas) {
    g.chunk_size = (uint32_t)tile;
    g.chunks = (uint32_t)(tiles ? tiles : 1);
  } else {
    uint64_t cs = (n + max_ctas - 1) / max_ctas;
    cs = (cs + 31) / 32 * 32;
    g.chunk_size = (uint32_t)cs;
    g.chunks = (uint32_t)((n + cs - 1) / cs);
  }
  return g;
}
namespace {
ChunkGeom make_geom(uint64_t n, uint32_t key_bits) { return sort_geometry(n, key_bits); }

struct PassIO {
  const void* keys_in = nullptr;   // KeyT keys (kInSplit / kInIota)
  const uint32_t* vals_in = nullptr;
  const uint2* recs_in = nullptr;  // packed records (kInRec)
  void* keys_out = nullptr;
  uint32_t* vals_out = nullptr;
  uint2* recs_out = nullptr;
  int in = kInSplit, out = kOutSplit;
};

template <typename KeyT, int BITS, int THREADS, int IPT, int MIN_BLOCKS, int IN, int OUT, bool TMA = false>
int launch_scatter_io(const PassIO& io, const uint32_t* offsets, const ChunkGeom& g, int shift,
                      cudaStream_t s) {
  using Smem = SortSmem<KeyT, BITS, THREADS, IPT, TMA>;
  const size_t smem = sizeof(Smem);
  auto kern = scatter_kernel<KeyT, BITS, THREADS, IPT, MIN_BLOCKS, IN, OUT, TMA>;
  RT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  RT_CUDA(launch_pdl(kern, g.chunks, THREADS, smem, s, (const KeyT*)io.keys_in, io.vals_in,
                     io.recs_in, (KeyT*)io.keys_out, io.vals_out, io.recs_out, offsets, g, shift));
  RT_LAUNCHED();
  return 0;
}

template <typename KeyT, int BITS, int THREADS, int IPT, int MIN_BLOCKS, bool TMA = false>
int launch_scatter(const PassIO& io, const uint32_t* offsets, const ChunkGeom& g, int shift,
                   cudaStream_t s) {
  if (sizeof(KeyT) == 8) {
    if (io.in == kInIota)
      return launch_scatter_io<KeyT, BITS, THREADS, IPT, MIN_BLOCKS, kInIota, kOutSplit>(io, offsets, g, shift, s);
    return launch_scatter_io<KeyT, BITS, THREADS, IPT, MIN_BLOCKS, kInSplit, kOutSplit>(io, offsets, g, shift, s);
  } else {
    // 32-bit keys: first pass builds records, last pass splits them again
    constexpr bool k32 = sizeof(KeyT) == 4;
    using K = typename std::conditional<k32, KeyT, uint32_t>::type;
    if (io.in == kInIota)
      return launch_scatter_io<K, BITS, THREADS, IPT, MIN_BLOCKS, kInIota, kOutRec>(io, offsets, g, shift, s);
    if (io.in == kInSplit)
      return launch_scatter_io<K, BITS, THREADS, IPT, MIN_BLOCKS, kInSplit, kOutRec>(io, offsets, g, shift, s);
    if (io.out == kOutRec)
      return launch_scatter_io<K, BITS, THREADS, IPT, MIN_BLOCKS, kInRec, kOutRec, TMA>(io, offsets, g, shift, s);
    return launch_scatter_io<K, BITS, THREADS, IPT, MIN_BLOCKS, kInRec, kOutSplit, TMA>(io, offsets, g, shift, s);
  }
}

template <typename KeyT, int BITS>
int run_pass(int variant, const PassIO& io, uint32_t* chunk_hist, uint32_t* totals,
             const ChunkGeom& g, int shift, cudaStream_t s, bool hist_ready) {
  constexpr bool k64 = sizeof(KeyT) == 8;
  constexpr int kBins = 1 << BITS;
  if (hist_ready) {
    // the producer of the keys (morton_hist_kernel) already counted this digit
  } else if (io.in == kInRec)  // the key is the high word of the record
    RT_CUDA(launch_pdl(chunk_hist_kernel<uint64_t, BITS>, g.chunks, kHistThreads, 0, s,
                       (const uint64_t*)io.recs_in, g, shift + 32, chunk_hist, totals));
  else
    RT_CUDA(launch_pdl(chunk_hist_kernel<KeyT, BITS>, g.chunks, kHistThreads, 0, s,
                       (const KeyT*)io.keys_in, g, shift, chunk_hist, totals));
  if (!hist_ready) RT_LAUNCHED();
  RT_CUDA(launch_pdl(chunk_offsets_kernel<BITS>, kBins / kOffWarps, kOffWarps * 32, 0, s, chunk_hist,
                     totals, g.chunks));
  RT_LAUNCHED();
  if constexpr (BITS == 10 && !k64) {
    if (variant == 2) return launch_scatter<KeyT, BITS, 512, 9, 2>(io, chunk_hist, g, shift, s);
    if (variant == 3) return launch_scatter<KeyT, BITS, 512, 9, 2, true>(io, chunk_hist, g, shift, s);
  }
  if (variant == 1)
    return launch_scatter<KeyT, BITS, 256, k64 ? 11 : 16, 4>(io, chunk_hist, g, shift, s);
  return launch_scatter<KeyT, BITS, 512, k64 ? 10 : 15, 2>(io, chunk_hist, g, shift, s);
}

template <typename KeyT>
int dispatch_pass(int bits, int variant, const PassIO& io, uint32_t* chunk_hist, uint32_t* totals,
                  const ChunkGeom& g, int shift, cudaStream_t s, bool hist_ready) {
  switch (bits) {
    case 8: return run_pass<KeyT, 8>(variant, io, chunk_hist, totals, g, shift, s, hist_ready);
    case 9: return run_pass<KeyT, 9>(variant, io, chunk_hist, totals, g, shift, s, hist_ready);
    case 10: return run_pass<KeyT, 10>(variant, io, chunk_hist, totals, g, shift, s, hist_ready);
    case 11: return run_pass<KeyT, 11>(variant, io, chunk_hist, totals, g, shift, s, hist_ready);
  }
  return fail(-2, "sort: unsupported digit width %d", bits);
}

}  // namespace

SortPlan sort_plan(uint32_t key_bits) {
  int bits = params().sort_bits;
  if (bits < 8 || bits > 11) bits = (key_bits == 30) ? 10 : 9;
  SortPlan p;
  p.bits = bits;
  p.passes = ((int)key_bits + bits - 1) / bits;
  return p;
}

// global bin totals of every pass: [passes][bins]
uint64_t sort_hist_words(uint32_t key_bits) {
  const SortPlan p = sort_plan(key_bits);
  return (uint64_t)p.passes << p.bits;
}

// per-chunk histograms / offsets of one pass (reused by every pass): [chunks][bins]
uint64_t sort_chunk_words(uint64_t n, uint32_t key_bits) {
  const SortPlan p = sort_plan(key_bits);
  return (uint64_t)make_geom(n, key_bits).chunks << p.bits;
}

int launch_sort(void* keys, uint64_t n, uint32_t key_bits, const SortWorkspace& ws,
                bool iota_payload, void** keys_out, uint32_t** vals_out, cudaStream_t s,
                bool first_hist_ready, int max_passes) {
  *keys_out = keys;
  *vals_out = ws.vals;
  if (n > ((1ull << 30) - 1)) return fail(-5, "sort: n = %llu beyond 2^30-1", (unsigned long long)n);
  if (n == 0) return 0;
  SortPlan plan = sort_plan(key_bits);
  // keys known to be narrower than key_bits (the ray coherence keys): only the low digits
  if (max_passes > 0 && max_passes < plan.passes) plan.passes = max_passes;
  const int variant = clamp_variant();
  const uint64_t bins = 1ull << plan.bits;
  const ChunkGeom g = make_geom(n, key_bits);
  if (ws.chunk_words < (uint64_t)g.chunks * bins)
    return fail(-2, "sort: chunk workspace too small");
  if (g.chunks > 32u * kOffMaxRounds) return fail(-2, "sort: %u chunks exceed the offsets scan", g.chunks);
  if (!first_hist_ready)
    RT_CUDA(cudaMemsetAsync(ws.hist, 0, sort_hist_words(key_bits) * sizeof(uint32_t), s));
  for (int p = 0; p < plan.passes; ++p) {
    const bool ready = first_hist_ready && p == 0;
    const bool fwd = (p & 1) == 0;
    uint32_t* totals = ws.hist + (size_t)p * bins;
    PassIO io;
    if (key_bits == 30) {
      // keys -> rec_a -> rec_b -> ... -> (keys, vals): records between the passes
      const bool first = p == 0, last = p == plan.passes - 1;
      if (first) {
        io.in = iota_payload ? kInIota : kInSplit;
        io.keys_in = keys;
        io.vals_in = ws.vals;
      } else {
        io.in = kInRec;
        io.recs_in = (const uint2*)(fwd ? ws.rec_b : ws.rec_a);
      }
      if (last) {
        io.out = kOutSplit;
        io.keys_out = keys;
        io.vals_out = ws.vals;
      } else {
        io.out = kOutRec;
        io.recs_out = (uint2*)(fwd ? ws.rec_a : ws.rec_b);
      }
      RT_TRY(dispatch_pass<uint32_t>(plan.bits, variant, io, ws.chunk_hist, totals, g,
                                     plan.bits * p, s, ready));
    } else {
      io.in = (iota_payload && p == 0) ? kInIota : kInSplit;
      io.keys_in = fwd ? keys : ws.keys_alt;
      io.keys_out = fwd ? ws.keys_alt : keys;
      io.vals_in = fwd ? ws.vals : ws.vals_alt;
      io.vals_out = fwd ? ws.vals_alt : ws.vals;
      RT_TRY(dispatch_pass<uint64_t>(plan.bits, variant, io, ws.chunk_hist, totals, g,
                                     plan.bits * p, s, ready));
    }
  }
  if (key_bits != 30 && (plan.passes & 1)) {  // odd pass count: result in the partner buffers
    *keys_out = ws.keys_alt;
    *vals_out = ws.vals_alt;
  }
  return 0;
}


// ---- 63-bit keys, hybrid order: LSD on the top 30 bits, ranks inside the runs -----------
// Seven 12-byte passes are what the 63-bit sort costs (8.0 ms at 100M).  The top 30 bits of
// a 63-bit Morton code are the 30-bit code, which the packed-record path sorts in three
// 8-byte passes; what is left is to order, inside every run of equal top bits, by the full
// key (the index order inside a run is already ascending, LSD being stable).  Runs are short
// where a scene needs 63 bits at all (terrain 100M: median 12, longest 98), so every element
// counts the members of its run that go before it.  A scene with a run longer than kMaxRun
// (e.g. all centroids equal) takes the seven-pass sort instead: same result either way.
constexpr uint32_t kMaxRun = 256;
constexpr int kRankThreads = 256;

__global__ void __launch_bounds__(256)
hi30_kernel(const uint64_t* __restrict__ k64, uint32_t n, uint32_t* __restrict__ k32) {
  pdl_enter();
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) k32[i] = (uint32_t)(k64[i] >> 33);
}

// keys into top-30 order; flags a run longer than kMaxRun (sorted: equal ends = equal run)
__global__ void __launch_bounds__(256)
gather_runs_kernel(const uint64_t* __restrict__ k64, const uint32_t* __restrict__ idx,
                   const uint32_t* __restrict__ k32s, uint32_t n, uint64_t* __restrict__ k64g,
                   uint32_t* __restrict__ long_run) {
  pdl_enter();
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  k64g[i] = k64[idx[i]];
  if (i + kMaxRun < n && k32s[i] == k32s[i + kMaxRun]) *long_run = 1u;
}

// One thread per element; the CTA's elements and kMaxRun neighbours on either side sit in
// shared memory.  Position inside the run = members with a smaller key, or an equal key
// and an earlier place.
__global__ void __launch_bounds__(kRankThreads)
rank_runs_kernel(const uint64_t* __restrict__ k64g, const uint32_t* __restrict__ idx,
                 const uint32_t* __restrict__ k32s, uint32_t n, uint64_t* __restrict__ keys_out,
                 uint32_t* __restrict__ vals_out, const uint32_t* __restrict__ long_run) {
  constexpr int kSpan = kRankThreads + 2 * (int)kMaxRun;
  constexpr int kWords = kSpan / 32;
  __shared__ uint64_t s_key[kSpan];
  __shared__ uint32_t s_head[kWords];  // bit = this slot starts a run (or lies outside the array)
  pdl_enter();
  if (*long_run != 0u) return;  // a run longer than the window: the caller sorts all 63 bits
  const long long base = (long long)blockIdx.x * kRankThreads - (long long)kMaxRun;
  for (int t = threadIdx.x; t < kSpan; t += kRankThreads) {  // kSpan is a multiple of the CTA: whole warps
    const long long g = base + t;
    const bool ok = g >= 0 && g < (long long)n;
    bool head = true;
    if (ok) {
      s_key[t] = k64g[g];
      head = (g == 0) || (k32s[g - 1] != k32s[g]);
    }
    const uint32_t word = __ballot_sync(0xFFFFFFFFu, head);
    if ((threadIdx.x & 31) == 0) s_head[t >> 5] = word;
  }
  __syncthreads();
  const uint32_t i = blockIdx.x * kRankThreads + threadIdx.x;
  if (i >= n) return;
  // the run's ends from the head bits: the nearest one at or before the slot, the next one
  // after it.  A run is at most kMaxRun long, so both lie inside the span; one or two words
  // each, where bisecting the top bits took nine dependent steps per end.
  const int c = (int)threadIdx.x + (int)kMaxRun;
  int w = c >> 5;
  const int bit = c & 31;
  uint32_t m = s_head[w] & (0xFFFFFFFFu >> (31 - bit));
  while (m == 0u) m = s_head[--w];
  const int lo = (w << 5) + 31 - __clz((int)m);
  w = c >> 5;
  m = (bit == 31) ? 0u : (s_head[w] & (0xFFFFFFFFu << (bit + 1)));
  while (m == 0u && w + 1 < kWords) m = s_head[++w];
  const int hi = m ? (w << 5) + __ffs((int)m) - 1 : kSpan;
  const uint64_t mine = s_key[c];
  // (key, place) < (mine, c) as the borrow of one 96-bit subtraction: branch-free, five
  // instructions per member where two 64-bit compares and the tie rule took eight
  const uint32_t mlo = (uint32_t)mine, mhi = (uint32_t)(mine >> 32);
  uint32_t before = 0;
#pragma unroll 4
  for (int j = lo; j < hi; ++j) {
    const uint64_t k = s_key[j];
    uint32_t t0, t1, t2, less;
    asm("sub.cc.u32 %0, %4, %5;\n\t"
        "subc.cc.u32 %1, %6, %7;\n\t"
        "subc.cc.u32 %2, %8, %9;\n\t"
        "subc.u32 %3, 0, 0;"
        : "=r"(t0), "=r"(t1), "=r"(t2), "=r"(less)
        : "r"((uint32_t)j), "r"((uint32_t)c), "r"((uint32_t)k), "r"(mlo), "r"((uint32_t)(k >> 32)), "r"(mhi));
    before -= less;  // less = 0xFFFFFFFF when (k, j) < (mine, c)
  }
  const uint32_t dst = i - (uint32_t)(c - lo) + before;
  keys_out[dst] = mine;
  vals_out[dst] = idx[i];
}

int launch_sort63_hybrid(uint64_t* k64, uint64_t n, const SortWorkspace& ws, uint32_t* k32,
                         uint32_t* long_run, void** keys_out, uint32_t** vals_out, cudaStream_t s,
                         bool k32_ready) {
  *keys_out = k64;
  *vals_out = ws.vals_alt;
  if (n > ((1ull << 30) - 1)) return fail(-5, "sort: n = %llu beyond 2^30-1", (unsigned long long)n);
  if (n == 0) return 0;
  const unsigned grid = (unsigned)((n + 255) / 256);
  // k32_ready: the Morton kernel stored the top bits and counted their first digit; long_run
  // was zeroed by the caller before the first kernel
  if (!k32_ready) {
    RT_CUDA(cudaMemsetAsync(long_run, 0, 4, s));
    RT_CUDA(launch_pdl(hi30_kernel, grid, 256, 0, s, (const uint64_t*)k64, (uint32_t)n, k32));
    RT_LAUNCHED();
  }
  void* sk32 = nullptr;
  uint32_t* sidx = nullptr;
  RT_TRY(launch_sort(k32, n, 30, ws, /*iota_payload=*/true, &sk32, &sidx, s, k32_ready));
  RT_CUDA(launch_pdl(gather_runs_kernel, grid, 256, 0, s, (const uint64_t*)k64, (const uint32_t*)sidx,
                     (const uint32_t*)sk32, (uint32_t)n, (uint64_t*)ws.keys_alt, long_run));
  RT_LAUNCHED();
  // no read-back here: rank_runs_kernel (and the caller's kernels after it) return at once
  // when gather_runs_kernel flagged a long run; the caller reads the flag when it next
  // synchronises and sorts all 63 bits instead
  RT_CUDA(launch_pdl(rank_runs_kernel, (unsigned)((n + kRankThreads - 1) / kRankThreads), kRankThreads,
                     0, s, (const uint64_t*)ws.keys_alt, (const uint32_t*)sidx, (const uint32_t*)sk32,
                     (uint32_t)n, k64, ws.vals_alt, (const uint32_t*)long_run));
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
