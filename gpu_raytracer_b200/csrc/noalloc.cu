// noalloc.cu — rt_sort_pairs_static: the build's radix sort of (key, payload) pairs with its
// scratch memory in a file-scope device array instead of pool allocations.
//
// No counterpart in the reference; added for callers that
// cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  The entry point calls no cudaMalloc / cudaFree / cudaMallocAsync and no library
// routine that allocates; it enqueues memsets and kernels on the caller's stream and returns.
//
// The scratch (record ping-pong buffers, bin totals, per-chunk histograms of radix_sort.cu)
// lives in g_scratch below: kScratchBytes of device memory that EVERY GPU which loads the
// library holds until the process exits.  That is what bounds the size: at most
// RT_SORT_STATIC_MAX_PAIRS pairs of 30-bit keys.  All calls share the array and it keeps the
// last call's contents; the sort clears the one part that must start at zero (the bin totals)
// itself.  Calls that overlap on one GPU (different streams, or host threads) race on it.
//
// In host code the array's name is a host-side stand-in, not a device address: its device
// address is asked for once per device, with that device current, through
// cudaGetSymbolAddress, and that pointer is what the kernels get.
#include <mutex>

#include "host.h"

namespace rt {
namespace {

constexpr uint64_t kMaxPairs = RT_SORT_STATIC_MAX_PAIRS;
constexpr uint64_t kMaxChunks = 768;                      // radix_sort.cu: chunk_offsets_kernel's limit
constexpr size_t kRecBytes = kMaxPairs * 8;               // one buffer of packed (key << 32 | payload) records
constexpr size_t kHistBytes = 3 * 1024 * 4;               // bin totals: 3 passes x 1024 bins
constexpr size_t kChunkBytes = kMaxChunks * 1024 * 4;     // per-chunk digit counts of one pass
constexpr size_t kOffRecA = 0, kOffRecB = kRecBytes, kOffHist = 2 * kRecBytes,
                 kOffChunk = kOffHist + ((kHistBytes + 255) & ~(size_t)255);
constexpr size_t kScratchBytes = kOffChunk + kChunkBytes;  // 19.9 MB

static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

constexpr int kMaxDev = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[kMaxDev] = {};

// device address of g_scratch on the CURRENT device (asked for once per device)
int scratch_address(unsigned char** out) {
  int dev = 0;
  RT_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) return fail(RT_ERR_INVALID, "rt_sort_pairs_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    RT_CUDA(cudaGetSymbolAddress(&p, g_scratch));
    g_scratch_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_ptr[dev];
  return 0;
}

}  // namespace
}  // namespace rt

using namespace rt;

extern "C" int rt_sort_pairs_static(uint32_t* keys, uint32_t* vals, uint64_t n, void* stream) {
  if (n > kMaxPairs)
    return fail(RT_ERR_LIMIT, "rt_sort_pairs_static: %llu pairs exceed the static scratch (RT_SORT_STATIC_MAX_PAIRS = %llu)",
                (unsigned long long)n, (unsigned long long)kMaxPairs);
  if (n == 0) return RT_OK;
  if (!keys || !vals) return fail(RT_ERR_INVALID, "rt_sort_pairs_static: NULL buffer");
  RT_TRY(ensure_device());
  int dev = 0, dk = -1, dv = -1;
  RT_CUDA(cudaGetDevice(&dev));
  if (!classify(keys, &dk) || !classify(vals, &dv) || dk != dev || dv != dev)
    return fail(RT_ERR_INVALID, "rt_sort_pairs_static: keys and vals must be device memory of the current device %d", dev);
  if (sort_plan(30).bits != 10 || sort_plan(30).passes != 3)
    return fail(RT_ERR_INVALID, "rt_sort_pairs_static: needs the default digit schedule (rt_set_param \"sort.bits\" 0)");
  unsigned char* base = nullptr;
  RT_TRY(scratch_address(&base));
  SortWorkspace ws;
  ws.rec_a = base + kOffRecA;
  ws.rec_b = base + kOffRecB;
  ws.hist = reinterpret_cast<uint32_t*>(base + kOffHist);
  ws.chunk_hist = reinterpret_cast<uint32_t*>(base + kOffChunk);
  ws.chunk_words = kChunkBytes / 4;
  ws.vals = vals;
  if (sort_chunk_words(n, 30) > ws.chunk_words)
    return fail(RT_ERR_LIMIT, "rt_sort_pairs_static: this device needs %llu chunk words, the scratch holds %llu",
                (unsigned long long)sort_chunk_words(n, 30), (unsigned long long)ws.chunk_words);
  void* ko = nullptr;
  uint32_t* vo = nullptr;
  // three packed passes: keys -> rec_a -> rec_b -> (keys, vals); the result is in place
  RT_TRY(launch_sort(keys, n, 30, ws, /*iota_payload=*/false, &ko, &vo, (cudaStream_t)stream));
  if (ko != keys || vo != vals) return fail(RT_ERR_CUDA, "rt_sort_pairs_static: result not in place");
  return RT_OK;
}
