// hamming_static.cu -- ydorb_hamming_top2_device_noalloc: the dense Hamming top-2 scan (k_top2_sharded, hamming.cu) for
// callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// ydorb_hamming_top2_device keeps its per-query fold state ((best, second) u64 pairs + one ticket counter per query
// block) in a grow-only cudaMalloc'ed buffer of the matcher handle.  Here that state lives in a file-scope device array
// of fixed size (YDORB_NOALLOC_MAX_QUERIES queries, ~256 KB, held by every GPU that loads the library until exit), so
// the call reaches no cudaMalloc / cudaFree / allocating library routine.  Every call and every handle on a device
// share the array: calls that overlap on the GPU (two streams, two threads) race.  The array keeps the last call's
// contents, so every call first resets the part it uses (k_static_reset, on the same stream).
//
// In host code `g_static_pairs` is only the host-side stand-in of the device array, not an address a kernel can use:
// its device address is fetched once per device with cudaGetSymbolAddress (that device current) and passed on.
#include <mutex>

#include "hamming.cuh"

namespace {

constexpr int kStaticMaxQ = YDORB_NOALLOC_MAX_QUERIES;
constexpr int kStaticMaxQb = (kStaticMaxQ + ydorb::kTopQblock - 1) / ydorb::kTopQblock;
// [0, MaxQ) best | [MaxQ, 2 MaxQ) second | then MaxQb u32 ticket counters (two per u64 word)
constexpr int kStaticWords = 2 * kStaticMaxQ + (kStaticMaxQb + 1) / 2;

}  // namespace

static __device__ __align__(256) unsigned long long g_static_pairs[kStaticWords];

namespace {

// device code may name the array directly: the fold state of the first nq queries back to "nothing seen yet"
__global__ void k_static_reset(int nq, int qblocks) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nq) {
    g_static_pairs[i] = ~0ull;
    g_static_pairs[kStaticMaxQ + i] = ~0ull;
  }
  if (i < qblocks) reinterpret_cast<unsigned int*>(g_static_pairs + 2 * kStaticMaxQ)[i] = 0u;
}

constexpr int kMaxDevices = 64;
std::mutex g_addr_mutex;
unsigned long long* g_addr[kMaxDevices];  // device address of g_static_pairs on device d, fetched on first use

// the caller has made `device` current
int static_pairs_address(int device, unsigned long long** out) {
  YD_REQUIRE(device >= 0 && device < kMaxDevices, "device index out of range");
  std::lock_guard<std::mutex> lock(g_addr_mutex);
  if (!g_addr[device]) {
    unsigned long long* p = nullptr;
    YD_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_pairs));
    YD_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
    g_addr[device] = p;
  }
  *out = g_addr[device];
  return YDORB_OK;
}

}  // namespace

using namespace ydorb;

extern "C" int ydorb_hamming_top2_device_noalloc(YdMatcher* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, int ndb,
                                                 int idx_base, YdTop2* d_out) {
  YD_REQUIRE(m && d_out, "null argument");
  YD_REQUIRE(nq >= 0 && ndb >= 0, "negative size");
  if (nq > kStaticMaxQ) {
    set_error("ydorb_hamming_top2_device_noalloc: nq = " + std::to_string(nq) + " exceeds YDORB_NOALLOC_MAX_QUERIES = " +
              std::to_string(kStaticMaxQ) + " (the static scratch array); use ydorb_hamming_top2_device or split the queries");
    return YDORB_E_CAPACITY;
  }
  if (nq == 0) return YDORB_OK;
  YD_REQUIRE(d_queries && (d_db || ndb == 0), "null descriptor pointer");
  YD_REQUIRE(((uintptr_t)d_queries & 15) == 0 && ((uintptr_t)d_db & 15) == 0, "descriptor matrices must be 16-byte aligned");
  YD_CUDA(cudaSetDevice(m->device));
  unsigned long long* pairs = nullptr;
  int rc = static_pairs_address(m->device, &pairs);
  if (rc != YDORB_OK) return rc;
  const int qblocks = (nq + kTopQblock - 1) / kTopQblock;
  k_static_reset<<<(nq + 255) / 256, 256, 0, m->stream>>>(nq, qblocks);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  ShardDev a;
  memset(&a, 0, sizeof(a));
  a.world = 1;
  a.mode = 0;
  a.idx_base = idx_base;
  a.best = pairs;
  a.second = pairs + kStaticMaxQ;
  a.tickets = reinterpret_cast<unsigned int*>(pairs + 2 * kStaticMaxQ);
  a.out = d_out;
  return launch_top2_sharded(m, d_queries, nq, d_db, ndb, a);
}
