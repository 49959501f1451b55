// Multi-GPU Hamming top-2 (BASELINE config 4): the map-descriptor database is split into contiguous row
// shards, one process per GPU scans its shard, and the per-shard records (16 B per query) are merged with the
// order-free rule that equals ONE sequential scan with `if d<best .. else if d<second` (src/orbMatcher.cpp:327-333):
// best = min (distance, global row), second = 2nd order statistic of {best_g, second_g}.
//
// One kernel per call and GPU (exchange mode "p2p"):
//   scan     every CTA streams its chunk of the shard (same inner loop as k_top2_chunks) and folds its packed top-2
//            into one (best, second) pair per query with 64-bit atomicMin (order-free, see the kernel);
//   finish   the last CTA of a query block (ticket counter) turns the pairs into YdTop2 records with global row
//            indices, stores them into its own slot of the gather buffer AND into the same slot on every peer GPU
//            (plain stores through CUDA-IPC mapped peer memory over NVLink), then releases a per-source arrival
//            counter on every peer (red.release.sys);
//   merge    the same CTA acquires the peers' counters (bounded spin, default 10 s), merges the world x 512 records of
//            its query block and writes the merged YdTop2 + the ratio-test flag (src/orbMatcher.cpp:336).
// No second launch, no NCCL call on the data path.  NCCL (dlopen'ed, so that libydorb.so also loads where NCCL is
// absent) creates the communicator from the caller's ncclUniqueId and is used at set-up time to exchange the IPC
// handles; it is also the fallback exchange (mode "nccl": scan+finish kernel -> in-place ncclAllGather of 32 KB per
// rank on the matcher's stream -> merge+accept kernel) when peer mapping is unavailable or YDORB_SHARD_EXCHANGE=nccl.
//
// Gather buffers and counters are double-buffered by call parity: a rank can run at most one call ahead of a peer
// (its call k+1 cannot complete before the peer has pushed call k+1, which the peer does after finishing call k).
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "hamming.cuh"

namespace ydorb {

// ---- NCCL through dlopen ------------------------------------------------------------------------------------------
struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
};

static NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.handle ? &api : nullptr;
  tried = true;
  const char* names[] = {getenv("YDORB_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    if (!n || !*n) continue;
    api.handle = dlopen(n, RTLD_NOW | RTLD_LOCAL);  // an already-loaded libnccl.so.2 (e.g. torch's) is reused
    if (api.handle) break;
  }
  if (!api.handle) {
    set_error(std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : ""));
    return nullptr;
  }
#define YD_SYM(field, name)                                                  \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, name)); \
  if (!api.field) {                                                          \
    set_error(std::string("NCCL symbol missing: ") + name);                  \
    api.handle = nullptr;                                                    \
    return nullptr;                                                          \
  }
  YD_SYM(GetUniqueId, "ncclGetUniqueId")
  YD_SYM(CommInitRank, "ncclCommInitRank")
  YD_SYM(CommDestroy, "ncclCommDestroy")
  YD_SYM(AllGather, "ncclAllGather")
  YD_SYM(AllReduce, "ncclAllReduce")
  YD_SYM(GetErrorString, "ncclGetErrorString")
  YD_SYM(GetVersion, "ncclGetVersion")
#undef YD_SYM
  return &api;
}

static int nccl_fail(NcclApi* api, ncclResult_t r, const char* what) {
  set_error(std::string("NCCL error in ") + what + ": " + (api && api->GetErrorString ? api->GetErrorString(r) : "?"));
  return YDORB_E_CUDA;
}
#define YD_NCCL(api, expr)                                        \
  do {                                                            \
    ncclResult_t _r = (expr);                                     \
    if (_r != ncclSuccess) return nccl_fail(api, _r, #expr);      \
  } while (0)

// NCCL exchange mode: merge + ratio test after the all-gather
__global__ void k_top2_merge_accept(const YdTop2* __restrict__ shards, int nshards, int nq, YdTop2* __restrict__ out,
                                    uint8_t* __restrict__ accept, float ratio, int th) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  unsigned long long b = ~0ull, s = ~0ull;
  for (int c = 0; c < nshards; c++) {
    const YdTop2 t = shards[(size_t)c * nq + qi];
    top2_push(top2_key(t.best_dist, t.best_idx), b, s);
    top2_push(top2_key(t.second_dist, t.second_idx), b, s);
  }
  const YdTop2 o = top2_unpack(b, s);
  out[qi] = o;
  if (accept) accept[qi] = ratio_accept(o, ratio, th);
}

}  // namespace ydorb

using namespace ydorb;

struct YdShard {   // one rank's rows of the map-descriptor database, resident on its GPU
  YdMatcher* m;
  uint8_t* d_rows;
  int n_rows, idx_base;
};

struct YdComm {
  YdMatcher* m;
  NcclApi* api;
  ncclComm_t comm;
  int rank, world, max_queries;
  int mode;                       // 0 single, 1 nccl, 2 p2p
  uint8_t* d_block;               // one allocation: counters | gather[2][world][max_queries] | tickets | status
  size_t block_bytes, gather_off, tickets_off, status_off, pairs_off;
  uint8_t* peer_block[kShardMaxWorld];  // peers' d_block mapped into this process (mode 2); own entry = d_block
  uint32_t arrivals_expected[2];  // per parity: counter value after the calls made so far
  unsigned long long calls;
  unsigned long long timeout_ns;
  YdTop2* d_out;                  // staging of the host entry point
  uint8_t* d_accept;
};

namespace {

constexpr size_t kCounterBytes = 2 * kShardMaxWorld * sizeof(uint32_t);  // [parity][source]

YdTop2* gather_of(uint8_t* block, const YdComm* c, int parity) {
  return reinterpret_cast<YdTop2*>(block + c->gather_off) + (size_t)parity * c->world * c->max_queries;
}
uint32_t* arrivals_of(uint8_t* block, int parity) { return reinterpret_cast<uint32_t*>(block) + parity * kShardMaxWorld; }

void comm_free(YdComm* c) {
  if (!c) return;
  cudaSetDevice(c->m->device);
  cudaStreamSynchronize(c->m->stream);
  for (int r = 0; r < c->world; r++)
    if (r != c->rank && c->peer_block[r]) cudaIpcCloseMemHandle(c->peer_block[r]);
  if (c->d_block) cudaFree(c->d_block);
  if (c->d_out) cudaFree(c->d_out);
  if (c->comm && c->api) c->api->CommDestroy(c->comm);
  delete c;
}

}  // namespace

extern "C" {

int ydorb_comm_unique_id(uint8_t* id, int bytes) {
  YD_REQUIRE(id && bytes >= (int)sizeof(ncclUniqueId), "need a buffer of YDORB_COMM_ID_BYTES bytes");
  NcclApi* api = nccl_api();
  if (!api) return YDORB_E_CUDA;
  ncclUniqueId u;
  YD_NCCL(api, api->GetUniqueId(&u));
  memset(id, 0, (size_t)bytes);
  memcpy(id, &u, sizeof(u));
  return YDORB_OK;
}

int ydorb_comm_create(YdMatcher* m, const uint8_t* id, int rank, int world_size, int max_queries, YdComm** out) {
  YD_REQUIRE(m && out && max_queries > 0, "bad argument");
  YD_REQUIRE(world_size >= 1 && world_size <= kShardMaxWorld && rank >= 0 && rank < world_size, "bad rank / world size");
  YD_REQUIRE(world_size == 1 || id, "a communicator over more than one rank needs rank 0's id (ydorb_comm_unique_id)");
  YD_CUDA(cudaSetDevice(m->device));
  YdComm* c = new YdComm();
  memset(static_cast<void*>(c), 0, sizeof(*c));
  c->m = m;
  c->rank = rank;
  c->world = world_size;
  c->max_queries = max_queries;
  c->timeout_ns = 10ull * 1000000000ull;
  if (const char* e = getenv("YDORB_SHARD_TIMEOUT_MS")) c->timeout_ns = (unsigned long long)std::max(1, atoi(e)) * 1000000ull;
  const int qblocks = (max_queries + kTopQblock - 1) / kTopQblock;
  c->gather_off = (kCounterBytes + 255) / 256 * 256;
  c->tickets_off = c->gather_off + (sizeof(YdTop2) * 2 * (size_t)world_size * max_queries + 255) / 256 * 256;
  c->status_off = c->tickets_off + (sizeof(unsigned int) * qblocks + 255) / 256 * 256;
  c->pairs_off = c->status_off + 256;
  c->block_bytes = c->pairs_off + 2 * sizeof(unsigned long long) * (size_t)max_queries;
  cudaError_t e = cudaMalloc(&c->d_block, c->block_bytes);
  if (e == cudaSuccess) e = cudaMemset(c->d_block, 0, c->pairs_off);
  if (e == cudaSuccess) e = cudaMemset(c->d_block + c->pairs_off, 0xFF, c->block_bytes - c->pairs_off);
  if (e == cudaSuccess) e = cudaMalloc(&c->d_out, (sizeof(YdTop2) + 1) * (size_t)max_queries + 256);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { comm_free(c); return cuda_fail(e, "cudaMalloc (comm buffers)", __FILE__, __LINE__); }
  c->d_accept = reinterpret_cast<uint8_t*>(c->d_out) + sizeof(YdTop2) * (size_t)max_queries;
  c->peer_block[rank] = c->d_block;
  if (world_size == 1) {
    c->mode = 0;
    *out = c;
    return YDORB_OK;
  }
  c->api = nccl_api();
  if (!c->api) { comm_free(c); return YDORB_E_CUDA; }
  ncclUniqueId u;
  memcpy(&u, id, sizeof(u));
  ncclResult_t r = c->api->CommInitRank(&c->comm, world_size, u, rank);
  if (r != ncclSuccess) { c->comm = nullptr; comm_free(c); return nccl_fail(nccl_api(), r, "ncclCommInitRank"); }
  c->mode = 1;
  // ---- try to map every peer's block (CUDA IPC); all ranks must agree, so the outcome is all-reduced (min)
  const char* want = getenv("YDORB_SHARD_EXCHANGE");
  int ok = !(want && strcmp(want, "nccl") == 0);
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (ok && cudaIpcGetMemHandle(&mine, c->d_block) != cudaSuccess) { ok = 0; cudaGetLastError(); }
  cudaIpcMemHandle_t* d_handles = nullptr;
  int* d_ok = nullptr;
  std::vector<cudaIpcMemHandle_t> handles((size_t)world_size);
  e = cudaMalloc(&d_handles, sizeof(cudaIpcMemHandle_t) * world_size);
  if (e == cudaSuccess) e = cudaMalloc(&d_ok, sizeof(int));
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_handles + rank, &mine, sizeof(mine), cudaMemcpyHostToDevice, m->stream);
  if (e != cudaSuccess) { cudaFree(d_handles); cudaFree(d_ok); comm_free(c); return cuda_fail(e, "comm set-up", __FILE__, __LINE__); }
  r = c->api->AllGather(d_handles + rank, d_handles, sizeof(cudaIpcMemHandle_t), ncclInt8, c->comm, m->stream);
  if (r == ncclSuccess) {
    cudaMemcpyAsync(handles.data(), d_handles, sizeof(cudaIpcMemHandle_t) * world_size, cudaMemcpyDeviceToHost, m->stream);
    e = cudaStreamSynchronize(m->stream);
  }
  if (r != ncclSuccess || e != cudaSuccess) {
    cudaFree(d_handles); cudaFree(d_ok); comm_free(c);
    return r != ncclSuccess ? nccl_fail(nccl_api(), r, "ncclAllGather (set-up)") : cuda_fail(e, "comm set-up", __FILE__, __LINE__);
  }
  for (int p = 0; ok && p < world_size; p++) {
    if (p == rank) continue;
    void* ptr = nullptr;
    if (cudaIpcOpenMemHandle(&ptr, handles[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      ok = 0;
      cudaGetLastError();
      break;
    }
    c->peer_block[p] = static_cast<uint8_t*>(ptr);
  }
  cudaMemcpyAsync(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice, m->stream);
  r = c->api->AllReduce(d_ok, d_ok, 1, ncclInt32, ncclMin, c->comm, m->stream);
  int all_ok = 0;
  cudaMemcpyAsync(&all_ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, m->stream);
  e = cudaStreamSynchronize(m->stream);
  cudaFree(d_handles);
  cudaFree(d_ok);
  if (r != ncclSuccess || e != cudaSuccess) {
    comm_free(c);
    return r != ncclSuccess ? nccl_fail(nccl_api(), r, "ncclAllReduce (set-up)") : cuda_fail(e, "comm set-up", __FILE__, __LINE__);
  }
  if (all_ok) {
    c->mode = 2;
  } else {
    for (int p = 0; p < world_size; p++)
      if (p != rank && c->peer_block[p]) { cudaIpcCloseMemHandle(c->peer_block[p]); c->peer_block[p] = nullptr; }
  }
  *out = c;
  return YDORB_OK;
}

void ydorb_comm_destroy(YdComm* c) { comm_free(c); }
int ydorb_comm_rank(const YdComm* c) { return c ? c->rank : -1; }
int ydorb_comm_size(const YdComm* c) { return c ? c->world : 0; }
/* 0 single GPU, 1 NCCL all-gather exchange, 2 peer-memory push inside the scan kernel */
int ydorb_comm_exchange_mode(const YdComm* c) { return c ? c->mode : -1; }

int ydorb_hamming_top2_sharded_device(YdComm* c, const uint8_t* d_queries, int nq, const uint8_t* d_db_shard, int n_rows,
                                      int idx_base, float ratio, int th, YdTop2* d_out, uint8_t* d_accept) {
  YD_REQUIRE(c && d_out, "null argument");
  YD_REQUIRE(nq >= 0 && nq <= c->max_queries && n_rows >= 0, "query count exceeds the communicator's max_queries");
  if (nq == 0) return YDORB_OK;
  YD_REQUIRE(d_queries && (d_db_shard || n_rows == 0), "null descriptor pointer");
  YD_REQUIRE(((uintptr_t)d_queries & 15) == 0 && ((uintptr_t)d_db_shard & 15) == 0, "descriptor matrices must be 16-byte aligned");
  YdMatcher* m = c->m;
  YD_CUDA(cudaSetDevice(m->device));
  int qblocks, nchunks, chunk_rows;
  int rc = top2_grid(m, nq, n_rows, &qblocks, &nchunks, &chunk_rows);
  if (rc != YDORB_OK) return rc;
  const int parity = (int)(c->calls & 1);
  ShardDev a;
  memset(&a, 0, sizeof(a));
  a.rank = c->rank;
  a.world = c->world;
  a.mode = c->mode;
  a.idx_base = idx_base;
  a.best = reinterpret_cast<unsigned long long*>(c->d_block + c->pairs_off);
  a.second = a.best + c->max_queries;
  a.tickets = reinterpret_cast<unsigned int*>(c->d_block + c->tickets_off);
  a.gather = gather_of(c->d_block, c, parity);
  a.arrivals = arrivals_of(c->d_block, parity);
  a.status = reinterpret_cast<int*>(c->d_block + c->status_off);
  a.timeout_ns = c->timeout_ns;
  a.out = d_out;
  a.accept = d_accept;
  a.ratio = ratio;
  a.th = th;
  if (c->mode == 2) {
    c->arrivals_expected[parity] += (uint32_t)qblocks;
    a.target = c->arrivals_expected[parity];
    for (int p = 0; p < c->world; p++) {
      a.peer_gather[p] = gather_of(c->peer_block[p], c, parity);
      a.peer_arrivals[p] = arrivals_of(c->peer_block[p], parity);
    }
  }
  c->calls++;
  rc = launch_top2_sharded(m, d_queries, nq, d_db_shard, n_rows, a);
  if (rc != YDORB_OK) return rc;
  if (c->mode == 1) {
    // in-place all-gather: this rank's records already sit in slot `rank` of the gather buffer
    YdTop2* g = a.gather;
    YD_NCCL(c->api, c->api->AllGather(g + (size_t)c->rank * nq, g, (size_t)nq * 4, ncclInt32, c->comm, m->stream));
    k_top2_merge_accept<<<(nq + 127) / 128, 128, 0, m->stream>>>(g, c->world, nq, d_out, d_accept, ratio, th);
    count_launches(1);
    YD_CUDA(cudaGetLastError());
  }
  return YDORB_OK;
}

int ydorb_comm_check(YdComm* c) {
  YD_REQUIRE(c, "null communicator");
  YD_CUDA(cudaSetDevice(c->m->device));
  int st = 0;
  YD_CUDA(cudaMemcpyAsync(&st, c->d_block + c->status_off, sizeof(int), cudaMemcpyDeviceToHost, c->m->stream));
  YD_CUDA(cudaStreamSynchronize(c->m->stream));
  if (st & kShardStatusTimeout) {
    set_error("sharded top-2: a peer's records did not arrive within the time limit (YDORB_SHARD_TIMEOUT_MS); "
              "every rank must make the same sequence of sharded calls");
    return YDORB_E_STATE;
  }
  return YDORB_OK;
}

int ydorb_shard_create(YdMatcher* m, const uint8_t* rows, int n_rows, int idx_base, YdShard** out) {
  YD_REQUIRE(m && out && n_rows >= 0 && (rows || n_rows == 0), "bad argument");
  YD_CUDA(cudaSetDevice(m->device));
  YdShard* s = new YdShard();
  s->m = m;
  s->n_rows = n_rows;
  s->idx_base = idx_base;
  s->d_rows = nullptr;
  cudaError_t e = cudaMalloc(&s->d_rows, (size_t)std::max(n_rows, 1) * 32);
  if (e == cudaSuccess && n_rows > 0) e = cudaMemcpyAsync(s->d_rows, rows, (size_t)n_rows * 32, cudaMemcpyHostToDevice, m->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
  if (e != cudaSuccess) {
    if (s->d_rows) cudaFree(s->d_rows);
    delete s;
    return cuda_fail(e, "shard upload", __FILE__, __LINE__);
  }
  *out = s;
  return YDORB_OK;
}

void ydorb_shard_destroy(YdShard* s) {
  if (!s) return;
  cudaSetDevice(s->m->device);
  if (s->d_rows) cudaFree(s->d_rows);
  delete s;
}

int ydorb_hamming_top2_sharded(YdComm* c, const YdShard* shard, const uint8_t* queries, int nq, float ratio, int th,
                               YdTop2* out, uint8_t* accept) {
  YD_REQUIRE(c && shard && out && queries, "null argument");
  YD_REQUIRE(shard->m == c->m, "shard and communicator belong to different matchers");
  const uint8_t* d_db_shard = shard->d_rows;
  const int n_rows = shard->n_rows, idx_base = shard->idx_base;
  YD_REQUIRE(nq >= 0 && nq <= c->max_queries, "query count exceeds the communicator's max_queries");
  if (nq == 0) return YDORB_OK;
  YdMatcher* m = c->m;
  YD_CUDA(cudaSetDevice(m->device));
  int rc = matcher_reserve_io(m, (size_t)nq * 32);
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(m->d_io, queries, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream));
  rc = ydorb_hamming_top2_sharded_device(c, static_cast<const uint8_t*>(m->d_io), nq, d_db_shard, n_rows, idx_base, ratio,
                                         th, c->d_out, accept ? c->d_accept : nullptr);
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(out, c->d_out, sizeof(YdTop2) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
  if (accept) YD_CUDA(cudaMemcpyAsync(accept, c->d_accept, (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
  return ydorb_comm_check(c);
}

}  // extern "C"
