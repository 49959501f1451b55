// M0  OrbMatcher::computeDescriptorsDistance (src/orbMatcher.cpp:11-23) and the dense top-2 scan that every
// BoW / brute-force matcher runs around it (src/orbMatcher.cpp:319-334: `if d<best {second=best; best=d}
// else if d<second {second=d}`).
//
// Descriptors are 32 B = 2 x uint4.  A thread keeps 4 query descriptors in registers (32 words), the CTA streams
// its slice of the database through shared memory with coalesced uint4 loads and every thread reads each database
// row as two broadcast LDS.128.  Distance = 8 x (XOR, POPC).  The running top-2 is two packed words
// (distance << 23 | row-in-chunk): best = min, second = min(second, max(best_before, key)) -- with the row in the
// low bits this reproduces the sequential rule exactly, including "first candidate wins ties".
// The database is split into chunks across blockIdx.y so the grid fills 148 SMs; every CTA folds its chunk's top-2 into
// one (best, second) pair per query with 64-bit atomicMin (order-free), and the last CTA of a query block unpacks the
// pairs -- one launch (k_top2_sharded).  The same kernel carries the cross-GPU exchange and merge (shard.cu).
#include <time.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "hamming.cuh"

namespace ydorb {

__global__ void k_hamming_pairs(const uint4* __restrict__ a, const uint4* __restrict__ b, int n, int32_t* __restrict__ dist) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 a0 = a[2 * i], a1 = a[2 * i + 1], b0 = b[2 * i], b1 = b[2 * i + 1];
  dist[i] = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
            __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(kTopThreads)
k_top2_sharded(const uint4* __restrict__ q, const uint4* __restrict__ db, int ndb, const ShardDev a) {
  __shared__ uint4 s_db[2 * kTopTile];
  __shared__ int s_flag;
  const int tid = threadIdx.x;
  const int nq = a.nq;
  const int q0 = blockIdx.x * kTopQblock;
  const int chunk = blockIdx.y;
  const int row_begin = chunk * a.chunk_rows;
  const int row_end = min(row_begin + a.chunk_rows, ndb);
  {
    uint32_t qa[kTopQpt][8];
    uint32_t best[kTopQpt], second[kTopQpt];
#pragma unroll
    for (int k = 0; k < kTopQpt; k++) {
      const int qi = q0 + k * kTopThreads + tid;
      uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
      if (qi < nq) { lo = q[2 * qi]; hi = q[2 * qi + 1]; }
      qa[k][0] = lo.x; qa[k][1] = lo.y; qa[k][2] = lo.z; qa[k][3] = lo.w;
      qa[k][4] = hi.x; qa[k][5] = hi.y; qa[k][6] = hi.z; qa[k][7] = hi.w;
      best[k] = kNoKey;
      second[k] = kNoKey;
    }
    for (int t0 = row_begin; t0 < row_end; t0 += kTopTile) {
      const int rows = min(kTopTile, row_end - t0);
      __syncthreads();
      for (int i = tid; i < 2 * rows; i += kTopThreads) s_db[i] = db[2 * (size_t)t0 + i];
      __syncthreads();
      const uint32_t local0 = (uint32_t)(t0 - row_begin);
      int r = 0;
#pragma unroll 2
      for (; r + 1 < rows; r += 2) {  // rows in pairs: 5 min/max per two keys (top2_push2)
        const uint4 lo = s_db[2 * r], hi = s_db[2 * r + 1], lo2 = s_db[2 * r + 2], hi2 = s_db[2 * r + 3];
#pragma unroll
        for (int k = 0; k < kTopQpt; k++)
          top2_push2(ham256_key(qa[k], lo, hi, local0 + (uint32_t)r), ham256_key(qa[k], lo2, hi2, local0 + (uint32_t)r + 1u),
                     best[k], second[k]);
      }
      if (r < rows) {
        const uint4 lo = s_db[2 * r], hi = s_db[2 * r + 1];
#pragma unroll
        for (int k = 0; k < kTopQpt; k++) {
          const uint32_t key = ham256_key(qa[k], lo, hi, local0 + (uint32_t)r);
          second[k] = min(second[k], max(best[k], key));
          best[k] = min(best[k], key);
        }
      }
    }
    // fold this chunk's top-2 into the per-query global pair (B, S) with 64-bit atomicMin on (distance << 32 | global row):
    // inserting x as  old = atomicMin(B, x); atomicMin(S, max(old, x))  leaves B = minimum and S = second order statistic
    // of everything inserted, in any order (every value pushed to S lost against a smaller one, and the runner-up loses
    // against the minimum whichever of the two arrives first).  A stale read of S prunes keys that cannot matter.
    const uint32_t row0 = (uint32_t)(a.idx_base + row_begin);
#pragma unroll
    for (int k = 0; k < kTopQpt; k++) {
      const int qi = q0 + k * kTopThreads + tid;
      if (qi >= nq) continue;
      unsigned long long* B = a.best + qi;
      unsigned long long* S = a.second + qi;
      const unsigned long long cut = *reinterpret_cast<volatile unsigned long long*>(S);
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const uint32_t kk = h ? second[k] : best[k];
        if (kk == kNoKey) continue;
        const unsigned long long x = ((unsigned long long)(kk >> 23) << 32) | (unsigned long long)(row0 + (kk & 0x7FFFFFu));
        if (x >= cut) continue;
        const unsigned long long old = atomicMin(B, x);
        atomicMin(S, max(old, x));
      }
    }
  }
  // ---- last CTA of this query block finishes it
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned int t = atomicAdd(&a.tickets[blockIdx.x], 1u);
    s_flag = (t == gridDim.y - 1);
    if (s_flag) a.tickets[blockIdx.x] = 0;  // ready for the next call (stream order)
  }
  __syncthreads();
  if (!s_flag) return;
  __threadfence();
  YdTop2 mine[kTopQpt];
#pragma unroll
  for (int k = 0; k < kTopQpt; k++) {
    const int qi = q0 + k * kTopThreads + tid;
    if (qi >= nq) continue;
    const unsigned long long b = __ldcg(a.best + qi), s = __ldcg(a.second + qi);
    a.best[qi] = ~0ull;    // ready for the next call
    a.second[qi] = ~0ull;
    mine[k] = top2_unpack(b, s);
    if (a.mode == 0) {
      a.out[qi] = mine[k];
      if (a.accept) a.accept[qi] = ratio_accept(mine[k], a.ratio, a.th);
    } else {
      reinterpret_cast<int4*>(a.gather)[(size_t)a.rank * nq + qi] = *reinterpret_cast<const int4*>(&mine[k]);
    }
  }
  if (a.mode != 2) return;
  // ---- push this query block's records to every peer, then release the arrival counters
  for (int p = 1; p < a.world; p++) {
    const int peer = (a.rank + p) % a.world;  // staggered so that the ranks do not all hit the same GPU first
    int4* dst = reinterpret_cast<int4*>(a.peer_gather[peer]) + (size_t)a.rank * nq;
#pragma unroll
    for (int k = 0; k < kTopQpt; k++) {
      const int qi = q0 + k * kTopThreads + tid;
      if (qi < nq) dst[qi] = *reinterpret_cast<const int4*>(&mine[k]);
    }
  }
  __threadfence_system();
  __syncthreads();
  if (tid > 0 && tid < a.world) red_release_sys_add(a.peer_arrivals[(a.rank + tid) % a.world] + a.rank, 1u);
  // ---- wait for the peers' records of this call (bounded)
  if (tid == 0) s_flag = 0;
  __syncthreads();
  if (tid > 0 && tid < a.world) {
    const uint32_t* ctr = a.arrivals + (a.rank + tid) % a.world;
    const unsigned long long t0 = global_ns();
    while ((int)(ld_acquire_sys(ctr) - a.target) < 0) {
      if (global_ns() - t0 > a.timeout_ns) {
        atomicOr(&s_flag, 1);
        break;
      }
      __nanosleep(200);
    }
  }
  __syncthreads();
  if (s_flag) {
    if (tid == 0) atomicOr(a.status, kShardStatusTimeout);
    return;
  }
  // ---- merge: one sequential scan over the concatenation of the shards
#pragma unroll
  for (int k = 0; k < kTopQpt; k++) {
    const int qi = q0 + k * kTopThreads + tid;
    if (qi >= nq) continue;
    unsigned long long b = ~0ull, s = ~0ull;
    for (int c = 0; c < a.world; c++) {
      YdTop2 t = mine[k];
      if (c != a.rank) {
        const int4 v = __ldcg(reinterpret_cast<const int4*>(a.gather) + (size_t)c * nq + qi);
        t = *reinterpret_cast<const YdTop2*>(&v);
      }
      top2_push(top2_key(t.best_dist, t.best_idx), b, s);
      top2_push(top2_key(t.second_dist, t.second_idx), b, s);
    }
    const YdTop2 o = top2_unpack(b, s);
    a.out[qi] = o;
    if (a.accept) a.accept[qi] = ratio_accept(o, a.ratio, a.th);
  }
}

// shard merge: result of one sequential scan over the concatenation of the shards
__global__ void k_top2_merge(const YdTop2* __restrict__ shards, int nshards, int nq, YdTop2* __restrict__ out) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  unsigned long long b = ~0ull, s = ~0ull;
  for (int c = 0; c < nshards; c++) {
    const YdTop2 t = shards[(size_t)c * nq + qi];
    top2_push(top2_key(t.best_dist, t.best_idx), b, s);
    top2_push(top2_key(t.second_dist, t.second_idx), b, s);
  }
  out[qi] = top2_unpack(b, s);
}

__global__ void k_ratio_accept(const YdTop2* __restrict__ t, int nq, float ratio, int th, uint8_t* __restrict__ accept) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  const YdTop2 v = t[qi];
  // bestDist<=lowThd && (float)bestDist < ratio*(float)secondBestDist   (src/orbMatcher.cpp:336)
  accept[qi] = (v.best_idx >= 0 && v.best_dist <= th && (float)v.best_dist < __fmul_rn(ratio, (float)v.second_dist)) ? 1 : 0;
}

// Batched variant for many small problems (frame t vs frame t-1): blockIdx.y = pair, the whole database frame
// (<= capacity rows) is streamed by every CTA; results are written unpacked, optionally with the ratio test.
template <int QPT, int THR>  // queries per thread: 4 amortises a database row best, 2 doubles the warps when the batch is small
__global__ void __launch_bounds__(THR)
k_top2_frames(const uint4* __restrict__ desc, const int32_t* __restrict__ n_of, int capacity,
              const int32_t* __restrict__ q_frame, const int32_t* __restrict__ db_frame, YdTop2* __restrict__ out,
              uint8_t* __restrict__ accept, float ratio, int th) {
  __shared__ uint4 s_db[2 * kTopTile];
  const int tid = threadIdx.x;
  const int pair = blockIdx.y;
  const int qf = q_frame[pair], df = db_frame[pair];
  const int nq = min(n_of[qf], capacity), ndb = min(n_of[df], capacity);
  const int q0 = blockIdx.x * (THR * QPT);
  if (q0 >= nq) return;
  const uint4* q = desc + (size_t)qf * capacity * 2;
  const uint4* db = desc + (size_t)df * capacity * 2;
  uint32_t qa[QPT][8];
  uint32_t best[QPT], second[QPT];
#pragma unroll
  for (int k = 0; k < QPT; k++) {
    const int qi = q0 + k * THR + tid;
    uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
    if (qi < nq) { lo = q[2 * qi]; hi = q[2 * qi + 1]; }
    qa[k][0] = lo.x; qa[k][1] = lo.y; qa[k][2] = lo.z; qa[k][3] = lo.w;
    qa[k][4] = hi.x; qa[k][5] = hi.y; qa[k][6] = hi.z; qa[k][7] = hi.w;
    best[k] = kNoKey;
    second[k] = kNoKey;
  }
  for (int t0 = 0; t0 < ndb; t0 += kTopTile) {
    const int rows = min(kTopTile, ndb - t0);
    __syncthreads();
    for (int i = tid; i < 2 * rows; i += THR) s_db[i] = db[2 * (size_t)t0 + i];
    __syncthreads();
    int r = 0;
#pragma unroll 2
    for (; r + 1 < rows; r += 2) {  // rows in pairs: 5 min/max per two keys (top2_push2)
      const uint4 lo = s_db[2 * r], hi = s_db[2 * r + 1], lo2 = s_db[2 * r + 2], hi2 = s_db[2 * r + 3];
#pragma unroll
      for (int k = 0; k < QPT; k++)
        top2_push2(ham256_key(qa[k], lo, hi, (uint32_t)(t0 + r)), ham256_key(qa[k], lo2, hi2, (uint32_t)(t0 + r + 1)), best[k],
                   second[k]);
    }
    if (r < rows) {
      const uint4 lo = s_db[2 * r], hi = s_db[2 * r + 1];
#pragma unroll
      for (int k = 0; k < QPT; k++) {
        const uint32_t key = ham256_key(qa[k], lo, hi, (uint32_t)(t0 + r));
        second[k] = min(second[k], max(best[k], key));
        best[k] = min(best[k], key);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < QPT; k++) {
    const int qi = q0 + k * THR + tid;
    if (qi >= nq) continue;
    YdTop2 o;
    o.best_idx = best[k] == kNoKey ? -1 : (int)(best[k] & 0x7FFFFFu);
    o.best_dist = best[k] == kNoKey ? 256 : (int)(best[k] >> 23);
    o.second_idx = second[k] == kNoKey ? -1 : (int)(second[k] & 0x7FFFFFu);
    o.second_dist = second[k] == kNoKey ? 256 : (int)(second[k] >> 23);
    out[(size_t)pair * capacity + qi] = o;
    if (accept)
      accept[(size_t)pair * capacity + qi] =
          (o.best_idx >= 0 && o.best_dist <= th && (float)o.best_dist < __fmul_rn(ratio, (float)o.second_dist)) ? 1 : 0;
  }
}

// Scratch is allocated when the handle is created (kInitScratch / kInitIo cover the headline batch) and only ever
// grows, geometrically: steady-state calls never reach cudaMalloc / cudaFree.
constexpr size_t kInitScratch = 8u << 20, kInitIo = 32u << 20;

static int grow(void** p, size_t* have, size_t bytes) {
  if (bytes <= *have) return YDORB_OK;
  const size_t want = std::max(bytes, 2 * *have);
  if (*p) YD_CUDA(cudaFree(*p));  // (synchronises the device: happens only while a workload is still growing)
  *p = nullptr;
  *have = 0;
  YD_CUDA(cudaMalloc(p, want));
  *have = want;
  return YDORB_OK;
}

int matcher_reserve(YdMatcher* m, size_t bytes) { return grow(&m->d_scratch, &m->scratch_bytes, bytes); }
int matcher_reserve_io(YdMatcher* m, size_t bytes) { return grow(&m->d_io, &m->io_bytes, bytes); }

// Waiting policy of the batched host entry points.  Measured on a B200, 256 frames per call, one calling thread:
// spinning in cudaStreamSynchronize 91.1 k frames/s; cudaEventSynchronize on a cudaEventBlockingSync event 73.8 k (the
// interrupt-driven wake-up is ~0.3 ms late, twice per step), also when only the kernels are slept through (74.7 k).
// But spinning keeps one core per calling thread busy for the whole 2-3 ms call, which starves 8 ranks x 4 threads on a
// 32-core host.  So: POLL the event recorded where the last kernel ends with short timed sleeps (the core is free in
// between), then SPIN through the remaining device-to-host copy, which hides the last sleep.
// YDORB_SPIN_WAIT=1 spins all the way, YDORB_SPIN_WAIT=0 (default) is the hybrid.
bool wait_sleeps() {
  static const bool spin = getenv("YDORB_SPIN_WAIT") && atoi(getenv("YDORB_SPIN_WAIT")) != 0;
  return !spin;
}
int wait_stream(cudaStream_t s, cudaEvent_t blocking_event, bool sleep) {
  if (sleep && wait_sleeps() && blocking_event) {  // recorded by the caller
    const timespec nap = {0, 30 * 1000};
    cudaError_t e;
    while ((e = cudaEventQuery(blocking_event)) == cudaErrorNotReady) nanosleep(&nap, nullptr);
    YD_CUDA(e);
  }
  YD_CUDA(cudaStreamSynchronize(s));
  return YDORB_OK;
}


int top2_grid(const YdMatcher* m, int nq, int n_rows, int* qblocks, int* nchunks, int* chunk_rows) {
  const int qb = (nq + kTopQblock - 1) / kTopQblock;
  int nc = std::max(1, (m->top_ctas_per_sm * m->sm_count + qb - 1) / qb);
  nc = std::min(nc, std::max(1, (n_rows + 255) / 256));  // at least one shared-memory tile of rows per chunk
  // rows per chunk in steps of 32, not of whole shared-memory tiles (the tile loop takes partial tiles): with 256-row
  // steps a 125 k-row shard got 768-row chunks where 563 fill the GPU evenly, and the busiest SM carried 14 % more
  // rows than the average -- 2000 x 125 k took 0.416 ms instead of the 0.33 ms an eighth of 1 M rows should take
  int cr = (n_rows + nc - 1) / nc;
  cr = std::max((cr + 31) / 32 * 32, 32);
  YD_REQUIRE(cr < (1 << 23), "database shard too large for one call (max 8M rows per chunk)");
  *qblocks = qb;
  *nchunks = std::max(1, (n_rows + cr - 1) / cr);
  *chunk_rows = cr;
  return YDORB_OK;
}

int matcher_reserve_pairs(YdMatcher* m, int nq) {
  if (nq <= m->pairs_cap) return YDORB_OK;
  const int cap = std::max(nq, 4096);
  const int qb = (cap + kTopQblock - 1) / kTopQblock;
  if (m->d_pairs) YD_CUDA(cudaFree(m->d_pairs));
  m->d_pairs = nullptr;
  m->pairs_cap = 0;
  const size_t pair_bytes = 2 * sizeof(unsigned long long) * (size_t)cap;
  YD_CUDA(cudaMalloc(&m->d_pairs, pair_bytes + sizeof(unsigned int) * qb));
  YD_CUDA(cudaMemsetAsync(m->d_pairs, 0xFF, pair_bytes, m->stream));
  YD_CUDA(cudaMemsetAsync(static_cast<uint8_t*>(m->d_pairs) + pair_bytes, 0, sizeof(unsigned int) * qb, m->stream));
  m->pairs_cap = cap;
  return YDORB_OK;
}

int launch_top2_sharded(YdMatcher* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, int n_rows, ShardDev a) {
  int qblocks, nchunks, chunk_rows;
  const int rc = top2_grid(m, nq, n_rows, &qblocks, &nchunks, &chunk_rows);
  if (rc != YDORB_OK) return rc;
  a.nq = nq;
  a.chunk_rows = chunk_rows;
  k_top2_sharded<<<dim3(qblocks, nchunks), kTopThreads, 0, m->stream>>>(
      reinterpret_cast<const uint4*>(d_queries), reinterpret_cast<const uint4*>(d_db), n_rows, a);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}
}  // namespace ydorb

using namespace ydorb;

extern "C" {

int ydorb_matcher_create(int device, YdMatcher** out) {
  YD_REQUIRE(out, "null argument");
  int ndev = 0;
  YD_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) { set_error("no such CUDA device"); return YDORB_E_CUDA; }
  YD_CUDA(cudaSetDevice(device));
  YdMatcher* m = new YdMatcher();
  m->device = device;
  m->d_scratch = nullptr;
  m->scratch_bytes = 0;
  m->d_io = nullptr;
  m->io_bytes = 0;
  m->d_pairs = nullptr;
  m->pairs_cap = 0;
  cudaError_t e = cudaStreamCreateWithFlags(&m->own_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete m; return cuda_fail(e, "cudaStreamCreate", __FILE__, __LINE__); }
  m->stream = m->own_stream;
  m->ev_dep = m->ev_block = nullptr;
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_dep, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_block, cudaEventDisableTiming | cudaEventBlockingSync);
  if (e == cudaSuccess) e = cudaMalloc(&m->d_scratch, kInitScratch);
  if (e == cudaSuccess) m->scratch_bytes = kInitScratch;
  if (e == cudaSuccess) e = cudaMalloc(&m->d_io, kInitIo);
  if (e == cudaSuccess) m->io_bytes = kInitIo;
  if (e != cudaSuccess) {
    const int rc = cuda_fail(e, "matcher allocation", __FILE__, __LINE__);
    ydorb_matcher_destroy(m);
    return rc;
  }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  m->sm_count = prop.multiProcessorCount;
  m->top_ctas_per_sm = kTopCtasPerSm;
  if (const char* e = getenv("YDORB_TOP_CTAS_PER_SM")) m->top_ctas_per_sm = std::max(1, atoi(e));  // tuning knob
  *out = m;
  return YDORB_OK;
}

void ydorb_matcher_destroy(YdMatcher* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  if (m->ev_dep) cudaEventDestroy(m->ev_dep);
  if (m->ev_block) cudaEventDestroy(m->ev_block);
  if (m->d_scratch) cudaFree(m->d_scratch);
  if (m->d_io) cudaFree(m->d_io);
  if (m->d_pairs) cudaFree(m->d_pairs);
  if (m->own_stream) cudaStreamDestroy(m->own_stream);
  delete m;
}

void* ydorb_matcher_stream(const YdMatcher* m) { return m ? (void*)m->stream : nullptr; }

int ydorb_matcher_sync(YdMatcher* m) {
  YD_REQUIRE(m, "null matcher");
  YD_CUDA(cudaSetDevice(m->device));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

int ydorb_matcher_set_stream(YdMatcher* m, void* cuda_stream) {
  YD_REQUIRE(m, "null matcher");
  YD_CUDA(cudaSetDevice(m->device));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  m->stream = cuda_stream ? (cudaStream_t)cuda_stream : m->own_stream;
  return YDORB_OK;
}

int ydorb_hamming_top2_frames_device(YdMatcher* m, const uint8_t* d_desc, const int32_t* d_n, int capacity,
                                     const int32_t* d_q_frame, const int32_t* d_db_frame, int npairs, YdTop2* d_out,
                                     uint8_t* d_accept, float ratio, int th) {
  YD_REQUIRE(m && d_desc && d_n && d_q_frame && d_db_frame && d_out, "null argument");
  YD_REQUIRE(capacity > 0 && capacity < (1 << 23) && npairs >= 0, "bad size");
  YD_REQUIRE(((uintptr_t)d_desc & 15) == 0, "descriptor block must be 16-byte aligned");
  if (npairs == 0) return YDORB_OK;
  YD_CUDA(cudaSetDevice(m->device));
  // Queries per thread: 4 amortise a database row over the most distance computations, but a small batch then has too
  // few warps to hide the kernel's latencies (256 pairs of 1000 x 1000 = 13.5 warps per SM): halve it until the launch
  // has ~20 warps per SM (measured: 0.410 -> 0.358 ms per 256 pairs, 731 Gpairs/s).  YDORB_MATCH_QPT overrides.
  const long long warps4 = (long long)npairs * ((capacity + kTopQblock - 1) / kTopQblock) * (kTopThreads / 32);
  int qpt = warps4 >= 20ll * m->sm_count ? 4 : (2 * warps4 >= 20ll * m->sm_count ? 2 : 1);
  if (const char* e = getenv("YDORB_MATCH_QPT")) qpt = atoi(e);
#define YD_LAUNCH_FRAMES(QPT, THR)                                                                                   \
  k_top2_frames<QPT, THR><<<dim3((capacity + THR * QPT - 1) / (THR * QPT), npairs), THR, 0, m->stream>>>(           \
      reinterpret_cast<const uint4*>(d_desc), d_n, capacity, d_q_frame, d_db_frame, d_out, d_accept, ratio, th)
  switch (qpt) {
    case 1: YD_LAUNCH_FRAMES(1, kTopThreads); break;
    case 2: YD_LAUNCH_FRAMES(2, kTopThreads); break;
    default: YD_LAUNCH_FRAMES(4, kTopThreads); break;
  }
#undef YD_LAUNCH_FRAMES
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

int ydorb_match_frames(YdMatcher* m, const YdExtractor* ex, const int32_t* q_frame, const int32_t* db_frame, int npairs,
                       float ratio, int th, YdTop2* out, uint8_t* accept) {
  YD_REQUIRE(m && ex && q_frame && db_frame && out && npairs >= 0, "bad argument");
  if (ex->cur_w <= 0) { set_error("no extract call yet"); return YDORB_E_STATE; }
  YD_REQUIRE(ex->device == m->device, "handles live on different devices");
  for (int p = 0; p < npairs; p++)
    YD_REQUIRE(q_frame[p] >= 0 && q_frame[p] < ex->cur_batch && db_frame[p] >= 0 && db_frame[p] < ex->cur_batch,
               "frame index out of range");
  if (npairs == 0) return YDORB_OK;
  YD_CUDA(cudaSetDevice(m->device));
  const int cap = ex->out_cap;
  const size_t nres = (size_t)npairs * cap;
  const size_t o_q = 0, o_db = (sizeof(int32_t) * npairs + 255) / 256 * 256, o_out = 2 * o_db,
               o_acc = o_out + (sizeof(YdTop2) * nres + 255) / 256 * 256, total = o_acc + nres;
  int rc = matcher_reserve_io(m, total);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  // descriptors were produced on the extractor's stream: order the two streams on the device, no host wait
  YD_CUDA(cudaEventRecord(m->ev_dep, ex->stream));
  YD_CUDA(cudaStreamWaitEvent(m->stream, m->ev_dep, 0));
  YD_CUDA(cudaMemcpyAsync(io + o_q, q_frame, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, m->stream));
  YD_CUDA(cudaMemcpyAsync(io + o_db, db_frame, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, m->stream));
  YD_CUDA(cudaMemsetAsync(io + o_out, 0xFF, sizeof(YdTop2) * nres, m->stream));  // rows >= n_out[q_frame]: all fields -1
  if (accept) YD_CUDA(cudaMemsetAsync(io + o_acc, 0, nres, m->stream));           // ... and never accepted
  rc = ydorb_hamming_top2_frames_device(m, ex->d_desc, ex->d_nout, cap, reinterpret_cast<int32_t*>(io + o_q),
                                        reinterpret_cast<int32_t*>(io + o_db), npairs, reinterpret_cast<YdTop2*>(io + o_out),
                                        accept ? io + o_acc : nullptr, ratio, th);
  if (rc != YDORB_OK) return rc;
  const bool sleep = npairs >= 8 && wait_sleeps();
  if (sleep) YD_CUDA(cudaEventRecord(m->ev_block, m->stream));  // where the kernel ends: sleep until here, spin through the copies
  YD_CUDA(cudaMemcpyAsync(out, io + o_out, sizeof(YdTop2) * nres, cudaMemcpyDeviceToHost, m->stream));
  if (accept) YD_CUDA(cudaMemcpyAsync(accept, io + o_acc, nres, cudaMemcpyDeviceToHost, m->stream));
  return wait_stream(m->stream, m->ev_block, sleep);
}

int ydorb_hamming_top2_device(YdMatcher* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, int ndb, int idx_base,
                              YdTop2* d_out) {
  YD_REQUIRE(m && d_out, "null argument");
  YD_REQUIRE(nq >= 0 && ndb >= 0, "negative size");
  if (nq == 0) return YDORB_OK;
  YD_REQUIRE(d_queries && (d_db || ndb == 0), "null descriptor pointer");
  YD_REQUIRE(((uintptr_t)d_queries & 15) == 0 && ((uintptr_t)d_db & 15) == 0, "descriptor matrices must be 16-byte aligned");
  YD_CUDA(cudaSetDevice(m->device));
  // one launch: chunked scan, order-free fold into one (best, second) pair per query, finish by the last CTA of each
  // query block (k_top2_sharded in single-GPU mode)
  int rc = matcher_reserve_pairs(m, nq);
  if (rc != YDORB_OK) return rc;
  ShardDev a;
  memset(&a, 0, sizeof(a));
  a.world = 1;
  a.mode = 0;
  a.idx_base = idx_base;
  a.best = static_cast<unsigned long long*>(m->d_pairs);
  a.second = a.best + m->pairs_cap;
  a.tickets = reinterpret_cast<unsigned int*>(a.second + m->pairs_cap);
  a.out = d_out;
  return launch_top2_sharded(m, d_queries, nq, d_db, ndb, a);
}

int ydorb_hamming_top2(YdMatcher* m, const uint8_t* queries, int nq, const uint8_t* db, int ndb, int idx_base,
                       YdTop2* out) {
  YD_REQUIRE(m && out, "null argument");
  YD_REQUIRE(nq >= 0 && ndb >= 0, "negative size");
  if (nq == 0) return YDORB_OK;
  YD_REQUIRE(queries && (db || ndb == 0), "null descriptor pointer");
  YD_CUDA(cudaSetDevice(m->device));
  const size_t qb = (size_t)nq * 32, dbb = (size_t)std::max(ndb, 1) * 32, ob = sizeof(YdTop2) * (size_t)nq;
  const size_t o_q = 0, o_db = (qb + 255) / 256 * 256, o_out = o_db + (dbb + 255) / 256 * 256;
  int rc = matcher_reserve_io(m, o_out + ob);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YD_CUDA(cudaMemcpyAsync(io + o_q, queries, qb, cudaMemcpyHostToDevice, m->stream));
  if (ndb > 0) YD_CUDA(cudaMemcpyAsync(io + o_db, db, (size_t)ndb * 32, cudaMemcpyHostToDevice, m->stream));
  rc = ydorb_hamming_top2_device(m, io + o_q, nq, io + o_db, ndb, idx_base, reinterpret_cast<YdTop2*>(io + o_out));
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(out, io + o_out, ob, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

int ydorb_hamming_pairs(YdMatcher* m, const uint8_t* a, const uint8_t* b, int n, int32_t* dist) {
  YD_REQUIRE(m && dist, "null argument");
  YD_REQUIRE(n >= 0, "negative size");
  if (n == 0) return YDORB_OK;
  YD_REQUIRE(a && b, "null descriptor pointer");
  YD_CUDA(cudaSetDevice(m->device));
  const size_t nb = ((size_t)n * 32 + 255) / 256 * 256;
  int rc = matcher_reserve_io(m, 2 * nb + sizeof(int32_t) * (size_t)n);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YD_CUDA(cudaMemcpyAsync(io, a, (size_t)n * 32, cudaMemcpyHostToDevice, m->stream));
  YD_CUDA(cudaMemcpyAsync(io + nb, b, (size_t)n * 32, cudaMemcpyHostToDevice, m->stream));
  k_hamming_pairs<<<(n + 127) / 128, 128, 0, m->stream>>>(reinterpret_cast<const uint4*>(io), reinterpret_cast<const uint4*>(io + nb), n,
                                                          reinterpret_cast<int32_t*>(io + 2 * nb));
  YD_CUDA(cudaGetLastError());
  YD_CUDA(cudaMemcpyAsync(dist, io + 2 * nb, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

int ydorb_top2_merge_device(YdMatcher* m, const YdTop2* d_shards, int nshards, int nq, YdTop2* d_out) {
  YD_REQUIRE(m && d_shards && d_out && nshards >= 1 && nq >= 0, "bad argument");
  if (nq == 0) return YDORB_OK;
  YD_CUDA(cudaSetDevice(m->device));
  k_top2_merge<<<(nq + 127) / 128, 128, 0, m->stream>>>(d_shards, nshards, nq, d_out);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

int ydorb_ratio_accept_device(YdMatcher* m, const YdTop2* d_top2, int nq, float ratio, int th, uint8_t* d_accept) {
  YD_REQUIRE(m && d_top2 && d_accept && nq >= 0, "bad argument");
  if (nq == 0) return YDORB_OK;
  YD_CUDA(cudaSetDevice(m->device));
  k_ratio_accept<<<(nq + 127) / 128, 128, 0, m->stream>>>(d_top2, nq, ratio, th, d_accept);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // extern "C"
