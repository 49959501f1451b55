// M0  OrbMatcher::computeDescriptorsDistance (src/orbMatcher.cpp:11-23) and the dense top-2 scan that every
// BoW / brute-force matcher runs around it (src/orbMatcher.cpp:319-334: `if d<best {second=best; best=d}
// else if d<second {second=d}`).
//
// Descriptors are 32 B = 2 x uint4.  A thread keeps 4 query descriptors in registers (32 words), the CTA streams
// its slice of the database through shared memory with coalesced uint4 loads and every thread reads each database
// row as two broadcast LDS.128.  Distance = 8 x (XOR, POPC).  The running top-2 is two packed words
// (distance << 23 | row-in-chunk): best = min, second = min(second, max(best_before, key)) -- with the row in the
// low bits this reproduces the sequential rule exactly, including "first candidate wins ties".
// The database is split into chunks across blockIdx.y so the grid fills 148 SMs; per-chunk top-2 pairs are merged
// by a second kernel with the same order-free rule, which is also the cross-GPU shard merge.
#include <algorithm>
#include <cstdlib>

#include "matcher.cuh"

namespace ydorb {

constexpr int kTopThreads = 128;
constexpr int kTopQpt = 4;                          // queries per thread
constexpr int kTopQblock = kTopThreads * kTopQpt;   // queries per CTA
constexpr int kTopTile = 256;                       // database rows per shared-memory tile
constexpr uint32_t kNoKey = 0xFFFFFFFFu;
#ifndef YD_TOP_CTAS_PER_SM
#define YD_TOP_CTAS_PER_SM 6
#endif
constexpr int kTopCtasPerSm = YD_TOP_CTAS_PER_SM;   // 6 x 4 warps per SM hide the POPC (XU) latency

// 256-bit Hamming distance with a carry-save adder tree: the 8 XOR words are compressed by four 3:2 compressors
// (sum = a^b^c, carry = maj(a,b,c), one LOP3 each) into words of weight 1,1,2,4, so a pair costs 4 POPC (XU pipe,
// 16 thread-ops/clk/SM) + 16 LOP3 (ALU pipe, 64/clk/SM) instead of 8 POPC + 8 LOP3.  Same integer result as
// src/orbMatcher.cpp:11-23.
__device__ __forceinline__ uint32_t csa_sum(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;  // inline PTX keeps the compiler from re-associating the XOR tree into more, narrower LOP3s
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint32_t csa_carry(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
// returns distance << 23, ready to be added to a row index below 2^23
__device__ __forceinline__ uint32_t ham256_key(const uint32_t (&q)[8], const uint4& lo, const uint4& hi, uint32_t row) {
  const uint32_t x0 = q[0] ^ lo.x, x1 = q[1] ^ lo.y, x2 = q[2] ^ lo.z, x3 = q[3] ^ lo.w;
  const uint32_t x4 = q[4] ^ hi.x, x5 = q[5] ^ hi.y, x6 = q[6] ^ hi.z, x7 = q[7] ^ hi.w;
  const uint32_t sA = csa_sum(x0, x1, x2), cA = csa_carry(x0, x1, x2);
  const uint32_t sB = csa_sum(x3, x4, x5), cB = csa_carry(x3, x4, x5);
  const uint32_t sC = csa_sum(sA, sB, x6), cC = csa_carry(sA, sB, x6);
  const uint32_t sD = csa_sum(cA, cB, cC), cD = csa_carry(cA, cB, cC);
  // weights 1,1,2,4 folded with the << 23 into multiply-adds (FMA pipe, off the ALU pipe the LOP3s use)
  uint32_t d = __popc(sC) + __popc(x7);
  d = __popc(sD) * 2u + d;
  d = __popc(cD) * 4u + d;
  return d * 0x800000u + row;
}
__device__ __forceinline__ uint32_t ham256(const uint32_t (&q)[8], const uint4& lo, const uint4& hi) {
  return ham256_key(q, lo, hi, 0) >> 23;
}

__global__ void k_hamming_pairs(const uint4* __restrict__ a, const uint4* __restrict__ b, int n, int32_t* __restrict__ dist) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 a0 = a[2 * i], a1 = a[2 * i + 1], b0 = b[2 * i], b1 = b[2 * i + 1];
  dist[i] = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
            __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(kTopThreads)
k_top2_chunks(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db, int ndb, int chunk_rows,
              uint2* __restrict__ partial) {
  __shared__ uint4 s_db[2 * kTopTile];
  const int tid = threadIdx.x;
  const int q0 = blockIdx.x * kTopQblock;
  const int chunk = blockIdx.y;
  const int row_begin = chunk * chunk_rows;
  const int row_end = min(row_begin + chunk_rows, ndb);
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
#pragma unroll 2
    for (int r = 0; r < rows; r++) {
      const uint4 lo = s_db[2 * r], hi = s_db[2 * r + 1];
#pragma unroll
      for (int k = 0; k < kTopQpt; k++) {
        const uint32_t key = ham256_key(qa[k], lo, hi, local0 + (uint32_t)r);
        second[k] = min(second[k], max(best[k], key));
        best[k] = min(best[k], key);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < kTopQpt; k++) {
    const int qi = q0 + k * kTopThreads + tid;
    if (qi < nq) partial[(size_t)chunk * nq + qi] = make_uint2(best[k], second[k]);
  }
}

__device__ __forceinline__ unsigned long long top2_key(int dist, int idx) {
  return idx < 0 ? ~0ull : (((unsigned long long)(uint32_t)dist << 32) | (uint32_t)idx);
}
__device__ __forceinline__ void top2_push(unsigned long long key, unsigned long long& b, unsigned long long& s) {
  s = min(s, max(b, key));
  b = min(b, key);
}
__device__ __forceinline__ YdTop2 top2_unpack(unsigned long long b, unsigned long long s) {
  YdTop2 o;
  o.best_idx = (b == ~0ull) ? -1 : (int)(uint32_t)b;
  o.best_dist = (b == ~0ull) ? 256 : (int)(b >> 32);
  o.second_idx = (s == ~0ull) ? -1 : (int)(uint32_t)s;
  o.second_dist = (s == ~0ull) ? 256 : (int)(s >> 32);
  return o;
}

__global__ void k_top2_finish(const uint2* __restrict__ partial, int nchunks, int nq, int chunk_rows, int idx_base,
                              YdTop2* __restrict__ out) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  unsigned long long b = ~0ull, s = ~0ull;
  for (int c = 0; c < nchunks; c++) {
    const uint2 p = partial[(size_t)c * nq + qi];
    const int base = idx_base + c * chunk_rows;
    if (p.x != kNoKey) top2_push(top2_key((int)(p.x >> 23), base + (int)(p.x & 0x7FFFFFu)), b, s);
    if (p.y != kNoKey) top2_push(top2_key((int)(p.y >> 23), base + (int)(p.y & 0x7FFFFFu)), b, s);
  }
  out[qi] = top2_unpack(b, s);
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
#pragma unroll 2
    for (int r = 0; r < rows; r++) {
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

int matcher_reserve(YdMatcher* m, size_t bytes) {
  if (bytes <= m->scratch_bytes) return YDORB_OK;
  if (m->d_scratch) YD_CUDA(cudaFree(m->d_scratch));
  m->d_scratch = nullptr;
  m->scratch_bytes = 0;
  YD_CUDA(cudaMalloc(&m->d_scratch, bytes));
  m->scratch_bytes = bytes;
  return YDORB_OK;
}

int matcher_reserve_io(YdMatcher* m, size_t bytes) {
  if (bytes <= m->io_bytes) return YDORB_OK;
  if (m->d_io) YD_CUDA(cudaFree(m->d_io));
  m->d_io = nullptr;
  m->io_bytes = 0;
  YD_CUDA(cudaMalloc(&m->d_io, bytes));
  m->io_bytes = bytes;
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
  cudaError_t e = cudaStreamCreateWithFlags(&m->own_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete m; return cuda_fail(e, "cudaStreamCreate", __FILE__, __LINE__); }
  m->stream = m->own_stream;
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
  if (m->d_scratch) cudaFree(m->d_scratch);
  if (m->d_io) cudaFree(m->d_io);
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
  YD_CUDA(cudaStreamSynchronize(ex->stream));  // descriptors were produced on the extractor's stream
  YD_CUDA(cudaMemcpyAsync(io + o_q, q_frame, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, m->stream));
  YD_CUDA(cudaMemcpyAsync(io + o_db, db_frame, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, m->stream));
  YD_CUDA(cudaMemsetAsync(io + o_out, 0xFF, sizeof(YdTop2) * nres, m->stream));  // unused rows read as idx -1
  rc = ydorb_hamming_top2_frames_device(m, ex->d_desc, ex->d_nout, cap, reinterpret_cast<int32_t*>(io + o_q),
                                        reinterpret_cast<int32_t*>(io + o_db), npairs, reinterpret_cast<YdTop2*>(io + o_out),
                                        accept ? io + o_acc : nullptr, ratio, th);
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(out, io + o_out, sizeof(YdTop2) * nres, cudaMemcpyDeviceToHost, m->stream));
  if (accept) YD_CUDA(cudaMemcpyAsync(accept, io + o_acc, nres, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

int ydorb_hamming_top2_device(YdMatcher* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, int ndb, int idx_base,
                              YdTop2* d_out) {
  YD_REQUIRE(m && d_out, "null argument");
  YD_REQUIRE(nq >= 0 && ndb >= 0, "negative size");
  if (nq == 0) return YDORB_OK;
  YD_REQUIRE(d_queries && (d_db || ndb == 0), "null descriptor pointer");
  YD_REQUIRE(((uintptr_t)d_queries & 15) == 0 && ((uintptr_t)d_db & 15) == 0, "descriptor matrices must be 16-byte aligned");
  YD_CUDA(cudaSetDevice(m->device));
  const int qblocks = (nq + kTopQblock - 1) / kTopQblock;
  int nchunks = std::max(1, (m->top_ctas_per_sm * m->sm_count + qblocks - 1) / qblocks);
  nchunks = std::min(nchunks, std::max(1, (ndb + 1023) / 1024));
  int chunk_rows = (ndb + nchunks - 1) / nchunks;
  chunk_rows = std::max((chunk_rows + kTopTile - 1) / kTopTile * kTopTile, kTopTile);
  YD_REQUIRE(chunk_rows < (1 << 23), "database shard too large for one call (max 8M rows per chunk)");
  nchunks = std::max(1, (ndb + chunk_rows - 1) / chunk_rows);
  int rc = matcher_reserve(m, sizeof(uint2) * (size_t)nchunks * nq);
  if (rc != YDORB_OK) return rc;
  uint2* partial = reinterpret_cast<uint2*>(m->d_scratch);
  dim3 grid(qblocks, nchunks);
  k_top2_chunks<<<grid, kTopThreads, 0, m->stream>>>(reinterpret_cast<const uint4*>(d_queries), nq,
                                                     reinterpret_cast<const uint4*>(d_db), ndb, chunk_rows, partial);
  k_top2_finish<<<(nq + 127) / 128, 128, 0, m->stream>>>(partial, nchunks, nq, chunk_rows, idx_base, d_out);
  count_launches(2);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
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
