// Device helpers shared by the Hamming kernels (hamming.cu) and the sharded multi-GPU scan (shard.cu).
#pragma once
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

__device__ __forceinline__ unsigned long long top2_key(int dist, int idx) {
  return idx < 0 ? ~0ull : (((unsigned long long)(uint32_t)dist << 32) | (uint32_t)idx);
}
__device__ __forceinline__ void top2_push(unsigned long long key, unsigned long long& b, unsigned long long& s) {
  s = min(s, max(b, key));
  b = min(b, key);
}
// two new keys into a running (best, second) pair of 32-bit keys (all keys distinct: the row index is in the low bits):
// 5 min/max instead of 2 x 3.  best' = min(best, m); second' = second smallest of {best, second, m, M} with m <= M,
// best <= second, which is min(second, max(best, m), M).
__device__ __forceinline__ void top2_push2(uint32_t ka, uint32_t kb, uint32_t& best, uint32_t& second) {
  const uint32_t m = min(ka, kb), M = max(ka, kb);
  second = __vimin3_u32(second, max(best, m), M);
  best = min(best, m);
}
__device__ __forceinline__ YdTop2 top2_unpack(unsigned long long b, unsigned long long s) {
  YdTop2 o;
  o.best_idx = (b == ~0ull) ? -1 : (int)(uint32_t)b;
  o.best_dist = (b == ~0ull) ? 256 : (int)(b >> 32);
  o.second_idx = (s == ~0ull) ? -1 : (int)(uint32_t)s;
  o.second_dist = (s == ~0ull) ? 256 : (int)(s >> 32);
  return o;
}

__device__ __forceinline__ uint8_t ratio_accept(const YdTop2& v, float ratio, int th) {
  // bestDist<=lowThd && (float)bestDist < ratio*(float)secondBestDist   (src/orbMatcher.cpp:336)
  return (v.best_idx >= 0 && v.best_dist <= th && (float)v.best_dist < __fmul_rn(ratio, (float)v.second_dist)) ? 1 : 0;
}

constexpr int kShardMaxWorld = 64;
constexpr int kShardStatusTimeout = 1;

struct ShardDev {          // everything the kernel needs besides the scan arguments
  int rank, world;
  int mode;                // 0 single GPU, 1 NCCL exchange (kernel stops after the finish), 2 peer push + wait + merge
  int nq, idx_base, chunk_rows;
  unsigned long long* best;    // [nq] per-query running minimum (distance << 32 | global row), ~0 between calls
  unsigned long long* second;  // [nq] per-query runner-up
  unsigned int* tickets;   // [qblocks], zero between calls
  YdTop2* gather;          // local [world][nq] of this call's parity
  YdTop2* peer_gather[kShardMaxWorld];      // rank r's gather buffer of this parity, mapped here (mode 2)
  uint32_t* peer_arrivals[kShardMaxWorld];  // rank r's arrival counters of this parity: [world] u32 (mode 2)
  uint32_t* arrivals;      // local counters of this parity
  uint32_t target;         // value each source's counter has reached once all query blocks of this call arrived
  unsigned long long timeout_ns;
  int* status;
  YdTop2* out;             // merged result (modes 0, 2)
  uint8_t* accept;         // may be null
  float ratio;
  int th;
};


// grid of the chunked scan for nq queries against n_rows rows: returns YDORB_OK and fills qblocks / nchunks / chunk_rows
int top2_grid(const YdMatcher* m, int nq, int n_rows, int* qblocks, int* nchunks, int* chunk_rows);
// makes sure the matcher's per-query (best, second) pairs and ticket counters cover nq queries (all-ones / zero)
int matcher_reserve_pairs(YdMatcher* m, int nq);
// launches k_top2_sharded on the matcher's stream; `a` complete except for chunk_rows
int launch_top2_sharded(YdMatcher* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, int n_rows, ShardDev a);

}  // namespace ydorb
