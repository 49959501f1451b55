// tools/umma_probe.cu -- standalone check of the tcgen05 (kind::i8) operand layout used by csrc/hamming_umma.cu:
// 128 x 256 x 256 AND-popcount product of 256-bit rows expanded to {0,1} bytes in shared memory (K-major, no swizzle:
// 8 x 16 B core matrices, SBO = 128 B between row groups, LBO = rows x 16 B between 16-byte K chunks), accumulators in
// TMEM, read back with tcgen05.ld.32x32b.  Prints PASS when D[i][j] == popc(a_i & b_j) for every i, j.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/_bin/umma_probe tools/umma_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int kM = 128, kN = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFFu);
  d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  return d;                // layout type 0 (no swizzle), base offset 0
}

// row `r` of a `rows`-row operand: 32 descriptor bytes -> 256 bytes of {0,1}, 16 chunks of 16 B
__device__ __forceinline__ void expand_row(uint8_t* tile, int rows, int r, const uint8_t* desc) {
  const uint4* d4 = reinterpret_cast<const uint4*>(desc);
  uint32_t w[8];
  *reinterpret_cast<uint4*>(w) = d4[0];
  *reinterpret_cast<uint4*>(w + 4) = d4[1];
#pragma unroll
  for (int c = 0; c < 16; c++) {  // chunk c = descriptor bytes 2c, 2c + 1
    const uint32_t h = (w[c >> 1] >> (16 * (c & 1))) & 0xFFFFu;
    uint4 o;
    o.x = ((h & 15u) * 0x00204081u) & 0x01010101u;
    o.y = (((h >> 4) & 15u) * 0x00204081u) & 0x01010101u;
    o.z = (((h >> 8) & 15u) * 0x00204081u) & 0x01010101u;
    o.w = (((h >> 12) & 15u) * 0x00204081u) & 0x01010101u;
    *reinterpret_cast<uint4*>(tile + (size_t)c * rows * 16 + (r >> 3) * 128 + (r & 7) * 16) = o;
  }
}

__global__ void __launch_bounds__(128) k_probe(const uint8_t* a, const uint8_t* b, int32_t* d) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sa = smem;                  // 128 x 256 B
  uint8_t* sb = smem + kM * 256;       // 256 x 256 B
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t s_bar;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  expand_row(sa, kM, tid, a + tid * 32);
  expand_row(sb, kN, tid, b + tid * 32);
  expand_row(sb, kN, tid + 128, b + (tid + 128) * 32);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = s_tmem;
  if (tid == 0) {
    const uint32_t idesc = (2u << 4) | ((uint32_t)(kN >> 3) << 17) | ((uint32_t)(kM >> 4) << 24);  // u8 x u8 -> s32, K-major both
    for (int k = 0; k < 8; k++) {
      const uint64_t da = make_desc(smem_u32(sa) + 2 * k * kM * 16, kM * 16, 128);
      const uint64_t db = make_desc(smem_u32(sb) + 2 * k * kN * 16, kN * 16, 128);
      const uint32_t acc = k > 0;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
          ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_bar)) : "memory");
  }
  {  // wait for the MMAs
    uint32_t done = 0;
    while (!done) {
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done) : "r"(smem_u32(&s_bar)), "r"(0));
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  for (int c0 = 0; c0 < kN; c0 += 32) {
    uint32_t v[32];
    const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, "
        "%18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 32; j++) d[tid * kN + c0 + j] = (int32_t)v[j];
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256));
}

int main() {
  std::vector<uint8_t> a(kM * 32), b(kN * 32);
  srand(7);
  for (auto& x : a) x = (uint8_t)(rand() >> 7);
  for (auto& x : b) x = (uint8_t)(rand() >> 7);
  uint8_t *da, *db;
  int32_t* dd;
  cudaMalloc(&da, a.size());
  cudaMalloc(&db, b.size());
  cudaMalloc(&dd, sizeof(int32_t) * kM * kN);
  cudaMemcpy(da, a.data(), a.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(db, b.data(), b.size(), cudaMemcpyHostToDevice);
  cudaMemset(dd, 0xFF, sizeof(int32_t) * kM * kN);
  const int smem = (kM + kN) * 256;
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k_probe<<<1, 128, smem>>>(da, db, dd);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("FAIL cuda: %s\n", cudaGetErrorString(e));
    return 1;
  }
  std::vector<int32_t> d(kM * kN);
  cudaMemcpy(d.data(), dd, sizeof(int32_t) * kM * kN, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int i = 0; i < kM; i++)
    for (int j = 0; j < kN; j++) {
      int want = 0;
      for (int k = 0; k < 32; k++) want += __builtin_popcount(a[i * 32 + k] & b[j * 32 + k]);
      if (d[i * kN + j] != want && bad++ < 8) printf("D[%d][%d] = %d, want %d\n", i, j, d[i * kN + j], want);
    }
  printf(bad ? "FAIL %d mismatches\n" : "PASS\n", bad);
  return bad != 0;
}
