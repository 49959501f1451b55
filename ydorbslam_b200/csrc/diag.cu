// Diagnostics: measured integer-pipe issue rates of the device, the denominators of the integer roofline
// ("INT32/popc peak must be measured on the box by a micro-kernel").  Not on the product path.
//
// Every kernel runs kChains independent dependent-chains per thread of ONE instruction class (inline PTX, volatile,
// so nothing is folded), 256 threads x 8 CTAs per SM, and reports thread-level operations per second.
#include "common.cuh"

namespace ydorb {

constexpr int kChains = 8;
constexpr int kUnroll = 8;

enum DiagOp { kOpPopc = 0, kOpLop3 = 1, kOpVimnmx3 = 2, kOpImad = 3, kOpMix = 4, kOpCount = 5,
              kOpHmnmx2 = 5, kOpMixHmnmxVimnmx = 6, kOpCountAll = 7 };  // 5, 6: only when the caller has room for 7

template <int OP>
__global__ void __launch_bounds__(256) k_diag_rate(int iters, uint32_t seed, uint32_t* __restrict__ sink) {
  uint32_t a[kChains];
#pragma unroll
  for (int i = 0; i < kChains; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u + blockIdx.x;
  const uint32_t b = seed ^ 0x5bd1e995u, c = seed + 0x01000193u;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < kUnroll; u++) {
#pragma unroll
      for (int i = 0; i < kChains; i++) {
        if (OP == kOpPopc) {
          asm volatile("popc.b32 %0, %0;" : "+r"(a[i]));
        } else if (OP == kOpLop3) {
          asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
        } else if (OP == kOpVimnmx3) {
          a[i] = (u & 1) ? __vimax3_s16x2(a[i], b, a[(i + 1) % kChains]) : __vimin3_s16x2(a[i], c, a[(i + 1) % kChains]);
        } else if (OP == kOpImad) {
          asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
        } else if (OP == kOpHmnmx2) {  // packed half min / max: which pipe, what rate?
          if (u & 1) asm volatile("max.f16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
          else asm volatile("min.f16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
        } else if (OP == kOpMixHmnmxVimnmx) {  // does packed half min / max co-issue with the integer one?
          if (i & 1) asm volatile("max.f16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
          else a[i] = __vimax3_s16x2(a[i], b, c);
        } else {  // alternate the alu and fma pipes: is the limit the pipe or the issue slot?
          if (i & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
          else asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
        }
      }
    }
  }
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < kChains; i++) r ^= a[i];
  if (r == 0x12345678u) sink[0] = r;  // never true in practice; keeps the chains alive
}

template <int OP>
static int run_rate(int sm_count, int iters, uint32_t* d_sink, double* ops_per_s) {
  const int ctas = sm_count * 8;
  cudaEvent_t e0, e1;
  YD_CUDA(cudaEventCreate(&e0));
  YD_CUDA(cudaEventCreate(&e1));
  k_diag_rate<OP><<<ctas, 256>>>(iters / 8 + 1, 12345u, d_sink);  // warm-up
  float best = 1e30f;
  for (int rep = 0; rep < 5; rep++) {
    YD_CUDA(cudaEventRecord(e0));
    k_diag_rate<OP><<<ctas, 256>>>(iters, 12345u + rep, d_sink);
    YD_CUDA(cudaEventRecord(e1));
    YD_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    YD_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    best = ms < best ? ms : best;
  }
  count_launches(6);
  YD_CUDA(cudaGetLastError());
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *ops_per_s = (double)ctas * 256.0 * (double)iters * kUnroll * kChains / (best * 1e-3);
  return YDORB_OK;
}

// glibc_sincosf over consecutive float bit patterns: out[2*i] = sin, out[2*i+1] = cos of the float with bits first+i
__global__ void __launch_bounds__(256) k_diag_sincosf(uint32_t first_bits, uint32_t count, float* __restrict__ out) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    float s, c;
    glibc_sincosf(__uint_as_float(first_bits + i), s, c);
    reinterpret_cast<float2*>(out)[i] = make_float2(s, c);
  }
}

}  // namespace ydorb

using namespace ydorb;

extern "C" int ydorb_diag_sincosf(int device, uint32_t first_bits, uint32_t count, float* out_sin_cos) {
  YD_REQUIRE(out_sin_cos && count > 0, "need an output buffer of 2*count floats");
  int ndev = 0;
  YD_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) { set_error("no such CUDA device"); return YDORB_E_CUDA; }
  YD_CUDA(cudaSetDevice(device));
  float* d = nullptr;
  YD_CUDA(cudaMalloc(&d, (size_t)count * 8));
  k_diag_sincosf<<<148 * 8, 256>>>(first_bits, count, d);
  count_launches(1);
  cudaError_t e = cudaMemcpy(out_sin_cos, d, (size_t)count * 8, cudaMemcpyDeviceToHost);
  cudaFree(d);
  YD_CUDA(e);
  return YDORB_OK;
}

extern "C" int ydorb_diag_int_peaks(int device, double* ops_per_s, int n) {
  YD_REQUIRE(ops_per_s && n >= kOpCount, "need room for 5 rates: popc, lop3, vimnmx3.s16x2, imad, lop3+imad mix");
  int ndev = 0;
  YD_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) { set_error("no such CUDA device"); return YDORB_E_CUDA; }
  YD_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  YD_CUDA(cudaGetDeviceProperties(&prop, device));
  uint32_t* d_sink = nullptr;
  YD_CUDA(cudaMalloc(&d_sink, 64));
  const int iters = 4096;
  int rc = run_rate<kOpPopc>(prop.multiProcessorCount, iters, d_sink, &ops_per_s[kOpPopc]);
  if (rc == YDORB_OK) rc = run_rate<kOpLop3>(prop.multiProcessorCount, iters, d_sink, &ops_per_s[kOpLop3]);
  if (rc == YDORB_OK) rc = run_rate<kOpVimnmx3>(prop.multiProcessorCount, iters, d_sink, &ops_per_s[kOpVimnmx3]);
  if (rc == YDORB_OK) rc = run_rate<kOpImad>(prop.multiProcessorCount, iters, d_sink, &ops_per_s[kOpImad]);
  if (rc == YDORB_OK) rc = run_rate<kOpMix>(prop.multiProcessorCount, iters, d_sink, &ops_per_s[kOpMix]);
  if (n >= kOpCountAll) {
    if (rc == YDORB_OK) rc = run_rate<kOpHmnmx2>(prop.multiProcessorCount, iters, d_sink, &ops_per_s[kOpHmnmx2]);
    if (rc == YDORB_OK) rc = run_rate<kOpMixHmnmxVimnmx>(prop.multiProcessorCount, iters, d_sink, &ops_per_s[kOpMixHmnmxVimnmx]);
  }
  cudaFree(d_sink);
  return rc;
}
