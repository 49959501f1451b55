// Probe: 3-D u8 TMA box load variants (dynamic index into a __grid_constant__ array of maps, PDL attribute on/off).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ydorbslam_b200/csrc -o /tmp/tma_probe tools/tma_probe.cu ydorbslam_b200/csrc/tma.cu -DTMA_PROBE
#include <cstdio>
#include <vector>
#include "tma.cuh"
namespace ydorb {
void set_error(const std::string& m) { printf("error: %s\n", m.c_str()); }
}
using namespace ydorb;
struct Maps { CUtensorMap lv[8]; };
template <int MODE>
__global__ void __launch_bounds__(256) k(const __grid_constant__ Maps maps, const int* levels, int box_w, int box_h, int x0, int y0,
                                         uint32_t* out, int pdl) {
  if (pdl) { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;"); }
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  uint8_t* tile = smem + 1280;
  const int level = MODE == 0 ? 0 : levels[blockIdx.x];
  if (threadIdx.x == 0) mbar_init(bar, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, box_w * box_h);
    tma_load_3d(tile, &maps.lv[level], bar, x0, y0, blockIdx.y);
  }
  mbar_wait(bar, 0);
  uint32_t s = 0;
  for (int i = threadIdx.x; i < box_w * box_h; i += 256) s += tile[i];
  atomicAdd(out + blockIdx.x + gridDim.x * blockIdx.y, s);
}
int main() {
  const int W = 688, H = 518, F = 4, bw = 144, bh = 70;
  std::vector<uint8_t> h((size_t)W * H * F);
  for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)(i * 7 + i / 688);
  uint8_t* d; cudaMalloc(&d, h.size()); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  Maps maps;
  for (int l = 0; l < 8; l++)
    if (tma_encode_u8_3d(&maps.lv[l], d, W, H, F, W, (uint64_t)W * H, bw, bh)) return 1;
  int hl[8] = {0, 1, 2, 3, 4, 5, 6, 7}; int* dl; cudaMalloc(&dl, sizeof hl); cudaMemcpy(dl, hl, sizeof hl, cudaMemcpyHostToDevice);
  uint32_t* out; cudaMalloc(&out, 4 * 64);
  for (int mode = 0; mode < 2; mode++)
    for (int pdl = 0; pdl < 2; pdl++) {
      cudaMemset(out, 0, 4 * 64);
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(8, F); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 1280 + bw * bh; cfg.stream = 0;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = at; cfg.numAttrs = pdl;
      cudaError_t e = mode == 0 ? cudaLaunchKernelEx(&cfg, k<0>, maps, (const int*)dl, bw, bh, 35, 19, out, pdl)
                                : cudaLaunchKernelEx(&cfg, k<1>, maps, (const int*)dl, bw, bh, 35, 19, out, pdl);
      cudaError_t e2 = cudaDeviceSynchronize();
      uint32_t ho[32]; cudaMemcpy(ho, out, sizeof ho, cudaMemcpyDeviceToHost);
      uint32_t ref = 0; for (int r = 0; r < bh; r++) for (int c = 0; c < bw; c++) ref += h[(size_t)(19 + r) * W + 35 + c];
      printf("mode %d pdl %d launch %s sync %s out0 %u ref %u\n", mode, pdl, cudaGetErrorString(e), cudaGetErrorString(e2), ho[0], ref);
      if (e2 != cudaSuccess) return 2;
    }
  return 0;
}
