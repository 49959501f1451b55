// TMA (cp.async.bulk.tensor) plumbing shared by the tile stagers (fast2.cu, pyramid.cu, describe.cu).
//
// Every pyramid level of a handle is described by one CUtensorMap over u8 [frame][row][byte]: the level's padded
// buffer has a 16-byte-multiple row stride, a 256-byte-aligned offset inside the frame's pyramid block and the frame
// pitch (pyr_bytes) is a multiple of 256, which is all cuTensorMapEncodeTiled asks for.  A box may start at ANY byte
// coordinate (the element is one byte), so a kernel gets its ROI with column 0 at smem column 0 whatever the ROI's
// alignment in HBM; rows/bytes outside the tensor are zero-filled.  The driver entry point is fetched at run time
// through cudaGetDriverEntryPoint, so libydorb.so keeps linking the CUDA runtime only.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace ydorb {

// host: encode a 3-D u8 map {inner bytes, rows, frames}; box = {box_w (multiple of 16, <= 256), box_h (<= 256), 1}
int tma_encode_u8_3d(CUtensorMap* out, const void* base, uint64_t width, uint64_t rows, uint64_t frames,
                     uint64_t row_stride, uint64_t frame_stride, uint32_t box_w, uint32_t box_h);

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// box load: coordinates {byte, row, frame}; completes `box bytes` on the mbarrier
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
#endif

}  // namespace ydorb
