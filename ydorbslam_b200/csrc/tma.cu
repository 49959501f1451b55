// Host side of tma.cuh: cuTensorMapEncodeTiled fetched through the runtime (no -lcuda at link time).
#include "tma.cuh"

#include <mutex>

namespace ydorb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int tma_encode_u8_3d(CUtensorMap* out, const void* base, uint64_t width, uint64_t rows, uint64_t frames,
                     uint64_t row_stride, uint64_t frame_stride, uint32_t box_w, uint32_t box_h) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return YDORB_E_CUDA;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (row_stride & 15u) || (frame_stride & 15u) || (box_w & 15u) ||
      box_w > 256 || box_h > 256) {
    set_error("internal: tensor map alignment");
    return YDORB_E_INVALID;
  }
  const cuuint64_t dims[3] = {width, rows, frames};
  const cuuint64_t strides[2] = {row_stride, frame_stride};  // bytes, dims 1..2
  const cuuint32_t box[3] = {box_w, box_h, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
    return YDORB_E_CUDA;
  }
  return YDORB_OK;
}

}  // namespace ydorb
