// Launch wrappers of the extraction stages (one translation unit each).
#pragma once
#include "common.cuh"

namespace ydorb {

// E2 computePyramid (src/orbExtractor.cpp:605-621)
int launch_pyramid(const YdExtractor* ex, const uint8_t* d_images, int stride, size_t frame_stride, int batch,
                   cudaStream_t s);
// appends the resize tile descriptors to the int table array (host side of the geometry set-up)
int pyramid_tile_tables(FrameGeomDev* g, std::vector<int>* tab);
void pyramid_box(const FrameGeomDev& g, int l, int* box_w, int* box_h);
void pyramid_tile_starts(const FrameGeomDev& g, const std::vector<int>& tab, PyrStart* out);
// E3 per-cell FAST with threshold fallback (src/orbExtractor.cpp:562-593)
int launch_fast(const YdExtractor* ex, int batch, cudaStream_t s);
// CTA-per-cell-block form (fast2.cu) and its host-side set-up
int launch_fast_blocks(const YdExtractor* ex, int batch, cudaStream_t s);
int fast2_build_blocks(const FrameGeomDev& g, const std::vector<int4>& cells, std::vector<FastBlock>* out,
                       std::vector<uint32_t>* seams, int* max_w, int* max_h);
size_t fast2_smem_bytes(int max_w, int max_h);
void fast2_box(int max_w, int max_h, int* box_w, int* box_h);
// E4 distributeQuadTree (src/orbExtractor.cpp:455-544)
int launch_quadtree(const YdExtractor* ex, int batch, cudaStream_t s);
// E5+E6+E7 orientation, blur, descriptors, output assembly (src/orbExtractor.cpp:379-454)
int launch_describe(const YdExtractor* ex, int batch, YdKeyPoint* d_kps, uint8_t* d_desc, int capacity,
                    int32_t* d_nout, cudaStream_t s);

void describe_box(int* box_w, int* box_h);  // TMA box of k_describe's patch loads

// status bits written by kernels into YdExtractor::d_status
constexpr int kStatusRawOverflow = 1;   // candidate buffer too small
constexpr int kStatusSortDepth = 2;     // std::sort emulation hit an internal limit
constexpr int kStatusOutOverflow = 4;   // caller capacity too small

}  // namespace ydorb
