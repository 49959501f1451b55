// Internal definitions shared by the libydorb.so translation units (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/ydorb.h"

namespace ydorb {

constexpr int kPad = 19;       // m_int_maxPadSize    src/orbExtractor.hpp:72
constexpr int kBorder = 16;    // kPad - 3            src/orbExtractor.cpp:550
constexpr int kHalfPatch = 15; // m_int_halfPatchSize src/orbExtractor.hpp:71
constexpr int kPatch = 31;     // m_int_patchSize     src/orbExtractor.hpp:70
constexpr int kMaxLevels = YDORB_MAX_LEVELS;
constexpr int kQtDigits = 16;  // quadtree path digits per code (2 bits each)
constexpr int kQtMapWords = 2048;  // 4^8 bits: occupancy bitmap of the 8 leading path digits (quadtree.cu)

void set_error(const std::string& msg);
void count_launches(int n);  // bookkeeping for ydorb_kernel_launches()
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);
// wait for everything enqueued on `s`.  sleep: first block in the driver on `blocking_event` (created with
// cudaEventBlockingSync and already recorded by the caller where the last kernel ends), then spin through the rest --
// batch calls issued from many host threads leave the cores to each other (policy and measurements: hamming.cu)
bool wait_sleeps();
int wait_stream(cudaStream_t s, cudaEvent_t blocking_event, bool sleep);

// Programmatic dependent launch: the kernels of one extract call form a chain in one stream.  Launched with the
// stream-serialization attribute, the next kernel's CTAs are placed while the previous kernel drains and block in
// pdl_wait() until it has completed and flushed, which takes the launch latency off a single frame's critical path.
// Every kernel launched through launch_chain MUST call pdl_wait() before anything else (also before an early return:
// a kernel that finishes without waiting would release its successor early).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // the successor may be placed as soon as every CTA of this grid is resident; it blocks in its own pdl_wait()
  asm volatile("griddepcontrol.launch_dependents;");
}
#endif
#ifdef __CUDACC__
// sinf/cosf exactly as the reference's host libm computes them (src/orbExtractor.cpp:426: `cos(angle)` / `sin(angle)`
// on a float resolve to the float overloads, fused by g++ -O3 into one sincosf call -- NOT double cos/sin).  This is
// glibc 2.39's sincosf in its x86-64 FMA variant (ARM optimized-routines algorithm): reduction by pi/2 and two short
// polynomials in double, every a + b*c fused, one final rounding to float.  Branches for |y| < 120 only (angles are
// degrees in [0, 360) times pi/180).  Model + exhaustive check against libm: oracle/ydorb_oracle.cpp orc_sincosf_model;
// the device function itself is checked over the same domain by tests/test_gpu_extract.py through ydorb_diag_sincosf.
__device__ __forceinline__ void glibc_sincosf(float y, float& sn, float& cs) {
  const uint32_t top12 = (__float_as_uint(y) >> 20) & 0x7ffu;
  double x = (double)y, xs = x, neg = 1.0;
  int n = 0;
  if (top12 < 0x3f4u) {      // |y| < 0.75
    if (top12 < 0x398u) {    // |y| < 2^-12
      sn = y;
      cs = 1.0f;
      return;
    }
  } else {
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    n = (__double2int_rz(r) + 0x800000) >> 24;
    x = __fma_rn(-(double)n, 0x1.921FB54442D18p0, x);
    xs = ((n + 1) & 2) ? -x : x;     // sign[n & 3] = {1, -1, -1, 1}
    if (n & 2) neg = -1.0;           // second table: cosine coefficients negated
  }
  const double x2 = __dmul_rn(x, x);
  const double x3 = __dmul_rn(x2, xs), x4 = __dmul_rn(x2, x2);
  const double s1 = __fma_rn(x2, -0x1.994eb3774cf24p-13, 0x1.1107605230bc4p-7);
  const double c2 = __fma_rn(x2, neg * 0x1.99343027bf8c3p-16, neg * -0x1.6c087e89a359dp-10);
  const double c1 = __fma_rn(x2, neg * -0x1.ffffffd0c621cp-2, neg);
  const double x5 = __dmul_rn(x2, x3), x6 = __dmul_rn(x2, x4);
  const double s = __fma_rn(x3, -0x1.555545995a603p-3, xs), c = __fma_rn(x4, neg * 0x1.55553e1068f19p-5, c1);
  const float sv = __double2float_rn(__fma_rn(s1, x5, s)), cv = __double2float_rn(__fma_rn(c2, x6, c));
  sn = (n & 1) ? cv : sv;
  cs = (n & 1) ? sv : cv;
}
#endif
bool pdl_enabled();  // YDORB_PDL=0 switches the attribute off
template <class... P, class... A>
inline cudaError_t launch_chain(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, P(std::forward<A>(args))...);
}

#define YD_CUDA(expr)                                                         \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) return ::ydorb::cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define YD_REQUIRE(cond, msg)             \
  do {                                    \
    if (!(cond)) {                        \
      ::ydorb::set_error(msg);            \
      return YDORB_E_INVALID;             \
    }                                     \
  } while (0)

// Geometry of one pyramid level for the current image size (device copy lives in constant-like global memory).
struct LevelGeom {
  int w, h;            // level size (E2, src/orbExtractor.cpp:608-609)
  int stride;          // row stride of the padded buffer (multiple of 16)
  int64_t offset;      // byte offset of the padded buffer inside a frame's pyramid block
  int cell_cols, cell_rows, cell_w, cell_h;  // FAST grid (E3, src/orbExtractor.cpp:556-561)
  int cell_base;       // index of this level's first cell in the frame's cell table
  int quota;           // m_v_keyPointsNumsPerLevel[level]
  int quota_base;      // prefix sum of quotas
  int root_w, root_h;  // quadtree root extents (maxBorder - minBorder)
  int tab_base;        // offset of this level's resize tables (x then y) in the table arrays; level 0 unused
  int tile_base;       // offset of this level's resize tile descriptors in the int table array (pyramid.cu)
  float sf;            // m_v_scaleFactors[level]
  float size;          // (float)(int)(31*sf)
};

struct FrameGeomDev {
  int n_levels;
  int n_cells;          // total cells over all levels
  int n_quota;          // sum of quotas
  int64_t pyr_bytes;    // bytes of one frame's pyramid block
  int raw_cap;          // candidate capacity per frame
  LevelGeom lv[kMaxLevels];
};

// fast2.cu: a block of up to 4 x 2 FAST cells handled by one CTA
struct FastBlock {
  int level;
  int x0, y0, w, h;          // union of the valid cells' ROIs (level coordinates); w = h = 0: no valid cell
  int cell0, ncx, ncy, cols; // first grid cell (index in the frame's cell table), cells covered, the level's cell_cols
  int conv_n, conv_magic;    // bytes -> pairs conversion: 16-byte groups per row, 65536 / conv_n + 1 (i / conv_n = i * magic >> 16)
  int pad;
};

// device pointer to the padded level origin (pixel (0,0) of the level, pad readable around it)
__host__ __device__ inline int64_t level_origin(const LevelGeom& g) {
  return g.offset + (int64_t)kPad * g.stride + kPad;
}

}  // namespace ydorb

// One captured extract pipeline (the kernels of a sub-batch): replayed with a single cudaGraphLaunch while the call's
// arguments stay the same (extractor_api.cu: run_stages)
struct YdGraphEntry {
  int f0, n, stride, capacity, fast_choice;
  size_t frame_stride;
  const void* d_images;
  const void* d_kps;
  const void* d_desc;
  const void* d_nout;
  cudaGraphExec_t exec;
  int launches;               // kernels in the graph (bookkeeping of ydorb_kernel_launches)
  unsigned long long stamp;   // last use, for eviction
};
constexpr int kYdGraphSlots = 24;

// The extractor handle (opaque to C callers).
namespace ydorb {
// first staged source column / row of every tile column / row of a resize launch, as launch parameters (constant
// bank): the thread that issues the TMA box load does not have to wait for a global-memory table read first
constexpr int kPyStartX = 40, kPyStartY = 120;
struct PyrStart {
  short sxa[kPyStartX];
  short sy_lo[kPyStartY];
  int in_params;  // 0: the grid is larger than the arrays, the kernel reads the tables
};
}  // namespace ydorb

struct YdExtractor {
  YdExtractorParams params;
  int device;
  int max_w, max_h, max_batch;
  cudaStream_t stream;      // stream the stages are enqueued on (own_stream unless the caller set one)
  cudaStream_t own_stream;
  cudaStream_t aux_stream;  // second stream of the host entry points: chunk i+1's H2D overlaps chunk i's kernels
  cudaEvent_t ev_fork, ev_join;
  cudaEvent_t ev_block;     // blocking-sync event of the batched host entry point
  cudaStream_t in_stream, out_stream;   // H2D and D2H copy streams of the chunked host entry point
  cudaEvent_t ev_in[16], ev_done[16];   // per chunk: input resident / results ready
  YdGraphEntry graphs[kYdGraphSlots];  // captured pipelines of the current geometry
  unsigned long long graph_clock;
  int graphs_off;           // capture failed once (or YDORB_GRAPHS=0): launch kernel by kernel from now on
  int profiling;            // record per-stage events
  cudaEvent_t ev_ring[64][5];  // per call: pyramid | fast | quadtree | describe boundaries (ring of 64 calls)
  cudaEvent_t* ev;           // slot of the current call
  long long prof_calls;      // calls recorded since profiling was enabled
  float sf[ydorb::kMaxLevels], inv_sf[ydorb::kMaxLevels], sf2[ydorb::kMaxLevels], inv_sf2[ydorb::kMaxLevels];
  int quota[ydorb::kMaxLevels];
  int umax[16];
  // geometry for the current (w,h)
  int cur_w, cur_h, cur_batch;
  int max_cell_w, max_cell_h;     // largest FAST cell ROI of the current geometry (sizes the smem tile)
  ydorb::FrameGeomDev geom;       // host copy
  ydorb::FrameGeomDev* d_geom;    // device copy
  // capacities computed for (max_w,max_h)
  int64_t cap_pyr_bytes;
  int cap_cells, cap_raw, cap_tab;
  // device buffers
  uint8_t* d_pyr;        // [max_batch][cap_pyr_bytes]
  int* d_tab_ofs;        // resize source offsets
  ydorb::PyrStart pyr_start[ydorb::kMaxLevels];  // per destination level: tile start coordinates, passed as launch parameters
  short* d_tab_a;        // resize weights (a0,a1 interleaved)
  ydorb::FastBlock* d_fast_blocks;  // [cap_cells] cell blocks of the current geometry (fast2.cu)
  uint32_t* d_fast_seams;           // [kMaxLevels][276] per-level cell-seam masks of the NMS (fast2.cu)
  int n_fast_blocks, fb_max_w, fb_max_h;
  uint8_t* pyr_base;     // d_pyr of the handle itself (frame_view() shifts d_pyr; TMA coordinates count from here)
  CUtensorMap tmap_level[ydorb::kMaxLevels];  // u8 [frame][row][byte] per pyramid level, box = FAST block tile
  CUtensorMap tmap_pyr[ydorb::kMaxLevels];    // [l]: level l-1 with the box of k_pyr_resize's source rows (l >= 1)
  CUtensorMap tmap_desc[ydorb::kMaxLevels];   // same tensors, box = the keypoint neighbourhood of k_describe
  int4* d_cells;         // per-cell ROI {x0,y0,x1,y1} (level coords), [cap_cells]
  int* d_cell_count;     // [max_batch][cap_cells]
  int* d_cell_start;     // [max_batch][cap_cells]
  int* d_cell_rank;      // [max_batch][cap_cells+1] exclusive scan of counts (candidate order)
  uint32_t* d_raw;       // [max_batch][cap_raw] packed candidates (y:12 | x:12 | response:8, border-relative)
  int* d_raw_counter;    // [max_batch]
  uint32_t* d_qt_keys[3];  // [max_batch][cap_raw] quadtree sort scratch
  uint32_t* d_qt_vals[3];
  uint8_t* d_qt_xpath;   // per level and border-relative column: the 8 leading x-decisions of divideNode (quadtree.cu)
  int qt_xoff[ydorb::kMaxLevels], qt_xlen[ydorb::kMaxLevels], cap_xpath;
  uint32_t* d_qt_bitmap; // [max_batch][kMaxLevels][kQtMapWords] occupancy of 8-digit path prefixes (zero between calls)
  uint32_t* d_qt_fold;   // [max_batch][kMaxLevels][2][kQtMapWords] D-digit occupancy + exclusive popcount prefix per word
  int* d_qt_meta;        // [max_batch][kMaxLevels][8] per segment: n, D, M, deep, level_base
  int* d_level_ncand;    // [max_batch][kMaxLevels]
  int4* d_sel;           // [max_batch][n_quota] selected keypoints {x,y (level coords), response bits, level}
  int* d_sel_n;          // [max_batch][kMaxLevels]
  int* d_status;         // device-side error flags
  // staging for the host entry points
  uint8_t* d_images;     // [max_batch][max_w*max_h]
  YdKeyPoint* d_kps;
  uint8_t* d_desc;
  int32_t* d_nout;
  int out_cap;           // per-frame capacity of d_kps/d_desc
};
