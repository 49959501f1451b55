// E5 computeOrientation (src/orbExtractor.cpp:400-421), E6 GaussianBlur 7x7 sigma 2 (:385-386),
// E7 computeDescriptors (:422-454) and the output assembly of extractAndCompute (:372-398).
//
// One warp per selected keypoint.  The warp stages the 43 x 53 px neighbourhood of the keypoint from the padded,
// UNBLURRED level (rows y-21..y+21, columns x-26..x+26) in shared memory and derives everything from it:
//   * IC angle: integer moments over the reference's footprint (half-width table of SURVEY F4: rows +-1..+-10 only
//     their centre column, rows +-11..+-15 are 53/51/45/39/31 px wide), lane-parallel + shuffle reduction, then
//     cv::fastAtan2's polynomial with individually rounded fp32 operations;
//   * blur: OpenCV's 8.8 fixed point 7-tap kernel [18 34 48 56 48 34 18]; the horizontal pass (exact in 16 bits) is
//     done once for the 43 x 37 window, the vertical pass only at the 512 sampled positions, rounded once
//     ((v + 32768) >> 16) -- bit-identical to blurring the whole level, without writing a blurred level to HBM;
//   * rBRIEF: lane j produces byte j (8 comparisons), coordinates rotated in fp32 without contraction.
#include "kernels.cuh"

namespace ydorb {

constexpr int kDescWarps = 4;
constexpr int kPatchRows = 43;   // y-21 .. y+21  (18 px pattern reach + 3 px blur reach)
constexpr int kPatchCols = 53;   // x-26 .. x+26  (IC-angle reach 26)
constexpr int kPatchStride = 56;
constexpr int kHbCols = 37;      // x-18 .. x+18
constexpr int kHbStride = 38;

__constant__ int8_t c_pattern[1024] = {
#include "brief_pattern.inc"
};

__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  // cv::fastAtan2 scalar path; every operation rounded separately (no FMA contraction)
  const float scale = (float)(180.0 / 3.14159265358979323846);
  const float p1 = __fmul_rn(0.9997878412794807f, scale), p3 = __fmul_rn(-0.3258083974640975f, scale),
              p5 = __fmul_rn(0.1555786518463281f, scale), p7 = __fmul_rn(-0.04432655554792128f, scale);
  const float eps = 2.2204460492503131e-16f;  // (float)DBL_EPSILON
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

__global__ void __launch_bounds__(kDescWarps * 32)
k_describe(const FrameGeomDev* __restrict__ g, const uint8_t* __restrict__ pyr, const int4* __restrict__ sel,
           const int* __restrict__ sel_n, YdKeyPoint* __restrict__ kps, uint8_t* __restrict__ desc, int capacity,
           int32_t* __restrict__ n_out, int* __restrict__ status) {
  __shared__ __align__(16) uint8_t s_patch[kDescWarps][kPatchRows * kPatchStride];
  __shared__ __align__(16) uint16_t s_hb[kDescWarps][kPatchRows * kHbStride];
  __shared__ int8_t s_pat[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_pat[i] = c_pattern[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = blockIdx.x * kDescWarps + warp;
  const int frame = blockIdx.y;
  if (slot >= g->n_quota) return;
  int level = 0;
  for (int l = 1; l < g->n_levels; l++)
    if (slot >= g->lv[l].quota_base) level = l;
  const LevelGeom lv = g->lv[level];
  const int j = slot - lv.quota_base;
  const int* cnt = sel_n + (size_t)frame * kMaxLevels;
  int out_base = 0, total = 0;
  for (int l = 0; l < g->n_levels; l++) {
    const int c = cnt[l];
    if (l < level) out_base += c;
    total += c;
  }
  if (slot == 0 && lane == 0) {
    n_out[frame] = min(total, capacity);
    if (total > capacity) atomicOr(status, kStatusOutOverflow);
  }
  if (j >= cnt[level]) return;
  const int oi = out_base + j;
  if (oi >= capacity) return;
  const int4 kp = sel[(size_t)frame * g->n_quota + slot];  // x, y (level coords), response, level
  const uint8_t* img = pyr + (size_t)frame * g->pyr_bytes + level_origin(lv);
  uint8_t* patch = s_patch[warp];
  uint16_t* hb = s_hb[warp];
  // stage the neighbourhood (the 19 px reflect-101 pad makes every read legal: keypoints are >= 19 px from the edge)
  {
    const uint8_t* src = img + (ptrdiff_t)(kp.y - 21) * lv.stride + (kp.x - 26);
    for (int i = lane; i < kPatchRows * kPatchCols; i += 32) {
      const int r = i / kPatchCols, c = i - r * kPatchCols;
      patch[r * kPatchStride + c] = src[(ptrdiff_t)r * lv.stride + c];
    }
  }
  __syncwarp();
  // ---- E5 IC angle
  const uint8_t* ctr = patch + 21 * kPatchStride + 26;
  int m10 = 0, m01 = 0;
  if (lane < 31) m10 += (lane - 15) * ctr[lane - 15];                       // v = 0 row, u = -15..15  (:404-406)
  if (lane < 10) {                                                          // v = 1..10: half width 0 (F4)
    const int v = lane + 1;
    m01 += v * ((int)ctr[v * kPatchStride] - (int)ctr[-v * kPatchStride]);
  }
  {
    const int hw[5] = {26, 25, 22, 19, 15};                                 // v = 11..15 (F4)
#pragma unroll
    for (int k = 0; k < 5; k++) {
      const int v = 11 + k, d = hw[k];
      int vs = 0;
      for (int u = -d + lane; u <= d; u += 32) {
        const int pos = ctr[v * kPatchStride + u], neg = ctr[-v * kPatchStride + u];
        vs += pos - neg;
        m10 += u * (pos + neg);
      }
      m01 += v * vs;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m10 += __shfl_xor_sync(0xffffffffu, m10, o);
    m01 += __shfl_xor_sync(0xffffffffu, m01, o);
  }
  const float angle = fast_atan2_deg((float)m01, (float)m10);               // :419
  // ---- E6 horizontal blur pass over rows y-21..y+21, columns x-18..x+18 (exact, <= 65280)
  for (int i = lane; i < kPatchRows * kHbCols; i += 32) {
    const int r = i / kHbCols, c = i - r * kHbCols;
    const uint8_t* p = patch + r * kPatchStride + (c + 26 - 18 - 3);
    const int acc = 18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3];
    hb[r * kHbStride + c] = (uint16_t)acc;
  }
  __syncwarp();
  // ---- E7 rBRIEF
  const float a_rad = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.0));  // :425
  const float cosA = (float)cos((double)a_rad), sinB = (float)sin((double)a_rad);    // :426
  uint32_t byte = 0;
#pragma unroll
  for (int b = 0; b < 8; b++) {
    int val[2];
#pragma unroll
    for (int e = 0; e < 2; e++) {
      const int idx = lane * 16 + 2 * b + e;
      const float px = (float)s_pat[2 * idx], py = (float)s_pat[2 * idx + 1];
      const int rr = __float2int_rn(__fadd_rn(__fmul_rn(px, sinB), __fmul_rn(py, cosA)));  // :428
      const int cc = __float2int_rn(__fsub_rn(__fmul_rn(px, cosA), __fmul_rn(py, sinB)));  // :429
      const uint16_t* h = hb + (rr + 21 - 3) * kHbStride + (cc + 18);
      const uint32_t acc = 18u * (h[0] + h[6 * kHbStride]) + 34u * (h[kHbStride] + h[5 * kHbStride]) +
                           48u * (h[2 * kHbStride] + h[4 * kHbStride]) + 56u * h[3 * kHbStride];
      val[e] = (int)((acc + 32768u) >> 16);
    }
    byte |= (uint32_t)(val[0] < val[1]) << b;
  }
  // gather 4 bytes per lane group and store the 32-byte descriptor as 8 x u32
  uint32_t word = byte;
  word |= __shfl_down_sync(0xffffffffu, byte, 1) << 8;
  word |= __shfl_down_sync(0xffffffffu, byte, 2) << 16;
  word |= __shfl_down_sync(0xffffffffu, byte, 3) << 24;
  uint32_t* dout = reinterpret_cast<uint32_t*>(desc + ((size_t)frame * capacity + oi) * 32);
  if ((lane & 3) == 0) dout[lane >> 2] = word;
  if (lane == 0) {
    YdKeyPoint k;
    k.x = (float)kp.x;
    k.y = (float)kp.y;
    if (level != 0) {  // keyPoint.pt *= scale  (:391-396)
      k.x = __fmul_rn(k.x, lv.sf);
      k.y = __fmul_rn(k.y, lv.sf);
    }
    k.size = lv.size;             // :600
    k.angle = angle;
    k.response = (float)kp.z;
    k.octave = level;             // :599
    k.class_id = -1;
    kps[(size_t)frame * capacity + oi] = k;
  }
}

int launch_describe(const YdExtractor* ex, int batch, YdKeyPoint* d_kps, uint8_t* d_desc, int capacity,
                    int32_t* d_nout, cudaStream_t s) {
  dim3 grid((ex->geom.n_quota + kDescWarps - 1) / kDescWarps, batch);
  k_describe<<<grid, kDescWarps * 32, 0, s>>>(ex->d_geom, ex->d_pyr, ex->d_sel, ex->d_sel_n, d_kps, d_desc, capacity,
                                              d_nout, ex->d_status);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
