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
constexpr int kPatchRows = 43;    // y-21 .. y+21  (18 px pattern reach + 3 px blur reach)
constexpr int kPatchStride = 64;  // bytes per staged row: 15 aligned words cover x-26-a .. x+26 (IC-angle reach 26)
constexpr int kPatchWords = 15;
constexpr int kHbGroups = 11;     // horizontal-blur outputs per row in groups of 4: patch columns 7 .. 50
constexpr int kHbStride = 4 * kHbGroups;  // u16 per row
constexpr int kHbCol0 = 7;        // patch column of hb[.][0]

__constant__ __align__(16) int8_t c_pattern[1024] = {
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
  __shared__ uint32_t s_pat[256];  // [bit b][lane]: the two pattern points (x0,y0,x1,y1 as s8) of bit b of byte `lane`
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    const int b = i >> 5, l = i & 31;
    s_pat[i] = reinterpret_cast<const uint32_t*>(c_pattern)[l * 8 + b];
  }
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
  // stage the neighbourhood with aligned 32-bit loads (the 19 px reflect-101 pad makes every read legal: keypoints
  // are >= 19 px from the edge and rows are padded to 16 B); pixel (dx, dy) lands at patch[(dy+21)*64 + dx+26+a]
  const uint8_t* src = img + (ptrdiff_t)(kp.y - 21) * lv.stride + (kp.x - 26);
  const int a = (int)(reinterpret_cast<uintptr_t>(src) & 3u);
  {
    const uint32_t* src_al = reinterpret_cast<const uint32_t*>(src - a);
    const int w = lane & 15, stride_w = lv.stride >> 2;
    for (int r = lane >> 4; r < kPatchRows; r += 2)
      if (w < kPatchWords) reinterpret_cast<uint32_t*>(patch)[r * (kPatchStride / 4) + w] = __ldg(src_al + (ptrdiff_t)r * stride_w + w);
  }
  __syncwarp();
  // ---- E5 IC angle
  const uint8_t* ctr = patch + 21 * kPatchStride + 26 + a;
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
  // ---- E6 horizontal blur pass over rows y-21..y+21 (exact, <= 65280).  A task = 4 adjacent outputs at patch
  // columns c0..c0+3 with c0 = 7 + 4*grp, i.e. c0-3 word aligned: taps [18 34 48 56 | 48 34 18 0] as two DP4A.
  for (int t = lane; t < kPatchRows * kHbGroups; t += 32) {
    const int r = t / kHbGroups, grp = t - r * kHbGroups;
    const uint32_t* pw = reinterpret_cast<const uint32_t*>(patch) + r * (kPatchStride / 4) + 1 + grp;
    const uint32_t A = pw[0], B = pw[1], C = pw[2];
    const uint32_t k_lo = 0x38302212u, k_hi = 0x00122230u;  // (18,34,48,56), (48,34,18,0) little endian
    const uint32_t o0 = __dp4a(A, k_lo, __dp4a(B, k_hi, 0u));
    const uint32_t o1 = __dp4a(__funnelshift_r(A, B, 8), k_lo, __dp4a(__funnelshift_r(B, C, 8), k_hi, 0u));
    const uint32_t o2 = __dp4a(__funnelshift_r(A, B, 16), k_lo, __dp4a(__funnelshift_r(B, C, 16), k_hi, 0u));
    const uint32_t o3 = __dp4a(__funnelshift_r(A, B, 24), k_lo, __dp4a(__funnelshift_r(B, C, 24), k_hi, 0u));
    *reinterpret_cast<uint2*>(hb + r * kHbStride + 4 * grp) = make_uint2(o0 | (o1 << 16), o2 | (o3 << 16));
  }
  __syncwarp();
  // ---- E7 rBRIEF
  const float a_rad = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.0));  // :425
  const float cosA = (float)cos((double)a_rad), sinB = (float)sin((double)a_rad);    // :426
  const uint16_t* hctr = hb + (21 - 3) * kHbStride + (26 + a - kHbCol0);  // hb position of (dx = 0, dy = -3)
  uint32_t byte = 0;
#pragma unroll
  for (int b = 0; b < 8; b++) {
    const uint32_t pp = s_pat[b * 32 + lane];
    int val[2];
#pragma unroll
    for (int e = 0; e < 2; e++) {
      const float px = (float)(int8_t)(pp >> (16 * e)), py = (float)(int8_t)(pp >> (16 * e + 8));
      const int rr = __float2int_rn(__fadd_rn(__fmul_rn(px, sinB), __fmul_rn(py, cosA)));  // :428
      const int cc = __float2int_rn(__fsub_rn(__fmul_rn(px, cosA), __fmul_rn(py, sinB)));  // :429
      const uint16_t* h = hctr + rr * kHbStride + cc;
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
