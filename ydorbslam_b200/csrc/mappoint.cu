// Map-point side of the front-end (SURVEY 8f "next" rows):
//   N3  MapPoint::computeDistinctiveDescriptors  (src/mapPoint.cpp:169-218)   -> k_distinctive
//   N2  Frame::isInCameraFrustum + MapPoint::predictScaleLevel (src/frame.cpp:295-326, src/mapPoint.cpp:265-277)
//                                                                               -> k_frustum
#include <algorithm>

#include "matcher.cuh"

namespace ydorb {

__device__ __forceinline__ int ham256_rows(const uint32_t (&a)[8], const uint4* __restrict__ b) {
  const uint4 lo = __ldg(b), hi = __ldg(b + 1);
  return __popc(a[0] ^ lo.x) + __popc(a[1] ^ lo.y) + __popc(a[2] ^ lo.z) + __popc(a[3] ^ lo.w) + __popc(a[4] ^ hi.x) +
         __popc(a[5] ^ hi.y) + __popc(a[6] ^ hi.z) + __popc(a[7] ^ hi.w);
}

// One warp per map point, lane = candidate row i (looping when the point has more than 32 observations).  The value
// the reference reads, sorted(distances[i])[k] with k = int(0.5 n), is the smallest v with #{j : d_ij <= v} >= k+1:
// found by bisection over v in [0, 256] (9 rounds), re-evaluating the n distances of the row each round (descriptor j
// is the same address for every lane: one broadcast load).  Nothing is stored, so n is unbounded.
constexpr int kDistWarps = 4;
__global__ void __launch_bounds__(kDistWarps * 32) k_distinctive(const uint4* __restrict__ desc, const int32_t* __restrict__ start,
                                                                 int n_points, int32_t* __restrict__ best) {
  const int p = blockIdx.x * kDistWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (p >= n_points) return;
  const int s0 = start[p], n = start[p + 1] - s0;
  if (n <= 0) {
    if (lane == 0) best[p] = -1;
    return;
  }
  const uint4* d = desc + 2 * (size_t)s0;
  const int k = (int)(0.5 * n);  // :208
  uint32_t best_key = 0xFFFFFFFFu;  // median << 20 | row: minimum = least median, first row on ties (:209 strict <)
  for (int i = lane; i < n; i += 32) {
    uint32_t a[8];
    {
      const uint4 lo = __ldg(d + 2 * i), hi = __ldg(d + 2 * i + 1);
      a[0] = lo.x; a[1] = lo.y; a[2] = lo.z; a[3] = lo.w; a[4] = hi.x; a[5] = hi.y; a[6] = hi.z; a[7] = hi.w;
    }
    int lo_v = 0, hi_v = 256;
    while (lo_v < hi_v) {
      const int mid = (lo_v + hi_v) >> 1;
      int cnt = 0;
      for (int j = 0; j < n; j++) cnt += ham256_rows(a, d + 2 * j) <= mid;
      if (cnt >= k + 1) hi_v = mid; else lo_v = mid + 1;
    }
    best_key = min(best_key, ((uint32_t)lo_v << 20) | (uint32_t)i);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best_key = min(best_key, __shfl_xor_sync(0xffffffffu, best_key, o));
  if (lane == 0) best[p] = (int)(best_key & 0xFFFFFu);
}

// 3x3 * 3x1 + 3x1 the way cv::gemm's small-matrix CV_32F path rounds it (float products summed left to right, then
// (float)(t0*alpha + c*beta) in double with alpha = beta = 1); pinned against cv2.gemm in tests/test_oracle_map.py.
__device__ __forceinline__ float gemm3(const float* __restrict__ a, const float* __restrict__ b, float c) {
  const float t0 = __fadd_rn(__fadd_rn(__fmul_rn(a[0], b[0]), __fmul_rn(a[1], b[1])), __fmul_rn(a[2], b[2]));
  return (float)__dadd_rn((double)t0, (double)c);
}

__global__ void __launch_bounds__(256) k_frustum(const YdCamera cam, const YdMapPoint3d* __restrict__ mps, int n,
                                                 float view_cos_limit, YdMapPointProj* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const YdMapPoint3d mp = mps[i];
  YdMapPointProj o;
  o.proj_x = o.proj_y = o.proj_right_x = o.view_cos = 0.f;
  o.level = 0;
  o.active = 0;  // m_b_isTrackInView = false (:296)
  float pc[3];
#pragma unroll
  for (int r = 0; r < 3; r++) pc[r] = gemm3(cam.R + 3 * r, mp.pos, cam.t[r]);            // :300
  const float x = __fadd_rn(__fdiv_rn(__fmul_rn(cam.fx, pc[0]), pc[2]), cam.cx);          // :302
  const float y = __fadd_rn(__fdiv_rn(__fmul_rn(cam.fy, pc[1]), pc[2]), cam.cy);          // :303
  const float max_d = __fmul_rn(1.2f, mp.max_distance), min_d = __fmul_rn(0.8f, mp.min_distance);  // mapPoint.hpp:67-74
  float v[3];
#pragma unroll
  for (int r = 0; r < 3; r++) v[r] = __fsub_rn(mp.pos[r], cam.origin[r]);                 // :307
  double s = 0, dot = 0;
#pragma unroll
  for (int r = 0; r < 3; r++) {
    s = __dadd_rn(s, __dmul_rn((double)v[r], (double)v[r]));                              // cv::norm, double accumulator
    dot = __dadd_rn(dot, __dmul_rn((double)v[r], (double)mp.normal[r]));                  // Mat::dot, double accumulator
  }
  const float distance = (float)__dsqrt_rn(s);                                            // :308
  const float view_cos = (float)__ddiv_rn(dot, (double)distance);                         // :311
  const bool in_image = x >= cam.min_x && x < cam.max_x && y >= cam.min_y && y < cam.max_y;  // :292-294
  if (pc[2] >= 0.0f && in_image && distance >= min_d && distance <= max_d && view_cos >= view_cos_limit) {  // :312
    const float ratio = __fdiv_rn(mp.max_distance, distance);                             // mapPoint.cpp:269
    int level = (int)ceil(__ddiv_rn(log((double)ratio), (double)cam.log_scale_factor));   // ::log(double), :271
    if (level < 0) level = 0;
    else if (level >= cam.n_levels) level = cam.n_levels - 1;
    o.active = 1;
    o.proj_x = x;
    o.proj_right_x = __fsub_rn(x, __fdiv_rn(cam.bf, pc[2]));                              // :318
    o.proj_y = y;
    o.level = level;
    o.view_cos = view_cos;
  }
  out[i] = o;
}

}  // namespace ydorb

using namespace ydorb;

extern "C" {

int ydorb_distinctive_descriptors_device(YdMatcher* m, const uint8_t* d_desc, const int32_t* d_start, int n_points,
                                         int32_t* d_best) {
  YD_REQUIRE(m && d_start && d_best && n_points >= 0, "bad argument");
  if (n_points == 0) return YDORB_OK;
  YD_REQUIRE(((uintptr_t)d_desc & 15) == 0, "descriptor block must be 16-byte aligned");
  YD_CUDA(cudaSetDevice(m->device));
  k_distinctive<<<(n_points + kDistWarps - 1) / kDistWarps, kDistWarps * 32, 0, m->stream>>>(
      reinterpret_cast<const uint4*>(d_desc), d_start, n_points, d_best);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

int ydorb_distinctive_descriptors(YdMatcher* m, const uint8_t* desc, const int32_t* start, int n_points, int32_t* best) {
  YD_REQUIRE(m && start && best && n_points >= 0, "bad argument");
  if (n_points == 0) return YDORB_OK;
  YD_REQUIRE(start[0] == 0, "start[0] must be 0");
  for (int p = 0; p < n_points; p++) YD_REQUIRE(start[p + 1] >= start[p], "start must be non-decreasing");
  const int total = start[n_points];
  YD_REQUIRE(total < (1 << 20) * 1024 && (total == 0 || desc), "bad descriptor block");
  for (int p = 0; p < n_points; p++) YD_REQUIRE(start[p + 1] - start[p] < (1 << 20), "more than 2^20 observations of one point");
  YD_CUDA(cudaSetDevice(m->device));
  const size_t db = ((size_t)std::max(total, 1) * 32 + 255) / 256 * 256, sb = (sizeof(int32_t) * (n_points + 1) + 255) / 256 * 256;
  int rc = matcher_reserve_io(m, db + sb + sizeof(int32_t) * (size_t)n_points);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  if (total > 0) YD_CUDA(cudaMemcpyAsync(io, desc, (size_t)total * 32, cudaMemcpyHostToDevice, m->stream));
  YD_CUDA(cudaMemcpyAsync(io + db, start, sizeof(int32_t) * (n_points + 1), cudaMemcpyHostToDevice, m->stream));
  rc = ydorb_distinctive_descriptors_device(m, io, reinterpret_cast<int32_t*>(io + db), n_points,
                                            reinterpret_cast<int32_t*>(io + db + sb));
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(best, io + db + sb, sizeof(int32_t) * (size_t)n_points, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

int ydorb_frustum_device(YdMatcher* m, const YdCamera* cam, const YdMapPoint3d* d_map_points, int n, float view_cos_limit,
                         YdMapPointProj* d_out) {
  YD_REQUIRE(m && cam && n >= 0, "bad argument");
  if (n == 0) return YDORB_OK;
  YD_REQUIRE(d_map_points && d_out, "null pointer");
  YD_REQUIRE(cam->n_levels >= 1 && cam->n_levels <= YDORB_MAX_LEVELS, "n_levels out of range");
  YD_CUDA(cudaSetDevice(m->device));
  k_frustum<<<(n + 255) / 256, 256, 0, m->stream>>>(*cam, d_map_points, n, view_cos_limit, d_out);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

int ydorb_frustum(YdMatcher* m, const YdCamera* cam, const YdMapPoint3d* map_points, int n, float view_cos_limit,
                  YdMapPointProj* out) {
  YD_REQUIRE(m && cam && n >= 0, "bad argument");
  if (n == 0) return YDORB_OK;
  YD_REQUIRE(map_points && out, "null pointer");
  YD_CUDA(cudaSetDevice(m->device));
  const size_t ib = (sizeof(YdMapPoint3d) * (size_t)n + 255) / 256 * 256;
  int rc = matcher_reserve_io(m, ib + sizeof(YdMapPointProj) * (size_t)n);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YD_CUDA(cudaMemcpyAsync(io, map_points, sizeof(YdMapPoint3d) * (size_t)n, cudaMemcpyHostToDevice, m->stream));
  rc = ydorb_frustum_device(m, cam, reinterpret_cast<const YdMapPoint3d*>(io), n, view_cos_limit,
                            reinterpret_cast<YdMapPointProj*>(io + ib));
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(out, io + ib, sizeof(YdMapPointProj) * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

}  // extern "C"
