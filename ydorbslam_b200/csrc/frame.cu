// G1 Frame::assignKeyPointsToGrid (src/frame.cpp:249-264, computeLocationInGrid :327-336)
// G2 Frame::getKeyPointsInArea    (src/frame.cpp:337-361)
// M1 OrbMatcher::searchByProjectionInFrameAndMapPoint (src/orbMatcher.cpp:24-64)
//
// The 64 x 48 grid is a CSR table on the device (cell id = ix*48+iy, the order the reference walks it: ix outer,
// iy inner; items inside a cell in ascending keypoint index = insertion order).  The reference's quirks are kept
// in one place each: grid_cell() uses minX for the y coordinate (:329) and in_window() has the inverted x test
// and the `octave < maxLevel` filter (:348-353).
//
// The projection search is greedy in the reference: a keypoint claimed by an earlier map point with
// observations>0 is skipped by later ones (:40).  The device evaluates all map points in parallel against a
// "blocker" table B[idx] (index of the first blocking claimant) and iterates to the fixed point; by induction on
// the map point index the fixed point is unique and equals the sequential result.
#include <algorithm>
#include <vector>

#include "frame.cuh"

namespace ydorb {

__global__ void k_grid_assign(GridParams gp, const YdKeyPoint* __restrict__ kps, int n, int* __restrict__ cell_of,
                              int* __restrict__ counts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = grid_cell(gp, kps[i].x, kps[i].y);
  cell_of[i] = c;
  if (c >= 0) atomicAdd(&counts[c + 1], 1);
}

// single CTA: inclusive scan of counts[1..] so that cell_start[c] = first slot of cell c
__global__ void __launch_bounds__(1024) k_grid_scan(int* __restrict__ cell_start) {
  __shared__ int s[1024];
  const int tid = threadIdx.x;
  // 3072 cells = 3 per thread
  int v[3], sum = 0;
  for (int k = 0; k < 3; k++) { v[k] = cell_start[1 + tid * 3 + k]; sum += v[k]; }
  s[tid] = sum;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    const int t = tid >= o ? s[tid - o] : 0;
    __syncthreads();
    s[tid] += t;
    __syncthreads();
  }
  int run = s[tid] - sum;
  for (int k = 0; k < 3; k++) { run += v[k]; cell_start[1 + tid * 3 + k] = run; }
  if (tid == 0) cell_start[0] = 0;
}

// one thread per cell: items of the cell in ascending keypoint index (= push_back order of :258-261)
__global__ void k_grid_fill(const int* __restrict__ cell_of, int n, const int* __restrict__ cell_start, int* __restrict__ items,
                            int* __restrict__ cursor) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = cell_of[i];
  if (c < 0) return;
  items[cell_start[c] + atomicAdd(&cursor[c], 1)] = i;
}
__global__ void k_grid_sort(const int* __restrict__ cell_start, int* __restrict__ items) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= kGridCells) return;
  const int b = cell_start[c], e = cell_start[c + 1];
  for (int i = b + 1; i < e; i++) {
    const int v = items[i];
    int j = i - 1;
    while (j >= b && items[j] > v) { items[j + 1] = items[j]; j--; }
    items[j + 1] = v;
  }
}

__global__ void k_area_query(GridParams gp, const YdKeyPoint* __restrict__ kps, const int* __restrict__ cell_start,
                             const int* __restrict__ items, float x, float y, float r, int minL, int maxL,
                             int* __restrict__ out, int cap, int* __restrict__ n_out) {
  // single thread: this entry point exists for parity tests of G2, the matcher kernels inline the same walk
  int n = 0;
  CellRange cr;
  if (cell_range(gp, x, y, r, cr)) {
    for (int ix = cr.x0; ix <= cr.x1; ix++)
      for (int iy = cr.y0; iy <= cr.y1; iy++) {
        const int c = ix * kGridRows + iy;
        for (int t = cell_start[c]; t < cell_start[c + 1]; t++) {
          const int idx = items[t];
          if (in_window(kps[idx], x, y, r, minL, maxL)) {
            if (n < cap) out[n] = idx;
            n++;
          }
        }
      }
  }
  *n_out = n;
}

struct SearchArgs {
  GridParams gp;
  const YdKeyPoint* kps;
  const uint4* desc;
  const float* right_x;
  const int* cell_start;
  const int* items;
  const YdMapPointProj* mps;
  const uint4* mp_desc;
  const uint8_t* mp_obs;
  int nmp;
  float th, ratio;
  const int* blocker;   // B[idx]: first blocking claimant (-1: occupied before the call)
  int* blocker_next;    // B'[idx]
  int* last_writer;     // max accepted claimant per keypoint
  int* best_of;         // [nmp] best keypoint of each map point or -1
  int* flags;           // [0] changed, [1] match count
};

// one thread per map point: body of the loop at src/orbMatcher.cpp:26-61
__global__ void __launch_bounds__(128) k_search_projection(SearchArgs A) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A.nmp) return;
  const YdMapPointProj mp = A.mps[j];
  int best_idx = -1;
  if (mp.active) {                                                                   // :27
    const int L = mp.level;
    const float radius = __fmul_rn(A.th, ((double)mp.view_cos > 0.998) ? 2.5f : 4.0f);  // :30, :820-826
    const float r = __fmul_rn(radius, A.gp.sf[L]);
    const uint4 q0 = A.mp_desc[2 * j], q1 = A.mp_desc[2 * j + 1];
    int bestDist = 256, bestLevel = -1, secondDist = 256, secondLevel = -1;
    CellRange cr;
    if (cell_range(A.gp, mp.proj_x, mp.proj_y, r, cr)) {
      for (int ix = cr.x0; ix <= cr.x1; ix++)
        for (int iy = cr.y0; iy <= cr.y1; iy++) {
          const int c = ix * kGridRows + iy;
          const int e = A.cell_start[c + 1];
          for (int t = A.cell_start[c]; t < e; t++) {
            const int idx = A.items[t];
            const YdKeyPoint kp = A.kps[idx];
            if (!in_window(kp, mp.proj_x, mp.proj_y, r, L - 1, L)) continue;         // :31
            if (A.blocker[idx] < j) continue;                                        // :40 slot held, observations>0
            const float rx = A.right_x[idx];
            if (!(rx <= 0.f || fabsf(__fsub_rn(mp.proj_right_x, rx)) <= r)) continue;  // :41
            const int d = hamming256(q0, q1, A.desc[2 * idx], A.desc[2 * idx + 1]);
            if (d < bestDist) {                                                      // :43-52
              secondDist = bestDist; bestDist = d; secondLevel = bestLevel; bestLevel = kp.octave; best_idx = idx;
            } else if (d < secondDist) {
              secondLevel = kp.octave; secondDist = d;
            }
          }
        }
    }
    const bool accept = best_idx >= 0 && bestDist <= 100 &&
                        (bestLevel != secondLevel || (float)bestDist <= __fmul_rn(A.ratio, (float)secondDist));  // :56
    if (!accept) best_idx = -1;
  }
  A.best_of[j] = best_idx;
  if (best_idx >= 0) {
    atomicAdd(&A.flags[1], 1);
    atomicMax(&A.last_writer[best_idx], j);
    if (A.mp_obs[j]) atomicMin(&A.blocker_next[best_idx], j);
  }
}

__global__ void k_blocker_diff(const int* __restrict__ a, const int* __restrict__ b, int n, int* __restrict__ flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && a[i] != b[i]) flags[0] = 1;
}

__global__ void k_blocker_init(const uint8_t* __restrict__ occupied, int n, int* __restrict__ b) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) b[i] = (occupied && occupied[i]) ? -1 : kNoBlocker;
}

__global__ void k_round_init(const int* __restrict__ b0, int n, int* __restrict__ bnext, int* __restrict__ last, int* __restrict__ flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    bnext[i] = (b0[i] == -1) ? -1 : kNoBlocker;  // pre-occupied slots stay blocked
    last[i] = -1;
  }
  if (i == 0) { flags[0] = 0; flags[1] = 0; }
}

}  // namespace ydorb

using namespace ydorb;

extern "C" {

int ydorb_frame_create(YdMatcher* m, const YdFrameGeom* geom, const YdKeyPoint* keypoints, const uint8_t* descriptors,
                       const float* right_x, int n, YdFrame** out) {
  YD_REQUIRE(m && geom && out && n >= 0, "bad argument");
  YD_REQUIRE(n == 0 || (keypoints && descriptors), "null keypoints/descriptors");
  YD_REQUIRE(geom->n_levels >= 1 && geom->n_levels <= kMaxLevels, "n_levels out of range");
  YD_REQUIRE(geom->max_x > geom->min_x && geom->max_y > geom->min_y, "empty image bounds");
  YD_CUDA(cudaSetDevice(m->device));
  YdFrame* f = new YdFrame();
  memset(f, 0, sizeof *f);
  f->m = m;
  f->n = n;
  f->gp.min_x = geom->min_x;
  f->gp.min_y = geom->min_y;
  f->gp.inv_w = (float)kGridCols / (geom->max_x - geom->min_x);  // src/frame.cpp:99
  f->gp.inv_h = (float)kGridRows / (geom->max_y - geom->min_y);  // src/frame.cpp:100
  for (int i = 0; i < kMaxLevels; i++) f->gp.sf[i] = i < geom->n_levels ? geom->scale_factors[i] : 1.f;
  const size_t nn = (size_t)std::max(n, 1);
  cudaError_t e = cudaSuccess;
  auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
  alloc((void**)&f->d_kps, sizeof(YdKeyPoint) * nn);
  alloc((void**)&f->d_desc, 32 * nn);
  alloc((void**)&f->d_right_x, sizeof(float) * nn);
  alloc((void**)&f->d_cell_start, sizeof(int) * (kGridCells + 1) * 2);
  alloc((void**)&f->d_cell_items, sizeof(int) * nn);
  alloc((void**)&f->d_cell_of, sizeof(int) * nn);
  alloc((void**)&f->d_blocker, sizeof(int) * nn);
  alloc((void**)&f->d_blocker2, sizeof(int) * nn);
  alloc((void**)&f->d_last, sizeof(int) * nn);
  alloc((void**)&f->d_flags, sizeof(int) * 4);
  if (e != cudaSuccess) { ydorb_frame_destroy(f); return cuda_fail(e, "frame allocation", __FILE__, __LINE__); }
  cudaStream_t s = m->stream;
  int* cursor = f->d_cell_start + (kGridCells + 1);
  YD_CUDA(cudaMemsetAsync(f->d_cell_start, 0, sizeof(int) * (kGridCells + 1) * 2, s));
  if (n > 0) {
    YD_CUDA(cudaMemcpyAsync(f->d_kps, keypoints, sizeof(YdKeyPoint) * n, cudaMemcpyHostToDevice, s));
    YD_CUDA(cudaMemcpyAsync(f->d_desc, descriptors, (size_t)32 * n, cudaMemcpyHostToDevice, s));
    if (right_x) {
      YD_CUDA(cudaMemcpyAsync(f->d_right_x, right_x, sizeof(float) * n, cudaMemcpyHostToDevice, s));
    } else {
      std::vector<float> neg((size_t)n, -1.0f);
      YD_CUDA(cudaMemcpyAsync(f->d_right_x, neg.data(), sizeof(float) * n, cudaMemcpyHostToDevice, s));
      YD_CUDA(cudaStreamSynchronize(s));
    }
    k_grid_assign<<<(n + 255) / 256, 256, 0, s>>>(f->gp, f->d_kps, n, f->d_cell_of, f->d_cell_start);
  }
  k_grid_scan<<<1, 1024, 0, s>>>(f->d_cell_start);
  if (n > 0) {
    k_grid_fill<<<(n + 255) / 256, 256, 0, s>>>(f->d_cell_of, n, f->d_cell_start, f->d_cell_items, cursor);
    k_grid_sort<<<(kGridCells + 127) / 128, 128, 0, s>>>(f->d_cell_start, f->d_cell_items);
  }
  YD_CUDA(cudaGetLastError());
  YD_CUDA(cudaStreamSynchronize(s));
  *out = f;
  return YDORB_OK;
}

void ydorb_frame_destroy(YdFrame* f) {
  if (!f) return;
  cudaSetDevice(f->m->device);
  void* ptrs[] = {f->d_kps, f->d_desc, f->d_right_x, f->d_cell_start, f->d_cell_items, f->d_cell_of, f->d_blocker,
                  f->d_blocker2, f->d_last, f->d_flags};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  delete f;
}

int ydorb_frame_keypoints_in_area(YdFrame* f, float x, float y, float r, int min_level, int max_level, int32_t* out,
                                  int capacity, int32_t* n_out) {
  YD_REQUIRE(f && n_out && capacity >= 0, "bad argument");
  YdMatcher* m = f->m;
  YD_CUDA(cudaSetDevice(m->device));
  int rc = matcher_reserve(m, sizeof(int) * ((size_t)capacity + 1));
  if (rc != YDORB_OK) return rc;
  int* d_out = reinterpret_cast<int*>(m->d_scratch);
  k_area_query<<<1, 1, 0, m->stream>>>(f->gp, f->d_kps, f->d_cell_start, f->d_cell_items, x, y, r, min_level, max_level,
                                       d_out + 1, capacity, d_out);
  YD_CUDA(cudaGetLastError());
  YD_CUDA(cudaMemcpyAsync(n_out, d_out, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  const int n = std::min(*n_out, capacity);
  if (n > 0 && out) YD_CUDA(cudaMemcpy(out, d_out + 1, sizeof(int) * n, cudaMemcpyDeviceToHost));
  return YDORB_OK;
}

int ydorb_search_by_projection(YdFrame* f, const YdMapPointProj* map_points, const uint8_t* mp_descriptors,
                               const uint8_t* mp_obs_positive, int n_map_points, const uint8_t* kp_occupied, float th,
                               float ratio, int32_t* match_out, int32_t* n_matches) {
  YD_REQUIRE(f && match_out && n_matches && n_map_points >= 0, "bad argument");
  YD_REQUIRE(n_map_points == 0 || (map_points && mp_descriptors && mp_obs_positive), "null map point arrays");
  YdMatcher* m = f->m;
  YD_CUDA(cudaSetDevice(m->device));
  cudaStream_t s = m->stream;
  const int n = f->n, nmp = n_map_points;
  for (int i = 0; i < n; i++) match_out[i] = -1;
  *n_matches = 0;
  if (n == 0 || nmp == 0) return YDORB_OK;
  // staging: map points | descriptors | obs | occupied | best_of
  const size_t o_mp = 0, o_desc = (sizeof(YdMapPointProj) * (size_t)nmp + 255) / 256 * 256,
               o_obs = o_desc + (size_t)32 * nmp, o_occ = o_obs + ((size_t)nmp + 255) / 256 * 256,
               o_best = o_occ + ((size_t)n + 255) / 256 * 256, total = o_best + sizeof(int) * (size_t)nmp;
  int rc = matcher_reserve_io(m, total);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YD_CUDA(cudaMemcpyAsync(io + o_mp, map_points, sizeof(YdMapPointProj) * (size_t)nmp, cudaMemcpyHostToDevice, s));
  YD_CUDA(cudaMemcpyAsync(io + o_desc, mp_descriptors, (size_t)32 * nmp, cudaMemcpyHostToDevice, s));
  YD_CUDA(cudaMemcpyAsync(io + o_obs, mp_obs_positive, (size_t)nmp, cudaMemcpyHostToDevice, s));
  if (kp_occupied) YD_CUDA(cudaMemcpyAsync(io + o_occ, kp_occupied, (size_t)n, cudaMemcpyHostToDevice, s));
  const int gb = (n + 255) / 256;
  k_blocker_init<<<gb, 256, 0, s>>>(kp_occupied ? io + o_occ : nullptr, n, f->d_blocker);
  SearchArgs A;
  A.gp = f->gp;
  A.kps = f->d_kps;
  A.desc = f->d_desc;
  A.right_x = f->d_right_x;
  A.cell_start = f->d_cell_start;
  A.items = f->d_cell_items;
  A.mps = reinterpret_cast<const YdMapPointProj*>(io + o_mp);
  A.mp_desc = reinterpret_cast<const uint4*>(io + o_desc);
  A.mp_obs = io + o_obs;
  A.nmp = nmp;
  A.th = th;
  A.ratio = ratio;
  A.last_writer = f->d_last;
  A.best_of = reinterpret_cast<int*>(io + o_best);
  A.flags = f->d_flags;
  int* cur = f->d_blocker;
  int* nxt = f->d_blocker2;
  int flags[2] = {1, 0};
  // fixed-point iteration: round t makes at least the first t claimants final, in practice 2-5 rounds
  for (int round = 0; round <= nmp + 1 && flags[0]; round++) {
    k_round_init<<<gb, 256, 0, s>>>(f->d_blocker, n, nxt, f->d_last, f->d_flags);
    A.blocker = cur;
    A.blocker_next = nxt;
    k_search_projection<<<(nmp + 127) / 128, 128, 0, s>>>(A);
    k_blocker_diff<<<gb, 256, 0, s>>>(cur, nxt, n, f->d_flags);
    YD_CUDA(cudaGetLastError());
    YD_CUDA(cudaMemcpyAsync(flags, f->d_flags, sizeof flags, cudaMemcpyDeviceToHost, s));
    YD_CUDA(cudaStreamSynchronize(s));
    std::swap(cur, nxt);
    if (cur == f->d_blocker) {  // keep the pristine "occupied before the call" table in d_blocker
      // after an even number of swaps cur is d_blocker again: we must not have overwritten its -1 entries;
      // k_round_init re-derives them from d_blocker, and d_blocker's -1 entries are never changed by atomicMin.
    }
  }
  *n_matches = flags[1];
  YD_CUDA(cudaMemcpy(match_out, f->d_last, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost));
  return YDORB_OK;
}

}  // extern "C"
