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
#include <chrono>
#include <cstdio>
#include <cstdlib>

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
  const uint8_t* occupied;  // [n] or nullptr: slot held before the call (blocks every map point)
  // candidate lists built by the one window walk
  uint32_t* pool;           // packed {idx:19, octave:4, dist:9} in walk order
  int pool_cap;
  int3* live;               // {map point, first pool slot, count}; count < 0: list did not fit, walk again each round
  int* best_prev;           // [live slot] best keypoint in the previous round (-2 before the first)
  int* blockers;            // 3 x [n]: round r reads r%3, claims into (r+1)%3, resets (r+2)%3
  int n;
  int* last_writer;         // [n] largest accepted claimant per keypoint
  int* ctr;                 // [0] live count, [1] pool cursor, [2] matches, [3] done, [4..19] changed flag of round r&15
};

constexpr int kCtrChanged = 4;

__device__ __forceinline__ float search_radius(const SearchArgs& A, const YdMapPointProj& mp) {
  const float radius = __fmul_rn(A.th, ((double)mp.view_cos > 0.998) ? 2.5f : 4.0f);  // :30, :820-826
  // (the device entry point cannot validate levels that live in device memory: clamp the table index)
  return __fmul_rn(radius, A.gp.sf[min(max(mp.level, 0), kMaxLevels - 1)]);
}

// the window walk of map point j (:31 getKeyPointsInArea order, then the gates of :41): f(idx, octave, dist)
template <class F>
__device__ __forceinline__ void search_walk(const SearchArgs& A, int j, const YdMapPointProj& mp, F&& f) {
  const int L = mp.level;
  const float r = search_radius(A, mp);
  const uint4 q0 = A.mp_desc[2 * j], q1 = A.mp_desc[2 * j + 1];
  CellRange cr;
  if (!cell_range(A.gp, mp.proj_x, mp.proj_y, r, cr)) return;
  for (int ix = cr.x0; ix <= cr.x1; ix++) {
    // cells (ix, y0..y1) are adjacent in the CSR table: one item range per column
    const int e = A.cell_start[ix * kGridRows + cr.y1 + 1];
    for (int t = A.cell_start[ix * kGridRows + cr.y0]; t < e; t++) {
      const int idx = A.items[t];
      const YdKeyPoint kp = A.kps[idx];
      if (!in_window(kp, mp.proj_x, mp.proj_y, r, L - 1, L)) continue;             // :31
      const float rx = A.right_x[idx];
      if (!(rx <= 0.f || fabsf(__fsub_rn(mp.proj_right_x, rx)) <= r)) continue;    // :41
      f(idx, kp.octave, hamming256(q0, q1, A.desc[2 * idx], A.desc[2 * idx + 1]));
    }
  }
}

struct Top2 {
  int bestDist = 256, bestLevel = -1, secondDist = 256, secondLevel = -1, best_idx = -1;
  __device__ __forceinline__ void push(int idx, int octave, int d) {  // :43-52
    if (d < bestDist) {
      secondDist = bestDist; bestDist = d; secondLevel = bestLevel; bestLevel = octave; best_idx = idx;
    } else if (d < secondDist) {
      secondLevel = octave; secondDist = d;
    }
  }
  __device__ __forceinline__ int result(float ratio) const {  // :56
    const bool accept = best_idx >= 0 && bestDist <= 100 &&
                        (bestLevel != secondLevel || (float)bestDist <= __fmul_rn(ratio, (float)secondDist));
    return accept ? best_idx : -1;
  }
};

// one warp per map point of [j0, j1): walk the window (lanes over the items of a grid column), keep what does not
// depend on the claims.  Pass 1 counts, lane 0 reserves exactly that much of the pool, pass 2 fills in walk order.
__global__ void __launch_bounds__(256) k_search_build(SearchArgs A, int j0, int j1) {
  const int j = j0 + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (j >= j1) return;
  const YdMapPointProj mp = A.mps[j];
  if (!mp.active) return;  // :27
  const int L = mp.level;
  const float r = search_radius(A, mp);
  CellRange cr;
  if (!cell_range(A.gp, mp.proj_x, mp.proj_y, r, cr)) return;
  auto passes = [&](int idx, int* octave) {
    const YdKeyPoint kp = A.kps[idx];
    *octave = kp.octave;
    if (!in_window(kp, mp.proj_x, mp.proj_y, r, L - 1, L)) return false;             // :31
    if (A.occupied && A.occupied[idx]) return false;                                 // :40, slot held before the call
    const float rx = A.right_x[idx];
    return rx <= 0.f || fabsf(__fsub_rn(mp.proj_right_x, rx)) <= r;                  // :41
  };
  // cells (ix, y0..y1) are adjacent in the CSR table: one item range per grid column.  Lanes take the columns'
  // ranges (32 columns at a time), then the items of all those columns as one flat index space.
  struct Group { int b, incl, total, ncol; };
  auto load_group = [&](int cg) {
    Group g;
    g.ncol = min(32, cr.x1 - cg + 1);
    int len = 0;
    g.b = 0;
    if (lane < g.ncol) {
      g.b = A.cell_start[(cg + lane) * kGridRows + cr.y0];
      len = A.cell_start[(cg + lane) * kGridRows + cr.y1 + 1] - g.b;
    }
    g.incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, g.incl, o);
      if (lane >= o) g.incl += v;
    }
    g.total = __shfl_sync(0xffffffffu, g.incl, 31);
    return g;
  };
  auto locate = [&](const Group& g, int fi) {  // flat index -> CSR slot (all lanes call)
    int t = -1, prev = 0;
    for (int c = 0; c < g.ncol; c++) {
      const int inc = __shfl_sync(0xffffffffu, g.incl, c), bc = __shfl_sync(0xffffffffu, g.b, c);
      if (fi >= prev && fi < inc) t = bc + fi - prev;
      prev = inc;
    }
    return t;
  };
  const bool one_group = cr.x1 - cr.x0 < 32;
  int cnt = 0, first_idx = 0, first_oct = 0, first_total = 0;
  bool first_ok = false;
  for (int cg = cr.x0; cg <= cr.x1; cg += 32) {
    const Group g = load_group(cg);
    if (cg == cr.x0) first_total = g.total;
    for (int base = 0; base < g.total; base += 32) {
      const int t = locate(g, base + lane);
      int octave = 0, idx = 0;
      bool ok = false;
      if (t >= 0) { idx = A.items[t]; ok = passes(idx, &octave); }
      if (cg == cr.x0 && base == 0) { first_ok = ok; first_idx = idx; first_oct = octave; }
      cnt += __popc(__ballot_sync(0xffffffffu, ok));
    }
  }
  if (cnt == 0) return;
  int start = 0;
  if (lane == 0) {
    start = atomicAdd(&A.ctr[1], cnt);
    const int slot = atomicAdd(&A.ctr[0], 1);
    const bool fits = start >= 0 && start + cnt <= A.pool_cap;
    A.live[slot] = fits ? make_int3(j, start, cnt) : make_int3(j, 0, -1);
    A.best_prev[slot] = -2;
    if (!fits) start = -1;
  }
  start = __shfl_sync(0xffffffffu, start, 0);
  if (start < 0) return;
  const uint4 q0 = A.mp_desc[2 * j], q1 = A.mp_desc[2 * j + 1];
  auto emit = [&](bool ok, int idx, int octave, int pos) {
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const int d = hamming256(q0, q1, A.desc[2 * idx], A.desc[2 * idx + 1]);
      A.pool[pos + __popc(bal & ((1u << lane) - 1u))] = ((uint32_t)idx << 13) | ((uint32_t)octave << 9) | (uint32_t)d;
    }
    return pos + __popc(bal);
  };
  if (one_group && first_total <= 32) {  // the usual case: what pass 1 looked at is still in registers
    emit(first_ok, first_idx, first_oct, start);
    return;
  }
  int pos = start;
  for (int cg = cr.x0; cg <= cr.x1; cg += 32) {
    const Group g = load_group(cg);
    for (int base = 0; base < g.total; base += 32) {
      const int t = locate(g, base + lane);
      int octave = 0, idx = 0;
      bool ok = false;
      if (t >= 0) { idx = A.items[t]; ok = passes(idx, &octave); }
      pos = emit(ok, idx, octave, pos);
    }
  }
}

__global__ void k_search_init(SearchArgs A) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A.n) {
    const int v = (A.occupied && A.occupied[i]) ? -1 : kNoBlocker;
    A.blockers[i] = v;
    A.blockers[A.n + i] = v;
    A.last_writer[i] = -1;
  }
  if (i < 4 + 16) A.ctr[i] = 0;
}

// claim round r: every live map point re-decides against the claims of round r-1.  Rounds are launched ahead of the
// host; once a round changes nothing the rest return at once.
__global__ void __launch_bounds__(256) k_search_round(SearchArgs A, int r) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
  if (tid == 0) A.ctr[kCtrChanged + ((r + 1) & 15)] = 0;
  if (r > 0 && A.ctr[kCtrChanged + ((r - 1) & 15)] == 0) return;
  const int* cur = A.blockers + (size_t)(r % 3) * A.n;
  int* nxt = A.blockers + (size_t)((r + 1) % 3) * A.n;
  int* clr = A.blockers + (size_t)((r + 2) % 3) * A.n;
  for (int i = tid; i < A.n; i += nthreads) clr[i] = (A.occupied && A.occupied[i]) ? -1 : kNoBlocker;
  const int nlive = A.ctr[0];
  bool changed = false;
  for (int e = tid; e < nlive; e += nthreads) {
    const int3 lv = A.live[e];
    const int j = lv.x;
    Top2 t2;
    if (lv.z >= 0) {
      for (int i = 0; i < lv.z; i++) {
        const uint32_t c = A.pool[lv.y + i];
        const int idx = (int)(c >> 13);
        if (cur[idx] < j) continue;  // :40 claimed by an earlier map point with observations > 0
        t2.push(idx, (int)((c >> 9) & 15u), (int)(c & 511u));
      }
    } else {
      search_walk(A, j, A.mps[j], [&](int idx, int octave, int d) {
        if (cur[idx] < j) return;
        t2.push(idx, octave, d);
      });
    }
    const int best = t2.result(A.ratio);
    if (best >= 0 && A.mp_obs[j]) atomicMin(&nxt[best], j);
    if (best != A.best_prev[e]) { A.best_prev[e] = best; changed = true; }
  }
  if (changed) A.ctr[kCtrChanged + (r & 15)] = 1;
}

// after `rounds` rounds: if the last one changed nothing the claims are final (:57-60)
__global__ void __launch_bounds__(256) k_search_final(SearchArgs A, int rounds) {
  if (A.ctr[kCtrChanged + ((rounds - 1) & 15)] != 0) return;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
  if (tid == 0) A.ctr[3] = 1;
  const int nlive = A.ctr[0];
  for (int e = tid; e < nlive; e += nthreads) {
    const int best = A.best_prev[e];
    if (best >= 0) {
      atomicAdd(&A.ctr[2], 1);
      atomicMax(&A.last_writer[best], A.live[e].x);
    }
  }
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
  YD_REQUIRE(octaves_in_range(keypoints, n, geom->n_levels), "keypoint octave outside [0, n_levels)");
  YD_CUDA(cudaSetDevice(m->device));
  YdFrame* f = new YdFrame();
  memset(f, 0, sizeof *f);
  f->m = m;
  f->n = n;
  f->n_levels = geom->n_levels;
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
  const int rc = [&]() -> int {
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
  return YDORB_OK;
  }();
  if (rc != YDORB_OK) {  // no leak on a failed copy / launch
    ydorb_frame_destroy(f);
    return rc;
  }
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

// shared body of the host and device entry points; device == true: every pointer is a device pointer
static int search_by_projection_impl(YdFrame* f, const YdMapPointProj* map_points, const uint8_t* mp_descriptors,
                                     const uint8_t* mp_obs_positive, int nmp, const uint8_t* kp_occupied, float th,
                                     float ratio, int32_t* match_out, int32_t* n_matches, bool device) {
  YdMatcher* m = f->m;
  cudaStream_t s = m->stream;
  const int n = f->n;
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  // staging (host entry only): map points | descriptors | obs | occupied
  const size_t o_mp = 0, o_desc = al(sizeof(YdMapPointProj) * (size_t)nmp), o_obs = o_desc + (size_t)32 * nmp,
               o_occ = o_obs + al((size_t)nmp), total = o_occ + al((size_t)n);
  uint8_t* io = nullptr;
  if (!device) {
    const int rc = matcher_reserve_io(m, total);
    if (rc != YDORB_OK) return rc;
    io = reinterpret_cast<uint8_t*>(m->d_io);
  }
  // work: candidate pool | live list | previous best | three blocker tables | counters
  // YDORB_SEARCH_POOL_CAP (debug knob, read per call): a tiny pool forces the candidate lists that do not fit onto the
  // walk-again-every-round path, 0 forces all of them (what a frame with more than 2^19 key points takes) -- the parity
  // tests run both so that those paths are pinned against the oracle too
  int pool_cap = n <= (1 << 19) ? (int)std::min<size_t>((size_t)32 * nmp + 65536, (size_t)1 << 30) : 0;
  if (const char* e = getenv("YDORB_SEARCH_POOL_CAP")) pool_cap = std::max(0, std::min(pool_cap, atoi(e)));
  const size_t w_pool = 0, w_live = w_pool + al(sizeof(uint32_t) * (size_t)std::max(pool_cap, 1)),
               w_prev = w_live + al(sizeof(int3) * (size_t)nmp), w_blk = w_prev + al(sizeof(int) * (size_t)nmp),
               w_ctr = w_blk + al(sizeof(int) * 3 * (size_t)n), w_total = w_ctr + 256;
  const int rc = matcher_reserve(m, w_total);
  if (rc != YDORB_OK) return rc;
  uint8_t* wk = reinterpret_cast<uint8_t*>(m->d_scratch);
  static const bool trace = getenv("YDORB_SEARCH_TRACE") != nullptr;  // diagnostic: per-phase wall times on stderr
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto t_prev = now();
  auto lap = [&](const char* what, int v) {
    if (!trace) return;
    cudaStreamSynchronize(s);
    const auto t = now();
    fprintf(stderr, "[ydorb search] %s %d: %.1f us\n", what, v, std::chrono::duration<double, std::micro>(t - t_prev).count());
    t_prev = t;
  };
  SearchArgs A;
  A.gp = f->gp;
  A.kps = f->d_kps;
  A.desc = f->d_desc;
  A.right_x = f->d_right_x;
  A.cell_start = f->d_cell_start;
  A.items = f->d_cell_items;
  A.mps = device ? map_points : reinterpret_cast<const YdMapPointProj*>(io + o_mp);
  A.mp_desc = reinterpret_cast<const uint4*>(device ? mp_descriptors : io + o_desc);
  A.mp_obs = device ? mp_obs_positive : io + o_obs;
  A.nmp = nmp;
  A.th = th;
  A.ratio = ratio;
  A.occupied = !kp_occupied ? nullptr : device ? kp_occupied : io + o_occ;
  A.pool = reinterpret_cast<uint32_t*>(wk + w_pool);
  A.pool_cap = pool_cap;
  A.live = reinterpret_cast<int3*>(wk + w_live);
  A.best_prev = reinterpret_cast<int*>(wk + w_prev);
  A.blockers = reinterpret_cast<int*>(wk + w_blk);
  A.n = n;
  A.last_writer = device ? match_out : f->d_last;
  A.ctr = reinterpret_cast<int*>(wk + w_ctr);
  if (kp_occupied && !device) YD_CUDA(cudaMemcpyAsync(io + o_occ, kp_occupied, (size_t)n, cudaMemcpyHostToDevice, s));
  k_search_init<<<(std::max(n, 32) + 255) / 256, 256, 0, s>>>(A);
  count_launches(1);
  // host entry: the window walk of a slice runs while the next slice is staged from the caller's pageable arrays
  const int slice = device ? nmp : 32768;
  for (int j0 = 0; j0 < nmp; j0 += slice) {
    const int j1 = std::min(nmp, j0 + slice), c = j1 - j0;
    if (!device) {
      YD_CUDA(cudaMemcpyAsync(io + o_mp + sizeof(YdMapPointProj) * (size_t)j0, map_points + j0,
                              sizeof(YdMapPointProj) * (size_t)c, cudaMemcpyHostToDevice, s));
      YD_CUDA(cudaMemcpyAsync(io + o_desc + (size_t)32 * j0, mp_descriptors + (size_t)32 * j0, (size_t)32 * c,
                              cudaMemcpyHostToDevice, s));
      YD_CUDA(cudaMemcpyAsync(io + o_obs + j0, mp_obs_positive + j0, (size_t)c, cudaMemcpyHostToDevice, s));
    }
    k_search_build<<<(c + 7) / 8, 256, 0, s>>>(A, j0, j1);
    count_launches(1);
  }
  lap("stage+walk", nmp);
  // claim rounds, launched ahead in batches: round t makes at least the first t claimants final (in practice < 12
  // rounds); a round that changes nothing turns the rest of its batch into no-ops
  int h_ctr[4] = {0, 0, 0, 0};
  int rounds = 0;
  const int grid = 148 * 2;
  while (!h_ctr[3]) {
    for (int b = 0; b < 12; b++) k_search_round<<<grid, 256, 0, s>>>(A, rounds++);
    k_search_final<<<grid, 256, 0, s>>>(A, rounds);
    count_launches(13);
    YD_CUDA(cudaGetLastError());
    YD_CUDA(cudaMemcpyAsync(h_ctr, A.ctr, sizeof h_ctr, cudaMemcpyDeviceToHost, s));
    YD_CUDA(cudaStreamSynchronize(s));
    lap("rounds", rounds);
    if (trace) fprintf(stderr, "[ydorb search] live %d pool %d of %d\n", h_ctr[0], h_ctr[1], pool_cap);
    if (rounds > nmp + 16) { set_error("projection search did not converge"); return YDORB_E_STATE; }
  }
  *n_matches = h_ctr[2];
  if (!device) {
    YD_CUDA(cudaMemcpyAsync(match_out, f->d_last, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, s));
    YD_CUDA(cudaStreamSynchronize(s));
    lap("d2h", n);
  }
  return YDORB_OK;
}

int ydorb_search_by_projection(YdFrame* f, const YdMapPointProj* map_points, const uint8_t* mp_descriptors,
                               const uint8_t* mp_obs_positive, int n_map_points, const uint8_t* kp_occupied, float th,
                               float ratio, int32_t* match_out, int32_t* n_matches) {
  YD_REQUIRE(f && match_out && n_matches && n_map_points >= 0, "bad argument");
  YD_REQUIRE(n_map_points == 0 || (map_points && mp_descriptors && mp_obs_positive), "null map point arrays");
  for (int i = 0; i < n_map_points; i++)
    YD_REQUIRE(map_points[i].level >= 0 && map_points[i].level < f->n_levels, "map point level outside [0, n_levels)");
  YD_CUDA(cudaSetDevice(f->m->device));
  for (int i = 0; i < f->n; i++) match_out[i] = -1;
  *n_matches = 0;
  if (f->n == 0 || n_map_points == 0) return YDORB_OK;
  return search_by_projection_impl(f, map_points, mp_descriptors, mp_obs_positive, n_map_points, kp_occupied, th, ratio,
                                   match_out, n_matches, false);
}

int ydorb_search_by_projection_device(YdFrame* f, const YdMapPointProj* map_points, const uint8_t* mp_descriptors,
                                      const uint8_t* mp_obs_positive, int n_map_points, const uint8_t* kp_occupied,
                                      float th, float ratio, int32_t* match_out, int32_t* n_matches) {
  YD_REQUIRE(f && n_matches && n_map_points >= 0 && (f->n == 0 || match_out), "bad argument");
  YD_REQUIRE(n_map_points == 0 || (map_points && mp_descriptors && mp_obs_positive), "null map point arrays");
  YD_REQUIRE((reinterpret_cast<uintptr_t>(mp_descriptors) & 15) == 0, "map point descriptors must be 16-byte aligned");
  YD_CUDA(cudaSetDevice(f->m->device));
  *n_matches = 0;
  if (f->n == 0) return YDORB_OK;
  if (n_map_points == 0) {
    YD_CUDA(cudaMemsetAsync(match_out, 0xFF, sizeof(int32_t) * (size_t)f->n, f->m->stream));
    YD_CUDA(cudaStreamSynchronize(f->m->stream));
    return YDORB_OK;
  }
  return search_by_projection_impl(f, map_points, mp_descriptors, mp_obs_positive, n_map_points, kp_occupied, th, ratio,
                                   match_out, n_matches, true);
}

}  // extern "C"
