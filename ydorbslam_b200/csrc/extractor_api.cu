// C ABI of the extractor (include/ydorb.h): handle life cycle, per-size geometry and OpenCV-compatible resize
// tables built on the host, stage orchestration on the handle's stream.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "kernels.cuh"
#include "tma.cuh"

namespace ydorb {

static thread_local std::string g_last_error;
static std::atomic<long long> g_launches{0};

static thread_local long long t_launches = 0;  // this thread's share (graph capture counts the kernels of one capture)
void count_launches(int n) {
  g_launches.fetch_add(n, std::memory_order_relaxed);
  t_launches += n;
}

bool pdl_enabled() {
  static const bool on = !(getenv("YDORB_PDL") && atoi(getenv("YDORB_PDL")) == 0);
  return on;
}

void set_error(const std::string& msg) { g_last_error = msg; }

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  g_last_error = buf;
  return YDORB_E_CUDA;
}

static inline int cv_round(float v) { return (int)lrintf(v); }  // cvRound(float): half to even
static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// OrbExtractor ctor tables, src/orbExtractor.cpp:325-353
static void ctor_tables(YdExtractor* ex) {
  const YdExtractorParams& p = ex->params;
  const float sf = p.scale_factor;
  int per_level = (int)round(p.n_features * (1 - 1.0 / sf) / (1.0 - pow(1.0 / sf, p.n_levels)));
  int sum = 0;
  for (int i = 0; i < p.n_levels; i++) {
    ex->sf[i] = (float)pow((double)sf, (double)i);
    ex->sf2[i] = (float)pow((double)sf, (double)(2 * i));
    ex->inv_sf[i] = (float)pow((double)sf, (double)(-i));
    ex->inv_sf2[i] = (float)pow((double)sf, (double)(-2 * i));
    if (i < p.n_levels - 1) {
      ex->quota[i] = per_level;
      sum += per_level;
      per_level = (int)round((double)((float)per_level / sf));
    } else {
      ex->quota[i] = std::max(p.n_features - sum, 0);
    }
  }
  // m_v_maxXcords as the reference computes it (resize + push_back, "+v^2"): {0 x11, 26,25,22,19,15}; the kernel
  // hard-codes this footprint, recompute here only to assert the constant stays what describe.cu assumes.
  static const int expect[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 26, 25, 22, 19, 15};
  memcpy(ex->umax, expect, sizeof expect);
}

// one axis of cv::resize INTER_LINEAR 8U: source offset and 11-bit weights
static void linear_axis(int sn, int dn, int* ofs, short* a) {
  const double scale = 1.0 / ((double)dn / (double)sn);
  for (int d = 0; d < dn; d++) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= sn - 1) { s = sn - 1; f = 0.f; }
    ofs[d] = s;
    a[2 * d] = (short)cv_round((1.f - f) * 2048.f);
    a[2 * d + 1] = (short)cv_round(f * 2048.f);
  }
}

struct HostGeom {
  FrameGeomDev g;
  std::vector<int> tab_ofs;
  std::vector<short> tab_a;
  std::vector<int4> cells;
  int64_t tested_px;
};

// Geometry for an image size: level sizes (:608-609), padded layout, FAST cell grid (:550-579), quadtree roots.
static int build_geom(const YdExtractor* ex, int w, int h, HostGeom* out) {
  FrameGeomDev& g = out->g;
  memset(&g, 0, sizeof g);
  const int L = ex->params.n_levels;
  g.n_levels = L;
  int64_t off = 0;
  int cell_base = 0, quota_base = 0, tab_base = 0;
  out->tested_px = 0;
  for (int l = 0; l < L; l++) {
    LevelGeom& lv = g.lv[l];
    lv.w = cv_round((float)w * ex->inv_sf[l]);
    lv.h = cv_round((float)h * ex->inv_sf[l]);
    if (lv.w <= kPad || lv.h <= kPad) {
      set_error("image too small: every pyramid level must be larger than the 19 px border");
      return YDORB_E_INVALID;
    }
    if (lv.w - 2 * kBorder > 4095 || lv.h - 2 * kBorder > 4095) {
      set_error("image too large: level extents above 4127 px are not supported by the candidate packing");
      return YDORB_E_INVALID;
    }
    lv.stride = round_up(lv.w + 2 * kPad, 16);
    lv.offset = off;
    off += (int64_t)round_up(lv.stride * (lv.h + 2 * kPad), 256);
    const int minB = kBorder, maxBX = lv.w - kBorder, maxBY = lv.h - kBorder;
    const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
    const int cols = width > 0 ? (int)(width / 30.0f) : 0, rows = height > 0 ? (int)(height / 30.0f) : 0;
    lv.cell_base = cell_base;
    lv.root_w = maxBX - minB;
    lv.root_h = maxBY - minB;
    if (cols > 0 && rows > 0) {
      lv.cell_cols = cols;
      lv.cell_rows = rows;
      lv.cell_w = (int)ceilf(width / cols);
      lv.cell_h = (int)ceilf(height / rows);
      for (int i = 0; i < rows; i++) {
        const float iniY = (float)(minB + i * lv.cell_h);
        float maxY = iniY + lv.cell_h + 6;
        const bool skip_row = iniY >= maxBY - 3;
        if (maxY > maxBY) maxY = (float)maxBY;
        for (int j = 0; j < cols; j++) {
          const float iniX = (float)(minB + j * lv.cell_w);
          float maxX = iniX + lv.cell_w + 6;
          const bool skip = skip_row || iniX >= maxBX - 6;
          if (maxX > maxBX) maxX = (float)maxBX;
          int4 c = make_int4((int)iniX, (int)iniY, (int)maxX, (int)maxY);
          if (skip || c.z - c.x < 7 || c.w - c.y < 7) c = make_int4(0, 0, 0, 0);
          else out->tested_px += (int64_t)(c.z - c.x - 6) * (c.w - c.y - 6);
          if (c.z - c.x > 66 || c.w - c.y > 66) {
            set_error("internal: FAST cell larger than the shared-memory tile");
            return YDORB_E_INVALID;
          }
          out->cells.push_back(c);
        }
      }
      cell_base += cols * rows;
    }
    lv.quota = ex->quota[l];
    lv.quota_base = quota_base;
    quota_base += lv.quota;
    lv.sf = ex->sf[l];
    lv.size = (float)(int)(kPatch * ex->sf[l]);
    lv.tab_base = tab_base;
    if (l > 0) {
      out->tab_ofs.resize(tab_base + lv.w + lv.h);
      out->tab_a.resize(2 * (size_t)(tab_base + lv.w + lv.h));
      linear_axis(g.lv[l - 1].w, lv.w, out->tab_ofs.data() + tab_base, out->tab_a.data() + 2 * (size_t)tab_base);
      linear_axis(g.lv[l - 1].h, lv.h, out->tab_ofs.data() + tab_base + lv.w,
                  out->tab_a.data() + 2 * (size_t)(tab_base + lv.w));
      tab_base += lv.w + lv.h;
    }
  }
  {
    const int rc = pyramid_tile_tables(&g, &out->tab_ofs);
    if (rc != YDORB_OK) return rc;
  }
  g.n_cells = cell_base;
  g.n_quota = quota_base;
  g.pyr_bytes = off;
  // candidate capacity: NMS keeps at most one pixel per 2x2 block; default provision is one per 8 tested pixels
  g.raw_cap = (int)std::max<int64_t>(out->tested_px / 8 + 1024, 4096);
  return YDORB_OK;
}

static void drop_graphs(YdExtractor* ex);

static int set_size(YdExtractor* ex, int w, int h) {
  if (ex->cur_w == w && ex->cur_h == h) return YDORB_OK;
  HostGeom hg;
  int rc = build_geom(ex, w, h, &hg);
  if (rc != YDORB_OK) return rc;
  if (hg.g.pyr_bytes > ex->cap_pyr_bytes || hg.g.n_cells > ex->cap_cells || hg.g.raw_cap > ex->cap_raw ||
      (int)hg.tab_ofs.size() > ex->cap_tab) {
    set_error("image larger than the max_width/max_height the extractor was created with");
    return YDORB_E_CAPACITY;
  }
  drop_graphs(ex);  // captured for the previous geometry
  ex->geom = hg.g;
  ex->max_cell_w = ex->max_cell_h = 8;
  for (const int4& c : hg.cells) {
    ex->max_cell_w = std::max(ex->max_cell_w, c.z - c.x);
    ex->max_cell_h = std::max(ex->max_cell_h, c.w - c.y);
  }
  YD_CUDA(cudaMemcpyAsync(ex->d_geom, &ex->geom, sizeof(FrameGeomDev), cudaMemcpyHostToDevice, ex->stream));
  pyramid_tile_starts(ex->geom, hg.tab_ofs, ex->pyr_start);
  if (!hg.tab_ofs.empty()) {
    YD_CUDA(cudaMemcpyAsync(ex->d_tab_ofs, hg.tab_ofs.data(), hg.tab_ofs.size() * sizeof(int), cudaMemcpyHostToDevice, ex->stream));
    YD_CUDA(cudaMemcpyAsync(ex->d_tab_a, hg.tab_a.data(), hg.tab_a.size() * sizeof(short), cudaMemcpyHostToDevice, ex->stream));
  }
  {
    // x-decisions of divideNode per column (src/orbExtractor.cpp:7-33: which half a point falls into horizontally depends
    // on its x alone): bit 7 - d of entry x = "right" at split d, d = 0..7
    std::vector<uint8_t> xp;
    for (int l = 0; l < kMaxLevels; l++) {
      ex->qt_xoff[l] = (int)xp.size();
      ex->qt_xlen[l] = 0;
      if (l >= hg.g.n_levels) continue;
      const int rw = hg.g.lv[l].root_w, len = std::max(rw, 0) + 64;
      ex->qt_xlen[l] = len;
      for (int x = 0; x < len; x++) {
        int tlx = 0, trx = rw;
        uint32_t bits = 0;
        for (int d = 0; d < 8; d++) {
          const int cx = (trx + tlx + 1) >> 1;
          const bool right = x >= cx;
          if (right) tlx = cx; else trx = cx;
          bits = (bits << 1) | (right ? 1u : 0u);
        }
        xp.push_back((uint8_t)bits);
      }
    }
    if ((int)xp.size() > ex->cap_xpath) {
      set_error("image larger than the max_width the extractor was created with");
      return YDORB_E_CAPACITY;
    }
    YD_CUDA(cudaMemcpyAsync(ex->d_qt_xpath, xp.data(), xp.size(), cudaMemcpyHostToDevice, ex->stream));
  }
  if (!hg.cells.empty())
    YD_CUDA(cudaMemcpyAsync(ex->d_cells, hg.cells.data(), hg.cells.size() * sizeof(int4), cudaMemcpyHostToDevice, ex->stream));
  // fast2.cu: cell blocks and one tensor map per level whose box is the largest block's tile
  std::vector<FastBlock> blocks;
  std::vector<uint32_t> seams;
  rc = fast2_build_blocks(hg.g, hg.cells, &blocks, &seams, &ex->fb_max_w, &ex->fb_max_h);
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(ex->d_fast_seams, seams.data(), seams.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ex->stream));
  ex->n_fast_blocks = (int)blocks.size();
  if (!blocks.empty())
    YD_CUDA(cudaMemcpyAsync(ex->d_fast_blocks, blocks.data(), blocks.size() * sizeof(FastBlock), cudaMemcpyHostToDevice, ex->stream));
  {
    int box_w = 0, box_h = 0;
    fast2_box(ex->fb_max_w, ex->fb_max_h, &box_w, &box_h);
    for (int l = 0; l < hg.g.n_levels; l++) {
      const LevelGeom& lv = hg.g.lv[l];
      rc = tma_encode_u8_3d(&ex->tmap_level[l], ex->pyr_base + lv.offset, (uint64_t)lv.stride, (uint64_t)(lv.h + 2 * kPad),
                            (uint64_t)ex->max_batch, (uint64_t)lv.stride, (uint64_t)hg.g.pyr_bytes, (uint32_t)box_w,
                            (uint32_t)box_h);
      if (rc != YDORB_OK) return rc;
    }
    for (int l = 1; l < hg.g.n_levels; l++) {
      const LevelGeom& sv = hg.g.lv[l - 1];
      pyramid_box(hg.g, l, &box_w, &box_h);
      if (box_w == 0) continue;
      rc = tma_encode_u8_3d(&ex->tmap_pyr[l], ex->pyr_base + sv.offset, (uint64_t)sv.stride, (uint64_t)(sv.h + 2 * kPad),
                            (uint64_t)ex->max_batch, (uint64_t)sv.stride, (uint64_t)hg.g.pyr_bytes, (uint32_t)box_w,
                            (uint32_t)box_h);
      if (rc != YDORB_OK) return rc;
    }
    describe_box(&box_w, &box_h);
    for (int l = 0; l < hg.g.n_levels; l++) {
      const LevelGeom& lv = hg.g.lv[l];
      rc = tma_encode_u8_3d(&ex->tmap_desc[l], ex->pyr_base + lv.offset, (uint64_t)lv.stride, (uint64_t)(lv.h + 2 * kPad),
                            (uint64_t)ex->max_batch, (uint64_t)lv.stride, (uint64_t)hg.g.pyr_bytes, (uint32_t)box_w,
                            (uint32_t)box_h);
      if (rc != YDORB_OK) return rc;
    }
  }
  YD_CUDA(cudaStreamSynchronize(ex->stream));  // host vectors go out of scope
  ex->cur_w = w;
  ex->cur_h = h;
  return YDORB_OK;
}

static int check_status(YdExtractor* ex) {
  int st = 0;
  YD_CUDA(cudaMemcpyAsync(&st, ex->d_status, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
  YD_CUDA(cudaStreamSynchronize(ex->stream));
  if (st == 0) return YDORB_OK;
  YD_CUDA(cudaMemsetAsync(ex->d_status, 0, sizeof(int), ex->stream));
  if (st & kStatusRawOverflow) { set_error("FAST candidate buffer overflow (more than 1 corner per 8 px)"); return YDORB_E_CAPACITY; }
  if (st & kStatusSortDepth) { set_error("std::sort emulation stack overflow"); return YDORB_E_CAPACITY; }
  if (st & kStatusOutOverflow) { set_error("keypoint capacity too small"); return YDORB_E_CAPACITY; }
  return YDORB_OK;
}

// A handle whose per-frame buffers start at frame f0 and whose work goes to stream s: lets the stage launchers run
// on a sub-batch without knowing about it (every kernel indexes its buffers by blockIdx = frame within the launch).
static YdExtractor frame_view(const YdExtractor* ex, int f0, cudaStream_t s) {
  YdExtractor v = *ex;
  const FrameGeomDev& g = ex->geom;
  v.stream = s;
  v.d_pyr += (size_t)f0 * g.pyr_bytes;
  v.d_cell_count += (size_t)f0 * g.n_cells;
  v.d_cell_start += (size_t)f0 * g.n_cells;
  v.d_cell_rank += (size_t)f0 * (g.n_cells + 1);
  v.d_raw += (size_t)f0 * g.raw_cap;
  v.d_raw_counter += f0;
  for (int i = 0; i < 3; i++) { v.d_qt_keys[i] += (size_t)f0 * g.raw_cap; v.d_qt_vals[i] += (size_t)f0 * g.raw_cap; }
  v.d_qt_bitmap += (size_t)f0 * kMaxLevels * kQtMapWords;
  v.d_qt_fold += (size_t)f0 * kMaxLevels * 2 * kQtMapWords;
  v.d_qt_meta += (size_t)f0 * kMaxLevels * 8;
  v.d_level_ncand += (size_t)f0 * kMaxLevels;
  v.d_sel += (size_t)f0 * g.n_quota;
  v.d_sel_n += (size_t)f0 * kMaxLevels;
  return v;
}

// stages of frames [f0, f0+n) on stream s, kernel by kernel
static int run_stages_direct(const YdExtractor* ex, int f0, int n, cudaStream_t s, const uint8_t* d_images, int stride,
                             size_t frame_stride, YdKeyPoint* d_kps, uint8_t* d_desc, int capacity, int32_t* d_nout) {
  const YdExtractor v = frame_view(ex, f0, s);
  int rc;
  if ((rc = launch_pyramid(&v, d_images, stride, frame_stride, n, s)) != YDORB_OK) return rc;
  if (ex->geom.n_cells == 0 || ex->geom.n_quota == 0) {
    YD_CUDA(cudaMemsetAsync(d_nout, 0, sizeof(int32_t) * n, s));
    YD_CUDA(cudaMemsetAsync(v.d_sel_n, 0, sizeof(int) * kMaxLevels * n, s));
    YD_CUDA(cudaMemsetAsync(v.d_level_ncand, 0, sizeof(int) * kMaxLevels * n, s));
    return YDORB_OK;
  }
  if ((rc = launch_fast(&v, n, s)) != YDORB_OK) return rc;
  if ((rc = launch_quadtree(&v, n, s)) != YDORB_OK) return rc;
  return launch_describe(&v, n, d_kps, d_desc, capacity, d_nout, s);
}

static void drop_graphs(YdExtractor* ex) {
  for (YdGraphEntry& e : ex->graphs)
    if (e.exec) {
      cudaGraphExecDestroy(e.exec);
      e.exec = nullptr;
    }
}

// The ~20 launches of a sub-batch are captured into a CUDA graph the first time a (frames, buffers) combination is
// seen and replayed with ONE cudaGraphLaunch afterwards: the kernels' arguments are all known on the host (no stage reads
// anything back), the programmatic-dependent-launch edges are kept by the capture.  A caller that keeps handing in the
// same buffers (a SLAM front-end does; so does the chunked host entry point, whose buffers belong to the handle) pays
// the per-kernel launch cost once.  YDORB_GRAPHS=0 switches this off; a failed capture switches it off for the handle.
static int run_stages(YdExtractor* ex, int f0, int n, cudaStream_t s, const uint8_t* d_images, int stride,
                      size_t frame_stride, YdKeyPoint* d_kps, uint8_t* d_desc, int capacity, int32_t* d_nout) {
  static const bool env_off = getenv("YDORB_GRAPHS") && atoi(getenv("YDORB_GRAPHS")) == 0;
  if (env_off || ex->graphs_off)
    return run_stages_direct(ex, f0, n, s, d_images, stride, frame_stride, d_kps, d_desc, capacity, d_nout);
  const char* fk = getenv("YDORB_FAST_KERNEL");  // read per call by launch_fast: part of the key
  const int fast_choice = fk ? fk[0] : 0;
  YdGraphEntry* hit = nullptr;
  YdGraphEntry* victim = &ex->graphs[0];
  for (YdGraphEntry& e : ex->graphs) {
    if (e.exec && e.f0 == f0 && e.n == n && e.d_images == d_images && e.stride == stride && e.frame_stride == frame_stride &&
        e.d_kps == d_kps && e.d_desc == d_desc && e.capacity == capacity && e.d_nout == d_nout && e.fast_choice == fast_choice) {
      hit = &e;
      break;
    }
    if (!e.exec || (victim->exec && e.stamp < victim->stamp)) victim = &e;
  }
  if (hit) {
    hit->stamp = ++ex->graph_clock;
    YD_CUDA(cudaGraphLaunch(hit->exec, s));
    count_launches(hit->launches);
    return YDORB_OK;
  }
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &st) != cudaSuccess || st != cudaStreamCaptureStatusNone)  // the caller captures already
    return run_stages_direct(ex, f0, n, s, d_images, stride, frame_stride, d_kps, d_desc, capacity, d_nout);
  const long long l0 = t_launches;
  if (cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed) != cudaSuccess) {
    cudaGetLastError();
    ex->graphs_off = 1;
    return run_stages_direct(ex, f0, n, s, d_images, stride, frame_stride, d_kps, d_desc, capacity, d_nout);
  }
  const int rc = run_stages_direct(ex, f0, n, s, d_images, stride, frame_stride, d_kps, d_desc, capacity, d_nout);
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(s, &graph);
  cudaGraphExec_t exec = nullptr;
  if (rc == YDORB_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&exec, graph, 0);
  if (graph) cudaGraphDestroy(graph);
  if (rc != YDORB_OK || e != cudaSuccess || !exec) {
    cudaGetLastError();
    g_launches.fetch_sub(t_launches - l0);
    t_launches = l0;
    if (rc != YDORB_OK) return rc;
    ex->graphs_off = 1;  // this driver / stream cannot capture the chain: plain launches from now on
    return run_stages_direct(ex, f0, n, s, d_images, stride, frame_stride, d_kps, d_desc, capacity, d_nout);
  }
  if (victim->exec) cudaGraphExecDestroy(victim->exec);
  *victim = YdGraphEntry{f0, n, stride, capacity, fast_choice, frame_stride, d_images, d_kps, d_desc, d_nout, exec,
                         (int)(t_launches - l0), ++ex->graph_clock};
  YD_CUDA(cudaGraphLaunch(exec, s));  // (the capture counted the launches already)
  return YDORB_OK;
}

static int run_pipeline(YdExtractor* ex, const uint8_t* d_images, int w, int h, int stride, size_t frame_stride,
                        int batch, YdKeyPoint* d_kps, uint8_t* d_desc, int capacity, int32_t* d_nout) {
  int rc = set_size(ex, w, h);
  if (rc != YDORB_OK) return rc;
  ex->cur_batch = batch;
  cudaStream_t s = ex->stream;
  const bool prof = ex->profiling != 0;
  if (!prof) return run_stages(ex, 0, batch, s, d_images, stride, frame_stride, d_kps, d_desc, capacity, d_nout);
  if (prof) ex->ev = ex->ev_ring[ex->prof_calls++ % 64];
  if (prof) YD_CUDA(cudaEventRecord(ex->ev[0], s));
  if ((rc = launch_pyramid(ex, d_images, stride, frame_stride, batch, s)) != YDORB_OK) return rc;
  if (prof) YD_CUDA(cudaEventRecord(ex->ev[1], s));
  if (ex->geom.n_cells == 0 || ex->geom.n_quota == 0) {
    if (prof) for (int i = 2; i < 5; i++) YD_CUDA(cudaEventRecord(ex->ev[i], s));
    YD_CUDA(cudaMemsetAsync(d_nout, 0, sizeof(int32_t) * batch, s));
    YD_CUDA(cudaMemsetAsync(ex->d_sel_n, 0, sizeof(int) * kMaxLevels * batch, s));
    YD_CUDA(cudaMemsetAsync(ex->d_level_ncand, 0, sizeof(int) * kMaxLevels * batch, s));
    return YDORB_OK;
  }
  if ((rc = launch_fast(ex, batch, s)) != YDORB_OK) return rc;
  if (prof) YD_CUDA(cudaEventRecord(ex->ev[2], s));
  if ((rc = launch_quadtree(ex, batch, s)) != YDORB_OK) return rc;
  if (prof) YD_CUDA(cudaEventRecord(ex->ev[3], s));
  if ((rc = launch_describe(ex, batch, d_kps, d_desc, capacity, d_nout, s)) != YDORB_OK) return rc;
  if (prof) YD_CUDA(cudaEventRecord(ex->ev[4], s));
  return YDORB_OK;
}

}  // namespace ydorb

using namespace ydorb;

extern "C" {

const char* ydorb_last_error(void) { return g_last_error.c_str(); }

long long ydorb_kernel_launches(void) { return g_launches.load(); }

int ydorb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

int ydorb_host_alloc(size_t bytes, void** out) {
  YD_REQUIRE(out, "null argument");
  *out = nullptr;
  YD_CUDA(cudaHostAlloc(out, std::max<size_t>(bytes, 1), cudaHostAllocPortable));
  return YDORB_OK;
}

void ydorb_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

int ydorb_extractor_create(const YdExtractorParams* params, int max_width, int max_height, int max_batch, int device,
                           YdExtractor** out) {
  YD_REQUIRE(params && out, "null argument");
  YD_REQUIRE(params->n_levels >= 1 && params->n_levels <= kMaxLevels, "n_levels out of range");
  YD_REQUIRE(params->scale_factor > 1.0f, "scale_factor must be > 1");
  YD_REQUIRE(params->n_features >= 0, "n_features must be >= 0");
  YD_REQUIRE(max_width > 0 && max_height > 0 && max_batch > 0, "bad capacity");
  int ndev = 0;
  YD_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) { set_error("no such CUDA device"); return YDORB_E_CUDA; }
  YD_CUDA(cudaSetDevice(device));
  YdExtractor* ex = new YdExtractor();
  memset(ex, 0, sizeof *ex);
  ex->params = *params;
  ex->device = device;
  ex->max_w = max_width;
  ex->max_h = max_height;
  ex->max_batch = max_batch;
  ctor_tables(ex);
  HostGeom hg;
  int rc = build_geom(ex, max_width, max_height, &hg);
  if (rc != YDORB_OK) { delete ex; return rc; }
  // capacities: geometry of the largest size, with slack so smaller sizes with different rounding still fit
  ex->cap_pyr_bytes = hg.g.pyr_bytes + 64 * 1024;
  ex->cap_cells = hg.g.n_cells + 64;
  ex->cap_raw = hg.g.raw_cap;
  ex->cap_tab = (int)hg.tab_ofs.size() + 64;
  const int nq = std::max(hg.g.n_quota, 1);
  ex->out_cap = nq;
  const size_t B = (size_t)max_batch;
  cudaError_t e = cudaStreamCreateWithFlags(&ex->own_stream, cudaStreamNonBlocking);
  ex->stream = ex->own_stream;
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ex->aux_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ex->ev_fork, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ex->ev_join, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ex->ev_block, cudaEventDisableTiming | cudaEventBlockingSync);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ex->in_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ex->out_stream, cudaStreamNonBlocking);
  for (int i = 0; i < 16 && e == cudaSuccess; i++) {
    e = cudaEventCreateWithFlags(&ex->ev_in[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ex->ev_done[i], cudaEventDisableTiming);
  }
  for (int i = 0; i < 64 * 5 && e == cudaSuccess; i++) e = cudaEventCreate(&ex->ev_ring[i / 5][i % 5]);
  ex->ev = ex->ev_ring[0];
  auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, std::max<size_t>(bytes, 16)); };
  alloc((void**)&ex->d_geom, sizeof(FrameGeomDev));
  alloc((void**)&ex->d_pyr, B * ex->cap_pyr_bytes);
  alloc((void**)&ex->d_tab_ofs, sizeof(int) * ex->cap_tab);
  alloc((void**)&ex->d_tab_a, sizeof(short) * 2 * ex->cap_tab);
  alloc((void**)&ex->d_cells, sizeof(int4) * ex->cap_cells);
  alloc((void**)&ex->d_fast_blocks, sizeof(FastBlock) * ex->cap_cells);
  alloc((void**)&ex->d_fast_seams, sizeof(uint32_t) * kMaxLevels * 512);
  alloc((void**)&ex->d_cell_count, sizeof(int) * B * ex->cap_cells);
  alloc((void**)&ex->d_cell_start, sizeof(int) * B * ex->cap_cells);
  alloc((void**)&ex->d_cell_rank, sizeof(int) * B * (ex->cap_cells + 1));
  alloc((void**)&ex->d_raw, sizeof(uint32_t) * B * ex->cap_raw);
  alloc((void**)&ex->d_raw_counter, sizeof(int) * B);
  for (int i = 0; i < 3; i++) {
    alloc((void**)&ex->d_qt_keys[i], sizeof(uint32_t) * B * ex->cap_raw);
    alloc((void**)&ex->d_qt_vals[i], sizeof(uint32_t) * B * ex->cap_raw);
  }
  ex->cap_xpath = (ex->max_w + 64) * kMaxLevels;
  alloc((void**)&ex->d_qt_xpath, (size_t)ex->cap_xpath);
  alloc((void**)&ex->d_qt_bitmap, sizeof(uint32_t) * B * kMaxLevels * kQtMapWords);
  alloc((void**)&ex->d_qt_fold, sizeof(uint32_t) * B * kMaxLevels * 2 * kQtMapWords);
  alloc((void**)&ex->d_qt_meta, sizeof(int) * B * kMaxLevels * 8);
  if (e == cudaSuccess) e = cudaMemset(ex->d_qt_bitmap, 0, sizeof(uint32_t) * B * kMaxLevels * kQtMapWords);
  alloc((void**)&ex->d_level_ncand, sizeof(int) * B * kMaxLevels);
  alloc((void**)&ex->d_sel, sizeof(int4) * B * nq);
  alloc((void**)&ex->d_sel_n, sizeof(int) * B * kMaxLevels);
  alloc((void**)&ex->d_status, sizeof(int));
  alloc((void**)&ex->d_images, B * (size_t)max_width * max_height);
  alloc((void**)&ex->d_kps, sizeof(YdKeyPoint) * B * nq);
  alloc((void**)&ex->d_desc, 32 * B * nq);
  alloc((void**)&ex->d_nout, sizeof(int32_t) * B);
  if (e == cudaSuccess) e = cudaMemset(ex->d_status, 0, sizeof(int));
  // the memsets above ran on the legacy stream; every stream of the handle is non-blocking, so order them here
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  ex->pyr_base = ex->d_pyr;
  if (e != cudaSuccess) {
    int r = cuda_fail(e, "extractor allocation", __FILE__, __LINE__);
    ydorb_extractor_destroy(ex);
    return r;
  }
  ex->cur_w = ex->cur_h = -1;
  *out = ex;
  return YDORB_OK;
}

void ydorb_extractor_destroy(YdExtractor* ex) {
  if (!ex) return;
  cudaSetDevice(ex->device);
  void* ptrs[] = {ex->d_geom, ex->d_pyr, ex->d_tab_ofs, ex->d_tab_a, ex->d_cells, ex->d_fast_blocks, ex->d_fast_seams, ex->d_cell_count, ex->d_cell_start,
                  ex->d_cell_rank, ex->d_raw, ex->d_raw_counter, ex->d_qt_keys[0], ex->d_qt_keys[1], ex->d_qt_keys[2],
                  ex->d_qt_vals[0], ex->d_qt_vals[1], ex->d_qt_vals[2], ex->d_qt_xpath, ex->d_qt_bitmap, ex->d_qt_fold, ex->d_qt_meta, ex->d_level_ncand, ex->d_sel, ex->d_sel_n,
                  ex->d_status, ex->d_images, ex->d_kps, ex->d_desc, ex->d_nout};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  for (int i = 0; i < 64 * 5; i++)
    if (ex->ev_ring[i / 5][i % 5]) cudaEventDestroy(ex->ev_ring[i / 5][i % 5]);
  drop_graphs(ex);
  if (ex->ev_fork) cudaEventDestroy(ex->ev_fork);
  if (ex->ev_join) cudaEventDestroy(ex->ev_join);
  if (ex->ev_block) cudaEventDestroy(ex->ev_block);
  if (ex->aux_stream) cudaStreamDestroy(ex->aux_stream);
  if (ex->in_stream) cudaStreamDestroy(ex->in_stream);
  if (ex->out_stream) cudaStreamDestroy(ex->out_stream);
  for (int i = 0; i < 16; i++) {
    if (ex->ev_in[i]) cudaEventDestroy(ex->ev_in[i]);
    if (ex->ev_done[i]) cudaEventDestroy(ex->ev_done[i]);
  }
  if (ex->own_stream) cudaStreamDestroy(ex->own_stream);
  delete ex;
}

int ydorb_extractor_levels(const YdExtractor* ex) { return ex ? ex->params.n_levels : 0; }

int ydorb_extractor_scale_tables(const YdExtractor* ex, float* sf, float* inv_sf, float* sf2, float* inv_sf2) {
  YD_REQUIRE(ex, "null extractor");
  const size_t n = sizeof(float) * ex->params.n_levels;
  if (sf) memcpy(sf, ex->sf, n);
  if (inv_sf) memcpy(inv_sf, ex->inv_sf, n);
  if (sf2) memcpy(sf2, ex->sf2, n);
  if (inv_sf2) memcpy(inv_sf2, ex->inv_sf2, n);
  return YDORB_OK;
}

int ydorb_extractor_quotas(const YdExtractor* ex, int32_t* per_level) {
  YD_REQUIRE(ex && per_level, "null argument");
  for (int i = 0; i < ex->params.n_levels; i++) per_level[i] = ex->quota[i];
  return YDORB_OK;
}

void* ydorb_extractor_stream(const YdExtractor* ex) { return ex ? (void*)ex->stream : nullptr; }

int ydorb_extractor_set_stream(YdExtractor* ex, void* cuda_stream) {
  YD_REQUIRE(ex, "null extractor");
  YD_CUDA(cudaSetDevice(ex->device));
  YD_CUDA(cudaStreamSynchronize(ex->stream));
  ex->stream = cuda_stream ? (cudaStream_t)cuda_stream : ex->own_stream;
  return YDORB_OK;
}

int ydorb_extractor_set_profiling(YdExtractor* ex, int enable) {
  YD_REQUIRE(ex, "null extractor");
  ex->profiling = enable ? 1 : 0;
  ex->prof_calls = 0;
  return YDORB_OK;
}

int ydorb_extractor_stage_ms(YdExtractor* ex, float* ms4) {
  YD_REQUIRE(ex && ms4, "null argument");
  if (!ex->profiling || ex->prof_calls <= 0) { set_error("profiling not enabled or no extract call since"); return YDORB_E_STATE; }
  YD_CUDA(cudaSetDevice(ex->device));
  YD_CUDA(cudaEventSynchronize(ex->ev[4]));
  // mean over the calls recorded since profiling was switched on (at most the last 64)
  const int calls = (int)std::min<long long>(ex->prof_calls, 64);
  for (int i = 0; i < 4; i++) ms4[i] = 0.f;
  for (int c = 0; c < calls; c++)
    for (int i = 0; i < 4; i++) {
      float t = 0.f;
      YD_CUDA(cudaEventElapsedTime(&t, ex->ev_ring[c][i], ex->ev_ring[c][i + 1]));
      ms4[i] += t / (float)calls;
    }
  return YDORB_OK;
}

int ydorb_extractor_sync(YdExtractor* ex) {
  YD_REQUIRE(ex, "null extractor");
  YD_CUDA(cudaSetDevice(ex->device));
  YD_CUDA(cudaStreamSynchronize(ex->stream));
  return check_status(ex);
}

int ydorb_extract_batch_device(YdExtractor* ex, const uint8_t* d_images, int width, int height, int stride,
                               size_t frame_stride, int batch, YdKeyPoint* d_keypoints, uint8_t* d_descriptors,
                               int capacity, int32_t* d_n_out) {
  YD_REQUIRE(ex && d_images && d_keypoints && d_descriptors && d_n_out, "null argument");
  YD_REQUIRE(batch >= 1 && batch <= ex->max_batch, "batch exceeds max_batch");
  YD_REQUIRE(width > 0 && height > 0 && stride >= width && capacity > 0, "bad image geometry");
  YD_CUDA(cudaSetDevice(ex->device));
  int split = (batch >= 64) ? 2 : 1;
  if (const char* e = getenv("YDORB_DEVICE_SPLIT")) split = std::max(1, std::min(atoi(e), std::min(batch, 16)));  // tuning knob
  if (ex->profiling || split == 1)
    return run_pipeline(ex, d_images, width, height, stride, frame_stride, batch, d_keypoints, d_descriptors, capacity,
                        d_n_out);
  // Sub-batches alternate between the handle's stream and its second stream: the latency-bound stages of one
  // (quadtree, small pyramid levels) overlap the issue-bound stages of the other (FAST, describe).  Per-stage events
  // make no sense here, so profiling mode keeps the serial path above.
  int rc = set_size(ex, width, height);
  if (rc != YDORB_OK) return rc;
  ex->cur_batch = batch;
  cudaStream_t s = ex->stream;
  YD_CUDA(cudaEventRecord(ex->ev_fork, s));
  YD_CUDA(cudaStreamWaitEvent(ex->aux_stream, ex->ev_fork, 0));
  const int chunk = (batch + split - 1) / split;
  int k = 0;
  for (int f0 = 0; f0 < batch; f0 += chunk, k++) {
    const int n = std::min(chunk, batch - f0);
    rc = run_stages(ex, f0, n, (k & 1) ? ex->aux_stream : s, d_images + (size_t)f0 * frame_stride, stride, frame_stride,
                    d_keypoints + (size_t)f0 * capacity, d_descriptors + (size_t)f0 * capacity * 32, capacity, d_n_out + f0);
    if (rc != YDORB_OK) return rc;
  }
  YD_CUDA(cudaEventRecord(ex->ev_join, ex->aux_stream));
  YD_CUDA(cudaStreamWaitEvent(s, ex->ev_join, 0));
  return YDORB_OK;
}

int ydorb_extract_batch(YdExtractor* ex, const uint8_t* images, int width, int height, int stride, size_t frame_stride,
                        int batch, YdKeyPoint* keypoints, uint8_t* descriptors, int capacity, int32_t* n_out) {
  YD_REQUIRE(ex && n_out, "null argument");
  YD_REQUIRE(batch >= 1 && batch <= ex->max_batch, "batch exceeds max_batch");
  if (!images || width <= 0 || height <= 0) {  // empty image: silent return (src/orbExtractor.cpp:356-358)
    for (int b = 0; b < batch; b++) n_out[b] = 0;
    return YDORB_OK;
  }
  YD_REQUIRE(keypoints && descriptors && capacity > 0, "null output");
  YD_REQUIRE(stride >= width, "stride smaller than width");
  YD_REQUIRE(width <= ex->max_w && height <= ex->max_h, "image larger than max_width/max_height");
  YD_CUDA(cudaSetDevice(ex->device));
  cudaStream_t s = ex->stream;
  const size_t dpitch = (size_t)width;
  const int dcap = ex->out_cap;
  int rc = set_size(ex, width, height);
  if (rc != YDORB_OK) return rc;
  ex->cur_batch = batch;
  const bool bulk_in = stride == width && (batch == 1 || frame_stride == (size_t)width * height);
  auto upload = [&](int f0, int n, cudaStream_t st) -> int {
    if (bulk_in) {
      YD_CUDA(cudaMemcpyAsync(ex->d_images + (size_t)f0 * width * height, images + (size_t)f0 * width * height,
                              (size_t)n * width * height, cudaMemcpyHostToDevice, st));
    } else {
      for (int b = f0; b < f0 + n; b++)
        YD_CUDA(cudaMemcpy2DAsync(ex->d_images + (size_t)b * width * height, dpitch, images + (size_t)b * frame_stride,
                                  (size_t)stride, (size_t)width, (size_t)height, cudaMemcpyHostToDevice, st));
    }
    return YDORB_OK;
  };
  if (!ex->profiling && capacity == dcap && batch >= 8) {
    // Chunked, four streams: a copy-in stream uploads sub-batch after sub-batch without waiting for any kernel, two
    // compute streams take the sub-batches alternately (one's small tail kernels overlap the other's wide ones), a
    // copy-out stream returns each sub-batch's results as soon as they exist.  Outputs have the caller's layout
    // (capacity == quota sum), so results are copied back in bulk without knowing the counts.
    // Sub-batch plan (measured on a B200, tools/chunk_probe.py and tools/e2e_threads_probe.py): sub-batches of 32-64
    // frames keep the grids wide enough while the first upload stays short; a handle that is one of several driven
    // from concurrent host threads (small batch each) does best with two sub-batches.
    // YDORB_HOST_PLAN="16,48,64,..." (frames per sub-batch, last value repeats) and YDORB_HOST_CHUNKS=n (n equal
    // sub-batches) are tuning knobs.
    int plan[16], nplan = 0;
    if (const char* e = getenv("YDORB_HOST_PLAN")) {
      for (const char* q = e; *q && nplan < 16;) {
        const int v = atoi(q);
        if (v > 0) plan[nplan++] = v;
        while (*q && *q != ',') q++;
        if (*q == ',') q++;
      }
    }
    if (nplan == 0) {
      int nchunks = batch >= 192 ? 6 : 2;
      if (const char* e = getenv("YDORB_HOST_CHUNKS")) nchunks = std::min(16, std::max(1, atoi(e)));
      plan[nplan++] = (batch + nchunks - 1) / nchunks;
    }
    static const bool trace = getenv("YDORB_HOST_TRACE") != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    YD_CUDA(cudaEventRecord(ex->ev_fork, s));
    YD_CUDA(cudaStreamWaitEvent(ex->aux_stream, ex->ev_fork, 0));
    YD_CUDA(cudaStreamWaitEvent(ex->in_stream, ex->ev_fork, 0));
    YD_CUDA(cudaStreamWaitEvent(ex->out_stream, ex->ev_fork, 0));
    int k = 0;
    // an error inside the loop stops the enqueueing but still joins the four streams below: D2H copies already queued
    // must not write into the caller's buffers after this call has returned
    auto enqueue = [&](int f0, int n) -> int {
      cudaStream_t st = (k & 1) ? ex->aux_stream : s;
      int r = upload(f0, n, ex->in_stream);
      if (r != YDORB_OK) return r;
      YD_CUDA(cudaEventRecord(ex->ev_in[k], ex->in_stream));
      YD_CUDA(cudaStreamWaitEvent(st, ex->ev_in[k], 0));
      r = run_stages(ex, f0, n, st, ex->d_images + (size_t)f0 * width * height, width, (size_t)width * height,
                     ex->d_kps + (size_t)f0 * dcap, ex->d_desc + (size_t)f0 * dcap * 32, dcap, ex->d_nout + f0);
      if (r != YDORB_OK) return r;
      if (f0 + n >= batch && wait_sleeps()) YD_CUDA(cudaEventRecord(ex->ev_block, st));  // last sub-batch's kernels end here
      YD_CUDA(cudaEventRecord(ex->ev_done[k], st));
      YD_CUDA(cudaStreamWaitEvent(ex->out_stream, ex->ev_done[k], 0));
      YD_CUDA(cudaMemcpyAsync(n_out + f0, ex->d_nout + f0, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, ex->out_stream));
      YD_CUDA(cudaMemcpyAsync(keypoints + (size_t)f0 * dcap, ex->d_kps + (size_t)f0 * dcap, sizeof(YdKeyPoint) * (size_t)n * dcap,
                              cudaMemcpyDeviceToHost, ex->out_stream));
      YD_CUDA(cudaMemcpyAsync(descriptors + (size_t)f0 * dcap * 32, ex->d_desc + (size_t)f0 * dcap * 32, (size_t)32 * n * dcap,
                              cudaMemcpyDeviceToHost, ex->out_stream));
      return YDORB_OK;
    };
    int err = YDORB_OK;
    for (int f0 = 0; f0 < batch && err == YDORB_OK; k++) {
      int n = std::min(plan[std::min(k, nplan - 1)], batch - f0);
      if (k == 15) n = batch - f0;  // event slots: at most 16 sub-batches
      err = enqueue(f0, n);
      f0 += n;
    }
    // join: the copy-out stream waits for the copy-in stream and both compute streams, the handle's stream for it
    cudaEventRecord(ex->ev_join, ex->in_stream);
    cudaStreamWaitEvent(ex->out_stream, ex->ev_join, 0);
    cudaEventRecord(ex->ev_join, ex->aux_stream);
    cudaStreamWaitEvent(ex->out_stream, ex->ev_join, 0);
    cudaEventRecord(ex->ev_join, ex->out_stream);
    cudaStreamWaitEvent(s, ex->ev_join, 0);
    const auto t_issued = std::chrono::steady_clock::now();
    const int wrc = wait_stream(s, ex->ev_block, err == YDORB_OK);
    if (err != YDORB_OK) {
      // a quadtree launch that failed half-way may have left the occupancy bitmap dirty for the next call
      cudaMemsetAsync(ex->d_qt_bitmap, 0, sizeof(uint32_t) * (size_t)ex->max_batch * kMaxLevels * kQtMapWords, s);
      cudaStreamSynchronize(s);
      return err;
    }
    if (wrc != YDORB_OK) return wrc;
    if (trace) {
      const auto t_end = std::chrono::steady_clock::now();
      fprintf(stderr, "[ydorb] extract_batch: %d sub-batches, issue %.3f ms, total %.3f ms\n", k,
              std::chrono::duration<double, std::milli>(t_issued - t_begin).count(),
              std::chrono::duration<double, std::milli>(t_end - t_begin).count());
    }
    return check_status(ex);
  }
  if ((rc = upload(0, batch, s)) != YDORB_OK) return rc;
  rc = run_pipeline(ex, ex->d_images, width, height, width, (size_t)width * height, batch, ex->d_kps, ex->d_desc,
                    dcap, ex->d_nout);
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(n_out, ex->d_nout, sizeof(int32_t) * batch, cudaMemcpyDeviceToHost, s));
  if (capacity == dcap) {  // same layout on both sides: two bulk copies instead of two per frame, one wait for everything
    YD_CUDA(cudaMemcpyAsync(keypoints, ex->d_kps, sizeof(YdKeyPoint) * (size_t)batch * dcap, cudaMemcpyDeviceToHost, s));
    YD_CUDA(cudaMemcpyAsync(descriptors, ex->d_desc, (size_t)32 * batch * dcap, cudaMemcpyDeviceToHost, s));
    YD_CUDA(cudaStreamSynchronize(s));
    return check_status(ex);
  }
  YD_CUDA(cudaStreamSynchronize(s));
  rc = check_status(ex);
  if (rc != YDORB_OK) return rc;
  int ret = YDORB_OK;
  for (int b = 0; b < batch; b++) {
    int n = n_out[b];
    if (n > capacity) { n = capacity; ret = YDORB_E_CAPACITY; set_error("keypoint capacity too small"); }
    n_out[b] = n;
    if (n == 0) continue;
    YD_CUDA(cudaMemcpyAsync(keypoints + (size_t)b * capacity, ex->d_kps + (size_t)b * dcap, sizeof(YdKeyPoint) * n,
                            cudaMemcpyDeviceToHost, s));
    YD_CUDA(cudaMemcpyAsync(descriptors + (size_t)b * capacity * 32, ex->d_desc + (size_t)b * dcap * 32, (size_t)32 * n,
                            cudaMemcpyDeviceToHost, s));
  }
  YD_CUDA(cudaStreamSynchronize(s));
  return ret;
}

int ydorb_extract(YdExtractor* ex, const uint8_t* image, int width, int height, int stride, YdKeyPoint* keypoints,
                  uint8_t* descriptors, int capacity, int32_t* n_out) {
  return ydorb_extract_batch(ex, image, width, height, stride, 0, 1, keypoints, descriptors, capacity, n_out);
}

int ydorb_pyramid_level(const YdExtractor* ex, int frame, int level, const uint8_t** d_level, int* width, int* height,
                        int* stride) {
  YD_REQUIRE(ex && d_level, "null argument");
  if (ex->cur_w <= 0) { set_error("no extract call yet"); return YDORB_E_STATE; }
  YD_REQUIRE(frame >= 0 && frame < ex->cur_batch && level >= 0 && level < ex->geom.n_levels, "frame/level out of range");
  const LevelGeom& lv = ex->geom.lv[level];
  *d_level = ex->d_pyr + (size_t)frame * ex->geom.pyr_bytes + level_origin(lv);
  if (width) *width = lv.w;
  if (height) *height = lv.h;
  if (stride) *stride = lv.stride;
  return YDORB_OK;
}

int ydorb_pyramid_download(YdExtractor* ex, int frame, int level, int padded, uint8_t* dst, int dst_stride) {
  const uint8_t* p;
  int w, h, st;
  int rc = ydorb_pyramid_level(ex, frame, level, &p, &w, &h, &st);
  if (rc != YDORB_OK) return rc;
  YD_REQUIRE(dst, "null dst");
  YD_CUDA(cudaSetDevice(ex->device));
  if (padded) { p -= (size_t)kPad * st + kPad; w += 2 * kPad; h += 2 * kPad; }
  YD_REQUIRE(dst_stride >= w, "dst_stride too small");
  YD_CUDA(cudaStreamSynchronize(ex->stream));
  YD_CUDA(cudaMemcpy2D(dst, (size_t)dst_stride, p, (size_t)st, (size_t)w, (size_t)h, cudaMemcpyDeviceToHost));
  return YDORB_OK;
}

int ydorb_debug_candidates(YdExtractor* ex, int frame, int level, int32_t* xy, float* resp, int capacity, int32_t* n_out) {
  YD_REQUIRE(ex && n_out, "null argument");
  if (ex->cur_w <= 0) { set_error("no extract call yet"); return YDORB_E_STATE; }
  YD_REQUIRE(frame >= 0 && frame < ex->cur_batch && level >= 0 && level < ex->geom.n_levels, "frame/level out of range");
  YD_CUDA(cudaSetDevice(ex->device));
  YD_CUDA(cudaStreamSynchronize(ex->stream));
  const FrameGeomDev& g = ex->geom;
  const LevelGeom& lv = g.lv[level];
  const int nc = lv.cell_cols * lv.cell_rows;
  std::vector<int> cnt(std::max(nc, 1)), st(std::max(nc, 1));
  if (nc > 0) {
    YD_CUDA(cudaMemcpy(cnt.data(), ex->d_cell_count + (size_t)frame * g.n_cells + lv.cell_base, sizeof(int) * nc, cudaMemcpyDeviceToHost));
    YD_CUDA(cudaMemcpy(st.data(), ex->d_cell_start + (size_t)frame * g.n_cells + lv.cell_base, sizeof(int) * nc, cudaMemcpyDeviceToHost));
  }
  int n = 0;
  std::vector<uint32_t> buf;
  for (int c = 0; c < nc; c++) {
    if (cnt[c] == 0) continue;
    buf.resize(cnt[c]);
    YD_CUDA(cudaMemcpy(buf.data(), ex->d_raw + (size_t)frame * g.raw_cap + st[c], sizeof(uint32_t) * cnt[c], cudaMemcpyDeviceToHost));
    for (int e = 0; e < cnt[c]; e++) {
      if (n < capacity) {
        if (xy) { xy[2 * n] = (int)((buf[e] >> 8) & 4095u); xy[2 * n + 1] = (int)(buf[e] >> 20); }
        if (resp) resp[n] = (float)(buf[e] & 255u);
      }
      n++;
    }
  }
  *n_out = n;
  return YDORB_OK;
}

}  // extern "C"
