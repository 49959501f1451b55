// E5 computeOrientation (src/orbExtractor.cpp:400-421), E6 GaussianBlur 7x7 sigma 2 (:385-386),
// E7 computeDescriptors (:422-454) and the output assembly of extractAndCompute (:372-398).
//
// One warp per selected keypoint.  The 43 x 53 px neighbourhood of the keypoint in the padded, UNBLURRED level (rows
// y-21..y+21, columns x-26..x+26) is fetched by ONE TMA box load (cp.async.bulk.tensor.3d, 80 bytes x 43 rows starting
// on the 16-byte boundary at or left of column x-26; an mbarrier per buffer, double buffered so that the next
// keypoint's box is in flight while this one computes) and everything is derived from it:
//   * IC angle: integer moments over the reference's footprint (half-width table of SURVEY F4: rows +-1..+-10 only
//     their centre column, rows +-11..+-15 are 53/51/45/39/31 px wide), lane-parallel + shuffle reduction, then
//     cv::fastAtan2's polynomial with individually rounded fp32 operations;
//   * blur: OpenCV's 8.8 fixed point 7-tap kernel [18 34 48 56 48 34 18]; the horizontal pass (exact in 16 bits) is
//     done once for the 43 x 37 window, the vertical pass only at the 512 sampled positions, rounded once
//     ((v + 32768) >> 16) -- bit-identical to blurring the whole level, without writing a blurred level to HBM;
//   * rBRIEF: lane j produces byte j (8 comparisons), coordinates rotated in fp32 without contraction.
#include <cstdlib>
#include "kernels.cuh"
#include "tma.cuh"

namespace ydorb {

constexpr int kDescWarps = 4;
constexpr int kPatchRows = 43;    // y-21 .. y+21  (18 px pattern reach + 3 px blur reach)
constexpr int kPatchStride = 80;  // bytes per staged row = TMA box width: 53 columns + up to 15 bytes of alignment (a box
                                  // starts on a 16-byte boundary).  20 words: 8 consecutive rows x 4 consecutive words hit
                                  // 32 different banks, which is how the horizontal blur walks it
constexpr int kPatchBytes = (kPatchRows * kPatchStride + 127) / 128 * 128;  // TMA destinations are 128-byte aligned
constexpr int kHbGroups = 11;     // horizontal-blur outputs per row in groups of 4: patch columns 7 .. 50
constexpr int kHbCols = 4 * kHbGroups;    // horizontal-blur columns kept
constexpr int kHbColStride = 44;          // u16 per COLUMN: the buffer is stored transposed, hb[col][row], rows 0..42 (+1 pad)
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

struct DescMaps {
  CUtensorMap lv[kMaxLevels];  // u8 [frame][row][byte] per level, box = kPatchStride x kPatchRows
};

struct DescLevel {      // per-level facts of this frame, built once per CTA
  int quota_base, cnt, out_base, stride;
  long long origin;     // byte offset of level pixel (0,0) in the frame's pyramid block
  float sf, size;
};

struct DescKp {
  int valid, level, oi, a;
  int4 kp;              // x, y (level coords), response, level
  int bx, by;           // TMA box origin in the padded level (byte column on a 16-byte boundary, row)
};

// qb: lane l holds level l's first slot (INT_MAX for l = 0 and l >= n_levels); kp: the slot's quadtree record, loaded by
// the caller for all of the warp's slots at once.  Everything here is uniform over the warp.
__device__ __forceinline__ DescKp desc_setup(const DescLevel* __restrict__ lvl, int qb, int n_quota, int slot, int4 kp,
                                             int capacity) {
  DescKp d;
  d.valid = 0;
  if (slot >= n_quota) return d;
  const int level = __popc(__ballot_sync(0xffffffffu, slot >= qb));  // quota bases ascend
  const DescLevel L = lvl[level];
  const int j = slot - L.quota_base;
  if (j >= L.cnt) return d;
  d.oi = L.out_base + j;
  if (d.oi >= capacity) return d;
  d.level = level;
  d.kp = kp;
  // the 19 px reflect-101 pad makes every byte the kernel reads a pixel of the level: keypoints are >= 16 px from the
  // edge; the box may reach past the row / buffer end, TMA zero-fills there and nothing reads it
  const int px = kPad + d.kp.x - 26;  // >= 9
  d.a = px & 15;
  d.bx = px - d.a;
  d.by = kPad + d.kp.y - 21;
  d.valid = 1;
  return d;
}

// One elected lane arms the buffer's mbarrier with the box size and issues the TMA load; everybody waits on the barrier.
__device__ __forceinline__ void desc_prefetch(const DescKp& d, int lane, uint8_t* __restrict__ patch, uint64_t* bar,
                                              const CUtensorMap* map, int frame) {
  if (lane == 0) {
    mbar_expect_tx(bar, (uint32_t)(kPatchRows * kPatchStride));
    tma_load_3d(patch, map, bar, d.bx, d.by, frame);
  }
}

// nxt / bar / nxt_map / tma_frame: the next key point of the warp, whose box is requested into the SAME patch buffer as
// soon as this key point's last read of it (the horizontal blur) is over
__device__ void desc_compute(const DescKp& d, const DescLevel& lv, int frame, int lane, uint8_t* __restrict__ patch,
                             uint16_t* __restrict__ hb, const uint32_t* __restrict__ s_pat, YdKeyPoint* __restrict__ kps,
                             uint8_t* __restrict__ desc, int capacity, const DescKp& nxt, uint64_t* bar,
                             const CUtensorMap* nxt_map, int tma_frame);

// KPW = consecutive keypoint slots a warp works through (8 for batches that fill the GPU; fewer spreads a small batch)
template <int KPW>
__global__ void __launch_bounds__(kDescWarps * 32, 7)
k_describe(const __grid_constant__ DescMaps maps, const FrameGeomDev* __restrict__ g, int frame0,
           const int4* __restrict__ sel, const int* __restrict__ sel_n, YdKeyPoint* __restrict__ kps,
           uint8_t* __restrict__ desc, int capacity, int32_t* __restrict__ n_out, int* __restrict__ status) {
  __shared__ __align__(128) uint8_t s_patch[kDescWarps][kPatchBytes];  // TMA destination, one per warp
  __shared__ __align__(8) uint64_t s_bar[kDescWarps];
  __shared__ __align__(16) uint16_t s_hb[kDescWarps][kHbCols * kHbColStride];
  __shared__ uint32_t s_pat[256];  // [bit b][lane]: the two pattern points (x0,y0,x1,y1 as s8) of bit b of byte `lane`
  __shared__ DescLevel s_lvl[kMaxLevels];
  const int frame = blockIdx.y;
  const int n_levels = g->n_levels, n_quota = g->n_quota;
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    const int b = i >> 5, l = i & 31;
    s_pat[i] = reinterpret_cast<const uint32_t*>(c_pattern)[l * 8 + b];
  }
  if ((threadIdx.x & 31) == 0) {
    mbar_init(&s_bar[threadIdx.x >> 5], 1);
  }
  pdl_wait();  // above: constants and geometry only; below: what the quadtree wrote
  // the warp's quadtree records first: their address depends on nothing that is loaded below
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot0 = (blockIdx.x * kDescWarps + warp) * KPW;
  const int4* sel_frame = sel + (size_t)frame * n_quota;
  int4 mine = make_int4(0, 0, 0, 0);  // lane i: the record of slot0 + i (slots past a level's count hold stale records, unused)
  if (lane < KPW && slot0 + lane < n_quota) mine = sel_frame[slot0 + lane];
  if (threadIdx.x < n_levels) {
    const int l = threadIdx.x;
    const int* cnt = sel_n + (size_t)frame * kMaxLevels;
    DescLevel L;
    L.quota_base = g->lv[l].quota_base;
    L.cnt = cnt[l];
    L.out_base = 0;
    for (int k = 0; k < l; k++) L.out_base += cnt[k];
    L.stride = g->lv[l].stride;
    L.origin = level_origin(g->lv[l]);
    L.sf = g->lv[l].sf;
    L.size = g->lv[l].size;
    s_lvl[l] = L;
    if (l == n_levels - 1 && blockIdx.x == 0) {
      const int total = L.out_base + L.cnt;
      n_out[frame] = min(total, capacity);
      if (total > capacity) atomicOr(status, kStatusOutOverflow);
    }
  }
  __syncthreads();
  uint16_t* hb = s_hb[warp];
  // pipeline over the warp's key points with ONE patch buffer per warp (7.3 KB of shared memory per warp instead of 10.8:
  // more resident warps): a key point reads the patch for its angle and horizontal blur only, so the next one's box is
  // requested right after the blur and lands while the descriptor bits are sampled from the blurred buffer.  Every
  // request completes one phase of the warp's mbarrier; `phase` counts them (key points without work request nothing).
  const int qb = (lane >= 1 && lane < n_levels) ? s_lvl[lane].quota_base : 0x7FFFFFFF;
  auto record = [&](int k) {
    return make_int4(__shfl_sync(0xffffffffu, mine.x, k), __shfl_sync(0xffffffffu, mine.y, k), __shfl_sync(0xffffffffu, mine.z, k),
                     __shfl_sync(0xffffffffu, mine.w, k));
  };
  uint32_t phase = 0;
  DescKp cur = desc_setup(s_lvl, qb, n_quota, slot0, record(0), capacity);
  if (cur.valid) desc_prefetch(cur, lane, s_patch[warp], &s_bar[warp], &maps.lv[cur.level], frame0 + frame);
  for (int k = 0; k < KPW; k++) {
    DescKp nxt;
    nxt.valid = 0;
    nxt.level = 0;
    if (k + 1 < KPW) nxt = desc_setup(s_lvl, qb, n_quota, slot0 + k + 1, record(k + 1), capacity);
    const CUtensorMap* nxt_map = &maps.lv[nxt.valid ? nxt.level : 0];
    if (cur.valid) {
      mbar_wait(&s_bar[warp], phase);
      phase ^= 1u;
      desc_compute(cur, s_lvl[cur.level], frame, lane, s_patch[warp], hb, s_pat, kps, desc, capacity, nxt, &s_bar[warp], nxt_map,
                   frame0 + frame);
    } else if (nxt.valid) {  // nothing in flight, nobody reads the buffer
      desc_prefetch(nxt, lane, s_patch[warp], &s_bar[warp], nxt_map, frame0 + frame);
    }
    __syncwarp();
    cur = nxt;
  }
}

// pixel (dx, dy) of the keypoint's neighbourhood sits at patch[(dy+21)*kPatchStride + dx+26+a], a = 0..15
__device__ void desc_compute(const DescKp& d, const DescLevel& lv, int frame, int lane, uint8_t* __restrict__ patch,
                             uint16_t* __restrict__ hb, const uint32_t* __restrict__ s_pat, YdKeyPoint* __restrict__ kps,
                             uint8_t* __restrict__ desc, int capacity, const DescKp& nxt, uint64_t* bar,
                             const CUtensorMap* nxt_map, int tma_frame) {
  const int a = d.a, oi = d.oi, level = d.level;
  const int4 kp = d.kp;
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
#pragma unroll
      for (int step = 0; step < 2; step++) {  // 2d + 1 <= 53 columns: two predicated steps, no loop control
        const int u = lane - d + 32 * step;
        if (u <= d) {
          const int pos = ctr[v * kPatchStride + u], neg = ctr[-v * kPatchStride + u];
          vs += pos - neg;
          m10 += u * (pos + neg);
        }
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
  // ---- E6 horizontal blur pass over rows y-21..y+21 (exact, <= 65280).  The whole words of the alignment offset
  // (a >> 2) move the window, the rest (a & 3) is applied when the samples are read.  A task = 4 adjacent outputs at
  // window columns c0..c0+3 with c0 = 7 + 4*grp, i.e. c0-3 word aligned: taps [18 34 48 56 | 48 34 18 0] as two DP4A.
  // A warp step covers 8 rows x 4 groups: with the 20-word row stride of the TMA box those 32 loads hit 32 different
  // banks, and so do the stores of the results, which go TRANSPOSED, hb[col][row], so that the 7 vertical taps of a
  // sample are 7 consecutive u16.
  {
    const uint32_t* pbase = reinterpret_cast<const uint32_t*>(patch) + 1 + (a >> 2);
    const int r8 = lane & 7, g4 = lane >> 3;
#pragma unroll
    for (int gb = 0; gb < kHbGroups; gb += 4) {  // fully unrolled: every load / store offset is an immediate
      const int grp = gb + g4;
#pragma unroll
      for (int rb = 0; rb < kPatchRows; rb += 8) {
        const int r = rb + r8;
        if (grp < kHbGroups && r < kPatchRows) {
          const uint32_t* pw = pbase + r * (kPatchStride / 4) + grp;
          const uint32_t A = pw[0], B = pw[1], C = pw[2];
          // the window of output k starts k bytes into (A, B, C): the TAPS are shifted, not the pixels -- 10 DP4A and no
          // funnel shifts (taps 18 34 48 56 48 34 18, little endian)
          uint16_t* o = hb + (4 * grp) * kHbColStride + r;
          o[0] = (uint16_t)__dp4a(A, 0x38302212u, __dp4a(B, 0x00122230u, 0u));
          o[kHbColStride] = (uint16_t)__dp4a(A, 0x30221200u, __dp4a(B, 0x12223038u, 0u));
          o[2 * kHbColStride] = (uint16_t)__dp4a(A, 0x22120000u, __dp4a(B, 0x22303830u, __dp4a(C, 0x00000012u, 0u)));
          o[3 * kHbColStride] = (uint16_t)__dp4a(A, 0x12000000u, __dp4a(B, 0x30383022u, __dp4a(C, 0x00001222u, 0u)));
        }
      }
    }
  }
  __syncwarp();
  // the patch is dead from here (rBRIEF samples the blurred buffer): the next key point's box lands under the sampling
  if (nxt.valid) desc_prefetch(nxt, lane, patch, bar, nxt_map, tma_frame);
  // ---- E7 rBRIEF
  const float a_rad = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.0));  // :425
  float cosA, sinB;
  glibc_sincosf(a_rad, sinB, cosA);                                                  // :426 (float overloads: sincosf)
  // 32-bit view of hb; a sample at (rr, cc) reads rows 18+rr .. 24+rr of column 19+a+cc: four consecutive words,
  // 4 x DP2A with the tap pairs (18,34)(48,56)(48,34)(18,0), moved by one slot when the first row is odd
  const uint32_t* hw32 = reinterpret_cast<const uint32_t*>(hb) + (26 + (a & 3) - kHbCol0) * (kHbColStride / 2);
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
      const int row0 = rr + 18;
      const uint32_t* h = hw32 + cc * (kHbColStride / 2) + (row0 >> 1);
      // an odd first row starts in the high half of w0: the TAPS move by one slot, not the data (18 34 48 56 48 34 18)
      const bool odd = (row0 & 1) != 0;
      const uint32_t t01 = odd ? 0x30221200u : 0x38302212u, t23 = odd ? 0x12223038u : 0x00122230u;
      const uint32_t w0 = h[0], w1 = h[1], w2 = h[2], w3 = h[3];
      uint32_t acc = __dp2a_lo(w0, t01, 0u);
      acc = __dp2a_hi(w1, t01, acc);
      acc = __dp2a_lo(w2, t23, acc);
      acc = __dp2a_hi(w3, t23, acc);
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
  // warps at 8 slots each; a launch with few of them is latency-bound: spread the slots over more warps instead
  const long long warps8 = (long long)((ex->geom.n_quota + 7) / 8) * batch;
  static const int force = getenv("YDORB_DESC_KPW") ? atoi(getenv("YDORB_DESC_KPW")) : 0;  // tuning knob
  const int kpw = force ? force : warps8 >= 148 * 32 ? 8 : 1;  // measured: bench.py --batch 16/64/256, tools/latency_probe.py
  DescMaps maps;
  for (int l = 0; l < kMaxLevels; l++) maps.lv[l] = ex->tmap_desc[l];
  const int frame0 = (int)((ex->d_pyr - ex->pyr_base) / ex->geom.pyr_bytes);
  auto launch = [&](auto kernel, int k) -> int {
    const int per_cta = kDescWarps * k;
    dim3 grid((ex->geom.n_quota + per_cta - 1) / per_cta, batch);
    YD_CUDA(launch_chain(kernel, grid, dim3(kDescWarps * 32), 0, s, maps, ex->d_geom, frame0, ex->d_sel, ex->d_sel_n, d_kps, d_desc,
                         capacity, d_nout, ex->d_status));
    return YDORB_OK;
  };
  const int rc = kpw >= 8 ? launch(k_describe<8>, 8) : kpw >= 4 ? launch(k_describe<4>, 4) : kpw >= 2 ? launch(k_describe<2>, 2)
                                                                                                      : launch(k_describe<1>, 1);
  if (rc != YDORB_OK) return rc;
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

void describe_box(int* box_w, int* box_h) {
  *box_w = kPatchStride;
  *box_h = kPatchRows;
}

}  // namespace ydorb
