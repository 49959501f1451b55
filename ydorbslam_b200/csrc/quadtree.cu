// E4  OrbExtractor::distributeQuadTree + QuadTreeNode::divideNode  (src/orbExtractor.cpp:4-54, 455-544)
//
// The reference grows a std::list of nodes by splitting EVERY divisible node each pass (the "biggest first"
// branch is dead code), until the list holds >= want nodes or stops growing, keeps the best-response
// point of every node (std::sort + front()) and truncates the list to `want` in list order.
//
// Restated without the list (verified against a literal std::list transcription in oracle/):
//   * the child a point falls into depends only on its coordinates, so every point has a fixed path
//     q1 q2 ... (2 bits per split, child-4 geometry bug F6 included) -> a 28-bit path code;
//   * #nodes after d passes = number of distinct d-digit prefixes  => depth D from a prefix histogram;
//   * final nodes = groups of equal D-digit prefix; a node that became a singleton at depth s<D was frozen there;
//   * list order = (s descending, then digits with alternating complement ascending);
//   * per node the winner is what libstdc++ std::sort(a.response > b.response) leaves in front: the first maximum
//     in candidate order for <= 16 points, the emulated introsort result otherwise.
//
// One CTA per (level, frame) segment.  Phases: (1) path codes in candidate order, (2) stable LSD radix sort by
// code -> prefix histogram -> D, (3) stable radix sort of candidate-ordered points by the D-digit prefix so node
// members stay in candidate order, (4) node records, (5) radix sort of node order keys, (6) emit.
#include "kernels.cuh"

namespace ydorb {

constexpr int kQtThreads = 512;
constexpr int kQtWarps = kQtThreads / 32;
constexpr int kQtItems = 8;                  // elements per thread and chunk in the scatter step
constexpr int kQtDig = 14;                   // digits per path code (28 bits): roots up to 16384 px
constexpr int kQtBits = 2 * kQtDig;

constexpr int kQtMapDig = 8;                 // depth covered by the occupancy bitmap (4^8 cells)
static_assert(kQtMapWords == (1 << (2 * kQtMapDig)) / 32, "kQtMapWords (common.cuh) must cover 4^kQtMapDig bits");

struct QtSmem {
  alignas(16) uint32_t bitmap[kQtMapWords];  // bit p set <=> some candidate's path starts with the 8-digit prefix p
  int cnt[kQtMapDig + 1];        // number of distinct d-digit prefixes, d = 0..8
  uint32_t ballots[kQtThreads / 32];
  int hist[256];
  alignas(8) int warp_cnt[kQtWarps][256];
  int scan_tmp[kQtWarps];
  int lcp_hist[kQtDig + 2];
  int misc[8];
};

// path code of a point (coordinates relative to the 16 px border), divideNode geometry :7-33.  NDIG leading digits are
// computed; the result is always aligned like a full kQtDig-digit code (missing low digits are zero).
template <int NDIG = kQtDig>
__device__ __forceinline__ uint32_t qt_code(int x, int y, int rootw, int rooth) {
  int tlx = 0, trx = rootw, tly = 0, bly = rooth;
  uint32_t code = 0;
#pragma unroll
  for (int d = 0; d < NDIG; d++) {
    const int cx = (trx + tlx + 1) >> 1;  // ceil(float(a+b)/2.0), a+b >= 0
    const int cy = (bly + tly + 1) >> 1;
    const bool right = x >= cx, low = y >= cy;
    uint32_t q;
    if (!right && !low) { q = 0; trx = cx; bly = cy; }
    else if (right && !low) { q = 1; tlx = cx; bly = cy; }
    else if (!right && low) { q = 2; trx = cx; tly = cy; }
    else { q = 3; const int t = tly; tlx = cx; tly = cy; bly = t; }  // F6: child-4 bl.y := parent tl.y
    code = (code << 2) | q;
  }
  return code << (2 * (kQtDig - NDIG));
}

// number of common leading digits of two ndig-digit values
__device__ __forceinline__ int lcp_digits(uint32_t a, uint32_t b, int ndig) {
  const uint32_t x = a ^ b;
  if (x == 0) return ndig;
  const int hb = 31 - __clz(x);
  return ndig - 1 - (hb >> 1);
}

// ---- CTA-wide stable LSD radix pass on bits [shift, shift+bits), bits <= 8.  Returns false (nothing moved) if every
// key has the same digit.  All threads must call it.  Counting is warp-aggregated (__match_any_sync, one shared-memory
// atomic per distinct digit and warp instruction): the digits of quadtree prefixes are few and heavily repeated, so
// per-element atomics would serialise on a handful of addresses.
__device__ bool qt_radix_pass(QtSmem& sm, const uint32_t* __restrict__ kin, const uint32_t* __restrict__ vin,
                              uint32_t* __restrict__ kout, uint32_t* __restrict__ vout, int n, int shift, int bits) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t dmask = (1u << bits) - 1u;
  for (int b = tid; b < 256; b += kQtThreads) sm.hist[b] = 0;
  if (tid == 0) sm.misc[0] = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 4 * kQtThreads) {
    uint32_t k4[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {  // four independent loads in flight before the first vote
      const int i = base + u * kQtThreads + tid;
      k4[u] = i < n ? kin[i] : 0u;
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = base + u * kQtThreads + tid;
      if (base + u * kQtThreads >= n) break;  // uniform over the CTA
      const uint32_t dgt = i < n ? ((k4[u] >> shift) & dmask) : 256u + lane;
      const uint32_t peers = __match_any_sync(0xffffffffu, dgt);
      if (i < n && lane == __ffs(peers) - 1) atomicAdd(&sm.hist[dgt], __popc(peers));
    }
  }
  __syncthreads();
  if (tid < 256 && sm.hist[tid] == n) sm.misc[0] = 1;
  __syncthreads();
  const int uniform = sm.misc[0];
  __syncthreads();  // everyone has read the flag before a following pass may reset it
  if (uniform) return false;
  // exclusive scan of the 256 bins by warp 0
  if (warp == 0) {
    int v[8], s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) { v[k] = sm.hist[lane * 8 + k]; s += v[k]; }
    int incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    int run = incl - s;
#pragma unroll
    for (int k = 0; k < 8; k++) { sm.hist[lane * 8 + k] = run; run += v[k]; }
  }
  __syncthreads();
  const int nbins = 1 << bits;
  constexpr int kChunk = kQtThreads * kQtItems;
  for (int base = 0; base < n; base += kChunk) {
    const int wbase = base + warp * 32 * kQtItems;
    for (int b = lane; b < nbins; b += 32) sm.warp_cnt[warp][b] = 0;
    __syncwarp();
    uint32_t key[kQtItems], val[kQtItems], peers[kQtItems];
#pragma unroll
    for (int it = 0; it < kQtItems; it++) {  // all loads of the chunk in flight before the first vote
      const int i = wbase + it * 32 + lane;
      key[it] = i < n ? kin[i] : 0u;
      val[it] = i < n ? vin[i] : 0u;
    }
#pragma unroll
    for (int it = 0; it < kQtItems; it++) {
      const int i = wbase + it * 32 + lane;
      const uint32_t dgt = i < n ? ((key[it] >> shift) & dmask) : 256u + lane;  // invalid lanes: singleton groups
      peers[it] = __match_any_sync(0xffffffffu, dgt);
      if (i < n && lane == __ffs(peers[it]) - 1) atomicAdd(&sm.warp_cnt[warp][dgt], __popc(peers[it]));
    }
    __syncthreads();
    if (tid < nbins) {
      int run = sm.hist[tid];
#pragma unroll
      for (int w = 0; w < kQtWarps; w++) {
        const int c = sm.warp_cnt[w][tid];
        sm.warp_cnt[w][tid] = run;
        run += c;
      }
      sm.hist[tid] = run;
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < kQtItems; it++) {
      const int i = wbase + it * 32 + lane;
      const bool valid = i < n;
      const uint32_t dgt = (key[it] >> shift) & dmask;
      int off = 0;
      if (valid) off = sm.warp_cnt[warp][dgt];
      __syncwarp();
      if (valid) {
        const int rank = __popc(peers[it] & ((1u << lane) - 1u));
        if (rank == 0) sm.warp_cnt[warp][dgt] = off + __popc(peers[it]);
        kout[off + rank] = key[it];
        vout[off + rank] = val[it];
      }
      __syncwarp();
    }
    // next chunk: each warp zeroes only its own counters, and the cross-warp read happens after a barrier
  }
  __syncthreads();
  return true;
}

// CTA-wide stable sort of (keys, vals) on bits [lo, nbits) (bits at and above nbits must be zero), in the fewest
// passes of at most 8 bits, digit widths balanced.  Buffers ping-pong; returns index (0/1) of the buffer holding the
// result.
__device__ int qt_sort(QtSmem& sm, uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, int n, int nbits, int lo = 0) {
  int cur = 0;
  const int total = nbits - lo;
  if (total <= 0) return 0;
  const int npass = (total + 7) / 8, width = (total + npass - 1) / npass;
  for (int p = 0; p < npass; p++) {
    const int shift = lo + p * width, bits = min(width, nbits - shift);
    const bool moved = cur == 0 ? qt_radix_pass(sm, k0, v0, k1, v1, n, shift, bits) : qt_radix_pass(sm, k1, v1, k0, v0, n, shift, bits);
    if (moved) cur ^= 1;
  }
  return cur;
}

// CTA-wide exclusive scan of one int per thread; returns the exclusive prefix, *total = block sum.
__device__ __forceinline__ int qt_block_scan(QtSmem& sm, int v, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  __syncthreads();  // protect scan_tmp reuse
  if (lane == 31) sm.scan_tmp[warp] = incl;
  __syncthreads();
  int wsum = (lane < kQtWarps) ? sm.scan_tmp[lane] : 0;
  int wincl = wsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, wincl, o);
    if (lane >= o) wincl += t;
  }
  const int wexcl = __shfl_sync(0xffffffffu, wincl - wsum, warp);
  *total = __shfl_sync(0xffffffffu, wincl, kQtWarps - 1);
  return wexcl + incl - v;
}

// ---- libstdc++ std::sort emulation (bits/stl_algo.h) on packed entries: response in bits 31..24, payload below.
// comp(a,b) = a.response > b.response, exactly the reference's lambda at src/orbExtractor.cpp:537-538.
struct StdSortEmu {
  uint32_t* a;
  __device__ __forceinline__ static bool comp(uint32_t x, uint32_t y) { return (x >> 24) > (y >> 24); }
  __device__ __forceinline__ void swp(int i, int j) { const uint32_t t = a[i]; a[i] = a[j]; a[j] = t; }

  __device__ void move_median_to_first(int res, int x, int y, int z) {
    if (comp(a[x], a[y])) {
      if (comp(a[y], a[z])) swp(res, y);
      else if (comp(a[x], a[z])) swp(res, z);
      else swp(res, x);
    } else if (comp(a[x], a[z])) swp(res, x);
    else if (comp(a[y], a[z])) swp(res, z);
    else swp(res, y);
  }
  __device__ int unguarded_partition(int first, int last, int pivot) {
    while (true) {
      while (comp(a[first], a[pivot])) ++first;
      --last;
      while (comp(a[pivot], a[last])) --last;
      if (!(first < last)) return first;
      swp(first, last);
      ++first;
    }
  }
  __device__ void adjust_heap(int first, int hole, int len, uint32_t value) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
      child = 2 * (child + 1);
      if (comp(a[first + child], a[first + child - 1])) child--;
      a[first + hole] = a[first + child];
      hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
      child = 2 * (child + 1);
      a[first + hole] = a[first + child - 1];
      hole = child - 1;
    }
    int parent = (hole - 1) / 2;
    while (hole > top && comp(a[first + parent], value)) {
      a[first + hole] = a[first + parent];
      hole = parent;
      parent = (hole - 1) / 2;
    }
    a[first + hole] = value;
  }
  __device__ void heap_sort(int first, int last) {  // __partial_sort(first,last,last)
    const int len = last - first;
    if (len >= 2) {
      int parent = (len - 2) / 2;
      while (true) {
        adjust_heap(first, parent, len, a[first + parent]);
        if (parent == 0) break;
        parent--;
      }
    }
    int l = last;
    while (l - first > 1) {
      --l;
      const uint32_t v = a[l];
      a[l] = a[first];
      adjust_heap(first, 0, l - first, v);
    }
  }
  __device__ void unguarded_linear_insert(int last) {
    const uint32_t val = a[last];
    int next = last - 1;
    while (comp(val, a[next])) { a[last] = a[next]; last = next; --next; }
    a[last] = val;
  }
  __device__ void insertion_sort(int first, int last) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
      if (comp(a[i], a[first])) {
        const uint32_t val = a[i];
        for (int j = i; j > first; --j) a[j] = a[j - 1];
        a[first] = val;
      } else {
        unguarded_linear_insert(i);
      }
    }
  }
  // What std::sort leaves in FRONT depends only on the leftmost chain of introsort partitions: the recursion into a
  // right part [cut, last) never touches anything left of cut, every element right of the leftmost segment is <= all
  // of its members, and the final insertion sort moves an element forward only past strictly smaller ones.  So the
  // front is the first maximum, in array order, of the leftmost segment once it is <= 16 long (or a[0] after the
  // depth-limit heap sort).  Runs the chain on a[0..n) and returns the length of that segment.
  __device__ int leftmost_segment(int n) {
    int first = 0, last = n, depth = 2 * (31 - __clz(n));
    while (last - first > 16) {
      if (depth == 0) { heap_sort(first, last); return 1; }
      --depth;
      const int mid = first + (last - first) / 2;
      move_median_to_first(first, first + 1, mid, last - 1);
      last = unguarded_partition(first + 1, last, first);
    }
    return last;
  }
  // returns false if the explicit stack overflowed (never for n < 2^31)
  __device__ bool sort(int n) {
    if (n <= 0) return true;
    int st_first[64], st_last[64], st_depth[64];
    int sp = 0;
    st_first[0] = 0; st_last[0] = n; st_depth[0] = 2 * (31 - __clz(n)); sp = 1;
    while (sp > 0) {
      --sp;
      int first = st_first[sp], last = st_last[sp], depth = st_depth[sp];
      while (last - first > 16) {
        if (depth == 0) { heap_sort(first, last); break; }
        --depth;
        const int mid = first + (last - first) / 2;
        move_median_to_first(first, first + 1, mid, last - 1);
        const int cut = unguarded_partition(first + 1, last, first);
        if (sp >= 64) return false;
        st_first[sp] = cut; st_last[sp] = last; st_depth[sp] = depth; ++sp;  // right part: disjoint, order-free
        last = cut;
      }
    }
    if (n > 16) {
      insertion_sort(0, 16);
      for (int i = 16; i < n; ++i) unguarded_linear_insert(i);
    } else {
      insertion_sort(0, n);
    }
    return true;
  }
};

// ---- per-frame exclusive scan of the per-cell candidate counts (candidate order = cell-major)
__global__ void __launch_bounds__(256) k_cell_scan(int nc, const int* __restrict__ cell_count,
                                                   int* __restrict__ cell_rank) {
  pdl_wait();
  __shared__ int s_warp[8];
  __shared__ int s_carry;
  const int frame = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int* cnt = cell_count + (size_t)frame * nc;
  int* rank = cell_rank + (size_t)frame * (nc + 1);
  if (tid == 0) s_carry = 0;
  __syncthreads();
  constexpr int kPer = 8;  // consecutive cells per thread and chunk: 2048 cells per CTA-wide scan
  for (int base = 0; base < nc; base += 256 * kPer) {
    const int i0 = base + tid * kPer;
    int v[kPer], sum = 0;
#pragma unroll
    for (int k = 0; k < kPer; k++) {
      v[k] = (i0 + k < nc) ? cnt[i0 + k] : 0;
      sum += v[k];
    }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < warp; w++) woff += s_warp[w];
    const int carry = s_carry;
    int run = carry + woff + incl - sum;
#pragma unroll
    for (int k = 0; k < kPer; k++) {
      if (i0 + k < nc) rank[i0 + k] = run;
      run += v[k];
    }
    __syncthreads();
    if (tid == 255) s_carry = carry + woff + incl;
    __syncthreads();
  }
  if (tid == 0) rank[nc] = s_carry;
}

// what the per-node walk needs of a level's geometry
struct QtLevel { int cell_base, cell_cols, cell_rows, cell_w, cell_h, root_w, root_h; };
struct QtArgs {
  const FrameGeomDev* g;
  const int4* cells;
  const int* cell_count;
  const int* cell_start;
  const int* cell_rank;
  const uint32_t* raw;
  uint32_t* keys[3];
  uint32_t* vals[3];
  uint32_t* bitmap;  // [frame][kMaxLevels][kQtMapWords]
  uint32_t* fold;    // [frame][kMaxLevels][2][kQtMapWords]
  int* meta;         // [frame][kMaxLevels][8]
  int4* sel;
  int* sel_n;
  int* level_ncand;
  int* status;
  int cell_base[kMaxLevels], cell_end[kMaxLevels], quota[kMaxLevels];  // per level: first cell, one past its last, keypoints wanted
  QtLevel lv[kMaxLevels];
  const uint8_t* xpath;  // x-decisions of the 8 leading splits per level and column (set_size)
  int xoff[kMaxLevels], xlen[kMaxLevels];
  int quota_base[kMaxLevels];  // first output slot of the level in a frame's selection
  int n_cells, raw_cap, n_levels, n_quota;  // of the current geometry: launch parameters, so that no kernel starts with a
                                   // global load that everything else depends on
};

// writes (key, packed candidate) of every candidate of the level at its candidate-order position
__device__ void qt_fill(const QtArgs& A, const LevelGeom& lv, int frame, int level_base, uint32_t* keys, uint32_t* vals,
                        uint32_t* bitmap) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ncell = lv.cell_cols * lv.cell_rows;
  const size_t cbase = (size_t)frame * A.n_cells + lv.cell_base;
  const int* rank = A.cell_rank + (size_t)frame * (A.n_cells + 1) + lv.cell_base;
  const uint32_t* raw = A.raw + (size_t)frame * A.raw_cap;
  for (int c = warp; c < ncell; c += kQtWarps) {
    const int cnt = A.cell_count[cbase + c];
    if (cnt == 0) continue;
    const int start = A.cell_start[cbase + c];
    const int pos0 = rank[c] - level_base;
    for (int e = lane; e < cnt; e += 32) {
      const uint32_t r = raw[start + e];  // y:12 | x:12 | response:8, border-relative coordinates
      const uint32_t code = qt_code<kQtDig>((int)((r >> 8) & 4095u), (int)(r >> 20), lv.root_w, lv.root_h);
      keys[pos0 + e] = code;
      vals[pos0 + e] = r;
      const uint32_t p8 = code >> (kQtBits - 2 * kQtMapDig);
      const uint32_t bit = 1u << (p8 & 31u);
      if (!(bitmap[p8 >> 5] & bit)) atomicOr(&bitmap[p8 >> 5], bit);
    }
  }
}

// Path codes of every candidate of every frame, written at the candidate-order position the quadtree CTA expects
// (keys[0] / vals[0]).  One warp per (frame, cell); only the kQtMapDig leading digits are computed here -- all the
// common case needs (depth <= 8) -- so this instruction-heavy step runs across the whole GPU instead of inside the one
// CTA that owns a (level, frame) segment.  The rare deeper case recomputes full codes inside k_quadtree (qt_fill).
constexpr int kCodeWarps = 8;
// kCodeLanes = lanes that share one FAST cell: 8 for launches that fill the GPU anyway, 32 spreads a small launch
template <int kCodeLanes>
__global__ void __launch_bounds__(kCodeWarps * 32) k_qt_codes(QtArgs A) {
  constexpr int kCodeCells = kCodeWarps * 32 / kCodeLanes;  // cells per CTA
  pdl_wait();
  // a cell holds a handful of candidates (~6 on a 640x480 frame): kCodeLanes lanes per cell, several cells per warp
  // (no warp-level collectives below, the groups of a warp simply diverge)
  const int frame = blockIdx.y, lane = threadIdx.x & (kCodeLanes - 1);
  const int c = blockIdx.x * kCodeCells + threadIdx.x / kCodeLanes;  // cell index within the frame
  if (c >= A.n_cells) return;
  const int cnt = A.cell_count[(size_t)frame * A.n_cells + c];
  if (cnt == 0) return;
  const FrameGeomDev* g = A.g;
  int level = 0;
  for (int l = 1; l < A.n_levels && c >= g->lv[l].cell_base; l++) level = l;  // cell bases ascend with the level
  const int root_h = g->lv[level].root_h;
  const int start = A.cell_start[(size_t)frame * A.n_cells + c];
  const size_t pos0 = (size_t)frame * A.raw_cap + A.cell_rank[(size_t)frame * (A.n_cells + 1) + c];
  const uint32_t* raw = A.raw + (size_t)frame * A.raw_cap;
  const uint8_t* xp = A.xpath + A.xoff[level];
  const int xmax = A.xlen[level] - 1;
  uint32_t* bitmap = A.bitmap + ((size_t)frame * kMaxLevels + level) * kQtMapWords;
  for (int e = lane; e < cnt; e += kCodeLanes) {
    const uint32_t r = raw[start + e];  // y:12 | x:12 | response:8, border-relative coordinates
    // divideNode splits x by x alone, so the 8 x-decisions come from a per-column table; only the y side (which the
    // child-4 quirk F6 couples to the x-decisions) is replayed here.  Same codes as qt_code<kQtMapDig>.
    const uint32_t xb = xp[min((int)((r >> 8) & 4095u), xmax)];
    const int y = (int)(r >> 20);
    int tly = 0, bly = root_h;
    uint32_t code = 0;
#pragma unroll
    for (int d = 0; d < kQtMapDig; d++) {
      const int cy = (bly + tly + 1) >> 1;
      const bool low = y >= cy, right = ((xb >> (kQtMapDig - 1 - d)) & 1u) != 0u;
      const int t = tly;
      tly = low ? cy : tly;
      bly = low ? (right ? t : bly) : cy;  // child 4: bl.y := parent tl.y (F6)
      code = (code << 2) | (low ? 2u : 0u) | (right ? 1u : 0u);
    }
    code <<= 2 * (kQtDig - kQtMapDig);
    A.keys[0][pos0 + e] = code;
    A.vals[0][pos0 + e] = r;
    const uint32_t p8 = code >> (kQtBits - 2 * kQtMapDig);
    const uint32_t bit = 1u << (p8 & 31u);
    if (!(bitmap[p8 >> 5] & bit)) atomicOr(&bitmap[p8 >> 5], bit);  // occupancy of the 8-digit prefixes (k_qt_depth)
  }
}

// ======================================================================================================================
// Node-table path (the common case, depth <= kQtMapDig).  Only the winner of every node matters, and which node a
// candidate falls into is a rank query into the occupancy bitmap, so the per-candidate work needs no sort at all and
// runs GPU-wide with order-free atomics; the CTA that owns a (level, frame) segment only touches the M nodes:
//   k_qt_codes   (GPU-wide)   path codes + occupancy bitmap of the 8-digit prefixes
//   k_qt_depth   (CTA/segment) depth D from the bitmap, D-digit occupancy + popcount prefix = node index of a prefix
//   k_qt_pass1/2 (GPU-wide)   per node: member count, first maximum in candidate order (atomicMax on
//                             response << 24 | ~position), number of members that tie with it
//   k_qt_records (CTA/segment) node records (order key + winner); nodes whose winner std::sort decides are queued
//   k_qt_ties    (GPU-wide)   warp per queued node: members gathered in candidate order through the FAST cell grid,
//                             libstdc++ std::sort emulation
//   k_qt_emit    (CTA/segment) list-order sort, truncate, emit
// Segments that need more than kQtMapDig passes (tightly clustered points) keep the sort-based k_quadtree below.
enum { kMetaN = 0, kMetaD = 1, kMetaM = 2, kMetaDeep = 3, kMetaBase = 4, kMetaFlagged = 5, kMetaBig = 6 };

__global__ void __launch_bounds__(kQtThreads) k_qt_depth(QtArgs A) {
  pdl_wait();
  __shared__ alignas(16) uint32_t s_map[kQtMapWords];
  __shared__ alignas(16) uint32_t s_fold[kQtMapWords];
  __shared__ int s_cnt[kQtMapDig + 1];
  __shared__ uint32_t s_ballots[kQtThreads / 32];
  __shared__ int s_scan[kQtWarps];
  __shared__ int s_misc[4];
  const int level = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // the segment's occupancy bitmap: one 16-byte load per thread, requested first -- its address depends on nothing, while
  // the candidate count below sits behind two dependent loads (geometry, then the cell ranks)
  static_assert(kQtMapWords == 4 * kQtThreads, "one uint4 per thread");
  uint4* gm4 = reinterpret_cast<uint4*>(A.bitmap + ((size_t)frame * kMaxLevels + level) * kQtMapWords);
  const uint4 map_v = gm4[tid];
  const int* rank = A.cell_rank + (size_t)frame * (A.n_cells + 1);
  const int level_base = rank[A.cell_base[level]];
  const int n = rank[A.cell_end[level]] - level_base;
  const int want = A.quota[level];
  int* meta = A.meta + ((size_t)frame * kMaxLevels + level) * 8;
  uint32_t* gmap = A.bitmap + ((size_t)frame * kMaxLevels + level) * kQtMapWords;
  if (tid == 0) {
    A.level_ncand[(size_t)frame * kMaxLevels + level] = n;
    meta[kMetaBase] = level_base;
  }
  if (n == 0 || want <= 0) {
    if (tid == 0) {
      A.sel_n[(size_t)frame * kMaxLevels + level] = 0;
      meta[kMetaN] = 0; meta[kMetaD] = 0; meta[kMetaM] = 0; meta[kMetaDeep] = 0;
    }
    // the bitmap is all zero when n == 0; with n > 0 and no quota it must be cleared for the next call
    if (n != 0)
      for (int i = tid; i < kQtMapWords; i += kQtThreads) gmap[i] = 0;
    return;
  }
  {
    // the whole bitmap arrived in a single round trip; zero it for the next call
    gm4[tid] = make_uint4(0u, 0u, 0u, 0u);
    reinterpret_cast<uint4*>(s_map)[tid] = map_v;
    reinterpret_cast<uint4*>(s_fold)[tid] = make_uint4(0u, 0u, 0u, 0u);
  }
  if (tid <= kQtMapDig) s_cnt[tid] = 0;
  __syncthreads();
  // #nodes after d passes = #distinct d-digit prefixes (d = 1..8): fold the bitmap.  Thread t < 256 owns the 8 words
  // under the 4-digit prefix t; deeper levels fold inside a word, shallower ones across lanes by ballot.
  {
    int c8 = 0, c7 = 0, c6 = 0, c5 = 0;
    bool any = false;
    if (tid < 256) {
      uint32_t pair = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const uint32_t w = s_map[tid * 8 + k];
        c8 += __popc(w);
        const uint32_t m7 = (w | (w >> 1) | (w >> 2) | (w >> 3)) & 0x11111111u;
        c7 += __popc(m7);
        c6 += __popc((m7 | (m7 >> 4) | (m7 >> 8) | (m7 >> 12)) & 0x00010001u);
        pair |= w;
        if (k & 1) { c5 += pair != 0; any |= pair != 0; pair = 0; }
      }
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, any);
    c8 = __reduce_add_sync(0xffffffffu, c8);
    c7 = __reduce_add_sync(0xffffffffu, c7);
    c6 = __reduce_add_sync(0xffffffffu, c6);
    c5 = __reduce_add_sync(0xffffffffu, c5);
    if (lane == 0) {
      if (c8) { atomicAdd(&s_cnt[8], c8); atomicAdd(&s_cnt[7], c7); atomicAdd(&s_cnt[6], c6); atomicAdd(&s_cnt[5], c5); }
      if (bal) {
        atomicAdd(&s_cnt[4], __popc(bal));
        atomicAdd(&s_cnt[3], __popc((bal | (bal >> 1) | (bal >> 2) | (bal >> 3)) & 0x11111111u));
        atomicAdd(&s_cnt[2], ((bal & 0xffffu) != 0) + ((bal >> 16) != 0));
      }
      s_ballots[warp] = bal;
    }
  }
  __syncthreads();
  if (tid == 0) {
    int c1 = 0;
    for (int k = 0; k < 4; k++) c1 += (s_ballots[2 * k] | s_ballots[2 * k + 1]) != 0;
    s_cnt[1] = c1;
    // the reference's loop condition (:479): split again while the list grew and is still shorter than `want`
    int D = 0, last = 0, size = 1, deep = 0;
    while (size > last && size < want && D < kQtDig) {
      last = size;
      D++;
      if (D > kQtMapDig) { deep = 1; break; }
      size = s_cnt[D];
    }
    s_misc[0] = D;
    s_misc[1] = deep;
  }
  __syncthreads();
  const int D = s_misc[0];
  if (s_misc[1]) {  // k_quadtree handles the segment
    if (tid == 0) { meta[kMetaN] = n; meta[kMetaD] = D; meta[kMetaM] = 0; meta[kMetaDeep] = 1; }
    return;
  }
  // D-digit occupancy: bit p set <=> some 8-digit prefix below p is occupied
  const int drop = 2 * (kQtMapDig - D);
  for (int w = tid; w < kQtMapWords; w += kQtThreads) {
    uint32_t bits = s_map[w];
    if (drop == 0) { s_fold[w] = bits; continue; }
    while (bits) {
      const int b = __ffs(bits) - 1;
      // all 8-digit prefixes that share the D-digit prefix of this one sit in the next (1 << drop) positions: skip them
      const uint32_t p8 = (uint32_t)w * 32u + (uint32_t)b, pd = p8 >> drop;
      atomicOr(&s_fold[pd >> 5], 1u << (pd & 31u));
      const uint32_t next = ((pd + 1u) << drop);  // first 8-digit prefix of the next D-digit group
      if (next >= (uint32_t)(w + 1) * 32u) break;
      bits &= ~((1u << (next - (uint32_t)w * 32u)) - 1u);
    }
  }
  __syncthreads();
  // exclusive popcount prefix per word: node index of prefix p = prefix[p >> 5] + popc(fold[p >> 5] & below(p))
  uint32_t* gfold = A.fold + ((size_t)frame * kMaxLevels + level) * 2 * kQtMapWords;
  {
    constexpr int kPer = kQtMapWords / kQtThreads;  // 4 consecutive words per thread
    int c[kPer], sum = 0;
#pragma unroll
    for (int k = 0; k < kPer; k++) { c[k] = __popc(s_fold[tid * kPer + k]); sum += c[k]; }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < warp; w++) woff += s_scan[w];
    int run = woff + incl - sum;
#pragma unroll
    for (int k = 0; k < kPer; k++) {
      gfold[tid * kPer + k] = s_fold[tid * kPer + k];
      gfold[kQtMapWords + tid * kPer + k] = (uint32_t)run;
      run += c[k];
    }
    if (tid == kQtThreads - 1) s_misc[2] = run;
  }
  __syncthreads();
  const int M = s_misc[2];
  // node tables of the segment: best key, member count, tie count
  const size_t seg = (size_t)frame * A.raw_cap + level_base;
  uint32_t *best = A.keys[1] + seg, *cnt = A.vals[1] + seg, *ties = A.keys[2] + seg;
  for (int m = tid; m < M; m += kQtThreads) { best[m] = 0; cnt[m] = 0; ties[m] = 0; }
  if (tid == 0) { meta[kMetaN] = n; meta[kMetaD] = D; meta[kMetaM] = M; meta[kMetaDeep] = 0; }
}

// segment of a candidate position and its node index; returns false for positions past the frame's candidates or in
// segments the node-table path does not handle
struct QtWhere { int level, i, D; size_t seg; uint32_t m; };
__device__ __forceinline__ bool qt_locate(const QtArgs& A, int frame, int pos, int total, uint32_t code, QtWhere* w) {
  if (pos >= total) return false;
  const int* meta = A.meta + (size_t)frame * kMaxLevels * 8;
  int level = 0;
  for (int l = 1; l < A.n_levels; l++)
    if (pos >= meta[l * 8 + kMetaBase]) level = l;
  meta += level * 8;
  if (meta[kMetaDeep] || meta[kMetaN] == 0) return false;
  const int base = meta[kMetaBase], D = meta[kMetaD];
  const uint32_t p = D == 0 ? 0u : code >> (kQtBits - 2 * D);
  const uint32_t* fold = A.fold + ((size_t)frame * kMaxLevels + level) * 2 * kQtMapWords;
  w->level = level;
  w->i = pos - base;
  w->D = D;
  w->seg = (size_t)frame * A.raw_cap + base;
  w->m = fold[kQtMapWords + (p >> 5)] + __popc(fold[p >> 5] & ((1u << (p & 31u)) - 1u));
  return true;
}

// Candidates that follow each other in candidate order come from the same FAST cell and mostly fall into the same node:
// the lanes of a warp that hit the same node combine their contribution (MATCH.ANY + REDUX) and one of them issues the
// atomics, instead of up to 32 same-address atomics serialising in L2.
__global__ void __launch_bounds__(256) k_qt_pass1(QtArgs A) {
  pdl_wait();
  const int frame = blockIdx.y, lane = threadIdx.x & 31;
  const int total = A.cell_rank[(size_t)frame * (A.n_cells + 1) + A.n_cells];
  for (int base = blockIdx.x * 256 + (threadIdx.x & ~31); base < total; base += gridDim.x * 256) {  // warp-uniform
    const int pos = base + lane;
    // the candidate's own words first: their addresses depend on nothing but pos, the segment lookup is a chain of loads
    uint32_t code = 0, r = 0;
    if (pos < total) {
      code = A.keys[0][(size_t)frame * A.raw_cap + pos];
      r = A.vals[0][(size_t)frame * A.raw_cap + pos];
    }
    QtWhere w;
    const bool ok = qt_locate(A, frame, pos, total, code, &w);
    uint32_t key = 0;
    if (ok) {
      A.vals[2][(size_t)frame * A.raw_cap + pos] = w.m;  // node index, reused by pass 2
      key = ((r & 255u) << 24) | (0xFFFFFFu - (uint32_t)w.i);  // max response, then first position
    }
    const uint32_t group = ok ? ((uint32_t)w.level << 16) | w.m : 0x80000000u | (uint32_t)lane;
    const uint32_t peers = __match_any_sync(0xffffffffu, group);
    const uint32_t top = __reduce_max_sync(peers, key);
    if (ok && lane == __ffs(peers) - 1) {
      atomicMax(&A.keys[1][w.seg + w.m], top);
      atomicAdd(&A.vals[1][w.seg + w.m], (uint32_t)__popc(peers));
    }
  }
}

__global__ void __launch_bounds__(256) k_qt_pass2(QtArgs A) {
  pdl_wait();
  const int frame = blockIdx.y, lane = threadIdx.x & 31;
  const int total = A.cell_rank[(size_t)frame * (A.n_cells + 1) + A.n_cells];
  const int* fmeta = A.meta + (size_t)frame * kMaxLevels * 8;
  for (int base = blockIdx.x * 256 + (threadIdx.x & ~31); base < total; base += gridDim.x * 256) {  // warp-uniform
    const int pos = base + lane;
    bool tie = false;
    uint32_t group = 0x80000000u | (uint32_t)lane;
    size_t slot = 0;
    if (pos < total) {
      const uint32_t r = A.vals[0][(size_t)frame * A.raw_cap + pos];  // requested before the segment lookup
      const uint32_t m = A.vals[2][(size_t)frame * A.raw_cap + pos];
      int level = 0;
      for (int l = 1; l < A.n_levels; l++)
        if (pos >= fmeta[l * 8 + kMetaBase]) level = l;
      const int* meta = fmeta + level * 8;
      if (!meta[kMetaDeep] && meta[kMetaN] != 0) {
        slot = (size_t)frame * A.raw_cap + meta[kMetaBase] + m;
        tie = (r & 255u) == (A.keys[1][slot] >> 24);
        group = ((uint32_t)level << 16) | m;
      }
    }
    const uint32_t peers = __match_any_sync(0xffffffffu, group);
    const uint32_t tb = __ballot_sync(0xffffffffu, tie) & peers;  // lanes of this node that tie with its maximum
    if (tie && lane == __ffs(tb) - 1) atomicAdd(&A.keys[2][slot], (uint32_t)__popc(tb));
  }
}

// previous / next occupied D-digit prefix around bit b of word w (or -1)
__device__ __forceinline__ int qt_prev_prefix(const uint32_t* fold, int w, int b) {
  uint32_t below = fold[w] & ((1u << b) - 1u);
  while (below == 0u) {
    if (--w < 0) return -1;
    below = fold[w];
  }
  return w * 32 + 31 - __clz(below);
}
__device__ __forceinline__ int qt_next_prefix(const uint32_t* fold, int nwords, int w, int b) {
  uint32_t above = b == 31 ? 0u : (fold[w] & ~((2u << b) - 1u));
  while (above == 0u) {
    if (++w >= nwords) return -1;
    above = fold[w];
  }
  return w * 32 + __ffs(above) - 1;
}

struct QtFinSmem {
  QtSmem q;                     // radix-sort workspace (q.bitmap holds the D-digit occupancy here)
  alignas(16) uint32_t prefix[kQtMapWords]; // exclusive popcount prefix per word
};

__global__ void __launch_bounds__(kQtThreads) k_qt_records(QtArgs A) {
  pdl_wait();
  __shared__ QtFinSmem fs;
  QtSmem& sm = fs.q;
  const int level = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x;
  int* meta = A.meta + ((size_t)frame * kMaxLevels + level) * 8;
  // both tables of the segment are full kQtMapWords arrays: one 16-byte load each per thread, requested before the
  // segment's meta data is known (their addresses depend on nothing; an empty segment just drops them)
  static_assert(kQtMapWords == 4 * kQtThreads, "one uint4 per thread and table");
  const uint4* gf4 = reinterpret_cast<const uint4*>(A.fold + ((size_t)frame * kMaxLevels + level) * 2 * kQtMapWords);
  const uint4 tab_bm = gf4[tid], tab_pf = gf4[kQtMapWords / 4 + tid];
  const int n = meta[kMetaN];
  if (n == 0 || meta[kMetaDeep]) return;
  const int D = meta[kMetaD], M = meta[kMetaM], level_base = meta[kMetaBase];
  const size_t seg = (size_t)frame * A.raw_cap + level_base;
  const uint32_t* CV = A.vals[0] + seg;   // packed candidates, candidate order
  uint32_t* best = A.keys[1] + seg;       // node winners (NV) once the records are written
  uint32_t* cnt = A.vals[1] + seg;
  uint32_t* ties = A.keys[2] + seg;
  uint32_t* NK = A.vals[2] + seg;         // order keys [0, M), then the flagged list (3 words per entry)
  uint32_t* NV = best;
  uint32_t* flagged = NK + M;             // M + 3 * #(nodes with > 16 members) <= n
  const int nbits = 1 << (2 * D), nwords = (nbits + 31) >> 5;
  if (4 * tid < nwords) {
    reinterpret_cast<uint4*>(sm.bitmap)[tid] = tab_bm;
    reinterpret_cast<uint4*>(fs.prefix)[tid] = tab_pf;
  }
  if (tid == 0) sm.misc[4] = 0;
  __syncthreads();

  // node records: order key + winner (first maximum in candidate order); nodes std::sort decides are queued.
  // Thread = one D-digit prefix; its node index is a rank query.
  for (int p = tid; p < nbits; p += kQtThreads) {
    const int w = p >> 5, b = p & 31;
    const uint32_t word = sm.bitmap[w];
    if (!((word >> b) & 1u)) continue;
    const int m = (int)fs.prefix[w] + __popc(word & ((1u << b) - 1u));
    const uint32_t pre = (uint32_t)p;
    const int len = (int)cnt[m];
    int s = D;
    if (len == 1) {
      int l = 0;
      const int pp = qt_prev_prefix(sm.bitmap, w, b), pn = qt_next_prefix(sm.bitmap, nwords, w, b);
      if (pp >= 0) l = max(l, lcp_digits(pre, (uint32_t)pp, D));
      if (pn >= 0) l = max(l, lcp_digits(pre, (uint32_t)pn, D));
      s = (n == 1) ? 0 : min(D, l + 1);
    }
    const uint32_t bk = best[m];
    const int first_max = (int)(0xFFFFFFu - (bk & 0xFFFFFFu));
    if (len > 16 && ties[m] > 1u) {  // introsort decides among tied maxima
      const int f = atomicAdd(&sm.misc[4], 1);
      flagged[3 * f] = (uint32_t)m;
      flagged[3 * f + 1] = pre;
      flagged[3 * f + 2] = (uint32_t)len;
    }
    const uint32_t ps = (s == 0) ? 0u : (pre >> (2 * (D - s)));
    uint32_t mask = 0;
    for (int k = s; k >= 1; k -= 2) mask |= 3u << (2 * (s - k));
    NK[m] = ((uint32_t)(D - s) << (2 * D)) | (ps ^ mask);  // 2D + 4 significant bits
    NV[m] = CV[first_max];
  }
  __syncthreads();
  if (tid == 0) { meta[kMetaFlagged] = sm.misc[4]; meta[kMetaBig] = 0; }
}

// nodes whose winner std::sort decides (tied maxima among more than 16 members): one warp per node, the
// flagged nodes of all levels of a frame spread over the frame's CTAs (a 4K frame has hundreds of them on level 0).
// The members, in candidate order, are the candidates with the node's prefix found by walking the FAST cells
// (candidate order = cell-major) that overlap the node's rectangle.
constexpr int kTieWarps = 8;
__global__ void __launch_bounds__(kTieWarps * 32) k_qt_ties(QtArgs A) {
  pdl_wait();
  __shared__ alignas(8) uint32_t s_buf[kTieWarps][256];
  const int frame = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int* fmeta = A.meta + (size_t)frame * kMaxLevels * 8;
  const int n_levels = A.n_levels;
  // lane l holds level l's flagged count and the inclusive prefix over the levels (most warps find nothing to do:
  // keep that path to one round trip)
  static_assert(kMaxLevels <= 32, "one lane per level");
  int lc = 0;
  if (lane < n_levels) {
    const int* mm = fmeta + lane * 8;
    lc = (mm[kMetaN] != 0 && !mm[kMetaDeep]) ? mm[kMetaFlagged] : 0;
  }
  int lincl = lc;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, lincl, o);
    if (lane >= o) lincl += t;
  }
  const int total = __shfl_sync(0xffffffffu, lincl, 31);
  const int nwarps = gridDim.x * kTieWarps;
  for (int item = blockIdx.x * kTieWarps + warp; item < total; item += nwarps) {
    const int level = __popc(__ballot_sync(0xffffffffu, lincl <= item));  // lanes past n_levels hold `total` > item
    const int f = item - (__shfl_sync(0xffffffffu, lincl, level) - __shfl_sync(0xffffffffu, lc, level));
    int* meta = fmeta + level * 8;
    const QtLevel lv = A.lv[level];  // launch parameters: no load between the meta data and the node's record
    const int D = meta[kMetaD], M = meta[kMetaM], level_base = meta[kMetaBase];
    const int sh = kQtBits - 2 * D;
    const size_t seg = (size_t)frame * A.raw_cap + level_base;
    const uint32_t* CK = A.keys[0] + seg;
    const uint32_t* CV = A.vals[0] + seg;
    uint32_t* NV = A.keys[1] + seg;
    const uint32_t* flagged = A.vals[2] + seg + M;
    uint32_t* big_buf = A.keys[2] + seg;  // node tables are dead by now: scratch for nodes with more than 128 members
    uint32_t* big_pos = A.vals[1] + seg;
    const size_t cbase = (size_t)frame * A.n_cells + lv.cell_base;
    const int* rank = A.cell_rank + (size_t)frame * (A.n_cells + 1) + lv.cell_base;
      const int m = (int)flagged[3 * f], len = (int)flagged[3 * f + 2];
      const uint32_t pre = flagged[3 * f + 1];
      // rectangle of the node: replay divideNode on the prefix digits, tracking the true bounds of the members
      int tlx = 0, trx = lv.root_w, tly = 0, bly = lv.root_h;
      int xlo = 0, xhi = lv.root_w, ylo = 0, yhi = lv.root_h;
      for (int d = D - 1; d >= 0; d--) {
        const uint32_t q = (pre >> (2 * d)) & 3u;
        const int cx = (trx + tlx + 1) >> 1, cy = (bly + tly + 1) >> 1;
        if (q & 1u) xlo = max(xlo, cx); else xhi = min(xhi, cx);
        if (q & 2u) ylo = max(ylo, cy); else yhi = min(yhi, cy);
        if (q == 0u) { trx = cx; bly = cy; }
        else if (q == 1u) { tlx = cx; bly = cy; }
        else if (q == 2u) { trx = cx; tly = cy; }
        else { const int t = tly; tlx = cx; tly = cy; bly = t; }
      }
      // a candidate at border-relative (x, y) was found by cell ((x - 3) / cell_w, (y - 3) / cell_h): the detection
      // areas [j * cell_w + 3, (j + 1) * cell_w + 3) tile the level exactly (src/orbExtractor.cpp:563-579)
      const int j0 = max(0, (max(xlo, 3) - 3) / lv.cell_w), j1 = min(lv.cell_cols - 1, (max(xhi - 1, 3) - 3) / lv.cell_w);
      const int i0 = max(0, (max(ylo, 3) - 3) / lv.cell_h), i1 = min(lv.cell_rows - 1, (max(yhi - 1, 3) - 3) / lv.cell_h);
      uint32_t *buf, *pos;
      if (len <= 128) {
        buf = s_buf[warp];
        pos = buf + 128;
      } else {
        int o = 0;
        if (lane == 0) o = atomicAdd(&meta[kMetaBig], len);
        o = __shfl_sync(0xffffffffu, o, 0);
        buf = big_buf + o;
        pos = big_pos + o;
      }
      // the cells of the rectangle in cell-major order, 32 at a time: lane = cell (count, first position), then the
      // candidates of those cells as one flat list, 32 per step (lane finds its cell by bisection over the prefix sums)
      const int ncx = j1 - j0 + 1, ncell = max(0, i1 - i0 + 1) * max(0, ncx);
      int k = 0;
      for (int cb = 0; cb < ncell; cb += 32) {
        const int ci = cb + lane;
        int cc = 0, p0 = 0;
        if (ci < ncell) {
          const int c = (i0 + ci / ncx) * lv.cell_cols + j0 + ci % ncx;
          cc = A.cell_count[cbase + c];
          p0 = rank[c] - level_base;
        }
        int incl = cc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        const int excl = incl - cc, T = __shfl_sync(0xffffffffu, incl, 31);
        for (int t0 = 0; t0 < T; t0 += 32) {
          const int t = t0 + lane;
          int lo = 0;
#pragma unroll
          for (int step = 16; step >= 1; step >>= 1) {
            const int e = __shfl_sync(0xffffffffu, excl, (lo + step) & 31);
            if (lo + step < 32 && e <= t) lo += step;
          }
          const int pc = __shfl_sync(0xffffffffu, p0, lo), ec = __shfl_sync(0xffffffffu, excl, lo);
          const int idx = pc + (t - ec);
          const bool is = t < T && (D == 0 || (CK[idx] >> sh) == pre);
          const uint32_t bal = __ballot_sync(0xffffffffu, is);
          if (is) {
            const int o = k + __popc(bal & ((1u << lane) - 1u));
            if (o < len) {
              buf[o] = ((CV[idx] & 255u) << 24) | (uint32_t)o;
              pos[o] = (uint32_t)idx;
            }
          }
          k += __popc(bal);
        }
      }
      __syncwarp();
      if (lane == 0) {
        if (k != len) atomicOr(A.status, kStatusSortDepth);  // the cell walk missed a member: never expected
        StdSortEmu emu{buf};
        const int sg = emu.leftmost_segment(len);
        uint32_t bst = buf[0];
        for (int i = 1; i < sg; i++)
          if (StdSortEmu::comp(buf[i], bst)) bst = buf[i];
        NV[m] = CV[pos[bst & 0xFFFFFFu]];
      }
      __syncwarp();
  }
}

// list order, then the first min(M, want) winners (:541-543), coordinates back in level space (:597-598)
__global__ void __launch_bounds__(kQtThreads, 3) k_qt_emit(QtArgs A) {
  pdl_wait();
  __shared__ QtSmem sm;
  const int level = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x;
  const int* meta = A.meta + ((size_t)frame * kMaxLevels + level) * 8;
  const int n = meta[kMetaN];
  if (n == 0 || meta[kMetaDeep]) return;
  const int D = meta[kMetaD], M = meta[kMetaM], level_base = meta[kMetaBase], want = A.quota[level];
  const size_t seg = (size_t)frame * A.raw_cap + level_base;
  uint32_t* NV = A.keys[1] + seg;
  uint32_t* cnt = A.vals[1] + seg;
  uint32_t* ties = A.keys[2] + seg;
  uint32_t* NK = A.vals[2] + seg;
  const int m_out = min(M, want);
  int4* sel = A.sel + (size_t)frame * A.n_quota + A.quota_base[level];
  constexpr int kSmemSort = kQtWarps * 256 / 2;  // 64-bit (order key, winner) pairs that fit the radix counters' space
  if (M <= kSmemSort) {
    // few nodes: bitonic sort of the pairs in shared memory (order keys are unique)
    unsigned long long* sk = reinterpret_cast<unsigned long long*>(&sm.warp_cnt[0][0]);
    int P = 2;
    while (P < M) P <<= 1;
    {
      // all of a thread's loads first (P <= 4 * kQtThreads): one round trip instead of one per element
      constexpr int kPer = kSmemSort / kQtThreads;
      uint32_t k4[kPer], v4[kPer];
#pragma unroll
      for (int k = 0; k < kPer; k++) {
        const int i = tid + k * kQtThreads;
        k4[k] = 0xFFFFFFFFu; v4[k] = 0xFFFFFFFFu;
        if (i < M) { k4[k] = NK[i]; v4[k] = NV[i]; }
      }
#pragma unroll
      for (int k = 0; k < kPer; k++) {
        const int i = tid + k * kQtThreads;
        if (i < P) sk[i] = ((unsigned long long)k4[k] << 32) | v4[k];
      }
    }
    __syncthreads();
    for (int k = 2; k <= P; k <<= 1)
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = tid; t < (P >> 1); t += kQtThreads) {
          const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), q = i | j;
          const unsigned long long a = sk[i], b = sk[q];
          if ((a > b) == ((i & k) == 0)) { sk[i] = b; sk[q] = a; }
        }
        // strides up to 32 keep a warp inside its own 64 elements (for every round of the t loop): the next such
        // stride only needs the warp's own writes
        if (j > 32 || j == 1) __syncthreads(); else __syncwarp();
      }
    for (int i = tid; i < m_out; i += kQtThreads) {
      const uint32_t r = (uint32_t)sk[i];
      sel[i] = make_int4((int)((r >> 8) & 4095u) + kBorder, (int)(r >> 20) + kBorder, (int)(r & 255u), level);
    }
  } else {
    const int ncur = qt_sort(sm, NK, NV, cnt, ties, M, 2 * D + 4);
    const uint32_t* OV = (ncur == 0) ? NV : ties;
    for (int i = tid; i < m_out; i += kQtThreads) {
      const uint32_t r = OV[i];
      sel[i] = make_int4((int)((r >> 8) & 4095u) + kBorder, (int)(r >> 20) + kBorder, (int)(r & 255u), level);
    }
  }
  if (tid == 0) A.sel_n[(size_t)frame * kMaxLevels + level] = m_out;
}

__global__ void __launch_bounds__(kQtThreads, 3) k_quadtree(QtArgs A) {
  pdl_wait();
  __shared__ QtSmem sm;
  const int level = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x;
  const FrameGeomDev* g = A.g;
  const LevelGeom lv = g->lv[level];
  const int* rank = A.cell_rank + (size_t)frame * (A.n_cells + 1);
  const int level_base = rank[lv.cell_base];
  const int level_end = rank[lv.cell_base + lv.cell_cols * lv.cell_rows];
  const int n = level_end - level_base;
  const int want = lv.quota;
  int* sel_n = A.sel_n + (size_t)frame * kMaxLevels + level;
  if (!A.meta[((size_t)frame * kMaxLevels + level) * 8 + kMetaDeep]) return;  // the node-table path did this segment
  if (n == 0 || want <= 0) {
    if (tid == 0) *sel_n = 0;
    return;
  }
  const size_t seg = (size_t)frame * A.raw_cap + level_base;
  uint32_t* K[3] = {A.keys[0] + seg, A.keys[1] + seg, A.keys[2] + seg};
  uint32_t* V[3] = {A.vals[0] + seg, A.vals[1] + seg, A.vals[2] + seg};

  // (1) path codes in candidate order + occupancy bitmap of their 8-digit prefixes
  for (int i = tid; i < kQtMapWords; i += kQtThreads) sm.bitmap[i] = 0;
  if (tid <= kQtMapDig) sm.cnt[tid] = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 4 * kQtThreads) {  // codes come from k_qt_codes
    uint32_t c4[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = base + u * kQtThreads + tid;
      c4[u] = i < n ? K[0][i] : 0xffffffffu;
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      if (c4[u] == 0xffffffffu) continue;  // codes have kQtBits = 28 significant bits
      const uint32_t p8 = c4[u] >> (kQtBits - 2 * kQtMapDig);
      const uint32_t bit = 1u << (p8 & 31u);
      if (!(sm.bitmap[p8 >> 5] & bit)) atomicOr(&sm.bitmap[p8 >> 5], bit);
    }
  }
  __syncthreads();
  // (2) #nodes after d passes = #distinct d-digit prefixes (d = 1..8): fold the bitmap.  Thread t < 256 owns the
  // 8 words under the 4-digit prefix t; deeper levels fold inside a word, shallower ones across lanes by ballot.
  {
    int c8 = 0, c7 = 0, c6 = 0, c5 = 0;
    bool any = false;
    if (tid < 256) {
      uint32_t pair = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const uint32_t w = sm.bitmap[tid * 8 + k];
        c8 += __popc(w);
        const uint32_t m7 = (w | (w >> 1) | (w >> 2) | (w >> 3)) & 0x11111111u;
        c7 += __popc(m7);
        c6 += __popc((m7 | (m7 >> 4) | (m7 >> 8) | (m7 >> 12)) & 0x00010001u);
        pair |= w;
        if (k & 1) { c5 += pair != 0; any |= pair != 0; pair = 0; }
      }
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, any);
    c8 = __reduce_add_sync(0xffffffffu, c8);
    c7 = __reduce_add_sync(0xffffffffu, c7);
    c6 = __reduce_add_sync(0xffffffffu, c6);
    c5 = __reduce_add_sync(0xffffffffu, c5);
    if ((tid & 31) == 0) {
      if (c8) { atomicAdd(&sm.cnt[8], c8); atomicAdd(&sm.cnt[7], c7); atomicAdd(&sm.cnt[6], c6); atomicAdd(&sm.cnt[5], c5); }
      if (bal) {
        atomicAdd(&sm.cnt[4], __popc(bal));
        atomicAdd(&sm.cnt[3], __popc((bal | (bal >> 1) | (bal >> 2) | (bal >> 3)) & 0x11111111u));
        atomicAdd(&sm.cnt[2], ((bal & 0xffffu) != 0) + ((bal >> 16) != 0));
      }
      sm.ballots[tid >> 5] = bal;
    }
  }
  __syncthreads();
  if (tid == 0) {
    int c1 = 0;
    for (int k = 0; k < 4; k++) c1 += (sm.ballots[2 * k] | sm.ballots[2 * k + 1]) != 0;
    sm.cnt[1] = c1;
    // the reference's loop condition (:479): split again while the list grew and is still shorter than `want`
    int D = 0, last = 0, size = 1, deep = 0;
    while (size > last && size < want && D < kQtDig) {
      last = size;
      D++;
      if (D > kQtMapDig) { deep = 1; break; }
      size = sm.cnt[D];
    }
    sm.misc[1] = D;
    sm.misc[3] = deep;
  }
  __syncthreads();
  int cur;
  if (sm.misc[3]) {
    // rare: more than 8 passes (tightly clustered points).  Sort the full codes, count distinct prefixes of every
    // depth from the common-prefix lengths of neighbours, then restore candidate order.
    __syncthreads();
    qt_fill(A, lv, frame, level_base, K[0], V[0], sm.bitmap);  // full kQtDig-digit codes
    __syncthreads();
    cur = qt_sort(sm, K[0], V[0], K[1], V[1], n, kQtBits);
    for (int d = tid; d < kQtDig + 2; d += kQtThreads) sm.lcp_hist[d] = 0;
    __syncthreads();
    {
      const uint32_t* ks = K[cur];
      for (int i = tid + 1; i < n; i += kQtThreads) atomicAdd(&sm.lcp_hist[lcp_digits(ks[i - 1], ks[i], kQtDig)], 1);
    }
    __syncthreads();
    if (tid == 0) {
      // cnt(d) = 1 + #adjacent pairs with lcp < d
      int D = 0, last = 0, size = 1, below = 0;
      while (size > last && size < want && D < kQtDig) {
        last = size;
        below += sm.lcp_hist[D];
        D++;
        size = 1 + below;
      }
      sm.misc[1] = D;
    }
    __syncthreads();
    qt_fill(A, lv, frame, level_base, K[0], V[0], sm.bitmap);
    __syncthreads();
  }
  const int D = sm.misc[1];
  const int sh = kQtBits - 2 * D;  // D-digit prefix of a code = code >> sh

  // (3) stable sort of the candidate-ordered points by their D-digit prefix (members of a node stay in candidate order)
  cur = (D == 0) ? 0 : qt_sort(sm, K[0], V[0], K[1], V[1], n, kQtBits, sh);
  const uint32_t* PK = K[cur];
  const uint32_t* PV = V[cur];
  uint32_t* node_start = K[cur ^ 1];   // free ping-pong buffer
  uint32_t* scratch = V[cur ^ 1];      // std::sort emulation workspace, indexed like the points

  // (4a) node boundaries
  if (tid == 0) sm.misc[2] = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 4 * kQtThreads) {  // thread = 4 consecutive points per CTA-wide scan
    const int i0 = base + 4 * tid;
    uint32_t pk[5];
#pragma unroll
    for (int u = 0; u < 5; u++) {
      const int i = i0 - 1 + u;
      pk[u] = (i >= 0 && i < n) ? (PK[i] >> sh) : 0xffffffffu;
    }
    int heads = 0;
    uint32_t hm = 0;
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = i0 + u;
      const bool head = (i < n) && (i == 0 || pk[u + 1] != pk[u]);
      hm |= (uint32_t)head << u;
      heads += head;
    }
    int total;
    int ex = qt_block_scan(sm, heads, &total);
    const int off = sm.misc[2];
#pragma unroll
    for (int u = 0; u < 4; u++)
      if ((hm >> u) & 1u) node_start[off + ex++] = (uint32_t)(i0 + u);
    __syncthreads();
    if (tid == 0) sm.misc[2] = off + total;
    __syncthreads();
  }
  const int M = sm.misc[2];

  // (4b) node records: order key + winner
  uint32_t* NK = K[2];
  uint32_t* NV = V[2];
  uint32_t* flagged = node_start + M;  // M + #nodes with > 16 members <= n, so the list fits behind node_start
  if (tid == 0) sm.misc[4] = 0;
  __syncthreads();
  for (int m = tid; m < M; m += kQtThreads) {
    const int beg = (int)node_start[m];
    const int end = (m + 1 < M) ? (int)node_start[m + 1] : n;
    const uint32_t pre = PK[beg] >> sh;
    int s = D;
    if (end - beg == 1) {
      int l = 0;
      if (m > 0) l = max(l, lcp_digits(pre, PK[node_start[m - 1]] >> sh, D));
      if (m + 1 < M) l = max(l, lcp_digits(pre, PK[end] >> sh, D));
      s = (n == 1) ? 0 : min(D, l + 1);
    }
    // best response, first maximum in candidate order, number of maxima
    uint32_t best_resp = 0, best_val = PV[beg];
    int ties = 0;
    for (int i = beg; i < end; i++) {
      const uint32_t v = PV[i];
      const uint32_t r = v & 255u;
      if (i == beg || r > best_resp) { best_resp = r; best_val = v; ties = 1; }
      else if (r == best_resp) ties++;
    }
    if (end - beg > 16 && ties > 1)  // introsort decides among tied maxima: queued for the warps below
      flagged[atomicAdd(&sm.misc[4], 1)] = (uint32_t)m;
    const uint32_t ps = (s == 0) ? 0u : (pre >> (2 * (D - s)));
    uint32_t mask = 0;
    for (int k = s; k >= 1; k -= 2) mask |= 3u << (2 * (s - k));
    NK[m] = ((uint32_t)(D - s) << (2 * D)) | (ps ^ mask);  // 2D + 4 significant bits
    NV[m] = best_val;
  }
  __syncthreads();
  // (4c) nodes whose winner std::sort decides: one warp per node, members packed (response << 24 | position) into the
  // warp's slice of shared memory (the radix counters are idle here), lane 0 runs the leftmost partition chain.
  {
    const int nflag = sm.misc[4];
    const int lane = tid & 31, warp = tid >> 5;
    for (int f = warp; f < nflag; f += kQtWarps) {
      const int m = (int)flagged[f];
      const int beg = (int)node_start[m];
      const int end = (m + 1 < M) ? (int)node_start[m + 1] : n;
      const int len = end - beg;
      uint32_t* buf = (len <= 256) ? reinterpret_cast<uint32_t*>(sm.warp_cnt[warp]) : scratch + beg;
      for (int i = lane; i < len; i += 32) buf[i] = ((PV[beg + i] & 255u) << 24) | (uint32_t)i;
      __syncwarp();
      if (lane == 0) {
        StdSortEmu emu{buf};
        const int seg = emu.leftmost_segment(len);
        uint32_t best = buf[0];
        for (int i = 1; i < seg; i++)
          if (StdSortEmu::comp(buf[i], best)) best = buf[i];
        NV[m] = PV[beg + (int)(best & 0xFFFFFFu)];
      }
      __syncwarp();
    }
  }
  __syncthreads();

  // (5) list order
  const int ncur = qt_sort(sm, NK, NV, K[cur ^ 1], V[cur ^ 1], M, 2 * D + 4);
  const uint32_t* OV = (ncur == 0) ? NV : V[cur ^ 1];

  // (6) emit the first min(M, want) winners (:541-543), coordinates back in level space (:597-598)
  const int m_out = min(M, want);
  int4* sel = A.sel + (size_t)frame * g->n_quota + lv.quota_base;
  for (int i = tid; i < m_out; i += kQtThreads) {
    const uint32_t r = OV[i];
    sel[i] = make_int4((int)((r >> 8) & 4095u) + kBorder, (int)(r >> 20) + kBorder, (int)(r & 255u), level);
  }
  if (tid == 0) *sel_n = m_out;
}

int launch_quadtree(const YdExtractor* ex, int batch, cudaStream_t s) {
  YD_CUDA(launch_chain(k_cell_scan, dim3(batch), dim3(256), 0, s, ex->geom.n_cells, ex->d_cell_count, ex->d_cell_rank));
  QtArgs A;
  A.g = ex->d_geom;
  A.cells = ex->d_cells;
  A.cell_count = ex->d_cell_count;
  A.cell_start = ex->d_cell_start;
  A.cell_rank = ex->d_cell_rank;
  A.raw = ex->d_raw;
  for (int i = 0; i < 3; i++) { A.keys[i] = ex->d_qt_keys[i]; A.vals[i] = ex->d_qt_vals[i]; }
  A.bitmap = ex->d_qt_bitmap;
  A.fold = ex->d_qt_fold;
  A.meta = ex->d_qt_meta;
  A.sel = ex->d_sel;
  A.sel_n = ex->d_sel_n;
  A.level_ncand = ex->d_level_ncand;
  A.status = ex->d_status;
  A.n_cells = ex->geom.n_cells;
  A.raw_cap = ex->geom.raw_cap;
  A.n_levels = ex->geom.n_levels;
  A.n_quota = ex->geom.n_quota;
  A.xpath = ex->d_qt_xpath;
  for (int l = 0; l < kMaxLevels; l++) {
    A.cell_base[l] = ex->geom.lv[l].cell_base;
    A.cell_end[l] = ex->geom.lv[l].cell_base + ex->geom.lv[l].cell_cols * ex->geom.lv[l].cell_rows;
    A.quota[l] = ex->geom.lv[l].quota;
    const LevelGeom& hl = ex->geom.lv[l];
    A.lv[l] = QtLevel{hl.cell_base, hl.cell_cols, hl.cell_rows, hl.cell_w, hl.cell_h, hl.root_w, hl.root_h};
    A.quota_base[l] = ex->geom.lv[l].quota_base;
    A.xoff[l] = ex->qt_xoff[l];
    A.xlen[l] = ex->qt_xlen[l];
  }
  if (ex->geom.n_cells > 0)
  {
    if ((long long)ex->geom.n_cells * batch >= 148 * 8 * 32) {  // measured: 0.276 -> 0.266 ms per 256 frames; a single frame prefers 32
      constexpr int cells = kCodeWarps * 32 / 8;
      YD_CUDA(launch_chain(k_qt_codes<8>, dim3((ex->geom.n_cells + cells - 1) / cells, batch), dim3(kCodeWarps * 32), 0, s, A));
    } else {
      YD_CUDA(launch_chain(k_qt_codes<32>, dim3((ex->geom.n_cells + kCodeWarps - 1) / kCodeWarps, batch), dim3(kCodeWarps * 32), 0, s, A));
    }
  }
  // candidate passes: grid-stride over the frame's candidates (their number is only known on the device); enough CTAs per
  // frame to fill the GPU when the batch is small, few when it is large
  const int per_frame = std::max(1, std::min((ex->geom.raw_cap + 255) / 256, (148 * 8 + batch - 1) / batch));
  dim3 grid(ex->geom.n_levels, batch), cgrid(per_frame, batch);
  YD_CUDA(launch_chain(k_qt_depth, grid, dim3(kQtThreads), 0, s, A));
  YD_CUDA(launch_chain(k_qt_pass1, cgrid, dim3(256), 0, s, A));
  YD_CUDA(launch_chain(k_qt_pass2, cgrid, dim3(256), 0, s, A));
  YD_CUDA(launch_chain(k_qt_records, grid, dim3(kQtThreads), 0, s, A));
  YD_CUDA(launch_chain(k_qt_ties, cgrid, dim3(kTieWarps * 32), 0, s, A));
  YD_CUDA(launch_chain(k_qt_emit, grid, dim3(kQtThreads), 0, s, A));
  YD_CUDA(launch_chain(k_quadtree, grid, dim3(kQtThreads), 0, s, A));  // segments deeper than kQtMapDig only
  count_launches(9);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
