// DBoW3::Vocabulary::transform on the device (the step Frame::computeBoW runs right before both BoW
// matchers, src/frame.cpp:265-272).
//   descent   thirdParty/DBow3/src/Vocabulary.cpp:836-873  -> k_voc_descend   (thread = descriptor)
//   assembly  thirdParty/DBow3/src/Vocabulary.cpp:752-816, BowVector.cpp:34-84, FeatureVector.cpp:31-45
//                                                            -> k_voc_assemble  (CTA = frame)
// The two std::map containers become sorted arrays: BowVector = (word ids ascending, values), FeatureVector = CSR
// (node ids ascending, features in push_back order = ascending feature index), the form ydorb_search_by_bow consumes.
// Word values are doubles accumulated in exactly the reference's order (feature order inside a word, ascending word
// order for the norm), so they are bit-identical, not merely close.
#include "vocabulary_stream.h"
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

#include "matcher.cuh"

struct YdVocabulary {
  YdMatcher* m;
  int32_t n_nodes, n_children, k, L, weighting, scoring, n_words;
  int32_t* d_child_start;  // [n_nodes + 1]
  int32_t* d_child_node;   // [n_children] node id of the j-th child slot (CSR order)
  uint4* d_child_desc;     // [n_children][2]  descriptor of the j-th child slot: siblings are contiguous (k x 32 B)
  double* d_weight;        // [n_nodes]
  int32_t* d_word;         // [n_nodes]
};

namespace ydorb {

struct VocDev {
  const int32_t* child_start;
  const int32_t* child_node;
  const uint4* child_desc;
  const double* weight;
  const int32_t* word;
  int L, weighting, norm;  // norm: 0 none, 1 L1, 2 L2 (ScoringObject.h:72-88)
};

// One thread walks one descriptor from the root to a leaf.  Distances are the 256-bit Hamming distance
// (DescManip::distance, DescManip.cpp:92-119, returns it as a double; comparing the integers is the same order).
__global__ void __launch_bounds__(128) k_voc_descend(const VocDev v, const uint4* __restrict__ desc, const int32_t* __restrict__ n_per_frame,
                                                     int capacity, int batch, int nid_level, int32_t* __restrict__ word_id,
                                                     int32_t* __restrict__ node_id, double* __restrict__ weight) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= capacity * batch) return;
  const int b = t / capacity, i = t - b * capacity;
  if (i >= n_per_frame[b]) return;
  const uint4 lo = __ldg(desc + 2 * (size_t)t), hi = __ldg(desc + 2 * (size_t)t + 1);
  int node = 0, level = 0, nid = 0;  // root; *nid = 0 when nid_level <= 0 (:847)
  int c0 = __ldg(v.child_start), c1 = __ldg(v.child_start + 1);
  while (c1 > c0) {  // do { ... } while (!isLeaf)  (:852-868); a childless root never occurs (empty() returns earlier)
    ++level;
    int best_d = 0x7fffffff, best_j = c0;
    for (int j = c0; j < c1; j++) {
      const uint4 a = __ldg(v.child_desc + 2 * (size_t)j), c = __ldg(v.child_desc + 2 * (size_t)j + 1);
      const int d = __popc(a.x ^ lo.x) + __popc(a.y ^ lo.y) + __popc(a.z ^ lo.z) + __popc(a.w ^ lo.w) + __popc(c.x ^ hi.x) +
                    __popc(c.y ^ hi.y) + __popc(c.z ^ hi.z) + __popc(c.w ^ hi.w);
      if (d < best_d) {  // strict: the first child wins ties (:859)
        best_d = d;
        best_j = j;
      }
    }
    node = __ldg(v.child_node + best_j);
    if (level == nid_level) nid = node;  // :866
    c0 = __ldg(v.child_start + node);
    c1 = __ldg(v.child_start + node + 1);
  }
  word_id[t] = __ldg(v.word + node);  // :870
  node_id[t] = nid;
  weight[t] = __ldg(v.weight + node);  // :871
}

constexpr int kAsmThreads = 1024;

// in-place bitonic sort of P (a power of two) 64-bit keys in shared memory; keys are unique, so stability is moot
__device__ void bitonic_sort(unsigned long long* keys, int P) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (P >> 1); t += kAsmThreads) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), p = i | j;
        const unsigned long long a = keys[i], b = keys[p];
        const bool up = (i & k) == 0;
        if ((a > b) == up) {
          keys[i] = b;
          keys[p] = a;
        }
      }
      __syncthreads();
    }
  }
}

// exclusive scan of one int per thread over the CTA; returns the prefix, *total = sum over all threads
__device__ int block_excl_scan(int x, int* warp_tot, int* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = x;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += y;
  }
  __syncthreads();  // warp_tot may still be read from the previous call
  if (lane == 31) warp_tot[w] = inc;
  __syncthreads();
  if (w == 0) {
    int s = warp_tot[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += y;
    }
    warp_tot[lane] = s;  // inclusive
  }
  __syncthreads();
  *total = warp_tot[31];
  return inc - x + (w ? warp_tot[w - 1] : 0);
}

__device__ __forceinline__ uint32_t key_hi(unsigned long long k) { return (uint32_t)(k >> 32); }
__device__ __forceinline__ uint32_t key_lo(unsigned long long k) { return (uint32_t)k; }

// CTA = one frame.  Two passes over the same machinery: sort (id << 32 | feature) keys, find the runs of equal ids,
// number them with a CTA-wide scan.  Pass A (ids = node ids) writes the FeatureVector, pass B (ids = word ids) the
// BowVector.  P = keys buffer length (power of two >= capacity); dynamic smem = P * 8 B.
__global__ void __launch_bounds__(kAsmThreads) k_voc_assemble(const VocDev v, const int32_t* __restrict__ n_per_frame, int capacity, int P,
                                                              const int32_t* __restrict__ word_id, const int32_t* __restrict__ node_id,
                                                              const double* __restrict__ weight, int32_t* __restrict__ bow_ids,
                                                              double* __restrict__ bow_values, int32_t* __restrict__ fv_node_ids,
                                                              int32_t* __restrict__ fv_node_start, int32_t* __restrict__ fv_items,
                                                              int32_t* __restrict__ counts) {
  extern __shared__ unsigned long long keys[];
  __shared__ int warp_tot[32];
  __shared__ int s_kept;
  __shared__ double s_norm;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int n = n_per_frame[b];
  const size_t base = (size_t)b * capacity;
  word_id += base; node_id += base; weight += base;
  bow_ids += base; bow_values += base; fv_node_ids += base; fv_items += base;
  fv_node_start += (size_t)b * (capacity + 1);
  const int C = P > kAsmThreads ? P / kAsmThreads : 1;
  const int p0 = min(tid * C, P), p1 = min(p0 + C, P);
  const unsigned long long kNone = ~0ull;

  int n_seg[2];
  for (int pass = 0; pass < 2; pass++) {
    const int32_t* ids = pass == 0 ? node_id : word_id;
    if (tid == 0) s_kept = 0;
    __syncthreads();
    int kept = 0;
    for (int i = tid; i < P; i += kAsmThreads) {
      const bool keep = i < n && weight[i] > 0.0;  // if (w > 0) // not stopped  (Vocabulary.cpp:781, :811)
      keys[i] = keep ? ((unsigned long long)(uint32_t)ids[i] << 32) | (uint32_t)i : kNone;
      kept += keep;
    }
    if (kept) atomicAdd(&s_kept, kept);
    __syncthreads();
    const int m = s_kept;
    bitonic_sort(keys, P);
    int heads = 0;
    for (int p = p0; p < p1; p++) heads += p < m && (p == 0 || key_hi(keys[p]) != key_hi(keys[p - 1]));
    int total;
    int seg = block_excl_scan(heads, warp_tot, &total);
    n_seg[pass] = total;
    for (int p = p0; p < p1; p++) {
      if (p >= m) break;
      const unsigned long long k = keys[p];
      const bool head = p == 0 || key_hi(k) != key_hi(keys[p - 1]);
      if (pass == 0) {
        fv_items[p] = (int32_t)key_lo(k);  // FeatureVector::addFeature push_back order = ascending feature index
        if (head) {
          fv_node_ids[seg] = (int32_t)key_hi(k);
          fv_node_start[seg] = p;
          seg++;
        }
      } else if (head) {
        // BowVector::addWeight (TF, TF_IDF: first insert w, then += w per further feature, in feature order) or
        // addIfNotExist (IDF, BINARY: the first value stays)
        double acc = weight[key_lo(k)];
        if (v.weighting <= 1)
          for (int q = p + 1; q < m && key_hi(keys[q]) == key_hi(k); q++) acc = __dadd_rn(acc, weight[key_lo(keys[q])]);
        bow_ids[seg] = (int32_t)key_hi(k);
        bow_values[seg] = acc;
        seg++;
      }
    }
    if (pass == 0 && tid == 0) fv_node_start[total] = m;
    __syncthreads();
  }

  // normalisation: BowVector::normalize (BowVector.cpp:62-84) sums in ascending word order; one thread does the same
  const int nb = n_seg[1];
  double* vals = reinterpret_cast<double*>(keys);
  for (int s = tid; s < nb; s += kAsmThreads) vals[s] = bow_values[s];
  __syncthreads();
  if (v.norm != 0) {
    if (tid == 0) {
      double norm = 0.0;
      if (v.norm == 1) {
        for (int s = 0; s < nb; s++) norm = __dadd_rn(norm, fabs(vals[s]));
      } else {
        for (int s = 0; s < nb; s++) norm = __dadd_rn(norm, __dmul_rn(vals[s], vals[s]));
        norm = __dsqrt_rn(norm);
      }
      s_norm = norm;
    }
    __syncthreads();
    const double norm = s_norm;
    if (norm > 0.0)
      for (int s = tid; s < nb; s += kAsmThreads) bow_values[s] = __ddiv_rn(vals[s], norm);
  } else if (v.weighting <= 1 && nb > 0) {
    const double nd = (double)nb;  // Vocabulary.cpp:787-793: unnecessary when normalizing
    for (int s = tid; s < nb; s += kAsmThreads) bow_values[s] = __ddiv_rn(vals[s], nd);
  }
  if (tid == 0) {
    counts[2 * b] = nb;
    counts[2 * b + 1] = n_seg[0];
  }
}

}  // namespace ydorb

using namespace ydorb;

namespace {

int next_pow2(int x) {
  int p = 1;
  while (p < x) p <<= 1;
  return p;
}

VocDev dev_view(const YdVocabulary* v) {
  VocDev d;
  d.child_start = v->d_child_start;
  d.child_node = v->d_child_node;
  d.child_desc = v->d_child_desc;
  d.weight = v->d_weight;
  d.word = v->d_word;
  d.L = v->L;
  d.weighting = v->weighting;
  d.norm = v->scoring == 5 ? 0 : (v->scoring == 1 ? 2 : 1);  // ScoringObject.h:72-88
  return d;
}

void free_vocabulary(YdVocabulary* v) {
  if (!v) return;
  cudaFree(v->d_child_start);
  cudaFree(v->d_child_node);
  cudaFree(v->d_child_desc);
  cudaFree(v->d_weight);
  cudaFree(v->d_word);
  delete v;
}

}  // namespace

extern "C" {

int ydorb_vocabulary_create(YdMatcher* m, const YdVocabularyDesc* desc, YdVocabulary** out) {
  YD_REQUIRE(m && desc && out, "null argument");
  *out = nullptr;
  const int nn = desc->n_nodes;
  YD_REQUIRE(nn >= 2, "a vocabulary needs a root and at least one word (DBoW3 empty() vocabularies transform to nothing)");
  YD_REQUIRE(desc->child_start && desc->children && desc->descriptors && desc->weights && desc->word_ids, "null array");
  YD_REQUIRE(desc->weighting >= 0 && desc->weighting <= 3, "weighting must be a DBoW3::WeightingType (0..3)");
  YD_REQUIRE(desc->scoring >= 0 && desc->scoring <= 5, "scoring must be a DBoW3::ScoringType (0..5)");
  YD_REQUIRE(desc->child_start[0] == 0 && desc->child_start[nn] == nn - 1, "child_start must cover exactly n_nodes-1 children");
  YD_REQUIRE(desc->child_start[1] > 0, "the root has no children");
  // every node but the root must be some node's child exactly once, and the child graph must be a tree below the root
  std::vector<uint8_t> seen(nn, 0);
  int n_words = 0;
  for (int i = 0; i < nn; i++) {
    YD_REQUIRE(desc->child_start[i + 1] >= desc->child_start[i], "child_start must be non-decreasing");
    if (desc->child_start[i + 1] == desc->child_start[i]) n_words++;
  }
  for (int j = 0; j < nn - 1; j++) {
    const int c = desc->children[j];
    YD_REQUIRE(c > 0 && c < nn && !seen[c], "children must list every non-root node exactly once");
    seen[c] = 1;
  }
  {  // reachability from the root (rules out cycles detached from it)
    std::vector<int> stack(1, 0);
    int reached = 0;
    while (!stack.empty()) {
      const int p = stack.back();
      stack.pop_back();
      reached++;
      for (int j = desc->child_start[p]; j < desc->child_start[p + 1]; j++) stack.push_back(desc->children[j]);
    }
    YD_REQUIRE(reached == nn, "some nodes are not reachable from the root");
  }
  YD_CUDA(cudaSetDevice(m->device));
  YdVocabulary* v = new YdVocabulary();
  std::memset(v, 0, sizeof(*v));
  v->m = m;
  v->n_nodes = nn;
  v->n_children = nn - 1;
  v->k = desc->k;
  v->L = desc->L;
  v->weighting = desc->weighting;
  v->scoring = desc->scoring;
  v->n_words = n_words;
  // sibling descriptors gathered into child-slot order: one node's k children are k x 32 contiguous bytes
  std::vector<uint8_t> slot_desc((size_t)(nn - 1) * 32);
  for (int j = 0; j < nn - 1; j++) std::memcpy(&slot_desc[(size_t)j * 32], desc->descriptors + (size_t)desc->children[j] * 32, 32);
  cudaError_t e = cudaSuccess;
  auto up = [&](void** d, const void* h, size_t bytes) {
    if (e != cudaSuccess) return;
    e = cudaMalloc(d, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(*d, h, bytes, cudaMemcpyHostToDevice);
  };
  up((void**)&v->d_child_start, desc->child_start, sizeof(int32_t) * (size_t)(nn + 1));
  up((void**)&v->d_child_node, desc->children, sizeof(int32_t) * (size_t)(nn - 1));
  up((void**)&v->d_child_desc, slot_desc.data(), slot_desc.size());
  up((void**)&v->d_weight, desc->weights, sizeof(double) * (size_t)nn);
  up((void**)&v->d_word, desc->word_ids, sizeof(int32_t) * (size_t)nn);
  if (e != cudaSuccess) {
    free_vocabulary(v);
    return cuda_fail(e, "vocabulary upload", __FILE__, __LINE__);
  }
  *out = v;
  return YDORB_OK;
}

void ydorb_vocabulary_destroy(YdVocabulary* v) { free_vocabulary(v); }

int ydorb_vocabulary_info(const YdVocabulary* v, int32_t* info6) {
  YD_REQUIRE(v && info6, "null argument");
  info6[0] = v->n_nodes;
  info6[1] = v->k;
  info6[2] = v->L;
  info6[3] = v->weighting;
  info6[4] = v->scoring;
  info6[5] = v->n_words;
  return YDORB_OK;
}

int ydorb_vocabulary_stream_inflate(const uint8_t* stream, size_t n, uint8_t* out, size_t capacity, size_t* out_len) {
  YD_REQUIRE(stream && out_len, "null argument");
  *out_len = 0;
  uint64_t sig = 0;
  YD_REQUIRE(n >= 13, "truncated vocabulary header");
  std::memcpy(&sig, stream, 8);
  YD_REQUIRE(sig == 88877711233ull, "not a DBoW3 binary vocabulary (bad signature)");
  std::vector<uint8_t> payload;
  const uint8_t* body = stream + 13;
  size_t body_len = n - 13;
  if (stream[8]) {
    std::string why;
    YD_REQUIRE(dbow3_inflate_chunks(body, body_len, payload, &why), "compressed vocabulary stream: " + why);
    body = payload.data();
    body_len = payload.size();
  }
  *out_len = 13 + body_len;
  if (!out) return YDORB_OK;
  YD_REQUIRE(capacity >= *out_len, "output buffer too small");
  std::memcpy(out, stream, 13);
  out[8] = 0;
  if (body_len) std::memcpy(out + 13, body, body_len);
  return YDORB_OK;
}

int ydorb_vocabulary_load(YdMatcher* m, const char* path, YdVocabulary** out) {
  YD_REQUIRE(m && path && out, "null argument");
  *out = nullptr;
  FILE* f = std::fopen(path, "rb");
  YD_REQUIRE(f, std::string("cannot open vocabulary file ") + path);
  std::vector<uint8_t> buf;
  {
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    buf.resize(sz > 0 ? (size_t)sz : 0);
    const size_t got = buf.empty() ? 0 : std::fread(buf.data(), 1, buf.size(), f);
    std::fclose(f);
    YD_REQUIRE(got == buf.size(), "short read of the vocabulary file");
  }
  const std::vector<uint8_t>* src = &buf;
  size_t pos = 0;
  bool ok = true;
  auto rd = [&](void* dst, size_t n) {
    if (!ok || pos + n > src->size()) {
      ok = false;
      std::memset(dst, 0, n);
      return;
    }
    std::memcpy(dst, src->data() + pos, n);
    pos += n;
  };
  uint64_t sig = 0;
  uint8_t compressed = 0;
  uint32_t nnodes = 0;
  rd(&sig, 8);                                                                          // Vocabulary.cpp:1340-1342
  YD_REQUIRE(ok && sig == 88877711233ull, "not a DBoW3 binary vocabulary (bad signature)");
  rd(&compressed, 1);                                                                   // :1343-1344 (sizeof(bool) = 1)
  rd(&nnodes, 4);                                                                       // :1345-1346
  YD_REQUIRE(ok, "truncated vocabulary header");
  YD_REQUIRE(nnodes >= 2 && nnodes < (1u << 30), "vocabulary has no words");
  std::vector<uint8_t> payload;
  if (compressed) {  // :1350-1367: 10 000-byte chunks through QuickLZ (level 1) -- what save() writes by default
    std::string why;
    YD_REQUIRE(dbow3_inflate_chunks(buf.data() + pos, buf.size() - pos, payload, &why), "compressed vocabulary stream: " + why);
    src = &payload;
    pos = 0;
  }
  int32_t k = 0, L = 0, scoring = 0, weighting = 0;
  rd(&k, 4); rd(&L, 4); rd(&scoring, 4); rd(&weighting, 4);                             // :1371-1374
  std::vector<int32_t> parent(nnodes, -1), word(nnodes, 0), order;                      // Node(): word_id(0)
  std::vector<double> weight(nnodes, 0.0);
  std::vector<uint8_t> descs((size_t)nnodes * 32, 0);
  order.reserve(nnodes - 1);
  for (uint32_t i = 1; i < nnodes && ok; i++) {                                         // :1382-1393
    uint32_t nid = 0, par = 0;
    int32_t cols = 0, rows = 0, type = 0;
    double w = 0;
    rd(&nid, 4); rd(&par, 4); rd(&w, 8);
    rd(&cols, 4); rd(&rows, 4); rd(&type, 4);                                           // DescManip.cpp:260-267
    YD_REQUIRE(ok, "truncated vocabulary node table");
    YD_REQUIRE(nid > 0 && nid < nnodes && par < nnodes && parent[nid] < 0, "corrupt node record");
    YD_REQUIRE(type == 0 && cols == 32 && rows == 1, "only 1x32 CV_8U (ORB) node descriptors are supported");
    rd(&descs[(size_t)nid * 32], 32);
    parent[nid] = (int32_t)par;
    weight[nid] = w;
    order.push_back((int32_t)nid);
  }
  uint32_t nwords = 0;
  rd(&nwords, 4);                                                                       // :1395-1397
  for (uint32_t i = 0; i < nwords && ok; i++) {                                         // :1398-1405
    uint32_t wid = 0, nid = 0;
    rd(&wid, 4); rd(&nid, 4);
    YD_REQUIRE(ok && nid < nnodes, "corrupt word table");
    word[nid] = (int32_t)wid;
  }
  YD_REQUIRE(ok, "truncated vocabulary file");
  // children in push_back order (= stream order, :1392)
  std::vector<int32_t> child_start(nnodes + 1, 0), children(nnodes - 1);
  for (int32_t c : order) child_start[parent[c] + 1]++;
  for (uint32_t i = 0; i < nnodes; i++) child_start[i + 1] += child_start[i];
  {
    std::vector<int32_t> fill(child_start.begin(), child_start.end() - 1);
    for (int32_t c : order) children[fill[parent[c]]++] = c;
  }
  YdVocabularyDesc d;
  d.n_nodes = (int32_t)nnodes;
  d.k = k; d.L = L; d.weighting = weighting; d.scoring = scoring;
  d.child_start = child_start.data();
  d.children = children.data();
  d.descriptors = descs.data();
  d.weights = weight.data();
  d.word_ids = word.data();
  return ydorb_vocabulary_create(m, &d, out);
}

int ydorb_vocabulary_transform_batch_device(YdVocabulary* v, const uint8_t* d_desc, const int32_t* d_n, int capacity,
                                            int batch, int levels_up, int32_t* d_word_id, int32_t* d_node_id,
                                            double* d_weight, int32_t* d_bow_ids, double* d_bow_values,
                                            int32_t* d_fv_node_ids, int32_t* d_fv_node_start, int32_t* d_fv_items,
                                            int32_t* d_counts) {
  YD_REQUIRE(v && d_n && batch >= 0 && capacity >= 0, "bad argument");
  if (batch == 0) return YDORB_OK;
  YD_REQUIRE(capacity >= 1 && capacity <= YDORB_VOC_MAX_FEATURES, "capacity must be in [1, YDORB_VOC_MAX_FEATURES]");
  YD_REQUIRE((long long)capacity * batch < (1ll << 31), "batch x capacity overflows");
  YD_REQUIRE(d_desc && d_bow_ids && d_bow_values && d_fv_node_ids && d_fv_node_start && d_fv_items && d_counts, "null output array");
  YD_REQUIRE(((uintptr_t)d_desc & 15) == 0, "descriptor block must be 16-byte aligned");
  YdMatcher* m = v->m;
  YD_CUDA(cudaSetDevice(m->device));
  const size_t total = (size_t)capacity * batch;
  if (!d_word_id || !d_node_id || !d_weight) {  // per-feature results not wanted by the caller: keep them in scratch
    const size_t wb = (sizeof(double) * total + 255) / 256 * 256, ib = (sizeof(int32_t) * total + 255) / 256 * 256;
    int rc = matcher_reserve(m, wb + 2 * ib);
    if (rc != YDORB_OK) return rc;
    uint8_t* s = reinterpret_cast<uint8_t*>(m->d_scratch);
    if (!d_weight) d_weight = reinterpret_cast<double*>(s);
    if (!d_word_id) d_word_id = reinterpret_cast<int32_t*>(s + wb);
    if (!d_node_id) d_node_id = reinterpret_cast<int32_t*>(s + wb + ib);
  }
  const VocDev dv = dev_view(v);
  const int P = next_pow2(capacity);
  const size_t smem = sizeof(unsigned long long) * (size_t)P;
  static thread_local int attr_device = -1;
  if (attr_device != m->device) {
    YD_CUDA(cudaFuncSetAttribute(k_voc_assemble, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)(sizeof(unsigned long long) * YDORB_VOC_MAX_FEATURES)));
    attr_device = m->device;
  }
  k_voc_descend<<<(unsigned)((total + 127) / 128), 128, 0, m->stream>>>(dv, reinterpret_cast<const uint4*>(d_desc), d_n, capacity, batch,
                                                                         v->L - levels_up, d_word_id, d_node_id, d_weight);
  k_voc_assemble<<<batch, kAsmThreads, smem, m->stream>>>(dv, d_n, capacity, P, d_word_id, d_node_id, d_weight, d_bow_ids,
                                                          d_bow_values, d_fv_node_ids, d_fv_node_start, d_fv_items, d_counts);
  count_launches(2);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

int ydorb_vocabulary_transform(YdVocabulary* v, const uint8_t* descriptors, int n, int levels_up, YdBowResult* out) {
  YD_REQUIRE(v && out && n >= 0, "bad argument");
  out->n_bow = 0;
  out->n_fv_nodes = 0;
  if (n == 0) {  // v.clear(); fv.clear(); nothing added
    if (out->fv_node_start) out->fv_node_start[0] = 0;
    return YDORB_OK;
  }
  YD_REQUIRE(n <= YDORB_VOC_MAX_FEATURES, "more than YDORB_VOC_MAX_FEATURES descriptors in one frame");
  YD_REQUIRE(descriptors && out->bow_ids && out->bow_values && out->fv_node_ids && out->fv_node_start && out->fv_items,
             "null array in YdBowResult");
  YdMatcher* m = v->m;
  YD_CUDA(cudaSetDevice(m->device));
  // io block: desc | n | word | node | weight | bow_ids | bow_values | fv_ids | fv_start | fv_items | counts
  auto al = [](size_t b) { return (b + 255) / 256 * 256; };
  const size_t N = (size_t)n;
  size_t o[12];
  const size_t sizes[11] = {N * 32, 4, N * 4, N * 4, N * 8, N * 4, N * 8, N * 4, (N + 1) * 4, N * 4, 8};
  o[0] = 0;
  for (int i = 0; i < 11; i++) o[i + 1] = o[i] + al(sizes[i]);
  int rc = matcher_reserve_io(m, o[11]);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YD_CUDA(cudaMemcpyAsync(io + o[0], descriptors, N * 32, cudaMemcpyHostToDevice, m->stream));
  YD_CUDA(cudaMemcpyAsync(io + o[1], &n, 4, cudaMemcpyHostToDevice, m->stream));
  rc = ydorb_vocabulary_transform_batch_device(
      v, io + o[0], reinterpret_cast<int32_t*>(io + o[1]), n, 1, levels_up, reinterpret_cast<int32_t*>(io + o[2]),
      reinterpret_cast<int32_t*>(io + o[3]), reinterpret_cast<double*>(io + o[4]), reinterpret_cast<int32_t*>(io + o[5]),
      reinterpret_cast<double*>(io + o[6]), reinterpret_cast<int32_t*>(io + o[7]), reinterpret_cast<int32_t*>(io + o[8]),
      reinterpret_cast<int32_t*>(io + o[9]), reinterpret_cast<int32_t*>(io + o[10]));
  if (rc != YDORB_OK) return rc;
  int32_t counts[2] = {0, 0};
  YD_CUDA(cudaMemcpyAsync(counts, io + o[10], 8, cudaMemcpyDeviceToHost, m->stream));
  if (out->word_id) YD_CUDA(cudaMemcpyAsync(out->word_id, io + o[2], N * 4, cudaMemcpyDeviceToHost, m->stream));
  if (out->node_id) YD_CUDA(cudaMemcpyAsync(out->node_id, io + o[3], N * 4, cudaMemcpyDeviceToHost, m->stream));
  if (out->weight) YD_CUDA(cudaMemcpyAsync(out->weight, io + o[4], N * 8, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaMemcpyAsync(out->bow_ids, io + o[5], N * 4, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaMemcpyAsync(out->bow_values, io + o[6], N * 8, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaMemcpyAsync(out->fv_node_ids, io + o[7], N * 4, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaMemcpyAsync(out->fv_node_start, io + o[8], (N + 1) * 4, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaMemcpyAsync(out->fv_items, io + o[9], N * 4, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  out->n_bow = counts[0];
  out->n_fv_nodes = counts[1];
  return YDORB_OK;
}

}  // extern "C"
