// CPU oracle, part 3: SURVEY 8(f) row N1, DBoW3::Vocabulary::transform.  TEST INFRASTRUCTURE ONLY: imported by tests/,
// __graft_entry__.smoke() and bench.py's cpu legs, never by the product path.
//
// Follows thirdParty/DBow3/src/Vocabulary.cpp:752-816 (batch transform), :836-873 (tree descent),
// BowVector.cpp:34-84 (addWeight / addIfNotExist / normalize), FeatureVector.cpp:31-45 (addFeature) and
// DescManip.cpp:92-119 (binary distance), with the same std::map containers the reference uses, so iteration order
// and double-precision accumulation order are the reference's by construction.
// Parity status: DBoW3 cannot be compiled here (it needs OpenCV C++ for cv::Mat) and ships no golden vectors for
// transform => "parity unpinned"; an independent pure-Python restatement in tests/test_oracle_bow.py cross-checks it.
#include <climits>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <vector>

#include "ydorb_oracle.h"

namespace {

// DescManip::distance for CV_8U rows: SWAR popcount over uint64 words, returned as a double (DescManip.cpp:102-118)
double dbow_distance(const uint8_t* a, const uint8_t* b) {
  uint64_t ret = 0;
  for (size_t i = 0; i < 32 / sizeof(uint64_t); ++i) {
    uint64_t pa, pb;
    std::memcpy(&pa, a + 8 * i, 8);
    std::memcpy(&pb, b + 8 * i, 8);
    uint64_t v = pa ^ pb;
    v = v - ((v >> 1) & (uint64_t) ~(uint64_t)0 / 3);
    v = (v & (uint64_t) ~(uint64_t)0 / 15 * 3) + ((v >> 2) & (uint64_t) ~(uint64_t)0 / 15 * 3);
    v = (v + (v >> 4)) & (uint64_t) ~(uint64_t)0 / 255 * 15;
    ret += (uint64_t)(v * ((uint64_t) ~(uint64_t)0 / 255)) >> (sizeof(uint64_t) - 1) * CHAR_BIT;
  }
  return (double)ret;
}

}  // namespace

extern "C" {

// Vocabulary::transform(feature, word_id, weight, nid, levelsup)  (:836-873)
void orc_voc_descend(const OrcVocabulary* voc, const uint8_t* feature, int levels_up, int32_t* word_id, int32_t* node_id,
                     double* weight) {
  const int nid_level = voc->L - levels_up;  // :846
  int nid = 0;                               // caller's NodeId is uninitialised; :847 sets 0 when nid_level <= 0
  int final_id = 0, current_level = 0;       // :849-850
  do {
    ++current_level;
    double best_d = std::numeric_limits<double>::max();  // :856
    const int c0 = voc->child_start[final_id], c1 = voc->child_start[final_id + 1];
    for (int j = c0; j < c1; j++) {                      // :859
      const int id = voc->children[j];
      const double d = dbow_distance(feature, voc->descriptors + (size_t)id * 32);
      if (d < best_d) {
        best_d = d;
        final_id = id;
      }
    }
    if (current_level == nid_level) nid = final_id;      // :866-867
  } while (voc->child_start[final_id + 1] > voc->child_start[final_id]);  // !isLeaf()
  *word_id = voc->word_ids[final_id];  // :870
  *weight = voc->weights[final_id];    // :871
  *node_id = nid;
}

// Vocabulary::transform(features, v, fv, levelsup)  (:752-816).  Outputs as in include/ydorb.h YdBowResult.
void orc_voc_transform(const OrcVocabulary* voc, const uint8_t* desc, int n, int levels_up, int32_t* word_id, int32_t* node_id,
                       double* weight, int32_t* bow_ids, double* bow_values, int32_t* n_bow, int32_t* fv_node_ids,
                       int32_t* fv_node_start, int32_t* fv_items, int32_t* n_fv_nodes) {
  std::map<unsigned, double> v;                   // BowVector
  std::map<unsigned, std::vector<unsigned>> fv;   // FeatureVector
  const bool must = voc->scoring != 5;            // DotProductScoring is the only one that does not normalise
  const bool l2 = voc->scoring == 1;              // L2Scoring; all others L1 (ScoringObject.h:72-88)
  const bool tf_like = voc->weighting == 0 || voc->weighting == 1;  // TF_IDF || TF  (:771)
  for (int i = 0; i < n; i++) {
    int32_t id, nid;
    double w;
    orc_voc_descend(voc, desc + (size_t)i * 32, levels_up, &id, &nid, &w);
    if (word_id) word_id[i] = id;
    if (node_id) node_id[i] = nid;
    if (weight) weight[i] = w;
    if (w > 0) {  // not stopped (:781, :811)
      if (tf_like) {
        auto vit = v.lower_bound((unsigned)id);  // BowVector::addWeight
        if (vit != v.end() && !(v.key_comp()((unsigned)id, vit->first))) vit->second += w;
        else v.insert(vit, std::make_pair((unsigned)id, w));
      } else {
        auto vit = v.lower_bound((unsigned)id);  // BowVector::addIfNotExist
        if (vit == v.end() || v.key_comp()((unsigned)id, vit->first)) v.insert(vit, std::make_pair((unsigned)id, w));
      }
      fv[(unsigned)nid].push_back((unsigned)i);  // FeatureVector::addFeature
    }
  }
  if (tf_like && !v.empty() && !must) {  // :787-793
    const double nd = (double)v.size();
    for (auto& kv : v) kv.second /= nd;
  }
  if (must) {  // BowVector::normalize (BowVector.cpp:62-84)
    double norm = 0.0;
    if (!l2) {
      for (auto& kv : v) norm += std::fabs(kv.second);
    } else {
      for (auto& kv : v) norm += kv.second * kv.second;
      norm = std::sqrt(norm);
    }
    if (norm > 0.0)
      for (auto& kv : v) kv.second /= norm;
  }
  int k = 0;
  for (auto& kv : v) {
    bow_ids[k] = (int32_t)kv.first;
    bow_values[k] = kv.second;
    k++;
  }
  *n_bow = k;
  int s = 0, p = 0;
  for (auto& kv : fv) {
    fv_node_ids[s] = (int32_t)kv.first;
    fv_node_start[s] = p;
    for (unsigned f : kv.second) fv_items[p++] = (int32_t)f;
    s++;
  }
  fv_node_start[s] = p;
  *n_fv_nodes = s;
}

}  // extern "C"
