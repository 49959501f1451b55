/* oracle/_ref: C entry points around the REFERENCE's own DBoW3::Vocabulary (thirdParty/DBow3/src/Vocabulary.cpp and
 * DescManip.cpp, compiled unmodified against the OpenCV stand-in).  TEST INFRASTRUCTURE: only tests/ call this.
 * It builds a vocabulary with the reference's own k-means (Vocabulary::create), writes / reads DBoW3's binary stream
 * (Vocabulary::save / load -> toStream / fromStream) and runs Vocabulary::transform the way Frame::computeBoW does
 * (src/frame.cpp:265-272: one 1 x 32 cv::Mat per descriptor, levelsup = 4). */
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "DBoW3/DBoW3.h"

namespace {
std::vector<cv::Mat> rows(const uint8_t* desc, int n) {
  std::vector<cv::Mat> out(n);
  for (int i = 0; i < n; i++) {
    out[i].create(1, 32, CV_8UC1);
    memcpy(out[i].ptr(0), desc + (size_t)32 * i, 32);
  }
  return out;
}
/* the per-feature overload is protected: reachable from a derived class; nothing is overridden */
struct ProbeVocabulary : DBoW3::Vocabulary {
  ProbeVocabulary() {}
  ProbeVocabulary(int k, int L, DBoW3::WeightingType w, DBoW3::ScoringType s) : DBoW3::Vocabulary(k, L, w, s) {}
  void one(const cv::Mat& f, DBoW3::WordId& id, DBoW3::WordValue& w, DBoW3::NodeId* nid, int levelsup) const {
    transform(f, id, w, nid, levelsup);
  }
};
}  // namespace

extern "C" {

/* Vocabulary(k, L, weighting, scoring) + create(training features): image i owns descriptors [start[i], start[i+1]) */
void* ref_voc_create(int k, int L, int weighting, int scoring, const uint8_t* desc, const int32_t* image_start, int n_images,
                     unsigned seed) {
  srand(seed);  // the k-means++ seeding draws from rand() (Vocabulary.cpp:431, :466)
  ProbeVocabulary* v = new ProbeVocabulary(k, L, (DBoW3::WeightingType)weighting, (DBoW3::ScoringType)scoring);
  std::vector<std::vector<cv::Mat>> training(n_images);
  for (int i = 0; i < n_images; i++) training[i] = rows(desc + (size_t)32 * image_start[i], image_start[i + 1] - image_start[i]);
  v->create(training);
  return v;
}

/* Vocabulary::load(filename): binary stream first, as the reference's constructor from a file name does */
void* ref_voc_load(const char* path) {
  ProbeVocabulary* v = new ProbeVocabulary();
  try {
    v->load(std::string(path));
  } catch (...) {
    delete v;
    return nullptr;
  }
  return v;
}

int ref_voc_save(void* p, const char* path, int compressed) {
  try {
    static_cast<ProbeVocabulary*>(p)->save(std::string(path), compressed != 0);
  } catch (...) {
    return -1;
  }
  return 0;
}

void ref_voc_info(void* p, int32_t* out) {  // k, L, weighting, scoring, words
  ProbeVocabulary* v = static_cast<ProbeVocabulary*>(p);
  out[0] = v->getBranchingFactor();
  out[1] = v->getDepthLevels();
  out[2] = (int)v->getWeightingType();
  out[3] = (int)v->getScoringType();
  out[4] = (int)v->size();
}

/* transform(features, BowVector, FeatureVector, levelsup): both maps in key order; the FeatureVector in CSR form.
 * Also the per-feature transform(feature, word id, weight, node id, levelsup).  Buffers hold n (+1) entries. */
void ref_voc_transform(void* p, const uint8_t* desc, int n, int levelsup, int32_t* word, int32_t* node, double* weight,
                       int32_t* bow_ids, double* bow_values, int32_t* n_bow, int32_t* fv_ids, int32_t* fv_start,
                       int32_t* fv_items, int32_t* n_fv) {
  ProbeVocabulary* v = static_cast<ProbeVocabulary*>(p);
  std::vector<cv::Mat> f = rows(desc, n);
  DBoW3::BowVector bow;
  DBoW3::FeatureVector fv;
  v->transform(f, bow, fv, levelsup);
  int b = 0;
  for (DBoW3::BowVector::const_iterator it = bow.begin(); it != bow.end(); ++it, ++b) {
    bow_ids[b] = (int32_t)it->first;
    bow_values[b] = it->second;
  }
  *n_bow = b;
  int g = 0, at = 0;
  for (DBoW3::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++g) {
    fv_ids[g] = (int32_t)it->first;
    fv_start[g] = at;
    for (size_t j = 0; j < it->second.size(); j++) fv_items[at++] = (int32_t)it->second[j];
  }
  fv_start[g] = at;
  *n_fv = g;
  for (int i = 0; i < n; i++) {
    DBoW3::WordId id = 0;
    DBoW3::WordValue w = 0;
    DBoW3::NodeId nid = 0;
    v->one(f[i], id, w, &nid, levelsup);
    word[i] = (int32_t)id;
    node[i] = (int32_t)nid;
    weight[i] = w;
  }
}

void ref_voc_destroy(void* p) { delete static_cast<ProbeVocabulary*>(p); }

}  // extern "C"
