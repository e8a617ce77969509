// TEST INFRASTRUCTURE.  C entry points around DBoW2 as vendored by the reference (Thirdparty/DBoW2: the
// TemplatedVocabulary template, FORB, BowVector, FeatureVector, ScoringObject, DUtils) compiled VERBATIM from
// /root/reference over oracle/slam_shim's cv::Mat.  Built by oracle/Makefile (target `refvoc`) into
// oracle/_ref/libref_vocabulary.so.  Pins the restated vocabulary descent of orb_oracle.cc (row N3:
// TemplatedVocabulary::transform, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1259, as Frame::ComputeBoW
// calls it, src/Frame.cc:395-402).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "Thirdparty/DBoW2/DBoW2/FORB.h"
#include "Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;  // include/ORBVocabulary.h:31-32

extern "C" {

void* refv_load_text(const char* path) {
  ORBVocabulary* v = new ORBVocabulary();
  if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
  return v;
}
void refv_destroy(void* v) { delete (ORBVocabulary*)v; }
int refv_size(void* v) { return (int)((ORBVocabulary*)v)->size(); }

// transform(features, BowVector, FeatureVector, levelsup): bow_id / bow_val in key order (returns the size),
// feature_node[i] = node of feature i in the FeatureVector, 0xffffffff when it is in none (stopped word)
int refv_transform(void* voc, const uint8_t* features, int n, int levelsup, uint32_t* bow_id, double* bow_val,
                   uint32_t* feature_node) {
  std::vector<cv::Mat> f;
  for (int i = 0; i < n; ++i) {
    cv::Mat d(1, 32, CV_8U);
    std::memcpy(d.data, features + (size_t)i * 32, 32);
    f.push_back(d);
  }
  DBoW2::BowVector bow;
  DBoW2::FeatureVector fv;
  ((ORBVocabulary*)voc)->transform(f, bow, fv, levelsup);
  int k = 0;
  for (const auto& e : bow) { bow_id[k] = e.first; bow_val[k] = e.second; ++k; }
  for (int i = 0; i < n; ++i) feature_node[i] = 0xffffffffu;
  for (const auto& e : fv)
    for (unsigned i : e.second) feature_node[i] = e.first;
  return k;
}

}  // extern "C"
