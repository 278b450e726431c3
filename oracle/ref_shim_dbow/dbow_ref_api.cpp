// dbow_ref_api.cpp — C entry points of oracle/_ref/libref_dbow.so: the REFERENCE's own Thirdparty/DBoW2 (BowVector.cpp,
// FeatureVector.cpp, FORB.cpp, ScoringObject.cpp, TemplatedVocabulary.h, DUtils/Random.cpp, Timestamp.cpp) compiled
// UNMODIFIED where it lies over a header shim for OpenCV core and boost::serialization (oracle/ref_shim_dbow).
// TEST INFRASTRUCTURE ONLY: tests/test_ref_dbow.py compares loadFromTextFile, transform (BowVector + FeatureVector),
// saveToTextFile and score with the oracle and with the product's host-side text parser; that pins row (f).3.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;  // include/ORBVocabulary.h:32

namespace {
struct Voc : ORBVocabulary {  // the node table is protected: a derived class may read it
  int n_nodes() const { return (int)m_nodes.size(); }
  int n_words() const { return (int)m_words.size(); }
  int word_node(int w) const { return (int)m_words[w]->id; }
  void dump(int* parent, int* word_id, double* weight, uint8_t* desc, int* child_off, int* child_idx) const {
    int c = 0;
    for (int i = 0; i < n_nodes(); i++) {
      const Node& nd = m_nodes[i];
      parent[i] = (int)nd.parent;
      word_id[i] = nd.isLeaf() ? (int)nd.word_id : -1;
      weight[i] = nd.weight;
      memset(desc + (size_t)i * 32, 0, 32);
      if (!nd.descriptor.empty()) memcpy(desc + (size_t)i * 32, nd.descriptor.data, 32);
      child_off[i] = c;
      for (size_t j = 0; j < nd.children.size(); j++) child_idx[c++] = (int)nd.children[j];
    }
    child_off[n_nodes()] = c;
  }
  void feature(const cv::Mat& f, int levelsup, int* word, double* weight, int* node) const {
    DBoW2::WordId w; DBoW2::WordValue wt; DBoW2::NodeId nid;
    transform(f, w, wt, &nid, levelsup);  // protected per-feature descent (TemplatedVocabulary.h:1201-1259)
    *word = (int)w; *weight = wt; *node = (int)nid;
  }
};
std::vector<cv::Mat> to_descriptors(const uint8_t* d, int n) {
  std::vector<cv::Mat> v(n);
  for (int i = 0; i < n; i++) { v[i].create(1, 32, CV_8U); memcpy(v[i].data, d + (size_t)i * 32, 32); }
  return v;
}
}  // namespace

extern "C" {

void* ref_voc_load_text(const char* path) {
  Voc* v = new Voc();
  if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
  return v;
}
void ref_voc_free(void* h) { delete (Voc*)h; }
int ref_voc_save_text(void* h, const char* path) { ((Voc*)h)->saveToTextFile(path); return 0; }

// k, L, scoring, weighting, number of nodes (root included), number of words
void ref_voc_info(void* h, int* out6) {
  Voc* v = (Voc*)h;
  out6[0] = v->getBranchingFactor(); out6[1] = v->getDepthLevels(); out6[2] = (int)v->getScoringType();
  out6[3] = (int)v->getWeightingType(); out6[4] = v->n_nodes(); out6[5] = v->n_words();
}

// per node: parent, word id (-1 for inner nodes), weight, 32 descriptor bytes (zeros for a node without one), the
// children in order as a CSR (child_off [n+1], child_idx); per word: its node
void ref_voc_nodes(void* h, int* parent, int* word_id, double* weight, uint8_t* desc, int* child_off, int* child_idx,
                   int* word_node) {
  Voc* v = (Voc*)h;
  v->dump(parent, word_id, weight, desc, child_off, child_idx);
  for (int w = 0; w < v->n_words(); w++) word_node[w] = v->word_node(w);
}

// transform(features, BowVector&, FeatureVector&, levelsup) (TemplatedVocabulary.h:1127-1176).  BowVector in map order
// (bow_word / bow_val, returns its size); FeatureVector as a CSR in map order (fv_node [nn], fv_off [nn+1], fv_idx [n]).
int ref_voc_transform(void* h, const uint8_t* desc, int n, int levelsup, int* bow_word, double* bow_val, int* fv_node,
                      int* fv_off, int* fv_idx, int* n_fv_nodes) {
  Voc* v = (Voc*)h;
  const std::vector<cv::Mat> feats = to_descriptors(desc, n);
  DBoW2::BowVector bv;
  DBoW2::FeatureVector fv;
  v->transform(feats, bv, fv, levelsup);
  int k = 0;
  for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++k) { bow_word[k] = (int)it->first; bow_val[k] = it->second; }
  int a = 0, e = 0;
  for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++a) {
    fv_node[a] = (int)it->first;
    fv_off[a] = e;
    for (size_t j = 0; j < it->second.size(); j++) fv_idx[e++] = (int)it->second[j];
  }
  fv_off[a] = e;
  *n_fv_nodes = a;
  return k;
}

// per-feature transform(feature, word id, weight, node id, levelsup) (TemplatedVocabulary.h:1201-1259)
void ref_voc_transform_features(void* h, const uint8_t* desc, int n, int levelsup, int* word, double* weight, int* node) {
  Voc* v = (Voc*)h;
  const std::vector<cv::Mat> feats = to_descriptors(desc, n);
  for (int i = 0; i < n; i++) v->feature(feats[i], levelsup, word + i, weight + i, node + i);
}

// score(v1, v2) of the vocabulary's scoring object between the BowVectors of two feature sets
double ref_voc_score(void* h, const uint8_t* desc1, int n1, const uint8_t* desc2, int n2) {
  Voc* v = (Voc*)h;
  DBoW2::BowVector a, b;
  v->transform(to_descriptors(desc1, n1), a);
  v->transform(to_descriptors(desc2, n2), b);
  return v->score(a, b);
}

}  // extern "C"
