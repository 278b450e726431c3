// voc.cu — DBoW2 vocabulary I/O for orbm_bow_transform: the reference's on-disk text format
// (TemplatedVocabulary::loadFromTextFile / saveToTextFile, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1464,
// FORB::fromString / toString Thirdparty/DBoW2/DBoW2/FORB.cpp:103-135; loaded at src/System.cc:118 as ORBvoc.txt)
// parsed on the host into the flat tree the kernels walk, and its upload to the device.
//
// Format: line 1 = "k L scoring weighting"; every further line = one node, ids 1.. in file order:
//   "parent isLeaf d0 ... d31 weight".  Children keep file order (push_back), leaves get word ids in file order.
// One documented difference: the reference's `while(!f.eof())` loop turns the file's trailing newline into one more
// node — child of node 0, no children, word 0, weight 0 and a descriptor of UNINITIALISED memory (cv::Mat::create).
// That node has no defined content to reproduce, so blank lines are skipped here.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/orbx.h"
#include <cuda_runtime.h>

static void voc_host_release(orbm_vocabulary_host* v) {
  if (!v) return;
  free(v->child_off); free(v->child_idx); free(v->node_desc); free(v->node_word); free(v->node_weight);
  free(v->node_parent);
  memset(v, 0, sizeof(*v));
}

extern "C" void orbm_vocabulary_host_free(orbm_vocabulary_host* v) { voc_host_release(v); }

extern "C" int orbm_vocabulary_parse_text(const char* path, orbm_vocabulary_host* out) {
  if (!path || !out) return ORBX_E_INVALID;
  memset(out, 0, sizeof(*out));
  FILE* f = fopen(path, "rb");
  if (!f) return ORBX_E_INVALID;
  std::string line;
  auto getline = [&](std::string& s) -> bool {
    s.clear();
    int c;
    bool any = false;
    while ((c = fgetc(f)) != EOF) {
      any = true;
      if (c == '\n') break;
      s.push_back((char)c);
    }
    return any;
  };
  if (!getline(line)) { fclose(f); return ORBX_E_INVALID; }
  int k = -1, L = -1, n1 = -1, n2 = -1;
  if (sscanf(line.c_str(), "%d %d %d %d", &k, &L, &n1, &n2) != 4 || k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 ||
      n1 > 5 || n2 < 0 || n2 > 3) {  // :1362-1366
    fclose(f);
    return ORBX_E_INVALID;
  }
  std::vector<int32_t> parent(1, 0), word(1, -1);
  std::vector<double> weight(1, 0.0);
  std::vector<uint8_t> desc(32, 0);
  std::vector<std::vector<int32_t>> children(1);
  int n_words = 0;
  bool bad = false;
  while (getline(line)) {
    const char* p = line.c_str();
    char* e = nullptr;
    while (*p == ' ' || *p == '\t' || *p == '\r') p++;
    if (!*p) continue;  // blank line (see header)
    const long pid = strtol(p, &e, 10);
    if (e == p) { bad = true; break; }
    p = e;
    const long leaf = strtol(p, &e, 10);
    if (e == p) { bad = true; break; }
    p = e;
    const int nid = (int)parent.size();
    if (pid < 0 || pid >= nid) { bad = true; break; }
    uint8_t d[32];
    for (int i = 0; i < 32; i++) {
      const long v = strtol(p, &e, 10);
      if (e == p) { bad = true; break; }
      d[i] = (uint8_t)v;  // FORB::fromString: *p = (unsigned char)n
      p = e;
    }
    if (bad) break;
    const double w = strtod(p, &e);
    if (e == p) { bad = true; break; }
    parent.push_back((int32_t)pid);
    children.emplace_back();
    children[pid].push_back(nid);
    desc.insert(desc.end(), d, d + 32);
    weight.push_back(w);
    word.push_back(leaf > 0 ? n_words++ : -1);
  }
  fclose(f);
  if (bad) return ORBX_E_INVALID;
  const int n = (int)parent.size();
  // a node marked as a leaf must not have children, and the kernels treat "no children" as a leaf (isLeaf(), :1243)
  for (int i = 1; i < n; i++)
    if ((word[i] >= 0) != children[i].empty()) return ORBX_E_INVALID;
  if (children[0].empty()) return ORBX_E_INVALID;
  out->k = k; out->L = L; out->scoring = n1; out->weighting = n2; out->n_nodes = n; out->n_words = n_words;
  out->child_off = (int32_t*)malloc(sizeof(int32_t) * (n + 1));
  out->child_idx = (int32_t*)malloc(sizeof(int32_t) * (n > 1 ? n - 1 : 1));
  out->node_desc = (uint8_t*)malloc((size_t)n * 32);
  out->node_word = (int32_t*)malloc(sizeof(int32_t) * n);
  out->node_weight = (double*)malloc(sizeof(double) * n);
  out->node_parent = (int32_t*)malloc(sizeof(int32_t) * n);
  if (!out->child_off || !out->child_idx || !out->node_desc || !out->node_word || !out->node_weight ||
      !out->node_parent) {
    voc_host_release(out);
    return ORBX_E_NOMEM;
  }
  int o = 0;
  for (int i = 0; i < n; i++) {
    out->child_off[i] = o;
    for (int c : children[i]) out->child_idx[o++] = c;
  }
  out->child_off[n] = o;
  memcpy(out->node_desc, desc.data(), (size_t)n * 32);
  memcpy(out->node_word, word.data(), sizeof(int32_t) * n);
  memcpy(out->node_weight, weight.data(), sizeof(double) * n);
  memcpy(out->node_parent, parent.data(), sizeof(int32_t) * n);
  return ORBX_OK;
}

extern "C" int orbm_vocabulary_save_text(const char* path, const orbm_vocabulary_host* v) {
  if (!path || !v || v->n_nodes < 1) return ORBX_E_INVALID;
  FILE* f = fopen(path, "wb");
  if (!f) return ORBX_E_INVALID;
  fprintf(f, "%d %d  %d %d\n", v->k, v->L, v->scoring, v->weighting);  // :1432 (two spaces, as the reference writes)
  for (int i = 1; i < v->n_nodes; i++) {
    fprintf(f, "%d %s", v->node_parent[i], v->node_word[i] >= 0 ? "1 " : "0 ");
    for (int j = 0; j < 32; j++) fprintf(f, "%d ", (int)v->node_desc[(size_t)i * 32 + j]);  // FORB::toString
    fprintf(f, " %.17g\n", v->node_weight[i]);  // (the reference streams the double at the default 6 digits)
  }
  fclose(f);
  return ORBX_OK;
}

extern "C" void orbm_vocabulary_release(orbm_vocabulary* d) {
  if (!d) return;
  cudaFree((void*)d->d_child_off); cudaFree((void*)d->d_child_idx); cudaFree((void*)d->d_node_desc);
  cudaFree((void*)d->d_node_word); cudaFree((void*)d->d_node_weight);
  memset(d, 0, sizeof(*d));
}

extern "C" int orbm_vocabulary_upload(const orbm_vocabulary_host* v, orbm_vocabulary* out) {
  if (!v || !out || v->n_nodes < 2) return ORBX_E_INVALID;
  memset(out, 0, sizeof(*out));
  const int n = v->n_nodes;
  void *a = nullptr, *b = nullptr, *c = nullptr, *d = nullptr, *e = nullptr;
  bool ok = cudaMalloc(&a, sizeof(int32_t) * (n + 1)) == cudaSuccess && cudaMalloc(&b, sizeof(int32_t) * n) == cudaSuccess &&
            cudaMalloc(&c, (size_t)n * 32) == cudaSuccess && cudaMalloc(&d, sizeof(int32_t) * n) == cudaSuccess &&
            cudaMalloc(&e, sizeof(double) * n) == cudaSuccess;
  ok = ok && cudaMemcpy(a, v->child_off, sizeof(int32_t) * (n + 1), cudaMemcpyHostToDevice) == cudaSuccess &&
       cudaMemcpy(b, v->child_idx, sizeof(int32_t) * (n - 1), cudaMemcpyHostToDevice) == cudaSuccess &&
       cudaMemcpy(c, v->node_desc, (size_t)n * 32, cudaMemcpyHostToDevice) == cudaSuccess &&
       cudaMemcpy(d, v->node_word, sizeof(int32_t) * n, cudaMemcpyHostToDevice) == cudaSuccess &&
       cudaMemcpy(e, v->node_weight, sizeof(double) * n, cudaMemcpyHostToDevice) == cudaSuccess;
  out->d_child_off = (const int32_t*)a; out->d_child_idx = (const int32_t*)b; out->d_node_desc = (const uint8_t*)c;
  out->d_node_word = (const int32_t*)d; out->d_node_weight = (const double*)e;
  out->n_nodes = n; out->L = v->L;
  if (!ok) { cudaGetLastError(); orbm_vocabulary_release(out); return ORBX_E_CUDA; }
  return ORBX_OK;
}

extern "C" int orbm_vocabulary_load_text(const char* path, orbm_vocabulary* out, int32_t* k, int32_t* n_words) {
  orbm_vocabulary_host h;
  int rc = orbm_vocabulary_parse_text(path, &h);
  if (rc) return rc;
  if (k) *k = h.k;
  if (n_words) *n_words = h.n_words;
  rc = orbm_vocabulary_upload(&h, out);
  voc_host_release(&h);
  return rc;
}
