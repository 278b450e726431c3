// octree_core.h — node-level (serial) half of the keypoint distribution
// (ORBextractor::DistributeOctTree, src/ORBextractor.cc:555-779; ExtractorNode::DivideNode
// :480-536; compareNodes :538-553), written so that one GPU thread can run it while the rest
// of the CTA does the per-key work (quadrant counting and relabelling) in parallel.
//
// Re-formulation: keys never move.  Every key carries the id of the list
// node that owns it.  A "round" is
//   (1) parallel: for keys whose node is a candidate for division, count keys per quadrant;
//   (2) serial  : oct_round() — replays the reference's list surgery for this round from the
//                 quadrant counts alone (push_front of non-empty children in n1..n4 order, erase
//                 of the parent, the size>=N / no-growth / "careful" decisions, and the
//                 libstdc++ std::sort replica that orders the careful phase);
//   (3) parallel: keys of divided nodes take their child's id.
// The HD functions compile for the host too; tests/cpu_harness drives them against the oracle.
#pragma once
#include <stdint.h>

#include "orbx_core.h"

struct OctNode {
  int16_t x0, y0, x1, y1;  // UL=(x0,y0) UR=(x1,y0) BL=(x0,y1) BR=(x1,y1)
  int32_t cnt;             // vKeys.size()
  int16_t prev, next;      // std::list links, -1 = end
  int32_t c[4];            // quadrant counts (step 1); after division: child node ids or -1
  uint8_t cand;            // member of vSizeAndPointerToNode for the coming round
  uint8_t divided;         // divided in the last round (keys need relabelling), then recycled
  uint8_t nomore;          // bNoMore
  uint8_t pad;
};

struct OctState {
  OctNode* nodes;
  int node_cap;
  int head, tail, size;  // lNodes
  int free_head;         // recycled node ids chained through .next
  int nalloc;            // high-water mark
  uint64_t* vec;         // vSizeAndPointerToNode of the coming round: cnt<<32 | x0<<16 | id
  uint64_t* vec2;
  int nvec, nvec2;
  int* divlist;          // nodes divided in the last round (to recycle)
  int ndiv;
  int N;
  int careful;   // inside the inner while(!bFinish) of :691-754
  int finished;  // bFinish
  int overflow;  // node pool exhausted (cannot happen with node_cap >= 2*max(N+3,4*nIni)+8)
};

ORBX_HD int oct_alloc(OctState& S) {
  int id;
  if (S.free_head >= 0) {
    id = S.free_head;
    S.free_head = S.nodes[id].next;
  } else if (S.nalloc < S.node_cap) {
    id = S.nalloc++;
  } else {
    S.overflow = 1;
    return -1;
  }
  return id;
}

ORBX_HD void oct_push_front(OctState& S, int id) {
  OctNode& n = S.nodes[id];
  n.prev = -1;
  n.next = (int16_t)S.head;
  if (S.head >= 0) S.nodes[S.head].prev = (int16_t)id; else S.tail = id;
  S.head = id;
  S.size++;
}
ORBX_HD void oct_push_back(OctState& S, int id) {
  OctNode& n = S.nodes[id];
  n.next = -1;
  n.prev = (int16_t)S.tail;
  if (S.tail >= 0) S.nodes[S.tail].next = (int16_t)id; else S.head = id;
  S.tail = id;
  S.size++;
}
ORBX_HD void oct_unlink(OctState& S, int id) {
  OctNode& n = S.nodes[id];
  if (n.prev >= 0) S.nodes[n.prev].next = n.next; else S.head = n.next;
  if (n.next >= 0) S.nodes[n.next].prev = n.prev; else S.tail = n.prev;
  S.size--;
}

// Quadrant of a key inside a node: 0=n1(UL) 1=n2(UR) 2=n3(BL) 3=n4(BR)   (:511-524)
ORBX_HD int oct_quadrant(const OctNode& n, int x, int y) {
  const int mx = n.x0 + ((n.x1 - n.x0 + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2)
  const int my = n.y0 + ((n.y1 - n.y0 + 1) >> 1);
  return (x < mx ? 0 : 1) | (y < my ? 0 : 2);
}

// DivideNode + the push_front block (:622-676 / :703-741).  Uses nodes[id].c[] as the child key
// counts and overwrites it with the child ids.  Returns number of children with >1 key.
ORBX_HD int oct_divide(OctState& S, int id) {
  OctNode P = S.nodes[id];
  const int hx = (P.x1 - P.x0 + 1) >> 1, hy = (P.y1 - P.y0 + 1) >> 1;
  int expand = 0;
  for (int q = 0; q < 4; q++) {
    const int cnt = P.c[q];
    int cid = -1;
    if (cnt > 0) {
      cid = oct_alloc(S);
      if (cid < 0) return expand;
      OctNode& c = S.nodes[cid];
      c.x0 = (int16_t)((q & 1) ? P.x0 + hx : P.x0);
      c.x1 = (int16_t)((q & 1) ? P.x1 : P.x0 + hx);
      c.y0 = (int16_t)((q & 2) ? P.y0 + hy : P.y0);
      c.y1 = (int16_t)((q & 2) ? P.y1 : P.y0 + hy);
      c.cnt = cnt;
      c.nomore = cnt == 1;
      c.divided = 0;
      c.cand = 0;
      c.c[0] = c.c[1] = c.c[2] = c.c[3] = 0;
      oct_push_front(S, cid);
      if (cnt > 1) {
        c.cand = 1;
        S.vec2[S.nvec2++] = ((uint64_t)(uint32_t)cnt << 32) | ((uint64_t)(uint16_t)c.x0 << 16) | (uint32_t)cid;
        expand++;
      }
    }
    S.nodes[id].c[q] = cid;
  }
  oct_unlink(S, id);
  S.nodes[id].divided = 1;
  S.nodes[id].cand = 0;
  S.divlist[S.ndiv++] = id;
  return expand;
}

// ---- libstdc++ std::sort replica on the record vector --------------------------------------
// compareNodes orders by (count, UL.x) and leaves ties unordered, so the permutation among ties
// is whatever std::sort (GCC libstdc++: introsort, _S_threshold 16, heap fallback) produces.
ORBX_HD bool oct_less(uint64_t a, uint64_t b) { return (a >> 16) < (b >> 16); }

ORBX_HD void oct_adjust_heap(uint64_t* first, int hole, int len, uint64_t value) {
  const int top = hole;
  int child = hole;
  while (child < (len - 1) / 2) {
    child = 2 * (child + 1);
    if (oct_less(first[child], first[child - 1])) child--;
    first[hole] = first[child];
    hole = child;
  }
  if ((len & 1) == 0 && child == (len - 2) / 2) {
    child = 2 * (child + 1);
    first[hole] = first[child - 1];
    hole = child - 1;
  }
  int parent = (hole - 1) / 2;
  while (hole > top && oct_less(first[parent], value)) {
    first[hole] = first[parent];
    hole = parent;
    parent = (hole - 1) / 2;
  }
  first[hole] = value;
}

ORBX_HD void oct_heap_sort(uint64_t* first, int len) {  // __partial_sort(first,last,last)
  if (len >= 2) {
    int parent = (len - 2) / 2;
    while (true) {
      oct_adjust_heap(first, parent, len, first[parent]);
      if (parent == 0) break;
      parent--;
    }
  }
  for (int last = len - 1; last >= 1; last--) {
    uint64_t value = first[last];
    first[last] = first[0];
    oct_adjust_heap(first, 0, last, value);
  }
}

ORBX_HD void oct_std_sort(uint64_t* v, int n) {
  if (n < 2) return;
  int lg = 0;
  while ((1 << (lg + 1)) <= n) lg++;
  // explicit stack instead of recursion; sub-ranges are disjoint, so order does not matter
  int stk_lo[64], stk_hi[64], stk_d[64];
  int sp = 0;
  stk_lo[0] = 0; stk_hi[0] = n; stk_d[0] = 2 * lg; sp = 1;
  while (sp > 0) {
    sp--;
    int lo = stk_lo[sp], hi = stk_hi[sp], depth = stk_d[sp];
    while (hi - lo > 16) {
      if (depth == 0) {
        oct_heap_sort(v + lo, hi - lo);
        break;
      }
      depth--;
      const int mid = lo + (hi - lo) / 2;
      const int a = lo + 1, b = mid, c = hi - 1;
      int m;  // __move_median_to_first(lo, a, b, c)
      if (oct_less(v[a], v[b])) {
        if (oct_less(v[b], v[c])) m = b;
        else if (oct_less(v[a], v[c])) m = c;
        else m = a;
      } else if (oct_less(v[a], v[c])) m = a;
      else if (oct_less(v[b], v[c])) m = c;
      else m = b;
      uint64_t t = v[lo]; v[lo] = v[m]; v[m] = t;
      int first = lo + 1, last = hi;  // __unguarded_partition(lo+1, hi, lo)
      while (true) {
        while (oct_less(v[first], v[lo])) first++;
        last--;
        while (oct_less(v[lo], v[last])) last--;
        if (!(first < last)) break;
        t = v[first]; v[first] = v[last]; v[last] = t;
        first++;
      }
      stk_lo[sp] = first; stk_hi[sp] = hi; stk_d[sp] = depth; sp++;  // right part
      hi = first;                                                     // loop on the left part
    }
  }
  // __final_insertion_sort
  const int guarded_end = n > 16 ? 16 : n;
  for (int i = 1; i < guarded_end; i++) {
    uint64_t val = v[i];
    if (oct_less(val, v[0])) {
      for (int j = i; j > 0; j--) v[j] = v[j - 1];
      v[0] = val;
    } else {
      int j = i;
      while (oct_less(val, v[j - 1])) { v[j] = v[j - 1]; j--; }
      v[j] = val;
    }
  }
  for (int i = guarded_end; i < n; i++) {  // __unguarded_insertion_sort
    uint64_t val = v[i];
    int j = i;
    while (oct_less(val, v[j - 1])) { v[j] = v[j - 1]; j--; }
    v[j] = val;
  }
}

// Root nodes (:564-601).  root_cnt[i] = number of keys whose (int)(x/hX) == i.
ORBX_HD void oct_init(OctState& S, int nIni, float hX, int height, const int* root_cnt) {
  S.head = S.tail = -1; S.size = 0; S.free_head = -1; S.nalloc = 0;
  S.nvec = S.nvec2 = 0; S.ndiv = 0; S.careful = 0; S.finished = 0; S.overflow = 0;
  for (int i = 0; i < nIni; i++) {
    // ids 0..nIni-1 are reserved for the roots so that keys can be labelled before this runs
    OctNode& n = S.nodes[i];
    n.x0 = (int16_t)(int)f_mul(hX, (float)i);
    n.x1 = (int16_t)(int)f_mul(hX, (float)(i + 1));
    n.y0 = 0; n.y1 = (int16_t)height;
    n.cnt = root_cnt[i];
    n.nomore = root_cnt[i] == 1;
    n.divided = 0;
    n.cand = root_cnt[i] > 1;
    n.c[0] = n.c[1] = n.c[2] = n.c[3] = 0;
    n.prev = n.next = -1;
  }
  S.nalloc = nIni;
  for (int i = 0; i < nIni; i++) {
    if (root_cnt[i] > 0) oct_push_back(S, i);
    else {  // never linked: recycle
      S.nodes[i].next = (int16_t)S.free_head;
      S.free_head = i;
    }
  }
}

// One round of the outer/inner while loops (:610-755).  Quadrant counts of every candidate
// node must be in nodes[].c[] on entry.  On exit `finished` is set, or the next round's
// candidates carry cand=1 and nodes divided in this round carry divided=1 (+ child ids in c[]).
ORBX_HD void oct_round(OctState& S) {
  // recycle the parents divided in the previous round (their keys were relabelled since)
  for (int i = 0; i < S.ndiv; i++) {
    const int id = S.divlist[i];
    S.nodes[id].divided = 0;
    S.nodes[id].next = (int16_t)S.free_head;
    S.free_head = id;
  }
  S.ndiv = 0;
  const int prevSize = S.size;
  S.nvec2 = 0;
  if (!S.careful) {
    int nToExpand = 0;
    int it = S.head;
    while (it >= 0) {
      const int nx = S.nodes[it].next;  // children go to the front: never revisited
      if (!S.nodes[it].nomore) nToExpand += oct_divide(S, it);
      it = nx;
    }
    if (S.size >= S.N || S.size == prevSize) S.finished = 1;
    else if (S.size + nToExpand * 3 > S.N) S.careful = 1;
  } else {
    oct_std_sort(S.vec, S.nvec);
    for (int j = S.nvec - 1; j >= 0; j--) {
      oct_divide(S, (int)(S.vec[j] & 0xffff));
      if (S.size >= S.N) break;
    }
    if (S.size >= S.N || S.size == prevSize) S.finished = 1;
  }
  if (S.overflow) S.finished = 1;
  uint64_t* t = S.vec; S.vec = S.vec2; S.vec2 = t;
  S.nvec = S.nvec2;
}

// =================================================================================================
// Array ("scan") form of the same procedure, used by the CUDA kernel.
//
// lNodes is kept as an ARRAY in list order and rebuilt every round, so a node's id is its list
// position.  A round processes a sequence of nodes proc_0 .. proc_{m-1} (first loop: every node
// with more than one key, in list order; "careful" loop: the std::sort'ed candidates from the back,
// up to and including the first one that brings the list to N).  Because every push goes to the
// FRONT and every processed node is erased, the list after the round is
//     kids(proc_{m-1}) , ... , kids(proc_1) , kids(proc_0) , <unprocessed nodes in their old order>
// with kids(p) = its non-empty children in the order n4,n3,n2,n1, and the next round's candidate
// vector holds the children with more than one key in creation order (proc order, n1..n4).  All
// positions are therefore prefix sums over the proc order (children) and over the list order
// (survivors); octp_emit below is the per-node part, the sums are block scans on the GPU and plain
// loops in the CPU harness (tests/cpu_harness, h_distribute_scan).
// =================================================================================================
struct OctPNode {
  int16_t x0, y0, x1, y1;
  int32_t cnt;  // vKeys.size(); bNoMore <=> cnt == 1
};

ORBX_HD int octp_quadrant(const OctPNode& n, int x, int y) {
  const int mx = n.x0 + ((n.x1 - n.x0 + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2)
  const int my = n.y0 + ((n.y1 - n.y0 + 1) >> 1);
  return (x < mx ? 0 : 1) | (y < my ? 0 : 2);
}

// vSizeAndPointerToNode record: compareNodes looks at (count, UL.x); the low 16 bits carry the node
ORBX_HD uint64_t octp_rec(int cnt, int x0, int pos) {
  return ((uint64_t)(uint32_t)cnt << 32) | ((uint64_t)(uint16_t)x0 << 16) | (uint32_t)pos;
}

// Children of processed node P with quadrant counts c[0..3] (n1..n4).  They take next[blk, blk+K)
// front to back as n4,n3,n2,n1; those with more than one key are appended to vec2 from index cnd
// in n1..n4 order; c[] is overwritten with each child's position (-1 if empty) for relabelling.
ORBX_HD void octp_emit(const OctPNode& P, int* c, int blk, int cnd, OctPNode* next, uint64_t* vec2) {
  const int hx = (P.x1 - P.x0 + 1) >> 1, hy = (P.y1 - P.y0 + 1) >> 1;
  int pos = blk + (c[0] > 0) + (c[1] > 0) + (c[2] > 0) + (c[3] > 0);
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int cnt = c[q];
    if (cnt > 0) {
      pos--;
      OctPNode k;
      k.x0 = (int16_t)((q & 1) ? P.x0 + hx : P.x0);
      k.x1 = (int16_t)((q & 1) ? P.x1 : P.x0 + hx);
      k.y0 = (int16_t)((q & 2) ? P.y0 + hy : P.y0);
      k.y1 = (int16_t)((q & 2) ? P.y1 : P.y0 + hy);
      k.cnt = cnt;
      next[pos] = k;
      if (cnt > 1) vec2[cnd++] = octp_rec(cnt, k.x0, pos);
      c[q] = pos;
    } else {
      c[q] = -1;
    }
  }
}

// Root rectangle i of nIni (:564-583)
ORBX_HD OctPNode octp_root(int i, float hX, int height, int cnt) {
  OctPNode n;
  n.x0 = (int16_t)(int)f_mul(hX, (float)i);
  n.x1 = (int16_t)(int)f_mul(hX, (float)(i + 1));
  n.y0 = 0;
  n.y1 = (int16_t)height;
  n.cnt = cnt;
  return n;
}
