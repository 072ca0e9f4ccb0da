// kdreplay_ref.h -- TEST SCAFFOLDING: the eager, readable forms of the kd-tree replay (whole-tree build, plain
// traversal, sequential lazy tree) that the product's cooperative lazy variant (slam_b200/csrc/kdreplay.h) is checked
// against on the CPU (tests/host/kdreplay_host.cpp, tests/test_kdreplay_cpu.py).  No kernel uses them.
// They restate the reference's deterministic kd-tree (kdtree.cpp:114-233, 235-269, 294-317) and its n_nearest(q, 1)
// traversal (kdtree.cpp:515-683).
#pragma once
#include "../../slam_b200/csrc/kdreplay.h"

namespace kdreplay {

// 48-byte node.  box = {x.lower, x.upper, y.lower, y.upper}.
struct Node {
  float cut_val, cut_val_left, cut_val_right;
  int32_t cut_dim;           // -1 marks a terminal node
  int32_t l, u;
  float box[4];
  int32_t left, right;       // node ids, -1 = NULL
};


// Builds the tree over n points.  ind[n] and nodes[max_nodes] are outputs;
// returns OK or an ERR_* code; *n_nodes receives the node count; root is node 0.
template <class P, class I>
KD_FN int build(const P& pts, int n, I* ind, Node* nodes, int max_nodes, int* n_nodes) {
  struct Frame { int l, u, parent, node, m, stage; };
  Frame st[STACK];
  int sp = 0, nn = 0, ret = -1;
  for (int i = 0; i < n; i++) ind[i] = (I)i;
  *n_nodes = 0;
  if (n <= 0) return OK;
  st[sp].l = 0; st[sp].u = n - 1; st[sp].parent = -1; st[sp].node = -1; st[sp].m = 0; st[sp].stage = 0; sp++;
  while (sp > 0) {
    const int fi = sp - 1;
    const int l = st[fi].l, u = st[fi].u;
    if (st[fi].stage == 0) {
      if (u < l) { ret = -1; sp--; continue; }                    // kdtree.cpp:125-127
      if (nn >= max_nodes) return ERR_NODES;
      const int id = nn++;
      Node nd;
      nd.l = l; nd.u = u; nd.left = nd.right = -1;
      nd.cut_val = nd.cut_val_left = nd.cut_val_right = 0.0f;
      if (u - l <= BUCKET) {                                      // terminal node, true box (:129-142)
        spread(pts, ind, 0, l, u, &nd.box[0], &nd.box[1]);
        spread(pts, ind, 1, l, u, &nd.box[2], &nd.box[3]);
        nd.cut_dim = -1;
        nodes[id] = nd;
        ret = id; sp--; continue;
      }
      // approximate box: recompute only the parent's cut dimension (:153-168)
      const int par = st[fi].parent;
      int c = -1;
      float maxspread = 0.0f;
      for (int i = 0; i < 2; i++) {
        if (par < 0 || nodes[par].cut_dim == i) {
          spread(pts, ind, i, l, u, &nd.box[2 * i], &nd.box[2 * i + 1]);
        } else {
          nd.box[2 * i] = nodes[par].box[2 * i];
          nd.box[2 * i + 1] = nodes[par].box[2 * i + 1];
        }
        float sprd = nd.box[2 * i + 1] - nd.box[2 * i];
        if (sprd > maxspread) { maxspread = sprd; c = i; }
      }
      if (c < 0) return ERR_DEGENERATE;     // > 13 coincident points: the reference indexes [-1] here
      // split value = sequential float mean (:181-186)
      float sum = 0.0f;
      for (int k = l; k <= u; k++) sum += coord(pts, ind[k], c);
      const float average = sum / (float)(u - l + 1);
      // partition <= average to the left (:294-317)
      int lb = l, ub = u;
      while (lb < ub) {
        if (coord(pts, ind[lb], c) <= average) {
          lb++;
        } else {
          I tmp = ind[lb]; ind[lb] = ind[ub]; ind[ub] = tmp;
          ub--;
        }
      }
      const int m = (coord(pts, ind[lb], c) <= average) ? lb : lb - 1;
      nd.cut_dim = c;
      nodes[id] = nd;
      st[fi].node = id; st[fi].m = m; st[fi].stage = 1;
      if (sp >= STACK) return ERR_DEPTH;
      st[sp].l = l; st[sp].u = m; st[sp].parent = id; st[sp].node = -1; st[sp].m = 0; st[sp].stage = 0; sp++;
    } else if (st[fi].stage == 1) {
      nodes[st[fi].node].left = ret;
      st[fi].stage = 2;
      if (sp >= STACK) return ERR_DEPTH;
      const int id = st[fi].node, m = st[fi].m;
      st[sp].l = m + 1; st[sp].u = u; st[sp].parent = id; st[sp].node = -1; st[sp].m = 0; st[sp].stage = 0; sp++;
    } else {
      Node& nd = nodes[st[fi].node];
      nd.right = ret;
      const int c = nd.cut_dim;
      if (nd.right < 0 && nd.left < 0) return ERR_DEGENERATE;
      if (nd.right < 0) {                                          // :206-210
        for (int i = 0; i < 4; i++) nd.box[i] = nodes[nd.left].box[i];
        nd.cut_val = nodes[nd.left].box[2 * c + 1];
        nd.cut_val_left = nd.cut_val_right = nd.cut_val;
      } else if (nd.left < 0) {                                    // :211-215
        for (int i = 0; i < 4; i++) nd.box[i] = nodes[nd.right].box[i];
        nd.cut_val = nodes[nd.right].box[2 * c + 1];
        nd.cut_val_left = nd.cut_val_right = nd.cut_val;
      } else {                                                     // :216-231
        nd.cut_val_right = nodes[nd.right].box[2 * c];
        nd.cut_val_left = nodes[nd.left].box[2 * c + 1];
        nd.cut_val = (float)((double)(nd.cut_val_left + nd.cut_val_right) / 2.0);
        for (int i = 0; i < 2; i++) {
          float lu = nodes[nd.left].box[2 * i + 1], ru = nodes[nd.right].box[2 * i + 1];
          float ll = nodes[nd.left].box[2 * i], rl = nodes[nd.right].box[2 * i];
          nd.box[2 * i + 1] = lu > ru ? lu : ru;       // std::max(left, right)
          nd.box[2 * i] = rl < ll ? rl : ll;           // std::min(left, right)
        }
      }
      ret = st[fi].node; sp--;
    }
  }
  *n_nodes = nn;
  return OK;
}


// kdtree.cpp:568-581
KD_HD bool box_in_range(const Node& nd, float qx, float qy, float ballsize) {
  float d0 = dis_from_bnd(qx, nd.box[0], nd.box[1]);
  float dis2 = 0.0f;
  dis2 += d0 * d0;
  if (dis2 > ballsize) return false;
  float d1 = dis_from_bnd(qy, nd.box[2], nd.box[3]);
  dis2 += d1 * d1;
  if (dis2 > ballsize) return false;
  return true;
}

// Exact replay of n_nearest(q, 1): returns the ORIGINAL model index the
// reference would report (or -1 when nothing lies within 1e38) and its float
// squared distance.  Returns OK or ERR_DEPTH.
template <class P, class I>
KD_FN int nearest(const P& pts, const I* ind, const Node* nodes, int n_nodes,
                  float qx, float qy, int32_t* idx_out, float* dis_out) {
  struct Pending { int node; float extra; };
  Pending st[STACK];
  int sp = 0;
  float ballsize = 1.0e38f;            // kdtree.cpp:325
  bool have = false;
  int32_t best = -1;
  float bestd = 0.0f;
  *idx_out = -1; *dis_out = 0.0f;
  if (n_nodes <= 0) return OK;
  int cur = 0;
  for (;;) {
    while (cur >= 0 && nodes[cur].cut_dim >= 0) {                 // descend (:528-548)
      const Node& nd = nodes[cur];
      float qval = nd.cut_dim == 0 ? qx : qy;
      int closer, farther;
      float extra;
      if (qval < nd.cut_val) { closer = nd.left; farther = nd.right; extra = nd.cut_val_right - qval; }
      else { closer = nd.right; farther = nd.left; extra = qval - nd.cut_val_left; }
      if (sp >= STACK) return ERR_DEPTH;
      st[sp].node = farther; st[sp].extra = extra; sp++;
      cur = closer;
    }
    if (cur >= 0) {                                               // leaf scan (:604-681, nn == 1, rearranged data)
      const Node& nd = nodes[cur];
      for (int i = nd.l; i <= nd.u; i++) {
        int o = (int)ind[i];
        const Pt po = pts(o);
        float dx = po.x - qx;
        float dis = 0.0f;
        dis += dx * dx;
        if (dis > ballsize) continue;
        float dy = po.y - qy;
        dis += dy * dy;
        if (dis > ballsize) continue;
        have = true; best = o; bestd = dis; ballsize = dis;       // push (first) or replace (later, dis <= best)
      }
    }
    bool found = false;
    while (sp > 0) {                                              // unwind: far side test (:550-554)
      sp--;
      int far_ = st[sp].node;
      float extra = st[sp].extra;
      if (far_ >= 0 && extra * extra < ballsize && box_in_range(nodes[far_], qx, qy, ballsize)) {
        cur = far_; found = true; break;
      }
    }
    if (!found) break;
  }
  if (have) { *idx_out = best; *dis_out = bestd; }
  return OK;
}


template <class P, class I>
KD_FN int lazy_init(const P& pts, int n, I* ind, LNode* nodes, int* n_nodes) {
  for (int i = 0; i < n; i++) ind[i] = (I)i;
  *n_nodes = 0;
  if (n <= 0) return OK;
  LNode r;
  r.l = 0; r.u = n - 1; r.left = r.right = -1;
  r.cut_val = r.cut_val_left = r.cut_val_right = 0.0f;
  spread(pts, ind, 0, 0, n - 1, &r.tbox[0], &r.tbox[1]);
  spread(pts, ind, 1, 0, n - 1, &r.tbox[2], &r.tbox[3]);
  for (int i = 0; i < 4; i++) r.abox[i] = r.tbox[i];
  r.cut_dim = (n - 1 <= BUCKET) ? LEAF : UNSPLIT;
  nodes[0] = r;
  *n_nodes = 1;
  return OK;
}


// Exact replay of n_nearest(q, 1) on the lazily materialised tree (same traversal as nearest()).
template <class P, class I>
KD_FN int lazy_nearest(const P& pts, I* ind, LNode* nodes, int max_nodes, int* n_nodes,
                       float qx, float qy, int32_t* idx_out, float* dis_out) {
  struct Pending { int node; float extra; };
  Pending st[STACK];
  int sp = 0;
  float ballsize = 1.0e38f;
  bool have = false;
  int32_t best = -1;
  float bestd = 0.0f;
  *idx_out = -1; *dis_out = 0.0f;
  if (*n_nodes <= 0) return OK;
  int cur = 0;
  for (;;) {
    while (cur >= 0 && nodes[cur].cut_dim != LEAF) {
      if (nodes[cur].cut_dim == UNSPLIT) {
        int rc = lazy_split(pts, ind, nodes, max_nodes, n_nodes, cur);
        if (rc != OK) return rc;
      }
      const LNode& nd = nodes[cur];
      float qval = nd.cut_dim == 0 ? qx : qy;
      int closer, farther;
      float extra;
      if (qval < nd.cut_val) { closer = nd.left; farther = nd.right; extra = nd.cut_val_right - qval; }
      else { closer = nd.right; farther = nd.left; extra = qval - nd.cut_val_left; }
      if (sp >= STACK) return ERR_DEPTH;
      st[sp].node = farther; st[sp].extra = extra; sp++;
      cur = closer;
    }
    if (cur >= 0) {
      const int l = nodes[cur].l, u = nodes[cur].u;
      for (int i = l; i <= u; i++) {
        int o = (int)ind[i];
        const Pt po = pts(o);
        float dx = po.x - qx;
        float dis = 0.0f;
        dis += dx * dx;
        if (dis > ballsize) continue;
        float dy = po.y - qy;
        dis += dy * dy;
        if (dis > ballsize) continue;
        have = true; best = o; bestd = dis; ballsize = dis;
      }
    }
    bool found = false;
    while (sp > 0) {
      sp--;
      int far_ = st[sp].node;
      float extra = st[sp].extra;
      if (far_ >= 0 && extra * extra < ballsize && lbox_in_range(nodes[far_], qx, qy, ballsize)) {
        cur = far_; found = true; break;
      }
    }
    if (!found) break;
  }
  if (have) { *idx_out = best; *dis_out = bestd; }
  return OK;
}


}  // namespace kdreplay
