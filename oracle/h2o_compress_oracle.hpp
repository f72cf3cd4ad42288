// TEST INFRASTRUCTURE ONLY (see oracle/h2o_tree_oracle.hpp). CPU restatement of the reference's tree serialisation:
//   DecidedNode.size / compress, LeafNode.compress, DTree.compress   h2o-algos/src/main/java/hex/tree/DTree.java:845-933, 950-956, 972-982
//   IcedBitSet.compress2 / compress3                                  h2o-core/src/main/java/water/util/IcedBitSet.java:92-111
// and of the reader the MOJO / in-cluster scoring runs over those bytes:
//   SharedTreeMojoModel.scoreTree                                     h2o-genmodel/src/main/java/hex/genmodel/algos/tree/SharedTreeMojoModel.java:129-240
//   GenmodelBitSet.fill2 / fill3 / isInRange / contains               h2o-genmodel/src/main/java/hex/genmodel/utils/GenmodelBitSet.java
// Byte order: AutoBuffer and ByteBufferWrapper both use the native order (little endian here).
// Parity status: CompressedTreeTest holds no byte vectors (it trains on smalldata/logreg/prostate.csv, which is not shipped), so
// the byte stream is pinned by the reference's own invariants instead: size() == bytes written (the assert at DTree.java:929),
// and scoreTree over the bytes == the node-table walk for every row (what CompressedTreeTest.testNodeIdAssignment checks).
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

#include "h2o_tree_oracle.hpp"

namespace h2o_oracle {

struct CompressedTreeBytes {
  std::vector<uint8_t> bits;  // CompressedTree._bits
  std::vector<uint8_t> aux;   // DTree._abAux: per decided node {nid, #nodes left, n0, n1, p0, p1, se0, se1, nid0, nid1}
};

namespace compress_detail {

inline void put1(std::vector<uint8_t>& b, int v) { b.push_back((uint8_t)v); }
inline void put2(std::vector<uint8_t>& b, int v) { b.push_back((uint8_t)v); b.push_back((uint8_t)(v >> 8)); }
inline void put3(std::vector<uint8_t>& b, int v) { put2(b, v); b.push_back((uint8_t)(v >> 16)); }
inline void put4(std::vector<uint8_t>& b, int32_t v) { put2(b, v); put2(b, v >> 16); }
inline void put4f(std::vector<uint8_t>& b, float f) { int32_t v; memcpy(&v, &f, 4); put4(b, v); }

struct Sizer {
  const TreeOut& t;
  std::vector<int> size, nnodes;
  std::vector<uint8_t> type;
  explicit Sizer(const TreeOut& tree) : t(tree), size(tree.nodes.size(), 0), nnodes(tree.nodes.size(), -1), type(tree.nodes.size(), 0) {}
  // DecidedNode.size (:845-881) / LeafNode.size (:921)
  int size_of(int nid) {
    const Node& n = t.nodes[nid];
    if (n.kind == LEAF) return 4;
    if (size[nid]) return size[nid];
    const Split& sp = *n._split;
    uint8_t ty = 0;
    if (sp._equal != 0) ty |= sp._equal == 1 ? 4 : (sp._equal == 2 ? 8 : 12);
    int res = sp._equal == 3 ? 9 + sp._bs->numBytes() : 7;
    res++;  // NA split direction
    if (sp._nasplit == NAvsREST) res -= 4;
    const int lsz = size_of(n._nids[0]);
    res += lsz;
    if (t.nodes[n._nids[0]].kind == LEAF) ty |= 48;
    else {
      const int slen = lsz < 256 ? 0 : (lsz < 65535 ? 1 : (lsz < (1 << 24) ? 2 : 3));
      ty |= (uint8_t)slen;
      res += slen + 1;
    }
    if (t.nodes[n._nids[1]].kind == LEAF) ty |= (uint8_t)(48 << 2);
    res += size_of(n._nids[1]);
    type[nid] = ty;
    return size[nid] = res;
  }
  int num_nodes(int nid) {  // :883-889 (leaves count 0)
    const Node& n = t.nodes[nid];
    if (n.kind == LEAF) return 0;
    if (nnodes[nid] >= 0) return nnodes[nid];
    return nnodes[nid] = 1 + num_nodes(n._nids[0]) + num_nodes(n._nids[1]);
  }
};

inline void compress_node(Sizer& z, int nid, CompressedTreeBytes& out) {
  const Node& n = z.t.nodes[nid];
  if (n.kind == LEAF) {  // LeafNode.compress (:918-920)
    put4f(out.bits, n._pred);
    return;
  }
  const size_t pos = out.bits.size();
  const int sz = z.size_of(nid);
  const Split& sp = *n._split;
  put1(out.bits, z.type[nid]);
  put2(out.bits, sp._col);
  put1(out.bits, sp._nasplit);
  if (sp._nasplit != NAvsREST) {
    if (sp._equal == 0 || sp._equal == 1) put4f(out.bits, n._splat);
    else if (sp._equal == 2) {  // IcedBitSet.compress2: exactly 4 bytes
      for (int i = 0; i < 4; i++) put1(out.bits, sp._bs->_val[i]);
    } else {  // compress3: 2 B bit offset, 4 B number of bits, the bytes
      put2(out.bits, sp._bs->_bitoff);
      put4(out.bits, sp._bs->_nbits);
      for (int i = 0; i < sp._bs->numBytes(); i++) put1(out.bits, sp._bs->_val[i]);
    }
  }
  put4(out.aux, n._nid);
  put4(out.aux, z.num_nodes(n._nids[0]));
  put4f(out.aux, (float)sp._n0);
  put4f(out.aux, (float)sp._n1);
  put4f(out.aux, (float)sp._p0);
  put4f(out.aux, (float)sp._p1);
  put4f(out.aux, (float)sp._se0);
  put4f(out.aux, (float)sp._se1);
  put4(out.aux, n._nids[0]);
  put4(out.aux, n._nids[1]);
  if ((z.type[nid] & 48) == 0) {  // size of the left subtree, so that going right can skip it
    const int lsz = z.size_of(n._nids[0]);
    if (lsz < 256) put1(out.bits, lsz);
    else if (lsz < 65535) put2(out.bits, lsz);
    else if (lsz < (1 << 24)) put3(out.bits, lsz);
    else put4(out.bits, lsz);
  }
  compress_node(z, n._nids[0], out);
  compress_node(z, n._nids[1], out);
  if ((int)(out.bits.size() - pos) != sz) throw std::runtime_error("DecidedNode.compress: reported size differs from the bytes written");
}

}  // namespace compress_detail

// DTree.compress (:972-982)
inline CompressedTreeBytes compressTree(const TreeOut& t) {
  using namespace compress_detail;
  CompressedTreeBytes out;
  Sizer z(t);
  if (t.nodes[0].kind == LEAF) {  // tree stump: flagged so that the reader does not look for a decision
    put1(out.bits, 0);
    put2(out.bits, 65535);
  }
  compress_node(z, 0, out);
  return out;
}

// SharedTreeMojoModel.scoreTree (:129-240), computeLeafAssignment == false, domains == null
inline float scoreCompressedTree(const std::vector<uint8_t>& tree, const std::function<double(int)>& x) {
  size_t pos = 0;
  auto get1U = [&]() { return (int)tree[pos++]; };
  auto get2 = [&]() { int v = tree[pos] | (tree[pos + 1] << 8); pos += 2; return v; };
  auto get3 = [&]() { int v = tree[pos] | (tree[pos + 1] << 8) | (tree[pos + 2] << 16); pos += 3; return v; };
  auto get4 = [&]() { int32_t v; memcpy(&v, &tree[pos], 4); pos += 4; return v; };
  auto get4f = [&]() { float v; memcpy(&v, &tree[pos], 4); pos += 4; return v; };
  while (true) {
    const int nodeType = get1U();
    const int colId = get2();
    if (colId == 65535) return get4f();
    const int naSplitDir = get1U();
    const bool naVsRest = naSplitDir == NAvsREST;
    const bool leftward = naSplitDir == NALeft || naSplitDir == Left;
    int lmask = nodeType & 51;
    const int equal = nodeType & 12;
    float splitVal = -1;
    size_t bs_pos = 0;
    int bs_nbits = 0, bs_bitoff = 0;
    if (!naVsRest) {
      if (equal == 0) splitVal = get4f();
      else if (equal == 8) {  // fill2
        bs_pos = pos; bs_nbits = 32; bs_bitoff = 0; pos += 4;
      } else {  // fill3
        bs_bitoff = get2();
        bs_nbits = get4();
        bs_pos = pos;
        pos += (size_t)(((bs_nbits - 1) >> 3) + 1);
      }
    }
    const double d = x(colId);
    auto in_range = [&](int idx) { idx -= bs_bitoff; return idx >= 0 && idx < bs_nbits; };
    auto contains = [&](int idx) { idx -= bs_bitoff; return (tree[bs_pos + (size_t)(idx >> 3)] & (uint8_t)(1u << (idx & 7))) != 0; };
    const bool has_bs = !naVsRest && equal != 0;
    const bool right = (std::isnan(d) || (equal != 0 && has_bs && !in_range(j_d2i(d))))
                           ? !leftward
                           : (!naVsRest && (equal == 0 ? d >= (double)splitVal : contains(j_d2i(d))));
    if (right) {
      switch (lmask) {
        case 0: pos += (size_t)get1U(); break;
        case 1: pos += (size_t)get2(); break;
        case 2: pos += (size_t)get3(); break;
        case 3: pos += (size_t)get4(); break;
        case 48: pos += 4; break;
        default: throw std::runtime_error("illegal lmask in compressed tree");
      }
      lmask = (nodeType & 0xC0) >> 2;
    } else if (lmask <= 3) {
      pos += (size_t)(lmask + 1);
    }
    if (lmask & 16) return get4f();
  }
}

}  // namespace h2o_oracle
