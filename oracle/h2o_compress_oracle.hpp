// TEST INFRASTRUCTURE ONLY (see oracle/h2o_tree_oracle.hpp). CPU restatement of the reference's tree serialisation:
//   DecidedNode.size / compress, LeafNode.compress, DTree.compress   h2o-algos/src/main/java/hex/tree/DTree.java:845-933, 950-956, 972-982
//   IcedBitSet.compress2 / compress3                                  h2o-core/src/main/java/water/util/IcedBitSet.java:92-111
// and of the reader the MOJO / in-cluster scoring runs over those bytes:
//   SharedTreeMojoModel.scoreTree                                     h2o-genmodel/src/main/java/hex/genmodel/algos/tree/SharedTreeMojoModel.java:129-240
//   GenmodelBitSet.fill2 / fill3 / isInRange / contains               h2o-genmodel/src/main/java/hex/genmodel/utils/GenmodelBitSet.java
// Byte order: AutoBuffer and ByteBufferWrapper both use the native order (little endian here).
// Parity status: pinned by bytes the reference wrote itself. The 770 GBM / DRF / isolation-forest trees among the reference's MOJO
// test fixtures (h2o-genmodel/src/test/resources/hex/genmodel: mojo.zip, algos/ensemble/*.zip, algos/isofor/trees; 28 260 decided
// nodes: 23 452 numeric, 3 965 32-bit bitsets, 664 long bitsets, 179 NA-vs-rest) are decoded with decompressTree below and serialised
// again with compressTree: bits and auxiliary records come out byte for byte (tests/test_oracle_compress_auc.py, run where the
// reference tree is present); the scoreTree restatement reproduces GbmMojoModelTest.testScore0's known answer on the reference's
// calibrated GBM MOJO. On top of that the reference's own invariants: size() == bytes written (the assert at
// DTree.java:929) and scoreTree over the bytes == the node-table walk for every row (CompressedTreeTest.testNodeIdAssignment).
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

// The inverse of compressTree: DTree.compress's two byte streams back into a node table (ids from the auxiliary records, which
// follow the decided nodes in the same pre-order). It exists so that trees serialised BY THE REFERENCE (the MOJO fixtures under
// h2o-genmodel/src/test/resources) can be sent through compressTree again: decode, re-encode, compare byte for byte.
// `reserved` receives every auxiliary record's second word, which old MOJOs fill with the parent id instead of the node count
// (SharedTreeMojoModel.AuxInfo: "depending on a MOJO version").
inline TreeOut decompressTree(const std::vector<uint8_t>& bits, const std::vector<uint8_t>& aux, std::vector<int>* reserved) {
  struct D {
    const std::vector<uint8_t>& b;
    const std::vector<uint8_t>& a;
    std::vector<int>* reserved;
    size_t pos = 0, apos = 0;
    TreeOut t;
    int u1() { if (pos + 1 > b.size()) throw std::runtime_error("compressed tree: truncated"); return b[pos++]; }
    int u2() { const int v = u1(); return v | (u1() << 8); }
    int u3() { const int v = u2(); return v | (u1() << 16); }
    int32_t i4() { const int lo = u2(); return (int32_t)((uint32_t)lo | ((uint32_t)u2() << 16)); }
    float f4() { const int32_t v = i4(); float f; memcpy(&f, &v, 4); return f; }
    int32_t a4() { if (apos + 4 > a.size()) throw std::runtime_error("auxiliary tree info: truncated"); int32_t v; memcpy(&v, &a[apos], 4); apos += 4; return v; }
    float af() { const int32_t v = a4(); float f; memcpy(&f, &v, 4); return f; }
    Node& at(int nid) {
      if (nid < 0 || nid > (1 << 24)) throw std::runtime_error("compressed tree: bad node id");
      if ((size_t)nid >= t.nodes.size()) t.nodes.resize((size_t)nid + 1);
      return t.nodes[(size_t)nid];
    }
    void leaf(int nid, int pid) {
      const float v = f4();
      Node& n = at(nid);
      n.kind = LEAF;
      n._nid = nid;
      n._pid = pid;
      n._pred = v;
    }
    int decided(int pid) {
      const int type = u1();
      Split sp;
      sp._col = u2();
      sp._nasplit = u1();
      const int eq = type & 12;
      sp._equal = (int8_t)(eq == 0 ? 0 : (eq == 4 ? 1 : (eq == 8 ? 2 : 3)));
      float splat = NAN;
      if (sp._nasplit != NAvsREST) {
        if (sp._equal <= 1) splat = f4();
        else {
          IcedBitSet bs;
          if (sp._equal == 2) { bs._bitoff = 0; bs._nbits = 32; }
          else { bs._bitoff = u2(); bs._nbits = i4(); }
          if (bs._nbits < 1 || bs._nbits > (1 << 24)) throw std::runtime_error("compressed tree: bad bitset");
          bs._val.resize((size_t)IcedBitSet::bytes(bs._nbits));
          for (uint8_t& v : bs._val) v = (uint8_t)u1();
          sp._bs = bs;
        }
      }
      const int nid = a4();
      if (reserved) reserved->push_back(a4()); else a4();
      sp._n0 = af(); sp._n1 = af(); sp._p0 = af(); sp._p1 = af(); sp._se0 = af(); sp._se1 = af();
      const int nid0 = a4(), nid1 = a4();
      if ((type & 48) == 0) {  // size of the left subtree: checked against what the left subtree turns out to take
        const int k = type & 3;
        const int lsz = k == 0 ? u1() : (k == 1 ? u2() : (k == 2 ? u3() : i4()));
        const size_t before = pos;
        if (decided(nid) != nid0) throw std::runtime_error("compressed tree: left child id differs from the auxiliary record");
        if ((int)(pos - before) != lsz) throw std::runtime_error("compressed tree: left subtree size field is wrong");
      } else if ((type & 48) == 48) leaf(nid0, nid);
      else throw std::runtime_error("compressed tree: illegal left mask");
      if ((type & 0xC0) == 0xC0) leaf(nid1, nid);
      else if ((type & 0xC0) == 0) {
        if (decided(nid) != nid1) throw std::runtime_error("compressed tree: right child id differs from the auxiliary record");
      } else throw std::runtime_error("compressed tree: illegal right mask");
      Node& n = at(nid);
      n.kind = DECIDED;
      n._nid = nid;
      n._pid = pid;
      n._split = sp;
      n._splat = splat;
      n._nids[0] = nid0;
      n._nids[1] = nid1;
      return nid;
    }
  } d{bits, aux, reserved};
  if (bits.size() >= 3 && bits[1] == 0xFF && bits[2] == 0xFF) {  // stump: {0, 65535, leaf value}
    d.pos = 3;
    d.leaf(0, NO_PARENT);
  } else {
    d.decided(NO_PARENT);
  }
  if (d.pos != bits.size() || d.apos != aux.size()) throw std::runtime_error("compressed tree: bytes left over");
  return d.t;
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
