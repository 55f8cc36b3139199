// Self-contained HDF5 writer/reader for Keras-compatible weight files.
//
// libhdf5 / h5py are not available on this box, yet checkpoints must use the layout Keras'
// `load_weights` understands: root attributes
// `layer_names` / `backend` / `keras_version`, one group per layer with a `weight_names`
// attribute, datasets `<layer>/<layer>/kernel:0` ...  This file writes the classic on-disk format
// h5py produces with libver='earliest':
//   superblock v0 - v1 object headers - "old style" groups (v1 B-tree + symbol-table nodes +
//   local heap) - contiguous little-endian datasets - v1 attribute messages (numeric arrays and
//   fixed-length, null-padded strings).
// The reader understands the same subset plus object-header continuation blocks, compact
// datasets and v2/v3 attribute messages, which is what real Keras weight files contain.
//
// Reference: the reference delegates all of this to Keras/h5py (e.g.
// core/tests/examples/call_run_within_script_with_autokeras.py:48, testdata/save_and_load.py:92-125).
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "rt_common.h"
#include "cloud_b200_runtime.h"   // the public declarations must match the definitions below

namespace {

constexpr uint64_t UNDEF = 0xFFFFFFFFFFFFFFFFull;

enum DType { F32 = 0, F64 = 1, I32 = 2, I64 = 3, U8 = 4, STR = 5 };

size_t dtype_size(int dt, size_t strlen_) {
  switch (dt) {
    case F32: return 4;
    case F64: return 8;
    case I32: return 4;
    case I64: return 8;
    case U8: return 1;
    case STR: return strlen_;
  }
  return 0;
}

struct Buf {
  std::vector<uint8_t> d;
  void u8(uint8_t v) { d.push_back(v); }
  void u16(uint16_t v) { for (int i = 0; i < 2; ++i) d.push_back((v >> (8 * i)) & 0xff); }
  void u32(uint32_t v) { for (int i = 0; i < 4; ++i) d.push_back((v >> (8 * i)) & 0xff); }
  void u64(uint64_t v) { for (int i = 0; i < 8; ++i) d.push_back((v >> (8 * i)) & 0xff); }
  void bytes(const void* p, size_t n) {
    const uint8_t* b = static_cast<const uint8_t*>(p);
    d.insert(d.end(), b, b + n);
  }
  void zeros(size_t n) { d.insert(d.end(), n, 0); }
  void pad8() { while (d.size() % 8) d.push_back(0); }
  size_t size() const { return d.size(); }
};

// ---- message encoders --------------------------------------------------------------------------
Buf enc_datatype(int dt, size_t slen) {
  Buf b;
  if (dt == F32 || dt == F64) {
    const bool dbl = dt == F64;
    b.u8(0x11); b.u8(0x20); b.u8(dbl ? 0x3F : 0x1F); b.u8(0);
    b.u32(dbl ? 8 : 4);
    b.u16(0); b.u16(dbl ? 64 : 32);
    b.u8(dbl ? 52 : 23); b.u8(dbl ? 11 : 8); b.u8(0); b.u8(dbl ? 52 : 23);
    b.u32(dbl ? 1023 : 127);
  } else if (dt == I32 || dt == I64 || dt == U8) {
    const uint32_t sz = dt == I32 ? 4 : (dt == I64 ? 8 : 1);
    b.u8(0x10); b.u8(dt == U8 ? 0x00 : 0x08); b.u8(0); b.u8(0);
    b.u32(sz);
    b.u16(0); b.u16(static_cast<uint16_t>(sz * 8));
  } else {  // fixed-length string, null padded, ASCII
    b.u8(0x13); b.u8(0x01); b.u8(0); b.u8(0);
    b.u32(static_cast<uint32_t>(slen));
  }
  b.pad8();
  return b;
}

Buf enc_dataspace(const std::vector<uint64_t>& dims) {
  Buf b;
  b.u8(1); b.u8(static_cast<uint8_t>(dims.size())); b.u8(0); b.u8(0); b.u32(0);
  for (uint64_t d : dims) b.u64(d);
  b.pad8();
  return b;
}

struct Attr {
  std::string name;
  int dtype;
  size_t slen;
  std::vector<uint64_t> dims;  // empty = scalar
  std::vector<uint8_t> data;
};

Buf enc_attr(const Attr& a) {
  Buf dt = enc_datatype(a.dtype, a.slen), ds = enc_dataspace(a.dims);
  // the size fields hold the unpadded sizes; the library rounds them up itself
  Buf b;
  b.u8(1); b.u8(0);
  b.u16(static_cast<uint16_t>(a.name.size() + 1));
  size_t dt_raw = (a.dtype == F32 || a.dtype == F64) ? 20 : (a.dtype == STR ? 8 : 12);
  size_t ds_raw = 8 + 8 * a.dims.size();
  b.u16(static_cast<uint16_t>(dt_raw));
  b.u16(static_cast<uint16_t>(ds_raw));
  b.bytes(a.name.c_str(), a.name.size() + 1);
  b.pad8();
  b.bytes(dt.d.data(), dt.size());
  b.bytes(ds.d.data(), ds.size());
  b.bytes(a.data.data(), a.data.size());
  b.pad8();
  return b;
}

struct Msg { uint16_t type; Buf body; };

Buf enc_object_header(const std::vector<Msg>& msgs) {
  Buf body;
  for (const Msg& m : msgs) {
    body.u16(m.type);
    body.u16(static_cast<uint16_t>(m.body.size()));
    body.u8(0); body.u8(0); body.u8(0); body.u8(0);
    body.bytes(m.body.d.data(), m.body.size());
  }
  Buf h;
  h.u8(1); h.u8(0); h.u16(static_cast<uint16_t>(msgs.size())); h.u32(1);
  h.u32(static_cast<uint32_t>(body.size()));
  h.u32(0);  // pad the 12-byte prefix to 16
  h.bytes(body.d.data(), body.size());
  return h;
}

// ---- in-memory tree ----------------------------------------------------------------------------
struct Node {
  bool is_group = true;
  std::map<std::string, std::unique_ptr<Node>> children;  // std::map keeps strcmp order
  std::vector<Attr> attrs;
  int dtype = F32;
  std::vector<uint64_t> dims;
  std::vector<uint8_t> data;
  uint64_t header_addr = 0;
  uint64_t btree_addr = 0, heap_addr = 0;
};

struct Writer {
  Node root;
  std::string err;
};

std::vector<std::string> split_path(const std::string& p) {
  std::vector<std::string> out;
  size_t i = 0;
  while (i < p.size()) {
    while (i < p.size() && p[i] == '/') ++i;
    size_t j = i;
    while (j < p.size() && p[j] != '/') ++j;
    if (j > i) out.push_back(p.substr(i, j - i));
    i = j;
  }
  return out;
}

Node* find_node(Node* root, const std::string& path, bool create_groups) {
  Node* cur = root;
  for (const std::string& part : split_path(path)) {
    if (!cur->is_group) return nullptr;
    auto it = cur->children.find(part);
    if (it == cur->children.end()) {
      if (!create_groups) return nullptr;
      auto n = std::make_unique<Node>();
      Node* raw = n.get();
      cur->children[part] = std::move(n);
      cur = raw;
    } else {
      cur = it->second.get();
    }
  }
  return cur;
}

struct Layout {
  std::vector<uint8_t> file;
  uint16_t leaf_k = 4, internal_k = 16;

  uint64_t alloc(size_t n) {
    while (file.size() % 8) file.push_back(0);
    uint64_t a = file.size();
    file.resize(file.size() + n, 0);
    return a;
  }
  void put(uint64_t addr, const Buf& b) { memcpy(file.data() + addr, b.d.data(), b.size()); }
};

size_t max_fanout(const Node& n) {
  size_t m = n.children.size();
  for (auto& kv : n.children)
    if (kv.second->is_group) m = std::max(m, max_fanout(*kv.second));
  return m;
}

void write_node(Layout& L, Node& n);

void write_group(Layout& L, Node& g) {
  for (auto& kv : g.children) write_node(L, *kv.second);

  // local heap data: [0] empty string, then names (8-byte padded), then one free block
  Buf heap;
  heap.zeros(8);
  std::vector<uint64_t> name_off;
  for (auto& kv : g.children) {
    name_off.push_back(heap.size());
    heap.bytes(kv.first.c_str(), kv.first.size() + 1);
    heap.pad8();
  }
  const uint64_t free_off = heap.size();
  heap.u64(1);   // next free block: H5HL_FREE_NULL
  heap.u64(16);  // size of this free block
  const uint64_t heap_data_addr = L.alloc(heap.size());
  L.put(heap_data_addr, heap);
  Buf hh;
  hh.bytes("HEAP", 4); hh.u8(0); hh.zeros(3);
  hh.u64(heap.size()); hh.u64(free_off); hh.u64(heap_data_addr);
  g.heap_addr = L.alloc(hh.size());
  L.put(g.heap_addr, hh);

  // symbol table nodes, 2*leaf_k entries each
  const size_t per = 2 * L.leaf_k;
  const size_t n = g.children.size();
  const size_t n_snod = std::max<size_t>(1, (n + per - 1) / per);
  std::vector<uint64_t> snod_addr;
  std::vector<uint64_t> last_name_off;
  auto it = g.children.begin();
  size_t idx = 0;
  for (size_t s = 0; s < n_snod; ++s) {
    const size_t cnt = std::min(per, n - idx);
    Buf sn;
    sn.bytes("SNOD", 4); sn.u8(1); sn.u8(0); sn.u16(static_cast<uint16_t>(cnt));
    uint64_t last = 0;
    for (size_t e = 0; e < per; ++e) {
      if (e < cnt) {
        Node& c = *it->second;
        sn.u64(name_off[idx]);
        sn.u64(c.header_addr);
        if (c.is_group) { sn.u32(1); sn.u32(0); sn.u64(c.btree_addr); sn.u64(c.heap_addr); }
        else            { sn.u32(0); sn.u32(0); sn.zeros(16); }
        last = name_off[idx];
        ++it; ++idx;
      } else {
        sn.zeros(40);
      }
    }
    const uint64_t a = L.alloc(sn.size());
    L.put(a, sn);
    snod_addr.push_back(a);
    last_name_off.push_back(last);
  }

  // one leaf-level B-tree node pointing at the SNODs (<= 2*internal_k children guaranteed)
  Buf bt;
  bt.bytes("TREE", 4); bt.u8(0); bt.u8(0); bt.u16(static_cast<uint16_t>(n ? n_snod : 0));
  bt.u64(UNDEF); bt.u64(UNDEF);
  const size_t cap = 2 * L.internal_k;
  bt.u64(0);  // key 0: the empty string
  for (size_t c = 0; c < cap; ++c) {
    if (c < n_snod && n) { bt.u64(snod_addr[c]); bt.u64(last_name_off[c]); }
    else                 { bt.u64(0); bt.u64(0); }
  }
  g.btree_addr = L.alloc(bt.size());
  L.put(g.btree_addr, bt);

  std::vector<Msg> msgs;
  Msg st; st.type = 0x0011; st.body.u64(g.btree_addr); st.body.u64(g.heap_addr);
  msgs.push_back(std::move(st));
  for (const Attr& a : g.attrs) { Msg m; m.type = 0x000C; m.body = enc_attr(a); msgs.push_back(std::move(m)); }
  Buf oh = enc_object_header(msgs);
  g.header_addr = L.alloc(oh.size());
  L.put(g.header_addr, oh);
}

void write_dataset(Layout& L, Node& d) {
  uint64_t data_addr = UNDEF;
  if (!d.data.empty()) {
    data_addr = L.alloc(d.data.size());
    memcpy(L.file.data() + data_addr, d.data.data(), d.data.size());
  }
  std::vector<Msg> msgs;
  { Msg m; m.type = 0x0001; m.body = enc_dataspace(d.dims); msgs.push_back(std::move(m)); }
  { Msg m; m.type = 0x0003; m.body = enc_datatype(d.dtype, 0); msgs.push_back(std::move(m)); }
  { Msg m; m.type = 0x0005; m.body.u8(2); m.body.u8(2); m.body.u8(2); m.body.u8(0); m.body.pad8();
    msgs.push_back(std::move(m)); }   // fill value v2: late alloc, fill if set, undefined
  { Msg m; m.type = 0x0008; m.body.u8(3); m.body.u8(1); m.body.u64(data_addr);
    m.body.u64(d.data.size()); m.body.pad8(); msgs.push_back(std::move(m)); }
  for (const Attr& a : d.attrs) { Msg m; m.type = 0x000C; m.body = enc_attr(a); msgs.push_back(std::move(m)); }
  Buf oh = enc_object_header(msgs);
  d.header_addr = L.alloc(oh.size());
  L.put(d.header_addr, oh);
}

void write_node(Layout& L, Node& n) {
  if (n.is_group) write_group(L, n);
  else write_dataset(L, n);
}

int save(Writer& w, const char* filename) {
  Layout L;
  const size_t fan = max_fanout(w.root);
  // 2*internal_k SNODs of 2*leaf_k entries must hold the widest group in one B-tree node
  while (static_cast<size_t>(2 * L.internal_k) * 2 * L.leaf_k < fan) L.leaf_k *= 2;
  L.alloc(96);  // superblock
  write_group(L, w.root);
  const uint64_t eof = L.alloc(0);
  Buf sb;
  const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
  sb.bytes(sig, 8);
  sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(8); sb.u8(8); sb.u8(0);
  sb.u16(L.leaf_k); sb.u16(L.internal_k); sb.u32(0);
  sb.u64(0); sb.u64(UNDEF); sb.u64(eof); sb.u64(UNDEF);
  sb.u64(0); sb.u64(w.root.header_addr); sb.u32(1); sb.u32(0);
  sb.u64(w.root.btree_addr); sb.u64(w.root.heap_addr);
  L.put(0, sb);
  std::string tmp = std::string(filename) + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) { cb2rt::set_error("cannot open %s for writing", tmp.c_str()); return 1; }
  const size_t wr = fwrite(L.file.data(), 1, L.file.size(), f);
  fclose(f);
  if (wr != L.file.size()) { cb2rt::set_error("short write to %s", tmp.c_str()); return 1; }
  if (rename(tmp.c_str(), filename) != 0) { cb2rt::set_error("cannot rename %s", tmp.c_str()); return 1; }
  return 0;
}

// ---- reader ------------------------------------------------------------------------------------
struct RAttr {
  std::string name;
  int dtype = -1;     // -1: unsupported type
  size_t slen = 0;
  std::vector<uint64_t> dims;
  std::vector<uint8_t> data;
};

struct RObject {
  bool is_group = false;
  uint64_t btree = UNDEF, heap = UNDEF;
  int dtype = -1;
  size_t elem_size = 0;
  std::vector<uint64_t> dims;
  uint64_t data_addr = UNDEF, data_size = 0;
  std::vector<uint8_t> compact;
  bool is_compact = false;
  std::vector<RAttr> attrs;
};

struct Reader {
  std::vector<uint8_t> f;
  uint64_t root_header = 0;
  uint64_t base = 0;
  uint16_t leaf_k = 4, internal_k = 16;

  // Every access of the reader goes through these: a truncated or corrupted file must produce
  // "no such object" / an error, never an out-of-bounds read (csrc/runtime/test_hdf5.cc feeds it
  // hundreds of broken files under AddressSanitizer).
  bool ok(uint64_t a, uint64_t n) const { return a != UNDEF && n <= f.size() && a <= f.size() - n; }
  uint8_t at(uint64_t a) const { return a < f.size() ? f[a] : 0; }
  uint64_t rd(uint64_t a, int n) const {
    if (!ok(a, static_cast<uint64_t>(n))) return 0;
    uint64_t v = 0;
    for (int i = 0; i < n; ++i) v |= static_cast<uint64_t>(f[a + i]) << (8 * i);
    return v;
  }
  // NUL-terminated string at `a`, cut at the end of the file
  std::string cstr(uint64_t a) const {
    std::string out;
    while (a < f.size() && f[a] != 0 && out.size() < 4096) out.push_back(static_cast<char>(f[a++]));
    return out;
  }
};

bool parse_datatype(const Reader& r, uint64_t a, int* dtype, size_t* size) {
  const uint8_t cv = r.at(a);
  const int cls = cv & 0x0f;
  const uint32_t sz = static_cast<uint32_t>(r.rd(a + 4, 4));
  *size = sz;
  if (cls == 1) { *dtype = sz == 4 ? F32 : (sz == 8 ? F64 : -1); return true; }
  if (cls == 0) {
    const bool sign = (r.at(a + 1) & 0x08) != 0;
    if (sz == 4 && sign) *dtype = I32;
    else if (sz == 8 && sign) *dtype = I64;
    else if (sz == 1) *dtype = U8;
    else *dtype = -1;
    return true;
  }
  if (cls == 3) { *dtype = STR; return true; }
  *dtype = -1;
  return true;
}

size_t datatype_msg_len(const Reader& r, uint64_t a) {
  const int cls = r.at(a) & 0x0f;
  if (cls == 1) return 20;
  if (cls == 0) return 12;
  if (cls == 3) return 8;
  return 8;
}

bool parse_dataspace(const Reader& r, uint64_t a, std::vector<uint64_t>* dims, size_t* len) {
  const int ver = r.at(a);
  const int rank = r.at(a + 1);
  const int flags = r.at(a + 2);
  if (rank > 32) return false;
  uint64_t p;
  if (ver == 1) p = a + 8;
  else if (ver == 2) p = a + 4;
  else return false;
  dims->clear();
  for (int i = 0; i < rank; ++i) dims->push_back(r.rd(p + 8 * i, 8));
  size_t l = (p - a) + 8 * rank;
  if (flags & 1) l += 8 * rank;
  *len = l;
  return true;
}

void parse_attr(const Reader& r, uint64_t a, size_t msg_size, RObject* obj) {
  const int ver = r.at(a);
  RAttr at;
  uint64_t p;
  const size_t name_sz = r.rd(a + 2, 2), dt_sz = r.rd(a + 4, 2), ds_sz = r.rd(a + 6, 2);
  auto pad = [&](size_t n) { return ver == 1 ? (n + 7) / 8 * 8 : n; };
  p = a + 8;
  if (ver == 3) p += 1;  // name character-set encoding
  if (ver < 1 || ver > 3) return;
  at.name = r.cstr(p);
  p += pad(name_sz);
  size_t esz = 0;
  parse_datatype(r, p, &at.dtype, &esz);
  at.slen = esz;
  p += pad(dt_sz);
  size_t dsl = 0;
  if (!parse_dataspace(r, p, &at.dims, &dsl)) return;
  p += pad(ds_sz);
  uint64_t n = 1;
  for (uint64_t d : at.dims) n = (d != 0 && n > (1ull << 40) / d) ? (1ull << 40) : n * d;   // no overflow
  const uint64_t bytes = n * esz;
  if (bytes > msg_size + 8 || !r.ok(p, bytes) || p + bytes > a + msg_size + 8) { at.dtype = -1; }
  else at.data.assign(r.f.begin() + static_cast<long>(p), r.f.begin() + static_cast<long>(p + bytes));
  obj->attrs.push_back(std::move(at));
}

void parse_messages(const Reader& r, uint64_t a, uint64_t end, int* remaining, RObject* obj, int depth) {
  while (a + 8 <= end && *remaining > 0) {
    const uint16_t type = static_cast<uint16_t>(r.rd(a, 2));
    const uint16_t size = static_cast<uint16_t>(r.rd(a + 2, 2));
    const uint64_t body = a + 8;
    --*remaining;
    if (!r.ok(body, size)) return;
    switch (type) {
      case 0x0011:
        obj->is_group = true;
        obj->btree = r.rd(body, 8) + r.base;
        obj->heap = r.rd(body + 8, 8) + r.base;
        break;
      case 0x0001: { size_t l; parse_dataspace(r, body, &obj->dims, &l); break; }
      case 0x0003: parse_datatype(r, body, &obj->dtype, &obj->elem_size); break;
      case 0x0008: {
        const int ver = r.at(body);
        if (ver == 1 || ver == 2) {
          // libhdf5 <= 1.6 (e.g. MATLAB R2008): version, dimensionality (= rank + 1), class, 5
          // reserved bytes, data address, then `dimensionality` 32-bit sizes, the last one being the
          // element size.  Only the contiguous class is read.
          const int ndim = r.at(body + 1), cls = r.at(body + 2);
          if (cls == 1 && ndim >= 1 && ndim <= 33) {
            obj->data_addr = r.rd(body + 8, 8);
            if (obj->data_addr != UNDEF) obj->data_addr += r.base;
            uint64_t bytes = 1;
            for (int i = 0; i < ndim; ++i) {
              const uint64_t d = r.rd(body + 16 + 4 * i, 4);
              bytes = (d != 0 && bytes > (1ull << 40) / d) ? (1ull << 40) : bytes * d;
            }
            obj->data_size = bytes;
          }
        } else if (ver == 3) {
          const int cls = r.at(body + 1);
          if (cls == 1) { obj->data_addr = r.rd(body + 2, 8); obj->data_size = r.rd(body + 10, 8);
                          if (obj->data_addr != UNDEF) obj->data_addr += r.base; }
          else if (cls == 0) { const size_t n = r.rd(body + 2, 2);
                               if (r.ok(body + 4, n)) {
                                 obj->is_compact = true;
                                 obj->compact.assign(r.f.begin() + static_cast<long>(body + 4),
                                                     r.f.begin() + static_cast<long>(body + 4 + n));
                               } }
        }
        break;
      }
      case 0x000C: parse_attr(r, body, size, obj); break;
      case 0x0010: {  // continuation
        const uint64_t ca = r.rd(body, 8) + r.base, cl = r.rd(body + 8, 8);
        if (depth < 16 && r.ok(ca, cl)) parse_messages(r, ca, ca + cl, remaining, obj, depth + 1);
        break;
      }
      default: break;
    }
    a = body + size;
  }
}

bool read_object(const Reader& r, uint64_t addr, RObject* obj) {
  if (!r.ok(addr, 16) || r.at(addr) != 1) return false;
  int nmsg = static_cast<int>(r.rd(addr + 2, 2));
  const uint64_t size = r.rd(addr + 8, 4);
  parse_messages(r, addr + 16, addr + 16 + size, &nmsg, obj, 0);
  return true;
}

void walk_btree(const Reader& r, uint64_t node, uint64_t heap_data,
                std::vector<std::pair<std::string, uint64_t>>* out, int depth) {
  if (!r.ok(node, 24) || depth > 8) return;
  if (memcmp(&r.f[node], "TREE", 4) == 0) {
    const int level = r.at(node + 5);
    const int used = static_cast<int>(r.rd(node + 6, 2));
    uint64_t p = node + 24 + 8;
    for (int i = 0; i < used; ++i) {
      const uint64_t child = r.rd(p, 8) + r.base;
      (void)level;
      walk_btree(r, child, heap_data, out, depth + 1);
      p += 16;
    }
  } else if (memcmp(&r.f[node], "SNOD", 4) == 0) {
    const int cnt = static_cast<int>(r.rd(node + 6, 2));
    uint64_t p = node + 8;
    for (int i = 0; i < cnt; ++i) {
      const uint64_t noff = r.rd(p, 8), oh = r.rd(p + 8, 8) + r.base;
      if (r.ok(heap_data, 1) && noff < r.f.size() && r.ok(heap_data + noff, 1) && out->size() < 100000)
        out->push_back({r.cstr(heap_data + noff), oh});
      p += 40;
    }
  }
}

bool list_group(const Reader& r, const RObject& g, std::vector<std::pair<std::string, uint64_t>>* out) {
  if (!g.is_group || !r.ok(g.heap, 32) || memcmp(&r.f[g.heap], "HEAP", 4) != 0) return false;
  const uint64_t heap_data = r.rd(g.heap + 24, 8) + r.base;
  walk_btree(r, g.btree, heap_data, out, 0);
  return true;
}

bool resolve(const Reader& r, const std::string& path, RObject* obj) {
  uint64_t addr = r.root_header;
  RObject cur;
  if (!read_object(r, addr, &cur)) return false;
  for (const std::string& part : split_path(path)) {
    std::vector<std::pair<std::string, uint64_t>> kids;
    if (!list_group(r, cur, &kids)) return false;
    bool found = false;
    for (auto& kv : kids)
      if (kv.first == part) { addr = kv.second; found = true; break; }
    if (!found) return false;
    cur = RObject();
    if (!read_object(r, addr, &cur)) return false;
  }
  *obj = std::move(cur);
  return true;
}

int copy_out(const std::string& s, char* buf, size_t buflen) {
  if (s.size() + 1 > buflen) { cb2rt::set_error("buffer too small (%zu needed)", s.size() + 1); return -static_cast<int>(s.size() + 1); }
  memcpy(buf, s.c_str(), s.size() + 1);
  return static_cast<int>(s.size());
}

}  // namespace

extern "C" {

void* cb2h5_writer_new() { return new Writer(); }
void cb2h5_writer_free(void* w) { delete static_cast<Writer*>(w); }

int cb2h5_writer_add_group(void* wp, const char* path) {
  Writer* w = static_cast<Writer*>(wp);
  Node* n = find_node(&w->root, path, true);
  if (!n || !n->is_group) { cb2rt::set_error("cannot create group %s", path); return 1; }
  return 0;
}

int cb2h5_writer_add_dataset(void* wp, const char* path, int dtype, int rank, const uint64_t* dims,
                             const void* data) {
  Writer* w = static_cast<Writer*>(wp);
  std::vector<std::string> parts = split_path(path);
  if (parts.empty()) { cb2rt::set_error("empty dataset path"); return 1; }
  std::string parent;
  for (size_t i = 0; i + 1 < parts.size(); ++i) parent += "/" + parts[i];
  Node* g = find_node(&w->root, parent, true);
  if (!g || !g->is_group) { cb2rt::set_error("bad parent for %s", path); return 1; }
  auto n = std::make_unique<Node>();
  n->is_group = false;
  n->dtype = dtype;
  uint64_t count = 1;
  for (int i = 0; i < rank; ++i) { n->dims.push_back(dims[i]); count *= dims[i]; }
  const size_t bytes = count * dtype_size(dtype, 0);
  n->data.assign(static_cast<const uint8_t*>(data), static_cast<const uint8_t*>(data) + bytes);
  g->children[parts.back()] = std::move(n);
  return 0;
}

// strings: `n` fixed-length slots of `slen` bytes each (null padded); rank 0 => scalar string.
int cb2h5_writer_add_attr(void* wp, const char* obj_path, const char* name, int dtype, size_t slen,
                          int rank, const uint64_t* dims, const void* data) {
  Writer* w = static_cast<Writer*>(wp);
  Node* n = find_node(&w->root, obj_path, false);
  if (!n) { cb2rt::set_error("no such object %s", obj_path); return 1; }
  Attr a;
  a.name = name;
  a.dtype = dtype;
  a.slen = slen;
  uint64_t count = 1;
  for (int i = 0; i < rank; ++i) { a.dims.push_back(dims[i]); count *= dims[i]; }
  const size_t bytes = count * dtype_size(dtype, slen);
  if (bytes + 64 + strlen(name) > 64000) {
    cb2rt::set_error("attribute %s is too large for an object-header message (%zu bytes)", name, bytes);
    return 2;
  }
  a.data.assign(static_cast<const uint8_t*>(data), static_cast<const uint8_t*>(data) + bytes);
  for (auto& e : n->attrs)
    if (e.name == a.name) { e = std::move(a); return 0; }
  n->attrs.push_back(std::move(a));
  return 0;
}

int cb2h5_writer_save(void* wp, const char* filename) { return save(*static_cast<Writer*>(wp), filename); }

void* cb2h5_reader_open(const char* filename) {
  FILE* f = fopen(filename, "rb");
  if (!f) { cb2rt::set_error("cannot open %s", filename); return nullptr; }
  auto r = std::make_unique<Reader>();
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  r->f.resize(n > 0 ? n : 0);
  if (n > 0 && fread(r->f.data(), 1, n, f) != static_cast<size_t>(n)) { fclose(f); cb2rt::set_error("short read"); return nullptr; }
  fclose(f);
  const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
  // the superblock may sit at 0, 512, 1024, ... (user block)
  uint64_t sb = 0;
  bool found = false;
  for (uint64_t off = 0; off + 96 <= r->f.size(); off = off ? off * 2 : 512) {
    if (memcmp(&r->f[off], sig, 8) == 0) { sb = off; found = true; break; }
  }
  if (!found) { cb2rt::set_error("%s is not an HDF5 file", filename); return nullptr; }
  const int ver = r->f[sb + 8];
  if (ver > 1 || r->f[sb + 13] != 8 || r->f[sb + 14] != 8) {
    cb2rt::set_error("unsupported HDF5 superblock version %d (only the classic v0/v1 layout)", ver);
    return nullptr;
  }
  r->leaf_k = static_cast<uint16_t>(r->rd(sb + 16, 2));
  r->internal_k = static_cast<uint16_t>(r->rd(sb + 18, 2));
  uint64_t p = sb + 24 + (ver == 1 ? 4 : 0);
  r->base = r->rd(p, 8);
  if (r->base == UNDEF) r->base = 0;
  r->base += 0;
  // End-of-file address: relative to the base address by the format specification, but files with a
  // user block written by libhdf5 1.6/1.8 (e.g. MATLAB v7.3 files: 512-byte header, base = 512)
  // store the absolute file size here - accept either reading, reject only a file that is shorter
  // than both
  const uint64_t eof = r->rd(p + 16, 8);
  if (eof != UNDEF && eof > r->f.size()) {
    cb2rt::set_error("%s is truncated: the superblock says %llu bytes, the file has %zu", filename,
                     static_cast<unsigned long long>(eof + r->base), r->f.size());
    return nullptr;
  }
  p += 32;  // base, free-space, eof, driver
  r->root_header = r->rd(p + 8, 8) + r->base;
  RObject root;
  if (!read_object(*r, r->root_header, &root) || !root.is_group) {
    cb2rt::set_error("%s: the root group's object header is unreadable", filename);
    return nullptr;
  }
  return r.release();
}

void cb2h5_reader_close(void* r) { delete static_cast<Reader*>(r); }

// newline separated "g:<name>" / "d:<name>" in stored (sorted) order
int cb2h5_reader_list(void* rp, const char* group_path, char* buf, size_t buflen) {
  Reader* r = static_cast<Reader*>(rp);
  RObject g;
  if (!resolve(*r, group_path, &g) || !g.is_group) { cb2rt::set_error("no such group %s", group_path); return -1; }
  std::vector<std::pair<std::string, uint64_t>> kids;
  list_group(*r, g, &kids);
  std::string out;
  for (auto& kv : kids) {
    RObject o;
    read_object(*r, kv.second, &o);
    out += (o.is_group ? "g:" : "d:") + kv.first + "\n";
  }
  return copy_out(out, buf, buflen);
}

int cb2h5_reader_dataset_info(void* rp, const char* path, int* dtype, int* rank, uint64_t* dims /*[8]*/) {
  Reader* r = static_cast<Reader*>(rp);
  RObject o;
  if (!resolve(*r, path, &o) || o.is_group) { cb2rt::set_error("no such dataset %s", path); return 1; }
  *dtype = o.dtype;
  *rank = static_cast<int>(o.dims.size());
  for (size_t i = 0; i < o.dims.size() && i < 8; ++i) dims[i] = o.dims[i];
  return 0;
}

int cb2h5_reader_read_dataset(void* rp, const char* path, void* out, size_t nbytes) {
  Reader* r = static_cast<Reader*>(rp);
  RObject o;
  if (!resolve(*r, path, &o) || o.is_group) { cb2rt::set_error("no such dataset %s", path); return 1; }
  uint64_t n = 1;
  for (uint64_t d : o.dims) n *= d;
  const size_t need = n * o.elem_size;
  if (need != nbytes) { cb2rt::set_error("dataset %s holds %zu bytes, buffer has %zu", path, need, nbytes); return 1; }
  if (o.is_compact) {
    if (o.compact.size() < need) { cb2rt::set_error("dataset %s: compact data shorter than its shape", path); return 1; }
    memcpy(out, o.compact.data(), need);
    return 0;
  }
  if (need == 0) return 0;
  if (!r->ok(o.data_addr, need)) { cb2rt::set_error("dataset %s is not contiguous (chunked/unallocated data is not supported)", path); return 1; }
  memcpy(out, &r->f[o.data_addr], need);
  return 0;
}

int cb2h5_reader_attr_names(void* rp, const char* obj_path, char* buf, size_t buflen) {
  Reader* r = static_cast<Reader*>(rp);
  RObject o;
  if (!resolve(*r, obj_path, &o)) { cb2rt::set_error("no such object %s", obj_path); return -1; }
  std::string out;
  for (auto& a : o.attrs) out += a.name + "\n";
  return copy_out(out, buf, buflen);
}

int cb2h5_reader_attr_info(void* rp, const char* obj_path, const char* name, int* dtype, size_t* slen,
                           int* rank, uint64_t* dims /*[8]*/, size_t* nbytes) {
  Reader* r = static_cast<Reader*>(rp);
  RObject o;
  if (!resolve(*r, obj_path, &o)) { cb2rt::set_error("no such object %s", obj_path); return 1; }
  for (auto& a : o.attrs) {
    if (a.name != name) continue;
    *dtype = a.dtype; *slen = a.slen; *rank = static_cast<int>(a.dims.size());
    for (size_t i = 0; i < a.dims.size() && i < 8; ++i) dims[i] = a.dims[i];
    *nbytes = a.data.size();
    return 0;
  }
  cb2rt::set_error("object %s has no attribute %s", obj_path, name);
  return 1;
}

int cb2h5_reader_attr_read(void* rp, const char* obj_path, const char* name, void* out, size_t nbytes) {
  Reader* r = static_cast<Reader*>(rp);
  RObject o;
  if (!resolve(*r, obj_path, &o)) { cb2rt::set_error("no such object %s", obj_path); return 1; }
  for (auto& a : o.attrs) {
    if (a.name != name) continue;
    if (a.data.size() != nbytes) { cb2rt::set_error("attribute size mismatch"); return 1; }
    memcpy(out, a.data.data(), nbytes);
    return 0;
  }
  cb2rt::set_error("object %s has no attribute %s", obj_path, name);
  return 1;
}

}  // extern "C"
