// Minimal HDF5 writer + reader (no libhdf5 on this image).
//
// Capability parity: the reference checkpoints through Keras -> h5py -> libhdf5
// (mnist_keras.py:101-104,118-119,124; tensorflow2_keras_mnist.py:86-88).  This file
// writes the subset of the HDF5 1.8 "earliest" file format those files use, which is
// also what h5py emits by default:
//   * superblock version 0, 8-byte offsets/lengths
//   * version-1 object headers (single chunk on write; continuation blocks on read)
//   * old-style groups: symbol-table message -> v1 B-tree ("TREE") -> symbol nodes
//     ("SNOD") + local heap ("HEAP") for names
//   * contiguous, unfiltered datasets; float32/float64/int32/int64 little-endian
//   * attributes: fixed-length (null-padded) strings, scalar or 1-D, and numeric arrays
// The writer builds the object tree in memory and serialises post-order so every child
// address is known before its parent's symbol node is emitted.
//
// C ABI only (ctypes binds it); declarations live in include/hvd_b200.h.

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace {

thread_local std::string g_err;
void set_err(const std::string& s) { g_err = s; }

constexpr uint64_t UNDEF = 0xFFFFFFFFFFFFFFFFull;
constexpr int LEAF_K = 32;       // symbol node holds up to 2*LEAF_K entries
constexpr int INTERNAL_K = 16;   // B-tree node holds up to 2*INTERNAL_K children

enum DType { DT_F32 = 0, DT_I64 = 1, DT_F64 = 2, DT_I32 = 3, DT_STR = 4 };

int64_t dtype_size(int dt, int64_t strw) {
  switch (dt) {
    case DT_F32: return 4; case DT_I64: return 8; case DT_F64: return 8; case DT_I32: return 4;
    case DT_STR: return strw;
  }
  return 0;
}

struct Buf {
  std::vector<uint8_t> d;
  void u8(uint8_t v) { d.push_back(v); }
  void u16(uint16_t v) { for (int i = 0; i < 2; i++) d.push_back(uint8_t(v >> (8 * i))); }
  void u32(uint32_t v) { for (int i = 0; i < 4; i++) d.push_back(uint8_t(v >> (8 * i))); }
  void u64(uint64_t v) { for (int i = 0; i < 8; i++) d.push_back(uint8_t(v >> (8 * i))); }
  void bytes(const void* p, size_t n) { const uint8_t* b = (const uint8_t*)p; d.insert(d.end(), b, b + n); }
  void zeros(size_t n) { d.insert(d.end(), n, 0); }
  void pad8() { while (d.size() % 8) d.push_back(0); }
  size_t size() const { return d.size(); }
};

// ---- message encoders --------------------------------------------------------------
void enc_datatype(Buf& b, int dt, int64_t strw) {
  switch (dt) {
    case DT_F32:
      b.u8(0x11); b.u8(0x20); b.u8(0x1f); b.u8(0x00); b.u32(4);
      b.u16(0); b.u16(32); b.u8(23); b.u8(8); b.u8(0); b.u8(23); b.u32(127);
      break;
    case DT_F64:
      b.u8(0x11); b.u8(0x20); b.u8(0x3f); b.u8(0x00); b.u32(8);
      b.u16(0); b.u16(64); b.u8(52); b.u8(11); b.u8(0); b.u8(52); b.u32(1023);
      break;
    case DT_I64:
      b.u8(0x10); b.u8(0x08); b.u8(0x00); b.u8(0x00); b.u32(8); b.u16(0); b.u16(64);
      break;
    case DT_I32:
      b.u8(0x10); b.u8(0x08); b.u8(0x00); b.u8(0x00); b.u32(4); b.u16(0); b.u16(32);
      break;
    case DT_STR:
      b.u8(0x13); b.u8(0x01); b.u8(0x00); b.u8(0x00); b.u32(uint32_t(strw));   // null-padded ASCII
      break;
  }
}

void enc_dataspace(Buf& b, const std::vector<int64_t>& dims) {
  b.u8(1); b.u8(uint8_t(dims.size())); b.u8(0); b.u8(0); b.u32(0);
  for (auto v : dims) b.u64(uint64_t(v));
}

struct Attr {
  std::string name; int dt; int64_t strw; std::vector<int64_t> dims; std::vector<uint8_t> data;
};

struct Node {
  bool is_group = true;
  std::map<std::string, std::unique_ptr<Node>> children;     // sorted by name (strcmp order)
  std::vector<Attr> attrs;
  // dataset
  int dt = DT_F32; std::vector<int64_t> dims; std::vector<uint8_t> data;
  // filled during serialisation
  uint64_t header_addr = 0, btree_addr = 0, heap_addr = 0;
};

struct Writer {
  Node root;
  Buf out;
  Node* find(const std::string& path, bool create_groups) {
    Node* cur = &root;
    size_t i = 0;
    while (i < path.size()) {
      while (i < path.size() && path[i] == '/') i++;
      if (i >= path.size()) break;
      size_t j = path.find('/', i);
      if (j == std::string::npos) j = path.size();
      std::string part = path.substr(i, j - i);
      auto it = cur->children.find(part);
      if (it == cur->children.end()) {
        if (!create_groups) return nullptr;
        if (!cur->is_group) return nullptr;
        cur->children[part] = std::make_unique<Node>();
        it = cur->children.find(part);
      }
      cur = it->second.get();
      i = j;
    }
    return cur;
  }

  uint64_t alloc(const Buf& b) {     // append a block, 8-byte aligned; return its address
    out.pad8();
    uint64_t a = out.size();
    out.bytes(b.d.data(), b.d.size());
    return a;
  }

  void msg(Buf& hdr, uint16_t type, const Buf& body, int& count) {
    Buf p = body; p.pad8();
    hdr.u16(type); hdr.u16(uint16_t(p.size())); hdr.u8(0); hdr.u8(0); hdr.u8(0); hdr.u8(0);
    hdr.bytes(p.d.data(), p.d.size());
    count++;
  }

  bool attr_msgs(Buf& hdr, const Node& n, int& count) {
    for (const auto& a : n.attrs) {
      Buf dtb, dsb; enc_datatype(dtb, a.dt, a.strw); enc_dataspace(dsb, a.dims);
      Buf m; m.u8(1); m.u8(0); m.u16(uint16_t(a.name.size() + 1)); m.u16(uint16_t(dtb.size())); m.u16(uint16_t(dsb.size()));
      m.bytes(a.name.c_str(), a.name.size() + 1); m.pad8();
      m.bytes(dtb.d.data(), dtb.size()); m.pad8();
      m.bytes(dsb.d.data(), dsb.size()); m.pad8();
      m.bytes(a.data.data(), a.data.size());
      if (m.size() + 8 > 65528) { set_err("attribute '" + a.name + "' exceeds the 64 KiB header-message limit"); return false; }
      msg(hdr, 0x000C, m, count);
    }
    return true;
  }

  uint64_t write_header(const Buf& msgs, int count) {
    Buf h; h.u8(1); h.u8(0); h.u16(uint16_t(count)); h.u32(1); h.u32(uint32_t(msgs.size())); h.u32(0);
    h.bytes(msgs.d.data(), msgs.size());
    return alloc(h);
  }

  bool serialise(Node& n) {
    if (!n.is_group) {
      Buf raw; raw.bytes(n.data.data(), n.data.size());
      uint64_t data_addr = n.data.empty() ? UNDEF : alloc(raw);
      Buf msgs; int count = 0;
      { Buf b; enc_dataspace(b, n.dims); msg(msgs, 0x0001, b, count); }
      { Buf b; enc_datatype(b, n.dt, 0); msg(msgs, 0x0003, b, count); }
      { Buf b; b.u8(2); b.u8(2); b.u8(2); b.u8(0); msg(msgs, 0x0005, b, count); }      // fill value v2, undefined
      { Buf b; b.u8(3); b.u8(1); b.u64(data_addr); b.u64(n.data.size()); msg(msgs, 0x0008, b, count); }
      if (!attr_msgs(msgs, n, count)) return false;
      n.header_addr = write_header(msgs, count);
      return true;
    }
    for (auto& kv : n.children) if (!serialise(*kv.second)) return false;
    // local heap: offset 0 = "", then each child name (8-byte padded)
    Buf heapdata; heapdata.zeros(8);
    std::vector<uint64_t> name_off;
    for (auto& kv : n.children) {
      name_off.push_back(heapdata.size());
      heapdata.bytes(kv.first.c_str(), kv.first.size() + 1); heapdata.pad8();
    }
    uint64_t heapdata_addr = alloc(heapdata);
    { Buf h; h.bytes("HEAP", 4); h.u8(0); h.zeros(3); h.u64(heapdata.size()); h.u64(1 /*H5HL_FREE_NULL*/); h.u64(heapdata_addr);
      n.heap_addr = alloc(h); }
    // symbol nodes
    std::vector<Node*> kids; for (auto& kv : n.children) kids.push_back(kv.second.get());
    size_t per = 2 * LEAF_K;
    std::vector<uint64_t> snod_addr, last_name_off;
    for (size_t s = 0; s < kids.size() || (s == 0 && kids.empty()); s += per) {
      size_t e = std::min(kids.size(), s + per);
      Buf sn; sn.bytes("SNOD", 4); sn.u8(1); sn.u8(0); sn.u16(uint16_t(e - s));
      for (size_t i = s; i < e; i++) {
        sn.u64(name_off[i]); sn.u64(kids[i]->header_addr);
        if (kids[i]->is_group) { sn.u32(1); sn.u32(0); sn.u64(kids[i]->btree_addr); sn.u64(kids[i]->heap_addr); }
        else { sn.u32(0); sn.u32(0); sn.zeros(16); }
      }
      sn.zeros((per - (e - s)) * 40);
      snod_addr.push_back(alloc(sn));
      last_name_off.push_back(e > s ? name_off[e - 1] : 0);
      if (kids.empty()) break;
    }
    if (snod_addr.size() > size_t(2 * INTERNAL_K)) { set_err("group has too many entries for a single-level B-tree"); return false; }
    { Buf t; t.bytes("TREE", 4); t.u8(0); t.u8(0); t.u16(uint16_t(kids.empty() ? 0 : snod_addr.size())); t.u64(UNDEF); t.u64(UNDEF);
      t.u64(0);                                             // key 0: the empty string
      size_t used = kids.empty() ? 0 : snod_addr.size();
      for (size_t i = 0; i < used; i++) { t.u64(snod_addr[i]); t.u64(last_name_off[i]); }
      size_t full = 24 + (2 * INTERNAL_K + 1) * 8 + (2 * INTERNAL_K) * 8;
      if (t.size() < full) t.zeros(full - t.size());
      n.btree_addr = alloc(t); }
    Buf msgs; int count = 0;
    { Buf b; b.u64(n.btree_addr); b.u64(n.heap_addr); msg(msgs, 0x0011, b, count); }
    if (!attr_msgs(msgs, n, count)) return false;
    n.header_addr = write_header(msgs, count);
    return true;
  }

  bool write(const char* fn) {
    out.d.clear();
    out.zeros(96);                                          // superblock placeholder
    if (!serialise(root)) return false;
    out.pad8();
    Buf sb; sb.bytes("\x89HDF\r\n\x1a\n", 8);
    sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(8); sb.u8(8); sb.u8(0);
    sb.u16(LEAF_K); sb.u16(INTERNAL_K); sb.u32(0);
    sb.u64(0); sb.u64(UNDEF); sb.u64(out.size()); sb.u64(UNDEF);
    sb.u64(0); sb.u64(root.header_addr); sb.u32(1); sb.u32(0); sb.u64(root.btree_addr); sb.u64(root.heap_addr);
    std::memcpy(out.d.data(), sb.d.data(), 96);
    std::string tmp = std::string(fn) + ".tmp";
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) { set_err(std::string("cannot open ") + tmp); return false; }
    size_t w = std::fwrite(out.d.data(), 1, out.size(), f);
    std::fclose(f);
    if (w != out.size()) { set_err("short write"); return false; }
    if (std::rename(tmp.c_str(), fn) != 0) { set_err("rename failed"); return false; }   // atomic replace
    return true;
  }
};

// ------------------------------------------------------------------------ reader
struct RObj {
  bool is_group = false;
  uint64_t btree = UNDEF, heap = UNDEF;
  int dt = -1; int64_t elsize = 0; std::vector<int64_t> dims; uint64_t data_addr = UNDEF, data_size = 0;
  std::vector<Attr> attrs;
};

struct Reader {
  std::vector<uint8_t> f;
  uint64_t root_header = 0;
  int leaf_k = 4, internal_k = 16;

  bool ok(uint64_t off, uint64_t n) const { return off != UNDEF && off + n <= f.size(); }
  uint64_t rd(uint64_t off, int n) const { uint64_t v = 0; for (int i = 0; i < n; i++) v |= uint64_t(f[off + i]) << (8 * i); return v; }

  bool open(const char* fn) {
    FILE* fp = std::fopen(fn, "rb");
    if (!fp) { set_err(std::string("cannot open ") + fn); return false; }
    std::fseek(fp, 0, SEEK_END); long sz = std::ftell(fp); std::fseek(fp, 0, SEEK_SET);
    f.resize(size_t(sz));
    size_t r = std::fread(f.data(), 1, f.size(), fp); std::fclose(fp);
    if (r != f.size() || f.size() < 96 || std::memcmp(f.data(), "\x89HDF\r\n\x1a\n", 8) != 0) { set_err("not an HDF5 file"); return false; }
    if (f[8] != 0 && f[8] != 1) { set_err("unsupported superblock version (only 0/1)"); return false; }
    if (f[13] != 8 || f[14] != 8) { set_err("only 8-byte offsets/lengths supported"); return false; }
    leaf_k = int(rd(16, 2)); internal_k = int(rd(18, 2));
    uint64_t p = 24 + (f[8] == 1 ? 4 : 0);
    p += 32;                                                  // base, freespace, eof, driver
    root_header = rd(p + 8, 8);
    return true;
  }

  bool parse_datatype(uint64_t p, int& dt, int64_t& elsize) const {
    int cls = f[p] & 0x0f; uint32_t size = uint32_t(rd(p + 4, 4)); elsize = size;
    if (cls == 1) dt = (size == 4) ? DT_F32 : DT_F64;
    else if (cls == 0) dt = (size == 8) ? DT_I64 : DT_I32;
    else if (cls == 3) dt = DT_STR;
    else { set_err("unsupported datatype class " + std::to_string(cls)); return false; }
    return true;
  }

  void parse_dataspace(uint64_t p, std::vector<int64_t>& dims) const {
    int ver = f[p], rank = f[p + 1], flags = f[p + 2];
    (void)flags;
    uint64_t q = p + (ver == 1 ? 8 : 4);
    dims.clear();
    for (int i = 0; i < rank; i++) dims.push_back(int64_t(rd(q + 8 * i, 8)));
  }

  bool parse_msgs(uint64_t p, uint64_t end, RObj& o, std::vector<std::pair<uint64_t, uint64_t>>& conts) const {
    while (p + 8 <= end) {
      uint16_t type = uint16_t(rd(p, 2)), size = uint16_t(rd(p + 2, 2));
      uint64_t b = p + 8;
      if (b + size > f.size()) { set_err("truncated header message"); return false; }
      switch (type) {
        case 0x0001: parse_dataspace(b, o.dims); break;
        case 0x0003: if (!parse_datatype(b, o.dt, o.elsize)) return false; break;
        case 0x0008: {
          int ver = f[b];
          if (ver == 3) {
            if (f[b + 1] != 1) { set_err("only contiguous datasets are supported"); return false; }
            o.data_addr = rd(b + 2, 8); o.data_size = rd(b + 10, 8);
          } else { set_err("unsupported data-layout version"); return false; }
          break; }
        case 0x000B: set_err("filtered datasets are not supported"); return false;
        case 0x000C: {
          Attr a; int ver = f[b];
          if (ver != 1) { set_err("unsupported attribute message version " + std::to_string(ver)); return false; }
          uint64_t ns = rd(b + 2, 2), ts = rd(b + 4, 2), ss = rd(b + 6, 2);
          auto r8 = [](uint64_t v) { return (v + 7) & ~uint64_t(7); };
          uint64_t q = b + 8;
          a.name = std::string((const char*)&f[q]); q += r8(ns);
          if (!parse_datatype(q, a.dt, a.strw)) return false; q += r8(ts);
          parse_dataspace(q, a.dims); q += r8(ss);
          int64_t n = 1; for (auto d : a.dims) n *= d;
          uint64_t nb = uint64_t(n * a.strw);
          if (!ok(q, nb)) { set_err("truncated attribute"); return false; }
          a.data.assign(f.begin() + q, f.begin() + q + nb);
          o.attrs.push_back(std::move(a));
          break; }
        case 0x0010: conts.emplace_back(rd(b, 8), rd(b + 8, 8)); break;
        case 0x0011: o.is_group = true; o.btree = rd(b, 8); o.heap = rd(b + 8, 8); break;
        default: break;
      }
      p = b + size;
    }
    return true;
  }

  bool read_obj(uint64_t addr, RObj& o) const {
    if (!ok(addr, 16)) { set_err("bad object header address"); return false; }
    if (f[addr] != 1) { set_err("only version-1 object headers are supported"); return false; }
    uint64_t hsize = rd(addr + 8, 4);
    std::vector<std::pair<uint64_t, uint64_t>> conts;
    if (!parse_msgs(addr + 16, addr + 16 + hsize, o, conts)) return false;
    for (size_t i = 0; i < conts.size(); i++) {
      auto c = conts[i];
      if (!ok(c.first, c.second)) { set_err("bad continuation block"); return false; }
      if (!parse_msgs(c.first, c.first + c.second, o, conts)) return false;
    }
    return true;
  }

  bool children(const RObj& g, std::vector<std::pair<std::string, uint64_t>>& out) const {
    if (!ok(g.heap, 32) || std::memcmp(&f[g.heap], "HEAP", 4) != 0) { set_err("bad local heap"); return false; }
    uint64_t heapdata = rd(g.heap + 24, 8);
    return walk_btree(g.btree, heapdata, out);
  }

  bool walk_btree(uint64_t addr, uint64_t heapdata, std::vector<std::pair<std::string, uint64_t>>& out) const {
    if (!ok(addr, 24) || std::memcmp(&f[addr], "TREE", 4) != 0) { set_err("bad B-tree node"); return false; }
    int level = f[addr + 5]; int used = int(rd(addr + 6, 2));
    uint64_t p = addr + 24;
    for (int i = 0; i < used; i++) {
      uint64_t child = rd(p + 8, 8);
      if (level > 0) { if (!walk_btree(child, heapdata, out)) return false; }
      else {
        if (!ok(child, 8) || std::memcmp(&f[child], "SNOD", 4) != 0) { set_err("bad symbol node"); return false; }
        int ns = int(rd(child + 6, 2));
        for (int k = 0; k < ns; k++) {
          uint64_t e = child + 8 + 40 * uint64_t(k);
          out.emplace_back(std::string((const char*)&f[heapdata + rd(e, 8)]), rd(e + 8, 8));
        }
      }
      p += 16;
    }
    return true;
  }

  bool resolve(const std::string& path, RObj& o) const {
    uint64_t addr = root_header;
    if (!read_obj(addr, o)) return false;
    size_t i = 0;
    while (i < path.size()) {
      while (i < path.size() && path[i] == '/') i++;
      if (i >= path.size()) break;
      size_t j = path.find('/', i); if (j == std::string::npos) j = path.size();
      std::string part = path.substr(i, j - i);
      if (!o.is_group) { set_err("'" + part + "': parent is not a group"); return false; }
      std::vector<std::pair<std::string, uint64_t>> ch;
      if (!children(o, ch)) return false;
      bool found = false;
      for (auto& c : ch) if (c.first == part) { addr = c.second; found = true; break; }
      if (!found) { set_err("no such object: " + path); return false; }
      o = RObj();
      if (!read_obj(addr, o)) return false;
      i = j;
    }
    return true;
  }
};

int64_t pack_names(const std::vector<std::string>& names, char* out, int64_t cap) {
  std::string s;
  for (size_t i = 0; i < names.size(); i++) { if (i) s += '\n'; s += names[i]; }
  if (out && int64_t(s.size()) + 1 <= cap) std::memcpy(out, s.c_str(), s.size() + 1);
  return int64_t(s.size()) + 1;
}

}  // namespace

extern "C" {

const char* hvd_h5_last_error() { return g_err.c_str(); }

void* hvd_h5w_create() { return new Writer(); }
void hvd_h5w_free(void* w) { delete (Writer*)w; }

int hvd_h5w_group(void* w, const char* path) {
  Node* n = ((Writer*)w)->find(path, true);
  if (!n || !n->is_group) { set_err(std::string("cannot create group ") + path); return -1; }
  return 0;
}

int hvd_h5w_dataset(void* w, const char* path, int dtype, int rank, const int64_t* dims, const void* data) {
  std::string p(path); size_t k = p.rfind('/');
  std::string parent = (k == std::string::npos) ? "" : p.substr(0, k);
  std::string leaf = (k == std::string::npos) ? p : p.substr(k + 1);
  if (dtype == DT_STR || dtype_size(dtype, 0) == 0) { set_err("bad dataset dtype"); return -1; }
  Node* g = ((Writer*)w)->find(parent, true);
  if (!g || !g->is_group || leaf.empty()) { set_err(std::string("bad dataset path ") + path); return -1; }
  auto n = std::make_unique<Node>();
  n->is_group = false; n->dt = dtype; n->dims.assign(dims, dims + rank);
  int64_t cnt = 1; for (int i = 0; i < rank; i++) cnt *= dims[i];
  size_t nb = size_t(cnt * dtype_size(dtype, 0));
  n->data.assign((const uint8_t*)data, (const uint8_t*)data + nb);
  g->children[leaf] = std::move(n);
  return 0;
}

static int add_attr(void* w, const char* obj, Attr&& a) {
  Node* n = ((Writer*)w)->find(obj, false);
  if (!n) { set_err(std::string("no such object for attribute: ") + obj); return -1; }
  for (auto& e : n->attrs) if (e.name == a.name) { e = std::move(a); return 0; }
  n->attrs.push_back(std::move(a));
  return 0;
}

int hvd_h5w_attr_str(void* w, const char* obj, const char* name, const char* data, int64_t len) {
  Attr a; a.name = name; a.dt = DT_STR; a.strw = std::max<int64_t>(len, 1);
  a.data.assign(size_t(a.strw), 0); std::memcpy(a.data.data(), data, size_t(len));
  return add_attr(w, obj, std::move(a));
}

int hvd_h5w_attr_strs(void* w, const char* obj, const char* name, const char* packed, int64_t n, int64_t width) {
  Attr a; a.name = name; a.dt = DT_STR; a.strw = std::max<int64_t>(width, 1); a.dims = {n};
  a.data.assign((const uint8_t*)packed, (const uint8_t*)packed + size_t(n * a.strw));
  return add_attr(w, obj, std::move(a));
}

int hvd_h5w_attr_num(void* w, const char* obj, const char* name, int dtype, int rank, const int64_t* dims, const void* data) {
  Attr a; a.name = name; a.dt = dtype; a.strw = dtype_size(dtype, 0); a.dims.assign(dims, dims + rank);
  int64_t cnt = 1; for (int i = 0; i < rank; i++) cnt *= dims[i];
  a.data.assign((const uint8_t*)data, (const uint8_t*)data + size_t(cnt * a.strw));
  return add_attr(w, obj, std::move(a));
}

int hvd_h5w_write(void* w, const char* filename) { return ((Writer*)w)->write(filename) ? 0 : -1; }

void* hvd_h5r_open(const char* filename) {
  auto* r = new Reader();
  if (!r->open(filename)) { delete r; return nullptr; }
  return r;
}
void hvd_h5r_close(void* r) { delete (Reader*)r; }

// 0 = missing, 1 = group, 2 = dataset
int hvd_h5r_kind(void* r, const char* path) {
  RObj o; if (!((Reader*)r)->resolve(path, o)) return 0;
  return o.is_group ? 1 : 2;
}

int64_t hvd_h5r_list(void* r, const char* path, char* out, int64_t cap) {
  RObj o; if (!((Reader*)r)->resolve(path, o)) return -1;
  if (!o.is_group) { set_err("not a group"); return -1; }
  std::vector<std::pair<std::string, uint64_t>> ch;
  if (!((Reader*)r)->children(o, ch)) return -1;
  std::vector<std::string> names; for (auto& c : ch) names.push_back(c.first);
  return pack_names(names, out, cap);
}

int hvd_h5r_dataset_info(void* r, const char* path, int* dtype, int* rank, int64_t* dims, int64_t* elsize) {
  RObj o; if (!((Reader*)r)->resolve(path, o)) return -1;
  if (o.is_group || o.dt < 0) { set_err("not a dataset"); return -1; }
  *dtype = o.dt; *rank = int(o.dims.size()); *elsize = o.elsize;
  for (size_t i = 0; i < o.dims.size() && i < 8; i++) dims[i] = o.dims[i];
  return 0;
}

int hvd_h5r_dataset_read(void* r, const char* path, void* out, int64_t nbytes) {
  Reader* rd = (Reader*)r; RObj o; if (!rd->resolve(path, o)) return -1;
  int64_t cnt = 1; for (auto d : o.dims) cnt *= d;
  int64_t need = cnt * o.elsize;
  if (need != nbytes) { set_err("dataset size mismatch"); return -1; }
  if (need == 0) return 0;
  if (!rd->ok(o.data_addr, uint64_t(need))) { set_err("dataset data out of file bounds"); return -1; }
  std::memcpy(out, &rd->f[o.data_addr], size_t(need));
  return 0;
}

int64_t hvd_h5r_attr_list(void* r, const char* obj, char* out, int64_t cap) {
  RObj o; if (!((Reader*)r)->resolve(obj, o)) return -1;
  std::vector<std::string> names; for (auto& a : o.attrs) names.push_back(a.name);
  return pack_names(names, out, cap);
}

int hvd_h5r_attr_info(void* r, const char* obj, const char* name, int* dtype, int* rank, int64_t* dims, int64_t* elsize) {
  RObj o; if (!((Reader*)r)->resolve(obj, o)) return -1;
  for (auto& a : o.attrs) if (a.name == name) {
    *dtype = a.dt; *rank = int(a.dims.size()); *elsize = a.strw;
    for (size_t i = 0; i < a.dims.size() && i < 8; i++) dims[i] = a.dims[i];
    return 0;
  }
  set_err(std::string("no such attribute: ") + name); return -1;
}

int hvd_h5r_attr_read(void* r, const char* obj, const char* name, void* out, int64_t nbytes) {
  RObj o; if (!((Reader*)r)->resolve(obj, o)) return -1;
  for (auto& a : o.attrs) if (a.name == name) {
    if (int64_t(a.data.size()) != nbytes) { set_err("attribute size mismatch"); return -1; }
    std::memcpy(out, a.data.data(), a.data.size());
    return 0;
  }
  set_err(std::string("no such attribute: ") + name); return -1;
}

}  // extern "C"
