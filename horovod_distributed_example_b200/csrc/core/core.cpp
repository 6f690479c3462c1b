// Host-side runtime core: bucket planner, TF event-file writer, Chrome-trace timeline,
// env/log helpers.  C ABI, bound from Python with ctypes; declared in include/hvd_b200.h.
//
// What each piece stands in for:
//  * hvd_plan_buckets   - Horovod's per-cycle tensor fusion (controller.cc + fusion buffer):
//                         here a one-time static plan over the arena layout.
//  * hvd_ev_*           - the TF event files Keras' TensorBoard callback writes
//                         (mnist_keras.py:105, tensorflow2_keras_mnist.py:89): TFRecord framing
//                         with masked CRC32C + hand-encoded Event/Summary protobufs.
//  * hvd_tl_*           - HOROVOD_TIMELINE (timeline.cc): Chrome-trace JSON, one "process"
//                         per tensor/bucket name, B/E phase events.

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

// --------------------------------------------------------------------- bucket planner
extern "C" int64_t hvd_plan_buckets(const int64_t* offsets, const int64_t* numels, int64_t n,
                                    int64_t slab_elems, int64_t threshold_bytes,
                                    const int64_t* cuts, int64_t n_cuts, int64_t* out /* 4*n */) {
  if (n <= 0) return 0;
  std::vector<char> is_cut(size_t(n), 0);
  for (int64_t i = 0; i < n_cuts; i++) if (cuts[i] >= 0 && cuts[i] < n) is_cut[size_t(cuts[i])] = 1;
  int64_t nb = 0, first = 0;
  while (first < n) {
    int64_t last = first, start = offsets[first];
    while (last + 1 < n) {
      if (is_cut[size_t(last)]) break;
      int64_t nxt_end = offsets[last + 1] + numels[last + 1];
      if ((nxt_end - start) * 4 > threshold_bytes) break;
      last++;
    }
    int64_t end = (last + 1 < n) ? offsets[last + 1] : slab_elems;
    out[4 * nb + 0] = start; out[4 * nb + 1] = end; out[4 * nb + 2] = first; out[4 * nb + 3] = last;
    nb++;
    first = last + 1;
  }
  return nb;
}

extern "C" uint64_t hvd_fnv1a(uint64_t h, const uint8_t* data, int64_t n) {
  for (int64_t i = 0; i < n; i++) { h ^= data[i]; h *= 0x100000001B3ull; }
  return h;
}

// --------------------------------------------------------------------- CRC32C (Castagnoli)
namespace {
uint32_t crc_table[256];
std::once_flag crc_once;
void crc_init() {
  for (uint32_t i = 0; i < 256; i++) {
    uint32_t c = i;
    for (int k = 0; k < 8; k++) c = (c & 1) ? (0x82F63B78u ^ (c >> 1)) : (c >> 1);
    crc_table[i] = c;
  }
}
uint32_t crc32c(const uint8_t* p, size_t n) {
  std::call_once(crc_once, crc_init);
  uint32_t c = 0xFFFFFFFFu;
  for (size_t i = 0; i < n; i++) c = crc_table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}
uint32_t masked_crc(const uint8_t* p, size_t n) {
  uint32_t c = crc32c(p, n);
  return ((c >> 15) | (c << 17)) + 0xa282ead8u;
}

// protobuf wire helpers
void pb_varint(std::string& s, uint64_t v) { while (v >= 0x80) { s.push_back(char(v | 0x80)); v >>= 7; } s.push_back(char(v)); }
void pb_tag(std::string& s, int field, int wt) { pb_varint(s, uint64_t(field) << 3 | uint64_t(wt)); }
void pb_double(std::string& s, int field, double v) { pb_tag(s, field, 1); s.append((const char*)&v, 8); }
void pb_float(std::string& s, int field, float v) { pb_tag(s, field, 5); s.append((const char*)&v, 4); }
void pb_bytes(std::string& s, int field, const std::string& b) { pb_tag(s, field, 2); pb_varint(s, b.size()); s += b; }
void pb_int(std::string& s, int field, int64_t v) { pb_tag(s, field, 0); pb_varint(s, uint64_t(v)); }

struct EventFile {
  FILE* f = nullptr;
  void record(const std::string& data) {
    uint64_t len = data.size();
    uint8_t hdr[12];
    std::memcpy(hdr, &len, 8);
    uint32_t c = masked_crc(hdr, 8); std::memcpy(hdr + 8, &c, 4);
    std::fwrite(hdr, 1, 12, f);
    std::fwrite(data.data(), 1, data.size(), f);
    c = masked_crc((const uint8_t*)data.data(), data.size());
    std::fwrite(&c, 1, 4, f);
  }
};
double now_s() {
  using namespace std::chrono;
  return duration<double>(system_clock::now().time_since_epoch()).count();
}
}  // namespace

extern "C" uint32_t hvd_crc32c(const uint8_t* p, int64_t n) { return crc32c(p, size_t(n)); }

extern "C" void* hvd_ev_open(const char* path) {
  FILE* f = std::fopen(path, "wb");
  if (!f) return nullptr;
  auto* e = new EventFile(); e->f = f;
  std::string ev; pb_double(ev, 1, now_s()); pb_bytes(ev, 3, "brain.Event:2");      // file_version
  e->record(ev);
  return e;
}

extern "C" int hvd_ev_scalar(void* h, const char* tag, double value, int64_t step, double wall_time) {
  auto* e = (EventFile*)h;
  if (!e || !e->f) return -1;
  std::string val; pb_bytes(val, 1, tag); pb_float(val, 2, float(value));               // Summary.Value{tag, simple_value}
  std::string summary; pb_bytes(summary, 1, val);                                       // Summary{value}
  std::string ev; pb_double(ev, 1, wall_time > 0 ? wall_time : now_s()); pb_int(ev, 2, step); pb_bytes(ev, 5, summary);
  e->record(ev);
  return 0;
}

extern "C" void hvd_ev_flush(void* h) { auto* e = (EventFile*)h; if (e && e->f) std::fflush(e->f); }
extern "C" void hvd_ev_close(void* h) { auto* e = (EventFile*)h; if (e) { if (e->f) std::fclose(e->f); delete e; } }

// --------------------------------------------------------------------- timeline
namespace {
struct Timeline {
  FILE* f = nullptr;
  std::mutex mu;
  std::map<std::string, int> pids;
  bool first = true;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  int64_t us() { return std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count(); }
  void sep() { if (!first) std::fputs(",\n", f); first = false; }
  int pid(const std::string& name) {
    auto it = pids.find(name);
    if (it != pids.end()) return it->second;
    int id = int(pids.size()) + 1; pids[name] = id;
    sep();
    std::fprintf(f, "{\"name\": \"process_name\", \"ph\": \"M\", \"pid\": %d, \"args\": {\"name\": \"%s\"}}", id, name.c_str());
    sep();
    std::fprintf(f, "{\"name\": \"process_sort_index\", \"ph\": \"M\", \"pid\": %d, \"args\": {\"sort_index\": %d}}", id, id);
    return id;
  }
};
}  // namespace

extern "C" void* hvd_tl_open(const char* path) {
  FILE* f = std::fopen(path, "w");
  if (!f) return nullptr;
  auto* t = new Timeline(); t->f = f;
  std::fputs("[\n", f);
  return t;
}

// ph: 'B' begin, 'E' end, 'i' instant, 'X' complete (dur_us used).  ts_us < 0 -> host clock now.
extern "C" int hvd_tl_event(void* h, const char* tensor, const char* activity, char ph, int64_t ts_us, int64_t dur_us) {
  auto* t = (Timeline*)h;
  if (!t || !t->f) return -1;
  std::lock_guard<std::mutex> lk(t->mu);
  int pid = t->pid(tensor);
  if (ts_us < 0) ts_us = t->us();
  t->sep();
  if (ph == 'X')
    std::fprintf(t->f, "{\"name\": \"%s\", \"ph\": \"X\", \"ts\": %lld, \"dur\": %lld, \"pid\": %d, \"tid\": 0}", activity, (long long)ts_us, (long long)dur_us, pid);
  else if (ph == 'i')
    std::fprintf(t->f, "{\"name\": \"%s\", \"ph\": \"i\", \"ts\": %lld, \"pid\": %d, \"tid\": 0, \"s\": \"g\"}", activity, (long long)ts_us, pid);
  else
    std::fprintf(t->f, "{\"name\": \"%s\", \"ph\": \"%c\", \"ts\": %lld, \"pid\": %d, \"tid\": 0}", activity, ph, (long long)ts_us, pid);
  return 0;
}

extern "C" int64_t hvd_tl_now_us(void* h) { auto* t = (Timeline*)h; return t ? t->us() : 0; }

extern "C" void hvd_tl_close(void* h) {
  auto* t = (Timeline*)h;
  if (!t) return;
  if (t->f) { std::fputs("\n]\n", t->f); std::fclose(t->f); }
  delete t;
}

extern "C" const char* hvd_core_version() { return "hvd_b200 core 0.1"; }
