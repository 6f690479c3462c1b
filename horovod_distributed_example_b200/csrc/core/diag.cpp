// Diagnostics for the native runtime: levelled logging and NVTX ranges.
//
// Role of horovod/common/logging.cc (HOROVOD_LOG_LEVEL) and of the NVTX half of SURVEY.md section 5.1.
// The level is read once from HVD_LOG_LEVEL (alias HOROVOD_LOG_LEVEL): TRACE, DEBUG, INFO, WARNING,
// ERROR, FATAL; lines go to stderr as "[hvd <rank>] LEVEL message" so interleaved ranks stay readable.
//
// NVTX uses the header-only NVTX3 that ships with the CUDA toolkit: the calls are no-ops costing one
// branch unless a profiler injected itself into the process, so ranges can stay in release builds.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <strings.h>

#include <nvtx3/nvToolsExt.h>

#include "hvd_b200.h"

namespace {

const char* kNames[] = {"TRACE", "DEBUG", "INFO", "WARNING", "ERROR", "FATAL"};

int parse_level(const char* s) {
  if (!s || !*s) return HVD_LOG_WARNING;
  for (int i = 0; i < 6; ++i)
    if (strcasecmp(s, kNames[i]) == 0) return i;
  if (strcasecmp(s, "WARN") == 0) return HVD_LOG_WARNING;
  char* end = nullptr;
  long v = std::strtol(s, &end, 10);
  if (end != s && v >= 0 && v <= 5) return int(v);
  return HVD_LOG_WARNING;
}

std::atomic<int> g_level{-1};
std::atomic<int> g_rank{-1};
std::mutex g_mu;

int level() {
  int l = g_level.load(std::memory_order_relaxed);
  if (l >= 0) return l;
  const char* s = std::getenv("HVD_LOG_LEVEL");
  if (!s) s = std::getenv("HOROVOD_LOG_LEVEL");
  l = parse_level(s);
  g_level.store(l, std::memory_order_relaxed);
  return l;
}

}  // namespace

extern "C" void hvd_log_set_level(int lvl) { g_level.store(lvl < 0 ? 0 : (lvl > 5 ? 5 : lvl)); }
extern "C" int hvd_log_level(void) { return level(); }
extern "C" void hvd_log_set_rank(int rank) { g_rank.store(rank); }
extern "C" int hvd_log_parse_level(const char* s) { return parse_level(s); }

extern "C" void hvd_log(int lvl, const char* msg) {
  if (lvl < level() || !msg) return;
  if (lvl > 5) lvl = 5;
  std::lock_guard<std::mutex> lk(g_mu);
  int r = g_rank.load();
  if (r >= 0) std::fprintf(stderr, "[hvd %d] %s %s\n", r, kNames[lvl], msg);
  else        std::fprintf(stderr, "[hvd] %s %s\n", kNames[lvl], msg);
  if (lvl >= HVD_LOG_ERROR) std::fflush(stderr);
}

// printf-style helper for the other translation units of the core library.
void hvd_logf(int lvl, const char* fmt, ...) {
  if (lvl < level()) return;
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  std::vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  hvd_log(lvl, buf);
}

// ------------------------------------------------------------------ NVTX
extern "C" int hvd_nvtx_push(const char* name) { return nvtxRangePushA(name ? name : "hvd"); }
extern "C" int hvd_nvtx_pop(void) { return nvtxRangePop(); }
extern "C" void hvd_nvtx_mark(const char* name) { nvtxMarkA(name ? name : "hvd"); }
