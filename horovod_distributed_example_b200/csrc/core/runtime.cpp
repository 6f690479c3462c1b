// Process-level runtime state of the native core: the C API Horovod exposes from operations.cc
// (horovod_init / horovod_rank / horovod_size / horovod_local_rank / horovod_shutdown; reference call sites
// mnist_keras.py:30,35,42 and tensorflow2_keras_mnist.py:25,32,35) and the registry of peer-mapped symmetric
// buffers that the collective kernels address.
//
// There is no background thread to start here (the bucket plan is static), so init is:
// read the launcher's environment contract, remember rank / size / local rank, hand them to C or C++ callers of
// the C ABI the same way Python gets them.  Contracts, in priority order: HVD_*, Open MPI (the reference's
// `mpirun` line, README.md:57), PMI, torchrun.  No launcher at all means a world of one (the reference's YAML
// actually runs the script outside mpirun).
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "hvd_b200.h"

namespace {
struct State {
  bool initialized = false;
  int rank = 0, size = 1, local_rank = 0, local_size = 1;
  std::string contract = "singleton";
};
State g_state;
std::mutex g_mu;

struct Contract { const char *rank, *size, *local_rank, *local_size, *name; };
const Contract kContracts[] = {
    {"HVD_RANK", "HVD_SIZE", "HVD_LOCAL_RANK", "HVD_LOCAL_SIZE", "hvdrun"},
    {"OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_SIZE", "openmpi"},
    {"PMI_RANK", "PMI_SIZE", "MPI_LOCALRANKID", "MPI_LOCALNRANKS", "pmi"},
    {"RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE", "torchrun"},
};

bool env_int(const char* name, int* out) {
  const char* v = std::getenv(name);
  if (!v || !*v) return false;
  char* end = nullptr;
  long x = std::strtol(v, &end, 10);
  if (end == v) return false;
  *out = (int)x;
  return true;
}

// One registered symmetric buffer: where every rank's copy is mapped in THIS process.
struct SymmEntry {
  int id;
  uint64_t bytes;
  uint64_t mc_ptr;
  int local_index;
  std::vector<uint64_t> ptrs;
};
std::vector<SymmEntry> g_registry;
int g_next_id = 1;
}  // namespace

// Discover (rank, size, local_rank, local_size) from the environment.  Returns 0, or -1 when a contract is
// present but inconsistent (rank outside [0, size)).  out[4]; *contract (may be NULL) names the launcher.
extern "C" int hvd_core_discover_env(int* out4, const char** contract) {
  for (const Contract& c : kContracts) {
    int r, s;
    if (!env_int(c.rank, &r) || !env_int(c.size, &s)) continue;
    if (s < 1 || r < 0 || r >= s) return -1;
    int lr = r, ls = s;
    env_int(c.local_rank, &lr);
    env_int(c.local_size, &ls);
    out4[0] = r; out4[1] = s; out4[2] = lr; out4[3] = ls;
    if (contract) *contract = c.name;
    return 0;
  }
  out4[0] = 0; out4[1] = 1; out4[2] = 0; out4[3] = 1;
  if (contract) *contract = "singleton";
  return 0;
}

// Idempotent.  Returns 0 on success, -1 on an inconsistent launcher environment.
extern "C" int hvd_core_init(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_state.initialized) return 0;
  int v[4];
  const char* name = nullptr;
  if (hvd_core_discover_env(v, &name) != 0) return -1;
  g_state.rank = v[0]; g_state.size = v[1]; g_state.local_rank = v[2]; g_state.local_size = v[3];
  g_state.contract = name ? name : "singleton";
  g_state.initialized = true;
  hvd_log_set_rank(g_state.rank);
  return 0;
}

extern "C" int hvd_core_is_initialized(void) { return g_state.initialized ? 1 : 0; }
// -1 before hvd_core_init (Horovod's C API returns -1 too; the Python layer turns that into ValueError).
extern "C" int hvd_core_rank(void) { return g_state.initialized ? g_state.rank : -1; }
extern "C" int hvd_core_size(void) { return g_state.initialized ? g_state.size : -1; }
extern "C" int hvd_core_local_rank(void) { return g_state.initialized ? g_state.local_rank : -1; }
extern "C" int hvd_core_local_size(void) { return g_state.initialized ? g_state.local_size : -1; }
extern "C" const char* hvd_core_launcher(void) { return g_state.contract.c_str(); }

extern "C" void hvd_core_shutdown(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_state = State();
  g_registry.clear();
}

// ------------------------------------------------------------------ registry of symmetric buffers
// ptrs[r] = address of rank r's copy in this process (n = world size, or the emulated rank count).
extern "C" int hvd_registry_add(const uint64_t* ptrs, int n, uint64_t bytes, uint64_t mc_ptr, int local_index) {
  if (!ptrs || n < 1 || local_index < 0 || local_index >= n) return -1;
  std::lock_guard<std::mutex> lk(g_mu);
  SymmEntry e{g_next_id++, bytes, mc_ptr, local_index, std::vector<uint64_t>(ptrs, ptrs + n)};
  g_registry.push_back(e);
  return e.id;
}

// Which registered buffer does the LOCAL address `p` fall into?  Returns its id (and the byte offset), 0 if none.
extern "C" int hvd_registry_find(uint64_t p, uint64_t* byte_offset) {
  std::lock_guard<std::mutex> lk(g_mu);
  // newest first: an address range can be registered again after its previous owner was released
  for (auto it = g_registry.rbegin(); it != g_registry.rend(); ++it) {
    const SymmEntry& e = *it;
    const uint64_t base = e.ptrs[(size_t)e.local_index];
    if (p >= base && p < base + e.bytes) {
      if (byte_offset) *byte_offset = p - base;
      return e.id;
    }
  }
  return 0;
}

extern "C" int hvd_registry_remove(int id) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (size_t i = 0; i < g_registry.size(); i++)
    if (g_registry[i].id == id) { g_registry.erase(g_registry.begin() + (long)i); return 0; }
  return -1;
}

// Address of rank r's copy of buffer `id` (0 if unknown); *mc_ptr / *bytes may be NULL.
extern "C" uint64_t hvd_registry_peer_ptr(int id, int r, uint64_t* mc_ptr, uint64_t* bytes) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (const SymmEntry& e : g_registry)
    if (e.id == id) {
      if (mc_ptr) *mc_ptr = e.mc_ptr;
      if (bytes) *bytes = e.bytes;
      return (r >= 0 && r < (int)e.ptrs.size()) ? e.ptrs[(size_t)r] : 0;
    }
  return 0;
}

extern "C" int hvd_registry_count(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  return (int)g_registry.size();
}

extern "C" void hvd_registry_clear(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_registry.clear();
}
