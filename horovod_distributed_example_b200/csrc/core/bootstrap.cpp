// Bootstrap channel of the peer-memory runtime: a fixed-layout message and POSIX file-descriptor passing
// (SCM_RIGHTS) between the ranks of one host over unix datagram sockets.
//
// Capability parity: what Horovod 0.18.1 does at hvd.init() with MPI_Bcast of the ncclUniqueId and its
// Request/Response flatbuffers (common/message.cc, wire/message.fbs; reference call site mnist_keras.py:30).
// Here nothing is negotiated per step, so the only control traffic that ever exists is this start-up exchange:
// every rank sends every peer the shareable handle (an fd) of each symmetric allocation, tagged with the
// allocation's ordinal, plus the multicast handle from rank 0.  The payload is a C struct, not a pickle.
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <sys/un.h>
#include <unistd.h>

#include <string>

#include "hvd_b200.h"

namespace {
constexpr uint32_t kMagic = 0x48564442;        // "HVDB"
constexpr uint16_t kVersion = 1;
thread_local std::string g_err;

struct Endpoint {
  int sock = -1;
  int rank = 0, size = 1;
  std::string dir, path;
};

std::string peer_path(const std::string& dir, int r) { return dir + "/r" + std::to_string(r) + ".sock"; }

int fail(const std::string& what) {
  g_err = what + ": " + strerror(errno);
  return -1;
}
}  // namespace

extern "C" const char* hvd_boot_last_error() { return g_err.c_str(); }

extern "C" void* hvd_boot_open(const char* dir, int rank, int size, double timeout_s) {
  if (!dir || rank < 0 || size < 1 || rank >= size) { g_err = "bad arguments"; return nullptr; }
  auto* e = new Endpoint();
  e->rank = rank; e->size = size; e->dir = dir; e->path = peer_path(e->dir, rank);
  mkdir(dir, 0700);
  if (e->path.size() >= sizeof(sockaddr_un::sun_path)) { g_err = "socket path too long: " + e->path; delete e; return nullptr; }
  unlink(e->path.c_str());
  e->sock = socket(AF_UNIX, SOCK_DGRAM, 0);
  if (e->sock < 0) { fail("socket"); delete e; return nullptr; }
  sockaddr_un a{};
  a.sun_family = AF_UNIX;
  strncpy(a.sun_path, e->path.c_str(), sizeof(a.sun_path) - 1);
  if (bind(e->sock, (sockaddr*)&a, sizeof(a)) != 0) { fail("bind " + e->path); close(e->sock); delete e; return nullptr; }
  if (timeout_s > 0) {
    timeval tv{(time_t)timeout_s, (suseconds_t)((timeout_s - (time_t)timeout_s) * 1e6)};
    setsockopt(e->sock, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  }
  return e;
}

// Send one message to rank `dst`; fd >= 0 rides along as SCM_RIGHTS ancillary data.
extern "C" int hvd_boot_send(void* h, int dst, int kind, int tag, uint64_t payload, int fd) {
  auto* e = (Endpoint*)h;
  if (!e || dst < 0 || dst >= e->size) { g_err = "bad destination"; return -1; }
  HvdBootMsg m{kMagic, kVersion, (uint16_t)kind, e->rank, tag, payload};
  iovec iov{&m, sizeof(m)};
  sockaddr_un a{};
  a.sun_family = AF_UNIX;
  const std::string p = peer_path(e->dir, dst);
  strncpy(a.sun_path, p.c_str(), sizeof(a.sun_path) - 1);
  msghdr mh{};
  mh.msg_name = &a; mh.msg_namelen = sizeof(a);
  mh.msg_iov = &iov; mh.msg_iovlen = 1;
  alignas(cmsghdr) char ctl[CMSG_SPACE(sizeof(int))];
  if (fd >= 0) {
    memset(ctl, 0, sizeof(ctl));
    mh.msg_control = ctl; mh.msg_controllen = sizeof(ctl);
    cmsghdr* c = CMSG_FIRSTHDR(&mh);
    c->cmsg_level = SOL_SOCKET; c->cmsg_type = SCM_RIGHTS; c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
  }
  // the peer may not have bound its socket yet: retry for a while on ENOENT / ECONNREFUSED
  for (int attempt = 0; attempt < 600; attempt++) {
    if (sendmsg(e->sock, &mh, 0) == (ssize_t)sizeof(m)) return 0;
    if (errno != ENOENT && errno != ECONNREFUSED && errno != EAGAIN) break;
    usleep(100000);
  }
  return fail("sendmsg to rank " + std::to_string(dst));
}

// Receive one message (blocking up to the endpoint's timeout).  *fd = the passed descriptor or -1.
extern "C" int hvd_boot_recv(void* h, HvdBootMsg* out, int* fd) {
  auto* e = (Endpoint*)h;
  if (!e || !out || !fd) { g_err = "bad arguments"; return -1; }
  *fd = -1;
  iovec iov{out, sizeof(*out)};
  msghdr mh{};
  mh.msg_iov = &iov; mh.msg_iovlen = 1;
  alignas(cmsghdr) char ctl[CMSG_SPACE(sizeof(int))];
  mh.msg_control = ctl; mh.msg_controllen = sizeof(ctl);
  const ssize_t n = recvmsg(e->sock, &mh, 0);
  if (n < 0) return fail("recvmsg");
  if (n != (ssize_t)sizeof(*out) || out->magic != kMagic || out->version != kVersion) {
    g_err = "malformed bootstrap message (" + std::to_string((long long)n) + " bytes)";
    return -2;
  }
  for (cmsghdr* c = CMSG_FIRSTHDR(&mh); c; c = CMSG_NXTHDR(&mh, c))
    if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) memcpy(fd, CMSG_DATA(c), sizeof(int));
  return 0;
}

extern "C" void hvd_boot_close(void* h) {
  auto* e = (Endpoint*)h;
  if (!e) return;
  if (e->sock >= 0) close(e->sock);
  unlink(e->path.c_str());
  delete e;
}
