// fdxchg.cpp -- moves one file descriptor per rank between the processes of a job.
//
// A rank context (one process per GPU: hvd.init() imagenet-resnet50-hvd.py:16, one MWMS
// worker imagenet-resnet50-multiworkers.py:24-25) shares its peer-mappable block with the
// other ranks as a POSIX fd.  fds only travel over SCM_RIGHTS, so each rank binds an
// abstract unix datagram socket named after (session, rank), sends its fd to every peer
// and collects one from each.  Host-only code: it works for any fd and is tested on CPU.
#include <cerrno>
#include <chrono>
#include <cstring>
#include <thread>

#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include "internal.h"

using namespace b2dp;

namespace {

socklen_t make_addr(sockaddr_un* a, const char* session, int rank) {
  memset(a, 0, sizeof *a);
  a->sun_family = AF_UNIX;
  // abstract namespace: sun_path[0] == 0, no filesystem entry to clean up
  int n = snprintf(a->sun_path + 1, sizeof(a->sun_path) - 1, "b2dp-%s-%d", session, rank);
  return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

struct Msg {
  int32_t magic;
  int32_t rank;
};
constexpr int32_t kMagic = 0x62326470;  // "b2dp"

int send_fd(int sock, const sockaddr_un* to, socklen_t tolen, int rank, int fd) {
  Msg m{kMagic, rank};
  iovec iov{&m, sizeof m};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof ctrl);
  msghdr h{};
  h.msg_name = (void*)to;
  h.msg_namelen = tolen;
  h.msg_iov = &iov;
  h.msg_iovlen = 1;
  h.msg_control = ctrl;
  h.msg_controllen = sizeof ctrl;
  cmsghdr* c = CMSG_FIRSTHDR(&h);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  return (int)sendmsg(sock, &h, 0);
}

int recv_fd(int sock, int* rank, int* fd) {
  Msg m{};
  iovec iov{&m, sizeof m};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  msghdr h{};
  h.msg_iov = &iov;
  h.msg_iovlen = 1;
  h.msg_control = ctrl;
  h.msg_controllen = sizeof ctrl;
  ssize_t n = recvmsg(sock, &h, 0);
  if (n < 0) return -1;
  if (n != (ssize_t)sizeof m || m.magic != kMagic) return -2;
  cmsghdr* c = CMSG_FIRSTHDR(&h);
  if (!c || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) return -2;
  memcpy(fd, CMSG_DATA(c), sizeof(int));
  *rank = m.rank;
  return 0;
}

}  // namespace

extern "C" int b2dp_fd_exchange(int32_t rank, int32_t nranks, const char* session, int my_fd,
                                int* peer_fds, int32_t timeout_ms) {
  if (!session || !peer_fds || nranks < 1 || rank < 0 || rank >= nranks || my_fd < 0)
    return fail(B2DP_ERR_INVALID, "fd_exchange: bad arguments");
  if (strlen(session) > 64) return fail(B2DP_ERR_INVALID, "fd_exchange: session name too long");
  for (int r = 0; r < nranks; ++r) peer_fds[r] = -1;
  peer_fds[rank] = dup(my_fd);
  if (peer_fds[rank] < 0) return fail(B2DP_ERR_SYS, "fd_exchange: dup: %s", strerror(errno));
  if (nranks == 1) return B2DP_OK;

  int sock = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
  if (sock < 0) return fail(B2DP_ERR_SYS, "fd_exchange: socket: %s", strerror(errno));
  sockaddr_un me;
  socklen_t melen = make_addr(&me, session, rank);
  if (bind(sock, (sockaddr*)&me, melen) != 0) {
    int e = errno;
    close(sock);
    return fail(B2DP_ERR_SYS, "fd_exchange: bind(%s/%d): %s", session, rank, strerror(e));
  }
  timeval tv{0, 50 * 1000};  // 50 ms receive slices so the deadline is honoured
  setsockopt(sock, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);

  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  int rc = B2DP_OK;
  // send to every peer; a peer that has not bound yet answers ECONNREFUSED -> retry
  for (int r = 0; r < nranks && rc == B2DP_OK; ++r) {
    if (r == rank) continue;
    sockaddr_un to;
    socklen_t tolen = make_addr(&to, session, r);
    for (;;) {
      if (send_fd(sock, &to, tolen, rank, my_fd) >= 0) break;
      if (errno != ECONNREFUSED && errno != ENOENT && errno != EAGAIN && errno != ENOBUFS) {
        rc = fail(B2DP_ERR_SYS, "fd_exchange: sendmsg to rank %d: %s", r, strerror(errno));
        break;
      }
      if (std::chrono::steady_clock::now() > deadline) {
        rc = fail(B2DP_ERR_TIMEOUT, "fd_exchange: rank %d never bound its socket", r);
        break;
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(5));
    }
  }
  int got = 0;
  while (rc == B2DP_OK && got < nranks - 1) {
    int from = -1, fd = -1;
    int r = recv_fd(sock, &from, &fd);
    if (r == 0) {
      if (from < 0 || from >= nranks || from == rank || peer_fds[from] >= 0) {
        close(fd);
        rc = fail(B2DP_ERR_SYS, "fd_exchange: unexpected message from rank %d", from);
      } else {
        peer_fds[from] = fd;
        ++got;
      }
    } else if (r == -2) {
      rc = fail(B2DP_ERR_SYS, "fd_exchange: malformed message");
    } else if (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR) {
      rc = fail(B2DP_ERR_SYS, "fd_exchange: recvmsg: %s", strerror(errno));
    } else if (std::chrono::steady_clock::now() > deadline) {
      rc = fail(B2DP_ERR_TIMEOUT, "fd_exchange: got %d of %d peer fds", got, nranks - 1);
    }
  }
  close(sock);
  if (rc != B2DP_OK)
    for (int r = 0; r < nranks; ++r)
      if (peer_fds[r] >= 0) {
        close(peer_fds[r]);
        peer_fds[r] = -1;
      }
  return rc;
}
