// File-descriptor exchange between the ranks of one node over Unix datagram sockets
// (SCM_RIGHTS). CUDA VMM allocations and multicast objects are shared across processes as
// POSIX fds (cuMemExportToShareableHandle); an fd number means nothing in another process,
// so the kernel has to duplicate it into the peer — this is that step.
//
// No counterpart in the reference (MPI did its own bootstrap, /root/reference/src/common.jl:22).
#define FMPI_HOST_ONLY
#include "../include/fmpi.h"

#include <cerrno>
#include <cstddef>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>

#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

namespace {

int fdx_fail(const char* fmt, ...);

// Abstract-namespace address (leading NUL): no filesystem entry to clean up.
socklen_t make_addr(const char* job, int rank, sockaddr_un* a) {
  memset(a, 0, sizeof(*a));
  a->sun_family = AF_UNIX;
  int n = snprintf(a->sun_path + 1, sizeof(a->sun_path) - 1, "fmpi-fdx-%s-%d", job, rank);
  return socklen_t(offsetof(sockaddr_un, sun_path) + 1 + n);
}

double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return double(ts.tv_sec) + 1e-9 * double(ts.tv_nsec);
}

int send_fd(int sock, const sockaddr_un* to, socklen_t tolen, int fd, int32_t tag, double deadline) {
  msghdr msg{};
  iovec iov{&tag, sizeof(tag)};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_name = const_cast<sockaddr_un*>(to);
  msg.msg_namelen = tolen;
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* cm = CMSG_FIRSTHDR(&msg);
  cm->cmsg_level = SOL_SOCKET;
  cm->cmsg_type = SCM_RIGHTS;
  cm->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(cm), &fd, sizeof(int));
  for (;;) {
    if (sendmsg(sock, &msg, 0) >= 0) return 0;
    // Peer not bound yet / queue full: retry until the deadline.
    if (errno != ECONNREFUSED && errno != ENOENT && errno != EAGAIN && errno != ENOBUFS &&
        errno != EINTR)
      return fdx_fail("sendmsg: %s", strerror(errno));
    if (now_s() > deadline) return fdx_fail("sendmsg timed out: %s", strerror(errno));
    usleep(1000);
  }
}

int recv_fd(int sock, int* fd_out, int32_t* tag_out, double deadline) {
  for (;;) {
    pollfd p{sock, POLLIN, 0};
    int ms = int((deadline - now_s()) * 1000.0);
    if (ms <= 0) return fdx_fail("recvmsg timed out waiting for a peer's fd");
    int pr = poll(&p, 1, ms > 1000 ? 1000 : ms);
    if (pr < 0 && errno != EINTR) return fdx_fail("poll: %s", strerror(errno));
    if (pr <= 0) continue;
    msghdr msg{};
    int32_t tag = -1;
    iovec iov{&tag, sizeof(tag)};
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    ssize_t n = recvmsg(sock, &msg, 0);
    if (n < 0) {
      if (errno == EINTR || errno == EAGAIN) continue;
      return fdx_fail("recvmsg: %s", strerror(errno));
    }
    cmsghdr* cm = CMSG_FIRSTHDR(&msg);
    if (!cm || cm->cmsg_level != SOL_SOCKET || cm->cmsg_type != SCM_RIGHTS)
      return fdx_fail("recvmsg: datagram carried no fd");
    memcpy(fd_out, CMSG_DATA(cm), sizeof(int));
    *tag_out = tag;
    return 0;
  }
}

thread_local std::string g_fdx_err;
int fdx_fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_fdx_err = buf;
  return 1;
}

}  // namespace

extern "C" {

// Shares the error slot with the host backend from the caller's point of view: Python asks
// fmpi_fdx_last_error() after a failing fmpi_fdx_* call.
const char* fmpi_fdx_last_error(void) { return g_fdx_err.c_str(); }

int fmpi_fdx_open(const char* job, int rank, int* sock_out) {
  int s = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
  if (s < 0) return fdx_fail("socket: %s", strerror(errno));
  sockaddr_un a;
  socklen_t len = make_addr(job, rank, &a);
  if (bind(s, reinterpret_cast<sockaddr*>(&a), len) != 0) {
    int e = errno;
    close(s);
    return fdx_fail("bind(fmpi-fdx-%s-%d): %s", job, rank, strerror(e));
  }
  *sock_out = s;
  return 0;
}

int fmpi_fdx_exchange(int sock, const char* job, int rank, int world, int fd, int* fds_out,
                      double timeout_s) {
  const double deadline = now_s() + (timeout_s > 0 ? timeout_s : 60.0);
  for (int r = 0; r < world; ++r) fds_out[r] = -1;
  fds_out[rank] = fd;
  for (int r = 0; r < world; ++r) {
    if (r == rank) continue;
    sockaddr_un a;
    socklen_t len = make_addr(job, r, &a);
    if (send_fd(sock, &a, len, fd, rank, deadline)) return 1;
  }
  for (int i = 0; i < world - 1; ++i) {
    int got = -1;
    int32_t from = -1;
    if (recv_fd(sock, &got, &from, deadline)) return 1;
    if (from < 0 || from >= world || from == rank || fds_out[from] != -1) {
      if (got >= 0) close(got);
      return fdx_fail("unexpected sender tag %d", from);
    }
    fds_out[from] = got;
  }
  return 0;
}

int fmpi_fdx_bcast(int sock, const char* job, int rank, int world, int root, int* fd_inout,
                   double timeout_s) {
  const double deadline = now_s() + (timeout_s > 0 ? timeout_s : 60.0);
  if (rank == root) {
    for (int r = 0; r < world; ++r) {
      if (r == rank) continue;
      sockaddr_un a;
      socklen_t len = make_addr(job, r, &a);
      if (send_fd(sock, &a, len, *fd_inout, 1000 + root, deadline)) return 1;
    }
    return 0;
  }
  int32_t tag = -1;
  int got = -1;
  if (recv_fd(sock, &got, &tag, deadline)) return 1;
  if (tag != 1000 + root) {
    if (got >= 0) close(got);
    return fdx_fail("bcast: unexpected tag %d", tag);
  }
  *fd_inout = got;
  return 0;
}

int fmpi_fdx_close(int sock, const char* job, int rank) {
  (void)job;
  (void)rank;
  if (sock >= 0) close(sock);
  return 0;
}

}  // extern "C"
