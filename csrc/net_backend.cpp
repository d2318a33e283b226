// Rail mesh: host collectives between the nodes of a job, over TCP.
//
// The reference reaches other nodes through whatever fabric its MPI was built for (SURVEY §2.2,
// "Multi-node: implicitly, via whatever mpiexec launches"). This image has no MPI, so local
// rank l of every node keeps a socket to local rank l of every other node (one such mesh per
// rail, csrc/host_backend.cpp hands each rail its slice of a chunk) and the few
// collectives the host world needs are written on top of one primitive, `exchange`, which
// progresses any number of sends and receives at once with poll(2): nobody can dead-lock on a
// full socket buffer and every wait is bounded by the job's timeout.
//
//   bring-up   node 0 listens on master_addr:port. Every other leader opens its own listening
//              socket on a free port, dials node 0 (retrying while it is not up yet) and tells
//              it that port; node 0 answers with the whole address table. The dial-in sockets
//              stay as the 0<->q links; q then dials every p with 0 < p < q.
//   allreduce  slice k of the message belongs to node k: everybody sends slice k to k, k folds
//              the N copies in node order (the same routine the shared-memory level uses) and
//              sends the result back. One owner + fixed order => the same bits on every node,
//              the replica-consistency rule of SURVEY §5.1 across nodes too.
//   reduce     the same scatter-reduce, then the slices are collected by the root only.
//   bcast      the root streams to everybody (node counts are small; the links run in parallel).
//
// Device tensors do not come through here: across nodes they take the NCCL path
// (parallel/device_world.py).
#define FMPI_HOST_ONLY
#include "../include/fmpi.h"
#include "host_internal.h"

#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <time.h>
#include <unistd.h>

using fmpi_host_detail::dtype_size;
using fmpi_host_detail::fail;
using fmpi_host_detail::now_s;
using fmpi_host_detail::reduce_slots;

namespace {

struct Addr {
  uint32_t ip = 0;    // network byte order
  uint16_t port = 0;  // host byte order
};

// One direction of traffic with one peer.
struct Xfer {
  int peer;
  const uint8_t* src;
  size_t src_n;
  uint8_t* dst;
  size_t dst_n;
};

void set_nonblocking(int fd, bool on) {
  const int fl = fcntl(fd, F_GETFL, 0);
  fcntl(fd, F_SETFL, on ? (fl | O_NONBLOCK) : (fl & ~O_NONBLOCK));
}

void tune(int fd) {
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  int buf = 4 << 20;
  setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &buf, sizeof(buf));
  setsockopt(fd, SOL_SOCKET, SO_RCVBUF, &buf, sizeof(buf));
}

// Blocking helpers for the bring-up handshake (sockets still in blocking mode, bounded by
// SO_RCVTIMEO / SO_SNDTIMEO).
bool send_all(int fd, const void* p, size_t n) {
  const uint8_t* b = static_cast<const uint8_t*>(p);
  while (n) {
    const ssize_t k = send(fd, b, n, MSG_NOSIGNAL);
    if (k < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    b += k;
    n -= size_t(k);
  }
  return true;
}

bool recv_all(int fd, void* p, size_t n) {
  uint8_t* b = static_cast<uint8_t*>(p);
  while (n) {
    const ssize_t k = recv(fd, b, n, 0);
    if (k == 0) return false;
    if (k < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    b += k;
    n -= size_t(k);
  }
  return true;
}

void set_io_timeout(int fd, double s) {
  timeval tv;
  tv.tv_sec = long(s);
  tv.tv_usec = long((s - double(tv.tv_sec)) * 1e6);
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof(tv));
}

int listen_on(uint16_t port, uint16_t* bound) {
  const int fd = socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) return -1;
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in a{};
  a.sin_family = AF_INET;
  a.sin_addr.s_addr = htonl(INADDR_ANY);
  a.sin_port = htons(port);
  if (bind(fd, reinterpret_cast<sockaddr*>(&a), sizeof(a)) != 0 || listen(fd, 64) != 0) {
    const int e = errno;
    close(fd);
    errno = e;
    return -1;
  }
  socklen_t len = sizeof(a);
  getsockname(fd, reinterpret_cast<sockaddr*>(&a), &len);
  *bound = ntohs(a.sin_port);
  return fd;
}

bool resolve(const char* host, uint32_t* ip) {
  in_addr a;
  if (inet_pton(AF_INET, host, &a) == 1) {
    *ip = a.s_addr;
    return true;
  }
  addrinfo hints{};
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  addrinfo* res = nullptr;
  if (getaddrinfo(host, nullptr, &hints, &res) != 0 || !res) return false;
  *ip = reinterpret_cast<sockaddr_in*>(res->ai_addr)->sin_addr.s_addr;
  freeaddrinfo(res);
  return true;
}

// Dial until the peer is listening or the deadline passes.
int dial(uint32_t ip, uint16_t port, double deadline) {
  for (;;) {
    const int fd = socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) return -1;
    sockaddr_in a{};
    a.sin_family = AF_INET;
    a.sin_addr.s_addr = ip;
    a.sin_port = htons(port);
    if (connect(fd, reinterpret_cast<sockaddr*>(&a), sizeof(a)) == 0) return fd;
    close(fd);
    if (now_s() > deadline) return -1;
    timespec ts{0, 20 * 1000 * 1000};
    nanosleep(&ts, nullptr);
  }
}

int accept_until(int lfd, double deadline, uint32_t* peer_ip) {
  for (;;) {
    pollfd p{lfd, POLLIN, 0};
    const double left = deadline - now_s();
    if (left <= 0) return -1;
    const int rc = poll(&p, 1, int(left * 1000.0 > 1000.0 ? 1000.0 : left * 1000.0 + 1));
    if (rc < 0 && errno != EINTR) return -1;
    if (rc > 0) {
      sockaddr_in a{};
      socklen_t len = sizeof(a);
      const int fd = accept(lfd, reinterpret_cast<sockaddr*>(&a), &len);
      if (fd >= 0) {
        if (peer_ip) *peer_ip = a.sin_addr.s_addr;
        return fd;
      }
    }
  }
}

}  // namespace

struct fmpi_net {
  int node = 0, nnodes = 1;
  double timeout_s = 300.0;
  std::vector<int> sock;          // sock[q], -1 for q == node
  std::vector<uint8_t> scratch;   // slice staging for the reductions

  ~fmpi_net() {
    for (int fd : sock)
      if (fd >= 0) close(fd);
  }

  // Run all transfers to completion. Sends and receives with one peer share its socket; the
  // byte streams of successive collectives cannot mix because every rank issues the same
  // sequence of collectives and each one consumes exactly what its counterpart produced.
  int exchange(std::vector<Xfer>& xs) {
    std::vector<size_t> sent(xs.size(), 0), got(xs.size(), 0);
    std::vector<pollfd> pf(xs.size());
    const double t0 = now_s();
    for (;;) {
      size_t open = 0;
      for (size_t i = 0; i < xs.size(); ++i) {
        pf[i].fd = sock[xs[i].peer];
        pf[i].events = 0;
        pf[i].revents = 0;
        if (sent[i] < xs[i].src_n) pf[i].events |= POLLOUT;
        if (got[i] < xs[i].dst_n) pf[i].events |= POLLIN;
        if (pf[i].events) ++open; else pf[i].fd = -1;
      }
      if (open == 0) return 0;
      const int rc = poll(pf.data(), nfds_t(pf.size()), 1000);
      if (rc < 0 && errno != EINTR) return fail("node %d: poll: %s", node, strerror(errno));
      if (now_s() - t0 > timeout_s)
        return fail("node %d: inter-node transfer timed out after %.0f s (a node is dead or out of step)",
                    node, timeout_s);
      if (rc <= 0) continue;
      for (size_t i = 0; i < xs.size(); ++i) {
        if (pf[i].fd < 0) continue;
        if (pf[i].revents & (POLLIN | POLLHUP | POLLERR)) {
          while (got[i] < xs[i].dst_n) {
            const ssize_t k = recv(pf[i].fd, xs[i].dst + got[i], xs[i].dst_n - got[i], 0);
            if (k > 0) { got[i] += size_t(k); continue; }
            if (k == 0) return fail("node %d: node %d closed the connection", node, xs[i].peer);
            if (errno == EAGAIN || errno == EWOULDBLOCK) break;
            if (errno == EINTR) continue;
            return fail("node %d: recv from node %d: %s", node, xs[i].peer, strerror(errno));
          }
        }
        if (pf[i].revents & (POLLOUT | POLLHUP | POLLERR)) {  // on error, send() names it
          while (sent[i] < xs[i].src_n) {
            const ssize_t k = send(pf[i].fd, xs[i].src + sent[i], xs[i].src_n - sent[i], MSG_NOSIGNAL);
            if (k > 0) { sent[i] += size_t(k); continue; }
            if (k < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) break;
            if (k < 0 && errno == EINTR) continue;
            return fail("node %d: send to node %d: %s", node, xs[i].peer, strerror(errno));
          }
        }
      }
    }
  }

  // Element range of the slice node k owns in a message of `count` elements.
  void slice(size_t count, int k, size_t* lo, size_t* hi) const {
    const size_t per = (count + size_t(nnodes) - 1) / size_t(nnodes);
    *lo = per * size_t(k) < count ? per * size_t(k) : count;
    *hi = *lo + per < count ? *lo + per : count;
  }

  // After this, buf[my slice] holds the reduction over all nodes; other slices are untouched.
  int scatter_reduce(uint8_t* buf, size_t count, int dtype, int op) {
    const size_t es = dtype_size(dtype);
    size_t lo, hi;
    slice(count, node, &lo, &hi);
    const size_t mine = (hi - lo) * es;
    const size_t stride = (mine + 63) & ~size_t(63);
    if (scratch.size() < stride * size_t(nnodes)) scratch.resize(stride * size_t(nnodes));
    std::vector<Xfer> xs;
    for (int q = 0; q < nnodes; ++q) {
      if (q == node) continue;
      size_t qlo, qhi;
      slice(count, q, &qlo, &qhi);
      xs.push_back({q, buf + qlo * es, (qhi - qlo) * es, scratch.data() + stride * size_t(q), mine});
    }
    if (exchange(xs)) return 1;
    if (mine) {
      memcpy(scratch.data() + stride * size_t(node), buf + lo * es, mine);
      reduce_slots(dtype, scratch.data(), stride, nnodes, buf + lo * es, 0, hi - lo, op);
    }
    return 0;
  }
};

extern "C" {

int fmpi_net_init(const char* master_addr, int port, int node, int nnodes, double timeout_s,
                  fmpi_net** out) {
  if (!out || nnodes < 1 || node < 0 || node >= nnodes || port <= 0 || port > 65535)
    return fail("fmpi_net_init: bad arguments (node %d of %d, port %d)", node, nnodes, port);
  auto* n = new fmpi_net;
  n->node = node;
  n->nnodes = nnodes;
  n->timeout_s = timeout_s > 0 ? timeout_s : 300.0;
  n->sock.assign(size_t(nnodes), -1);
  if (nnodes == 1) { *out = n; return 0; }
  const double deadline = now_s() + n->timeout_s;
  std::vector<Addr> table;
  table.resize(size_t(nnodes));
  int lfd = -1;
  auto bail = [&](int rc) { if (lfd >= 0) close(lfd); delete n; return rc; };

  if (node == 0) {
    uint16_t bound = 0;
    lfd = listen_on(uint16_t(port), &bound);
    if (lfd < 0) return bail(fail("node 0: cannot listen on port %d: %s", port, strerror(errno)));
    for (int k = 1; k < nnodes; ++k) {
      uint32_t ip = 0;
      const int fd = accept_until(lfd, deadline, &ip);
      if (fd < 0) return bail(fail("node 0: only %d of %d nodes dialled in before the timeout", k, nnodes));
      set_io_timeout(fd, n->timeout_s);
      uint32_t hello[3];  // magic, node, listening port
      if (!recv_all(fd, hello, sizeof(hello)) || hello[0] != 0x464d4e54u || hello[1] == 0 ||
          hello[1] >= uint32_t(nnodes) || n->sock[hello[1]] >= 0) {
        close(fd);
        return bail(fail("node 0: bad hello on the rendezvous port (is port %d shared with another job?)", port));
      }
      n->sock[hello[1]] = fd;
      table[hello[1]].ip = ip;
      table[hello[1]].port = uint16_t(hello[2]);
    }
    for (int q = 1; q < nnodes; ++q)
      if (!send_all(n->sock[size_t(q)], table.data(), table.size() * sizeof(Addr)))
        return bail(fail("node 0: could not send the address table to node %d", q));
  } else {
    uint32_t ip0 = 0;
    if (!master_addr || !resolve(master_addr, &ip0))
      return bail(fail("node %d: cannot resolve master address '%s'", node, master_addr ? master_addr : ""));
    uint16_t my_port = 0;
    lfd = listen_on(0, &my_port);
    if (lfd < 0) return bail(fail("node %d: cannot open a listening socket: %s", node, strerror(errno)));
    const int fd0 = dial(ip0, uint16_t(port), deadline);
    if (fd0 < 0) return bail(fail("node %d: node 0 is not reachable at %s:%d", node, master_addr, port));
    n->sock[0] = fd0;
    set_io_timeout(fd0, n->timeout_s);
    const uint32_t hello[3] = {0x464d4e54u, uint32_t(node), uint32_t(my_port)};
    if (!send_all(fd0, hello, sizeof(hello)) || !recv_all(fd0, table.data(), table.size() * sizeof(Addr)))
      return bail(fail("node %d: rendezvous with node 0 failed", node));
    // lower-numbered leaders (other than 0) are dialled, higher-numbered ones dial us
    for (int p = 1; p < node; ++p) {
      const int fd = dial(table[size_t(p)].ip, table[size_t(p)].port, deadline);
      if (fd < 0) return bail(fail("node %d: node %d is not reachable", node, p));
      const uint32_t me = uint32_t(node);
      set_io_timeout(fd, n->timeout_s);
      if (!send_all(fd, &me, sizeof(me))) { close(fd); return bail(fail("node %d: hello to node %d failed", node, p)); }
      n->sock[size_t(p)] = fd;
    }
    for (int k = node + 1; k < nnodes; ++k) {
      const int fd = accept_until(lfd, deadline, nullptr);
      if (fd < 0) return bail(fail("node %d: timed out waiting for higher-numbered nodes", node));
      set_io_timeout(fd, n->timeout_s);
      uint32_t who = 0;
      if (!recv_all(fd, &who, sizeof(who)) || who <= uint32_t(node) || who >= uint32_t(nnodes) ||
          n->sock[who] >= 0) {
        close(fd);
        return bail(fail("node %d: unexpected connection on the mesh port", node));
      }
      n->sock[who] = fd;
    }
  }
  close(lfd);
  lfd = -1;
  for (int q = 0; q < nnodes; ++q) {
    if (q == node) continue;
    tune(n->sock[size_t(q)]);
    set_nonblocking(n->sock[size_t(q)], true);
  }
  *out = n;
  return 0;
}

int fmpi_net_finalize(fmpi_net* n) {
  delete n;
  return 0;
}

int fmpi_net_barrier(fmpi_net* n) {
  if (!n) return fail("null mesh");
  if (n->nnodes == 1) return 0;
  std::vector<uint8_t> in(size_t(n->nnodes), 0);
  const uint8_t token = 1;
  std::vector<Xfer> xs;
  for (int q = 0; q < n->nnodes; ++q)
    if (q != n->node) xs.push_back({q, &token, 1, &in[size_t(q)], 1});
  return n->exchange(xs);
}

int fmpi_net_allreduce(fmpi_net* n, void* buf, size_t count, int dtype, int op) {
  if (!n) return fail("null mesh");
  const size_t es = dtype_size(dtype);
  if (es == 0) return fail("unsupported dtype %d", dtype);
  if (n->nnodes == 1 || count == 0) return 0;
  uint8_t* b = static_cast<uint8_t*>(buf);
  if (n->scatter_reduce(b, count, dtype, op)) return 1;
  size_t lo, hi;
  n->slice(count, n->node, &lo, &hi);
  std::vector<Xfer> xs;
  for (int q = 0; q < n->nnodes; ++q) {
    if (q == n->node) continue;
    size_t qlo, qhi;
    n->slice(count, q, &qlo, &qhi);
    xs.push_back({q, b + lo * es, (hi - lo) * es, b + qlo * es, (qhi - qlo) * es});
  }
  return n->exchange(xs);
}

int fmpi_net_reduce(fmpi_net* n, void* buf, size_t count, int dtype, int op, int root_node) {
  if (!n) return fail("null mesh");
  const size_t es = dtype_size(dtype);
  if (es == 0) return fail("unsupported dtype %d", dtype);
  if (root_node < 0 || root_node >= n->nnodes) return fail("bad root node %d", root_node);
  if (n->nnodes == 1 || count == 0) return 0;
  uint8_t* b = static_cast<uint8_t*>(buf);
  if (n->scatter_reduce(b, count, dtype, op)) return 1;
  std::vector<Xfer> xs;
  if (n->node == root_node) {
    for (int q = 0; q < n->nnodes; ++q) {
      if (q == n->node) continue;
      size_t qlo, qhi;
      n->slice(count, q, &qlo, &qhi);
      xs.push_back({q, nullptr, 0, b + qlo * es, (qhi - qlo) * es});
    }
  } else {
    size_t lo, hi;
    n->slice(count, n->node, &lo, &hi);
    xs.push_back({root_node, b + lo * es, (hi - lo) * es, nullptr, 0});
  }
  return n->exchange(xs);
}

int fmpi_net_bcast(fmpi_net* n, void* buf, size_t nbytes, int root_node) {
  if (!n) return fail("null mesh");
  if (root_node < 0 || root_node >= n->nnodes) return fail("bad root node %d", root_node);
  if (n->nnodes == 1 || nbytes == 0) return 0;
  uint8_t* b = static_cast<uint8_t*>(buf);
  std::vector<Xfer> xs;
  if (n->node == root_node) {
    for (int q = 0; q < n->nnodes; ++q)
      if (q != n->node) xs.push_back({q, b, nbytes, nullptr, 0});
  } else {
    xs.push_back({root_node, nullptr, 0, b, nbytes});
  }
  return n->exchange(xs);
}

int fmpi_net_allgather(fmpi_net* n, const void* send, void* recv, size_t nbytes_each) {
  if (!n) return fail("null mesh");
  uint8_t* r = static_cast<uint8_t*>(recv);
  uint8_t* mine = r + size_t(n->node) * nbytes_each;
  if (mine != send) memmove(mine, send, nbytes_each);
  if (n->nnodes == 1 || nbytes_each == 0) return 0;
  std::vector<Xfer> xs;
  for (int q = 0; q < n->nnodes; ++q)
    if (q != n->node) xs.push_back({q, mine, nbytes_each, r + size_t(q) * nbytes_each, nbytes_each});
  return n->exchange(xs);
}

}  // extern "C"
