// Launcher-based bootstrap for C/C++ programs: one process per GPU, no Python and no MPI in the process.
//
// The reference is a C++ program started by an MPI launcher (allreduce_over_mpi/benchmark.cpp:48-52:
// MPI_Init_thread / MPI_Comm_size / MPI_Comm_rank; hostfile mpi_config_file:1-16). The equivalent here is
//     ftrun -np N ./program args...
// (csrc/tools/ftrun.cc), which forks N copies with FT_RANK / FT_SIZE / FT_LOCAL_RANK / FT_BOOT_TOKEN in the
// environment; ft_comm_create_from_env() then builds the communicator. The only thing a communicator needs from the
// outside world is a byte all-gather (ft_allgather_fn) for exchanging memory handles - here it runs over abstract
// unix datagram sockets named from the token, the same kind of socket the VMM provider already uses to pass file
// descriptors (fd_channel.h), so a launcher-started C program gets VMM segments, the multicast object and the NVLS
// kernels exactly like a torchrun-started Python program.
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <sys/un.h>
#include <unistd.h>

#include <deque>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/flextree.h"

namespace ft {
void set_last_error(const std::string& s);  // planner/c_api.cc
}

namespace {

constexpr size_t kChunk = 32 << 10;  // payload bytes per datagram

struct Header {
  uint32_t seq, chunk;
  int32_t src;
  uint32_t len;
};

struct Stashed {
  Header h;
  std::vector<char> data;
};

struct EnvBoot {
  int rank = -1, size = 0, sock = -1;
  unsigned long long token = 0;
  uint32_t seq = 0;
  std::deque<Stashed> stash;  // datagrams of a LATER collective that arrived early (a fast peer is already there)
  std::mutex mu;
  bool ready = false;

  void addr(int r, sockaddr_un* a, socklen_t* len) const {
    memset(a, 0, sizeof *a);
    a->sun_family = AF_UNIX;
    int n = snprintf(a->sun_path + 1, sizeof(a->sun_path) - 1, "flextree.boot.%llx.%d", token, r);
    *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
  }
  bool init(std::string* err) {
    if (ready) return true;
    const char* r = getenv("FT_RANK");
    const char* s = getenv("FT_SIZE");
    const char* t = getenv("FT_BOOT_TOKEN");
    if (!r || !s || !t) { *err = "FT_RANK / FT_SIZE / FT_BOOT_TOKEN are not set: start the program with ftrun -np N"; return false; }
    rank = atoi(r);
    size = atoi(s);
    token = strtoull(t, nullptr, 16);
    if (size < 1 || rank < 0 || rank >= size) { *err = "bad FT_RANK / FT_SIZE"; return false; }
    sock = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
    if (sock < 0) { *err = std::string("socket: ") + strerror(errno); return false; }
    sockaddr_un a;
    socklen_t len;
    addr(rank, &a, &len);
    if (bind(sock, (sockaddr*)&a, len) != 0) { *err = std::string("bind: ") + strerror(errno); return false; }
    int buf = 4 << 20;
    setsockopt(sock, SOL_SOCKET, SO_RCVBUF, &buf, sizeof buf);
    setsockopt(sock, SOL_SOCKET, SO_SNDBUF, &buf, sizeof buf);
    timeval tv{120, 0};
    setsockopt(sock, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
    ready = true;
    return true;
  }
  // One non-blocking attempt: 1 sent, 0 try again later (peer not bound yet, or its queue is full), -1 error.
  int try_send(int peer, const Header& h, const char* data) {
    sockaddr_un a;
    socklen_t len;
    addr(peer, &a, &len);
    iovec iov[2] = {{(void*)&h, sizeof h}, {(void*)data, h.len}};
    msghdr msg{};
    msg.msg_name = &a;
    msg.msg_namelen = len;
    msg.msg_iov = iov;
    msg.msg_iovlen = h.len ? 2 : 1;
    if (sendmsg(sock, &msg, MSG_DONTWAIT) >= 0) return 1;
    if (errno == ECONNREFUSED || errno == ENOENT || errno == EAGAIN || errno == EWOULDBLOCK || errno == ENOBUFS || errno == EINTR) return 0;
    return -1;
  }
  // every rank contributes `bytes`, everybody receives size * bytes in rank order.
  // Sends and receives are interleaved: a unix datagram socket queues only a handful of datagrams
  // (net.unix.max_dgram_qlen, 10 by default), so "send everything, then receive" deadlocks as soon as every rank has
  // more to send than its peers' queues hold.
  int allgather(const void* send, void* recv, size_t bytes) {
    std::lock_guard<std::mutex> lk(mu);
    const uint32_t s = seq++;
    char* out = (char*)recv;
    memcpy(out + (size_t)rank * bytes, send, bytes);
    const uint32_t nchunks = bytes ? (uint32_t)((bytes + kChunk - 1) / kChunk) : 1;
    const size_t want = (size_t)(size - 1) * nchunks, to_send = want;
    size_t got = 0, sent = 0;
    auto place = [&](const Header& h, const char* d) {
      if (h.src < 0 || h.src >= size || (size_t)h.chunk * kChunk + h.len > bytes) return false;
      memcpy(out + (size_t)h.src * bytes + (size_t)h.chunk * kChunk, d, h.len);
      return true;
    };
    for (auto it = stash.begin(); it != stash.end();) {
      if (it->h.seq == s) {
        if (!place(it->h, it->data.data())) return 1;
        ++got;
        it = stash.erase(it);
      } else {
        ++it;
      }
    }
    std::vector<char> buf(sizeof(Header) + kChunk);
    timeval t0;
    gettimeofday(&t0, nullptr);
    while (got < want || sent < to_send) {
      bool progress = false;
      if (sent < to_send) {  // message number `sent`: peers round-robin starting behind me, chunks in order
        const int pi = (int)(sent % (size_t)(size - 1));
        const uint32_t c = (uint32_t)(sent / (size_t)(size - 1));
        const int p = (rank + 1 + pi) % size;
        Header h{s, c, rank, (uint32_t)(bytes - (size_t)c * kChunk < kChunk ? bytes - (size_t)c * kChunk : kChunk)};
        const int r = try_send(p, h, (const char*)send + (size_t)c * kChunk);
        if (r < 0) return 1;
        if (r > 0) { ++sent; progress = true; }
      }
      ssize_t n = ::recv(sock, buf.data(), buf.size(), MSG_DONTWAIT);
      if (n >= (ssize_t)sizeof(Header)) {
        Header h;
        memcpy(&h, buf.data(), sizeof h);
        if ((size_t)n != sizeof(Header) + h.len) return 1;
        if (h.seq == s) {
          if (!place(h, buf.data() + sizeof h)) return 1;
          ++got;
        } else if (h.seq > s) {
          Stashed st;
          st.h = h;
          st.data.assign(buf.data() + sizeof h, buf.data() + sizeof h + h.len);
          stash.push_back(std::move(st));
        } else {
          return 1;  // a datagram of a collective that is already complete: protocol error
        }
        progress = true;
      } else if (n >= 0 || (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR)) {
        return 1;
      }
      if (!progress) {
        timeval t1;
        gettimeofday(&t1, nullptr);
        if (t1.tv_sec - t0.tv_sec > 180) return 1;  // a peer never showed up
        usleep(200);
      } else {
        gettimeofday(&t0, nullptr);
      }
    }
    return 0;
  }
};

EnvBoot g_boot;

int env_allgather(void* ctx, const void* send, void* recv, size_t bytes) { return ((EnvBoot*)ctx)->allgather(send, recv, bytes); }

}  // namespace

extern "C" {

int ft_env_rank(void) {
  const char* r = getenv("FT_RANK");
  return (r && getenv("FT_SIZE")) ? atoi(r) : -1;
}
int ft_env_size(void) {
  const char* s = getenv("FT_SIZE");
  return (s && getenv("FT_RANK")) ? atoi(s) : -1;
}
int ft_env_local_rank(void) {
  const char* s = getenv("FT_LOCAL_RANK");
  return s ? atoi(s) : ft_env_rank();
}

int ft_env_allgather(const void* send, void* recv, size_t bytes) {
  std::string err;
  if (!g_boot.init(&err)) { ft::set_last_error(err); return 1; }
  return g_boot.allgather(send, recv, bytes);
}
int ft_env_barrier(void) {
  char x = 0;
  std::vector<char> all((size_t)(ft_env_size() > 0 ? ft_env_size() : 1));
  return ft_env_allgather(&x, all.data(), 1);
}

ft_result_t ft_comm_create_from_env(ft_comm_t** out, int device, const ft_comm_options_t* opts) {
  std::string err;
  if (!g_boot.init(&err)) { ft::set_last_error(err); return FT_ERR_INVALID; }
  if (device < 0) device = ft_env_local_rank();
  return ft_comm_create(out, g_boot.rank, g_boot.size, device, env_allgather, &g_boot, opts);
}

}  // extern "C"
