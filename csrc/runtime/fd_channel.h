// File-descriptor passing between the ranks of one host (used to share cuMem allocation handles).
#pragma once
#include <errno.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <sys/un.h>
#include <unistd.h>

#include <string>

namespace ft {

// Every rank binds an abstract unix datagram socket named from (token, rank); an fd is sent with
// SCM_RIGHTS together with the sender's rank and a user tag.
struct FdChannel {
  int sock = -1;
  unsigned long long token = 0;
  int rank = 0;

  static void make_addr(unsigned long long token, int rank, sockaddr_un* a, socklen_t* len) {
    memset(a, 0, sizeof *a);
    a->sun_family = AF_UNIX;
    int n = snprintf(a->sun_path + 1, sizeof(a->sun_path) - 1, "flextree.%llx.%d", token, rank);
    *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
  }
  bool open(unsigned long long tok, int r, std::string* err) {
    token = tok;
    rank = r;
    sock = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
    if (sock < 0) { if (err) *err = std::string("socket: ") + strerror(errno); return false; }
    sockaddr_un a;
    socklen_t len;
    make_addr(token, rank, &a, &len);
    if (bind(sock, (sockaddr*)&a, len) != 0) {
      if (err) *err = std::string("bind: ") + strerror(errno);
      return false;
    }
    timeval tv{60, 0};
    setsockopt(sock, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
    return true;
  }
  void close_() {
    if (sock >= 0) ::close(sock);
    sock = -1;
  }
  bool send_fd(int peer, int fd, int tag, std::string* err) {
    sockaddr_un a;
    socklen_t len;
    make_addr(token, peer, &a, &len);
    int payload[2] = {rank, tag};
    iovec iov{payload, sizeof payload};
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof ctrl);
    msghdr msg{};
    msg.msg_name = &a;
    msg.msg_namelen = len;
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof ctrl;
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    c->cmsg_level = SOL_SOCKET;
    c->cmsg_type = SCM_RIGHTS;
    c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
    for (int tries = 0; tries < 200; ++tries) {
      if (sendmsg(sock, &msg, 0) >= 0) return true;
      if (errno != ECONNREFUSED && errno != ENOENT && errno != EAGAIN && errno != ENOBUFS) break;
      usleep(10000);
    }
    if (err) *err = std::string("sendmsg(SCM_RIGHTS): ") + strerror(errno);
    return false;
  }
  bool recv_fd(int* src, int* tag, int* fd, std::string* err) {
    int payload[2] = {-1, -1};
    iovec iov{payload, sizeof payload};
    char ctrl[CMSG_SPACE(sizeof(int))];
    msghdr msg{};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof ctrl;
    ssize_t n = recvmsg(sock, &msg, MSG_CMSG_CLOEXEC);
    if (n < 0) { if (err) *err = std::string("recvmsg: ") + strerror(errno); return false; }
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    if (!c || c->cmsg_type != SCM_RIGHTS) { if (err) *err = "recvmsg: no fd attached"; return false; }
    memcpy(fd, CMSG_DATA(c), sizeof(int));
    *src = payload[0];
    *tag = payload[1];
    return true;
  }
};


}  // namespace ft
