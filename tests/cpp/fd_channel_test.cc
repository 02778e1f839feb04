// CPU test of the SCM_RIGHTS fd exchange used by the VMM symmetric-memory provider: N forked "ranks"
// each create a memfd holding their rank, send it to every peer and read every peer's value back.
#include <stdio.h>
#include <stdlib.h>
#include <sys/mman.h>
#include <sys/wait.h>

#include <vector>

#include "../../csrc/runtime/fd_channel.h"

int main(int argc, char** argv) {
  int n = argc > 1 ? atoi(argv[1]) : 4;
  unsigned long long tok = ((unsigned long long)getpid() << 20) ^ 0xabcdefull;
  std::vector<pid_t> pids;
  for (int r = 0; r < n; ++r) {
    pid_t pid = fork();
    if (pid == 0) {
      std::string err;
      ft::FdChannel ch;
      if (!ch.open(tok, r, &err)) { fprintf(stderr, "rank %d open: %s\n", r, err.c_str()); _exit(1); }
      int fd = memfd_create("x", 0);
      if (ftruncate(fd, 4096) != 0) _exit(1);
      int* p = (int*)mmap(nullptr, 4096, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
      p[0] = 1000 + r;
      usleep(1000 * (r * 37 % 50));  // skew: peers may not be bound yet -> send_fd retries
      for (int q = 0; q < n; ++q)
        if (q != r && !ch.send_fd(q, fd, 0, &err)) { fprintf(stderr, "rank %d send: %s\n", r, err.c_str()); _exit(1); }
      int seen = 0;
      for (int i = 0; i < n - 1; ++i) {
        int src, tag, rfd;
        if (!ch.recv_fd(&src, &tag, &rfd, &err)) { fprintf(stderr, "rank %d recv: %s\n", r, err.c_str()); _exit(1); }
        int* q = (int*)mmap(nullptr, 4096, PROT_READ, MAP_SHARED, rfd, 0);
        if (q == MAP_FAILED || q[0] != 1000 + src || tag != 0) { fprintf(stderr, "rank %d bad payload from %d\n", r, src); _exit(1); }
        seen |= 1 << src;
        close(rfd);
      }
      if (seen != (((1 << n) - 1) & ~(1 << r))) _exit(1);
      ch.close_();
      _exit(0);
    }
    pids.push_back(pid);
  }
  int bad = 0;
  for (pid_t p : pids) {
    int st;
    waitpid(p, &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st)) bad = 1;
  }
  printf("fd_channel %d ranks: %s\n", n, bad ? "FAIL" : "ok");
  return bad;
}
