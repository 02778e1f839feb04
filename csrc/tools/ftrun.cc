// ftrun -np N [--gpus G] program [args...] : start N copies of a program, one per GPU, for the C API of the library.
//
// Capability parity: the reference is started by an MPI launcher with a hostfile (allreduce_over_mpi/
// mpi_config_file:1-16, Makefile:8-24 pushes the binary to the hosts). One NVSwitch box needs neither: the launcher
// forks the ranks and hands each one FT_RANK, FT_SIZE, FT_LOCAL_RANK (GPU ordinal, rank mod G) and FT_BOOT_TOKEN, from
// which ft_comm_create_from_env() (csrc/runtime/env_bootstrap.cc) finds its peers over abstract unix sockets.
// Exit status: the first nonzero status of a rank (128 + signal for a killed rank); when a rank fails the others get
// SIGTERM so that nobody waits forever on a peer that is gone.
#include <errno.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <sys/wait.h>
#include <unistd.h>

#include <string>
#include <vector>

int main(int argc, char** argv) {
  int np = 0, gpus = 0;
  int i = 1;
  for (; i < argc; ++i) {
    if ((!strcmp(argv[i], "-np") || !strcmp(argv[i], "-n") || !strcmp(argv[i], "--np")) && i + 1 < argc) np = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--gpus") && i + 1 < argc) gpus = atoi(argv[++i]);
    else break;
  }
  if (np < 1 || np > 64 || i >= argc) {
    fprintf(stderr, "usage: ftrun -np N [--gpus G] program [args...]\n");
    return 2;
  }
  if (gpus <= 0) gpus = np;
  timeval tv;
  gettimeofday(&tv, nullptr);
  char token[32];
  snprintf(token, sizeof token, "%llx", ((unsigned long long)getpid() << 24) ^ (unsigned long long)tv.tv_usec ^ ((unsigned long long)tv.tv_sec << 8));
  std::vector<pid_t> pids;
  for (int r = 0; r < np; ++r) {
    pid_t pid = fork();
    if (pid < 0) { perror("fork"); return 2; }
    if (pid == 0) {
      setenv("FT_RANK", std::to_string(r).c_str(), 1);
      setenv("FT_SIZE", std::to_string(np).c_str(), 1);
      setenv("FT_LOCAL_RANK", std::to_string(r % gpus).c_str(), 1);
      setenv("FT_BOOT_TOKEN", token, 1);
      execvp(argv[i], argv + i);
      perror("execvp");
      _exit(127);
    }
    pids.push_back(pid);
  }
  int rc = 0;
  size_t left = pids.size();
  while (left > 0) {
    int st = 0;
    pid_t p = wait(&st);
    if (p < 0) {
      if (errno == EINTR) continue;
      break;
    }
    --left;
    const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
    if (code && !rc) {
      rc = code;
      for (pid_t q : pids)
        if (q != p) kill(q, SIGTERM);
    }
  }
  return rc;
}
