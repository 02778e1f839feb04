// ftmpirun -np N <program> [args...] : launcher for the single-host MPI shim (see mpi.h).
// Creates the shared control block and one sparse arena per rank as memfds, then fork+execs N copies
// of the program with FTMPI_* in the environment. Exit status: first nonzero child status.
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <string>
#include <vector>

#ifndef FTMPI_CONTROL_BYTES
#define FTMPI_CONTROL_BYTES (1100ull << 20)  // >= sizeof(Control) in mpi_shim.cc (sparse)
#endif

int main(int argc, char** argv) {
  int np = 0;
  int i = 1;
  for (; i < argc; ++i) {
    if ((!strcmp(argv[i], "-np") || !strcmp(argv[i], "-n")) && i + 1 < argc) np = atoi(argv[++i]);
    else break;
  }
  if (np < 1 || np > 64 || i >= argc) {
    fprintf(stderr, "usage: ftmpirun -np N program [args...]\n");
    return 2;
  }
  int ctrl = memfd_create("ftmpi-ctrl", 0);
  if (ctrl < 0 || ftruncate(ctrl, (off_t)FTMPI_CONTROL_BYTES) != 0) { perror("memfd ctrl"); return 2; }
  std::vector<int> arenas;
  for (int r = 0; r < np; ++r) {
    int fd = memfd_create("ftmpi-arena", 0);
    if (fd < 0 || ftruncate(fd, (off_t)(16ull << 30)) != 0) { perror("memfd arena"); return 2; }
    if (r && fd != arenas[0] + r) { fprintf(stderr, "ftmpirun: arena fds not consecutive\n"); return 2; }
    arenas.push_back(fd);
  }
  std::vector<pid_t> pids;
  for (int r = 0; r < np; ++r) {
    pid_t pid = fork();
    if (pid < 0) { perror("fork"); return 2; }
    if (pid == 0) {
      setenv("FTMPI_RANK", std::to_string(r).c_str(), 1);
      setenv("FTMPI_SIZE", std::to_string(np).c_str(), 1);
      setenv("FTMPI_CTRL_FD", std::to_string(ctrl).c_str(), 1);
      setenv("FTMPI_ARENA_FD0", std::to_string(arenas[0]).c_str(), 1);
      execvp(argv[i], argv + i);
      perror("execvp");
      _exit(127);
    }
    pids.push_back(pid);
  }
  int rc = 0;
  for (pid_t p : pids) {
    int st = 0;
    waitpid(p, &st, 0);
    int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
    if (code && !rc) rc = code;
  }
  return rc;
}
