// Launcher bootstrap without a GPU: `ftrun -np N env_boot_test` - every rank all-gathers payloads of several sizes
// (one byte, a handle-sized blob, more than one datagram) back to back and checks every slot; rank-ordered output
// through ft_env_barrier. Exercises csrc/runtime/env_bootstrap.cc + csrc/tools/ftrun.cc (the stand-in for mpirun).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "flextree.h"

int main() {
  const int rank = ft_env_rank(), size = ft_env_size();
  if (rank < 0 || size < 1) { fprintf(stderr, "not started by ftrun\n"); return 2; }
  const size_t sizes[] = {1, 64, 8, 100000, 3, 32768, 65537};
  for (int round = 0; round < 3; ++round) {
    for (size_t bytes : sizes) {
      std::vector<unsigned char> mine(bytes), all(bytes * (size_t)size);
      for (size_t i = 0; i < bytes; ++i) mine[i] = (unsigned char)((rank * 131 + i * 7 + round) & 0xff);
      if (ft_env_allgather(mine.data(), all.data(), bytes) != 0) { fprintf(stderr, "rank %d: allgather failed: %s\n", rank, ft_last_error()); return 1; }
      for (int r = 0; r < size; ++r)
        for (size_t i = 0; i < bytes; ++i)
          if (all[(size_t)r * bytes + i] != (unsigned char)((r * 131 + i * 7 + round) & 0xff)) {
            fprintf(stderr, "rank %d: wrong byte from rank %d at %zu (size %zu)\n", rank, r, i, bytes);
            return 1;
          }
    }
  }
  for (int r = 0; r < size; ++r) {
    fflush(stdout);
    if (ft_env_barrier() != 0) return 1;
    if (r == rank) printf("ENV_BOOT_OK %d/%d local %d\n", rank, size, ft_env_local_rank());
  }
  fflush(stdout);
  ft_env_barrier();
  return 0;
}
