// An ordinary MPI program moved to this library by changing the include path (no source change):
//
//   g++ -Icompat -Iinclude examples/mpi_program.cc -Lbuild -lflextree_host -lpthread -lrt  $(your MPI's flags)
//   mpirun -np 6 ./a.out            # FT_TOPO=2,3 picks the tree; unset = flat; any 1 = ring
//
// On this image (no MPI installed): add -Ibaseline/mpi_shim baseline/mpi_shim/mpi_shim.cc and launch with
// baseline/_ref/bin/ftmpirun -np 6 ./a.out.  compat/mpi_mod.hpp captures MPI_Allreduce exactly like the reference's
// header does (allreduce_over_mpi/mpi_mod.hpp:1167-1172) and runs it over shared memory.
#include <mpi.h>
#include <stdio.h>

#include <vector>

#include "mpi_mod.hpp"

int main(int argc, char** argv) {
  int provided = 0, rank = 0, n = 1;
  MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided);
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &n);
  std::vector<float> grad(1 << 20, (float)(rank + 1));
  MPI_Allreduce(MPI_IN_PLACE, grad.data(), (int)grad.size(), MPI_FLOAT, MPI_SUM, MPI_COMM_WORLD);
  if (rank == 0) printf("sum over %d ranks: %.0f (expected %d)\n", n, grad[12345], n * (n + 1) / 2);
  MPI_Finalize();
  return 0;
}
