// Host-side, table-form view of a plan: who sends which blocks to whom in every stage.
// Only used for printing (`ft_plan --print-ops`), tests and the byte/message accounting; kernels use
// the closed forms in core/plan.h.
//
// Capability parity: Operation / Send_Ops / Recv_Ops::generate_ops and print_ops
// (allreduce_over_mpi/mpi_mod.hpp:45-214), including the lonely-rank ops (:172-178, :208-211).
#pragma once
#include <string>
#include <vector>

#include "../core/plan.h"
#include "factorize.h"

namespace ft {

struct PeerBlocks {
  int peer;
  std::vector<int> blocks;
};

struct Schedule {
  Plan plan;
  int rank;
  int lonely;  // number of lonely ranks attached to the tree (ranks plan.nranks .. plan.nranks+lonely-1)
  // [stage][j] - reduce-scatter direction. send[k][j].blocks are what `rank` sends to send[k][j].peer;
  // recv[k][j].blocks are what it receives from recv[k][j].peer. Self entries are included (as in the
  // reference tables) and skipped by the data plane.
  std::vector<std::vector<PeerBlocks>> send, recv;
  std::vector<PeerBlocks> lonely_send, lonely_recv;
};

// rank may be >= plan.nranks to build the schedule of a lonely rank.
Schedule build_schedule(const Plan& plan, int rank, int lonely = 0);
std::string print_ops(const Schedule& s);

struct Traffic {
  long long messages = 0;  // block-granular messages, as the reference sends them (one per block)
  long long elements = 0;  // elements sent by this rank over RS + AG
};
Traffic traffic(const Schedule& s, long long count);

}  // namespace ft
