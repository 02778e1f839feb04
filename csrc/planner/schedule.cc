#include "schedule.h"

#include <sstream>

namespace ft {

Schedule build_schedule(const Plan& plan, int rank, int lonely) {
  Schedule s;
  s.plan = plan;
  s.rank = rank;
  s.lonely = lonely;
  const int N = plan.nranks;
  if (rank < N) {
    s.send.resize(plan.nstages);
    s.recv.resize(plan.nstages);
    for (int k = 0; k < plan.nstages; ++k) {
      const int gn = plan.gap[k + 1];
      for (int j = 0; j < plan.width[k]; ++j) {
        int peer = stage_peer(plan, rank, k, j);
        PeerBlocks snd{peer, {}}, rcv{peer, {}};
        for (int b = peer % gn; b < N; b += gn) snd.blocks.push_back(b);
        for (int b = rank % gn; b < N; b += gn) rcv.blocks.push_back(b);
        s.send[k].push_back(snd);
        s.recv[k].push_back(rcv);
      }
    }
    // tree ranks receive their fully-owned block from every lonely rank
    for (int l = 0; l < lonely; ++l) s.lonely_recv.push_back({N + l, {rank}});
  } else {
    // a lonely rank sends block i to tree rank i
    for (int i = 0; i < N; ++i) s.lonely_send.push_back({i, {i}});
  }
  return s;
}

std::string print_ops(const Schedule& s) {
  std::ostringstream o;
  auto dump = [&](const char* title, const std::vector<std::vector<PeerBlocks>>& ops) {
    o << title << " (rank " << s.rank << ")\n";
    for (size_t k = 0; k < ops.size(); ++k) {
      o << "  stage " << k << " [width " << s.plan.width[k] << ", gap " << s.plan.gap[k] << "]\n";
      for (auto& pb : ops[k]) {
        o << "    |- peer " << pb.peer << (pb.peer == s.rank ? " (self)" : "") << ": blocks";
        for (int b : pb.blocks) o << ' ' << b;
        o << '\n';
      }
    }
  };
  dump("send ops", s.send);
  dump("recv ops", s.recv);
  if (!s.lonely_send.empty() || !s.lonely_recv.empty()) {
    o << "lonely ops (rank " << s.rank << ")\n";
    for (auto& pb : s.lonely_send) o << "    |- send block " << pb.blocks[0] << " to " << pb.peer << '\n';
    for (auto& pb : s.lonely_recv) o << "    |- recv block " << pb.blocks[0] << " from " << pb.peer << '\n';
  }
  return o.str();
}

Traffic traffic(const Schedule& s, long long count) {
  Traffic t;
  const int N = s.plan.nranks;
  const long long B = ceil_div_ll(count, N);
  auto blen = [&](int b) {
    long long lo = (long long)b * B, hi = lo + B;
    if (hi > count) hi = count;
    return hi > lo ? hi - lo : 0;
  };
  for (size_t k = 0; k < s.send.size(); ++k) {
    for (auto& pb : s.send[k]) {  // reduce-scatter: send what the peer owns
      if (pb.peer == s.rank) continue;
      for (int b : pb.blocks)
        if (blen(b)) { t.messages++; t.elements += blen(b); }
    }
    for (auto& pb : s.recv[k]) {  // all-gather: send what I own to every group peer
      if (pb.peer == s.rank) continue;
      for (int b : pb.blocks)
        if (blen(b)) { t.messages++; t.elements += blen(b); }
    }
  }
  return t;
}

}  // namespace ft
