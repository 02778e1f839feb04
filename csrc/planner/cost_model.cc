#include "cost_model.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

namespace ft {

LegacyTerms legacy_cost(const Widths& tree, int total_nodes, double chunk, const LegacyCoeffs& c) {
  LegacyTerms t;
  const double n = total_nodes;
  for (int w : tree) {
    t.latency += 2.0 * c.lo;
    if (w > c.incast_knee) t.latency += c.latency_chunk * (w - c.incast_knee) * c.co;
  }
  // steps = N + 1 + 2 * (tree[0] + tree[0]*tree[1] + ... up to height-1 prefixes)
  double steps = n + 1.0;
  double prefix = 1.0;
  for (size_t k = 0; k + 1 < tree.size(); ++k) {
    prefix *= tree[k];
    steps += 2.0 * prefix;
  }
  t.memory = n > 0 ? (steps * chunk / n) * c.o : 0.0;
  t.bandwidth = n > 0 ? ((n - 1.0) / n) * chunk * c.bo : 0.0;
  return t;
}

int legacy_choose(const std::vector<Widths>& trees, int total_nodes, double chunk, double* best_cost,
                  const LegacyCoeffs& c) {
  int best = -1;
  double bc = 1e9;  // same sentinel as the reference
  for (size_t i = 0; i < trees.size(); ++i) {
    double cost = legacy_cost(trees[i], total_nodes, chunk, c).total();
    if (cost < bc) {
      bc = cost;
      best = (int)i;
    }
  }
  if (best_cost) *best_cost = bc;
  return best;
}

const char* algo_name(int a) {
  switch (a) {
    case ALGO_AUTO: return "auto";
    case ALGO_ONESHOT: return "oneshot";
    case ALGO_STAGED: return "stages";
    case ALGO_NVLS: return "nvls";
    case ALGO_RING: return "ring";
    case ALGO_TWOSHOT: return "twoshot";
  }
  return "?";
}

int algo_from_name(const std::string& s) {
  if (s == "auto" || s.empty()) return ALGO_AUTO;
  if (s == "oneshot" || s == "ll") return ALGO_ONESHOT;
  if (s == "stages" || s == "staged" || s == "tree" || s == "flextree") return ALGO_STAGED;
  if (s == "nvls" || s == "multicast") return ALGO_NVLS;
  if (s == "ring") return ALGO_RING;
  if (s == "twoshot" || s == "ll2") return ALGO_TWOSHOT;
  return -1;
}

// Minimal reader for a flat {"key": number|true|false, ...} JSON object.
bool load_link_model(const std::string& path, LinkModel* m, std::string* err) {
  std::ifstream f(path);
  if (!f) {
    if (err) *err = "cannot open " + path;
    return false;
  }
  std::stringstream ss;
  ss << f.rdbuf();
  std::string s = ss.str();
  size_t pos = 0;
  int found = 0;
  while ((pos = s.find('"', pos)) != std::string::npos) {
    size_t end = s.find('"', pos + 1);
    if (end == std::string::npos) break;
    std::string key = s.substr(pos + 1, end - pos - 1);
    size_t colon = s.find_first_not_of(" \t\r\n", end + 1);
    pos = end + 1;
    if (colon == std::string::npos || s[colon] != ':') continue;
    size_t v = s.find_first_not_of(" \t\r\n", colon + 1);
    if (v == std::string::npos) break;
    double val;
    if (!s.compare(v, 4, "true")) val = 1;
    else if (!s.compare(v, 5, "false")) val = 0;
    else {
      char* e = nullptr;
      val = std::strtod(s.c_str() + v, &e);
      if (e == s.c_str() + v) continue;  // not a number (string/nested) - skip
    }
    pos = v;
#define FT_KEY(name, cast) if (key == #name) { m->name = (cast)val; ++found; continue; }
    FT_KEY(alpha_launch_us, double) FT_KEY(alpha_sync_us, double) FT_KEY(alpha_ll_us, double)
    FT_KEY(alpha_peer_us, double) FT_KEY(alpha_nvls_us, double) FT_KEY(beta_ll_self_GBps, double)
    FT_KEY(beta_p2p_GBps, double) FT_KEY(beta_p2p_1peer_GBps, double) FT_KEY(fanin_knee, int)
    FT_KEY(beta_nvls_GBps, double) FT_KEY(beta_ll_GBps, double) FT_KEY(gamma_hbm_GBps, double)
    FT_KEY(half_bw_bytes, double) FT_KEY(has_nvls, bool) FT_KEY(max_oneshot_bytes, int) FT_KEY(max_twoshot_bytes, int)
#undef FT_KEY
  }
  if (!found && err) *err = "no known keys in " + path;
  return found > 0;
}

std::string dump_link_model(const LinkModel& m) {
  char buf[2048];
  std::snprintf(buf, sizeof buf,
                "{\"alpha_launch_us\": %.4g, \"alpha_sync_us\": %.4g, \"alpha_peer_us\": %.4g, \"alpha_ll_us\": %.4g, "
                "\"alpha_nvls_us\": %.4g, \"beta_p2p_GBps\": %.5g, \"beta_p2p_1peer_GBps\": %.5g, \"fanin_knee\": %d, "
                "\"beta_nvls_GBps\": %.5g, \"beta_ll_GBps\": %.5g, \"beta_ll_self_GBps\": %.5g, \"gamma_hbm_GBps\": %.5g, "
                "\"half_bw_bytes\": %.6g, \"has_nvls\": %s, \"max_oneshot_bytes\": %d, \"max_twoshot_bytes\": %d}",
                m.alpha_launch_us, m.alpha_sync_us, m.alpha_peer_us, m.alpha_ll_us, m.alpha_nvls_us, m.beta_p2p_GBps,
                m.beta_p2p_1peer_GBps, m.fanin_knee, m.beta_nvls_GBps, m.beta_ll_GBps, m.beta_ll_self_GBps,
                m.gamma_hbm_GBps, m.half_bw_bytes, m.has_nvls ? "true" : "false", m.max_oneshot_bytes, m.max_twoshot_bytes);
  return buf;
}

namespace {
// Pull bandwidth as a function of how many peers are read concurrently.
double beta_fanin(const LinkModel& m, int peers) {
  if (peers <= 1) return m.beta_p2p_1peer_GBps;
  if (peers >= m.fanin_knee) return m.beta_p2p_GBps;
  double t = double(peers - 1) / double(m.fanin_knee - 1);
  return m.beta_p2p_1peer_GBps + t * (m.beta_p2p_GBps - m.beta_p2p_1peer_GBps);
}
// A transfer of `bytes` at peak `GBps` with a half-bandwidth point (latency-bound ramp).
double xfer_us(double bytes, double GBps, double half_bytes) {
  if (bytes <= 0) return 0;
  double eff = GBps * bytes / (bytes + half_bytes);
  return bytes / (eff * 1e3);  // GB/s = 1e3 bytes/us
}
}  // namespace

double predict_us(const LinkModel& m, int algo, const Widths& widths, int n, double S,
                  bool registered) {
  if (n <= 1) return m.alpha_launch_us + S / (m.gamma_hbm_GBps * 1e3);
  const double N = n;
  auto hop = [&](int w) { return m.alpha_sync_us + m.alpha_peer_us * (w - 1); };
  switch (algo) {
    case ALGO_ONESHOT: {
      // every rank pushes its whole message to N-1 peers as LL lines, then polls and reduces them
      return m.alpha_launch_us + m.alpha_ll_us + S / (m.beta_ll_self_GBps * 1e3) +
             (N - 1) * S / (m.beta_ll_GBps * 1e3);
    }
    case ALGO_TWOSHOT: {
      // shot 1: block p of my message to rank p; reduce my block; shot 2: my reduced block to everybody. Two LL
      // hops of latency, 2 (N-1)/N S of payload pushed per rank instead of the one-shot's (N-1) S.
      return m.alpha_launch_us + 2.0 * m.alpha_ll_us + S / (m.beta_ll_self_GBps * 1e3) +
             2.0 * (N - 1) / N * S / (m.beta_ll_GBps * 1e3);
    }
    case ALGO_STAGED: {
      double t = m.alpha_launch_us;
      // private buffers: an extra publish phase + local copies (measured: ~22 us + S / 4.5 TB/s)
      if (!registered) t += 4.0 * hop(widths.empty() ? n : widths[0]) + S / (0.75 * m.gamma_hbm_GBps * 1e3);
      t += hop(widths.empty() ? n : widths[0]);  // input-ready hop
      double g = 1;
      for (int w : widths) {
        double gn = g * w;
        double stage_bytes = (w - 1) / gn * S;  // inbound per rank, same for RS and AG
        double b = beta_fanin(m, w - 1);
        t += 2.0 * (hop(w) + xfer_us(stage_bytes, b, m.half_bw_bytes));
        g = gn;
      }
      if (registered) t += hop(widths.empty() ? n : widths[0]);  // exit barrier
      return t;
    }
    case ALGO_NVLS: {
      if (!m.has_nvls) return 1e30;
      // three all-to-all flag hops (input ready, reduced, exit) + a constant for the multicast stores in flight
      double t = m.alpha_launch_us + 3.0 * hop(n) + m.alpha_nvls_us;
      // private buffers: copy-in and copy-out roles + two more flag hops per tile
      // (round-2 fit on 4 and 8 GPUs, 16 - 256 MiB: +21..28 us at 16 MiB, +48..92 us at 64 MiB, +83..127 us at 256 MiB over
      // the zero-copy kernel - the copy passes run at about half the HBM copy rate and head / tail of the tile pipeline
      // are not hidden)
      if (!registered) t += 4.0 * hop(n) + 12.0 + S / (0.5 * m.gamma_hbm_GBps * 1e3);
      // per-GPU link traffic (each direction): every rank's copy of all N blocks goes up to the switch once
      // (S, including the requester's own copy) plus its reduced block broadcast (S/N)  ->  (1 + 1/N) S,
      // against 2 (N-1)/N S for the P2P stages: NVLS loses at N = 2 and wins from N = 4 up.
      t += S * (1.0 + 1.0 / N) / (m.beta_nvls_GBps * 1e3);
      return t;
    }
    case ALGO_RING: {
      double step = hop(2) + xfer_us(S / N, beta_fanin(m, 1), m.half_bw_bytes / 4);
      return m.alpha_launch_us + hop(2) + 2 * (N - 1) * step;
    }
  }
  return 1e30;
}

Choice choose(const LinkModel& m, int n, double S, bool registered, int max_width) {
  Choice best;
  best.est_us = 1e30;
  if (n <= 1) {
    best.algo = ALGO_STAGED;
    best.widths = {1};
    best.est_us = predict_us(m, ALGO_STAGED, best.widths, n, S, registered);
    return best;
  }
  auto consider = [&](int algo, const Widths& w) {
    double t = predict_us(m, algo, w, n, S, registered);
    if (t < best.est_us) {
      best.est_us = t;
      best.algo = algo;
      best.widths = w;
    }
  };
  if (S <= m.max_oneshot_bytes) consider(ALGO_ONESHOT, {n});
  if (S <= m.max_twoshot_bytes) consider(ALGO_TWOSHOT, {n});
  for (auto& w : enumerate_widths(n, false)) {
    bool ok = true;
    for (int x : w) ok = ok && x <= max_width;
    if (ok) consider(ALGO_STAGED, w);
  }
  consider(ALGO_NVLS, {n});
  // The ring is kept for parity with the reference (FT_TOPO containing a 1) and as the fallback for world
  // sizes with no usable factorization; on an NVSwitch fabric it never beats the direct exchanges, so it
  // is not an automatic candidate.
  if (best.est_us >= 1e30) {  // e.g. prime n above max_width: fall back to ring
    best.algo = ALGO_RING;
    best.widths = {1};
    best.est_us = predict_us(m, ALGO_RING, best.widths, n, S, registered);
  }
  return best;
}

}  // namespace ft
