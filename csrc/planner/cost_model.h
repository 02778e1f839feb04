// Cost models that rank stage-width factorizations (and, for the GPU model, algorithms).
//
// 1. LegacyCost: the reference's hand-calibrated formula, kept as an oracle so that the planner can
//    reproduce the reference's picks (cost_model/CostModel.h:1-120). Behaviour re-implemented:
//      latency  = sum over layers of 2*lo (+ 100*(w-9)*co if w > 9)      (CostModel.h:1-20, :92 -
//                 the reference passes a literal 100 instead of the chunk size; reproduced)
//      bandwidth= (N-1)/N * chunk * bo                                    (CostModel.h:22-30)
//      memory   = (N + 1 + 2*sum_{k=1}^{h-1} prod_{j<k} tree[j]) * chunk/N * o   (CostModel.h:32-79)
//    The reference has undefined behaviour for tree heights 0 and >9 and an uninitialised
//    accumulator; this version is total (any height) and initialised.
//
// 2. LinkModel: an alpha-beta-gamma model of one NVSwitch box, re-fit from measurements taken on the
//    box (bench/fit_cost_model.py writes the JSON that load_link_model reads). It predicts the time of
//    every algorithm this library implements and drives FT_ALGO=auto.
#pragma once
#include <string>
#include <vector>

#include "factorize.h"

namespace ft {

struct LegacyTerms {
  double latency = 0, memory = 0, bandwidth = 0;
  double total() const { return latency + memory + bandwidth; }
};

struct LegacyCoeffs {
  double lo = 0.004;         // per-layer latency (x2)
  double co = 0.0002;        // incast/control penalty per chunk-unit per width above 9
  double bo = 0.0068;        // bandwidth + compute
  double o = 0.0004;         // memory read/write
  double latency_chunk = 100;  // the reference hard-codes 100 here (CostModel.h:92)
  int incast_knee = 9;
};

LegacyTerms legacy_cost(const Widths& tree, int total_nodes, double chunk,
                        const LegacyCoeffs& c = LegacyCoeffs());

// arg-min over candidates (first minimum wins, like the reference's strict '<').
// Returns the index into `trees`, or -1 if empty.
int legacy_choose(const std::vector<Widths>& trees, int total_nodes, double chunk,
                  double* best_cost = nullptr, const LegacyCoeffs& c = LegacyCoeffs());

// ---------------------------------------------------------------------------------------------
enum Algo { ALGO_AUTO = 0, ALGO_ONESHOT = 1, ALGO_STAGED = 2, ALGO_NVLS = 3, ALGO_RING = 4 };
// internal kernel selector (not a user-visible algorithm): the P2P stages with TMA-staged tiles
enum { ALGO_STAGED_TMA = 5, ALGO_TWOSHOT = 6, ALGO_STAGED_CLUSTER = 7 };  // 7: TMA stages, two-CTA cluster per tile (internal)  // 6: two-shot LL (user-visible as FT_ALGO_TWOSHOT / "twoshot")
const char* algo_name(int a);
int algo_from_name(const std::string& s);  // -1 if unknown

struct LinkModel {
  // Defaults are the round-1 fit on 8xB200 / 2xB200 (profiles/link_model_b200.json, docs/TUNING.md).
  // latencies, microseconds
  double alpha_launch_us = 8.35;  // launch + first instruction + exit, back-to-back calls
  double alpha_sync_us = 1.16;    // one release->acquire flag hop between two GPUs
  double alpha_peer_us = 0.087;   // extra per additional peer signalled/polled in the same hop
  double alpha_ll_us = 3.63;      // one LL (data+flag) hop
  double alpha_nvls_us = 5.3;     // extra fixed cost of the NVLS path on top of its three flag hops
  // bandwidths, GB/s
  // TMA-staged pull bandwidth. Round-2 fit (profiles/cost_model_validation_r2.md, 1 GiB on 8 GPUs): the flat tree pulls from 7
  // peers at once and reaches 572 GB/s, 2x2x2 pulls from one peer per stage and reaches 633 GB/s (2 GPUs: 627, 4 GPUs: 629-637)
  double beta_p2p_GBps = 575.0;        // ... with >= fanin_knee concurrent peers
  double beta_p2p_1peer_GBps = 635.0;  // ... from a single peer
  int fanin_knee = 7;
  double beta_nvls_GBps = 539.0;  // link rate of multimem ld_reduce/st: bytes = (1 + 1/N) * S
  double beta_ll_GBps = 400.0;    // LL: rate at which the N-1 peer copies are pushed/polled (two-pass kernel) ...
  double beta_ll_self_GBps = 2000.0;  // ... plus a per-message term independent of N (negligible in the fit)
  double gamma_hbm_GBps = 6000.0; // local HBM copy/reduce traffic rate
  double half_bw_bytes = 4.0e6;   // per-stage bytes at which a P2P stage reaches half its peak rate
  bool has_nvls = false;
  int max_oneshot_bytes = 2 << 20;  // further limited by the LL buffer; at 4 MiB the two-shot wins even on 2 GPUs (22 vs 29 us)
  // two-shot LL as an automatic candidate up to this size. Measured on 8 GPUs with private buffers (profiles/
  // sweep_n8_staging_mid): 17-25 us from 64 KiB to 2 MiB and 36 us at 4 MiB, where the copy-in NVLS path takes 52-62 us
  // and NCCL 29-87 us; beyond 4 MiB the doubled wire bytes of LL lines lose to NVLS.
  int max_twoshot_bytes = 4 << 20;
};

bool load_link_model(const std::string& json_path, LinkModel* m, std::string* err = nullptr);
std::string dump_link_model(const LinkModel& m);

struct Choice {
  int algo = ALGO_STAGED;
  Widths widths;     // for ALGO_STAGED
  double est_us = 0;
};

// Predicted time of one algorithm. `registered` = zero-copy path (no copy-in pass).
double predict_us(const LinkModel& m, int algo, const Widths& widths, int nranks, double bytes,
                  bool registered);

// Best (algo, widths) for a message. Considers every ordered factorization with width <= max_width.
Choice choose(const LinkModel& m, int nranks, double bytes, bool registered, int max_width = 32);

}  // namespace ft
