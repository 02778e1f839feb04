// ft_plan - command-line face of the planner.
//
// Capability parity:
//   cost_model/main.cpp:18-26   sweep N = 1..999: "count,microseconds" of enumerate + cost  -> --sweep [--out f.csv]
//   cost_model/ChooseWidth.h    candidate trees of N, N+1 ("-1"), N-1 ("+1")               -> --candidates N
//   cost_model/CostModel.h      per-term cost of every tree + the optimum                   -> --legacy N [--chunk C]
//   mpi_mod.hpp:107-144         print_ops schedule dump                                     -> --print-ops --widths 2,4 --rank r
//   topo_count/factor_count.py  H(n)                                                        -> --count N
// plus the B200 link model:  --choose N --bytes S [--nvls] [--model file.json]
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <string>

#include "../../include/flextree.h"
#include "../planner/cost_model.h"
#include "../planner/factorize.h"
#include "../planner/schedule.h"
#include "../core/version.h"

using namespace ft;

static void usage() {
  fprintf(stderr,
          "ft_plan --count N | --enumerate N [--legacy-flat] | --candidates N | --legacy N [--chunk C]\n"
          "        | --print-ops --widths a,b,c --rank r [--lonely k] | --sweep [--max N] [--chunk C] [--out file]\n"
          "        | --choose N --bytes S [--nvls] [--registered] [--model file.json] | --topo TEXT --ranks N | --version\n");
}

int main(int argc, char** argv) {
  std::string mode, widths_s, out, model_path;
  int n = 0, rank = 0, lonely = 0, maxn = 999;
  double chunk = 100, bytes = 1 << 20;
  bool legacy_flat = false, nvls = false, registered = false;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto val = [&]() { if (i + 1 >= argc) { usage(); exit(1); } return argv[++i]; };
    if (a == "--count" || a == "--enumerate" || a == "--candidates" || a == "--legacy" || a == "--choose") { mode = a; n = atoi(val()); }
    else if (a == "--print-ops" || a == "--sweep" || a == "--version") mode = a;
    else if (a == "--widths") widths_s = val();
    else if (a == "--topo") { mode = a; widths_s = val(); }
    else if (a == "--ranks") n = atoi(val());
    else if (a == "--rank") rank = atoi(val());
    else if (a == "--lonely") lonely = atoi(val());
    else if (a == "--chunk") chunk = atof(val());
    else if (a == "--bytes") bytes = atof(val());
    else if (a == "--max") maxn = atoi(val());
    else if (a == "--out") out = val();
    else if (a == "--model") model_path = val();
    else if (a == "--legacy-flat") legacy_flat = true;
    else if (a == "--nvls") nvls = true;
    else if (a == "--registered") registered = true;
    else { usage(); return 1; }
  }
  if (mode == "--version") { printf("%s\n", FT_VERSION_STRING); return 0; }
  if (mode == "--topo") {  // how FT_TOPO=TEXT is read on N ranks (incl. the '+k' lonely / '-k' phantom suffixes)
    int w[64], lone = 0, ph = 0;
    const int k = ft_parse_topo_ex(widths_s.c_str(), n, w, 64, &lone, &ph);
    if (k < 0) { fprintf(stderr, "%s\n", ft_last_error()); return 1; }
    Widths ws(w, w + k);
    if (k == 1 && w[0] == 1 && n > 1) printf("ring over %d ranks\n", n);
    else printf("tree %s: %d positions on %d ranks, %d lonely, %d phantom\n", to_string(ws, lone ? lone : -ph).c_str(), n - lone + ph, n, lone, ph);
    return 0;
  }
  if (mode == "--count") { printf("%llu\n", (unsigned long long)count_factorizations((uint64_t)n)); return 0; }
  if (mode == "--enumerate") {
    for (auto& w : enumerate_widths(n, legacy_flat)) printf("%s\n", to_string(w).c_str());
    return 0;
  }
  if (mode == "--candidates") {
    for (auto& c : candidates(n, true)) printf("%s\n", to_string(c.widths, c.lonely).c_str());
    return 0;
  }
  if (mode == "--legacy") {
    auto trees = enumerate_widths(n, true);
    for (auto& w : trees) {
      LegacyTerms t = legacy_cost(w, n, chunk);
      printf("%-16s latency %.6g  memory %.6g  bandwidth %.6g  total %.6g\n", to_string(w).c_str(), t.latency, t.memory,
             t.bandwidth, t.total());
    }
    double c = 0;
    int i = legacy_choose(trees, n, chunk, &c);
    if (i >= 0)
      printf("the optimized tree structure for %d total nodes should be: %s\nthe optimized all_reduce time for %d total nodes should be: %g\n",
             n, to_string(trees[(size_t)i]).c_str(), n, c);
    return 0;
  }
  if (mode == "--print-ops") {
    Widths w;
    Plan p;
    if (!parse_widths(widths_s, &w) || !make_plan(w.data(), (int)w.size(), &p)) { fprintf(stderr, "bad --widths\n"); return 1; }
    if (rank < 0 || rank >= p.nranks + lonely) { fprintf(stderr, "bad --rank\n"); return 1; }
    printf("%s", print_ops(build_schedule(p, rank, lonely)).c_str());
    return 0;
  }
  if (mode == "--sweep") {
    FILE* f = out.empty() ? stdout : fopen(out.c_str(), "w");
    if (!f) { perror("fopen"); return 1; }
    fprintf(f, "n,structures,microseconds,best,cost\n");
    for (int i = 1; i <= maxn; ++i) {
      auto t0 = std::chrono::steady_clock::now();
      auto trees = enumerate_widths(i, true);
      double c = 0;
      int b = legacy_choose(trees, i, chunk, &c);
      auto t1 = std::chrono::steady_clock::now();
      double us = std::chrono::duration<double, std::micro>(t1 - t0).count();
      fprintf(f, "%d,%zu,%.1f,%s,%.6g\n", i, trees.size(), us, b >= 0 ? to_string(trees[(size_t)b]).c_str() : "", b >= 0 ? c : 0.0);
    }
    if (f != stdout) fclose(f);
    return 0;
  }
  if (mode == "--choose") {
    LinkModel m;
    if (!model_path.empty()) {
      std::string err;
      if (!load_link_model(model_path, &m, &err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    }
    if (nvls) m.has_nvls = true;
    printf("model: %s\n", dump_link_model(m).c_str());
    auto show = [&](int algo, const Widths& w) {
      printf("  %-8s %-12s %10.2f us\n", algo_name(algo), to_string(w).c_str(), predict_us(m, algo, w, n, bytes, registered));
    };
    show(ALGO_ONESHOT, {n});
    for (auto& w : enumerate_widths(n, false)) show(ALGO_STAGED, w);
    if (m.has_nvls) show(ALGO_NVLS, {n});
    show(ALGO_RING, {1});
    Choice c = choose(m, n, bytes, registered);
    printf("choice for N=%d, %.0f bytes: %s %s (%.2f us)\n", n, bytes, algo_name(c.algo), to_string(c.widths).c_str(), c.est_us);
    return 0;
  }
  usage();
  return 1;
}
