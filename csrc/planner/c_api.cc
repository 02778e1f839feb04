// extern "C" surface of the planner (declared in include/flextree.h). Pure CPU.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>

#include "../../include/flextree.h"
#include "../core/version.h"
#include "cost_model.h"
#include "factorize.h"
#include "schedule.h"

using namespace ft;

namespace ft {
thread_local std::string g_last_error;
void set_last_error(const std::string& s) { g_last_error = s; }
}  // namespace ft

struct ft_link_model { LinkModel m; };

static long long copy_out(const std::string& s, char* out, long long cap) {
  if (out && cap > 0) {
    long long n = (long long)s.size() < cap - 1 ? (long long)s.size() : cap - 1;
    memcpy(out, s.data(), (size_t)n);
    out[n] = 0;
  }
  return (long long)s.size();
}

extern "C" {

const char* ft_last_error(void) { return g_last_error.c_str(); }
const char* ft_version_string(void) { return FT_VERSION_STRING; }
// Probe that the library is the one in use (parity: FT_enabled(), mpi_mod.hpp:8-12).
int ft_enabled(void) {
  fputs("FlexTree enabled", stdout);
  return 0;
}

uint64_t ft_count_factorizations(uint64_t n) { return count_factorizations(n); }

long long ft_enumerate_widths(int n, int legacy_flat, int* out, long long cap, long long* n_lists) {
  auto lists = enumerate_widths(n, legacy_flat != 0);
  long long need = 0;
  for (auto& w : lists) need += 1 + (long long)w.size();
  if (n_lists) *n_lists = (long long)lists.size();
  if (need > cap || !out) return -need;
  long long i = 0;
  for (auto& w : lists) {
    out[i++] = (int)w.size();
    for (int x : w) out[i++] = x;
  }
  return need;
}

long long ft_candidates_text(int n, char* out, long long cap) {
  std::string s;
  for (auto& c : candidates(n, true)) s += to_string(c.widths, c.lonely) + "\n";
  return copy_out(s, out, cap);
}

int ft_prime_factors(int n, int* out, int cap) {
  auto f = prime_factors(n);
  for (int i = 0; i < (int)f.size() && i < cap; ++i) out[i] = f[i];
  return (int)f.size();
}
int ft_is_prime(int n) { return is_prime(n); }
int ft_is_prime_legacy(int n) { return is_prime_legacy(n); }

double ft_legacy_cost(const int* widths, int nstages, int total_nodes, double chunk, double* terms) {
  LegacyTerms t = legacy_cost(Widths(widths, widths + nstages), total_nodes, chunk);
  if (terms) { terms[0] = t.latency; terms[1] = t.memory; terms[2] = t.bandwidth; }
  return t.total();
}

int ft_legacy_choose(int n, double chunk, int* widths_out, int cap, double* cost_out) {
  auto lists = enumerate_widths(n, true);
  double c = 0;
  int i = legacy_choose(lists, n, chunk, &c);
  if (i < 0) return 0;
  if (cost_out) *cost_out = c;
  int len = (int)lists[i].size();
  for (int k = 0; k < len && k < cap; ++k) widths_out[k] = lists[i][k];
  return len;
}

int ft_plan_traffic(const int* widths, int nstages, int rank, long long count, long long* messages,
                    long long* elements) {
  Plan p;
  if (!make_plan(widths, nstages, &p) || rank < 0 || rank >= p.nranks) return FT_ERR_INVALID;
  Traffic t = traffic(build_schedule(p, rank, 0), count);
  if (messages) *messages = t.messages;
  if (elements) *elements = t.elements;
  return FT_SUCCESS;
}

long long ft_print_ops(const int* widths, int nstages, int rank, int lonely, char* out, long long cap) {
  Plan p;
  if (!make_plan(widths, nstages, &p) || rank < 0 || rank >= p.nranks + lonely) return -1;
  return copy_out(print_ops(build_schedule(p, rank, lonely)), out, cap);
}

ft_link_model_t* ft_link_model_create(const char* json_path) {
  auto* m = new ft_link_model;
  const char* path = json_path && *json_path ? json_path : getenv("FT_COSTMODEL");
  if (path && *path) {
    std::string err;
    if (!load_link_model(path, &m->m, &err)) {
      set_last_error(err);
      if (json_path && *json_path) { delete m; return nullptr; }
    }
  }
  return m;
}
void ft_link_model_destroy(ft_link_model_t* m) { delete m; }
long long ft_link_model_dump(const ft_link_model_t* m, char* out, long long cap) {
  return copy_out(dump_link_model(m->m), out, cap);
}
int ft_link_model_set(ft_link_model_t* m, const char* key, double v) {
  std::string k = key;
#define FT_KEY(name, cast) if (k == #name) { m->m.name = (cast)v; return FT_SUCCESS; }
  FT_KEY(alpha_launch_us, double) FT_KEY(alpha_sync_us, double) FT_KEY(alpha_ll_us, double)
  FT_KEY(alpha_peer_us, double) FT_KEY(alpha_nvls_us, double) FT_KEY(beta_ll_self_GBps, double)
  FT_KEY(beta_p2p_GBps, double) FT_KEY(beta_p2p_1peer_GBps, double) FT_KEY(fanin_knee, int)
  FT_KEY(beta_nvls_GBps, double) FT_KEY(beta_ll_GBps, double) FT_KEY(gamma_hbm_GBps, double)
  FT_KEY(half_bw_bytes, double) FT_KEY(has_nvls, bool) FT_KEY(max_oneshot_bytes, int)
#undef FT_KEY
  return FT_ERR_INVALID;
}
double ft_predict_us(const ft_link_model_t* m, int algo, const int* widths, int nstages, int nranks,
                     double bytes, int registered) {
  return predict_us(m->m, algo, Widths(widths, widths + nstages), nranks, bytes, registered != 0);
}
int ft_choose(const ft_link_model_t* m, int nranks, double bytes, int registered, int* algo_out,
              int* widths_out, int cap, double* est_us) {
  Choice c = choose(m->m, nranks, bytes, registered != 0);
  if (algo_out) *algo_out = c.algo;
  if (est_us) *est_us = c.est_us;
  for (int k = 0; k < (int)c.widths.size() && k < cap; ++k) widths_out[k] = c.widths[k];
  return (int)c.widths.size();
}

int ft_parse_topo(const char* text, int nranks, int* widths_out, int cap) {
  std::string s = text ? text : (getenv("FT_TOPO") ? getenv("FT_TOPO") : "");
  Widths w;
  bool blank = true;
  for (char ch : s) if (ch != ' ' && ch != '\t') blank = false;
  if (blank) {
    w = {nranks};
  } else {
    if (!parse_widths(s, &w)) { set_last_error("invalid FT_TOPO " + s); return -1; }
    bool ring = false;
    for (int x : w) if (x == 1) ring = true;
    if (ring) w = {1};
    else if (product(w) != nranks) { set_last_error("invalid FT_TOPO " + s); return -1; }
  }
  for (int k = 0; k < (int)w.size() && k < cap; ++k) widths_out[k] = w[k];
  return (int)w.size();
}

// FT_TOPO with the tree-size suffixes: "2,3+1" = a tree of 6 plus one lonely rank, "2,4-1" = a tree of 8 positions of
// which one is a phantom (the notation of the reference's tree listings, cost_model/PrintTreeStructure.h:20-53; its
// own parser, mpi_mod.hpp:882-929, has no syntax for them). lonely_out / phantoms_out receive k (0 for a ring).
int ft_parse_topo_ex(const char* text, int nranks, int* widths_out, int cap, int* lonely_out, int* phantoms_out) {
  std::string s = text ? text : (getenv("FT_TOPO") ? getenv("FT_TOPO") : "");
  int lonely = 0, phantoms = 0;
  const size_t plus = s.find('+'), minus = s.find('-');
  if (plus != std::string::npos) {
    lonely = atoi(s.c_str() + plus + 1);
    s = s.substr(0, plus);
    if (lonely <= 0) { set_last_error("invalid FT_TOPO: '+k' needs k >= 1"); return -1; }
  } else if (minus != std::string::npos) {
    phantoms = atoi(s.c_str() + minus + 1);
    s = s.substr(0, minus);
    if (phantoms <= 0) { set_last_error("invalid FT_TOPO: '-k' needs k >= 1"); return -1; }
  }
  const int tree = nranks - lonely + phantoms;
  if (tree < 1) { set_last_error("invalid FT_TOPO: more lonely ranks than ranks"); return -1; }
  const int k = ft_parse_topo(s.c_str(), tree, widths_out, cap);
  if (k < 0) return -1;
  const bool ring = k == 1 && widths_out[0] == 1 && nranks > 1;
  if (lonely_out) *lonely_out = ring ? 0 : lonely;
  if (phantoms_out) *phantoms_out = ring ? 0 : phantoms;
  return k;
}

}  // extern "C"
