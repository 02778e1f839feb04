// Ordered-factorization enumeration, counting, prime utilities and tree printing.
//
// Capability parity (behaviour only, re-implemented):
//   cost_model/GetWidth.h:10-47        getWidth      -> enumerate_widths(n, /*legacy_flat=*/true)
//   cost_model/GetWidth.h:51-227       getWidth2     -> same set (see tests), no 9-stage limit here
//   cost_model/GetPrimeFactor.h:5-19   getPrimeFactor-> prime_factors
//   cost_model/IsPrimeNumber.h:4-21    isPrimeNumber -> is_prime / is_prime_legacy
//   cost_model/PrintTreeStructure.h    printTree*    -> to_string(widths, lonely)
//   cost_model/ChooseWidth.h:8-38      chooseWidth   -> candidates(n)
//   topo_count/factor_count.py:1-11    H(n)          -> count_factorizations
//
// Unlike the reference enumerator (file-scope globals, not re-entrant) everything here is
// re-entrant and allocation is bounded by the output size.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace ft {

using Widths = std::vector<int>;

// H(n): number of ordered factorizations of n into factors >= 2. H(0)=0, H(1)=1.
uint64_t count_factorizations(uint64_t n);

// All ordered factorizations of n with every factor >= 2, in depth-first order with the
// leading factor ascending (so for 8: 2*2*2, 2*4, 4*2, 8).
//   legacy_flat=false : the single-stage tree is emitted as {n}.
//   legacy_flat=true  : reference listing convention - {n} is replaced by the pair {1,n},{n,1}
//                       (so the list has H(n)+1 entries for n >= 2).
// n == 1 -> { {} } ; n <= 0 -> {}.
std::vector<Widths> enumerate_widths(int n, bool legacy_flat = false);

// Visit factorizations without materialising the list (used by the planner for large n).
// The callback may return false to stop early.
void for_each_widths(int n, bool (*cb)(const int* w, int len, void* user), void* user);

std::vector<int> prime_factors(int n);  // ascending, with multiplicity; n<2 -> {}
bool is_prime(int n);                    // mathematically correct (1 is not prime)
bool is_prime_legacy(int n);             // reference quirk: 1 reports prime (n<=0 does not)

// Lonely-rank annotation of a candidate tree for a world of N ranks:
//   lonely = 0  : tree covers exactly N ranks
//   lonely = +k : tree covers N-k ranks, k ranks are "lonely" (reference prints "+1")
//   lonely = -k : tree covers N+k slots, k slots are phantom (reference prints "-1")
struct Candidate {
  Widths widths;
  int lonely = 0;
};
std::string to_string(const Widths& w, int lonely = 0);

// Reference chooseWidth(N) listing: trees of N (unless N is "legacy prime"), then trees of N+1
// tagged -1, then trees of N-1 tagged +1.
std::vector<Candidate> candidates(int n, bool legacy_flat = true);

// Parse "2,3" / "2 3" / "2*3" / "2x3" into widths. Returns false on garbage.
bool parse_widths(const std::string& s, Widths* out);

long long product(const Widths& w);

}  // namespace ft
