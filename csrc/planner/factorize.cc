#include "factorize.h"

#include <cctype>
#include <unordered_map>

namespace ft {

namespace {
uint64_t count_memo(uint64_t n, std::unordered_map<uint64_t, uint64_t>& memo) {
  if (n == 0) return 0;
  if (n == 1) return 1;
  auto it = memo.find(n);
  if (it != memo.end()) return it->second;
  // Sum over divisors d >= 2 of H(n/d); walk divisor pairs (d, n/d) up to sqrt(n).
  uint64_t total = 1;  // d == n contributes H(1) = 1
  for (uint64_t d = 2; d * d <= n; ++d) {
    if (n % d) continue;
    uint64_t q = n / d;
    total += count_memo(q, memo);
    if (q != d) total += count_memo(d, memo);
  }
  memo.emplace(n, total);
  return total;
}

struct Dfs {
  bool (*cb)(const int*, int, void*);
  void* user;
  std::vector<int> cur;
  bool stop = false;
  void run(int rest) {
    if (stop) return;
    if (rest == 1) {
      if (!cb(cur.data(), (int)cur.size(), user)) stop = true;
      return;
    }
    for (int d = 2; d <= rest && !stop; ++d) {
      if (rest % d) continue;
      cur.push_back(d);
      run(rest / d);
      cur.pop_back();
    }
  }
};
}  // namespace

uint64_t count_factorizations(uint64_t n) {
  std::unordered_map<uint64_t, uint64_t> memo;
  return count_memo(n, memo);
}

void for_each_widths(int n, bool (*cb)(const int*, int, void*), void* user) {
  if (n <= 0) return;
  Dfs d{cb, user, {}, false};
  d.run(n);
}

std::vector<Widths> enumerate_widths(int n, bool legacy_flat) {
  std::vector<Widths> out;
  if (n <= 0) return out;
  struct Ctx {
    std::vector<Widths>* out;
    bool legacy;
  } ctx{&out, legacy_flat};
  for_each_widths(
      n,
      [](const int* w, int len, void* u) {
        auto* c = static_cast<Ctx*>(u);
        if (c->legacy && len == 1) {
          c->out->push_back({1, w[0]});
          c->out->push_back({w[0], 1});
        } else {
          c->out->emplace_back(w, w + len);
        }
        return true;
      },
      &ctx);
  return out;
}

std::vector<int> prime_factors(int n) {
  std::vector<int> f;
  if (n < 2) return f;
  for (int p = 2; (long long)p * p <= n; ++p)
    while (n % p == 0) {
      f.push_back(p);
      n /= p;
    }
  if (n > 1) f.push_back(n);
  return f;
}

bool is_prime(int n) {
  if (n < 2) return false;
  for (int p = 2; (long long)p * p <= n; ++p)
    if (n % p == 0) return false;
  return true;
}

bool is_prime_legacy(int n) { return n <= 0 ? false : (n < 4 ? true : is_prime(n)); }

std::string to_string(const Widths& w, int lonely) {
  std::string s;
  for (size_t i = 0; i < w.size(); ++i) {
    if (i) s += '*';
    s += std::to_string(w[i]);
  }
  if (lonely > 0) s += "+" + std::to_string(lonely);
  if (lonely < 0) s += std::to_string(lonely);
  return s;
}

std::vector<Candidate> candidates(int n, bool legacy_flat) {
  std::vector<Candidate> out;
  auto add = [&](int m, int lonely) {
    for (auto& w : enumerate_widths(m, legacy_flat)) out.push_back({w, lonely});
  };
  if (!is_prime_legacy(n)) add(n, 0);
  add(n + 1, -1);
  add(n - 1, +1);
  return out;
}

bool parse_widths(const std::string& s, Widths* out) {
  out->clear();
  long long cur = -1;
  for (char ch : s) {
    if (std::isdigit((unsigned char)ch)) {
      cur = (cur < 0 ? 0 : cur) * 10 + (ch - '0');
      if (cur > 1000000000LL) return false;
    } else if (ch == ',' || ch == ' ' || ch == '\t' || ch == '*' || ch == 'x' || ch == 'X' || ch == '-') {
      if (cur >= 0) out->push_back((int)cur);
      cur = -1;
    } else {
      return false;
    }
  }
  if (cur >= 0) out->push_back((int)cur);
  return !out->empty();
}

long long product(const Widths& w) {
  long long p = 1;
  for (int x : w) p *= x;
  return p;
}

}  // namespace ft
