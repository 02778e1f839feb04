// ThreadSanitizer driver for the CPU world: runs the kernel algorithm templates with one thread per
// "CTA" under TSAN to validate the hazard analysis (data accesses are plain, flags are acq/rel).
// Build+run: make tsan   (see Makefile)
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

extern "C" {
void* ft_host_world_create(int, size_t, size_t);
void ft_host_world_destroy(void*);
void* ft_host_registered(void*, int);
long long ft_host_allreduce(void*, const void* const*, void* const*, long long, int, int, int, int,
                            const int*, int, int, long long, int, double, int, unsigned long long, int, int, int);
}

static int run(int n, std::vector<int> widths, int algo, int mode, long long count, int ctas, long long tile, int lonely = 0) {
  void* w = ft_host_world_create(n, 1 << 18, 1 << 18);
  std::vector<std::vector<float>> in(n, std::vector<float>(count + 8)), out(n, std::vector<float>(count + 8));
  std::vector<const void*> s(n);
  std::vector<void*> r(n);
  for (int i = 0; i < n; ++i) {
    float* base = mode ? (float*)ft_host_registered(w, i) : (float*)(((uintptr_t)in[i].data() + 15) & ~(uintptr_t)15);
    for (long long j = 0; j < count; ++j) base[j] = (float)(i + 1) * (float)(j % 7);
    s[i] = base;
    r[i] = mode ? (void*)base : (void*)(((uintptr_t)out[i].data() + 15) & ~(uintptr_t)15);
  }
  int bad = 0;
  for (int rep = 0; rep < 3; ++rep) {
    if (mode) for (int i = 0; i < n; ++i) for (long long j = 0; j < count; ++j) ((float*)s[i])[j] = (float)(i + 1) * (float)(j % 7);
    long long rc = ft_host_allreduce(w, s.data(), r.data(), count, 6, 6, 0, algo, widths.data(), (int)widths.size(),
                                     ctas, tile, mode, 1.0, 0, 20000000000ull, -1, lonely, 0);
    if (rc) { printf("rc=%lld\n", rc); bad = 1; }
    float tot = (float)n * (n + 1) / 2;
    for (int i = 0; i < n; ++i)
      for (long long j = 0; j < count; ++j)
        if (((float*)r[i])[j] != tot * (float)(j % 7)) { bad = 1; }
  }
  ft_host_world_destroy(w);
  printf("n=%d stages=%zu algo=%d mode=%d count=%lld: %s\n", n, widths.size(), algo, mode, count, bad ? "FAIL" : "ok");
  return bad;
}

// The two halves as collectives of their own (phases 1 and 2), fp32, checked against each other.
static int run_halves(int n, std::vector<int> widths, long long per_rank, int ctas, long long tile, int algo) {
  void* w = ft_host_world_create(n, 1 << 18, 1 << 18);
  const long long total = per_rank * n;
  std::vector<std::vector<float>> in(n, std::vector<float>(total + 8)), part(n, std::vector<float>(per_rank + 8)),
      full(n, std::vector<float>(total + 8));
  auto al = [](std::vector<float>& v) { return (float*)(((uintptr_t)v.data() + 15) & ~(uintptr_t)15); };
  std::vector<const void*> s(n);
  std::vector<void*> r(n);
  for (int i = 0; i < n; ++i) {
    for (long long j = 0; j < total; ++j) al(in[i])[j] = (float)(i + 1) * (float)(j % 5);
    s[i] = al(in[i]);
    r[i] = al(part[i]);
  }
  int bad = 0;
  long long rc = ft_host_allreduce(w, s.data(), r.data(), total, 6, 6, 0, algo, widths.data(), (int)widths.size(), ctas, tile, 0,
                                   1.0, 0, 20000000000ull, -1, 0, /*phases=*/1);
  if (rc) { printf("rs rc=%lld\n", rc); bad = 1; }
  for (int i = 0; i < n; ++i) {
    s[i] = al(part[i]);
    r[i] = al(full[i]);
  }
  rc = ft_host_allreduce(w, s.data(), r.data(), total, 6, 6, 0, algo, widths.data(), (int)widths.size(), ctas, tile, 0, 1.0, 0,
                         20000000000ull, -1, 0, /*phases=*/2);
  if (rc) { printf("ag rc=%lld\n", rc); bad = 1; }
  const float tot = (float)n * (n + 1) / 2;
  for (int i = 0; i < n; ++i)
    for (long long j = 0; j < total; ++j)
      if (al(full[i])[j] != tot * (float)(j % 5)) bad = 1;
  ft_host_world_destroy(w);
  printf("halves n=%d stages=%zu algo=%d per_rank=%lld: %s\n", n, widths.size(), algo, per_rank, bad ? "FAIL" : "ok");
  return bad;
}

// fp32 data with a narrower wire (13 = bf16, 14 = block-scaled fp8) and the fused 1/N scale; small integers, so
// that the expected value is exact in both wire formats.
static int run_wire(int n, std::vector<int> widths, int wire, long long count, int ctas, long long tile) {
  void* w = ft_host_world_create(n, 1 << 18, 1 << 18);
  std::vector<std::vector<float>> buf(n, std::vector<float>(count + 8));
  std::vector<const void*> s(n);
  std::vector<void*> r(n);
  for (int i = 0; i < n; ++i) {
    float* base = (float*)(((uintptr_t)buf[i].data() + 15) & ~(uintptr_t)15);
    for (long long j = 0; j < count; ++j) base[j] = (float)((j % 3) * n);  // sum = (j%3)*n*n, scaled by 1/n
    s[i] = base;
    r[i] = base;
  }
  int bad = 0;
  long long rc = ft_host_allreduce(w, s.data(), r.data(), count, wire, 6, 0, 2, widths.data(), (int)widths.size(), ctas, tile, 0,
                                   1.0 / n, 1, 20000000000ull, -1, 0, 0);
  if (rc) { printf("wire rc=%lld\n", rc); bad = 1; }
  for (int i = 0; i < n; ++i)
    for (long long j = 0; j < count; ++j)
      if (((float*)r[i])[j] != (float)((j % 3) * n)) bad = 1;
  ft_host_world_destroy(w);
  printf("wire=%d n=%d stages=%zu count=%lld: %s\n", wire, n, widths.size(), count, bad ? "FAIL" : "ok");
  return bad;
}

int main() {
  int bad = 0;
  bad |= run(4, {2, 2}, 2, 0, 5000, 3, 256);
  bad |= run(6, {2, 3}, 2, 0, 7001, 2, 512);
  bad |= run(8, {2, 2, 2}, 2, 1, 6000, 2, 256);
  bad |= run(8, {8}, 2, 1, 6000, 3, 128);
  bad |= run(4, {4}, 3, 0, 3000, 2, 256);
  bad |= run(4, {4}, 3, 0, 30000, 5, 256);
  bad |= run(8, {8}, 3, 0, 20000, 8, 128);
  bad |= run(4, {4}, 3, 1, 3000, 2, 256);
  bad |= run(5, {1}, 4, 0, 3000, 2, 256);
  bad |= run(5, {1}, 4, 1, 3000, 2, 256);
  bad |= run(4, {4}, 1, 0, 999, 2, 256);
  bad |= run(4, {4}, 6, 0, 3001, 3, 256);  // two-shot LL
  bad |= run(7, {7}, 6, 0, 5, 2, 256);     // ... with empty blocks
  bad |= run(7, {2, 3}, 2, 0, 5000, 2, 256, 1);
  bad |= run(5, {3}, 2, 0, 3001, 3, 128, 2);
  // bulk tiles (algo 5, cpu_backend.h): private and registered buffers
  bad |= run(6, {2, 3}, 5, 0, 7000, 2, 512);
  bad |= run(8, {2, 2, 2}, 5, 1, 6000, 3, 256);
  bad |= run(4, {4}, 5, 0, 40000, 1, 4096);
  // the halves (reduce-scatter / all-gather as collectives), per-vector text and bulk tiles
  bad |= run_halves(6, {2, 3}, 1024, 2, 256, 2);
  bad |= run_halves(4, {4}, 512, 3, 128, 5);
#ifdef FT_WITH_F16
  // narrower wire formats under fp32 user buffers, fused scale
  bad |= run_wire(4, {2, 2}, 13, 4096, 2, 256);
  bad |= run_wire(8, {2, 4}, 14, 4096, 3, 256);
#endif
  return bad;
}
