// POD plan + layout shared by host code, the CPU model of the kernels and the CUDA kernels.
//
// Math being implemented (re-derived from the reference's schedule generator, see SURVEY.md 2.7;
// reference code: allreduce_over_mpi/mpi_mod.hpp:152-179 (send groups), :187-213 (recv groups),
// :216-243 (split sizing)):
//   N = w_0 * w_1 * ... * w_{m-1},  g_0 = 1,  g_{k+1} = g_k * w_k
//   stage-k group of rank r : { r - d_k*g_k + j*g_k : 0 <= j < w_k },  d_k = (r / g_k) mod w_k
//   before RS stage k rank r holds partial sums of blocks { b : b == r (mod g_k) };
//   after it, of blocks { b : b == r (mod g_{k+1}) }.
// The reference materialises per-call tables of (peer, block list); here everything is closed-form
// so a kernel needs only this ~80-byte struct.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define FT_HD __host__ __device__ __forceinline__
#else
#define FT_HD inline
#endif

#define FT_MAX_STAGES 8
#define FT_MAX_RANKS 64
#define FT_MAX_WIDTH 32   // reference limit is 20 (mpi_mod.hpp:245); we allow up to 32 per stage
#define FT_MAX_LOCAL 32   // logical ranks served by one launch (single-GPU emulation mode)
#define FT_MAX_CTAS 256   // CTAs per rank that own flag slots

namespace ft {

struct Plan {
  int nranks;   // ranks in the tree (product of widths)
  int nstages;  // m
  int width[FT_MAX_STAGES];
  int gap[FT_MAX_STAGES + 1];  // g_0 .. g_m
};

// Returns false if the widths are unusable (empty, w<2 for multi-rank, product overflow, too many).
inline bool make_plan(const int* widths, int nstages, Plan* p) {
  if (nstages < 1 || nstages > FT_MAX_STAGES) return false;
  long long g = 1;
  p->nstages = nstages;
  for (int k = 0; k < FT_MAX_STAGES; ++k) p->width[k] = 1;
  p->gap[0] = 1;
  for (int k = 0; k < nstages; ++k) {
    int w = widths[k];
    if (w < 1 || w > FT_MAX_WIDTH) return false;
    if (w == 1 && nstages != 1) return false;  // 1 is only meaningful as the whole plan (N == 1)
    g *= w;
    if (g > FT_MAX_RANKS) return false;
    p->width[k] = w;
    p->gap[k + 1] = (int)g;
  }
  for (int k = nstages + 1; k <= FT_MAX_STAGES; ++k) p->gap[k] = (int)g;
  p->nranks = (int)g;
  return true;
}

FT_HD int stage_digit(const Plan& p, int r, int k) { return (r / p.gap[k]) % p.width[k]; }
// lowest rank of r's stage-k group; members are base + j*gap[k], j < width[k]
FT_HD int stage_base(const Plan& p, int r, int k) { return r - stage_digit(p, r, k) * p.gap[k]; }
FT_HD int stage_peer(const Plan& p, int r, int k, int j) { return stage_base(p, r, k) + j * p.gap[k]; }
// number of blocks a rank owns after RS stage k (== blocks exchanged per peer in that stage)
FT_HD int owned_after(const Plan& p, int k) { return p.nranks / p.gap[k + 1]; }

// Element layout of one message. Blocks are the reference's ceil(count/N) split
// (mpi_mod.hpp:231-236) rounded up so that every block starts on a 16-byte boundary; the tail
// block(s) may be short or empty exactly as in the reference (:679-696, :791-805).
struct Layout {
  long long count;   // elements in the message
  long long block;   // B: elements per block (multiple of vec)
  long long tile;    // C: elements of each block handled per tile (multiple of vec)
  int ntiles;        // ceil(B / C)
  int vec;           // elements per 16-byte vector (1 when the scalar path is used)
  int esize;         // bytes per element
};

inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

inline Layout make_layout(long long count, int nranks, int esize, long long tile_bytes, bool vectorized) {
  Layout l;
  l.count = count;
  l.esize = esize;
  l.vec = vectorized ? (16 / esize) : 1;
  long long b = ceil_div_ll(count, nranks);
  if (b < 1) b = 1;
  l.block = ceil_div_ll(b, l.vec) * l.vec;
  long long c = tile_bytes / esize;
  c = (c / l.vec) * l.vec;
  if (c < l.vec) c = l.vec;
  if (c > l.block) c = l.block;
  l.tile = c;
  l.ntiles = (int)ceil_div_ll(l.block, c);
  return l;
}

// [lo, hi) element range of tile t inside block b, clamped to the message; hi <= lo means empty.
FT_HD void tile_range(const Layout& l, long long b, int t, long long* lo, long long* hi) {
  long long s = b * l.block + (long long)t * l.tile;
  long long e = s + l.tile;
  long long bend = (b + 1) * l.block;
  if (e > bend) e = bend;
  if (e > l.count) e = l.count;
  *lo = s;
  *hi = e;
}

}  // namespace ft
