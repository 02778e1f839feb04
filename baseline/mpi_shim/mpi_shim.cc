// Single-host MPI subset over shared memory (see mpi.h). Processes are started by ftmpirun, which
// creates one memfd for the control block and one sparse memfd "arena" per rank and passes them down
// as inherited file descriptors. Every process maps every arena at the same fixed address, and this
// library replaces operator new[] so that large arrays (the reference allocates its data, recvbuf and
// staging buffer with new[]) live in the caller's arena. A send is then just a descriptor
// {pointer, bytes}: the receiver copies straight out of the sender's arena (rendezvous, zero
// intermediate copy) - so the measured time is the reference's algorithm + one memcpy per message,
// which is the most favourable transport an MPI library could offer it on one host.
#include "mpi.h"

#include <errno.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <new>

namespace {

constexpr int kMaxRanks = 64;
constexpr int kQueue = 8192;                       // outstanding sends per (src, dst)
constexpr uintptr_t kArenaBase = 0x600000000000ull;
constexpr size_t kArenaSize = 16ull << 30;         // virtual, sparse
constexpr size_t kArenaThreshold = 4096;           // new[] of at least this goes to the arena

struct SendDesc {
  const void* ptr;
  size_t bytes;
  std::atomic<uint32_t> done;
};
struct Channel {  // src -> dst
  std::atomic<uint64_t> posted;  // number of descriptors published by src
  char pad[56];
  SendDesc q[kQueue];
};
struct Control {
  std::atomic<int> barrier_count;
  std::atomic<int> barrier_sense;
  std::atomic<uint64_t> arena_top[kMaxRanks];
  const void* ar_send[kMaxRanks];  // scratch for the native allreduce
  void* ar_recv[kMaxRanks];
  Channel ch[kMaxRanks][kMaxRanks];
};

int g_rank = 0, g_size = 1;
Control* g_ctl = nullptr;
bool g_init = false;
uint64_t g_recv_seq[kMaxRanks];  // receives posted so far from each source
uint64_t g_send_seq[kMaxRanks];
int g_local_sense = 0;

void die(const char* msg) {
  fprintf(stderr, "[mpi_shim rank %d] %s (errno %d: %s)\n", g_rank, msg, errno, strerror(errno));
  _exit(3);
}

inline void relax(unsigned& spins) {
  if (++spins > 200) { sched_yield(); spins = 0; }
}

bool in_arena(const void* p) {
  uintptr_t a = (uintptr_t)p;
  return a >= kArenaBase && a < kArenaBase + (uintptr_t)kMaxRanks * kArenaSize;
}

void* arena_alloc(size_t n) {
  n = (n + 63) & ~(size_t)63;
  uint64_t off = g_ctl->arena_top[g_rank].fetch_add(n);
  if (off + n > kArenaSize) die("arena exhausted");
  return (void*)(kArenaBase + (uintptr_t)g_rank * kArenaSize + off);
}

void setup() {
  const char* r = getenv("FTMPI_RANK");
  const char* s = getenv("FTMPI_SIZE");
  const char* c = getenv("FTMPI_CTRL_FD");
  const char* a0 = getenv("FTMPI_ARENA_FD0");
  if (!r || !s || !c || !a0) {  // not under ftmpirun: singleton world, private "arena"
    g_rank = 0;
    g_size = 1;
    g_ctl = (Control*)mmap(nullptr, sizeof(Control), PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (g_ctl == MAP_FAILED) die("mmap control");
    void* m = mmap((void*)kArenaBase, kArenaSize, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE | MAP_FIXED_NOREPLACE, -1, 0);
    if (m == MAP_FAILED) die("mmap arena");
    g_init = true;
    return;
  }
  g_rank = atoi(r);
  g_size = atoi(s);
  if (g_size > kMaxRanks) die("too many ranks");
  g_ctl = (Control*)mmap(nullptr, sizeof(Control), PROT_READ | PROT_WRITE, MAP_SHARED, atoi(c), 0);
  if (g_ctl == MAP_FAILED) die("mmap control");
  int fd0 = atoi(a0);
  for (int i = 0; i < g_size; ++i) {
    void* want = (void*)(kArenaBase + (uintptr_t)i * kArenaSize);
    void* m = mmap(want, kArenaSize, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_NORESERVE | MAP_FIXED_NOREPLACE, fd0 + i, 0);
    if (m != want) die("mmap arena at fixed address");
  }
  g_init = true;
}

struct AutoInit {  // arenas must exist before main() allocates anything with new[]
  AutoInit() { if (!g_init) setup(); }
} g_auto_init __attribute__((init_priority(101)));

size_t type_size(MPI_Datatype dt) {
  switch (dt) {
    case MPI_UINT8_T: case MPI_INT8_T: case MPI_C_BOOL: return 1;
    case MPI_UINT16_T: case MPI_INT16_T: return 2;
    case MPI_INT32_T: case MPI_UINT32_T: case MPI_FLOAT: return 4;
    case MPI_INT64_T: case MPI_UINT64_T: case MPI_DOUBLE: case MPI_LONG_LONG_INT: return 8;
  }
  return 0;
}

template <typename T> void reduce_into(T* dst, const T* src, size_t n, MPI_Op op, bool first) {
  if (first) { memcpy(dst, src, n * sizeof(T)); return; }
  switch (op) {
    case MPI_SUM: for (size_t i = 0; i < n; ++i) dst[i] = (T)(dst[i] + src[i]); break;
    case MPI_MAX: for (size_t i = 0; i < n; ++i) dst[i] = dst[i] > src[i] ? dst[i] : src[i]; break;
    case MPI_MIN: for (size_t i = 0; i < n; ++i) dst[i] = dst[i] < src[i] ? dst[i] : src[i]; break;
    default: break;
  }
}
template <typename T> void band_into(T* dst, const T* src, size_t n) {
  for (size_t i = 0; i < n; ++i) dst[i] = (T)(dst[i] & src[i]);
}

}  // namespace

// ---- operator new[] / delete[]: large arrays live in the shared arena -------------------------------
void* operator new[](size_t n) {
  if (!g_init) setup();
  if (n >= kArenaThreshold) return arena_alloc(n);
  void* p = malloc(n ? n : 1);
  if (!p) throw std::bad_alloc();
  return p;
}
void operator delete[](void* p) noexcept {
  if (p && !in_arena(p)) free(p);  // arena memory is bump-allocated and never reused
}
void operator delete[](void* p, size_t) noexcept { operator delete[](p); }

extern "C" {

int MPI_Init_thread(int*, char***, int required, int* provided) {
  if (!g_init) setup();
  if (provided) *provided = required;
  MPI_Barrier(MPI_COMM_WORLD);
  return MPI_SUCCESS;
}
int MPI_Finalize(void) {
  MPI_Barrier(MPI_COMM_WORLD);
  return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm, int* size) { *size = g_size; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm, int* rank) { *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_split(MPI_Comm comm, int, int, MPI_Comm* newcomm) { *newcomm = comm; return MPI_SUCCESS; }
int MPI_Type_size(MPI_Datatype dt, int* size) { *size = (int)type_size(dt); return MPI_SUCCESS; }
int MPI_Type_get_name(MPI_Datatype dt, char* name, int* len) {
  *len = snprintf(name, 19, "MPI_TYPE_%d", dt);
  return MPI_SUCCESS;
}
double MPI_Wtime(void) {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int MPI_Isend(const void* buf, int count, MPI_Datatype dt, int dest, int, MPI_Comm, MPI_Request* req) {
  size_t bytes = (size_t)count * type_size(dt);
  const void* src = buf;
  if (!in_arena(buf) && bytes) {  // stack/static/small-heap data: bounce through my arena
    void* b = arena_alloc(bytes);
    memcpy(b, buf, bytes);
    src = b;
  }
  Channel& ch = g_ctl->ch[g_rank][dest];
  uint64_t seq = g_send_seq[dest]++;
  SendDesc& d = ch.q[seq % kQueue];
  unsigned spins = 0;
  // slot reuse: wait until the receive that used this slot kQueue sends ago has completed
  while (seq >= kQueue && d.done.load(std::memory_order_acquire) == 0) relax(spins);
  d.ptr = src;
  d.bytes = bytes;
  d.done.store(0, std::memory_order_relaxed);
  ch.posted.store(seq + 1, std::memory_order_release);
  req->kind = 1;
  req->peer = dest;
  req->seq = seq;
  req->buf = nullptr;
  req->bytes = bytes;
  return MPI_SUCCESS;
}

int MPI_Irecv(void* buf, int count, MPI_Datatype dt, int source, int, MPI_Comm, MPI_Request* req) {
  req->kind = 2;
  req->peer = source;
  req->seq = g_recv_seq[source]++;  // k-th receive from `source` matches its k-th send (non-overtaking)
  req->buf = buf;
  req->bytes = (size_t)count * type_size(dt);
  return MPI_SUCCESS;
}

int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status*) {
  // receives first (the receiver does the copy), then wait for my sends to be consumed
  for (int i = 0; i < count; ++i) {
    MPI_Request& r = reqs[i];
    if (r.kind != 2) continue;
    Channel& ch = g_ctl->ch[r.peer][g_rank];
    unsigned spins = 0;
    while (ch.posted.load(std::memory_order_acquire) <= r.seq) relax(spins);
    SendDesc& d = ch.q[r.seq % kQueue];
    size_t n = d.bytes < r.bytes ? d.bytes : r.bytes;
    if (n) memcpy(r.buf, d.ptr, n);
    d.done.store(1, std::memory_order_release);
    r.kind = 0;
  }
  for (int i = 0; i < count; ++i) {
    MPI_Request& r = reqs[i];
    if (r.kind != 1) continue;
    SendDesc& d = g_ctl->ch[g_rank][r.peer].q[r.seq % kQueue];
    unsigned spins = 0;
    while (d.done.load(std::memory_order_acquire) == 0) relax(spins);
    r.kind = 0;
  }
  return MPI_SUCCESS;
}

int MPI_Barrier(MPI_Comm) {
  if (g_size == 1) return MPI_SUCCESS;
  g_local_sense ^= 1;
  if (g_ctl->barrier_count.fetch_add(1, std::memory_order_acq_rel) == g_size - 1) {
    g_ctl->barrier_count.store(0, std::memory_order_relaxed);
    g_ctl->barrier_sense.store(g_local_sense, std::memory_order_release);
  } else {
    unsigned spins = 0;
    while (g_ctl->barrier_sense.load(std::memory_order_acquire) != g_local_sense) relax(spins);
  }
  return MPI_SUCCESS;
}

// "Native" allreduce of the shim (what `--comm-type mpi` measures): slice-owner reduce + gather over
// the shared arenas. Buffers outside the arena are bounced.
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm) {
  const size_t es = type_size(dt), bytes = (size_t)count * es;
  if (sendbuf == MPI_IN_PLACE) sendbuf = recvbuf;
  if (g_size == 1) {
    if (sendbuf != recvbuf) memcpy(recvbuf, sendbuf, bytes);
    return MPI_SUCCESS;
  }
  const void* s = sendbuf;
  void* r = recvbuf;
  if (!in_arena(sendbuf)) { void* b = arena_alloc(bytes); memcpy(b, sendbuf, bytes); s = b; }
  if (!in_arena(recvbuf)) r = arena_alloc(bytes);
  g_ctl->ar_send[g_rank] = s;
  g_ctl->ar_recv[g_rank] = r;
  MPI_Barrier(0);
  const size_t per = ((size_t)count + g_size - 1) / g_size;
  const size_t lo = per * g_rank < (size_t)count ? per * g_rank : (size_t)count;
  const size_t hi = lo + per < (size_t)count ? lo + per : (size_t)count;
  void* tmp = arena_alloc((hi - lo) * es + 64);
  for (int p = 0; p < g_size; ++p) {
    const char* src = (const char*)g_ctl->ar_send[p] + lo * es;
    const size_t n = hi - lo;
    const bool first = p == 0;
    if (op == MPI_BAND && !first) {
      switch (es) {
        case 1: band_into((uint8_t*)tmp, (const uint8_t*)src, n); break;
        case 2: band_into((uint16_t*)tmp, (const uint16_t*)src, n); break;
        case 4: band_into((uint32_t*)tmp, (const uint32_t*)src, n); break;
        default: band_into((uint64_t*)tmp, (const uint64_t*)src, n); break;
      }
      continue;
    }
    switch (dt) {
      case MPI_FLOAT: reduce_into((float*)tmp, (const float*)src, n, op, first); break;
      case MPI_DOUBLE: reduce_into((double*)tmp, (const double*)src, n, op, first); break;
      case MPI_UINT8_T: case MPI_C_BOOL: reduce_into((uint8_t*)tmp, (const uint8_t*)src, n, op, first); break;
      case MPI_INT8_T: reduce_into((int8_t*)tmp, (const int8_t*)src, n, op, first); break;
      case MPI_UINT16_T: reduce_into((uint16_t*)tmp, (const uint16_t*)src, n, op, first); break;
      case MPI_INT16_T: reduce_into((int16_t*)tmp, (const int16_t*)src, n, op, first); break;
      case MPI_INT32_T: reduce_into((int32_t*)tmp, (const int32_t*)src, n, op, first); break;
      case MPI_UINT32_T: reduce_into((uint32_t*)tmp, (const uint32_t*)src, n, op, first); break;
      case MPI_UINT64_T: reduce_into((uint64_t*)tmp, (const uint64_t*)src, n, op, first); break;
      default: reduce_into((int64_t*)tmp, (const int64_t*)src, n, op, first); break;
    }
  }
  MPI_Barrier(0);  // everyone has finished reading the send buffers (they may alias recv)
  for (int p = 0; p < g_size; ++p) memcpy((char*)g_ctl->ar_recv[p] + lo * es, tmp, (hi - lo) * es);
  MPI_Barrier(0);
  if (r != recvbuf) memcpy(recvbuf, r, bytes);
  return MPI_SUCCESS;
}

}  // extern "C"
