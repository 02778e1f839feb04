// ft_bench - standalone benchmark / smoke check with the reference's command line.
//
// Capability parity: allreduce_over_mpi/benchmark.cpp:30-219 - flags --size --repeat --to-file
// --comm-type --tag --version (:67-116), data[i] = i (:119-123), configuration banner (:128-143), timed
// in-place allreduce per repeat (:147-178), rank-ordered "CHECK r: data[9..19]" lines (:180-189), the
// per-repeat time file "[tag.]N.len.{w-w-|nccl}.ar_test.<unixtime>.txt" (:194-213) and the
// "average time / min time" summary (:215). FT_TOPO selects the stage widths exactly as in the
// reference (mpi_mod.hpp:882-929).
//
// Differences: started as `ftrun -np N ft_bench ...` it is one process per GPU like the reference under mpirun
// (ft_comm_create_from_env: VMM segments + multicast, the same path as the torch API); started directly, ranks are
// the GPUs of this one process (--np N, one communicator per device) or, with --emulate N, N logical ranks on one GPU; the comparison library for `--comm-type nccl` is
// NCCL (the reference compares against the MPI library's own MPI_Allreduce); time is device time
// (CUDA events, max over ranks) instead of rank-0 MPI_Wtime; --sweep / --algo / --dtype are additions.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/flextree.h"
#include "../core/version.h"

#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                                \
    }                                                                                         \
  } while (0)
#define FT(call)                                                                         \
  do {                                                                                   \
    ft_result_t r_ = (call);                                                             \
    if (r_ != FT_SUCCESS) {                                                              \
      fprintf(stderr, "flextree error %d: %s (%s:%d)\n", (int)r_, ft_last_error(), __FILE__, __LINE__); \
      exit(1);                                                                           \
    }                                                                                    \
  } while (0)

// ---- NCCL through dlopen (baseline only; keeps libflextree free of any NCCL dependency) ----------
struct Nccl {
  void* lib = nullptr;
  int (*CommInitAll)(void**, int, const int*) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*GetUniqueId)(void*) = nullptr;                 // ncclUniqueId is 128 bytes
  bool load() {
    for (const char* n : {"libnccl.so.2", "libnccl.so"}) {
      lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (lib) break;
    }
    if (!lib) return false;
    CommInitAll = (decltype(CommInitAll))dlsym(lib, "ncclCommInitAll");
    AllReduce = (decltype(AllReduce))dlsym(lib, "ncclAllReduce");
    GroupStart = (decltype(GroupStart))dlsym(lib, "ncclGroupStart");
    GroupEnd = (decltype(GroupEnd))dlsym(lib, "ncclGroupEnd");
    CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
    GetUniqueId = (decltype(GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    return CommInitAll && AllReduce && GroupStart && GroupEnd;
  }
};

static void show_version() {
  fprintf(stderr, "\n----------\nFlexTree-B200 Standalone Benchmark\n%s\n", ft_version_string());
}

int main(int argc, char** argv) {
  int repeat = 1, np = 0, emulate = 0;
  long long data_len = 35;
  bool to_file = false, sweep = false, symmetric = false;
  std::string comm_type = "flextree", tag, algo = "auto", dtype = "fp32";
  for (int i = 1; i < argc; ++i) {
    auto need = [&](const char* f) { if (i + 1 >= argc) { fprintf(stderr, "%s needs a value\n", f); exit(1); } return argv[++i]; };
    if (!strcmp(argv[i], "--size")) data_len = atoll(need("--size"));
    else if (!strcmp(argv[i], "--repeat")) repeat = atoi(need("--repeat"));
    else if (!strcmp(argv[i], "--to-file")) to_file = true;
    else if (!strcmp(argv[i], "--comm-type")) comm_type = need("--comm-type");
    else if (!strcmp(argv[i], "--tag")) tag = need("--tag");
    else if (!strcmp(argv[i], "--version")) { show_version(); return 0; }
    else if (!strcmp(argv[i], "--np") || !strcmp(argv[i], "-np")) np = atoi(need("--np"));
    else if (!strcmp(argv[i], "--emulate")) emulate = atoi(need("--emulate"));
    else if (!strcmp(argv[i], "--algo")) algo = need("--algo");
    else if (!strcmp(argv[i], "--dtype")) dtype = need("--dtype");
    else if (!strcmp(argv[i], "--sweep")) sweep = true;
    else if (!strcmp(argv[i], "--symmetric")) symmetric = true;  // buffers from ft_symm_alloc: the zero-copy (NVLS / P2P in place) path
    else { fprintf(stderr, "unknown parameter: %s\n", argv[i]); return 1; }
  }
  if (comm_type == "mpi") comm_type = "nccl";  // the reference's "library allreduce" slot
  if (comm_type != "flextree" && comm_type != "nccl") { fprintf(stderr, "unknown comm type: %s\n", comm_type.c_str()); return 1; }
  if (dtype != "fp32") { fprintf(stderr, "ft_bench: only fp32 is wired here (the reference benchmark is MPI_FLOAT only); use bench/sweep.py for bf16\n"); return 1; }

  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  const bool mp = ft_env_size() > 0 && !emulate;   // started by ftrun: one process per GPU
  const int me = mp ? ft_env_rank() : 0;
  const int N = mp ? ft_env_size() : emulate > 0 ? emulate : (np > 0 ? np : ndev);
  if (!emulate && !mp && N > ndev) { fprintf(stderr, "need %d GPUs, have %d (use --emulate %d for one GPU)\n", N, ndev, N); return 1; }
  int widths[16];
  int nst = ft_parse_topo(nullptr, N, widths, 16);  // exits like the reference on a bad FT_TOPO
  if (nst < 0) { fprintf(stderr, "%s\n", ft_last_error()); return 1; }

  // communicators
  std::vector<ft_comm_t*> comms;
  const int L = mp ? 1 : N;                         // ranks driven by this process
  std::vector<int> devs(L);
  for (int i = 0; i < L; ++i) devs[i] = emulate ? 0 : mp ? ft_env_local_rank() % ndev : i;
  ft_comm_t* emu = nullptr;
  if (emulate) {
    FT(ft_comm_create_emulated(&emu, N, 0, nullptr));
  } else if (mp) {
    comms.resize(1);
    CK(cudaSetDevice(devs[0]));
    FT(ft_comm_create_from_env(&comms[0], devs[0], nullptr));
  } else {
    comms.resize(N);
    FT(ft_comm_create_local(comms.data(), N, devs.data(), nullptr));
  }
  auto each_comm = [&](auto fn) { if (emulate) fn(emu); else for (auto* c : comms) fn(c); };
  int algo_code = algo == "auto" ? FT_ALGO_AUTO : algo == "oneshot" ? FT_ALGO_ONESHOT : algo == "stages" ? FT_ALGO_STAGES
                : algo == "nvls" ? FT_ALGO_NVLS : algo == "ring" ? FT_ALGO_RING : algo == "twoshot" ? FT_ALGO_TWOSHOT : -1;
  if (algo_code < 0) { fprintf(stderr, "unknown --algo %s\n", algo.c_str()); return 1; }
  each_comm([&](ft_comm_t* c) { FT(ft_comm_set_algo(c, (ft_algo_t)algo_code)); });

  Nccl nccl;
  std::vector<void*> ncomms(L, nullptr);
  if (comm_type == "nccl") {
    if (emulate) { fprintf(stderr, "--comm-type nccl needs real GPUs\n"); return 1; }
    if (!nccl.load()) { fprintf(stderr, "NCCL unavailable\n"); return 1; }
    if (mp) {
      typedef struct { char b[128]; } Uid;
      auto init_rank = (int (*)(void**, int, Uid, int))dlsym(nccl.lib, "ncclCommInitRank");
      Uid id;
      memset(&id, 0, sizeof id);
      if (me == 0 && (!nccl.GetUniqueId || nccl.GetUniqueId(&id) != 0)) { fprintf(stderr, "ncclGetUniqueId failed\n"); return 1; }
      std::vector<Uid> all((size_t)N);
      if (ft_env_allgather(&id, all.data(), sizeof id) != 0 || !init_rank || init_rank(&ncomms[0], N, all[0], me) != 0) {
        fprintf(stderr, "NCCL init failed\n");
        return 1;
      }
    } else if (nccl.CommInitAll(ncomms.data(), N, devs.data()) != 0) {
      fprintf(stderr, "NCCL unavailable\n");
      return 1;
    }
  }

  const long long max_len = sweep ? std::max(data_len, (long long)(1 << 28)) : data_len;
  std::vector<float*> data(L);
  std::vector<cudaStream_t> streams(L);
  std::vector<cudaEvent_t> e0(L), e1(L);
  std::vector<float*> align_buf(L, nullptr);  // 1 KiB scratch per rank for the device-side start alignment, see timed()
  std::vector<float> host(max_len);
  for (long long i = 0; i < max_len; ++i) host[i] = (float)i;
  for (int r = 0; r < L; ++r) {
    CK(cudaSetDevice(devs[r]));
    if (symmetric && comm_type == "flextree" && !emulate) {
      void* p = nullptr;
      FT(ft_symm_alloc(comms[r], sizeof(float) * (size_t)max_len, &p));
      data[r] = (float*)p;
    } else {
      CK(cudaMalloc(&data[r], sizeof(float) * (size_t)max_len));
    }
    CK(cudaMemcpy(data[r], host.data(), sizeof(float) * (size_t)max_len, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&align_buf[r], 1024));
    CK(cudaMemset(align_buf[r], 0, 1024));
    CK(cudaStreamCreate(&streams[r]));
    CK(cudaEventCreate(&e0[r]));
    CK(cudaEventCreate(&e1[r]));
  }
  if (emulate) FT(ft_comm_reserve(emu, sizeof(float) * (size_t)max_len));
  else if (!symmetric) FT(ft_comm_reserve(comms[0], sizeof(float) * (size_t)max_len));  // zero-copy buffers need no staging window
  const bool talk = me == 0;  // rank 0 prints the banner, the summary and the time file

  {
    std::ostringstream ss;
    ss << "configuration: \n  - total_peers: " << N << "\n  - data_size: " << data_len << "\n  - repeat: " << repeat
       << "\n  - to_file: " << (to_file ? "true" : "false");
    if (to_file && !tag.empty()) ss << "\n  - file tag: " << tag;
    ss << "\n  - communication method: " << comm_type;
    if (comm_type == "flextree") {
      ss << "\n  - And FlexTree topo is ";
      for (int k = 0; k < nst; ++k) ss << widths[k] << " ";
      ss << "\n  - algorithm: " << algo << (emulate ? " (ranks emulated on one GPU)" : "");
      ss << "\n  - buffers: " << (symmetric && !emulate ? "symmetric (zero-copy)" : "private (cudaMalloc)")
         << "\n  - multicast (NVLS): " << (emulate ? 0 : ft_comm_has_multicast(comms[0]));
    }
    if (talk) fprintf(stderr, "\n%s\n", ss.str().c_str());
  }

  auto sync_all = [&]() {
    for (int r = 0; r < L; ++r) { CK(cudaSetDevice(devs[r])); CK(cudaDeviceSynchronize()); }
    if (mp && ft_env_barrier() != 0) { fprintf(stderr, "bootstrap barrier failed\n"); exit(1); }
  };
  auto one_allreduce = [&](long long len) {
    if (comm_type == "nccl") {
      nccl.GroupStart();
      for (int r = 0; r < L; ++r) nccl.AllReduce(data[r], data[r], (size_t)len, /*ncclFloat*/ 7, /*ncclSum*/ 0, ncomms[r], streams[r]);
      nccl.GroupEnd();
    } else if (emulate) {
      std::vector<void*> p(data.begin(), data.end());
      FT(ft_allreduce_multi(nullptr, p.data(), (size_t)len, FT_FLOAT, FT_FLOAT, FT_SUM, 1.0, emu, streams[0]));
    } else {
      for (int r = 0; r < L; ++r) {
        CK(cudaSetDevice(devs[r]));
        FT(ft_allreduce(FT_IN_PLACE, data[r], (size_t)len, FT_FLOAT, FT_SUM, comms[r], streams[r]));
      }
    }
  };
  // A tiny collective on scratch memory right before the timed one: host-side barriers (sockets, or one thread
  // walking over the devices) leave the ranks' launches tens of microseconds apart, and a rank that launched early
  // would bill that skew to the collective. The end of the aligning kernel is a device-side common start line.
  auto align_start = [&]() {
    if (emulate) return;
    if (comm_type == "nccl") {
      nccl.GroupStart();
      for (int r = 0; r < L; ++r) nccl.AllReduce(align_buf[r], align_buf[r], 256, 7, 0, ncomms[r], streams[r]);
      nccl.GroupEnd();
    } else {
      for (int r = 0; r < L; ++r) {
        CK(cudaSetDevice(devs[r]));
        FT(ft_allreduce(FT_IN_PLACE, align_buf[r], 256, FT_FLOAT, FT_SUM, comms[r], streams[r]));
      }
    }
  };
  // device time of one call: max over ranks of (event stop - event start)
  auto timed = [&](long long len) {
    sync_all();
    align_start();
    const int T = emulate ? 1 : L;
    for (int r = 0; r < T; ++r) { CK(cudaSetDevice(devs[r])); CK(cudaEventRecord(e0[r], streams[r])); }
    one_allreduce(len);
    for (int r = 0; r < T; ++r) { CK(cudaSetDevice(devs[r])); CK(cudaEventRecord(e1[r], streams[r])); }
    sync_all();
    float worst = 0;
    for (int r = 0; r < T; ++r) { float ms; CK(cudaEventElapsedTime(&ms, e0[r], e1[r])); worst = std::max(worst, ms); }
    if (mp) {  // max over the ranks' device times
      std::vector<float> all((size_t)N);
      if (ft_env_allgather(&worst, all.data(), sizeof worst) != 0) { fprintf(stderr, "bootstrap allgather failed\n"); exit(1); }
      for (float v : all) worst = std::max(worst, v);
    }
    return (double)worst * 1e-3;
  };

  if (sweep) {
    if (talk) printf("bytes,comm,algo,widths,avg_us,min_us,algbw_GBps,busbw_GBps,pct_of_770GBps\n");
    for (long long bytes = 1024; bytes <= (long long)sizeof(float) * max_len; bytes *= 4) {
      long long len = bytes / 4;
      for (int w = 0; w < 3; ++w) timed(len);
      double sum = 0, mn = 1e9;
      const int reps = std::max(3, repeat);
      for (int i = 0; i < reps; ++i) { double t = timed(len); sum += t; mn = std::min(mn, t); }
      ft_call_info_t info{};
      if (comm_type == "flextree") ft_comm_last_call(emulate ? emu : comms[0], &info);
      std::string ws;
      for (int k = 0; k < info.nstages; ++k) ws += (k ? "x" : "") + std::to_string(info.widths[k]);
      static const char* an[] = {"auto", "oneshot", "stages", "nvls", "ring", "stages", "twoshot"};
      double avg = sum / reps, alg = bytes / avg / 1e9, bus = alg * 2.0 * (N - 1) / N;
      if (talk) printf("%lld,%s,%s,%s,%.2f,%.2f,%.3f,%.3f,%.1f\n", bytes, comm_type.c_str(), comm_type == "flextree" ? an[info.algo] : "nccl",
             ws.c_str(), avg * 1e6, mn * 1e6, alg, bus, 100.0 * bus / 770.0);
      fflush(stdout);
    }
    return 0;
  }

  std::vector<double> repeat_time;
  double sum_time = 0, min_time = 1e30;
  for (int i = 0; i < repeat; ++i) {
    double t = timed(data_len);
    repeat_time.push_back(t);
    sum_time += t;
    min_time = std::min(min_time, t);
    if (talk) fprintf(stderr, "repeat %d finished\n", i);
  }
  unsigned err = 0;
  each_comm([&](ft_comm_t* c) { err |= ft_comm_error(c); });
  if (err) fprintf(stderr, "DEVICE ERROR %#x\n", err);

  // CHECK lines, in rank order (benchmark.cpp:180-189; reads indices 9..19, so sizes < 20 print fewer)
  for (int r = 0; r < N; ++r) {
    if (mp) {  // one rank at a time, in rank order, like the reference's barrier loop (benchmark.cpp:180-189)
      fflush(stdout);
      ft_env_barrier();
      if (r != me) continue;
    }
    const int l = mp ? 0 : r;
    CK(cudaSetDevice(devs[l]));
    float buf[11] = {0};
    long long cnt = std::min<long long>(11, std::max<long long>(0, data_len - 9));
    if (cnt > 0) CK(cudaMemcpy(buf, data[l] + 9, sizeof(float) * (size_t)cnt, cudaMemcpyDeviceToHost));
    std::ostringstream ss;
    ss << "CHECK " << r << ": ";
    for (long long i = 0; i < cnt; ++i) ss << buf[i] << " ";
    printf("%s\n", ss.str().c_str());
    fflush(stdout);
  }
  if (mp) ft_env_barrier();

  if (to_file && talk) {
    std::ostringstream ss;
    if (!tag.empty()) ss << tag << ".";
    ss << N << "." << data_len << ".";
    if (comm_type == "flextree") for (int k = 0; k < nst; ++k) ss << widths[k] << "-";
    else ss << "nccl";
    ss << ".ar_test." << time(nullptr) << ".txt";
    std::ofstream f(ss.str());
    for (double t : repeat_time) f << t << std::endl;
  }
  if (talk) fprintf(stderr, "\nDONE, average time: %g, min time: %g\n", sum_time / repeat, min_time);
  each_comm([&](ft_comm_t* c) { ft_comm_destroy(c); });
  return err ? 2 : 0;
}
