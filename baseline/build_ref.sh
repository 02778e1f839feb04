#!/bin/bash
# Builds the UNMODIFIED reference (allreduce_over_mpi/benchmark.cpp + mpi_mod.hpp and the cost_model
# driver) against the in-repo MPI / glog shims into baseline/_ref/. Sources are copied verbatim from
# $REF (default /root/reference) into baseline/_ref/src so the build also works on the GPU box, where
# only the repo snapshot exists. Flags follow allreduce_over_mpi/CMakeLists.txt:5.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/_ref"
CXX=/usr/bin/g++; [ -x "$CXX" ] || CXX=g++
mkdir -p "$OUT/src" "$OUT/bin"
if [ -d "$REF/allreduce_over_mpi" ]; then
  rm -rf "$OUT/src/allreduce_over_mpi" "$OUT/src/cost_model" "$OUT/src/topo_count"
  cp -r "$REF/allreduce_over_mpi" "$REF/cost_model" "$REF/topo_count" "$OUT/src/"
fi
[ -f "$OUT/src/allreduce_over_mpi/benchmark.cpp" ] || { echo "reference sources not found" >&2; exit 1; }
FLAGS="-std=c++11 -O2 -fPIC -fopenmp -mavx2 -w"
$CXX $FLAGS '-DGIT_REPO_VERSION="reference"' '-DGIT_REPO_DATE="unknown"' '-DGIT_REPO_HASH="unmodified"' \
  -I"$HERE/mpi_shim" -I"$HERE/glog_shim" \
  "$OUT/src/allreduce_over_mpi/benchmark.cpp" "$HERE/mpi_shim/mpi_shim.cc" -o "$OUT/bin/allreduce_over_mpi" -lpthread
$CXX -std=c++17 -O2 "$HERE/mpi_shim/ftmpirun.cc" -o "$OUT/bin/ftmpirun"
# oracle driver: the reference header, unmodified, on caller-supplied data (tests/test_reference_oracle.py)
$CXX $FLAGS -I"$HERE/mpi_shim" -I"$HERE/glog_shim" -I"$OUT/src/allreduce_over_mpi" \
  "$HERE/ref_oracle.cc" "$HERE/mpi_shim/mpi_shim.cc" -o "$OUT/bin/ref_oracle" -lpthread
# the offline planner of the reference (cost_model/main.cpp), unmodified
$CXX -std=c++14 -O0 -w "$OUT/src/cost_model/main.cpp" -o "$OUT/bin/cost_model" || echo "cost_model build failed (non-fatal)" >&2
echo "reference built in $OUT/bin"
# the drop-in build: the SAME unmodified benchmark.cpp, but `#include "mpi_mod.hpp"` resolves to compat/mpi_mod.hpp
# (this library's CPU data plane). A quoted include looks next to the including file first, so benchmark.cpp is
# copied alone into its own directory.
ROOT="$(cd "$HERE/.." && pwd)"
if [ -f "$ROOT/build/libflextree_host.so" ]; then
  mkdir -p "$OUT/dropin"
  cp "$OUT/src/allreduce_over_mpi/benchmark.cpp" "$OUT/dropin/benchmark.cpp"
  $CXX -std=c++11 -O2 -fPIC -w -DSTANDALONE_TEST '-DGIT_REPO_VERSION="flextree-b200 drop-in"' '-DGIT_REPO_DATE="unknown"' '-DGIT_REPO_HASH="unmodified benchmark.cpp"' \
    -I"$ROOT/compat" -I"$ROOT/include" -I"$HERE/mpi_shim" -I"$HERE/glog_shim" \
    "$OUT/dropin/benchmark.cpp" "$HERE/mpi_shim/mpi_shim.cc" -o "$OUT/bin/allreduce_over_mpi_dropin" \
    -L"$ROOT/build" -lflextree_host -Wl,-rpath,"$ROOT/build" -lpthread -lrt
  echo "drop-in built: $OUT/bin/allreduce_over_mpi_dropin"
else
  echo "build/libflextree_host.so missing: run 'make host' first to get the drop-in binary" >&2
fi
if [ -f "$ROOT/build/libflextree_host.so" ] && [ -f "$ROOT/tests/cpp/dropin_user.cc" ]; then
  # a plain MPI program whose MPI_Allreduce calls are captured by the header (no STANDALONE_TEST)
  $CXX -std=c++11 -O2 -fPIC -w -I"$ROOT/compat" -I"$ROOT/include" -I"$HERE/mpi_shim" \
    "$ROOT/tests/cpp/dropin_user.cc" "$HERE/mpi_shim/mpi_shim.cc" -o "$OUT/bin/dropin_user" \
    -L"$ROOT/build" -lflextree_host -Wl,-rpath,"$ROOT/build" -lpthread -lrt
fi
