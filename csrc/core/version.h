// Version stamping (parity with the GIT_REPO_VERSION/DATE/HASH injection of
// allreduce_over_mpi/CMakeLists.txt:10-31, shown by `ft_bench --version`, benchmark.cpp:13-16).
// The Makefile passes -DFT_GIT_DESCRIBE / -DFT_GIT_DATE / -DFT_GIT_HASH when git is available.
#pragma once
#ifndef FT_GIT_DESCRIBE
#define FT_GIT_DESCRIBE "unknown"
#endif
#ifndef FT_GIT_DATE
#define FT_GIT_DATE "unknown"
#endif
#ifndef FT_GIT_HASH
#define FT_GIT_HASH "unknown"
#endif
#define FT_VERSION_STRING "flextree-b200 0.1 (" FT_GIT_DESCRIBE ", " FT_GIT_DATE ", " FT_GIT_HASH ")"
