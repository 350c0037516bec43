// oracle/build_info.cpp — the four symbols common/build-info.cpp.in declares (LLAMA_BUILD_NUMBER, LLAMA_COMMIT, LLAMA_COMPILER,
// LLAMA_BUILD_TARGET), written by hand: the reference's generator (common/cmake/build-info-gen-cpp.cmake) is missing from
// the snapshot and would write into the read-only source tree.  TEST / MEASUREMENT INFRASTRUCTURE ONLY.
int LLAMA_BUILD_NUMBER = 0;
char const *LLAMA_COMMIT = "snapshot";
char const *LLAMA_COMPILER = "g++ (oracle/Makefile refbench)";
char const *LLAMA_BUILD_TARGET = "x86_64-linux-gnu";
