// Shared host-side helpers for the NDS B200 engine (error plumbing, CUDA checks).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include <cuda_runtime.h>

namespace nds {

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

[[noreturn]] inline void fail(const std::string& msg) { throw Error(msg); }

#define NDS_CUDA(expr)                                                                  \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess)                                                              \
      ::nds::fail(std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " +      \
                  __FILE__ + ":" + std::to_string(__LINE__) + " (" #expr ")");         \
  } while (0)

#define NDS_REQUIRE(cond, msg)                \
  do {                                        \
    if (!(cond)) ::nds::fail(std::string(msg)); \
  } while (0)

inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

}  // namespace nds
