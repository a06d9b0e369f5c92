// Shared declarations of the B200 SevenNet hot-path kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/sevennet_b200.h"

namespace b200 {

// error channel of the C-ABI (SURVEY 8b "Errors": status code + b200_last_error()).
void set_error(const char* fmt, ...);
int check_launch(const char* what, int n_launched = 1);  // also counts kernels for b200_launch_count()

#define B200_REQUIRE(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      b200::set_error(__VA_ARGS__);        \
      return B200_ERR_INVALID;             \
    }                                      \
  } while (0)

constexpr float kSilu2Mom = 1.6791767923989418f;  // e3nn normalize2mom(silu), frozen-graph constant c5

__device__ __forceinline__ float silu_f(float z) { return __fdividef(z, 1.0f + __expf(-z)); }
// value and derivative of c*silu(z)
__device__ __forceinline__ void silu_vd(float z, float& v, float& d) {
  float s = __fdividef(1.0f, 1.0f + __expf(-z));
  v = kSilu2Mom * z * s;
  d = kSilu2Mom * s * (1.0f + z * (1.0f - s));
}

// canonical path ids of x(0e,1e,2e) (x) Y(0e,1e,2e) -> l3<=2, reference instruction order
// (sevenn/nn/convolution.py:72-86): 000 011 022 101 110 111 112 121 122 202 211 212 220 221 222
enum PathId { P000 = 0, P011, P022, P101, P110, P111, P112, P121, P122, P202, P211, P212, P220, P221, P222, kNumPaths };

struct ConvDesc {
  int m0, m1, m2;        // multiplicities of the 0e / 1e / 2e blocks of x
  int lout;              // 0 or 2: highest l kept in the output
  int xdim, wn, odim;    // row widths of x, of the per-edge weights, of the output
  int woff[kNumPaths];   // offset of each path in the weight row (-1: path absent)
  int ooff[kNumPaths];   // offset of each path's slot in the output row
};

inline int round_up(int a, int b) { return (a + b - 1) / b * b; }

// The stream-ordered allocator returns freed memory to the driver at the next synchronisation unless the pool's
// release threshold is raised; the operand packs (a few hundred KB, allocated and freed around every tensor-core
// launch) would otherwise cost a driver allocation per call.  Once per device.
inline void keep_async_pool() {
  static bool done[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || done[dev]) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
    unsigned long long thr = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
  }
  done[dev] = true;
}

}  // namespace b200
