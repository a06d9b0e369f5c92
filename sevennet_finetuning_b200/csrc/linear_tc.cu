// Per-node irreps linear on the tcgen05 tensor cores (3xTF32, fp32 accumulation in TMEM).
//
// Reference: sevenn/nn/linear.py:46-49 (IrrepsLinear -> e3nn o3.Linear), sevenn/nn/self_connection.py:60-62.
//   y[n, o, m] = scale * sum_k x[n, k, m] W[k, o]     per l-block, scale = 1/sqrt(mul_in)
// GEMM view: rows = (node, m) pairs, K = mul_in, N = mul_out.  One CTA owns 128 rows of one block:
//   A = x tile (rows, K-major): gathered from the irreps-blocked rows ([mul, 2l+1], m fastest), split into tf32
//       hi/lo and written in the canonical operand layout by all threads, 16 k at a time (two stages),
//   B = W^T (rows = output channel, K-major): prepacked hi/lo per 16-k chunk, streamed with cp.async.bulk,
//   D = 128 lanes (rows) x N columns in TMEM (N <= 256 per pass).
// Epilogue: a warp transposes 32x32 blocks of D through shared memory so that its stores run along the output
// channels of one (node, m) row.  transpose = 1 is the x-gradient (roles of in/out swapped, W read transposed).
#include "common.cuh"
#include "tc.cuh"

namespace b200 {

using namespace tc;

constexpr int LK = 16;                 // k per chunk
constexpr int L_ACH = 128 * 16;        // bytes of one 16-byte K-chunk column of the 128-row A tile
constexpr int L_APART = (LK / 4) * L_ACH;   // 8 KB (hi or lo)
constexpr int L_MAXN = 64;             // output columns per pass
constexpr int L_BPART = (LK / 4) * L_MAXN * 16;   // 4 KB (hi or lo)
constexpr int L_T = 256;
constexpr int L_GPT = (128 * LK + L_T - 1) / L_T;   // gather slots per thread

struct LinTcBlock {
  int D, K, N, Npad;          // 2l+1, contraction length, real / padded (multiple of 16) output columns of this pass
  int xoff, yoff, o0;         // block offsets in the input / output rows; first output channel of the pass
  int pack_off;               // float offset of the pass's operand pack
  int w_off, ldw;             // block offset in the flat weight vector, its row length (mul_out)
  int nchunk;                 // ceil(K / 16)
  float scale;
};
struct LinTcPlan {
  int npass;
  int ldx, ldy;
  LinTcBlock p[16];
};

struct LinTcSmem {
  uint8_t A[2][2 * L_APART];
  uint8_t B[2][2 * L_BPART];
  float T[L_T / 32][32][33];   // per-warp transposition scratch of the epilogue
  uint64_t full_b[2], mma_done[2];
  uint32_t tmem_base;
};

// pack[pass][chunk][hi|lo][k4][Npad rows][4]: row o, k -> W[k][o0+o] (or W[o0+o][k] for the transposed product)
__global__ void linear_prepack_kernel(LinTcPlan pl, const float* __restrict__ W, int transpose, float* __restrict__ pack) {
  const LinTcBlock& b = pl.p[blockIdx.y];
  const int per_chunk = (LK / 4) * b.Npad;
  const int total = b.nchunk * per_chunk;
  for (int it = blockIdx.x * blockDim.x + threadIdx.x; it < total; it += gridDim.x * blockDim.x) {
    const int c = it / per_chunk, rem = it % per_chunk, k4 = rem / b.Npad, o = rem % b.Npad;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (o < b.N) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = c * LK + k4 * 4 + j;
        if (k < b.K)
          v[j] = transpose ? W[b.w_off + (size_t)(b.o0 + o) * b.ldw + k] : W[b.w_off + (size_t)k * b.ldw + b.o0 + o];
      }
    }
    float4 hi, lo;
    split4(make_float4(v[0], v[1], v[2], v[3]), hi, lo);
    float* dst = pack + b.pack_off + (size_t)c * (2 * per_chunk * 4) + (size_t)(k4 * b.Npad + o) * 4;
    *reinterpret_cast<float4*>(dst) = hi;
    *reinterpret_cast<float4*>(dst + per_chunk * 4) = lo;
  }
}

// The tensor core adds into its fp32 accumulator with truncation: the bias grows with the number of MMAs an
// accumulator receives (measured: -2.4e-8 relative per MMA).  So every K=8 step of the main product (hi*hi) gets a
// fresh TMEM accumulator (two per chunk, ping-pong over chunks) that the threads drain into fp32 registers with
// round-to-nearest adds; the two small cross products (lo*hi, hi*lo) of the step are issued first into the same
// accumulator, so the large product is added (one truncation) to a sum that is 2^-11 of its size.  Four accumulators
// of 64 columns = 256 TMEM columns per CTA: two CTAs share an SM and hide each other's gather and drain latency.
__global__ void __launch_bounds__(L_T, 2)
irreps_linear_tc_kernel(LinTcPlan pl, int n_nodes, const float* __restrict__ x, const float* __restrict__ pack,
                        float* __restrict__ y, int accumulate) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  LinTcSmem& s = *reinterpret_cast<LinTcSmem*>(smem_raw);
  const LinTcBlock b = pl.p[blockIdx.y];
  const int D = b.D, TN = 128 / D;               // nodes per tile
  const int n0 = blockIdx.x * TN;
  if (n0 >= n_nodes) return;
  const int nn = min(TN, n_nodes - n0);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, half = warp >> 2;      // TMEM lane quarter, 32-column half of the pass
  const uint32_t bpart = (uint32_t)(LK / 4) * b.Npad * 16;   // bytes of one part (hi or lo) of a B chunk

  if (warp == 0) tmem_alloc(&s.tmem_base, 256);
  if (threadIdx.x == 32) {
    for (int k = 0; k < 2; ++k) { mbar_init(&s.full_b[k], 1); mbar_init(&s.mma_done[k], 1); }
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;
  const uint32_t idesc = make_idesc_tf32(128, b.Npad);
  const uint32_t tlane = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(32 * half);
  const bool have_cols = 32 * half < b.Npad;
  float acc[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = 0.f;
  auto drain = [&](int c) {   // add the two accumulators of chunk c (its MMAs have retired) into the registers
    if (have_cols) {
      float v[32];
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        tmem_ld32(tlane + (uint32_t)((2 * (c & 1) + kk) * L_MAXN), v);
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] += v[j];
      }
    }
    tc_fence_before();
  };

  // per-thread gather slots (fixed over the chunks).  A warp fills one 8-row x 16-byte core matrix group per slot
  // (lane -> row = lane / 4, k = lane % 4 inside a 4-k unit): its 32 stores land in 128 contiguous bytes of the operand
  // tile - no bank conflicts (the former mapping, lanes along the contiguous (k, m) index of a node, was 4-way conflicted
  // and spent 30 % of the kernel on the shared-memory store pipe: profiles/r01n_finetune_linear_tc_ncu_raw.csv).  The four
  // k-units of the same rows are consecutive slots of the warp, so the 32-byte sectors of the gathered rows are reused from L1.
  const int span = LK * D;
  int gnl[L_GPT];            // node of the slot, -1: unused
  uint32_t goff[L_GPT];      // byte offset of the element in the operand tile
  const float* gptr[L_GPT];  // its address in chunk 0
  int gk[L_GPT];
#pragma unroll
  for (int j = 0; j < L_GPT; ++j) {
    const int U = warp + j * (L_T / 32);
    const int kc = U & 3, rb = U >> 2;
    const int row = rb * 8 + (lane >> 2), k = kc * 4 + (lane & 3);
    const int nl = row / D, m = row % D;
    gnl[j] = row < TN * D ? nl : -1;
    gk[j] = k;
    goff[j] = (uint32_t)kc * L_ACH + (uint32_t)row * 16 + (uint32_t)(lane & 3) * 4;
    gptr[j] = x + (size_t)(n0 + min(nl, nn - 1)) * pl.ldx + b.xoff + k * D + m;
  }
  auto gload = [&](int c, int j) -> float {
    return (gnl[j] >= 0 && gnl[j] < nn && c * LK + gk[j] < b.K) ? __ldg(gptr[j] + (size_t)c * span) : 0.f;
  };
  float gv[L_GPT];
#pragma unroll
  for (int j = 0; j < L_GPT; ++j) gv[j] = gload(0, j);

  for (int c = 0; c < b.nchunk; ++c) {
    const int st = c & 1;
    if (c >= 2) {   // the MMAs of chunk c-2 have consumed the operand stage and filled its two accumulators
      mbar_wait(&s.mma_done[st], ((c - 2) >> 1) & 1);
      tc_fence_after();
      drain(c - 2);
    }
    if (warp == 0 && elect_one()) {
      mbar_expect_tx(&s.full_b[st], 2 * bpart);
      const float* src = pack + b.pack_off + (size_t)c * (2 * bpart / 4);
      bulk_copy_g2s(s.B[st], src, bpart, &s.full_b[st]);
      bulk_copy_g2s(s.B[st] + L_BPART, src + bpart / 4, bpart, &s.full_b[st]);
    }
    // gather the A chunk: element (node nl, k, m) -> row nl*D+m, column k; lanes run along the contiguous (k, m) index.
    // The values of the NEXT chunk are requested before this chunk's are split and stored (one round trip hidden).
#pragma unroll
    for (int j = 0; j < L_GPT; ++j) {
      const float v = gv[j];
      if (c + 1 < b.nchunk) gv[j] = gload(c + 1, j);
      if (gnl[j] >= 0) {
        float hi, lo;
        split_tf32(v, hi, lo);
        *reinterpret_cast<float*>(s.A[st] + goff[j]) = hi;
        *reinterpret_cast<float*>(s.A[st] + L_APART + goff[j]) = lo;
      }
    }
    if (TN * D < 128) {   // rows of the tile that belong to no node (128 is not a multiple of 2l+1)
      for (int it = threadIdx.x; it < (128 - TN * D) * LK; it += L_T) {
        const int r = TN * D + it / LK, k = it % LK;
        const uint32_t off = (uint32_t)(k >> 2) * L_ACH + (uint32_t)r * 16 + (uint32_t)(k & 3) * 4;
        *reinterpret_cast<float*>(s.A[st] + off) = 0.f;
        *reinterpret_cast<float*>(s.A[st] + L_APART + off) = 0.f;
      }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (warp == 0 && elect_one()) {
      mbar_wait(&s.full_b[st], (c >> 1) & 1);
      tc_fence_after();
      const uint32_t ah = smem_u32(s.A[st]), al = ah + L_APART, bh = smem_u32(s.B[st]), bl = bh + L_BPART;
      const uint32_t bch = (uint32_t)b.Npad * 16;
      const uint32_t dmain = tmem + (uint32_t)(2 * (c & 1) * L_MAXN);
#pragma unroll
      for (int kk = 0; kk < LK / 8; ++kk) {
        const uint64_t dah = make_desc(ah + kk * 2 * L_ACH, L_ACH, 128), dal = make_desc(al + kk * 2 * L_ACH, L_ACH, 128);
        const uint64_t dbh = make_desc(bh + kk * 2 * bch, bch, 128), dbl = make_desc(bl + kk * 2 * bch, bch, 128);
        mma_tf32(dmain + (uint32_t)(kk * L_MAXN), dal, dbh, idesc, 0u);
        mma_tf32(dmain + (uint32_t)(kk * L_MAXN), dah, dbl, idesc, 1u);
        mma_tf32(dmain + (uint32_t)(kk * L_MAXN), dah, dbh, idesc, 1u);
      }
      mma_commit(&s.mma_done[st]);
    }
  }
  // the last two chunks (MMAs retire in order)
  for (int c = max(0, b.nchunk - 2); c < b.nchunk; ++c) {
    mbar_wait(&s.mma_done[c & 1], (c >> 1) & 1);
    tc_fence_after();
    drain(c);
  }
  // transpose the warp's 32x32 block so that the stores run along the output channels of one (node, m) row
  float (*T)[33] = s.T[warp];
#pragma unroll
  for (int j = 0; j < 32; ++j) T[lane][j] = acc[j];
  __syncwarp();
  const int o = 32 * half + lane;
  if (o < b.N) {
    for (int r = 0; r < 32; ++r) {
      const int row = 32 * q + r, nl = row / D, m = row % D;
      if (nl >= nn) break;
      float* p = y + (size_t)(n0 + nl) * pl.ldy + b.yoff + (size_t)(b.o0 + o) * D + m;
      const float val = T[r][lane] * b.scale;
      *p = accumulate ? (*p + val) : val;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// plan: one pass per (block, <=256 output columns); returns false when a block is too small for the tensor-core path
static bool build_lin_plan(const LinDesc& d, int transpose, LinTcPlan& pl, size_t* pack_floats) {
  pl.npass = 0;
  pl.ldx = transpose ? d.out_dim : d.in_dim;
  pl.ldy = transpose ? d.in_dim : d.out_dim;
  size_t off = 0;
  for (int i = 0; i < d.nblk; ++i) {
    const LinBlock& lb = d.blk[i];
    const int K = transpose ? lb.mul_out : lb.mul_in, N = transpose ? lb.mul_in : lb.mul_out;
    if (K < 16 || N < 16) return false;
    const int npass = (N + L_MAXN - 1) / L_MAXN;
    const int per = ((N + npass - 1) / npass + 15) / 16 * 16;
    for (int p = 0; p < npass; ++p) {
      if (pl.npass >= 16) return false;
      LinTcBlock& b = pl.p[pl.npass];
      b.D = lb.d; b.K = K;
      b.o0 = p * per;
      b.N = min(per, N - b.o0);
      b.Npad = (b.N + 15) / 16 * 16;
      b.xoff = transpose ? lb.out_off : lb.in_off;
      b.yoff = transpose ? lb.in_off : lb.out_off;
      b.nchunk = (K + LK - 1) / LK;
      b.scale = lb.scale;
      b.w_off = lb.w_off;
      b.ldw = lb.mul_out;
      b.pack_off = (int)off;
      off += (size_t)b.nchunk * 2 * (LK / 4) * b.Npad * 4;
      ++pl.npass;
    }
  }
  *pack_floats = off;
  return pl.npass > 0;
}

bool irreps_linear_tc_supported(const LinDesc& d, int transpose) {
  LinTcPlan pl;
  size_t n;
  if (d.nblk < 1 || d.nblk > 4) return false;
  return build_lin_plan(d, transpose, pl, &n);
}

int irreps_linear_tc(const LinDesc& d, int n_nodes, const float* x, const float* W, float* y, int transpose, int accumulate,
                     cudaStream_t st) {
  LinTcPlan pl;
  size_t pack_floats;
  B200_REQUIRE(build_lin_plan(d, transpose, pl, &pack_floats), "irreps_linear_tc: unsupported shape");
  if (n_nodes == 0) return B200_OK;
  float* pack;
  keep_async_pool();
  if (cudaMallocAsync((void**)&pack, pack_floats * sizeof(float), st) != cudaSuccess) {
    set_error("irreps_linear_tc: cudaMallocAsync of the operand pack failed");
    return B200_ERR_CUDA;
  }
  linear_prepack_kernel<<<dim3(16, pl.npass), 256, 0, st>>>(pl, W, transpose, pack);
  cudaFuncSetAttribute(irreps_linear_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LinTcSmem));
  int mind = 5;
  for (int p = 0; p < pl.npass; ++p) mind = pl.p[p].D < mind ? pl.p[p].D : mind;
  int maxd = 1;
  for (int p = 0; p < pl.npass; ++p) maxd = pl.p[p].D > maxd ? pl.p[p].D : maxd;
  const int tn = 128 / maxd;   // the widest irrep has the fewest nodes per tile; surplus CTAs of the others exit
  dim3 grid((n_nodes + tn - 1) / tn, pl.npass);
  irreps_linear_tc_kernel<<<grid, L_T, sizeof(LinTcSmem), st>>>(pl, n_nodes, x, pack, y, accumulate);
  cudaFreeAsync(pack, st);
  return check_launch("irreps_linear_tc", 2);
}

}  // namespace b200
