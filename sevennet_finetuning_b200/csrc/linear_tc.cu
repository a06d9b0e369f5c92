// Per-node irreps linear on the tcgen05 tensor cores (3xTF32, fp32 accumulation in TMEM).
//
// Reference: sevenn/nn/linear.py:46-49 (IrrepsLinear -> e3nn o3.Linear), sevenn/nn/self_connection.py:60-62.
//   y[n, o, m] = scale * sum_k x[n, k, m] W[k, o]     per l-block, scale = 1/sqrt(mul_in)
// GEMM view: rows = (node, m) pairs, K = mul_in, N = mul_out.  One CTA owns 128 rows of one block:
//   A = x tile (rows, K-major): the raw fp32 rows of a 16-k chunk ([mul, 2l+1], m fastest: 16 (2l+1) contiguous floats
//       per node) arrive by cp.async two chunks ahead; the eight worker warps read them back in (row, k) order, split
//       them into tf32 hi/lo and write the canonical operand layout (two operand stages),
//   B = W^T (rows = output channel, K-major): prepacked hi/lo per 16-k chunk, streamed with cp.async.bulk by the issuer
//       warp, whose elected thread also issues the MMAs (the workers never wait for it),
//   D = 128 lanes (rows) x N columns in TMEM (N <= 64 per pass, four accumulators per CTA, two CTAs per SM); a CTA
//       streams the passes of its tile one after the other,
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
constexpr int L_T = 256;               // worker threads
constexpr int L_TT = L_T + 32;         // + the issuer warp
constexpr int L_NS = 3;                // raw row staging slots (two chunks in flight)
constexpr int L_SFLOATS = 128 * (LK + 4);   // floats of one slot: TN nodes x (16 (2l+1) + 4) <= 2560

struct LinTcBlock {
  int D, K, N, Npad;          // 2l+1, contraction length, real / padded (multiple of 16) output columns of this pass
  int xoff, yoff, o0;         // block offsets in the input / output rows; first output channel of the pass
  int pack_off;               // float offset of the pass's operand pack
  int w_off, ldw;             // block offset in the flat weight vector, its row length (mul_out)
  int nchunk;                 // ceil(K / 16)
  int Nblk;                   // output channels of the whole l-block (row length of a component in a component-major block)
  float scale;
};
struct LinTcPlan {
  int npass;
  int ldx, ldy;
  int x_cm, y_cm;             // component-major blocks ([2l+1][mul] instead of [mul][2l+1]) on the input / output side
  int nblk, blk_first[16], blk_npass[16];   // groups of consecutive passes of one l-block: one CTA runs a group back to back
  LinTcBlock p[16];
};

struct LinTcSmem {
  uint8_t A[2][2 * L_APART];
  uint8_t B[2][2 * L_BPART];
  float S[L_NS][L_SFLOATS];    // raw fp32 rows of the coming chunks, filled by cp.async
  float T[L_T / 32][32][33];   // per-warp transposition scratch of the pass epilogues
  uint64_t full_b[2], mma_done[2], a_full[2];
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
template <int D>
__device__ __forceinline__ void lin_tile(LinTcSmem& s, const LinTcPlan& pl, int first, int npass, int n_nodes,
                                         const float* __restrict__ x, const float* __restrict__ pack,
                                         float* __restrict__ y, int accumulate) {
  constexpr int TN = 128 / D;                    // nodes per tile
  const int n0 = blockIdx.x * TN;
  if (n0 >= n_nodes) return;
  const int nn = min(TN, n_nodes - n0);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, half = warp >> 2;      // TMEM lane quarter, 32-column half of the pass
  const LinTcBlock& b = pl.p[first];             // K, offsets and chunk count are the block's; N, o0, pack_off the pass's
  const int nchunk = b.nchunk, G = npass * nchunk;   // the CTA's chunk stream: pass after pass, no pipeline refill between

  if (warp == 0) tmem_alloc(&s.tmem_base, 256);
  if (threadIdx.x == 32) {
    for (int k = 0; k < 2; ++k) { mbar_init(&s.full_b[k], 1); mbar_init(&s.mma_done[k], 1); mbar_init(&s.a_full[k], L_T / 32); }
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;
  // Warp 8 is the issuer: one elected thread requests the weight chunks, waits for the operand stage the eight worker warps
  // fill, and issues the MMAs.  The workers never wait for it (no CTA-wide barrier in the loop): a cycle-counter trace of the
  // first version, where warp 0 both worked and issued, showed the whole CTA running at warp 0's pace - its own gather and
  // drain (990 cycles) plus the wait for the weights and the issue of six MMAs behind the other CTA's (710 cycles) per chunk,
  // with the other seven warps parked at the barrier for 840 of every 1850 cycles.
  if (warp == L_T / 32) {
    if (elect_one()) {
      for (int g = 0; g < G; ++g) {
        const int st = g & 1, ps = g / nchunk, c = g - ps * nchunk;
        const LinTcBlock& bp = pl.p[first + ps];
        const uint32_t bpart = (uint32_t)(LK / 4) * bp.Npad * 16;   // bytes of one part (hi or lo) of a B chunk
        if (g >= 2) mbar_wait(&s.mma_done[st], ((g - 2) >> 1) & 1);   // the weight stage is free
        mbar_expect_tx(&s.full_b[st], 2 * bpart);
        const float* src = pack + bp.pack_off + (size_t)c * (2 * bpart / 4);
        bulk_copy_g2s(s.B[st], src, bpart, &s.full_b[st]);
        bulk_copy_g2s(s.B[st] + L_BPART, src + bpart / 4, bpart, &s.full_b[st]);
        mbar_wait(&s.a_full[st], (g >> 1) & 1);    // operand written, accumulators of this stage drained
        mbar_wait(&s.full_b[st], (g >> 1) & 1);
        tc_fence_after();
        const uint32_t idesc = make_idesc_tf32(128, bp.Npad);
        const uint32_t ah = smem_u32(s.A[st]), al = ah + L_APART, bh = smem_u32(s.B[st]), bl = bh + L_BPART;
        const uint32_t bch = (uint32_t)bp.Npad * 16;
        const uint32_t dmain = tmem + (uint32_t)(2 * st * L_MAXN);
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
    __syncwarp();
    tc_fence_before();
    __syncthreads();      // the workers' last barrier (before the TMEM is released)
    return;
  }
  const uint32_t tlane = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(32 * half);
  float acc[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = 0.f;
  float (*T)[33] = s.T[warp];
  // add the two accumulators of stream position g (its MMAs have retired) into the registers; at the last chunk of a
  // pass, write the pass out: the warp transposes its 32x32 block through shared memory so that its stores run along the
  // output channels of one (node, m) row
  auto drain = [&](int g) {
    const int ps = g / nchunk;
    const LinTcBlock& bp = pl.p[first + ps];
    if (32 * half < bp.Npad) {
      float v[32];
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        tmem_ld32(tlane + (uint32_t)((2 * (g & 1) + kk) * L_MAXN), v);
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] += v[j];
      }
    }
    tc_fence_before();
    if (g - ps * nchunk == nchunk - 1) {
#pragma unroll
      for (int j = 0; j < 32; ++j) { T[lane][j] = acc[j]; acc[j] = 0.f; }
      __syncwarp();
      const int o = 32 * half + lane;
      if (o < bp.N) {
        for (int r = 0; r < 32; ++r) {
          const int orow = 32 * q + r, nl = orow / D, m = orow % D;
          if (nl >= nn) break;
          float* p = y + (size_t)(n0 + nl) * pl.ldy + bp.yoff +
                     (pl.y_cm ? (size_t)m * bp.Nblk + (bp.o0 + o) : (size_t)(bp.o0 + o) * D + m);
          const float val = T[r][lane] * bp.scale;
          *p = accumulate ? (*p + val) : val;
        }
      }
      __syncwarp();
    }
  };

  // Operand A.  (1) Raw rows: the chunk's 16 (2l+1) floats of a node are contiguous, 4 (2l+1) pieces of 16 bytes; the
  // tile's <= 512 pieces are two cp.async per worker thread, issued two chunks ahead into a 3-slot ring (node stride
  // padded by 4 floats so that the reads below spread over the banks).  Pieces past the block's K are zero-filled
  // (src-size operand); nodes past the end of a short last tile re-read the last valid node (tile rows are independent in
  // the product and those rows are never stored).  Direct 4-byte loads of (row, k) elements, the former scheme, cost the
  // L1 data pipe one sector per lane and load: 128M sector reads for 39M sectors of data, 72 % of the L1 cycles
  // (profiles/r01q_linear_tc_raw.csv); the pieces move every sector once.  Every pass of the block re-reads the tile's rows
  // (L1 / L2 hits after the first).
  // (2) Split: thread -> one row (node, m) and 8 of the 16 k; a warp covers 32 consecutive rows, so its 16-byte operand
  // stores are one contiguous 512-byte run.
  constexpr int RS = LK * D + 4;                 // floats per node in a slot
  constexpr int PPN = 4 * D;                     // 16-byte pieces per node and chunk
  const float* psrc[2];
  uint32_t pdst[2];
  int pfl[2];                                    // float offset of the piece inside the node's chunk (huge: no piece)
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int id = threadIdx.x + u * L_T, nl = id / PPN, p = id % PPN;
    // component-major input: piece p = (component p / 4, quarter p % 4 of its 16 k); the slot then holds [m][16 k]
    pfl[u] = id < TN * PPN ? (pl.x_cm ? (p & 3) * 4 : p * 4) : (1 << 20);
    psrc[u] = x + (size_t)(n0 + min(nl, nn - 1)) * pl.ldx + b.xoff + (pl.x_cm ? (p >> 2) * b.K + (p & 3) * 4 : p * 4);
    pdst[u] = smem_u32(&s.S[0][0]) + (uint32_t)(nl * RS + p * 4) * 4;
  }
  const int cstep = pl.x_cm ? LK : LK * D;       // floats between the chunks of a piece
  auto stage = [&](int g) {                      // request stream position g (nothing past the end; the group is still committed)
    if (g < G) {
      const int c = g % nchunk;
      const int kfl = min(LK, b.K - c * LK) * (pl.x_cm ? 1 : D);   // valid floats of a node's chunk (of one component of it)
      const uint32_t slot = (uint32_t)(g % L_NS) * (L_SFLOATS * 4);
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int nb = min(16, max(0, (kfl - pfl[u]) * 4));
        if (pfl[u] < (1 << 20))
          asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(pdst[u] + slot), "l"(psrc[u] + (size_t)c * cstep), "r"(nb) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const int row = threadIdx.x & 127, kp = threadIdx.x >> 7;
  // this thread's first element inside a slot, and the distance between its k's
  const int soff = min(row / D, TN - 1) * RS + (pl.x_cm ? (row % D) * LK + kp * 8 : kp * 8 * D + row % D);
  const int sk = pl.x_cm ? 1 : D;
  const uint32_t aoff = (uint32_t)(2 * kp) * L_ACH + (uint32_t)row * 16;
  stage(0);
  stage(1);
  asm volatile("cp.async.wait_group 1;" ::: "memory");
  bar_sync(1, L_T);

  for (int g = 0; g < G; ++g) {
    const int st = g & 1;
    if (g >= 2) {   // the MMAs of position g-2 have consumed the operand stage and filled its two accumulators
      mbar_wait(&s.mma_done[st], ((g - 2) >> 1) & 1);
      tc_fence_after();
      drain(g - 2);
    }
    stage(g + L_NS - 1);   // into the slot position g-1 was read from (all workers are past that iteration's barrier)
    {
      const float* sp = &s.S[g % L_NS][soff];
      float4 v0, v1, h0, l0, h1, l1;
      if (D == 1 || sk == 1) {
        v0 = *reinterpret_cast<const float4*>(sp);
        v1 = *reinterpret_cast<const float4*>(sp + 4);
      } else {
        v0 = make_float4(sp[0], sp[D], sp[2 * D], sp[3 * D]);
        v1 = make_float4(sp[4 * D], sp[5 * D], sp[6 * D], sp[7 * D]);
      }
      split4(v0, h0, l0);
      split4(v1, h1, l1);
      uint8_t* a = s.A[st] + aoff;
      *reinterpret_cast<float4*>(a) = h0;
      *reinterpret_cast<float4*>(a + L_ACH) = h1;
      *reinterpret_cast<float4*>(a + L_APART) = l0;
      *reinterpret_cast<float4*>(a + L_APART + L_ACH) = l1;
    }
    fence_proxy_async();
    tc_fence_before();
    warp_arrive(&s.a_full[st]);
    asm volatile("cp.async.wait_group 1;" ::: "memory");   // own pieces of position g+1 have landed ...
    bar_sync(1, L_T);                                       // ... and every worker's: the slot may be read
  }
  // the last two positions (MMAs retire in order)
  for (int g = max(0, G - 2); g < G; ++g) {
    mbar_wait(&s.mma_done[g & 1], (g >> 1) & 1);
    tc_fence_after();
    drain(g);
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

__global__ void __launch_bounds__(L_TT, 2)
irreps_linear_tc_kernel(LinTcPlan pl, int n_nodes, const float* __restrict__ x, const float* __restrict__ pack,
                        float* __restrict__ y, int accumulate) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  LinTcSmem& s = *reinterpret_cast<LinTcSmem*>(smem_raw);
  const int first = pl.blk_first[blockIdx.y], npass = pl.blk_npass[blockIdx.y];
  switch (pl.p[first].D) {   // 2l+1 as a compile-time constant: row <-> (node, m) without integer division
    case 1: lin_tile<1>(s, pl, first, npass, n_nodes, x, pack, y, accumulate); break;
    case 3: lin_tile<3>(s, pl, first, npass, n_nodes, x, pack, y, accumulate); break;
    default: lin_tile<5>(s, pl, first, npass, n_nodes, x, pack, y, accumulate); break;
  }
}

// plan: one pass per (block, <=256 output columns); returns false when a block is too small for the tensor-core path
static bool build_lin_plan(const LinDesc& d, int transpose, int n_nodes, LinTcPlan& pl, size_t* pack_floats, int layout = 0) {
  pl.x_cm = layout & 1;
  pl.y_cm = (layout >> 1) & 1;
  pl.npass = 0;
  pl.nblk = 0;
  pl.ldx = transpose ? d.out_dim : d.in_dim;
  pl.ldy = transpose ? d.in_dim : d.out_dim;
  size_t off = 0;
  for (int i = 0; i < d.nblk; ++i) {
    const LinBlock& lb = d.blk[i];
    const int K = transpose ? lb.mul_out : lb.mul_in, N = transpose ? lb.mul_in : lb.mul_out;
    if (K < 16 || N < 16) return false;
    if ((pl.ldx & 3) || ((transpose ? lb.out_off : lb.in_off) & 3)) return false;   // 16-byte cp.async pieces of the rows
    if (pl.x_cm && (K & 3)) return false;   // component rows of a component-major block start on 16-byte boundaries
    const int npass = (N + L_MAXN - 1) / L_MAXN;
    // One CTA streams several passes of its tile (no TMEM allocation, pipeline refill and barrier setup per pass: with
    // K = 32 a pass is two chunks long) as long as that leaves at least two CTAs per resident slot; small inputs (the
    // 3,200-node batches of fine-tuning) keep one pass per CTA and their parallelism.
    const int tiles = (n_nodes + 128 / lb.d - 1) / (128 / lb.d);
    const long long want = (long long)tiles * npass / (2 * 2 * 148);
    const int gp = (int)(want < 1 ? 1 : (want > npass ? npass : want));
    for (int p0 = 0; p0 < npass; p0 += gp) {
      if (pl.nblk >= 16) return false;
      pl.blk_first[pl.nblk] = pl.npass + p0;
      pl.blk_npass[pl.nblk] = gp < npass - p0 ? gp : npass - p0;
      ++pl.nblk;
    }
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
      b.Nblk = N;
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

bool irreps_linear_tc_supported(const LinDesc& d, int transpose, int layout) {
  LinTcPlan pl;
  size_t n;
  if (d.nblk < 1 || d.nblk > 4) return false;
  return build_lin_plan(d, transpose, 1 << 30, pl, &n, layout);
}

// pack_cache (optional): where the caller keeps the hi / lo operand pack of THIS (W, transpose) between calls - the weights
// of an inference engine are constants, splitting them again in every call is 30 launches per force call for nothing.  The
// pack layout does not depend on n_nodes.  nullptr: the pack lives for the call (stream-ordered allocation).
int irreps_linear_tc(const LinDesc& d, int n_nodes, const float* x, const float* W, float* y, int transpose, int accumulate,
                     cudaStream_t st, int layout, float** pack_cache) {
  LinTcPlan pl;
  size_t pack_floats;
  B200_REQUIRE(build_lin_plan(d, transpose, n_nodes, pl, &pack_floats, layout), "irreps_linear_tc: unsupported shape");
  if (n_nodes == 0) return B200_OK;
  float* pack;
  bool fresh = true;
  if (pack_cache && *pack_cache) {
    pack = *pack_cache;
    fresh = false;
  } else if (pack_cache) {
    if (cudaMalloc((void**)&pack, pack_floats * sizeof(float)) != cudaSuccess) {
      set_error("irreps_linear_tc: cudaMalloc of the operand pack failed");
      return B200_ERR_CUDA;
    }
    *pack_cache = pack;
  } else {
    keep_async_pool();
    if (cudaMallocAsync((void**)&pack, pack_floats * sizeof(float), st) != cudaSuccess) {
      set_error("irreps_linear_tc: cudaMallocAsync of the operand pack failed");
      return B200_ERR_CUDA;
    }
  }
  if (fresh) linear_prepack_kernel<<<dim3(16, pl.npass), 256, 0, st>>>(pl, W, transpose, pack);
  cudaFuncSetAttribute(irreps_linear_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LinTcSmem));
  int mind = 5;
  for (int p = 0; p < pl.npass; ++p) mind = pl.p[p].D < mind ? pl.p[p].D : mind;
  int maxd = 1;
  for (int p = 0; p < pl.npass; ++p) maxd = pl.p[p].D > maxd ? pl.p[p].D : maxd;
  const int tn = 128 / maxd;   // the widest irrep has the fewest nodes per tile; surplus CTAs of the others exit
  dim3 grid((n_nodes + tn - 1) / tn, pl.nblk);
  irreps_linear_tc_kernel<<<grid, L_TT, sizeof(LinTcSmem), st>>>(pl, n_nodes, x, pack, y, accumulate);
  if (!pack_cache) cudaFreeAsync(pack, st);
  return check_launch("irreps_linear_tc", fresh ? 2 : 1);
}

}  // namespace b200
