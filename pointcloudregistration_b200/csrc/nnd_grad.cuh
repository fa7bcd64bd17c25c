// nnd_grad.cuh -- included by nnd_kernels.cu inside its anonymous namespace.
//
// Backward of nnd (my_lib.cpp:91-130 / nnd_cuda.cu:164-183).  Per row r of a term with neighbour j2 = idx[r]:
//     g = fl(2 * graddist[r]);  t_c = fl(g * fl(q[r][c] - target[j2][c]))         (each op separately rounded)
//     grad_q[r][c] += t_c ;  grad_target[j2][c] -= t_c
// The reference CUDA kernel issues 6 scalar atomicAdd per row with one row in flight per thread.  Measured on B200
// (profiles/r01_bwd_variants.txt, r01_ncu_grad_2kernel_summary.txt) the op is bound by the NUMBER of L2 requests
// (random 12-byte gathers and reductions, ~150-200 G requests/s), not by bytes and not by latency: keeping 2/4/8 rows
// in flight per thread (NND_GRAD_ROWS) is slower than 1, and a two-kernel variant with plain own-row stores that
// gathers twice is slower too.  Hence:
//   * a [*,3] row is fetched and reduced with TWO requests instead of three: it starts 8-byte aligned at an even
//     float offset ({x,y} + z) or at an odd one (x + {y,z})  -> LDG.64 + LDG.32, REDG.F32x2 + REDG.F32;
//   * lanes of a warp that hit the same neighbour row are combined first (match.any) so clustered clouds do not
//     serialise on one address.
// Gradients are accumulated into zero-initialised buffers; nnd_backward_cuda_f32 zeroes them on the same stream.
//
// Global fp32 reductions flush denormal sums to zero on this part (SASS: REDG.E.ADD.F32.FTZ.RN and
// REDG.E.ADD.F32x2.FTZ.RN alike, i.e. the reference's scalar atomicAdd does the same).  The effect is far below the
// 1e-5 gradient contract (it needs |grad| < 1.2e-38).

struct GradTerm {
  const float* q;      // [b][nq][3]  rows of this term
  const float* t;      // [b][nt][3]  the other cloud
  const float* gd;     // [b][nq]
  const int32_t* idx;  // [b][nq]
  float* gq;           // [b][nq][3]
  float* gt;           // [b][nt][3]
  int nq, nt;
  long long rows;  // b * nq
};
struct GradParams {
  GradTerm t[2];
  int nterms;
};

constexpr int kGradThreads = 256;
#ifndef NND_GRAD_ROWS
#define NND_GRAD_ROWS 1
#endif
constexpr int kGradRows = NND_GRAD_ROWS;  // rows in flight per thread

__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
  asm volatile("red.global.v2.f32.add [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

// 3 floats into row `row` of a [*,3] array with two reductions (8-byte aligned base) or three (otherwise).
__device__ __forceinline__ void red_add_row(float* base, long long row, float x, float y, float z, bool al8) {
  float* p = base + row * 3;
  if (!al8) {
    atomicAdd(p, x);
    atomicAdd(p + 1, y);
    atomicAdd(p + 2, z);
  } else if ((row & 1) == 0) {
    red_add_v2(p, x, y);
    atomicAdd(p + 2, z);
  } else {
    atomicAdd(p, x);
    red_add_v2(p + 1, y, z);
  }
}

// Row `row` of a [*,3] array with two loads (8-byte aligned base) or three.
__device__ __forceinline__ void load_row(const float* base, long long row, float& x, float& y, float& z, bool al8) {
  const float* p = base + row * 3;
  if (!al8) {
    x = __ldg(p);
    y = __ldg(p + 1);
    z = __ldg(p + 2);
  } else if ((row & 1) == 0) {
    const float2 v = __ldg(reinterpret_cast<const float2*>(p));
    x = v.x;
    y = v.y;
    z = __ldg(p + 2);
  } else {
    x = __ldg(p);
    const float2 v = __ldg(reinterpret_cast<const float2*>(p + 1));
    y = v.x;
    z = v.y;
  }
}

__global__ void __launch_bounds__(kGradThreads) nnd_grad_kernel(GradParams p) {
  const unsigned lane = threadIdx.x & 31u;
  constexpr int SPAN = kGradThreads * kGradRows;
  for (int ti = 0; ti < p.nterms; ++ti) {
    const GradTerm t = p.t[ti];
    const bool q_al = (reinterpret_cast<uintptr_t>(t.q) & 7u) == 0;
    const bool t_al = (reinterpret_cast<uintptr_t>(t.t) & 7u) == 0;
    const bool gq_al = (reinterpret_cast<uintptr_t>(t.gq) & 7u) == 0;
    const bool gt_al = (reinterpret_cast<uintptr_t>(t.gt) & 7u) == 0;
    const long long nspans = (t.rows + SPAN - 1) / SPAN;
    for (long long sp = blockIdx.x; sp < nspans; sp += gridDim.x) {
      const long long r0 = sp * SPAN + threadIdx.x;
      long long trow[kGradRows];
      float g[kGradRows], ax[kGradRows], ay[kGradRows], az[kGradRows];
      float tx[kGradRows], ty[kGradRows], tz[kGradRows];
#pragma unroll
      for (int k = 0; k < kGradRows; ++k) {  // independent loads first: neighbour index, upstream gradient, own row
        const long long r = r0 + k * kGradThreads;
        trow[k] = -1 - (long long)lane;  // dead lanes never match anybody
        g[k] = ax[k] = ay[k] = az[k] = 0.f;
        if (r < t.rows) {
          trow[k] = (r / t.nq) * t.nt + __ldg(t.idx + r);
          g[k] = __fmul_rn(__ldg(t.gd + r), 2.0f);
          load_row(t.q, r, ax[k], ay[k], az[k], q_al);
        }
      }
#pragma unroll
      for (int k = 0; k < kGradRows; ++k) {  // dependent gathers, all in flight together
        tx[k] = ty[k] = tz[k] = 0.f;
        if (trow[k] >= 0) {
          float bx, by, bz;
          load_row(t.t, trow[k], bx, by, bz, t_al);
          tx[k] = __fmul_rn(g[k], __fsub_rn(ax[k], bx));
          ty[k] = __fmul_rn(g[k], __fsub_rn(ay[k], by));
          tz[k] = __fmul_rn(g[k], __fsub_rn(az[k], bz));
        }
      }
#pragma unroll
      for (int k = 0; k < kGradRows; ++k) {
        const bool live = trow[k] >= 0;
        if (live) red_add_row(t.gq, r0 + k * kGradThreads, tx[k], ty[k], tz[k], gq_al);
        // whole warps reach this point together (the loop bounds are block-uniform)
        const unsigned peers = __match_any_sync(0xffffffffu, trow[k]);
        const unsigned leader = __ffs(peers) - 1;
        float sx = tx[k], sy = ty[k], sz = tz[k];
        if (peers != (1u << lane)) {  // somebody shares my neighbour row: every member sums the group
          sx = sy = sz = 0.f;
          unsigned rest = peers;
          while (rest) {  // same trip count for all lanes of a group; other groups diverge harmlessly
            const int src = __ffs(rest) - 1;
            rest &= rest - 1;
            sx += __shfl_sync(peers, tx[k], src);
            sy += __shfl_sync(peers, ty[k], src);
            sz += __shfl_sync(peers, tz[k], src);
          }
        }
        if (live && lane == leader) red_add_row(t.gt, trow[k], -sx, -sy, -sz, gt_al);
      }
    }
  }
}

// ----------------------------------------------------------------------------------------------------------------
// Fused truncated-chamfer epilogue (c2p-net/deformationpyramid/model/loss.py:150-215 with point_reduction = batch_reduction
// = "mean" and no weights, the form registration.py:236 uses):
//     loss = scale1 * sum_{dist1 < trunc} dist1 + scale2 * sum_{dist2 < trunc} dist2
//     graddist1 = dist1 < trunc ? scale1 : 0        (d loss / d dist1; the mask is reused by the backward)
// One launch instead of ~10 elementwise/reduction launches.  Deterministic: per-block partial sums are combined in
// block order by the last block to finish.
struct ReduceParams {
  const float* dist[2];
  float* graddist[2];  // may be null
  long long count[2];
  float scale[2];
  float trunc;
  float* loss;       // [1]
  double* partial;   // [gridDim.x]
  unsigned* ticket;  // [1], zeroed by the launcher on the same stream when there is more than one block
};

__global__ void __launch_bounds__(256) nnd_chamfer_reduce_kernel(ReduceParams p) {
  __shared__ double warp_sum[8];
  __shared__ bool last;
  double acc = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    double a = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.count[k]; i += stride) {
      const float d = p.dist[k][i];
      const bool keep = !(d >= p.trunc);  // loss.py:151-154: cham[cham >= trunc] = 0
      if (keep) a += (double)d;
      if (p.graddist[k] != nullptr) p.graddist[k][i] = keep ? p.scale[k] : 0.0f;
    }
    acc += a * (double)p.scale[k];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 8; ++w) s += warp_sum[w];
    p.partial[blockIdx.x] = s;
    __threadfence();
    last = gridDim.x == 1 || atomicAdd(p.ticket, 1u) == gridDim.x - 1;  // one block: no ticket at all
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double s = 0.0;
    for (unsigned b = 0; b < gridDim.x; ++b) s += reinterpret_cast<volatile double*>(p.partial)[b];
    *p.loss = (float)s;
  }
}
