// nnd_grad.cuh -- included by nnd_kernels.cu inside its anonymous namespace.
//
// Backward of nnd (my_lib.cpp:91-130 / nnd_cuda.cu:164-183).  Per row r of a term with neighbour j2 = idx[r]:
//     g = fl(2 * graddist[r]);  t_c = fl(g * fl(q[r][c] - target[j2][c]))         (each op separately rounded)
//     grad_q[r][c] += t_c ;  grad_target[j2][c] -= t_c
// The reference CUDA kernel issues 6 scalar atomicAdd per row into caller-zeroed buffers.  Here the full backward is
// two launches and no memset:
//   nnd_grad_own_kernel      every gradient row receives exactly ONE own-row contribution (row r of term 1 -> gradxyz1[r],
//                            row r of term 2 -> gradxyz2[r]), so these are PLAIN coalesced 16-byte stores that overwrite
//                            both outputs (the zero-init of my_lib.cpp:91-94 is implied);
//   nnd_grad_scatter_kernel  the neighbour-side updates land anywhere in the other cloud: warp-aggregated
//                            (match.any on the neighbour row) vector reductions, 2 per row instead of 3 scalar ones
//                            (a [*,3] row starts 8-byte aligned at even or odd float offset: {x,y}+z or x+{y,z}).
// Rows are staged through shared memory so that the stride-3 layout is read and written with 16-byte transactions.
// The one-term entry point used by the multi-GPU path keeps the accumulate-into-caller-zeroed-buffers contract and
// uses the scatter kernel with OWN = true (own-row side as vector reductions as well).
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

__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
  asm volatile("red.global.v2.f32.add [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

// 3 floats into row `row` of a [*,3] array with two reductions; needs an 8-byte aligned base pointer.
__device__ __forceinline__ void red_add_row(float* base, long long row, float x, float y, float z) {
  float* p = base + row * 3;
  if ((row & 1) == 0) {
    red_add_v2(p, x, y);
    atomicAdd(p + 2, z);
  } else {
    atomicAdd(p, x);
    red_add_v2(p + 1, y, z);
  }
}

// Copies rows [r0, r0 + count) of a [*,3] array into shared memory with 16-byte loads when the base is aligned.
__device__ __forceinline__ void stage_rows(const float* src, long long r0, int count, float* sm, bool aligned) {
  const float* g = src + r0 * 3;
  const int nfl = count * 3;
  if (aligned) {  // r0 is a multiple of 256, so g is 16-byte aligned whenever src is
    const int nv = nfl >> 2;
    for (int i = threadIdx.x; i < nv; i += kGradThreads)
      reinterpret_cast<float4*>(sm)[i] = __ldg(reinterpret_cast<const float4*>(g) + i);
    for (int i = (nv << 2) + threadIdx.x; i < nfl; i += kGradThreads) sm[i] = __ldg(g + i);
  } else {
    for (int i = threadIdx.x; i < nfl; i += kGradThreads) sm[i] = __ldg(g + i);
  }
}

// t = fl(2*gd) * fl(q - target[idx]) per component, exactly the oracle's rounding sequence.
__device__ __forceinline__ void grad_row(const GradTerm& t, long long r, const float* qrow, float& tx, float& ty,
                                         float& tz, long long& trow) {
  const long long batch = r / t.nq;
  trow = batch * t.nt + t.idx[r];
  const float* b = t.t + trow * 3;
  const float g = __fmul_rn(t.gd[r], 2.0f);
  tx = __fmul_rn(g, __fsub_rn(qrow[0], __ldg(b + 0)));
  ty = __fmul_rn(g, __fsub_rn(qrow[1], __ldg(b + 1)));
  tz = __fmul_rn(g, __fsub_rn(qrow[2], __ldg(b + 2)));
}

// Own-row side: plain stores.  One block iteration = 256 consecutive rows of one term.
__global__ void __launch_bounds__(kGradThreads) nnd_grad_own_kernel(GradParams p) {
  __shared__ __align__(16) float sm[kGradThreads * 3];
  for (int ti = 0; ti < p.nterms; ++ti) {
    const GradTerm t = p.t[ti];
    const bool in_al = (reinterpret_cast<uintptr_t>(t.q) & 15u) == 0;
    const bool out_al = (reinterpret_cast<uintptr_t>(t.gq) & 15u) == 0;
    const long long nblocks = (t.rows + kGradThreads - 1) / kGradThreads;
    for (long long blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
      const long long r0 = blk * kGradThreads;
      const int count = (int)(t.rows - r0 < kGradThreads ? t.rows - r0 : kGradThreads);
      stage_rows(t.q, r0, count, sm, in_al);
      __syncthreads();
      float tx = 0.f, ty = 0.f, tz = 0.f;
      if ((int)threadIdx.x < count) {
        long long trow;
        grad_row(t, r0 + threadIdx.x, sm + 3 * threadIdx.x, tx, ty, tz, trow);
      }
      __syncthreads();  // everyone has read its row: reuse the buffer for the results
      sm[3 * threadIdx.x + 0] = tx;
      sm[3 * threadIdx.x + 1] = ty;
      sm[3 * threadIdx.x + 2] = tz;
      __syncthreads();
      float* out = t.gq + r0 * 3;
      const int nfl = count * 3;
      if (out_al) {
        const int nv = nfl >> 2;
        for (int i = threadIdx.x; i < nv; i += kGradThreads)
          reinterpret_cast<float4*>(out)[i] = reinterpret_cast<const float4*>(sm)[i];
        for (int i = (nv << 2) + threadIdx.x; i < nfl; i += kGradThreads) out[i] = sm[i];
      } else {
        for (int i = threadIdx.x; i < nfl; i += kGradThreads) out[i] = sm[i];
      }
      __syncthreads();
    }
  }
}

// Neighbour side (and, with OWN, the own-row side as reductions too).  Lanes of a warp that hit the same target row
// are combined first (match.any), one lane issues the reductions.
template <bool OWN>
__global__ void __launch_bounds__(kGradThreads) nnd_grad_scatter_kernel(GradParams p) {
  __shared__ __align__(16) float sm[kGradThreads * 3];
  const unsigned lane = threadIdx.x & 31u;
  for (int ti = 0; ti < p.nterms; ++ti) {
    const GradTerm t = p.t[ti];
    const bool in_al = (reinterpret_cast<uintptr_t>(t.q) & 15u) == 0;
    const bool gt_al = (reinterpret_cast<uintptr_t>(t.gt) & 7u) == 0;
    const bool gq_al = (reinterpret_cast<uintptr_t>(t.gq) & 7u) == 0;
    const long long nblocks = (t.rows + kGradThreads - 1) / kGradThreads;
    for (long long blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
      const long long r0 = blk * kGradThreads;
      const int count = (int)(t.rows - r0 < kGradThreads ? t.rows - r0 : kGradThreads);
      stage_rows(t.q, r0, count, sm, in_al);
      __syncthreads();
      const bool live = (int)threadIdx.x < count;
      float tx = 0.f, ty = 0.f, tz = 0.f;
      long long trow = -1 - (long long)lane;  // dead lanes never match anybody
      if (live) {
        grad_row(t, r0 + threadIdx.x, sm + 3 * threadIdx.x, tx, ty, tz, trow);
        if (OWN) {
          const long long r = r0 + threadIdx.x;
          if (gq_al) red_add_row(t.gq, r, tx, ty, tz);
          else {
            atomicAdd(t.gq + r * 3 + 0, tx);
            atomicAdd(t.gq + r * 3 + 1, ty);
            atomicAdd(t.gq + r * 3 + 2, tz);
          }
        }
      }
      __syncthreads();  // the staging buffer is free again; whole warps are converged for match/shfl
      const unsigned peers = __match_any_sync(0xffffffffu, trow);
      const unsigned leader = __ffs(peers) - 1;
      float sx = tx, sy = ty, sz = tz;
      if (peers != (1u << lane)) {  // somebody shares my target row: every member sums the group
        sx = sy = sz = 0.f;
        unsigned rest = peers;
        while (rest) {  // same trip count for all lanes of a group; other groups diverge harmlessly
          const int src = __ffs(rest) - 1;
          rest &= rest - 1;
          sx += __shfl_sync(peers, tx, src);
          sy += __shfl_sync(peers, ty, src);
          sz += __shfl_sync(peers, tz, src);
        }
      }
      if (live && lane == leader) {
        if (gt_al) red_add_row(t.gt, trow, -sx, -sy, -sz);
        else {
          atomicAdd(t.gt + trow * 3 + 0, -sx);
          atomicAdd(t.gt + trow * 3 + 1, -sy);
          atomicAdd(t.gt + trow * 3 + 2, -sz);
        }
      }
    }
  }
}
