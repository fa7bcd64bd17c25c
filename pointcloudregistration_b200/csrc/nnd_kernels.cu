// nnd_kernels.cu -- hand-written sm_100a kernels + C ABI for the brute-force 3-D
// nearest-neighbour / chamfer ("nnd") hot path.
//
// Replaces (reference file:line, semantics only -- the design is new):
//   NmDistanceKernel      dip/torch-nndistance/src/nnd_cuda.cu:8-130    -> nnd_prepare_kernel + nnd_scan_kernel + nnd_finalize_kernel
//   NmDistanceGradKernel  dip/torch-nndistance/src/nnd_cuda.cu:164-183  -> nnd_grad_kernel
//   launchers             dip/torch-nndistance/src/nnd_cuda.cu:132-162,185-222 -> extern "C" entry points at the bottom
// Numerics follow the reference CPU path dip/torch-nndistance/src/my_lib.cpp:3-25
// (d = fl(fl(fl(dx*dx)+fl(dy*dy))+fl(dz*dz)), strict <, lowest index on ties), NOT
// the FMA-contracted result nvcc produces from nnd_cuda.cu:28-31.
//
// Forward design (B200), three paths with bit-identical results:
//   * nnd_prepare_kernel rewrites each [n,3] cloud once into structure-of-arrays tiles  x[TT] | y[TT] | z[TT]
//     (last tile padded by repeating the last point: a duplicate at a higher index can never win the tie-break).
//   * clouds below 8192 points: nnd_scan_kernel (this file), one scan per direction.  Persistent CTAs over the
//     linear list of (direction, batch, query block, target tile) units, every CTA an equal contiguous share
//     (balanced to one tile); target tiles arrive by 1-D TMA bulk copies (cp.async.bulk + mbarrier ring).
//   * from 8192 points: nnd_scan_sym_kernel (nnd_sym.cuh): every distance is evaluated ONCE and serves both
//     directions (row minima in registers, column minima through a filtered, warp-aggregated slow path).
//   * row ranges from 200 000 rows (or any range against a cloud that large): nnd_scan_symc_kernel (nnd_symc.cuh),
//     the same algorithm with the column points in UNIFORM registers, fed from the constant bank, one launch per
//     panel of columns on rotating streams.
//   Common to all: each thread keeps QPT rows in registers, negated, so one packed instruction evaluates two pairs:
//         FADD2 x3 (t - q), FMUL2 x3 (squares), FFMA2 x2 with a run-time 1.0
//         (fma(a,1,b) == fl(a+b); ptxas would contract mul.rn.f32x2 + add.rn.f32x2 into
//          one FFMA2 and change the rounding, the run-time 1.0 makes that impossible)
//     = 8 packed instructions per 2 pairs, each product and sum separately rounded; minimum with one 3-input FMNMX
//     per 2 pairs; the index is recovered lazily: per chunk of C targets the thread keeps only the chunk minimum and,
//     when it beats the running best (strict <, ascending order), remembers WHICH chunk.
//   * partial results of different CTAs / launches merge through atomicMin on a packed (dist_bits << 32 | chunk_base)
//     key: d >= +0 so uint order == float order, and the earlier chunk wins exact ties.  nnd_finalize_kernel
//     re-evaluates that single chunk per query (same instruction sequence => same bits) and takes the lowest k with
//     d == dist.
//
// Backward: nnd_grad_kernel, one thread per (term, batch, row); warp-aggregated
// (match.any on the neighbour index) fp32 atomics into zero-initialised grads.
// Deterministic backward (opt-in): nnd_grad_exact_kernel (nnd_grad_exact.cuh), bit-identical to my_lib.cpp:91-130.

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>

#include <cub/device/device_radix_sort.cuh>  // the deterministic backward sorts (neighbour row, own row) keys

#include "nnd_b200.h"

#define NND_VERSION_STRING "nnd_b200 0.1.0 sm_100a"

namespace {

// ----------------------------------------------------------------------------------
// packed FP32 helpers (64-bit register pair = {lo, hi} floats)
// ----------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t pk2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ float fmin3(float a, float b, float c) {
  float r;
  asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

// Two squared distances at once.  T* = {t[k], t[k+1]}, nq* = {-q, -q}, one = {1.0f, 1.0f}
// passed at run time.  Every operation rounds separately (my_lib.cpp:12-15).
__device__ __forceinline__ uint64_t sqdist2(uint64_t TX, uint64_t TY, uint64_t TZ, uint64_t nqx,
                                            uint64_t nqy, uint64_t nqz, uint64_t one) {
  const uint64_t dx = add2(TX, nqx), dy = add2(TY, nqy), dz = add2(TZ, nqz);
  const uint64_t xx = mul2(dx, dx), yy = mul2(dy, dy), zz = mul2(dz, dz);
  return fma2(fma2(xx, one, yy), one, zz);
}

// ----------------------------------------------------------------------------------
// mbarrier + 1-D TMA bulk copy
// ----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// ----------------------------------------------------------------------------------
// parameters
// ----------------------------------------------------------------------------------
struct ScanDir {
  const float* qsoa;         // [b][qtiles][3][TT]   queries of this direction
  const float* tsoa;         // [b][ttiles][3][TT]   targets of this direction
  unsigned long long* keys;  // [b][nq]  packed (dist_bits << 32 | chunk_base), pre-set to ~0
  int nq, nt;
  int qtiles, ttiles;
  int qblocks;      // ceil(nq / QB)
  long long units;  // b * qblocks * ttiles   (one unit = one query block x one target tile)
};
struct ScanParams {
  ScanDir d[2];
  int ndir;
  long long total_units;
  float one;  // 1.0f, deliberately opaque to the compiler (see header comment)
};

struct PrepCloud {
  const float* src;  // [b][n][3]
  float* soa;        // [b][tiles][3][TT]
  int n, tiles;
  long long padded_total;  // b * tiles * TT
  long long src_stride;    // points between batch elements of src (n, or more when src is a row range of a cloud)
};
struct PrepParams {
  PrepCloud c[2];
  int nclouds;
  unsigned long long* fill64[3];  // key arrays set to kKeyInit
  long long count64[3];
  unsigned* fill32;  // column filter set to 0xffffffff
  long long count32;
};

// Initial value of every packed key.  Larger than any (dist_bits << 32 | x) with dist_bits <= 0x7fffffff, and
// positive as a SIGNED 64-bit integer, so per-rank column keys can be merged with an int64 MIN all-reduce.
constexpr unsigned long long kKeyInit = 0x7fffffffffffffffull;

struct FinDir {
  const unsigned long long* keys;  // [b][nq]
  const float* q;                  // [b][nq][3] original query cloud (AoS)
  const float* tsoa;               // [b][ttiles][3][TT]
  float* dist;                     // [b][nq]
  int32_t* idx;                    // [b][nq]
  int nq, ttiles;
  int row_lo, nrows;  // rows [row_lo, row_lo + nrows) of every batch element are finalized
  long long count;    // b * nrows
  // Partial results instead of dist/idx (role-swapped row ranges): keyout[batch * key_stride + row] =
  // (dist_bits << 32 | (idx_offset + idx) / key_group), the column-key format of nnd_sym.cuh.  No NaN-first rule here:
  // it is applied once, after the partial keys of all ranges have been merged (nnd_finalize_cols_kernel).
  unsigned long long* keyout;
  long long key_stride;
  int idx_offset, key_group;
};
struct FinParams {
  FinDir d[2];
  int ndir;
  float one;
};

// ----------------------------------------------------------------------------------
// prepare: AoS [n,3] -> SoA tiles, keys = ~0
// ----------------------------------------------------------------------------------
template <int TT>
__global__ void __launch_bounds__(256) nnd_prepare_kernel(PrepParams p) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (int ci = 0; ci < p.nclouds; ++ci) {
    const PrepCloud c = p.c[ci];
    const long long per_batch = (long long)c.tiles * TT;
    for (long long g = first; g < c.padded_total; g += stride) {
      const long long batch = g / per_batch;
      const int j = (int)(g - batch * per_batch);
      const int tile = j / TT, off = j - tile * TT;
      const int js = j < c.n ? j : c.n - 1;  // pad by repeating the last point
      const float* s = c.src + (batch * c.src_stride + js) * 3;
      float* d = c.soa + ((batch * c.tiles + tile) * 3) * (long long)TT + off;
      d[0] = s[0];
      d[TT] = s[1];
      d[2 * TT] = s[2];
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k)
    for (long long g = first; g < p.count64[k]; g += stride) p.fill64[k][g] = kKeyInit;
  for (long long g = first; g < p.count32; g += stride) p.fill32[g] = 0xffffffffu;
}

// ----------------------------------------------------------------------------------
// finalize: key = (dist_bits << 32 | first index of the C-target chunk that attains the minimum).
// The chunk is re-evaluated with the same instruction sequence as the scan (bit-identical distances) and the lowest
// k with d == dist is taken (my_lib.cpp:16: first wins).
// Warp-cooperative: a warp owns 32 consecutive rows (keys and query points loaded coalesced, handed round by
// shuffles); C / 4 lanes share one row, each re-evaluating 4 consecutive targets of its chunk, so the chunk is read
// with fully coalesced 16-byte loads -- one thread per row walking its own chunk touched 32 different lines per
// load and made the kernel L2-bound at 3x the bytes (0.35 ms for the 10^6 x 128-target chunks behind a panel scan).
// ----------------------------------------------------------------------------------
template <int TT, int C>
__global__ void __launch_bounds__(256) nnd_finalize_kernel(FinParams p) {
  static_assert(C % 4 == 0 && (C / 4 == 32 || C / 4 == 16 || C / 4 == 8 || C / 4 == 4), "chunk size");
  constexpr int G = C / 4;   // lanes per row
  constexpr int R = 32 / G;  // rows in flight per warp
  constexpr unsigned FULL = 0xffffffffu;
  const uint64_t one = pk2(p.one, p.one);
  const int lane = threadIdx.x & 31;
  const int sub = lane % G, grp = lane / G;
  const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (int di = 0; di < p.ndir; ++di) {
    const FinDir d = p.d[di];
    const long long nblocks = (d.count + 31) / 32;
    for (long long wb = (((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5); wb < nblocks; wb += warps) {
      // ---- this lane's own row
      const long long w = wb * 32 + lane;
      const bool live = w < d.count;
      const long long wc = live ? w : d.count - 1;
      const long long batch = wc / d.nrows;
      const long long g = batch * d.nq + d.row_lo + (wc - batch * d.nrows);  // flat [b][nq] index of this row
      const unsigned long long key = d.keys[g];
      const float* qp = d.q + g * 3;
      const float x = -qp[0], y = -qp[1], z = -qp[2];
      const int mybatch = (int)batch;
      int myfound = 0;
#pragma unroll 4  // four rows' loads in flight: the loop is latency-bound (ncu: long scoreboard 4.2 per issue at unroll 1)
      for (int it = 0; it < 32 / R; ++it) {
        const int r = it * R + grp;  // the row (lane index inside the warp) this lane group works on
        const unsigned long long rkey = __shfl_sync(FULL, key, r);
        const float rx = __shfl_sync(FULL, x, r), ry = __shfl_sync(FULL, y, r), rz = __shfl_sync(FULL, z, r);
        const int rbatch = __shfl_sync(FULL, mybatch, r);
        const float bq = __uint_as_float((unsigned)(rkey >> 32));
        const int base = (int)(unsigned)(rkey & 0xffffffffull);
        const uint64_t nqx = pk2(rx, rx), nqy = pk2(ry, ry), nqz = pk2(rz, rz);
        const int tile = base / TT, off = base - tile * TT;
        const float* sx = d.tsoa + (((long long)rbatch * d.ttiles + tile) * 3) * (long long)TT + off + 4 * sub;
        const ulonglong2 X = __ldg(reinterpret_cast<const ulonglong2*>(sx));
        const ulonglong2 Y = __ldg(reinterpret_cast<const ulonglong2*>(sx + TT));
        const ulonglong2 Z = __ldg(reinterpret_cast<const ulonglong2*>(sx + 2 * TT));
        float a0, a1, b0, b1;
        upk2(sqdist2(X.x, Y.x, Z.x, nqx, nqy, nqz, one), a0, a1);
        upk2(sqdist2(X.y, Y.y, Z.y, nqx, nqy, nqz, one), b0, b1);
        int found = 0x7fffffff;  // lowest matching target of this lane's four
        if (b1 == bq) found = 4 * sub + 3;
        if (b0 == bq) found = 4 * sub + 2;
        if (a1 == bq) found = 4 * sub + 1;
        if (a0 == bq) found = 4 * sub;
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) found = min(found, __shfl_xor_sync(FULL, found, o));
        // no match at all (a NaN minimum never compares equal): the chunk's first target, as before
        found = found == 0x7fffffff ? 0 : found;
        const int mine = __shfl_sync(FULL, found, (lane % R) * G);  // the group that worked on MY row in iteration lane / R
        if (lane / R == it) myfound = mine;
      }
      if (live) {
        const float bq = __uint_as_float((unsigned)(key >> 32));
        const int base = (int)(unsigned)(key & 0xffffffffull);
        const uint64_t nqx = pk2(x, x), nqy = pk2(y, y), nqz = pk2(z, z);
        // my_lib.cpp:16 accepts k == 0 unconditionally and later candidates only if d < best: a NaN first distance
        // (NaN query, or NaN first target) is therefore the answer.  The scan's minimum ignores NaNs, which gives the
        // reference's result in every other non-finite case.
        const float* t0 = d.tsoa + (batch * d.ttiles * 3) * (long long)TT;
        float first, unused;
        upk2(sqdist2(pk2(t0[0], t0[0]), pk2(t0[TT], t0[TT]), pk2(t0[2 * TT], t0[2 * TT]), nqx, nqy, nqz, one), first, unused);
        const bool nan0 = first != first;
        if (d.keyout != nullptr) {
          d.keyout[batch * d.key_stride + d.row_lo + (w - batch * d.nrows)] =
              (key & 0xffffffff00000000ull) | (unsigned long long)(unsigned)((d.idx_offset + base + myfound) / d.key_group);
        } else {
          d.dist[g] = nan0 ? first : bq;
          d.idx[g] = nan0 ? 0 : base + myfound;
        }
      }
    }
  }
}

// ----------------------------------------------------------------------------------
// scan: the hot kernel
// ----------------------------------------------------------------------------------
// Work is the linear sequence of units (direction, batch, query block, target tile), tile fastest.
// CTA c owns the contiguous range [U*c/G, U*(c+1)/G): perfectly balanced to one tile, and the tile
// stream of a CTA is contiguous in memory within a query block.  A CTA's range may start or end in the
// middle of a query block's tile list; partial results merge through atomicMin on the packed keys.
struct UnitCoords {
  int dir, batch, qblock, tile;
};
__device__ __forceinline__ UnitCoords decode_unit(const ScanParams& p, long long u) {
  UnitCoords c;
  c.dir = (p.ndir > 1 && u >= p.d[0].units) ? 1 : 0;
  const ScanDir& d = p.d[c.dir];
  if (c.dir) u -= p.d[0].units;
  const long long qb = u / d.ttiles;
  c.tile = (int)(u - qb * d.ttiles);
  c.batch = (int)(qb / d.qblocks);
  c.qblock = (int)(qb - (long long)c.batch * d.qblocks);
  return c;
}

template <int THREADS, int QPT, int TT, int C, int STAGES, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) nnd_scan_kernel(const __grid_constant__ ScanParams p) {
  static_assert(TT % C == 0 && C % 4 == 0, "tile/chunk shape");
  constexpr int QB = THREADS * QPT;
  constexpr uint32_t TILE_BYTES = 3 * TT * sizeof(float);
  __shared__ __align__(128) float tiles[STAGES][3 * TT];
  __shared__ __align__(8) uint64_t full[STAGES];

  const int tid = threadIdx.x;
  const long long u_begin = p.total_units * blockIdx.x / gridDim.x;
  const long long u_end = p.total_units * (blockIdx.x + 1) / gridDim.x;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // producer (thread 0): unit `pu` is the next tile to fetch
  long long pu = u_begin;
  auto produce = [&](int stage) {
    const UnitCoords c = decode_unit(p, pu);
    const ScanDir& d = p.d[c.dir];
    const float* src = d.tsoa + ((long long)c.batch * d.ttiles + c.tile) * (3 * TT);
    mbar_expect_tx(&full[stage], TILE_BYTES);
    tma_load_1d(&tiles[stage][0], src, TILE_BYTES, &full[stage]);
    ++pu;
  };
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s)
      if (pu < u_end) produce(s);
  }

  const uint64_t one = pk2(p.one, p.one);
  uint32_t cnt = 0;  // tiles consumed by this CTA so far

  long long u = u_begin;
  while (u < u_end) {
    const UnitCoords uc = decode_unit(p, u);
    const ScanDir& d = p.d[uc.dir];
    const long long rest = (long long)d.ttiles - uc.tile;  // tiles left in this query block's list
    const int seg_tiles = (int)(u_end - u < rest ? u_end - u : rest);

    // ---- queries -> registers (negated, duplicated in both halves)
    uint64_t nqx[QPT], nqy[QPT], nqz[QPT];
    float best[QPT];
    int bchunk[QPT];  // first index of the first chunk that attains best
    const float* qbase = d.qsoa + (long long)uc.batch * d.qtiles * (3 * TT);
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
      int j = uc.qblock * QB + q * THREADS + tid;
      j = j < d.nq ? j : d.nq - 1;
      const int qt = j / TT, qo = j - qt * TT;
      const float* s = qbase + (long long)qt * (3 * TT) + qo;
      const float x = -s[0], y = -s[TT], z = -s[2 * TT];
      nqx[q] = pk2(x, x);
      nqy[q] = pk2(y, y);
      nqz[q] = pk2(z, z);
      best[q] = __int_as_float(0x7f800000);  // +inf: an all-inf row keeps idx 0 like the reference
      bchunk[q] = 0;
    }

    for (int tile = uc.tile; tile < uc.tile + seg_tiles; ++tile) {
      const int stage = cnt % STAGES;
      mbar_wait(&full[stage], (cnt / STAGES) & 1);
      const float* sx = &tiles[stage][0];
      const float* sy = sx + TT;
      const float* sz = sx + 2 * TT;
      const int valid = min(TT, d.nt - tile * TT);
      const int nchunks = (valid + C - 1) / C;

      for (int ch = 0; ch < nchunks; ++ch) {
        const ulonglong2* cx = reinterpret_cast<const ulonglong2*>(sx + ch * C);
        const ulonglong2* cy = reinterpret_cast<const ulonglong2*>(sy + ch * C);
        const ulonglong2* cz = reinterpret_cast<const ulonglong2*>(sz + ch * C);
        float cm[QPT];  // minimum over this chunk only
#pragma unroll
        for (int q = 0; q < QPT; ++q) cm[q] = __int_as_float(0x7f800000);
#pragma unroll
        for (int s4 = 0; s4 < C / 4; ++s4) {
          const ulonglong2 X = cx[s4], Y = cy[s4], Z = cz[s4];  // 3 x LDS.128 broadcast = 4 targets
#pragma unroll
          for (int q = 0; q < QPT; ++q) {
            const uint64_t d0 = sqdist2(X.x, Y.x, Z.x, nqx[q], nqy[q], nqz[q], one);
            const uint64_t d1 = sqdist2(X.y, Y.y, Z.y, nqx[q], nqy[q], nqz[q], one);
            float a0, a1, b0, b1;
            upk2(d0, a0, a1);
            upk2(d1, b0, b1);
            cm[q] = fmin3(cm[q], a0, a1);
            cm[q] = fmin3(cm[q], b0, b1);
          }
        }
        // strict < in ascending chunk order: the first chunk that attains the minimum is remembered
        const int cbase = tile * TT + ch * C;
#pragma unroll
        for (int q = 0; q < QPT; ++q) {
          if (cm[q] < best[q]) bchunk[q] = cbase;
          best[q] = fminf(best[q], cm[q]);
        }
      }

      __syncthreads();  // every thread is done reading this stage
      if (tid == 0 && pu < u_end) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        produce(stage);
      }
      ++cnt;
    }

    // ---- merge this segment's result: smaller dist wins, then the earlier chunk
    unsigned long long* keys = d.keys + (long long)uc.batch * d.nq;
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
      const int j = uc.qblock * QB + q * THREADS + tid;
      if (j < d.nq) {
        const unsigned long long key =
            ((unsigned long long)__float_as_uint(best[q]) << 32) | (unsigned long long)(unsigned)bchunk[q];
        atomicMin(keys + j, key);
      }
    }
    u += seg_tiles;
  }
}

#include "nnd_sym.cuh"
#include "nnd_symc.cuh"
#include "nnd_hint.cuh"
#include "nnd_pruned.cuh"

// ----------------------------------------------------------------------------------
// backward
// ----------------------------------------------------------------------------------
#include "nnd_grad.cuh"
#include "nnd_grad_exact.cuh"
#include "nnd_small.cuh"

// ----------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------
#ifndef NND_THREADS
#define NND_THREADS 256
#endif
#ifndef NND_QPT
#define NND_QPT 8
#endif
#ifndef NND_MINB
#define NND_MINB 2
#endif
#ifndef NND_STAGES
#define NND_STAGES 2
#endif
constexpr int kThreads = NND_THREADS;
constexpr int kQPT = NND_QPT;
#ifndef NND_TT
#define NND_TT 512
#endif
constexpr int kTT = NND_TT;
#ifndef NND_CHUNK
#define NND_CHUNK 32
#endif
constexpr int kChunk = NND_CHUNK;
#ifndef NND_SYMC_CHUNK
#define NND_SYMC_CHUNK NND_CHUNK
#endif
constexpr int kSymcChunk = NND_SYMC_CHUNK;  // lazy-index chunk of the panel scan
#ifndef NND_SYMC_SKIP
#define NND_SYMC_SKIP 4
#endif
constexpr int kSymcFinChunk = NND_SYMC_SKIP * NND_SYMC_CHUNK;  // ... and of the finalize that follows it
constexpr int kStages = NND_STAGES;
constexpr int kMinBlocks = NND_MINB;
constexpr int kQB = kThreads * kQPT;
constexpr int kRowAlign = 8;  // row ranges of the multi-GPU path start on multiples of this (a multiple of every QPT)

// Work shapes.  The large shape is the throughput kernel; the two small ones exist because a (query block x target
// tile) unit of 2048 x 512 leaves most of the 148 SMs idle on the clouds the reference's callers really use
// (DIP patches B=256,N=M=256; NDP clouds ~3 k points): fewer rows per thread -> more, smaller units.
enum ScanShape { kShapeLarge = 0, kShapeMid = 1, kShapeSmall = 2 };
struct ShapeInfo {
  int threads, qpt, minb;
};
constexpr ShapeInfo kShapes[3] = {{kThreads, kQPT, kMinBlocks}, {256, 2, 4}, {128, 2, 8}};

// Chosen from the GLOBAL problem size only (never from a rank's row range): every rank of the multi-GPU path must
// use the same row-group size, because column keys carry row-group ids.
inline ScanShape choose_shape(int b, int n, int m, int sms) {
  const long long units_large = (long long)b * ((n + kQB - 1) / kQB) * ((m + kTT - 1) / kTT);
  if (units_large >= 4ll * sms) return kShapeLarge;
  if (n <= 256) return kShapeSmall;
  return kShapeMid;
}

thread_local int g_launches = 0;

// optional per-kernel event timing of the forward (bench.py's roofline block)
constexpr int kProfSlots = 64;
struct Profiler {
  bool enabled = false, created = false;
  int used = 0;
  cudaEvent_t ev[kProfSlots * 4];
};
thread_local Profiler g_prof;

struct DeviceInfo {
  int sms = 0;
  int major = 0;
};
int device_info(DeviceInfo* out) {
  // per-device cache, filled once under a mutex (calls may come from any host thread)
  static DeviceInfo cache[64];
  static std::atomic<bool> ready[64];
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return NND_EINVAL;
  if (!ready[dev].load(std::memory_order_acquire)) {
    std::lock_guard<std::mutex> lock(mu);
    if (!ready[dev].load(std::memory_order_relaxed)) {
      int sms = 0, major = 0;
      e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (e != cudaSuccess) return (int)e;
      e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
      if (e != cudaSuccess) return (int)e;
      cache[dev].major = major;
      cache[dev].sms = sms;
      ready[dev].store(true, std::memory_order_release);
    }
  }
  *out = cache[dev];
  return out->major == 10 ? 0 : NND_EDEVICE;
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel instantiation, device); thread-safe.
struct SmemOptIn {
  std::atomic<int> have[64];
  SmemOptIn() {
    for (auto& h : have) h.store(0);
  }
  template <typename K>
  int ensure(K kern, size_t smem) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    dev &= 63;
    if ((size_t)have[dev].load(std::memory_order_acquire) >= smem) return 0;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    int cur = have[dev].load();
    while ((size_t)cur < smem && !have[dev].compare_exchange_weak(cur, (int)smem)) {
    }
    return 0;
  }
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
inline int tiles_of(int n) { return (n + kTT - 1) / kTT; }
inline size_t soa_bytes(int b, int n) { return align_up((size_t)b * tiles_of(n) * 3 * kTT * sizeof(float), 256); }
inline size_t key_bytes(int b, int n) { return align_up((size_t)b * n * sizeof(unsigned long long), 256); }
inline size_t colkey_count(int b, int m) { return (size_t)b * tiles_of(m) * kTT; }
inline size_t colkey_bytes(int b, int m) { return align_up(colkey_count(b, m) * sizeof(unsigned long long), 256); }
inline size_t filter_bytes(int b, int m) { return align_up(colkey_count(b, m) * sizeof(unsigned), 256); }

// Workspace of the row-range forward: SoA copies of both clouds plus either orientation's keys and filter
// (normal: row keys of xyz1 + filter of xyz2; role-swapped: row keys of xyz2 + column keys and filter of xyz1's range).
// Scratch of the automatic hints (nnd_hint.cuh), placed behind everything else in the row-range workspace: one hint per
// column, a hash table of 2^k >= 2 * rows cells, the bounding boxes.  Sized for either orientation.
inline int hint_table_size(int rows) {
  int t = 1024;  // kHintLevels entries per row, load factor <= 3/4
  while (t < (long long)(kHintLevels + 1) * rows && t < (1 << 30)) t <<= 1;
  return t;
}
inline size_t hint_bytes(int b, int n, int m) {
  const int big = n > m ? n : m;
  return align_up((size_t)b * big * sizeof(int32_t), 256) + align_up((size_t)b * hint_table_size(big) * sizeof(int), 256) +
         256;
}
inline size_t rows_core_need(int b, int n, int m) {
  const size_t normal = key_bytes(b, n) + filter_bytes(b, m);
  const size_t swapped = key_bytes(b, m) + colkey_bytes(b, n) + filter_bytes(b, n);
  return soa_bytes(b, n) + soa_bytes(b, m) + (normal > swapped ? normal : swapped);
}
inline size_t rows_need(int b, int n, int m) { return rows_core_need(b, n, m) + hint_bytes(b, n, m); }

// 0 = off, 1 = in front of the panel scans (default), 2 = in front of the persistent symmetric kernel too.
std::atomic<int> g_auto_hint{1};

// Fills `hint` ([b][m] int32, inside the workspace) with a nearby row of `rows` for every point of `cols`
// (nnd_hint.cuh); returns the hint pointer, or nullptr when the clouds are too small to bother.
const int32_t* make_hints(const float* rows, long long row_stride, int n, const float* cols, long long col_stride, int m,
                          int b, char* scratch, int sms, cudaStream_t st) {
  HintParams hp{};
  hp.rows = rows;
  hp.cols = cols;
  hp.row_stride = row_stride;
  hp.col_stride = col_stride;
  hp.n = n;
  hp.m = m;
  hp.b = b;
  const int big = n > m ? n : m;
  hp.hint = (int32_t*)scratch;
  scratch += align_up((size_t)b * big * sizeof(int32_t), 256);
  hp.tsize = hint_table_size(n);
  hp.table = (int*)scratch;
  scratch += align_up((size_t)b * hint_table_size(big) * sizeof(int), 256);
  hp.bbox = (unsigned*)scratch;
  int g = (int)cbrt((double)n / 2.0);
  hp.g = g < 1 ? 1 : g > 1024 ? 1024 : g;
  // coarse levels below level 0 for the descent of far columns: down to at most 6 cells per axis (all of them probed);
  // NND_B200_HINT_COARSE=0 switches the descent off (A/B measurements)
  static const int coarse_on = [] {
    const char* e = getenv("NND_B200_HINT_COARSE");
    return e ? atoi(e) : 1;
  }();
  hp.coarse = 0;
  if (coarse_on)
    while ((hp.g >> hp.coarse) > 6) ++hp.coarse;
  auto blocks = [&](long long items) {
    long long bl = (items + 255) / 256;
    return (unsigned)(bl > 8ll * sms ? 8ll * sms : bl < 1 ? 1 : bl);
  };
  nnd_hint_clear_kernel<<<blocks((long long)b * hp.tsize), 256, 0, st>>>(hp);
  nnd_hint_bbox_kernel<<<dim3(blocks(n), (unsigned)b), 256, 0, st>>>(hp);
  nnd_hint_build_kernel<<<dim3(blocks(n), (unsigned)b), 256, 0, st>>>(hp);
  nnd_hint_query_kernel<<<dim3(blocks(m), (unsigned)b), 256, 0, st>>>(hp);
  g_launches += 4;
  return hp.hint;
}

int check_last(cudaError_t e) { return e == cudaSuccess ? 0 : (int)e; }

constexpr int kReduceMaxBlocks = 1024;  // partial sums of nnd_chamfer_reduce_kernel

// NND_B200_DEBUG_SYNC=1: synchronise after every kernel of the forward and report it on stderr (triage aid).
void debug_sync(const char* what, cudaStream_t st) {
  static const int on = [] {
    const char* e = getenv("NND_B200_DEBUG_SYNC");
    return (e && e[0] == '1') ? 1 : 0;
  }();
  if (!on) return;
  fprintf(stderr, "[nnd] %s launched...\n", what);
  fflush(stderr);
  cudaError_t e = cudaStreamSynchronize(st);
  fprintf(stderr, "[nnd] %s done: %s\n", what, cudaGetErrorString(e));
  fflush(stderr);
}

inline unsigned grid_for(long long items, int sms) {
  long long blocks = (items + 255) / 256;
  if (blocks > (long long)sms * 16) blocks = (long long)sms * 16;
  return (unsigned)(blocks < 1 ? 1 : blocks);
}

template <int THREADS, int QPT, int MINB>
int launch_sym(const SymParams& yp, int grid, size_t smem, cudaStream_t st) {
  auto kern = nnd_scan_sym_kernel<THREADS, QPT, kTT, kChunk, kStages, MINB>;
  static SmemOptIn optin;
  int rc = optin.ensure(kern, smem);
  if (rc) return rc;
  kern<<<grid, THREADS, smem, st>>>(yp);
  return 0;
}

// ---- panel path (nnd_symc.cuh): one launch per panel of 1, 2 or 4 target tiles, rotating over 8, 4 or 2 internal
// streams / constant buffers.  The caller's stream forks into them and joins them at the end, so the whole thing is
// ordered like a single operation on `st` (also under stream capture).
struct PanelCtx {
  cudaStream_t s[kPanelMaxBufs] = {};
  cudaEvent_t fork = nullptr, join[kPanelMaxBufs] = {};
  cudaEvent_t done = nullptr;  // end of the last panel scan that was enqueued outside stream capture
  bool have_done = false;
  bool ready = false;
};
std::mutex g_panel_mu;  // the constant window is per device: panel scans of one process are enqueued one at a time

int panel_ctx(PanelCtx** out) {
  static PanelCtx ctx[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  PanelCtx& c = ctx[dev & 63];
  if (!c.ready) {
    for (int i = 0; i < kPanelMaxBufs; ++i) {
      if ((e = cudaStreamCreateWithFlags(&c.s[i], cudaStreamNonBlocking)) != cudaSuccess) return (int)e;
      if ((e = cudaEventCreateWithFlags(&c.join[i], cudaEventDisableTiming)) != cudaSuccess) return (int)e;
    }
    if ((e = cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming)) != cudaSuccess) return (int)e;
    if ((e = cudaEventCreateWithFlags(&c.done, cudaEventDisableTiming)) != cudaSuccess) return (int)e;
    c.ready = true;
  }
  *out = &c;
  return 0;
}

// Rows per launch below which the panel path does not pay (each launch must be long against its launch, copy and
// ramp): NND_B200_PANEL_MIN_ROWS overrides (tests use 1 to push every symmetric scan through the panels).
// Tuning knobs are atomics: nnd_set_panel_* may be called while other host threads enqueue forwards.
std::atomic<long long> g_panel_min_rows{-1};
std::atomic<int> g_panel_tiles{-1};  // 0 = choose from the row count; 1, 2, 4 = forced (tests, tuning)
long long panel_min_rows() {
  long long v = g_panel_min_rows.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("NND_B200_PANEL_MIN_ROWS");
    long long init = e ? atoll(e) : 200000;
    if (init < 1) init = 1;
    long long expect = -1;
    g_panel_min_rows.compare_exchange_strong(expect, init);  // somebody else's value (or a set) wins
    v = g_panel_min_rows.load(std::memory_order_relaxed);
  }
  return v;
}
int panel_tiles() {
  int v = g_panel_tiles.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("NND_B200_PANEL_TILES");
    int init = e ? atoi(e) : 0;
    if (!(init == 1 || init == 2 || init == 4)) init = 0;
    int expect = -1;
    g_panel_tiles.compare_exchange_strong(expect, init);
    v = g_panel_tiles.load(std::memory_order_relaxed);
  }
  return v;
}

// Shortest row range that takes the role-swapped panel path (each panel launch streams <= 2048 of its rows).
long long panel_swap_min_rows() { return panel_min_rows() < 8192 ? panel_min_rows() : 8192; }

template <int PT>
int scan_panels_pt(SymCParams yp, const float* soa2, int b, int qblocks, PanelCtx* c) {
  constexpr int THREADS = kThreads, QPT = kQPT;
  constexpr int NBUF = 8 / PT;
  const size_t smem = (size_t)PT * kTT * sizeof(unsigned) + (size_t)THREADS * QPT * 8;
  cudaError_t e;
  int turn = 0;
  for (int batch = 0; batch < b; ++batch) {
    for (int tile_lo = 0; tile_lo < yp.ttiles; tile_lo += PT, ++turn) {
      const int which = turn % NBUF;
      yp.batch = batch;
      yp.tile_lo = tile_lo;
      yp.ntiles = yp.ttiles - tile_lo < PT ? yp.ttiles - tile_lo : PT;
      yp.cofs = which * PT * 3 * kTT;
      const float* src = soa2 + ((long long)batch * yp.ttiles + tile_lo) * (3 * kTT);
      e = cudaMemcpyToSymbolAsync(c_panel, src, sizeof(float) * yp.ntiles * 3 * kTT, sizeof(float) * yp.cofs,
                                  cudaMemcpyDeviceToDevice, c->s[which]);
      if (e != cudaSuccess) return (int)e;
      const bool full = yp.ntiles == PT && (tile_lo + PT) * kTT <= yp.m;
      if (full)
        nnd_scan_symc_kernel<THREADS, QPT, kTT, kSymcChunk, kMinBlocks, PT, true><<<qblocks, THREADS, smem, c->s[which]>>>(yp);
      else
        nnd_scan_symc_kernel<THREADS, QPT, kTT, kSymcChunk, kMinBlocks, PT, false><<<qblocks, THREADS, smem, c->s[which]>>>(yp);
      ++g_launches;
    }
  }
  return 0;
}

int scan_panels(const SymCParams& yp, const float* soa2, int b, int qblocks, int sms, cudaStream_t st) {
  std::lock_guard<std::mutex> lock(g_panel_mu);
  PanelCtx* c = nullptr;
  int rc = panel_ctx(&c);
  if (rc) return rc;
  cudaError_t e;
  if ((e = cudaEventRecord(c->fork, st)) != cudaSuccess) return (int)e;
  // From here on the internal streams depend on `st`: whatever fails below, they are joined back into `st` before
  // returning, so the caller's stream stays ordered after every kernel that reads the workspace and the constant
  // window.  The first error is kept and returned after the join.
  int first_err = 0;
  auto keep = [&](cudaError_t x) {
    if (x != cudaSuccess && first_err == 0) first_err = (int)x;
  };
  for (int i = 0; i < kPanelMaxBufs; ++i) keep(cudaStreamWaitEvent(c->s[i], c->fork, 0));
  // Scans enqueued from DIFFERENT caller streams share the constant window.  Buffer i is only ever written and read on
  // internal stream i, which orders scans with the same panel size; scans with different panel sizes cut the window
  // differently, so every internal stream also waits for the whole previous scan.  (Panel scans are never captured
  // into CUDA graphs: forward_rows_impl falls back to the persistent kernel while `st` is capturing, so every panel
  // scan of the process passes through here and is ordered by these events.)
  if (c->have_done)
    for (int i = 0; i < kPanelMaxBufs; ++i) keep(cudaStreamWaitEvent(c->s[i], c->done, 0));
  if (first_err == 0) {
    // launches in flight x CTAs per launch should cover the 2 * sms resident CTAs about twice
    int pt = qblocks * 2 >= 3 * sms ? 4 : qblocks * 4 >= 3 * sms ? 2 : 1;
    const int forced = panel_tiles();
    if (forced == 1 || forced == 2 || forced == 4) pt = forced;
    rc = pt == 4 ? scan_panels_pt<4>(yp, soa2, b, qblocks, c)
       : pt == 2 ? scan_panels_pt<2>(yp, soa2, b, qblocks, c)
                 : scan_panels_pt<1>(yp, soa2, b, qblocks, c);
    if (rc && first_err == 0) first_err = rc;
  }
  for (int i = 0; i < kPanelMaxBufs; ++i) {
    keep(cudaEventRecord(c->join[i], c->s[i]));
    keep(cudaStreamWaitEvent(st, c->join[i], 0));
  }
  e = cudaEventRecord(c->done, st);
  keep(e);
  c->have_done = e == cudaSuccess;  // a failed record leaves no stale event for the next scan to wait on
  if (first_err) return first_err;
  return check_last(cudaGetLastError());
}

template <int THREADS, int QPT, int MINB>
void launch_scan(const ScanParams& sp, int grid, cudaStream_t st) {
  nnd_scan_kernel<THREADS, QPT, kTT, kChunk, kStages, MINB><<<grid, THREADS, 0, st>>>(sp);
}

inline bool use_symmetric(int n, int m);
// Sample tiles (512 rows each) of the column-filter seed run in front of the persistent symmetric kernel.  One tile
// costs 512 / rows of the scan and removes the cold-filter start (every first-wave warp step taking the column slow
// path): measured forward, profiles/r02_sym_seed_ab.txt / r02_panel_threshold_ab.txt: N=M=150 000 8.78 -> 6.51 ms,
// 32 768 0.482 -> 0.429 ms, 65 536 1.404 -> 1.285 ms, C2 2.75 -> 2.69 ms, C3 2.91 -> 2.88 ms; more tiles lose again.
std::atomic<int> g_sym_seed{1};
// Shape of the symmetric kernel: the large one whenever the single-GPU entry point would pick the symmetric path
// (more rows per warp step = fewer column slow paths); small clouds only get here through the row-range entry point.
inline ScanShape sym_shape(int b, int n, int m, int sms) {
  return use_symmetric(n, m) ? kShapeLarge : choose_shape(b, n, m, sms);
}

int scan_panels(const SymCParams& yp, const float* soa2, int b, int qblocks, int sms, cudaStream_t st);

// The one place that decides which scan a row range takes; forward_rows_impl and nnd_rows_scan_kind both ask here.
// While the caller's stream is being captured into a CUDA graph the persistent kernel is used whatever the size: the
// panel scan keeps per-device state (the constant window and the events that order its users), which a graph replay
// would bypass.  Results are bit-identical on every path.
enum RowsPath { kRowsNone = 0, kRowsPersistent = 1, kRowsPanel = 2, kRowsSwapped = 3 };
inline RowsPath rows_path(int b, int n, int m, int row_lo, int row_hi, int sms, bool capturing) {
  const int nrows = row_hi - row_lo;
  if (b <= 0 || n <= 0 || m <= 0 || nrows <= 0) return kRowsNone;
  if (capturing) return kRowsPersistent;
  if (nrows < m && m >= panel_min_rows() && m >= 2 * kTT && nrows >= panel_swap_min_rows()) return kRowsSwapped;
  if (sym_shape(b, n, m, sms) == kShapeLarge && nrows >= panel_min_rows()) return kRowsPanel;
  return kRowsPersistent;
}

// Role-swapped row range: when the range [row_lo, row_hi) of xyz1 is shorter than xyz2 (a rank's shard of the
// multi-GPU path, or simply n < m), the panel scan runs with xyz2 in the registers and the RANGE as the streamed
// columns: every launch then has ceil(m / 2048) CTAs however short the range is, and each column sees m / 256 warp
// visits, so the slow path stays rare.  The kernel's row direction (xyz2 points -> nearest range row) is finalized
// to exact indices and written out as column keys in the usual format (dist_bits << 32 | xyz1 row / 8); its column
// direction (range rows -> nearest xyz2 point) is complete, because every range row has seen all of xyz2, and goes
// straight to dist1 / idx1.
int forward_rows_swapped_impl(const float* c1, int n1, const float* c2, int n2, int b, int row_lo, int row_hi,
                              float* dist1, int32_t* idx1, unsigned long long* colkeys, void* workspace,
                              const DeviceInfo& di, cudaEvent_t* ev, cudaStream_t st) {
  const int nrows = row_hi - row_lo;
  char* w = (char*)workspace;
  float* soaR = (float*)w;  // xyz2: the register side
  w += soa_bytes(b, n2);
  float* soaS = (float*)w;  // rows [row_lo, row_hi) of xyz1 as a cloud of its own: the streamed side
  w += soa_bytes(b, n1);
  unsigned long long* keysR = (unsigned long long*)w;
  w += key_bytes(b, n2);
  unsigned long long* ckeysS = (unsigned long long*)w;
  w += colkey_bytes(b, nrows);
  unsigned* cdistS = (unsigned*)w;
  const float* sub = c1 + (long long)row_lo * 3;

  PrepParams pp{};
  pp.nclouds = 2;
  pp.c[0] = PrepCloud{c2, soaR, n2, tiles_of(n2), (long long)b * tiles_of(n2) * kTT, n2};
  pp.c[1] = PrepCloud{sub, soaS, nrows, tiles_of(nrows), (long long)b * tiles_of(nrows) * kTT, n1};
  pp.fill64[0] = keysR;
  pp.count64[0] = (long long)b * n2;
  pp.fill64[1] = ckeysS;
  pp.count64[1] = (long long)colkey_count(b, nrows);
  pp.fill64[2] = colkeys;  // padding entries of the output stay at the initial value
  pp.count64[2] = (long long)colkey_count(b, n2);
  pp.fill32 = cdistS;
  pp.count32 = (long long)colkey_count(b, nrows);
  {
    long long total = pp.c[0].padded_total > pp.count64[2] ? pp.c[0].padded_total : pp.count64[2];
    if (ev) cudaEventRecord(ev[0], st);
    nnd_prepare_kernel<kTT><<<grid_for(total, di.sms), 256, 0, st>>>(pp);
    ++g_launches;
    if (ev) cudaEventRecord(ev[1], st);
    debug_sync("prepare(swapped)", st);
  }
  const int qblocks = (n2 + kThreads * kQPT - 1) / (kThreads * kQPT);
  {  // seed the filter of the range's points from whole sample tiles of xyz2
    int ns = n2 / kTT;
    const int want = n2 >= 500000 ? 4 : n2 >= 200000 ? 2 : 1;
    ns = ns > want ? want : ns;
    // automatic hints (nnd_hint.cuh): here the register side ("rows") is xyz2 and the columns are the range's points
    const int32_t* hints = nullptr;
    if (g_auto_hint.load(std::memory_order_relaxed) >= 1)
      hints = make_hints(c2, n2, n2, sub, n1, nrows, b, (char*)workspace + rows_core_need(b, n1, n2), di.sms, st);
    if (ns > 0 || hints != nullptr) {
      dim3 g((unsigned)((nrows + 511) / 512), (unsigned)b);
      nnd_seed_filter_kernel<kTT><<<g, 256, 0, st>>>(sub, soaR, cdistS, nrows, tiles_of(nrows), tiles_of(n2), 0, ns, 1.0f, n1,
                                                     hints, c2, n2, 0, n2);
      ++g_launches;
    }
  }
  SymCParams cp{};
  cp.qsoa = soaR;
  cp.rkeys = keysR;
  cp.ckeys = ckeysS;
  cp.cdist = cdistS;
  cp.n = n2;
  cp.m = nrows;
  cp.row_lo = 0;
  cp.row_hi = n2;
  cp.qtiles = tiles_of(n2);
  cp.ttiles = tiles_of(nrows);
  cp.one = 1.0f;
  int rc = scan_panels(cp, soaS, b, qblocks, di.sms, st);
  if (rc) return rc;
  if (ev) cudaEventRecord(ev[2], st);
  debug_sync("scan_panels(swapped)", st);

  // xyz2 points -> exact nearest row of the range -> column keys
  FinParams fp{};
  fp.ndir = 1;
  fp.one = 1.0f;
  fp.d[0] = FinDir{keysR, c2, soaS, nullptr, nullptr, n2, tiles_of(nrows), 0, n2, (long long)b * n2,
                   colkeys, (long long)tiles_of(n2) * kTT, row_lo,
                   kShapes[sym_shape(b, n1, n2, di.sms)].qpt};  // the row-group size finalize_cols_impl will assume
  nnd_finalize_kernel<kTT, kSymcFinChunk><<<grid_for(fp.d[0].count, di.sms), 256, 0, st>>>(fp);
  ++g_launches;
  // range rows -> nearest xyz2 point (complete): dist1 / idx1 of rows [row_lo, row_hi)
  FinColParams fc{ckeysS, c2, c1, dist1, idx1, n2, nrows, tiles_of(nrows), (long long)b * nrows, 1.0f, n1, row_lo};
  nnd_finalize_cols_kernel<kTT, kQPT><<<grid_for(fc.count, di.sms), 256, 0, st>>>(fc);
  ++g_launches;
  if (ev) cudaEventRecord(ev[3], st);
  debug_sync("finalize(swapped)", st);
  return check_last(cudaGetLastError());
}

// Symmetric forward over the row range [row_lo, row_hi) of xyz1 (every batch element):
//   prepare (SoA copies, key/filter init) -> nnd_scan_sym_kernel -> nnd_finalize_kernel for those rows.
// colkeys receives this launch's column keys; finalize_cols_impl turns (merged) column keys into dist2/idx2.
int forward_rows_impl(const float* c1, int n1, const float* c2, int n2, int b, int row_lo, int row_hi, float* dist1,
                      int32_t* idx1, unsigned long long* colkeys, void* workspace, size_t workspace_bytes,
                      cudaStream_t st, const int32_t* hint2 = nullptr) {
  cudaEvent_t* ev = nullptr;
  if (g_prof.enabled && g_prof.used < kProfSlots) ev = &g_prof.ev[4 * g_prof.used++];
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  if (((uintptr_t)workspace & 255u) != 0) return NND_EWORKSPACE;
  const size_t need = rows_need(b, n1, n2);
  if (workspace_bytes < need) return NND_EWORKSPACE;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaError_t ce = cudaStreamIsCapturing(st, &cap);
  if (ce != cudaSuccess) return (int)ce;
  const RowsPath path = rows_path(b, n1, n2, row_lo, row_hi, di.sms, cap != cudaStreamCaptureStatusNone);
  if (path == kRowsSwapped)
    return forward_rows_swapped_impl(c1, n1, c2, n2, b, row_lo, row_hi, dist1, idx1, colkeys, workspace, di, ev, st);
  char* w = (char*)workspace;
  float* soa1 = (float*)w;
  w += soa_bytes(b, n1);
  float* soa2 = (float*)w;
  w += soa_bytes(b, n2);
  unsigned long long* rkeys = (unsigned long long*)w;
  w += key_bytes(b, n1);
  unsigned* cdist = (unsigned*)w;
  char* hint_scratch = (char*)workspace + rows_core_need(b, n1, n2);
  const int nrows = row_hi - row_lo;

  PrepParams pp{};
  pp.nclouds = nrows > 0 ? 2 : 0;
  pp.c[0] = PrepCloud{c1, soa1, n1, tiles_of(n1), (long long)b * tiles_of(n1) * kTT, n1};
  pp.c[1] = PrepCloud{c2, soa2, n2, tiles_of(n2), (long long)b * tiles_of(n2) * kTT, n2};
  pp.fill64[0] = rkeys;
  pp.count64[0] = nrows > 0 ? (long long)b * n1 : 0;
  pp.fill64[1] = colkeys;
  pp.count64[1] = (long long)colkey_count(b, n2);
  pp.fill32 = cdist;
  pp.count32 = nrows > 0 ? (long long)colkey_count(b, n2) : 0;
  {
    long long total = pp.c[0].padded_total > pp.c[1].padded_total ? pp.c[0].padded_total : pp.c[1].padded_total;
    if (ev) cudaEventRecord(ev[0], st);
    nnd_prepare_kernel<kTT><<<grid_for(total, di.sms), 256, 0, st>>>(pp);
    ++g_launches;
    if (ev) cudaEventRecord(ev[1], st);
    debug_sync("prepare", st);
  }
  if (nrows <= 0) {  // nothing to scan: the column keys stay at their initial value
    if (ev) {
      cudaEventRecord(ev[2], st);
      cudaEventRecord(ev[3], st);
    }
    return check_last(cudaGetLastError());
  }

  SymParams yp{};
  yp.qsoa = soa1;
  yp.tsoa = soa2;
  yp.rkeys = rkeys;
  yp.ckeys = colkeys;
  yp.cdist = cdist;
  yp.n = n1;
  yp.m = n2;
  yp.row_lo = row_lo;
  yp.row_hi = row_hi;
  yp.qtiles = tiles_of(n1);
  yp.ttiles = tiles_of(n2);
  const ScanShape shape = sym_shape(b, n1, n2, di.sms);
  const ShapeInfo si = kShapes[shape];
  const int qb = si.threads * si.qpt;
  yp.qblocks = (nrows + qb - 1) / qb;
  yp.total_units = (long long)b * yp.qblocks * yp.ttiles;
  yp.one = 1.0f;
  const int grid_full = di.sms * si.minb;
  const int grid = yp.total_units < grid_full ? (int)yp.total_units : grid_full;
  const size_t smem = sizeof(SymStage<kTT>) * kStages + (size_t)si.threads * si.qpt * 8;  // ring + row bookkeeping
  const bool panels = path == kRowsPanel;
  // Cold start: automatic hints from a hash grid of the rows (nnd_hint.cuh), unless the caller brought its own.
  // Hints name rows of the WHOLE cloud; the seed pass ignores those outside [row_lo, row_hi).
  const int auto_hint = g_auto_hint.load(std::memory_order_relaxed);
  if (hint2 == nullptr && shape == kShapeLarge && (auto_hint >= 2 || (auto_hint == 1 && panels)))
    hint2 = make_hints(c1, n1, n1, c2, n2, n2, b, hint_scratch, di.sms, st);
  // The panel path's filter seed in front of the persistent kernel as well (nnd_set_sym_seed_tiles; default 1 tile).
  const int sym_seed = g_sym_seed.load(std::memory_order_relaxed);
  if (!panels && (sym_seed > 0 || hint2 != nullptr) && shape == kShapeLarge) {
    const int t_first = (row_lo + kTT - 1) / kTT;
    int ns = row_hi / kTT - t_first;
    ns = ns > sym_seed ? sym_seed : ns;
    if (ns > 0 || hint2 != nullptr) {  // with a hint (warm start) the seed runs even without a whole sample tile
      dim3 g((unsigned)((n2 + 511) / 512), (unsigned)b);
      nnd_seed_filter_kernel<kTT><<<g, 256, 0, st>>>(c2, soa1, cdist, n2, yp.ttiles, yp.qtiles, t_first, ns < 0 ? 0 : ns,
                                                     1.0f, n2, hint2, c1, n1, row_lo, row_hi);
      ++g_launches;
    }
  }
  if (panels) {
    SymCParams cp{};
    cp.qsoa = soa1;
    cp.rkeys = rkeys;
    cp.ckeys = colkeys;
    cp.cdist = cdist;
    cp.n = n1;
    cp.m = n2;
    cp.row_lo = row_lo;
    cp.row_hi = row_hi;
    cp.qtiles = yp.qtiles;
    cp.ttiles = yp.ttiles;
    cp.one = 1.0f;
    {  // seed the column filter from whole sample tiles of rows inside [row_lo, row_hi): ~0.2-0.4 % of the scan
      const int t_first = (row_lo + kTT - 1) / kTT;
      int ns = row_hi / kTT - t_first;
      const int want = nrows >= 500000 ? 4 : nrows >= 200000 ? 2 : 1;
      ns = ns > want ? want : ns;
      if (ns > 0 || hint2 != nullptr) {
        dim3 g((unsigned)((n2 + 511) / 512), (unsigned)b);
        nnd_seed_filter_kernel<kTT><<<g, 256, 0, st>>>(c2, soa1, cdist, n2, yp.ttiles, yp.qtiles, t_first, ns < 0 ? 0 : ns,
                                                       1.0f, n2, hint2, c1, n1, row_lo, row_hi);
        ++g_launches;
      }
    }
    rc = scan_panels(cp, soa2, b, yp.qblocks, di.sms, st);
  } else
  switch (shape) {
    case kShapeLarge: rc = launch_sym<kThreads, kQPT, kMinBlocks>(yp, grid, smem, st); break;
    case kShapeMid: rc = launch_sym<256, 2, 4>(yp, grid, smem, st); break;
    default: rc = launch_sym<128, 2, 8>(yp, grid, smem, st); break;
  }
  if (rc) return rc;
  if (!panels) ++g_launches;
  if (ev) cudaEventRecord(ev[2], st);
  debug_sync("scan_sym", st);

  FinParams fp{};
  fp.ndir = 1;
  fp.one = 1.0f;
  fp.d[0] = FinDir{rkeys, c1, soa2, dist1, idx1, n1, tiles_of(n2), row_lo, nrows, (long long)b * nrows};
  if (panels)
    nnd_finalize_kernel<kTT, kSymcFinChunk><<<grid_for(fp.d[0].count, di.sms), 256, 0, st>>>(fp);
  else
    nnd_finalize_kernel<kTT, kChunk><<<grid_for(fp.d[0].count, di.sms), 256, 0, st>>>(fp);
  ++g_launches;
  if (ev) cudaEventRecord(ev[3], st);
  debug_sync("finalize_rows", st);
  return check_last(cudaGetLastError());
}

int finalize_cols_impl(const float* c1, int n1, const float* c2, int n2, int b, const unsigned long long* colkeys,
                       float* dist2, int32_t* idx2, cudaStream_t st) {
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  FinColParams cp{colkeys, c1, c2, dist2, idx2, n1, n2, tiles_of(n2), (long long)b * n2, 1.0f, n2, 0};
  if (kShapes[sym_shape(b, n1, n2, di.sms)].qpt == kQPT)
    nnd_finalize_cols_kernel<kTT, kQPT><<<grid_for(cp.count, di.sms), 256, 0, st>>>(cp);
  else
    nnd_finalize_cols_kernel<kTT, 2><<<grid_for(cp.count, di.sms), 256, 0, st>>>(cp);
  ++g_launches;
  debug_sync("finalize_cols", st);
  return check_last(cudaGetLastError());
}

// One direction (nnd_query_cuda_f32): prepare + nnd_scan_kernel + finalize.
// Two-pass forward: every direction scans the other cloud separately (nnd_scan_kernel; no column bookkeeping).
// ndir == 1: for every q point its nearest t point.  ndir == 2: additionally for every t point its nearest q point
// (dist2/idx2); used for clouds too small for the symmetric kernel to pay off (see use_symmetric()).
int twopass_impl(const float* q, int nq, const float* t, int nt, int b, int ndir, float* dist, int32_t* idx,
                 float* dist2, int32_t* idx2, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  cudaEvent_t* ev = nullptr;
  if (g_prof.enabled && g_prof.used < kProfSlots) ev = &g_prof.ev[4 * g_prof.used++];
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  if (((uintptr_t)workspace & 255u) != 0) return NND_EWORKSPACE;
  const size_t need = soa_bytes(b, nq) + soa_bytes(b, nt) + key_bytes(b, nq) + (ndir == 2 ? key_bytes(b, nt) : 0);
  if (workspace_bytes < need) return NND_EWORKSPACE;
  char* w = (char*)workspace;
  float* soaq = (float*)w;
  w += soa_bytes(b, nq);
  float* soat = (float*)w;
  w += soa_bytes(b, nt);
  unsigned long long* keys = (unsigned long long*)w;
  w += key_bytes(b, nq);
  unsigned long long* keys2 = ndir == 2 ? (unsigned long long*)w : nullptr;

  PrepParams pp{};
  pp.nclouds = 2;
  pp.c[0] = PrepCloud{q, soaq, nq, tiles_of(nq), (long long)b * tiles_of(nq) * kTT, nq};
  pp.c[1] = PrepCloud{t, soat, nt, tiles_of(nt), (long long)b * tiles_of(nt) * kTT, nt};
  pp.fill64[0] = keys;
  pp.count64[0] = (long long)b * nq;
  pp.fill64[1] = keys2;
  pp.count64[1] = ndir == 2 ? (long long)b * nt : 0;
  {
    long long total = pp.c[0].padded_total > pp.c[1].padded_total ? pp.c[0].padded_total : pp.c[1].padded_total;
    if (ev) cudaEventRecord(ev[0], st);
    nnd_prepare_kernel<kTT><<<grid_for(total, di.sms), 256, 0, st>>>(pp);
    ++g_launches;
    if (ev) cudaEventRecord(ev[1], st);
  }
  const ScanShape shape = choose_shape(b, nq, nt, di.sms);
  const ShapeInfo si = kShapes[shape];
  const int qb = si.threads * si.qpt;
  ScanParams sp{};
  sp.ndir = ndir;
  sp.one = 1.0f;
  sp.d[0].qsoa = soaq;
  sp.d[0].tsoa = soat;
  sp.d[0].keys = keys;
  sp.d[0].nq = nq;
  sp.d[0].nt = nt;
  sp.d[1].qsoa = soat;
  sp.d[1].tsoa = soaq;
  sp.d[1].keys = keys2;
  sp.d[1].nq = nt;
  sp.d[1].nt = nq;
  sp.total_units = 0;
  for (int i = 0; i < ndir; ++i) {
    ScanDir& d = sp.d[i];
    d.qtiles = tiles_of(d.nq);
    d.ttiles = tiles_of(d.nt);
    d.qblocks = (d.nq + qb - 1) / qb;
    d.units = (long long)b * d.qblocks * d.ttiles;
    sp.total_units += d.units;
  }
  const int grid_full = di.sms * si.minb;
  const int grid = sp.total_units < grid_full ? (int)sp.total_units : grid_full;
  switch (shape) {
    case kShapeLarge: launch_scan<kThreads, kQPT, kMinBlocks>(sp, grid, st); break;
    case kShapeMid: launch_scan<256, 2, 4>(sp, grid, st); break;
    default: launch_scan<128, 2, 8>(sp, grid, st); break;
  }
  ++g_launches;
  if (ev) cudaEventRecord(ev[2], st);
  FinParams fp{};
  fp.ndir = ndir;
  fp.one = 1.0f;
  fp.d[0] = FinDir{keys, q, soat, dist, idx, nq, tiles_of(nt), 0, nq, (long long)b * nq};
  if (ndir == 2) fp.d[1] = FinDir{keys2, t, soaq, dist2, idx2, nt, tiles_of(nq), 0, nt, (long long)b * nt};
  const long long most = ndir == 2 && fp.d[1].count > fp.d[0].count ? fp.d[1].count : fp.d[0].count;
  nnd_finalize_kernel<kTT, kChunk><<<grid_for(most, di.sms), 256, 0, st>>>(fp);
  ++g_launches;
  if (ev) cudaEventRecord(ev[3], st);
  return check_last(cudaGetLastError());
}

// The symmetric kernel evaluates every distance once but pays a warp-collective slow path whenever a column
// minimum improves: about ln(rows / rows-per-warp-step) times per column.  That is < 1 % of the steps for large clouds
// and most of them for small ones (measured: 86 us at N=M=2048 against 13 us of arithmetic), so small clouds take the
// two-pass kernel (twice the arithmetic, no bookkeeping).
inline bool use_symmetric(int n, int m) { return n >= 8192 && m >= 8192; }

// ---- small clouds: single-launch forward / backward (nnd_small.cuh) ---------------------------------------------
// Clouds of at most this many points (both of them) take nnd_small_kernel / nnd_grad_small_kernel.
// NND_B200_SMALL_MAX overrides; nnd_set_small_max_points() is the run-time knob tests use to force either path.
constexpr int kSmallHardMax = 16384;  // 3 * 16384 * 4 B = 192 KB of the 227 KB a CTA may have
std::atomic<int> g_small_max{-1};
int small_max_points() {
  int v = g_small_max.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("NND_B200_SMALL_MAX");
    int init = e ? atoi(e) : 4096;
    init = init < 0 ? 0 : init > kSmallHardMax ? kSmallHardMax : init;
    int expect = -1;
    g_small_max.compare_exchange_strong(expect, init);
    v = g_small_max.load(std::memory_order_relaxed);
  }
  return v;
}
// Measured crossover (profiles/r02_small_time.txt): up to 4096 points the single launch wins at every batch size; a
// single cloud pair (b * points small) keeps winning up to ~10 k points, where the general path still runs at a
// tenth of the machine (64 work units for 148 SMs).
inline bool use_small(int b, int n, int m) {
  const long long mx = n > m ? n : m, lim = small_max_points();
  if (mx <= lim) return true;
  const long long lim2 = lim * 5 / 2 < kSmallHardMax ? lim * 5 / 2 : kSmallHardMax;
  return mx <= lim2 && (long long)b * mx <= lim * 3;
}
// Ragged batches and the one-launch backward key on the cloud sizes alone (no batch term): <= small_max_points().
inline bool use_small_nm(int n, int m) { return (n > m ? n : m) <= small_max_points(); }
// CTAs per batch element of the shared-memory backward: one when the batch alone fills the machine (every extra part
// re-reads the whole index arrays), up to 16 for a lone pair.
inline int grad_small_parts(int b, int n, int m, int sms) {
  long long parts = sms / (b > 0 ? b : 1);
  const long long by_rows = ((long long)n + m) / 256;  // at least ~256 owned rows per CTA
  if (parts > by_rows) parts = by_rows;
  if (parts > 16) parts = 16;
  return parts < 1 ? 1 : (int)parts;
}

struct SmallCfg {
  int warps, qpt, tsplit;
};
std::atomic<int> g_small_cfg{0};  // 0 = choose from the problem size; else warps << 16 | qpt << 8 | tsplit
// total_q: queries of the whole launch; nq_max / nt_max: the larger cloud in the query / target role.
SmallCfg small_config(long long total_q, int nq_max, int nt_max, int sms) {
  const int forced = g_small_cfg.load(std::memory_order_relaxed);  // nnd_set_small_config: tests and tuning only
  if (forced > 0) return SmallCfg{forced >> 16, (forced >> 8) & 0xff, forced & 0xff};
  SmallCfg c;
  // two queries per lane halve the shared-memory reads per pair; worth it once every scheduler still gets >= 2 warps
  c.qpt = total_q >= (long long)sms * 4 * 4 * 64 ? 2 : 1;
  const long long qwarps = (total_q + 32 * c.qpt - 1) / (32 * c.qpt);
  const int chunks = (nt_max + kSmallChunk - 1) / kSmallChunk;
  int ts = 1;  // split the targets over warps until the machine has ~2 warps per scheduler (each split keeps >= 2 chunks)
  while (ts < 16 && qwarps * ts < 2ll * sms * 4 && chunks / (ts * 2) >= 2) ts *= 2;
  c.tsplit = ts;
  const int qgroups = (nq_max + 32 * c.qpt - 1) / (32 * c.qpt);  // query groups that cover the larger query cloud
  // 4-warp CTAs when they cover the query cloud anyway, or when the target cloud is so small that staging it per CTA
  // costs nothing and more, smaller CTAs balance better over the SMs (DIP shape: 14.6 us against 15.6, r02_small_time_1.txt)
  c.warps = ts == 16 ? 16 : ((long long)ts * qgroups <= 4 || (nt_max <= 512 && ts <= 4)) ? 4 : 8;
  return c;
}

template <int WARPS, int QPT>
int launch_small_t(const SmallParams& sp, int grid, size_t smem, cudaStream_t st) {
  auto kern = nnd_small_kernel<WARPS, QPT>;
  if (smem > 48 * 1024) {
    static SmemOptIn optin;
    int rc = optin.ensure(kern, smem);
    if (rc) return rc;
  }
  kern<<<grid, WARPS * 32, smem, st>>>(sp);
  return 0;
}

// ndir == 2: both directions (nnd_forward);  ndir == 1: cloud 0 queries cloud 1 only (nnd_query).
int small_forward_impl(const float* c0, int n0, const float* c1, int n1, int b, int ndir, const long long* len0,
                       const long long* len1, float* dist0, int32_t* idx0, float* dist1, int32_t* idx1,
                       cudaStream_t st) {
  cudaEvent_t* ev = nullptr;
  if (g_prof.enabled && g_prof.used < kProfSlots) ev = &g_prof.ev[4 * g_prof.used++];
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  const long long total_q = (long long)b * (n0 + (ndir == 2 ? n1 : 0));
  const int nq_max = ndir == 2 && n1 > n0 ? n1 : n0;
  const int nt_max = ndir == 2 && n0 > n1 ? n0 : n1;
  const SmallCfg cfg = small_config(total_q, nq_max, nt_max, di.sms);
  const int qc = (cfg.warps / cfg.tsplit) * 32 * cfg.qpt;
  SmallParams sp{};
  sp.xyz[0] = c0;
  sp.xyz[1] = c1;
  sp.dist[0] = dist0;
  sp.dist[1] = dist1;
  sp.idx[0] = idx0;
  sp.idx[1] = idx1;
  sp.len[0] = len0;
  sp.len[1] = len1;
  sp.cnt[0] = n0;
  sp.cnt[1] = n1;
  sp.qblocks[0] = (n0 + qc - 1) / qc;
  sp.qblocks[1] = ndir == 2 ? (n1 + qc - 1) / qc : 0;
  sp.blocks0 = b * sp.qblocks[0];
  sp.tsplit = cfg.tsplit;
  sp.mp_max = (nt_max + kSmallChunk - 1) / kSmallChunk * kSmallChunk;
  sp.one = 1.0f;
  const long long grid = (long long)b * (sp.qblocks[0] + sp.qblocks[1]);
  if (grid > 0x7fffffffLL) return NND_EINVAL;
  const size_t smem = (size_t)3 * sp.mp_max * sizeof(float) + (cfg.tsplit > 1 ? (size_t)qc * 8 : 0);
  if (ev) {
    cudaEventRecord(ev[0], st);
    cudaEventRecord(ev[1], st);
  }
#define NND_SMALL_CASE(W, Q) \
  if (cfg.warps == W && cfg.qpt == Q) rc = launch_small_t<W, Q>(sp, (int)grid, smem, st); else
  NND_SMALL_CASE(2, 1) NND_SMALL_CASE(2, 2) NND_SMALL_CASE(2, 4) NND_SMALL_CASE(2, 8)
  NND_SMALL_CASE(4, 1) NND_SMALL_CASE(4, 2) NND_SMALL_CASE(4, 4) NND_SMALL_CASE(4, 8)
  NND_SMALL_CASE(8, 1) NND_SMALL_CASE(8, 2) NND_SMALL_CASE(8, 4) NND_SMALL_CASE(8, 8)
  NND_SMALL_CASE(16, 1) NND_SMALL_CASE(16, 2) NND_SMALL_CASE(16, 4) NND_SMALL_CASE(16, 8)
  rc = NND_EINVAL;
#undef NND_SMALL_CASE
  if (rc) return rc;
  ++g_launches;
  if (ev) {
    cudaEventRecord(ev[2], st);
    cudaEventRecord(ev[3], st);
  }
  debug_sync("small_forward", st);
  return check_last(cudaGetLastError());
}

template <int THREADS>
int launch_grad_small_t(const GradSmallParams& gp, long long grid, size_t smem, cudaStream_t st) {
  auto kern = nnd_grad_small_kernel<THREADS>;
  if (smem > 48 * 1024) {
    static SmemOptIn optin;
    int rc = optin.ensure(kern, smem);
    if (rc) return rc;
  }
  kern<<<(unsigned)grid, THREADS, smem, st>>>(gp);
  return 0;
}

int small_backward_impl(GradSmallParams gp, int b, cudaStream_t st) {
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  gp.parts = grad_small_parts(b, gp.cnt[0], gp.cnt[1], di.sms);
  const int own = (gp.cnt[0] + gp.parts - 1) / gp.parts + (gp.cnt[1] + gp.parts - 1) / gp.parts;  // rows a CTA owns
  const size_t smem = (size_t)3 * own * sizeof(float);
  const long long grid = (long long)b * gp.parts;
  if (grid > 0x7fffffffLL || smem > 220 * 1024) return NND_EINVAL;
  // every CTA walks ALL rows of both clouds in phase 1 (dependent gathers behind each index): enough threads that a
  // thread sees at most ~4 rows, so its loads are one batch in flight instead of a chain of batches
  const long long scan = (long long)gp.cnt[0] + gp.cnt[1];
  rc = scan > 2048 ? launch_grad_small_t<1024>(gp, grid, smem, st)
     : scan > 1024 ? launch_grad_small_t<512>(gp, grid, smem, st)
                   : launch_grad_small_t<256>(gp, grid, smem, st);
  if (rc) return rc;
  ++g_launches;
  debug_sync("small_backward", st);
  return check_last(cudaGetLastError());
}

int zero_outputs(float* dist, int32_t* idx, long long count, cudaStream_t st) {
  if (count <= 0) return 0;
  cudaError_t e = cudaMemsetAsync(dist, 0, (size_t)count * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemsetAsync(idx, 0, (size_t)count * sizeof(int32_t), st);
  return check_last(e);
}

int launch_grad(const GradParams& gp, cudaStream_t st) {
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  long long rows = 0;
  for (int i = 0; i < gp.nterms; ++i) rows = gp.t[i].rows > rows ? gp.t[i].rows : rows;
  if (rows == 0) return 0;
  const long long span = (long long)kGradThreads * kGradRows;
  long long blocks = (rows + span - 1) / span;
  if (blocks > (long long)di.sms * 8) blocks = (long long)di.sms * 8;
  nnd_grad_kernel<<<(unsigned)blocks, kGradThreads, 0, st>>>(gp);
  ++g_launches;
  return check_last(cudaGetLastError());
}

}  // namespace

// ----------------------------------------------------------------------------------
// C ABI
// ----------------------------------------------------------------------------------
// ---- exact search with spatial pruning (nnd_pruned.cuh; opt-in entry point nnd_forward_pruned_cuda_f32) ------------
namespace {
inline int pr_tiles(int n) { return (n + kPrTile - 1) / kPrTile; }
inline int pr_stiles(int n) { return (pr_tiles(n) + kPrSuper - 1) / kPrSuper; }
inline int pr_ltiles(int n) { return (pr_stiles(n) + kPrSuper - 1) / kPrSuper; }
struct PrSizes {
  size_t keys, pts, tbox, sbox, lbox, tcode;
  size_t total() const { return 2 * keys + pts + tbox + sbox + lbox + tcode; }
};
inline PrSizes pr_sizes(int b, int n) {
  PrSizes z;
  z.keys = align_up((size_t)b * n * sizeof(unsigned long long), 256);
  z.pts = align_up((size_t)b * pr_tiles(n) * kPrTile * sizeof(float4), 256);
  z.tbox = align_up((size_t)b * pr_tiles(n) * sizeof(PrBox), 256);
  z.sbox = align_up((size_t)b * pr_stiles(n) * sizeof(PrBox), 256);
  z.lbox = align_up((size_t)b * pr_ltiles(n) * sizeof(PrBox), 256);
  z.tcode = align_up((size_t)b * pr_tiles(n) * sizeof(unsigned), 256);
  return z;
}
inline size_t pr_temp_bytes(int b, int n, int m) {
  const size_t t1 = exact_sort_temp_bytes((long long)b * n), t2 = exact_sort_temp_bytes((long long)b * m);
  return align_up(t1 > t2 ? t1 : t2, 256) + 256;
}
inline size_t pr_need(int b, int n, int m) {
  return pr_sizes(b, n).total() + pr_sizes(b, m).total() + 256 /* joint boxes */ + align_up((size_t)b * 24, 256) +
         pr_temp_bytes(b, n, m);
}

// Query groups (b * ceil(n / 32)) below which the warps of a CTA share a group (nnd_pr_search_kernel<kPrWarps>).
constexpr long long kPrShareBelow = 16384;
std::atomic<int> g_pr_share{0};  // nnd_set_pruned_share: 0 = by size (default), 1 = never, 2 = always

// Sort one cloud along the curve and build its tiles; `keys_in` is scratch.  Returns the first CUDA error.
// keep_order: the sorted keys of a previous call stay as they are (nnd_forward_pruned_reuse_cuda_f32); the points and
// every box are rebuilt from the cloud's CURRENT coordinates, so a stale order costs pruning power, never exactness.
int pr_prepare(PrCloud& c, int b, unsigned long long* keys_in, const unsigned* blo, const unsigned* bhi, int kb, int end_bit,
               void* temp, size_t temp_bytes, int sms, cudaStream_t st, bool keep_order) {
  const long long count = (long long)b * c.n;
  long long bl = ((long long)c.n + 255) / 256;
  const unsigned blocks = (unsigned)(bl > 8ll * sms ? 8ll * sms : bl < 1 ? 1 : bl);
  if (!keep_order) {
    nnd_pr_code_kernel<<<dim3(blocks, (unsigned)b), 256, 0, st>>>(c.xyz, c.n, blo, bhi, kb, keys_in);
    ++g_launches;
    cudaError_t e = cub::DeviceRadixSort::SortKeys(temp, temp_bytes, (const unsigned long long*)keys_in, c.keys, (int)count,
                                                   32, end_bit, st);
    if (e != cudaSuccess) return (int)e;
  }
  nnd_pr_gather_kernel<<<dim3((unsigned)c.tiles, (unsigned)b), kPrTile, 0, st>>>(c);
  nnd_pr_super_kernel<<<dim3((unsigned)((c.stiles + 3) / 4), (unsigned)b), 128, 0, st>>>(c.tbox, c.tiles, c.sbox, c.stiles);
  nnd_pr_super_kernel<<<dim3((unsigned)((c.ltiles + 3) / 4), (unsigned)b), 128, 0, st>>>(c.sbox, c.stiles, c.lbox, c.ltiles);
  g_launches += 3;  // + the library's sort passes
  return check_last(cudaGetLastError());
}

// ndir == 2: both directions; ndir == 1: only cloud 1's points against cloud 2 (nnd_query_pruned_cuda_f32).
// reuse (nnd_forward_pruned_reuse_cuda_f32): bit 0 = keep cloud 1's curve order, bit 1 = cloud 2 unchanged (all of its
// preparation in the workspace is still valid).
int pruned_forward_impl(const float* xyz1, int n, const float* xyz2, int m, int b, int ndir, float* dist1, int32_t* idx1,
                        float* dist2, int32_t* idx2, void* workspace, size_t workspace_bytes, cudaStream_t st,
                        int reuse = 0) {
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  if (((uintptr_t)workspace & 255u) != 0 || workspace_bytes < pr_need(b, n, m)) return NND_EWORKSPACE;
  const int batch_bits = bits_for((unsigned long long)(b - 1));
  int kb = (32 - batch_bits) / 3;
  kb = kb > 10 ? 10 : kb;
  if (kb < 1 || b > 65535) return NND_EINVAL;  // grid.y carries the batch
  const int end_bit = 32 + 3 * kb + batch_bits;

  char* w = (char*)workspace;
  PrCloud c[2];
  const float* src[2] = {xyz1, xyz2};
  const int cnt[2] = {n, m};
  for (int i = 0; i < 2; ++i) {
    const PrSizes z = pr_sizes(b, cnt[i]);
    c[i].xyz = src[i];
    c[i].n = cnt[i];
    c[i].tiles = pr_tiles(cnt[i]);
    c[i].stiles = pr_stiles(cnt[i]);
    c[i].ltiles = pr_ltiles(cnt[i]);
    c[i].keys = (unsigned long long*)w;
    w += z.keys;
    c[i].pts = (float4*)w;
    w += z.pts;
    c[i].tbox = (PrBox*)w;
    w += z.tbox;
    c[i].sbox = (PrBox*)w;
    w += z.sbox;
    c[i].lbox = (PrBox*)w;
    w += z.lbox;
    c[i].tcode = (unsigned*)w;
    w += z.tcode;
  }
  const size_t kmax = pr_sizes(b, n > m ? n : m).keys;
  unsigned long long* keys_in = (unsigned long long*)w;  // shared by both clouds (one sorts after the other)
  w += kmax;
  // (the second cloud's "2 * keys" share covers the smaller cloud's unused input half: total() counts 2 per cloud)
  unsigned* blo = (unsigned*)w;
  unsigned* bhi = blo + (size_t)b * 3;
  w += 256 + align_up((size_t)b * 24, 256);
  void* temp = w;
  const size_t temp_bytes = pr_temp_bytes(b, n, m);

  const bool keep1 = (reuse & 1) != 0, same2 = (reuse & 2) != 0;
  if (!same2) {  // a new joint box; with cloud 2 reused the old one stays (its codes were cut with it; points of cloud 1
                 // that left it clamp to the border cells)
    cudaError_t e = cudaMemsetAsync(blo, 0xff, (size_t)b * 3 * sizeof(unsigned), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(bhi, 0, (size_t)b * 3 * sizeof(unsigned), st);
    if (e != cudaSuccess) return (int)e;
    for (int i = 0; i < 2; ++i) {
      long long bl = ((long long)cnt[i] + 255) / 256;
      const unsigned blocks = (unsigned)(bl > 8ll * di.sms ? 8ll * di.sms : bl < 1 ? 1 : bl);
      nnd_pr_bbox_kernel<<<dim3(blocks, (unsigned)b), 256, 0, st>>>(src[i], cnt[i], blo, bhi);
      ++g_launches;
    }
  }
  rc = pr_prepare(c[0], b, keys_in, blo, bhi, kb, end_bit, temp, temp_bytes, di.sms, st, keep1);
  if (rc) return rc;
  if (!same2) {
    rc = pr_prepare(c[1], b, keys_in, blo, bhi, kb, end_bit, temp, temp_bytes, di.sms, st, false);
    if (rc) return rc;
  }
  for (int dir = 0; dir < ndir; ++dir) {
    PrSearch sp{c[dir], c[1 - dir], dir == 0 ? dist1 : dist2, dir == 0 ? idx1 : idx2, 1.0f};
    // The 4 warps of a CTA share a group of 32 queries when one warp per group would not fill the machine, and
    // whenever the target cloud is large (many tiles per group: sharing evens out the warps' very different tile
    // counts); small clouds in large batches keep one warp per group (a few tiles each: the shared seed tile and the
    // repeated box tests would dominate).  Measured, profiles/r02_pruned_split_ab.txt: C3 0.67 -> 0.48 ms, C5
    // 0.87-1.42 -> 0.74-0.88 ms, C4 2.62 / 2.86 -> 2.41 / 2.24 ms; C2 0.95 -> 1.02 ms if it shared.
    const long long groups = ((long long)sp.q.n + 31) / 32;
    const int share_mode = g_pr_share.load(std::memory_order_relaxed);
    const bool shared_groups = share_mode == 2 || (share_mode == 0 && (groups * b < kPrShareBelow || sp.t.n >= 65536));
    if (!shared_groups) {
      const unsigned gx = (unsigned)((groups + kPrWarps - 1) / kPrWarps);
      nnd_pr_search_kernel<1><<<dim3(gx, (unsigned)b), 32 * kPrWarps, 0, st>>>(sp);
    } else {
      nnd_pr_search_kernel<kPrWarps><<<dim3((unsigned)groups, (unsigned)b), 32 * kPrWarps, 0, st>>>(sp);
    }
    ++g_launches;
  }
  debug_sync("pruned_forward", st);
  return check_last(cudaGetLastError());
}
}  // namespace

// ---- no-allocation entry point (nnd_query_noalloc_cuda_f32) ------------------------------------------------------
// Scratch of the two-pass query in a file-scope device array instead of caller memory: every device that loads the
// library holds these NND_NOALLOC_SCRATCH_BYTES until the process exits.  In host code `g_noalloc_scratch` is only a
// host-side stand-in for the symbol, NOT a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, looked up once per device with that device current (noalloc_scratch() below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[NND_NOALLOC_SCRATCH_BYTES];

namespace {
int noalloc_scratch(void** out) {
  static void* addr[64];
  static std::atomic<bool> ready[64];
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return NND_EINVAL;
  if (!ready[dev].load(std::memory_order_acquire)) {
    std::lock_guard<std::mutex> lock(mu);
    if (!ready[dev].load(std::memory_order_relaxed)) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // the current device's copy of the array
      if (e != cudaSuccess) return (int)e;
      if (p == nullptr || ((uintptr_t)p & 255u) != 0) return NND_EWORKSPACE;
      addr[dev] = p;
      ready[dev].store(true, std::memory_order_release);
    }
  }
  *out = addr[dev];
  return 0;
}
}  // namespace

extern "C" {

const char* nnd_version(void) { return NND_VERSION_STRING; }

const char* nnd_error_string(int code) {
  switch (code) {
    case 0: return "success";
    case NND_EINVAL: return "nnd: invalid argument";
    case NND_EWORKSPACE: return "nnd: workspace too small or not 256-byte aligned";
    case NND_EDEVICE: return "nnd: current device is not compute capability 10.x (B200)";
    case NND_ETOOBIG:
      return "nnd: problem needs more scratch than the built-in NND_NOALLOC_SCRATCH_BYTES of the no-allocation entry "
             "point; use nnd_query_cuda_f32 with a caller-owned workspace";
    default: return cudaGetErrorString((cudaError_t)code);
  }
}

size_t nnd_workspace_bytes(int b, int n, int m) {
  if (b <= 0 || n <= 0 || m <= 0) return 256;
  if (use_small(b, n, m)) return 0;  // single-launch path: no workspace (the pointer may be null)
  return nnd_rows_workspace_bytes(b, n, m) + colkey_bytes(b, m);
}

size_t nnd_rows_workspace_bytes(int b, int n, int m) {
  if (b <= 0 || n <= 0 || m <= 0) return 256;
  return rows_need(b, n, m);
}

size_t nnd_colkeys_count(int b, int m) {
  if (b <= 0 || m <= 0) return 0;
  return colkey_count(b, m);
}

int nnd_row_group(void) { return kRowAlign; }

int nnd_forward_is_symmetric(int n, int m) { return !use_small(1 << 20, n, m) && use_symmetric(n, m) ? 1 : 0; }

int nnd_forward_kind(int b, int n, int m) { return use_small(b, n, m) ? 0 : use_symmetric(n, m) ? 2 : 1; }

int nnd_small_max_points(void) { return small_max_points(); }

int nnd_set_small_config(int warps, int qpt, int tsplit) {
  if (warps == 0 && qpt == 0 && tsplit == 0) {
    g_small_cfg.store(0);
    return 0;
  }
  if (!(warps == 2 || warps == 4 || warps == 8 || warps == 16) || !(qpt == 1 || qpt == 2 || qpt == 4 || qpt == 8) ||
      tsplit < 1 || warps % tsplit != 0)
    return NND_EINVAL;
  g_small_cfg.store(warps << 16 | qpt << 8 | tsplit);
  return 0;
}

int nnd_set_small_max_points(int points) {
  const int before = small_max_points();
  if (points >= 0) g_small_max.store(points > kSmallHardMax ? kSmallHardMax : points);
  return before;
}

int nnd_rows_scan_kind(int b, int n, int m, int row_lo, int row_hi) {
  DeviceInfo di;
  const int sms = device_info(&di) == 0 ? di.sms : 148;  // callable without a device (tests, planning): B200 has 148 SMs
  return (int)rows_path(b, n, m, row_lo, row_hi, sms, false);
}

int nnd_set_auto_hint(int mode) {
  const int before = g_auto_hint.load();
  if (mode >= 0 && mode <= 2) g_auto_hint.store(mode);
  return before;
}

int nnd_set_sym_seed_tiles(int tiles) {
  const int before = g_sym_seed.load();
  if (tiles >= 0) g_sym_seed.store(tiles > 64 ? 64 : tiles);
  return before;
}

size_t nnd_query_workspace_bytes(int b, int nq, int nt) {
  if (b <= 0 || nq <= 0 || nt <= 0) return 256;
  if (use_small(b, nq, nt)) return 0;
  return soa_bytes(b, nq) + soa_bytes(b, nt) + key_bytes(b, nq);
}

int nnd_last_launch_count(void) { return g_launches; }

static bool sizes_ok(int b, int n, int m) {
  if (b < 0 || n < 0 || m < 0) return false;
  return (long long)b * n <= 0x7fffffffLL / 3 * 256 && (long long)b * m <= 0x7fffffffLL / 3 * 256;
}

int nnd_forward_cuda_f32(const float* xyz1, const float* xyz2, int b, int n, int m, float* dist1, float* dist2,
                         int32_t* idx1, int32_t* idx2, void* workspace, size_t workspace_bytes, void* stream) {
  return nnd_forward_warm_cuda_f32(xyz1, xyz2, b, n, m, dist1, dist2, idx1, idx2, nullptr, workspace, workspace_bytes,
                                   stream);
}

int nnd_forward_warm_cuda_f32(const float* xyz1, const float* xyz2, int b, int n, int m, float* dist1, float* dist2,
                              int32_t* idx1, int32_t* idx2, const int32_t* hint_idx2, void* workspace,
                              size_t workspace_bytes, void* stream) {
  g_launches = 0;
  if (!sizes_ok(b, n, m)) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0 || (n == 0 && m == 0)) return 0;
  if (n == 0) return zero_outputs(dist2, idx2, (long long)b * m, st);  // my_lib.cpp:9-10,21-22 with an empty inner loop
  if (m == 0) return zero_outputs(dist1, idx1, (long long)b * n, st);
  if (!xyz1 || !xyz2 || !dist1 || !dist2 || !idx1 || !idx2) return NND_EINVAL;
  if (use_small(b, n, m))
    return small_forward_impl(xyz1, n, xyz2, m, b, 2, nullptr, nullptr, dist1, idx1, dist2, idx2, st);
  if (!workspace) return NND_EINVAL;
  if (workspace_bytes < nnd_workspace_bytes(b, n, m)) return NND_EWORKSPACE;
  if (!use_symmetric(n, m))
    return twopass_impl(xyz1, n, xyz2, m, b, 2, dist1, idx1, dist2, idx2, workspace, workspace_bytes, st);
  const size_t rows_bytes = nnd_rows_workspace_bytes(b, n, m);
  unsigned long long* colkeys = (unsigned long long*)((char*)workspace + rows_bytes);
  // hint_idx2 may be idx2 itself (a loop's own previous result): it is read by the seed pass at the start and idx2 is
  // written by the column finalize at the end, in stream order
  int rc = forward_rows_impl(xyz1, n, xyz2, m, b, 0, n, dist1, idx1, colkeys, workspace, rows_bytes, st, hint_idx2);
  if (rc) return rc;
  return finalize_cols_impl(xyz1, n, xyz2, m, b, colkeys, dist2, idx2, st);
}

int nnd_set_pruned_share(int mode) {
  const int before = g_pr_share.load();
  if (mode >= 0 && mode <= 2) g_pr_share.store(mode);
  return before;
}

size_t nnd_pruned_workspace_bytes(int b, int n, int m) {
  if (b <= 0 || n <= 0 || m <= 0) return 256;
  if ((long long)b * n > 0x7fffffffLL - 1024 || (long long)b * m > 0x7fffffffLL - 1024) return 0;
  return pr_need(b, n, m);
}

int nnd_forward_pruned_cuda_f32(const float* xyz1, const float* xyz2, int b, int n, int m, float* dist1, float* dist2,
                                int32_t* idx1, int32_t* idx2, void* workspace, size_t workspace_bytes, void* stream) {
  g_launches = 0;
  if (!sizes_ok(b, n, m)) return NND_EINVAL;
  if ((long long)b * n > 0x7fffffffLL - 1024 || (long long)b * m > 0x7fffffffLL - 1024) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0 || (n == 0 && m == 0)) return 0;
  if (n == 0) return zero_outputs(dist2, idx2, (long long)b * m, st);  // my_lib.cpp:9-10,21-22 with an empty inner loop
  if (m == 0) return zero_outputs(dist1, idx1, (long long)b * n, st);
  if (!xyz1 || !xyz2 || !dist1 || !dist2 || !idx1 || !idx2 || !workspace) return NND_EINVAL;
  return pruned_forward_impl(xyz1, n, xyz2, m, b, 2, dist1, idx1, dist2, idx2, workspace, workspace_bytes, st);
}

int nnd_forward_pruned_reuse_cuda_f32(const float* xyz1, const float* xyz2, int b, int n, int m, float* dist1, float* dist2,
                                      int32_t* idx1, int32_t* idx2, void* workspace, size_t workspace_bytes, int reuse,
                                      void* stream) {
  g_launches = 0;
  if (reuse < 0 || reuse > 3) return NND_EINVAL;
  if (!sizes_ok(b, n, m) || b <= 0 || n <= 0 || m <= 0) return NND_EINVAL;  // a loop over fixed, non-empty clouds
  if ((long long)b * n > 0x7fffffffLL - 1024 || (long long)b * m > 0x7fffffffLL - 1024) return NND_EINVAL;
  if (!xyz1 || !xyz2 || !dist1 || !dist2 || !idx1 || !idx2 || !workspace) return NND_EINVAL;
  return pruned_forward_impl(xyz1, n, xyz2, m, b, 2, dist1, idx1, dist2, idx2, workspace, workspace_bytes,
                             (cudaStream_t)stream, reuse);
}

int nnd_query_pruned_cuda_f32(const float* query, const float* target, int b, int nq, int nt, float* dist, int32_t* idx,
                              void* workspace, size_t workspace_bytes, void* stream) {
  g_launches = 0;
  if (!sizes_ok(b, nq, nt)) return NND_EINVAL;
  if ((long long)b * nq > 0x7fffffffLL - 1024 || (long long)b * nt > 0x7fffffffLL - 1024) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0 || nq == 0) return 0;
  if (nt == 0) return zero_outputs(dist, idx, (long long)b * nq, st);
  if (!query || !target || !dist || !idx || !workspace) return NND_EINVAL;
  return pruned_forward_impl(query, nq, target, nt, b, 1, dist, idx, nullptr, nullptr, workspace, workspace_bytes, st);
}

int nnd_forward_rows_cuda_f32(const float* xyz1, const float* xyz2, int b, int n, int m, int row_lo, int row_hi,
                              float* dist1, int32_t* idx1, unsigned long long* colkeys, void* workspace,
                              size_t workspace_bytes, void* stream) {
  g_launches = 0;
  if (!sizes_ok(b, n, m) || b == 0 || n == 0 || m == 0) return NND_EINVAL;
  if (row_lo < 0 || row_hi < row_lo || row_hi > n) return NND_EINVAL;
  // whole row groups; an empty range at the very end (a rank left without rows) is fine
  if ((row_lo % kRowAlign != 0 && row_lo != n) || (row_hi % kRowAlign != 0 && row_hi != n)) return NND_EINVAL;
  if (!xyz1 || !xyz2 || !dist1 || !idx1 || !colkeys || !workspace) return NND_EINVAL;
  return forward_rows_impl(xyz1, n, xyz2, m, b, row_lo, row_hi, dist1, idx1, colkeys, workspace, workspace_bytes,
                           (cudaStream_t)stream);
}

int nnd_peer_max_ranks(void) { return kMaxPeers; }

int nnd_exchange_cols_p2p_cuda_f32(const void* const* peer_colkeys, void* const* peer_dist1, void* const* peer_idx1,
                                   void* const* peer_dist2, void* const* peer_idx2, int world, int rank, const float* xyz1,
                                   const float* xyz2, int b, int n, int m, int row_lo, int row_hi, void* stream) {
  g_launches = 0;
  if (!sizes_ok(b, n, m) || b == 0 || n == 0 || m == 0) return NND_EINVAL;
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return NND_EINVAL;
  if (row_lo < 0 || row_hi < row_lo || row_hi > n) return NND_EINVAL;
  if (!peer_colkeys || !peer_dist1 || !peer_idx1 || !peer_dist2 || !peer_idx2 || !xyz1 || !xyz2) return NND_EINVAL;
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  ExchangeParams ep{};
  for (int r = 0; r < world; ++r) {
    if (!peer_colkeys[r] || !peer_dist1[r] || !peer_idx1[r] || !peer_dist2[r] || !peer_idx2[r]) return NND_EINVAL;
    ep.keys[r] = (const unsigned long long*)peer_colkeys[r];
    ep.dist1[r] = (float*)peer_dist1[r];
    ep.idx1[r] = (int32_t*)peer_idx1[r];
    ep.dist2[r] = (float*)peer_dist2[r];
    ep.idx2[r] = (int32_t*)peer_idx2[r];
  }
  ep.fc = FinColParams{nullptr, xyz1, xyz2, nullptr, nullptr, n, m, tiles_of(m), (long long)b * m, 1.0f, m, 0};
  ep.world = world;
  ep.rank = rank;
  const int cs = (m + world - 1) / world;  // columns per rank
  ep.col_lo = rank * cs < m ? rank * cs : m;
  ep.col_hi = ep.col_lo + cs < m ? ep.col_lo + cs : m;
  ep.row_lo = row_lo;
  ep.row_hi = row_hi;
  const long long work = (long long)b * ((ep.col_hi - ep.col_lo) > (row_hi - row_lo) ? (ep.col_hi - ep.col_lo) : (row_hi - row_lo));
  cudaStream_t st = (cudaStream_t)stream;
  if (kShapes[sym_shape(b, n, m, di.sms)].qpt == kQPT)
    nnd_exchange_cols_kernel<kTT, kQPT><<<grid_for(work, di.sms), 256, 0, st>>>(ep);
  else
    nnd_exchange_cols_kernel<kTT, 2><<<grid_for(work, di.sms), 256, 0, st>>>(ep);
  ++g_launches;
  debug_sync("exchange_cols", st);
  return check_last(cudaGetLastError());
}

int nnd_sum_p2p_cuda_f32(void* const* peer_bufs, int world, int rank, long long count, void* stream) {
  g_launches = 0;
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || count < 0 || !peer_bufs) return NND_EINVAL;
  if (count == 0) return 0;
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  SumParams sp{};
  for (int r = 0; r < world; ++r) {
    if (!peer_bufs[r] || ((uintptr_t)peer_bufs[r] & 15u) != 0) return NND_EINVAL;
    sp.buf[r] = (float*)peer_bufs[r];
  }
  sp.world = world;
  sp.rank = rank;
  const long long n4 = count / 4, per = (n4 + world - 1) / world;
  sp.lo4 = rank * per < n4 ? rank * per : n4;
  sp.hi4 = sp.lo4 + per < n4 ? sp.lo4 + per : n4;
  sp.tail_lo = rank == world - 1 ? n4 * 4 : count;
  sp.tail_hi = count;
  const long long work = sp.hi4 - sp.lo4 > 0 ? sp.hi4 - sp.lo4 : 1;
  nnd_sum_peers_kernel<<<grid_for(work, di.sms), 256, 0, (cudaStream_t)stream>>>(sp);
  ++g_launches;
  return check_last(cudaGetLastError());
}

int nnd_finalize_cols_cuda_f32(const float* xyz1, const float* xyz2, int b, int n, int m,
                               const unsigned long long* colkeys, float* dist2, int32_t* idx2, void* stream) {
  g_launches = 0;
  if (!sizes_ok(b, n, m) || b == 0 || n == 0 || m == 0) return NND_EINVAL;
  if (!xyz1 || !xyz2 || !colkeys || !dist2 || !idx2) return NND_EINVAL;
  return finalize_cols_impl(xyz1, n, xyz2, m, b, colkeys, dist2, idx2, (cudaStream_t)stream);
}

int nnd_profile_enable(int on) {
  g_prof.enabled = on != 0;
  g_prof.used = 0;
  if (on && !g_prof.created) {
    for (int i = 0; i < kProfSlots * 4; ++i) {
      cudaError_t e = cudaEventCreate(&g_prof.ev[i]);
      if (e != cudaSuccess) return (int)e;
    }
    g_prof.created = true;
  }
  return 0;
}

int nnd_profile_collect(float* kernel_ms, int* calls) {
  if (!kernel_ms || !calls) return NND_EINVAL;
  kernel_ms[0] = kernel_ms[1] = kernel_ms[2] = 0.f;
  *calls = g_prof.used;
  for (int c = 0; c < g_prof.used; ++c) {
    cudaEvent_t* ev = &g_prof.ev[c * 4];
    cudaError_t e = cudaEventSynchronize(ev[3]);
    if (e != cudaSuccess) return (int)e;
    for (int i = 0; i < 3; ++i) {
      float ms = 0.f;
      e = cudaEventElapsedTime(&ms, ev[i], ev[i + 1]);
      if (e != cudaSuccess) return (int)e;
      kernel_ms[i] += ms;
    }
  }
  g_prof.used = 0;
  return 0;
}

int nnd_query_cuda_f32(const float* query, const float* target, int b, int nq, int nt, float* dist, int32_t* idx,
                       void* workspace, size_t workspace_bytes, void* stream) {
  g_launches = 0;
  if (b < 0 || nq < 0 || nt < 0) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0 || nq == 0) return 0;
  if (nt == 0) return zero_outputs(dist, idx, (long long)b * nq, st);
  if (!query || !target || !dist || !idx) return NND_EINVAL;
  if (use_small(b, nq, nt))
    return small_forward_impl(query, nq, target, nt, b, 1, nullptr, nullptr, dist, idx, nullptr, nullptr, st);
  if (!workspace) return NND_EINVAL;
  return twopass_impl(query, nq, target, nt, b, 1, dist, idx, nullptr, nullptr, workspace, workspace_bytes, st);
}

size_t nnd_query_noalloc_max_bytes(void) { return (size_t)NND_NOALLOC_SCRATCH_BYTES; }

// nnd_query_cuda_f32 with the scratch in g_noalloc_scratch: no cudaMalloc / cudaFree / allocating library call on this
// path (device attribute queries, cudaFuncSetAttribute and kernel launches only).  Nothing in the scratch has to
// survive between calls or start at zero: nnd_prepare_kernel rewrites the SoA copies and the packed keys of every call
// before the scan reads them, so the previous call's contents are harmless.  Calls that overlap on the GPU (two
// streams, or a graph replay beside an eager call) share the array and race.
int nnd_query_noalloc_cuda_f32(const float* query, const float* target, int b, int nq, int nt, float* dist,
                               int32_t* idx, void* stream) {
  g_launches = 0;
  if (b < 0 || nq < 0 || nt < 0) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0 || nq == 0) return 0;
  if (nt == 0) return zero_outputs(dist, idx, (long long)b * nq, st);
  if (!query || !target || !dist || !idx) return NND_EINVAL;
  if (!sizes_ok(b, nq, nt)) return NND_ETOOBIG;
  if (use_small(b, nq, nt))  // single-launch kernel: needs no scratch at all
    return small_forward_impl(query, nq, target, nt, b, 1, nullptr, nullptr, dist, idx, nullptr, nullptr, st);
  const size_t need = soa_bytes(b, nq) + soa_bytes(b, nt) + key_bytes(b, nq);
  if (need > (size_t)NND_NOALLOC_SCRATCH_BYTES) return NND_ETOOBIG;
  void* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc) return rc;
  return twopass_impl(query, nq, target, nt, b, 1, dist, idx, nullptr, nullptr, scratch, need, st);
}

int nnd_backward_term_cuda_f32(const float* query, const float* target, const float* graddist, const int32_t* idx,
                               int b, int nq, int nt, float* grad_query, float* grad_target, void* stream) {
  g_launches = 0;
  if (b < 0 || nq < 0 || nt < 0) return NND_EINVAL;
  if (b == 0 || nq == 0) return 0;
  if (nt == 0) return NND_EINVAL;  // idx would point nowhere
  if (!query || !target || !graddist || !idx || !grad_query || !grad_target) return NND_EINVAL;
  GradParams gp{};
  gp.nterms = 1;
  gp.t[0] = GradTerm{query, target, graddist, idx, grad_query, grad_target, nq, nt, (long long)b * nq};
  return launch_grad(gp, (cudaStream_t)stream);
}

#ifdef NND_PR_STATS
int nnd_debug_pruned_stats(unsigned long long* out4) {  // {tiles scanned, super-tiles passed, tiles passed, warps}; resets
  cudaError_t e = cudaMemcpyFromSymbol(out4, g_pr_stats, 4 * sizeof(unsigned long long));
  if (e != cudaSuccess) return (int)e;
  const unsigned long long z[4] = {0, 0, 0, 0};
  return (int)cudaMemcpyToSymbol(g_pr_stats, z, sizeof(z));
}
#endif

#ifdef NND_SYM_COUNT
int nnd_debug_sym_counters(unsigned long long* out2) {  // {slow-path entries, warp steps}; resets both
  unsigned long long z = 0;
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(&out2[0], g_sym_slow, 8);
  cudaMemcpyFromSymbol(&out2[1], g_sym_steps, 8);
  cudaMemcpyToSymbol(g_sym_slow, &z, 8);
  cudaMemcpyToSymbol(g_sym_steps, &z, 8);
  return 0;
}
#endif

long long nnd_set_panel_min_rows(long long rows) {
  const long long before = panel_min_rows();
  if (rows > 0) g_panel_min_rows.store(rows);
  return before;
}

int nnd_set_panel_tiles(int tiles) {
  const int before = panel_tiles();
  if (tiles == 0 || tiles == 1 || tiles == 2 || tiles == 4) g_panel_tiles.store(tiles);
  return before;
}

size_t nnd_chamfer_reduce_workspace_bytes(void) { return (size_t)kReduceMaxBlocks * sizeof(double) + 256; }

int nnd_chamfer_reduce_cuda_f32(const float* dist1, const float* dist2, int b, int n, int m, float trunc, float scale1,
                                float scale2, float* loss, float* graddist1, float* graddist2, void* workspace,
                                size_t workspace_bytes, void* stream) {
  g_launches = 0;
  if (b < 0 || n < 0 || m < 0 || !loss || !workspace) return NND_EINVAL;
  if (((uintptr_t)workspace & 255u) != 0 || workspace_bytes < nnd_chamfer_reduce_workspace_bytes()) return NND_EWORKSPACE;
  if (((long long)b * n > 0 && !dist1) || ((long long)b * m > 0 && !dist2)) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  ReduceParams rp{};
  rp.dist[0] = dist1;
  rp.dist[1] = dist2;
  rp.graddist[0] = graddist1;
  rp.graddist[1] = graddist2;
  rp.count[0] = (long long)b * n;
  rp.count[1] = (long long)b * m;
  rp.scale[0] = scale1;
  rp.scale[1] = scale2;
  rp.trunc = trunc;
  rp.loss = loss;
  rp.ticket = (unsigned*)workspace;  // first 256 bytes: the ticket
  rp.partial = (double*)((char*)workspace + 256);
  const long long most = rp.count[0] > rp.count[1] ? rp.count[0] : rp.count[1];
  // clouds of up to 8192 points (the NDP loop's real size is ~3 k): ONE block, no ticket, no cross-call state
  long long blocks = most <= 8192 ? 1 : (most + 1023) / 1024;
  if (blocks > kReduceMaxBlocks) blocks = kReduceMaxBlocks;
  if (blocks > 1) {  // the ticket is zeroed here, on the same stream: the workspace needs no preparation by the caller
    cudaError_t e = cudaMemsetAsync(rp.ticket, 0, sizeof(unsigned), st);
    if (e != cudaSuccess) return (int)e;
  }
  nnd_chamfer_reduce_kernel<<<(unsigned)blocks, 256, 0, st>>>(rp);
  ++g_launches;
  return check_last(cudaGetLastError());
}

int nnd_backward_cuda_f32(const float* xyz1, const float* xyz2, const float* graddist1, const float* graddist2,
                          const int32_t* idx1, const int32_t* idx2, int b, int n, int m, float* gradxyz1,
                          float* gradxyz2, void* stream) {
  g_launches = 0;
  if (b < 0 || n < 0 || m < 0) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0) return 0;
  if ((n > 0 && !gradxyz1) || (m > 0 && !gradxyz2)) return NND_EINVAL;
  if (n > 0 && m > 0 && use_small(b, n, m)) {  // one launch: zero-init, both terms and the stores
    if (!xyz1 || !xyz2 || !graddist1 || !graddist2 || !idx1 || !idx2) return NND_EINVAL;
    GradSmallParams sp{{xyz1, xyz2}, {graddist1, graddist2}, {idx1, idx2}, {gradxyz1, gradxyz2}, {nullptr, nullptr},
                       {n, m}, 2, 1};
    return small_backward_impl(sp, b, st);
  }
  cudaError_t e = cudaSuccess;  // zero-init fused into the call (my_lib.cpp:91-94)
  if (n > 0) e = cudaMemsetAsync(gradxyz1, 0, (size_t)b * n * 3 * sizeof(float), st);
  if (m > 0 && e == cudaSuccess) e = cudaMemsetAsync(gradxyz2, 0, (size_t)b * m * 3 * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  if (n == 0 || m == 0) return 0;  // no pairs: both gradients are zero
  if (!xyz1 || !xyz2 || !graddist1 || !graddist2 || !idx1 || !idx2) return NND_EINVAL;
  GradParams gp{};
  gp.nterms = 2;
  gp.t[0] = GradTerm{xyz1, xyz2, graddist1, idx1, gradxyz1, gradxyz2, n, m, (long long)b * n};
  gp.t[1] = GradTerm{xyz2, xyz1, graddist2, idx2, gradxyz2, gradxyz1, m, n, (long long)b * m};
  return launch_grad(gp, st);
}

// ---- ragged batches (per-cloud lengths), single-launch kernels only
int nnd_ragged_supported(int n, int m) { return use_small_nm(n, m) ? 1 : 0; }

int nnd_forward_ragged_cuda_f32(const float* xyz1, const float* xyz2, int b, int n, int m, const long long* len1,
                                const long long* len2, float* dist1, float* dist2, int32_t* idx1, int32_t* idx2,
                                void* stream) {
  g_launches = 0;
  if (!sizes_ok(b, n, m)) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0 || (n == 0 && m == 0)) return 0;
  if (n == 0) return zero_outputs(dist2, idx2, (long long)b * m, st);
  if (m == 0) return zero_outputs(dist1, idx1, (long long)b * n, st);
  if (!use_small_nm(n, m)) return NND_EINVAL;
  if (!xyz1 || !xyz2 || !dist1 || !dist2 || !idx1 || !idx2) return NND_EINVAL;
  return small_forward_impl(xyz1, n, xyz2, m, b, 2, len1, len2, dist1, idx1, dist2, idx2, st);
}

int nnd_query_ragged_cuda_f32(const float* query, const float* target, int b, int nq, int nt, const long long* lenq,
                              const long long* lent, float* dist, int32_t* idx, void* stream) {
  g_launches = 0;
  if (b < 0 || nq < 0 || nt < 0) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0 || nq == 0) return 0;
  if (nt == 0) return zero_outputs(dist, idx, (long long)b * nq, st);
  if (!use_small_nm(nq, nt)) return NND_EINVAL;
  if (!query || !target || !dist || !idx) return NND_EINVAL;
  return small_forward_impl(query, nq, target, nt, b, 1, lenq, lent, dist, idx, nullptr, nullptr, st);
}

int nnd_backward_ragged_cuda_f32(const float* xyz1, const float* xyz2, const float* graddist1, const float* graddist2,
                                 const int32_t* idx1, const int32_t* idx2, int b, int n, int m, const long long* len1,
                                 const long long* len2, float* gradxyz1, float* gradxyz2, void* stream) {
  g_launches = 0;
  if (b < 0 || n <= 0 || m <= 0) return NND_EINVAL;
  if (b == 0) return 0;
  if (!use_small_nm(n, m)) return NND_EINVAL;
  if (!xyz1 || !xyz2 || !graddist1 || !graddist2 || !idx1 || !idx2 || !gradxyz1 || !gradxyz2) return NND_EINVAL;
  GradSmallParams sp{{xyz1, xyz2}, {graddist1, graddist2}, {idx1, idx2}, {gradxyz1, gradxyz2}, {len1, len2}, {n, m}, 2, 1};
  return small_backward_impl(sp, b, (cudaStream_t)stream);
}

int nnd_knn_backward_ragged_cuda_f32(const float* query, const float* target, const float* graddist, const int32_t* idx,
                                     int b, int nq, int nt, const long long* lenq, const long long* lent,
                                     float* grad_query, float* grad_target, void* stream) {
  g_launches = 0;
  if (b < 0 || nq <= 0 || nt <= 0) return NND_EINVAL;
  if (b == 0) return 0;
  if (!use_small_nm(nq, nt)) return NND_EINVAL;
  if (!query || !target || !graddist || !idx || !grad_query || !grad_target) return NND_EINVAL;
  GradSmallParams sp{{query, target}, {graddist, nullptr}, {idx, nullptr}, {grad_query, grad_target}, {lenq, lent},
                     {nq, nt}, 1, 1};
  return small_backward_impl(sp, b, (cudaStream_t)stream);
}

size_t nnd_backward_exact_workspace_bytes(int b, int n, int m) {
  if (b <= 0 || n <= 0 || m <= 0) return 256;
  const long long c1 = (long long)b * n, c2 = (long long)b * m;
  if (c1 > 0x7fffffffLL || c2 > 0x7fffffffLL) return 0;  // rows are 32-bit halves of the sort keys
  const size_t t1 = exact_sort_temp_bytes(c1), t2 = exact_sort_temp_bytes(c2);
  return 2 * key_bytes(b, n) + 2 * key_bytes(b, m) + align_up(t1 > t2 ? t1 : t2, 256) + 256;
}

int nnd_backward_exact_cuda_f32(const float* xyz1, const float* xyz2, const float* graddist1, const float* graddist2,
                                const int32_t* idx1, const int32_t* idx2, int b, int n, int m, float* gradxyz1,
                                float* gradxyz2, void* workspace, size_t workspace_bytes, void* stream) {
  g_launches = 0;
  if (b < 0 || n < 0 || m < 0) return NND_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (b == 0) return 0;
  if ((n > 0 && !gradxyz1) || (m > 0 && !gradxyz2)) return NND_EINVAL;
  if (n == 0 || m == 0) {  // no pairs: both gradients are zero
    cudaError_t e = cudaSuccess;
    if (n > 0) e = cudaMemsetAsync(gradxyz1, 0, (size_t)b * n * 3 * sizeof(float), st);
    if (m > 0 && e == cudaSuccess) e = cudaMemsetAsync(gradxyz2, 0, (size_t)b * m * 3 * sizeof(float), st);
    return check_last(e);
  }
  if (!xyz1 || !xyz2 || !graddist1 || !graddist2 || !idx1 || !idx2 || !workspace) return NND_EINVAL;
  const long long c1 = (long long)b * n, c2 = (long long)b * m;
  if (c1 > 0x7fffffffLL || c2 > 0x7fffffffLL) return NND_EINVAL;
  if (((uintptr_t)workspace & 255u) != 0) return NND_EWORKSPACE;
  const size_t t1 = exact_sort_temp_bytes(c1), t2 = exact_sort_temp_bytes(c2);
  const size_t temp_bytes = align_up(t1 > t2 ? t1 : t2, 256) + 256;
  if (workspace_bytes < 2 * key_bytes(b, n) + 2 * key_bytes(b, m) + temp_bytes) return NND_EWORKSPACE;
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  char* w = (char*)workspace;
  unsigned long long* k1_in = (unsigned long long*)w;   // from idx1: (cloud-2 row << 32 | cloud-1 row), c1 keys
  w += key_bytes(b, n);
  unsigned long long* k1_out = (unsigned long long*)w;
  w += key_bytes(b, n);
  unsigned long long* k2_in = (unsigned long long*)w;   // from idx2: (cloud-1 row << 32 | cloud-2 row), c2 keys
  w += key_bytes(b, m);
  unsigned long long* k2_out = (unsigned long long*)w;
  w += key_bytes(b, m);
  void* temp = (void*)w;

  ExactKeyParams kp{};
  kp.s[0] = ExactKeySide{idx1, k1_in, n, m, c1};
  kp.s[1] = ExactKeySide{idx2, k2_in, m, n, c2};
  const long long most = c1 > c2 ? c1 : c2;
  nnd_exact_keys_kernel<<<grid_for(most, di.sms), 256, 0, st>>>(kp);
  ++g_launches;
  int end1 = 32 + bits_for((unsigned long long)c2), end2 = 32 + bits_for((unsigned long long)c1);
  end1 = end1 > 64 ? 64 : end1;
  end2 = end2 > 64 ? 64 : end2;
  size_t tb = temp_bytes;
  cudaError_t e = cub::DeviceRadixSort::SortKeys(temp, tb, (const unsigned long long*)k1_in, k1_out, (int)c1, 0, end1, st);
  if (e != cudaSuccess) return (int)e;
  tb = temp_bytes;
  e = cub::DeviceRadixSort::SortKeys(temp, tb, (const unsigned long long*)k2_in, k2_out, (int)c2, 0, end2, st);
  if (e != cudaSuccess) return (int)e;

  ExactParams ep{};
  // cloud 1: own term (term 1) first, then the term-2 contributions of the cloud-2 rows that chose it (keys from idx2)
  ep.s[0] = ExactSide{xyz1, xyz2, graddist1, graddist2, idx1, k2_out, gradxyz1, n, m, c1, c2, 1};
  // cloud 2: the term-1 contributions of the cloud-1 rows that chose it (keys from idx1) first, then its own term
  ep.s[1] = ExactSide{xyz2, xyz1, graddist2, graddist1, idx2, k1_out, gradxyz2, m, n, c2, c1, 0};
  nnd_grad_exact_kernel<<<grid_for(most, di.sms), 256, 0, st>>>(ep);
  ++g_launches;
  return check_last(cudaGetLastError());
}

}  // extern "C"
