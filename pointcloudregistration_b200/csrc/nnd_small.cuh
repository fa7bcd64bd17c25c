// nnd_small.cuh -- included by nnd_kernels.cu inside its anonymous namespace.
//
// Small clouds: the whole forward in ONE launch, no workspace; the whole backward in ONE launch, no memsets.
//
// The reference's callers mostly run nnd on small clouds (DIP: B=256 patches of 256 points, dip/train.py:84 with
// dip/preprocess_lrf.py:27-29; NDP: ~3 k points, c2p-net/deformationpyramid/model/registration.py:236), where the
// reference launches 2 forward kernels (nnd_cuda.cu:152-153) and the general path here needs prepare + scan +
// finalize + two memsets + grad: launch latency, not arithmetic.  Here:
//
//  nnd_small_kernel: one CTA = a block of queries of one (direction, batch element) against the WHOLE target cloud,
//    which it first copies AoS -> SoA into shared memory (padded to a multiple of the chunk by repeating the last
//    point: a duplicate at a higher index never wins the tie-break).  The CTA's warps form (query groups) x (target
//    splits): warps of one query group hold the same queries in registers and each scans 1/tsplit of the 16-target
//    chunks with the library's packed sequence (FADD2 x3, FMUL2 x3, FFMA2 x2 with a run-time 1.0: separately rounded,
//    my_lib.cpp:12-15), keeps chunk minima with FMNMX3 and the first chunk that attains the minimum (strict <),
//    resolves the lowest index inside that chunk from shared memory and -- when the targets are split -- merges
//    (dist_bits << 32 | index) into a shared-memory u64 key with atomicMin (lowest index wins ties across warps).
//    The k == 0 rule of my_lib.cpp:16 (a NaN first distance sticks) is applied on the way out.  Both directions are
//    in the same grid.  Optional per-cloud lengths (ragged batches, loss.py:143-144): rows past the length get
//    dist = 0, idx = 0, neighbours are searched among the first len points only.
//
//  nnd_grad_small_kernel: every CTA owns a range of rows of both gradients of one batch element and accumulates
//    them in shared memory (own-row terms by plain stores, which doubles as the zero-init; neighbour-row terms by
//    shared-memory atomics after reading the whole neighbour-index arrays: no global memset, no global atomics) and
//    writes them out once with plain coalesced stores.

constexpr int kSmallChunk = 16;  // targets per lazy-index chunk (the re-scan that resolves the index costs one chunk)

struct SmallParams {
  const float* xyz[2];      // [b][cnt[0]][3], [b][cnt[1]][3]
  float* dist[2];           // [b][cnt[d]]
  int32_t* idx[2];
  const long long* len[2];  // optional per-batch-element valid lengths (null: cnt)
  int cnt[2];
  int qblocks[2];           // CTAs per batch element of direction 0 / 1
  int blocks0;              // b * qblocks[0]: CTAs of direction 0, the rest belong to direction 1
  int tsplit;               // warps sharing one set of queries, each scanning 1/tsplit of the target chunks
  int mp_max;               // padded size of the larger target cloud (component stride in shared memory)
  float one;
};

// lengths come from the caller's device tensor unchecked (no host read): clamp to [0, cnt]
__device__ __forceinline__ int clamp_len(long long len, int cnt) { return len < 0 ? 0 : len > cnt ? cnt : (int)len; }

template <int WARPS, int QPT>
__global__ void __launch_bounds__(WARPS * 32) nnd_small_kernel(const __grid_constant__ SmallParams p) {
  constexpr int THREADS = WARPS * 32, C = kSmallChunk;
  extern __shared__ __align__(16) unsigned char small_smem[];
  const int dir = blockIdx.x >= (unsigned)p.blocks0 ? 1 : 0;
  const int local = (int)blockIdx.x - (dir ? p.blocks0 : 0);
  const int batch = local / p.qblocks[dir], qblock = local - batch * p.qblocks[dir];
  const int cq = p.cnt[dir], ct = p.cnt[1 - dir];
  int nq = cq, nt = ct;
  if (p.len[dir] != nullptr) nq = clamp_len(p.len[dir][batch], cq);
  if (p.len[1 - dir] != nullptr) nt = clamp_len(p.len[1 - dir][batch], ct);
  const int tsplit = p.tsplit;
  const int QC = (WARPS / tsplit) * 32 * QPT;  // queries of this CTA
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float* dist = p.dist[dir] + (long long)batch * cq;
  int32_t* idx = p.idx[dir] + (long long)batch * cq;

  if (nt <= 0 || qblock * QC >= nq) {  // my_lib.cpp:9-10,21-22 with an empty inner loop; rows past the length
    for (int i = tid; i < QC; i += THREADS) {
      const int j = qblock * QC + i;
      if (j < cq) {
        dist[j] = 0.f;
        idx[j] = 0;
      }
    }
    return;
  }

  const int MP = (nt + C - 1) / C * C;
  float* sx = reinterpret_cast<float*>(small_smem);
  float* sy = sx + p.mp_max;
  float* sz = sy + p.mp_max;
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(sz + p.mp_max);  // [QC], used when tsplit > 1
  const float* q = p.xyz[dir] + (long long)batch * cq * 3;
  const float* t = p.xyz[1 - dir] + (long long)batch * ct * 3;

  for (int pt = tid; pt < MP; pt += THREADS) {
    const float* s = t + (pt < nt ? pt : nt - 1) * 3;  // pad by repeating the last point
    sx[pt] = __ldg(s);
    sy[pt] = __ldg(s + 1);
    sz[pt] = __ldg(s + 2);
  }
  if (tsplit > 1)
    for (int i = tid; i < QC; i += THREADS) keys[i] = ~0ull;

  const uint64_t one = pk2(p.one, p.one);
  const float INF = __int_as_float(0x7f800000);
  const int qg = warp / tsplit, tp = warp - qg * tsplit;
  const int NC = MP / C;
  const int c_lo = tp * NC / tsplit, c_hi = (tp + 1) * NC / tsplit;  // this warp's slice of chunks
  const int qbase = qblock * QC + qg * (32 * QPT);
  uint64_t nqx[QPT], nqy[QPT], nqz[QPT];
  float best[QPT];
  int bchunk[QPT];
#pragma unroll
  for (int k = 0; k < QPT; ++k) {
    int j = qbase + k * 32 + lane;
    j = j < nq ? j : nq - 1;
    const float x = -__ldg(q + j * 3), y = -__ldg(q + j * 3 + 1), z = -__ldg(q + j * 3 + 2);
    nqx[k] = pk2(x, x);
    nqy[k] = pk2(y, y);
    nqz[k] = pk2(z, z);
    best[k] = INF;
    bchunk[k] = c_lo * C;
  }
  __syncthreads();

  for (int ch = c_lo; ch < c_hi; ++ch) {
    const ulonglong2* cx = reinterpret_cast<const ulonglong2*>(sx + ch * C);
    const ulonglong2* cy = reinterpret_cast<const ulonglong2*>(sy + ch * C);
    const ulonglong2* cz = reinterpret_cast<const ulonglong2*>(sz + ch * C);
    float cm[QPT];
#pragma unroll
    for (int k = 0; k < QPT; ++k) cm[k] = INF;
#pragma unroll
    for (int s4 = 0; s4 < C / 4; ++s4) {
      const ulonglong2 X = cx[s4], Y = cy[s4], Z = cz[s4];  // broadcast LDS.128: 4 targets
#pragma unroll
      for (int k = 0; k < QPT; ++k) {
        float a0, a1, b0, b1;
        upk2(sqdist2(X.x, Y.x, Z.x, nqx[k], nqy[k], nqz[k], one), a0, a1);
        upk2(sqdist2(X.y, Y.y, Z.y, nqx[k], nqy[k], nqz[k], one), b0, b1);
        cm[k] = fmin3(cm[k], a0, a1);
        cm[k] = fmin3(cm[k], b0, b1);
      }
    }
#pragma unroll
    for (int k = 0; k < QPT; ++k) {  // strict <, chunks ascend: the first chunk that attains the minimum stays
      if (cm[k] < best[k]) bchunk[k] = ch * C;
      best[k] = fminf(best[k], cm[k]);
    }
  }

  // my_lib.cpp:16: k == 0 is accepted unconditionally -> a NaN first distance is the answer
  const uint64_t T0x = pk2(sx[0], sx[0]), T0y = pk2(sy[0], sy[0]), T0z = pk2(sz[0], sz[0]);
  if (c_lo < c_hi) {
#pragma unroll
    for (int k = 0; k < QPT; ++k) {  // index inside the winning chunk: lowest target with d == best (descending scan)
      const ulonglong2* cx = reinterpret_cast<const ulonglong2*>(sx + bchunk[k]);
      const ulonglong2* cy = reinterpret_cast<const ulonglong2*>(sy + bchunk[k]);
      const ulonglong2* cz = reinterpret_cast<const ulonglong2*>(sz + bchunk[k]);
      int found = 0;
#pragma unroll
      for (int s4 = C / 4 - 1; s4 >= 0; --s4) {
        const ulonglong2 X = cx[s4], Y = cy[s4], Z = cz[s4];
        float a0, a1, b0, b1;
        upk2(sqdist2(X.x, Y.x, Z.x, nqx[k], nqy[k], nqz[k], one), a0, a1);
        upk2(sqdist2(X.y, Y.y, Z.y, nqx[k], nqy[k], nqz[k], one), b0, b1);
        if (b1 == best[k]) found = 4 * s4 + 3;
        if (b0 == best[k]) found = 4 * s4 + 2;
        if (a1 == best[k]) found = 4 * s4 + 1;
        if (a0 == best[k]) found = 4 * s4;
      }
      const int hit = bchunk[k] + found;
      if (tsplit == 1) {  // this lane saw every target: final
        const int j = qbase + k * 32 + lane;
        float first, unused;
        upk2(sqdist2(T0x, T0y, T0z, nqx[k], nqy[k], nqz[k], one), first, unused);
        const bool nan0 = first != first;
        if (j < nq) {
          dist[j] = nan0 ? first : best[k];
          idx[j] = nan0 ? 0 : hit;
        } else if (j < cq) {
          dist[j] = 0.f;
          idx[j] = 0;
        }
      } else {
        const unsigned long long key =
            ((unsigned long long)__float_as_uint(best[k]) << 32) | (unsigned long long)(unsigned)hit;
        atomicMin(&keys[qg * (32 * QPT) + k * 32 + lane], key);
      }
    }
  }
  if (tsplit == 1) return;
  __syncthreads();
  for (int i = tid; i < QC; i += THREADS) {
    const int j = qblock * QC + i;
    if (j < nq) {
      const unsigned long long key = keys[i];
      const float x = -__ldg(q + j * 3), y = -__ldg(q + j * 3 + 1), z = -__ldg(q + j * 3 + 2);
      float first, unused;
      upk2(sqdist2(T0x, T0y, T0z, pk2(x, x), pk2(y, y), pk2(z, z), one), first, unused);
      const bool nan0 = first != first;
      dist[j] = nan0 ? first : __uint_as_float((unsigned)(key >> 32));
      idx[j] = nan0 ? 0 : (int32_t)(unsigned)(key & 0xffffffffull);
    } else if (j < cq) {
      dist[j] = 0.f;
      idx[j] = 0;
    }
  }
}

// ----------------------------------------------------------------------------------------------------------------
// Backward for small clouds (my_lib.cpp:91-130 / nnd_cuda.cu:164-183): t_c = fl(fl(2 g) * fl(q_c - target_c)),
// own row += t, neighbour row -= t.
// One batch element is cut into `parts` CTAs; CTA (batch, part) OWNS the rows [c*part/parts, c*(part+1)/parts) of both
// gradients and keeps them in shared memory:
//   phase 0: the own-row terms of its rows, plain stores (every owned entry is written exactly once: this is also the
//            zero-init; rows at or past the cloud's length get 0);
//   phase 1: it reads the WHOLE neighbour-index array of both terms (coalesced, 4 B per row) and, for the rows whose
//            neighbour lies in its range, recomputes the term and subtracts it with a shared-memory atomic
//            (fp32 atomicAdd on shared memory is a compare-and-swap loop on this part: SASS ATOMS.CAST.SPIN);
//   then it writes its rows out with plain coalesced stores.
// No global memset, no global atomics, one launch; `parts` > 1 only spreads a small batch over more SMs.
struct GradSmallParams {
  const float* xyz[2];      // [b][cnt[s]][3]
  const float* gd[2];       // [b][cnt[s]]   upstream gradient of dist_s
  const int32_t* idx[2];    // [b][cnt[s]]   neighbour of row r of cloud s, an index into cloud 1-s
  float* grad[2];           // [b][cnt[s]][3]  OVERWRITTEN
  const long long* len[2];  // optional valid lengths
  int cnt[2];
  int nterms;  // 2: both terms (nnd backward).  1: term 0 only (one-direction knn backward); grad[1] still overwritten
  int parts;   // CTAs per batch element
};

__device__ __forceinline__ void grad_small_own(const float* __restrict__ q, const float* __restrict__ t,
                                               const float* __restrict__ gd, const int32_t* __restrict__ ix, float* acc,
                                               int lo, int hi, int live, bool has_term, int tid, int threads) {
  for (int r = lo + tid; r < hi; r += threads) {
    float tx = 0.f, ty = 0.f, tz = 0.f;
    if (has_term && r < live) {
      const int j = __ldg(ix + r);
      const float g = __fmul_rn(__ldg(gd + r), 2.0f);
      tx = __fmul_rn(g, __fsub_rn(__ldg(q + 3 * r), __ldg(t + 3 * j)));
      ty = __fmul_rn(g, __fsub_rn(__ldg(q + 3 * r + 1), __ldg(t + 3 * j + 1)));
      tz = __fmul_rn(g, __fsub_rn(__ldg(q + 3 * r + 2), __ldg(t + 3 * j + 2)));
    }
    float* a = acc + 3 * (r - lo);
    a[0] = tx;
    a[1] = ty;
    a[2] = tz;
  }
}

// rows r of cloud s (all of them) whose neighbour j = ix[r] lies in [lo, hi) of the other cloud: acc[j - lo] -= t
__device__ __forceinline__ void grad_small_scatter(const float* __restrict__ q, const float* __restrict__ t,
                                                   const float* __restrict__ gd, const int32_t* __restrict__ ix,
                                                   float* acc, int lo, int hi, int live, int tid, int threads) {
#pragma unroll 4
  for (int r = tid; r < live; r += threads) {
    const int j = __ldg(ix + r);
    if (j >= lo && j < hi) {
      const float g = __fmul_rn(__ldg(gd + r), 2.0f);
      const float tx = __fmul_rn(g, __fsub_rn(__ldg(q + 3 * r), __ldg(t + 3 * j)));
      const float ty = __fmul_rn(g, __fsub_rn(__ldg(q + 3 * r + 1), __ldg(t + 3 * j + 1)));
      const float tz = __fmul_rn(g, __fsub_rn(__ldg(q + 3 * r + 2), __ldg(t + 3 * j + 2)));
      float* a = acc + 3 * (j - lo);
      atomicAdd(a, -tx);
      atomicAdd(a + 1, -ty);
      atomicAdd(a + 2, -tz);
    }
  }
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1) nnd_grad_small_kernel(const __grid_constant__ GradSmallParams p) {
  extern __shared__ __align__(16) float gsmall_acc[];
  const int tid = threadIdx.x;
  const int batch = blockIdx.x / p.parts, part = blockIdx.x - batch * p.parts;
  const int c0 = p.cnt[0], c1 = p.cnt[1];
  const int lo0 = (int)((long long)c0 * part / p.parts), hi0 = (int)((long long)c0 * (part + 1) / p.parts);
  const int lo1 = (int)((long long)c1 * part / p.parts), hi1 = (int)((long long)c1 * (part + 1) / p.parts);
  float* acc0 = gsmall_acc;
  float* acc1 = gsmall_acc + 3 * (hi0 - lo0);
  int l0 = c0, l1 = c1;
  if (p.len[0] != nullptr) l0 = clamp_len(p.len[0][batch], c0);
  if (p.len[1] != nullptr) l1 = clamp_len(p.len[1][batch], c1);
  if (l0 <= 0 || l1 <= 0) l0 = l1 = 0;  // no pairs: both gradients are zero
  const float* x0 = p.xyz[0] + (long long)batch * c0 * 3;
  const float* x1 = p.xyz[1] + (long long)batch * c1 * 3;
  const float* gd0 = p.gd[0] + (long long)batch * c0;
  const int32_t* ix0 = p.idx[0] + (long long)batch * c0;
  const bool two = p.nterms == 2;
  const float* gd1 = two ? p.gd[1] + (long long)batch * c1 : nullptr;
  const int32_t* ix1 = two ? p.idx[1] + (long long)batch * c1 : nullptr;
  grad_small_own(x0, x1, gd0, ix0, acc0, lo0, hi0, l0, true, tid, THREADS);
  grad_small_own(x1, x0, gd1, ix1, acc1, lo1, hi1, l1, two, tid, THREADS);
  __syncthreads();
  grad_small_scatter(x0, x1, gd0, ix0, acc1, lo1, hi1, l0, tid, THREADS);          // term 1 lands in cloud 2's rows
  if (two) grad_small_scatter(x1, x0, gd1, ix1, acc0, lo0, hi0, l1, tid, THREADS);  // term 2 lands in cloud 1's rows
  __syncthreads();
  float* g0 = p.grad[0] + ((long long)batch * c0 + lo0) * 3;
  float* g1 = p.grad[1] + ((long long)batch * c1 + lo1) * 3;
  for (int i = tid; i < 3 * (hi0 - lo0); i += THREADS) g0[i] = acc0[i];
  for (int i = tid; i < 3 * (hi1 - lo1); i += THREADS) g1[i] = acc1[i];
}
