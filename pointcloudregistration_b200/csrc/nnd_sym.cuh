// nnd_sym.cuh -- included by nnd_kernels.cu inside its anonymous namespace.
//
// Symmetric forward: BOTH directions of nnd from ONE evaluation of every (xyz1[j], xyz2[k]) distance.
//   dist1 direction computes fl(..(x2-x1)^2..) (my_lib.cpp:12-15 with query = xyz1), dist2 direction computes
//   fl(..(x1-x2)^2..) (my_lib.cpp:57, roles swapped).  x1-x2 == -(x2-x1) exactly in IEEE arithmetic and the
//   squares of a and -a are the same bits, so the two directions see bit-identical distance matrices: the
//   row minima give dist1/idx1 and the column minima give dist2/idx2.
//
// Rows (xyz1 points, in registers) are handled exactly like nnd_scan_kernel: per-chunk minimum, first chunk wins,
// key = (dist_bits << 32 | chunk_base), resolved by nnd_finalize_kernel.  A launch covers the row range
// [row_lo, row_hi) of every batch element (the multi-GPU path gives each rank one range; row_lo % QPT == 0).
// Columns (xyz2 points, streamed through shared memory): every thread owns QPT CONSECUTIVE rows
// j = g*QPT .. g*QPT+QPT-1 ("group" g, numbered globally), so a smaller g means smaller row indices.  Per target the
// thread reduces its QPT distances to one value c (3-input FMNMX tree).  A 32-bit FILTER array cdist[k] (an upper
// bound of the current column minimum's bit pattern) travels with the tile (second TMA copy); only when some lane's
// c is <= the filter value (rare once a target has seen a few thousand rows) the warp takes the slow path:
// CREDUX.MIN over the lanes, lowest lane holding the minimum (= lowest group) issues atomicMin(key = dist_bits << 32
// | g) on the global 64-bit column key and lowers the filter (global and the shared-memory copy).  A stale filter is
// only ever too large, so no improvement or tie is missed.  Column keys of different launches (ranks) merge by a
// plain integer MIN (all keys are < 2^63), and nnd_finalize_cols_kernel resolves the lowest row inside the winning
// group.

struct SymParams {
  const float* qsoa;          // xyz1 as SoA tiles [b][qtiles][3][TT]  (rows)
  const float* tsoa;          // xyz2 as SoA tiles [b][ttiles][3][TT]  (columns)
  unsigned long long* rkeys;  // [b][n]            (dist_bits << 32 | chunk_base)   pre-set to kKeyInit
  unsigned long long* ckeys;  // [b][ttiles * TT]  (dist_bits << 32 | row group)    pre-set to kKeyInit
  unsigned* cdist;            // [b][ttiles * TT]  filter: >= dist_bits of ckeys    pre-set to 0xffffffff
  int n, m;
  int row_lo, row_hi;  // rows of every batch element this launch covers; row_lo % QPT == 0
  int qtiles, ttiles, qblocks;
  long long total_units;  // b * qblocks * ttiles
  float one;
};

struct FinColParams {
  const unsigned long long* ckeys;  // [b][ttiles * TT]
  const float* rows;                // xyz1 [b][n][3]
  const float* cols;                // xyz2 [b][m][3]
  float* dist;                      // dist2 [b][m]
  int32_t* idx;                     // idx2  [b][m]
  int n, m, ttiles;
  long long count;  // b * m
  float one;
  // cols / dist / idx may be a row range of a larger [b][col_stride] array starting at col_offset (role-swapped row
  // ranges: the "columns" are rows [row_lo, row_hi) of xyz1); col_stride == m, col_offset == 0 otherwise.
  long long col_stride;
  int col_offset;
};

#ifdef NND_SYM_COUNT  // dev instrumentation (tools/slow_path_rate.py): slow-path entries / 4-target warp steps
__device__ unsigned long long g_sym_slow, g_sym_steps;
#endif


template <int TT>
struct __align__(128) SymStage {
  float xyz[3 * TT];
  unsigned filt[TT];
};

// Column slow path of nnd_scan_sym_kernel, entered by ALL lanes of a warp when some lane's column minimum passed
// its filter value.  Deliberately not inlined: it runs for well under 1 % of the 4-target steps, and keeping it out
// of line keeps its address arithmetic and registers out of the packed-FP32 loop.
__device__ __forceinline__ void sym_column_update_body(unsigned u0, unsigned u1, unsigned u2, unsigned u3, unsigned h0,
                                                       unsigned h1, unsigned h2, unsigned h3, unsigned lane,
                                                       unsigned group, unsigned long long* ckeys, unsigned* cdist,
                                                       unsigned* filt, int tile_base, int koff) {
  constexpr unsigned FULL = 0xffffffffu;
  const unsigned uu[4] = {u0, u1, u2, u3};
  const unsigned hh[4] = {h0, h1, h2, h3};
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    // The lanes of a warp need not be converged when they load the filter (independent thread scheduling, e.g.
    // after the mbarrier spin), so hh[t] may differ between lanes while another warp lowers the filter.  Every
    // decision that guards a collective must therefore itself come from a collective: the ballot is identical in
    // all lanes.
    if (__ballot_sync(FULL, uu[t] <= hh[t]) != 0u) {
      // The lowest value over ALL lanes is a true distance of its group, so posting it is always valid, and it
      // is <= the value of every lane that passed its filter: nothing is missed.
      const unsigned mn = __reduce_min_sync(FULL, uu[t]);
      const unsigned holders = __ballot_sync(FULL, uu[t] == mn);
      if (lane == (unsigned)(__ffs(holders) - 1)) {  // lowest lane = lowest row group
        const unsigned long long key = ((unsigned long long)mn << 32) | group;
        unsigned long long* kp = ckeys + (long long)tile_base + koff + t;
        // An exact tie with the filter can only matter if this group is lower than the recorded one: look before
        // posting, so that degenerate clouds (many identical points) do not turn every step into three atomics.
        if (mn < hh[t] || key < *reinterpret_cast<volatile unsigned long long*>(kp)) {
          atomicMin(kp, key);
          atomicMin(cdist + (long long)tile_base + koff + t, mn);
          atomicMin(filt + koff + t, mn);
        }
      }
    }
  }
}

__device__ __noinline__ void sym_column_update(unsigned u0, unsigned u1, unsigned u2, unsigned u3, unsigned h0,
                                               unsigned h1, unsigned h2, unsigned h3, unsigned lane, unsigned group,
                                               unsigned long long* ckeys, unsigned* cdist, unsigned* filt,
                                               int tile_base, int koff) {
  sym_column_update_body(u0, u1, u2, u3, h0, h1, h2, h3, lane, group, ckeys, cdist, filt, tile_base, koff);
}

template <int THREADS, int QPT, int TT, int C, int STAGES, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) nnd_scan_sym_kernel(const __grid_constant__ SymParams p) {
  static_assert(TT % C == 0 && C % 4 == 0 && QPT % 2 == 0 && TT % QPT == 0, "tile/chunk shape");
  constexpr int QB = THREADS * QPT;
  constexpr uint32_t XYZ_BYTES = 3 * TT * sizeof(float);
  constexpr uint32_t FILT_BYTES = TT * sizeof(unsigned);
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ __align__(128) unsigned char sym_smem[];
  SymStage<TT>* stages = reinterpret_cast<SymStage<TT>*>(sym_smem);
  // Row bookkeeping lives in shared memory and is touched once per C-target chunk: previous best and the base of the
  // chunk that produced it, per (row slot, thread).  Keeps 2*QPT registers free for the packed stream
  // (272.4 -> 270.4 ms at 1M; the run-time 1.0 stays a kernel parameter = uniform register: as a vector register it
  // costs +3 %, profiles/r01_ubench_step.txt).
  float* sprev = reinterpret_cast<float*>(sym_smem + STAGES * sizeof(SymStage<TT>)) + threadIdx.x;
  int* sbase = reinterpret_cast<int*>(sprev - threadIdx.x + QPT * THREADS) + threadIdx.x;
  __shared__ __align__(8) uint64_t full[STAGES];

  const int tid = threadIdx.x;
  const unsigned lane = tid & 31u;
  const long long u_begin = p.total_units * blockIdx.x / gridDim.x;
  const long long u_end = p.total_units * (blockIdx.x + 1) / gridDim.x;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  long long pu = u_begin;  // producer cursor (thread 0)
  auto produce = [&](int stage) {
    const long long qb = pu / p.ttiles;
    const int tile = (int)(pu - qb * p.ttiles);
    const long long batch = qb / p.qblocks;
    const long long t = batch * p.ttiles + tile;
    mbar_expect_tx(&full[stage], XYZ_BYTES + FILT_BYTES);
    tma_load_1d(stages[stage].xyz, p.tsoa + t * (3 * TT), XYZ_BYTES, &full[stage]);
    tma_load_1d(stages[stage].filt, p.cdist + t * TT, FILT_BYTES, &full[stage]);
    ++pu;
  };
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s)
      if (pu < u_end) produce(s);
  }

  const uint64_t one = pk2(p.one, p.one);
  const float INF = __int_as_float(0x7f800000);
  uint32_t cnt = 0;

  long long u = u_begin;
  while (u < u_end) {
    const long long qbg = u / p.ttiles;
    const int tile0 = (int)(u - qbg * p.ttiles);
    const int batch = (int)(qbg / p.qblocks);
    const int qblock = (int)(qbg - (long long)batch * p.qblocks);
    const long long rest = (long long)p.ttiles - tile0;
    const int seg_tiles = (int)(u_end - u < rest ? u_end - u : rest);

    // ---- rows -> registers: QPT consecutive rows per thread
    uint64_t nqx[QPT], nqy[QPT], nqz[QPT];
    float cm[QPT];  // running row minima of this segment
    const unsigned group = (unsigned)(p.row_lo / QPT + qblock * THREADS + tid);  // global row group id
    const float* qbase = p.qsoa + (long long)batch * p.qtiles * (3 * TT);
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
      int j = p.row_lo + qblock * QB + tid * QPT + q;
      j = j < p.row_hi ? j : p.row_hi - 1;  // a repeated row sits in a higher group: it never wins a tie
      const int qt = j / TT, qo = j - qt * TT;
      const float* s = qbase + (long long)qt * (3 * TT) + qo;
      const float x = -s[0], y = -s[TT], z = -s[2 * TT];
      nqx[q] = pk2(x, x);
      nqy[q] = pk2(y, y);
      nqz[q] = pk2(z, z);
      cm[q] = INF;
      sprev[q * THREADS] = INF;
      sbase[q * THREADS] = 0;
    }
    unsigned long long* ckeys = p.ckeys + (long long)batch * p.ttiles * TT;
    unsigned* cdist = p.cdist + (long long)batch * p.ttiles * TT;

    // A warp whose 32*QPT rows all lie beyond row_hi (ragged last query block of a row range) only keeps the
    // barriers company: its rows are repeats that can never win.  Warp 0 holds the producer and always has rows.
    const bool warp_live = p.row_lo + qblock * QB + (tid & ~31) * QPT < p.row_hi;

    for (int tile = tile0; tile < tile0 + seg_tiles; ++tile) {
      const int stage = cnt % STAGES;
      const int nchunks_live = warp_live ? (min(TT, p.m - tile * TT) + C - 1) / C : 0;
      if (warp_live) mbar_wait(&full[stage], (cnt / STAGES) & 1);
      const float* sx = stages[stage].xyz;
      const float* sy = sx + TT;
      const float* sz = sx + 2 * TT;
      unsigned* filt = stages[stage].filt;
      const int nchunks = nchunks_live;

      for (int ch = 0; ch < nchunks; ++ch) {
        const ulonglong2* cx = reinterpret_cast<const ulonglong2*>(sx + ch * C);
        const ulonglong2* cy = reinterpret_cast<const ulonglong2*>(sy + ch * C);
        const ulonglong2* cz = reinterpret_cast<const ulonglong2*>(sz + ch * C);
        // one 4-target step: 32 distances, row minima into cm[], column minima of this thread's rows out as bit patterns
        auto step = [&](int s4, unsigned& u0, unsigned& u1, unsigned& u2, unsigned& u3) {
          const ulonglong2 X = cx[s4], Y = cy[s4], Z = cz[s4];
          float c0 = INF, c1 = INF, c2 = INF, c3 = INF;  // column minima over this thread's rows
#pragma unroll
          for (int q = 0; q < QPT; q += 2) {
            float a0, a1, b0, b1, e0, e1, f0, f1;
            upk2(sqdist2(X.x, Y.x, Z.x, nqx[q], nqy[q], nqz[q], one), a0, a1);
            upk2(sqdist2(X.y, Y.y, Z.y, nqx[q], nqy[q], nqz[q], one), b0, b1);
            upk2(sqdist2(X.x, Y.x, Z.x, nqx[q + 1], nqy[q + 1], nqz[q + 1], one), e0, e1);
            upk2(sqdist2(X.y, Y.y, Z.y, nqx[q + 1], nqy[q + 1], nqz[q + 1], one), f0, f1);
            cm[q] = fmin3(cm[q], a0, a1);
            cm[q] = fmin3(cm[q], b0, b1);
            cm[q + 1] = fmin3(cm[q + 1], e0, e1);
            cm[q + 1] = fmin3(cm[q + 1], f0, f1);
            c0 = fmin3(c0, a0, e0);
            c1 = fmin3(c1, a1, e1);
            c2 = fmin3(c2, b0, f0);
            c3 = fmin3(c3, b1, f1);
          }
          u0 = __float_as_uint(c0);
          u1 = __float_as_uint(c1);
          u2 = __float_as_uint(c2);
          u3 = __float_as_uint(c3);
        };
        // (One filter test per TWO steps, which gains 0.6 % in the panel kernel, LOSES here: C2 2.68 -> 2.78 ms, C3 2.87 ->
        // 2.91 ms, profiles/r02_ab_sym_pair.txt -- a quarter of the steps take the slow path at C2, and it is called twice.)
#pragma unroll
        for (int s4 = 0; s4 < C / 4; ++s4) {
          unsigned u0, u1, u2, u3;
          step(s4, u0, u1, u2, u3);
          // ---- columns: compare with the filter (an upper bound of the column minimum's bit pattern)
          const int koff = ch * C + 4 * s4;
          const uint4 H = *reinterpret_cast<const uint4*>(filt + koff);
          const unsigned h0 = H.x, h1 = H.y, h2 = H.z, h3 = H.w;
          // (The slow path inline instead of the out-of-line call LOSES at every size -- behind the common test C2 2.68 ->
          // 2.79 ms, C3 2.87 -> 3.00 ms; as four per-target tests without the common one C2 3.00 ms, C3 3.33 ms:
          // profiles/r02_ab_sym_inline.txt.)
          const bool ev = (u0 <= h0) | (u1 <= h1) | (u2 <= h2) | (u3 <= h3);
          if (__any_sync(FULL, ev)) {
#ifdef NND_SYM_COUNT
            if (lane == 0) atomicAdd(&g_sym_slow, 1ull);
#endif
            sym_column_update(u0, u1, u2, u3, h0, h1, h2, h3, lane, group, ckeys, cdist, filt, tile * TT, koff);
          }
        }
#ifdef NND_SYM_COUNT
        if (lane == 0) atomicAdd(&g_sym_steps, (unsigned long long)(C / 4));
#endif
        // end of the chunk: did it improve a row?  (strict <, chunks ascend: the earliest chunk with the minimum stays)
        const int cbase = tile * TT + ch * C;
#pragma unroll
        for (int q = 0; q < QPT; ++q) {
          if (cm[q] < sprev[q * THREADS]) {
            sprev[q * THREADS] = cm[q];
            sbase[q * THREADS] = cbase;
          }
        }
      }
      __syncthreads();
      if (tid == 0 && pu < u_end) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        produce(stage);
      }
      ++cnt;
    }

    unsigned long long* rkeys = p.rkeys + (long long)batch * p.n;
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
      const int j = p.row_lo + qblock * QB + tid * QPT + q;
      if (j < p.row_hi) {
        const unsigned long long key =
            ((unsigned long long)__float_as_uint(cm[q]) << 32) | (unsigned long long)(unsigned)sbase[q * THREADS];
        atomicMin(rkeys + j, key);
      }
    }
    u += seg_tiles;
  }
}

// Column finalize: key = (dist_bits << 32 | row group).  Re-evaluate the QPT rows of the group against this
// column point with the same packed sequence and take the lowest row with d == dist.
template <int QPT>
__device__ __forceinline__ void sym_finalize_col(const FinColParams& p, long long batch, int k, unsigned long long key,
                                                 uint64_t one, float& dist, int32_t& idx) {
  const float bq = __uint_as_float((unsigned)(key >> 32));
  const long long j0 = (long long)(unsigned)(key & 0xffffffffull) * QPT;
  const long long go = batch * p.col_stride + p.col_offset + k;  // position in cols
  const float* t = p.cols + go * 3;
  const uint64_t TX = pk2(t[0], t[0]), TY = pk2(t[1], t[1]), TZ = pk2(t[2], t[2]);
  const float* rows = p.rows + batch * p.n * 3;
  int found = 0;
#pragma unroll
  for (int q = QPT - 2; q >= 0; q -= 2) {  // descending: the last hit is the lowest row
    long long ja = j0 + q, jb = j0 + q + 1;
    ja = ja < p.n ? ja : p.n - 1;
    jb = jb < p.n ? jb : p.n - 1;
    const float* ra = rows + ja * 3;
    const float* rb = rows + jb * 3;
    const uint64_t dd = sqdist2(TX, TY, TZ, pk2(-ra[0], -rb[0]), pk2(-ra[1], -rb[1]), pk2(-ra[2], -rb[2]), one);
    float lo, hi;
    upk2(dd, lo, hi);
    if (hi == bq) found = q + 1;
    if (lo == bq) found = q;
  }
  long long j = j0 + found;
  j = j < p.n ? j : p.n - 1;
  // my_lib.cpp:16: the first candidate (row 0) is accepted unconditionally, so a NaN first distance is the answer
  float first, unused;
  upk2(sqdist2(TX, TY, TZ, pk2(-rows[0], -rows[0]), pk2(-rows[1], -rows[1]), pk2(-rows[2], -rows[2]), one), first, unused);
  const bool nan0 = first != first;
  dist = nan0 ? first : bq;
  idx = nan0 ? 0 : (int32_t)j;
}

template <int TT, int QPT>
__global__ void __launch_bounds__(256) nnd_finalize_cols_kernel(FinColParams p) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const uint64_t one = pk2(p.one, p.one);
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < p.count; g += stride) {
    const long long batch = g / p.m;
    const int k = (int)(g - batch * p.m);
    const unsigned long long key = p.ckeys[batch * p.ttiles * TT + k];
    const long long go = batch * p.col_stride + p.col_offset + k;  // position in dist / idx
    sym_finalize_col<QPT>(p, batch, k, key, one, p.dist[go], p.idx[go]);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Multi-GPU exchange over PEER MEMORY (NVLink / NVSwitch P2P, buffers from torch's symmetric-memory allocator):
// the merge of the per-rank column keys, the column finalize and the gather of all results in ONE kernel.
// Rank r owns a slice of the columns: for each of them it loads the partial key of EVERY rank straight from that
// rank's memory, takes the integer minimum (order-independent: bit-identical for any world size), resolves it to
// (dist2, idx2) with the finalize above and stores the result into every rank's output.  The same launch pushes the
// rank's own rows of dist1 / idx1 to every rank.  Replaces all_reduce(MIN) + replicated finalize + all_gather.
// The caller brackets the launch with two barriers over all ranks (before: every rank's keys are complete; after:
// every store has landed).
constexpr int kMaxPeers = 16;
struct ExchangeParams {
  const unsigned long long* keys[kMaxPeers];  // [b][ttiles * TT] per rank
  float* dist2[kMaxPeers];                    // [b][m]
  int32_t* idx2[kMaxPeers];
  float* dist1[kMaxPeers];                    // [b][n]
  int32_t* idx1[kMaxPeers];
  FinColParams fc;                            // rows = xyz1, cols = xyz2 (local copies); ckeys / dist / idx unused
  int world, rank;
  int col_lo, col_hi;                         // this rank's slice of the columns
  int row_lo, row_hi;                         // this rank's rows of xyz1 (already final in ITS dist1 / idx1)
};

template <int TT, int QPT>
__global__ void __launch_bounds__(256) nnd_exchange_cols_kernel(const __grid_constant__ ExchangeParams p) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t one = pk2(p.fc.one, p.fc.one);
  const int b = (int)(p.fc.count / p.fc.m);
  const int ncols = p.col_hi - p.col_lo;
  for (long long g = first; g < (long long)b * ncols; g += stride) {
    const long long batch = g / ncols;
    const int k = p.col_lo + (int)(g - batch * ncols);
    const long long ko = batch * p.fc.ttiles * TT + k;
    // all peer loads in flight together (each is a ~2 us round trip over NVLink): predicated, fully unrolled
    unsigned long long kk[kMaxPeers];
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r)
      kk[r] = r < p.world ? __ldcv(p.keys[r] + ko) : 0xffffffffffffffffull;
    unsigned long long key = kk[0];
#pragma unroll
    for (int r = 1; r < kMaxPeers; ++r) key = kk[r] < key ? kk[r] : key;
    float d;
    int32_t i;
    sym_finalize_col<QPT>(p.fc, batch, k, key, one, d, i);
    const long long go = batch * p.fc.m + k;
    for (int r = 0; r < p.world; ++r) {
      p.dist2[r][go] = d;
      p.idx2[r][go] = i;
    }
  }
  const int nrows = p.row_hi - p.row_lo;
  const float* own_d = p.dist1[p.rank];
  const int32_t* own_i = p.idx1[p.rank];
  for (long long g = first; g < (long long)b * nrows; g += stride) {
    const long long batch = g / nrows;
    const long long go = batch * p.fc.n + p.row_lo + (g - batch * nrows);
    const float d = own_d[go];
    const int32_t i = own_i[go];
    for (int r = 0; r < p.world; ++r)
      if (r != p.rank) {
        p.dist1[r][go] = d;
        p.idx1[r][go] = i;
      }
  }
}

// Sum of the per-rank gradient buffers over peer memory, two-shot: rank r adds up its slice of EVERY rank's buffer in
// rank order (the same bits on every rank, run to run) and stores the sum into every rank's buffer.  Slice r is read and
// written by rank r only, so the exchange is in place.  Bracketed by two barriers like the column exchange.
struct SumParams {
  float* buf[kMaxPeers];
  int world, rank;
  long long lo4, hi4;  // this rank's slice in float4 units
  long long tail_lo, tail_hi;  // scalar leftovers (the last rank takes them)
};
__global__ void __launch_bounds__(256) nnd_sum_peers_kernel(const __grid_constant__ SumParams p) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (long long i = p.lo4 + first; i < p.hi4; i += stride) {
    float4 v[kMaxPeers];  // all peer loads in flight together; summed in rank order
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r)
      v[r] = r < p.world ? __ldcv(reinterpret_cast<const float4*>(p.buf[r]) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 acc = v[0];
#pragma unroll
    for (int r = 1; r < kMaxPeers; ++r)
      if (r < p.world) {
        acc.x += v[r].x;
        acc.y += v[r].y;
        acc.z += v[r].z;
        acc.w += v[r].w;
      }
    for (int r = 0; r < p.world; ++r) reinterpret_cast<float4*>(p.buf[r])[i] = acc;
  }
  for (long long i = p.tail_lo + first; i < p.tail_hi; i += stride) {
    float acc = p.buf[0][i];
    for (int r = 1; r < p.world; ++r) acc += p.buf[r][i];
    for (int r = 0; r < p.world; ++r) p.buf[r][i] = acc;
  }
}
