// nnd_pruned.cuh -- included by nnd_kernels.cu inside its anonymous namespace, after nnd_hint.cuh (f2ord / ord2f).
//
// EXACT nearest neighbours with spatial pruning (opt-in: nnd_forward_pruned_cuda_f32).  New design: the reference only
// has the brute-force scan (my_lib.cpp:3-25, nnd_cuda.cu:8-130), and so does the default path of this library -- the
// metric of this project is defined on it.  This path returns THE SAME BITS (dist bit-exact, lowest index on ties, the
// same NaN / inf rules) while evaluating only the pairs that can matter:
//
//   * both clouds are sorted along a Morton curve (<= 30-bit codes over the joint bounding box; cub radix sort of
//     (batch : code : index) keys) and cut into TILES of kPrTile consecutive points with their bounding boxes,
//     SUPER-TILES of 32 tiles and LARGE TILES of 32 super-tiles;
//   * a GROUP of 32 consecutive sorted queries (a compact box) belongs to one warp -- or, for small problems and large
//     target clouds, to the 4 warps of a CTA, each taking a quarter of the target tiles.  A warp first scans the target
//     tile at the group's own position on the curve (good upper bounds), then walks large tiles -> super-tiles ->
//     tiles, skipping every box whose distance lower bound exceeds the largest current best distance of the group's
//     queries, and scans a surviving tile only if some query's own point-to-box bound does not exceed that query's best;
//   * a skipped box satisfies  fl-distance(q, t) >= bound > best(q)  for every target t in it and every query q of the
//     warp, so it holds neither the minimum nor a tie; the bound is the exact box distance evaluated in fp32 and then
//     shrunk by 2^-20 relative and 2^-100 absolute, which covers the rounding of both sides;
//   * inside a tile every distance is evaluated with the reference's separately rounded sequence
//     (target - query, (xx + yy) + zz) and candidates compete on (distance, ORIGINAL index), so the visiting order does
//     not matter: the result is the lowest original index among the minima, exactly what the ascending scan gives;
//   * my_lib.cpp:16 accepts k = 0 unconditionally: a NaN first distance sticks, and with no finite distance at all the
//     answer is (distance to point 0, 0).  The epilogue re-evaluates the distance to ORIGINAL target 0 and applies that.
//
// Non-finite coordinates are safe: boxes are built with NaN-dropping min / max, every prune test is written so that a
// NaN bound does not prune, NaN distances never win a comparison.

#ifndef NND_PR_TILE
#define NND_PR_TILE 64
#endif
constexpr int kPrTile = NND_PR_TILE;  // points per tile
constexpr int kPrSuper = 32;   // tiles per super-tile (one lane per tile)
constexpr int kPrWarps = 4;    // warps (of 32 queries) per CTA of the search kernel
// A warp's 32 queries sit kPrQPL to a lane in 32 / kPrQPL lanes; the 32 / (32 / kPrQPL) = kPrQPL lane GROUPS each hold a copy
// of the queries and take every kPrQPL-th target pair of a tile.  One broadcast LDG.128 then feeds kPrQPL evaluations per
// lane instead of one: with one query per lane the scan was bound by the register-file return bandwidth of the loads
// (512 bytes per warp load), 3.4 ms instead of 2.0 ms for the two searches at C4 (profiles/r02_pruned_time_v*.txt).
constexpr int kPrQPL = 4;
static_assert(kPrQPL == 4, "the lane mapping below is written for 4 queries per lane");
static_assert((kPrTile / 2) % kPrQPL == 0 && kPrTile % 32 == 0, "tile size");

struct __align__(32) PrBox {  // lo.xyz, hi.xyz (+ padding to 32 bytes: two 16-byte loads)
  float lx, ly, lz, hx, hy, hz, pad0, pad1;
};

struct PrCloud {
  const float* xyz;          // [b][n][3] original order
  int n;                     // points per batch element
  int tiles, stiles, ltiles; // tiles / super-tiles (32 tiles) / large tiles (32 super-tiles) per batch element
  unsigned long long* keys;  // [b * n] sorted keys: (batch : morton) << 32 | original index
  float4* pts;               // [b][tiles * kPrTile]  sorted points in PAIR layout: points 2p, 2p+1 are the two float4
                             //   {x0, x1, y0, y1} {z0, z1, w0, w1}, w = original index (int bits) -- operands of the
                             //   packed f32x2 sequence; padded with copies of the last point carrying index INT_MAX
  PrBox* tbox;               // [b][tiles]
  PrBox* sbox;               // [b][stiles]
  PrBox* lbox;               // [b][ltiles]   (65 536 points each: keeps the walk O(N/65536 + log) per warp at 10^7..10^8 points)
  unsigned* tcode;           // [b][tiles]  high key word of the first point of the tile
};

__device__ __forceinline__ unsigned pr_expand10(unsigned v) {  // 10 bits -> every third bit
  v &= 0x3ffu;
  v = (v | (v << 16)) & 0x030000ffu;
  v = (v | (v << 8)) & 0x0300f00fu;
  v = (v | (v << 4)) & 0x030c30c3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

// blo[b][3] / bhi[b][3]: ordered-uint min / max x,y,z over the finite coordinates (atomics: called once per cloud, so the
// box is the JOINT box of both clouds and their codes are comparable); pre-set to 0xffffffff / 0 by the launcher
__global__ void __launch_bounds__(256) nnd_pr_bbox_kernel(const float* __restrict__ xyz, int n, unsigned* blo, unsigned* bhi) {
  const int batch = blockIdx.y;
  const float* r = xyz + (long long)batch * n * 3;
  unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = __ldg(r + 3 * (long long)j + c);
      if (v == v && fabsf(v) < 3e38f) {
        const unsigned o = f2ord(v);
        lo[c] = min(lo[c], o);
        hi[c] = max(hi[c], o);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    lo[c] = __reduce_min_sync(0xffffffffu, lo[c]);
    hi[c] = __reduce_max_sync(0xffffffffu, hi[c]);
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      atomicMin(blo + batch * 3 + c, lo[c]);
      atomicMax(bhi + batch * 3 + c, hi[c]);
    }
  }
}

// keys[batch * n + j] = ((batch << 3*kb) | morton(kb bits per axis)) << 32 | j
__global__ void __launch_bounds__(256) nnd_pr_code_kernel(const float* __restrict__ xyz, int n, const unsigned* __restrict__ blo,
                                                          const unsigned* __restrict__ bhi, int kb,
                                                          unsigned long long* __restrict__ keys) {
  const int batch = blockIdx.y;
  const float* r = xyz + (long long)batch * n * 3;
  float lo[3], inv[3];
  const float cells = (float)(1 << kb);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const unsigned ulo = blo[batch * 3 + c], uhi = bhi[batch * 3 + c];
    const float l = ulo <= uhi ? ord2f(ulo) : 0.f, h = ulo <= uhi ? ord2f(uhi) : 0.f;
    const float ext = h - l;
    lo[c] = l;
    inv[c] = ext > 0.f && ext < 3e38f ? cells / ext : 0.f;
  }
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    unsigned code = 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = __ldg(r + 3 * (long long)j + c);
      // fmaxf / fminf drop NaNs; +-inf clamp to the border cells
      const unsigned cell = (unsigned)fminf(fmaxf((v - lo[c]) * inv[c], 0.f), cells - 1.f);
      code |= pr_expand10(cell) << c;
    }
    const unsigned hiw = ((unsigned)batch << (3 * kb)) | code;
    keys[(long long)batch * n + j] = ((unsigned long long)hiw << 32) | (unsigned)j;
  }
}

__device__ __forceinline__ float pr_wmin(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float pr_wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// One CTA of kPrTile threads per tile: sorted points (float4, w = original index) + the tile's box and first code.
__global__ void __launch_bounds__(kPrTile) nnd_pr_gather_kernel(PrCloud c) {
  static_assert(kPrTile % 32 == 0 && kPrTile <= 1024, "tile size");
  const int batch = blockIdx.y, tile = blockIdx.x;
  const int s = tile * kPrTile + threadIdx.x;
  const int sc = s < c.n ? s : c.n - 1;
  const unsigned long long key = c.keys[(long long)batch * c.n + sc];
  const unsigned j = (unsigned)(key & 0xffffffffull);
  const float* p = c.xyz + ((long long)batch * c.n + j) * 3;
  const float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2);
  {
    float* o = reinterpret_cast<float*>(c.pts + ((long long)batch * c.tiles + tile) * kPrTile + (threadIdx.x & ~1u)) +
               (threadIdx.x & 1u);
    o[0] = x;
    o[2] = y;
    o[4] = z;
    o[6] = __int_as_float(s < c.n ? (int)j : 0x7fffffff);
  }
  __shared__ float red[6][kPrTile / 32];
  const float INF = __int_as_float(0x7f800000);
  float v[6] = {x, y, z, x, y, z};
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float w = v[k] == v[k] ? v[k] : (k < 3 ? INF : -INF);  // NaN: neutral element
    w = k < 3 ? pr_wmin(w) : pr_wmax(w);
    if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = w;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float o[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      o[k] = red[k][0];
#pragma unroll
      for (int w = 1; w < kPrTile / 32; ++w) o[k] = k < 3 ? fminf(o[k], red[k][w]) : fmaxf(o[k], red[k][w]);
    }
    c.tbox[(long long)batch * c.tiles + tile] = PrBox{o[0], o[1], o[2], o[3], o[4], o[5], 0.f, 0.f};
    c.tcode[(long long)batch * c.tiles + tile] = (unsigned)(c.keys[(long long)batch * c.n + tile * kPrTile] >> 32);
  }
}

// One warp per output box: union of (up to) 32 consecutive input boxes (tiles -> super-tiles -> large tiles).
__global__ void __launch_bounds__(128) nnd_pr_super_kernel(const PrBox* __restrict__ in, int nin, PrBox* __restrict__ out,
                                                           int nout) {
  const int batch = blockIdx.y;
  const int o = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (o >= nout) return;
  const int lane = threadIdx.x & 31;
  const int i = o * kPrSuper + lane;
  const float INF = __int_as_float(0x7f800000);
  PrBox bx{INF, INF, INF, -INF, -INF, -INF, 0.f, 0.f};
  if (i < nin) bx = in[(long long)batch * nin + i];
  bx.lx = pr_wmin(bx.lx);
  bx.ly = pr_wmin(bx.ly);
  bx.lz = pr_wmin(bx.lz);
  bx.hx = pr_wmax(bx.hx);
  bx.hy = pr_wmax(bx.hy);
  bx.hz = pr_wmax(bx.hz);
  if (lane == 0) out[(long long)batch * nout + o] = bx;
}

// Lower bound of the fl-evaluated squared distance between any point of box a and any point of box b, made safe
// against rounding (header comment).  NaN-tolerant: a NaN gap is dropped by fmaxf, which only lowers the bound.
__device__ __forceinline__ float pr_safe(float lb) { return lb * (1.0f - 9.5367431640625e-7f) - 7.8886090522101181e-31f; }
__device__ __forceinline__ float pr_box_box(const PrBox& a, const PrBox& b) {
  const float gx = fmaxf(fmaxf(a.lx - b.hx, b.lx - a.hx), 0.f);
  const float gy = fmaxf(fmaxf(a.ly - b.hy, b.ly - a.hy), 0.f);
  const float gz = fmaxf(fmaxf(a.lz - b.hz, b.lz - a.hz), 0.f);
  return pr_safe(gx * gx + gy * gy + gz * gz);
}
__device__ __forceinline__ float pr_point_box(float x, float y, float z, const PrBox& b) {
  const float gx = fmaxf(fmaxf(b.lx - x, x - b.hx), 0.f);
  const float gy = fmaxf(fmaxf(b.ly - y, y - b.hy), 0.f);
  const float gz = fmaxf(fmaxf(b.lz - z, z - b.hz), 0.f);
  return pr_safe(gx * gx + gy * gy + gz * gz);
}

#ifdef NND_PR_STATS  // dev instrumentation (tools/pruned_stats.py): {tiles scanned, super-tiles passed, tiles passed, warps}
__device__ unsigned long long g_pr_stats[4];
#endif

struct PrSearch {
  PrCloud q, t;    // queries, targets
  float* dist;     // [b][q.n]  original query order
  int32_t* idx;    // [b][q.n]
  float one;       // 1.0f at run time (sqdist2's additions are x * 1.0 + y, like every scan of this library)
};

// This lane group's share of one tile (every kPrQPL-th pair of targets) against the lane's kPrQPL queries, two targets at
// a time with the library's packed sequence (sqdist2: target - query, (xx + yy) + zz, every operation rounded
// separately).  Candidates compete on the 64-bit key (distance bits << 32 | original index): for the non-negative
// distances of this op unsigned key order is (distance, index) order, NaN patterns sort above +inf and never win.
__device__ __forceinline__ unsigned long long pr_key(float d, float w) {
  unsigned long long k;
  asm("mov.b64 %0, {%1, %2};" : "=l"(k) : "f"(w), "f"(d));
  return k;
}
__device__ __forceinline__ void pr_scan_tile(const float4* __restrict__ tile, int grp, const uint64_t (&nqx)[kPrQPL],
                                             const uint64_t (&nqy)[kPrQPL], const uint64_t (&nqz)[kPrQPL], uint64_t one,
                                             unsigned long long (&bk)[kPrQPL]) {
  const ulonglong2* tp = reinterpret_cast<const ulonglong2*>(tile) + 2 * grp;
#pragma unroll 4
  for (int i = 0; i < kPrTile / 2 / kPrQPL; ++i) {
    const ulonglong2 A = __ldg(tp + 2 * kPrQPL * i);      // {x0, x1} {y0, y1}   one address per lane group
    const ulonglong2 B = __ldg(tp + 2 * kPrQPL * i + 1);  // {z0, z1} {w0, w1}
    float w0, w1;
    upk2(B.y, w0, w1);
#pragma unroll
    for (int q = 0; q < kPrQPL; ++q) {
      float d0, d1;
      upk2(sqdist2(A.x, A.y, B.x, nqx[q], nqy[q], nqz[q], one), d0, d1);
      const unsigned long long k0 = pr_key(d0, w0), k1 = pr_key(d1, w1);
      const unsigned long long k = k0 < k1 ? k0 : k1;
      bk[q] = k < bk[q] ? k : bk[q];
    }
  }
  // every lane group has seen a quarter of the tile: combine, so that all copies of a query agree again
#pragma unroll
  for (int q = 0; q < kPrQPL; ++q) {
    unsigned long long o = __shfl_xor_sync(0xffffffffu, bk[q], 8);
    bk[q] = o < bk[q] ? o : bk[q];
    o = __shfl_xor_sync(0xffffffffu, bk[q], 16);
    bk[q] = o < bk[q] ? o : bk[q];
  }
}

// SPLIT == 1: every warp of the CTA owns its own group of 32 queries.  SPLIT == kPrWarps: the warps of the CTA SHARE one
// group and each takes the target tiles t with t % SPLIT == its number -- for clouds too small to fill the machine with
// one warp per group (C3: 3 125 groups for 592 schedulers; the search is then a chain of dependent tile visits and
// latency-bound).  The warps publish their current best distances in shared memory (atomicMin on the bit patterns: any
// published value is a real distance of that query, so pruning with it is as valid as pruning with the own one) and
// merge their keys there at the end.
template <int SPLIT>
__global__ void __launch_bounds__(32 * kPrWarps) nnd_pr_search_kernel(PrSearch p) {
  static_assert(SPLIT == 1 || SPLIT == kPrWarps, "one group per warp or one per CTA");
  constexpr unsigned FULL = 0xffffffffu;
  const int batch = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int slot = lane & 7, grp = lane >> 3;  // queries 4 * slot .. 4 * slot + 3 of the group; lane group
  const int warp = threadIdx.x >> 5;
  const int part = SPLIT == 1 ? 0 : warp;  // which share of the tiles this warp takes
  const int wq = (SPLIT == 1 ? blockIdx.x * kPrWarps + warp : blockIdx.x) * 32;  // first sorted query of the group
  const int nq = p.q.n;
  if (wq >= nq) return;  // SPLIT > 1: the whole CTA leaves together
  __shared__ unsigned best_s[32];
  __shared__ unsigned long long key_s[32];
  const int qi = slot * kPrQPL + grp;  // "my" query inside the group
  if (SPLIT > 1) {
    if (warp == 0) {
      best_s[lane] = 0x7f800000u;
      key_s[lane] = 0x7f8000007fffffffull;
    }
    __syncthreads();
  }
  const float INF = __int_as_float(0x7f800000);
  const uint64_t one = pk2(p.one, p.one);
  uint64_t nqx[kPrQPL], nqy[kPrQPL], nqz[kPrQPL];
  unsigned long long bk[kPrQPL];
  float mx = 0.f, my = 0.f, mz = 0.f;  // the query this lane tests boxes for and writes out: number grp of its slot
  int mw = 0;
  PrBox qb{INF, INF, INF, -INF, -INF, -INF, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < kPrQPL; ++i) {
    const int s = wq + slot * kPrQPL + i < nq ? wq + slot * kPrQPL + i : nq - 1;  // dead slots repeat the last query
    const float* o = reinterpret_cast<const float*>(p.q.pts + (long long)batch * p.q.tiles * kPrTile + (s & ~1)) + (s & 1);
    const float x = o[0], y = o[2], z = o[4];
    nqx[i] = pk2(-x, -x);
    nqy[i] = pk2(-y, -y);
    nqz[i] = pk2(-z, -z);
    bk[i] = 0x7f8000007fffffffull;  // (+inf, INT_MAX)
    if (i == grp) {
      mx = x;
      my = y;
      mz = z;
      mw = __float_as_int(o[6]);
    }
    // the warp's query box (NaN coordinates dropped by fminf / fmaxf)
    qb.lx = fminf(qb.lx, x);
    qb.ly = fminf(qb.ly, y);
    qb.lz = fminf(qb.lz, z);
    qb.hx = fmaxf(qb.hx, x);
    qb.hy = fmaxf(qb.hy, y);
    qb.hz = fmaxf(qb.hz, z);
  }
  qb.lx = pr_wmin(qb.lx);
  qb.ly = pr_wmin(qb.ly);
  qb.lz = pr_wmin(qb.lz);
  qb.hx = pr_wmax(qb.hx);
  qb.hy = pr_wmax(qb.hy);
  qb.hz = pr_wmax(qb.hz);

  const int T = p.t.tiles, ST = p.t.stiles, LT = p.t.ltiles;
  const PrBox* lbox = p.t.lbox + (long long)batch * LT;
  const float4* tpts = p.t.pts + (long long)batch * T * kPrTile;
  const PrBox* tbox = p.t.tbox + (long long)batch * T;
  const PrBox* sbox = p.t.sbox + (long long)batch * ST;
  const unsigned* tcode = p.t.tcode + (long long)batch * T;

  float best;  // current best distance of "my" query (its bit pattern is the high key word)
  auto mine = [&]() {
    const unsigned long long k = grp == 0 ? bk[0] : grp == 1 ? bk[1] : grp == 2 ? bk[2] : bk[3];
    best = __uint_as_float((unsigned)(k >> 32));
    return k;
  };
  // ---- seed: the target tile at this warp's position on the curve (last tile whose first code is <= ours)
  int t0;
  {
    const unsigned qcode = (unsigned)(p.q.keys[(long long)batch * nq + wq] >> 32);
    int lo = 0, hi = T - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (__ldg(tcode + mid) <= qcode) lo = mid; else hi = mid - 1;
    }
    t0 = lo;
  }
  pr_scan_tile(tpts + (long long)t0 * kPrTile, grp, nqx, nqy, nqz, one, bk);  // (every warp of a shared group: good bounds)
  mine();
  auto share = [&]() {  // publish my query's best distance, adopt the other warps' if lower
    if (SPLIT > 1) best = fminf(best, __uint_as_float(atomicMin(&best_s[qi], __float_as_uint(best))));
  };
  share();
#ifdef NND_PR_STATS
  unsigned long long n_scan = 1, n_super = 0, n_tile = 0;
#endif
  float bmax = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(best)));  // best >= +0 or +inf: bit order = value order

  // ---- large tiles -> super-tiles -> tiles -> per-query test -> scan
  for (int lb = 0; lb < LT; lb += 32) {
   bool lpass = false;
   if (lb + lane < LT) lpass = !(pr_box_box(qb, lbox[lb + lane]) > bmax);
   unsigned lmask = __ballot_sync(FULL, lpass);
   while (lmask) {
    const int sb = (lb + __ffs(lmask) - 1) * kPrSuper;  // first super-tile of this large tile
    lmask &= lmask - 1;
    const int st = sb + lane;
    bool pass = false;
    if (st < ST) pass = !(pr_box_box(qb, sbox[st]) > bmax);
    unsigned smask = __ballot_sync(FULL, pass);
#ifdef NND_PR_STATS
    n_super += __popc(smask);
#endif
    while (smask) {
      const int sst = sb + __ffs(smask) - 1;
      smask &= smask - 1;
      if (SPLIT > 1) {  // what the other warps of the group found meanwhile
        best = fminf(best, __uint_as_float(*reinterpret_cast<volatile unsigned*>(&best_s[qi])));
        bmax = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(best)));
      }
      const int tile = sst * kPrSuper + lane;
      bool pass2 = false;
      if (tile < T && tile != t0 && (SPLIT == 1 || tile % SPLIT == part)) pass2 = !(pr_box_box(qb, tbox[tile]) > bmax);
      unsigned tmask = __ballot_sync(FULL, pass2);
#ifdef NND_PR_STATS
      n_tile += __popc(tmask);
#endif
      while (tmask) {
        const int t = sst * kPrSuper + __ffs(tmask) - 1;
        tmask &= tmask - 1;
        const PrBox bx = tbox[t];
        if (__any_sync(FULL, !(pr_point_box(mx, my, mz, bx) > best))) {  // every query of the warp is some lane's
          pr_scan_tile(tpts + (long long)t * kPrTile, grp, nqx, nqy, nqz, one, bk);
          mine();
          share();
          bmax = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(best)));
#ifdef NND_PR_STATS
          ++n_scan;
#endif
        }
      }
    }
   }
  }
#ifdef NND_PR_STATS
  if (lane == 0) {
    atomicAdd(&g_pr_stats[0], n_scan);
    atomicAdd(&g_pr_stats[1], n_super);
    atomicAdd(&g_pr_stats[2], n_tile);
    atomicAdd(&g_pr_stats[3], 1ull);
  }
#endif
  // ---- epilogue: my_lib.cpp:16 (k = 0 is accepted unconditionally) and the scatter to the original query order
  if (SPLIT > 1) {  // merge the warps' keys; warp 0 writes
    atomicMin(&key_s[qi], mine());
    __syncthreads();
    if (warp != 0) return;
  }
  if (wq + slot * kPrQPL + grp < nq) {
    unsigned long long k = mine();
    if (SPLIT > 1) {
      k = key_s[qi];
      best = __uint_as_float((unsigned)(k >> 32));
    }
    const float* t0p = p.t.xyz + (long long)batch * p.t.n * 3;
    const float dx = __fsub_rn(__ldg(t0p), mx), dy = __fsub_rn(__ldg(t0p + 1), my), dz = __fsub_rn(__ldg(t0p + 2), mz);
    const float d0 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    const bool first = d0 != d0 || !(best < INF);
    const long long o = (long long)batch * nq + mw;
    p.dist[o] = first ? d0 : best;
    p.idx[o] = first ? 0 : (int32_t)(unsigned)(k & 0xffffffffull);
  }
}
