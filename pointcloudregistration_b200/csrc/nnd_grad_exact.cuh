// nnd_grad_exact.cuh -- included by nnd_kernels.cu inside its anonymous namespace (needs <cub/device/device_radix_sort.cuh>).
//
// Deterministic backward: gradients BIT-IDENTICAL to the reference CPU path my_lib.cpp:91-130, run to run and
// against the oracle.  The atomic kernel (nnd_grad.cuh) sums the scatter side in arbitrary order, so it only meets
// the 1e-5 contract; this one reproduces the oracle's sequential order ("sort-by-idx segmented
// reduction"):
//     my_lib.cpp:96-113   for j ascending over cloud 1:  g1[j]        += t1(j) ;  g2[idx1[j]] -= t1(j)
//     my_lib.cpp:114-129  for j ascending over cloud 2:  g2[j]        += t2(j) ;  g1[idx2[j]] -= t2(j)
// hence, per element (all fp32, every operation rounded on its own):
//     g1[r] = ((0 + t1(r)) - t2(j_a)) - t2(j_b) ...      j_a < j_b < ... the cloud-2 rows whose neighbour is r
//     g2[k] = (((0 - t1(j_a)) - t1(j_b)) ...) + t2(k)    j_a < j_b < ... the cloud-1 rows whose neighbour is k
// The "rows whose neighbour is r, ascending" lists come from sorting the 64-bit keys (neighbour row << 32 | own row)
// with cub::DeviceRadixSort (library call: plain sort, not on the forward's hot path); one thread per gradient row
// then finds its segment by binary search and accumulates it in order (the first 8 entries on its own; the rest of a
// long segment with the whole warp fetching 32 entries at a time while the sum itself stays sequential).  No atomics,
// no zero-init pass: every row is written exactly once.  Unlike the REDG path there is no flush-to-zero of denormal sums.

struct ExactSide {
  const float* a;                    // [b][na][3]  the cloud whose gradient this side produces
  const float* o;                    // [b][no][3]  the other cloud
  const float* gd_a;                 // [b][na]     upstream gradient of this cloud's own term
  const float* gd_o;                 // [b][no]     upstream gradient of the other cloud's term (the scatter source)
  const int32_t* idx_a;              // [b][na]     neighbour (in the other cloud) of every row of this cloud
  const unsigned long long* keys;    // [b*no] sorted (global row of this cloud << 32 | global row of the other cloud)
  float* grad;                       // [b][na][3]
  int na, no;
  long long rows, nkeys;             // b*na, b*no
  int own_first;                     // 1: own term first, then the scatter terms (cloud 1); 0: scatter terms first (cloud 2)
};
struct ExactParams {
  ExactSide s[2];
};

struct ExactKeySide {
  const int32_t* idx;        // [b][nq]
  unsigned long long* keys;  // [b*nq]
  int nq, nt;
  long long rows;
};
struct ExactKeyParams {
  ExactKeySide s[2];
};

// keys[r] = (global neighbour row << 32) | global own row
__global__ void __launch_bounds__(256) nnd_exact_keys_kernel(ExactKeyParams p) {
  const long long stride = (long long)gridDim.x * blockDim.x;
#pragma unroll
  for (int si = 0; si < 2; ++si) {
    const ExactKeySide s = p.s[si];
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < s.rows; r += stride) {
      const long long batch = r / s.nq;
      const unsigned long long nb = (unsigned long long)(batch * s.nt + s.idx[r]);
      s.keys[r] = (nb << 32) | (unsigned long long)r;
    }
  }
}

// t_c = fl(fl(2 * gd) * fl(q_c - t_c))   (my_lib.cpp:105-109)
__device__ __forceinline__ void exact_term(const float* __restrict__ q, const float* __restrict__ t, float gd, float& x,
                                           float& y, float& z) {
  const float g = __fmul_rn(gd, 2.0f);
  x = __fmul_rn(g, __fsub_rn(q[0], t[0]));
  y = __fmul_rn(g, __fsub_rn(q[1], t[1]));
  z = __fmul_rn(g, __fsub_rn(q[2], t[2]));
}

// One entry of a segment as seen by the cooperative phase: is key `idx` a contribution to row `row`, and if so its term.
__device__ __forceinline__ bool exact_entry(const ExactSide& s, long long idx, unsigned long long row, float mx, float my,
                                            float mz, float& tx, float& ty, float& tz) {
  const unsigned long long k = idx < s.nkeys ? s.keys[idx] : ~0ull;
  const bool ok = (k >> 32) == row;
  tx = ty = tz = 0.0f;
  if (ok) {
    const long long src = (long long)(k & 0xffffffffull);  // global row of the other cloud
    const float* q = s.o + src * 3;
    const float g = __fmul_rn(s.gd_o[src], 2.0f);
    tx = __fmul_rn(g, __fsub_rn(q[0], mx));
    ty = __fmul_rn(g, __fsub_rn(q[1], my));
    tz = __fmul_rn(g, __fsub_rn(q[2], mz));
  }
  return ok;
}

constexpr int kExactSerial = 8;  // entries a thread sums on its own before the warp takes over the rest of its segment

__global__ void __launch_bounds__(256) nnd_grad_exact_kernel(ExactParams p) {
  constexpr unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31u;
  const long long stride = (long long)gridDim.x * blockDim.x;  // a multiple of 32: the lanes of a warp iterate together
#pragma unroll 1
  for (int si = 0; si < 2; ++si) {
    const ExactSide s = p.s[si];
    const long long rows32 = (s.rows + 31) / 32 * 32;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rows32; r += stride) {
      const bool live = r < s.rows;
      float mx = 0.0f, my = 0.0f, mz = 0.0f;  // this row
      float ox = 0.0f, oy = 0.0f, oz = 0.0f;  // own term: this row against its neighbour
      float ax = 0.0f, ay = 0.0f, az = 0.0f;
      bool more = false;  // the segment goes on beyond what this thread has summed
      long long e = 0;    // next entry of the segment
      if (live) {
        const long long batch = r / s.na;
        const float* me = s.a + r * 3;
        mx = me[0];
        my = me[1];
        mz = me[2];
        exact_term(me, s.o + (batch * s.no + s.idx_a[r]) * 3, s.gd_a[r], ox, oy, oz);
        if (s.own_first) {  // 0 + t keeps the reference's sign of zero (+0 + -0 = +0)
          ax = __fadd_rn(ax, ox);
          ay = __fadd_rn(ay, oy);
          az = __fadd_rn(az, oz);
        }
        // first key >= (r << 32): start of this row's segment
        const unsigned long long want = (unsigned long long)r << 32;
        long long lo = 0, hi = s.nkeys;
        while (lo < hi) {
          const long long mid = (lo + hi) >> 1;
          if (s.keys[mid] < want) lo = mid + 1; else hi = mid;
        }
        // the first entries, in ascending order of the contributing row; 4 entries' loads in flight, adds strictly in order
        more = true;
        e = lo;
#pragma unroll 1
        for (int bt = 0; bt < kExactSerial / 4 && more; ++bt, e += 4) {
          float tx[4], ty[4], tz[4];
          bool ok[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) ok[u] = exact_entry(s, e + u, (unsigned long long)r, mx, my, mz, tx[u], ty[u], tz[u]);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (ok[u]) {
              ax = __fsub_rn(ax, tx[u]);
              ay = __fsub_rn(ay, ty[u]);
              az = __fsub_rn(az, tz[u]);
            } else {
              more = false;  // keys of one row are contiguous: nothing follows
            }
          }
        }
      }
      // Long segments (a point that thousands of rows chose): the warp loads 32 entries at a time, the sum itself stays
      // strictly sequential (every lane runs the same chain on shuffled terms; the owner keeps the result).  The next
      // 32 entries are fetched while the current ones are being summed.
      unsigned pending = __ballot_sync(FULL, more);
      while (pending) {
        const int owner = __ffs(pending) - 1;
        pending &= pending - 1;
        const unsigned long long row = (unsigned long long)__shfl_sync(FULL, r, owner);
        long long at = __shfl_sync(FULL, e, owner) + lane;
        const float qx = __shfl_sync(FULL, mx, owner), qy = __shfl_sync(FULL, my, owner), qz = __shfl_sync(FULL, mz, owner);
        float cx = __shfl_sync(FULL, ax, owner), cy = __shfl_sync(FULL, ay, owner), cz = __shfl_sync(FULL, az, owner);
        float tx, ty, tz;
        bool ok = exact_entry(s, at, row, qx, qy, qz, tx, ty, tz);
        while (true) {
          float nx, ny, nz;
          const bool nok = exact_entry(s, at + 32, row, qx, qy, qz, nx, ny, nz);  // prefetch
          const unsigned okm = __ballot_sync(FULL, ok);
          const int cnt = okm == FULL ? 32 : __ffs(~okm) - 1;  // leading run of entries that belong to the row
          if (cnt == 32) {
#pragma unroll
            for (int u = 0; u < 32; ++u) {
              cx = __fsub_rn(cx, __shfl_sync(FULL, tx, u));
              cy = __fsub_rn(cy, __shfl_sync(FULL, ty, u));
              cz = __fsub_rn(cz, __shfl_sync(FULL, tz, u));
            }
          } else {
            for (int u = 0; u < cnt; ++u) {
              cx = __fsub_rn(cx, __shfl_sync(FULL, tx, u));
              cy = __fsub_rn(cy, __shfl_sync(FULL, ty, u));
              cz = __fsub_rn(cz, __shfl_sync(FULL, tz, u));
            }
            break;
          }
          ok = nok;
          tx = nx;
          ty = ny;
          tz = nz;
          at += 32;
        }
        if ((int)lane == owner) {
          ax = cx;
          ay = cy;
          az = cz;
        }
      }
      if (live) {
        if (!s.own_first) {
          ax = __fadd_rn(ax, ox);
          ay = __fadd_rn(ay, oy);
          az = __fadd_rn(az, oz);
        }
        float* g = s.grad + r * 3;
        g[0] = ax;
        g[1] = ay;
        g[2] = az;
      }
    }
  }
}

inline int bits_for(unsigned long long v) {  // number of bits needed to represent v
  int b = 0;
  while (v) {
    ++b;
    v >>= 1;
  }
  return b;
}

inline size_t exact_sort_temp_bytes(long long count) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                 (int)count, 0, 64, (cudaStream_t)0);
  return bytes;
}
