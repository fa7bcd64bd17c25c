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
// then finds its segment by binary search and accumulates it in order.  No atomics, no zero-init pass: every row
// is written exactly once.  Unlike the REDG path there is no flush-to-zero of denormal sums.

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

__global__ void __launch_bounds__(256) nnd_grad_exact_kernel(ExactParams p) {
  const long long stride = (long long)gridDim.x * blockDim.x;
#pragma unroll 1
  for (int si = 0; si < 2; ++si) {
    const ExactSide s = p.s[si];
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < s.rows; r += stride) {
      const long long batch = r / s.na;
      const float* me = s.a + r * 3;
      float ox, oy, oz;  // own term: this row against its neighbour
      exact_term(me, s.o + (batch * s.no + s.idx_a[r]) * 3, s.gd_a[r], ox, oy, oz);
      float ax = 0.0f, ay = 0.0f, az = 0.0f;
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
      // the segment, in ascending order of the contributing row; 4 entries' loads in flight, adds strictly in order
      bool more = true;
      for (long long e = lo; more; e += 4) {
        float tx[4], ty[4], tz[4];
        bool ok[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const unsigned long long k = e + u < s.nkeys ? s.keys[e + u] : ~0ull;
          ok[u] = (k >> 32) == (unsigned long long)r;
          tx[u] = ty[u] = tz[u] = 0.0f;
          if (ok[u]) {
            const long long src = (long long)(k & 0xffffffffull);  // global row of the other cloud
            exact_term(s.o + src * 3, me, s.gd_o[src], tx[u], ty[u], tz[u]);
          }
        }
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
