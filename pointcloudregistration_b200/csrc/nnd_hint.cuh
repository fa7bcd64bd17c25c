// nnd_hint.cuh -- included by nnd_kernels.cu inside its anonymous namespace, before the launchers.
//
// Automatic hints for the column filter of the symmetric scans (the cold-start twin of nnd_forward_warm_cuda_f32).
//
// The symmetric kernels keep, per column (xyz2 point), an upper bound of its current minimum and take a slow path
// whenever a warp's rows reach that bound.  In the panel scan all ~300 resident CTAs of a launch work on the same
// columns at the same time, each from the bound it found in global memory when it started, so with a bound that is only
// "the best of S sample rows" every warp visit still has a 256 / (S + 256) chance of an event: at S = 2048 that is a tenth
// of the visits of the first wave.  A bound from a row that is actually NEAR the column removes almost all of them
// (measured with perfect hints, profiles/r02_warm_ab.txt: C4 251.2 -> 242.5 ms, 262 144 points 18.6 -> 16.9 ms).
//
// So, before the seed pass: hash the rows into uniform grids of kHintLevels resolutions (cell edge halving per level; one
// representative row per cell, last writer wins -- any row is a valid hint, the results never depend on which one),
// then for every column look at the 27 cells around it, finest level first, and keep the nearest representative of the
// first level that has any (clustered clouds are hundreds of points per coarse cell: the fine levels find a row at
// the local spacing; uniform clouds fall through to the coarse level).  The seed pass evaluates the distance to that
// row with the library's exact packed sequence, which makes it a true entry of the column, hence a valid bound.
// O(N + M) work, a few tenths of a percent of the scan at 1M.
//
// Measured (profiles/r02_hint_ab2.txt, forward): C4 uniform 251.4 -> 243.1 ms (85.5 -> 88.4 % of the FP32 issue peak),
// 262 144 points uniform / crosscopy / lattice 18.6-18.9 -> 17.1 ms (79 -> 86 %).
//
// The "clustered" generator's two clouds have DIFFERENT cluster centres: a column lies in empty space ~0.1 away from the
// nearest cluster of rows and no cell around it holds a row at any level (5 levels change nothing, r02_hint_ab_l5.txt).
// Such columns take a coarse-to-fine DESCENT instead (end of nnd_hint_query_kernel): the table also holds coarser levels
// (cells doubling per level down to <= 6 per axis); the column takes the nearest representative of all cells of the
// coarsest one, then per finer level the nearest representative of the 27 cells around the current best row.
// Measured (profiles/r02_hint_descent_ab.txt): C4 clustered 253.6 -> 244.5 ms (84.8 -> 87.9 %; true neighbours as hints
// would give 88.6 %, r02_warm_ab_clustered.txt), C5 clustered 18.44 -> 17.49 ms (80.1 -> 84.4 %), everything else unchanged.
// In front of the persistent kernel the hints cost more than they save at C2 / C3 (mode 2), so they are on for the
// panel scans only.

#ifndef NND_HINT_LEVELS
#define NND_HINT_LEVELS 3
#endif
constexpr int kHintLevels = NND_HINT_LEVELS;

struct HintParams {
  const float* rows;     // [b][row_stride][3]  the cloud held in registers by the scan ("rows")
  const float* cols;     // [b][col_stride][3]  the streamed cloud ("columns")
  long long row_stride;  // points between batch elements of `rows`
  long long col_stride;
  int n, m, b;           // rows / columns per batch element
  unsigned* bbox;        // [b][6] ordered-uint min x,y,z then max x,y,z
  int* table;            // [b][tsize] representative row per hashed cell, -1 = empty
  int tsize;             // power of two
  int g;                 // cells per axis at level 0 (level l >= 0 has g << l, level -l has g >> l)
  int coarse;            // coarse levels -1 .. -coarse below level 0 for the descent of far columns (0 = none)
  int32_t* hint;         // [b][m] out
};

__device__ __forceinline__ unsigned f2ord(float f) {  // monotone float -> uint
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// bbox[b][0..2] = 0xffffffff, bbox[b][3..5] = 0 and table = -1 (one pass, grid-stride)
__global__ void __launch_bounds__(256) nnd_hint_clear_kernel(HintParams p) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (long long i = first; i < (long long)p.b * p.tsize; i += stride) p.table[i] = -1;
  for (long long i = first; i < (long long)p.b * 6; i += stride) p.bbox[i] = (i % 6) < 3 ? 0xffffffffu : 0u;
}

__global__ void __launch_bounds__(256) nnd_hint_bbox_kernel(HintParams p) {
  const int batch = blockIdx.y;
  const float* r = p.rows + (long long)batch * p.row_stride * 3;
  unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < p.n; j += gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = __ldg(r + 3 * (long long)j + c);
      if (v == v && fabsf(v) < 3e38f) {  // finite values only: a stray inf / NaN must not blow the grid up
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
      atomicMin(p.bbox + batch * 6 + c, lo[c]);
      atomicMax(p.bbox + batch * 6 + 3 + c, hi[c]);
    }
  }
}

struct HintGrid {
  float lo[3], inv[3];
};
__device__ __forceinline__ HintGrid hint_grid(const HintParams& p, int batch) {
  HintGrid g;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const unsigned ulo = p.bbox[batch * 6 + c], uhi = p.bbox[batch * 6 + 3 + c];
    const float lo = ulo <= uhi ? ord2f(ulo) : 0.f, hi = ulo <= uhi ? ord2f(uhi) : 0.f;
    const float ext = hi - lo;
    g.lo[c] = lo;
    g.inv[c] = ext > 0.f ? (float)p.g / ext : 0.f;
  }
  return g;
}
__device__ __forceinline__ int hint_cell(float v, float lo, float inv, int g) {
  return (int)fminf(fmaxf((v - lo) * inv, 0.f), (float)(g - 1));  // fmaxf / fminf drop NaNs: always a valid cell
}
__device__ __forceinline__ unsigned hint_hash(int ix, int iy, int iz, int level) {
  return (unsigned)ix * 73856093u ^ (unsigned)iy * 19349663u ^ (unsigned)iz * 83492791u ^
         (unsigned)(level + 16) * 2654435761u;
}
// cells per axis at `level` (negative: the coarse levels of the descent) and that level's scale relative to level 0
// (exactly 2^level for level >= 0)
__device__ __forceinline__ int hint_cells(int g, int level) { return level >= 0 ? g << level : max(g >> -level, 1); }
__device__ __forceinline__ float hint_scale(int g, int level) { return (float)hint_cells(g, level) / (float)g; }

__global__ void __launch_bounds__(256) nnd_hint_build_kernel(HintParams p) {
  const int batch = blockIdx.y;
  const HintGrid g = hint_grid(p, batch);
  const float* r = p.rows + (long long)batch * p.row_stride * 3;
  int* table = p.table + (long long)batch * p.tsize;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < p.n; j += gridDim.x * blockDim.x) {
    const float x = __ldg(r + 3 * (long long)j), y = __ldg(r + 3 * (long long)j + 1), z = __ldg(r + 3 * (long long)j + 2);
    for (int l = -p.coarse; l < kHintLevels; ++l) {
      const int gl = hint_cells(p.g, l);
      const float sc = hint_scale(p.g, l);
      const unsigned h = hint_hash(hint_cell(x, g.lo[0], g.inv[0] * sc, gl), hint_cell(y, g.lo[1], g.inv[1] * sc, gl),
                                   hint_cell(z, g.lo[2], g.inv[2] * sc, gl), l);
      table[h & (unsigned)(p.tsize - 1)] = j;  // last writer wins: any row of the cell will do
    }
  }
}

__global__ void __launch_bounds__(256) nnd_hint_query_kernel(HintParams p) {
  const int batch = blockIdx.y;
  const HintGrid g = hint_grid(p, batch);
  const float* r = p.rows + (long long)batch * p.row_stride * 3;
  const float* c = p.cols + (long long)batch * p.col_stride * 3;
  const int* table = p.table + (long long)batch * p.tsize;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < p.m; k += gridDim.x * blockDim.x) {
    const float x = __ldg(c + 3 * (long long)k), y = __ldg(c + 3 * (long long)k + 1), z = __ldg(c + 3 * (long long)k + 2);
    float best = __int_as_float(0x7f800000);
    int bj = -1;
    for (int l = kHintLevels - 1; l >= 0 && bj < 0; --l) {  // finest level first; stop at the first level with a row nearby
      const int gl = p.g << l;
      const float sc = (float)(1 << l);
      const int cx = hint_cell(x, g.lo[0], g.inv[0] * sc, gl), cy = hint_cell(y, g.lo[1], g.inv[1] * sc, gl),
                cz = hint_cell(z, g.lo[2], g.inv[2] * sc, gl);
#pragma unroll 1
      for (int dz = -1; dz <= 1; ++dz)
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
          for (int dx = -1; dx <= 1; ++dx) {
            const int ix = cx + dx, iy = cy + dy, iz = cz + dz;
            if (ix < 0 || iy < 0 || iz < 0 || ix >= gl || iy >= gl || iz >= gl) continue;
            const int j = table[hint_hash(ix, iy, iz, l) & (unsigned)(p.tsize - 1)];
            if (j < 0) continue;
            const float ex = __ldg(r + 3 * (long long)j) - x, ey = __ldg(r + 3 * (long long)j + 1) - y,
                        ez = __ldg(r + 3 * (long long)j + 2) - z;
            const float d = ex * ex + ey * ey + ez * ez;  // only ranks the candidates; the bound is evaluated exactly later
            // a hash collision can return a row of a far-away cell (which would end the search at this level with a
            // useless bound): accept only rows within two cell edges of this level
            const float ux = ex * g.inv[0] * sc, uy = ey * g.inv[1] * sc, uz = ez * g.inv[2] * sc;
            if (ux * ux + uy * uy + uz * uz <= 4.0f && d < best) {
              best = d;
              bj = j;
            }
          }
    }
    // Far columns (no row within two cell edges at any level: the clouds do not overlap here, e.g. clustered clouds with
    // different centres).  Coarse-to-fine descent instead: the nearest representative of ALL cells of the coarsest level
    // (at most 6^3), then per finer level the nearest representative of the 27 cells around the CURRENT BEST ROW's cell --
    // each step moves at most one cell towards the column, and the cells halve, so the walk ends on the near face of
    // the nearest cluster, at the finest level's resolution.  A colliding table entry is just another candidate row.
    if (bj < 0 && p.coarse > 0) {
      float px = 0.f, py = 0.f, pz = 0.f;  // position of the current best row
      auto probe = [&](int ix, int iy, int iz, int lv) {
        const int j = table[hint_hash(ix, iy, iz, lv) & (unsigned)(p.tsize - 1)];
        if (j < 0) return;
        const float rx = __ldg(r + 3 * (long long)j), ry = __ldg(r + 3 * (long long)j + 1), rz = __ldg(r + 3 * (long long)j + 2);
        const float ex = rx - x, ey = ry - y, ez = rz - z;
        const float d = ex * ex + ey * ey + ez * ez;
        if (d < best) {
          best = d;
          bj = j;
          px = rx;
          py = ry;
          pz = rz;
        }
      };
      int lv = -p.coarse;
      {
        const int gl = hint_cells(p.g, lv);
#pragma unroll 1
        for (int iz = 0; iz < gl; ++iz)
#pragma unroll 1
          for (int iy = 0; iy < gl; ++iy)
#pragma unroll 1
            for (int ix = 0; ix < gl; ++ix) probe(ix, iy, iz, lv);
      }
#pragma unroll 1
      for (++lv; lv < kHintLevels && bj >= 0; ++lv) {
        const int gl = hint_cells(p.g, lv);
        const float sc = hint_scale(p.g, lv);
        const int cx = hint_cell(px, g.lo[0], g.inv[0] * sc, gl), cy = hint_cell(py, g.lo[1], g.inv[1] * sc, gl),
                  cz = hint_cell(pz, g.lo[2], g.inv[2] * sc, gl);
#pragma unroll 1
        for (int dz = -1; dz <= 1; ++dz)
#pragma unroll 1
          for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
              const int ix = cx + dx, iy = cy + dy, iz = cz + dz;
              if (ix < 0 || iy < 0 || iz < 0 || ix >= gl || iy >= gl || iz >= gl) continue;
              probe(ix, iy, iz, lv);
            }
      }
    }
    p.hint[(long long)batch * p.m + k] = bj;
  }
}
