// nnd_symc.cuh -- included by nnd_kernels.cu inside its anonymous namespace, after nnd_sym.cuh.
//
// Symmetric scan with the TARGETS IN UNIFORM REGISTERS (large row counts only).
//
// What limits nnd_scan_sym_kernel is not the FP32 pipe itself but register-file operand traffic: every packed
// FADD2 of the distance reads a 64-bit target pair AND a 32-bit query from the vector register file, next to the
// three-operand FMNMX3s (profiles/r01_ubench_step.txt: the same 4-target step runs at 82 % of the FP32 issue peak
// with the targets in vector registers and at 88 % with the targets in uniform registers).  Targets are the same
// for all lanes, so they can live in UNIFORM registers:  FADD2 R, -Rq.F32, UR.F32x2  reads one vector register
// instead of three.  The only way to load uniform registers from memory on sm_100a is LDCU from the constant bank,
// and a kernel cannot refill the constant bank for itself; so the scan becomes a sequence of launches, each over
// one PANEL of 1, 2 or 4 target tiles that the host copies (device -> constant, cudaMemcpyToSymbolAsync) into one
// of 8, 4 or 2 constant buffers.  Consecutive panels rotate over as many internal streams, so the next panel's CTAs
// back-fill the SMs as the previous panel's CTAs drain and the copy of panel i+2 waits only for panel i.
//
// Everything else is the symmetric algorithm of nnd_sym.cuh: rows in registers (QPT per thread), row minima in
// registers with the lazy 32-target chunk index in shared memory, column minima through the filtered slow path
// (sym_column_update), row keys merged with atomicMin.  No TMA ring and no block-wide barrier inside the loop: the
// panel is static for the whole launch, the column filter of the panel sits in shared memory.

#ifndef NND_TT
#define NND_TT 512
#endif
// Row bookkeeping (did this chunk improve a row? then remember where) runs after every NND_SYMC_SKIP-th 32-target chunk
// only; nnd_finalize_kernel then re-scans SKIP * 32 targets per row, which is nothing at the sizes this kernel serves.
// Measured at C4 (profiles/r02_ab_symc_skip.txt): 254.22 ms with 1, 253.53 with 2, 253.17 with 4, 253.2-253.8 with 8.
// Also measured, all worse (r02_ab_symc_chunk.txt, r02_ab_symc_chunk2.txt): a 64-target chunk (+16 %: the unrolled body
// no longer fits the instruction cache), a 16-target chunk (+2 %), the chunk loop unrolled 4 or 2 steps instead of 8
// (+0.7 / +1 %), the previous minima in registers instead of shared memory (+0.6 %).
#ifndef NND_SYMC_SKIP
#define NND_SYMC_SKIP 4
#endif
// The 48 KB constant window is cut into 2 buffers of 4 tiles, 4 of 2 or 8 of 1: fewer rows per launch (one CTA per
// row block) want more launches in flight to keep 2 CTAs on every SM.
constexpr int kPanelFloats = 8 * 3 * NND_TT;  // 8 tiles x 6 KB = 48 KB of the 64 KB constant bank
constexpr int kPanelMaxBufs = 8;
__constant__ float c_panel[kPanelFloats];

struct SymCParams {
  const float* qsoa;          // xyz1 as SoA tiles [b][qtiles][3][TT]
  unsigned long long* rkeys;  // [b][n]
  unsigned long long* ckeys;  // [b][ttiles * TT]
  unsigned* cdist;            // [b][ttiles * TT]
  int n, m;
  int row_lo, row_hi;
  int qtiles, ttiles;
  int batch;            // batch element this launch works on
  int tile_lo, ntiles;  // panel = target tiles [tile_lo, tile_lo + ntiles) of that batch element
  int cofs;             // float offset of the panel inside c_panel
  float one;
};

// One CTA = one row block (THREADS * QPT rows) against one panel; grid = row blocks.  Deliberately NOT persistent and
// free of data-dependent control flow around the target loads: ptxas only keeps the panel addresses in uniform
// registers (LDCU + UR operands) when the loops around them are plain counted loops -- a work-unit loop with
// divisions, min() or uniform `if`s silently falls back to LDC into vector registers (tools/uniform_probe.cu).
// PT: tiles per panel.  FULL: the panel has PT complete tiles (constant trip count).  !FULL: the last, ragged panel of a cloud.
template <int THREADS, int QPT, int TT, int C, int MINB, int PT, bool FULL>
__global__ void __launch_bounds__(THREADS, MINB) nnd_scan_symc_kernel(const __grid_constant__ SymCParams p) {
  static_assert(TT % C == 0 && C % 8 == 0 && QPT % 2 == 0 && TT % QPT == 0, "tile/chunk shape");
  constexpr int QB = THREADS * QPT;
  constexpr unsigned FULLMASK = 0xffffffffu;
  extern __shared__ __align__(128) unsigned char symc_smem[];
  unsigned* filt = reinterpret_cast<unsigned*>(symc_smem);  // [PT * TT]
  float* sprev = reinterpret_cast<float*>(filt + PT * TT) + threadIdx.x;
  int* sbase = reinterpret_cast<int*>(sprev - threadIdx.x + QPT * THREADS) + threadIdx.x;

  const int tid = threadIdx.x;
  const unsigned lane = tid & 31u;
  const int qblock = blockIdx.x;
  const int panel_base = p.tile_lo * TT;  // first target of the panel
  const int panel_live = FULL ? PT * TT : min(p.m - panel_base, p.ntiles * TT);

  unsigned long long* ckeys = p.ckeys + (long long)p.batch * p.ttiles * TT;
  unsigned* cdist = p.cdist + (long long)p.batch * p.ttiles * TT;
  for (int i = tid; i < p.ntiles * TT; i += THREADS) filt[i] = cdist[panel_base + i];

  // ---- rows -> registers: QPT consecutive rows per thread
  const uint64_t one = pk2(p.one, p.one);
  const float INF = __int_as_float(0x7f800000);
  uint64_t nqx[QPT], nqy[QPT], nqz[QPT];
  float cm[QPT];
  const unsigned group = (unsigned)(p.row_lo / QPT + qblock * THREADS + tid);
  const float* qbase = p.qsoa + (long long)p.batch * p.qtiles * (3 * TT);
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
  __syncthreads();

  const float* cpanel = c_panel + p.cofs;
  // Warps of a ragged last row block whose rows all lie beyond row_hi scan too (repeats of the last row: higher
  // group, never win a tie; their row keys are not posted): a thread-dependent branch here would cost the LDCUs.
  // kv: the same counter as k0, but opaque to the compiler and living in a vector register; every per-thread use
  // (filter address, slow-path arguments, lazy index) goes through kv so that k0 is left with the constant-bank
  // address as its only consumer and can stay in a uniform register.
  int kv = 0;
  asm volatile("mov.s32 %0, 0;" : "=r"(kv));
#pragma unroll 1
  for (int k0 = 0; k0 < panel_live; k0 += C) {          // one 32-target chunk per iteration
    const int tl = k0 / TT, ko = k0 - tl * TT;          // local tile, offset inside the tile
    const float* cx = cpanel + tl * (3 * TT) + ko;
    // one 4-target step: 32 distances, row minima into cm[], column minima of this thread's rows out as bit patterns
    auto step = [&](int s4, unsigned& u0, unsigned& u1, unsigned& u2, unsigned& u3) {
      const ulonglong2 X = *reinterpret_cast<const ulonglong2*>(cx + 4 * s4);
      const ulonglong2 Y = *reinterpret_cast<const ulonglong2*>(cx + TT + 4 * s4);
      const ulonglong2 Z = *reinterpret_cast<const ulonglong2*>(cx + 2 * TT + 4 * s4);
      float c0 = INF, c1 = INF, c2 = INF, c3 = INF;
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
    // Two steps share ONE filter test (vote + branch): 251.5 ms against 253.1 at C4 (profiles/r02_ab_symc_pair.txt); the
    // 8 column minima stay live across the test, which takes the kernel to its 128-register budget, no spills.
#pragma unroll
    for (int s4 = 0; s4 < C / 4; s4 += 2) {
      unsigned a0, a1, a2, a3, b0, b1, b2, b3;
      step(s4, a0, a1, a2, a3);
      step(s4 + 1, b0, b1, b2, b3);
      const int kl = kv + 4 * s4;  // local index of the first of these 8 columns
      const uint4 HA = *reinterpret_cast<const uint4*>(filt + kl);
      const uint4 HB = *reinterpret_cast<const uint4*>(filt + kl + 4);
      const bool ev = (a0 <= HA.x) | (a1 <= HA.y) | (a2 <= HA.z) | (a3 <= HA.w) | (b0 <= HB.x) | (b1 <= HB.y) |
                      (b2 <= HB.z) | (b3 <= HB.w);
      if (__any_sync(FULLMASK, ev)) {
        sym_column_update(a0, a1, a2, a3, HA.x, HA.y, HA.z, HA.w, lane, group, ckeys, cdist, filt, panel_base, kl);
        sym_column_update(b0, b1, b2, b3, HB.x, HB.y, HB.z, HB.w, lane, group, ckeys, cdist, filt, panel_base, kl + 4);
      }
    }
    // end of the chunk: did it improve a row?  (strict <, chunks ascend: the earliest chunk with the minimum stays)
#if NND_SYMC_SKIP > 1  // row bookkeeping after every NND_SYMC_SKIP-th chunk (the finalize re-scans SKIP * C targets)
    if ((kv & ((NND_SYMC_SKIP - 1) * C)) == (NND_SYMC_SKIP - 1) * C || kv + C >= panel_live)  // kv, not k0: a condition
#endif                                                                            // on the uniform counter costs the LDCUs
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
      if (cm[q] < sprev[q * THREADS]) {
        sprev[q * THREADS] = cm[q];
        sbase[q * THREADS] = panel_base + (kv & ~((NND_SYMC_SKIP - 1) * C));  // start of the group of chunks
      }
    }
    kv += C;
  }

  unsigned long long* rkeys = p.rkeys + (long long)p.batch * p.n;
#pragma unroll
  for (int q = 0; q < QPT; ++q) {
    const int j = p.row_lo + qblock * QB + tid * QPT + q;
    if (j < p.row_hi) {
      const unsigned long long key =
          ((unsigned long long)__float_as_uint(cm[q]) << 32) | (unsigned long long)(unsigned)sbase[q * THREADS];
      atomicMin(rkeys + j, key);
    }
  }
}

// Filter seed for the panel path.  All resident CTAs of a panel launch work on the same few thousand columns at the
// same time and each one starts from the filter values it finds in global memory, so on a cold filter (0xffffffff)
// every warp step of the first wave would take the column slow path.  This pre-pass gives every column an upper
// bound first: its exact distance (same packed sequence, (row - col)^2 == (col - row)^2 bit for bit) to the nearest
// of ns * TT sample rows.  Only the filter is written, not the column keys: the scan's `<=` test lets the sample
// rows through again when their turn comes, so nothing is lost.  Cost: ns * TT / N of the scan (0.2 % at 1M rows).
template <int TT>
__global__ void __launch_bounds__(256) nnd_seed_filter_kernel(const float* __restrict__ cols, const float* __restrict__ qsoa,
                                                              unsigned* __restrict__ cdist, int m, int ttiles,
                                                              int qtiles, int tile_first, int ns, float onef,
                                                              long long col_stride,
                                                              const int32_t* __restrict__ hint = nullptr,
                                                              const float* __restrict__ rows_aos = nullptr, int n = 0,
                                                              int row_lo = 0, int row_hi = 0) {
  __shared__ __align__(16) float rows[3 * TT];
  const int batch = blockIdx.y;
  const uint64_t one = pk2(onef, onef);
  const float INF = __int_as_float(0x7f800000);
  uint64_t ncx[2], ncy[2], ncz[2];
  float best[2] = {INF, INF};
  int col[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    col[k] = blockIdx.x * 512 + k * 256 + threadIdx.x;
    const int c = col[k] < m ? col[k] : m - 1;
    const float* s = cols + ((long long)batch * col_stride + c) * 3;  // cols may be a row range of a larger cloud
    const float x = -s[0], y = -s[1], z = -s[2];
    ncx[k] = pk2(x, x);
    ncy[k] = pk2(y, y);
    ncz[k] = pk2(z, z);
  }
  for (int t = 0; t < ns; ++t) {
    const float* src = qsoa + ((long long)batch * qtiles + tile_first + t) * (3 * TT);
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * TT; i += 256) rows[i] = src[i];
    __syncthreads();
#pragma unroll 4
    for (int g = 0; g < TT / 4; ++g) {
      const ulonglong2 X = *reinterpret_cast<const ulonglong2*>(rows + 4 * g);
      const ulonglong2 Y = *reinterpret_cast<const ulonglong2*>(rows + TT + 4 * g);
      const ulonglong2 Z = *reinterpret_cast<const ulonglong2*>(rows + 2 * TT + 4 * g);
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        float a0, a1, b0, b1;
        upk2(sqdist2(X.x, Y.x, Z.x, ncx[k], ncy[k], ncz[k], one), a0, a1);
        upk2(sqdist2(X.y, Y.y, Z.y, ncx[k], ncy[k], ncz[k], one), b0, b1);
        best[k] = fmin3(best[k], a0, a1);
        best[k] = fmin3(best[k], b0, b1);
      }
    }
  }
  // Warm start (nnd_forward_warm_cuda_f32): hint[k] is a row that was column k's neighbour before (the previous
  // iteration of a registration loop).  Its distance NOW, evaluated with the same packed sequence, is a real entry of
  // column k, hence a valid upper bound of the column minimum -- and almost always the minimum itself, so the scan
  // finds next to no column to improve.  Any index is safe: out-of-range hints are ignored.
  if (hint != nullptr) {
#pragma unroll
    for (int k = 0; k < 2; ++k)
      if (col[k] < m) {
        const int j = __ldg(hint + (long long)batch * m + col[k]);
        if (j >= row_lo && j < row_hi) {
          const float* r = rows_aos + ((long long)batch * n + j) * 3;
          const float rx = __ldg(r), ry = __ldg(r + 1), rz = __ldg(r + 2);
          float d0, d1;
          upk2(sqdist2(pk2(rx, rx), pk2(ry, ry), pk2(rz, rz), ncx[k], ncy[k], ncz[k], one), d0, d1);
          best[k] = fmin3(best[k], d0, d1);
        }
      }
  }
#pragma unroll
  for (int k = 0; k < 2; ++k)
    if (col[k] < m) cdist[(long long)batch * ttiles * TT + col[k]] = __float_as_uint(best[k]);
}
