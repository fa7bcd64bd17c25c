// Dev probe (compile only: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -cubin, then cuobjdump -sass | grep "c\[0x3\]"):
// when does ptxas keep __constant__ operands in UNIFORM registers (LDCU + "FADD2 R, -R.F32, UR.F32x2") and when does it
// fall back to LDC into vector registers?  Findings that shaped nnd_symc.cuh (CUDA 12.9, sm_100a):
//   * the address induction variable must have NO per-thread consumer (index arithmetic with threadIdx, a value
//     stored by a thread, an argument of the slow-path call): keep a second, opaque counter for those (kv in nnd_symc.cuh);
//   * no runtime division / modulo / min() in the chain that feeds the address (64-bit division is a call, 32-bit
//     division by a non-power-of-two uses IMAD.HI, min() is IMNMX: none exist on the uniform datapath);
//   * no uniform `if` / `break` inside the loops around the loads (first/last-unit logic of a persistent work loop);
//   * a warp vote + conditional call is fine; plain counted loops are fine.
// V bits below switch the offending constructs on one at a time.
#include <cstdint>
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b){uint64_t r; asm("add.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b){uint64_t r; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ uint64_t pk2(float lo, float hi){uint64_t r; asm("mov.b64 %0, {%1, %2};":"=l"(r):"f"(lo),"f"(hi)); return r;}
__constant__ float cp[2][7680];
struct P { int which, per, rem, m, live; };
constexpr int NS = 20;
template<int V> __global__ void k(const __grid_constant__ P p, const float* q, uint64_t* out) {
  float s = -q[threadIdx.x];
  uint64_t a = pk2(s, s);
  uint64_t acc = 0;
  const float* c = cp[p.which];
  const unsigned bx = blockIdx.x;
  unsigned ub = bx * p.per, ue = ub + p.per;
  if (V & 1) { ub = bx * p.per + (bx < (unsigned)p.rem ? bx : (unsigned)p.rem); ue = ub + p.per + (bx < (unsigned)p.rem ? 1u : 0u); }
  for (unsigned u = ub; u < ue;) {
    unsigned qb = u / NS; int s0 = u - qb * NS;
    int seg = 1;
    if (V & 2) { int rest = NS - s0; seg = (int)(ue-u) < rest ? (int)(ue-u) : rest; }
    if (V & 4) a = pk2(-q[qb * 256 + threadIdx.x], s);
    int kend = (s0 + seg) * 128;
    if (V & 8) kend = min(kend, p.live);
    for (int k0 = s0 * 128; k0 < kend; k0 += 32) {
      int tl = k0 >> 9, ko = k0 & 511;
      const float* cx = c + tl * 1536 + ko;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        ulonglong2 X = *reinterpret_cast<const ulonglong2*>(cx + 4 * i);
        uint64_t d = add2(X.x, a); acc = add2(acc, mul2(d, d)); d = add2(X.y, a); acc = add2(acc, mul2(d, d));
        if (V & 16) { if (__any_sync(0xffffffffu, (unsigned)acc < 5u)) atomicAdd((unsigned*)out, 1u); }
      }
    }
    if (V & 32) { if (qb * 256 + threadIdx.x < p.m) atomicMin((unsigned long long*)out + threadIdx.x, acc); }
    u += seg;
  }
  out[threadIdx.x] = acc;
}
#define I(v) template __global__ void k<v>(const __grid_constant__ P, const float*, uint64_t*);
I(0) I(1) I(2) I(4) I(8) I(16) I(32) I(63)
