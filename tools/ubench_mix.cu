// Micro-benchmark 2: cost of one extra instruction of each kind inside a packed-FP32 stream on sm_100a.
// Each thread runs 8 independent chains; per loop body: 8 x (FADD2,FMUL2,FFMA2-ish) packed ops plus K extra ops.
// Reports time relative to the pure packed stream => cost of the extra op in "packed-op slots".
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); return 1;}}while(0)
__device__ __forceinline__ uint64_t pk(float a,float b){uint64_t r; asm("mov.b64 %0,{%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(uint64_t v,float&a,float&b){asm("mov.b64 {%0,%1},%2;":"=f"(a),"=f"(b):"l"(v));}
constexpr int ITERS=20000;
// MODE: which extra op; NX: how many extra ops per 8 packed ops
template<int MODE,int NX> __global__ void __launch_bounds__(256,2) kern(float* out, float a, float b, const float* smemsrc){
  __shared__ __align__(16) float sm[1024];
  for(int i=threadIdx.x;i<1024;i+=256) sm[i]=smemsrc[i];
  __syncthreads();
  uint64_t w[8]; float e[8]; int ei[8];
  #pragma unroll
  for(int i=0;i<8;i++){ w[i]=pk(threadIdx.x*0.001f+i,i+0.5f); e[i]=3e38f-i; ei[i]=i; }
  uint64_t A=pk(a,a), B=pk(b,b);
  #pragma unroll 1
  for(int it=0;it<ITERS;it++){
    #pragma unroll
    for(int i=0;i<8;i++){
      // one packed op per chain, rotating the opcode like the real loop (3 add : 3 mul : 2 fma)
      if(i<3) asm volatile("add.rn.f32x2 %0,%0,%1;":"+l"(w[i]):"l"(B));
      else if(i<6) asm volatile("mul.rn.f32x2 %0,%0,%1;":"+l"(w[i]):"l"(A));
      else asm volatile("fma.rn.f32x2 %0,%0,%1,%2;":"+l"(w[i]):"l"(A),"l"(B));
    }
    #pragma unroll
    for(int x=0;x<NX;x++){
      float lo,hi; upk(w[x&7],lo,hi);
      if(MODE==1) asm volatile("min.f32 %0,%0,%1,%2;":"+f"(e[x&7]):"f"(lo),"f"(hi));          // FMNMX3
      if(MODE==2) asm volatile("min.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(lo));                      // FMNMX
      if(MODE==3) asm volatile("min.s32 %0,%0,%1;":"+r"(ei[x&7]):"r"(__float_as_int(lo)));     // IMNMX
      if(MODE==4) asm volatile("{.reg .pred p; setp.lt.f32 p,%1,%2; selp.s32 %0,%3,%0,p;}":"+r"(ei[x&7]):"f"(lo),"f"(e[x&7]),"r"(it)); // FSETP+SEL
      if(MODE==5) asm volatile("lop3.b32 %0,%0,%1,%2,0x96;":"+r"(ei[x&7]):"r"(__float_as_int(lo)),"r"(__float_as_int(hi))); // LOP3
      if(MODE==6) asm volatile("add.s32 %0,%0,%1;":"+r"(ei[x&7]):"r"(__float_as_int(lo)));     // IADD
      if(MODE==7) asm volatile("add.rn.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(lo));                   // scalar FADD (fma pipe)
      if(MODE==8){ float4 v=*reinterpret_cast<const float4*>(&sm[((it+x)&255)*4]); asm volatile("":"+f"(e[x&7]):"f"(v.x),"f"(v.y),"f"(v.z),"f"(v.w)); e[x&7]=v.x; } // LDS.128 (broadcast)
      if(MODE==9) asm volatile("{.reg .s32 t; min.s32 t,%1,%2; min.s32 %0,%0,t;}":"+r"(ei[x&7]):"r"(__float_as_int(lo)),"r"(__float_as_int(hi))); // 2x IMNMX (maybe VIMNMX3)
      if(MODE==10) asm volatile("max.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(lo));  // FMNMX max
      if(MODE==11) asm volatile("min.xorsign.abs.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(lo));  // FMNMX.XORSIGN
      if(MODE==12) asm volatile("min.NaN.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(lo));          // FMNMX.NAN
      if(MODE==13) asm volatile("{.reg .f32 t; neg.f32 t,%1; max.f32 %0,%0,t;}":"+f"(e[x&7]):"f"(lo)); // FMNMX with -src
      if(MODE==14){ // chain on ONE accumulator alternating flavours (cannot merge into FMNMX3): 2 inputs per extra op
        asm volatile("min.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(lo));
        asm volatile("min.NaN.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(hi)); }
      if(MODE==16) asm volatile("min.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(hi));   // FMNMX on the ODD half only
      if(MODE==17){ float lo2,hi2; upk(w[(x+1)&7],lo2,hi2);                     // FMNMX3 on two EVEN halves
        asm volatile("min.f32 %0,%0,%1,%2;":"+f"(e[x&7]):"f"(lo),"f"(lo2)); }
      if(MODE==18){ float lo2,hi2; upk(w[(x+1)&7],lo2,hi2);                     // dependent pair, even halves only
        float t; asm volatile("min.f32 %0,%1,%2;":"=f"(t):"f"(lo),"f"(lo2));
        asm volatile("min.NaN.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(t)); }
      if(MODE==19){ float lo2,hi2; upk(w[(x+1)&7],lo2,hi2);                     // FMNMX3 on two ODD halves
        asm volatile("min.f32 %0,%0,%1,%2;":"+f"(e[x&7]):"f"(hi),"f"(hi2)); }
      if(MODE==15){ // tree: t=min(lo,hi) plain, e=min.NaN(e,t)
        float t; asm volatile("min.f32 %0,%1,%2;":"=f"(t):"f"(lo),"f"(hi));
        asm volatile("min.NaN.f32 %0,%0,%1;":"+f"(e[x&7]):"f"(t)); }
    }
  }
  float s=0; int si=0;
  #pragma unroll
  for(int i=0;i<8;i++){ float lo,hi; upk(w[i],lo,hi); s+=lo+hi+e[i]; si+=ei[i]; }
  out[blockIdx.x*blockDim.x+threadIdx.x]=s+si;
}
static float base_ms=0;
template<int MODE,int NX> int run(const char* name,int nsm){
  float* out; float* src; int nb=nsm*2;
  CK(cudaMalloc(&out,sizeof(float)*nb*256)); CK(cudaMalloc(&src,4096)); CK(cudaMemset(src,0,4096));
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  kern<MODE,NX><<<nb,256>>>(out,1.0000001f,1e-9f,src);
  CK(cudaDeviceSynchronize());
  float best=1e9;
  for(int r=0;r<3;r++){
    cudaEventRecord(e0); kern<MODE,NX><<<nb,256>>>(out,1.0000001f,1e-9f,src); cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms,e0,e1); if(ms<best) best=ms;
  }
  if(MODE==0) base_ms=best;
  double laneops=(double)ITERS*8*2*256*nb; // packed lane-ops
  printf("%-22s extra/8packed=%d  %.3f ms  packed-rate %.2f Tlane-ops/s  rel-time %.3f  => cost per extra op = %.2f packed-op slots\n",
    name,NX,best,laneops/(best*1e-3)/1e12,best/base_ms, NX? (best/base_ms-1.0)*8.0/NX : 0.0);
  cudaFree(out); cudaFree(src); return 0;
}
int main(){
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,0)); int nsm=p.multiProcessorCount;
  // warm the clocks for ~1 s
  for(int i=0;i<30;i++) run<0,0>("warm",nsm);
  printf("(warm-up lines above)\n");
  printf("---- measurements (256 thr x 2 CTA/SM, 8 chains/thread)\n");
  run<0,0>("packed only",nsm);
  run<1,1>("FMNMX3",nsm); run<1,2>("FMNMX3",nsm); run<1,4>("FMNMX3",nsm);
  run<2,1>("FMNMX",nsm);  run<2,2>("FMNMX",nsm);  run<2,4>("FMNMX",nsm);
  run<10,2>("FMNMX(max)",nsm);
  run<3,1>("IMNMX",nsm);  run<3,2>("IMNMX",nsm);  run<3,4>("IMNMX",nsm);
  run<9,1>("2xIMNMX/VIMNMX3",nsm); run<9,2>("2xIMNMX/VIMNMX3",nsm);
  run<4,1>("FSETP+SEL",nsm); run<4,2>("FSETP+SEL",nsm);
  run<5,2>("LOP3",nsm); run<5,4>("LOP3",nsm);
  run<6,2>("IADD",nsm); run<6,4>("IADD",nsm);
  run<7,2>("FADD scalar",nsm); run<7,4>("FADD scalar",nsm);
  run<8,1>("LDS.128",nsm); run<8,2>("LDS.128",nsm);
  run<11,2>("FMNMX.XORSIGN",nsm); run<11,4>("FMNMX.XORSIGN",nsm);
  run<12,2>("FMNMX.NAN",nsm); run<12,4>("FMNMX.NAN",nsm); run<12,8>("FMNMX.NAN",nsm);
  run<2,8>("FMNMX",nsm);
  run<13,2>("FMNMX(-src)",nsm); run<13,4>("FMNMX(-src)",nsm);
  run<14,2>("min+min.NaN chain (2 ops)",nsm); run<14,4>("min+min.NaN chain (2 ops)",nsm);
  run<15,2>("min tree + min.NaN (2 ops)",nsm); run<15,4>("min tree + min.NaN (2 ops)",nsm);
  run<16,2>("FMNMX odd half",nsm); run<16,4>("FMNMX odd half",nsm);
  run<17,2>("FMNMX3 even+even",nsm); run<17,4>("FMNMX3 even+even",nsm);
  run<19,2>("FMNMX3 odd+odd",nsm);
  run<18,2>("dep. pair even only",nsm); run<18,4>("dep. pair even only",nsm);
  run<1,2>("FMNMX3 (again)",nsm);
  run<0,0>("packed only (again)",nsm);
  return 0;
}
