// Micro-benchmark: FP32 pipe issue rates on sm_100a (scalar vs packed f32x2, FMNMX3 mix).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_fp32 ubench_fp32.cu
// Prints lane-ops per clock per SM for each instruction stream (128 = full FP32 rate).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); return 1;}}while(0)
__device__ __forceinline__ uint64_t pk(float a,float b){uint64_t r; asm("mov.b64 %0,{%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(uint64_t v,float&a,float&b){asm("mov.b64 {%0,%1},%2;":"=f"(a),"=f"(b):"l"(v));}
constexpr int NCH=16;   // independent chains per thread
constexpr int ITERS=4096;
template<int MODE> __global__ void __launch_bounds__(256) kern(float* out, float a, float b, long long* cyc){
  float v[NCH]; uint64_t w[NCH];
  #pragma unroll
  for(int i=0;i<NCH;i++){ v[i]=threadIdx.x*0.001f+i; w[i]=pk(v[i],v[i]+0.5f);}
  uint64_t A=pk(a,a), B=pk(b,b);
  float mn=3e38f;
  long long t0=clock64();
  #pragma unroll 1
  for(int it=0;it<ITERS;it++){
    #pragma unroll
    for(int i=0;i<NCH;i++){
      if(MODE==0) asm volatile("fma.rn.f32 %0,%0,%1,%2;":"+f"(v[i]):"f"(a),"f"(b));
      if(MODE==1) asm volatile("add.rn.f32 %0,%0,%1;":"+f"(v[i]):"f"(b));
      if(MODE==2) asm volatile("mul.rn.f32 %0,%0,%1;":"+f"(v[i]):"f"(a));
      if(MODE==3) asm volatile("fma.rn.f32x2 %0,%0,%1,%2;":"+l"(w[i]):"l"(A),"l"(B));
      if(MODE==4) asm volatile("add.rn.f32x2 %0,%0,%1;":"+l"(w[i]):"l"(B));
      if(MODE==5) asm volatile("mul.rn.f32x2 %0,%0,%1;":"+l"(w[i]):"l"(A));
      if(MODE==6){ // packed + min3 every packed op (issue mix of the real loop: 8 FP2 : 1 FMNMX3)
        asm volatile("fma.rn.f32x2 %0,%0,%1,%2;":"+l"(w[i]):"l"(A),"l"(B));
        if((i&7)==7){ float lo,hi; upk(w[i],lo,hi); asm volatile("min.f32 %0,%0,%1,%2;":"+f"(mn):"f"(lo),"f"(hi)); }
      }
      if(MODE==7){ // scalar + min every 8
        asm volatile("fma.rn.f32 %0,%0,%1,%2;":"+f"(v[i]):"f"(a),"f"(b));
        if((i&7)==7){ asm volatile("min.f32 %0,%0,%1;":"+f"(mn):"f"(v[i])); }
      }
      if(MODE==8){ // packed, 2 FP2 : 1 FMNMX3  (heavier ALU mix)
        asm volatile("fma.rn.f32x2 %0,%0,%1,%2;":"+l"(w[i]):"l"(A),"l"(B));
        if((i&1)==1){ float lo,hi; upk(w[i],lo,hi); asm volatile("min.f32 %0,%0,%1,%2;":"+f"(mn):"f"(lo),"f"(hi)); }
      }
      if(MODE==9){ // packed 1:1 with FMNMX3
        asm volatile("fma.rn.f32x2 %0,%0,%1,%2;":"+l"(w[i]):"l"(A),"l"(B));
        { float lo,hi; upk(w[i],lo,hi); asm volatile("min.f32 %0,%0,%1,%2;":"+f"(mn):"f"(lo),"f"(hi)); }
      }
    }
  }
  long long t1=clock64();
  float s=mn;
  #pragma unroll
  for(int i=0;i<NCH;i++){ float lo,hi; upk(w[i],lo,hi); s+=v[i]+lo+hi; }
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
  if(threadIdx.x==0) cyc[blockIdx.x]=t1-t0;
}
template<int MODE> int run(const char* name,int lanes_per_op,int ctas_per_sm,int nsm){
  float* out; long long* cyc; int nb=nsm*ctas_per_sm;
  CK(cudaMalloc(&out,sizeof(float)*nb*256)); CK(cudaMalloc(&cyc,sizeof(long long)*nb));
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for(int w=0;w<2;w++) kern<MODE><<<nb,256>>>(out,1.0000001f,1e-9f,cyc);
  CK(cudaDeviceSynchronize());
  cudaEventRecord(e0); kern<MODE><<<nb,256>>>(out,1.0000001f,1e-9f,cyc); cudaEventRecord(e1);
  CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms,e0,e1);
  long long* h=(long long*)malloc(sizeof(long long)*nb); cudaMemcpy(h,cyc,sizeof(long long)*nb,cudaMemcpyDeviceToHost);
  double avg=0; for(int i=0;i<nb;i++) avg+=h[i]; avg/=nb;
  double ops=(double)ITERS*NCH*256*ctas_per_sm*lanes_per_op; // lane-ops per SM
  printf("%-28s ctas/sm=%d  lane-ops/clk/SM=%7.2f  (cycles %.0f, %.3f ms, eff clk %.0f MHz, chip %.2f Tlane-ops/s)\n",
    name,ctas_per_sm,ops/avg,avg,ms,avg/ms/1e3,ops*nsm/(ms*1e-3)/1e12);
  cudaFree(out); cudaFree(cyc); free(h); return 0;
}
int main(){
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,0));
  printf("device %s  SMs %d  clock %d kHz  smem/SM %zu\n",p.name,p.multiProcessorCount,p.clockRate,p.sharedMemPerMultiprocessor);
  int nsm=p.multiProcessorCount;
  for(int c=1;c<=4;c*=2){
    run<0>("FFMA scalar",1,c,nsm); run<1>("FADD scalar",1,c,nsm); run<2>("FMUL scalar",1,c,nsm);
    run<3>("FFMA2 packed",2,c,nsm); run<4>("FADD2 packed",2,c,nsm); run<5>("FMUL2 packed",2,c,nsm);
    run<6>("FFMA2 + FMNMX3 (8:1)",2,c,nsm); run<7>("FFMA + FMNMX (8:1)",1,c,nsm);
    run<8>("FFMA2 + FMNMX3 (2:1)",2,c,nsm); run<9>("FFMA2 + FMNMX3 (1:1)",2,c,nsm);
  }
  return 0;
}
