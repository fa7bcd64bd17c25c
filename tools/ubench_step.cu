// Micro-benchmark 3: the symmetric scan's 4-target step (16 packed squared distances, 32 FMNMX3, filter compare + vote)
// with the target coordinates delivered (A) from shared memory by LDS.128 into vector registers, as the kernel does,
// or (B) from the constant bank by LDCU into UNIFORM registers (FADD2 R, -Rq.F32, UR.F32x2: one register-file read
// instead of three).  Answers: how much of the 21 % gap to the FP32 issue peak is register-file operand traffic?
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); return 1;}}while(0)
__device__ __forceinline__ uint64_t pk2(float a,float b){uint64_t r; asm("mov.b64 %0,{%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk2(uint64_t v,float&a,float&b){asm("mov.b64 {%0,%1},%2;":"=f"(a),"=f"(b):"l"(v));}
__device__ __forceinline__ uint64_t add2(uint64_t a,uint64_t b){uint64_t r; asm("add.rn.f32x2 %0,%1,%2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ uint64_t mul2(uint64_t a,uint64_t b){uint64_t r; asm("mul.rn.f32x2 %0,%1,%2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ uint64_t fma2(uint64_t a,uint64_t b,uint64_t c){uint64_t r; asm("fma.rn.f32x2 %0,%1,%2,%3;":"=l"(r):"l"(a),"l"(b),"l"(c)); return r;}
__device__ __forceinline__ float fmin3(float a,float b,float c){float r; asm("min.f32 %0,%1,%2,%3;":"=f"(r):"f"(a),"f"(b),"f"(c)); return r;}
__device__ __forceinline__ uint64_t sqdist2(uint64_t TX,uint64_t TY,uint64_t TZ,uint64_t nqx,uint64_t nqy,uint64_t nqz,uint64_t one){
  const uint64_t dx=add2(TX,nqx),dy=add2(TY,nqy),dz=add2(TZ,nqz);
  const uint64_t xx=mul2(dx,dx),yy=mul2(dy,dy),zz=mul2(dz,dz);
  return fma2(fma2(xx,one,yy),one,zz);
}
constexpr int TT=512, QPT=8, GROUPS=TT/4;
__constant__ ulonglong2 cT[3*GROUPS*8];  // 8 tiles of 512 targets: [tile][group][x|y|z] as 2 packed pairs = 48 KB
template<int SRC> __global__ void __launch_bounds__(256,2) kern(float* out,const float* q,const ulonglong2* gT,float onef,int iters){
  __shared__ ulonglong2 sT[3*GROUPS];
  __shared__ __align__(16) unsigned filt[TT];
  for(int i=threadIdx.x;i<3*GROUPS;i+=256) sT[i]=gT[i];
  for(int i=threadIdx.x;i<TT;i+=256) filt[i]=0u;   // never passes: distances are > 0
  __syncthreads();
  uint64_t nqx[QPT],nqy[QPT],nqz[QPT]; float cm[QPT];
  #pragma unroll
  for(int k=0;k<QPT;k++){ const float* s=q+((blockIdx.x*256+threadIdx.x)*QPT+k)*3; float x=-s[0],y=-s[1],z=-s[2];
    nqx[k]=pk2(x,x); nqy[k]=pk2(y,y); nqz[k]=pk2(z,z); cm[k]=__int_as_float(0x7f800000); }
  const uint64_t one=pk2(onef,onef);
  const float INF=__int_as_float(0x7f800000);
  unsigned hits=0;
  #pragma unroll 1
  for(int it=0;it<iters;it++){
    const ulonglong2* ct=cT+(it&7)*3*GROUPS;
    #pragma unroll 8
    for(int g=0;g<GROUPS;g++){
      ulonglong2 X,Y,Z;
      if(SRC==0){ X=sT[g*3]; Y=sT[g*3+1]; Z=sT[g*3+2]; } else { X=ct[g*3]; Y=ct[g*3+1]; Z=ct[g*3+2]; }
      float c0=INF,c1=INF,c2=INF,c3=INF;
      #pragma unroll
      for(int k=0;k<QPT;k+=2){
        float a0,a1,b0,b1,e0,e1,f0,f1;
        upk2(sqdist2(X.x,Y.x,Z.x,nqx[k],nqy[k],nqz[k],one),a0,a1);
        upk2(sqdist2(X.y,Y.y,Z.y,nqx[k],nqy[k],nqz[k],one),b0,b1);
        upk2(sqdist2(X.x,Y.x,Z.x,nqx[k+1],nqy[k+1],nqz[k+1],one),e0,e1);
        upk2(sqdist2(X.y,Y.y,Z.y,nqx[k+1],nqy[k+1],nqz[k+1],one),f0,f1);
        cm[k]=fmin3(cm[k],a0,a1); cm[k]=fmin3(cm[k],b0,b1);
        cm[k+1]=fmin3(cm[k+1],e0,e1); cm[k+1]=fmin3(cm[k+1],f0,f1);
        c0=fmin3(c0,a0,e0); c1=fmin3(c1,a1,e1); c2=fmin3(c2,b0,f0); c3=fmin3(c3,b1,f1);
      }
      const uint4 H=*reinterpret_cast<const uint4*>(filt+4*g);
      const bool ev=(__float_as_uint(c0)<=H.x)|(__float_as_uint(c1)<=H.y)|(__float_as_uint(c2)<=H.z)|(__float_as_uint(c3)<=H.w);
      if(__any_sync(0xffffffffu,ev)) hits++;
    }
  }
  float s=0;
  #pragma unroll
  for(int k=0;k<QPT;k++) s+=cm[k];
  out[blockIdx.x*256+threadIdx.x]=s+hits;
}
template<int SRC> int run(const char* name,int nsm,float* out,const float* q,const ulonglong2* gT,int iters){
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int nb=nsm*2;
  kern<SRC><<<nb,256>>>(out,q,gT,1.0f,iters); CK(cudaDeviceSynchronize());
  float best=1e9;
  for(int r=0;r<3;r++){ cudaEventRecord(e0); kern<SRC><<<nb,256>>>(out,q,gT,1.0f,iters); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms,e0,e1); if(ms<best) best=ms; }
  double evals=(double)nb*256*QPT*TT*iters;
  printf("%-44s %8.3f ms  %.2f Tevals/s  %.2f Tlane-ops/s = %.1f %% of 37.22\n",name,best,evals/(best*1e-3)/1e12,evals*8/(best*1e-3)/1e12,evals*8/(best*1e-3)/1e12/37.22*100);
  return 0;
}
int main(){
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,0)); int nsm=p.multiProcessorCount; int nb=nsm*2;
  float *out,*q; ulonglong2* gT; 
  CK(cudaMalloc(&out,4*nb*256)); CK(cudaMalloc(&q,4*nb*256*QPT*3)); CK(cudaMalloc(&gT,sizeof(ulonglong2)*3*GROUPS*8));
  float* h=(float*)malloc(4*nb*256*QPT*3); for(int i=0;i<nb*256*QPT*3;i++) h[i]=(i*2654435761u%1000003)/1000003.0f;
  CK(cudaMemcpy(q,h,4*nb*256*QPT*3,cudaMemcpyHostToDevice));
  float* ht=(float*)malloc(sizeof(ulonglong2)*3*GROUPS*8); for(int i=0;i<3*GROUPS*8*4;i++) ht[i]=2.0f+(i*40503u%65521)/65521.0f;
  CK(cudaMemcpy(gT,ht,sizeof(ulonglong2)*3*GROUPS*8,cudaMemcpyHostToDevice));
  CK(cudaMemcpyToSymbol(cT,ht,sizeof(ulonglong2)*3*GROUPS*8));
  for(int i=0;i<3;i++){ run<0>("warm",nsm,out,q,gT,600); }
  printf("---- 4-target step, 256 thr x 8 rows, 2 CTA/SM\n");
  for(int r=0;r<2;r++){
    run<0>("targets: LDS.128 -> vector registers",nsm,out,q,gT,600);
    run<1>("targets: LDCU -> uniform registers",nsm,out,q,gT,600);
  }
  return 0;
}
