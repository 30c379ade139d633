// Sweep 2: grid/block shape for fill (write-only) and verify (read-only); not part of the product.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d: %s\n",#x,__LINE__,cudaGetErrorString(e)); exit(1);} }while(0)
constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;
struct alignas(32) U64x4 { uint64_t a,b,c,d; };
__device__ __forceinline__ void st256(void*p,const U64x4&v){asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};"::"l"(p),"l"(v.a),"l"(v.b),"l"(v.c),"l"(v.d):"memory");}
__device__ __forceinline__ U64x4 ld256(const void*p){U64x4 v;asm volatile("ld.global.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];":"=l"(v.a),"=l"(v.b),"=l"(v.c),"=l"(v.d):"l"(p));return v;}
__device__ __forceinline__ uint64_t splitmix64(uint64_t x){uint64_t z=x+kGolden;z=(z^(z>>30))*0xBF58476D1CE4E5B9ull;z=(z^(z>>27))*0x94D049BB133111EBull;return z^(z>>31);}
template<int PAT> __device__ __forceinline__ uint64_t word(uint64_t i,uint64_t key){ if(PAT==0) return 1ull<<((i+key)&63); return splitmix64(key+i);}
template<int PAT,int U,int MINB>
__global__ void __launch_bounds__(1024,MINB) k_fill(uint64_t* buf,uint64_t nvec,uint64_t key){
  const uint64_t tid=(uint64_t)blockIdx.x*blockDim.x+threadIdx.x, nth=(uint64_t)gridDim.x*blockDim.x;
  for(uint64_t v0=tid;v0<nvec;v0+=nth*U){
    #pragma unroll
    for(int u=0;u<U;++u){uint64_t v=v0+(uint64_t)u*nth; if(v<nvec){uint64_t i=4*v;U64x4 w;w.a=word<PAT>(i,key);w.b=word<PAT>(i+1,key);w.c=word<PAT>(i+2,key);w.d=word<PAT>(i+3,key);st256(buf+4*v,w);}}
  }
}
template<int PAT,int U,int MINB>
__global__ void __launch_bounds__(1024,MINB) k_verify(const uint64_t* buf,uint64_t nvec,uint64_t key,unsigned long long* res){
  unsigned long long bad=0;
  const uint64_t tid=(uint64_t)blockIdx.x*blockDim.x+threadIdx.x, nth=(uint64_t)gridDim.x*blockDim.x;
  for(uint64_t v0=tid;v0<nvec;v0+=nth*U){ U64x4 w[U];
    #pragma unroll
    for(int u=0;u<U;++u){uint64_t v=v0+(uint64_t)u*nth; if(v<nvec) w[u]=ld256(buf+4*v);}
    #pragma unroll
    for(int u=0;u<U;++u){uint64_t v=v0+(uint64_t)u*nth; if(v<nvec){const uint64_t i=4*v; uint64_t e0=word<PAT>(i,key),e1=word<PAT>(i+1,key),e2=word<PAT>(i+2,key),e3=word<PAT>(i+3,key);
      if((w[u].a^e0)|(w[u].b^e1)|(w[u].c^e2)|(w[u].d^e3)) bad+=(w[u].a!=e0)+(w[u].b!=e1)+(w[u].c!=e2)+(w[u].d!=e3);}}
  }
  if(bad) atomicAdd(res,bad);
}
struct Ctx{uint64_t*buf;uint64_t nvec;uint64_t key;unsigned long long*res;int block;};
typedef void(*Fn)(int,Ctx*);
static float time_it(Fn fn,int grid,Ctx*c,int reps=4){
  cudaEvent_t a,b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); fn(grid,c); float best=1e30f;
  for(int r=0;r<reps;++r){CK(cudaEventRecord(a)); fn(grid,c); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms,a,b)); if(ms<best)best=ms;}
  CK(cudaGetLastError()); CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b)); return best;
}
#define F(name,P,U,M) static void name(int g,Ctx*x){k_fill<P,U,M><<<g,x->block>>>(x->buf,x->nvec,x->key);}
#define V(name,P,U,M) static void name(int g,Ctx*x){k_verify<P,U,M><<<g,x->block>>>(x->buf,x->nvec,x->key,x->res);}
F(fw1,0,1,1) F(fw4,0,4,1) F(fr4,1,4,1) F(fr2,1,2,1) F(fr1,1,1,1)
V(vw4,0,4,1) V(vw8,0,8,1) V(vr4,1,4,1) V(vr4m2,1,4,2) V(vr2,1,2,1) V(vr2m2,1,2,2) V(vw8m2,0,8,2)
int main(int argc,char**argv){
  double gib= argc>1? atof(argv[1]):8.0; uint64_t bytes=(uint64_t)(gib*(1ull<<30)); uint64_t nvec=bytes/32;
  Ctx c; CK(cudaMalloc(&c.buf,bytes)); CK(cudaMalloc(&c.res,8)); CK(cudaMemset(c.res,0,8)); c.nvec=nvec; c.key=3;
  int sms; CK(cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0));
  int mult[]={4,8,12,16,20,24,32,48,64,128,512};
  int blocks[]={128,256,512,1024};
  struct E{const char*n;Fn f;int pat;}; 
  E fills[]={{"fill walk U1",fw1,0},{"fill walk U4",fw4,0},{"fill rand U4",fr4,1},{"fill rand U2",fr2,1},{"fill rand U1",fr1,1}};
  E vers[]={{"ver walk U4",vw4,0},{"ver walk U8",vw8,0},{"ver walk U8 minb2",vw8m2,0},{"ver rand U4",vr4,1},{"ver rand U4 minb2",vr4m2,1},{"ver rand U2",vr2,1},{"ver rand U2 minb2",vr2m2,1}};
  printf("%-26s","kernel/block \\ CTAs per SM"); for(int m:mult)printf("%7d",m); printf("   (GB/s)\n");
  for(auto&e:fills) for(int b:blocks){ c.block=b; c.key= e.pat? (kGolden^3):3; printf("%-20s b%-5d",e.n,b); for(int m:mult){ float ms=time_it(e.f,sms*m,&c); printf("%7.0f",bytes/(ms*1e6)); } printf("\n"); }
  for(auto&e:vers) for(int b:blocks){ c.block=b; c.key= e.pat? (kGolden^3):3; (e.pat? fr4:fw4)(sms*8,&c); CK(cudaDeviceSynchronize());
    printf("%-20s b%-5d",e.n,b); for(int m:mult){ float ms=time_it(e.f,sms*m,&c); printf("%7.0f",bytes/(ms*1e6)); } printf("\n"); }
  unsigned long long bad=1; CK(cudaMemcpy(&bad,c.res,8,cudaMemcpyDeviceToHost)); printf("mismatches during timing: %llu (must be 0)\n",bad);
  return 0;
}
