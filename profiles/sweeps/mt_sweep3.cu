// Sweep 3: non-persistent, CTA-contiguous tiles (threads x U vectors), all four memtest kernels.
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
template<int PAT,int T,int U>
__global__ void __launch_bounds__(T) k_fill(uint64_t* buf,uint64_t nvec,uint64_t key){
  const uint64_t base=(uint64_t)blockIdx.x*(T*U)+threadIdx.x;
  #pragma unroll
  for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec){uint64_t i=4*v;U64x4 w;w.a=word<PAT>(i,key);w.b=word<PAT>(i+1,key);w.c=word<PAT>(i+2,key);w.d=word<PAT>(i+3,key);st256(buf+4*v,w);}}
}
template<int PAT,int T,int U>
__global__ void __launch_bounds__(T) k_verify(const uint64_t* buf,uint64_t nvec,uint64_t key,unsigned long long* res){
  unsigned long long bad=0; U64x4 w[U];
  const uint64_t base=(uint64_t)blockIdx.x*(T*U)+threadIdx.x;
  #pragma unroll
  for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec) w[u]=ld256(buf+4*v);}
  #pragma unroll
  for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec){const uint64_t i=4*v; uint64_t e0=word<PAT>(i,key),e1=word<PAT>(i+1,key),e2=word<PAT>(i+2,key),e3=word<PAT>(i+3,key);
    if((w[u].a^e0)|(w[u].b^e1)|(w[u].c^e2)|(w[u].d^e3)) bad+=(w[u].a!=e0)+(w[u].b!=e1)+(w[u].c!=e2)+(w[u].d!=e3);}}
  if(bad) atomicAdd(res,bad);
}
struct Ctx{uint64_t*buf;uint64_t nvec;uint64_t key;unsigned long long*res;};
template<int PAT,int T,int U> void run_fill(Ctx*c){ unsigned g=(unsigned)((c->nvec+T*U-1)/(T*U)); k_fill<PAT,T,U><<<g,T>>>(c->buf,c->nvec,c->key);}
template<int PAT,int T,int U> void run_ver(Ctx*c){ unsigned g=(unsigned)((c->nvec+T*U-1)/(T*U)); k_verify<PAT,T,U><<<g,T>>>(c->buf,c->nvec,c->key,c->res);}
typedef void(*Fn)(Ctx*);
static float time_it(Fn fn,Ctx*c,int reps=5){
  cudaEvent_t a,b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); fn(c); float best=1e30f;
  for(int r=0;r<reps;++r){CK(cudaEventRecord(a)); fn(c); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms,a,b)); if(ms<best)best=ms;}
  CK(cudaGetLastError()); CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b)); return best;
}
template<int T,int U> void row(Ctx*c,uint64_t bytes){
  c->key=3; float fw=time_it(run_fill<0,T,U>,c); float vw=time_it(run_ver<0,T,U>,c);
  c->key=kGolden^3; float fr=time_it(run_fill<1,T,U>,c); float vr=time_it(run_ver<1,T,U>,c);
  auto g=[&](float ms){return bytes/(ms*1e6);};
  printf("T=%4d U=%d tile=%6d B | fill_walk %6.0f  ver_walk %6.0f  fill_rand %6.0f  ver_rand %6.0f | pass walk %6.0f rand %6.0f\n",T,U,T*U*32,g(fw),g(vw),g(fr),g(vr),2*bytes/((fw+vw)*1e6),2*bytes/((fr+vr)*1e6));
}
int main(int argc,char**argv){
  double gib= argc>1? atof(argv[1]):8.0; uint64_t bytes=(uint64_t)(gib*(1ull<<30)); uint64_t nvec=bytes/32;
  Ctx c; CK(cudaMalloc(&c.buf,bytes)); CK(cudaMalloc(&c.res,8)); CK(cudaMemset(c.res,0,8)); c.nvec=nvec;
  row<128,1>(&c,bytes); row<128,2>(&c,bytes); row<128,4>(&c,bytes); row<128,8>(&c,bytes);
  row<256,1>(&c,bytes); row<256,2>(&c,bytes); row<256,4>(&c,bytes); row<256,8>(&c,bytes);
  row<512,1>(&c,bytes); row<512,2>(&c,bytes); row<512,4>(&c,bytes);
  row<1024,1>(&c,bytes); row<1024,2>(&c,bytes); row<1024,4>(&c,bytes);
  unsigned long long bad=1; CK(cudaMemcpy(&bad,c.res,8,cudaMemcpyDeviceToHost)); printf("mismatches during timing: %llu (must be 0)\n",bad);
  return 0;
}
