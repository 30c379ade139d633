// Sweep 4: ALU diet for the random-pattern verify (and fill), non-persistent T x U tiles.
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
struct Hid { uint32_t m30, m27, m31; };
__device__ __forceinline__ void mul64c(uint32_t &lo, uint32_t &hi, uint32_t clo, uint32_t chi){
  uint64_t p=(uint64_t)lo*clo; uint32_t ph=(uint32_t)(p>>32); ph=lo*chi+ph; ph=hi*clo+ph; lo=(uint32_t)p; hi=ph; }
// MODE 0: plain 64-bit code. MODE 1: halves + folded compare, ALU shifts. MODE 2: + hi shifts via IMAD.HI (hidden consts).
// MODE 3: + lo funnel shift via IMAD.WIDE+IMAD too (all shifts on the fma pipe)
template<int MODE>
__device__ __forceinline__ void sm_core(uint64_t z0, const Hid&h, uint32_t &lo, uint32_t &hi, uint32_t &slo, uint32_t &shi){
  lo=(uint32_t)z0; hi=(uint32_t)(z0>>32); uint32_t t,u;
  auto sh=[&](int s,uint32_t m,uint32_t&tl,uint32_t&th){
    if(MODE==3){ uint64_t p=(uint64_t)lo*m; tl=(uint32_t)(p>>32); tl=hi*m+tl; th=__umulhi(hi,m);} 
    else { tl=__funnelshift_r(lo,hi,s); th= MODE>=2? __umulhi(hi,m):(hi>>s);} };
  sh(30,h.m30,t,u); lo^=t; hi^=u; mul64c(lo,hi,0x1CE4E5B9u,0xBF58476Du);
  sh(27,h.m27,t,u); lo^=t; hi^=u; mul64c(lo,hi,0x133111EBu,0x94D049BBu);
  sh(31,h.m31,slo,shi);
}
template<int MODE> __device__ __forceinline__ uint32_t diff(uint64_t z0,const Hid&h,uint64_t got,uint32_t acc){
  if(MODE==0){ uint64_t e=splitmix64(z0-kGolden); uint64_t d=e^got; return acc|(uint32_t)d|(uint32_t)(d>>32);} 
  uint32_t lo,hi,slo,shi; sm_core<MODE>(z0,h,lo,hi,slo,shi); uint32_t d0=lo^slo^(uint32_t)got, d1=hi^shi^(uint32_t)(got>>32); return acc|d0|d1; }
template<int MODE> __device__ __forceinline__ uint64_t gen(uint64_t z0,const Hid&h){
  if(MODE==0) return splitmix64(z0-kGolden);
  uint32_t lo,hi,slo,shi; sm_core<MODE>(z0,h,lo,hi,slo,shi); return ((uint64_t)(hi^shi)<<32)|(lo^slo); }

__device__ __noinline__ void slow_count(const uint64_t*buf,uint64_t nvec,uint64_t key,uint64_t base,int T,int U,unsigned long long*res){
  unsigned long long bad=0; for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec) for(int k=0;k<4;++k) bad+= buf[4*v+k]!=splitmix64(key+4*v+k);} if(bad) atomicAdd(res,bad); }

template<int MODE,int T,int U>
__global__ void __launch_bounds__(T) k_verify(const uint64_t* buf,uint64_t nvec,uint64_t key,Hid h,unsigned long long* res){
  U64x4 w[U]; const uint64_t base=(uint64_t)blockIdx.x*(T*U)+threadIdx.x; uint32_t acc=0; const uint64_t kz=key+kGolden;
  #pragma unroll
  for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec) w[u]=ld256(buf+4*v); else {w[u].a=gen<0>(kz+4*v,h);w[u].b=gen<0>(kz+4*v+1,h);w[u].c=gen<0>(kz+4*v+2,h);w[u].d=gen<0>(kz+4*v+3,h);} }
  #pragma unroll
  for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; const uint64_t z=kz+4*v;
    acc=diff<MODE>(z,h,w[u].a,acc); acc=diff<MODE>(z+1,h,w[u].b,acc); acc=diff<MODE>(z+2,h,w[u].c,acc); acc=diff<MODE>(z+3,h,w[u].d,acc); }
  if(acc) slow_count(buf,nvec,key,base,T,U,res);
}
template<int MODE,int T,int U>
__global__ void __launch_bounds__(T) k_fill(uint64_t* buf,uint64_t nvec,uint64_t key,Hid h){
  const uint64_t base=(uint64_t)blockIdx.x*(T*U)+threadIdx.x; const uint64_t kz=key+kGolden;
  #pragma unroll
  for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec){const uint64_t z=kz+4*v;U64x4 w;w.a=gen<MODE>(z,h);w.b=gen<MODE>(z+1,h);w.c=gen<MODE>(z+2,h);w.d=gen<MODE>(z+3,h);st256(buf+4*v,w);}}
}
struct Ctx{uint64_t*buf;uint64_t nvec;uint64_t key;Hid h;unsigned long long*res;};
template<int M,int T,int U> void run_fill(Ctx*c){ unsigned g=(unsigned)((c->nvec+T*U-1)/(T*U)); k_fill<M,T,U><<<g,T>>>(c->buf,c->nvec,c->key,c->h);}
template<int M,int T,int U> void run_ver(Ctx*c){ unsigned g=(unsigned)((c->nvec+T*U-1)/(T*U)); k_verify<M,T,U><<<g,T>>>(c->buf,c->nvec,c->key,c->h,c->res);}
typedef void(*Fn)(Ctx*);
static float time_it(Fn fn,Ctx*c,int reps=5){
  cudaEvent_t a,b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); fn(c); float best=1e30f;
  for(int r=0;r<reps;++r){CK(cudaEventRecord(a)); fn(c); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms,a,b)); if(ms<best)best=ms;}
  CK(cudaGetLastError()); CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b)); return best;
}
template<int M,int T,int U> void row(Ctx*c,uint64_t bytes){
  float fr=time_it(run_fill<M,T,U>,c); float vr=time_it(run_ver<M,T,U>,c);
  printf("MODE=%d T=%4d U=%d | fill_rand %6.0f  ver_rand %6.0f GB/s\n",M,T,U,bytes/(fr*1e6),bytes/(vr*1e6)); }
int main(int argc,char**argv){
  double gib= argc>1? atof(argv[1]):8.0; uint64_t bytes=(uint64_t)(gib*(1ull<<30)); uint64_t nvec=bytes/32;
  Ctx c; CK(cudaMalloc(&c.buf,bytes)); CK(cudaMalloc(&c.res,8)); CK(cudaMemset(c.res,0,8)); c.nvec=nvec; c.key=kGolden^3; c.h=Hid{4u,32u,2u};
  // cross-check: every fill mode verified by mode 0 and every verify mode on a mode-0 fill
  run_fill<1,128,4>(&c); run_ver<0,128,4>(&c); run_fill<2,128,4>(&c); run_ver<0,128,4>(&c); run_fill<3,128,4>(&c); run_ver<0,128,4>(&c);
  run_fill<0,128,4>(&c); run_ver<1,128,4>(&c); run_ver<2,128,4>(&c); run_ver<3,128,4>(&c);
  unsigned long long bad=1; CK(cudaMemcpy(&bad,c.res,8,cudaMemcpyDeviceToHost)); printf("cross-check mismatches: %llu (must be 0)\n",bad);
  c.key^=1; run_ver<2,128,4>(&c); CK(cudaMemcpy(&bad,c.res,8,cudaMemcpyDeviceToHost)); printf("wrong-key mismatches: %llu (must be %llu)\n",bad,(unsigned long long)nvec*4); c.key^=1; CK(cudaMemset(c.res,0,8));
  row<0,128,4>(&c,bytes); row<1,128,4>(&c,bytes); row<2,128,4>(&c,bytes); row<3,128,4>(&c,bytes);
  row<0,256,4>(&c,bytes); row<1,256,4>(&c,bytes); row<2,256,4>(&c,bytes); row<3,256,4>(&c,bytes);
  row<0,128,2>(&c,bytes); row<1,128,2>(&c,bytes); row<2,128,2>(&c,bytes); row<3,128,2>(&c,bytes);
  row<1,128,8>(&c,bytes); row<2,128,8>(&c,bytes); row<2,64,4>(&c,bytes); row<2,64,8>(&c,bytes);
  CK(cudaMemcpy(&bad,c.res,8,cudaMemcpyDeviceToHost)); printf("mismatches during timing: %llu (must be 0)\n",bad);
  return 0;
}
