// Microbenchmark sweep for the memtest kernels (not part of the product).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d: %s\n",#x,__LINE__,cudaGetErrorString(e)); exit(1);} }while(0)
constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;
struct alignas(32) U64x4 { uint64_t a,b,c,d; };
__device__ __forceinline__ void st256_na(void*p,const U64x4&v){asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};"::"l"(p),"l"(v.a),"l"(v.b),"l"(v.c),"l"(v.d):"memory");}
__device__ __forceinline__ void st256_def(void*p,const U64x4&v){asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};"::"l"(p),"l"(v.a),"l"(v.b),"l"(v.c),"l"(v.d):"memory");}
__device__ __forceinline__ void st256_cs(void*p,const U64x4&v){asm volatile("st.global.cs.v4.u64 [%0], {%1,%2,%3,%4};"::"l"(p),"l"(v.a),"l"(v.b),"l"(v.c),"l"(v.d):"memory");}
__device__ __forceinline__ void st256_wt(void*p,const U64x4&v){asm volatile("st.global.wt.v4.u64 [%0], {%1,%2,%3,%4};"::"l"(p),"l"(v.a),"l"(v.b),"l"(v.c),"l"(v.d):"memory");}
__device__ __forceinline__ void st128x2(void*p,const U64x4&v){asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1,%2};"::"l"(p),"l"(v.a),"l"(v.b):"memory");asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1,%2};"::"l"((char*)p+16),"l"(v.c),"l"(v.d):"memory");}
__device__ __forceinline__ U64x4 ld256(const void*p){U64x4 v;asm volatile("ld.global.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];":"=l"(v.a),"=l"(v.b),"=l"(v.c),"=l"(v.d):"l"(p));return v;}
__device__ __forceinline__ uint64_t splitmix64(uint64_t x){uint64_t z=x+kGolden;z=(z^(z>>30))*0xBF58476D1CE4E5B9ull;z=(z^(z>>27))*0x94D049BB133111EBull;return z^(z>>31);}

// ---- optimized splitmix: 32-bit halves, hi shifts on the fma pipe (IMAD.HI with hidden constants)
struct Hid { uint32_t m30, m27, m31; };  // 1<<2, 1<<5, 1<<1 : umulhi(hi, m) == hi >> s
__device__ __forceinline__ void mul64c(uint32_t &lo, uint32_t &hi, uint32_t clo, uint32_t chi){
  uint64_t p=(uint64_t)lo*clo; uint32_t ph=(uint32_t)(p>>32); ph=lo*chi+ph; ph=hi*clo+ph; lo=(uint32_t)p; hi=ph; }
template<bool HID>
__device__ __forceinline__ void sm_core(uint64_t z0, const Hid&h, uint32_t &lo, uint32_t &hi, uint32_t &slo, uint32_t &shi){
  lo=(uint32_t)z0; hi=(uint32_t)(z0>>32);
  uint32_t t=__funnelshift_r(lo,hi,30); uint32_t u= HID? __umulhi(hi,h.m30) : (hi>>30); lo^=t; hi^=u;
  mul64c(lo,hi,0x1CE4E5B9u,0xBF58476Du);
  t=__funnelshift_r(lo,hi,27); u= HID? __umulhi(hi,h.m27):(hi>>27); lo^=t; hi^=u;
  mul64c(lo,hi,0x133111EBu,0x94D049BBu);
  slo=__funnelshift_r(lo,hi,31); shi= HID? __umulhi(hi,h.m31):(hi>>31);
}
template<bool HID> __device__ __forceinline__ uint64_t sm_opt(uint64_t x,const Hid&h){uint32_t lo,hi,slo,shi; sm_core<HID>(x+kGolden,h,lo,hi,slo,shi); return ((uint64_t)(hi^shi)<<32)|(lo^slo);}
// returns got ^ expected folded (nonzero iff mismatch) as 32-bit OR of both halves
template<bool HID> __device__ __forceinline__ uint32_t sm_diff(uint64_t x,const Hid&h,uint64_t got){uint32_t lo,hi,slo,shi; sm_core<HID>(x+kGolden,h,lo,hi,slo,shi); return (lo^slo^(uint32_t)got)|(hi^shi^(uint32_t)(got>>32));}

template<int PAT,int GEN> __device__ __forceinline__ uint64_t word(uint64_t i,uint64_t key,const Hid&h){
  if(PAT==0) return 1ull<<((i+key)&63);
  if(GEN==0) return splitmix64(key+i);
  if(GEN==1) return sm_opt<false>(key+i,h);
  return sm_opt<true>(key+i,h);
}
// LAYOUT 0: thread-strided (current). LAYOUT 1: CTA-contiguous tiles of 256*32*U bytes, tiles grid-strided
template<int PAT,int GEN,int ST,int LAYOUT,int U>
__global__ void __launch_bounds__(256) k_fill(uint64_t* buf,uint64_t nvec,uint64_t key,Hid h){
  if(LAYOUT==0){
    const uint64_t tid=(uint64_t)blockIdx.x*256+threadIdx.x, nth=(uint64_t)gridDim.x*256;
    for(uint64_t v0=tid;v0<nvec;v0+=nth*U){
      #pragma unroll
      for(int u=0;u<U;++u){uint64_t v=v0+(uint64_t)u*nth; if(v<nvec){uint64_t i=4*v;U64x4 w;w.a=word<PAT,GEN>(i,key,h);w.b=word<PAT,GEN>(i+1,key,h);w.c=word<PAT,GEN>(i+2,key,h);w.d=word<PAT,GEN>(i+3,key,h);
        if(ST==0)st256_na(buf+4*v,w);else if(ST==1)st256_def(buf+4*v,w);else if(ST==2)st256_cs(buf+4*v,w);else if(ST==3)st256_wt(buf+4*v,w);else st128x2(buf+4*v,w);}}
    }
  } else {
    const uint64_t tilev=256ull*U; const uint64_t ntiles=(nvec+tilev-1)/tilev;
    for(uint64_t t=blockIdx.x;t<ntiles;t+=gridDim.x){
      #pragma unroll
      for(int u=0;u<U;++u){uint64_t v=t*tilev+(uint64_t)u*256+threadIdx.x; if(v<nvec){uint64_t i=4*v;U64x4 w;w.a=word<PAT,GEN>(i,key,h);w.b=word<PAT,GEN>(i+1,key,h);w.c=word<PAT,GEN>(i+2,key,h);w.d=word<PAT,GEN>(i+3,key,h);
        if(ST==0)st256_na(buf+4*v,w);else if(ST==1)st256_def(buf+4*v,w);else if(ST==2)st256_cs(buf+4*v,w);else if(ST==3)st256_wt(buf+4*v,w);else st128x2(buf+4*v,w);}}
    }
  }
}
// verify: GEN 0 baseline compare; 1 opt no-hid; 2 opt hid; folded diff accumulate
template<int PAT,int GEN,int LAYOUT,int U>
__global__ void __launch_bounds__(256) k_verify(const uint64_t* buf,uint64_t nvec,uint64_t key,Hid h,unsigned long long* res){
  unsigned long long bad=0;
  auto body=[&](uint64_t v,const U64x4&w){
    const uint64_t i=4*v;
    if(PAT==1 && GEN>0){
      uint32_t d = GEN==1 ? (sm_diff<false>(key+i,h,w.a)|sm_diff<false>(key+i+1,h,w.b)|sm_diff<false>(key+i+2,h,w.c)|sm_diff<false>(key+i+3,h,w.d))
                          : (sm_diff<true>(key+i,h,w.a)|sm_diff<true>(key+i+1,h,w.b)|sm_diff<true>(key+i+2,h,w.c)|sm_diff<true>(key+i+3,h,w.d));
      if(d) bad+= (w.a!=splitmix64(key+i))+(w.b!=splitmix64(key+i+1))+(w.c!=splitmix64(key+i+2))+(w.d!=splitmix64(key+i+3));
    } else {
      uint64_t e0=word<PAT,0>(i,key,h),e1=word<PAT,0>(i+1,key,h),e2=word<PAT,0>(i+2,key,h),e3=word<PAT,0>(i+3,key,h);
      if((w.a^e0)|(w.b^e1)|(w.c^e2)|(w.d^e3)) bad+=(w.a!=e0)+(w.b!=e1)+(w.c!=e2)+(w.d!=e3);
    }
  };
  if(LAYOUT==0){
    const uint64_t tid=(uint64_t)blockIdx.x*256+threadIdx.x, nth=(uint64_t)gridDim.x*256;
    for(uint64_t v0=tid;v0<nvec;v0+=nth*U){ U64x4 w[U];
      #pragma unroll
      for(int u=0;u<U;++u){uint64_t v=v0+(uint64_t)u*nth; if(v<nvec) w[u]=ld256(buf+4*v);}
      #pragma unroll
      for(int u=0;u<U;++u){uint64_t v=v0+(uint64_t)u*nth; if(v<nvec) body(v,w[u]);}
    }
  } else {
    const uint64_t tilev=256ull*U; const uint64_t ntiles=(nvec+tilev-1)/tilev;
    for(uint64_t t=blockIdx.x;t<ntiles;t+=gridDim.x){ U64x4 w[U];
      #pragma unroll
      for(int u=0;u<U;++u){uint64_t v=t*tilev+(uint64_t)u*256+threadIdx.x; if(v<nvec) w[u]=ld256(buf+4*v);}
      #pragma unroll
      for(int u=0;u<U;++u){uint64_t v=t*tilev+(uint64_t)u*256+threadIdx.x; if(v<nvec) body(v,w[u]);}
    }
  }
  if(bad) atomicAdd(res,bad);
}

static float time_it(void(*fn)(int,void*),int grid,void*ctx,int reps=5){
  cudaEvent_t a,b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); fn(grid,ctx); fn(grid,ctx); float best=1e30f;
  for(int r=0;r<reps;++r){CK(cudaEventRecord(a)); fn(grid,ctx); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms,a,b)); if(ms<best)best=ms;}
  CK(cudaGetLastError()); return best;
}
struct Ctx{uint64_t*buf;uint64_t nvec;uint64_t key;Hid h;unsigned long long*res;};
#define FILLFN(name,P,G,S,L,U) static void name(int g,void*c){Ctx*x=(Ctx*)c;k_fill<P,G,S,L,U><<<g,256>>>(x->buf,x->nvec,x->key,x->h);}
#define VERFN(name,P,G,L,U) static void name(int g,void*c){Ctx*x=(Ctx*)c;k_verify<P,G,L,U><<<g,256>>>(x->buf,x->nvec,x->key,x->h,x->res);}
FILLFN(f_w_na_l0,0,0,0,0,4) FILLFN(f_w_def_l0,0,0,1,0,4) FILLFN(f_w_cs_l0,0,0,2,0,4) FILLFN(f_w_wt_l0,0,0,3,0,4) FILLFN(f_w_128_l0,0,0,4,0,4)
FILLFN(f_w_na_l1,0,0,0,1,4) FILLFN(f_w_def_l1,0,0,1,1,4) FILLFN(f_w_cs_l1,0,0,2,1,4) FILLFN(f_w_na_l1u8,0,0,0,1,8) FILLFN(f_w_na_l1u2,0,0,0,1,2) FILLFN(f_w_na_l1u1,0,0,0,1,1) FILLFN(f_w_na_l0u1,0,0,0,0,1)
FILLFN(f_r_g0,1,0,0,0,4) FILLFN(f_r_g1,1,1,0,0,4) FILLFN(f_r_g2,1,2,0,0,4) FILLFN(f_r_g2_l1,1,2,0,1,4)
VERFN(v_w_l0,0,0,0,4) VERFN(v_w_l1,0,0,1,4) VERFN(v_w_l0u8,0,0,0,8) VERFN(v_w_l0u2,0,0,0,2)
VERFN(v_r_g0,1,0,0,4) VERFN(v_r_g1,1,1,0,4) VERFN(v_r_g2,1,2,0,4) VERFN(v_r_g2_l1,1,2,1,4) VERFN(v_r_g2_u2,1,2,0,2) VERFN(v_r_g2_u8,1,2,0,8)
int main(int argc,char**argv){
  double gib= argc>1? atof(argv[1]):8.0; uint64_t bytes=(uint64_t)(gib*(1ull<<30)); uint64_t nvec=bytes/32;
  Ctx c; CK(cudaMalloc(&c.buf,bytes)); CK(cudaMalloc(&c.res,8)); CK(cudaMemset(c.res,0,8)); c.nvec=nvec; c.key=kGolden^5; c.h=Hid{4u,32u,2u};
  int sms; CK(cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0));
  // correctness of optimized generators: fill with g2, verify with g0 and vice versa
  f_r_g2(sms*8,&c); v_r_g0(sms*8,&c); f_r_g0(sms*8,&c); v_r_g2(sms*8,&c); v_r_g1(sms*8,&c); f_r_g1(sms*8,&c); v_r_g0(sms*8,&c);
  unsigned long long bad=1; CK(cudaMemcpy(&bad,c.res,8,cudaMemcpyDeviceToHost)); printf("cross-check mismatches: %llu (must be 0)\n",bad);
  c.key=kGolden^6; v_r_g2(sms*8,&c); CK(cudaMemcpy(&bad,c.res,8,cudaMemcpyDeviceToHost)); printf("wrong-key mismatches: %llu (must be %llu)\n",bad,(unsigned long long)nvec*4); c.key=kGolden^5; CK(cudaMemset(c.res,0,8));
  struct E{const char*n;void(*f)(int,void*);}; 
  std::vector<E> fills={{"fill walk NA  L0 U4",f_w_na_l0},{"fill walk def L0 U4",f_w_def_l0},{"fill walk cs  L0 U4",f_w_cs_l0},{"fill walk wt  L0 U4",f_w_wt_l0},{"fill walk 128 L0 U4",f_w_128_l0},
    {"fill walk NA  L1 U4",f_w_na_l1},{"fill walk def L1 U4",f_w_def_l1},{"fill walk cs  L1 U4",f_w_cs_l1},{"fill walk NA  L1 U8",f_w_na_l1u8},{"fill walk NA  L1 U2",f_w_na_l1u2},{"fill walk NA  L1 U1",f_w_na_l1u1},{"fill walk NA  L0 U1",f_w_na_l0u1},
    {"fill rand gen0",f_r_g0},{"fill rand gen1",f_r_g1},{"fill rand gen2(hid)",f_r_g2},{"fill rand gen2 L1",f_r_g2_l1}};
  std::vector<E> vers={{"ver walk L0 U4",v_w_l0},{"ver walk L1 U4",v_w_l1},{"ver walk L0 U8",v_w_l0u8},{"ver walk L0 U2",v_w_l0u2},{"ver rand gen0",v_r_g0},{"ver rand gen1",v_r_g1},{"ver rand gen2(hid)",v_r_g2},{"ver rand gen2 L1",v_r_g2_l1},{"ver rand gen2 U2",v_r_g2_u2},{"ver rand gen2 U8",v_r_g2_u8}};
  int grids[]={sms*1,sms*2,sms*4,sms*8,sms*16,32*8};
  printf("%-24s","kernel \\ grid"); for(int g:grids)printf("%9d",g); printf("   (GB/s)\n");
  for(auto&e:fills){printf("%-24s",e.n); for(int g:grids){float ms=time_it(e.f,g,&c); printf("%9.0f",bytes/(ms*1e6));} printf("\n");}
  // make buffer valid for walk verify then rand verify
  f_w_na_l0(sms*8,&c); CK(cudaDeviceSynchronize());
  for(auto&e:vers){ if(e.n[4]=='r'&&e.n[5]=='a'&&e.f==v_r_g0){f_r_g0(sms*8,&c);CK(cudaDeviceSynchronize());}
    printf("%-24s",e.n); for(int g:grids){float ms=time_it(e.f,g,&c); printf("%9.0f",bytes/(ms*1e6));} printf("\n");}
  CK(cudaMemcpy(&bad,c.res,8,cudaMemcpyDeviceToHost)); printf("mismatches during timing: %llu (must be 0)\n",bad);
  return 0;
}
