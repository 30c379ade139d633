// Sweep 5 (round 2): memtest kernel shapes INSIDE green-context SM partitions.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mt_sweep5 mt_sweep5.cu -lcuda
//   ./mt_sweep5 <nparts> <percent> <fine 0|1> <GiB per partition>
// For every kernel shape: GB/s with ONE partition busy (and per SM), and the aggregate with ALL partitions busy.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d: %s\n",#x,__LINE__,cudaGetErrorString(e)); exit(1);} }while(0)
#define CU(x) do{CUresult e=(x); if(e!=CUDA_SUCCESS){const char*s=0;cuGetErrorString(e,&s);printf("ERR %s line %d: %d %s\n",#x,__LINE__,(int)e,s?s:"?"); exit(1);} }while(0)
constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;
struct alignas(32) U64x4 { uint64_t a,b,c,d; };
__device__ __forceinline__ void st256(void*p,const U64x4&v){asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};"::"l"(p),"l"(v.a),"l"(v.b),"l"(v.c),"l"(v.d):"memory");}
__device__ __forceinline__ U64x4 ld256(const void*p){U64x4 v;asm volatile("ld.global.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];":"=l"(v.a),"=l"(v.b),"=l"(v.c),"=l"(v.d):"l"(p));return v;}
__device__ __forceinline__ uint64_t splitmix64(uint64_t x){uint64_t z=x+kGolden;z=(z^(z>>30))*0xBF58476D1CE4E5B9ull;z=(z^(z>>27))*0x94D049BB133111EBull;return z^(z>>31);}
template<int PAT> __device__ __forceinline__ uint64_t word(uint64_t i,uint64_t key){ if(PAT==0) return 1ull<<((i+key)&63); return splitmix64(key+i);}

// PERSIST = 0: one tile per CTA (grid = tiles).  PERSIST = 1: grid-stride over tiles with a bounded grid.
template<int PAT,int T,int U,int PERSIST>
__global__ void __launch_bounds__(T) k_fill(uint64_t* buf,uint64_t nvec,uint64_t key){
  const uint64_t ntiles=(nvec+(uint64_t)T*U-1)/((uint64_t)T*U);
  for(uint64_t tile=blockIdx.x; tile<ntiles; tile+=gridDim.x){
    const uint64_t base=tile*(T*U)+threadIdx.x;
    #pragma unroll
    for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec){uint64_t i=4*v;U64x4 w;w.a=word<PAT>(i,key);w.b=word<PAT>(i+1,key);w.c=word<PAT>(i+2,key);w.d=word<PAT>(i+3,key);st256(buf+4*v,w);}}
    if(!PERSIST) break;
  }
}
template<int PAT,int T,int U,int PERSIST>
__global__ void __launch_bounds__(T) k_verify(const uint64_t* buf,uint64_t nvec,uint64_t key,unsigned long long* res){
  const uint64_t ntiles=(nvec+(uint64_t)T*U-1)/((uint64_t)T*U);
  uint64_t acc=0;
  for(uint64_t tile=blockIdx.x; tile<ntiles; tile+=gridDim.x){
    U64x4 w[U];
    const uint64_t base=tile*(T*U)+threadIdx.x;
    #pragma unroll
    for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec) w[u]=ld256(buf+4*v);}
    #pragma unroll
    for(int u=0;u<U;++u){uint64_t v=base+(uint64_t)u*T; if(v<nvec){const uint64_t i=4*v;
      acc|=(w[u].a^word<PAT>(i,key))|(w[u].b^word<PAT>(i+1,key))|(w[u].c^word<PAT>(i+2,key))|(w[u].d^word<PAT>(i+3,key));}}
    if(!PERSIST) break;
  }
  if(acc) atomicAdd(res,1ull);
}
struct Part{CUgreenCtx g; CUstream s; unsigned sms; uint64_t* buf; unsigned long long* res;};
struct Shape{const char*name; int T,U,persist; void(*fill)(Part&,uint64_t,uint64_t,unsigned,int); void(*ver)(Part&,uint64_t,uint64_t,unsigned,int);};
template<int PAT,int T,int U,int P> void run_fill(Part&p,uint64_t nvec,uint64_t key,unsigned cap,int){
  uint64_t tiles=(nvec+(uint64_t)T*U-1)/((uint64_t)T*U); unsigned g=(unsigned)(P? (tiles<cap?tiles:cap):tiles);
  k_fill<PAT,T,U,P><<<g,T,0,(cudaStream_t)p.s>>>(p.buf,nvec,key);}
template<int PAT,int T,int U,int P> void run_ver(Part&p,uint64_t nvec,uint64_t key,unsigned cap,int){
  uint64_t tiles=(nvec+(uint64_t)T*U-1)/((uint64_t)T*U); unsigned g=(unsigned)(P? (tiles<cap?tiles:cap):tiles);
  k_verify<PAT,T,U,P><<<g,T,0,(cudaStream_t)p.s>>>(p.buf,nvec,key,p.res);}

static std::vector<Part> parts;
static cudaEvent_t e0,e1; static std::vector<cudaEvent_t> ends;
// time `fn` on partitions [0,k): best of reps, ms
template<class F> float time_on(int k,F fn,int reps=3){
  float best=1e30f;
  for(int r=0;r<reps+1;++r){
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0,0));
    for(int i=0;i<k;++i) CK(cudaStreamWaitEvent((cudaStream_t)parts[i].s,e0,0));
    for(int i=0;i<k;++i) fn(parts[i]);
    for(int i=0;i<k;++i){CK(cudaEventRecord(ends[i],(cudaStream_t)parts[i].s)); CK(cudaStreamWaitEvent(0,ends[i],0));}
    CK(cudaEventRecord(e1,0)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms,e0,e1)); if(r>0 && ms<best) best=ms;
  }
  CK(cudaGetLastError());
  return best;
}
template<int T,int U,int P> void row(uint64_t bytes,int ctas_per_sm){
  const uint64_t nvec=bytes/32; const int n=(int)parts.size(); const unsigned sms=parts[0].sms;
  const unsigned cap=sms*(unsigned)ctas_per_sm;
  auto g=[&](float ms,int k){return k*(double)bytes/(ms*1e6);};
  float t[4][2];
  for(int k_i=0;k_i<2;++k_i){ int k=k_i?n:1;
    t[0][k_i]=time_on(k,[&](Part&p){run_fill<0,T,U,P>(p,nvec,3,cap,0);});
    t[1][k_i]=time_on(k,[&](Part&p){run_ver<0,T,U,P>(p,nvec,3,cap,0);});
    t[2][k_i]=time_on(k,[&](Part&p){run_fill<1,T,U,P>(p,nvec,kGolden^3,cap,0);});
    t[3][k_i]=time_on(k,[&](Part&p){run_ver<1,T,U,P>(p,nvec,kGolden^3,cap,0);});
  }
  char shape[64]; if(P) snprintf(shape,sizeof shape,"%4dx%d persist %2d/SM",T,U,ctas_per_sm); else snprintf(shape,sizeof shape,"%4dx%d tile/CTA      ",T,U);
  double step1=4.0/(1/g(t[0][0],1)+1/g(t[1][0],1)+1/g(t[2][0],1)+1/g(t[3][0],1));
  double stepn=4.0/(1/g(t[0][1],n)+1/g(t[1][1],n)+1/g(t[2][1],n)+1/g(t[3][1],n));
  printf("%s | one part (GB/s per SM): fw %5.1f vw %5.1f fr %5.1f vr %5.1f step %5.1f | all %d parts (GB/s): fw %6.0f vw %6.0f fr %6.0f vr %6.0f step %6.0f\n",
    shape,g(t[0][0],1)/sms,g(t[1][0],1)/sms,g(t[2][0],1)/sms,g(t[3][0],1)/sms,step1/sms,n,g(t[0][1],n),g(t[1][1],n),g(t[2][1],n),g(t[3][1],n),stepn);
  fflush(stdout);
}
int main(int argc,char**argv){
  int n= argc>1? atoi(argv[1]):10; int pct= argc>2? atoi(argv[2]):10; int fine= argc>3? atoi(argv[3]):1; double gib= argc>4? atof(argv[4]):4.0;
  uint64_t bytes=(uint64_t)(gib*(1ull<<30));
  CK(cudaSetDevice(0)); CK(cudaFree(0));
  CUdevice dev; CU(cuDeviceGet(&dev,0));
  CUdevResource all,rest,groups[64]; CU(cuDeviceGetDevResource(dev,&all,CU_DEV_RESOURCE_TYPE_SM));
  unsigned total=all.sm.smCount, want=total*pct/100, nb=0;
  for(;;--want){ nb=n; CUresult r=cuDevSmResourceSplitByCount(groups,&nb,&all,&rest,fine?CU_DEV_SM_RESOURCE_SPLIT_IGNORE_SM_COSCHEDULING:0,want);
    if(r==CUDA_SUCCESS && nb>=(unsigned)n && (groups[0].sm.smCount<=total*pct/100||want<=1)) break; if(want<=1){printf("cannot split\n");return 1;} }
  parts.resize(n); ends.resize(n);
  for(int i=0;i<n;++i){ CUdevResourceDesc d; CU(cuDevResourceGenerateDesc(&d,&groups[i],1)); CU(cuGreenCtxCreate(&parts[i].g,d,dev,CU_GREEN_CTX_DEFAULT_STREAM));
    CU(cuGreenCtxStreamCreate(&parts[i].s,parts[i].g,CU_STREAM_NON_BLOCKING,0)); parts[i].sms=groups[i].sm.smCount;
    CK(cudaMalloc(&parts[i].buf,bytes)); CK(cudaMalloc(&parts[i].res,8)); CK(cudaMemset(parts[i].res,0,8)); CK(cudaEventCreateWithFlags(&ends[i],cudaEventDisableTiming)); }
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  printf("# %d partitions x %u SMs (of %u, %s), %.1f GiB each\n",n,parts[0].sms,total,fine?"SM-granular":"8-SM granule",gib);
  row<128,4,0>(bytes,0); row<256,4,0>(bytes,0); row<256,2,0>(bytes,0); row<512,2,0>(bytes,0); row<512,4,0>(bytes,0); row<1024,2,0>(bytes,0); row<256,8,0>(bytes,0); row<128,8,0>(bytes,0);
  row<256,4,1>(bytes,8); row<256,4,1>(bytes,16); row<256,4,1>(bytes,32); row<512,4,1>(bytes,4); row<1024,2,1>(bytes,2); row<1024,4,1>(bytes,2); row<128,4,1>(bytes,16); row<128,4,1>(bytes,64);
  unsigned long long bad=0,b; for(auto&p:parts){CK(cudaMemcpy(&b,p.res,8,cudaMemcpyDeviceToHost)); bad+=b;} printf("mismatch flags raised during timing by cross-pattern verifies are expected only for the first call: %llu\n",bad);
  return 0;
}
