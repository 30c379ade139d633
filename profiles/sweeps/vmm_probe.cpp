#include <cuda.h>
#include <chrono>
#include <cstdio>
#include <vector>
#define CK(x) do{CUresult r=(x); if(r!=CUDA_SUCCESS){const char*s; cuGetErrorString(r,&s); printf("ERR %s: %s\n",#x,s); return 1;} }while(0)
static double now(){return std::chrono::duration<double,std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();}
int main(){
  CK(cuInit(0)); CUdevice d; CK(cuDeviceGet(&d,0)); CUcontext c; CK(cuDevicePrimaryCtxRetain(&c,d)); CK(cuCtxSetCurrent(c));
  CUmemAllocationProp prop{}; prop.type=CU_MEM_ALLOCATION_TYPE_PINNED; prop.location.type=CU_MEM_LOCATION_TYPE_DEVICE; prop.location.id=0;
  CUmemAccessDesc acc{}; acc.location=prop.location; acc.flags=CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  size_t total = 8ull<<30;
  for (size_t sz : {2ull<<20, 16ull<<20, 64ull<<20, 256ull<<20, 1024ull<<20}) {
    size_t n = total/sz; CUdeviceptr base; CK(cuMemAddressReserve(&base,total,2<<20,0,0));
    std::vector<CUmemGenericAllocationHandle> h(n);
    double t0=now(); for(size_t i=0;i<n;++i) CK(cuMemCreate(&h[i],sz,&prop,0));
    double t1=now(); for(size_t i=0;i<n;++i) CK(cuMemMap(base+i*sz,sz,0,h[i],0));
    double t2=now(); for(size_t i=0;i<n;++i) CK(cuMemSetAccess(base+i*sz,sz,&acc,1));
    double t3=now(); for(size_t i=0;i<n;++i) CK(cuMemUnmap(base+i*sz,sz));
    double t4=now(); for(size_t i=0;i<n;++i) CK(cuMemMap(base+i*sz,sz,0,h[i],0));
    double t5=now(); CK(cuMemSetAccess(base,total,&acc,1));
    double t6=now(); CK(cuMemUnmap(base,total));
    double t7=now(); for(size_t i=0;i<n;++i) CK(cuMemRelease(h[i]));
    double t8=now(); CK(cuMemAddressFree(base,total));
    printf("handle %5zu MiB x %5zu: create %8.1f us/op  map %6.1f  setaccess %8.1f  unmap %7.1f  remap %6.1f  setaccess(all) %9.1f us total  unmap(all) %9.1f  release %7.1f us/op | per-GiB ms: create %.2f map+access %.2f unmap %.2f\n",
      sz>>20,n,(t1-t0)/n,(t2-t1)/n,(t3-t2)/n,(t4-t3)/n,(t5-t4)/n,(t6-t5),(t7-t6),(t8-t7)/n, (t1-t0)/8e3,(t3-t1)/8e3,(t4-t3)/8e3);
  }
  return 0;
}
