#include <cuda_runtime.h>
#include <sys/mman.h>
#include <chrono>
#include <cstdio>
#include <cstring>
static double now(){return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();}
int main(){
  cudaFree(0);
  size_t n = 8ull<<30;
  void*p; double t0=now(); cudaError_t e=cudaHostAlloc(&p,n,cudaHostAllocPortable|cudaHostAllocMapped); double t1=now();
  printf("cudaHostAlloc 8 GiB: %.2f s (%s) -> %.2f GiB/s\n",t1-t0,cudaGetErrorString(e),8/(t1-t0)); cudaFreeHost(p);
  for (int huge=0; huge<2; ++huge){
    t0=now(); void*m=mmap(nullptr,n,PROT_READ|PROT_WRITE,MAP_PRIVATE|MAP_ANONYMOUS,-1,0);
    if(huge) madvise(m,n,MADV_HUGEPAGE);
    memset(m,0,n); double t2=now();
    e=cudaHostRegister(m,n,cudaHostRegisterPortable|cudaHostRegisterMapped); double t3=now();
    printf("mmap%s + touch: %.2f s, cudaHostRegister: %.2f s (%s) -> total %.2f GiB/s\n",huge?"+MADV_HUGEPAGE":"",t2-t0,t3-t2,cudaGetErrorString(e),8/(t3-t0));
    if(e==cudaSuccess) cudaHostUnregister(m); munmap(m,n);
  }
  FILE*f=fopen("/sys/kernel/mm/transparent_hugepage/enabled","r"); char b[128]={0}; if(f){fgets(b,127,f);fclose(f);} printf("THP: %s",b);
  return 0;
}
