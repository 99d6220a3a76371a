#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e){printf("err %s line %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)
int main(){
  const int S=256, H=480, W=640, step=4, h=H/step;
  for (int elem : {1, 4}) {
    size_t row = (size_t)W*elem, full=(size_t)S*H*row, part=(size_t)S*h*row;
    void *hp, *dp; CK(cudaMallocHost(&hp, full)); CK(cudaMalloc(&dp, full));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int mode=0; mode<4; mode++) {
      float best=1e9;
      for (int it=0; it<6; it++) {
        cudaEventRecord(a, st);
        if (mode==0) CK(cudaMemcpy2DAsync(dp,row,hp,row*step,row,(size_t)S*h,cudaMemcpyHostToDevice,st));
        if (mode==1) CK(cudaMemcpyAsync(dp,hp,part,cudaMemcpyHostToDevice,st));
        if (mode==2) CK(cudaMemcpyAsync(dp,hp,full,cudaMemcpyHostToDevice,st));
        if (mode==3) { // one 2D copy per stream image: 256 calls of 120 rows
          for (int s=0;s<S;s++) CK(cudaMemcpy2DAsync((char*)dp+(size_t)s*h*row,row,(char*)hp+(size_t)s*H*row,row*step,row,h,cudaMemcpyHostToDevice,st));
        }
        cudaEventRecord(b, st); CK(cudaStreamSynchronize(st));
        float ms; cudaEventElapsedTime(&ms,a,b); if (ms<best) best=ms;
      }
      size_t bytes = mode==2?full:part;
      printf("elem %d mode %d (%s): %.3f ms  %.1f GB/s\n", elem, mode, mode==0?"2D rows, one call":mode==1?"contiguous, same bytes":mode==2?"contiguous, full frames":"2D rows, per image", best, bytes/best/1e6);
    }
    cudaFreeHost(hp); cudaFree(dp);
  }
  return 0;
}
