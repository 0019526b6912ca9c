#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define LD(k) (c - (int)p[DY[k]*pitch+DX[k]])
__host__ __device__ inline int vA(const uint8_t* p,int pitch){
    const int DX[16]={0,1,2,3,3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1}, DY[16]={3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1,0,1,2,3};
    int c=p[0]; int d[16];
#pragma unroll
    for(int k=0;k<16;k++) d[k]=LD(k);
    int best=0;
#pragma unroll
    for(int k=0;k<16;k++){ int mn=d[k],mx=d[k];
#pragma unroll
      for(int j=1;j<9;j++){int v=d[(k+j)&15]; mn=min(mn,v); mx=max(mx,v);} best=max(best,mn); best=max(best,-mx);}
    return best; }
// B: no negation: separate negated array
__host__ __device__ inline int vB(const uint8_t* p,int pitch){
    const int DX[16]={0,1,2,3,3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1}, DY[16]={3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1,0,1,2,3};
    int c=p[0]; int d[16], e[16];
#pragma unroll
    for(int k=0;k<16;k++){ d[k]=LD(k); e[k]=(int)p[DY[k]*pitch+DX[k]]-c; }
    int best=0;
#pragma unroll
    for(int k=0;k<16;k++){ int mn=d[k],me=e[k];
#pragma unroll
      for(int j=1;j<9;j++){ mn=min(mn,d[(k+j)&15]); me=min(me,e[(k+j)&15]);} best=max(best,mn); best=max(best,me);}
    return best; }
// C: returns only the bright part and dark part separately packed
__host__ __device__ inline int vC(const uint8_t* p,int pitch){
    const int DX[16]={0,1,2,3,3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1}, DY[16]={3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1,0,1,2,3};
    int c=p[0]; int d[16];
#pragma unroll
    for(int k=0;k<16;k++) d[k]=LD(k);
    int bmn=-1000, bmx=1000;
#pragma unroll
    for(int k=0;k<16;k++){ int mn=d[k],mx=d[k];
#pragma unroll
      for(int j=1;j<9;j++){int v=d[(k+j)&15]; mn=min(mn,v); mx=max(mx,v);} bmn=max(bmn,mn); bmx=min(bmx,mx);}
    return (bmn+1000)*4096+(bmx+1000); }
__global__ void kern(const uint8_t* imgs,int* out,int n){int t=blockIdx.x*blockDim.x+threadIdx.x; if(t<n){ out[3*t]=vA(imgs+t*49+24,7); out[3*t+1]=vB(imgs+t*49+24,7); out[3*t+2]=vC(imgs+t*49+24,7);} }
int main(){
  const int N=20000; static uint8_t imgs[N*49]; unsigned s=12345;
  for(int t=0;t<N;t++){ uint8_t* img=imgs+t*49; for(int i=0;i<49;i++){ s=s*1664525u+1013904223u; img[i]=(s>>24);} if(t%2){for(int i=0;i<49;i++) img[i]=img[i]/4+(i%7<3?150:20);} }
  uint8_t* d; int* o; cudaMalloc(&d,N*49); cudaMalloc(&o,N*12); cudaMemcpy(d,imgs,N*49,cudaMemcpyHostToDevice);
  kern<<<(N+255)/256,256>>>(d,o,N); static int out[N*3]; cudaMemcpy(out,o,N*12,cudaMemcpyDeviceToHost);
  int ba=0,bb=0,bc=0; for(int t=0;t<N;t++){ const uint8_t* p=imgs+t*49+24; int a=vA(p,7),b=vB(p,7),c=vC(p,7);
    if(a!=out[3*t]) ba++; if(b!=out[3*t+1]) bb++; if(c!=out[3*t+2]){ if(bc<4) printf("C t=%d host mn %d mx %d dev mn %d mx %d\n",t,c/4096-1000,c%4096-1000,out[3*t+2]/4096-1000,out[3*t+2]%4096-1000); bc++;} }
  printf("badA=%d badB=%d badC=%d\n",ba,bb,bc); }
