#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#ifndef XC
#define XC 16
#endif
#ifndef L2P
#define L2P CU_TENSOR_MAP_L2_PROMOTION_NONE
#endif
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ void waitbar(uint64_t* bar){ uint32_t done; do { asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p; }" : "=r"(done) : "r"(smem_u32(bar)), "r"(0) : "memory"); } while(!done); }
// V0: barrier only
__global__ void k0(uint8_t* out){ __shared__ __align__(8) uint64_t bar; if(threadIdx.x==0){ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1)); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar)) : "memory"); } __syncthreads(); waitbar(&bar); out[threadIdx.x]=1; }
// V1: 1-D bulk copy (no tensor map)
__global__ void k1(const uint8_t* src, uint8_t* out){ __shared__ __align__(128) uint8_t sm[1024]; __shared__ __align__(8) uint64_t bar; if(threadIdx.x==0){ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1)); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1024) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sm)), "l"(src), "r"(1024), "r"(smem_u32(&bar)) : "memory"); } __syncthreads(); waitbar(&bar); for(int i=threadIdx.x;i<1024;i+=blockDim.x) out[i]=sm[i]; }
// V2: tensor 2D, grid_constant
__global__ void k2(const __grid_constant__ CUtensorMap tmap, uint8_t* out){ __shared__ __align__(128) uint8_t sm[1024]; __shared__ __align__(8) uint64_t bar; if(threadIdx.x==0){ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1)); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1024) : "memory");
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(sm)), "l"(&tmap), "r"(XC), "r"(8), "r"(smem_u32(&bar)) : "memory"); } __syncthreads(); waitbar(&bar); for(int i=threadIdx.x;i<1024;i+=blockDim.x) out[i]=sm[i]; }
// V3: tensor 2D, descriptor in global memory
__global__ void k3(const CUtensorMap* tmap, uint8_t* out){ __shared__ __align__(128) uint8_t sm[1024]; __shared__ __align__(8) uint64_t bar; if(threadIdx.x==0){ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1)); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1024) : "memory");
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(sm)), "l"(tmap), "r"(XC), "r"(8), "r"(smem_u32(&bar)) : "memory"); } __syncthreads(); waitbar(&bar); for(int i=threadIdx.x;i<1024;i+=blockDim.x) out[i]=sm[i]; }
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc,char**argv){ int v=atoi(argv[1]);
  void* fn=nullptr; cudaDriverEntryPointQueryResult q; cudaGetDriverEntryPoint("cuTensorMapEncodeTiled",&fn,cudaEnableDefault,&q); Enc enc=(Enc)fn;
  int W=640,H=480,pitch=640; uint8_t* d; cudaMalloc(&d,(size_t)pitch*H); static uint8_t h[640*480]; for(size_t i=0;i<sizeof(h);i++) h[i]=(uint8_t)(i*7+i/640); cudaMemcpy(d,h,sizeof(h),cudaMemcpyHostToDevice);
  uint8_t* out; cudaMalloc(&out,65536); static uint8_t ho[65536];
  CUtensorMap t; cuuint64_t dims[2]={(cuuint64_t)W,(cuuint64_t)H}; cuuint64_t str[1]={(cuuint64_t)pitch}; cuuint32_t box[2]={64,16}; cuuint32_t es[2]={1,1};
  int dt = (v>=10)? (int)CU_TENSOR_MAP_DATA_TYPE_UINT8 : (int)CU_TENSOR_MAP_DATA_TYPE_UINT8;
  CUresult r=enc(&t,(CUtensorMapDataType)dt,2,d,dims,str,box,es,CU_TENSOR_MAP_INTERLEAVE_NONE,CU_TENSOR_MAP_SWIZZLE_NONE,L2P,CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CUtensorMap* dt_; cudaMalloc(&dt_,sizeof(t)); cudaMemcpy(dt_,&t,sizeof(t),cudaMemcpyHostToDevice);
  if(v==0) k0<<<1,128>>>(out); if(v==1) k1<<<1,128>>>(d,out); if(v==2) k2<<<1,128>>>(t,out); if(v==3) k3<<<1,128>>>(dt_,out);
  cudaError_t e=cudaDeviceSynchronize(); int bad=-1;
  if(e==cudaSuccess && v>=2){ bad=0; cudaMemcpy(ho,out,1024,cudaMemcpyDeviceToHost); for(int r2=0;r2<16;r2++) for(int c=0;c<64;c++) if(ho[r2*64+c]!=h[(8+r2)*pitch+XC+c]) bad++; }
  printf("variant %d enc=%d run=%s bad=%d\n",v,(int)r,cudaGetErrorString(e),bad); }
