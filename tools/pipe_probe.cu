#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "fp.cuh"
using namespace ctg;
#define ITER 4096
template<int V> __global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t x, uint32_t y) {
  uint32_t a0=threadIdx.x, a1=a0+1, a2=a0+2, a3=a0+3, a4=a0+4, a5=a0+5, a6=a0+6, a7=a0+7;
  uint64_t w0=a0,w1=a1,w2=a2,w3=a3;
  x += threadIdx.x; y ^= blockIdx.x;
  #pragma unroll 1
  for (int i = 0; i < ITER; i++) {
    if (V == 0) { // IMAD lo, 8 independent
      asm volatile("mad.lo.u32 %0,%0,%8,%9; mad.lo.u32 %1,%1,%8,%9; mad.lo.u32 %2,%2,%8,%9; mad.lo.u32 %3,%3,%8,%9;"
                   "mad.lo.u32 %4,%4,%8,%9; mad.lo.u32 %5,%5,%8,%9; mad.lo.u32 %6,%6,%8,%9; mad.lo.u32 %7,%7,%8,%9;"
        : "+r"(a0),"+r"(a1),"+r"(a2),"+r"(a3),"+r"(a4),"+r"(a5),"+r"(a6),"+r"(a7) : "r"(x),"r"(y));
    } else if (V == 1) { // IMAD.HI
      asm volatile("mad.hi.u32 %0,%0,%8,%9; mad.hi.u32 %1,%1,%8,%9; mad.hi.u32 %2,%2,%8,%9; mad.hi.u32 %3,%3,%8,%9;"
                   "mad.hi.u32 %4,%4,%8,%9; mad.hi.u32 %5,%5,%8,%9; mad.hi.u32 %6,%6,%8,%9; mad.hi.u32 %7,%7,%8,%9;"
        : "+r"(a0),"+r"(a1),"+r"(a2),"+r"(a3),"+r"(a4),"+r"(a5),"+r"(a6),"+r"(a7) : "r"(x),"r"(y));
    } else if (V == 2) { // IMAD.WIDE, 8 (4 accumulators x2)
      asm volatile("{ .reg .u32 t0,t1,t2,t3; cvt.u32.u64 t0,%0; cvt.u32.u64 t1,%1; cvt.u32.u64 t2,%2; cvt.u32.u64 t3,%3;"
                   "mad.wide.u32 %0,t0,%4,%0; mad.wide.u32 %1,t1,%4,%1; mad.wide.u32 %2,t2,%4,%2; mad.wide.u32 %3,t3,%4,%3;"
                   "cvt.u32.u64 t0,%0; cvt.u32.u64 t1,%1; cvt.u32.u64 t2,%2; cvt.u32.u64 t3,%3;"
                   "mad.wide.u32 %0,t0,%5,%0; mad.wide.u32 %1,t1,%5,%1; mad.wide.u32 %2,t2,%5,%2; mad.wide.u32 %3,t3,%5,%3; }"
        : "+l"(w0),"+l"(w1),"+l"(w2),"+l"(w3) : "r"(x),"r"(y));
    } else if (V == 3) { // carry chain lo/hi pairs: 4 MACs (8 PTX instr)
      asm volatile("mad.lo.cc.u32 %0,%1,%8,%0; madc.hi.cc.u32 %1,%1,%8,%1; madc.lo.cc.u32 %2,%3,%9,%2; madc.hi.cc.u32 %3,%3,%9,%3;"
                   "madc.lo.cc.u32 %4,%5,%8,%4; madc.hi.cc.u32 %5,%5,%8,%5; madc.lo.cc.u32 %6,%7,%9,%6; madc.hi.u32 %7,%7,%9,%7;"
        : "+r"(a0),"+r"(a1),"+r"(a2),"+r"(a3),"+r"(a4),"+r"(a5),"+r"(a6),"+r"(a7) : "r"(x),"r"(y));
    } else if (V == 4) { // IADD3 x8
      asm volatile("add.u32 %0,%0,%8; add.u32 %1,%1,%9; add.u32 %2,%2,%8; add.u32 %3,%3,%9;"
                   "add.u32 %4,%4,%8; add.u32 %5,%5,%9; add.u32 %6,%6,%8; add.u32 %7,%7,%9;"
        : "+r"(a0),"+r"(a1),"+r"(a2),"+r"(a3),"+r"(a4),"+r"(a5),"+r"(a6),"+r"(a7) : "r"(x),"r"(y));
    } else if (V == 5) { // 4 IMAD + 4 LOP (xor) interleaved: dual pipe
      asm volatile("mad.lo.u32 %0,%0,%8,%9; xor.b32 %1,%1,%8; mad.lo.u32 %2,%2,%8,%9; xor.b32 %3,%3,%9;"
                   "mad.lo.u32 %4,%4,%8,%9; xor.b32 %5,%5,%8; mad.lo.u32 %6,%6,%8,%9; xor.b32 %7,%7,%9;"
        : "+r"(a0),"+r"(a1),"+r"(a2),"+r"(a3),"+r"(a4),"+r"(a5),"+r"(a6),"+r"(a7) : "r"(x),"r"(y));
    } else if (V == 6) { // 4 IMAD.WIDE + 4 xor
      asm volatile("{ .reg .u32 t0,t1,t2,t3; cvt.u32.u64 t0,%0; cvt.u32.u64 t1,%1; cvt.u32.u64 t2,%2; cvt.u32.u64 t3,%3;"
                   "mad.wide.u32 %0,t0,%8,%0; xor.b32 %4,%4,%8; mad.wide.u32 %1,t1,%9,%1; xor.b32 %5,%5,%9;"
                   "mad.wide.u32 %2,t2,%8,%2; xor.b32 %6,%6,%8; mad.wide.u32 %3,t3,%9,%3; xor.b32 %7,%7,%9; }"
        : "+l"(w0),"+l"(w1),"+l"(w2),"+l"(w3),"+r"(a4),"+r"(a5),"+r"(a6),"+r"(a7) : "r"(x),"r"(y));
    }
  }
  out[blockIdx.x*blockDim.x+threadIdx.x] = a0^a1^a2^a3^a4^a5^a6^a7^(uint32_t)(w0^w1^w2^w3)^(uint32_t)((w0^w1^w2^w3)>>32);
}
template<int V> void run(const char* name, int per_iter) {
  int blocks = 148*8; uint32_t* out; cudaMalloc(&out, blocks*256*4);
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<V><<<blocks,256>>>(out,3,5); cudaDeviceSynchronize();
  cudaEventRecord(e0); for (int r=0;r<5;r++) k<V><<<blocks,256>>>(out,3,5); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms,e0,e1); ms/=5;
  double ops = (double)blocks*256*ITER*per_iter;
  printf("%-28s %8.3f ms  %8.2f Tinstr/s (thread-level)  %6.2f lanes/clk/SM @1.965GHz\n", name, ms, ops/ms/1e9, ops/(ms*1e-3)/148/1.965e9);
  cudaFree(out);
}
template<class F, int U> __global__ void __launch_bounds__(256) kmul(F* io, int iters) {
  int i = blockIdx.x*blockDim.x+threadIdx.x;
  F x[U]; F y = io[i];
  #pragma unroll
  for (int u=0;u<U;u++) { x[u] = y; x[u].v[0] ^= u; }
  #pragma unroll 1
  for (int k=0;k<iters;k++) {
    #pragma unroll
    for (int u=0;u<U;u++) x[u] = x[u]*y;
  }
  F acc = x[0];
  #pragma unroll
  for (int u=1;u<U;u++) acc = acc + x[u];
  io[i] = acc;
}
template<class F, int U> void runmul(const char* name, int blocks_per_sm, int macs) {
  int blocks = 148*blocks_per_sm; F* io; cudaMalloc(&io, (size_t)blocks*256*sizeof(F)); cudaMemset(io, 1, (size_t)blocks*256*sizeof(F));
  int iters = 512;
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  kmul<F,U><<<blocks,256>>>(io,iters); cudaDeviceSynchronize();
  cudaEventRecord(e0); for (int r=0;r<3;r++) kmul<F,U><<<blocks,256>>>(io,iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms,e0,e1); ms/=3;
  double muls = (double)blocks*256*iters*U;
  printf("%-28s %8.3f ms  %8.2f Gmul/s  %8.2f T MAC32/s  (%d blocks/SM req)  err=%d\n", name, ms, muls/ms/1e6, muls*macs/ms/1e9, blocks_per_sm, (int)cudaGetLastError());
  cudaFree(io);
}
int main() {
  run<0>("IMAD lo x8",8); run<1>("IMAD.HI x8",8); run<2>("IMAD.WIDE x8",8); run<3>("lo/hi carry chain (4 MAC)",4);
  run<4>("IADD x8",8); run<5>("4 IMAD + 4 XOR",8); run<6>("4 IMAD.WIDE + 4 XOR",8);
  runmul<Fq,1>("Fq mul U=1 occ8", 8, 300); runmul<Fq,1>("Fq mul U=1 occ2", 2, 300); runmul<Fq,1>("Fq mul U=1 occ1", 1, 300);
  runmul<Fq,2>("Fq mul U=2 occ2", 2, 300);
  runmul<Fr,1>("Fr mul U=1 occ8", 8, 136); runmul<Fr,2>("Fr mul U=2 occ4", 4, 136);
  return 0;
}
