// Carry-chain primitives for 32-bit-limb big-integer arithmetic on sm_100a.
//
// On the device every helper is one PTX instruction (add.cc / madc.lo.cc / madc.hi.cc ...);
// ptxas pairs `mad.lo.cc` + `madc.hi.cc` on the same operands into one IMAD.WIDE.U32(.X)
// on the FMA pipe; that instruction occupies TWO issue slots of the pipe (measured:
// profiles/r01_pipe_probe.txt), so a 32x32->64 multiply-accumulate costs two IMAD slots
// whichever way it is written -- the MAC32 peak is half the IMAD issue rate.  When the
// header is compiled for the host (g++ or the host pass of nvcc) the same helpers are
// emulated with an explicit carry word so the field/curve code above them can be unit
// tested without a GPU (tests/test_device_math_host.py).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CTG_HD __host__ __device__ __forceinline__
#define CTG_D __device__ __forceinline__
// cold group operations: one out-of-line copy per kernel keeps code size and compile time down
#define CTG_HD_COLD __host__ __device__ __noinline__
#else
#define CTG_HD inline
#define CTG_D inline
#define CTG_HD_COLD inline
#endif

namespace ctg {
namespace ptx {

#if !defined(__CUDA_ARCH__)
// host emulation of the PTX condition-code register
static thread_local uint32_t g_cc = 0;
static inline uint32_t emu_add(uint32_t a, uint32_t b, uint32_t cin, bool set) {
  uint64_t t = (uint64_t)a + b + cin;
  if (set) g_cc = (uint32_t)(t >> 32);
  return (uint32_t)t;
}
static inline uint32_t emu_sub(uint32_t a, uint32_t b, uint32_t bin, bool set) {
  uint64_t t = (uint64_t)a - b - bin;
  if (set) g_cc = (uint32_t)(t >> 63);  // borrow
  return (uint32_t)t;
}
#endif

CTG_HD uint32_t add_cc(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r;
#else
  return emu_add(a, b, 0, true);
#endif
}
CTG_HD uint32_t addc_cc(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r;
#else
  return emu_add(a, b, g_cc, true);
#endif
}
CTG_HD uint32_t addc(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r;
#else
  return emu_add(a, b, g_cc, false);
#endif
}
CTG_HD uint32_t sub_cc(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r;
#else
  return emu_sub(a, b, 0, true);
#endif
}
CTG_HD uint32_t subc_cc(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r;
#else
  return emu_sub(a, b, g_cc, true);
#endif
}
CTG_HD uint32_t subc(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r;
#else
  return emu_sub(a, b, g_cc, false);
#endif
}
CTG_HD uint32_t mul_lo(uint32_t a, uint32_t b) { return a * b; }
CTG_HD uint32_t mul_hi(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}
// r = lo(a*b) + c, sets carry
CTG_HD uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
#else
  return emu_add((uint32_t)((uint64_t)a * b), c, 0, true);
#endif
}
CTG_HD uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
#else
  return emu_add((uint32_t)((uint64_t)a * b), c, g_cc, true);
#endif
}
CTG_HD uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
#else
  return emu_add((uint32_t)(((uint64_t)a * b) >> 32), c, g_cc, true);
#endif
}
CTG_HD uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  uint32_t r; asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
#else
  return emu_add((uint32_t)(((uint64_t)a * b) >> 32), c, g_cc, false);
#endif
}

}  // namespace ptx
}  // namespace ctg
