// Element-wise field kernels (tests + ncu integer-pipe captures) and the integer-pipe probes
// that give the roofline denominators.  Replaces nothing by itself: it exposes the device
// versions of ark-ff 0.3.0 Fp256/Fp384 mul_assign/add_assign/sub_assign and Fp2 mul for
// bit-exact checks by the test-suite.
#include "ctg_internal.h"

namespace ctg {

template <class F, int OP> __global__ void field_op_kernel(const F* __restrict__ a, const F* __restrict__ b, F* __restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  F x = a[i], y = b[i];
  out[i] = OP == CTG_OP_MUL ? x * y : (OP == CTG_OP_ADD ? x + y : x - y);
}
template <class F> __global__ void __launch_bounds__(128) field_inv_kernel(const F* __restrict__ a, F* __restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = a[i].inv_bingcd();
}

// The lane-paired Fq2 arithmetic of the G2 bucket-reduction tail (msm_pair.cuh), element-wise, for the test-suite:
//   which = 0: a * b     1: a^2     2: a * b - b * a^2  (Fq2L::mul2_sub(a, b, b, a^2))     3: (a == 0) ? 1 : 0 in c0 (is_zero agreement)
}  // namespace ctg
#include "msm.cuh"   // msm_pair.cuh needs copy16 / HalveArgs
namespace ctg {
__global__ void __launch_bounds__(128) fq2_paired_op_kernel(const Fq2* __restrict__ a, const Fq2* __restrict__ b, Fq2* __restrict__ out, size_t n, int which) {
  const size_t gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t i = gt >> 1;
  const unsigned par = (unsigned)(gt & 1);
  if (i >= n) return;   // pairs leave together
  const Fq* fa = reinterpret_cast<const Fq*>(a + i);
  const Fq* fb = reinterpret_cast<const Fq*>(b + i);
  Fq2L x{fa[par]}, y{fb[par]}, r;
  if (which == 0) r = x * y;
  else if (which == 1) r = x.sqr();
  else if (which == 2) r = Fq2L::mul2_sub(x, y, y, x.sqr());
  else { const bool z = x.is_zero(); r = z ? Fq2L::one() : Fq2L::zero(); }
  reinterpret_cast<Fq*>(out + i)[par] = r.v;
}

template <class F> static int32_t launch_op(ctg_ctx* ctx, int op, const void* a, const void* b, void* out, size_t n, cudaStream_t st) {
  unsigned blocks = (unsigned)((n + 127) / 128);
  const F* pa = (const F*)a; const F* pb = (const F*)b; F* po = (F*)out;
  switch (op) {
    case CTG_OP_MUL: field_op_kernel<F, CTG_OP_MUL><<<blocks, 128, 0, st>>>(pa, pb, po, n); break;
    case CTG_OP_ADD: field_op_kernel<F, CTG_OP_ADD><<<blocks, 128, 0, st>>>(pa, pb, po, n); break;
    case CTG_OP_SUB: field_op_kernel<F, CTG_OP_SUB><<<blocks, 128, 0, st>>>(pa, pb, po, n); break;
    case CTG_OP_INV: field_inv_kernel<F><<<blocks, 128, 0, st>>>(pa, po, n); break;
    default: return fail(ctx, CTG_ERR_INVALID, "field op: unknown op");
  }
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

int32_t field_op_dev(ctg_ctx* ctx, int field, int op, const void* a, const void* b, void* out, size_t n, cudaStream_t st) {
  if (n == 0) return CTG_OK;
  switch (field) {
    case CTG_FIELD_FR: return launch_op<Fr>(ctx, op, a, b, out, n, st);
    case CTG_FIELD_FQ: return launch_op<Fq>(ctx, op, a, b, out, n, st);
    case CTG_FIELD_FQ2:
      if (op >= CTG_OP_PAIRED_MUL && op <= CTG_OP_PAIRED_IS_ZERO) {
        fq2_paired_op_kernel<<<(unsigned)((2 * n + 127) / 128), 128, 0, st>>>((const Fq2*)a, (const Fq2*)b, (Fq2*)out, n, op - CTG_OP_PAIRED_MUL);
        CTG_KERNEL_CHECK(ctx);
        return CTG_OK;
      }
      return launch_op<Fq2>(ctx, op, a, b, out, n, st);
  }
  return fail(ctx, CTG_ERR_INVALID, "field op: unknown field");
}

size_t field_elem_bytes(int field) { return field == CTG_FIELD_FR ? 32 : (field == CTG_FIELD_FQ ? 48 : (field == CTG_FIELD_FQ2 ? 96 : 0)); }

// ---- probes ----------------------------------------------------------------------------------
static constexpr int kProbeIters = 4096;
// 8 independent dependent-chains of 32-bit IMAD per thread: the FMA-heavy pipe's issue peak.
__global__ void __launch_bounds__(256) imad_probe_kernel(uint32_t* out, uint32_t x, uint32_t y) {
  uint32_t a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  x += threadIdx.x;
  y ^= blockIdx.x;
#pragma unroll 1
  for (int i = 0; i < kProbeIters; i++) {
    asm volatile(
        "mad.lo.u32 %0,%0,%8,%9; mad.lo.u32 %1,%1,%8,%9; mad.lo.u32 %2,%2,%8,%9; mad.lo.u32 %3,%3,%8,%9;"
        "mad.lo.u32 %4,%4,%8,%9; mad.lo.u32 %5,%5,%8,%9; mad.lo.u32 %6,%6,%8,%9; mad.lo.u32 %7,%7,%8,%9;"
        : "+r"(a0), "+r"(a1), "+r"(a2), "+r"(a3), "+r"(a4), "+r"(a5), "+r"(a6), "+r"(a7)
        : "r"(x), "r"(y));
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
}
// dependent Fq Montgomery products, one chain per thread
__global__ void __launch_bounds__(256) fq_chain_probe_kernel(Fq* io, int iters) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  Fq y = io[i], x = y;
#pragma unroll 1
  for (int k = 0; k < iters; k++) x = x * y;
  io[i] = x;
}

int32_t imad_probe(ctg_ctx* ctx, double* imad_per_s, double* fq_mul_per_s) {
  cudaStream_t st = ctx->stream;
  const int blocks = ctx->sm_count * 8;
  uint32_t* d_out = nullptr;
  Fq* d_io = nullptr;
  CTG_CUDA(ctx, cudaMalloc(&d_out, (size_t)blocks * 256 * 4));
  CTG_CUDA(ctx, cudaMalloc(&d_io, (size_t)blocks * 256 * sizeof(Fq)));
  CTG_CUDA(ctx, cudaMemsetAsync(d_io, 1, (size_t)blocks * 256 * sizeof(Fq), st));
  cudaEvent_t e0, e1;
  CTG_CUDA(ctx, cudaEventCreate(&e0));
  CTG_CUDA(ctx, cudaEventCreate(&e1));
  float ms = 0;
  const int reps = 5, iters = 512;
  imad_probe_kernel<<<blocks, 256, 0, st>>>(d_out, 3, 5);  // warm-up
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaEventRecord(e0, st));
  for (int r = 0; r < reps; r++) { imad_probe_kernel<<<blocks, 256, 0, st>>>(d_out, 3, 5); CTG_KERNEL_CHECK(ctx); }
  CTG_CUDA(ctx, cudaEventRecord(e1, st));
  CTG_CUDA(ctx, cudaEventSynchronize(e1));
  CTG_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
  if (imad_per_s) *imad_per_s = (double)blocks * 256 * kProbeIters * 8 * reps / (ms * 1e-3);
  fq_chain_probe_kernel<<<blocks, 256, 0, st>>>(d_io, iters);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaEventRecord(e0, st));
  for (int r = 0; r < reps; r++) { fq_chain_probe_kernel<<<blocks, 256, 0, st>>>(d_io, iters); CTG_KERNEL_CHECK(ctx); }
  CTG_CUDA(ctx, cudaEventRecord(e1, st));
  CTG_CUDA(ctx, cudaEventSynchronize(e1));
  CTG_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
  if (fq_mul_per_s) *fq_mul_per_s = (double)blocks * 256 * iters * reps / (ms * 1e-3);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d_out);
  cudaFree(d_io);
  return CTG_OK;
}

}  // namespace ctg
