// ctg_field_op_batch_noalloc: the element-wise field kernel of ctg_field_op_batch for callers that must not allocate (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the
// reference.
//
// ctg_field_op_batch stages its operands and its result in the context's growable scratch buffer (ensure_scratch ->
// cudaMalloc).  This variant stages them in a file-scope __device__ array of a fixed, modest size instead:
//   * kNoAllocScratchBytes = 3 MiB, held by every GPU that loads the library until the process exits;
//   * layout  a[n] | b[n] | out[n], so 3 * n * element_bytes must fit: n <= 32 768 (Fr), 21 845 (Fq), 10 922 (Fq2);
//   * larger calls are rejected with CTG_ERR_INVALID and a message; nothing falls back to an allocating path;
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine: two copies in, one memset, one kernel, one
//     copy out, one stream synchronisation;
//   * ALL calls on a device share the one array, and it keeps the previous call's contents: the result region is cleared at
//     the start of every call, the operand regions are overwritten by the copies.  Calls that overlap on the GPU (two contexts
//     or two streams on one device, two host threads) RACE on it -- serialise them.
// In host code the array's name is only a host-side stand-in for the symbol: its device address is taken once per device with
// cudaGetSymbolAddress (that device current) and that pointer is what the copies and the kernel get.
#include <mutex>

#include "ctg_internal.h"

namespace ctg {

static constexpr size_t kNoAllocScratchBytes = (size_t)3 << 20;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

static constexpr int kMaxCudaDevices = 64;
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[kMaxCudaDevices] = {};   // device address of g_noalloc_scratch on each device, once known

template <class F> __global__ void __launch_bounds__(128) field_op_noalloc_kernel(const F* __restrict__ a, const F* __restrict__ b, F* __restrict__ out, size_t n, int op) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const F x = a[i], y = b[i];
  out[i] = op == CTG_OP_MUL ? x * y : (op == CTG_OP_ADD ? x + y : x - y);
}

// set-up, once per device, with that device current
static int32_t noalloc_scratch(ctg_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= kMaxCudaDevices) return fail(ctx, CTG_ERR_INVALID, "field_op_batch_noalloc: device index out of range");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  int cur = -1;
  CTG_CUDA(ctx, cudaGetDevice(&cur));
  if (cur != ctx->device) CTG_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!g_noalloc_ptr[ctx->device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(ctx, CTG_ERR_CUDA, std::string("field_op_batch_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
    }
    g_noalloc_ptr[ctx->device] = (unsigned char*)p;
  }
  *out = g_noalloc_ptr[ctx->device];
  return CTG_OK;
}

template <class F> static int32_t run_noalloc(ctg_ctx* ctx, int op, const void* a, const void* b, void* out, size_t n, unsigned char* d) {
  const size_t bytes = n * sizeof(F);
  F* da = (F*)d;
  F* db = (F*)(d + bytes);
  F* dout = (F*)(d + 2 * bytes);
  cudaStream_t st = ctx->stream;
  CTG_CUDA(ctx, cudaMemsetAsync(dout, 0, bytes, st));   // the array keeps the last call's contents: the result region starts at zero
  CTG_CUDA(ctx, cudaMemcpyAsync(da, a, bytes, cudaMemcpyHostToDevice, st));
  CTG_CUDA(ctx, cudaMemcpyAsync(db, b, bytes, cudaMemcpyHostToDevice, st));
  field_op_noalloc_kernel<F><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(da, db, dout, n, op);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaMemcpyAsync(out, dout, bytes, cudaMemcpyDeviceToHost, st));
  CTG_CUDA(ctx, cudaStreamSynchronize(st));
  return CTG_OK;
}

}  // namespace ctg

using namespace ctg;

extern "C" int32_t ctg_field_op_batch_noalloc(ctg_ctx* ctx, int field, int op, const void* a, const void* b, void* out, size_t n) {
  if (!ctx) return CTG_ERR_INVALID;
  const size_t eb = field == CTG_FIELD_FR ? sizeof(Fr) : (field == CTG_FIELD_FQ ? sizeof(Fq) : (field == CTG_FIELD_FQ2 ? sizeof(Fq2) : 0));
  if (!eb) return fail(ctx, CTG_ERR_INVALID, "field_op_batch_noalloc: unknown field");
  if (op != CTG_OP_MUL && op != CTG_OP_ADD && op != CTG_OP_SUB) return fail(ctx, CTG_ERR_INVALID, "field_op_batch_noalloc: op must be CTG_OP_MUL, CTG_OP_ADD or CTG_OP_SUB");
  if (n == 0) return CTG_OK;
  if (!a || !b || !out) return fail(ctx, CTG_ERR_INVALID, "field_op_batch_noalloc: null buffer");
  if (n > kNoAllocScratchBytes / (3 * eb))
    return fail(ctx, CTG_ERR_INVALID, "field_op_batch_noalloc: " + std::to_string(n) + " elements of " + std::to_string(eb) + " bytes exceed the fixed 3 MiB scratch (at most " +
                                          std::to_string(kNoAllocScratchBytes / (3 * eb)) + "); use ctg_field_op_batch");
  unsigned char* d = nullptr;
  CTG_TRY(noalloc_scratch(ctx, &d));
  switch (field) {
    case CTG_FIELD_FR: return run_noalloc<Fr>(ctx, op, a, b, out, n, d);
    case CTG_FIELD_FQ: return run_noalloc<Fq>(ctx, op, a, b, out, n, d);
    default: return run_noalloc<Fq2>(ctx, op, a, b, out, n, d);
  }
}
