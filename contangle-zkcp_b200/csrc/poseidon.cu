// Bulk Poseidon over BLS12-381 Fr: many independent sponge hashes and the Poseidon Merkle tree of a ciphertext.
// SURVEY.md §8(f) row 4 -- the witness-side native crypto upstream of the prover:
//   /root/reference circuits-ark/src/sample_entries.rs:127-134  build_merkle_tree: MerkleTree::new(params, params, leaves)
//       with leaves = the ciphertext's field elements, one element per leaf (leaf digest = CRH([e]),
//       inner digest = TwoToOneCRH(left, right) = CRH([left, right]))
//   /root/reference circuits-ark/src/encryption.rs:154-172      encrypt: PoseidonSponge::new / absorb / squeeze(1)
// The permutation and the duplex-sponge bookkeeping restate ark-sponge ^0.3.0 poseidon/mod.rs (a crates.io dependency,
// absent from /root/reference -- parity unpinned): width 3 = rate 2 + capacity 1, rate cells first (absorbed element i goes to
// state[i]), every round = add round constants, S-box x^alpha (all cells in the full rounds, cell 0 in the partial ones),
// MDS product new[i] = sum_j mds[i][j] * state[j]; full_rounds/2 full + partial_rounds partial + full_rounds/2 full rounds;
// squeeze after absorb = one permutation, output state[0].  Parameters are the caller's (the reference reads its own from
// circuits-ark/src/parameters.rs: alpha 17, 8 full and 31 partial rounds); nothing here fixes them beyond the width.
//
// Wide levels: one thread owns one hash, 3 state cells in registers, round constants and the MDS matrix staged once per CTA in shared
// memory (a warp reads the same constant at the same time, i.e. a broadcast).  A permutation is
// 9 * rounds + (sbox muls) * (3 * full + partial) Fr products -- 626 for the reference's parameters -- against 64..96 bytes of
// HBM traffic per hash, so the kernels are bound by the integer pipe like the NTT butterflies, not by HBM.
#include <vector>

#include "ctg_internal.h"

struct ctg_poseidon {
  uint32_t full_rounds = 0, partial_rounds = 0, alpha = 0;
  ctg::Fr* d_consts = nullptr;   // [rounds][3] round constants, then [3][3] MDS (row-major), Montgomery form
};

namespace ctg {

static constexpr int kPosThreads = 128;
static constexpr int kMaxRounds = 128;

struct PosCfg {
  const Fr* consts;
  uint32_t full_rounds, partial_rounds, alpha;
};

__device__ __forceinline__ Fr pos_sbox(const Fr& x, uint32_t alpha) {
  // left-to-right square and multiply on the (small, warp-uniform) exponent
  Fr acc = x;
  for (int b = 30 - __clz(alpha); b >= 0; b--) {
    acc = acc.sqr();
    if ((alpha >> b) & 1) acc = acc * x;
  }
  return acc;
}

__device__ __forceinline__ void pos_permute(Fr s[3], const Fr* sc, const PosCfg& c) {
  const uint32_t rounds = c.full_rounds + c.partial_rounds;
  const uint32_t half = c.full_rounds / 2;
  const Fr* mds = sc + 3 * rounds;
  for (uint32_t r = 0; r < rounds; r++) {
    const bool full = r < half || r >= half + c.partial_rounds;
    Fr t0 = s[0] + sc[3 * r], t1 = s[1] + sc[3 * r + 1], t2 = s[2] + sc[3 * r + 2];
    t0 = pos_sbox(t0, c.alpha);
    if (full) {
      t1 = pos_sbox(t1, c.alpha);
      t2 = pos_sbox(t2, c.alpha);
    }
#pragma unroll
    for (int i = 0; i < 3; i++) s[i] = mds[3 * i] * t0 + mds[3 * i + 1] * t1 + mds[3 * i + 2] * t2;
  }
}

__device__ __forceinline__ void pos_stage(Fr* sc, const PosCfg& c) {
  const uint32_t words = (3 * (c.full_rounds + c.partial_rounds) + 9) * 2;   // in uint4
  const uint4* g = reinterpret_cast<const uint4*>(c.consts);
  uint4* s = reinterpret_cast<uint4*>(sc);
  for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) s[i] = __ldg(g + i);
  __syncthreads();
}

__device__ __forceinline__ Fr pos_load(const Fr* p) {
  const uint4* q = reinterpret_cast<const uint4*>(p);
  uint4 a = q[0], b = q[1];
  Fr r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ void pos_store(Fr* p, const Fr& v) {
  uint4* q = reinterpret_cast<uint4*>(p);
  q[0] = make_uint4(v.v[0], v.v[1], v.v[2], v.v[3]);
  q[1] = make_uint4(v.v[4], v.v[5], v.v[6], v.v[7]);
}

// out[i] = squeeze_1(absorb(in[i*k .. i*k + k)))  -- ark-sponge absorb_internal / squeeze_internal with rate 2
__global__ void __launch_bounds__(kPosThreads) poseidon_hash_kernel(PosCfg c, const Fr* __restrict__ in, uint64_t n, uint32_t k, Fr* __restrict__ out) {
  extern __shared__ uint4 pos_smem[];
  Fr* sc = reinterpret_cast<Fr*>(pos_smem);
  pos_stage(sc, c);
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Fr s[3] = {Fr::zero(), Fr::zero(), Fr::zero()};
  const Fr* src = in + i * k;
  uint32_t fill = 0;
  for (uint32_t e = 0; e < k; e++) {
    if (fill == 2) {            // rate full and more to absorb: permute, restart at cell 0
      pos_permute(s, sc, c);
      fill = 0;
    }
    const Fr x = pos_load(src + e);
    if (fill == 0) s[0] = s[0] + x;     // (no dynamic index: keeps the state in registers)
    else s[1] = s[1] + x;
    fill++;
  }
  pos_permute(s, sc, c);        // squeeze from the absorbing state
  pos_store(out + i, s[0]);
}

// parents[i] = CRH([children[2i], children[2i+1]])
__global__ void __launch_bounds__(kPosThreads) poseidon_compress_kernel(PosCfg c, const Fr* __restrict__ children, uint64_t n_parents, Fr* __restrict__ parents) {
  extern __shared__ uint4 pos_smem[];
  Fr* sc = reinterpret_cast<Fr*>(pos_smem);
  pos_stage(sc, c);
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_parents) return;
  Fr s[3] = {pos_load(children + 2 * i), pos_load(children + 2 * i + 1), Fr::zero()};
  pos_permute(s, sc, c);
  pos_store(parents + i, s[0]);
}

// Narrow tree levels are latency-bound (a permutation is one dependent chain of products), so below kSplitBelow nodes a hash
// is spread over a group of four lanes, one state cell per lane (the fourth idles): the three S-boxes of a full round and the
// three rows of the MDS product run side by side and the chain drops from 14..24 to 8 products per round.  The cells meet
// through warp shuffles.  arity 1: digest of one element (leaf level); arity 2: TwoToOneCRH.
static constexpr uint64_t kSplitBelow = 1ull << 15;

__device__ __forceinline__ Fr pos_shfl(const Fr& v, unsigned src) {
  Fr r;
#pragma unroll
  for (int w = 0; w < 8; w++) r.v[w] = __shfl_sync(0xffffffffu, v.v[w], src);
  return r;
}

__global__ void __launch_bounds__(kPosThreads) poseidon_split_kernel(PosCfg c, const Fr* __restrict__ in, uint64_t n, uint32_t arity, Fr* __restrict__ out) {
  extern __shared__ uint4 pos_smem[];
  Fr* sc = reinterpret_cast<Fr*>(pos_smem);
  pos_stage(sc, c);
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t i = t >> 2;
  const uint32_t lane_cell = (uint32_t)(t & 3);
  const uint32_t cell = lane_cell < 2 ? lane_cell : 2;     // the idle fourth lane shadows cell 2; nothing reads it
  const bool active = i < n;                                // whole groups are active or not; every lane stays for the shuffles
  Fr s = Fr::zero();
  if (active && lane_cell < arity) s = pos_load(in + (uint64_t)arity * i + lane_cell);
  const unsigned base = (threadIdx.x & 31u) & ~3u;
  const uint32_t rounds = c.full_rounds + c.partial_rounds;
  const uint32_t half = c.full_rounds / 2;
  const Fr* row = sc + 3 * rounds + 3 * cell;
  for (uint32_t r = 0; r < rounds; r++) {
    const bool full = r < half || r >= half + c.partial_rounds;
    Fr x = s + sc[3 * r + cell];
    if (lane_cell == 0 || (full && lane_cell < 3)) x = pos_sbox(x, c.alpha);
    const Fr t0 = pos_shfl(x, base), t1 = pos_shfl(x, base + 1), t2 = pos_shfl(x, base + 2);
    s = row[0] * t0 + row[1] * t1 + row[2] * t2;
  }
  if (active && lane_cell == 0) pos_store(out + i, s);
}

static size_t pos_smem_bytes(const ctg_poseidon* p) { return (size_t)(3 * (p->full_rounds + p->partial_rounds) + 9) * sizeof(Fr); }
static PosCfg pos_cfg(const ctg_poseidon* p) { return PosCfg{p->d_consts, p->full_rounds, p->partial_rounds, p->alpha}; }

static int32_t pos_split_dev(ctg_ctx* ctx, const ctg_poseidon* p, const Fr* d_in, uint64_t n, uint32_t arity, Fr* d_out, cudaStream_t st) {
  const uint64_t blocks = (4 * n + kPosThreads - 1) / kPosThreads;
  poseidon_split_kernel<<<(unsigned)blocks, kPosThreads, pos_smem_bytes(p), st>>>(pos_cfg(p), d_in, n, arity, d_out);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

static int32_t pos_hash_dev(ctg_ctx* ctx, const ctg_poseidon* p, const Fr* d_in, uint64_t n, uint32_t k, Fr* d_out, cudaStream_t st) {
  if (n == 0) return CTG_OK;
  const uint64_t blocks = (n + kPosThreads - 1) / kPosThreads;
  poseidon_hash_kernel<<<(unsigned)blocks, kPosThreads, pos_smem_bytes(p), st>>>(pos_cfg(p), d_in, n, k, d_out);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

// d_leaf_digests: n entries; d_inner: n - 1 entries in heap order (root at 0, children of i at 2i+1 and 2i+2)
static int32_t pos_merkle_dev(ctg_ctx* ctx, const ctg_poseidon* p, const Fr* d_leaves, uint64_t n, Fr* d_leaf_digests, Fr* d_inner, cudaStream_t st) {
  if (n < kSplitBelow) CTG_TRY(pos_split_dev(ctx, p, d_leaves, n, 1, d_leaf_digests, st));
  else CTG_TRY(pos_hash_dev(ctx, p, d_leaves, n, 1, d_leaf_digests, st));
  const Fr* children = d_leaf_digests;
  for (uint64_t np = n / 2; np >= 1; np /= 2) {
    Fr* parents = d_inner + (np - 1);       // level with np nodes starts at heap index np - 1
    if (np < kSplitBelow) {
      CTG_TRY(pos_split_dev(ctx, p, children, np, 2, parents, st));
    } else {
      const uint64_t blocks = (np + kPosThreads - 1) / kPosThreads;
      poseidon_compress_kernel<<<(unsigned)blocks, kPosThreads, pos_smem_bytes(p), st>>>(pos_cfg(p), children, np, parents);
      CTG_KERNEL_CHECK(ctx);
    }
    children = parents;
  }
  return CTG_OK;
}

}  // namespace ctg

using namespace ctg;

extern "C" {

int32_t ctg_poseidon_new(ctg_ctx* ctx, uint32_t full_rounds, uint32_t partial_rounds, uint32_t alpha, const ctg_fr* mds, const ctg_fr* ark,
                         ctg_poseidon** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  if (!mds || !ark) return fail(ctx, CTG_ERR_INVALID, "poseidon: null parameters");
  if (full_rounds == 0 || (full_rounds & 1) || full_rounds + partial_rounds > kMaxRounds)
    return fail(ctx, CTG_ERR_INVALID, "poseidon: full_rounds must be even and > 0, full + partial <= 128");
  if (alpha < 2) return fail(ctx, CTG_ERR_INVALID, "poseidon: alpha must be >= 2");
  const uint32_t rounds = full_rounds + partial_rounds;
  ctg_poseidon* p = new ctg_poseidon();
  p->full_rounds = full_rounds;
  p->partial_rounds = partial_rounds;
  p->alpha = alpha;
  const size_t bytes = (size_t)(3 * rounds + 9) * sizeof(Fr);
  cudaError_t e = cudaMalloc(&p->d_consts, bytes);
  if (e != cudaSuccess) { cudaGetLastError(); delete p; return fail(ctx, CTG_ERR_OOM, "poseidon: out of device memory"); }
  std::vector<Fr> host(3 * rounds + 9);
  memcpy(host.data(), ark, (size_t)3 * rounds * sizeof(Fr));
  memcpy(host.data() + 3 * rounds, mds, 9 * sizeof(Fr));
  e = cudaMemcpyAsync(p->d_consts, host.data(), bytes, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) { cudaGetLastError(); cudaFree(p->d_consts); delete p; return fail(ctx, CTG_ERR_CUDA, "poseidon: parameter upload failed"); }
  if (bytes > 48 * 1024) {
    cudaFuncSetAttribute(poseidon_hash_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    cudaFuncSetAttribute(poseidon_compress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    cudaFuncSetAttribute(poseidon_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  }
  *out = p;
  return CTG_OK;
}

void ctg_poseidon_free(ctg_ctx* ctx, ctg_poseidon* p) {
  (void)ctx;
  if (!p) return;
  cudaFree(p->d_consts);
  delete p;
}

int32_t ctg_poseidon_hash_batch_dev(ctg_ctx* ctx, const ctg_poseidon* p, const void* d_in, uint64_t n, uint32_t k, void* d_out) {
  if (!ctx || !p) return ctx ? fail(ctx, CTG_ERR_INVALID, "poseidon: null handle") : CTG_ERR_INVALID;
  if (n && (!d_in || !d_out)) return fail(ctx, CTG_ERR_INVALID, "poseidon: null buffer");
  if (k == 0) return fail(ctx, CTG_ERR_INVALID, "poseidon: inputs per hash must be >= 1");
  return pos_hash_dev(ctx, p, (const Fr*)d_in, n, k, (Fr*)d_out, ctx->stream);
}

int32_t ctg_poseidon_hash_batch(ctg_ctx* ctx, const ctg_poseidon* p, const ctg_fr* in, uint64_t n, uint32_t k, ctg_fr* out) {
  if (!ctx || !p) return ctx ? fail(ctx, CTG_ERR_INVALID, "poseidon: null handle") : CTG_ERR_INVALID;
  if (k == 0) return fail(ctx, CTG_ERR_INVALID, "poseidon: inputs per hash must be >= 1");
  if (n == 0) return CTG_OK;
  if (!in || !out) return fail(ctx, CTG_ERR_INVALID, "poseidon: null buffer");
  const size_t in_bytes = (size_t)n * k * sizeof(Fr), out_bytes = (size_t)n * sizeof(Fr);
  CTG_TRY(ensure_scratch(ctx, in_bytes + out_bytes));
  Fr* d_in = (Fr*)ctx->scratch;
  Fr* d_out = d_in + n * k;
  CTG_CUDA(ctx, cudaMemcpyAsync(d_in, in, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
  CTG_TRY(pos_hash_dev(ctx, p, d_in, n, k, d_out, ctx->stream));
  CTG_CUDA(ctx, cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}

int32_t ctg_poseidon_merkle_tree_dev(ctg_ctx* ctx, const ctg_poseidon* p, const void* d_leaves, uint64_t n, void* d_leaf_digests, void* d_inner) {
  if (!ctx || !p) return ctx ? fail(ctx, CTG_ERR_INVALID, "poseidon: null handle") : CTG_ERR_INVALID;
  if (n < 2 || (n & (n - 1))) return fail(ctx, CTG_ERR_INVALID, "merkle tree: number of leaves must be a power of two >= 2");
  if (!d_leaves || !d_leaf_digests || !d_inner) return fail(ctx, CTG_ERR_INVALID, "merkle tree: null buffer");
  return pos_merkle_dev(ctx, p, (const Fr*)d_leaves, n, (Fr*)d_leaf_digests, (Fr*)d_inner, ctx->stream);
}

int32_t ctg_poseidon_merkle_tree(ctg_ctx* ctx, const ctg_poseidon* p, const ctg_fr* leaves, uint64_t n, ctg_fr* leaf_digests, ctg_fr* inner) {
  if (!ctx || !p) return ctx ? fail(ctx, CTG_ERR_INVALID, "poseidon: null handle") : CTG_ERR_INVALID;
  if (n < 2 || (n & (n - 1))) return fail(ctx, CTG_ERR_INVALID, "merkle tree: number of leaves must be a power of two >= 2");
  if (!leaves || !inner) return fail(ctx, CTG_ERR_INVALID, "merkle tree: null buffer");
  const size_t nb = (size_t)n * sizeof(Fr);
  CTG_TRY(ensure_scratch(ctx, 3 * nb));
  Fr* d_leaves = (Fr*)ctx->scratch;
  Fr* d_dig = d_leaves + n;
  Fr* d_inner = d_dig + n;
  CTG_CUDA(ctx, cudaMemcpyAsync(d_leaves, leaves, nb, cudaMemcpyHostToDevice, ctx->stream));
  CTG_TRY(pos_merkle_dev(ctx, p, d_leaves, n, d_dig, d_inner, ctx->stream));
  if (leaf_digests) CTG_CUDA(ctx, cudaMemcpyAsync(leaf_digests, d_dig, nb, cudaMemcpyDeviceToHost, ctx->stream));
  CTG_CUDA(ctx, cudaMemcpyAsync(inner, d_inner, nb - sizeof(Fr), cudaMemcpyDeviceToHost, ctx->stream));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}

}  // extern "C"
