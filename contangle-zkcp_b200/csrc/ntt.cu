// Radix-2 NTT over BLS12-381 Fr, shared-memory staged.
// Replaces ark-poly 0.3.0 Radix2EvaluationDomain::{fft,ifft,coset_fft,coset_ifft}_in_place
// (domain/radix2/fft.rs), reached from ark-groth16 R1CStoQAP::witness_map, i.e. from
// Groth16::prove at /root/reference lib/src/zk/encryption.rs:76.
//
// A transform of 2^k points is split into passes of up to 8 (strided) or 11 (contiguous)
// butterfly levels.  One CTA owns a tile of 2^11 elements (64 KB of shared memory, stored
// as two 16-byte planes so that consecutive lanes hit consecutive banks), reads it from HBM
// once with 128-bit loads in runs of >= 256 contiguous bytes, runs all levels of the pass in
// shared memory, and writes it back once.  HBM traffic is therefore 64 B per element per
// pass (+ the twiddle reads, which come from a per-domain table that mostly lives in L2).
//
// Two kernels' worth of dataflow are provided by one template:
//   DIF (Gentleman-Sande): natural order in  -> bit-reversed out
//   DIT (Cooley-Tukey):    bit-reversed in   -> natural order out
// so the witness map (ifft -> coset_fft -> pointwise -> coset_ifft) never runs a separate
// bit-reversal pass: ifft is DIF, the following coset_fft is DIT, the final coset_ifft is DIF
// and its bit-reversed output is consumed by an MSM whose bases were permuted at key upload.
// Coset shifts (g = 7), the 1/n of the inverse transform and their product are fused into the
// first load / last store of a pass through per-domain power tables.
#include <cuda.h>   // CUtensorMap and the cuTensorMapEncodeTiled prototype (resolved at run time, libcuda is not linked)

#include <algorithm>
#include <cstdlib>

#include "ctg_internal.h"

namespace ctg {

static constexpr int kTileLog = 11;     // elements per CTA tile
static constexpr int kStridedLevels = 8;
static constexpr int kThreads = 256;

struct PowParams { Fr sq[32]; };  // base^(2^j)

__global__ void pow_table_kernel(Fr* out, uint64_t n, PowParams p, Fr scale) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Fr acc = scale;
  for (int j = 0; j < 32; j++)
    if ((i >> j) & 1) acc = acc * p.sq[j];
  out[i] = acc;
}

struct NttPass {
  int log_n;
  int log_blk;   // log2 of the sub-transform size at this pass
  int S;         // butterfly levels in this pass
  int log_tile;  // log2 elements per CTA
  int dit;
  const Fr* tw;        // w^k table of the full domain, k < n/2
  const Fr* pre_tab;   // optional multiplier applied while loading (indexed by natural position)
  const Fr* post_tab;  // optional multiplier applied while storing
  int pre_bitrev, post_bitrev;  // data at that point is bit-reversed: table index = bitrev(address)
  int post_const;      // multiply by post_c while storing
  Fr post_c;
};

__device__ __forceinline__ Fr ld_fr(const uint4* lo, const uint4* hi, int i) {
  Fr r;
  uint4 a = lo[i], b = hi[i];
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ void st_fr(uint4* lo, uint4* hi, int i, const Fr& r) {
  lo[i] = make_uint4(r.v[0], r.v[1], r.v[2], r.v[3]);
  hi[i] = make_uint4(r.v[4], r.v[5], r.v[6], r.v[7]);
}
__device__ __forceinline__ Fr ldg_fr(const Fr* p) {
  const uint4* q = reinterpret_cast<const uint4*>(p);
  uint4 a = __ldg(q), b = __ldg(q + 1);
  Fr r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ uint64_t bitrev_idx(uint64_t a, int log_n) {
  return log_n == 0 ? 0 : (uint64_t)(__brev((uint32_t)a) >> (32 - log_n));
}

// blockIdx.y = index of the transform in a batch of equal-size transforms stored back to back
__global__ void __launch_bounds__(kThreads) ntt_pass_kernel(Fr* __restrict__ data, NttPass p) {
  extern __shared__ uint4 smem[];
  data += (uint64_t)blockIdx.y << p.log_n;
  const int tile = 1 << p.log_tile;
  uint4* lo = smem;
  uint4* hi = smem + tile;
  const int log_cpb = p.log_blk - p.S;    // columns per sub-transform
  const int log_C = p.log_tile - p.S;     // columns per tile
  const bool strided = log_cpb > log_C;
  const int C = 1 << log_C;
  const uint64_t tile_id = blockIdx.x;
  uint64_t base;  // address of tile-local index 0
  uint32_t j0 = 0;
  if (strided) {
    uint64_t q0 = tile_id << log_C;
    uint64_t b = q0 >> log_cpb;
    j0 = (uint32_t)(q0 & ((1ull << log_cpb) - 1));
    base = (b << p.log_blk) + j0;
  } else {
    base = tile_id << p.log_tile;
  }
  auto addr_of = [&](int idx) -> uint64_t {
    return strided ? base + ((uint64_t)(idx >> log_C) << log_cpb) + (idx & (C - 1)) : base + idx;
  };

  // ---- load tile (128-bit loads; runs of C*32 >= 256 contiguous bytes)
  for (int e = threadIdx.x; e < tile; e += kThreads) {
    uint64_t a = addr_of(e);
    const uint4* g = reinterpret_cast<const uint4*>(data + a);
    uint4 x = g[0], y = g[1];
    if (p.pre_tab) {
      Fr v;
      v.v[0] = x.x; v.v[1] = x.y; v.v[2] = x.z; v.v[3] = x.w;
      v.v[4] = y.x; v.v[5] = y.y; v.v[6] = y.z; v.v[7] = y.w;
      uint64_t ti = p.pre_bitrev ? bitrev_idx(a, p.log_n) : a;
      v = v * ldg_fr(p.pre_tab + ti);
      st_fr(lo, hi, e, v);
    } else {
      lo[e] = x;
      hi[e] = y;
    }
  }
  __syncthreads();

  // ---- butterflies
  for (int t = 0; t < p.S; t++) {
    const int log_h = p.dit ? t : (p.S - 1 - t);
    const int log_D = log_h + (strided ? log_C : log_cpb);
    const int log_half = log_h + log_cpb;
    const int tw_shift = p.log_n - 1 - log_half;
    const int D = 1 << log_D;
    for (int bf = threadIdx.x; bf < (tile >> 1); bf += kThreads) {
      const int w = bf & (D - 1);
      const int iu = ((bf >> log_D) << (log_D + 1)) + w;
      const int iv = iu + D;
      const uint64_t pos = strided ? (((uint64_t)(w >> log_C) << log_cpb) + j0 + (w & (C - 1))) : (uint64_t)w;
      const Fr tw = ldg_fr(p.tw + (pos << tw_shift));
      Fr u = ld_fr(lo, hi, iu), v = ld_fr(lo, hi, iv);
      if (p.dit) {
        Fr tv = v * tw;
        st_fr(lo, hi, iu, u + tv);
        st_fr(lo, hi, iv, u - tv);
      } else {
        st_fr(lo, hi, iu, u + v);
        st_fr(lo, hi, iv, (u - v) * tw);
      }
    }
    __syncthreads();
  }

  // ---- store tile
  for (int e = threadIdx.x; e < tile; e += kThreads) {
    uint64_t a = addr_of(e);
    uint4* g = reinterpret_cast<uint4*>(data + a);
    if (p.post_tab || p.post_const) {
      Fr v = ld_fr(lo, hi, e);
      if (p.post_tab) {
        uint64_t ti = p.post_bitrev ? bitrev_idx(a, p.log_n) : a;
        v = v * ldg_fr(p.post_tab + ti);
      } else {
        v = v * p.post_c;
      }
      g[0] = make_uint4(v.v[0], v.v[1], v.v[2], v.v[3]);
      g[1] = make_uint4(v.v[4], v.v[5], v.v[6], v.v[7]);
    } else {
      g[0] = lo[e];
      g[1] = hi[e];
    }
  }
}

// ======================================================================================================================
// Blackwell-native pass kernel (the default for tiles of >= 32 elements): the tile moves between HBM and shared memory as ONE
// tensor-map copy in each direction (cp.async.bulk.tensor = UTMALDG / UTMASTG; the copy engine computes the strided addresses:
// a strided pass is a 3-D box of 256 rows x C columns, a contiguous pass a 2-D box), completion through an mbarrier.  The tensor
// map asks for the 128-byte swizzle, so that four consecutive elements (one 128-byte line) land XOR-permuted by the line number
// and the warp-wide 16-byte accesses of the butterfly rounds are bank-conflict free.  The butterflies run in REGISTERS: a
// thread owns 8 elements for three levels (radix-8 round: 12 butterflies, 7 distinct twiddles) between two shared-memory
// exchanges, instead of one shared-memory round trip, one twiddle load and one index computation per butterfly and level.
// Two CTAs (2 x 64 KB) are resident per SM: while one computes, the other's tile is in flight -- the double buffer.
// Results are bit-identical to ntt_pass_kernel (same butterflies, same twiddles, exact field arithmetic).
static constexpr int kTmaMinLogTile = 5;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* tm, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(src), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

// byte offset of half `h` (16 bytes) of tile element e in the 128B-swizzled tile: line = e / 4, chunk ^= line % 8
__device__ __forceinline__ uint32_t sw_off(uint32_t e, uint32_t h) {
  const uint32_t line = e >> 2;
  return (line << 7) | (((((e & 3u) << 1) | h) ^ (line & 7u)) << 4);
}
__device__ __forceinline__ Fr lds_fr(const unsigned char* sm, uint32_t e) {
  const uint4 a = *reinterpret_cast<const uint4*>(sm + sw_off(e, 0));
  const uint4 b = *reinterpret_cast<const uint4*>(sm + sw_off(e, 1));
  Fr r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ void sts_fr(unsigned char* sm, uint32_t e, const Fr& r) {
  *reinterpret_cast<uint4*>(sm + sw_off(e, 0)) = make_uint4(r.v[0], r.v[1], r.v[2], r.v[3]);
  *reinterpret_cast<uint4*>(sm + sw_off(e, 1)) = make_uint4(r.v[4], r.v[5], r.v[6], r.v[7]);
}

struct TileGeom {
  int strided, log_C, log_cpb;
  uint32_t j0;
};
// table position of the twiddle of the butterfly whose upper operand sits at tile-local offset w inside its butterfly block
__device__ __forceinline__ uint64_t tw_pos(const TileGeom& g, uint32_t w) {
  return g.strided ? ((uint64_t)(w >> g.log_C) << g.log_cpb) + g.j0 + (w & ((1u << g.log_C) - 1)) : (uint64_t)w;
}

// One round of K butterfly levels on register-resident groups of 2^K elements spaced 2^g apart in the tile.
// DIF: levels run from log_h = log_h0 down; DIT: from log_h0 up.  `tile` elements, all threads of the CTA take part.
template <int K, bool DIT, int THREADS>
__device__ __forceinline__ void ntt_round(unsigned char* sm, const NttPass& p, const TileGeom& geo, int g, int log_h0, int tile) {
  constexpr int R = 1 << K;
  for (uint32_t grp = threadIdx.x; grp < (uint32_t)(tile >> K); grp += THREADS) {
    const uint32_t low = grp & ((1u << g) - 1);
    const uint32_t i0 = ((grp >> g) << (g + K)) | low;
    Fr x[R];
#pragma unroll
    for (int m = 0; m < R; m++) x[m] = lds_fr(sm, i0 + ((uint32_t)m << g));
#pragma unroll
    for (int l = 0; l < K; l++) {
      const int md = DIT ? (1 << l) : (1 << (K - 1 - l));           // register distance of the pair
      const int log_h = DIT ? log_h0 + l : log_h0 - l;
      const int tw_shift = p.log_n - 1 - (log_h + geo.log_cpb);
#pragma unroll
      for (int q = 0; q < md; q++) {
        const Fr tw = ldg_fr(p.tw + (tw_pos(geo, low | ((uint32_t)q << g)) << tw_shift));
#pragma unroll
        for (int hb = 0; hb < R / (2 * md); hb++) {
          const int m = hb * 2 * md + q;
          if (DIT) {
            const Fr tv = x[m + md] * tw;
            const Fr u = x[m];
            x[m] = u + tv;
            x[m + md] = u - tv;
          } else {
            const Fr u = x[m], v = x[m + md];
            x[m] = u + v;
            x[m + md] = (u - v) * tw;
          }
        }
      }
    }
#pragma unroll
    for (int m = 0; m < R; m++) sts_fr(sm, i0 + ((uint32_t)m << g), x[m]);
  }
}

// Instantiations.  <3, 128> (default): radix-8 rounds, 128 registers, tiles of 2^10 elements (32 KB), 128 threads -- a CTA is
// 16 K registers + 33 KB, so four are resident per SM alone (16 warps) AND one still fits beside two bucket-accumulation CTAs
// of a concurrent MSM (2 x 23.5 K registers): inside a proof the witness map runs on a high-priority stream while the MSM
// kernels hold the SMs, and a 32 K-register CTA would have to wait for one of those long-lived CTAs to retire (measured: the
// 2^22 proof took 169-175 ms with <3, 256> against 162 ms with the 58-register round-1 kernel).  <3, 256>: the same with
// tiles of 2^11 (CTG_NTT_TILE=11).  <2, 256>: radix-4 rounds, 80 registers, three CTAs per SM (CTG_NTT_RADIX4=1) -- an A/B arm.
template <int MAXK, int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS == 128 ? 4 : (MAXK == 3 ? 2 : 3)) ntt_pass_tma_kernel(const __grid_constant__ CUtensorMap tmap, NttPass p) {
  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t mbar;
  unsigned char* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // the swizzle pattern is a function of address bits 7..9
  const int tile = 1 << p.log_tile;
  TileGeom geo;
  geo.log_cpb = p.log_blk - p.S;
  geo.log_C = p.log_tile - p.S;
  geo.strided = geo.log_cpb > geo.log_C;
  geo.j0 = 0;
  const uint64_t tile_id = blockIdx.x;
  uint64_t base;   // element address (inside this transform) of tile-local index 0
  int c1, c2;      // tensor coordinates of the tile
  if (geo.strided) {
    const uint64_t q0 = tile_id << geo.log_C;
    const uint64_t b = q0 >> geo.log_cpb;
    geo.j0 = (uint32_t)(q0 & ((1ull << geo.log_cpb) - 1));
    base = (b << p.log_blk) + geo.j0;
    c1 = (int)(geo.j0 >> 2);
    c2 = (int)(((uint64_t)blockIdx.y << (p.log_n - geo.log_cpb)) + (b << p.S));
  } else {
    base = tile_id << p.log_tile;
    c1 = (int)((((uint64_t)blockIdx.y << p.log_n) + base) >> 2);
    c2 = 0;
  }
  const int lines = tile >> 2, box_lines = lines < 256 ? lines : 256;
  const uint32_t sm_a = smem_u32(sm);
  if (threadIdx.x == 0) mbar_init(&mbar, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(&mbar, (uint32_t)tile * 32u);
    if (geo.strided) tma_load_3d(sm_a, &tmap, &mbar, 0, c1, c2);
    else
      for (int l = 0; l < lines; l += box_lines) tma_load_2d(sm_a + (uint32_t)l * 128u, &tmap, &mbar, 0, c1 + l);
  }
  mbar_wait(&mbar, 0);

  auto addr_of = [&](uint32_t e) -> uint64_t {
    return geo.strided ? base + ((uint64_t)(e >> geo.log_C) << geo.log_cpb) + (e & ((1u << geo.log_C) - 1)) : base + e;
  };
  if (p.pre_tab) {
    for (uint32_t e = threadIdx.x; e < (uint32_t)tile; e += THREADS) {
      const uint64_t a = addr_of(e);
      sts_fr(sm, e, lds_fr(sm, e) * ldg_fr(p.pre_tab + (p.pre_bitrev ? bitrev_idx(a, p.log_n) : a)));
    }
    __syncthreads();
  }

  // levels t = 0 .. S-1 in rounds of up to three (rounds of two at the end so that no round is a single level when S >= 2)
  const int dbase = geo.strided ? geo.log_C : geo.log_cpb;   // log2 of the tile-local butterfly distance at log_h = 0
  int t = 0;
  while (t < p.S) {
    const int rem = p.S - t;
    const int K = MAXK == 3 ? (rem > 4 ? 3 : rem == 4 ? 2 : rem) : (rem >= 2 ? 2 : 1);
    // DIF walks log_h = S-1-t downwards (spacing = the smallest distance of the round); DIT walks log_h = t upwards
    const int log_h0 = p.dit ? t : p.S - 1 - t;
    const int g = dbase + (p.dit ? log_h0 : log_h0 - (K - 1));
    if (p.dit) {
      if (MAXK == 3 && K == 3) ntt_round<MAXK, true, THREADS>(sm, p, geo, g, log_h0, tile);
      else if (K == 2) ntt_round<2, true, THREADS>(sm, p, geo, g, log_h0, tile);
      else ntt_round<1, true, THREADS>(sm, p, geo, g, log_h0, tile);
    } else {
      if (MAXK == 3 && K == 3) ntt_round<MAXK, false, THREADS>(sm, p, geo, g, log_h0, tile);
      else if (K == 2) ntt_round<2, false, THREADS>(sm, p, geo, g, log_h0, tile);
      else ntt_round<1, false, THREADS>(sm, p, geo, g, log_h0, tile);
    }
    t += K;
    __syncthreads();
  }

  if (p.post_tab || p.post_const) {
    for (uint32_t e = threadIdx.x; e < (uint32_t)tile; e += THREADS) {
      Fr v = lds_fr(sm, e);
      if (p.post_tab) {
        const uint64_t a = addr_of(e);
        v = v * ldg_fr(p.post_tab + (p.post_bitrev ? bitrev_idx(a, p.log_n) : a));
      } else {
        v = v * p.post_c;
      }
      sts_fr(sm, e, v);
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes above -> visible to the copy engine
  __syncthreads();
  if (threadIdx.x == 0) {
    if (geo.strided) tma_store_3d(&tmap, sm_a, 0, c1, c2);
    else
      for (int l = 0; l < lines; l += box_lines) tma_store_2d(&tmap, sm_a + (uint32_t)l * 128u, 0, c1 + l);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // shared memory must outlive the copy
  }
}

__global__ void bitrev_permute_kernel(Fr* data, int log_n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (1ull << log_n)) return;
  uint64_t j = bitrev_idx(i, log_n);
  if (i < j) {
    uint4* a = reinterpret_cast<uint4*>(data + i);
    uint4* b = reinterpret_cast<uint4*>(data + j);
    uint4 a0 = a[0], a1 = a[1], b0 = b[0], b1 = b[1];
    a[0] = b0; a[1] = b1; b[0] = a0; b[1] = a1;
  }
}

// Tiled bit-reversal permutation for log_n >= 8.  Index i = (h : 4 bits | m : log_n - 8 bits | l : 4 bits) maps to
// (rev l | rev m | rev h): the 16 x 16 tile of all (h, l) for one m lands on the tile of rev m, transposed.  One CTA swaps
// the pair (m, rev m) through shared memory, so both the loads and the stores are runs of 16 consecutive elements (512 B)
// instead of single scattered 32-byte sectors.
__global__ void __launch_bounds__(256) bitrev_tiled_kernel(Fr* data, int log_n) {
  __shared__ uint4 tile[2][16][17][2];   // row stride 17 elements = 136 words: rows fall into distinct bank groups
  const int mid_bits = log_n - 8;
  const uint32_t m = blockIdx.x;
  const uint32_t rm = mid_bits ? (__brev(m) >> (32 - mid_bits)) : 0u;
  if (m > rm) return;
  const int r = threadIdx.x >> 4, c = threadIdx.x & 15;
  const int rr = __brev((uint32_t)r) >> 28, rc = __brev((uint32_t)c) >> 28;
  const uint64_t ia = ((uint64_t)r << (log_n - 4)) | ((uint64_t)m << 4) | (uint64_t)c;
  const uint64_t ib = ((uint64_t)r << (log_n - 4)) | ((uint64_t)rm << 4) | (uint64_t)c;
  uint4* pa = reinterpret_cast<uint4*>(data + ia);
  uint4* pb = reinterpret_cast<uint4*>(data + ib);
  tile[0][r][c][0] = pa[0]; tile[0][r][c][1] = pa[1];
  if (m != rm) { tile[1][r][c][0] = pb[0]; tile[1][r][c][1] = pb[1]; }
  __syncthreads();
  // position (h' = r, l' = c) of the destination tile holds source (h = rev c, l = rev r)
  if (m != rm) {
    pb[0] = tile[0][rc][rr][0]; pb[1] = tile[0][rc][rr][1];
    pa[0] = tile[1][rc][rr][0]; pa[1] = tile[1][rc][rr][1];
  } else {
    pa[0] = tile[0][rc][rr][0]; pa[1] = tile[0][rc][rr][1];
  }
}

__global__ void scale_table_kernel(Fr* data, uint64_t n, const Fr* tab, Fr c, int use_c) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  data[i] = use_c ? data[i] * c : data[i] * tab[i];
}

// ---------------------------------------------------------------------------- host side
static Fr host_fr_from_u64(uint64_t x) {
  Fr a = Fr::zero();
  a.v[0] = (uint32_t)x;
  a.v[1] = (uint32_t)(x >> 32);
  return a.to_mont();
}
static Fr host_root_of_unity(int log_n) {
  // ark-bls12-381 FrParameters::TWO_ADIC_ROOT_OF_UNITY = 7^((r-1)/2^32), canonical limbs
  static const uint32_t root[8] = {0x439f0d2bu, 0x3829971fu, 0x8c2280b9u, 0xb6368350u, 0x22c813b4u, 0xd09b6819u, 0xdfe81f20u, 0x16a2a19eu};
  Fr w;
  for (int i = 0; i < 8; i++) w.v[i] = root[i];
  w = w.to_mont();
  for (int i = log_n; i < 32; i++) w = w.sqr();
  return w;
}
static Fr host_pow_u64(Fr b, uint64_t e) {
  Fr acc = Fr::one();
  for (int i = 63; i >= 0; i--) {
    acc = acc.sqr();
    if ((e >> i) & 1) acc = acc * b;
  }
  return acc;
}

static int32_t fill_pow_table(ctg_ctx* ctx, Fr* d_out, uint64_t n, Fr base, Fr scale, cudaStream_t st) {
  if (n == 0) return CTG_OK;
  PowParams pp;
  Fr b = base;
  for (int j = 0; j < 32; j++) { pp.sq[j] = b; b = b.sqr(); }
  unsigned blocks = (unsigned)((n + 255) / 256);
  pow_table_kernel<<<blocks, 256, 0, st>>>(d_out, n, pp, scale);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

int32_t ntt_get_domain(ctg_ctx* ctx, int log_n, NttDomain** out) {
  if (log_n < 0 || log_n > 28) return fail(ctx, CTG_ERR_INVALID, "ntt: log_n out of range [0, 28]");
  auto it = ctx->domains.find(log_n);
  if (it != ctx->domains.end()) { *out = &it->second; return CTG_OK; }
  NttDomain d;
  d.log_n = log_n;
  const uint64_t n = 1ull << log_n;
  const uint64_t half = n > 1 ? n / 2 : 1;
  Fr w = host_root_of_unity(log_n), wi = w.inv();
  Fr g = host_fr_from_u64(7), gi = g.inv();
  d.n_inv = host_fr_from_u64(n).inv();
  d.zinv = (host_pow_u64(g, n) - Fr::one()).inv();
  CTG_CUDA(ctx, cudaMalloc(&d.tw_fwd, half * sizeof(Fr)));
  CTG_CUDA(ctx, cudaMalloc(&d.tw_inv, half * sizeof(Fr)));
  CTG_CUDA(ctx, cudaMalloc(&d.gpow, n * sizeof(Fr)));
  CTG_CUDA(ctx, cudaMalloc(&d.gpow_n, n * sizeof(Fr)));
  CTG_CUDA(ctx, cudaMalloc(&d.gipow_n, n * sizeof(Fr)));
  cudaStream_t st = ctx->stream;
  CTG_TRY(fill_pow_table(ctx, d.tw_fwd, half, w, Fr::one(), st));
  CTG_TRY(fill_pow_table(ctx, d.tw_inv, half, wi, Fr::one(), st));
  CTG_TRY(fill_pow_table(ctx, d.gpow, n, g, Fr::one(), st));
  CTG_TRY(fill_pow_table(ctx, d.gpow_n, n, g, d.n_inv, st));
  CTG_TRY(fill_pow_table(ctx, d.gipow_n, n, gi, d.n_inv, st));
  // a per-device attribute: once per context, not once per process (a multi-device context has one context per GPU)
  if (ctx->domains.empty())
  {
    CTG_CUDA(ctx, cudaFuncSetAttribute(ntt_pass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (1 << kTileLog) * 32));
    CTG_CUDA(ctx, cudaFuncSetAttribute(ntt_pass_tma_kernel<3, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (1 << 10) * 32 + 1024));
    CTG_CUDA(ctx, cudaFuncSetAttribute(ntt_pass_tma_kernel<3, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (1 << kTileLog) * 32 + 1024));
    CTG_CUDA(ctx, cudaFuncSetAttribute(ntt_pass_tma_kernel<2, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (1 << kTileLog) * 32 + 1024));
  }
  CTG_CUDA(ctx, cudaStreamSynchronize(st));  // other streams (prover lanes) may use the tables next
  ctx->domains[log_n] = d;
  *out = &ctx->domains[log_n];
  return CTG_OK;
}

// exported for setup.cu (key generation): power tables and the domain generator
int32_t fr_pow_table(ctg_ctx* ctx, Fr* d_out, uint64_t n, const Fr& base, const Fr& scale, cudaStream_t st) { return fill_pow_table(ctx, d_out, n, base, scale, st); }
Fr fr_root_of_unity(int log_n) { return host_root_of_unity(log_n); }
Fr fr_from_u64(uint64_t x) { return host_fr_from_u64(x); }

// ---- tensor maps for ntt_pass_tma_kernel.  cuTensorMapEncodeTiled is a pure host function of the driver; it is resolved through
// the runtime (cudaGetDriverEntryPoint), so the library still links against libcudart only.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) f = nullptr;
    cudaGetLastError();
    return (EncodeTiledFn)f;
  }();
  return fn;
}
static bool ntt_radix4() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("CTG_NTT_RADIX4"); v = e && *e == '1'; }
  return v != 0;
}
static int ntt_tma_tile_log() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("CTG_NTT_TILE"); v = e && atoi(e) == 11 ? 11 : 10; }
  return v;
}
static bool ntt_legacy() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("CTG_NTT_LEGACY"); v = e && *e == '1'; }
  return v != 0;
}
// The transform(s) at d seen as lines of 128 bytes (four elements).  Strided pass: dims (32 words, cpb / 4 line columns, rows of all
// sub-transforms of all batch entries), box = (32, C / 4, 2^S); contiguous pass: dims (32 words, lines), box = (32, min(256, tile / 4)).
static int32_t make_tile_map(ctg_ctx* ctx, CUtensorMap* tm, Fr* d, const NttPass& p, uint32_t batch) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return fail(ctx, CTG_ERR_CUDA, "ntt: the driver does not export cuTensorMapEncodeTiled");
  const int log_cpb = p.log_blk - p.S, log_C = p.log_tile - p.S;
  const cuuint32_t ones[3] = {1, 1, 1};
  CUresult r;
  if (log_cpb > log_C) {
    const cuuint64_t dims[3] = {32, 1ull << (log_cpb - 2), (cuuint64_t)batch << (p.log_n - log_cpb)};
    const cuuint64_t strides[2] = {128, 32ull << log_cpb};
    const cuuint32_t box[3] = {32, 1u << (log_C - 2), 1u << p.S};
    r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, d, dims, strides, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    const cuuint64_t lines = ((cuuint64_t)batch << p.log_n) >> 2;
    const cuuint64_t dims[2] = {32, lines};
    const cuuint64_t strides[1] = {128};
    const cuuint32_t tl = 1u << (p.log_tile - 2);
    const cuuint32_t box[2] = {32, tl < 256 ? tl : 256};
    r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d, dims, strides, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) return fail(ctx, CTG_ERR_CUDA, "ntt: cuTensorMapEncodeTiled failed with " + std::to_string((int)r));
  return CTG_OK;
}

struct PassGeom { int log_blk, S; };
// DIF order (large sub-transforms first).  DIT uses the same list reversed.
static std::vector<PassGeom> plan_passes(int log_n, int tile_log) {
  std::vector<PassGeom> v;
  int rem = log_n;
  while (rem > tile_log) { v.push_back({rem, kStridedLevels}); rem -= kStridedLevels; }
  if (rem > 0) v.push_back({rem, rem});
  return v;
}

struct Fuse {
  const Fr* pre_tab = nullptr; int pre_bitrev = 0;
  const Fr* post_tab = nullptr; int post_bitrev = 0;
  bool post_const = false; Fr post_c;
};

// Runs all passes of one transform.  dit = 0: natural -> bit-reversed; dit = 1: the reverse.
static int32_t run_passes(ctg_ctx* ctx, Fr* d, int log_n, const Fr* tw, int dit, const Fuse& f, cudaStream_t st, uint32_t batch = 1) {
  // kernel and tile size first: the pass structure depends on the tile
  const bool aligned = (reinterpret_cast<uintptr_t>(d) & 15) == 0;
  int tile_log = kTileLog;
  bool use_tma = false;
  if (!ntt_legacy() && aligned) {
    const int t = ntt_radix4() ? kTileLog : ntt_tma_tile_log();
    if ((log_n < t ? log_n : t) >= kTmaMinLogTile) { use_tma = true; tile_log = t; }
  }
  std::vector<PassGeom> plan = plan_passes(log_n, tile_log);
  if (dit) std::reverse(plan.begin(), plan.end());
  const uint64_t n = 1ull << log_n;
  if (plan.empty()) {  // n == 1: only the scale factors act
    if (batch != 1) return fail(ctx, CTG_ERR_INVALID, "ntt: batched transforms need n >= 2");
    if (f.pre_tab) { scale_table_kernel<<<1, 32, 0, st>>>(d, 1, f.pre_tab, Fr::one(), 0); CTG_KERNEL_CHECK(ctx); }
    if (f.post_tab) { scale_table_kernel<<<1, 32, 0, st>>>(d, 1, f.post_tab, Fr::one(), 0); CTG_KERNEL_CHECK(ctx); }
    if (f.post_const) { scale_table_kernel<<<1, 32, 0, st>>>(d, 1, nullptr, f.post_c, 1); CTG_KERNEL_CHECK(ctx); }
    return CTG_OK;
  }
  for (size_t i = 0; i < plan.size(); i++) {
    NttPass p;
    p.log_n = log_n;
    p.log_blk = plan[i].log_blk;
    p.S = plan[i].S;
    p.log_tile = log_n < tile_log ? log_n : tile_log;
    p.dit = dit;
    p.tw = tw;
    const bool first = i == 0, last = i + 1 == plan.size();
    p.pre_tab = first ? f.pre_tab : nullptr;
    p.pre_bitrev = f.pre_bitrev;
    p.post_tab = last ? f.post_tab : nullptr;
    p.post_bitrev = f.post_bitrev;
    p.post_const = last && f.post_const;
    p.post_c = f.post_c;
    unsigned blocks = (unsigned)(n >> p.log_tile);
    size_t smem = ((size_t)1 << p.log_tile) * 32;
    if (use_tma) {
      CUtensorMap tm;
      CTG_TRY(make_tile_map(ctx, &tm, d, p, batch));
      if (ntt_radix4()) ntt_pass_tma_kernel<2, 256><<<dim3(blocks, batch), 256, smem + 1024, st>>>(tm, p);
      else if (tile_log == 10) ntt_pass_tma_kernel<3, 128><<<dim3(blocks, batch), 128, smem + 1024, st>>>(tm, p);
      else ntt_pass_tma_kernel<3, 256><<<dim3(blocks, batch), 256, smem + 1024, st>>>(tm, p);
    } else {
      ntt_pass_kernel<<<dim3(blocks, batch), kThreads, smem, st>>>(d, p);
    }
    CTG_KERNEL_CHECK(ctx);
  }
  return CTG_OK;
}

int32_t ntt_run(ctg_ctx* ctx, Fr* d, int log_n, int inverse, int coset, cudaStream_t st) {
  NttDomain* dom;
  CTG_TRY(ntt_get_domain(ctx, log_n, &dom));
  Fuse f;
  if (!inverse) {
    if (coset) f.pre_tab = dom->gpow;  // x_i * g^i on the natural-order input
    CTG_TRY(run_passes(ctx, d, log_n, dom->tw_fwd, 0, f, st));
  } else {
    if (coset) { f.post_tab = dom->gipow_n; f.post_bitrev = 1; }  // output still bit-reversed here
    else { f.post_const = true; f.post_c = dom->n_inv; }
    CTG_TRY(run_passes(ctx, d, log_n, dom->tw_inv, 0, f, st));
  }
  if (log_n > 1) {
    uint64_t n = 1ull << log_n;
    if (log_n >= 8) bitrev_tiled_kernel<<<(unsigned)(n >> 8), 256, 0, st>>>(d, log_n);
    else bitrev_permute_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d, log_n);
    CTG_KERNEL_CHECK(ctx);
  }
  return CTG_OK;
}

int32_t ntt_ifft_then_coset_fft(ctg_ctx* ctx, Fr* d, int log_n, cudaStream_t st, uint32_t batch) {
  NttDomain* dom;
  CTG_TRY(ntt_get_domain(ctx, log_n, &dom));
  Fuse f;
  f.post_tab = dom->gpow_n;  // (1/n) * g^i, i = natural index of the bit-reversed slot
  f.post_bitrev = 1;
  CTG_TRY(run_passes(ctx, d, log_n, dom->tw_inv, 0, f, st, batch));
  Fuse none;
  return run_passes(ctx, d, log_n, dom->tw_fwd, 1, none, st, batch);
}

int32_t ntt_coset_ifft_bitrev(ctg_ctx* ctx, Fr* d, int log_n, cudaStream_t st, uint32_t batch) {
  NttDomain* dom;
  CTG_TRY(ntt_get_domain(ctx, log_n, &dom));
  Fuse f;
  f.post_tab = dom->gipow_n;
  f.post_bitrev = 1;
  return run_passes(ctx, d, log_n, dom->tw_inv, 0, f, st, batch);
}

}  // namespace ctg
