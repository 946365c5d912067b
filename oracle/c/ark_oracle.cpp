// TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the arkworks 0.3.0 algorithms
// behind ark_groth16::create_proof for BLS12-381.  Nothing in the product
// (contangle-zkcp_b200/) may include, link or call this file; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
//
// PARITY UNPINNED at byte level: /root/reference (nulltea/contangle-zkcp) holds no golden
// proof / MSM / FFT vectors (circuits-ark/src/encryption.rs:403-411 only asserts accept),
// no Rust toolchain exists here, and the arithmetic lives in crates.io dependencies that
// are NOT in the tree: ark-ff, ark-ec, ark-poly, ark-groth16, ark-relations,
// ark-bls12-381 -- all "^0.3.0" (circuits-ark/Cargo.toml:10-25; Cargo.lock git-ignored,
// .gitignore:7).  This file restates their published algorithms:
//   ark-ff   fields/models/fp_{256,384}.rs   64-bit-limb Montgomery (R = 2^256 / 2^384)
//   ark-ff   fields/models/fp2.rs            Fq2 = Fq[u]/(u^2+1), Karatsuba
//   ark-ec   models/short_weierstrass_jacobian.rs  Jacobian add / mixed add / double
//   ark-ec   msm/variable_base.rs            unsigned-window Pippenger, one task per window
//   ark-poly domain/radix2/{mod,fft}.rs      in-place radix-2 FFT, coset shift g = 7
//   ark-groth16 r1cs_to_qap.rs, prover.rs, generator.rs
// Reference call sites that reach it: lib/src/zk/encryption.rs:76,
// verifiable_encryption.rs:92, sample_entries.rs:86, property.rs:133.
// It is cross-checked against oracle/py (independent big-int code) and against the
// Groth16 trapdoor closed form in tests/.
//
// Wire layout (shared with include/ctg_groth16.h): field elements are little-endian
// Montgomery limb images (Fr 32 B, Fq 48 B), G1 affine = x||y (96 B), G2 affine =
// x.c0||x.c1||y.c0||y.c1 (192 B), the point at infinity is the all-zero image.
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

typedef unsigned __int128 u128;
typedef uint64_t u64;

// ------------------------------------------------------------------ threads
static int g_threads = 0;
static int n_threads() {
  if (g_threads > 0) return g_threads;
  unsigned h = std::thread::hardware_concurrency();
  return h ? (int)h : 1;
}
// rayon-style: split [0,n) into contiguous chunks, one per worker.
static void parallel_chunks(size_t n, const std::function<void(size_t, size_t)>& fn, size_t min_chunk = 1) {
  size_t t = (size_t)n_threads();
  if (n / (min_chunk ? min_chunk : 1) < t) t = std::max<size_t>(1, n / (min_chunk ? min_chunk : 1));
  if (t <= 1) { fn(0, n); return; }
  std::vector<std::thread> th;
  size_t per = (n + t - 1) / t;
  for (size_t i = 0; i < t; i++) {
    size_t lo = i * per, hi = std::min(n, lo + per);
    if (lo >= hi) break;
    th.emplace_back([=, &fn] { fn(lo, hi); });
  }
  for (auto& x : th) x.join();
}
// one task per index, dynamic
static void parallel_tasks(size_t n, const std::function<void(size_t)>& fn) {
  size_t t = std::min<size_t>(n, (size_t)n_threads());
  if (t <= 1) { for (size_t i = 0; i < n; i++) fn(i); return; }
  std::atomic<size_t> next(0);
  std::vector<std::thread> th;
  for (size_t k = 0; k < t; k++) th.emplace_back([&] { for (;;) { size_t i = next++; if (i >= n) break; fn(i); } });
  for (auto& x : th) x.join();
}

// ------------------------------------------------------------------ prime fields
template <int N> struct Big { u64 v[N]; };

template <int N> static inline bool geq(const u64* a, const u64* b) {
  for (int i = N - 1; i >= 0; i--) { if (a[i] != b[i]) return a[i] > b[i]; }
  return true;
}
template <int N> static inline u64 add_n(u64* r, const u64* a, const u64* b) {
  u64 c = 0;
  for (int i = 0; i < N; i++) { u128 t = (u128)a[i] + b[i] + c; r[i] = (u64)t; c = (u64)(t >> 64); }
  return c;
}
template <int N> static inline u64 sub_n(u64* r, const u64* a, const u64* b) {
  u64 br = 0;
  for (int i = 0; i < N; i++) { u128 t = (u128)a[i] - b[i] - br; r[i] = (u64)t; br = (u64)(t >> 64) & 1; }
  return br;
}

template <int N, int ID> struct Fp {
  u64 v[N];
  static u64 MOD[N], R1[N], R2[N], INV;  // R1 = R mod p (Montgomery one), R2 = R^2 mod p
  static void init(const u64* mod) {
    memcpy(MOD, mod, sizeof(MOD));
    u64 inv = 1;
    for (int i = 0; i < 63; i++) { inv *= inv; inv *= MOD[0]; }
    INV = (u64)0 - inv;
    // R mod p by repeated doubling of 1
    u64 t[N] = {1};
    for (int i = 0; i < 64 * N * 2; i++) {
      u64 c = add_n<N>(t, t, t);
      if (c || geq<N>(t, MOD)) sub_n<N>(t, t, MOD);
      if (i == 64 * N - 1) memcpy(R1, t, sizeof(t));
    }
    memcpy(R2, t, sizeof(t));
  }
  static Fp zero() { Fp r; memset(r.v, 0, sizeof(r.v)); return r; }
  static Fp one() { Fp r; memcpy(r.v, R1, sizeof(r.v)); return r; }
  bool is_zero() const { u64 o = 0; for (int i = 0; i < N; i++) o |= v[i]; return o == 0; }
  bool operator==(const Fp& b) const { return memcmp(v, b.v, sizeof(v)) == 0; }
  bool operator!=(const Fp& b) const { return !(*this == b); }
  Fp operator+(const Fp& b) const {
    Fp r; add_n<N>(r.v, v, b.v);  // p < 2^(64N-1): no carry out
    if (geq<N>(r.v, MOD)) sub_n<N>(r.v, r.v, MOD);
    return r;
  }
  Fp operator-(const Fp& b) const {
    Fp r; if (sub_n<N>(r.v, v, b.v)) add_n<N>(r.v, r.v, MOD);
    return r;
  }
  Fp neg() const { if (is_zero()) return *this; Fp r; sub_n<N>(r.v, MOD, v); return r; }
  Fp dbl() const { return *this + *this; }
  // CIOS Montgomery product, fully reduced (ark-ff Fp*::mul_assign semantics)
  Fp operator*(const Fp& b) const {
    u64 t[N + 2];
    memset(t, 0, sizeof(t));
    for (int i = 0; i < N; i++) {
      u64 c = 0;
      for (int j = 0; j < N; j++) { u128 x = (u128)v[j] * b.v[i] + t[j] + c; t[j] = (u64)x; c = (u64)(x >> 64); }
      u128 x = (u128)t[N] + c; t[N] = (u64)x; t[N + 1] = (u64)(x >> 64);
      u64 m = t[0] * INV;
      x = (u128)m * MOD[0] + t[0]; c = (u64)(x >> 64);
      for (int j = 1; j < N; j++) { x = (u128)m * MOD[j] + t[j] + c; t[j - 1] = (u64)x; c = (u64)(x >> 64); }
      x = (u128)t[N] + c; t[N - 1] = (u64)x; t[N] = t[N + 1] + (u64)(x >> 64);
    }
    Fp r; memcpy(r.v, t, sizeof(r.v));
    if (t[N] || geq<N>(r.v, MOD)) sub_n<N>(r.v, r.v, MOD);
    return r;
  }
  // dedicated squaring as ark-ff's square_in_place: off-diagonal products once, doubled, plus the diagonal; then one
  // Montgomery reduction of the 2N-limb square (N(N+1)/2 + N^2 word products instead of 2 N^2)
  Fp sqr() const {
    u64 t[2 * N];
    memset(t, 0, sizeof(t));
    for (int i = 0; i < N - 1; i++) {
      u64 c = 0;
      for (int j = i + 1; j < N; j++) { u128 x = (u128)v[i] * v[j] + t[i + j] + c; t[i + j] = (u64)x; c = (u64)(x >> 64); }
      t[i + N] = c;
    }
    t[2 * N - 1] = t[2 * N - 2] >> 63;
    for (int k = 2 * N - 2; k >= 2; k--) t[k] = (t[k] << 1) | (t[k - 1] >> 63);
    t[1] <<= 1;
    u64 c = 0;
    for (int i = 0; i < N; i++) {
      u128 x = (u128)v[i] * v[i] + t[2 * i] + c; t[2 * i] = (u64)x; c = (u64)(x >> 64);
      x = (u128)t[2 * i + 1] + c; t[2 * i + 1] = (u64)x; c = (u64)(x >> 64);
    }
    u64 carry2 = 0;
    for (int i = 0; i < N; i++) {
      u64 m = t[i] * INV, cc = 0;
      for (int j = 0; j < N; j++) { u128 x = (u128)m * MOD[j] + t[i + j] + cc; t[i + j] = (u64)x; cc = (u64)(x >> 64); }
      u128 x = (u128)t[i + N] + cc + carry2; t[i + N] = (u64)x; carry2 = (u64)(x >> 64);
    }
    Fp r; memcpy(r.v, t + N, sizeof(r.v));
    if (carry2 || geq<N>(r.v, MOD)) sub_n<N>(r.v, r.v, MOD);
    return r;
  }
  Fp pow(const u64* e, int n) const {
    Fp acc = one();
    for (int i = n * 64 - 1; i >= 0; i--) { acc = acc.sqr(); if ((e[i / 64] >> (i % 64)) & 1) acc = acc * *this; }
    return acc;
  }
  Fp inv() const { u64 e[N]; u64 two[N] = {2}; sub_n<N>(e, MOD, two); return pow(e, N); }
  static Fp from_canonical(const u64* c) { Fp a, r2; memcpy(a.v, c, sizeof(a.v)); memcpy(r2.v, R2, sizeof(r2.v)); return a * r2; }
  static Fp from_u64(u64 x) { u64 c[N] = {x}; return from_canonical(c); }
  void to_canonical(u64* out) const { Fp o; memset(o.v, 0, sizeof(o.v)); o.v[0] = 1; Fp r = *this * o; memcpy(out, r.v, sizeof(r.v)); }
};
template <int N, int ID> u64 Fp<N, ID>::MOD[N];
template <int N, int ID> u64 Fp<N, ID>::R1[N];
template <int N, int ID> u64 Fp<N, ID>::R2[N];
template <int N, int ID> u64 Fp<N, ID>::INV;

typedef Fp<4, 0> Fr;
typedef Fp<6, 1> Fq;

static const u64 FR_MOD[4] = {0xffffffff00000001ULL, 0x53bda402fffe5bfeULL, 0x3339d80809a1d805ULL, 0x73eda753299d7d48ULL};
static const u64 FQ_MOD[6] = {0xb9feffffffffaaabULL, 0x1eabfffeb153ffffULL, 0x6730d2a0f6b0f624ULL,
                              0x64774b84f38512bfULL, 0x4b1ba7b6434bacd7ULL, 0x1a0111ea397fe69aULL};

struct Fq2 {
  Fq c0, c1;
  static Fq2 zero() { return {Fq::zero(), Fq::zero()}; }
  static Fq2 one() { return {Fq::one(), Fq::zero()}; }
  bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
  bool operator==(const Fq2& b) const { return c0 == b.c0 && c1 == b.c1; }
  bool operator!=(const Fq2& b) const { return !(*this == b); }
  Fq2 operator+(const Fq2& b) const { return {c0 + b.c0, c1 + b.c1}; }
  Fq2 operator-(const Fq2& b) const { return {c0 - b.c0, c1 - b.c1}; }
  Fq2 neg() const { return {c0.neg(), c1.neg()}; }
  Fq2 dbl() const { return {c0.dbl(), c1.dbl()}; }
  Fq2 operator*(const Fq2& b) const {  // Karatsuba, nonresidue -1
    Fq v0 = c0 * b.c0, v1 = c1 * b.c1;
    return {v0 - v1, (c0 + c1) * (b.c0 + b.c1) - v0 - v1};
  }
  Fq2 sqr() const { Fq t = c0 * c1; return {(c0 + c1) * (c0 - c1), t + t}; }
  Fq2 inv() const { Fq n = (c0.sqr() + c1.sqr()).inv(); return {c0 * n, (c1 * n).neg()}; }
};

// ------------------------------------------------------------------ curves (a = 0)
template <class F> struct Affine { F x, y; bool is_inf() const { return x.is_zero() && y.is_zero(); } };
template <class F> struct Jac {
  F X, Y, Z;
  static Jac inf() { return {F::one(), F::one(), F::zero()}; }
  bool is_inf() const { return Z.is_zero(); }
};

// ark-ec double_in_place (dbl-2009-l)
template <class F> static Jac<F> jdbl(const Jac<F>& p) {
  if (p.is_inf()) return p;
  F A = p.X.sqr(), B = p.Y.sqr(), C = B.sqr();
  F D = ((p.X + B).sqr() - A - C).dbl();
  F E = A.dbl() + A, Fv = E.sqr();
  Jac<F> r;
  r.X = Fv - D.dbl();
  r.Y = E * (D - r.X) - C.dbl().dbl().dbl();
  r.Z = (p.Y * p.Z).dbl();
  return r;
}
// ark-ec add_assign_mixed (madd-2007-bl shape)
template <class F> static Jac<F> jadd_mixed(const Jac<F>& p, const Affine<F>& q) {
  if (q.is_inf()) return p;
  if (p.is_inf()) return {q.x, q.y, F::one()};
  F Z1Z1 = p.Z.sqr(), U2 = q.x * Z1Z1, S2 = q.y * p.Z * Z1Z1;
  if (p.X == U2) { if (p.Y == S2) return jdbl(p); return Jac<F>::inf(); }
  F H = U2 - p.X, HH = H.sqr(), I = HH.dbl().dbl(), J = H * I, r = (S2 - p.Y).dbl(), V = p.X * I;
  Jac<F> o;
  o.X = r.sqr() - J - V.dbl();
  o.Y = r * (V - o.X) - (p.Y * J).dbl();
  o.Z = (p.Z + H).sqr() - Z1Z1 - HH;
  return o;
}
// ark-ec add_assign (add-2007-bl shape)
template <class F> static Jac<F> jadd(const Jac<F>& p, const Jac<F>& q) {
  if (p.is_inf()) return q;
  if (q.is_inf()) return p;
  F Z1Z1 = p.Z.sqr(), Z2Z2 = q.Z.sqr(), U1 = p.X * Z2Z2, U2 = q.X * Z1Z1;
  F S1 = p.Y * q.Z * Z2Z2, S2 = q.Y * p.Z * Z1Z1;
  if (U1 == U2) { if (S1 == S2) return jdbl(p); return Jac<F>::inf(); }
  F H = U2 - U1, I = H.dbl().sqr(), J = H * I, r = (S2 - S1).dbl(), V = U1 * I;
  Jac<F> o;
  o.X = r.sqr() - J - V.dbl();
  o.Y = r * (V - o.X) - (S1 * J).dbl();
  o.Z = ((p.Z + q.Z).sqr() - Z1Z1 - Z2Z2) * H;
  return o;
}
template <class F> static Jac<F> jneg(const Jac<F>& p) { return {p.X, p.Y.neg(), p.Z}; }
template <class F> static Affine<F> to_affine(const Jac<F>& p) {
  if (p.is_inf()) return {F::zero(), F::zero()};
  F zi = p.Z.inv(), zi2 = zi.sqr();
  return {p.X * zi2, p.Y * zi2 * zi};
}
template <class F> static Jac<F> from_affine(const Affine<F>& a) { return a.is_inf() ? Jac<F>::inf() : Jac<F>{a.x, a.y, F::one()}; }
// scalar given as canonical 4 x u64
template <class F> static Jac<F> jmul(const Jac<F>& p, const u64* k) {
  Jac<F> acc = Jac<F>::inf();
  for (int i = 255; i >= 0; i--) { acc = jdbl(acc); if ((k[i / 64] >> (i % 64)) & 1) acc = jadd(acc, p); }
  return acc;
}
template <class F> static void batch_to_affine(const Jac<F>* in, Affine<F>* out, size_t n) {
  std::vector<F> pref(n);
  F acc = F::one();
  for (size_t i = 0; i < n; i++) { pref[i] = acc; if (!in[i].is_inf()) acc = acc * in[i].Z; }
  F inv = acc.inv();
  for (size_t i = n; i-- > 0;) {
    if (in[i].is_inf()) { out[i] = {F::zero(), F::zero()}; continue; }
    F zi = inv * pref[i];
    inv = inv * in[i].Z;
    F zi2 = zi.sqr();
    out[i] = {in[i].X * zi2, in[i].Y * zi2 * zi};
  }
}

// ------------------------------------------------------------------ MSM (ark-ec 0.3.0 variable_base.rs)
static int ark_log2_ceil(size_t x) { int l = 0; while (((size_t)1 << l) < x) l++; return l; }
static int ark_window(size_t n) { return n < 32 ? 3 : ark_log2_ceil(n) * 69 / 100 + 2; }

template <class F> static Jac<F> msm_ark(const Affine<F>* bases, const u64* scalars /*canonical, 4 limbs each*/, size_t n) {
  const int c = ark_window(n), num_bits = 255;
  std::vector<int> starts;
  for (int w = 0; w < num_bits; w += c) starts.push_back(w);
  std::vector<Jac<F>> sums(starts.size());
  parallel_tasks(starts.size(), [&](size_t wi) {
    const int w_start = starts[wi];
    Jac<F> res = Jac<F>::inf();
    std::vector<Jac<F>> buckets(((size_t)1 << c) - 1, Jac<F>::inf());
    for (size_t i = 0; i < n; i++) {
      const u64* s = scalars + 4 * i;
      if ((s[0] | s[1] | s[2] | s[3]) == 0) continue;
      if (s[0] == 1 && (s[1] | s[2] | s[3]) == 0) { if (w_start == 0) res = jadd_mixed(res, bases[i]); continue; }
      int limb = w_start / 64, off = w_start % 64;
      u64 d = s[limb] >> off;
      if (off + c > 64 && limb + 1 < 4) d |= s[limb + 1] << (64 - off);
      d &= ((u64)1 << c) - 1;
      if (d) buckets[d - 1] = jadd_mixed(buckets[d - 1], bases[i]);
    }
    Jac<F> running = Jac<F>::inf();
    for (size_t b = buckets.size(); b-- > 0;) { running = jadd(running, buckets[b]); res = jadd(res, running); }
    sums[wi] = res;
  });
  Jac<F> total = sums.back();
  for (size_t wi = sums.size() - 1; wi-- > 0;) {
    for (int k = 0; k < c; k++) total = jdbl(total);
    total = jadd(total, sums[wi]);
  }
  return total;
}

// ------------------------------------------------------------------ radix-2 FFT (ark-poly 0.3.0)
static const u64 FR_TWO_ADIC_ROOT_CANON[4] = {0x3829971f439f0d2bULL, 0xb63683508c2280b9ULL, 0xd09b681922c813b4ULL, 0x16a2a19edfe81f20ULL};
static Fr fr_root_of_unity(int log_n) {
  Fr w = Fr::from_canonical(FR_TWO_ADIC_ROOT_CANON);
  for (int i = log_n; i < 32; i++) w = w.sqr();
  return w;
}
static void distribute_powers(Fr* a, size_t n, Fr g) {
  parallel_chunks(n, [&](size_t lo, size_t hi) {
    u64 e[1] = {lo};
    Fr p = g.pow(e, 1);
    for (size_t i = lo; i < hi; i++) { a[i] = a[i] * p; p = p * g; }
  }, 1024);
}
static void fft_in_place(Fr* a, int log_n, Fr omega) {
  size_t n = (size_t)1 << log_n;
  parallel_chunks(n, [&](size_t lo, size_t hi) {   // bit-reversal permutation (each swap is owned by its smaller index)
    for (size_t i = lo; i < hi; i++) {
      size_t j = 0;
      for (int b = 0; b < log_n; b++) j |= ((i >> b) & 1) << (log_n - 1 - b);
      if (i < j) std::swap(a[i], a[j]);
    }
  }, 4096);
  std::vector<Fr> roots(n / 2 ? n / 2 : 1);
  roots[0] = Fr::one();
  parallel_chunks(n / 2, [&](size_t lo, size_t hi) {   // the root table in parallel chunks, as ark-poly's roots_of_unity does
    u64 e[1] = {lo};
    Fr p = omega.pow(e, 1);
    for (size_t i = lo; i < hi; i++) { roots[i] = p; p = p * omega; }
  }, 4096);
  for (size_t m = 1; m < n; m *= 2) {
    size_t stride = n / (2 * m);
    parallel_chunks(n / 2, [&](size_t lo, size_t hi) {
      for (size_t t = lo; t < hi; t++) {
        size_t k = (t / m) * 2 * m, j = t % m;
        Fr u = a[k + j], v = a[k + j + m] * roots[j * stride];
        a[k + j] = u + v;
        a[k + j + m] = u - v;
      }
    }, 4096);
  }
}
static void ntt_fr(Fr* a, int log_n, bool inverse, bool coset) {
  size_t n = (size_t)1 << log_n;
  Fr g = Fr::from_u64(7);
  if (coset && !inverse) distribute_powers(a, n, g);
  Fr w = fr_root_of_unity(log_n);
  if (inverse) w = w.inv();
  fft_in_place(a, log_n, w);
  if (inverse) {
    Fr ninv = Fr::from_u64(n).inv();
    parallel_chunks(n, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) a[i] = a[i] * ninv; }, 4096);
    if (coset) distribute_powers(a, n, g.inv());
  }
}

// ------------------------------------------------------------------ R1CS -> QAP witness map (ark-groth16 r1cs_to_qap.rs)
struct Csr { const u64* row_ptr; const uint32_t* col; const Fr* val; };
static Fr row_dot(const Csr& m, size_t i, const Fr* z) {
  Fr acc = Fr::zero();
  Fr one = Fr::one();
  for (u64 k = m.row_ptr[i]; k < m.row_ptr[i + 1]; k++) {
    if (m.val[k] == one) acc = acc + z[m.col[k]];
    else acc = acc + z[m.col[k]] * m.val[k];
  }
  return acc;
}
static size_t domain_size(u64 nc, u64 ni) { size_t n = 1; while (n < nc + ni) n *= 2; return n; }
static void witness_map(u64 ni, u64 nc, const Csr& A, const Csr& B, const Csr& C, const Fr* z, Fr* h /* n */) {
  size_t n = domain_size(nc, ni);
  int log_n = ark_log2_ceil(n);
  std::vector<Fr> a(n, Fr::zero()), b(n, Fr::zero()), c(n, Fr::zero());
  parallel_chunks(nc, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) { a[i] = row_dot(A, i, z); b[i] = row_dot(B, i, z); } }, 1024);
  for (u64 j = 0; j < ni; j++) a[nc + j] = z[j];
  ntt_fr(a.data(), log_n, true, false);
  ntt_fr(b.data(), log_n, true, false);
  ntt_fr(a.data(), log_n, false, true);
  ntt_fr(b.data(), log_n, false, true);
  parallel_chunks(n, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) a[i] = a[i] * b[i]; }, 4096);
  parallel_chunks(nc, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) c[i] = row_dot(C, i, z); }, 1024);
  ntt_fr(c.data(), log_n, true, false);
  ntt_fr(c.data(), log_n, false, true);
  u64 e[1] = {n};
  Fr zinv = (Fr::from_u64(7).pow(e, 1) - Fr::one()).inv();
  parallel_chunks(n, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) a[i] = (a[i] - c[i]) * zinv; }, 4096);
  ntt_fr(a.data(), log_n, true, true);
  memcpy((void*)h, a.data(), n * sizeof(Fr));
}

// ------------------------------------------------------------------ prover (ark-groth16 prover.rs)
typedef Affine<Fq> G1A;
typedef Affine<Fq2> G2A;
typedef Jac<Fq> G1J;
typedef Jac<Fq2> G2J;

struct PkView {  // mirrors ctg_pk_view in include/ctg_groth16.h
  const G1A *alpha_g1, *beta_g1, *delta_g1;
  const G2A *beta_g2, *delta_g2;
  const G1A* a_query; u64 a_len;
  const G1A* b_g1_query; u64 b_g1_len;
  const G2A* b_g2_query; u64 b_g2_len;
  const G1A* h_query; u64 h_len;
  const G1A* l_query; u64 l_len;
};

static void to_canon_vec(const Fr* in, size_t n, std::vector<u64>& out) {
  out.resize(4 * n);
  parallel_chunks(n, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) in[i].to_canonical(&out[4 * i]); }, 4096);
}
template <class F> static Jac<F> calculate_coeff(const Jac<F>& initial, const Affine<F>* query, u64 qlen, const Affine<F>& vk_param,
                                                 const std::vector<u64>& z1 /* canonical z[1..] */) {
  size_t n = std::min<size_t>(qlen - 1, z1.size() / 4);
  Jac<F> acc = msm_ark<F>(query + 1, z1.data(), n);
  Jac<F> res = jadd_mixed(initial, query[0]);
  res = jadd(res, acc);
  res = jadd_mixed(res, vk_param);
  return res;
}
static void create_proof(const PkView& pk, u64 ni, u64 nw, u64 nc, const Csr& A, const Csr& B, const Csr& C, const Fr* z, const Fr& r,
                         const Fr& s, G1A* out_a, G2A* out_b, G1A* out_c) {
  size_t n = domain_size(nc, ni);
  std::vector<Fr> h(n);
  witness_map(ni, nc, A, B, C, z, h.data());
  std::vector<u64> hc, aux, z1;
  to_canon_vec(h.data(), n, hc);
  G1J h_acc = msm_ark<Fq>(pk.h_query, hc.data(), std::min<size_t>(pk.h_len, n));
  to_canon_vec(z + ni, nw, aux);
  G1J l_acc = msm_ark<Fq>(pk.l_query, aux.data(), std::min<size_t>(pk.l_len, nw));
  to_canon_vec(z + 1, ni + nw - 1, z1);
  u64 rc[4], sc[4], rsc[4];
  r.to_canonical(rc); s.to_canonical(sc); (r * s).to_canonical(rsc);
  G1J delta1 = from_affine(*pk.delta_g1);
  G1J g_a = calculate_coeff<Fq>(jmul(delta1, rc), pk.a_query, pk.a_len, *pk.alpha_g1, z1);
  G1J g1_b = r.is_zero() ? G1J::inf() : calculate_coeff<Fq>(jmul(delta1, sc), pk.b_g1_query, pk.b_g1_len, *pk.beta_g1, z1);
  G2J g2_b = calculate_coeff<Fq2>(jmul(from_affine(*pk.delta_g2), sc), pk.b_g2_query, pk.b_g2_len, *pk.beta_g2, z1);
  G1J g_c = jadd(jmul(g_a, sc), jmul(g1_b, rc));
  g_c = jadd(g_c, jneg(jmul(delta1, rsc)));
  g_c = jadd(g_c, l_acc);
  g_c = jadd(g_c, h_acc);
  *out_a = to_affine(g_a);
  *out_b = to_affine(g2_b);
  *out_c = to_affine(g_c);
}

// ------------------------------------------------------------------ key material helpers (generator.rs restated + synthetic progressions)
// fixed-base: out[i] = k_i * G  (4-bit windows over a precomputed table; cold path)
template <class F> static void fixed_base_mul(const Affine<F>& g, const Fr* k, size_t n, Affine<F>* out) {
  const int W = 8, NW = 32;  // 8-bit windows
  std::vector<Affine<F>> table((size_t)NW * 255);
  {
    Jac<F> base = from_affine(g);
    std::vector<Jac<F>> tj((size_t)NW * 255);
    for (int w = 0; w < NW; w++) {
      Jac<F> acc = base;
      for (int d = 1; d <= 255; d++) { tj[(size_t)w * 255 + d - 1] = acc; acc = jadd(acc, base); }
      for (int b = 0; b < W; b++) base = jdbl(base);
    }
    batch_to_affine(tj.data(), table.data(), tj.size());
  }
  parallel_chunks(n, [&](size_t lo, size_t hi) {
    std::vector<Jac<F>> tmp(hi - lo);
    for (size_t i = lo; i < hi; i++) {
      u64 c[4]; k[i].to_canonical(c);
      Jac<F> acc = Jac<F>::inf();
      for (int w = 0; w < NW; w++) {
        unsigned d = (c[w / 8] >> (8 * (w % 8))) & 0xff;
        if (d) acc = jadd_mixed(acc, table[(size_t)w * 255 + d - 1]);
      }
      tmp[i - lo] = acc;
    }
    batch_to_affine(tmp.data(), out + lo, hi - lo);
  }, 64);
}
// out[i] = (a + i*b) * G via chunked scalar-mul + repeated mixed addition of b*G
template <class F> static void progression(const Affine<F>& g, const Fr& a, const Fr& b, size_t n, Affine<F>* out) {
  u64 bc[4]; b.to_canonical(bc);
  Affine<F> step = to_affine(jmul(from_affine(g), bc));
  parallel_chunks(n, [&](size_t lo, size_t hi) {
    const size_t CH = 1024;
    std::vector<Jac<F>> tmp(CH);
    for (size_t s0 = lo; s0 < hi; s0 += CH) {
      size_t e0 = std::min(hi, s0 + CH);
      Fr k = a + b * Fr::from_u64(s0);
      u64 kc[4]; k.to_canonical(kc);
      Jac<F> acc = jmul(from_affine(g), kc);
      for (size_t i = s0; i < e0; i++) { tmp[i - s0] = acc; acc = jadd_mixed(acc, step); }
      batch_to_affine(tmp.data(), out + s0, e0 - s0);
    }
  }, 1024);
}

// ------------------------------------------------------------------ C ABI (ctypes)
static bool g_init = false;
static void ensure_init() { if (!g_init) { Fr::init(FR_MOD); Fq::init(FQ_MOD); g_init = true; } }

extern "C" {
void ark_set_threads(int t) { g_threads = t; }
int ark_get_threads() { return n_threads(); }
void ark_fr_mul_batch(const Fr* a, const Fr* b, Fr* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i] * b[i]; }
void ark_fq_mul_batch(const Fq* a, const Fq* b, Fq* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i] * b[i]; }
void ark_fr_add_batch(const Fr* a, const Fr* b, Fr* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i] + b[i]; }
void ark_fr_sub_batch(const Fr* a, const Fr* b, Fr* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i] - b[i]; }
void ark_fq_add_batch(const Fq* a, const Fq* b, Fq* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i] + b[i]; }
void ark_fq_sub_batch(const Fq* a, const Fq* b, Fq* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i] - b[i]; }
void ark_fq2_mul_batch(const Fq2* a, const Fq2* b, Fq2* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i] * b[i]; }
void ark_fr_to_mont(const u64* canon, Fr* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = Fr::from_canonical(canon + 4 * i); }
void ark_fr_from_mont(const Fr* in, u64* canon, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) in[i].to_canonical(canon + 4 * i); }
void ark_fq_to_mont(const u64* canon, Fq* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = Fq::from_canonical(canon + 6 * i); }
void ark_fq_from_mont(const Fq* in, u64* canon, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) in[i].to_canonical(canon + 6 * i); }
void ark_fr_sqr_batch(const Fr* a, Fr* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i].sqr(); }
void ark_fq_sqr_batch(const Fq* a, Fq* out, u64 n) { ensure_init(); for (u64 i = 0; i < n; i++) out[i] = a[i].sqr(); }
// out[j] = sum_i coeffs[i] * points[j]^i by Horner's rule, one thread per point: a check of the NTTs that is not an NTT
void ark_fr_poly_eval(const Fr* coeffs, u64 n, const Fr* points, u64 k, Fr* out) {
  ensure_init();
  parallel_tasks(k, [&](size_t j) {
    Fr acc = Fr::zero();
    const Fr x = points[j];
    for (u64 i = n; i-- > 0;) acc = acc * x + coeffs[i];
    out[j] = acc;
  });
}
void ark_ntt_fr(Fr* data, int log_n, int inverse, int coset) { ensure_init(); ntt_fr(data, log_n, inverse != 0, coset != 0); }
// scalars canonical (BigInteger256 images); out affine
void ark_msm_g1(const G1A* bases, const u64* scalars, u64 n, G1A* out) { ensure_init(); *out = to_affine(msm_ark<Fq>(bases, scalars, n)); }
void ark_msm_g2(const G2A* bases, const u64* scalars, u64 n, G2A* out) { ensure_init(); *out = to_affine(msm_ark<Fq2>(bases, scalars, n)); }
void ark_g1_mul(const G1A* p, const u64* k_canon, G1A* out) { ensure_init(); *out = to_affine(jmul(from_affine(*p), k_canon)); }
void ark_g2_mul(const G2A* p, const u64* k_canon, G2A* out) { ensure_init(); *out = to_affine(jmul(from_affine(*p), k_canon)); }
void ark_g1_add(const G1A* p, const G1A* q, G1A* out) { ensure_init(); *out = to_affine(jadd_mixed(from_affine(*p), *q)); }
void ark_g2_add(const G2A* p, const G2A* q, G2A* out) { ensure_init(); *out = to_affine(jadd_mixed(from_affine(*p), *q)); }
int ark_g1_on_curve(const G1A* p) { ensure_init(); if (p->is_inf()) return 1; return p->y.sqr() == p->x.sqr() * p->x + Fq::from_u64(4); }
int ark_g2_on_curve(const G2A* p) { ensure_init(); if (p->is_inf()) return 1; Fq2 b = {Fq::from_u64(4), Fq::from_u64(4)}; return p->y.sqr() == p->x.sqr() * p->x + b; }
void ark_g1_fixed_base(const G1A* g, const Fr* k, u64 n, G1A* out) { ensure_init(); fixed_base_mul<Fq>(*g, k, n, out); }
void ark_g2_fixed_base(const G2A* g, const Fr* k, u64 n, G2A* out) { ensure_init(); fixed_base_mul<Fq2>(*g, k, n, out); }
void ark_g1_progression(const G1A* g, const Fr* a, const Fr* b, u64 n, G1A* out) { ensure_init(); progression<Fq>(*g, *a, *b, n, out); }
void ark_g2_progression(const G2A* g, const Fr* a, const Fr* b, u64 n, G2A* out) { ensure_init(); progression<Fq2>(*g, *a, *b, n, out); }
void ark_witness_map(u64 ni, u64 nc, const u64* a_ptr, const uint32_t* a_col, const Fr* a_val, const u64* b_ptr, const uint32_t* b_col,
                     const Fr* b_val, const u64* c_ptr, const uint32_t* c_col, const Fr* c_val, const Fr* z, Fr* h) {
  ensure_init();
  witness_map(ni, nc, Csr{a_ptr, a_col, a_val}, Csr{b_ptr, b_col, b_val}, Csr{c_ptr, c_col, c_val}, z, h);
}
void ark_create_proof(const PkView* pk, u64 ni, u64 nw, u64 nc, const u64* a_ptr, const uint32_t* a_col, const Fr* a_val, const u64* b_ptr,
                      const uint32_t* b_col, const Fr* b_val, const u64* c_ptr, const uint32_t* c_col, const Fr* c_val, const Fr* z,
                      const Fr* r, const Fr* s, G1A* out_a, G2A* out_b, G1A* out_c) {
  ensure_init();
  create_proof(*pk, ni, nw, nc, Csr{a_ptr, a_col, a_val}, Csr{b_ptr, b_col, b_val}, Csr{c_ptr, c_col, c_val}, z, *r, *s, out_a, out_b, out_c);
}
// lagrange coefficients at tau (evaluate_all_lagrange_coefficients), all Montgomery
void ark_lagrange_coeffs(int log_n, const Fr* tau, Fr* out) {
  ensure_init();
  size_t n = (size_t)1 << log_n;
  u64 e[1] = {n};
  Fr zt = tau->pow(e, 1) - Fr::one();
  Fr w = fr_root_of_unity(log_n), c = zt * Fr::from_u64(n).inv();
  // out[i] = zt/n * w^i / (tau - w^i); batch inversion
  std::vector<Fr> den(n), pref(n);
  Fr wi = Fr::one(), acc = Fr::one();
  for (size_t i = 0; i < n; i++) { den[i] = *tau - wi; pref[i] = acc; acc = acc * den[i]; out[i] = c * wi; wi = wi * w; }
  Fr inv = acc.inv();
  for (size_t i = n; i-- > 0;) { out[i] = out[i] * (inv * pref[i]); inv = inv * den[i]; }
}
// ---- Poseidon (ark-sponge 0.3.0 PoseidonSponge, rate 2 / capacity 1) and the ciphertext Merkle tree -- the CPU side of
// SURVEY 8(f) row 4 (circuits-ark/src/sample_entries.rs:127-134, encryption.rs:154-172); mirrors oracle/py/poseidon.py, which
// holds the conventions, with machine-word Montgomery arithmetic and all host threads (rayon would do the same per level).
struct PoseidonCfg { int full, partial; u64 alpha; const Fr* mds; const Fr* ark; };
static inline Fr poseidon_sbox(const Fr& x, u64 alpha) {   // x^alpha, square-and-multiply from the top set bit (alpha >= 1)
  Fr acc = x;
  for (int b = 62 - __builtin_clzll(alpha); b >= 0; b--) { acc = acc.sqr(); if ((alpha >> b) & 1) acc = acc * x; }
  return acc;
}
static void poseidon_permute(const PoseidonCfg& c, Fr st[3]) {
  const int half = c.full / 2, rounds = c.full + c.partial;
  for (int r = 0; r < rounds; r++) {
    for (int i = 0; i < 3; i++) st[i] = st[i] + c.ark[3 * r + i];
    if (r < half || r >= half + c.partial) { for (int i = 0; i < 3; i++) st[i] = poseidon_sbox(st[i], c.alpha); }
    else st[0] = poseidon_sbox(st[0], c.alpha);
    Fr nw[3];
    for (int i = 0; i < 3; i++) nw[i] = c.mds[3 * i] * st[0] + c.mds[3 * i + 1] * st[1] + c.mds[3 * i + 2] * st[2];
    for (int i = 0; i < 3; i++) st[i] = nw[i];
  }
}
// fresh sponge, absorb k elements, squeeze one
static Fr poseidon_crh(const PoseidonCfg& c, const Fr* in, u64 k) {
  Fr st[3] = {Fr::zero(), Fr::zero(), Fr::zero()};
  u64 pos = 0;
  for (u64 e = 0; e < k; e++) {
    if (pos == 2) { poseidon_permute(c, st); pos = 0; }
    st[pos] = st[pos] + in[e];
    pos++;
  }
  poseidon_permute(c, st);
  return st[0];
}
void ark_poseidon_hash_batch(int full, int partial, u64 alpha, const Fr* mds, const Fr* ark, const Fr* in, u64 n, u64 k, Fr* out) {
  ensure_init();
  PoseidonCfg c{full, partial, alpha, mds, ark};
  parallel_chunks(n, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) out[i] = poseidon_crh(c, in + i * k, k); });
}
void ark_poseidon_merkle_tree(int full, int partial, u64 alpha, const Fr* mds, const Fr* ark, const Fr* leaves, u64 n, Fr* leaf_nodes, Fr* non_leaf) {
  ensure_init();
  PoseidonCfg c{full, partial, alpha, mds, ark};
  parallel_chunks(n, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) leaf_nodes[i] = poseidon_crh(c, leaves + i, 1); });
  const Fr* children = leaf_nodes;
  for (u64 width = n / 2; width >= 1; width /= 2) {
    Fr* parents = non_leaf + (width - 1);
    parallel_chunks(width, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) parents[i] = poseidon_crh(c, children + 2 * i, 2); });
    children = parents;
  }
}
}
