"""TEST INFRASTRUCTURE ONLY -- big-int restatement of the Poseidon duplex sponge, the Poseidon CRHs and the Merkle
tree the reference builds over a ciphertext.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg
may import this.

PARITY UNPINNED: ark-sponge ^0.3.0 (crates.io) and ark-crypto-primitives (git, branch main, no pinned revision:
circuits-ark/Cargo.toml:18-19) are absent from /root/reference, and the reference holds no hash or root value to
compare with (its Merkle test, circuits-ark/src/sample_entries.rs:263-290, only asserts satisfiability).  Restated:

  ark-sponge 0.3.0 poseidon/mod.rs
    PoseidonParameters::new(full_rounds, partial_rounds, alpha, mds, ark): rate 2, capacity 1 (state of 3 cells)
    permute: for each round r: state[i] += ark[r][i]; S-box x^alpha on every cell (full round) or on cell 0 only
      (partial round); state = mds * state; rounds are full_rounds/2 full, partial_rounds partial, full_rounds/2 full
    absorb_internal(start, elems): add elems into state[start..], permuting and restarting at cell 0 whenever the
      rate is used up and elements remain; mode becomes Absorbing{next = ...}
    squeeze_internal(start, out): copy state[start..rate] out, permuting whenever the rate is used up and more
      is wanted; mode becomes Squeezing{next = ...}
    absorb when Absorbing at next == rate: permute first; absorb when Squeezing: permute, start at 0
    squeeze when Absorbing: permute, start at 0; squeeze when Squeezing at next == rate: permute, start at 0
  ark-crypto-primitives crh/poseidon: CRH::evaluate(params, input) = new sponge, absorb(input), squeeze 1;
    TwoToOneCRH::evaluate(l, r) = CRH::evaluate([l, r])
  ark-crypto-primitives merkle_tree: MerkleTree::new(leaf_params, two_to_one_params, leaves): leaf digests
    CRH(leaf), parents TwoToOneCRH(left, right), non-leaf nodes stored in heap order (root at 0), leaf count a
    power of two >= 2; Path = {leaf_sibling_hash, auth_path (root side first), leaf_index}
Reference call sites: circuits-ark/src/sample_entries.rs:127-134 (build_merkle_tree, one element per leaf),
circuits-ark/src/encryption.rs:154-172 (encrypt: absorb the affine point p_ra = (x, y), squeeze 1).
"""
from __future__ import annotations

import random
from dataclasses import dataclass

from .bls12_381 import R_MOD

RATE, CAPACITY = 2, 1


@dataclass
class PoseidonParameters:
    full_rounds: int
    partial_rounds: int
    alpha: int
    mds: list      # 3 x 3
    ark: list      # (full + partial) x 3


def random_parameters(seed, full_rounds=8, partial_rounds=31, alpha=17):
    """Parameters of the reference's shape (circuits-ark/src/parameters.rs:226-229) with seeded random constants --
    the caller of the library supplies the real ones; nothing depends on their values."""
    rnd = random.Random(seed)
    mds = [[rnd.randrange(R_MOD) for _ in range(3)] for _ in range(3)]
    ark = [[rnd.randrange(R_MOD) for _ in range(3)] for _ in range(full_rounds + partial_rounds)]
    return PoseidonParameters(full_rounds, partial_rounds, alpha, mds, ark)


class PoseidonSponge:
    def __init__(self, params):
        self.p = params
        self.state = [0] * (RATE + CAPACITY)
        self.mode = ("absorbing", 0)

    def permute(self):
        p, st = self.p, list(self.state)
        half = p.full_rounds // 2
        for r in range(p.full_rounds + p.partial_rounds):
            st = [(s + c) % R_MOD for s, c in zip(st, p.ark[r])]
            if r < half or r >= half + p.partial_rounds:
                st = [pow(s, p.alpha, R_MOD) for s in st]
            else:
                st[0] = pow(st[0], p.alpha, R_MOD)
            st = [sum(m * s for m, s in zip(row, st)) % R_MOD for row in p.mds]
        self.state = st

    def _absorb_internal(self, start, elems):
        elems = list(elems)
        while True:
            if start + len(elems) <= RATE:
                for i, e in enumerate(elems):
                    self.state[start + i] = (self.state[start + i] + e) % R_MOD
                self.mode = ("absorbing", start + len(elems))
                return
            take = RATE - start
            for i, e in enumerate(elems[:take]):
                self.state[start + i] = (self.state[start + i] + e) % R_MOD
            self.permute()
            elems = elems[take:]
            start = 0

    def _squeeze_internal(self, start, count):
        out = []
        while True:
            if start + count - len(out) <= RATE:
                need = count - len(out)
                out += self.state[start:start + need]
                self.mode = ("squeezing", start + need)
                return out
            out += self.state[start:RATE]
            self.permute()
            start = 0

    def absorb(self, elems):
        elems = list(elems)
        if not elems:
            return
        kind, nxt = self.mode
        if kind == "absorbing":
            if nxt == RATE:
                self.permute()
                nxt = 0
            self._absorb_internal(nxt, elems)
        else:
            self.permute()
            self._absorb_internal(0, elems)

    def squeeze(self, count):
        kind, nxt = self.mode
        if kind == "absorbing" or nxt == RATE:
            self.permute()
            nxt = 0
        return self._squeeze_internal(nxt, count)


def crh(params, inputs):
    s = PoseidonSponge(params)
    s.absorb(inputs)
    return s.squeeze(1)[0]


def two_to_one(params, left, right):
    return crh(params, [left, right])


class MerkleTree:
    def __init__(self, params, leaves):
        n = len(leaves)
        assert n >= 2 and n & (n - 1) == 0, "leaf count must be a power of two >= 2"
        self.params = params
        self.leaf_nodes = [crh(params, [e]) for e in leaves]
        self.non_leaf_nodes = [0] * (n - 1)
        level = self.leaf_nodes
        width = n // 2
        while width >= 1:
            parents = [two_to_one(params, level[2 * i], level[2 * i + 1]) for i in range(width)]
            self.non_leaf_nodes[width - 1:2 * width - 1] = parents
            level = parents
            width //= 2

    def root(self):
        return self.non_leaf_nodes[0]

    def generate_proof(self, index):
        """(leaf_sibling_hash, auth_path root-side first, index) as arkworks' Path."""
        n = len(self.leaf_nodes)
        sibling = self.leaf_nodes[index ^ 1]
        node = (n - 1 + index - 1) // 2          # heap index of the leaf's parent
        path = []
        while node != 0:
            sib = node + 1 if node % 2 == 1 else node - 1
            path.append(self.non_leaf_nodes[sib])
            node = (node - 1) // 2
        return sibling, path[::-1], index


def verify_path(params, root, leaf, proof):
    """Path::verify: hash the leaf, fold with the sibling by the index bits, compare with the root."""
    sibling, path, index = proof
    cur = crh(params, [leaf])
    cur = two_to_one(params, cur, sibling) if index % 2 == 0 else two_to_one(params, sibling, cur)
    index //= 2
    for sib in reversed(path):
        cur = two_to_one(params, cur, sib) if index % 2 == 0 else two_to_one(params, sib, cur)
        index //= 2
    return cur == root
