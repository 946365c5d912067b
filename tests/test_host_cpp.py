"""The compiled host-side mirror of the ark-groth16 API (contangle-zkcp_b200/host/ctg_groth16.hpp) exercised by a C++
program (tests/host_cpp/roundtrip.cpp): setup -> circuit.zkey / verification.key -> read key -> prove -> proof bytes.
CPU suite: it builds, links the C ABI and refuses to run without a device.  GPU suite: its outputs are checked against the
ctypes path and the pairing oracle."""
import os
import struct
import subprocess

import numpy as np
import pytest

import ctg
from ctg import synthetic as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HERE = os.path.join(ROOT, "tests", "host_cpp")
R_WORDS, S_WORDS = S.fr_mont([0x1234567])[0], S.fr_mont([0x7654321])[0]


def _build():
    subprocess.check_call(["make", "-C", HERE], stdout=subprocess.DEVNULL)
    return os.path.join(HERE, "roundtrip")


def _write_input(path, r1cs, zm, td_words):
    ni, nw, nc = r1cs["num_instance"], r1cs["num_witness"], r1cs["num_constraints"]
    with open(path, "wb") as f:
        f.write(struct.pack("<QQQ", ni, nw, nc))
        for k in "abc":
            m = r1cs[k]
            f.write(struct.pack("<Q", len(m["col"])))
            f.write(np.ascontiguousarray(m["row_ptr"], dtype=np.uint64).tobytes())
            f.write(np.ascontiguousarray(m["col"], dtype=np.uint32).tobytes())
            f.write(np.ascontiguousarray(m["val"], dtype=np.uint32).tobytes())
        f.write(np.ascontiguousarray(zm, dtype=np.uint32).tobytes())
        for k in ("alpha", "beta", "gamma", "delta", "tau", "g1", "g2"):
            f.write(np.ascontiguousarray(td_words[k], dtype=np.uint32).tobytes())
        f.write(R_WORDS.tobytes()); f.write(S_WORDS.tobytes())


def test_host_cpp_builds_and_refuses_without_device(tmp_path):
    exe = _build()
    r1cs, zm, _ = S.mul_chain(4)
    _write_input(tmp_path / "in.bin", r1cs, zm, S.trapdoor()[0])
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    p = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path)], capture_output=True, text=True, timeout=300)
    if has_gpu:
        assert p.returncode == 0, p.stderr
    else:
        assert p.returncode == 3 and "no usable" in p.stderr, (p.returncode, p.stderr)   # fails loudly: no CPU fallback


@pytest.mark.gpu
@pytest.mark.parametrize("shape", ["mulchain7", "poe1"])
def test_host_cpp_round_trip_matches_ctypes_path_and_verifies(tmp_path, shape):
    from oracle import ark
    from oracle.py import serialize as SER
    from oracle.py import pairing as PR
    exe = _build()
    r1cs, zm, z = S.mul_chain(7) if shape == "mulchain7" else S.poe_shape(1, gates=150, bits=16)
    td_words, _ = S.trapdoor(seed=0xBEEF)
    _write_input(tmp_path / "in.bin", r1cs, zm, td_words)
    p = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr + p.stdout
    zkey, vk, proof_bytes = ((tmp_path / n).read_bytes() for n in ("circuit.zkey", "verification.key", "proof.bin"))
    # same objects through the ctypes path
    ctx = ctg.Context(0)
    dp = ctx.generate_parameters(r1cs, td_words)
    assert dp.serialize_unchecked() == zkey and dp.serialize_vk() == vk
    dpk, dr = dp.to_device_pk(), ctx.upload_r1cs(r1cs)
    proof = ctx.prove(dpk, dr, zm, R_WORDS, S_WORDS)
    assert ctg.proof_to_bytes(proof) == proof_bytes
    dpk.free(); dr.free(); dp.free(); ctx.close()
    # the verifier's view: only verification.key and the 192 proof bytes, both parsed by the oracle
    ni = r1cs["num_instance"]
    alpha = SER.point_from_compressed(vk[:48], 1)
    beta, gamma, delta = (SER.point_from_compressed(vk[48 + 96 * k:144 + 96 * k], 2) for k in range(3))
    (n_abc,) = struct.unpack("<Q", vk[336:344])
    assert n_abc == ni and len(vk) == 344 + 48 * ni
    abc = [SER.point_from_compressed(vk[344 + 48 * k:392 + 48 * k], 1) for k in range(ni)]
    assert PR.groth16_verify(alpha, beta, gamma, delta, abc, SER.proof_from_compressed(proof_bytes), z[1:ni])
    assert not PR.groth16_verify(alpha, beta, gamma, delta, abc, SER.proof_from_compressed(proof_bytes), [(z[1] + 1)] + list(z[2:ni]))
