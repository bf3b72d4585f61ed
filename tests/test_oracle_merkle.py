"""not-gpu: pins the Merkle / spot-check oracle (oracle/merkle_ref.py, oracle/sha256_ref.c) against
tests/golden/proof_golden.json — outputs of the reference's own proof.py and Miner._build_merkle_tree executed by
oracle/gen_golden.py — and against the FIPS 180-4 known answers."""
import base64
import ctypes
import hashlib
import json
import os
import subprocess

import pytest
import torch

from oracle import merkle_ref as M

HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "proof_golden.json")))


def d64(s):
    return base64.b64decode(s)


@pytest.fixture(scope="module")
def cref():
    root = os.path.dirname(HERE)
    so = os.path.join(root, "oracle", "libsha256_ref.so")
    src = os.path.join(root, "oracle", "sha256_ref.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O3", "-shared", "-fPIC", "-o", so, src])
    return ctypes.CDLL(so)


def c_sha(lib, prefix, a, b):
    out = ctypes.create_string_buffer(32)
    lib.ref_sha256_concat(prefix, ctypes.c_size_t(len(prefix)), a, ctypes.c_size_t(len(a)), b, ctypes.c_size_t(len(b)), out)
    return out.raw


def test_c_sha256_fips_vectors(cref):
    assert c_sha(cref, b"", b"abc", b"").hex() == "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"
    assert c_sha(cref, b"", b"", b"").hex() == "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"
    m = b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq"
    assert c_sha(cref, b"", m, b"").hex() == "248d6a61d20638b8e5c026930c3e6039a33ce45964ff2167f6ecedd419db06c1"
    assert c_sha(cref, b"", b"a" * 1000000, b"").hex() == "cdc76e5c9914fb9281a1c7e284d73e67f1809a48a497200e046d39ccc7112cd0"


@pytest.mark.parametrize("n", [0, 1, 55, 56, 63, 64, 65, 119, 120, 1000, 4099])
def test_c_sha256_vs_hashlib(cref, n):
    data = bytes((i * 7 + 3) % 256 for i in range(n))
    for cut in (0, n // 3, n):
        assert c_sha(cref, b"\x03\xd5", data[:cut], data[cut:]) == hashlib.sha256(b"\x03\xd5" + data).digest()


def test_c_merkle_levels_match_golden(cref):
    for case in G["merkle"]:
        leaves = [d64(x) for x in case["leaves"]]
        lh = b"".join(hashlib.sha256(x).digest() for x in leaves)
        buf = ctypes.create_string_buffer(32 * (2 * len(leaves) + 8))
        total = cref.ref_merkle_levels(lh, len(leaves), buf)
        assert buf.raw[32 * (total - 1):32 * total].hex() == case["root"]


def test_merkle_root_paths_verify_golden():
    for case in G["merkle"]:
        leaves = [d64(x) for x in case["leaves"]]
        root, proofs = M.build_merkle_root(leaves)
        assert root.hex() == case["root"]
        assert [[p.hex() for p in pr] for pr in proofs] == case["proofs"]
        assert [M.verify_merkle_leaf(hashlib.sha256(l).digest(), proofs[i], root) for i, l in enumerate(leaves)] == case["verify"]
        assert [M.verify_merkle_leaf(hashlib.sha256(l + b"x").digest(), proofs[i], root)
                for i, l in enumerate(leaves)] == case["verify_tampered"]
    assert M.build_merkle_root([]) == (bytes.fromhex(G["merkle_empty"]["root"]), G["merkle_empty"]["proofs"])


def test_survey_known_answers():
    # SURVEY.md §8c vectors (recorded from the reference in the survey session)
    assert M.build_merkle_root([bytes([i]) * 32 for i in range(8)])[0].hex() == \
        "888e60e20c7ec4899e1704f189074caef0c346ff8632b817f9a1b3b9d64bb5db"
    r5, p5 = M.build_merkle_root([bytes([i]) * 32 for i in range(5)])
    assert r5.hex() == "c05df5873e2f049edde60e779650076569c01a2a16fd00c8a07dcbed2fd97519"
    assert all(len(p) == 3 for p in p5)
    assert M.derive_seed(b"\x01" * 32, "5Fwallet") == 8605131804522081524
    assert M.spotcheck_indices(50, 5, 1234) == [28, 29, 7, 8, 0, 1, 5, 6, 37, 38]


def test_seed_and_spotcheck_golden():
    for c in G["derive_seed"]:
        assert M.derive_seed(d64(c["challenge"]), c["wallet"]) == c["seed"]
    for c in G["spotcheck"]:
        assert M.spotcheck_indices(c["n"], c["k"], c["seed"]) == c["indices"]


def test_miner_commit_golden():
    for c in G["miner_commit"]:
        g = torch.Generator().manual_seed(c["seed"])
        shape = tuple(c["shape"])
        zs = [torch.randn(shape, generator=g).half() for _ in range(c["n"])]
        es = [torch.randn(shape, generator=g).half() for _ in range(c["n"])]
        root, leaf = M.commit_trajectory(zs, es, c["timesteps"])
        assert root.hex() == c["root"]
        assert {str(t): [p.hex() for p in leaf[t][2]] for t in c["timesteps"]} == c["paths"]
        assert hashlib.sha256(leaf[c["timesteps"][0]][0]).hexdigest() == c["z0_sha"]
        # P1 == P2 on the concatenated raw leaves
        raw = [int(t).to_bytes(2, "big") + leaf[t][0] + leaf[t][1] for t in c["timesteps"]]
        assert M.build_merkle_root(raw)[0] == root
        assert M.commit_trajectory(zs, es, c["timesteps"], threads=4)[0] == root


def test_ddim_verdicts_golden_both_branches():
    cfg = G["unet_config"]
    for rec in G["ddim_checks"]:
        zb, eb = [d64(x) for x in rec["z"]], [d64(x) for x in rec["eps"]]
        c = dict(cfg, alphas=rec["alphas"])
        assert [M.run_unet_step_scheduler(zb[i], eb[i], c, i) for i in range(8)] == rec["run_unet_step"]["scheduler"]
        assert [M.run_unet_step_fallback(zb[i], eb[i], c, i) for i in range(8)] == rec["run_unet_step"]["fallback"]
        ts = rec["coherence"]["all"]["indices"]
        leaves = {str(t): (rec["z"][i], rec["eps"][i], []) for i, t in enumerate(ts)}
        t2s = {t: i for i, t in enumerate(ts)}
        for name, v in rec["coherence"].items():
            assert M.temporal_coherence(v["indices"], leaves, t2s, rec["alphas"], cfg) == v["scheduler"], (rec["kind"], name)
    for m in G["ddim_malformed"]:
        c = dict(cfg, alphas=m["alphas"])
        assert M.run_unet_step_scheduler(d64(m["z"]), d64(m["eps"]), c, m["step"]) == m["scheduler"], m["name"]
        assert M.run_unet_step_fallback(d64(m["z"]), d64(m["eps"]), c, m["step"]) == m["fallback"], m["name"]
