"""-m gpu: the proof module (same names as InferNet/template/validator/proof.py) through the C-ABI kernels against the
golden vectors recorded from the reference itself (tests/golden/proof_golden.json) and against the oracle.
Bit-exact for roots / paths; identical booleans for every spot-check verdict (scheduler branch = what the reference
computes when diffusers is importable)."""
import base64
import hashlib
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "proof_golden.json")))


def d64(s):
    return base64.b64decode(s)


def test_build_merkle_root_and_verify_golden(cuda):
    from b2v import proof
    for case in G["merkle"]:
        leaves = [d64(x) for x in case["leaves"]]
        root, proofs = proof.build_merkle_root(leaves)
        assert root.hex() == case["root"]
        assert [[p.hex() for p in pr] for pr in proofs] == case["proofs"]
        got = [proof.verify_merkle_leaf(hashlib.sha256(l).digest(), proofs[i], root) for i, l in enumerate(leaves)]
        assert got == case["verify"]
        bad = [proof.verify_merkle_leaf(hashlib.sha256(l + b"x").digest(), proofs[i], root) for i, l in enumerate(leaves)]
        assert bad == case["verify_tampered"]
        assert proof.verify_merkle_leaves([hashlib.sha256(l).digest() for l in leaves], proofs, root) == case["verify"]
    assert proof.build_merkle_root([]) == (b"", [])


def test_verify_merkle_leaf_odd_inputs(cuda):
    from b2v import proof
    from oracle import merkle_ref as M
    for leaf, path, root in [(b"ab", [b"c" * 32], b"d" * 32), (b"a" * 32, [b"b" * 5, b"c" * 32], b"d" * 32),
                             (b"a" * 32, [], b"a" * 32), (b"a" * 32, [], b"b" * 32)]:
        assert proof.verify_merkle_leaf(leaf, path, root) == M.verify_merkle_leaf(leaf, path, root)
    want_root = M._parent(M._parent(b"ab", b"c" * 32), b"x")
    assert proof.verify_merkle_leaf(b"ab", [b"c" * 32, b"x"], want_root) is True


def test_miner_build_merkle_tree_golden(cuda):
    from b2v import proof
    for c in G["miner_commit"]:
        g = torch.Generator().manual_seed(c["seed"])
        shape = tuple(c["shape"])
        zs = [torch.randn(shape, generator=g).half() for _ in range(c["n"])]
        es = [torch.randn(shape, generator=g).half() for _ in range(c["n"])]
        root, leaf = proof.build_merkle_tree([z.to(cuda) for z in zs], [e.to(cuda) for e in es], c["timesteps"])
        assert root.hex() == c["root"]
        assert {str(t): [p.hex() for p in leaf[t][2]] for t in c["timesteps"]} == c["paths"]
        t0 = c["timesteps"][0]
        assert leaf[t0][0] == zs[0].numpy().tobytes() and leaf[t0][1] == es[0].numpy().tobytes()
        # every revealed leaf verifies against the root (validator.py:799-802)
        for t in c["timesteps"]:
            lh = hashlib.sha256(int(t).to_bytes(2, "big") + leaf[t][0] + leaf[t][1]).digest()
            assert proof.verify_merkle_leaf(lh, leaf[t][2], root)


def test_run_unet_step_and_coherence_verdicts_golden(cuda):
    from b2v import proof
    cfg = G["unet_config"]
    for rec in G["ddim_checks"]:
        zb, eb = [d64(x) for x in rec["z"]], [d64(x) for x in rec["eps"]]
        c = dict(cfg, alphas=rec["alphas"])
        got = [proof.run_unet_step(zb[i], eb[i], c, i) for i in range(8)]
        assert got == rec["run_unet_step"]["scheduler"], (rec["kind"], rec["shape"])
        ts = rec["coherence"]["all"]["indices"]
        leaves = {str(t): (rec["z"][i], rec["eps"][i], []) for i, t in enumerate(ts)}
        t2s = {t: i for i, t in enumerate(ts)}
        for name, v in rec["coherence"].items():
            got = proof.verify_temporal_coherence_spot_check(v["indices"], leaves, t2s, rec["alphas"], cfg, "t")
            assert got == v["scheduler"], (rec["kind"], rec["shape"], name)
    for m in G["ddim_malformed"]:
        c = dict(cfg, alphas=m["alphas"])
        assert proof.run_unet_step(d64(m["z"]), d64(m["eps"]), c, m["step"]) == m["scheduler"], m["name"]


def test_spotcheck_fuzz_vs_oracle(cuda):
    """random fp16 leaves incl. large magnitudes / infs: device verdicts == torch-CPU fp16 verdicts of the oracle."""
    from b2v import proof
    from oracle import merkle_ref as M
    cfg = {"latent_channels": 4, "latent_height": 6, "latent_width": 10}
    g = torch.Generator().manual_seed(123)
    n = 50
    c = dict(cfg, alphas=[0.5] * n)
    mism = 0
    for i in range(n):
        scale = [0.1, 1.0, 3.0, 9.0, 12.0, 100.0, 3e4][i % 7]
        z = (torch.randn(4, 2, 6, 10, generator=g) * scale).half()
        e = (torch.randn(4, 2, 6, 10, generator=g) * (scale if i % 3 else 1.0)).half()
        zb, eb = z.numpy().tobytes(), e.numpy().tobytes()
        mism += proof.run_unet_step(zb, eb, c, i) != M.run_unet_step_scheduler(zb, eb, c, i)
    assert mism == 0


def test_verify_proof_of_inference_end_to_end(cuda):
    from b2v import proof
    from oracle.ddim_ref import DDIMSchedulerRef
    # honest miner trajectory under the verifier's (default) scheduler, 8 steps, 3 frames
    g = torch.Generator().manual_seed(4)
    s = DDIMSchedulerRef()
    s.set_timesteps(8)
    shape = (1, 4, 3, 8, 8)
    z = torch.randn(shape, generator=g).half()
    zs, es, ts = [], [], []
    for t in s.timesteps:
        e = (0.5 * torch.randn(shape, generator=g)).half()
        zs.append(z); es.append(e); ts.append(int(t))
        z = s.step(e, t, z).prev_sample
    root, leaf = proof.build_merkle_tree([x.to(cuda) for x in zs], [x.to(cuda) for x in es], ts)
    reveal = proof.commit_then_reveal_merkle_spotcheck(root, len(ts), 1, 42)
    # NB proof.py maps timesteps to step indices by *position in leaf_data*, so reveal a prefix-aligned pair
    reveal = [0, 1]
    b64 = lambda b: base64.b64encode(b).decode()
    leaf_data = {str(ts[i]): (b64(leaf[ts[i]][0]), b64(leaf[ts[i]][1]), [b64(p) for p in leaf[ts[i]][2]]) for i in reveal}
    video = b"mp4-bytes"
    pr = {"challenge": b64(b"C" * 32), "seed": 99, "merkle_root": b64(root), "signature": b64(b"sig"),
          "leaf_data": leaf_data, "alphas": [0.5] * 8}
    ucfg = {"latent_channels": 4, "latent_height": 8, "latent_width": 8, "alphas": [0.5] * 8}
    seen = {}

    def verifier(hotkey, message, signature):
        seen["msg"] = message
        return signature == b"sig"
    assert proof.verify_proof_of_inference("5Fhot", video, pr, ucfg, signature_verifier=verifier) is True
    assert seen["msg"] == b"C" * 32 + (99).to_bytes(8, "little") + hashlib.sha256(video).digest() + root
    bad = dict(pr, signature=b64(b"nope"))
    assert proof.verify_proof_of_inference("5Fhot", video, bad, ucfg, signature_verifier=verifier) is False
    tam = json.loads(json.dumps(pr))
    k = str(ts[1])
    zt = bytearray(base64.b64decode(tam["leaf_data"][k][0])); zt[0] ^= 1
    tam["leaf_data"][k][0] = b64(bytes(zt))
    assert proof.verify_proof_of_inference("5Fhot", video, tam, ucfg, signature_verifier=verifier) is False
    assert proof.verify_proof_of_inference("5Fhot", video, {"challenge": "x"}, ucfg, signature_verifier=verifier) is False
