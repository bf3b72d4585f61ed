"""-m gpu: integer / byte kernels of the Proof-of-Inference path through the C-ABI: bit-exact against hashlib
(the reference's own primitive, InferNet/neurons/miner.py:207-233, template/validator/proof.py:250-304)."""
import hashlib

import pytest
import torch

pytestmark = pytest.mark.gpu


def _sorted_pair(l, r):
    return hashlib.sha256(l + r if l < r else r + l).digest()


def _tree(leaf_hashes):
    tree = [list(leaf_hashes)]
    while len(tree[-1]) > 1:
        cur = tree[-1]
        tree.append([_sorted_pair(cur[i], cur[i + 1] if i + 1 < len(cur) else cur[i]) for i in range(0, len(cur), 2)])
    paths = []
    for idx in range(len(leaf_hashes)):
        p, i = [], idx
        for level in tree[:-1]:
            s = i ^ 1
            p.append(level[s] if s < len(level) else level[i])
            i //= 2
        paths.append(p)
    return tree[-1][0], paths


@pytest.mark.parametrize("mode", [1, 2, 3])
@pytest.mark.parametrize("lens", [[0], [1], [55], [56], [63], [64], [65], [119], [120], [128], [1000, 3, 64, 0, 4096],
                                  [2048] * 7, [100003]])
def test_sha256_raw(cuda, mode, lens):
    from b2v import ops
    g = torch.Generator().manual_seed(sum(lens) + len(lens))
    bufs = [torch.randint(0, 256, (n,), dtype=torch.uint8, generator=g) for n in lens]
    dev = [b.to(cuda) if b.numel() else torch.empty(0, dtype=torch.uint8, device=cuda) for b in bufs]
    dig = ops.sha256_leaves(dev, mode=mode).cpu().numpy()
    for i, b in enumerate(bufs):
        assert dig[i].tobytes() == hashlib.sha256(b.numpy().tobytes()).digest(), (mode, lens[i])


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
@pytest.mark.parametrize("shape", [(1, 4, 2, 8, 8), (1, 4, 5, 32, 32), (1, 4, 3, 5, 7)])
def test_sha256_leaf_prefix_z_eps(cuda, mode, shape):
    """leaf = t.to_bytes(2,'big') + z_bytes + eps_bytes (miner.py:207), z/eps hashed in place on the device."""
    from b2v import ops
    g = torch.Generator().manual_seed(shape[2])
    ts = [981, 1, 65535, 0, 256]
    zs = [torch.randn(shape, generator=g).half() for _ in ts]
    es = [torch.randn(shape, generator=g).half() for _ in ts]
    dig = ops.sha256_leaves([z.to(cuda) for z in zs], [e.to(cuda) for e in es], prefixes=ts, mode=mode).cpu().numpy()
    for i, t in enumerate(ts):
        want = hashlib.sha256(t.to_bytes(2, "big") + zs[i].numpy().tobytes() + es[i].numpy().tobytes()).digest()
        assert dig[i].tobytes() == want


def test_sha256_unaligned_views(cuda):
    from b2v import ops
    g = torch.Generator().manual_seed(9)
    big = torch.randint(0, 256, (5000,), dtype=torch.uint8, generator=g)
    d = big.to(cuda)
    views = [(1, 1001), (3, 130), (2, 2), (7, 4999)]
    dig = ops.sha256_leaves([d[a:b] for a, b in views], [d[a + 1:b] for a, b in views], mode=1).cpu().numpy()
    for i, (a, b) in enumerate(views):
        raw = big.numpy().tobytes()
        assert dig[i].tobytes() == hashlib.sha256(raw[a:b] + raw[a + 1:b]).digest()


def test_sha256_prefix_range_error(cuda):
    from b2v import ops
    z = torch.zeros(8, dtype=torch.float16, device=cuda)
    with pytest.raises(OverflowError):
        ops.sha256_leaves([z], [z], prefixes=[70000])


@pytest.mark.parametrize("n", [1, 2, 3, 5, 8, 50, 51, 400, 1000])
def test_merkle_build_verify(cuda, n):
    from b2v import ops
    leaves = [hashlib.sha256(bytes([i % 256, i // 256]) * 16).digest() for i in range(n)]
    lh = torch.tensor(list(b"".join(leaves)), dtype=torch.uint8).reshape(n, 32).to(cuda)
    root, paths = ops.merkle_build(lh)
    want_root, want_paths = _tree(leaves)
    assert bytes(root.cpu().numpy().tobytes()) == want_root
    p = paths.cpu().numpy()
    for i in range(n):
        assert [p[i, k].tobytes() for k in range(p.shape[1])] == want_paths[i]
    ok = ops.merkle_verify(lh, paths, root).cpu()
    assert bool(ok.all())
    if n > 1:
        bad = lh.clone()
        bad[n // 2, 5] ^= 1
        ok = ops.merkle_verify(bad, paths, root).cpu()
        assert int(ok.sum()) == n - 1 and ok[n // 2] == 0


def test_merkle_duplicate_leaves_and_ties(cuda):
    """equal siblings (left == right) take the `else` branch of `if left < right` (proof.py:266): right+left."""
    from b2v import ops
    leaves = [hashlib.sha256(b"same").digest()] * 6
    lh = torch.tensor(list(b"".join(leaves)), dtype=torch.uint8).reshape(6, 32).to(cuda)
    root, _ = ops.merkle_build(lh)
    assert root.cpu().numpy().tobytes() == _tree(leaves)[0]


@pytest.mark.parametrize("row_bytes", [1, 2, 3, 4, 5, 11, 12, 13, 23, 24, 36, 47, 48, 1000, 4096, 1769472])
def test_base64_rows_match_stdlib(cuda, row_bytes):
    """b2v_base64_rows == base64.b64encode (miner.py:643-657) for every padding case, unaligned and vector paths."""
    import base64
    from b2v import ops
    n = 3 if row_bytes > 100000 else 5
    g = torch.Generator().manual_seed(row_bytes)
    host = torch.randint(0, 256, (n, row_bytes), dtype=torch.uint8, generator=g)
    enc = ops.base64_rows(host.to(cuda)).cpu().numpy()
    for i in range(n):
        assert enc[i].tobytes() == base64.b64encode(host[i].numpy().tobytes()), (row_bytes, i)


def test_base64_rows_fp16_latent(cuda):
    import base64
    from b2v import ops
    z = torch.randn(4, 1, 4, 3, 8, 8, generator=torch.Generator().manual_seed(1)).half()
    enc = ops.base64_rows(z.to(cuda)).cpu().numpy()
    for i in range(4):
        assert enc[i].tobytes() == base64.b64encode(z[i].numpy().tobytes())


def test_merkle_verify_roots_many_trees(cuda):
    """b2v_merkle_verify_roots: leaves of different trees (a validator round over many miners) in one launch, each
    against its own root; bit-exact vs the oracle fold."""
    import hashlib
    from b2v import ops
    from oracle import merkle_ref as M
    trees = []
    for k in range(5):
        hashes = [hashlib.sha256(bytes([k, i])).digest() for i in range(9 + k)]   # depth 4 for 9..13 leaves
        levels = M.tree_levels(hashes)
        trees.append((hashes, M.sibling_paths(levels), levels[-1][0]))
    rows = [(k, i) for k in range(5) for i in (0, 3, len(trees[k][0]) - 1)]
    lh = torch.tensor(list(b"".join(trees[k][0][i] for k, i in rows)), dtype=torch.uint8, device=cuda).view(-1, 32)
    pt = torch.tensor(list(b"".join(b"".join(trees[k][1][i]) for k, i in rows)), dtype=torch.uint8, device=cuda).view(len(rows), 4, 32)
    roots = [trees[k][2] for k, _ in rows]
    roots[4] = trees[0][2]                       # a leaf of tree 1 presented against tree 0's root
    rt = torch.tensor(list(b"".join(roots)), dtype=torch.uint8, device=cuda).view(-1, 32)
    ok = ops.merkle_verify_roots(lh, pt, rt).cpu().tolist()
    want = [int(M.verify_merkle_leaf(trees[k][0][i], trees[k][1][i], roots[j])) for j, (k, i) in enumerate(rows)]
    assert ok == want and want.count(0) == 1 and want[4] == 0


@pytest.mark.parametrize("nbytes", [0, 1, 2, 3, 4, 11, 12, 13, 47, 48, 1000, 3 * 4096 + 1, 1769472])
def test_base64_decode_rows_bit_exact(cuda, nbytes):
    """b2v_base64_decode_rows == base64.b64decode on canonical text, every byte, with the pad count reported."""
    import base64
    import numpy as np
    from b2v import ops
    rng = np.random.default_rng(nbytes)
    rows = [rng.integers(0, 256, nbytes, dtype=np.uint8).tobytes() for _ in range(3)]
    txt = [base64.b64encode(r) for r in rows]
    chars = len(txt[0])
    stride = max(16, (chars + 15) // 16 * 16)
    host = np.zeros((3, stride), dtype=np.uint8)
    for i, t in enumerate(txt):
        host[i, :chars] = np.frombuffer(t, dtype=np.uint8)
    out, status = ops.base64_decode_rows(torch.from_numpy(host).to(cuda), chars)
    st = status.cpu().tolist()
    assert st == [(3 - nbytes % 3) % 3] * 3
    o = out.cpu().numpy()
    for i, r in enumerate(rows):
        assert o[i, :nbytes].tobytes() == r == base64.b64decode(txt[i])


def test_base64_decode_flags_non_canonical_rows(cuda):
    """Rows Python's permissive decoder would treat specially (foreign characters, '=' before the end) are flagged
    0x100 and left to the host routine; trailing bits of a padded quad are ignored like binascii does."""
    import numpy as np
    from b2v import ops
    texts = [b"QUJDRA==", b"QUJD\nA==", b"QU=DRA==", b"QUJDRB==", b"QUJDR===", b"QUJDRE-G", b"QUJDREVG"]
    host = np.zeros((len(texts), 16), dtype=np.uint8)
    for i, t in enumerate(texts):
        host[i, :8] = np.frombuffer(t, dtype=np.uint8)
    out, status = ops.base64_decode_rows(torch.from_numpy(host).to(cuda), 8)
    st = status.cpu().tolist()
    assert st[0] == 2 and st[3] == 2 and st[6] == 0
    assert all(st[i] & 0x100 for i in (1, 2, 4, 5))
    o = out.cpu().numpy()
    assert o[0, :4].tobytes() == b"ABCD" and o[3, :4].tobytes() == b"ABCD" and o[6, :6].tobytes() == b"ABCDEF"
    # unaligned source rows take the byte path
    big = torch.from_numpy(np.frombuffer(b"x" + b"QUJDREVG" * 5, dtype=np.uint8).copy()).to(cuda)
    from b2v._lib import check, load
    dst = torch.zeros(32, dtype=torch.uint8, device=cuda)
    stt = torch.zeros(1, dtype=torch.int32, device=cuda)
    check(load().b2v_base64_decode_rows(big.data_ptr() + 1, 1, 40, 40, ops.ptr(dst), 32, ops.ptr(stt), ops.stream()))
    assert dst[:30].cpu().numpy().tobytes() == b"ABCDEF" * 5 and int(stt[0]) == 0
