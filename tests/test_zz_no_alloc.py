"""-m gpu: b2v_merkle_build_static, the Merkle build that allocates nothing.

Some hosts of this library cannot give b2v_merkle_build the scratch buffer it asks for: a miner that commits its
trajectory while the denoising loop is still running on the same GPU would stall that loop on cudaMalloc, and a serving
process that has handed every byte of HBM to its own pool has nothing left to allocate from. For them the library keeps
the tree levels in an array of its own (include/b2v.h). This file checks, through the C-ABI, that the entry point gives
the same root and sibling paths as b2v_merkle_build — bit for bit, the bar for all byte work in this repo — on two calls
in a row that share the array, that free device memory does not move across a call, and that trees beyond the stated
limit are refused with a message that names it.

The file name sorts after every other test file on purpose: pytest -x stops at the first failure, and a failure here
must not hide the tests of the reference-facing path. It holds no state of its own, so it passes alone in a fresh
process (where its first call is also the library's first call on the device) and after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _hashes(n, seed, dev):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (n, 32), dtype=torch.uint8, generator=g).to(dev)


def _static(lib, ops, h, fill):
    """Outputs are pre-filled with `fill` so a kernel that skipped a byte would show."""
    n = h.shape[0]
    depth = lib.b2v_merkle_depth(n)
    root = torch.full((32,), fill, dtype=torch.uint8, device=h.device)
    paths = torch.full((n, depth, 32), fill, dtype=torch.uint8, device=h.device)
    rc = lib.b2v_merkle_build_static(ops.ptr(h), n, ops.ptr(root), ops.ptr(paths) if depth else C.c_void_p(0),
                                     ops.stream())
    assert rc == 0, lib.b2v_last_error()
    return root, paths


def test_static_scratch_matches_merkle_build_twice_in_a_row(cuda):
    from b2v import ops
    lib = ops._lib.load()                                    # the handle and error type ops itself uses
    limit = lib.b2v_merkle_static_max_leaves()
    assert limit == 4096
    # Back-to-back calls on one stream, sizes chosen so that each call finds the previous, differently shaped tree in the
    # shared array: large then small (stale upper levels), odd counts (self-paired tail nodes), one leaf, the limit.
    sizes = [limit, 50, 777, 3, 1, 2, 4095, 50]
    for k, n in enumerate(sizes):
        h = _hashes(n, 1000 + k, cuda)
        want_root, want_paths = ops.merkle_build(h)          # existing entry point, caller-provided scratch
        first = _static(lib, ops, h, 0xAA)
        second = _static(lib, ops, h, 0x55)                  # same inputs again: the array now holds this very tree
        torch.cuda.synchronize()
        for root, paths in (first, second):
            assert torch.equal(root, want_root), n
            assert torch.equal(paths, want_paths), n
    # the root-only form (paths_dev = NULL)
    h = _hashes(50, 7, cuda)
    root = torch.zeros(32, dtype=torch.uint8, device=cuda)
    ops.merkle_build_static(h, root)
    assert torch.equal(root, ops.merkle_build(h, want_paths=False)[0])


def test_static_scratch_allocates_nothing_and_enforces_its_limit(cuda):
    from b2v import ops
    lib = ops._lib.load()                                    # the handle and error type ops itself uses
    limit = lib.b2v_merkle_static_max_leaves()
    h = _hashes(limit + 1, 3, cuda)
    root = torch.zeros(32, dtype=torch.uint8, device=cuda)
    paths = torch.zeros((limit + 1, lib.b2v_merkle_depth(limit + 1), 32), dtype=torch.uint8, device=cuda)
    paths_ok = torch.zeros((limit, lib.b2v_merkle_depth(limit), 32), dtype=torch.uint8, device=cuda)
    _static(lib, ops, h[:64], 0)                             # set-up call: module load + symbol lookup happen here
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    ops.merkle_build_static(h[:limit], root, paths_ok)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, (free_before, free_after)
    want_root, want_paths = ops.merkle_build(h[:limit])
    assert torch.equal(root, want_root) and torch.equal(paths_ok, want_paths)

    before = root.clone()
    rc = lib.b2v_merkle_build_static(ops.ptr(h), limit + 1, ops.ptr(root), ops.ptr(paths), ops.stream())
    assert rc == -1                                          # B2V_ERR_ARG
    msg = lib.b2v_last_error().decode()
    assert str(limit) in msg and "b2v_merkle_build" in msg, msg
    with pytest.raises(ops._lib.B2VError):
        ops.merkle_build_static(h, root, paths)
    assert lib.b2v_merkle_build_static(ops.ptr(h), 0, ops.ptr(root), None, ops.stream()) == -1
    torch.cuda.synchronize()
    assert torch.equal(root, before)                         # a refused call launches nothing
