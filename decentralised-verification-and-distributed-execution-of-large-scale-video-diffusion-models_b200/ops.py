"""torch.Tensor -> libb2v plumbing: pointers, dtype tags, current stream. torch supplies device memory and streams
only; every computation below happens in the CUDA kernels of csrc/."""
import ctypes as C

import torch

from . import _lib
from ._lib import BF16, F16, F32, check

_DT = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16}


def dtype_tag(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError("libb2v supports float32/float16/bfloat16 tensors, got %s" % t.dtype)


def _cuda(t, name="tensor"):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise _lib.B2VError("%s must be a CUDA tensor (libb2v has no CPU path)" % name)
    if not t.is_contiguous():
        raise ValueError("%s must be contiguous" % name)
    return t


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count():
    return int(_lib.load().b2v_launch_count())


# ---- step glue -------------------------------------------------------------------------------------------------
def cfg_ddim_step(lat, eps2, a_t, a_prev, guidance_scale=None, ctx=None, ctx_w=0.0, clip_sample=False,
                  make_next_input=0, leaf_z=None, leaf_eps=None, out=None, xnext=None):
    """lat [1,C,F,h,w]; eps2 [2,C,F,h,w] (uncond, cond) when guidance_scale is not None else [1,C,F,h,w].
    Returns (lat', x_next or None)."""
    lib = _lib.load()
    _cuda(lat, "lat"); _cuda(eps2, "eps2")
    _, Cc, F, h, w = lat.shape
    cfg = guidance_scale is not None
    if eps2.shape[0] != (2 if cfg else 1) or eps2.dtype != lat.dtype:
        raise ValueError("eps2 batch/dtype mismatch")
    if out is None:
        out = torch.empty_like(lat)
    if make_next_input and xnext is None:
        xnext = torch.empty((make_next_input,) + tuple(lat.shape[1:]), dtype=lat.dtype, device=lat.device)
    check(lib.b2v_cfg_ddim_step(ptr(lat), ptr(eps2), ptr(ctx), float(ctx_w), int(cfg),
                                float(guidance_scale or 0.0), float(a_t), float(a_prev), int(bool(clip_sample)),
                                ptr(out), ptr(xnext), int(make_next_input), ptr(leaf_z), ptr(leaf_eps),
                                dtype_tag(lat), Cc, F, h * w, stream()))
    return out, xnext


def make_unet_input(lat, n_dup, ctx=None, ctx_w=0.0):
    lib = _lib.load()
    _cuda(lat, "lat")
    _, Cc, F, h, w = lat.shape
    x = torch.empty((n_dup,) + tuple(lat.shape[1:]), dtype=lat.dtype, device=lat.device)
    check(lib.b2v_make_unet_input(ptr(lat), ptr(ctx), float(ctx_w), ptr(x), n_dup, dtype_tag(lat), Cc, F, h * w, stream()))
    return x


def context_mean(base):
    lib = _lib.load()
    _cuda(base, "base")
    _, Cc, T, h, w = base.shape
    ctx = torch.empty((1, Cc, 1, h, w), dtype=base.dtype, device=base.device)
    check(lib.b2v_context_mean(ptr(base), ptr(ctx), dtype_tag(base), Cc, T, h * w, stream()))
    return ctx


def blend_overlap_ptrs(ptrs, ranges, weights, T, Cc, h, w, dtype, device, return_full=False):
    """b2v_blend_overlap on raw device pointers: ptrs[i] -> [1,C,len_i,h,w] of `dtype` in THIS device's address space —
    local memory or a peer GPU's memory mapped over NVLink (exchange.PeerWindows): the kernel then pulls the windows out
    of the owners' HBM while it blends. Returns fp32 [1,C,T,h,w] (and the `dtype` accumulator when return_full)."""
    import numpy as np
    lib = _lib.load()
    n = len(ptrs)
    if n == 0:
        raise ValueError("no chunks")
    cp = (C.c_void_p * n)(*[int(p) for p in ptrs])
    starts = (C.c_int * n)(*[int(s) for s, _ in ranges])
    ends = (C.c_int * n)(*[int(e) for _, e in ranges])
    wflat = np.concatenate([np.asarray(wi, dtype=np.float32).reshape(-1) for wi in weights]).astype(np.float32)
    if wflat.size != sum(e - s for s, e in ranges):
        raise ValueError("one weight per frame of every chunk is required")
    out = torch.empty((1, Cc, T, h, w), dtype=torch.float32, device=device)
    full = torch.empty((1, Cc, T, h, w), dtype=dtype, device=device) if return_full else None
    check(lib.b2v_blend_overlap(cp, starts, ends, n, wflat.ctypes.data_as(C.c_void_p), Cc, T, h * w, _DT[dtype],
                                ptr(out), ptr(full), stream()))
    return (out, full) if return_full else out


def blend_overlap(chunks, ranges, weights, T, return_full=False):
    """chunks[i]: CUDA [1,C,len_i,h,w]; ranges[i] = (s,e); weights[i]: python/torch fp32 per-frame weights.
    Returns fp32 [1,C,T,h,w] (and the accumulator in the chunks' dtype when return_full)."""
    if len(chunks) == 0:
        raise ValueError("no chunks")
    for c in chunks:
        _cuda(c, "chunk")
    _, Cc, _, h, w = chunks[0].shape
    dev, dt = chunks[0].device, chunks[0].dtype
    for (s, e), c in zip(ranges, chunks):
        if c.shape[2] != e - s or c.dtype != dt:
            raise ValueError("chunk shape/dtype does not match its range")
    return blend_overlap_ptrs([c.data_ptr() for c in chunks], ranges, weights, T, Cc, h, w, dt, dev, return_full)


def base64_rows(buf, n_rows=None):
    """RFC 4648 base64 of the rows of a contiguous CUDA tensor [n_rows, ...] (any dtype; a row is its raw bytes), encoded
    on the device. Returns a CUDA uint8 tensor [n_rows, 4*ceil(row_bytes/3)] of ASCII (miner.py:643-657)."""
    lib = _lib.load()
    _cuda(buf, "buffer")
    if not buf.is_contiguous():
        raise ValueError("base64_rows needs a contiguous tensor")
    n = buf.shape[0] if n_rows is None else int(n_rows)
    row_bytes = (buf.numel() // max(buf.shape[0], 1)) * buf.element_size()
    out_row = int(lib.b2v_base64_len(row_bytes))
    out = torch.empty((n, out_row), dtype=torch.uint8, device=buf.device)
    check(lib.b2v_base64_rows(ptr(buf), n, row_bytes, row_bytes, ptr(out), stream()))
    return out


def base64_decode_rows(buf, row_chars=None):
    """Rows of base64 text (CUDA uint8 [n, stride], the first row_chars characters of each row) -> (bytes [n, dst_stride]
    CUDA uint8 with dst_stride = 3*row_chars/4 rounded up to 256, status int32 [n]: pad count | 0x100 if the row is not
    canonical base64 and must be decoded by the caller's permissive host routine)."""
    lib = _lib.load()
    _cuda(buf)
    n = buf.shape[0]
    row_chars = buf.shape[1] if row_chars is None else int(row_chars)
    nbytes = row_chars // 4 * 3
    stride = max(256, (nbytes + 255) // 256 * 256)
    out = torch.empty((n, stride), dtype=torch.uint8, device=buf.device)
    status = torch.empty(n, dtype=torch.int32, device=buf.device)
    check(lib.b2v_base64_decode_rows(ptr(buf), n, row_chars, buf.stride(0), ptr(out), stride, ptr(status), stream()))
    return out, status


# ---- proof -----------------------------------------------------------------------------------------------------
def sha256_leaves(a_list, b_list=None, prefixes=None, mode=0):
    """digest[i] = SHA256(prefix_i(2B BE, optional) ‖ bytes(a_i) ‖ bytes(b_i)); tensors are hashed in place on the
    device as raw bytes. Returns a CUDA uint8 tensor [n,32]."""
    lib = _lib.load()
    n = len(a_list)
    dev = a_list[0].device if n else torch.device("cuda")
    dig = torch.empty((n, 32), dtype=torch.uint8, device=dev)
    if n == 0:
        return dig
    for t in a_list:
        _cuda(t, "leaf buffer")
    ap = (C.c_void_p * n)(*[t.data_ptr() for t in a_list])
    al = (C.c_uint64 * n)(*[t.numel() * t.element_size() for t in a_list])
    bp = bl = None
    if b_list is not None:
        for t in b_list:
            _cuda(t, "leaf buffer")
        bp = (C.c_void_p * n)(*[t.data_ptr() for t in b_list])
        bl = (C.c_uint64 * n)(*[t.numel() * t.element_size() for t in b_list])
    pf = None
    if prefixes is not None:
        for p in prefixes:
            if not 0 <= int(p) < 65536:
                raise OverflowError("int too big to convert")  # same failure as int.to_bytes(2,'big') (miner.py:207)
        pf = (C.c_uint16 * n)(*[int(p) for p in prefixes])
    check(lib.b2v_sha256_leaves(pf, ap, al, bp, bl, n, ptr(dig), int(mode), stream()))
    return dig


def merkle_build(leaf_hashes, want_paths=True):
    """leaf_hashes: CUDA uint8 [n,32] -> (root uint8[32], paths uint8[n,depth,32] or None)."""
    lib = _lib.load()
    _cuda(leaf_hashes, "leaf_hashes")
    n = leaf_hashes.shape[0]
    depth = lib.b2v_merkle_depth(n)
    dev = leaf_hashes.device
    root = torch.empty(32, dtype=torch.uint8, device=dev)
    paths = torch.empty((n, depth, 32), dtype=torch.uint8, device=dev) if want_paths else None
    scratch = torch.empty(lib.b2v_merkle_scratch_bytes(n), dtype=torch.uint8, device=dev)
    check(lib.b2v_merkle_build(ptr(leaf_hashes), n, ptr(root), ptr(paths) if depth > 0 else C.c_void_p(0),
                               ptr(scratch), stream()))
    return root, paths


def merkle_build_static(leaf_hashes, root, paths=None):
    """b2v_merkle_build_static: the Merkle build for hosts that may not allocate device memory at the call site (a commit
    issued beside a running denoise loop, a process whose HBM is fully owned by its own pool). The caller brings the
    outputs — root uint8[32], paths uint8[n,depth,32] or None — and the tree levels go to an array inside libb2v.so;
    nothing is allocated here or in the library. n <= lib.b2v_merkle_static_max_leaves(), else B2VError. Calls share
    that array: keep them on one stream."""
    lib = _lib.load()
    _cuda(leaf_hashes, "leaf_hashes"); _cuda(root, "root")
    if paths is not None:
        _cuda(paths, "paths")
    check(lib.b2v_merkle_build_static(ptr(leaf_hashes), leaf_hashes.shape[0], ptr(root), ptr(paths), stream()))
    return root, paths


def merkle_verify(leaf_hashes, paths, root):
    lib = _lib.load()
    _cuda(leaf_hashes); _cuda(paths); _cuda(root)
    m, depth = paths.shape[0], paths.shape[1]
    ok = torch.empty(m, dtype=torch.uint8, device=leaf_hashes.device)
    check(lib.b2v_merkle_verify(ptr(leaf_hashes), ptr(paths), m, depth, ptr(root), ptr(ok), stream()))
    return ok


def merkle_verify_roots(leaf_hashes, paths, roots):
    """Leaves of different trees in one launch: leaf i against roots[i]. All CUDA uint8: [m,32], [m,depth,32], [m,32]."""
    lib = _lib.load()
    _cuda(leaf_hashes); _cuda(paths); _cuda(roots)
    m, depth = paths.shape[0], paths.shape[1]
    if roots.shape != (m, 32) or leaf_hashes.shape != (m, 32):
        raise ValueError("merkle_verify_roots: expected [m,32] hashes and roots")
    ok = torch.empty(m, dtype=torch.uint8, device=leaf_hashes.device)
    check(lib.b2v_merkle_verify_roots(ptr(leaf_hashes), ptr(paths), m, depth, ptr(roots), ptr(ok), stream()))
    return ok


def spotcheck_ddim(z_list, eps_list, znext_list, a_t, a_prev, C_, F, hw):
    import numpy as np
    lib = _lib.load()
    m = len(z_list)
    dev = z_list[0].device
    zp = (C.c_void_p * m)(*[t.data_ptr() for t in z_list])
    ep = (C.c_void_p * m)(*[t.data_ptr() for t in eps_list])
    np_ = (C.c_void_p * m)(*[t.data_ptr() for t in znext_list]) if znext_list is not None else None
    at = np.asarray(a_t, dtype=np.float32)
    ap = np.asarray(a_prev, dtype=np.float32)
    flags = torch.empty(m, dtype=torch.uint8, device=dev)
    check(lib.b2v_spotcheck_ddim(zp, ep, np_, at.ctypes.data_as(C.c_void_p), ap.ctypes.data_as(C.c_void_p), m,
                                 int(C_), int(F), int(hw), ptr(flags), stream()))
    return flags


# ---- building-block ops (tests) --------------------------------------------------------------------------------
def act_tag(t):
    """Activation dtype tag of a 2-byte operand tensor: bf16 (default arithmetic) or fp16 (the reference's)."""
    if t.dtype == torch.float16:
        return F16
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError("activation tensors are bfloat16 or float16, got %s" % t.dtype)


def gemm(a, w, bias=None, rowbias=None, rows_per_group=0, residual=None, taps=1, tap_row_stride=0, rows_per_batch=0,
         flags=0, out=None):
    lib = _lib.load()
    _cuda(a); _cuda(w)
    M, Kc = a.shape
    N = w.shape[0]
    assert w.shape[1] == taps * Kc
    n_out = N // 2 if flags & _lib.EPI_GEGLU else N
    if out is None:
        out = torch.empty((M, n_out), dtype=a.dtype, device=a.device)
    check(lib.b2v_op_gemm(ptr(a), ptr(w), ptr(bias), ptr(rowbias), int(rows_per_group), ptr(residual), ptr(out), M, N,
                          Kc, taps, int(tap_row_stride), int(rows_per_batch), int(flags), act_tag(a), stream()))
    return out


def gemm_rowstats(a, w, bias=None, residual=None):
    """gemm() that also emits the per-row {sum, sumsq} partials of its output: (out, stats [M, slots, 2] fp32)."""
    lib = _lib.load()
    _cuda(a); _cuda(w)
    M, Kc = a.shape
    N = w.shape[0]
    slots = lib.b2v_row_stats_slots(M, N, 0)
    out = torch.empty((M, N), dtype=a.dtype, device=a.device)
    stats = torch.zeros((M, max(slots, 1), 2), dtype=torch.float32, device=a.device)
    check(lib.b2v_op_gemm_rowstats(ptr(a), ptr(w), ptr(bias), ptr(residual), ptr(out), M, N, Kc, act_tag(a), ptr(stats),
                                   stream()))
    return out, stats


def gemm_lnin(a_raw, w_gamma, bias_folded, csum, ln_stats, eps=1e-5, flags=0):
    """LayerNorm(a_raw) @ W^T + b with the normalisation finished in the epilogue (see include/b2v.h)."""
    lib = _lib.load()
    _cuda(a_raw); _cuda(w_gamma)
    M, Kc = a_raw.shape
    N = w_gamma.shape[0]
    n_out = N // 2 if flags & _lib.EPI_GEGLU else N
    out = torch.empty((M, n_out), dtype=a_raw.dtype, device=a_raw.device)
    check(lib.b2v_op_gemm_lnin(ptr(a_raw), ptr(w_gamma), ptr(bias_folded), ptr(csum), ptr(ln_stats), ln_stats.shape[1],
                               float(eps), ptr(out), M, N, Kc, int(flags), act_tag(a_raw), stream()))
    return out


def gn_stats_ok(M, N, rows_per_batch=0, groups=32):
    return bool(_lib.load().b2v_gn_stats_ok(int(M), int(N), int(rows_per_batch), int(groups)))


def gemm_gnstats(a, w, bias=None, rowbias=None, rows_per_group=0, residual=None, taps=1, tap_row_stride=0,
                 rows_per_batch=0, groups=32):
    """gemm() whose epilogue also emits the GroupNorm statistics of its output: (out, stats[M/32, groups, 2] fp32)."""
    lib = _lib.load()
    _cuda(a); _cuda(w)
    M, Kc = a.shape
    N = w.shape[0]
    out = torch.empty((M, N), dtype=a.dtype, device=a.device)
    stats = torch.zeros((M // 32, groups, 2), dtype=torch.float32, device=a.device)
    check(lib.b2v_op_gemm_gnstats(ptr(a), ptr(w), ptr(bias), ptr(rowbias), int(rows_per_group), ptr(residual), ptr(out), M,
                                  N, Kc, taps, int(tap_row_stride), int(rows_per_batch), act_tag(a), ptr(stats), groups,
                                  stream()))
    return out, stats


def conv3x3_gnstats(x, w, bias=None, rowbias=None, rows_per_group=0, residual=None, groups=32):
    lib = _lib.load()
    _cuda(x); _cuda(w)
    N, H, W, Cin = x.shape
    Cout = w.shape[0]
    out = torch.empty((N, H, W, Cout), dtype=x.dtype, device=x.device)
    stats = torch.zeros((N * H * W // 32, groups, 2), dtype=torch.float32, device=x.device)
    check(lib.b2v_op_conv3x3_gnstats(ptr(x), ptr(w), ptr(bias), ptr(rowbias), int(rows_per_group), ptr(residual), ptr(out),
                                     N, H, W, Cin, Cout, act_tag(x), ptr(stats), groups, stream()))
    return out, stats


def group_norm_pre(x, gamma, beta, n_items, groups, eps, silu, pre):
    """group_norm() fed with statistics emitted by the producing GEMM (gemm_gnstats / conv3x3_gnstats)."""
    lib = _lib.load()
    _cuda(x)
    rows, C = x.shape
    y = torch.empty((rows, C), dtype=x.dtype, device=x.device)
    scratch = torch.empty(lib.b2v_op_group_norm_scratch_bytes(n_items, rows // n_items, groups) + 256, dtype=torch.uint8,
                          device=x.device)
    check(lib.b2v_op_group_norm_pre(ptr(x), C, ptr(gamma), ptr(beta), ptr(y), n_items, rows // n_items, groups, float(eps),
                                    int(silu), ptr(scratch), ptr(pre), act_tag(x), stream()))
    return y


def conv3x3(x, w, bias=None, rowbias=None, rows_per_group=0, residual=None, stride=1, mode=0):
    """x: [N,H,W,Cin] bf16; w: [Cout,3,3,Cin] bf16 -> [N,P,Q,Cout]."""
    lib = _lib.load()
    _cuda(x); _cuda(w)
    N, H, W, Cin = x.shape
    Cout = w.shape[0]
    P, Q = (H - 1) // stride + 1, (W - 1) // stride + 1
    out = torch.empty((N, P, Q, Cout), dtype=x.dtype, device=x.device)
    sb = lib.b2v_op_conv3x3_scratch_bytes(N, H, W, Cin, stride, mode)
    scratch = torch.empty(max(sb, 1), dtype=torch.uint8, device=x.device)
    check(lib.b2v_op_conv3x3(ptr(x), ptr(w), ptr(bias), ptr(rowbias), int(rows_per_group), ptr(residual), ptr(out), N,
                             H, W, Cin, Cout, stride, mode, ptr(scratch), act_tag(x), stream()))
    return out


def group_norm(x, gamma, beta, n_items, groups, eps, silu, x2=None):
    lib = _lib.load()
    _cuda(x)
    rows, C1 = x.shape
    C2 = x2.shape[1] if x2 is not None else 0
    y = torch.empty((rows, C1 + C2), dtype=x.dtype, device=x.device)
    stats = torch.empty(lib.b2v_op_group_norm_scratch_bytes(n_items, rows // n_items, groups) + 256, dtype=torch.uint8,
                        device=x.device)
    check(lib.b2v_op_group_norm(ptr(x), C1, ptr(x2), C2, ptr(gamma), ptr(beta), ptr(y), n_items, rows // n_items,
                                groups, float(eps), int(silu), ptr(stats), act_tag(x), stream()))
    return y


def layer_norm(x, gamma, beta, eps=1e-5):
    lib = _lib.load()
    _cuda(x)
    rows, Cc = x.shape
    y = torch.empty_like(x)
    check(lib.b2v_op_layer_norm(ptr(x), ptr(gamma), ptr(beta), ptr(y), rows, Cc, float(eps), act_tag(x), stream()))
    return y


def attention(q, k, v, heads, kv_group=1, scale=0.125):
    """q: [nb,Sq,heads*64] (may be a column slice view base: pass contiguous tensors), k/v: [nb/kv_group,Skv,heads*64]."""
    lib = _lib.load()
    nb, Sq, D = q.shape
    Skv = k.shape[1]
    out = torch.empty((nb, Sq, D), dtype=q.dtype, device=q.device)
    check(lib.b2v_op_attention(ptr(q), q.stride(1), ptr(k), ptr(v), k.stride(1), ptr(out), D, nb, heads, Sq, Skv,
                               kv_group, float(scale), act_tag(q), stream()))
    return out


def temporal_attention(qkv, B, F, HW, heads, scale=0.125):
    """qkv: [B*F*HW, 3*heads*64] fused rows in (b,f,p) order -> [B*F*HW, heads*64]."""
    lib = _lib.load()
    D = heads * 64
    out = torch.empty((qkv.shape[0], D), dtype=qkv.dtype, device=qkv.device)
    check(lib.b2v_op_temporal_attention(ptr(qkv), qkv.shape[1], 0, D, 2 * D, ptr(out), D, B, F, HW, heads,
                                        float(scale), act_tag(qkv), stream()))
    return out


# ---- frame decode pieces (composed by vae.py) --------------------------------------------------
def softmax_rows_(x, scale):
    """In-place row softmax of bf16 scores [rows, cols] (the decoder's single 512-wide attention head)."""
    lib = _lib.load()
    _cuda(x)
    rows, cols = x.shape
    check(lib.b2v_op_softmax_rows(ptr(x), rows, cols, x.stride(0), float(scale), stream()))
    return x


def frames_u8(rows, n_px, out=None):
    """Decoder output rows [n_px, ld] bf16 (RGB in channels 0..2) -> uint8 [n_px, 3]."""
    lib = _lib.load()
    _cuda(rows)
    if out is None:
        out = torch.empty((n_px, 3), dtype=torch.uint8, device=rows.device)
    check(lib.b2v_op_frames_u8(ptr(rows), rows.stride(0), int(n_px), ptr(out), stream()))
    return out


def upsample2x(x):
    """[N,H,W,C] bf16 -> [N,2H,2W,C] (nearest)."""
    lib = _lib.load()
    _cuda(x)
    N, H, W, Cc = x.shape
    y = torch.empty((N, 2 * H, 2 * W, Cc), dtype=torch.bfloat16, device=x.device)
    check(lib.b2v_op_upsample2x(ptr(x), ptr(y), N, H, W, Cc, stream()))
    return y


def latent_rows(x, cpad=64, act_dtype=torch.bfloat16):
    """[B,C,F,h,w] (fp32/fp16/bf16) -> channels-last rows [B*F*h*w, cpad] of act_dtype, channels >= C zero."""
    lib = _lib.load()
    _cuda(x)
    B, Cc, F, h, w = x.shape
    y = torch.empty((B * F * h * w, cpad), dtype=act_dtype, device=x.device)
    check(lib.b2v_op_latent_rows(ptr(x), dtype_tag(x), ptr(y), B, Cc, F, h * w, cpad, act_tag(y), stream()))
    return y
