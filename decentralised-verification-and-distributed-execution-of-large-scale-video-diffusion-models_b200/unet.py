"""``pipe.unet`` of the reference's hot loop on libb2v.

Mirrors the call surface the reference uses on diffusers' ``UNet3DConditionModel``:
    noise = pipe.unet(x, t, encoder_hidden_states=emb).sample        (chunk_only.py:71, miner.py:581)
    pipe.unet.config.in_channels                                     (chunk_only.py:111)
    unet.enable_xformers_memory_efficient_attention() / enable_gradient_checkpointing()   (miner.py:94-100; no-ops)
Weights are fed under their diffusers ``state_dict`` names (b2v_unet3d_load_tensor) so a real checkpoint loads.
"""
import ctypes as C
import threading
from types import SimpleNamespace

import torch

from . import _lib, ops
from ._lib import check


class UNet3DOutput:
    """diffusers' UNet3DConditionOutput has one field the reference reads: ``.sample``."""
    __slots__ = ("sample",)

    def __init__(self, sample):
        self.sample = sample

    def __getitem__(self, i):  # return_dict=False style access
        return (self.sample,)[i]


ZEROSCOPE_CONFIG = dict(in_channels=4, out_channels=4, block_out_channels=(320, 640, 1280, 1280), layers_per_block=2,
                        cross_attention_dim=1024, attention_head_dim=64, norm_num_groups=32, norm_eps=1e-5,
                        transformer_in_heads=8, text_len=77)


class UNet3DConditionModel:
    def __init__(self, config=None, device="cuda", dtype=torch.bfloat16):
        """dtype: element type of the activations and GEMM operands INSIDE the network — torch.bfloat16 (default) or
        torch.float16 (the reference's ``torch_dtype=torch.float16``, chunk_only.py:44-47). Accumulation is fp32 either
        way; I/O tensors of a call may be fp32/fp16/bf16 independently."""
        if dtype not in (torch.bfloat16, torch.float16):
            raise ValueError("UNet3DConditionModel arithmetic dtype must be bfloat16 or float16, got %s" % dtype)
        cfg = dict(ZEROSCOPE_CONFIG)
        if config:
            cfg.update({k: v for k, v in dict(config).items() if k in cfg})
        self.config = SimpleNamespace(**cfg)
        dev = device if isinstance(device, torch.device) else torch.device(device)
        if dev.type != "cuda":
            raise _lib.B2VError("UNet3DConditionModel runs on an sm_100 CUDA device only (no CPU path)")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        self.dtype = dtype  # arithmetic type of the kernels; I/O tensors may be fp32/fp16/bf16
        lib = _lib.load()
        c = _lib.UNet3DConfigC()
        c.in_channels, c.out_channels = cfg["in_channels"], cfg["out_channels"]
        for i, v in enumerate(cfg["block_out_channels"]):
            c.block_out_channels[i] = int(v)
        c.layers_per_block = cfg["layers_per_block"]
        c.cross_attention_dim = cfg["cross_attention_dim"]
        c.attention_head_dim = cfg["attention_head_dim"]
        c.norm_num_groups = cfg["norm_num_groups"]
        c.norm_eps = cfg["norm_eps"]
        c.transformer_in_heads = cfg["transformer_in_heads"]
        c.text_len = cfg["text_len"]
        c.act_dtype = _lib.F16 if dtype == torch.float16 else _lib.BF16
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(lib.b2v_unet3d_create(C.byref(c), C.byref(h)))
        self._h = h
        self._lib = lib
        self._finalized = False
        self._ws = None
        self._fwd_lock = threading.Lock()

    # ---- weights ----
    def load_state_dict(self, state_dict, strict=True):
        """state_dict: diffusers names -> tensors (CPU or CUDA, fp32/fp16/bf16)."""
        with torch.cuda.device(self.device):
            for name, t in state_dict.items():
                t = t.detach()
                if t.dtype != torch.float32:
                    t = t.float()  # the library keeps fp32 masters until finalize: hand them over as fp32 (a plain copy)
                t = t.contiguous()
                shape = (C.c_int64 * t.ndim)(*t.shape)
                check(self._lib.b2v_unet3d_load_tensor(self._h, name.encode(), C.c_void_p(t.data_ptr()),
                                                       ops._DT[t.dtype], t.ndim, shape))
            check(self._lib.b2v_unet3d_finalize(self._h))
        self._finalized = True
        return self

    @classmethod
    def from_module(cls, module, device="cuda", dtype=torch.bfloat16):
        """Build from an object exposing ``cfg``/``config`` and ``state_dict()`` (the oracle restatement or a real
        diffusers model)."""
        cfg = getattr(module, "cfg", None)
        cfg = cfg.to_dict() if hasattr(cfg, "to_dict") else dict(getattr(module, "config", {}))
        return cls(cfg, device=device, dtype=dtype).load_state_dict(module.state_dict())

    def param_count(self):
        return int(self._lib.b2v_unet3d_param_count(self._h))

    def flops(self, B, F, H, W):
        return float(self._lib.b2v_unet3d_flops(self._h, B, F, H, W))

    def workspace_bytes(self, B, F, H, W):
        return int(self._lib.b2v_unet3d_workspace_bytes(self._h, B, F, H, W))

    # ---- diffusers no-ops the miner calls (miner.py:94-100) ----
    def enable_xformers_memory_efficient_attention(self, *a, **k):
        return None

    def enable_gradient_checkpointing(self, *a, **k):
        return None

    def eval(self):
        return self

    def to(self, *a, **k):
        return self

    # ---- forward ----
    def _workspace(self, B, F, H, W):
        need = self.workspace_bytes(B, F, H, W)
        if need == 0:
            raise _lib.B2VError("unsupported UNet shape B=%d F=%d H=%d W=%d: %s" % (B, F, H, W, _lib.last_error()))
        if self._ws is None or self._ws.numel() < need:
            self._ws = self._ws_raw = None
            self._ws_raw = torch.empty(need + 1024, dtype=torch.uint8, device=self.device)
            off = (-self._ws_raw.data_ptr()) % 1024  # the C-ABI wants a 1 KiB aligned workspace
            self._ws = self._ws_raw[off:off + need]
        return self._ws

    def forward_into(self, sample, timestep, encoder_hidden_states, out):
        if not self._finalized:
            raise _lib.B2VError("UNet weights not loaded")
        ops._cuda(sample, "sample"); ops._cuda(encoder_hidden_states, "encoder_hidden_states")
        B, Cin, F, H, W = sample.shape
        if encoder_hidden_states.dtype != sample.dtype:
            encoder_hidden_states = encoder_hidden_states.to(sample.dtype)
        if encoder_hidden_states.shape[0] != B:
            raise ValueError("encoder_hidden_states batch %d != sample batch %d" % (encoder_hidden_states.shape[0], B))
        t = float(timestep.item()) if isinstance(timestep, torch.Tensor) else float(timestep)
        # the miner calls from several axon worker threads with no lock of its own (miner.py:571-600): one forward at a
        # time may enqueue onto this handle's workspace
        with self._fwd_lock:
            ws = self._workspace(B, F, H, W)
            check(self._lib.b2v_unet3d_forward(self._h, ops.ptr(sample), t, ops.ptr(encoder_hidden_states), ops.ptr(out),
                                               ops.dtype_tag(sample), B, F, H, W, ops.ptr(ws), ws.numel(), ops.stream()))
        return out

    def __call__(self, sample, timestep, encoder_hidden_states=None, return_dict=True, **_ignored):
        out = torch.empty((sample.shape[0], self.config.out_channels) + tuple(sample.shape[2:]), dtype=sample.dtype,
                          device=sample.device)
        self.forward_into(sample, timestep, encoder_hidden_states, out)
        return UNet3DOutput(out) if return_dict else (out,)

    forward = __call__

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.b2v_unet3d_destroy(self._h)
                self._h = None
        except Exception:
            pass
