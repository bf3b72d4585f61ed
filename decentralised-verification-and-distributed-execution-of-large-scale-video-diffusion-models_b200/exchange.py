"""End-of-denoise window exchange (chunk_only.py:123-135 ``all_gather_object`` of pickled CPU tensors) over NVLink.

Two transports, same result bit for bit:

* ``PeerWindows`` (default on CUDA, world > 1): every rank writes its denoised windows into a buffer that all peers
  have mapped (CUDA VMM / NVLink peer mapping set up once through ``torch.distributed._symmetric_memory``; torch is
  the plumbing for the mapping and the cross-rank signal, nothing else). After one device-side barrier the blend
  kernel (``b2v_blend_overlap``) is launched with the PEERS' pointers as its chunk pointers: it pulls the halo and
  interior frames straight out of the neighbours' HBM through NVSwitch and applies the ramp while they arrive — the
  exchange and the overlap smoothing are ONE kernel, there is no staging copy and no NCCL collective on the data path.
* ``nccl_all_gather`` (``--exchange nccl``, and the gloo path of the CPU tests): one ``all_gather_into_tensor`` of a
  padded device buffer, then the same blend kernel on local memory.

The reference performs no per-step communication and parity forbids adding any: this is the only
exchange of the run (<= 1.2 MB per window at ZeroscopeXL).
"""
import logging
import time

import torch
import torch.distributed as dist

LOG = logging.getLogger("b2v.exchange")


def window_layout(plan, world, C_, h, w, itemsize):
    """Per rank: [(s, e, byte offset inside that rank's buffer)], and the largest per-rank byte count."""
    lay, cap = [], 0
    for r in range(world):
        off, lst = 0, []
        for s, e in plan.of_rank(r):
            lst.append((s, e, off))
            off += C_ * (e - s) * h * w * itemsize
        lay.append(lst)
        cap = max(cap, off)
    return lay, cap


class PeerWindows:
    """Symmetric (peer-mapped) window buffers of one process group. Build once per run; ``gather_blend`` may be called
    repeatedly (buffers are reused; the second barrier keeps a fast rank from overwriting what a slow one still reads)."""

    def __init__(self, plan, world, rank, C_, h, w, dtype, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.plan, self.world, self.rank = plan, world, rank
        self.C, self.h, self.w, self.dtype, self.device = C_, h, w, dtype, device
        self.itemsize = torch.empty((), dtype=dtype).element_size()
        self.layout, cap = window_layout(plan, world, C_, h, w, self.itemsize)
        self.cap = max(cap, 16)
        group = group if group is not None else dist.group.WORLD
        self.buf = symm_mem.empty(self.cap, dtype=torch.uint8, device=device)
        self.hdl = symm_mem.rendezvous(self.buf, group)
        self.peer_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        if len(self.peer_ptrs) != world or any(p == 0 for p in self.peer_ptrs):
            raise RuntimeError("peer mapping failed: %r" % (self.peer_ptrs,))

    def stage(self, my_windows):
        """Copy this rank's denoised windows ([1,C,len,h,w]) into its peer-visible buffer (device-local copies)."""
        for (s, e, t), (ls, le, off) in zip(my_windows, self.layout[self.rank]):
            if (s, e) != (ls, le):
                raise ValueError("window order does not match the plan: %r vs %r" % ((s, e), (ls, le)))
            n = t.numel() * t.element_size()
            dst = self.buf[off:off + n].view(self.dtype).view(t.shape[1:])
            dst.copy_(t[0].to(self.dtype))

    def chunk_pointers(self):
        """(pointers, ranges) of every window of every rank in the reference's gather order (rank 0's list, rank
        1's, ...), pointing into the owners' memory."""
        ptrs, ranges = [], []
        for r in range(self.world):
            for s, e, off in self.layout[r]:
                ptrs.append(self.peer_ptrs[r] + off)
                ranges.append((s, e))
        return ptrs, ranges

    def gather_blend(self, my_windows, T, ov):
        """stage -> barrier -> ONE blend kernel reading all windows through the peer mappings -> barrier.
        Returns (fp32 [1,C,T,h,w], seconds spent in the exchange+blend on the device)."""
        from . import ops, planner
        self.stage(my_windows)
        torch.cuda.synchronize(self.device)
        t0 = time.time()
        self.hdl.barrier(channel=0)           # every rank's windows are in place and visible
        ptrs, ranges = self.chunk_pointers()
        weights = [planner.window_weights(e - s, ov) for s, e in ranges]
        out = ops.blend_overlap_ptrs(ptrs, ranges, weights, T, self.C, self.h, self.w, self.dtype, self.device)
        self.hdl.barrier(channel=1)           # nobody restages before every reader is done
        torch.cuda.synchronize(self.device)
        return out, time.time() - t0


def nccl_all_gather(my_windows, plan, rank, world, device):
    """All ranks end up with every window, in gather order (rank 0's list, rank 1's, ...): one all_gather of a
    device buffer padded to the longest per-rank payload. Returns (list of (s, e, tensor), seconds)."""
    if world == 1:
        return [(s, e, t) for (s, e, t) in my_windows], 0.0
    shape = my_windows[0][2].shape  # [1,C,len,h,w]
    C_, h, w = shape[1], shape[3], shape[4]
    dtype = my_windows[0][2].dtype
    per_rank = [plan.of_rank(r) for r in range(world)]
    frames = [sum(e - s for s, e in lst) for lst in per_rank]
    cap = max(frames)
    send = torch.zeros((C_, cap, h, w), dtype=dtype, device=device)
    off = 0
    for s, e, t in my_windows:
        send[:, off:off + (e - s)] = t[0]
        off += e - s
    recv = torch.empty((world, C_, cap, h, w), dtype=dtype, device=device)
    if device.type == "cuda":
        torch.cuda.synchronize(device)
    t0 = time.time()
    dist.all_gather_into_tensor(recv.view(-1), send.view(-1))
    if device.type == "cuda":
        torch.cuda.synchronize(device)
    dt = time.time() - t0
    out = []
    for r in range(world):
        off = 0
        for s, e in per_rank[r]:
            out.append((s, e, recv[r, :, off:off + (e - s)].unsqueeze(0).contiguous()))
            off += e - s
    return out, dt
