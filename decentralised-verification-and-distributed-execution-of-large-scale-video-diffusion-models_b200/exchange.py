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


# ---- per-step halo exchange ---------------------------------------------------------------
def halo_neighbours(plan, world):
    """Global window order and, for every window, its overlap (in frames) with the previous / next window.
    Returns (windows, halos): windows = [(s, e, rank, index on that rank)] sorted by start; halos[i] = (ov_left, ov_right).
    Raises ValueError when a frame is covered by more than two windows or two windows share a start (padded plans):
    the per-step exchange is defined between adjacent windows only."""
    wins = sorted(((s, e, r, j) for r in range(world) for j, (s, e) in enumerate(plan.of_rank(r))),
                  key=lambda x: (x[0], x[1]))
    for a, b in zip(wins[:-1], wins[1:]):
        if b[0] <= a[0] or b[1] <= a[1]:
            raise ValueError("halo exchange needs strictly advancing windows, got %r then %r" % (a[:2], b[:2]))
    for a, c in zip(wins[:-2], wins[2:]):
        if c[0] < a[1]:
            raise ValueError("halo exchange needs at most two windows per frame, %r and %r overlap" % (a[:2], c[:2]))
    halos = []
    for i, wdw in enumerate(wins):
        left = max(0, wins[i - 1][1] - wdw[0]) if i > 0 else 0
        right = max(0, wdw[1] - wins[i + 1][0]) if i + 1 < len(wins) else 0
        halos.append((left, right))
    return wins, halos


def crossfade(ov):
    """Partition-of-unity weights over an overlap of `ov` frames: (earlier window's, later window's), fp32; the later
    window's weight of overlap frame k is (k + 0.5) / ov, the earlier one's is 1 minus that — computed once so both
    sides multiply by identical numbers."""
    later = torch.tensor([(k + 0.5) / ov for k in range(ov)], dtype=torch.float32)
    return (1.0 - later), later


class HaloSync:
    """After EVERY denoising step adjacent windows exchange their overlapping frames and both replace them with the same
    crossfade, so the windows evolve as one video instead of drifting apart until the final blend (the reference's
    windows never talk during the loop: this is the N4 variant, opt-in, with its own single-process oracle).

    Each rank publishes the first / last halo frames of its windows into a peer-mapped buffer (two parity slots, one
    device-side barrier per step); the blend kernel then reads the NEIGHBOUR's published frames straight out of its
    HBM over NVLink together with the local window: exchange and crossfade are one kernel, as in PeerWindows.
    world == 1 uses a plain device buffer and no barrier. Chunks are always accumulated earlier-window-first, so the
    two owners of an overlap frame compute bit-identical values."""

    def __init__(self, plan, world, rank, C_, h, w, dtype, device, group=None):
        self.world, self.rank, self.device, self.dtype = world, rank, device, dtype
        self.C, self.h, self.w = C_, h, w
        self.itemsize = torch.empty((), dtype=dtype).element_size()
        self.wins, self.halos = halo_neighbours(plan, world)
        self.n_local = [len(plan.of_rank(r)) for r in range(world)]
        ov_max = max([max(l, r) for l, r in self.halos] + [1])
        self.slot = (C_ * ov_max * h * w * self.itemsize + 255) // 256 * 256
        cap = 2 * 2 * max(self.n_local) * self.slot                       # parity x (head, tail) x windows
        if world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            self.buf = symm_mem.empty(cap, dtype=torch.uint8, device=device)
            self.hdl = symm_mem.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
            self.base_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
            if len(self.base_ptrs) != world or any(p == 0 for p in self.base_ptrs):
                raise RuntimeError("peer mapping failed: %r" % (self.base_ptrs,))
        else:
            self.buf, self.hdl = torch.empty(cap, dtype=torch.uint8, device=device), None
            self.base_ptrs = [self.buf.data_ptr()]
        self.mine = [i for i, wdw in enumerate(self.wins) if wdw[2] == rank]  # global indices of my windows, by start
        self.mine.sort(key=lambda i: self.wins[i][3])                       # ... in this rank's own window order
        self.step = 0

    def _offset(self, parity, j, side):
        return ((parity * max(self.n_local) + j) * 2 + side) * self.slot

    def _slot_view(self, parity, j, side, ov):
        off = self._offset(parity, j, side)
        n = self.C * ov * self.h * self.w * self.itemsize
        return self.buf[off:off + n].view(self.dtype).view(self.C, ov, self.h, self.w)

    def sync(self, lats):
        """lats: this rank's window latents [1,C,len,h,w] (the rank's window order). Returns the synchronised latents."""
        from . import ops
        parity = self.step & 1
        for gi, lat in zip(self.mine, lats):
            (s, e, _, j), (ovl, ovr) = self.wins[gi], self.halos[gi]
            if lat.shape[2] != e - s or lat.dtype != self.dtype:
                raise ValueError("window %r: latent %s does not match the plan" % ((s, e), tuple(lat.shape)))
            if ovl:
                self._slot_view(parity, j, 0, ovl).copy_(lat[0, :, :ovl])
            if ovr:
                self._slot_view(parity, j, 1, ovr).copy_(lat[0, :, e - s - ovr:])
        if self.hdl is not None:
            self.hdl.barrier(channel=parity)      # every rank's halo frames of this step are published
        out = []
        for gi, lat in zip(self.mine, lats):
            (s, e, _, _), (ovl, ovr) = self.wins[gi], self.halos[gi]
            n = e - s
            own_w = torch.ones(n, dtype=torch.float32)
            ptrs, ranges, weights = [], [], []
            if ovl:
                _, _, lr, lj = self.wins[gi - 1]
                earlier, later = crossfade(ovl)
                ptrs.append(self.base_ptrs[lr] + self._offset(parity, lj, 1))      # the left window's published tail
                ranges.append((0, ovl))
                weights.append(earlier)
                own_w[:ovl] = later
            ptrs.append(lat.data_ptr())
            ranges.append((0, n))
            if ovr:
                earlier, later = crossfade(ovr)
                own_w[n - ovr:] = earlier
            weights.append(own_w)
            if ovr:
                _, _, rr, rj = self.wins[gi + 1]
                ptrs.append(self.base_ptrs[rr] + self._offset(parity, rj, 0))      # the right window's published head
                ranges.append((n - ovr, n))
                weights.append(later)
            if len(ptrs) == 1:
                out.append(lat)
                continue
            _, full = ops.blend_overlap_ptrs(ptrs, ranges, weights, n, self.C, self.h, self.w, self.dtype, self.device,
                                             return_full=True)
            out.append(full)
        self.step += 1
        return out
