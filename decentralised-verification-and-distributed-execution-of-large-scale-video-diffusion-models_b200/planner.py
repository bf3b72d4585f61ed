"""Frame-window planner and blend weights: the host-side integer logic of the strategies.

  plan(T, world, chunk_size, overlap, variant)
      variant "chunk_only"  — Distribution/strategies/chunk_only.py:80-105 (== fsdp_chunked.py:136-171)
      variant "coherent"    — Distribution/strategies/fsdp_chunked_coherent.py:149-177
  window_weights(length, ov)  — the per-frame ramp of chunk_only.py:138-147 (computed with torch.linspace so the fp32
                                values are the reference's)
  assignment(ranges, world)   — window i -> rank i % world (chunk_only.py:115)

Reference quirks kept on purpose: the ramp starts at exactly 0 (Q1); the coherent variant does
not clamp the overlap to cs // 3, and where the reference would loop forever (cs <= ov, Q3) this raises PlanError.
"""
import torch


class PlanError(ValueError):
    pass


class Plan:
    __slots__ = ("T", "world", "chunk_size", "overlap", "ranges")

    def __init__(self, T, world, chunk_size, overlap, ranges):
        self.T, self.world, self.chunk_size, self.overlap, self.ranges = T, world, chunk_size, overlap, ranges

    def of_rank(self, rank):
        return [r for i, r in enumerate(self.ranges) if i % self.world == rank]

    def gather_order(self):
        """Windows in the order the reference's blend visits them: rank 0's list, rank 1's list, ..."""
        out = []
        for r in range(self.world):
            out += self.of_rank(r)
        return out

    def __repr__(self):
        return "Plan(T=%d, world=%d, cs=%d, ov=%d, ranges=%s)" % (self.T, self.world, self.chunk_size, self.overlap,
                                                                 self.ranges)


def auto_chunk_size(T, world):
    smallest = max(4, T // (2 * world))
    largest = min(16, T // world)
    return min(largest, max(smallest, T // world))


def _windows(T, size, ov):
    stride = size - ov
    if stride <= 0:
        raise PlanError("chunk_size=%d <= overlap=%d: windows would never advance (the reference hangs here; pass "
                        "--chunk_size/--overlap explicitly)" % (size, ov))
    return [(s, min(s + size, T)) for s in range(0, T, stride)]


def plan(T, world, chunk_size=0, overlap=4, variant="chunk_only", no_chunking=False):
    if T <= 0 or world <= 0:
        raise PlanError("num_frames and world size must be positive")
    if variant == "coherent" and no_chunking:
        cs, ov = T, 0
    else:
        cs = chunk_size if chunk_size > 0 else auto_chunk_size(T, world)
        if variant == "chunk_only":
            ov = min(overlap, cs // 3)
        elif variant == "coherent":
            ov = overlap if overlap > 0 else max(4, cs // 3)
        else:
            raise PlanError("unknown planner variant %r" % (variant,))
    if cs <= 0:
        raise PlanError("chunk size %d (num_frames=%d, world=%d)" % (cs, T, world))
    ranges = _windows(T, cs, ov)
    if len(ranges) % world != 0:
        # grow the window until the count divides the world size ...
        for grown in range(cs + 1, 2 * cs):
            cand = _windows(T, grown, ov)
            if len(cand) % world == 0:
                cs, ranges = grown, cand
                break
    if len(ranges) % world != 0:
        # ... else pad with copies of the last window so every rank runs the same number of UNet forwards
        ranges = ranges + [ranges[-1]] * (world - len(ranges) % world)
    return Plan(T, world, cs, ov, ranges)


def window_weights(length, ov):
    """fp32 [length] blend weights of one window."""
    w = torch.ones(length, dtype=torch.float32)
    if ov > 0:
        k = min(ov, length)
        ramp = torch.linspace(0, 1, ov)[:k]
        if k > 0:
            w[:k] = ramp
            w[length - k:] = torch.flip(ramp, [0])
    return w


def payload_bytes(ranges, in_channels):
    """`network_bytes` CSV column exactly as the reference computes it (chunk_only.py:126; omits h*w, Q4)."""
    return sum((e - s) * in_channels * 2 for (s, e) in ranges)


def true_payload_bytes(ranges, in_channels, h, w, elem_size=2):
    return sum((e - s) * in_channels * h * w * elem_size for (s, e) in ranges)
