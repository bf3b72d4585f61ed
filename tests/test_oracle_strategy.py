"""not-gpu: pins the strategy oracle (oracle/strategy_ref.py) AND the product's host-side planner (b2v.planner) against
tests/golden/strategy_golden.json — the reference's own planner / blend source lines executed by oracle/gen_golden.py."""
import hashlib
import json
import os

import pytest
import torch

import b2v  # noqa: F401
from b2v import planner
from oracle import strategy_ref as S

HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "strategy_golden.json")))


def test_planner_oracle_and_product_match_reference():
    n_ok = n_hang = 0
    for p in G["plans"]:
        fn = S.plan_chunk_only if p["variant"] == "chunk_only" else S.plan_coherent
        if p.get("hangs"):
            n_hang += 1
            with pytest.raises(ValueError):
                fn(p["T"], p["world"], p["chunk_size"], p["overlap"])
            with pytest.raises(planner.PlanError):
                planner.plan(p["T"], p["world"], p["chunk_size"], p["overlap"], variant=p["variant"])
            continue
        assert "raises" not in p
        cs, ov, ranges = fn(p["T"], p["world"], p["chunk_size"], p["overlap"])
        assert (cs, ov, [list(r) for r in ranges]) == (p["cs"], p["ov"], p["ranges"]), p
        q = planner.plan(p["T"], p["world"], p["chunk_size"], p["overlap"], variant=p["variant"])
        assert (q.chunk_size, q.overlap, [list(r) for r in q.ranges]) == (p["cs"], p["ov"], p["ranges"]), p
        assert len(q.ranges) % p["world"] == 0 or p["world"] == 1
        n_ok += 1
    assert n_ok > 400 and n_hang > 0


def test_survey_plans():
    # SURVEY.md §8a S3 "verified plans"
    assert S.plan_chunk_only(8, 2)[1:] == (1, [(0, 5), (4, 8)])
    assert S.plan_chunk_only(24, 1)[2] == [(0, 16), (12, 24)]
    assert S.plan_chunk_only(24, 4) == (8, 2, [(0, 8), (6, 14), (12, 20), (18, 24)])
    assert S.plan_chunk_only(24, 8)[2] == [(3 * k, 3 * k + 4) for k in range(7)] + [(21, 24)]
    assert S.plan_chunk_only(96, 8)[2] == [(12 * k, min(12 * k + 16, 96)) for k in range(8)]
    with pytest.raises(ValueError):
        S.plan_coherent(24, 8)
    with pytest.raises(ValueError):
        S.plan_coherent(8, 2)


def test_blend_oracle_matches_reference():
    for b in G["blends"]:
        dtype = getattr(torch, b["dtype"])
        g = torch.Generator().manual_seed(b["seed"])
        T, world = b["T"], b["world"]
        base = torch.randn(1, 4, T, 5, 7, generator=g).to(dtype)
        gathered = []
        for r in range(world):
            gathered.append([(s, e, (3 * torch.randn(1, 4, e - s, 5, 7, generator=g)).to(dtype))
                             for i, (s, e) in enumerate(b["ranges"]) if i % world == r])
        lat, full, weight = S.blend(gathered, T, b["ov"], base)
        assert str(lat.dtype) == b["lat_dtype"]
        assert weight.flatten().tolist() == b["weight"]
        assert hashlib.sha256(lat.numpy().tobytes()).hexdigest() == b["lat_sha256"]


@pytest.mark.parametrize("length,ov", [(1, 4), (2, 4), (3, 2), (5, 1), (8, 2), (16, 4), (6, 4), (4, 0), (7, 7)])
def test_product_window_weights_equal_oracle(length, ov):
    assert torch.equal(planner.window_weights(length, ov), S.ramp_weights(length, ov))


def test_gather_order_and_assignment():
    p = planner.plan(24, 4, variant="coherent")
    assert p.of_rank(1) == [p.ranges[1], p.ranges[5], p.ranges[9]]
    assert p.gather_order() == [r for k in range(4) for r in S.ranges_of_rank(p.ranges, 4, k)]
    assert planner.payload_bytes(p.of_rank(0), 4) == sum((e - s) * 8 for s, e in p.of_rank(0))
