"""Temporal-segment partitioner (comfyui-seedvr2_b200/segments.py): integer split arithmetic vs the oracle restatement of
inference_cli.py:1087-1135, and the halo exchange + gather over a world-size-2/3 gloo group on CPU (host logic only: the
crossfade itself is injected — the CUDA kernel svr2_halo_blend is checked in tests/test_gpu_kernels.py)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import partition


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("total,n,overlap", [(121, 8, 4), (121, 2, 4), (33, 4, 0), (7, 8, 2), (100, 3, 3), (1, 1, 4), (240, 8, 4)])
def test_segment_ranges_match_oracle(svr2, total, n, overlap):
    from seedvr2_b200 import segments
    assert segments.segment_ranges(total, n, overlap, 5) == partition.segment_ranges(total, n, overlap, 5)
    r = segments.segment_ranges(total, n, overlap)
    owned = [b - a - (overlap if i < n - 1 and b - a > 0 else 0) for i, (a, b) in enumerate(r)]
    assert r[0][0] == 0 and r[-1][1] == total
    assert all(r[i + 1][0] <= r[i][1] for i in range(n - 1))


def test_preloaded_chunks(svr2):
    from seedvr2_b200 import segments
    # inference_cli.py:1118-1135 by hand: total 100, 4 ranks, overlap 4, batch 5 -> chunk = ceil((25+4)/5)*5 = 30, stride 26
    assert segments.chunk_ranges_preloaded(100, 4, 4, 5) == [(0, 30), (26, 56), (52, 82), (78, 100)]
    assert segments.chunk_ranges_preloaded(10, 3, 0) == [(0, 4), (4, 8), (8, 10)]          # torch.chunk
    assert [segments.padded_batch_frames(n) for n in (1, 2, 5, 19, 20, 21, 33)] == [1, 5, 5, 21, 21, 21, 33]


@pytest.mark.parametrize("total,bs,ov", [(232, 33, 4), (120, 33, 4), (116, 33, 4), (62, 33, 4), (33, 33, 4), (29, 33, 4), (121, 33, 4),
                                         (121, 121, 4), (37, 33, 4), (38, 33, 4), (10, 5, 0), (10, 4, 4), (10, 4, 7), (0, 33, 4), (1, 33, 4)])
def test_batch_ranges_match_oracle(svr2, total, bs, ov):
    from seedvr2_b200 import segments
    got = segments.batch_ranges(total, bs, ov)
    assert got == partition.batch_ranges(total, bs, ov)
    if total:
        assert got[0][0] == 0 and got[-1][1] == total
        assert all(b - a <= bs for a, b in got)


def test_stream_workload_is_balanced(svr2):
    """bench.py's default job: 232 frames at 1080p, batch_size 33, overlap 4 — the reference's own split arithmetic gives every
    rank the same number of 33-frame forwards at 1, 2, 4 and 8 ranks (the last batch of the last rank holds 29 frames)."""
    from seedvr2_b200 import segments
    for n in (1, 2, 4, 8):
        per_rank = []
        for (a, b) in segments.segment_ranges(232, n, 4):
            bt = segments.batch_ranges(b - a, 33, 4)
            per_rank.append([segments.latent_frames(e - s) for s, e in bt])
        assert all(len(x) == 8 // n for x in per_rank), per_rank
        assert sorted(set(sum(per_rank, []))) == [8, 9]
    assert [segments.latent_frames(f) for f in (1, 5, 21, 29, 33, 121)] == [1, 2, 6, 8, 9, 31]


def _worker(rank, world, port, lens, overlap, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import seedvr2_b200  # noqa: F401
    from seedvr2_b200 import segments
    g = torch.Generator().manual_seed(100 + rank)
    frames = torch.rand(lens[rank], 6, 8, 3, generator=g)
    out = segments.halo_exchange_and_gather(frames.clone(), overlap, blend=partition.blend)
    if rank == 0:
        torch.save(out, out_path)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("lens,overlap", [([12, 9], 4), ([10, 8, 9], 3), ([12, 9], 0), ([5, 9, 6], 4), ([9, 2, 9], 4)])
def test_halo_exchange_gloo(svr2, tmp_path, lens, overlap):
    """World size 2 and 3 on CPU/gloo: the distributed stitch equals the reference parent loop (oracle.partition.stitch) on
    the same per-rank chunks, including the short-segment cases that fall back to the sequential order."""
    world = len(lens)
    ctx = mp.get_context("spawn")
    out_path = str(tmp_path / "stitched.pt")
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, lens, overlap, out_path)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    got = torch.load(out_path)
    chunks = [torch.rand(n, 6, 8, 3, generator=torch.Generator().manual_seed(100 + r)) for r, n in enumerate(lens)]
    ref = partition.stitch(chunks, overlap)
    assert got.shape == ref.shape
    assert torch.equal(got, ref)


@pytest.mark.parametrize("lens,overlap", [((20, 19, 19, 15), 4), ((8, 8), 4), ((5, 7, 6), 2), ((3, 3, 3), 0), ((9,), 3),
                                          ((16, 15, 15, 15, 15, 15, 15, 15), 4)])
def test_stitch_layout_places_frames_like_the_sequential_loop(svr2, lens, overlap):
    """Host arithmetic of the peer-memory stitch (segments.stitch_layout) against the parent loop restated in
    segments.stitch_sequential (inference_cli.py:1163-1198): same length, copied frames in the same places, crossfades where
    the loop blends."""
    import torch
    from seedvr2_b200 import segments
    chunks = [torch.arange(l, dtype=torch.float32).view(l, 1, 1, 1).expand(l, 1, 1, 4) + 1000.0 * r for r, l in enumerate(lens)]
    ref = segments.stitch_sequential(chunks, overlap, blend=lambda p, c, o: torch.full_like(c, -1.0))
    lay = segments.stitch_layout(lens, overlap)
    total = lay[-1][0] + lay[-1][1] + lay[-1][3]
    assert total == ref.shape[0]
    for r, (off, head, first_body, n_body) in enumerate(lay):
        assert bool((ref[off:off + head] == -1.0).all())
        assert torch.equal(ref[off + head: off + head + n_body], chunks[r][first_body:first_body + n_body])
    with pytest.raises(ValueError):
        segments.stitch_layout((6, 3, 6), 2)           # a segment shorter than 2 * overlap: order-dependent, not offered


# ---- the launcher (comfyui-seedvr2_b200/launch.py): frame ranges per rank, per-rank pipeline, stitch, prepend handling -------------
def _launch_worker(rank, world, port, total, overlap, prepend, preloaded, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import seedvr2_b200  # noqa: F401
    from seedvr2_b200 import launch
    r, w, dev = launch.init_distributed("gloo")
    assert (r, w) == (rank, world) and dev.type == "cpu"
    seen = {}

    def pipeline(a, b, pre):       # a deterministic stand-in for the per-rank upscaling pipeline: frame f -> value f (+ rank tint)
        seen["range"], seen["prepend"] = (a, b), pre
        f = torch.arange(a, b, dtype=torch.float32).view(-1, 1, 1, 1).expand(-1, 4, 6, 3).clone()
        return f + 0.001 * rank

    out = launch.run_segments(total, pipeline, overlap=overlap, prepend_frames=prepend, preloaded=preloaded, batch_size=5,
                              blend=partition.blend)
    torch.save({"out": out, "range": seen["range"], "prepend": seen["prepend"]}, f"{out_path}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("total,overlap,prepend,preloaded", [(41, 4, 0, False), (41, 4, 3, False), (30, 2, 0, True), (17, 0, 0, False)])
def test_launcher_runs_segments_and_stitches(svr2, tmp_path, total, overlap, prepend, preloaded):
    from seedvr2_b200 import segments
    world = 2
    out_path = str(tmp_path / "launch")
    mp.spawn(_launch_worker, args=(world, _free_port(), total, overlap, prepend, preloaded, out_path), nprocs=world, join=True)
    res = [torch.load(f"{out_path}.{r}") for r in range(world)]
    ranges = (segments.chunk_ranges_preloaded(total, world, overlap, 5) if preloaded else segments.segment_ranges(total, world, overlap))
    assert [tuple(r["range"]) for r in res] == [tuple(x) for x in ranges]
    assert [r["prepend"] for r in res] == [prepend, 0]                      # only the first segment starts the video
    assert res[1]["out"] is None
    chunks = [torch.arange(a, b, dtype=torch.float32).view(-1, 1, 1, 1).expand(-1, 4, 6, 3) + 0.001 * r for r, (a, b) in enumerate(ranges)]
    want = partition.stitch(chunks, overlap)
    if prepend:
        want = want[prepend:]
    assert torch.equal(res[0]["out"], want)
