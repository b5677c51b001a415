"""NCCL check of the temporal-segment halo exchange on real GPUs (run under torchrun):
every rank makes its deterministic output segment, halo_exchange_and_gather stitches on rank 0, and rank 0 compares with the
sequential parent loop (segments.stitch_sequential) on the same chunks — the reference's order, inference_cli.py:1163-1198."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import seedvr2_b200 as svr2
from seedvr2_b200 import segments

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True
for total, overlap, hw in ((41, 4, (36, 64)), (23, 2, (20, 20)), (16, 0, (8, 8)), (9, 3, (8, 8))):
    ranges = segments.segment_ranges(total, world, overlap)
    def chunk(r):
        s, e = ranges[r]
        g = torch.Generator().manual_seed(100 + r)
        return torch.rand(e - s, hw[0], hw[1], 3, generator=g)
    mine = chunk(rank).to(dev)
    out = segments.halo_exchange_and_gather(mine, overlap)
    if rank == 0:
        ref = segments.stitch_sequential([chunk(r).to(dev) for r in range(world)], overlap)
        same = out.shape == ref.shape and torch.equal(out, ref)
        ok &= bool(same)
        print(f"total {total} overlap {overlap} world {world}: stitched {tuple(out.shape)} {'== sequential parent loop (bit-exact)' if same else 'MISMATCH'}")
# timing at BASELINE config C3: 121 frames of 1920x1080 RGB fp32 split over the ranks with overlap 4: neighbour halo
# over NVLink + blend kernel + gather to rank 0, device-timed, max over ranks
total, overlap = 121, 4
ranges = segments.segment_ranges(total, world, overlap)
s_, e_ = ranges[rank]
mine = torch.rand(e_ - s_, 1080, 1920, 3, device=dev)
for it in range(3):
    dist.barrier(); torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    out = segments.halo_exchange_and_gather(mine, overlap)
    t1.record(); torch.cuda.synchronize()
    ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0 and it == 2:
        halo_mb = overlap * 1080 * 1920 * 3 * 4 / 1e6
        print(f"C3 stitch (121 x 1080p fp32, overlap 4, world {world}): {float(ms):.2f} ms = halo {halo_mb:.0f} MB per neighbour pair + blend + "
              f"gather of {out.shape[0]} frames ({out.numel() * 4 / 1e9:.2f} GB) to rank 0")
    del out
# ---- the same stitch as ONE kernel per rank over peer memory (segments.PeerStitcher -> svr2_segment_stitch_peer)
try:
    lens = [e - s for s, e in ranges]
    ps = segments.PeerStitcher(max(lens), total, (1080, 1920, 3), dev)
    ps.segment_buffer()[:lens[rank]].copy_(mine)
    ref = segments.halo_exchange_and_gather(mine, overlap)
    for it in range(4):
        dist.barrier(); torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        out = ps.stitch(lens, overlap)
        t1.record(); torch.cuda.synchronize()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        same = out.shape == ref.shape and torch.equal(out, ref)
        ok &= bool(same)
        print(f"C3 stitch as one peer-memory kernel per rank (world {world}): {float(ms):.2f} ms, "
              f"{'== NCCL halo exchange + gather (bit-exact)' if same else 'MISMATCH'}")
    # small ragged cases against the sequential parent loop
    for total2, overlap2, hw in ((41, 4, (36, 64)), (23, 2, (20, 20)), (16, 0, (8, 8))):
        r2 = segments.segment_ranges(total2, world, overlap2)
        l2 = [e - s for s, e in r2]
        if any(l < max(2 * overlap2, 1) for l in l2):
            if rank == 0:
                print(f"peer stitch total {total2} overlap {overlap2} world {world}: a segment is shorter than 2 x overlap "
                      f"(order-dependent stitch) -> this split stays on the NCCL gather path")
            continue
        def chunk2(r):
            g = torch.Generator().manual_seed(200 + r)
            return torch.rand(l2[r], hw[0], hw[1], 3, generator=g)
        ps2 = segments.PeerStitcher(max(l2), total2, (hw[0], hw[1], 3), dev)
        ps2.segment_buffer()[:l2[rank]].copy_(chunk2(rank).to(dev))
        o2 = ps2.stitch(l2, overlap2)
        if rank == 0:
            ref2 = segments.stitch_sequential([chunk2(r).to(dev) for r in range(world)], overlap2)
            same = o2.shape == ref2.shape and torch.equal(o2, ref2)
            ok &= bool(same)
            print(f"peer stitch total {total2} overlap {overlap2} world {world}: {'== sequential parent loop (bit-exact)' if same else 'MISMATCH'}")
        dist.barrier()
except Exception as e:      # symmetric memory needs CUDA VMM handle exchange between the ranks' processes
    ok = False
    print(f"rank {rank}: peer-memory stitch FAILED: {type(e).__name__}: {e}")
dist.barrier()
if rank == 0:
    print("NCCL segments check", "PASSED" if ok else "FAILED")
dist.destroy_process_group()
sys.exit(0 if ok else 1)
