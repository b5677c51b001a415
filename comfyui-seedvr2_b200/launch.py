"""One process per GPU launcher for the multi-GPU path.

Replaces the parent / worker fan-out of the reference CLI (inference_cli.py:936-1209): ``mp.Process`` per GPU pinned with
``CUDA_VISIBLE_DEVICES``, every worker loading the models again, results returned through ``share_memory_()`` + ``mp.Queue`` and
cross-faded sequentially on the parent's CPU.  Here the ranks are started by ``torchrun`` (or any launcher that sets RANK /
LOCAL_RANK / WORLD_SIZE / MASTER_*), each rank owns one GPU for the life of the job, and the only inter-rank traffic is the halo
exchange + gather of ``segments.py`` over NVLink — one peer-memory kernel per rank where symmetric memory is available, NCCL
send/recv otherwise.  The frame split is the reference's own (streaming mode :1087-1114, pre-loaded mode :1118-1135), only rank 0
keeps ``prepend_frames`` (:983-984) and rank 0 removes them from the stitched result (:1200-1207).

    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 -m seedvr2_b200.launch  <inference_cli.py arguments>

``run_segments`` is the library entry: it takes the per-rank pipeline as a callable, so the same code drives the reference's
``_stream_video_chunks`` / ``_process_frames_core`` (``main`` below, where the reference CLI is importable) and the stand-in
pipelines of the tests."""
import argparse
import os
import sys
from typing import Callable, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from . import segments


def init_distributed(backend: Optional[str] = None) -> Tuple[int, int, torch.device]:
    """(rank, world, device) from the torchrun environment; initialises the default process group when world > 1."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
    else:
        dev = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = backend or ("nccl" if dev.type == "cuda" else "gloo")
        if backend == "nccl":
            dist.init_process_group(backend, device_id=dev)
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world, dev


def rank_frame_range(total_frames: int, overlap: int, start_frame: int = 0, preloaded: bool = False, batch_size: int = 1,
                     rank: Optional[int] = None, world: Optional[int] = None) -> Tuple[int, int]:
    """[start, end) frames THIS rank reads: streaming split (inference_cli.py:1087-1114) or, for frames already in memory, the
    batch-aligned chunks of :1118-1135."""
    rank = dist.get_rank() if rank is None and dist.is_initialized() else (rank or 0)
    world = dist.get_world_size() if world is None and dist.is_initialized() else (world or 1)
    if preloaded:
        r = segments.chunk_ranges_preloaded(total_frames, world, overlap, batch_size)
        r += [(total_frames, total_frames)] * (world - len(r))        # torch.chunk may return fewer chunks than ranks
        a, b = r[rank]
        return start_frame + a, start_frame + b
    return segments.segment_ranges(total_frames, world, overlap, start_frame)[rank]


def run_segments(total_frames: int, process_segment: Callable[[int, int, int], torch.Tensor], *, overlap: int, start_frame: int = 0,
                 prepend_frames: int = 0, preloaded: bool = False, batch_size: int = 1, peer_memory: bool = True,
                 blend: Optional[Callable] = None, group: Optional[dist.ProcessGroup] = None) -> Optional[torch.Tensor]:
    """Run ``process_segment(start, end, prepend_frames_for_this_rank)`` on this rank's frame range and stitch the ranks' results.

    ``process_segment`` returns the rank's output frames ``[F, H, W, C]`` (fp32 in [0, 1] in the reference) on this rank's device.
    Returns the stitched video on rank 0 (prepended frames removed), ``None`` on the other ranks.  ``blend`` overrides the
    crossfade (tests on CPU); the default is the CUDA kernel ``svr2_halo_blend`` / the fused peer-memory stitch."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    a, b = rank_frame_range(total_frames, overlap, start_frame, preloaded, batch_size, rank, world)
    mine = process_segment(a, b, prepend_frames if rank == 0 else 0)          # only the first segment starts the video
    if mine.dim() != 4:
        raise RuntimeError("run_segments: process_segment must return [frames, H, W, C]")
    mine = mine.float().contiguous()
    if world == 1:
        out = mine
    else:
        out = _stitch(mine, overlap if world > 1 else 0, peer_memory and blend is None, blend, group)
    if rank != 0:
        return None
    if prepend_frames > 0 and prepend_frames < out.shape[0]:
        out = out[prepend_frames:]
    return out


def _stitch(mine: torch.Tensor, overlap: int, peer_memory: bool, blend: Optional[Callable], group) -> Optional[torch.Tensor]:
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if peer_memory and mine.is_cuda and overlap > 0:
        # every rank must take the same path: agree on the segment lengths first
        lens_t = torch.zeros(world, dtype=torch.int64, device=mine.device)
        lens_t[rank] = mine.shape[0]
        dist.all_reduce(lens_t, group=group)
        lens = [int(v) for v in lens_t.tolist()]
        frame_shape = tuple(mine.shape[1:])
        ok = all(n >= max(2 * overlap, 1) for n in lens) and (mine[0].numel() % 4 == 0)
        if ok:
            try:
                total = sum(lens) - overlap * (world - 1)
                st = segments.PeerStitcher(max(lens), total, frame_shape, mine.device, group)
                st.segment_buffer()[:lens[rank]].copy_(mine)
                out = st.stitch(lens, overlap)
                return out.clone() if out is not None else None      # the symmetric buffer dies with the stitcher
            except (RuntimeError, ImportError, AttributeError):
                pass              # no symmetric memory between these ranks: the NCCL form of the same exchange
    kw = {} if blend is None else {"blend": blend}
    return segments.halo_exchange_and_gather(mine, overlap, group, **kw)


# ----------------------------------------------------------------------------------------------------------------------------
# CLI: the reference's own argument parser and per-segment pipeline, one rank per GPU
# ----------------------------------------------------------------------------------------------------------------------------
def main(argv: Optional[Sequence[str]] = None) -> int:
    """``torchrun ... -m seedvr2_b200.launch <inference_cli.py arguments>``.  Needs the reference repository (with its dependencies)
    on PYTHONPATH: the launcher installs the B200 blocks, lets the reference parse its own arguments, gives every rank its frame
    range of the input video and hands the stitched frames back to the reference's writer on rank 0."""
    rank, world, dev = init_distributed()
    try:
        import inference_cli as ref          # the reference's CLI module (repository root on PYTHONPATH)
    except Exception as e:                   # noqa: BLE001 — missing module or one of its third-party imports
        raise SystemExit(f"seedvr2_b200.launch: the reference CLI is not importable here ({type(e).__name__}: {e}); put the "
                         f"ComfyUI-SeedVR2 checkout on PYTHONPATH") from e
    import cv2
    import seedvr2_b200
    seedvr2_b200.install()                   # before the reference builds its DiT (model_loader.py:437-438)
    if argv is not None:                     # parse_arguments() (inference_cli.py:1216) reads sys.argv
        sys.argv = [sys.argv[0]] + list(argv)
    args = ref.parse_arguments()
    args.cuda_device = None                  # the rank owns LOCAL_RANK's GPU; the reference's own fan-out stays off
    debug = ref.Debug(enabled=getattr(args, "debug", False))
    video_path = args.input
    cap = cv2.VideoCapture(video_path)
    n_total = int(cap.get(cv2.CAP_PROP_FRAME_COUNT))
    cap.release()
    start = int(getattr(args, "skip_first_frames", 0))
    n = n_total - start
    if getattr(args, "load_cap", 0):
        n = min(n, int(args.load_cap))

    def process_segment(a: int, b: int, prepend: int) -> torch.Tensor:
        cap = cv2.VideoCapture(video_path)
        cap.set(cv2.CAP_PROP_POS_FRAMES, a)
        worker_args = argparse.Namespace(**vars(args))
        worker_args.prepend_frames = prepend
        seg = b - a
        chunk = args.chunk_size if args.chunk_size > 0 else seg
        cache = {} if (args.cache_dit or args.cache_vae or seg > chunk) else None      # one rank, one model: keep it across chunks
        outs = [r for r in ref._stream_video_chunks(cap=cap, frames_to_process=seg, chunk_size=chunk, overlap=args.temporal_overlap,
                                                     args=worker_args, device_id=str(dev.index or 0), debug=debug, runner_cache=cache,
                                                     log_prefix=f"[rank {rank}] ")]
        cap.release()
        return torch.cat(outs, dim=0).to(dev) if outs else torch.empty(0, 1, 1, 3, device=dev)

    out = run_segments(n, process_segment, overlap=int(args.temporal_overlap), start_frame=start,
                       prepend_frames=int(getattr(args, "prepend_frames", 0)))
    if rank == 0:
        fps = cv2.VideoCapture(video_path).get(cv2.CAP_PROP_FPS) or 30.0
        out_path = args.output or os.path.splitext(video_path)[0] + "_upscaled.mp4"
        w = ref.save_frames_to_video(out.cpu(), out_path, fps)          # inference_cli.py:666-708: uint8 + BGR + mp4v writer
        if w is not None:
            w.release()
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
