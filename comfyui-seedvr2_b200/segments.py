"""Temporal-segment partitioner: spreads frames with temporal overlap across ranks, exchanges the overlap halo between
neighbouring ranks and gathers the stitched result on rank 0.

Replaces the reference's multi-GPU fan-out (inference_cli.py:1048-1209): ``mp.Process`` per GPU + ``CUDA_VISIBLE_DEVICES``
pinning, frame-range split (:1087-1114 streaming mode, :1118-1135 pre-loaded mode), ``share_memory_()`` + ``mp.Queue`` return
(:1021, :1148-1153) and the parent's sequential CPU crossfade (:1163-1198, src/core/generation_utils.py:284-312).

B200 design: one process per GPU under ``torchrun``; the only messages are (1) the halo — rank i sends its last ``overlap``
output frames to rank i+1 (grouped NCCL send/recv over NVLink, neighbours only), the receiver crossfades them into its own
first ``overlap`` frames with ``svr2_halo_blend`` and the sender drops them; (2) the final gather of every rank's owned
frames to rank 0.  There is no collective inside the DiT forward.  The split arithmetic is integer and identical to the
reference's; the stitched video is identical to the reference parent loop's result for the same split whenever every
segment has at least 2*overlap frames (the reference's chained edge cases for shorter segments are handled by falling back
to its sequential order on rank 0, still with the CUDA blend).
"""
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def segment_ranges(total_frames: int, num_ranks: int, overlap: int, start_frame: int = 0) -> List[Tuple[int, int]]:
    """[start, end) frames every rank READS (streaming mode, inference_cli.py:1087-1114): ``base = total // n`` frames,
    +1 for the first ``total % n`` ranks; every rank but the last reads ``overlap`` extra frames, clamped to the end."""
    if num_ranks <= 0 or total_frames < 0 or overlap < 0:
        raise ValueError("segment_ranges: bad arguments")
    base, rem = divmod(total_frames, num_ranks)
    out, cur = [], start_frame
    for idx in range(num_ranks):
        n = base + (1 if idx < rem else 0)
        end = cur + n
        if idx < num_ranks - 1 and overlap > 0:
            end = min(end + overlap, start_frame + total_frames)
        out.append((cur, end))
        cur += n
    return out


def chunk_ranges_preloaded(total_frames: int, num_ranks: int, overlap: int, batch_size: int = 1) -> List[Tuple[int, int]]:
    """Pre-loaded-frames mode (inference_cli.py:1118-1135): chunk = total // n + overlap rounded up to a multiple of
    ``batch_size``; stride = chunk - overlap; the last rank takes everything that is left.  Without overlap (or n == 1) it
    is ``torch.chunk``."""
    if overlap > 0 and num_ranks > 1:
        chunk = total_frames // num_ranks + overlap
        if batch_size > 1:
            chunk = ((chunk + batch_size - 1) // batch_size) * batch_size
        stride = chunk - overlap
        out = []
        for i in range(num_ranks):
            s = i * stride
            e = total_frames if i == num_ranks - 1 else min(s + chunk, total_frames)
            out.append((min(s, total_frames), max(e, min(s, total_frames))))
        return out
    size = -(-total_frames // num_ranks) if total_frames else 0       # torch.chunk: ceil-sized chunks
    out, s = [], 0
    while s < total_frames:
        out.append((s, min(s + size, total_frames)))
        s += size
    return out


def padded_batch_frames(n_frames: int) -> int:
    """Frames after the pipeline pads a batch to the VAE's 4n+1 constraint (generation_phases.py:109-124)."""
    if n_frames <= 0:
        return 0
    return n_frames if n_frames % 4 == 1 else ((n_frames - 1 + 3) // 4) * 4 + 1


def batch_ranges(total_frames: int, batch_size: int, overlap: int) -> List[Tuple[int, int]]:
    """[start, end) frames of every latent batch INSIDE one rank's segment (generation_phases.py:267-282): batches of
    ``batch_size`` frames advance by ``batch_size - overlap``; a trailing batch that would hold nothing but overlap frames is
    dropped; ``overlap >= batch_size`` resets the overlap to 0 as the reference does.  Each batch is then padded to 4n+1
    frames (``padded_batch_frames``) and is one DiT forward."""
    if total_frames < 0 or batch_size <= 0 or overlap < 0:
        raise ValueError("batch_ranges: bad arguments")
    step = batch_size - overlap if overlap > 0 else batch_size
    if step <= 0:
        step, overlap = batch_size, 0
    out = []
    for idx in range(0, total_frames, step):
        end = min(idx + batch_size, total_frames)
        if idx > 0 and end - idx <= overlap:
            break
        out.append((idx, end))
    return out


def latent_frames(n_frames: int) -> int:
    """Latent frames of a batch of ``n_frames`` pixel frames: padded to 4n+1, then the VAE's 4x temporal compression with the
    first frame kept (T = (F - 1) / 4 + 1)."""
    return (padded_batch_frames(n_frames) - 1) // 4 + 1 if n_frames > 0 else 0


def _default_blend(prev_tail: torch.Tensor, cur_head: torch.Tensor, overlap: int) -> torch.Tensor:
    from . import ops
    return ops.halo_blend(prev_tail.contiguous(), cur_head.contiguous(), overlap)


def stitch_sequential(chunks: Sequence[torch.Tensor], overlap: int, blend: Callable = _default_blend) -> torch.Tensor:
    """The parent loop of inference_cli.py:1163-1198 on one device (all chunks local)."""
    chunks = [c.float() for c in chunks]
    if overlap <= 0 or len(chunks) == 1:
        return torch.cat(chunks, dim=0)
    result = chunks[0]
    for c in chunks[1:]:
        if c.shape[0] > overlap and result.shape[0] >= overlap:
            result = torch.cat([result[:-overlap], blend(result[-overlap:], c[:overlap], overlap), c[overlap:]], dim=0)
        elif c.shape[0] > overlap:
            result = torch.cat([result, c[overlap:]], dim=0)
    return result


def halo_exchange_and_gather(frames: torch.Tensor, overlap: int, group: Optional[dist.ProcessGroup] = None,
                             blend: Callable = _default_blend) -> Optional[torch.Tensor]:
    """``frames`` = this rank's output segment ``[F_r, H, W, C]`` (fp32 after the pipeline's post-processing), ranks ordered
    by time.  Returns the stitched video on rank 0, ``None`` elsewhere.

    Step 1 (halo): rank r>0 receives rank r-1's last ``overlap`` frames, blends them into its own first ``overlap`` frames;
    rank r<n-1 drops its last ``overlap`` frames.  Step 2: ragged gather of the owned frames to rank 0.
    If any segment is shorter than 2*overlap the reference's result depends on the sequential order (the blended head of a
    chunk is itself the next blend's tail), so everything is gathered first and stitched on rank 0 in that order."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    frames = frames.float().contiguous()
    if world == 1:
        return frames
    dev = frames.device
    lens_t = torch.zeros(world, dtype=torch.int64, device=dev)
    lens_t[rank] = frames.shape[0]
    dist.all_reduce(lens_t, group=group)
    lens = [int(v) for v in lens_t.tolist()]
    frame_shape = tuple(frames.shape[1:])

    def g(r):     # global rank of group rank r
        return dist.get_global_rank(group, r) if group is not None else r

    simple = overlap > 0 and all(n >= 2 * overlap for n in lens)
    if overlap > 0 and not simple:
        parts = _gather_ragged(frames, lens, frame_shape, rank, world, group, g)
        return stitch_sequential(parts, overlap, blend) if rank == 0 else None

    if overlap > 0:
        ops_ = []
        recv = None
        if rank < world - 1:
            ops_.append(dist.P2POp(dist.isend, frames[-overlap:].contiguous(), g(rank + 1), group))
        if rank > 0:
            recv = torch.empty((overlap,) + frame_shape, dtype=frames.dtype, device=dev)
            ops_.append(dist.P2POp(dist.irecv, recv, g(rank - 1), group))
        for w in dist.batch_isend_irecv(ops_):
            w.wait()
        if rank > 0:
            frames[:overlap] = blend(recv, frames[:overlap], overlap)
        if rank < world - 1:
            frames = frames[:-overlap]
            lens = [n - overlap if r < world - 1 else n for r, n in enumerate(lens)]
    parts = _gather_ragged(frames.contiguous(), lens, frame_shape, rank, world, group, g)
    return torch.cat(parts, dim=0) if rank == 0 else None


def _gather_ragged(frames, lens, frame_shape, rank, world, group, g):
    """Rank 0 receives every rank's [lens[r], ...] tensor (grouped point-to-point: NCCL has no ragged gather)."""
    if rank == 0:
        parts = [frames] + [torch.empty((lens[r],) + frame_shape, dtype=frames.dtype, device=frames.device) for r in range(1, world)]
        ops_ = [dist.P2POp(dist.irecv, parts[r], g(r), group) for r in range(1, world) if lens[r] > 0]
    else:
        parts = None
        ops_ = [dist.P2POp(dist.isend, frames, g(0), group)] if lens[rank] > 0 else []
    if ops_:
        for w in dist.batch_isend_irecv(ops_):
            w.wait()
    return parts


# ---- the same stitch as ONE kernel per rank over peer memory -----------------------------------------------------------------
def stitch_layout(lens: Sequence[int], overlap: int) -> List[Tuple[int, int, int, int]]:
    """Per rank (out_offset, head, first_body, n_body) of the stitched video for segments of ``lens`` frames:
    ``head`` frames are crossfades with the previous segment's tail, then ``n_body`` frames are copied from
    ``first_body`` on; every segment but the last drops its final ``overlap`` frames (inference_cli.py:1163-1198).
    Only defined where the sequential loop's result does not depend on its order: every segment >= 2 * overlap frames."""
    n = len(lens)
    if overlap < 0 or any(l < max(2 * overlap, 1) for l in lens):
        raise ValueError("stitch_layout: every segment needs at least max(2 * overlap, 1) frames")
    out, off = [], 0
    for r, l in enumerate(lens):
        head = overlap if r > 0 else 0
        body = l - head - (overlap if r < n - 1 else 0)
        out.append((off, head, head, body))
        off += head + body
    return out


class PeerStitcher:
    """Halo exchange + crossfade + gather as one kernel per rank (``svr2_segment_stitch_peer``): every rank reads its
    predecessor's tail and writes its frames into the root's video through NVLink peer mappings of two symmetric-memory
    buffers; the only other device work is a signal-pad barrier on either side.  The producer writes its segment into
    ``segment_buffer()``; ``stitch(lens)`` returns the stitched video on rank 0 (a view of the root buffer), None elsewhere.
    Symmetric memory is torch.distributed plumbing (CUDA VMM handles exchanged through the process group's store)."""

    def __init__(self, max_segment_frames: int, total_out_frames: int, frame_shape: Tuple[int, ...], device: torch.device,
                 group: Optional[dist.ProcessGroup] = None):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.frame_shape = tuple(frame_shape)
        self.frame_elems = 1
        for d in self.frame_shape:
            self.frame_elems *= d
        if self.frame_elems % 4:
            raise ValueError("PeerStitcher: frame element count must be a multiple of 4")
        self.seg = symm.empty((max_segment_frames,) + self.frame_shape, dtype=torch.float32, device=device)
        self.out = symm.empty((total_out_frames,) + self.frame_shape, dtype=torch.float32, device=device)
        self.h_seg = symm.rendezvous(self.seg, self.group)
        self.h_out = symm.rendezvous(self.out, self.group)

    def segment_buffer(self) -> torch.Tensor:
        return self.seg

    def stitch(self, lens: Sequence[int], overlap: int) -> Optional[torch.Tensor]:
        from . import _lib, ops
        lay = stitch_layout(lens, overlap)
        off, head, first_body, n_body = lay[self.rank]
        total = lay[-1][0] + lay[-1][1] + lay[-1][3]
        if lens[self.rank] > self.seg.shape[0] or total > self.out.shape[0]:
            raise ValueError("PeerStitcher: buffers too small for these segments")
        fb = self.frame_elems * 4
        prev_tail = None
        if head:
            prev_tail = int(self.h_seg.buffer_ptrs[self.rank - 1]) + (lens[self.rank - 1] - head) * fb
        out_ptr = int(self.h_out.buffer_ptrs[0]) + off * fb
        self.h_seg.barrier(channel=0)                  # every rank's segment is complete before anyone reads a tail
        ops._call("svr2_segment_stitch_peer", 1, self.seg.device, _lib.lib().svr2_segment_stitch_peer, self.seg.data_ptr(),
                  prev_tail, out_ptr, head, first_body, n_body, self.frame_elems, ops._stream(self.seg))
        self.h_out.barrier(channel=0)                  # every rank's frames have landed in the root's buffer
        return self.out[:total] if self.rank == 0 else None
