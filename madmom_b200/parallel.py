"""Multi-GPU sharding of the front end: one process per GPU (torchrun), torch.distributed for the plumbing.

Two cases:

* a batch of tracks: independent units, rank r takes a contiguous block of tracks, NO collective
  (`shard_tracks`) -- the reference's equivalent is one worker process per file (madmom/processors.py:637-713);
* one long signal: frames are independent except for the `diff_frames` lag, so rank r takes the contiguous
  frame range [f_r, f_r+1) with equal sizes ceil(T / R), loads the PCM samples those frames (and their
  `diff_frames` lead-in frames) touch -- a halo of frame_size/2 + diff_frames*hop samples on the left and
  frame_size/2 on the right -- recomputes the lead-in rows locally and contributes its (rows, cols) slice to
  ONE all-gather (NCCL over NVLink on GPUs).  Frame starts always use the GLOBAL frame index in int(i * hop)
  (madmom/audio/signal.py:917), so the gathered matrix is bit-identical to the unsharded one; this is what
  `FramedSignal.__getitem__(slice)` (signal.py:1181-1194) means for an integer hop.

The shard arithmetic is pure Python (testable on CPU with the gloo backend); only `LongSignalFrontEnd.run`
touches the GPU.
"""
import os

import numpy as np

from . import recipe as R


def bind_host_thread_to_gpu(device):
    """Pin the calling host thread to the CPUs next to GPU `device` (NVML's ideal affinity), so that the pinned
    staging buffers allocated afterwards and the copy threads sit on the GPU's NUMA node.  With one process per GPU
    on a two-socket host the PCIe transfers of the host path otherwise cross the socket interconnect.  Returns True
    when the affinity was set."""
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(int(device)))
        return True
    except Exception:
        return False


def shard_tracks(n_tracks, world_size, rank):
    """Contiguous block [lo, hi) of tracks for `rank`; blocks differ by at most one track."""
    base, extra = divmod(int(n_tracks), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class FrameShard(object):
    """What one rank needs for its frame range of a long signal."""

    def __init__(self, rank, frame_first, frame_count, rows_per_rank, sample_lo, sample_hi):
        self.rank = rank
        self.frame_first = frame_first      # global index of the first output frame
        self.frame_count = frame_count      # real output frames of this rank (0 for ranks past the end)
        self.rows_per_rank = rows_per_rank  # rows every rank contributes to the all-gather (zero padded)
        self.sample_lo = sample_lo          # PCM slice [sample_lo, sample_hi) of the global signal
        self.sample_hi = sample_hi

    def unit(self):
        """Unit-table row (include/madmom_b200.h: mmb_track) for the local PCM slice."""
        return (0, self.sample_hi - self.sample_lo, self.sample_lo, self.frame_first, self.frame_count, 0)

    def __repr__(self):
        return "FrameShard(rank=%d frames [%d, %d) samples [%d, %d))" % (
            self.rank, self.frame_first, self.frame_first + self.frame_count, self.sample_lo, self.sample_hi)


def frame_shards(num_samples, hop_size, frame_size, world_size, origin=0, diff_frames=0, end="normal"):
    """[FrameShard] for every rank of `world_size`."""
    T = R.num_frames(num_samples, hop_size, end)
    per = -(-T // world_size) if T else 0
    # frame_size 1024: the kernel packs frames (2m, 2m + 1) into one FFT, so ranges start at even frames and the
    # halo reaches one frame further back; that keeps the shards bit-identical to the unsharded launch
    pair = 1 if frame_size == 1024 else 0
    per += per & pair
    shards = []
    for r in range(world_size):
        f0 = min(r * per, T)
        f1 = min(f0 + per, T)
        if f1 > f0:
            lead = max(f0 - diff_frames - pair, 0)
            s_lo = max(R.frame_start(lead, frame_size, hop_size, origin), 0)
            s_hi = min(R.frame_start(f1 - 1, frame_size, hop_size, origin) + frame_size, num_samples)
            s_hi = max(s_hi, s_lo)
        else:
            s_lo = s_hi = 0
        shards.append(FrameShard(r, f0, f1 - f0, per, s_lo, s_hi))
    return shards


def gather_rows(local, total_rows, group=None):
    """All-gather equal-sized (rows_per_rank, cols) slices along the rows and drop the padding of the last
    ranks.  `local` is a torch tensor (CUDA with NCCL, CPU with gloo)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    full = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(full, local.contiguous(), group=group)
    return full[:total_rows]


class LongSignalFrontEnd(object):
    """One long signal split into frame ranges over the ranks of the default process group."""

    def __init__(self, rec, pcm_dtype=np.int16, device=None):
        import torch
        from .batch import plan_for
        self.recipe = rec
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.plan = plan_for(rec, pcm_dtype, self.device)
        self.pcm_dtype = np.dtype(pcm_dtype)

    def shard(self, num_samples, world_size, rank):
        rec = self.recipe
        return frame_shards(num_samples, rec.hop_size, rec.frame_size, world_size,
                            R.resolve_origin(rec.origin, rec.frame_size), self.plan.diff_frames, rec.end)[rank]

    def run_local(self, pcm_slice, shard, outputs=("diff",)):
        """Fused kernel on this rank's PCM slice (CUDA tensor holding samples [sample_lo, sample_hi)).
        Returns {name: (rows_per_rank, cols) tensor}, rows past the rank's frame_count are zero."""
        import torch
        from . import _lib
        units = np.zeros(1, dtype=_lib.TRACK_DTYPE)
        units[0] = shard.unit()
        dev = "cuda:%d" % self.device
        out = {}
        for n in outputs:
            shape = (shard.rows_per_rank,) if n == "odf" else (shard.rows_per_rank, self.plan.width(n))
            out[n] = torch.zeros(shape, dtype=torch.float32, device=dev)
        if shard.frame_count > 0:
            self.plan.run(pcm_slice, units, {n: t[:shard.frame_count] for n, t in out.items()})
        return out

    def run_fused(self, pcm_slice, num_samples, outputs=("diff",), group=None, multicast=None):
        """Like run(), but without a separate collective: the front-end kernel stores every output row straight
        into the result matrices of ALL ranks (peer memory over NVLink, torch symmetric memory for the mapping;
        mmb_run_gather), so the transfer overlaps the computation row by row and the step ends with one barrier.
        Returns {name: (T, cols) tensor}: views of this rank's symmetric buffers (row stride padded to a multiple
        of 32 floats), valid until the next call with the same shapes.  Bit-identical to run().
        multicast: store through the NVLS multicast mapping of the buffers (one multimem.st per value, replicated
        by the NVSwitch) instead of one store per peer; None = when the symmetric memory has one (and
        MMB_GATHER_MULTICAST is not "0")."""
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        group = dist.group.WORLD if group is None else group
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        shard = self.shard(num_samples, world, rank)
        T = R.num_frames(num_samples, self.recipe.hop_size, self.recipe.end)
        per = shard.rows_per_rank
        dev = torch.device("cuda", self.device)
        if not hasattr(self, "_symm"):
            self._symm = {}
        bufs, handles = {}, {}
        for n in outputs:
            if n not in ("log", "diff", "odf"):
                raise ValueError("run_fused gathers the outputs log, diff and odf, not %r" % n)
            # rows padded to whole 128-byte lines: every row then starts on a line, and the kernel's remote stores
            # are full NVLink write packets instead of two partial ones each (the padding is never written)
            cols = 1 if n == "odf" else -(-self.plan.width(n) // 32) * 32
            key = (n, world, per, cols)
            if key not in self._symm:
                buf = symm.empty((world * max(per, 1), cols), dtype=torch.float32, device=dev)
                buf.zero_()
                self._symm[key] = (buf, symm.rendezvous(buf, group))
            bufs[n], handles[n] = self._symm[key]
        first = handles[outputs[0]]
        first.barrier(channel=0)             # every rank is done with the results of the previous call
        if shard.frame_count > 0:
            units = np.zeros(1, dtype=_lib.TRACK_DTYPE)
            units[0] = shard.unit()
            mc_ptrs = {n: int(getattr(handles[n], "multicast_ptr", 0) or 0) for n in outputs}
            if multicast is None:
                multicast = all(mc_ptrs.values()) and os.environ.get("MMB_GATHER_MULTICAST", "1") != "0"
            if multicast and not all(mc_ptrs.values()):
                raise RuntimeError("the symmetric memory of this group has no multicast (NVLS) mapping")
            dsts = []
            if multicast:
                dsts.append({n: (mc_ptrs[n] + rank * per * bufs[n].shape[1] * 4, bufs[n].shape[1]) for n in outputs})
            for p in ([] if multicast else [rank] + [q for q in range(world) if q != rank]):
                d = {}
                for n in outputs:
                    cols = bufs[n].shape[1]
                    d[n] = (int(handles[n].buffer_ptrs[p]) + rank * per * cols * 4, cols)
                dsts.append(d)
            self.plan.run_gather(pcm_slice, units, dsts, multicast=bool(multicast))
        first.barrier(channel=0)             # all peers have finished storing into this rank's buffers
        return {n: (bufs[n][:T, 0] if n == "odf" else bufs[n][:T, :self.plan.width(n)]) for n in outputs}

    def run(self, pcm_slice, num_samples, outputs=("diff",), group=None):
        """Compute this rank's slice and all-gather: every rank returns the full (T, cols) matrices."""
        import torch.distributed as dist
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        shard = self.shard(num_samples, world, rank)
        local = self.run_local(pcm_slice, shard, outputs)
        T = R.num_frames(num_samples, self.recipe.hop_size, self.recipe.end)
        return {n: gather_rows(t, T, group) for n, t in local.items()}
