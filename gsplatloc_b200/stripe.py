"""Tile-striped tracking of ONE frame over several GPUs (BASELINE.json configs[4]: the 2 M-Gaussian
1920x1080 map on 8 B200).  The shape of gsplat's ``rasterization(distributed=True)`` (``IDX:14882-14926``) re-cut for
GsplatLoc's path, where the only gradient is the 6-DoF pose and the loss couples neighbouring pixels:

* every rank holds the whole map and the observation and projects ALL Gaussians (replicating the 73 us projection of
  2 M Gaussians is cheaper than gathering 66 MB of records over NVLink, ~95 us at the measured 770 GB/s);
* rank r bins, sorts and rasterises only the horizontal tile pairs (32x16 pixels) whose index is r modulo ``world``
  (interleaved: balanced; a pair holds four 32x8 loss blocks, so the loss kernel is striped the same way);
* the rendered depth image is summed over ranks (``all_reduce``: each pixel is non-zero on one rank, so the sum is
  exact), every rank evaluates the loss blocks of its own tile pairs (their 2-pixel halo comes from the summed image),
  replays its own tiles, and the per-tile pose partials and per-block loss partials are summed the same way; ``pose_finish`` then takes the identical Adam step on every rank.

Two exchanges per iteration, both exact, so the trajectory is bit-identical to the one-GPU tracker's.  Two
implementations: (1) all-reduces through ``torch.distributed`` (NCCL over NVLink on the GPU box) on the tracker's
stream; (2) ``fuse()``: the kernels that produce the data store it straight into every rank's buffers (peer stores over
NVLink, CUDA IPC) and a barrier kernel follows -- compute and collective in one kernel, no library call in the loop.
Neither synchronises with the host inside the loop.
"""
from __future__ import annotations

import ctypes
from typing import Callable, Optional

import torch

from . import _lib
from .tracker import PoseTracker


def tile_owner(tx, ty, tile_w: int, world: int):
    """Rank that owns raster tile (tx, ty) of a striped frame: host mirror of ``stripe_owns`` (csrc/common.cuh).  Horizontal
    tile pairs (32x16 pixels = four 32x8 loss blocks), interleaved over the ranks; ``world`` a power of two.  Works on ints
    and on numpy arrays."""
    if world & (world - 1):
        raise ValueError("world must be a power of two")
    return (ty * ((tile_w + 1) >> 1) + (tx >> 1)) & (world - 1)


def loss_block_owner(bx, by, tile_w: int, world: int):
    """Rank that evaluates the 32x8 loss block (bx, by): the owner of the tile pair it lies in."""
    return tile_owner(2 * bx, by >> 1, tile_w, world)


class _DevArray:
    """Zero-copy view of a raw device pointer for ``torch.as_tensor`` (CUDA array interface, float32)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class StripedTracker:
    """``PoseTracker`` of one rank of a tile-striped frame.  ``all_reduce(tensor)`` sums a float32 CUDA tensor in place
    over the ranks (default: ``torch.distributed.all_reduce``); tests pass their own to emulate several ranks on one GPU."""

    def __init__(self, tracker: PoseTracker, rank: int, world: int, all_reduce: Optional[Callable] = None):
        if world & (world - 1) or not 0 <= rank < world:
            raise ValueError("world must be a power of two and 0 <= rank < world")
        self.trk, self.rank, self.world = tracker, rank, world
        self._L = tracker._L
        _lib.check(self._L.gsb_tracker_set_stripe(tracker._h, rank, world), "tracker_set_stripe")
        d, p = ctypes.c_void_p(), ctypes.c_void_p()
        nd, npar = _lib.c_i64(0), _lib.c_i64(0)
        _lib.check(self._L.gsb_tracker_stripe_buffers(tracker._h, ctypes.byref(d), ctypes.byref(p), ctypes.byref(nd),
                                                      ctypes.byref(npar)), "tracker_stripe_buffers")
        with torch.cuda.device(tracker.device):
            self.depth = torch.as_tensor(_DevArray(d.value, nd.value), device=tracker.device)
            self.partials = torch.as_tensor(_DevArray(p.value, npar.value), device=tracker.device)
        if all_reduce is None:
            import torch.distributed as dist
            all_reduce = lambda t: dist.all_reduce(t)   # noqa: E731  (sum, on the current stream)
        self._all_reduce = all_reduce
        self.nvlink_bytes_per_iter = 2 * (world - 1) / world * 4 * (nd.value + npar.value) if world > 1 else 0

    # ------------------------------------------------------------------ fused exchange (peer stores instead of all-reduce)
    def _own_exchange_buffers(self):
        d, p, f = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p()
        _lib.check(self._L.gsb_tracker_stripe_exchange_buffers(self.trk._h, ctypes.byref(d), ctypes.byref(p), ctypes.byref(f)),
                   "tracker_stripe_exchange_buffers")
        return d.value, p.value, f.value

    def _set_peers(self, ed, part, flags):
        n = len(ed)
        arr = lambda v: (ctypes.c_void_p * n)(*v)   # noqa: E731
        _lib.check(self._L.gsb_tracker_stripe_set_peers(self.trk._h, n, arr(ed), arr(part), arr(flags) if flags else None),
                   "tracker_stripe_set_peers")
        self.fused = n > 0
        self._graph = None

    def fuse_with_local_ranks(self, ranks):
        """Test aid: all ranks of the stripe are ``StripedTracker`` objects of THIS process on one device and one stream;
        the kernels store into each other's buffers directly and stream order replaces the barrier kernels."""
        bufs = [r._own_exchange_buffers() for r in ranks]
        self._set_peers([b[0] for b in bufs], [b[1] for b in bufs], None)

    def fuse(self, group=None):
        """Switch to the fused exchange: raster_fwd / loss / raster_bwd store their results straight into every rank's
        buffers over NVLink (CUDA IPC handles exchanged once through ``torch.distributed``), two barrier kernels per
        iteration, NO collective call.  One process per GPU on one node."""
        import torch.distributed as dist
        if self.world == 1:
            return
        own = self._own_exchange_buffers()
        handles = []
        for ptr in own:
            h = (ctypes.c_ubyte * 64)()
            _lib.check(self._L.gsb_ipc_export(ctypes.c_void_p(ptr), h), "ipc_export")
            handles.append(bytes(h))
        gathered = [None] * self.world
        dist.all_gather_object(gathered, handles, group=group)
        tables = ([], [], [])
        self._opened = []
        for r, hs in enumerate(gathered):
            for k in range(3):
                if r == self.rank:
                    tables[k].append(own[k])
                else:
                    out = ctypes.c_void_p()
                    _lib.check(self._L.gsb_ipc_open((ctypes.c_ubyte * 64).from_buffer_copy(hs[k]), ctypes.byref(out)), "ipc_open")
                    self._opened.append(out.value)
                    tables[k].append(out.value)
        self._set_peers(*tables)
        dist.barrier(group=group)      # every rank has zeroed its flags before anybody's first barrier kernel runs

    def unfuse(self):
        self._set_peers([], [], None)
        for ptr in getattr(self, "_opened", []):
            self._L.gsb_ipc_close(ctypes.c_void_p(ptr))
        self._opened = []

    def close(self):
        """Leave the fused exchange (unmap the peers' buffers) and close the tracker."""
        if getattr(self, "fused", False) or getattr(self, "_opened", None):
            self.unfuse()
        self.trk.close()

    def segment(self, k: int):
        _lib.check(self._L.gsb_tracker_stripe_segment(self.trk._h, int(k)), "tracker_stripe_segment")

    def _one_iteration(self):
        exchange = self.world > 1 and not getattr(self, "fused", False)
        self.segment(0)
        if exchange:
            self._all_reduce(self.depth)
        self.segment(1)
        if exchange:
            self._all_reduce(self.partials)
        self.segment(2)

    def capture(self):
        """Record one iteration -- the six kernels and the two exchanges (NCCL all-reduces, or the barrier kernels of the
        fused exchange) -- as ONE CUDA graph on the tracker's stream (torch's graph capture; NCCL collectives are
        capturable).  Call after at least one eager iteration at the current frame size (the first segment sizes the
        tile bins, which synchronises).  The graph is dropped when the bins are re-sized."""
        self._graph, self._graph_cap = None, self.trk.counters()["capacity"]
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize(self.trk.device)
        with torch.cuda.graph(g, stream=self.trk.stream):
            self._one_iteration()
        self._graph = g

    def iterate(self, n: int):
        """``n`` iterations, nothing but enqueues on the tracker's stream: one graph launch each when ``capture`` has
        been called, else 3 segment calls + 2 collectives."""
        self.segment(-1)   # after set_poses / set_map: viewmats and bin sizes (a no-op otherwise)
        g = getattr(self, "_graph", None)
        if g is not None and self.trk.counters()["capacity"] != self._graph_cap:
            g = self._graph = None
        with torch.cuda.stream(self.trk.stream):
            for _ in range(n):
                if g is not None:
                    g.replay()
                else:
                    self._one_iteration()

    def finish(self) -> dict:
        over, stopped = ctypes.c_int(0), ctypes.c_int(0)
        _lib.check(self._L.gsb_tracker_stripe_finish(self.trk._h, ctypes.byref(over), ctypes.byref(stopped)),
                   "tracker_stripe_finish")
        if over.value:
            raise RuntimeError("a tile bin overflowed during the striped run: set the poses again (the bins are re-sized) "
                               "and repeat; the iterations since the overflow are invalid")
        return {"all_stopped": bool(stopped.value)}

    def run(self, n_iters: int, poll_every: int = 25) -> dict:
        """Up to ``n_iters`` iterations; the early-stop flag (identical on every rank) is polled every ``poll_every``."""
        done = 0
        while done < n_iters:
            n = min(poll_every if poll_every > 0 else n_iters, n_iters - done)
            self.iterate(n)
            done += n
            if self.finish()["all_stopped"]:
                break
        return {"iters": done}
