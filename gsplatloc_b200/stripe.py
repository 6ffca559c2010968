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

Two collectives per iteration, both exact, so the trajectory is bit-identical to the one-GPU tracker's.  The
collectives run through ``torch.distributed`` (NCCL over NVLink on the GPU box) on the tracker's stream, with no host
synchronisation inside the loop.
"""
from __future__ import annotations

import ctypes
from typing import Callable, Optional

import torch

from . import _lib
from .tracker import PoseTracker


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

    def segment(self, k: int):
        _lib.check(self._L.gsb_tracker_stripe_segment(self.trk._h, int(k)), "tracker_stripe_segment")

    def _one_iteration(self):
        self.segment(0)
        if self.world > 1:
            self._all_reduce(self.depth)
        self.segment(1)
        if self.world > 1:
            self._all_reduce(self.partials)
        self.segment(2)

    def capture(self):
        """Record one iteration -- five kernel segments' launches and the two NCCL collectives -- as ONE CUDA graph on
        the tracker's stream (torch's graph capture; NCCL collectives are capturable).  Call after at least one eager
        iteration at the current frame size (the first segment sizes the tile bins, which synchronises).  The graph is
        dropped when the bins are re-sized."""
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
