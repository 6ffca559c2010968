"""gsplatloc_b200 — B200-native (sm_100a) depth-splat renderer and pose optimiser behind gsplat's
``rasterization()`` signature, for the hot path GsplatLoc drives in ``src/my_gsplat``.

Importing the package never touches the GPU; the first op call loads ``lib/libgsplatloc_b200.so`` and raises if
it is missing (no CPU / PyTorch fallback exists).
"""
from . import ops
from .rendering import rasterization
from .datatypes import AlignData, TensorWrapper
from .tracker import PoseTracker, TrackerConfig

__all__ = ["rasterization", "ops", "PoseTracker", "TrackerConfig", "AlignData", "TensorWrapper"]
__version__ = "0.1.0"
