"""B200-native ArcFace head: a drop-in for ``ArcMarginProduct`` of
XiaoyuWant/DDP_Classification_pytorch (ARCFACE/arc_main.py:130-176) whose arithmetic runs in
hand-written sm_100a CUDA kernels behind a C ABI (``include/arcface_b200.h``)."""
from ._lib import LIB_PATH, build, lib  # noqa: F401
from .head import ArcMarginProduct, shard_bounds  # noqa: F401
from .optim import HeadAdam, HeadSGD  # noqa: F401

__all__ = ["ArcMarginProduct", "HeadSGD", "HeadAdam", "shard_bounds", "build", "lib", "LIB_PATH"]
