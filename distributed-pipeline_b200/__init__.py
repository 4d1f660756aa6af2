"""B200-native data-parallel update path for the ``TrainLoop`` scaffold of
YAIconvalley/distributed-pipeline: flat state buffers, hand-written sm_100a
kernels behind a C ABI (``include/dpk.h``), NCCL bucket all-reduce.

The directory is named ``distributed-pipeline_b200`` (as the project is); import
it as ``distributed_pipeline_b200`` -- the one-file loader at the repository
root maps the importable name onto this directory.
"""
from . import _lib, dist_util, logger
from ._lib import DpkError
from .engine import Engine, ema_update
from .flat import FlatLayout, FlatState
from .optim import FusedAdamW
from .reducer import BucketReducer, FlatDDP
from .trainer import TrainLoop, update_ema

__all__ = ["TrainLoop", "update_ema", "FusedAdamW", "Engine", "FlatLayout", "FlatState", "BucketReducer",
           "FlatDDP", "DpkError", "dist_util", "logger", "ema_update"]
