"""B200-native drop-in for the DiffMimic sampling hot path (sample -> condition -> score).

Public names mirror the reference modules they replace:
  diffuser.models.temporal.TemporalUnet            -> TemporalUnet
  diffuser.models.diffusion.GaussianDiffusion      -> GaussianDiffusion (sampling half), Sample, InjectedNoise
  diffuser.models.sampling_config.apply_conditioning* -> sampling_config
  data_loaders.motion_dataset.MotionDataset        -> MotionDataset
  data_loaders.motion_dataset_v2.MotionDataset     -> MotionDatasetV2 (69-wide qpos | qvel rows)
  metrics.fid_score.MotionFID                      -> MotionFID
  diffuser.models.diffusion_v4.DiffusionV4 (sample) -> DiffusionV4
"""
from ._lib import DmError, build, lib  # noqa: F401
from .diffusion import GaussianDiffusion, InjectedNoise, Sample, default_sample_fn  # noqa: F401
from .sampling_config import (apply_conditioning, apply_conditioning_motion_editing,  # noqa: F401
                              trace_conditioning)
from .temporal import TemporalUnet  # noqa: F401
from .motion_dataset import Batch, MotionDataset, MotionDatasetV2, save_motion_npy, save_motions  # noqa: F401
from .scoring import MotionFID  # noqa: F401
from .diffusion_v4 import DiffusionV4  # noqa: F401
