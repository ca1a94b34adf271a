"""ramp_b200 — B200-native (sm_100a) implementation of RAMP's reverse-diffusion planning hot path.

Drop-in names of the reference's `mpd.models` for this path; the arithmetic lives in
libramp_b200.so (hand-written CUDA behind a C ABI, include/ramp_b200.h).  There is no CPU fallback."""
from .unet import TemporalUnetInference, UNET_DIM_MULTS  # noqa: F401
from .diffusion import (StaticGaussianDiffusionModel, GaussianDiffusionModel3d, make_timesteps,  # noqa: F401
                        cosine_beta_schedule, exponential_beta_schedule)
from .dynamic import DynamicGaussianDiffusionModel  # noqa: F401
from .cost import compute_trajectory_costs, compute_collision_with_pointcloud, compute_path_length, compute_smoothness  # noqa: F401
from .apf import ObstacleField, avoidance  # noqa: F401
from .sample_functions import apply_hard_conditioning, ddpm_sample_fn, extract, guide_gradient_steps  # noqa: F401
from .metrics import Metrics, DynamicMetrics  # noqa: F401
