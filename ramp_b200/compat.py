"""Register ramp_b200 under the reference's module names, so `inference_static.py`, `inference_dynamic.py` and
`inference3d.py` resolve the hot path here without an edit:

    import ramp_b200.compat as compat; compat.install()
    from mpd.models import TemporalUnetInference, UNET_DIM_MULTS                      # inference_static.py:6
    from mpd.models.diffusion_models.sample_functions import ddpm_sample_fn           # inference_static.py:7
    model = getattr(mpd.models, 'StaticGaussianDiffusionModel')(model=unet, ...)      # train_loaders.py:11-25

If the real `mpd` package is importable it is extended in place (its other sub-packages — datasets, trainer, utils —
keep working); otherwise light module objects are created.  Only names on the reverse-diffusion path are rebound."""
import importlib
import sys
import types

from . import apf, apf_dynamic, cost, datasets, diffusion, dynamic, metrics, sample_functions, unet

_HOT = {
    "mpd.models": dict(
        TemporalUnetInference=unet.TemporalUnetInference, UNET_DIM_MULTS=unet.UNET_DIM_MULTS,
        StaticGaussianDiffusionModel=diffusion.StaticGaussianDiffusionModel,
        GaussianDiffusionModel3d=diffusion.GaussianDiffusionModel3d,
        DynamicGaussianDiffusionModel=dynamic.DynamicGaussianDiffusionModel,
        ddpm_sample_fn=sample_functions.ddpm_sample_fn, apply_hard_conditioning=sample_functions.apply_hard_conditioning,
        extract=sample_functions.extract, compute_trajectory_costs=cost.compute_trajectory_costs),
    "mpd.models.diffusion_models": dict(
        StaticGaussianDiffusionModel=diffusion.StaticGaussianDiffusionModel,
        GaussianDiffusionModel3d=diffusion.GaussianDiffusionModel3d,
        DynamicGaussianDiffusionModel=dynamic.DynamicGaussianDiffusionModel,
        TemporalUnetInference=unet.TemporalUnetInference, UNET_DIM_MULTS=unet.UNET_DIM_MULTS),
    "mpd.models.diffusion_models.sample_functions": dict(
        ddpm_sample_fn=sample_functions.ddpm_sample_fn, apply_hard_conditioning=sample_functions.apply_hard_conditioning,
        extract=sample_functions.extract),
    "mpd.models.diffusion_models.sample_functionsdynamic": dict(
        ddpm_sample_fn=sample_functions.ddpm_sample_fn, apply_hard_conditioning=sample_functions.apply_hard_conditioning,
        extract=sample_functions.extract, guide_gradient_steps=sample_functions.guide_gradient_steps),
    "mpd.models.diffusion_models.cost": dict(
        compute_trajectory_costs=cost.compute_trajectory_costs,
        compute_collision_with_pointcloud=cost.compute_collision_with_pointcloud,
        compute_path_length=cost.compute_path_length, compute_smoothness=cost.compute_smoothness),
    "mpd.models.diffusion_models.APFhelper": dict(ObstacleField=apf.ObstacleField, avoidance=apf.avoidance),
    "mpd.models.diffusion_models.APFhelper_dynamic": dict(
        ObstacleField=apf_dynamic.ObstacleField, avoidance=apf_dynamic.avoidance,
        generate_sphere_points=apf_dynamic.generate_sphere_points, generate_box_points=apf_dynamic.generate_box_points),
    "mpd.models.diffusion_models.UnetInference": dict(
        TemporalUnetInference=unet.TemporalUnetInference, UNET_DIM_MULTS=unet.UNET_DIM_MULTS),
    "mpd.models.diffusion_models.diffusion_model_static": dict(
        StaticGaussianDiffusionModel=diffusion.StaticGaussianDiffusionModel, make_timesteps=diffusion.make_timesteps),
    "mpd.models.diffusion_models.diffusion_model_3d": dict(
        GaussianDiffusionModel3d=diffusion.GaussianDiffusionModel3d, make_timesteps=diffusion.make_timesteps),
    "mpd.models.diffusion_models.diffusion_model_dynamic": dict(
        DynamicGaussianDiffusionModel=dynamic.DynamicGaussianDiffusionModel, make_timesteps=diffusion.make_timesteps),
    # `from core.metrics import Metrics` / `DynamicMetrics` (inference_static.py:16, inference_dynamic.py:17)
    "core.metrics": dict(Metrics=metrics.Metrics, DynamicMetrics=metrics.DynamicMetrics),
    # formats either side of the loop: `getattr(datasets, dataset_class)` (train_loaders.py:84,118),
    # `from core.utils import StateGenerator, ContextManager` (inference_static.py:15)
    "mpd.datasets": dict(TrajectoryDataset=datasets.TrajectoryDataset, ExpDirectoryDataset=datasets.ExpDirectoryDataset,
                         TrajectoryDataset3d=datasets.TrajectoryDataset3d),
    "mpd.datasets.normalization": dict(
        DatasetNormalizer=datasets.DatasetNormalizer, Normalizer=datasets.Normalizer, Identity=datasets.Identity,
        GaussianNormalizer=datasets.GaussianNormalizer, LimitsNormalizer=datasets.LimitsNormalizer,
        SafeLimitsNormalizer=datasets.SafeLimitsNormalizer, FixedLimitsNormalizer=datasets.FixedLimitsNormalizer),
    "core.utils": dict(StateGenerator=datasets.StateGenerator, ContextManager=datasets.ContextManager,
                       DynamicsGenerator=datasets.DynamicsGenerator),
}


def _module(name, use_real):
    if name in sys.modules:
        return sys.modules[name]
    if use_real:
        try:
            return importlib.import_module(name)
        except Exception:
            pass
    m = types.ModuleType(name)
    m.__path__ = []                      # behaves as a package for sub-module imports
    sys.modules[name] = m
    parent, _, leaf = name.rpartition(".")
    if parent:
        setattr(_module(parent, use_real), leaf, m)
    return m


def install(use_real_mpd=True):
    """Rebind the hot-path names.  Returns the list of (module, name) pairs that were bound."""
    bound = []
    for mod_name, names in _HOT.items():
        m = _module(mod_name, use_real_mpd)
        for k, v in names.items():
            setattr(m, k, v)
            bound.append((mod_name, k))
    return bound
