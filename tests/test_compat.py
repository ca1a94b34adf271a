"""The reference-name shim: the imports the three inference scripts make resolve to ramp_b200 (no GPU needed)."""
import sys


def test_reference_imports_resolve_to_ramp_b200():
    for k in [k for k in sys.modules if k == "mpd" or k.startswith("mpd.") or k == "core" or k.startswith("core.")]:
        del sys.modules[k]
    import ramp_b200
    import ramp_b200.compat as compat
    bound = compat.install(use_real_mpd=False)
    assert ("mpd.models", "StaticGaussianDiffusionModel") in bound
    from mpd.models import TemporalUnetInference, UNET_DIM_MULTS                         # inference_static.py:6
    from mpd.models.diffusion_models.sample_functions import ddpm_sample_fn              # inference_static.py:7
    from mpd.models.diffusion_models.sample_functionsdynamic import ddpm_sample_fn as f2  # inference_dynamic.py:7
    from mpd.models.diffusion_models.cost import compute_trajectory_costs
    from mpd.models.diffusion_models.APFhelper import ObstacleField, avoidance
    import mpd.models
    assert TemporalUnetInference is ramp_b200.TemporalUnetInference and UNET_DIM_MULTS[1] == (1, 2, 4, 8)
    assert ddpm_sample_fn is f2 is ramp_b200.ddpm_sample_fn
    for cls in ("StaticGaussianDiffusionModel", "DynamicGaussianDiffusionModel", "GaussianDiffusionModel3d"):
        assert getattr(mpd.models, cls) is getattr(ramp_b200, cls)     # get_model's lookup, train_loaders.py:11-25
    assert compute_trajectory_costs is ramp_b200.compute_trajectory_costs and ObstacleField is ramp_b200.ObstacleField
    assert callable(avoidance)
    from core.metrics import DynamicMetrics, Metrics                                     # inference_static.py:16, inference_dynamic.py:17
    import ramp_b200.metrics
    assert Metrics is ramp_b200.metrics.Metrics and issubclass(DynamicMetrics, Metrics)
    from core.utils import ContextManager, DynamicsGenerator, StateGenerator            # inference_dynamic.py:16
    from mpd.datasets.normalization import LimitsNormalizer
    import mpd.datasets
    import ramp_b200.datasets as D
    assert (ContextManager, DynamicsGenerator, StateGenerator) == (D.ContextManager, D.DynamicsGenerator, D.StateGenerator)
    assert LimitsNormalizer is D.LimitsNormalizer
    for cls in ("TrajectoryDataset", "ExpDirectoryDataset", "TrajectoryDataset3d"):     # get_dataset's lookup, train_loaders.py:84
        assert getattr(mpd.datasets, cls) is getattr(D, cls)
    # constructor contract of the scripts (inference_static.py:85-105): kwargs accepted, state-dict keys of the reference
    net = TemporalUnetInference(state_dim=4, n_support_points=64, unet_input_dim=32, dim_mults=UNET_DIM_MULTS[1])
    m = mpd.models.StaticGaussianDiffusionModel(model=net, variance_schedule='exponential', n_diffusion_steps=100,
                                                predict_epsilon=True, compose=True, use_apf=True, training=False,
                                                state_dim=4, n_support_points=64, unet_input_dim=32, dim_mults=(1, 2, 4, 8))
    keys = set(m.state_dict().keys())
    assert {"betas", "alphas_cumprod", "posterior_mean_coef2", "model.time_mlp.encoder.1.weight",
            "model.downs.0.3.transformer_blocks.1.attn2.to_k.weight", "model.final_conv.1.bias"} <= keys
    assert m.ddim_num_inference_steps == 8 and m.ddim is True
    for k in [k for k in sys.modules if k == "mpd" or k.startswith("mpd.")]:
        del sys.modules[k]
    for k in [k for k in sys.modules if k == "core" or k.startswith("core.")]:
        del sys.modules[k]
