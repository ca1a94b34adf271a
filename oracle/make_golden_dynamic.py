"""Generate tests/golden/dynamic_tick.pt: the UNMODIFIED reference DynamicGaussianDiffusionModel
(diffusion_model_dynamic.py) run on CPU with a synthetic pursuit environment, a noise tape and name-seeded weights.
Stage I + the first N_TICKS replan ticks are recorded (inputs, clouds, the batch handed to
compute_trajectory_costs, the selected trajectory).  Run in the build container only."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.make_golden import hc2d, ref_model, save  # noqa: E402
from oracle.ref_harness import NoiseTape  # noqa: E402
from ramp_b200.testing import SyntheticPursuitDataset, pursuit_cloud_2d  # noqa: E402

N_TICKS = 3
B = 5        # odd on purpose: the stacked CFG rows then alternate proper / inverted guidance (UnetInference.py:192-195)


class _Stop(Exception):
    pass


def main():
    torch.set_num_threads(8)
    model = ref_model("dynamic", 100)
    import mpd.models.diffusion_models.diffusion_model_dynamic as dmd
    from mpd.models.diffusion_models.sample_functionsdynamic import ddpm_sample_fn
    ds = SyntheticPursuitDataset()
    cloud = pursuit_cloud_2d(ds)
    calls = []
    orig = dmd.compute_trajectory_costs

    def recording(trajs, pts, **kw):
        out = orig(trajs, pts, **kw)
        print('cost call', len(calls), 'cloud', tuple(pts.shape), 'thr', kw.get('collision_threshold'), 'free', int(out[3].sum()), 'max|x|', float(trajs[..., :2].abs().max()))
        field = ctx.get('obstacle_field')
        calls.append(dict(batch=trajs.clone(), cloud=pts.clone().float(), thr=kw.get('collision_threshold'),
                          best=None if out[0] is None else out[0].clone(), free=out[3].clone(),
                          disc=None if field is None or field.dynamic_kdtree is None else
                          torch.from_numpy(np.asarray(field.dynamic_kdtree.data)).float(),
                          center=None if field is None or field.dynamic_kdtree is None else
                          torch.from_numpy(np.asarray(field.dynamic_center)).float(),
                          static_pts=None if field is None else torch.from_numpy(field.static_obstacle_points).float(),
                          draws=tape.k))
        if len(calls) == 1 + N_TICKS:
            raise _Stop()
        return out

    dmd.compute_trajectory_costs = recording
    ctx = {'dataset': ds}
    orig_copy = dmd.copy
    dmd.copy = lambda o: ctx if o is ctx else orig_copy(o)     # run_inference shallow-copies the context; keep ours
    np.random.seed(0)
    tape = NoiseTape(500)
    try:
        with tape.patched():
            model.run_inference(ctx, hc2d(), n_samples=B, horizon=64, return_chain=True,
                                traj_normalized=torch.zeros(64, 4), obstacle_pts=cloud, sample_fn=ddpm_sample_fn)
    except _Stop:
        pass
    finally:
        dmd.compute_trajectory_costs = orig
        dmd.copy = orig_copy
    assert len(calls) == 1 + N_TICKS and all(c['best'] is not None for c in calls)
    out = dict(cloud=cloud, noise_seed=500, B=B, T=100, n_ticks=N_TICKS, stage1_batch=calls[0]['batch'],
               stage1_best=calls[0]['best'], static_pts=calls[1]['static_pts'])
    for k in range(N_TICKS):
        c = calls[1 + k]
        out['tick%d' % k] = dict(batch=c['batch'], cost_cloud=c['cloud'], best=c['best'], free=c['free'], disc=c['disc'],
                                 center=c['center'], draw_index=calls[k]['draws'])
    save("dynamic_tick", out)


if __name__ == "__main__":
    main()
