"""Generate tests/golden/datasets.pt with the UNMODIFIED reference dataset / normaliser / script-helper code
(mpd/datasets/*, scripts/inference/core/utils.py under /root/reference) on small synthetic dataset trees.  The fixture
holds the tree contents themselves (so a test can write the same files into a temp dir) and what the reference made of
them.  Build container only — matplotlib / gitpython are not installed, so they are stubbed (neither is used by the
code that runs here).  Test infrastructure: nothing under ramp_b200/ imports this."""
import contextlib
import importlib.util
import io
import os
import sys
import tempfile
import types

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
REF = "/root/reference"


def _stub_absent_modules():
    import ref_harness
    ref_harness._stub_matplotlib()

    class _S(types.ModuleType):
        def __getattr__(self, n):
            if n.startswith("__"):
                raise AttributeError(n)
            return _S(self.__name__ + "." + n)

        def __call__(self, *a, **k):
            return _S("call")

    for n in ("git", "mpl_toolkits", "mpl_toolkits.mplot3d", "mpl_toolkits.mplot3d.art3d"):
        sys.modules.setdefault(n, _S(n))


def smooth_trajs(n, H, q, seed):
    g = torch.Generator("cpu").manual_seed(seed)
    t = torch.linspace(0, 1, H)[None, :, None]
    a = torch.rand(n, 1, q, generator=g) * 1.6 - 0.8
    b = torch.rand(n, 1, q, generator=g) * 1.6 - 0.8
    pos = a + (b - a) * t + 0.2 * torch.cumsum(torch.randn(n, H, q, generator=g), 1) / H ** 0.5 * (t * (1 - t) * 4)
    vel = torch.diff(pos, dim=1, prepend=pos[:, :1]) * 10
    return torch.cat([pos, vel], -1).contiguous()


def make_trees():
    """{subdir: {entry id: {file name: content}}} — tensors, numpy arrays, or dicts (yaml)."""
    g = torch.Generator("cpu").manual_seed(3)
    trees = {"static": {}, "pursuit": {}, "env3d": {}}
    for e in range(3):
        centers = (torch.rand(6, 2, generator=g) * 1.4 - 0.7).numpy().astype(np.float64)
        sizes = (torch.rand(6, 2, generator=g) * 0.2 + 0.1).tolist()
        trees["static"][str(e)] = {
            "trajs-free.pt": smooth_trajs(4 + e, 64, 2, 10 + e),
            "obstacle_points.pt": torch.rand(6, 64, 2, generator=g) * 2 - 1,
            "box_centers.npy": centers,
            "metadata.yaml": {"env_id": "EnvParameterizedBox", "robot_id": "RobotPointMass", "box_sizes": sizes},
        }
    for e in range(2):
        cfg = [{"type": "box", "center": [0.5 * (-1) ** k, 0.4 * (-1) ** (k // 2)], "size": [0.26, 0.26]} for k in range(4)]
        cfg.append({"type": "sphere", "center": [0.0, 0.0], "radius": 0.05})
        trees["pursuit"][str(e)] = {
            "trajs-free.pt": smooth_trajs(3 + e, 48, 2, 20 + e),
            "obstacle_pointsORG.pt": torch.rand(6, 64, 2, generator=g) * 2 - 1,
            "obstacle_config.npy": np.array(cfg, dtype=object),
            "args.yaml": {"threshold_start_goal_pos": 1.0},
            "metadata.yaml": {"env_id": "EnvPredator", "robot_id": "RobotPointMass"},
        }
    for e in range(2):
        trees["env3d"][str(e)] = {
            "trajs-free.pt": smooth_trajs(3 + e, 64, 3, 30 + e),
            "obstacle_points.pt": torch.rand(25, 64, 3, generator=g) * 2 - 1,
            "args.yaml": {"threshold_start_goal_pos": 1.5},
            "metadata.yaml": {"env_id": "EnvParameterized3D", "robot_id": "RobotPointMass3D",
                              "box_centers": (torch.rand(5, 3, generator=g) * 1.4 - 0.7).tolist(),
                              "box_sizes": (torch.rand(5, 3, generator=g) * 0.2 + 0.1).tolist(),
                              "sphere_centers": (torch.rand(4, 3, generator=g) * 1.4 - 0.7).tolist(),
                              "sphere_radii": (torch.rand(4, generator=g) * 0.1 + 0.05).tolist()},
        }
    return trees


def write_trees(trees, base):
    for sub, entries in trees.items():
        for eid, files in entries.items():
            d = os.path.join(base, sub, eid)
            os.makedirs(d, exist_ok=True)
            for name, content in files.items():
                p = os.path.join(d, name)
                if name.endswith(".pt"):
                    torch.save(content, p)
                elif name.endswith(".npy"):
                    np.save(p, content, allow_pickle=True)
                else:
                    with open(p, "w") as fh:
                        yaml.safe_dump(content, fh)


def snapshot(ds, items=(0, -1)):
    """What a caller can observe of a dataset object."""
    out = dict(n_trajs=ds.n_trajs, n_support_points=ds.n_support_points, state_dim=ds.state_dim,
               fields={k: v.clone() for k, v in ds.fields.items()},
               mins={k: n.mins.clone() for k, n in ds.normalizer.normalizers.items()},
               maxs={k: n.maxs.clone() for k, n in ds.normalizer.normalizers.items()},
               obstacle_points=ds.obstacle_points.clone(), box_centers=ds.box_centers.clone(), box_sizes=ds.box_sizes.clone(),
               task_of_traj=[int(ds.map_trajectory_id_to_task_id[j]) for j in range(ds.n_trajs)]
               if ds.map_trajectory_id_to_task_id else None,
               items={}, hard_conds={})
    n = len(ds)
    for i in items:
        i = i % n
        out["items"][i] = ds[i]
    t = ds.fields["traj"][0]
    for norm in (False, True):
        try:
            out["hard_conds"][norm] = ds.get_hard_conditions(t, horizon=None, normalize=norm)
        except NotImplementedError:                      # ExpDirectoryDataset inherits the abstract method
            out["hard_conds"] = None
            break
    g = torch.Generator("cpu").manual_seed(99)
    probe = torch.rand(7, ds.n_support_points, ds.state_dim, generator=g) * 2.4 - 1.2       # leaves [-1, 1]: clip branch
    out["probe"] = probe
    out["probe_unnormalized"] = ds.unnormalize_trajectories(probe)
    out["probe_roundtrip"] = ds.normalize_trajectories(ds.unnormalize_trajectories(probe * 0.8))
    fixed = ds.env.obj_fixed_list[0].fields[0]
    out["env_box_centers"] = fixed.centers.clone()
    out["env_box_sizes"] = fixed.sizes.clone()
    return out


def main():
    _stub_absent_modules()
    sys.path.insert(0, REF)
    with contextlib.redirect_stdout(io.StringIO()):
        from mpd.datasets import ExpDirectoryDataset, TrajectoryDataset, TrajectoryDataset3d
        import mpd.datasets.normalization as N
    spec = importlib.util.spec_from_file_location("ref_core_utils", os.path.join(REF, "scripts/inference/core/utils.py"))
    U = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(U)

    ta = dict(device="cpu", dtype=torch.float32)
    trees = make_trees()
    gold = dict(trees=trees, datasets={})
    with tempfile.TemporaryDirectory() as base, contextlib.redirect_stdout(io.StringIO()):
        write_trees(trees, base)
        dyn_fn, vel = U.DynamicsGenerator.create_pursuit_dynamics(velocity_max=0.5)
        cases = {
            "exp_dir_1": (ExpDirectoryDataset, dict(dataset_subdir="static", current_dir_id=1, include_velocity=True,
                                                    tensor_args=ta, dataset_base_dir=base)),
            "exp_dir_0_pos_only": (ExpDirectoryDataset, dict(dataset_subdir="static", current_dir_id=0, include_velocity=False,
                                                             tensor_args=ta, dataset_base_dir=base)),
            "static_all": (TrajectoryDataset, dict(dataset_subdir="static", include_velocity=True, tensor_args=ta,
                                                   dataset_base_dir=base, training=True)),
            "pursuit_all": (TrajectoryDataset, dict(dataset_subdir="pursuit", include_velocity=True, tensor_args=ta,
                                                    dataset_base_dir=base, static=False, use_extra_objects=True,
                                                    dynamics_fn=dyn_fn, velocity=vel, pursuer_pos=[-0.35, -0.95])),
            "env3d_all": (TrajectoryDataset3d, dict(dataset_subdir="env3d", include_velocity=True, tensor_args=ta,
                                                    base_dataset_dir=base)),
        }
        for name, (cls, kw) in cases.items():
            ds = cls(**kw)
            snap = snapshot(ds)
            snap["kwargs"] = {k: v for k, v in kw.items() if k not in ("tensor_args", "dataset_base_dir", "base_dataset_dir",
                                                                      "dynamics_fn", "velocity")}
            snap["cls"] = cls.__name__
            if name == "pursuit_all":
                sphere = ds.env.obj_extra_list[0].fields[0]
                snap["pursuer_radii"] = sphere.radii.clone()
                track = [sphere.centers.clone()]
                robot = torch.tensor([[0.8, 0.7]])
                for t in range(12):
                    sphere.update_centers(t, robot - 0.05 * t)
                    track.append(sphere.centers.clone())
                snap["pursuer_track"] = torch.stack(track)
            if name == "env3d_all":
                sph = ds.env.obj_fixed_list[0].fields[1]
                snap["env_sphere_centers"], snap["env_sphere_radii"] = sph.centers.clone(), sph.radii.clone()
            gold["datasets"][name] = snap

    # single-field normalisers on one [N, d] matrix with a constant column (SafeLimits) and out-of-range probes
    g = torch.Generator("cpu").manual_seed(5)
    X = torch.randn(200, 4, generator=g) * torch.tensor([1.0, 0.3, 2.0, 1.0]) + torch.tensor([0.0, 1.0, -2.0, 0.5])
    Xc = X.clone()
    Xc[:, 3] = 0.25
    probe = torch.rand(9, 4, generator=g) * 2.6 - 1.3
    norms = {}
    with contextlib.redirect_stdout(io.StringIO()):
        for cname, data in (("Identity", X), ("GaussianNormalizer", X), ("LimitsNormalizer", X),
                            ("SafeLimitsNormalizer", Xc), ("FixedLimitsNormalizer", X)):
            n = getattr(N, cname)(data)
            norms[cname] = dict(X=data, normalized=n.normalize(data), unnormalized=n.unnormalize(probe), mins=n.mins.clone(),
                                maxs=n.maxs.clone())
    gold["normalizers"] = dict(probe=probe, cases=norms)

    # pursuer motion law, float32 and float64 inputs
    dyn = []
    for vmax, ps, rs in ((0.5, 0.8, 0.2), (0.3, 1.0, 0.0), (0.9, 0.5, 0.5)):
        fn, vel = U.DynamicsGenerator.create_pursuit_dynamics(vmax, ps, rs)
        for dt_ in (np.float32, np.float64):
            c = np.array([[-0.35, -0.95]], dtype=dt_)
            r = np.array([[0.6, 0.7]], dtype=dt_)
            track = [c.copy()]
            for t in range(40):
                c = fn(t, c.astype(dt_), r - 0.01 * t * np.array([[1.0, 0.5]], dtype=dt_), vel.astype(dt_))
                track.append(np.array(c))
            dyn.append(dict(args=(vmax, ps, rs), dtype=np.dtype(dt_).name, velocity=vel, track=np.concatenate(track)))
        same = fn(3, np.array([[0.2, 0.2]], dtype=np.float32), np.array([[0.2, 0.2]], dtype=np.float32), vel.astype(np.float32))
        dyn.append(dict(args=(vmax, ps, rs), dtype="coincident", velocity=vel, track=np.array(same)))
    gold["dynamics"] = dyn

    # hard conditions helper and the context file format
    traj = smooth_trajs(1, 64, 2, 77)[0, :, :2]
    gold["state_generator"] = dict(traj=traj, with_vel=U.StateGenerator.get_hard_cond_custom(traj, 64, True),
                                   without_vel=U.StateGenerator.get_hard_cond_custom(traj, 64, False))
    with tempfile.TemporaryDirectory() as d, contextlib.redirect_stdout(io.StringIO()):
        p = U.ContextManager.save_context(traj[0], traj[-1], d, "static", 7)
        gold["context"] = dict(relpath=os.path.relpath(p, d), content=torch.load(p), start=traj[0], goal=traj[-1])

    path = os.path.join(ROOT, "tests", "golden", "datasets.pt")
    torch.save(gold, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
