"""SURVEY §8f f3 — dataset trees, normalisers, hard conditions, contexts and the pursuer's motion law, against what the
UNMODIFIED reference classes produced from the same files (tests/golden/datasets.pt, made by
oracle/make_golden_datasets.py in the build container).  Host-side code: runs without a GPU."""
import os

import numpy as np
import pytest
import torch
import yaml

GOLD = os.path.join(os.path.dirname(__file__), "golden", "datasets.pt")


@pytest.fixture(scope="module")
def gold():
    return torch.load(GOLD, weights_only=False)


@pytest.fixture(scope="module")
def tree_root(gold, tmp_path_factory):
    base = tmp_path_factory.mktemp("dataset")
    for sub, entries in gold["trees"].items():
        for eid, files in entries.items():
            d = base / sub / eid
            d.mkdir(parents=True)
            for name, content in files.items():
                p = str(d / name)
                if name.endswith(".pt"):
                    torch.save(content, p)
                elif name.endswith(".npy"):
                    np.save(p, content, allow_pickle=True)
                else:
                    with open(p, "w") as fh:
                        yaml.safe_dump(content, fh)
    return str(base)


def same(a, b, path=""):
    if torch.is_tensor(a):
        assert torch.is_tensor(b) and a.shape == b.shape and a.dtype == b.dtype, path
        assert torch.equal(a, b), (path, (a.double() - b.double()).abs().max().item())
    elif isinstance(a, dict):
        assert set(a) == set(b), (path, set(a) ^ set(b))
        for k in a:
            same(a[k], b[k], "%s[%r]" % (path, k))
    else:
        assert a == b, path


def _build(name, snap, base):
    from ramp_b200 import datasets as D
    ta = dict(device="cpu", dtype=torch.float32)
    kw = dict(snap["kwargs"], tensor_args=ta)
    kw["base_dataset_dir" if snap["cls"] == "TrajectoryDataset3d" else "dataset_base_dir"] = base
    if name == "pursuit_all":
        kw["dynamics_fn"], kw["velocity"] = D.DynamicsGenerator.create_pursuit_dynamics(velocity_max=0.5)
    return getattr(D, snap["cls"])(**kw)


@pytest.mark.parametrize("name", ["exp_dir_1", "exp_dir_0_pos_only", "static_all", "pursuit_all", "env3d_all"])
def test_dataset_matches_reference(gold, tree_root, name):
    snap = gold["datasets"][name]
    ds = _build(name, snap, tree_root)
    assert (ds.n_trajs, ds.n_support_points, ds.state_dim, len(ds)) == (snap["n_trajs"], snap["n_support_points"],
                                                                      snap["state_dim"], snap["n_trajs"])
    same(snap["fields"], ds.fields, "fields")
    for k in snap["mins"]:
        same(snap["mins"][k], ds.normalizer.normalizers[k].mins, "mins")
        same(snap["maxs"][k], ds.normalizer.normalizers[k].maxs, "maxs")
    same(snap["obstacle_points"], ds.obstacle_points.contiguous(), "obstacle_points")
    same(snap["box_centers"], ds.box_centers.contiguous(), "box_centers")
    same(snap["box_sizes"], ds.box_sizes.contiguous(), "box_sizes")
    if snap["task_of_traj"] is not None:
        assert [int(ds.map_trajectory_id_to_task_id[j]) for j in range(ds.n_trajs)] == snap["task_of_traj"]
    for i, item in snap["items"].items():
        same(item, ds[i], "item %d" % i)
    if snap["hard_conds"] is not None:
        for norm, hc in snap["hard_conds"].items():
            same(hc, ds.get_hard_conditions(ds.fields["traj"][0], horizon=None, normalize=norm), "hard_conds")
    same(snap["probe_unnormalized"], ds.unnormalize_trajectories(snap["probe"]), "unnormalize")
    same(snap["probe_roundtrip"], ds.normalize_trajectories(ds.unnormalize_trajectories(snap["probe"] * 0.8)), "roundtrip")
    fixed = ds.env.obj_fixed_list[0].fields[0]
    same(snap["env_box_centers"], fixed.centers, "env centers")
    same(snap["env_box_sizes"], fixed.sizes, "env sizes")
    if name == "env3d_all":
        sph = ds.env.obj_fixed_list[0].fields[1]
        same(snap["env_sphere_centers"], sph.centers)
        same(snap["env_sphere_radii"], sph.radii)
        assert ds.threshold_start_goal_pos == 1.5
    if name == "pursuit_all":
        sphere = ds.env.obj_extra_list[0].fields[0]
        same(snap["pursuer_radii"], sphere.radii)
        track = [sphere.centers.clone()]
        robot = torch.tensor([[0.8, 0.7]])
        for t in range(12):
            sphere.update_centers(t, robot - 0.05 * t)
            track.append(sphere.centers.clone())
        same(snap["pursuer_track"], torch.stack(track), "pursuer track")


def test_single_field_normalisers(gold):
    from ramp_b200 import datasets as D
    probe = gold["normalizers"]["probe"]
    for cname, c in gold["normalizers"]["cases"].items():
        n = getattr(D, cname)(c["X"])
        same(c["mins"], n.mins, cname)
        same(c["maxs"], n.maxs, cname)
        same(c["normalized"], n.normalize(c["X"]), cname)
        same(c["unnormalized"], n.unnormalize(probe), cname)
    with pytest.raises(D.RampError):
        D.DatasetNormalizer({"traj": torch.zeros(2, 3, 4)}, "os.system")       # names are looked up, never evaluated
    with pytest.raises(NotImplementedError):
        D.DatasetNormalizer({"traj": torch.zeros(3)}, "LimitsNormalizer")


def test_pursuit_dynamics_law(gold):
    from ramp_b200.datasets import DynamicsGenerator
    for case in gold["dynamics"]:
        fn, vel = DynamicsGenerator.create_pursuit_dynamics(*case["args"])
        assert np.array_equal(vel, case["velocity"])
        if case["dtype"] == "coincident":                     # pursuer on top of the robot: no division by zero
            dt_ = np.float32
            out = fn(3, np.array([[0.2, 0.2]], dtype=dt_), np.array([[0.2, 0.2]], dtype=dt_), vel.astype(dt_))
            assert np.array_equal(out, case["track"])
            continue
        dt_ = np.dtype(case["dtype"]).type
        c = np.array([[-0.35, -0.95]], dtype=dt_)
        r = np.array([[0.6, 0.7]], dtype=dt_)
        track = [c.copy()]
        for t in range(40):
            c = fn(t, c.astype(dt_), r - 0.01 * t * np.array([[1.0, 0.5]], dtype=dt_), vel.astype(dt_))
            track.append(np.array(c))
        got = np.concatenate(track)
        assert got.dtype == case["track"].dtype
        assert np.array_equal(got, case["track"]), np.abs(got - case["track"]).max()


def test_state_generator_and_context_files(gold, tmp_path):
    from ramp_b200.datasets import ContextManager, StateGenerator
    sg = gold["state_generator"]
    same(sg["with_vel"], StateGenerator.get_hard_cond_custom(sg["traj"], 64, True))
    same(sg["without_vel"], StateGenerator.get_hard_cond_custom(sg["traj"], 64, False))
    ctx = gold["context"]
    p = ContextManager.save_context(ctx["start"], ctx["goal"], str(tmp_path), "static", 7)
    assert os.path.relpath(p, str(tmp_path)) == ctx["relpath"]
    same(ctx["content"], torch.load(p))
    s, g = ContextManager.load_context(os.path.dirname(p), 7)
    same(ctx["start"], s)
    same(ctx["goal"], g)
    with pytest.raises(ValueError):
        ContextManager.save_context([0.0, 0.0], ctx["goal"], str(tmp_path), "static", 1)
    with pytest.raises(ValueError):
        ContextManager.save_context(ctx["start"], ctx["goal"], str(tmp_path), "static", -1)


def test_explicit_dir_loader_and_errors(gold, tree_root, tmp_path):
    from ramp_b200 import datasets as D
    ta = dict(device="cpu", dtype=torch.float32)
    train, tl, val, vl = D.get_dataset_explicit_dir("ExpDirectoryDataset", "static", batch_size=2, val_set_size=0.25,
                                                    results_dir=str(tmp_path), save_indices=True, current_dir=2,
                                                    include_velocity=True, tensor_args=ta, dataset_base_dir=tree_root)
    n = gold["trees"]["static"]["2"]["trajs-free.pt"].shape[0]
    assert len(train) + len(val) == n and len(val) == int(0.25 * n)
    assert torch.load(os.path.join(str(tmp_path), "dir_2", "train_subset_indices.pt")) == list(range(len(train)))
    batch = next(iter(tl))
    assert batch["traj_normalized"].shape == (2, 64, 4) and batch["obstacle_points"].shape == (2, 6, 64, 2)
    assert train[0]["traj_normalized"].abs().max() <= 1.0
    with pytest.raises(D.RampError):
        D.TrajectoryDataset(dataset_subdir="static", tensor_args=ta)                       # no base directory
    (tmp_path / "empty" / "0").mkdir(parents=True)
    with open(str(tmp_path / "empty" / "0" / "metadata.yaml"), "w") as fh:
        yaml.safe_dump({"env_id": "EnvParameterizedBox", "robot_id": "RobotPointMass"}, fh)
    with pytest.raises(D.RampError):
        D.TrajectoryDataset(dataset_subdir="empty", tensor_args=ta, dataset_base_dir=str(tmp_path))
