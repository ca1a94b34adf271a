"""On-disk formats either side of the planning loop: the dataset trees the three inference scripts read,
the normaliser that maps states to [-1, 1] before the loop and back after it, hard conditions, and saved contexts.

Mirrors, by name and argument meaning:
  mpd/datasets/normalization.py:13-195   DatasetNormalizer, Identity, GaussianNormalizer, LimitsNormalizer,
                                         SafeLimitsNormalizer, FixedLimitsNormalizer
  mpd/datasets/trajectories.py:33-354    TrajectoryDataset, ExpDirectoryDataset (2-D static + pursuit trees)
  mpd/datasets/trajectories3d.py:10-215  TrajectoryDataset3d
  mpd/trainer/train_loaders.py:106-160   get_dataset_explicit_dir
  scripts/inference/core/utils.py:6-137  StateGenerator, ContextManager, DynamicsGenerator

A dataset tree is `<base>/<subdir>/<id>/` holding
  static 2-D : trajs-free.pt [n, H, 2q] · obstacle_points.pt [n_obs, n_pts, 2] · box_centers.npy [n_obs, 2] ·
               metadata.yaml {env_id, robot_id, box_sizes}
  pursuit 2-D: trajs-free.pt · obstacle_pointsORG.pt · obstacle_config.npy (pickled list of {type, center, size}) ·
               args.yaml · metadata.yaml
  3-D        : trajs-free.pt [n, H, 6] · obstacle_points.pt · args.yaml {threshold_start_goal_pos} ·
               metadata.yaml {env_id, robot_id, box_centers, box_sizes, sphere_centers, sphere_radii}

The reference builds `dataset.env` / `dataset.robot` from torch_robotics classes named in metadata.yaml.  The planning
loop only touches a narrow protocol of them — `env.obj_fixed_list[0].fields[0].{centers, sizes}`,
`env.obj_extra_list[0].fields[0].{centers, radii, update_centers(t, state)}` (diffusion_model_dynamic.py:509-513) and
`robot.get_position / get_velocity` — so this module carries exactly that (`PlanningScene`, `PointMassRobot`) instead
of an SDF / rendering stack.  All of this is host-side glue: nothing here launches a kernel."""
import os

import numpy as np
import torch
import yaml
from torch.utils.data import DataLoader, Dataset, Subset

from ._lib import RampError


# --------------------------------------------------------------------------------------------- normalisers
class Normalizer:
    """normalization.py:83-107 — per-dimension min / max of a flattened [N, d] field."""

    def __init__(self, X):
        self.X = X
        self.mins = X.min(dim=0).values
        self.maxs = X.max(dim=0).values

    def __call__(self, x):
        return self.normalize(x)

    def __repr__(self):
        return "[ %s ] dim: %d\n    -: %s\n    +: %s\n" % (type(self).__name__, self.mins.numel(),
                                                         torch.round(self.mins, decimals=2), torch.round(self.maxs, decimals=2))

    def normalize(self, x):
        raise NotImplementedError

    def unnormalize(self, x):
        raise NotImplementedError


class Identity(Normalizer):
    def normalize(self, x):
        return x

    def unnormalize(self, x):
        return x


class GaussianNormalizer(Normalizer):
    """normalization.py:117-141 — zero mean, unit (unbiased) standard deviation."""

    def __init__(self, X):
        super().__init__(X)
        self.means = X.mean(dim=0)
        self.stds = X.std(dim=0)
        self.z = 1

    def normalize(self, x):
        return (x - self.means) / self.stds

    def unnormalize(self, x):
        return x * self.stds + self.means


class LimitsNormalizer(Normalizer):
    """normalization.py:144-168 — [min, max] -> [-1, 1], same operation order as the reference so results agree to the
    bit; `unnormalize` clips inputs that leave [-1, 1] by more than eps first."""

    def normalize(self, x):
        unit = (x - self.mins) / (self.maxs - self.mins)
        return 2 * unit - 1

    def unnormalize(self, x, eps=1e-4):
        if x.max() > 1 + eps or x.min() < -1 - eps:
            x = torch.clip(x, -1, 1)
        unit = (x + 1) / 2.
        return unit * (self.maxs - self.mins) + self.mins


class SafeLimitsNormalizer(LimitsNormalizer):
    """normalization.py:171-185 — a constant dimension widens the limits; as in the reference the widening applies to
    ALL dimensions (`self.mins -= eps` on the whole vector) once per constant dimension found."""

    def __init__(self, X, eps=1):
        super().__init__(X)
        for i in range(len(self.mins)):
            if self.mins[i] == self.maxs[i]:
                self.mins -= eps
                self.maxs += eps


class FixedLimitsNormalizer(LimitsNormalizer):
    """normalization.py:188-195 — limits given, not measured."""

    def __init__(self, X, min=-1, max=1):
        super().__init__(X)
        self.mins = torch.ones_like(self.mins) * min
        self.maxs = torch.ones_like(self.maxs) * max


_NORMALIZERS = {c.__name__: c for c in (Identity, GaussianNormalizer, LimitsNormalizer, SafeLimitsNormalizer,
                                        FixedLimitsNormalizer)}


def _rows(x):
    """normalization.py:62-81 — fields of 2, 3 or 4 dims flatten to [N, d]."""
    if x.ndim not in (2, 3, 4):
        raise NotImplementedError("field with %d dims" % x.ndim)
    return x.reshape(-1, x.shape[-1])


class DatasetNormalizer:
    """normalization.py:13-45 — one single-field normaliser per key.  `normalizer` is a class or its name (the reference
    `eval`s the string; here it is looked up in the table of normalisers this module defines)."""

    def __init__(self, dataset, normalizer):
        if isinstance(normalizer, str):
            if normalizer not in _NORMALIZERS:
                raise RampError("unknown normalizer %r (have %s)" % (normalizer, sorted(_NORMALIZERS)))
            normalizer = _NORMALIZERS[normalizer]
        self.normalizers = {key: normalizer(_rows(val)) for key, val in dataset.items()}

    def __repr__(self):
        return "".join("%s: %s]\n" % kv for kv in self.normalizers.items())

    def __call__(self, x, key):
        return self.normalize(x, key)

    def normalize(self, x, key):
        return self.normalizers[key].normalize(x)

    def unnormalize(self, x, key):
        return self.normalizers[key].unnormalize(x)

    def get_field_normalizers(self):
        return self.normalizers


# --------------------------------------------------------------------------------------------- scene protocol
class _Field:
    def __init__(self, **tensors):
        self.__dict__.update(tensors)


class _ObjectField:
    """`ObjectField(fields, name)` of torch_robotics: the loop reads `.fields[0]`."""

    def __init__(self, fields, name):
        self.fields = list(fields)
        self.name = name


class PursuerField:
    """deps/torch_robotics/.../primitives.py:90-107 (MultiSphereFieldDynamics): sphere centres moved by
    `dynamics_fn(t, prev_centers, current_state, velocity)` on the host, in numpy, exactly as the reference calls it."""

    def __init__(self, centers, radii, dynamics_fn=None, velocity=None, tensor_args=None):
        self.tensor_args = tensor_args or dict(device="cpu", dtype=torch.float32)
        self.initial_centers = centers
        self.centers = torch.as_tensor(np.asarray(centers), **self.tensor_args)
        self.radii = torch.as_tensor(np.asarray(radii), **self.tensor_args)
        self.velocity = None if velocity is None else torch.as_tensor(np.asarray(velocity), **self.tensor_args)
        self.dynamics_fn = dynamics_fn

    def update_centers(self, t, current_state=None):
        if self.dynamics_fn is None or t is None or current_state is None:
            return
        state = current_state.detach().cpu().numpy() if torch.is_tensor(current_state) else np.asarray(current_state)
        new = self.dynamics_fn(t, self.centers.detach().cpu().numpy(), state, self.velocity.detach().cpu().numpy())
        self.centers = torch.as_tensor(new, **self.tensor_args)


class PlanningScene:
    """What the loop needs of a torch_robotics environment: fixed boxes (and spheres in 3-D), optional moving sphere."""

    def __init__(self, name, box_centers, box_sizes, sphere_centers=None, sphere_radii=None, pursuer=None, dim=2,
                 tensor_args=None):
        self.name = name
        self.dim = dim
        self.tensor_args = tensor_args
        lim = torch.ones(dim, **(tensor_args or {}))
        self.limits = torch.stack([-lim, lim])
        fixed = [_Field(centers=box_centers, sizes=box_sizes)]
        if sphere_centers is not None:
            fixed.append(_Field(centers=sphere_centers, radii=sphere_radii))
        self.obj_fixed_list = [_ObjectField(fixed, name)]
        self.obj_extra_list = [] if pursuer is None else [_ObjectField([pursuer], name + "Obstacles")]


class PointMassRobot:
    """robot_base.py:147-156 of torch_robotics: state = (q, q̇); a missing velocity half is a central finite difference."""

    def __init__(self, q_dim, dt=1.0, name="RobotPointMass", tensor_args=None):
        self.q_dim = q_dim
        self.dt = dt
        self.name = name
        self.tensor_args = tensor_args

    def get_position(self, x):
        return x[..., :self.q_dim]

    def get_velocity(self, x):
        vel = x[..., self.q_dim:2 * self.q_dim]
        if x.nelement() != 0 and vel.nelement() == 0:
            pos = x[..., :self.q_dim]
            vel = torch.zeros_like(pos)
            vel[..., 1:-1, :] = (pos[..., 2:, :] - pos[..., :-2, :]) / (2 * self.dt)
        return vel


def _yaml(path, tolerant=False):
    """`load_params_from_yaml` (mpd/utils/loading.py:4-6, FullLoader) for static and 3-D trees; `safe_load_yaml` for
    pursuit trees (trajectories.py:17-30: SafeLoader on which unknown tags read as None)."""
    loader = yaml.FullLoader
    if tolerant:
        class loader(yaml.SafeLoader):
            pass
        loader.add_constructor(None, lambda self, node: None)
    with open(path, "r") as fh:
        return yaml.load(fh, Loader=loader)


# --------------------------------------------------------------------------------------------- dataset trees
def _load_pt(path, device):
    return torch.load(path, map_location=device)


def _read_static_2d(d, ta):
    """trajectories.py:128-137 — boxes from box_centers.npy + metadata.yaml['box_sizes']."""
    meta = _yaml(os.path.join(d, "metadata.yaml"))
    return dict(trajs=_load_pt(os.path.join(d, "trajs-free.pt"), ta["device"]),
                cloud=_load_pt(os.path.join(d, "obstacle_points.pt"), ta["device"]),
                box_centers=torch.from_numpy(np.load(os.path.join(d, "box_centers.npy"))).to(**ta),
                box_sizes=torch.tensor(meta["box_sizes"], **ta))


def _read_pursuit_2d(d, ta):
    """trajectories.py:138-156 — boxes are the 'box' entries of the pickled obstacle_config.npy."""
    cfg = np.load(os.path.join(d, "obstacle_config.npy"), allow_pickle=True)
    boxes = [o for o in cfg if o["type"] == "box"]
    return dict(trajs=_load_pt(os.path.join(d, "trajs-free.pt"), ta["device"]),
                cloud=_load_pt(os.path.join(d, "obstacle_pointsORG.pt"), ta["device"]),
                box_centers=torch.tensor(np.array([o["center"] for o in boxes]), **ta),
                box_sizes=torch.tensor(np.array([o["size"] for o in boxes]), **ta))


def _read_3d(d, ta):
    """trajectories3d.py:78-87 — every primitive comes from metadata.yaml."""
    meta = _yaml(os.path.join(d, "metadata.yaml"))
    return dict(trajs=_load_pt(os.path.join(d, "trajs-free.pt"), ta["device"]),
                cloud=_load_pt(os.path.join(d, "obstacle_points.pt"), ta["device"]),
                box_centers=torch.tensor(meta["box_centers"], **ta), box_sizes=torch.tensor(meta["box_sizes"], **ta),
                sphere_centers=torch.tensor(meta["sphere_centers"], **ta),
                sphere_radii=torch.tensor(meta["sphere_radii"], **ta))


# EnvPredator.py:22-38 — six square pillars on a 3 x 2 grid; the dynamic loop plans around the first four (:510-511)
_PX, _PY = 0.5575312972068787, 0.5508843064308167
PREDATOR_ARENA_BOX_CENTERS = [[_PX, -_PY], [_PX, _PY], [-_PX, _PY], [-_PX, -_PY], [0.0, -_PY], [0.0, _PY]]
PREDATOR_ARENA_BOX_SIDE = 0.26

_LAYOUTS = {
    "static2d": (("trajs-free.pt", "obstacle_points.pt", "box_centers.npy"), _read_static_2d),
    "pursuit2d": (("trajs-free.pt", "obstacle_pointsORG.pt", "obstacle_config.npy"), _read_pursuit_2d),
    "3d": (("trajs-free.pt", "obstacle_points.pt", "metadata.yaml"), _read_3d),
}
_SCENE_KEYS = ("box_centers", "box_sizes", "sphere_centers", "sphere_radii")


class _TrajectoryTree(Dataset):
    """Shared body of the three dataset classes: gather entries, build scene + robot, normalise `traj` and `task`."""

    layout = "static2d"
    scene_dim = 2

    def __init__(self, dataset_subdir=None, include_velocity=False, normalizer="LimitsNormalizer",
                 use_extra_objects=False, obstacle_cutoff_margin=None, tensor_args=None, dynamics_fn=None, velocity=None,
                 training=False, static=True, dataset_base_dir=None, base_dataset_dir=None, **kwargs):
        base = dataset_base_dir or base_dataset_dir
        if base is None:
            raise RampError("dataset_base_dir is required (the reference defaults to its author's home directory)")
        self.tensor_args = tensor_args or dict(device="cpu", dtype=torch.float32)
        self.dataset_subdir = dataset_subdir
        self.base_dir = os.path.join(base, dataset_subdir)
        self.training = training
        self.static = static
        self.include_velocity = include_velocity
        self.use_extra_objects = use_extra_objects
        self.dynamics_fn = dynamics_fn
        self.velocity = velocity
        self.field_key_traj, self.field_key_task = "traj", "task"
        self.fields = {}
        self.map_task_id_to_trajectories_id = {}
        self.map_trajectory_id_to_task_id = {}
        self._read_head_files()
        self._gather(self._entry_dirs())
        self.env = self._make_scene(kwargs)
        q_dim = self.scene_dim
        self.robot = PointMassRobot(q_dim, name=str(self.metadata.get("robot_id", "RobotPointMass")), tensor_args=self.tensor_args)
        self._make_fields()
        self.min_x = self.fields["traj"].min()
        self.max_x = self.fields["traj"].max()
        b, h, d = self.dataset_shape = self.fields["traj"].shape
        self.n_trajs, self.n_support_points, self.state_dim = b, h, d
        self.trajectory_dim = (h, d)
        self.normalizer = DatasetNormalizer(self.fields, normalizer=normalizer)
        self.normalizer_keys = [self.field_key_traj, self.field_key_task]
        self.normalize_all_data(*self.normalizer_keys)

    # ---- reading
    def _read_head_files(self):
        """trajectories.py:55-60 — entry '0' names the environment and the robot."""
        head = os.path.join(self.base_dir, "0")
        tolerant = self.layout == "pursuit2d"
        self.metadata = _yaml(os.path.join(head, "metadata.yaml"), tolerant)
        args = os.path.join(head, "args.yaml")
        self.args = _yaml(args, tolerant) if os.path.exists(args) else None

    def _entry_dirs(self):
        """`os.walk(base_dir, topdown=True)` order, keeping the directories that hold all required files."""
        need, _ = _LAYOUTS[self.layout]
        return [d for d, _, files in os.walk(self.base_dir, topdown=True) if all(f in files for f in need)]

    def _gather(self, dirs):
        _, read = _LAYOUTS[self.layout]
        trajs, per_traj = [], {k: [] for k in ("cloud",) + _SCENE_KEYS}
        n = 0
        for task_id, d in enumerate(dirs):
            e = read(d, self.tensor_args)
            m = e["trajs"].shape[0]
            idx = n + np.arange(m)
            self.map_task_id_to_trajectories_id[task_id] = idx
            self.map_trajectory_id_to_task_id.update({j: task_id for j in idx})
            n += m
            trajs.append(e["trajs"])
            for k in per_traj:
                if k in e:
                    per_traj[k].append(e[k].unsqueeze(0).expand(m, *e[k].shape))
        if not trajs:
            raise RampError("no dataset entries with %s under %s" % (_LAYOUTS[self.layout][0], self.base_dir))
        self.trajs_free = torch.cat(trajs)
        self.obstacle_points = torch.cat(per_traj["cloud"])
        self.box_centers = torch.cat(per_traj["box_centers"])
        self.box_sizes = torch.cat(per_traj["box_sizes"])
        if per_traj["sphere_centers"]:
            self.sphere_centers = torch.cat(per_traj["sphere_centers"])
            self.sphere_radi = torch.cat(per_traj["sphere_radii"])
        self.n_trajs = n

    def _make_scene(self, kwargs):
        """trajectories.py:80-93: `<env_id>Obstacles` adds the moving sphere (EnvPredatorExtra.py:14-42: one sphere of
        radius 0.05 at `pursuer_pos`, default the origin)."""
        env_id = str(self.metadata.get("env_id", "Env"))
        pursuer = None
        if self.use_extra_objects and self.scene_dim == 2:
            pos = kwargs.get("pursuer_pos") if "Predator" in env_id else None
            pursuer = PursuerField(np.array([pos if pos is not None else [0.0, 0.0]]), np.array([0.05]),
                                   dynamics_fn=self.dynamics_fn, velocity=self.velocity, tensor_args=self.tensor_args)
        spheres = getattr(self, "sphere_centers", None)
        if env_id.startswith("EnvPredator"):
            # the pursuit arena is fixed geometry, whatever the entry's obstacle_config says (EnvPredator.py:22-38)
            ta = self.tensor_args
            return PlanningScene(env_id + ("Obstacles" if self.use_extra_objects else ""),
                                 torch.tensor(PREDATOR_ARENA_BOX_CENTERS, **ta),
                                 torch.full((len(PREDATOR_ARENA_BOX_CENTERS), 2), PREDATOR_ARENA_BOX_SIDE, **ta),
                                 pursuer=pursuer, dim=2, tensor_args=ta)
        return PlanningScene(env_id + ("Obstacles" if self.use_extra_objects else ""), self.box_centers[0], self.box_sizes[0],
                             None if spheres is None else spheres[0], None if spheres is None else self.sphere_radi[0],
                             pursuer=pursuer, dim=self.scene_dim, tensor_args=self.tensor_args)

    def _make_fields(self):
        """trajectories.py:182-198 — `traj` (positions, or positions + velocities) and `task` (start ++ goal position)."""
        pos = self.robot.get_position(self.trajs_free)
        self.fields["traj"] = self.trajs_free if self.include_velocity else pos
        self.fields["task"] = torch.cat((pos[..., 0, :], pos[..., -1, :]), dim=-1)
        self.fields["box_centers"] = self.box_centers
        if not (len(self.fields["traj"]) == len(self.obstacle_points) == len(self.box_centers)):
            raise RampError("trajectories, obstacle points and box centers disagree in count")

    def normalize_all_data(self, *keys):
        for key in keys:
            self.fields[key + "_normalized"] = self.normalizer(self.fields[key], key)

    # ---- torch Dataset protocol
    def __repr__(self):
        return "TrajectoryDataset\nn_trajs: %d\ntrajectory_dim: %s\n" % (self.n_trajs, (self.trajectory_dim,))

    def __len__(self):
        return self.n_trajs

    def _item_extras(self, index):
        return {"box_sizes": self.box_sizes[index]} if not self.training else {}

    def _item_has_hard_conds(self):
        return not self.static

    def __getitem__(self, index):
        traj = self.fields["traj_normalized"][index]
        data = {"traj_normalized": traj, "task_normalized": self.fields["task_normalized"][index],
                "obstacle_points": self.obstacle_points[index], "box_centers": self.box_centers[index]}
        data.update(self._item_extras(index))
        if self._item_has_hard_conds():
            data["hard_conds"] = self.get_hard_conditions(traj, horizon=len(traj))
        return data

    # ---- hard conditions and (un)normalisation
    def get_hard_conditions(self, traj, horizon=None, normalize=False):
        """trajectories.py:282-307 — {0: start, horizon-1: goal}; with velocities in the state both ends are at rest."""
        ends = [self.robot.get_position(traj[0]), self.robot.get_position(traj[-1])]
        if self.include_velocity:
            ends = [torch.cat((p, torch.zeros_like(p)), dim=-1) for p in ends]
        if normalize:
            ends = [self.normalizer.normalize(s, key=self.field_key_traj) for s in ends]
        horizon = self.n_support_points if horizon is None else horizon
        return {0: ends[0], horizon - 1: ends[1]}

    def unnormalize(self, x, key):
        return self.normalizer.unnormalize(x, key)

    def normalize(self, x, key):
        return self.normalizer.normalize(x, key)

    def unnormalize_trajectories(self, x):
        return self.unnormalize(x, self.field_key_traj)

    def normalize_trajectories(self, x):
        return self.normalize(x, self.field_key_traj)

    def unnormalize_tasks(self, x):
        return self.unnormalize(x, self.field_key_task)

    def normalize_tasks(self, x):
        return self.normalize(x, self.field_key_task)


class TrajectoryDataset(_TrajectoryTree):
    """mpd/datasets/trajectories.py:277-307 — every entry under the tree; `static=False` selects the pursuit layout."""

    def __init__(self, static=True, **kwargs):
        self.layout = "static2d" if static else "pursuit2d"
        super().__init__(static=static, **kwargs)


class ExpDirectoryDataset(_TrajectoryTree):
    """mpd/datasets/trajectories.py:310-354 — ONE entry (`current_dir_id`) of a static tree: the environment an
    experiment runs in.  Its trajectories only size the normaliser."""

    def __init__(self, current_dir_id=0, **kwargs):
        self.specified_dir_id = int(current_dir_id)
        super().__init__(**kwargs)

    def _entry_dirs(self):
        return [os.path.join(self.base_dir, str(self.specified_dir_id))]

    def get_random_traj_from_current_dir(self):
        return {"dir_id": self.specified_dir_id, "n_trajs": self.n_trajs}


class TrajectoryDataset3d(_TrajectoryTree):
    """mpd/datasets/trajectories3d.py:10-215 — boxes and spheres; items always carry hard conditions (:160-161)."""

    layout = "3d"
    scene_dim = 3

    def __init__(self, training=True, **kwargs):
        super().__init__(training=training, **kwargs)
        self.threshold_start_goal_pos = (self.args or {}).get("threshold_start_goal_pos")

    def _item_extras(self, index):
        return {"box_sizes": self.box_sizes[index], "sphere_centers": self.sphere_centers[index],
                "sphere_radii": self.sphere_radi[index]}

    def _item_has_hard_conds(self):
        return True


def get_dataset_explicit_dir(dataset_class=None, dataset_subdir=None, batch_size=2, val_set_size=0.0, results_dir=None,
                             save_indices=False, current_dir=0, **kwargs):
    """mpd/trainer/train_loaders.py:106-160 — (train_subset, train_loader, val_subset, val_loader), unshuffled."""
    cls = {c.__name__: c for c in (TrajectoryDataset, ExpDirectoryDataset, TrajectoryDataset3d)}[dataset_class]
    full = cls(current_dir_id=current_dir, dataset_subdir=dataset_subdir, **kwargs)
    n = len(full)
    n_train = n - int(val_set_size * n)
    idx = list(range(n))
    train, val = Subset(full, idx[:n_train]), Subset(full, idx[n_train:])
    if save_indices and results_dir:
        out = os.path.join(results_dir, "dir_%s" % current_dir)
        os.makedirs(out, exist_ok=True)
        torch.save(idx[:n_train], os.path.join(out, "train_subset_indices.pt"))
        torch.save(idx[n_train:], os.path.join(out, "val_subset_indices.pt"))
    return (train, DataLoader(train, batch_size=batch_size, shuffle=False),
            val, DataLoader(val, batch_size=batch_size, shuffle=False))


# --------------------------------------------------------------------------------------------- script-side helpers
class StateGenerator:
    """scripts/inference/core/utils.py:6-27."""

    @staticmethod
    def get_hard_cond_custom(traj, horizon, include_velocity=True):
        ends = [traj[0], traj[-1]]
        if include_velocity:
            ends = [torch.cat((p, torch.zeros_like(p)), dim=-1) for p in ends]
        return {0: ends[0], horizon - 1: ends[1]}


class ContextManager:
    """scripts/inference/core/utils.py:29-79 — `<env_dir>/contexts/context_###.pt` =
    {start_pos, goal_pos, metadata: {context_idx, dataset_id}}."""

    @staticmethod
    def save_context(start_state_pos, goal_state_pos, env_dir, dataset_id, context_idx):
        if not torch.is_tensor(start_state_pos) or not torch.is_tensor(goal_state_pos):
            raise ValueError("Start and goal positions must be torch tensors")
        if not isinstance(context_idx, int) or context_idx < 0:
            raise ValueError("Invalid context_idx: %s" % (context_idx,))
        contexts_dir = os.path.join(env_dir, "contexts")
        os.makedirs(contexts_dir, exist_ok=True)
        path = os.path.join(contexts_dir, "context_%03d.pt" % context_idx)
        torch.save({"start_pos": start_state_pos.cpu(), "goal_pos": goal_state_pos.cpu(),
                    "metadata": {"context_idx": context_idx, "dataset_id": dataset_id}}, path)
        return path

    @staticmethod
    def load_context(contexts_dir, context_idx, device="cpu"):
        cfg = torch.load(os.path.join(contexts_dir, "context_%03d.pt" % context_idx), map_location=device)
        return cfg["start_pos"].to(device), cfg["goal_pos"].to(device)


class DynamicsGenerator:
    """scripts/inference/core/utils.py:81-137 — the pursuer's motion law.  `dynamics_fn(t, prev_center, robot_position,
    velocity_input)` -> new centre [1, 2] (float64 numpy): a unit vector toward the robot
    weighted `pursuit_strength`, plus `random_strength` · (sin 2πt, cos 2πt), times the per-axis speed, times dt = 0.1,
    clipped to the arena [-1, 1]².  `ramp_b200.pursuit.DevicePursuer` is the same law as a CUDA kernel."""

    dt = 0.1

    @staticmethod
    def create_pursuit_dynamics(velocity_max=0.5, pursuit_strength=0.8, random_strength=0.2):
        velocity = np.array([[velocity_max / np.sqrt(2), velocity_max / np.sqrt(2)]])

        def dynamics_fn(t, prev_center, robot_position, velocity_input):
            # dtypes flow as in the reference's scalar code: the direction to the robot is computed in the inputs' own
            # precision (float32 when the centres come from float32 tensors), the wobble and the update in float64
            c = np.asarray(prev_center)[0]
            to_robot = np.asarray(robot_position)[0][:2] - c
            dist = np.sqrt(to_robot[0] ** 2 + to_robot[1] ** 2)
            if dist > 0:
                to_robot = to_robot / dist
            wobble = np.array([np.sin(2 * np.pi * t), np.cos(2 * np.pi * t)])
            heading = pursuit_strength * to_robot + random_strength * wobble
            new = c + heading * np.asarray(velocity_input)[0] * DynamicsGenerator.dt
            return np.clip(new, -1, 1)[None, :]

        return dynamics_fn, velocity
