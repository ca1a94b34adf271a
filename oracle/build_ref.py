"""Recipe for oracle/_ref/: an UNTRACKED, unmodified copy of the reference's Python sources for the hot path, so that
`bench.py --impl reference` and `bench.py`'s `reference_gpu_eager` leg can run the reference ITSELF on the GPU box
(where /root/reference does not exist).  oracle/_ref/ is listed in .gitignore (never committed) but not in
.gpurunignore (it travels with the snapshot, like the built .so).

What is copied: `mpd/` (the models, samplers, APF / cost helpers the path uses) and the three `torch_robotics` files
`mpd.models` imports (`torch_utils/torch_utils.py` + package markers).  Nothing is edited; matplotlib is stubbed at import
time by oracle/ref_harness.py exactly as for the golden fixtures.

Called from `__graft_entry__.build()` in the build container.  TEST / BENCH INFRASTRUCTURE — never imported by the product.
"""
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("RAMP_REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")
PARTS = ["mpd",
         "deps/torch_robotics/torch_robotics/__init__.py",
         "deps/torch_robotics/torch_robotics/torch_utils/__init__.py",
         "deps/torch_robotics/torch_robotics/torch_utils/torch_utils.py"]


def build_ref(verbose=False):
    """Copies the reference sources into oracle/_ref/.  Returns the path, or None when the reference tree is absent
    (the GPU box: the copy made in the build container is used as it is)."""
    if not os.path.isdir(os.path.join(SRC, "mpd")):
        return DST if os.path.isdir(os.path.join(DST, "mpd")) else None
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    n = 0
    for part in PARTS:
        s, d = os.path.join(SRC, part), os.path.join(DST, part)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        if os.path.isdir(s):
            shutil.copytree(s, d, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
            n += sum(len(f) for _, _, f in os.walk(d))
        else:
            shutil.copy2(s, d)
            n += 1
    with open(os.path.join(DST, "README"), "w") as f:
        f.write("Unmodified copy of %s (wondmgezahu/RAMP) made by oracle/build_ref.py; untracked, bench/test use only.\n" % SRC)
    if verbose:
        print("oracle/_ref: %d files copied from %s" % (n, SRC))
    return DST


def ref_root():
    """Where an importable reference tree is: oracle/_ref if it was built, else the build container's /root/reference."""
    if os.path.isdir(os.path.join(DST, "mpd")):
        return DST
    if os.path.isdir(os.path.join(SRC, "mpd")):
        return SRC
    return None


if __name__ == "__main__":
    print(build_ref(verbose=True))
