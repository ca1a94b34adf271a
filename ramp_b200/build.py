"""In-tree build of the CUDA library (nvcc, sm_100a only).  `python -m ramp_b200.build`."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libramp_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math=false",
         "-Xcompiler", "-fPIC", "-cudart", "static"]
FLAGS = [f for f in FLAGS if f != "--use_fast_math=false"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "ramp_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, planes_bf16=False, variant=None, defines=()):
    """planes_bf16: round 1's operand-plane format (bf16 + bf16, csrc/common.cuh) as a SECOND library,
    libramp_b200_bf16.so, selected at run time with RAMP_B200_LIB=<that path> for A/B runs; the default library
    (fp16 + fp16 planes) is untouched."""
    if planes_bf16:
        variant, defines = "bf16", ("RAMP_PLANES_FP16=0",)
    if os.environ.get("RAMP_B200_LIB") and not variant:
        return os.environ["RAMP_B200_LIB"]
    if not force and not variant and not needs_build():
        return OUT
    # variant: a SECOND library libramp_b200_<variant>.so compiled with extra -D defines, for same-box A/B runs
    target = os.path.join(HERE, "libramp_b200_%s.so" % variant) if variant else OUT
    bdir = os.path.join(HERE, "build", variant) if variant else os.path.join(HERE, "build")
    extra = ["-D" + d for d in defines]
    objs = []
    procs = []
    os.makedirs(bdir, exist_ok=True)
    for src in sources():
        obj = os.path.join(bdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [NVCC] + FLAGS + extra + ["-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd))
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (src, out.decode()))
    tmp = target + ".tmp"          # link beside the target, then rename: a snapshot never sees a half-written library
    cmd = [NVCC, "-shared", "-cudart", "static", "-o", tmp] + objs + ["-ldl", "-lpthread", "-lrt"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout.decode())
    os.replace(tmp, target)
    return target


if __name__ == "__main__":
    var = sys.argv[sys.argv.index("--variant") + 1] if "--variant" in sys.argv else None
    defs = tuple(a[2:] for a in sys.argv if a.startswith("-D"))
    print(build(force="--force" in sys.argv, verbose=True, planes_bf16="--planes-bf16" in sys.argv, variant=var, defines=defs))
