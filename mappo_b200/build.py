"""In-tree build of the C-ABI library: nvcc, sm_100a only, one .so next to this file.

    python -m mappo_b200.build [--force] [-v]

The .so is git-ignored but travels to the GPU box with the repo snapshot.  No torch dependency: the
library's ABI is plain C (include/mappo_b200.h)."""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libmappo_b200.so")
BUILD = os.path.join(CSRC, "build")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]
# Sources whose arithmetic must round like numpy's separately-rounded ops (no FMA contraction):
# the f64 physics, the f32 GAE scan / ValueNorm and the f64 rasteriser (bit-exact vs oracle/render.py).
NO_FMA = {"env_mpe.cu", "gae.cu", "render.cu"}


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(paths, flags):
    h = hashlib.sha256(" ".join(flags).encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def build(force=False, verbose=False):
    os.makedirs(BUILD, exist_ok=True)
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(os.path.dirname(HERE), "include", "mappo_b200.h"))
    objs, rebuilt, procs = [], False, []
    for src in sources():
        sp = os.path.join(CSRC, src)
        obj = os.path.join(BUILD, src[:-3] + ".o")
        stamp = obj + ".sha"
        flags = ARCH + COMMON + (["-fmad=false"] if src in NO_FMA else [])
        dig = _digest([sp] + hdrs, flags)
        objs.append(obj)
        if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
            continue
        cmd = ["nvcc"] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", sp, "-o", obj]
        procs.append((src, stamp, dig,
                      subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, stamp, dig, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (src, out))
        if verbose and out:
            print("== %s\n%s" % (src, out))
        with open(stamp, "w") as f:
            f.write(dig)
        rebuilt = True
    if rebuilt or force or not os.path.exists(OUT):
        cmd = ["nvcc"] + ARCH + ["-shared", "-o", OUT] + objs + ["-lcudart"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s" % r.stdout)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
