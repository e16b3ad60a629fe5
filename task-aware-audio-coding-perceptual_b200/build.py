"""Build libtac_spectral.so (hand-written sm_100a CUDA + the C ABI of include/tac_spectral.h) in-tree.

    python task-aware-audio-coding-perceptual_b200/build.py [--force]

nvcc cross-compiles without a GPU; the .so lands next to the Python host package
(tac_spectral/libtac_spectral.so) so that it travels with the source tree."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "tac_spectral", "libtac_spectral.so")
SOURCES = ["tables.cu", "stft.cu", "stft_ex.cu", "loss.cu", "loss_stream.cu", "dpca.cu", "istft.cu", "latent.cu", "exchange.cu", "tail.cu", "chain.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--use_fast_math=false"]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "tac_spectral.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, extra=None, out=None):
    """extra: additional nvcc flags (experiments, e.g. ["-DTAC_BWD_OCC=2"]); out: alternative .so path."""
    global OUT
    if out:
        OUT = out
        force = True
    if not force and not needs_build():
        return OUT
    objs = []
    procs = []
    flags = [f for f in NVCC_FLAGS if not f.startswith("--use_fast_math")] + list(extra or [])
    tag = str(abs(hash(tuple(extra or [])))) if extra else ""
    for src in SOURCES:
        obj = os.path.join("/tmp", "tac_" + tag + src.replace(".cu", ".o"))
        cmd = [_nvcc()] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{out}")
        if verbose:
            print(out)
    cmd = [_nvcc(), "-shared", "-o", OUT] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}")
    return OUT


if __name__ == "__main__":
    extra = [a for a in sys.argv[1:] if a.startswith("-D")]
    out = next((a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--out=")), None)
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, extra=extra, out=out))
