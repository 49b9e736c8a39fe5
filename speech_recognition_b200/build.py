"""Build lib/libb200mel.so (sm_100a only) with nvcc.  nvcc cross-compiles without a GPU."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "b200mel_api.cu")
OUT = os.path.join(HERE, "lib", "libb200mel.so")


def nvcc_path():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libb200mel.so")


def sources():
    d = os.path.join(HERE, "csrc")
    files = [os.path.join(d, f) for f in sorted(os.listdir(d))]
    files.append(os.path.join(HERE, "..", "include", "b200mel.h"))
    return files


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(f) > t for f in sources())


def regenerate_tc_tables():
    """tc_tables_generated.inc (the tensor-core kernel's straight-line mel epilogue) is generated
    from plan.h by gen_tc_tables.cc; the result is committed, and refreshed here when a host
    compiler is available and the inputs are newer."""
    d = os.path.join(HERE, "csrc")
    out = os.path.join(d, "tc_tables_generated.inc")
    srcs = [os.path.join(d, "gen_tc_tables.cc"), os.path.join(d, "plan.h")]
    if os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(f) for f in srcs):
        return
    gxx = shutil.which("g++")
    if gxx is None:
        if os.path.exists(out):
            return
        raise RuntimeError("g++ not found and tc_tables_generated.inc is missing")
    exe = os.path.join(HERE, "lib", "gen_tc_tables")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.run([gxx, "-O1", "-std=c++17", "-o", exe, srcs[0]], check=True)
    text = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    with open(out, "w") as f:
        f.write(text)
    os.remove(exe)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = [
        nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
        "-Xcompiler", "-fPIC", "-shared", "-cudart", "static", "-o", OUT, SRC,
    ]
    if os.path.exists(os.path.join(HERE, "csrc", "tc_kernel.cuh")):
        regenerate_tc_tables()
        cmd += ["-DB200MEL_WITH_TC"]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stdout + res.stderr)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
