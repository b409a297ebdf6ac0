"""TEST INFRASTRUCTURE ONLY -- builds the *unmodified* reference CUDA extensions.

Compiles lizuoyue/torch-ngp's three hot-path extensions straight from the sources where they
lie under /root/reference (nothing is copied into this repository) into ``oracle/_ref/``:

    _gridencoder  <- gridencoder/src/{gridencoder.cu,bindings.cpp}
    _raymarching  <- raymarching/src/{raymarching.cu,bindings.cpp}
    _ffmlp        <- ffmlp/src/{ffmlp.cu,bindings.cpp} (+ vendored CUTLASS 2.8 headers)
    _shencoder    <- shencoder/src/{shencoder.cu,bindings.cpp}

The only deviation from the reference's own flags (gridencoder/backend.py:6-9) is
``-std=c++17`` (torch 2.11's ATen refuses C++14) and an explicit sm_100a ``-gencode``.
The resulting .so files are git-ignored but travel to the GPU box with the gpurun snapshot,
where tests/ use them as the strongest parity oracle (the reference itself, on the same
inputs).  /root/reference does not exist on the GPU box, so this script is a no-op there.

Usage:  python oracle/build_ref_ext.py [gridencoder raymarching ffmlp shencoder]
"""
import os
import subprocess
import sys
import sysconfig
import time

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("NGP_REFERENCE_ROOT", "/root/reference")

EXTS = {
    "gridencoder": dict(name="_gridencoder", srcs=["gridencoder/src/gridencoder.cu", "gridencoder/src/bindings.cpp"], inc=[]),
    "raymarching": dict(name="_raymarching", srcs=["raymarching/src/raymarching.cu", "raymarching/src/bindings.cpp"], inc=[]),
    "shencoder": dict(name="_shencoder", srcs=["shencoder/src/shencoder.cu", "shencoder/src/bindings.cpp"], inc=[]),
    "ffmlp": dict(name="_ffmlp", srcs=["ffmlp/src/ffmlp.cu", "ffmlp/src/bindings.cpp"],
                  inc=["ffmlp/dependencies/cutlass/include", "ffmlp/dependencies/cutlass/tools/util/include"]),
}


def _torch_flags():
    import torch
    from torch.utils import cpp_extension as ce
    incs = ce.include_paths("cuda") if "device_type" in ce.include_paths.__code__.co_varnames else ce.include_paths(True)
    incs = list(incs) + [sysconfig.get_paths()["include"]]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    return incs, libdir, abi


def build_one(key):
    spec = EXTS[key]
    name = spec["name"]
    so = os.path.join(OUT, name + ".so")
    if os.path.exists(so):
        print(f"[ref-ext] {so} exists, skipping")
        return so
    if not os.path.isdir(REF):
        print(f"[ref-ext] {REF} not present (GPU box?) -- nothing to build for {name}")
        return None
    os.makedirs(OUT, exist_ok=True)
    incs, libdir, abi = _torch_flags()
    inc_flags = [f"-I{p}" for p in incs] + [f"-I{os.path.join(REF, p)}" for p in spec["inc"]]
    common = ["-O3", "-std=c++17", f"-DTORCH_EXTENSION_NAME={name}", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={abi}"]
    objs = []
    t0 = time.time()
    for src in spec["srcs"]:
        src_path = os.path.join(REF, src)
        obj = os.path.join(OUT, f"{name}_{os.path.basename(src)}.o")
        if src.endswith(".cu"):
            cmd = ["nvcc", "-c", src_path, "-o", obj, *common, *inc_flags,
                   "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__", "-U__CUDA_NO_HALF2_OPERATORS__",
                   "--expt-relaxed-constexpr", "--expt-extended-lambda",
                   "-gencode", "arch=compute_100a,code=sm_100a",
                   "-Xcompiler", "-fPIC", "-w"]
            if key == "ffmlp":
                cmd += ["-Xcompiler=-mf16c", "-Xcompiler=-Wno-float-conversion", "-Xcompiler=-fno-strict-aliasing"]
        else:
            cmd = ["g++", "-c", src_path, "-o", obj, *common, *inc_flags, "-fPIC", "-w",
                   "-I/usr/local/cuda/include"]
        print("[ref-ext]", " ".join(cmd[:6]), "...", flush=True)
        subprocess.check_call(cmd)
        objs.append(obj)
    cmd = ["g++", "-shared", "-o", so, *objs, f"-L{libdir}", "-L/usr/local/cuda/lib64",
           "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart",
           f"-Wl,-rpath,{libdir}"]
    subprocess.check_call(cmd)
    for o in objs:
        os.remove(o)
    print(f"[ref-ext] built {so} in {time.time() - t0:.0f}s", flush=True)
    return so


# The reference's own L1 Python wrappers (autograd Functions + nn.Modules above the pybind11 ops).  They are staged -- byte
# for byte, as build OUTPUT next to the compiled extensions, git-ignored like them -- so that the GPU box (which has no
# /root/reference) can run them on top of this repository's backend.py shims (tests/test_reference_wrappers.py): the one
# place where forward_buffer, the [L,B,C] layout and fp16 grad_weights are passed exactly the way the reference passes them.
WRAPPERS = {
    "gridencoder": ["gridencoder/grid.py"],
    "raymarching": ["raymarching/raymarching.py"],
    "ffmlp": ["ffmlp/ffmlp.py"],
    "shencoder": ["shencoder/sphere_harmonics.py"],
}


def stage_wrappers():
    """-> list of staged paths under oracle/_ref/py/ (empty when /root/reference is absent)."""
    import shutil
    if not os.path.isdir(REF):
        return []
    out = []
    for pkg, files in WRAPPERS.items():
        dst_dir = os.path.join(OUT, "py", pkg)
        os.makedirs(dst_dir, exist_ok=True)
        for f in files:
            dst = os.path.join(dst_dir, os.path.basename(f))
            shutil.copyfile(os.path.join(REF, f), dst)
            out.append(dst)
    print(f"[ref-ext] staged {len(out)} reference wrapper modules under {os.path.join(OUT, 'py')}", flush=True)
    return out


def build_all(keys=None, jobs=None):
    """The extensions side by side: the wall clock is the slowest one (ffmlp with its CUTLASS headers, ~8 min on 8 cores)
    instead of the sum (~25 min).  Returns {key: path or exception}."""
    import concurrent.futures as cf
    keys = list(keys or EXTS)
    try:
        stage_wrappers()
    except OSError as e:  # the checker's convenience must not fail the build
        print(f"[ref-ext] could not stage the reference wrappers: {e}", flush=True)
    out = {}
    with cf.ThreadPoolExecutor(max_workers=jobs or len(keys)) as ex:
        futs = {ex.submit(build_one, k): k for k in keys}
        for f in cf.as_completed(futs):
            try:
                out[futs[f]] = f.result()
            except Exception as e:  # noqa: BLE001 -- reported by the caller
                out[futs[f]] = e
    return out


def main(argv):
    res = build_all(argv or None)
    bad = {k: v for k, v in res.items() if isinstance(v, Exception)}
    if bad:
        raise SystemExit(f"[ref-ext] failed: {bad}")


if __name__ == "__main__":
    main(sys.argv[1:])
