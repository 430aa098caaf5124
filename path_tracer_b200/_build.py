"""In-tree builds of the native libraries.

* ``libptb.so``        CUDA C ABI (include/ptb.h), sm_100a only      -> build_cuda()
* ``libptb_scene.so``  host scene layer (CPU, mirrors scene.hpp API)  -> build_scene()

The built files are git-ignored but travel to the GPU box with gpurun.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(ROOT)
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

HOST_SOURCES = ["scene.cpp", "pack.cpp", "spectrum.cpp", "procedural.cpp", "model.cpp", "image.cpp", "serializer.cpp", "capi.cpp"]
CUDA_SOURCES = ["ptb_api.cu", "ptb_kernels.cu", "ptb_gpu_build.cu", "ptb_spectrum.cu", "ptb_group.cu", "ptb_noalloc.cu", "ptb_wide_build.cpp"]

SPECTRUM_TABLE = os.path.join(ROOT, "data", "sRGBSpectrumTable.dat")


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd, cwd):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd))
    return r.stdout


def build_scene(force=False):
    d = os.path.join(ROOT, "host")
    out = os.path.join(ROOT, "libptb_scene.so")
    deps = [os.path.join(d, f) for f in os.listdir(d)] + [os.path.join(REPO, "include", "ptb.h")]
    if force or _newer(out, deps):
        _run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-pthread", "-shared", "-o", out] + HOST_SOURCES + ["-lz"], d)
    return out


def cuda_flags():
    return ["-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
            "--fmad=false",            # numerical contract with the oracle: no implicit contraction
            "-Xcompiler", "-fPIC"]


def _includes(path, seen):
    """Transitive closure of the quoted #includes of one source (for per-object rebuilds)."""
    if path in seen or not os.path.exists(path):
        return
    seen.add(path)
    with open(path, errors="replace") as f:
        for line in f:
            if line.startswith("#include \""):
                name = line.split('"')[1]
                for base in (os.path.dirname(path), os.path.join(REPO, "include")):
                    _includes(os.path.normpath(os.path.join(base, name)), seen)


def source_hash():
    """sha256 (12 hex digits) over the sources the trace / scatter kernels are compiled from (ptb_kernels.cu and every
    header it includes, transitively) and the compile flags: names the build that profiles/traffic.json's ncu figures
    were captured from, so bench.py never quotes counters of another kernel.  (Host-side files of the library - the API,
    the group code, the builders - do not change those kernels and are not part of it.)"""
    import hashlib
    d = os.path.join(ROOT, "csrc")
    deps = set()
    _includes(os.path.join(d, "ptb_kernels.cu"), deps)
    h = hashlib.sha256(" ".join(cuda_flags()).encode())
    for f in sorted(deps):
        h.update(os.path.relpath(f, REPO).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:12]


def build_cuda(force=False, verbose=False):
    """One object per source (rebuilt only when it or a header it includes changed), then one link."""
    d = os.path.join(ROOT, "csrc")
    out = os.path.join(ROOT, "libptb.so")
    objdir = os.path.join(ROOT, "build")
    os.makedirs(objdir, exist_ok=True)
    objs, procs = [], []
    for src in CUDA_SOURCES:
        obj = os.path.join(objdir, src + ".o")
        objs.append(obj)
        deps = set()
        _includes(os.path.join(d, src), deps)
        if force or _newer(obj, sorted(deps) + [os.path.abspath(__file__)]):
            cmd = [NVCC] + cuda_flags() + (["-Xptxas", "-v"] if verbose else []) + ["-I", os.path.join(REPO, "include"), "-I", d, "-c", "-o", obj, src]
            procs.append((cmd, subprocess.Popen(cmd, cwd=d, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        log = p.communicate()[0]
        if p.returncode != 0:
            sys.stderr.write(log)
            raise RuntimeError("build failed: " + " ".join(cmd))
        if verbose:
            print(log)
    if procs or force or _newer(out, objs):
        _run([NVCC, "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs + ["-ldl"], d)
    return out


def build_experiment(tag, defines):
    """A/B builds: build_exp/libptb_<tag>.so from the same sources with extra -D macros (loaded through PTB_LIB)."""
    d = os.path.join(ROOT, "csrc")
    objdir = os.path.join(REPO, "build_exp", tag)
    os.makedirs(objdir, exist_ok=True)
    out = os.path.join(REPO, "build_exp", "libptb_%s.so" % tag)
    objs, procs = [], []
    for src in CUDA_SOURCES:
        obj = os.path.join(objdir, src + ".o")
        objs.append(obj)
        cmd = [NVCC] + cuda_flags() + ["-D" + x for x in defines] + ["-I", os.path.join(REPO, "include"), "-I", d, "-c", "-o", obj, src]
        procs.append((cmd, subprocess.Popen(cmd, cwd=d, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        log = p.communicate()[0]
        if p.returncode != 0:
            sys.stderr.write(log)
            raise RuntimeError("build failed: " + " ".join(cmd))
    _run([NVCC, "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs + ["-ldl"], d)
    return out


def build_headless(force=False):
    """bin/ptb_headless: the reference application's frame loop without a window, in C++ over host/integrator.hpp
    (the reference's integrator interface on top of the C ABI).  Links libptb.so and libptb_scene.so next to it."""
    build_scene(force)
    build_cuda(force)
    out = os.path.join(ROOT, "bin", "ptb_headless")
    srcs = [os.path.join(ROOT, "app", "headless.cpp"), os.path.join(ROOT, "host", "integrator.cpp")]
    deps = srcs + [os.path.join(ROOT, "host", f) for f in ("integrator.hpp", "scene.hpp", "procedural.hpp")] + \
        [os.path.join(REPO, "include", "ptb.h"), os.path.join(ROOT, "libptb.so"), os.path.join(ROOT, "libptb_scene.so")]
    if force or _newer(out, deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        _run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-pthread", "-I", os.path.join(REPO, "include"), "-o", out] + srcs +
             ["-L", ROOT, "-lptb", "-lptb_scene", "-Wl,-rpath,$ORIGIN/.."], ROOT)
    return out


def build_all(force=False):
    return build_scene(force), build_cuda(force), build_headless(force)
