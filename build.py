#!/usr/bin/env python
"""Build libb200ps.so (sm_100a) in-tree with nvcc.  No GPU is needed to build.

    python build.py            # incremental
    python build.py --force    # rebuild everything
    python build.py --verbose  # also print ptxas resource usage (-Xptxas -v)
"""
import argparse
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(ROOT, "b200ps", "_lib")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(OUT_DIR, "libb200ps.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
          "--expt-relaxed-constexpr", "-I", os.path.join(ROOT, "include")]


def sources():
    out = []
    for d, _, files in os.walk(CSRC):
        if os.path.basename(d) == "python":          # CPython extension sources are built separately
            continue
        for f in sorted(files):
            if f.endswith((".cu", ".cpp")):
                out.append(os.path.join(d, f))
    return sorted(out)


def build_python_ext(quiet=False):
    """Optional CPython extension with the per-step fast paths (csrc/python/fastfeed.c) -> b200ps/_lib/_fastfeed.so.
    Skipped (ctypes fallback) when Python.h or a C compiler is missing."""
    import sysconfig
    src = os.path.join(CSRC, "python", "fastfeed.c")
    out = os.path.join(OUT_DIR, "_fastfeed.so")
    inc = sysconfig.get_paths()["include"]
    if not os.path.exists(os.path.join(inc, "Python.h")):
        return None
    if os.path.exists(out) and os.path.getmtime(out) >= max(os.path.getmtime(src), os.path.getmtime(LIB)):
        return out
    cmd = [os.environ.get("CC", "gcc"), "-O2", "-shared", "-fPIC", "-Wall", "-I", inc, src, "-o", out,
           "-L", OUT_DIR, "-lb200ps", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        print("[build] _fastfeed extension skipped:\n" + r.stdout + r.stderr, file=sys.stderr)
        return None
    if not quiet:
        print(f"[build] built {os.path.relpath(out, ROOT)}")
    return out


def headers_digest():
    h = hashlib.sha1()
    for d, _, files in os.walk(CSRC):
        for f in sorted(files):
            if f.endswith((".h", ".cuh")):
                h.update(open(os.path.join(d, f), "rb").read())
    h.update(open(os.path.join(ROOT, "include", "b200ps.h"), "rb").read())
    return h.hexdigest()


def compile_one(src, obj, verbose, extra=()):
    cmd = [NVCC, *ARCH, *COMMON, *extra, "-c", src, "-o", obj]
    if src.endswith(".cpp"):
        cmd = [NVCC, *ARCH, *COMMON, *extra, "-x", "cu", "-c", src, "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    return src, r.returncode, r.stdout + r.stderr


def build(force=False, verbose=False, quiet=False, checked=False):
    """``checked=True`` builds the same sources with -DB200PS_CHECKED (device-side protocol assertions) into
    ``libb200ps_checked.so`` (loaded instead of the release library when B200PS_CHECKED=1)."""
    obj_dir = OBJ_DIR + ("_checked" if checked else "")
    LIB = os.path.join(OUT_DIR, "libb200ps_checked.so" if checked else "libb200ps.so")
    extra = ("-DB200PS_CHECKED",) if checked else ()
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(obj_dir, exist_ok=True)
    hdig = headers_digest()
    jobs, objs = [], []
    for src in sources():
        rel = os.path.relpath(src, CSRC).replace(os.sep, "_")
        obj = os.path.join(obj_dir, rel + ".o")
        stamp = obj + ".stamp"
        key = hashlib.sha1(open(src, "rb").read() + hdig.encode()).hexdigest()
        objs.append(obj)
        fresh = (not force and os.path.exists(obj) and os.path.exists(stamp)
                 and open(stamp).read() == key)
        if not fresh:
            jobs.append((src, obj, stamp, key))
    failed = False
    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        futs = {ex.submit(compile_one, s, o, verbose, extra): (s, o, st, k) for s, o, st, k in jobs}
        for fut in cf.as_completed(futs):
            src, rc, log = fut.result()
            s, o, st, k = futs[fut]
            if rc != 0:
                failed = True
                print(f"[build] FAILED {os.path.relpath(src, ROOT)}\n{log}", file=sys.stderr)
            else:
                open(st, "w").write(k)
                if not quiet:
                    print(f"[build] compiled {os.path.relpath(src, ROOT)}")
                if verbose or "warning" in log:
                    print(log)
    if failed:
        raise RuntimeError("nvcc failed")
    if jobs or not os.path.exists(LIB):
        cmd = [NVCC, *ARCH, "-shared", "-o", LIB, *objs, "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stdout + r.stderr, file=sys.stderr)
            raise RuntimeError("link failed")
        if not quiet:
            print(f"[build] linked {os.path.relpath(LIB, ROOT)}")
    if not checked:
        build_python_ext(quiet)
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--checked", action="store_true", help="build libb200ps_checked.so (-DB200PS_CHECKED)")
    a = ap.parse_args()
    build(force=a.force, verbose=a.verbose, checked=a.checked)
