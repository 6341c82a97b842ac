"""In-tree native builds.

Two shared libraries are produced under ``distributed_tensorflow_b200/_lib``:

* ``libdtf_kernels.so`` -- every ``csrc/kernels/*.cu`` compiled by nvcc for sm_100a only
  (``-gencode arch=compute_100a,code=sm_100a -lineinfo``); the CUDA runtime is linked statically so
  the library loads on a GPU-less box (nvcc cross-compiles there).
* ``libdtf_native.so`` -- the CPU runtime in ``csrc/native/*.cc`` (checkpoint-V2 bundle writer /
  reader, crc32c, SSTable, TCP parameter-server transport, named queues, optimizer apply).

Objects are cached by source mtime so that a rebuild only recompiles what changed.  The ``.so``
files stay in the tree (git-ignored) so that they travel with the repository snapshot.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parents[2]
CSRC = REPO / "csrc"
LIB_DIR = REPO / "distributed_tensorflow_b200" / "_lib"
OBJ_DIR = REPO / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC",
    "-Xcompiler", "-Wno-unused-function",
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread", "-msse4.2"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _newest_header_mtime(d: Path) -> float:
    m = 0.0
    for pat in ("*.cuh", "*.h", "*.hpp"):
        for p in d.glob(pat):
            m = max(m, p.stat().st_mtime)
    for p in (REPO / "include").glob("*.h"):
        m = max(m, p.stat().st_mtime)
    return m


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))


def _compile_all(srcs: list[Path], mk_cmd, tag: str, jobs: int) -> list[Path]:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    objs, todo = [], []
    for s in srcs:
        o = OBJ_DIR / (tag + "_" + s.stem + ".o")
        objs.append(o)
        dep_m = max(s.stat().st_mtime, _newest_header_mtime(s.parent))
        if not o.exists() or o.stat().st_mtime < dep_m:
            todo.append((s, o))
    if todo:
        with cf.ThreadPoolExecutor(max_workers=jobs) as ex:
            futs = [ex.submit(_run, mk_cmd(s, o)) for s, o in todo]
            for f in futs:
                f.result()
    return objs


def source_hash(subdir: str) -> str:
    """sha256 over the names and contents of csrc/<subdir>/* and include/*.h."""
    import hashlib
    h = hashlib.sha256()
    files = sorted(p for p in (CSRC / subdir).iterdir() if p.is_file()) + sorted((REPO / "include").glob("*.h"))
    for p in files:
        h.update(p.name.encode() + b"\0")
        h.update(p.read_bytes())
        h.update(b"\0")
    return h.hexdigest()[:32]


class _BuildLock:
    """Inter-process lock around a build: several tasks of one job start at the same moment (ps, worker,
    master, or N torchrun ranks) and each checks the libraries; without the lock two of them could decide to
    rebuild at once and delete each other's objects, or load a library that another one is relinking."""

    def __init__(self, name: str):
        LIB_DIR.mkdir(parents=True, exist_ok=True)
        self.path = LIB_DIR / (name + ".lock")
        self.f = None

    def __enter__(self):
        import fcntl
        self.f = open(self.path, "w")
        fcntl.flock(self.f, fcntl.LOCK_EX)
        return self

    def __exit__(self, *a):
        import fcntl
        fcntl.flock(self.f, fcntl.LOCK_UN)
        self.f.close()


def _stamp_path(out: Path) -> Path:
    return out.with_suffix(out.suffix + ".srchash")


def built_hash(out: Path) -> str:
    p = _stamp_path(out)
    return p.read_text().strip() if p.exists() and out.exists() else ""


def build_kernels(force: bool = False, jobs: int | None = None, verbose: bool = False) -> Path:
    """Compile csrc/kernels/*.cu -> _lib/libdtf_kernels.so (sm_100a).

    Objects are cached by mtime for quick edits, but the library as a whole is tied to a content hash of
    ``csrc/kernels`` + ``include``: when the hash stamped next to (and compiled into) the existing library
    differs from the tree's, everything is rebuilt from scratch, so a prebuilt ``.so`` that travelled with
    a snapshot can never silently stand in for different sources."""
    with _BuildLock("kernels"):
        return _build_kernels_locked(force, jobs, verbose)


def _build_kernels_locked(force: bool, jobs, verbose: bool) -> Path:
    srcs = sorted((CSRC / "kernels").glob("*.cu"))
    out = LIB_DIR / "libdtf_kernels.so"
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    want = source_hash("kernels")
    if built_hash(out) != want:
        force = True
    if force and OBJ_DIR.exists():
        for o in OBJ_DIR.glob("k_*.o"):
            o.unlink()
    nvcc = _nvcc()
    extra = ["-Xptxas", "-v"] if verbose else []

    def cmd(s, o):
        stamp = ["-DDTF_SOURCE_HASH=\"%s\"" % want] if s.name == "common.cu" else []
        return [nvcc, *NVCC_FLAGS, *extra, *stamp, "-I", str(REPO / "include"), "-c", str(s), "-o", str(o)]
    objs = _compile_all(srcs, cmd, "k", jobs or min(8, os.cpu_count() or 4))
    if force or not out.exists() or any(o.stat().st_mtime > out.stat().st_mtime for o in objs):
        tmp = out.with_suffix(".so.tmp%d" % os.getpid())
        _run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static",
              "-Xcompiler", "-fPIC", "-o", str(tmp), *map(str, objs), "-lpthread", "-ldl", "-lrt"])
        os.replace(tmp, out)                 # readers never see a half-written library
    _stamp_path(out).write_text(want + "\n")
    return out


def build_native(force: bool = False, jobs: int | None = None) -> Path:
    """Compile csrc/native/*.cc -> _lib/libdtf_native.so (CPU runtime)."""
    with _BuildLock("native"):
        return _build_native_locked(force, jobs)


def _build_native_locked(force: bool, jobs) -> Path:
    srcs = [p for p in sorted((CSRC / "native").glob("*.cc")) if p.name != "tsan_stress.cc"]   # that one has a main()
    out = LIB_DIR / "libdtf_native.so"
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    want = source_hash("native")
    if built_hash(out) != want:
        force = True                       # same rule as the kernels: the library follows the content of its sources
    if force and OBJ_DIR.exists():
        for o in OBJ_DIR.glob("n_*.o"):
            o.unlink()
    cxx = os.environ.get("CXX", "g++")
    objs = _compile_all(
        srcs, lambda s, o: [cxx, *CXX_FLAGS, "-I", str(REPO / "include"), "-c", str(s), "-o", str(o)],
        "n", jobs or min(8, os.cpu_count() or 4))
    if force or not out.exists() or any(o.stat().st_mtime > out.stat().st_mtime for o in objs):
        tmp = out.with_suffix(".so.tmp%d" % os.getpid())
        _run([cxx, "-shared", "-pthread", "-o", str(tmp), *map(str, objs)])
        os.replace(tmp, out)
    _stamp_path(out).write_text(want + "\n")
    return out


def build_tsan_stress(out: Path | None = None) -> Path:
    """ThreadSanitizer build of the C++ runtime + its stress driver (csrc/native/tsan_stress.cc).
    Run it with TSAN_OPTIONS=suppressions=csrc/native/tsan.supp; tests/test_tsan_native.py does."""
    out = out or (OBJ_DIR / "dtf_tsan_stress")
    out.parent.mkdir(parents=True, exist_ok=True)
    srcs = sorted((CSRC / "native").glob("*.cc"))
    last = None
    for cxx in dict.fromkeys([os.environ.get("CXX", "g++"), "g++", "/usr/bin/g++"]):   # not every g++ ships libtsan
        try:
            _run([cxx, "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-pthread", "-msse4.2", "-I", str(REPO / "include"),
                  *map(str, srcs), "-o", str(out)])
            return out
        except Exception as e:
            last = e
    raise last


def build_all(force: bool = False, verbose: bool = False) -> dict:
    return {"kernels": str(build_kernels(force=force, verbose=verbose)), "native": str(build_native(force=force))}


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
