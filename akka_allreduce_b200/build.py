"""In-tree builder for the native libraries.

Two artefacts land in ``akka_allreduce_b200/_native/`` (git-ignored, shipped to the GPU box
by the snapshot):

* ``_aar_host*.so``   - C++17 host library (protocol engines, buffers, bus, TCP nodes),
                        pybind11, built with g++.
* ``libaar_cuda.so``  - sm_100a CUDA kernels + symmetric heap behind a C ABI
                        (``include/aar_b200.h``), built with nvcc, loaded through ctypes.

``python -m akka_allreduce_b200.build`` builds both; ``--host`` / ``--cuda`` pick one; ``--tools`` builds the
stand-alone binaries ``build/aar_cpu_cluster`` and ``build/aar_shm_cluster``.
Rebuilds only when a source or header is newer than the artefact.
"""
from __future__ import annotations

import fcntl
import hashlib
import os
import shutil
import subprocess
import sys
import sysconfig
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
NATIVE = PKG / "_native"
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"

HOST_SOURCES = [
    "host/buffers.cc",
    "host/worker_engine.cc",
    "host/master_engine.cc",
    "host/local_bus.cc",
    "host/node.cc",
    "host/shm_cluster.cc",
    "host/shm_bindings.cc",
    "host/node_bindings.cc",
    "host/bindings.cc",
]
CUDA_SOURCES = [
    "cuda/fused_allreduce.cu",
    "cuda/sym_heap.cu",
    "cuda/probe.cu",
]
CUDA_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _digest(deps: list[Path], extra: str = "") -> str:
    h = hashlib.sha256(extra.encode())
    for d in sorted(set(deps)):
        if d.exists():
            h.update(str(d.relative_to(ROOT)).encode())
            h.update(d.read_bytes())
    return h.hexdigest()


def _stale(target: Path, deps: list[Path], extra: str = "") -> bool:
    """A library is stale when the digest of its sources differs from the one recorded at build time.
    (Content, not mtimes: a snapshot copied to another machine keeps its libraries.)"""
    stamp = target.with_suffix(target.suffix + ".srchash")
    if target.exists() and not CSRC.exists():
        return False  # installed package: the libraries were built at install time, there is nothing to rebuild from
    if not target.exists() or not stamp.exists():
        return True
    return stamp.read_text().strip() != _digest(deps, extra)


def _record(target: Path, deps: list[Path], extra: str = "") -> None:
    target.with_suffix(target.suffix + ".srchash").write_text(_digest(deps, extra))


class _BuildLock:
    """One builder at a time per checkout: ranks started together must not write the same .so."""

    def __enter__(self):
        NATIVE.mkdir(exist_ok=True)
        self._f = open(NATIVE / ".build.lock", "w")
        fcntl.flock(self._f, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        fcntl.flock(self._f, fcntl.LOCK_UN)
        self._f.close()
        return False


def _headers() -> list[Path]:
    return list(INCLUDE.rglob("*.h")) + list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.cuh"))


def _run(cmd: list[str]) -> None:
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + proc.stdout + proc.stderr)
        raise RuntimeError(f"native build failed: {cmd[0]} exited {proc.returncode}")
    if proc.stderr.strip() and os.environ.get("AAR_BUILD_VERBOSE"):
        sys.stderr.write(proc.stderr)


def host_sanitizer() -> str:
    """``AAR_HOST_SANITIZE=address|undefined|thread`` builds (and loads) an instrumented copy of the host library
    in its own directory, next to the normal one; see scripts/sanitize_host.sh."""
    san = os.environ.get("AAR_HOST_SANITIZE", "").strip().lower()
    if san and san not in ("address", "undefined", "thread"):
        raise RuntimeError(f"AAR_HOST_SANITIZE={san!r}: expected address, undefined or thread")
    return san


def host_native_dir() -> Path:
    san = host_sanitizer()
    return PKG / f"_native_{san}" if san else NATIVE


def host_library_path() -> Path:
    return host_native_dir() / ("_aar_host" + sysconfig.get_config_var("EXT_SUFFIX"))


def cuda_library_path() -> Path:
    return NATIVE / "libaar_cuda.so"


def build_host(force: bool = False) -> Path:
    import pybind11

    out = host_library_path()
    srcs = [CSRC / s for s in HOST_SOURCES]
    deps = srcs + _headers()
    if not force and not _stale(out, deps):
        return out
    with _BuildLock():
        if not force and not _stale(out, deps):  # somebody else built it while we waited
            return out
        san = host_sanitizer()
        objdir = ROOT / "build" / ("host_" + san if san else "host")
        objdir.mkdir(parents=True, exist_ok=True)
        out.parent.mkdir(exist_ok=True)
        (out.parent / "__init__.py").touch()
        opt = ["-O1", "-g", "-fno-omit-frame-pointer", f"-fsanitize={san}"] if san else ["-O2"]
        common = [
            "g++", "-std=c++17", *opt, "-fPIC", "-fvisibility=hidden", "-Wall", "-Wno-unused-parameter",
            f"-I{INCLUDE}", f"-I{CSRC / 'host'}", f"-I{pybind11.get_include()}",
            f"-I{sysconfig.get_paths()['include']}",
        ]
        objs, procs = [], []
        for s in srcs:
            o = objdir / (s.stem + ".o")
            objs.append(o)
            procs.append((s, subprocess.Popen(common + ["-c", str(s), "-o", str(o)],
                                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        for s, p in procs:
            log, _ = p.communicate()
            if p.returncode != 0:
                sys.stderr.write(log)
                raise RuntimeError(f"native build failed compiling {s}")
        tmp = out.with_suffix(".tmp%d" % os.getpid())
        _run(["g++", "-shared", "-o", str(tmp)] + ([f"-fsanitize={san}"] if san else []) + [str(o) for o in objs] +
             ["-lpthread"])
        os.replace(tmp, out)
        _record(out, deps)
    return out


def build_cpu_cluster(force: bool = False) -> Path:
    """Stand-alone binary of the CPU backend (coordinator + N workers in one process, no Python):
    ``build/aar_cpu_cluster [numOfWorkers dataSize maxChunkSize maxRound maxLag thAllreduce thReduce thComplete]``."""
    out = ROOT / "build" / "aar_cpu_cluster"
    core = ["host/buffers.cc", "host/worker_engine.cc", "host/master_engine.cc", "host/local_bus.cc", "host/node.cc",
            "host/cpu_cluster_main.cc"]
    srcs = [CSRC / s for s in core]
    deps = srcs + _headers()
    if not force and not _stale(out, deps):
        return out
    with _BuildLock():
        (ROOT / "build").mkdir(exist_ok=True)
        tmp = out.with_suffix(".tmp%d" % os.getpid())
        _run(["g++", "-std=c++17", "-O2", "-Wall", "-Wno-unused-parameter", f"-I{INCLUDE}", f"-I{CSRC / 'host'}",
              "-o", str(tmp)] + [str(s) for s in srcs] + ["-lpthread"])
        os.replace(tmp, out)
        _record(out, deps)
    return out


def build_shm_cluster(force: bool = False) -> Path:
    """Stand-alone binary of the shared-memory restatement of the device protocol (csrc/host/shm_cluster.h):
    ``build/aar_shm_cluster [world n maxChunkSize maxLag rounds thReduce thComplete thAllreduce stallProb stallUs timeoutMs]``."""
    out = ROOT / "build" / "aar_shm_cluster"
    srcs = [CSRC / "host/shm_cluster.cc", CSRC / "host/shm_cluster_main.cc"]
    deps = srcs + _headers()
    if not force and not _stale(out, deps):
        return out
    with _BuildLock():
        (ROOT / "build").mkdir(exist_ok=True)
        tmp = out.with_suffix(".tmp%d" % os.getpid())
        _run(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", f"-I{INCLUDE}", f"-I{CSRC / 'host'}", "-o", str(tmp)] +
             [str(s) for s in srcs] + ["-lpthread"])
        os.replace(tmp, out)
        _record(out, deps)
    return out


def nvcc_path() -> str | None:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return cand if Path(cand).exists() else None


def build_cuda(force: bool = False) -> Path:
    out = cuda_library_path()
    srcs = [CSRC / s for s in CUDA_SOURCES if (CSRC / s).exists()]
    deps = srcs + _headers()
    if not force and not _stale(out, deps):
        return out
    with _BuildLock():
        if not force and not _stale(out, deps):
            return out
        nvcc = nvcc_path()
        if nvcc is None:
            raise RuntimeError("nvcc not found; cannot build the sm_100a kernels")
        tmp = out.with_suffix(".tmp%d" % os.getpid())
        cmd = [nvcc, "-std=c++17", "-O3", "-lineinfo", *CUDA_ARCH, "-Xcompiler", "-fPIC", "-shared",
               "-Xptxas", "-v", f"-I{INCLUDE}", f"-I{CSRC / 'cuda'}", "-o", str(tmp)] + [str(s) for s in srcs]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        (ROOT / "build").mkdir(exist_ok=True)
        (ROOT / "build" / "ptxas_v.log").write_text(proc.stdout + proc.stderr)
        if proc.returncode != 0:
            sys.stderr.write(" ".join(cmd) + "\n" + proc.stdout + proc.stderr)
            raise RuntimeError("nvcc failed")
        os.replace(tmp, out)
        _record(out, deps)
    return out


def build_all(force: bool = False) -> None:
    build_host(force)
    build_cuda(force)


if __name__ == "__main__":
    args = set(sys.argv[1:])
    force = "--force" in args
    if "--host" in args or not (args - {"--force"}):
        print(build_host(force))
    if "--cuda" in args or not (args - {"--force"}):
        print(build_cuda(force))
    if "--tools" in args:  # stand-alone native binaries: TCP cluster and shared-memory cluster
        print(build_cpu_cluster(force))
        print(build_shm_cluster(force))
