"""Worker processes for cold policy compilation (translation + NVRTC).

NVRTC serialises much of its work inside one process (about 3x speed-up on 8 threads, measured), and the
Python translator holds the GIL, so a generation of fresh policies compiles fastest in separate processes.
A generation goes through the pool in two phases:

  1. translate: the fresh SOURCES are cut into one slice per worker; each worker returns, per source, the
     structure it translates to (CUDA text, symbol, flags) and the policy's float literals;
  2. nvrtc + link: the parent keeps one copy of every structure it has not compiled yet -- a generation of 1,024
     mutated sources is a few hundred structures, often far fewer -- forms groups of `batch` structures (one NVRTC
     program each) and deals the groups to the workers, which leave the cubins in the shared cache directory
     (ks_set_cache_dir) and link what they compiled with the replay kernel into one loadable module per worker
     (nvJitLink serialises inside a process, so linking in the workers is what makes it parallel); the parent then
     registers the same groups, which are cache hits, and loads the modules.

No CUDA context is touched in the workers.  They are started when the evaluator is constructed (start()), not
inside the first evaluation, and warm themselves up (imports, mapping libnvrtc and its builtins, one trivial
program): on a fresh box that first touch is the slowest part of the whole pipeline.

The workers are plain `python -m funsearch_kubernetes_simulator_b200.compile_pool <cache_dir> <batch>`
subprocesses talking length-prefixed pickles over stdin/stdout: unlike multiprocessing's spawn they do not
re-import the caller's __main__, so they work from a REPL, a notebook or `python -`.  A worker that dies is
simply dropped: the parent compiles whatever is missing in-process.
"""
from __future__ import annotations

import atexit
import ctypes as C
import os
import pickle
import shutil
import struct
import subprocess
import sys
import tempfile
import time
from typing import List, Optional, Sequence, Tuple

_WORKERS: List["_Worker"] = []
_CACHE_DIR: Optional[str] = None
_BATCH = 0

# (cuda_source, symbol, flags) of one structure
Structure = Tuple[str, str, int]

_WARMUP_SRC = ('extern "C" __device__ long long ks_warmup_fn(KS_POLICY_PARAMS) { return ks_wrap_f(ks_div((double)p_cpu, 3.0)); }\n',
               "ks_warmup_fn", 0)


class _Worker:
    def __init__(self, proc: subprocess.Popen):
        self.proc = proc
        self.ready = False
        self.warmup_s: Optional[float] = None

    def alive(self) -> bool:
        return self.proc.poll() is None


def _translate_slice(sources: Sequence[str]):
    """Per source: (cuda_source, symbol, flags, consts) or None when it is outside the supported subset."""
    from .policy_compiler import PolicyCompileError, compile_policy
    out = []
    for s in sources:
        try:
            cp = compile_policy(s)
            out.append((cp.cuda_source, cp.symbol, cp.flags, cp.consts))
        except PolicyCompileError as e:
            out.append(("error", str(e)))
    return out


def compile_group(lib, group: Sequence[Structure]) -> Tuple[List[int], List[int]]:
    """One NVRTC program for the structures of `group` (ks_policy_compile_many with the batch size >= len(group));
    returns (per-structure status, per-structure handle).  Shared by the workers and by the parent's registration pass, so that both
    form byte-identical programs and the second one is a disk-cache hit."""
    n = len(group)
    srcs = (C.c_char_p * n)(*[g[0].encode() for g in group])
    syms = (C.c_char_p * n)(*[g[1].encode() for g in group])
    flags = (C.c_longlong * n)(*[g[2] for g in group])
    handles = (C.c_longlong * n)()
    status = (C.c_int * n)()
    rc = lib.ks_policy_compile_many(n, srcs, syms, flags, 1, handles, status)
    if rc != 0:
        return [rc] * n, [-1] * n
    return [int(status[i]) for i in range(n)], [int(handles[i]) for i in range(n)]


def _nvrtc_link(lib, groups, image_path: str) -> dict:
    """Compile `groups` (one NVRTC program each, cubins left in the shared cache) and link all of their structures, in
    order, with the replay kernel into a loadable module written to `image_path` (ks_program_link_image; no GPU is
    touched).  The parent registers the same groups and loads the image (ks_program_load_image)."""
    status, handles = [], []
    for g in groups:
        st, hs = compile_group(lib, g)
        status.append(st)
        handles.extend(hs)
    linked = False
    if all(x == 0 for st in status for x in st):
        arr = (C.c_longlong * len(handles))(*handles)
        linked = lib.ks_program_link_image(len(handles), arr, image_path.encode()) == 0
    return {"status": status, "linked": linked}


def _send(pipe, obj) -> None:
    data = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
    pipe.write(struct.pack("<Q", len(data)))
    pipe.write(data)
    pipe.flush()


def _recv(pipe):
    hdr = pipe.read(8)
    if len(hdr) < 8:
        raise EOFError
    (n,) = struct.unpack("<Q", hdr)
    data = pipe.read(n)
    if len(data) < n:
        raise EOFError
    return pickle.loads(data)


def _worker_main(cache_dir: str, batch: int) -> None:
    inp, out = sys.stdin.buffer, sys.stdout.buffer
    sys.stdout = sys.stderr  # nothing but the protocol may reach the pipe
    t0 = time.perf_counter()
    from . import _native as nat
    from . import policy_compiler  # noqa: F401  (imported now, not inside the first job)
    lib = nat.lib()
    lib.ks_set_compile_batch(max(batch, 1))
    # warm-up: maps libnvrtc and its builtins and runs the compiler once.  Its cubin is not cached on disk.
    compile_group(lib, [_WARMUP_SRC])
    lib.ks_set_cache_dir(cache_dir.encode())
    _send(out, ("ready", time.perf_counter() - t0))
    while True:
        try:
            kind, payload = _recv(inp)
        except EOFError:
            return
        try:
            t0 = time.perf_counter()
            if kind == "translate":
                res = _translate_slice(payload)
            elif kind == "nvrtc":
                res = [compile_group(lib, g)[0] for g in payload]
            elif kind == "nvrtc_link":
                res = _nvrtc_link(lib, *payload)
            else:
                raise ValueError(kind)
            _send(out, ("ok", res, time.perf_counter() - t0))
        except Exception as exc:  # noqa: BLE001
            _send(out, ("error", repr(exc), 0.0))


def cache_dir() -> str:
    """The cubin cache shared by the parent and its workers: RAM-backed when the box has /dev/shm."""
    global _CACHE_DIR
    if _CACHE_DIR is None:
        base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
        _CACHE_DIR = tempfile.mkdtemp(prefix="ks_cubin_cache_", dir=base)
        atexit.register(_cleanup)
    return _CACHE_DIR


def _cleanup() -> None:
    shutdown()
    if _CACHE_DIR and os.path.isdir(_CACHE_DIR):
        shutil.rmtree(_CACHE_DIR, ignore_errors=True)


def start(processes: int, batch: int) -> str:
    """Spawn workers up to `processes` (non-blocking: they warm up on their own) and return the cache directory."""
    global _BATCH
    d = cache_dir()
    if _BATCH != batch:
        shutdown()
        _BATCH = batch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ)
    env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
    # On a machine with the NVIDIA driver NVRTC keeps its results in the driver's compute cache (~/.nv/ComputeCache).  A
    # generation's programs never repeat (structures that do are found in ks_set_cache_dir's cache before NVRTC is asked),
    # so that cache would only be written (about 20 MB per generation of 1,024 sources), never hit.
    env.setdefault("CUDA_CACHE_DISABLE", "1")
    _WORKERS[:] = [w for w in _WORKERS if w.alive()]
    while len(_WORKERS) < processes:
        _WORKERS.append(_Worker(subprocess.Popen(
            [sys.executable, "-m", "funsearch_kubernetes_simulator_b200.compile_pool", d, str(batch)],
            stdin=subprocess.PIPE, stdout=subprocess.PIPE, env=env)))
    return d


def wait_ready(timeout_s: float = 300.0) -> dict:
    """Block until every live worker has finished its warm-up; returns {"workers", "warmup_s_max"}."""
    t_end = time.monotonic() + timeout_s
    for w in list(_WORKERS):
        if w.ready:
            continue
        try:
            if time.monotonic() > t_end:
                raise EOFError
            tag, secs = _recv(w.proc.stdout)
            if tag != "ready":
                raise EOFError
            w.ready, w.warmup_s = True, float(secs)
        except (EOFError, OSError, pickle.UnpicklingError, ValueError):
            _drop(w)
    ws = [w.warmup_s for w in _WORKERS if w.warmup_s is not None]
    return {"workers": len(_WORKERS), "warmup_s_max": max(ws) if ws else 0.0}


def _drop(w: "_Worker") -> None:
    if w in _WORKERS:
        _WORKERS.remove(w)
    try:
        w.proc.kill()
    except Exception:  # noqa: BLE001
        pass


def run_jobs(kind: str, jobs: List[list]):
    """Run `jobs` (one payload each) on the pool, as many at a time as there are workers.  Returns
    (results, worker_seconds): results[i] is the worker's answer for jobs[i] or None when its worker failed."""
    wait_ready()
    results: List[Optional[list]] = [None] * len(jobs)
    secs: List[float] = [0.0] * len(jobs)
    pending = list(range(len(jobs)))
    while pending:
        wave = []
        for w in list(_WORKERS):
            if not pending:
                break
            i = pending.pop(0)
            try:
                _send(w.proc.stdin, (kind, jobs[i]))
                wave.append((w, i))
            except (BrokenPipeError, OSError):
                _drop(w)
        if not wave:
            break  # no live worker: the caller does the work in-process
        for w, i in wave:
            try:
                tag, payload, dt = _recv(w.proc.stdout)
                if tag == "ok":
                    results[i] = payload
                    secs[i] = dt
            except (EOFError, OSError, pickle.UnpicklingError, ValueError):
                _drop(w)
    return results, secs


def n_workers() -> int:
    return len([w for w in _WORKERS if w.alive()])


def shutdown() -> None:
    for w in _WORKERS:
        try:
            w.proc.stdin.close()
            w.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            w.proc.kill()
    _WORKERS.clear()


if __name__ == "__main__":
    _worker_main(sys.argv[1], int(sys.argv[2]))
