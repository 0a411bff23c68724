"""Worker processes for cold policy compilation (translation + NVRTC).

NVRTC serialises much of its work inside one process (about 3x speed-up on 8 threads, measured), and the
Python translator holds the GIL, so a population of fresh policies compiles fastest in separate processes.
Workers translate their slice, compile it with NVRTC in groups (ks_policy_compile_many) and leave the cubins in
the shared on-disk cache (ks_set_cache_dir); the parent then registers the same groups, which are cache hits.
No CUDA context is touched in the workers.

The workers are plain `python -m funsearch_kubernetes_simulator_b200.compile_pool <cache_dir> <batch>`
subprocesses talking length-prefixed pickles over stdin/stdout: unlike multiprocessing's spawn they do not
re-import the caller's __main__, so they work from a REPL, a notebook or `python -`.  A worker that dies is
simply dropped: the parent compiles whatever is missing in-process.
"""
from __future__ import annotations

import ctypes as C
import os
import pickle
import struct
import subprocess
import sys
import tempfile
from typing import List, Optional, Tuple

_WORKERS: List[subprocess.Popen] = []
_CACHE_DIR: Optional[str] = None
_BATCH = 0


def _compile_slice(sources: List[str]) -> List[Optional[Tuple[str, str, int]]]:
    """Translate and NVRTC-compile one slice; returns (cuda_source, symbol, flags) or None per policy."""
    from . import _native as nat
    from .policy_compiler import PolicyCompileError, compile_policy
    out: List[Optional[Tuple[str, str, int]]] = []
    ok = []
    for s in sources:
        try:
            cp = compile_policy(s)
            out.append((cp.cuda_source, cp.symbol, cp.flags))
            ok.append(cp)
        except PolicyCompileError:
            out.append(None)
    seen, uniq = set(), []
    for cp in ok:  # distinct symbols only, in order: the parent rebuilds exactly these groups
        if cp.symbol not in seen:
            seen.add(cp.symbol)
            uniq.append(cp)
    n = len(uniq)
    if n:
        srcs = (C.c_char_p * n)(*[cp.cuda_source.encode() for cp in uniq])
        syms = (C.c_char_p * n)(*[cp.symbol.encode() for cp in uniq])
        flags = (C.c_longlong * n)(*[cp.flags for cp in uniq])
        handles = (C.c_longlong * n)()
        status = (C.c_int * n)()
        nat.check(nat.lib().ks_policy_compile_many(n, srcs, syms, flags, 1, handles, status))
    return out


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
    from . import _native as nat
    lib = nat.lib()
    lib.ks_set_cache_dir(cache_dir.encode())
    lib.ks_set_compile_batch(batch)
    inp, out = sys.stdin.buffer, sys.stdout.buffer
    sys.stdout = sys.stderr  # nothing but the protocol may reach the pipe
    while True:
        try:
            job = _recv(inp)
        except EOFError:
            return
        try:
            _send(out, ("ok", _compile_slice(job)))
        except Exception as exc:  # noqa: BLE001
            _send(out, ("error", repr(exc)))


def _start(processes: int, batch: int) -> str:
    global _CACHE_DIR, _BATCH
    if _CACHE_DIR is None:
        _CACHE_DIR = tempfile.mkdtemp(prefix="ks_cubin_cache_")
    if _BATCH != batch:
        shutdown()
        _BATCH = batch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ)
    env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
    alive = [w for w in _WORKERS if w.poll() is None]
    _WORKERS[:] = alive
    while len(_WORKERS) < processes:
        _WORKERS.append(subprocess.Popen([sys.executable, "-m", "funsearch_kubernetes_simulator_b200.compile_pool",
                                          _CACHE_DIR, str(batch)], stdin=subprocess.PIPE, stdout=subprocess.PIPE, env=env))
    return _CACHE_DIR


def compile_slices(slices: List[List[str]], processes: int, batch: int):
    """Compile `slices` on up to `processes` workers.  Returns (cache_dir, results) where results[i] is the list
    _compile_slice returns for slices[i], or None when its worker failed."""
    cache_dir = _start(min(processes, len(slices)), batch)
    results: List[Optional[list]] = [None] * len(slices)
    pending = list(range(len(slices)))
    while pending:
        wave = []
        for w in list(_WORKERS):
            if not pending:
                break
            i = pending.pop(0)
            try:
                _send(w.stdin, slices[i])
                wave.append((w, i))
            except (BrokenPipeError, OSError):
                _WORKERS.remove(w)
        if not wave:
            break  # no live worker: the caller compiles in-process
        for w, i in wave:
            try:
                tag, payload = _recv(w.stdout)
                if tag == "ok":
                    results[i] = payload
            except (EOFError, OSError, pickle.UnpicklingError):
                if w in _WORKERS:
                    _WORKERS.remove(w)
    return cache_dir, results


def shutdown() -> None:
    for w in _WORKERS:
        try:
            w.stdin.close()
            w.wait(timeout=5)
        except Exception:  # noqa: BLE001
            w.kill()
    _WORKERS.clear()


if __name__ == "__main__":
    _worker_main(sys.argv[1], int(sys.argv[2]))
