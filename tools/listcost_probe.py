"""Scratch: what in a list-heavy policy costs the cycles?  Variants of the slowest population policy, 148 episodes each (1 warp/SM)
and 1,036 (7 warps/SM, same policy: no instruction-cache contention).  Needs KS_PHASE_CLOCKS=1."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, build
build.build()
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
PRE = '''def priority_function(pod, node):
    if (pod.cpu_milli > node.cpu_milli_left or
        pod.memory_mib > node.memory_mib_left or
        pod.num_gpu > node.gpu_left):
        return 0

    if pod.num_gpu > 0:
        available_gpus = 0
        for gpu in node.gpus:
            if gpu.gpu_milli_left >= pod.gpu_milli:
                available_gpus += 1
        if available_gpus < pod.num_gpu:
            return 0

    score = 0.0
    cpu_util = (node.cpu_milli_total - node.cpu_milli_left + pod.cpu_milli) / max(1, node.cpu_milli_total)
    mem_util = (node.memory_mib_total - node.memory_mib_left + pod.memory_mib) / max(1, node.memory_mib_total)
    score += (1 - abs(cpu_util - mem_util)) ** 1.401347 * 277.765166
'''
POST = "\n    return max(1, int(score))\n"
V = {
 "full": PRE + '''    if pod.num_gpu > 0:
        fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]
        tight = sorted(fits, key=lambda g: g.gpu_milli_left)[:pod.num_gpu]
        score += sum(1 - (g.gpu_milli_left - pod.gpu_milli) / g.gpu_milli_total for g in tight) / pod.num_gpu * 561.759576
''' + POST,
 "no gpu part": PRE + POST,
 "no sorted": PRE + '''    if pod.num_gpu > 0:
        fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]
        tight = fits[:pod.num_gpu]
        score += sum(1 - (g.gpu_milli_left - pod.gpu_milli) / g.gpu_milli_total for g in tight) / pod.num_gpu * 561.759576
''' + POST,
 "no sum (plain loop)": PRE + '''    if pod.num_gpu > 0:
        fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]
        tight = sorted(fits, key=lambda g: g.gpu_milli_left)[:pod.num_gpu]
        acc = 0.0
        for g in tight:
            acc += 1 - (g.gpu_milli_left - pod.gpu_milli) / g.gpu_milli_total
        score += acc / pod.num_gpu * 561.759576
''' + POST,
 "listcomp+slice+len": PRE + '''    if pod.num_gpu > 0:
        fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]
        tight = fits[:pod.num_gpu]
        score += len(tight) / pod.num_gpu * 561.759576
''' + POST,
 "listcomp+sum(fits)": PRE + '''    if pod.num_gpu > 0:
        fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]
        score += sum(1 - (g.gpu_milli_left - pod.gpu_milli) / g.gpu_milli_total for g in fits) / pod.num_gpu * 561.759576
''' + POST,
 "listcomp+sum(left)": PRE + '''    if pod.num_gpu > 0:
        fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]
        score += sum(g.gpu_milli_left for g in fits) / pod.num_gpu * 0.561759576
''' + POST,
 "only listcomp+len": PRE + '''    if pod.num_gpu > 0:
        fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]
        score += len(fits) / pod.num_gpu * 561.759576
''' + POST,
}
names = list(V)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile([V[n] for n in names])
for k, n in enumerate(names):
    for n_ep in (148,):
        rb = ev.prepare_resident(compiled, [k] * n_ep, [0] * n_ep)
        rb.launch(); torch.cuda.synchronize()
        r = rb.results()
        ev_ = r["events"].astype(float); at = r["attempts"].astype(float); placed = r["placed"].astype(float)
        sc = r["reserved"][:, 1].astype(float); cyc = r["device_cycles"].astype(float)
        fails = at - placed
        print(f"{n:22s} n={n_ep:5d}: events {ev_.mean():6.0f} attempts {at.mean():6.0f} fails {fails.mean():6.0f} cyc/event {np.mean(cyc/ev_):5.0f} score/attempt {np.mean(sc/at):5.0f} "
              f"score/placement(if fails cost 150) {np.mean((sc - 150*fails)/placed):6.0f}", flush=True)
for n in names:
    c = compile_policy(V[n]); print(n, "CUDA lines", c.cuda_source.count("\n"))
