"""Scratch: run the translator fuzz policy by policy and name the first one that faults or disagrees."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from policy_fuzz import Gen, cpython_scores, random_cases
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.policy_compiler import PolicyCompileError, compile_policy
s0, n = int(sys.argv[1]), int(sys.argv[2])
srcs = []
for seed in range(s0, s0 + n):
    src = Gen(seed).policy()
    try:
        compile_policy(src); srcs.append((seed, src))
    except PolicyCompileError:
        pass
pods, ns, nc, objs = random_cases(192, seed=7)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default").head(10)])
compiled = ev.compile([s for _, s in srcs], processes=0)
bad = 0
for slot, (seed, src) in enumerate(srcs):
    try:
        got = ev.debug_scores(compiled, slot, pods, ns, nc)
        torch.cuda.synchronize()
    except Exception as e:
        print("FAULT at seed", seed, type(e).__name__); print(src); print(compiled[slot].cuda_source); sys.exit(3)
    want = np.array(cpython_scores(src, objs), dtype=np.int64)
    d = np.nonzero((got != want) & (got != -2))[0]
    if d.size:
        bad += 1
        i = int(d[0]); pod, node = objs[i]
        print(f"MISMATCH seed {seed} case {i}: device {int(got[i])} cpython {int(want[i])} ({d.size} cases)"); print(src)
        print("pod", vars(pod)); print("node", node.cpu_milli_left, node.memory_mib_left, node.gpu_left, node.cpu_milli_total, node.memory_mib_total, [g.gpu_milli_left for g in node.gpus])
        for j in d[:12]:
            pj, nj = objs[int(j)]
            print("   case", int(j), "dev", int(got[j]), "py", int(want[j]), "pod", pj.cpu_milli, pj.memory_mib, pj.num_gpu, pj.gpu_milli, "node", nj.cpu_milli_left, nj.memory_mib_left, nj.gpu_left, [g.gpu_milli_left for g in nj.gpus])
        if bad >= 3: break
print("done", len(srcs), "policies,", bad, "mismatching")
