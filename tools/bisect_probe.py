"""Scratch: evaluate hand-written policy variants through ks_debug_score and compare with CPython."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from policy_fuzz import cpython_scores, random_cases
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
H = "def priority_function(pod, node):\n    score = 0\n"
V = {
 "C1 int(max3)": H + "    t = -0.25 / 1e6\n    w = int(max(t, 1e6, node.memory_mib_left))\n    return score\n",
 "C2 max3": H + "    t = -0.25 / 1e6\n    w = max(t, 1e6, node.memory_mib_left)\n    return score\n",
 "C3 int(max2)": H + "    w = int(max(1e6, node.memory_mib_left))\n    return score\n",
 "C4 1/max(1,c)": H + "    w = 1.0 / max(1, 1000000)\n    return score\n",
 "C5 t": H + "    t = -0.25 / 1e6\n    return score\n",
 "C6 return int(max3)": H + "    t = -0.25 / 1e6\n    return int(max(t, 1e6, node.memory_mib_left))\n",
 "C7 1/max(1,int(f))": H + "    f = node.memory_mib_left * 1.5\n    w = 1.0 / max(1, int(f))\n    return score\n",
 "C8 1/int(max2)": H + "    w = 1.0 / int(max(1e6, node.memory_mib_left))\n    return score\n",
}
pods, ns, nc, objs = random_cases(192, seed=7)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default").head(10)])
names = list(V)
compiled = ev.compile([V[k] for k in names], processes=0)
for slot, k in enumerate(names):
    got = ev.debug_scores(compiled, slot, pods, ns, nc)
    want = np.array(cpython_scores(V[k], objs), dtype=np.int64)
    d = np.nonzero(got != want)[0]
    print(f"{k:22s} mismatches {d.size:3d}", [(int(i), int(got[i]), int(want[i])) for i in d[:4]])
