"""Differential test of the policy translator: random policy sources (tests/policy_fuzz.py) are translated,
compiled and evaluated on the GPU for random (pod, node) cases through ks_debug_score, and every wrapped score
int(max(0, f(pod, node))) is compared with what CPython itself returns for the same source and case -- including
which cases raise.  Integer results must be identical; the only tolerated disagreement is the device declining
(KS_RET_UNSUPPORTED) where CPython used an int beyond int64 or a complex number."""
import os

import numpy as np
import pytest

from policy_fuzz import Gen, cpython_scores, random_cases

pytestmark = pytest.mark.gpu

N_POLICIES = int(os.environ.get("KS_FUZZ_POLICIES", "240"))   # raise for a longer hunt
FUZZ_SEED0 = int(os.environ.get("KS_FUZZ_SEED0", "1000"))
N_CASES = 192


def test_random_policies_match_cpython(built_lib):
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    from funsearch_kubernetes_simulator_b200.policy_compiler import PolicyCompileError, compile_policy
    sources = []
    for seed in range(N_POLICIES):
        src = Gen(FUZZ_SEED0 + seed).policy()
        try:
            compile_policy(src)
            sources.append(src)
        except PolicyCompileError:
            pass
    assert len(sources) >= N_POLICIES * 0.9, "the generator should stay inside the supported subset"
    pods, ns, nc, objs = random_cases(N_CASES, seed=7)
    ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default").head(10)])
    compiled = ev.compile(sources, processes=0)
    declined = total = 0
    for slot, src in enumerate(sources):
        got = ev.debug_scores(compiled, slot, pods, ns, nc)
        want = np.array(cpython_scores(src, objs), dtype=np.int64)
        total += len(want)
        soft = got == -2          # the device declined: allowed only as documented, and rarely
        declined += int(soft.sum())
        bad = np.nonzero((got != want) & ~soft)[0]
        if bad.size:
            i = int(bad[0])
            pod, node = objs[i]
            raise AssertionError(f"policy #{slot} case {i}: device {int(got[i])} vs CPython {int(want[i])}\n{src}\n"
                                 f"pod={vars(pod)}\nnode cpu_left={node.cpu_milli_left} mem_left={node.memory_mib_left} "
                                 f"gpu_left={node.gpu_left} totals=({node.cpu_milli_total},{node.memory_mib_total}) "
                                 f"gpus={[g.gpu_milli_left for g in node.gpus]}")
    assert declined <= 0.03 * total, f"device declined {declined} of {total} cases"


def test_shipped_policies_match_cpython_case_by_case(built_lib):
    """The same comparison for every shipped policy source (tests/golden/policies.json)."""
    from helpers import load_golden
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    _g, pol = load_golden()
    names = sorted(pol)
    pods, ns, nc, objs = random_cases(512, seed=11)
    ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default").head(10)])
    compiled = ev.compile([pol[n] for n in names], processes=0)
    for slot, n in enumerate(names):
        got = ev.debug_scores(compiled, slot, pods, ns, nc)
        want = np.array(cpython_scores(pol[n], objs), dtype=np.int64)
        bad = np.nonzero(got != want)[0]
        assert bad.size == 0, f"{n}: case {int(bad[0])}: device {int(got[bad[0]])} vs CPython {int(want[bad[0]])}"
