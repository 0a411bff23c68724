"""B200-native batched fitness evaluator for FunSearch Kubernetes scheduling policies.

Drop-in for the reference's episode-evaluation path (simulator/main.py run_schedule +
event_simulator.py + evaluator.py as driven by funsearch/funsearch_integration.py:30-64,535-540).
"""
from .policy_compiler import CompiledPolicy, PolicyCompileError, compile_policy  # noqa: F401
from .workload import ClusterSpec, Trace, available_traces, bootstrap_trace, synthetic_trace  # noqa: F401


def __getattr__(name):
    # the evaluator pulls in torch; keep `import funsearch_kubernetes_simulator_b200` light
    if name in ("BatchEvaluator", "EvaluationResults", "EpisodeOutcome", "evaluate_population"):
        from . import evaluator
        return getattr(evaluator, name)
    raise AttributeError(name)
