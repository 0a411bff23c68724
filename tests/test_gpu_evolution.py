"""The generation loop (evolution.BatchedFunSearch, the population bookkeeping of funsearch_integration.py:487-572) driven by
the GPU evaluator against the same loop driven by the CPU oracle: same generator seed -> the same populations, generation
after generation (sources identical, scores within 1e-12)."""
import pytest

pytestmark = pytest.mark.gpu


def test_generation_loop_on_gpu_equals_the_loop_on_the_oracle(built_lib):
    from helpers import oracle_specs
    from oracle import pysim
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    from funsearch_kubernetes_simulator_b200.evolution import BatchedFunSearch, mutation_generator
    from funsearch_kubernetes_simulator_b200.variants import base_policies
    cluster, trace = ClusterSpec.load("gpu16"), Trace.load("default").head(1500)
    base = [base_policies()[k] for k in ("fs_first_fit", "fs_best_fit", "ts_funsearch_4816", "template_cpu_mem")]
    ev = BatchEvaluator(cluster, [trace])
    cs, ts = oracle_specs(cluster, trace)

    def cpu_population(codes):
        out = []
        for c in codes:
            try:
                out.append(pysim.run_episode(cs, ts, pysim.scorer_from_source(c)).score)
            except Exception:  # noqa: BLE001  (evaluate_policy_standalone: any exception -> 0)
                out.append(0)
        return out
    calls = []

    def gpu_population(codes):
        calls.append(len(codes))
        return ev.evaluate_population(codes)
    gpu = BatchedFunSearch(gpu_population, mutation_generator(7), evaluate_full=ev.evaluate_policy_full, population_size=14, elite_size=4)
    cpu = BatchedFunSearch(cpu_population, mutation_generator(7), population_size=14, elite_size=4)
    gpu.initialize_population(base)
    cpu.initialize_population(base)
    for _ in range(3):
        gi, ci = gpu.evolve_generation(), cpu.evolve_generation()
        assert (gi["requested"], gi["generated"], gi["kept"], gi["too_similar"]) == (ci["requested"], ci["generated"], ci["kept"], ci["too_similar"])
        assert [c for c, _s in gpu.population] == [c for c, _s in cpu.population]
        assert all(abs(a[1] - b[1]) <= 1e-12 for a, b in zip(gpu.population, cpu.population))
        assert not ev.last_report.not_evaluated      # nothing of this run left the translated subset
    assert calls == [10, 10, 10] and gpu.best_policy == cpu.best_policy and gpu.best_score > 0.3
