"""Host logic of the generation loop (evolution.BatchedFunSearch) with a fake evaluator: elite selection, the lifted cap on
new policies, similarity dedupe against equal-or-better members only, truncation, None handling."""
import hashlib


def _fake_score(code: str) -> float:
    return int(hashlib.sha256(code.encode()).hexdigest()[:6], 16) / 0xFFFFFF


def test_generation_bookkeeping_matches_the_reference_rules():
    from funsearch_kubernetes_simulator_b200.evolution import BatchedFunSearch, mutation_generator
    from funsearch_kubernetes_simulator_b200.variants import base_policies
    calls = []

    def evaluate(codes):
        calls.append(len(codes))
        return [_fake_score(c) for c in codes]
    base = list(base_policies().values())[:4]
    fs = BatchedFunSearch(evaluate, mutation_generator(3), evaluate_full=lambda c: None if "worst" in c.lower() else _fake_score(c),
                          population_size=12, elite_size=3, similarity_threshold=0.999)
    fs.initialize_population(base + ["def priority_function(pod, node):\n    return 'worst'\n"])
    assert len(fs.population) == 4 and fs.population == sorted(fs.population, key=lambda x: -x[1])   # the None one is dropped
    info = fs.evolve_generation()
    assert calls == [9] and info["requested"] == 9                 # population_size - elites, not min(8, .): ONE call
    assert len(fs.population) <= 12 and fs.population == sorted(fs.population, key=lambda x: -x[1])
    assert fs.best_score == fs.population[0][1] == max(s for _c, s in fs.population)
    elite_before = sorted(fs.population, key=lambda x: -x[1])[:3]
    fs.evolve_generation()
    assert all(e in fs.population or e[1] < fs.population[-1][1] for e in elite_before)   # elites survive unless outscored
    # reference cap: new_per_generation=8 reproduces min(8, .)
    fs8 = BatchedFunSearch(evaluate, mutation_generator(3), population_size=40, elite_size=3, new_per_generation=8)
    fs8.initialize_population(base)
    assert fs8.evolve_generation()["requested"] == 8


def test_similarity_dedupe_only_against_equal_or_better():
    from funsearch_kubernetes_simulator_b200.evolution import BatchedFunSearch
    a = "def priority_function(pod, node):\n    return 100\n"
    near = "def priority_function(pod, node):\n    return 101\n"
    scores = {a: 0.5, near: 0.4}
    fs = BatchedFunSearch(lambda cs: [scores[c] for c in cs], lambda i, e, f: near, population_size=5, elite_size=1,
                          similarity_threshold=0.85)
    fs.population = [(a, 0.5)]
    assert fs.evolve_generation()["too_similar"] == 4          # near-duplicate of a better member: skipped
    scores[near] = 0.6
    info = fs.evolve_generation()                               # better than everything it resembles: kept (first one; the
    assert info["kept"] == 4 and fs.best_policy == near        # dedupe looks at the OLD population only, like the reference)
