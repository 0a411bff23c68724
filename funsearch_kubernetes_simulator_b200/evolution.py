"""The generation loop around the batched evaluator: SimpleFunSearch's population bookkeeping
(funsearch/funsearch_integration.py:175-215 initialize_population / _is_too_similar, :487-572 evolve_generation,
:574-599 run_evolution) with its ProcessPoolExecutor block (:535-562) replaced by ONE evaluate_population call per
generation and the cap of 8 new policies per generation (:499) lifted to the population's size.

The policy generator is a plug-in: the reference asks an LLM (safe_execution.py:283-317, network); anything with the
signature generate(index, elite_policies, feedback) -> source | None fits, e.g. `mutation_generator` below, the seeded
stand-in used on the GPU box.  Differences from the reference that are deliberate and visible:
  * new policies are processed in index order (the reference takes them in completion order, which only matters for
    which of two equal scores is called "best");
  * `new_per_generation` replaces the literal 8.
"""
from __future__ import annotations

import difflib
import random
from typing import Callable, List, Optional, Sequence, Tuple

Policy = Tuple[str, float]
FEEDBACK = ("Elite policies achieve good performance by balancing resource utilization "
            "and considering GPU/CPU workload separation. "
            "Focus on: CPU/mem/GPU util, efficiency, GPU placement strategies, fragmentation reduction.")


class BatchedFunSearch:
    def __init__(self, evaluate_population: Callable[[List[str]], List[float]],
                 generate: Callable[[int, List[Policy], str], Optional[str]],
                 evaluate_full: Optional[Callable[[str], Optional[float]]] = None,
                 population_size: int = 20, elite_size: int = 5, similarity_threshold: float = 0.85,
                 new_per_generation: Optional[int] = None, early_stop_threshold: float = 1.0):
        self.evaluate_population = evaluate_population
        self.generate = generate
        # _evaluate_policy_full's convention: None when evaluating the policy raised (:457-459)
        self.evaluate_full = evaluate_full or (lambda code: self.evaluate_population([code])[0])
        self.population_size = population_size
        self.elite_size = elite_size
        self.similarity_threshold = similarity_threshold
        self.new_per_generation = new_per_generation
        self.early_stop_threshold = early_stop_threshold
        self.population: List[Policy] = []
        self.generation = 0
        self.best_policy: Optional[str] = None
        self.best_score = float("-inf")
        self.history: List[dict] = []

    def initialize_population(self, baseline_policies: Sequence[str]) -> None:
        """funsearch_integration.py:175-206: baselines are evaluated one by one, failures (None) are dropped."""
        for code in baseline_policies:
            score = self.evaluate_full(code)
            if score is not None:
                self.population.append((code, score))
                if score > self.best_score:
                    self.best_score, self.best_policy = score, code
        self.population.sort(key=lambda x: x[1], reverse=True)
        self.population = self.population[:self.population_size]

    def _is_too_similar(self, new_code: str, new_score: float) -> bool:
        """:208-215: against policies of the current population with an equal or better score."""
        for existing_code, existing_score in self.population:
            if existing_score >= new_score:
                if difflib.SequenceMatcher(None, new_code.strip(), existing_code.strip()).ratio() >= self.similarity_threshold:
                    return True
        return False

    def evolve_generation(self) -> dict:
        self.generation += 1
        self.population.sort(key=lambda x: x[1], reverse=True)
        elite = self.population[:self.elite_size]
        room = self.population_size - len(elite)
        n_new = room if self.new_per_generation is None else min(self.new_per_generation, room)
        info = {"generation": self.generation, "requested": n_new, "generated": 0, "kept": 0, "too_similar": 0}
        if n_new <= 0:
            self.history.append(info)
            return info
        generated = [(i, self.generate(i, elite, FEEDBACK)) for i in range(n_new)]
        generated = [(i, c) for i, c in generated if c]
        info["generated"] = len(generated)
        if not generated:
            self.history.append(info)
            return info
        scores = self.evaluate_population([c for _i, c in generated])      # the whole generation: one batched call
        new_policies: List[Policy] = []
        for (_i, code), score in zip(generated, scores):
            if score is None:
                continue
            if self._is_too_similar(code, score):
                info["too_similar"] += 1
                continue
            new_policies.append((code, score))
            if score > self.best_score:
                self.best_score, self.best_policy = score, code
        info["kept"] = len(new_policies)
        merged = elite + new_policies
        merged.sort(key=lambda x: x[1], reverse=True)
        self.population = merged[:self.population_size]
        info["best_score"] = self.best_score
        self.history.append(info)
        return info

    def run_evolution(self, generations: int, baseline_policies: Optional[Sequence[str]] = None) -> Tuple[Optional[str], float]:
        if not self.population:
            self.initialize_population(baseline_policies or [])
        for _ in range(generations):
            self.evolve_generation()
            if self.best_score >= self.early_stop_threshold:
                break
        return self.best_policy, self.best_score


def mutation_generator(seed: int = 0) -> Callable[[int, List[Policy], str], Optional[str]]:
    """Seeded stand-in for LLMCodeGenerator.generate_policy: picks up to two parents like _generate_single_policy
    (:466) and returns a mutant of the first (variants.mutate_source), or None when the mutant would not pass the
    reference's validators (the LLM path drops such sources too, safe_execution.py:307-313)."""
    from .variants import admitted_by_reference_validators, mutate_source
    rng = random.Random(seed)

    def generate(index: int, elite: List[Policy], feedback: str) -> Optional[str]:
        parents = rng.sample(elite, min(2, len(elite)))
        if not parents:
            return None
        k = 1
        while k < 4 and rng.random() < 0.5:
            k += 1
        try:
            src = mutate_source(parents[0][0], rng, k)
        except (SyntaxError, ValueError, IndexError):
            return None
        return src if admitted_by_reference_validators(src) else None
    return generate
