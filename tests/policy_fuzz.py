"""Random policy sources and random (pod, node) cases for differential testing of the Python -> CUDA translator
against CPython itself.  Everything is seeded; the grammar stays inside what funsearch/safe_execution.py admits."""
import random
from types import SimpleNamespace as NS

import numpy as np

INT_ATOMS = ["pod.cpu_milli", "pod.memory_mib", "pod.num_gpu", "pod.gpu_milli", "node.cpu_milli_left", "node.memory_mib_left",
             "node.gpu_left", "node.cpu_milli_total", "node.memory_mib_total", "len(node.gpus)", "pod.duration_time",
             "pod.creation_time"]
INT_LITS = ["0", "1", "2", "3", "7", "10", "100", "1000", "-1", "-3", "12345"]
FLOAT_LITS = ["0.0", "0.5", "1.5", "2.0", "-0.25", "0.001", "100.0", "0.3", "1e6", "-7.5"]
POW_EXP = ["2", "0.5", "3", "-1", "0.6", "2.5", "2.0", "0", "-0.5", "1.5"]
PREFIX = '''    if (pod.cpu_milli > node.cpu_milli_left or
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
'''


class Gen:
    def __init__(self, seed):
        self.r = random.Random(seed)
        self.vars = []       # numeric locals defined so far
        self.lists = []      # gpu-list locals defined so far

    def atom(self):
        r = self.r
        k = r.random()
        if self.vars and k < 0.25:
            return r.choice(self.vars)
        if k < 0.6:
            return r.choice(INT_ATOMS)
        if k < 0.75:
            return r.choice(INT_LITS)
        if k < 0.9:
            return r.choice(FLOAT_LITS)
        return self.gpu_reduce()

    def gpu_iter(self):
        r = self.r
        base = r.choice(["node.gpus"] + self.lists) if self.lists else "node.gpus"
        k = r.random()
        if k < 0.15:
            return f"{base}[:pod.num_gpu]"
        if k < 0.25:
            return f"{base}[{r.choice(['1', '-2', '0'])}:]"
        if k < 0.35:
            return f"sorted({base}, key=lambda q: q.gpu_milli_left{r.choice(['', ' * -1', ' % 300'])})"
        return base

    def gpu_term(self, v="g"):
        r = self.r
        return r.choice([f"{v}.gpu_milli_left", f"{v}.gpu_milli_total", f"({v}.gpu_milli_total - {v}.gpu_milli_left)",
                         f"{v}.gpu_milli_left / {v}.gpu_milli_total", f"({v}.gpu_milli_left - pod.gpu_milli) ** 2",
                         f"abs({v}.gpu_milli_left - pod.gpu_milli)", f"{v}.memory_mib_total", f"1"])

    def gpu_reduce(self):
        r = self.r
        cond = r.choice(["", " if g.gpu_milli_left >= pod.gpu_milli", " if g.gpu_milli_left > 0", " if 100 < g.gpu_milli_left < 900"])
        k = r.random()
        if k < 0.45:
            return f"sum({self.gpu_term()} for g in {self.gpu_iter()}{cond})"
        if k < 0.6:
            return f"max({self.gpu_term()} for g in {self.gpu_iter()}{cond})"
        if k < 0.7:
            return f"min([{self.gpu_term()} for g in {self.gpu_iter()}{cond}])"
        if k < 0.85:
            return f"len([g for g in {self.gpu_iter()}{cond}])"
        return f"node.gpus[{r.choice(['0', '1', '-1', 'pod.num_gpu', '7'])}].gpu_milli_left"

    def expr(self, d=0):
        r = self.r
        if d >= 3 or r.random() < 0.25:
            return self.atom()
        k = r.random()
        a = self.expr(d + 1)
        if k < 0.45:
            return f"({a} {r.choice(['+', '-', '*', '+', '-', '*', '/', '//', '%'])} {self.expr(d + 1)})"
        if k < 0.53:
            return f"({a} ** {r.choice(POW_EXP)})"
        if k < 0.6:
            return f"{r.choice(['abs', 'int', 'float', 'round', '-'])}({a})"
        if k < 0.72:
            n = r.choice([2, 2, 3])
            return f"{r.choice(['max', 'min'])}({', '.join([a] + [self.expr(d + 1) for _ in range(n - 1)])})"
        if k < 0.82:
            return f"({a} if {self.cond(d + 1)} else {self.expr(d + 1)})"
        if k < 0.88:
            return f"({a} {r.choice(['and', 'or'])} {self.expr(d + 1)})"
        if k < 0.895:
            return f"({a} if pod.num_gpu in {r.choice(['{1, 2, 4}', '(0, 8)', '[1]'])} else {r.choice(INT_LITS)})"
        if k < 0.91:
            return f"math.sqrt(abs({a}))"
        if k < 0.94:
            return f"math.log({r.choice(['1 + ', '0.5 + ', '1e-9 + '])}abs({a}))"
        if k < 0.96:
            return f"math.exp({r.choice(['-', ''])}abs({a}) / {r.choice(['1000.0', '100000', '7.0'])})"
        return f"({a} / max(1, {self.expr(d + 1)}))"

    def cond(self, d=0):
        r = self.r
        k = r.random()
        a, b = self.expr(d + 1), self.expr(d + 1)
        op = r.choice(["<", "<=", ">", ">=", "==", "!="])
        if k < 0.6:
            return f"{a} {op} {b}"
        if k < 0.75:
            return f"{r.choice(FLOAT_LITS)} < {a} {r.choice(['<', '<='])} {b}"
        if k < 0.88:
            return f"({a} {op} {b}) {r.choice(['and', 'or'])} ({self.expr(d + 1)} {r.choice(['<', '>'])} {self.expr(d + 1)})"
        return f"not ({a} {op} {b})"

    def stmt(self, ind, d=0):
        r = self.r
        pad = "    " * ind
        k = r.random()
        if k < 0.45 or d >= 2:
            name = r.choice(["a", "b", "c", "score", "w", "t"])
            if name in self.vars and r.random() < 0.4:
                line = f"{pad}{name} {r.choice(['+=', '-=', '*=', '/='])} {self.expr(1)}\n"
            else:
                line = f"{pad}{name} = {self.expr(0)}\n"
                if name not in self.vars:
                    self.vars.append(name)
            return line
        if k < 0.7:
            vars_before = list(self.vars)
            s = f"{pad}if {self.cond()}:\n" + self.block(ind + 1, d + 1)
            if r.random() < 0.3:
                s += f"{pad}elif {self.cond()}:\n" + self.block(ind + 1, d + 1)
            if r.random() < 0.6:
                s += f"{pad}else:\n" + self.block(ind + 1, d + 1)
            self.vars = vars_before  # names first bound inside a branch may be unbound afterwards
            return s
        if k < 0.82:
            acc = r.choice(["acc", "cnt"])
            s = f"{pad}{acc} = {r.choice(['0', '0.0', '1'])}\n{pad}for gg in {self.gpu_iter()}:\n"
            s += f"{pad}    if gg.gpu_milli_left {r.choice(['>=', '<', '>'])} {r.choice(['pod.gpu_milli', '500', '0'])}:\n"
            s += f"{pad}        {acc} {r.choice(['+=', '-='])} {self.gpu_term('gg')}\n"
            if acc not in self.vars:
                self.vars.append(acc)
            return s
        if k < 0.9:
            name = r.choice(["fits", "tight"])
            s = f"{pad}{name} = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]\n"
            if name not in self.lists and ind == 1:
                self.lists.append(name)
                return s
            return f"{pad}pass\n"
        if k < 0.93:
            return f"{pad}for i in range({r.choice(['3', 'pod.num_gpu', 'len(node.gpus)', '1, 4'])}):\n{pad}    score = score + i * {self.expr(2)}\n"
        if k < 0.97 and d < 2:
            # try / except: only a bare handler catches in the reference's sandbox (its builtins hold no exception classes)
            vars_before = list(self.vars)
            s = f"{pad}try:\n" + self.block(ind + 1, d + 1)
            if r.random() < 0.5:
                s += f"{pad}    score = score + node.gpus[{r.choice(['0', '5', 'pod.num_gpu'])}].gpu_milli_left / {r.choice(['pod.num_gpu', '(node.gpu_left - 1)', '2'])}\n"
            s += f"{pad}except{r.choice(['', '', '', ' ZeroDivisionError', ' Exception'])}:\n" + self.block(ind + 1, d + 1)
            if r.random() < 0.3:
                s += f"{pad}else:\n" + self.block(ind + 1, d + 1)
            if r.random() < 0.2 and "return" not in s:
                s += f"{pad}finally:\n{pad}    score = score + {r.choice(['1', '0.5', 'pod.num_gpu'])}\n"
            self.vars = vars_before
            return s
        return f"{pad}return {self.expr(0)}\n"

    def block(self, ind, d):
        return "".join(self.stmt(ind, d) for _ in range(self.r.choice([1, 1, 2])))

    def policy(self):
        self.vars, self.lists = ["score"], []
        src = "def priority_function(pod, node):\n"
        if self.r.random() < 0.5:
            src += PREFIX
        src += "    score = " + self.r.choice(["0.0", "0", "1"]) + "\n"
        for _ in range(self.r.choice([1, 2, 3, 4])):
            src += self.stmt(1)
        src += "    return " + self.r.choice(["max(1, int(score))", "score", self.expr(1), "int(score)", "max(0, score)"]) + "\n"
        return src


def random_cases(n, seed):
    """n independent (pod, node) cases: arrays for the device hook plus entity-shaped objects for CPython."""
    rng = np.random.default_rng(seed)
    pods = np.zeros((n, 6), dtype=np.int64)
    ns = np.zeros((n, 12), dtype=np.int32)   # rows of KS_NS_FIELDS (include/ks_abi.h)
    nc = np.zeros((n, 20), dtype=np.int32)   # rows of KS_NC_FIELDS
    objs = []
    for i in range(n):
        ngpus = int(rng.choice([0, 1, 2, 4, 8, 3]))
        cpu_t = int(rng.choice([32000, 64000, 96000, 128000]))
        mem_t = int(rng.choice([131072, 262144, 393216, 786432]))
        cpu_l = int(rng.integers(-2000, cpu_t + 1)) if rng.random() < 0.9 else cpu_t
        mem_l = int(rng.integers(0, mem_t + 1))
        gl = int(rng.integers(0, ngpus + 1))
        gmt = [1000] * ngpus
        gml = [int(rng.choice([0, 1000, int(rng.integers(0, 1001))])) for _ in range(ngpus)]
        gmem = [int(rng.choice([16280, 32510, 81251]))] * ngpus
        p_cpu = int(rng.choice([0, int(rng.integers(0, 64001))]))
        p_mem = int(rng.integers(0, 262145))
        p_ng = int(rng.choice([0, 1, 1, 1, 2, 4, 8]))
        p_gm = int(rng.choice([0, 1000, int(rng.integers(1, 1001))])) if p_ng else 0
        ct, du = int(rng.integers(0, 10 ** 7)), int(rng.integers(0, 10 ** 6))
        pods[i] = (p_cpu, p_mem, p_ng, p_gm, ct, du)
        ns[i, 0], ns[i, 1], ns[i, 2] = cpu_l, mem_l, gl
        nc[i, 0], nc[i, 1], nc[i, 2] = cpu_t, mem_t, ngpus
        for g in range(ngpus):
            ns[i, 4 + g] = gml[g]
            nc[i, 4 + g] = gmt[g]
            nc[i, 12 + g] = gmem[g]
        pod = NS(pod_id=f"p{i}", cpu_milli=p_cpu, memory_mib=p_mem, num_gpu=p_ng, gpu_milli=p_gm, gpu_spec="", creation_time=ct,
                 duration_time=du, assigned_node="", assigned_gpus=[])
        node = NS(node_id=f"n{i}", cpu_milli_left=cpu_l, cpu_milli_total=cpu_t, memory_mib_left=mem_l, memory_mib_total=mem_t,
                  gpu_left=gl, gpus=[NS(memory_mib_left=gmem[g], memory_mib_total=gmem[g], gpu_milli_left=gml[g], gpu_milli_total=gmt[g])
                                     for g in range(ngpus)])
        objs.append((pod, node))
    return pods, ns, nc, objs


def cpython_scores(src, objs):
    """What FunSearchScheduler.__call__ returns for each case; -1 where it raises, -2 where the int leaves int64."""
    import sys
    import os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    from oracle import pysim
    env = pysim.sandbox_env()
    exec(src, env)
    fn = env["priority_function"]
    out = []
    for pod, node in objs:
        try:
            v = int(max(0, fn(pod, node)))
            out.append(v if v < 2 ** 63 else -2)
        except Exception:  # noqa: BLE001
            out.append(-1)
    return out
