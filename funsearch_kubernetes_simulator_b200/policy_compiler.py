"""Policy compiler: restricted-Python `priority_function(pod, node)` source -> CUDA C device function.

Input language = what the reference's validators admit (funsearch/safe_execution.py:15-124) and its
policy template produces (:171-224); the wrapper semantics int(max(0, score)) of
FunSearchScheduler.__call__ (funsearch/funsearch_integration.py:91-101) are folded into the
generated function's return value.  The output is compiled by NVRTC (--fmad=false) inside
libks_b200.so and linked with the replay kernel; arithmetic helpers live in csrc/ks_pyrt.cuh.

Translation scheme
  * a policy = a STRUCTURE + a vector of float literals: every non-zero float literal of the source becomes a read of
    the argument `kf` (in source order), so that sources which differ only in those constants -- what most mutations
    of a FunSearch generation change -- translate to the same CUDA text, are compiled once and share their code in the
    SM's instruction cache; the values travel as fp64 bit patterns, so nothing is rounded or folded differently from
    CPython (which folds nothing the translator does not).  Int literals stay structural: the bit-range tracking
    depends on them;
  * static types: int (int64 + overflow => 'unsupported', never a silent wrap), float (fp64), bool,
    num (int-or-float decided at run time, e.g. max(0, score)), gpu (index into node.gpus), small
    lists (GPU lists, int/float lists, literal lists);
  * variable types are flow-insensitive joins over all assignments, found by iterating the
    translation to a fixed point; integer ranges are tracked in bits so that provably small
    arithmetic needs no overflow checks and converts to fp64 exactly;
  * every sub-expression is evaluated into a C temporary in Python's left-to-right order; anything
    that raises in CPython becomes `return KS_RET_EXCEPTION`, anything outside the subset that can
    only be detected at run time `return KS_RET_UNSUPPORTED`;
  * constructs outside the subset raise PolicyCompileError (the policy's episodes then report
    status 'unsupported' instead of a wrong score).
"""
from __future__ import annotations

import ast
import hashlib
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

INT, FLOAT, BOOL, NUM, GPU, GPULIST, ILIST, FLIST, STATIC, POD, NODE, NONE, GPUS, LAMBDA, MODULE, STR, ELIST = (
    "int", "float", "bool", "num", "gpu", "gpulist", "ilist", "flist", "static", "pod", "node", "none", "gpus",
    "lambda", "module", "str", "elist")   # elist: `xs = []` whose element type the .append() calls decide
NUMERIC = (INT, FLOAT, BOOL, NUM)
CTYPE = {INT: "ks_int", FLOAT: "double", BOOL: "int", NUM: "KsNum", GPU: "int", GPULIST: "KsGpuList",
         ILIST: "KsIntList", FLIST: "KsFloatList", ELIST: "KsIntList"}

ALLOWED_BUILTINS = {"abs", "min", "max", "sum", "len", "range", "enumerate", "int", "float", "bool", "str", "round",
                    "sorted"}
MATH_FUNCS = {"sqrt", "log", "exp", "pow", "sin", "cos", "tan"}
OPERATOR_FUNCS = {"add": ast.Add, "sub": ast.Sub, "mul": ast.Mult, "truediv": ast.Div, "mod": ast.Mod}

POD_FIELDS = {"cpu_milli": ("p_cpu", 31), "memory_mib": ("p_mem", 31), "num_gpu": ("p_ngpu", 5),
              "gpu_milli": ("p_gmilli", 25), "creation_time": ("p_ctime", 40), "duration_time": ("p_dur", 40)}
NODE_STATE = {"cpu_milli_left": "KS_NS_CPU_LEFT", "memory_mib_left": "KS_NS_MEM_LEFT", "gpu_left": "KS_NS_GPU_LEFT"}
NODE_CONST = {"cpu_milli_total": "KS_NC_CPU_TOTAL", "memory_mib_total": "KS_NC_MEM_TOTAL"}
GPU_FIELDS = {"gpu_milli_left": ("ns", "KS_NS_GPU_MILLI_LEFT"), "gpu_milli_total": ("nc", "KS_NC_GPU_MILLI_TOTAL"),
              "memory_mib_total": ("nc", "KS_NC_GPU_MEM_TOTAL"), "memory_mib_left": ("nc", "KS_NC_GPU_MEM_TOTAL")}

LOOP_BUDGET = 200000


class PolicyCompileError(Exception):
    """The source is outside the subset this compiler translates."""


@dataclass
class Val:
    ty: str
    code: str = ""
    bits: int = 64                      # |value| < 2**bits for INT/BOOL
    items: Optional[list] = None        # STATIC: element Vals
    node: Optional[ast.AST] = None      # LAMBDA
    const: object = None                # compile-time constant value if known (int/float/bool/str)
    lit: Optional[tuple] = None         # (ast.Constant node, sign) when this is an int literal of the source


class _Lifted:
    """A float literal that lives in the policy's constant vector, not in the structure's code."""
    __slots__ = ("index", "value")

    def __init__(self, index: int, value: float):
        self.index = index
        self.value = value


# handle of each compiled structure (ks_policy_compile*), shared by all policies with that structure
_STRUCT_HANDLES: Dict[str, Optional[int]] = {}


@dataclass
class CompiledPolicy:
    source: str
    cuda_source: str                 # the structure's device function (identical for all policies sharing it)
    symbol: str                      # the structure's symbol: "ks_st_" + hash of the generated code
    var_types: Dict[str, str] = field(default_factory=dict)
    flags: int = 0   # KS_PF_* (include/ks_abi.h); bit 0: the template's feasibility prefix is evaluated by the kernel
    consts: Tuple[float, ...] = ()   # this policy's float literals, in the order the structure reads them (kf[i])

    @property
    def handle(self) -> Optional[int]:
        return _STRUCT_HANDLES.get(self.symbol)

    @handle.setter
    def handle(self, h: Optional[int]) -> None:
        _STRUCT_HANDLES[self.symbol] = h


def _join(a: Optional[str], b: str) -> str:
    if a is None or a == b:
        return b
    if ELIST in (a, b) and {a, b} <= {ELIST, GPULIST, ILIST, FLIST}:
        return b if a == ELIST else a
    s = {a, b}
    if s <= {INT, BOOL}:
        return INT
    if s <= set(NUMERIC):
        return NUM
    raise PolicyCompileError(f"a variable holds both {a} and {b}")


def _flit(x: float) -> str:
    if x != x:
        return "__longlong_as_double(0x7ff8000000000000LL)"
    if x in (float("inf"), float("-inf")):
        return "__longlong_as_double(0x7ff0000000000000LL)" if x > 0 else "__longlong_as_double((long long)0xfff0000000000000ULL)"
    if x == 0.0:
        return "-0.0" if str(x).startswith("-") else "0.0"
    return float(x).hex()


class _Returned(Exception):
    pass


def _def_as_lambda(fn: ast.FunctionDef) -> Optional[ast.Lambda]:
    """A helper `def f(a, b): [docstring] return <expr>` is the lambda `lambda a, b: <expr>` (inlined at its call sites
    like any lambda); anything longer is not translated."""
    body = [st for st in fn.body if not (isinstance(st, ast.Expr) and isinstance(st.value, ast.Constant))]
    a = fn.args
    if len(body) != 1 or not isinstance(body[0], ast.Return) or body[0].value is None or fn.decorator_list or a.vararg or a.kwarg \
            or a.kwonlyargs or a.defaults or a.posonlyargs:
        return None
    return ast.copy_location(ast.Lambda(args=a, body=body[0].value), fn)


class _Translator:
    def __init__(self, fn: ast.FunctionDef, consts: Dict[str, object], var_types: Dict[str, str],
                 var_bits: Dict[str, int], flagged: set, base_bits: Optional[Dict[str, int]] = None, skip: int = 0,
                 lift: Optional[Dict[int, int]] = None, kvalues: Optional[List[float]] = None,
                 ilift: Optional[Dict[tuple, int]] = None):
        self.ilift = ilift if ilift is not None else {}   # (id(int literal node), sign) -> index, assigned on first use
        self.lift = lift if lift is not None else {}   # id(ast.Constant node) -> index in the policy's constant vector
        self.kvalues = kvalues if kvalues is not None else []   # the vector itself (grows when an int literal is lifted)
        self.base_bits = dict(base_bits or {})  # bits of each int variable's non-additive assignments (input)
        self.seen_base: Dict[str, int] = {}
        self.skip = skip                        # leading statements handled by the kernel (hoisted prefix)
        self.fn = fn
        self.consts = consts
        self.var_types = dict(var_types)      # declared (joined) types, input of this pass
        self.var_bits = dict(var_bits)
        self.flagged = set(flagged)           # variables that need a run-time 'bound' flag
        self.seen_types: Dict[str, str] = {}  # joins observed in this pass
        self.seen_bits: Dict[str, int] = {}
        self.need_flag: set = set()
        self.lines: List[str] = []
        self.ind = 1
        self.ntmp = 0
        self.scopes: List[Dict[str, Val]] = []  # comprehension / lambda bindings
        self.static_vars: Dict[str, Val] = {}
        self.defined: set = set()
        self.uses_budget = False
        self.try_labels: List[str] = []         # handler labels of the enclosing try bodies
        self.ntry = 0
        args = [a.arg for a in fn.args.args]
        if len(args) != 2 or fn.args.vararg or fn.args.kwarg or fn.args.kwonlyargs or fn.args.defaults:
            raise PolicyCompileError("priority_function must take exactly (pod, node)")
        self.pod_name, self.node_name = args
        # what the function's text alone decides is scanned once per function, not once per pass of the type fixed point
        scan = getattr(fn, "_ks_scan", None)
        if scan is None:
            scan = fn._ks_scan = self._scan(fn)
        self.appended, counts = scan
        self.assign_count: Dict[str, int] = dict(counts)

    @staticmethod
    def _scan(fn: ast.FunctionDef):
        assign_count: Dict[str, int] = {}
        # names that are built up with .append(): `xs = []` then `xs.append(v)` in loops / branches
        appended = set()
        for n in (getattr(fn, "_ks_nodes", None) or ast.walk(fn)):
            if isinstance(n, ast.Call):
                if (isinstance(n.func, ast.Attribute) and n.func.attr in ("append", "extend")
                        and isinstance(n.func.value, ast.Name)):
                    appended.add(n.func.value.id)
            elif isinstance(n, (ast.Assign, ast.AugAssign, ast.AnnAssign)):
                tg = n.targets if isinstance(n, ast.Assign) else [n.target]
                for t in tg:
                    for nm in ([t] if isinstance(t, ast.Name) else ast.walk(t)):
                        if isinstance(nm, ast.Name):
                            assign_count[nm.id] = assign_count.get(nm.id, 0) + 1
            elif isinstance(n, (ast.For,)):
                for nm in ([n.target] if isinstance(n.target, ast.Name) else ast.walk(n.target)):
                    if isinstance(nm, ast.Name):
                        assign_count[nm.id] = assign_count.get(nm.id, 0) + 2
            elif isinstance(n, ast.FunctionDef) and n is not fn:
                assign_count[n.name] = assign_count.get(n.name, 0) + 1
        return appended, assign_count

    # ------------------------------------------------------------------ emission helpers
    def emit(self, s: str):
        self.lines.append("  " * self.ind + s)

    def new(self, prefix="t") -> str:
        self.ntmp += 1
        return f"{prefix}{self.ntmp}"

    def tmp(self, ty: str, expr: str, bits: int = 64, const=None) -> Val:
        name = self.new()
        self.emit(f"const {CTYPE[ty]} {name} = {expr};")
        return Val(ty, name, bits=bits, const=const)

    def exc(self) -> str:
        """The statement that raises: leave the function, or -- inside a try body -- jump to the handler."""
        if self.try_labels:
            return f"{{ if (!st) st = 1; goto {self.try_labels[-1]}; }}"
        return "KS_RETURN(KS_RET_EXCEPTION);"

    def raise_exc(self):
        self.emit(self.exc())

    def unsupported_rt(self):
        self.emit("KS_RETURN(KS_RET_UNSUPPORTED);")

    def check_st(self):
        # The status word is sticky (ks_pyrt.cuh): outside a try body it is tested where the function returns.  Inside
        # one, control must reach the handler before anything after the failing operation takes effect.
        if self.try_labels:
            self.emit(f"if (st) goto {self.try_labels[-1]};")

    def fail(self, node, msg):
        line = getattr(node, "lineno", "?")
        raise PolicyCompileError(f"line {line}: {msg}")

    # ------------------------------------------------------------------ conversions
    def as_int(self, v: Val) -> Val:
        if v.ty == INT:
            return v
        if v.ty == BOOL:
            return Val(INT, f"(ks_int){v.code}", bits=1, const=None if v.const is None else int(v.const))
        raise PolicyCompileError(f"expected an int, got {v.ty}")

    def as_double(self, v: Val) -> str:
        if v.ty == FLOAT:
            return v.code
        if v.ty in (INT, BOOL):
            if v.const is not None:
                if v.ty == INT and v.lit is not None and 2 < abs(int(v.const)) < (1 << 53):
                    # An int literal used where a float is needed (`u * 100`, `score -= 20`, `x > 50` with float u,
                    # score, x): CPython converts it exactly, so float(literal) can live in the constant vector like
                    # a float literal, and policies that differ in such coefficients still share a structure.
                    # (Ints that stay ints are structure: the bit-range tracking depends on their values.)
                    key = (id(v.lit[0]), v.lit[1])   # kept apart from the float literals' ids: an int literal is
                    k = self.ilift.get(key)          # only lifted where it is used as a float
                    if k is None:
                        k = self.ilift[key] = len(self.kvalues)
                        self.kvalues.append(float(int(v.const)))
                    return f"ks_kf(kf, {k})"
                return _flit(float(int(v.const)))
            return f"ks_i2d((ks_int){v.code})" if v.bits > 53 else f"(double){v.code}"
        if v.ty == NUM:
            return f"ks_num_d({v.code})"
        raise PolicyCompileError(f"expected a number, got {v.ty}")

    def as_num(self, v: Val) -> str:
        if v.ty == NUM:
            return v.code
        if v.ty == FLOAT:
            return f"ks_num_f({v.code})"
        if v.ty in (INT, BOOL):
            return f"ks_num_i((ks_int){v.code})"
        raise PolicyCompileError(f"expected a number, got {v.ty}")

    def convert(self, v: Val, ty: str) -> str:
        if v.ty == ty:
            return v.code
        if ty == INT and v.ty == BOOL:
            return f"(ks_int){v.code}"
        if ty == NUM:
            return self.as_num(v)
        raise PolicyCompileError(f"cannot store {v.ty} into a {ty} variable")

    @staticmethod
    def is_set(v: Val) -> bool:
        return v.ty == STATIC and v.const == "<set>"

    @staticmethod
    def is_dict(v: Val) -> bool:
        return v.ty == STATIC and isinstance(v.const, tuple) and len(v.const) == 2 and v.const[0] == "<dict>"

    def truth(self, v: Val) -> str:
        if v.ty == BOOL:
            return v.code
        if v.ty == INT:
            return f"({v.code} != 0)"
        if v.ty == FLOAT:
            return f"({v.code} != 0.0)"
        if v.ty == NUM:
            return f"ks_num_truth({v.code})"
        if v.ty in (GPULIST, ILIST, FLIST):
            return f"({v.code}.n != 0)"
        if v.ty == GPUS:
            return "(ngpus != 0)"
        if v.ty == STATIC:
            return "1" if v.items else "0"
        if v.ty == NONE:
            return "0"
        if v.ty in (GPU, POD, NODE, LAMBDA, MODULE):
            return "1"
        if v.ty == STR:
            return "1" if v.const else "0"
        raise PolicyCompileError(f"truth value of {v.ty}")

    # ------------------------------------------------------------------ names
    def lookup(self, name: str, node) -> Val:
        for sc in reversed(self.scopes):
            if name in sc:
                return sc[name]
        if name == self.pod_name:
            return Val(POD)
        if name == self.node_name:
            return Val(NODE)
        if name in self.static_vars:
            return self.static_vars[name]
        if name in self.assign_count:
            ty = self.var_types.get(name)
            if ty is None:
                # read before any assignment was translated in this pass -> UnboundLocalError unless a later
                # pass learns the type (loops); treat as not-yet-known
                self.need_flag.add(name)
                self.raise_exc()
                raise _Returned()
            if name not in self.defined:
                self.need_flag.add(name)
                if name in self.flagged:
                    self.emit(f"if (!b_{name}) {self.exc()}  // UnboundLocalError")
            return Val(ty, f"v_{name}", bits=self.var_bits.get(name, 64))
        if name in self.consts:
            if isinstance(self.consts[name], ast.Lambda):
                return Val(LAMBDA, node=self.consts[name])
            return self.const_val(self.consts[name], node)
        if name == "math" or name == "operator":
            return Val(MODULE, name)
        # NameError at run time (bare sqrt(), print, any, ...)
        self.raise_exc()
        raise _Returned()

    def const_val(self, c, node) -> Val:
        if isinstance(c, _Lifted):
            return Val(FLOAT, f"ks_kf(kf, {c.index})")   # a run-time value: nothing may be folded on it
        if isinstance(c, bool):
            return Val(BOOL, "1" if c else "0", bits=1, const=c)
        if isinstance(c, int):
            if abs(c) >= 1 << 63:
                self.fail(node, "integer literal beyond int64")
            code = f"{c}LL" if c > -(1 << 63) else "(-9223372036854775807LL-1)"
            lit = (node, 1) if isinstance(node, ast.Constant) and node.value is c else None
            return Val(INT, f"({code})" if c < 0 else code, bits=max(1, abs(c).bit_length()), const=c, lit=lit)
        if isinstance(c, float):
            return Val(FLOAT, f"({_flit(c)})", const=c)
        if c is None:
            return Val(NONE)
        if isinstance(c, str):
            return Val(STR, const=c)
        if isinstance(c, (list, tuple)):
            return Val(STATIC, items=[self.const_val(x, node) for x in c])
        self.fail(node, f"constant of type {type(c).__name__}")

    def store(self, name: str, v: Val, node, additive: bool = False):
        if v.ty == STATIC and not v.items and name in self.appended and not self.scopes:
            # `xs = []` of a list that .append() fills: a real (bounded) list variable; its element type comes from the
            # appends (found by the type-inference iteration), the assignment itself just empties it
            prev = self.seen_types.get(name)
            self.seen_types[name] = _join(prev, ELIST)
            ty = self.var_types.get(name)
            if ty is None:
                ty = self.var_types[name] = ELIST
            if ty not in (ELIST, GPULIST, ILIST, FLIST):
                self.fail(node, f"variable {name!r} holds both a list and {ty}")
            if ty != ELIST:
                self.emit(f"v_{name}.n = 0;")
            if name in self.flagged:
                self.emit(f"b_{name} = 1;")
            self.defined.add(name)
            return
        if v.ty in (STATIC, LAMBDA, NONE, STR, POD, NODE, GPUS, MODULE):
            if v.ty == GPUS:
                v = self.gpus_to_list()
            else:
                if self.assign_count.get(name, 0) != 1 or self.scopes or self.ind != 1:
                    self.fail(node, f"variable {name!r} of type {v.ty} must be assigned exactly once, at function level")
                self.static_vars[name] = v
                return
        if not additive and v.ty in (INT, BOOL):
            self.base_bits[name] = max(self.base_bits.get(name, 0), v.bits)
            self.seen_base[name] = max(self.seen_base.get(name, 0), v.bits)
        prev = self.seen_types.get(name)
        self.seen_types[name] = _join(prev, v.ty)
        self.seen_bits[name] = max(self.seen_bits.get(name, 0), v.bits if v.ty in (INT, BOOL) else 0)
        ty = self.var_types.get(name)
        if ty is None:
            ty = self.var_types[name] = v.ty
            self.var_bits[name] = v.bits
        try:
            ty = _join(ty, v.ty)
        except PolicyCompileError as e:
            self.fail(node, str(e))
        self.var_types[name] = ty
        self.var_bits[name] = max(self.var_bits.get(name, 0), v.bits if v.ty in (INT, BOOL) else 0)
        self.emit(f"v_{name} = {self.convert(v, ty)};")
        if name in self.flagged:
            self.emit(f"b_{name} = 1;")
        self.defined.add(name)

    # ------------------------------------------------------------------ attribute access
    def gpu_field(self, g: Val, attr: str, node) -> Val:
        if attr not in GPU_FIELDS:
            self.raise_exc()  # AttributeError
            raise _Returned()
        arr, f = GPU_FIELDS[attr]
        if arr == "ns":
            return self.tmp(INT, f"(ks_int)ks_ns(ns, {f} + {g.code})", bits=25)
        return self.tmp(INT, f"(ks_int)ks_nc(nc, {f} + {g.code})", bits=31)

    def attribute(self, node: ast.Attribute) -> Val:
        base = self.expr(node.value)
        a = node.attr
        if base.ty == POD:
            if a in POD_FIELDS:
                code, bits = POD_FIELDS[a]
                return self.tmp(INT, f"(ks_int){code}", bits=bits)
            if a in ("pod_id", "gpu_spec", "assigned_node", "assigned_gpus"):
                self.fail(node, f"pod.{a} (non-numeric field) is not supported")
            self.raise_exc()
            raise _Returned()
        if base.ty == NODE:
            if a in NODE_STATE:
                return self.tmp(INT, f"(ks_int)ks_ns(ns, {NODE_STATE[a]})", bits=32)
            if a in NODE_CONST:
                return self.tmp(INT, f"(ks_int)ks_nc(nc, {NODE_CONST[a]})", bits=31)
            if a == "gpus":
                return Val(GPUS)
            if a == "node_id":
                self.fail(node, "node.node_id (string) is not supported")
            self.raise_exc()
            raise _Returned()
        if base.ty == GPU:
            return self.gpu_field(base, a, node)
        if base.ty == MODULE:
            self.fail(node, f"{base.code}.{a} used as a value")
        if base.ty in (GPULIST, ILIST, FLIST, STATIC, GPUS):
            self.fail(node, f"list method .{a} is not supported")
        self.raise_exc()  # AttributeError on numbers etc.
        raise _Returned()

    def gpus_to_list(self) -> Val:
        name = self.new("L")
        self.emit(f"KsGpuList {name}; {name}.n = ngpus;")
        self.emit("#pragma unroll 1")
        self.emit(f"for (int q = 0; q < KS_MAX_GPUS_PER_NODE; ++q) {name}.g[q] = (signed char)q;")
        return Val(GPULIST, name)

    # ------------------------------------------------------------------ arithmetic
    def arith(self, op, a: Val, b: Val, node) -> Val:
        for v in (a, b):
            if v.ty not in NUMERIC:
                if v.ty in (NONE, STR, GPU, POD, NODE, LAMBDA, MODULE) or isinstance(op, (ast.Sub, ast.Div, ast.Mod, ast.Pow, ast.FloorDiv)):
                    if v.ty in (STATIC, GPULIST, ILIST, FLIST, GPUS) and isinstance(op, (ast.Add, ast.Mult)):
                        self.fail(node, "list concatenation/repetition is not supported")
                    self.raise_exc()  # TypeError
                    raise _Returned()
                self.fail(node, f"operator on {v.ty}")
        if isinstance(op, (ast.BitAnd, ast.BitOr, ast.BitXor)):
            if a.ty in (INT, BOOL) and b.ty in (INT, BOOL):
                sym = {ast.BitAnd: "&", ast.BitOr: "|", ast.BitXor: "^"}[type(op)]
                if a.ty == BOOL and b.ty == BOOL:
                    return self.tmp(BOOL, f"({a.code} {sym} {b.code})", bits=1)
                return self.tmp(INT, f"({self.as_int(a).code} {sym} {self.as_int(b).code})", bits=64)
            if FLOAT in (a.ty, b.ty):
                self.raise_exc()
                raise _Returned()
            self.fail(node, "bitwise operator on dynamically typed numbers")
        if isinstance(op, (ast.LShift, ast.RShift, ast.MatMult)):
            self.fail(node, "shift / matmul operators are not supported")
        if NUM in (a.ty, b.ty):
            return self.arith_num(op, a, b, node)
        if a.ty in (INT, BOOL) and b.ty in (INT, BOOL):
            return self.arith_int(op, self.as_int(a), self.as_int(b), node)
        return self.arith_float(op, self.as_double(a), self.as_double(b), node)

    def arith_int(self, op, a: Val, b: Val, node) -> Val:
        if isinstance(op, ast.Add):
            bits = max(a.bits, b.bits) + 1
            if bits <= 62:
                return self.tmp(INT, f"{a.code} + {b.code}", bits=bits)
            v = self.tmp(INT, f"ks_iadd({a.code}, {b.code}, &st)")
            self.check_st()
            return v
        if isinstance(op, ast.Sub):
            bits = max(a.bits, b.bits) + 1
            if bits <= 62:
                return self.tmp(INT, f"{a.code} - {b.code}", bits=bits)
            v = self.tmp(INT, f"ks_isub({a.code}, {b.code}, &st)")
            self.check_st()
            return v
        if isinstance(op, ast.Mult):
            bits = a.bits + b.bits
            if bits <= 62:
                return self.tmp(INT, f"{a.code} * {b.code}", bits=bits)
            v = self.tmp(INT, f"ks_imul({a.code}, {b.code}, &st)")
            self.check_st()
            return v
        if isinstance(op, ast.Div):
            self.emit(f"if ({b.code} == 0) {self.exc()}  // ZeroDivisionError")
            if a.bits <= 53 and b.bits <= 53:
                return self.tmp(FLOAT, f"ks_div((double){a.code}, (double){b.code})")
            v = self.tmp(FLOAT, f"ks_itruediv({a.code}, {b.code}, &st)")
            self.check_st()
            return v
        if isinstance(op, ast.FloorDiv):
            v = self.tmp(INT, f"ks_ifloordiv({a.code}, {b.code}, &st)", bits=min(64, a.bits + 1))
            self.check_st()
            return v
        if isinstance(op, ast.Mod):
            v = self.tmp(INT, f"ks_imod({a.code}, {b.code}, &st)", bits=b.bits)
            self.check_st()
            return v
        if isinstance(op, ast.Pow):
            if b.const is not None and int(b.const) >= 0:
                e = int(b.const)
                bits = a.bits * e if e > 0 else 1
                if bits <= 62 and e <= 4:
                    if e == 0:
                        return self.const_val(1, node)
                    return self.tmp(INT, " * ".join([a.code] * e), bits=bits)
                v = self.tmp(INT, f"ks_ipow({a.code}, {b.code}, &st)", bits=min(64, max(bits, 1)))
                self.check_st()
                return v
            if b.const is not None:  # negative literal exponent: float result
                v = self.tmp(FLOAT, f"ks_ipow_neg({a.code}, {b.code}, &st)")
                self.check_st()
                return v
            v = self.tmp(NUM, f"ks_npow(ks_num_i({a.code}), ks_num_i({b.code}), &st)")
            self.check_st()
            return v
        self.fail(node, f"operator {type(op).__name__}")

    def arith_float(self, op, a: str, b: str, node) -> Val:
        if isinstance(op, ast.Add):
            return self.tmp(FLOAT, f"__dadd_rn({a}, {b})")
        if isinstance(op, ast.Sub):
            return self.tmp(FLOAT, f"__dsub_rn({a}, {b})")
        if isinstance(op, ast.Mult):
            return self.tmp(FLOAT, f"__dmul_rn({a}, {b})")
        fn = {ast.Div: "ks_ftruediv", ast.FloorDiv: "ks_ffloordiv", ast.Mod: "ks_fmod", ast.Pow: "ks_fpow"}.get(type(op))
        if fn is None:
            self.fail(node, f"operator {type(op).__name__}")
        v = self.tmp(FLOAT, f"{fn}({a}, {b}, &st)")
        self.check_st()
        return v

    def arith_num(self, op, a: Val, b: Val, node) -> Val:
        fn = {ast.Add: "ks_nadd", ast.Sub: "ks_nsub", ast.Mult: "ks_nmul", ast.FloorDiv: "ks_nfloordiv",
              ast.Mod: "ks_nmod", ast.Pow: "ks_npow"}.get(type(op))
        if isinstance(op, ast.Div):
            v = self.tmp(FLOAT, f"ks_ntruediv({self.as_num(a)}, {self.as_num(b)}, &st)")
            self.check_st()
            return v
        if fn is None:
            self.fail(node, f"operator {type(op).__name__}")
        # one float operand fixes the result type
        if FLOAT in (a.ty, b.ty) and not isinstance(op, ast.Pow):
            return self.arith_float(op, self.as_double(a), self.as_double(b), node)
        if FLOAT in (a.ty, b.ty):
            return self.arith_float(op, self.as_double(a), self.as_double(b), node)
        v = self.tmp(NUM, f"{fn}({self.as_num(a)}, {self.as_num(b)}, &st)")
        self.check_st()
        return v

    # ------------------------------------------------------------------ comparisons
    def compare_pair(self, op, a: Val, b: Val, node) -> str:
        """C expression (int 0/1) for one comparison."""
        if isinstance(op, (ast.Is, ast.IsNot)):
            if a.ty == NONE or b.ty == NONE:
                same = a.ty == b.ty
                return ("1" if same else "0") if isinstance(op, ast.Is) else ("0" if same else "1")
            self.fail(node, "'is' between non-None values")
        if isinstance(op, (ast.In, ast.NotIn)):
            if b.ty != STATIC:
                self.fail(node, "'in' is only supported against a literal list/tuple/set/dict")
            if self.is_dict(b):
                if a.const is None or a.ty not in (STR, INT):
                    self.fail(node, "'in' on a dict with a non-constant key")
                hit = a.const in b.const[1]
                return ("1" if hit else "0") if isinstance(op, ast.In) else ("0" if hit else "1")
            parts = [self.compare_pair(ast.Eq(), a, it, node) for it in b.items]
            e = "(" + " || ".join(parts) + ")" if parts else "0"
            return e if isinstance(op, ast.In) else f"(!{e})"
        if isinstance(op, (ast.Eq, ast.NotEq)) and (a.ty not in NUMERIC or b.ty not in NUMERIC):
            if a.ty == STR and b.ty == STR:
                r = a.const == b.const
            elif a.ty == NONE or b.ty == NONE:
                r = a.ty == b.ty
            else:
                self.fail(node, f"== between {a.ty} and {b.ty}")
            return ("1" if r else "0") if isinstance(op, ast.Eq) else ("0" if r else "1")
        if a.ty not in NUMERIC or b.ty not in NUMERIC:
            self.raise_exc()  # TypeError: '<' not supported
            raise _Returned()
        sym = {ast.Lt: "<", ast.LtE: "<=", ast.Gt: ">", ast.GtE: ">=", ast.Eq: "==", ast.NotEq: "!="}[type(op)]
        ai = a.ty in (INT, BOOL)
        bi = b.ty in (INT, BOOL)
        if ai and bi:
            return f"({self.as_int(a).code} {sym} {self.as_int(b).code})"
        if a.ty == FLOAT and b.ty == FLOAT:
            return f"({a.code} {sym} {b.code})"
        if (ai and b.ty == FLOAT and a.bits <= 53) or (bi and a.ty == FLOAT and b.bits <= 53):
            return f"({self.as_double(a)} {sym} {self.as_double(b)})"
        if ai and b.ty == FLOAT:
            c = f"ks_cmp_if({self.as_int(a).code}, {b.code})"
        elif bi and a.ty == FLOAT:
            c = f"(-ks_cmp_if({self.as_int(b).code}, {a.code}))"
            t = self.tmp(INT, c, bits=2)
            # nan: ks_cmp_if gives 2, negated -2 -> every comparison but != is false
            tests = {"<": f"({t.code} == -1)", "<=": f"({t.code} == -1 || {t.code} == 0)", ">": f"({t.code} == 1)",
                     ">=": f"({t.code} == 1 || {t.code} == 0)", "==": f"({t.code} == 0)", "!=": f"({t.code} != 0)"}
            return tests[sym]
        else:
            c = f"ks_cmp_nn({self.as_num(a)}, {self.as_num(b)})"
        t = self.tmp(INT, c, bits=2)
        tests = {"<": f"({t.code} == -1)", "<=": f"({t.code} == -1 || {t.code} == 0)", ">": f"({t.code} == 1)",
                 ">=": f"({t.code} == 1 || {t.code} == 0)", "==": f"({t.code} == 0)", "!=": f"({t.code} != 0)"}
        return tests[sym]

    def compare(self, node: ast.Compare) -> Val:
        left = self.expr(node.left)
        if len(node.ops) == 1:
            right = self.expr(node.comparators[0])
            return self.tmp(BOOL, self.compare_pair(node.ops[0], left, right, node), bits=1)
        res = self.new("c")
        self.emit(f"int {res} = 0;")
        depth = 0
        try:
            for i, (op, cmpn) in enumerate(zip(node.ops, node.comparators)):
                try:
                    right = self.expr(cmpn)
                    e = self.compare_pair(op, left, right, node)
                except _Returned:
                    if i == 0:
                        raise  # the first comparison always runs
                    break      # later comparators are only evaluated while the chain holds
                if i == len(node.ops) - 1:
                    self.emit(f"{res} = {e};")
                else:
                    self.emit(f"if ({e}) {{")
                    self.ind += 1
                    depth += 1
                left = right
        finally:
            for _ in range(depth):
                self.ind -= 1
                self.emit("}")
        return Val(BOOL, res, bits=1)

    # ------------------------------------------------------------------ boolean operators / conditionals
    def boolop(self, node: ast.BoolOp) -> Val:
        # result type = join of the operand types; found by a dry translation of the operands
        tys = []
        for v in node.values:
            tys.append(self.dry_type(v))
        if any(t not in NUMERIC for t in tys):
            if all(t == tys[0] for t in tys) and tys[0] in (GPULIST,):
                self.fail(node, "and/or over lists is not supported")
            self.fail(node, f"and/or over {tys}")
        rty = None
        for t in tys:
            rty = _join(rty, t)
        res = self.new("o")
        self.emit(f"{CTYPE[rty]} {res};")
        is_and = isinstance(node.op, ast.And)
        depth = 0
        bits = 0
        try:
            for i, vn in enumerate(node.values):
                try:
                    v = self.expr(vn)
                except _Returned:
                    if i == 0:
                        raise  # the first operand always runs: the whole expression always raises
                    break      # a later operand raises only when evaluation gets that far: code after the expression stays reachable
                bits = max(bits, v.bits)
                self.emit(f"{res} = {self.convert(v, rty)};")
                if i < len(node.values) - 1:
                    cond = self.truth(Val(rty, res))
                    self.emit(f"if ({cond if is_and else '!' + cond}) {{")
                    self.ind += 1
                    depth += 1
        finally:
            for _ in range(depth):
                self.ind -= 1
                self.emit("}")
        return Val(rty, res, bits=bits if rty in (INT, BOOL) else 64)

    def dry_type(self, node: ast.AST) -> str:
        """Static type of an expression without emitting code (run the translator on a scratch buffer)."""
        saved = (self.lines, self.ind, self.ntmp, dict(self.seen_types), dict(self.seen_bits), set(self.need_flag),
                 dict(self.var_types), dict(self.var_bits), set(self.defined), dict(self.static_vars))
        self.lines = []
        try:
            v = self.expr(node)
            return v.ty
        except _Returned:
            return BOOL  # the expression always raises; any numeric type will do
        finally:
            (self.lines, self.ind, self.ntmp, self.seen_types, self.seen_bits, self.need_flag, self.var_types,
             self.var_bits, self.defined, self.static_vars) = saved

    def ifexp(self, node: ast.IfExp) -> Val:
        ta, tb = self.dry_type(node.body), self.dry_type(node.orelse)
        if ta not in NUMERIC or tb not in NUMERIC:
            if ta == tb and ta in (GPULIST, ILIST, FLIST, GPU):
                rty = ta
            else:
                self.fail(node, f"conditional expression over {ta} / {tb}")
        else:
            rty = _join(ta, tb)
        c = self.expr(node.test)
        res = self.new("x")
        self.emit(f"{CTYPE[rty]} {res};")
        self.emit(f"if ({self.truth(c)}) {{")
        self.ind += 1
        bits = 0
        try:
            v = self.expr(node.body)
            bits = max(bits, v.bits)
            self.emit(f"{res} = {self.convert(v, rty)};")
        except _Returned:
            pass
        self.ind -= 1
        self.emit("} else {")
        self.ind += 1
        try:
            v = self.expr(node.orelse)
            bits = max(bits, v.bits)
            self.emit(f"{res} = {self.convert(v, rty)};")
        except _Returned:
            pass
        self.ind -= 1
        self.emit("}")
        return Val(rty, res, bits=bits if rty in (INT, BOOL) else 64)

    # ------------------------------------------------------------------ iteration
    def iterate(self, it_node: ast.AST, target: ast.AST, body_fn, node):
        """Emit a loop over `it_node`, binding `target` for each element; body_fn() emits the body.
        Bindings go to the innermost scope when inside a comprehension, else to real variables."""
        enum = False
        if isinstance(it_node, ast.Call) and isinstance(it_node.func, ast.Name) and it_node.func.id == "enumerate" \
                and not self.shadowed("enumerate"):
            if len(it_node.args) != 1 or it_node.keywords:
                self.fail(node, "enumerate() with a start value is not supported")
            enum = True
            it_node = it_node.args[0]
            if not (isinstance(target, (ast.Tuple, ast.List)) and len(target.elts) == 2):
                self.fail(node, "enumerate() must be unpacked into two names")
            idx_t, target = target.elts
        if isinstance(it_node, ast.Call) and isinstance(it_node.func, ast.Name) and it_node.func.id == "range" \
                and not self.shadowed("range"):
            args = [self.as_int(self.expr(a)) for a in it_node.args]
            if it_node.keywords or not 1 <= len(args) <= 3:
                self.fail(node, "range() arguments")
            lo, hi, step = "0", None, "1"
            if len(args) == 1:
                hi = args[0].code
                bits = args[0].bits
            else:
                lo, hi = args[0].code, args[1].code
                bits = max(args[0].bits, args[1].bits)
            if len(args) == 3:
                step = args[2].code
                self.emit(f"if ({step} == 0) {self.exc()}  // ValueError: range() arg 3 must not be zero")
            i = self.new("i")
            self.uses_budget = True
            self.emit("#pragma unroll 1")
            self.emit(f"for (ks_int {i} = {lo}; ({step}) > 0 ? {i} < {hi} : {i} > {hi}; {i} += {step}) {{")
            self.ind += 1
            self.emit("if (--budget < 0) KS_RETURN(KS_RET_UNSUPPORTED);")
            elem = Val(INT, i, bits=min(64, bits + 1))
            cnt = Val(INT, f"(({i} - ({lo})) / ({step}))", bits=min(64, bits + 1))
        else:
            src = self.expr(it_node)
            if src.ty == GPUS:
                i = self.new("g")
                self.emit("#pragma unroll 1")
                self.emit(f"for (int {i} = 0; {i} < ngpus; ++{i}) {{")
                self.ind += 1
                elem = Val(GPU, i)
                cnt = Val(INT, f"(ks_int){i}", bits=4)
            elif src.ty == GPULIST:
                i = self.new("k")
                self.emit("#pragma unroll 1")
                self.emit(f"for (int {i} = 0; {i} < {src.code}.n; ++{i}) {{")
                self.ind += 1
                elem = self.tmp(GPU, f"((int){src.code}.g[{i}] & 7)")
                cnt = Val(INT, f"(ks_int){i}", bits=4)
            elif src.ty in (ILIST, FLIST):
                i = self.new("k")
                self.emit("#pragma unroll 1")
                self.emit(f"for (int {i} = 0; {i} < {src.code}.n; ++{i}) {{")
                self.ind += 1
                elem = self.tmp(INT if src.ty == ILIST else FLOAT, f"{src.code}.v[{i}]")
                cnt = Val(INT, f"(ks_int){i}", bits=6)
            elif src.ty == STATIC:
                if self.is_set(src) or self.is_dict(src):
                    self.fail(node, "iteration over a set / dict literal")
                # unrolled
                for k, item in enumerate(src.items):
                    self.emit("{")
                    self.ind += 1
                    if enum:
                        self.bind(idx_t, self.const_val(k, node), node)
                    try:
                        self.bind(target, item, node)
                        # break/continue inside an unrolled literal loop are not representable
                        body_fn(unrolled=True)
                    except _Returned:
                        pass  # conservative: what follows is translated as reachable
                    finally:
                        self.ind -= 1
                        self.emit("}")
                return
            else:
                if src.ty in NUMERIC or src.ty in (NONE, GPU, POD, NODE):
                    self.raise_exc()  # TypeError: not iterable
                    raise _Returned()
                self.fail(node, f"iteration over {src.ty}")
        try:
            if enum:
                self.bind(idx_t, self.tmp(INT, cnt.code, bits=cnt.bits), node)
            self.bind(target, elem, node)
            body_fn(unrolled=False)
        except _Returned:
            pass  # the body raises, but the loop may run zero times: code after it stays reachable
        finally:
            self.ind -= 1
            self.emit("}")

    def shadowed(self, name: str) -> bool:
        return name in self.assign_count or any(name in sc for sc in self.scopes)

    def bind(self, target: ast.AST, v: Val, node):
        if isinstance(target, ast.Name):
            if self.scopes:
                self.scopes[-1][target.id] = v
            else:
                self.store(target.id, v, node)
            return
        if isinstance(target, (ast.Tuple, ast.List)) and v.ty == STATIC and len(v.items) == len(target.elts):
            for t, it in zip(target.elts, v.items):
                self.bind(t, it, node)
            return
        self.fail(node, "unsupported assignment target")

    # ------------------------------------------------------------------ comprehensions
    def comprehension_loop(self, comp_node, elt_fn, node):
        """Run elt_fn() for every element produced by a (single- or multi-generator) comprehension."""
        gens = comp_node.generators
        self.scopes.append({})

        def level(k):
            if k == len(gens):
                elt_fn()
                return
            g = gens[k]
            if g.is_async:
                self.fail(node, "async comprehension")

            def body(unrolled=False):
                depth = 0
                for cond in g.ifs:
                    c = self.expr(cond)
                    self.emit(f"if ({self.truth(c)}) {{")
                    self.ind += 1
                    depth += 1
                try:
                    level(k + 1)
                finally:
                    for _ in range(depth):
                        self.ind -= 1
                        self.emit("}")
            try:
                self.iterate(g.iter, g.target, body, node)
            except _Returned:
                # the iterable (or a filter) always raises: nothing after it in the comprehension is reachable
                raise
        try:
            level(0)
        finally:
            self.scopes.pop()

    def listcomp(self, node) -> Val:
        ety = self.comp_elt_type(node)
        if ety == GPU:
            name = self.new("L")
            self.emit(f"KsGpuList {name}; {name}.n = 0;")

            def add():
                v = self.expr(node.elt)
                self.emit(f"if ({name}.n >= KS_MAX_GPUS_PER_NODE) KS_RETURN(KS_RET_UNSUPPORTED);")
                self.emit(f"{name}.g[{name}.n++] = (signed char){v.code};")
            self.comprehension_loop(node, add, node)
            return Val(GPULIST, name)
        if ety in (INT, BOOL, FLOAT):
            lty = FLIST if ety == FLOAT else ILIST
            name = self.new("L")
            self.emit(f"{CTYPE[lty]} {name}; {name}.n = 0;")

            def add():
                v = self.expr(node.elt)
                self.emit(f"if ({name}.n >= KS_LIST_CAP) KS_RETURN(KS_RET_UNSUPPORTED);")
                code = v.code if lty == FLIST else self.as_int(v).code
                self.emit(f"{name}.v[{name}.n++] = {code};")
            self.comprehension_loop(node, add, node)
            return Val(lty, name)
        self.fail(node, f"list of {ety} elements is not supported")

    def comp_elt_type(self, node) -> str:
        saved = (self.lines, self.ind, self.ntmp, dict(self.seen_types), dict(self.seen_bits), set(self.need_flag),
                 dict(self.var_types), dict(self.var_bits), set(self.defined), dict(self.static_vars))
        self.lines = []
        out = []

        def probe():
            out.append(self.expr(node.elt).ty)
        try:
            self.comprehension_loop(node, probe, node)
        except _Returned:
            pass
        finally:
            (self.lines, self.ind, self.ntmp, self.seen_types, self.seen_bits, self.need_flag, self.var_types,
             self.var_bits, self.defined, self.static_vars) = saved
        if not out:
            return INT
        ty = None
        for t in out:
            ty = t if ty is None else (_join(ty, t) if t in NUMERIC and ty in NUMERIC else ty)
        return ty

    # ------------------------------------------------------------------ reductions over iterables
    def each_element(self, arg: ast.AST, fn, node):
        """fn(Val) for each element of a generator expression, comprehension or list-valued expression."""
        if isinstance(arg, (ast.GeneratorExp, ast.ListComp)):
            self.comprehension_loop(arg, lambda: fn(self.expr(arg.elt)), node)
            return
        src = self.expr(arg)
        if src.ty == STATIC:
            if self.is_set(src) or self.is_dict(src):
                self.fail(node, "iteration over a set / dict literal")
            for it in src.items:
                fn(it)
            return
        if src.ty == GPUS:
            src = self.gpus_to_list()
        if src.ty == GPULIST:
            i = self.new("k")
            self.emit("#pragma unroll 1")
            self.emit(f"for (int {i} = 0; {i} < {src.code}.n; ++{i}) {{")
            self.ind += 1
            try:
                fn(self.tmp(GPU, f"((int){src.code}.g[{i}] & 7)"))
            except _Returned:
                pass  # the list may be empty: code after the loop stays reachable
            finally:
                self.ind -= 1
                self.emit("}")
            return
        if src.ty in (ILIST, FLIST):
            i = self.new("k")
            self.emit("#pragma unroll 1")
            self.emit(f"for (int {i} = 0; {i} < {src.code}.n; ++{i}) {{")
            self.ind += 1
            try:
                fn(self.tmp(INT if src.ty == ILIST else FLOAT, f"{src.code}.v[{i}]"))
            except _Returned:
                pass
            finally:
                self.ind -= 1
                self.emit("}")
            return
        if src.ty in NUMERIC or src.ty in (NONE, GPU, POD, NODE):
            self.raise_exc()  # TypeError: object is not iterable
            raise _Returned()
        self.fail(node, f"iteration over {src.ty}")

    def call_sum(self, node: ast.Call) -> Val:
        if len(node.args) not in (1, 2) or node.keywords:
            self.fail(node, "sum() arguments")
        s = self.new("s")
        self.emit(f"KsSum {s}; ks_sum_init(&{s});")
        if len(node.args) == 2:
            st = self.expr(node.args[1])
            if st.ty in (INT, BOOL):
                self.emit(f"{s}.i = {self.as_int(st).code};")
            elif st.ty == FLOAT:
                self.emit(f"{s}.f = {st.code}; {s}.mode = 1;")
            else:
                self.fail(node, "sum() start value")
        tys = set()

        def add(v: Val):
            tys.add(v.ty)
            if v.ty in (INT, BOOL):
                self.emit(f"ks_sum_add_i(&{s}, {self.as_int(v).code}, &st);")
                self.check_st()
            elif v.ty == FLOAT:
                self.emit(f"ks_sum_add_f(&{s}, {v.code});")
            elif v.ty == NUM:
                self.emit(f"ks_sum_add_n(&{s}, {v.code}, &st);")
                self.check_st()
            else:
                self.raise_exc()  # TypeError: unsupported operand type(s) for +
        self.each_element(node.args[0], add, node)
        if tys <= {INT, BOOL} and (len(node.args) == 1 or st.ty in (INT, BOOL)):
            return self.tmp(INT, f"{s}.i", bits=64)
        return self.tmp(NUM, f"ks_sum_result(&{s})")

    def call_minmax(self, node: ast.Call, is_max: bool) -> Val:
        key = None
        for kw in node.keywords:
            if kw.arg == "key":
                key = kw.value
            else:
                self.fail(node, f"{'max' if is_max else 'min'}() keyword {kw.arg}")
        if not node.args:
            self.raise_exc()  # TypeError
            raise _Returned()
        cmpop = ast.Gt() if is_max else ast.Lt()
        if len(node.args) >= 2:
            vals = [self.expr(a) for a in node.args]
            if key is not None:
                self.fail(node, "max/min of several arguments with key=")
            if any(v.ty not in NUMERIC for v in vals):
                self.raise_exc()  # TypeError
                raise _Returned()
            rty = None
            for v in vals:
                rty = _join(rty, v.ty)
            if all(v.const is not None for v in vals) and rty != NUM:
                best = vals[0]
                for v in vals[1:]:
                    if (v.const > best.const) if is_max else (v.const < best.const):
                        best = v
                return best
            res = self.new("m")
            bits = max(v.bits for v in vals)
            self.emit(f"{CTYPE[rty]} {res} = {self.convert(vals[0], rty)};")
            for v in vals[1:]:
                c = self.compare_pair(cmpop, v, Val(rty, res, bits=bits), node)
                self.emit(f"if ({c}) {res} = {self.convert(v, rty)};")
            return Val(rty, res, bits=bits if rty in (INT, BOOL) else 64)
        # single iterable argument
        found = self.new("f")
        self.emit(f"int {found} = 0;")
        state = {}

        def visit(v: Val):
            if key is not None:
                kv = self.apply_lambda(key, [v], node)
            else:
                kv = v
            if kv.ty not in NUMERIC:
                self.raise_exc()
                raise _Returned()
            if "res" not in state:
                state["kty"] = kv.ty if kv.ty != BOOL else INT
                state["ety"] = v.ty if v.ty != BOOL else INT
                state["res"] = self.new("m")
                state["kres"] = self.new("mk")
                state["bits"] = v.bits
                # declarations must precede the loop: patch them in at the recorded position
                zero = {INT: "0", FLOAT: "0.0", NUM: "ks_num_i(0)", GPU: "0"}
                decl = (f"{CTYPE[state['ety']]} {state['res']} = {zero[state['ety']]}; "
                        f"{CTYPE[state['kty']]} {state['kres']} = {zero[state['kty']]};")
                self.lines.insert(state["decl_at"], "  " * state["decl_ind"] + decl)
            state["bits"] = max(state["bits"], v.bits)
            kcur = Val(state["kty"], state["kres"], bits=kv.bits)
            if kv.ty != state["kty"] and not (kv.ty == BOOL and state["kty"] == INT):
                self.fail(node, "max/min over elements of varying type")
            c = self.compare_pair(cmpop, kv, kcur, node)
            self.emit(f"if (!{found} || {c}) {{ {state['res']} = {self.convert(v, state['ety'])}; "
                      f"{state['kres']} = {self.convert(kv, state['kty'])}; {found} = 1; }}")
        state["decl_at"] = len(self.lines)
        state["decl_ind"] = self.ind
        self.each_element(node.args[0], visit, node)
        if "res" not in state:
            self.raise_exc()  # empty literal: ValueError
            raise _Returned()
        self.emit(f"if (!{found}) {self.exc()}  // ValueError: empty sequence")
        return Val(state["ety"], state["res"], bits=state["bits"] if state["ety"] in (INT, BOOL) else 64)

    def apply_lambda(self, fn_node: ast.AST, args: List[Val], node) -> Val:
        lam = fn_node
        if isinstance(fn_node, ast.Name) and fn_node.id in self.static_vars and self.static_vars[fn_node.id].ty == LAMBDA:
            lam = self.static_vars[fn_node.id].node
        elif isinstance(fn_node, ast.Name) and not self.shadowed(fn_node.id) and isinstance(self.consts.get(fn_node.id), ast.Lambda):
            lam = self.consts[fn_node.id]
        if not isinstance(lam, ast.Lambda):
            self.fail(node, "only lambda expressions can be used as key functions")
        params = [a.arg for a in lam.args.args]
        if len(params) != len(args) or lam.args.vararg or lam.args.kwarg or lam.args.defaults:
            self.fail(node, "lambda signature")
        self.scopes.append(dict(zip(params, args)))
        try:
            return self.expr(lam.body)
        finally:
            self.scopes.pop()

    def call_sorted(self, node: ast.Call) -> Val:
        key = None
        reverse = None
        for kw in node.keywords:
            if kw.arg == "key":
                key = kw.value
            elif kw.arg == "reverse":
                reverse = self.expr(kw.value)
            else:
                self.fail(node, f"sorted() keyword {kw.arg}")
        if len(node.args) != 1:
            self.fail(node, "sorted() arguments")
        arg = node.args[0]
        if isinstance(arg, ast.GeneratorExp):
            lst = self.listcomp(arg)
        else:
            lst = self.expr(arg)
        if lst.ty == GPUS:
            lst = self.gpus_to_list()
        if lst.ty == STATIC:
            self.fail(node, "sorted() of a literal list")
        if lst.ty not in (GPULIST, ILIST, FLIST):
            self.raise_exc()
            raise _Returned()
        if lst.ty == GPULIST and key is None:
            self.raise_exc()  # TypeError: '<' not supported between GPU objects
            raise _Returned()
        out = self.new("L")
        self.emit(f"{CTYPE[lst.ty]} {out} = {lst.code};")
        arr = "g" if lst.ty == GPULIST else "v"
        ety = {GPULIST: GPU, ILIST: INT, FLIST: FLOAT}[lst.ty]
        cap = "KS_MAX_GPUS_PER_NODE" if lst.ty == GPULIST else "KS_LIST_CAP"
        # keys
        kname = self.new("K")
        i = self.new("k")
        if key is not None:
            kty = self.dry_type_lambda(key, ety, node)
            if kty == BOOL:
                kty = INT
            if kty not in (INT, FLOAT):
                self.fail(node, "sort key must be int or float")
            self.emit(f"{CTYPE[kty]} {kname}[{cap}];")
            self.emit("#pragma unroll 1")
            self.emit(f"for (int {i} = 0; {i} < {out}.n; ++{i}) {{")
            self.ind += 1
            try:
                el = self.tmp(ety, f"((int){out}.{arr}[{i}] & 7)" if ety == GPU else f"{out}.{arr}[{i}]")
                kv = self.apply_lambda(key, [el], node)
                self.emit(f"{kname}[{i}] = {self.convert(kv, kty) if kv.ty != BOOL else self.as_int(kv).code};")
            except _Returned:
                pass  # the key raises for every element, but the list may be empty
            finally:
                self.ind -= 1
                self.emit("}")
        else:
            kty = ety
        # stable insertion sort (list.sort is stable; reverse=True keeps equal elements in original order)
        a, b = self.new("a"), self.new("b")
        rev = self.truth(reverse) if reverse is not None else "0"
        kt, et = CTYPE[kty], CTYPE[ety] if lst.ty != GPULIST else "signed char"
        self.emit("#pragma unroll 1")
        self.emit(f"for (int {a} = 1; {a} < {out}.n; ++{a}) {{")
        self.ind += 1
        if key is not None:
            self.emit(f"const {kt} kk = {kname}[{a}]; const {et} ee = {out}.{arr}[{a}]; int {b} = {a} - 1;")
            self.emit(f"while ({b} >= 0 && ({rev} ? ({kname}[{b}] < kk) : (kk < {kname}[{b}]))) {{ "
                      f"{kname}[{b} + 1] = {kname}[{b}]; {out}.{arr}[{b} + 1] = {out}.{arr}[{b}]; --{b}; }}")
            self.emit(f"{kname}[{b} + 1] = kk; {out}.{arr}[{b} + 1] = ee;")
        else:
            self.emit(f"const {et} ee = {out}.{arr}[{a}]; int {b} = {a} - 1;")
            self.emit(f"while ({b} >= 0 && ({rev} ? ({out}.{arr}[{b}] < ee) : (ee < {out}.{arr}[{b}]))) {{ "
                      f"{out}.{arr}[{b} + 1] = {out}.{arr}[{b}]; --{b}; }}")
            self.emit(f"{out}.{arr}[{b} + 1] = ee;")
        self.ind -= 1
        self.emit("}")
        return Val(lst.ty, out)

    def return_max_int_of_float(self, c) -> bool:
        """`return max(<int literal>, int(<float expression>))` -- how the reference's template ends every policy -- is
        finished by one shared runtime routine (ks_rt_ret_max_f2i) instead of ~25 inlined instructions per policy:
        what is resident per policy is what the instruction cache has to hold per warp."""
        if not (isinstance(c, ast.Call) and isinstance(c.func, ast.Name) and c.func.id == "max" and not c.keywords
                and len(c.args) == 2 and isinstance(c.args[0], ast.Constant) and type(c.args[0].value) is int
                and abs(c.args[0].value) < 2 ** 62 and isinstance(c.args[1], ast.Call)
                and isinstance(c.args[1].func, ast.Name) and c.args[1].func.id == "int" and not c.args[1].keywords
                and len(c.args[1].args) == 1 and not self.shadowed("max") and not self.shadowed("int")) or self.try_labels:
            return False
        saved = (list(self.lines), self.ind, self.ntmp, dict(self.seen_types), dict(self.seen_bits), set(self.need_flag),
                 dict(self.var_types), dict(self.var_bits), set(self.defined), dict(self.static_vars))
        x = self.expr(c.args[1].args[0])
        if x.ty != FLOAT:   # int(int), int(bool), int(num): the general path
            (self.lines, self.ind, self.ntmp, self.seen_types, self.seen_bits, self.need_flag, self.var_types,
             self.var_bits, self.defined, self.static_vars) = saved
            return False
        self.emit(f"return ks_rt_ret_max_f2i({x.code}, {c.args[0].value}LL, st);")
        return True

    def dry_type_lambda(self, key, ety, node) -> str:
        saved = (self.lines, self.ind, self.ntmp, dict(self.seen_types), dict(self.seen_bits), set(self.need_flag),
                 dict(self.var_types), dict(self.var_bits), set(self.defined), dict(self.static_vars))
        self.lines = []
        try:
            return self.apply_lambda(key, [Val(ety, "0")], node).ty
        except _Returned:
            return INT
        finally:
            (self.lines, self.ind, self.ntmp, self.seen_types, self.seen_bits, self.need_flag, self.var_types,
             self.var_bits, self.defined, self.static_vars) = saved

    # ------------------------------------------------------------------ calls
    def call(self, node: ast.Call) -> Val:
        f = node.func
        if isinstance(f, ast.Attribute) and isinstance(f.value, ast.Name) and f.value.id in ("math", "operator") \
                and not self.shadowed(f.value.id):
            return self.call_module(f.value.id, f.attr, node)
        if isinstance(f, ast.Attribute) and f.attr == "append" and isinstance(f.value, ast.Name) and f.value.id in self.appended \
                and f.value.id in self.assign_count and not self.scopes:
            return self.list_append(f.value.id, node)
        if isinstance(f, ast.Attribute) and f.attr == "extend" and isinstance(f.value, ast.Name) and f.value.id in self.appended \
                and f.value.id in self.assign_count and not self.scopes and len(node.args) == 1 and not node.keywords \
                and isinstance(node.args[0], (ast.List, ast.Tuple)):
            # xs.extend([a, b]) with a literal: the elements are evaluated first, then appended one by one
            for el in node.args[0].elts:
                self.list_append(f.value.id, ast.copy_location(ast.Call(func=f, args=[el], keywords=[]), node))
            return Val(NONE)
        if isinstance(f, ast.Attribute):
            self.fail(node, f"method call .{f.attr}()")
        if isinstance(f, ast.Lambda):
            return self.apply_lambda(f, [self.expr(a) for a in node.args], node)
        if not isinstance(f, ast.Name):
            self.fail(node, "call of a computed function")
        name = f.id
        if name in self.static_vars and self.static_vars[name].ty == LAMBDA:
            return self.apply_lambda(f, [self.expr(a) for a in node.args], node)
        if not self.shadowed(name) and isinstance(self.consts.get(name), ast.Lambda):
            if node.keywords:
                self.fail(node, "keyword arguments to a helper function")
            return self.apply_lambda(self.consts[name], [self.expr(a) for a in node.args], node)
        if self.shadowed(name):
            self.fail(node, f"call of local name {name!r}")
        if name not in ALLOWED_BUILTINS:
            # not in the sandbox's builtins (safe_execution.py:98-124): NameError at run time
            self.raise_exc()
            raise _Returned()
        if name == "sum":
            return self.call_sum(node)
        if name in ("max", "min"):
            return self.call_minmax(node, name == "max")
        if name == "sorted":
            return self.call_sorted(node)
        if name == "len":
            if len(node.args) != 1 or node.keywords:
                self.raise_exc()
                raise _Returned()
            v = self.expr(node.args[0])
            if v.ty == GPUS:
                return self.tmp(INT, "(ks_int)ngpus", bits=4)
            if v.ty in (GPULIST, ILIST, FLIST):
                return self.tmp(INT, f"(ks_int){v.code}.n", bits=6)
            if self.is_set(v):
                if any(it.const is None for it in v.items):
                    self.fail(node, "len() of a set literal with computed elements")
                return self.const_val(len({it.const for it in v.items}), node)
            if v.ty == STATIC:
                return self.const_val(len(v.items), node)
            if v.ty == STR:
                return self.const_val(len(v.const), node)
            self.raise_exc()  # TypeError: object of type 'int' has no len()
            raise _Returned()
        if name in ("range", "enumerate"):
            self.fail(node, f"{name}() outside a for loop / comprehension")
        if name == "str":
            self.fail(node, "str() is not supported")
        if node.keywords:
            self.fail(node, f"{name}() keywords")
        args = [self.expr(a) for a in node.args]
        if name == "abs":
            if len(args) != 1:
                self.raise_exc()
                raise _Returned()
            v = args[0]
            if v.ty in (INT, BOOL):
                v = self.as_int(v)
                if v.bits <= 62:
                    return self.tmp(INT, f"({v.code} < 0 ? -{v.code} : {v.code})", bits=v.bits)
                r = self.tmp(INT, f"ks_iabs({v.code}, &st)")
                self.check_st()
                return r
            if v.ty == FLOAT:
                return self.tmp(FLOAT, f"fabs({v.code})")
            if v.ty == NUM:
                r = self.tmp(NUM, f"ks_nabs({v.code}, &st)")
                self.check_st()
                return r
            self.raise_exc()
            raise _Returned()
        if name == "int":
            if not args:
                return self.const_val(0, node)
            if len(args) != 1:
                self.fail(node, "int() with a base")
            v = args[0]
            if v.ty in (INT, BOOL):
                return self.as_int(v)
            if v.ty == FLOAT:
                r = self.tmp(INT, f"ks_f2i({v.code}, &st)", bits=64)
                self.check_st()
                return r
            if v.ty == NUM:
                r = self.tmp(INT, f"ks_n2i({v.code}, &st)", bits=64)
                self.check_st()
                return r
            if v.ty == STR:
                self.fail(node, "int() of a string")
            self.raise_exc()
            raise _Returned()
        if name == "float":
            if not args:
                return self.const_val(0.0, node)
            v = args[0]
            if v.ty == STR:
                try:
                    return self.const_val(float(v.const), node)
                except ValueError:
                    self.raise_exc()
                    raise _Returned()
            if v.ty in NUMERIC:
                return self.tmp(FLOAT, self.as_double(v))
            self.raise_exc()
            raise _Returned()
        if name == "bool":
            if not args:
                return self.const_val(False, node)
            return self.tmp(BOOL, self.truth(args[0]), bits=1)
        if name == "round":
            if len(args) == 1 or (len(args) == 2 and args[1].ty == NONE):
                v = args[0]
                if v.ty in (INT, BOOL):
                    return self.as_int(v)
                if v.ty == FLOAT:
                    r = self.tmp(INT, f"ks_fround({v.code}, &st)", bits=64)
                    self.check_st()
                    return r
                if v.ty == NUM:
                    r = self.tmp(INT, f"ks_nround({v.code}, &st)", bits=64)
                    self.check_st()
                    return r
                self.raise_exc()
                raise _Returned()
            if len(args) == 2 and args[1].ty in (INT, BOOL):
                v, nd = args[0], self.as_int(args[1])
                if v.ty in (INT, BOOL):
                    if nd.const is not None and int(nd.const) >= 0:
                        return self.as_int(v)      # round(int, n >= 0) is the int itself
                    self.fail(node, "round(int, negative or dynamic ndigits)")
                if v.ty == FLOAT:
                    if nd.const is not None and not 0 <= int(nd.const) <= 15:
                        self.fail(node, "round(x, ndigits) with ndigits outside 0..15")
                    r = self.tmp(FLOAT, f"ks_fround_n({v.code}, {nd.code}, &st)")   # a float, not an int, in CPython too
                    self.check_st()
                    return r
                self.fail(node, "round(x, ndigits) of a dynamically typed number")
            self.raise_exc()   # TypeError: ndigits must be an integer
            raise _Returned()
        self.fail(node, f"builtin {name}()")

    def list_append(self, name: str, node: ast.Call) -> Val:
        """xs.append(v) on a local list that started as `xs = []` (bounded: KS_LIST_CAP elements, 8 for GPU lists)."""
        if len(node.args) != 1 or node.keywords:
            self.raise_exc()  # TypeError: append() takes exactly one argument
            raise _Returned()
        cur = self.lookup(name, node)            # UnboundLocalError when appended to before `xs = []`
        v = self.expr(node.args[0])
        want = {INT: ILIST, BOOL: ILIST, FLOAT: FLIST, GPU: GPULIST}.get(v.ty)
        if want is None:
            self.fail(node, f"append of a {v.ty} value")
        if cur.ty not in (ELIST, want):
            self.fail(node, f"list {name!r} holds elements of more than one type")
        self.seen_types[name] = _join(self.seen_types.get(name), want)
        self.var_types[name] = _join(self.var_types.get(name), want)
        if cur.ty == ELIST:
            return Val(NONE)   # element type learnt in this pass; the next pass emits the code
        cap = "KS_MAX_GPUS_PER_NODE" if want == GPULIST else "KS_LIST_CAP"
        self.emit(f"if (v_{name}.n >= {cap}) KS_RETURN(KS_RET_UNSUPPORTED);")
        if want == GPULIST:
            self.emit(f"v_{name}.g[v_{name}.n++] = (signed char){v.code};")
        else:
            code = v.code if want == FLIST else self.as_int(v).code
            self.emit(f"v_{name}.v[v_{name}.n++] = {code};")
        return Val(NONE)

    def call_module(self, mod: str, attr: str, node: ast.Call) -> Val:
        if node.keywords:
            self.fail(node, "keywords in math/operator call")
        if mod == "operator":
            if attr not in OPERATOR_FUNCS:
                self.raise_exc()  # AttributeError
                raise _Returned()
            if len(node.args) != 2:
                self.raise_exc()
                raise _Returned()
            a, b = self.expr(node.args[0]), self.expr(node.args[1])
            return self.arith(OPERATOR_FUNCS[attr](), a, b, node)
        if attr not in MATH_FUNCS:
            self.raise_exc()  # AttributeError: the sandbox math object has only 7 functions
            raise _Returned()
        args = [self.expr(a) for a in node.args]
        if any(a.ty not in NUMERIC for a in args):
            self.raise_exc()
            raise _Returned()
        if attr == "pow":
            if len(args) != 2:
                self.raise_exc()
                raise _Returned()
            r = self.tmp(FLOAT, f"ks_mpow({self.as_double(args[0])}, {self.as_double(args[1])}, &st)")
            self.check_st()
            return r
        if len(args) != 1:
            if attr == "log" and len(args) == 2:
                self.fail(node, "math.log with a base")
            self.raise_exc()
            raise _Returned()
        if attr == "log" and args[0].ty in (INT, BOOL, NUM) and args[0].bits > 53:
            self.fail(node, "math.log of a possibly huge int")
        r = self.tmp(FLOAT, f"ks_m{attr}({self.as_double(args[0])}, &st)")
        self.check_st()
        return r

    # ------------------------------------------------------------------ subscripts
    def subscript(self, node: ast.Subscript) -> Val:
        base = self.expr(node.value)
        sl = node.slice
        if isinstance(sl, ast.Slice):
            if sl.step is not None:
                self.fail(node, "slice step")
            if base.ty == GPUS:
                base = self.gpus_to_list()
            if self.is_dict(base) or self.is_set(base):
                self.raise_exc()  # TypeError
                raise _Returned()
            if base.ty == STATIC:
                lo = self.expr(sl.lower).const if sl.lower is not None else None
                hi = self.expr(sl.upper).const if sl.upper is not None else None
                if (sl.lower is not None and lo is None) or (sl.upper is not None and hi is None):
                    self.fail(node, "dynamic slice of a literal list")
                return Val(STATIC, items=base.items[lo:hi])
            if base.ty not in (GPULIST, ILIST, FLIST):
                self.raise_exc()
                raise _Returned()
            arr = "g" if base.ty == GPULIST else "v"
            lo = self.as_int(self.expr(sl.lower)).code if sl.lower is not None else "0"
            hi = self.as_int(self.expr(sl.upper)).code if sl.upper is not None else f"(ks_int){base.code}.n"
            out = self.new("L")
            a, b, q = self.new("a"), self.new("b"), self.new("q")
            self.emit(f"{CTYPE[base.ty]} {out}; const int {a} = ks_slice_bound({lo}, {base.code}.n), "
                      f"{b} = ks_slice_bound({hi}, {base.code}.n);")
            self.emit(f"{out}.n = 0;")
            self.emit("#pragma unroll 1")
            self.emit(f"for (int {q} = {a}; {q} < {b}; ++{q}) {out}.{arr}[{out}.n++] = {base.code}.{arr}[{q}];")
            return Val(base.ty, out)
        idx = self.expr(sl)
        if self.is_dict(base):
            if idx.const is None or idx.ty not in (STR, INT):
                self.fail(node, "dict lookup with a non-constant key")
            if idx.const not in base.const[1]:
                self.raise_exc()  # KeyError
                raise _Returned()
            return base.items[base.const[1].index(idx.const)]
        if self.is_set(base):
            self.raise_exc()  # TypeError: 'set' object is not subscriptable
            raise _Returned()
        if idx.ty not in (INT, BOOL):
            if base.ty in (GPUS, GPULIST, ILIST, FLIST, STATIC):
                self.raise_exc()  # TypeError: list indices must be integers
                raise _Returned()
            self.fail(node, "subscript")
        idx = self.as_int(idx)
        if base.ty == STATIC:
            if idx.const is None:
                tys = {it.ty for it in base.items}
                if not base.items or not tys <= set(NUMERIC):
                    self.fail(node, "dynamic index into a literal list of non-numbers")
                rty = None
                for t in tys:
                    rty = _join(rty, t)
                res = self.new("e")
                n = len(base.items)
                k = self.tmp(INT, f"({idx.code} < 0 ? {idx.code} + {n} : {idx.code})", bits=idx.bits)
                self.emit(f"{CTYPE[rty]} {res};")
                self.emit(f"switch ({k.code}) {{")
                for j, it in enumerate(base.items):
                    self.emit(f"  case {j}: {res} = {self.convert(it, rty)}; break;")
                self.emit(f"  default: {self.exc()}  // IndexError")
                self.emit("}")
                return Val(rty, res, bits=max(it.bits for it in base.items) if rty in (INT, BOOL) else 64)
            k = int(idx.const)
            if not -len(base.items) <= k < len(base.items):
                self.raise_exc()
                raise _Returned()
            return base.items[k]
        if base.ty == GPUS:
            n = "ngpus"
        elif base.ty in (GPULIST, ILIST, FLIST):
            n = f"{base.code}.n"
        else:
            self.raise_exc()  # TypeError: not subscriptable
            raise _Returned()
        k = self.new("j")
        self.emit(f"ks_int {k} = {idx.code}; if ({k} < 0) {k} += {n};")
        self.emit(f"if ({k} < 0 || {k} >= {n}) {self.exc()}  // IndexError")
        if base.ty == GPUS:
            # "& 7" is the identity on a checked index; it keeps the shared-memory load in range when the compiler
            # hoists it (ks_ns is a plain, non-volatile ld.shared) above the IndexError test
            return Val(GPU, f"((int){k} & 7)")
        if base.ty == GPULIST:
            return self.tmp(GPU, f"((int){base.code}.g[{k}] & 7)")
        return self.tmp(INT if base.ty == ILIST else FLOAT, f"{base.code}.v[{k}]")

    # ------------------------------------------------------------------ expressions
    def expr(self, node: ast.AST) -> Val:
        if isinstance(node, ast.Constant):
            k = self.lift.get(id(node))
            if k is not None:
                return Val(FLOAT, f"ks_kf(kf, {k})")
            return self.const_val(node.value, node)
        if isinstance(node, ast.Name):
            return self.lookup(node.id, node)
        if isinstance(node, ast.Attribute):
            return self.attribute(node)
        if isinstance(node, ast.BinOp):
            a = self.expr(node.left)
            b = self.expr(node.right)
            return self.arith(node.op, a, b, node)
        if isinstance(node, ast.UnaryOp):
            v = self.expr(node.operand)
            if isinstance(node.op, ast.Not):
                return self.tmp(BOOL, f"(!{self.truth(v)})", bits=1)
            if v.ty not in NUMERIC:
                self.raise_exc()
                raise _Returned()
            if isinstance(node.op, ast.UAdd):
                return self.as_int(v) if v.ty == BOOL else v
            if isinstance(node.op, ast.USub):
                if v.ty in (INT, BOOL):
                    v = self.as_int(v)
                    if v.const is not None:
                        r = self.const_val(-int(v.const), node)
                        if v.lit is not None and v.lit[1] == 1:   # `-5`: still a literal of the source
                            r.lit = (v.lit[0], -1)
                        return r
                    if v.bits <= 62:
                        return self.tmp(INT, f"-{v.code}", bits=v.bits)
                    r = self.tmp(INT, f"ks_ineg({v.code}, &st)")
                    self.check_st()
                    return r
                if v.ty == FLOAT:
                    if v.const is not None:
                        return self.const_val(-float(v.const), node)
                    return self.tmp(FLOAT, f"-{v.code}")
                r = self.tmp(NUM, f"ks_nneg({v.code}, &st)")
                self.check_st()
                return r
            if isinstance(node.op, ast.Invert):
                if v.ty in (INT, BOOL):
                    return self.tmp(INT, f"(~{self.as_int(v).code})", bits=min(64, v.bits + 1))
                self.raise_exc()
                raise _Returned()
        if isinstance(node, ast.BoolOp):
            return self.boolop(node)
        if isinstance(node, ast.Compare):
            return self.compare(node)
        if isinstance(node, ast.IfExp):
            return self.ifexp(node)
        if isinstance(node, ast.Call):
            return self.call(node)
        if isinstance(node, ast.Subscript):
            return self.subscript(node)
        if isinstance(node, (ast.List, ast.Tuple)):
            if any(isinstance(e, ast.Starred) for e in node.elts):
                self.fail(node, "starred expression")
            return Val(STATIC, items=[self.expr(e) for e in node.elts])
        if isinstance(node, ast.Set):
            # a set literal: only membership (`x in {1, 2, 4}`), len() and truth are order-free; anything that would
            # iterate it is refused (the iteration order of a set is an implementation detail)
            if any(isinstance(e, ast.Starred) for e in node.elts):
                self.fail(node, "starred expression")
            return Val(STATIC, items=[self.expr(e) for e in node.elts], const="<set>")
        if isinstance(node, ast.Dict):
            # a dict literal with constant keys (the usual `weights = {'cpu': 0.6, 'mem': 0.4}`): read by constant key
            keys = []
            for k in node.keys:
                if not (isinstance(k, ast.Constant) and isinstance(k.value, (str, int)) and not isinstance(k.value, bool)):
                    self.fail(node, "dict literal with non-constant keys")
                keys.append(k.value)
            vals = [self.expr(v) for v in node.values]
            merged = {}
            for k, v in zip(keys, vals):   # a repeated key keeps its first position and its last value, like CPython
                merged[k] = v
            return Val(STATIC, items=list(merged.values()), const=("<dict>", tuple(merged.keys())))
        if isinstance(node, ast.ListComp):
            return self.listcomp(node)
        if isinstance(node, ast.Lambda):
            return Val(LAMBDA, node=node)
        if isinstance(node, ast.GeneratorExp):
            self.fail(node, "generator expression outside sum/min/max/sorted")
        self.fail(node, f"expression {type(node).__name__}")

    # ------------------------------------------------------------------ statements
    def block(self, stmts, loop_depth=0) -> bool:
        """Translate a statement list; returns True when control can fall out of its end."""
        for s in stmts:
            try:
                if not self.stmt(s, loop_depth):
                    return False
            except _Returned:
                return False
        return True

    def stmt(self, s: ast.stmt, loop_depth: int) -> bool:
        if isinstance(s, ast.Expr):
            if isinstance(s.value, ast.Constant):
                return True
            self.expr(s.value)
            return True
        if isinstance(s, ast.Pass):
            return True
        if isinstance(s, ast.Return):
            if s.value is None:
                self.raise_exc()  # int(max(0, None)): TypeError
                return False
            if self.return_max_int_of_float(s.value):
                return False
            v = self.expr(s.value)
            # (the scheduler's wrapper int(max(0, .)) runs in the caller, outside any try of the policy: what it raises
            # -- int(inf) -- is not catchable by the policy, so ks_wrap_* may return the exception code directly)
            if v.ty in (INT, BOOL):
                self.emit(f"KS_RETURN(ks_wrap_i({self.as_int(v).code}));")
            elif v.ty == FLOAT:
                self.emit(f"KS_RETURN(ks_wrap_f({v.code}));")
            elif v.ty == NUM:
                self.emit(f"KS_RETURN(ks_wrap_n({v.code}));")
            else:
                self.emit("KS_RETURN(KS_RET_EXCEPTION);")  # max(0, <list/None/...>) in the wrapper: TypeError, not catchable
            return False
        if isinstance(s, ast.Assign):
            v = self.expr(s.value)
            for t in s.targets:
                if isinstance(t, (ast.Attribute, ast.Subscript)):
                    self.fail(s, "assignment to an attribute or subscript (policies must not mutate pod/node state)")
                if isinstance(t, (ast.Tuple, ast.List)):
                    if v.ty != STATIC or len(v.items) != len(t.elts):
                        self.fail(s, "tuple unpacking of a non-literal")
                    # materialise the right-hand side first (a, b = b, a)
                    items = [self.tmp(it.ty, it.code, bits=it.bits) if it.ty in CTYPE and it.ty in NUMERIC else it
                             for it in v.items]
                    for tt, it in zip(t.elts, items):
                        if not isinstance(tt, ast.Name):
                            self.fail(s, "nested unpacking")
                        self.store(tt.id, it, s)
                else:
                    self.store(t.id, v, s)
            return True
        if isinstance(s, ast.AnnAssign):
            if s.value is None or not isinstance(s.target, ast.Name):
                return True
            self.store(s.target.id, self.expr(s.value), s)
            return True
        if isinstance(s, ast.AugAssign):
            if not isinstance(s.target, ast.Name):
                self.fail(s, "augmented assignment to an attribute or subscript")
            name = s.target.id
            cur = self.lookup(name, s)
            if cur.ty in CTYPE and cur.ty in NUMERIC:
                cur = self.tmp(cur.ty, cur.code, bits=cur.bits)
            v = self.expr(s.value)
            self_ref = any(isinstance(n, ast.Name) and n.id == name for n in ast.walk(s.value))
            if (isinstance(s.op, (ast.Add, ast.Sub)) and cur.ty in (INT, BOOL) and v.ty in (INT, BOOL) and not self_ref
                    and not self.scopes):
                # x += e with e independent of x: the function executes at most LOOP_BUDGET * 8**3 < 2**27 such
                # updates, so |x| < 2**(max(bits of x's other assignments, bits of e) + 27): no overflow check
                grown = max(self.base_bits.get(name, 0), v.bits) + 27
                if grown <= 62:
                    r = self.tmp(INT, f"{self.as_int(cur).code} {'+' if isinstance(s.op, ast.Add) else '-'} {self.as_int(v).code}",
                                 bits=grown)
                    self.store(name, r, s, additive=True)
                    return True
            self.store(name, self.arith(s.op, cur, v, s), s)
            return True
        if isinstance(s, ast.If):
            try:
                c = self.expr(s.test)
            except _Returned:
                return False
            self.emit(f"if ({self.truth(c)}) {{")
            self.ind += 1
            before = set(self.defined)
            a_falls = self.block(s.body, loop_depth)
            after_a = set(self.defined)
            self.ind -= 1
            self.defined = set(before)
            if s.orelse:
                self.emit("} else {")
                self.ind += 1
                b_falls = self.block(s.orelse, loop_depth)
                after_b = set(self.defined)
                self.ind -= 1
            else:
                b_falls = True
                after_b = set(before)
            self.emit("}")
            if a_falls and b_falls:
                self.defined = after_a & after_b
            elif a_falls:
                self.defined = after_a
            elif b_falls:
                self.defined = after_b
            else:
                return False
            return True
        if isinstance(s, ast.For):
            if s.orelse:
                self.fail(s, "for ... else")
            before = set(self.defined)

            def body(unrolled=False):
                if unrolled and any(isinstance(n, (ast.Break, ast.Continue)) for n in ast.walk(s)):
                    self.fail(s, "break/continue in a loop over a literal list")
                self.block(s.body, loop_depth + 1)
            self.iterate(s.iter, s.target, body, s)
            self.defined = before  # the body may run zero times
            return True
        if isinstance(s, ast.While):
            if s.orelse:
                self.fail(s, "while ... else")
            before = set(self.defined)
            self.uses_budget = True
            self.emit("for (;;) {")
            self.ind += 1
            self.emit("if (--budget < 0) KS_RETURN(KS_RET_UNSUPPORTED);")
            try:
                c = self.expr(s.test)
                self.emit(f"if (!{self.truth(c)}) break;")
                self.block(s.body, loop_depth + 1)
            except _Returned:
                pass
            self.ind -= 1
            self.emit("}")
            self.defined = before
            return True
        if isinstance(s, ast.FunctionDef):
            lam = _def_as_lambda(s)
            if lam is None:
                self.fail(s, f"helper function {s.name!r} is more than a single return statement")
            self.store(s.name, Val(LAMBDA, node=lam), s)
            return True
        if isinstance(s, ast.Try):
            return self.try_stmt(s, loop_depth)
        if isinstance(s, ast.Break):
            self.emit("break;")
            return False
        if isinstance(s, ast.Continue):
            self.emit("continue;")
            return False
        if isinstance(s, (ast.Import, ast.ImportFrom)):
            self.fail(s, "import")
        if isinstance(s, ast.Assert):
            c = self.expr(s.test)
            self.emit(f"if (!{self.truth(c)}) {self.exc()}  // AssertionError")
            return True
        if isinstance(s, ast.Raise):
            self.raise_exc()
            return False
        self.fail(s, f"statement {type(s).__name__}")

    def try_stmt(self, s: ast.Try, loop_depth: int) -> bool:
        """try / except / else.  In the reference's sandbox only a bare `except:` can catch anything: the restricted
        builtins (safe_execution.py:98-124) hold no exception classes, so `except ZeroDivisionError:` raises NameError
        while the exception is being matched, and that NameError leaves the try statement (later clauses are not tried).
        A try body therefore either has a bare first handler, which catches everything the policy can raise, or a first
        handler that turns any exception into another exception -- i.e. no handler at all."""
        if s.finalbody:
            return self.try_finally(s, loop_depth)
        bare = bool(s.handlers) and s.handlers[0].type is None
        if not bare:
            # no effective handler: the body runs as if it were not wrapped (else: follows on normal completion)
            falls = self.block(s.body, loop_depth)
            if falls and s.orelse:
                falls = self.block(s.orelse, loop_depth)
            return falls
        self.ntry += 1
        k = self.ntry
        label, done, flag = f"ks_try{k}", f"ks_try{k}_ok", f"tr{k}"
        before = set(self.defined)
        if not self.try_labels:
            # outside a try body failures are sticky and only tested at the return: one that happened BEFORE this try
            # statement must not be mistaken for one of its body's (the handler would swallow it)
            self.emit("if (st) KS_RETURN(0);")
        self.emit(f"int {flag} = 0;")
        self.emit("{")
        self.ind += 1
        self.try_labels.append(label)
        try:
            body_falls = self.block(s.body, loop_depth)
        finally:
            self.try_labels.pop()
        after_body = set(self.defined)
        self.ind -= 1
        self.emit("}")
        self.emit(f"goto {done};")
        self.emit(f"{label}: {flag} = 1;")
        self.emit(f"{done}:;")
        self.emit(f"if ({flag}) {{")
        self.ind += 1
        # leaving the translated subset is not an exception the policy could have caught
        self.emit("if (st != 1) KS_RETURN(KS_RET_UNSUPPORTED);")
        self.emit("st = 0;")
        self.defined = set(before)   # the body may have stopped anywhere
        h = s.handlers[0]
        if h.name:
            self.fail(s, "except ... as name")
        h_falls = self.block(h.body, loop_depth)
        after_h = set(self.defined)
        self.ind -= 1
        if s.orelse or True:
            self.emit("} else {")
            self.ind += 1
            self.defined = set(after_body)
            e_falls = self.block(s.orelse, loop_depth) if (s.orelse and body_falls) else body_falls
            after_e = set(self.defined)
            self.ind -= 1
        self.emit("}")
        if h_falls and e_falls:
            self.defined = after_h & after_e
        elif h_falls:
            self.defined = after_h
        elif e_falls:
            self.defined = after_e
        else:
            return False
        return True

    def try_finally(self, s: ast.Try, loop_depth: int) -> bool:
        """try [except] [else] finally == try: (try [except] [else]) finally: F.  F runs on the way out of the inner
        statement, whether it completed or an exception is leaving it; afterwards that exception continues.  Only without
        return / break / continue anywhere inside (they would have to run F on their way out too)."""
        for part in (s.body, s.handlers, s.orelse, s.finalbody):
            for st_ in part:
                if any(isinstance(n, (ast.Return, ast.Break, ast.Continue)) for n in ast.walk(st_)):
                    self.fail(s, "return / break / continue inside try ... finally")
        inner = ast.copy_location(ast.Try(body=s.body, handlers=s.handlers, orelse=s.orelse, finalbody=[]), s)
        self.ntry += 1
        k = self.ntry
        label, done, flag = f"ks_fin{k}", f"ks_fin{k}_ok", f"fin{k}"
        before = set(self.defined)
        if not self.try_labels:
            self.emit("if (st) KS_RETURN(0);")   # a sticky failure from before the statement is not its body's
        self.emit(f"int {flag} = 0;")
        self.emit("{")
        self.ind += 1
        self.try_labels.append(label)
        try:
            if s.handlers or s.orelse:
                inner_falls = self.stmt(inner, loop_depth)
            else:
                inner_falls = self.block(s.body, loop_depth)
        except _Returned:
            inner_falls = False
        finally:
            self.try_labels.pop()
        after_inner = set(self.defined)
        self.ind -= 1
        self.emit("}")
        self.emit(f"goto {done};")
        self.emit(f"{label}: {flag} = 1;")
        self.emit(f"{done}:;")
        self.emit(f"if ({flag} && st != 1) KS_RETURN(KS_RET_UNSUPPORTED);")   # leaving the subset is not an exception
        self.emit(f"if ({flag}) st = 0;")                                      # F computes normally; re-raised below
        self.defined = set(before)
        f_falls = self.block(s.finalbody, loop_depth)
        after_f = set(self.defined)
        if f_falls:
            self.emit(f"if ({flag}) {self.exc()}")   # the exception that was leaving continues (to an outer handler or out)
        self.defined = (after_inner | after_f) if inner_falls else after_f
        return f_falls and inner_falls

    def run(self) -> bool:
        return self.block(self.fn.body[self.skip:])


def _module_constants(tree: ast.Module, fn: ast.FunctionDef) -> Dict[str, object]:
    consts: Dict[str, object] = {}
    for s in tree.body:
        if s is fn:
            continue
        if isinstance(s, ast.Expr) and isinstance(s.value, ast.Constant):
            continue
        if isinstance(s, ast.Assign) and len(s.targets) == 1 and isinstance(s.targets[0], ast.Name):
            try:
                consts[s.targets[0].id] = ast.literal_eval(s.value)
                continue
            except Exception:  # noqa: BLE001
                pass
        if isinstance(s, ast.FunctionDef):
            lam = _def_as_lambda(s)
            if lam is None:
                raise PolicyCompileError(f"line {s.lineno}: helper function {s.name!r} is more than a single return statement")
            # inlined at the call site, where names resolve in priority_function's scope: the helper may only read its
            # own parameters, module-level constants and builtins
            params = {a.arg for a in s.args.args}
            local_names = {n.id for n in ast.walk(fn) if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Store)} | \
                          {a.arg for a in fn.args.args}
            free = {n.id for n in ast.walk(lam.body) if isinstance(n, ast.Name)} - params
            if free & local_names:
                raise PolicyCompileError(f"line {s.lineno}: helper {s.name!r} reads a name that is local to priority_function")
            consts[s.name] = lam
            continue
        raise PolicyCompileError(f"line {getattr(s, 'lineno', '?')}: unsupported module-level statement")
    return consts


_PREFIX_SRC = """
def priority_function(pod, node):
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
"""
_PREFIX_DUMP = [ast.dump(st) for st in ast.parse(_PREFIX_SRC).body[0].body]


class _Rename(ast.NodeTransformer):
    def __init__(self, mapping):
        self.mapping = mapping

    def visit_Name(self, node):
        return ast.copy_location(ast.Name(id=self.mapping.get(node.id, node.id), ctx=node.ctx), node)


def _reads_outside_own_binding(stmts, names) -> bool:
    """True if any of `names` is read in `stmts` other than inside a comprehension / lambda / for-loop that
    binds it first."""
    def visit(node, bound):
        if isinstance(node, ast.Name):
            return isinstance(node.ctx, ast.Load) and node.id in names and node.id not in bound
        if isinstance(node, (ast.ListComp, ast.GeneratorExp, ast.SetComp, ast.DictComp)):
            inner = set(bound)
            for g in node.generators:
                if visit(g.iter, inner):
                    return True
                inner |= {n.id for n in ast.walk(g.target) if isinstance(n, ast.Name)}
                if any(visit(c, inner) for c in g.ifs):
                    return True
            elts = [node.elt] if hasattr(node, "elt") else [node.key, node.value]
            return any(visit(e, inner) for e in elts)
        if isinstance(node, ast.Lambda):
            return visit(node.body, bound | {a.arg for a in node.args.args})
        if isinstance(node, ast.For):
            if visit(node.iter, bound):
                return True
            inner = bound | {n.id for n in ast.walk(node.target) if isinstance(n, ast.Name)}
            # after the loop the name may be unbound again (zero iterations): only the body is covered
            return any(visit(c, inner) for c in node.body) or any(visit(c, bound) for c in node.orelse)
        return any(visit(c, bound) for c in ast.iter_child_nodes(node))
    return any(visit(st, set()) for st in stmts)


def _hoistable_prefix(fn: ast.FunctionDef) -> int:
    """Number of leading statements (docstring included) that are exactly the reference template's
    feasibility prefix (funsearch/safe_execution.py:204-216) and can be evaluated by the kernel instead."""
    body = fn.body
    k = 1 if body and isinstance(body[0], ast.Expr) and isinstance(getattr(body[0], "value", None), ast.Constant) else 0
    if len(body) < k + 2 or len(fn.args.args) != 2:
        return 0
    pod, node = fn.args.args[0].arg, fn.args.args[1].arg
    a, b = body[k], body[k + 1]
    try:
        cnt = b.body[0].targets[0].id
        gpu = b.body[1].target.id
    except (AttributeError, IndexError):
        return 0
    mapping = {pod: "pod", node: "node", cnt: "available_gpus", gpu: "gpu"}
    if len(set(mapping)) != 4 or {cnt, gpu} & {"pod", "node"}:
        return 0
    # compare under the renaming without copying the statements: the names are mapped in place for the dump and put back
    names = [n for x in (a, b) for n in ast.walk(x) if isinstance(n, ast.Name)]
    saved = [n.id for n in names]
    try:
        for n in names:
            n.id = mapping.get(n.id, n.id)
        same = [ast.dump(x) for x in (a, b)] == _PREFIX_DUMP
    finally:
        for n, old in zip(names, saved):
            n.id = old
    if not same:
        return 0
    if _reads_outside_own_binding(body[k + 2:], {cnt, gpu}):
        return 0
    return k + 2


_CACHE: Dict[str, CompiledPolicy] = {}


def cached(source: str) -> Optional[CompiledPolicy]:
    """The translation of `source` if this process already has it."""
    return _CACHE.get(hashlib.sha256(source.encode()).hexdigest())


def adopt(source: str, cuda_source: str, symbol: str, flags: int, consts: Tuple[float, ...] = ()) -> CompiledPolicy:
    """Register a translation produced by another process (compile_pool) under this process's cache."""
    digest = hashlib.sha256(source.encode()).hexdigest()
    cp = _CACHE.get(digest)
    if cp is None:
        cp = CompiledPolicy(source, cuda_source, symbol, {}, flags, tuple(consts))
        _CACHE[digest] = cp
    return cp


def _is_lifted_value(c) -> bool:
    """Float literals that go to the constant vector: all but +-0.0 (the usual `score = 0.0`, which costs nothing as
    an immediate and which mutations leave alone)."""
    return type(c) is float and c != 0.0


def _lift_literals(fn: ast.FunctionDef, consts: Dict[str, object]):
    """Numbers the float literals of the function body (document order) and of the module-level constants the
    function can see (in order of definition).  Returns (id(node) -> index, module constants with their floats
    replaced by _Lifted markers, the values)."""
    values: List[float] = []
    lift: Dict[int, int] = {}

    def visit(node):   # document order, like ast.NodeVisitor
        if type(node) is ast.Constant:
            if _is_lifted_value(node.value):
                lift[id(node)] = len(values)
                values.append(node.value)
            return
        for f in node._fields:
            v = getattr(node, f, None)
            if type(v) is list:
                for x in v:
                    if isinstance(x, ast.AST):
                        visit(x)
            elif isinstance(v, ast.AST):
                visit(v)
    visit(fn)

    def mark(c):
        if _is_lifted_value(c):
            values.append(c)
            return _Lifted(len(values) - 1, c)
        if isinstance(c, (list, tuple)):
            return type(c)(mark(x) for x in c)
        return c
    return lift, {k: mark(v) for k, v in consts.items()}, values


def compile_policy(source: str) -> CompiledPolicy:
    """Translate a policy source to CUDA C.  Raises PolicyCompileError outside the subset; a
    SyntaxError in the source is reported the same way (the reference scores such a policy 0)."""
    digest = hashlib.sha256(source.encode()).hexdigest()
    hit = _CACHE.get(digest)
    if hit is not None:
        return hit
    try:
        tree = ast.parse(source)
    except (SyntaxError, ValueError, RecursionError, MemoryError) as e:
        raise PolicyCompileError(f"syntax error: {e}") from e
    fns = [s for s in tree.body if isinstance(s, ast.FunctionDef) and s.name == "priority_function"]
    if not fns:
        raise PolicyCompileError("no priority_function defined")
    fn = fns[-1]
    fn._ks_nodes = list(ast.walk(fn))   # walked once: the checks below and _Translator._scan go over the list
    for n in fn._ks_nodes:
        if isinstance(n, (ast.AsyncFunctionDef, ast.ClassDef)):
            raise PolicyCompileError(f"line {n.lineno}: nested class / async definitions are not supported")
        if isinstance(n, (ast.With, ast.Global, ast.Nonlocal, ast.Yield, ast.YieldFrom, ast.Await, ast.Delete,
                          ast.NamedExpr, ast.JoinedStr, ast.DictComp, ast.SetComp, ast.Starred)):
            raise PolicyCompileError(f"line {getattr(n, 'lineno', '?')}: {type(n).__name__} is not supported")
    consts = _module_constants(tree, fn)
    lift, consts, kvalues = _lift_literals(fn, consts)
    ilift: Dict[tuple, int] = {}

    skip = _hoistable_prefix(fn)
    var_types: Dict[str, str] = {}
    var_bits: Dict[str, int] = {}
    base_bits: Dict[str, int] = {}
    flagged: set = set()
    tr = None
    for it in range(8):
        tr = _Translator(fn, consts, var_types, var_bits, flagged, base_bits, skip, lift, kvalues, ilift)
        falls = tr.run()
        new_base = dict(base_bits)
        for k, b in tr.seen_base.items():
            new_base[k] = max(new_base.get(k, 0), b)
        new_types = dict(var_types)
        for k, t in tr.var_types.items():
            new_types[k] = _join(new_types.get(k), t)
        new_bits = dict(var_bits)
        for k, b in tr.var_bits.items():
            nb = max(new_bits.get(k, 0), b)
            if it >= 3 and nb > new_bits.get(k, 0):
                nb = 64  # a loop-carried integer keeps growing: give up on the range and check at run time
            new_bits[k] = min(64, nb)
        new_flagged = flagged | tr.need_flag
        if new_types == var_types and new_bits == var_bits and new_flagged == flagged and new_base == base_bits:
            break
        var_types, var_bits, flagged, base_bits = new_types, new_bits, new_flagged, new_base
    else:
        raise PolicyCompileError("type inference did not converge")

    flags = 1 if skip else 0
    body = ["  int st = 0; (void)st;",
            "  const int ngpus = ks_nc(nc, KS_NC_NUM_GPUS); (void)ngpus;"]
    if tr.uses_budget:
        body.append(f"  int budget = {LOOP_BUDGET};")
    for name in sorted(tr.var_types):
        ty = tr.var_types[name]
        if ty == ELIST:
            ty = ILIST   # never appended to on any path the translator saw: an empty list of whatever
        init = {INT: "0", FLOAT: "0.0", BOOL: "0", NUM: "ks_num_i(0)", GPU: "0"}.get(ty)
        if init is None:
            body.append(f"  {CTYPE[ty]} v_{name}; v_{name}.n = 0;")
        else:
            body.append(f"  {CTYPE[ty]} v_{name} = {init};")
        if name in flagged:
            body.append(f"  int b_{name} = 0;")
    body.extend(tr.lines)
    # reached only by falling off the end of the Python function: int(max(0, None)) raises TypeError
    body.append("  KS_RETURN(KS_RET_EXCEPTION);")
    body.append("}")
    text = "\n".join(body) + "\n"
    if "p_ctime" in text:
        flags |= 2   # KS_PF_READS_TIME: the one pod field a retry changes (main.py:118-121); the kernel then rescans all nodes
    if "p_dur" in text:
        flags |= 4   # KS_PF_READS_DURATION: pods of one request shape may score differently; no per-shape score rows
    # the structure's name is the hash of its code: equal code <=> one device function, whatever source it came from
    symbol = "ks_st_" + hashlib.sha256(f"{flags}|{len(kvalues)}|".encode() + text.encode()).hexdigest()[:20]
    head = (f"// policy structure {symbol}: a priority_function(pod, node) with its {len(kvalues)} float literals in kf[]\n"
            f'extern "C" __device__ long long {symbol}(KS_POLICY_PARAMS) {{\n')
    cp = CompiledPolicy(source, head + text, symbol, dict(tr.var_types), flags, tuple(kvalues))
    _CACHE[digest] = cp
    return cp


def structure_of(source: str) -> str:
    """Name of the structure `source` translates to (raises PolicyCompileError outside the subset)."""
    return compile_policy(source).symbol
