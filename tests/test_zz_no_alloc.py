"""ks_eval_batch_static: the entry point for embedders that may not allocate device memory for the
kernel's temporaries (fixed device-memory budget, or a no-cudaMalloc-after-start-up rule because
cudaMalloc synchronises the device).  It must give exactly what ks_eval_batch gives with caller-owned
scratch, call after call, while the amount of free device memory stays where it was.

Collected last on purpose (file name): if it fails, a run that stops at the first failure has already
executed every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import load_golden

pytestmark = pytest.mark.gpu

COMPARED = ("status", "placed", "num_snapshots", "num_fragmentation_events", "max_nodes", "max_waiting", "events",
            "attempts", "last_event_time", "avg_cpu_utilization", "avg_memory_utilization",
            "avg_gpu_count_utilization", "avg_gpu_memory_utilization", "gpu_fragmentation_score", "policy_score",
            "scorer_calls")


def _setup(built_lib, n_pods, n_ep):
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    _gold, pol = load_golden()
    names = ["ts_funsearch_4901", "ts_best_fit", "ts_funsearch_4800", "fs_worst_fit"]
    ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default").head(n_pods)])
    ev.pack_by_cost = False   # the default cut of the table: any block count replays every episode (explicit per-CTA
                              # ranges, KsBatch.range_begin, are tied to the plan's block count and cannot be shrunk)
    compiled = ev.compile([pol[n] for n in names])
    rb = ev.prepare_resident(compiled, [i % len(names) for i in range(n_ep)], [0] * n_ep)
    return ev, rb


def _static_call(ev, rb, plan=None):
    """Straight through the C ABI: same program / plan / inputs / outputs, scratch left null."""
    import torch
    from funsearch_kubernetes_simulator_b200 import _native as nat
    b = nat.KsBatch()
    C.memmove(C.byref(b), C.byref(rb.batch), C.sizeof(nat.KsBatch))
    b.scratch = None
    st = torch.cuda.current_stream(ev.device)
    return nat.lib().ks_eval_batch_static(rb.prog, C.byref(plan or rb.plan), C.byref(b), C.c_void_p(st.cuda_stream))


def test_static_scratch_entry_point_matches_caller_scratch_twice(built_lib):
    import torch
    from funsearch_kubernetes_simulator_b200 import _native as nat
    ev, rb = _setup(built_lib, 1500, 48)
    assert rb.plan.scratch_bytes <= nat.KS_STATIC_SCRATCH_BYTES
    rb.launch()                       # existing entry point, scratch allocated by the caller (torch)
    want = rb.results().copy()
    assert (want["status"] == 0).all() and (want["placed"] == 1500).all()
    free_before = None
    for call in range(2):             # two calls in a row share the built-in array
        rb.d_res.zero_()
        torch.cuda.synchronize()
        if call == 1:
            free_before = torch.cuda.mem_get_info()[0]
        rc = _static_call(ev, rb)
        assert rc == 0, nat.lib().ks_last_error().decode()
        torch.cuda.synchronize()
        if call == 1:
            assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the amount of free device memory"
        got = rb.results()
        for f in COMPARED:            # integers bit-exact; the fp64 metrics come from the same kernel: bit-exact too
            assert np.array_equal(got[f], want[f]), (call, f)


def test_static_scratch_limit_and_shrunk_plan(built_lib):
    import torch
    from funsearch_kubernetes_simulator_b200 import _native as nat
    ev, rb = _setup(built_lib, 8152, 600)
    assert rb.plan.scratch_bytes > nat.KS_STATIC_SCRATCH_BYTES
    rc = _static_call(ev, rb)
    msg = nat.lib().ks_last_error().decode()
    assert rc != 0 and str(rb.plan.scratch_bytes) in msg and str(nat.KS_STATIC_SCRATCH_BYTES) in msg
    # the documented way out: fewer resident blocks
    rb.launch()
    want = rb.results().copy()
    small = nat.KsPlan()
    C.memmove(C.byref(small), C.byref(rb.plan), C.sizeof(nat.KsPlan))
    small.blocks = max(1, int((nat.KS_STATIC_SCRATCH_BYTES - nat.KS_SCRATCH_HEADER) // (small.scratch_per_team * small.teams_per_block)))
    small.scratch_bytes = nat.KS_SCRATCH_HEADER + small.scratch_per_team * small.blocks * small.teams_per_block
    assert small.blocks < rb.plan.blocks and small.scratch_bytes <= nat.KS_STATIC_SCRATCH_BYTES
    rb.d_res.zero_()
    rc = _static_call(ev, rb, small)
    assert rc == 0, nat.lib().ks_last_error().decode()
    torch.cuda.synchronize()
    got = rb.results()
    for f in COMPARED:
        assert np.array_equal(got[f], want[f]), f


def test_static_scratch_rejects_plans_that_understate_their_need(built_lib):
    """The size check uses the plan's geometry, not the hand-editable scratch_bytes field: lowering scratch_bytes while
    leaving blocks alone (half of the documented shrink recipe) must be refused, as must nonsense geometry; the call
    never launches in these cases, so the result buffer keeps its zeros."""
    import torch
    from funsearch_kubernetes_simulator_b200 import _native as nat
    ev, rb = _setup(built_lib, 8152, 600)
    need = nat.KS_SCRATCH_HEADER + rb.plan.scratch_per_team * rb.plan.blocks * rb.plan.teams_per_block
    assert need > nat.KS_STATIC_SCRATCH_BYTES
    rb.d_res.zero_()
    torch.cuda.synchronize()

    def edited(**kw):
        p = nat.KsPlan()
        C.memmove(C.byref(p), C.byref(rb.plan), C.sizeof(nat.KsPlan))
        for k, v in kw.items():
            setattr(p, k, v)
        return p
    for tag, plan in (("scratch_bytes lowered, blocks unchanged", edited(scratch_bytes=nat.KS_STATIC_SCRATCH_BYTES)),
                      ("negative scratch_bytes", edited(scratch_bytes=-1)),
                      ("zero scratch_per_team", edited(scratch_per_team=0, scratch_bytes=nat.KS_SCRATCH_HEADER)),
                      ("zero blocks", edited(blocks=0, scratch_bytes=nat.KS_SCRATCH_HEADER))):
        rc = _static_call(ev, rb, plan)
        assert rc != 0, tag
        assert "ks_eval_batch_static" in nat.lib().ks_last_error().decode(), tag
    torch.cuda.synchronize()
    assert not rb.d_res.any().item()
    # the caller-scratch entry point checks the same relation
    rc = nat.lib().ks_eval_batch(rb.prog, C.byref(edited(scratch_bytes=4096)), C.byref(rb.batch),
                                 C.c_void_p(torch.cuda.current_stream(ev.device).cuda_stream))
    assert rc != 0 and "scratch_bytes" in nat.lib().ks_last_error().decode()


def test_static_scratch_checks_the_device(built_lib):
    """The array is per GPU: a call made while another device is current is refused (needs two visible GPUs)."""
    import torch
    from funsearch_kubernetes_simulator_b200 import _native as nat
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    ev, rb = _setup(built_lib, 500, 8)
    try:
        torch.cuda.set_device(1)
        rc = _static_call(ev, rb)
        assert rc != 0 and "current device" in nat.lib().ks_last_error().decode()
    finally:
        torch.cuda.set_device(0)
