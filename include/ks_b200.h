/* ks_b200.h -- C ABI of libks_b200.so: the B200 (sm_100a) batched fitness evaluator that stands
 * in for the reference's per-policy CPU replay.
 *
 * Every entry point is plain C: pointers and sizes only, no torch / CUDA-runtime types.  Device
 * memory is owned by the caller (the Python host uses torch tensors and passes data_ptr());
 * the library itself never allocates device memory.  All functions return 0 on success and a
 * non-zero code otherwise; ks_last_error() then describes the failure (thread-local string).
 *
 * Each declaration cites the reference interface it replaces (paths relative to the reference
 * repository ttanv/funsearch-kubernetes-simulator).
 */
#ifndef KS_B200_H_
#define KS_B200_H_

#include <stddef.h>

#include "ks_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

#define KS_API_VERSION 3

/* Error text of the last failing call on this thread.  Replaces the Python exception text the
 * reference's worker swallows (funsearch/funsearch_integration.py:63-64). */
const char* ks_last_error(void);

int ks_api_version(void);

/* Build variant of the loaded library: bit 0 = diagnostic phase clocks, bit 1 = KS_DEBUG_INVARIANTS (the kernel runs
 * _validate_cluster_invariants, simulator/main.py:201-272, after every mutation and ends the episode with status
 * KS_ST_INVARIANT on the first violation -- the counterpart of KubernetesSimulator(..., validate_invariants=True)). */
int ks_build_flags(void);

/* Bind the calling thread to `device` (its primary context, shared with torch) and load the CUDA
 * driver.  Fails loudly when no driver / no sm_100 device is present: there is no CPU fallback.
 * Replaces nothing in the reference (it has no device); precondition of everything below. */
int ks_init(int device);

/* ---- workload packing (host only, no GPU needed) ------------------------------------------ */

/* Pack a pod list into the device layout: pod records indexed by id_rank, the heapified initial
 * event heap and the initial placement image.
 * Replaces: DiscreteEventSimulator.__init__/_populate_heap + heapq.heapify
 * (simulator/event_simulator.py:23-34) over Pod objects (simulator/entities.py:30-43) produced by
 * TraceParser.parse_pods (benchmarks/parser.py:68-101).
 *   id_rank[i] = rank of pod i's pod_id in string order (Event.__lt__, event_simulator.py:16-17)
 *   podrec  [n_pods * KS_POD_INTS] ints, heap0 [n_pods], place0 [n_pods]: caller-allocated host arrays
 * A pod record also carries the rank of the pod's request shape (cpu_milli, memory_mib, num_gpu, gpu_milli) among the
 * trace's distinct shapes, commonest first (KS_POD_SHAPE): on clusters replayed by teams of warps the kernel keeps each
 * node's score per shape and rescores only the nodes that changed since the shape's last pod. */
int ks_trace_pack(long long n_pods, const int* cpu_milli, const int* memory_mib, const int* num_gpu,
                  const int* gpu_milli, const long long* creation_time, const long long* duration_time,
                  const int* id_rank, int* podrec, unsigned long long* heap0, unsigned long long* place0);

/* Pack a cluster into the constant node table (node-major rows of KS_NC_FIELDS ints, n_pad = n_nodes rounded
 * up to 32 rows, rows past n_nodes zero) and its evaluator totals {cpu, mem, gpu count, gpu milli}.
 * Replaces: Cluster/Node/GPU objects (simulator/entities.py:4-26) from TraceParser.parse_nodes
 * (benchmarks/parser.py:23-61) and SchedulingEvaluator.__init__'s totals (simulator/evaluator.py:35-38).
 *   gpu_milli_total, gpu_mem_total: [n_nodes * 8] row-major, unused slots ignored
 *   nconst: [KS_NC_FIELDS * n_pad] ints, totals: [4] */
int ks_cluster_pack(int n_nodes, const int* cpu_total, const int* mem_total, const int* num_gpus,
                    const int* gpu_milli_total, const int* gpu_mem_total, int* nconst, long long* totals);

/* ---- policies --------------------------------------------------------------------------------- */

/* Compile one policy STRUCTURE, given as CUDA C source of a device function
 *     extern "C" __device__ long long <symbol>(KS_POLICY_PARAMS)          (include/ks_abi.h)
 * (what the Python-AST translator emits: the policy with its float literals lifted into the argument `kf`, so that
 * all policies that differ only in those constants share one function), with NVRTC for sm_100a into a relocatable
 * cubin.  `flags` are the KS_PF_* bits the dispatch table will carry for this function (KS_PF_READS_TIME and
 * KS_PF_READS_DURATION are added by the library when the source mentions p_ctime / p_dur: the kernel's retry shortcut and
 * its per-shape score rows are only valid for functions that do not see those fields).  Results
 * are cached by source hash (in memory, and on disk when ks_set_cache_dir was called).
 * Replaces: FunSearchScheduler._compile_policy, exec of the policy source
 * (funsearch/funsearch_integration.py:77-89).  *policy receives a handle valid until ks_shutdown. */
int ks_policy_compile(const char* cuda_source, const char* symbol, long long flags, long long* policy);

/* Same for n policies, compiled on `threads` host threads (0 = hardware concurrency).  status[i] is 0
 * or the error code of policy i (its handle is then -1); the call itself fails only on bad arguments. */
int ks_policy_compile_many(int n, const char* const* cuda_sources, const char* const* symbols, const long long* flags,
                           int threads, long long* policies, int* status);

/* How many policies ks_policy_compile_many puts into one NVRTC program (default 16).  NVRTC start-up
 * and header parsing dominate the cost of a small policy, so grouping cuts compile time about 3x. */
int ks_set_compile_batch(int policies_per_program);

/* Log of the last failed compilation on this thread (NVRTC diagnostics). */
const char* ks_compile_log(void);

int ks_set_cache_dir(const char* path);

/* Link the replay kernel with a table of policy structures (slot i of the table = policies[i], at most KS_MAX_TABLE)
 * and load the module on the current device.  Nothing is compiled here: the cubins are linked (nvJitLink) and the
 * table is filled from the function addresses the loaded module reports.  Programs are cached by their policy list.
 * Replaces: handing a scorer to KubernetesSimulator(..., scheduler=...) (simulator/main.py:29-41). */
int ks_program_link(int n_policies, const long long* policies, long long* program);
int ks_program_release(long long program);
/* The two halves of ks_program_link for hosts that link in helper processes while the GPU process does other work (a
 * fresh generation's structures are compiled AND linked by the compile workers, one module per worker): the first
 * writes the linked, loadable image to a file and needs no GPU; the second loads an image.  `policies` must name the
 * same structures in the same order on both sides (handles are per process).  Replaces the same step of the reference
 * as ks_program_link (simulator/main.py:29-41). */
int ks_program_link_image(int n_policies, const long long* policies, const char* path);
int ks_program_load_image(const void* image, long long bytes, int n_policies, const long long* policies, long long* program);
/* Link only (no device needed): proves every policy symbol resolves against the replay kernel. */
int ks_program_link_check(int n_policies, const long long* policies, long long* linked_bytes);

/* ---- evaluation ----------------------------------------------------------------------------- */

/* Launch geometry and scratch requirement for a batch.  Fill n_episodes, n_nodes, max_pods,
 * want_placements and (optionally, 0 = choose) warps_per_team / teams_per_block / ctas_per_sm; the call fills the
 * rest.  Purely arithmetic apart from reading the device's SM count and shared-memory limit. */
typedef struct KsPlan {
  int n_episodes;
  int n_nodes;
  int max_pods;
  int want_placements;
  int warps_per_team;    /* warps replaying one episode together (0 = choose: 1 up to 64 nodes, else one lane per node) */
  int teams_per_block;   /* episodes in flight per CTA */
  int heap_smem_entries; /* reserved, returned as 0: the event heap lives in the team's scratch (L1/L2-resident) */
  int wait_smem_entries; /* reserved, returned as 0 */
  int ctas_per_sm;       /* 0 = choose; 2 = throughput variant of the kernel (128 registers, two CTAs per SM) */
  int blocks;
  int smem_per_team;
  int smem_per_block;
  long long scratch_per_team;
  long long scratch_bytes; /* total device scratch the caller must provide (KS_SCRATCH_HEADER bytes of work counters included) */
} KsPlan;
int ks_plan(KsPlan* plan);

/* One batch of episodes, everything already resident in device memory.
 * Replaces: the ProcessPoolExecutor block over evaluate_policy_standalone
 * (funsearch/funsearch_integration.py:30-64, 535-540), i.e. per episode
 * KubernetesSimulator.run_schedule() (simulator/main.py:50-72) + SchedulingEvaluator
 * (simulator/evaluator.py:27-164) + get_policy_score (:101-127). */
typedef struct KsBatch {
  const KsTraceDev* traces;       /* device: table of traces */
  const int* ep_policy;           /* device [n_episodes]: slot in the program's policy table (the policy's structure) */
  const int* ep_const;            /* device [n_episodes], optional: where the policy's float literals start in consts */
  const double* consts;           /* device, optional: the float literals of the batch's policies, back to back */
  const int* ep_trace;            /* device [n_episodes] */
  const int* range_begin;         /* device [plan.blocks + 1], optional: CTA b replays table entries [range_begin[b],
                                     range_begin[b + 1]) and nothing else (a table packed by predicted cost); null = the
                                     kernel cuts the table into equal ranges and idle teams steal from other ranges.
                                     Explicit ranges belong to one block count: do not edit plan.blocks with them */
  const int* nconst;              /* device: packed cluster (ks_cluster_pack) */
  KsResult* results;              /* device [n_episodes] */
  unsigned long long* placements; /* device, optional */
  const long long* place_off;     /* device [n_episodes], required with placements */
  void* scratch;                  /* device, plan.scratch_bytes, 256-byte aligned */
  long long totals[4];            /* from ks_cluster_pack */
  long long max_events;           /* 0 = no event budget (the reference has none) */
  double snapshot_interval;       /* SchedulingEvaluator default 0.05 (evaluator.py:30) */
} KsBatch;
int ks_eval_batch(long long program, const KsPlan* plan, const KsBatch* batch, void* stream);

/* ks_eval_batch for hosts that must not allocate device memory for the kernel's temporaries.
 * (No counterpart in the reference: its worker is plain CPython and has no device.)
 *
 * Who needs it: ks_eval_batch expects the caller to hand in `batch->scratch`, which in practice means
 * one more device allocation sized by ks_plan.  Some embedders cannot make that allocation: a service
 * that received its input/output buffers from another component and runs under a fixed device-memory
 * budget, or a host whose rules forbid cudaMalloc after start-up because it synchronises the whole
 * device and would stall batches in flight on other streams.  For them the library carries its own
 * scratch: a file-scope `static __device__` array of KS_STATIC_SCRATCH_BYTES (32 MiB) per GPU that
 * loads the library, held until the process exits.  This entry point uses that array instead of
 * `batch->scratch` (which is ignored and may be null); inputs and outputs are still the caller's.
 * It calls no cudaMalloc/cudaFree and no allocating library routine.
 *
 * Limit: what the plan's geometry needs -- KS_SCRATCH_HEADER + scratch_per_team x blocks x teams_per_block,
 * recomputed inside the call, and plan->scratch_bytes, which must not understate it -- must not exceed
 * KS_STATIC_SCRATCH_BYTES, otherwise the call fails with a message stating both numbers.  The call also fails when
 * the current device is not the one the program was loaded on (the array lives per GPU).
 * A plan that is too large can be shrunk by lowering plan->blocks
 * (fewer resident teams; the kernel hands out the episode table dynamically, so any block count is correct) and
 * setting plan->scratch_bytes = KS_SCRATCH_HEADER + plan->scratch_per_team * plan->blocks * plan->teams_per_block.
 *
 * Set-up: the first call on each device looks up the array's device address (cudaGetSymbolAddress,
 * with that device current) and remembers it; make that first call before any stream capture begins.
 * Sharing: every call on a device uses the same array, which keeps the previous call's contents.  The
 * work counters at its start are cleared on `stream` before each launch and the kernel rewrites every
 * other byte it reads, so back-to-back calls on one stream are fine, but two calls that can overlap
 * on the GPU (different streams, different threads) race on the array and must be serialised by the
 * caller. */
#define KS_STATIC_SCRATCH_BYTES (32ll << 20)
int ks_eval_batch_static(long long program, const KsPlan* plan, const KsBatch* batch, void* stream);

/* The same evaluation with HOST buffers: the call copies the workload (traces, cluster, episode
 * table) from host memory into the caller-provided device workspace, launches the replay kernel,
 * copies the results (and placements, if requested) back and waits for the stream.  Use pinned host
 * memory for truly asynchronous copies.  ks_host_workspace_bytes() tells how large the device
 * workspace must be; *h2d_bytes / *d2h_bytes (optional) receive the bytes moved.
 * Replaces, per call, what one pool worker of the reference does for every episode: parse_workload +
 * deepcopy + run_schedule + get_policy_score (funsearch/funsearch_integration.py:36-59). */
typedef struct KsHostTrace {
  const int* podrec;                /* host, from ks_trace_pack */
  const unsigned long long* heap0;  /* host */
  const unsigned long long* place0; /* host */
  long long n_pods;
} KsHostTrace;
typedef struct KsHostBatch {
  const KsHostTrace* traces;
  int n_traces;
  int reserved;
  const int* ep_policy;           /* host [n_episodes] */
  const int* ep_const;            /* host [n_episodes], optional */
  const double* consts;           /* host [n_consts], optional */
  long long n_consts;
  const int* ep_trace;            /* host [n_episodes] */
  const int* range_begin;         /* host [plan.blocks + 1], optional (see KsBatch) */
  const int* nconst;              /* host, from ks_cluster_pack */
  KsResult* results;              /* host [n_episodes], out */
  unsigned long long* placements; /* host, optional out: episode e at place_off[e] */
  const long long* place_off;     /* host [n_episodes], even offsets */
  long long placements_len;       /* entries in placements */
  long long totals[4];
  long long max_events;
  double snapshot_interval;
} KsHostBatch;
int ks_host_workspace_bytes(const KsPlan* plan, const KsHostBatch* batch, long long* bytes);
int ks_eval_batch_host(long long program, const KsPlan* plan, const KsHostBatch* batch, void* workspace,
                       long long workspace_bytes, void* stream, long long* h2d_bytes, long long* d2h_bytes);

/* Test hook: out[i] = pow(x[i], y[i]) as the policies' `**` computes it on the device (glibc-2.39-exact fast
 * path, see csrc/ks_glibc_pow.h).  Device pointers.  Replaces CPython's call into libm (Objects/floatobject.c
 * float_pow), which is what makes `**` in a reference policy platform-specific. */
int ks_debug_pow(long long program, const double* x, const double* y, double* out, int n, void* stream);

/* Test hook, same shape: op 0 = pow(x[i], y[i]), 1 = math.exp(x[i]), 2 = math.log(x[i]) (y unused for 1 and 2),
 * 3 = x[i] / y[i] (the shared fp64 division routine; must equal IEEE division bit for bit) as the
 * policies compute them on the device (csrc/ks_glibc_explog.h: glibc 2.39's exp and log, whole domain).  Replaces
 * CPython's calls into libm from Modules/mathmodule.c (math_exp, math_log). */
int ks_debug_math(long long program, int op, const double* x, const double* y, double* out, int n, void* stream);

/* Test hook: the wrapped score int(max(0, priority_function(pod, node))) of policy-table slot `slot` for n independent
 * (pod, node) cases, obtained exactly as the replay kernel obtains it; KS_RET_EXCEPTION / KS_RET_UNSUPPORTED where the policy
 * raises / leaves the subset.  Replaces a direct call of FunSearchScheduler.__call__ (funsearch_integration.py:91-101).
 * Device pointers: pods [n][6] int64 (cpu, mem, num_gpu, gpu_milli, creation_time, duration), ns [n][KS_NS_FIELDS],
 * nc [n][KS_NC_FIELDS] (rows as in ks_abi.h), kf = the policy's float literals (may be null if it has none), out [n]. */
int ks_debug_score(long long program, int slot, int n, const long long* pods, const int* ns, const int* nc, const double* kf,
                   long long* out, void* stream);

/* Number of replay-kernel launches issued by this process so far (bench.py's gpu_launches). */
long long ks_launch_count(void);

int ks_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* KS_B200_H_ */
