// ks_abi.h -- structures shared by the host library, the episode-replay kernel and the
// NVRTC-compiled policy functions.  Plain C layout, no CUDA types, so the same file is
// given to NVRTC as an in-memory header.
//
// Vocabulary follows the reference: pods, nodes, GPUs, episodes (one episode = one
// (policy, trace) replay = one KubernetesSimulator.run_schedule(), simulator/main.py:50-72).
#ifndef KS_ABI_H_
#define KS_ABI_H_

// ---- per-episode status words (SURVEY.md section 8b error convention) --------------------
#define KS_ST_OK 0                // episode ran to the end of the event heap
#define KS_ST_POLICY_EXCEPTION 1  // the policy raised (ZeroDivisionError, OverflowError ...)
#define KS_ST_GPU_SHORTFALL 2     // ValueError("Not enough GPUs available"), simulator/main.py:164-165
#define KS_ST_STEP_BUDGET 3       // caller-supplied event budget exhausted (the reference has none)
#define KS_ST_UNSUPPORTED 4       // policy left the supported subset at run time (int64 overflow, complex ...)
#define KS_ST_CAPACITY 5          // a per-slot scratch capacity was exceeded
#define KS_ST_INVARIANT 6         // debug build only: _validate_cluster_invariants failed (simulator/main.py:201-272);
                                  // KsResult.reserved[0] = (number of the failing check << 32) | node index

// ---- what a policy function returns instead of a score when it cannot -----------------
// Scores handed to the simulator are int(max(0, score)) >= 0 (funsearch_integration.py:96),
// so negative values are free to carry the exception channel.
#define KS_RET_EXCEPTION (-1LL)
#define KS_RET_UNSUPPORTED (-2LL)

// ---- event-heap entry: (time:39 | id_rank:23 | is_deletion:1) -------------------------
// Order of two entries == order of the reference's (time, Event) tuples: time first, then the
// pod_id string order (event_simulator.py:16-17) which the host pre-computes as id_rank.
// At most one event per pod is ever queued, so the low bit never decides a comparison.
#define KS_RANK_BITS 23
#define KS_KEY_SHIFT (KS_RANK_BITS + 1)
#define KS_MAX_PODS (1 << KS_RANK_BITS)
#define KS_MAX_TIME ((1LL << 39) - 2)
#define KS_HEAP_INF 0xFFFFFFFFFFFFFFFFull

// ---- placement record: (final creation time:40 | node:16 | gpu mask:8) -----------------
#define KS_NODE_NONE 0xFFFFu     // never attempted / never placed
#define KS_NODE_WAITING 0xFFFEu  // failed at least once and not placed (member of waiting_pods)
#define KS_MAX_NODES 0xFFF0
#define KS_MAX_GPUS_PER_NODE 8
#define KS_SCRATCH_HEADER 4096   // bytes in front of the per-team scratch: one work counter per CTA (<= 1024 CTAs)

// ---- mutable per-node state: KS_NS_FIELDS ints per node, node-major (48 bytes, 16-byte aligned rows) -----
// Rows are contiguous so that the kernel reads a node with three 16-byte loads and a policy addresses a field
// as base + constant.  Slot 3 is padding; GPU slots a node does not have hold 0 for the whole episode.
#define KS_NS_CPU_LEFT 0
#define KS_NS_MEM_LEFT 1
#define KS_NS_GPU_LEFT 2
#define KS_NS_GPU_MILLI_LEFT 4  // .. +7
#define KS_NS_FIELDS 12

// ---- constant per-node data: KS_NC_FIELDS ints per node, node-major (80 bytes) ----------------------------
#define KS_NC_CPU_TOTAL 0
#define KS_NC_MEM_TOTAL 1
#define KS_NC_NUM_GPUS 2
#define KS_NC_TYPE 3             // index of the first node with an identical constant row ("node type"), set by ks_cluster_pack
#define KS_NC_GPU_MILLI_TOTAL 4  // .. +7
#define KS_NC_GPU_MEM_TOTAL 12   // .. +7   (GPU.memory_mib_total == memory_mib_left: never mutated)
#define KS_NC_FIELDS 20

// ---- packed pod record, indexed by id_rank, 32 bytes -------------------------------------
#define KS_POD_CPU 0
#define KS_POD_MEM 1
#define KS_POD_NUM_GPU 2
#define KS_POD_GPU_MILLI 3
#define KS_POD_DUR_LO 4
#define KS_POD_DUR_HI 5
#define KS_POD_ORIG_INDEX 6
#define KS_POD_SHAPE 7  // 1 + rank of the pod's (cpu_milli, memory_mib, num_gpu, gpu_milli) among the trace's distinct request
                        // shapes, most frequent first (ks_trace_pack); 0 = not ranked (the kernel then keeps no scores for it)
#define KS_POD_INTS 8

// ---- policy function ABI (v3) ----------------------------------------------------------
// One translated priority_function(pod, node) STRUCTURE: the source with its float literals lifted out.  Policies
// that differ only in their float constants (what most mutations of a FunSearch generation change) share one device
// function and differ in the constant vector `kf` they are called with.  The pod's fields (entities.py:30-43) travel
// in registers (p_ctime is the retry-mutated creation_time == time of the creation event being handled); `ns` is the
// SHARED-memory byte address of this node's state row (field f at ns + 4 * f), `nc` points at this node's
// row of constants in global memory (field f at nc[f]), `kf` at the policy's float literals in source order.
// Returns int(max(0, score)) >= 0, or KS_RET_EXCEPTION / KS_RET_UNSUPPORTED.
#define KS_POLICY_PARAMS                                                                                             \
  int p_cpu, int p_mem, int p_ngpu, int p_gmilli, long long p_ctime, long long p_dur, unsigned ns, const int *__restrict__ nc, \
      const double *__restrict__ kf
typedef long long (*ks_policy_fn)(KS_POLICY_PARAMS);
#define KS_SHAPE_SLOTS 256  // request shapes per episode whose scores per node are kept (teams of warps)
#define KS_MAX_TABLE 4096  // policy structures one linked module can dispatch to
// flags of a dispatch-table entry
#define KS_PF_PREFIX_HOISTED 1  // the source began with the reference template's feasibility prefix
                                // (funsearch/safe_execution.py:204-216); the kernel evaluates it, the
                                // function holds only what follows and is called for feasible nodes only
#define KS_PF_READS_TIME 2      // the function reads pod.creation_time, which a retry changes: its retries are scored on
                                // every node again (otherwise only on the nodes that changed since the pod last failed)
#define KS_PF_READS_DURATION 4  // the function reads pod.duration_time: pods of one request shape may score differently, so
                                // the kernel keeps no per-shape scores for it (teams of warps, see KS_SHAPE_SLOTS)
struct KsPolicyEntry {
  ks_policy_fn fn;
  long long flags;
};

// One trace resident in HBM.
struct KsTraceDev {
  const int* podrec;                  // [n_pods * KS_POD_INTS], indexed by id_rank
  const unsigned long long* heap0;    // [n_pods] heapified creation events (event_simulator.py:23-34)
  const unsigned long long* place0;   // [n_pods] initial placement records (orig creation time, NONE, 0)
  long long n_pods;
};

// Per-episode output, 128 bytes.  Mirrors EvaluationResults (evaluator.py:16-25) plus counters.
struct KsResult {
  int status;
  int placed;
  int num_snapshots;
  int num_fragmentation_events;
  int max_nodes;
  int max_waiting;
  long long events;
  long long attempts;
  long long last_event_time;
  double avg_cpu_utilization;
  double avg_memory_utilization;
  double avg_gpu_count_utilization;
  double avg_gpu_memory_utilization;
  double gpu_fragmentation_score;
  double policy_score;  // evaluator.py:101-127, 0 when status != OK
  long long scorer_calls;
  long long device_cycles;  // clock64() span of the episode on its SM (diagnostic)
  long long reserved[2];
};

// Kernel argument block.
struct KsLaunch {
  const KsTraceDev* traces;
  const int* ep_policy;  // [n_episodes] slot in the linked policy table (the policy's structure)
  const int* ep_const;   // [n_episodes] offset (in doubles) of the policy's float literals in `consts`; null = 0 everywhere
  const double* consts;  // float literals of all policies of the batch (may be null when no policy has any)
  const int* ep_trace;   // [n_episodes]
  const int* nconst;     // [n_pad][KS_NC_FIELDS]
  int* work_counter;     // [gridDim.x] zeroed by the host before launch: next entry of each CTA's range of the table
  const int* range_begin;  // optional [gridDim.x + 1]: explicit table ranges per CTA (cost-packed table, no stealing)
  KsResult* results;     // [n_episodes]
  unsigned long long* placements;  // optional; episode e writes [place_off[e], +n_pods)
  const long long* place_off;      // [n_episodes] or null
  unsigned char* scratch;          // per resident team: event heap + (placements if not requested) + waiting list
  long long scratch_per_team;
  long long scratch_place_off;  // byte offset of the placement records inside a team's scratch (if not an output)
  long long scratch_wait_off;   // byte offset of the waiting list inside a team's scratch
  long long scratch_shape_off;  // teams of warps: byte offset of the per-shape stamps and score rows inside a team's scratch
  long long tot_cpu, tot_mem, tot_gpu_count, tot_gpu_milli;
  long long max_events;  // 0 = unlimited
  double snapshot_interval;
  int n_episodes;
  int n_nodes;
  int n_pad;
  int shape_slots;        // teams of warps: request shapes whose scores per node are kept between attempts (0 = none)
  int class_slots;        // teams of warps: entries of the node-class table in shared memory (a power of two >= 2 * n_pad)
  int heap_smem_entries;  // reserved (0): the heap lives in the team's scratch, L1-resident
  int wait_smem_entries;  // reserved (0)
  int max_pods;           // largest n_pods over the traces of this batch (sizes the scratch regions)
  int smem_per_team;      // bytes
  int warps_per_team;     // warps cooperating on one episode (1 = warp per episode)
};

#ifdef __cplusplus
static_assert(sizeof(KsResult) == 128, "KsResult is 128 bytes");
#endif

#endif  // KS_ABI_H_
