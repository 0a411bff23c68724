/* ksim_oracle.c -- CPU oracle in plain C (TEST INFRASTRUCTURE, never the product path).
 *
 * A flat-array restatement of the reference's episode replay, fast enough for the 1M-pod x 1,024-node
 * shape that the Python oracle (oracle/pysim.py) cannot reach.  Only tests/, smoke() and bench.py's
 * cpu_baseline leg may load it.
 *
 * Parity status: PINNED through oracle/pysim.py -- tests/test_c_oracle.py requires this file to
 * reproduce the committed golden fixtures (tests/golden/golden_runs.json, recorded from the real
 * reference) bit for bit: placement logs, counters, and metrics.
 *
 * What it follows (reference file:line):
 *   heap           CPython Lib/heapq.py:132-145 (heappush/heappop), :207-219 (_siftdown), :260-278 (_siftup),
 *                  as used by simulator/event_simulator.py:19-59; keys ordered by (time, pod_id string rank)
 *   retry          event_simulator.py:51-59: first DELETION in heap ARRAY order, new time = its time + 1
 *   event loop     simulator/main.py:50-72 ; creation :101-148 ; deletion :74-99 ; GPU best fit :150-177
 *   evaluator      simulator/evaluator.py:55-75 (snapshots, fragmentation), :77-127 (means, fitness)
 *   policies       hand-translated from the policy SOURCES in tests/golden/policies.json (the reference's
 *                  tests/test_scheduler.py:20-218 and funsearch_integration.py:217-431 families), wrapped as
 *                  FunSearchScheduler does: int(max(0, score)) (funsearch_integration.py:96).
 *                  They are written by hand, independently of the Python->CUDA translator they check.
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared (oracle/build_oracle.py); -ffp-contract=off because
 * CPython never fuses a*b+c.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef long long ll;
typedef unsigned long long u64;

enum { ST_OK = 0, ST_POLICY_EXCEPTION = 1, ST_GPU_SHORTFALL = 2, ST_STEP_BUDGET = 3 };

typedef struct {
  int n;
  const int *cpu_total, *mem_total, *ngpus;
  const int* gpu_milli_total; /* [n*8] */
  int *cpu_left, *mem_left, *gpu_left;
  int* gml; /* [n*8] */
} Cluster;

typedef struct {
  int cpu, mem, ngpu, gmilli;
  ll ctime, dur;
} Pod;

typedef struct {
  int id;      /* which built-in policy */
  double p[4]; /* parameters of the random family */
} Policy;

/* ---- Python arithmetic helpers ---------------------------------------------------------------------- */
static int py_raise; /* set when CPython would raise */

static ll py_imod(ll a, ll b) {
  if (b == 0) { py_raise = 1; return 0; }
  ll r = a % b;
  if (r != 0 && ((r < 0) != (b < 0))) r += b;
  return r;
}
static double py_div(double a, double b) {
  if (b == 0.0) { py_raise = 1; return 0.0; }
  return a / b;
}
static int is_odd_int(double y) { return fmod(fabs(y), 2.0) == 1.0; }
/* float ** float: Objects/floatobject.c float_pow (complex results count as a raise here) */
static double py_pow(double x, double y) {
  if (y == 0.0) return 1.0;
  if (isnan(x)) return x;
  if (isnan(y)) return x == 1.0 ? 1.0 : y;
  if (isinf(y)) {
    double ax = fabs(x);
    if (ax == 1.0) return 1.0;
    return ((y > 0.0) == (ax > 1.0)) ? fabs(y) : 0.0;
  }
  if (isinf(x)) {
    int odd = is_odd_int(y);
    if (y > 0.0) return odd ? x : fabs(x);
    return odd ? copysign(0.0, x) : 0.0;
  }
  if (x == 0.0) {
    if (y < 0.0) { py_raise = 1; return 0.0; }
    return is_odd_int(y) ? x : 0.0;
  }
  int neg = 0;
  if (x < 0.0) {
    if (y != floor(y)) { py_raise = 1; return 0.0; }
    neg = is_odd_int(y);
    x = -x;
  }
  if (x == 1.0) return neg ? -1.0 : 1.0;
  double r = pow(x, y);
  if (isinf(r) || isnan(r)) { py_raise = 1; return 0.0; }
  return neg ? -r : r;
}
static ll py_int(double x) { /* int(float) */
  if (isnan(x) || isinf(x)) { py_raise = 1; return 0; }
  return (ll)x;
}
/* builtin sum() over floats starting from int 0 (Python/bltinmodule.c, 3.12: Neumaier) */
typedef struct { int started; double f, c; } FSum;
static void fsum_add(FSum* s, double x) {
  if (!s->started) { s->started = 1; s->f = 0.0 + x; s->c = 0.0; return; }
  double t = s->f + x;
  if (fabs(s->f) >= fabs(x)) s->c += (s->f - t) + x; else s->c += (x - t) + s->f;
  s->f = t;
}
static double fsum_result(const FSum* s) { /* caller handles the empty (int 0) case */
  double r = s->f;
  if (s->c != 0.0 && isfinite(s->c)) r += s->c;
  return r;
}
static ll imaxll(ll a, ll b) { return b > a ? b : a; } /* max(a, b): first maximal */
static ll iminll(ll a, ll b) { return b < a ? b : a; }
static double dmax2(double a, double b) { return b > a ? b : a; }
static double dmin2(double a, double b) { return b < a ? b : a; }

/* ---- policies ------------------------------------------------------------------------------------------ */
/* the template prefix (safe_execution.py:204-216); returns 0 when the node is infeasible */
static int feasible(const Pod* p, const Cluster* c, int n) {
  if (p->cpu > c->cpu_left[n] || p->mem > c->mem_left[n] || p->ngpu > c->gpu_left[n]) return 0;
  if (p->ngpu > 0) {
    int avail = 0;
    for (int g = 0; g < c->ngpus[n]; ++g)
      if (c->gml[n * 8 + g] >= p->gmilli) avail++;
    if (avail < p->ngpu) return 0;
  }
  return 1;
}

/* each returns the policy's own return value converted by the wrapper int(max(0, v)) */
static ll wrap_i(ll v) { return v > 0 ? v : 0; }

static ll pol_first_fit(const Pod* p, const Cluster* c, int n) {
  if (!feasible(p, c, n)) return 0;
  return 1000;
}

static ll pol_best_fit(const Pod* p, const Cluster* c, int n) { /* tests/test_scheduler.py best_fit_scheduler */
  if (!feasible(p, c, n)) return 0;
  ll rem_cpu = (ll)c->cpu_left[n] - p->cpu, rem_mem = (ll)c->mem_left[n] - p->mem, rem_g = (ll)c->gpu_left[n] - p->ngpu;
  double norm_cpu = py_div((double)rem_cpu, (double)c->cpu_total[n]);
  double norm_mem = py_div((double)rem_mem, (double)c->mem_total[n]);
  double norm_g = py_div((double)rem_g, (double)imaxll(c->ngpus[n], 1));
  double nr = (norm_cpu * 0.33 + norm_mem * 0.33) + norm_g * 0.34;
  ll score = py_int((1 - nr) * 10000);
  return wrap_i(imaxll(1, score));
}

static ll pol_worst_fit(const Pod* p, const Cluster* c, int n) { /* funsearch_integration.py _create_worst_fit_policy */
  if (!feasible(p, c, n)) return 0;
  double rc = py_div((double)((ll)c->cpu_left[n] - p->cpu), (double)c->cpu_total[n]);
  double rm = py_div((double)((ll)c->mem_left[n] - p->mem), (double)c->mem_total[n]);
  double rg = py_div((double)((ll)c->gpu_left[n] - p->ngpu), (double)imaxll(c->ngpus[n], 1));
  ll score = py_int(((rc * 0.33 + rm * 0.33) + rg * 0.34) * 10000);
  return wrap_i(imaxll(1, score));
}

static ll pol_random_family(const Pod* p, const Cluster* c, int n, const double* q) { /* _create_random_policy */
  if (!feasible(p, c, n)) return 0;
  /* score = base + cpu_left * cpu_factor + mem_left * mem_factor */
  double score = (q[3] + (double)c->cpu_left[n] * q[0]) + (double)c->mem_left[n] * q[1];
  if (p->ngpu > 0 && c->gpu_left[n] > 0) score += (double)c->gpu_left[n] * q[2];
  return wrap_i(imaxll(1, py_int(score)));
}

static ll pol_4901(const Pod* p, const Cluster* c, int n) {
  if (!feasible(p, c, n)) return 0;
  const int* gml = c->gml + n * 8;
  const int ng = c->ngpus[n];
  double cpu_u = py_div((double)((ll)c->cpu_total[n] - c->cpu_left[n]), (double)c->cpu_total[n]);
  double cpu_s = cpu_u < 0.7 ? (1.0 - cpu_u) * 100 : (1.0 - cpu_u) * 50;
  double mem_u = py_div((double)((ll)c->mem_total[n] - c->mem_left[n]), (double)c->mem_total[n]);
  double mem_s = mem_u < 0.7 ? (1.0 - mem_u) * 100 : (1.0 - mem_u) * 50;
  double gpu_s = 0.0; /* int 0 in the source: adds the same */
  ll sum_left = 0;
  for (int g = 0; g < ng; ++g) sum_left += gml[g];
  if (p->ngpu > 0) {
    if (ng == 0) { py_raise = 1; return 0; } /* node.gpus[0]: IndexError */
    ll denom = (ll)c->gpu_left[n] * c->gpu_milli_total[n * 8];
    double gu = py_div((double)(denom - sum_left), (double)denom);
    gpu_s = gu < 0.7 ? (1.0 - gu) * 200 : (1.0 - gu) * 100;
  }
  double score = (cpu_s + mem_s) + gpu_s;
  if (p->ngpu > 0) {
    ll frag = py_imod(sum_left, p->gmilli);
    score -= (double)frag * 0.2;
  }
  if (c->cpu_total[n] < 2000 || c->mem_total[n] < 12) {
    score -= (double)(2000 - (ll)c->cpu_total[n]) * 0.01;
    score -= (double)(12 - (ll)c->mem_total[n]) * 0.1;
  }
  double bal = fabs(py_div((double)c->cpu_left[n], (double)imaxll(1, c->mem_left[n])) -
                    py_div((double)p->cpu, (double)imaxll(1, p->mem)));
  score -= bal * 0.5;
  if ((ll)c->cpu_left[n] > (ll)p->cpu * 2 && (ll)c->mem_left[n] > (ll)p->mem * 2) score += 25;
  if (p->ngpu > 0) {
    if (ng == 0) { py_raise = 1; return 0; } /* max() of an empty sequence */
    ll mx = gml[0], mn = gml[0];
    for (int g = 1; g < ng; ++g) { if (gml[g] > mx) mx = gml[g]; if (gml[g] < mn) mn = gml[g]; }
    score -= (double)(mx - mn) * 0.05;
  }
  if (c->cpu_total[n] > 10000 && c->mem_total[n] > 64) score += 15;
  if (cpu_u > 0.9 || mem_u > 0.9) score -= 20;
  return wrap_i(imaxll(1, py_int(score)));
}

static ll pol_4816(const Pod* p, const Cluster* c, int n) {
  if (!feasible(p, c, n)) return 0;
  const int* gml = c->gml + n * 8;
  const int* gmt = c->gpu_milli_total + n * 8;
  const int ng = c->ngpus[n];
  double cpu_u = py_div((double)((ll)c->cpu_total[n] - c->cpu_left[n] + p->cpu), (double)imaxll(1, c->cpu_total[n]));
  double mem_u = py_div((double)((ll)c->mem_total[n] - c->mem_left[n] + p->mem), (double)imaxll(1, c->mem_total[n]));
  double balance = 1 - fabs(cpu_u - mem_u);
  double eff = py_pow(cpu_u * mem_u, 0.5);
  double score;
  if (p->ngpu > 0) {
    int el[8], ne = 0;
    for (int g = 0; g < ng; ++g)
      if (gml[g] >= p->gmilli) el[ne++] = g;
    if (ne > p->ngpu) ne = p->ngpu; /* [:pod.num_gpu] */
    ll s_used = 0, s_tot = 0, s_sq = 0, s_left = 0;
    for (int k = 0; k < ne; ++k) {
      int g = el[k];
      s_used += (ll)gmt[g] - gml[g] + p->gmilli;
      s_tot += gmt[g];
      ll d = (ll)gml[g] - p->gmilli;
      s_sq += d * d;
      s_left += gml[g];
    }
    double gpu_u = py_div((double)s_used, (double)imaxll(1, s_tot));
    double gpu_frag = py_div((double)s_sq, (double)imaxll(1, s_left));
    double iso = 0.5 - fabs(0.5 - py_pow(gpu_frag, 0.5));
    score = ((((((cpu_u * 0.25 + mem_u * 0.15) + gpu_u * 0.45) + balance * 0.05) + eff * 0.05) - gpu_frag * 0.05) + iso * 0.1) * 10000;
  } else {
    double f1 = py_div((double)py_imod(c->cpu_left[n], imaxll(1, p->cpu)), (double)c->cpu_total[n]);
    double f2 = py_div((double)py_imod(c->mem_left[n], imaxll(1, p->mem)), (double)c->mem_total[n]);
    double frag = dmin2(f1, f2);
    score = ((((cpu_u * 0.45 + mem_u * 0.35) + balance * 0.1) + eff * 0.1) - frag * 0.1) * 10000;
  }
  return wrap_i(imaxll(1, py_int(score)));
}

static ll pol_4800(const Pod* p, const Cluster* c, int n) {
  if (!feasible(p, c, n)) return 0;
  const int* gml = c->gml + n * 8;
  const int* gmt = c->gpu_milli_total + n * 8;
  const int ng = c->ngpus[n];
  double cpu_u = py_div((double)((ll)c->cpu_total[n] - c->cpu_left[n] + p->cpu), (double)c->cpu_total[n]);
  double mem_u = py_div((double)((ll)c->mem_total[n] - c->mem_left[n] + p->mem), (double)c->mem_total[n]);
  double balance = py_pow(1 - fabs(cpu_u - mem_u), 2.5) * 300;
  double gpu_score = 0.0;
  if (p->ngpu > 0) {
    int v[8], nv = 0;
    for (int g = 0; g < ng; ++g)
      if (gml[g] >= p->gmilli) v[nv++] = g;
    for (int a = 1; a < nv; ++a) { /* stable sort by gpu_milli_left */
      int e = v[a], b = a - 1;
      while (b >= 0 && gml[e] < gml[v[b]]) { v[b + 1] = v[b]; --b; }
      v[b + 1] = e;
    }
    if (nv >= p->ngpu) {
      FSum s = {0, 0.0, 0.0};
      for (int k = 0; k < p->ngpu; ++k) {
        int g = v[k];
        fsum_add(&s, 1 - py_div((double)((ll)gml[g] - p->gmilli), (double)gmt[g]));
      }
      double gpu_eff = py_div(fsum_result(&s), (double)p->ngpu);
      gpu_score = py_pow(gpu_eff, 2.0) * 450;
    }
  }
  ll slack = iminll((ll)c->cpu_left[n] - p->cpu, (ll)c->mem_left[n] - p->mem);
  double frag = py_div(py_pow((double)slack, 0.6), (double)imaxll(c->cpu_total[n], c->mem_total[n])) * 300;
  double util = (dmin2(cpu_u, mem_u) * 0.6 + dmax2(cpu_u, mem_u) * 0.4) * 600;
  double score = ((util + balance) + gpu_score) + frag;
  return wrap_i(imaxll(1, py_int(score)));
}

static ll score_node(const Policy* pol, const Pod* p, const Cluster* c, int n) {
  switch (pol->id) {
    case 0: return pol_first_fit(p, c, n);
    case 1: return pol_best_fit(p, c, n);
    case 2: return pol_4901(p, c, n);
    case 3: return pol_4816(p, c, n);
    case 4: return pol_4800(p, c, n);
    case 5: return pol_worst_fit(p, c, n);
    case 6: return pol_random_family(p, c, n, pol->p);
    default: py_raise = 1; return 0;
  }
}

/* ---- heapq, literally ---------------------------------------------------------------------------------- */
typedef struct { u64* a; ll n; } Heap;
static void siftdown(u64* h, ll startpos, ll pos) { /* heapq.py:207-219 */
  u64 item = h[pos];
  while (pos > startpos) {
    ll parent = (pos - 1) >> 1;
    if (item < h[parent]) { h[pos] = h[parent]; pos = parent; continue; }
    break;
  }
  h[pos] = item;
}
static void siftup(u64* h, ll n, ll pos) { /* heapq.py:260-278: bubble the smaller child up, then sift down */
  ll startpos = pos;
  u64 item = h[pos];
  ll child = 2 * pos + 1;
  while (child < n) {
    ll right = child + 1;
    if (right < n && !(h[child] < h[right])) child = right;
    h[pos] = h[child];
    pos = child;
    child = 2 * pos + 1;
  }
  h[pos] = item;
  siftdown(h, startpos, pos);
}
static void heap_push(Heap* H, u64 item) { H->a[H->n++] = item; siftdown(H->a, 0, H->n - 1); }
static u64 heap_pop(Heap* H) {
  u64 last = H->a[--H->n];
  if (H->n > 0) { u64 top = H->a[0]; H->a[0] = last; siftup(H->a, H->n, 0); return top; }
  return last;
}

/* ---- outputs --------------------------------------------------------------------------------------------- */
typedef struct {
  int status, placed, num_snapshots, num_frag, max_nodes, max_waiting;
  ll events, attempts, last_event_time, scorer_calls;
  double avg_cpu, avg_mem, avg_gpu_count, avg_gpu_milli, fragmentation, score;
} OracleResult;

/* exact-enough mean: long double accumulation (the series have < 1e3 terms of magnitude <= 1) */
static double mean_ld(const double* v, ll n) {
  long double s = 0;
  for (ll i = 0; i < n; ++i) s += v[i];
  return (double)(s / (long double)n);
}

#define KEY(t, r, d) (((u64)(t) << 24) | ((u64)(r) << 1) | (u64)(d))

/* One episode.  Pods in pod_list order; id_rank[i] = rank of pod i's id in string order.
 * out_time/out_node/out_mask: per pod (pod_list order): final creation time, node index (-1 unplaced), GPU bitmask. */
int ks_oracle_run(int n_nodes, const int* cpu_total, const int* mem_total, const int* ngpus, const int* gpu_milli_total,
                  ll n_pods, const int* cpu, const int* mem, const int* ngpu, const int* gmilli, const ll* ctime,
                  const ll* dur, const int* id_rank, int policy_id, const double* policy_params, double snapshot_interval,
                  ll max_events, OracleResult* out, ll* out_time, int* out_node, int* out_mask) {
  memset(out, 0, sizeof *out);
  Cluster C;
  C.n = n_nodes; C.cpu_total = cpu_total; C.mem_total = mem_total; C.ngpus = ngpus; C.gpu_milli_total = gpu_milli_total;
  C.cpu_left = malloc(sizeof(int) * n_nodes); C.mem_left = malloc(sizeof(int) * n_nodes);
  C.gpu_left = malloc(sizeof(int) * n_nodes); C.gml = calloc((size_t)n_nodes * 8, sizeof(int));
  ll tot_cpu = 0, tot_mem = 0, tot_gc = 0, tot_gm = 0;
  for (int i = 0; i < n_nodes; ++i) {
    C.cpu_left[i] = cpu_total[i]; C.mem_left[i] = mem_total[i]; C.gpu_left[i] = ngpus[i];
    tot_cpu += cpu_total[i]; tot_mem += mem_total[i]; tot_gc += ngpus[i];
    for (int g = 0; g < ngpus[i]; ++g) { C.gml[i * 8 + g] = gpu_milli_total[i * 8 + g]; tot_gm += gpu_milli_total[i * 8 + g]; }
  }
  Policy pol; pol.id = policy_id;
  for (int k = 0; k < 4; ++k) pol.p[k] = policy_params ? policy_params[k] : 0.0;
  int* by_rank = malloc(sizeof(int) * (n_pods > 0 ? n_pods : 1));
  for (ll i = 0; i < n_pods; ++i) by_rank[id_rank[i]] = (int)i;
  Heap H; H.a = malloc(sizeof(u64) * (n_pods + 1)); H.n = 0;
  for (ll i = 0; i < n_pods; ++i) H.a[H.n++] = KEY(ctime[i], id_rank[i], 0); /* _populate_heap, pod_list order */
  for (ll i = H.n / 2 - 1; i >= 0; --i) siftup(H.a, H.n, i);                 /* heapq.heapify */
  char* waiting = calloc(n_pods > 0 ? n_pods : 1, 1);
  int* wl = malloc(sizeof(int) * (n_pods > 0 ? n_pods : 1));
  ll n_wait = 0;
  ll cap = 256; double *s_cpu = malloc(cap * 8), *s_mem = malloc(cap * 8), *s_gc = malloc(cap * 8), *s_gm = malloc(cap * 8);
  ll fcap = 1024; double* frag = malloc(fcap * 8);
  ll n_snap = 0, n_frag = 0, processed = 0;
  double thr = snapshot_interval;
  for (ll i = 0; i < n_pods; ++i) { out_time[i] = ctime[i]; out_node[i] = -1; out_mask[i] = 0; }
  int status = ST_OK;
  ll used_cpu = 0, used_mem = 0, used_gc = 0, used_gm = 0;
  int active = 0, max_nodes = 0;
  py_raise = 0;
  while (H.n > 0) {
    if (max_events && processed >= max_events) { status = ST_STEP_BUDGET; break; }
    u64 key = heap_pop(&H);
    ll now = (ll)(key >> 24);
    int pi = by_rank[(key >> 1) & 0x7FFFFF];
    out->last_event_time = now;
    Pod p = {cpu[pi], mem[pi], ngpu[pi], gmilli[pi], now, dur[pi]};
    if (key & 1) { /* deletion, main.py:74-99 */
      int nd = out_node[pi];
      int was = C.cpu_left[nd] < cpu_total[nd] || C.mem_left[nd] < mem_total[nd] || C.gpu_left[nd] < ngpus[nd];
      C.cpu_left[nd] += p.cpu; C.mem_left[nd] += p.mem; C.gpu_left[nd] += p.ngpu;
      for (int g = 0; g < 8; ++g) if ((out_mask[pi] >> g) & 1) { C.gml[nd * 8 + g] += p.gmilli; used_gm -= p.gmilli; }
      int is = C.cpu_left[nd] < cpu_total[nd] || C.mem_left[nd] < mem_total[nd] || C.gpu_left[nd] < ngpus[nd];
      active += is - was;
      used_cpu -= p.cpu; used_mem -= p.mem; used_gc -= p.ngpu;
    } else { /* creation, main.py:101-148 */
      out->attempts++;
      ll best = 0; int best_node = -1;
      for (int nd = 0; nd < n_nodes; ++nd) {
        out->scorer_calls++;
        ll s = score_node(&pol, &p, &C, nd);
        if (py_raise) break;
        if (s > best) { best = s; best_node = nd; }
      }
      if (py_raise) { status = ST_POLICY_EXCEPTION; break; }
      if (best_node < 0) {
        if (!waiting[pi]) { waiting[pi] = 1; wl[n_wait++] = pi; if (n_wait > out->max_waiting) out->max_waiting = (int)n_wait; }
        ll need = -1;
        for (ll k = 0; k < n_wait; ++k) if (ngpu[wl[k]] > 0 && (need < 0 || gmilli[wl[k]] < need)) need = gmilli[wl[k]];
        double fr = 0.0;
        if (need >= 0) {
          ll lost = 0;
          for (int nd = 0; nd < n_nodes; ++nd) for (int g = 0; g < ngpus[nd]; ++g) {
            int left = C.gml[nd * 8 + g];
            if (left > 0 && left < need) lost += left;
          }
          fr = tot_gm > 0 ? (double)lost / (double)tot_gm : 0.0;
        }
        if (n_frag == fcap) { fcap *= 2; frag = realloc(frag, fcap * 8); }
        frag[n_frag++] = fr;
        for (ll k = 0; k < H.n; ++k) /* first DELETION in array order */
          if (H.a[k] & 1) {
            ll nt = (ll)(H.a[k] >> 24) + 1;
            out_time[pi] = nt;
            heap_push(&H, KEY(nt, id_rank[pi], 0));
            break;
          }
      } else {
        int nd = best_node;
        int was = C.cpu_left[nd] < cpu_total[nd] || C.mem_left[nd] < mem_total[nd] || C.gpu_left[nd] < ngpus[nd];
        C.cpu_left[nd] -= p.cpu; C.mem_left[nd] -= p.mem; C.gpu_left[nd] -= p.ngpu;
        int mask = 0;
        if (p.ngpu != 0) { /* main.py:150-177 */
          int cand[8], nc = 0;
          for (int g = 0; g < ngpus[nd]; ++g) if (C.gml[nd * 8 + g] >= p.gmilli) cand[nc++] = g;
          if (nc < p.ngpu) { status = ST_GPU_SHORTFALL; break; }
          for (int a = 1; a < nc; ++a) { /* stable sort by gpu_milli_left */
            int e = cand[a], b = a - 1;
            while (b >= 0 && C.gml[nd * 8 + e] < C.gml[nd * 8 + cand[b]]) { cand[b + 1] = cand[b]; --b; }
            cand[b + 1] = e;
          }
          for (int k = 0; k < p.ngpu; ++k) { mask |= 1 << cand[k]; C.gml[nd * 8 + cand[k]] -= p.gmilli; used_gm += p.gmilli; }
        }
        int is = C.cpu_left[nd] < cpu_total[nd] || C.mem_left[nd] < mem_total[nd] || C.gpu_left[nd] < ngpus[nd];
        active += is - was;
        used_cpu += p.cpu; used_mem += p.mem; used_gc += p.ngpu;
        out_node[pi] = nd; out_mask[pi] = mask; out_time[pi] = now;
        if (waiting[pi]) {
          waiting[pi] = 0;
          for (ll k = 0; k < n_wait; ++k) if (wl[k] == pi) { wl[k] = wl[--n_wait]; break; }
        }
        out->placed++;
        heap_push(&H, KEY(now + p.dur, id_rank[pi], 1));
      }
    }
    processed++;
    double progress = n_pods > 0 ? (double)processed / (double)n_pods : 0.0;
    if (progress >= thr) {
      if (n_snap == cap) { cap *= 2; s_cpu = realloc(s_cpu, cap * 8); s_mem = realloc(s_mem, cap * 8); s_gc = realloc(s_gc, cap * 8); s_gm = realloc(s_gm, cap * 8); }
      s_cpu[n_snap] = tot_cpu > 0 ? (double)used_cpu / (double)tot_cpu : 0.0;
      s_mem[n_snap] = tot_mem > 0 ? (double)used_mem / (double)tot_mem : 0.0;
      s_gc[n_snap] = tot_gc > 0 ? (double)used_gc / (double)tot_gc : 0.0;
      s_gm[n_snap] = tot_gm > 0 ? (double)used_gm / (double)tot_gm : 0.0;
      n_snap++;
      thr += snapshot_interval;
    }
    if (active > max_nodes) max_nodes = active;
  }
  out->status = status; out->events = processed; out->num_snapshots = (int)n_snap; out->num_frag = (int)n_frag;
  out->max_nodes = max_nodes;
  if (n_snap > 0) {
    out->avg_cpu = mean_ld(s_cpu, n_snap); out->avg_mem = mean_ld(s_mem, n_snap);
    out->avg_gpu_count = mean_ld(s_gc, n_snap); out->avg_gpu_milli = mean_ld(s_gm, n_snap);
    out->fragmentation = n_frag > 0 ? mean_ld(frag, n_frag) : 0.0;
    if (status == ST_OK && out->placed == n_pods) {
      double overall = (out->avg_cpu + out->avg_mem + out->avg_gpu_count + out->avg_gpu_milli) / 4.0;
      double pen = out->fragmentation < 0.1 ? out->fragmentation : 0.1;
      double sc = overall - pen;
      out->score = sc < 0.0 ? 0.0 : (sc > 1.0 ? 1.0 : sc);
    }
  }
  free(C.cpu_left); free(C.mem_left); free(C.gpu_left); free(C.gml); free(by_rank); free(H.a); free(waiting); free(wl);
  free(s_cpu); free(s_mem); free(s_gc); free(s_gm); free(frag);
  return 0;
}
