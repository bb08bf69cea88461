/*
 * oracle/tppr_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Plain-C, single-threaded CPU restatement of the reference's T-PPR algorithms, used only as
 * the checker in tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs.  The product path (zebra_b200/) never links, imports or calls anything in oracle/.
 *
 * What it restates (reference = /root/reference, Python + numba jitclass):
 *   - streaming top-k T-PPR ("SANTA")          utils/util.py:473-576  (zo_stream)
 *   - row extraction                           utils/util.py:447-469  (emit_row)
 *   - update-only fill pass                    utils/util.py:787-873  (zo_stream with rows==0)
 *   - pruned power iteration query ("PPI")     utils/util.py:185-276  (zo_pruned_topk)
 *   - time-sorted adjacency prefix search      utils/util.py:152-154  (lower_bound)
 *
 * Third-party behaviour the reference inherits from numba 0.65.0 / llvmlite 0.47.0 (not vendored
 * in /root/reference, no requirements file pins it; version = the one installed in this image):
 *   - typed.Dict iterates in insertion order, assignment to an existing key keeps its position;
 *     restated here as a small ordered array with linear key search (<= 2k+1 entries).
 *   - np.argsort on a float64 array = numba.misc.quicksort (median-of-three quicksort, insertion
 *     sort below 15 elements, explicit stack, larger partition pushed); restated in
 *     numba_argsort() from its published algorithm so that the surviving entry among exactly
 *     tied scores and the resulting slot order match the reference bit for bit.
 *   - arithmetic is IEEE fp64 with separate multiply/add roundings (numba emits no FMA for this
 *     code): build with -ffp-contract=off.
 *
 * Parity pinning: the reference ships no tests or golden vectors.  This oracle is pinned against
 * the reference itself, executed in the build container (tests/golden/make_golden.py imports
 * /root/reference/utils/util.py under numba and records inputs/outputs to tests/golden/ as .npz files);
 * tests/test_oracle.py replays those fixtures through this file.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

typedef struct {
    int64_t num_nodes;
    int32_t k;
    int32_t *cnt;    /* [N]   entries in use (dict length) */
    int64_t *eidx;   /* [N*k] key part 0 */
    int64_t *node;   /* [N*k] key part 1 */
    double  *ts;     /* [N*k] key part 2 */
    double  *score;  /* [N*k] value */
    double  *norm;   /* [N]   sum_{z=1..deg} beta^z ; 0 <=> never seen */
} zo_table;

typedef struct {
    int64_t prunes;          /* updates where |T| > k */
    int64_t prunes_with_tie; /* ... and two candidates had exactly equal scores */
    int64_t boundary_ties;   /* ... and the k-th and (k+1)-th largest were equal */
    int64_t relevant_ties;   /* ... and a tied pair includes a survivor (order/selection depends on the sort) */
} zo_stats;

static zo_stats g_stats;

zo_table *zo_table_new(int64_t num_nodes, int32_t k)
{
    zo_table *t = (zo_table *)calloc(1, sizeof(zo_table));
    t->num_nodes = num_nodes;
    t->k = k;
    t->cnt = (int32_t *)calloc((size_t)num_nodes, sizeof(int32_t));
    t->eidx = (int64_t *)calloc((size_t)num_nodes * k, sizeof(int64_t));
    t->node = (int64_t *)calloc((size_t)num_nodes * k, sizeof(int64_t));
    t->ts = (double *)calloc((size_t)num_nodes * k, sizeof(double));
    t->score = (double *)calloc((size_t)num_nodes * k, sizeof(double));
    t->norm = (double *)calloc((size_t)num_nodes, sizeof(double));
    return t;
}

void zo_table_free(zo_table *t)
{
    if (!t) return;
    free(t->cnt); free(t->eidx); free(t->node); free(t->ts); free(t->score); free(t->norm);
    free(t);
}

zo_table *zo_table_clone(const zo_table *s)
{
    zo_table *t = zo_table_new(s->num_nodes, s->k);
    size_t n = (size_t)s->num_nodes, nk = n * (size_t)s->k;
    memcpy(t->cnt, s->cnt, n * sizeof(int32_t));
    memcpy(t->eidx, s->eidx, nk * sizeof(int64_t));
    memcpy(t->node, s->node, nk * sizeof(int64_t));
    memcpy(t->ts, s->ts, nk * sizeof(double));
    memcpy(t->score, s->score, nk * sizeof(double));
    memcpy(t->norm, s->norm, n * sizeof(double));
    return t;
}

int32_t zo_table_get(const zo_table *t, int64_t v, int64_t *eidx, int64_t *node, double *ts,
                     double *score, double *norm)
{
    int32_t c = t->cnt[v];
    size_t o = (size_t)v * t->k;
    for (int32_t j = 0; j < c; ++j) {
        eidx[j] = t->eidx[o + j]; node[j] = t->node[o + j];
        ts[j] = t->ts[o + j]; score[j] = t->score[o + j];
    }
    *norm = t->norm[v];
    return c;
}

void zo_stats_get(int64_t *out4)
{
    out4[0] = g_stats.prunes; out4[1] = g_stats.prunes_with_tie; out4[2] = g_stats.boundary_ties;
    out4[3] = g_stats.relevant_ties;
}
void zo_stats_reset(void) { memset(&g_stats, 0, sizeof g_stats); }

/* ---- numba.misc.quicksort restated (argsort flavour, lt = a < b, no NaNs on this path) ---- */

#define ZO_SMALL 15
#define ZO_SWAP(a, b) do { int32_t _t = (a); (a) = (b); (b) = _t; } while (0)

static void ins_sort(const double *A, int32_t *R, int low, int high)
{
    for (int i = low + 1; i <= high; ++i) {
        int32_t kk = R[i];
        double v = A[kk];
        int j = i;
        while (j > low && v < A[R[j - 1]]) { R[j] = R[j - 1]; --j; }
        R[j] = kk;
    }
}

static int qs_partition(const double *A, int32_t *R, int low, int high)
{
    int mid = (low + high) >> 1;
    if (A[R[mid]] < A[R[low]]) ZO_SWAP(R[low], R[mid]);
    if (A[R[high]] < A[R[mid]]) ZO_SWAP(R[high], R[mid]);
    if (A[R[mid]] < A[R[low]]) ZO_SWAP(R[low], R[mid]);
    double pivot = A[R[mid]];
    ZO_SWAP(R[high], R[mid]);
    int i = low, j = high - 1;
    for (;;) {
        while (i < high && A[R[i]] < pivot) ++i;
        while (j >= low && pivot < A[R[j]]) --j;
        if (i >= j) break;
        ZO_SWAP(R[i], R[j]);
        ++i; --j;
    }
    ZO_SWAP(R[i], R[high]);
    return i;
}

void zo_numba_argsort(const double *A, int32_t n, int32_t *R)
{
    for (int32_t i = 0; i < n; ++i) R[i] = i;
    if (n < 2) return;
    int lo_stack[100], hi_stack[100];
    int sp = 0;
    lo_stack[0] = 0; hi_stack[0] = n - 1; sp = 1;
    while (sp > 0) {
        --sp;
        int low = lo_stack[sp], high = hi_stack[sp];
        while (high - low >= ZO_SMALL) {
            int i = qs_partition(A, R, low, high);
            if (high - i > i - low) {
                if (high > i) { lo_stack[sp] = i + 1; hi_stack[sp] = high; ++sp; }
                high = i - 1;
            } else {
                if (i > low) { lo_stack[sp] = low; hi_stack[sp] = i - 1; ++sp; }
                low = i + 1;
            }
        }
        ins_sort(A, R, low, high);
    }
}

/* stable ascending argsort: ties keep insertion order (the documented simplification) */
static void stable_argsort(const double *A, int32_t n, int32_t *R)
{
    for (int32_t i = 0; i < n; ++i) R[i] = i;
    ins_sort(A, R, 0, n - 1);
}

/* ---- one endpoint update: new vector of s1 from the OLD vectors of s1 and s2 ---- */

#define ZO_MAXC 4100 /* 2k+1 candidates; k <= 2049 */

typedef struct {
    int32_t n;
    int64_t eidx[ZO_MAXC];
    int64_t node[ZO_MAXC];
    double ts[ZO_MAXC];
    double score[ZO_MAXC];
} zo_vec;

static int find_key(const zo_vec *T, int64_t e, int64_t nd, double ts)
{
    for (int32_t j = 0; j < T->n; ++j)
        if (T->eidx[j] == e && T->node[j] == nd && T->ts[j] == ts) return j;
    return -1;
}

static void update_one(const zo_table *tb, double alpha, double beta, int64_t s1, int64_t s2,
                       int64_t edge_idx, double timestamp, int tie_mode, zo_vec *T, zo_vec *out)
{
    const int32_t k = tb->k;
    double scale_s2;
    T->n = 0;
    if (tb->norm[s1] == 0) {                       /* util.py:514-519 */
        scale_s2 = 1 - alpha;
    } else {                                       /* util.py:521-527 */
        double last_norm = tb->norm[s1];
        double new_norm = last_norm * beta + beta;
        double scale_s1 = last_norm / new_norm * beta;
        scale_s2 = beta / new_norm * (1 - alpha);
        size_t o = (size_t)s1 * k;
        for (int32_t j = 0; j < tb->cnt[s1]; ++j) {
            T->eidx[j] = tb->eidx[o + j]; T->node[j] = tb->node[o + j]; T->ts[j] = tb->ts[o + j];
            T->score[j] = tb->score[o + j] * scale_s1;
        }
        T->n = tb->cnt[s1];
    }
    if (tb->norm[s2] != 0) {                       /* util.py:533-538 */
        size_t o = (size_t)s2 * k;
        for (int32_t j = 0; j < tb->cnt[s2]; ++j) {
            int p = find_key(T, tb->eidx[o + j], tb->node[o + j], tb->ts[o + j]);
            double add = tb->score[o + j] * scale_s2;
            if (p >= 0) {
                T->score[p] = T->score[p] + add;
            } else {
                p = T->n++;
                T->eidx[p] = tb->eidx[o + j]; T->node[p] = tb->node[o + j]; T->ts[p] = tb->ts[o + j];
                T->score[p] = add;
            }
        }
    }
    {                                              /* util.py:531,540-541: assignment, not += */
        double w = (alpha != 0) ? scale_s2 * alpha : scale_s2;
        int p = find_key(T, edge_idx, s2, timestamp);
        if (p < 0) {
            p = T->n++;
            T->eidx[p] = edge_idx; T->node[p] = s2; T->ts[p] = timestamp;
        }
        T->score[p] = w;
    }
    if (T->n <= k) {                               /* util.py:549-551 */
        out->n = T->n;
        memcpy(out->eidx, T->eidx, (size_t)T->n * sizeof(int64_t));
        memcpy(out->node, T->node, (size_t)T->n * sizeof(int64_t));
        memcpy(out->ts, T->ts, (size_t)T->n * sizeof(double));
        memcpy(out->score, T->score, (size_t)T->n * sizeof(double));
        return;
    }
    /* util.py:553-559: k largest, re-inserted in ascending score order */
    int32_t R[ZO_MAXC];
    if (tie_mode == 0) zo_numba_argsort(T->score, T->n, R);
    else stable_argsort(T->score, T->n, R);
    g_stats.prunes++;
    {
        int32_t S[ZO_MAXC];
        stable_argsort(T->score, T->n, S);
        int tie = 0;
        for (int32_t j = 1; j < T->n; ++j) if (T->score[S[j]] == T->score[S[j - 1]]) { tie = 1; break; }
        if (tie) g_stats.prunes_with_tie++;
        if (T->score[S[T->n - k]] == T->score[S[T->n - k - 1]]) g_stats.boundary_ties++;
        for (int32_t j = T->n - k; j < T->n; ++j)
            if (T->score[S[j]] == T->score[S[j - 1]]) { g_stats.relevant_ties++; break; }
    }
    out->n = k;
    for (int32_t j = 0; j < k; ++j) {
        int32_t ind = R[T->n - k + j];
        out->eidx[j] = T->eidx[ind]; out->node[j] = T->node[ind]; out->ts[j] = T->ts[ind];
        out->score[j] = T->score[ind];
    }
}

static void commit(zo_table *tb, int64_t v, const zo_vec *nv)
{
    size_t o = (size_t)v * tb->k;
    for (int32_t j = 0; j < nv->n; ++j) {
        tb->eidx[o + j] = nv->eidx[j]; tb->node[o + j] = nv->node[j];
        tb->ts[o + j] = nv->ts[j]; tb->score[o + j] = nv->score[j];
    }
    tb->cnt[v] = nv->n;
}

/* util.py:447-469 — one output row in dict order; dt = f32(t_now - f64(f32(ts))), pads carry
 * dt = f32(t_now) when the vector is non-empty, an empty vector leaves the zero row untouched. */
static void emit_row(const zo_table *tb, int64_t v, double t_now, int64_t row, int32_t *o_node,
                     int32_t *o_eidx, float *o_dt, float *o_w)
{
    const int32_t k = tb->k;
    int32_t c = tb->cnt[v];
    if (c == 0) return;
    size_t o = (size_t)v * k, r = (size_t)row * k;
    for (int32_t j = 0; j < k; ++j) {
        float tsf = 0.0f;
        if (j < c) {
            o_node[r + j] = (int32_t)tb->node[o + j];
            o_eidx[r + j] = (int32_t)tb->eidx[o + j];
            o_w[r + j] = (float)tb->score[o + j];
            tsf = (float)tb->ts[o + j];
        } else {
            o_node[r + j] = 0; o_eidx[r + j] = 0; o_w[r + j] = 0.0f;
        }
        o_dt[r + j] = (float)(t_now - (double)tsf);
    }
}

/*
 * Streaming pass over n edges for one (alpha,beta) instance on one table.
 *   nodes: [rows_per_edge * n] = [src ; dst ; (neg)], rows_per_edge in {0 (update only: nodes
 *          holds [src;dst] and nothing is emitted), 2 (no fake), 3}.
 *   outputs: [rows_per_edge*n, k] arrays, pre-zeroed by the caller.
 */
void zo_stream(zo_table *tb, double alpha, double beta, const int64_t *nodes, const double *ts,
               const int64_t *eidx, int64_t n, int32_t rows_per_edge, int32_t tie_mode,
               int32_t *o_node, int32_t *o_eidx, float *o_dt, float *o_w)
{
    zo_vec *T = (zo_vec *)malloc(sizeof(zo_vec));
    zo_vec *n1 = (zo_vec *)malloc(sizeof(zo_vec));
    zo_vec *n2 = (zo_vec *)malloc(sizeof(zo_vec));
    for (int64_t i = 0; i < n; ++i) {
        int64_t source = nodes[i], target = nodes[i + n];
        double timestamp = ts[i];
        int64_t edge_idx = eidx[i];
        if (rows_per_edge >= 2) {                  /* util.py:504-506: read BEFORE update */
            emit_row(tb, source, timestamp, i, o_node, o_eidx, o_dt, o_w);
            emit_row(tb, target, timestamp, i + n, o_node, o_eidx, o_dt, o_w);
            if (rows_per_edge == 3)
                emit_row(tb, nodes[i + 2 * n], timestamp, i + 2 * n, o_node, o_eidx, o_dt, o_w);
        }
        update_one(tb, alpha, beta, source, target, edge_idx, timestamp, tie_mode, T, n1);
        if (source != target)
            update_one(tb, alpha, beta, target, source, edge_idx, timestamp, tie_mode, T, n2);
        commit(tb, source, n1);                    /* util.py:567-574 */
        tb->norm[source] = tb->norm[source] * beta + beta;
        if (source != target) {
            commit(tb, target, n2);
            tb->norm[target] = tb->norm[target] * beta + beta;
        }
    }
    free(T); free(n1); free(n2);
}

/* ---- pruned power iteration (util.py:185-276) over a CSR time-sorted adjacency ---- */

static int64_t lower_bound(const double *a, int64_t n, double x)
{   /* np.searchsorted(a, x) side='left': entries strictly earlier than x (util.py:152-154) */
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (a[mid] < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

typedef struct { int64_t node; double ts; double w; } zo_query;

/* numba lowers pow(float64, int64) to exponentiation by squaring (numba/cpython/numbers.py
 * int_power_impl), not libm pow: verified against numba 0.65.0 for beta=0.95, n=1..399. */
static double pow_int(double a, int64_t n)
{
    /* beyond 0x10000 numba's int_power_impl hands over to libm: math.pow(a, float(n)) (checked against
     * numba 0.65.0: beta = 0.999 / 0.9999, n = 65536 squaring, n = 65537, 70000, 200000 libm) */
    if (n > 0x10000) return pow(a, (double)n);
    double r = 1.0;
    while (n != 0) {
        if (n & 1) r *= a;
        n >>= 1;
        a *= a;
    }
    return r;
}

void zo_pruned_topk(const int64_t *indptr, const int32_t *adj_node, const int32_t *adj_eidx,
                    const double *adj_ts, const int32_t *q_nodes, const double *q_ts, int64_t nq,
                    int32_t width, int32_t depth, double alpha, double beta, int32_t k,
                    int32_t tie_mode, int32_t *o_node, int32_t *o_eidx, float *o_dt, float *o_w)
{
    size_t cap = 1;
    for (int d = 0; d < depth; ++d) cap *= (size_t)width;
    size_t dict_cap = 0, lvl = 1;
    for (int d = 0; d < depth; ++d) { lvl *= (size_t)width; dict_cap += lvl; }
    zo_query *cur = (zo_query *)malloc((cap + 1) * sizeof(zo_query));
    zo_query *nxt = (zo_query *)malloc((cap + 1) * sizeof(zo_query));
    int64_t *d_e = (int64_t *)malloc((dict_cap + 1) * sizeof(int64_t));
    int64_t *d_n = (int64_t *)malloc((dict_cap + 1) * sizeof(int64_t));
    double *d_t = (double *)malloc((dict_cap + 1) * sizeof(double));
    double *d_w = (double *)malloc((dict_cap + 1) * sizeof(double));
    int32_t *R = (int32_t *)malloc((dict_cap + 1) * sizeof(int32_t));
    for (int64_t i = 0; i < nq; ++i) {
        size_t dn = 0, ncur = 1;
        cur[0].node = q_nodes[i]; cur[0].ts = q_ts[i]; cur[0].w = 1.0;
        for (int dep = 0; dep < depth; ++dep) {
            size_t nn = 0;
            for (size_t q = 0; q < ncur; ++q) {
                int64_t v = cur[q].node;
                int64_t base = indptr[v];
                int64_t n_ngh = lower_bound(adj_ts + base, indptr[v + 1] - base, cur[q].ts);
                if (n_ngh == 0) continue;
                double norm = beta / (1 - beta) * (1 - pow_int(beta, n_ngh));
                double weight = (alpha != 0 && dep == 0)
                                    ? cur[q].w * (1 - alpha) * beta / norm * alpha
                                    : cur[q].w * (1 - alpha) * beta / norm;
                int64_t lim = n_ngh < width ? n_ngh : width;
                for (int64_t z = 0; z < lim; ++z) {
                    int64_t p = base + n_ngh - (z + 1);
                    int64_t e = adj_eidx[p], nd = adj_node[p];
                    double t = adj_ts[p];
                    size_t f = 0;
                    for (; f < dn; ++f)
                        if (d_e[f] == e && d_n[f] == nd && d_t[f] == t) break;
                    if (f < dn) d_w[f] = d_w[f] + weight;
                    else { d_e[dn] = e; d_n[dn] = nd; d_t[dn] = t; d_w[dn] = weight; ++dn; }
                    nxt[nn].node = nd; nxt[nn].ts = t; nxt[nn].w = weight; ++nn;
                    weight = weight * beta;
                }
            }
            if (nn == 0) break;
            zo_query *tmp = cur; cur = nxt; nxt = tmp;
            ncur = nn;
        }
        if (dn == 0) continue;
        double t_now = q_ts[i];
        size_t r = (size_t)i * k;
        int32_t m = dn <= (size_t)k ? (int32_t)dn : k;
        if (dn <= (size_t)k) for (int32_t j = 0; j < m; ++j) R[j] = j;
        else {
            if (tie_mode == 0) zo_numba_argsort(d_w, (int32_t)dn, R);
            else stable_argsort(d_w, (int32_t)dn, R);
            for (int32_t j = 0; j < k; ++j) R[j] = R[dn - k + j];
        }
        for (int32_t j = 0; j < k; ++j) {
            float tsf = 0.0f;
            if (j < m) {
                int32_t ind = R[j];
                o_node[r + j] = (int32_t)d_n[ind]; o_eidx[r + j] = (int32_t)d_e[ind];
                o_w[r + j] = (float)d_w[ind]; tsf = (float)d_t[ind];
            } else { o_node[r + j] = 0; o_eidx[r + j] = 0; o_w[r + j] = 0.0f; }
            o_dt[r + j] = (float)(t_now - (double)tsf);
        }
    }
    free(cur); free(nxt); free(d_e); free(d_n); free(d_t); free(d_w); free(R);
}
