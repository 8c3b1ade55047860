/*
 * CPU oracle for the K-hop propagation (TEST INFRASTRUCTURE ONLY -- never linked into or
 * called from the product library; see oracle/__init__.py).
 *
 * Restates /root/reference/utils.py:92-97 (sgc_precompute: K x torch.spmm(adj, features)).
 * The arithmetic lives in a third-party dependency absent from /root/reference (torch, pinned
 * torch==1.6.0 in requirements.txt:9).  Its CPU sparse-dense kernel was established by probe
 *  to be, for every output row i and feature f,
 *
 *     acc = 0.0f;  for p in row i, ascending column order:  acc = fmaf(val[p], X[col[p]][f], acc)
 *
 * i.e. one fused multiply-add per stored entry, sequential, parallel across rows only.  This
 * file is that loop in plain C.  Pinned against torch.spmm outputs of the unmodified reference
 * in tests/golden/ (sha256 of the full Y for Cora / Citeseer / Pubmed-graph).
 *
 * Build: see oracle/Makefile (gcc -O3 -pthread -ffp-contract=off -mfma).
 */
#include <math.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

/* Minimal parallel-for over rows (libgomp is not in the image): worker threads pull blocks of
 * `grain` iterations from an atomic counter.  Parallelism is across rows only, exactly like the
 * reference's CPU kernel, so results do not depend on the thread count. */
typedef void (*range_fn)(int64_t begin, int64_t end, void *ctx);
typedef struct { range_fn fn; void *ctx; int64_t n, grain; atomic_llong next; } pf_t;

static void *pf_worker(void *arg)
{
    pf_t *pf = (pf_t *)arg;
    for (;;) {
        int64_t b = atomic_fetch_add(&pf->next, pf->grain);
        if (b >= pf->n) break;
        int64_t e = b + pf->grain < pf->n ? b + pf->grain : pf->n;
        pf->fn(b, e, pf->ctx);
    }
    return NULL;
}

static int oracle_threads(void)
{
    const char *env = getenv("SGC_ORACLE_THREADS");
    long t = env ? atol(env) : sysconf(_SC_NPROCESSORS_ONLN);
    if (t < 1) t = 1;
    if (t > 256) t = 256;
    return (int)t;
}

static void parallel_for(int64_t n, int64_t grain, range_fn fn, void *ctx)
{
    pf_t pf = { fn, ctx, n, grain, 0 };
    int nt = oracle_threads();
    if (n <= grain || nt == 1) { fn(0, n, ctx); return; }
    pthread_t th[256];
    int started = 0;
    for (int t = 0; t < nt - 1; ++t)
        if (pthread_create(&th[started], NULL, pf_worker, &pf) == 0) ++started;
    pf_worker(&pf);
    for (int t = 0; t < started; ++t) pthread_join(th[t], NULL);
}

/* One hop: Y[n_rows, f] = S[n_rows, *] * X[*, f].  CSR with int64 row_ptr, int32 col. */
typedef struct {
    const int64_t *row_ptr; const int32_t *col; const float *val; const float *x; int64_t ldx;
    const float *mean, *std; float *y; int64_t ldy, f;
} spmm_ctx;

static void spmm_rows(int64_t begin, int64_t end, void *vctx)
{
    const spmm_ctx *c = (const spmm_ctx *)vctx;
    const int64_t f = c->f;
    for (int64_t i = begin; i < end; ++i) {
        float *restrict yi = c->y + i * c->ldy;
        memset(yi, 0, (size_t)f * sizeof(float));
        for (int64_t p = c->row_ptr[i]; p < c->row_ptr[i + 1]; ++p) {
            const float v = c->val[p];
            const float *restrict xr = c->x + (int64_t)c->col[p] * c->ldx;
            for (int64_t j = 0; j < f; ++j)
                yi[j] = fmaf(v, xr[j], yi[j]);
        }
    }
}

void sgc_oracle_spmm_f32(int64_t n_rows, const int64_t *row_ptr, const int32_t *col,
                         const float *val, const float *x, int64_t ldx, float *y, int64_t ldy,
                         int64_t f)
{
    spmm_ctx c = { row_ptr, col, val, x, ldx, NULL, NULL, y, ldy, f };
    parallel_for(n_rows, 64, spmm_rows, &c);
}

/* Same hop with the first-hop feature standardisation of utils.py:119 applied to the gathered
 * operand exactly as the reference materialises it: xs = (x - mean[f]) / std[f] in fp32. */
static void spmm_std_rows(int64_t begin, int64_t end, void *vctx)
{
    const spmm_ctx *c = (const spmm_ctx *)vctx;
    const int64_t f = c->f;
    for (int64_t i = begin; i < end; ++i) {
        float *restrict yi = c->y + i * c->ldy;
        memset(yi, 0, (size_t)f * sizeof(float));
        for (int64_t p = c->row_ptr[i]; p < c->row_ptr[i + 1]; ++p) {
            const float v = c->val[p];
            const float *restrict xr = c->x + (int64_t)c->col[p] * c->ldx;
            for (int64_t j = 0; j < f; ++j) {
                const float xs = (xr[j] - c->mean[j]) / c->std[j];
                yi[j] = fmaf(v, xs, yi[j]);
            }
        }
    }
}

void sgc_oracle_spmm_standardized_f32(int64_t n_rows, const int64_t *row_ptr, const int32_t *col,
                                      const float *val, const float *x, int64_t ldx,
                                      const float *mean, const float *std, float *y, int64_t ldy,
                                      int64_t f)
{
    spmm_ctx c = { row_ptr, col, val, x, ldx, mean, std, y, ldy, f };
    parallel_for(n_rows, 64, spmm_std_rows, &c);
}

/* fp64 r^-1/2 the way the product library computes it on the GPU (1/sqrt(r) followed by one
 * FMA-exact Newton correction).  Kept here so the CPU suite can measure how often it differs
 * from libm's pow(r, -0.5), which is what normalization.py:9 calls through numpy. */
double sgc_oracle_rsqrt_refined(double r)
{
    if (!(r > 0.0)) {
        if (r == 0.0) return 0.0;        /* pow(0,-0.5) = inf -> 0 (normalization.py:10) */
        return NAN;                      /* negative or NaN row sum -> NaN, as pow does   */
    }
    if (isinf(r)) return 0.0;
    double y0 = 1.0 / sqrt(r);
    double t = y0 * y0;
    double h = fma(y0, y0, -t);          /* y0^2 = t + h exactly */
    double e = fma(r, t, -1.0) + r * h;  /* r*y0^2 - 1 */
    return fma(-0.5 * y0, e, y0);
}

void sgc_oracle_rsqrt_refined_array(int64_t n, const double *r, double *out)
{
    for (int64_t i = 0; i < n; ++i) out[i] = sgc_oracle_rsqrt_refined(r[i]);
}

/* Column mean and unbiased std in fp64 (utils.py:119 statistics, exact-value reference). */
typedef struct { int64_t n; const float *x; int64_t ldx; double *mean, *std; } stats_ctx;

static void stats_cols(int64_t begin, int64_t end, void *vctx)
{
    const stats_ctx *c = (const stats_ctx *)vctx;
    const int64_t n = c->n, ldx = c->ldx;
    const float *x = c->x;
    double *mean = c->mean, *std = c->std;
    for (int64_t j = begin; j < end; ++j) {
        double s = 0.0;
        for (int64_t i = 0; i < n; ++i) s += (double)x[i * ldx + j];
        double m = s / (double)n;
        double q = 0.0;
        for (int64_t i = 0; i < n; ++i) {
            double dlt = (double)x[i * ldx + j] - m;
            q += dlt * dlt;
        }
        mean[j] = m;
        std[j] = sqrt(q / (double)(n - 1));
    }
}

void sgc_oracle_col_stats(int64_t n, int64_t f, const float *x, int64_t ldx, double *mean,
                          double *std)
{
    stats_ctx c = { n, x, ldx, mean, std };
    parallel_for(f, 8, stats_cols, &c);
}
