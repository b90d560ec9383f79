/*
 * cpu_baseline.c -- TEST / BENCH INFRASTRUCTURE ONLY (never on the product path).
 *
 * The reference's CPU implementation of the KKT linear solve, as its QDLDL slot
 * (include/qdldl/sqp.cuh) runs it per SQP iteration:
 *
 *   mode "qdldl"  qdldl_solve_schur (include/qdldl/sqp.cuh:22-49): QDLDL_factor -> x = b ->
 *                 QDLDL_solve on PRECOMPUTED CSC values and gamma (the h_val / h_gamma the
 *                 reference copies back from the GPU, :268-271), QDLDL_etree once (:193).
 *                 This is the whole of the reference-authored CPU work on the path: the
 *                 reference forms S on the GPU (form_schur_qdl_kernel).
 *   mode "kkt"    the same preceded by the oracle's restatement of that Schur formation
 *                 (oracle_schur_f32 with Pinv = NULL: S, gamma, G^-1, no preconditioner) and
 *                 the CSC packing -- an end-to-end CPU figure, labelled a restatement.
 *
 * Built by oracle/build.py into oracle/_ref/libcpu_baseline.so together with the REFERENCE's
 * own qdldl/src/qdldl.c (compiled from /root/reference where it lies, float/int
 * configuration of the reference Makefile:16).
 *
 * Threads: a persistent pthread pool (created once, reused by every step): each thread keeps
 * its QDLDL workspace, the CSC pattern and the elimination tree for the pool's lifetime;
 * instances are split into contiguous slabs (SURVEY 8(d) "CPU baseline timing" (2)).
 */
#define _POSIX_C_SOURCE 200809L
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "qdldl.h"

void oracle_schur_f32(int s, int c, int N, float *G, const float *C, const float *g,
                      const float *cv, float rho, float *S, float *Pinv, float *gamma);
void oracle_csc_values_f32(int s, int N, const float *S, float *val);
void oracle_csc_pattern(int s, int N, int *col_ptr, int *row_ind);

static double now_s(void)
{
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);   /* include/utils/experiment.cuh:14 */
    return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec;
}

typedef struct pool pool_t;

typedef struct {
    pool_t *pool;
    int id;
    pthread_t th;
    /* per-thread QDLDL state, alive for the pool's lifetime */
    QDLDL_int *Ap, *Ai, *etree, *Lnz, *Lp, *Li, *iwork;
    QDLDL_float *Ax, *Lx, *D, *Dinv, *fwork;
    QDLDL_bool *bwork;
    float *Gw, *S, *gamma;
    /* the current job's slab */
    int first, count;
    double t_schur, t_qdldl;
    int status;
} worker_t;

struct pool {
    int s, c, N, nthreads, nnz;
    worker_t *w;
    pthread_mutex_t mu;
    pthread_cond_t go, done;
    long generation;
    int pending, quit;
    /* the current job */
    int mode;                      /* 0: qdldl on precomputed (val, gamma); 1: Schur restatement + CSC + qdldl;
                                      2: Schur restatement + CSC only (untimed preparation of mode 0's inputs) */
    const float *G, *C, *g, *cv, *val, *gamma_in;
    float rho;
    float *lambda;
    float *val_out, *gamma_out;
};

static void run_slab(worker_t *W)
{
    pool_t *P = W->pool;
    const int s = P->s, c = P->c, N = P->N, An = s * N, ss = s * s, cc = c * c;
    const size_t nG = (size_t)(ss + cc) * N - cc, nC = (size_t)(ss + s * c) * (N - 1);
    const size_t ng = (size_t)(s + c) * N - c, nc = (size_t)s * N;
    int b, i;
    W->t_schur = W->t_qdldl = 0.0;
    for (b = W->first; b < W->first + W->count; b++) {
        double t0 = now_s(), t1;
        const QDLDL_float *Ax;
        const float *gb;
        if (P->mode == 2) {
            memcpy(W->Gw, P->G + (size_t)b * nG, sizeof(float) * nG);
            oracle_schur_f32(s, c, N, W->Gw, P->C + (size_t)b * nC, P->g + (size_t)b * ng,
                             P->cv + (size_t)b * nc, P->rho, W->S, NULL, P->gamma_out + (size_t)b * nc);
            oracle_csc_values_f32(s, N, W->S, P->val_out + (size_t)b * P->nnz);
            continue;
        }
        if (P->mode == 1) {
            memcpy(W->Gw, P->G + (size_t)b * nG, sizeof(float) * nG);
            oracle_schur_f32(s, c, N, W->Gw, P->C + (size_t)b * nC, P->g + (size_t)b * ng,
                             P->cv + (size_t)b * nc, P->rho, W->S, NULL, W->gamma);
            oracle_csc_values_f32(s, N, W->S, W->Ax);
            Ax = W->Ax; gb = W->gamma;
        } else {
            Ax = P->val + (size_t)b * P->nnz; gb = P->gamma_in + (size_t)b * nc;
        }
        t1 = now_s();
        /* qdldl_solve_schur, include/qdldl/sqp.cuh:22-49 */
        if (QDLDL_factor(An, W->Ap, W->Ai, Ax, W->Lp, W->Li, W->Lx, W->D, W->Dinv, W->Lnz, W->etree,
                         W->bwork, W->iwork, W->fwork) < 0)
            W->status = -2;
        {
            float *x = P->lambda + (size_t)b * nc;
            for (i = 0; i < An; i++) x[i] = gb[i];
            QDLDL_solve(An, W->Lp, W->Li, W->Lx, W->Dinv, x);
        }
        W->t_schur += t1 - t0;
        W->t_qdldl += now_s() - t1;
    }
}

static void *worker_main(void *arg)
{
    worker_t *W = (worker_t *)arg;
    pool_t *P = W->pool;
    long seen = 0;
    for (;;) {
        pthread_mutex_lock(&P->mu);
        while (!P->quit && P->generation == seen) pthread_cond_wait(&P->go, &P->mu);
        if (P->quit) { pthread_mutex_unlock(&P->mu); return NULL; }
        seen = P->generation;
        pthread_mutex_unlock(&P->mu);
        run_slab(W);
        pthread_mutex_lock(&P->mu);
        if (--P->pending == 0) pthread_cond_signal(&P->done);
        pthread_mutex_unlock(&P->mu);
    }
}

static int worker_init(worker_t *W)
{
    pool_t *P = W->pool;
    const int s = P->s, c = P->c, N = P->N, An = s * N, ss = s * s, cc = c * c;
    QDLDL_int sumLnz;
    W->Ap = malloc(sizeof(QDLDL_int) * (An + 1)); W->Ai = malloc(sizeof(QDLDL_int) * P->nnz);
    W->Ax = malloc(sizeof(QDLDL_float) * P->nnz);
    W->etree = malloc(sizeof(QDLDL_int) * An); W->Lnz = malloc(sizeof(QDLDL_int) * An);
    W->Lp = malloc(sizeof(QDLDL_int) * (An + 1));
    W->D = malloc(sizeof(QDLDL_float) * An); W->Dinv = malloc(sizeof(QDLDL_float) * An);
    W->iwork = malloc(sizeof(QDLDL_int) * 3 * An);
    W->bwork = malloc(sizeof(QDLDL_bool) * An); W->fwork = malloc(sizeof(QDLDL_float) * An);
    W->Gw = malloc(sizeof(float) * ((size_t)(ss + cc) * N - cc));
    W->S = malloc(sizeof(float) * (size_t)3 * ss * N); W->gamma = malloc(sizeof(float) * An);
    oracle_csc_pattern(s, N, W->Ap, W->Ai);
    sumLnz = QDLDL_etree(An, W->Ap, W->Ai, W->iwork, W->Lnz, W->etree);     /* include/qdldl/sqp.cuh:193 */
    if (sumLnz < 0) return -1;
    W->Li = malloc(sizeof(QDLDL_int) * sumLnz);
    W->Lx = malloc(sizeof(QDLDL_float) * sumLnz);
    return 0;
}

int cpu_pool_nnz(int s, int N) { return (N - 1) * s * s + N * ((s + 1) * s / 2); }   /* include/qdldl/sqp.cuh:148 */

void *cpu_pool_create(int s, int c, int N, int nthreads)
{
    pool_t *P = calloc(1, sizeof(pool_t));
    int t;
    if (nthreads < 1) nthreads = 1;
    P->s = s; P->c = c; P->N = N; P->nthreads = nthreads; P->nnz = cpu_pool_nnz(s, N);
    pthread_mutex_init(&P->mu, NULL); pthread_cond_init(&P->go, NULL); pthread_cond_init(&P->done, NULL);
    P->w = calloc(nthreads, sizeof(worker_t));
    for (t = 0; t < nthreads; t++) {
        P->w[t].pool = P; P->w[t].id = t;
        if (worker_init(&P->w[t]) != 0) return NULL;
    }
    for (t = 1; t < nthreads; t++) pthread_create(&P->w[t].th, NULL, worker_main, &P->w[t]);   /* slab 0 runs on the caller */
    return P;
}

static double pool_run(pool_t *P, int batch, double *sum_schur_s, double *sum_qdldl_s)
{
    int t, first = 0, status = 0, active = P->nthreads < batch ? P->nthreads : batch;
    double t0 = now_s(), wall;
    pthread_mutex_lock(&P->mu);
    for (t = 0; t < P->nthreads; t++) {
        int cnt = t < active ? batch / active + (t < batch % active) : 0;
        P->w[t].first = first; P->w[t].count = cnt; P->w[t].status = 0;
        first += cnt;
    }
    P->pending = P->nthreads - 1;
    P->generation++;
    pthread_cond_broadcast(&P->go);
    pthread_mutex_unlock(&P->mu);
    run_slab(&P->w[0]);
    pthread_mutex_lock(&P->mu);
    while (P->pending > 0) pthread_cond_wait(&P->done, &P->mu);
    pthread_mutex_unlock(&P->mu);
    wall = now_s() - t0;
    if (sum_schur_s) *sum_schur_s = 0;
    if (sum_qdldl_s) *sum_qdldl_s = 0;
    for (t = 0; t < P->nthreads; t++) {
        if (sum_schur_s) *sum_schur_s += P->w[t].t_schur;
        if (sum_qdldl_s) *sum_qdldl_s += P->w[t].t_qdldl;
        if (P->w[t].status) status = P->w[t].status;
    }
    return status ? (double)status : wall;
}

/* QDLDL factor + solve of `batch` systems given as CSC values val[batch][nnz] and gamma[batch][s*N]
 * (what the reference's QDLDL slot holds after its D2H copies).  Returns wall seconds (< 0: QDLDL failure). */
double cpu_pool_qdldl(void *pool, int batch, const float *val, const float *gamma, float *lambda)
{
    pool_t *P = (pool_t *)pool;
    P->mode = 0; P->val = val; P->gamma_in = gamma; P->lambda = lambda;
    return pool_run(P, batch, NULL, NULL);
}

/* Schur restatement (no preconditioner) + CSC packing + QDLDL factor + solve from (G, C, g, cv, rho). */
double cpu_pool_kkt(void *pool, int batch, const float *G, const float *C, const float *g, const float *cv,
                    float rho, float *lambda, double *sum_schur_s, double *sum_qdldl_s)
{
    pool_t *P = (pool_t *)pool;
    P->mode = 1; P->G = G; P->C = C; P->g = g; P->cv = cv; P->rho = rho; P->lambda = lambda;
    return pool_run(P, batch, sum_schur_s, sum_qdldl_s);
}

/* Preparation (not part of any timed figure): CSC values and gamma of every instance, as the reference's
 * form_schur_qdl_kernel would leave them in h_val / h_gamma. */
double cpu_pool_schur_csc(void *pool, int batch, const float *G, const float *C, const float *g, const float *cv,
                          float rho, float *val_out, float *gamma_out)
{
    pool_t *P = (pool_t *)pool;
    P->mode = 2; P->G = G; P->C = C; P->g = g; P->cv = cv; P->rho = rho; P->val_out = val_out; P->gamma_out = gamma_out;
    return pool_run(P, batch, NULL, NULL);
}

void cpu_pool_destroy(void *pool)
{
    pool_t *P = (pool_t *)pool;
    int t;
    if (!P) return;
    pthread_mutex_lock(&P->mu);
    P->quit = 1;
    pthread_cond_broadcast(&P->go);
    pthread_mutex_unlock(&P->mu);
    for (t = 1; t < P->nthreads; t++) pthread_join(P->w[t].th, NULL);
    for (t = 0; t < P->nthreads; t++) {
        worker_t *W = &P->w[t];
        free(W->Ap); free(W->Ai); free(W->Ax); free(W->etree); free(W->Lnz); free(W->Lp); free(W->D); free(W->Dinv);
        free(W->iwork); free(W->bwork); free(W->fwork); free(W->Gw); free(W->S); free(W->gamma); free(W->Li); free(W->Lx);
    }
    free(P->w); free(P);
}

/* One-shot form kept for the tests: a pool for the duration of the call. */
double cpu_kkt_solve_batch(int s, int c, int N, int batch, int nthreads, int with_schur,
                           const float *G, const float *C, const float *g, const float *cv,
                           float rho, const float *S_in, const float *gamma_in,
                           float *lambda, double *sum_schur_s, double *sum_qdldl_s)
{
    void *P;
    double wall;
    if (nthreads > batch) nthreads = batch;
    P = cpu_pool_create(s, c, N, nthreads);
    if (!P) return -1.0;
    if (with_schur) {
        wall = cpu_pool_kkt(P, batch, G, C, g, cv, rho, lambda, sum_schur_s, sum_qdldl_s);
    } else {
        const int nnz = cpu_pool_nnz(s, N);
        float *val = malloc(sizeof(float) * (size_t)nnz * batch);
        int b;
        for (b = 0; b < batch; b++) oracle_csc_values_f32(s, N, S_in + (size_t)b * 3 * s * s * N, val + (size_t)b * nnz);
        wall = cpu_pool_qdldl(P, batch, val, gamma_in, lambda);
        if (sum_schur_s) *sum_schur_s = 0;
        if (sum_qdldl_s) *sum_qdldl_s = wall;
        free(val);
    }
    cpu_pool_destroy(P);
    return wall;
}

static int cmp_double(const void *a, const void *b) { const double x = *(const double *)a, y = *(const double *)b; return (x > y) - (x < y); }

/* Single-core latency of the reference's host solve for ONE system (SURVEY 8(d) CPU baseline (1)): median over
 * `reps` timed repetitions (after `warmup` untimed ones) of qdldl_solve_schur = QDLDL_factor + copy +
 * QDLDL_solve on precomputed CSC values, elimination tree computed once (include/qdldl/sqp.cuh:22-49,193),
 * clock_gettime(CLOCK_MONOTONIC) like include/utils/experiment.cuh:14.  Returns microseconds (< 0: failure). */
double cpu_qdldl_single_us(int s, int N, const float *val, const float *gamma, int reps, int warmup, float *lambda)
{
    pool_t P;
    worker_t W;
    double *t, med;
    int r, i;
    const int An = s * N;
    memset(&P, 0, sizeof P); memset(&W, 0, sizeof W);
    P.s = s; P.c = 0; P.N = N; P.nthreads = 1; P.nnz = cpu_pool_nnz(s, N);
    W.pool = &P;
    if (reps < 1 || worker_init(&W) != 0) return -1.0;
    t = malloc(sizeof(double) * reps);
    for (r = -warmup; r < reps; r++) {
        const double t0 = now_s();
        if (QDLDL_factor(An, W.Ap, W.Ai, val, W.Lp, W.Li, W.Lx, W.D, W.Dinv, W.Lnz, W.etree, W.bwork, W.iwork, W.fwork) < 0) { free(t); return -2.0; }
        for (i = 0; i < An; i++) lambda[i] = gamma[i];
        QDLDL_solve(An, W.Lp, W.Li, W.Lx, W.Dinv, lambda);
        if (r >= 0) t[r] = (now_s() - t0) * 1e6;
    }
    qsort(t, reps, sizeof(double), cmp_double);
    med = t[reps / 2];
    free(t);
    free(W.Ap); free(W.Ai); free(W.Ax); free(W.etree); free(W.Lnz); free(W.Lp); free(W.D); free(W.Dinv);
    free(W.iwork); free(W.bwork); free(W.fwork); free(W.Gw); free(W.S); free(W.gamma); free(W.Li); free(W.Lx);
    return med;
}
