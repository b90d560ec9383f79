/*
 * cpu_baseline.c -- TEST / BENCH INFRASTRUCTURE ONLY (never on the product path).
 *
 * The reference's CPU implementation of the KKT linear solve, as
 * include/qdldl/sqp.cuh does it per SQP iteration:
 *     Schur formation (oracle restatement of include/qdldl/linsys_setup.cuh:12-336,
 *     == oracle_schur_f32 without Pinv use) -> CSC values (include/utils/csr.cuh:10-36)
 *     -> QDLDL_factor -> x=b -> QDLDL_solve      (include/qdldl/sqp.cuh:22-49,271)
 * with QDLDL_etree computed once (include/qdldl/sqp.cuh:193).
 *
 * Built by oracle/build.py into oracle/_ref/libcpu_baseline.so together with the
 * REFERENCE's own qdldl/src/qdldl.c (compiled from /root/reference where it lies,
 * float/int configuration of the reference Makefile:16).
 *
 * Threads: a plain pthread pool, instances split into contiguous slabs, one QDLDL
 * workspace per thread (SURVEY 8(d) "CPU baseline timing" (2)).
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

typedef struct {
    int s, c, N, first, count, with_schur;
    const float *G, *C, *g, *cv;          /* batch-major inputs */
    const float *S_in, *gamma_in;         /* used when with_schur == 0 */
    float rho;
    float *lambda;                        /* batch * s*N */
    double t_schur, t_qdldl;
    int status;
} job_t;

static void *worker(void *arg)
{
    job_t *J = (job_t *)arg;
    const int s = J->s, c = J->c, N = J->N, An = s * N;
    const int ss = s * s, cc = c * c;
    const size_t nG = (size_t)(ss + cc) * N - cc, nC = (size_t)(ss + s * c) * (N - 1);
    const size_t ng = (size_t)(s + c) * N - c, nc = (size_t)s * N, nS = (size_t)3 * ss * N;
    const int nnz = (N - 1) * ss + N * ((s + 1) * s / 2);    /* include/qdldl/sqp.cuh:148 */
    QDLDL_int *Ap = malloc(sizeof(QDLDL_int) * (An + 1)), *Ai = malloc(sizeof(QDLDL_int) * nnz);
    QDLDL_float *Ax = malloc(sizeof(QDLDL_float) * nnz);
    QDLDL_int *etree = malloc(sizeof(QDLDL_int) * An), *Lnz = malloc(sizeof(QDLDL_int) * An);
    QDLDL_int *Lp = malloc(sizeof(QDLDL_int) * (An + 1));
    QDLDL_float *D = malloc(sizeof(QDLDL_float) * An), *Dinv = malloc(sizeof(QDLDL_float) * An);
    QDLDL_int *iwork = malloc(sizeof(QDLDL_int) * 3 * An);
    QDLDL_bool *bwork = malloc(sizeof(QDLDL_bool) * An);
    QDLDL_float *fwork = malloc(sizeof(QDLDL_float) * An);
    float *Gw = malloc(sizeof(float) * nG), *S = malloc(sizeof(float) * nS);
    float *Pinv = malloc(sizeof(float) * nS), *gamma = malloc(sizeof(float) * nc);
    QDLDL_int sumLnz, *Li;
    QDLDL_float *Lx;
    int b, i;

    oracle_csc_pattern(s, N, Ap, Ai);
    sumLnz = QDLDL_etree(An, Ap, Ai, iwork, Lnz, etree);     /* include/qdldl/sqp.cuh:193 */
    if (sumLnz < 0) { J->status = -1; return NULL; }
    Li = malloc(sizeof(QDLDL_int) * sumLnz);
    Lx = malloc(sizeof(QDLDL_float) * sumLnz);

    for (b = J->first; b < J->first + J->count; b++) {
        double t0 = now_s(), t1;
        const float *Sb, *gb;
        if (J->with_schur) {
            memcpy(Gw, J->G + (size_t)b * nG, sizeof(float) * nG);
            oracle_schur_f32(s, c, N, Gw, J->C + (size_t)b * nC, J->g + (size_t)b * ng,
                             J->cv + (size_t)b * nc, J->rho, S, Pinv, gamma);
            Sb = S; gb = gamma;
        } else {
            Sb = J->S_in + (size_t)b * nS; gb = J->gamma_in + (size_t)b * nc;
        }
        t1 = now_s();
        oracle_csc_values_f32(s, N, Sb, Ax);
        /* qdldl_solve_schur, include/qdldl/sqp.cuh:22-49 */
        if (QDLDL_factor(An, Ap, Ai, Ax, Lp, Li, Lx, D, Dinv, Lnz, etree, bwork, iwork, fwork) < 0)
            J->status = -2;
        {
            float *x = J->lambda + (size_t)b * nc;
            for (i = 0; i < An; i++) x[i] = gb[i];
            QDLDL_solve(An, Lp, Li, Lx, Dinv, x);
        }
        J->t_schur += t1 - t0;
        J->t_qdldl += now_s() - t1;
    }
    free(Ap); free(Ai); free(Ax); free(etree); free(Lnz); free(Lp); free(D); free(Dinv);
    free(iwork); free(bwork); free(fwork); free(Gw); free(S); free(Pinv); free(gamma);
    free(Li); free(Lx);
    return NULL;
}

/* Solve `batch` KKT systems on `nthreads` host threads.  Returns wall seconds, or <0 on
 * a QDLDL failure.  with_schur=1: inputs are (G,C,g,cv,rho) and the Schur formation is
 * part of the timed work; with_schur=0: inputs are precomputed (S_in, gamma_in).
 * sum_schur_s / sum_qdldl_s: per-thread CPU seconds summed over threads. */
double cpu_kkt_solve_batch(int s, int c, int N, int batch, int nthreads, int with_schur,
                           const float *G, const float *C, const float *g, const float *cv,
                           float rho, const float *S_in, const float *gamma_in,
                           float *lambda, double *sum_schur_s, double *sum_qdldl_s)
{
    pthread_t *th;
    job_t *jobs;
    int t, first = 0, status = 0;
    double t0, wall;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > batch) nthreads = batch;
    th = malloc(sizeof(pthread_t) * nthreads);
    jobs = calloc(nthreads, sizeof(job_t));
    t0 = now_s();
    for (t = 0; t < nthreads; t++) {
        int cnt = batch / nthreads + (t < batch % nthreads);
        job_t *J = &jobs[t];
        J->s = s; J->c = c; J->N = N; J->first = first; J->count = cnt;
        J->with_schur = with_schur; J->G = G; J->C = C; J->g = g; J->cv = cv;
        J->S_in = S_in; J->gamma_in = gamma_in; J->rho = rho; J->lambda = lambda;
        first += cnt;
        if (nthreads == 1) worker(J);
        else pthread_create(&th[t], NULL, worker, J);
    }
    if (nthreads > 1) for (t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    wall = now_s() - t0;
    *sum_schur_s = 0; *sum_qdldl_s = 0;
    for (t = 0; t < nthreads; t++) {
        *sum_schur_s += jobs[t].t_schur; *sum_qdldl_s += jobs[t].t_qdldl;
        if (jobs[t].status) status = jobs[t].status;
    }
    free(th); free(jobs);
    return status ? (double)status : wall;
}
