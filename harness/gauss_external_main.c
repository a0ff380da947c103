/*
 * gauss_external_main -- host harness for the reference's EXTERNAL-input program
 * (OpenMP_and_MPI/gauss_openmp/gauss_external_input.c) over include/gauss_cuda.h.
 *
 * Kept from the reference: CLI  <matrixfile> [threads]  (:214-230; threads accepted and ignored),
 * initMatrix(file) / initRHS / initResult before the timer (:233-237), only computeGauss inside it
 * (:238-240), solveGauss after it (:242), and the prints
 *   "\nMatrix File: %s; Matrix Size: %d ; Threads: %d\n" (:234), "Time:  %f seconds\n" (:244), "Error: %e\n" (:253);
 * a missing file prints "The matrix file open error" and exits -1 (:42-45), a singular matrix prints
 * "The matrix is singular" on stderr and exits -1 (:133-136).  Elimination uses partial pivoting.
 * Extra: a trailing  -r  solves in the reference's own precision (FP32 pivoted elimination + iterative refinement
 * with double residuals, gauss_cuda_solve_refined); the timer then covers the whole refined solve.
 */
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <sys/time.h>
#include "gauss_cuda.h"

double **matrix, *X, *R, *X__;
static gauss_cuda_ctx *ctx;
static float *A32, *R32, *X32;

static void fail(int st, const char *where)
{
    if (st == GAUSS_CUDA_OK) return;
    if (st == GAUSS_CUDA_ZERO_PIVOT) { fprintf(stderr, "The matrix is singular\n"); exit(-1); }
    fprintf(stderr, "%s: %s (%s)\n", where, gauss_cuda_strerror(st), gauss_cuda_last_error());
    exit(2);
}

int initMatrix(const char *fname)
{
    double *flat = NULL;
    int n = gauss_host_read_dat(fname, &flat);
    if (n < 0) { fprintf(stderr, "The matrix file open error\n"); exit(-1); }
    matrix = malloc((size_t)n * sizeof *matrix);
    for (int i = 0; i < n; ++i) matrix[i] = flat + (size_t)i * n;
    return n;
}

void initRHS(int nsize)
{
    X__ = malloc((size_t)nsize * sizeof *X__);
    R = malloc((size_t)nsize * sizeof *R);
    gauss_host_init_rhs(matrix[0], R, X__, nsize);
}

void initResult(int nsize)
{
    X = calloc((size_t)nsize, sizeof *X);
}

void computeGauss(int nsize)
{
    fail(gauss_cuda_upload(ctx, A32, R32), "gauss_cuda_upload");
    fail(gauss_cuda_compute_gauss(ctx, NULL), "gauss_cuda_compute_gauss");
}

void solveGauss(int nsize)
{
    fail(gauss_cuda_solve_gauss(ctx, NULL), "gauss_cuda_solve_gauss");
    fail(gauss_cuda_download(ctx, X32, NULL), "gauss_cuda_download");
    for (int i = 0; i < nsize; ++i) X[i] = X32[i];
}

int main(int argc, char *argv[])
{
    int refine = 0;
    if (argc >= 3 && argv[argc - 1][0] == '-' && argv[argc - 1][1] == 'r' && argv[argc - 1][2] == 0) { refine = 1; --argc; }
    if (argc != 2 && argc != 3) {
        fprintf(stderr, "usage: %s <matrixfile> <number_of_threads (optional)>\n", argv[0]);
        exit(-1);
    }
    if (argc == 3 && atoi(argv[2]) <= 0) {
        fprintf(stderr, "Error: Threads count should be a natural number. \n");
        fprintf(stderr, "usage: %s <matrixfile> <number_of_threads (optional)>\n", argv[0]);
        exit(-1);
    }
    int nsize = initMatrix(argv[1]);
    printf("\nMatrix File: %s; Matrix Size: %d ; Threads: %d\n", argv[1], nsize, 1);
    initRHS(nsize);
    initResult(nsize);

    /* device context and FP32 staging are set up outside the timer, like the reference's allocation */
    size_t n = (size_t)nsize;
    A32 = malloc(n * n * sizeof *A32); R32 = malloc(n * sizeof *R32); X32 = malloc(n * sizeof *X32);
    for (size_t i = 0; i < n * n; ++i) A32[i] = (float)matrix[0][i];
    for (size_t i = 0; i < n; ++i) R32[i] = (float)R[i];
    fail(gauss_cuda_init(&ctx, nsize, 1, 0, GAUSS_CUDA_MODE_3XTF32), "gauss_cuda_init");
    fail(gauss_cuda_set_pivoting(ctx, 1), "gauss_cuda_set_pivoting");

    struct timeval start, finish;
    gauss_cuda_refine_info info;
    gettimeofday(&start, 0);
    if (refine) fail(gauss_cuda_solve_refined(ctx, matrix[0], R, X, 30, 1e-14, &info), "gauss_cuda_solve_refined");
    else computeGauss(nsize);
    gettimeofday(&finish, 0);
    if (!refine) solveGauss(nsize);
    fprintf(stdout, "Time:  %f seconds\n", (finish.tv_sec - start.tv_sec) + (finish.tv_usec - start.tv_usec) * 0.000001);

    double error = 0.0;
    for (int i = 0; i < nsize; ++i) {
        double e = (X__[i] == 0.0) ? 1.0 : fabs((X[i] - X__[i]) / X__[i]);
        if (error < e) error = e;
    }
    fprintf(stdout, "Error: %e\n", error);
    if (refine) fprintf(stdout, "Refinement: %d sweeps, %s, last correction %.2e\n", info.iterations,
                        info.converged == 1 ? "converged" : (info.converged == 2 ? "stagnated at working precision" : "NOT converged"), info.last_correction);
    gauss_cuda_free(ctx);
    return 0;
}
