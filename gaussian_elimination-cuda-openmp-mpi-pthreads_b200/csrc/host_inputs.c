/* Host-side input generators shared by libgauss_cuda.so and the C harness (include/gauss_cuda.h).
 * Plain C, no CUDA. */
#include "gauss_cuda.h"
#include <math.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>

/* The reference's built-in fixture, gauss_internal_input.c:50-60. */
void gauss_host_init_matrix(double *A, double *B, int n)
{
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j)
            A[(size_t)i * n + j] = (j < i) ? 2.0 * (j + 1) : 2.0 * (i + 1);
        B[i] = (double)i;
    }
}

/* Seeded mode: uniform rand()/32768.0 entries made strictly diagonally dominant (a no-pivot FP32
 * solver cannot meet 1e-3 on un-dominated random matrices), preset solution i+1 and
 * B = A * X_true in double as the reference's initRHS does (gauss_external_input.c:88-106). */
void gauss_host_initialize_inputs(double *A, double *B, double *X_true, int n, unsigned seed)
{
    srand(seed);
    for (int i = 0; i < n; ++i) {
        double rowsum = 0.0;
        for (int j = 0; j < n; ++j) {
            const double v = (double)rand() / 32768.0;
            A[(size_t)i * n + j] = v;
            rowsum += fabs(v);
        }
        A[(size_t)i * n + i] += rowsum;
    }
    for (int i = 0; i < n; ++i) X_true[i] = (double)(i + 1);
    for (int i = 0; i < n; ++i) {
        double r = 0.0;
        for (int j = 0; j < n; ++j) r += A[(size_t)i * n + j] * X_true[j];
        B[i] = r;
    }
}

/* cuda_matmul.cu:126-132 */
void gauss_host_matmul_fill(float *A, float *B, int n)
{
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            const int idx = (int)(float)(j + i * n);
            A[idx] = idx + 1.0f;
            B[idx] = 1.0f / (idx + 1.0f);
        }
}

/* gauss_external_input.c:32-84 */
int gauss_host_read_dat(const char *fname, double **A_out)
{
    if (!fname || !A_out) return -1;
    FILE *fp = fopen(fname, "r");
    if (!fp) return -1;
    char buf[1024];
    int n = 0, ncols = 0, nnz = 0;
    if (!fgets(buf, sizeof buf, fp) || sscanf(buf, "%d %d %d", &n, &ncols, &nnz) < 1 || n <= 0) { fclose(fp); return -1; }
    double *A = calloc((size_t)n * (size_t)n, sizeof *A);
    if (!A) { fclose(fp); return -1; }
    while (fgets(buf, sizeof buf, fp)) {
        int i = 0, j = 0;
        double v = 0.0;
        if (sscanf(buf, "%d %d %lf", &i, &j, &v) < 1 || i == 0) break;
        if (i >= 1 && i <= n && j >= 1 && j <= n) A[(size_t)(i - 1) * n + (j - 1)] = v;
    }
    fclose(fp);
    *A_out = A;
    return n;
}

/* gauss_external_input.c:88-106 */
void gauss_host_init_rhs(const double *A, double *R, double *X_pre, int n)
{
    for (int i = 0; i < n; ++i) X_pre[i] = (double)(i + 1);
    for (int i = 0; i < n; ++i) {
        double r = 0.0;
        for (int j = 0; j < n; ++j) r += A[(size_t)i * n + j] * X_pre[j];
        R[i] = r;
    }
}

void gauss_host_free(void *p) { free(p); }
