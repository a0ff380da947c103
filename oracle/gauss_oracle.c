/*
 * oracle/gauss_oracle.c -- TEST INFRASTRUCTURE ONLY (never shipped, never the
 * thing measured, except as bench.py's `cpu_baseline` leg).
 *
 * CPU restatement, in double precision, of the reference's Gaussian-elimination
 * path and of its FP32 matrix multiply.  Citations are to files under
 * /root/reference:
 *   OI  = OpenMP_and_MPI/gauss_openmp/gauss_internal_input.c
 *   OE  = OpenMP_and_MPI/gauss_openmp/gauss_external_input.c
 *   CU2 = CUDA_and_OpenMP/Version-2/cuda_matmul.cu
 *
 * Parity is PINNED: tests/test_oracle.py checks this file against
 *   (1) the analytic known answer of the built-in fixture (OI:45-48),
 *   (2) the reference's own object code (oracle/_ref/libgauss_ref.so, built by
 *       oracle/Makefile from the unmodified OI source) on the fixture and on
 *       seeded inputs, bit for bit,
 *   (3) committed golden vectors in tests/golden/ produced by that reference
 *       build (tests/golden/make_golden.py).
 *
 * Unlike the reference (file-scope globals, OI:15-16) every routine here takes
 * flat caller-owned buffers: A is row-major n*n with leading dimension n.
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_OK        0
#define ORACLE_SINGULAR  1

/* ---- inputs ----------------------------------------------------------- */

/* Built-in fixture, OI:50-60: A[i][j] = 2*(min(i,j)+1), B[i] = i, swap[i] = i. */
void oracle_init_matrix(double *A, double *B, int *swap, int n)
{
    for (int r = 0; r < n; ++r) {
        double *row = A + (size_t)r * n;
        for (int c = 0; c < n; ++c)
            row[c] = 2.0 * ((c < r ? c : r) + 1);
        B[r] = (double)r;
        if (swap) swap[r] = r;
    }
}

/* Seeded mode (north_star's initialize_inputs).  srand(seed);
 * A[i][j] = rand()/32768.0 in row-major order; the diagonal then gets the row's
 * absolute sum added (strict dominance, needed for a no-pivot FP32 solver);
 * the preset solution is X__[i] = i+1 and B = A * X__ accumulated in double in
 * ascending j, which is the reference's initRHS (OE:88-106). */
void oracle_initialize_inputs(double *A, double *B, double *Xtrue, int n, unsigned seed)
{
    srand(seed);
    for (int r = 0; r < n; ++r) {
        double *row = A + (size_t)r * n;
        double s = 0.0;
        for (int c = 0; c < n; ++c) {
            row[c] = (double)rand() / 32768.0;
            s += fabs(row[c]);
        }
        row[r] += s;
    }
    for (int r = 0; r < n; ++r) Xtrue[r] = (double)(r + 1);
    for (int r = 0; r < n; ++r) {
        const double *row = A + (size_t)r * n;
        double acc = 0.0;
        for (int c = 0; c < n; ++c) acc += row[c] * Xtrue[c];
        B[r] = acc;
    }
}

/* ---- elimination ------------------------------------------------------- */

/* getPivot, OI:66-112.  Only an exactly-zero pivot triggers a search for the
 * first lower row with a non-zero entry in this column (OI:75-84); the tails
 * [cur, n) of the two rows, B and swap are exchanged (OI:91-97); then the row
 * is divided by its pivot (true division, OI:100-111) unless the pivot is 1.
 * The reference prints "The matrix is singular" and exits (OI:86-89); here the
 * caller gets ORACLE_SINGULAR. */
int oracle_get_pivot(double *A, double *B, int *swap, int n, int cur)
{
    double *prow = A + (size_t)cur * n;
    double lead = prow[cur];
    int pick = cur;
    if (lead == 0.0) {
        for (int r = cur; r < n; ++r) {
            double v = A[(size_t)r * n + cur];
            if (v != 0.0) { lead = v; pick = r; break; }
        }
    }
    if (lead == 0.0) return ORACLE_SINGULAR;
    if (pick != cur) {
        double *other = A + (size_t)pick * n;
        for (int c = cur; c < n; ++c) { double t = other[c]; other[c] = prow[c]; prow[c] = t; }
        { double t = B[pick]; B[pick] = B[cur]; B[cur] = t; }
        if (swap) { int t = swap[pick]; swap[pick] = swap[cur]; swap[cur] = t; }
    }
    double piv = prow[cur];
    if (piv != 1.0) {
        prow[cur] = 1.0;
        for (int c = cur + 1; c < n; ++c) prow[c] /= piv;
        B[cur] /= piv;
    }
    return ORACLE_OK;
}

/* Pivot steps [first, last) of computeGauss, OI:118-138: after getPivot, every
 * row j > i does  m = A[j][i]; A[j][i] = 0; A[j][i+1:] -= m*A[i][i+1:];
 * B[j] -= m*B[i]  (rows in parallel, OI:128).  The reference is built without
 * -ffp-contract control; this file is compiled with -ffp-contract=off so the
 * multiply and subtract round separately like the reference's x86-64 build. */
int oracle_compute_gauss_steps(double *A, double *B, int *swap, int n,
                               int first, int last, int nthreads)
{
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#else
    (void)nthreads;
#endif
    for (int i = first; i < last; ++i) {
        int rc = oracle_get_pivot(A, B, swap, n, i);
        if (rc != ORACLE_OK) return rc;
        const double *pi = A + (size_t)i * n;
        const double bi = B[i];
#pragma omp parallel for schedule(static)
        for (int j = i + 1; j < n; ++j) {
            double *pj = A + (size_t)j * n;
            const double m = pj[i];
            pj[i] = 0.0;
            for (int k = i + 1; k < n; ++k) pj[k] -= m * pi[k];
            B[j] -= m * bi;
        }
    }
    return ORACLE_OK;
}

int oracle_compute_gauss(double *A, double *B, int *swap, int n, int nthreads)
{
    return oracle_compute_gauss_steps(A, B, swap, n, 0, n, nthreads);
}

/* solveGauss, OI:144-159: unit-diagonal back substitution; the inner sum runs
 * over DESCENDING j (OI:151-153); C is a plain copy of V (OI:156-158). */
void oracle_solve_gauss(const double *A, const double *B, double *V, double *C, int n)
{
    V[n - 1] = B[n - 1];
    for (int i = n - 2; i >= 0; --i) {
        const double *row = A + (size_t)i * n;
        double v = B[i];
        for (int j = n - 1; j > i; --j) v -= row[j] * V[j];
        V[i] = v;
    }
    for (int i = 0; i < n; ++i) C[i] = V[i];
}

/* gauss() of the north_star = computeGauss + solveGauss on the same buffers. */
int oracle_gauss(double *A, double *B, double *X, int n, int nthreads)
{
    int rc = oracle_compute_gauss(A, B, NULL, n, nthreads);
    if (rc != ORACLE_OK) return rc;
    double *V = (double *)malloc((size_t)n * sizeof(double));
    if (!V) return -1;
    oracle_solve_gauss(A, B, V, X, n);
    free(V);
    return ORACLE_OK;
}

/* ---- acceptance metrics ------------------------------------------------ */

/* OE:246-253: max_i |X[i]-Xref[i]| / |Xref[i]|, a term counting 1.0 where the
 * reference value is exactly zero. */
double oracle_max_rel_error(const double *X, const double *Xref, int n)
{
    double worst = 0.0;
    for (int i = 0; i < n; ++i) {
        double e = (Xref[i] == 0.0) ? 1.0 : fabs((X[i] - Xref[i]) / Xref[i]);
        if (e > worst) worst = e;
    }
    return worst;
}

/* Norm-wise variant for the fixture, whose solution has n-2 exact zeros
 * (SURVEY 8c): max|dX| / max|Xref|. */
double oracle_normwise_error(const double *X, const double *Xref, int n)
{
    double num = 0.0, den = 0.0;
    for (int i = 0; i < n; ++i) {
        double d = fabs(X[i] - Xref[i]);
        if (d > num) num = d;
        if (fabs(Xref[i]) > den) den = fabs(Xref[i]);
    }
    return den == 0.0 ? num : num / den;
}

/* north_star's scaled residual ||Ax-b||inf / (||A||inf ||x||inf n eps), eps = 2^-24,
 * evaluated in double from the FLOAT inputs the GPU path was given. */
double oracle_scaled_residual_f32(const float *A, const float *B, const float *X, int n)
{
    double rmax = 0.0, anorm = 0.0, xnorm = 0.0;
    for (int i = 0; i < n; ++i) if (fabs((double)X[i]) > xnorm) xnorm = fabs((double)X[i]);
#pragma omp parallel for schedule(static) reduction(max:rmax,anorm)
    for (int i = 0; i < n; ++i) {
        const float *row = A + (size_t)i * n;
        double acc = 0.0, rs = 0.0;
        for (int j = 0; j < n; ++j) { acc += (double)row[j] * (double)X[j]; rs += fabs((double)row[j]); }
        double r = fabs(acc - (double)B[i]);
        if (r > rmax) rmax = r;
        if (rs > anorm) anorm = rs;
    }
    const double eps = 1.0 / 16777216.0;
    double den = anorm * xnorm * (double)n * eps;
    return den == 0.0 ? rmax : rmax / den;
}

/* ---- FP32 matmul (CU2) ------------------------------------------------- */

/* CU2:126-132: idx = j + i*n; A[idx] = idx + 1, B[idx] = 1/(idx + 1), FP32. */
void oracle_matmul_fill(float *A, float *B, int n)
{
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            int idx = (int)(float)(j + i * n);          /* CU2:128 round-trips through float */
            A[idx] = idx + 1.0f;
            B[idx] = 1.0f / (idx + 1.0f);
        }
}

/* seq_matmul CU2:28-39 / omp_matmul CU2:42-57 / gpu_matmul CU2:89-101 all form
 * C[i][j] = sum_k A[i][k]*B[k][j] with ONE FP32 accumulator and ascending k.
 * `fused` selects what the multiply-add rounds like: 0 = separate multiply and
 * add (the CPU loops on baseline x86-64), 1 = fmaf (nvcc contracts the device
 * loop to FFMA).  nthreads <= 1 is the sequential loop. */
void oracle_matmul(const float *A, const float *B, float *C, int n, int fused, int nthreads)
{
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(static) if (nthreads != 1)
    for (int i = 0; i < n; ++i) {
        const float *arow = A + (size_t)i * n;
        for (int j = 0; j < n; ++j) {
            float acc = 0.0f;
            if (fused) {
                for (int k = 0; k < n; ++k) acc = fmaf(arow[k], B[(size_t)k * n + j], acc);
            } else {
                for (int k = 0; k < n; ++k) { float p = arow[k] * B[(size_t)k * n + j]; acc = acc + p; }
            }
            C[(size_t)i * n + j] = acc;
        }
    }
}

/* Double-precision matmul of the float inputs: the yardstick both the GPU
 * result and the reference loops are measured against in the 1e-4 check. */
void oracle_matmul_f64(const float *A, const float *B, double *C, int n)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; ++i) {
        double *crow = C + (size_t)i * n;
        for (int j = 0; j < n; ++j) crow[j] = 0.0;
        for (int k = 0; k < n; ++k) {
            const double a = (double)A[(size_t)i * n + k];
            const float *brow = B + (size_t)k * n;
            for (int j = 0; j < n; ++j) crow[j] += a * (double)brow[j];
        }
    }
}

double oracle_max_rel_diff_f32(const float *X, const float *Y, size_t count)
{
    double worst = 0.0;
    for (size_t i = 0; i < count; ++i) {
        double d = fabs((double)X[i] - (double)Y[i]);
        double s = fabs((double)Y[i]);
        double e = s > 0.0 ? d / s : d;
        if (e > worst) worst = e;
    }
    return worst;
}

/* ---- external-input variant: coordinate files, partial pivoting (OE) ----------------------- */
#include <stdio.h>

/* initMatrix(fname), OE:32-84: first line "rows cols nnz" (only rows is used), then "row col value"
 * lines, 1-based, until a line whose row is 0.  Returns the order n and a malloc'd dense row-major
 * matrix in *A_out (caller frees), or -1 if the file cannot be opened (the reference exits). */
int oracle_read_dat(const char *fname, double **A_out)
{
    FILE *f = fopen(fname, "r");
    if (!f) return -1;
    char line[1024];
    int a = 0, b = 0, c = 0;
    if (!fgets(line, sizeof line, f) || sscanf(line, "%d %d %d", &a, &b, &c) < 1 || a <= 0) { fclose(f); return -1; }
    const int n = a;
    double *A = (double *)calloc((size_t)n * n, sizeof(double));
    if (!A) { fclose(f); return -1; }
    while (fgets(line, sizeof line, f)) {
        int r = 0, col = 0;
        double v = 0.0;
        sscanf(line, "%d %d %lf", &r, &col, &v);
        if (r == 0) break;
        A[(size_t)(r - 1) * n + (col - 1)] = v;
    }
    fclose(f);
    *A_out = A;
    return n;
}

/* initRHS, OE:88-106: X__[i] = i+1, R = A * X__ accumulated in ascending j. */
void oracle_init_rhs(const double *A, double *R, double *Xpre, int n)
{
    for (int i = 0; i < n; ++i) Xpre[i] = (double)(i + 1);
    for (int i = 0; i < n; ++i) {
        double acc = 0.0;
        for (int j = 0; j < n; ++j) acc += A[(size_t)i * n + j] * Xpre[j];
        R[i] = acc;
    }
}

/* getPivot with partial pivoting, OE:123-148: the row with the largest |a[i][cur]|, i >= cur (first one wins
 * ties, strict >), is swapped in (tails [cur, n) and R); a zero column is singular. */
int oracle_get_pivot_partial(double *A, double *R, int n, int cur, int *chosen)
{
    int pr = cur;
    for (int i = cur + 1; i < n; ++i)
        if (fabs(A[(size_t)i * n + cur]) > fabs(A[(size_t)pr * n + cur])) pr = i;
    if (chosen) *chosen = pr;
    if (fabs(A[(size_t)pr * n + cur]) == 0.0) return ORACLE_SINGULAR;
    if (pr != cur) {
        for (int c = cur; c < n; ++c) { double t = A[(size_t)pr * n + c]; A[(size_t)pr * n + c] = A[(size_t)cur * n + c]; A[(size_t)cur * n + c] = t; }
        double t = R[pr]; R[pr] = R[cur]; R[cur] = t;
    }
    return ORACLE_OK;
}

/* computeGauss of the pivoting variant, OE:153-182: pivot, scale the pivot row (OE:161-169), eliminate
 * (OE:172-181).  perm (may be NULL) receives the pivot row chosen at every step. */
int oracle_compute_gauss_pivot(double *A, double *R, int *perm, int n, int nthreads)
{
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#else
    (void)nthreads;
#endif
    for (int i = 0; i < n; ++i) {
        int rc = oracle_get_pivot_partial(A, R, n, i, perm ? perm + i : NULL);
        if (rc != ORACLE_OK) return rc;
        double *pi = A + (size_t)i * n;
        const double piv = pi[i];
        if (piv != 1.0) {
            pi[i] = 1.0;
            for (int j = i + 1; j < n; ++j) pi[j] /= piv;
            R[i] /= piv;
        }
        const double ri = R[i];
#pragma omp parallel for schedule(static)
        for (int j = i + 1; j < n; ++j) {
            double *pj = A + (size_t)j * n;
            const double m = pj[i];
            pj[i] = 0.0;
            for (int k = i + 1; k < n; ++k) pj[k] -= m * pi[k];
            R[j] -= m * ri;
        }
    }
    return ORACLE_OK;
}

/* Whole external-input solve: elimination with pivoting + solveGauss (OE:186-205, same recurrence as OI's). */
int oracle_gauss_pivot(double *A, double *R, double *X, int n, int nthreads)
{
    int rc = oracle_compute_gauss_pivot(A, R, NULL, n, nthreads);
    if (rc != ORACLE_OK) return rc;
    X[n - 1] = R[n - 1];
    for (int i = n - 2; i >= 0; --i) {
        double v = R[i];
        for (int j = n - 1; j > i; --j) v -= A[(size_t)i * n + j] * X[j];
        X[i] = v;
    }
    return ORACLE_OK;
}
