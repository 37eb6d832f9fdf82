/* lapack_jacobi.c — fallback sgesvd_/dgesvd_ (Fortran ABI) for the oracle/_ref build when no LAPACK is found.
 * TEST INFRASTRUCTURE (see oracle/README). The reference reaches LAPACK through TooN::SVD
 * (TooN/TooN/SVD.h:122-163, call sites rebvio/src/core.cpp:247-248 and sab_estimator.cpp:31-32); LAPACK itself
 * is an un-vendored system dependency of the reference (rebvio/CMakeLists.txt:19,163). Only the shape TooN uses
 * is implemented: square n x n, JOBU='S' (U returned in u), JOBVT='O' (V^T overwrites a), column major, plus the
 * LWORK=-1 workspace query. Algorithm: one-sided Jacobi (Hestenes) on the columns in double precision,
 * singular values sorted descending. */
#include <math.h>
#include <stdlib.h>
#include <string.h>

static void jacobi_svd(int n, double* A /* col-major n x n, in: matrix, out: U*S */, double* V /* n x n col-major */) {
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) V[i + j * n] = (i == j);
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < n - 1; ++p) {
      for (int q = p + 1; q < n; ++q) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int i = 0; i < n; ++i) {
          alpha += A[i + p * n] * A[i + p * n];
          beta += A[i + q * n] * A[i + q * n];
          gamma += A[i + p * n] * A[i + q * n];
        }
        if (gamma == 0.0) continue;
        double lim = fabs(gamma) / sqrt(alpha * beta);
        if (lim > off) off = lim;
        if (lim < 1e-17) continue;
        double zeta = (beta - alpha) / (2.0 * gamma);
        double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
        for (int i = 0; i < n; ++i) {
          double ap = A[i + p * n], aq = A[i + q * n];
          A[i + p * n] = c * ap - s * aq;
          A[i + q * n] = s * ap + c * aq;
          double vp = V[i + p * n], vq = V[i + q * n];
          V[i + p * n] = c * vp - s * vq;
          V[i + q * n] = s * vp + c * vq;
        }
      }
    }
    if (off < 1e-15) break;
  }
}

static void gesvd_core(int n, const double* Ain, int lda, double* S, double* U, int ldu, double* VTout, int ldvt) {
  double* A = (double*)malloc(sizeof(double) * n * n);
  double* V = (double*)malloc(sizeof(double) * n * n);
  for (int j = 0; j < n; ++j) for (int i = 0; i < n; ++i) A[i + j * n] = Ain[i + j * lda];
  jacobi_svd(n, A, V);
  int* order = (int*)malloc(sizeof(int) * n);
  double* sv = (double*)malloc(sizeof(double) * n);
  for (int j = 0; j < n; ++j) {
    double s = 0;
    for (int i = 0; i < n; ++i) s += A[i + j * n] * A[i + j * n];
    sv[j] = sqrt(s);
    order[j] = j;
  }
  for (int a = 0; a < n; ++a) for (int b = a + 1; b < n; ++b) if (sv[order[b]] > sv[order[a]]) { int t = order[a]; order[a] = order[b]; order[b] = t; }
  for (int k = 0; k < n; ++k) {
    int j = order[k];
    S[k] = sv[j];
    for (int i = 0; i < n; ++i) {
      U[i + k * ldu] = sv[j] > 0 ? A[i + j * n] / sv[j] : (i == k);
      VTout[k + i * ldvt] = V[i + j * n];
    }
  }
  free(A); free(V); free(order); free(sv);
}

void dgesvd_(const char* jobu, const char* jobvt, int* m, int* n, double* a, int* lda, double* s, double* u, int* ldu,
             double* vt, int* ldvt, double* work, int* lwork, int* info) {
  (void)vt; (void)ldvt;
  *info = 0;
  if (*lwork == -1) { work[0] = 1; return; }
  if (*m != *n || *jobu != 'S' || *jobvt != 'O') { *info = -1; return; }
  int N = *n;
  double* vtmp = (double*)malloc(sizeof(double) * N * N);
  gesvd_core(N, a, *lda, s, u, *ldu, vtmp, N);
  for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i) a[i + j * (*lda)] = vtmp[i + j * N];
  free(vtmp);
}

void sgesvd_(const char* jobu, const char* jobvt, int* m, int* n, float* a, int* lda, float* s, float* u, int* ldu,
             float* vt, int* ldvt, float* work, int* lwork, int* info) {
  (void)vt; (void)ldvt;
  *info = 0;
  if (*lwork == -1) { work[0] = 1; return; }
  if (*m != *n || *jobu != 'S' || *jobvt != 'O') { *info = -1; return; }
  int N = *n;
  double* ad = (double*)malloc(sizeof(double) * N * N * 3 + sizeof(double) * N);
  double* ud = ad + N * N; double* vd = ud + N * N; double* sd = vd + N * N;
  for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i) ad[i + j * N] = a[i + j * (*lda)];
  gesvd_core(N, ad, N, sd, ud, N, vd, N);
  for (int j = 0; j < N; ++j) {
    s[j] = (float)sd[j];
    for (int i = 0; i < N; ++i) { u[i + j * (*ldu)] = (float)ud[i + j * N]; a[i + j * (*lda)] = (float)vd[i + j * N]; }
  }
  free(ad);
}
