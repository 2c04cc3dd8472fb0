/* miniblas.c -- the eight BLAS/LAPACK Fortran symbols the reference links
 * against (src/LinearAlgebra/d4est_blas.h:4-19), written from their published
 * interface semantics (column-major, all arguments by pointer).  The reference
 * normally gets them from OpenBLAS 0.3.20, whose tarball is absent.
 * TEST INFRASTRUCTURE for oracle/_ref only. */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

static int
is_t (const char *c)
{
  return (*c == 'T' || *c == 't' || *c == 'C' || *c == 'c');
}

/* C := alpha*op(A)*op(B) + beta*C */
void
dgemm_ (const char *transa, const char *transb, const int *M, const int *N, const int *K,
        const double *alpha, const double *A, const int *lda, const double *B,
        const int *ldb, const double *beta, double *C, const int *ldc)
{
  int m = *M, n = *N, k = *K, ta = is_t (transa), tb = is_t (transb);
  for (int j = 0; j < n; j++) {
    for (int i = 0; i < m; i++) {
      double s = 0.;
      for (int l = 0; l < k; l++) {
        double a = ta ? A[l + (size_t) i * *lda] : A[i + (size_t) l * *lda];
        double b = tb ? B[j + (size_t) l * *ldb] : B[l + (size_t) j * *ldb];
        s += a * b;
      }
      double *c = &C[i + (size_t) j * *ldc];
      *c = (*beta == 0.) ? (*alpha) * s : (*alpha) * s + (*beta) * (*c);
    }
  }
}

/* y := alpha*op(A)*x + beta*y */
void
dgemv_ (const char *trans, const int *M, const int *N, const double *alpha, const double *A,
        const int *lda, const double *x, const int *incx, const double *beta, double *y,
        const int *incy)
{
  int m = *M, n = *N, t = is_t (trans);
  int leny = t ? n : m, lenx = t ? m : n;
  for (int i = 0; i < leny; i++) {
    double s = 0.;
    for (int l = 0; l < lenx; l++) {
      double a = t ? A[l + (size_t) i * *lda] : A[i + (size_t) l * *lda];
      s += a * x[(size_t) l * *incx];
    }
    double *yy = &y[(size_t) i * *incy];
    *yy = (*beta == 0.) ? (*alpha) * s : (*alpha) * s + (*beta) * (*yy);
  }
}

void
daxpy_ (const int *N, const double *alpha, const double *x, const int *incx, double *y,
        const int *incy)
{
  for (int i = 0; i < *N; i++) y[(size_t) i * *incy] += (*alpha) * x[(size_t) i * *incx];
}

void
dscal_ (const int *N, const double *alpha, double *x, const int *incx)
{
  for (int i = 0; i < *N; i++) x[(size_t) i * *incx] *= (*alpha);
}

double
ddot_ (const int *N, const double *x, const int *incx, const double *y, const int *incy)
{
  double s = 0.;
  for (int i = 0; i < *N; i++) s += x[(size_t) i * *incx] * y[(size_t) i * *incy];
  return s;
}

/* LU with partial pivoting, column-major, 1-based pivots */
void
dgetrf_ (const int *M, const int *N, double *A, const int *lda, int *ipiv, int *info)
{
  int m = *M, n = *N, ld = *lda;
  int mn = m < n ? m : n;
  *info = 0;
  for (int j = 0; j < mn; j++) {
    int p = j;
    double big = fabs (A[j + (size_t) j * ld]);
    for (int i = j + 1; i < m; i++)
      if (fabs (A[i + (size_t) j * ld]) > big) { big = fabs (A[i + (size_t) j * ld]); p = i; }
    ipiv[j] = p + 1;
    if (big == 0.) { if (!*info) *info = j + 1; continue; }
    if (p != j)
      for (int c = 0; c < n; c++) {
        double t = A[j + (size_t) c * ld];
        A[j + (size_t) c * ld] = A[p + (size_t) c * ld];
        A[p + (size_t) c * ld] = t;
      }
    double inv = 1. / A[j + (size_t) j * ld];
    for (int i = j + 1; i < m; i++) A[i + (size_t) j * ld] *= inv;
    for (int c = j + 1; c < n; c++) {
      double ajc = A[j + (size_t) c * ld];
      for (int i = j + 1; i < m; i++) A[i + (size_t) c * ld] -= A[i + (size_t) j * ld] * ajc;
    }
  }
}

/* inverse from the LU factors of dgetrf_ */
void
dgetri_ (const int *N, double *A, const int *lda, const int *ipiv, double *work,
         const int *lwork, int *info)
{
  int n = *N, ld = *lda;
  (void) work; (void) lwork;
  *info = 0;
  double *inv = (double *) malloc (sizeof (double) * (size_t) n * n);
  double *col = (double *) malloc (sizeof (double) * (size_t) n);
  for (int c = 0; c < n; c++) {
    /* solve A x = e_c with P A = L U */
    for (int i = 0; i < n; i++) col[i] = (i == c) ? 1. : 0.;
    for (int i = 0; i < n; i++) {
      int p = ipiv[i] - 1;
      if (p != i) { double t = col[i]; col[i] = col[p]; col[p] = t; }
    }
    for (int i = 0; i < n; i++)
      for (int l = 0; l < i; l++) col[i] -= A[i + (size_t) l * ld] * col[l];
    for (int i = n - 1; i >= 0; i--) {
      for (int l = i + 1; l < n; l++) col[i] -= A[i + (size_t) l * ld] * col[l];
      col[i] /= A[i + (size_t) i * ld];
    }
    for (int i = 0; i < n; i++) inv[i + (size_t) c * n] = col[i];
  }
  for (int c = 0; c < n; c++)
    for (int i = 0; i < n; i++) A[i + (size_t) c * ld] = inv[i + (size_t) c * n];
  free (inv);
  free (col);
}

/* symmetric eigenproblem by cyclic Jacobi rotations (only used by debugging
 * utilities of the reference, never on the apply / smoother path) */
void
dsyev_ (const char *jobz, const char *uplo, const int *N, double *A, const int *lda,
        double *w, double *work, const int *lwork, int *info)
{
  int n = *N, ld = *lda;
  (void) work;
  *info = 0;
  if (*lwork == -1) { work[0] = 1; return; }
  double *V = (double *) calloc ((size_t) n * n, sizeof (double));
  for (int i = 0; i < n; i++) V[i + (size_t) i * n] = 1.;
  int up = (*uplo == 'U' || *uplo == 'u');
  for (int j = 0; j < n; j++)
    for (int i = 0; i < j; i++) {
      double v = up ? A[i + (size_t) j * ld] : A[j + (size_t) i * ld];
      A[i + (size_t) j * ld] = A[j + (size_t) i * ld] = v;
    }
  for (int sweep = 0; sweep < 100; sweep++) {
    double off = 0.;
    for (int j = 0; j < n; j++) for (int i = 0; i < j; i++) off += A[i + (size_t) j * ld] * A[i + (size_t) j * ld];
    if (off < 1e-300) break;
    for (int p = 0; p < n; p++)
      for (int q = p + 1; q < n; q++) {
        double apq = A[p + (size_t) q * ld];
        if (apq == 0.) continue;
        double theta = (A[q + (size_t) q * ld] - A[p + (size_t) p * ld]) / (2. * apq);
        double t = (theta >= 0 ? 1. : -1.) / (fabs (theta) + sqrt (theta * theta + 1.));
        double c = 1. / sqrt (t * t + 1.), s = t * c;
        for (int k = 0; k < n; k++) {
          double akp = A[k + (size_t) p * ld], akq = A[k + (size_t) q * ld];
          A[k + (size_t) p * ld] = c * akp - s * akq;
          A[k + (size_t) q * ld] = s * akp + c * akq;
        }
        for (int k = 0; k < n; k++) {
          double apk = A[p + (size_t) k * ld], aqk = A[q + (size_t) k * ld];
          A[p + (size_t) k * ld] = c * apk - s * aqk;
          A[q + (size_t) k * ld] = s * apk + c * aqk;
        }
        for (int k = 0; k < n; k++) {
          double vkp = V[k + (size_t) p * n], vkq = V[k + (size_t) q * n];
          V[k + (size_t) p * n] = c * vkp - s * vkq;
          V[k + (size_t) q * n] = s * vkp + c * vkq;
        }
      }
  }
  /* sort ascending */
  int *idx = (int *) malloc (sizeof (int) * (size_t) n);
  for (int i = 0; i < n; i++) { w[i] = A[i + (size_t) i * ld]; idx[i] = i; }
  for (int i = 0; i < n; i++)
    for (int j = i + 1; j < n; j++)
      if (w[idx[j]] < w[idx[i]]) { int t = idx[i]; idx[i] = idx[j]; idx[j] = t; }
  double *ws = (double *) malloc (sizeof (double) * (size_t) n);
  for (int i = 0; i < n; i++) ws[i] = w[idx[i]];
  if (*jobz == 'V' || *jobz == 'v')
    for (int j = 0; j < n; j++)
      for (int i = 0; i < n; i++) A[i + (size_t) j * ld] = V[i + (size_t) idx[j] * n];
  memcpy (w, ws, sizeof (double) * (size_t) n);
  free (ws); free (idx); free (V);
}
