/* Extended-precision (x87 long double, 64-bit mantissa) restatement of the exact-GPR path.
 *
 * TEST INFRASTRUCTURE ONLY (oracle/): the reference of record for the conditioning sweep
 * (tests/test_gpu_parity.py::test_i8_precision_envelope, tools/precision_envelope.py).  It answers the question
 * "what can ANY float64 implementation attain at this condition number": the NumPy/LAPACK float64 oracle
 * (oracle/gpr.py), the CUDA FP64 mma.sync path and the int8-sliced default path are all measured against it.
 *
 * Same formulas as oracle/gpr.py, i.e. what the reference asks GPflow for:
 *   kernels   gpflow.kernels.{Matern12,Matern32,Matern52,SquaredExponential}, ARD (gpflow_models.py:116-135)
 *   lml       GPR.log_marginal_likelihood  (gpflow_models.py:337)
 *   gradient  d lml / d theta_p = 1/2 tr((a a^T - Ks^-1) dKs/d theta_p)   (TF autodiff of :317-321)
 *   predict   GPR.predict_f  (gpflow_models.py:229-230)
 * Inputs and outputs are float64; every intermediate is long double.  Build: oracle/Makefile (gcc -O2 -fopenmp).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef long double ld;

static ld k_of_r2(int kind, ld r2, ld var) {
  if (kind == 3) return var * expl(-0.5L * r2);
  ld r = sqrtl(r2 > 1e-36L ? r2 : 1e-36L);
  if (kind == 0) return var * expl(-r);
  if (kind == 1) { ld s = sqrtl(3.0L) * r; return var * (1.0L + s) * expl(-s); }
  ld s = sqrtl(5.0L) * r;
  return var * (1.0L + s + 5.0L / 3.0L * r * r) * expl(-s);
}

/* g(r) with dk/dl_d = g(r) (x_d - x'_d)^2 / l_d^3 (zero on the clamped branch of Matern12) */
static ld g_of_r2(int kind, ld r2, ld var) {
  if (kind == 3) return var * expl(-0.5L * r2);
  ld r = sqrtl(r2 > 1e-36L ? r2 : 1e-36L);
  if (kind == 0) return r2 > 1e-36L ? var * expl(-r) / r : 0.0L;
  if (kind == 1) return 3.0L * var * expl(-sqrtl(3.0L) * r);
  ld s = sqrtl(5.0L) * r;
  return 5.0L / 3.0L * var * (1.0L + s) * expl(-s);
}

static ld sqdist(const double* a, const double* b, const ld* inv_ls, int D) {
  ld r2 = 0.0L;
  for (int d = 0; d < D; ++d) { ld t = ((ld)a[d] - (ld)b[d]) * inv_ls[d]; r2 += t * t; }
  return r2;
}

/* kind: 0 Matern12, 1 Matern32, 2 Matern52, 3 RBF.  theta = [ls_1..ls_D, var, noise].
 * grad (may be NULL) [D+2]; fmean / fvar (may be NULL when m == 0) [m].
 * Returns 0, 1 when Ks is not positive definite, 2 on allocation failure. */
int ld_gpr(int kind, int n, int D, int m, const double* X, const double* y, const double* Xs, const double* theta,
           double* lml, double* grad, double* fmean, double* fvar) {
  ld inv_ls[16];
  for (int d = 0; d < D; ++d) inv_ls[d] = 1.0L / (ld)theta[d];
  const ld var = theta[D], noise = theta[D + 1];
  const size_t N = (size_t)n;
  ld* L = (ld*)malloc(N * N * sizeof(ld));
  ld* a = (ld*)malloc(N * sizeof(ld));
  ld* alpha = (ld*)malloc(N * sizeof(ld));
  if (!L || !a || !alpha) { free(L); free(a); free(alpha); return 2; }
#pragma omp parallel for schedule(dynamic, 16)
  for (int i = 0; i < n; ++i)
    for (int j = 0; j <= i; ++j)
      L[i * N + j] = k_of_r2(kind, sqdist(X + (size_t)i * D, X + (size_t)j * D, inv_ls, D), var) + (i == j ? noise : 0.0L);
  /* left-looking Cholesky, rows are contiguous so every inner product is unit-stride */
  int bad = 0;
  for (int j = 0; j < n && !bad; ++j) {
    ld s = L[j * N + j];
    for (int k = 0; k < j; ++k) s -= L[j * N + k] * L[j * N + k];
    if (!(s > 0.0L)) { bad = 1; break; }
    const ld ljj = sqrtl(s);
    L[j * N + j] = ljj;
#pragma omp parallel for schedule(static)
    for (int i = j + 1; i < n; ++i) {
      ld t = L[i * N + j];
      const ld* li = L + i * N; const ld* lj = L + j * N;
      for (int k = 0; k < j; ++k) t -= li[k] * lj[k];
      L[i * N + j] = t / ljj;
    }
  }
  if (bad) { free(L); free(a); free(alpha); return 1; }
  ld quad = 0.0L, logdet = 0.0L;
  for (int i = 0; i < n; ++i) {
    ld t = y[i];
    for (int k = 0; k < i; ++k) t -= L[i * N + k] * a[k];
    a[i] = t / L[i * N + i];
    quad += a[i] * a[i];
    logdet += logl(L[i * N + i]);
  }
  *lml = (double)(-0.5L * quad - logdet - 0.5L * n * logl(2.0L * acosl(-1.0L)));
  for (int i = n - 1; i >= 0; --i) {
    ld t = a[i];
    for (int k = i + 1; k < n; ++k) t -= L[k * N + i] * alpha[k];
    alpha[i] = t / L[i * N + i];
  }
  if (grad) {
    /* Xt[j][i] = (L^-1)[i][j], i >= j: columns of L^-1 stored as rows */
    ld* Xt = (ld*)calloc(N * N, sizeof(ld));
    if (!Xt) { free(L); free(a); free(alpha); return 2; }
#pragma omp parallel for schedule(dynamic, 8)
    for (int j = 0; j < n; ++j) {
      ld* xj = Xt + j * N;
      xj[j] = 1.0L / L[j * N + j];
      for (int i = j + 1; i < n; ++i) {
        ld t = 0.0L;
        const ld* li = L + i * N;
        for (int k = j; k < i; ++k) t -= li[k] * xj[k];
        xj[i] = t / li[i];
      }
    }
    ld g[18];
    for (int p = 0; p < D + 2; ++p) g[p] = 0.0L;
#pragma omp parallel
    {
      ld gl[18];
      for (int p = 0; p < D + 2; ++p) gl[p] = 0.0L;
#pragma omp for schedule(dynamic, 8)
      for (int i = 0; i < n; ++i) {
        for (int j = 0; j <= i; ++j) {
          /* Ks^-1[i][j] = sum_{k >= i} Linv[k][i] Linv[k][j] */
          ld kin = 0.0L;
          const ld* xi = Xt + i * N; const ld* xj = Xt + j * N;
          for (int k = i; k < n; ++k) kin += xi[k] * xj[k];
          const ld q = alpha[i] * alpha[j] - kin;
          const ld w = i == j ? q : 2.0L * q;
          const ld r2 = sqdist(X + (size_t)i * D, X + (size_t)j * D, inv_ls, D);
          const ld gg = g_of_r2(kind, r2, var);
          for (int d = 0; d < D; ++d) {
            ld dx = (ld)X[(size_t)i * D + d] - (ld)X[(size_t)j * D + d];
            gl[d] += w * gg * dx * dx;
          }
          gl[D] += w * k_of_r2(kind, r2, var);
          if (i == j) gl[D + 1] += q;
        }
      }
#pragma omp critical
      for (int p = 0; p < D + 2; ++p) g[p] += gl[p];
    }
    for (int d = 0; d < D; ++d) grad[d] = (double)(0.5L * g[d] * inv_ls[d] * inv_ls[d] * inv_ls[d]);
    grad[D] = (double)(0.5L * g[D] / var);
    grad[D + 1] = (double)(0.5L * g[D + 1]);
    free(Xt);
  }
  if (m > 0 && fmean && fvar) {
#pragma omp parallel
    {
      ld* c = (ld*)malloc(N * sizeof(ld));
#pragma omp for schedule(dynamic, 1)
      for (int s = 0; s < m; ++s) {
        ld mean = 0.0L, ss = 0.0L;
        for (int i = 0; i < n; ++i) {
          ld t = k_of_r2(kind, sqdist(X + (size_t)i * D, Xs + (size_t)s * D, inv_ls, D), var);
          for (int k = 0; k < i; ++k) t -= L[i * N + k] * c[k];
          c[i] = t / L[i * N + i];
          mean += c[i] * a[i];
          ss += c[i] * c[i];
        }
        fmean[s] = (double)mean;
        fvar[s] = (double)(var - ss);
      }
      free(c);
    }
  }
  free(L); free(a); free(alpha);
  return 0;
}

/* ---- sparse GPR (gpflow.models.SGPR.elbo / predict_f; call sites gpflow_models.py:825-830, 860-862) in extended
 * precision.  Same formulas as oracle/sgpr.py:
 *   Kuu + jitter I = L L^T;  A = L^-1 Kuf / sigma;  B = A A^T + I = LB LB^T;  c = LB^-1 A y / sigma
 *   elbo = -n/2 log 2pi - sum log diag LB - n/2 log sigma^2 - y^T y / (2 sigma^2) + c^T c / 2
 *          - n var / (2 sigma^2) + tr(A A^T) / 2
 *   f* = Kus^T L^-T LB^-T c;   f*_var = var + |LB^-1 L^-1 k_s|^2 - |L^-1 k_s|^2
 * Returns 0, 1 when Kuu + jitter I or B is not positive definite, 2 on allocation failure. */
static int ld_chol(ld* A, int n) {          /* in place, lower, row-major [n][n] */
  const size_t N = (size_t)n;
  for (int j = 0; j < n; ++j) {
    ld s = A[j * N + j];
    for (int k = 0; k < j; ++k) s -= A[j * N + k] * A[j * N + k];
    if (!(s > 0.0L)) return 1;
    const ld ljj = sqrtl(s);
    A[j * N + j] = ljj;
#pragma omp parallel for schedule(static)
    for (int i = j + 1; i < n; ++i) {
      ld t = A[i * N + j];
      const ld* ai = A + i * N; const ld* aj = A + j * N;
      for (int k = 0; k < j; ++k) t -= ai[k] * aj[k];
      A[i * N + j] = t / ljj;
    }
  }
  return 0;
}

int ld_sgpr(int kind, int n, int D, int M, int m, const double* X, const double* y, const double* Z, const double* Xs,
            const double* theta, double jitter, double* elbo, double* fmean, double* fvar) {
  ld inv_ls[16];
  for (int d = 0; d < D; ++d) inv_ls[d] = 1.0L / (ld)theta[d];
  const ld var = theta[D], noise = theta[D + 1], sigma = sqrtl(noise);
  const size_t MM = (size_t)M, NN = (size_t)n;
  ld* L = (ld*)malloc(MM * MM * sizeof(ld));
  ld* At = (ld*)malloc(NN * MM * sizeof(ld));      /* [n][M]: column j of A */
  ld* A = (ld*)malloc(MM * NN * sizeof(ld));       /* [M][n] */
  ld* B = (ld*)malloc(MM * MM * sizeof(ld));
  ld* Ay = (ld*)malloc(MM * sizeof(ld));
  ld* c = (ld*)malloc(MM * sizeof(ld));
  if (!L || !At || !A || !B || !Ay || !c) { free(L); free(At); free(A); free(B); free(Ay); free(c); return 2; }
  int rc = 0;
#pragma omp parallel for schedule(dynamic, 16)
  for (int i = 0; i < M; ++i)
    for (int j = 0; j <= i; ++j)
      L[i * MM + j] = k_of_r2(kind, sqdist(Z + (size_t)i * D, Z + (size_t)j * D, inv_ls, D), var) + (i == j ? (ld)jitter : 0.0L);
  rc = ld_chol(L, M);
  if (!rc) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int j = 0; j < n; ++j) {
      ld* a = At + (size_t)j * MM;
      for (int i = 0; i < M; ++i) {
        ld t = k_of_r2(kind, sqdist(Z + (size_t)i * D, X + (size_t)j * D, inv_ls, D), var);
        const ld* li = L + i * MM;
        for (int k = 0; k < i; ++k) t -= li[k] * a[k];
        a[i] = t / li[i];
      }
      for (int i = 0; i < M; ++i) a[i] /= sigma;
    }
#pragma omp parallel for schedule(static)
    for (int i = 0; i < M; ++i)
      for (int j = 0; j < n; ++j) A[i * NN + j] = At[(size_t)j * MM + i];
    ld trAAt = 0.0L;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : trAAt)
    for (int p = 0; p < M; ++p) {
      const ld* ap = A + p * NN;
      for (int q = 0; q <= p; ++q) {
        const ld* aq = A + q * NN;
        ld s = 0.0L;
        for (int j = 0; j < n; ++j) s += ap[j] * aq[j];
        if (p == q) { trAAt += s; s += 1.0L; }
        B[p * MM + q] = s;
      }
      ld t = 0.0L;
      for (int j = 0; j < n; ++j) t += ap[j] * (ld)y[j];
      Ay[p] = t / sigma;
    }
    rc = ld_chol(B, M);
    if (!rc) {
      ld logdet = 0.0L, cc = 0.0L, yy = 0.0L;
      for (int i = 0; i < M; ++i) {
        ld t = Ay[i];
        for (int k = 0; k < i; ++k) t -= B[i * MM + k] * c[k];
        c[i] = t / B[i * MM + i];
        cc += c[i] * c[i];
        logdet += logl(B[i * MM + i]);
      }
      for (int j = 0; j < n; ++j) yy += (ld)y[j] * (ld)y[j];
      const ld log2pi = logl(2.0L * acosl(-1.0L));
      *elbo = (double)(-0.5L * n * log2pi - logdet - 0.5L * n * logl(noise) - 0.5L * yy / noise + 0.5L * cc
                       - 0.5L * n * var / noise + 0.5L * trAAt);
      if (m > 0 && fmean && fvar) {
        /* w = L^-T LB^-T c, so that f* = k_s^T w */
        ld* w = (ld*)malloc(MM * sizeof(ld));
        ld* u = (ld*)malloc(MM * sizeof(ld));
        for (int i = M - 1; i >= 0; --i) {
          ld t = c[i];
          for (int k = i + 1; k < M; ++k) t -= B[k * MM + i] * u[k];
          u[i] = t / B[i * MM + i];
        }
        for (int i = M - 1; i >= 0; --i) {
          ld t = u[i];
          for (int k = i + 1; k < M; ++k) t -= L[k * MM + i] * w[k];
          w[i] = t / L[i * MM + i];
        }
#pragma omp parallel
        {
          ld* t1 = (ld*)malloc(MM * sizeof(ld));
          ld* t2 = (ld*)malloc(MM * sizeof(ld));
          ld* ks = (ld*)malloc(MM * sizeof(ld));
#pragma omp for schedule(dynamic, 4)
          for (int s = 0; s < m; ++s) {
            ld mean = 0.0L, s1 = 0.0L, s2 = 0.0L;
            for (int i = 0; i < M; ++i) {
              ks[i] = k_of_r2(kind, sqdist(Z + (size_t)i * D, Xs + (size_t)s * D, inv_ls, D), var);
              mean += ks[i] * w[i];
              ld t = ks[i];
              for (int k = 0; k < i; ++k) t -= L[i * MM + k] * t1[k];
              t1[i] = t / L[i * MM + i];
              s1 += t1[i] * t1[i];
            }
            for (int i = 0; i < M; ++i) {
              ld t = t1[i];
              for (int k = 0; k < i; ++k) t -= B[i * MM + k] * t2[k];
              t2[i] = t / B[i * MM + i];
              s2 += t2[i] * t2[i];
            }
            fmean[s] = (double)mean;
            fvar[s] = (double)(var + s2 - s1);
          }
          free(t1); free(t2); free(ks);
        }
        free(w); free(u);
      }
    }
  }
  free(L); free(At); free(A); free(B); free(Ay); free(c);
  return rc;
}
