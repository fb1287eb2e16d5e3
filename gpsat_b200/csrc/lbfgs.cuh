// Per-expert L-BFGS state machine (reverse communication), host+device.
//
// Replaces the reference's per-expert SciPy loop:
//   GPSat/models/gpflow_models.py:317-321  gpflow.optimizers.Scipy().minimize(...)
//     -> scipy.optimize.minimize(method="L-BFGS-B", jac=True) on the *unconstrained* vector
//        (constraints live in the bijectors, so SciPy sees no bounds).
// For an unbounded problem L-BFGS-B (Byrd, Lu, Nocedal, Zhu 1995; v3.0 2011) reduces to:
//   first direction  d = -g with trial step min(1/|d|, 1e10); afterwards d = -H g with the
//   limited-memory BFGS inverse H (m = 10 pairs, H0 = (s'y / y'y) I) and trial step 1;
//   More'-Thuente line search `dcsrch` (ftol 1e-3, gtol 0.9, xtol 0.1, stpmin 0, stpmax 1e10),
//   at most `maxls` = 20 trial points per iteration; pair skipped when s'y <= eps * (-g'd * stp);
//   memory reset + steepest-descent restart when the line search fails; stop when
//   max|g| <= pgtol (1e-5) or (f_k - f_{k+1}) <= ftol * max(|f_k|, |f_{k+1}|, 1) (ftol 2.22e-9),
//   or after maxiter iterations / maxfun evaluations (SciPy wrapper semantics,
//   scipy/optimize/_lbfgsb_py.py).
// dcsrch/dcstep follow MINPACK-2 (More' & Thuente 1994) as published.
//
// One state per expert ("slot"); `lbfgs_step` consumes (f, g) at `xtrial` and either asks for
// another evaluation at a new `xtrial` or terminates.  Pure arithmetic on <= 8 variables: it
// runs as one thread per expert on the device (see engine.cu) and is unit-tested on the host.
#pragma once
#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define GPSAT_HD __host__ __device__ __forceinline__
#else
#define GPSAT_HD inline
#endif

#define GPSAT_PMAX 8       // max number of optimised variables per expert (D + 2 <= 8)
#define GPSAT_LBFGS_M 10   // SciPy maxcor default

enum LbfgsReturn : int {
  LBFGS_CONTINUE = 0,          // evaluate f, g at st->xtrial and call lbfgs_step again
  LBFGS_CONV_PGTOL = 1,        // CONVERGENCE: NORM_OF_PROJECTED_GRADIENT_<=_PGTOL
  LBFGS_CONV_FTOL = 2,         // CONVERGENCE: REL_REDUCTION_OF_F_<=_FACTR*EPSMCH
  LBFGS_STOP_MAXITER = 3,      // STOP: TOTAL NO. of ITERATIONS REACHED LIMIT
  LBFGS_STOP_MAXFUN = 4,       // STOP: TOTAL NO. of f AND g EVALUATIONS EXCEEDS LIMIT
  LBFGS_ABNORMAL = 5,          // ABNORMAL_TERMINATION_IN_LNSRCH
  LBFGS_NOT_FINITE = 6         // objective not finite at the starting point (K + s2 I not PD)
};

struct LbfgsOptions {
  double pgtol;    // 1e-5
  double ftol;     // 2.220446049250313e-09  (= factr * epsmch)
  int maxiter;     // 10000 in the reference (gpflow_models.py:291)
  int maxfun;      // 15000
  int maxls;       // 20
};

struct LbfgsState {
  int np;                 // number of free variables
  int started;            // 0 until the first (f, g) has been consumed
  int iter;               // completed iterations
  int nfev;               // evaluations consumed
  int col;                // stored correction pairs
  int head;               // ring index of the oldest pair
  int ifun, iback;        // line-search counters (lnsrlb)
  // accepted iterate
  double x[GPSAT_PMAX], g[GPSAT_PMAX], f;
  // point to evaluate next
  double xtrial[GPSAT_PMAX];
  // line search bookkeeping: t = x_k, r = g_k (later y_k), d = direction (later s_k)
  double t[GPSAT_PMAX], r[GPSAT_PMAX], d[GPSAT_PMAX];
  double fold, stp, dnorm, gd, gdold;
  // limited-memory pairs
  double S[GPSAT_LBFGS_M][GPSAT_PMAX], Y[GPSAT_LBFGS_M][GPSAT_PMAX];
  double rho[GPSAT_LBFGS_M];
  double theta;           // y'y / s'y of the newest pair
  // dcsrch state
  int brackt, stage, ls_started;
  double ginit, gtest, gx, gy, finit, fx, fy, stx, sty, stmin, stmax, width, width1;
};

GPSAT_HD double lbfgs_dot(const double* a, const double* b, int n) {
  double s = 0.0;
  for (int i = 0; i < n; ++i) s += a[i] * b[i];
  return s;
}

GPSAT_HD double lbfgs_clip(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }

// MINPACK-2 dcstep: safeguarded cubic/quadratic step and interval update.
GPSAT_HD void lbfgs_dcstep(double& stx, double& fx, double& dx, double& sty, double& fy, double& dy,
                           double& stp, double fp, double dp, int& brackt, double stpmin,
                           double stpmax) {
  double sgnd = dp * (dx / fabs(dx));
  double stpf, stpc, stpq, theta, s, gamma, p, q, r;
  if (fp > fx) {
    theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
    gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
    if (stp < stx) gamma = -gamma;
    p = (gamma - dx) + theta;
    q = ((gamma - dx) + gamma) + dp;
    r = p / q;
    stpc = stx + r * (stp - stx);
    stpq = stx + ((dx / ((fx - fp) / (stp - stx) + dx)) / 2.0) * (stp - stx);
    if (fabs(stpc - stx) <= fabs(stpq - stx)) stpf = stpc;
    else stpf = stpc + (stpq - stpc) / 2.0;
    brackt = 1;
  } else if (sgnd < 0.0) {
    theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
    gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
    if (stp > stx) gamma = -gamma;
    p = (gamma - dp) + theta;
    q = ((gamma - dp) + gamma) + dx;
    r = p / q;
    stpc = stp + r * (stx - stp);
    stpq = stp + (dp / (dp - dx)) * (stx - stp);
    if (fabs(stpc - stp) > fabs(stpq - stp)) stpf = stpc;
    else stpf = stpq;
    brackt = 1;
  } else if (fabs(dp) < fabs(dx)) {
    theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
    gamma = s * sqrt(fmax(0.0, (theta / s) * (theta / s) - (dx / s) * (dp / s)));
    if (stp > stx) gamma = -gamma;
    p = (gamma - dp) + theta;
    q = (gamma + (dx - dp)) + gamma;
    r = p / q;
    if (r < 0.0 && gamma != 0.0) stpc = stp + r * (stx - stp);
    else if (stp > stx) stpc = stpmax;
    else stpc = stpmin;
    stpq = stp + (dp / (dp - dx)) * (stx - stp);
    if (brackt) {
      if (fabs(stpc - stp) < fabs(stpq - stp)) stpf = stpc;
      else stpf = stpq;
      if (stp > stx) stpf = fmin(stp + 0.66 * (sty - stp), stpf);
      else stpf = fmax(stp + 0.66 * (sty - stp), stpf);
    } else {
      if (fabs(stpc - stp) > fabs(stpq - stp)) stpf = stpc;
      else stpf = stpq;
      stpf = lbfgs_clip(stpf, stpmin, stpmax);
    }
  } else {
    if (brackt) {
      theta = 3.0 * (fp - fy) / (sty - stp) + dy + dp;
      s = fmax(fabs(theta), fmax(fabs(dy), fabs(dp)));
      gamma = s * sqrt((theta / s) * (theta / s) - (dy / s) * (dp / s));
      if (stp > sty) gamma = -gamma;
      p = (gamma - dp) + theta;
      q = ((gamma - dp) + gamma) + dy;
      r = p / q;
      stpc = stp + r * (sty - stp);
      stpf = stpc;
      if (!isfinite(stpf)) stpf = stp + 0.5 * (sty - stp);   // sty was a non-finite trial
    } else if (stp > stx) stpf = stpmax;
    else stpf = stpmin;
  }
  if (fp > fx) {
    sty = stp; fy = fp; dy = dp;
  } else {
    if (sgnd < 0.0) { sty = stx; fy = fx; dy = dx; }
    stx = stp; fx = fp; dx = dp;
  }
  stp = stpf;
}

enum DcsrchTask : int { DCSRCH_FG = 0, DCSRCH_CONV = 1, DCSRCH_WARN = 2, DCSRCH_ERROR = 3 };

// MINPACK-2 dcsrch with the constants L-BFGS-B passes (lnsrlb): ftol 1e-3, gtol 0.9, xtol 0.1,
// stpmin 0, stpmax = stpmx.  `st->ls_started == 0` plays the role of task == 'START'.
GPSAT_HD int lbfgs_dcsrch(LbfgsState* st, double f, double g, double& stp, double stpmx) {
  const double ftol = 1e-3, gtol = 0.9, xtol = 0.1, stpmin = 0.0, stpmax = stpmx;
  const double p5 = 0.5, p66 = 0.66, xtrapl = 1.1, xtrapu = 4.0;
  if (!st->ls_started) {
    if (stp < stpmin || stp > stpmax || g >= 0.0) return DCSRCH_ERROR;
    st->ls_started = 1;
    st->brackt = 0;
    st->stage = 1;
    st->finit = f;
    st->ginit = g;
    st->gtest = ftol * g;
    st->width = stpmax - stpmin;
    st->width1 = st->width / p5;
    st->stx = 0.0; st->fx = f; st->gx = g;
    st->sty = 0.0; st->fy = f; st->gy = g;
    st->stmin = 0.0;
    st->stmax = stp + xtrapu * stp;
    return DCSRCH_FG;
  }
  if (!isfinite(f) || !isfinite(g)) {
    // Not in MINPACK: the reference would raise inside TensorFlow's Cholesky here.  Treat the
    // trial as an infinitely bad upper end of the bracket and bisect towards the best point.
    st->brackt = 1;
    st->sty = stp; st->fy = DBL_MAX; st->gy = DBL_MAX;
    st->width1 = st->width;
    st->width = fabs(st->sty - st->stx);
    st->stmin = fmin(st->stx, st->sty);
    st->stmax = fmax(st->stx, st->sty);
    if (st->width <= 1e-14 * fmax(1.0, st->stmax)) { stp = st->stx; return DCSRCH_WARN; }
    stp = st->stx + p5 * (st->sty - st->stx);
    return DCSRCH_FG;
  }
  double ftest = st->finit + stp * st->gtest;
  if (st->stage == 1 && f <= ftest && g >= 0.0) st->stage = 2;
  int task = DCSRCH_FG;
  if (st->brackt && (stp <= st->stmin || stp >= st->stmax)) task = DCSRCH_WARN;
  if (st->brackt && st->stmax - st->stmin <= xtol * st->stmax) task = DCSRCH_WARN;
  if (stp == stpmax && f <= ftest && g <= st->gtest) task = DCSRCH_WARN;
  if (stp == stpmin && (f > ftest || g >= st->gtest)) task = DCSRCH_WARN;
  if (f <= ftest && fabs(g) <= gtol * (-st->ginit)) task = DCSRCH_CONV;
  if (task != DCSRCH_FG) return task;
  if (st->stage == 1 && f <= st->fx && f > ftest) {
    double fm = f - stp * st->gtest;
    double fxm = st->fx - st->stx * st->gtest;
    double fym = st->fy - st->sty * st->gtest;
    double gm = g - st->gtest;
    double gxm = st->gx - st->gtest;
    double gym = st->gy - st->gtest;
    lbfgs_dcstep(st->stx, fxm, gxm, st->sty, fym, gym, stp, fm, gm, st->brackt, st->stmin, st->stmax);
    st->fx = fxm + st->stx * st->gtest;
    st->fy = fym + st->sty * st->gtest;
    st->gx = gxm + st->gtest;
    st->gy = gym + st->gtest;
  } else {
    lbfgs_dcstep(st->stx, st->fx, st->gx, st->sty, st->fy, st->gy, stp, f, g, st->brackt,
                 st->stmin, st->stmax);
  }
  if (st->brackt) {
    if (fabs(st->sty - st->stx) >= p66 * st->width1) stp = st->stx + p5 * (st->sty - st->stx);
    st->width1 = st->width;
    st->width = fabs(st->sty - st->stx);
  }
  if (st->brackt) {
    st->stmin = fmin(st->stx, st->sty);
    st->stmax = fmax(st->stx, st->sty);
  } else {
    st->stmin = stp + xtrapl * (stp - st->stx);
    st->stmax = stp + xtrapu * (stp - st->stx);
  }
  stp = lbfgs_clip(stp, stpmin, stpmax);
  if ((st->brackt && (stp <= st->stmin || stp >= st->stmax)) ||
      (st->brackt && st->stmax - st->stmin <= xtol * st->stmax))
    stp = st->stx;
  return DCSRCH_FG;
}

GPSAT_HD void lbfgs_init(LbfgsState* st, int np, const double* x0) {
  st->np = np;
  st->started = 0;
  st->iter = 0; st->nfev = 0; st->col = 0; st->head = 0; st->ifun = 0; st->iback = 0;
  st->theta = 1.0;
  st->f = 0.0; st->fold = 0.0; st->stp = 0.0; st->dnorm = 0.0; st->gd = 0.0; st->gdold = 0.0;
  st->ls_started = 0; st->brackt = 0; st->stage = 1;
  for (int i = 0; i < GPSAT_PMAX; ++i) {
    double v = i < np ? x0[i] : 0.0;
    st->x[i] = v; st->xtrial[i] = v; st->g[i] = 0.0; st->t[i] = 0.0; st->r[i] = 0.0; st->d[i] = 0.0;
  }
}

// d = -H g (two-loop recursion; equals the L-BFGS-B subspace-minimisation step when no
// variable is bounded).
GPSAT_HD void lbfgs_direction(LbfgsState* st) {
  const int n = st->np;
  double q[GPSAT_PMAX], a[GPSAT_LBFGS_M];
  for (int i = 0; i < n; ++i) q[i] = st->g[i];
  for (int k = st->col - 1; k >= 0; --k) {
    int j = (st->head + k) % GPSAT_LBFGS_M;
    a[k] = st->rho[j] * lbfgs_dot(st->S[j], q, n);
    for (int i = 0; i < n; ++i) q[i] -= a[k] * st->Y[j][i];
  }
  if (st->col > 0) {
    double h0 = 1.0 / st->theta;
    for (int i = 0; i < n; ++i) q[i] *= h0;
  }
  for (int k = 0; k < st->col; ++k) {
    int j = (st->head + k) % GPSAT_LBFGS_M;
    double b = st->rho[j] * lbfgs_dot(st->Y[j], q, n);
    for (int i = 0; i < n; ++i) q[i] += (a[k] - b) * st->S[j][i];
  }
  for (int i = 0; i < n; ++i) st->d[i] = -q[i];
}

GPSAT_HD double lbfgs_maxabs(const double* g, int n) {
  double m = 0.0;
  for (int i = 0; i < n; ++i) m = fmax(m, fabs(g[i]));
  return m;
}

// Consume (f, g) evaluated at st->xtrial.  Returns LBFGS_CONTINUE (evaluate the new
// st->xtrial) or a terminal code; on termination st->x / st->f / st->g hold the result.
GPSAT_HD int lbfgs_step(LbfgsState* st, double f, const double* g, const LbfgsOptions* opt) {
  const int n = st->np;
  const double epsmch = DBL_EPSILON;
  const double stpmx = 1e10;
  st->nfev++;
  bool new_direction = false;
  if (!st->started) {
    st->started = 1;
    if (!isfinite(f)) { st->f = f; return LBFGS_NOT_FINITE; }
    st->f = f;
    for (int i = 0; i < n; ++i) { st->x[i] = st->xtrial[i]; st->g[i] = g[i]; }
    if (lbfgs_maxabs(st->g, n) <= opt->pgtol) return LBFGS_CONV_PGTOL;
    new_direction = true;
  }
  // f_cur / g_cur: function and gradient at the point the line search looks at
  double f_cur = f;
  const double* g_cur = g;
  for (;;) {
    if (new_direction) {
      lbfgs_direction(st);
      double dtd = lbfgs_dot(st->d, st->d, n);
      st->dnorm = sqrt(dtd);
      st->stp = (st->iter == 0) ? fmin(1.0 / st->dnorm, stpmx) : 1.0;
      for (int i = 0; i < n; ++i) { st->t[i] = st->x[i]; st->r[i] = st->g[i]; }
      st->fold = st->f;
      st->ifun = 0; st->iback = 0; st->ls_started = 0;
      f_cur = st->f; g_cur = st->g;
      new_direction = false;
    }
    // ---- lnsrlb, label 556 ----
    int info = 0;
    st->gd = isfinite(f_cur) ? lbfgs_dot(g_cur, st->d, n) : NAN;
    int task = DCSRCH_FG;
    if (st->ifun == 0) {
      st->gdold = st->gd;
      if (st->gd >= 0.0) info = -4;   // ascent direction
    }
    if (info == 0) {
      task = lbfgs_dcsrch(st, f_cur, st->gd, st->stp, stpmx);
      if (task == DCSRCH_ERROR) info = -4;
    }
    if (info == 0 && task == DCSRCH_FG) {
      st->ifun++;
      st->iback = st->ifun - 1;
      if (st->iback < opt->maxls) {
        if (st->stp == 1.0) for (int i = 0; i < n; ++i) st->xtrial[i] = st->t[i] + st->d[i];
        else for (int i = 0; i < n; ++i) st->xtrial[i] = st->stp * st->d[i] + st->t[i];
        return LBFGS_CONTINUE;
      }
    }
    if (info != 0 || (task == DCSRCH_FG && st->iback >= opt->maxls)) {
      // line search failed: restore the previous iterate
      for (int i = 0; i < n; ++i) { st->x[i] = st->t[i]; st->g[i] = st->r[i]; }
      st->f = st->fold;
      if (st->col == 0) return LBFGS_ABNORMAL;
      st->col = 0; st->head = 0; st->theta = 1.0;   // refresh the memory and restart
      new_direction = true;
      continue;
    }
    // ---- line search finished (CONV or WARN): xtrial is the new iterate ----
    for (int i = 0; i < n; ++i) { st->x[i] = st->xtrial[i]; st->g[i] = g_cur[i]; }
    st->f = f_cur;
    st->iter++;
    // SciPy wrapper checks its limits when it sees NEW_X, before the convergence tests run
    if (st->iter >= opt->maxiter) return LBFGS_STOP_MAXITER;
    if (st->nfev > opt->maxfun) return LBFGS_STOP_MAXFUN;
    if (lbfgs_maxabs(st->g, n) <= opt->pgtol) return LBFGS_CONV_PGTOL;
    double ddum = fmax(fmax(fabs(st->fold), fabs(st->f)), 1.0);
    if ((st->fold - st->f) <= opt->ftol * ddum) return LBFGS_CONV_FTOL;
    // ---- BFGS pair ----
    for (int i = 0; i < n; ++i) st->r[i] = st->g[i] - st->r[i];
    double rr = lbfgs_dot(st->r, st->r, n);
    double dr, dd;
    if (st->stp == 1.0) { dr = st->gd - st->gdold; dd = -st->gdold; }
    else {
      dr = (st->gd - st->gdold) * st->stp;
      for (int i = 0; i < n; ++i) st->d[i] *= st->stp;
      dd = -st->gdold * st->stp;
    }
    if (dr > epsmch * dd) {
      int slot;
      if (st->col < GPSAT_LBFGS_M) { slot = (st->head + st->col) % GPSAT_LBFGS_M; st->col++; }
      else { slot = st->head; st->head = (st->head + 1) % GPSAT_LBFGS_M; }
      for (int i = 0; i < n; ++i) { st->S[slot][i] = st->d[i]; st->Y[slot][i] = st->r[i]; }
      st->rho[slot] = 1.0 / dr;
      st->theta = rr / dr;
    }
    new_direction = true;
  }
}
