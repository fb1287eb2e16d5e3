// Test-only host build of the device L-BFGS state machine (gpsat_b200/csrc/lbfgs.cuh) so its
// logic can be compared with scipy.optimize.minimize(method="L-BFGS-B") on the CPU.
// Not part of the product path: the engine runs the same header on the device.
#include "../../gpsat_b200/csrc/lbfgs.cuh"
#include <cstring>

extern "C" {
typedef void (*eval_cb)(const double* x, double* f, double* g, void* user);

int lbfgs_host_minimize(int np, double* x, eval_cb cb, void* user, double pgtol, double ftol,
                        int maxiter, int maxfun, int maxls, double* f_out, int* nfev, int* nit,
                        double* trace, int trace_cap) {
  LbfgsState st;
  LbfgsOptions opt{pgtol, ftol, maxiter, maxfun, maxls};
  lbfgs_init(&st, np, x);
  int rc = LBFGS_CONTINUE;
  int k = 0;
  while (rc == LBFGS_CONTINUE) {
    double f, g[GPSAT_PMAX];
    cb(st.xtrial, &f, g, user);
    if (trace && k < trace_cap) { std::memcpy(trace + (size_t)k * np, st.xtrial, sizeof(double) * np); }
    ++k;
    rc = lbfgs_step(&st, f, g, &opt);
  }
  std::memcpy(x, st.x, sizeof(double) * np);
  *f_out = st.f; *nfev = st.nfev; *nit = st.iter;
  return rc;
}
int lbfgs_state_bytes() { return (int)sizeof(LbfgsState); }
}
