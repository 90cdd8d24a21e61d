/* c1_lorentz.c — BASELINE config C1 from plain C through the C ABI alone (include/tci_b200.h):
 *   f(v) = 1 / (1 + sum v_i^2), localdims = fill(10, 8), tolerance = 1e-8      (reference README.md:21-29)
 * crossinterpolate2 (src/tensorci2.jl:1313-1323) = TensorCI2 constructor + optimize!, then sum(tci) and one evaluation.
 *
 *   gcc -std=c99 -Iinclude examples/c1_lorentz.c -Ltensorcrossinterpolation.jl_b200 -ltci_b200 \
 *       -Wl,-rpath,$PWD/tensorcrossinterpolation.jl_b200 -o c1_lorentz
 *
 * There is no CPU fallback: without a CUDA device tci_create fails and the program says so and exits with status 2. */
#include <stdio.h>
#include <stdlib.h>

#include "tci_b200.h"

#define N 8

static int die(const char* what, const char* msg) {
    fprintf(stderr, "%s: %s\n", what, msg ? msg : "(no message)");
    return 2;
}

int main(int argc, char** argv) {
    tci_handle h = NULL;
    if (tci_create(&h, 0) != TCI_OK) return die("tci_create", tci_last_error(NULL));

    int32_t localdims[N], pivot[N];
    for (int b = 0; b < N; ++b) { localdims[b] = 10; pivot[b] = 1; }
    const double coeff = 1.0;
    tci_functor f = NULL;
    if (tci_functor_builtin(h, TCI_F_LORENTZ_ID, &coeff, 1, &f) != TCI_OK) return die("tci_functor_builtin", tci_last_error(h));

    tci_tci2 t = NULL;
    if (tci_tci2_create(h, localdims, N, &t) != TCI_OK) return die("tci_tci2_create", tci_last_error(h));
    /* argv[1] = "percall": one C-ABI call per bond phase; default: half-sweeps as single launches (csrc/sweep.cu) */
    if (tci_tci2_set_engine(t, (argc > 1 && argv[1][0] == 'p') ? 0 : 1) != TCI_OK) return die("tci_tci2_set_engine", tci_tci2_last_error(t));
    if (tci_tci2_addglobalpivots(t, pivot, 1) != TCI_OK || tci_tci2_init_maxsample(t, f, pivot, 1) != TCI_OK)
        return die("initial pivot", tci_tci2_last_error(t));

    int64_t ranks[20];
    double errors[20];
    int niter = 0;
    /* tolerance 1e-8, unbounded bond dimension, maxiter 20, :backandforth, normalised errors, ncheckhistory 3,
       no global pivot search (the parity configuration of SURVEY.md 8d), not strictly nested */
    if (tci_tci2_optimize(t, f, 1e-8, 0, 20, 1, 1, 3, 0, 5, 10.0, 0, 1, ranks, errors, &niter) != TCI_OK)
        return die("tci_tci2_optimize", tci_tci2_last_error(t));

    double total = 0.0, value = 0.0;
    int32_t point[N] = {1, 2, 3, 4, 5, 6, 7, 8};
    if (tci_tci2_sum(t, &total) != TCI_OK || tci_tci2_evaluate(t, point, 1, &value) != TCI_OK)
        return die("sum / evaluate", tci_tci2_last_error(t));

    printf("iterations %d, rank %lld, error %.3e\n", niter, (long long)ranks[niter - 1], errors[niter - 1]);
    printf("sum over 10^8 grid points = %.12e\n", total);
    printf("tt(1..8) = %.12e   f(1..8) = %.12e\n", value, 1.0 / (1.0 + 204.0));

    tci_tci2_destroy(t);
    tci_functor_destroy(h, f);
    tci_destroy(h);
    return 0;
}
