/*
 * abi_demo.c -- the C ABI from plain C: no Python, no torch, no C++.
 *
 * A resting layer of uniform depth (rotating sphere, no topography) must stay exactly at rest, and its mass
 * must be 4 pi a^2 h up to the O(1/N^2) quadrature error of the cell areas.  The program builds that
 * state on the host, hands it to cs_step_host (upload, halo fill, steps, diagnostics, download) and
 * checks both facts.  It is the smallest complete caller of include/cubesphere.h:
 *
 *   gcc -std=c99 -Iinclude examples/abi_demo.c -L<pkg>/lib -lcubesphere_b200 -Wl,-rpath,<pkg>/lib -lm -o abi_demo
 *
 * Exit code 0: checks passed; 2: no usable CUDA device (the library has no CPU path); 1: anything else.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "cubesphere.h"

int main(int argc, char** argv) {
    const int N = argc > 1 ? atoi(argv[1]) : 48, g = 2, steps = argc > 2 ? atoi(argv[2]) : 10;
    const int m = N + 2 * g;
    const size_t cells = (size_t)6 * m * m;
    const double depth = 1000.0, a = 6.37122e6, pi = 3.14159265358979323846;

    cs_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.abi_version = CS_ABI_VERSION;
    cfg.N = N, cfg.tiles_per_edge = 1, cfg.halo = g;
    cfg.dtype = CS_F64, cfg.rk = CS_RK_SSPRK3, cfg.limiter = CS_LIM_MC;

    cs_ctx* ctx = NULL;
    int rc = cs_ctx_create(&cfg, &ctx);
    if (rc == CS_E_CUDA) {
        fprintf(stderr, "abi_demo: %s\n", cs_last_error(NULL));
        return 2;
    }
    if (rc != CS_OK) {
        fprintf(stderr, "abi_demo: cs_ctx_create failed (%d): %s\n", rc, cs_last_error(NULL));
        return 1;
    }

    double* h = malloc(cells * sizeof(double));
    double* u = calloc(cells, sizeof(double));
    double* v = calloc(cells, sizeof(double));
    if (!h || !u || !v) return 1;
    for (size_t k = 0; k < cells; ++k) h[k] = depth;      /* ghosts included: the fill overwrites them */

    double diag[2] = {0.0, 0.0};
    rc = cs_step_host(ctx, h, u, v, /*dt=*/60.0, steps, diag, /*copy_back=*/1);
    if (rc != CS_OK) {
        fprintf(stderr, "abi_demo: cs_step_host failed (%d): %s\n", rc, cs_last_error(ctx));
        return 1;
    }

    double dev = 0.0, speed = 0.0;
    for (int f = 0; f < 6; ++f)
        for (int j = g; j < g + N; ++j)
            for (int i = g; i < g + N; ++i) {
                const size_t k = ((size_t)f * m + j) * m + i;
                dev = fmax(dev, fabs(h[k] - depth));
                speed = fmax(speed, fmax(fabs(u[k]), fabs(v[k])));
            }
    const double mass_exact = 4.0 * pi * a * a * depth;
    const double mass_err = fabs(diag[0] - mass_exact) / mass_exact;
    printf("C%d, %d steps: max |h - h0| = %.3e m, max |u| = %.3e rad/s, mass = %.12e (4 pi a^2 h: rel. diff %.2e), "
           "energy = %.6e, kernel launches = %lld\n",
           N, steps, dev, speed, diag[0], mass_err, diag[1], (long long)cs_launch_count(ctx));
    const int ok = dev <= 1e-9 && speed <= 1e-18 && mass_err <= 2.0 / ((double)N * N);
    cs_ctx_destroy(ctx);
    free(h), free(u), free(v);
    return ok ? 0 : 1;
}
