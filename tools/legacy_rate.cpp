// Rate of the reference's backend interface as implemented by libbrp_b200.so, called from C the way demod_binary.c calls it
// (set_up_* once, then run_resampling / run_fft / run_harmonic_summing per template; demod_binary.c:1123-1141,1245-1292),
// without the driver around it: the standalone driver writes its checkpoint file after every template, which would be what is
// measured otherwise.  Synthetic noise series; the thresholds are the base thresholds of -A 0.08.
//
//   g++ -O2 -o tools/bin/legacy_rate tools/legacy_rate.cpp -Lboinc-app-eah-brp_b200 -lbrp_b200 -Wl,-rpath,'$ORIGIN/../../boinc-app-eah-brp_b200'
//   tools/bin/legacy_rate [n_templates] [padding]
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include <vector>

#include "../include/brp_b200.h"

static double now_s()
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int main(int argc, char **argv)
{
    const int n_t = argc > 1 ? atoi(argv[1]) : 1000;
    const float padding = argc > 2 ? (float)atof(argv[2]) : 1.0f;
    const uint32_t nu = 1u << 22;
    brp_geometry g;
    if (brp_derive_geometry(nu, 65.47619047619051, padding, 400.0f, 1000, &g)) return 1;
    std::vector<float> series(nu);
    unsigned long long x = 88172645463325252ull;
    for (uint32_t i = 0; i < nu; i++) { // sum of four uniforms: close enough to Gaussian noise for a rate measurement
        float v = 0.f;
        for (int k = 0; k < 4; k++) {
            x ^= x << 13; x ^= x >> 7; x ^= x << 17;
            v += (float)(x >> 40) * (1.0f / 16777216.0f) - 0.5f;
        }
        series[i] = v * 1.7320508f;
    }
    int dev = 0;
    if (initialize_cuda(1, &dev)) return 2;
    DIfloatPtr in, resamp, ps;
    in.host_ptr = series.data();
    RESAMP_PARAMS p;
    p.nsamples = g.nsamples; p.nsamples_unpadded = nu; p.fft_size = g.fft_size; p.dt = g.dt; p.step_inv = g.step_inv;
    float *sumspec[5];
    int32_t *dirty[5];
    unsigned int nr_pages = 0;
    if (set_up_resampling(in, &resamp, &p, NULL, NULL) || set_up_fft(resamp, &ps, g.nsamples, g.fft_size) ||
        set_up_harmonic_summing(sumspec, dirty, &nr_pages, g.fundamental_idx_hi, g.harmonic_idx_hi))
        return 3;
    float thr[5];
    brp_base_thresholds(0.08f, g.fft_size, thr);
    double t0 = 0;
    unsigned long hits = 0;
    for (int it = -20; it < n_t; it++) {
        if (it == 0) t0 = now_s();
        brp_template t;
        brp_derive_template(660.0 + 1.37 * (it + 20), 0.05 + 0.0002 * (it + 20), 0.001 * (it + 20), g.step_inv, &t);
        p.tau = t.tau; p.Omega = t.Omega; p.Psi0 = t.Psi0; p.S0 = t.S0;
        if (run_resampling(in, resamp, &p) || run_fft(resamp, ps, g.nsamples, g.fft_size, 1.0f / (float)g.nsamples) ||
            run_harmonic_summing(sumspec, dirty, nr_pages, ps, g.window_2, g.fundamental_idx_hi, g.harmonic_idx_hi, thr))
            return 4;
        for (unsigned int pg = 0; pg < nr_pages; pg++) hits += dirty[4][pg];
    }
    const double dt = now_s() - t0;
    printf("legacy backend functions from C, 2^22-sample series, -P %g: %d templates in %.3f s = %.0f templates/s (%.1f us per template; %lu dirty pages seen)\n",
           padding, n_t, dt, n_t / dt, 1e6 * dt / n_t, hits);
    tear_down_harmonic_summing(sumspec, dirty);
    tear_down_fft(ps);
    tear_down_resampling(resamp);
    shutdown_cuda();
    return 0;
}
