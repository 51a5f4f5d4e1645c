// dev_main.cpp — developer driver for the oracle physics (not part of any test).
#include "oracle.h"
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
int main(int argc, char **argv) {
    const char *path = argc > 1 ? argv[1] : "build/gm_ls.bin";
    double targetRpm = argc > 2 ? atof(argv[2]) : 2000;
    double seconds = argc > 3 ? atof(argv[3]) : 3;
    double thr = argc > 4 ? atof(argv[4]) : 1.0;
    static EsbEngineParams P;
    FILE *f = fopen(path, "rb"); if (!f || fread(&P, sizeof(P), 1, f) != 1) { fprintf(stderr, "cannot read %s\n", path); return 1; }
    fclose(f);
    oracle::Engine e;
    e.initialize(P, 10000, 8, 0);
    if (argc > 5) e.solver.maxIterations = atoi(argv[5]);
    if (argc > 6) e.solver.minDelta = atof(argv[6]);
    e.ctl.dynoEnabled = 1; e.ctl.dynoHold = 1; e.ctl.dynoSpeed = targetRpm * oracle::kPi / 30; e.ctl.dynoMaxTorque = 5000;
    e.ctl.starter = 1; e.ctl.ignition = 1;
    int steps = (int)(seconds * 10000);
    long itSum = 0; double maxC = 0;
    auto t0 = std::chrono::steady_clock::now();
    for (int s = 0; s < steps; ++s) {
        e.ctl.throttle = thr * std::fmin(1.0, s * 1e-4 / 0.5);
        e.step();
        itSum += e.lastIterations;
        if (s % 2000 == 1999) {
            printf("t=%.2f rpm=%8.2f dynoT=%8.2f it=%5.1f Pman=%7.0f Pch0=%9.0f Tch0=%6.0f exh0=%9.3f fuel0=%.4f pistonAx=%.5f\n", (s + 1) * 1e-4, e.rpm(), e.dynoTorque, itSum / 2000.0,
                   e.manifoldPressure(), e.chamber[0].pressure(), e.chamber[0].temperature(), e.exhaustOut[0], e.chamber[0].mix.p_fuel, e.chamberVolume(0));
            itSum = 0;
        }
    }
    double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("%.3f s wall for %.1f engine-s => %.3f x realtime\n", el, seconds, seconds / el);
    return 0;
}
