// main.cpp — drives the oracle physics on a few instances of the bench grid and counts executed
// basic blocks during Engine::step() only (TEST INFRASTRUCTURE; see count_flops.py).
#include "oracle.h"
#include <cstdio>
#include <cstdlib>
extern "C" int flopcount_enabled;
extern "C" void flopcount_dump(const char *path);
int main(int argc, char **argv) {
    const char *blobPath = argv[1];
    const int warm = atoi(argv[2]), measure = atoi(argv[3]), nInst = atoi(argv[4]);
    static EsbEngineParams P;
    FILE *f = fopen(blobPath, "rb");
    if (!f || fread(&P, sizeof(P), 1, f) != 1) return 1;
    fclose(f);
    long long steps = 0;
    for (int k = 0; k < nInst; ++k) {
        const int i = (int)((long long)k * 4095 / (nInst > 1 ? nInst - 1 : 1)); // spread over the 64 x 64 grid
        const double rpm = 800.0 + 6200.0 * (i % 64) / 63.0, thr = 0.1 + 0.9 * (i / 64) / 63.0;
        oracle::Engine e;
        e.initialize(P, 10000, 8, (uint64_t)i);
        e.ctl.dynoEnabled = 1; e.ctl.dynoHold = 1; e.ctl.dynoSpeed = rpm * oracle::kPi / 30; e.ctl.dynoMaxTorque = 5000;
        e.ctl.starter = 1; e.ctl.ignition = 1; e.ctl.throttle = thr;
        for (int s = 0; s < warm; ++s) e.step();
        flopcount_enabled = 1;
        for (int s = 0; s < measure; ++s) e.step();
        flopcount_enabled = 0;
        steps += measure;
    }
    flopcount_dump(argv[5]);
    printf("%lld\n", steps);
    return 0;
}
