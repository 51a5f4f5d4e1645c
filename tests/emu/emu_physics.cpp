// emu_physics.cpp — host emulation of the physics kernel's lane choreography (TEST ONLY).
//
// Compiles engine-sim-cli_b200/csrc/physics_run.cuh for the CPU with one std::thread per lane
// and a barrier-based stand-in for warp shuffles, so the CPU test-suite (no GPU) can check that
// the lane-parallel formulation reproduces the sequential oracle bit for bit.  Never linked
// into the product library; the product has no CPU path.
#define ESB_EMULATE 1
#include <barrier>
#include <cstring>
#include <thread>
#include <vector>

#include "physics_run.cuh"
#include "host.h"

namespace esbk {
struct EmuShared {
    std::barrier<> bar;
    double slots[32];
    explicit EmuShared(int g) : bar(g) {}
};
void emuBarrier(EmuShared *s) { s->bar.arrive_and_wait(); }
double *emuSlots(EmuShared *s) { return s->slots; }
} // namespace esbk

template <int G> static void runGroup(const esbk::RunArgs &a, int inst) {
    bool slaves = false;
    for (int i = 0; i < a.D->P.cylinderCount; ++i) slaves = slaves || a.D->P.cyl[i].masterCylinder >= 0;
    esbk::EmuShared sh(G);
    // the block's "shared memory": same layout as on the device, one lane group
    const EsbGasNet &gn = a.D->gas;
    const esbk::GasSmemLayout L = esbk::gasSmemLayout(gn, a.D->P.cylinderCount, a.D->P.intakeCount, a.D->fluidSteps, 1);
    std::vector<double> mem(L.total / sizeof(double) + 2, 0.0);
    unsigned char *base = reinterpret_cast<unsigned char *>(mem.data());
    esbk::gasTablesToShared(base, L, a.D, 0, 1);
    const esbk::GasView view = esbk::gasViewAt(base, L, gn, 0);
    std::vector<std::thread> th;
    for (int lane = 0; lane < G; ++lane)
        th.emplace_back([&, lane] {
            esbk::Group<G> g; g.lane = lane; g.mask = 0; g.sh = &sh;
            if (slaves) esbk::runInstance<G, true>(a, inst, true, g, view);
            else esbk::runInstance<G, false>(a, inst, true, g, view);
        });
    for (auto &t : th) t.join();
}

extern "C" {

// Runs nSteps for nInst instances of one engine type starting from `state` ([fields][nInst], in/out).
// ctl: [ESB_CTL_COUNT][nInst]. keys optional. stream out: [nSteps][channels][nInst].
int emu_physics_run(const EsbEngineParams *P, double simFreq, int fluidSteps, int maxIt, double minDelta, double beta, uint64_t seedBase,
                    int nInst, double *state, const double *ctl, int nSteps, long long stepBase, int nKeys, const long long *keyStep,
                    const double *keyThrottle, const double *keyDyno, double *stream, double *telemetry, int telemetryEvery, int lanesOverride) {
    static EsbDeviceParams D;
    esb::deriveDeviceParams(*P, simFreq, fluidSteps, maxIt, minDelta, beta, seedBase, D);
    esbk::RunArgs a;
    a.D = &D; a.S = state; a.ctl = ctl; a.nInst = nInst; a.nSteps = nSteps; a.stepBase = stepBase;
    a.nKeys = nKeys; a.keyStep = keyStep; a.keyThrottle = keyThrottle; a.keyDynoSpeed = keyDyno;
    a.stream = stream; a.channels = P->exhaustSystemCount; a.telemetry = telemetry; a.telemetryEvery = telemetryEvery; a.telemetryBase = stepBase / (telemetryEvery > 0 ? telemetryEvery : 1);
    const int G = lanesOverride > 0 ? lanesOverride : D.lanesPerInstance;
    if (G != D.lanesPerInstance) { D.lanesPerInstance = G; esb::buildGasNet(*P, D); }   // the slot schedule is built for the lane count
    for (int inst = 0; inst < nInst; ++inst) {
        switch (G) {
        case 1: runGroup<1>(a, inst); break;
        case 2: runGroup<2>(a, inst); break;
        case 4: runGroup<4>(a, inst); break;
        case 8: runGroup<8>(a, inst); break;
        case 16: runGroup<16>(a, inst); break;
        default: return -1;
        }
    }
    return 0;
}

int emu_state_fields() { return ESB_STATE_FIELDS; }

// initial state replicated for nInst instances
void emu_initial_state(const EsbEngineParams *P, double simFreq, int nInst, double *state) {
    static EsbDeviceParams D;
    esb::deriveDeviceParams(*P, simFreq, 8, 24, 1e-7, 0.2, 0, D);
    std::vector<double> col;
    esb::initialStateColumn(D, col);
    for (int f = 0; f < ESB_STATE_FIELDS; ++f)
        for (int i = 0; i < nInst; ++i) state[(size_t)f * nInst + i] = col[f];
}

} // extern "C"
