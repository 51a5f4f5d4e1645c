// capi.cpp — C interface of the oracle for ctypes (TEST INFRASTRUCTURE ONLY).
// Called from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
#include "oracle.h"
#include "esb_math.h"

#include <chrono>
#include <cstring>

using namespace oracle;

extern "C" {

void *orc_sim_create(const EsbEngineParams *p, double simFreq, int fluidSteps, uint64_t seed, int sineMode) {
    Simulator *s = new Simulator();
    s->simFreq = simFreq;
    s->sine = sineMode != 0;
    s->engine.initialize(*p, simFreq, fluidSteps, seed);
    s->synth.initialize(p->exhaustSystemCount, simFreq, seed);
    // applyLoadedEngineSettings (SimulatorFactory.cpp:63-71)
    s->synth.ap.inputSampleNoise = (float)p->initialJitter;
    s->synth.ap.airNoise = (float)p->initialNoise;
    s->synth.ap.dF_F_mix = (float)p->initialHighFrequencyGain;
    return s;
}
void orc_sim_destroy(void *h) { delete (Simulator *)h; }

void orc_sim_set_ir(void *h, int channel, const int16_t *samples, unsigned n, float volume, int maxTaps) {
    ((Simulator *)h)->synth.initializeImpulseResponse(samples, n, volume, channel, maxTaps);
}
int orc_sim_ir_taps(void *h, int channel) { return (int)((Simulator *)h)->synth.ch[channel].ir.size(); }
void orc_sim_set_audio_params(void *h, float convolution, float dF_F_mix, float inputSampleNoise, float airNoise, float volume) {
    AudioParameters &a = ((Simulator *)h)->synth.ap;
    a.convolution = convolution; a.dF_F_mix = dF_F_mix; a.inputSampleNoise = inputSampleNoise; a.airNoise = airNoise; a.volume = volume;
}
// output sample rate of the synthesizer (before the first step; the reference's tests also run at 48 kHz)
void orc_sim_set_sample_rate(void *h, int rate) { ((Simulator *)h)->synth.audioSampleRate = rate; ((Simulator *)h)->sampleRate = rate; }
void orc_sim_set_fir_mode(void *h, int mode) { ((Simulator *)h)->synth.firMode = mode; }
void orc_sim_set_solver(void *h, int maxIterations, double minDelta, double beta) {
    SolverConfig &c = ((Simulator *)h)->engine.solver;
    c.maxIterations = maxIterations; c.minDelta = minDelta; c.beta = beta;
}
void orc_sim_set_controls(void *h, double throttle, int ignition, int starter, int dynoEnabled, int dynoHold, double dynoSpeed, double dynoMaxTorque) {
    Controls &c = ((Simulator *)h)->engine.ctl;
    c.throttle = throttle; c.ignition = ignition; c.starter = starter; c.dynoEnabled = dynoEnabled; c.dynoHold = dynoHold;
    c.dynoSpeed = dynoSpeed; c.dynoMaxTorque = dynoMaxTorque;
}

void orc_sim_set_drivetrain(void *h, double clutchPressure, double brake) {
    Controls &c = ((Simulator *)h)->engine.ctl;
    c.clutchPressure = clutchPressure; c.brake = brake;
}
void orc_sim_change_gear(void *h, int gear) { ((Simulator *)h)->engine.changeGear(gear); }
double orc_sim_vehicle_inv_inertia(void *h) { const Engine &e = ((Simulator *)h)->engine; return e.bodies[e.nb - 1].invI; }

#define ORC_TELEMETRY_STRIDE 8
static void telemetry(const Simulator &s, double *t) {
    const Engine &e = s.engine;
    t[0] = e.rpm(); t[1] = e.dynoTorque; t[2] = e.manifoldPressure(); t[3] = e.totalExhaustFlow;
    t[4] = e.chamber[0].pressure(); t[5] = e.cycleAngle(); t[6] = e.throttlePlate; t[7] = (double)e.lastIterations;
}
void orc_sim_get_telemetry(void *h, double *out) { telemetry(*(Simulator *)h, out); }

// Full physics state, for bit-exact comparison with the device state dump.
// Layout: per body {px,py,th,vx,vy,w}; then per cylinder chamber, intake runner, exhaust runner
// {n,E_k,V,mom0,mom1,p_fuel,p_inert,p_o2}; then plenums, collectors; returns doubles written.
int orc_sim_get_state(void *h, double *out, int maxDoubles) {
    const Engine &e = ((Simulator *)h)->engine;
    int k = 0;
    auto put = [&](double v) { if (k < maxDoubles) out[k] = v; ++k; };
    for (const Body &b : e.bodies) { put(b.px); put(b.py); put(b.th); put(b.vx); put(b.vy); put(b.w); }
    auto gas = [&](const GasSystem &g) { put(g.n_mol); put(g.E_k); put(g.V); put(g.mom[0]); put(g.mom[1]); put(g.mix.p_fuel); put(g.mix.p_inert); put(g.mix.p_o2); };
    for (int i = 0; i < e.N; ++i) { gas(e.chamber[i]); gas(e.intakeRunner[i]); gas(e.exhaustRunner[i]); }
    for (const GasSystem &g : e.plenum) gas(g);
    for (const GasSystem &g : e.collector) gas(g);
    return k;
}

// Runs nSteps in batches of stepsPerBatch: {steps; endFrame; renderAudioOnDemand; readAudioOutput(readCap)}
// — the schedule of the reference's offline spike (RealEngineAudioSpike.cpp:197-219).
// throttle / dynoSpeed: per-step arrays or NULL (keep current controls).
// exhaustStream (optional): nSteps * channels doubles, the synthesizer input of every step.
// Returns the number of PCM samples produced.
int orc_sim_run(void *h, int nSteps, const double *throttle, const double *dynoSpeed, int stepsPerBatch, int readCap,
                int16_t *pcm, float *pcmFloat, int maxSamples, double *telem, int telemetryEvery, double *exhaustStream) {
    Simulator &s = *(Simulator *)h;
    int produced = 0, done = 0, tcount = 0;
    const int C = s.synth.channels;
    std::vector<int16_t> tmp(readCap);
    std::vector<float> tmpf(readCap);
    while (done < nSteps) {
        const int n = (nSteps - done < stepsPerBatch) ? nSteps - done : stepsPerBatch;
        for (int k = 0; k < n; ++k, ++done) {
            if (throttle) s.engine.ctl.throttle = throttle[done];
            if (dynoSpeed) s.engine.ctl.dynoSpeed = dynoSpeed[done];
            s.stepOnce();
            if (exhaustStream && !s.sine) for (int c = 0; c < C; ++c) exhaustStream[(size_t)done * C + c] = s.engine.exhaustOut[c];
            if (telem && telemetryEvery > 0 && (done + 1) % telemetryEvery == 0) { telemetry(s, telem + (size_t)tcount * ORC_TELEMETRY_STRIDE); ++tcount; }
        }
        s.endFrame();
        s.synth.renderAudio();
        const int got = s.synth.readAudioOutput(readCap, tmp.data(), tmpf.data());
        const int keep = (produced + got <= maxSamples) ? got : maxSamples - produced;
        if (keep > 0) {
            if (pcm) std::memcpy(pcm + produced, tmp.data(), sizeof(int16_t) * keep);
            if (pcmFloat) std::memcpy(pcmFloat + produced, tmpf.data(), sizeof(float) * keep);
            produced += keep;
        }
    }
    return produced;
}

// Stand-alone synthesizer run on a given input stream (channels x nSteps doubles, step-major).
int orc_synth_run(void *h, int nSteps, const double *stream, int stepsPerBatch, int readCap, int16_t *pcm, float *pcmFloat, int maxSamples) {
    Simulator &s = *(Simulator *)h;
    const int C = s.synth.channels;
    int produced = 0, done = 0;
    std::vector<int16_t> tmp(readCap);
    std::vector<float> tmpf(readCap);
    while (done < nSteps) {
        const int n = (nSteps - done < stepsPerBatch) ? nSteps - done : stepsPerBatch;
        for (int k = 0; k < n; ++k, ++done) s.synth.writeInput(stream + (size_t)done * C);
        s.synth.endInputBlock();
        s.synth.renderAudio();
        const int got = s.synth.readAudioOutput(readCap, tmp.data(), tmpf.data());
        const int keep = (produced + got <= maxSamples) ? got : maxSamples - produced;
        if (keep > 0) {
            if (pcm) std::memcpy(pcm + produced, tmp.data(), sizeof(int16_t) * keep);
            if (pcmFloat) std::memcpy(pcmFloat + produced, tmpf.data(), sizeof(float) * keep);
            produced += keep;
        }
    }
    return produced;
}

// FP64 direct convolution y[c][n] = sum_k h[k] x[c][n-k] (zero history) — config 5 oracle.
void orc_fir_f64(const float *x, int channels, int n, const float *h, int L, int hStride, double *y) {
    for (int c = 0; c < channels; ++c) {
        const float *xc = x + (size_t)c * n;
        const float *hc = h + (size_t)c * hStride;
        for (int i = 0; i < n; ++i) {
            double acc = 0;
            const int kmax = (i < L - 1) ? i : L - 1;
            for (int k = 0; k <= kmax; ++k) acc += (double)hc[k] * (double)xc[i - k];
            y[(size_t)c * n + i] = acc;
        }
    }
}

// deterministic math of include/esb_math.h, for tests/test_math.py: 0 sin, 1 cos, 2 log, 3 exp, 4 pow
double orc_math(int fn, double x, double y) {
    switch (fn) { case 0: return esb_sin(x); case 1: return esb_cos(x); case 2: return esb_log(x); case 3: return esb_exp(x); default: return esb_pow(x, y); }
}
double orc_rand_u01(uint64_t seed, uint64_t stream, uint64_t counter) { return esb_rand_u01(seed, stream, counter); }

double orc_sample_function(const EsbEngineParams *p, int fn, double x) { return sampleTriangle(*p, fn, x); }
int orc_params_size() { return (int)sizeof(EsbEngineParams); }

} // extern "C"
