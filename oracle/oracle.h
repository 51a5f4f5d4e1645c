// oracle.h — CPU restatement of the engine-sim hot path.  TEST INFRASTRUCTURE ONLY.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, link or call anything under oracle/.  The product
// (engine-sim-cli_b200/) never includes or links it.
//
// PARITY UNPINNED: the reference's implementation of this path lives in the absent
// submodule github.com/danieljsinclair/engine-sim @ 1b3ba11 and the
// reference holds no golden vector for it.  This is a restatement of
// the published upstream algorithm (ange-yaghi/engine-sim + simple-2d-constraint-solver)
// anchored on the reference's call sites and the code fragments quoted in its docs; every
// function cites what it follows and oracle/ASSUMPTIONS.md lists what had to be assumed.
#pragma once
#include "esb_params.h"

#include <cstdint>
#include <vector>

namespace oracle {

constexpr double kR = 8.31446261815324;        // J/(mol K)
constexpr double kAirMolarMass = 0.02897;      // kg/mol (upstream units::AirMolecularMass)
constexpr double kAtm = 101325.0;              // PresetSerializer.cpp:214
constexpr double kT25C = 298.15;               // PresetSerializer.cpp:215
constexpr double kPi = 3.14159265358979323846;

struct Mix { double p_fuel = 0, p_inert = 1, p_o2 = 0; };

// upstream gas_system.{h,cpp} (engine-sim-esp32/main/CMakeLists.txt:48)
struct GasSystem {
    double n_mol = 0, E_k = 0, V = 0;
    double mom[2] = {0, 0};
    Mix mix;
    double width = 0, height = 0, dx = 1, dy = 0;
    int dof = 5;
    double chokedFlowLimit = 0, chokedFlowFactor = 0;

    void initialize(double P, double V, double T, const Mix &m, int dof = 5);
    void reset(double P, double T, const Mix &m);
    void setGeometry(double w, double h, double dx_, double dy_) { width = w; height = h; dx = dx_; dy = dy_; }
    double hcr() const { return 1.0 + 2.0 / dof; }
    double pressure() const { return (V != 0) ? E_k / (0.5 * dof * V) : 0; }
    double temperature() const { return (n_mol != 0) ? E_k / (0.5 * dof * n_mol * kR) : 0; }
    double mass() const { return kAirMolarMass * n_mol; }
    double kineticEnergyPerMol() const { return (n_mol != 0) ? E_k / n_mol : 0; }
    double c() const;
    double bulkKineticEnergy() const;
    double dynamicPressure(double dirx, double diry) const;
    void changeVolume(double dV);
    void gainN(double dn, double E_k_per_mol, const Mix &m);
    void loseN(double dn, double E_k_per_mol);
    double pressureEquilibriumMaxFlow(const GasSystem &sink) const;
    void dissipateExcessVelocity();
    void updateVelocity(double dt, double beta);
    static double flowRate(double k, double P0, double P1, double T0, double T1, double hcr, double chokedLimit, double chokedFactor);
};

struct FlowParameters {
    GasSystem *system_0 = nullptr, *system_1 = nullptr;
    double k_flow = 0, dt = 0, direction_x = 1, direction_y = 0;
    double crossSectionArea_0 = 0, crossSectionArea_1 = 0;
};
double gasFlow(const FlowParameters &p); // GasSystem::flow(const FlowParameters&)

struct Body { double px = 0, py = 0, th = 0, vx = 0, vy = 0, w = 0, invM = 0, invI = 0; };

struct Row {
    int b0 = -1, b1 = -1;
    double J0[3] = {0, 0, 0}, J1[3] = {0, 0, 0};
    double rhs = 0, lo = 0, hi = 0, invA = 0;
};

struct FlameEvent {
    bool lit = false;
    double travel_x = 0, travel_y = 0, lastVolume = 0, flameSpeed = 0, efficiency = 1;
};

struct Controls {
    double throttle = 0;       // ISimulator::setThrottle -> Engine::setSpeedControl (BridgeSimulator.cpp:183-189)
    int ignition = 1;          // BridgeSimulator.cpp:191-195
    int starter = 0;           // m_starterMotor.m_enabled
    int dynoEnabled = 0, dynoHold = 0;
    double dynoSpeed = 0;      // rad/s (BridgeSimulator.cpp:328)
    double dynoMaxTorque = 0;  // N m
    double clutchPressure = 0, brake = 0;
    int gear = -1;             // engine-sim convention: -1 neutral, 0.. forward gears (GearConventions.h); Transmission::changeGear
    double dynoKs = 10.0, dynoKd = 1.0; // Dynamometer::m_ks / m_kd (RealEngineAudioSpike.cpp:163-164): carried, see ASSUMPTIONS.md #3a
};

struct SolverConfig {
    int maxIterations = 24;
    double minDelta = 1e-7;   // largest row velocity change in an iteration (m/s or rad/s)
    double beta = 0.2;        // position-error feedback per step
};

constexpr int kMaxDelay = 32;

// One engine instance: rigid bodies + gas cells + per-cylinder state.
struct Engine {
    EsbEngineParams P;
    int N = 0, nb = 0;
    double dt = 1e-4;
    int fluidSteps = 8;
    uint64_t seed = 0;
    SolverConfig solver;
    Controls ctl;

    std::vector<Body> bodies; // 0 crank, 1+2i piston i, 2+2i rod i, last vehicle mass
    std::vector<Row> rows;
    std::vector<double> lambda; // warm-start impulses, one per row
    int rowDyno = -1, rowStarter = -1, rowClutch = -1;

    std::vector<GasSystem> chamber, intakeRunner, exhaustRunner, plenum, collector;
    std::vector<FlameEvent> flame;
    std::vector<double> lastVolume;
    std::vector<uint32_t> igniteCount;
    std::vector<double> delayRing; // N * kMaxDelay
    std::vector<int> delayLen, delayPos;
    double lastCycleAngle = 0, revLimitTimer = 0, filteredSpeed = 0;
    double throttlePlate = 1.0; // Engine::getThrottle(): 1 = closed plate (upstream convention)
    double dynoTorque = 0, clutchTorque = 0;
    int lastIterations = 0;
    uint64_t stepIndex = 0;
    double exhaustOut[ESB_MAX_EXHAUST] = {0, 0, 0, 0};
    double totalExhaustFlow = 0;

    void initialize(const EsbEngineParams &p, double simFreq, int fluidSteps, uint64_t seed);
    void step();
    void changeGear(int newGear);   // Transmission::changeGear
    double rpm() const { return -bodies[0].w * (30.0 / kPi); }
    double cycleAngle() const;
    double chamberVolume(int i) const;
    double manifoldPressure() const { return plenum[0].pressure(); }

  private:
    void placeBodies();
    void buildRows();
    void solve();
    void ignitionUpdate();
    void chamberUpdate(int i);
    void chamberFlow(int i, double sdt);
    void writeAudio();
};

double sampleTriangle(const EsbEngineParams &P, int fn, double x); // Function::sampleTriangle

// ---------------------------------------------------------------------------------------
// Synthesizer (upstream synthesizer.cpp + filters; quotes at
// archive/PHYSICS_ANALYSIS.md:39-55,291-311,337-383,424-465)
struct AudioParameters { // defaults confirmed at SineSimulator.cpp:43-46
    float volume = 1.0f, convolution = 1.0f, dF_F_mix = 0.01f, inputSampleNoise = 0.5f, airNoise = 1.0f;
    float inputSampleNoiseFrequencyCutoff = 10000.0f, airNoiseFrequencyCutoff = 2000.0f;
    float levelerTarget = 30000.0f, levelerMaxGain = 1.9f, levelerMinGain = 1e-5f;
};

struct SynthChannel {
    double lastInputSample = 0;
    float antialiasY = 0, jitterNoiseY = 0, dcY = 0, derivPrev = 0, airY = 0;
    float jitterHistory[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    int jitterOffset = 0;
    std::vector<float> input;     // pending resampled input (ring buffer contents)
    std::vector<float> ir;        // h[k] = volume * s16[k] / 32767
    std::vector<float> shift;     // convolution shift register (circular)
    int shiftOffset = 0;
    uint64_t sampleCounter = 0;   // noise counter
};

struct Synth {
    int audioSampleRate = 44100, inputBufferSize = 44100, channels = 0;
    double inputSampleRate = 10000;
    AudioParameters ap;
    uint64_t seed = 0;
    int firMode = 1;              // 0 = FP32 serial tap order (reference order), 1 = FP64 accumulation
    double inputWriteOffset = 0, lastInputSampleOffset = 0;
    long writeIndex = 0;          // ring write index of the input buffers
    int inputSamplesRead = 0;
    std::vector<SynthChannel> ch;
    float masterY = 0, levelerPeak = 0;
    std::vector<int16_t> audio;   // audio ring contents (int16, as readAudioOutput returns them)
    std::vector<float> audioFloat;// same samples before rounding/clamping (for SNR)

    void initialize(int channels, double inputSampleRate, uint64_t seed);
    // Synthesizer::initializeImpulseResponse(s16*, n, volume, channel); maxTaps<=0 -> upstream cap 10000
    void initializeImpulseResponse(const int16_t *s, unsigned n, float volume, int channel, int maxTaps);
    void writeInput(const double *data);
    void endInputBlock();
    int renderAudio();            // renderAudioOnDemand(): returns samples rendered
    int readAudioOutput(int samples, int16_t *out, float *outFloat);
};

// One simulator = Engine + Synth with BridgeSimulator's stepping arithmetic
// (BridgeSimulator.cpp:61-98,525-548).
struct Simulator {
    Engine engine;
    Synth synth;
    double simFreq = 10000;
    int sampleRate = 44100;
    bool sine = false;            // SineSimulator path (SineSimulator.cpp:84-105)
    double sinePhase = 0;
    void stepOnce();              // simulateStep + writeToSynthesizer
    void endFrame() { synth.endInputBlock(); }
    void advance(int steps);      // startFrame; steps x simulateStep; endFrame
};

} // namespace oracle
