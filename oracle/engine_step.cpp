// engine_step.cpp — oracle: Engine::step() and the per-step engine parts
// (TEST INFRASTRUCTURE ONLY; see oracle.h).
#include "oracle.h"
#include "esb_math.h"

#include <cmath>

namespace oracle {

static double positiveMod(double x, double m) {
    double r = std::fmod(x, m);
    return (r < 0) ? r + m : r;
}

// How complete a burn can get, 0..1 -> [lowEfficiencyAttenuation, 1].  The three parameters are required by the
// reference's fuel schema and handed to Fuel::initialize (src/preset/FuelDeserializer.cpp:26-39; defaults 0.6 / 2.0 / 10.0 in
// es/objects/objects.mr:59-63); the body that consumes them is in the absent submodule.  Restated from their names,
// ranges and defaults (ASSUMPTIONS.md #7): turbulence (the mean-piston-speed function's output) saturates its benefit at
// maxTurbulenceEffect; dilution = inert gas in the charge relative to the inert gas its oxygen arrived with as air
// (1 = no residuals), harmful up to maxDilutionEffect.
static double burnQuality(double turbulence, double p_inert, double p_o2, double lowAttenuation, double maxTurbulence, double maxDilution) {
    double t = (maxTurbulence > 0) ? turbulence / maxTurbulence : 1.0;
    if (t < 0) t = 0;
    if (t > 1) t = 1;
    const double inertFresh = (0.79 / 0.21) * p_o2;
    double d = 0;
    if (maxDilution > 1.0) {
        const double dilution = (inertFresh > 0) ? p_inert / inertFresh : maxDilution;
        d = (dilution - 1.0) / (maxDilution - 1.0);
        if (d < 0) d = 0;
        if (d > 1) d = 1;
    }
    const double quality = t * (1.0 - d);
    return lowAttenuation + (1.0 - lowAttenuation) * quality;
}

// IgnitionModule::update: spark when the cycle angle passes
// firingAngle - timingCurve(|omega|); suppressed while the rev limiter timer runs or
// ignition is off (BridgeSimulator.cpp:191-195).
void Engine::ignitionUpdate() {
    const double fourPi = 4.0 * kPi;
    const double now = cycleAngle();
    const double speed = -bodies[0].w;
    if (ctl.ignition && revLimitTimer == 0 && speed > 0) {
        const double advance = sampleTriangle(P, P.fnTimingCurve, speed);
        const double travelled = positiveMod(now - lastCycleAngle, fourPi);
        if (travelled < 2.0 * kPi) {
            for (int i = 0; i < N; ++i) {
                const double target = positiveMod(P.cyl[i].firingAngle - advance, fourPi);
                const double off = positiveMod(target - lastCycleAngle, fourPi);
                if (off > 0 && off <= travelled) {
                    // CombustionChamber::ignite (docs/FIX_PRESET_ENGINE_TESTS.md:71-77)
                    GasSystem &g = chamber[i];
                    FlameEvent &f = flame[i];
                    if (f.lit || g.mix.p_fuel == 0) continue;
                    const double afr = g.mix.p_o2 / g.mix.p_fuel;
                    const double equivalence = P.fuelMolecularAfr / afr;
                    if (equivalence < 0.5 || equivalence > 1.9) continue;
                    const double u = esb_rand_u01(seed, ESB_STREAM_BURN + i, igniteCount[i]);
                    igniteCount[i] += 1;
                    const double rnd = P.burningEfficiencyRandomness;
                    const double meanPistonSpeed = 2.0 * (2.0 * P.crankThrow) * (speed / (2.0 * kPi));
                    const double turbulence = sampleTriangle(P, P.cyl[i].fnMeanPistonSpeedToTurbulence, meanPistonSpeed);
                    // Fuel parameters the reference requires and forwards (FuelDeserializer.cpp:26-39): a burn in a slow,
                    // residual-laden charge is less complete (ASSUMPTIONS.md #7)
                    f.efficiency = P.maxBurningEfficiency * (rnd * u + (1.0 - rnd)) *
                                   burnQuality(turbulence, g.mix.p_inert, g.mix.p_o2, P.lowEfficiencyAttenuation, P.maxTurbulenceEffect, P.maxDilutionEffect);
                    const double ratio = sampleTriangle(P, P.fnTurbulenceToFlameSpeedRatio, turbulence);
                    f.flameSpeed = 2.0 * ratio; // laminar speed 2 m/s (ASSUMPTIONS.md #7)
                    f.lit = true; f.travel_x = 0; f.travel_y = 0; f.lastVolume = 0;
                }
            }
        }
    }
    revLimitTimer -= dt;
    if (revLimitTimer < 0) revLimitTimer = 0;
    if (std::fabs(bodies[0].w) > P.revLimit) revLimitTimer = P.limiterDuration;
    lastCycleAngle = now;
}

// CombustionChamber::update(dt): piston work on the gas, flame progress, heat release.
void Engine::chamberUpdate(int i) {
    const EsbCylinder &c = P.cyl[i];
    GasSystem &g = chamber[i];
    const double V = chamberVolume(i);
    g.changeVolume(V - g.V);
    g.V = V;
    const double area = kPi * (0.5 * c.bore) * (0.5 * c.bore);
    g.height = V / area;
    FlameEvent &f = flame[i];
    if (!f.lit) return;
    const double maxX = 0.5 * c.bore, maxY = g.height;
    f.travel_x = std::fmin(f.travel_x + dt * f.flameSpeed, maxX);
    f.travel_y = std::fmin(f.travel_y + dt * f.flameSpeed, maxY);
    const double fx = f.travel_x / maxX, fy = f.travel_y / maxY;
    const double frac = fx * fx * fy;
    const double last = f.lastVolume; // burned fraction so far
    if (frac > last && last < 1.0) {
        const double share = (frac - last) / (1.0 - last); // of what is still unburned
        const double fuel = g.mix.p_fuel * g.n_mol * share * f.efficiency;
        const double oxygen = g.mix.p_o2 * g.n_mol;
        const double burned = std::fmin(fuel, oxygen / P.fuelMolecularAfr);
        if (burned > 0 && g.n_mol > 0) {
            const double dFuel = burned / g.n_mol, dO2 = burned * P.fuelMolecularAfr / g.n_mol;
            g.mix.p_fuel -= dFuel;
            g.mix.p_o2 -= dO2;
            g.mix.p_inert += dFuel + dO2;
            if (g.mix.p_fuel < 0) g.mix.p_fuel = 0;
            if (g.mix.p_o2 < 0) g.mix.p_o2 = 0;
            g.E_k += burned * P.fuelMolecularMass * P.fuelEnergyDensity;
        }
        f.lastVolume = frac;
    }
    if (f.travel_x >= maxX && f.travel_y >= maxY) f.lit = false;
}

// CombustionChamber::flow(dt): the four exchanges of one cylinder in upstream order.
void Engine::chamberFlow(int i, double sdt) {
    const EsbCylinder &c = P.cyl[i];
    const EsbIntake &in = P.intakes[c.intake];
    const EsbExhaustSystem &ex = P.exhaust[c.exhaustSystem];
    const double boreArea = kPi * (0.5 * c.bore) * (0.5 * c.bore);

    // valve lifts from the cam angle (Camshaft::valveLift)
    const double cyc = cycleAngle();
    auto lift = [&](double advance, double centerline, int fn) {
        double cam = positiveMod((cyc + advance) * 0.5, 2.0 * kPi);
        double a = cam - 0.5 * centerline;
        a = positiveMod(a + kPi, 2.0 * kPi) - kPi;
        return sampleTriangle(P, fn, a);
    };
    const double kIntake = sampleTriangle(P, c.fnIntakePortFlow, lift(c.intakeCamAdvance, c.intakeLobeCenterline, c.fnIntakeLobe));
    const double kExhaust = sampleTriangle(P, c.fnExhaustPortFlow, lift(c.exhaustCamAdvance, c.exhaustLobeCenterline, c.fnExhaustLobe));

    if (c.blowbyK != 0) { // blow-by to the crankcase reservoir
        GasSystem &g = chamber[i];
        double flow = sdt * GasSystem::flowRate(c.blowbyK, g.pressure(), c.crankcasePressure, g.temperature(), kT25C, g.hcr(), g.chokedFlowLimit, g.chokedFlowFactor);
        if (flow > 0) {
            flow = std::fmin(flow, 0.5 * g.n_mol);
            const double before = g.n_mol;
            g.loseN(flow, g.kineticEnergyPerMol());
            if (before > 0) { const double k = g.n_mol / before; g.mom[0] *= k; g.mom[1] *= k; }
        }
    }

    FlowParameters fp;
    fp.dt = sdt;
    // exhaust runner -> collector
    fp.system_0 = &exhaustRunner[i]; fp.system_1 = &collector[c.exhaustSystem];
    fp.crossSectionArea_0 = c.exhaustRunnerCrossSectionArea; fp.crossSectionArea_1 = ex.collectorCrossSectionArea;
    fp.direction_x = 1.0; fp.direction_y = 0.0; fp.k_flow = ex.primaryFlowRate;
    gasFlow(fp);
    // plenum -> intake runner
    fp.system_0 = &plenum[c.intake]; fp.system_1 = &intakeRunner[i];
    fp.crossSectionArea_0 = in.plenumCrossSectionArea; fp.crossSectionArea_1 = c.intakeRunnerCrossSectionArea;
    fp.direction_x = 0.0; fp.direction_y = -1.0; fp.k_flow = in.runnerFlowRate;
    gasFlow(fp);
    // intake runner -> chamber (intake valve)
    fp.system_0 = &intakeRunner[i]; fp.system_1 = &chamber[i];
    fp.crossSectionArea_0 = c.intakeRunnerCrossSectionArea; fp.crossSectionArea_1 = boreArea;
    fp.k_flow = kIntake;
    gasFlow(fp);
    // chamber -> exhaust runner (exhaust valve)
    fp.system_0 = &chamber[i]; fp.system_1 = &exhaustRunner[i];
    fp.crossSectionArea_0 = boreArea; fp.crossSectionArea_1 = c.exhaustRunnerCrossSectionArea;
    fp.direction_x = 1.0; fp.direction_y = 0.0; fp.k_flow = kExhaust;
    const double exFlow = gasFlow(fp);
    totalExhaustFlow += exFlow;

    intakeRunner[i].dissipateExcessVelocity();
    intakeRunner[i].updateVelocity(sdt, in.velocityDecay);
    exhaustRunner[i].dissipateExcessVelocity();
    exhaustRunner[i].updateVelocity(sdt, ex.velocityDecay);
}

// PistonEngineSimulator::writeToSynthesizer — formula quoted at archive/TA2_FINDINGS.md:200-213
// and archive/PHYSICS_ANALYSIS.md:64-82.
void Engine::writeAudio() {
    for (int k = 0; k < ESB_MAX_EXHAUST; ++k) exhaustOut[k] = 0;
    const double att = std::fmin(std::fabs(filteredSpeed) / (40.0 * kPi / 30.0), 1.0);
    const double att3 = att * att * att;
    for (int i = 0; i < N; ++i) {
        const EsbCylinder &c = P.cyl[i];
        const GasSystem &r = exhaustRunner[i];
        const double exhaustFlow = att3 * 1600.0 * (1.0 * (r.pressure() - kAtm) + 0.1 * r.dynamicPressure(1.0, 0.0) + 0.1 * r.dynamicPressure(-1.0, 0.0));
        // DelayFilter::fast_f
        double *ring = &delayRing[(size_t)i * kMaxDelay];
        const double delayed = ring[delayPos[i]];
        ring[delayPos[i]] = exhaustFlow;
        delayPos[i] = (delayPos[i] + 1 == delayLen[i]) ? 0 : delayPos[i] + 1;
        exhaustOut[c.exhaustSystem] += P.exhaust[c.exhaustSystem].audioVolume * (c.soundAttenuation * delayed) / N;
    }
}

void Engine::step() {
    // Engine::update: DirectThrottleLinkage (docs/FIX_PRESET_ENGINE_TESTS.md:60-64)
    {
        double s = ctl.throttle; if (s < 0) s = 0; if (s > 1) s = 1;
        throttlePlate = 1.0 - esb_pow(s, P.throttleGamma);
    }
    solve(); // m_system->process(dt, 1)
    // updateFilteredEngineSpeed
    {
        const double alpha = dt / (0.01 + dt);
        filteredSpeed += alpha * ((-bodies[0].w) - filteredSpeed);
    }
    // simulateStep_()
    ignitionUpdate();
    for (int i = 0; i < N; ++i) chamberUpdate(i);
    const double sdt = dt / fluidSteps;
    totalExhaustFlow = 0;
    Mix air; air.p_fuel = 0; air.p_inert = 0.79; air.p_o2 = 0.21;
    for (int s = 0; s < fluidSteps; ++s) {
        for (int k = 0; k < P.intakeCount; ++k) { // Intake::process
            const EsbIntake &in = P.intakes[k];
            const double airPerFuel = in.molecularAfr / 0.21;
            Mix fuelAir; fuelAir.p_fuel = 1.0 / (1.0 + airPerFuel);
            fuelAir.p_o2 = 0.21 * (1.0 - fuelAir.p_fuel); fuelAir.p_inert = 0.79 * (1.0 - fuelAir.p_fuel);
            GasSystem atmosphere;
            atmosphere.initialize(kAtm, 1000.0, kT25C, fuelAir);
            const double plate = in.idleThrottlePlatePosition * throttlePlate;
            const double attenuation = esb_cos(plate * kPi / 2.0);
            FlowParameters fp; fp.dt = sdt;
            fp.system_0 = &atmosphere; fp.system_1 = &plenum[k];
            fp.crossSectionArea_0 = 10.0; fp.crossSectionArea_1 = in.plenumCrossSectionArea;
            fp.direction_x = 0.0; fp.direction_y = -1.0;
            fp.k_flow = attenuation * in.inputFlowK;
            gasFlow(fp);
            if (in.idleFlowK != 0) {
                atmosphere.reset(kAtm, kT25C, fuelAir);
                fp.k_flow = in.idleFlowK;
                gasFlow(fp);
            }
            plenum[k].dissipateExcessVelocity();
            plenum[k].updateVelocity(sdt, in.velocityDecay);
        }
        for (int i = 0; i < N; ++i) chamberFlow(i, sdt);
        for (int k = 0; k < P.exhaustSystemCount; ++k) { // ExhaustSystem::process
            const EsbExhaustSystem &ex = P.exhaust[k];
            GasSystem atmosphere;
            atmosphere.initialize(kAtm, 1000.0, kT25C, air);
            FlowParameters fp; fp.dt = sdt;
            fp.system_0 = &collector[k]; fp.system_1 = &atmosphere;
            fp.crossSectionArea_0 = ex.collectorCrossSectionArea; fp.crossSectionArea_1 = 10.0;
            fp.direction_x = 1.0; fp.direction_y = 0.0;
            fp.k_flow = ex.outletFlowRate;
            gasFlow(fp);
            collector[k].dissipateExcessVelocity();
            collector[k].updateVelocity(sdt, ex.velocityDecay);
        }
    }
    totalExhaustFlow /= dt;
    writeAudio();
    ++stepIndex;
}

} // namespace oracle
