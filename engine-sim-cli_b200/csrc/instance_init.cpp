// instance_init.cpp — host side: derived engine-type constants and the initial state of an
// instance (bodies placed on the slider-crank geometry, gas cells at rest).
//
// What the reference does once per simulator in PistonEngineSimulator::loadSimulation /
// placeAndInitialize (called from SimulatorFactory.cpp:73-93) and Simulator::initialize
// (SimulatorFactory.cpp:164-172): one rigid body per crankshaft / rod / piston plus the
// vehicle mass (SimulatorInitHelpers.cpp:25-28), one gas system per chamber, runner, plenum
// and collector, intake reset to air at 1 atm / 25 C (IntakeDeserializer.cpp, last lines).
// Here it produces one column of the SoA state block, which the batch replicates.
#include "host.h"
#include "esb_math.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

namespace esb {

namespace {
constexpr double kR = 8.31446261815324;
constexpr double kAtm = 101325.0;
constexpr double kT25C = 298.15;
constexpr double kPi = 3.14159265358979323846;

struct CellInit { double n, E, V, mx, my, pf, pi, po; };
CellInit makeCell(double P, double V, double T, double pf, double pi, double po) {
    CellInit c;
    c.V = V;
    c.n = P * V / (kR * T);
    c.E = T * (0.5 * 5 * c.n * kR);
    c.mx = c.my = 0;
    c.pf = pf; c.pi = pi; c.po = po;
    return c;
}
void putCell(std::vector<double> &col, int field, int stride, const CellInit &c) {
    const double v[8] = {c.n, c.E, c.V, c.mx, c.my, c.pf, c.pi, c.po};
    for (int i = 0; i < 8; ++i) col[field + i * stride] = v[i];
}
} // namespace

// Lanes per instance: a power of two >= the cylinder count (lane c owns cylinder c's rigid bodies), and at least minLanes: the gas
// network is scheduled over lane PAIRS whatever the cylinder count (gas_schedule.cpp), so a small engine on more lanes has more
// exchanges in flight per slot and a shorter step, at the price of fewer instances per warp.
int lanesForCylinders(int n, int minLanes) {
    int g = 1;
    while (g < n || g < minLanes) g <<= 1;
    return g;
}

void deriveDeviceParams(const EsbEngineParams &P, double simFreq, int fluidSteps, int maxIterations, double minDelta, double beta,
                        uint64_t seedBase, EsbDeviceParams &D, int minLanes) {
    std::memset(&D, 0, sizeof(D));
    D.P = P;
    if (P.cylinderCount < 1 || P.cylinderCount > ESB_MAX_CYL) throw std::runtime_error("cylinderCount out of range");
    D.dt = 1.0 / simFreq;
    D.invDt = 1.0 / D.dt;
    D.fluidSteps = fluidSteps;
    D.fluidDt = D.dt / fluidSteps;
    D.speedFilterAlpha = D.dt / (0.01 + D.dt);
    D.attenuationSpeed = 40.0 * kPi / 30.0;
    D.solverBeta = beta; D.solverMinDelta = minDelta; D.solverMaxIterations = maxIterations;
    if (const char *e = std::getenv("ESB_MIN_LANES")) minLanes = std::atoi(e);   // developer override (tools/bench_engines_quick.sh)
    D.lanesPerInstance = lanesForCylinders(P.cylinderCount, minLanes < 1 ? 1 : (minLanes > 16 ? 16 : minLanes));
    D.seedBase = seedBase;
    D.invCrankMass = 1.0 / (P.crankMass + P.flywheelMass);
    D.invCrankInertia = 1.0 / P.crankMomentOfInertia;
    D.invVehMass = 1.0; D.invVehInertia = 1.0; // SimulatorInitHelpers.cpp:25-28
    const int dof = 5;
    const double g = 1.0 + 2.0 / dof;
    D.hcr = g;
    D.chokedFlowLimit = esb_pow(2.0 / (g + 1.0), g / (g - 1.0));
    D.chokedFlowFactor = std::sqrt(g) * esb_pow(2.0 / (g + 1.0), (g + 1.0) / (2.0 * (g - 1.0)));
    {
        const CellInit atm = makeCell(kAtm, 1000.0, kT25C, 0, 0, 0);
        D.atmN_air = atm.n; D.atmE_air = atm.E;
    }
    for (int i = 0; i < P.intakeCount; ++i) {
        const EsbIntake &in = P.intakes[i];
        D.plenumWidth[i] = in.plenumVolume / in.plenumCrossSectionArea;
        D.plenumHeight[i] = std::sqrt(in.plenumCrossSectionArea);
        const double airPerFuel = in.molecularAfr / 0.21;
        const double pf = 1.0 / (1.0 + airPerFuel);
        D.fuelAirFuel[i] = pf;
        D.fuelAirO2[i] = 0.21 * (1.0 - pf);
        D.fuelAirInert[i] = 0.79 * (1.0 - pf);
    }
    for (int i = 0; i < P.exhaustSystemCount; ++i) {
        D.collectorWidth[i] = P.exhaust[i].length;
        D.collectorHeight[i] = std::sqrt(P.exhaust[i].collectorCrossSectionArea);
    }
    for (int i = 0; i < P.cylinderCount; ++i) {
        const EsbCylinder &c = P.cyl[i];
        EsbCylDerived &d = D.cd[i];
        esb_sincos(c.bankAngle, &d.sinBank, &d.cosBank);
        double sJ, cJ; esb_sincos(P.journalAngle[c.journal], &sJ, &cJ);
        d.journalLX = P.crankThrow * cJ; d.journalLY = P.crankThrow * sJ;
        d.parentInvMass = D.invCrankMass; d.parentInvInertia = D.invCrankInertia;
        if (c.masterCylinder >= 0) {   // slave rod: the journal is a point of the master rod (oracle/solver.cpp, same expressions)
            const EsbCylinder &mc = P.cyl[c.masterCylinder];
            double sA, cA; esb_sincos(c.slaveJournalAngle, &sA, &cA);
            const double masterBigEnd = -(0.5 * mc.rodLength + mc.rodCenterOfMass);
            d.journalLX = -sA * mc.slaveThrow; d.journalLY = masterBigEnd + cA * mc.slaveThrow;
            d.parentInvMass = 1.0 / mc.rodMass; d.parentInvInertia = 1.0 / mc.rodMomentOfInertia;
        }
        d.littleEnd = 0.5 * c.rodLength - c.rodCenterOfMass;
        d.bigEnd = -(0.5 * c.rodLength + c.rodCenterOfMass);
        d.boreArea = kPi * (0.5 * c.bore) * (0.5 * c.bore);
        d.invPistonMass = 1.0 / c.pistonMass;
        d.invRodMass = 1.0 / c.rodMass;
        d.invRodInertia = 1.0 / c.rodMomentOfInertia;
        d.frictionCoulomb = c.frictionCoeff * 500.0;
        d.frictionVSt = c.breakawayFrictionVelocity * 1.4142135623730951;
        d.frictionVCoul = c.breakawayFrictionVelocity / 10.0;
        d.frictionBreakawayScale = 1.4142135623730951 * 2.718281828459045 * (c.breakawayFriction - d.frictionCoulomb);
        d.intakeRunnerHeight = std::sqrt(c.intakeRunnerCrossSectionArea);
        d.exhaustRunnerHeight = std::sqrt(c.exhaustRunnerCrossSectionArea);
        // DelayFilter length: header primary length over the speed of sound, in steps
        int len = (int)((c.headerPrimaryLength / 343.0) * simFreq + 0.5);
        if (len < 1) len = 1;
        if (len > ESB_MAX_DELAY) len = ESB_MAX_DELAY;
        d.delayLen = len;
    }
    buildGasNet(P, D);
}

// One column (ESB_STATE_FIELDS doubles) of the state block: the state every instance starts from.
void initialStateColumn(const EsbDeviceParams &D, std::vector<double> &col) {
    const EsbEngineParams &P = D.P;
    col.assign(ESB_STATE_FIELDS, 0.0);
    const int N = P.cylinderCount;
    // crankshaft at theta = 0
    col[ESB_GF_CRANK + 0] = P.crankPosX; col[ESB_GF_CRANK + 1] = P.crankPosY;
    double crankTh = 0;
    (void)crankTh;
    for (int i = 0; i < N; ++i) {
        const EsbCylinder &c = P.cyl[i];
        const double phi = P.journalAngle[c.journal];
        double jx = P.crankPosX + P.crankThrow * esb_cos(phi), jy = P.crankPosY + P.crankThrow * esb_sin(phi);
        if (c.masterCylinder >= 0) {   // slave rod (radial engines): the journal is a point of the master rod, which is already placed
            const EsbCylinder &mc = P.cyl[c.masterCylinder];
            const int mf = ESB_CYL_FIELD(ESB_CF_ROD, c.masterCylinder);
            const double mpx = col[mf], mpy = col[mf + ESB_MAX_CYL], mth = col[mf + 2 * ESB_MAX_CYL];
            const double lx = -esb_sin(c.slaveJournalAngle) * mc.slaveThrow;
            const double ly = -(0.5 * mc.rodLength + mc.rodCenterOfMass) + esb_cos(c.slaveJournalAngle) * mc.slaveThrow;
            const double sM = esb_sin(mth), cM = esb_cos(mth);
            jx = mpx + (cM * lx - sM * ly); jy = mpy + (sM * lx + cM * ly);
        }
        const double dx = -esb_sin(c.bankAngle), dy = esb_cos(c.bankAngle);
        const double nx = dy, ny = -dx;
        const double rx = jx - c.bankX, ry = jy - c.bankY;
        const double along = rx * dx + ry * dy, perp = rx * nx + ry * ny;
        const double L = c.rodLength;
        const double s = along + std::sqrt(L * L - perp * perp);
        const double pinx = c.bankX + s * dx, piny = c.bankY + s * dy;
        const double ux = (pinx - jx) / L, uy = (piny - jy) / L;
        const double littleEnd = 0.5 * L - c.rodCenterOfMass;
        const int st = ESB_MAX_CYL;
        const int pf = ESB_CYL_FIELD(ESB_CF_PISTON, i), rf = ESB_CYL_FIELD(ESB_CF_ROD, i);
        const double ppx = pinx, ppy = piny - c.wristPinPosition;
        col[pf] = ppx; col[pf + st] = ppy; col[pf + 2 * st] = 0;
        col[rf] = pinx - ux * littleEnd; col[rf + st] = piny - uy * littleEnd; col[rf + 2 * st] = std::atan2(-ux, uy);
        // combustion chamber at the starting pressure / temperature, all inert
        const double sB = esb_sin(c.bankAngle), cB = esb_cos(c.bankAngle);
        const double ax = -sB, ay = cB;
        const double axial = (ppx - c.bankX) * ax + (ppy - c.bankY) * ay;
        const double area = kPi * (0.5 * c.bore) * (0.5 * c.bore);
        double h = c.deckHeight - (axial + c.compressionHeight);
        if (h < 0) h = 0;
        const double V = area * h + c.combustionChamberVolume;
        putCell(col, ESB_CYL_FIELD(ESB_CF_CHAMBER, i), st, makeCell(c.startingPressure, V, c.startingTemperature, 0, 1, 0));
        putCell(col, ESB_CYL_FIELD(ESB_CF_INTAKE_RUNNER, i), st, makeCell(kAtm, c.intakeRunnerVolume, kT25C, 0, 0.79, 0.21));
        putCell(col, ESB_CYL_FIELD(ESB_CF_EXHAUST_RUNNER, i), st, makeCell(kAtm, c.exhaustRunnerVolume, kT25C, 0, 1, 0));
    }
    for (int k = 0; k < P.intakeCount; ++k)
        putCell(col, ESB_GF_PLENUM + 8 * k, 1, makeCell(kAtm, P.intakes[k].plenumVolume, kT25C, 0, 0.79, 0.21));
    for (int k = 0; k < P.exhaustSystemCount; ++k)
        putCell(col, ESB_GF_COLLECTOR + 8 * k, 1, makeCell(kAtm, P.exhaust[k].collectorCrossSectionArea * P.exhaust[k].length, kT25C, 0, 1, 0));
    col[ESB_GF_THROTTLE_PLATE] = 1.0;
    col[ESB_GF_VEH_INV_INERTIA] = D.invVehInertia;
    // Crankshaft::getCycleAngle at theta = 0
    {
        double r = std::fmod(-(0.0 - P.tdc), 4.0 * kPi);
        if (r < 0) r += 4.0 * kPi;
        col[ESB_GF_LAST_CYCLE] = r;
    }
}

} // namespace esb
