// gas.cpp — oracle restatement of upstream gas_system.cpp (TEST INFRASTRUCTURE ONLY).
// Evidence: SURVEY.md Appendix A.4 [UPSTREAM-RECALL]; helper use in the reference at
// engine-sim-bridge/test/spikes/ClutchParameterSweepSpike.cpp:104-112,138-156; the
// dynamicPressure() call in the exhaust-audio formula quoted at archive/TA2_FINDINGS.md:200-213.
#include "oracle.h"
#include "esb_math.h"

#include <cmath>

namespace oracle {

void GasSystem::initialize(double P, double V_, double T, const Mix &m, int dof_) {
    dof = dof_;
    V = V_;
    n_mol = P * V_ / (kR * T);
    E_k = T * (0.5 * dof * n_mol * kR);
    mix = m;
    mom[0] = mom[1] = 0;
    const double g = hcr();
    chokedFlowLimit = esb_pow(2.0 / (g + 1.0), g / (g - 1.0));
    chokedFlowFactor = std::sqrt(g) * esb_pow(2.0 / (g + 1.0), (g + 1.0) / (2.0 * (g - 1.0)));
}

void GasSystem::reset(double P, double T, const Mix &m) {
    n_mol = P * V / (kR * T);
    E_k = T * (0.5 * dof * n_mol * kR);
    mix = m;
    mom[0] = mom[1] = 0;
}

double GasSystem::c() const {
    if (n_mol == 0 || V == 0) return 0;
    const double rho = mass() / V;
    return std::sqrt(hcr() * pressure() / rho);
}

double GasSystem::bulkKineticEnergy() const {
    const double m = mass();
    if (m == 0) return 0;
    return (mom[0] * mom[0] + mom[1] * mom[1]) / (2.0 * m);
}

// stagnation minus static pressure of the bulk flow component along (dirx, diry)
double GasSystem::dynamicPressure(double dirx, double diry) const {
    if (n_mol == 0) return 0;
    const double invM = 1.0 / mass();
    const double v = (mom[0] * invM) * dirx + (mom[1] * invM) * diry;
    if (v <= 0) return 0;
    const double g = hcr();
    const double cs = c();
    const double mach = v / cs;
    const double x = 1.0 + ((g - 1.0) / 2.0) * mach * mach;
    return pressure() * (esb_pow(x, g / (g - 1.0)) - 1.0);
}

void GasSystem::changeVolume(double dV) {
    const double W = pressure() * dV; // work done by the gas
    V += dV;
    E_k -= W;
    if (E_k < 0) E_k = 0;
}

void GasSystem::gainN(double dn, double E_k_per_mol, const Mix &m) {
    const double next_n = n_mol + dn;
    const double current_n = n_mol;
    E_k += dn * E_k_per_mol;
    n_mol = next_n;
    if (next_n != 0) {
        mix.p_fuel = (mix.p_fuel * current_n + dn * m.p_fuel) / next_n;
        mix.p_inert = (mix.p_inert * current_n + dn * m.p_inert) / next_n;
        mix.p_o2 = (mix.p_o2 * current_n + dn * m.p_o2) / next_n;
    } else {
        mix.p_fuel = mix.p_inert = mix.p_o2 = 0;
    }
}

void GasSystem::loseN(double dn, double E_k_per_mol) {
    E_k -= E_k_per_mol * dn;
    n_mol -= dn;
    if (n_mol < 0) n_mol = 0;
    if (E_k < 0) E_k = 0;
}

// Amount that equalises static pressure when it leaves *this (the higher-pressure side)
// carrying this system's energy per mol.
double GasSystem::pressureEquilibriumMaxFlow(const GasSystem &sink) const {
    const double e = kineticEnergyPerMol();
    if (e == 0) return 0;
    const double maxFlow = (sink.V * E_k - V * sink.E_k) / (e * (sink.V + V));
    return std::fmax(0.0, std::fmin(maxFlow, n_mol));
}

double GasSystem::flowRate(double k, double P0, double P1, double T0, double T1, double g, double chokedLimit, double chokedFactor) {
    if (k == 0) return 0;
    double direction, T_up, p_up, p_dn;
    if (P0 > P1) { direction = 1.0; T_up = T0; p_up = P0; p_dn = P1; }
    else { direction = -1.0; T_up = T1; p_up = P1; p_dn = P0; }
    if (p_up <= 0 || T_up <= 0) return 0;
    const double ratio = p_dn / p_up;
    double rate;
    if (ratio <= chokedLimit) {
        rate = chokedFactor / std::sqrt(kR * T_up);
    } else {
        const double s = esb_pow(ratio, 1.0 / g);
        rate = (2.0 * g) / (g - 1.0);
        rate *= s * (s - ratio);
        rate = std::sqrt(std::fmax(rate, 0.0) / (kR * T_up));
    }
    rate *= direction * p_up;
    return rate * k;
}

void GasSystem::dissipateExcessVelocity() {
    const double m = mass();
    if (m == 0) return;
    const double vx = mom[0] / m, vy = mom[1] / m;
    const double v2 = vx * vx + vy * vy;
    const double cs = c();
    const double c2 = cs * cs;
    if (c2 >= v2 || v2 == 0) return;
    const double k = std::sqrt(c2 / v2);
    mom[0] *= k;
    mom[1] *= k;
    E_k += 0.5 * m * (v2 - c2);
    if (E_k < 0) E_k = 0;
}

void GasSystem::updateVelocity(double dt, double beta) {
    if (n_mol == 0) return;
    const double depth = V / (width * height);
    const double p0 = dynamicPressure(dx, dy);
    const double p1 = dynamicPressure(-dx, -dy);
    const double p2 = dynamicPressure(dy, dx);
    const double p3 = dynamicPressure(-dy, -dx);
    const double sa0 = p0 * (height * depth);
    const double sa1 = p1 * (height * depth);
    const double sa2 = p2 * (width * depth);
    const double sa3 = p3 * (width * depth);
    double dmx = 0, dmy = 0;
    dmx += sa0 * dx; dmy += sa0 * dy;
    dmx -= sa1 * dx; dmy -= sa1 * dy;
    dmx += sa2 * dy; dmy += sa2 * dx;
    dmx -= sa3 * dy; dmy -= sa3 * dx;
    const double m = mass();
    const double invM = 1.0 / m;
    const double v0x = mom[0] * invM, v0y = mom[1] * invM;
    mom[0] -= dmx * dt * beta;
    mom[1] -= dmy * dt * beta;
    const double v1x = mom[0] * invM, v1y = mom[1] * invM;
    E_k -= 0.5 * m * ((v1x * v1x - v0x * v0x) + (v1y * v1y - v0y * v0y));
    if (E_k < 0) E_k = 0;
}

// GasSystem::flow(const FlowParameters&): exchange between two cells through an orifice.
double gasFlow(const FlowParameters &p) {
    GasSystem &s0 = *p.system_0, &s1 = *p.system_1;
    const double P_0 = s0.pressure() + s0.dynamicPressure(p.direction_x, p.direction_y);
    const double P_1 = s1.pressure() + s1.dynamicPressure(-p.direction_x, -p.direction_y);
    GasSystem *source, *sink;
    double dx, dy, sourceP, sinkP, sourceArea, sinkArea, direction;
    if (P_0 > P_1) {
        dx = p.direction_x; dy = p.direction_y;
        source = &s0; sink = &s1; sourceP = P_0; sinkP = P_1;
        sourceArea = p.crossSectionArea_0; sinkArea = p.crossSectionArea_1; direction = 1.0;
    } else {
        dx = -p.direction_x; dy = -p.direction_y;
        source = &s1; sink = &s0; sourceP = P_1; sinkP = P_0;
        sourceArea = p.crossSectionArea_1; sinkArea = p.crossSectionArea_0; direction = -1.0;
    }
    double flow = p.dt * GasSystem::flowRate(p.k_flow, sourceP, sinkP, source->temperature(), sink->temperature(),
                                             source->hcr(), source->chokedFlowLimit, source->chokedFlowFactor);
    const double maxFlow = source->pressureEquilibriumMaxFlow(*sink);
    flow = std::fmax(0.0, std::fmin(flow, maxFlow));
    if (source->n_mol == 0) return 0;

    const double fraction = flow / source->n_mol;
    const double fractionVolume = fraction * source->V;
    const double fractionMass = fraction * source->mass();

    if (flow != 0) {
        const double bulk0 = source->bulkKineticEnergy() + sink->bulkKineticEnergy();
        const double e = source->kineticEnergyPerMol();
        sink->gainN(flow, e, source->mix);
        source->loseN(flow, e);
        const double dpx = source->mom[0] * fraction, dpy = source->mom[1] * fraction;
        source->mom[0] -= dpx; source->mom[1] -= dpy;
        sink->mom[0] += dpx; sink->mom[1] += dpy;
        const double bulk1 = source->bulkKineticEnergy() + sink->bulkKineticEnergy();
        sink->E_k -= (bulk1 - bulk0);
    }

    const double sourceMass = source->mass(), sinkMass = sink->mass();
    const double cSource = source->c(), cSink = sink->c();
    const double srcMom0x = source->mom[0], srcMom0y = source->mom[1];
    const double snkMom0x = sink->mom[0], snkMom0y = sink->mom[1];

    if (sinkArea != 0) {
        const double v = std::fmax(0.0, std::fmin((fractionVolume / sinkArea) / p.dt, cSink));
        sink->mom[0] += (v * dx) * fractionMass;
        sink->mom[1] += (v * dy) * fractionMass;
    }
    if (sourceArea != 0 && sourceMass != 0) {
        const double v = std::fmax(0.0, std::fmin((fractionVolume / sourceArea) / p.dt, cSource));
        source->mom[0] += (v * dx) * fractionMass;
        source->mom[1] += (v * dy) * fractionMass;
    }
    if (sourceMass != 0) {
        const double inv = 1.0 / sourceMass;
        const double v0x = srcMom0x * inv, v0y = srcMom0y * inv, v1x = source->mom[0] * inv, v1y = source->mom[1] * inv;
        source->E_k -= 0.5 * sourceMass * ((v1x * v1x - v0x * v0x) + (v1y * v1y - v0y * v0y));
    }
    if (sinkMass > 0) {
        const double inv = 1.0 / sinkMass;
        const double v0x = snkMom0x * inv, v0y = snkMom0y * inv, v1x = sink->mom[0] * inv, v1y = sink->mom[1] * inv;
        sink->E_k -= 0.5 * sinkMass * ((v1x * v1x - v0x * v0x) + (v1y * v1y - v0y * v0y));
    }
    if (sink->E_k < 0) sink->E_k = 0;
    if (source->E_k < 0) source->E_k = 0;
    return flow * direction;
}

} // namespace oracle
