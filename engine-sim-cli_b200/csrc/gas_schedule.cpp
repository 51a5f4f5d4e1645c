// gas_schedule.cpp — host side: the gas network of an engine type (cells, exchange / relax tables) and the
// slot schedule of one simulation step (see EsbGasNet in esb_device.h).
//
// The reference runs the fluid part of a step strictly in sequence (PistonEngineSimulator::simulateStep_,
// quoted at archive/TA2_FINDINGS.md:194-199; restated in oracle/engine_step.cpp Engine::step):
//     for each of fluidSimulationSteps substeps:
//         Intake::process            for every intake   (atmosphere -> plenum, idle circuit, plenum velocity update)
//         CombustionChamber::flow    for every cylinder (blow-by, runner -> collector, plenum -> runner, the two valve
//                                                        exchanges, the two runner velocity updates)
//         ExhaustSystem::process     for every exhaust  (collector -> atmosphere, collector velocity update)
// Every one of these operations reads and writes one or two gas cells and nothing else, so operations on disjoint
// cells commute exactly; what has to be kept is each cell's own sequence.  buildGasNet() writes the step's
// operations down in the reference's order, takes "the previous operation on any of my cells" as the dependencies,
// and list-schedules them (longest remaining path first) into slots of one kind each: up to G/2 two-cell exchanges
// (a lane pair shares one), up to G relaxations, or the blow-by of all chambers of a substep.
#include "host.h"

#include <algorithm>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace esb {

namespace {
struct Op {
    int type;       // EsbGasSlotType
    int kind;       // index into exchange[] / relax[]; for blow-by: unused
    int sub;        // fluid substep
    int cells[2];   // real cells touched (-1: none); blow-by touches every chamber with a blow-by path
    std::vector<int> preds;
    double cost, rank = 0;
    int slot = -1;
};
} // namespace

void buildGasNet(const EsbEngineParams &P, EsbDeviceParams &D) {
    EsbGasNet &g = D.gas;
    std::memset(&g, 0, sizeof(g));
    const int N = P.cylinderCount, nIn = P.intakeCount, nEx = P.exhaustSystemCount;
    const int G = D.lanesPerInstance;
    g.nCells = 3 * N + nIn + nEx;
    g.nConst = nIn + 2;
    g.entriesPerSlot = G;
    g.split = G >= 2 ? 1 : 0;
    g.pairsPerSlot = G >= 2 ? G / 2 : 1;
    auto chamber = [&](int c) { return 3 * c; };
    auto intr = [&](int c) { return 3 * c + 1; };
    auto exhr = [&](int c) { return 3 * c + 2; };
    auto plenum = [&](int i) { return 3 * N + i; };
    auto collector = [&](int e) { return 3 * N + nIn + e; };
    const int atmIn0 = g.nCells, atmOut = g.nCells + nIn, rest = g.nCells + nIn + 1;

    // constant cells
    for (int i = 0; i < nIn; ++i) {
        const double v[8] = {D.atmN_air, D.atmE_air, 1000.0, 0, 0, D.fuelAirFuel[i], D.fuelAirInert[i], D.fuelAirO2[i]};
        std::memcpy(g.constCells[i], v, sizeof v);
    }
    {
        const double air[8] = {D.atmN_air, D.atmE_air, 1000.0, 0, 0, 0, 0.79, 0.21};
        std::memcpy(g.constCells[nIn], air, sizeof air);
        const double atRest[8] = {1.0, 6000.0, 0.024, 0, 0, 0, 0.79, 0.21};
        std::memcpy(g.constCells[nIn + 1], atRest, sizeof atRest);
    }

    // operation tables
    auto addExchange = [&](int c0, int c1, double k, int kSource, double dx, double dy, double a0, double a1, int flowCyl) {
        EsbGasExchange &e = g.exchange[g.nExchange];
        e.cell0 = c0; e.cell1 = c1; e.kConst = k; e.kSource = kSource; e.dirx = dx; e.diry = dy; e.area0 = a0; e.area1 = a1; e.flowCylinder = flowCyl;
        return g.nExchange++;
    };
    auto addRelax = [&](int cell, double beta, double w, double h, double dx, double dy) {
        EsbGasRelax &r = g.relax[g.nRelax];
        r.cell = cell; r.beta = beta; r.width = w; r.height = h; r.dx = dx; r.dy = dy;
        return g.nRelax++;
    };
    std::vector<int> exInlet(nIn), exIdle(nIn, -1), rxPlenum(nIn), exOutlet(nEx), rxCollector(nEx);
    std::vector<int> exC(N), exP(N), exVI(N), exVE(N), rxIntr(N), rxExhr(N);
    for (int i = 0; i < nIn; ++i) {
        const EsbIntake &in = P.intakes[i];
        exInlet[i] = addExchange(atmIn0 + i, plenum(i), 0.0, 2 * N + i, 0.0, -1.0, 10.0, in.plenumCrossSectionArea, -1);
        if (in.idleFlowK != 0) exIdle[i] = addExchange(atmIn0 + i, plenum(i), in.idleFlowK, -1, 0.0, -1.0, 10.0, in.plenumCrossSectionArea, -1);
        rxPlenum[i] = addRelax(plenum(i), in.velocityDecay, D.plenumWidth[i], D.plenumHeight[i], 0.0, -1.0);
    }
    bool anyBlowby = false;
    for (int c = 0; c < N; ++c) {
        const EsbCylinder &cy = P.cyl[c];
        const EsbIntake &in = P.intakes[cy.intake];
        const EsbExhaustSystem &ex = P.exhaust[cy.exhaustSystem];
        anyBlowby = anyBlowby || cy.blowbyK != 0;
        exC[c] = addExchange(exhr(c), collector(cy.exhaustSystem), ex.primaryFlowRate, -1, 1.0, 0.0, cy.exhaustRunnerCrossSectionArea, ex.collectorCrossSectionArea, -1);
        exP[c] = addExchange(plenum(cy.intake), intr(c), in.runnerFlowRate, -1, 0.0, -1.0, in.plenumCrossSectionArea, cy.intakeRunnerCrossSectionArea, -1);
        exVI[c] = addExchange(intr(c), chamber(c), 0.0, c, 0.0, -1.0, cy.intakeRunnerCrossSectionArea, D.cd[c].boreArea, -1);
        exVE[c] = addExchange(chamber(c), exhr(c), 0.0, N + c, 1.0, 0.0, D.cd[c].boreArea, cy.exhaustRunnerCrossSectionArea, c);
        rxIntr[c] = addRelax(intr(c), in.velocityDecay, in.runnerLength, D.cd[c].intakeRunnerHeight, 0.0, -1.0);
        rxExhr[c] = addRelax(exhr(c), ex.velocityDecay, ex.primaryTubeLength, D.cd[c].exhaustRunnerHeight, 1.0, 0.0);
    }
    for (int e = 0; e < nEx; ++e) {
        const EsbExhaustSystem &ex = P.exhaust[e];
        exOutlet[e] = addExchange(collector(e), atmOut, ex.outletFlowRate, -1, 1.0, 0.0, ex.collectorCrossSectionArea, 10.0, -1);
        rxCollector[e] = addRelax(collector(e), ex.velocityDecay, D.collectorWidth[e], D.collectorHeight[e], 1.0, 0.0);
    }
    {   // the idle exchange: two cells at rest, nothing flows, nothing is stored
        EsbGasExchange &e = g.exchange[g.nExchange];
        e.cell0 = rest; e.cell1 = rest; e.kConst = 0; e.kSource = -1; e.dirx = 1.0; e.diry = 0.0; e.area0 = 10.0; e.area1 = 10.0; e.flowCylinder = -1;
    }

    // the step's operations in the reference's order
    std::vector<Op> ops;
    auto push = [&](int type, int kind, int sub, int c0, int c1, double cost) {
        Op o; o.type = type; o.kind = kind; o.sub = sub; o.cells[0] = c0; o.cells[1] = c1; o.cost = cost;
        ops.push_back(o);
    };
    auto real = [&](int cell) { return cell < g.nCells ? cell : -1; };
    const double costExchange = 6.0, costRelax = 3.5, costBlowby = 1.5;
    for (int s = 0; s < D.fluidSteps; ++s) {
        for (int i = 0; i < nIn; ++i) {
            push(ESB_SLOT_EXCHANGE, exInlet[i], s, plenum(i), -1, costExchange);
            if (exIdle[i] >= 0) push(ESB_SLOT_EXCHANGE, exIdle[i], s, plenum(i), -1, costExchange);
            push(ESB_SLOT_RELAX, rxPlenum[i], s, plenum(i), -1, costRelax);
        }
        // Blow-by only touches its own chamber and sits first in the cylinder's sequence; the chambers' blow-bys of a
        // substep are taken together (one slot, lane c does cylinder c), which is the same order for every chamber.
        if (anyBlowby) push(ESB_SLOT_BLOWBY, 0, s, -2, -1, costBlowby);
        for (int c = 0; c < N; ++c) {
            const EsbGasExchange *e;
            e = &g.exchange[exC[c]]; push(ESB_SLOT_EXCHANGE, exC[c], s, real(e->cell0), real(e->cell1), costExchange);
            e = &g.exchange[exP[c]]; push(ESB_SLOT_EXCHANGE, exP[c], s, real(e->cell0), real(e->cell1), costExchange);
            e = &g.exchange[exVI[c]]; push(ESB_SLOT_EXCHANGE, exVI[c], s, real(e->cell0), real(e->cell1), costExchange);
            e = &g.exchange[exVE[c]]; push(ESB_SLOT_EXCHANGE, exVE[c], s, real(e->cell0), real(e->cell1), costExchange);
            push(ESB_SLOT_RELAX, rxIntr[c], s, intr(c), -1, costRelax);
            push(ESB_SLOT_RELAX, rxExhr[c], s, exhr(c), -1, costRelax);
        }
        for (int e = 0; e < nEx; ++e) {
            push(ESB_SLOT_EXCHANGE, exOutlet[e], s, collector(e), -1, costExchange);
            push(ESB_SLOT_RELAX, rxCollector[e], s, collector(e), -1, costRelax);
        }
    }

    // dependencies: the previous operation on each of my cells (blow-by: every chamber)
    {
        std::vector<int> last(g.nCells, -1);
        for (int j = 0; j < (int)ops.size(); ++j) {
            Op &o = ops[j];
            auto touch = [&](int cell) {
                if (last[cell] >= 0 && std::find(o.preds.begin(), o.preds.end(), last[cell]) == o.preds.end()) o.preds.push_back(last[cell]);
                last[cell] = j;
            };
            if (o.cells[0] == -2) { for (int c = 0; c < N; ++c) touch(chamber(c)); }
            else { for (int q = 0; q < 2; ++q) if (o.cells[q] >= 0) touch(o.cells[q]); }
        }
    }
    // rank = longest path from the operation to the end of the step
    {
        std::vector<std::vector<int>> succ(ops.size());
        for (int j = 0; j < (int)ops.size(); ++j) for (int p : ops[j].preds) succ[p].push_back(j);
        for (int j = (int)ops.size() - 1; j >= 0; --j) {
            double best = 0;
            for (int q : succ[j]) best = std::max(best, ops[q].rank);
            ops[j].rank = ops[j].cost + best;
        }
    }
    // list scheduling into typed slots
    int remaining = (int)ops.size();
    int nSlots = 0;
    while (remaining > 0) {
        if (nSlots >= ESB_GAS_MAX_SLOTS) throw std::runtime_error("gas schedule does not fit (too many fluid substeps for this engine)");
        std::vector<int> ready;
        for (int j = 0; j < (int)ops.size(); ++j) {
            if (ops[j].slot >= 0) continue;
            bool ok = true;
            for (int p : ops[j].preds) ok = ok && ops[p].slot >= 0 && ops[p].slot < nSlots;
            if (ok) ready.push_back(j);
        }
        std::stable_sort(ready.begin(), ready.end(), [&](int a, int b) { return ops[a].rank > ops[b].rank; });
        const int type = ops[ready[0]].type;
        const int cap = (type == ESB_SLOT_EXCHANGE) ? g.pairsPerSlot : (type == ESB_SLOT_RELAX ? G : 1);
        uint16_t *row = &g.slotOps[(size_t)nSlots * 16];
        for (int l = 0; l < 16; ++l) row[l] = ESB_GAS_IDLE;
        g.slotType[nSlots] = (uint8_t)type;
        int used = 0;
        for (int j : ready) {
            if (ops[j].type != type || used >= cap) continue;
            const uint16_t entry = (uint16_t)(ops[j].kind | (ops[j].sub << 8));
            if (type == ESB_SLOT_EXCHANGE) {
                row[used] = entry;
                if (g.split) row[used + G / 2] = entry;   // the partner lane of the pair
            } else if (type == ESB_SLOT_RELAX) {
                row[used] = entry;
            } else {
                for (int l = 0; l < N; ++l) row[l] = (P.cyl[l].blowbyK != 0) ? entry : (uint16_t)ESB_GAS_IDLE;   // lane c: cylinder c
            }
            ops[j].slot = nSlots;
            ++used; --remaining;
        }
        ++nSlots;
    }
    g.nSlots = nSlots;
    if (D.fluidSteps > 255) throw std::runtime_error("fluidSimulationSteps out of range");
}

} // namespace esb
