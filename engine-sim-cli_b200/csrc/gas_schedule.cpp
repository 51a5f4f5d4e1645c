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
// and list-schedules them into slots of one kind each: up to G/2 two-cell exchanges (a lane pair shares one), up to G
// relaxations, or the blow-by of all chambers of a substep.  The schedule is searched for (a few hundred list-scheduling
// passes over perturbed priorities, see below): a V8 step on 8 lanes goes from 81 exchange + 29 relax slots (longest
// remaining path first) to 74 + 21, which is 1 slot away from the bound its intake chain sets.
#include "host.h"

#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
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
    g.split = G >= 2 ? 2000 : 0;
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
    const int cellBytes = ESB_GAS_CELL_STRIDE * (int)sizeof(double);
    auto cellOffset = [&](int cell) { return cell < g.nCells ? cell * cellBytes : ~((cell - g.nCells) * cellBytes); };
    auto addExchange = [&](int c0, int c1, double k, int kSource, double dx, double dy, double a0, double a1, int flowCyl) {
        EsbGasExchange &e = g.exchange[g.nExchange];
        e.cell0 = c0; e.cell1 = c1; e.kConst = k; e.kSource = kSource; e.dirx = dx; e.diry = dy; e.area0 = a0; e.area1 = a1; e.flowCylinder = flowCyl;
        e.off0 = cellOffset(c0); e.off1 = cellOffset(c1);
        return g.nExchange++;
    };
    auto addRelax = [&](int cell, double beta, double w, double h, double dx, double dy) {
        EsbGasRelax &r = g.relax[g.nRelax];
        r.cell = cell; r.off = cell * cellBytes; r.beta = beta; r.width = w; r.height = h; r.dx = dx; r.dy = dy;
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
        e.off0 = cellOffset(rest); e.off1 = cellOffset(rest);
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
        std::vector<std::vector<int>> after(ops.size());
        for (int j = 0; j < (int)ops.size(); ++j) for (int p : ops[j].preds) after[p].push_back(j);
        for (int j = (int)ops.size() - 1; j >= 0; --j) {
            double best = 0;
            for (int q : after[j]) best = std::max(best, ops[q].rank);
            ops[j].rank = ops[j].cost + best;
        }
    }
    // List scheduling into typed slots.  A slot costs its kind's time whatever its fill, so the aim is few slots, weighted
    // by kind.  One pass = (priorities, relax threshold): each slot takes the kind of the ready operation with the highest
    // priority, except that a relax slot is also opened as soon as `relaxThreshold` relaxations are ready (relaxations
    // picked up one at a time, only when the chain stalls on them, leave the other lanes of their slot empty).  The first
    // pass uses the plain ranks; the others perturb them (fixed seed) and vary the threshold; the cheapest schedule wins.
    // Every pass keeps each cell's own operation order, so they all compute the same bits.
    const int nOps = (int)ops.size();
    std::vector<std::vector<int>> succ(nOps);
    for (int j = 0; j < nOps; ++j) for (int p : ops[j].preds) succ[p].push_back(j);
    auto capOf = [&](int type) { return (type == ESB_SLOT_EXCHANGE) ? g.pairsPerSlot : (type == ESB_SLOT_RELAX ? G : 1); };
    auto costOf = [&](int type) { return (type == ESB_SLOT_EXCHANGE) ? costExchange : (type == ESB_SLOT_RELAX ? costRelax : costBlowby); };
    std::vector<int> slotOf(nOps), bestSlotOf;
    std::vector<int> slotKind, bestSlotKind;
    double bestCost = 1e300;
    // the search depends on the shape of the network only: one search per shape and process
    std::string shape = std::to_string(G) + "/" + std::to_string(g.pairsPerSlot) + ":";
    for (const Op &o : ops) shape += std::to_string(o.type) + "," + std::to_string(o.cells[0]) + "," + std::to_string(o.cells[1]) + ";";
    static std::mutex cacheLock;
    static std::map<std::string, std::pair<std::vector<int>, std::vector<int>>> cache;
    {
        std::lock_guard<std::mutex> hold(cacheLock);
        auto hit = cache.find(shape);
        if (hit != cache.end()) { bestSlotOf = hit->second.first; bestSlotKind = hit->second.second; }
    }
    uint64_t rng = 0x9E3779B97F4A7C15ull;
    auto uniform = [&]() {   // xorshift64*, [0, 1)
        rng ^= rng >> 12; rng ^= rng << 25; rng ^= rng >> 27;
        return (double)((rng * 0x2545F4914F6CDD1Dull) >> 11) * (1.0 / 9007199254740992.0);
    };
    std::vector<double> prio(nOps);
    std::vector<int> missing(nOps), ready, chosen;
    const int passes = bestSlotKind.empty() ? 2000 : 0;
    for (int pass = 0; pass < passes; ++pass) {
        int relaxThreshold = 1 << 30;
        for (int j = 0; j < nOps; ++j) prio[j] = ops[j].rank;
        if (pass > 0) {
            static const double spread[6] = {0.0, 0.5, 1.0, 2.0, 4.0, 8.0};
            const double a = spread[(int)(uniform() * 6)];
            relaxThreshold = 2 + (int)(uniform() * (G - 1));
            for (int j = 0; j < nOps; ++j) prio[j] += a * (2.0 * uniform() - 1.0);
        }
        for (int j = 0; j < nOps; ++j) missing[j] = (int)ops[j].preds.size();
        ready.clear();
        for (int j = 0; j < nOps; ++j) if (missing[j] == 0) ready.push_back(j);
        slotKind.clear();
        int remaining = nOps;
        double cost = 0;
        while (remaining > 0 && cost < bestCost) {
            std::stable_sort(ready.begin(), ready.end(), [&](int a, int b) { return prio[a] > prio[b]; });
            int type = ops[ready[0]].type;
            if (type != ESB_SLOT_RELAX) {
                int nRelax = 0;
                for (int j : ready) nRelax += ops[j].type == ESB_SLOT_RELAX;
                if (nRelax >= relaxThreshold) type = ESB_SLOT_RELAX;
            }
            const int cap = capOf(type);
            chosen.clear();
            for (int j : ready) if (ops[j].type == type && (int)chosen.size() < cap) chosen.push_back(j);
            for (int j : chosen) {
                slotOf[j] = (int)slotKind.size();
                ready.erase(std::find(ready.begin(), ready.end(), j));
            }
            for (int j : chosen) for (int q : succ[j]) if (--missing[q] == 0) ready.push_back(q);
            remaining -= (int)chosen.size();
            slotKind.push_back(type);
            cost += costOf(type);
        }
        if (remaining == 0 && cost < bestCost) { bestCost = cost; bestSlotOf = slotOf; bestSlotKind = slotKind; }
    }
    if (passes > 0) {
        std::lock_guard<std::mutex> hold(cacheLock);
        cache[shape] = std::make_pair(bestSlotOf, bestSlotKind);
    }
    const int nSlots = (int)bestSlotKind.size();
    if (nSlots > ESB_GAS_MAX_SLOTS) throw std::runtime_error("gas schedule does not fit (too many fluid substeps for this engine)");
    for (int sl = 0; sl < nSlots; ++sl) {
        const int type = bestSlotKind[sl];
        uint32_t *row = &g.slotOps[(size_t)sl * 16];
        for (int l = 0; l < 16; ++l) row[l] = ESB_GAS_IDLE;
        g.slotType[sl] = (uint8_t)type;
        int used = 0;
        for (int j = 0; j < nOps; ++j) {   // operations of a slot in the reference's order
            if (bestSlotOf[j] != sl) continue;
            const uint32_t entry = (uint32_t)ops[j].kind;
            if (type == ESB_SLOT_EXCHANGE) {
                const int flowCyl = g.exchange[ops[j].kind].flowCylinder;
                row[used] = entry | (flowCyl >= 0 ? (uint32_t)(ops[j].sub * N + flowCyl + 1) << 16 : 0u);   // the first lane of the pair records the flow
                if (g.split) row[used + G / 2] = entry;   // the partner lane of the pair
            } else if (type == ESB_SLOT_RELAX) {
                row[used] = entry;
            } else {
                for (int l = 0; l < N; ++l) row[l] = (P.cyl[l].blowbyK != 0) ? entry : (uint32_t)ESB_GAS_IDLE;   // lane c: cylinder c
            }
            ++used;
        }
    }
    g.nSlots = nSlots;
    if (D.fluidSteps * N > 65534 || D.fluidSteps > 255) throw std::runtime_error("fluidSimulationSteps out of range");
}

// Checks a built schedule against the sequential form it stands for: every operation of the step appears exactly once, no two
// operations of a slot touch the same (stored) cell, and each cell meets its operations in the reference's order.  Returns an empty
// string when the schedule is valid, else what is wrong (used by EsbGasScheduleInfo and the tests).
std::string checkGasSchedule(const EsbEngineParams &P, const EsbDeviceParams &D) {
    const EsbGasNet &g = D.gas;
    const int N = P.cylinderCount, nIn = P.intakeCount, nEx = P.exhaustSystemCount, G = D.lanesPerInstance;
    // the reference's sequence, per cell: (type, kind, substep) in order
    struct Ref { int type, kind, sub; };
    std::vector<std::vector<Ref>> want(g.nCells), got(g.nCells);
    auto touch = [&](std::vector<std::vector<Ref>> &seq, int type, int kind, int sub) {
        if (type == ESB_SLOT_EXCHANGE) {
            const EsbGasExchange &e = g.exchange[kind];
            if (e.cell0 < g.nCells) seq[e.cell0].push_back({type, kind, sub});
            if (e.cell1 < g.nCells && e.cell1 != e.cell0) seq[e.cell1].push_back({type, kind, sub});
        } else if (type == ESB_SLOT_RELAX) {
            seq[g.relax[kind].cell].push_back({type, kind, sub});
        }
    };
    // exchange / relax indices as buildGasNet lays them out
    int ex = 0, rx = 0;
    std::vector<int> exInlet(nIn), exIdle(nIn, -1), rxPlenum(nIn), exC(N), exP(N), exVI(N), exVE(N), rxIntr(N), rxExhr(N), exOutlet(nEx), rxCollector(nEx);
    for (int i = 0; i < nIn; ++i) { exInlet[i] = ex++; if (P.intakes[i].idleFlowK != 0) exIdle[i] = ex++; rxPlenum[i] = rx++; }
    for (int c = 0; c < N; ++c) { exC[c] = ex++; exP[c] = ex++; exVI[c] = ex++; exVE[c] = ex++; rxIntr[c] = rx++; rxExhr[c] = rx++; }
    for (int e = 0; e < nEx; ++e) { exOutlet[e] = ex++; rxCollector[e] = rx++; }
    if (ex != g.nExchange || rx != g.nRelax) return "operation tables do not match the engine";
    long long wantOps = 0;
    for (int s = 0; s < D.fluidSteps; ++s) {
        for (int i = 0; i < nIn; ++i) {
            touch(want, ESB_SLOT_EXCHANGE, exInlet[i], s); ++wantOps;
            if (exIdle[i] >= 0) { touch(want, ESB_SLOT_EXCHANGE, exIdle[i], s); ++wantOps; }
            touch(want, ESB_SLOT_RELAX, rxPlenum[i], s); ++wantOps;
        }
        for (int c = 0; c < N; ++c) {
            touch(want, ESB_SLOT_EXCHANGE, exC[c], s); touch(want, ESB_SLOT_EXCHANGE, exP[c], s);
            touch(want, ESB_SLOT_EXCHANGE, exVI[c], s); touch(want, ESB_SLOT_EXCHANGE, exVE[c], s);
            touch(want, ESB_SLOT_RELAX, rxIntr[c], s); touch(want, ESB_SLOT_RELAX, rxExhr[c], s);
            wantOps += 6;
        }
        for (int e = 0; e < nEx; ++e) { touch(want, ESB_SLOT_EXCHANGE, exOutlet[e], s); touch(want, ESB_SLOT_RELAX, rxCollector[e], s); wantOps += 2; }
    }
    // the schedule: which substep an entry belongs to is fixed by how often its operation has appeared before
    std::vector<int> seenEx(g.nExchange, 0), seenRx(g.nRelax, 0);
    long long gotOps = 0;
    int blowbySlots = 0;
    for (int sl = 0; sl < g.nSlots; ++sl) {
        const int type = g.slotType[sl];
        const uint32_t *row = &g.slotOps[(size_t)sl * 16];
        std::vector<int> cellsInSlot;
        if (type == ESB_SLOT_BLOWBY) { ++blowbySlots; continue; }
        const int lanes = (type == ESB_SLOT_EXCHANGE && g.split) ? G / 2 : G;
        for (int l = 0; l < lanes; ++l) {
            const int kind = (int)(row[l] & 0xffu);
            if (kind == ESB_GAS_IDLE) continue;
            if (type == ESB_SLOT_EXCHANGE) {
                if (kind >= g.nExchange) return "slot " + std::to_string(sl) + ": exchange index out of range";
                if (g.split && (row[l + G / 2] & 0xffu) != (uint32_t)kind) return "slot " + std::to_string(sl) + ": the partner lane has a different operation";
                const int sub = seenEx[kind]++;
                const int flowCyl = g.exchange[kind].flowCylinder;
                const uint32_t wantFlow = flowCyl >= 0 ? (uint32_t)(sub * N + flowCyl + 1) : 0u;
                if ((row[l] >> 16) != wantFlow) return "slot " + std::to_string(sl) + ": flow record index";
                touch(got, type, kind, sub);
                const EsbGasExchange &e = g.exchange[kind];
                if (e.cell0 < g.nCells) cellsInSlot.push_back(e.cell0);
                if (e.cell1 < g.nCells) cellsInSlot.push_back(e.cell1);
            } else {
                if (kind >= g.nRelax) return "slot " + std::to_string(sl) + ": relax index out of range";
                touch(got, type, kind, seenRx[kind]++);
                cellsInSlot.push_back(g.relax[kind].cell);
            }
            ++gotOps;
        }
        std::sort(cellsInSlot.begin(), cellsInSlot.end());
        if (std::adjacent_find(cellsInSlot.begin(), cellsInSlot.end()) != cellsInSlot.end()) return "slot " + std::to_string(sl) + ": two operations share a cell";
    }
    if (gotOps != wantOps) return "the schedule holds " + std::to_string(gotOps) + " operations, the step has " + std::to_string(wantOps);
    bool anyBlowby = false;
    for (int c = 0; c < N; ++c) anyBlowby = anyBlowby || P.cyl[c].blowbyK != 0;
    if (blowbySlots != (anyBlowby ? D.fluidSteps : 0)) return "blow-by slots";
    for (int c = 0; c < g.nCells; ++c) {
        if (want[c].size() != got[c].size()) return "cell " + std::to_string(c) + ": operation count";
        for (size_t j = 0; j < want[c].size(); ++j)
            if (want[c][j].type != got[c][j].type || want[c][j].kind != got[c][j].kind || want[c][j].sub != got[c][j].sub)
                return "cell " + std::to_string(c) + ": operation " + std::to_string(j) + " out of the reference's order";
    }
    return std::string();
}

} // namespace esb
