// preset.cpp — *.preset.json -> EsbEngineParams (host side).
//
// B200-native counterpart of the reference's Piranha-free loader:
//   engine-sim-bridge/src/preset/PresetDeserializer.cpp and the per-part
//   deserializers (EngineDeserializer.cpp:26-345, CylinderHeadDeserializer.cpp,
//   CamshaftDeserializer.cpp, FunctionDeserializer.cpp, ...).  Schema written by
//   engine-sim-bridge/src/common/PresetSerializer.cpp:35-385.
// Instead of building an object graph of parts it fills one flat, pointer-free
// struct that is uploaded to HBM once per engine type.
#include "host.h"
#include "json.h"

#include <cstring>
#include <fstream>
#include <sstream>

namespace esb {

namespace {

// Function tables are pooled and de-duplicated by content: both banks of a V8
// carry identical lobe profiles / port-flow tables (4 x 511 samples otherwise).
int addFunction(EsbEngineParams &p, int &poolUsed, const Json &fj, const std::string &ctx) {
    if (!fj.isObject()) throw JsonError("Expected function object in " + ctx);
    const Json &samples = fj.at("samples", ctx);
    if (!samples.isArray() || samples.size() == 0)
        throw JsonError("Function 'samples' must be a non-empty array in " + ctx);
    EsbFunction f;
    f.filterRadius = fj.number("filterRadius", ctx);
    f.inputScale = fj.numberOr("inputScale", 1.0);
    f.outputScale = fj.numberOr("outputScale", 1.0);
    std::vector<double> xs, ys;
    for (size_t i = 0; i < samples.size(); ++i) {
        const Json &pt = samples[i];
        if (pt.isArray() && pt.size() >= 2) { xs.push_back(pt[0].num); ys.push_back(pt[1].num); }
    }
    // Function::addSample keeps samples ordered by x (insertion sort upstream)
    for (size_t i = 1; i < xs.size(); ++i)
        for (size_t j = i; j > 0 && xs[j - 1] > xs[j]; --j) { std::swap(xs[j - 1], xs[j]); std::swap(ys[j - 1], ys[j]); }
    f.count = (int)xs.size();
    for (int k = 0; k < p.functionCount; ++k) {
        const EsbFunction &g = p.fn[k];
        if (g.count == f.count && g.filterRadius == f.filterRadius && g.inputScale == f.inputScale && g.outputScale == f.outputScale &&
            std::memcmp(&p.fn_x[g.offset], xs.data(), sizeof(double) * xs.size()) == 0 &&
            std::memcmp(&p.fn_y[g.offset], ys.data(), sizeof(double) * ys.size()) == 0)
            return k;
    }
    if (p.functionCount >= ESB_MAX_FUNCS) throw JsonError("Too many function tables in preset (" + ctx + ")");
    if (poolUsed + f.count > ESB_FN_POOL) throw JsonError("Function sample pool exhausted (" + ctx + ")");
    f.offset = poolUsed;
    std::memcpy(&p.fn_x[poolUsed], xs.data(), sizeof(double) * xs.size());
    std::memcpy(&p.fn_y[poolUsed], ys.data(), sizeof(double) * ys.size());
    poolUsed += f.count;
    p.fn[p.functionCount] = f;
    return p.functionCount++;
}

// PathNormalizer::normalizeImpulseResponsePath: keep the part from "sound-library/".
std::string normalizeIrPath(const std::string &raw) {
    size_t k = raw.find("sound-library/");
    return k == std::string::npos ? raw : raw.substr(k);
}

} // namespace

void parsePreset(const std::string &text, LoadedPreset &out) {
    Json root = Json::parse(text);
    if (!root.isObject()) throw JsonError("Preset root must be an object");
    const Json &e = root.at("engine", "preset");
    EsbEngineParams &p = out.params;
    std::memset(&p, 0, sizeof(p));
    int pool = 0;
    const std::string ctx = "engine";

    out.name = e.stringOr("name", "");
    p.cylinderBankCount = (int)e.number("cylinderBankCount", ctx);
    p.cylinderCount = (int)e.number("cylinderCount", ctx);
    int crankCount = (int)e.number("crankshaftCount", ctx);
    p.exhaustSystemCount = (int)e.number("exhaustSystemCount", ctx);
    p.intakeCount = (int)e.number("intakeCount", ctx);
    if (p.cylinderCount <= 0 || p.cylinderBankCount <= 0 || crankCount <= 0)
        throw JsonError("Invalid engine configuration in " + ctx);
    if (crankCount != 1) throw JsonError("Only single-crankshaft engines are supported (crankshaftCount=" + std::to_string(crankCount) + ")");
    if (p.cylinderCount > ESB_MAX_CYL) throw JsonError("Too many cylinders (max 16)");
    if (p.exhaustSystemCount > ESB_MAX_EXHAUST || p.intakeCount > ESB_MAX_INTAKE) throw JsonError("Too many exhaust systems / intakes");
    if (p.exhaustSystemCount <= 0 || p.intakeCount <= 0) throw JsonError("Engine needs at least one intake and one exhaust system");

    p.starterTorque = e.number("starterTorque", ctx);
    p.starterSpeed = e.number("starterSpeed", ctx);
    p.redline = e.number("redline", ctx);
    p.dynoMinSpeed = e.number("dynoMinSpeed", ctx);
    p.dynoMaxSpeed = e.number("dynoMaxSpeed", ctx);
    p.dynoHoldStep = e.number("dynoHoldStep", ctx);
    p.simulationFrequency = e.number("simulationFrequency", ctx);
    p.initialHighFrequencyGain = e.number("initialHighFrequencyGain", ctx);
    p.initialNoise = e.number("initialNoise", ctx);
    p.initialJitter = e.number("initialJitter", ctx);
    p.throttleGamma = e.number("throttleGamma", ctx);

    // crankshaft 0
    const Json &cranks = e.at("crankshafts", ctx);
    if (!cranks.isArray() || cranks.size() == 0) throw JsonError("Missing required field 'crankshafts' in " + ctx);
    {
        const Json &c = cranks[0]; const std::string cc = ctx + ".crankshafts[0]";
        p.crankMass = c.number("mass", cc);
        p.flywheelMass = c.number("flywheelMass", cc);
        p.crankMomentOfInertia = c.number("momentOfInertia", cc);
        p.crankThrow = c.number("crankThrow", cc);
        p.crankPosX = c.number("posX", cc);
        p.crankPosY = c.number("posY", cc);
        p.tdc = c.number("tdc", cc);
        p.frictionTorque = c.number("frictionTorque", cc);
        const Json &js = c.at("rodJournals", cc);
        p.journalCount = (int)js.size();
        if (p.journalCount > ESB_MAX_JOURNALS) throw JsonError("Too many rod journals");
        for (int i = 0; i < p.journalCount; ++i) p.journalAngle[i] = js[i].number("angle", cc);
    }

    const Json &exh = e.at("exhaustSystems", ctx);
    out.irFiles.assign(p.exhaustSystemCount, "");
    for (int i = 0; i < p.exhaustSystemCount; ++i) {
        const Json &x = exh[i]; const std::string xc = ctx + ".exhaustSystems[" + std::to_string(i) + "]";
        EsbExhaustSystem &d = p.exhaust[i];
        d.length = x.number("length", xc);
        d.collectorCrossSectionArea = x.number("collectorCrossSectionArea", xc);
        d.outletFlowRate = x.number("outletFlowRate", xc);
        d.primaryTubeLength = x.number("primaryTubeLength", xc);
        d.primaryFlowRate = x.number("primaryFlowRate", xc);
        d.velocityDecay = x.number("velocityDecay", xc);
        d.audioVolume = x.number("audioVolume", xc);
        if (x.has("impulseResponseFilename")) {
            out.irFiles[i] = normalizeIrPath(x.stringOr("impulseResponseFilename", ""));
            d.impulseResponseVolume = x.number("impulseResponseVolume", xc);
        }
    }
    const Json &ins = e.at("intakes", ctx);
    for (int i = 0; i < p.intakeCount; ++i) {
        const Json &x = ins[i]; const std::string xc = ctx + ".intakes[" + std::to_string(i) + "]";
        EsbIntake &d = p.intakes[i];
        d.runnerFlowRate = x.number("runnerFlowRate", xc);
        d.runnerLength = x.number("runnerLength", xc);
        d.plenumCrossSectionArea = x.number("plenumCrossSectionArea", xc);
        d.velocityDecay = x.number("velocityDecay", xc);
        d.inputFlowK = x.number("inputFlowK", xc);
        d.plenumVolume = x.number("plenumVolume", xc);
        d.idleFlowK = x.number("idleFlowK", xc);
        d.idleThrottlePlatePosition = x.number("idleThrottlePlatePosition", xc);
        d.molecularAfr = x.number("molecularAfr", xc);
    }

    // banks -> flattened cylinders (EngineDeserializer.cpp:201-238)
    const Json &banks = e.at("cylinderBanks", ctx);
    int g = 0;
    for (size_t bi = 0; bi < banks.size() && (int)bi < p.cylinderBankCount; ++bi) {
        const Json &b = banks[bi]; const std::string bc = ctx + ".cylinderBanks[" + std::to_string(bi) + "]";
        int cylCount = (int)b.number("cylinderCount", bc);
        const Json &cyls = b.at("cylinders", bc);
        const Json &head = b.at("cylinderHead", bc); const std::string hc = bc + ".cylinderHead";
        const Json &hcyl = head.at("cylinders", hc);
        const Json &icam = head.at("intakeCamshaft", hc), &ecam = head.at("exhaustCamshaft", hc);
        const Json &ilobes = icam.at("lobeCenterlines", hc), &elobes = ecam.at("lobeCenterlines", hc);
        if (ilobes.size() == 0 || elobes.size() == 0) throw JsonError("Missing or empty 'lobeCenterlines' in " + hc);
        int fnIL = addFunction(p, pool, icam.at("lobeProfile", hc), hc + ".intakeCamshaft.lobeProfile");
        int fnEL = addFunction(p, pool, ecam.at("lobeProfile", hc), hc + ".exhaustCamshaft.lobeProfile");
        int fnIP = addFunction(p, pool, head.at("intakePortFlow", hc), hc + ".intakePortFlow");
        int fnEP = addFunction(p, pool, head.at("exhaustPortFlow", hc), hc + ".exhaustPortFlow");
        for (int ci = 0; ci < cylCount && ci < (int)cyls.size(); ++ci, ++g) {
            if (g >= p.cylinderCount) break;
            const Json &c = cyls[ci]; const std::string cc = bc + ".cylinders[" + std::to_string(ci) + "]";
            EsbCylinder &d = p.cyl[g];
            d.bankAngle = b.number("angle", bc);
            d.bankX = b.number("positionX", bc);
            d.bankY = b.number("positionY", bc);
            d.bore = b.number("bore", bc);
            d.deckHeight = b.number("deckHeight", bc);
            d.bankIndex = (int)bi; d.indexInBank = ci;
            d.pistonMass = c.number("pistonMass", cc);
            d.compressionHeight = c.number("compressionHeight", cc);
            d.wristPinPosition = c.number("wristPinPosition", cc);
            d.blowbyK = c.number("blowbyK", cc);
            c.number("displacement", cc);
            const Json &r = c.at("connectingRod", cc); const std::string rc = cc + ".connectingRod";
            d.rodMass = r.number("mass", rc);
            d.rodMomentOfInertia = r.number("momentOfInertia", rc);
            d.rodCenterOfMass = r.number("centerOfMass", rc);
            d.rodLength = r.number("length", rc);
            d.slaveThrow = r.numberOr("slaveThrow", 0.0);
            // The reference's schema has no master link (PresetSerializer.cpp:59-67 writes "journal" and "slaveThrow" only);
            // presets compiled by this repo's script front-end add "masterCylinder" / "slaveJournalAngle" to the rod.
            d.masterCylinder = (int)r.numberOr("masterCylinder", -1.0);
            d.slaveJournalAngle = r.numberOr("slaveJournalAngle", 0.0);
            if (d.masterCylinder >= g) throw JsonError("masterCylinder must name an earlier cylinder in " + rc);
            if (d.masterCylinder >= 0 && p.cyl[d.masterCylinder].masterCylinder >= 0) throw JsonError("a slave rod cannot be a master in " + rc);
            d.journal = c.has("rodJournalIndex") ? (int)c.number("rodJournalIndex", cc) : g;
            if (d.masterCylinder >= 0) d.journal = 0;
            if (d.journal < 0 || d.journal >= p.journalCount) throw JsonError("rodJournalIndex out of range in " + cc);
            d.combustionChamberVolume = head.number("combustionChamberVolume", hc);
            d.intakeRunnerVolume = head.number("intakeRunnerVolume", hc);
            d.intakeRunnerCrossSectionArea = head.number("intakeRunnerCrossSectionArea", hc);
            d.exhaustRunnerVolume = head.number("exhaustRunnerVolume", hc);
            d.exhaustRunnerCrossSectionArea = head.number("exhaustRunnerCrossSectionArea", hc);
            const Json &hcj = hcyl[ci];
            d.soundAttenuation = hcj.number("soundAttenuation", hc);
            d.headerPrimaryLength = hcj.number("headerPrimaryLength", hc);
            d.intakeLobeCenterline = ilobes[ci % ilobes.size()].num;
            d.exhaustLobeCenterline = elobes[ci % elobes.size()].num;
            d.intakeCamAdvance = icam.number("advance", hc);
            d.exhaustCamAdvance = ecam.number("advance", hc);
            d.fnIntakeLobe = fnIL; d.fnExhaustLobe = fnEL; d.fnIntakePortFlow = fnIP; d.fnExhaustPortFlow = fnEP;
            // CylinderHeadDeserializer.cpp:86-99: bank i -> intake/exhaust i when it exists, else 0
            d.intake = (p.intakeCount > (int)bi) ? (int)bi : 0;
            d.exhaustSystem = (p.exhaustSystemCount > (int)bi) ? (int)bi : 0;
            // Documents written by the .mr front-end (mr_frontend.py) carry the script's own plumbing per cylinder
            // (add_cylinder(exhaust_system: ..., intake: ...), es/actions/actions.mr): e.g. an inline four with two collectors.
            if (c.has("exhaustSystemIndex")) d.exhaustSystem = (int)c.number("exhaustSystemIndex", cc);
            if (c.has("intakeIndex")) d.intake = (int)c.number("intakeIndex", cc);
            if (d.exhaustSystem < 0 || d.exhaustSystem >= p.exhaustSystemCount) throw JsonError("exhaustSystemIndex out of range in " + cc);
            if (d.intake < 0 || d.intake >= p.intakeCount) throw JsonError("intakeIndex out of range in " + cc);
        }
    }
    if (g != p.cylinderCount) throw JsonError("cylinderCount does not match the cylinders listed in cylinderBanks");

    if (e.has("fuel")) {
        const Json &f = e.at("fuel", ctx); const std::string fc = ctx + ".fuel";
        p.maxBurningEfficiency = f.number("maxBurningEfficiency", fc);
        p.burningEfficiencyRandomness = f.number("burningEfficiencyRandomness", fc);
        p.lowEfficiencyAttenuation = f.number("lowEfficiencyAttenuation", fc);
        p.maxTurbulenceEffect = f.number("maxTurbulenceEffect", fc);
        p.maxDilutionEffect = f.number("maxDilutionEffect", fc);
        p.fuelMolecularAfr = f.number("molecularAfr", fc);
        p.fuelMolecularMass = f.number("molecularMass", fc);
        p.fuelEnergyDensity = f.number("energyDensity", fc);
        p.fuelDensity = f.number("density", fc);
        p.fnTurbulenceToFlameSpeedRatio = addFunction(p, pool, f.at("turbulenceToFlameSpeedRatio", fc), fc + ".turbulenceToFlameSpeedRatio");
    } else throw JsonError("Missing required field 'fuel' in " + ctx);

    {
        const Json &ig = e.at("ignitionModule", ctx); const std::string ic = ctx + ".ignitionModule";
        p.revLimit = ig.number("revLimit", ic);
        p.limiterDuration = ig.number("limiterDuration", ic);
        p.fnTimingCurve = addFunction(p, pool, ig.at("timingCurve", ic), ic + ".timingCurve");
        const Json &fo = ig.at("firingOrder", ic);
        for (int i = 0; i < p.cylinderCount && i < (int)fo.size(); ++i) p.cyl[i].firingAngle = fo[i].num;
    }

    const Json &ch = e.at("combustionChambers", ctx);
    if (!ch.isArray()) throw JsonError("Missing required field 'combustionChambers' in " + ctx);
    for (int i = 0; i < p.cylinderCount && i < (int)ch.size(); ++i) {
        const Json &c = ch[i]; const std::string cc = ctx + ".combustionChambers[" + std::to_string(i) + "]";
        EsbCylinder &d = p.cyl[i];
        d.fnMeanPistonSpeedToTurbulence = addFunction(p, pool, c.at("meanPistonSpeedToTurbulence", cc), cc + ".meanPistonSpeedToTurbulence");
        d.crankcasePressure = c.number("crankcasePressure", cc);
        d.startingPressure = c.number("startingPressure", cc);
        d.startingTemperature = c.number("startingTemperature", cc);
        d.frictionCoeff = c.number("frictionCoeff", cc);
        d.breakawayFriction = c.number("breakawayFriction", cc);
        d.breakawayFrictionVelocity = c.number("breakawayFrictionVelocity", cc);
        d.viscousFrictionCoefficient = c.number("viscousFrictionCoefficient", cc);
    }

    if (root.has("vehicle")) {
        const Json &v = root.at("vehicle", "preset");
        p.vehicleMass = v.number("mass", "vehicle");
        p.dragCoefficient = v.number("dragCoefficient", "vehicle");
        p.crossSectionArea = v.number("crossSectionArea", "vehicle");
        p.diffRatio = v.number("diffRatio", "vehicle");
        p.tireRadius = v.number("tireRadius", "vehicle");
        p.rollingResistance = v.number("rollingResistance", "vehicle");
    } else { p.vehicleMass = 1000; p.diffRatio = 3.42; p.tireRadius = 0.254; p.dragCoefficient = 0.3; p.crossSectionArea = 2.0; }
    if (root.has("transmission")) {
        const Json &t = root.at("transmission", "preset");
        const Json &gr = t.at("gearRatios", "transmission");
        p.gearCount = (int)gr.size();
        if (p.gearCount > ESB_MAX_GEARS) throw JsonError("Too many gears");
        for (int i = 0; i < p.gearCount; ++i) p.gearRatios[i] = gr[i].num;
        p.currentGear = (int)t.numberOr("currentGear", -1);
        p.maxClutchTorque = t.number("maxClutchTorque", "transmission");
        p.clutchPressure = t.numberOr("clutchPressure", 0.0);
    } else { p.gearCount = 0; p.currentGear = -1; }
}

void loadEngineFile(const std::string &path, LoadedPreset &out) {
    if (path.size() >= 5 && path.compare(path.size() - 5, 5, ".json") == 0) { loadPresetFile(path, out); return; }
    parsePreset(compileScriptToPresetJson(path, {}), out);
}

void loadPresetFile(const std::string &path, LoadedPreset &out) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw JsonError("Cannot open preset file: " + path);
    std::stringstream ss; ss << f.rdbuf();
    parsePreset(ss.str(), out);
}

} // namespace esb
