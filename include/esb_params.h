/* esb_params.h — flat, SI-unit, pointer-free description of one engine type.
 *
 * This is the parameter contract between the preset loader (host), the CUDA
 * kernels (device, read-only, shared by every instance of that engine type)
 * and the CPU oracle.  Field names follow the reference's preset JSON schema
 * (engine-sim-bridge/src/common/PresetSerializer.cpp:35-385) and its
 * per-part deserializers (engine-sim-bridge/src/preset/*Deserializer.cpp).
 * Cylinders are flattened bank-major exactly as the reference numbers them
 * (EngineDeserializer.cpp:142-149, globalCylIdx + ci).
 *
 * Plain C, no pointers: the whole struct is copied to HBM as one blob.
 */
#ifndef ESB_PARAMS_H
#define ESB_PARAMS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ESB_MAX_CYL 16
#define ESB_MAX_EXHAUST 4
#define ESB_MAX_INTAKE 4
#define ESB_MAX_JOURNALS 16
#define ESB_MAX_GEARS 12
#define ESB_MAX_FUNCS 48
#define ESB_FN_POOL 6144
#define ESB_MAX_IR_TAPS 98304 /* longest in-tree IR is 96,392 frames */

/* A `Function` lookup table (FunctionDeserializer.cpp: n samples, filterRadius,
 * input/output scale).  Samples live in EsbEngineParams::fn_x / fn_y. */
typedef struct EsbFunction {
    int32_t offset; /* first sample in the pool */
    int32_t count;  /* number of samples */
    double filterRadius;
    double inputScale;
    double outputScale;
} EsbFunction;

typedef struct EsbCylinder {
    /* bank (CylinderBankDeserializer.cpp) */
    double bankAngle, bankX, bankY, bore, deckHeight;
    int32_t bankIndex, indexInBank;
    /* piston (EngineDeserializer.cpp:142-199) */
    double pistonMass, compressionHeight, wristPinPosition, blowbyK;
    /* connecting rod (ConnectingRodDeserializer.cpp) */
    double rodMass, rodMomentOfInertia, rodCenterOfMass, rodLength;
    int32_t journal; /* rod journal on crankshaft 0 (a rod on the crankshaft) */
    /* master / slave rods of radial engines (es/engines/atg-video-2/09_radial_9.mr:145-165; ConnectingRod "slaveThrow",
     * PresetSerializer.cpp:65): a slave rod's big end rides on a journal of its master rod instead of the crankshaft.
     * masterCylinder = -1 for a rod on the crankshaft, else the cylinder whose rod carries this rod's journal (it must
     * come earlier in cylinder order); slaveJournalAngle is that journal's angle on the master (same convention as a
     * bank angle: from the master rod's own axis, counter-clockwise), at radius slaveThrow OF THE MASTER from its big end. */
    int32_t masterCylinder;
    double slaveThrow, slaveJournalAngle;
    /* head (CylinderHeadDeserializer.cpp:126-154) */
    double combustionChamberVolume;
    double intakeRunnerVolume, intakeRunnerCrossSectionArea;
    double exhaustRunnerVolume, exhaustRunnerCrossSectionArea;
    double soundAttenuation, headerPrimaryLength;
    double intakeLobeCenterline, exhaustLobeCenterline; /* crank radians as stored in JSON */
    double intakeCamAdvance, exhaustCamAdvance;
    int32_t fnIntakeLobe, fnExhaustLobe, fnIntakePortFlow, fnExhaustPortFlow;
    int32_t exhaustSystem, intake; /* CylinderHeadDeserializer.cpp:86-99 */
    /* combustion chamber (EngineDeserializer.cpp:240-306) */
    double crankcasePressure, startingPressure, startingTemperature;
    double frictionCoeff, breakawayFriction, breakawayFrictionVelocity, viscousFrictionCoefficient;
    int32_t fnMeanPistonSpeedToTurbulence;
    int32_t pad1;
    double firingAngle; /* IgnitionModuleDeserializer.cpp: firingOrder[i], cycle radians */
} EsbCylinder;

typedef struct EsbExhaustSystem {
    double length, collectorCrossSectionArea, outletFlowRate, primaryTubeLength;
    double primaryFlowRate, velocityDecay, audioVolume, impulseResponseVolume;
} EsbExhaustSystem;

typedef struct EsbIntake {
    double runnerFlowRate, runnerLength, plenumCrossSectionArea, velocityDecay;
    double inputFlowK, plenumVolume, idleFlowK, idleThrottlePlatePosition, molecularAfr;
} EsbIntake;

typedef struct EsbEngineParams {
    int32_t cylinderCount, cylinderBankCount, exhaustSystemCount, intakeCount;
    int32_t journalCount, gearCount, currentGear, functionCount;
    /* engine (EngineDeserializer.cpp:26-87) */
    double starterTorque, starterSpeed, redline;
    double dynoMinSpeed, dynoMaxSpeed, dynoHoldStep;
    double simulationFrequency;
    double initialHighFrequencyGain, initialNoise, initialJitter;
    double throttleGamma;
    /* crankshaft 0 (CrankshaftDeserializer.cpp) */
    double crankMass, flywheelMass, crankMomentOfInertia, crankThrow;
    double crankPosX, crankPosY, tdc, frictionTorque;
    double journalAngle[ESB_MAX_JOURNALS];
    /* fuel (FuelDeserializer.cpp) */
    double maxBurningEfficiency, burningEfficiencyRandomness, lowEfficiencyAttenuation;
    double maxTurbulenceEffect, maxDilutionEffect;
    double fuelMolecularAfr, fuelMolecularMass, fuelEnergyDensity, fuelDensity;
    int32_t fnTurbulenceToFlameSpeedRatio, fnTimingCurve;
    /* ignition (IgnitionModuleDeserializer.cpp) */
    double revLimit, limiterDuration;
    /* vehicle / transmission (VehicleDeserializer.cpp, TransmissionDeserializer.cpp) */
    double vehicleMass, dragCoefficient, crossSectionArea, diffRatio, tireRadius, rollingResistance;
    double maxClutchTorque, clutchPressure;
    double gearRatios[ESB_MAX_GEARS];
    EsbCylinder cyl[ESB_MAX_CYL];
    EsbExhaustSystem exhaust[ESB_MAX_EXHAUST];
    EsbIntake intakes[ESB_MAX_INTAKE];
    EsbFunction fn[ESB_MAX_FUNCS];
    double fn_x[ESB_FN_POOL];
    double fn_y[ESB_FN_POOL];
} EsbEngineParams;

/* Run-time configuration of a simulator (reference: ISimulatorConfig,
 * engine-sim-bridge/include/simulator/EngineSimTypes.h:97-105; the historical
 * C struct EngineSimConfig, archive/TA1_FINDINGS.md:237-266, carried the same
 * values plus inputBufferSize/audioBufferSize/airNoise). */
typedef struct EngineSimConfig {
    int32_t sampleRate;           /* 44100 */
    int32_t inputBufferSize;      /* 44100 (upstream simulator.cpp:226-231) */
    int32_t audioBufferSize;      /* 44100 */
    int32_t simulationFrequency;  /* 0 = engine decides (CLIMain.cpp:240-246) */
    int32_t fluidSimulationSteps; /* 8 */
    int32_t maxChunkFrames;       /* 4096 */
    double targetSynthesizerLatency; /* 0.02 */
    float volume;                 /* 0.5 */
    float convolutionLevel;       /* 0.5 */
    float airNoise;               /* <0 = engine decides (initialNoise) */
    uint32_t seed;                /* noise seed of this instance (batch: seed + instance) */
} EngineSimConfig;

/* Telemetry snapshot; field-for-field the reference's EngineSimStats
 * (EngineSimTypes.h:108-133) flattened to C. */
typedef struct EsbSimStats {
    double currentRPM;
    double currentLoad;
    double exhaustFlow;
    double manifoldPressure;
    int32_t activeChannels;
    int32_t gear;            /* bridge convention: 0 = neutral (GearConventions.h) */
    double processingTimeMs;
    double dynoTorque;       /* ft*lbs, as the reference reports it */
    double dynoTargetRPM;
    double dynoTorqueScale;
    double vehicleSpeedKmh;
    double engineTorqueNm;
    double drivetrainTorqueNm;
    int32_t gearSelector;
    int32_t gearAutoMode;
} EsbSimStats;
/* The reference's own EngineSimTypes.h (include guard ATG_ENGINE_SIM_BRIDGE_H) defines a C++ struct of this
 * name with a different member order; when both headers meet in one translation unit (include/GpuSimulator.h)
 * only the Esb name exists and the adapter converts field by field. */
#ifndef ATG_ENGINE_SIM_BRIDGE_H
typedef EsbSimStats EngineSimStats;
#endif

#ifdef __cplusplus
}
#endif
#endif
