// host.h — host-side declarations shared by the C-ABI translation units.
#pragma once
#include "esb_params.h"
#include "esb_device.h"

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

namespace esb {

struct LoadedPreset {
    EsbEngineParams params;
    std::string name;
    std::vector<std::string> irFiles; // per exhaust system, normalised to "sound-library/..."
};

// preset.cpp
void parsePreset(const std::string &jsonText, LoadedPreset &out);
void loadPresetFile(const std::string &path, LoadedPreset &out);
// "*.json" -> preset, anything else -> engine script (SimulatorFactory.cpp:111-156 createPistonEngineSimulator)
void loadEngineFile(const std::string &path, LoadedPreset &out);

// mr_script.cpp — engine scripts without Piranha: script -> preset document (JSON text of PresetSerializer's schema)
struct ScriptError : std::runtime_error { using std::runtime_error::runtime_error; };
std::string compileScriptToPresetJson(const std::string &scriptPath, const std::vector<std::string> &searchPaths);

// wav.cpp — 16-bit PCM RIFF reader/writer (replaces dr_wav, wav_loader.h:41-63)
struct WavData {
    int sampleRate = 0;
    int channels = 0;
    std::vector<int16_t> samples; // interleaved
};
bool readWav(const std::string &path, WavData &out, std::string &err);
bool writeWavMono16(const std::string &path, const int16_t *samples, size_t n, int sampleRate, std::string &err);

// instance_init.cpp — derived constants and the initial state column of an instance
int lanesForCylinders(int n, int minLanes = 1);
void deriveDeviceParams(const EsbEngineParams &P, double simFreq, int fluidSteps, int maxIterations, double minDelta, double beta,
                        uint64_t seedBase, EsbDeviceParams &D, int minLanes = 1);
void initialStateColumn(const EsbDeviceParams &D, std::vector<double> &col);
// gas_schedule.cpp — gas-network tables and the slot schedule of one step (called by deriveDeviceParams)
void buildGasNet(const EsbEngineParams &P, EsbDeviceParams &D);
std::string checkGasSchedule(const EsbEngineParams &P, const EsbDeviceParams &D);   // "" = valid

} // namespace esb
