// synth.cpp — oracle restatement of upstream synthesizer.cpp and its filters
// (TEST INFRASTRUCTURE ONLY; see oracle.h).
//
// writeInput / endInputBlock / renderAudio block loop / readAudioOutput follow the code the
// reference quotes at archive/PHYSICS_ANALYSIS.md:39-55,337-383,424-465,291-311.  The
// per-sample chain, filter laws and the impulse-response clip follow SURVEY.md A.9
// [UPSTREAM-RECALL]; parameter names: SimulatorFactory.cpp:63-71, SineSimulator.cpp:43-52.
// Noise comes from esb_math.h's counter-based generator instead of libc rand().
#include "oracle.h"
#include "esb_math.h"

#include <cmath>
#include <cstring>

namespace oracle {

static float lowPassAlpha(double cutoff, double sampleRate) {
    const double dt = 1.0 / sampleRate;
    const double rc = 1.0 / (2.0 * kPi * cutoff);
    return (float)(dt / (rc + dt));
}

void Synth::initialize(int channels_, double inputSampleRate_, uint64_t seed_) {
    channels = channels_;
    inputSampleRate = inputSampleRate_;
    seed = seed_;
    ch.assign(channels, SynthChannel());
    for (auto &c : ch) { c.ir.assign(1, 1.0f); c.shift.assign(1, 0.0f); }
    inputWriteOffset = 0; lastInputSampleOffset = 0; writeIndex = 0; inputSamplesRead = 0;
    masterY = 0; levelerPeak = 0;
    audio.clear(); audioFloat.clear();
}

void Synth::initializeImpulseResponse(const int16_t *s, unsigned n, float volume, int channel, int maxTaps) {
    unsigned clipped = 0;
    for (unsigned i = 0; i < n; ++i) if (std::abs((int)s[i]) > 100) clipped = i + 1;
    unsigned cap = (maxTaps > 0) ? (unsigned)maxTaps : 10000u;
    unsigned L = clipped < cap ? clipped : cap;
    if (L == 0) L = 1;
    SynthChannel &c = ch[channel];
    c.ir.assign(L, 0.0f);
    for (unsigned i = 0; i < L && i < n; ++i) c.ir[i] = volume * (float)s[i] / 32767.0f;
    c.shift.assign(L, 0.0f);
    c.shiftOffset = 0;
}

void Synth::writeInput(const double *data) {
    const double bufferSize = (double)inputBufferSize;
    inputWriteOffset += (double)audioSampleRate / inputSampleRate;
    if (inputWriteOffset >= bufferSize) inputWriteOffset -= bufferSize;
    auto inputDistance = [&](double s1, double s0) { return (s1 < s0) ? (bufferSize - s0) + s1 : s1 - s0; };
    const float aaAlpha = lowPassAlpha(1900.0, audioSampleRate);
    const double distance = inputDistance(inputWriteOffset, lastInputSampleOffset);
    long written = 0;
    for (int i = 0; i < channels; ++i) {
        SynthChannel &c = ch[i];
        const double last = c.lastInputSample;
        double s = inputDistance((double)writeIndex, lastInputSampleOffset);
        written = 0;
        for (; s <= distance; s += 1.0) {
            if (s >= bufferSize) s -= bufferSize;
            const double f = s / distance;
            const double sample = last * (1.0 - f) + data[i] * f;
            c.antialiasY = aaAlpha * (float)sample + (1.0f - aaAlpha) * c.antialiasY;
            c.input.push_back(c.antialiasY);
            ++written;
        }
        c.lastInputSample = data[i];
    }
    writeIndex = (writeIndex + written) % inputBufferSize;
    lastInputSampleOffset = inputWriteOffset;
}

void Synth::endInputBlock() {
    for (auto &c : ch) c.input.erase(c.input.begin(), c.input.begin() + inputSamplesRead);
    inputSamplesRead = 0;
}

int Synth::renderAudio() {
    if (channels == 0) return 0;
    const int room = 2000 - (int)audio.size();
    int n = room > 0 ? room : 0;
    // samples already consumed by a previous render in this frame stay in the ring until endInputBlock
    const int available = (int)ch[0].input.size() - inputSamplesRead;
    if (available < n) n = available;
    if (n <= 0) return 0;
    const float jitterAlpha = lowPassAlpha(ap.inputSampleNoiseFrequencyCutoff, audioSampleRate);
    const float airAlpha = lowPassAlpha(ap.airNoiseFrequencyCutoff, audioSampleRate);
    const float dcAlpha = lowPassAlpha(10.0, audioSampleRate);
    const float masterAlpha = lowPassAlpha(1900.0, audioSampleRate);
    const float invDt = (float)audioSampleRate;
    for (int k = 0; k < n; ++k) {
        float signal = 0.0f;
        for (int i = 0; i < channels; ++i) {
            SynthChannel &c = ch[i];
            const float x = c.input[inputSamplesRead + k];
            // JitterFilter::fast_f
            c.jitterHistory[c.jitterOffset] = x;
            c.jitterOffset = (c.jitterOffset + 1 == 10) ? 0 : c.jitterOffset + 1;
            const float u = esb_rand_u01f(seed, ESB_STREAM_JITTER + i, c.sampleCounter);
            c.jitterNoiseY = jitterAlpha * u + (1.0f - jitterAlpha) * c.jitterNoiseY;
            int d = (int)(c.jitterNoiseY * ap.inputSampleNoise * 9.0f + 0.5f);
            if (d < 0) d = 0;
            if (d > 9) d = 9;
            int idx = c.jitterOffset - 1 - d;
            if (idx < 0) idx += 10;
            const float f_in = c.jitterHistory[idx];
            // DC removal, derivative
            c.dcY = dcAlpha * f_in + (1.0f - dcAlpha) * c.dcY;
            const float f = f_in - c.dcY;
            const float f_p = (f_in - c.derivPrev) * invDt;
            c.derivPrev = f_in;
            // air noise
            const float noise = 2.0f * esb_rand_u01f(seed, ESB_STREAM_AIR + i, c.sampleCounter) - 1.0f;
            c.airY = airAlpha * noise + (1.0f - airAlpha) * c.airY;
            const float r_mixed = ap.airNoise * c.airY + (1.0f - ap.airNoise);
            c.sampleCounter += 1;
            float v_in = f_p * ap.dF_F_mix + (f * r_mixed) * (1.0f - ap.dF_F_mix);
            if (std::fabs(v_in) < 1.17549435e-38f) v_in = 0.0f;
            // ConvolutionFilter::f — circular shift register, direct form
            const int L = (int)c.ir.size();
            c.shift[c.shiftOffset] = v_in;
            float y;
            if (firMode == 0) {
                float acc = 0.0f;
                for (int t = 0; t < L; ++t) { int p = c.shiftOffset + t; if (p >= L) p -= L; acc += c.ir[t] * c.shift[p]; }
                y = acc;
            } else {
                double acc = 0.0;
                for (int t = 0; t < L; ++t) { int p = c.shiftOffset + t; if (p >= L) p -= L; acc += (double)c.ir[t] * (double)c.shift[p]; }
                y = (float)acc;
            }
            c.shiftOffset = (c.shiftOffset == 0) ? L - 1 : c.shiftOffset - 1;
            const float v = ap.convolution * y + (1.0f - ap.convolution) * v_in;
            signal += v;
        }
        masterY = masterAlpha * signal + (1.0f - masterAlpha) * masterY;
        // LevelingFilter::f
        levelerPeak = 0.9999f * levelerPeak;
        const float a = std::fabs(masterY);
        if (a > levelerPeak) levelerPeak = a;
        float out = 0.0f;
        if (levelerPeak != 0.0f) {
            float gain = ap.levelerTarget / levelerPeak;
            if (gain < ap.levelerMinGain) gain = ap.levelerMinGain;
            if (gain > ap.levelerMaxGain) gain = ap.levelerMaxGain;
            out = masterY * gain;
        }
        out *= ap.volume;
        long r = std::lround(out);
        if (r > 32767) r = 32767;
        if (r < -32768) r = -32768;
        audio.push_back((int16_t)r);
        audioFloat.push_back(out);
    }
    inputSamplesRead += n;
    return n;
}

int Synth::readAudioOutput(int samples, int16_t *out, float *outFloat) {
    const int have = (int)audio.size();
    const int n = have < samples ? have : samples;
    if (out) {
        std::memcpy(out, audio.data(), sizeof(int16_t) * n);
        if (n < samples) std::memset(out + n, 0, sizeof(int16_t) * (samples - n));
    }
    if (outFloat) {
        std::memcpy(outFloat, audioFloat.data(), sizeof(float) * n);
        if (n < samples) std::memset(outFloat + n, 0, sizeof(float) * (samples - n));
    }
    audio.erase(audio.begin(), audio.begin() + n);
    audioFloat.erase(audioFloat.begin(), audioFloat.begin() + n);
    return n;
}

// One simulateStep (+ writeToSynthesizer)
void Simulator::stepOnce() {
    if (sine) {
        // SineSimulator::simulateStep_ / writeToSynthesizer (SineSimulator.cpp:84-105);
        // rpm = 800 + 5200 * throttle (SineEngine.h:35-38)
        const double rpm = 800.0 + 5200.0 * engine.ctl.throttle;
        const double frequency = rpm / 6.0;
        sinePhase += (2.0 * kPi) / (simFreq / frequency);
        if (sinePhase >= 2.0 * kPi) sinePhase -= 2.0 * kPi;
        double staging[ESB_MAX_EXHAUST] = {0, 0, 0, 0};
        staging[0] = esb_sin(sinePhase) * 28000.0;
        synth.writeInput(staging);
    } else {
        engine.step();
        synth.writeInput(engine.exhaustOut);
    }
}

// startFrame(dt); steps x { simulateStep }; endFrame()   (BridgeSimulator.cpp:525-534)
void Simulator::advance(int steps) {
    for (int s = 0; s < steps; ++s) stepOnce();
    endFrame();
}

} // namespace oracle
