// synth_kernels.cu — K3 and K5: the Synthesizer chain around the convolution (sm_100a).
//
// Reference path: Synthesizer::writeInput (quoted at archive/PHYSICS_ANALYSIS.md:39-55,337-350),
// Synthesizer::renderAudio block loop (archive/PHYSICS_ANALYSIS.md:424-465) and its per-sample
// chain (JitterFilter, DC low-pass, DerivativeFilter, air-noise low-pass, dF_F mix ->
// ConvolutionFilter -> master low-pass, LevelingFilter, int16), parameters wired
// at SimulatorFactory.cpp:63-71 and SineSimulator.cpp:43-52.
//
// The chain is a causal per-sample recurrence, so it is evaluated eagerly for every input
// sample; how many samples a render/read call RELEASES (the <=2000-per-render rule, zero fill)
// is integer bookkeeping on the host (batch.cpp).  One thread owns one channel row (K3) or one
// instance (K5); a warp owns 32 consecutive instances and transposes 32x32 sample tiles through
// shared memory so that every HBM access is a full 128-byte row segment.
//
// FP32 exactly as the reference computes it (FP64 only inside the resampler), built with
// -fmad=false: bit-identical to the oracle up to the convolution.
#include "kernels.h"
#include "esb_math.h"

#include <cmath>

namespace {

__device__ __forceinline__ double inputDistance(double s1, double s0, double bufferSize) {
    return (s1 < s0) ? (bufferSize - s0) + s1 : s1 - s0;
}

constexpr int TILE = 32;
// staged samples per row: fewer than TILE left over from the last flush + the samples one physics step writes.
// ESB_MAX_SAMPLES_PER_STEP (kernels.h) bounds the latter; Batch::reset() rejects sampleRate / simFreq ratios beyond it.
constexpr int STAGE = TILE + ESB_MAX_SAMPLES_PER_STEP;

// K3 -------------------------------------------------------------------------------------
// grid: (nInstPad / 32, channels); block: 32 threads (one warp); lane = instance within the block
__global__ void __launch_bounds__(32) esb_synth_pre_kernel(EsbPreArgs a) {
    __shared__ float hist[10][TILE];          // JitterFilter history, [slot][lane]
    __shared__ float stage[TILE][STAGE + 1];  // [lane][sample]
    const int lane = threadIdx.x;
    const int ch = blockIdx.y;
    const int instRaw = blockIdx.x * TILE + lane;
    const bool valid = instRaw < a.nInst;
    const int inst = valid ? instRaw : a.nInst - 1;
    const long long row = (long long)ch * a.nInstPad + instRaw;   // padded rows exist in every buffer
    const long long rows = (long long)a.channels * a.nInstPad;
    const long long rowBlock = (long long)ch * a.nInstPad + blockIdx.x * TILE;

    // channel state
    double lastInput = a.chanLast[row];
    float antialiasY = a.chanState[0 * rows + row];
    float jitterNoiseY = a.chanState[1 * rows + row];
    float dcY = a.chanState[2 * rows + row];
    float derivPrev = a.chanState[3 * rows + row];
    float airY = a.chanState[4 * rows + row];
#pragma unroll
    for (int i = 0; i < 10; ++i) hist[i][lane] = a.chanState[(5 + i) * rows + row];

    const EsbSynthParams sp = a.sp;
    const double bufferSize = a.clock.bufferSize;
    double inputWriteOffset = a.clock.inputWriteOffset;
    double lastInputSampleOffset = a.clock.lastInputSampleOffset;
    long long writeIndex = a.clock.writeIndex;
    long long sampleIndex = a.sampleBase;            // global index of the next sample
    int jitterOffset = (int)(sampleIndex % 10);
    const uint64_t seed = a.seedBase + (uint64_t)inst;
    int staged = 0;
    long long flushed = sampleIndex;                 // global index of stage[.][0]

    for (int step = 0; step < a.nSteps; ++step) {
        const double data = a.stream[((size_t)step * a.channels + ch) * a.nInst + inst];
        // Synthesizer::writeInput
        inputWriteOffset += a.clock.samplesPerStep;
        if (inputWriteOffset >= bufferSize) inputWriteOffset -= bufferSize;
        const double distance = inputDistance(inputWriteOffset, lastInputSampleOffset, bufferSize);
        double s = inputDistance((double)writeIndex, lastInputSampleOffset, bufferSize);
        int written = 0;
        for (; s <= distance; s += 1.0) {
            if (s >= bufferSize) s -= bufferSize;
            const double f = s / distance;
            const double sample = lastInput * (1.0 - f) + data * f;
            antialiasY = sp.antialiasAlpha * (float)sample + (1.0f - sp.antialiasAlpha) * antialiasY;
            const float x = antialiasY;
            // ---- renderAudio(i), up to the convolution input ----
            hist[jitterOffset][lane] = x;
            jitterOffset = (jitterOffset + 1 == 10) ? 0 : jitterOffset + 1;
            const float u = esb_rand_u01f(seed, ESB_STREAM_JITTER + ch, (uint64_t)sampleIndex);
            jitterNoiseY = sp.jitterAlpha * u + (1.0f - sp.jitterAlpha) * jitterNoiseY;
            int d = (int)(jitterNoiseY * sp.inputSampleNoise * 9.0f + 0.5f);
            if (d < 0) d = 0;
            if (d > 9) d = 9;
            int idx = jitterOffset - 1 - d;
            if (idx < 0) idx += 10;
            const float f_in = hist[idx][lane];
            dcY = sp.dcAlpha * f_in + (1.0f - sp.dcAlpha) * dcY;
            const float fd = f_in - dcY;
            const float f_p = (f_in - derivPrev) * sp.sampleRate;
            derivPrev = f_in;
            const float noise = 2.0f * esb_rand_u01f(seed, ESB_STREAM_AIR + ch, (uint64_t)sampleIndex) - 1.0f;
            airY = sp.airAlpha * noise + (1.0f - sp.airAlpha) * airY;
            const float r_mixed = sp.airNoise * airY + (1.0f - sp.airNoise);
            float v_in = f_p * sp.dF_F_mix + (fd * r_mixed) * (1.0f - sp.dF_F_mix);
            if (fabsf(v_in) < 1.17549435e-38f) v_in = 0.0f;
            stage[lane][staged + written] = v_in;
            ++sampleIndex;
            ++written;
        }
        lastInput = data;
        writeIndex = (writeIndex + written) % (long long)bufferSize;
        lastInputSampleOffset = inputWriteOffset;
        staged += written;
        while (staged >= TILE) { // flush 32-sample tiles, transposed: lane = sample
            __syncwarp();
            const long long col = a.hist + (flushed - a.xOrigin) + lane;
            for (int r = 0; r < TILE; ++r) a.X[(rowBlock + r) * a.xStride + col] = stage[r][lane];
            __syncwarp();
            // move the remainder to the front of this lane's stage row
            const int rem = staged - TILE;
            for (int i = 0; i < rem; ++i) stage[lane][i] = stage[lane][TILE + i];
            staged = rem;
            flushed += TILE;
            __syncwarp();
        }
    }
    if (staged > 0) { // tail
        __syncwarp();
        const long long col = a.hist + (flushed - a.xOrigin) + lane;
        if (lane < staged)
            for (int r = 0; r < TILE; ++r) a.X[(rowBlock + r) * a.xStride + col] = stage[r][lane];
    }
    a.chanLast[row] = lastInput;
    a.chanState[0 * rows + row] = antialiasY;
    a.chanState[1 * rows + row] = jitterNoiseY;
    a.chanState[2 * rows + row] = dcY;
    a.chanState[3 * rows + row] = derivPrev;
    a.chanState[4 * rows + row] = airY;
#pragma unroll
    for (int i = 0; i < 10; ++i) a.chanState[(5 + i) * rows + row] = hist[i][lane];
}

// K5 -------------------------------------------------------------------------------------
// grid: nInstPad / 32; block: 32 threads; lane = instance within the block
constexpr int MAXCH = 4;
__global__ void __launch_bounds__(32) esb_synth_post_kernel(EsbPostArgs a) {
    __shared__ float tile[MAXCH][2][TILE][TILE + 1]; // [channel][y | x][instance][sample]
    __shared__ float outF[TILE][TILE + 1];
    __shared__ short outS[TILE][TILE + 2];
    const int lane = threadIdx.x;
    const int instRaw = blockIdx.x * TILE + lane;
    const bool valid = instRaw < a.nInst;
    const EsbSynthParams sp = a.sp;
    float masterY = valid ? a.instState[instRaw] : 0.0f;
    float levelerPeak = valid ? a.instState[a.nInst + instRaw] : 0.0f;

    for (long long base = a.first; base < a.first + a.count; base += TILE) {
        const int m = (int)((a.first + a.count - base < TILE) ? (a.first + a.count - base) : TILE);
        // load: lane = sample
        __syncwarp();
        for (int c = 0; c < a.channels; ++c) {
            const long long rowBlock = (long long)c * a.nInstPad + blockIdx.x * TILE;
            for (int r = 0; r < TILE; ++r) {
                float y = 0.0f, x = 0.0f;
                if (lane < m) {
                    y = a.Y[(rowBlock + r) * a.yStride + (base - a.n0) + lane];
                    x = a.X[(rowBlock + r) * a.xStride + a.hist + (base - a.xOrigin) + lane];
                }
                tile[c][0][r][lane] = y;
                tile[c][1][r][lane] = x;
            }
        }
        __syncwarp();
        // compute: lane = instance, samples in order
        for (int k = 0; k < m; ++k) {
            float signal = 0.0f;
            for (int c = 0; c < a.channels; ++c) {
                const float y = tile[c][0][lane][k], v_in = tile[c][1][lane][k];
                const float v = sp.convolution * y + (1.0f - sp.convolution) * v_in;
                signal += v;
            }
            masterY = sp.masterAlpha * signal + (1.0f - sp.masterAlpha) * masterY;
            // LevelingFilter::f
            levelerPeak = 0.9999f * levelerPeak;
            const float av = fabsf(masterY);
            if (av > levelerPeak) levelerPeak = av;
            float out = 0.0f;
            if (levelerPeak != 0.0f) {
                float gain = sp.levelerTarget / levelerPeak;
                if (gain < sp.levelerMinGain) gain = sp.levelerMinGain;
                if (gain > sp.levelerMaxGain) gain = sp.levelerMaxGain;
                out = masterY * gain;
            }
            out *= sp.volume;
            long r = lroundf(out);
            if (r > 32767) r = 32767;
            if (r < -32768) r = -32768;
            outF[lane][k] = out;
            outS[lane][k] = (short)r;
        }
        __syncwarp();
        // store: lane = sample
        if (lane < m) {
            for (int r = 0; r < TILE; ++r) {
                const int instR = blockIdx.x * TILE + r;
                if (instR >= a.nInst) break;
                const long long o = (long long)instR * a.pcmStride + (base - a.pcmOrigin) + lane;
                a.pcm[o] = outS[r][lane];
                if (a.pcmFloat) a.pcmFloat[o] = outF[r][lane];
            }
        }
    }
    if (valid) {
        a.instState[instRaw] = masterY;
        a.instState[a.nInst + instRaw] = levelerPeak;
    }
}

// SineSimulator::simulateStep_ / writeToSynthesizer (SineSimulator.cpp:84-105): channel 0 carries
// sin(phase) * 28000 at f = rpm / 6 Hz, rpm = 800 + 5200 * throttle (SineEngine.h:35-38).
__global__ void esb_sine_kernel(double *stream, int nSteps, int channels, int nInst, const double *ctl, double *phase, double simFreq) {
    const int inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= nInst) return;
    const double kPi = 3.14159265358979323846;
    const double throttle = ctl[inst]; // ESB_CTL_THROTTLE row
    const double rpm = 800.0 + 5200.0 * throttle;
    const double frequency = rpm / 6.0;
    double ph = phase[inst];
    for (int step = 0; step < nSteps; ++step) {
        ph += (2.0 * kPi) / (simFreq / frequency);
        if (ph >= 2.0 * kPi) ph -= 2.0 * kPi;
        for (int c = 0; c < channels; ++c) stream[((size_t)step * channels + c) * nInst + inst] = (c == 0) ? esb_sin(ph) * 28000.0 : 0.0;
    }
    phase[inst] = ph;
}

} // namespace

cudaError_t esbLaunchSine(double *stream, int nSteps, int channels, int nInst, const double *ctl, double *phase, double simFreq, cudaStream_t st) {
    esb_sine_kernel<<<(nInst + 127) / 128, 128, 0, st>>>(stream, nSteps, channels, nInst, ctl, phase, simFreq);
    return cudaGetLastError();
}

cudaError_t esbLaunchSynthPre(const EsbPreArgs &a, cudaStream_t stream) {
    dim3 grid(a.nInstPad / TILE, a.channels);
    esb_synth_pre_kernel<<<grid, TILE, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t esbLaunchSynthPost(const EsbPostArgs &a, cudaStream_t stream) {
    if (a.channels > MAXCH) return cudaErrorInvalidValue;
    if (a.count <= 0) return cudaSuccess;
    esb_synth_post_kernel<<<a.nInstPad / TILE, TILE, 0, stream>>>(a);
    return cudaGetLastError();
}
