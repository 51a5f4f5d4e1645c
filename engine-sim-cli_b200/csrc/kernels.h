// kernels.h — launch entry points of the CUDA kernels (host-callable, nvcc-compiled TUs).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace esbk { struct RunArgs; }
struct EsbDeviceParams;

// K1+K2: physics (physics_kernel.cu).  hostD: the host copy of the parameter block a.D points to (lane count, shared-memory layout)
cudaError_t esbLaunchPhysics(const esbk::RunArgs &a, const EsbDeviceParams &hostD, int blockThreads, int instPerWarp, int variant, cudaStream_t stream);

// Synthesizer clock: the 10 kHz -> 44.1 kHz resampler bookkeeping of Synthesizer::writeInput
// (archive/PHYSICS_ANALYSIS.md:39-55), identical for every instance, so the host tracks it.
struct EsbSynthClock {
    double inputWriteOffset, lastInputSampleOffset;
    long long writeIndex;
    double samplesPerStep; // audioSampleRate / inputSampleRate
    double bufferSize;     // inputBufferSize (44100)
};
// Advances the clock by nSteps and returns the number of 44.1 kHz samples those steps produce.
long long esbSynthClockAdvance(EsbSynthClock &c, int nSteps);

struct EsbSynthParams {
    float antialiasAlpha, jitterAlpha, airAlpha, dcAlpha, masterAlpha;
    float inputSampleNoise, airNoise, dF_F_mix, convolution, volume;
    float levelerTarget, levelerMaxGain, levelerMinGain;
    float sampleRate;
};

// K3: resample + per-channel filter chain up to the convolution input (synth_kernels.cu)
struct EsbPreArgs {
    const double *stream;     // [nSteps][channels][nInst]
    int nSteps, channels, nInst, nInstPad;
    EsbSynthClock clock;      // clock BEFORE these steps
    float *chanState;         // [ESB_PRE_STATE_FIELDS][rows] floats (+ doubles packed, see synth_kernels.cu)
    double *chanLast;         // [rows] last input sample (double)
    long long sampleBase;     // number of samples rendered before this launch (noise counter / jitter phase)
    float *X;                 // [rows][xStride], sample m of a row at X[row*xStride + hist + m - xOrigin]
    long long xStride;
    long long hist;           // offset of sample index xOrigin in a row
    long long xOrigin;
    uint64_t seedBase;
    EsbSynthParams sp;
};
#define ESB_PRE_STATE_FIELDS 15 /* antialiasY, jitterNoiseY, dcY, derivPrev, airY, jitterHistory[10] */
cudaError_t esbLaunchSynthPre(const EsbPreArgs &a, cudaStream_t stream);

// K4: FIR convolution (fir_kernel.cu)
struct EsbFirArgs {
    const float *X;           // [rows][xStride]
    long long xStride;
    long long hist;           // row offset of sample index xOrigin; (hist + 1) % 4 == 0
    long long xOrigin;
    const float *hr;          // [nIr][Lp] reversed, zero-padded impulse responses
    int Lp;                   // padded length, multiple of ESB_FIR_KT
    int rows;                 // multiple of 32
    int rowsPerIr;            // rows [i*rowsPerIr, (i+1)*rowsPerIr) use IR (i % nIr)
    int nIr;
    long long n0;             // first output sample index (multiple of 4)
    int count;                // number of outputs per row
    float *Y;                 // [rows][yStride], output sample n at Y[row*yStride + (n - n0)]
    long long yStride;
};
#define ESB_MAX_SAMPLES_PER_STEP 64 /* K3 stage size: ceil(sampleRate / simFreq) + 1 must not exceed it */
#define ESB_FIR_KT 128
#define ESB_FIR_T 192
cudaError_t esbLaunchFir(const EsbFirArgs &a, cudaStream_t stream);
int esbFirSmemBytes();

// K5: channel mix + master low-pass + leveler + int16 (synth_kernels.cu)
struct EsbPostArgs {
    const float *X; long long xStride, hist, xOrigin;
    const float *Y; long long yStride; long long n0; // Y sample n at (n - n0)
    int channels, nInst, nInstPad;
    long long first; int count; // samples [first, first+count)
    float *instState;         // [2][nInst]: masterY, levelerPeak
    int16_t *pcm;             // [nInst][pcmStride], sample n at (n - pcmOrigin)
    float *pcmFloat;          // same layout or null
    long long pcmStride, pcmOrigin;
    EsbSynthParams sp;
};
cudaError_t esbLaunchSynthPost(const EsbPostArgs &a, cudaStream_t stream);

// SineSimulator input (synth_kernels.cu)
cudaError_t esbLaunchSine(double *stream, int nSteps, int channels, int nInst, const double *ctl, double *phase, double simFreq, cudaStream_t st);

// FP32 / FP64 pipe peak probes (peak_probe.cu): register-resident FMA loops, returns ops issued
cudaError_t esbLaunchPeakProbe(int fp64, int fma, int blocks, int threads, int iters, float *sink, cudaStream_t stream);

// Self-check of the physics kernel's straight-line division / sqrt (physics_kernel.cu; device pointers)
cudaError_t esbLaunchFastArCheck(int op, int n, const double *a, const double *b, double *out, unsigned *flagged, cudaStream_t stream);
