// peak_probe.cu — register-resident arithmetic loops that measure the FP32 / FP64 pipe peaks of
// the device the batch runs on (the roofline denominators for K4 and K1/K2; SURVEY.md §8(d):
// "measure them on the box with a register-resident FMA microkernel").
#include "kernels.h"

namespace {

template <typename T, bool FMA>
__global__ void esb_peak_probe_kernel(int iters, float *sink) {
    T a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = (T)(threadIdx.x + i) * (T)1e-3;
    const T m = (T)0.999999, c = (T)1e-7;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (FMA) {
                    if (sizeof(T) == 8) a[i] = (T)__fma_rn((double)a[i], (double)m, (double)c);
                    else a[i] = (T)__fmaf_rn((float)a[i], (float)m, (float)c);
                } else {
                    // separate multiply and add (the physics kernels run with contraction off)
                    if (sizeof(T) == 8) a[i] = (T)__dadd_rn(__dmul_rn((double)a[i], (double)m), (double)c);
                    else a[i] = (T)__fadd_rn(__fmul_rn((float)a[i], (float)m), (float)c);
                }
            }
        }
    }
    T s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    if (s == (T)12345.678) sink[0] = (float)s;
}

} // namespace

// Each thread executes iters * 64 FMAs (or 64 mul + 64 add): 128 * iters flops per thread.
cudaError_t esbLaunchPeakProbe(int fp64, int fma, int blocks, int threads, int iters, float *sink, cudaStream_t stream) {
    if (fp64) {
        if (fma) esb_peak_probe_kernel<double, true><<<blocks, threads, 0, stream>>>(iters, sink);
        else esb_peak_probe_kernel<double, false><<<blocks, threads, 0, stream>>>(iters, sink);
    } else {
        if (fma) esb_peak_probe_kernel<float, true><<<blocks, threads, 0, stream>>>(iters, sink);
        else esb_peak_probe_kernel<float, false><<<blocks, threads, 0, stream>>>(iters, sink);
    }
    return cudaGetLastError();
}
