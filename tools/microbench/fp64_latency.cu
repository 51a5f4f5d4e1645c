// fp64_latency.cu — dependent-issue latency and per-scheduler throughput of the FP64 instructions the physics kernel is
// made of (DFMA / DADD / DMUL / MUFU.RCP64H / SHFL / FSEL-pair select), measured with clock64() on one warp.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_latency fp64_latency.cu && ./fp64_latency
// Prints one JSON object: cycles per instruction for a chain of 1 / 2 / 4 / 8 independent accumulators per thread, and for
// 1 / 2 / 4 warps on one scheduler (block of 32 / 32*4*k threads).  DESIGN.md quotes these next to the kernel's ncu figures.
#include <cstdio>
#include <cuda_runtime.h>

template <int OP, int CHAINS>
__global__ void chain(double *out, long long *cycles, double seed, int iters) {
    double a[CHAINS];
    for (int c = 0; c < CHAINS; ++c) a[c] = seed + c + threadIdx.x * 1e-3;
    const double m = 1.0000001, b = 1e-9;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
#pragma unroll
            for (int c = 0; c < CHAINS; ++c) {
                if (OP == 0) a[c] = __fma_rn(a[c], m, b);
                else if (OP == 1) a[c] = __dadd_rn(a[c], b);
                else if (OP == 2) a[c] = __dmul_rn(a[c], m);
                else if (OP == 3) {   // MUFU.RCP64H on the high word, as in the division sequence
                    double r;
                    asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a[c]));
                    a[c] = r;
                } else if (OP == 4) a[c] = __shfl_xor_sync(0xffffffffu, a[c], 1);
                else if (OP == 5) a[c] = (a[c] > 0.5) ? __dadd_rn(a[c], b) : m;   // DSETP + select + DADD
            }
        }
    }
    const long long t1 = clock64();
    double s = 0;
    for (int c = 0; c < CHAINS; ++c) s += a[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP, int CHAINS>
static double run(int threads, double *out, long long *cyc) {
    const int iters = 2000;
    chain<OP, CHAINS><<<1, threads>>>(out, cyc, 1.5, iters);
    chain<OP, CHAINS><<<1, threads>>>(out, cyc, 1.5, iters);
    long long h = 0;
    cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    return (double)h / (iters * 16.0 * CHAINS);
}

template <int OP> static void row(const char *name, double *out, long long *cyc, bool last) {
    std::printf("\"%s\": {\"1warp_chains_1_2_4_8\": [%.2f, %.2f, %.2f, %.2f], \"chains1_warps_per_scheduler_1_2_4\": [%.2f, %.2f, %.2f]}%s\n", name,
                run<OP, 1>(32, out, cyc), run<OP, 2>(32, out, cyc), run<OP, 4>(32, out, cyc), run<OP, 8>(32, out, cyc),
                run<OP, 1>(128, out, cyc), run<OP, 1>(256, out, cyc), run<OP, 1>(512, out, cyc), last ? "" : ",");
}

int main() {
    double *out; long long *cyc;
    cudaMalloc(&out, 1024 * sizeof(double)); cudaMalloc(&cyc, 64 * sizeof(long long));
    std::printf("{\"unit\": \"cycles per instruction per warp (dependent chain)\",\n");
    row<0>("DFMA", out, cyc, false); row<1>("DADD", out, cyc, false); row<2>("DMUL", out, cyc, false);
    row<3>("MUFU.RCP64H", out, cyc, false); row<4>("SHFL.64 (2x SHFL)", out, cyc, false); row<5>("DSETP+DADD+select", out, cyc, true);
    std::printf("}\n");
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}
