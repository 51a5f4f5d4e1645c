// physics_kernel.cu — K1+K2 fused: the batched physics kernel (sm_100a).
//
// One lane group per engine instance (G = 1, 2, 4, 8 or 16 lanes, the next power of two >= the
// cylinder count), 32/G instances per warp, rigid-body state in registers and gas cells in shared
// memory for the whole launch; see physics_core.cuh for the algorithm and its mapping onto lanes
// and physics_run.cuh for the launch-level loop and the shared-memory layout.  FP64 throughout, built with -fmad=false so every operation rounds as in
// the reference's operation order.
#include "physics_run.cuh"
#include "kernels.h"

namespace esbk {

// SLAVES: engines with master / slave connecting rods (radial engines) take the general form of the rod-journal rows.
template <int G, bool SLAVES>
__global__ void __launch_bounds__(256, 1) esb_physics_kernel(RunArgs a, int instPerWarp) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int laneInWarp = threadIdx.x & 31;
    const int slot = laneInWarp / G;
    // No thread leaves early (every thread of the block takes part in the lockstep barriers): lane groups
    // without an instance of their own shadow the last instance and store nothing.
    // With instPerWarp < 32 / G the warps are deliberately left part-empty: a batch too small to fill the machine is
    // latency-bound, and more warps hide more of the FP64 dependency latency than full warps do.
    int inst = warp * instPerWarp + slot;
    const bool live = slot < instPerWarp && inst < a.nInst;
    if (!live) inst = a.nInst - 1;
    // shared memory: the gas-network tables once per block, then one region per lane group
    extern __shared__ __align__(16) unsigned char esbSmem[];
    const EsbGasNet &gn = a.D->gas;
    const GasSmemLayout L = gasSmemLayout(gn, a.D->P.cylinderCount, a.D->P.intakeCount, a.D->fluidSteps, blockDim.x / G);
    gasTablesToShared(esbSmem, L, a.D, threadIdx.x, blockDim.x);
    __syncthreads();
    const GasView view = gasViewAt(esbSmem, L, gn, threadIdx.x / G);
    Group<G> g;
    g.lane = laneInWarp % G;
    g.mask = (G >= 32) ? 0xffffffffu : (((1u << G) - 1u) << (laneInWarp & ~(G - 1)));
#ifdef ESB_PHASE_CLOCKS
    g.clkOn = blockIdx.x == 0 && threadIdx.x == 0;
    g.clkLast = clock64();
#endif
    runInstance<G, SLAVES>(a, inst, live, g, view);
}

} // namespace esbk

// Self-check of the kernel's straight-line arithmetic (FastAr): out[i] = a[i] / b[i] (op 0) or sqrt(a[i]) (op 1),
// flagged[i] = 1 where FastAr would hand the exchange to the exact path instead.
__global__ void esb_fastar_check_kernel(int op, int n, const double *a, const double *b, double *out, unsigned *flagged) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    esbk::FastAr ar;
    out[i] = (op == 0) ? ar.div(a[i], b[i]) : ar.sqrt(a[i]);
    flagged[i] = ar.failed() ? 1u : 0u;
}
cudaError_t esbLaunchFastArCheck(int op, int n, const double *a, const double *b, double *out, unsigned *flagged, cudaStream_t stream) {
    esb_fastar_check_kernel<<<(n + 255) / 256, 256, 0, stream>>>(op, n, a, b, out, flagged);
    return cudaGetLastError();
}

template <int G, bool SLAVES>
static cudaError_t launchKernel(const esbk::RunArgs &a, int grid, int blockThreads, int instPerWarp, size_t smem, cudaStream_t stream) {
    // per (function, device): set before every launch, a process may drive several devices
    cudaError_t e = cudaFuncSetAttribute(esbk::esb_physics_kernel<G, SLAVES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    esbk::esb_physics_kernel<G, SLAVES><<<grid, blockThreads, smem, stream>>>(a, instPerWarp);
    return cudaGetLastError();
}

template <int G>
static cudaError_t launchG(const esbk::RunArgs &a, const EsbDeviceParams &hostD, int blockThreads, int instPerWarp, cudaStream_t stream) {
    const int full = 32 / G;
    if (instPerWarp <= 0 || instPerWarp > full) instPerWarp = full;
    const long long warps = ((long long)a.nInst + instPerWarp - 1) / instPerWarp;
    const int wpb = blockThreads / 32;
    const int grid = (int)((warps + wpb - 1) / wpb);
    const esbk::GasSmemLayout L = esbk::gasSmemLayout(hostD.gas, hostD.P.cylinderCount, hostD.P.intakeCount, hostD.fluidSteps, blockThreads / G);
    if (L.total > 227u * 1024u) return cudaErrorInvalidConfiguration;
    bool slaves = false;
    for (int i = 0; i < hostD.P.cylinderCount; ++i) slaves = slaves || hostD.P.cyl[i].masterCylinder >= 0;
    if (slaves) {
        // master / slave rods exist in the radial engines (5, 7, 9 cylinders): built for 8 and 16 lanes
        if constexpr (G == 8 || G == 16) return launchKernel<G, true>(a, grid, blockThreads, instPerWarp, L.total, stream);
        else return cudaErrorNotSupported;
    }
    return launchKernel<G, false>(a, grid, blockThreads, instPerWarp, L.total, stream);
}

cudaError_t esbLaunchPhysics(const esbk::RunArgs &a, const EsbDeviceParams &hostD, int blockThreads, int instPerWarp, int /*variant*/, cudaStream_t stream) {
    if (blockThreads % 32 != 0 || blockThreads < 32 || blockThreads > 256) return cudaErrorInvalidValue;
    switch (hostD.lanesPerInstance) {
    case 1: return launchG<1>(a, hostD, blockThreads, instPerWarp, stream);
    case 2: return launchG<2>(a, hostD, blockThreads, instPerWarp, stream);
    case 4: return launchG<4>(a, hostD, blockThreads, instPerWarp, stream);
    case 8: return launchG<8>(a, hostD, blockThreads, instPerWarp, stream);
    case 16: return launchG<16>(a, hostD, blockThreads, instPerWarp, stream);
    default: return cudaErrorInvalidValue;
    }
}

#ifdef ESB_PHASE_CLOCKS
extern "C" int EsbDebugPhaseClocks(unsigned long long *out, int reset) {
    if (cudaMemcpyFromSymbol(out, esbk::esbPhaseClk, sizeof(unsigned long long) * 32) != cudaSuccess) return -1;
    if (reset) { unsigned long long z[32] = {0}; cudaMemcpyToSymbol(esbk::esbPhaseClk, z, sizeof z); }
    return 0;
}
#endif
