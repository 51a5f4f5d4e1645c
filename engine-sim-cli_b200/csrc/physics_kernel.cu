// physics_kernel.cu — K1+K2 fused: the batched physics kernel (sm_100a).
//
// One lane group per engine instance (G = 1, 2, 4, 8 or 16 lanes, the next power of two >= the
// cylinder count), 32/G instances per warp, state in registers for the whole launch; see
// physics_core.cuh for the algorithm and its mapping onto lanes and physics_run.cuh for the
// launch-level loop.  FP64 throughout, built with -fmad=false so every operation rounds as in
// the reference's operation order.
#include "physics_run.cuh"
#include "kernels.h"

namespace esbk {

template <int G>
__global__ void __launch_bounds__(128) esb_physics_kernel(RunArgs a) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int inst = tid / G;
    if (inst >= a.nInst) return; // groups are lane-aligned, so a whole group leaves together
    Group<G> g;
    g.lane = tid % G;
    const int laneInWarp = threadIdx.x & 31;
    g.mask = (G >= 32) ? 0xffffffffu : (((1u << G) - 1u) << (laneInWarp & ~(G - 1)));
    runInstance<G>(a, inst, g);
}

} // namespace esbk

cudaError_t esbLaunchPhysics(const esbk::RunArgs &a, int lanes, int blockThreads, cudaStream_t stream) {
    const long long threads = (long long)a.nInst * lanes;
    const int grid = (int)((threads + blockThreads - 1) / blockThreads);
    switch (lanes) {
    case 1: esbk::esb_physics_kernel<1><<<grid, blockThreads, 0, stream>>>(a); break;
    case 2: esbk::esb_physics_kernel<2><<<grid, blockThreads, 0, stream>>>(a); break;
    case 4: esbk::esb_physics_kernel<4><<<grid, blockThreads, 0, stream>>>(a); break;
    case 8: esbk::esb_physics_kernel<8><<<grid, blockThreads, 0, stream>>>(a); break;
    case 16: esbk::esb_physics_kernel<16><<<grid, blockThreads, 0, stream>>>(a); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}
