// fir_kernel.cu — K4: batched impulse-response convolution (ConvolutionFilter::f for every
// channel of every instance), FP32 FMA on the CUDA cores of sm_100a, TMA-staged.
//
// Reference: upstream convolution_filter.cpp as driven from Synthesizer::renderAudio — a
// direct-form FIR over a circular shift register, y[n] = sum_k h[k] x[n-k], one sample at a
// time per channel (call sites ScriptLoadHelpers.cpp:73-78, SimulatorInitHelpers.cpp:72-85,
// SineSimulator.cpp:37-41).  Here the whole time axis of a launch is tiled.
//
// Formulation.  The IR is stored reversed and zero-padded to Lp (multiple of KT):
// hr[j] = h[Lp-1-j], so  y[n] = sum_{j<Lp} hr[j] * x[n - Lp + 1 + j]  walks x FORWARD as j
// grows.  Rows keep sample m at column hist+m with (hist+1) % 4 == 0, which makes every window
// start 16-byte aligned.
//
// CTA = 32 rows (lane = row) x T=192 outputs (16 warps x R=12 outputs per thread), all rows of
// a CTA share one IR.  Taps are consumed in chunks of KT=128: chunk c needs x columns
// [W0 + c*KT, W0 + c*KT + KT + T) of each row.  That window of all 32 rows (plus 4 columns, so that
// the shared-memory row stride is 324 floats = 4 mod 32 banks and the lane-per-row float4 loads are
// conflict-free) is ONE TMA tensor copy: the rows are described to the TMA unit as a 3-D tensor
// [row][column / 4][4] and the box is 32 x 81 x 4, which lands densely as [32][324] floats.  Two
// such buffers; chunk c+2's box and IR chunk c+2 are fetched (cp.async.bulk.tensor + cp.async.bulk
// onto an mbarrier, issued by one thread) as soon as every warp is done with chunk c.  No CTA-wide
// barrier in the loop: a warp that finishes chunk c bumps a shared counter, the LAST of the 16 to do
// so issues the copies, and warps meet again only through the "data has landed" mbarriers - a fast
// warp runs up to a chunk ahead of a slow one.  Every shared-memory address in the tap loop is
// (per-chunk base + immediate).  Each thread slides a 16-float register window over its row: per 4
// taps one LDS.128 of x, one broadcast LDS.128 of hr and 48 FFMA.  Chunk sums are folded into a
// second accumulator (two-level summation keeps the FP32 error ~sqrt(KT)+sqrt(Lp/KT) ulp).
// Neighbouring chunks' windows overlap by T columns, so a CTA pulls each x value 2.5 times from L2
// into shared memory (3 TB/s over the GPU at the FP32 peak, all L2 hits); DRAM traffic is unchanged.
#include "kernels.h"

#include <cuda.h>
#include <cstdint>
#include <mutex>

namespace {

constexpr int KT = ESB_FIR_KT;      // taps per chunk
constexpr int T = ESB_FIR_T;        // outputs per CTA tile
constexpr int R = 12;               // outputs per thread
constexpr int WARPS = T / R;        // 16
constexpr int ROWS = 32;
constexpr int NBUF = 2;
constexpr int XROW = KT + T + 4;    // 324 floats per row of a chunk's window: stride = 4 (mod 32) banks
constexpr int BUF_BYTES = ROWS * XROW * 4;
constexpr int SMEM_X = NBUF * BUF_BYTES;
constexpr int SMEM_H = NBUF * KT * 4;
constexpr int SMEM_BYTES = SMEM_X + SMEM_H + 64; // + NBUF mbarriers + NBUF counters
static_assert(XROW % 32 == 4 && XROW % 4 == 0 && XROW / 4 <= 256 && BUF_BYTES % 128 == 0, "TMA box 4 x (XROW/4) x 32, buffers 128-byte aligned");
static_assert(T % R == 0 && R % 4 == 0 && KT % 16 == 0, "tile shape");

__device__ __forceinline__ uint32_t smemAddr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbarExpectTx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulkLoad(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ void tensorLoad3d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst), "l"(map), "r"(bar),
                 "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}

#ifndef ESB_FIR_UNROLL
#define ESB_FIR_UNROLL 8
#endif
constexpr int UNROLL = ESB_FIR_UNROLL; // 16-tap groups per loop body

// grid: (tiles over n, rows / 32); block: 512 threads.  xmap: a.X as [rows][xStride / 4][4] floats.
__global__ void __launch_bounds__(WARPS * 32, 2) esb_fir_kernel(EsbFirArgs a, const __grid_constant__ CUtensorMap xmap) {
    extern __shared__ __align__(128) unsigned char smem[];
    float *xs = reinterpret_cast<float *>(smem);                    // [NBUF][ROWS][XROW]
    float *hs = reinterpret_cast<float *>(smem + SMEM_X);           // [NBUF][KT]
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + SMEM_X + SMEM_H); // [NBUF] "landed" barriers
    unsigned *done = reinterpret_cast<unsigned *>(bars + NBUF);     // [NBUF] warps that have finished chunk c, at c % NBUF

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rowBlock = blockIdx.y * ROWS;
    const long long tileN0 = a.n0 + (long long)blockIdx.x * T;     // first output sample of this tile
    const int ir = (rowBlock / a.rowsPerIr) % a.nIr;
    const float *hr = a.hr + (size_t)ir * a.Lp;
    const int nChunks = a.Lp / KT;
    // column of x[n - Lp + 1] for n = tileN0 : hist + (tileN0 - xOrigin) - Lp + 1 (a multiple of 4)
    const int W0q = (int)((a.hist + (tileN0 - a.xOrigin) - a.Lp + 1) >> 2);

    const uint32_t barBase = smemAddr(bars);
    if (threadIdx.x == 0) {
        for (int i = 0; i < NBUF; ++i) { mbarInit(barBase + 8 * i, 1); done[i] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // one thread: the window of chunk k and IR chunk k into buffer k % NBUF, onto that buffer's barrier
    auto fetch = [&](int k) {
        const int b = k & (NBUF - 1);
        const uint32_t bar = barBase + 8 * b;
        mbarExpectTx(bar, BUF_BYTES + KT * 4);
        tensorLoad3d(smemAddr(xs) + b * BUF_BYTES, &xmap, 0, W0q + k * (KT / 4), rowBlock, bar);
        bulkLoad(smemAddr(hs + b * KT), hr + (size_t)k * KT, KT * 4, bar);
    };
    if (threadIdx.x == 0) {
        fetch(0);
        if (nChunks > 1) fetch(1);
    }

    float total[R];
#pragma unroll
    for (int i = 0; i < R; ++i) total[i] = 0.0f;
    const float *myRow = xs + lane * XROW + warp * R;

    for (int c = 0; c < nChunks; ++c) {
        const int b = c & (NBUF - 1);
        mbarWait(barBase + 8 * b, (c / NBUF) & 1); // window and IR chunk c have landed

        const float *h = hs + b * KT;
        const float *xw = myRow + b * (ROWS * XROW); // x of (output 0, tap 0) of this thread
        float acc[R];
#pragma unroll
        for (int i = 0; i < R; ++i) acc[i] = 0.0f;
        float win[16];
        // preload the first 12 window entries (3 float4)
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const float4 v = *reinterpret_cast<const float4 *>(xw + 4 * q);
            win[4 * q + 0] = v.x; win[4 * q + 1] = v.y; win[4 * q + 2] = v.z; win[4 * q + 3] = v.w;
        }
#pragma unroll UNROLL
        for (int jj = 0; jj < KT; jj += 16) {
#pragma unroll
            for (int u = 0; u < 4; ++u) { // 4 taps per u; window slot rotation has period 4
                const float4 v = *reinterpret_cast<const float4 *>(xw + 12 + jj + 4 * u);
                win[(12 + 4 * u + 0) & 15] = v.x; win[(12 + 4 * u + 1) & 15] = v.y;
                win[(12 + 4 * u + 2) & 15] = v.z; win[(12 + 4 * u + 3) & 15] = v.w;
                const float4 hv = *reinterpret_cast<const float4 *>(h + jj + 4 * u);
#pragma unroll
                for (int i = 0; i < R; ++i) acc[i] = fmaf(hv.x, win[(4 * u + 0 + i) & 15], acc[i]);
#pragma unroll
                for (int i = 0; i < R; ++i) acc[i] = fmaf(hv.y, win[(4 * u + 1 + i) & 15], acc[i]);
#pragma unroll
                for (int i = 0; i < R; ++i) acc[i] = fmaf(hv.z, win[(4 * u + 2 + i) & 15], acc[i]);
#pragma unroll
                for (int i = 0; i < R; ++i) acc[i] = fmaf(hv.w, win[(4 * u + 3 + i) & 15], acc[i]);
            }
        }
#pragma unroll
        for (int i = 0; i < R; ++i) total[i] += acc[i];

        // done with chunk c: the last warp to say so refills this buffer with chunk c + NBUF
        if (c + NBUF < nChunks) {
            __syncwarp();
            if (lane == 0) {
                __threadfence_block();
                if (atomicAdd(&done[b], 1u) == WARPS - 1) {
                    done[b] = 0; // next bumped after chunk c + NBUF, whose data is only now being asked for
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // the warps' reads of the buffer, then the copies' writes
                    fetch(c + NBUF);
                }
            }
        }
    }

    // epilogue: each lane owns 12 consecutive outputs of its row
    const long long nRel = (long long)blockIdx.x * T + warp * R; // relative to a.n0
    float *yrow = a.Y + (size_t)(rowBlock + lane) * a.yStride + nRel;
#pragma unroll
    for (int q = 0; q < R / 4; ++q) {
        if (nRel + 4 * q + 3 < a.count) {
            *reinterpret_cast<float4 *>(yrow + 4 * q) = make_float4(total[4 * q], total[4 * q + 1], total[4 * q + 2], total[4 * q + 3]);
        } else {
            for (int i = 0; i < 4; ++i)
                if (nRel + 4 * q + i < a.count) yrow[4 * q + i] = total[4 * q + i];
        }
    }
}

} // namespace

int esbFirSmemBytes() { return SMEM_BYTES; }

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// cuTensorMapEncodeTiled through the runtime's driver entry-point lookup (the library does not link libcuda)
EncodeTiledFn encodeTiled() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}
} // namespace

cudaError_t esbLaunchFir(const EsbFirArgs &a, cudaStream_t stream) {
    if (a.count <= 0) return cudaSuccess;
    if (a.Lp % KT != 0 || a.rows % ROWS != 0 || (a.hist + 1) % 4 != 0 || (a.n0 - a.xOrigin) % 4 != 0 || a.xStride % 4 != 0 || a.yStride % 4 != 0 ||
        a.rowsPerIr % ROWS != 0 || (reinterpret_cast<uintptr_t>(a.X) & 15) != 0 || (reinterpret_cast<uintptr_t>(a.hr) & 15) != 0)
        return cudaErrorInvalidValue;
    // the rows as the TMA unit sees them: [rows][xStride / 4][4] floats; a box is 32 rows x 81 quads = one chunk's window.
    // Encoded per launch (host arithmetic, no device work): X and its stride are the caller's.
    const EncodeTiledFn encode = encodeTiled();
    if (!encode) return cudaErrorNotSupported;
    CUtensorMap xmap;
    const cuuint64_t dims[3] = {4, (cuuint64_t)(a.xStride / 4), (cuuint64_t)a.rows};
    const cuuint64_t strides[2] = {16, (cuuint64_t)a.xStride * 4};
    const cuuint32_t box[3] = {4, XROW / 4, ROWS};
    const cuuint32_t elem[3] = {1, 1, 1};
    if (encode(&xmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float *>(a.X), dims, strides, box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorInvalidValue;
    // The attribute belongs to the (function, device) pair and a process may hold batches on several devices,
    // from several host threads: set it before every launch (a host-side table update, no device work).
    cudaError_t e = cudaFuncSetAttribute(esb_fir_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    dim3 grid((unsigned)((a.count + T - 1) / T), (unsigned)(a.rows / ROWS));
    esb_fir_kernel<<<grid, WARPS * 32, SMEM_BYTES, stream>>>(a, xmap);
    return cudaGetLastError();
}
