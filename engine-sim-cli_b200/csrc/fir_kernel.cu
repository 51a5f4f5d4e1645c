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
// a CTA share one IR.  Taps are consumed in chunks of KT=256: chunk c needs x columns
// [W0 + c*KT, W0 + c*KT + KT + T), i.e. 256-float segments c and c+1 of each row.  Segments live
// in a 3-slot ring per row in shared memory (row stride 772 floats = 4 mod 32 banks, so the
// lane-per-row float4 window loads are conflict-free); segment c+2 and IR chunk c+1 are
// fetched by cp.async.bulk (TMA bulk copy, one 1 KB copy per row) onto an mbarrier while chunk
// c is being multiplied.  Each thread slides a 16-float register window over its row: per 4
// taps one LDS.128 of x, one broadcast LDS.128 of hr and 48 FFMA.  Chunk sums are folded into
// a second accumulator (two-level summation keeps the FP32 error ~sqrt(KT)+sqrt(Lp/KT) ulp).
#include "kernels.h"

#include <cstdint>

namespace {

constexpr int KT = ESB_FIR_KT;      // taps per chunk
constexpr int T = ESB_FIR_T;        // outputs per CTA tile
constexpr int R = 12;               // outputs per thread
constexpr int WARPS = T / R;        // 16
constexpr int ROWS = 32;
constexpr int NSLOT = 3;
constexpr int RING = NSLOT * KT;    // 768 floats per row
constexpr int XROW = RING + 4;      // 772: stride = 4 (mod 32) banks
constexpr int SMEM_X = ROWS * XROW * 4;
constexpr int SMEM_H = 2 * KT * 4;
constexpr int SMEM_BYTES = SMEM_X + SMEM_H + 64;

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

// grid: (tiles over n, rows / 32); block: 512 threads
__global__ void __launch_bounds__(WARPS * 32, 2) esb_fir_kernel(EsbFirArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    float *xs = reinterpret_cast<float *>(smem);                    // [ROWS][XROW]
    float *hs = reinterpret_cast<float *>(smem + SMEM_X);           // [2][KT]
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + SMEM_X + SMEM_H); // [NSLOT]

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rowBlock = blockIdx.y * ROWS;
    const long long tileN0 = a.n0 + (long long)blockIdx.x * T;     // first output sample of this tile
    const int ir = (rowBlock / a.rowsPerIr) % a.nIr;
    const float *hr = a.hr + (size_t)ir * a.Lp;
    const int nChunks = a.Lp / KT;
    // column of x[n - Lp + 1] for n = tileN0 : hist + (tileN0 - xOrigin) - Lp + 1
    const long long W0 = a.hist + (tileN0 - a.xOrigin) - a.Lp + 1;
    const float *xrow = a.X + (size_t)(rowBlock + lane) * a.xStride + W0;

    const uint32_t barBase = smemAddr(bars);
    if (threadIdx.x == 0) {
        for (int i = 0; i < NSLOT; ++i) mbarInit(barBase + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // producer = warp 0: lane r copies row r's segment; lane 0 also copies the IR chunk
    auto issue = [&](int seg, int hchunk) {
        const uint32_t bar = barBase + 8 * (seg % NSLOT);
        if (lane == 0) mbarExpectTx(bar, ROWS * KT * 4 + (hchunk >= 0 ? KT * 4 : 0));
        __syncwarp();
        bulkLoad(smemAddr(xs + lane * XROW + (seg % NSLOT) * KT), xrow + (size_t)seg * KT, KT * 4, bar);
        if (lane == 0 && hchunk >= 0) bulkLoad(smemAddr(hs + (hchunk & 1) * KT), hr + (size_t)hchunk * KT, KT * 4, bar);
    };
    if (warp == 0) {
        issue(0, -1);
        issue(1, 0);
    }

    float total[R];
#pragma unroll
    for (int i = 0; i < R; ++i) total[i] = 0.0f;
    const float *myRow = xs + lane * XROW;

    mbarWait(barBase + 0, 0); // segment 0
    for (int c = 0; c < nChunks; ++c) {
        // everyone is done with chunk c-1, so slot (c+2)%3 (segment c-1) and IR buffer (c+1)&1 are free
        __syncthreads();
        if (warp == 0 && c + 1 < nChunks) issue(c + 2, c + 1); // the last chunk needs no further data
        // segment c+1 and IR chunk c have landed?
        mbarWait(barBase + 8 * ((c + 1) % NSLOT), ((c + 1) / NSLOT) & 1);

        const float *h = hs + (c & 1) * KT;
        int pos = (c % NSLOT) * KT + warp * R;  // ring column of x for (output 0, tap 0) of this thread
        float acc[R];
#pragma unroll
        for (int i = 0; i < R; ++i) acc[i] = 0.0f;
        float win[16];
        // preload the first 12 window entries (3 float4)
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const float4 v = *reinterpret_cast<const float4 *>(myRow + pos);
            win[4 * q + 0] = v.x; win[4 * q + 1] = v.y; win[4 * q + 2] = v.z; win[4 * q + 3] = v.w;
            pos += 4; if (pos >= RING) pos -= RING;
        }
        for (int jj = 0; jj < KT; jj += 16) {
#pragma unroll
            for (int u = 0; u < 4; ++u) { // 4 taps per u; window slot rotation has period 4
                const float4 v = *reinterpret_cast<const float4 *>(myRow + pos);
                pos += 4; if (pos >= RING) pos -= RING;
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

cudaError_t esbLaunchFir(const EsbFirArgs &a, cudaStream_t stream) {
    if (a.count <= 0) return cudaSuccess;
    if (a.Lp % KT != 0 || a.rows % ROWS != 0 || (a.hist + 1) % 4 != 0 || (a.n0 - a.xOrigin) % 4 != 0 || a.xStride % 4 != 0 || a.yStride % 4 != 0 ||
        a.rowsPerIr % ROWS != 0)
        return cudaErrorInvalidValue;
    // The attribute belongs to the (function, device) pair and a process may hold batches on several devices,
    // from several host threads: set it before every launch (a host-side table update, no device work).
    cudaError_t e = cudaFuncSetAttribute(esb_fir_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    dim3 grid((unsigned)((a.count + T - 1) / T), (unsigned)(a.rows / ROWS));
    esb_fir_kernel<<<grid, WARPS * 32, SMEM_BYTES, stream>>>(a);
    return cudaGetLastError();
}
