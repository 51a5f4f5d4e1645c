"""BASELINE config 5 — ConvolutionFilter-only microbenchmark: 16384 audio channels x 48,000-tap impulse response,
10 s (441,000-sample) streams, FP32 on the CUDA cores (K4, csrc/fir_kernel.cu) through the C ABI (EsbFirRun).

    python bench_fir.py [--channels 16384] [--taps 48000] [--samples 441000] [--irs 1|16] [--check-channels 8]

Prints ONE JSON line: achieved FP32 TFLOP/s (2*L*N*C / CUDA-event time) against the FMA-pipe peak measured live,
and the SNR of a subset of channels against FP64 (the oracle's direct FP64 convolution for a few channels, an FP64
FFT convolution for 64).  Inputs are synthetic: x ~ U(-1,1), h[k] = exp(-6k/L) * U(-1,1) * 1e-3.
The time axis is processed in chunks so that the output buffer stays small; the input rows stay resident in HBM."""
import argparse
import json
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--channels", type=int, default=16384)
    ap.add_argument("--taps", type=int, default=48000)
    ap.add_argument("--samples", type=int, default=441000)
    ap.add_argument("--irs", type=int, default=1, help="distinct impulse responses (1 or 16)")
    ap.add_argument("--chunk", type=int, default=44100, help="output samples per launch")
    ap.add_argument("--check-channels", type=int, default=8, help="channels checked against the oracle's direct FP64 convolution")
    ap.add_argument("--check-samples", type=int, default=44100)
    ap.add_argument("--reps", type=int, default=1)
    args = ap.parse_args()

    import torch
    import engine_sim_cli_b200 as esb

    if not torch.cuda.is_available():
        raise SystemExit("bench_fir.py: no CUDA device (the renderer has no CPU path)")
    dev = torch.device("cuda", 0)
    C, L, N = args.channels, args.taps, args.samples
    assert C % 32 == 0 and C % (32 * args.irs) == 0
    Lp = (L + 255) // 256 * 256
    hist = Lp + 3
    xs = (hist + N + 1024 + 3) // 4 * 4
    g = torch.Generator(device=dev)
    g.manual_seed(1234)
    X = torch.zeros((C, xs), dtype=torch.float32, device=dev)
    for r0 in range(0, C, 1024):  # generate in slabs (keeps the temporary small)
        X[r0:r0 + 1024, hist:hist + N] = torch.rand((min(1024, C - r0), N), generator=g, device=dev) * 2 - 1
    g.manual_seed(5678)
    k = torch.arange(L, device=dev, dtype=torch.float64)
    h = (torch.exp(-6 * k / L)[None, :] * (torch.rand((args.irs, L), generator=g, device=dev, dtype=torch.float64) * 2 - 1) * 1e-3).float()
    HR = torch.zeros((args.irs, Lp), dtype=torch.float32, device=dev)
    HR[:, Lp - L:] = torch.flip(h, dims=[1])
    chunk = (args.chunk + 3) // 4 * 4
    Y = torch.zeros((C, chunk), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    rows_per_ir = C // args.irs

    other_ir = 5 % args.irs                      # with several IRs: a few rows of another response are checked as well
    other_row = other_ir * rows_per_ir
    keep2 = torch.zeros((8, min(args.check_samples, N)), dtype=torch.float32, device=dev)

    def run_all(keep=None):
        n0 = 0
        while n0 < N:
            cnt = min(chunk, N - n0)
            # EsbFirRun convolves `count` outputs starting at the row origin: advance the origin by offsetting the row pointer
            esb.fir_run(X.data_ptr() + 4 * n0, xs, hist, HR.data_ptr(), Lp, args.irs, rows_per_ir, C, cnt, Y.data_ptr(), chunk, stream)
            if keep is not None and n0 < keep.shape[1]:
                m = min(cnt, keep.shape[1] - n0)
                keep[:, n0:n0 + m] = Y[:keep.shape[0], :m]
                keep2[:, n0:n0 + m] = Y[other_row:other_row + 8, :m]
            n0 += cnt

    # correctness on a subset (first launch doubles as warm-up)
    nchk = min(args.check_samples, N)
    keep = torch.zeros((64, nchk), dtype=torch.float32, device=dev)
    run_all(keep)
    torch.cuda.synchronize()
    got = keep.cpu().numpy().astype(np.float64)
    x64 = X[:64, hist:hist + nchk].cpu().numpy().astype(np.float64)
    h64 = h[0].cpu().numpy().astype(np.float64)
    from scipy.signal import fftconvolve
    ref_fft = fftconvolve(x64, h64[None, :], mode="full", axes=1)[:, :nchk]

    def snr(ref, test):
        return 10 * math.log10((ref ** 2).sum() / max(((test - ref) ** 2).sum(), 1e-300))
    snr_fft = min(snr(ref_fft[r], got[r]) for r in range(64))
    x2 = X[other_row:other_row + 8, hist:hist + nchk].cpu().numpy().astype(np.float64)
    ref2 = fftconvolve(x2, h[other_ir].cpu().numpy().astype(np.float64)[None, :], mode="full", axes=1)[:, :nchk]
    snr_other = min(snr(ref2[r], keep2.cpu().numpy().astype(np.float64)[r]) for r in range(8))
    snr_direct = None
    if args.check_channels > 0:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        nd = min(args.check_channels, 64)
        ns = min(nchk, 8192)  # direct FP64 convolution is O(L) per sample: keep it to seconds
        ref_d = ol.fir_f64(x64[:nd, :ns].astype(np.float32), h64.astype(np.float32))
        snr_direct = min(snr(ref_d[r], got[r, :ns]) for r in range(nd))

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = None
    for _ in range(args.reps):
        torch.cuda.synchronize()
        e0.record()
        run_all()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    flop = 2.0 * L * N * C
    peak = esb.probe_peak(False, True) * 1e-3
    line = {
        "metric": "ConvolutionFilter microbench (config 5) FP32 throughput", "value": flop / (best * 1e-3) * 1e-12, "unit": "TFLOP/s",
        "config": {"workload": f"{C} channels x {L}-tap IR x {N} samples, {args.irs} distinct IR(s), FP32", "chunk": chunk,
                   "l2": "input rows (%.1f GB) exceed the 126 MB L2" % (C * xs * 4 / 1e9)},
        "ms": best, "algorithmic_flop": flop, "dtype": "f32", "data": "synthetic",
        "roofline": {"bound": "fp32", "achieved": flop / (best * 1e-3) * 1e-12, "peak": peak, "unit": "TFLOP/s",
                     "frac": flop / (best * 1e-3) * 1e-12 / peak, "peak_source": "measured live: register-resident FFMA loop (EsbProbePeak)"},
        "audio_seconds_per_wall_second": C * (N / 44100.0) / (best * 1e-3),
        "snr_db_vs_fp64_fft_64ch": snr_fft, "snr_db_vs_fp64_fft_rows_of_ir_%d" % other_ir: snr_other, "snr_db_vs_oracle_direct_fp64": snr_direct,
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
