"""GPU diagnostic (not a pytest): sustained FP32 throughput of the non-tensor FMA pipes, the roofline denominator of the
blend kernels (SURVEY.md 8d asks for it; MEASURED_PEAKS.json has HBM and bf16 only).

  python tests/gpu_fma_peak.py [out.json]

Dependent-free FFMA chains (8 per thread, 8 resident warps per CTA, CTAs = 148 SMs x 8), all SMs, long enough
(~40 ms per launch) that the clock under load is what nvidia-smi reports; the median SM clock during the run is
recorded beside the figure. Also measured: Blackwell's packed FFMA2 (fma.rn.f32x2) and unfused FMUL+FADD.
bench.py reads profiles/r02_fp32_peak.json for `roofline.peak` of the blend kernels."""
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "segment-any-gaussian-splatting-slam_b200", "lib", "libgsr_microbench.so")


def measure(blocks_per_sm=8, iters=20000, reps=5):
    L = C.CDLL(LIB)
    L.gsr_mb_fma.restype = C.c_float
    L.gsr_mb_fma.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    scratch = torch.zeros(16, device="cuda")
    rows, stop = [], [False]

    def sample():
        p = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,"
                              "clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,"
                              "clocks_event_reasons.sw_thermal_slowdown",
                              "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE, text=True)
        for line in p.stdout:
            rows.append([c.strip() for c in line.split(",")])
            if stop[0]:
                break
        p.terminate()

    th = threading.Thread(target=sample, daemon=True)
    th.start()
    t0 = time.time()
    while not rows and time.time() - t0 < 10:
        time.sleep(0.05)
    out = {}
    names = {0: "ffma", 1: "ffma2_packed", 2: "fmul_fadd"}
    blocks = sms * blocks_per_sm
    for mode in (0, 1, 2):
        first = len(rows)
        ms = float(L.gsr_mb_fma(mode, blocks, iters, reps, scratch.data_ptr(), torch.cuda.current_stream().cuda_stream))
        if ms <= 0:
            raise RuntimeError("microbenchmark launch failed")
        time.sleep(0.06)
        flops = float(L.gsr_mb_flops_per_thread_iter(mode)) * iters * 256 * blocks
        clk = sorted(float(r[0]) for r in rows[first:] if r and r[0].replace(".", "").isdigit())
        out[names[mode]] = dict(tflops=round(flops / ms / 1e9, 2), ms_per_launch=round(ms, 3),
                                sm_mhz_median=clk[len(clk) // 2] if clk else None,
                                warp_inst_per_clk_per_sm=None)
        if clk:
            per_inst = 64.0 if mode != 1 else 128.0   # flops per warp-instruction (32 lanes x 2, x2 packed)
            f = out[names[mode]]
            f["warp_inst_per_clk_per_sm"] = round(flops / per_inst / (ms * 1e-3) / (clk[len(clk) // 2] * 1e6) / sms, 3)
    stop[0] = True
    reasons = sorted({n for r in rows for n, v in zip(("sw_power_cap", "hw_slowdown", "sw_thermal_slowdown"), r[3:6])
                      if v.lower().startswith("active")})
    nominal = sms * 128 * 2 * float(rows[-1][1]) * 1e6 / 1e12 if rows else None
    return dict(gpu=torch.cuda.get_device_name(0), sms=sms, blocks=blocks, threads_per_block=256, iters=iters,
                chains_per_thread=8, modes=out, fp32_tflops=out["ffma"]["tflops"], nominal_tflops_at_max_clock=nominal,
                sm_max_mhz=float(rows[-1][1]) if rows else None, reasons=reasons,
                how="dependent-free FFMA chains, 8 per thread, 8 warps per CTA, %d CTAs per SM, CUDA events around %d "
                    "back-to-back launches; clocks sampled by nvidia-smi during the run" % (blocks_per_sm, reps))


if __name__ == "__main__":
    r = measure()
    s = json.dumps(r, indent=1)
    print(s)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(s + "\n")
