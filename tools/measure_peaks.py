"""Measure the FP64/FP32 ceilings used as roofline denominators on this B200.

Run on the GPU box:  python tools/measure_peaks.py > gpurun_out/peaks.json
 - tools/peaks.cu : register-only DMMA.8x8x4 / DFMA / FFMA / mma.sync-tf32 issue-rate loops
 - torch.matmul   : cuBLAS DGEMM / SGEMM / TF32 GEMM, burst (best of 10) and sustained (3 s loop)
MEASURED_PEAKS.json (driver-written) has only HBM copy and bf16; FP64 is what this path is bound by.
"""
import json, os, subprocess, sys, time

HERE = os.path.dirname(os.path.abspath(__file__))


def build_and_run_micro():
    exe = os.path.join(HERE, "peaks_bin")
    subprocess.check_call(["nvcc", "-O3", "-gencode", "arch=compute_100a,code=sm_100a",
                           "-o", exe, os.path.join(HERE, "peaks.cu")])
    out = subprocess.check_output([exe], timeout=300).decode()
    return json.loads(out)


def gemm_peaks():
    import torch
    res = {}
    dev = torch.device("cuda:0")
    for name, dt, n, tf32 in (("dgemm", torch.float64, 8192, False), ("sgemm", torch.float32, 8192, False),
                              ("tf32gemm", torch.float32, 8192, True)):
        torch.backends.cuda.matmul.allow_tf32 = tf32
        a = torch.randn(n, n, device=dev, dtype=dt); b = torch.randn(n, n, device=dev, dtype=dt)
        c = torch.empty_like(a)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(10):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(); torch.matmul(a, b, out=c); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        res[name + "_burst_tflops"] = 2.0 * n ** 3 / best * 1e-9
        # sustained: back to back for ~3 s
        t0 = time.time(); cnt = 0
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        while time.time() - t0 < 3.0:
            for _ in range(5):
                torch.matmul(a, b, out=c); cnt += 1
            torch.cuda.synchronize()
        e1.record(); torch.cuda.synchronize()
        res[name + "_sustained_tflops"] = 2.0 * n ** 3 * cnt / e0.elapsed_time(e1) * 1e-9
        del a, b, c
    torch.backends.cuda.matmul.allow_tf32 = False
    return res


if __name__ == "__main__":
    out = {"when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}
    out.update(build_and_run_micro())
    out.update(gemm_peaks())
    try:
        q = subprocess.check_output(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw",
                                     "--format=csv,noheader"]).decode().strip()
        out["nvidia_smi_after"] = q
    except Exception as e:  # noqa
        out["nvidia_smi_after"] = str(e)
    json.dump(out, sys.stdout, indent=1); print()
