"""FP64 GEMM peak of this GPU the way MEASURED_PEAKS.json records the bf16 one: cuBLAS (torch.matmul) 8192^3,
best of 10 launches timed alone ("burst") and the mean over about 4 s of back-to-back launches ("sustained"), with the
SM clock sampled during the sustained part.  Writes one JSON object to stdout; profiles/fp64_peak_r2.json is a copy of a
run on the pool's B200 and is what bench.py's roofline divides by (MEASURED_PEAKS.json has no FP64 entry).

    python parsec-1_b200/tools/fp64_peak.py > gpurun_out/fp64_peak.json
"""
import json
import subprocess
import time

import torch


def clocks():
    try:
        out = subprocess.run(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_throttle_reasons.active",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout.strip()
        f = [x.strip() for x in out.split(",")]
        return {"sm_mhz": float(f[0]), "sm_max_mhz": float(f[1]), "power_w": float(f[2]), "throttle_reasons": f[3]}
    except Exception:
        return None


def main():
    n = 8192
    a = torch.rand(n, n, dtype=torch.float64, device="cuda") - 0.5
    b = torch.rand(n, n, dtype=torch.float64, device="cuda") - 0.5
    c = torch.empty(n, n, dtype=torch.float64, device="cuda")
    for _ in range(3):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    flop = 2.0 * n ** 3
    best = 1e30
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b, out=c)
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    reps = max(10, int(4000.0 / best))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    samples = []
    for i in range(reps):
        torch.matmul(a, b, out=c)
        if i == reps // 2:
            torch.cuda.synchronize()   # let the queue drain to the middle, then sample the clock under load
            for _ in range(8):
                torch.matmul(a, b, out=c)
            samples.append(clocks())
    e1.record()
    e1.synchronize()
    sustained_ms = e0.elapsed_time(e1) / (reps + 8)
    props = torch.cuda.get_device_properties(0)
    sm_mhz = samples[0]["sm_mhz"] if samples and samples[0] else None
    print(json.dumps({
        "gpu": props.name, "sms": props.multi_processor_count, "n": n,
        "fp64_tflops_burst": flop / (best * 1e-3) * 1e-12, "fp64_tflops_sustained": flop / (sustained_ms * 1e-3) * 1e-12,
        "burst_ms": best, "sustained_ms": sustained_ms, "sustained_launches": reps + 8,
        "clocks_under_load": samples[0] if samples else None,
        # 4 sub-partitions x one DMMA.8x8x4 (512 flop) per 16 cycles = 128 flop per clock per SM
        "dmma_issue_peak_tflops_at_clock": (props.multi_processor_count * 128 * sm_mhz * 1e6 * 1e-12) if sm_mhz else None,
        "how": "torch.matmul fp64 (cuBLAS) 8192^3: best of 10 single launches; mean of ~4 s of back-to-back launches",
        "when": time.strftime("%Y-%m-%d %H:%M:%S"),
    }))


if __name__ == "__main__":
    main()
