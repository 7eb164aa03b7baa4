"""Measure sustained cuBLAS DGEMM on this box: the FP64 roofline denominator
(MEASURED_PEAKS.json carries no FP64 figure).  Prints one JSON line."""
import json, sys, torch
def main(n=8192, secs=3.0):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    c = torch.empty_like(a)
    for _ in range(3): torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); torch.matmul(a, b, out=c); e1.record(); e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    # sustained: back-to-back for `secs`
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    reps = max(3, int(secs * 1e3 / best))
    e0.record()
    for _ in range(reps): torch.matmul(a, b, out=c)
    e1.record(); e1.synchronize()
    sus = e0.elapsed_time(e1) / reps
    fl = 2.0 * n ** 3
    print(json.dumps({"n": n, "dgemm_tflops_burst": fl / best * 1e-9, "dgemm_tflops_sustained": fl / sus * 1e-9, "reps": reps}))
if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 8192)
