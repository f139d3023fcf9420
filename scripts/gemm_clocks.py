"""SM clock / power while one GEMM mode runs back to back for ~3 s (is the tensor work power-limited?)."""
import os, sys, subprocess, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from graph_representation_learning_b200 import _lib, ops
dev = torch.device("cuda:0")
M, N, K = 128000, 256, 1792
a = torch.randn(M, K, device=dev); b = torch.randn(N, K, device=dev) / K ** 0.5
ab, bb = a.bfloat16(), b.bfloat16()
def sample(stop, rows):
    while not stop.is_set():
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True).stdout.strip()
        rows.append(out); stop.wait(0.25)
cases = [("tf32x3", lambda: ops.gemm_nt([a], [b], fp32_mode="tf32x3")), ("tf32bc", lambda: ops.gemm_nt([a], [b], fp32_mode="tf32bc")),
         ("tf32", lambda: ops.gemm_nt([a], [b], fp32_mode="tf32")), ("bf16", lambda: ops.gemm_nt([ab], [bb])), ("cublas_bf16", lambda: ab @ bb.t())]
for name, fn in cases:
    for _ in range(3): fn()
    torch.cuda.synchronize()
    stop, rows = threading.Event(), []
    th = threading.Thread(target=sample, args=(stop, rows)); th.start()
    t0 = time.time(); n = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < 3.0:
        for _ in range(50): fn()
        n += 50
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    stop.set(); th.join()
    ms = e0.elapsed_time(e1) / n
    print(f"{name}: {ms:.4f} ms/GEMM {2.0*M*N*K/ms/1e9:.0f} TF/s useful | smi (MHz, W, power_cap): {rows[2:-1]}")
    time.sleep(1.0)
