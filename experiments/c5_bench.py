"""Config-5 shape on one GPU: Pearson of n x 1024 float32 (default 25 M rows = 102 GB).  C5_ROWS / C5_COLS / FSX_GRAM_WIDE_SLAB."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import fsx
ctx = fsx.Context(0)
n, d = int(os.environ.get("C5_ROWS", 25_000_000)), int(os.environ.get("C5_COLS", 1024))
cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=5, seed2=2000 + j, p0=0.01 * j, p1=1.0 + 0.001 * j, p2=1.5) for j in range(d)]
import subprocess, threading
clk = []
def poll():
    p = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown", "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE, text=True)
    for line in p.stdout:
        clk.append(line.strip())
threading.Thread(target=poll, daemon=True).start()
for rep in range(int(os.environ.get("C5_REPS", 2))):
    t0 = time.perf_counter()
    ctx.pearson(cols)
    wall = (time.perf_counter() - t0) * 1e3
    ms = ctx.timing().main_kernel_ms
    print(f"n={n} d={d}: call {wall:.1f} ms, gram passes {ms:.1f} ms -> {n * d * (d + 1) / 2 * 2 / ms / 1e9:.1f} TFLOP/s symmetric-equivalent, "
          f"{n * (d // 64) * (d // 64 + 1) / 2 * 128 * 448 * 2 / ms / 1e9:.0f} TOP/s int8 issued, {n * d * 4 / ms / 1e6:.0f} GB/s", flush=True)
print("clock samples (sm MHz, W, sw_power_cap, hw_slowdown):", clk[-12:], flush=True)
ctx.close()
