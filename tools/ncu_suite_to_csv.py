"""`ncu --set full` report of tools/gemm_suite.py (ITERS=1: one launch per shape) -> the summary CSV
that bench.py's `roofline.traffic` reads:  python tools/ncu_suite_to_csv.py gpurun_out/r02_gemm_suite.ncu-rep > profiles/r02_ncu_gemm_suite.csv"""
import csv
import subprocess
import sys

LABELS = ["rn2 N512 K512 (+res,LN)", "rn1 N512 K512 (LN,SiLU)", "plain N512 K512", "relu N2048 K512", "rn1 N512 K2048 (LN,SiLU)",
          "plain N1536 K512"]
KEEP = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__cluster_size", "launch__registers_per_thread",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max"]
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = [hdr.index(k) for k in KEEP if k in hdr]
w = csv.writer(sys.stdout)
print('"# ncu --set full --clock-control none, tools/gemm_suite.py (M = 90112), round-2 kernels (elected-lane producer / issuer); one launch per shape"')
w.writerow(["shape"] + [hdr[i] for i in idx])
w.writerow([""] + [units[i] for i in idx])
for n, label in enumerate(LABELS):
    r = data[n]
    w.writerow([label] + [r[i] for i in idx])
