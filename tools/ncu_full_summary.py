"""
Text summary of a `ncu --set full --import-source on` report (tools/gpu_profile.sh): one metric table over all captured
launches (raw page), then the stall-sampling hot spots of every launch (tools/ncu_hotspots.py).

    python tools/ncu_full_summary.py gpurun_out/r02_v9_convs.ncu-rep "<one line saying what was captured>" > profiles/r02_ncu_convs_v9_summary.txt
"""
import csv
import subprocess
import sys
from pathlib import Path

METRICS = [
    "Kernel Name", "Block Size", "Grid Size", "launch__cluster_size", "gpu__time_duration.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "launch__registers_per_thread",
]


def main() -> None:
    report, what = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
    raw = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    header, units, launches = rows[0], rows[1], rows[2:]
    print(what)
    print()
    for name in METRICS:
        column = header.index(name)
        cells = [launch[column][:24] for launch in launches]
        print(f"{name:<80} {units[column]:<14} " + " | ".join(f"{cell:>24}" for cell in cells))
    hotspots = Path(__file__).resolve().parent / "ncu_hotspots.py"
    for index in range(len(launches)):
        print(f"\n=== launch {index}")
        sys.stdout.flush()
        subprocess.run([sys.executable, str(hotspots), report, str(index)], check=True)


if __name__ == "__main__":
    main()
