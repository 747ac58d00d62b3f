"""
Summarise an ncu launch list (`ncu --metrics ... --csv --log-file <csv>`, tools/gpu_profile.sh) of one bench.py run:
finds the LAST step (the launches between the last two `advance_counter_kernel` launches, i.e. one CUDA-graph replay),
prints per-kernel totals (time, share, DRAM MB, tensor-pipe %, issue %, shared-memory wavefronts) and every conv launch
in launch order, and writes the conv-group DRAM traffic JSON bench.py's `roofline.traffic` reads.

    python tools/ncu_launch_summary.py gpurun_out/r02_v3_launches.csv profiles/r02_conv_traffic.json > profiles/r02_launches_v3_summary.txt
"""
import csv
import json
import re
import sys
from collections import OrderedDict, defaultdict

CONV = re.compile(r"modconv3x3_(tc|pair|halo|halo2|halo_pair)_kernel|upconv_fused(_pair)?_kernel|upconv_phase|modconv3x3_tf32x3")
TIME, READ, WRITE = "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum"
TENSOR = "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"
ISSUE = "smsp__issue_active.avg.pct_of_peak_sustained_elapsed"
SHARED = "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"
INST = "sm__inst_executed.sum"

UNIT_SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def short(name: str) -> str:
    name = re.sub(r"\(.*", "", name)
    return name.replace("void ", "").replace("gance::", "").replace("tc::", "")


def read_launches(path: str):
    lines = open(path, encoding="utf-8", errors="replace").read().splitlines()
    start = next(i for i, line in enumerate(lines) if line.startswith('"ID"'))
    launches = OrderedDict()
    for row in csv.DictReader(lines[start:]):
        launch = launches.setdefault(int(row["ID"]), {"name": row["Kernel Name"], "grid": row["Grid Size"], "block": row["Block Size"]})
        value = float(row["Metric Value"].replace(",", "")) * UNIT_SCALE.get(row["Metric Unit"], 1.0)
        launch[row["Metric Name"]] = value
    return list(launches.values())


def main() -> None:
    path = sys.argv[1]
    traffic_path = sys.argv[2] if len(sys.argv) > 2 else None
    launches = read_launches(path)
    marks = [i for i, launch in enumerate(launches) if "advance_counter_kernel" in launch["name"]]
    if len(marks) < 2:
        raise SystemExit("need two graph replays in the launch list")
    # one replay = everything after the previous replay's counter advance up to and including this one's
    step = [l for l in launches[marks[-2] + 1: marks[-1] + 1] if not l["name"].startswith("void at::")]
    total_us = sum(l[TIME] for l in step)
    total_bytes = sum(l.get(READ, 0) + l.get(WRITE, 0) for l in step)
    print(f"ncu launch list {path}: {len(launches)} launches; last step = launches {marks[-2] + 1}..{marks[-1]} "
          f"({len(step)} of our kernels, serialised, cold cache): {total_us:.1f} us, DRAM {total_bytes / 1e6:.0f} MB")
    groups = defaultdict(list)
    for launch in step:
        groups[short(launch["name"])].append(launch)
    print(f"{'kernel':58} {'n':>2} {'us':>8} {'share':>6} {'DRAM MB':>8} {'tensor%':>7} {'issue%':>6}")
    for name, members in sorted(groups.items(), key=lambda kv: -sum(l[TIME] for l in kv[1])):
        us = sum(l[TIME] for l in members)
        weighted = lambda key: sum(l.get(key, 0) * l[TIME] for l in members) / max(us, 1e-9)  # noqa: E731
        print(f"{name[:58]:58} {len(members):2d} {us:8.1f} {100 * us / total_us:5.1f}% "
              f"{sum(l.get(READ, 0) + l.get(WRITE, 0) for l in members) / 1e6:8.1f} {weighted(TENSOR):7.1f} {weighted(ISSUE):6.1f}")
    print()
    print("modulated-convolution launches of the step, in launch order:")
    print(f"{'kernel':52} {'grid':>12} {'us':>8} {'rd MB':>7} {'wr MB':>7} {'tensor%':>7} {'issue%':>6} {'Minst':>7} {'smem Mwave':>10}")
    conv = [l for l in step if CONV.search(l["name"])]
    for launch in conv:
        print(f"{short(launch['name'])[:52]:52} {launch['grid']:>12} {launch[TIME]:8.1f} {launch.get(READ, 0) / 1e6:7.1f} "
              f"{launch.get(WRITE, 0) / 1e6:7.1f} {launch.get(TENSOR, 0):7.1f} {launch.get(ISSUE, 0):6.1f} "
              f"{launch.get(INST, 0) / 1e6:7.1f} {launch.get(SHARED, 0) / 1e6:10.1f}")
    conv_us = sum(l[TIME] for l in conv)
    conv_bytes = sum(l.get(READ, 0) + l.get(WRITE, 0) for l in conv)
    print(f"conv group: {len(conv)} launches, {conv_us:.1f} us = {100 * conv_us / total_us:.1f}% of the step, DRAM {conv_bytes / 1e6:.0f} MB")
    if traffic_path:
        with open(traffic_path, "w", encoding="utf-8") as handle:
            json.dump({
                "what": "DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of every modulated-convolution launch of ONE "
                        f"bench step (32 frames, 1024^2 config-f, fp16, fresh noise), summed; from {path} "
                        "(tools/gpu_profile.sh: ncu --metrics ... --clock-control none, bench.py --steps 2 --warmup 3)",
                "conv_launches_per_step": len(conv),
                "conv_dram_bytes_per_step": conv_bytes,
                "all_kernels_dram_bytes_per_step": total_bytes,
                "conv_time_share_under_ncu": conv_us / total_us,
                "step_time_under_ncu_us": total_us,
            }, handle, indent=1)


if __name__ == "__main__":
    main()
