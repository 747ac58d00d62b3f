"""Summarise one kernel of an .ncu-rep: per-role stall samples and the hottest SASS instructions (source page)."""
import csv
import subprocess
import sys


def main() -> None:
    report, skip = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "0"
    out = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"],
                         stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    print(rows[0][1][:160])
    header, data = rows[1], rows[2:]
    samples, executed = header.index("Warp Stall Sampling (All Samples)"), header.index("Instructions Executed")
    total = sum(int(r[samples]) for r in data)
    print(f"SASS instructions: {len(data)}; stall samples: {total}")
    reasons = [i for i, h in enumerate(header) if h.startswith("stall_") and "Not Issued" not in h]
    totals = {header[i]: sum(int(r[i] or 0) for r in data) for i in reasons}
    print("stall reasons:", ", ".join(f"{k[6:]} {v}" for k, v in sorted(totals.items(), key=lambda kv: -kv[1]) if v))
    print(f"{'idx':>5} {'samples':>8} {'executed':>10}  instruction")
    for index, row in enumerate(data):
        if int(row[samples]) >= total * 0.006 or "SYNCS.PHASECHK" in row[1] or "UTCBAR" in row[1]:
            print(f"{index:5d} {row[samples]:>8} {row[executed]:>10}  {row[1].strip()[:110]}")


if __name__ == "__main__":
    main()
