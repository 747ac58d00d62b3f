"""
Tensor-core / TMA / TMEM instruction census of the shipped library, per kernel, from `cuobjdump -sass` (runs without a GPU):

    python tools/sass_census.py > profiles/r02_sass_tensor_tma_tmem.txt
"""
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

LIBRARY = Path(__file__).resolve().parents[1] / "gance_b200" / "csrc" / "libgance_b200.so"
WANTED = re.compile(r"\b(UTCHMMA|UTCQMMA|UTMALDG|UTMASTG|LDTM|STTM|UTCBAR|UTCATOMSWS|UCGABAR_\w+|SYNCS)\b")


def mnemonic(line: str) -> str:
    """The counted key of one SASS line: opcode plus the modifiers that matter (.2CTA, .2D/.4D, MULTICAST, LDTM width)."""
    op = re.search(r"\b((?:UTCHMMA|UTCQMMA|UTMALDG|UTMASTG|LDTM|STTM|UTCBAR|UTCATOMSWS|UCGABAR_\w+|SYNCS)[\w.]*)", line).group(1)
    parts = op.split(".")
    keep = [parts[0]]
    for part in parts[1:]:
        if part in ("2CTA", "MULTICAST", "2D", "3D", "4D", "5D") or (parts[0] == "LDTM" and part.startswith("16dp")):
            keep.append("16" if part.startswith("16dp") else part)
    return ".".join(keep) if parts[0] != "SYNCS" else "SYNCS"


def main() -> None:
    sass = subprocess.run(["cuobjdump", "-sass", str(LIBRARY)], stdout=subprocess.PIPE, text=True, check=True).stdout
    kernels = OrderedDict()
    current = None
    for line in sass.splitlines():
        found = re.search(r"Function : (\S+)", line)
        if found:
            current = found.group(1)
            kernels[current] = []
        elif current is not None and WANTED.search(line) and "/*" in line:
            kernels[current].append(line)
    names = list(kernels)
    demangled = subprocess.run(["cu++filt", *names], stdout=subprocess.PIPE, text=True, check=True).stdout.splitlines()
    print(f"cuobjdump -sass gance_b200/csrc/{LIBRARY.name} (sm_100a), tensor-core / TMA / TMEM instructions per kernel (tools/sass_census.py).")
    print("UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), UTMALDG = cp.async.bulk.tensor load (.2CTA = signalling the pair leader's barrier),")
    print("UTMASTG = cp.async.bulk.tensor store, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit (.2CTA.MULTICAST = to both CTAs of a pair),")
    print("UTCATOMSWS = tcgen05.alloc/dealloc, UCGABAR = barrier.cluster, SYNCS = mbarrier operations.\n")
    totals = Counter()
    for mangled, pretty in zip(names, demangled):
        lines = kernels[mangled]
        if not any("UTCHMMA" in line or "UTCQMMA" in line for line in lines):
            continue
        counts = Counter(mnemonic(line) for line in lines)
        totals.update(counts)
        short = re.sub(r"^void gance::tc::|\(CUtensorMap_st.*$|\(int\)|\(bool\)", "", pretty).replace("(bool)", "")
        print(short.strip())
        print("    " + ", ".join(f"{key} x{counts[key]}" for key in sorted(counts)))
        shown = set()
        for line in lines:
            key = mnemonic(line)
            if key in shown or key == "SYNCS" or key.startswith("UTCATOMSWS") or key.startswith("UCGABAR"):
                continue
            shown.add(key)
            print("      e.g. " + line.strip()[:150])
    print("\ntotals over the tcgen05 kernels: " + ", ".join(f"{key} x{totals[key]}" for key in sorted(totals)))


if __name__ == "__main__":
    sys.exit(main())
