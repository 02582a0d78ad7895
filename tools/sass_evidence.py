"""Counts the Blackwell-specific SASS mnemonics per kernel of libdv3b200.so (cuobjdump -sass) and prints a markdown table.

    python tools/sass_evidence.py [title] > profiles/rNN_sass_evidence.md
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "sheeprl_prey_b200", "libdv3b200.so")

COLS = [("tcgen05.mma (UTCHMMA)", r"\bUTCHMMA"), ("tcgen05.commit (UTCBAR)", r"\bUTCBAR"), ("TMA tensor load (UTMALDG)", r"\bUTMALDG"),
        ("bulk copy global/shared/DSMEM (UBLKCP)", r"\bUBLKCP"), ("TMEM load (LDTM)", r"\bLDTM"), ("st.async / store to peer smem (STAS)", r"\bSTAS"),
        ("mbarrier ops (SYNCS)", r"\bSYNCS"), ("mma.sync (HMMA)", r"\bHMMA"), ("cluster barrier (UCGABAR)", r"\bUCGABAR"),
        ("cp.async (LDGSTS)", r"\bLDGSTS")]


def main():
    title = sys.argv[1] if len(sys.argv) > 1 else "SASS evidence"
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = {}
    counts, order, cur = {}, [], None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            order.append(cur)
            counts[cur] = [0] * len(COLS)
            continue
        if cur is None:
            continue
        for i, (_, pat) in enumerate(COLS):
            if re.search(pat, line):
                counts[cur][i] += 1
    names = subprocess.run(["cu++filt"] + order, capture_output=True, text=True).stdout.splitlines() if order else []
    for raw, nice in zip(order, names):
        nice = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", nice)
        nice = re.sub(r"\((int|bool|unsigned int)\)", "", nice)
        demangle[raw] = re.sub(r"\(.*", "", nice)
    print(f"# {title}\n")
    print("`cuobjdump -sass sheeprl_prey_b200/libdv3b200.so` (sm_100a), static instruction counts per kernel "
          "(tools/sass_evidence.py). Only kernels with at least one of the first six columns are listed.\n")
    print("| kernel | " + " | ".join(c for c, _ in COLS) + " |")
    print("|---|" + "---:|" * len(COLS))
    for raw in order:
        c = counts[raw]
        if sum(c[:6]) == 0:
            continue
        print(f"| `{demangle.get(raw, raw)}` | " + " | ".join(str(x) for x in c) + " |")


if __name__ == "__main__":
    main()
