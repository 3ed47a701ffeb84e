"""Per-kernel SASS instruction-mix table of libperseus_b200.so (cuobjdump -sass): the Blackwell-native evidence
(B200_PROFILING.md: tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG / UBLKCP) plus the plain global accesses.

    python tools/sass_summary.py > profiles/r02_sass_summary.txt
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "perseus_b200", "libperseus_b200.so")
COLS = ["UTCHMMA", "UTCHMMA.2CTA", "UTMALDG", "UTMALDG.MULTICAST", "UTMALDG.2CTA", "UBLKCP", "UTMASTG", "LDTM", "STTM", "UTCBAR",
        "SYNCS", "LDG", "STG", "LDS", "STS", "SHFL", "HMMA", "FFMA", "DFMA", "R2UR"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def short(name):
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    name = re.sub(r"pb200::", "", name)
    name = re.sub(r"^void ", "", name)
    return re.sub(r"\(.*$", "", name)


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, cur = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = dict.fromkeys(COLS, 0)
            counts[cur]["instructions"] = 0
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        c = counts[cur]
        c["instructions"] += 1
        base = op.split(".")[0]
        if base in c:
            c[base] += 1
        if op.startswith("UTCHMMA") and ".2CTA" in op:
            c["UTCHMMA.2CTA"] += 1
        if op.startswith("UTMALDG"):
            if ".MULTICAST" in op:
                c["UTMALDG.MULTICAST"] += 1
            if ".2CTA" in op:
                c["UTMALDG.2CTA"] += 1
    names = demangle(list(counts))
    show = [c for c in COLS if any(v[c] for v in counts.values())]
    print("# SASS instruction mix per kernel of perseus_b200/libperseus_b200.so (sm_100a), static counts from `cuobjdump -sass`")
    print("# UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), UTMALDG = cp.async.bulk.tensor (TMA tile load; .MULTICAST / .2CTA variants),")
    print("# UBLKCP = cp.async.bulk (1-D TMA copy), LDTM = tcgen05.ld (TMEM -> registers), UTCBAR = tcgen05.commit, SYNCS = mbarrier ops.")
    print("# No UTMASTG / STTM: every epilogue stores with plain STG (16-byte vectors arranged into complete 128-byte lines).")
    print("kernel | instructions | " + " | ".join(show))
    for k in sorted(counts, key=lambda k: short(names[k])):
        c = counts[k]
        print(f"{short(names[k])} | {c['instructions']} | " + " | ".join(str(c[x]) for x in show))


if __name__ == "__main__":
    sys.exit(main())
