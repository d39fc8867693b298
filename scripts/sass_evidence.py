"""Machine-code evidence (no GPU needed): per kernel of impala_b200/_C/libb200rl.so, the count of SASS instructions and of the
mnemonics that show which hardware paths a kernel uses (TMA bulk / tensor copies, mbarriers, cp.async, multimem, tensor
cores).  python scripts/sass_evidence.py > profiles/r02_sass_evidence.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "impala_b200", "_C", "libb200rl.so")
MARK = ["UBLKCP", "UBLKPF", "UTMALDG", "UTMASTG", "UTMAPF", "UTMACMDFLUSH", "SYNCS", "LDGSTS", "LDGDEPBAR", "ACQBULK", "MULTIMEM", "LDGMC", "REDG", "ATOMG", "ATOMS",
        "HMMA", "IMMA", "UTCHMMA", "UTCQMMA", "UTCIMMA", "LDTM", "STTM", "UCGABAR", "BAR", "MEMBAR", "ERRBAR", "CCTL", "MUFU", "SHFL", "ELECT",
        "LDG", "STG", "LDS", "STS", "NANOSLEEP", "WARPSYNC"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
    cur, counts, total = None, collections.OrderedDict(), collections.Counter()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur:
            total[cur] += 1
            op, mods = m.group(1), m.group(2)
            for k in MARK:
                if op == k or (k in ("MULTIMEM", "LDGMC") and (k in op)):
                    counts[cur][op + (mods if op in ("UBLKCP", "UBLKPF", "SYNCS", "LDG", "STG", "REDG", "ATOMG", "HMMA") and ("SYS" in mods or "S.G" in mods or "G.S" in mods or "TRANS" in mods or op == "HMMA" or "L2" in mods) else "")] += 1
                    break
    print("# cuobjdump -sass impala_b200/_C/libb200rl.so -- instruction counts per kernel (static), marker mnemonics only")
    print("# UBLKCP = cp.async.bulk (TMA 1-D), UBLKPF = cp.async.bulk.prefetch.L2, UTMALDG/UTMASTG = cp.async.bulk.tensor (TMA tiled), SYNCS = mbarrier ops,")
    print("# LDGSTS = cp.async, HMMA = legacy mma.sync tensor path, UTC*MMA/LDTM/STTM = tcgen05 (none in this library: nothing on the hot path is a contraction;")
    print("# the CNN around it uses mma.sync TF32, see DESIGN.md K10/K11).")
    for fn, c in counts.items():
        name = demangle(fn)
        name = re.sub(r"\((?:int|bool|unsigned int)\)", "", name)
        name = re.sub(r"\(.*", "", name).replace("b200rl::", "")[:110]
        keep = ["%s %d" % (k, v) for k, v in sorted(c.items(), key=lambda kv: -kv[1])]
        print("%-112s %6d instr | %s" % (name, total[fn], ", ".join(keep)))


if __name__ == "__main__":
    main()
