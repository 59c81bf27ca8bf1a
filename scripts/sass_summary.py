"""Per-kernel SASS opcode counts of the built library (the B200_PROFILING.md mnemonics that prove tcgen05 / TMA use):

    python scripts/sass_summary.py [muscle_b200/lib/libmuscle_b200.so] > profiles/r2_sass_summary.txt

UTCHMMA = tcgen05.mma (kind::f16), UTCBAR = tcgen05.commit, LDTM/STTM = tcgen05.ld/st, UTMALDG = TMA tensor load,
UTMASTG = TMA tensor store, SYNCS = mbarrier ops, MUFU.EX2 = exp2 on the XU pipe, REDG/RED/ATOMG = global atomics."""
import collections
import re
import subprocess
import sys

WATCH = ("UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "MUFU.EX2", "MUFU.RCP",
         "MUFU.LG2", "HMMA", "FFMA", "REDG", "RED.", "ATOMG", "ATOM.", "LDG", "STG", "LDS", "STS", "SHFL", "MATCH", "REDUX")


def main():
    so = sys.argv[1] if len(sys.argv) > 1 else "muscle_b200/lib/libmuscle_b200.so"
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
    demangle = lambda s: subprocess.run(["cu++filt", s], capture_output=True, text=True).stdout.strip() or s
    cur, counts, total = None, collections.OrderedDict(), collections.Counter()
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = demangle(m.group(1))
            cur = re.sub(r"\((int|bool|unsigned int)\)", "", cur).replace("void ", "")
            cur = re.sub(r"\(.*$", "", cur)
            counts[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            op = m.group(1)
            total[cur] += 1
            for w in WATCH:
                if op.startswith(w):
                    counts[cur][w] += 1
                    break
    print("# %s : %d kernels" % (so, len(counts)))
    for k, c in counts.items():
        tags = "  ".join("%s=%d" % (w, c[w]) for w in WATCH if c[w])
        print("%-62s instr=%-6d %s" % (k[:62], total[k], tags))


if __name__ == "__main__":
    main()
