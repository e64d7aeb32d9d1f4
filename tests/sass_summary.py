"""SASS opcode summary of the product library (run here, no GPU needed):
  python tests/sass_summary.py > profiles/r02_sass_opcodes.txt
Per kernel: registers, static instruction count, and the counts of the opcodes that show what the kernel is built on
(B200_PROFILING.md "What proves a Blackwell-native kernel"): UBLKCP / SYNCS (TMA bulk copies completing on mbarriers),
UTCHMMA / LDTM / STTM (tcgen05.mma, tcgen05.ld / st: TMEM), FFMA2 (Blackwell packed fp32 FMA), HMMA (warp-level mma.sync), REDG (vector red.global.add), REDUX, MATCH, LDSM, ..."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "segment-any-gaussian-splatting-slam_b200", "lib", "libgsraster_b200.so")
KEY = ("UBLKCP", "LDGSTS", "SYNCS", "UTMALDG", "UTCHMMA", "UTCBAR", "LDTM", "STTM", "FFMA2", "FMUL2", "FADD2", "HMMA", "REDG", "RED", "REDUX", "MATCH",
       "VOTE", "SHFL", "LDSM", "LDS", "STS", "LDG", "STG", "ATOMG", "ATOMS", "MUFU", "FFMA", "FMUL", "FADD", "BAR", "LDL", "STL")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    regs = {}
    for m in re.finditer(r"Function (\S+):\n\s+REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", res):
        regs[m.group(1)] = (int(m.group(2)), int(m.group(3)), int(m.group(4)), int(m.group(5)))
    cur, ops = None, collections.OrderedDict()
    for line in sass.splitlines():
        m = re.match(r"\s+Function : (\S+)", line)
        if m:
            cur = m.group(1)
            ops[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            ops[cur][m.group(1).split(".")[0]] += 1
    demangle = subprocess.run(["c++filt"] + list(ops), capture_output=True, text=True).stdout.splitlines()
    print("# %s (sm_100a), cuobjdump -sass: static opcode counts per kernel" % os.path.relpath(LIB, ROOT))
    for (fn, c), name in zip(ops.items(), demangle):
        r = regs.get(fn, (0, 0, 0, 0))
        short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).replace("gsr::", "").replace("void ", "")
        keys = " ".join("%s=%d" % (k, c[k]) for k in KEY if c.get(k))
        print("%-52s regs %3d stack %3d  inst %5d | %s" % (short[:52], r[0], r[1], sum(c.values()), keys))


if __name__ == "__main__":
    main()
