"""Static evidence from the built library (no GPU needed): for every kernel in libstanmath_cuda.so the resource
usage (registers, stack = spills, static shared memory) and the count of the SASS mnemonics that prove what it is
made of: UBLKCP (cp.async.bulk, the TMA 1-D copies), SYNCS (mbarrier), DMMA (FP64 tensor-core mma.sync m8n8k4),
DFMA/DMUL/DADD (FP64 pipe), MUFU (special-function unit), BAR, LDG/STG, RED/ATOM.

    python tools/sass_summary.py > profiles/r02_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "math_b200", "lib", "libstanmath_cuda.so")
MNEMONICS = ["UBLKCP", "SYNCS", "DMMA", "DFMA", "DMUL", "DADD", "MUFU", "BAR", "LDG", "STG", "LDS", "STS", "RED", "ATOM", "SHFL"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), stdout=subprocess.PIPE, text=True).stdout.splitlines()
    return dict(zip(names, out))


def short(name):
    name = re.sub(r"^void ", "", name)
    name = name.replace("(anonymous namespace)::", "")
    name = re.sub(r"\(.*\)$", "", name)
    return name.replace("(anonymous namespace)::", "").replace("stan_cuda::", "")


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], stdout=subprocess.PIPE, text=True).stdout
    usage = {}
    for m in re.finditer(r"Function (\S+):\n\s+(.*)", res):
        usage[m.group(1)] = dict(kv.split(":") for kv in m.group(2).split() if ":" in kv)
    sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True).stdout
    counts, cur = collections.defaultdict(collections.Counter), None
    for line in sass.splitlines():
        m = re.match(r"\s+Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            op = m.group(1)
            counts[cur]["total"] += 1
            for mn in MNEMONICS:
                if op == mn or op.startswith(mn + "."):
                    counts[cur][mn] += 1
    names = demangle(sorted(usage))
    print("# %s, sm_100a: %d kernels" % (os.path.basename(LIB), len(usage)))
    print("# regs / stack (spill bytes) / static smem / SASS instructions, then mnemonic counts (zeros omitted)")
    for k in sorted(usage, key=lambda k: short(names[k])):
        u, c = usage[k], counts[k]
        mn = " ".join("%s=%d" % (m, c[m]) for m in MNEMONICS if c[m])
        print("%-110s REG=%-4s STACK=%-4s SHARED=%-6s SASS=%-6d %s" % (short(names[k])[:110], u.get("REG"), u.get("STACK"),
                                                                     u.get("SHARED"), c["total"], mn))
    spilled = [short(names[k]) for k in usage if int(usage[k].get("STACK", 0)) > 0]
    print("# kernels with a stack frame (spills or local arrays): %s" % (", ".join(spilled) if spilled else "none"))


if __name__ == "__main__":
    sys.exit(main())
