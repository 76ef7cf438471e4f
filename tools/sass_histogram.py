#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of libcoherent_cuda.so (cuobjdump -sass), so that the instruction mix behind
the pipe-utilisation counters is on record: how much of a kernel is butterflies (FFMA2 / FMUL2 / FADD2 / FFMA / FMUL /
FADD), how much is twiddle / address arithmetic (MOV / IMAD / IADD3 / LEA / LOP3 / SHF), shared-memory traffic
(LDS / STS), global traffic (LDG / STG / UBLKCP = bulk TMA) and barriers (BAR / SYNCS / WARPSYNC).

    python tools/sass_histogram.py [lib.so] > profiles/rN_sass_histogram.md
Static counts (every instruction once, loops not weighted): read them as the mix of the unrolled item loop, which is
>95 % of each of these kernels' SASS."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "coherent-rtlsdr_b200", "lib", "libcoherent_cuda.so")
GROUPS = [
    ("fp32 math (packed)", r"^(FFMA2|FMUL2|FADD2)"),
    ("fp32 math (scalar)", r"^(FFMA|FMUL|FADD|FSEL|FMNMX|FSETP|MUFU|FCHK)"),
    ("integer dot / int math", r"^(IDP|IMAD|IADD3|IADD|LEA|ISETP|IMNMX|IABS|VIADD|VIMNMX)"),
    ("logic / shift / move", r"^(LOP3|SHF|MOV|PRMT|SEL|I2F|F2I|I2FP|F2FP|CS2R|S2R|S2UR|R2UR|UMOV|ULOP|UIADD|UIMAD|ULEA|USHF|USEL|UISETP|PLOP3|P2R|R2P|SGXT|BMSK|POPC|FLO|BREV)"),
    ("shared memory", r"^(LDS|STS|LDSM|STSM)"),
    ("global / local memory", r"^(LDG|STG|LD\.|ST\.|LDL|STL|ATOM|RED|LDC|ULDC|CCTL)"),
    ("bulk TMA / mbarrier", r"^(UBLKCP|UTMALDG|UTMASTG|SYNCS|FENCE|UBLKPF)"),
    ("barrier / shuffle / vote", r"^(BAR|WARPSYNC|SHFL|VOTE|MATCH|REDUX|NANOSLEEP|DEPBAR|MEMBAR|ERRBAR)"),
    ("tensor core", r"^(IMMA|HMMA|UTC|QMMA|BMMA)"),
    ("control flow", r"^(BRA|BSSY|BSYNC|EXIT|RET|CALL|NOP|BREAK|YIELD|BPT|JMP|BRX|WARPSYNC)"),
]


def main():
    txt = subprocess.check_output(["cuobjdump", "-sass", LIB], text=True, errors="replace")
    kern, name = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.check_output(["c++filt", m.group(1)], text=True).strip()
            name = re.sub(r"\(.*", "", name).replace("coh::", "").replace("void ", "")
            kern[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
        if m and name:
            kern[name][m.group(1)] += 1
    print(f"# SASS opcode histogram of `{os.path.relpath(LIB, ROOT)}` (static counts, `cuobjdump -sass`)\n")
    for name, c in kern.items():
        total = sum(c.values())
        if total < 200:
            continue
        print(f"## `{name}` — {total} instructions\n")
        print("| group | count | share | opcodes |\n|---|---|---|---|")
        seen = set()
        for g, pat in GROUPS:
            ops = {o: n for o, n in c.items() if re.match(pat, o) and o not in seen}
            seen |= set(ops)
            if ops:
                n = sum(ops.values())
                detail = ", ".join(f"{o} {k}" for o, k in sorted(ops.items(), key=lambda kv: -kv[1])[:8])
                print(f"| {g} | {n} | {100.0 * n / total:.1f} % | {detail} |")
        rest = {o: n for o, n in c.items() if o not in seen}
        if rest:
            n = sum(rest.values())
            print(f"| other | {n} | {100.0 * n / total:.1f} % | " + ", ".join(f"{o} {k}" for o, k in sorted(rest.items(), key=lambda kv: -kv[1])[:8]) + " |")
        print()


if __name__ == "__main__":
    main()
