#!/usr/bin/env python3
"""Per-opcode histogram of the SASS of selected kernels of libxsg.so (cuobjdump -sass), split by the pipe the
instruction is issued to on sm_100 (ALU: logic / shifts / integer adds / compares / PRMT; FMA: IMAD, IDP, integer
multiplies and IMAD-form moves and shifts; LSU; control / uniform / other).  The count is STATIC (instructions in the
binary); the hot loop of count_kernel is fully unrolled over the 8 rows of a granule, so the straight-line part of a
kernel divided by 8 is the per-row figure DESIGN.md 4.1 argues with.

  python tools/sass_hist.py 'count_kernel<5, true, true, false, false>' 'count_kernel<3, false, true, false, false>' emit_kernel
"""
import collections
import re
import subprocess
import sys

ALU = {"LOP3", "SHF", "IADD3", "IADD", "VIADD", "ISETP", "PRMT", "SEL", "LEA", "POPC", "FLO", "BREV", "IABS", "IMNMX", "VIMNMX",
       "SHL", "SHR", "LOP", "PLOP3", "FSEL", "BMSK", "SGXT", "VABSDIFF", "VABSDIFF4", "ICMP", "I2I", "CS2R", "P2R", "R2P", "VOTE",
       "VOTEU", "MOV", "UFLO", "ULOP3", "UIADD3", "USHF", "UMOV", "USEL", "UISETP", "ULEA", "UPRMT", "UPOPC", "UIMAD", "R2UR", "S2UR"}
FMA = {"IMAD", "IDP", "IDP4A", "IMUL", "FFMA", "FMUL", "FADD"}
LSU = {"LDG", "STG", "LDS", "STS", "LD", "ST", "ATOM", "ATOMG", "RED", "LDC", "LDCU", "ULDC", "SHFL", "MATCH", "REDUX", "LDL", "STL",
       "CCTL", "MEMBAR", "ERRBAR", "LDSM", "ATOMS"}


def pipe(op):
    base = op.split(".")[0]
    if base in FMA:
        return "FMA"
    if base.startswith("U") and base not in ("ULDC",):
        return "UNIFORM"
    if base in LSU:
        return "LSU"
    if base in ALU:
        return "ALU"
    return "OTHER"


def main():
    lib = "xsearch_b200/libxsg.so"
    wanted = sys.argv[1:] or ["count_kernel<5, true, true, false, false>", "count_kernel<3, false, true, false, false>", "emit_kernel"]
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    parts = re.split(r"\n\s*Function : (\S+)\n", sass)
    for mangled, body in zip(parts[1::2], parts[2::2]):
        name = subprocess.run(["c++filt", mangled], capture_output=True, text=True).stdout.strip()
        hit = [w for w in wanted if w in name]
        if not hit:
            continue
        ops = collections.Counter()
        for m in re.finditer(r"^\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", body, re.M):
            ops[m.group(1)] += 1
        total = sum(ops.values())
        by_pipe = collections.Counter()
        by_base = collections.Counter()
        for op, c in ops.items():
            by_pipe[pipe(op)] += c
            by_base[op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("IMAD.", "LDG.", "STG.")) and "." in op else "")] += c
        print("## `%s`\n" % name.replace("xsg::(anonymous namespace)::", ""))
        print("%d instructions: %s\n" % (total, ", ".join("%s %d" % kv for kv in by_pipe.most_common())))
        print("| opcode | count | pipe |\n|---|---|---|")
        for op, c in by_base.most_common(28):
            print("| %s | %d | %s |" % (op, c, pipe(op)))
        print()


if __name__ == "__main__":
    main()
