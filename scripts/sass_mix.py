#!/usr/bin/env python
"""SASS instruction mix of dumped specialised kernels (development tool, no GPU needed).

    python scripts/sass_mix.py gpurun_out/jit_dump/k_pass_*.cubin

Prints per kernel: registers, total instructions and the share of each class (FP64 pipe, shared-memory,
global-memory / async copy, integer + logic, moves / selects, control).  Straight-line kernels: the static
mix of the tile loop is, up to the prologue, the dynamic mix."""
import collections
import re
import subprocess
import sys

CLASSES = [
    ("fp64", r"^(DFMA|DMUL|DADD|DSETP|DMNMX)"),
    ("lds/sts", r"^(LDS|STS|LDSM)"),
    ("global", r"^(LDG|STG|LDGSTS|LDGDEPBAR|DEPBAR|UBLKCP|UTMA|SYNCS|LD\.|ST\.|LDC|ULDC|LDCU)"),
    ("int/logic", r"^(IADD|IMAD|LOP|SHF|LEA|POPC|ISETP|SEL|PRMT|I2F|F2I|IABS|FLO|BREV|UIADD|ULOP|USHF|ULEA|UISETP|USEL|UPOPC|UIMAD|BMSK|SGXT|PLOP|UPLOP|R2P|P2R|VOTE)"),
    ("mov", r"^(MOV|UMOV|FSEL|CS2R|S2R|S2UR|R2UR|SHFL|IMAD\.MOV)"),
    ("control", r"^(BRA|BAR|EXIT|BSYNC|BSSY|WARPSYNC|NOP|CALL|RET|MEMBAR|FENCE|ERRBAR|YIELD|UBRA)"),
]


def mix(path):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    regs = subprocess.run(["cuobjdump", "-res-usage", path], capture_output=True, text=True).stdout
    m = re.search(r"REG:(\d+)", regs)
    sm = re.search(r"SHARED:(\d+)", regs)
    cnt = collections.Counter()
    ops = collections.Counter()
    total = 0
    for line in out.splitlines():
        mm = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not mm:
            continue
        op = mm.group(1)
        total += 1
        ops[op.split(".")[0]] += 1
        for name, pat in CLASSES:
            if re.match(pat, op):
                cnt[name] += 1
                break
        else:
            cnt["other"] += 1
    return int(m.group(1)) if m else -1, total, cnt, ops


def main():
    for path in sys.argv[1:]:
        regs, total, cnt, ops = mix(path)
        parts = ", ".join(f"{k} {v} ({100.0 * v / max(1, total):.0f}%)" for k, v in cnt.most_common())
        print(f"{path.split('/')[-1]}: {regs} regs, {total} instr: {parts}")
        if "-v" in sys.argv:
            print("   top ops:", ", ".join(f"{k} {v}" for k, v in ops.most_common(14)))


if __name__ == "__main__":
    main()
