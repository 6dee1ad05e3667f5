#!/usr/bin/env python
"""tools/sass_excerpt.py -- `cuobjdump -sass` evidence for profiles/: per kernel, the target arch, register
count, an opcode histogram and the memory / TMA / barrier instructions in program order.  Runs in the build
container (no GPU): python tools/sass_excerpt.py > profiles/r2_sass_excerpts.txt"""
import collections
import re
import subprocess
import sys

LIB = "pint_b200/lib/libpint_cuda.so"
WANT = [
    ("C1 op pair, RGBA8 (bench.py's timed kernel)", r"binary_pair_stream256_kernel.*pair_fnILi1ELi2E.*static_layoutIjJLm8ELm8ELm8ELm8E"),
    ("C2 single op, u64 16x4 add_signed_saturate", r"binary_stream256_kernel.*swar_fnILi2E.*static_layoutImJLm16ELm16ELm16ELm16E"),
    ("C4 fused unpack->add_signed_saturate->repack, irregular u64", r"binary_stream256_kernel.*fused_fnILi2E.*static_layoutImJLm3ELm7ELm11ELm13ELm17ELm13E"),
    ("C3 unpack to int32 planes, staged + bulk (TMA) stores", r"unpack_bulk_kernelImLb1E.*static_layoutImJLm3ELm7ELm11ELm13ELm17ELm13E.*Li2048ELi256E"),
    ("C3 repack from int32 planes", r"repack_quad_kernelImLi4E.*static_layoutImJLm3ELm7ELm11ELm13ELm17ELm13E"),
    ("dense RGB24 add_signed_saturate (TMA-staged tiles)", r"dense_binary_kernel.*swar_fnILi2E.*Lm8ELm8ELm8E"),
    ("dense RGB24 -> three uint8 planes (byte permutes)", r"dense_bytes_to_planes_kernelILi3ELi1E"),
    ("dense shift as a masked funnel shift over the word stream", r"dense_shift_stream_kernelILi4E"),
    ("the reference benchmark's pair reduction, type-2 layout, add_signed_saturate", r"pairs_reduce_kernelIjLi2ELb1E.*static_layoutIjJLm1ELm2ELm3ELm4ELm5ELm6ELm11E"),
]
MEM = re.compile(r"\b(LDG|STG|LDS|STS|UBLKCP|UTMA\w*|SYNCS|BAR|FENCE|ATOM|RED|LDC|LDGSTS|PRMT|SHFL|ACQBULK|CCTL)\b")


def main():
    text = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", text)))
    print(f"{LIB}: code for {arch}\n")
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    regs = {}
    cur = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+)", line)
        if m and cur:
            regs[cur] = int(m.group(1))
    funcs = re.split(r"\n\s*Function : ", text)
    for title, pat in WANT:
        hit = None
        for f in funcs[1:]:
            name = f.split("\n", 1)[0].strip()
            if re.search(pat, name):
                hit = (name, f)
                break
        print("=" * 100)
        print(title)
        if not hit:
            print("   (no kernel matches %s)\n" % pat)
            continue
        name, body = hit
        ops = re.findall(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", body)
        print("   " + name[:180])
        print("   registers/thread: %s; instructions: %d" % (regs.get(name, "?"), len(ops)))
        hist = collections.Counter(o for o in ops)
        print("   opcode histogram: " + ", ".join(f"{o} x{c}" for o, c in hist.most_common(24)))
        seq = [o for o in ops if MEM.search(o.split(".")[0]) and not o.startswith("LDC")]
        print("   memory / TMA / barrier / permute instructions in order:")
        line, out = "      ", []
        for o in seq[:80]:
            if len(line) + len(o) > 110:
                out.append(line)
                line = "      "
            line += o + " "
        out.append(line)
        print("\n".join(out))
        print()


if __name__ == "__main__":
    sys.exit(main())
