#!/usr/bin/env python
"""Per-kernel SASS evidence for the built library: which Blackwell-only instructions each kernel contains.

    python tools/sass_evidence.py > profiles/r2/sass_per_kernel.txt

Runs ``cuobjdump -sass`` on ``accelerate_hpc_test_b200/lib/libb200.so`` (no GPU needed) and prints, for every kernel, the counts of
the mnemonics that prove the hardware path: UTCHMMA / UTCQMMA (tcgen05.mma f16 / f8 incl. block scaled), .2CTA (cta_group::2),
UTCCP (tcgen05.cp), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG (TMA tensor load / store), UBLKCP / UBLKRED (cp.async.bulk /
cp.reduce.async.bulk), UTCBAR (tcgen05.commit), LDGMC.*.HPADD (multimem.ld_reduce: the NVSwitch adds; multimem.st is a plain STG.E.128.STRONG.SYS to the multicast mapping), SYNCS (mbarrier), ELECT.
"""
import collections
import hashlib
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "accelerate_hpc_test_b200", "lib", "libb200.so")
WATCH = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCOMMA|UTCCP|LDTM|STTM|UTMALDG|UTMASTG|UTMAREDG|UBLKCP|UBLKRED|UTCBAR|UTCATOMSWS|LDGMC|STG\.E\.\w*MULTIMEM|REDG?MC|"
                   r"MULTIMEM|SYNCS|ELECT|UCGABAR_ARV|UCGABAR_WAIT|HMMA|MUFU\.EX2)[\w.]*")


def main():
    sha = hashlib.sha256(open(LIB, "rb").read()).hexdigest()
    text = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    current = None
    for line in text.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            current = m.group(1)
            kernels[current] = collections.Counter()
            continue
        if current is None:
            continue
        for hit in WATCH.finditer(line.split("/*")[1] if line.count("/*") >= 1 and ";" in line else ""):
            name = hit.group(0)
            # fold operand-ish suffixes: keep the op and its structural modifiers
            kernels[current][name] += 1
    demangled = {}
    try:
        out = subprocess.run(["cu++filt"] + list(kernels), capture_output=True, text=True, check=True).stdout.splitlines()
        demangled = dict(zip(kernels, out))
    except Exception:
        pass
    print(f"# {os.path.relpath(LIB, ROOT)}  sha256 {sha}")
    print(f"# {len(kernels)} kernels; counts of Blackwell / NVLS mnemonics per kernel (cuobjdump -sass)")
    for mangled, counts in kernels.items():
        name = demangled.get(mangled, mangled)
        name = re.sub(r"\((int|bool)\)", "", name)
        name = re.sub(r"\(.*", "", name)[:110]
        if not counts:
            print(f"{name}: -")
            continue
        print(f"{name}: " + ", ".join(f"{k} x{v}" for k, v in sorted(counts.items(), key=lambda kv: (-kv[1], kv[0]))))


if __name__ == "__main__":
    sys.exit(main())
