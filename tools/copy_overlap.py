#!/usr/bin/env python
"""Token-shingle overlap of every package file against the same-named reference file.

Method (the one the round-1 review used): tokenise with ``tokenize``, drop comments, docstrings and layout
tokens, build the set of 10-token shingles of both files, report |ours ∩ ref| / |ours|.  Blind to re-wrapping
and docstring stripping, so it measures what is left of the reference's statement sequence.

    python tools/copy_overlap.py [--threshold 0.35] [--ref /root/reference/src/accelerate]
"""

from __future__ import annotations

import argparse
import io
import os
import sys
import tokenize

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "accelerate_hpc_test_b200")
SKIP = {tokenize.COMMENT, tokenize.NL, tokenize.NEWLINE, tokenize.INDENT, tokenize.DEDENT, tokenize.ENCODING, tokenize.ENDMARKER}


def tokens(path: str) -> list[str]:
    out = []
    try:
        src = open(path, "rb").read()
        prev_sig = None
        for tok in tokenize.tokenize(io.BytesIO(src).readline):
            if tok.type in SKIP:
                if tok.type in (tokenize.NEWLINE, tokenize.INDENT, tokenize.DEDENT):
                    prev_sig = "stmt"
                continue
            if tok.type == tokenize.STRING and prev_sig in (None, "stmt"):
                prev_sig = "stmt"  # an expression statement that is just a string: docstring
                continue
            prev_sig = "tok"
            out.append(tok.string)
    except (tokenize.TokenError, SyntaxError, OSError):
        pass
    return out


def shingles(toks: list[str], n: int = 10) -> set:
    return {tuple(toks[i : i + n]) for i in range(max(0, len(toks) - n + 1))}


def overlap(ours: str, ref: str) -> float:
    a, b = shingles(tokens(ours)), shingles(tokens(ref))
    return len(a & b) / len(a) if a else 0.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference/src/accelerate")
    ap.add_argument("--threshold", type=float, default=0.35)
    ap.add_argument("files", nargs="*")
    args = ap.parse_args()
    rows = []
    for dirpath, _, names in os.walk(PKG):
        for nm in names:
            if not nm.endswith(".py"):
                continue
            p = os.path.join(dirpath, nm)
            rel = os.path.relpath(p, PKG)
            if args.files and rel not in args.files:
                continue
            r = os.path.join(args.ref, rel)
            if not os.path.exists(r):
                continue
            rows.append((overlap(p, r), rel, sum(1 for _ in open(p))))
    rows.sort(reverse=True)
    bad = 0
    for ov, rel, n in rows:
        flag = "  <-- over" if ov >= args.threshold else ""
        bad += ov >= args.threshold
        print(f"{ov:5.2f}  {n:5d}  {rel}{flag}")
    print(f"{bad} file(s) at or above {args.threshold}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
