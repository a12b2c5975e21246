#!/usr/bin/env python3
"""Generates collision tables in the format of src/cross_section.txt (10 000 rows) -- used for the synthetic
supercritical tables of BASELINE cfg3.  src/cross_section.txt itself is the REFERENCE'S FILE, byte for byte, and is not
produced by this script: regenerating it gives the same fp32 values (pinned by tests/test_table.py) but 16 rows differ in
the last printed digit with this libm -- only the fp32 equality is guaranteed.

Row i holds `split_chance remove_chance` in percent, both equal to
(sin(i**2.6 / 5e7) + 1.01) * i**0.1 -- the formula the upstream project documents
for its shipped table (reference: cross_section_gen.py:6-7).  The file is *input
data*: the engine parses it into fp32, and tests/test_table.py pins the SHA-256 of
that fp32 image (f1b1701b...) so a libm that rounds the last decimal digit
differently cannot silently change results.

Usage: python data/make_cross_section.py [out_path] [--supercritical SPLIT REMOVE]
"""
import math
import sys


def rows():
    for i in range(10000):
        v = (math.sin(pow(i, 2.6) / 50000000) + 1.01) * pow(i, 0.1)
        yield v, v


def main(argv):
    out = "src/cross_section.txt"
    const = None
    args = list(argv[1:])
    if "--supercritical" in args:
        k = args.index("--supercritical")
        const = (float(args[k + 1]), float(args[k + 2]))
        del args[k:k + 3]
    if args:
        out = args[0]
    with open(out, "w") as f:
        for s, r in rows():
            if const is not None:
                s, r = const
            f.write(f"{s} {r}\n")


if __name__ == "__main__":
    main(sys.argv)
