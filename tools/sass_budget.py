#!/usr/bin/env python3
"""Per-section instruction budget of the fused kernel's step loop, from the SASS of the built library.

usage: sass_budget.py [kernel-substring] [--all-loops]
  cuobjdump -xelf + nvdisasm -g give every SASS instruction its source line (-lineinfo build).  The step loop of
  lockstep_push (csrc/pic_mover.cuh) is found as the innermost backward-branch loop that contains the leapfrog lines;
  its instructions are attributed to sections by source line.  No GPU needed.
"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "particle-simulation_b200", "libpicb200.so")
MOVER = "pic_mover.cuh"
DEVICE = "pic_device.cuh"


def source_sections():
    """line ranges of the sections, located by marker text so that the tool survives edits of the source"""
    mv = open(os.path.join(ROOT, "particle-simulation_b200", "csrc", MOVER)).read().splitlines()
    dv = open(os.path.join(ROOT, "particle-simulation_b200", "csrc", DEVICE)).read().splitlines()

    def find(lines, text, start=0):
        for i in range(start, len(lines)):
            if text in lines[i]:
                return i + 1
        raise SystemExit(f"marker not found: {text}")

    f0 = find(mv, "__device__ __forceinline__ void lockstep_push(")
    loop = find(mv, "for (int t = t_first; t <= c.pt; ++t)", f0)
    sec = {}
    sec["philox"] = [(MOVER, find(mv, "if ((t != t_first && (g & 3u) == 0u)", loop), find(mv, "refresh = false;", loop) + 1),
                     (DEVICE, find(dv, "uint4 philox4x32_10("), find(dv, "return c;", find(dv, "uint4 philox4x32_10(uint4 c, const PhiloxKeys& ks)")))]
    sec["leapfrog"] = [(MOVER, find(mv, "const bool act = kStaged", loop), find(mv, "if (!kStaged || act)", loop))]
    sec["bounds"] = [(MOVER, find(mv, "const bool oob =", loop), find(mv, "const bool oob =", loop))]
    sec["draw"] = [(MOVER, find(mv, "const uint32_t w01 =", loop), find(mv, "const float r = __uint2float_rn", loop))]
    e0 = find(dv, "int energy_row_estimate(float e")
    r0 = find(mv, "float2 collision_row(const MoverCtx& c, float en)")
    sec["energy + table row"] = [(MOVER, find(mv, "const float en =", loop), find(mv, "const float2 th =", loop)), (DEVICE, e0, find(dv, "return j - ((e < s_edge[j])", e0)),
                                 (MOVER, r0, find(mv, "return c.thr[idx];", r0))]
    sec["collide / die / vote"] = [(MOVER, find(mv, "const bool coll =", loop), find(mv, "unsigned m = __ballot_sync(kFull, spawn);", loop) + 1)]
    sec["loop control"] = [(MOVER, loop - 1, loop + 1)]
    end = find(mv, "if (kAdopt) {\n".strip(), find(mv, "const uint32_t flip =", loop))
    # the vote "somebody spawned" is rarely taken: the staging of children, and the overflow path inlined into it
    sp0 = find(mv, "__device__ __forceinline__ int reserve_slots(")
    sp1 = find(mv, "// Push one lane's electron through mobility steps")
    sec["spawn path (cold)"] = [(MOVER, find(mv, "bool stage_it = spawn;", loop), end), (MOVER, sp0, sp1), ("device_atomic_functions.hpp", 0, 10**9),
                                ("sm_32_intrinsics.hpp", 0, 10**9), ("sm_30_intrinsics.hpp", 0, 10**9)]
    return sec, (loop, end)


def disassemble(kernel_filter):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, check=True, capture_output=True)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    funcs, cur, name = {}, None, None
    for ln in txt.splitlines():
        m = re.match(r"^\.text\.(\S+):", ln)
        if m:
            name = m.group(1)
            cur = funcs.setdefault(name, [])
            continue
        if ln.startswith("//-") or ln.startswith("\t.section"):
            if ln.startswith("//-"):
                cur = None
            continue
        if cur is not None:
            cur.append(ln)
    return {k: v for k, v in funcs.items() if kernel_filter in k}


def parse(lines):
    """-> list of (kind, payload): ('label', name) | ('inst', (text, file, line))"""
    out, f, l = [], None, None
    for ln in lines:
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            f, l = os.path.basename(m.group(1)), int(m.group(2))
            continue
        m = re.match(r"^(\.L_x_\d+):", ln)
        if m:
            out.append(("label", m.group(1)))
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(.*?);", ln)
        if m:
            out.append(("inst", (m.group(1).strip(), f, l)))
    return out


def loops(items):
    pos = {p: i for i, (k, p) in enumerate(items) if k == "label"}
    res = []
    for i, (k, p) in enumerate(items):
        if k == "inst":
            m = re.search(r"BRA\S*\s.*`\((\.L_x_\d+)\)", p[0])
            if m and m.group(1) in pos and pos[m.group(1)] < i:
                res.append((pos[m.group(1)], i))
    return res


def main():
    flt = next((a for a in sys.argv[1:] if not a.startswith("--")), "k_poisson_fusedILb0ELi0")
    sec, (loop_line, end_line) = source_sections()
    funcs = disassemble(flt)
    for name, lines in funcs.items():
        items = parse(lines)
        n_inst = sum(1 for k, _ in items if k == "inst")
        print(f"== {name}: {n_inst} SASS instructions")
        cand = []
        for a, b in loops(items):
            body = [p for k, p in items[a:b + 1] if k == "inst"]
            lf = sum(1 for t, f, l in body if f == MOVER and any(lo <= (l or 0) <= hi for ff, lo, hi in sec["leapfrog"] if ff == MOVER))
            if lf >= 9:
                cand.append((len(body), a, b, body))
        cand.sort()
        seen = 0
        for size, a, b, body in cand:
            # innermost loops only: skip loops that contain an already reported one
            if any(a <= a2 and b2 <= b and (a, b) != (a2, b2) for _, a2, b2, _ in cand[:seen]):
                continue
            seen += 1
            counts = collections.Counter()
            other = collections.Counter()
            for t, f, l in body:
                hit = None
                for s, ranges in sec.items():
                    if any(f == ff and lo <= (l or 0) <= hi for ff, lo, hi in ranges):
                        hit = s
                        break
                if hit is None:
                    hit = "other"
                    other[(f, l)] += 1
                counts[hit] += 1
            hot = size - counts["spawn path (cold)"]
            print(f"  step loop at items {a}..{b}: {size} instructions, {hot} outside the spawn path")
            for s in list(sec) + ["other"]:
                if counts[s]:
                    print(f"    {s:24s} {counts[s]:4d}")
            if other:
                print("    other, by line:", ", ".join(f"{f}:{l} x{n}" for (f, l), n in other.most_common(8)))
            ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for t, f, l in body
                                      if not any(f == ff and lo <= (l or 0) <= hi for ff, lo, hi in sec["spawn path (cold)"]))
            print("    opcodes outside the spawn path:", ", ".join(f"{k} {v}" for k, v in ops.most_common(14)))
            if "--all-loops" not in sys.argv and seen >= 3:
                break


if __name__ == "__main__":
    main()
