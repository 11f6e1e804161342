#!/usr/bin/env python
"""Static SASS statistics of one kernel from `nvdisasm -g -c <cubin>` output (no GPU needed).

usage: sass_static.py one.dis [kernel-substring] [--lines] [--files afx_step.cuh,fft_gen.cuh,...]

Prints the opcode mix of the whole kernel and, per source file, instruction counts; with --lines the
instructions attributed to each source line (the -lineinfo mapping), largest first.  The warp step is
straight-line code, so the static count of its lines is what one step executes (the bounds-checked
load variant is listed separately by its own lines)."""
import collections
import re
import sys


def parse(path, kernel_sub=None):
    cur_kernel, cur_file, cur_line = None, None, None
    rows = []
    for ln in open(path, errors="replace"):
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            cur_kernel = m.group(1)
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur_file, cur_line = m.group(1).split("/")[-1], int(m.group(2))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", ln)
        if m and cur_kernel and (kernel_sub is None or kernel_sub in cur_kernel):
            ins = m.group(2).strip()
            toks = ins.split()
            op = toks[1] if toks[0].startswith("@") else toks[0]
            rows.append((cur_kernel, cur_file, cur_line, op, ins))
    return rows


def opclass(op):
    base = op.split(".")[0]
    if base in ("LDS", "STS", "LDG", "STG", "LDL", "STL", "SHFL", "ATOMG", "ATOMS", "RED"):
        w = [p for p in op.split(".")[1:] if p in ("64", "128", "U16", "U8")]
        return base + ("." + w[0] if w else "")
    return base


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    flags = [a for a in sys.argv[1:] if a.startswith("--")]
    rows = parse(args[0], args[1] if len(args) > 1 else None)
    files = None
    for f in flags:
        if f.startswith("--files="):
            files = set(f.split("=", 1)[1].split(","))
    if files:
        rows = [r for r in rows if r[1] in files]
    print(f"instructions: {len(rows)}")
    mix = collections.Counter(opclass(r[3]) for r in rows)
    print("opcode mix:", ", ".join(f"{k} {v}" for k, v in mix.most_common(40)))
    byfile = collections.Counter(r[1] for r in rows)
    print("by file:", ", ".join(f"{k} {v}" for k, v in byfile.most_common()))
    if "--lines" in flags:
        byline = collections.Counter((r[1], r[2]) for r in rows)
        for (f, l), c in byline.most_common(60):
            ops = collections.Counter(opclass(r[3]) for r in rows if r[1] == f and r[2] == l)
            print(f"  {f}:{l:4d} {c:5d}  " + " ".join(f"{k}:{v}" for k, v in ops.most_common(8)))


if __name__ == "__main__":
    main()
