#!/usr/bin/env python
"""Joins the per-SASS-instruction execution counts of an ncu report (source page) with the inline call paths that
nvdisasm recovers from the -lineinfo of the SAME build, and prints the share of warp instructions, the average
active threads and the static size per call path -- the table DESIGN.md's "where the instructions go" quotes.

    python profiles/line_profile.py gpurun_out/prof.ncu-rep '_Z18mega_render_kernelILi8EEv6DScene9RenderJob8MegaPlanPf' \
        [depth=3] > profiles/rNN_paths_<kernel>.txt

The library under ray-tracing-c_b200/_build must be the one that was profiled.  Function names come from the nearest
preceding declaration in the source file of each line record: exact for the kernel's own inlined code, approximate
inside the out-of-line functions (bracketed regions), whose bracket label is the symbol nvdisasm prints.
"""
import collections
import csv
import glob
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "ray-tracing-c_b200", "csrc", "device")


def disassemble():
    lib = os.path.join(ROOT, "ray-tracing-c_b200", "_build", "librtc_b200.so")
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
        cubin = glob.glob(os.path.join(tmp, "rtc_mega*.cubin"))[0]
        return subprocess.run(["nvdisasm", "-gi", cubin], capture_output=True, text=True).stdout.split("\n")


_cache = {}


def func_of(path, line):
    key = (path, line)
    if key not in _cache:
        name = os.path.basename(path)
        local = os.path.join(SRC, name)
        if os.path.exists(local):
            src = open(local).read().split("\n")
            for i in range(min(line, len(src)) - 1, -1, -1):
                s = src[i]
                if re.match(r"^(  )?(RTC_DEV|template|static|__device__|__global__|RTC_CHAIN_ATTR)", s) and "(" in s:
                    m = re.search(r"(\w+)\s*\(", s)
                    name = m.group(1) if m else s[:30]
                    break
        _cache[key] = "kernel" if name == "__launch_bounds__" else name
    return _cache[key]


def main():
    report, kern = sys.argv[1], sys.argv[2]
    depth = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    lines = disassemble()
    start = next(i for i, l in enumerate(lines) if l.startswith(".text." + kern + ":"))
    end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith("\t.section")), len(lines))
    ipat = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(.*?);")
    fpat = re.compile(r'File "([^"]+)", line (\d+)')
    keys, cur, blk, region = [], [], [], ""
    lpat = re.compile(r"^\$[^:]*\$(\w+):")
    for l in lines[start:end]:
        m = lpat.match(l)
        if m:  # an out-of-line function (or a libdevice slow path) cloned into this kernel's section
            name = m.group(1)
            for known in ("lights_pdf", "lights_rand", "trace_media", "texture_value_slow", "chain_apply_call",
                          "trace_group_call", "chain_unapply_call", "rcp_rn_f32_slowpath", "sqrt_rn_f32_slowpath",
                          "div_rn_noftz_f32_slowpath"):
                if known in name:
                    name = known
            region, cur, blk = "[" + name + "] ", [], []
            continue
        if "//## File" in l:
            blk.append(l)
            continue
        if ipat.match(l):
            if blk:
                cur = [func_of(f, int(n)) for b in blk for f, n in fpat.findall(b)]
                blk = []
            path = []
            for c in cur:
                if not path or path[-1] != c:
                    path.append(c)
            keys.append(region + " > ".join(path[::-1][:depth]))
    out = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ie, it = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
    data = [r for r in rows[2:] if len(r) > it and r[ie].isdigit()]
    if len(data) != len(keys):
        sys.exit(f"the library is not the profiled build: {len(keys)} SASS instructions here, {len(data)} in the report")
    agg, total = collections.defaultdict(lambda: [0, 0, 0]), 0
    for k, r in zip(keys, data):
        e, t = int(r[ie]), int(r[it])
        agg[k][0] += e
        agg[k][1] += t
        agg[k][2] += 1
        total += e
    print(f"# {report}: {kern}\n# {total / 1e9:.2f} G warp instructions, {len(keys)} SASS instructions")
    print(f"# {'share':>6s} {'threads':>7s} {'static':>6s}  inline path (outermost first)")
    for k, (e, t, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:40]:
        print(f"  {100 * e / total:5.1f}% {t / max(e, 1):7.1f} {n:6d}  {k}")


if __name__ == "__main__":
    main()
