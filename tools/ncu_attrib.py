#!/usr/bin/env python
"""Attribute an ncu per-instruction profile to source lines of the OUTERMOST frame in a chosen file.

ncu's CSV source page is per SASS instruction; `nvdisasm -gi` gives, for the same instruction
sequence of the same cubin, the inline chain of every instruction. Joining the two by instruction
order tells how many warp-instructions, thread-instructions and stall samples each line of the
kernel body (e.g. each phase of nb_k_tile) is responsible for, everything it inlines included.

  python tools/ncu_attrib.py report.ncu-rep libnorib200.so nb_mega <mangled-kernel-substring> [file=nb_mega.cu] [top=40]
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def sass_lines(lib, cubin_key, kernel_sub):
    d = tempfile.mkdtemp(prefix="cub_")
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, check=True, stdout=subprocess.DEVNULL)
    cub = [f for f in os.listdir(d) if cubin_key in f][0]
    out = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(d, cub)], stdout=subprocess.PIPE, text=True).stdout
    insts, chain, active = [], [], False
    pending = []
    for line in out.splitlines():
        if line.startswith("\t.section"):
            active = ".text." in line and kernel_sub in line
            continue
        if not active:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', line)
        if m:
            pending.append((m.group(1), int(m.group(2))))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            if pending:
                chain = pending
                pending = []
            insts.append((int(m.group(1), 16), m.group(2).strip(), list(chain)))
    return insts


def main():
    rep, lib, cubin_key, kernel_sub = sys.argv[1:5]
    opt = dict(a.split("=", 1) for a in sys.argv[5:])
    want_file = opt.get("file", "nb_mega.cu")
    top = int(opt.get("top", 40))
    insts = sass_lines(lib, cubin_key, kernel_sub)
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    body = [r for r in rows[h + 1:] if len(r) == len(hdr)]
    if len(body) != len(insts):
        print("warning: %d profiled instructions vs %d disassembled" % (len(body), len(insts)), file=sys.stderr)
    col = {n: i for i, n in enumerate(hdr)}
    agg = defaultdict(lambda: [0, 0, 0])
    tot = [0, 0, 0]
    for r, (_, _, chain) in zip(body, insts):
        outer = next(((f, l) for f, l in reversed(chain) if f.endswith(want_file)), ("?", 0))
        v = (int(r[col["Instructions Executed"]]), int(r[col["Thread Instructions Executed"]]), int(r[col["# Samples"]]))
        for k in range(3):
            agg[outer[1]][k] += v[k]
            tot[k] += v[k]
    src = open(next(f for _, _, c in insts for f, _ in c if f.endswith(want_file))).read().splitlines() if os.path.exists("/root/repo") else []
    print("line   warp-inst%%  thr/warp  samples%%   source (%s), totals: %.3g warp-inst, %.3g thread-inst, %d samples" % (want_file, tot[0], tot[1], tot[2]))
    for line, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
        text = src[line - 1].strip()[:90] if 0 < line <= len(src) else ""
        print("%4d   %6.2f     %5.1f    %6.2f    %s" % (line, 100.0 * v[0] / tot[0], v[1] / max(1, v[0]), 100.0 * v[2] / max(1, tot[2]), text))


if __name__ == "__main__":
    main()
