#!/usr/bin/env python
"""Static code size of one kernel by source line: SASS instructions (16 B each) attributed to the outermost and the
second-outermost frame in nb_mega.cu (nvdisasm -gi inline chains), to find what is inlined more than once.

  python tools/sass_static.py libnorib200.so nb_mega <mangled-kernel-substring> [top=15]"""
import os
import sys
from collections import Counter

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ncu_attrib import sass_lines

lib, key, ksub = sys.argv[1:4]
top = int(dict(a.split("=", 1) for a in sys.argv[4:]).get("top", 15))
insts = sass_lines(lib, key, ksub)
print(len(insts), "instructions =", len(insts) * 16 // 1024, "KB")
outer, second, innermost = Counter(), Counter(), Counter()
for _, _, chain in insts:
    mega = [(os.path.basename(f), l) for f, l in chain if f.endswith("nb_mega.cu")]
    outer[mega[-1] if mega else ("?", 0)] += 1
    if len(mega) >= 2:
        second[mega[-2]] += 1
    if chain:
        innermost[(os.path.basename(chain[0][0]), chain[0][1])] += 1
src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "optix-renderer_b200", "csrc", "nb_mega.cu")).read().splitlines()
for title, c in (("outermost frame", outer), ("second frame", second)):
    print("---", title)
    for (f, l), n in c.most_common(top):
        print("%5d  %4d  %s" % (n, l, src[l - 1].strip()[:110] if 0 < l <= len(src) else ""))
