#!/usr/bin/env python
"""Build the device SAH tree of the C3 scene once (for an ncu launch list of the builder's kernels):
   python tools/sah_build_once.py /path/to/c3_path_mis.xml"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import norib200 as nb
s = nb.load_xml(sys.argv[1], host_bvh=False)
ctx = nb.Context(0)
ctx.upload_scene(s)
ctx.build_bvh_device(sah=True)
print(ctx.bvh_stats())
