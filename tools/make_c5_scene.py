#!/usr/bin/env python
"""C5 of SURVEY 8(d): the synthetic multi-mesh scene of the multi-GPU configuration.

64 copies of the reference's scenes/pa1/camelhead.obj (22 704 triangles each -> 1 453 056) on an
8 x 8 grid through `toWorld` translations (baked into the vertices by the OBJ loader, obj.cpp:77),
BSDF cycle {diffuse, microfacet alpha 0.1, dielectric, mirror} by copy index, a floor quad, four
quad area lights (radiance 10) and one point light (power 50), path_mis, 3840 x 2160, 4096 spp,
gaussian filter (the default) or box (filter=box).

    python tools/make_c5_scene.py out_dir [width=3840 height=2160 spp=4096 filter=gaussian copies=8]
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CAMEL = os.path.join(HERE, "..", "tests", "golden", "scenes", "c5", "camelhead.obj")

BSDFS = ['<bsdf type="diffuse"><color name="albedo" value="0.65 0.45 0.3"/></bsdf>',
         '<bsdf type="microfacet"><float name="alpha" value="0.1"/><color name="kd" value="0.2 0.35 0.5"/></bsdf>',
         '<bsdf type="dielectric"/>',
         '<bsdf type="mirror"/>']
QUAD = "v -1 0 -1\nv 1 0 -1\nv 1 0 1\nv -1 0 1\nf 1 3 2\nf 1 4 3\n"          # normal +y
QUAD_DOWN = "v -1 0 -1\nv 1 0 -1\nv 1 0 1\nv -1 0 1\nf 1 2 3\nf 1 3 4\n"     # normal -y (lights face the floor)


def main():
    out = sys.argv[1]
    opt = dict(a.split("=", 1) for a in sys.argv[2:])
    w, h, spp = int(opt.get("width", 3840)), int(opt.get("height", 2160)), int(opt.get("spp", 4096))
    n = int(opt.get("copies", 8))
    os.makedirs(out, exist_ok=True)
    shutil.copyfile(CAMEL, os.path.join(out, "camelhead.obj"))
    open(os.path.join(out, "quad_up.obj"), "w").write(QUAD)
    open(os.path.join(out, "quad_down.obj"), "w").write(QUAD_DOWN)
    pitch = 0.8
    half = (n - 1) * pitch / 2
    shapes = []
    for i in range(n):
        for j in range(n):
            k = i * n + j
            shapes.append('\t<shape type="obj">\n\t\t<string name="filename" value="camelhead.obj"/>\n'
                          '\t\t<transform name="toWorld"><translate value="%g,0.33,%g"/></transform>\n\t\t%s\n\t</shape>'
                          % (j * pitch - half, i * pitch - half, BSDFS[k % 4]))
    ext = half + 1.5
    shapes.append('\t<shape type="obj">\n\t\t<string name="filename" value="quad_up.obj"/>\n'
                  '\t\t<transform name="toWorld"><scale value="%g,1,%g"/></transform>\n'
                  '\t\t<bsdf type="diffuse"><color name="albedo" value="0.7 0.7 0.7"/></bsdf>\n\t</shape>' % (ext, ext))
    for sx, sz in ((-1, -1), (1, -1), (-1, 1), (1, 1)):
        shapes.append('\t<shape type="obj">\n\t\t<string name="filename" value="quad_down.obj"/>\n'
                      '\t\t<transform name="toWorld"><scale value="0.6,1,0.6"/><translate value="%g,2.5,%g"/></transform>\n'
                      '\t\t<emitter type="area"><color name="radiance" value="10 10 10"/></emitter>\n\t</shape>' % (sx * half * 0.55, sz * half * 0.55))
    rf = '\t\t<rfilter type="box"/>\n' if opt.get("filter", "gaussian") == "box" else ""
    xml = ('<?xml version="1.0" encoding="utf-8"?>\n<scene>\n\t<integrator type="path_mis"/>\n'
           '\t<sampler type="independent"><integer name="sampleCount" value="%d"/></sampler>\n'
           '\t<camera type="perspective">\n\t\t<float name="fov" value="42"/>\n'
           '\t\t<transform name="toWorld"><lookat target="0,0.2,0" origin="%g,%g,%g" up="0,1,0"/></transform>\n'
           '\t\t<integer name="width" value="%d"/>\n\t\t<integer name="height" value="%d"/>\n%s\t</camera>\n'
           '\t<emitter type="point"><point name="position" value="0,3.5,0"/><color name="power" value="50 50 50"/></emitter>\n'
           '%s\n</scene>\n') % (spp, 0.0, half * 1.6 + 1.5, half * 2.2 + 2.5, w, h, rf, "\n".join(shapes))
    name = "c5_path_mis%s.xml" % ("_box" if rf else "")
    open(os.path.join(out, name), "w").write(xml)
    print(os.path.join(out, name), "triangles", n * n * 22704 + 2 + 8)


if __name__ == "__main__":
    main()
