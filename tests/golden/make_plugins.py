#!/usr/bin/env python
"""Fixtures for the SURVEY 8(f) rank-1/2 rows: disney BSDF, spot + directional emitters, normal maps,
and the normals / av / preview integrators. Run in the build container (needs /root/reference and
oracle/_ref/nori_ref); the GPU box only reads what this writes.

  scenes/plugins/*.xml        small scenes written here (data) over the reference's own plane / cube /
                              cone meshes and normal-map PNG, copied verbatim from scenes/project
  plug_<name>_ref_<spp>.npy   reference CPU render of each scene (`nori_ref render`)
  plug_preview_normal_*.npy   the Normal render layer of the preview scene (`nori_ref render ... normal=`)
  bsdf_<name>_{in,out}.bin    BSDF::eval / pdf / sample records from the reference's plugins (`nori_ref bsdf`)
  emit_<name>_{in,out}.bin    Emitter::sample records for the spot and directional lights (`nori_ref emitter`)
"""
import os
import shutil
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("NORI_REFERENCE", "/root/reference")
NORI_REF = os.path.join(ROOT, "oracle", "_ref", "nori_ref")
DST = os.path.join(HERE, "scenes", "plugins")

CAMERA = """
	<camera type="perspective">
		<transform name="toWorld">
			<rotate angle="30" axis="1,0,0"/>
			<translate value="0,6,-11"/>
		</transform>
		<float name="fov" value="40"/>
		<integer name="width" value="%d"/>
		<integer name="height" value="%d"/>
	</camera>
"""

DISNEY_A = """<bsdf type="disney">
			<color name="albedo" value=".819,.67,.157"/>
			<float name="metallic" value="0.8"/>
			<float name="roughness" value="0.3"/>
			<float name="specular" value="0.7"/>
			<float name="anisotropic" value="0.6"/>
			<float name="specularTint" value="0.4"/>
		</bsdf>"""
DISNEY_B = """<bsdf type="disney">
			<color name="albedo" value=".2,.3,.8"/>
			<float name="subsurface" value="0.6"/>
			<float name="sheen" value="0.9"/>
			<float name="sheenTint" value="0.3"/>
			<float name="clearcoat" value="1"/>
			<float name="clearcoatGloss" value="0.7"/>
			<float name="roughness" value="0.6"/>
		</bsdf>"""
DISNEY_TEX = """<bsdf type="disney">
			<texture type="checkerboard_color" name="albedo">
				<color name="value1" value="0.8,0.8,0.8"/>
				<color name="value2" value="0.25,0.1,0.1"/>
				<vector name="scale" value="0.5,0.5"/>
			</texture>
			<float name="roughness" value="0.8"/>
		</bsdf>"""

GEOMETRY = """
	<shape type="obj">
		<string name="filename" value="meshes/plane.obj"/>
		%(plane_extra)s
		<transform name="toWorld"><scale value="5,5,5"/></transform>
	</shape>
	<shape type="sphere">
		<point name="center" value="0,1,0"/>
		<float name="radius" value="1"/>
		%(sphere_extra)s
	</shape>
	<shape type="obj">
		<string name="filename" value="meshes/cube.obj"/>
		%(cube_extra)s
		<transform name="toWorld"><translate value="2,.5,-2"/></transform>
	</shape>
	<shape type="obj">
		<string name="filename" value="meshes/cone.obj"/>
		%(cone_extra)s
		<transform name="toWorld"><translate value="-2,.5,-3"/></transform>
	</shape>
"""
NORMAL_MAP = """<texture type="png_texture" name="normal">
			<string name="filename" value="res/normal-primitives.png"/>%s
		</texture>"""
LIGHTS_SPOT_DIR = """
	<emitter type="spot">
		<point name="position" value="0,6,-3"/>
		<vector name="direction" value="0.1,-1,0.3"/>
		<color name="power" value="600,500,400"/>
		<float name="falloffstart" value="20"/>
		<float name="totalwidth" value="70"/>
	</emitter>
	<emitter type="directional">
		<color name="radiance" value="40,40,60"/>
		<vector name="direction" value="0.4,-1,0.5"/>
		<float name="angle" value="3"/>
	</emitter>
	<emitter type="envmap">
		<texture type="constant_color" name="albedo"><color name="value" value="0.05,0.05,0.05"/></texture>
	</emitter>
"""
POINT = """
	<emitter type="point">
		<point name="position" value="0,2,-3"/>
		<color name="power" value="1000,1000,1000"/>
	</emitter>
"""


def scene(integrator, size, spp, lights, **extra):
    e = {"plane_extra": "", "sphere_extra": "", "cube_extra": "", "cone_extra": ""}
    e.update(extra)
    return ("<scene>\n\t<sampler type=\"independent\"><integer name=\"sampleCount\" value=\"%d\"/></sampler>\n\t%s\n" % (spp, integrator)
            + CAMERA % size + GEOMETRY % e + lights + "</scene>\n")


VOL_LIGHT = """<medium type="homog">
			<color name="sigma_a" value="0.4 0.4 0.4"/>
			<color name="sigma_s" value="1.5 1.5 1.5"/>%s
			<emitter type="volumelight"><color name="radiance" value="%s"/></emitter>
		</medium>"""
SURFACE_VOL_LIGHT = '<emitter type="volumelight"><color name="radiance" value="0.5 1 0.5"/></emitter>\n\t\t'

W = (160, 120)
SCENES = {
    # name: (xml text, golden spp)
    "disney_spot_dir_path_mis": (scene('<integrator type="path_mis"/>', W, 128, LIGHTS_SPOT_DIR, plane_extra=DISNEY_TEX,
                                       sphere_extra=DISNEY_A, cube_extra=DISNEY_B), 128),
    "disney_spot_dir_direct_mis": (scene('<integrator type="direct_mis"/>', W, 128, LIGHTS_SPOT_DIR, plane_extra=DISNEY_TEX,
                                         sphere_extra=DISNEY_A, cube_extra=DISNEY_B), 128),
    "disney_path_mats_area": (scene('<integrator type="path_mats"/>', W, 128,
                                    '\n\t<shape type="sphere"><point name="center" value="0,5,-2"/><float name="radius" value="1"/>'
                                    '<emitter type="area"><color name="radiance" value="12,12,12"/></emitter></shape>\n',
                                    plane_extra=DISNEY_TEX, sphere_extra=DISNEY_A, cube_extra=DISNEY_B), 128),
    "normalmap_direct_mis": (scene('<integrator type="direct_mis"/>', W, 64, POINT, plane_extra=NORMAL_MAP % "",
                                   sphere_extra=NORMAL_MAP % '\n\t\t\t<float name="intensity" value="0.6"/>', cube_extra=NORMAL_MAP % ""), 64),
    "normalmap_normals": (scene('<integrator type="normals"/>', W, 32, "", plane_extra=NORMAL_MAP % "",
                                sphere_extra=NORMAL_MAP % "", cube_extra=NORMAL_MAP % ""), 32),
    "normals_direction": (scene('<integrator type="normals"><point name="direction" value="0.6,0,0.8"/></integrator>', W, 32, ""), 32),
    "av": (scene('<integrator type="av"><float name="length" value="1.5"/></integrator>', W, 64, ""), 64),
    "preview": (scene('<integrator type="preview"/>', W, 64, LIGHTS_SPOT_DIR, sphere_extra=DISNEY_A), 64),
    # volume lights (volumelight.cpp): inside the ball of a sphere and inside the bounding box of a mesh, seen by both
    # volume integrators; one of them also attached to the shape itself ("with-surface" case of the reference's scenes)
    "volumelight_mats": (scene('<integrator type="path_vol_mats"/>', W, 128, "", sphere_extra=VOL_LIGHT % ("", "3 2 1"),
                               cube_extra=VOL_LIGHT % ('\n\t\t\t<phase type="anisophase"/>', "1 2 4")), 128),
    "volumelight_mis": (scene('<integrator type="path_vol_mis"/>', W, 128, "", sphere_extra=VOL_LIGHT % ("", "3 2 1"),
                              cube_extra=VOL_LIGHT % ('\n\t\t\t<phase type="anisophase"/>', "1 2 4")), 128),
    "volumelight_surface_mis": (scene('<integrator type="path_vol_mis"/>', W, 128, POINT.replace("1000,1000,1000", "60,60,60"),
                                      sphere_extra=SURFACE_VOL_LIGHT + VOL_LIGHT % ("", "3 2 1")), 128),
    "preview_nolight": (scene('<integrator type="preview"/>', W, 16, ""), 16),
}
BSDF_PROBES = {  # name: (scene, shape index)
    "disney_metal": ("disney_spot_dir_path_mis", 1), "disney_coat": ("disney_spot_dir_path_mis", 2),
    "disney_checker": ("disney_spot_dir_path_mis", 0), "diffuse_default": ("normalmap_direct_mis", 3),
}
EMITTER_PROBES = {"spot": ("disney_spot_dir_path_mis", 0), "directional": ("disney_spot_dir_path_mis", 1)}


def run(*args):
    r = subprocess.run([NORI_REF] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    if r.returncode != 0:
        raise RuntimeError(" ".join(args) + "\n" + r.stderr[-2000:])
    return r.stdout.strip().splitlines()[-1]


def main():
    os.makedirs(os.path.join(DST, "meshes"), exist_ok=True)
    os.makedirs(os.path.join(DST, "res"), exist_ok=True)
    proj = os.path.join(REF, "scenes", "project")
    for m in ("plane.obj", "cube.obj", "cone.obj"):
        shutil.copyfile(os.path.join(proj, "meshes", m), os.path.join(DST, "meshes", m))
    shutil.copyfile(os.path.join(proj, "res", "normal-primitives.png"), os.path.join(DST, "res", "normal-primitives.png"))
    for name, (text, spp) in SCENES.items():
        xml = os.path.join(DST, name + ".xml")
        open(xml, "w").write(text)
        extra = ["normal=" + os.path.join(HERE, "plug_%s_normal_%dspp.npy" % (name, spp))] if name == "preview" else []
        print(name, run("render", xml, os.path.join(HERE, "plug_%s_ref_%dspp.npy" % (name, spp)), "spp=%d" % spp, *extra))
    rng = np.random.default_rng(77)
    n = 4096

    def hemi(k):
        v = rng.normal(size=(k, 3)); v /= np.linalg.norm(v, axis=1, keepdims=True)
        v[:, 2] = np.abs(v[:, 2])
        return v
    for name, (sc, index) in BSDF_PROBES.items():
        wi, wo = hemi(n), hemi(n)
        wo[: n // 16, 2] *= -1                                      # some directions below the surface
        wi[n // 16: n // 8, 2] *= -1
        rec = np.concatenate([wi, wo, rng.uniform(0, 1, (n, 2)), rng.uniform(0, 4, (n, 2))], axis=1).astype("<f4")
        fin, fout = os.path.join(HERE, "bsdf_%s_in.bin" % name), os.path.join(HERE, "bsdf_%s_out.bin" % name)
        rec.tofile(fin)
        print(name, run("bsdf", os.path.join(DST, sc + ".xml"), fin, fout, "index=%d" % index))
    for name, (sc, index) in EMITTER_PROBES.items():
        rec = np.concatenate([rng.uniform([-5, 0, -8], [5, 3, 3], (n, 3)), rng.uniform(0, 1, (n, 2))], axis=1).astype("<f4")
        fin, fout = os.path.join(HERE, "emit_%s_in.bin" % name), os.path.join(HERE, "emit_%s_out.bin" % name)
        rec.tofile(fin)
        print(name, run("emitter", os.path.join(DST, sc + ".xml"), fin, fout, "index=%d" % index))


if __name__ == "__main__":
    main()
