# big tree scenes: C3 (1.31 M triangles) and a reduced C5 (1.45 M triangles) under both strategies
python tools/make_c5_scene.py /tmp/c5p width=1920 height=1080 spp=8 > /dev/null
python tools/make_c3_scene.py /tmp/c3p > /dev/null 2>&1
ls /tmp/c3p | head -3
for st in tile wavefront; do
  echo "== strategy $st"
  python tools/render_xml.py /tmp/c5p/c5_path_mis.xml spp=8 device_bvh=1 repeat=2 normalize=0 strategy=$st | tail -1 | cut -c1-160
  python tools/render_xml.py /tmp/c5p/c5_path_mis.xml spp=8 repeat=2 normalize=0 strategy=$st | tail -1 | cut -c1-160
  python tools/render_xml.py /tmp/c3p/c3_path_mis.xml spp=16 device_bvh=1 repeat=2 normalize=0 strategy=$st | tail -1 | cut -c1-160
  python tools/render_xml.py /tmp/c3p/c3_path_mis.xml spp=16 repeat=2 normalize=0 strategy=$st | tail -1 | cut -c1-160
done
