python tools/make_c3_scene.py /tmp/c3p > /dev/null 2>&1
for f in 0 1; do echo "== fast_hits=$f"
for x in tests/golden/scenes/cbox/cbox_path_mis.xml tests/golden/scenes/cbox/cbox_path_mats.xml tests/golden/scenes/vol/cbox_val_mis.xml tests/golden/scenes/table/table_path_mis.xml tests/golden/scenes/veach/veach_mis.xml /tmp/c3p/c3_path_mis.xml; do
python tools/render_xml.py $x spp=64 repeat=3 fast_hits=$f 2>&1 | tail -2 | cut -c1-72 | tr '\n' ' '; echo; done; done
python -m pytest tests -m gpu -q -x --deselect tests/test_c3_gpu.py 2>&1 | tail -4
