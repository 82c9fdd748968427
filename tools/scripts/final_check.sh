# full GPU test suite, big-scene rows, smoke and the bench line
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python tools/make_c3_scene.py /tmp/c3p > /dev/null 2>&1
python tools/make_c5_scene.py /tmp/c5p width=1920 height=1080 spp=8 > /dev/null 2>&1
echo "C3 $(timeout 200 python tools/render_xml.py /tmp/c3p/c3_path_mis.xml spp=32 repeat=3 normalize=0 | tail -1 | cut -c1-170)"
echo "C5 $(timeout 300 python tools/render_xml.py /tmp/c5p/c5_path_mis.xml spp=8 repeat=3 normalize=0 | tail -1 | cut -c1-170)"
echo "table $(timeout 200 python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=64 repeat=3 normalize=0 | tail -1 | cut -c1-170)"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1


