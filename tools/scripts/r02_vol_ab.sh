for lib in optix-renderer_b200/variants/*.so; do echo "== $lib"
for x in cbox_heterog_fluid_mis cbox_heterog_fluid; do NORIB200_LIB=$lib python tools/render_xml.py tests/golden/scenes/vol/$x.xml spp=64 repeat=3 normalize=0 2>&1 | tail -1 | cut -c1-75; done; done
