/* oracle/ref_vdb_stub.cpp — TEST INFRASTRUCTURE. src/textures/NvdbVolume.vdb.cpp needs OpenVDB
 * (absent in this image); the one symbol the rest of the reference links against
 * is NvdbVolume::loadFromFile. Volume grids are not loadable through the reference here. */
#include <nori/NvdbVolume.h>
NORI_NAMESPACE_BEGIN
void NvdbVolume::loadFromFile() {
    throw NoriException("oracle harness: OpenVDB/NanoVDB file loading is unavailable (%s)", filename.string());
}
NORI_NAMESPACE_END
