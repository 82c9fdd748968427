/* nb_scene_file.h — the ".nbs" container: a flattened scene (the arrays of nb_scene_desc plus
 * the BVH and render settings) as named sections, so fixtures can be written by the host
 * loader, by the oracle harness, and read from Python with numpy alone.
 *
 *   u32 magic 'NBS1' | u32 n_sections | n_sections x nbs_section | payload (16-byte aligned)
 *
 * Section names used: "shapes" (nb_shape), "positions" "normals" "uvs" "tangents" "bitangents" (f32), "faces" (u32),
 * "bsdfs" (nb_bsdf), "emitters" (nb_emitter), "emitter_cdf" "area_cdf" (f32), "camera"
 * (nb_camera), "render" (nb_render_params), "spp" (u32), "bvh_nodes" (nb_bvh_node),
 * "bvh_indices" (u32). A section that does not apply is simply absent.
 */
#ifndef NB_SCENE_FILE_H
#define NB_SCENE_FILE_H
#include <stdint.h>

#define NBS_MAGIC 0x3153424eu /* "NBS1" little endian */

typedef struct nbs_section {
    char     name[16];
    uint32_t elem_size;   /* bytes per element */
    uint32_t reserved;
    uint64_t count;       /* elements */
    uint64_t offset;      /* from file start */
} nbs_section;            /* 40 bytes */

#endif
