/* b200state.h — the backend object with OptixState's call shape
 * (include/nori/optix/OptixState.h:88-109), forwarding to the C ABI of include/norib200.h. */
#pragma once
#include "plugins.h"

namespace nori {

class B200State {
public:
    explicit B200State(int device = 0);
    /* several GPUs of one box behind the same object (nb_create_multi): every device holds the scene, renders the
     * sample slice (stride n, offset g) of each subframe, and download() sums the films on the first device with one
     * NCCL reduce (nb_film_reduce) */
    explicit B200State(const std::vector<int> &devices);
    ~B200State();
    B200State(const B200State &) = delete;
    /* OptixState::preRender (OptixState.render.cpp:19-85): flatten + upload + BVH + film setup.
     * false = the scene uses a plugin without a GPU kernel (the caller renders on its CPU path). */
    bool preRender(const Scene &scene, uint64_t seed = 0, uint32_t poolPaths = 0, uint32_t layers = 0);   /* layers: NB_LAYER_* bits */
    /* which tree the GPUs traverse: NB_BVH_DEVICE_SAH (default: the reference's SAH criterion built on the device in
     * milliseconds, no host build at all), NB_BVH_DEVICE (linear BVH), or NB_BVH_IMPORT (the reference's own tree from the
     * host builder — what the bit-parity mode of the traversal needs). NB_BVH_BUILDER=host|lbvh|sah overrides. */
    void setTreeBuilder(int mode) { m_treeMode = mode; }
    /* every filter weight 1 over the pixel's own square: a pixel of the film is then the plain mean of its samples
     * (what the t-test harness averages, src/utils/ttest.cpp:205-222) */
    void setBoxFilter(bool on) { m_boxFilter = on; }
    /* OptixState::renderSubframe (OptixState.cpp:485-510) */
    void renderSubframe(uint32_t currentSample, uint32_t samplesPerLaunch, uint32_t stride = 1, uint32_t offset = 0);
    /* one round of the sample loop for an `adaptive` sampler (render.cpp:283-360 with :323-335 inside): the blocks' variance
     * loops on the device. One device: the feedback needs a block's own samples in one place. */
    nb_adaptive_info renderAdaptiveRound(uint32_t blockSize, uint32_t initialUniform, uint32_t roundCap = 0);
    /* accumulators in ImageBlock layout (rows x cols x {sum w*rgb, sum w}) or the normalised bitmap */
    void download(std::vector<float> &out, bool normalize, uint32_t *rows = nullptr, uint32_t *cols = nullptr, int layer = NB_LAYER_COMPOSITE);
    void clear();                                   /* OptixState::clear */
    nb_counters counters() const;
    const FlatScene &flat() const { return m_flat; }
    const std::string &whyUnsupported() const { return m_why; }
    nb_ctx *handle() const { return m_ctx; }
    size_t deviceCount() const { return m_all.size(); }
private:
    void check(int rc) const;
    nb_ctx *m_ctx = nullptr;                        /* the root: m_all[0] */
    std::vector<nb_ctx *> m_all;
    FlatScene m_flat;
    std::string m_why;
    int m_treeMode = NB_BVH_DEVICE_SAH;
    bool m_boxFilter = false;
};

} // namespace nori
