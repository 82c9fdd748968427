/* oracle/ref_main.cpp — TEST INFRASTRUCTURE, not product.
 *
 * Headless `main` for the reference's own CPU renderer (rogerbarton/optix-renderer @ be2533f).
 * The reference can only be driven through its GLFW window (src/utils/main.cpp:81-105 is
 * stale), so this file restates the ~60 lines of RenderThread::loadScene
 * (src/utils/render.cpp:118-178), RenderThread::renderThreadMain (:232-360) and the static
 * renderBlock (:421-459) with std::thread instead of TBB, and writes .npy instead of EXR.
 * Everything else — parser, plugins, BVH build and traversal, integrators, film — is the
 * reference's unmodified code, compiled from /root/reference by oracle/Makefile.
 *
 * Modes (see usage()):
 *   render   scene.xml out.npy [spp=N] [threads=T] [film=raw.npy] [split=K out2=b.npy [seg=M]]
 *            (split: passes [0,K) go to out.npy, passes [K,spp) to out2 — independent renders of one run; with seg=M the
 *            tail is cut into M-pass pieces written to b.npy.0, b.npy.1, ...)
 *   genrays  scene.xml rays.bin n_primary seed
 *   hits     scene.xml rays.bin hits.bin
 *   bvh      scene.xml bvh.bin
 *   flat     scene.xml scene.nbs
 *   test     test.xml                 (runs <test type="ttest|chi2test"> roots)
 *   ttest    test.xml [samples=N]     (scene-mode t-tests; restated, see modeTTestScenes)
 *   rng      n                        (pcg32 known-answer stream)
 */
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <filesystem>
#include <functional>
#include <algorithm>
#include <cmath>
#include <Eigen/Core>
#include <Eigen/Geometry>
#include <tinyformat.h>
#include <pcg32.h>
#include <imgui/imgui.h>
#include <filesystem/resolver.h>

/* The harness needs the BVH's node/index arrays (to dump them and to report which primitive
 * was hit — Intersection does not carry the index). Access control is lifted for the nori
 * headers of THIS translation unit only; layout is unchanged. */
#define private public
#define protected public
#include <nori/bvh.h>
#include <nori/scene.h>
#include <nori/mesh.h>
#include <nori/sphere.h>
#include <nori/camera.h>
#include <nori/perspective.h>
#include <nori/emitter.h>
#include <nori/dpdf.h>
#undef private
#undef protected
#include <nori/parser.h>
#include <nori/integrator.h>
#include <nori/sampler.h>
#include <nori/bsdf.h>
#include <nori/block.h>
#include <nori/bitmap.h>
#include <nori/rfilter.h>
#include <nori/timer.h>
#include <nori/warp.h>
#include <pugixml.hpp>
#include <hypothesis.h>
#include <unistd.h>
#include "../include/norib200.h"
#include "../include/nb_scene_file.h"

using namespace nori;

/* ---- OpenEXR is absent: Bitmap I/O is never reached by the harness -------------------- */
NORI_NAMESPACE_BEGIN
Bitmap::Bitmap(const std::string &f) { throw NoriException("oracle harness: EXR loading unavailable (%s)", f); }
void Bitmap::save(const std::string &) { throw NoriException("oracle harness: EXR saving unavailable"); }
void Bitmap::saveToLDR(const std::string &) { throw NoriException("oracle harness: PNG saving unavailable"); }
NORI_NAMESPACE_END

/* ---- BVH query counters through ld --wrap (oracle/Makefile) ---------------------------- */
struct RayCounters { uint64_t closest = 0, any = 0, dup = 0; float last[8] = {0}; bool lastValid = false; };
static thread_local RayCounters tlsCounters;
extern "C" bool __real__ZNK4nori3BVH12rayIntersectERKNS_4TRayINS_6TPointIfLi3EEENS_7TVectorIfLi3EEEEERNS_12IntersectionEb(
    const BVH *, const Ray3f &, Intersection &, bool);
extern "C" bool __wrap__ZNK4nori3BVH12rayIntersectERKNS_4TRayINS_6TPointIfLi3EEENS_7TVectorIfLi3EEEEERNS_12IntersectionEb(
    const BVH *self, const Ray3f &ray, Intersection &its, bool shadow) {
    RayCounters &c = tlsCounters;
    float cur[8] = { ray.o.x(), ray.o.y(), ray.o.z(), ray.d.x(), ray.d.y(), ray.d.z(), ray.mint, ray.maxt };
    /* the MIS probe ray (path_mis.cpp:119) is traced again as the next main ray (:34): count once */
    if (!shadow && c.lastValid && !memcmp(cur, c.last, sizeof(cur))) c.dup++;
    if (shadow) c.any++; else { c.closest++; memcpy(c.last, cur, sizeof(cur)); c.lastValid = true; }
    return __real__ZNK4nori3BVH12rayIntersectERKNS_4TRayINS_6TPointIfLi3EEENS_7TVectorIfLi3EEEEERNS_12IntersectionEb(self, ray, its, shadow);
}

/* ---- .npy writer ----------------------------------------------------------------------- */
static void writeNpy(const std::string &path, const float *data, const std::vector<size_t> &shape) {
    std::string dims;
    size_t n = 1;
    for (size_t i = 0; i < shape.size(); ++i) { dims += std::to_string(shape[i]) + (shape.size() == 1 || i + 1 < shape.size() ? "," : ""); n *= shape[i]; }
    std::string hdr = "{'descr': '<f4', 'fortran_order': False, 'shape': (" + dims + "), }";
    while ((10 + hdr.size() + 1) % 64) hdr += ' ';
    hdr += '\n';
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw NoriException("cannot write %s", path);
    const unsigned char magic[8] = { 0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0 };
    fwrite(magic, 1, 8, f);
    uint16_t hl = (uint16_t) hdr.size();
    fwrite(&hl, 2, 1, f);
    fwrite(hdr.data(), 1, hdr.size(), f);
    fwrite(data, sizeof(float), n, f);
    fclose(f);
}

/* ---- scene loading: RenderThread::loadScene, render.cpp:118-178 ------------------------ */
struct LoadedScene {
    std::unique_ptr<NoriObject> guiRoot;
    Scene *gui = nullptr;
    std::unique_ptr<Scene> render;
    double loadSeconds = 0;
};

static LoadedScene loadScene(const std::string &filename) {
    auto t0 = std::chrono::steady_clock::now();
    filesystem::path path(filename);
    getFileResolver()->prepend(path.parent_path());
    LoadedScene ls;
    /* the reference prints progress to stdout; keep stdout clean for our JSON line */
    std::streambuf *old = std::cout.rdbuf(std::cerr.rdbuf());
    ls.guiRoot.reset(loadFromXML(filename));
    if (ls.guiRoot->getClassType() != NoriObject::EScene) {
        std::cout.rdbuf(old);
        throw NoriException("root of %s is not a scene", filename);
    }
    ls.gui = static_cast<Scene *>(ls.guiRoot.get());
    ls.render.reset(static_cast<Scene *>(ls.gui->cloneAndInit()));
    ls.render->update(ls.gui);
    std::cout.rdbuf(old);
    ls.loadSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return ls;
}

static std::map<std::string, std::string> parseKv(int argc, char **argv, int first) {
    std::map<std::string, std::string> kv;
    for (int i = first; i < argc; ++i) {
        std::string a(argv[i]);
        size_t eq = a.find('=');
        if (eq == std::string::npos) kv[a] = "1"; else kv[a.substr(0, eq)] = a.substr(eq + 1);
    }
    return kv;
}

/* ---- render: renderThreadMain (render.cpp:232-360) + renderBlock (:421-459) ------------ */
static void renderBlockRef(const Scene *scene, Integrator *integrator, Sampler *sampler, ImageBlock &block,
                           ImageBlock *blockNormal = nullptr) {
    const Camera *camera = scene->getCamera();
    Point2i offset = block.getOffset();
    std::vector<std::pair<int, int>> sampleIndices = sampler->getSampleIndices(block);
    block.clear();
    if (blockNormal) blockNormal->clear();
    for (size_t i = 0; i < sampleIndices.size(); i++) {
        int x = sampleIndices[i].first, y = sampleIndices[i].second;
        Point2f pixelSample = Point2f((float) (x + offset.x()), (float) (y + offset.y())) + sampler->next2D();
        Point2f apertureSample = sampler->next2D();
        Ray3f ray;
        Color3f value = camera->sampleRay(ray, pixelSample, apertureSample);
        Color3f albedo = 0.f, normal = 0.f;
        value *= integrator->Li(scene, sampler, ray, albedo, normal);
        block.put(pixelSample, value);
        if (blockNormal) {                      /* render.cpp:456-457; the albedo layer is 0 in every integrator */
            blockNormal->put(pixelSample, normal);   /* put() rejects (and reports) every normal with a negative component */
        }
    }
}

static int modeRender(int argc, char **argv) {
    if (argc < 4) return 2;
    auto kv = parseKv(argc, argv, 4);
    LoadedScene ls = loadScene(argv[2]);
    Scene *scene = ls.render.get();
    const Camera *camera = scene->getCamera();
    const Vector2i outputSize = camera->getOutputSize();
    int spp = kv.count("spp") ? std::stoi(kv["spp"]) : scene->getSampler()->getSampleCount();
    int nThreads = kv.count("threads") ? std::stoi(kv["threads"]) : (int) std::thread::hardware_concurrency();
    if (nThreads < 1) nThreads = 1;

    ImageBlock result(outputSize, camera->getReconstructionFilter());
    result.clear();
    const bool wantNormal = kv.count("normal") != 0;
    ImageBlock resultNormal(outputSize, camera->getReconstructionFilter());
    resultNormal.clear();
    /* split=K: a second film takes the passes from K on. The per-block samplers run on through both parts, so the two
     * images are renders of the SAME reference run over disjoint stretches of its random streams: independent, which
     * two separate invocations (identical per-block seeds, independent.cpp:55-60) would not be. */
    const int split = kv.count("split") ? std::stoi(kv["split"]) : spp;
    const int seg = kv.count("seg") ? std::stoi(kv["seg"]) : std::max(1, spp - split);
    const int nSeg = spp > split ? (spp - split + seg - 1) / seg : 0;
    std::vector<std::unique_ptr<ImageBlock>> result2;
    for (int i = 0; i < nSeg; ++i) {
        result2.emplace_back(new ImageBlock(outputSize, camera->getReconstructionFilter()));
        result2.back()->clear();
    }
    /* render.cpp:266: the adaptive sampler works on 4 x 4 blocks, and each block runs the loop of render.cpp:323-335 */
    const bool adaptive = scene->getSampler()->isAdaptive();
    const int blockSize = adaptive ? NORI_BLOCK_SIZE_ADAPTIVE : NORI_BLOCK_SIZE;
    BlockGenerator blockGenerator(outputSize, blockSize);
    Integrator *integrator = scene->getIntegrator(false);
    {
        std::streambuf *old = std::cout.rdbuf(std::cerr.rdbuf());
        integrator->preprocess(scene);
        std::cout.rdbuf(old);
    }
    const int numBlocks = blockGenerator.getBlockCount();
    std::vector<std::unique_ptr<Sampler>> samplers(numBlocks);
    std::vector<float> blockFlat(numBlocks, 0.f);        /* adaptive: rounds whose first placed decision saw a constant variance map */
    std::vector<float> blockPasses(numBlocks, 0.f);      /* adaptive: passes each block rendered, summed over the rounds */

    std::vector<RayCounters> counters(nThreads);
    std::streambuf *cerrBuf = wantNormal ? std::cerr.rdbuf(nullptr) : nullptr;   /* one warning per rejected normal otherwise */
    auto t0 = std::chrono::steady_clock::now();
    for (int currentSample = 0; currentSample < spp; ++currentSample) {
        scene->getSampler()->setSampleRound(currentSample);
        std::atomic<int> nextBlock(0);
        auto worker = [&](int tid) {
            ImageBlock block(Vector2i(blockSize), camera->getReconstructionFilter());
            ImageBlock blockN(Vector2i(blockSize), camera->getReconstructionFilter());
            while (true) {
                if (nextBlock.fetch_add(1) >= numBlocks) break;
                blockGenerator.next(block);
                blockN.setOffset(block.getOffset()); blockN.setSize(block.getSize());
                uint32_t blockId = block.getBlockId();
                if (currentSample == 0) {
                    std::unique_ptr<Sampler> sampler(scene->getSampler()->clone());
                    sampler->prepare(block);
                    samplers.at(blockId) = std::move(sampler);
                }
                if (adaptive) {
                    int count = 0;
                    samplers.at(blockId)->setSampleRound(count);
                    for (;;) {
                        /* what the sampler is about to see at its first placed decision of this round (count == 3 with the
                         * default initialUniform = 2): is the variance map constant? */
                        if (count == 3 && computeVarianceFromImage(block).maxCoeff() == 0.f) ++blockFlat[blockId];
                        if (!samplers.at(blockId)->computeVariance(block)) break;
                        renderBlockRef(scene, integrator, samplers.at(blockId).get(), block, nullptr);
                        result.put(block);
                        samplers.at(blockId)->setSampleRound(count++);
                        ++blockPasses[blockId];
                    }
                    continue;
                }
                renderBlockRef(scene, integrator, samplers.at(blockId).get(), block, wantNormal ? &blockN : nullptr);
                if (currentSample < split) result.put(block); else result2[(currentSample - split) / seg]->put(block);
                if (wantNormal) resultNormal.put(blockN);
            }
            RayCounters &c = tlsCounters;
            counters[tid].closest += c.closest; counters[tid].any += c.any; counters[tid].dup += c.dup;
            c = RayCounters();
        };
        std::vector<std::thread> pool;
        for (int t = 1; t < nThreads; ++t) pool.emplace_back(worker, t);
        worker(0);
        for (auto &t : pool) t.join();
        blockGenerator.reset();
    }
    double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (cerrBuf) std::cerr.rdbuf(cerrBuf);

    std::unique_ptr<Bitmap> bitmap(result.toBitmap());
    writeNpy(argv[3], (const float *) bitmap->data(), { (size_t) outputSize.y(), (size_t) outputSize.x(), 3 });
    if (wantNormal) {                           /* the Normal render layer (integrator.h:29-39), as toBitmap() gives it */
        std::unique_ptr<Bitmap> nb(resultNormal.toBitmap());
        writeNpy(kv["normal"], (const float *) nb->data(), { (size_t) outputSize.y(), (size_t) outputSize.x(), 3 });
    }
    for (int i = 0; kv.count("out2") && i < nSeg; ++i) {
        std::unique_ptr<Bitmap> b2(result2[i]->toBitmap());
        writeNpy(nSeg == 1 && !kv.count("seg") ? kv["out2"] : kv["out2"] + "." + std::to_string(i), (const float *) b2->data(),
                 { (size_t) outputSize.y(), (size_t) outputSize.x(), 3 });
    }
    if (kv.count("film"))
        writeNpy(kv["film"], (const float *) result.data(), { (size_t) result.rows(), (size_t) result.cols(), 4 });
    if (adaptive && kv.count("passes")) writeNpy(kv["passes"], blockPasses.data(), { (size_t) numBlocks });
    if (adaptive && kv.count("flat")) writeNpy(kv["flat"], blockFlat.data(), { (size_t) numBlocks });

    uint64_t closest = 0, any = 0, dup = 0;
    for (auto &c : counters) { closest += c.closest; any += c.any; dup += c.dup; }
    double samples = (double) outputSize.x() * outputSize.y() * spp;
    printf("{\"mode\": \"render\", \"scene\": \"%s\", \"width\": %d, \"height\": %d, \"spp\": %d, \"threads\": %d, "
           "\"seconds\": %.6f, \"load_seconds\": %.3f, \"samples\": %.0f, \"samples_per_s\": %.1f, "
           "\"rays_closest\": %llu, \"rays_any\": %llu, \"rays_dup\": %llu, \"rays_raw_per_s\": %.1f, \"rays_dedup_per_s\": %.1f}\n",
           argv[2], outputSize.x(), outputSize.y(), spp, nThreads, seconds, ls.loadSeconds, samples, samples / seconds,
           (unsigned long long) closest, (unsigned long long) any, (unsigned long long) dup,
           (closest + any) / seconds, (closest + any - dup) / seconds);
    return 0;
}

/* ---- traversal that also reports WHICH primitive was hit --------------------------------
 * Same loop as BVH::rayIntersect (src/utils/bvh.cpp:402-460) run over the reference's own
 * node/index arrays and calling the reference's own bbox and Shape::rayIntersect code; it
 * exists only because Intersection carries no primitive index. modeHits cross-checks its
 * t/shape against the real BVH::rayIntersect for every ray. */
struct HitRecord {          /* 128 bytes, little endian; documented in tests/golden/README.md */
    int32_t shape;          /* index into Scene::getShapes(), -1 = miss */
    int32_t prim;           /* primitive index local to the shape */
    float t, u, v;          /* as returned by Shape::rayIntersect (before setHitInformation) */
    uint32_t nodesVisited, primsTested;
    uint32_t anyHit;        /* result of the shadow-ray query on the same ray */
    float p[3], uv[2];      /* Intersection after setHitInformation */
    float shN[3], shS[3], shT[3], geoN[3];
    uint32_t pad[7];
};
static_assert(sizeof(HitRecord) == 128, "HitRecord layout");

static bool traverseWithIndex(const BVH *bvh, const Ray3f &_ray, HitRecord &rec) {
    uint32_t node_idx = 0, stack_idx = 0, stack[64];
    Ray3f ray(_ray);
    rec.shape = -1; rec.prim = -1; rec.t = std::numeric_limits<float>::infinity(); rec.u = rec.v = 0;
    rec.nodesVisited = rec.primsTested = 0;
    if (ray.mint == Epsilon)
        ray.mint = std::max(ray.mint, ray.mint * ray.o.array().abs().maxCoeff());
    if (bvh->m_nodes.empty() || ray.maxt < ray.mint)
        return false;
    bool found = false;
    while (true) {
        const BVH::BVHNode &node = bvh->m_nodes[node_idx];
        rec.nodesVisited++;
        if (!node.bbox.rayIntersect(ray)) {
            if (stack_idx == 0) break;
            node_idx = stack[--stack_idx];
            continue;
        }
        if (node.isInner()) {
            stack[stack_idx++] = node.inner.rightChild;
            node_idx++;
        } else {
            for (uint32_t i = node.start(), end = node.end(); i < end; ++i) {
                uint32_t idx = bvh->m_indices[i];
                uint32_t shapeIdx = bvh->findShape(idx);
                const Shape *shape = bvh->m_shapes[shapeIdx];
                float u = 0, v = 0, t;
                rec.primsTested++;
                if (shape->rayIntersect(idx, ray, u, v, t)) {
                    found = true;
                    ray.maxt = rec.t = t;
                    rec.u = u; rec.v = v;
                    rec.shape = (int32_t) shapeIdx; rec.prim = (int32_t) idx;
                }
            }
            if (stack_idx == 0) break;
            node_idx = stack[--stack_idx];
        }
    }
    return found;
}

static std::vector<float> readFloats(const std::string &path) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) throw NoriException("cannot read %s", path);
    size_t bytes = (size_t) f.tellg();
    f.seekg(0);
    std::vector<float> v(bytes / 4);
    f.read((char *) v.data(), bytes);
    return v;
}

static int modeHits(int argc, char **argv) {
    if (argc < 5) return 2;
    LoadedScene ls = loadScene(argv[2]);
    const Scene *scene = ls.render.get();
    std::vector<float> rays = readFloats(argv[3]);
    size_t n = rays.size() / 8;
    std::vector<HitRecord> out(n);
    size_t mismatches = 0;
    for (size_t i = 0; i < n; ++i) {
        const float *r = &rays[8 * i];
        Ray3f ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]), r[6], r[7]);
        HitRecord &rec = out[i];
        memset(&rec, 0, sizeof(rec));
        bool found = traverseWithIndex(scene->getBVH(), ray, rec);
        Intersection its;
        bool foundRef = scene->rayIntersect(ray, its);
        rec.anyHit = scene->rayIntersect(ray) ? 1u : 0u;
        if (found != foundRef || (found && (its.t != rec.t || its.shape != scene->getShapes()[rec.shape])))
            mismatches++;
        if (foundRef) {
            for (int k = 0; k < 3; ++k) {
                rec.p[k] = its.p[k]; rec.shN[k] = its.shFrame.n[k]; rec.shS[k] = its.shFrame.s[k];
                rec.shT[k] = its.shFrame.t[k]; rec.geoN[k] = its.geoFrame.n[k];
            }
            rec.uv[0] = its.uv.x(); rec.uv[1] = its.uv.y();
        }
    }
    FILE *f = fopen(argv[4], "wb");
    fwrite(out.data(), sizeof(HitRecord), n, f);
    fclose(f);
    printf("{\"mode\": \"hits\", \"rays\": %zu, \"harness_vs_bvh_mismatches\": %zu}\n", n, mismatches);
    return mismatches ? 1 : 0;
}

/* A fixed, realistic ray population: primary rays, bounce rays leaving the primary hits with
 * mint = Epsilon (exercises the adaptive-epsilon branch, bvh.cpp:409-410), and the integrators'
 * own NEE shadow rays (finite maxt) from Emitter::sample. */
static int modeGenRays(int argc, char **argv) {
    if (argc < 6) return 2;
    LoadedScene ls = loadScene(argv[2]);
    const Scene *scene = ls.render.get();
    const Camera *camera = scene->getCamera();
    size_t nPrimary = std::stoull(argv[4]);
    pcg32 rng((uint64_t) std::stoull(argv[5]), 7u);
    std::vector<float> out;
    auto push = [&](const Ray3f &r) {
        const float v[8] = { r.o.x(), r.o.y(), r.o.z(), r.d.x(), r.d.y(), r.d.z(), r.mint, r.maxt };
        out.insert(out.end(), v, v + 8);
    };
    Vector2i size = camera->getOutputSize();
    for (size_t i = 0; i < nPrimary; ++i) {
        Point2f px(rng.nextFloat() * size.x(), rng.nextFloat() * size.y());
        Ray3f ray;
        camera->sampleRay(ray, px, Point2f(rng.nextFloat(), rng.nextFloat()));
        push(ray);
        Intersection its;
        for (int bounce = 0; bounce < 3; ++bounce) {
            if (!scene->rayIntersect(ray, its)) break;
            /* NEE shadow ray exactly as path_mis.cpp:75-89 builds it */
            if (!scene->getLights().empty()) {
                const Emitter *em = scene->getRandomEmitter(rng.nextFloat());
                if (em) {
                    EmitterQueryRecord eqr(its.p);
                    em->sample(eqr, Point2f(rng.nextFloat(), rng.nextFloat()));
                    if (std::isfinite(eqr.shadowRay.maxt) && eqr.shadowRay.d.allFinite()) push(eqr.shadowRay);
                }
            }
            Vector3f d = Warp::squareToUniformSphere(Point2f(rng.nextFloat(), rng.nextFloat()));
            if (d.dot(its.geoFrame.n) < 0 && rng.nextFloat() < 0.8f) d = -d;
            ray = Ray3f(its.p, d);
            push(ray);
        }
    }
    FILE *f = fopen(argv[3], "wb");
    fwrite(out.data(), 4, out.size(), f);
    fclose(f);
    printf("{\"mode\": \"genrays\", \"rays\": %zu}\n", out.size() / 8);
    return 0;
}

/* bvh.bin: u32 magic 'NBVH', u32 nNodes, u32 nIndices, u32 nShapes, then nodes (32 B each, the
 * reference's in-memory BVHNode, bvh.h:127-164), indices (u32), shapeOffset (nShapes+1 u32). */
static int modeBvh(int argc, char **argv) {
    if (argc < 4) return 2;
    LoadedScene ls = loadScene(argv[2]);
    const BVH *bvh = ls.render->getBVH();
    FILE *f = fopen(argv[3], "wb");
    uint32_t hdr[4] = { 0x4856424eu, (uint32_t) bvh->m_nodes.size(), (uint32_t) bvh->m_indices.size(), (uint32_t) bvh->m_shapes.size() };
    fwrite(hdr, 4, 4, f);
    fwrite(bvh->m_nodes.data(), sizeof(BVH::BVHNode), bvh->m_nodes.size(), f);
    fwrite(bvh->m_indices.data(), 4, bvh->m_indices.size(), f);
    fwrite(bvh->m_shapeOffset.data(), 4, bvh->m_shapeOffset.size(), f);
    fclose(f);
    printf("{\"mode\": \"bvh\", \"nodes\": %u, \"indices\": %u, \"shapes\": %u}\n", hdr[1], hdr[2], hdr[3]);
    return 0;
}


/* ---- flat: dump what the reference holds after Scene::update() as an .nbs container -------
 * (include/nb_scene_file.h). Geometry, area CDFs, emitter CDF, camera matrices and the BVH are
 * read out of the reference's own objects; BSDF parameters live in classes private to their
 * .cpp files and are therefore NOT dumped (the product's own loader reads them from the XML).
 * Used as the golden for the product loader and as the first GPU fixtures. */
struct NbsWriter {
    struct Sec { std::string name; uint32_t elem; std::vector<char> bytes; };
    std::vector<Sec> secs;
    void add(const char *name, uint32_t elem, const void *data, size_t count) {
        Sec s; s.name = name; s.elem = elem;
        s.bytes.assign((const char *) data, (const char *) data + (size_t) elem * count);
        secs.push_back(std::move(s));
    }
    void write(const std::string &path) {
        FILE *f = fopen(path.c_str(), "wb");
        if (!f) throw NoriException("cannot write %s", path);
        uint32_t hdr[2] = { NBS_MAGIC, (uint32_t) secs.size() };
        fwrite(hdr, 4, 2, f);
        uint64_t off = 8 + sizeof(nbs_section) * secs.size();
        std::vector<nbs_section> table(secs.size());
        for (size_t i = 0; i < secs.size(); ++i) {
            off = (off + 15) & ~(uint64_t) 15;
            memset(&table[i], 0, sizeof(nbs_section));
            strncpy(table[i].name, secs[i].name.c_str(), 15);
            table[i].elem_size = secs[i].elem;
            table[i].count = secs[i].bytes.size() / secs[i].elem;
            table[i].offset = off;
            off += secs[i].bytes.size();
        }
        fwrite(table.data(), sizeof(nbs_section), table.size(), f);
        for (size_t i = 0; i < secs.size(); ++i) {
            long cur = ftell(f);
            for (uint64_t k = cur; k < table[i].offset; ++k) fputc(0, f);
            fwrite(secs[i].bytes.data(), 1, secs[i].bytes.size(), f);
        }
        fclose(f);
    }
};

static int modeFlat(int argc, char **argv) {
    if (argc < 4) return 2;
    LoadedScene ls = loadScene(argv[2]);
    Scene *scene = ls.render.get();
    const std::vector<Shape *> &shapes = scene->getShapes();
    const std::vector<Emitter *> &lights = scene->getLights();
    std::vector<nb_shape> nshapes;
    std::vector<float> positions, normals, uvs, tangents, bitangents, areaCdf;
    std::vector<uint32_t> faces;
    bool anyN = false, anyUV = false, anyT = false;
    for (Shape *sh : shapes) {
        if (auto *m = dynamic_cast<Mesh *>(sh)) { anyN |= m->m_N.size() > 0; anyUV |= m->m_UV.size() > 0; anyT |= m->m_T.size() > 0; }
    }
    for (Shape *sh : shapes) {
        nb_shape ns; memset(&ns, 0, sizeof(ns));
        ns.bsdf = -1; ns.emitter = -1; ns.medium = -1;
        for (size_t e = 0; e < lights.size(); ++e) if (lights[e] == sh->getEmitter()) ns.emitter = (int32_t) e;
        ns.prim_count = sh->getPrimitiveCount();
        if (auto *m = dynamic_cast<Mesh *>(sh)) {
            ns.type = NB_SHAPE_MESH;
            ns.vertex_offset = (uint32_t) (positions.size() / 3);
            ns.vertex_count = (uint32_t) m->m_V.cols();
            ns.face_offset = (uint32_t) (faces.size() / 3);
            for (int v = 0; v < m->m_V.cols(); ++v) {
                for (int k = 0; k < 3; ++k) positions.push_back(m->m_V(k, v));
                if (anyN) for (int k = 0; k < 3; ++k) normals.push_back(m->m_N.size() > 0 ? m->m_N(k, v) : 0.f);
                if (anyUV) for (int k = 0; k < 2; ++k) uvs.push_back(m->m_UV.size() > 0 ? m->m_UV(k, v) : 0.f);
                if (anyT) for (int k = 0; k < 3; ++k) {
                    tangents.push_back(m->m_T.size() > 0 ? m->m_T(k, v) : 0.f);
                    bitangents.push_back(m->m_BT.size() > 0 ? m->m_BT(k, v) : 0.f);
                }
            }
            for (int t = 0; t < m->m_F.cols(); ++t) for (int k = 0; k < 3; ++k) faces.push_back(m->m_F(k, t));
            if (m->m_N.size() > 0) ns.flags |= NB_SHAPE_HAS_NORMALS;
            if (m->m_UV.size() > 0) ns.flags |= NB_SHAPE_HAS_UVS;
            ns.cdf_offset = (uint32_t) areaCdf.size();
            areaCdf.insert(areaCdf.end(), m->m_pdf.m_cdf.begin(), m->m_pdf.m_cdf.end());
            ns.pdf_norm = m->m_pdf.m_normalization;
        } else if (auto *s = dynamic_cast<Sphere *>(sh)) {
            ns.type = NB_SHAPE_SPHERE;
            ns.radius = s->m_radius;
            for (int k = 0; k < 3; ++k) ns.center[k] = s->m_position[k];
        } else throw NoriException("flat: unknown shape class");
        nshapes.push_back(ns);
    }
    NbsWriter w;
    w.add("shapes", sizeof(nb_shape), nshapes.data(), nshapes.size());
    w.add("positions", 4, positions.data(), positions.size());
    if (anyN) w.add("normals", 4, normals.data(), normals.size());
    if (anyUV) w.add("uvs", 4, uvs.data(), uvs.size());
    if (anyT) { w.add("tangents", 4, tangents.data(), tangents.size()); w.add("bitangents", 4, bitangents.data(), bitangents.size()); }
    w.add("faces", 4, faces.data(), faces.size());
    w.add("area_cdf", 4, areaCdf.data(), areaCdf.size());
    w.add("emitter_cdf", 4, scene->emitterDpdf.m_cdf.data(), scene->emitterDpdf.m_cdf.size());
    std::vector<nb_emitter> nem;
    for (Emitter *e : lights) {
        nb_emitter ne; memset(&ne, 0, sizeof(ne));
        ne.type = e->m_shape ? NB_EMITTER_AREA : NB_EMITTER_POINT;
        ne.shape = -1;
        for (size_t i = 0; i < shapes.size(); ++i) if (shapes[i] == e->m_shape) ne.shape = (int32_t) i;
        for (int k = 0; k < 3; ++k) ne.radiance[k] = e->m_radiance[k];
        nem.push_back(ne);
    }
    w.add("emitters", sizeof(nb_emitter), nem.data(), nem.size());
    if (auto *cam = dynamic_cast<const PerspectiveCamera *>(scene->getCamera())) {
        nb_camera c; memset(&c, 0, sizeof(c));
        for (int r = 0; r < 4; ++r) for (int col = 0; col < 4; ++col) {
            c.sample_to_camera[r * 4 + col] = cam->m_sampleToCamera.getMatrix()(r, col);
            c.camera_to_world[r * 4 + col] = cam->m_cameraToWorld.getMatrix()(r, col);
        }
        c.inv_output_size[0] = cam->m_invOutputSize.x(); c.inv_output_size[1] = cam->m_invOutputSize.y();
        c.near_clip = cam->m_nearClip; c.far_clip = cam->m_farClip;
        c.lens_radius = cam->m_lensRadius; c.focal_distance = cam->m_focalDistance;
        c.width = cam->getOutputSize().x(); c.height = cam->getOutputSize().y();
        w.add("camera", sizeof(nb_camera), &c, 1);
        /* the tabulated filter exactly as ImageBlock::init builds it (block.cpp:58-64) */
        nb_render_params rp; memset(&rp, 0, sizeof(rp));
        const ReconstructionFilter *rf = cam->getReconstructionFilter();
        rp.filter_radius = rf->getRadius();
        for (int i = 0; i < NB_FILTER_RESOLUTION; ++i) rp.filter_table[i] = rf->eval((rp.filter_radius * i) / NB_FILTER_RESOLUTION);
        rp.filter_table[NB_FILTER_RESOLUTION] = 0.f;
        w.add("render", sizeof(nb_render_params), &rp, 1);
    }
    uint32_t spp = (uint32_t) scene->getSampler()->getSampleCount();
    w.add("spp", 4, &spp, 1);
    const BVH *bvh = scene->getBVH();
    static_assert(sizeof(BVH::BVHNode) == sizeof(nb_bvh_node), "node layout");
    w.add("bvh_nodes", sizeof(nb_bvh_node), bvh->m_nodes.data(), bvh->m_nodes.size());
    w.add("bvh_indices", 4, bvh->m_indices.data(), bvh->m_indices.size());
    w.write(argv[3]);
    printf("{\"mode\": \"flat\", \"shapes\": %zu, \"vertices\": %zu, \"faces\": %zu, \"nodes\": %zu, \"emitters\": %zu}\n",
           nshapes.size(), positions.size() / 3, faces.size() / 3, bvh->m_nodes.size(), nem.size());
    return 0;
}

/* <test> roots never execute in this fork (render.cpp:133-137); drive them the only way the
 * classes allow: cloneAndInit() then update(gui), which calls execute() (ttest.cpp:93,:120). */
static int modeTest(int argc, char **argv) {
    if (argc < 3) return 2;
    filesystem::path path(argv[2]);
    getFileResolver()->prepend(path.parent_path());
    std::unique_ptr<NoriObject> root(loadFromXML(argv[2]));
    if (root->getClassType() != NoriObject::ETest) throw NoriException("root is not a <test>");
    std::ostringstream sink;
    std::streambuf *old = std::cout.rdbuf(sink.rdbuf());
    std::unique_ptr<NoriObject> clone(root->cloneAndInit());   /* runs once on un-updated children: discarded */
    sink.str("");
    clone->update(root.get());
    std::cout.rdbuf(old);
    std::cout << sink.str();
    return 0;
}


/* Scene-mode t-tests. StudentsTTest::execute (src/utils/ttest.cpp:191-239) prints
 * scene->toString() first, and that throws in this fork (tinyformat "Too many conversion
 * specifiers"), so the scene branch is restated here: same sampler ("independent", default
 * seed), same pixelSample/apertureSample draws, same Knuth mean/variance, same
 * hypothesis::students_t_test call. Each <scene> child of the <test> root is written to a
 * temporary XML (pugixml) and loaded through the reference parser. */
static int modeTTestScenes(int argc, char **argv) {
    if (argc < 3) return 2;
    auto kv = parseKv(argc, argv, 3);
    filesystem::path path(argv[2]);
    getFileResolver()->prepend(path.parent_path());
    pugi::xml_document doc;
    if (!doc.load_file(argv[2])) throw NoriException("cannot parse %s", argv[2]);
    pugi::xml_node test = doc.child("test");
    std::vector<float> references;
    int sampleCount = 100000;
    float significance = 0.01f;
    for (pugi::xml_node ch : test.children()) {
        std::string name = ch.attribute("name").value();
        if (name == "references") for (auto &tok : tokenize(ch.attribute("value").value())) references.push_back(toFloat(tok));
        if (name == "sampleCount") sampleCount = std::stoi(ch.attribute("value").value());
        if (name == "significanceLevel") significance = toFloat(ch.attribute("value").value());
    }
    if (kv.count("samples")) sampleCount = std::stoi(kv["samples"]);
    int idx = 0, passed = 0;
    for (pugi::xml_node sc : test.children("scene")) {
        std::string tmp = "/tmp/nori_ref_ttest_" + std::to_string(getpid()) + "_" + std::to_string(idx) + ".xml";
        pugi::xml_document one;
        one.append_copy(sc);
        one.save_file(tmp.c_str());
        std::streambuf *old = std::cout.rdbuf(std::cerr.rdbuf());
        std::unique_ptr<NoriObject> root(loadFromXML(tmp));
        std::unique_ptr<Scene> scene(static_cast<Scene *>(static_cast<Scene *>(root.get())->cloneAndInit()));
        scene->update(root.get());
        std::cout.rdbuf(old);
        std::remove(tmp.c_str());
        std::unique_ptr<Sampler> sampler(static_cast<Sampler *>(NoriObjectFactory::createInstance("independent")));
        const Integrator *integrator = scene->getIntegrator(false);
        const Camera *camera = scene->getCamera();
        double mean = 0, variance = 0;
        for (int k = 0; k < sampleCount; ++k) {
            Ray3f ray;
            Point2f pixelSample = (sampler->next2D().array() * camera->getOutputSize().cast<float>().array()).matrix();
            Color3f value = camera->sampleRay(ray, pixelSample, sampler->next2D());
            Color3f tmpA(0.f), tmpB(0, 0, 1);
            value *= integrator->Li(scene.get(), sampler.get(), ray, tmpA, tmpB);
            double result = (double) value.getLuminance();
            double delta = result - mean;
            mean += delta / (double) (k + 1);
            variance += delta * (result - mean);
        }
        variance /= sampleCount - 1;
        float reference = idx < (int) references.size() ? references[idx] : 0.f;
        auto res = hypothesis::students_t_test(mean, variance, reference, sampleCount, significance, (int) references.size());
        if (res.first) ++passed;
        printf("{\"mode\": \"ttest\", \"scene\": %d, \"integrator\": \"%s\", \"mean\": %.9g, \"variance\": %.9g, \"reference\": %.9g, \"samples\": %d, \"accepted\": %s}\n",
               idx, sc.child("integrator").attribute("type").value(), mean, variance, reference, sampleCount, res.first ? "true" : "false");
        ++idx;
    }
    printf("{\"mode\": \"ttest\", \"passed\": %d, \"total\": %d}\n", passed, idx);
    return passed == idx ? 0 : 1;
}

static int modeRng(int argc, char **argv) {
    int n = argc > 2 ? std::stoi(argv[2]) : 8;
    pcg32 a;                       /* default seed */
    for (int i = 0; i < n; ++i) printf("%08x ", a.nextUInt());
    printf("\n");
    pcg32 b; b.seed(32, 64);       /* Independent::prepare(block at offset (32,64)) */
    for (int i = 0; i < n; ++i) printf("%.9g ", b.nextFloat());
    printf("\n");
    pcg32 c; c.seed(42u, 54u);     /* ext/pcg32/pcg32-demo.cpp:54 -> pcg32-demo.out "Round 1" */
    for (int i = 0; i < n; ++i) printf("%08x ", c.nextUInt());
    printf("\n");
    return 0;
}

/* ---- medium: Medium::sampleFreePath / getTransmittance / (heterogeneous) density on fixed rays ----
 * rays.bin: n x 8 f32 (o, d, mint, maxt). Output n x 4 f32: density at o (0 for media without a
 * grid), fraction of `samples` free paths shorter than maxt, their mean length, mean transmittance
 * (channel 0) from o to o + maxt*d. The medium is that of the first shape that has one, else the
 * scene's ambient medium. */
NORI_NAMESPACE_BEGIN
float oracleHeterogDensity(const Medium *m, const Point3f &p, bool &isHeterog);   /* ref_heterog.cpp */
NORI_NAMESPACE_END
static int modeMedium(int argc, char **argv) {
    if (argc < 5) return 2;
    auto kv = parseKv(argc, argv, 5);
    const int samples = kv.count("samples") ? std::stoi(kv["samples"]) : 1024;
    LoadedScene ls = loadScene(argv[2]);
    const Medium *medium = ls.render->getAmbientMedium();
    for (Shape *sh : ls.render->getShapes()) if (sh->getMedium()) { medium = sh->getMedium(); break; }
    if (!medium) throw NoriException("scene has no medium");
    FILE *f = fopen(argv[3], "rb");
    if (!f) throw NoriException("cannot read %s", argv[3]);
    fseek(f, 0, SEEK_END); const size_t n = (size_t) ftell(f) / 32; fseek(f, 0, SEEK_SET);
    std::vector<float> rays(n * 8), out(n * 4);
    if (fread(rays.data(), 32, n, f) != n) throw NoriException("short read on %s", argv[3]);
    fclose(f);
    std::unique_ptr<Sampler> sampler(static_cast<Sampler *>(NoriObjectFactory::createInstance("independent")));
    for (size_t i = 0; i < n; ++i) {
        const float *r = &rays[i * 8];
        Ray3f ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]), r[6], r[7]);
        bool het = false;
        out[i * 4] = oracleHeterogDensity(medium, ray.o, het);
        double hit = 0, len = 0, tr = 0;
        for (int k = 0; k < samples; ++k) {
            MediumQueryRecord mRec(Ray3f(ray.o, ray.d, Epsilon, INFINITY));
            const float t = medium->sampleFreePath(mRec, *sampler);
            if (t < ray.maxt) { hit += 1; len += t; }
            tr += medium->getTransmittance(ray.o, ray(ray.maxt), false, *sampler)(0);
        }
        out[i * 4 + 1] = (float) (hit / samples); out[i * 4 + 2] = hit > 0 ? (float) (len / hit) : 0.f; out[i * 4 + 3] = (float) (tr / samples);
    }
    f = fopen(argv[4], "wb");
    if (!f) throw NoriException("cannot write %s", argv[4]);
    fwrite(out.data(), 16, n, f);
    fclose(f);
    printf("{\"mode\": \"medium\", \"rays\": %zu, \"samples\": %d}\n", n, samples);
    return 0;
}

/* ---- emitter: Emitter::sample / pdf / eval on fixed (reference point, 2D sample) pairs ----------
 * in.bin: n x 5 f32 (ref.xyz, sample.xy). out.bin: n x 12 f32: lRec.wi, the returned value
 * (radiance / pdf), lRec.pdf, shadowRay.maxt, emitter->eval(lRec) after sampling, 0. */
static int modeEmitter(int argc, char **argv) {
    if (argc < 5) return 2;
    auto kv = parseKv(argc, argv, 5);
    const size_t index = kv.count("index") ? (size_t) std::stoi(kv["index"]) : 0;
    LoadedScene ls = loadScene(argv[2]);
    const std::vector<Emitter *> &lights = ls.render->getLights();
    if (index >= lights.size()) throw NoriException("scene has %d emitters", (int) lights.size());
    const Emitter *em = lights[index];
    FILE *f = fopen(argv[3], "rb");
    if (!f) throw NoriException("cannot read %s", argv[3]);
    fseek(f, 0, SEEK_END); const size_t n = (size_t) ftell(f) / 20; fseek(f, 0, SEEK_SET);
    std::vector<float> in(n * 5), out(n * 12, 0.f);
    if (fread(in.data(), 20, n, f) != n) throw NoriException("short read on %s", argv[3]);
    fclose(f);
    for (size_t i = 0; i < n; ++i) {
        EmitterQueryRecord lRec(Point3f(in[i * 5], in[i * 5 + 1], in[i * 5 + 2]));
        const Color3f v = em->sample(lRec, Point2f(in[i * 5 + 3], in[i * 5 + 4]));
        const Color3f e = em->eval(lRec);
        float *o = &out[i * 12];
        o[0] = lRec.wi.x(); o[1] = lRec.wi.y(); o[2] = lRec.wi.z(); o[3] = v(0); o[4] = v(1); o[5] = v(2);
        o[6] = lRec.pdf; o[7] = lRec.shadowRay.maxt; o[8] = e(0); o[9] = e(1); o[10] = e(2);
    }
    f = fopen(argv[4], "wb");
    if (!f) throw NoriException("cannot write %s", argv[4]);
    fwrite(out.data(), 48, n, f);
    fclose(f);
    printf("{\"mode\": \"emitter\", \"index\": %zu, \"emitters\": %zu, \"records\": %zu}\n", index, lights.size(), n);
    return 0;
}

/* BSDF::eval / pdf / sample of the BSDF of shape `index` on fixed records (include/nori/bsdf.h:33-76):
 * in = 10 floats {wi, wo, sample, uv}; out = 12 floats {eval rgb, pdf, sample weight rgb, sampled wo,
 * measure == EDiscrete, pdf of the sampled direction} — the layout of nb_bsdf_probe. */
static int modeBsdf(int argc, char **argv) {
    if (argc < 5) return 2;
    auto kv = parseKv(argc, argv, 5);
    const size_t index = kv.count("index") ? (size_t) std::stoi(kv["index"]) : 0;
    LoadedScene ls = loadScene(argv[2]);
    const auto &shapes = ls.render->getShapes();
    if (index >= shapes.size() || !shapes[index]->getBSDF()) throw NoriException("shape %d has no bsdf", (int) index);
    const BSDF *bsdf = shapes[index]->getBSDF();
    FILE *f = fopen(argv[3], "rb");
    if (!f) throw NoriException("cannot read %s", argv[3]);
    fseek(f, 0, SEEK_END); const size_t n = (size_t) ftell(f) / 40; fseek(f, 0, SEEK_SET);
    std::vector<float> in(n * 10), out(n * 12, 0.f);
    if (fread(in.data(), 40, n, f) != n) throw NoriException("short read on %s", argv[3]);
    fclose(f);
    for (size_t i = 0; i < n; ++i) {
        const float *r = &in[i * 10];
        const Vector3f wi(r[0], r[1], r[2]), wo(r[3], r[4], r[5]);
        BSDFQueryRecord bRec(wi, wo, ESolidAngle);
        bRec.uv = Point2f(r[8], r[9]);
        const Color3f e = bsdf->eval(bRec);
        const float pdf = bsdf->pdf(bRec);
        BSDFQueryRecord sRec(wi);
        sRec.uv = Point2f(r[8], r[9]);
        sRec.measure = EUnknownMeasure;
        const Color3f w = bsdf->sample(sRec, Point2f(r[6], r[7]));
        const bool discrete = sRec.measure == EDiscrete;
        BSDFQueryRecord pRec(wi, sRec.wo, ESolidAngle);
        pRec.uv = sRec.uv;
        float *o = &out[i * 12];
        o[0] = e(0); o[1] = e(1); o[2] = e(2); o[3] = pdf; o[4] = w(0); o[5] = w(1); o[6] = w(2);
        o[7] = sRec.wo.x(); o[8] = sRec.wo.y(); o[9] = sRec.wo.z(); o[10] = discrete ? 1.f : 0.f;
        o[11] = discrete ? 0.f : bsdf->pdf(pRec);
    }
    f = fopen(argv[4], "wb");
    if (!f) throw NoriException("cannot write %s", argv[4]);
    fwrite(out.data(), 48, n, f);
    fclose(f);
    printf("{\"mode\": \"bsdf\", \"shape\": %zu, \"records\": %zu}\n", index, n);
    return 0;
}

static int usage() {
    fprintf(stderr,
        "nori_ref — headless harness around the reference CPU renderer (test infrastructure)\n"
        "  nori_ref render  scene.xml out.npy [spp=N] [threads=T] [film=raw.npy] [normal=layer.npy]\n"
        "  nori_ref genrays scene.xml rays.bin n_primary seed\n"
        "  nori_ref hits    scene.xml rays.bin hits.bin\n"
        "  nori_ref bvh     scene.xml bvh.bin\n"
        "  nori_ref flat    scene.xml scene.nbs      (flattened scene container, include/nb_scene_file.h)\n"
        "  nori_ref test    test.xml\n"
        "  nori_ref ttest   test.xml [samples=N]   (scene-mode t-tests, restated: see modeTTestScenes)\n"
        "  nori_ref rng     n\n"
        "  nori_ref emitter scene.xml in.bin out.bin [index=K]      (Emitter::sample/pdf/eval probes)\n"
        "  nori_ref bsdf    scene.xml in.bin out.bin [index=K]      (BSDF::eval/pdf/sample probes of shape K's BSDF)\n"
        "  nori_ref medium  scene.xml rays.bin out.bin [samples=K]   (free path / transmittance / density probes)\n");
    return 2;
}

int main(int argc, char **argv) {
    if (argc < 2) return usage();
    std::string mode(argv[1]);
    try {
        int rc = 2;
        if (mode == "render") rc = modeRender(argc, argv);
        else if (mode == "genrays") rc = modeGenRays(argc, argv);
        else if (mode == "hits") rc = modeHits(argc, argv);
        else if (mode == "bvh") rc = modeBvh(argc, argv);
        else if (mode == "flat") rc = modeFlat(argc, argv);
        else if (mode == "test") rc = modeTest(argc, argv);
        else if (mode == "ttest") rc = modeTTestScenes(argc, argv);
        else if (mode == "rng") rc = modeRng(argc, argv);
        else if (mode == "medium") rc = modeMedium(argc, argv);
        else if (mode == "emitter") rc = modeEmitter(argc, argv);
        else if (mode == "bsdf") rc = modeBsdf(argc, argv);
        if (rc == 2) return usage();
        return rc;
    } catch (const std::exception &e) {
        fprintf(stderr, "nori_ref: %s\n", e.what());
        return 1;
    }
}
