/* main.cpp — nori_b200: command-line renderer. Plays the role of RenderThread's render loop
 * (src/utils/render.cpp:232-419 with the GPU thread of :462-574): load XML, preRender, subframes
 * of samplesPerLaunch until the sampler's sampleCount is reached, save. Output is .npy (H x W x 3
 * float32) or .pfm — the reference's EXR writer needs OpenEXR, which this build does not carry. */
#include "b200state.h"
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>

using namespace nori;

static void writeNpy(const std::string &path, const float *data, size_t h, size_t w, size_t c) {
    std::string hdr = "{'descr': '<f4', 'fortran_order': False, 'shape': (" + std::to_string(h) + ", " + std::to_string(w) + ", " + std::to_string(c) + "), }";
    while ((10 + hdr.size() + 1) % 64) hdr += ' ';
    hdr += '\n';
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw NoriException("cannot write %s", path);
    const unsigned char magic[8] = { 0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0 };
    fwrite(magic, 1, 8, f);
    const uint16_t hl = (uint16_t) hdr.size();
    fwrite(&hl, 2, 1, f);
    fwrite(hdr.data(), 1, hdr.size(), f);
    fwrite(data, 4, h * w * c, f);
    fclose(f);
}

/* the accumulator file of checkpoint= / resume=: a 3-d '<f4' array in C order, as writeNpy makes it */
static std::vector<float> readNpy(const std::string &path, size_t shape[3]) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw NoriException("cannot open %s", path);
    unsigned char head[10];
    if (fread(head, 1, 10, f) != 10 || memcmp(head, "\x93NUMPY", 6) != 0 || head[6] != 1) { fclose(f); throw NoriException("%s: not a version-1 .npy file", path); }
    const size_t hl = head[8] | ((size_t) head[9] << 8);
    std::string hdr(hl, ' ');
    if (fread(&hdr[0], 1, hl, f) != hl) { fclose(f); throw NoriException("%s: truncated", path); }
    if (hdr.find("'<f4'") == std::string::npos || hdr.find("False") == std::string::npos) { fclose(f); throw NoriException("%s: expected a C-ordered float32 array", path); }
    const size_t a = hdr.find("'shape': ("), b = hdr.find(')', a);
    unsigned long long d[3] = { 0, 0, 0 };
    if (a == std::string::npos || b == std::string::npos || sscanf(hdr.c_str() + a + 10, "%llu, %llu, %llu", &d[0], &d[1], &d[2]) != 3) { fclose(f); throw NoriException("%s: expected a 3-d array", path); }
    for (int k = 0; k < 3; ++k) shape[k] = (size_t) d[k];
    std::vector<float> data(shape[0] * shape[1] * shape[2]);
    const bool ok = fread(data.data(), 4, data.size(), f) == data.size();
    fclose(f);
    if (!ok) throw NoriException("%s: truncated", path);
    return data;
}

static void writePfm(const std::string &path, const float *data, size_t h, size_t w) {
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw NoriException("cannot write %s", path);
    fprintf(f, "PF\n%zu %zu\n-1.0\n", w, h);
    for (size_t y = h; y-- > 0;) fwrite(data + y * w * 3, 4, w * 3, f);      /* PFM is bottom-up */
    fclose(f);
}

namespace nori { void writePng(const std::string &path, const float *data, size_t h, size_t w); }
using nori::writePng;

static void writeImage(const std::string &path, const float *data, size_t h, size_t w) {
    const std::string ext = path.size() > 4 ? path.substr(path.size() - 4) : "";
    if (ext == ".pfm") writePfm(path, data, h, w);
    else if (ext == ".png") writePng(path, data, h, w);
    else writeNpy(path, data, h, w, 3);
}

namespace nori { bool runTTest(NoriObject *root, int device); }

int main(int argc, char **argv) {
    if (argc < 2 || (argc < 3 && !(argc == 2 && strstr(argv[1], ".xml")))) {
        fprintf(stderr, "usage: nori_b200 test.xml | nori_b200 scene.xml out.{npy,pfm,png} [spp=N] [device=D | devices=0,1,...] [seed=S] [launch=N] [snapshot=K] [flat=scene.nbs] [adaptive_cap=N] [bvh=sah|lbvh|host] [checkpoint=film.npy] [resume=film.npy first=K]\n");
        return 2;
    }
    try {
        int spp = -1, device = 0, launch = 0, snapshot = 0, adaptiveCap = 0, treeMode = NB_BVH_DEVICE_SAH;   /* adaptive_cap: AdaptiveSampler's 1000-decision limit (black blocks reach it) */
        uint64_t seed = 0;
        std::string flatOut, checkpointPath, resumePath;     /* checkpoint=: also write the film's accumulators; resume=: add a film written that way */
        int firstSample = 0;                                  /* first=K: the samples [K, spp) — what a resumed render still has to draw */
        std::vector<int> devices;
        for (int i = (argc > 2 && strchr(argv[2], '=')) ? 2 : 3; i < argc; ++i) {      /* a test file takes no output path */
            if (!strncmp(argv[i], "spp=", 4)) spp = atoi(argv[i] + 4);
            else if (!strncmp(argv[i], "device=", 7)) device = atoi(argv[i] + 7);
            else if (!strncmp(argv[i], "devices=", 8)) {
                for (const char *p = argv[i] + 8; *p;) {
                    char *end = nullptr;
                    const long d = strtol(p, &end, 10);
                    if (end == p) throw NoriException("devices=: expected a comma-separated list of device numbers");
                    devices.push_back((int) d);
                    p = *end == ',' ? end + 1 : end;
                }
            }
            else if (!strncmp(argv[i], "seed=", 5)) seed = strtoull(argv[i] + 5, nullptr, 10);
            else if (!strncmp(argv[i], "launch=", 7)) launch = atoi(argv[i] + 7);
            else if (!strncmp(argv[i], "snapshot=", 9)) snapshot = atoi(argv[i] + 9);
            else if (!strncmp(argv[i], "flat=", 5)) flatOut = argv[i] + 5;
            else if (!strncmp(argv[i], "adaptive_cap=", 13)) adaptiveCap = atoi(argv[i] + 13);
            else if (!strncmp(argv[i], "checkpoint=", 11)) checkpointPath = argv[i] + 11;
            else if (!strncmp(argv[i], "resume=", 7)) resumePath = argv[i] + 7;
            else if (!strncmp(argv[i], "first=", 6)) firstSample = atoi(argv[i] + 6);
            else if (!strncmp(argv[i], "bvh=", 4)) {
                const std::string b(argv[i] + 4);
                if (b == "host") treeMode = NB_BVH_IMPORT; else if (b == "lbvh") treeMode = NB_BVH_DEVICE; else if (b == "sah") treeMode = NB_BVH_DEVICE_SAH;
                else throw NoriException("bvh=%s: host (the reference's tree, built on the host), lbvh or sah (built on the device; sah is the default)", b);
            }
        }
        auto t0 = std::chrono::steady_clock::now();
        std::unique_ptr<NoriObject> root(loadFromXML(argv[1]));
        if (root->getClassType() == NoriObject::ETest) {              /* <test type="ttest">: the reference's known-answer harness, on the GPU */
            return runTTest(root.get(), devices.empty() ? device : devices[0]) ? 0 : 1;
        }
        if (root->getClassType() != NoriObject::EScene) throw NoriException("root of %s is not a scene", argv[1]);
        if (argc < 3) throw NoriException("no output file given for the scene %s", argv[1]);
        Scene *scene = static_cast<Scene *>(root.get());
        if (devices.empty()) devices.push_back(device);
        B200State state(devices);
        state.setTreeBuilder(treeMode);
        if (!state.preRender(*scene, seed)) {
            fprintf(stderr, "nori_b200: this scene cannot run on the GPU backend: %s\n", state.whyUnsupported().c_str());
            return 3;
        }
        if (!flatOut.empty()) state.flat().writeNbs(flatOut);
        if (spp < 0) spp = (int) scene->getSampler()->getSampleCount();
        if (launch <= 0) launch = spp;
        double load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        auto t1 = std::chrono::steady_clock::now();
        /* progressive output without a window (the reference shows the running image through GL and saves on demand,
         * render.cpp:725-747): every `snapshot` launches the accumulators so far are normalised and written next to
         * the final image as <out>.<samples>spp.<ext>. With several devices each snapshot reduces the films onto the
         * first one, which is safe to repeat: the reduce moves the samples (nb_film_reduce). */
        const std::string outPath(argv[2]);
        int launches = 0, snapshots = 0;
        const bool adaptive = scene->getSampler()->isAdaptive();
        if (adaptive) launch = 1;                                   /* one round of the sample loop per call */
        uint64_t adaptivePasses = 0;
        if (firstSample < 0 || firstSample > spp) throw NoriException("first=%d outside 0..%d", firstSample, spp);
        for (int s = firstSample; s < spp; s += launch) {
            const int n = std::min(launch, spp - s);
            if (adaptive) adaptivePasses += state.renderAdaptiveRound(4, (uint32_t) scene->getSampler()->getInitialUniform(), (uint32_t) std::max(0, adaptiveCap)).block_passes;   /* block.h:31 */
            else state.renderSubframe((uint32_t) s, (uint32_t) n);
            if (snapshot > 0 && ++launches % snapshot == 0 && s + n < spp) {
                std::vector<float> snap;
                uint32_t sh = 0, sw = 0;
                state.download(snap, true, &sh, &sw);
                const std::string name = outPath.substr(0, outPath.size() - 4) + "." + std::to_string(s + n) + "spp" + outPath.substr(outPath.size() - 4);
                writeImage(name, snap.data(), sh, sw);
                ++snapshots;
            }
        }
        double render_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
        std::vector<float> img;
        uint32_t h = 0, w = 0;
        if (checkpointPath.empty() && resumePath.empty()) state.download(img, true, &h, &w);
        else {
            /* checkpoint / resume (the reference can only interrupt a render, render.cpp:285; SURVEY 5 notes that the film is
             * linear): the accumulators (sum w rgb, sum w) in ImageBlock layout ARE the state of a render, and sample s of
             * pixel p has its own random stream, so samples [0, K) saved + samples [K, n) drawn later = one render of n */
            std::vector<float> film;
            uint32_t rows = 0, cols = 0;
            state.download(film, false, &rows, &cols);
            if (!resumePath.empty()) {
                size_t shape[3];
                const std::vector<float> saved = readNpy(resumePath, shape);
                if (shape[0] != rows || shape[1] != cols || shape[2] != 4) throw NoriException("%s: a %zux%zux%zu film, this render's is %ux%ux4", resumePath, shape[0], shape[1], shape[2], rows, cols);
                for (size_t k = 0; k < film.size(); ++k) film[k] += saved[k];
            }
            if (!checkpointPath.empty()) writeNpy(checkpointPath, film.data(), rows, cols, 4);
            std::vector<float> plain;                                /* ImageBlock::toBitmap: crop the border, divide by the weight */
            state.download(plain, true, &h, &w);
            const uint32_t border = (rows - h) / 2;
            img.assign((size_t) h * w * 3, 0.f);
            for (uint32_t y = 0; y < h; ++y)
                for (uint32_t x = 0; x < w; ++x) {
                    const float *a = film.data() + ((size_t) (y + border) * cols + x + border) * 4;
                    if (std::fabs(a[3]) > 1e-4f) for (int k = 0; k < 3; ++k) img[((size_t) y * w + x) * 3 + k] = a[k] / a[3];
                }
        }
        writeImage(outPath, img.data(), h, w);
        nb_counters c = state.counters();
        printf("{\"scene\": \"%s\", \"width\": %u, \"height\": %u, \"spp\": %d, \"load_seconds\": %.3f, \"seconds\": %.6f, "
               "\"samples\": %llu, \"samples_per_s\": %.1f, \"rays_closest\": %llu, \"rays_any\": %llu, \"rays_per_s\": %.1f, "
               "\"device_ms\": %.3f, \"invalid_samples\": %llu, \"devices\": %zu, \"snapshots\": %d, \"adaptive_block_passes\": %llu}\n",
               argv[1], w, h, spp, load_s, render_s, (unsigned long long) c.samples, c.samples / render_s,
               (unsigned long long) c.rays_closest, (unsigned long long) c.rays_any, (c.rays_closest + c.rays_any) / render_s,
               c.device_ms, (unsigned long long) c.invalid_samples, state.deviceCount(), snapshots, (unsigned long long) adaptivePasses);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "nori_b200: %s\n", e.what());
        return 1;
    }
}
