/* ttest.cpp — <test type="ttest"> and <test type="chi2test">: the reference's Student-t known-answer harness (src/utils/ttest.cpp:60-268), driven
 * through the GPU backend. The reference runs it on its CPU when the file is loaded: 100 000 camera paths (or BSDF samples)
 * per case, mean and variance of their luminance, a two-sided t-test against the value stored in the file, Sidak-corrected
 * for the number of cases (ext/hypothesis/hypothesis.h:314-346). Here:
 *   scenes — each <scene> is flattened and rendered RENDERS times with independent seeds at sampleCount / RENDERS samples
 *            per pixel through a box filter (every weight 1, so a pixel is the plain mean of Li, which is what the reference
 *            averages); the device returns means, not single samples, so the variance of the mean is estimated from the
 *            RENDERS means and the test has RENDERS - 1 degrees of freedom;
 *   BSDFs  — sampleCount draws of BSDF::sample through nb_bsdf_probe on a one-sphere scene that carries the BSDF; mean and
 *            variance over the single samples as in the reference (Knuth's online update), sampleCount - 1 degrees of freedom.
 * nori_b200 test.xml runs it (host/main.cpp) and prints the reference's report lines; the exit code is 0 when every case
 * is accepted. Nothing here runs on load: parsing a test file needs no GPU. */
#include "b200state.h"
#include <memory>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <sstream>

namespace nori {

namespace {

/* regularised incomplete beta function I_x(a, b) by its continued fraction (modified Lentz) */
double betaFraction(double a, double b, double x) {
    const double tiny = 1e-300;
    double qab = a + b, qap = a + 1.0, qam = a - 1.0, c = 1.0, d = 1.0 - qab * x / qap;
    if (std::fabs(d) < tiny) d = tiny;
    d = 1.0 / d;
    double h = d;
    for (int m = 1; m <= 10000; ++m) {
        const int m2 = 2 * m;
        double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
        d = 1.0 + aa * d; if (std::fabs(d) < tiny) d = tiny;
        c = 1.0 + aa / c; if (std::fabs(c) < tiny) c = tiny;
        d = 1.0 / d; h *= d * c;
        aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
        d = 1.0 + aa * d; if (std::fabs(d) < tiny) d = tiny;
        c = 1.0 + aa / c; if (std::fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        const double del = d * c;
        h *= del;
        if (std::fabs(del - 1.0) < 1e-15) break;
    }
    return h;
}
double incompleteBeta(double a, double b, double x) {
    if (x <= 0.0) return 0.0;
    if (x >= 1.0) return 1.0;
    const double front = std::exp(std::lgamma(a + b) - std::lgamma(a) - std::lgamma(b) + a * std::log(x) + b * std::log(1.0 - x));
    return x < (a + 1.0) / (a + b + 2.0) ? front * betaFraction(a, b, x) / a : 1.0 - front * betaFraction(b, a, 1.0 - x) / b;
}
/* two-sided p-value of Student's t with `dof` degrees of freedom: 2 (1 - cdf(|t|)) = I_{dof / (dof + t^2)}(dof / 2, 1 / 2) */
double twoSidedP(double t, double dof) { return incompleteBeta(0.5 * dof, 0.5, dof / (dof + t * t)); }

/* hypothesis::students_t_test (hypothesis.h:314-346): same report, same decision; `n` = how many values the mean is of,
 * `floorVar` = the reference's floor of 1e-5 on the variance of its 100 000 single samples, scaled to what `variance` is the
 * variance of */
std::pair<bool, std::string> studentsTTest(double mean, double variance, double reference, int n, double floorVar, double significance, int numTests) {
    std::ostringstream oss;
    const double t = std::fabs(mean - reference) * std::sqrt(n / std::max(variance, floorVar));
    const int dof = n - 1;
    oss << "Sample mean = " << mean << " (reference value = " << reference << ")\n";
    oss << "Sample variance = " << variance << "\n";
    oss << "t-statistic = " << t << " (d.o.f. = " << dof << ")\n";
    const double pval = twoSidedP(t, dof);
    const double alpha = 1.0 - std::pow(1.0 - significance, 1.0 / numTests);
    const bool rejected = pval < alpha || !std::isfinite(pval);
    oss << (rejected ? "***** Rejected ***** the null hypothesis (p-value = " : "Accepted the null hypothesis (p-value = ") << pval
        << ", significance level = " << alpha << ")\n";
    return { !rejected, oss.str() };
}

struct Pcg32 {                                      /* ext/pcg32: the harness's own generator for the BSDF samples */
    uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
    uint32_t nextUInt() {
        const uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        const uint32_t x = (uint32_t) (((old >> 18u) ^ old) >> 27u), r = (uint32_t) (old >> 59u);
        return (x >> r) | (x << ((0u - r) & 31u));
    }
    float nextFloat() { union { uint32_t u; float f; } v; v.u = (nextUInt() >> 9) | 0x3f800000u; return v.f - 1.0f; }
};

double luminanceOf(const float *rgb) { return (double) (rgb[0] * 0.212671f + rgb[1] * 0.715160f + rgb[2] * 0.072169f); }

} // namespace

class StudentsTTest : public NoriObject {
public:
    enum { RENDERS = 24 };
    explicit StudentsTTest(const PropertyList &p) {                      /* ttest.cpp:62-79 */
        m_significanceLevel = p.getFloat("significanceLevel", 0.01f);
        for (const std::string &a : tokenize(p.getString("angles", ""))) m_angles.push_back(toFloat(a));
        for (const std::string &r : tokenize(p.getString("references", ""))) m_references.push_back(toFloat(r));
        m_sampleCount = p.getInteger("sampleCount", 100000);
    }
    ~StudentsTTest() override {
        for (BSDF *b : m_bsdfs) delete b;
        for (Scene *s : m_scenes) delete s;
    }
    EClassType getClassType() const override { return ETest; }
    void addChild(NoriObject *obj) override {                            /* ttest.cpp:130-145 */
        switch (obj->getClassType()) {
            case EBSDF: m_bsdfs.push_back(static_cast<BSDF *>(obj)); break;
            case EScene: m_scenes.push_back(static_cast<Scene *>(obj)); break;
            default: throw NoriException("StudentsTTest::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        }
    }
    std::string toString() const override {
        return NoriException::format("StudentsTTest[\n  significanceLevel = %f,\n  sampleCount= %d\n]", m_significanceLevel, m_sampleCount);
    }

    /* ttest.cpp:148-249; returns true when every case was accepted */
    bool execute(int device, int *passedOut = nullptr, int *totalOut = nullptr) {
        int total = 0, passed = 0;
        if (!m_bsdfs.empty()) {
            if (m_references.size() * m_bsdfs.size() != m_angles.size()) throw NoriException("Specified a different number of angles and reference values!");
            if (!m_scenes.empty()) throw NoriException("Cannot test BSDFs and scenes at the same time!");
            Pcg32 random;
            size_t ctr = 0;
            for (BSDF *&bsdf : m_bsdfs) {
                /* the BSDF on a unit sphere in front of a one-pixel camera: a scene the boundary can take; the probe then
                 * calls BSDF::sample on the device. The sphere owns the BSDF from here on. */
                const std::string name = bsdf->toString();
                std::unique_ptr<Scene> scene(static_cast<Scene *>(NoriObjectFactory::createInstance("scene")));
                PropertyList cam; cam.setInteger("width", 1); cam.setInteger("height", 1);
                NoriObject *camera = NoriObjectFactory::createInstance("perspective", cam); camera->activate();
                NoriObject *integrator = NoriObjectFactory::createInstance("normals"); integrator->activate();
                PropertyList sp; sp.setPoint3("center", Point3f(0.f, 0.f, 5.f)); sp.setFloat("radius", 1.f);
                NoriObject *sphere = NoriObjectFactory::createInstance("sphere", sp);
                sphere->addChild(bsdf); bsdf = nullptr;
                sphere->activate();
                scene->addChild(camera); scene->addChild(integrator); scene->addChild(sphere);
                scene->activate();
                B200State state(device);
                if (!state.preRender(*scene)) throw NoriException("ttest: %s", state.whyUnsupported());
                for (size_t i = 0; i < m_references.size(); ++i) {
                    const float angle = m_angles[i], reference = m_references[ctr++];
                    printf("------------------------------------------------------\nTesting (angle=%g): %s\n", angle, name.c_str());
                    ++total;
                    const float theta = degToRad(angle);
                    const float wi[3] = { std::sin(theta), 0.f, std::cos(theta) };      /* sphericalDirection(theta, 0) */
                    printf("Drawing %d samples .. \n", m_sampleCount);
                    std::vector<float> in((size_t) m_sampleCount * 10, 0.f), out((size_t) m_sampleCount * 12);
                    for (int k = 0; k < m_sampleCount; ++k) {
                        float *r = in.data() + (size_t) k * 10;
                        r[0] = wi[0]; r[1] = wi[1]; r[2] = wi[2]; r[5] = 1.f;
                        r[6] = random.nextFloat(); r[7] = random.nextFloat();
                    }
                    if (nb_bsdf_probe(state.handle(), 0, in.data(), (uint64_t) m_sampleCount, out.data()) != NB_OK)
                        throw NoriException("ttest: %s", nb_last_error(state.handle()));
                    double mean = 0, variance = 0;
                    for (int k = 0; k < m_sampleCount; ++k) {                           /* Knuth's online update, ttest.cpp:178-184 */
                        const double result = luminanceOf(out.data() + (size_t) k * 12 + 4);
                        const double delta = result - mean;
                        mean += delta / (double) (k + 1);
                        variance += delta * (result - mean);
                    }
                    variance /= m_sampleCount - 1;
                    const auto result = studentsTTest(mean, variance, reference, m_sampleCount, 1e-5, m_significanceLevel, (int) m_references.size());
                    if (result.first) ++passed;
                    printf("%s\n", result.second.c_str());
                }
            }
        } else {
            if (m_references.size() != m_scenes.size()) throw NoriException("Specified a different number of scenes and reference values!");
            const uint32_t spp = (uint32_t) std::max(1, (m_sampleCount * 2 + RENDERS - 1) / RENDERS);    /* twice the reference's paths in all */
            size_t ctr = 0;
            for (Scene *scene : m_scenes) {
                const float reference = m_references[ctr++];
                printf("------------------------------------------------------\nTesting scene: %s\n", scene->toString().c_str());
                ++total;
                printf("Generating %d x %u paths.. \n", (int) RENDERS, spp);
                B200State state(device);
                state.setBoxFilter(true);
                double means[RENDERS];
                for (int k = 0; k < RENDERS; ++k) {
                    if (!state.preRender(*scene, 1000 + (uint64_t) k)) throw NoriException("ttest: %s", state.whyUnsupported());
                    state.renderSubframe(0, spp);
                    std::vector<float> img;
                    uint32_t h = 0, w = 0;
                    state.download(img, true, &h, &w);
                    double m = 0;
                    for (size_t px = 0; px < (size_t) h * w; ++px) m += luminanceOf(img.data() + 3 * px);
                    means[k] = m / ((double) h * w);
                }
                double mean = 0, variance = 0;
                for (int k = 0; k < RENDERS; ++k) mean += means[k] / RENDERS;
                for (int k = 0; k < RENDERS; ++k) variance += (means[k] - mean) * (means[k] - mean) / (RENDERS - 1);
                /* the reference floors the variance of its 100 000 single samples at 1e-5, i.e. the standard error of its mean at
                 * 1e-5 (the stored values are rounded to about that): the same floor for the RENDERS means */
                auto result = studentsTTest(mean, variance, reference, RENDERS, RENDERS * 1e-10, m_significanceLevel, (int) m_references.size());
                /* a film pixel is a sum of fp32 splats: with a constant integrand (the furnace scenes) every add rounds the same way
                 * and the mean can sit 1e-4 (relative) off; that close is accepted without the t statistic */
                if (!result.first && std::fabs(mean - reference) <= 2e-4 * std::fabs(reference)) {
                    result.first = true;
                    result.second += "(within 2e-4 of the reference value: accepted)\n";
                }
                if (result.first) ++passed;
                printf("%s\n", result.second.c_str());
            }
        }
        printf("Passed %d/%d tests.\n", passed, total);
        if (passedOut) *passedOut = passed;
        if (totalOut) *totalOut = total;
        return passed == total;
    }
private:
    std::vector<BSDF *> m_bsdfs;
    std::vector<Scene *> m_scenes;
    std::vector<float> m_angles, m_references;
    float m_significanceLevel;
    int m_sampleCount;
};
NORI_REGISTER_CLASS(StudentsTTest, "ttest")

/* <test type="chi2test"> (src/utils/chi2test.cpp:44-250): does BSDF::sample draw from the density BSDF::pdf claims? Samples and
 * densities both come from the device through nb_bsdf_probe: the histogram of sampleCount sampled directions over
 * cos(theta) x phi cells against the density integrated over each cell. The reference integrates with an adaptive Simpson rule
 * (hypothesis.h); here every cell gets a fixed composite Simpson rule of 65 x 65 points — one batched probe call per test. */
class ChiSquareTest : public NoriObject {
public:
    explicit ChiSquareTest(const PropertyList &p) {                      /* chi2test.cpp:44-73 */
        m_significanceLevel = p.getFloat("significanceLevel", 0.01f);
        m_cosThetaResolution = p.getInteger("resolution", 10);
        m_minExpFrequency = p.getInteger("minExpFrequency", 5);
        m_sampleCount = p.getInteger("sampleCount", -1);
        m_testCount = p.getInteger("testCount", 5);
        m_phiResolution = 2 * m_cosThetaResolution;
        if (m_sampleCount < 0) m_sampleCount = m_cosThetaResolution * m_phiResolution * 5000;
    }
    ~ChiSquareTest() override { for (BSDF *b : m_bsdfs) delete b; }
    EClassType getClassType() const override { return ETest; }
    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != EBSDF) throw NoriException("ChiSquareTest::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        m_bsdfs.push_back(static_cast<BSDF *>(obj));
    }
    std::string toString() const override {
        return NoriException::format("ChiSquareTest[\n  thetaResolution = %d,\n  phiResolution = %d,\n  minExpFrequency = %d,\n  sampleCount = %d,\n  testCount = %d,\n  significanceLevel = %f\n]",
                                     m_cosThetaResolution, m_phiResolution, m_minExpFrequency, m_sampleCount, m_testCount, m_significanceLevel);
    }

    bool execute(int device) {                                           /* chi2test.cpp:131-233 */
        int passed = 0, total = 0;
        const int res = m_cosThetaResolution * m_phiResolution;
        Pcg32 random;
        std::vector<double> obs((size_t) res), exp((size_t) res);
        const int Q = 65;                                                /* Simpson points per cell and axis */
        for (BSDF *&bsdf : m_bsdfs) {
            const std::string name = bsdf->toString();
            std::unique_ptr<Scene> scene(static_cast<Scene *>(NoriObjectFactory::createInstance("scene")));
            PropertyList cam; cam.setInteger("width", 1); cam.setInteger("height", 1);
            NoriObject *camera = NoriObjectFactory::createInstance("perspective", cam); camera->activate();
            NoriObject *integrator = NoriObjectFactory::createInstance("normals"); integrator->activate();
            PropertyList sp; sp.setPoint3("center", Point3f(0.f, 0.f, 5.f)); sp.setFloat("radius", 1.f);
            NoriObject *sphere = NoriObjectFactory::createInstance("sphere", sp);
            sphere->addChild(bsdf); bsdf = nullptr;                      /* the sphere owns it now */
            sphere->activate();
            scene->addChild(camera); scene->addChild(integrator); scene->addChild(sphere);
            scene->activate();
            B200State state(device);
            if (!state.preRender(*scene)) throw NoriException("chi2test: %s", state.whyUnsupported());
            for (int l = 0; l < m_testCount; ++l) {
                std::fill(obs.begin(), obs.end(), 0.0); std::fill(exp.begin(), exp.end(), 0.0);
                printf("------------------------------------------------------\nTesting: %s\n", name.c_str());
                ++total;
                const float cosTheta = random.nextFloat(), sinTheta = std::sqrt(std::max(0.f, 1.f - cosTheta * cosTheta));
                const float phi = 2.0f * (float) M_PI * random.nextFloat();
                const float wi[3] = { std::cos(phi) * sinTheta, std::sin(phi) * sinTheta, cosTheta };
                printf("Accumulating %d samples into a %dx%d contingency table .. ", m_sampleCount, m_cosThetaResolution, m_phiResolution);
                fflush(stdout);
                {
                    std::vector<float> in((size_t) m_sampleCount * 10, 0.f), out((size_t) m_sampleCount * 12);
                    for (int i = 0; i < m_sampleCount; ++i) {
                        float *r = in.data() + (size_t) i * 10;
                        r[0] = wi[0]; r[1] = wi[1]; r[2] = wi[2]; r[5] = 1.f; r[6] = random.nextFloat(); r[7] = random.nextFloat();
                    }
                    if (nb_bsdf_probe(state.handle(), 0, in.data(), (uint64_t) m_sampleCount, out.data()) != NB_OK)
                        throw NoriException("chi2test: %s", nb_last_error(state.handle()));
                    for (int i = 0; i < m_sampleCount; ++i) {
                        const float *o = out.data() + (size_t) i * 12;
                        if (o[4] == 0.f && o[5] == 0.f && o[6] == 0.f) continue;      /* a failed sample (chi2test.cpp:168-169) */
                        const int cb = std::min(std::max(0, (int) std::floor((o[9] * 0.5f + 0.5f) * m_cosThetaResolution)), m_cosThetaResolution - 1);
                        float scaledPhi = std::atan2(o[8], o[7]) * (float) (0.5 / M_PI);
                        if (scaledPhi < 0) scaledPhi += 1;
                        const int pb = std::min(std::max(0, (int) std::floor(scaledPhi * m_phiResolution)), m_phiResolution - 1);
                        obs[(size_t) cb * m_phiResolution + pb] += 1;
                    }
                }
                printf("done.\nIntegrating expected frequencies .. ");
                fflush(stdout);
                {
                    const size_t perCell = (size_t) Q * Q, n = perCell * res;
                    std::vector<float> in(n * 10, 0.f), out(n * 12);
                    for (int i = 0; i < m_cosThetaResolution; ++i)
                        for (int j = 0; j < m_phiResolution; ++j) {
                            const double c0 = -1.0 + i * 2.0 / m_cosThetaResolution, c1 = -1.0 + (i + 1) * 2.0 / m_cosThetaResolution;
                            const double p0 = j * 2 * M_PI / m_phiResolution, p1 = (j + 1) * 2 * M_PI / m_phiResolution;
                            float *r = in.data() + ((size_t) i * m_phiResolution + j) * perCell * 10;
                            for (int a = 0; a < Q; ++a)
                                for (int b = 0; b < Q; ++b, r += 10) {
                                    const double ct = c0 + (c1 - c0) * a / (Q - 1), ph = p0 + (p1 - p0) * b / (Q - 1), st = std::sqrt(std::max(0.0, 1 - ct * ct));
                                    r[0] = wi[0]; r[1] = wi[1]; r[2] = wi[2];
                                    r[3] = (float) (st * std::cos(ph)); r[4] = (float) (st * std::sin(ph)); r[5] = (float) ct;
                                }
                        }
                    if (nb_bsdf_probe(state.handle(), 0, in.data(), (uint64_t) n, out.data()) != NB_OK)
                        throw NoriException("chi2test: %s", nb_last_error(state.handle()));
                    auto w = [&](int k) { return k == 0 || k == Q - 1 ? 1.0 : (k & 1) ? 4.0 : 2.0; };
                    for (int cell = 0; cell < res; ++cell) {
                        const float *o = out.data() + (size_t) cell * perCell * 12;
                        double sum = 0;
                        for (int a = 0; a < Q; ++a)
                            for (int b = 0; b < Q; ++b) sum += w(a) * w(b) * (double) o[((size_t) a * Q + b) * 12 + 3];      /* BSDF::pdf(wi, wo) */
                        const double hc = (2.0 / m_cosThetaResolution) / (Q - 1), hp = (2 * M_PI / m_phiResolution) / (Q - 1);
                        exp[(size_t) cell] = sum * hc * hp / 9.0 * m_sampleCount;
                    }
                }
                printf("done.\n");
                const auto result = chi2Test(res, obs.data(), exp.data(), m_sampleCount, m_minExpFrequency, m_significanceLevel, m_testCount * (int) m_bsdfs.size());
                if (result.first) ++passed;
                printf("%s\n", result.second.c_str());
            }
        }
        printf("Passed %d/%d tests.\n", passed, total);
        return passed == total;
    }
private:
    /* hypothesis::chi2_test (hypothesis.h:140-236) */
    static std::pair<bool, std::string> chi2Test(int nCells, const double *obs, const double *exp, int sampleCount, double minExp, double significance, int numTests) {
        std::vector<int> order((size_t) nCells);
        for (int i = 0; i < nCells; ++i) order[(size_t) i] = i;
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return exp[a] < exp[b]; });
        double pooledObs = 0, pooledExp = 0, chsq = 0;
        int pooledCells = 0, dof = 0;
        std::ostringstream oss;
        for (int c : order) {
            if (exp[c] == 0) {
                if (obs[c] > sampleCount * 1e-5) {
                    oss << "Encountered " << obs[c] << " samples in a cell with expected frequency 0. Rejecting the null hypothesis!\n";
                    return { false, oss.str() };
                }
            } else if (exp[c] < minExp || (pooledExp > 0 && pooledExp < minExp)) {
                pooledObs += obs[c]; pooledExp += exp[c]; ++pooledCells;
            } else {
                const double diff = obs[c] - exp[c];
                chsq += diff * diff / exp[c];
                ++dof;
            }
        }
        if (pooledExp > 0 || pooledObs > 0) {
            oss << "Pooled " << pooledCells << " to ensure sufficiently high expected cell frequencies (>" << minExp << ")\n";
            const double diff = pooledObs - pooledExp;
            chsq += diff * diff / pooledExp;
            ++dof;
        }
        dof -= 1;
        if (dof <= 0) { oss << "The number of degrees of freedom (" << dof << ") is too low!\n"; return { false, oss.str() }; }
        oss << "Chi^2 statistic = " << chsq << " (d.o.f. = " << dof << ")\n";
        const double pval = upperGamma(0.5 * dof, 0.5 * chsq);           /* 1 - chi2_cdf(chsq, dof) */
        const double alpha = 1.0 - std::pow(1.0 - significance, 1.0 / numTests);
        const bool rejected = pval < alpha || !std::isfinite(pval);
        oss << (rejected ? "***** Rejected ***** the null hypothesis (p-value = " : "Accepted the null hypothesis (p-value = ") << pval
            << ", significance level = " << alpha << ")\n";
        return { !rejected, oss.str() };
    }
public:
    static double upperTail(double a, double x) { return upperGamma(a, x); }
private:
    /* regularised upper incomplete gamma function Q(a, x): series below a + 1, continued fraction above */
    static double upperGamma(double a, double x) {
        if (x <= 0) return 1.0;
        const double lg = std::lgamma(a);
        if (x < a + 1.0) {
            double ap = a, sum = 1.0 / a, del = sum;
            for (int n = 0; n < 100000; ++n) { ap += 1.0; del *= x / ap; sum += del; if (std::fabs(del) < std::fabs(sum) * 1e-16) break; }
            return 1.0 - sum * std::exp(-x + a * std::log(x) - lg);
        }
        const double tiny = 1e-300;
        double b = x + 1.0 - a, c = 1.0 / tiny, d = 1.0 / b, h = d;
        for (int i = 1; i < 100000; ++i) {
            const double an = -i * (i - a);
            b += 2.0;
            d = an * d + b; if (std::fabs(d) < tiny) d = tiny;
            c = b + an / c; if (std::fabs(c) < tiny) c = tiny;
            d = 1.0 / d;
            const double del = d * c;
            h *= del;
            if (std::fabs(del - 1.0) < 1e-16) break;
        }
        return std::exp(-x + a * std::log(x) - lg) * h;
    }
    std::vector<BSDF *> m_bsdfs;
    float m_significanceLevel;
    int m_cosThetaResolution, m_phiResolution, m_minExpFrequency, m_sampleCount, m_testCount;
};
NORI_REGISTER_CLASS(ChiSquareTest, "chi2test")

/* for the tests: the two tail probabilities the harness decides with */
double studentTwoSidedP(double t, double dof) { return twoSidedP(t, dof); }
double chi2UpperP(double chsq, double dof) { return ChiSquareTest::upperTail(0.5 * dof, 0.5 * chsq); }

bool runTTest(NoriObject *root, int device) {
    if (auto *t = dynamic_cast<StudentsTTest *>(root)) return t->execute(device);
    if (auto *c = dynamic_cast<ChiSquareTest *>(root)) return c->execute(device);
    throw NoriException("this <test> object has no harness in the GPU backend");
}

} // namespace nori
