/* capi.cpp — a small C surface over the host library so the Python tests can drive it:
 * load a scene XML, flatten it (plugins' hooks + BVH build), write the result as .nbs. No GPU. */
#include "plugins.h"
#include <memory>

static thread_local std::string g_error;

extern "C" {

const char *nbh_last_error(void) { return g_error.c_str(); }

/* XML -> flattened scene file. Returns 0, or -1 (error text in nbh_last_error), or -5 when a
 * plugin of the scene has no GPU kernel (NB_E_UNSUPPORTED). */
int nbh_flatten_xml(const char *xml_path, const char *nbs_path) {
    try {
        std::unique_ptr<nori::NoriObject> root(nori::loadFromXML(xml_path));
        if (root->getClassType() != nori::NoriObject::EScene) { g_error = "root element is not a <scene>"; return -1; }
        nori::FlatScene flat;
        std::string why;
        if (!static_cast<nori::Scene *>(root.get())->flatten(flat, &why)) { g_error = why; return NB_E_UNSUPPORTED; }
        flat.writeNbs(nbs_path);
        return 0;
    } catch (const std::exception &e) {
        g_error = e.what();
        return -1;
    }
}

/* Parse only: reports the object tree's summary or the parser's error */
int nbh_parse_xml(const char *xml_path, char *summary, int summary_len) {
    try {
        std::unique_ptr<nori::NoriObject> root(nori::loadFromXML(xml_path));
        if (summary && summary_len > 0) snprintf(summary, summary_len, "%s", root->toString().c_str());
        return 0;
    } catch (const std::exception &e) {
        g_error = e.what();
        return -1;
    }
}

}
