/* parser.cpp — scene XML -> object tree. Same grammar, tag table and construction order as the
 * reference's loadFromXML (src/utils/parser.cpp:28-378): children are parsed first into a
 * PropertyList, the object is created through NoriObjectFactory from its `type` attribute, child
 * objects are attached with addChild/setParent, then activate() runs. Transform tags compose on
 * the left (`op * transform`), lookat as parser.cpp:345-361. The XML reader below is a small
 * hand-written one (elements, attributes, comments, declarations, entities). */
#include "plugins.h"
#include <cctype>
#include <cstring>
#include <fstream>
#include <set>
#include <sstream>

namespace nori {

namespace {

struct XmlNode {
    std::string name;
    std::vector<std::pair<std::string, std::string>> attrs;
    std::vector<XmlNode> children;
    size_t offset = 0;
    const char *attr(const char *n) const {
        for (auto &a : attrs) if (a.first == n) return a.second.c_str();
        return "";
    }
};

class XmlReader {
public:
    XmlReader(const std::string &text, const std::string &file) : s(text), file(file) {}
    XmlNode parseDocument() {
        skipMisc();
        XmlNode root = parseElement();
        skipMisc();
        if (p < s.size()) fail("trailing content after the root element");
        return root;
    }
private:
    const std::string &s, &file;
    size_t p = 0;
    [[noreturn]] void fail(const char *what) const {
        size_t line = 1, col = 1;
        for (size_t i = 0; i < p && i < s.size(); ++i) { if (s[i] == '\n') { ++line; col = 1; } else ++col; }
        throw NoriException("Error while parsing \"%s\": %s (at line %d, col %d)", file, what, (int) line, (int) col);
    }
    bool starts(const char *t) const { return s.compare(p, strlen(t), t) == 0; }
    void skipWs() { while (p < s.size() && isspace((unsigned char) s[p])) ++p; }
    void skipMisc() {                    /* whitespace, comments, <?...?>, <!DOCTYPE ...> */
        for (;;) {
            skipWs();
            if (starts("<!--")) { size_t e = s.find("-->", p); if (e == std::string::npos) fail("unterminated comment"); p = e + 3; }
            else if (starts("<?")) { size_t e = s.find("?>", p); if (e == std::string::npos) fail("unterminated declaration"); p = e + 2; }
            else if (starts("<!")) { size_t e = s.find('>', p); if (e == std::string::npos) fail("unterminated declaration"); p = e + 1; }
            else return;
        }
    }
    std::string parseName() {
        size_t b = p;
        while (p < s.size() && (isalnum((unsigned char) s[p]) || s[p] == '_' || s[p] == '-' || s[p] == ':' || s[p] == '.')) ++p;
        if (p == b) fail("expected a name");
        return s.substr(b, p - b);
    }
    static std::string unescape(const std::string &v) {
        std::string o;
        for (size_t i = 0; i < v.size(); ++i) {
            if (v[i] != '&') { o += v[i]; continue; }
            if (!v.compare(i, 4, "&lt;")) { o += '<'; i += 3; } else if (!v.compare(i, 4, "&gt;")) { o += '>'; i += 3; }
            else if (!v.compare(i, 5, "&amp;")) { o += '&'; i += 4; } else if (!v.compare(i, 6, "&quot;")) { o += '"'; i += 5; }
            else if (!v.compare(i, 6, "&apos;")) { o += '\''; i += 5; } else o += v[i];
        }
        return o;
    }
    XmlNode parseElement() {
        if (p >= s.size() || s[p] != '<') fail("expected '<'");
        XmlNode n;
        n.offset = p;
        ++p;
        n.name = parseName();
        for (;;) {
            skipWs();
            if (p >= s.size()) fail("unterminated element");
            if (s[p] == '/') { if (!starts("/>")) fail("expected '/>'"); p += 2; return n; }
            if (s[p] == '>') { ++p; break; }
            std::string an = parseName();
            skipWs();
            if (p >= s.size() || s[p] != '=') fail("expected '=' after attribute name");
            ++p; skipWs();
            if (p >= s.size() || (s[p] != '"' && s[p] != '\'')) fail("expected a quoted attribute value");
            const char q = s[p++];
            size_t e = s.find(q, p);
            if (e == std::string::npos) fail("unterminated attribute value");
            std::string value = unescape(s.substr(p, e - p));
            for (char &c : value) if (c == '\t' || c == '\n' || c == '\r') c = ' ';      /* attribute-value normalisation (XML 1.0, 3.3.3; pugixml's parse_wconv_attribute) */
            n.attrs.emplace_back(an, value);
            p = e + 1;
        }
        for (;;) {                        /* content: child elements; text is ignored as pugixml's node iteration does */
            skipMisc();
            if (p >= s.size()) fail("unterminated element");
            if (starts("</")) {
                p += 2;
                std::string cn = parseName();
                if (cn != n.name) fail("mismatched closing tag");
                skipWs();
                if (p >= s.size() || s[p] != '>') fail("expected '>'");
                ++p;
                return n;
            }
            if (s[p] == '<') n.children.push_back(parseElement());
            else { while (p < s.size() && s[p] != '<') ++p; }
        }
    }
};

enum ETag { EScene = NoriObject::EScene, EShape = NoriObject::EShape, ETexture = NoriObject::ETexture,
            EVolume = NoriObject::EVolume, EBSDF = NoriObject::EBSDF, EPhaseFunction = NoriObject::EPhaseFunction,
            EEmitter = NoriObject::EEmitter, EMedium = NoriObject::EMedium, ECamera = NoriObject::ECamera,
            EIntegrator = NoriObject::EIntegrator, ESampler = NoriObject::ESampler,
            EPixelSampler = NoriObject::EPixelSampler, EDenoiser = NoriObject::EDenoiser, ETest = NoriObject::ETest,
            EReconstructionFilter = NoriObject::EReconstructionFilter, ERenderer = NoriObject::ERenderer,
            EBoolean = NoriObject::EClassTypeCount, EInteger, EFloat, EString, EPoint, EVector, EColor, ETransform,
            ETranslate, EMatrix, ERotate, EScale, ELookAt, EInvalid };

const std::map<std::string, ETag> &tagTable() {         /* parser.cpp:100-130 */
    static const std::map<std::string, ETag> t = {
        {"scene", EScene}, {"shape", EShape}, {"mesh", EShape}, {"texture", ETexture}, {"volume", EVolume}, {"bsdf", EBSDF},
        {"phase", EPhaseFunction}, {"emitter", EEmitter}, {"medium", EMedium}, {"camera", ECamera},
        {"integrator", EIntegrator}, {"sampler", ESampler}, {"pxsampler", EPixelSampler}, {"denoiser", EDenoiser},
        {"test", ETest}, {"rfilter", EReconstructionFilter}, {"renderer", ERenderer},
        {"boolean", EBoolean}, {"integer", EInteger}, {"float", EFloat}, {"string", EString}, {"point", EPoint},
        {"vector", EVector}, {"color", EColor}, {"transform", ETransform}, {"translate", ETranslate},
        {"matrix", EMatrix}, {"rotate", ERotate}, {"scale", EScale}, {"lookat", ELookAt} };
    return t;
}

void checkAttributes(const XmlNode &node, std::set<std::string> attrs) {
    for (auto &a : node.attrs) {
        auto it = attrs.find(a.first);
        if (it == attrs.end()) throw NoriException("unexpected attribute \"%s\" in \"%s\"", a.first, node.name);
        attrs.erase(it);
    }
    if (!attrs.empty()) throw NoriException("missing attribute \"%s\" in \"%s\"", *attrs.begin(), node.name);
}

struct Parser {
    std::string filename;
    Matrix4f transform = Matrix4f::identity();

    NoriObject *parseTag(const XmlNode &node, PropertyList &list, int parentTag) {
        auto it = tagTable().find(node.name);
        if (it == tagTable().end()) throw NoriException("unexpected tag \"%s\"", node.name);
        const int tag = it->second;
        const bool hasParent = parentTag != EInvalid;
        const bool parentIsObject = hasParent && parentTag < NoriObject::EClassTypeCount;
        const bool currentIsObject = tag < NoriObject::EClassTypeCount;
        const bool parentIsTransform = parentTag == ETransform;
        const bool currentIsTransformOp = tag == ETranslate || tag == ERotate || tag == EScale || tag == ELookAt || tag == EMatrix;
        if (!hasParent && !currentIsObject) throw NoriException("root element \"%s\" must be a Nori object", node.name);
        if (parentIsTransform != currentIsTransformOp) throw NoriException("transform nodes can only contain transform operations");
        if (hasParent && !parentIsObject && !(parentIsTransform && currentIsTransformOp))
            throw NoriException("node \"%s\" requires a Nori object as parent", node.name);
        if (tag == ETransform) transform = Matrix4f::identity();

        PropertyList props;
        std::vector<NoriObject *> children;
        for (const XmlNode &ch : node.children) {
            NoriObject *child = parseTag(ch, props, tag);
            if (child) children.push_back(child);
        }
        NoriObject *result = nullptr;
        if (currentIsObject) {
            std::string type = node.attr("type");
            if (tag == EScene && type.empty()) type = "scene";
            /* like the reference (parser.cpp:207-227) an object's `name` attribute reaches its constructor as the
             * string property "name" and its parent through getIdName() */
            props.setString("name", node.attr("name"));
            result = NoriObjectFactory::createInstance(type, props);
            result->setIdName(node.attr("name"));
            if ((int) result->getClassType() != tag)
                throw NoriException("Unexpectedly constructed an object of type <%s> (expected type <%s>): %s",
                                    NoriObject::classTypeName(result->getClassType()),
                                    NoriObject::classTypeName((NoriObject::EClassType) tag), result->toString());
            for (NoriObject *ch : children) { result->addChild(ch); ch->setParent(result); }
            result->activate();
            return result;
        }
        switch (tag) {
            case EString: checkAttributes(node, {"name", "value"}); list.setString(node.attr("name"), node.attr("value")); break;
            case EFloat: checkAttributes(node, {"name", "value"}); list.setFloat(node.attr("name"), toFloat(node.attr("value"))); break;
            case EInteger: checkAttributes(node, {"name", "value"}); list.setInteger(node.attr("name"), toInt(node.attr("value"))); break;
            case EBoolean: checkAttributes(node, {"name", "value"}); list.setBoolean(node.attr("name"), toBool(node.attr("value"))); break;
            case EPoint: case EVector: {
                checkAttributes(node, {"name", "value"});
                const size_t n = tokenize(node.attr("value")).size();
                if (n == 2) { Vector2f v = toVector2f(node.attr("value")); tag == EPoint ? list.setPoint2(node.attr("name"), v) : list.setVector2(node.attr("name"), v); }
                else if (n == 3) { Vector3f v = toVector3f(node.attr("value")); tag == EPoint ? list.setPoint3(node.attr("name"), v) : list.setVector3(node.attr("name"), v); }
                else throw NoriException("%s %s (value: %s) is not of size 2 or 3", tag == EPoint ? "Point" : "Vector", node.attr("name"), node.attr("value"));
            } break;
            case EColor: checkAttributes(node, {"name", "value"}); list.setColor(node.attr("name"), toVector3f(node.attr("value"))); break;
            case ETransform: checkAttributes(node, {"name"}); list.setTransform(node.attr("name"), Transform(transform)); break;
            case ETranslate: {
                checkAttributes(node, {"value"});
                Vector3f v = toVector3f(node.attr("value"));
                Matrix4f t = Matrix4f::identity(); t(0, 3) = v.x(); t(1, 3) = v.y(); t(2, 3) = v.z();
                transform = t * transform;
            } break;
            case EMatrix: {
                checkAttributes(node, {"value"});
                std::vector<std::string> tok = tokenize(node.attr("value"));
                if (tok.size() != 16) throw NoriException("Expected 16 values");
                Matrix4f m;
                for (int i = 0; i < 16; ++i) m.m[i] = toFloat(tok[i]);
                transform = m * transform;
            } break;
            case EScale: {
                checkAttributes(node, {"value"});
                Vector3f v = toVector3f(node.attr("value"));
                Matrix4f t = Matrix4f::identity(); t(0, 0) = v.x(); t(1, 1) = v.y(); t(2, 2) = v.z();
                transform = t * transform;
            } break;
            case ERotate: {                                   /* Eigen::AngleAxis::toRotationMatrix */
                checkAttributes(node, {"angle", "axis"});
                const float angle = degToRad(toFloat(node.attr("angle")));
                const Vector3f a = toVector3f(node.attr("axis"));
                const float s = std::sin(angle), c = std::cos(angle);
                const Vector3f sin_axis = a * s, cos1_axis = a * (1.f - c);
                Matrix4f r = Matrix4f::identity();
                float tmp;
                tmp = cos1_axis.x() * a.y(); r(0, 1) = tmp - sin_axis.z(); r(1, 0) = tmp + sin_axis.z();
                tmp = cos1_axis.x() * a.z(); r(0, 2) = tmp + sin_axis.y(); r(2, 0) = tmp - sin_axis.y();
                tmp = cos1_axis.y() * a.z(); r(1, 2) = tmp - sin_axis.x(); r(2, 1) = tmp + sin_axis.x();
                r(0, 0) = cos1_axis.x() * a.x() + c; r(1, 1) = cos1_axis.y() * a.y() + c; r(2, 2) = cos1_axis.z() * a.z() + c;
                transform = r * transform;
            } break;
            case ELookAt: {                                   /* parser.cpp:345-361 */
                checkAttributes(node, {"origin", "target", "up"});
                const Vector3f origin = toVector3f(node.attr("origin")), target = toVector3f(node.attr("target")), up = toVector3f(node.attr("up"));
                const Vector3f dir = (target - origin).normalized();
                const Vector3f left = up.normalized().cross(dir).normalized();
                const Vector3f newUp = dir.cross(left).normalized();
                Matrix4f t = Matrix4f::identity();
                for (int i = 0; i < 3; ++i) { t(i, 0) = left[i]; t(i, 1) = newUp[i]; t(i, 2) = dir[i]; t(i, 3) = origin[i]; }
                transform = t * transform;
            } break;
            default: throw NoriException("Unhandled element \"%s\"", node.name);
        }
        return nullptr;
    }
};

} // namespace

NoriObject *loadFromXML(const std::string &filename) {
    std::ifstream is(filename, std::ios::binary);
    if (!is) throw NoriException("Error while parsing \"%s\": cannot open the file", filename);
    std::stringstream ss;
    ss << is.rdbuf();
    const std::string text = ss.str();
    XmlReader reader(text, filename);
    XmlNode root = reader.parseDocument();
    Parser parser;
    parser.filename = filename;
    const size_t slash = filename.find_last_of('/');
    setResolveRoot(slash == std::string::npos ? std::string(".") : filename.substr(0, slash));   /* render.cpp:129 */
    PropertyList list;
    try {
        return parser.parseTag(root, list, EInvalid);
    } catch (const NoriException &e) {
        const std::string msg = e.what();
        if (msg.rfind("Error while parsing", 0) == 0) throw;
        throw NoriException("Error while parsing \"%s\": %s", filename, msg);
    }
}

} // namespace nori
