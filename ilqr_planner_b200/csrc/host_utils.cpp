// host_utils.cpp — host-side builders behind the C ABI: URDF -> KDL chain descriptor and the control-primitive
// bases.  They run once per robot / solver and do no per-problem arithmetic.
//
//   ilqrb_chain_from_urdf  replaces tup::TinyURDFParser (lib/tinyurdfparser/TinyURDFParser/TinyURDFParser.cpp:35-160,
//                          163-218,270-304) + KDL::Tree::getChain + the chain part of sim::KDLRobot's constructor
//                          (src/sim/KDLRobot.cpp:17-70).
//   ilqrb_build_psi        replaces src/utils/primitives.cpp:13-96.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/ilqrb.h"

namespace {

// ---- a minimal XML reader: elements, attributes, comments, declarations; text content is skipped -------------
struct XmlNode {
    std::string name;
    std::map<std::string, std::string> attr;
    std::vector<std::unique_ptr<XmlNode>> children;
    const XmlNode* child(const char* n) const {
        for (auto& c : children)
            if (c->name == n) return c.get();
        return nullptr;
    }
    const std::string* get(const char* k) const {
        auto it = attr.find(k);
        return it == attr.end() ? nullptr : &it->second;
    }
};

struct XmlParser {
    const std::string& s;
    size_t p = 0;
    explicit XmlParser(const std::string& text) : s(text) {}
    [[noreturn]] void bad() { throw std::runtime_error("Unable to open urdf file"); }  // TinyURDFParser.cpp:38
    void skipWs() {
        while (p < s.size() && std::isspace((unsigned char)s[p])) p++;
    }
    bool starts(const char* lit) const { return s.compare(p, std::strlen(lit), lit) == 0; }
    void skipMisc() {  // whitespace, comments, processing instructions, doctype, text
        for (;;) {
            while (p < s.size() && s[p] != '<') p++;
            if (p >= s.size()) return;
            if (starts("<!--")) {
                size_t e = s.find("-->", p + 4);
                if (e == std::string::npos) bad();
                p = e + 3;
            } else if (starts("<?")) {
                size_t e = s.find("?>", p + 2);
                if (e == std::string::npos) bad();
                p = e + 2;
            } else if (starts("<!")) {
                size_t e = s.find('>', p + 2);
                if (e == std::string::npos) bad();
                p = e + 1;
            } else {
                return;
            }
        }
    }
    std::string name() {
        size_t b = p;
        while (p < s.size() && (std::isalnum((unsigned char)s[p]) || s[p] == '_' || s[p] == '-' || s[p] == ':' || s[p] == '.')) p++;
        if (p == b) bad();
        return s.substr(b, p - b);
    }
    std::unique_ptr<XmlNode> element() {
        if (p >= s.size() || s[p] != '<') bad();
        p++;
        auto node = std::make_unique<XmlNode>();
        node->name = name();
        for (;;) {
            skipWs();
            if (p >= s.size()) bad();
            if (s[p] == '/') {
                if (p + 1 >= s.size() || s[p + 1] != '>') bad();
                p += 2;
                return node;
            }
            if (s[p] == '>') {
                p++;
                break;
            }
            std::string k = name();
            skipWs();
            if (p >= s.size() || s[p] != '=') bad();
            p++;
            skipWs();
            if (p >= s.size() || (s[p] != '"' && s[p] != '\'')) bad();
            char q = s[p++];
            size_t e = s.find(q, p);
            if (e == std::string::npos) bad();
            node->attr[k] = s.substr(p, e - p);
            p = e + 1;
        }
        for (;;) {
            skipMisc();
            if (p >= s.size()) bad();
            if (starts("</")) {
                p += 2;
                std::string n = name();
                if (n != node->name) bad();
                skipWs();
                if (p >= s.size() || s[p] != '>') bad();
                p++;
                return node;
            }
            node->children.push_back(element());
        }
    }
    std::unique_ptr<XmlNode> document() {
        skipMisc();
        if (p >= s.size()) throw std::runtime_error("[urdf_parsing] Given XML file does not contain a root element.");
        return element();
    }
};

struct V3 {
    double v[3] = {0, 0, 0};
};

// TinyURDFParser::rawArrayToEigenVector (TinyURDFParser.cpp:270-304): single-space separated, exactly 3 numbers
V3 parseVec3(std::string raw) {
    raw.erase(std::find_if(raw.rbegin(), raw.rend(), [](unsigned char ch) { return !std::isspace(ch); }).base(), raw.end());
    V3 out;
    size_t index = 0;
    for (;;) {
        size_t pos = raw.find(' ');
        if (index > 2) throw std::runtime_error("[urdf_parsing] Expected a 3d array got more, check arrays format.");
        if (pos > 0) {
            out.v[index] = std::stod(raw.substr(0, pos));
            index++;
        }
        if (pos == std::string::npos) break;
        raw.erase(0, pos + 1);
    }
    if (index != 3) throw std::runtime_error("[urdf_parsing] Expected a 3d array got less, check arrays format.");
    return out;
}

// KDL::Rotation::RPY (orocos_kdl frames.cpp), row-major
void rotRPY(double roll, double pitch, double yaw, double* R) {
    double ca1 = std::cos(yaw), sa1 = std::sin(yaw), cb1 = std::cos(pitch), sb1 = std::sin(pitch), cc1 = std::cos(roll), sc1 = std::sin(roll);
    R[0] = ca1 * cb1;
    R[1] = ca1 * sb1 * sc1 - sa1 * cc1;
    R[2] = ca1 * sb1 * cc1 + sa1 * sc1;
    R[3] = sa1 * cb1;
    R[4] = sa1 * sb1 * sc1 + ca1 * cc1;
    R[5] = sa1 * sb1 * cc1 - ca1 * sc1;
    R[6] = -sb1;
    R[7] = cb1 * sc1;
    R[8] = cb1 * cc1;
}

struct UJoint {
    std::string name, type, parent, child;
    V3 xyz, rpy, axis;
    UJoint() { axis.v[0] = 1; }  // TinyURDFParser.hpp: default axis (1,0,0)
};

int jointTypeOf(const std::string& t, bool& ok) {
    ok = true;
    if (t == "prismatic") return ILQRB_JOINT_TRANS;
    if (t == "revolute" || t == "continuous") return ILQRB_JOINT_ROT;
    if (t == "fixed") return ILQRB_JOINT_FIXED;
    ok = false;
    return ILQRB_JOINT_FIXED;
}

void chainFromUrdf(const std::string& text, const std::string& base, const std::string& tip, const double* rpy, const double* xyz, std::vector<ilqrb_segment>& out) {
    XmlParser parser(text);
    auto root = parser.document();
    if (root->name != "robot") throw std::runtime_error("[urdf_parsing] Root element should be named <robot> not <" + root->name + ">");
    std::map<std::string, UJoint> joints;  // std::map: name order, as in the reference (Q17)
    std::vector<std::string> links;
    for (auto& c : root->children) {
        if (c->name == "joint") {
            UJoint j;
            const std::string* nm = c->get("name");
            if (!nm) throw std::runtime_error("[urdf_parsing] Mandatory attribute \"name\" is missing for one of the joint");
            j.name = *nm;
            const std::string* ty = c->get("type");
            if (!ty) throw std::runtime_error("[urdf_parsing] Mandatory attribute \"type\" is missing for joint " + j.name);
            j.type = *ty;
            bool ok;
            jointTypeOf(j.type, ok);
            if (!ok) throw std::runtime_error("[urdf_parsing] Unknown joint type \"" + j.type + "\"");
            const XmlNode* pa = c->child("parent");
            if (!pa || !pa->get("link")) throw std::runtime_error("[urdf_parsing] Mandatory element <parent> is missing for joint " + j.name);
            j.parent = *pa->get("link");
            const XmlNode* ch = c->child("child");
            if (!ch || !ch->get("link")) throw std::runtime_error("[urdf_parsing] Mandatory element <child> is missing for joint " + j.name);
            j.child = *ch->get("link");
            if (const XmlNode* o = c->child("origin")) {
                if (o->get("rpy")) j.rpy = parseVec3(*o->get("rpy"));
                if (o->get("xyz")) j.xyz = parseVec3(*o->get("xyz"));
            }
            const XmlNode* ax = c->child("axis");
            if ((j.type == "revolute" || j.type == "continuous" || j.type == "prismatic" || j.type == "planar") && ax && ax->get("xyz")) j.axis = parseVec3(*ax->get("xyz"));
            joints[j.name] = j;
        } else if (c->name == "link") {
            if (c->get("name")) links.push_back(*c->get("name"));
        }
    }
    if (joints.empty() || links.empty()) throw std::runtime_error("[urdf_parsing] URDF file does not have <joint> or <link> tags");
    // links_[parent].parent = &joint : the last joint in name order wins (TinyURDFParser.cpp:91)
    std::map<std::string, const UJoint*> linkJoint, childJoint;
    for (auto& kv : joints) {
        linkJoint[kv.second.parent] = &kv.second;
        childJoint[kv.second.child] = &kv.second;
    }
    const std::string msg = "[KDLRobot] Unable to build kinematic chain from " + base + " to " + tip;  // KDLRobot.cpp:49,56
    std::vector<const UJoint*> path;
    std::string cur = tip;
    size_t guard = 0;
    while (cur != base) {
        auto it = childJoint.find(cur);
        if (it == childJoint.end() || ++guard > joints.size()) throw std::runtime_error(msg);
        path.push_back(it->second);
        cur = it->second->parent;
    }
    if (std::find(links.begin(), links.end(), base) == links.end()) throw std::runtime_error(msg);
    std::reverse(path.begin(), path.end());
    out.clear();
    for (const UJoint* j : path) {
        ilqrb_segment sg;
        std::memset(&sg, 0, sizeof(sg));
        double R[9];
        rotRPY(j->rpy.v[0], j->rpy.v[1], j->rpy.v[2], R);
        const UJoint* prev = linkJoint[j->parent];  // TinyURDFParser.cpp:135
        bool ok;
        const int jt = jointTypeOf(prev->type, ok);
        double origin[3] = {0, 0, 0};
        if (jt != ILQRB_JOINT_FIXED) {
            // KDL::Joint(name, origin = xyz, axis = RPY * axis, type): the axis is normalised by the Joint ctor
            double a[3];
            for (int i = 0; i < 3; i++) a[i] = R[3 * i] * prev->axis.v[0] + R[3 * i + 1] * prev->axis.v[1] + R[3 * i + 2] * prev->axis.v[2];
            const double nrm = std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
            for (int i = 0; i < 3; i++) {
                sg.axis[i] = a[i] / nrm;
                origin[i] = j->xyz.v[i];
                sg.origin[i] = origin[i];
            }
        }
        sg.joint_type = jt;
        // KDL::Segment: tip = joint.pose(0)^-1 * f_tip = (R_rpy, xyz - origin)
        for (int i = 0; i < 9; i++) sg.tip_R[i] = R[i];
        for (int i = 0; i < 3; i++) sg.tip_p[i] = j->xyz.v[i] - origin[i];
        out.push_back(sg);
    }
    // KDLRobot.cpp:60-66: fixed "robot_custom_tip" with Rotation::EulerZYX(rpy0, rpy1, rpy2) = RPY(rpy2, rpy1, rpy0)  (Q11)
    ilqrb_segment vt;
    std::memset(&vt, 0, sizeof(vt));
    vt.joint_type = ILQRB_JOINT_FIXED;
    const double z3[3] = {0, 0, 0};
    const double* r = rpy ? rpy : z3;
    const double* t = xyz ? xyz : z3;
    rotRPY(r[2], r[1], r[0], vt.tip_R);
    for (int i = 0; i < 3; i++) vt.tip_p[i] = t[i];
    out.push_back(vt);
}

int binomialCoefficients(int n, int k) {  // primitives.cpp:13-17
    if (k == 0 || k == n) return 1;
    return binomialCoefficients(n - 1, k - 1) + binomialCoefficients(n - 1, k);
}

}  // namespace

extern "C" {

void ilqrb_set_error_(const char* msg);  // defined in ilqrb_api.cu

int ilqrb_chain_from_urdf(const char* urdf, int is_path, const char* base, const char* tip, const double* transform_rpy, const double* transform_xyz,
                          ilqrb_segment* out, int cap, int* n_out, int* dof_out) {
    try {
        if (!urdf || !base || !tip || !out || !n_out) throw std::runtime_error("null argument");
        std::string text;
        if (is_path) {  // TinyURDFParser::fromFile (TinyURDFParser.cpp:23-33): an unreadable file parses as empty
            std::ifstream f(urdf);
            std::stringstream buffer;
            buffer << f.rdbuf();
            text = buffer.str();
        } else {
            text = urdf;
        }
        std::vector<ilqrb_segment> segs;
        chainFromUrdf(text, base, tip, transform_rpy, transform_xyz, segs);
        if ((int)segs.size() > cap) throw std::runtime_error("chain has more segments than the output capacity");
        int dof = 0;
        for (size_t i = 0; i < segs.size(); i++) {
            out[i] = segs[i];
            if (segs[i].joint_type != ILQRB_JOINT_FIXED) dof++;
        }
        *n_out = (int)segs.size();
        if (dof_out) *dof_out = dof;
        return 0;
    } catch (std::exception& e) {
        ilqrb_set_error_(e.what());
        return ILQRB_EINVAL;
    }
}

int ilqrb_build_psi(const char* kind, int dim, int K, double* out) {
    try {
        if (!kind || !out || dim < 1 || K < 1) throw std::runtime_error("bad arguments");
        const std::string k = kind;
        auto at = [&](int j, int i, int cols) -> double& { return out[(size_t)j * cols + i]; };
        if (k == "rbf") {  // primitives.cpp:19-33
            double bw = ((double)dim) / K, avg = bw / 2, sig = bw;
            for (int i = 0; i < K; i++) {
                for (int j = 0; j < dim; j++) at(j, i, K) = 1 / (2 * M_PI * sig) * std::exp(-1 * ((double)j - avg) * ((double)j - avg) / (2 * sig * sig));
                avg += bw;
            }
        } else if (k == "bernstein") {  // :35-50
            int order = K - 1;
            double mx = (double)(dim - 1);
            for (int i = 0; i < K; i++) {
                int binom = binomialCoefficients(order, i);
                for (int j = 0; j < dim; j++) {
                    double t = (double)j / mx;
                    at(j, i, K) = binom * std::pow(t, i) * std::pow(1 - t, order - i);
                }
            }
        } else if (k == "unitstep") {  // :52-69
            int bw = (int)std::round(((double)dim) / K), low = 0, high = bw;
            for (int i = 0; i < K; i++) {
                for (int j = 0; j < dim; j++) at(j, i, K) = (j >= low && j < high) ? 1.0 / bw : 0;
                low += bw;
                high += bw;
            }
        } else if (k == "sawtooth") {  // :71-88
            int bw = (int)std::ceil(((double)dim) / K);
            double low = 0, high = bw;
            for (int i = 0; i < K; i++) {
                for (int j = 0; j < dim; j++) at(j, i, K) = (j >= low && j < high) ? ((j - low) / (bw - 1) - 0.5) : 0;
                low += bw;
                high += bw;
            }
        } else if (k == "linear") {  // :90-96  [unitstep, sawtooth]
            std::vector<double> a((size_t)dim * K), b((size_t)dim * K);
            ilqrb_build_psi("unitstep", dim, K, a.data());
            ilqrb_build_psi("sawtooth", dim, K, b.data());
            for (int j = 0; j < dim; j++)
                for (int i = 0; i < K; i++) {
                    at(j, i, 2 * K) = a[(size_t)j * K + i];
                    at(j, K + i, 2 * K) = b[(size_t)j * K + i];
                }
        } else {
            throw std::runtime_error("unknown primitive kind: " + k);
        }
        return 0;
    } catch (std::exception& e) {
        ilqrb_set_error_(e.what());
        return ILQRB_EINVAL;
    }
}
}
