// sspp_model.cpp — MJCF-subset loader and model specialisation (host, FP64).  See sspp_model.h.
#include "sspp_model.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>

namespace sspp {

void quat_to_mat(const double* qi, double* R);

// ------------------------------------------------------------------------------------------
// minimal XML reader: elements, attributes, comments, declarations; character data ignored
// ------------------------------------------------------------------------------------------
namespace {

struct XmlNode {
  std::string tag;
  std::vector<std::pair<std::string, std::string>> attr;
  std::vector<std::unique_ptr<XmlNode>> child;
  const std::string* get(const std::string& k) const {
    for (auto& a : attr) if (a.first == k) return &a.second;
    return nullptr;
  }
};

struct XmlReader {
  const std::string& s;
  size_t i = 0;
  explicit XmlReader(const std::string& t) : s(t) {}
  [[noreturn]] void fail(const std::string& m) const {
    throw std::runtime_error("XML parse error at byte " + std::to_string(i) + ": " + m);
  }
  void skip_ws() { while (i < s.size() && std::isspace((unsigned char)s[i])) ++i; }
  bool starts(const char* p) const { return s.compare(i, std::strlen(p), p) == 0; }
  void skip_misc() {
    for (;;) {
      while (i < s.size() && s[i] != '<') ++i;   // character data
      if (i >= s.size()) return;
      if (starts("<!--")) { size_t e = s.find("-->", i + 4); if (e == std::string::npos) fail("unterminated comment"); i = e + 3; }
      else if (starts("<![CDATA[")) { size_t e = s.find("]]>", i + 9); if (e == std::string::npos) fail("unterminated CDATA"); i = e + 3; }
      else if (starts("<?")) { size_t e = s.find("?>", i + 2); if (e == std::string::npos) fail("unterminated declaration"); i = e + 2; }
      else if (starts("<!")) { size_t e = s.find('>', i); if (e == std::string::npos) fail("unterminated <!"); i = e + 1; }
      else return;
    }
  }
  std::string decode_entities(const std::string& v) const {
    if (v.find('&') == std::string::npos) return v;
    static const std::pair<const char*, char> ent[] = {{"&lt;", '<'}, {"&gt;", '>'}, {"&amp;", '&'}, {"&quot;", '"'}, {"&apos;", '\''}};
    std::string o;
    for (size_t k = 0; k < v.size();) {
      bool hit = false;
      if (v[k] == '&')
        for (auto& e : ent) { const size_t n = std::strlen(e.first); if (v.compare(k, n, e.first) == 0) { o += e.second; k += n; hit = true; break; } }
      if (!hit) o += v[k++];
    }
    return o;
  }
  std::string name() {
    size_t b = i;
    while (i < s.size() && (std::isalnum((unsigned char)s[i]) || s[i] == '_' || s[i] == '-' || s[i] == ':' || s[i] == '.')) ++i;
    if (b == i) fail("expected a name");
    return s.substr(b, i - b);
  }
  std::unique_ptr<XmlNode> element() {
    if (i >= s.size() || s[i] != '<') fail("expected '<'");
    ++i;
    auto n = std::make_unique<XmlNode>();
    n->tag = name();
    for (;;) {
      skip_ws();
      if (i >= s.size()) fail("unterminated tag");
      if (s[i] == '/') { if (!starts("/>")) fail("expected '/>'"); i += 2; return n; }
      if (s[i] == '>') { ++i; break; }
      std::string k = name();
      skip_ws();
      if (i >= s.size() || s[i] != '=') fail("expected '='");
      ++i; skip_ws();
      if (i >= s.size() || (s[i] != '"' && s[i] != '\'')) fail("expected a quoted value");
      char q = s[i++];
      size_t e = s.find(q, i);
      if (e == std::string::npos) fail("unterminated attribute value");
      n->attr.emplace_back(k, decode_entities(s.substr(i, e - i)));
      i = e + 1;
    }
    for (;;) {
      skip_misc();
      if (i >= s.size()) fail("missing </" + n->tag + ">");
      if (starts("</")) {
        i += 2;
        std::string c = name();
        if (c != n->tag) fail("mismatched </" + c + "> for <" + n->tag + ">");
        skip_ws();
        if (i >= s.size() || s[i] != '>') fail("expected '>'");
        ++i;
        return n;
      }
      n->child.push_back(element());
    }
  }
};

std::vector<double> nums(const std::string& v) {
  std::vector<double> out;
  std::istringstream is(v);
  double d;
  while (is >> d) out.push_back(d);
  return out;
}

using AttrMap = std::map<std::string, std::string>;
struct DefaultClass { AttrMap geom, joint; };

void mat_mul(const double* A, const double* B, double* C) {
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j)
    C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
void mat_vec(const double* A, const double* v, double* o) {
  for (int i = 0; i < 3; ++i) o[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
}
void mat_to_quat(const double* R, double* q) {
  double tr = R[0] + R[4] + R[8];
  if (tr > 0) { double s = std::sqrt(tr + 1.0) * 2; q[0] = 0.25 * s; q[1] = (R[7] - R[5]) / s; q[2] = (R[2] - R[6]) / s; q[3] = (R[3] - R[1]) / s; }
  else if (R[0] > R[4] && R[0] > R[8]) { double s = std::sqrt(1.0 + R[0] - R[4] - R[8]) * 2; q[0] = (R[7] - R[5]) / s; q[1] = 0.25 * s; q[2] = (R[1] + R[3]) / s; q[3] = (R[2] + R[6]) / s; }
  else if (R[4] > R[8]) { double s = std::sqrt(1.0 + R[4] - R[0] - R[8]) * 2; q[0] = (R[2] - R[6]) / s; q[1] = (R[1] + R[3]) / s; q[2] = 0.25 * s; q[3] = (R[5] + R[7]) / s; }
  else { double s = std::sqrt(1.0 + R[8] - R[0] - R[4]) * 2; q[0] = (R[3] - R[1]) / s; q[1] = (R[2] + R[6]) / s; q[2] = (R[5] + R[7]) / s; q[3] = 0.25 * s; }
}
void normalize3(double* v) {
  double n = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (n < 1e-14) throw std::runtime_error("zero-length axis in MJCF");
  v[0] /= n; v[1] /= n; v[2] /= n;
}
void zaxis_to_quat(const double* zin, double* q) {   // minimal rotation taking (0,0,1) to z
  double z[3] = {zin[0], zin[1], zin[2]};
  normalize3(z);
  double ax[3] = {-z[1], z[0], 0.0};                 // (0,0,1) x z
  double s = std::sqrt(ax[0] * ax[0] + ax[1] * ax[1]);
  double ang = std::atan2(s, z[2]);
  if (s < 1e-10) { ax[0] = 1; ax[1] = 0; ax[2] = 0; } else { ax[0] /= s; ax[1] /= s; }
  q[0] = std::cos(ang / 2); q[1] = ax[0] * std::sin(ang / 2); q[2] = ax[1] * std::sin(ang / 2); q[3] = ax[2] * std::sin(ang / 2);
}

struct Compiler { bool degree = true; std::string eulerseq = "xyz"; };

// rigid transform of a <frame> (or the identity); composed onto the bodies and geoms it wraps
struct Pose {
  double pos[3] = {0, 0, 0}, quat[4] = {1, 0, 0, 0};
  bool identity = true;
};
void quat_mul(const double* a, const double* b, double* o) {
  o[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  o[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  o[2] = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  o[3] = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
}
void quat_rotate(const double* q, const double* v, double* o) {
  double R[9];
  quat_to_mat(q, R);
  mat_vec(R, v, o);
}
// child pose (pos, quat) expressed in the frame's parent: outer o inner
void compose(const Pose& outer, double* pos, double* quat) {
  if (outer.identity) return;
  double rp[3], q[4];
  quat_rotate(outer.quat, pos, rp);
  for (int k = 0; k < 3; ++k) pos[k] = outer.pos[k] + rp[k];
  quat_mul(outer.quat, quat, q);
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int k = 0; k < 4; ++k) quat[k] = q[k] / n;
}

// orientation attributes of body / geom elements -> quaternion
void orientation(const AttrMap& a, const Compiler& c, double* q) {
  q[0] = 1; q[1] = q[2] = q[3] = 0;
  auto it = a.find("quat");
  if (it != a.end()) {
    auto v = nums(it->second);
    if (v.size() != 4) throw std::runtime_error("quat needs 4 numbers");
    double n = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3]);
    if (n < 1e-14) throw std::runtime_error("zero quaternion");
    for (int k = 0; k < 4; ++k) q[k] = v[k] / n;
    return;
  }
  const double ang = c.degree ? M_PI / 180.0 : 1.0;
  if ((it = a.find("euler")) != a.end()) {
    auto v = nums(it->second);
    if (v.size() != 3) throw std::runtime_error("euler needs 3 numbers");
    double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int k = 0; k < 3; ++k) {
      const char ch = c.eulerseq[k];
      const double t = v[k] * ang, cs = std::cos(t), sn = std::sin(t);
      double E[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
      const int ax = std::tolower(ch) - 'x';
      const int i1 = (ax + 1) % 3, i2 = (ax + 2) % 3;
      E[3 * i1 + i1] = cs; E[3 * i1 + i2] = -sn; E[3 * i2 + i1] = sn; E[3 * i2 + i2] = cs;
      double T[9];
      if (std::islower(ch)) mat_mul(R, E, T); else mat_mul(E, R, T);   // intrinsic / extrinsic
      std::memcpy(R, T, sizeof(R));
    }
    mat_to_quat(R, q);
    return;
  }
  if ((it = a.find("axisangle")) != a.end()) {
    auto v = nums(it->second);
    if (v.size() != 4) throw std::runtime_error("axisangle needs 4 numbers");
    double ax[3] = {v[0], v[1], v[2]};
    normalize3(ax);
    const double t = v[3] * ang;
    q[0] = std::cos(t / 2); for (int k = 0; k < 3; ++k) q[1 + k] = ax[k] * std::sin(t / 2);
    return;
  }
  if ((it = a.find("zaxis")) != a.end()) {
    auto v = nums(it->second);
    if (v.size() != 3) throw std::runtime_error("zaxis needs 3 numbers");
    zaxis_to_quat(v.data(), q);
    return;
  }
  if ((it = a.find("xyaxes")) != a.end()) {
    auto v = nums(it->second);
    if (v.size() != 6) throw std::runtime_error("xyaxes needs 6 numbers");
    double x[3] = {v[0], v[1], v[2]}, y[3] = {v[3], v[4], v[5]};
    normalize3(x);
    const double d = x[0] * y[0] + x[1] * y[1] + x[2] * y[2];
    for (int k = 0; k < 3; ++k) y[k] -= d * x[k];
    normalize3(y);
    const double z[3] = {x[1] * y[2] - x[2] * y[1], x[2] * y[0] - x[0] * y[2], x[0] * y[1] - x[1] * y[0]};
    const double R[9] = {x[0], y[0], z[0], x[1], y[1], z[1], x[2], y[2], z[2]};
    mat_to_quat(R, q);
  }
}

int geom_type_of(const std::string& t) {
  if (t == "plane") return ssppm::GT_PLANE;
  if (t == "sphere") return ssppm::GT_SPHERE;
  if (t == "capsule") return ssppm::GT_CAPSULE;
  if (t == "cylinder") return ssppm::GT_CYLINDER;
  if (t == "box") return ssppm::GT_BOX;
  if (t == "mesh") return 7;
  if (t == "hfield") return 1;
  if (t == "ellipsoid") return 4;
  throw std::runtime_error("unknown geom type '" + t + "'");
}

struct Builder {
  FlatModel m;
  Compiler comp;
  std::map<std::string, DefaultClass> defaults;

  void read_defaults(const XmlNode& n, const DefaultClass& inherited, bool top) {
    DefaultClass d = inherited;
    std::string cls = "main";
    if (const std::string* c = n.get("class")) cls = *c; else if (!top) throw std::runtime_error("nested <default> needs a class");
    for (auto& ch : n.child) {
      if (ch->tag == "geom") for (auto& a : ch->attr) d.geom[a.first] = a.second;
      else if (ch->tag == "joint") for (auto& a : ch->attr) d.joint[a.first] = a.second;
    }
    defaults[cls] = d;
    for (auto& ch : n.child) if (ch->tag == "default") read_defaults(*ch, d, false);
  }

  AttrMap resolve(const XmlNode& e, const std::string& childclass, bool is_geom) const {
    std::string cls = "main";
    if (const std::string* c = e.get("class")) cls = *c; else if (!childclass.empty()) cls = childclass;
    AttrMap a;
    auto it = defaults.find(cls);
    if (it != defaults.end()) a = is_geom ? it->second.geom : it->second.joint;
    else if (cls != "main") throw std::runtime_error("unknown default class '" + cls + "'");
    for (auto& kv : e.attr) a[kv.first] = kv.second;
    return a;
  }

  void add_geom(const XmlNode& e, int body, const std::string& childclass, const Pose& frame) {
    AttrMap a = resolve(e, childclass, true);
    auto geti = [&](const char* k, int dflt) { auto it = a.find(k); return it == a.end() ? dflt : (int)std::lround(nums(it->second).at(0)); };
    const int contype = geti("contype", 1), conaff = geti("conaffinity", 1);
    std::string tname = a.count("type") ? a["type"] : std::string(a.count("mesh") ? "mesh" : "sphere");
    const int type = geom_type_of(tname);
    const std::string gname = a.count("name") ? a["name"] : std::string();
    if (type == 7 || type == 1 || type == 4) {
      if (contype == 0 && conaff == 0) return;   // visual only: not on the scoring path
      throw std::runtime_error("geom '" + gname + "': collidable " + tname + " geoms are not supported");
    }
    double size[3] = {0, 0, 0}, pos[3] = {0, 0, 0}, quat[4];
    if (a.count("size")) { auto v = nums(a["size"]); for (size_t k = 0; k < v.size() && k < 3; ++k) size[k] = v[k]; }
    if (a.count("pos")) { auto v = nums(a["pos"]); if (v.size() != 3) throw std::runtime_error("geom pos needs 3 numbers"); for (int k = 0; k < 3; ++k) pos[k] = v[k]; }
    orientation(a, comp, quat);
    if (a.count("fromto")) {
      auto v = nums(a["fromto"]);
      if (v.size() != 6) throw std::runtime_error("fromto needs 6 numbers");
      if (type != ssppm::GT_CAPSULE && type != ssppm::GT_CYLINDER && type != ssppm::GT_BOX) throw std::runtime_error("fromto on unsupported geom type");
      double d[3] = {v[0] - v[3], v[1] - v[4], v[2] - v[5]};
      const double len = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
      for (int k = 0; k < 3; ++k) pos[k] = 0.5 * (v[k] + v[3 + k]);
      zaxis_to_quat(d, quat);
      if (type == ssppm::GT_BOX) size[2] = 0.5 * len; else size[1] = 0.5 * len;
    }
    compose(frame, pos, quat);
    const bool ok = (type == ssppm::GT_PLANE) || (type == ssppm::GT_SPHERE && size[0] > 0) ||
                    ((type == ssppm::GT_CAPSULE || type == ssppm::GT_CYLINDER) && size[0] > 0 && size[1] > 0) ||
                    (type == ssppm::GT_BOX && size[0] > 0 && size[1] > 0 && size[2] > 0);
    if (!ok) throw std::runtime_error("geom '" + gname + "': size is incomplete for type " + tname);
    if (type == ssppm::GT_PLANE && body != 0 && m.body_weld[body] != 0)
      throw std::runtime_error("geom '" + gname + "': planes must belong to a static body");
    const double margin = a.count("margin") ? nums(a["margin"]).at(0) : 0.0;
    if (!(margin >= 0.0)) throw std::runtime_error("geom '" + gname + "': margin must be non-negative");
    m.geom_name.push_back(gname);
    m.geom_body.push_back(body);
    m.geom_type.push_back(type);
    m.geom_contype.push_back(contype);
    m.geom_conaffinity.push_back(conaff);
    m.geom_margin.push_back(margin);
    for (int k = 0; k < 3; ++k) { m.geom_size.push_back(size[k]); m.geom_pos.push_back(pos[k]); }
    for (int k = 0; k < 4; ++k) m.geom_quat.push_back(quat[k]);
    ++m.ngeom;
  }

  // children of a body (or of the world), <frame> wrappers flattened in document order
  struct Item { const XmlNode* node; Pose frame; std::string childclass; };
  void collect(const XmlNode& e, const Pose& frame, const std::string& childclass, std::vector<Item>* out) const {
    for (auto& ch : e.child) {
      if (ch->tag == "frame") {
        AttrMap fa;
        for (auto& kv : ch->attr) fa[kv.first] = kv.second;
        Pose inner;
        if (fa.count("pos")) { auto v = nums(fa["pos"]); if (v.size() != 3) throw std::runtime_error("frame pos needs 3 numbers"); for (int k = 0; k < 3; ++k) inner.pos[k] = v[k]; }
        orientation(fa, comp, inner.quat);
        compose(frame, inner.pos, inner.quat);
        inner.identity = false;
        collect(*ch, inner, fa.count("childclass") ? fa["childclass"] : childclass, out);
      } else if (ch->tag == "geom" || ch->tag == "body" || ch->tag == "joint" || ch->tag == "freejoint") {
        out->push_back(Item{ch.get(), frame, childclass});
      }
    }
  }

  int push_body(const std::string& name, int parent, const double* pos, const double* quat) {
    const int id = m.nbody++;
    m.body_name.push_back(name);
    m.body_parent.push_back(parent < 0 ? 0 : parent);
    for (int k = 0; k < 3; ++k) m.body_pos.push_back(pos[k]);
    for (int k = 0; k < 4; ++k) m.body_quat.push_back(quat[k]);
    m.body_jtype.push_back(MJ_NONE);
    m.body_qadr.push_back(-1);
    for (int k = 0; k < 3; ++k) { m.jnt_axis.push_back(k == 2 ? 1.0 : 0.0); m.jnt_pos.push_back(0.0); }
    m.jnt_range.push_back(0.0); m.jnt_range.push_back(0.0);
    m.body_weld.push_back(0);
    return id;
  }

  void set_joint(int id, int parent, const Item& it, const std::string& owner) {
    const XmlNode& j = *it.node;
    if (j.tag == "freejoint") { m.body_jtype[id] = MJ_FREE; }
    else {
      AttrMap ja = resolve(j, it.childclass, false);
      const std::string t = ja.count("type") ? ja["type"] : "hinge";
      if (t == "hinge") m.body_jtype[id] = MJ_HINGE;
      else if (t == "slide") m.body_jtype[id] = MJ_SLIDE;
      else if (t == "free") m.body_jtype[id] = MJ_FREE;
      else throw std::runtime_error("body '" + owner + "': joint type '" + t + "' is not supported");
      double axis[3] = {0, 0, 1}, jpos[3] = {0, 0, 0};
      if (ja.count("axis")) { auto v = nums(ja["axis"]); if (v.size() != 3) throw std::runtime_error("joint axis needs 3 numbers"); for (int k = 0; k < 3; ++k) axis[k] = v[k]; }
      if (ja.count("pos")) { auto v = nums(ja["pos"]); if (v.size() != 3) throw std::runtime_error("joint pos needs 3 numbers"); for (int k = 0; k < 3; ++k) jpos[k] = v[k]; }
      normalize3(axis);
      if (!it.frame.identity) {   // joint declared inside a <frame>: anchor and axis are given in that frame
        double t3[3];
        quat_rotate(it.frame.quat, axis, t3); for (int k = 0; k < 3; ++k) axis[k] = t3[k];
        quat_rotate(it.frame.quat, jpos, t3); for (int k = 0; k < 3; ++k) jpos[k] = it.frame.pos[k] + t3[k];
      }
      for (int k = 0; k < 3; ++k) { m.jnt_axis[3 * id + k] = axis[k]; m.jnt_pos[3 * id + k] = jpos[k]; }
      if (ja.count("range")) { auto v = nums(ja["range"]); if (v.size() == 2) { const double sc = (comp.degree && m.body_jtype[id] == MJ_HINGE) ? M_PI / 180.0 : 1.0; m.jnt_range[2 * id] = v[0] * sc; m.jnt_range[2 * id + 1] = v[1] * sc; } }
      // qpos is used as the joint displacement itself; a reference offset would shift every planned coordinate
      if (ja.count("ref") && m.body_jtype[id] != MJ_FREE && nums(ja["ref"]).at(0) != 0.0)
        throw std::runtime_error("body '" + owner + "': joint ref other than 0 is not supported");
    }
    m.body_qadr[id] = m.nq;
    if (m.body_jtype[id] == MJ_FREE) {
      if (parent != 0) throw std::runtime_error("free joint below a non-world body");
      for (int k = 0; k < 3; ++k) m.qpos0.push_back(m.body_pos[3 * id + k]);
      for (int k = 0; k < 4; ++k) m.qpos0.push_back(m.body_quat[4 * id + k]);
      m.nq += 7;
    } else { m.qpos0.push_back(0.0); m.nq += 1; }
  }

  // One MJCF body.  A body with several hinge / slide joints becomes a chain of bodies with one
  // joint each (MuJoCo applies a body's joints in order, each in the frame the earlier ones left,
  // which is what the chain does); the links of the chain share one weld id, so the collision
  // filters (same body, parent-child) see them as the single body they stand for.
  int add_body(const XmlNode* e, int parent, std::string childclass, const Pose& frame) {
    AttrMap a;
    if (e) for (auto& kv : e->attr) a[kv.first] = kv.second;
    if (a.count("childclass")) childclass = a["childclass"];
    double pos[3] = {0, 0, 0}, quat[4] = {1, 0, 0, 0};
    if (a.count("pos")) { auto v = nums(a["pos"]); if (v.size() != 3) throw std::runtime_error("body pos needs 3 numbers"); for (int k = 0; k < 3; ++k) pos[k] = v[k]; }
    orientation(a, comp, quat);
    compose(frame, pos, quat);
    const std::string name = a.count("name") ? a["name"] : (m.nbody == 0 ? "world" : "");
    if (!name.empty() && m.body_id(name) >= 0) throw std::runtime_error("repeated body name '" + name + "'");
    std::vector<Item> items;
    if (e) collect(*e, Pose(), childclass, &items);
    std::vector<const Item*> joints;
    for (auto& it : items) if (it.node->tag == "joint" || it.node->tag == "freejoint") joints.push_back(&it);
    if (!joints.empty() && m.nbody == 0) throw std::runtime_error("joint in the world body");
    const double zero[3] = {0, 0, 0}, ident[4] = {1, 0, 0, 0};
    int id = -1, top = -1;
    const size_t nlinks = joints.empty() ? 1 : joints.size();
    for (size_t l = 0; l < nlinks; ++l) {
      const bool last = l + 1 == nlinks;
      id = push_body(last ? name : std::string(), l == 0 ? parent : id, l == 0 ? pos : zero, l == 0 ? quat : ident);
      if (l == 0) top = id;
      if (!joints.empty()) {
        set_joint(id, parent, *joints[l], name);
        if (nlinks > 1 && m.body_jtype[id] == MJ_FREE) throw std::runtime_error("body '" + name + "': a free joint cannot be combined with other joints");
      }
      m.body_weld[id] = (m.body_jtype[id] != MJ_NONE) ? top : m.body_weld[m.body_parent[id]];
    }
    for (auto& it : items) if (it.node->tag == "geom") add_geom(*it.node, id, it.childclass, it.frame);
    for (auto& it : items) if (it.node->tag == "body") add_body(it.node, id, it.childclass, it.frame);
    return id;
  }
};

// <include file="..."/> anywhere in the tree: the children of the included root take its place
void expand_includes(XmlNode* n, const std::string& base_dir, int depth) {
  if (depth > 16) throw std::runtime_error("<include> nesting deeper than 16 (a cycle?)");
  std::vector<std::unique_ptr<XmlNode>> out;
  for (auto& ch : n->child) {
    if (ch->tag != "include") { expand_includes(ch.get(), base_dir, depth); out.push_back(std::move(ch)); continue; }
    const std::string* f = ch->get("file");
    if (!f) throw std::runtime_error("<include> needs a file attribute");
    if (base_dir.empty() && (f->empty() || (*f)[0] != '/')) throw std::runtime_error("<include file='" + *f + "'> in a model given as text: no directory to resolve it against");
    const std::string path = (!f->empty() && (*f)[0] == '/') ? *f : base_dir + "/" + *f;
    std::ifstream in(path);
    if (!in) throw std::runtime_error("cannot open included file '" + path + "'");
    std::stringstream ss; ss << in.rdbuf();
    const std::string text = ss.str();
    XmlReader rd(text);
    rd.skip_misc();
    if (rd.i >= text.size()) throw std::runtime_error("included file '" + path + "' is empty");
    std::unique_ptr<XmlNode> root = rd.element();
    const size_t slash = path.find_last_of('/');
    expand_includes(root.get(), slash == std::string::npos ? std::string(".") : path.substr(0, slash), depth + 1);
    for (auto& g : root->child) out.push_back(std::move(g));
  }
  n->child = std::move(out);
}

}  // namespace

int FlatModel::body_id(const std::string& name) const {
  for (int i = 0; i < nbody; ++i) if (body_name[i] == name) return i;
  return -1;
}

void quat_to_mat(const double* qi, double* R) {
  double n = std::sqrt(qi[0] * qi[0] + qi[1] * qi[1] + qi[2] * qi[2] + qi[3] * qi[3]);
  const double w = qi[0] / n, x = qi[1] / n, y = qi[2] / n, z = qi[3] / n;
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - w * z);     R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z);     R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y);     R[7] = 2 * (y * z + w * x);     R[8] = 1 - 2 * (x * x + y * y);
}

FlatModel parse_mjcf(const std::string& text, const std::string& base_dir) {
  XmlReader rd(text);
  rd.skip_misc();
  if (rd.i >= text.size()) throw std::runtime_error("empty MJCF document");
  std::unique_ptr<XmlNode> root = rd.element();
  if (root->tag != "mujoco") throw std::runtime_error("root element is <" + root->tag + ">, expected <mujoco>");
  expand_includes(root.get(), base_dir, 0);
  Builder b;
  for (auto& ch : root->child) if (ch->tag == "compiler") {
    if (const std::string* a = ch->get("angle")) {
      if (*a != "radian" && *a != "degree") throw std::runtime_error("compiler angle must be 'degree' or 'radian'");
      b.comp.degree = (*a != "radian");
    }
    if (const std::string* a = ch->get("eulerseq")) {
      if (a->size() != 3) throw std::runtime_error("eulerseq needs 3 letters");
      for (char c : *a) if (std::tolower(c) < 'x' || std::tolower(c) > 'z') throw std::runtime_error("eulerseq letters must be x, y or z");
      b.comp.eulerseq = *a;
    }
    if (const std::string* a = ch->get("coordinate")) if (*a != "local") throw std::runtime_error("compiler coordinate='" + *a + "' is not supported");
  }
  for (auto& ch : root->child) if (ch->tag == "default") b.read_defaults(*ch, DefaultClass(), true);
  std::vector<const XmlNode*> worlds;          // several <worldbody> sections are merged, as MuJoCo does
  for (auto& ch : root->child) if (ch->tag == "worldbody") worlds.push_back(ch.get());
  if (worlds.empty()) throw std::runtime_error("no <worldbody>");
  b.add_body(nullptr, -1, "", Pose());
  std::vector<Builder::Item> items;
  for (const XmlNode* w : worlds) b.collect(*w, Pose(), "", &items);
  for (auto& it : items) if (it.node->tag == "joint" || it.node->tag == "freejoint") throw std::runtime_error("joint in the world body");
  for (auto& it : items) if (it.node->tag == "geom") b.add_geom(*it.node, 0, it.childclass, it.frame);
  for (auto& it : items) if (it.node->tag == "body") b.add_body(it.node, 0, it.childclass, it.frame);
  for (auto& ch : root->child) if (ch->tag == "contact")
    for (auto& ex : ch->child) {
      if (ex->tag == "pair") throw std::runtime_error("<contact><pair> (explicit geom pairs) is not supported");
      if (ex->tag != "exclude") continue;
      const std::string* b1 = ex->get("body1"); const std::string* b2 = ex->get("body2");
      if (!b1 || !b2) throw std::runtime_error("<exclude> needs body1 and body2");
      const int i1 = b.m.body_id(*b1), i2 = b.m.body_id(*b2);
      if (i1 < 0 || i2 < 0) throw std::runtime_error("<exclude> names an unknown body");
      b.m.exclude.push_back(i1); b.m.exclude.push_back(i2);
    }
  return b.m;
}

FlatModel load_mjcf(const std::string& path) {
  std::ifstream f(path);
  if (!f) throw std::runtime_error("Failed to load MuJoCo model from XML: cannot open '" + path + "'");
  std::stringstream ss; ss << f.rdbuf();
  const size_t slash = path.find_last_of('/');
  try { return parse_mjcf(ss.str(), slash == std::string::npos ? std::string(".") : path.substr(0, slash)); }
  catch (const std::exception& e) { throw std::runtime_error(std::string("Failed to load MuJoCo model from XML: ") + e.what()); }
}

bool pair_active(const FlatModel& m, int g1, int g2) {
  const int b1 = m.geom_body[g1], b2 = m.geom_body[g2];
  const int w1 = m.body_weld[b1], w2 = m.body_weld[b2];
  if (w1 == w2) return false;
  if (w1 != 0 && w2 != 0 && (m.body_weld[m.body_parent[w1]] == w2 || m.body_weld[m.body_parent[w2]] == w1)) return false;
  for (size_t k = 0; k + 1 < m.exclude.size(); k += 2)
    if ((m.exclude[k] == b1 && m.exclude[k + 1] == b2) || (m.exclude[k] == b2 && m.exclude[k + 1] == b1)) return false;
  if (m.geom_type[g1] == ssppm::GT_PLANE && m.geom_type[g2] == ssppm::GT_PLANE) return false;
  return (m.geom_contype[g1] & m.geom_conaffinity[g2]) || (m.geom_contype[g2] & m.geom_conaffinity[g1]);
}

namespace {

bool pair_supported(int ta, int tb) {
  if (ta > tb) std::swap(ta, tb);
  using namespace ssppm;
  if (ta == GT_PLANE) return tb != GT_PLANE;
  if (ta == GT_SPHERE) return true;
  if (ta == GT_CAPSULE) return tb == GT_CAPSULE || tb == GT_BOX;
  if (ta == GT_CYLINDER) return tb == GT_BOX;
  return ta == GT_BOX && tb == GT_BOX;
}

double bound_radius(int type, const double* s) {
  using namespace ssppm;
  switch (type) {
    case GT_SPHERE: return s[0];
    case GT_CAPSULE: return s[0] + s[1];
    case GT_CYLINDER: return std::sqrt(s[0] * s[0] + s[1] * s[1]);
    case GT_BOX: return std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
    default: return -1.0;
  }
}

struct Xf { double R[9]; double p[3]; };

}  // namespace

// Spheres inscribed in a primitive.  Boxes: radius rho = ratio * smallest half extent, centres on a
// lattice with spacing <= rho that reaches h - rho on every axis.  Capsules: radius r along the axis,
// spacing <= ratio * r.  Cylinders: rho = ratio * min(r, half length) on axial layers, a centre line plus
// rings out to r - rho.  Spheres: themselves.
void sample_spheres(int type, const double* size, double ratio, std::vector<double>* out) {
  using namespace ssppm;
  auto lattice = [](double h, double rho) {     // coordinates covering [-(h-rho), h-rho] with spacing <= rho
    std::vector<double> v;
    const double span = std::max(0.0, h - rho);
    const int n = (span <= 1e-12) ? 1 : 1 + (int)std::ceil(2.0 * span / rho - 1e-9);
    for (int i = 0; i < n; ++i) v.push_back(n == 1 ? 0.0 : -span + 2.0 * span * i / (n - 1));
    return v;
  };
  auto push = [&](double x, double y, double z, double r) { out->push_back(x); out->push_back(y); out->push_back(z); out->push_back(r); };
  ratio = std::min(1.0, std::max(0.05, ratio));
  if (type == GT_SPHERE) { push(0, 0, 0, size[0]); return; }
  if (type == GT_CAPSULE) {
    const double r = size[0], h = size[1];
    const int n = 1 + (int)std::ceil(2.0 * h / (ratio * r) - 1e-9);
    for (int i = 0; i < n; ++i) push(0, 0, n == 1 ? 0.0 : -h + 2.0 * h * i / (n - 1), r);
    return;
  }
  if (type == GT_CYLINDER) {
    const double r = size[0], h = size[1], rho = ratio * std::min(r, h);
    for (double z : lattice(h, rho)) {
      push(0, 0, z, rho);
      for (double rr = rho; rr <= r - rho + 1e-12; rr += rho) {
        const int n = std::max(3, (int)std::ceil(2.0 * M_PI * rr / rho));
        for (int i = 0; i < n; ++i) push(rr * std::cos(2.0 * M_PI * i / n), rr * std::sin(2.0 * M_PI * i / n), z, rho);
      }
    }
    return;
  }
  if (type == GT_BOX) {
    const double rho = ratio * std::min(size[0], std::min(size[1], size[2]));
    for (double x : lattice(size[0], rho)) for (double y : lattice(size[1], rho)) for (double z : lattice(size[2], rho)) push(x, y, z, rho);
    return;
  }
  throw std::runtime_error("SDF mode cannot sample this geom type with spheres");
}

ssppm::DevModel compile_model(const FlatModel& m, PlanMode mode, int dof, const std::string& body, CompileInfo* info,
                              const CompileOptions& opt) {
  using namespace ssppm;
  // --- which joints are planned ---------------------------------------------------------
  std::vector<int> planned(m.nbody, 0), qidx(m.nbody, -1);
  if (mode == PLAN_JOINT) {
    if (dof < 1 || dof > MAX_DOF) throw std::runtime_error("dof out of range");
    int covered = 0;
    for (int b = 1; b < m.nbody; ++b) {
      if (m.body_jtype[b] == MJ_NONE || m.body_qadr[b] >= dof) continue;
      if (m.body_jtype[b] != MJ_HINGE && m.body_jtype[b] != MJ_SLIDE)
        throw std::runtime_error("planned coordinate " + std::to_string(m.body_qadr[b]) + " belongs to a joint that is neither hinge nor slide");
      planned[b] = 1; qidx[b] = m.body_qadr[b]; ++covered;
    }
    if (covered != dof) throw std::runtime_error("model has fewer than " + std::to_string(dof) + " scalar joint coordinates at the start of qpos");
  } else {
    const int b = m.body_id(body);
    if (b < 0) throw std::runtime_error("Body with name '" + body + "' not found.");
    if (m.body_jtype[b] != MJ_FREE) throw std::runtime_error("Joint of body '" + body + "' is not a free joint.");
    planned[b] = 1; qidx[b] = 0;
  }
  // --- world frames at qpos0 and rigid clusters -------------------------------------------
  std::vector<Xf> xw(m.nbody);
  std::vector<int> cluster(m.nbody, -1);   // nearest planned ancestor-or-self, -1: static
  std::vector<Xf> xc(m.nbody);             // frame relative to the cluster root frame
  for (int b = 0; b < m.nbody; ++b) {
    Xf loc;
    quat_to_mat(&m.body_quat[4 * b], loc.R);
    for (int k = 0; k < 3; ++k) loc.p[k] = m.body_pos[3 * b + k];
    if (b == 0) { xw[0] = loc; continue; }
    const int p = m.body_parent[b];
    if (m.body_jtype[b] == MJ_FREE) {      // qpos0 holds the XML pose: world frame = local frame
      xw[b] = loc;
    } else {                               // hinge / slide at qpos0 = 0: identity joint transform
      mat_mul(xw[p].R, loc.R, xw[b].R);
      mat_vec(xw[p].R, loc.p, xw[b].p);
      for (int k = 0; k < 3; ++k) xw[b].p[k] += xw[p].p[k];
    }
    if (planned[b]) { cluster[b] = b; std::memset(&xc[b], 0, sizeof(Xf)); xc[b].R[0] = xc[b].R[4] = xc[b].R[8] = 1; }
    else if (cluster[p] >= 0) {
      cluster[b] = cluster[p];
      mat_mul(xc[p].R, loc.R, xc[b].R);
      mat_vec(xc[p].R, loc.p, xc[b].p);
      for (int k = 0; k < 3; ++k) xc[b].p[k] += xc[p].p[k];
    }
  }
  DevModel D;
  std::memset(&D, 0, sizeof(D));
  D.rule = (mode == PLAN_JOINT) ? CR_COUNT : ((opt.use_center_dist && !opt.sdf) ? CR_CENTER : CR_DEPTH);
  // --- moving bodies = cluster roots, in id order (parents precede children) ---------------
  std::vector<int> mb_of(m.nbody, -1);
  for (int b = 1; b < m.nbody; ++b) {
    if (!planned[b]) continue;
    if (D.n_mbody >= MAX_MBODY) throw std::runtime_error("too many moving bodies");
    DevBody& B = D.body[D.n_mbody];
    mb_of[b] = D.n_mbody++;
    const int p = m.body_parent[b];
    Xf loc;
    quat_to_mat(&m.body_quat[4 * b], loc.R);
    for (int k = 0; k < 3; ++k) loc.p[k] = m.body_pos[3 * b + k];
    Xf rel;   // frame of b relative to the parent cluster root (or the world when the parent is static)
    const Xf& base = (cluster[p] >= 0) ? xc[p] : xw[p];
    mat_mul(base.R, loc.R, rel.R);
    mat_vec(base.R, loc.p, rel.p);
    for (int k = 0; k < 3; ++k) rel.p[k] += base.p[k];
    B.parent = (cluster[p] >= 0) ? mb_of[cluster[p]] : -1;
    for (int k = 0; k < 3; ++k) { B.pos[k] = (float)rel.p[k]; B.axis[k] = (float)m.jnt_axis[3 * b + k]; B.anchor[k] = (float)m.jnt_pos[3 * b + k]; }
    for (int k = 0; k < 9; ++k) B.R0[k] = (float)rel.R[k];
    B.has_anchor = (m.jnt_pos[3 * b] != 0 || m.jnt_pos[3 * b + 1] != 0 || m.jnt_pos[3 * b + 2] != 0);
    B.axis_kind = 3; B.axis_sign = 1.0f;
    for (int k = 0; k < 3; ++k)
      if (std::fabs(m.jnt_axis[3 * b + k]) == 1.0 && m.jnt_axis[3 * b + (k + 1) % 3] == 0.0 && m.jnt_axis[3 * b + (k + 2) % 3] == 0.0) {
        B.axis_kind = k; B.axis_sign = (float)m.jnt_axis[3 * b + k];
      }
    B.r0_identity = 1;
    for (int k = 0; k < 9; ++k) if (B.R0[k] != ((k % 4 == 0) ? 1.0f : 0.0f)) B.r0_identity = 0;
    B.jtype = (m.body_jtype[b] == MJ_HINGE) ? JT_HINGE : (m.body_jtype[b] == MJ_SLIDE) ? JT_SLIDE : JT_FREE4;
    B.qidx = qidx[b];
  }
  // --- geoms -------------------------------------------------------------------------------
  std::vector<int> mg_of(m.ngeom, -1), sg_of(m.ngeom, -1), mgeom_src, sgeom_src;
  for (int mb = 0; mb < D.n_mbody; ++mb) {
    D.body[mb].geom_begin = D.n_mgeom;
    for (int g = 0; g < m.ngeom; ++g) {
      const int gb = m.geom_body[g];
      if (cluster[gb] < 0 || mb_of[cluster[gb]] != mb) continue;
      bool any = false;   // keep only geoms that can touch something
      for (int h = 0; h < m.ngeom && !any; ++h) if (h != g && cluster[m.geom_body[h]] != cluster[gb] && pair_active(m, std::min(g, h), std::max(g, h))) any = true;
      if (!any) continue;
      if (D.n_mgeom >= MAX_MGEOM) throw std::runtime_error("too many moving geoms");
      DevMGeom& G = D.mg[D.n_mgeom];
      mg_of[g] = D.n_mgeom++;
      mgeom_src.push_back(g);
      double Rg[9], Rl[9], pl[3];
      quat_to_mat(&m.geom_quat[4 * g], Rg);
      mat_mul(xc[gb].R, Rg, Rl);
      mat_vec(xc[gb].R, &m.geom_pos[3 * g], pl);
      for (int k = 0; k < 3; ++k) { G.pos[k] = (float)(pl[k] + xc[gb].p[k]); G.size[k] = (float)m.geom_size[3 * g + k]; }
      for (int k = 0; k < 9; ++k) G.R[k] = (float)Rl[k];
      G.type = m.geom_type[g];
      G.rbound = (float)(bound_radius(G.type, &m.geom_size[3 * g]) * (1.0 + 1e-6));
      G.body = mb;
      if (G.type == GT_PLANE) throw std::runtime_error("a plane geom on a moving body is not supported");
    }
    D.body[mb].geom_end = D.n_mgeom;
  }
  auto static_geom = [&](int g) {
    if (sg_of[g] >= 0) return sg_of[g];
    if (D.n_sgeom >= MAX_SGEOM) throw std::runtime_error("too many static geoms");
    DevSGeom& S = D.sg[D.n_sgeom];
    const int gb = m.geom_body[g];
    double Rg[9], Rw[9], pw[3];
    quat_to_mat(&m.geom_quat[4 * g], Rg);
    mat_mul(xw[gb].R, Rg, Rw);
    mat_vec(xw[gb].R, &m.geom_pos[3 * g], pw);
    for (int k = 0; k < 3; ++k) { S.pos[k] = (float)(pw[k] + xw[gb].p[k]); S.size[k] = (float)m.geom_size[3 * g + k]; }
    for (int k = 0; k < 9; ++k) S.R[k] = (float)Rw[k];
    S.type = m.geom_type[g];
    const double rb = bound_radius(S.type, &m.geom_size[3 * g]);
    S.rbound = (float)(rb < 0 ? -1.0 : rb * (1.0 + 1e-6));
    sgeom_src.push_back(g);
    return sg_of[g] = D.n_sgeom++;
  };
  // Workspace bound of every moving body: its frame origin stays within reach[b] of the world point centre[b]
  // whatever the planned coordinates are (hinges turn about their own anchor; slides and free bodies are unbounded).  A pair whose static partner lies beyond that bound can never generate a
  // contact and is dropped here, once, instead of being culled per collocation point.
  const double kInf = 1e30;
  std::vector<double> reach(D.n_mbody, kInf);
  std::vector<std::array<double, 3>> centre(D.n_mbody, std::array<double, 3>{0, 0, 0});
  for (int mb = 0; mb < D.n_mbody; ++mb) {
    const DevBody& B = D.body[mb];
    const double plen = std::sqrt((double)B.pos[0] * B.pos[0] + (double)B.pos[1] * B.pos[1] + (double)B.pos[2] * B.pos[2]);
    const double alen = std::sqrt((double)B.anchor[0] * B.anchor[0] + (double)B.anchor[1] * B.anchor[1] + (double)B.anchor[2] * B.anchor[2]);
    double own = kInf;
    if (B.jtype == JT_HINGE) own = 2.0 * alen;
    // a slide joint is UNBOUNDED here even when the MJCF gives it a range: nothing on the path clamps the planned coordinate
    // to jnt_range (the noise is unbounded, the spline may overshoot, limit_penalty is a soft cost) and the reference writes
    // qpos as it is (sspp.h:139-142) - so a partner beyond the range is still reachable and must stay in the pair table
    if (B.parent < 0) { centre[mb] = {B.pos[0], B.pos[1], B.pos[2]}; reach[mb] = own; }
    else { centre[mb] = centre[B.parent]; reach[mb] = std::min(kInf, reach[B.parent] + plen + own); }
  }
  auto unreachable = [&](int ig, int h) {     // moving geom ig can never come within the margin of model geom h
    if (opt.sdf) return false;                // the grid needs the same static set for every moving geom
    const DevMGeom& G = D.mg[ig];
    const double r = reach[G.body];
    if (r >= kInf) return false;
    const double glen = std::sqrt((double)G.pos[0] * G.pos[0] + (double)G.pos[1] * G.pos[1] + (double)G.pos[2] * G.pos[2]);
    const int hb = m.geom_body[h];
    double Rg[9], Rw[9], pw[3];
    quat_to_mat(&m.geom_quat[4 * h], Rg);
    mat_mul(xw[hb].R, Rg, Rw);
    mat_vec(xw[hb].R, &m.geom_pos[3 * h], pw);
    for (int k = 0; k < 3; ++k) pw[k] += xw[hb].p[k];
    const double margin = std::max(0.0, std::max(m.geom_margin[mgeom_src[ig]], m.geom_margin[h]));
    const double need = r + glen + G.rbound + margin + 2e-3;      // 2 mm of slack over every FP32 rounding in the kernel
    const double dx = centre[G.body][0] - pw[0], dy = centre[G.body][1] - pw[1], dz = centre[G.body][2] - pw[2];
    if (m.geom_type[h] == GT_PLANE) return Rw[2] * dx + Rw[5] * dy + Rw[8] * dz > need;
    if (m.geom_type[h] == GT_BOX) {
      // exact distance from the workspace centre to the box, not to its bounding sphere: a long thin box (a wall) has a
      // bounding sphere that swallows the whole workspace, which kept every link's pair with it alive - and sent each of
      // them through the narrow phase's early exit at every collocation point
      const double d3[3] = {dx, dy, dz};
      double out2 = 0.0;
      for (int a = 0; a < 3; ++a) {
        const double l = Rw[a] * d3[0] + Rw[3 + a] * d3[1] + Rw[6 + a] * d3[2];          // Rw^T (centre - box centre)
        const double e = std::max(std::fabs(l) - m.geom_size[3 * h + a], 0.0);
        out2 += e * e;
      }
      return std::sqrt(out2) > need;
    }
    const double rb = bound_radius(m.geom_type[h], &m.geom_size[3 * h]);
    return rb >= 0 && std::sqrt(dx * dx + dy * dy + dz * dz) > need + rb;
  };
  int n_static_pairs = 0, n_unreachable = 0;
  for (int ig = 0; ig < D.n_mgeom; ++ig) {
    const int g = mgeom_src[ig];
    D.mg[ig].pair_begin = D.n_pair;
    for (int h = 0; h < m.ngeom; ++h) {
      if (h == g || cluster[m.geom_body[h]] >= 0) continue;
      if (!pair_active(m, std::min(g, h), std::max(g, h))) continue;
      if (!pair_supported(m.geom_type[g], m.geom_type[h]))
        throw std::runtime_error("collision between geoms '" + m.geom_name[g] + "' and '" + m.geom_name[h] + "' is of an unsupported type pairing");
      if (unreachable(ig, h)) { ++n_unreachable; continue; }
      if (D.n_pair >= MAX_PAIR) throw std::runtime_error("too many geom pairs");
      D.pair[D.n_pair].s = static_geom(h);
      D.pair[D.n_pair].margin = (float)std::max(m.geom_margin[g], m.geom_margin[h]);
      ++D.n_pair;
    }
    D.mg[ig].pair_end = D.n_pair;
    // plane partners first (the kernel's broad phase runs the two kinds in separate branch-free loops)
    std::stable_partition(D.pair + D.mg[ig].pair_begin, D.pair + D.mg[ig].pair_end,
                          [&](const DevPair& pr) { return D.sg[pr.s].rbound < 0.0f; });
    D.mg[ig].pair_mid = D.mg[ig].pair_begin;
    while (D.mg[ig].pair_mid < D.mg[ig].pair_end && D.sg[D.pair[D.mg[ig].pair_mid].s].rbound < 0.0f) ++D.mg[ig].pair_mid;
  }
  // broad-phase test of every pair, precomputed in FP32 (see DevPair)
  for (int ig = 0; ig < D.n_mgeom; ++ig)
    for (int k = D.mg[ig].pair_begin; k < D.mg[ig].pair_end; ++k) {
      DevPair& pr = D.pair[k];
      const DevSGeom& S = D.sg[pr.s];
      pr.thr = std::max(pr.margin, 0.0f) + 1e-4f;
      if (S.rbound >= 0.0f) {
        // FP64 here, rounded once.  The kernel's FP32 expansion of |p - c|^2 errs by a few ulp of |p|^2 + |c|^2: the
        // squared radius is widened by 16 ulp of that magnitude (|p| bounded by the workspace bound, else 4 m)
        const DevMGeom& G = D.mg[ig];
        const double rr = (double)G.rbound + S.rbound + pr.thr;
        const double c2 = (double)S.pos[0] * S.pos[0] + (double)S.pos[1] * S.pos[1] + (double)S.pos[2] * S.pos[2];
        const double cen = std::sqrt(centre[G.body][0] * centre[G.body][0] + centre[G.body][1] * centre[G.body][1] + centre[G.body][2] * centre[G.body][2]);
        const double glen = std::sqrt((double)G.pos[0] * G.pos[0] + (double)G.pos[1] * G.pos[1] + (double)G.pos[2] * G.pos[2]);
        const double pmax = reach[G.body] < kInf ? cen + reach[G.body] + glen : 4.0;
        const double widen = 16.0 * 1.1920929e-7 * (pmax * pmax + c2);
        pr.plane = 0; pr.c[0] = -2.0f * S.pos[0]; pr.c[1] = -2.0f * S.pos[1]; pr.c[2] = -2.0f * S.pos[2];
        pr.c[3] = (float)(c2 - (rr * rr + widen));
      } else {
        const v3 n = mk(S.R[2], S.R[5], S.R[8]);
        pr.plane = 1; pr.c[0] = n.x; pr.c[1] = n.y; pr.c[2] = n.z; pr.c[3] = -(dot(n, ld3(S.pos)) + D.mg[ig].rbound + pr.thr);
      }
    }
  for (int ig = 0; ig < D.n_mgeom; ++ig)
    for (int ih = ig + 1; ih < D.n_mgeom; ++ih) {
      const int g = mgeom_src[ig], h = mgeom_src[ih];
      if (D.mg[ig].body == D.mg[ih].body) continue;   // rigid with respect to each other
      if (!pair_active(m, std::min(g, h), std::max(g, h))) continue;
      if (!pair_supported(m.geom_type[g], m.geom_type[h]))
        throw std::runtime_error("collision between geoms '" + m.geom_name[g] + "' and '" + m.geom_name[h] + "' is of an unsupported type pairing");
      if (D.n_mmpair >= MAX_MMPAIR) throw std::runtime_error("too many moving-moving geom pairs");
      D.mm[D.n_mmpair].a = ig; D.mm[D.n_mmpair].b = ih;
      D.mm[D.n_mmpair].margin = (float)std::max(m.geom_margin[g], m.geom_margin[h]);
      ++D.n_mmpair;
    }
  // a moving geom carries at most MAX_GEOM_PAIRS pair records (one mask word in the kernel's broad phase): geoms
  // with more partners are emitted again with the next slice of their pairs
  {
    std::vector<char> in_mm(D.n_mgeom, 0);
    for (int k = 0; k < D.n_mmpair; ++k) in_mm[D.mm[k].a] = in_mm[D.mm[k].b] = 1;
    bool rebuild = false;
    for (int ig = 0; ig < D.n_mgeom; ++ig) {
      const int np = D.mg[ig].pair_end - D.mg[ig].pair_begin;
      if (np > MAX_GEOM_PAIRS || (np == 0 && !in_mm[ig])) rebuild = true;
    }
    if (rebuild && !opt.sdf) {
      std::vector<DevMGeom> out;
      std::vector<int> src_out, first_of(D.n_mgeom, 0);
      for (int mb = 0; mb < D.n_mbody; ++mb) {
        const int gb = D.body[mb].geom_begin, ge = D.body[mb].geom_end;
        D.body[mb].geom_begin = (int)out.size();
        for (int ig = gb; ig < ge; ++ig) {
          first_of[ig] = (int)out.size();
          const DevMGeom& G = D.mg[ig];
          if (G.pair_end == G.pair_begin && !in_mm[ig]) continue;      // every partner is out of reach
          for (int k = G.pair_begin; k < G.pair_end || k == G.pair_begin; k += MAX_GEOM_PAIRS) {
            DevMGeom c = G;
            c.pair_begin = k; c.pair_end = std::min(k + (int)MAX_GEOM_PAIRS, G.pair_end);
            c.pair_mid = std::min(std::max(G.pair_mid, c.pair_begin), c.pair_end);
            out.push_back(c); src_out.push_back(mgeom_src[ig]);
          }
        }
        D.body[mb].geom_end = (int)out.size();
      }
      if ((int)out.size() > MAX_MGEOM) throw std::runtime_error("too many moving geoms after splitting their pair lists");
      for (int k = 0; k < D.n_mmpair; ++k) { D.mm[k].a = first_of[D.mm[k].a]; D.mm[k].b = first_of[D.mm[k].b]; }
      std::copy(out.begin(), out.end(), D.mg);
      D.n_mgeom = (int)out.size();
      mgeom_src = src_out;
    }
  }
  // Body-level broad phase.  A body that carries several geoms (a tool with its fingers, pads and couplers) gets one bounding
  // sphere around all of them and one affine record per static partner of any of them, stored behind the geom pairs; the
  // kernel tests those first and skips the body's geoms - their placement and their own pair tests - when the sphere is clear
  // of every partner.  Exact: every geom lies inside the body sphere.  Not for the planned free body (it is what the planner
  // steers among the obstacles: the test would rarely cull) and not worth it below three geoms.
  D.n_bpair = 0;
  for (int mb = 0; mb < D.n_mbody; ++mb) D.body[mb].bpair_begin = D.body[mb].bpair_mid = D.body[mb].bpair_end = 0;
  if (!opt.sdf)
    for (int mb = 0; mb < D.n_mbody; ++mb) {
      DevBody& B = D.body[mb];
      const int ng = B.geom_end - B.geom_begin;
      if (ng < 3 || B.jtype == JT_FREE4) continue;
      double cb[3] = {0, 0, 0};
      for (int ig = B.geom_begin; ig < B.geom_end; ++ig) for (int a = 0; a < 3; ++a) cb[a] += D.mg[ig].pos[a] / ng;
      double rb = 0.0;
      for (int ig = B.geom_begin; ig < B.geom_end; ++ig) {
        const DevMGeom& G = D.mg[ig];
        const double dx = G.pos[0] - cb[0], dy = G.pos[1] - cb[1], dz = G.pos[2] - cb[2];
        rb = std::max(rb, std::sqrt(dx * dx + dy * dy + dz * dz) + G.rbound);
      }
      rb *= 1.0 + 1e-6;
      std::map<int, float> partner;                       // static geom -> largest cull threshold among this body's pairs with it
      for (int ig = B.geom_begin; ig < B.geom_end; ++ig)
        for (int k = D.mg[ig].pair_begin; k < D.mg[ig].pair_end; ++k) {
          auto it = partner.find(D.pair[k].s);
          if (it == partner.end()) partner[D.pair[k].s] = D.pair[k].thr; else it->second = std::max(it->second, D.pair[k].thr);
        }
      if (partner.empty() || D.n_pair + D.n_bpair + (int)partner.size() > MAX_PAIR) continue;
      for (int a = 0; a < 3; ++a) B.bound[a] = (float)cb[a];
      B.bound[3] = (float)rb;
      B.bpair_begin = D.n_pair + D.n_bpair;
      const double cen = std::sqrt(centre[mb][0] * centre[mb][0] + centre[mb][1] * centre[mb][1] + centre[mb][2] * centre[mb][2]);
      const double blen = std::sqrt(cb[0] * cb[0] + cb[1] * cb[1] + cb[2] * cb[2]);
      for (int pass = 0; pass < 2; ++pass) {              // planes first, as in the geom tables
        for (const auto& kv : partner) {
          const DevSGeom& S = D.sg[kv.first];
          if ((S.rbound < 0.0f) != (pass == 0)) continue;
          DevPair& pr = D.pair[D.n_pair + D.n_bpair++];
          pr.s = kv.first; pr.margin = 0.0f; pr.thr = kv.second;
          if (S.rbound >= 0.0f) {
            const double rr = rb + S.rbound + pr.thr;
            const double c2 = (double)S.pos[0] * S.pos[0] + (double)S.pos[1] * S.pos[1] + (double)S.pos[2] * S.pos[2];
            const double pmax = reach[mb] < kInf ? cen + reach[mb] + blen : 4.0;
            const double widen = 16.0 * 1.1920929e-7 * (pmax * pmax + c2);
            pr.plane = 0; pr.c[0] = -2.0f * S.pos[0]; pr.c[1] = -2.0f * S.pos[1]; pr.c[2] = -2.0f * S.pos[2];
            pr.c[3] = (float)(c2 - (rr * rr + widen));
          } else {
            const v3 n = mk(S.R[2], S.R[5], S.R[8]);
            pr.plane = 1; pr.c[0] = n.x; pr.c[1] = n.y; pr.c[2] = n.z; pr.c[3] = -(dot(n, ld3(S.pos)) + (float)rb + pr.thr);
          }
        }
        if (pass == 0) B.bpair_mid = D.n_pair + D.n_bpair;
      }
      B.bpair_end = D.n_pair + D.n_bpair;
    }
  if (opt.sdf) {
    // The grid is the union of the static partners; it cannot express a per-geom collision mask, so every
    // sampled moving geom must see exactly the same static geoms, and nothing may collide moving-vs-moving.
    if (D.n_mmpair) throw std::runtime_error("SDF collision mode does not support pairs of geoms that move relative to each other");
    std::vector<int> ref;
    double margin = 0.0;
    for (int ig = 0; ig < D.n_mgeom; ++ig) {
      std::vector<int> mine;
      for (int k = D.mg[ig].pair_begin; k < D.mg[ig].pair_end; ++k) { mine.push_back(D.pair[k].s); margin = std::max(margin, (double)D.pair[k].margin); }
      std::sort(mine.begin(), mine.end());
      if (ig == 0) ref = mine;
      else if (mine != ref) throw std::runtime_error("SDF collision mode needs a uniform collision mask: geom '" + m.geom_name[mgeom_src[ig]] +
                                                     "' collides with a different set of static geoms than '" + m.geom_name[mgeom_src[0]] + "'");
    }
    D.sdf_mode = 1;
    for (int mb = 0; mb < D.n_mbody; ++mb) {
      D.body[mb].sph_begin = D.n_sphere;
      for (int ig = D.body[mb].geom_begin; ig < D.body[mb].geom_end; ++ig) {
        const int g = mgeom_src[ig];
        std::vector<double> sp;
        sample_spheres(m.geom_type[g], &m.geom_size[3 * g], opt.sphere_ratio, &sp);
        for (size_t k = 0; k + 3 < sp.size(); k += 4) {
          if (D.n_sphere >= MAX_SPHERE) throw std::runtime_error("SDF collision mode: more than " + std::to_string((int)MAX_SPHERE) +
                                                                 " sample spheres; raise sdf_sphere_ratio");
          DevSphere& S = D.sph[D.n_sphere++];
          const v3 c = add(ld3(D.mg[ig].pos), mulv(ldm(D.mg[ig].R), mk((float)sp[k], (float)sp[k + 1], (float)sp[k + 2])));
          S.c[0] = c.x; S.c[1] = c.y; S.c[2] = c.z; S.r = (float)sp[k + 3]; S.margin = (float)margin; S.body = mb;
        }
      }
      D.body[mb].sph_end = D.n_sphere;
    }
    D.n_pair = 0;        // the narrow phase is the grid
  }
  // --- pairs that never move: evaluated once with the same FP32 narrow phase ------------------
  float static_cost = 0.0f;
  for (int g = 0; g < m.ngeom; ++g)
    for (int h = g + 1; h < m.ngeom; ++h) {
      if (cluster[m.geom_body[g]] >= 0 || cluster[m.geom_body[h]] >= 0) continue;
      if (!pair_active(m, g, h)) continue;
      if (!pair_supported(m.geom_type[g], m.geom_type[h])) continue;
      ++n_static_pairs;
      DevSGeom A, B;
      auto fill = [&](int gg, DevSGeom& S) {
        const int gb = m.geom_body[gg];
        double Rg[9], Rw[9], pw[3];
        quat_to_mat(&m.geom_quat[4 * gg], Rg);
        mat_mul(xw[gb].R, Rg, Rw);
        mat_vec(xw[gb].R, &m.geom_pos[3 * gg], pw);
        for (int k = 0; k < 3; ++k) { S.pos[k] = (float)(pw[k] + xw[gb].p[k]); S.size[k] = (float)m.geom_size[3 * gg + k]; }
        for (int k = 0; k < 9; ++k) S.R[k] = (float)Rw[k];
        S.type = m.geom_type[gg]; S.rbound = -1.0f;
      };
      fill(g, A); fill(h, B);
      bool ok;
      const Prim P = static_prim(A), Q = static_prim(B);
      const float d = prim_sd(P, Q, &ok);
      static_cost += contact_cost(D.rule, d, (float)std::max(m.geom_margin[g], m.geom_margin[h]), P.p, Q.p);
    }
  D.static_cost = static_cost;
  bool chain = (D.n_mmpair == 0);
  for (int b = 0; b < D.n_mbody; ++b) if (D.body[b].parent != b - 1) chain = false;
  D.chain = chain ? 1 : 0;
  if (D.n_mbody == 0) throw std::runtime_error("nothing moves with the planned coordinates");
  if (info) {
    info->n_static_pairs = n_static_pairs;
    info->n_active_pairs = D.n_pair;
    info->n_unreachable_pairs = n_unreachable;
    info->mgeom_src = mgeom_src;
    info->sgeom_src = sgeom_src;
  }
  return D;
}

}  // namespace sspp
