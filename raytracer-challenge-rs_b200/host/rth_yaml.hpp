// rth_yaml.hpp — the YAML scene front end (caller of the hot path).
//
// Mirrors scene-parser/src/lib.rs: `define` elements first, in file order (named transforms = arrays,
// named materials = hashes, optional `extend`), then `add` elements in file order (camera | light |
// sphere | plane | cube; anything else is skipped like the reference's "unhandled element"). Quirks
// kept: a transform list is applied first-to-last as item * acc (lib.rs:337-341); the pattern
// `transform:` key is ignored and only stripes | checkers are known, anything else is the default
// (test) pattern (lib.rs:448-483); `field-of-view` must be a real, not an integer (lib.rs:395-397).
// The reference parses with yaml-rust 0.4.5; this reader handles the subset its scene files use:
// a top-level block sequence of block mappings, nested block mappings/sequences by indentation,
// flow sequences `[a, b, c]`, plain scalars and `#` comments.
// `extended = true` additionally accepts `add: cylinder | cone | group | csg | obj`, `shadow: false`
// and `divide: N`, which the reference's parser lacks (lib.rs:126-132); see Builder::shape.
#pragma once
#include <array>
#include <cstdlib>
#include <fstream>
#include <iterator>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "rth_obj.hpp"
#include "rth_scene.hpp"

namespace rth {
namespace yaml {

struct Node {
  enum Kind { SCALAR, SEQ, MAP } kind = SCALAR;
  std::string scalar;
  std::vector<Node> seq;
  std::vector<std::pair<std::string, Node>> map;  // insertion order
  const Node* get(const std::string& k) const {
    for (auto& kv : map)
      if (kv.first == k) return &kv.second;
    return nullptr;
  }
};

struct Line {
  int indent;
  std::string text;  // without indentation, comment and trailing blanks
  int number;
};

inline std::string trim(const std::string& s) {
  size_t a = s.find_first_not_of(" \t\r"), b = s.find_last_not_of(" \t\r");
  return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
}

inline std::vector<Line> split_lines(const std::string& text) {
  std::vector<Line> out;
  size_t pos = 0;
  int number = 0;
  while (pos <= text.size()) {
    size_t nl = text.find('\n', pos);
    std::string raw = text.substr(pos, nl == std::string::npos ? std::string::npos : nl - pos);
    pos = nl == std::string::npos ? text.size() + 1 : nl + 1;
    number++;
    // strip comments: '#' at line start or preceded by a blank (no quoted strings in scene files)
    for (size_t i = 0; i < raw.size(); i++)
      if (raw[i] == '#' && (i == 0 || raw[i - 1] == ' ' || raw[i - 1] == '\t')) { raw.erase(i); break; }
    size_t ind = raw.find_first_not_of(' ');
    if (ind == std::string::npos) continue;
    std::string t = trim(raw);
    if (t.empty() || t == "---") continue;
    out.push_back({(int)ind, t, number});
  }
  return out;
}

[[noreturn]] inline void fail_at(int line, const std::string& what) {
  throw std::runtime_error("yaml line " + std::to_string(line) + ": " + what);
}

inline Node parse_flow(const std::string& s, size_t& i, int line) {
  while (i < s.size() && s[i] == ' ') i++;
  Node n;
  if (i < s.size() && s[i] == '[') {
    n.kind = Node::SEQ;
    i++;
    for (;;) {
      while (i < s.size() && s[i] == ' ') i++;
      if (i >= s.size()) fail_at(line, "unterminated flow sequence");
      if (s[i] == ']') { i++; break; }
      n.seq.push_back(parse_flow(s, i, line));
      while (i < s.size() && s[i] == ' ') i++;
      if (i < s.size() && s[i] == ',') i++;
    }
    return n;
  }
  size_t b = i;
  while (i < s.size() && s[i] != ',' && s[i] != ']') i++;
  n.scalar = trim(s.substr(b, i - b));
  return n;
}

inline Node parse_inline(const std::string& s, int line) {
  size_t i = 0;
  std::string t = trim(s);
  if (!t.empty() && t[0] == '[') return parse_flow(t, i, line);
  Node n;
  n.scalar = t;
  return n;
}

// "key: rest" → true (a ':' followed by a blank or end of line, outside brackets)
inline bool split_key(const std::string& t, std::string& key, std::string& rest) {
  if (!t.empty() && t[0] == '[') return false;
  for (size_t i = 0; i < t.size(); i++)
    if (t[i] == ':' && (i + 1 == t.size() || t[i + 1] == ' ')) {
      key = trim(t.substr(0, i));
      rest = trim(t.substr(i + 1));
      return !key.empty();
    }
  return false;
}

struct Reader {
  std::vector<Line> L;
  size_t p = 0;

  Node block(int indent) {
    if (p >= L.size()) return Node();
    if (L[p].text[0] == '-' && (L[p].text.size() == 1 || L[p].text[1] == ' ')) return sequence(indent);
    return mapping(indent);
  }

  Node sequence(int indent) {
    Node n;
    n.kind = Node::SEQ;
    while (p < L.size() && L[p].indent == indent && L[p].text[0] == '-' && (L[p].text.size() == 1 || L[p].text[1] == ' ')) {
      Line cur = L[p];
      std::string rest = cur.text.size() > 1 ? cur.text.substr(2) : std::string();
      size_t lead = rest.find_first_not_of(' ');
      int item_indent = indent + 2 + (lead == std::string::npos ? 0 : (int)lead);
      rest = trim(rest);
      std::string k, v;
      if (rest.empty()) {
        p++;
        if (p < L.size() && L[p].indent > indent) n.seq.push_back(block(L[p].indent));
        else n.seq.push_back(Node());
      } else if (split_key(rest, k, v)) {
        // "- key: value" opens a mapping whose further keys sit at item_indent
        L[p].indent = item_indent;
        L[p].text = rest;
        n.seq.push_back(mapping(item_indent));
      } else {
        p++;
        n.seq.push_back(parse_inline(rest, cur.number));
      }
    }
    return n;
  }

  Node mapping(int indent) {
    Node n;
    n.kind = Node::MAP;
    while (p < L.size() && L[p].indent == indent) {
      std::string k, v;
      if (!split_key(L[p].text, k, v)) fail_at(L[p].number, "expected 'key: value'");
      int number = L[p].number;
      p++;
      if (!v.empty()) {
        n.map.push_back({k, parse_inline(v, number)});
      } else if (p < L.size() && (L[p].indent > indent || (L[p].indent == indent && L[p].text[0] == '-' &&
                                                           (L[p].text.size() == 1 || L[p].text[1] == ' ')))) {
        n.map.push_back({k, block(L[p].indent)});  // a block sequence may sit at the key's own indent
      } else {
        n.map.push_back({k, Node()});
      }
    }
    if (p < L.size() && L[p].indent > indent) fail_at(L[p].number, "bad indentation");
    return n;
  }
};

inline Node parse(const std::string& text) {
  Reader r;
  r.L = split_lines(text);
  if (r.L.empty()) return Node();
  Node root = r.block(r.L[0].indent);
  if (r.p != r.L.size()) fail_at(r.L[r.p].number, "unexpected content");
  return root;
}

}  // namespace yaml

// ---- scene builder (scene-parser/src/lib.rs) -------------------------------------------------------

struct YamlScene {
  World world;
  std::unique_ptr<Camera> camera;
  int skipped = 0;  // `add:` kinds the parser does not handle (lib.rs:132)
};

namespace yaml {

[[noreturn]] inline void bad(const std::string& what) { throw std::runtime_error("scene parser: " + what); }

// yaml-rust scalar typing: i64 first, then f64, else string (yaml.rs Yaml::from_str)
inline bool as_integer(const Node& n, long long& out) {
  if (n.kind != Node::SCALAR || n.scalar.empty()) return false;
  const std::string& s = n.scalar;
  size_t i = (s[0] == '-' || s[0] == '+') ? 1 : 0;
  if (i >= s.size()) return false;
  for (size_t k = i; k < s.size(); k++)
    if (s[k] < '0' || s[k] > '9') return false;
  out = std::strtoll(s.c_str(), nullptr, 10);
  return true;
}
inline bool as_real(const Node& n, double& out) {
  long long i;
  if (n.kind != Node::SCALAR || n.scalar.empty() || as_integer(n, i)) return false;
  const std::string& s = n.scalar;
  for (char c : s)  // plain decimal / exponent notation only (what str::parse::<f64> shares with strtod)
    if (!((c >= '0' && c <= '9') || c == '.' || c == '-' || c == '+' || c == 'e' || c == 'E')) return false;
  char* end = nullptr;
  out = std::strtod(s.c_str(), &end);
  return end && *end == 0;
}
// to_f64, lib.rs:488-497
inline double to_f64(const Node& n, const char* what) {
  long long i;
  double d;
  if (as_integer(n, i)) return (double)i;
  if (as_real(n, d)) return d;
  bad(std::string("expected a number for ") + what);
}
inline void to_vec3(const Node* n, const char* what, double out[3]) {  // to_point/to_vector/to_color
  if (!n || n->kind != Node::SEQ || n->seq.size() != 3) bad(std::string("expected [x, y, z] for ") + what);
  for (int k = 0; k < 3; k++) out[k] = to_f64(n->seq[k], what);
}
inline const Node& required(const Node& h, const char* key) {  // get_required_attribute, lib.rs:481-485
  const Node* n = h.get(key);
  if (!n) bad(std::string("missing required key ") + key);
  return *n;
}

struct Builder {
  std::map<std::string, Mat4> transforms;
  std::map<std::string, Material> materials;
  bool extended = false;
  std::string base_dir;  // directory `obj` file paths are relative to

  Mat4 transform_item(const Node& n) {  // lib.rs:349-373
    if (n.kind == Node::SEQ) {
      if (n.seq.empty() || n.seq[0].kind != Node::SCALAR) bad("bad transform item");
      const std::string& kind = n.seq[0].scalar;
      std::vector<double> a;
      for (size_t k = 1; k < n.seq.size(); k++) a.push_back(to_f64(n.seq[k], "transform argument"));
      auto need = [&](size_t c) { if (a.size() < c) bad("too few arguments for " + kind); };
      if (kind == "scale") { need(3); return scaling(a[0], a[1], a[2]); }
      if (kind == "translate") { need(3); return translation(a[0], a[1], a[2]); }
      if (kind == "rotate-x") { need(1); return rotation_x(a[0]); }
      if (kind == "rotate-y") { need(1); return rotation_y(a[0]); }
      if (kind == "rotate-z") { need(1); return rotation_z(a[0]); }
      bad("unknown transform " + kind);
    }
    if (n.kind == Node::SCALAR) {
      auto it = transforms.find(n.scalar);
      if (it == transforms.end()) bad("undefined transform " + n.scalar);
      return it->second;
    }
    bad("bad transform item");
  }
  Mat4 transform(const Node& n) {  // lib.rs:333-347: item * acc, first to last
    if (n.kind != Node::SEQ) bad("transform must be a list");
    Mat4 acc = Mat4::identity();
    for (const Node& item : n.seq) acc = transform_item(item) * acc;
    return acc;
  }

  static Pattern pattern(const Node& n) {  // lib.rs:448-479; `transform:` is not read
    if (n.kind != Node::MAP) bad("pattern must be a mapping");
    const Node* type = n.get("type");
    if (!type || type->kind != Node::SCALAR) bad("pattern needs a type");
    const Node* colors = n.get("colors");
    if (!colors || colors->kind != Node::SEQ) bad("pattern needs colors");
    std::vector<std::array<double, 3>> c;
    for (const Node& cn : colors->seq) {
      std::array<double, 3> v;
      to_vec3(&cn, "pattern color", v.data());
      c.push_back(v);
    }
    int kind = RTC_PAT_TEST;
    if (type->scalar == "stripes") kind = RTC_PAT_STRIPE;
    else if (type->scalar == "checkers") kind = RTC_PAT_CHECKERS;
    if (kind == RTC_PAT_TEST) return Pattern();
    if (c.size() < 2) bad("pattern needs two colors");  // the reference indexes colors[1] and panics
    return make_pattern(kind, c[0][0], c[0][1], c[0][2], c[1][0], c[1][1], c[1][2]);
  }

  static void apply_material_keys(Material& m, const Node& h) {  // lib.rs:212-253 == 286-327
    if (const Node* n = h.get("color")) to_vec3(n, "color", m.color);
    if (const Node* n = h.get("pattern")) m.set_pattern(pattern(*n));
    if (const Node* n = h.get("ambient")) m.ambient = to_f64(*n, "ambient");
    if (const Node* n = h.get("diffuse")) m.diffuse = to_f64(*n, "diffuse");
    if (const Node* n = h.get("specular")) m.specular = to_f64(*n, "specular");
    if (const Node* n = h.get("shininess")) m.shininess = to_f64(*n, "shininess");
    if (const Node* n = h.get("reflective")) m.reflective = to_f64(*n, "reflective");
    if (const Node* n = h.get("transparency")) m.transparency = to_f64(*n, "transparency");
    if (const Node* n = h.get("refractive-index")) m.refractive_index = to_f64(*n, "refractive-index");
  }
  Material material(const Node& n) {  // lib.rs:200-260
    if (n.kind == Node::SCALAR) {
      auto it = materials.find(n.scalar);
      if (it == materials.end()) bad("undefined material " + n.scalar);
      return it->second;
    }
    if (n.kind != Node::MAP) bad("bad material");
    Material m;
    apply_material_keys(m, n);
    return m;
  }

  void define(const Node& el) {  // lib.rs:139-182
    const Node& name = required(el, "define");
    if (name.kind != Node::SCALAR) bad("define needs a name");
    const Node* value = el.get("value");
    if (!value) bad("define needs a value");
    if (value->kind == Node::SEQ) {
      transforms[name.scalar] = transform(*value);
    } else if (value->kind == Node::MAP) {
      if (const Node* ext = el.get("extend")) {
        if (ext->kind != Node::SCALAR) bad("extend needs a name");
        auto it = materials.find(ext->scalar);
        if (it == materials.end()) bad("undefined material " + ext->scalar);
        Material m = it->second;
        apply_material_keys(m, *value);
        materials[name.scalar] = m;
      } else {
        materials[name.scalar] = material(*value);
      }
    } else {
      bad("define value must be a list or a mapping");
    }
  }

  // parse_shape (lib.rs:184-198): sphere | plane | cube, transform first, then material. nullptr = a kind
  // the parser does not handle. With `extended`: cylinder/cone (min, max, closed), group (children:
  // list of shapes, optional divide: N), csg (operation, left, right), obj (file: path relative to the
  // scene file; optional divide: N) and `shadow: false` on any shape — the constructs config 3b and 4
  // need to be expressible as data. Order inside a shape follows the reference's
  // bins: build children, then set_transform, then set_material, then divide.
  ShapeP shape(const Node& el, int depth) {
    if (depth > 64) bad("shapes nested too deeply");
    const Node& kind = required(el, "add");
    if (kind.kind != Node::SCALAR) bad("invalid add element");
    const std::string& k = kind.scalar;
    ShapeP s;
    if (k == "sphere") s = Shape::sphere();
    else if (k == "plane") s = Shape::plane();
    else if (k == "cube") s = Shape::cube();
    else if (!extended) return nullptr;
    else if (k == "cylinder" || k == "cone") {
      double mn = -INF, mx = INF;
      bool closed = false;
      if (const Node* n = el.get("min")) mn = to_f64(*n, "min");
      if (const Node* n = el.get("max")) mx = to_f64(*n, "max");
      if (const Node* n = el.get("closed")) closed = n->scalar == "true";
      s = k == "cylinder" ? Shape::cylinder(mn, mx, closed) : Shape::cone(mn, mx, closed);
    } else if (k == "group") {
      s = Shape::group();
      if (const Node* ch = el.get("children")) {
        if (ch->kind != Node::SEQ) bad("children must be a list");
        for (const Node& c : ch->seq) {
          if (c.kind != Node::MAP) bad("a child must be a mapping");
          ShapeP cs = shape(c, depth + 1);
          if (!cs) bad("unknown shape kind in group");
          s->add_child(std::move(cs));
        }
      }
    } else if (k == "csg") {
      const Node& op = required(el, "operation");
      int o = op.scalar == "union" ? RTC_CSG_UNION : op.scalar == "intersection" ? RTC_CSG_INTERSECTION
              : op.scalar == "difference" ? RTC_CSG_DIFFERENCE : -1;
      if (o < 0) bad("unknown csg operation " + op.scalar);
      const Node& l = required(el, "left");
      const Node& r = required(el, "right");
      if (l.kind != Node::MAP || r.kind != Node::MAP) bad("csg operands must be mappings");
      ShapeP ls = shape(l, depth + 1), rs = shape(r, depth + 1);
      if (!ls || !rs) bad("unknown shape kind in csg");
      s = Shape::csg(o, std::move(ls), std::move(rs));
    } else if (k == "obj") {
      const Node& file = required(el, "file");
      std::string path = file.scalar;
      if (!path.empty() && path[0] != '/' && !base_dir.empty()) path = base_dir + "/" + path;
      std::ifstream in(path, std::ios::binary);
      if (!in) bad("cannot open obj file " + path);
      std::string text((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
      ObjParser p;
      p.parse(text.data(), text.size());
      s = p.as_group();
    } else {
      return nullptr;
    }
    if (const Node* n = el.get("transform")) s->set_transform(transform(*n));
    if (const Node* n = el.get("material")) s->set_material(material(*n));
    if (extended) {
      if (const Node* n = el.get("shadow"))
        if (n->scalar == "false") s->no_shadow();
      if (const Node* n = el.get("divide")) {
        long long t;
        if (!as_integer(*n, t) || t < 1) bad("divide needs a positive integer");
        if (s->is_group()) s->divide((size_t)t);
      }
    }
    return s;
  }

  void add(const Node& el, YamlScene& out) {  // lib.rs:121-137
    const Node& kind = required(el, "add");
    if (kind.kind != Node::SCALAR) bad("invalid add element");
    const std::string& k = kind.scalar;
    if (k == "camera") {  // lib.rs:391-431
      long long w, h;
      double fov, from[3], to[3], up[3];
      if (!as_integer(required(el, "width"), w)) bad("width must be an integer");
      if (!as_integer(required(el, "height"), h)) bad("height must be an integer");
      if (!as_real(required(el, "field-of-view"), fov)) bad("field-of-view must be a real");
      to_vec3(&required(el, "from"), "from", from);
      to_vec3(&required(el, "to"), "to", to);
      to_vec3(&required(el, "up"), "up", up);
      out.camera.reset(new Camera((size_t)w, (size_t)h, fov));
      out.camera->set_transform(view_transform(Vec3{from[0], from[1], from[2]}, Vec3{to[0], to[1], to[2]},
                                               Vec3{up[0], up[1], up[2]}));
      return;
    }
    if (k == "light") {  // lib.rs:433-446
      PointLight l;
      to_vec3(&required(el, "at"), "at", l.position);
      to_vec3(&required(el, "intensity"), "intensity", l.intensity);
      out.world.add_light(l);
      return;
    }
    ShapeP s = shape(el, 0);
    if (!s) {
      out.skipped++;
      return;
    }
    out.world.add_object(std::move(s));
  }
};

}  // namespace yaml

// SceneParser::load_file (lib.rs:90-119): all `define` elements first, then all `add` elements.
inline std::unique_ptr<YamlScene> load_yaml_scene(const std::string& text, bool extended = false,
                                                  const std::string& base_dir = std::string()) {
  yaml::Node root = yaml::parse(text);
  if (root.kind != yaml::Node::SEQ) yaml::bad("the scene file must be a list of elements");
  yaml::Builder b;
  b.extended = extended;
  b.base_dir = base_dir;
  std::unique_ptr<YamlScene> out(new YamlScene);
  for (const yaml::Node& el : root.seq)
    if (el.kind == yaml::Node::MAP && el.get("define")) b.define(el);
  for (const yaml::Node& el : root.seq)
    if (el.kind == yaml::Node::MAP && el.get("add")) b.add(el, *out);
  return out;
}

}  // namespace rth
