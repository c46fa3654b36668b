// rth_obj.hpp — Wavefront OBJ front end of the B200 backend (SURVEY §8-f "next" #1): one pass over
// the text with a pointer scanner (no per-line strings), same accepted subset and the same quirks
// as the reference's parser, raytracer/src/obj_parser/mod.rs:23-149:
//   v / vn / f / g records, everything else counted as ignored (:97-99); 1-based indices with a dummy
//   element 0 (:31-32); faces containing '/' take the FIRST and the LAST slash field as
//   vertex / normal index (:70-83); fan triangulation (:103-134); `g name` (re)creates an empty
//   group under that name (:90-95); as_group returns the default group alone if no `g` was seen,
//   else a group of the non-empty named groups (:136-149) — in first-insertion order here, where
//   the reference's HashMap::drain order is random.
#pragma once
#include <cstdlib>
#include <string>
#include <utility>
#include <vector>

#include "rth_scene.hpp"

namespace rth {

class ObjParser {
 public:
  size_t ignored = 0;
  std::vector<Vec3> vertices{Vec3{0, 0, 0}};
  std::vector<Vec3> normals{Vec3{0, 0, 0}};

  ObjParser() { groups_.emplace_back("default", Shape::group()); }

  void parse(const char* text, size_t len) {
    const char* p = text;
    const char* end = text + len;
    while (p < end) {
      const char* eol = p;
      while (eol < end && *eol != '\n') eol++;
      line(p, eol);
      p = eol + 1;
    }
  }
  ShapeP as_group() {
    if (groups_.size() == 1) return std::move(groups_[0].second);
    ShapeP g = Shape::group();
    for (auto& e : groups_)
      if (!e.second->children.empty()) g->add_child(std::move(e.second));
    return g;
  }

 private:
  std::vector<std::pair<std::string, ShapeP>> groups_;
  size_t selected_ = 0;

  static bool ws(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\f' || c == '\v' || c == '\n'; }
  static void skip_ws(const char*& p, const char* e) { while (p < e && ws(*p)) p++; }
  static bool token(const char*& p, const char* e, const char*& b, const char*& t) {
    skip_ws(p, e);
    if (p >= e) return false;
    b = p;
    while (p < e && !ws(*p)) p++;
    t = p;
    return true;
  }
  static double number(const char* b, const char* t) {
    std::string s(b, t);
    char* endp = nullptr;
    double v = std::strtod(s.c_str(), &endp);
    if (endp == s.c_str() || *endp) throw std::runtime_error("obj: bad number '" + s + "'");
    return v;
  }
  static size_t index(const char* b, const char* t) {
    if (b == t) throw std::runtime_error("obj: empty index");
    size_t v = 0;
    for (const char* q = b; q < t; q++) {
      if (*q < '0' || *q > '9') throw std::runtime_error("obj: bad index '" + std::string(b, t) + "'");
      v = v * 10 + (size_t)(*q - '0');
    }
    return v;
  }
  Vec3 three(const char*& p, const char* e) {
    const char *b, *t;
    double v[3];
    for (int i = 0; i < 3; i++) {
      if (!token(p, e, b, t)) throw std::runtime_error("obj: record needs 3 numbers");
      v[i] = number(b, t);
    }
    return Vec3{v[0], v[1], v[2]};
  }
  const Vec3& vert(size_t i) const {
    if (i >= vertices.size()) throw std::runtime_error("obj: vertex index out of range");
    return vertices[i];
  }
  const Vec3& norm(size_t i) const {
    if (i >= normals.size()) throw std::runtime_error("obj: normal index out of range");
    return normals[i];
  }
  void line(const char* p, const char* e) {
    const char *b, *t;
    if (!token(p, e, b, t)) return;
    size_t kl = (size_t)(t - b);
    if (kl == 1 && b[0] == 'v') {
      vertices.push_back(three(p, e));
    } else if (kl == 2 && b[0] == 'v' && b[1] == 'n') {
      normals.push_back(three(p, e));
    } else if (kl == 1 && b[0] == 'f') {
      bool slashes = false;
      for (const char* q = b; q < e; q++) slashes |= (*q == '/');  // `line.contains('/')`
      std::vector<std::pair<size_t, size_t>> idx;
      while (token(p, e, b, t)) {
        if (!slashes) {
          idx.emplace_back(index(b, t), 0);
        } else {
          const char* first_end = b;
          while (first_end < t && *first_end != '/') first_end++;
          const char* last_begin = t;
          while (last_begin > b && last_begin[-1] != '/') last_begin--;
          idx.emplace_back(index(b, first_end), index(last_begin, t));
        }
      }
      Shape& g = *groups_[selected_].second;
      for (size_t i = 1; i + 1 < idx.size(); i++) {
        if (!slashes)
          g.add_child(Shape::triangle(vert(idx[0].first), vert(idx[i].first), vert(idx[i + 1].first)));
        else
          g.add_child(Shape::smooth_triangle(vert(idx[0].first), vert(idx[i].first), vert(idx[i + 1].first),
                                             norm(idx[0].second), norm(idx[i].second), norm(idx[i + 1].second)));
      }
    } else if (kl == 1 && b[0] == 'g') {
      if (!token(p, e, b, t)) throw std::runtime_error("obj: g needs a name");
      std::string name(b, t);
      for (size_t i = 0; i < groups_.size(); i++)
        if (groups_[i].first == name) {
          groups_[i].second = Shape::group();
          selected_ = i;
          return;
        }
      groups_.emplace_back(name, Shape::group());
      selected_ = groups_.size() - 1;
    } else {
      ignored++;
    }
  }
};

}  // namespace rth
