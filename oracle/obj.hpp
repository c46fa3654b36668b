// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of the reference's Wavefront-OBJ subset parser,
// /root/reference/raytracer/src/obj_parser/mod.rs:15-191.
//
// Parity note ("unpinned" in the reference): Parser::as_group drains a HashMap
// (obj_parser/mod.rs:143), so the child order of named groups is random per run in the
// reference. The oracle fixes it to first-insertion order of the group names.
#pragma once
#include <cstdlib>
#include <sstream>
#include <string>
#include <utility>

#include "scene.hpp"

namespace orc {

struct ObjParser {
  size_t ignored = 0;
  std::vector<Point> vertices{Point::origin()};         // 1-based: dummy element 0 (mod.rs:31)
  std::vector<Vector> vertex_normals{Vector(0, 0, 0)};  // mod.rs:32
  std::vector<std::pair<std::string, std::unique_ptr<Group>>> groups;
  std::string selected_group = "default";

  ObjParser() { groups.emplace_back("default", std::make_unique<Group>()); }

  Group* find_group(const std::string& name) {
    for (auto& g : groups)
      if (g.first == name) return g.second.get();
    return nullptr;
  }
  static std::vector<std::string> split_ws(const std::string& line) {
    std::vector<std::string> out;
    size_t i = 0, n = line.size();
    while (i < n) {
      while (i < n && (line[i] == ' ' || line[i] == '\t' || line[i] == '\r' || line[i] == '\n' || line[i] == '\f' ||
                       line[i] == '\v'))
        i++;
      size_t j = i;
      while (j < n && !(line[j] == ' ' || line[j] == '\t' || line[j] == '\r' || line[j] == '\n' || line[j] == '\f' ||
                        line[j] == '\v'))
        j++;
      if (j > i) out.push_back(line.substr(i, j - i));
      i = j;
    }
    return out;
  }
  static double parse_f64(const std::string& s) {
    char* end = nullptr;
    double v = std::strtod(s.c_str(), &end);
    if (end == s.c_str() || *end != '\0') throw std::runtime_error("obj: bad float '" + s + "'");
    return v;
  }
  static size_t parse_usize(const std::string& s) {
    if (s.empty()) throw std::runtime_error("obj: empty index");
    size_t v = 0;
    for (char c : s) {
      if (c < '0' || c > '9') throw std::runtime_error("obj: bad index '" + s + "'");
      v = v * 10 + (size_t)(c - '0');
    }
    return v;
  }
  void parse(const std::string& contents) {  // mod.rs:38-42
    size_t pos = 0;
    while (pos <= contents.size()) {
      size_t nl = contents.find('\n', pos);
      if (nl == std::string::npos) nl = contents.size();
      if (nl > pos || nl < contents.size()) parse_line(contents.substr(pos, nl - pos));
      pos = nl + 1;
    }
  }
  void parse_line(const std::string& line) {  // mod.rs:44-101
    std::vector<std::string> items = split_ws(line);
    if (items.empty()) return;
    const std::string& kind = items[0];
    if (kind == "v") {
      vertices.emplace_back(parse_f64(items.at(1)), parse_f64(items.at(2)), parse_f64(items.at(3)));
    } else if (kind == "vn") {
      vertex_normals.emplace_back(parse_f64(items.at(1)), parse_f64(items.at(2)), parse_f64(items.at(3)));
    } else if (kind == "f") {
      Group* group = find_group(selected_group);
      if (line.find('/') == std::string::npos) {
        std::vector<size_t> idx;
        for (size_t i = 1; i < items.size(); i++) idx.push_back(parse_usize(items[i]));
        // fan_triangulation, mod.rs:103-116
        for (size_t i = 1; i + 1 < idx.size(); i++)
          group->add_child(std::make_unique<Triangle>(vertices.at(idx[0]), vertices.at(idx[i]), vertices.at(idx[i + 1])));
      } else {
        // mod.rs:70-83: first and LAST '/'-separated field of every item
        std::vector<std::pair<size_t, size_t>> faces;
        for (size_t i = 1; i < items.size(); i++) {
          const std::string& it = items[i];
          size_t a = it.find('/');
          size_t b = it.rfind('/');
          std::string first = it.substr(0, a);
          std::string last = (b == std::string::npos) ? it : it.substr(b + 1);
          faces.emplace_back(parse_usize(first), parse_usize(last));
        }
        // smooth_fan_triangulation, mod.rs:118-134
        for (size_t i = 1; i + 1 < faces.size(); i++)
          group->add_child(std::make_unique<SmoothTriangle>(
              vertices.at(faces[0].first), vertices.at(faces[i].first), vertices.at(faces[i + 1].first),
              vertex_normals.at(faces[0].second), vertex_normals.at(faces[i].second),
              vertex_normals.at(faces[i + 1].second)));
      }
    } else if (kind == "g") {  // mod.rs:90-95: (re)creates an EMPTY group under that name
      const std::string& name = items.at(1);
      selected_group = name;
      bool found = false;
      for (auto& g : groups)
        if (g.first == name) { g.second = std::make_unique<Group>(); found = true; }
      if (!found) groups.emplace_back(name, std::make_unique<Group>());
    } else {
      ignored++;
    }
  }
  // mod.rs:136-149
  std::unique_ptr<Group> as_group() {
    if (groups.size() == 1) {
      std::unique_ptr<Group> g = std::move(groups[0].second);
      groups.clear();
      return g;
    }
    auto group = std::make_unique<Group>();
    for (auto& g : groups)
      if (!g.second->children.empty()) group->add_child(std::move(g.second));
    groups.clear();
    return group;
  }
};

inline ObjParser parse_obj_string(const std::string& contents) {
  ObjParser p;
  p.parse(contents);
  return p;
}

}  // namespace orc
