// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of the reference's render driver. Paths relative to /root/reference/raytracer/src.
//   world.rs:16-151          World (intersect, shade_hit, color_at, is_shadowed, reflected_color,
//                            refracted_color, MAX_RECURSION_DEPTH, Default)
//   camera.rs:20-259         Camera (new, ray_for_pixel, rays_for_pixel, get_offsets, render,
//                            render_multithreaded), RenderOpts, AASamples
//   canvas.rs:8-48           Canvas (row-major y*width+x)
//   image/ppm.rs:24-75       canvas_to_ppm, scale_color_component  (png.rs:29-31 is the same quantiser)
#pragma once
#include <string>
#include <thread>

#include "scene.hpp"

namespace orc {

constexpr size_t MAX_RECURSION_DEPTH = 5;  // world.rs:16

struct World {
  std::vector<ShapeP> objects;
  std::vector<PointLight> lights;

  // world.rs:137-151
  static World default_world() {
    World w;
    w.lights.emplace_back(Point(-10, 10, -10), Color(1.0, 1.0, 1.0));
    auto s1 = std::make_unique<Sphere>();
    Material m;
    m.color = Color(0.8, 1.0, 0.6);
    m.diffuse = 0.7;
    m.specular = 0.2;
    s1->set_material(m);
    auto s2 = std::make_unique<Sphere>();
    s2->set_transform(scaling(0.5, 0.5, 0.5));
    w.objects.push_back(std::move(s1));
    w.objects.push_back(std::move(s2));
    return w;
  }
  void add_light(const PointLight& l) { lights.push_back(l); }  // world.rs:83
  void add_object(ShapeP s) { objects.push_back(std::move(s)); }  // world.rs:87-93

  // world.rs:31-38 — every top-level object, concatenated, then sorted
  Intersections intersect(const Ray& ray) const {
    Intersections xs;
    for (const ShapeP& o : objects) o->intersect(ray, xs);
    return intersections(xs);
  }
  // world.rs:95-105
  bool is_shadowed(Point point, const PointLight& light) const {
    counters().shadow_setup++;
    counters().rays_shadow++;
    Vector v = light.position - point;
    double distance = v.magnitude();
    Vector direction = v.normalize();
    Ray r(point, direction);
    Intersections xs = intersect(r);
    const Intersection* h = shadow_hit(xs);
    return h != nullptr && h->t < distance;
  }
  // world.rs:107-114
  Color reflected_color(const Computations& comps, size_t remaining) const {
    if (equal(comps.object->material().reflective, 0.0) || remaining == 0) return Color::black();
    counters().rays_reflect++;
    Ray reflect_ray(comps.over_point, comps.reflectv);
    Color color = color_at(reflect_ray, remaining - 1);
    return color * comps.object->material().reflective;
  }
  // world.rs:116-134
  Color refracted_color(const Computations& comps, size_t remaining) const {
    if (equal(comps.object->material().transparency, 0.0) || remaining == 0) return Color::black();
    counters().refract++;
    double n_ratio = comps.n1 / comps.n2;
    double cos_i = dot(comps.eyev, comps.normalv);
    double sin2_t = n_ratio * n_ratio * (1.0 - cos_i * cos_i);
    if (sin2_t > 1.0) return Color::black();
    double cos_t = std::sqrt(1.0 - sin2_t);
    Vector direction = comps.normalv * (n_ratio * cos_i - cos_t) - comps.eyev * n_ratio;  // not normalised
    Ray refract_ray(comps.under_point, direction);
    counters().rays_refract++;
    return color_at(refract_ray, remaining - 1) * comps.object->material().transparency;
  }
  // world.rs:40-68
  Color shade_hit(const Computations& comps, size_t remaining) const {
    counters().shade++;
    Color surface(0.0, 0.0, 0.0);  // Sum: fold from (0,0,0) in light order
    for (const PointLight& light : lights) {
      bool shadowed = is_shadowed(comps.over_point, light);
      surface = surface + comps.object->material().lighting(*comps.object, light, comps.over_point, comps.eyev,
                                                            comps.normalv, shadowed);
    }
    Color reflected = reflected_color(comps, remaining);
    Color refracted = refracted_color(comps, remaining);
    const Material& material = comps.object->material();
    if (material.reflective > 0.0 && material.transparency > 0.0) {
      double reflectance = comps.schlick();
      return surface + reflected * reflectance + refracted * (1.0 - reflectance);
    }
    return surface + reflected + refracted;
  }
  // world.rs:70-81
  Color color_at(const Ray& ray, size_t remaining) const {
    Intersections xs = intersect(ray);
    const Intersection* h = hit(xs);
    if (!h) return Color::black();
    Computations comps = prepare_computations(*h, ray, xs);
    return shade_hit(comps, remaining);
  }
};

// canvas.rs
struct Canvas {
  size_t width = 0, height = 0;
  std::vector<Color> pixels;
  Canvas() = default;
  Canvas(size_t w, size_t h) : width(w), height(h), pixels(w * h, Color(0.0, 0.0, 0.0)) {}
  size_t pixel_idx(size_t x, size_t y) const {  // canvas.rs:44-48
    if (!(x < width) || !(y < height)) throw std::out_of_range("Canvas::pixel_idx");
    return y * width + x;
  }
  Color get_pixel(size_t x, size_t y) const { return pixels[pixel_idx(x, y)]; }
  void set_pixel(size_t x, size_t y, Color c) { pixels[pixel_idx(x, y)] = c; }
};

// ppm.rs:73-75 / png.rs:29-31 — (v*255).round() as u8: round half away from zero, saturating, NaN→0
inline uint8_t scale_color_component(double value) {
  double r = std::round(value * 255.0);
  if (std::isnan(r)) return 0;
  if (r <= 0.0) return 0;
  if (r >= 255.0) return 255;
  return (uint8_t)r;
}

// ppm.rs:24-51 — P3 text; a value is moved to the next line when it would push the line over 70
// characters (the line is right-trimmed); every image row ends its line.
inline std::string canvas_to_ppm(const Canvas& canvas) {
  std::string ppm = "P3\n" + std::to_string(canvas.width) + " " + std::to_string(canvas.height) + "\n255\n";
  for (size_t j = 0; j < canvas.height; j++) {
    std::string line;
    for (size_t i = 0; i < canvas.width; i++) {
      Color c = canvas.get_pixel(i, j);
      std::string pixel[3] = {std::to_string((int)scale_color_component(c.red)),
                              std::to_string((int)scale_color_component(c.green)),
                              std::to_string((int)scale_color_component(c.blue))};
      for (int idx = 0; idx < 3; idx++) {
        const std::string& val = pixel[idx];
        if (line.size() + val.size() > 70) {
          size_t e = line.find_last_not_of(" \t\n\r");
          ppm += (e == std::string::npos) ? std::string() : line.substr(0, e + 1);
          ppm += "\n";
          line.clear();
        }
        line += val;
        if (idx < 2) line += " ";
      }
      if (i + 1 < canvas.width) line += " ";
    }
    ppm += line;
    ppm += "\n";
  }
  return ppm;
}

// camera.rs
enum class AASamples : int { X1 = 1, X2 = 2, X4 = 4, X8 = 8, X16 = 16 };

struct RenderOpts {  // camera.rs:220-252
  size_t num_threads = 1;
  AASamples aa_samples = AASamples::X1;
};

struct Camera {
  size_t hsize, vsize;
  double field_of_view;
  Matrix transform = Matrix::identity(4, 4);
  Matrix transform_inverse = Matrix::identity(4, 4);
  double pixel_size, half_width, half_height;
  RenderOpts render_opts;

  Camera(size_t h, size_t v, double fov) : hsize(h), vsize(v), field_of_view(fov) {  // camera.rs:33-55
    double half_view = std::tan(fov / 2.0);
    double aspect = (double)hsize / (double)vsize;
    if (aspect >= 1.0) {
      half_width = half_view;
      half_height = half_view / aspect;
    } else {
      half_width = half_view * aspect;
      half_height = half_view;
    }
    pixel_size = half_width * 2.0 / (double)hsize;
  }
  void set_transform(const Matrix& t) {  // camera.rs:128-131
    transform = t;
    transform_inverse = transform.inverse();
  }
  Ray ray_with_offset(size_t px, size_t py, double ox, double oy) const {
    counters().ray_gen++;
    double xoffset = ((double)px + ox) * pixel_size;
    double yoffset = ((double)py + oy) * pixel_size;
    double world_x = half_width - xoffset;
    double world_y = half_height - yoffset;
    Point pixel = transform_inverse * Point(world_x, world_y, -1.0);
    Point origin = transform_inverse * Point::origin();
    Vector direction = (pixel - origin).normalize();
    return Ray(origin, direction);
  }
  Ray ray_for_pixel(size_t px, size_t py) const { return ray_with_offset(px, py, 0.5, 0.5); }  // camera.rs:57-69

  // camera.rs:92-126
  static std::vector<std::pair<double, double>> get_offsets(AASamples s) {
    switch (s) {
      case AASamples::X1: return {{0.5, 0.5}};
      case AASamples::X2: return {{0.25, 0.5}, {0.75, 0.5}};
      case AASamples::X4: return {{0.25, 0.25}, {0.75, 0.25}, {0.25, 0.75}, {0.75, 0.75}};
      case AASamples::X8:
        return {{0.25, 0.25}, {0.5, 0.25}, {0.75, 0.25}, {0.25, 0.5}, {0.75, 0.5}, {0.25, 0.75}, {0.5, 0.75}, {0.75, 0.75}};
      case AASamples::X16:
        return {{0.125, 0.125}, {0.375, 0.125}, {0.625, 0.125}, {0.875, 0.125}, {0.125, 0.375}, {0.375, 0.375},
                {0.625, 0.375}, {0.875, 0.375}, {0.125, 0.625}, {0.375, 0.625}, {0.625, 0.625}, {0.875, 0.625},
                {0.125, 0.875}, {0.375, 0.875}, {0.625, 0.875}, {0.875, 0.875}};
    }
    return {};
  }
  std::vector<Ray> rays_for_pixel(size_t px, size_t py) const {  // camera.rs:71-90
    std::vector<Ray> rays;
    for (auto& o : get_offsets(render_opts.aa_samples)) rays.push_back(ray_with_offset(px, py, o.first, o.second));
    return rays;
  }

  // camera.rs:133-148 (the println! progress lines are dropped)
  Canvas render(const World& world) const {
    Canvas image(hsize, vsize);
    for (size_t y = 0; y < vsize; y++)
      for (size_t x = 0; x < hsize; x++) {
        Ray ray = ray_for_pixel(x, y);
        counters().rays_primary++;
        counters().pixels++;
        image.set_pixel(x, y, world.color_at(ray, MAX_RECURSION_DEPTH));
      }
    return image;
  }
  // The per-pixel body of render_multithreaded, camera.rs:178-187
  Color pixel_aa(const World& world, size_t x, size_t y) const {
    std::vector<Ray> rays = rays_for_pixel(x, y);
    std::vector<Color> colors;
    for (const Ray& ray : rays) {
      counters().rays_primary++;
      colors.push_back(world.color_at(ray, MAX_RECURSION_DEPTH));
    }
    counters().pixels++;
    return color_average(colors);
  }
  // Rows [y0,y1) only, into a full-size canvas: what one thread of render_multithreaded does
  // (camera.rs:168-189). Used by tests to sample big frames.
  void render_rows(const World& world, size_t y0, size_t y1, Canvas& image) const {
    for (size_t y = y0; y < y1; y++)
      for (size_t x = 0; x < hsize; x++) image.set_pixel(x, y, pixel_aa(world, x, y));
  }
  // camera.rs:150-217 — contiguous row blocks, rows/num_threads each, last thread takes the rest.
  // `totals` receives the summed per-thread counters.
  Canvas render_multithreaded(const World& world, Counters* totals = nullptr) const {
    Canvas image(hsize, vsize);
    size_t rows = vsize;
    size_t num_threads = render_opts.num_threads;
    size_t rows_per_thread = rows / num_threads;
    std::vector<std::thread> handles;
    std::vector<Counters> per_thread(num_threads);
    for (size_t i = 0; i < num_threads; i++) {
      handles.emplace_back([&, i]() {
        size_t start = i * rows_per_thread, end = i * rows_per_thread + rows_per_thread;
        if (i == num_threads - 1) end = rows;
        counters() = Counters();
        render_rows(world, start, end, image);
        per_thread[i] = counters();
      });
    }
    for (auto& h : handles) h.join();
    if (totals)
      for (auto& c : per_thread) totals->add(c);
    return image;
  }
};

}  // namespace orc
