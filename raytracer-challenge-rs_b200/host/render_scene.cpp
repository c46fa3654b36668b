// render_scene <scene-file> <output-file> — scene-parser/src/bin/render_scene.rs on the CUDA backend.
// Options after the two paths: --extended (cylinder/cone/group/csg/obj `add:` kinds), --aa N (1|4|8|16),
// --device D. Output format by extension: .png, .ppm (P3 text as image/ppm.rs writes it), else binary P6.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/rtc.h"

extern "C" {
const char* rth_last_error();
int rth_render_scene_file(const char* scene_path, const char* out_path, uint32_t flags, uint32_t aa, int device,
                          rtc_stats* stats);
}

int main(int argc, char** argv) {
  if (argc < 3) {
    std::printf("usage: render_scene <scene-file> <output-file> [--extended] [--aa N] [--device D]\n");
    return 2;
  }
  uint32_t flags = 0, aa = 1;
  int device = 0;
  for (int i = 3; i < argc; i++) {
    if (!std::strcmp(argv[i], "--extended")) flags |= 1u;
    else if (!std::strcmp(argv[i], "--aa") && i + 1 < argc) aa = (uint32_t)std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--device") && i + 1 < argc) device = std::atoi(argv[++i]);
    else { std::fprintf(stderr, "unknown option %s\n", argv[i]); return 2; }
  }
  rtc_stats st;
  std::memset(&st, 0, sizeof st);
  int rc = rth_render_scene_file(argv[1], argv[2], flags, aa, device, &st);
  if (rc) {
    std::fprintf(stderr, "render_scene: %s\n", rth_last_error());
    return 1;
  }
  unsigned long long rays = st.rays_primary + st.rays_shadow + st.rays_reflect + st.rays_refract;
  std::printf("scene saved to %s (%llu rays, kernel %.3f ms, render %.3f ms)\n", argv[2], rays, st.kernel_ms, st.total_ms);
  return 0;
}
