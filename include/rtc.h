/* rtc.h — C-ABI drop-in boundary of the B200 render backend (librtc_cuda.so).
 *
 * The reference (tlinford/raytracer-challenge-rs) has no FFI or plugin interface; the seam this
 * replaces is the Rust method pair
 *     Camera::render(&mut self, world: &World) -> Canvas              raytracer/src/camera.rs:133
 *     Camera::render_multithreaded(Arc<Camera>, Arc<World>) -> Canvas raytracer/src/camera.rs:150
 * called from scene-parser/src/lib.rs:278 and raytracer/src/bin/{first_scene,teapot,glass_sphere,
 * glass_sphere2,hexagon}.rs. A Rust `raytracer-cuda` crate binds exactly the entry points below
 * (see INTEGRATION.md for the extern "C" block and build.rs).
 *
 * Conventions: plain C structs of double / int32_t / uint64_t and pointers to caller-owned arrays;
 * no C++ types, no exceptions across the boundary. Every function returns 0 on success or a
 * negative rtc_status; rtc_last_error() gives the message for the calling thread. The caller owns
 * all host buffers; an rtc_scene owns its device memory. One render per scene handle at a time;
 * distinct handles are independent. There is NO CPU fallback: without a CUDA device every compute
 * entry point fails with RTC_ERR_CUDA.
 *
 * All arithmetic is IEEE f64 without FMA contraction, in the operation order of the reference.
 */
#ifndef RTC_H
#define RTC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTC_ABI_VERSION 1

/* Only the entry points below are exported; everything else in the library has hidden visibility. */
#if defined(__GNUC__)
#define RTC_API __attribute__((visibility("default")))
#else
#define RTC_API
#endif

typedef enum rtc_status {
  RTC_OK = 0,
  RTC_ERR_INVALID = -1,   /* malformed description / argument */
  RTC_ERR_CUDA = -2,      /* CUDA runtime error or no device */
  RTC_ERR_OVERFLOW = -3,  /* a bounded device list (CSG hits, containers) overflowed; image invalid */
  RTC_ERR_NAN = -4,       /* a NaN t reached the intersection sort (the reference panics there,
                             geometry/intersection.rs:113) */
  RTC_ERR_UNSUPPORTED = -5
} rtc_status;

/* Shape kinds — one per `impl Shape` in raytracer/src/geometry/shape/ (mod.rs:12-22). */
typedef enum rtc_kind {
  RTC_SPHERE = 0,          /* sphere.rs */
  RTC_PLANE = 1,           /* plane.rs */
  RTC_CUBE = 2,            /* cube.rs */
  RTC_CYLINDER = 3,        /* cylinder.rs */
  RTC_CONE = 4,            /* cone.rs */
  RTC_TRIANGLE = 5,        /* triangle.rs */
  RTC_SMOOTH_TRIANGLE = 6, /* smooth_triangle.rs */
  RTC_GROUP = 7,           /* group.rs */
  RTC_CSG = 8              /* csg.rs */
} rtc_kind;

typedef enum rtc_pattern_kind { /* pattern/mod.rs:53-60 */
  RTC_PAT_TEST = 0, RTC_PAT_STRIPE = 1, RTC_PAT_GRADIENT = 2, RTC_PAT_RING = 3, RTC_PAT_CHECKERS = 4
} rtc_pattern_kind;

typedef enum rtc_csg_op { /* csg.rs:15-19 */
  RTC_CSG_UNION = 0, RTC_CSG_INTERSECTION = 1, RTC_CSG_DIFFERENCE = 2
} rtc_csg_op;

#define RTC_FLAG_SHADOW 1u /* Shape::has_shadow(), geometry/mod.rs:101 */
#define RTC_FLAG_CLOSED 2u /* Cylinder/Cone.closed */

/* One node per shape, in PRE-ORDER over World.objects (world.rs:19): a Group is followed by its
 * children in order (group.rs:17), a Csg by its left subtree then its right subtree (csg.rs:47-48).
 * `skip` is the index of the first node after this node's subtree (leaf: own index + 1), which
 * makes the reference's depth-first, in-order traversal a stackless loop. */
typedef struct rtc_node {
  int32_t kind;     /* rtc_kind */
  int32_t skip;
  int32_t xform;    /* leaf, CSG: index into xforms (the shape's own inverse); GROUP: -1 */
  int32_t material; /* leaf: index into materials; GROUP/CSG: -1 */
  int32_t aux;      /* GROUP/CSG: index into boxes; TRIANGLE/SMOOTH_TRIANGLE: index into triangles;
                       CYLINDER/CONE: index into limits; else -1 */
  uint32_t flags;   /* RTC_FLAG_* ; bits 8..9: rtc_csg_op for CSG */
  int32_t right;    /* CSG: node index of the right subtree's root; else -1 */
  int32_t eq_class; /* value-equality class of the shape (`impl PartialEq for dyn Shape`,
                       geometry/mod.rs:110-114), used by the n1/n2 containers walk
                       (geometry/intersection.rs:63-90) */
} rtc_node;

typedef struct rtc_material { /* material.rs:11-21 */
  double color[3];
  double ambient, diffuse, specular, shininess, reflective, transparency, refractive_index;
  int32_t pattern; /* index into patterns, or -1 */
  int32_t _pad;
} rtc_material;

typedef struct rtc_pattern { /* pattern/mod.rs:17-21 */
  int32_t kind;    /* rtc_pattern_kind */
  int32_t _pad;
  double a[3], b[3];
  double inverse[12]; /* rows 0..2 of Pattern.transform_inverse, row-major */
} rtc_pattern;

typedef struct rtc_light { /* light.rs:4-8 */
  double position[3];
  double intensity[3];
} rtc_light;

/* Flattened World. All pointers are host memory, read during rtc_scene_create only. */
typedef struct rtc_scene_desc {
  uint32_t abi_version;      /* RTC_ABI_VERSION */
  uint32_t n_nodes, n_xforms, n_boxes, n_limits, n_triangles, n_materials, n_patterns, n_lights;
  const rtc_node* nodes;
  const double* xforms;      /* n_xforms x 12: rows 0..2 of BaseShape.transform_inverse (geometry/mod.rs:15).
                                The inverse-transpose used for normals (mod.rs:16,54) is read from the
                                same 3x3 block transposed — bit-identical to Matrix::transpose(). */
  const double* boxes;       /* n_boxes x 6: BoundingBox min xyz, max xyz (bounding_box.rs:5-9) */
  const double* limits;      /* n_limits x 2: minimum, maximum (cylinder.rs:15-16, cone.rs:15-16) */
  const double* tri_p1e1e2;  /* n_triangles x 9: p1, e1, e2 (triangle.rs:14-21) */
  const double* tri_normals; /* n_triangles x 9: flat → normal,0,0 ; smooth → n1,n2,n3 */
  const rtc_material* materials;
  const rtc_pattern* patterns;
  const rtc_light* lights;   /* World.lights in insertion order (world.rs:20) */
} rtc_scene_desc;

typedef struct rtc_camera { /* camera.rs:20-29 */
  uint32_t hsize, vsize;
  double pixel_size, half_width, half_height;
  double inverse[16]; /* Camera.transform_inverse, row-major 4x4 (rows 0..2 are used) */
} rtc_camera;

typedef struct rtc_opts {
  uint32_t aa_samples;  /* 1,2,4,8,16 — AASamples, camera.rs:227-233; 1 = Camera::render */
  uint32_t max_depth;   /* MAX_RECURSION_DEPTH = 5, world.rs:16 */
  uint32_t row_begin, row_end; /* rows [row_begin,row_end) of the frame are rendered; 0,0 = all */
  uint32_t tile_rows;   /* multi-GPU: rows are grouped in tiles of this many rows ...            */
  uint32_t rank, nranks;/* ... and this call renders the tiles with tile_index % nranks == rank.  */
                        /* nranks <= 1 renders every tile. Untouched pixels are left as they are. */
  uint32_t flags;       /* RTC_OPT_* */
} rtc_opts;

#define RTC_OPT_FAITHFUL_TRAVERSAL 1u /* disable the t-range box culling; visit exactly what
                                         Group::intersect (group.rs:49-58) visits */
#define RTC_OPT_WAVEFRONT 4u          /* force the level-synchronous wavefront renderer (one work item per
                                         color_at call; aa_samples == 1 only) ...                          */
#define RTC_OPT_RECURSIVE 8u          /* ... or the one-thread-per-pixel kernel; neither: the library picks.
                                         Both produce the same bits. */
#define RTC_OPT_WAVE_STREAM 16u       /* with the wavefront: ONE barrier-free persistent launch (rt_stream.cuh; experimental:
                                         measured slower than the passes at every share of the headline frame) instead of
                                         the 1 + 2 * max_depth level-synchronous passes (rt_wave.cuh); same bits */
#define RTC_OPT_AVERAGE 2u            /* render_multithreaded semantics even for aa_samples == 1:
                                         Color::average = (0 + c) * (1.0/1) (camera.rs:186) */

typedef struct rtc_stats {
  uint64_t rays_primary, rays_shadow, rays_reflect, rays_refract;
  uint64_t overflow_count; /* pixels whose bounded lists overflowed (→ RTC_ERR_OVERFLOW) */
  uint64_t nan_count;      /* pixels that met a NaN t (→ RTC_ERR_NAN) */
  double kernel_ms;        /* render kernel(s), CUDA events on the launch stream */
  double total_ms;         /* whole call, incl. copies, host clock */
  uint32_t launches;       /* kernels launched by this call */
  uint32_t flags;          /* RTC_STAT_* */
} rtc_stats;

#define RTC_STAT_WAVEFRONT 1u     /* the frame was rendered by a wavefront renderer */
#define RTC_STAT_WAVE_FALLBACK 2u /* ... whose queues overflowed: rtc_scene_finish (or the synchronous entry points) rendered
                                     the frame again with the recursive kernel. A caller that copied or pushed the canvas
                                     between rtc_render_device and rtc_scene_finish must do so again. */
#define RTC_STAT_WAVE_OVERFLOWED 4u /* some wavefront frame queued on this scene since the previous rtc_scene_finish overflowed
                                     its queues (only the LAST frame is rendered again; earlier ones were left incomplete) */

typedef struct rtc_scene rtc_scene;

RTC_API const char* rtc_last_error(void);
RTC_API int rtc_abi_version(void);
RTC_API int rtc_device_count(void);

/* Upload a flattened world to `device`. Replaces building `World` (world.rs:24,83-93). */
RTC_API int rtc_scene_create(const rtc_scene_desc* desc, int device, rtc_scene** out);
RTC_API void rtc_scene_destroy(rtc_scene* scene);

/* Camera::render / render_multithreaded (camera.rs:133,150): render into a HOST canvas of
 * hsize*vsize*3 doubles, row-major y*hsize+x like Canvas (canvas.rs:44-48). The device→host copy
 * of the rendered rows is part of the call. */
RTC_API int rtc_render(rtc_scene* scene, const rtc_camera* camera, const rtc_opts* opts, double* rgb_host,
               rtc_stats* stats);

/* Same, into a DEVICE canvas (same layout; may be a peer-mapped pointer on another GPU — the
 * multi-GPU gather is the kernel's own stores over NVLink). Asynchronous on `stream`
 * (a cudaStream_t, NULL = default stream) unless `stats` is non-NULL, in which case the call
 * synchronises the stream to fill it. */
RTC_API int rtc_render_device(rtc_scene* scene, const rtc_camera* camera, const rtc_opts* opts, double* rgb_dev,
                      void* stream, rtc_stats* stats);

/* Wait for the work queued on `stream` by rtc_render_device(…, stats = NULL), check the device
 * error flags and fill `stats` (ray counts; kernel_ms of the last launch). */
RTC_API int rtc_scene_finish(rtc_scene* scene, void* stream, rtc_stats* stats);

/* Output stage (image/ppm.rs:73-75, png.rs:29-31): rgb8 = (v*255).round() as u8, saturating,
 * NaN→0, for n doubles already on the device → n bytes on the device. */
RTC_API int rtc_quantize_device(int device, const double* rgb_dev, size_t n, uint8_t* rgb8_dev, void* stream);

/* Camera::render for a caller that wants an image file (scene-parser/src/lib.rs:278-281 →
 * image/png.rs:13-31, image/ppm.rs:24-75): render into a device canvas, quantise there, and copy
 * 3 bytes per pixel to `rgb8_host` (hsize*vsize*3 bytes, rows row_begin..row_end written) instead of
 * 24. nranks must be 1. */
RTC_API int rtc_render_rgb8(rtc_scene* scene, const rtc_camera* camera, const rtc_opts* opts, uint8_t* rgb8_host,
                    rtc_stats* stats);

/* Debug / known-answer entry: World::color_at(ray, remaining) (world.rs:70) for a batch of HOST
 * rays. rgb_out (3n) and t_out (n; hit t or -1) may each be NULL. */
RTC_API int rtc_trace_rays(rtc_scene* scene, size_t n, const double* origins, const double* directions,
                   uint32_t remaining, double* rgb_out, double* t_out);

/* No counterpart in the reference. rtc_trace_rays for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the
 * device staging of the batch is a file-scope static __device__ array (RTC_TRACE_STATIC_MAX rays x 10 doubles =
 * 320 KB per GPU that loads the library, held until exit), so the call allocates nothing — no cudaMalloc, cudaFree or
 * allocating library routine. n > RTC_TRACE_STATIC_MAX is rejected with RTC_ERR_INVALID. Results are bit-identical
 * to rtc_trace_rays. All calls share the one array per GPU: calls that overlap on a GPU (two host threads) race. */
#define RTC_TRACE_STATIC_MAX 4096
RTC_API int rtc_trace_rays_static(rtc_scene* scene, size_t n, const double* origins, const double* directions,
                                  uint32_t remaining, double* rgb_out, double* t_out);

/* Debug / known-answer entry: Intersection::prepare_computations (geometry/intersection.rs:53)
 * for one host ray. out[20] = hit(0/1), t, point[3], normalv[3], over_point[3], under_point[3],
 * reflectv[3], n1, n2, inside. */
RTC_API int rtc_probe_hit(rtc_scene* scene, const double* origin, const double* direction, double* out20);

/* Debug / known-answer entry: World::is_shadowed(point, light) (world.rs:95). Returns 0/1 or <0. */
RTC_API int rtc_is_shadowed(rtc_scene* scene, const double* point, uint32_t light_index);

/* Device canvases for the multi-GPU gather: rank 0 creates one, exports its CUDA IPC handle
 * (64 bytes), the other ranks open it and pass rtc_canvas_ptr() to rtc_render_device. */
typedef struct rtc_canvas rtc_canvas;
RTC_API int rtc_canvas_create(int device, size_t n_doubles, rtc_canvas** out);
RTC_API int rtc_canvas_export(rtc_canvas* canvas, unsigned char handle[64]);
RTC_API int rtc_canvas_open(int device, const unsigned char handle[64], size_t n_doubles, rtc_canvas** out);
RTC_API double* rtc_canvas_ptr(rtc_canvas* canvas);
RTC_API int rtc_canvas_download(rtc_canvas* canvas, double* host, size_t offset_doubles, size_t n_doubles);
/* Copy the tiles `rank` of `nranks` owns (tile t = rows [t*tile_rows, (t+1)*tile_rows), t % nranks == rank) from a
 * local device canvas of the same shape into `canvas` (usually rank 0's, opened over IPC): one strided DMA over NVLink,
 * queued on `stream`. The alternative to letting the kernels store into the peer canvas themselves: the wavefront
 * renderer completes pixels in 24-byte pieces from many passes, which a peer mapping pays for piece by piece. */
RTC_API int rtc_canvas_push_tiles(rtc_canvas* canvas, const double* local_dev, uint32_t hsize, uint32_t vsize,
                                  uint32_t tile_rows, uint32_t rank, uint32_t nranks, void* stream);
RTC_API void rtc_canvas_destroy(rtc_canvas* canvas);

/* The HOST canvas of a multi-GPU render. What Camera::render_multithreaded hands back is the host Canvas its row
 * blocks were gathered into (camera.rs:195-207); here that is ONE POSIX shared-memory segment (`name`, as for
 * shm_open: "/something"), mapped by every rank's process and page-locked + mapped for that process's GPU
 * (cudaHostRegisterPortable | cudaHostRegisterMapped). Every rank then calls
 *     rtc_render(scene, camera, opts{rank, nranks, tile_rows}, rtc_host_canvas_ptr(canvas), stats)
 * and ITS tiles reach the canvas over ITS OWN PCIe link (kernel stores straight into the mapping, or one strided DMA
 * of the rank's tiles for the wavefront renderer) — N links instead of rank 0's one. `create` != 0: create and size the
 * segment (rank 0; removed again by its rtc_host_canvas_destroy); 0: open an existing one. Nothing is touched or
 * page-locked yet (see rtc_host_canvas_place). */
typedef struct rtc_host_canvas rtc_host_canvas;
RTC_API int rtc_host_canvas_create(const char* name, size_t n_doubles, int device, int create, rtc_host_canvas** out);
/* Optional, between create/open and the first rtc_host_canvas_ptr, on every rank, followed by a barrier of the caller's:
 * put the pages of THIS rank's tiles on the NUMA node its GPU hangs off (mbind + first touch), so that eight GPUs do not
 * all write across one socket's memory controllers and the inter-socket link (measured at 8 GPUs: 14 GB/s per rank into
 * a canvas that lives on one node, against 50 GB/s at 4). Best effort: without NUMA (or without the right) it only
 * zero-fills the tiles. The segment is page-locked for the GPU at the first rtc_host_canvas_ptr. */
RTC_API int rtc_host_canvas_place(rtc_host_canvas* canvas, uint32_t hsize, uint32_t vsize, uint32_t tile_rows, uint32_t rank,
                                  uint32_t nranks);
RTC_API double* rtc_host_canvas_ptr(rtc_host_canvas* canvas);
/* Frame hand-over through the segment itself (the mpsc channel of camera.rs:195-207): a rank that has rendered its
 * tiles of frame `frame` (1, 2, 3 …; rtc_render has returned, so they are in host memory) arrives; wait returns
 * once ranks 0 … nranks-1 have all arrived at `frame` (or RTC_ERR_INVALID after timeout_s seconds). A few
 * microseconds instead of a collective. */
RTC_API int rtc_host_canvas_arrive(rtc_host_canvas* canvas, uint32_t rank, uint32_t frame);
RTC_API int rtc_host_canvas_wait(rtc_host_canvas* canvas, uint32_t nranks, uint32_t frame, double timeout_s);
RTC_API void rtc_host_canvas_destroy(rtc_host_canvas* canvas);

/* Debug / known-answer entry: Vector::normalize (vector.rs:25-28) of n host vectors (xyz triples) on
 * the device, through the same normalize() the kernels use (its divisions share one reciprocal). */
RTC_API int rtc_debug_normalize(int device, size_t n, const double* xyz_host, double* out_host);

/* Roofline denominator: measured FP64 throughput of `device` in GFLOP/s.
 * fused=1: dependent-chain-free DFMA stream (2 flops each); fused=0: alternating DADD/DMUL
 * (1 flop each) — the attainable roof for this library, which is built with --fmad=false. */
RTC_API int rtc_measure_fp64(int device, int fused, double* gflops_out);

#ifdef __cplusplus
}
#endif
#endif /* RTC_H */
