/* gdpt.h — C ABI of the B200 gradient-domain path tracer (libgdpt.so).
 *
 * Drop-in boundary for the `gradpath` arm of lajolla's
 *     Image3 render(const Scene &scene)            /root/reference/src/render.h:9
 *                                                  /root/reference/src/render.cpp:387-388
 * i.e. for gradient_path_render (render.cpp:257-370), grad_path_tracing
 * (path_tracing.h:354-1050), the Embree scene build + rtcIntersect1 it sits on
 * (scene.cpp:19-34, intersection.cpp:7-65) and fourierSolve (render.cpp:172-254).
 * The reference-side binding is shown in INTEGRATION.md.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a
 * negative error code (message via gdpt_last_error()); nothing throws across the ABI; the
 * caller owns every host buffer; a context owns all device memory it allocates.  Images
 * are interleaved-RGB doubles, row-major, y down, W*H*3 — exactly the layout of the
 * c / cx / cy / out arrays at render.cpp:336-353.  There is no CPU fallback: every entry
 * point that computes fails with GDPT_ERR_CUDA when no CUDA device is usable.
 */
#ifndef GDPT_H
#define GDPT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GDPT_OK 0
#define GDPT_ERR_ARG (-1)
#define GDPT_ERR_CUDA (-2)
#define GDPT_ERR_UNSUPPORTED (-3)

#define GDPT_PRECISION_F64 0 /* arithmetic in double like the reference (Real = double, lajolla.h:23) */
#define GDPT_PRECISION_F32 1 /* float shading/traversal, double accumulation and Poisson */

#define GDPT_OFFSETS_REFERENCE 0 /* grad_path_tracing as shipped (path_tracing.h:354-1050): the drop-in default */
#define GDPT_OFFSETS_SHIFT 1     /* shift-mapped offsets, the algorithm of small_gdpt.cpp:153-352 on lajolla scenes: random-number
                                  * replay, reconnection at the first diffuse-diffuse vertex pair, MIS-weighted gradients, y offsets
                                  * paired consistently (opt-in; gradpath integrator only) */

typedef struct gdpt_ctx gdpt_ctx;

/* One image of a TexturePool mip chain, level 0 only (src/texture.h:13-19, src/mipmap.h:28-44;
 * the mip levels are rebuilt by gdpt_create).  Either `data` (w*h*channels doubles) or
 * `palette` + `index` (w*h*channels bytes indexing palette) is set. */
typedef struct {
    int32_t width, height, channels;
    const double *data;
    const double *palette; int32_t palette_size;
    const uint8_t *index;
} gdpt_image;

/* POD mirror of the fields of lajolla's Scene that gradpath reads (src/scene.h:43-83). */
typedef struct {
    /* Camera (src/camera.h:10-24): row-major 4x4 */
    double sample_to_cam[16], cam_to_world[16];
    int32_t width, height;
    int32_t filter_kind;      /* 0 Box, 1 Tent, 2 Gaussian (src/filter.h:31-43) */
    double filter_param;      /* Box/Tent width, Gaussian stddev */
    /* Shapes (src/shape.h:26-53); shape id == index == Embree geomID */
    int32_t n_shapes;
    const int32_t *shapes;        /* [n][6]: kind(0 sphere,1 mesh), material_id, area_light_id, n_verts, n_tris, flags(1 normals|2 uvs) */
    const int64_t *shape_offsets; /* [n][4]: first vertex, first triangle, first normal, first uv of the mesh in the pools below */
    const double *spheres;        /* [n][4]: centre xyz, radius (zeros for meshes) */
    int64_t n_positions, n_triangles, n_normals, n_uvs;
    const double *positions;      /* [n_positions][3] */
    const int32_t *indices;       /* [n_triangles][3], mesh-local vertex indices */
    const double *normals;        /* [n_normals][3] */
    const double *uvs;            /* [n_uvs][2] */
    /* Materials (src/material.h:12-110) */
    int32_t n_materials;
    const double *materials;      /* [n][14]: type(0..8 in variant order), eta, 12 texture ids (-1 unused) */
    int32_t n_textures;
    const double *textures;       /* [n][14]: kind(0 const,1 image,2 checker), is_spectrum, v0[3], v1[3], image id, uscale, vscale, uoffset, voffset, pad */
    int32_t n_images3, n_images1;
    const gdpt_image *images3, *images1;
    /* Lights (src/light.h:15-34) */
    int32_t n_lights;
    const double *lights;         /* [n][5]: area light: 0, shape_id, intensity rgb;  environment map (src/light.h:23-30; ignored by
                                   * gradpath, used by path): 1, texture id of `values` (an image texture; -1: not provided), scale, 0, 0 */
    /* RenderOptions (src/scene.h:25-32) */
    int32_t max_depth, rr_depth, samples_per_pixel;
    /* Envmap::to_world / to_local, row-major 4x4 each; NULL when the scene has no environment map (or does not provide it:
     * the path integrator then refuses the scene) */
    const double *envmap_to_world, *envmap_to_local;
} gdpt_scene_desc;

typedef struct {
    int32_t spp;              /* samples per pixel (the estimator divides by this, render.cpp:313-317) */
    int32_t sample0;          /* index of the first sample: RNG stream = (sample0+s)*W*H + y*W + x */
    int32_t max_depth;        /* -2: keep the scene's value */
    int32_t rr_depth;         /* -2: keep the scene's value */
    uint64_t seed;            /* 0: the PCG default 0x31e241f862a1fb5e (src/pcg.h:33) */
    double data_cost;         /* Poisson alpha; <=0: 0.04 (render.cpp:353) */
    int32_t precision;        /* GDPT_PRECISION_* */
    int32_t y0, y1;           /* row band [y0,y1) to render; y1 <= 0: whole image */
    int32_t wave_samples;     /* samples per pixel per wave; 0: auto */
    int32_t offset_mode;      /* GDPT_OFFSETS_*; 0 = the reference's semantics */
    int32_t strip_count;      /* > 1: rows [y0,y1) are cut into 4-row strips and this call renders the strips whose index is */
    int32_t strip_phase;      /*      strip_phase (mod strip_count) — interleaved sharing of one film between GPUs (the reference's
                               *      dynamic tile queue, parallel.cpp:183-237, balances load the same way); <= 1: every row */
} gdpt_render_params;

/* Host output images; any pointer may be NULL.  primal/dx/dy/poisson = c/cx/cy/out of render.cpp:336-353. */
typedef struct { double *primal, *dx, *dy, *poisson; } gdpt_images;

typedef struct {
    double ms_total, ms_trace, ms_shade, ms_poisson, ms_copy;
    int64_t kernel_launches, rays_traced, paths, bounce_iterations;
    int64_t tail_paths;       /* paths finished by the single-launch tail kernel (their rays are not in rays_traced) */
} gdpt_stats;

const char *gdpt_last_error(void);
int gdpt_device_count(void);

#define GDPT_CREATE_NO_BVH 1 /* gdpt_create_ex: everything but the BVH build */
#define GDPT_CREATE_BVH_HOST 2   /* build the tree with the host's binned-SAH builder + triangle pre-splitting (best tree, ~45 ms for Sponza) */
#define GDPT_CREATE_BVH_DEVICE 4 /* build it on the device (PLOC, ~9 ms, 5-6 % slower traversal); neither flag: chosen from the planned work
                                  * width * height * samples_per_pixel (device build below 1e8 pixel-samples) */

/* Upload the scene to `device`, rebuild mip chains, build the BVH (replaces Scene::Scene's
 * Embree registration/commit, scene.cpp:19-34). */
int gdpt_create(const gdpt_scene_desc *scene, int device, gdpt_ctx **out);
/* gdpt_create with flags.  GDPT_CREATE_NO_BVH: the context gets everything but the tree and refuses to render until a BVH
 * built for the SAME scene is handed over with gdpt_bvh_adopt.  With one process per GPU the tree is then built once and
 * broadcast over NCCL (multi.DistributedRenderer) instead of N times on the same host cores (the reference builds one
 * Embree scene per process, scene.cpp:19-34; there is one process there). */
int gdpt_create_ex(const gdpt_scene_desc *scene, int device, int flags, gdpt_ctx **out);
/* The BVH of a context: node count (128 B each), primitive-record count (48 B each), depth, and the device arrays. */
int gdpt_bvh_info(gdpt_ctx *ctx, int64_t *n_nodes, int64_t *n_prims, int32_t *depth, void **d_nodes, void **d_prims);
/* Copy a BVH (device or host pointers) into a context created with GDPT_CREATE_NO_BVH. */
int gdpt_bvh_adopt(gdpt_ctx *ctx, const void *d_nodes, int64_t n_nodes, const void *d_prims, int64_t n_prims, int32_t depth);
void gdpt_destroy(gdpt_ctx *ctx);
/* Page-locked host memory for the caller's image buffers (cudaMallocHost / cudaFreeHost): gdpt_render copies its four images
 * into pinned buffers at PCIe speed and into pageable ones (what `new[]` at render.cpp:336-338 gives) several times slower. */
int gdpt_host_alloc(size_t bytes, void **out);
void gdpt_host_free(void *p);
/* scene bounds radius / centre as scene.cpp:30-34 derives them (for parity checks) */
int gdpt_scene_bounds(gdpt_ctx *ctx, double *radius_center4);

/* render(): gradient_path_render end to end with HOST outputs (accumulate, merge, Poisson, copy back). */
int gdpt_render(gdpt_ctx *ctx, const gdpt_render_params *params, gdpt_images *out);
/* Same accumulation into caller-provided DEVICE images (W*H*3 doubles each, already zeroed or
 * holding earlier passes): the five buffers of render.cpp:264-268.  No merge, no solve, no copy. */
int gdpt_render_accumulate_device(gdpt_ctx *ctx, const gdpt_render_params *params,
                                  double *d_img, double *d_cx0, double *d_cy0, double *d_cx1, double *d_cy1);
/* render.cpp:340-350 on device buffers */
int gdpt_merge_device(gdpt_ctx *ctx, int w, int h, const double *d_cx0, const double *d_cx1, const double *d_cy0,
                      const double *d_cy1, double *d_cx, double *d_cy);
/* fourierSolve (render.cpp:172-254): host buffers / device buffers.  `device` selects the GPU for the host form. */
int gdpt_poisson(int device, int w, int h, const double *u, const double *gx, const double *gy, double alpha, double *out);
int gdpt_poisson_device(gdpt_ctx *ctx, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  gdpt_poisson_device with every work
 * buffer in a fixed file-scope __device__ array (4.03 MiB, held by every GPU that loads the library until exit): the call
 * makes no cudaMalloc / cudaFree and needs no context.  Device pointers in and out (W*H*3 doubles each), launched on
 * `cuda_stream` (a cudaStream_t; NULL = the legacy default stream), asynchronous.  Images larger than
 * GDPT_NOALLOC_MAX_DIM in either dimension are refused with GDPT_ERR_ARG.  All calls share the array: calls that overlap on
 * the GPU race — serialise them.  The first call on a device resolves the array's address (cudaGetSymbolAddress); make it
 * outside graph capture. */
#define GDPT_NOALLOC_MAX_DIM 256
int gdpt_poisson_device_noalloc(int device, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha,
                                double *d_out, void *cuda_stream);

/* ---- distributed screened-Poisson solve (BASELINE configs[4]: 3840x2160 on 8 GPUs) ----
 * fourierSolve (render.cpp:172-254) over the ranks of one box.  Rank r owns the row band [y0,y1) of the
 * image and the column band [r*band_w, (r+1)*band_w) (the last band runs to w).  The three stages below
 * are the local work; between them the caller runs the two all-to-all transposes (NCCL) and, before the
 * first stage, the 3-double all-reduce of the DC sums (render.cpp:199-211):
 *
 *   halo:     cy1 row y0-1 from rank-1, cy0 row y1 from rank+1 (into the full-size images), then
 *             gdpt_merge_rows_device(rows [y0, min(y1+1,h)))                                  render.cpp:340-350
 *   stage 1:  gdpt_poisson_dist_rows_forward  -> d_send [3*(y1-y0)*w]: for every rank s, contiguous,
 *             the chunk [y1-y0][3][w_s] it must receive; d_dc_partial[3] = this band's DC sums
 *   exchange: all-to-all (send split (y1-y0)*3*w_s to rank s; the chunks received from ranks 0..n-1,
 *             stacked, ARE the row-major [h][3*w_me] matrix of this rank's column band); all-reduce(dc)
 *   stage 2:  gdpt_poisson_dist_columns       -> d_send_back [h][3*w_me] (rows [y0_r,y1_r) go back to rank r)
 *   exchange: all-to-all back (contiguous row slices; received chunk s = [y1-y0][3][w_s])
 *   stage 3:  gdpt_poisson_dist_rows_inverse  -> rows [y0,y1) of the full-size interleaved d_out
 * With n_ranks = 1 the three stages alone equal gdpt_poisson_device. */
typedef struct {
    int32_t w, h;             /* full image */
    int32_t n_ranks, rank;
    int32_t y0, y1;           /* this rank's row band */
    int32_t band_w;           /* column-band width; band_w*(n_ranks-1) < w */
} gdpt_dist_geom;
int gdpt_merge_rows_device(gdpt_ctx *ctx, int w, int h, int ya, int yb, const double *d_cx0, const double *d_cx1,
                           const double *d_cy0, const double *d_cy1, double *d_cx, double *d_cy);
int gdpt_poisson_dist_rows_forward(gdpt_ctx *ctx, const gdpt_dist_geom *geom, const double *d_u, const double *d_gx,
                                   const double *d_gy, double alpha, double *d_send, double *d_dc_partial);
int gdpt_poisson_dist_columns(gdpt_ctx *ctx, const gdpt_dist_geom *geom, double alpha, const double *d_dc,
                              const double *d_recv, double *d_send_back);
int gdpt_poisson_dist_rows_inverse(gdpt_ctx *ctx, const gdpt_dist_geom *geom, const double *d_recv_back, double *d_out);

/* ---- several GPUs of one box behind the ABI: gradient_path_render with the tile loop of render.cpp:271-331 spread over
 * n_devices GPUs inside this process (one host thread per device; scene replicated, BVH built once and copied device to
 * device; film shared as interleaved 4-row strips; the strips gathered on the first device over NVLink peer copies; merge and
 * solve there).  devices == NULL: 0 .. n_devices-1.  params: as gdpt_render, with y0 / y1 / strip_* left 0. ---- */
typedef struct gdpt_multi gdpt_multi;
int gdpt_multi_create(const gdpt_scene_desc *scene, int n_devices, const int *devices, gdpt_multi **out);
int gdpt_multi_render(gdpt_multi *m, const gdpt_render_params *params, gdpt_images *out);
int gdpt_multi_get_stats(gdpt_multi *m, gdpt_stats *out);      /* counters summed over the devices */
int gdpt_multi_device_count(gdpt_multi *m);
void gdpt_multi_destroy(gdpt_multi *m);

/* ---- L1 reconstruction.  No counterpart in the reference's code: fourierSolve (render.cpp:172-254,
 * small_gdpt.cpp:355-436) is the L2 screened-Poisson solve and the only reconstruction either program has; this is the robust
 * alternative of the paper its handout cites (Kettunen et al. 2015, section 5.2; handouts/final_project.tex:109-111) by
 * iteratively reweighted least squares:
 *   E(I) = alpha sum wU |I - u|^2 + sum_{x>=1} wX |I(x,y) - I(x-1,y) - gx(x,y)|^2 + sum_{y>=1} wY |I(x,y) - I(x,y-1) - gy(x,y)|^2,
 * wX, wY = 1 / (eps + L1 norm over the channels of the term's residual), eps = epsilon * mean|u|, each reweighting followed by a
 * Jacobi-preconditioned conjugate-gradient solve started from the current image.  u, gx, gy as for gdpt_poisson (the merged
 * gradients of render.cpp:340-350; column 0 of gx and row 0 of gy are not used).  info (may be NULL): [0] relative residual of
 * the last solve, [1] conjugate-gradient steps taken in all solves, [2] device milliseconds, [3] eps. ---- */
#define GDPT_L1_START_FOURIER 0   /* start from fourierSolve's image (gdpt_poisson_device) */
#define GDPT_L1_START_PRIMAL 1    /* start from u */
#define GDPT_L1_START_OUT 2       /* `out` holds the starting image */
typedef struct {
    int32_t irls_iterations;  /* reweightings; 0: the L2 problem (all weights 1) by conjugate gradients */
    int32_t cg_iterations;    /* conjugate-gradient steps per reweighting (upper bound: a solve stops at cg_tolerance) */
    int32_t start;            /* GDPT_L1_START_* */
    int32_t reweight_primal;  /* 0: wU = 1, only the gradient terms are reweighted (the image keeps the primal's mean exactly);
                               * 1: the primal term is reweighted too (the all-L1 energy of the paper; pulls towards medians) */
    double alpha;             /* <= 0: 0.04 (render.cpp:353) */
    double epsilon;           /* <= 0: 0.05 */
    double cg_tolerance;      /* relative residual |r| / |b|; <= 0: 1e-6 */
} gdpt_l1_params;
int gdpt_poisson_l1(int device, int w, int h, const double *u, const double *gx, const double *gy, const gdpt_l1_params *params,
                    double *out, double *info);
int gdpt_poisson_l1_device(gdpt_ctx *ctx, int w, int h, const double *d_u, const double *d_gx, const double *d_gy,
                           const gdpt_l1_params *params, double *d_out, double *info);

/* ---- the reference's second program, small_gdpt.cpp: main() (:439-563) on its built-in scene of eight
 * spheres (:60-69, camera :55-57) — generatePath (:153-197), shiftPath (:200-285: random-number replay, reconnection at the
 * first diffuse-diffuse edge), pathContrib / pathProb (:288-352), the balance-heuristic weights and the five running sums
 * (:470-527), then the merge (:530-539) and fourierSolve(.., 0.04, ..) (:541) by the product's own kernels.  It needs no
 * context and no scene description.  out: host images, height*width*3 doubles each, rows as the program stores them (top row
 * first); dx / dy are the SIGNED merged gradients (the program writes their absolute values to image_dx.pfm / image_dy.pfm,
 * :549-550); any pointer may be NULL.  rng_mode GDPT_SMALLGDPT_ROWS keeps the program's erand48 streams (one per film row,
 * :462): the images equal the compiled program's up to the device libm's last bit in sin / cos; GDPT_SMALLGDPT_PIXELS gives
 * every pixel a stream and a thread (the same estimator on other random numbers).  render_ms (may be NULL): device time of the
 * sampling kernel. ---- */
#define GDPT_SMALLGDPT_ROWS 0
#define GDPT_SMALLGDPT_PIXELS 1
int gdpt_smallgdpt_render(int device, int width, int height, int spp, int rng_mode, gdpt_images *out, double *render_ms);

int gdpt_get_stats(gdpt_ctx *ctx, gdpt_stats *out);
/* Scheduling / tuning knobs of a context (no effect on results; no counterpart in the reference, whose only knob is -t threads,
 * src/main.cpp:20-23).  "async_waves" 1|0: launch the waves of a render from the first wave's measured path counts without
 * per-bounce host read-backs (default 1); "wave_streams" 1..4: waves in flight on separate streams (default 3); "time_kernels"
 * 1|0: keep the host-synchronised loop and fill gdpt_stats.ms_trace / ms_shade with per-kernel event times (default 0: those two
 * fields stay 0); "deterministic" 1|0: gradpath accumulates through per-path
 * result records that are added to the images in sample order (bit-reproducible images, the summation order of render.cpp:303-327;
 * default 1) or with atomicAdd(double) straight into the images; "poisson_fft" 1|0: the O(N log N) form of the DCT-I in fourierSolve's
 * replacement or the dense-matrix form (default 1; films under 16 pixels a side always take the matrix form); "tail_paths" n; "sort_rays" 1|0; "trace_variant" v; "shade_variant" v; "tri_records" 1|0: a hit triangle's vertex data is read from one packed record built at creation (96 bytes / three 256-bit loads in f32, 192 / six in f64) or from the index / vertex pools (default 1; same values, same results). */
int gdpt_set_option(gdpt_ctx *ctx, const char *name, int value);
/* Run all of the context's work on the caller's CUDA stream (a cudaStream_t; NULL restores the context's own). */
int gdpt_set_stream(gdpt_ctx *ctx, void *cuda_stream);
/* Measurement aid: render `spp` samples with a counting traversal and return the mean number of BVH nodes
 * visited and primitives tested per ray (out[0], out[1]) and the rays traced (out[2]).  Not for timed runs. */
int gdpt_traversal_counts(gdpt_ctx *ctx, const gdpt_render_params *params, double *out3);
/* Measurement aid: L2 read bandwidth (GB/s) of `device`, streaming an `mbytes` buffer (< L2) `iters` times. */
int gdpt_measure_l2_read(int device, int mbytes, int iters, double *gbs);
int gdpt_synchronize(gdpt_ctx *ctx);

/* ---- parity / validation entry points (same device code as the renderer) ---- */
/* sample_primary (camera.cpp:23-47): screen_pos[n][2] -> rays[n][6] (origin, direction) */
int gdpt_sample_primary(gdpt_ctx *ctx, int precision, int n, const double *screen_pos, double *rays);
/* intersect (intersection.cpp:7-65): rays[n][8] = org, dir, tnear, tfar; rdiff[n][2] or NULL;
 * ids[n][3] = shape, primitive, material (-1 miss); vtx[n][24] = position3, geometric_normal3, frame x3 y3 n3,
 * st2, uv2, uv_screen_size, mean_curvature, ray_radius, t, pad */
int gdpt_intersect(gdpt_ctx *ctx, int precision, int n, const double *rays, const double *rdiff, int32_t *ids, double *vtx);
/* eval / pdf_sample_bsdf / sample_bsdf (material.cpp:90-119) for material `mat_id`:
 * vtx[n][15] = geometric_normal3, frame x3 y3 n3, uv2, uv_screen_size; dirs[n][6] = dir_in, dir_out; rnd[n][3];
 * out[n][10] = eval3, pdf, ok, sampled dir3, eta, roughness */
int gdpt_bsdf(gdpt_ctx *ctx, int precision, int mat_id, int n, const double *vtx, const double *dirs, const double *rnd, double *out);
/* grad_path_tracing (path_tracing.h:354-1050) for n explicit (x, y, rng stream) triples;
 * rec[n][23] = radiance3, contrib3, contribX0 3, contribX1 3, contribY0 3, contribY1 3, prob, wX0, wY0, wX1, wY1 */
int gdpt_grad_samples(gdpt_ctx *ctx, const gdpt_render_params *params, int n, const int32_t *xs, const int32_t *ys,
                      const uint64_t *streams, double *rec);
/* ---- the plain `path` integrator (render.cpp:74-117, path_tracing.h:13-348): NEE + MIS on the same kernels.
 * Area lights (spheres, meshes) and an image environment map (envmap.inl, TableDist2D); a scene whose environment map
 * was not provided in the description returns GDPT_ERR_UNSUPPORTED. ---- */
/* path_render end to end: out_image = W*H*3 doubles (host), the Image3 render() returns for Integrator::Path */
int gdpt_render_path(gdpt_ctx *ctx, const gdpt_render_params *params, double *out_image);
/* same accumulation into a caller-provided DEVICE image (already zeroed or holding earlier passes / other bands) */
int gdpt_render_path_accumulate_device(gdpt_ctx *ctx, const gdpt_render_params *params, double *d_img);
/* path_tracing for n explicit (x, y, rng stream) triples; rad[n][3] */
int gdpt_path_samples(gdpt_ctx *ctx, const gdpt_render_params *params, int n, const int32_t *xs, const int32_t *ys,
                      const uint64_t *streams, double *rad);
/* sample_light + sample_point_on_light + pdf_point_on_light + light_pmf (scene.cpp:61-67, light.cpp:57-71):
 * ref_points[n][3]; rnd[n][4] = light_w, uv0, uv1, shape_w; out[n][9] = light id, position3, normal3, pdf, pmf */
int gdpt_sample_light(gdpt_ctx *ctx, int precision, int n, const double *ref_points, const double *rnd, double *out);
/* occluded (intersection.cpp:66-85): rays[n][8] = org, dir, tnear, tfar; out[n] = 1 if anything is hit in (tnear, tfar] */
int gdpt_occluded(gdpt_ctx *ctx, int precision, int n, const double *rays, int32_t *out);
/* aux_render (render.cpp:15-72): the auxiliary integrators of render() — `integrator` is the Integrator enum value
 * (src/scene.h:14-23): 0 depth, 1 shadingNormal, 2 meanCurvature, 3 rayDifferential, 4 mipmapLevel; one ray through every
 * pixel centre; out_image = W*H*3 doubles (host).  Cheap always-available validation channels. */
int gdpt_render_aux(gdpt_ctx *ctx, int integrator, int precision, double *out_image);
/* PCG32 (pcg.h): first n doubles (and raw 32-bit outputs, may be NULL) of a stream, generated on the device */
int gdpt_pcg_fill(int device, uint64_t stream, uint64_t seed, int n, double *out, uint32_t *raw);

#ifdef __cplusplus
}
#endif
#endif
