/* vpt_host.h — C entry points of the host library (libvpt_host.so) for callers
 * without a C++ toolchain (the Python tests / bench drive the renderer through
 * these with ctypes). They wrap the C++ facade of
 * volumetric-path-tracer_b200/host/vpt_pathtrace.h, which in turn reaches the
 * GPU only through include/vpt_cuda.h. Host-only queries (bvh, lights, state
 * seeds) run without a GPU; everything that renders needs one.
 *
 * Reference call sequence mirrored: run_offline
 * (apps/ypathtrace/ypathtrace.cpp:41-87): load_scene -> make_bvh -> make_lights
 * -> make_state -> pathtrace_samples xN -> get_render.
 */
#ifndef VPT_HOST_H
#define VPT_HOST_H

#include "vpt_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vpth_scene  vpth_scene;  /* scene_data + bvh_scene + pathtrace_lights */
typedef struct vpth_render vpth_render; /* pathtrace_params + pathtrace_state */

const char* vpth_last_error(void);

/* load_scene + make_bvh + make_lights (host only) */
int  vpth_load_scene(const char* path, vpth_scene** out);
void vpth_free_scene(vpth_scene* scene);
/* counts: cameras, instances, shapes, materials, textures, environments,
 * volumes, vol_instances, sdfs, lights, scene bvh nodes, total shape bvh nodes */
int  vpth_scene_counts(const vpth_scene* scene, int32_t counts[12]);

/* host-built structures, for comparison with the reference's */
int  vpth_bvh_size(const vpth_scene* scene, int shape, int32_t* num_nodes, int32_t* num_prims);
int  vpth_bvh_copy(const vpth_scene* scene, int shape, vpt_bvh_node* nodes, int32_t* prims);
int  vpth_light_info(const vpth_scene* scene, int light, int32_t ids[3], int32_t* cdf_size);
int  vpth_light_cdf(const vpth_scene* scene, int light, float* cdf);
/* make_lights on the device (vpt_cuda_upload_scene_device_lights): (re)uploads the scene with the
 * light list and CDFs built in HBM; renders created afterwards use it */
/* subdivision surfaces: read the control cages of an archive (tools/vpt_export --dump-subdivs /
 * --keep-subdivs), run tesselate_surfaces with the numeric stages on the device, read a shape back */
int  vpth_scene_load_subdivs(vpth_scene* scene, const char* path);
int  vpth_scene_tesselate(vpth_scene* scene);
int  vpth_shape_sizes(const vpth_scene* scene, int shape, int32_t sizes[6]); /* positions, normals,
                                                         texcoords, triangles, quads, colors */
int  vpth_shape_copy(const vpth_scene* scene, int shape, float* positions, float* normals,
         float* texcoords, int32_t* triangles, int32_t* quads); /* null: skip */
/* rebuild the shape BVHs on the device (vpt_cuda_build_shape_bvh); `total` sums the per-shape stats */
int  vpth_scene_use_device_bvh(vpth_scene* scene, vpt_bvh_build_stats* total);
int  vpth_scene_use_device_lights(vpth_scene* scene, vpt_lights_stats* stats);
/* light `light` as it lies on the device (light < 0: *num_elements = number of lights) */
int  vpth_device_light(vpth_scene* scene, int light, int32_t ids[3], float* cdf, size_t capacity,
                       int32_t* num_elements);

/* make_state; rngs_out (optional) receives the initial per-pixel generators */
int  vpth_render_create(vpth_scene* scene, const vpt_params* params, vpth_render** out);
void vpth_render_free(vpth_render* render);
int  vpth_render_size(const vpth_render* render, int32_t* width, int32_t* height);
int  vpth_render_get_rngs(vpth_render* render, uint64_t* rngs); /* [w*h][2], synced from the device */
/* back to sample 0 with the make_state generators, advanced by rng_offset draws per pixel */
int  vpth_render_reset(vpth_render* render, uint64_t rng_offset);
/* page-lock the render's host mirrors now (otherwise: after its third host transfer). Worth it for
 * loops that move the generators / the image every step; a render that downloads once skips it */
int  vpth_render_pin(vpth_render* render);

/* pathtrace_samples with a batch of nsamples; blocking */
int  vpth_render_samples(vpth_render* render, int nsamples);
int  vpth_render_done_samples(const vpth_render* render);
/* normalized != 0: get_render (sum / samples); else the raw sums */
int  vpth_render_get_image(vpth_render* render, float* rgba, int normalized);
/* device-side output stage: 8-bit sRGB RGBA of the current render, [height][width][4] */
/* the un-normalised accumulation image WITHOUT a copy: brings pathtrace_state::image (the host mirror,
 * page-locked once pinned) up to date and returns its address, [h][w][4] floats, valid until the next
 * call on this render. What a reference caller gets by reading state.image */
int  vpth_render_image_view(vpth_render* render, const float** rgba);
int  vpth_render_get_srgb8(vpth_render* render, uint8_t* rgba);
/* get_render + tonemap(exposure, filmic) + 8 bit on the device (vpt_cuda_download_tonemapped8) */
int  vpth_render_get_tonemapped8(vpth_render* render, float exposure, int filmic, uint8_t* rgba);
/* render on `gpus` GPUs of this box from the next vpth_render_samples call on (pathtrace_params.gpus
 * of host/vpt_pathtrace.h: sample-range sharding, the sum happens inside get_image / get_srgb8) */
int  vpth_render_set_gpus(vpth_render* render, int gpus);
int  vpth_render_stats(const vpth_render* render, vpt_trace_stats* stats);
/* device pointer of the accumulation image (for a collective), or null */
void* vpth_render_device_image(vpth_render* render);
/* copy the device accumulation image into / from a caller-owned device buffer */
int  vpth_render_export_image(vpth_render* render, void* device_dst);
int  vpth_render_import_image(vpth_render* render, const void* device_src, int total_samples);

/* the vpt_scene handle of include/vpt_cuda.h behind a host scene (uploads it on first use), for
 * callers / tests that talk to libvpt_cuda.so directly; owned by `scene`; null on failure */
vpt_scene* vpth_scene_device_handle(vpth_scene* scene);
int  vpth_intersect(vpth_scene* scene, const float* rays, int n, int instance, int32_t* ids,
         float* uvd);
int  vpth_spheretrace(vpth_scene* scene, const float* rays, int n, int maxiter, int32_t* ids,
         float* dn);

#ifdef __cplusplus
}
#endif
#endif /* VPT_HOST_H */
