/* vpt_cuda.h — the C ABI of the B200 path-tracing hot path.
 *
 * This is the drop-in boundary named by BASELINE.json:north_star: host code
 * (C++, or anything with an FFI) reaches the GPU only through these plain-C
 * entry points.  Pointers + sizes only; no C++ / torch types.  Every call
 * returns 0 on success or a negative vpt_status; vpt_cuda_last_error() gives a
 * human-readable message for the calling thread.  Nothing here has a CPU
 * fallback: without a CUDA device every compute call fails with
 * VPT_ERR_NO_DEVICE.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   vpt_cuda_upload_scene    <- the const scene_data& / bvh_scene& / pathtrace_lights&
 *                               arguments of pathtrace_samples
 *                               (libs/yocto_pathtrace/yocto_pathtrace.cpp:1052-1054);
 *                               inputs are the outputs of make_bvh (:955) and
 *                               make_lights (:983)
 *   vpt_cuda_upload_scene_device_lights
 *                            <- the same, with make_lights (:983-1049) run on the device
 *   vpt_cuda_create_state    <- make_state (yocto_pathtrace.cpp:960-980): image / rngs
 *   vpt_cuda_trace_samples   <- pathtrace_samples (yocto_pathtrace.cpp:1052-1092),
 *                               nsamples calls folded into one
 *   vpt_cuda_download_image  <- reading pathtrace_state::image for get_render (:1105-1116)
 *   vpt_cuda_intersect       <- intersect_bvh (libs/yocto/yocto_bvh.cpp:1097-1113)
 *   vpt_cuda_spheretrace     <- spheretrace + eval_sdf_normal
 *                               (yocto_pathtrace.cpp:289-307, libs/yocto/yocto_sdfs.cpp:53-89)
 */
#ifndef VPT_CUDA_H
#define VPT_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum vpt_status {
  VPT_OK               = 0,
  VPT_ERR_INVALID      = -1, /* bad argument / malformed scene description */
  VPT_ERR_UNSUPPORTED  = -2, /* point / line primitives, BVH deeper than the device stack */
  VPT_ERR_NO_DEVICE    = -3, /* no CUDA device: there is no CPU fallback */
  VPT_ERR_CUDA         = -4, /* a CUDA runtime call or kernel failed */
  VPT_ERR_OUT_OF_MEMORY = -5
} vpt_status;

/* shader enum: order is ABI (yocto_pathtrace.h:74-84) */
typedef enum vpt_shader {
  VPT_SHADER_VOLPATHTRACE    = 0,
  VPT_SHADER_PATHTRACE       = 1,
  VPT_SHADER_NAIVE           = 2,
  VPT_SHADER_EYELIGHT        = 3,
  VPT_SHADER_NORMAL          = 4,
  VPT_SHADER_TEXCOORD        = 5,
  VPT_SHADER_COLOR           = 6,
  VPT_SHADER_IMPLICIT        = 7,
  VPT_SHADER_IMPLICIT_NORMAL = 8,
  VPT_SHADER_COUNT           = 9
} vpt_shader;

/* material_type order (libs/yocto/yocto_scene.h:107-110) */
typedef enum vpt_material_type {
  VPT_MAT_MATTE = 0, VPT_MAT_GLOSSY, VPT_MAT_REFLECTIVE, VPT_MAT_TRANSPARENT,
  VPT_MAT_REFRACTIVE, VPT_MAT_SUBSURFACE, VPT_MAT_VOLUMETRIC, VPT_MAT_GLTFPBR
} vpt_material_type;

/* sdf_type order (libs/yocto/yocto_sdfs.h:23) */
typedef enum vpt_sdf_type {
  VPT_SDF_BBOX = 0, VPT_SDF_BOX, VPT_SDF_CAPPED_CONE, VPT_SDF_PLANE,
  VPT_SDF_SPHERE, VPT_SDF_TORUS
} vpt_sdf_type;

/* how the sample loop is scheduled on the device */
typedef enum vpt_pipeline {
  VPT_PIPELINE_AUTO      = 0, /* wavefront for pathtrace/volpathtrace/naive/implicit, megakernel for the debug shaders */
  VPT_PIPELINE_MEGA      = 1, /* one thread per pixel runs whole paths */
  VPT_PIPELINE_WAVEFRONT = 2  /* raygen -> extend -> shade -> light-pdf queues */
} vpt_pipeline;

/* bvh_node, byte-compatible with yocto::bvh_node (yocto_bvh.h:73-79): 32 bytes,
 * so bvh_data::nodes.data() can be passed as is. */
typedef struct vpt_bvh_node {
  float   bbox_min[3];
  float   bbox_max[3];
  int32_t start;
  int16_t num;
  int8_t  axis;
  uint8_t internal;
} vpt_bvh_node;

typedef struct vpt_bvh_desc {
  const vpt_bvh_node* nodes;
  int32_t             num_nodes;
  const int32_t*      primitives;
  int32_t             num_primitives;
} vpt_bvh_desc;

/* frames are yocto frame3f: x, y, z, o column vectors, 12 floats */
typedef struct vpt_camera_desc {
  float   frame[12];
  int32_t orthographic;
  float   lens, film, aspect, focus, aperture;
} vpt_camera_desc;

typedef struct vpt_instance_desc {
  float   frame[12];
  int32_t shape, material;
} vpt_instance_desc;

typedef struct vpt_shape_desc {
  const float*   positions;  int32_t num_positions;  /* [n][3] */
  const float*   normals;    int32_t num_normals;    /* [n][3] or null */
  const float*   texcoords;  int32_t num_texcoords;  /* [n][2] or null */
  const float*   colors;     int32_t num_colors;     /* [n][4] or null */
  const int32_t* triangles;  int32_t num_triangles;  /* [n][3] */
  const int32_t* quads;      int32_t num_quads;      /* [n][4] */
  int32_t        num_points, num_lines;              /* must be 0: unsupported */
} vpt_shape_desc;

typedef struct vpt_material_desc {
  int32_t type;
  float   emission[3], color[3];
  float   roughness, metallic, ior;
  float   scattering[3];
  float   scanisotropy, trdepth, opacity;
  int32_t emission_tex, color_tex, roughness_tex, scattering_tex, normal_tex;
} vpt_material_desc;

typedef struct vpt_texture_desc {
  int32_t        width, height, linear;
  const float*   pixelsf; /* [h][w][4] or null */
  const uint8_t* pixelsb; /* [h][w][4] or null */
} vpt_texture_desc;

typedef struct vpt_environment_desc {
  float   frame[12];
  float   emission[3];
  int32_t emission_tex;
} vpt_environment_desc;

typedef struct vpt_volume_desc {
  int32_t      whd[3];
  float        res;
  const float* voxels; /* x fastest (yocto_scene.h:211) */
} vpt_volume_desc;

typedef struct vpt_volume_instance_desc {
  float   frame[12];
  float   scalef;
  int32_t volume, material;
} vpt_volume_instance_desc;

/* analytic sdf with typed parameters (the reference keeps a std::function
 * closure, yocto_sceneio.cpp:3685-3728). params: bbox {thickness, w, h, d};
 * capped_cone {height, r1, r2}; sphere {radius}; torus {r1, r2}; box uses whd. */
typedef struct vpt_sdf_desc {
  float   frame[12];
  int32_t type, material;
  float   params[4];
  float   whd[3];
} vpt_sdf_desc;

/* pathtrace_light (yocto_pathtrace.h:106-111) */
typedef struct vpt_light_desc {
  int32_t      instance, environment, sdf;
  const float* elements_cdf;
  int32_t      num_elements;
} vpt_light_desc;

typedef struct vpt_scene_desc {
  int32_t num_cameras;       const vpt_camera_desc*          cameras;
  int32_t num_instances;     const vpt_instance_desc*        instances;
  int32_t num_shapes;        const vpt_shape_desc*           shapes;
  int32_t num_materials;     const vpt_material_desc*        materials;
  int32_t num_textures;      const vpt_texture_desc*         textures;
  int32_t num_environments;  const vpt_environment_desc*     environments;
  int32_t num_volumes;       const vpt_volume_desc*          volumes;
  int32_t num_vol_instances; const vpt_volume_instance_desc* vol_instances;
  int32_t num_sdfs;          const vpt_sdf_desc*             sdfs;
  int32_t num_lights;        const vpt_light_desc*           lights;
  vpt_bvh_desc               scene_bvh;   /* over instances */
  const vpt_bvh_desc*        shape_bvhs;  /* [num_shapes] */
} vpt_scene_desc;

/* pathtrace_params (yocto_pathtrace.h:87-99) as plain C, plus device options */
typedef struct vpt_params {
  int32_t camera;
  int32_t resolution;
  int32_t shader;   /* vpt_shader */
  int32_t samples;  /* total target; samples == 1 selects the no-jitter preview branch */
  int32_t bounces;
  int32_t noparallel; /* accepted, ignored on the GPU */
  int32_t pratio;
  float   exposure;
  int32_t filmic;
  int32_t noimplicit_mis;
  int32_t spheretrace_maxiter;
  int32_t pipeline; /* vpt_pipeline */
} vpt_params;

typedef struct vpt_scene vpt_scene; /* device-resident flattened scene */

/* vpt_cuda_upload_scene_device_lights: timing of the device-side light construction */
typedef struct vpt_lights_stats {
  int32_t  num_lights;
  uint64_t cdf_elements;  /* floats written */
  float    weights_ms;    /* parallel pass: element areas / texel weights */
  float    scan_ms;       /* running sums in the reference's (sequential) order */
} vpt_lights_stats;
typedef struct vpt_state vpt_state; /* device-resident accumulation image + rngs */

/* per-call device statistics, filled by vpt_cuda_trace_samples */
typedef struct vpt_trace_stats {
  float    device_ms;       /* CUDA-event time of all kernels of the last call */
  float    top_kernel_ms;   /* summed CUDA-event time of the SAMPLED launches of the dominant
                               kernel (k_wf_extend: one launch per counter poll; the whole
                               call for the megakernel) */
  int32_t  kernel_launches; /* kernels launched by the last call */
  int32_t  top_kernel_launches; /* how many launches top_kernel_ms covers */
  uint64_t paths;           /* camera paths completed by the last call */
  uint64_t rays_extend;     /* closest-hit scene traversals */
  uint64_t rays_lightpdf;   /* instance-targeted traversals */
  uint64_t sdf_evals;       /* scene sdf evaluations */
  uint64_t top_kernel_units; /* rays (paths for the megakernel) those sampled launches processed */
  uint64_t graph_launches;   /* batches (8 iterations = 40 of the kernel_launches) that the host
                                submitted as ONE CUDA graph launch instead of kernel by kernel */
} vpt_trace_stats;

int  vpt_cuda_device_count(void);
int  vpt_cuda_set_device(int device);
int  vpt_cuda_current_device(void); /* of the calling thread; -1 without a device */
const char* vpt_cuda_last_error(void);

/* page-lock caller-owned host buffers so uploads / downloads run at full PCIe rate */
int  vpt_cuda_host_register(void* ptr, size_t bytes);
int  vpt_cuda_host_unregister(void* ptr);

int  vpt_cuda_upload_scene(const vpt_scene_desc* desc, vpt_scene** out);
/* make_lights on the device (yocto_pathtrace.cpp:983-1049). desc->lights is
 * ignored: the light list is enumerated from the description in the reference's order (emissive
 * instances, environments, emissive sdfs) and every elements_cdf is built in HBM -- element areas
 * / texel weights (max(rgba) * sin(theta), glibc-exact sinf) in parallel, then the running sums
 * in the reference's sequential float order, so the CDFs equal make_lights' bit for bit.
 * `stats` may be null. */
int  vpt_cuda_upload_scene_device_lights(const vpt_scene_desc* desc, vpt_scene** out,
                                         vpt_lights_stats* stats);
/* Shape BVH built on the device: an OPTIONAL replacement of make_bvh(shape)
 * (yocto_bvh.cpp:526-571 -> build_bvh :447-507, split_middle :411-443). LBVH: primitive boxes and
 * 30-bit Morton codes of their centroids, cub radix sort, Karras' hierarchy, boxes bottom-up, subtrees
 * of <= 4 primitives folded into leaves (bvh_max_prims, :444). The tree comes back in the
 * reference's own node format (children adjacent at start / start + 1) with its primitive order,
 * sized nodes[<= 2 n + 1], primitives[n], and goes into vpt_bvh_desc like a host-built one.
 * It is NOT the reference's tree: closest distances agree, but of two primitives at exactly the
 * same distance another one may be reported (the reference keeps the later-tested one), so the
 * bit-exact image gates hold only with the host builder; the first-hit gate records the ties. */
typedef struct vpt_bvh_build_stats {
  int32_t primitives, nodes;
  float   keys_ms;       /* boxes, centroid bounds, Morton keys */
  float   sort_ms;       /* cub::DeviceRadixSort, 64-bit keys */
  float   hierarchy_ms;  /* hierarchy, box fit, leaf folding, renumbering */
} vpt_bvh_build_stats;
int  vpt_cuda_build_shape_bvh(const vpt_shape_desc* shape, vpt_bvh_node* nodes, size_t node_capacity,
         int32_t* primitives, size_t prim_capacity, int32_t* num_nodes,
         vpt_bvh_build_stats* stats);
/* Subdivision surfaces on the device: the three numeric stages of
 * tesselate_surface (yocto_pathtrace.cpp:1230-1273). Results equal the reference's bit for bit.
 *
 * vpt_cuda_catmullclark: `levels` rounds of tesselate_catmullclark (:1119-1228) on a quad mesh
 *   (a triangle is a quad with z == w). dim = 3 with lock_boundary = 0 for positions, dim = 2
 *   with lock_boundary = 1 for texture coordinates, as tesselate_surface calls it. Outputs are
 *   malloc'ed by the library (quads [n][4], verts [n][dim]); release with vpt_cuda_free_host.
 * vpt_cuda_vertex_normals: quads_normals (verts_per = 4) / triangles_normals (verts_per = 3)
 *   (yocto_shape.cpp:1478-1510): area-weighted face normals summed per vertex in face order.
 * vpt_cuda_displace: positions[i] += normals[i] * displacement * d(texcoords[i]) with d = mean
 *   of the displacement texture's rgb (bilinear, sRGB decoded), minus 0.5 for 8-bit textures
 *   (:1256-1263). `positions` is updated in place. */
int  vpt_cuda_catmullclark(const int32_t* quads, int nquads, const float* verts, int nverts, int dim,
         int lock_boundary, int levels, int32_t** out_quads, int32_t* out_nquads,
         float** out_verts, int32_t* out_nverts);
void vpt_cuda_free_host(void* p);
int  vpt_cuda_vertex_normals(const int32_t* elems, int nelems, int verts_per, const float* positions,
         int nverts, float* normals);
int  vpt_cuda_displace(float* positions, const float* normals, const float* texcoords, int nverts,
         const vpt_texture_desc* texture, float displacement);
/* light `light` of an uploaded scene: ids = {instance, environment, sdf}; *num_elements floats of
 * CDF are copied to `cdf` if it is non-null and `capacity` suffices */
int  vpt_cuda_scene_num_lights(const vpt_scene* scene);
int  vpt_cuda_download_light(const vpt_scene* scene, int light, int32_t ids[3], float* cdf,
                             size_t capacity, int32_t* num_elements);
void vpt_cuda_free_scene(vpt_scene* scene);
size_t vpt_cuda_scene_bytes(const vpt_scene* scene);

/* rngs: [width*height][2] uint64 {state, inc} per pixel, or null for zero */
int  vpt_cuda_create_state(int width, int height, const uint64_t* rngs, vpt_state** out);
/* re-initialise an existing state from host rngs and clear the image */
int  vpt_cuda_reset_state(vpt_state* state, const uint64_t* rngs);
/* advance every pixel's generator by `delta` draws (PCG jump-ahead): sample-range sharding */
int  vpt_cuda_advance_rngs(vpt_state* state, uint64_t delta);
void vpt_cuda_free_state(vpt_state* state);
int  vpt_cuda_state_samples(const vpt_state* state);

/* runs min(nsamples, params->samples - state.samples) samples for every pixel; blocking */
int  vpt_cuda_trace_samples(vpt_scene* scene, vpt_state* state, const vpt_params* params,
         int nsamples);
int  vpt_cuda_get_trace_stats(const vpt_state* state, vpt_trace_stats* stats);

/* Output stage on the device: get_render (yocto_pathtrace.cpp:1105-1116:
 * sum * (1 / samples)) followed by what save_image does for 8-bit formats
 * (yocto_sceneio.cpp:523-529: rgb_to_srgb + float_to_byte, alpha linear); n = width*height*4
 * bytes; `samples` is the divisor (pass 0 for the state's own count; after a cross-GPU reduce it is
 * the total over ranks). A quarter of the float image crosses PCIe; bytes equal the reference's. */
int  vpt_cuda_download_srgb8(vpt_state* state, int samples, uint8_t* rgba, size_t n);

/* The same with the reference's display tonemap in front (tonemap(), yocto_color.h:306-316, driven by
 * pathtrace_params.exposure / .filmic, yocto_pathtrace.h:94-95, as the viewer's display loop does,
 * apps/ypathtrace/ypathtrace.cpp:159-189): rgb * exp2(exposure) if exposure != 0, the ACES-fitted
 * tonemap_filmic curve (yocto_color.h:274-281) if filmic, then rgb_to_srgb + float_to_byte.
 * Bytes equal the reference's. The device and pinned host buffers live with the state (no
 * allocation per frame of a progressive preview). */
int  vpt_cuda_download_tonemapped8(vpt_state* state, int samples, float exposure, int filmic,
         uint8_t* rgba, size_t n);

/* Multi-GPU accumulation (the sample loop of apps/ypathtrace/ypathtrace.cpp:77-81
 * is where it lands): `peer` - a state of the same size, normally on another device, rendering a
 * disjoint sample range of every pixel (vpt_cuda_advance_rngs) - joins `root`. From then on
 * vpt_cuda_download_image / _srgb8 / _tonemapped8 on `root` return the SUM of the root's and all
 * peers' accumulation images: one kernel on the root's device reads the peers' images through
 * P2P loads over NVLink (copied over first where P2P is unavailable); the root's own image is
 * not modified. Peers must outlive the root's downloads; remove_peers detaches them. */
int  vpt_cuda_state_add_peer(vpt_state* root, vpt_state* peer);
int  vpt_cuda_state_remove_peers(vpt_state* root);
/* samples of the root plus all peers: the divisor of the summed image */
int  vpt_cuda_state_total_samples(const vpt_state* state);

/* un-normalised sums, n = width*height*4 floats (get_render divides by samples); with peers:
 * the sum over root and peers (see vpt_cuda_state_add_peer) */
int  vpt_cuda_download_image(vpt_state* state, float* rgba, size_t n);
int  vpt_cuda_download_rngs(vpt_state* state, uint64_t* rngs, size_t n);
/* device pointer of the accumulation image, for a collective on the caller's side */
void* vpt_cuda_image_device_ptr(vpt_state* state);
/* device-to-device copy of the accumulation image into / from a caller-owned device buffer
 * (e.g. a torch tensor that takes part in an NCCL reduce); n = width*height*4 floats */
int  vpt_cuda_export_image(vpt_state* state, void* device_dst, size_t n);
int  vpt_cuda_import_image(vpt_state* state, const void* device_src, size_t n);

/* first-hit queries for ray batches. rays: [n][8] = o.xyz, d.xyz, tmin, tmax.
 * instance < 0: whole scene; otherwise the instance-targeted variant.
 * ids: [n][3] = hit, instance, element; uvd: [n][3] = u, v, distance. */
int  vpt_cuda_intersect(vpt_scene* scene, const float* rays, int n, int instance,
         int32_t* ids, float* uvd);
/* vpt_cuda_intersect for callers that cannot allocate. NO COUNTERPART IN THE REFERENCE: added
 * for callers that must not call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it). Same arguments and results as
 * vpt_cuda_intersect (host pointers in, host pointers out, blocking), but the device staging of
 * rays / ids / uvd is one static __device__ array inside the library instead of three
 * cudaMalloc'ed buffers: the call performs no cudaMalloc / cudaFree / allocating library call.
 *   - limit: n <= VPT_NOALLOC_MAX_RAYS (65 536 rays = 3.5 MiB of device memory, held by every
 *     GPU that loads the library until the process exits); larger n -> VPT_ERR_INVALID with a
 *     message, nothing is launched;
 *   - the scene must live on the CURRENT device (the array's address is taken per device with
 *     cudaGetSymbolAddress on first use);
 *   - every call shares the array and finds the last call's contents in it (all n records are
 *     overwritten, nothing needs clearing); calls that overlap on the GPU would race: host
 *     threads of one process are serialised inside the call, separate processes have separate
 *     copies. */
#define VPT_NOALLOC_MAX_RAYS 65536
int  vpt_cuda_intersect_noalloc(vpt_scene* scene, const float* rays, int n, int instance,
         int32_t* ids, float* uvd);
/* ids: [n][3] = hit, vol_instance, sdf; dn: [n][4] = distance, normal.xyz */
int  vpt_cuda_spheretrace(vpt_scene* scene, const float* rays, int n, int maxiter,
         int32_t* ids, float* dn);

/* sums `count` floats of every peer buffer into dst on the current device (P2P loads) */
int  vpt_cuda_reduce_images(float* dst, const float* const* peers, int npeers, size_t count);

#ifdef __cplusplus
}
#endif
#endif /* VPT_CUDA_H */
