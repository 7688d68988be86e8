/* drt.h -- C ABI of the B200-native path-tracing hot path (libdrt.so).
 *
 * Drop-in boundary for the de-facto library surface of LightShade12/DustRayTracer (the reference has
 * no formal FFI; its only caller is the editor, DustRayTracer/src/Editor/EditorLayer.cpp:35-55,317-318,424).
 * Each entry point names the reference interface it replaces. Plain pointers and sizes only; every
 * function returns 0 on success or a negative drt_status, and drt_last_error() gives the message of the
 * calling thread's last failure (the reference prints and exit(99)s: Editor/Common/CudaCommon.cu:4-13).
 * There is no CPU fallback: every call that touches the device fails with DRT_ERR_CUDA without one.
 */
#ifndef DRT_H
#define DRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct drt_scene drt_scene;       /* replaces struct Scene            (Core/Scene/Scene.cuh:41-57) */
typedef struct drt_renderer drt_renderer; /* replaces class Renderer          (Core/Renderer.hpp:14-47)    */

typedef enum drt_status {
	DRT_OK = 0,
	DRT_ERR_INVALID = -1, /* bad argument / call order */
	DRT_ERR_IO = -2,      /* file missing or unreadable */
	DRT_ERR_FORMAT = -3,  /* malformed glTF / image */
	DRT_ERR_CUDA = -4,    /* CUDA runtime error (message carries cudaGetErrorString) */
	DRT_ERR_NOMEM = -5
} drt_status;

/* Field-for-field the reference's Camera POD (Core/Scene/Camera.cuh:32-46), 88 bytes. */
typedef struct drt_camera {
	float exposure;        /* 1 */
	float vfov_rad;        /* deg2rad(60) */
	float zfar, znear;     /* unused by the path, kept for layout */
	float aspect_ratio;    /* m_AspectRatio, unused */
	float defocus_angle;   /* degrees; <= 0 disables the thin lens */
	float focus_dist;      /* 10 */
	float movement_speed;  /* host fly-cam only */
	float position[3];     /* (0,2,5) */
	float forward_dir[3];  /* (0,0,-1) */
	float up_dir[3];       /* host fly-cam only */
	float right_dir[3];    /* host fly-cam only */
	uint32_t viewheight, viewwidth;
} drt_camera;

/* Field-for-field the reference's RendererSettings (Core/Scene/RendererSettings.h:22-34), 64 bytes. */
typedef struct drt_settings {
	uint8_t gamma_correction; /* true */
	uint8_t tone_mapping;     /* true */
	uint8_t enable_sunlight;  /* false */
	uint8_t _pad0;
	int32_t max_samples;      /* 500: rendering stops when the frame index reaches it (Renderer.cu:82) */
	int32_t ray_bounce_limit; /* 2: path segments = limit + 1 */
	int32_t render_mode;      /* 0 NORMALMODE, 1 DEBUGMODE */
	int32_t debug_mode;       /* 0 albedo, 1 normal, 2 barycentric, 3 uvs, 4 mesh-bvh heat, 5 world-bvh */
	int32_t _pad1;
	float sunlight_dir[2];    /* radians (-0.803, 0.681) */
	float sunlight_color[3];
	float sunlight_intensity; /* 30 */
	float sky_color[3];
	float sky_intensity;      /* 20 */
} drt_settings;

/* Field-for-field the reference's Material (Core/Scene/Material.cuh:4-23), 44 bytes. */
typedef struct drt_material {
	float albedo[3];
	float emissive_factor[3];
	int32_t albedo_texture_index; /* -1: none. Indexes IMAGES, like the reference (Scene.cu:79, RayGen.cuh:117) */
	float roughness;
	uint8_t transmission, _pad0[3];
	float refractive_index;
	uint8_t metallic, _pad1[3];
} drt_material;

typedef struct drt_bvh_params {
	int32_t bin_count;   /* informational: the GPU builder uses 16 bins (8 for small nodes); ref m_BinCount = 8 */
	int32_t leaf_target; /* informational: wide-node leaf slots hold <= 3 triangles; ref m_TargetLeafPrimitivesCount */
	int32_t reserved[6]; /* reserved[0] bit 0: disable the opacity classification of alpha-tested triangles (A/B checks);
	                        bit 1: build scenes of <= 2048 triangles with the level-synchronous multi-launch builder too instead
	                        of the single-launch one (same tree; A/B checks) */
} drt_bvh_params;

typedef struct drt_scene_stats {
	uint64_t triangles, materials, textures, meshes;
	uint64_t wide_nodes, bvh2_nodes, bvh_levels;
	uint64_t device_bytes;   /* nodes + triangles + shading records + textures resident in HBM */
	double build_ms;         /* device time of the last drt_scene_build_bvh */
	double load_ms;
	/* alpha-tested triangles (albedo texture with 4 components, AnyHit.cuh:17), and how many of them the exact host-side
	 * opacity classification proved always-opaque (test skipped) / never-hittable (collapsed); set by drt_scene_build_bvh */
	uint64_t alpha_tris, alpha_opaque, alpha_hidden;
	uint64_t build_launches; /* kernel launches of the last build: 1 for scenes of <= 2048 triangles (single-launch builder) */
} drt_scene_stats;

typedef struct drt_hit {
	int32_t prim_id; /* canonical id = index of the triangle in load order; -1 = miss */
	float t, u, v;   /* hit distance and barycentrics (UVW = (1-u-v, u, v), Intersection.cu:31) */
} drt_hit;

typedef struct drt_counters {
	uint64_t closest_rays; /* TraceRay calls  (Kernel/TraceRay.cu:15) */
	uint64_t shadow_rays;  /* RayTest calls   (Kernel/TraceRay.cu:35) */
	uint64_t node_visits, tri_tests, tie_events, alpha_tests, max_stack, texel_clamps; /* stats launches only */
} drt_counters;

typedef enum drt_format {
	DRT_RGBA32F = 0,      /* display buffer: acc / samples, a = 1   (what the reference writes to its GL surface) */
	DRT_RGBA8 = 1,        /* display buffer clamped to 8 bit          (what the editor's "save png" reads back)   */
	DRT_ACCUM_RGB32F = 2  /* raw accumulation sums, 3 floats / pixel  (the reference's float3 buffer)              */
} drt_format;

typedef enum drt_flags {
	DRT_FLAG_STRICT = 1,       /* separately-rounded fp32 in the reference's operation order (parity mode)      */
	DRT_FLAG_HOST_ARG_ORDER = 2, /* draw randomUnitVec3 components right-to-left like the g++-compiled oracle     */
	DRT_FLAG_STATS = 4,        /* count node visits / triangle tests (slower; for the roofline's V and T terms) */
	DRT_FLAG_WATERTIGHT = 16,  /* production arithmetic only (ignored with DRT_FLAG_STRICT): the watertight ray / triangle test of Woop,
	                              Benthin, Wald (JCGT 2013) replaces the reference's Moller-Trumbore (Intersection.cu:4-36), whose
	                              absolute determinant epsilon (1e-6, physical_units.hpp:12) drops small / distant triangles and
	                              whose per-triangle edge tests can leak between triangles that share an edge. Changes which
	                              rays hit near edges, so it is OFF by default and off in parity mode. */
	DRT_FLAG_EXTENDED_MATERIALS = 8 /* opt-in, beyond the reference: shade the material fields the
	                              reference loads but ignores (Material.cuh:10-22; RayGen.cuh never reads them):
	                              emissive_factor, metallic + roughness (mirror / fuzzy reflection), transmission +
	                              refractive_index (dielectric, with the reference's own dead helpers refract /
	                              reflectance, Random.cu:26-40). Off: exactly the reference's shading. */,
	DRT_FLAG_NO_TILE_ORDER = 32 /* keep the row-major block order. Without this flag a renderer measures, per block of 4 tiles, how long
	                              the block was resident, and runs the blocks of the following batched launches of the same view
	                              (>= 16 samples, >= two waves of blocks; a launch of >= 64 samples is ordered itself, after a
	                              one-sample probe launch) most expensive first, so that a launch drains on its cheapest tiles.
	                              The results do not depend on the order (bit-identical images and counters); no counterpart in the
	                              reference, whose launch is row-major (Renderer.hpp:43-44). */
} drt_flags;

/* ---- scene: replaces Scene::loadGLTFmodel (Scene.cu:181-317) and BVHBuilder::build* (BVHBuilder.cu:11-145) ---- */
int drt_scene_load_gltf(const char* path, drt_scene** out);
/* The step before the path. load_flags = 0 is drt_scene_load_gltf (the reference loader's exact
 * behaviour, quirks included: Scene.cu:120-178,192-207). DRT_LOAD_ROBUST reads the file as the glTF 2.0 specification
 * says: node hierarchy + transforms of the default scene, accessor byteOffset / bufferView byteStride / any buffer,
 * u8/u16/u32 or no indices, strips and fans, missing NORMAL / TEXCOORD_0, normalised integer attributes,
 * texture -> image indirection, alphaMode OPAQUE / MASK(alphaCutoff) / BLEND(= the reference's cut-out rule). */
typedef enum drt_load_flags {
	DRT_LOAD_REFERENCE_COMPAT = 0,
	DRT_LOAD_ROBUST = 1
} drt_load_flags;
int drt_scene_load_gltf_ex(const char* path, uint32_t load_flags, drt_scene** out);
/* Synthetic soups: pos 9 floats/tri, uv 6 floats/tri or NULL, material index per tri or NULL. */
int drt_scene_from_triangles(const float* pos, const float* uv, const int32_t* material_index, size_t n_tris,
	const drt_material* materials, size_t n_materials, drt_scene** out);
/* Appends an 8-bit texture (comps = 1..4) to a scene made by drt_scene_from_triangles. */
int drt_scene_add_texture(drt_scene* scene, int32_t width, int32_t height, int32_t comps, const uint8_t* pixels);
/* The image decoder of the loader on its own: replaces stbi_load_from_memory(data, len, &w, &h, &comps, 0) (Texture.cu:21-30,
 * stb_image 2.29): PNG and JPEG (baseline and progressive), 8 bits per channel out, the file's own channel count (JPEG: 3, or 1
 * for grey), stb's bytes. pixels_or_null = NULL asks for the size only. Host only, no GPU needed. */
int drt_decode_image(const uint8_t* data, size_t len, int32_t* width, int32_t* height, int32_t* comps, uint8_t* pixels_or_null, size_t bytes);
int drt_scene_build_bvh(drt_scene* scene, const drt_bvh_params* params_or_null, int device);
int drt_scene_get_stats(const drt_scene* scene, drt_scene_stats* out);
/* Host copy of the triangle soup in load order, in the reference's 128-byte Triangle layout (Triangle.cuh:7-19). */
int drt_scene_get_triangles(const drt_scene* scene, void* out, size_t bytes);
int drt_scene_get_materials(const drt_scene* scene, drt_material* out, size_t count);
int drt_scene_get_texture(const drt_scene* scene, int32_t index, int32_t* width, int32_t* height, int32_t* comps,
	uint8_t* pixels_or_null, size_t bytes);
/* Device-built tree for inspection: 80-byte wide nodes and 48-byte leaf-order triangle records. */
int drt_scene_get_wide_nodes(const drt_scene* scene, void* out, size_t bytes);
int drt_scene_get_leaf_triangles(const drt_scene* scene, void* out, size_t bytes);
void drt_scene_destroy(drt_scene* scene);

/* ---- renderer: replaces Renderer::{ResizeBuffer,Render,resetAccumulationBuffer,getSampleCount} (Renderer.cu:29-136) ---- */
int drt_renderer_create(int device, drt_renderer** out);
int drt_renderer_set_stream(drt_renderer* r, void* cuda_stream); /* launch on a caller-owned stream (NULL = own stream) */
int drt_renderer_set_flags(drt_renderer* r, uint32_t flags);
/* Kernel selection for measurements. Results do not depend on it. variant -1 (the state of a new renderer) = automatic:
 * variant 4 (scenes of 1024 triangles or more: variant 15 = variant 4 with a TRI step that tests two pending triangles at once),
 * except that launches of at most 8 samples are timed on variant 4 and on the persistent pixel-queue kernel
 * (variant 9) and run on the faster one (re-probed after 256, 512, ... 4096 launches). Explicit values: 4 = per-warp phase voting over
 * NODE/TRI/SHADE steps, 6 blocks/SM (expanded fp32 nodes when the scene has them), 2 = the same without the register cap,
 * 3 / 7 / 8 = compressed nodes, 0 = flat per-thread state machine, 1 = per-sample bounce loop, 11-14 = variant 4 / 7 with
 * the SHADE-only lane state in shared memory (8 / 7 / 6 blocks per SM; 14 = compressed nodes), 9 / 10 = their persistent
 * pixel-queue form (fp32 / compressed nodes), 16 = block-level regrouping of path contexts in shared memory (measured alternative),
 * 17 / 18 = variants 4 / 15 with a shared-memory traversal stack (measured alternative), 22 / 23 = variants 4 / 15 with a SAMPLE POOL:
 * the last R of a pixel's samples are handed out dynamically inside the warp's tile, which removes the wait for the slowest lane
 * (what automatic selection runs for launches of >= 8 samples on scenes of >= 128 triangles; same bits as every other variant).
 * block_threads = 0 (default 128) or a multiple of 32 up to 256. */
int drt_renderer_set_tuning(drt_renderer* r, int32_t variant, int32_t block_threads);
void drt_renderer_destroy(drt_renderer* r);
int drt_renderer_kernel_launches(const drt_renderer* r); /* kernels launched by this renderer so far */
int drt_renderer_last_variant(const drt_renderer* r);    /* kernel variant the last render launch ran (automatic selection resolved) */
int drt_renderer_last_launch_ordered(const drt_renderer* r); /* 1 if the last render launch ran its blocks in cost order (DRT_FLAG_NO_TILE_ORDER) */
/* Checkpoint / resume: restore accumulation sums (3 floats / pixel) and the number of samples they hold. */
int drt_renderer_upload_accum(drt_renderer* r, const float* rgb32f, size_t bytes, uint32_t samples_done);
/* Sample-range sharding across GPUs: this renderer renders frame indices first, first+stride, ... (default 1, 1). */
int drt_renderer_set_frame_schedule(drt_renderer* r, uint32_t first_frame, uint32_t stride);
/* Image-tile sharding across GPUs (BASELINE config 3): this renderer renders only the 8x4-pixel tiles whose
 * row-major index is first_tile, first_tile+stride, ... (default 0, 1 = all) and leaves the other pixels of its buffers
 * untouched (zero after a reset), so that ONE sum-reduce of the accumulation buffers over the ranks yields the bits of
 * the single-GPU image. Every rank uses the same frame schedule. Resets the accumulation. */
int drt_renderer_set_tile_schedule(drt_renderer* r, uint32_t first_tile, uint32_t stride);
/* Image-tile sharding WITHOUT a collective: the renderers of the other ranks render their tiles directly into the root
 * renderer's accumulation and display buffers (peer stores over NVLink / NVSwitch from inside the render kernel), so a frame
 * is complete on the root as soon as every rank's launch has returned -- no reduce, no zeroing of foreign pixels, no resolve.
 *   export: the root (an ordinary, resized renderer) writes DRT_FRAME_HANDLE_BYTES bytes to send to the other PROCESSES;
 *   import: a renderer of the same size in another process (any device of the node) maps them (CUDA IPC + peer access);
 *   attach: the same for two renderers of ONE process (owner == NULL detaches).
 * An attached / importing renderer keeps its own frame schedule, tile schedule and counters; its reset leaves the shared
 * buffers to their owner. The caller orders the ranks (barrier after the owner's reset and after the launches). */
#define DRT_FRAME_HANDLE_BYTES 160
int drt_renderer_export_frame_buffers(drt_renderer* r, void* handle_out);
int drt_renderer_import_frame_buffers(drt_renderer* r, const void* handle);
int drt_renderer_attach_frame_buffers(drt_renderer* r, drt_renderer* owner);
/* Fused render + reduce: SAMPLE-RANGE sharding without a collective (no counterpart in the reference). With on != 0 the batched
 * launches of this renderer start every pixel sum at zero and ADD it to the accumulation buffer with fp32 atomics
 * (red.global.add.f32) instead of load + store, and leave the display buffer alone. Renderers that share ONE pair of frame buffers
 * (drt_renderer_attach_frame_buffers / _import_frame_buffers: the buffers of the root, mapped over NVLink) and own different frame
 * indices (drt_renderer_set_frame_schedule) then sum their samples in the root's buffer from inside the render kernels: compute and
 * exchange are one kernel. The caller zeroes the buffer first (the root's reset), orders the ranks (a barrier after the launches)
 * and calls drt_renderer_resolve(total samples) on the root. The sum is that of the per-renderer partial sums, as with the
 * reduce; with more than two renderers the ORDER of those additions, and with it the last bit, varies from run to run.
 * Fixed-tile kernels only: debug views, extended materials and explicit variants 0 / 1 / 9-14 / 16 return DRT_ERR_CUDA. */
int drt_renderer_set_accumulate_add(drt_renderer* r, int on);
int drt_renderer_resize(drt_renderer* r, uint32_t width, uint32_t height);
/* Renders up to n_samples further samples in one batch (the reference renders one per call) and updates the
 * display buffer; stops at settings->max_samples like Renderer.cu:82. ms = device time of the launch(es). */
int drt_renderer_render(drt_renderer* r, const drt_camera* cam, const drt_scene* scene, const drt_settings* settings,
	uint32_t n_samples, float* ms_or_null);
int drt_renderer_reset(drt_renderer* r);
uint32_t drt_renderer_sample_count(const drt_renderer* r); /* the reference's m_FrameIndex: samples done + 1 */
uint32_t drt_renderer_width(const drt_renderer* r);
uint32_t drt_renderer_height(const drt_renderer* r);
int drt_renderer_readback(drt_renderer* r, drt_format format, void* host, size_t bytes);
/* Device pointer of the W*H float4 accumulation buffer (for an NCCL reduce over NVLink) and the sample total
 * the display divides by; drt_renderer_resolve recomputes the display buffer after an external reduce. */
int drt_renderer_accum_device_ptr(drt_renderer* r, void** ptr, size_t* bytes);
int drt_renderer_resolve(drt_renderer* r, uint32_t total_samples);
int drt_renderer_get_counters(drt_renderer* r, drt_counters* out, int reset);
/* Parity probe: primary rays exactly as the path issues them for `frame_index`, closest hit only. */
int drt_trace_primary(drt_renderer* r, const drt_camera* cam, const drt_scene* scene, uint32_t frame_index,
	drt_hit* out_host, size_t count);

/* ---- N GPUs of one node behind one handle, NCCL inside the library (drt_renderer_create(device_ordinals[], n); 8e) ----
 * What the reference's only kind of caller -- a C++ host, Editor/EditorLayer.cpp:317-318 -- needs to use more than one GPU
 * without Python: every device gets a replica of the scene (built there) and a renderer; drt_multi_render splits a batch of
 * samples over the devices (mode 0, default: device g renders frame indices g+1, g+1+N, ...; mode 1: device g renders the 8x4
 * tiles g, g+N, ... of every frame, which reproduces the one-GPU bits), runs ONE ncclReduce(sum) over NVLink of the W*H float4
 * accumulation buffers to the first device and resolves the display buffer there. n = 1 is exactly the single-device renderer.
 * One host thread drives all devices (ncclCommInitAll); NCCL is loaded at run time (libnccl.so.2), so the library has no
 * link-time dependency on it and drt_multi_create fails with a message where it is missing. */
typedef struct drt_multi drt_multi;
int drt_multi_create(const int32_t* device_ordinals, int32_t n, drt_multi** out);
void drt_multi_destroy(drt_multi* m);
int drt_multi_device_count(const drt_multi* m);
int drt_multi_set_scene(drt_multi* m, const drt_scene* scene, const drt_bvh_params* params_or_null); /* replicates + builds on every device */
int drt_multi_set_sharding(drt_multi* m, int32_t mode); /* 0 = sample ranges, 1 = image tiles; resets the accumulation */
int drt_multi_set_flags(drt_multi* m, uint32_t flags);
int drt_multi_resize(drt_multi* m, uint32_t width, uint32_t height);
int drt_multi_render(drt_multi* m, const drt_camera* cam, const drt_settings* settings, uint32_t n_samples, float* ms_or_null);
int drt_multi_reset(drt_multi* m);
uint32_t drt_multi_sample_count(const drt_multi* m); /* samples in the frame + 1, like Renderer::getSampleCount */
int drt_multi_readback(drt_multi* m, drt_format format, void* host, size_t bytes);
int drt_multi_get_counters(drt_multi* m, drt_counters* out, int reset); /* summed over the devices */
int drt_multi_nccl_reduces(const drt_multi* m); /* collectives issued so far */

/* Ray-level parity probe: n_rays caller-given rays (origins / directions: 3 floats each, host memory; directions as the path
 * would pass them, not normalised) through the very step functions the render kernels run (wide-node slab test, Moller-Trumbore
 * in the reference's operation order, alpha cut-out). anyhit = 0: TraceRay closest hit (Kernel/TraceRay.cu:15-33 ->
 * BVH/BVHTraversal.cuh:14-73); anyhit != 0: RayTest (TraceRay.cu:35-38 -> BVHTraversal.cuh:76-134), which ignores the hit
 * interval and ends at the first accepted triangle (prim_id >= 0 means "occluded"; WHICH triangle depends on the tree).
 * Replaces nothing the editor calls: it exists so that the reference's Intersection.cu:4-36 edge cases (ray in the triangle's
 * plane, |det| at the epsilon, t below epsilon, hits on edges) can be pushed through the CUDA code one ray at a time. */
int drt_trace_rays(drt_renderer* r, const drt_scene* scene, const float* origins, const float* directions, size_t n_rays,
	int anyhit, drt_hit* out_host);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation). Same kernel, arguments and results
 * as drt_trace_primary, but the device-side staging of the hit records lives in a static device array inside libdrt.so
 * instead of a lazily allocated buffer: the call performs no cudaMalloc / cudaFree and calls no allocating library routine.
 * Limit: width * height <= DRT_NOALLOC_MAX_PIXELS (16-byte records: 4 MiB, held by every GPU that loads the library until
 * exit); larger renderers are rejected with DRT_ERR_INVALID. All calls share that array and it keeps the last call's
 * contents (every record copied out is overwritten first, nothing has to be cleared): calls that overlap on one GPU --
 * two renderers, two threads, two streams -- race on it, so serialise them. The first call on a device resolves the
 * array's device address (cudaGetSymbolAddress, with that device current); make it before starting a graph capture. */
#define DRT_NOALLOC_MAX_PIXELS 262144 /* = 512 x 512, BASELINE config 1 */
int drt_trace_primary_noalloc(drt_renderer* r, const drt_camera* cam, const drt_scene* scene, uint32_t frame_index,
	drt_hit* out_host, size_t count);

void drt_default_camera(drt_camera* out);     /* Camera.cuh:32-46 + EditorLayer.cpp:35-40 */
void drt_default_settings(drt_settings* out); /* RendererSettings.h:22-34 */
const char* drt_last_error(void);
const char* drt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DRT_H */
