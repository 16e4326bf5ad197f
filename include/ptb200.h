/* ptb200.h -- C ABI of the B200-native wavefront path tracer (libptb200.so).
 *
 * This is the drop-in boundary for the ONE hot path of HugoPeters1024/cuda_pathtracer:
 *   class Application { Init(); Render(const Camera&, float, float, bool shouldClear); Finish(); }
 *   (reference src/application.h:7-20, implemented by Pathtracer, src/pathtracer.h:46-310).
 * The reference has no FFI; its host code hands the kernels a by-value `SceneBuffers`
 * (src/types.h:689-703), a `Camera` (src/types.h:586-677) and a `RandState` (src/types.h:679-687).
 * Every POD struct below is layout-compatible with the reference struct it cites (sizes and
 * offsets are pinned against the reference's own headers by tests/test_abi.py), so reference-side
 * host code can pass its arrays straight through.  All entry points return 0 on success or a
 * negative pt_status; nothing here calls exit() (the reference's cudaSafe does, use_cuda.h:32-40).
 * There is no CPU fallback behind this ABI: every compute entry point needs a CUDA device.
 */
#ifndef PTB200_H
#define PTB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PT_ALIGN(n) __attribute__((aligned(n)))
#define PT_API __attribute__((visibility("default")))
#else
#define PT_ALIGN(n)
#define PT_API
#endif

typedef enum {
    PT_OK = 0,
    PT_ERR_INVALID = -1,   /* bad argument / call order */
    PT_ERR_CUDA = -2,      /* a CUDA runtime call failed; see pt_last_error */
    PT_ERR_NO_DEVICE = -3, /* no CUDA device: there is no CPU fallback */
    PT_ERR_OOM = -4,
    PT_ERR_IO = -5
} pt_status;

/* ---- POD mirrors of the reference structs (src/types.h) -------------------------------------- */
typedef struct { float x, y, z; } pt_float3;            /* CUDA float3: 12 B, align 4 */
typedef struct PT_ALIGN(8) { float x, y; } pt_float2;   /* CUDA float2: 8 B, align 8  */
typedef struct PT_ALIGN(16) { float x, y, z, w; } pt_float4;

/* Material, src/types.h:33-56 (88 B, align 8).  `texture`/`normal_texture` are CUDA texture
 * object handles valid in the calling process (cudaTextureObject_t, or pt_texture_create). */
typedef struct {
    pt_float3 diffuse_color;
    pt_float3 specular_color;
    pt_float3 emission;
    float reflect, glossy, transmit, refractive_index;
    pt_float3 absorption;
    uint64_t texture;
    uint64_t normal_texture;
    uint8_t hasTexture, hasNormalMap;
    uint8_t _pad[6];
} pt_material;

typedef struct { pt_float3 pos; float radius; uint32_t material; } pt_sphere;   /* types.h:58-69  */
typedef struct { pt_float3 normal; float d; uint32_t material; } pt_plane;      /* types.h:124-133 */

typedef struct PT_ALIGN(16) { pt_float3 v0, v1, v2; uint32_t _pad[3]; } pt_triangle_v;   /* types.h:79-83, 48 B */
typedef struct PT_ALIGN(16) {                                                            /* types.h:85-96, 80 B */
    pt_float3 normal, tangent, bitangent;
    pt_float2 uv0, uv1, uv2;
    uint32_t material;
} pt_triangle_d;

typedef struct PT_ALIGN(16) {          /* RadianceCache, types.h:98-104, 112 B */
    float radianceCache[8];
    float additionCache[8];
    float additionCacheCount[8];
    float radianceTotal;
} pt_radiance_cache;

typedef struct { uint32_t triangle_index, instance_index; } pt_triangle_light;   /* types.h:106-110 */

typedef struct PT_ALIGN(16) { pt_float4 origin; pt_float4 direction; } pt_ray;   /* RayPacked, types.h:273-294:
                                            origin.w = length (tmax), direction.w = pixel index (uint bits) */
typedef struct PT_ALIGN(16) {          /* HitInfo / HitInfoPacked, types.h:299-334 */
    uint32_t primitive_type;           /* 0 TRIANGLE, 1 SPHERE, 2 PLANE (types.h:296) */
    uint32_t primitive_id;             /* 0xffffffff = miss */
    uint32_t instance_id;
    float t;
} pt_hit;

typedef struct PT_ALIGN(16) {          /* SampleCache, types.h:339-345 */
    uint32_t sample_type;              /* 0 IGNORE, 1 TERMINATE, 2 BUCKET (types.h:337) */
    uint32_t triangle_id;
    int32_t cache_bucket_id;
    pt_float3 cum_mask;
} pt_sample_cache;

typedef struct PT_ALIGN(16) { pt_float4 vmin; pt_float4 vmax; } pt_bvh_node;   /* BVHNode, types.h:347-375:
                                            vmin.w = child1 | t_start (uint bits), vmax.w = t_count (leaf iff > 0) */
typedef struct PT_ALIGN(16) {          /* Model, types.h:377-383 */
    const pt_bvh_node* bvh;
    uint32_t triangleStart, nrTriangles, nrBvhNodes;
} pt_model;

typedef struct { pt_float4 r1, r2, r3; } pt_mat4x3;     /* types.h:385-406, rows of the 3x4 matrix */
typedef struct PT_ALIGN(16) {          /* Instance, types.h:408-414, 128 B */
    uint32_t model_id;
    pt_mat4x3 transform;
    pt_mat4x3 invTransform;
    uint32_t material_id;              /* 0xffffffff => per-triangle material */
} pt_instance;

typedef struct PT_ALIGN(16) {          /* TopLevelBVH, types.h:431-459, 32 B */
    pt_float3 vmin, vmax;
    uint32_t children;                 /* child1 | child2 << 16; 0 => leaf */
    uint32_t leaf;
} pt_tlas_node;

typedef struct {                       /* Camera, types.h:586-677, 76 B (first member is the private bool) */
    uint32_t _has_moved;
    pt_float3 eye, viewDir;
    pt_float3 lt, u, v;
    float d, focalLength, aperture;
} pt_camera;

/* Host-pointer scene description == what Pathtracer::Init uploads (src/pathtracer.h:129-190,216-217).
 * Mirrors SceneBuffers (types.h:689-703) + the host vectors of Scene (scene.h:124-133). */
typedef struct {
    const pt_triangle_v* vertices;     /* post-permute, global triangle order */
    const pt_triangle_d* vertexData;
    uint32_t num_triangles;
    const pt_model* models;            /* bvh = HOST pointers to the per-model BVH2 (createBVHBinned output) */
    uint32_t num_models;
    const pt_material* materials;
    uint32_t num_materials;
    const pt_instance* instances;
    uint32_t num_instances;
    const pt_tlas_node* topBvh;        /* 2*num_instances-1 nodes (scene.h:358) */
    uint32_t num_tlas_nodes;
    const pt_sphere* spheres;
    uint32_t num_spheres;
    const pt_plane* planes;
    uint32_t num_planes;
} pt_scene_desc;

/* Run-time replacement for the reference's compile-time constants (src/constants.h:4-14). */
typedef struct {
    uint32_t width, height;            /* WINDOW_WIDTH / WINDOW_HEIGHT */
    uint32_t max_ray_depth;            /* MAX_RAY_DEPTH (reference: 32) */
    uint32_t interactive_depth;        /* Scene::interactive_depth (scene.h:137) */
    uint32_t guiding_frames;           /* frames during which the radiance cache trains (pathtracer.h:292: 100) */
    uint32_t blue_noise_frames;        /* rand() uses blue noise while sampleIdx < this (kernels.h:23: 1) */
    int32_t device;                    /* CUDA device ordinal */
    uint32_t rank, world_size;         /* tile partition: this context renders tiles t with t % world_size == rank */
    uint32_t flags;                    /* PT_FLAG_* */
} pt_config;
#define PT_FLAG_DEFAULT 0u
/* Scheduling only (results are identical): in a sample that runs alone (clearing, cache-training frames) bounces >= the
 * tail bounce (default 3) run as ONE per-path launch instead of wavefront launches per bounce; samples that overlap
 * others stay wavefront.  PT_FLAG_TAIL_BOUNCE(n) forces the per-path launch from bounce n for every sample,
 * PT_FLAG_WAVEFRONT_ONLY keeps every bounce of every sample in the wavefront. */
#define PT_FLAG_WAVEFRONT_ONLY 1u
/* Converge samples normally rotate through four working sets on private streams (samples k+1.. start while the
 * short-queue end of sample k drains); results are identical.  This flag keeps one sample in flight. */
#define PT_FLAG_NO_PIPELINE 2u
#define PT_FLAG_TAIL_BOUNCE(n) ((((uint32_t)(n) + 1u) & 0xffu) << 8)
/* pt_upload_scene converts the BVH2 to the 8-wide layout on the device; this flag selects the host converter instead
 * (same bytes; kept as the checker of the device converter). */
#define PT_FLAG_HOST_WIDE_BVH 4u
/* Stage kernels are launched with programmatic dependent launch (the next kernel's blocks are scheduled while the previous
 * one drains; every kernel waits for its predecessor's results itself).  This flag restores plain stream order. */
#define PT_FLAG_NO_PDL 8u
/* Ray sorting between shade and trace on from the start (see pt_set_ray_sorting). */
#define PT_FLAG_SORT_RAYS 16u

typedef struct {                       /* device-side ray counters since the last pt_reset_stats */
    uint64_t primary_rays, extend_rays, shadow_rays, shade_calls, samples;
    uint64_t wave_extend_rays, wave_shadow_rays;   /* the part of extend/shadow rays traced by the wavefront launches (the rest: k_tail) */
    uint64_t kernel_launches;                      /* kernels pt_render launched (host-side count) */
} pt_stats;

typedef struct pt_ctx pt_ctx;

/* Called between the local radiance-cache update and the propagate step of a training frame
 * (north_star: NCCL all-reduce of additionCache||additionCacheCount).  `dev_additions` is a device
 * pointer to 16*num_triangles floats (SoA: [16][num_triangles]); the hook must enqueue an in-place
 * SUM all-reduce on `cuda_stream`. */
typedef int (*pt_allreduce_hook)(void* user, float* dev_additions, size_t count, void* cuda_stream);

PT_API const char* pt_version(void);
PT_API void pt_default_config(pt_config* cfg);
PT_API int  pt_create(pt_ctx** out, const pt_config* cfg);          /* ~ Pathtracer ctor + allocation half of Init() (pathtracer.h:192-209) */
PT_API void pt_destroy(pt_ctx* ctx);
PT_API const char* pt_last_error(const pt_ctx* ctx);                /* ctx may be NULL: last create error */

PT_API int  pt_upload_scene(pt_ctx*, const pt_scene_desc*);         /* ~ upload half of Init() (pathtracer.h:129-190,216-220) */
PT_API int  pt_update_instances(pt_ctx*, const pt_instance*, uint32_t n, const pt_tlas_node*, uint32_t n_nodes); /* pathtracer.h:245-246 */
/* Same re-sync with the top-level BVH rebuilt ON THE DEVICE from the instance matrices (BuildTopLevelBVH / FindBestMatch /
 * transformBox, scene.h:27-118: same nodes as the host code); stream-ordered, no host-side tree. */
PT_API int  pt_update_instances_device_tlas(pt_ctx*, const pt_instance*, uint32_t n);
PT_API int  pt_debug_read_tlas(pt_ctx*, pt_tlas_node* out, uint32_t n_nodes);   /* test hook: the top-level BVH as it sits on the device */
PT_API int  pt_set_toggles(pt_ctx*, int nee, int guiding);          /* HNEE / HCACHE (globals.h:19-20); takes effect on the next should_clear render (pathtracer.h:242-243) */
PT_API int  pt_set_guiding_frames(pt_ctx*, uint32_t frames);           /* run-time value of the training window (pathtracer.h:292: sampleIdx < 100) */
PT_API int  pt_set_skydome(pt_ctx*, const float* rgba, uint32_t w, uint32_t h);   /* ~ loadTextureHDR (use_cuda.h:202-263): rows already flipped */
PT_API int  pt_set_sky_constant(pt_ctx*, float r, float g, float b);            /* constant sky radiance (default 1,1,1) when no HDR dome is available */
PT_API int  pt_set_blue_noise(pt_ctx*, const float* lum, uint32_t w, uint32_t h); /* ~ loadTextureL (use_cuda.h:166-200) */
PT_API int  pt_texture_create(pt_ctx*, const float* rgba, uint32_t w, uint32_t h, uint64_t* out_handle); /* ~ loadTexture (use_cuda.h:108-164) */
/* SURVEY 8(f) row 4, both opt-in (defaults reproduce the reference's shipped behaviour):
 * skydome importance sampling through the CDF Pathtracer::Init builds (pathtracer.h:85-122, binarySearch kernels.h:65-79;
 * the reference never samples it): diffuse vertices send their shadow ray to a dome texel drawn from the CDF, sky radiance
 * is then counted by those rays instead of by diffuse paths that escape;
 * ray sorting (the reference's disabled experiment, src/raySorting.h:1-68): counting sort of the extension and shadow queues
 * by (origin cell, direction octant) between shade and trace; results are bit-identical, only the traversal order changes. */
PT_API int  pt_set_sky_sampling(pt_ctx*, int on);
PT_API int  pt_set_ray_sorting(pt_ctx*, int on, uint32_t bits_per_axis /* 1..5, 0 = keep */, uint32_t bounces /* 1..2, 0 = keep */);
PT_API int  pt_set_allreduce_hook(pt_ctx*, pt_allreduce_hook hook, void* user);
/* Multi-GPU exchange over NVLink peer memory, the default transport of the radiance-cache step on one box (replaces the
 * hook + propagate pair == what kernel_update_buckets -> [all-reduce] -> kernel_propagate_buckets needs, kernels.h:863-905):
 * every rank exports the cudaIpc handle of its radiance-cache block, the caller gathers the handles of all ranks in rank
 * order (any transport) and hands them to pt_peer_connect; training frames then run ONE fused reduce + propagate +
 * broadcast kernel per rank over the peers' HBM.  All ranks must render the same sequence of frames. */
#define PT_PEER_HANDLE_BYTES 128    /* two cudaIpc handles: radiance-cache block, framebuffer-gather block */
PT_API int  pt_peer_export(pt_ctx*, void* handle_out /* PT_PEER_HANDLE_BYTES */);
PT_API int  pt_peer_connect(pt_ctx*, const void* handles /* world_size x PT_PEER_HANDLE_BYTES, rank order */, uint32_t world_size);
PT_API int  pt_peer_disconnect(pt_ctx*);

PT_API int  pt_render(pt_ctx*, const pt_camera*, int should_clear);   /* == Pathtracer::Render (pathtracer.h:224-302); asynchronous */
PT_API int  pt_join(pt_ctx*);                                         /* asynchronous: orders every render enqueued so far before later work on pt_cuda_stream() (e.g. a timing event) */
PT_API int  pt_finish(pt_ctx*);                                       /* == Pathtracer::Finish (pathtracer.h:305-310); synchronises */
PT_API int  pt_read_image(pt_ctx*, float* host_rgbw, int which);      /* headless stand-in for the GL textures: which = 0 luminance (rgb sum, w = samples), 1 albedo */
PT_API int  pt_read_tiles(pt_ctx*, float* host_rgbw_slots, int which);  /* this rank's pixels only: pt_tile_num_slots() float4 in slot order (multi-GPU: 1/N of the image per rank) */
/* Streaming read-back: the image as it is after everything enqueued so far, copied to (pinned) host memory on a private
 * stream while later samples render (only their accumulate step waits for the copy, so the snapshot is exact).
 * pt_wait_read(ticket) blocks until that copy has landed; pt_finish / pt_join cover all of them. */
PT_API int  pt_read_image_async(pt_ctx*, float* host_rgbw, int which, uint64_t* ticket);
PT_API int  pt_read_tiles_async(pt_ctx*, float* host_rgbw_slots, int which, uint64_t* ticket);
PT_API int  pt_wait_read(pt_ctx*, uint64_t ticket);
/* N ranks, one framebuffer (SURVEY 8(e) "end of run: gather tiles"; main.cpp:308-319 has one consumer per frame): collective
 * over the connected peers.  Every rank's tiles travel over NVLink into an image ring in rank 0's HBM, rank 0 ships the
 * assembled W*H float4 image to host_full (pinned; ignored on other ranks) with one copy.  Asynchronous, tickets as above. */
PT_API int  pt_gather_image_async(pt_ctx*, float* host_full, int which, uint64_t* ticket);
/* N ranks, one framebuffer in HOST memory, every rank through its own PCIe link (the gather above funnels 16 B/pixel through rank
 * 0's link).  The caller provides a block of process-shared host memory (e.g. a /dev/shm file mapped by every rank's process),
 * zero-filled by its creator: PT_HOST_RING_HEADER_BYTES of header (uint32 arrived[rank] at byte 4*rank, uint32 consumed at byte
 * 256) followed by `slots` images of width*height float4.  attach maps it into this rank's GPU (cudaHostRegister); deliver (every
 * rank, collective in the sense that all ranks call it after the same Render calls) stores this rank's tiles into slot
 * epoch % slots and publishes the epoch; wait (the consumer, any attached context) returns the assembled image once all ranks
 * have published it; release hands the slot back.  deliver blocks (bounded) while the slot it needs is unreleased. */
#define PT_HOST_RING_HEADER_BYTES 4096u
PT_API int  pt_host_ring_attach(pt_ctx*, void* base, uint64_t bytes, uint32_t slots);
PT_API int  pt_host_ring_detach(pt_ctx*);
PT_API int  pt_deliver_image_async(pt_ctx*, int which, uint64_t* epoch);
PT_API int  pt_host_ring_wait(pt_ctx*, uint64_t epoch, const float** image);
PT_API int  pt_host_ring_release(pt_ctx*, uint64_t epoch);
PT_API int  pt_write_pfm(pt_ctx*, const char* path, int which);        /* headless render-to-PFM (display value rgb / w, main.cpp:58-62) in place of the GLFW window */
/* Display post-process (main.cpp:46-171,316-376: demodulated separable Gaussian, remodulate, gamma 2, vignette) as CUDA
 * kernels; host float4 in (or NULL, NULL = the context's accumulators), host float4 RGBA out. */
PT_API int  pt_postprocess(pt_ctx*, const float* lum_host, const float* albedo_host, float nr_samples, int blur, float* out_host_rgba);
/* Roofline helper: measured GB/s of random 16/32-byte reads over an array of `array_bytes` (L2-resident when it fits). */
PT_API int  pt_measure_gather_bandwidth(pt_ctx*, uint64_t array_bytes, uint32_t access_bytes, uint32_t repeats, double* out_gbs);
PT_API int  pt_read_radiance_cache(pt_ctx*, pt_radiance_cache* host_out);
/* createBVHBinned (bvhBuilder.h:35-274: 16-bin SAH, leaf <= 4 triangles, two-pointer partition, pre-order node ids) built
 * on the device, level by level; same nodes, same triangle order and same node numbering as the reference's host build
 * (a box may differ in the SIGN of a zero coordinate).  tris: n host TriangleV records; nodes: room for max(1, 2n-1);
 * order[i] = input index of the triangle that ends up at position i (apply it to the TriangleV / TriangleD arrays ==
 * permuteTriangles, bvhBuilder.h:17-27).  device_ms (optional): device time of the build without the copies;
 * levels (optional): tree levels processed.  No CPU fallback: PT_ERR_NO_DEVICE without a GPU. */
PT_API int  pt_build_bvh(int device, const pt_triangle_v* tris, uint32_t n, uint32_t tri_offset, pt_bvh_node* nodes, uint32_t* n_nodes,
                         uint32_t* order, float* device_ms, uint32_t* levels);
PT_API const char* pt_build_bvh_last_error(void);
/* Test hook: the 8-wide BVH as it sits on the device (80-byte nodes, 48-byte triangle records); NULL pointers: sizes only. */
PT_API int  pt_debug_read_wide_bvh(pt_ctx*, void* nodes, uint32_t* n_nodes, void* tris, uint32_t* n_tris);
/* Per-stage device time, CUDA events around every stage launch on the context's stream; resolved by pt_finish. */
#define PT_STAGE_EXTEND 0
#define PT_STAGE_SHADE 1
#define PT_STAGE_CONNECT 2
#define PT_STAGE_GUIDING 3
#define PT_STAGE_OTHER 4
#define PT_STAGE_TAIL 5     /* k_tail: extend + shade + connect of every bounce >= the tail bounce */
#define PT_NUM_STAGES 6
PT_API int  pt_set_profiling(pt_ctx*, int on);                         /* also resets the accumulators */
PT_API int  pt_get_profile(pt_ctx*, double* ms /*[PT_NUM_STAGES]*/, uint64_t* launches /*[PT_NUM_STAGES]*/);
/* Measurement mode for the roofline line: every bounce as separate wavefront launches of counting traversal kernels;
 * pt_get_tally (8 values): node steps, triangle tests, top-level steps of the extension rays [0..2] and of the shadow rays [3..5];
 * [6] / [7]: the most steps (node + triangle) any single extension / shadow ray took -- what a short launch is bounded by. */
PT_API int  pt_set_counting(pt_ctx*, int on);
PT_API int  pt_get_tally(pt_ctx*, uint64_t* out8);
PT_API int  pt_debug_read_counters(pt_ctx*, uint32_t* ext_out, uint32_t* shadow_out, uint32_t n_bounces);   /* per-bounce queue sizes of the last sample on working set 0 */
PT_API int  pt_get_stats(pt_ctx*, pt_stats* out);
PT_API int  pt_reset_stats(pt_ctx*);
PT_API int  pt_get_rand_state(pt_ctx*, uint32_t* sampleIdx, uint32_t* randIdx);

/* Stage-level parity hooks (host buffers in, host buffers out; pixel index = RayPacked.direction.w). */
PT_API int  pt_debug_primary_rays(pt_ctx*, const pt_camera*, uint32_t randIdx, pt_ray* host_out /* width*height, indexed by pixel */);
PT_API int  pt_debug_trace(pt_ctx*, const pt_ray* host_rays, uint32_t n, int any_hit, pt_hit* host_out);
/* Device-resident variant used for kernel-only timing: rays stay in HBM, returns elapsed ms (CUDA events). */
PT_API int  pt_debug_trace_resident(pt_ctx*, const pt_ray* host_rays, uint32_t n, int any_hit, uint32_t repeats, float* out_ms, pt_hit* host_out);

/* No counterpart in the reference.  pt_debug_trace for callers that may not
 * allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): same kernel, same
 * results, but its scratch (work cursor + staged rays + hit records) is a file-scope device array of
 * PT_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Calls neither cudaMalloc nor cudaFree nor any allocating
 * routine.  n > PT_STATIC_TRACE_MAX_RAYS (one 640x480 frame) is PT_ERR_INVALID with a message in pt_last_error.  All calls
 * share the array (it keeps the previous call's contents; the cursor is cleared per call): calls that overlap on one GPU race. */
#define PT_STATIC_TRACE_MAX_RAYS 307200u
#define PT_STATIC_SCRATCH_BYTES (256u + PT_STATIC_TRACE_MAX_RAYS * 48u)
PT_API int  pt_trace_static(pt_ctx*, const pt_ray* host_rays, uint32_t n, int any_hit, pt_hit* host_out);

/* Host-only, no GPU and no context: the BVH2 -> 8-wide conversion of one model exactly as pt_upload_scene's host converter does it (the device
 * converter produces the same bytes).  No counterpart in the reference (its traversal walks the BVH2).  nodes (80 B each) / wtris (48 B each)
 * may be NULL to ask for the counts; optimal = 0 selects round 1's largest-area collapse instead of the SAH-optimal one. */
PT_API int  pt_host_wide_bvh(const pt_bvh_node* bvh2, uint32_t n_bvh2, const pt_triangle_v* triangles_global, int optimal,
                             void* nodes, uint32_t* n_nodes, void* wtris, uint32_t* n_tris, double* sah_cost, uint32_t* max_depth);
/* Bounds-checked build only (make -C cuda_pathtracer_b200/csrc variant NAME=_chk DEFS=-DPT_BOUNDS_CHECK; compute-sanitizer is not
 * available on the GPU pool): every index the kernels form is compared with its array's extent; out4 = failures since the last
 * report, code / index / limit of the first.  Other builds answer PT_ERR_INVALID.  pt_version() names the build. */
PT_API int  pt_debug_check_report(pt_ctx*, uint32_t* out4);

PT_API void* pt_cuda_stream(pt_ctx*);

/* Tile partition, host-callable without a GPU: 32x32-pixel tiles dealt round-robin to
 * ranks; a rank's local slot = (local tile, 8x4 block, lane).  Returns 1 (pixel), 0 (slot off the image). */
PT_API uint32_t pt_tile_num_slots(const pt_config* cfg);
PT_API int  pt_tile_slot_to_pixel(const pt_config* cfg, uint32_t slot, uint32_t* x, uint32_t* y);   /* the stream every launch of this context goes to */

#ifdef __cplusplus
}
#endif
#endif /* PTB200_H */
