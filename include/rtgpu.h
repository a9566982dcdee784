/*
 * rtgpu.h — C ABI of the B200-native path-tracing backend for rust_raytracer.
 *
 * This is the drop-in boundary for the reference's ONE hot path,
 *     camera.render(world, lights, &mut buf)            (reference src/main.rs:74-75,
 *                                                        src/camera.rs:189-256)
 * A host (the Rust crate through `extern "C"`, the C++/Python harness in this repo)
 * describes the scene as flat POD tables, uploads it once, and asks for a render into
 * a caller-owned f64 RGBA buffer laid out exactly like the reference's `Buffer`
 * (src/buffer.rs:8-12,54-60: row-major, top row first, Vec4 = 4 x f64 per pixel).
 *
 * Conventions
 *  - plain pointers and sizes only; every array is owned by the caller and copied by
 *    rtgpu_scene_create(); nothing here throws or aborts across the ABI;
 *  - every function returns 0 on success or a negative RtgpuStatus; the message of the
 *    last failure on the calling thread is rtgpu_last_error();
 *  - all scene numbers are f64 like the reference (Vec4 = [f64;4], src/vec4.rs:10);
 *    the device keeps fp32 copies for traversal and f64 copies for the exact path.
 *  - there is NO CPU fallback: every compute entry point fails with
 *    RTGPU_ERR_NO_DEVICE / RTGPU_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef RTGPU_H
#define RTGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTGPU_ABI_VERSION 1

typedef enum RtgpuStatus {
    RTGPU_OK = 0,
    RTGPU_ERR_INVALID = -1,     /* bad argument / malformed scene description        */
    RTGPU_ERR_NO_DEVICE = -2,   /* no CUDA device (the product path never falls back) */
    RTGPU_ERR_CUDA = -3,        /* a CUDA runtime call or kernel failed               */
    RTGPU_ERR_UNSUPPORTED = -4, /* scene construct outside the accelerated path       */
    RTGPU_ERR_IO = -5,          /* file not found / unreadable (host front-end)       */
    RTGPU_ERR_PARSE = -6        /* scene DSL / OBJ / flag parse error                 */
} RtgpuStatus;

/* ------------------------------------------------------------------------------------
 * Scene description (what a `flatten()` visitor over Arc<dyn Hit> / Material / Sampler
 * emits). Indices are into the sibling arrays; -1 = none.
 * ---------------------------------------------------------------------------------- */

/* Object tree — one entry per reference `Hit` implementor (src/object.rs:107-115). */
typedef enum RtgpuObjectKind {
    RTGPU_OBJ_LIST = 0,      /* ObjectList               src/object/list.rs            */
    RTGPU_OBJ_BVH = 1,       /* BoundingVolumeHierarchyNode  src/object/bvh.rs         */
    RTGPU_OBJ_NULL = 2,      /* NullObject               src/object/bvh/null_obj.rs    */
    RTGPU_OBJ_TRANSFORM = 3, /* Transform                src/object/transform.rs       */
    RTGPU_OBJ_MESH = 4,      /* TriangleMesh             src/object/mesh.rs            */
    RTGPU_OBJ_SPHERE = 5,    /* Sphere                   src/object/sphere.rs          */
    RTGPU_OBJ_PLANE = 6,     /* Plane (finite quad)      src/object/plane.rs           */
    RTGPU_OBJ_VOLUME = 7,    /* Volume (constant medium) src/object/volume.rs          */
    RTGPU_OBJ_SKY = 8,       /* Sky                      src/object/sky.rs             */
    RTGPU_OBJ_SUN = 9        /* Sun                      src/object/sun.rs             */
} RtgpuObjectKind;

enum {
    RTGPU_OBJF_DISABLE_BOUNDS_CHECK = 1u << 0, /* ObjectList.disable_bounds_check list.rs:23 */
    RTGPU_OBJF_BACKFACE = 1u << 1,             /* Plane.render_backface     plane.rs:16      */
    RTGPU_OBJF_FLAT_SHADING = 1u << 2,         /* TriangleMesh.flat_shading mesh.rs:25       */
    RTGPU_OBJF_HIT_BACK_FACES = 1u << 3        /* TriangleMesh.hit_back_faces mesh.rs:26     */
};

typedef struct RtgpuObject {
    int32_t kind;        /* RtgpuObjectKind                                               */
    int32_t material;    /* MESH, SPHERE, PLANE, VOLUME: index into materials             */
    int32_t texture;     /* SKY, SUN: emission colour texture (Emissive::new(emission_map)) */
    int32_t mesh;        /* MESH: index into meshes                                        */
    int32_t first_child; /* LIST / BVH(2) / TRANSFORM(1) / VOLUME(1 = boundary): offset into `children` */
    int32_t n_children;
    uint32_t flags;      /* RTGPU_OBJF_*                                                   */
    int32_t reserved;
    /* SPHERE: p0 = centre (w = 1), p1[0] = radius (may be negative, golden_monkey.rs:113)
     * PLANE : p0 = centre, p1 = u half-extent, p2 = v half-extent   (plane.rs:29-62)
     * SUN   : p0 = direction AFTER to_unit(), all four components kept (sun.rs:27;
     *         the DSL builds it with Vec4::point so w = 1/len, loaders/scene.rs:843)
     * VOLUME: p1[0] = density                                       (volume.rs:22-29) */
    double p0[4];
    double p1[4];
    double p2[4];
    /* TRANSFORM: row-major 4x4 like Mat4 (mat4.rs:10); inv_transform is the matrix the
     * reference accumulates separately (transform.rs:44-92), NOT recomputed.           */
    double transform[16];
    double inv_transform[16];
} RtgpuObject;

/* TriangleMesh storage (mesh.rs:16-35): three independently indexed attribute arrays. */
typedef struct RtgpuMesh {
    const double* positions; /* n_positions x 3 (w = 1 implied)                          */
    const double* normals;   /* n_normals   x 3 (already normalised on load, obj.rs:48) */
    const double* uvs;       /* n_uvs       x 2 (may be NULL when n_uvs == 0)            */
    const int32_t* tri_pos;  /* n_triangles x 3                                          */
    const int32_t* tri_nrm;  /* n_triangles x 3                                          */
    const int32_t* tri_uv;   /* n_triangles x 3, all -1 for a triangle without uvs; may be NULL */
    int32_t n_positions, n_normals, n_uvs, n_triangles;
} RtgpuMesh;

typedef enum RtgpuMaterialKind {
    RTGPU_MAT_LAMBERTIAN = 0,  /* material/lambertian.rs   */
    RTGPU_MAT_METAL = 1,       /* material/metal.rs        */
    RTGPU_MAT_DIELECTRIC = 2,  /* material/dielectric.rs   */
    RTGPU_MAT_GLOSSY = 3,      /* material/glossy.rs       */
    RTGPU_MAT_EMISSIVE = 4,    /* material/emissive.rs     */
    RTGPU_MAT_ISOTROPIC = 5,   /* material/isotropic.rs    */
    RTGPU_MAT_NORMAL_DEBUG = 6 /* material/normal_debug.rs */
} RtgpuMaterialKind;

typedef struct RtgpuMaterial {
    int32_t kind;
    int32_t albedo;     /* colour texture: albedo (lambertian/metal/glossy/isotropic), emission map (emissive) */
    int32_t roughness;  /* float texture (metal, glossy)                                 */
    int32_t normal_map; /* colour texture or -1 (glossy, normal_debug)                   */
    double ior;         /* dielectric: index of refraction; glossy: ior (inv_ior = 1/ior) */
} RtgpuMaterial;

/* Texture expression nodes — one per reference `Sampler` implementor (src/texture/ directory).
 * A node yields a colour or a float depending on `is_float`.                            */
typedef enum RtgpuTextureKind {
    RTGPU_TEX_CONSTANT = 0,      /* value[0..2] (colour) or value[0] (float)  constant.rs   */
    RTGPU_TEX_CHECKER = 1,       /* a = even, b = odd, scale              checkerboard.rs:34 */
    RTGPU_TEX_CHECKER_SOLID = 2, /* a = even, b = odd, scale              checkerboard.rs:74 */
    RTGPU_TEX_IMAGE = 3,         /* image, repeat                         image.rs:40-53     */
    RTGPU_TEX_LERP = 4,          /* a = start, b = end, c = t (float)     interpolate.rs     */
    RTGPU_TEX_NOISE_SOLID = 5,   /* noise, scale, samples (float result)  noise.rs           */
    RTGPU_TEX_CHANNEL = 6,       /* a = colour texture, channel (float result) channel.rs    */
    RTGPU_TEX_UV_DEBUG = 7       /* (u, v, 0.5)                           uv_debug.rs        */
} RtgpuTextureKind;

typedef struct RtgpuTexture {
    int32_t kind;
    int32_t is_float;
    int32_t a, b, c;
    int32_t image;
    int32_t noise;
    int32_t channel;
    int32_t samples;  /* noise_solid octaves (default 7)      */
    int32_t repeat;   /* image: 1 = Repeat (default), 0 = Clamp */
    double scale;     /* checker / checker_solid / noise_solid  */
    double value[4];
} RtgpuTexture;

typedef enum RtgpuImageFormat {
    RTGPU_IMG_RGB8 = 0,   /* 3 x u8  per texel, value = byte / 255   (buffer.rs:31 into_rgb32f) */
    RTGPU_IMG_RGB16 = 1,  /* 3 x u16 per texel, value = word / 65535                         */
    RTGPU_IMG_RGBF32 = 2  /* 3 x f32 per texel                                               */
} RtgpuImageFormat;

typedef struct RtgpuImage {
    int32_t width, height;
    int32_t format; /* RtgpuImageFormat */
    int32_t reserved;
    const void* data; /* row-major, top row first, tightly packed RGB */
} RtgpuImage;

/* PerlinNoise3D tables (noise/perlin.rs:11-52). */
typedef struct RtgpuNoise {
    double gradients[256][3];
    int32_t perm_x[256], perm_y[256], perm_z[256];
} RtgpuNoise;

typedef struct RtgpuSceneDesc {
    uint32_t abi_version; /* RTGPU_ABI_VERSION */
    int32_t world_root;   /* object index of `world`  (scene.rs:30 SceneData.1) */
    int32_t lights_root;  /* object index of `lights` (SceneData.2)             */
    int32_t n_objects, n_children, n_meshes, n_materials, n_textures, n_images, n_noises;
    const RtgpuObject* objects;
    const int32_t* children;
    const RtgpuMesh* meshes;
    const RtgpuMaterial* materials;
    const RtgpuTexture* textures;
    const RtgpuImage* images;
    const RtgpuNoise* noises;
} RtgpuSceneDesc;

/* Camera state after Camera::init (src/camera.rs:86-130). */
typedef struct RtgpuCameraDesc {
    int32_t width, height;
    double position[3];
    double first_pixel[3];
    double pixel_delta_u[3], pixel_delta_v[3];
    double basis_u[3], basis_v[3];
    double aperture_radius; /* < 0 : no depth of field (aperture_radius == None) */
    double background[3];
} RtgpuCameraDesc;

enum {
    /* Exact mode: every primitive test of the closest-hit search runs in f64, in object
     * space, with the reference's operation order (used by rtgpu_trace_primary always). */
    RTGPU_FLAG_EXACT = 1u << 0,
    /* Use the guarded estimator only (paths with pdf <= 0 or non-finite weight are
     * terminated). Always on; bit kept for ABI symmetry.       */
    RTGPU_FLAG_GUARD_PDF = 1u << 1,
    /* Record CUDA events around every kernel launch of the render; rtgpu_get_stats then
     * reports the device time per kernel family (ms_raygen / ms_trace / ms_shade).          */
    RTGPU_FLAG_PROFILE = 1u << 2,
    /* `--fixed`: render with the two biased light samplers of the reference corrected (Plane::random covers the whole quad
     * instead of one quadrant, plane.rs:120-126; Sphere::random uses an orthonormal basis, sphere.rs:123-128; SURVEY.md
     * Appendix A Q1, Q2). Off by default: parity is defined against the reference, quirks included.                     */
    RTGPU_FLAG_FIXED = 1u << 3
};

typedef struct RtgpuRenderParams {
    int32_t replicas;     /* T: the reference's thread_count (config.rs:74,119)             */
    int32_t sqrt_spt;     /* k: sqrt_samples_per_thread (config.rs:154-155); spp = T*k*k    */
    int32_t max_depth;    /* config.rs:76                                                   */
    int32_t n_gpus;       /* devices to shard the sample strata over (<= rtgpu_init's list) */
    double light_bias;    /* config.rs:77                                                   */
    uint64_t seed;        /* Philox key; streams are (pixel, sample, bounce)-addressed      */
    uint32_t flags;       /* RTGPU_FLAG_*                                                   */
    int32_t sample_begin; /* shard: render global samples [sample_begin, sample_end)        */
    int32_t sample_end;   /*        0,0 = all T*k*k samples                                 */
    int32_t reserved;
} RtgpuRenderParams;

typedef struct RtgpuScene RtgpuScene; /* opaque: per-device uploaded scene */

/* First-hit record of the parity hook (HitRecord, src/object.rs:32-41). */
typedef struct RtgpuHit {
    int32_t object;     /* index into RtgpuSceneDesc.objects of the leaf hit, -1 = miss       */
    int32_t instance;   /* ordinal of that leaf in a DFS pre-order walk from world_root      */
    int32_t primitive;  /* triangle index inside the mesh, 0 otherwise                       */
    int32_t front_face; /* HitRecord.front_face                                              */
    double t;
    double pos[3];
    double normal[3];
    double uv[2];
} RtgpuHit;

typedef struct RtgpuStats {
    uint64_t paths;          /* camera paths started                                     */
    uint64_t rays;           /* world.test-equivalent closest-hit queries                */
    uint64_t rays_per_bounce[64];
    double ms_total;         /* device time of the last render (CUDA events)             */
    double ms_raygen, ms_trace, ms_shade, ms_resolve;
    uint64_t kernel_launches;
    uint64_t bvh_nodes, bvh_triangles, bvh_prims;
    double ms_build;         /* host time of flatten + BVH build in rtgpu_scene_create   */
    uint64_t rays_traversed; /* of `rays`, those that walked the BVH in the trace kernel; the others
                                were resolved by the broadcast pre-test where the ray was made */
    uint64_t rays_traversed_per_bounce[64];
    uint64_t trace_launches; /* launches of the traversal kernel so far (the kernel `roofline` in bench.py is quoted on) */
    uint64_t pool_bytes;     /* device memory the path pools of all lanes and devices hold now                         */
    uint32_t lanes, batch_paths, tail_group, tail_bounce; /* schedule of the last render (api.cu: enqueue_render)      */
    double ms_build_device;  /* RTGPU_BUILD=gpu: device time of the acceleration-structure build (0 = built on the host) */
} RtgpuStats;

/* ------------------------------------------------------------------------------------
 * Entry points
 * ---------------------------------------------------------------------------------- */

/* Select the CUDA devices to use (NULL, 0 = device 0 only). Replaces nothing in the
 * reference (single process, no device); called once before rtgpu_scene_create.       */
int rtgpu_init(const int* devices, int n_devices);
/* Same, but returns at once: driver initialisation and context creation (~1.2 s + ~0.2 s per device
 * on the B200 box) run on a helper thread while the caller parses flags, loads the scene and while
 * rtgpu_scene_create flattens it and builds the BVH (the host half of main.rs's "Ready" phase,
 * main.rs:28-61). Errors (no device, bad index) surface from the next call that needs a device.   */
int rtgpu_init_async(const int* devices, int n_devices);
int rtgpu_shutdown(void);
int rtgpu_device_count(void);
const char* rtgpu_last_error(void);

/* Build the wide BVH + tables and upload them to every selected device.
 * Replaces the scene ownership hand-over `Arc<dyn Hit>` world/lights of
 * Camera::render (src/camera.rs:189).                                                 */
int rtgpu_scene_create(const RtgpuSceneDesc* desc, RtgpuScene** out_scene);
int rtgpu_scene_destroy(RtgpuScene* scene);

/* Replaces Camera::render (src/camera.rs:189-256). ADDS the per-pixel mean radiance
 * (sum over all T*k*k samples / spp) into out_rgba[(y*W + x)*4 + {0,1,2}], leaves [3]
 * untouched-plus-zero, exactly like `buf.set_pixel(x, y, local + thread_color)`
 * (camera.rs:247-253). Host buffer; H2D/D2H copies are inside this call.               */
int rtgpu_render(RtgpuScene* scene, const RtgpuCameraDesc* cam,
                 const RtgpuRenderParams* params, double* out_rgba);

/* Same render, device-resident result for hosts that own device memory (torch / NCCL):
 * accumulates the UN-normalised fp32 RGBA radiance sums of this shard into
 * d_accum_rgba (device pointer on the current device of `device_index`, W*H*4 floats),
 * enqueued on `stream` (a cudaStream_t, may be NULL).                                  */
int rtgpu_render_device(RtgpuScene* scene, int device_index, const RtgpuCameraDesc* cam,
                        const RtgpuRenderParams* params, float* d_accum_rgba, void* stream);

/* K5: sums / spp -> f64 Buffer layout (camera.rs:229) on the device, then ADDED into the host
 * buffer out_rgba like Camera::render does (camera.rs:247-253).                          */
int rtgpu_resolve_device(int device_index, const float* d_accum_rgba, int width, int height,
                         int total_spp, double* out_rgba, void* stream);

/* Replaces Writer::save's per-pixel maths (src/output.rs:23-50): tonemap
 * (0 = clamp tonemapping/clamp.rs, 1 = ACES tonemapping/aces.rs:27-33) -> sRGB OETF ->
 * trunc(x*255.999). rgba: host, W*H*4 f64; out_rgb: host, W*H*3 bytes.                  */
int rtgpu_tonemap_srgb8(const double* rgba, int width, int height, int tonemap, uint8_t* out_rgb);

/* rtgpu_tonemap_srgb8 for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same kernel and
 * output; the device staging (f64 RGBA in, RGB8 out) is a file-scope static device array of
 * RTGPU_NOALLOC_MAX_PIXELS x 35 bytes (8.75 MiB, held by every GPU that loads the library until exit). Calls no
 * cudaMalloc / cudaFree. Images above the maximum are rejected with RTGPU_ERR_INVALID. All calls share the array (it
 * keeps the last call's contents; nothing in it needs clearing): calls that overlap on one GPU race — serialise them.  */
#define RTGPU_NOALLOC_MAX_PIXELS 262144 /* e.g. 512 x 512 */
int rtgpu_tonemap_srgb8_noalloc(const double* rgba, int width, int height, int tonemap, uint8_t* out_rgb);

/* Camera::render followed by Writer::save's per-pixel maths (main.rs:74-82, output.rs:23-50) WITHOUT leaving the
 * device: the fp32 sums of all devices -> / spp -> tonemap (0 clamp, 1 ACES) -> sRGB OETF -> trunc(x*255.999), then one
 * W*H*3-byte copy to out_rgb (host) instead of the W*H*32-byte f64 Buffer down and up again. Same pixels as
 * rtgpu_render into a zeroed buffer followed by rtgpu_tonemap_srgb8.                                               */
int rtgpu_render_srgb8(RtgpuScene* scene, const RtgpuCameraDesc* cam, const RtgpuRenderParams* params,
                       int tonemap, uint8_t* out_rgb);

/* Parity hook: closest hit of `n` rays against `world` with t in (0.001, +inf)
 * (camera.rs:294), always in exact mode. rays = n x 6 f64 (origin xyz, direction xyz).
 * Stochastic primitives (Volume) are skipped by this hook.                             */
int rtgpu_trace_primary(RtgpuScene* scene, const double* rays, int64_t n, RtgpuHit* out_hits);

/* Parity hook for the PRODUCTION traversal kernel (fp32 slab tests, fp32 world-space triangles
 * and quads, dynamic ray fetch): closest surface hit of `n` rays with t in (0.001, +inf) as the
 * render loop sees it. instance / primitive use the numbering of RtgpuHit; t is the value the
 * shading stage uses: fp32 for triangles, and for quads / spheres (which the fp32 traversal only
 * selects) the f64 re-evaluation the shade kernels apply. Volumes, sky and sun are not resolved
 * (instance = -1 on a miss).                                                               */
typedef struct RtgpuFastHit {
    int32_t instance, primitive;
    double t;
    float u, v;
} RtgpuFastHit;
int rtgpu_trace_fast(RtgpuScene* scene, const double* rays, int64_t n, RtgpuFastHit* out_hits);

/* Parity hook: the camera rays Camera::get_ray (camera.rs:260-280) generates for the
 * given pixels and injected random numbers. px: n x 2 int32 (x, y); sxy: n x 2 int32
 * stratum cell; xi: n x 4 f64 (jitter rx, ry, lens nx, ny — nx, ny are N(0,1) draws,
 * vec4.rs:35-40). out_rays: n x 6 f64.                                                  */
int rtgpu_generate_rays(const RtgpuCameraDesc* cam, int sqrt_spt, const int32_t* px,
                        const int32_t* sxy, const double* xi, int64_t n, double* out_rays);

/* Parity hook: ONE bounce of ray_color (camera.rs:294-328) for n rays with injected random
 * numbers (xi: n x n_xi f64, consumed in the reference's draw order; uniforms and N(0,1)
 * draws share the stream). kind: 0 miss, 1 ScatteredWithPDF, 2 ScatteredWithRay, 3 Absorbed,
 * 4 Emissive; weight = attenuation * scattering_pdf / pdf (kind 1) or attenuation (kind 2). */
typedef struct RtgpuBounce {
    int32_t kind, reserved;
    double emission[3], weight[3], next_origin[3], next_dir[3], pdf, scattering_pdf;
} RtgpuBounce;
int rtgpu_bounce(RtgpuScene* scene, double light_bias, const double* rays, const double* xi, int n_xi,
                 int64_t n, RtgpuBounce* out);

int rtgpu_get_stats(RtgpuScene* scene, RtgpuStats* out_stats);

#ifdef __cplusplus
}
#endif
#endif /* RTGPU_H */
