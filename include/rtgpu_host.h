/*
 * rtgpu_host.h — host front-end of the rtgpu backend: the reference's one-shot set-up
 * code that stays on the CPU (flags, scene DSL, OBJ loader, built-in scenes, camera
 * set-up). It produces the flat tables of rtgpu.h. No CUDA is touched here, so these
 * entry points work on a machine without a GPU.
 *
 * Mirrors (file:line in the reference):
 *   rtgpu_config_from_args   Config::from_args            src/config.rs:62-176
 *   rtgpu_scene_load         main.rs scene dispatch        src/main.rs:30-59
 *                            SceneLoader::load             src/loaders/scene.rs:80-156
 *                            load_mesh_from_file           src/loaders/obj.rs:13-107
 *                            GoldenMonkeyScene::init       src/scene/golden_monkey.rs:23-138
 *   rtgpu_camera_init        Camera::new + init            src/camera.rs:47-130
 */
#ifndef RTGPU_HOST_H
#define RTGPU_HOST_H

#include "rtgpu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* SceneConfig (config.rs:9-18): has_* mirrors Option<>. */
typedef struct RtgpuSceneConfig {
    int32_t has_output_width, output_width;
    int32_t has_aspect_ratio, has_focal_length, has_f_number, has_focus_distance;
    int32_t has_camera_pos, has_camera_target, has_background;
    double aspect_ratio, focal_length, f_number, focus_distance;
    double camera_pos[3], camera_target[3], background[3];
} RtgpuSceneConfig;

/* Config (config.rs:46-59) + the new backend switches. */
typedef struct RtgpuConfig {
    RtgpuSceneConfig scene;
    int32_t thread_count;            /* -t, default 1   */
    int32_t samples_per_pixel;       /* -s, default 250 (upper bound, README.md:43-49) */
    int32_t sqrt_samples_per_thread; /* floor(sqrt(floor(s / t)))                      */
    int32_t max_depth;               /* --max-depth, default 20                         */
    double light_bias;               /* --light-bias, default 0.25, must be in [0,1]    */
    char scene_name[1024];           /* positional argument ("" = default scene)        */
    /* additions of this backend (a bare `--gpu` has no '=' and is ignored by the
     * reference's regex, config.rs:63; it is recognised explicitly here)              */
    int32_t use_gpu;    /* --gpu      */
    int32_t n_gpus;     /* --gpus=N   */
    uint64_t seed;      /* --seed=N   sampling seed (default 1)            */
    uint64_t scene_seed;/* --scene-seed=N seed of the scene-construction RNG (default 42) */
    int32_t fixed;      /* --fixed    (bare flag) render with the biased light samplers corrected: RTGPU_FLAG_FIXED */
    int32_t reserved;
} RtgpuConfig;

/* Config::from_args (config.rs:62). argv[0] is skipped like `args.skip(1)`.
 * Returns RTGPU_ERR_PARSE where the reference would panic (`expect`, `assert!`).       */
int rtgpu_config_from_args(int argc, const char* const* argv, RtgpuConfig* out);

/* Image decoding (buffer.rs:30-48 `image::open(path)?.decode()?.into_rgb32f()`, up to the
 * integer texels). The front-end decodes PNG (zlib) and sequential Huffman JPEG (every .jpg
 * of the reference tree) itself. A registered plug-in takes over every format but PNG — the
 * Python front-end registers PIL — and is the only way to load what the built-in decoders
 * refuse (progressive JPEG, other formats). It must fill *out with malloc()ed data (freed
 * with free()) in one of the RtgpuImageFormat layouts and return 0; fn = NULL unregisters.  */
typedef int (*RtgpuImageDecodeFn)(const char* path, RtgpuImage* out, void* user);
int rtgpu_set_image_decoder(RtgpuImageDecodeFn fn, void* user);
/* Decode one file the way an `image <file>` texture does (texture/image.rs:21-33);
 * builtin_only != 0 bypasses the plug-in. out->data is malloc()ed: free() it.              */
int rtgpu_decode_image(const char* path, int builtin_only, RtgpuImage* out);

typedef struct RtgpuLoadedScene RtgpuLoadedScene; /* owns every array the desc points to */

/* main.rs:30-59: "" | "golden_monkey" and the other built-in names, or a DSL file path.
 * Built-in scenes open assets relative to `asset_root` (the reference uses the CWD).
 * The merged Config (scene defaults <- CLI, config.rs:32-43) is written back to *cfg.    */
int rtgpu_scene_load(RtgpuConfig* cfg, const char* asset_root, RtgpuLoadedScene** out);
int rtgpu_scene_free(RtgpuLoadedScene* scene);
const RtgpuSceneDesc* rtgpu_loaded_scene_desc(const RtgpuLoadedScene* scene);
/* warnings the reference prints and skips (loaders/scene.rs:127-135), '\n'-separated */
const char* rtgpu_loaded_scene_warnings(const RtgpuLoadedScene* scene);

/* Camera::new + init (camera.rs:47-130) from a merged Config. */
int rtgpu_camera_init(const RtgpuConfig* cfg, RtgpuCameraDesc* out_cam, RtgpuRenderParams* out_params);

/* Writer::save (output.rs:23-39) with the ACES tonemap main.rs:81 selects; tonemap
 * arithmetic runs on the GPU (rtgpu_tonemap_srgb8), PNG encoding on the host.          */
int rtgpu_save_png(const char* path, const uint8_t* rgb, int width, int height);

/* Deterministic ~1M-triangle stand-in for the missing resource/dragon_high.obj
 * (.MISSING_LARGE_BLOBS): `levels` rounds of 1->4 subdivision of an OBJ mesh with a
 * smooth procedural displacement, written as an OBJ file. target_width > 0 rescales the
 * result to that x-extent, centres it in x/z and rests it on y = 0 (the dragon fills about
 * 5.8 object units of the box in samples/sample2.png).                                  */
int rtgpu_make_proxy_mesh(const char* src_obj, const char* dst_obj, int levels, double displacement,
                          double target_width);

#ifdef __cplusplus
}
#endif
#endif
