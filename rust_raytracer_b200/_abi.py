"""ctypes mirror of include/rtgpu.h and include/rtgpu_host.h (field for field).

This is the reference-side binding a Python host would write; the Rust `#[repr(C)]`
equivalent is shown in INTEGRATION.md. Nothing here computes anything.
"""
import ctypes as C

ABI_VERSION = 1

# status codes (RtgpuStatus)
OK, ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_UNSUPPORTED, ERR_IO, ERR_PARSE = 0, -1, -2, -3, -4, -5, -6

# object kinds
OBJ_LIST, OBJ_BVH, OBJ_NULL, OBJ_TRANSFORM, OBJ_MESH, OBJ_SPHERE, OBJ_PLANE, OBJ_VOLUME, OBJ_SKY, OBJ_SUN = range(10)
OBJF_DISABLE_BOUNDS_CHECK, OBJF_BACKFACE, OBJF_FLAT_SHADING, OBJF_HIT_BACK_FACES = 1, 2, 4, 8
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_GLOSSY, MAT_EMISSIVE, MAT_ISOTROPIC, MAT_NORMAL_DEBUG = range(7)
(TEX_CONSTANT, TEX_CHECKER, TEX_CHECKER_SOLID, TEX_IMAGE, TEX_LERP, TEX_NOISE_SOLID, TEX_CHANNEL,
 TEX_UV_DEBUG) = range(8)
IMG_RGB8, IMG_RGB16, IMG_RGBF32 = range(3)
FLAG_EXACT, FLAG_GUARD_PDF, FLAG_PROFILE, FLAG_FIXED = 1, 2, 4, 8
NOALLOC_MAX_PIXELS = 262144


class Object(C.Structure):
    _fields_ = [("kind", C.c_int32), ("material", C.c_int32), ("texture", C.c_int32), ("mesh", C.c_int32),
                ("first_child", C.c_int32), ("n_children", C.c_int32), ("flags", C.c_uint32),
                ("reserved", C.c_int32),
                ("p0", C.c_double * 4), ("p1", C.c_double * 4), ("p2", C.c_double * 4),
                ("transform", C.c_double * 16), ("inv_transform", C.c_double * 16)]


class Mesh(C.Structure):
    _fields_ = [("positions", C.POINTER(C.c_double)), ("normals", C.POINTER(C.c_double)),
                ("uvs", C.POINTER(C.c_double)), ("tri_pos", C.POINTER(C.c_int32)),
                ("tri_nrm", C.POINTER(C.c_int32)), ("tri_uv", C.POINTER(C.c_int32)),
                ("n_positions", C.c_int32), ("n_normals", C.c_int32), ("n_uvs", C.c_int32),
                ("n_triangles", C.c_int32)]


class Material(C.Structure):
    _fields_ = [("kind", C.c_int32), ("albedo", C.c_int32), ("roughness", C.c_int32),
                ("normal_map", C.c_int32), ("ior", C.c_double)]


class Texture(C.Structure):
    _fields_ = [("kind", C.c_int32), ("is_float", C.c_int32), ("a", C.c_int32), ("b", C.c_int32),
                ("c", C.c_int32), ("image", C.c_int32), ("noise", C.c_int32), ("channel", C.c_int32),
                ("samples", C.c_int32), ("repeat", C.c_int32), ("scale", C.c_double),
                ("value", C.c_double * 4)]


class Image(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("format", C.c_int32),
                ("reserved", C.c_int32), ("data", C.c_void_p)]


class Noise(C.Structure):
    _fields_ = [("gradients", (C.c_double * 3) * 256), ("perm_x", C.c_int32 * 256),
                ("perm_y", C.c_int32 * 256), ("perm_z", C.c_int32 * 256)]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("world_root", C.c_int32), ("lights_root", C.c_int32),
                ("n_objects", C.c_int32), ("n_children", C.c_int32), ("n_meshes", C.c_int32),
                ("n_materials", C.c_int32), ("n_textures", C.c_int32), ("n_images", C.c_int32),
                ("n_noises", C.c_int32),
                ("objects", C.POINTER(Object)), ("children", C.POINTER(C.c_int32)),
                ("meshes", C.POINTER(Mesh)), ("materials", C.POINTER(Material)),
                ("textures", C.POINTER(Texture)), ("images", C.POINTER(Image)),
                ("noises", C.POINTER(Noise))]


class CameraDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32),
                ("position", C.c_double * 3), ("first_pixel", C.c_double * 3),
                ("pixel_delta_u", C.c_double * 3), ("pixel_delta_v", C.c_double * 3),
                ("basis_u", C.c_double * 3), ("basis_v", C.c_double * 3),
                ("aperture_radius", C.c_double), ("background", C.c_double * 3)]


class RenderParams(C.Structure):
    _fields_ = [("replicas", C.c_int32), ("sqrt_spt", C.c_int32), ("max_depth", C.c_int32),
                ("n_gpus", C.c_int32), ("light_bias", C.c_double), ("seed", C.c_uint64),
                ("flags", C.c_uint32), ("sample_begin", C.c_int32), ("sample_end", C.c_int32),
                ("reserved", C.c_int32)]


class FastHit(C.Structure):
    _fields_ = [("instance", C.c_int32), ("primitive", C.c_int32), ("t", C.c_double), ("u", C.c_float), ("v", C.c_float)]


class Hit(C.Structure):
    _fields_ = [("object", C.c_int32), ("instance", C.c_int32), ("primitive", C.c_int32),
                ("front_face", C.c_int32), ("t", C.c_double), ("pos", C.c_double * 3),
                ("normal", C.c_double * 3), ("uv", C.c_double * 2)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("rays_per_bounce", C.c_uint64 * 64),
                ("ms_total", C.c_double), ("ms_raygen", C.c_double), ("ms_trace", C.c_double),
                ("ms_shade", C.c_double), ("ms_resolve", C.c_double), ("kernel_launches", C.c_uint64),
                ("bvh_nodes", C.c_uint64), ("bvh_triangles", C.c_uint64), ("bvh_prims", C.c_uint64),
                ("ms_build", C.c_double), ("rays_traversed", C.c_uint64), ("rays_traversed_per_bounce", C.c_uint64 * 64),
                ("trace_launches", C.c_uint64), ("pool_bytes", C.c_uint64), ("lanes", C.c_uint32), ("batch_paths", C.c_uint32),
                ("tail_group", C.c_uint32), ("tail_bounce", C.c_uint32), ("ms_build_device", C.c_double)]


class Bounce(C.Structure):
    _fields_ = [("kind", C.c_int32), ("reserved", C.c_int32), ("emission", C.c_double * 3),
                ("weight", C.c_double * 3), ("next_origin", C.c_double * 3), ("next_dir", C.c_double * 3),
                ("pdf", C.c_double), ("scattering_pdf", C.c_double)]


class SceneConfig(C.Structure):
    _fields_ = [("has_output_width", C.c_int32), ("output_width", C.c_int32),
                ("has_aspect_ratio", C.c_int32), ("has_focal_length", C.c_int32),
                ("has_f_number", C.c_int32), ("has_focus_distance", C.c_int32),
                ("has_camera_pos", C.c_int32), ("has_camera_target", C.c_int32),
                ("has_background", C.c_int32),
                ("aspect_ratio", C.c_double), ("focal_length", C.c_double), ("f_number", C.c_double),
                ("focus_distance", C.c_double), ("camera_pos", C.c_double * 3),
                ("camera_target", C.c_double * 3), ("background", C.c_double * 3)]


class Config(C.Structure):
    _fields_ = [("scene", SceneConfig), ("thread_count", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("sqrt_samples_per_thread", C.c_int32), ("max_depth", C.c_int32),
                ("light_bias", C.c_double), ("scene_name", C.c_char * 1024),
                ("use_gpu", C.c_int32), ("n_gpus", C.c_int32), ("seed", C.c_uint64),
                ("scene_seed", C.c_uint64), ("fixed", C.c_int32), ("reserved", C.c_int32)]


IMAGE_DECODE_FN = C.CFUNCTYPE(C.c_int, C.c_char_p, C.POINTER(Image), C.c_void_p)

# every symbol the headers declare: name -> (restype, argtypes)
HOST_SYMBOLS = {
    "rtgpu_last_error": (C.c_char_p, []),
    "rtgpu_config_from_args": (C.c_int, [C.c_int, C.POINTER(C.c_char_p), C.POINTER(Config)]),
    "rtgpu_set_image_decoder": (C.c_int, [IMAGE_DECODE_FN, C.c_void_p]),
    "rtgpu_decode_image": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(Image)]),
    "rtgpu_scene_load": (C.c_int, [C.POINTER(Config), C.c_char_p, C.POINTER(C.c_void_p)]),
    "rtgpu_scene_free": (C.c_int, [C.c_void_p]),
    "rtgpu_loaded_scene_desc": (C.POINTER(SceneDesc), [C.c_void_p]),
    "rtgpu_loaded_scene_warnings": (C.c_char_p, [C.c_void_p]),
    "rtgpu_camera_init": (C.c_int, [C.POINTER(Config), C.POINTER(CameraDesc), C.POINTER(RenderParams)]),
    "rtgpu_save_png": (C.c_int, [C.c_char_p, C.POINTER(C.c_uint8), C.c_int, C.c_int]),
    "rtgpu_make_proxy_mesh": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, C.c_double, C.c_double]),
}
DEVICE_SYMBOLS = {
    "rtgpu_init": (C.c_int, [C.POINTER(C.c_int), C.c_int]),
    "rtgpu_init_async": (C.c_int, [C.POINTER(C.c_int), C.c_int]),
    "rtgpu_shutdown": (C.c_int, []),
    "rtgpu_device_count": (C.c_int, []),
    "rtgpu_scene_create": (C.c_int, [C.POINTER(SceneDesc), C.POINTER(C.c_void_p)]),
    "rtgpu_scene_destroy": (C.c_int, [C.c_void_p]),
    "rtgpu_render": (C.c_int, [C.c_void_p, C.POINTER(CameraDesc), C.POINTER(RenderParams), C.POINTER(C.c_double)]),
    "rtgpu_render_device": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(CameraDesc), C.POINTER(RenderParams), C.c_void_p, C.c_void_p]),
    "rtgpu_resolve_device": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.c_void_p]),
    "rtgpu_render_srgb8": (C.c_int, [C.c_void_p, C.POINTER(CameraDesc), C.POINTER(RenderParams), C.c_int, C.POINTER(C.c_uint8)]),
    "rtgpu_tonemap_srgb8_noalloc": (C.c_int, [C.POINTER(C.c_double), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint8)]),
    "rtgpu_tonemap_srgb8": (C.c_int, [C.POINTER(C.c_double), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint8)]),
    "rtgpu_trace_primary": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.c_int64, C.POINTER(Hit)]),
    "rtgpu_trace_fast": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.c_int64, C.POINTER(FastHit)]),
    "rtgpu_generate_rays": (C.c_int, [C.POINTER(CameraDesc), C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_double), C.c_int64, C.POINTER(C.c_double)]),
    "rtgpu_bounce": (C.c_int, [C.c_void_p, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int, C.c_int64, C.POINTER(Bounce)]),
    "rtgpu_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
}


def bind(lib, table):
    for name, (res, args) in table.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib
