/* oracle.h — TEST INFRASTRUCTURE. C entry points of the CPU oracle (liboracle.so).
 *
 * The oracle is an f64 CPU restatement of the reference's hot path (Camera::render ->
 * ray_color -> Hit::test / Material::scatter / PDF), consuming the same flat scene
 * tables as the product (include/rtgpu.h). Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * PARITY: the reference cannot be built here (no Rust toolchain) and ships no tests or
 * golden vectors; the only pins are the README spp rule and samples/sample{1,2}.png
 * (statistical). RNG streams are unpinned by construction (OS-seeded in the reference).
 */
#ifndef ORACLE_H
#define ORACLE_H
#include "rtgpu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OracleScene OracleScene;

const char* oracle_last_error(void);
int oracle_scene_create(const RtgpuSceneDesc* desc, OracleScene** out);
int oracle_scene_destroy(OracleScene* s);

/* Camera::render (camera.rs:189-256): `replicas` OS threads, each renders the whole
 * image with its own RNG over the k x k strata; thread buffers are summed into
 * out_rgba (W*H*4 f64, added). n_threads <= 0: one OS thread per replica (like the
 * reference); otherwise replicas are distributed over n_threads workers.
 * out_rays (optional): number of world.test calls made by ray_color.                     */
int oracle_render(const OracleScene* s, const RtgpuCameraDesc* cam, const RtgpuRenderParams* params,
                  int n_threads, double* out_rgba, uint64_t* out_rays);

/* Same, restricted to the rows y = row_offset (mod row_step): the bounded sample the timed
 * CPU baseline uses. */
int oracle_render_rows(const OracleScene* s, const RtgpuCameraDesc* cam, const RtgpuRenderParams* params,
                       int n_threads, int row_offset, int row_step, double* out_rgba, uint64_t* out_rays);

/* world.test(ray, Interval(0.001, inf)) for n rays (camera.rs:294). */
int oracle_trace_primary(const OracleScene* s, const double* rays, int64_t n, RtgpuHit* out_hits);

/* Camera::get_ray (camera.rs:260-280) with injected random numbers; see rtgpu_generate_rays. */
int oracle_generate_rays(const RtgpuCameraDesc* cam, int sqrt_spt, const int32_t* px, const int32_t* sxy,
                         const double* xi, int64_t n, double* out_rays);

/* One bounce of ray_color (camera.rs:294-328) for n rays with injected random numbers
 * (xi: n x n_xi f64, consumed in the reference's draw order; uniforms and N(0,1) draws
 * share the stream). Outputs per ray:
 *   kind: 0 miss, 1 ScatteredWithPDF, 2 ScatteredWithRay, 3 Absorbed, 4 Emissive
 *   emission[3], weight[3] (attenuation*scattering_pdf/pdf, or attenuation), next ray[6], pdf_mix */
typedef struct OracleBounce {
    int32_t kind, reserved;
    double emission[3], weight[3], next_origin[3], next_dir[3], pdf, scattering_pdf;
} OracleBounce;
int oracle_bounce(const OracleScene* s, double light_bias, const double* rays, const double* xi, int n_xi,
                  int64_t n, OracleBounce* out);

/* Writer::save per-pixel maths (output.rs:23-50, tonemapping/aces.rs, clamp.rs). */
int oracle_tonemap_srgb8(const double* rgba, int width, int height, int tonemap, uint8_t* out_rgb);

/* known-answer hooks */
int oracle_aabb_test(const double box_min[3], const double box_max[3], const double ray[6], double t_lo, double t_hi);
int oracle_octree_stats(const OracleScene* s, int mesh_index, int64_t out[4]); /* nodes, leaves, leaf entries, max depth */
int oracle_texture_sample(const OracleScene* s, int texture, double u, double v, const double p[3], double out[3]);
double oracle_reflectance(double cos_theta, double ior_ratio);

#ifdef __cplusplus
}
#endif
#endif
