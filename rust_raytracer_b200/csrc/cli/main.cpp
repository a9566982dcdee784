// rust_raytracer — the reference's driver (src/main.rs:25-88) as a native program on top of librtgpu.so:
//
//   rust_raytracer [<scene>] [-w= -r= -s= -a= -f= -t= -d= -c= -l= -b= --max-depth= --light-bias=] --gpu
//
// Same positional scene argument (built-in name or DSL file), same `-k=v` flags and defaults (config.rs:62-176), same three
// progress lines, same output file `out.png`, written through the Writer::save maths (ACES -> sRGB OETF -> trunc(x * 255.999),
// output.rs:23-50). The one call site `camera.render(world, lights, &mut buf)` (main.rs:74-75) is rtgpu_render_srgb8 /
// rtgpu_render. Only the C ABI of include/rtgpu.h and include/rtgpu_host.h is used: this file is also the worked example
// of INTEGRATION.md for a C or C++ host. Textures are decoded by the library's own PNG / JPEG decoders.
//
// Flags the reference does not have:
//   --gpu        accepted for symmetry with the north-star command line; this backend is GPU-only and fails loudly when no
//                CUDA device is present (there is no CPU fallback)
//   --gpus=N     shard the sample strata over N devices of this process
//   --seed=N     Philox key of the sampling streams (the reference seeds from OS entropy, camera.rs:208-209)
//   --fixed      render with the two biased light samplers corrected (RTGPU_FLAG_FIXED)
//   --out=PATH   output file (default out.png like main.rs:23)
//   --preview=N  progressive output: the PNG is rewritten after every N samples per pixel
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include <sys/stat.h>
#include <unistd.h>

#include "rtgpu.h"
#include "rtgpu_host.h"

namespace {

double seconds_since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

bool file_exists(const std::string& p) { struct stat st; return ::stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode); }
bool dir_exists(const std::string& p) { struct stat st; return ::stat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode); }

std::string dir_of(const std::string& p) {
    const size_t s = p.rfind('/');
    return s == std::string::npos ? std::string(".") : p.substr(0, s);
}

[[noreturn]] void die(const std::string& what) {
    const char* e = rtgpu_last_error();
    std::fprintf(stderr, "rust_raytracer: %s%s%s\n", what.c_str(), e && *e ? ": " : "", e && *e ? e : "");
    std::exit(1);
}

// built-in scenes open `scenes/resource/...` below the asset root (the reference: below the working directory)
std::string asset_root() {
    if (const char* e = std::getenv("RTGPU_ASSET_ROOT")) return std::string(e) + (e[0] && e[std::strlen(e) - 1] != '/' ? "/" : "");
    char exe[4096];
    const ssize_t n = ::readlink("/proc/self/exe", exe, sizeof exe - 1);
    if (n > 0) {
        exe[n] = 0;
        const std::string cand = dir_of(dir_of(exe)) + "/assets/";
        if (dir_exists(cand)) return cand;
    }
    return "";
}

// scenes/cornell_dragon names resource/dragon_high.obj, a large blob missing from the reference tree: generate the
// deterministic ~1.0 M-triangle stand-in next to the scene file once (tools/make_dragon_proxy.py's parameters) and say so
void ensure_dragon_proxy(const std::vector<std::string>& args) {
    for (const std::string& a : args) {
        if (a.empty() || a[0] == '-' || !file_exists(a)) continue;
        std::ifstream f(a);
        std::stringstream ss;
        ss << f.rdbuf();
        if (ss.str().find("resource/dragon_high.obj") == std::string::npos) continue;
        const std::string res = dir_of(a) + "/resource", dst = res + "/dragon_high.obj", src = res + "/monkey.obj";
        if (!file_exists(dst) && file_exists(src)) {
            std::fprintf(stderr, "note: %s is missing upstream; generating the proxy mesh (subdivided, displaced monkey.obj)\n", dst.c_str());
            if (rtgpu_make_proxy_mesh(src.c_str(), dst.c_str(), 3, 0.004, 5.8) != 0) die("proxy mesh");
        }
    }
}

} // namespace

int main(int argc, char** argv) {
    std::string out_path = "out.png";
    int preview = 0;
    std::vector<std::string> rest;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        if (a.rfind("--out=", 0) == 0) out_path = a.substr(6);
        else if (a.rfind("--preview=", 0) == 0) preview = std::max(0, std::atoi(a.c_str() + 10));
        else rest.push_back(a);
    }
    ensure_dragon_proxy(rest);
    const auto t0 = std::chrono::steady_clock::now();

    // Config::from_args (main.rs:26); argv[0] is skipped like `args.skip(1)`
    std::vector<const char*> cargv;
    cargv.push_back(argv[0]);
    for (const std::string& a : rest) cargv.push_back(a.c_str());
    RtgpuConfig cfg;
    if (rtgpu_config_from_args((int)cargv.size(), cargv.data(), &cfg) != 0) die("bad arguments");
    const int n_gpus = cfg.n_gpus > 0 ? cfg.n_gpus : 1;

    // CUDA bring-up on a helper thread, overlapped with scene loading and flattening
    std::vector<int> devices;
    for (int d = 0; d < n_gpus; d++) devices.push_back(d);
    rtgpu_init_async(devices.data(), n_gpus);

    RtgpuLoadedScene* loaded = nullptr; // main.rs:30-59: built-in scene or DSL file
    if (rtgpu_scene_load(&cfg, asset_root().c_str(), &loaded) != 0) die("scene");
    if (const char* w = rtgpu_loaded_scene_warnings(loaded)) if (*w) std::fprintf(stderr, "%s\n", w);
    RtgpuCameraDesc cam;
    RtgpuRenderParams prm;
    if (rtgpu_camera_init(&cfg, &cam, &prm) != 0) die("camera");
    RtgpuScene* scene = nullptr; // flatten + acceleration structure + upload: part of "Ready"
    if (rtgpu_scene_create(rtgpu_loaded_scene_desc(loaded), &scene) != 0) die(std::to_string(n_gpus) + " CUDA device(s) requested");
    std::printf("Ready: %.2fs\n", seconds_since(t0));

    const int spp = prm.replicas * prm.sqrt_spt * prm.sqrt_spt;
    std::printf("Rendering: %dx%d @%dspp on %d threads (%d samples/thread) [%d GPU(s)]\n", cam.width, cam.height, spp, prm.replicas,
                prm.sqrt_spt * prm.sqrt_spt, n_gpus);
    std::fflush(stdout);
    prm.n_gpus = n_gpus;
    prm.seed = cfg.seed;
    if (cfg.fixed) prm.flags |= RTGPU_FLAG_FIXED;
    const size_t npix = (size_t)cam.width * cam.height;
    std::vector<uint8_t> rgb(npix * 3);
    if (preview > 0 && preview < spp) {
        // Camera::render ADDS each block's share (sum of its samples / spp, camera.rs:229,247-253) into the same f64 Buffer;
        // after m of spp samples the preview is that buffer scaled by spp / m
        std::vector<double> buf(npix * 4, 0.0), scaled(npix * 4);
        for (int begin = 0; begin < spp; begin += preview) {
            const int end = std::min(spp, begin + preview);
            RtgpuRenderParams blk = prm;
            blk.sample_begin = begin; blk.sample_end = end;
            if (rtgpu_render(scene, &cam, &blk, buf.data()) != 0) die("render");
            const double k = (double)spp / (double)end;
            for (size_t i = 0; i < buf.size(); i++) scaled[i] = buf[i] * k;
            if (rtgpu_tonemap_srgb8(scaled.data(), cam.width, cam.height, 1, rgb.data()) != 0) die("tonemap");
            if (rtgpu_save_png(out_path.c_str(), rgb.data(), cam.width, cam.height) != 0) die("png");
            std::printf("  %d/%d spp: %.2fs\n", end, spp, seconds_since(t0));
            std::fflush(stdout);
        }
    } else {
        // camera.create_buffer + camera.render + Writer::save's maths (main.rs:74-82) in one device-resident call
        if (rtgpu_render_srgb8(scene, &cam, &prm, 1, rgb.data()) != 0) die("render");
    }
    std::printf("Done: %.2fs. Writing output to file...\n", seconds_since(t0));
    if (rtgpu_save_png(out_path.c_str(), rgb.data(), cam.width, cam.height) != 0) die("png");
    RtgpuStats st;
    std::memset(&st, 0, sizeof st);
    rtgpu_get_stats(scene, &st);
    std::printf("Done! Took %.2fs. Goodbye :)  [%.1f Mpaths, %.1f Mrays, %llu kernel launches]\n", seconds_since(t0), (double)st.paths / 1e6,
                (double)st.rays / 1e6, (unsigned long long)st.kernel_launches);
    rtgpu_scene_destroy(scene);
    rtgpu_scene_free(loaded);
    rtgpu_shutdown();
    return 0;
}
