// config.cpp — Config::from_args (reference src/config.rs:62-176), restated.
// The reference matches every argument that starts with '-' against
//   ^-([^=\s]+)=([^=\s]+)$
// (so `--width=5` yields key "-width"), silently ignores unknown keys and arguments
// without '=', and takes the last non-dash argument as the scene name.
#include <cctype>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>

#include "host_scene.h"

namespace rthost {

static bool has_space(const std::string& s) {
    for (char c : s) if (std::isspace((unsigned char)c)) return true;
    return false;
}

// the regex above, by hand: returns false when the argument does not match
static bool split_flag(const std::string& arg, std::string& key, std::string& value) {
    if (arg.size() < 4 || arg[0] != '-') return false;
    size_t eq = arg.find('=');
    if (eq == std::string::npos || eq == 1) return false;
    key = arg.substr(1, eq - 1);
    value = arg.substr(eq + 1);
    if (value.empty() || value.find('=') != std::string::npos) return false;
    if (has_space(key) || has_space(value)) return false;
    return true;
}

static bool parse_usize(const std::string& s, long long& out) { // str::parse::<usize>
    if (s.empty()) return false;
    size_t i = 0;
    if (s[0] == '+') i = 1;
    if (i >= s.size()) return false;
    long long v = 0;
    for (; i < s.size(); i++) {
        if (!std::isdigit((unsigned char)s[i])) return false;
        v = v * 10 + (s[i] - '0');
        if (v > 2000000000LL) return false;
    }
    out = v;
    return true;
}

bool parse_f64(const std::string& s, double& out) { // str::parse::<f64> (no leading/trailing junk)
    if (s.empty() || std::isspace((unsigned char)s[0])) return false;
    char* end = nullptr;
    out = std::strtod(s.c_str(), &end);
    return end && *end == 0;
}

// utils::parse_vec (utils.rs:39-49): 1 = ok, 0 = wrong component count (Err), -1 = panic
int parse_vec3(const std::string& s, double out[3]) {
    double comp[8]; int n = 0;
    size_t start = 0;
    for (;;) {
        size_t comma = s.find(',', start);
        std::string part = s.substr(start, comma == std::string::npos ? std::string::npos : comma - start);
        double v;
        if (!parse_f64(part, v)) return -1;
        if (n < 8) comp[n] = v;
        n++;
        if (comma == std::string::npos) break;
        start = comma + 1;
    }
    if (n != 3) return 0;
    out[0] = comp[0]; out[1] = comp[1]; out[2] = comp[2];
    return 1;
}

} // namespace rthost

using namespace rthost;

extern "C" int rtgpu_config_from_args(int argc, const char* const* argv, RtgpuConfig* out) {
    if (!out) { set_error("rtgpu_config_from_args: null output"); return RTGPU_ERR_INVALID; }
    RtgpuConfig c;
    std::memset(&c, 0, sizeof c);
    c.thread_count = 1; c.samples_per_pixel = 250; c.max_depth = 20; c.light_bias = 0.25; // config.rs:74-77
    c.seed = 1; c.scene_seed = 42; c.n_gpus = 1;
    auto fail = [&](const char* msg) { set_error(msg); return RTGPU_ERR_PARSE; };
    for (int i = 1; i < argc; i++) { // args.skip(1)
        std::string arg = argv[i];
        if (arg.rfind("-", 0) == 0) {
            if (arg == "--gpu") { c.use_gpu = 1; continue; } // new: bare flag
            if (arg == "--fixed") { c.fixed = 1; continue; } // new: bare flag (RTGPU_FLAG_FIXED)
            std::string key, value;
            if (!split_flag(arg, key, value)) continue; // regex does not match: ignored
            long long iv; double dv, vec[3];
            if (key == "w" || key == "-width") {
                if (!parse_usize(value, iv)) return fail("Output width must be a positive integer");
                c.scene.has_output_width = 1; c.scene.output_width = (int)iv;
            } else if (key == "r" || key == "-aspect-ratio") {
                if (!parse_f64(value, dv)) return fail("Aspect ratio must be a number");
                c.scene.has_aspect_ratio = 1; c.scene.aspect_ratio = dv;
            } else if (key == "f" || key == "-focal-length") {
                if (!parse_f64(value, dv)) return fail("Focal length must be a number");
                c.scene.has_focal_length = 1; c.scene.focal_length = dv;
            } else if (key == "a" || key == "-aperture") {
                if (!parse_f64(value, dv)) return fail("Aperture must be a number");
                c.scene.has_f_number = 1; c.scene.f_number = dv;
            } else if (key == "d" || key == "-focus-dist") {
                if (!parse_f64(value, dv)) return fail("Focus distance must be a number");
                c.scene.has_focus_distance = 1; c.scene.focus_distance = dv;
            } else if (key == "c" || key == "-camera-position") {
                if (parse_vec3(value, vec) != 1) return fail("Vector must have three components");
                c.scene.has_camera_pos = 1; for (int k = 0; k < 3; k++) c.scene.camera_pos[k] = vec[k];
            } else if (key == "l" || key == "-look-at") {
                if (parse_vec3(value, vec) != 1) return fail("Vector must have three components");
                c.scene.has_camera_target = 1; for (int k = 0; k < 3; k++) c.scene.camera_target[k] = vec[k];
            } else if (key == "t" || key == "-threads") {
                if (!parse_usize(value, iv)) return fail("Thread count must be a positive integer");
                c.thread_count = (int)iv;
            } else if (key == "s" || key == "-samples") {
                if (!parse_usize(value, iv)) return fail("Sample count must be a positive integer");
                c.samples_per_pixel = (int)iv;
            } else if (key == "b" || key == "-background-color") {
                if (parse_vec3(value, vec) != 1) return fail("Vector must have three components");
                c.scene.has_background = 1; for (int k = 0; k < 3; k++) c.scene.background[k] = vec[k];
            } else if (key == "-max-depth") {
                if (!parse_usize(value, iv)) return fail("Max ray depth must be a positive integer");
                c.max_depth = (int)iv;
            } else if (key == "-light-bias") {
                if (!parse_f64(value, dv)) return fail("Light bias must be a number");
                if (!(dv >= 0.0 && dv <= 1.0)) return fail("Light bias must be in range [0; 1]");
                c.light_bias = dv;
            } else if (key == "-gpus") { // new
                if (!parse_usize(value, iv) || iv < 1) return fail("GPU count must be a positive integer");
                c.n_gpus = (int)iv; c.use_gpu = 1;
            } else if (key == "-seed") { // new
                if (!parse_usize(value, iv)) return fail("Seed must be a non-negative integer");
                c.seed = (uint64_t)iv;
            } else if (key == "-scene-seed") { // new
                if (!parse_usize(value, iv)) return fail("Scene seed must be a non-negative integer");
                c.scene_seed = (uint64_t)iv;
            } // `_ => ()` : unknown keys are ignored (config.rs:146)
        } else {
            std::strncpy(c.scene_name, arg.c_str(), sizeof c.scene_name - 1);
        }
    }
    if (c.thread_count < 1) return fail("Thread count must be a positive integer"); // division by zero panic in the reference
    int spt = c.samples_per_pixel / c.thread_count;                 // config.rs:154
    c.sqrt_samples_per_thread = (int)std::sqrt((double)spt);        // :155
    *out = c;
    return 0;
}
