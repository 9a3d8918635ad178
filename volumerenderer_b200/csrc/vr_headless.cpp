// vr_headless.cpp -- headless frame driver on top of the C ABI.
//
// Stands where VolumeApp::mainLoop / drawFrame stand in the reference (src/main.cpp:601-607, 620-837) minus everything
// that needs a window: per frame it bumps frameCount, regenerates the virtual lights and gathers, exactly the compute
// section of drawFrame; at the end it writes the image instead of presenting it. The ImGui "Settings" panel
// (src/main.cpp:649-700) becomes a flat JSON object / --set name=value pairs with the same names and slider ranges.
//
//   vr_headless --grid cloud:256 --algo point --frames 16 --size 1920x1080 --set maxLights=1000 --lightgen 32 --ppm out.ppm
//   vr_headless --config scene.json            {"grid": "bunny.nvdb", "algo": "beam", "frames": 64, "beamRadius": 0.1, ...}
//   vr_headless --grid sphere:32 --cpu-test density.ppm      (CPU_test/main.cpp: fixed camera, 256x256, 1200 steps)
//   vr_headless --gpus 8 --grid cloud:256 --algo point --size 1920x1080 --lightgen 32 --frames 64 --bench 20
//                                                             (one process, 8 GPUs: vr_group_*, image bands, csrc/vr_group.cu)
//   vr_headless --gpus 8 --progressive 1 --algo beam --frames 64 ...   (sample-parallel: frame f on GPU f mod 8, one reduce)
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <map>
#include <string>
#include <vector>

#include "../../include/vr.h"

namespace {

struct Options {
    std::string grid = "cloud:64", algo = "point", ppm, pfm, snap_in, snap_out, cpu_test, grid_name, save_nvdb;
    uint32_t width = 1024, height = 1024;  // the reference window (src/main.cpp:40-41)
    uint32_t frames = 1, gx = 1, gy = 1, gz = 1;
    int device = 0, keep_origin = 0, quiet = 0, march_cache = 0;
    // several GPUs from this one process (vr_group_*): image bands, or whole frames round-robin (--progressive)
    int gpus = 1, progressive = 0, balance_rounds = 6;
    uint32_t bench = 0;  // time this many extra frames with CUDA events after the run (maximum over the GPUs)
    std::string exchange = "none", devices;
    // opt-in physics (vr_options) and the reference's commented-out radius update (src/main.cpp:609-618, :713)
    int trilinear = 0, light_transmittance = 0, phase_hg = 0, radius_shrink = 0;
    double hg_g = 0.0, radius_alpha = 0.0;
    std::map<std::string, std::vector<double>> sets;  // UBO overrides
};

[[noreturn]] void die(const std::string& msg) {
    fprintf(stderr, "vr_headless: %s\n", msg.c_str());
    exit(2);
}
void check(int rc, vr_ctx* ctx, const char* what) {
    if (rc == VR_OK) return;
    const char* e = vr_last_error(ctx);  // ctx == NULL: the message of a failed vr_create
    if (!e || !*e) e = vr_io_last_error();
    fprintf(stderr, "vr_headless: %s failed (status %d): %s\n", what, rc, e ? e : "");
    exit(1);
}

// ---- the settings surface: UBO member -> (offset in floats/uints, count, is_uint, slider min, slider max) src/main.cpp:662-691
struct Field {
    const char* name;
    int count;
    bool is_uint;
    double lo, hi;
};
const Field kFields[] = {
    {"cameraPos", 3, false, -200.0, 200.0},          {"fov", 1, false, 1.0, 179.0},
    {"photonInitialIntensity", 1, false, 0.0, 500.0}, {"scatteringProbability", 1, false, 0.0, 1.0},
    {"absorptionCoefficient", 1, false, 0.0, 1.0},    {"maxLights", 1, true, 0.0, 1000000.0},
    {"rayMaxDistance", 1, false, 0.0, 20000.0},       {"rayMarchingStepSize", 1, false, 0.0, 10.0},
    {"lightSourceWorldPos", 3, false, -100.0, 100.0}, {"beamRadius", 1, false, 0.0, 10.0},
    {"lightRayStepSize", 1, false, 0.0, 10.0},        {"radiusFalloff", 1, false, 0.0, 1.0},
};
void apply_field(vr_ubo& u, const std::string& name, const std::vector<double>& v, bool quiet) {
    const Field* f = nullptr;
    for (const Field& k : kFields)
        if (name == k.name) f = &k;
    if (!f) die("unknown setting '" + name + "'");
    if ((int)v.size() != f->count) die("setting '" + name + "' takes " + std::to_string(f->count) + " value(s)");
    for (double x : v)
        if (!quiet && (x < f->lo || x > f->hi))
            fprintf(stderr, "vr_headless: note: %s = %g is outside the reference UI range [%g, %g]\n", f->name, x, f->lo, f->hi);
    if (name == "cameraPos") for (int i = 0; i < 3; i++) u.cameraPos[i] = (float)v[i];
    else if (name == "fov") u.fov = (float)v[0];
    else if (name == "photonInitialIntensity") u.photonInitialIntensity = (float)v[0];
    else if (name == "scatteringProbability") u.scatteringProbability = (float)v[0];
    else if (name == "absorptionCoefficient") u.absorptionCoefficient = (float)v[0];
    else if (name == "maxLights") u.maxLights = (uint32_t)v[0];
    else if (name == "rayMaxDistance") u.rayMaxDistance = (float)v[0];
    else if (name == "rayMarchingStepSize") u.rayMarchingStepSize = (float)v[0];
    else if (name == "lightSourceWorldPos") for (int i = 0; i < 3; i++) u.lightSourceWorldPos[i] = (float)v[i];
    else if (name == "beamRadius") u.beamRadius = (float)v[0];
    else if (name == "lightRayStepSize") u.lightRayStepSize = (float)v[0];
    else if (name == "radiusFalloff") u.radiusFalloff = (float)v[0];
}

std::vector<double> parse_numbers(const std::string& s) {
    std::vector<double> out;
    const char* p = s.c_str();
    while (*p) {
        while (*p == ',' || *p == ' ' || *p == '[' || *p == ']' || *p == 'x') p++;
        if (!*p) break;
        char* end = nullptr;
        const double v = strtod(p, &end);
        if (end == p) die("cannot parse number in '" + s + "'");
        out.push_back(v);
        p = end;
    }
    return out;
}

// ---- flat JSON object: {"key": number | "string" | [numbers], ...}
void parse_json_config(const std::string& path, Options& o);

void set_option(Options& o, const std::string& key, const std::string& val) {
    if (key == "grid") o.grid = val;
    else if (key == "grid-name" || key == "gridName") o.grid_name = val;
    else if (key == "algo" || key == "algorithm") o.algo = val;
    else if (key == "ppm") o.ppm = val;
    else if (key == "pfm") o.pfm = val;
    else if (key == "snapshot-in" || key == "snapshotIn") o.snap_in = val;
    else if (key == "snapshot-out" || key == "snapshotOut") o.snap_out = val;
    else if (key == "cpu-test" || key == "cpuTest") o.cpu_test = val;
    else if (key == "save-nvdb" || key == "saveNvdb") o.save_nvdb = val;
    else if (key == "frames") o.frames = (uint32_t)parse_numbers(val).at(0);
    else if (key == "device") o.device = (int)parse_numbers(val).at(0);
    else if (key == "gpus") o.gpus = (int)parse_numbers(val).at(0);
    else if (key == "devices") o.devices = val;
    else if (key == "progressive") o.progressive = (int)parse_numbers(val).at(0);
    else if (key == "balance-rounds" || key == "balanceRounds") o.balance_rounds = (int)parse_numbers(val).at(0);
    else if (key == "bench") o.bench = (uint32_t)parse_numbers(val).at(0);
    else if (key == "exchange") o.exchange = val;
    else if (key == "width") o.width = (uint32_t)parse_numbers(val).at(0);
    else if (key == "height") o.height = (uint32_t)parse_numbers(val).at(0);
    else if (key == "march-cache" || key == "marchCache") o.march_cache = (int)parse_numbers(val).at(0);
    else if (key == "trilinear") o.trilinear = (int)parse_numbers(val).at(0);
    else if (key == "light-transmittance" || key == "lightTransmittance") o.light_transmittance = (int)parse_numbers(val).at(0);
    else if (key == "phase-hg" || key == "phaseHg") o.phase_hg = 1, o.hg_g = parse_numbers(val).at(0);
    else if (key == "radius-update" || key == "radiusUpdate") o.radius_shrink = 1, o.radius_alpha = parse_numbers(val).at(0);
    else if (key == "size") {
        auto v = parse_numbers(val);
        if (v.size() != 2) die("--size takes WxH");
        o.width = (uint32_t)v[0], o.height = (uint32_t)v[1];
    } else if (key == "lightgen") {
        auto v = parse_numbers(val);
        if (v.empty() || v.size() > 3) die("--lightgen takes gx[,gy[,gz]]");
        o.gx = (uint32_t)v[0], o.gy = v.size() > 1 ? (uint32_t)v[1] : 1, o.gz = v.size() > 2 ? (uint32_t)v[2] : 1;
    } else if (key == "lightgen-variant" || key == "lightgenVariant") {
        if (val == "reset" || val == "light_gen") o.keep_origin = 0;
        else if (val == "keep" || val == "point_light_gen") o.keep_origin = 1;
        else die("--lightgen-variant takes reset|keep");
    } else if (key == "config") parse_json_config(val, o);
    else o.sets[key] = parse_numbers(val);  // a UBO member; validated when applied
}

void parse_json_config(const std::string& path, Options& o) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) die("cannot read config " + path);
    std::string s;
    char buf[4096];
    size_t n;
    while ((n = fread(buf, 1, sizeof(buf), f)) > 0) s.append(buf, n);
    fclose(f);
    size_t i = s.find('{');
    if (i == std::string::npos) die(path + ": expected a JSON object");
    i++;
    for (;;) {
        while (i < s.size() && (isspace((unsigned char)s[i]) || s[i] == ',')) i++;
        if (i >= s.size()) die(path + ": unterminated object");
        if (s[i] == '}') break;
        if (s[i] != '"') die(path + ": expected a quoted key");
        const size_t k1 = s.find('"', i + 1);
        if (k1 == std::string::npos) die(path + ": unterminated key");
        const std::string key = s.substr(i + 1, k1 - i - 1);
        i = s.find(':', k1);
        if (i == std::string::npos) die(path + ": expected ':' after key " + key);
        i++;
        while (i < s.size() && isspace((unsigned char)s[i])) i++;
        std::string val;
        if (s[i] == '"') {
            const size_t e = s.find('"', i + 1);
            if (e == std::string::npos) die(path + ": unterminated string");
            val = s.substr(i + 1, e - i - 1);
            i = e + 1;
        } else if (s[i] == '[') {
            const size_t e = s.find(']', i);
            if (e == std::string::npos) die(path + ": unterminated array");
            val = s.substr(i, e - i + 1);
            i = e + 1;
        } else {
            size_t e = i;
            while (e < s.size() && s[e] != ',' && s[e] != '}' && !isspace((unsigned char)s[e])) e++;
            val = s.substr(i, e - i);
            if (val == "true") val = "1";
            if (val == "false") val = "0";
            i = e;
        }
        if (key == "config") die(path + ": nested config files are not supported");
        set_option(o, key, val);
    }
}

vr_algo parse_algo(const std::string& a) {
    // the combo box order of src/main.cpp:65-68, :652-658
    if (a == "beam" || a == "Beam") return VR_ALGO_BEAM;
    if (a == "ray" || a == "Ray") return VR_ALGO_RAY;
    if (a == "point" || a == "Point") return VR_ALGO_POINT;
    if (a == "sphere" || a == "Sphere") return VR_ALGO_SPHERE;
    if (a == "path" || a == "Path") return VR_ALGO_PATH;
    die("unknown algorithm '" + a + "' (beam|ray|point|sphere|path)");
}

void load_grid(const Options& o, void** bytes, uint64_t* size) {
    const double tr[3] = {-32.0, -12.0, -32.0};  // SURVEY 8(d) scene
    if (o.grid.rfind("cloud:", 0) == 0) {
        auto v = parse_numbers(o.grid.substr(6));
        if (v.empty()) die("--grid cloud:N[,seed]");
        const int n = (int)v[0];
        check(vr_grid_build_cloud(n, v.size() > 1 ? (uint32_t)v[1] : 1234u, 64.0 / n, tr, 0, bytes, size), nullptr, "vr_grid_build_cloud");
    } else if (o.grid.rfind("sphere:", 0) == 0) {
        auto v = parse_numbers(o.grid.substr(7));
        if (v.empty()) die("--grid sphere:R[,cx,cy,cz]");
        // default centre: in front of CPU_test's fixed camera (CPU_test/main.cpp:92)
        const int32_t c[3] = {v.size() > 3 ? (int32_t)v[1] : 0, v.size() > 3 ? (int32_t)v[2] : 250, v.size() > 3 ? (int32_t)v[3] : -200};
        const double zero[3] = {0, 0, 0};
        check(vr_grid_build_fog_sphere(v[0], c, 1.0, zero, 0, bytes, size), nullptr, "vr_grid_build_fog_sphere");
    } else {
        check(vr_grid_load_nvdb(o.grid.c_str(), o.grid_name.empty() ? nullptr : o.grid_name.c_str(), bytes, size), nullptr, "vr_grid_load_nvdb");
    }
}

void usage() {
    puts("vr_headless: headless driver of libvr (VolumeRenderer's many-light passes on B200)\n"
         "  --grid cloud:N[,seed] | sphere:R[,cx,cy,cz] | file.nvdb     volume (default cloud:64)\n"
         "  --grid-name NAME         float grid to pick from a multi-grid .nvdb (default: the first float grid)\n"
         "  --algo beam|ray|point|sphere|path                            (default point)\n"
         "  --size WxH | --width W --height H                            (default 1024x1024, the reference window)\n"
         "  --frames N               progressive frames (new lights every frame)\n"
         "  --lightgen gx[,gy,gz]    light-generation work-groups of 4x4 photons (reference: 1)\n"
         "  --lightgen-variant reset|keep   light_gen.comp | {point,ray,sphere}_light_gen.comp\n"
         "  --set NAME=V[,V,V]       UBO member (cameraPos, fov, photonInitialIntensity, scatteringProbability,\n"
         "                           absorptionCoefficient, maxLights, rayMaxDistance, rayMarchingStepSize,\n"
         "                           lightSourceWorldPos, beamRadius, lightRayStepSize); also --NAME V\n"
         "  --config FILE.json       flat JSON object with any of the options above\n"
         "  --ppm FILE  --pfm FILE   rgba8 image as binary PPM / float accumulator as PFM\n"
         "  --snapshot-in FILE --snapshot-out FILE   resume from / save the progressive state\n"
         "  --cpu-test FILE.ppm      CPU_test's density ray-march (256x256) instead of the many-light passes\n"
         "  --save-nvdb FILE         write the volume as an uncompressed .nvdb\n"
         "  --trilinear 0|1  --phase-hg G  --light-transmittance 0|1   opt-in physics the reference lacks (vr_options)\n"
         "  --radius-update ALPHA    beamRadius = calculateRadius(beamRadius, ALPHA) before every frame (src/main.cpp:609-618)\n"
         "  --gpus N | --devices a,b,..   several GPUs from this process (image bands balanced by measured feedback;\n"
         "                           a device may be listed twice: several bands on one GPU)\n"
         "  --exchange none|allgather     none: every GPU keeps the average of its band (default); allgather: the whole image\n"
         "                           on every GPU after every frame (peer stores from the resolve kernel + device barrier)\n"
         "  --progressive 1          sample-parallel progressive run: frame f on GPU f mod N, float sums, one reduce\n"
         "  --balance-rounds R       feedback rounds of the band balancer (default 6)\n"
         "  --bench K                afterwards time K more frames with CUDA events (prints ms_per_frame_device)\n"
         "  --march-cache 0|1  --device D  --quiet");
}

void gcheck(int rc, vr_group* g, const char* what) {
    if (rc == VR_OK) return;
    fprintf(stderr, "vr_headless: %s failed (status %d): %s\n", what, rc, vr_group_last_error(g));
    exit(1);
}

// the frame loop on several GPUs of this node (vr_group_*): what `--gpus N` / `--devices` select
int run_group(const Options& o, void* grid, uint64_t grid_size) {
    std::vector<int> devs;
    if (!o.devices.empty()) for (double d : parse_numbers(o.devices)) devs.push_back((int)d);
    else for (int i = 0; i < o.gpus; i++) devs.push_back(i);
    if (!o.cpu_test.empty() || !o.snap_in.empty() || !o.snap_out.empty() || o.radius_shrink)
        die("--cpu-test, snapshots and --radius-update are one-GPU options");
    const vr_algo algo = parse_algo(o.algo);
    vr_group* g = nullptr;
    gcheck(vr_group_create(devs.data(), (uint32_t)devs.size(), &g), nullptr, "vr_group_create");
    const uint32_t n = vr_group_size(g);
    gcheck(vr_group_set_grid(g, grid, grid_size), g, "vr_group_set_grid");
    vr_grid_free(grid);
    if (!o.quiet) fprintf(stderr, "%s, %u GPUs, grid %s: %.2f MB\n", vr_version(), n, o.grid.c_str(), grid_size / 1e6);
    gcheck(vr_group_resize(g, o.width, o.height), g, "vr_group_resize");
    vr_ubo u;
    vr_default_params(&u, o.width, o.height);
    for (auto& kv : o.sets) apply_field(u, kv.first, kv.second, o.quiet != 0);
    if (u.maxLights > 87500) gcheck(vr_group_set_max_lights_capacity(g, u.maxLights), g, "vr_group_set_max_lights_capacity");
    gcheck(vr_group_set_params(g, &u), g, "vr_group_set_params");
    gcheck(vr_group_set_light_gen_groups(g, o.gx, o.gy, o.gz, o.keep_origin ? VR_LG_KEEP_ORIGIN : VR_LG_RESET_ORIGIN), g, "vr_group_set_light_gen_groups");
    vr_options phys = {(uint32_t)(o.trilinear != 0), (uint32_t)o.phase_hg, (float)o.hg_g, (uint32_t)(o.light_transmittance != 0)};
    for (uint32_t i = 0; i < n; i++) check(vr_set_options(vr_group_ctx(g, i), &phys), vr_group_ctx(g, i), "vr_set_options");
    if (o.exchange == "allgather") gcheck(vr_group_set_exchange(g, VR_EXCHANGE_ALLGATHER), g, "vr_group_set_exchange");
    else if (o.exchange != "none") die("--exchange takes none|allgather");

    std::vector<uint32_t> y0(n), bh(n);
    const auto t0 = std::chrono::steady_clock::now();
    if (o.progressive) {
        gcheck(vr_group_render_progressive(g, algo, o.frames), g, "vr_group_render_progressive");
    } else {
        gcheck(vr_group_balance(g, algo, (uint32_t)o.balance_rounds), g, "vr_group_balance");
        for (uint32_t f = 0; f < o.frames; f++) gcheck(vr_group_render_frame(g, algo), g, "vr_group_render_frame");
        gcheck(vr_group_sync(g), g, "vr_group_sync");
    }
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    vr_group_get_bands(g, y0.data(), bh.data());
    vr_group_get_params(g, &u);

    const size_t npix = (size_t)o.width * o.height;
    if (!o.ppm.empty()) {
        std::vector<uint8_t> img(npix * 4);
        if (o.progressive) check(vr_read_image_rgba8(vr_group_ctx(g, 0), img.data()), vr_group_ctx(g, 0), "vr_read_image_rgba8");
        else gcheck(vr_group_read_image_rgba8(g, img.data()), g, "vr_group_read_image_rgba8");
        check(vr_image_write_ppm(o.ppm.c_str(), img.data(), o.width, o.height), nullptr, "vr_image_write_ppm");
    }
    if (!o.pfm.empty()) {
        std::vector<float> acc(npix * 4);
        if (o.progressive) check(vr_read_accum_f32(vr_group_ctx(g, 0), acc.data()), vr_group_ctx(g, 0), "vr_read_accum_f32");
        else gcheck(vr_group_read_accum_f32(g, acc.data()), g, "vr_group_read_accum_f32");
        check(vr_image_write_pfm(o.pfm.c_str(), acc.data(), o.width, o.height, 4), nullptr, "vr_image_write_pfm");
    }
    float dev_ms = 0.0f;
    std::vector<float> per(n, 0.0f);
    if (o.bench && !o.progressive) {
        for (int w = 0; w < 3; w++) gcheck(vr_group_render_frame(g, algo), g, "vr_group_render_frame");
        gcheck(vr_group_time_frames(g, algo, o.bench, &dev_ms, per.data()), g, "vr_group_time_frames");
    }
    if (!o.quiet) {
        printf("{\"gpus\": %u, \"mode\": \"%s\", \"exchange\": \"%s\", \"frames\": %u, \"frameCount\": %u, \"wall_ms_total\": %.3f, \"algo\": \"%s\", "
               "\"width\": %u, \"height\": %u, \"bands\": [", n, o.progressive ? "sample-parallel" : "image bands", o.exchange.c_str(), o.frames,
               u.frameCount, ms, o.algo.c_str(), o.width, o.height);
        for (uint32_t i = 0; i < n; i++) printf("%s[%u, %u]", i ? ", " : "", y0[i], bh[i]);
        printf("], \"ms_per_frame_device\": %.4f, \"per_gpu_ms\": [", dev_ms);
        for (uint32_t i = 0; i < n; i++) printf("%s%.4f", i ? ", " : "", per[i]);
        printf("]}\n");
    }
    vr_group_destroy(g);
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    Options o;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        if (a == "--help" || a == "-h") {
            usage();
            return 0;
        }
        if (a == "--quiet") {
            o.quiet = 1;
            continue;
        }
        if (a.rfind("--", 0) != 0) die("unexpected argument '" + a + "' (see --help)");
        a = a.substr(2);
        if (i + 1 >= argc) die("--" + a + " needs a value");
        std::string v = argv[++i];
        if (a == "set") {
            const size_t eq = v.find('=');
            if (eq == std::string::npos) die("--set takes NAME=VALUE");
            o.sets[v.substr(0, eq)] = parse_numbers(v.substr(eq + 1));
        } else {
            set_option(o, a, v);
        }
    }

    void* grid = nullptr;
    uint64_t grid_size = 0;
    load_grid(o, &grid, &grid_size);
    if (!o.save_nvdb.empty()) check(vr_grid_save_nvdb(o.save_nvdb.c_str(), grid, grid_size, 0), nullptr, "vr_grid_save_nvdb");

    if (o.gpus > 1 || !o.devices.empty()) return run_group(o, grid, grid_size);

    vr_ctx* ctx = nullptr;
    check(vr_create(o.device, &ctx), nullptr, "vr_create");
    check(vr_set_grid(ctx, grid, grid_size), ctx, "vr_set_grid");
    vr_grid_free(grid);
    if (!o.quiet) fprintf(stderr, "%s, grid %s: %.2f MB\n", vr_version(), o.grid.c_str(), grid_size / 1e6);

    if (!o.cpu_test.empty()) {
        const uint32_t w = 256, h = 256;  // CPU_test/main.cpp:90-91
        std::vector<float> dens((size_t)w * h);
        check(vr_cpu_test_render(ctx, w, h, dens.data()), ctx, "vr_cpu_test_render");
        std::vector<uint8_t> img((size_t)w * h * 4, 255);
        for (size_t p = 0; p < dens.size(); p++) {
            // unsigned char color = static_cast<unsigned char>(std::min(density * 5.0f, 255.0f));  CPU_test/main.cpp:118
            const uint8_t c = (uint8_t)fminf(dens[p] * 5.0f, 255.0f);
            img[4 * p] = img[4 * p + 1] = img[4 * p + 2] = c;
        }
        check(vr_image_write_ppm(o.cpu_test.c_str(), img.data(), w, h), nullptr, "vr_image_write_ppm");
        if (!o.quiet) fprintf(stderr, "CPU_test density image written to %s\n", o.cpu_test.c_str());
        vr_destroy(ctx);
        return 0;
    }

    const vr_algo algo = parse_algo(o.algo);
    if (!o.snap_in.empty()) {
        check(vr_snapshot_load(ctx, o.snap_in.c_str()), ctx, "vr_snapshot_load");
        vr_ubo u;
        vr_get_params(ctx, &u);
        o.width = u.framebufferDim[0], o.height = u.framebufferDim[1];
        if (!o.quiet) fprintf(stderr, "resumed %ux%u at frame %u from %s\n", o.width, o.height, u.frameCount, o.snap_in.c_str());
    } else {
        check(vr_resize(ctx, o.width, o.height), ctx, "vr_resize");
    }
    vr_ubo u;
    check(vr_get_params(ctx, &u), ctx, "vr_get_params");
    if (o.snap_in.empty()) {
        vr_default_params(&u, o.width, o.height);  // src/main.cpp:546-559
    }
    for (auto& kv : o.sets) apply_field(u, kv.first, kv.second, o.quiet != 0);
    if (u.maxLights > 87500) check(vr_set_max_lights_capacity(ctx, u.maxLights), ctx, "vr_set_max_lights_capacity");
    check(vr_set_params(ctx, &u), ctx, "vr_set_params");
    check(vr_set_light_gen_groups(ctx, o.gx, o.gy, o.gz, o.keep_origin ? VR_LG_KEEP_ORIGIN : VR_LG_RESET_ORIGIN), ctx, "vr_set_light_gen_groups");
    check(vr_set_march_cache(ctx, o.march_cache), ctx, "vr_set_march_cache");
    vr_options phys = {(uint32_t)(o.trilinear != 0), (uint32_t)o.phase_hg, (float)o.hg_g, (uint32_t)(o.light_transmittance != 0)};
    check(vr_set_options(ctx, &phys), ctx, "vr_set_options");

    const auto t0 = std::chrono::steady_clock::now();
    vr_counters cnt{};
    for (uint32_t f = 0; f < o.frames; f++) {
        if (o.radius_shrink) {  // uboData.beamRadius = calculateRadius(uboData.beamRadius, uboData.radiusFalloff); (:713, commented out there)
            check(vr_get_params(ctx, &u), ctx, "vr_get_params");
            u.beamRadius = (float)vr_progressive_radius(u.beamRadius, o.radius_alpha);
            check(vr_set_params(ctx, &u), ctx, "vr_set_params");
        }
        check(vr_render_frame(ctx, algo, nullptr), ctx, "vr_render_frame");  // the compute section of drawFrame
        check(vr_get_counters(ctx, &cnt), ctx, "vr_get_counters");
    }
    check(vr_sync(ctx), ctx, "vr_sync");
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    vr_timings tsum{};
    uint64_t nfr = 0;
    vr_get_timings_sum(ctx, &tsum, &nfr, 0);
    vr_get_params(ctx, &u);
    if (!o.quiet)
        printf("{\"frames\": %u, \"frameCount\": %u, \"ms_per_frame\": %.4f, \"algo\": \"%s\", \"width\": %u, \"height\": %u, "
               "\"light_count\": %llu, \"eval_lights\": %llu, \"S_ref\": %llu, \"S_eff\": %llu, \"E\": %llu, "
               "\"eval_ms_per_frame\": %.4f, \"march_ms_per_frame\": %.4f, \"light_gen_ms_per_frame\": %.4f}\n",
               o.frames, u.frameCount, o.frames ? ms / o.frames : 0.0, o.algo.c_str(), o.width, o.height, (unsigned long long)cnt.light_count,
               (unsigned long long)cnt.eval_lights, (unsigned long long)cnt.S_ref, (unsigned long long)cnt.S_eff, (unsigned long long)cnt.E,
               nfr ? tsum.eval_ms / nfr : 0.0, nfr ? tsum.march_ms / nfr : 0.0, nfr ? tsum.light_gen_ms / nfr : 0.0);

    if (!o.ppm.empty()) {
        std::vector<uint8_t> img((size_t)o.width * o.height * 4);
        check(vr_read_image_rgba8(ctx, img.data()), ctx, "vr_read_image_rgba8");
        check(vr_image_write_ppm(o.ppm.c_str(), img.data(), o.width, o.height), nullptr, "vr_image_write_ppm");
    }
    if (!o.pfm.empty()) {
        std::vector<float> acc((size_t)o.width * o.height * 4);
        check(vr_read_accum_f32(ctx, acc.data()), ctx, "vr_read_accum_f32");
        check(vr_image_write_pfm(o.pfm.c_str(), acc.data(), o.width, o.height, 4), nullptr, "vr_image_write_pfm");
    }
    if (!o.snap_out.empty()) check(vr_snapshot_save(ctx, o.snap_out.c_str()), ctx, "vr_snapshot_save");
    vr_destroy(ctx);
    return 0;
}
