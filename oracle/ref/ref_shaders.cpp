// ref_shaders.cpp -- runs the reference's own compute-shader source on the CPU (part of oracle/_ref).
//
// TEST INFRASTRUCTURE. Each shader of <reference>/shaders is compiled into a C++ class (the .comp text is #included
// inside the class body after the lexical clean-up of gen_ref_sources.py), so `main()` below IS the reference's main().
// This file only plays the role of the Vulkan dispatch: it binds the UBO / buffers, sets gl_GlobalInvocationID and
// calls main() once per invocation. Light generation runs its invocations one after another in the canonical
// photon-major order (a legal serialisation of the shader's atomicAdd race).
#include <atomic>
#include <thread>
#include <vector>

#include "../vr_oracle.h"
#include "glsl_shim.h"

namespace glsl {
pnv::pnanovdb_buf_t g_buf = {nullptr};

#define VR_SHADER_CLASS(NAME, FILE) /* documentation only: see the expanded blocks below */

struct S_light_gen : Builtins {
#include "light_gen.comp"
};
struct S_point_light_gen : Builtins {
#include "point_light_gen.comp"
};
struct S_ray_light_gen : Builtins {
#include "ray_light_gen.comp"
};
struct S_sphere_light_gen : Builtins {
#include "sphere_light_gen.comp"
};
struct S_beam_light_gen : Builtins {
#include "beam_light_gen.comp"
};
struct S_path_light_gen : Builtins {
#include "path_light_gen.comp"
};
struct S_point_cc : Builtins {
#include "point_compute_color.comp"
};
struct S_sphere_cc : Builtins {
#include "sphere_compute_color.comp"
};
struct S_ray_cc : Builtins {
#include "ray_compute_color.comp"
};
struct S_beam_cc : Builtins {
#include "beam_compute_color.comp"
};
struct S_path_cc : Builtins {
#include "path_compute_color.comp"
};

template <class S>
void bind_ubo(S& s, const vro_ubo& u) {
    s.frameCount = u.frameCount;
    s.framebufferDim = uvec2(u.framebufferDim[0], u.framebufferDim[1]);
    s.cameraPos = vec3(u.cameraPos[0], u.cameraPos[1], u.cameraPos[2]);
    s.fov = u.fov;
    s.photonInitialIntensity = u.photonInitialIntensity;
    s.scatteringProbability = u.scatteringProbability;
    s.absorptionCoefficient = u.absorptionCoefficient;
    s.maxLights = u.maxLights;
    s.rayMaxDistance = u.rayMaxDistance;
    s.rayMarchingStepSize = u.rayMarchingStepSize;
    s.lightSourceWorldPos = vec3(u.lightSourceWorldPos[0], u.lightSourceWorldPos[1], u.lightSourceWorldPos[2]);
    s.beamRadius = u.beamRadius;
    s.lightRayStepSize = u.lightRayStepSize;
    s.radiusFalloff = u.radiusFalloff;
}

template <class S>
int run_light_gen(const vro_ubo& u, uint32_t gx, uint32_t gy, uint32_t gz, vro_light* out, uint32_t* count) {
    S s;
    bind_ubo(s, u);
    std::vector<typename S::RayLight> buf((size_t)u.maxLights + 1);
    s.lights = buf.data();
    s.lightCount = 0;
    const uint32_t Nx = gx * 4u, Ny = gy * 4u, Nz = gz;
    for (uint32_t z = 0; z < Nz; z++)
        for (uint32_t y = 0; y < Ny; y++)
            for (uint32_t x = 0; x < Nx; x++) {
                s.gl_GlobalInvocationID = uvec3(x, y, z);
                s.main();
            }
    uint32_t n = s.lightCount < u.maxLights ? s.lightCount : u.maxLights;
    for (uint32_t i = 0; i < n; i++) {
        out[i].from[0] = buf[i].positionFrom.x, out[i].from[1] = buf[i].positionFrom.y, out[i].from[2] = buf[i].positionFrom.z;
        out[i]._pad0 = 0.0f;
        out[i].to[0] = buf[i].positionTo.x, out[i].to[1] = buf[i].positionTo.y, out[i].to[2] = buf[i].positionTo.z;
        out[i].intensity = buf[i].intensity;
    }
    *count = s.lightCount;
    return 0;
}

template <class S>
int run_gather(const vro_ubo& u, const vro_light* lights, uint32_t lightCount, uint32_t x0, uint32_t y0, uint32_t w,
               uint32_t h, uint32_t stride, int nthreads, float* image_rgba, int unorm8) {
    std::vector<typename S::RayLight> buf((size_t)lightCount + 1);
    for (uint32_t i = 0; i < lightCount; i++) {
        buf[i].positionFrom = vec3(lights[i].from[0], lights[i].from[1], lights[i].from[2]);
        buf[i].positionTo = vec3(lights[i].to[0], lights[i].to[1], lights[i].to[2]);
        buf[i].intensity = lights[i].intensity;
    }
    if (stride == 0) stride = 1;
    if (nthreads < 1) nthreads = 1;
    const uint32_t nrows = (h + stride - 1) / stride;
    std::atomic<uint32_t> next{0};
    auto worker = [&]() {
        S s;
        bind_ubo(s, u);
        s.lights = buf.data();
        s.lightCount = lightCount;
        s.img.w = (int)u.framebufferDim[0], s.img.h = (int)u.framebufferDim[1];
        s.img.rgba = image_rgba;
        s.img.unorm8 = unorm8 != 0;
        for (;;) {
            uint32_t r = next.fetch_add(1);
            if (r >= nrows) break;
            uint32_t py = y0 + r * stride;
            for (uint32_t px = x0; px < x0 + w; px += stride) {
                s.gl_GlobalInvocationID = uvec3(px, py, 0);
                s.main();
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; t++) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
    return 0;
}
}  // namespace glsl

extern "C" {

// binding 2
void vrr_set_grid(const void* grid) { glsl::g_buf.data = (uint32_t*)grid; }

// shader: 0 light_gen.comp, 1 point_light_gen, 2 ray_light_gen, 3 sphere_light_gen, 4 beam_light_gen, 5 path_light_gen
int vrr_light_gen(const vro_ubo* ubo, int shader, uint32_t gx, uint32_t gy, uint32_t gz, vro_light* lights_out,
                  uint32_t* count_out) {
    using namespace glsl;
    switch (shader) {
        case 0: return run_light_gen<S_light_gen>(*ubo, gx, gy, gz, lights_out, count_out);
        case 1: return run_light_gen<S_point_light_gen>(*ubo, gx, gy, gz, lights_out, count_out);
        case 2: return run_light_gen<S_ray_light_gen>(*ubo, gx, gy, gz, lights_out, count_out);
        case 3: return run_light_gen<S_sphere_light_gen>(*ubo, gx, gy, gz, lights_out, count_out);
        case 4: return run_light_gen<S_beam_light_gen>(*ubo, gx, gy, gz, lights_out, count_out);
        case 5: return run_light_gen<S_path_light_gen>(*ubo, gx, gy, gz, lights_out, count_out);
    }
    return 1;
}

// algo: enum order of src/main.cpp:65-68 (0 beam, 1 ray, 2 point, 3 sphere, 4 path). image_rgba is the storage image
// (float RGBA, W*H*4, read-modify-written by the shader's running average); unorm8 != 0 quantises stores like rgba8.
int vrr_gather(const vro_ubo* ubo, int algo, const vro_light* lights, uint32_t light_count, uint32_t x0, uint32_t y0,
               uint32_t w, uint32_t h, uint32_t stride, int nthreads, float* image_rgba, int unorm8) {
    using namespace glsl;
    switch (algo) {
        case VRO_BEAM: return run_gather<S_beam_cc>(*ubo, lights, light_count, x0, y0, w, h, stride, nthreads, image_rgba, unorm8);
        case VRO_RAY: return run_gather<S_ray_cc>(*ubo, lights, light_count, x0, y0, w, h, stride, nthreads, image_rgba, unorm8);
        case VRO_POINT: return run_gather<S_point_cc>(*ubo, lights, light_count, x0, y0, w, h, stride, nthreads, image_rgba, unorm8);
        case VRO_SPHERE: return run_gather<S_sphere_cc>(*ubo, lights, light_count, x0, y0, w, h, stride, nthreads, image_rgba, unorm8);
        case VRO_PATH: return run_gather<S_path_cc>(*ubo, lights, light_count, x0, y0, w, h, stride, nthreads, image_rgba, unorm8);
    }
    return 1;
}

// the shaders' hash / randf / randomDir, for known-answer tests
uint32_t vrr_hash(uint32_t x, uint32_t y, uint32_t z) {
    glsl::S_light_gen s;
    return s.hash(glsl::uvec3(x, y, z));
}
float vrr_randf(uint32_t x, uint32_t y, uint32_t z) {
    glsl::S_light_gen s;
    return s.randf(glsl::uvec3(x, y, z));
}
void vrr_random_dir(float r1, float r2, float out[3]) {
    glsl::S_light_gen s;
    glsl::vec3 d = s.randomDir(r1, r2);
    out[0] = d.x, out[1] = d.y, out[2] = d.z;
}

}  // extern "C"
