// glsl_shim.h -- minimal GLSL 4.5 vocabulary in C++ so the reference's compute shaders compile as C++ (oracle/_ref).
//
// TEST INFRASTRUCTURE. The shader bodies come from <reference>/shaders via gen_ref_sources.py; this file supplies only
// the language: vector types, constructors, operators, built-in functions, image/atomic stand-ins and the
// pnanovdb_* entry points the shaders call (forwarded to the reference's vendored PNanoVDB.h compiled in C mode).
// Built-ins whose GLSL precision is implementation-defined are taken from vr_glsl_math.h (see that header).
#pragma once
#include <stdint.h>
#include <string.h>

#include "../../volumerenderer_b200/csrc/vr_glsl_math.h"

// ---- the reference's PNanoVDB (C mode) in its own namespace
namespace pnv {
#define PNANOVDB_C
#include <nanovdb/PNanoVDB.h>
}  // namespace pnv

namespace glsl {

typedef uint32_t uint;

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    vec2(float a, float b) : x(a), y(b) {}
};
struct uvec2 {
    uint x, y;
    uvec2() : x(0), y(0) {}
    uvec2(uint a, uint b) : x(a), y(b) {}
};
struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(int a, int b) : x(a), y(b) {}
    explicit ivec2(uvec2 v) : x((int)v.x), y((int)v.y) {}
    // `ivec2 imageSize = imageSize(img);` -- in C++ the new local hides the built-in inside its own initialiser;
    // this call operator gives that spelling the built-in's meaning.
    template <class Image>
    ivec2 operator()(const Image& img) const { return ivec2(img.w, img.h); }
};
struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
    vec3& operator+=(vec3 b) { x = x + b.x; y = y + b.y; z = z + b.z; return *this; }
    vec3& operator*=(float s) { x = x * s; y = y * s; z = z * s; return *this; }
};
struct ivec3 {
    int x, y, z;
    ivec3() : x(0), y(0), z(0) {}
    ivec3(int a, int b, int c) : x(a), y(b), z(c) {}
    ivec3(float a, float b, float c) : x((int)a), y((int)b), z((int)c) {}
};
struct uvec3 {
    uint x, y, z;
    uvec3() : x(0), y(0), z(0) {}
    uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
    uvec2 xy() const { return uvec2(x, y); }
    uvec3& operator+=(uvec3 b) { x += b.x; y += b.y; z += b.z; return *this; }
};
struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
    vec3 rgb() const { return vec3(x, y, z); }
};

inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline uvec3 operator*(uvec3 a, uint s) { return uvec3(a.x * s, a.y * s, a.z * s); }

// image2D stand-in: float RGBA storage; `unorm8` reproduces the rgba8 format (UNORM load, round-half-even store)
struct image2D {
    int w = 0, h = 0;
    float* rgba = nullptr;
    bool unorm8 = false;
};

// GLSL built-ins. They are members of the class every shader is compiled into (class scope wins over ::floor etc.).
struct Builtins {
    static float exp(float x) { return vr_exp(x); }
    static float sin(float x) { return vr_sin(x); }
    static float cos(float x) { return vr_cos(x); }
    static float acos(float x) { return vr_acos(x); }
    static float tan(float x) { return vr_tan(x); }
    static float sqrt(float x) { return vr_sqrt(x); }
    static float floor(float x) { return vr_floor(x); }
    static float abs(float x) { return x < 0.0f ? -x : x; }
    static float min(float x, float y) { return vr_min(x, y); }
    static float max(float x, float y) { return vr_max(x, y); }
    static float clamp(float x, float lo, float hi) { return vr_clamp(x, lo, hi); }
    static vec3 clamp(vec3 v, float lo, float hi) {
        return vec3(vr_clamp(v.x, lo, hi), vr_clamp(v.y, lo, hi), vr_clamp(v.z, lo, hi));
    }
    static float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
    static float length(vec3 a) { return vr_sqrt(dot(a, a)); }
    static vec3 normalize(vec3 a) {
        float l = length(a);
        return vec3(a.x / l, a.y / l, a.z / l);
    }
    static uint atomicAdd(uint& mem, uint v) {
        uint old = mem;
        mem += v;
        return old;
    }
    static ivec2 imageSize(const image2D& img) { return ivec2(img.w, img.h); }
    static vec4 imageLoad(const image2D& img, ivec2 p) {
        const float* q = img.rgba + 4 * ((size_t)p.y * img.w + p.x);
        return vec4(q[0], q[1], q[2], q[3]);
    }
    static void imageStore(image2D& img, ivec2 p, vec4 c) {
        float* q = img.rgba + 4 * ((size_t)p.y * img.w + p.x);
        float v[4] = {c.x, c.y, c.z, c.w};
        for (int i = 0; i < 4; i++) {
            float f = v[i];
            if (img.unorm8) {
                f = vr_clamp(f, 0.0f, 1.0f);
                if (f != f) f = 0.0f;
                f = nearbyintf(f * 255.0f) / 255.0f;
            }
            q[i] = f;
        }
    }
};

// ---- pnanovdb_* as the shaders spell them (GLSL flavour: vec3/ivec3 arguments, inout by reference)
typedef vec3 pnanovdb_vec3_t;
typedef ivec3 pnanovdb_coord_t;
typedef uint pnanovdb_uint32_t;
struct pnanovdb_buf_t {
    uint unused;
};
typedef pnv::pnanovdb_address_t pnanovdb_address_t;
typedef pnv::pnanovdb_grid_handle_t pnanovdb_grid_handle_t;
typedef pnv::pnanovdb_tree_handle_t pnanovdb_tree_handle_t;
typedef pnv::pnanovdb_root_handle_t pnanovdb_root_handle_t;
typedef pnv::pnanovdb_readaccessor_t pnanovdb_readaccessor_t;

// the grid bytes bound at "binding 2"
extern pnv::pnanovdb_buf_t g_buf;

inline pnv::pnanovdb_vec3_t to_pnv(vec3 v) {
    pnv::pnanovdb_vec3_t r;
    r.x = v.x, r.y = v.y, r.z = v.z;
    return r;
}
inline vec3 from_pnv(pnv::pnanovdb_vec3_t v) { return vec3(v.x, v.y, v.z); }

inline pnanovdb_tree_handle_t pnanovdb_grid_get_tree(pnanovdb_buf_t, pnanovdb_grid_handle_t g) {
    return pnv::pnanovdb_grid_get_tree(g_buf, g);
}
inline pnanovdb_root_handle_t pnanovdb_tree_get_root(pnanovdb_buf_t, pnanovdb_tree_handle_t t) {
    return pnv::pnanovdb_tree_get_root(g_buf, t);
}
inline void pnanovdb_readaccessor_init(pnanovdb_readaccessor_t& acc, pnanovdb_root_handle_t root) {
    pnv::pnanovdb_readaccessor_init(&acc, root);
}
inline uint pnanovdb_grid_get_grid_type(pnanovdb_buf_t, pnanovdb_grid_handle_t g) {
    return pnv::pnanovdb_grid_get_grid_type(g_buf, g);
}
inline vec3 pnanovdb_grid_world_to_indexf(pnanovdb_buf_t, pnanovdb_grid_handle_t g, vec3 src) {
    pnv::pnanovdb_vec3_t s = to_pnv(src);
    return from_pnv(pnv::pnanovdb_grid_world_to_indexf(g_buf, g, &s));
}
inline vec3 pnanovdb_grid_world_to_index_dirf(pnanovdb_buf_t, pnanovdb_grid_handle_t g, vec3 src) {
    pnv::pnanovdb_vec3_t s = to_pnv(src);
    return from_pnv(pnv::pnanovdb_grid_world_to_index_dirf(g_buf, g, &s));
}
inline vec3 pnanovdb_grid_index_to_worldf(pnanovdb_buf_t, pnanovdb_grid_handle_t g, vec3 src) {
    pnv::pnanovdb_vec3_t s = to_pnv(src);
    return from_pnv(pnv::pnanovdb_grid_index_to_worldf(g_buf, g, &s));
}
inline ivec3 pnanovdb_root_get_bbox_min(pnanovdb_buf_t, pnanovdb_root_handle_t r) {
    pnv::pnanovdb_coord_t c = pnv::pnanovdb_root_get_bbox_min(g_buf, r);
    return ivec3(c.x, c.y, c.z);
}
inline ivec3 pnanovdb_root_get_bbox_max(pnanovdb_buf_t, pnanovdb_root_handle_t r) {
    pnv::pnanovdb_coord_t c = pnv::pnanovdb_root_get_bbox_max(g_buf, r);
    return ivec3(c.x, c.y, c.z);
}
inline pnanovdb_address_t pnanovdb_readaccessor_get_value_address(uint grid_type, pnanovdb_buf_t,
                                                                  pnanovdb_readaccessor_t& acc, ivec3 ijk) {
    pnv::pnanovdb_coord_t c;
    c.x = ijk.x, c.y = ijk.y, c.z = ijk.z;
    return pnv::pnanovdb_readaccessor_get_value_address(grid_type, g_buf, &acc, &c);
}
inline float pnanovdb_read_float(pnanovdb_buf_t, pnanovdb_address_t a) { return pnv::pnanovdb_read_float(g_buf, a); }

}  // namespace glsl
