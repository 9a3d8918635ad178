/* vr_oracle.h -- C interface of the CPU oracle (TEST INFRASTRUCTURE, not product code).
 *
 * The oracle is a CPU restatement of the reference's shader arithmetic for the many-light hot path and of
 * the CPU_test ray-march loop. Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl
 * reference legs) may load it. The product library (libvr) never links or calls it.
 *
 * Every function cites the reference lines it follows (paths relative to the reference checkout).
 */
#ifndef VR_ORACLE_H
#define VR_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* shaders/common_bindings.h:19-34 (std140), host mirror src/main.cpp:83-98 -- 96 bytes */
typedef struct vro_ubo {
    uint32_t frameCount;      /* @0  */
    uint32_t _pad0;           /* @4  */
    uint32_t framebufferDim[2]; /* @8  */
    float cameraPos[3];       /* @16 */
    float fov;                /* @28 */
    float photonInitialIntensity; /* @32 */
    float scatteringProbability;  /* @36 */
    float absorptionCoefficient;  /* @40 */
    uint32_t maxLights;       /* @44 */
    float rayMaxDistance;     /* @48 */
    float rayMarchingStepSize; /* @52 */
    uint32_t _pad1[2];        /* @56 */
    float lightSourceWorldPos[3]; /* @64 */
    float beamRadius;         /* @76 */
    float lightRayStepSize;   /* @80 */
    float radiusFalloff;      /* @84 */
    uint32_t _pad2[2];        /* @88 */
} vro_ubo;

/* shaders/common_structures.h:4-8 in std430 (stride 32) */
typedef struct vro_light {
    float from[3];
    float _pad0;
    float to[3];
    float intensity;
} vro_light;

/* enum order of src/main.cpp:65-68 */
enum { VRO_BEAM = 0, VRO_RAY = 1, VRO_POINT = 2, VRO_SPHERE = 3, VRO_PATH = 4 };
/* 0: light_gen.comp (origin reset on scatter), 1: {point,ray,sphere}_light_gen.comp (no reset) */
enum { VRO_LG_RESET = 0, VRO_LG_KEEP = 1 };

typedef struct vro_counters {
    uint64_t S_ref;   /* camera samples inside the root bbox (before the T cut ends the ray) */
    uint64_t S_eff;   /* those with val != 0 */
    uint64_t E;       /* point/sphere evaluations actually performed */
    uint64_t pixels;  /* pixels evaluated */
    uint64_t hit_pixels; /* pixels whose ray hit the bbox */
} vro_counters;

uint32_t vro_hash(uint32_t x, uint32_t y, uint32_t z);
float vro_randf(uint32_t x, uint32_t y, uint32_t z);
void vro_random_dir(float r1, float r2, float out[3]);
/* built-in layer probes (for the accuracy tests against libm) */
float vro_exp(float x);
float vro_sin(float x);
float vro_cos(float x);
float vro_acos(float x);
float vro_tan(float x);
int vro_builtin_array(int which, const float* in, uint64_t n, float* out);

/* grid header facts: returns 0 if the buffer looks like a float NanoVDB grid */
int vro_grid_info(const void* grid, uint64_t size, int32_t bbox_min[3], int32_t bbox_max[3], float* background,
                  uint32_t node_counts[3], uint32_t* root_tiles);
void vro_lookup(const void* grid, const int32_t* ijk, uint64_t n, float* out);
void vro_world_to_index(const void* grid, const float* w, float* idx);
void vro_index_to_world(const void* grid, const float* idx, float* w);

/* light generation in canonical (photon-major) order; lights_out has capacity ubo->maxLights.
 * total_emitted (optional) receives the uncapped number of scatter events, walk_steps the loop iterations. */
int vro_light_gen(const void* grid, const vro_ubo* ubo, int variant, uint32_t gx, uint32_t gy, uint32_t gz,
                  vro_light* lights_out, uint32_t* count_out, uint64_t* total_emitted, uint64_t* walk_steps);

/* gather for one frame: writes frameColor (one float per pixel, R=G=B) for pixels
 * x in [x0,x0+w) step stride, y in [y0,y0+h) step stride into frame_color[py*W+px] (W from the UBO).
 * skip_zero: 1 = skip light evaluation when val == 0 (exact), 0 = evaluate like the shader.
 * nthreads: worker threads over rows. */
int vro_gather(const void* grid, const vro_ubo* ubo, int algo, const vro_light* lights, uint32_t light_count,
               uint32_t x0, uint32_t y0, uint32_t w, uint32_t h, uint32_t stride, int skip_zero, int nthreads,
               float* frame_color, vro_counters* counters);

/* opt-in physics the reference's shaders lack; all zero = the reference's arithmetic.
 * Definitions: oracle/vr_oracle.cpp (trilinear, physFactor). Mirrors vr_options of include/vr.h. */
typedef struct vro_options {
    uint32_t trilinear;           /* density at a camera sample: nanovdb TrilinearSampler instead of the nearest voxel */
    uint32_t phase_hg;            /* Henyey-Greenstein phase factor (relative to isotropic) per light evaluation */
    float hg_g;                   /* its asymmetry parameter, |g| < 1 */
    uint32_t light_transmittance; /* exp(-sigma_a * optical depth) between the light and the sample */
} vro_options;
int vro_gather_ex(const void* grid, const vro_ubo* ubo, int algo, const vro_light* lights, uint32_t light_count,
                  uint32_t x0, uint32_t y0, uint32_t w, uint32_t h, uint32_t stride, int skip_zero, int nthreads,
                  const vro_options* options, float* frame_color, vro_counters* counters);
/* TrilinearSampler at n index-space positions xyz[3*n] */
void vro_trilinear(const void* grid, const float* xyz, uint64_t n, float* out);

/* progressive average, float and rgba8 (shaders/point_compute_color.comp:99-105); n = number of pixels.
 * accum: float RGBA; rgba8: 4 bytes per pixel. path_mode selects the integer form of path_compute_color.comp:121-125. */
void vro_accumulate(const float* frame_color, uint64_t n, uint32_t frame_count, int path_mode, float* accum_rgba,
                    uint8_t* rgba8);

/* CPU_test/main.cpp:25-43, 89-126: density[j*w+i] = sum of values along the ray */
int vro_cpu_test(const void* grid, int width, int height, int nthreads, float* density, uint8_t* grey);

#ifdef __cplusplus
}
#endif
#endif
