/* vr.h -- C ABI of libvr: headless B200 (sm_100a) implementation of VolumeRenderer's many-light hot path.
 *
 * This library stands where basalt's Vulkan compute dispatch stands in the reference (chrylt/VolumeRenderer).
 * There is no plugin API in the reference; the boundary is the descriptor-set contract of
 * shaders/common_bindings.h:1-34 plus the commands VolumeApp::drawFrame records around the two compute dispatches
 * (src/main.cpp:712-837). Each entry point below cites the reference mechanism it replaces (paths relative to the
 * reference checkout). Plain C types only; the caller owns every host pointer, the context owns all device memory.
 *
 * Threading: one thread per context at a time (the reference is single-threaded with one queue). Calls enqueue work
 * on the context's CUDA stream and return; vr_sync and the vr_read_* calls block. No exception crosses this ABI:
 * every function returns a vr_status and vr_last_error() gives the text (the reference throws std::runtime_error
 * and exits, src/main.cpp:1217-1227).
 */
#ifndef VR_H
#define VR_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum vr_status {
    VR_OK = 0,
    VR_E_ARG = 1,      /* invalid argument (incl. values the reference UI allows but its shaders cannot survive) */
    VR_E_CUDA = 2,     /* a CUDA runtime call failed */
    VR_E_BADGRID = 3,  /* buffer is not a NanoVDB float grid */
    VR_E_CAPACITY = 4, /* maxLights exceeds the light-buffer capacity */
    VR_E_STATE = 5     /* call order: no grid / no image yet */
} vr_status;

/* Algorithm enum, order of src/main.cpp:65-68 (selects the *_compute_color / *_light_gen pipeline pair, :192-224). */
typedef enum vr_algo { VR_ALGO_BEAM = 0, VR_ALGO_RAY = 1, VR_ALGO_POINT = 2, VR_ALGO_SPHERE = 3, VR_ALGO_PATH = 4 } vr_algo;

/* Which light-generation shader body: light_gen.comp (the one the app loads for Beam/Ray/Point/Sphere,
 * src/main.cpp:47-57; resets ray origin and t on scatter, light_gen.comp:91-92) or the never-loaded
 * {point,ray,sphere}_light_gen.comp bodies (no reset). PATH always uses path_light_gen.comp (empty main). */
typedef enum vr_lightgen_variant { VR_LG_RESET_ORIGIN = 0, VR_LG_KEEP_ORIGIN = 1 } vr_lightgen_variant;

/* binding 5: the std140 UBO of shaders/common_bindings.h:19-34, host mirror src/main.cpp:83-98. 96 bytes, bit-for-bit. */
typedef struct vr_ubo {
    uint32_t frameCount;          /* @0  */
    uint32_t _pad0;               /* @4  */
    uint32_t framebufferDim[2];   /* @8  */
    float cameraPos[3];           /* @16 */
    float fov;                    /* @28 */
    float photonInitialIntensity; /* @32 */
    float scatteringProbability;  /* @36 */
    float absorptionCoefficient;  /* @40 */
    uint32_t maxLights;           /* @44 */
    float rayMaxDistance;         /* @48 */
    float rayMarchingStepSize;    /* @52 */
    uint32_t _pad1[2];            /* @56 */
    float lightSourceWorldPos[3]; /* @64 */
    float beamRadius;             /* @76 */
    float lightRayStepSize;       /* @80 */
    float radiusFalloff;          /* @84 (unused by every shader) */
    uint32_t _pad2[2];            /* @88 */
} vr_ubo;

/* binding 3: struct RayLight (shaders/common_structures.h:4-8) in std430: stride 32 B. */
typedef struct vr_ray_light {
    float positionFrom[3]; /* @0  */
    float _pad0;           /* @12 */
    float positionTo[3];   /* @16 */
    float intensity;       /* @28 */
} vr_ray_light;

/* pixel rectangle for image-tiled multi-GPU rendering (no counterpart in the reference: it dispatches the whole
 * image, src/main.cpp:835-837) */
typedef struct vr_tile {
    uint32_t x0, y0, w, h;
} vr_tile;

/* work counters of the last vr_dispatch_compute_color */
typedef struct vr_counters {
    uint64_t S_ref;        /* camera samples inside the root bbox */
    uint64_t S_eff;        /* those with density != 0 (the ones that reach the light loop) */
    uint64_t E;            /* point/sphere evaluations performed = S_eff * (expanded light list length) */
    uint64_t eval_lights;  /* length of the expanded light list (= lightCount for point/sphere, sum n_i for ray/beam) */
    uint64_t light_count;  /* lightCount used for the 1/lightCount normalisation */
} vr_counters;

/* CUDA-event timings (ms) of the passes of the last frame / dispatches */
typedef struct vr_timings {
    float light_gen_ms; /* on the auxiliary stream in vr_render_frame: overlaps march_ms */
    float march_ms;    /* camera-ray march: count + scan + fill */
    float prepare_ms;  /* light list expansion, incl. any wait for light generation to finish */
    float eval_ms;     /* the sample x light kernel (dominant) */
    float resolve_ms;  /* per-pixel reduction + progressive average */
    float total_ms;    /* gather start to resolve end (light generation runs inside this window in vr_render_frame) */
} vr_timings;

typedef struct vr_ctx vr_ctx;

/* ---- device / context: replaces basalt::Instance/Device creation (src/main.cpp:265-270) */
int vr_create(int cuda_device, vr_ctx** out);
int vr_destroy(vr_ctx* ctx);
const char* vr_last_error(const vr_ctx* ctx); /* ctx may be NULL: error of a failed vr_create on this thread */

/* ---- binding 2: the NanoVDB buffer, device-local storage buffer filled through a staging copy
 * (src/main.cpp:1191-1212). Copies `size` bytes; validates magic, grid type FLOAT and 32-byte size granularity.
 * Grids >= 4 GiB are accepted (64-bit addressing) but cannot be parity-checked against 32-bit GLSL addressing. */
int vr_set_grid(vr_ctx* ctx, const void* nanovdb_bytes, uint64_t size);
/* The same from a buffer that already is in device memory (this GPU's, or a peer's reachable over NVLink): multi-GPU
 * hosts push the volume over PCIe once and replicate it GPU to GPU (e.g. an NCCL broadcast into `device_bytes`) instead
 * of uploading it once per GPU. The copy is enqueued on the context's stream (vr_set_stream): it runs behind the work
 * the caller enqueued there to produce `device_bytes`, and `device_bytes` may be overwritten once later work on that
 * stream has started. `header` is a HOST copy of the first `header_size` bytes of the grid, at least up to the end of
 * the root node's fixed part (1024 bytes cover every grid the NanoVDB writers lay out, where the root follows the tree);
 * it supplies what the host must know: map, bounding box, root offset, tile count. No host synchronisation: the grid
 * lives in two device buffers, the bytes go into the one the installed grid does not occupy, and the copy waits ON THE
 * DEVICE for that buffer's last readers (the frames and light generations enqueued before the previous install), so a
 * host that runs ahead may hand over the volume of frame k+1 while the GPU is still on frame k-1 (only a buffer that
 * has to grow is reallocated behind a host wait). */
int vr_set_grid_device(vr_ctx* ctx, const void* device_bytes, uint64_t size, const void* header, uint64_t header_size);
/* As vr_set_grid_device, for bytes that are produced in the order of ANOTHER stream of the caller (`ready_stream`, a
 * cudaStream_t: e.g. the side stream on which a replication of the volume runs): the copy into the grid buffer that no
 * enqueued frame reads runs on the library's upload stream behind that stream's current tail, i.e. next to the frames in
 * flight instead of behind them; the render stream and the light-generation stream wait for it. With it the next
 * frame's light generation (which needs the new grid, but not the previous frame) runs underneath the previous frame's
 * evaluation. `device_bytes` may be overwritten once work enqueued on the context's stream after this call has started. */
int vr_set_grid_device_from_stream(vr_ctx* ctx, const void* device_bytes, uint64_t size, const void* header, uint64_t header_size,
                                   void* ready_stream);

/* ---- binding 0: rgba8 storage image sized to the swapchain extent, recreated on resize (src/main.cpp:338-378, :950).
 * Allocates the rgba8 image plus a float4 accumulator and resets the progressive state (frameCount = 0, :937). */
int vr_resize(vr_ctx* ctx, uint32_t width, uint32_t height);

/* ---- binding 5: uniformBuffer->updateBuffer(&uboData) every frame (src/main.cpp:715). */
int vr_set_params(vr_ctx* ctx, const vr_ubo* params);
int vr_get_params(const vr_ctx* ctx, vr_ubo* params);
/* defaults of src/main.cpp:546-559 for a width x height image */
void vr_default_params(vr_ubo* params, uint32_t width, uint32_t height);

/* ---- bindings 3/4: light buffer + counter (src/main.cpp:380-398). The reference allocates 100000*28 B, i.e. 87500
 * lights at the 32 B GPU stride; capacity is explicit here. */
int vr_set_max_lights_capacity(vr_ctx* ctx, uint32_t n);
/* vkCmdFillBuffer(counter, 0, 8, 0) (src/main.cpp:722-728) */
int vr_reset_lights(vr_ctx* ctx);
/* upload a light list (tests; broadcast in multi-GPU runs). count <= capacity. */
int vr_write_lights(vr_ctx* ctx, const vr_ray_light* lights, uint32_t count);

/* ---- light generation without any device-memory allocation. NO COUNTERPART IN THE REFERENCE (Vulkan gives the shader
 * no scratch to manage; the need comes from how this library's light_gen pass is built). vr_dispatch_light_gen lets the
 * context allocate its working buffers (candidate lights per photon, counts, scan) with cudaMalloc. Some hosts cannot
 * accept that: one that gave the whole GPU to its own pool (a plug-in's cudaMalloc then fails or bypasses the pool's
 * accounting), or one that calls mid-frame, where cudaMalloc/cudaFree may stall every stream of the device. This call
 * needs no context and allocates nothing -- no cudaMalloc, no cudaFree, no allocating library routine: the working
 * memory is one fixed array inside the library image (about 2.1 MB on every GPU that loads libvr, held until exit).
 *   limit        at most VR_STATIC_MAX_PHOTONS photons per call (gx*4 * gy*4 * gz, work-groups of 4x4x1 as in
 *                shaders/light_gen.comp:7); more is rejected with VR_E_CAPACITY and a message, nothing is enqueued
 *   inputs       the volume in DEVICE memory (grid_device_bytes, 32-byte aligned, readable on cuda_device) plus a HOST copy
 *                of its first bytes up to the end of the root node's fixed part (grid_header; 1024 bytes cover the NanoVDB
 *                writers' layouts), exactly as for vr_set_grid_device; the UBO
 *   outputs      lights_device: room for params->maxLights entries of 32 B, 16-byte aligned; counter_device: 16 bytes,
 *                8-byte aligned: u32 lightCount, u32 contributionCounter, u64 scatter events before the maxLights cap.
 *                The counter is zeroed first (the per-frame reset, src/main.cpp:722-728), so the result equals
 *                vr_reset_lights + vr_dispatch_light_gen of a context with the same grid and UBO, byte for byte
 *   ordering     everything is enqueued on cuda_stream (a cudaStream_t of cuda_device; NULL = the default stream) and the
 *                call returns without waiting; it makes cuda_device the calling thread's current device
 *   sharing      all calls on one GPU use the same working memory, which keeps the previous call's contents (the call
 *                clears what has to start at zero). Calls that can overlap on the GPU -- different streams, other
 *                threads -- race on it and produce garbage; keep them on one stream or order them with events
 *   first call   the first call on each GPU looks up the array's device address and thereby loads this code onto that
 *                GPU; issue it at start-up, before any stream capture begins
 *   errors       vr_status as everywhere; the text goes to error_text (error_text_size bytes, may be NULL/0) because
 *                there is no context to keep it. */
#define VR_STATIC_MAX_PHOTONS 4096u
int vr_light_gen_static(int cuda_device, void* cuda_stream, const void* grid_device_bytes, uint64_t grid_size,
                        const void* grid_header, uint64_t grid_header_size, const vr_ubo* params, uint32_t gx, uint32_t gy,
                        uint32_t gz, vr_lightgen_variant variant, vr_ray_light* lights_device, void* counter_device,
                        char* error_text, uint64_t error_text_size);

/* ---- image clear to (0,0,0,1) when frameCount == 1 (src/main.cpp:731-808) */
int vr_clear_image(vr_ctx* ctx);

/* ---- vkCmdBindPipeline(lightGen[algo]) + vkCmdDispatch(1,1,1) (src/main.cpp:811-814). Work-groups are 4x4x1
 * photons (shaders/light_gen.comp:7); gx,gy,gz are the dispatch dimensions (reference: 1,1,1). Lights are appended in
 * the canonical photon-major order (one legal serialisation of the shader's atomicAdd race). */
int vr_dispatch_light_gen(vr_ctx* ctx, vr_algo algo, uint32_t gx, uint32_t gy, uint32_t gz, vr_lightgen_variant variant);

/* ---- vkCmdBindPipeline(computeColor[algo]) + vkCmdDispatch(ceil(w/16), ceil(h/16), 1) (src/main.cpp:832-837).
 * tile == NULL renders the whole image; otherwise only that pixel rectangle is updated. Stream order replaces the
 * compute->compute barrier of :817-829. */
int vr_dispatch_compute_color(vr_ctx* ctx, vr_algo algo, const vr_tile* tile_or_null);

/* ---- the compute section of VolumeApp::drawFrame (src/main.cpp:712-837): frameCount++, UBO update, counter reset,
 * clear when frameCount == 1, light generation (gx,gy,gz as set by vr_set_light_gen_groups, default 1,1,1), gather. */
int vr_set_light_gen_groups(vr_ctx* ctx, uint32_t gx, uint32_t gy, uint32_t gz, vr_lightgen_variant variant);
int vr_render_frame(vr_ctx* ctx, vr_algo algo, const vr_tile* tile_or_null);
/* frameCount = 0 (algorithm change :657, "Refresh" :695-698, resize :937) */
int vr_reset_accumulation(vr_ctx* ctx);

/* ---- read-back (the reference never reads back; the image goes to the swapchain blit) */
int vr_sync(vr_ctx* ctx);
int vr_read_lights(vr_ctx* ctx, vr_ray_light* out, uint32_t capacity, uint32_t* count);
int vr_read_light_counter(vr_ctx* ctx, uint32_t counter[2], uint64_t* total_emitted);
int vr_read_image_rgba8(vr_ctx* ctx, uint8_t* out);        /* width*height*4 bytes */
int vr_read_accum_f32(vr_ctx* ctx, float* out);            /* width*height*4 floats (RGBA, linear) */
/* the same read-back without blocking: the accumulator as it is after the work enqueued so far is snapshotted on the
 * device and copied to `out` (pinned host memory for a truly asynchronous copy) on a copy stream, underneath whatever is
 * enqueued next; `out` is valid after vr_read_wait (or the next vr_read_accum_f32_async, which waits for the previous one). */
int vr_read_accum_f32_async(vr_ctx* ctx, float* out);
int vr_read_wait(vr_ctx* ctx);
/* the same for one tile: in image-tiled multi-GPU runs every GPU returns the rows it rendered over its own PCIe link
 * instead of one GPU pulling the whole image. `out` addresses the tile's first pixel in the caller's image (pinned memory,
 * possibly shared between the ranks' processes), rows out_pitch_bytes apart. One read-back in flight per context. */
int vr_read_accum_f32_tile_async(vr_ctx* ctx, const vr_tile* tile_or_null, float* out, uint64_t out_pitch_bytes);
/* the rgba8 pixels of one tile (blocking); `out` addresses the tile's first pixel, rows out_pitch_bytes apart */
int vr_read_image_rgba8_tile(vr_ctx* ctx, const vr_tile* tile_or_null, uint8_t* out, uint64_t out_pitch_bytes);
int vr_read_frame_color_f32(vr_ctx* ctx, float* out);      /* width*height floats: the last frame's clamp()ed colour, R=G=B */
int vr_write_accum_f32(vr_ctx* ctx, const float* in, uint32_t frame_count); /* restore a snapshot (resume) */
int vr_write_image_rgba8(vr_ctx* ctx, const uint8_t* in);
int vr_get_counters(vr_ctx* ctx, vr_counters* out);
int vr_get_timings(vr_ctx* ctx, vr_timings* out);
/* per-pass times summed over every frame finished since the last reset, and the number of frames. Frames are timed
 * through a small ring of CUDA-event sets that is read lazily, so collecting these costs no host synchronisation per
 * frame (this call itself waits for the frames still in flight). reset != 0 clears the sums afterwards. */
int vr_get_timings_sum(vr_ctx* ctx, vr_timings* sum, uint64_t* frames, int reset);
/* device pointers for zero-copy interop (NCCL): float4 accumulator, rgba8 image, frame colour; CUDA stream */
int vr_get_device_buffers(vr_ctx* ctx, void** accum_f32, void** image_rgba8, void** frame_color_f32, void** stream);
/* multi-GPU image tiling: apply the progressive average (shaders/point_compute_color.comp:99-105, current frameCount) to
 * the pixels of `tile` from frame colours that are ALREADY in the frame-colour buffer (vr_get_device_buffers) -- rows
 * another GPU rendered and sent over NVLink. A rank renders its own tile with vr_render_frame(tile), receives the other
 * ranks' frame colours (4 B per pixel instead of the 16 B accumulator) and calls this for the rest of the image. */
int vr_accumulate_frame_color(vr_ctx* ctx, const vr_tile* tile_or_null);
/* fused tile all-gather: `local` (width*height floats, caller-owned; NULL = the context's private buffer) becomes the
 * frame-colour buffer, and peers[0..n) are the addresses of the SAME buffer on the GPUs of the group as mapped into this
 * process (CUDA peer / symmetric memory over NVLink; the entry equal to `local` is skipped). From then on the resolve
 * and path kernels store every pixel's frame colour to the local buffer and to all peers, so that after a cross-GPU
 * barrier each GPU holds the whole frame's colours without a separate collective; it then calls
 * vr_accumulate_frame_color for the rows it did not render. The call does not synchronise: alternate two buffers per
 * frame so that a fast GPU's stores for frame f+1 never land in a buffer a slow GPU still reads for frame f. */
int vr_set_frame_color_peers(vr_ctx* ctx, void* local, void* const* peers, uint32_t n);
/* per-row count of non-zero-density camera samples for the current params (tile balancing across GPUs) */
int vr_row_sample_counts(vr_ctx* ctx, uint64_t* rows_out /* height entries */);
/* 0: re-march camera rays every frame (default, what the reference does); 1: reuse the camera samples while grid,
 * camera, image size and march parameters are unchanged (progressive accumulation) */
int vr_set_march_cache(vr_ctx* ctx, int enabled);

/* 1 (default): with the ray-segment evaluation (variant 8) the camera march runs ONCE per frame: every pixel gets a slot
 * in the sample buffer sized by a lookup-free upper bound of its step count (the march loop's own t sequence), so the
 * counting pass of the two-pass march (count, scan, fill) is not needed; the sample buffer then has gaps and is about
 * S_ref/S_eff times larger. 0: always two passes and a compact buffer. Images, counters and vr_read_samples are identical
 * either way. The exact kernels (variants 0-3), the opt-in physics and frames whose bound total would not fit 32-bit
 * sample indices use the two-pass march regardless. */
int vr_set_march_single_pass(vr_ctx* ctx, int enabled);
/* layout of the last gather's sample buffer (diagnostic): out[0] = 1 single-pass layout with gaps / 0 compact two-pass
 * layout, out[1] = extent of the buffer in samples (slots), out[2] = samples in it */
int vr_get_march_info(vr_ctx* ctx, uint64_t out[3]);

/* 1 (default): once two consecutive vr_render_frame calls had the same inputs, each call also generates the NEXT frame's
 * lights (same UBO, frameCount + 1) into a second light buffer on the auxiliary stream, underneath its evaluation kernel;
 * the next call adopts them if its inputs are what was assumed and regenerates otherwise. Results are identical either
 * way; 0 switches the prefetch off. */
int vr_set_light_prefetch(vr_ctx* ctx, int enabled);

/* 1 (default): the small kernels of the ray-segment evaluation (odd 8-sample segments, the exact kernel over leftovers and
 * near-light segments) run on a second, low-priority stream next to the big segment kernel; 0: everything on the context's
 * stream, one kernel after the other. Results are identical bit for bit (tests/test_gpu_parity.py). */
int vr_set_eval_side_stream(vr_ctx* ctx, int enabled);

/* ---- sample-parallel progressive rendering: G GPUs each render every G-th frame of a progressive run and
 * keep a float SUM of their frame colours; one reduce(sum) and a division replace the sequential running average of
 * shaders/point_compute_color.comp:99-105 (equal up to float reassociation; the rgba8 running average, re-quantised every
 * frame, is sequential by construction and is not available in this mode).
 *   vr_set_frame_step(first, step)  the next vr_render_frame is frame number `first` (0: frameCount + step as usual) and
 *                                 every later one advances frameCount by `step` instead of 1 (src/main.cpp:712): rank r
 *                                 of G calls (r + 1, G) and renders frames r+1, r+1+G, ... with exactly the light
 *                                 lists the sequential run generates for them; (0, 1) restores the default
 *   vr_set_accumulate_mode(SUM)   the float accumulator adds each frame colour instead of averaging; rgba8 is not touched;
 *                                 clear with vr_clear_image (the implicit clear only happens at frameCount == 1)
 *   vr_finish_sum(total)          after the sums were reduced into this context's accumulator (vr_get_device_buffers +
 *                                 NCCL, or vr_group_render_progressive): accumulator = sum / total, rgba8 = UNORM8 of
 *                                 that average, frameCount = total */
enum { VR_ACCUM_AVERAGE = 0, VR_ACCUM_SUM = 1 };
int vr_set_frame_step(vr_ctx* ctx, uint32_t first, uint32_t step);
int vr_set_accumulate_mode(vr_ctx* ctx, int mode);
int vr_finish_sum(vr_ctx* ctx, uint32_t total_frames);

/* nearest-voxel grid values at n index coordinates (ijk[3*n]) through the device traversal: what
 * pnanovdb_readaccessor_get_value_address + pnanovdb_read_float return (shaders/PNanoVDB_GLSL.h:2202-2231, :365-368),
 * i.e. leaf values, tile values at lower/upper/root level, and the background outside every root tile. */
int vr_lookup_values(vr_ctx* ctx, const int32_t* ijk, uint64_t n, float* values_out);

/* ---- opt-in physics. None of these exist in the reference's shaders (its gather uses the nearest
 * voxel, no phase function and no light->sample attenuation: shaders/point_compute_color.comp:60-97,
 * shaders/common_functions.h:186-194); they are what the reference's description of the method mentions, defined here and
 * restated on the CPU in oracle/vr_oracle.cpp (trilinear, physFactor). Everything zero (the default) = the reference's
 * arithmetic. Path mode ignores them.
 *   trilinear            density at a camera sample = nanovdb::math::TrilinearSampler of the index-space position
 *                        (external/openvdb/nanovdb/nanovdb/math/SampleFromVoxels.h:174-260) instead of the nearest voxel
 *   phase_hg, hg_g       every light evaluation is multiplied by the Henyey-Greenstein phase function relative to
 *                        isotropic, (1-g^2)/(1+g^2-2g cos)^(3/2), cos = angle between the light's travel direction at
 *                        the sample and the direction to the camera; |hg_g| < 1
 *   light_transmittance  ... and by exp(-absorptionCoefficient * step * sum of the nearest-voxel values) over
 *                        n = uint(distance/step) index-space steps from the light to the sample (step =
 *                        rayMarchingStepSize; the convention of the path mode's shadow term, path_compute_color.comp:17-28)
 * phase_hg / light_transmittance switch the evaluation to a kernel that keeps every rounding of the CPU restatement; the
 * transmittance term marches the grid once per (sample, light) pair and is orders of magnitude slower than the gather. */
typedef struct vr_options {
    uint32_t trilinear;
    uint32_t phase_hg;
    float hg_g;
    uint32_t light_transmittance;
} vr_options;
int vr_set_options(vr_ctx* ctx, const vr_options* options);
int vr_get_options(vr_ctx* ctx, vr_options* options);
/* test hook: the device trilinear sampler at n index-space positions xyz[3*n] */
int vr_trilinear_values(vr_ctx* ctx, const float* xyz, uint64_t n, float* values_out);
/* calculateRadius of the reference (src/main.cpp:609-618; the call in drawFrame, :713, is commented out there): the
 * progressive radius update r' = sqrt((r^2 * prod_{k=1}^{15} (k + alpha)/k)^(1/16)); a caller that wants the
 * reference's per-frame radius update sets ubo.beamRadius = vr_progressive_radius(ubo.beamRadius, alpha) before each frame. Host arithmetic. */
double vr_progressive_radius(double previous_radius, double alpha);

/* evaluation-kernel variant for point-type light lists (point, ray): 0 = scalar FP32 kernel, 1 = packed FP32x2 kernel
 * with the d2 < 1e-4 guard in the loop, 2 = packed kernel that runs the loop unguarded and repeats a thread block with the
 * guard when one of its threads saw d2 < 1e-4 (same results as 1), 3 = as 2 with 6 samples per thread, 8 = below. All follow
 * shaders/common_functions.h:186-194; results differ by rounding only. The product runs 8 (the default: ray-segment
 * kernel, with variant 2's kernel for near-light segments and leftover samples); 2 is the exact reference for 8's error
 * bound; 0, 1 and 3 are kept as CROSS-CHECKS only (0 keeps every rounding of the shader, the tests compare the others
 * against it) and are not tuned any further. */
int vr_set_eval_variant(vr_ctx* ctx, int variant);
/* variant 8 = ray-segment kernel: the samples of one pixel lie on the camera ray, so eight consecutive samples share the
 * per-light geometry (distance of the light from the ray, foot point) and two lights share one reciprocal. The line
 * model drops the FP32 rounding of the sample positions (~1e-5 world units), a relative error of 2e-5/d on the term of a
 * light at distance d; every segment that comes closer than `distance` to a light (found geometrically before the
 * evaluation; default 0.25, scaled by sqrt(1000 / list length) down to at most 0.1 for long ray/beam sub-point lists) or
 * whose samples deviate from their line by more than distance / 5e3 goes to variant 2's exact kernel, as do the samples
 * that do not fill a segment. Measured pixel error against the exact variants: see tests/test_gpu_parity.py. */
int vr_set_eval_near_distance(vr_ctx* ctx, float distance);
/* bookkeeping of the last variant-8 evaluation: out[0] = full segments, out[1] = leftover samples, out[2] = samples the
 * exact kernel evaluated (leftovers + flagged) */
int vr_get_eval_segment_counts(vr_ctx* ctx, uint64_t out[3]);
/* the camera samples of the last vr_dispatch_compute_color (point/sphere/ray/beam) that reached the light loop, in pixel-major
 * march order: world x, y, z and weight T*val*step per sample (shaders/point_compute_color.comp:72-89). *count receives
 * the total; at most `capacity` samples (4 floats each) are copied. Diagnostic / test hook. */
int vr_read_samples(vr_ctx* ctx, float* out, uint64_t capacity, uint64_t* count);
/* thread blocks of variants >= 2 that had to repeat their light loop with the guard, since vr_create (diagnostic) */
int vr_get_eval_redo_count(vr_ctx* ctx, uint64_t* ctas);

/* ---- CPU_test/main.cpp:25-43, 89-126 as a kernel: fixed camera (0,250,-800) in index space, fov 45, 1200 unit
 * steps, density[j*w+i] = sum of grid values. */
int vr_cpu_test_render(vr_ctx* ctx, uint32_t width, uint32_t height, float* density_out);

/* ---- host-side NanoVDB float-grid writer (fixtures for configs without a .vdb asset; replaces the
 * openvdb -> nanovdb::tools::createNanoGrid step of src/main.cpp:1157-1193 for synthetic volumes).
 * Output is a raw grid buffer in the layout PNanoVDB reads; free with vr_grid_free. */
int vr_grid_build_cloud(int32_t n, uint32_t seed, double voxel_size, const double translation[3], int nthreads,
                        void** bytes_out, uint64_t* size_out);
int vr_grid_build_fog_sphere(double radius_vox, const int32_t center[3], double voxel_size, const double translation[3],
                             int nthreads, void** bytes_out, uint64_t* size_out);
/* dense[(i*dims[1]+j)*dims[2]+k] at index (origin+ijk); values equal to background are inactive */
int vr_grid_build_dense(const float* dense, const int32_t dims[3], const int32_t origin[3], float background,
                        double voxel_size, const double translation[3], int nthreads, void** bytes_out,
                        uint64_t* size_out);
void vr_grid_free(void* bytes);

/* ---- NanoVDB files. The reference reads an OpenVDB .vdb and converts it with createNanoGrid
 * (src/main.cpp:1157-1193); this library takes the converted volume: an uncompressed .nvdb file (segments of
 * FileHeader + FileMetaData + names + grids, external/openvdb/nanovdb/nanovdb/io/IO.h:381-437, NanoVDB.h:6286-6325) or a
 * raw grid buffer. Picks the first float grid, like src/main.cpp:1173-1181, or the float grid called grid_name.
 * ZIP/BLOSC-compressed files are rejected (VR_E_BADGRID). Free the buffer with vr_grid_free. */
int vr_grid_load_nvdb(const char* path, const char* grid_name_or_null, void** bytes_out, uint64_t* size_out);
/* write a grid buffer as a one-grid uncompressed .nvdb (raw == 0; readable by nanovdb::io::readGrid) or as the raw buffer */
int vr_grid_save_nvdb(const char* path, const void* bytes, uint64_t size, int raw);

/* ---- image files (row N1). PPM: binary P6 of the rgba8 image's RGB, the format CPU_test writes
 * (CPU_test/main.cpp:128-132). PFM: little-endian float, channels = 1 (Pf), 3 or 4 (PF, alpha dropped), rows bottom-up. */
int vr_image_write_ppm(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height);
int vr_image_write_pfm(const char* path, const float* data, uint32_t width, uint32_t height, uint32_t channels);

/* ---- snapshots of the progressive state (row N4): UBO incl. frameCount + float4 accumulator + rgba8 image. The
 * reference keeps this state only in the storage image and uboData.frameCount (src/main.cpp:657, :695-698, :937).
 * vr_snapshot_load resizes the context's image if needed; rendering continues at frameCount + 1. */
int vr_snapshot_save(vr_ctx* ctx, const char* path);
int vr_snapshot_load(vr_ctx* ctx, const char* path);
/* message of the last failed file call above on this thread (they have no context to attach it to) */
const char* vr_io_last_error(void);

/* run on a caller-owned CUDA stream (e.g. torch's current stream, so that NCCL collectives and these kernels are
 * ordered on one stream); NULL restores the context's own stream. The caller keeps the stream alive. */
int vr_set_stream(vr_ctx* ctx, void* cuda_stream);
/* number of kernel launches this context has issued since creation */
int vr_get_launch_count(vr_ctx* ctx, uint64_t* launches);

/* ---- measurement helpers: CUDA events on the context's stream (torch.cuda.Event only sees torch's stream) */
int vr_timer_start(vr_ctx* ctx);
int vr_timer_stop(vr_ctx* ctx, float* elapsed_ms); /* records, synchronises and returns the device time since start */
/* microbenchmarks that measure this GPU's non-tensor ceilings for the roofline (FP32 FMA issue, MUFU.RCP):
 * kind 0: dependent FFMA chains -> TFLOP/s (2 flop per FFMA); kind 1: MUFU.RCP -> 1e12 op/s;
 * kind 2: the point-light inner loop with operands in registers (3 FADD, FMUL, 2 FFMA, FSETP, MUFU.RCP, FFMA) -> 1e12 eval/s
 * kind 3: dependent FFMA2 (packed FP32x2) chains -> TFLOP/s; kind 4: the packed point-light inner loop -> 1e12 eval/s */
int vr_ubench(vr_ctx* ctx, int kind, double* result);

/* ---- several GPUs of one node from ONE process (the reference is single-GPU, src/main.cpp:265-270).
 * A group owns one context per GPU, each driven by its own host thread, with CUDA peer access between all of them.
 * Frames are image-tiled: GPU i renders a band of rows (grid, UBO and light list replicated; light generation is
 * deterministic and repeated on every GPU) and keeps the progressive accumulator of that band -- the path has no exchange
 * step; the bands meet where the image is consumed (vr_group_read_*: every GPU returns its rows over its own PCIe link).
 * VR_EXCHANGE_ALLGATHER additionally leaves the whole image on every GPU after every frame: the resolve kernel stores
 * the band's frame colours to all GPUs through peer mappings over NVLink, a flag-based device-side barrier orders them,
 * every GPU averages the rows it did not render. Results are bit-identical to a one-GPU render in both modes.
 * Calls are issued by one thread at a time; vr_group_ctx gives the per-GPU context for everything not wrapped here
 * (timings, counters, options): set those identically on every GPU. */
typedef struct vr_group vr_group;
enum { VR_EXCHANGE_NONE = 0, VR_EXCHANGE_ALLGATHER = 1 };
/* cuda_devices == NULL: devices 0..n-1. A device may be listed more than once (several bands on one GPU: how the group
 * logic is tested on a one-GPU box). */
int vr_group_create(const int* cuda_devices, uint32_t n, vr_group** out);
int vr_group_destroy(vr_group* g);
const char* vr_group_last_error(const vr_group* g); /* g may be NULL: error of a failed vr_group_create on this thread */
uint32_t vr_group_size(const vr_group* g);
vr_ctx* vr_group_ctx(vr_group* g, uint32_t i);
/* the volume crosses PCIe once (to the first GPU) and is replicated GPU to GPU over NVLink (vr_set_grid_device) */
int vr_group_set_grid(vr_group* g, const void* nanovdb_bytes, uint64_t size);
int vr_group_resize(vr_group* g, uint32_t width, uint32_t height);
int vr_group_set_params(vr_group* g, const vr_ubo* params);
int vr_group_get_params(vr_group* g, vr_ubo* params);
int vr_group_set_max_lights_capacity(vr_group* g, uint32_t n);
int vr_group_set_light_gen_groups(vr_group* g, uint32_t gx, uint32_t gy, uint32_t gz, vr_lightgen_variant variant);
int vr_group_set_exchange(vr_group* g, int mode);
int vr_group_reset_accumulation(vr_group* g);
/* row bands, one per GPU in GPU order: contiguous, non-empty, covering the image. vr_group_balance chooses them: a per-row
 * cost model from one measured frame (evaluation time follows the row's non-zero-density samples, the march's the row),
 * then `rounds` rounds of measured feedback (each GPU reports the device time of its band); the best split seen is kept.
 * It renders untimed frames, resets the accumulation and restores the parameters. */
int vr_group_set_bands(vr_group* g, const uint32_t* y0, const uint32_t* h);
int vr_group_get_bands(vr_group* g, uint32_t* y0, uint32_t* h);
int vr_group_balance(vr_group* g, vr_algo algo, uint32_t rounds);
/* the compute section of drawFrame (src/main.cpp:712-837) on every GPU, each for its band; returns once enqueued */
int vr_group_render_frame(vr_group* g, vr_algo algo);
int vr_group_sync(vr_group* g);
int vr_group_read_accum_f32(vr_group* g, float* out);       /* width*height*4 floats; every GPU copies its own rows */
int vr_group_read_accum_f32_async(vr_group* g, float* out); /* pinned memory; valid after vr_group_read_wait */
int vr_group_read_wait(vr_group* g);
int vr_group_read_image_rgba8(vr_group* g, uint8_t* out);
/* sample-parallel progressive run (vr_set_frame_step / vr_set_accumulate_mode above): frames 1..frames go round-robin to
 * the GPUs, each renders WHOLE images into a float sum; the first GPU adds the others' sums in GPU order (peer reads over
 * NVLink), divides, and holds the result (float accumulator, rgba8 of the average, frameCount = frames). Until the next
 * vr_group_reset_accumulation (or resize / balance) the vr_group_read_* calls return that image from the first GPU, and
 * vr_group_render_frame / vr_group_time_frames return VR_E_STATE: the other GPUs hold partial sums and other frame numbers,
 * tiled frames cannot continue from them. */
int vr_group_render_progressive(vr_group* g, vr_algo algo, uint32_t frames);
/* device time of `frames` tiled frames: CUDA events on every GPU's stream, maximum over the GPUs; per_gpu_ms may be NULL */
int vr_group_time_frames(vr_group* g, vr_algo algo, uint32_t frames, float* ms_per_frame_max, float* per_gpu_ms);

/* library facts */
const char* vr_version(void);

#ifdef __cplusplus
}
#endif
#endif /* VR_H */
