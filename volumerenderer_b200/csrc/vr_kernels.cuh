// vr_kernels.cuh -- launch interface between the C-ABI layer (vr_api.cu) and the sm_100a kernels (vr_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vr.h"
#include "vr_device.cuh"

// one camera-ray sample that reaches the light loop: world position + weight T*val*step
// (shaders/point_compute_color.comp:72-89)
struct VrSample {
    float x, y, z, w;
};
// one entry of the expanded light list the evaluation kernel loops over: position + intensity/(4*pi)
struct VrEvalLight {
    float x, y, z, i4pi;
};

struct VrLightCounter {  // binding 4 (shaders/common_bindings.h:14-17) + bookkeeping
    uint32_t lightCount;
    uint32_t contributionCounter;
    unsigned long long totalEmitted;  // scatter events before the maxLights cap
};

// where a pixel's frame colour goes: the local buffer and, in image-tiled multi-GPU runs, the same buffer on every
// other GPU, addressed through peer mappings over NVLink (the all-gather of the tiles is fused into the kernel's stores)
#define VR_MAX_PEERS 16
struct VrFrameColorOut {
    float* local;
    float* peer[VR_MAX_PEERS];  // other GPUs' buffers as mapped into this process
    uint32_t npeers;
};

// flag arrays of the in-process cross-GPU barrier (vr_group): flags[g] is GPU g's array of VR_MAX_PEERS epochs (+ one word
// that a timed-out wait sets), mapped into every GPU of the group through CUDA peer access
struct VrPeerFlags {
    uint32_t* flags[VR_MAX_PEERS];
};

struct VrWorkCounters {
    unsigned long long S_ref, S_eff;
};

// ---- light generation (shaders/light_gen.comp and variants): one walk per photon into a private slab of
// vrk_lightgen_local_capacity() lights, scan, compaction in canonical photon order, second walk only for photons that
// overflowed the slab, counter update. slab holds nphot * capacity * 2 float4; scan_tmp ceil(nphot/4096)+1 entries.
void vrk_lightgen(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, int keep_origin, uint32_t nx, uint32_t ny, uint32_t nz,
                  uint32_t* counts, unsigned long long* offsets, unsigned long long* total, unsigned long long* scan_tmp,
                  float4* slab, vr_ray_light* lights, VrLightCounter* counter);
uint32_t vrk_lightgen_local_capacity();
// ---- exclusive scan of u32 -> u64 (+ grand total), n up to 2^31; tmp holds ceil(n/4096)+1 entries
void vrk_scan_u32(cudaStream_t st, const uint32_t* in, unsigned long long* out, unsigned long long* total, uint64_t n,
                  unsigned long long* tmp);
// ---- light list expansion (ray/beam sub-points: shaders/ray_compute_color.comp:13-20). Every kernel is launched over
// the buffer capacity and reads lightCount on the device; out = VrEvalLight list, out2 = duplicated (x,x,y,y)(z,z,w,w)
// list for the packed kernels; *neval = list length (prepare_simple) / scan total (expand).
void vrk_prepare_simple(cudaStream_t st, const vr_ray_light* lights, const VrLightCounter* counter, int use_from,
                        VrEvalLight* out, float4* out2, uint32_t capacity, unsigned long long* neval);
void vrk_prepare_count_steps(cudaStream_t st, const vr_ray_light* lights, const VrLightCounter* counter, float lrs,
                             uint32_t* steps, uint32_t capacity);
void vrk_prepare_expand(cudaStream_t st, const vr_ray_light* lights, const VrLightCounter* counter, float lrs,
                        const unsigned long long* offsets, VrEvalLight* out, float4* out2, uint32_t capacity,
                        uint64_t out_capacity, uint32_t* overflow);
// ---- camera-ray march (shaders/point_compute_color.comp:60-94 minus the light loop)
// mode 0: count per pixel (+ work counters); 1: fill at offsets (compact buffer, offsets = scan of the counts);
// 2: single pass, counts + fill at offsets = scan of vrk_march_bound's per-pixel upper bounds (buffer with gaps)
void vrk_march(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, int mode, int trilinear,
               uint32_t* counts, const unsigned long long* offsets, VrSample* samples, VrWorkCounters* wc);
void vrk_march_bound(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, uint32_t* bound);
void vrk_compact_samples(cudaStream_t st, const uint32_t* counts, const unsigned long long* offsets, const unsigned long long* compact_offsets,
                         uint64_t npix, const VrSample* samples, VrSample* out);
// ---- opt-in physics (vr_options): evaluation with the phase / transmittance factor, and the trilinear test hook
void vrk_eval_phys(cudaStream_t st, int sphere, const VrGridDev& g, const vr_ubo& u, const VrPhysics& phys, const VrSample* samples,
                   uint64_t nsamples, const VrEvalLight* lights, const unsigned long long* nlights_dev, float* out);
void vrk_trilinear(cudaStream_t st, const VrGridDev& g, const float* xyz, uint64_t n, float* out);
// ---- the hot kernels: every sample against every entry of the expanded light list (length read from *nlights_dev)
// scalar FP32 kernels (point / sphere), every rounding of the shader kept
void vrk_eval(cudaStream_t st, int sphere, const VrSample* samples, uint64_t nsamples, const VrEvalLight* lights,
              const unsigned long long* nlights_dev, float radius, float* out);
// packed FP32x2 kernels for point-type lists. variant 1: guard in the loop (k_eval_point2); 2, 3: speculative
// k_eval_point3 with 8 / 6 samples per thread; redo_count (device, may be NULL) counts CTAs that repeated with the guard
void vrk_eval_point_packed(cudaStream_t st, int variant, const VrSample* samples, uint64_t nsamples, const float4* lights2,
                           const unsigned long long* nlights_dev, float* out, unsigned long long* redo_count);
// ray-segment kernel (variant 8; point-type lists, and sphere-type lists with sphere = 1): eight consecutive samples of one camera ray share the per-light
// geometry (k_eval_seg); what does not fill a segment, and every sample whose thread came closer than the segment's
// rho to a light (k_flag_near_segments), is evaluated by the exact kernel over an index list (k_eval_point3 / k_eval_sphere2 in LIST mode).
struct VrSegLists {
    // a pixel's q = count / 8 full segments are numbered consecutively (seg_base[pixel] + k): that number indexes `flags`.
    // Pairs of them (16 consecutive samples) are the entries of the seg16 table, an odd last one goes to the seg table.
    uint32_t *seg16_first, *seg16_pix, *seg16_flag;  // first sample / pixel (tile-local, row-major) / flag number
    uint32_t *seg_first, *seg_pix, *seg_flag;
    uint32_t* seg_base;           // per pixel: flag number of its first segment
    uint8_t* flags;               // per segment: 1 = comes near a light, evaluate exactly (k_flag_near_segments)
    uint32_t* flag_all;           // 1 = a light sits next to the camera plane: every segment goes to the exact kernel
    uint32_t* list;               // samples for the exact kernel: leftovers, then the near-light entries (k_list_flagged)
    uint32_t* list2;              // samples the segment kernels hand over themselves (residual check): evaluated after the join
    unsigned long long* counts;   // [0] seg entries, [1] leftovers, [2] list length (leftovers + near-light), [3] flag_all,
                                  // [4] seg16 entries, [5] segments (flag numbers), [6] unused, [7] list2 length
    uint32_t list_capacity;       // of list and of list2
    // capacities, read by the bounds checks of the VR_CHECKED build only: entries of the seg16 / seg tables, flags,
    // words per sched array, floats of part16 / part8, slots of the sample buffer
    uint32_t cap16, cap8, cap_flags, sched_words;
    uint64_t part16_floats, part8_floats, nslots;
    // work distribution of k_eval_seg: [1 + c] = slice units of chunk c that have delivered their partial sums (sched16[0] is
    // flag_all); one array per
    // table (the kernels of both run concurrently), allocated back to back (sched_bytes from sched16), zeroed per pass
    uint32_t *sched16, *sched8;
    // scratch of the slice units: [chunk of the last wave][slice][sample of the entry][entry]: vrk_seg_part_floats(len) floats
    float *part16, *part8;
    uint64_t sched_bytes;
};
// builds seg_first / seg_pix / the leftover part of list from the per-pixel sample counts and offsets (nsamples < 2^32)
// pairs != 0: pairs of segments go to the 16-sample table (point-type lists); 0: every segment to the 8-sample table
// single-pass march (mode 2 of vrk_march) that also writes the segment tables (= vrk_march + vrk_build_segments in one
// launch); the tables must have been sized for this march's sample count beforehand
void vrk_march_build(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, int trilinear,
                     uint32_t* counts, const unsigned long long* offsets, VrSample* samples, VrWorkCounters* wc, const VrSegLists& sl,
                     int pairs);
// floats of VrSegLists::part16 (len 16) / part8 (len 8) on the current device
uint64_t vrk_seg_part_floats(int len);
void vrk_build_segments(cudaStream_t st, const uint32_t* counts, const unsigned long long* offsets, uint64_t npix, const VrSegLists& sl,
                        int pairs);
// lights / lights2: the same list in both layouts; nlights_capacity sizes the flagging launch (one block per light)
// side, side2 (may be NULL = st): lower-priority streams for the slice units and the small kernels (they run next to the
// full units of the big kernel and fill its tail); ev_fork / ev_join / ev_join2 order the streams
void vrk_eval_point_seg(cudaStream_t st, cudaStream_t side, cudaStream_t side2, cudaEvent_t ev_fork, cudaEvent_t ev_join, cudaEvent_t ev_join2, const VrGridDev& g,
                        const vr_ubo& u, float tan_scale, vr_tile tile, const VrSample* samples,
                        uint64_t nsamples, uint64_t npix, const uint32_t* counts, const unsigned long long* offsets, const VrSegLists& sl,
                        const VrEvalLight* lights, const float4* lights2, uint64_t nlights_capacity,
                        const unsigned long long* nlights_dev, float rho, int sphere, float radius, float* out,
                        unsigned long long* redo_count);
// packed FP32x2 kernel for sphere lights / beam sub-points (common_functions.h:182-201)
void vrk_eval_sphere_packed(cudaStream_t st, const VrSample* samples, uint64_t nsamples, const float4* lights2,
                            const unsigned long long* nlights_dev, float radius, float* out);
// ---- per-pixel reduction + clamp + progressive average (shaders/point_compute_color.comp:97-105)
void vrk_resolve(cudaStream_t st, const vr_ubo& u, vr_tile tile, const uint32_t* counts, const unsigned long long* offsets,
                 const float* contrib, const VrLightCounter* counter, const VrFrameColorOut& frame_color, float4* accum, uchar4* rgba8,
                 int sum_mode);
// running average (same arithmetic as vrk_resolve) of frame colours that are already in frame_color for this tile
void vrk_accumulate(cudaStream_t st, const vr_ubo& u, vr_tile tile, const float* frame_color, float4* accum, uchar4* rgba8, int sum_mode);
// sample-parallel progressive mode (vr_set_accumulate_mode): accumulator = sum / total, rgba8 from that average; dst += src
void vrk_finish_sum(cudaStream_t st, float4* accum, uchar4* rgba8, uint64_t npix, float total);
void vrk_add_accum(cudaStream_t st, float4* dst, const float4* src, uint64_t npix);
// in-process cross-GPU barrier on `st` (GPU `self` of n, epoch counts up from 1)
void vrk_peer_barrier(cudaStream_t st, const VrPeerFlags& f, uint32_t self, uint32_t n, uint32_t epoch);
// ---- path mode (shaders/path_compute_color.comp), fully fused per pixel
// tab (27 * (nmax + 1) floats, may be NULL) receives the shadow-term table: light after n multiplies, per neighbour voxel
void vrk_path(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, const VrFrameColorOut& frame_color,
              float4* accum, uchar4* rgba8, VrWorkCounters* wc, float* tab, uint32_t nmax, int sum_mode);
// ---- CPU_test/main.cpp:25-43, 89-126
void vrk_cpu_test(cudaStream_t st, const VrGridDev& g, uint32_t w, uint32_t h, float aspect, float scale, float* density);
// ---- traversal test hook
void vrk_lookup(cudaStream_t st, const VrGridDev& g, const int32_t* ijk, uint64_t n, float* out);
// ---- helpers
void vrk_clear_image(cudaStream_t st, float4* accum, uchar4* rgba8, float* frame_color, uint64_t npix);
void vrk_row_sums(cudaStream_t st, const uint32_t* counts, uint32_t w, uint32_t h, unsigned long long* rows);
// ---- microbenchmarks (roofline denominators); returns the op count launched
double vrk_ubench(cudaStream_t st, int kind, float* sink, int iters);
// VR_CHECKED build (-DVR_CHECKED; tools/gpu_checked.sh): every index the segment path computes is checked against its
// buffer's capacity on the device; a failed check sets a bit (see VR_CHK_* in vr_kernels.cu). Returns the mask and clears it;
// always 0 in the product build, which carries no checks.
unsigned int vrk_check_flags();
// number of kernels launched by the vrk_* functions in this process (monotonic)
unsigned long long vrk_launch_count();
