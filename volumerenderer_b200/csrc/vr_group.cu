// vr_group.cu -- several GPUs of one node driven from ONE process through the C ABI (include/vr.h, vr_group_*).
//
// The reference is single-GPU, one queue (src/main.cpp:265-270); this is the C/C++ host side of SURVEY.md section 8e:
//   * image-tiled frames: GPU i renders a band of rows with the grid, UBO and light list replicated (light generation is
//     deterministic and simply repeated) and keeps the progressive accumulator of that band: pixels are independent, so
//     the path has no exchange step (VR_EXCHANGE_NONE, the default); the bands meet where the image is consumed
//     (vr_group_read_*: every GPU returns its rows over its own PCIe link).
//     VR_EXCHANGE_ALLGATHER is for consumers that want the whole image on every GPU after every frame (a display GPU):
//     k_resolve stores the band's frame colours into every GPU's frame-colour buffer through CUDA peer mappings over
//     NVLink (vr_set_frame_color_peers: the all-gather is the kernel's own epilogue), a flag-based device-side barrier
//     orders them (k_peer_barrier), and every GPU averages the rows it did not render from its local copy
//     (vr_accumulate_frame_color) -- no host synchronisation and no collective library per frame;
//   * band balancing: a per-row cost model from one measured frame, refined by measured feedback (same arithmetic as
//     volumerenderer_b200/tiles.py, which drives the one-process-per-GPU variant under torchrun);
//   * the volume crosses PCIe once (GPU 0) and is replicated GPU to GPU (vr_set_grid_device from a peer buffer);
//   * read-back: every GPU returns its own band over its own PCIe link;
//   * sample-parallel progressive runs: frame f on GPU f mod n, float sums, one reduce to GPU 0, one division.
// One worker thread per GPU issues that GPU's calls (a context is used by one thread at a time), so the host-side launch
// work of the n GPUs overlaps.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/vr.h"
#include "vr_kernels.cuh"

namespace {

thread_local std::string g_group_create_error;

// one thread per GPU; run() hands it a job and returns, wait() joins the job
struct Worker {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<int()> job;
    bool has_job = false, done = true, quit = false;
    bool threw = false;  // the last job ended in a C++ exception (out of host memory)
    int rc = 0;
    void loop() {
        for (;;) {
            std::function<int()> j;
            {
                std::unique_lock<std::mutex> l(m);
                cv.wait(l, [&] { return has_job || quit; });
                if (quit && !has_job) return;
                j = std::move(job);
                has_job = false;
            }
            int r = VR_OK;
            bool t = false;
            try {  // no exception leaves the thread (std::bad_alloc from a job's containers would end the process)
                r = j();
            } catch (...) {
                r = VR_E_STATE, t = true;
            }
            {
                std::lock_guard<std::mutex> l(m);
                rc = r;
                threw = t;
                done = true;
            }
            cv.notify_all();
        }
    }
    void run(std::function<int()> j) {
        {
            std::lock_guard<std::mutex> l(m);
            job = std::move(j);
            has_job = true;
            done = false;
        }
        cv.notify_all();
    }
    int wait() {
        std::unique_lock<std::mutex> l(m);
        cv.wait(l, [&] { return done; });
        return rc;
    }
    void stop() {
        {
            std::lock_guard<std::mutex> l(m);
            quit = true;
        }
        cv.notify_all();
        if (th.joinable()) th.join();
    }
};

}  // namespace

struct vr_group {
    uint32_t n = 0;
    std::vector<int> dev;
    std::vector<vr_ctx*> ctx;
    std::vector<Worker*> worker;
    std::vector<cudaStream_t> stream;
    std::string err;
    uint32_t width = 0, height = 0;
    std::vector<vr_tile> band;                 // rows GPU i renders
    std::vector<float*> fc[2];                 // two frame-colour buffers per GPU, alternating by frame
    std::vector<uint32_t*> flags;              // barrier flag arrays, one per GPU
    VrPeerFlags peer_flags{};
    uint32_t epoch = 0, parity = 0;
    int exchange = VR_EXCHANGE_NONE;
    // A device listed more than once (bands on one GPU): its contexts' streams may share a hardware queue, where a
    // spinning barrier kernel would block the very peer it waits for. Such groups order the exchange with CUDA events
    // (recorded after each GPU's resolve, waited for by every other GPU's stream) in two host phases instead.
    bool shared_device = false;
    std::vector<cudaEvent_t> stored[2];        // per parity, per GPU: its band's frame colours are stored everywhere
    uint8_t* staging[2] = {nullptr, nullptr};  // the volume on GPU 0, source of the GPU-to-GPU replication
    uint64_t staging_cap[2] = {0, 0};
    std::vector<cudaEvent_t> staged_read[2];   // per GPU: its copy out of staging[k] is done
    uint32_t staging_next = 0;
    bool have_grid = false;
    // vr_group_render_progressive leaves the finished image (average, rgba8, frameCount) on GPU 0 only; the other GPUs
    // hold raw partial sums and their own frame numbers. While set, the read-backs take the whole image from GPU 0 and
    // tiled frames are refused (they would continue from mixed states) until vr_group_reset_accumulation.
    bool progressive_result = false;
};

namespace {

int gfail(vr_group* g, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (g) g->err = buf;
    else g_group_create_error = buf;
    return code;
}

// fn(i) on every GPU's worker thread; the first failure is reported with that context's message
int on_all(vr_group* g, const std::function<int(uint32_t)>& fn) {
    for (uint32_t i = 0; i < g->n; i++) g->worker[i]->run([&fn, i] { return fn(i); });
    int first = VR_OK;
    uint32_t who = 0;
    for (uint32_t i = 0; i < g->n; i++) {
        const int rc = g->worker[i]->wait();
        if (rc != VR_OK && first == VR_OK) first = rc, who = i;
    }
    if (first != VR_OK) {
        if (g->worker[who]->threw) return gfail(g, first, "GPU %u (device %d): out of host memory (C++ exception in the worker)", who, g->dev[who]);
        const char* e = vr_last_error(g->ctx[who]);
        return gfail(g, first, "GPU %u (device %d): %s", who, g->dev[who], e && *e ? e : "failed");
    }
    return VR_OK;
}

void free_frame_colors(vr_group* g) {
    for (int p = 0; p < 2; p++) {
        for (uint32_t i = 0; i < g->fc[p].size(); i++)
            if (g->fc[p][i]) {
                cudaSetDevice(g->dev[i]);
                cudaFree(g->fc[p][i]);
            }
        g->fc[p].clear();
    }
}

// ---- band balancing: the arithmetic of volumerenderer_b200/tiles.py (row_costs, balanced_row_bands, refine_row_costs)
std::vector<vr_tile> split_rows(const std::vector<double>& cost_in, uint32_t parts, uint32_t width) {
    std::vector<double> cost = cost_in;
    const uint32_t H = (uint32_t)cost.size();
    double total = 0.0;
    for (double c : cost) total += c;
    if (!(total > 0.0)) {
        std::fill(cost.begin(), cost.end(), 1.0);
        total = H;
    }
    std::vector<double> prefix(H + 1, 0.0);
    for (uint32_t y = 0; y < H; y++) prefix[y + 1] = prefix[y] + cost[y];
    std::vector<uint32_t> cuts{0};
    for (uint32_t p = 1; p < parts; p++) {
        const double target = total * p / parts;
        uint32_t y = (uint32_t)(std::lower_bound(prefix.begin(), prefix.end(), target) - prefix.begin());
        if (y > 0 && fabs(prefix[y - 1] - target) <= fabs(prefix[std::min(y, H)] - target)) y -= 1;
        y = std::max(y, cuts.back() + 1);
        y = std::min(y, H - (parts - p));
        cuts.push_back(y);
    }
    cuts.push_back(H);
    std::vector<vr_tile> out;
    for (uint32_t p = 0; p < parts; p++) out.push_back(vr_tile{0, cuts[p], width, cuts[p + 1] - cuts[p]});
    return out;
}

void refine_costs(std::vector<double>& cost, const std::vector<vr_tile>& bands, const std::vector<double>& ms) {
    double total_model = 0.0, total_ms = 0.0;
    for (double c : cost) total_model += c;
    for (double t : ms) {
        if (!(t > 0.0)) return;
        total_ms += t;
    }
    for (size_t r = 0; r < bands.size(); r++) {
        double model = 0.0;
        for (uint32_t y = bands[r].y0; y < bands[r].y0 + bands[r].h; y++) model += cost[y];
        if (!(model > 0.0)) continue;
        const double scale = (ms[r] / total_ms) / (model / total_model);
        for (uint32_t y = bands[r].y0; y < bands[r].y0 + bands[r].h; y++) cost[y] *= scale;
    }
}

int require_image(vr_group* g) {
    if (!g->width) return gfail(g, VR_E_STATE, "no image: call vr_group_resize first");
    return VR_OK;
}

// the rows GPU i returns in a read-back: its band, or, after a sample-parallel progressive run, everything from GPU 0
vr_tile read_rows(const vr_group* g, uint32_t i) {
    if (g->progressive_result) return i == 0 ? vr_tile{0, 0, g->width, g->height} : vr_tile{0, 0, 0, 0};
    return g->n > 1 ? g->band[i] : vr_tile{0, 0, g->width, g->height};
}

// the rows GPU i did not render: at most one rectangle above and one below its band
void complement(const vr_group* g, uint32_t i, vr_tile out[2], int* n) {
    const vr_tile& b = g->band[i];
    *n = 0;
    if (b.y0 > 0) out[(*n)++] = vr_tile{0, 0, g->width, b.y0};
    if (b.y0 + b.h < g->height) out[(*n)++] = vr_tile{0, b.y0 + b.h, g->width, g->height - (b.y0 + b.h)};
}

// the rows GPU i received: progressive average from its local copy of the frame colours
int accumulate_rest(vr_group* g, uint32_t i) {
    vr_tile rest[2];
    int nrest = 0, rc = VR_OK;
    complement(g, i, rest, &nrest);
    for (int k = 0; k < nrest && rc == VR_OK; k++) rc = vr_accumulate_frame_color(g->ctx[i], &rest[k]);
    return rc;
}

// one tiled frame on GPU i (worker thread i): own band with peer stores, barrier, the rest of the image from the local copy
int frame_on(vr_group* g, uint32_t i, vr_algo algo, uint32_t parity, uint32_t epoch) {
    vr_ctx* c = g->ctx[i];
    int rc = VR_OK;
    const bool gather = g->n > 1 && g->exchange == VR_EXCHANGE_ALLGATHER;
    if (g->n > 1 && !gather) {
        rc = vr_set_frame_color_peers(c, nullptr, nullptr, 0);
        if (rc != VR_OK) return rc;
        return vr_render_frame(c, algo, &g->band[i]);  // owner computes: this GPU keeps the average of its band only
    }
    if (gather) {
        std::vector<void*> peers(g->n);
        for (uint32_t j = 0; j < g->n; j++) peers[j] = g->fc[parity][j];
        rc = vr_set_frame_color_peers(c, g->fc[parity][i], peers.data(), g->n);
        if (rc != VR_OK) return rc;
    }
    rc = vr_render_frame(c, algo, g->n > 1 ? &g->band[i] : nullptr);
    if (rc != VR_OK || g->n == 1) return rc;
    if (g->shared_device) {  // second phase (after every GPU recorded its event): vr_group_render_frame
        cudaSetDevice(g->dev[i]);
        return cudaEventRecord(g->stored[parity][i], g->stream[i]) == cudaSuccess ? VR_OK : VR_E_CUDA;
    }
    vrk_peer_barrier(g->stream[i], g->peer_flags, i, g->n, epoch);
    return accumulate_rest(g, i);
}

}  // namespace

extern "C" {

const char* vr_group_last_error(const vr_group* g) { return g ? g->err.c_str() : g_group_create_error.c_str(); }

int vr_group_destroy(vr_group* g) {
    if (!g) return VR_OK;
    for (uint32_t i = 0; i < g->ctx.size(); i++)
        if (g->ctx[i]) vr_sync(g->ctx[i]);
    for (Worker* w : g->worker)
        if (w) {
            w->stop();
            delete w;
        }
    for (uint32_t i = 0; i < g->ctx.size(); i++)
        if (g->ctx[i]) vr_destroy(g->ctx[i]);
    free_frame_colors(g);
    for (uint32_t i = 0; i < g->flags.size(); i++)
        if (g->flags[i]) {
            cudaSetDevice(g->dev[i]);
            cudaFree(g->flags[i]);
        }
    if (!g->dev.empty()) cudaSetDevice(g->dev[0]);
    for (int p = 0; p < 2; p++)
        for (cudaEvent_t e : g->stored[p])
            if (e) cudaEventDestroy(e);
    for (int k = 0; k < 2; k++) {
        if (g->staging[k]) cudaFree(g->staging[k]);
        for (cudaEvent_t e : g->staged_read[k])
            if (e) cudaEventDestroy(e);
    }
    delete g;
    return VR_OK;
}

int vr_group_create(const int* cuda_devices, uint32_t n, vr_group** out) {
    if (!out) return gfail(nullptr, VR_E_ARG, "vr_group_create: out is NULL");
    *out = nullptr;
    if (n == 0 || n > VR_MAX_PEERS) return gfail(nullptr, VR_E_ARG, "vr_group_create: 1..%d GPUs", VR_MAX_PEERS);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return gfail(nullptr, VR_E_CUDA, "no CUDA device available: libvr has no CPU fallback");
    vr_group* g = new (std::nothrow) vr_group();
    if (!g) return gfail(nullptr, VR_E_CUDA, "out of host memory");
    g->n = n;
    for (uint32_t i = 0; i < n; i++) {
        const int d = cuda_devices ? cuda_devices[i] : (int)i;
        if (d < 0 || d >= ndev) {  // a device may be listed more than once: several bands on one GPU (tests on a one-GPU box)
            gfail(nullptr, VR_E_ARG, "vr_group_create: device %d out of range [0,%d)", d, ndev);
            vr_group_destroy(g);
            return VR_E_ARG;
        }
        g->dev.push_back(d);
    }
    g->ctx.assign(n, nullptr);
    g->stream.assign(n, nullptr);
    g->flags.assign(n, nullptr);
    g->band.assign(n, vr_tile{0, 0, 0, 0});
    // every GPU maps every other GPU's memory (NVLink / NVSwitch): kernels store to peers with plain pointers
    for (uint32_t i = 0; i < n; i++) {
        cudaSetDevice(g->dev[i]);
        for (uint32_t j = 0; j < n; j++) {
            if (g->dev[i] == g->dev[j]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, g->dev[i], g->dev[j]);
            if (!can) {
                gfail(nullptr, VR_E_CUDA, "GPU %d cannot map GPU %d's memory (no peer access)", g->dev[i], g->dev[j]);
                vr_group_destroy(g);
                return VR_E_CUDA;
            }
            const cudaError_t e = cudaDeviceEnablePeerAccess(g->dev[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                gfail(nullptr, VR_E_CUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", g->dev[i], g->dev[j], cudaGetErrorString(e));
                vr_group_destroy(g);
                return VR_E_CUDA;
            }
            cudaGetLastError();
        }
    }
    for (uint32_t i = 0; i < n; i++) {
        int rc = vr_create(g->dev[i], &g->ctx[i]);
        if (rc != VR_OK) {
            gfail(nullptr, rc, "vr_create(device %d): %s", g->dev[i], vr_last_error(nullptr));
            vr_group_destroy(g);
            return rc;
        }
        void* st = nullptr;
        vr_get_device_buffers(g->ctx[i], nullptr, nullptr, nullptr, &st);
        g->stream[i] = (cudaStream_t)st;
        cudaSetDevice(g->dev[i]);
        if (cudaMalloc(&g->flags[i], (VR_MAX_PEERS + 1) * sizeof(uint32_t)) != cudaSuccess ||
            cudaMemset(g->flags[i], 0, (VR_MAX_PEERS + 1) * sizeof(uint32_t)) != cudaSuccess) {
            gfail(nullptr, VR_E_CUDA, "barrier flags: out of device memory");
            vr_group_destroy(g);
            return VR_E_CUDA;
        }
        g->peer_flags.flags[i] = g->flags[i];
    }
    for (int k = 0; k < 2; k++) g->staged_read[k].assign(n, nullptr), g->stored[k].assign(n, nullptr);
    for (uint32_t i = 0; i < n; i++) {
        cudaSetDevice(g->dev[i]);
        for (int k = 0; k < 2; k++) {
            cudaEventCreateWithFlags(&g->staged_read[k][i], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&g->stored[k][i], cudaEventDisableTiming);
        }
        for (uint32_t j = 0; j < i; j++) g->shared_device = g->shared_device || g->dev[i] == g->dev[j];
    }
    for (uint32_t i = 0; i < n; i++) {
        Worker* w = new (std::nothrow) Worker();
        if (!w) {
            gfail(nullptr, VR_E_CUDA, "out of host memory");
            vr_group_destroy(g);
            return VR_E_CUDA;
        }
        g->worker.push_back(w);
        w->th = std::thread([w] { w->loop(); });
    }
    *out = g;
    return VR_OK;
}

uint32_t vr_group_size(const vr_group* g) { return g ? g->n : 0; }
vr_ctx* vr_group_ctx(vr_group* g, uint32_t i) { return (g && i < g->n) ? g->ctx[i] : nullptr; }

int vr_group_set_grid(vr_group* g, const void* bytes, uint64_t size) {
    if (!g) return VR_E_ARG;
    if (!bytes || size < 1024) return gfail(g, VR_E_ARG, "vr_group_set_grid: NULL or tiny buffer");
    if (g->n == 1) {
        const int rc = vr_set_grid(g->ctx[0], bytes, size);
        if (rc != VR_OK) return gfail(g, rc, "%s", vr_last_error(g->ctx[0]));
        g->have_grid = true;
        return VR_OK;
    }
    // host -> GPU 0 once; the other GPUs pull it from there over NVLink, each on its own stream, in front of its next frame
    const uint32_t k = g->staging_next & 1u;
    cudaSetDevice(g->dev[0]);
    if (g->staging_next >= 2)
        for (uint32_t i = 0; i < g->n; i++) cudaEventSynchronize(g->staged_read[k][i]);  // two grids ago: every copy out of it is done
    if (g->staging_cap[k] < size) {
        if (g->staging[k]) cudaFree(g->staging[k]);
        g->staging[k] = nullptr;
        if (cudaMalloc(&g->staging[k], size) != cudaSuccess) return gfail(g, VR_E_CUDA, "vr_group_set_grid: out of device memory");
        g->staging_cap[k] = size;
    }
    if (cudaMemcpy(g->staging[k], bytes, size, cudaMemcpyHostToDevice) != cudaSuccess) return gfail(g, VR_E_CUDA, "vr_group_set_grid: upload failed");
    uint8_t* src = g->staging[k];
    const int rc = on_all(g, [=](uint32_t i) {
        int r = vr_set_grid_device(g->ctx[i], src, size, bytes, size);
        if (r == VR_OK) {
            cudaSetDevice(g->dev[i]);
            cudaEventRecord(g->staged_read[k][i], g->stream[i]);
        }
        return r;
    });
    if (rc != VR_OK) return rc;
    g->staging_next++;
    g->have_grid = true;
    return VR_OK;
}

int vr_group_resize(vr_group* g, uint32_t w, uint32_t h) {
    if (!g) return VR_E_ARG;
    if (h < g->n) return gfail(g, VR_E_ARG, "%u rows cannot be split between %u GPUs", h, g->n);
    // Not an image until everything below has succeeded: a failure on one GPU leaves contexts of different sizes behind,
    // and require_image then refuses frames, balancing and read-backs until a resize goes through.
    g->width = g->height = 0;
    g->progressive_result = false;  // vr_resize clears every accumulator
    int rc = on_all(g, [=](uint32_t i) { return vr_resize(g->ctx[i], w, h); });  // drains each GPU's streams first
    if (rc != VR_OK)  // a context that refused may not have drained: nothing may still store into the buffers freed next
        for (uint32_t i = 0; i < g->n; i++) {
            cudaSetDevice(g->dev[i]);
            cudaDeviceSynchronize();
        }
    free_frame_colors(g);
    if (rc != VR_OK) return rc;
    const size_t bytes = (size_t)w * h * sizeof(float);
    for (int p = 0; p < 2; p++) {
        g->fc[p].assign(g->n, nullptr);
        for (uint32_t i = 0; i < g->n && g->n > 1; i++) {
            cudaSetDevice(g->dev[i]);
            if (cudaMalloc(&g->fc[p][i], bytes) != cudaSuccess || cudaMemset(g->fc[p][i], 0, bytes) != cudaSuccess) {
                cudaGetLastError();
                free_frame_colors(g);
                return gfail(g, VR_E_CUDA, "frame-colour buffers: out of device memory");
            }
        }
    }
    // even split until vr_group_balance / vr_group_set_bands says otherwise
    std::vector<double> ones(h, 1.0);
    g->band = split_rows(ones, g->n, w);
    g->parity = 0;
    g->width = w, g->height = h;
    return VR_OK;
}

int vr_group_set_params(vr_group* g, const vr_ubo* p) {
    if (!g || !p) return VR_E_ARG;
    return on_all(g, [=](uint32_t i) { return vr_set_params(g->ctx[i], p); });
}

int vr_group_get_params(vr_group* g, vr_ubo* p) {
    if (!g || !p) return VR_E_ARG;
    return vr_get_params(g->ctx[0], p);
}

int vr_group_set_max_lights_capacity(vr_group* g, uint32_t n) {
    if (!g) return VR_E_ARG;
    return on_all(g, [=](uint32_t i) { return vr_set_max_lights_capacity(g->ctx[i], n); });
}

int vr_group_set_light_gen_groups(vr_group* g, uint32_t gx, uint32_t gy, uint32_t gz, vr_lightgen_variant variant) {
    if (!g) return VR_E_ARG;
    return on_all(g, [=](uint32_t i) { return vr_set_light_gen_groups(g->ctx[i], gx, gy, gz, variant); });
}

int vr_group_reset_accumulation(vr_group* g) {
    if (!g) return VR_E_ARG;
    g->progressive_result = false;
    return on_all(g, [=](uint32_t i) { return vr_reset_accumulation(g->ctx[i]); });
}

int vr_group_set_bands(vr_group* g, const uint32_t* y0, const uint32_t* h) {
    if (!g || !y0 || !h) return VR_E_ARG;
    int rc = require_image(g);
    if (rc != VR_OK) return rc;
    uint32_t next = 0;
    for (uint32_t i = 0; i < g->n; i++) {
        if (y0[i] != next || h[i] == 0) return gfail(g, VR_E_ARG, "bands must be non-empty, contiguous and in GPU order (band %u starts at row %u, expected %u)", i, y0[i], next);
        next += h[i];
    }
    if (next != g->height) return gfail(g, VR_E_ARG, "bands cover %u of %u rows", next, g->height);
    for (uint32_t i = 0; i < g->n; i++) g->band[i] = vr_tile{0, y0[i], g->width, h[i]};
    return VR_OK;
}

int vr_group_get_bands(vr_group* g, uint32_t* y0, uint32_t* h) {
    if (!g || !y0 || !h) return VR_E_ARG;
    for (uint32_t i = 0; i < g->n; i++) y0[i] = g->band[i].y0, h[i] = g->band[i].h;
    return VR_OK;
}

int vr_group_render_frame(vr_group* g, vr_algo algo) {
    if (!g) return VR_E_ARG;
    int rc = require_image(g);
    if (rc != VR_OK) return rc;
    if (g->progressive_result)
        return gfail(g, VR_E_STATE, "the accumulators hold a sample-parallel progressive run (result on GPU 0, partial sums elsewhere): "
                                    "call vr_group_reset_accumulation before tiled frames");
    const uint32_t parity = g->parity, epoch = ++g->epoch;
    g->parity ^= 1u;
    rc = on_all(g, [=](uint32_t i) { return frame_on(g, i, algo, parity, epoch); });
    if (rc != VR_OK || g->n == 1 || g->exchange != VR_EXCHANGE_ALLGATHER || !g->shared_device) return rc;
    return on_all(g, [=](uint32_t i) {
        cudaSetDevice(g->dev[i]);
        for (uint32_t j = 0; j < g->n; j++)
            if (j != i && cudaStreamWaitEvent(g->stream[i], g->stored[parity][j], 0) != cudaSuccess) return (int)VR_E_CUDA;
        return accumulate_rest(g, i);
    });
}

int vr_group_sync(vr_group* g) {
    if (!g) return VR_E_ARG;
    int rc = on_all(g, [=](uint32_t i) { return vr_sync(g->ctx[i]); });
    if (rc != VR_OK) return rc;
    // A barrier wait that gave up (k_peer_barrier) left the epoch it was waiting for in its GPU's mark word. Every mark is
    // read and cleared here (the streams are idle), so that a later sync reports later timeouts only; the frames since the
    // timeout averaged rows that had not arrived: the caller resets the accumulation before it goes on.
    int bad_gpu = -1;
    uint32_t bad_epoch = 0;
    for (uint32_t i = 0; i < g->n && g->exchange == VR_EXCHANGE_ALLGATHER; i++) {
        if (!g->flags[i]) continue;
        uint32_t mark = 0;
        cudaSetDevice(g->dev[i]);
        if (cudaMemcpy(&mark, g->flags[i] + VR_MAX_PEERS, sizeof(mark), cudaMemcpyDeviceToHost) != cudaSuccess)
            return gfail(g, VR_E_CUDA, "GPU %u: cannot read the barrier state", i);
        if (mark == 0) continue;
        if (cudaMemset(g->flags[i] + VR_MAX_PEERS, 0, sizeof(mark)) != cudaSuccess)
            return gfail(g, VR_E_CUDA, "GPU %u: cannot clear the barrier state", i);
        if (bad_gpu < 0) bad_gpu = (int)i, bad_epoch = mark;
    }
    if (bad_gpu >= 0)
        return gfail(g, VR_E_CUDA,
                     "GPU %d: the cross-GPU barrier of frame epoch %u timed out, a peer never arrived (the group is at epoch %u); the "
                     "accumulated image is incomplete: call vr_group_reset_accumulation before rendering on",
                     bad_gpu, bad_epoch, g->epoch);
    return VR_OK;
}

// Per-row cost from one measured full frame on GPU 0, split, then `rounds` rounds of measured feedback: every GPU renders
// its band a few times, reports the device time, the rows of slow bands get dearer. The best split seen is kept (the
// per-band time is a staircase in the number of CTA waves, so the iteration need not improve monotonically).
int vr_group_balance(vr_group* g, vr_algo algo, uint32_t rounds) {
    if (!g) return VR_E_ARG;
    int rc = require_image(g);
    if (rc != VR_OK) return rc;
    if (g->n == 1) return VR_OK;
    g->progressive_result = false;  // balancing renders into the accumulators and resets them at the end
    vr_ctx* c0 = g->ctx[0];
    vr_ubo saved;
    vr_get_params(c0, &saved);
    const uint32_t H = g->height;
    std::vector<double> cost(H, 1.0);
    {
        vr_timings t{};
        std::vector<uint64_t> rows(H);
        rc = vr_set_frame_color_peers(c0, nullptr, nullptr, 0);
        if (rc == VR_OK) rc = vr_render_frame(c0, algo, nullptr);
        if (rc == VR_OK) rc = vr_sync(c0);
        if (rc == VR_OK) rc = vr_get_timings(c0, &t);
        if (rc == VR_OK && algo != VR_ALGO_PATH) rc = vr_row_sample_counts(c0, rows.data());
        if (rc != VR_OK) return gfail(g, rc, "%s", vr_last_error(c0));
        double total = 0.0;
        for (uint64_t r : rows) total += (double)r;
        if (total > 0.0 && t.eval_ms > 0.0f) {
            const double per_row = ((double)t.march_ms / H) / ((double)t.eval_ms / total);
            for (uint32_t y = 0; y < H; y++) cost[y] = (double)rows[y] + per_row;
        }
    }
    g->band = split_rows(cost, g->n, g->width);
    std::vector<vr_tile> best = g->band;
    std::vector<double> best_per(g->n, 0.0);
    double best_ms = 1e300;
    // device time per GPU of `frames` frames of the current split (after one unmeasured frame, which re-sizes buffers)
    auto measure = [&](int frames, std::vector<double>& ms) -> int {
        int r0 = vr_group_reset_accumulation(g);
        if (r0 != VR_OK) return r0;
        ms.assign(g->n, 0.0);
        for (int f = 0; f < frames + 1 && r0 == VR_OK; f++) {
            if (f == 1) {
                r0 = on_all(g, [=](uint32_t i) {
                    int r = vr_sync(g->ctx[i]);
                    if (r == VR_OK) r = vr_get_timings_sum(g->ctx[i], nullptr, nullptr, 1);
                    return r;
                });
                if (r0 != VR_OK) break;
            }
            r0 = vr_group_render_frame(g, algo);
        }
        if (r0 != VR_OK) return r0;
        return on_all(g, [&](uint32_t i) {
            vr_timings s{};
            uint64_t nf = 0;
            int r = vr_sync(g->ctx[i]);
            if (r == VR_OK) r = vr_get_timings_sum(g->ctx[i], &s, &nf, 1);
            ms[i] = nf ? (double)s.total_ms / (double)nf : 0.0;
            return r;
        });
    };
    std::vector<double> ms;
    for (uint32_t round = 0; round <= rounds; round++) {
        rc = measure(3, ms);
        if (rc != VR_OK) return rc;
        const double worst = *std::max_element(ms.begin(), ms.end());
        if (worst < best_ms) best_ms = worst, best = g->band, best_per = ms;
        if (round == rounds) break;
        refine_costs(cost, g->band, ms);
        g->band = split_rows(cost, g->n, g->width);
    }
    // polish (same rule as tiles.polish_move): the slowest band gives one boundary row to the faster of its neighbours,
    // kept while the measured maximum goes down
    for (int move = 0; move < 8 && rounds > 0; move++) {
        const uint32_t w = (uint32_t)(std::max_element(best_per.begin(), best_per.end()) - best_per.begin());
        if (best[w].h <= 1) break;
        uint32_t t = w;
        if (w > 0) t = w - 1;
        if (w + 1 < g->n && (t == w || best_per[w + 1] < best_per[t])) t = w + 1;
        if (t == w) break;
        std::vector<vr_tile> cand = best;
        cand[w].h -= 1, cand[t].h += 1;
        if (t < w) cand[w].y0 += 1;
        else cand[t].y0 -= 1;
        g->band = cand;
        rc = measure(5, ms);
        if (rc != VR_OK) return rc;
        const double worst = *std::max_element(ms.begin(), ms.end());
        if (worst >= best_ms) break;
        best_ms = worst, best = cand, best_per = ms;
    }
    g->band = best;
    rc = vr_group_reset_accumulation(g);
    if (rc != VR_OK) return rc;
    return vr_group_set_params(g, &saved);
}

int vr_group_read_accum_f32(vr_group* g, float* out) {
    if (!g || !out) return VR_E_ARG;
    int rc = require_image(g);
    if (rc != VR_OK) return rc;
    if (g->n == 1 || g->progressive_result) return vr_read_accum_f32(g->ctx[0], out);
    // after the exchange every GPU holds the whole accumulator; each returns the rows it rendered over its own PCIe link
    rc = on_all(g, [=](uint32_t i) {
        const vr_tile& b = g->band[i];
        return vr_read_accum_f32_tile_async(g->ctx[i], &b, out + (size_t)b.y0 * g->width * 4, (uint64_t)g->width * 16);
    });
    if (rc != VR_OK) return rc;
    return on_all(g, [=](uint32_t i) { return vr_read_wait(g->ctx[i]); });
}

int vr_group_read_accum_f32_async(vr_group* g, float* out) {
    if (!g || !out) return VR_E_ARG;
    int rc = require_image(g);
    if (rc != VR_OK) return rc;
    return on_all(g, [=](uint32_t i) {
        const vr_tile b = read_rows(g, i);
        if (b.h == 0) return (int)VR_OK;
        return vr_read_accum_f32_tile_async(g->ctx[i], &b, out + (size_t)b.y0 * g->width * 4, (uint64_t)g->width * 16);
    });
}

int vr_group_read_wait(vr_group* g) {
    if (!g) return VR_E_ARG;
    return on_all(g, [=](uint32_t i) { return vr_read_wait(g->ctx[i]); });
}

int vr_group_read_image_rgba8(vr_group* g, uint8_t* out) {
    if (!g || !out) return VR_E_ARG;
    int rc = require_image(g);
    if (rc != VR_OK) return rc;
    return on_all(g, [=](uint32_t i) {
        const vr_tile b = read_rows(g, i);
        if (b.h == 0) return (int)VR_OK;
        return vr_read_image_rgba8_tile(g->ctx[i], &b, out + (size_t)b.y0 * g->width * 4, (uint64_t)g->width * 4);
    });
}

int vr_group_set_exchange(vr_group* g, int mode) {
    if (!g) return VR_E_ARG;
    if (mode != VR_EXCHANGE_NONE && mode != VR_EXCHANGE_ALLGATHER) return gfail(g, VR_E_ARG, "unknown exchange mode %d", mode);
    g->exchange = mode;
    return VR_OK;
}

// Sample-parallel progressive run: frames first+1 .. first+frames of the run go round-robin to the GPUs, each
// renders WHOLE images into a float sum, GPU 0 adds the other sums in GPU order over NVLink (peer reads) and divides.
int vr_group_render_progressive(vr_group* g, vr_algo algo, uint32_t frames) {
    if (!g) return VR_E_ARG;
    int rc = require_image(g);
    if (rc != VR_OK) return rc;
    if (frames == 0) return gfail(g, VR_E_ARG, "vr_group_render_progressive: frames must be >= 1");
    const uint32_t n = g->n;
    g->progressive_result = n > 1;  // from here on (also after a failure below) the GPUs' accumulators are out of step
    rc = on_all(g, [=](uint32_t i) {
        vr_ctx* c = g->ctx[i];
        int r = vr_set_frame_color_peers(c, nullptr, nullptr, 0);
        if (r != VR_OK) return r;
        if ((r = vr_set_frame_step(c, i + 1, n)) != VR_OK) return r;  // frames i + 1, i + 1 + n, ...
        if ((r = vr_set_accumulate_mode(c, VR_ACCUM_SUM)) != VR_OK) return r;
        if ((r = vr_clear_image(c)) != VR_OK) return r;
        for (uint32_t f = i; f < frames && r == VR_OK; f += n) r = vr_render_frame(c, algo, nullptr);
        const int r2 = vr_set_accumulate_mode(c, VR_ACCUM_AVERAGE);
        const int r3 = vr_set_frame_step(c, 0, 1);
        if (r == VR_OK) r = r2 != VR_OK ? r2 : r3;
        if (r == VR_OK) r = vr_sync(c);
        return r;
    });
    if (rc != VR_OK) return rc;
    // reduce on GPU 0: sum_0 += sum_j, j = 1 .. n-1 in this order (deterministic), reading the peers' accumulators directly
    void* acc0 = nullptr;
    vr_get_device_buffers(g->ctx[0], &acc0, nullptr, nullptr, nullptr);
    cudaSetDevice(g->dev[0]);
    const uint64_t npix = (uint64_t)g->width * g->height;
    for (uint32_t j = 1; j < n && j < frames; j++) {
        void* accj = nullptr;
        vr_get_device_buffers(g->ctx[j], &accj, nullptr, nullptr, nullptr);
        vrk_add_accum(g->stream[0], (float4*)acc0, (const float4*)accj, npix);
    }
    rc = vr_finish_sum(g->ctx[0], frames);
    if (rc == VR_OK) rc = vr_sync(g->ctx[0]);
    if (rc != VR_OK) return gfail(g, rc, "%s", vr_last_error(g->ctx[0]));
    return VR_OK;
}

// device time of `frames` tiled frames (CUDA events on every GPU's stream, the maximum over the GPUs), for benchmarks
int vr_group_time_frames(vr_group* g, vr_algo algo, uint32_t frames, float* ms_per_frame_max, float* per_gpu_ms /* n entries or NULL */) {
    if (!g || !ms_per_frame_max) return VR_E_ARG;
    int rc = require_image(g);
    if (rc != VR_OK) return rc;
    if (frames == 0) return gfail(g, VR_E_ARG, "frames must be >= 1");
    rc = on_all(g, [=](uint32_t i) { return vr_timer_start(g->ctx[i]); });  // synchronises each stream first
    for (uint32_t f = 0; f < frames && rc == VR_OK; f++) rc = vr_group_render_frame(g, algo);
    if (rc != VR_OK) return rc;
    std::vector<float> ms(g->n, 0.0f);
    rc = on_all(g, [&](uint32_t i) { return vr_timer_stop(g->ctx[i], &ms[i]); });
    if (rc != VR_OK) return rc;
    float worst = 0.0f;
    for (uint32_t i = 0; i < g->n; i++) {
        ms[i] /= (float)frames;
        if (per_gpu_ms) per_gpu_ms[i] = ms[i];
        worst = std::max(worst, ms[i]);
    }
    *ms_per_frame_max = worst;
    return VR_OK;
}

}  // extern "C"
