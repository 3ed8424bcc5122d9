// api.cu -- the C-ABI of libweaklabel_cuda.so (include/weaklabel_cuda.h): the reference's
// `e_step` entry point (estep/weaklabel.cpp:6-7) on host buffers, the device-pointer plan API,
// the explicit class-order stream, and the fused softmax-CE entry.  Host logic only; all
// arithmetic on scores happens in estep_kernel.cu / softmax_ce.cu.  No CPU fallback.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "../../include/weaklabel_cuda.h"
#include "estep_kernel.cuh"


using namespace wl;

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
#define CUDA_TRY(expr)                                                                        \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return fail(WEAKLABEL_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                   \
                        cudaGetErrorString(e__), __FILE__, __LINE__);                         \
    } while (0)

// The library works on the PLAN'S device and leaves the caller's current device as it found it (a process may
// drive several GPUs; a library call must not move it to another one behind its back).
struct DeviceScope {
    int prev = -1;
    bool moved = false;
    cudaError_t enter(int dev) {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) return e;
        if (prev != dev) {
            e = cudaSetDevice(dev);
            moved = (e == cudaSuccess);
        }
        return e;
    }
    ~DeviceScope() { if (moved) cudaSetDevice(prev); }
};
#define ON_DEVICE(dev) DeviceScope device_scope_; CUDA_TRY(device_scope_.enter(dev))

extern "C" WL_API const char *weaklabel_last_error(void) { return g_err.c_str(); }
extern "C" WL_API int weaklabel_version(void) { return 200; }
extern "C" WL_API int weaklabel_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// ---------------------------------------------------------------------------------------------
// class-order stream: glibc's rand() restated (TYPE_3 additive feedback generator of
// stdlib/random_r.c: r[i] = r[i-31] + r[i-3], output r[i] >> 1, Lehmer-seeded, 310 warm-up
// draws).  glibc is third-party to the reference, which reaches it through libstdc++'s
// std::random_shuffle at estep/weaklabel.cpp:83.
// ---------------------------------------------------------------------------------------------
struct weaklabel_rng {
    uint32_t ring[34];
    int head;
    long long draws;
};

static inline uint32_t rng_word(weaklabel_rng *g) {
    const int i = g->head;
    const uint32_t v = g->ring[(i + 3) % 34] + g->ring[(i + 31) % 34];   // i-31 and i-3 modulo 34
    g->ring[i] = v;
    g->head = (i + 1) % 34;
    return v;
}

static void rng_seed(weaklabel_rng *g, unsigned int seed) {
    if (seed == 0) seed = 1;
    int32_t x = (int32_t)seed;
    g->ring[0] = (uint32_t)x;
    for (int i = 1; i < 31; i++) {
        const long hi = x / 127773, lo = x % 127773;
        long w = 16807 * lo - 2836 * hi;       // Schrage: 16807*x mod (2^31 - 1)
        if (w < 0) w += 2147483647;
        x = (int32_t)w;
        g->ring[i] = (uint32_t)x;
    }
    for (int i = 31; i < 34; i++) g->ring[i] = g->ring[i - 31];
    g->head = 0;
    g->draws = 0;
    for (int i = 0; i < 310; i++) (void)rng_word(g);
}

extern "C" WL_API weaklabel_rng *weaklabel_rng_create(unsigned int seed) {
    weaklabel_rng *g = (weaklabel_rng *)malloc(sizeof(weaklabel_rng));
    if (g) rng_seed(g, seed);
    return g;
}
extern "C" WL_API void weaklabel_rng_destroy(weaklabel_rng *rng) { free(rng); }
extern "C" WL_API int weaklabel_rng_next(weaklabel_rng *rng) {
    rng->draws++;
    return (int)(rng_word(rng) >> 1);
}
extern "C" WL_API void weaklabel_rng_skip(weaklabel_rng *rng, long long n) {
    for (long long i = 0; i < n; i++) (void)weaklabel_rng_next(rng);
}
extern "C" WL_API long long weaklabel_rng_draws(const weaklabel_rng *rng) { return rng->draws; }
extern "C" WL_API long long weaklabel_draws_for(int n_present, int num_iter) {
    return (long long)(num_iter > 0 ? num_iter : 0) * (n_present > 2 ? n_present - 2 : 0);
}

static inline int draw(weaklabel_rng *rng) { return rng ? weaklabel_rng_next(rng) : rand(); }

// ---------------------------------------------------------------------------------------------
// label sets (weaklabel.cpp:19-31)
// ---------------------------------------------------------------------------------------------
extern "C" WL_API int weaklabel_label_set_host(const int *label, int B, int H, int W, int C,
                                        unsigned char *present) {
    if (!label || !present || B < 0 || H <= 0 || W <= 0 || C <= 0)
        return fail(WEAKLABEL_ERR_ARG, "weaklabel_label_set_host: bad argument");
    memset(present, 0, (size_t)B * C);
    for (int b = 0; b < B; b++) {
        const int *img = label + (size_t)b * H * W;
        for (int h = 0; h < H; h += 2)
            for (int w = 0; w < W; w += 2) {
                const int t = img[(size_t)h * W + w];
                if (t >= C) continue;                                  // void label
                if (t < 0) return fail(WEAKLABEL_ERR_ARG, "negative label %d in image %d", t, b);
                present[(size_t)b * C + t] = 1;
            }
    }
    return WEAKLABEL_OK;
}

// Class visiting order (weaklabel.cpp:36-43, 83) for B consecutive images on one stream:
// order[b][j][t] = class id visited at position t of iteration j, -1 padded.
// std::random_shuffle(label_single+1, label_single+P) is, in libstdc++ (bits/stl_algo.h),
// "for i = 1..n-1: swap(a[i], a[rand() % (i+1)])" on the n = P-1 tail entries; the permutation
// carries over from iteration to iteration and the first entry never moves.
extern "C" WL_API int weaklabel_draw_order_host(const unsigned char *present, int B, int C, int num_iter,
                                         weaklabel_rng *rng, int *order) {
    if (!present || !order || B < 0 || C <= 0 || C > WEAKLABEL_MAX_CLASSES || num_iter < 0)
        return fail(WEAKLABEL_ERR_ARG, "weaklabel_draw_order_host: bad argument");
    int cls[WEAKLABEL_MAX_CLASSES];
    for (int b = 0; b < B; b++) {
        int P = 0;
        for (int c = 0; c < C; c++)
            if (present[(size_t)b * C + c]) cls[P++] = c;
        if (P == 0)
            return fail(WEAKLABEL_ERR_EMPTY_SET, "image %d has an empty label set (the reference "
                        "crashes there, weaklabel.cpp:83)", b);
        for (int j = 0; j < num_iter; j++) {
            int *tail = cls + 1;
            const int n = P - 1;
            for (int i = 1; i < n; i++) {
                const int r = draw(rng) % (i + 1);
                if (r != i) { const int t = tail[i]; tail[i] = tail[r]; tail[r] = t; }
            }
            int *o = order + ((size_t)b * num_iter + j) * C;
            for (int t = 0; t < C; t++) o[t] = t < P ? cls[t] : -1;
        }
    }
    return WEAKLABEL_OK;
}

// ---------------------------------------------------------------------------------------------
// debugging aid (WEAKLABEL_TRAIL=1): progress words of every CTA of the warp-specialised kernel
// in mapped host memory, printed at process exit -- they survive a device-side exception
// ---------------------------------------------------------------------------------------------
static unsigned int *g_trail = nullptr;
static void trail_dump() {
    if (!g_trail) return;
    for (int cta = 0; cta < 160; cta++) {
        const unsigned int *w = g_trail + cta * 32;
        bool any = false;
        for (int k = 0; k < 32; k++) any = any || w[k];
        if (!any) continue;
        fprintf(stderr, "trail cta %3d:", cta);
        for (int k = 0; k < 32; k++) fprintf(stderr, " %08x", w[k]);
        fprintf(stderr, "\n");
    }
}
static volatile unsigned int *debug_trail() {
    static bool tried = false;
    if (!tried) {
        tried = true;
        if (getenv("WEAKLABEL_TRAIL")) {
            void *h = nullptr;
            if (cudaHostAlloc(&h, 160 * 32 * sizeof(unsigned int), cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
                memset(h, 0, 160 * 32 * sizeof(unsigned int));
                g_trail = (unsigned int *)h;
                atexit(trail_dump);
            } else {
                cudaGetLastError();
            }
        }
    }
    return g_trail;
}

// ---------------------------------------------------------------------------------------------
// plans
// ---------------------------------------------------------------------------------------------
struct weaklabel_plan {
    int B, H, W, C, N, device, sms;
    int G, TS, ntiles, L, Npad, slots, sub_px;      // generic kernel (estep_kernel.cu)
    int tune_G, tune_TS, tune_slots, tune_path;     // tune_path: 0 auto, 1 generic, 2 on-chip
    OnchipConfig oc; bool oc_ok; int oc_slots;      // on-chip kernel (estep_onchip.cu)
    int env_prefetch, env_flags;                    // WEAKLABEL_PREFETCH / WEAKLABEL_FLAGS, read once at creation
    int want_sums;                                  // outputs of run(): 1 = labels + sums of every image, 0 = labels only
    int last_sums;                                  // what the last run() produced
    unsigned long long *d_prof; bool prof_on;       // [B][32] phase time stamps
    bool geometry_ok;
    int scratch_slots;
    int *d_sched;                   // -> d_tables
    int last_path;
    float *d_planes;
    uint8_t *d_absidx;
    SlotCtl *d_ctl;
    int *d_counter;     // [0] image counter, [1] status
    ImageMeta *d_meta;              // -> d_tables
    uint8_t *d_order;               // -> d_tables
    // label-set tables [sched | meta | order]: ONE device block filled by ONE asynchronous copy out of a
    // pinned staging buffer; two staging buffers alternate (an event each) so that prepare() never waits
    // for the device unless it is two uploads ahead
    unsigned char *d_tables; size_t tables_cap;
    unsigned char *h_stage[2]; size_t h_stage_cap[2]; cudaEvent_t ev_stage[2]; bool stage_pending[2]; int stage_next;
    cudaEvent_t ev_tables, ev_run;  // tables uploaded / last kernel of the last run() queued
    cudaStream_t tables_stream; bool has_tables, has_run;
    uint32_t *d_masks; int *d_flags; uint32_t *h_masks;   // device label sets (label_set.cu); h_masks pinned [B + 1]
    std::vector<unsigned char> present_tmp;
    float *d_bias;
    size_t bias_cap;
    float *d_mean0;
    double *d_mean1, *d_shift;
    std::vector<ImageMeta> h_meta;
    std::vector<uint8_t> h_order;    // plane indices [B][num_iter][C]
    bool prepared;
    int num_iter, suppress, nth_bg, nth_fg;
    float margin;
    long long launches;
    cudaStream_t last_stream;
};

static void plan_free_geometry(weaklabel_plan *pl) {
    cudaFree(pl->d_planes); pl->d_planes = nullptr;
    cudaFree(pl->d_absidx); pl->d_absidx = nullptr;
    cudaFree(pl->d_ctl); pl->d_ctl = nullptr;
    pl->geometry_ok = false;
}

static int plan_geometry(weaklabel_plan *pl) {
    if (pl->geometry_ok) return WEAKLABEL_OK;
    const int N = pl->N;
    int G = pl->tune_G;
    if (G > 0 && G != 16) {                             // the generic kernel takes power-of-two clusters: round down
        int g2 = 1;                                     // (sizes 3, 5, 6, 7, 9 exist for the on-chip kernel only)
        while (g2 * 2 <= G && g2 < 8) g2 *= 2;
        G = g2;
    }
    if (G == 0) {
        G = 1;
        while (G < 8 && (N + G - 1) / G > TS_MAX) G *= 2;
        if ((N + G - 1) / G > TS_MAX && (N + 15) / 16 <= TS_MAX) G = 16;
        while (G < 8 && (long long)pl->B * G * 2 <= pl->sms && N / (2 * G) >= 4096) G *= 2;
    }
    if (G != 1 && G != 2 && G != 4 && G != 8 && G != 16)
        return fail(WEAKLABEL_ERR_ARG, "cluster size %d not in {1,2,4,8,16}", G);
    int TS = pl->tune_TS ? pl->tune_TS : ((N + G - 1) / G + 3) / 4 * 4;
    TS = (TS + 3) / 4 * 4;
    if (TS < 4) TS = 4;
    if (TS > TS_MAX) TS = TS_MAX;
    pl->G = G;
    pl->TS = TS;
    pl->ntiles = (N + TS - 1) / TS;
    pl->L = (TS + NT - 1) / NT;
    if (pl->L % 2 == 0) pl->L++;
    pl->Npad = (N + 3) / 4 * 4;
    int sub = (STREAM_BUF_FLOATS - 8) / pl->C;
    if (sub > NT) sub = NT;
    sub = sub / 32 * 32;
    pl->sub_px = sub;
    int maxc = estep_max_clusters(G);
    if (maxc <= 0)
        return fail(WEAKLABEL_ERR_CUDA, "cluster size %d cannot be scheduled on device %d (%s)", G,
                    pl->device, cudaGetErrorString(cudaPeekAtLastError()));
    int slots = pl->tune_slots ? pl->tune_slots : maxc;
    if (slots > maxc) slots = maxc;
    if (slots > pl->B) slots = pl->B;
    if (slots < 1) slots = 1;
    pl->slots = slots;
    pl->oc_ok = onchip_configure(N, pl->C, pl->B, pl->sms, pl->tune_G, pl->tune_TS, &pl->oc);
    if (pl->oc_ok) {
        int vs = pl->tune_slots ? pl->tune_slots : pl->oc.max_clusters;
        if (vs > pl->oc.max_clusters) vs = pl->oc.max_clusters;
        if (vs > pl->B) vs = pl->B;
        if (vs < 1) vs = 1;
        pl->oc_slots = vs;
    }
    if (pl->tune_path == 2 && !pl->oc_ok)
        return fail(WEAKLABEL_ERR_ARG, "the on-chip kernel cannot hold %d pixels with this cluster/tile tuning", N);
    int ss = slots;
    if (pl->oc_ok && 2 * pl->oc_slots > ss) ss = 2 * pl->oc_slots;      // two plane sets per resident cluster
    pl->scratch_slots = ss;
    CUDA_TRY(cudaMalloc(&pl->d_planes, (size_t)ss * (pl->C + 2) * pl->Npad * sizeof(float)));
    CUDA_TRY(cudaMalloc(&pl->d_absidx, (size_t)ss * pl->Npad));
    CUDA_TRY(cudaMalloc(&pl->d_ctl, (size_t)ss * sizeof(SlotCtl)));
    pl->geometry_ok = true;
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_plan_create(weaklabel_plan **out, int B, int H, int W, int C, int device) {
    if (!out) return fail(WEAKLABEL_ERR_ARG, "weaklabel_plan_create: out is NULL");
    *out = nullptr;
    if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || C > WEAKLABEL_MAX_CLASSES)
        return fail(WEAKLABEL_ERR_ARG, "bad shape B=%d H=%d W=%d C=%d (C <= %d)", B, H, W, C,
                    WEAKLABEL_MAX_CLASSES);
    if ((long long)H * W > 0x7fffffffLL / 64)
        return fail(WEAKLABEL_ERR_ARG, "image of %lld pixels is too large", (long long)H * W);
    if (weaklabel_device_count() <= 0)
        return fail(WEAKLABEL_ERR_CUDA, "no CUDA device: libweaklabel_cuda has no CPU fallback");
    ON_DEVICE(device);
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 9)
        return fail(WEAKLABEL_ERR_CUDA, "device %d (%s, sm_%d%d) has no thread-block clusters; "
                    "this library is built for sm_100a", device, prop.name, prop.major, prop.minor);
    weaklabel_plan *pl = new weaklabel_plan();
    pl->prepared = false;
    {
        const char *pf = getenv("WEAKLABEL_PREFETCH");      // tuning knob: chunks of L2 prefetch distance
        pl->env_prefetch = pf ? atoi(pf) : OC_PF_AHEAD;
        const char *fl = getenv("WEAKLABEL_FLAGS");         // A/B switches (estep_kernel.cuh: OC_F_*)
        pl->env_flags = fl ? atoi(fl) : 0;
    }
    pl->launches = 0;
    pl->want_sums = 1;
    pl->last_sums = 1;
    pl->last_stream = nullptr;
    pl->B = B; pl->H = H; pl->W = W; pl->C = C; pl->N = H * W;
    pl->device = device;
    pl->sms = prop.multiProcessorCount;
    cudaError_t e = cudaMalloc(&pl->d_counter, 2 * sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(pl->d_counter, 0, 2 * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_masks, (size_t)B * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_flags, sizeof(int));
    if (e == cudaSuccess) e = cudaHostAlloc((void **)&pl->h_masks, (size_t)(B + 1) * sizeof(uint32_t), cudaHostAllocDefault);
    for (int k = 0; k < 2 && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&pl->ev_stage[k], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&pl->ev_tables, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&pl->ev_run, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_mean0, (size_t)B * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_mean1, (size_t)B * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_shift, (size_t)B * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_prof, (size_t)B * 32 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(pl->d_prof, 0, (size_t)B * 32 * sizeof(unsigned long long));
    if (e != cudaSuccess) {
        weaklabel_plan_destroy(pl);
        return fail(WEAKLABEL_ERR_CUDA, "cudaMalloc failed: %s", cudaGetErrorString(e));
    }
    *out = pl;
    return WEAKLABEL_OK;
}

extern "C" WL_API void weaklabel_plan_destroy(weaklabel_plan *pl) {
    if (!pl) return;
    DeviceScope scope;
    scope.enter(pl->device);
    plan_free_geometry(pl);
    if (pl->has_run) cudaStreamSynchronize(pl->last_stream);
    if (pl->has_tables) cudaStreamSynchronize(pl->tables_stream);
    cudaFree(pl->d_counter); cudaFree(pl->d_tables); cudaFree(pl->d_bias); cudaFree(pl->d_masks); cudaFree(pl->d_flags);
    cudaFree(pl->d_mean0); cudaFree(pl->d_mean1); cudaFree(pl->d_shift); cudaFree(pl->d_prof);
    if (pl->h_masks) cudaFreeHost(pl->h_masks);
    for (int k = 0; k < 2; k++) {
        if (pl->h_stage[k]) cudaFreeHost(pl->h_stage[k]);
        if (pl->ev_stage[k]) cudaEventDestroy(pl->ev_stage[k]);
    }
    if (pl->ev_tables) cudaEventDestroy(pl->ev_tables);
    if (pl->ev_run) cudaEventDestroy(pl->ev_run);
    delete pl;
}

extern "C" WL_API int weaklabel_plan_tune(weaklabel_plan *pl, int cluster_size, int tile_pixels, int max_clusters) {
    if (!pl) return fail(WEAKLABEL_ERR_ARG, "plan is NULL");
    if (cluster_size < 0 || tile_pixels < 0 || max_clusters < 0 || tile_pixels > TS_MAX)
        return fail(WEAKLABEL_ERR_ARG, "bad tuning values");
    ON_DEVICE(pl->device);
    if (pl->has_run) CUDA_TRY(cudaStreamSynchronize(pl->last_stream));
    plan_free_geometry(pl);
    pl->tune_G = cluster_size; pl->tune_TS = tile_pixels; pl->tune_slots = max_clusters;
    return WEAKLABEL_OK;
}

extern "C" WL_API long long weaklabel_plan_launches(const weaklabel_plan *pl) { return pl ? pl->launches : 0; }

extern "C" WL_API int weaklabel_plan_set_path(weaklabel_plan *pl, int path) {
    if (!pl || path < 0 || path > 2) return fail(WEAKLABEL_ERR_ARG, "path must be 0 (auto), 1 (generic) or 2 (on-chip)");
    pl->tune_path = path;
    return WEAKLABEL_OK;
}
extern "C" WL_API int weaklabel_plan_last_path(const weaklabel_plan *pl) { return pl ? pl->last_path : 0; }

extern "C" WL_API int weaklabel_plan_profile(weaklabel_plan *pl, int enable, unsigned long long *out) {
    if (!pl) return fail(WEAKLABEL_ERR_ARG, "plan is NULL");
    ON_DEVICE(pl->device);
    if (pl->has_run) CUDA_TRY(cudaStreamSynchronize(pl->last_stream));
    if (out) CUDA_TRY(cudaMemcpy(out, pl->d_prof, (size_t)pl->B * 32 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    pl->prof_on = enable != 0;
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_plan_prepare_async(weaklabel_plan *pl, const unsigned char *present,
                                            bool suppress_others, int num_iter, float margin_others,
                                            float bg, float fg, weaklabel_rng *rng, void *stream) {
    if (!pl || !present) return fail(WEAKLABEL_ERR_ARG, "weaklabel_plan_prepare: NULL argument");
    if (num_iter < 0) return fail(WEAKLABEL_ERR_ARG, "num_iter %d < 0", num_iter);
    const int B = pl->B, C = pl->C, N = pl->N;
    cudaStream_t st = (cudaStream_t)stream;
    // weaklabel.cpp:88 -- float * int product truncated to int
    const int nth_bg = (int)(bg * N), nth_fg = (int)(fg * N);
    // validate every label set before drawing anything from the order stream
    for (int b = 0; b < B; b++) {
        int P = 0;
        bool has0 = present[(size_t)b * C] != 0;
        for (int c = 0; c < C; c++) P += present[(size_t)b * C + c] ? 1 : 0;
        if (P == 0)
            return fail(WEAKLABEL_ERR_EMPTY_SET, "image %d has an empty label set (the reference "
                        "crashes there, weaklabel.cpp:83)", b);
        if ((has0 && (nth_bg < 0 || nth_bg >= N)) || ((P > 1 || !has0) && (nth_fg < 0 || nth_fg >= N)))
            return fail(WEAKLABEL_ERR_ARG, "rank outside [0,N): bg=%g fg=%g N=%d (weaklabel.cpp:88-89 "
                        "needs nth < N)", (double)bg, (double)fg, N);
    }
    pl->h_meta.assign(B, ImageMeta());
    pl->h_order.assign((size_t)B * num_iter * C, 0);
    std::vector<int> ord_cls((size_t)B * num_iter * C + 1, -1);
    int rc0 = weaklabel_draw_order_host(present, B, C, num_iter, rng, ord_cls.data());
    if (rc0) return rc0;
    int pos_of[WEAKLABEL_MAX_CLASSES];
    for (int b = 0; b < B; b++) {
        ImageMeta &m = pl->h_meta[b];
        memset(&m, 0, sizeof m);
        m.first_absent = -1;
        int P = 0;
        for (int c = 0; c < C; c++) {
            if (present[(size_t)b * C + c]) {
                m.mask |= 1u << c;
                m.plist[P] = (uint8_t)c;
                pos_of[c] = P++;
            } else if (m.first_absent < 0) {
                m.first_absent = c;
            }
        }
        m.P = P;
        for (int j = 0; j < num_iter; j++)
            for (int t = 0; t < P; t++) {
                const size_t o = ((size_t)b * num_iter + j) * C + t;
                pl->h_order[o] = (uint8_t)pos_of[ord_cls[o]];
            }
    }
    ON_DEVICE(pl->device);
    // one block [sched | meta | order], staged in pinned memory and uploaded with one asynchronous copy
    const size_t osz = pl->h_order.size() > 0 ? pl->h_order.size() : 1;
    const size_t off_meta = ((size_t)B * sizeof(int) + 15) & ~(size_t)15;
    const size_t off_order = off_meta + (size_t)B * sizeof(ImageMeta);
    const size_t total = (off_order + osz + 15) & ~(size_t)15;
    const int k = pl->stage_next;
    pl->stage_next ^= 1;
    if (pl->stage_pending[k]) {                           // the upload this buffer fed two prepares ago
        CUDA_TRY(cudaEventSynchronize(pl->ev_stage[k]));
        pl->stage_pending[k] = false;
    }
    if (total > pl->h_stage_cap[k]) {
        if (pl->h_stage[k]) cudaFreeHost(pl->h_stage[k]);
        pl->h_stage[k] = nullptr; pl->h_stage_cap[k] = 0;
        CUDA_TRY(cudaHostAlloc((void **)&pl->h_stage[k], total + total / 2, cudaHostAllocDefault));
        pl->h_stage_cap[k] = total + total / 2;
    }
    unsigned char *h = pl->h_stage[k];
    {   // claim order: largest label sets first (longest jobs first keeps the tail short)
        int *sched = reinterpret_cast<int *>(h);
        for (int b = 0; b < B; b++) sched[b] = b;
        std::stable_sort(sched, sched + B, [&](int a, int b2) { return pl->h_meta[a].P > pl->h_meta[b2].P; });
    }
    memcpy(h + off_meta, pl->h_meta.data(), (size_t)B * sizeof(ImageMeta));
    if (!pl->h_order.empty()) memcpy(h + off_order, pl->h_order.data(), pl->h_order.size());
    if (total > pl->tables_cap || osz > pl->bias_cap) {   // growth (first prepare, larger num_iter): the old tables may be in use
        if (pl->has_run) CUDA_TRY(cudaStreamSynchronize(pl->last_stream));
        if (pl->has_tables) CUDA_TRY(cudaStreamSynchronize(pl->tables_stream));
        if (total > pl->tables_cap) {
            cudaFree(pl->d_tables); pl->d_tables = nullptr; pl->tables_cap = 0;
            CUDA_TRY(cudaMalloc(&pl->d_tables, total + total / 2));
            pl->tables_cap = total + total / 2;
        }
        if (osz > pl->bias_cap) {
            cudaFree(pl->d_bias); pl->d_bias = nullptr; pl->bias_cap = 0;
            CUDA_TRY(cudaMalloc(&pl->d_bias, (osz + osz / 2) * sizeof(float)));
            pl->bias_cap = osz + osz / 2;
        }
    }
    // a kernel of the previous run() on ANOTHER stream may still read the tables: order the upload behind it
    if (pl->has_run && pl->last_stream != st) CUDA_TRY(cudaStreamWaitEvent(st, pl->ev_run, 0));
    CUDA_TRY(cudaMemcpyAsync(pl->d_tables, h, total, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaEventRecord(pl->ev_stage[k], st));
    pl->stage_pending[k] = true;
    CUDA_TRY(cudaMemsetAsync(pl->d_bias, 0, osz * sizeof(float), st));
    CUDA_TRY(cudaEventRecord(pl->ev_tables, st));
    pl->tables_stream = st;
    pl->has_tables = true;
    pl->d_sched = reinterpret_cast<int *>(pl->d_tables);
    pl->d_meta = reinterpret_cast<ImageMeta *>(pl->d_tables + off_meta);
    pl->d_order = pl->d_tables + off_order;
    pl->num_iter = num_iter;
    pl->suppress = suppress_others ? 1 : 0;
    pl->margin = margin_others;
    pl->nth_bg = nth_bg;
    pl->nth_fg = nth_fg;
    pl->prepared = true;
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_plan_prepare(weaklabel_plan *pl, const unsigned char *present,
                                      bool suppress_others, int num_iter, float margin_others,
                                      float bg, float fg, weaklabel_rng *rng) {
    // the tables go up on the stream of the last run() (or the legacy default stream); run() on any other
    // stream waits for the upload through an event -- nothing here blocks the host
    void *st = (pl && pl->has_run) ? (void *)pl->last_stream : nullptr;
    return weaklabel_plan_prepare_async(pl, present, suppress_others, num_iter, margin_others, bg, fg, rng, st);
}

// ---------------------------------------------------------------------------------------------
// GT shrink + label sets on the device (deeplab.py:110 + weaklabel.cpp:19-31), label_set.cu
// ---------------------------------------------------------------------------------------------
extern "C" WL_API int weaklabel_label_set_device(const void *d_gt, int gt_is_int32, int B, int Hg, int Wg, int H, int W, int C,
                                          unsigned int *d_masks, int *d_flags, void *stream) {
    if (!d_gt || !d_masks || !d_flags || B <= 0 || Hg <= 0 || Wg <= 0 || H <= 0 || W <= 0 || C <= 0 || C > WEAKLABEL_MAX_CLASSES)
        return fail(WEAKLABEL_ERR_ARG, "weaklabel_label_set_device: bad argument");
    CUDA_TRY(launch_label_set(d_gt, gt_is_int32, B, Hg, Wg, H, W, C, d_masks, d_flags, (cudaStream_t)stream));
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_plan_prepare_device_gt(weaklabel_plan *pl, const void *d_gt, int gt_is_int32, int Hg, int Wg,
                                                bool suppress_others, int num_iter, float margin_others,
                                                float bg, float fg, weaklabel_rng *rng, void *stream) {
    if (!pl || !d_gt) return fail(WEAKLABEL_ERR_ARG, "weaklabel_plan_prepare_device_gt: NULL argument");
    if (Hg <= 0 || Wg <= 0) return fail(WEAKLABEL_ERR_ARG, "bad ground-truth size %dx%d", Hg, Wg);
    ON_DEVICE(pl->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int B = pl->B, C = pl->C;
    CUDA_TRY(launch_label_set(d_gt, gt_is_int32, B, Hg, Wg, pl->H, pl->W, C, pl->d_masks, pl->d_flags, st));
    pl->launches += 1;
    CUDA_TRY(cudaMemcpyAsync(pl->h_masks, pl->d_masks, (size_t)B * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(pl->h_masks + B, pl->d_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
    // the class order is drawn on the host from rand() and depends on every label-set size: B + 1 words
    // come back (the only synchronisation of the prepare path)
    CUDA_TRY(cudaStreamSynchronize(st));
    if (pl->h_masks[B] & 1u) return fail(WEAKLABEL_ERR_ARG, "negative label in the ground-truth map");
    pl->present_tmp.assign((size_t)B * C, 0);
    for (int b = 0; b < B; b++)
        for (int c = 0; c < C; c++) pl->present_tmp[(size_t)b * C + c] = (pl->h_masks[b] >> c) & 1u;
    return weaklabel_plan_prepare_async(pl, pl->present_tmp.data(), suppress_others, num_iter, margin_others, bg, fg, rng, stream);
}

// up_src != NULL: fused bilinear up-sampling -- d_map is ignored, the scores are interpolated from the
// low-resolution tensor up_src [B][up_h][up_w][C] inside the on-chip kernel (labels only)
static int run_estep(weaklabel_plan *pl, const float *d_map, long long batch_stride, long long pixel_stride,
                     unsigned char *d_labels, void *stream, int want_sums, const float *up_src = nullptr,
                     int up_h = 0, int up_w = 0) {
    if (!pl || (!d_map && !up_src)) return fail(WEAKLABEL_ERR_ARG, "weaklabel_estep_labels_device: NULL argument");
    if (!pl->prepared) return fail(WEAKLABEL_ERR_STATE, "plan not prepared");
    if (up_src) {
        d_map = up_src;                              // never dereferenced as a map; keeps the checks below uniform
        batch_stride = (long long)pl->N * pl->C;
        pixel_stride = pl->C;
    }
    if (pixel_stride < pl->C || (pl->B > 1 && batch_stride < (long long)(pl->N - 1) * pixel_stride + pl->C))
        return fail(WEAKLABEL_ERR_ARG, "strides overlap: batch %lld pixel %lld", batch_stride, pixel_stride);
    if (((uintptr_t)d_map & 3u) != 0) return fail(WEAKLABEL_ERR_ARG, "d_map is not 4-byte aligned");
    ON_DEVICE(pl->device);
    int rc = plan_geometry(pl);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    EStepParams p;
    memset(&p, 0, sizeof p);
    p.map = d_map; p.s0 = batch_stride; p.s2 = pixel_stride;
    p.B = pl->B; p.N = pl->N; p.C = pl->C;
    p.num_iter = pl->num_iter; p.suppress = pl->suppress; p.margin = pl->margin;
    p.nth_bg = pl->nth_bg; p.nth_fg = pl->nth_fg;
    p.meta = pl->d_meta; p.order = pl->d_order;
    p.planes = pl->d_planes; p.absent_idx = pl->d_absidx; p.ctl = pl->d_ctl;
    // the on-chip fast path covers suppress_others with a margin >= 0 on contiguous pixels
    // (weaklabel.cpp:45-65 then guarantees that no absent class exceeds a present one)
    const bool oc_ok = pl->oc_ok && pl->suppress && pl->margin >= 0.f && pixel_stride == pl->C;
    if (pl->tune_path == 2 && !oc_ok)
        return fail(WEAKLABEL_ERR_ARG, "on-chip path forced but not applicable (needs suppress_others, margin >= 0, "
                    "contiguous pixels)");
    const int path = pl->tune_path ? pl->tune_path : (oc_ok ? 2 : 1);
    if (up_src && path != 2)
        return fail(WEAKLABEL_ERR_STATE, "fused up-sampling runs in the on-chip kernel only (needs suppress_others, margin >= 0 "
                    "and a shape that kernel covers)");
    if (up_src) {
        p.up_src = up_src; p.up_h = up_h; p.up_w = up_w; p.up_W = pl->W;
        p.up_sy = (float)up_h / (float)pl->H;        // CalculateResizeScale: in_size / static_cast<float>(out_size)
        p.up_sx = (float)up_w / (float)pl->W;
    }
    int slots_used = pl->slots;
    p.Npad = pl->Npad; p.sub_px = pl->sub_px;
    if (path == 2) {
        p.TS = pl->oc.TS; p.ntiles = pl->oc.G; p.L = pl->oc.L; p.G = pl->oc.G; p.sub_px = pl->oc.cpx;
        p.se_floats = pl->oc.se_floats;
        slots_used = pl->oc_slots;
        p.pf_ahead = pl->env_prefetch;
        p.flags = pl->env_flags;
        p.want_sums = (want_sums || !d_labels) ? 1 : 0;
        p.trail = debug_trail();
    } else {
        p.TS = pl->TS; p.ntiles = pl->ntiles; p.L = pl->L; p.G = pl->G;
    }
    p.sched = pl->d_sched;
    p.image_counter = pl->d_counter; p.status = pl->d_counter + 1;
    p.labels = d_labels; p.bias_out = pl->d_bias;
    p.mean0 = pl->d_mean0; p.mean1 = pl->d_mean1; p.shift = pl->d_shift;
    p.prof = pl->prof_on ? pl->d_prof : nullptr;
    // the label-set tables were uploaded on another stream: wait for them on this one (device side)
    if (pl->has_tables && pl->tables_stream != st) CUDA_TRY(cudaStreamWaitEvent(st, pl->ev_tables, 0));
    CUDA_TRY(cudaMemsetAsync(pl->d_counter, 0, sizeof(int), st));      // image counter; the status word is sticky
    if (path == 2) {
        CUDA_TRY(launch_estep_onchip(p, pl->oc, slots_used, st));
    } else {
        CUDA_TRY(cudaMemsetAsync(pl->d_ctl, 0, (size_t)slots_used * sizeof(SlotCtl), st));
        CUDA_TRY(launch_estep(p, slots_used, st));
    }
    CUDA_TRY(cudaEventRecord(pl->ev_run, st));
    pl->last_path = path;
    pl->last_sums = (path == 2) ? p.want_sums : 1;      // the generic kernel always forms the sums
    pl->launches += 1;
    pl->last_stream = st;
    pl->has_run = true;
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_estep_labels_device(weaklabel_plan *pl, const float *d_map,
                                             long long batch_stride, long long pixel_stride,
                                             unsigned char *d_labels, void *stream) {
    return run_estep(pl, d_map, batch_stride, pixel_stride, d_labels, stream, pl ? pl->want_sums : 1);
}

// ---------------------------------------------------------------------------------------------
// bilinear up-sampling of the fc8 scores (network.py:40): materialised, and fused into the E-step
// ---------------------------------------------------------------------------------------------
extern "C" WL_API int weaklabel_upsample_bilinear_device(const float *d_src, int B, int h, int w, int C, int H, int W,
                                                  float *d_dst, void *stream) {
    if (!d_src || !d_dst) return fail(WEAKLABEL_ERR_ARG, "weaklabel_upsample_bilinear_device: NULL argument");
    if (B <= 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0 || C <= 0 || C > WEAKLABEL_MAX_CLASSES)
        return fail(WEAKLABEL_ERR_ARG, "bad shape");
    if ((long long)H * W * C >= (1ll << 31) || (long long)h * w * C >= (1ll << 31))
        return fail(WEAKLABEL_ERR_ARG, "image too large: H*W*C and h*w*C must stay below 2^31");
    if (weaklabel_device_count() <= 0)
        return fail(WEAKLABEL_ERR_CUDA, "no CUDA device: libweaklabel_cuda has no CPU fallback");
    CUDA_TRY(launch_upsample_bilinear(d_src, B, h, w, C, H, W, d_dst, (cudaStream_t)stream));
    return WEAKLABEL_OK;
}

// Labels of the E-step run on the UP-SAMPLED scores, which are never materialised: the plan is prepared for the
// full-resolution shape (B, H, W, C); d_fc8 is the contiguous low-resolution tensor [B][h][w][C].  Labels only:
// there is no full-resolution map to rewrite.  When the on-chip kernel does not cover the configuration the call
// fails (materialise with weaklabel_upsample_bilinear_device and use weaklabel_estep_labels_device instead).
extern "C" WL_API int weaklabel_estep_labels_upsampled_device(weaklabel_plan *pl, const float *d_fc8, int h, int w,
                                                       unsigned char *d_labels, void *stream) {
    if (!pl || !d_fc8 || !d_labels) return fail(WEAKLABEL_ERR_ARG, "weaklabel_estep_labels_upsampled_device: NULL argument");
    if (h <= 0 || w <= 0 || (long long)h * w * pl->C >= (1ll << 31)) return fail(WEAKLABEL_ERR_ARG, "bad low-resolution shape %d x %d", h, w);
    if (((uintptr_t)d_fc8 & 3u) != 0) return fail(WEAKLABEL_ERR_ARG, "d_fc8 is not 4-byte aligned");
    return run_estep(pl, nullptr, 0, 0, d_labels, stream, 0, d_fc8, h, w);
}

extern "C" WL_API int weaklabel_plan_set_outputs(weaklabel_plan *pl, int with_sums) {
    if (!pl) return fail(WEAKLABEL_ERR_ARG, "plan is NULL");
    pl->want_sums = with_sums ? 1 : 0;
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_estep_finalize_map_device(weaklabel_plan *pl, float *d_map,
                                                   long long batch_stride, long long pixel_stride,
                                                   void *stream) {
    if (!pl || !d_map) return fail(WEAKLABEL_ERR_ARG, "weaklabel_estep_finalize_map_device: NULL argument");
    if (!pl->prepared) return fail(WEAKLABEL_ERR_STATE, "plan not prepared");
    if (!pl->last_sums)
        return fail(WEAKLABEL_ERR_STATE, "the last run() was labels-only (weaklabel_plan_set_outputs(plan, 0)): the in-place "
                    "map needs the re-centring sums of every image");
    ON_DEVICE(pl->device);
    FinalizeParams f;
    memset(&f, 0, sizeof f);
    f.map = d_map; f.s0 = batch_stride; f.s2 = pixel_stride;
    f.B = pl->B; f.N = pl->N; f.C = pl->C; f.num_iter = pl->num_iter; f.suppress = pl->suppress;
    f.margin = pl->margin; f.meta = pl->d_meta; f.order = pl->d_order; f.bias = pl->d_bias;
    f.shift = pl->d_shift;
    if (pl->has_run && pl->last_stream != (cudaStream_t)stream)        // biases / shift come from run()'s stream
        CUDA_TRY(cudaStreamWaitEvent((cudaStream_t)stream, pl->ev_run, 0));
    if (finalize_tiles_applicable(f) && !getenv("WEAKLABEL_FINALIZE_SIMPLE")) CUDA_TRY(launch_finalize_tiles(f, pl->sms, (cudaStream_t)stream));
    else CUDA_TRY(launch_finalize_map(f, (cudaStream_t)stream));
    CUDA_TRY(cudaEventRecord(pl->ev_run, (cudaStream_t)stream));
    pl->launches += 1;
    pl->last_stream = (cudaStream_t)stream;
    pl->has_run = true;
    return WEAKLABEL_OK;
}

static int plan_check_status(weaklabel_plan *pl, int *code) {      // read and clear
    int st = 0;
    CUDA_TRY(cudaMemcpy(&st, pl->d_counter + 1, sizeof st, cudaMemcpyDeviceToHost));
    if (code) *code = st;
    if (st != 0) CUDA_TRY(cudaMemset(pl->d_counter + 1, 0, sizeof(int)));
    if (st != 0)
        return fail(WEAKLABEL_ERR_CUDA, "E-step kernel reported an inconsistent selection (code %d): "
                    "NaN or Inf in the score map?", st);
    return WEAKLABEL_OK;
}

// The kernels' status word, accumulated over every run() since it was last read: synchronises the plan's last stream
// and reads ONE int.  0 = fine; 2xx / 3xx = a selection found no consistent rank (NaN / Inf among the
// scores: the labels of that image are not to be trusted); 4xx = a bounded wait gave up.
extern "C" WL_API int weaklabel_plan_status(weaklabel_plan *pl, int *code) {
    if (!pl) return fail(WEAKLABEL_ERR_ARG, "plan is NULL");
    if (code) *code = 0;
    if (!pl->has_run) return WEAKLABEL_OK;
    ON_DEVICE(pl->device);
    CUDA_TRY(cudaStreamSynchronize(pl->last_stream));
    return plan_check_status(pl, code);
}

extern "C" WL_API int weaklabel_plan_probes(weaklabel_plan *pl, int *order, float *bias, float *mean0, double *mean1) {
    if (!pl || !pl->prepared) return fail(WEAKLABEL_ERR_STATE, "plan not prepared");
    ON_DEVICE(pl->device);
    if (pl->has_run) CUDA_TRY(cudaStreamSynchronize(pl->last_stream));
    int rc = plan_check_status(pl, nullptr);
    if (rc) return rc;
    const int B = pl->B, C = pl->C, J = pl->num_iter;
    if (order) {
        for (int b = 0; b < B; b++)
            for (int j = 0; j < J; j++)
                for (int t = 0; t < C; t++) {
                    const size_t o = ((size_t)b * J + j) * C + t;
                    order[o] = t < pl->h_meta[b].P ? (int)pl->h_meta[b].plist[pl->h_order[o]] : -1;
                }
    }
    if (bias && (size_t)B * J * C > 0)
        CUDA_TRY(cudaMemcpy(bias, pl->d_bias, (size_t)B * J * C * sizeof(float), cudaMemcpyDeviceToHost));
    if (mean0) CUDA_TRY(cudaMemcpy(mean0, pl->d_mean0, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost));
    if (mean1) CUDA_TRY(cudaMemcpy(mean1, pl->d_mean1, (size_t)B * sizeof(double), cudaMemcpyDeviceToHost));
    return WEAKLABEL_OK;
}

// ---------------------------------------------------------------------------------------------
// cached plans for the one-call entry points: a small most-recently-used list keyed by shape and
// device (train / val shapes alternate without re-creating scratch)
// ---------------------------------------------------------------------------------------------
static std::mutex g_mu;
static std::vector<weaklabel_plan *> g_plans;          // front = most recently used
static const size_t PLAN_CACHE = 6;
static std::vector<unsigned char> g_present;

static int cached_plan(int B, int H, int W, int C, weaklabel_plan **out) {
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    for (size_t i = 0; i < g_plans.size(); i++) {
        weaklabel_plan *pl = g_plans[i];
        if (pl->B == B && pl->H == H && pl->W == W && pl->C == C && pl->device == dev) {
            g_plans.erase(g_plans.begin() + i);
            g_plans.insert(g_plans.begin(), pl);
            *out = pl;
            return WEAKLABEL_OK;
        }
    }
    weaklabel_plan *pl = nullptr;
    int rc = weaklabel_plan_create(&pl, B, H, W, C, dev);
    if (rc) return rc;
    g_plans.insert(g_plans.begin(), pl);
    while (g_plans.size() > PLAN_CACHE) {
        weaklabel_plan_destroy(g_plans.back());
        g_plans.pop_back();
    }
    *out = pl;
    return WEAKLABEL_OK;
}

static int check_shape_stride(const int *shape, const int *stride, long long *s0, long long *s2) {
    if (!shape || !stride) return fail(WEAKLABEL_ERR_ARG, "shape/stride is NULL");
    // weaklabel.cpp:13-16: byte strides divided by sizeof(float); only [0] and [2] are used (:50)
    *s0 = stride[0] / (int)sizeof(float);
    *s2 = stride[2] / (int)sizeof(float);
    if (shape[0] <= 0 || shape[1] <= 0 || shape[2] <= 0 || shape[3] <= 0)
        return fail(WEAKLABEL_ERR_ARG, "bad shape {%d,%d,%d,%d}", shape[0], shape[1], shape[2], shape[3]);
    if (*s2 < shape[3])
        return fail(WEAKLABEL_ERR_ARG, "pixel stride %d bytes < C*4 = %d", stride[2], shape[3] * 4);
    return WEAKLABEL_OK;
}

// weaklabel.cpp:19-31 over a large batch: the stride-2 scan of a 256 x 513^2 int32 map is 270 MB of host
// memory traffic in front of the pipeline -- a few threads take blocks of images each
static int label_set_host_parallel(const int *label, int B, int H, int W, int C, unsigned char *present) {
    const size_t per = (size_t)H * W;
    int nt = (int)std::min<size_t>(8, (size_t)B * per / ((size_t)4 << 20));
    if (nt <= 1) return weaklabel_label_set_host(label, B, H, W, C, present);
    std::vector<int> rcs(nt, 0);
    std::vector<std::string> errs(nt);
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++) {
        const int b0 = (int)((long long)B * t / nt), b1 = (int)((long long)B * (t + 1) / nt);
        th.emplace_back([=, &rcs, &errs]() {
            rcs[t] = b1 > b0 ? weaklabel_label_set_host(label + (size_t)b0 * per, b1 - b0, H, W, C, present + (size_t)b0 * C) : 0;
            if (rcs[t]) errs[t] = weaklabel_last_error();          // the error string is thread-local
        });
    }
    for (auto &x : th) x.join();
    for (int t = 0; t < nt; t++)
        if (rcs[t]) return fail(rcs[t], "%s", errs[t].c_str());
    return WEAKLABEL_OK;
}

// ---------------------------------------------------------------------------------------------
// the legacy host-pointer path (estep.py:29-31 -> weaklabel.cpp:6-7), pipelined: the batch is cut into
// blocks of images; block k+1 goes up (H2D) while block k is worked on and block k-1 comes back (D2H),
// on three streams chained by events, so that PCIe runs full duplex.  Pinned / registered host memory is
// copied directly; pageable memory (what numpy hands over) goes through a pinned staging ring filled
// and drained by a few host threads.  The class order is drawn on the host block after block from the
// process-global rand(), which is exactly the sequence of B consecutive single-image calls.
// ---------------------------------------------------------------------------------------------
struct LegacyCtx {
    int device;
    cudaStream_t s_in, s_k, s_out;
    cudaEvent_t ev_in[3], ev_k[3], ev_out[3];
    float *d_buf[3]; size_t d_cap;
    unsigned char *h_in[2], *h_out[2]; size_t h_cap;
    bool ok;
};
static LegacyCtx g_lc[16];

static void par_memcpy(void *dst, const void *src, size_t bytes) {
    const size_t MINB = (size_t)8 << 20;
    int nt = (int)(bytes / MINB);
    if (nt > 6) nt = 6;
    if (nt <= 1) { memcpy(dst, src, bytes); return; }
    std::vector<std::thread> th;
    const size_t per = ((bytes / nt) + 4095) & ~(size_t)4095;
    for (int t = 1; t < nt; t++) {
        const size_t o = (size_t)t * per;
        if (o >= bytes) break;
        const size_t n = std::min(per, bytes - o);
        th.emplace_back([=]() { memcpy((char *)dst + o, (const char *)src + o, n); });
    }
    memcpy(dst, src, std::min(per, bytes));
    for (auto &t : th) t.join();
}

static int legacy_ctx(LegacyCtx **out) {
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 16) return fail(WEAKLABEL_ERR_ARG, "device %d out of range", dev);
    LegacyCtx *lc = &g_lc[dev];
    if (!lc->ok) {
        lc->device = dev;
        CUDA_TRY(cudaStreamCreateWithFlags(&lc->s_in, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&lc->s_k, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&lc->s_out, cudaStreamNonBlocking));
        for (int k = 0; k < 3; k++) {
            CUDA_TRY(cudaEventCreateWithFlags(&lc->ev_in[k], cudaEventDisableTiming));
            CUDA_TRY(cudaEventCreateWithFlags(&lc->ev_k[k], cudaEventDisableTiming));
            CUDA_TRY(cudaEventCreateWithFlags(&lc->ev_out[k], cudaEventDisableTiming));
        }
        lc->ok = true;
    }
    *out = lc;
    return WEAKLABEL_OK;
}

static bool host_pointer_is_pinned(const void *p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

static size_t legacy_block_bytes() {
    const char *e = getenv("WEAKLABEL_BLOCK_MB");         // tuning knob: size of a pipeline block
    const long mb = e ? atol(e) : 96;
    return (size_t)(mb > 0 ? mb : 96) << 20;
}

static int legacy_impl(float *map, const int *shape, const int *stride, const int *label,
                       bool suppress_others, int num_iter, float margin_others, float bg, float fg) {
    if (!map || !label) return fail(WEAKLABEL_ERR_ARG, "map/label is NULL");
    long long s0, s2;
    int rc = check_shape_stride(shape, stride, &s0, &s2);
    if (rc) return rc;
    const int B = shape[0], H = shape[1], W = shape[2], C = shape[3];
    if (C > WEAKLABEL_MAX_CLASSES) return fail(WEAKLABEL_ERR_ARG, "C=%d > %d", C, WEAKLABEL_MAX_CLASSES);
    if (weaklabel_device_count() <= 0)
        return fail(WEAKLABEL_ERR_CUDA, "no CUDA device: libweaklabel_cuda has no CPU fallback");
    const long long N = (long long)H * W;
    if (B > 1 && s0 < (N - 1) * s2 + C) return fail(WEAKLABEL_ERR_ARG, "batch stride %d overlaps images", stride[0]);
    std::lock_guard<std::mutex> lock(g_mu);
    LegacyCtx *lc = nullptr;
    rc = legacy_ctx(&lc);
    if (rc) return rc;
    g_present.resize((size_t)B * C);
    rc = label_set_host_parallel(label, B, H, W, C, g_present.data());
    if (rc) return rc;
    // validate every label set before a single rand() draw or copy (the reference has no partial results either)
    for (int b = 0; b < B; b++) {
        int P = 0;
        for (int c = 0; c < C; c++) P += g_present[(size_t)b * C + c] ? 1 : 0;
        if (P == 0) return fail(WEAKLABEL_ERR_EMPTY_SET, "image %d has an empty label set (the reference crashes there, weaklabel.cpp:83)", b);
    }
    // blocks of nb images (~96 MB): small enough to overlap, large enough to keep the copies efficient
    const size_t img_span = ((size_t)(N - 1) * (size_t)s2 + (size_t)C) * sizeof(float);
    const size_t img_step = (size_t)s0 * sizeof(float);
    int nb = (int)std::max<size_t>(1, legacy_block_bytes() / std::max<size_t>(img_step, 1));
    if (nb > B || B <= nb + nb / 2) nb = B;
    const int nblocks = (B + nb - 1) / nb;
    const size_t blk_span = (size_t)(nb - 1) * img_step + img_span;
    if (blk_span + 256 > lc->d_cap) {
        CUDA_TRY(cudaStreamSynchronize(lc->s_out));
        for (int k = 0; k < 3; k++) { cudaFree(lc->d_buf[k]); lc->d_buf[k] = nullptr; }
        lc->d_cap = 0;
        for (int k = 0; k < (nblocks > 1 ? 3 : 1); k++) CUDA_TRY(cudaMalloc(&lc->d_buf[k], blk_span + 256));
        lc->d_cap = blk_span + 256;
    }
    for (int k = 0; k < (nblocks > 1 ? 3 : 1); k++)
        if (!lc->d_buf[k]) CUDA_TRY(cudaMalloc(&lc->d_buf[k], lc->d_cap));
    const bool pinned = host_pointer_is_pinned(map) && host_pointer_is_pinned((const char *)map + (size_t)(B - 1) * img_step + img_span - 1);
    if (!pinned && blk_span > lc->h_cap) {
        CUDA_TRY(cudaStreamSynchronize(lc->s_out));
        for (int k = 0; k < 2; k++) {
            if (lc->h_in[k]) cudaFreeHost(lc->h_in[k]);
            if (lc->h_out[k]) cudaFreeHost(lc->h_out[k]);
            lc->h_in[k] = lc->h_out[k] = nullptr;
        }
        lc->h_cap = 0;
        for (int k = 0; k < (nblocks > 1 ? 2 : 1); k++) {
            CUDA_TRY(cudaHostAlloc((void **)&lc->h_in[k], blk_span, cudaHostAllocDefault));
            CUDA_TRY(cudaHostAlloc((void **)&lc->h_out[k], blk_span, cudaHostAllocDefault));
        }
        lc->h_cap = blk_span;
    }
    if (!pinned)
        for (int k = 0; k < (nblocks > 1 ? 2 : 1); k++)
            if (!lc->h_in[k]) {
                CUDA_TRY(cudaHostAlloc((void **)&lc->h_in[k], lc->h_cap, cudaHostAllocDefault));
                CUDA_TRY(cudaHostAlloc((void **)&lc->h_out[k], lc->h_cap, cudaHostAllocDefault));
            }
    std::vector<weaklabel_plan *> used;
    auto span_of = [&](int k) { const int n = std::min(nb, B - k * nb); return (size_t)(n - 1) * img_step + img_span; };
    auto drain = [&](int k) -> int {                      // pageable: block k's result from the staging buffer to the caller
        CUDA_TRY(cudaEventSynchronize(lc->ev_out[k % 3]));
        par_memcpy((char *)map + (size_t)k * nb * img_step, lc->h_out[k % 2], span_of(k));
        return WEAKLABEL_OK;
    };
    for (int k = 0; k < nblocks; k++) {
        const int slot = k % 3, b0 = k * nb, n = std::min(nb, B - b0);
        char *host = (char *)map + (size_t)b0 * img_step;
        const size_t span = span_of(k);
        weaklabel_plan *pl = nullptr;
        rc = cached_plan(n, H, W, C, &pl);
        if (rc) return rc;
        if (std::find(used.begin(), used.end(), pl) == used.end()) used.push_back(pl);
        // ---- up
        if (k >= 3) CUDA_TRY(cudaStreamWaitEvent(lc->s_in, lc->ev_out[slot], 0));      // d_buf[slot] read back (block k-3)
        if (pinned) {
            CUDA_TRY(cudaMemcpyAsync(lc->d_buf[slot], host, span, cudaMemcpyHostToDevice, lc->s_in));
        } else {
            if (k >= 2) CUDA_TRY(cudaEventSynchronize(lc->ev_in[(k - 2) % 3]));         // h_in[k % 2] consumed (block k-2)
            par_memcpy(lc->h_in[k % 2], host, span);
            CUDA_TRY(cudaMemcpyAsync(lc->d_buf[slot], lc->h_in[k % 2], span, cudaMemcpyHostToDevice, lc->s_in));
        }
        CUDA_TRY(cudaEventRecord(lc->ev_in[slot], lc->s_in));
        // ---- E-step + in-place map (weaklabel.cpp:108-112) on the compute stream
        CUDA_TRY(cudaStreamWaitEvent(lc->s_k, lc->ev_in[slot], 0));
        rc = weaklabel_plan_prepare_async(pl, g_present.data() + (size_t)b0 * C, suppress_others, num_iter, margin_others,
                                          bg, fg, nullptr, lc->s_k);
        if (rc) return rc;
        rc = weaklabel_estep_labels_device(pl, lc->d_buf[slot], s0, s2, nullptr, lc->s_k);
        if (rc) return rc;
        rc = weaklabel_estep_finalize_map_device(pl, lc->d_buf[slot], s0, s2, lc->s_k);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(lc->ev_k[slot], lc->s_k));
        // ---- down
        CUDA_TRY(cudaStreamWaitEvent(lc->s_out, lc->ev_k[slot], 0));
        if (pinned) {
            CUDA_TRY(cudaMemcpyAsync(host, lc->d_buf[slot], span, cudaMemcpyDeviceToHost, lc->s_out));
        } else {
            if (k >= 2) { rc = drain(k - 2); if (rc) return rc; }                       // frees h_out[k % 2]
            CUDA_TRY(cudaMemcpyAsync(lc->h_out[k % 2], lc->d_buf[slot], span, cudaMemcpyDeviceToHost, lc->s_out));
        }
        CUDA_TRY(cudaEventRecord(lc->ev_out[slot], lc->s_out));
    }
    if (!pinned)
        for (int k = std::max(0, nblocks - 2); k < nblocks; k++) { rc = drain(k); if (rc) return rc; }
    CUDA_TRY(cudaStreamSynchronize(lc->s_out));           // our three streams only: the caller's streams are left alone
    for (weaklabel_plan *pl : used) {
        rc = plan_check_status(pl, nullptr);
        if (rc) return rc;
    }
    return WEAKLABEL_OK;
}

extern "C" WL_API float *e_step(float *map, const int *shape, const int *stride, const int *label,
                         const bool suppress_others, const int num_iter, const float margin_others,
                         float bg, float fg) {
    const int rc = legacy_impl(map, shape, stride, label, suppress_others, num_iter, margin_others, bg, fg);
    if (rc != WEAKLABEL_OK) {
        // the reference signature has no error channel and its callers ignore the return
        // value (estep.py:30): never hand back an unmodified map silently.
        fprintf(stderr, "libweaklabel_cuda: e_step failed (%d): %s\n", rc, weaklabel_last_error());
        fflush(stderr);
        abort();
    }
    return map;
}

extern "C" WL_API int e_step_labels_device(const float *d_map, const int *shape, const int *stride_bytes,
                                    const int *h_label, bool suppress_others, int num_iter,
                                    float margin_others, float bg, float fg, weaklabel_rng *rng,
                                    unsigned char *d_labels, float *d_map_out, void *stream) {
    if (!d_map || !h_label) return fail(WEAKLABEL_ERR_ARG, "d_map/h_label is NULL");
    long long s0, s2;
    int rc = check_shape_stride(shape, stride_bytes, &s0, &s2);
    if (rc) return rc;
    const int B = shape[0], H = shape[1], W = shape[2], C = shape[3];
    if (C > WEAKLABEL_MAX_CLASSES) return fail(WEAKLABEL_ERR_ARG, "C=%d > %d", C, WEAKLABEL_MAX_CLASSES);
    std::lock_guard<std::mutex> lock(g_mu);
    weaklabel_plan *pl = nullptr;
    rc = cached_plan(B, H, W, C, &pl);
    if (rc) return rc;
    g_present.resize((size_t)B * C);
    rc = weaklabel_label_set_host(h_label, B, H, W, C, g_present.data());
    if (rc) return rc;
    rc = weaklabel_plan_prepare_async(pl, g_present.data(), suppress_others, num_iter, margin_others, bg, fg, rng, stream);
    if (rc) return rc;
    rc = weaklabel_estep_labels_device(pl, d_map, s0, s2, d_labels, stream);
    if (rc) return rc;
    if (d_map_out) {
        const long long N = (long long)H * W;
        const size_t span = ((size_t)(B - 1) * (size_t)s0 + (size_t)(N - 1) * (size_t)s2 + (size_t)C) * sizeof(float);
        if (d_map_out != d_map)
            CUDA_TRY(cudaMemcpyAsync(d_map_out, d_map, span, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
        rc = weaklabel_estep_finalize_map_device(pl, d_map_out, s0, s2, stream);
    }
    return rc;
}

// ---------------------------------------------------------------------------------------------
// softmax cross-entropy (deeplab.py:182) and the E-step + loss sequence (deeplab.py:120-123,182)
// ---------------------------------------------------------------------------------------------
static int device_sms() {
    static int sms[16] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return 148;
    if (!sms[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        sms[dev] = n;
    }
    return sms[dev];
}

extern "C" WL_API int weaklabel_softmax_ce_device(const float *d_logits, const unsigned char *d_labels,
                                           long long n_pix, int C, int ignore_label,
                                           float *d_loss_mean, int *d_valid, float *d_dlogits,
                                           void *d_work, void *stream) {
    if (!d_logits || !d_labels || !d_work) return fail(WEAKLABEL_ERR_ARG, "weaklabel_softmax_ce_device: NULL argument");
    if (n_pix < 0 || C <= 0 || C > WEAKLABEL_MAX_CLASSES) return fail(WEAKLABEL_ERR_ARG, "bad n_pix/C");
    if ((((uintptr_t)d_logits) & 3u) || (((uintptr_t)d_work) & 7u)) return fail(WEAKLABEL_ERR_ARG, "misaligned pointer");
    if (weaklabel_device_count() <= 0)
        return fail(WEAKLABEL_ERR_CUDA, "no CUDA device: libweaklabel_cuda has no CPU fallback");
    int n = 0;
    CUDA_TRY(launch_softmax_ce(d_logits, d_labels, n_pix, C, ignore_label, d_loss_mean, d_valid,
                               d_dlogits, d_work, -1, device_sms(), (cudaStream_t)stream, &n));
    return WEAKLABEL_OK;
}

// The loss for callers that cannot allocate: no d_work argument, the scratch is a static device array of
// softmax_ce.cu.  No counterpart in the reference.  Nothing on
// this path calls cudaMalloc / cudaFree or an allocating library routine.
extern "C" WL_API int weaklabel_softmax_ce_static(const float *d_logits, const unsigned char *d_labels,
                                           long long n_pix, int C, int ignore_label,
                                           float *d_loss_mean, int *d_valid, float *d_dlogits, void *stream) {
    if (!d_logits || !d_labels) return fail(WEAKLABEL_ERR_ARG, "weaklabel_softmax_ce_static: NULL argument");
    if (n_pix < 0 || C <= 0 || C > WEAKLABEL_MAX_CLASSES) return fail(WEAKLABEL_ERR_ARG, "bad n_pix/C");
    if (n_pix > WEAKLABEL_CE_STATIC_MAX_PIXELS)
        return fail(WEAKLABEL_ERR_ARG, "weaklabel_softmax_ce_static: n_pix exceeds WEAKLABEL_CE_STATIC_MAX_PIXELS (2^31 - 1: "
                                       "the kept-pixel counter in the static scratch is 32 bits wide); split the batch");
    if (((uintptr_t)d_logits) & 3u) return fail(WEAKLABEL_ERR_ARG, "misaligned pointer");
    if (weaklabel_device_count() <= 0)
        return fail(WEAKLABEL_ERR_CUDA, "no CUDA device: libweaklabel_cuda has no CPU fallback");
    int n = 0;
    CUDA_TRY(launch_softmax_ce_static(d_logits, d_labels, n_pix, C, ignore_label, d_loss_mean, d_valid, d_dlogits,
                                      device_sms(), (cudaStream_t)stream, &n));
    return WEAKLABEL_OK;
}

// E-step labels and the loss that consumes them, queued back to back on one stream: the label map goes
// from the E-step kernel to the loss kernel through the L2, every pixel is kept (the E-step never emits
// ignore_label), so the loss is ONE launch with the denominator B * H * W known beforehand.
extern "C" WL_API int weaklabel_estep_ce_device(weaklabel_plan *pl, const float *d_map, unsigned char *d_labels,
                                         float *d_loss_mean, int *d_valid, float *d_dlogits, void *d_work,
                                         void *stream) {
    if (!pl || !d_map || !d_labels || !d_work) return fail(WEAKLABEL_ERR_ARG, "weaklabel_estep_ce_device: NULL argument");
    if ((((uintptr_t)d_work) & 7u)) return fail(WEAKLABEL_ERR_ARG, "misaligned work pointer");
    int rc = run_estep(pl, d_map, (long long)pl->N * pl->C, pl->C, d_labels, stream, 0);     // labels only
    if (rc) return rc;
    const long long n_pix = (long long)pl->B * pl->N;
    int n = 0;
    CUDA_TRY(launch_softmax_ce(d_map, d_labels, n_pix, pl->C, 255, d_loss_mean, d_valid, d_dlogits, d_work, n_pix,
                               pl->sms, (cudaStream_t)stream, &n));
    pl->launches += n;
    CUDA_TRY(cudaEventRecord(pl->ev_run, (cudaStream_t)stream));
    return WEAKLABEL_OK;
}

// ---------------------------------------------------------------------------------------------
// EM-Fixed (the fixed-bias variant of the paper the reference's readme cites)
// ---------------------------------------------------------------------------------------------
extern "C" WL_API int weaklabel_fixed_bias_labels_device(const float *d_map, int B, int H, int W, int C,
                                                  const unsigned int *d_masks, float bias_bg, float bias_fg,
                                                  const unsigned char *d_strong, unsigned char *d_labels, void *stream) {
    if (!d_map || !d_masks || !d_labels) return fail(WEAKLABEL_ERR_ARG, "weaklabel_fixed_bias_labels_device: NULL argument");
    if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || C > WEAKLABEL_MAX_CLASSES) return fail(WEAKLABEL_ERR_ARG, "bad shape");
    if (((uintptr_t)d_map) & 3u) return fail(WEAKLABEL_ERR_ARG, "misaligned pointer");
    if (weaklabel_device_count() <= 0)
        return fail(WEAKLABEL_ERR_CUDA, "no CUDA device: libweaklabel_cuda has no CPU fallback");
    CUDA_TRY(launch_fixed_bias(d_map, (long long)B * H * W, H * W, C, d_masks, bias_bg, bias_fg, d_strong, d_labels, device_sms(),
                               (cudaStream_t)stream));
    return WEAKLABEL_OK;
}

// ---------------------------------------------------------------------------------------------
// DLPack front door: the score map (and the label map) as DLTensor / DLManagedTensor (dlpack.h of
// dmlc/dlpack, third-party; a DLManagedTensor starts with its DLTensor, so either pointer works)
// ---------------------------------------------------------------------------------------------
extern "C" WL_API int weaklabel_estep_labels_dlpack(weaklabel_plan *pl, const void *dl_scores, const void *dl_labels,
                                             void *stream) {
    if (!pl || !dl_scores || !dl_labels) return fail(WEAKLABEL_ERR_ARG, "weaklabel_estep_labels_dlpack: NULL argument");
    const weaklabel_dltensor *s = (const weaklabel_dltensor *)dl_scores, *l = (const weaklabel_dltensor *)dl_labels;
    if (s->device_type != 2 || l->device_type != 2)       // kDLCUDA
        return fail(WEAKLABEL_ERR_ARG, "DLPack tensors must live in CUDA device memory (device_type 2), got %d / %d",
                    s->device_type, l->device_type);
    if (s->device_id != pl->device || l->device_id != pl->device)
        return fail(WEAKLABEL_ERR_ARG, "DLPack tensors are on device %d / %d, the plan on %d", s->device_id, l->device_id, pl->device);
    if (s->dtype_code != 2 || s->dtype_bits != 32 || s->dtype_lanes != 1) return fail(WEAKLABEL_ERR_ARG, "scores must be float32");
    if (l->dtype_code != 1 || l->dtype_bits != 8 || l->dtype_lanes != 1) return fail(WEAKLABEL_ERR_ARG, "labels must be uint8");
    if (s->ndim != 4 || s->shape[0] != pl->B || s->shape[1] != pl->H || s->shape[2] != pl->W || s->shape[3] != pl->C)
        return fail(WEAKLABEL_ERR_ARG, "scores must have shape [%d,%d,%d,%d]", pl->B, pl->H, pl->W, pl->C);
    if (l->ndim != 3 || l->shape[0] != pl->B || l->shape[1] != pl->H || l->shape[2] != pl->W)
        return fail(WEAKLABEL_ERR_ARG, "labels must have shape [%d,%d,%d]", pl->B, pl->H, pl->W);
    long long bs = (long long)pl->N * pl->C, ps = pl->C;
    if (s->strides) {                                     // element strides; NULL = compact row-major
        if (s->strides[3] != 1 || s->strides[1] != (long long)pl->W * s->strides[2])
            return fail(WEAKLABEL_ERR_ARG, "scores: classes must be contiguous and rows dense");
        bs = s->strides[0]; ps = s->strides[2];
    }
    if (l->strides && (l->strides[2] != 1 || l->strides[1] != pl->W || l->strides[0] != (long long)pl->N))
        return fail(WEAKLABEL_ERR_ARG, "labels must be contiguous");
    const float *d_map = (const float *)((const char *)s->data + s->byte_offset);
    unsigned char *d_lab = (unsigned char *)l->data + l->byte_offset;
    return weaklabel_estep_labels_device(pl, d_map, bs, ps, d_lab, stream);
}
