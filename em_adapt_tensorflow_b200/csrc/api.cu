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

extern "C" WL_API const char *weaklabel_last_error(void) { return g_err.c_str(); }
extern "C" WL_API int weaklabel_version(void) { return 100; }
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
    unsigned long long *d_prof; bool prof_on;       // [B][32] phase time stamps
    bool geometry_ok;
    int scratch_slots;
    int *d_sched;
    int last_path;
    float *d_planes;
    uint8_t *d_absidx;
    SlotCtl *d_ctl;
    int *d_counter;     // [0] image counter, [1] status
    ImageMeta *d_meta;
    uint8_t *d_order;
    size_t order_cap;
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
    CUDA_TRY(cudaSetDevice(device));
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
    pl->last_stream = nullptr;
    pl->B = B; pl->H = H; pl->W = W; pl->C = C; pl->N = H * W;
    pl->device = device;
    pl->sms = prop.multiProcessorCount;
    cudaError_t e = cudaMalloc(&pl->d_counter, 2 * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_meta, (size_t)B * sizeof(ImageMeta));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_mean0, (size_t)B * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_mean1, (size_t)B * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_shift, (size_t)B * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_sched, (size_t)B * sizeof(int));
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
    cudaSetDevice(pl->device);
    plan_free_geometry(pl);
    cudaFree(pl->d_counter); cudaFree(pl->d_meta); cudaFree(pl->d_order); cudaFree(pl->d_bias);
    cudaFree(pl->d_mean0); cudaFree(pl->d_mean1); cudaFree(pl->d_shift); cudaFree(pl->d_sched); cudaFree(pl->d_prof);
    delete pl;
}

extern "C" WL_API int weaklabel_plan_tune(weaklabel_plan *pl, int cluster_size, int tile_pixels, int max_clusters) {
    if (!pl) return fail(WEAKLABEL_ERR_ARG, "plan is NULL");
    if (cluster_size < 0 || tile_pixels < 0 || max_clusters < 0 || tile_pixels > TS_MAX)
        return fail(WEAKLABEL_ERR_ARG, "bad tuning values");
    CUDA_TRY(cudaSetDevice(pl->device));
    if (pl->last_stream || pl->geometry_ok) CUDA_TRY(cudaDeviceSynchronize());
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
    CUDA_TRY(cudaSetDevice(pl->device));
    if (pl->last_stream) CUDA_TRY(cudaStreamSynchronize(pl->last_stream));
    if (out) CUDA_TRY(cudaMemcpy(out, pl->d_prof, (size_t)pl->B * 32 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    pl->prof_on = enable != 0;
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_plan_prepare(weaklabel_plan *pl, const unsigned char *present,
                                      bool suppress_others, int num_iter, float margin_others,
                                      float bg, float fg, weaklabel_rng *rng) {
    if (!pl || !present) return fail(WEAKLABEL_ERR_ARG, "weaklabel_plan_prepare: NULL argument");
    if (num_iter < 0) return fail(WEAKLABEL_ERR_ARG, "num_iter %d < 0", num_iter);
    const int B = pl->B, C = pl->C, N = pl->N;
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
    CUDA_TRY(cudaSetDevice(pl->device));
    if (pl->last_stream) CUDA_TRY(cudaStreamSynchronize(pl->last_stream));   // tables may be in use
    const size_t osz = pl->h_order.size() > 0 ? pl->h_order.size() : 1;
    if (osz > pl->order_cap) {
        cudaFree(pl->d_order); pl->d_order = nullptr;
        cudaFree(pl->d_bias); pl->d_bias = nullptr;
        CUDA_TRY(cudaMalloc(&pl->d_order, osz));
        CUDA_TRY(cudaMalloc(&pl->d_bias, osz * sizeof(float)));
        pl->order_cap = osz;
    }
    {   // claim order: largest label sets first (longest jobs first keeps the tail short)
        std::vector<int> sched(B);
        for (int b = 0; b < B; b++) sched[b] = b;
        std::stable_sort(sched.begin(), sched.end(),
                         [&](int a, int b2) { return pl->h_meta[a].P > pl->h_meta[b2].P; });
        CUDA_TRY(cudaMemcpy(pl->d_sched, sched.data(), (size_t)B * sizeof(int), cudaMemcpyHostToDevice));
    }
    CUDA_TRY(cudaMemcpy(pl->d_meta, pl->h_meta.data(), (size_t)B * sizeof(ImageMeta), cudaMemcpyHostToDevice));
    if (!pl->h_order.empty())
        CUDA_TRY(cudaMemcpy(pl->d_order, pl->h_order.data(), pl->h_order.size(), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemset(pl->d_bias, 0, osz * sizeof(float)));
    pl->num_iter = num_iter;
    pl->suppress = suppress_others ? 1 : 0;
    pl->margin = margin_others;
    pl->nth_bg = nth_bg;
    pl->nth_fg = nth_fg;
    pl->prepared = true;
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_estep_labels_device(weaklabel_plan *pl, const float *d_map,
                                             long long batch_stride, long long pixel_stride,
                                             unsigned char *d_labels, void *stream) {
    if (!pl || !d_map) return fail(WEAKLABEL_ERR_ARG, "weaklabel_estep_labels_device: NULL argument");
    if (!pl->prepared) return fail(WEAKLABEL_ERR_STATE, "plan not prepared");
    if (pixel_stride < pl->C || (pl->B > 1 && batch_stride < (long long)(pl->N - 1) * pixel_stride + pl->C))
        return fail(WEAKLABEL_ERR_ARG, "strides overlap: batch %lld pixel %lld", batch_stride, pixel_stride);
    if (((uintptr_t)d_map & 3u) != 0) return fail(WEAKLABEL_ERR_ARG, "d_map is not 4-byte aligned");
    CUDA_TRY(cudaSetDevice(pl->device));
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
    int slots_used = pl->slots;
    p.Npad = pl->Npad; p.sub_px = pl->sub_px;
    if (path == 2) {
        p.TS = pl->oc.TS; p.ntiles = pl->oc.G; p.L = pl->oc.L; p.G = pl->oc.G; p.sub_px = pl->oc.cpx;
        p.se_floats = pl->oc.se_floats;
        slots_used = pl->oc_slots;
        p.pf_ahead = pl->env_prefetch;
        p.flags = pl->env_flags;
        p.trail = debug_trail();
    } else {
        p.TS = pl->TS; p.ntiles = pl->ntiles; p.L = pl->L; p.G = pl->G;
    }
    p.sched = pl->d_sched;
    p.image_counter = pl->d_counter; p.status = pl->d_counter + 1;
    p.labels = d_labels; p.bias_out = pl->d_bias;
    p.mean0 = pl->d_mean0; p.mean1 = pl->d_mean1; p.shift = pl->d_shift;
    p.prof = pl->prof_on ? pl->d_prof : nullptr;
    CUDA_TRY(cudaMemsetAsync(pl->d_counter, 0, 2 * sizeof(int), st));
    if (path == 2) {
        CUDA_TRY(launch_estep_onchip(p, pl->oc, slots_used, st));
    } else {
        CUDA_TRY(cudaMemsetAsync(pl->d_ctl, 0, (size_t)slots_used * sizeof(SlotCtl), st));
        CUDA_TRY(launch_estep(p, slots_used, st));
    }
    pl->last_path = path;
    pl->launches += 1;
    pl->last_stream = st;
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_estep_finalize_map_device(weaklabel_plan *pl, float *d_map,
                                                   long long batch_stride, long long pixel_stride,
                                                   void *stream) {
    if (!pl || !d_map) return fail(WEAKLABEL_ERR_ARG, "weaklabel_estep_finalize_map_device: NULL argument");
    if (!pl->prepared) return fail(WEAKLABEL_ERR_STATE, "plan not prepared");
    CUDA_TRY(cudaSetDevice(pl->device));
    FinalizeParams f;
    memset(&f, 0, sizeof f);
    f.map = d_map; f.s0 = batch_stride; f.s2 = pixel_stride;
    f.B = pl->B; f.N = pl->N; f.C = pl->C; f.num_iter = pl->num_iter; f.suppress = pl->suppress;
    f.margin = pl->margin; f.meta = pl->d_meta; f.order = pl->d_order; f.bias = pl->d_bias;
    f.shift = pl->d_shift;
    CUDA_TRY(launch_finalize_map(f, (cudaStream_t)stream));
    pl->launches += 1;
    pl->last_stream = (cudaStream_t)stream;
    return WEAKLABEL_OK;
}

static int plan_check_status(weaklabel_plan *pl) {
    int st[2] = {0, 0};
    CUDA_TRY(cudaMemcpy(st, pl->d_counter, sizeof st, cudaMemcpyDeviceToHost));
    if (st[1] != 0)
        return fail(WEAKLABEL_ERR_CUDA, "E-step kernel reported an inconsistent selection (code %d): "
                    "NaN or Inf in the score map?", st[1]);
    return WEAKLABEL_OK;
}

extern "C" WL_API int weaklabel_plan_probes(weaklabel_plan *pl, int *order, float *bias, float *mean0, double *mean1) {
    if (!pl || !pl->prepared) return fail(WEAKLABEL_ERR_STATE, "plan not prepared");
    CUDA_TRY(cudaSetDevice(pl->device));
    CUDA_TRY(cudaStreamSynchronize(pl->last_stream));
    int rc = plan_check_status(pl);
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
// cached plan for the one-call entry points
// ---------------------------------------------------------------------------------------------
static std::mutex g_mu;
static weaklabel_plan *g_plan = nullptr;
static float *g_dmap = nullptr;
static size_t g_dmap_cap = 0;
static std::vector<unsigned char> g_present;

static int cached_plan(int B, int H, int W, int C, weaklabel_plan **out) {
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (g_plan && (g_plan->B != B || g_plan->H != H || g_plan->W != W || g_plan->C != C || g_plan->device != dev)) {
        weaklabel_plan_destroy(g_plan);
        g_plan = nullptr;
    }
    if (!g_plan) {
        int rc = weaklabel_plan_create(&g_plan, B, H, W, C, dev);
        if (rc) return rc;
    }
    *out = g_plan;
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
    std::lock_guard<std::mutex> lock(g_mu);
    weaklabel_plan *pl = nullptr;
    rc = cached_plan(B, H, W, C, &pl);
    if (rc) return rc;
    g_present.resize((size_t)B * C);
    rc = weaklabel_label_set_host(label, B, H, W, C, g_present.data());
    if (rc) return rc;
    rc = weaklabel_plan_prepare(pl, g_present.data(), suppress_others, num_iter, margin_others, bg, fg, nullptr);
    if (rc) return rc;
    const long long N = (long long)H * W;
    if (B > 1 && s0 < (N - 1) * s2 + C) return fail(WEAKLABEL_ERR_ARG, "batch stride %d overlaps images", stride[0]);
    const size_t span = ((size_t)(B - 1) * (size_t)s0 + (size_t)(N - 1) * (size_t)s2 + (size_t)C) * sizeof(float);
    if (span > g_dmap_cap) {
        cudaFree(g_dmap); g_dmap = nullptr; g_dmap_cap = 0;
        CUDA_TRY(cudaMalloc(&g_dmap, span + 256));
        g_dmap_cap = span;
    }
    CUDA_TRY(cudaMemcpy(g_dmap, map, span, cudaMemcpyHostToDevice));
    rc = weaklabel_estep_labels_device(pl, g_dmap, s0, s2, nullptr, nullptr);
    if (rc) return rc;
    rc = weaklabel_estep_finalize_map_device(pl, g_dmap, s0, s2, nullptr);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpy(map, g_dmap, span, cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaDeviceSynchronize());
    return plan_check_status(pl);
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
    rc = weaklabel_plan_prepare(pl, g_present.data(), suppress_others, num_iter, margin_others, bg, fg, rng);
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
                               d_dlogits, d_work, (cudaStream_t)stream, &n));
    return WEAKLABEL_OK;
}
