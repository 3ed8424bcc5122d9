// estep_kernel.cuh -- shared declarations between the E-step kernels and the C-ABI host code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace wl {

constexpr int NT = 1024;              // threads per CTA (one CTA per SM)
constexpr int NWARPS = NT / 32;
constexpr int HBINS = 2048;           // radix-select bins per pass
constexpr int MAXC = 32;              // classes (label sets are 32-bit masks)
constexpr int MAX_CLUSTER = 16;
constexpr int TS_MAX = 33024;         // pixels per tile: one tile's terms fit in shared memory
constexpr int STREAM_BUF_FLOATS = 21512;   // 1024 px * 21 classes + 8 floats of alignment slack

// Per-image label-set description, prepared on the host (weaklabel.cpp:19-43).
struct ImageMeta {
    uint32_t mask;            // bit c set <=> class c present
    int32_t P;                // number of present classes
    int32_t first_absent;     // lowest absent class id, -1 when P == C
    int32_t pad;
    uint8_t plist[MAXC];      // ascending present class ids; plane p holds class plist[p]
};

// Per-cluster control block in global memory (lives in L2): merged histograms and the
// hand-over cells of the order-exact sums.
struct SlotCtl {
    uint32_t hist[3][HBINS];  // rotating: pass n uses hist[n%3], zeroes hist[(n+1)%3]
    uint32_t below[3][4];     // [n%3][0] = #keys below the bracket
    double tot[2][MAX_CLUSTER];   // tile totals of the current round of an exact sum
    unsigned long long carry; // running-sum bits handed from CTA to CTA
    int32_t image;            // image this cluster is working on
    int32_t pad[13];
};

struct EStepParams {
    const float *map;         // scores, element (b,i,c) at map + b*s0 + i*s2 + c
    long long s0, s2;
    int B, N, C;
    int num_iter, suppress;
    float margin;
    int nth_bg, nth_fg;       // (int)(bg*N), (int)(fg*N)   weaklabel.cpp:88
    const ImageMeta *meta;    // [B]
    const uint8_t *order;     // [B][num_iter][C] plane index visited at (iteration, position)
    // per-cluster scratch
    float *planes;            // [slots][(C+1)][Npad]: P compact class planes, last = absent max
    uint8_t *absent_idx;      // [slots][Npad] first absent class achieving the absent max
    SlotCtl *ctl;             // [slots]
    int Npad, TS, ntiles, L, G;
    int sub_px;               // pixels per streaming sub-tile
    int pf_ahead;             // v3: chunks prefetched into the L2 ahead of the staging ring
    int *image_counter;
    const int *sched;         // [B] order in which images are claimed (largest label sets first), or NULL
    int *status;              // device-side error flag (0 = ok)
    // outputs
    uint8_t *labels;          // [B][N]
    float *bias_out;          // [B][num_iter][C] by (iteration, position)
    float *mean0;             // [B]
    double *mean1;            // [B]
    double *shift;            // [B]  (double)mean0 - mean1
    unsigned long long *prof; // [B][8] phase time stamps (globaltimer ns) or NULL
    volatile unsigned int *trail;  // v3 debugging: [CTA][16] progress words in mapped host memory, or NULL
    // v4
    int se_floats;            // floats of the running-max tile in shared memory (>= TS and >= NT * L)
    int flags;                // V4_F_* switches (A/B runs and bisection; 0 = everything on)
    unsigned long long *prof2; // [B][32] fine-grained phase accumulators (ns) or NULL
};

// v4 switches: a set bit turns the named optimisation OFF (falls back to the simpler code path)
constexpr int V4_F_NO_RANKCOUNT = 1;      // local selection always by histogram rounds
constexpr int V4_F_NO_RESIDENT = 2;       // evict the class plane after every step (no TMEM residency across steps)
constexpr int V4_F_NO_FASTLABEL = 4;      // every pixel takes the exact (double) label path
constexpr int V4_F_LOOSE_ZEROS = 8;       // first-generation zero-count bound (counts every pixel above the old max)
constexpr int V4_F_NO_DIGEST = 16;        // sums: the walker fetches every segment from its owner (no pushed digests)
constexpr int V4_F_CLASSIC_XCHG = 32;     // exchanges with plain DSMEM stores + arrive/wait instead of st.async

size_t estep_smem_bytes();
cudaError_t launch_estep(const EStepParams &p, int n_clusters, cudaStream_t stream);
int estep_max_clusters(int G);

// on-chip fast path (estep_fast.cu): one tile of at most FAST_PPT_MAX * NT pixels per CTA
constexpr int FAST_PPT_SMALL = 4;
constexpr int FAST_PPT_MAX = 33;
int fast_ppt_for(int tile_pixels);                     // 0 when the tile does not fit
cudaError_t launch_estep_fast(const EStepParams &p, int n_clusters, cudaStream_t stream);
int estep_fast_max_clusters(int G, int tile_pixels);

// second-generation kernel of the common configuration (estep_v2.cu)
constexpr int V2_PPT_SMALL = 4;
constexpr int V2_PPT_MAX = 33;
struct V2Config { int nt, ppt, G, TS, L, cpx, max_clusters; };
bool v2_configure(int N, int C, int B, int sms, int tune_G, int tune_TS, int tune_nt, V2Config *out);
cudaError_t launch_estep_v2(const EStepParams &p, const V2Config &cfg, int n_clusters, cudaStream_t stream);

// third-generation kernel (estep_v3.cu): warp-specialised, producers stream image t+1 while the
// compute threads work on image t; the scratch holds TWO plane sets per resident cluster
constexpr int V3_NT = 896;            // compute threads per CTA
constexpr int V3_NPW = 4;             // producer warps per CTA
constexpr int V3_NST = 8;             // stages of the staging ring: a multiple of V3_NPW, so that a stage is
                                      // always filled for, and drained by, the same warp (which then waits for
                                      // its phases in order -- a warp must never wait one phase ahead)
constexpr int V3_STAGE_BYTES = 7168;
constexpr int V3_PF_AHEAD = 16;       // chunks prefetched into the L2 ahead of the ring
constexpr int V3_MAXG = 9;
constexpr int V3_PPT_SMALL = 4;
constexpr int V3_PPT_MID = 12;
constexpr int V3_PPT_BIG = 33;
constexpr int V3_PPT_MAX = 37;
struct V3Config { int nt, ppt, G, TS, L, cpx, max_clusters; };
bool v3_configure(int N, int C, int B, int sms, int tune_G, int tune_TS, V3Config *out);
cudaError_t launch_estep_v3(const EStepParams &p, const V3Config &cfg, int n_clusters, cudaStream_t stream);

// fourth-generation kernel (estep_v4.cu): quad (float4) pixel ownership, two class planes resident in
// tensor memory, branch-free histogram pass, st.async exchanges, digest-based order-exact sums
constexpr int V4_NT = 896;
constexpr int V4_MAXG = 9;
struct V4Config { int nt, qpt, G, TS, L, cpx, se_floats, max_clusters; };
bool v4_configure(int N, int C, int B, int sms, int tune_G, int tune_TS, V4Config *out);
cudaError_t launch_estep_v4(const EStepParams &p, const V4Config &cfg, int n_clusters, cudaStream_t stream);

struct FinalizeParams {
    float *map; long long s0, s2;
    int B, N, C, num_iter, suppress;
    float margin;
    const ImageMeta *meta;
    const uint8_t *order;
    const float *bias;        // [B][num_iter][C]
    const double *shift;      // [B]
};
cudaError_t launch_finalize_map(const FinalizeParams &p, cudaStream_t stream);

cudaError_t launch_softmax_ce(const float *logits, const uint8_t *labels, long long n_pix, int C,
                              int ignore_label, float *loss_mean, int *valid, float *dlogits,
                              void *work, cudaStream_t stream, int *n_launches);

}  // namespace wl
