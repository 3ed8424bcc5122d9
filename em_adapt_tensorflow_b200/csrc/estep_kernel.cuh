// estep_kernel.cuh -- shared declarations between the E-step kernels and the C-ABI host code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>

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
    int pf_ahead;             // chunks prefetched into the L2 ahead of the staging ring
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
    volatile unsigned int *trail;  // debugging: [CTA][16] progress words in mapped host memory, or NULL
    // on-chip kernel
    int se_floats;            // floats of the running-max tile in shared memory (>= TS and >= NT * L)
    int flags;                // OC_F_* switches (A/B runs and bisection; 0 = everything on)
    int light_ok;             // the geometry holds two planes in tensor memory: label sets of <= 2 classes take the light path
    int want_sums;            // 1: mean_max_0 / mean_max_1 / shift of EVERY image are outputs (in-place map, probes);
                              // 0: labels only -- the two order-exact sums are formed only for images in which a
                              //    pixel's argmax can depend on the rounding of the shift (NaN elsewhere)
    unsigned long long *prof2; // [B][32] fine-grained phase accumulators (ns) or NULL
    // fused bilinear up-sampling (network.py:40): when up_src != NULL the scores of pixel (y, x) are NOT read from
    // `map` but interpolated from the low-resolution tensor up_src [B][up_h][up_w][C] (upsample.cuh); on-chip kernel only
    const float *up_src;
    int up_h, up_w, up_W;      // low-resolution extent; width of the full-resolution image (N = H * up_W)
    float up_sy, up_sx;        // up_h / (float)H, up_w / (float)W
};

// switches of the on-chip kernel: a set bit turns the named optimisation OFF (falls back to the simpler code path)
constexpr int OC_F_NO_RANKCOUNT = 1;      // local selection always by histogram rounds
constexpr int OC_F_NO_RESIDENT = 2;       // evict the class plane after every step (no TMEM residency across steps)
constexpr int OC_F_NO_FASTLABEL = 4;      // every pixel takes the exact (double) label path
constexpr int OC_F_LOOSE_ZEROS = 8;       // first-generation zero-count bound (counts every pixel above the old max)
constexpr int OC_F_NO_DIGEST = 16;        // sums: the walker fetches every segment from its owner (no pushed digests)
constexpr int OC_F_CLASSIC_XCHG = 32;
constexpr int OC_F_RAGGED_COPIES = 64;    // every chunk as aligned interior + cp.async head / tail floats (no whole-span copies)     // exchanges with plain DSMEM stores + arrive/wait instead of st.async

// Function attributes (dynamic shared memory, cluster size) are per DEVICE: a process that drives several GPUs
// sets them once on each.  `done` is the launcher's static bit mask by device ordinal.
inline bool first_launch_on_device(std::atomic<unsigned long long> &done) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
    const unsigned long long bit = 1ull << dev;
    return (done.fetch_or(bit) & bit) == 0;
}

size_t estep_smem_bytes();
cudaError_t launch_estep(const EStepParams &p, int n_clusters, cudaStream_t stream);
int estep_max_clusters(int G);

// on-chip kernel of the common configuration (estep_onchip.cu): warp-specialised (producer warps stream
// image t+1 while the compute threads run the selection chain of image t; the scratch holds TWO plane
// sets per resident cluster), quad (float4) pixel ownership, two class planes resident in tensor memory
constexpr int OC_NT = 896;            // compute threads per CTA
constexpr int OC_MAXG = 9;            // largest cluster (non-portable size; 15 of them are co-resident on B200)
constexpr int OC_PF_AHEAD = 0;        // chunks prefetched into the L2 ahead of the staging ring: 0 = off (measured: the producers
                                      // are bound by instruction latency, not by data; 16 ahead costs 5 % -- tools/pf_sweep.sh)
struct OnchipConfig { int nt, qpt, G, TS, L, cpx, se_floats, max_clusters; };
bool onchip_configure(int N, int C, int B, int sms, int tune_G, int tune_TS, OnchipConfig *out);
cudaError_t launch_estep_onchip(const EStepParams &p, const OnchipConfig &cfg, int n_clusters, cudaStream_t stream);

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
// the same on bulk tiles (stream_kernels.cu): contiguous batches of images with >= 256 pixels
bool finalize_tiles_applicable(const FinalizeParams &p);
cudaError_t launch_finalize_tiles(const FinalizeParams &p, int sms, cudaStream_t stream);
// EM-Fixed labels (stream_kernels.cu): argmax over the present classes of score + fixed bias, strong labels override
cudaError_t launch_fixed_bias(const float *map, long long n_pix, int N, int C, const uint32_t *masks, float bias_bg, float bias_fg,
                              const uint8_t *strong, uint8_t *labels, int sms, cudaStream_t stream);

// valid_known >= 0: the number of kept pixels is known (E-step labels never carry ignore_label): no count
// pass, a single launch; < 0: counted from the labels first.  work: 16 bytes of device scratch.
cudaError_t launch_softmax_ce(const float *logits, const uint8_t *labels, long long n_pix, int C,
                              int ignore_label, float *loss_mean, int *valid, float *dlogits,
                              void *work, long long valid_known, int sms, cudaStream_t stream, int *n_launches);

// the same with the scratch in a file-scope device array of softmax_ce.cu (no allocation anywhere; calls on one
// device share the array and must not overlap); the labels are counted first (valid_known = -1)
cudaError_t launch_softmax_ce_static(const float *logits, const uint8_t *labels, long long n_pix, int C, int ignore_label,
                                     float *loss_mean, int *valid, float *dlogits, int sms, cudaStream_t stream,
                                     int *n_launches);

// tf.image.resize_bilinear (TF1 defaults, network.py:40) of src [B][h][w][C] to dst [B][H][W][C], materialised (stream_kernels.cu)
cudaError_t launch_upsample_bilinear(const float *src, int B, int h, int w, int C, int H, int W, float *dst, cudaStream_t stream);

// GT shrink + image-level label sets on the device (deeplab.py:110, weaklabel.cpp:19-31): nearest-neighbour
// resize of a dense [B][Hg][Wg] ground-truth map (uint8 or int32) to [H][W], stride-2 scan, values >= C void;
// masks[b] = bit c set <=> class c present; flags[0] |= 1 when a negative label was seen.
cudaError_t launch_label_set(const void *gt, int gt_is_int32, int B, int Hg, int Wg, int H, int W, int C,
                             uint32_t *masks, int *flags, cudaStream_t stream);

}  // namespace wl
