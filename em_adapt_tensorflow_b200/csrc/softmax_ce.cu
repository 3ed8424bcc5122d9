// softmax_ce.cu -- softmax cross-entropy forward + backward with ignore_label for sm_100a.
//
// Replaces deeplab.py:182 (tf.nn.sparse_softmax_cross_entropy_with_logits + tf.reduce_mean over
// the fc8 logits and the E-step labels) with the ignore semantics of network.py:77-89: a pixel
// whose label is ignore_label (255) or >= C adds no loss, gets a zero gradient row and is left
// out of the mean's denominator.  Labels are constants (tf.py_func / argmax carry no gradient),
// so dlogits = (softmax(z) - onehot(y)) / n_kept.
//
// HBM-bound: 4*C bytes read + 4*C written + 1 label byte per pixel (169 B/pixel at C = 21).  One
// persistent launch: every CTA walks tiles of 256 pixels through a double-buffered shared-memory
// ring that the copy engine fills (cp.async.bulk global -> shared, completion on an mbarrier) and
// drains (cp.async.bulk shared -> global, bulk groups), so both directions move as 21.5 KB bulk
// transfers and no thread issues a 4-byte store; one thread per pixel works on its 84-byte row in
// place (row stride C words: conflict-free for odd C).  The loss is reduced in double; the last
// CTA to finish writes the mean.  When the number of kept pixels is known beforehand (the E-step
// never emits 255: every pixel is kept) the count pass is skipped and this is the only launch.
#include <float.h>
#include <math.h>
#include <mutex>
#include "estep_kernel.cuh"
#include "bulk_tile.cuh"
#include "../../include/weaklabel_cuda.h"

namespace wl {

constexpr int CE_NT = 256;          // threads per CTA = pixels per tile
constexpr int CE_STAGES = 2;

struct CeWork {                     // 16 bytes of device scratch, zeroed by the call
    double loss_sum;
    int valid;
    unsigned int done;
};

__global__ void __launch_bounds__(CE_NT) ce_count_kernel(const uint8_t *__restrict__ labels, long long n_pix,
                                                         int C, int ignore_label, CeWork *work) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    int cnt = 0;
    for (; i < n_pix; i += stride) {
        const int y = labels[i];
        cnt += (y != ignore_label && y < C) ? 1 : 0;
    }
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    __shared__ int ws[CE_NT / 32];
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < CE_NT / 32; w++) t += ws[w];
        if (t) atomicAdd(&work->valid, t);
    }
}

// one pixel: z (its C logits in shared memory) becomes the gradient row; returns the pixel's loss
__device__ __forceinline__ double ce_pixel(float *z, int C, int y, bool keep, float inv) {
    if (!keep) {
        for (int k = 0; k < C; k++) z[k] = 0.f;
        return 0.0;
    }
    float m = z[0];
    for (int k = 1; k < C; k++) m = fmaxf(m, z[k]);
    float se = 0.f, others = 0.f;
    const float zy = z[y];
    for (int k = 0; k < C; k++) {
        const float e = ce_exp(z[k] - m);
        se += e;
        if (k != y) others += e;
        z[k] = e;
    }
    const float r = inv / se;
    for (int k = 0; k < C; k++) z[k] = (k == y) ? -others * r : z[k] * r;     // p_y - 1 = -(sum of the others) / se
    return (double)(logf(se) + m) - (double)zy;
}

// valid_known >= 0: the number of kept pixels (no count pass ran); < 0: read it from work->valid
__global__ void __launch_bounds__(CE_NT) ce_fused_kernel(const float *__restrict__ logits, const uint8_t *__restrict__ labels,
                                                         long long n_pix, int C, int ignore_label, float *__restrict__ dlogits,
                                                         CeWork *work, float *loss_mean, int *valid_out, long long valid_known,
                                                         int bulk_ok) {
    extern __shared__ __align__(128) unsigned char ce_smem[];
    __shared__ uint64_t full_bar[CE_STAGES];
    __shared__ double ws[CE_NT / 32];
    const int tid = threadIdx.x;
    const size_t tile_bytes = (size_t)CE_NT * C * sizeof(float);
    const long long ntiles = (n_pix + CE_NT - 1) / CE_NT;
    const long long valid = valid_known >= 0 ? valid_known : (long long)work->valid;
    const float inv = valid > 0 ? 1.0f / (float)valid : 0.f;
    if (tid == 0) {
        for (int s = 0; s < CE_STAGES; s++) ce_mbar_init(&full_bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();
    // a tile is moved by the copy engine when it is a full tile (1024 * C bytes: a multiple of 16) of 16-byte
    // aligned tensors; the ragged last tile (and misaligned tensors) take plain coalesced loads / stores
    auto is_bulk = [&](long long t) { return bulk_ok && (t + 1) * CE_NT <= n_pix; };
    auto issue = [&](long long t, int s) {              // thread 0
        if (is_bulk(t)) {
            ce_mbar_expect_tx(&full_bar[s], (uint32_t)tile_bytes);
            ce_bulk_g2s(ce_smem + (size_t)s * tile_bytes, logits + t * CE_NT * C, (uint32_t)tile_bytes, &full_bar[s]);
        }
    };
    double loss = 0.0;
    long long t = blockIdx.x;
    if (tid == 0 && t < ntiles) issue(t, 0);
    uint32_t it = 0;
    for (; t < ntiles; t += gridDim.x, it++) {
        const int s = it & 1;
        float *buf = reinterpret_cast<float *>(ce_smem + (size_t)s * tile_bytes);
        const long long tn = t + gridDim.x;
        if (tid == 0 && tn < ntiles) {
            // the other buffer was handed to the copy engine (store of the previous tile): its reads must be
            // done before the next load lands in it
            asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
            issue(tn, s ^ 1);
        }
        const long long pix0 = t * CE_NT;
        const int npx = (int)min((long long)CE_NT, n_pix - pix0);
        const bool bulk = is_bulk(t);
        if (bulk) {
            ce_mbar_wait(&full_bar[s], (it >> 1) & 1u);
        } else {
            if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
            __syncthreads();
            const float *g = logits + pix0 * C;
            for (int f = tid; f < npx * C; f += CE_NT) buf[f] = g[f];
            __syncthreads();
        }
        if (tid < npx) {
            const int y = labels[pix0 + tid];
            loss += ce_pixel(buf + tid * C, C, y, y != ignore_label && y < C, inv);
        }
        if (dlogits) {
            if (bulk) {
                asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");     // my row -> visible to the copy engine
                __syncthreads();
                if (tid == 0) ce_bulk_s2g(dlogits + pix0 * C, buf, (uint32_t)tile_bytes);
            } else {
                __syncthreads();
                float *o = dlogits + pix0 * C;
                for (int f = tid; f < npx * C; f += CE_NT) o[f] = buf[f];
                __syncthreads();
            }
        } else {
            __syncthreads();                            // buffer free for the load after next
        }
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");      // my stores are complete
    for (int o = 16; o; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
    if ((tid & 31) == 0) ws[tid >> 5] = loss;
    __syncthreads();
    if (tid == 0) {
        double tsum = 0.0;
        for (int w = 0; w < CE_NT / 32; w++) tsum += ws[w];
        atomicAdd(&work->loss_sum, tsum);
        __threadfence();
        const unsigned int prev = atomicAdd(&work->done, 1u);
        if (prev == gridDim.x - 1) {                    // the last CTA: every partial sum is in
            __threadfence();
            const double total = atomicAdd(&work->loss_sum, 0.0);
            if (loss_mean) *loss_mean = valid > 0 ? (float)(total / (double)valid) : 0.f;
            if (valid_out) *valid_out = (int)valid;
        }
    }
}

cudaError_t launch_softmax_ce(const float *logits, const uint8_t *labels, long long n_pix, int C,
                              int ignore_label, float *loss_mean, int *valid, float *dlogits,
                              void *work, long long valid_known, int sms, cudaStream_t stream, int *n_launches) {
    CeWork *w = reinterpret_cast<CeWork *>(work);
    cudaError_t e = cudaMemsetAsync(w, 0, sizeof(CeWork), stream);
    if (e != cudaSuccess) return e;
    const long long nb = (n_pix + CE_NT - 1) / CE_NT;
    if (valid_known < 0 && n_pix > 0) {
        const int cgrid = (int)min(nb, (long long)sms * 8);
        ce_count_kernel<<<cgrid, CE_NT, 0, stream>>>(labels, n_pix, C, ignore_label, w);
        if (n_launches) *n_launches += 1;
    }
    const size_t tile_bytes = (size_t)CE_NT * C * sizeof(float);
    const size_t smem = CE_STAGES * tile_bytes;
    static std::atomic<unsigned long long> attr_set{0};
    if (first_launch_on_device(attr_set)) {
        e = cudaFuncSetAttribute(ce_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(CE_STAGES * CE_NT * MAXC * sizeof(float)));
        if (e != cudaSuccess) return e;
    }
    // resident CTAs per SM by shared memory (43 KB per CTA at C = 21 -> 5); one persistent wave
    int per_sm = (int)(220 * 1024 / (smem + 1024));
    if (per_sm > 8) per_sm = 8;
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sms * per_sm;
    if (grid > nb) grid = nb;
    if (grid < 1) grid = 1;
    const int bulk_ok = ((((uintptr_t)logits) & 15u) == 0 && (dlogits == nullptr || (((uintptr_t)dlogits) & 15u) == 0)) ? 1 : 0;
    ce_fused_kernel<<<(unsigned)grid, CE_NT, smem, stream>>>(logits, labels, n_pix, C, ignore_label, dlogits, w, loss_mean,
                                                             valid, valid_known, bulk_ok);
    if (n_launches) *n_launches += 1;
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// The same two kernels for callers that cannot allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc): the 16 bytes of CeWork live in a file-scope device array instead of a caller's
// buffer.  Every GPU that loads the library holds the array until exit; every call on a device shares it
// (it is cleared on the stream at the start of each call), so calls that overlap on one GPU race.
// In host code `ce_static_scratch` is only a stand-in for the symbol: its device address is resolved
// once per device, with that device current, by cudaGetSymbolAddress.
// ---------------------------------------------------------------------------------------------
constexpr size_t CE_STATIC_BYTES = WEAKLABEL_CE_STATIC_SCRATCH_BYTES;      // the stated maximum (the header documents it)
constexpr int CE_STATIC_MAX_DEVICES = 64;
static_assert(sizeof(CeWork) <= CE_STATIC_BYTES, "the static scratch holds one CeWork");
static __device__ __align__(256) unsigned char ce_static_scratch[CE_STATIC_BYTES];

static cudaError_t ce_static_scratch_address(void **out) {
    static void *addr[CE_STATIC_MAX_DEVICES];
    static bool have[CE_STATIC_MAX_DEVICES];
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= CE_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!have[dev]) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, ce_static_scratch);      // the copy of the array on the CURRENT device
        if (e != cudaSuccess) return e;
        if (!ptr) return cudaErrorInvalidSymbol;
        addr[dev] = ptr;
        have[dev] = true;
    }
    *out = addr[dev];
    return cudaSuccess;
}

cudaError_t launch_softmax_ce_static(const float *logits, const uint8_t *labels, long long n_pix, int C, int ignore_label,
                                     float *loss_mean, int *valid, float *dlogits, int sms, cudaStream_t stream,
                                     int *n_launches) {
    void *work = nullptr;
    cudaError_t e = ce_static_scratch_address(&work);
    if (e != cudaSuccess) return e;
    return launch_softmax_ce(logits, labels, n_pix, C, ignore_label, loss_mean, valid, dlogits, work, -1, sms, stream,
                             n_launches);
}

}  // namespace wl
