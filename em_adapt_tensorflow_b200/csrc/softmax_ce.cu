// softmax_ce.cu -- softmax cross-entropy forward + backward with ignore_label for sm_100a.
//
// Replaces deeplab.py:182 (tf.nn.sparse_softmax_cross_entropy_with_logits + tf.reduce_mean over
// the fc8 logits and the E-step labels) with the ignore semantics of network.py:77-89: a pixel
// whose label is ignore_label (255) or >= C adds no loss, gets a zero gradient row and is left
// out of the mean's denominator.  Labels are constants (tf.py_func / argmax carry no gradient),
// so dlogits = (softmax(z) - onehot(y)) / n_kept.
//
// HBM-bound: 4*C bytes read + 4*C written + 1 label byte per pixel.  The flat float stream is
// staged through shared memory with 16-byte cp.async so that the 84-byte pixels are read and
// written fully coalesced; the gradient overwrites the staged logits in place.
#include <float.h>
#include <math.h>
#include "estep_kernel.cuh"

namespace wl {

constexpr int CE_NT = 256;

struct CeWork {          // 16 bytes of device scratch
    double loss_sum;
    int valid;
    int pad;
};

__device__ __forceinline__ void ce_cp_async16(void *smem_dst, const void *gsrc) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void ce_cp_async4(void *smem_dst, const void *gsrc) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(gsrc) : "memory");
}

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

__global__ void __launch_bounds__(CE_NT) ce_fwd_bwd_kernel(const float *__restrict__ logits,
                                                           const uint8_t *__restrict__ labels,
                                                           long long n_pix, int C, int ignore_label,
                                                           float *__restrict__ dlogits, CeWork *work) {
    extern __shared__ __align__(16) float s_z[];       // [lead + CE_NT*C (+3)]
    __shared__ double ws[CE_NT / 32];
    const int tid = threadIdx.x;
    const long long pix0 = (long long)blockIdx.x * CE_NT;
    const int npx = (int)min((long long)CE_NT, n_pix - pix0);
    const int nfl = npx * C;
    const float *g0 = logits + pix0 * C;
    const int lead = (int)(((uintptr_t)g0 & 15u) >> 2);
    {
        const char *a0 = (const char *)g0 - lead * 4;
        const int nchunk = (lead + nfl + 3) >> 2;
        for (int q = tid; q < nchunk; q += CE_NT) {
            const int first = q * 4 - lead;
            const char *src = a0 + (size_t)q * 16;
            char *dst = (char *)s_z + (size_t)q * 16;
            if (first >= 0 && first + 4 <= nfl) ce_cp_async16(dst, src);
            else {
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const int fi = first + e;
                    if (fi >= 0 && fi < nfl) ce_cp_async4(dst + e * 4, src + e * 4);
                }
            }
        }
        asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
    }
    const int valid = work->valid;                     // written by ce_count_kernel (earlier launch)
    const float inv = valid > 0 ? 1.0f / (float)valid : 0.f;
    __syncthreads();
    double loss = 0.0;
    if (tid < npx) {
        float *z = s_z + lead + tid * C;
        const int y = labels[pix0 + tid];
        const bool keep = (y != ignore_label && y < C);
        if (keep) {
            float m = z[0];
            for (int k = 1; k < C; k++) m = fmaxf(m, z[k]);
            float se = 0.f, others = 0.f;
            const float zy = z[y];
            for (int k = 0; k < C; k++) {
                const float e = expf(z[k] - m);
                se += e;
                if (k != y) others += e;
                z[k] = e;
            }
            loss = (double)(logf(se) + m) - (double)zy;
            const float r = inv / se;
            for (int k = 0; k < C; k++) z[k] = (k == y) ? -others * r : z[k] * r;
        } else {
            for (int k = 0; k < C; k++) z[k] = 0.f;
        }
    }
    __syncthreads();
    if (dlogits) {
        float *o = dlogits + pix0 * C;
        for (int f = tid; f < nfl; f += CE_NT) o[f] = s_z[lead + f];
    }
    for (int o = 16; o; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
    if ((tid & 31) == 0) ws[tid >> 5] = loss;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < CE_NT / 32; w++) t += ws[w];
        atomicAdd(&work->loss_sum, t);
    }
}

__global__ void ce_finish_kernel(const CeWork *work, float *loss_mean, int *valid) {
    const int v = work->valid;
    if (loss_mean) *loss_mean = v > 0 ? (float)(work->loss_sum / (double)v) : 0.f;
    if (valid) *valid = v;
}

cudaError_t launch_softmax_ce(const float *logits, const uint8_t *labels, long long n_pix, int C,
                              int ignore_label, float *loss_mean, int *valid, float *dlogits,
                              void *work, cudaStream_t stream, int *n_launches) {
    CeWork *w = reinterpret_cast<CeWork *>(work);
    cudaError_t e = cudaMemsetAsync(w, 0, sizeof(CeWork), stream);
    if (e != cudaSuccess) return e;
    if (n_pix > 0) {
        const long long nb = (n_pix + CE_NT - 1) / CE_NT;
        const int cgrid = (int)min(nb, (long long)148 * 16);
        ce_count_kernel<<<cgrid, CE_NT, 0, stream>>>(labels, n_pix, C, ignore_label, w);
        const size_t smem = ((size_t)CE_NT * C + 8) * sizeof(float);
        ce_fwd_bwd_kernel<<<(unsigned)nb, CE_NT, smem, stream>>>(logits, labels, n_pix, C, ignore_label, dlogits, w);
        if (n_launches) *n_launches += 2;
    }
    ce_finish_kernel<<<1, 1, 0, stream>>>(w, loss_mean, valid);
    if (n_launches) *n_launches += 1;
    return cudaGetLastError();
}

}  // namespace wl
