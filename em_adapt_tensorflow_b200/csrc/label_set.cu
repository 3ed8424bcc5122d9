// label_set.cu -- ground-truth shrink + image-level label sets on the device (sm_100a).
//
// deeplab.py:110 shrinks the dense GT map to the fc8 size with tf.image.resize_nearest_neighbor
// (align_corners = False: source row = min(floor(row * Hg / H), Hg - 1), same for columns -- TensorFlow is
// third-party to the reference and version-unpinned; this is its documented kernel), and weaklabel.cpp:19-31
// turns the shrunk map into the image-level label set: a stride-2 scan in both directions, values >= C are
// void.  Here both happen in one pass over the (H/2) x (W/2) scanned positions, straight from the
// full-resolution map on the device -- no shrunk map is materialised and no host pass is needed.
#include "estep_kernel.cuh"

namespace wl {

template <typename T>
__global__ void __launch_bounds__(256) label_set_kernel(const T *__restrict__ gt, int Hg, int Wg, int H, int W, int C,
                                                        uint32_t *__restrict__ masks, int *__restrict__ flags) {
    __shared__ uint32_t wm[8];
    const int b = blockIdx.y;
    const int nw = (W + 1) >> 1, nh = (H + 1) >> 1;
    const float sh = (float)Hg / (float)H, sw = (float)Wg / (float)W;     // resize_nearest_neighbor scales
    const T *img = gt + (size_t)b * Hg * Wg;
    uint32_t m = 0u;
    int neg = 0;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nh * nw; idx += gridDim.x * blockDim.x) {
        const int h = 2 * (idx / nw), w = 2 * (idx % nw);
        const int y = min((int)floorf((float)h * sh), Hg - 1), x = min((int)floorf((float)w * sw), Wg - 1);
        const int t = (int)img[(size_t)y * Wg + x];
        if (t >= C) continue;                                       // void label (weaklabel.cpp:27)
        if (t < 0) { neg = 1; continue; }                           // the reference writes out of bounds here
        m |= 1u << t;
    }
    for (int o = 16; o; o >>= 1) m |= __shfl_xor_sync(0xffffffffu, m, o);
    neg = __any_sync(0xffffffffu, neg);
    if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t all = 0u;
        for (int k = 0; k < 8; k++) all |= wm[k];
        if (all) atomicOr(&masks[b], all);
    }
    if (neg && (threadIdx.x & 31) == 0) atomicOr(flags, 1);
}

cudaError_t launch_label_set(const void *gt, int gt_is_int32, int B, int Hg, int Wg, int H, int W, int C,
                             uint32_t *masks, int *flags, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(masks, 0, (size_t)B * sizeof(uint32_t), stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(flags, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    const int n = ((H + 1) / 2) * ((W + 1) / 2);
    int gx = (n + 255) / 256;
    if (gx > 64) gx = 64;
    dim3 grid((unsigned)gx, (unsigned)B);
    if (gt_is_int32) label_set_kernel<int><<<grid, 256, 0, stream>>>((const int *)gt, Hg, Wg, H, W, C, masks, flags);
    else label_set_kernel<uint8_t><<<grid, 256, 0, stream>>>((const uint8_t *)gt, Hg, Wg, H, W, C, masks, flags);
    return cudaGetLastError();
}

}  // namespace wl
