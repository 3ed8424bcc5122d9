// upsample.cuh -- bilinear up-sampling of a low-resolution score tensor, one value at a time.
//
// network.py:40 (`tf.image.resize_bilinear(output, (h, w))`) with the defaults of TensorFlow 1.x -- the
// reference's framework, a third-party dependency that is not part of /root/reference and has no pinned
// version there (it uses tf.py_func / tf.contrib, i.e. TF 1.x).  Published algorithm
// (tensorflow/core/kernels/resize_bilinear_op.cc, align_corners = False, no half-pixel centres):
//     scale = in_size / (float)out_size
//     in = out_index * scale;  lower = (int)in;  upper = min(lower + 1, in_size - 1);  lerp = in - lower
//     top = tl + (tr - tl) * x_lerp;  bottom = bl + (br - bl) * x_lerp;  value = top + (bottom - top) * y_lerp
// every step one IEEE float32 operation (the library is compiled with --fmad=false), so that the value of a
// pixel is the same bits wherever it is formed: in the materialising kernel, in the E-step's producers and in
// the rare exact-label paths that look at a pixel again.
#pragma once
#include <cuda_runtime.h>

namespace wl {

struct UpTap {
    int o00, o01, o10, o11;   // offsets (in floats) of class 0 of the four taps inside the low-resolution image
    float lx, ly;
};

__device__ __forceinline__ UpTap up_tap(int y, int x, int h, int w, int C, float sy, float sx) {
    const float fy = __fmul_rn((float)y, sy), fx = __fmul_rn((float)x, sx);
    const int y0 = (int)fy, x0 = (int)fx;                     // fy, fx >= 0: truncation is floor
    const int y1 = min(y0 + 1, h - 1), x1 = min(x0 + 1, w - 1);
    UpTap t;
    t.ly = __fsub_rn(fy, (float)y0);
    t.lx = __fsub_rn(fx, (float)x0);
    t.o00 = (y0 * w + x0) * C; t.o01 = (y0 * w + x1) * C;
    t.o10 = (y1 * w + x0) * C; t.o11 = (y1 * w + x1) * C;
    return t;
}

// GLOBAL: src is read-only global memory (ld.global.nc); else any address space (a staged copy in shared memory)
template <bool GLOBAL = true>
__device__ __forceinline__ float up_value(const float *__restrict__ src, const UpTap &t, int cls) {
    float tl, tr, bl, br;
    if (GLOBAL) {
        tl = __ldg(src + t.o00 + cls); tr = __ldg(src + t.o01 + cls);
        bl = __ldg(src + t.o10 + cls); br = __ldg(src + t.o11 + cls);
    } else {
        tl = src[t.o00 + cls]; tr = src[t.o01 + cls];
        bl = src[t.o10 + cls]; br = src[t.o11 + cls];
    }
    const float top = __fadd_rn(tl, __fmul_rn(__fsub_rn(tr, tl), t.lx));
    const float bot = __fadd_rn(bl, __fmul_rn(__fsub_rn(br, bl), t.lx));
    return __fadd_rn(top, __fmul_rn(__fsub_rn(bot, top), t.ly));
}

}  // namespace wl
