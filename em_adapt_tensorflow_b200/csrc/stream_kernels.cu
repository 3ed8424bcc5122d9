// stream_kernels.cu -- the two pure streaming kernels around the E-step (sm_100a), both on the bulk-tile
// pattern of softmax_ce.cu (a tile of 256 whole pixels moves global -> shared -> global as ONE bulk copy each
// way; one thread per pixel works on its 84-byte row in shared memory):
//
//   finalize_tile_kernel   the reference's in-place contract (weaklabel.cpp:45-65, 90-92, 108-112): the score
//                          map becomes the suppressed, bias-shifted, re-centred map.  168 B/pixel.
//   fixed_bias_tile_kernel EM-Fixed (Papandreou et al. 2015, the paper readme.md:3 cites; NOT implemented in the
//                          reference): labels = argmax over the PRESENT classes of score + a fixed bias
//                          (one for the background, one for foreground classes), optionally overridden by a
//                          strong (pixel-level) label where one is given.  85 B/pixel.
#include <float.h>
#include <math.h>
#include "estep_kernel.cuh"
#include "bulk_tile.cuh"
#include "upsample.cuh"

namespace wl {

constexpr int ST_NT = 256;          // threads per CTA = pixels per tile
constexpr int FB_J = 16;            // iterations of bias history cached per round

// Tiles are cut from the FLAT pixel sequence of the contiguous batch, so a tile may straddle two images
// (never more: N >= ST_NT is required); tables are cached for both.
__global__ void __launch_bounds__(ST_NT) finalize_tile_kernel(const __grid_constant__ FinalizeParams p, long long n_pix, int bulk_ok) {
    extern __shared__ __align__(128) unsigned char st_smem[];
    __shared__ uint64_t full_bar[2];
    __shared__ float cb[2][MAXC * FB_J];     // [image of the tile][class][iteration]: the biases in visit order
    const int tid = threadIdx.x, C = p.C;
    const size_t tile_bytes = (size_t)ST_NT * C * sizeof(float);
    const long long ntiles = (n_pix + ST_NT - 1) / ST_NT;
    if (tid == 0) {
        ce_mbar_init(&full_bar[0], 1);
        ce_mbar_init(&full_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();
    auto is_bulk = [&](long long t) { return bulk_ok && (t + 1) * ST_NT <= n_pix; };
    auto issue = [&](long long t, int s) {
        if (is_bulk(t)) {
            ce_mbar_expect_tx(&full_bar[s], (uint32_t)tile_bytes);
            ce_bulk_g2s(st_smem + (size_t)s * tile_bytes, p.map + t * ST_NT * C, (uint32_t)tile_bytes, &full_bar[s]);
        }
    };
    long long t = blockIdx.x;
    if (tid == 0 && t < ntiles) issue(t, 0);
    uint32_t it = 0;
    for (; t < ntiles; t += gridDim.x, it++) {
        const int s = it & 1;
        float *buf = reinterpret_cast<float *>(st_smem + (size_t)s * tile_bytes);
        const long long tn = t + gridDim.x;
        if (tid == 0 && tn < ntiles) {
            asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");      // the other buffer's store has read it
            issue(tn, s ^ 1);
        }
        const long long pix0 = t * ST_NT;
        const int npx = (int)min((long long)ST_NT, n_pix - pix0);
        const bool bulk = is_bulk(t);
        if (bulk) {
            ce_mbar_wait(&full_bar[s], (it >> 1) & 1u);
        } else {
            if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
            __syncthreads();
            const float *g = p.map + pix0 * C;
            for (int f = tid; f < npx * C; f += ST_NT) buf[f] = g[f];
            __syncthreads();
        }
        const int b0 = (int)(pix0 / p.N), b1 = (int)((pix0 + npx - 1) / p.N);
        const long long gp = pix0 + tid;
        const int b = (int)(gp / p.N), w = b - b0;
        const bool live = tid < npx;
        float *z = buf + tid * C;
        uint32_t mask = 0u;
        if (live) {
            mask = p.meta[b].mask;
            // suppression of the absent classes (weaklabel.cpp:45-65)
            if (p.suppress) {
                float m = FLT_MAX;
                for (int k = 0; k < C; k++)
                    if (((mask >> k) & 1u) && z[k] < m) m = z[k];
                const float low = __fsub_rn(m, p.margin);
                for (int k = 0; k < C; k++)
                    if (!((mask >> k) & 1u) && z[k] > m) z[k] = low;
            }
        }
        // the biases of every present class in visit order (weaklabel.cpp:90-92), FB_J iterations at a time
        for (int j0 = 0; j0 < p.num_iter; j0 += FB_J) {
            const int nj = min(FB_J, p.num_iter - j0);
            __syncthreads();
            for (int wi = 0; wi <= b1 - b0; wi++) {
                const ImageMeta *im = p.meta + b0 + wi;
                for (int e = tid; e < nj * im->P; e += ST_NT) {
                    const int j = j0 + e / im->P, pos = e % im->P;
                    const size_t o = ((size_t)(b0 + wi) * p.num_iter + j) * C + pos;
                    cb[wi][im->plist[p.order[o]] * FB_J + (j - j0)] = p.bias[o];
                }
            }
            __syncthreads();
            if (live)
                for (int k = 0; k < C; k++)
                    if ((mask >> k) & 1u) {
                        float v = z[k];
                        for (int j = 0; j < nj; j++) v = __fadd_rn(v, cb[w][k * FB_J + j]);
                        z[k] = v;
                    }
        }
        if (live) {                                                           // re-centring (weaklabel.cpp:108-112)
            const double shift = p.shift[b];
            for (int k = 0; k < C; k++) z[k] = __double2float_rn(__dadd_rn((double)z[k], shift));
        }
        if (bulk) {
            asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
            __syncthreads();
            if (tid == 0) ce_bulk_s2g(p.map + pix0 * C, buf, (uint32_t)tile_bytes);
        } else {
            __syncthreads();
            float *o = p.map + pix0 * C;
            for (int f = tid; f < npx * C; f += ST_NT) o[f] = buf[f];
            __syncthreads();
        }
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");
}

static int stream_grid(size_t smem, long long ntiles, int sms) {
    int per_sm = (int)(220 * 1024 / (smem + 2048));
    if (per_sm > 8) per_sm = 8;
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sms * per_sm;
    if (grid > ntiles) grid = ntiles;
    return (int)(grid < 1 ? 1 : grid);
}

// contiguous batches of images with at least ST_NT pixels only; the caller falls back to finalize_map_kernel
bool finalize_tiles_applicable(const FinalizeParams &p) {
    return p.s2 == p.C && (p.B == 1 || p.s0 == (long long)p.N * p.C) && p.N >= ST_NT;
}

cudaError_t launch_finalize_tiles(const FinalizeParams &p, int sms, cudaStream_t stream) {
    const long long n_pix = (long long)p.B * p.N;
    const size_t smem = 2 * (size_t)ST_NT * p.C * sizeof(float);
    static std::atomic<unsigned long long> attr_set{0};
    if (first_launch_on_device(attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(finalize_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * ST_NT * MAXC * sizeof(float)));
        if (e != cudaSuccess) return e;
    }
    const long long ntiles = (n_pix + ST_NT - 1) / ST_NT;
    const int bulk_ok = (((uintptr_t)p.map) & 15u) == 0 ? 1 : 0;
    finalize_tile_kernel<<<stream_grid(smem, ntiles, sms), ST_NT, smem, stream>>>(p, n_pix, bulk_ok);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// EM-Fixed labels
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ST_NT) fixed_bias_tile_kernel(const float *__restrict__ map, long long n_pix, int N, int C,
                                                                const uint32_t *__restrict__ masks, float bias_bg, float bias_fg,
                                                                const uint8_t *__restrict__ strong, uint8_t *__restrict__ labels, int bulk_ok) {
    extern __shared__ __align__(128) unsigned char st_smem[];
    __shared__ uint64_t full_bar[2];
    const int tid = threadIdx.x;
    const size_t tile_bytes = (size_t)ST_NT * C * sizeof(float);
    const long long ntiles = (n_pix + ST_NT - 1) / ST_NT;
    if (tid == 0) {
        ce_mbar_init(&full_bar[0], 1);
        ce_mbar_init(&full_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();
    auto is_bulk = [&](long long t) { return bulk_ok && (t + 1) * ST_NT <= n_pix; };
    auto issue = [&](long long t, int s) {
        if (is_bulk(t)) {
            ce_mbar_expect_tx(&full_bar[s], (uint32_t)tile_bytes);
            ce_bulk_g2s(st_smem + (size_t)s * tile_bytes, map + t * ST_NT * C, (uint32_t)tile_bytes, &full_bar[s]);
        }
    };
    long long t = blockIdx.x;
    if (tid == 0 && t < ntiles) issue(t, 0);
    uint32_t it = 0;
    for (; t < ntiles; t += gridDim.x, it++) {
        const int s = it & 1;
        float *buf = reinterpret_cast<float *>(st_smem + (size_t)s * tile_bytes);
        const long long tn = t + gridDim.x;
        __syncthreads();                                  // every thread is done with the other buffer (previous tile)
        if (tid == 0 && tn < ntiles) issue(tn, s ^ 1);
        const long long pix0 = t * ST_NT;
        const int npx = (int)min((long long)ST_NT, n_pix - pix0);
        if (is_bulk(t)) {
            ce_mbar_wait(&full_bar[s], (it >> 1) & 1u);
        } else {
            const float *g = map + pix0 * C;
            for (int f = tid; f < npx * C; f += ST_NT) buf[f] = g[f];
            __syncthreads();
        }
        if (tid < npx) {
            const long long gp = pix0 + tid;
            const uint32_t mask = masks[gp / N];
            const float *z = buf + tid * C;
            float best = -INFINITY;
            int bi = 255;                                 // no present class: ignore label
            for (int k = 0; k < C; k++)
                if ((mask >> k) & 1u) {
                    const float v = __fadd_rn(z[k], k == 0 ? bias_bg : bias_fg);
                    if (v > best || bi == 255) { best = v; bi = k; }
                }
            if (strong) { const int g = strong[gp]; if (g < C) bi = g; }     // pixel-level annotation where there is one
            labels[gp] = (uint8_t)bi;
        }
    }
}

cudaError_t launch_fixed_bias(const float *map, long long n_pix, int N, int C, const uint32_t *masks, float bias_bg, float bias_fg,
                              const uint8_t *strong, uint8_t *labels, int sms, cudaStream_t stream) {
    const size_t smem = 2 * (size_t)ST_NT * C * sizeof(float);
    static std::atomic<unsigned long long> attr_set{0};
    if (first_launch_on_device(attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(fixed_bias_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * ST_NT * MAXC * sizeof(float)));
        if (e != cudaSuccess) return e;
    }
    const long long ntiles = (n_pix + ST_NT - 1) / ST_NT;
    const int bulk_ok = (((uintptr_t)map) & 15u) == 0 ? 1 : 0;
    fixed_bias_tile_kernel<<<stream_grid(smem, ntiles, sms), ST_NT, smem, stream>>>(map, n_pix, N, C, masks, bias_bg, bias_fg, strong,
                                                                                   labels, bulk_ok);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// tf.image.resize_bilinear (network.py:40), materialised: dst [B][H][W][C] from src [B][h][w][C].
// A CTA owns a run of consecutive output rows.  The two low-resolution rows an output row interpolates between are
// staged in shared memory (and kept while the following output rows map to the same pair: H / h of them do); one
// thread per output FLOAT of the row, so the 4 * C-byte pixel rows leave fully coalesced.  Write-bound:
// 4 * C bytes per output pixel.  STAGED = false (a low-resolution row pair beyond shared memory): taps from global.
// ---------------------------------------------------------------------------------------------
constexpr int UP_NT = 256;
template <bool STAGED>
__global__ void __launch_bounds__(UP_NT) upsample_bilinear_kernel(const float *__restrict__ src, int h, int w, int C, int H, int W,
                                                                  float sy, float sx, float *__restrict__ dst, int rows, int rows_per_cta,
                                                                  unsigned int inv_c) {
    extern __shared__ __align__(16) float up_rows[];         // [2][w * C]
    const int wc = w * C, WC = W * C;
    const int r_begin = blockIdx.x * rows_per_cta, r_end = min(rows, r_begin + rows_per_cta);
    int have_b = -1, have_y0 = -1;
    for (int r = r_begin; r < r_end; r++) {
        const int b = r / H, y = r - b * H;
        const float fy = __fmul_rn((float)y, sy);
        const int y0 = (int)fy, y1 = min(y0 + 1, h - 1);
        const float ly = __fsub_rn(fy, (float)y0);
        const float *s0 = src + ((long long)b * h + y0) * wc, *s1 = src + ((long long)b * h + y1) * wc;
        if (STAGED && (b != have_b || y0 != have_y0)) {       // block-uniform
            __syncthreads();                                   // the readers of the previous pair are done
            for (int f = threadIdx.x; f < wc; f += UP_NT) { up_rows[f] = __ldg(s0 + f); up_rows[wc + f] = __ldg(s1 + f); }
            __syncthreads();
            have_b = b; have_y0 = y0;
        }
        float *drow = dst + (long long)r * WC;
        for (int f = threadIdx.x; f < WC; f += UP_NT) {
            const int x = (int)__umulhi((unsigned int)f, inv_c);      // f / C (exact for f * (C - 1) < 2^32)
            const int cls = f - x * C;
            const float fx = __fmul_rn((float)x, sx);
            const int x0 = (int)fx, x1 = min(x0 + 1, w - 1);
            UpTap t;
            t.lx = __fsub_rn(fx, (float)x0);
            t.ly = ly;
            if (STAGED) {
                t.o00 = x0 * C; t.o01 = x1 * C; t.o10 = wc + x0 * C; t.o11 = wc + x1 * C;
                drow[f] = up_value<false>(up_rows, t, cls);
            } else {
                t.o00 = x0 * C; t.o01 = x1 * C;
                t.o10 = (int)(s1 - s0) + x0 * C; t.o11 = (int)(s1 - s0) + x1 * C;
                drow[f] = up_value<true>(s0, t, cls);
            }
        }
    }
}

cudaError_t launch_upsample_bilinear(const float *src, int B, int h, int w, int C, int H, int W, float *dst, cudaStream_t stream) {
    const long long rows_ll = (long long)B * H;
    if (rows_ll == 0 || W == 0) return cudaSuccess;
    if (rows_ll >= (1ll << 31)) return cudaErrorInvalidValue;
    const int rows = (int)rows_ll;
    // enough CTAs to fill the GPU, few enough that a CTA keeps its staged row pair for several output rows
    int ctas = 148 * 8;
    if (ctas > rows) ctas = rows;
    const int rpc = (rows + ctas - 1) / ctas;
    ctas = (rows + rpc - 1) / rpc;
    const unsigned int inv_c = (unsigned int)(((1ull << 32) + (unsigned long long)C - 1ull) / (unsigned long long)C);
    const size_t smem = (size_t)2 * w * C * sizeof(float);
    const float sy = (float)h / (float)H, sx = (float)w / (float)W;
    if (smem <= 200 * 1024) {
        static std::atomic<unsigned long long> attr_set{0};
        if (first_launch_on_device(attr_set)) {
            cudaError_t e = cudaFuncSetAttribute(upsample_bilinear_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e != cudaSuccess) return e;
        }
        upsample_bilinear_kernel<true><<<ctas, UP_NT, smem, stream>>>(src, h, w, C, H, W, sy, sx, dst, rows, rpc, inv_c);
    } else {
        upsample_bilinear_kernel<false><<<ctas, UP_NT, 0, stream>>>(src, h, w, C, H, W, sy, sx, dst, rows, rpc, inv_c);
    }
    return cudaGetLastError();
}

}  // namespace wl
