// estep_kernel.cu -- the EM-Adapt E-step as ONE persistent thread-block-cluster kernel for
// sm_100a (B200).  One cluster works on one image at a time:
//
//   1. stream-in   : the image's NHWC scores are read from HBM exactly once (cp.async, 16-byte
//                    chunks of the flat float stream, staged in shared memory), the absent
//                    classes are suppressed (estep/weaklabel.cpp:45-65) and the P present classes
//                    are written as compact class-major planes into a per-cluster scratch that is
//                    small enough to stay in L2 (plus the per-pixel max of the absent classes).
//   2. mean_max_0  : order-exact FLOAT running sum of the floored per-pixel max
//                    (weaklabel.cpp:67-80) -- see seqsum.cuh.
//   3. iterations  : num_iter x P exact order statistics (the value std::nth_element leaves at
//                    diff[nth], weaklabel.cpp:85-99) by a bracketed radix select on the float bit
//                    pattern: shared-memory histograms per CTA, merged across the cluster through
//                    L2 atomics, one cluster barrier per pass.  The bias of step t is folded into
//                    the class plane by the first pass of step t+1.
//   4. mean_max_1  : order-exact DOUBLE running sum of the final per-pixel max (:102-106).
//   5. labels      : re-centre by (mean_max_0 - mean_max_1) as the reference does (float += double,
//                    :108-112) and take the arg-max over all C classes, lowest index on ties
//                    (deeplab.py:123); absent classes are represented by their maximum and only
//                    pixels where that could matter re-read the original scores.
//
// Nothing here is a dense contraction: no tensor cores.  HBM traffic is 4*C bytes read + 1 byte
// written per pixel; everything else is L2 / shared memory.
#include <float.h>
#include <math.h>
#include "estep_device.cuh"

namespace wl {

// ---------------------------------------------------------------------------------------------
// per-pixel state access shared by the select passes and the sums
// ---------------------------------------------------------------------------------------------
struct Pending {      // bias of the previous step, not yet folded into its plane
    int plane;        // -1: none
    float bias;
};

// Loads the 4 pixels idx..idx+3 of every present plane, folds the pending bias (writing the
// plane back when asked), and returns the floored max E (weaklabel.cpp:71-77, 94-99) and the
// values of plane `want` (or zeros when want < 0).
__device__ __forceinline__ void load_quad(const Ctx &c, int idx, const Pending &pend, bool write_back,
                                          int want, float4 &E, float4 &xc) {
    const EStepParams &p = *c.p;
    E = make_float4(FLT_MIN, FLT_MIN, FLT_MIN, FLT_MIN);
    xc = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((!p.suppress || p.margin < 0.f) && c.P < p.C) {   // with suppression and a margin >= 0 the absent
                                                         // classes never exceed the present ones
        float4 a = *reinterpret_cast<const float4 *>(c.absmax + idx);
        if (a.x > E.x) E.x = a.x;
        if (a.y > E.y) E.y = a.y;
        if (a.z > E.z) E.z = a.z;
        if (a.w > E.w) E.w = a.w;
    }
    for (int pl = 0; pl < c.P; pl++) {
        float4 *ptr = reinterpret_cast<float4 *>(c.planes + (size_t)pl * p.Npad + idx);
        float4 v = *ptr;
        if (pl == pend.plane) {
            v.x = __fadd_rn(v.x, pend.bias); v.y = __fadd_rn(v.y, pend.bias);
            v.z = __fadd_rn(v.z, pend.bias); v.w = __fadd_rn(v.w, pend.bias);
            if (write_back) *ptr = v;
        }
        if (v.x > E.x) E.x = v.x;
        if (v.y > E.y) E.y = v.y;
        if (v.z > E.z) E.z = v.z;
        if (v.w > E.w) E.w = v.w;
        if (pl == want) xc = v;
    }
}

// ---------------------------------------------------------------------------------------------
// phases 2 and 4: order-exact running sums of the floored per-pixel max
// ---------------------------------------------------------------------------------------------
// Sum over all pixels of the image, in pixel order, of max(FLT_MIN, max_c score) with the
// rounding of a sequential accumulation in T.  Folds `pend` into its plane on the way.
// Every CTA of the cluster returns the same value.
template <typename T>
__device__ __forceinline__ T exact_pixel_max_sum(Ctx &c, Pending pend, uint32_t &tot_parity) {
    typedef fp_traits<T> F;
    const EStepParams &p = *c.p;
    float *s_terms = c.s_dyn;                                             // [TS]
    ChunkFn<T> *s_fn = reinterpret_cast<ChunkFn<T> *>(c.s_dyn + TS_MAX);   // [NT]
    ChunkFn<T> *s_wfn = s_fn + NT;                                         // [NWARPS]
    const int nrounds = (p.ntiles + c.G - 1) / c.G;
    double carry_real = 0.0;
    for (int round = 0; round < nrounds; round++) {
        const int tile = round * c.G + c.rank;
        const int base = tile * p.TS;
        const int n_t = tile < p.ntiles ? min(p.TS, p.N - base) : 0;
        // a. terms of my tile -> shared memory
        const int nq = (n_t + 3) >> 2;
        for (int q = c.tid; q < nq; q += NT) {
            float4 E, xc;
            load_quad(c, base + 4 * q, pend, true, -1, E, xc);
            *reinterpret_cast<float4 *>(s_terms + 4 * q) = E;
        }
        __syncthreads();
        // b. ordinary (double) sums of the thread chunks, prefix over the tile
        const int c0 = c.tid * p.L;
        const int len = max(0, min(p.L, n_t - c0));
        double cs = 0.0;
        for (int i = 0; i < len; i++) cs += (double)s_terms[c0 + i];
        double tile_total;
        const double cpre = block_excl_scan_f64(c, cs, &tile_total);
        // c. exchange tile totals across the cluster
        double *tot = c.ctl->tot[tot_parity & 1u];
        if (c.tid == 0) tot[c.rank] = tile_total;
        cluster_barrier(c.G);
        double before = carry_real, all = carry_real;
        for (int r = 0; r < c.G; r++) {
            double t = __ldcg(tot + r);
            if (r < c.rank) before += t;
            all += t;
        }
        // d. chunk summaries on the predicted binade; e. in-order composition per warp
        ChunkFn<T> f = chunk_eval<T>(s_terms + c0, len, before + cpre);
        ChunkFn<T> w = f;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            ChunkFn<T> o = FnShfl<T>::down(w, d);
            if ((c.lane & (2 * d - 1)) == 0) w = chunk_compose(w, o);
        }
        s_fn[c.tid] = f;
        if (c.lane == 0) s_wfn[c.warp] = w;
        __syncthreads();
        // f. the walk: CTA after CTA, one thread, verifying every summary it uses
        for (int r = 0; r < c.G; r++) {
            if (c.rank == r && c.tid == 0) {
                T s = (round == 0 && r == 0) ? (T)0
                                             : F::from((typename F::U)__ldcg(&c.ctl->carry));
                for (int wv = 0; wv < NWARPS; wv++) {
                    if (chunk_apply(s_wfn[wv], s)) continue;
                    for (int l = 0; l < 32; l++) {
                        const int ch = wv * 32 + l;
                        if (chunk_apply(s_fn[ch], s)) continue;
                        const int cc0 = ch * p.L;
                        s = chunk_sequential<T>(s_terms + cc0, max(0, min(p.L, n_t - cc0)), s);
                    }
                }
                c.ctl->carry = (unsigned long long)F::bits(s);
            }
            cluster_barrier(c.G);
        }
        carry_real = all;
        tot_parity++;
    }
    T res = (p.ntiles == 0) ? (T)0 : F::from((typename F::U)__ldcg(&c.ctl->carry));
    __syncthreads();
    return res;
}

// ---------------------------------------------------------------------------------------------
// phase 3: exact order statistic of the margin of one class (weaklabel.cpp:85-89)
// ---------------------------------------------------------------------------------------------
// Returns the float whose rank among diff[i] = E[i] - x_plane[i] (i < N) is k (0-based).
__device__ __forceinline__ float select_margin(Ctx &c, int plane, int k, Pending &pend) {
    const EStepParams &p = *c.p;
    uint32_t lo = 1u, hi = 0x7f7fffffu;      // zeros are counted as "below" in the first pass
    uint32_t n_below = 0;
    bool first = true;
    float answer = 0.f;
    for (int guard = 0; guard < 8; guard++) {
        int sh = 0;
        while (((hi - lo) >> sh) >= (uint32_t)HBINS) sh++;
        const uint32_t buf = c.pass_no % 3u, nbuf = (c.pass_no + 1u) % 3u;
        uint32_t below = 0;
        for (int tile = c.rank; tile < p.ntiles; tile += c.G) {
            const int base = tile * p.TS;
            const int n_t = min(p.TS, p.N - base);
            const int nq = (n_t + 3) >> 2;
            for (int q = c.tid; q < nq; q += NT) {
                const int idx = base + 4 * q;
                float4 E, xc;
                load_quad(c, idx, pend, true, plane, E, xc);
                const float d[4] = {__fsub_rn(E.x, xc.x), __fsub_rn(E.y, xc.y),
                                    __fsub_rn(E.z, xc.z), __fsub_rn(E.w, xc.w)};
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    if (idx + e < p.N) {
                        const uint32_t key = __float_as_uint(d[e]);
                        if (key < lo) below++;
                        else if (key <= hi) atomicAdd(&c.s_hist[(key - lo) >> sh], 1u);
                    }
                }
            }
        }
        pend.plane = -1;                       // folded into its plane by this pass
        __syncthreads();
        // merge through L2, recycle the next buffer
        for (int bin = c.tid; bin < HBINS; bin += NT) {
            const uint32_t h = c.s_hist[bin];
            if (h) { atomicAdd(&c.ctl->hist[buf][bin], h); c.s_hist[bin] = 0; }
        }
        if (first) {
            const uint32_t bt = block_sum_u32(c, below);
            if (c.tid == 0 && bt) atomicAdd(&c.ctl->below[buf][0], bt);
        }
        {
            const int per = HBINS / c.G;
            for (int bin = c.rank * per + c.tid; bin < (c.rank + 1) * per; bin += NT)
                c.ctl->hist[nbuf][bin] = 0;
            if (c.rank == 0 && c.tid == 0) c.ctl->below[nbuf][0] = 0;
        }
        c.pass_no++;
        cluster_barrier(c.G);
        if (first) {
            n_below = __ldcg(&c.ctl->below[buf][0]);
            first = false;
            if ((uint32_t)k < n_below) { answer = 0.f; break; }   // rank falls among the zero margins
        }
        // locate the bin holding rank k
        const uint32_t h0 = __ldcg(&c.ctl->hist[buf][2 * c.tid]);
        const uint32_t h1 = __ldcg(&c.ctl->hist[buf][2 * c.tid + 1]);
        if (c.tid == 0) c.sm->sel_bin = -1;
        uint32_t total;
        const uint32_t ex = block_excl_scan_u32(c, h0 + h1, &total);
        const uint32_t r = (uint32_t)k - n_below;
        if (r >= ex && r < ex + h0) { c.sm->sel_bin = 2 * c.tid; c.sm->sel_pre = ex; c.sm->sel_cnt = h0; }
        else if (r >= ex + h0 && r < ex + h0 + h1) {
            c.sm->sel_bin = 2 * c.tid + 1; c.sm->sel_pre = ex + h0; c.sm->sel_cnt = h1;
        }
        __syncthreads();
        const int bin = c.sm->sel_bin;
        const uint32_t pre = c.sm->sel_pre;
        __syncthreads();
        if (bin < 0) {                         // cannot happen for NaN-free input
            if (c.tid == 0) atomicExch(c.p->status, 100 + guard);
            answer = 0.f;
            break;
        }
        n_below += pre;
        lo = lo + ((uint32_t)bin << sh);
        const uint32_t top = lo + ((1u << sh) - 1u);
        if (top < hi) hi = top;
        if (lo == hi) { answer = __uint_as_float(lo); break; }
    }
    return answer;
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
constexpr size_t SMEM_FIXED = HBINS * sizeof(uint32_t) + ((sizeof(SharedMisc) + 15) / 16) * 16;
constexpr size_t SMEM_STREAM = 2 * (size_t)STREAM_BUF_FLOATS * sizeof(float);
constexpr size_t SMEM_SUM = (size_t)TS_MAX * sizeof(float) + (NT + NWARPS) * sizeof(ChunkFn<double>);
constexpr size_t SMEM_DYN = SMEM_STREAM > SMEM_SUM ? SMEM_STREAM : SMEM_SUM;

size_t estep_smem_bytes() { return SMEM_FIXED + SMEM_DYN; }

__global__ void __launch_bounds__(NT, 1) estep_cluster_kernel(const __grid_constant__ EStepParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Ctx c;
    c.p = &p;
    c.tid = threadIdx.x;
    c.lane = c.tid & 31;
    c.warp = c.tid >> 5;
    c.G = p.G;
    c.rank = blockIdx.x % p.G;
    const int slot = blockIdx.x / p.G;
    c.ctl = p.ctl + slot;
    c.planes = p.planes + (size_t)slot * (size_t)(p.C + 1) * p.Npad;
    c.absmax = c.planes + (size_t)p.C * p.Npad;
    c.absidx = p.absent_idx + (size_t)slot * p.Npad;
    c.s_hist = reinterpret_cast<uint32_t *>(smem_raw);
    c.sm = reinterpret_cast<SharedMisc *>(smem_raw + HBINS * sizeof(uint32_t));
    c.s_dyn = reinterpret_cast<float *>(smem_raw + SMEM_FIXED);
    c.pass_no = 0;
    uint32_t tot_parity = 0;
    for (int i = c.tid; i < HBINS; i += NT) c.s_hist[i] = 0;
    __syncthreads();

    for (;;) {
        if (c.rank == 0 && c.tid == 0) {
            const int t = atomicAdd(p.image_counter, 1);
            c.ctl->image = (t < p.B && p.sched) ? p.sched[t] : t;
        }
        cluster_barrier(c.G);
        c.b = __ldcg(&c.ctl->image);
        if (c.b >= p.B) break;
        const ImageMeta *im = p.meta + c.b;
        c.mask = im->mask;
        c.P = im->P;
        c.first_absent = im->first_absent;
        if (c.tid < MAXC) c.sm->plist[c.tid] = im->plist[c.tid];
        __syncthreads();

        stream_in(c);
        cluster_barrier(c.G);       // planes complete (tile ownership is the same in every phase,
                                    // the barrier orders them for the sums' walk anyway)
        Pending pend;
        pend.plane = -1; pend.bias = 0.f;
        const float sum0 = exact_pixel_max_sum<float>(c, pend, tot_parity);
        const float mean0 = __fdiv_rn(sum0, (float)p.N);                    // weaklabel.cpp:80

        const uint8_t *ord = p.order + (size_t)c.b * p.num_iter * p.C;
        for (int j = 0; j < p.num_iter; j++) {
            for (int t = 0; t < c.P; t++) {
                const int plane = ord[j * p.C + t];
                const int cls = c.sm->plist[plane];
                const int k = (cls == 0) ? p.nth_bg : p.nth_fg;             // weaklabel.cpp:88
                const float bias = select_margin(c, plane, k, pend);
                pend.plane = plane;
                pend.bias = bias;
                if (c.rank == 0 && c.tid == 0)
                    p.bias_out[((size_t)c.b * p.num_iter + j) * p.C + t] = bias;
            }
        }
        const double sum1 = exact_pixel_max_sum<double>(c, pend, tot_parity);   // folds the last bias
        const double mean1 = __ddiv_rn(sum1, (double)p.N);                  // weaklabel.cpp:106
        const double shift = __dsub_rn((double)mean0, mean1);               // weaklabel.cpp:110
        if (c.rank == 0 && c.tid == 0) {
            p.mean0[c.b] = mean0;
            p.mean1[c.b] = mean1;
            p.shift[c.b] = shift;
        }
        if (p.labels) write_labels(c, shift);
    }
}

cudaError_t launch_estep(const EStepParams &p, int n_clusters, cudaStream_t stream) {
    static std::atomic<unsigned long long> attr_done{0};
    const size_t smem = estep_smem_bytes();
    cudaError_t e;
    if (first_launch_on_device(attr_done)) {
        e = cudaFuncSetAttribute(estep_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(estep_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(n_clusters * p.G));
    cfg.blockDim = dim3(NT);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.G;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, estep_cluster_kernel, p);
}

int estep_max_clusters(int G) {
    const size_t smem = estep_smem_bytes();
    if (cudaFuncSetAttribute(estep_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
    cudaFuncSetAttribute(estep_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(G * 64));
    cfg.blockDim = dim3(NT);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)G;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, estep_cluster_kernel, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// ---------------------------------------------------------------------------------------------
// the reference's in-place contract: rewrite the score map (weaklabel.cpp:45-65, 90-92, 108-112)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) finalize_map_kernel(const __grid_constant__ FinalizeParams p) {
    __shared__ float cb[MAXC * 16];     // per class, the biases of iterations [j0, j0+16)
    const int b = blockIdx.y;
    const ImageMeta *im = p.meta + b;
    const uint32_t mask = im->mask;
    const double shift = p.shift[b];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float v[MAXC];
    float *px = p.map + (long long)b * p.s0 + (long long)i * p.s2;
    float m = FLT_MAX;
    if (i < p.N) {
#pragma unroll
        for (int k = 0; k < MAXC; k++)
            if (k < p.C) { v[k] = px[k]; if (((mask >> k) & 1u) && v[k] < m) m = v[k]; }
        const float low = __fsub_rn(m, p.margin);
#pragma unroll
        for (int k = 0; k < MAXC; k++)
            if (k < p.C && !((mask >> k) & 1u) && p.suppress && v[k] > m) v[k] = low;
    }
    for (int j0 = 0; j0 < p.num_iter; j0 += 16) {
        const int nj = min(16, p.num_iter - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < nj * im->P; t += blockDim.x) {
            const int j = j0 + t / im->P, pos = t % im->P;
            const size_t o = ((size_t)b * p.num_iter + j) * p.C + pos;
            cb[im->plist[p.order[o]] * 16 + (j - j0)] = p.bias[o];
        }
        __syncthreads();
        if (i < p.N) {
#pragma unroll
            for (int k = 0; k < MAXC; k++)
                if (k < p.C && ((mask >> k) & 1u))
                    for (int j = 0; j < nj; j++) v[k] = __fadd_rn(v[k], cb[k * 16 + j]);
        }
    }
    if (i < p.N) {
#pragma unroll
        for (int k = 0; k < MAXC; k++)
            if (k < p.C) px[k] = __double2float_rn(__dadd_rn((double)v[k], shift));
    }
}

cudaError_t launch_finalize_map(const FinalizeParams &p, cudaStream_t stream) {
    dim3 grid((unsigned)((p.N + 255) / 256), (unsigned)p.B);
    finalize_map_kernel<<<grid, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace wl
