// estep_device.cuh -- device helpers shared by the generic E-step kernel (estep_kernel.cu) and the
// on-chip fast path (estep_fast.cu): cluster barrier, cp.async staging of the NHWC stream,
// absent-class suppression + plane compaction (estep/weaklabel.cpp:45-65), block scans, label
// writer (weaklabel.cpp:108-112 + deeplab.py:123).  Everything is static-inline device code.
#pragma once
#include <float.h>
#include <math.h>
#include "estep_kernel.cuh"
#include "seqsum.cuh"

namespace wl {


// ---------------------------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void cluster_barrier(int G) {
    if (G > 1) {
        asm volatile("barrier.cluster.arrive.release.aligned;\n"
                     "barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
    } else {
        __syncthreads();
    }
}

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gsrc) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

struct SharedMisc {
    double wscratch_d[NWARPS];
    uint32_t wscratch_u[NWARPS];
    double bcast_d;
    uint32_t bcast_u;
    int sel_bin;
    uint32_t sel_pre, sel_cnt;
    uint8_t plist[MAXC];
};

struct Ctx {
    const EStepParams *p;
    int tid, lane, warp, rank, G;
    SlotCtl *ctl;
    float *planes;        // this cluster's scratch
    float *absmax;        // plane C of the scratch
    uint8_t *absidx;
    uint32_t *s_hist;
    SharedMisc *sm;
    float *s_dyn;         // phase-shared dynamic region
    uint32_t pass_no;     // rotating histogram index, identical in every CTA of the cluster
    // per image
    int b, P;
    uint32_t mask;
    int first_absent;
};

__device__ __forceinline__ uint32_t block_sum_u32(Ctx &c, uint32_t v) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (c.lane == 0) c.sm->wscratch_u[c.warp] = v;
    __syncthreads();
    if (c.warp == 0) {
        uint32_t t = c.sm->wscratch_u[c.lane];
        for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (c.lane == 0) c.sm->bcast_u = t;
    }
    __syncthreads();
    uint32_t r = c.sm->bcast_u;
    __syncthreads();
    return r;
}

// exclusive prefix over the block in thread order; *total receives the block sum
__device__ __forceinline__ uint32_t block_excl_scan_u32(Ctx &c, uint32_t v, uint32_t *total) {
    uint32_t inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (c.lane >= o) inc += t;
    }
    if (c.lane == 31) c.sm->wscratch_u[c.warp] = inc;
    __syncthreads();
    if (c.warp == 0) {
        uint32_t w = c.sm->wscratch_u[c.lane], winc = w;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
            if (c.lane >= o) winc += t;
        }
        c.sm->wscratch_u[c.lane] = winc - w;          // exclusive warp offsets
        if (c.lane == 31) c.sm->bcast_u = winc;
    }
    __syncthreads();
    uint32_t r = c.sm->wscratch_u[c.warp] + inc - v;
    *total = c.sm->bcast_u;
    __syncthreads();
    return r;
}

__device__ __forceinline__ double block_excl_scan_f64(Ctx &c, double v, double *total) {
    double inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        double t = __shfl_up_sync(0xffffffffu, inc, o);
        if (c.lane >= o) inc += t;
    }
    if (c.lane == 31) c.sm->wscratch_d[c.warp] = inc;
    __syncthreads();
    if (c.warp == 0) {
        double w = c.sm->wscratch_d[c.lane], winc = w;
        for (int o = 1; o < 32; o <<= 1) {
            double t = __shfl_up_sync(0xffffffffu, winc, o);
            if (c.lane >= o) winc += t;
        }
        c.sm->wscratch_d[c.lane] = winc - w;
        if (c.lane == 31) c.sm->bcast_d = winc;
    }
    __syncthreads();
    double r = c.sm->wscratch_d[c.warp] + (inc - v);
    *total = c.sm->bcast_d;
    __syncthreads();
    return r;
}

// ---------------------------------------------------------------------------------------------
// phase 1: stream-in
// ---------------------------------------------------------------------------------------------
// One pixel: suppression (weaklabel.cpp:45-65) + compaction.  px -> the pixel's C scores.
__device__ __forceinline__ void ingest_pixel(const Ctx &c, const float *px, int i) {
    const EStepParams &p = *c.p;
    const uint32_t mask = c.mask;
    float m = FLT_MAX;
    for (int k = 0; k < p.C; k++)
        if ((mask >> k) & 1u) { float t = px[k]; if (t < m) m = t; }
    float amax = -INFINITY;
    int aidx = 255, pp = 0;
    const float low = __fsub_rn(m, p.margin);
    for (int k = 0; k < p.C; k++) {
        float t = px[k];
        if ((mask >> k) & 1u) {
            c.planes[(size_t)pp * p.Npad + i] = t;
            pp++;
        } else {
            if (p.suppress && t > m) t = low;
            if (t > amax) { amax = t; aidx = k; }
        }
    }
    c.absmax[i] = amax;
    c.absidx[i] = (uint8_t)aidx;
}

// cp.async the flat float range [f0, f0+nfl) of the image into buf; returns the number of
// leading floats (0..3) that precede f0 inside the first 16-byte chunk.
__device__ __forceinline__ int prefetch_floats(const Ctx &c, float *buf, const float *img,
                                               long long f0, int nfl) {
    const float *g0 = img + f0;
    const int lead = (int)(((uintptr_t)g0 & 15u) >> 2);
    const char *a0 = (const char *)g0 - lead * 4;
    const int nchunk = (lead + nfl + 3) >> 2;
    for (int q = c.tid; q < nchunk; q += NT) {
        const int first = q * 4 - lead;
        const char *src = a0 + (size_t)q * 16;
        char *dst = (char *)buf + (size_t)q * 16;
        if (first >= 0 && first + 4 <= nfl) {
            cp_async16(dst, src);
        } else {
#pragma unroll
            for (int e = 0; e < 4; e++) {
                int fi = first + e;
                if (fi >= 0 && fi < nfl) cp_async4(dst + e * 4, src + e * 4);
            }
        }
    }
    return lead;
}

__device__ __forceinline__ void stream_in(Ctx &c) {
    const EStepParams &p = *c.p;
    const float *img = p.map + (long long)c.b * p.s0;
    const bool contig = (p.s2 == (long long)p.C);
    float *buf0 = c.s_dyn, *buf1 = c.s_dyn + STREAM_BUF_FLOATS;
    for (int tile = c.rank; tile < p.ntiles; tile += c.G) {
        const int pix0 = tile * p.TS;
        const int npx = min(p.TS, p.N - pix0);
        if (!contig) {      // padded pixel stride: gather straight from global memory
            for (int j = c.tid; j < npx; j += NT)
                ingest_pixel(c, img + (long long)(pix0 + j) * p.s2, pix0 + j);
            continue;
        }
        const int SUB = p.sub_px;
        const int nsub = (npx + SUB - 1) / SUB;
        int lead_cur = prefetch_floats(c, buf0, img, (long long)pix0 * p.C, min(SUB, npx) * p.C);
        cp_async_commit();
        for (int s = 0; s < nsub; s++) {
            float *cur = (s & 1) ? buf1 : buf0;
            float *nxt = (s & 1) ? buf0 : buf1;
            int lead_nxt = 0;
            if (s + 1 < nsub) {
                const int q0 = pix0 + (s + 1) * SUB;
                lead_nxt = prefetch_floats(c, nxt, img, (long long)q0 * p.C,
                                           min(SUB, npx - (s + 1) * SUB) * p.C);
            }
            cp_async_commit();
            cp_async_wait<1>();
            __syncthreads();
            const int n_here = min(SUB, npx - s * SUB);
            if (c.tid < n_here)
                ingest_pixel(c, cur + lead_cur + c.tid * p.C, pix0 + s * SUB + c.tid);
            __syncthreads();
            lead_cur = lead_nxt;
        }
        cp_async_wait<0>();
    }
}


template <typename T> struct FnShfl;
template <> struct FnShfl<float> {
    static __device__ __forceinline__ ChunkFn<float> down(const ChunkFn<float> &f, int d) {
        ChunkFn<float> g;
        g.e = __shfl_down_sync(0xffffffffu, f.e, d);
        g.a0 = __shfl_down_sync(0xffffffffu, f.a0, d);
        g.a1 = __shfl_down_sync(0xffffffffu, f.a1, d);
        g.kind = __shfl_down_sync(0xffffffffu, f.kind, d);
        return g;
    }
};
template <> struct FnShfl<double> {
    static __device__ __forceinline__ ChunkFn<double> down(const ChunkFn<double> &f, int d) {
        ChunkFn<double> g;
        g.e = (uint64_t)__shfl_down_sync(0xffffffffu, (unsigned long long)f.e, d);
        g.a0 = (int64_t)__shfl_down_sync(0xffffffffu, (long long)f.a0, d);
        g.a1 = (int64_t)__shfl_down_sync(0xffffffffu, (long long)f.a1, d);
        g.kind = __shfl_down_sync(0xffffffffu, f.kind, d);
        return g;
    }
};


// ---------------------------------------------------------------------------------------------
// phase 5: labels
// ---------------------------------------------------------------------------------------------
// Exact per-pixel recomputation from the original scores (rare: an absent class ties with or
// beats the best present class and is not provably the lowest-index winner).
static __device__ __noinline__ int label_slow(const EStepParams *pp_, const float *planes, uint32_t mask,
                                       int b, int i, double shift) {
    const EStepParams &p = *pp_;
    const float *px = p.map + (long long)b * p.s0 + (long long)i * p.s2;
    float m = FLT_MAX;
    for (int k = 0; k < p.C; k++)
        if ((mask >> k) & 1u) { float t = px[k]; if (t < m) m = t; }
    const float low = __fsub_rn(m, p.margin);
    float best = -INFINITY;
    int bi = 0, pp = 0;
    for (int k = 0; k < p.C; k++) {
        float v;
        if ((mask >> k) & 1u) { v = __ldcg(planes + (size_t)pp * p.Npad + i); pp++; }
        else { v = px[k]; if (p.suppress && v > m) v = low; }
        const float fv = __double2float_rn(__dadd_rn((double)v, shift));
        if (fv > best) { best = fv; bi = k; }
    }
    return bi;
}

__device__ __forceinline__ void write_labels(Ctx &c, double shift) {
    const EStepParams &p = *c.p;
    uint8_t *out = p.labels + (size_t)c.b * p.N;
    const bool has_absent = c.P < p.C;
    for (int tile = c.rank; tile < p.ntiles; tile += c.G) {
        const int base = tile * p.TS;
        const int n_t = min(p.TS, p.N - base);
        const int nq = (n_t + 3) >> 2;
        for (int q = c.tid; q < nq; q += NT) {
            const int idx = base + 4 * q;
            float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
            int bi[4] = {0, 0, 0, 0};
            for (int pl = 0; pl < c.P; pl++) {
                const float4 v = __ldcg(reinterpret_cast<const float4 *>(c.planes + (size_t)pl * p.Npad + idx));
                const float vv[4] = {v.x, v.y, v.z, v.w};
                const int cls = c.sm->plist[pl];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float fv = __double2float_rn(__dadd_rn((double)vv[e], shift));
                    if (fv > best[e]) { best[e] = fv; bi[e] = cls; }
                }
            }
            if (has_absent) {
                const float4 a = __ldcg(reinterpret_cast<const float4 *>(c.absmax + idx));
                const uchar4 ai = __ldcg(reinterpret_cast<const uchar4 *>(c.absidx + idx));
                const float av[4] = {a.x, a.y, a.z, a.w};
                const int aidx[4] = {ai.x, ai.y, ai.z, ai.w};
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    if (idx + e >= p.N) continue;
                    const float fa = __double2float_rn(__dadd_rn((double)av[e], shift));
                    if (fa >= best[e]) {
                        if (aidx[e] == c.first_absent)
                            bi[e] = (fa > best[e]) ? aidx[e] : min(aidx[e], bi[e]);
                        else
                            bi[e] = label_slow(c.p, c.planes, c.mask, c.b, idx + e, shift);
                    }
                }
            }
#pragma unroll
            for (int e = 0; e < 4; e++)
                if (idx + e < p.N) out[idx + e] = (uint8_t)bi[e];
        }
    }
}


}  // namespace wl
