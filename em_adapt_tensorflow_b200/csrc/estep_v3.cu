// estep_v3.cu -- the E-step kernel of the common configuration, third generation (sm_100a).
//
// Contract: one thread-block cluster per image, persistent, images claimed from a counter
// (suppress_others = true with a margin >= 0, contiguous pixels, image fits the cluster).
//
// Warp-specialised: every CTA runs NT_ COMPUTE threads and NPW PRODUCER warps.
//   * Producers own the HBM stream.  While the compute threads work on image t, the producers pull
//     the 84-byte score vectors of image t+1 ONCE through private rings of shared-memory stages
//     (cp.async.bulk + mbarrier) and copy the present classes out as compact planes into an
//     L2-resident, double-buffered scratch (plus the suppressed absent maximum for single-class
//     label sets, weaklabel.cpp:45-65).  HBM therefore streams during the whole select chain
//     instead of only between chains.
//   * Compute threads never touch the score map (except for the rare exact recomputation of a
//     label): they form the floored running max E ("extrenum", weaklabel.cpp:67-77) in SHARED
//     memory from the planes of their tile and keep it there for the image.
//   * Cluster-wide synchronisation of the compute threads uses remote mbarrier arrivals (one
//     thread per CTA pair) instead of barrier.cluster, so that the producers are never coupled to
//     the select chain.
//   * Everything else follows estep_v2.cu: read-only planes with a bias history
//     (weaklabel.cpp:90-92), lower-bound zero counts to skip steps, windowed histogram + DSMEM
//     exchange + candidate gather for the order statistics (weaklabel.cpp:85-89), order-exact
//     running sums (weaklabel.cpp:69-80, 102-106; seqsum.cuh), bias-shifted re-centred argmax.
#include <float.h>
#include <math.h>
#include "estep_kernel.cuh"
#include "seqsum.cuh"

namespace wl {
namespace v3 {

constexpr int NB2 = 2048;               // fine bins per histogram round
constexpr int NB1 = 64;                 // super bins (32 fine bins each) exchanged through DSMEM
constexpr int SUBB = NB2 / NB1;
constexpr int GCAP = 1024;              // candidates gathered into every CTA
constexpr int MBW = 72;                 // mailbox words per CTA per exchange
constexpr int WIN_UNDER = 33;           // bin 0: zero margins; 1..32: below the window, by exponent group
constexpr int WIN_BINADES = 31;
constexpr int WIN_SH = 17;              // 64 bins per binade inside the window
constexpr int WIN_BINS = WIN_BINADES * 64;          // 1984
constexpr int WIN_OVER0 = WIN_UNDER + WIN_BINS;     // 2017 .. 2047: above the window
constexpr int EAGER_P = 4;              // planes whose zero counts are formed with E0
constexpr int HT = 512;                 // threads that scan a 2048-bin histogram (4 bins each)
constexpr int HCAP = 6;                 // biases remembered per class before its plane is rewritten
constexpr int MAXG = V3_MAXG;           // largest cluster
constexpr int NPW = V3_NPW;             // producer warps per CTA
constexpr int NST = V3_NST;             // stages of the staging ring (shared by the producer warps)
constexpr int STAGE_BYTES = V3_STAGE_BYTES;
constexpr int PF_AHEAD = V3_PF_AHEAD;   // chunks prefetched into the L2 ahead of the ring
static_assert(NST % NPW == 0, "a stage must always belong to the same producer warp");

// ---------------------------------------------------------------------------------------------
// PTX helpers: cluster, distributed shared memory, mbarrier, bulk copy
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void hw_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n"
                 "barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
template <int NT_> __device__ __forceinline__ void csync() {       // the compute threads of this CTA
    asm volatile("bar.sync 1, %0;\n" ::"n"(NT_) : "memory");
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t peer_addr(const void *local_smem, int rank) {
    uint32_t a = smem_u32(local_smem), r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(r) : "r"(a), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_peer(uint32_t addr, uint32_t v) {
    asm volatile("st.shared::cluster.u32 [%0], %1;\n" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void st_peer64(uint32_t addr, unsigned long long v) {
    asm volatile("st.shared::cluster.u64 [%0], %1;\n" ::"r"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_peer(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared::cluster.u32 %0, [%1];\n" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint4 ld_peer128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared::cluster.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {           // release.cta
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_peer(uint32_t cluster_addr) {   // release at cluster scope
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {  // acquire.cta
    asm volatile("{\n.reg .pred p;\nWL_WAIT_%=:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra WL_DONE_%=;\nbra WL_WAIT_%=;\nWL_DONE_%=:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t *bar, uint32_t parity) {  // acquire.cluster
    asm volatile("{\n.reg .pred p;\nWL_WAITC_%=:\n"
                 "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra WL_DONEC_%=;\nbra WL_WAITC_%=;\nWL_DONEC_%=:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (16-byte aligned on both sides, size a multiple of 16), completion on bar
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// ---- tensor memory (TMEM) as a per-thread array: 32x32b shape, thread i of the warp <-> TMEM lane
// (32 * (warp % 4) + i), one 32-bit column per element.  256 KB per SM that nothing else uses here.
template <int COLS> __device__ __forceinline__ void tm_alloc(uint32_t *smem_result) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(smem_result)), "n"(COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
template <int COLS> __device__ __forceinline__ void tm_dealloc(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void tm_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tm_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void tm_ld8(uint32_t taddr, float (&v)[8]) {
    uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7) : "r"(taddr) : "memory");
    tm_wait_ld();
    v[0] = __uint_as_float(r0); v[1] = __uint_as_float(r1); v[2] = __uint_as_float(r2); v[3] = __uint_as_float(r3);
    v[4] = __uint_as_float(r4); v[5] = __uint_as_float(r5); v[6] = __uint_as_float(r6); v[7] = __uint_as_float(r7);
}
__device__ __forceinline__ float tm_ld1(uint32_t taddr) {
    uint32_t r0;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];\n" : "=r"(r0) : "r"(taddr) : "memory");
    tm_wait_ld();
    return __uint_as_float(r0);
}
__device__ __forceinline__ void tm_st8(uint32_t taddr, const float (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};\n"
                 ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
                   "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
                   "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])) : "memory");
}
__device__ __forceinline__ void tm_st1(uint32_t taddr, float v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};\n" ::"r"(taddr), "r"(__float_as_uint(v)) : "memory");
}
// visit my PPT plane elements (element j <-> tile pixel j * NT + tid), eight at a time
template <int PPT, typename F> __device__ __forceinline__ void for_each_x(uint32_t tx, F &&f) {
    constexpr int NC = PPT / 8;
#pragma unroll
    for (int ch = 0; ch < NC; ch++) {
        float v[8];
        tm_ld8(tx + ch * 8, v);
#pragma unroll
        for (int e = 0; e < 8; e++) f(ch * 8 + e, v[e]);
    }
#pragma unroll
    for (int j = NC * 8; j < PPT; j++) f(j, tm_ld1(tx + j));
}
__host__ __device__ constexpr int tm_cols(int ppt) {          // 7 warps share a lane quarter: 7 * PPT columns, a power of two >= 32
    return 7 * ppt <= 32 ? 32 : 7 * ppt <= 64 ? 64 : 7 * ppt <= 128 ? 128 : 7 * ppt <= 256 ? 256 : 512;
}

// L2 policy for the score stream: every byte is read exactly once, so its lines should be the first to
// leave the L2 (the plane scratch, which is re-read for the whole select chain, should stay)
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;\n" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;\n"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
    return t;
}

// ---------------------------------------------------------------------------------------------
// shared memory
// ---------------------------------------------------------------------------------------------
struct Misc {
    double wscr_d[32];
    uint32_t wscr_u[32];
    uint32_t wred[32][8];
    uint32_t outbox[MBW];
    uint32_t zlb[MAXC];                 // zero-margin count per plane: exact or a lower bound
    float hbias[MAXC][HCAP];            // bias history per plane (weaklabel.cpp:90-92)
    int hn[MAXC];
    uint64_t full_bar[NST];             // producers: stage filled (bulk-copy bytes landed)
    uint64_t e0_bar;                    // compute: E0 of my tile landed in shared memory
    uint32_t tmem_base;                 // tcgen05.alloc result
    uint32_t pstat[2][NPW][8];          // producers -> compute: zero counts [0..3], largest margin [4] per warp
    uint64_t cmd_bar[2];                // compute -> producers: image of sequence parity q may be streamed
    uint64_t ready_bar[2];              // producers -> compute: planes of sequence parity q are in the scratch
    uint64_t xbar[2];                   // cluster sync of the compute threads (alternating)
    int cmd_img[2];                     // image index of the command (>= B: stop)
    unsigned long long sumres;          // written by the walker of an exact sum (any CTA's copy)
    double bc_d;
    uint32_t bc_u;
    int sel_bin;
    uint32_t sel_pre, sel_cnt;
    int sel_S;
    uint32_t sel_preS;
    uint32_t lcount;
    uint8_t plist[MAXC];
    uint8_t pplist[NPW][MAXC];          // the producers' copy (image t+1)
};

struct SelScratch {
    uint32_t hist[2][NB2];              // rotating per histogram round; peers read 32 fine bins
    uint32_t lhist[NB2];                // CTA-local selection among gathered candidates
    uint32_t gather[GCAP];
    uint32_t mylist[GCAP];
    uint32_t fine[MAXG][SUBB];
};

template <typename T, int NT_> struct SumScratch {
    ChunkFn<T> pf[NT_];                 // composite from the head of the chunk's segment to the chunk
    ChunkFn<T> wagg[32];                // per warp: carry between warps
    uint16_t heads[NT_];                // chunk index of every segment head, ascending
    int wany[32];
    int nheads;
    int pad[3];
    alignas(16) float terms[64];        // the walker's staging of a chunk's terms
};

template <int NT_> struct Layout {
    static constexpr int NW = NT_ / 32;
    static __host__ __device__ constexpr size_t A16(size_t x) { return (x + 15) / 16 * 16; }
    static __host__ __device__ constexpr size_t uni_bytes() {
        size_t b = A16(sizeof(SelScratch)), c = A16(sizeof(SumScratch<double, NT_>));
        return b > c ? b : c;
    }
    static constexpr size_t MBOX_BYTES = 2 * MAXG * MBW * 4;
    static constexpr size_t STAGING_BYTES = (size_t)NST * STAGE_BYTES;
    static size_t total(int ppt) {
        return A16((size_t)ppt * NT_ * 4) + MBOX_BYTES + A16(sizeof(Misc)) + uni_bytes() + STAGING_BYTES;
    }
};

template <int NT_> struct Cx {
    const EStepParams *p;
    int tid, lane, warp, rank, G;
    float *planes;        // scratch of the image in work: [C+1][Npad]
    float *absmax;        // plane C of the scratch (single-class label sets only)
    uint8_t *absidx;
    float *s_E;
    uint32_t (*mbox)[MAXG][MBW];
    Misc *ms;
    unsigned char *uni;
    int base, n_t;        // my tile: pixels [base, base + n_t) of the image
    uint32_t xround;      // exchanges done (mailbox parity), identical in every CTA
    uint32_t hround;      // histogram rounds done (hist buffer parity)
    uint32_t csyncs;      // cluster syncs done (xbar index / parity)
    uint32_t tx;          // TMEM address of my plane elements (lane quarter of my warp, my warp's columns)
    // per image
    int b, P;
    uint32_t mask;
    int first_absent;
};

template <int NT_> __device__ __forceinline__ void trc(const Cx<NT_> &c, int k, uint32_t v) {
    if (c.p->trail && c.tid == 0) c.p->trail[blockIdx.x * 32 + k] = v;
}

// Cluster-wide barrier of the COMPUTE threads (the producers never take part): every CTA arrives
// on the barrier of every CTA of the cluster (release at cluster scope, after a CTA-level barrier
// so that the arrival covers the writes of all its compute threads), then waits on its own.  Two
// barriers alternate: an arrival for sync n+1 can overtake a peer's arrival for sync n only onto
// the other barrier.
template <int NT_>
__device__ __forceinline__ void cluster_sync_compute(Cx<NT_> &c) {
    csync<NT_>();
    if (c.G > 1) {
        uint64_t *bar = &c.ms->xbar[c.csyncs & 1u];
        if (c.tid < c.G) mbar_arrive_peer(peer_addr(bar, c.tid));
        mbar_wait_cluster(bar, (c.csyncs >> 1) & 1u);
        c.csyncs++;
    }
}

// Every CTA hands the first n words of its outbox to every CTA of the cluster (itself included);
// returns the mailbox that now holds them: box[p][w] = word w of CTA p.
template <int NT_>
__device__ __forceinline__ const uint32_t (*exchange(Cx<NT_> &c, int n))[MBW] {
    Misc *s = c.ms;
    const uint32_t par = c.xround & 1u;
    csync<NT_>();                                   // outbox complete
    if (c.G > 1) {
        for (int i = c.tid; i < c.G * n; i += NT_) {
            const int peer = i / n, w = i - peer * n;
            st_peer(peer_addr(&c.mbox[par][c.rank][w], peer), s->outbox[w]);
        }
    } else {
        for (int w = c.tid; w < n; w += NT_) c.mbox[par][0][w] = s->outbox[w];
    }
    cluster_sync_compute(c);
    c.xround++;
    return c.mbox[par];
}

template <int NT_>
__device__ __forceinline__ uint32_t block_excl_scan_u32(Cx<NT_> &c, uint32_t v, uint32_t *total) {
    constexpr int NW = NT_ / 32;
    uint32_t inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (c.lane >= o) inc += t;
    }
    if (c.lane == 31) c.ms->wscr_u[c.warp] = inc;
    csync<NT_>();
    if (c.warp == 0) {
        uint32_t w = c.lane < NW ? c.ms->wscr_u[c.lane] : 0u, winc = w;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
            if (c.lane >= o) winc += t;
        }
        c.ms->wscr_u[c.lane] = winc - w;          // exclusive warp offsets
        if (c.lane == 31) c.ms->bc_u = winc;
    }
    csync<NT_>();
    uint32_t r = c.ms->wscr_u[c.warp] + inc - v;
    *total = c.ms->bc_u;
    csync<NT_>();
    return r;
}

template <int NT_>
__device__ __forceinline__ double block_excl_scan_f64(Cx<NT_> &c, double v, double *total) {
    constexpr int NW = NT_ / 32;
    double inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        double t = __shfl_up_sync(0xffffffffu, inc, o);
        if (c.lane >= o) inc += t;
    }
    if (c.lane == 31) c.ms->wscr_d[c.warp] = inc;
    csync<NT_>();
    if (c.warp == 0) {
        double w = c.lane < NW ? c.ms->wscr_d[c.lane] : 0.0, winc = w;
        for (int o = 1; o < 32; o <<= 1) {
            double t = __shfl_up_sync(0xffffffffu, winc, o);
            if (c.lane >= o) winc += t;
        }
        c.ms->wscr_d[c.lane] = winc - w;
        if (c.lane == 31) c.ms->bc_d = winc;
    }
    csync<NT_>();
    double r = c.ms->wscr_d[c.warp] + (inc - v);
    *total = c.ms->bc_d;
    csync<NT_>();
    return r;
}

// block reduction of up to 6 counters (sums) and one max into outbox[off .. off+6]
template <int NT_>
__device__ __forceinline__ void block_reduce_to_outbox(Cx<NT_> &c, uint32_t z0, uint32_t z1, uint32_t z2,
                                                       uint32_t z3, uint32_t mx, int off) {
    constexpr int NW = NT_ / 32;
    Misc *s = c.ms;
    for (int o = 16; o; o >>= 1) {
        z0 += __shfl_xor_sync(0xffffffffu, z0, o);
        z1 += __shfl_xor_sync(0xffffffffu, z1, o);
        z2 += __shfl_xor_sync(0xffffffffu, z2, o);
        z3 += __shfl_xor_sync(0xffffffffu, z3, o);
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if (c.lane == 0) {
        s->wred[c.warp][0] = z0; s->wred[c.warp][1] = z1; s->wred[c.warp][2] = z2;
        s->wred[c.warp][3] = z3; s->wred[c.warp][4] = mx;
    }
    csync<NT_>();
    if (c.warp == 0) {
        uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0, am = 0;
        if (c.lane < NW) {
            a0 = s->wred[c.lane][0]; a1 = s->wred[c.lane][1]; a2 = s->wred[c.lane][2];
            a3 = s->wred[c.lane][3]; am = s->wred[c.lane][4];
        }
        for (int o = 16; o; o >>= 1) {
            a0 += __shfl_xor_sync(0xffffffffu, a0, o);
            a1 += __shfl_xor_sync(0xffffffffu, a1, o);
            a2 += __shfl_xor_sync(0xffffffffu, a2, o);
            a3 += __shfl_xor_sync(0xffffffffu, a3, o);
            am = max(am, __shfl_xor_sync(0xffffffffu, am, o));
        }
        if (c.lane == 0) {
            s->outbox[off + 0] = a0; s->outbox[off + 1] = a1; s->outbox[off + 2] = a2;
            s->outbox[off + 3] = a3; s->outbox[off + 4] = am;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// bias history: the current value of a plane element = its original score with the biases the
// class received so far added one by one in float (weaklabel.cpp:90-92)
// ---------------------------------------------------------------------------------------------
struct Hist {
    float b[HCAP];
    int n;
};
__device__ __forceinline__ Hist load_hist(const Misc *ms, int plane) {
    Hist h;
    h.n = ms->hn[plane];
#pragma unroll
    for (int e = 0; e < HCAP; e++) h.b[e] = ms->hbias[plane][e];
    return h;
}
__device__ __forceinline__ float with_hist(float x, const Hist &h) {
#pragma unroll
    for (int e = 0; e < HCAP; e++)
        if (e < h.n) x = __fadd_rn(x, h.b[e]);
    return x;
}

// ---------------------------------------------------------------------------------------------
// producers: HBM -> (L2 prefetch) -> staging ring (bulk copy) -> compact planes + E0 plane in the
// L2 scratch, zero-margin counts and the largest possible margin of my tile
// ---------------------------------------------------------------------------------------------
struct Prod {
    const EStepParams *p;
    Misc *ms;
    unsigned char *staging;   // NST stages shared by the producer warps
    int pw, lane;
    int base, n_t;
    uint32_t nbase;           // chunks of earlier images (stage and parity of chunk q: from nbase + q)
    uint64_t pol;             // L2 evict-first policy of the score stream
};

__device__ __forceinline__ void prod_prefetch(const Prod &w, const float *img, int q) {
    const EStepParams &p = *w.p;
    const int cpx = p.sub_px;
    const int npx = min(cpx, w.n_t - q * cpx);
    const uintptr_t a0 = (uintptr_t)(img + (long long)(w.base + q * cpx) * p.C);
    const uintptr_t lo = (a0 + 15u) & ~(uintptr_t)15u, hi = (a0 + (uintptr_t)npx * p.C * 4u) & ~(uintptr_t)15u;
    if (hi > lo)
        asm volatile("cp.async.bulk.prefetch.L2.global.L2::cache_hint [%0], %1, %2;\n" ::"l"(lo), "r"((uint32_t)(hi - lo)), "l"(w.pol) : "memory");
}

// called by a whole warp: copy chunk q of the tile into its stage, prefetch chunk q + PF_AHEAD
__device__ __forceinline__ void prod_issue(const Prod &w, const float *img, int q, int nchunks) {
    const EStepParams &p = *w.p;
    const int cpx = p.sub_px;
    const int npx = min(cpx, w.n_t - q * cpx);
    const int nfl = npx * p.C;
    const float *g0 = img + (long long)(w.base + q * cpx) * p.C;
    const int lead = (int)(((uintptr_t)g0 & 15u) >> 2);
    const int d = (int)((w.nbase + (uint32_t)q) % NST);
    float *stage = reinterpret_cast<float *>(w.staging + (size_t)d * STAGE_BYTES);
    const int nh = lead ? min(4 - lead, nfl) : 0;               // floats before the first aligned 16 bytes
    const uint32_t bytes = (uint32_t)((nfl - nh) * 4) & ~15u;   // aligned interior
    const int t0 = nh + (int)(bytes >> 2);                      // first float of the ragged tail
    uint64_t *bar = &w.ms->full_bar[d];
    if (w.lane >= 1 && w.lane <= 3) {
        const int f = w.lane - 1;
        if (f < nh) stage[lead + f] = __ldg(g0 + f);
    } else if (w.lane >= 4 && w.lane <= 6) {
        const int f = t0 + (w.lane - 4);
        if (f < nfl) stage[lead + f] = __ldg(g0 + f);
    } else if (w.lane == 7) {
        if (q + p.pf_ahead < nchunks && p.pf_ahead > 0) prod_prefetch(w, img, q + p.pf_ahead);
    }
    __syncwarp();                                               // ragged stores before the arrival (release)
    if (w.lane == 0) {
        if (bytes) {
            mbar_expect_tx(bar, bytes);
            bulk_g2s_hint(stage + lead + nh, g0 + nh, bytes, bar, w.pol);
        } else {
            mbar_arrive(bar);
        }
    }
}

// one image: chunks q with (nbase + q) % NPW == pw of the CTA's tile
__device__ __forceinline__ void prod_image(Prod &w, int b, int par, float *planes, uint8_t *absidx) {
    const EStepParams &p = *w.p;
    const ImageMeta *im = p.meta + b;
    const int P = im->P;
    uint8_t *pl = w.ms->pplist[w.pw];
    if (w.lane < MAXC) pl[w.lane] = im->plist[w.lane];
    __syncwarp();
    const int pl0 = pl[0], pl1 = pl[1], pl2 = pl[2], pl3 = pl[3];
    const int first_absent = im->first_absent;
    const float *img = p.map + (long long)b * p.s0;
    const int cpx = p.sub_px;
    const int nchunks = (w.n_t + cpx - 1) / cpx;
    const bool need_abs = (P == 1) && (P < p.C);
    float *absmax = planes + (size_t)p.C * p.Npad;
    float *e0pl = planes + (size_t)(p.C + 1) * p.Npad;
    uint32_t zc0 = 0u, zc1 = 0u, zc2 = 0u, zc3 = 0u;
    float dmx = 0.f;
    const int q0 = (int)((w.pw + NPW - (w.nbase % NPW)) % NPW);   // my first chunk
    // every stage is free (the producers met after the previous image): fill the ring, warm the L2
    for (int q = q0; q < nchunks && q < max(NST, p.pf_ahead); q += NPW) {
        if (q < NST) prod_issue(w, img, q, nchunks);
        else if (w.lane == 0) prod_prefetch(w, img, q);
    }
    for (int q = q0; q < nchunks; q += NPW) {
        const uint32_t n = w.nbase + (uint32_t)q;
        const int d = (int)(n % NST);
        mbar_wait(&w.ms->full_bar[d], (n / NST) & 1u);
        if (p.trail && w.lane == 0) p.trail[blockIdx.x * 32 + 8 + w.pw] = ((uint32_t)par << 28) | ((uint32_t)(b & 0xfff) << 16) | (uint32_t)q;
        const float *stage = reinterpret_cast<const float *>(w.staging + (size_t)d * STAGE_BYTES);
        const int npx = min(cpx, w.n_t - q * cpx);
        const float *g0 = img + (long long)(w.base + q * cpx) * p.C;
        const int lead = (int)(((uintptr_t)g0 & 15u) >> 2);
        const int i0 = w.base + q * cpx;
        float *out = planes + i0;
        if (need_abs) {
            // single-class label set: keep the suppressed absent maximum (weaklabel.cpp:45-65) and
            // whether the lowest absent class attains it (else a tie is resolved from the scores)
            for (int gt = w.lane; gt < npx; gt += 32) {
                const float *px = stage + lead + gt * p.C;
                const float x = px[pl0];
                const float low = __fsub_rn(x, p.margin);
                float a0 = -INFINITY, a1 = -INFINITY;
                int k = 0;
                if (p.C == 21) {                        // the VOC class count: fully unrolled
                    float a2 = -INFINITY;
#pragma unroll
                    for (int kk = 0; kk < 21; kk += 3) {
                        const float t0 = px[kk], t1 = px[kk + 1], t2 = px[kk + 2];
                        const float u0 = (kk == pl0) ? -INFINITY : (t0 > x ? low : t0);
                        const float u1 = (kk + 1 == pl0) ? -INFINITY : (t1 > x ? low : t1);
                        const float u2 = (kk + 2 == pl0) ? -INFINITY : (t2 > x ? low : t2);
                        a0 = fmaxf(a0, u0);
                        a1 = fmaxf(a1, u1);
                        a2 = fmaxf(a2, u2);
                    }
                    a0 = fmaxf(a0, a2);
                    k = p.C;                            // both generic loops below are empty
                }
                for (; k + 1 < pl0; k += 2) {
                    const float t0 = px[k], t1 = px[k + 1];
                    a0 = fmaxf(a0, t0 > x ? low : t0);
                    a1 = fmaxf(a1, t1 > x ? low : t1);
                }
                if (k < pl0) { const float t0 = px[k]; a0 = fmaxf(a0, t0 > x ? low : t0); }
                k = max(k, pl0 + 1);
                for (; k + 1 < p.C; k += 2) {
                    const float t0 = px[k], t1 = px[k + 1];
                    a0 = fmaxf(a0, t0 > x ? low : t0);
                    a1 = fmaxf(a1, t1 > x ? low : t1);
                }
                if (k < p.C) { const float t0 = px[k]; a0 = fmaxf(a0, t0 > x ? low : t0); }
                const float amax = fmaxf(a0, a1);
                float tf = px[first_absent];
                if (tf > x) tf = low;
                float E = FLT_MIN;
                if (x > E) E = x;
                __stcg(out + gt, x);
                __stcg(e0pl + i0 + gt, E);
                __stcg(absmax + i0 + gt, amax);
                absidx[i0 + gt] = (uint8_t)((tf == amax) ? first_absent : 254);
                zc0 += (x == E) ? 1u : 0u;
                dmx = fmaxf(dmx, __fsub_rn(E, x));
            }
        } else if (P <= 4) {
            for (int gt = w.lane; gt < npx; gt += 32) {
                const float *px = stage + lead + gt * p.C;
                const float v0 = px[pl0];
                const float v1 = (P > 1) ? px[pl1] : v0;
                const float v2 = (P > 2) ? px[pl2] : v0;
                const float v3 = (P > 3) ? px[pl3] : v0;
                float E = FLT_MIN, mn = FLT_MAX;
                if (v0 > E) E = v0;
                if (v0 < mn) mn = v0;
                if (v1 > E) E = v1;
                if (v1 < mn) mn = v1;
                if (v2 > E) E = v2;
                if (v2 < mn) mn = v2;
                if (v3 > E) E = v3;
                if (v3 < mn) mn = v3;
                __stcg(out + gt, v0);
                if (P > 1) __stcg(out + (size_t)p.Npad + gt, v1);
                if (P > 2) __stcg(out + 2 * (size_t)p.Npad + gt, v2);
                if (P > 3) __stcg(out + 3 * (size_t)p.Npad + gt, v3);
                __stcg(e0pl + i0 + gt, E);
                zc0 += (v0 == E) ? 1u : 0u;
                zc1 += (P > 1 && v1 == E) ? 1u : 0u;
                zc2 += (P > 2 && v2 == E) ? 1u : 0u;
                zc3 += (P > 3 && v3 == E) ? 1u : 0u;
                dmx = fmaxf(dmx, __fsub_rn(E, mn));
            }
        } else {
            for (int gt = w.lane; gt < npx; gt += 32) {
                const float *px = stage + lead + gt * p.C;
                float E = FLT_MIN, mn = FLT_MAX;
                for (int k = 0; k < P; k++) {
                    const float v = px[pl[k]];
                    __stcg(out + (size_t)k * p.Npad + gt, v);
                    if (v > E) E = v;
                    if (v < mn) mn = v;
                }
                __stcg(e0pl + i0 + gt, E);
                zc0 += (px[pl0] == E) ? 1u : 0u;
                zc1 += (px[pl1] == E) ? 1u : 0u;
                zc2 += (px[pl2] == E) ? 1u : 0u;
                zc3 += (px[pl3] == E) ? 1u : 0u;
                dmx = fmaxf(dmx, __fsub_rn(E, mn));
            }
        }
        __syncwarp();                                  // stage drained by every lane
        if (q + NST < nchunks) prod_issue(w, img, q + NST, nchunks);
    }
    w.nbase += (uint32_t)nchunks;
    // my share of the tile's statistics
    uint32_t dm = __float_as_uint(dmx);               // margins are >= +0: the bit patterns order like the values
    for (int o = 16; o; o >>= 1) {
        zc0 += __shfl_xor_sync(0xffffffffu, zc0, o);
        zc1 += __shfl_xor_sync(0xffffffffu, zc1, o);
        zc2 += __shfl_xor_sync(0xffffffffu, zc2, o);
        zc3 += __shfl_xor_sync(0xffffffffu, zc3, o);
        dm = max(dm, __shfl_xor_sync(0xffffffffu, dm, o));
    }
    if (w.lane == 0) {
        uint32_t *st = w.ms->pstat[par][w.pw];
        st[0] = zc0; st[1] = zc1; st[2] = zc2; st[3] = zc3; st[4] = dm;
    }
}

template <int NT_>
__device__ __forceinline__ void producer_main(const EStepParams &p, Misc *ms, unsigned char *staging, int rank,
                                              int slot, int base, int n_t) {
    Prod w;
    w.p = &p;
    w.ms = ms;
    w.pw = (int)(threadIdx.x - NT_) >> 5;
    w.lane = threadIdx.x & 31;
    w.staging = staging;
    w.base = base;
    w.n_t = n_t;
    w.nbase = 0;
    w.pol = l2_evict_first_policy();
    for (uint32_t s = 0;; s++) {
        const int q = s & 1u;
        if (p.trail && w.lane == 0) p.trail[blockIdx.x * 32 + 12 + w.pw] = (s << 8) | 1u;
        mbar_wait(&ms->cmd_bar[q], (s >> 1) & 1u);
        const int b = *reinterpret_cast<volatile int *>(&ms->cmd_img[q]);
        if (b >= p.B) break;
        float *planes = p.planes + (size_t)(slot * 2 + q) * (size_t)(p.C + 2) * p.Npad;
        uint8_t *absidx = p.absent_idx + (size_t)(slot * 2 + q) * p.Npad;
        asm volatile("bar.sync 2, %0;\n" ::"n"(NPW * 32) : "memory");   // every stage of the ring is drained
        prod_image(w, b, q, planes, absidx);
        __threadfence();                                        // my plane/E0 stores have reached the L2 (the
        __syncwarp();                                           // E0 tile is fetched from there by the copy engine)
        if (w.lane == 0) mbar_arrive(&ms->ready_bar[q]);        // release: planes and statistics of my chunks
        if (p.trail && w.lane == 0) p.trail[blockIdx.x * 32 + 12 + w.pw] = (s << 8) | 2u;
    }
}

// ---------------------------------------------------------------------------------------------
// order-exact running sum of E over the whole image (weaklabel.cpp:69-80 float, :102-106 double)
// ---------------------------------------------------------------------------------------------
template <typename T> struct FnIO;
template <> struct FnIO<float> {
    static __device__ __forceinline__ ChunkFn<float> ld_peer_fn(const ChunkFn<float> *local, int rank) {
        const uint4 v = ld_peer128(peer_addr(local, rank));
        ChunkFn<float> f;
        f.e = v.x; f.a0 = (int32_t)v.y; f.a1 = (int32_t)v.z; f.kind = (int)v.w;
        return f;
    }
    static __device__ __forceinline__ ChunkFn<float> shfl(const ChunkFn<float> &f, int src) {
        ChunkFn<float> g;
        g.e = __shfl_sync(0xffffffffu, f.e, src);
        g.a0 = __shfl_sync(0xffffffffu, f.a0, src);
        g.a1 = __shfl_sync(0xffffffffu, f.a1, src);
        g.kind = __shfl_sync(0xffffffffu, f.kind, src);
        return g;
    }
    static __device__ __forceinline__ ChunkFn<float> shfl_up(const ChunkFn<float> &f, int d) {
        ChunkFn<float> g;
        g.e = __shfl_up_sync(0xffffffffu, f.e, d);
        g.a0 = __shfl_up_sync(0xffffffffu, f.a0, d);
        g.a1 = __shfl_up_sync(0xffffffffu, f.a1, d);
        g.kind = __shfl_up_sync(0xffffffffu, f.kind, d);
        return g;
    }
    static __device__ __forceinline__ ChunkFn<float> shfl_down(const ChunkFn<float> &f, int d) {
        ChunkFn<float> g;
        g.e = __shfl_down_sync(0xffffffffu, f.e, d);
        g.a0 = __shfl_down_sync(0xffffffffu, f.a0, d);
        g.a1 = __shfl_down_sync(0xffffffffu, f.a1, d);
        g.kind = __shfl_down_sync(0xffffffffu, f.kind, d);
        return g;
    }
};
template <> struct FnIO<double> {
    static __device__ __forceinline__ ChunkFn<double> ld_peer_fn(const ChunkFn<double> *local, int rank) {
        const uint32_t a = peer_addr(local, rank);
        const uint4 v = ld_peer128(a), w = ld_peer128(a + 16);
        ChunkFn<double> f;
        f.e = (uint64_t)v.x | ((uint64_t)v.y << 32);
        f.a0 = (int64_t)((uint64_t)v.z | ((uint64_t)v.w << 32));
        f.a1 = (int64_t)((uint64_t)w.x | ((uint64_t)w.y << 32));
        f.kind = (int)w.z;
        return f;
    }
    static __device__ __forceinline__ ChunkFn<double> shfl(const ChunkFn<double> &f, int src) {
        ChunkFn<double> g;
        g.e = (uint64_t)__shfl_sync(0xffffffffu, (unsigned long long)f.e, src);
        g.a0 = (int64_t)__shfl_sync(0xffffffffu, (long long)f.a0, src);
        g.a1 = (int64_t)__shfl_sync(0xffffffffu, (long long)f.a1, src);
        g.kind = __shfl_sync(0xffffffffu, f.kind, src);
        return g;
    }
    static __device__ __forceinline__ ChunkFn<double> shfl_up(const ChunkFn<double> &f, int d) {
        ChunkFn<double> g;
        g.e = (uint64_t)__shfl_up_sync(0xffffffffu, (unsigned long long)f.e, d);
        g.a0 = (int64_t)__shfl_up_sync(0xffffffffu, (long long)f.a0, d);
        g.a1 = (int64_t)__shfl_up_sync(0xffffffffu, (long long)f.a1, d);
        g.kind = __shfl_up_sync(0xffffffffu, f.kind, d);
        return g;
    }
    static __device__ __forceinline__ ChunkFn<double> shfl_down(const ChunkFn<double> &f, int d) {
        ChunkFn<double> g;
        g.e = (uint64_t)__shfl_down_sync(0xffffffffu, (unsigned long long)f.e, d);
        g.a0 = (int64_t)__shfl_down_sync(0xffffffffu, (long long)f.a0, d);
        g.a1 = (int64_t)__shfl_down_sync(0xffffffffu, (long long)f.a1, d);
        g.kind = __shfl_down_sync(0xffffffffu, f.kind, d);
        return g;
    }
};
static_assert(sizeof(ChunkFn<float>) == 16 && sizeof(ChunkFn<double>) == 32, "ChunkFn layout");

// per-thread chunk sums in double (the prediction of seqsum.cuh); returns my chunk's exclusive
// prefix inside the tile, *tile_total = the tile's sum
template <int NT_>
__device__ __forceinline__ double chunk_prefix(Cx<NT_> &c, double *tile_total, double *chunk_sum) {
    const int L = c.p->L;
    const int c0 = c.tid * L;
    const int len = max(0, min(L, c.n_t - c0));
    double cs = 0.0;
    for (int i = 0; i < len; i++) cs += (double)c.s_E[c0 + i];
    *chunk_sum = cs;
    return block_excl_scan_f64(c, cs, tile_total);
}

// Tolerance (relative to the binade) inside which a predicted binade boundary is not trusted: the
// float running sum drifts from the real sum by accumulated rounding (a few hundred ulps at most
// here), the double one practically not at all.  Only speed depends on it, never the result.
template <typename T> struct drift_tol;
template <> struct drift_tol<float> { static constexpr double value = 1.0 / 8192.0; };
template <> struct drift_tol<double> { static constexpr double value = 1.0 / 1073741824.0; };

__device__ __forceinline__ uint32_t ld_peer16(uint32_t addr) {
    uint16_t v;
    asm volatile("ld.shared::cluster.u16 %0, [%1];\n" : "=h"(v) : "r"(addr) : "memory");
    return v;
}

// `before` = the (double) sum of the tiles of lower rank; cpre / cs from chunk_prefix.
//
// Every chunk is summarised for its predicted binade (seqsum.cuh); maximal runs of consecutive
// summaries of the same binade ("segments") are composed by a segmented scan over the whole tile,
// so that the walker crosses a segment with ONE application of the composite stored at its last
// chunk.  Chunks that contain (or are suspiciously close to) a binade crossing form singleton
// segments and are added term by term.  A composite that does not apply to the actual state
// (mispredicted binade) falls back to term-by-term addition of its segment: exact in every case.
template <typename T, int NT_>
__device__ __forceinline__ T exact_sum(Cx<NT_> &c, double cpre, double cs, double before) {
    typedef fp_traits<T> F;
    typedef FnIO<T> IO;
    constexpr int NW = NT_ / 32;
    const EStepParams &p = *c.p;
    SumScratch<T, NT_> *ss = reinterpret_cast<SumScratch<T, NT_> *>(c.uni);
    const int L = p.L;
    const int c0 = c.tid * L;
    const int len = max(0, min(L, c.n_t - c0));
    const double pin = before + cpre;
    ChunkFn<T> f = chunk_eval<T>(c.s_E + c0, len, pin);
    if (f.kind == 1) {
        // distrust predictions that pass close to a boundary of the binade
        const double lo = (double)F::from(f.e), tol = lo * drift_tol<T>::value;
        if (pin - lo < tol || (lo + lo) - (pin + cs) < tol) f.kind = 0;
    }
    // segment heads
    ss->pf[c.tid] = f;
    csync<NT_>();
    bool head;
    if (f.kind == 2) head = false;
    else if (c.tid == 0 || f.kind == 0) head = true;
    else {
        const ChunkFn<T> pv = ss->pf[c.tid - 1];
        head = (pv.kind != 1) || (pv.e != f.e);
    }
    csync<NT_>();
    // segmented inclusive scan inside the warp
    ChunkFn<T> x = f;
    bool flag = head;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const ChunkFn<T> y = IO::shfl_up(x, d);
        const bool yf = __shfl_up_sync(0xffffffffu, flag ? 1 : 0, d) != 0;
        if (c.lane >= d && !flag) { x = chunk_compose(y, x); flag = yf; }
    }
    // warp aggregates -> carry into the next warps (segmented scan over the warps, by warp 0)
    ChunkFn<T> *wagg = ss->wagg;
    if (c.lane == 31) { wagg[c.warp] = x; ss->wany[c.warp] = flag ? 1 : 0; }
    csync<NT_>();
    if (c.warp == 0) {
        ChunkFn<T> v = c.lane < NW ? wagg[c.lane] : chunk_identity<T>();
        bool vf = c.lane < NW ? ss->wany[c.lane] != 0 : false;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const ChunkFn<T> y = IO::shfl_up(v, d);
            const bool yf = __shfl_up_sync(0xffffffffu, vf ? 1 : 0, d) != 0;
            if (c.lane >= d && !vf) { v = chunk_compose(y, v); vf = yf; }
        }
        // carry INTO warp w = inclusive result of warp w-1
        const ChunkFn<T> cin = IO::shfl_up(v, 1);
        if (c.lane < NW) wagg[c.lane] = (c.lane == 0) ? chunk_identity<T>() : cin;
    }
    csync<NT_>();
    if (!flag) x = chunk_compose(wagg[c.warp], x);     // no head at or before me inside my warp
    csync<NT_>();
    ss->pf[c.tid] = x;
    uint32_t nheads;
    const uint32_t hrank = block_excl_scan_u32(c, head ? 1u : 0u, &nheads);
    if (head) ss->heads[hrank] = (uint16_t)c.tid;
    if (c.tid == 0) ss->nheads = (int)nheads;
    cluster_sync_compute(c);                             // every tile's summaries are published
    if (c.rank == 0 && c.warp == 0) {
        // one warp walks the whole image; lanes fetch summaries / terms from the owners' shared
        // memory in parallel, the state s is identical in every lane
        int nh_r = 0;
        ChunkFn<T> cf_r = chunk_identity<T>();
        if (c.lane < c.G) {
            nh_r = (int)ld_peer(peer_addr(&ss->nheads, c.lane));
            cf_r = IO::ld_peer_fn(&ss->pf[NT_ - 1], c.lane);
        }
        T s = (T)0;
        for (int r = 0; r < c.G; r++) {
            const int n_t_r = max(0, min(p.TS, p.N - r * p.TS));
            if (n_t_r == 0) break;
            const int nh = __shfl_sync(0xffffffffu, nh_r, r);
            const ChunkFn<T> cf = IO::shfl(cf_r, r);
            if (nh == 1 && chunk_apply(cf, s)) continue;
            for (int k0 = 0; k0 < nh; k0 += 32) {
                const int kk = k0 + c.lane;
                int hk = NT_, tk = NT_ - 1;
                ChunkFn<T> pk = chunk_identity<T>();
                if (kk < nh) {
                    if (r == 0) {                           // my own tile
                        hk = (int)ss->heads[kk];
                        if (kk + 1 < nh) tk = (int)ss->heads[kk + 1] - 1;
                        pk = ss->pf[tk];
                    } else {
                        hk = (int)ld_peer16(peer_addr(&ss->heads[kk], r));
                        if (kk + 1 < nh) tk = (int)ld_peer16(peer_addr(&ss->heads[kk + 1], r)) - 1;
                        pk = IO::ld_peer_fn(&ss->pf[tk], r);
                    }
                }
                const int m = min(32, nh - k0);
                for (int k = 0; k < m; k++) {
                    const ChunkFn<T> fk = IO::shfl(pk, k);
                    const int h0 = __shfl_sync(0xffffffffu, hk, k), t0 = __shfl_sync(0xffffffffu, tk, k);
                    if (chunk_apply(fk, s)) continue;
                    // a crossing chunk (or a mispredicted segment): term by term
                    const int e0 = h0 * L, e1 = min((t0 + 1) * L, n_t_r);
                    // The adds form one dependent chain.  The terms are staged in an aligned buffer and read
                    // back four at a time (one shared-memory latency per four adds; a shuffle or a scalar
                    // load per term would put its latency into the chain).
                    for (int i0 = e0; i0 < e1; i0 += 64) {
                        const int m = min(64, e1 - i0);
                        __syncwarp();
                        if (r == 0) {                               // my own tile (the walker is rank 0)
                            if (c.lane < m) ss->terms[c.lane] = c.s_E[i0 + c.lane];
                            if (c.lane + 32 < m) ss->terms[c.lane + 32] = c.s_E[i0 + 32 + c.lane];
                        } else {
                            if (c.lane < m) ss->terms[c.lane] = __uint_as_float(ld_peer(peer_addr(c.s_E + i0 + c.lane, r)));
                            if (c.lane + 32 < m) ss->terms[c.lane + 32] = __uint_as_float(ld_peer(peer_addr(c.s_E + i0 + 32 + c.lane, r)));
                        }
                        __syncwarp();
                        const float4 *t4 = reinterpret_cast<const float4 *>(ss->terms);
                        int i = 0;
                        for (; i + 4 <= m; i += 4) {
                            const float4 v = t4[i >> 2];
                            s = F::add(s, (T)v.x);
                            s = F::add(s, (T)v.y);
                            s = F::add(s, (T)v.z);
                            s = F::add(s, (T)v.w);
                        }
                        for (; i < m; i++) s = F::add(s, (T)ss->terms[i]);
                    }
                }
            }
        }
        if (c.lane < c.G) st_peer64(peer_addr(&c.ms->sumres, c.lane), (unsigned long long)F::bits(s));
    }
    cluster_sync_compute(c);
    return F::from((typename F::U)c.ms->sumres);
}

// ---------------------------------------------------------------------------------------------
// binning of a margin key (bit pattern of a float >= +0)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t win_bin(uint32_t key, uint32_t lo_key) {
    if (key == 0u) return 0u;
    if (key < lo_key) return 1u + (key >> 26);
    const uint32_t o = key - lo_key, w = o >> WIN_SH;
    if (w < (uint32_t)WIN_BINS) return WIN_UNDER + w;
    return WIN_OVER0 + min((o - ((uint32_t)WIN_BINS << WIN_SH)) >> 26, 30u);
}
// inclusive key range of a window bin
__device__ __forceinline__ void win_range(uint32_t bin, uint32_t lo_key, uint32_t &klo, uint32_t &khi) {
    if (bin == 0u) { klo = khi = 0u; return; }
    if (bin < (uint32_t)WIN_UNDER) {
        const uint32_t g = bin - 1u;
        klo = max(g << 26, 1u);
        khi = min(((g + 1u) << 26) - 1u, lo_key - 1u);
        return;
    }
    if (bin < (uint32_t)WIN_OVER0) {
        klo = lo_key + ((bin - WIN_UNDER) << WIN_SH);
        khi = klo + ((1u << WIN_SH) - 1u);
        return;
    }
    const uint32_t ov = bin - WIN_OVER0;
    klo = lo_key + ((uint32_t)WIN_BINS << WIN_SH) + (ov << 26);
    khi = ov < 30u ? klo + ((1u << 26) - 1u) : 0xffffffffu;
}

struct Located {
    int bin;              // -1: inconsistent (NaN in the scores)
    uint32_t pre, cnt;    // keys of the histogram below the bin / inside it
};

template <int NT_> __device__ __forceinline__ SelScratch *sel(Cx<NT_> &c) { return reinterpret_cast<SelScratch *>(c.uni); }

// one histogram round over my tile's margins E - x.  Zero margins (usually the largest single
// bin by far) are counted with ballots instead of same-address shared atomics.
template <int NT_, int PPT, bool WINDOW>
__device__ __forceinline__ void hist_pass(Cx<NT_> &c, uint32_t lo, uint32_t hi, int sh, uint32_t *h) {
    for (int i = c.tid; i < NB2; i += NT_) h[i] = 0u;
    csync<NT_>();
    uint32_t zeros = 0u;
    for_each_x<PPT>(c.tx, [&](int j, float xv) {
        const int i = j * NT_ + c.tid;
        const bool valid = i < c.n_t;
        const uint32_t key = valid ? __float_as_uint(__fsub_rn(c.s_E[i], xv)) : 0xffffffffu;
        if (WINDOW) {
            zeros += (key == 0u) ? 1u : 0u;
            if (valid && key != 0u) atomicAdd(&h[win_bin(key, lo)], 1u);
        } else {
            if (valid && key >= lo && key <= hi) atomicAdd(&h[(key - lo) >> sh], 1u);
        }
    });
    if (WINDOW) {
        zeros = __reduce_add_sync(0xffffffffu, zeros);
        if (c.lane == 0 && zeros) atomicAdd(&h[0], zeros);
    }
    csync<NT_>();
}

// agree, across the cluster, on the bin of the current histogram that holds rank r
template <int NT_>
__device__ __forceinline__ Located locate(Cx<NT_> &c, uint32_t r) {
    constexpr int BPT = NB2 / HT;               // bins per scanning thread (the first HT threads scan)
    constexpr int TPS = SUBB / BPT;             // threads per super bin
    Misc *s = c.ms;
    SelScratch *ss = sel(c);
    uint32_t *h = ss->hist[c.hround & 1u];
    if (c.tid < HT) {   // 64 super-bin sums of my histogram -> outbox
        uint32_t v = 0u;
#pragma unroll
        for (int e = 0; e < BPT; e++) v += h[BPT * c.tid + e];
#pragma unroll
        for (int o = TPS / 2; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((c.tid & (TPS - 1)) == 0) s->outbox[c.tid / TPS] = v;
    }
    const uint32_t (*box)[MBW] = exchange(c, NB1);
    if (c.tid == 0) s->sel_S = -1;
    csync<NT_>();
    if (c.warp == 0) {
        uint32_t s0 = 0, s1 = 0;
        for (int p = 0; p < c.G; p++) { s0 += box[p][2 * c.lane]; s1 += box[p][2 * c.lane + 1]; }
        uint32_t inc = s0 + s1;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (c.lane >= o) inc += t;
        }
        const uint32_t ex = inc - (s0 + s1);
        if (r >= ex && r < ex + s0) { s->sel_S = 2 * c.lane; s->sel_preS = ex; }
        else if (r >= ex + s0 && r < ex + s0 + s1) { s->sel_S = 2 * c.lane + 1; s->sel_preS = ex + s0; }
    }
    csync<NT_>();
    const int S = s->sel_S;
    const uint32_t preS = s->sel_preS;
    Located L;
    L.bin = -1; L.pre = 0; L.cnt = 0;
    c.hround++;
    if (S < 0) return L;                               // uniform over the cluster
    if (c.tid < c.G * SUBB) {
        const int p = c.tid >> 5, i = c.tid & 31;
        ss->fine[p][i] = (p == c.rank) ? h[S * SUBB + i] : ld_peer(peer_addr(&h[S * SUBB + i], p));
    }
    if (c.tid == 0) s->sel_bin = -1;
    csync<NT_>();
    if (c.warp == 0) {
        uint32_t f = 0;
        for (int p = 0; p < c.G; p++) f += ss->fine[p][c.lane];
        uint32_t inc = f;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (c.lane >= o) inc += t;
        }
        const uint32_t ex = inc - f, rr = r - preS;
        if (rr >= ex && rr < ex + f) { s->sel_bin = S * SUBB + c.lane; s->sel_pre = preS + ex; s->sel_cnt = f; }
    }
    csync<NT_>();
    L.bin = s->sel_bin;
    L.pre = s->sel_pre;
    L.cnt = s->sel_cnt;
    return L;
}

// exact selection of rank r among cnt keys in [lo, hi] held in CTA-local shared memory
template <int NT_>
__device__ __forceinline__ uint32_t local_select(Cx<NT_> &c, const uint32_t *keys, uint32_t cnt, uint32_t r,
                                                 uint32_t lo, uint32_t hi) {
    constexpr int BPT = NB2 / HT;
    Misc *s = c.ms;
    uint32_t *h = sel(c)->lhist;
    for (int guard = 0; guard < 4; guard++) {
        if (lo == hi) break;
        int sh = 0;
        while (((hi - lo) >> sh) >= (uint32_t)NB2) sh++;
        for (int i = c.tid; i < NB2; i += NT_) h[i] = 0u;
        csync<NT_>();
        for (uint32_t i = c.tid; i < cnt; i += NT_) {
            const uint32_t key = keys[i];
            if (key >= lo && key <= hi) atomicAdd(&h[(key - lo) >> sh], 1u);
        }
        csync<NT_>();
        uint32_t hv[BPT], mine = 0u;
#pragma unroll
        for (int e = 0; e < BPT; e++) { hv[e] = (c.tid < HT) ? h[BPT * c.tid + e] : 0u; mine += hv[e]; }
        if (c.tid == 0) s->sel_bin = -1;
        uint32_t total;
        uint32_t ex = block_excl_scan_u32(c, mine, &total);
#pragma unroll
        for (int e = 0; e < BPT; e++) {
            if (r >= ex && r < ex + hv[e]) { s->sel_bin = BPT * c.tid + e; s->sel_pre = ex; }
            ex += hv[e];
        }
        csync<NT_>();
        const int bin = s->sel_bin;
        const uint32_t pre = s->sel_pre;
        csync<NT_>();
        if (bin < 0) { if (c.tid == 0) atomicExch(c.p->status, 300 + guard); return lo; }
        r -= pre;
        lo += (uint32_t)bin << sh;
        const uint32_t top = lo + ((1u << sh) - 1u);
        if (top < hi) hi = top;
    }
    return lo;
}

// The value of rank k (0-based) among the margins E[i] - x[i] of the whole image.  *zeros receives
// the exact number of zero margins when the first (windowed) round saw them (else 0xffffffff).
template <int NT_, int PPT>
__device__ __forceinline__ float select_margin(Cx<NT_> &c, uint32_t k, uint32_t lo_key,
                                               uint32_t *zeros) {
    SelScratch *ss = sel(c);
    uint32_t lo = lo_key, hi = 0u, r = k;
    int sh = 0;
    bool window = true;
    *zeros = 0xffffffffu;
    for (int guard = 0; guard < 6; guard++) {
        uint32_t *h = ss->hist[c.hround & 1u];
        if (window) hist_pass<NT_, PPT, true>(c, lo, hi, sh, h);
        else hist_pass<NT_, PPT, false>(c, lo, hi, sh, h);
        const Located L = locate(c, r);
        if (L.bin < 0) {
            if (c.tid == 0) atomicExch(c.p->status, 200 + guard);
            return 0.f;
        }
        uint32_t klo, khi;
        if (window) win_range((uint32_t)L.bin, lo_key, klo, khi);
        else {
            klo = lo + ((uint32_t)L.bin << sh);
            khi = min(hi, klo + ((1u << sh) - 1u));
        }
        if (window && L.bin == 0) *zeros = L.cnt;
        r -= L.pre;
        if (klo == khi) return __uint_as_float(klo);
        if (L.cnt <= (uint32_t)GCAP) {
            // gather the candidates of the bin into every CTA, in rank order of the CTAs
            if (c.tid == 0) c.ms->lcount = 0u;
            csync<NT_>();
            for_each_x<PPT>(c.tx, [&](int j, float xv) {
                const int i = j * NT_ + c.tid;
                if (i < c.n_t) {
                    const uint32_t key = __float_as_uint(__fsub_rn(c.s_E[i], xv));
                    if (key >= klo && key <= khi) {
                        const uint32_t pos = atomicAdd(&c.ms->lcount, 1u);
                        if (pos < (uint32_t)GCAP) ss->mylist[pos] = key;
                    }
                }
            });
            csync<NT_>();
            const int fb = L.bin & (SUBB - 1);
            const uint32_t mine = ss->fine[c.rank][fb];
            uint32_t off = 0;
            for (int p = 0; p < c.rank; p++) off += ss->fine[p][fb];
            if (c.G > 1) {
                for (uint32_t i = c.tid; i < (uint32_t)c.G * mine; i += NT_) {
                    const uint32_t peer = i / mine, w = i - peer * mine;
                    st_peer(peer_addr(&ss->gather[off + w], (int)peer), ss->mylist[w]);
                }
            } else {
                for (uint32_t w = c.tid; w < mine; w += NT_) ss->gather[w] = ss->mylist[w];
            }
            cluster_sync_compute(c);
            return __uint_as_float(local_select(c, ss->gather, L.cnt, r, klo, khi));
        }
        // heavy ties or a coarse (out-of-window) bin: narrow the bracket and bin again
        window = false;
        lo = klo; hi = khi;
        sh = 0;
        while (((hi - lo) >> sh) >= (uint32_t)NB2) sh++;
    }
    if (c.tid == 0) atomicExch(c.p->status, 299);
    return 0.f;
}

// ---------------------------------------------------------------------------------------------
// labels (weaklabel.cpp:108-112 + deeplab.py:123)
// ---------------------------------------------------------------------------------------------
// Exact per-pixel recomputation from the original scores (rare: an absent class may tie with or
// beat the best present class).
static __device__ __noinline__ int label_slow(const EStepParams *pp_, const Misc *ms, const float *planes,
                                              uint32_t mask, int b, int i, double shift) {
    const EStepParams &p = *pp_;
    const float *px = p.map + (long long)b * p.s0 + (long long)i * p.s2;
    float m = FLT_MAX;
    for (int k = 0; k < p.C; k++)
        if ((mask >> k) & 1u) { const float t = px[k]; if (t < m) m = t; }
    const float low = __fsub_rn(m, p.margin);
    float best = -INFINITY;
    int bi = 0, pp = 0;
    for (int k = 0; k < p.C; k++) {
        float v;
        if ((mask >> k) & 1u) {
            v = with_hist(__ldcg(planes + (size_t)pp * p.Npad + i), load_hist(ms, pp));
            pp++;
        } else {
            v = px[k];
            if (v > m) v = low;
        }
        const float fv = __double2float_rn(__dadd_rn((double)v, shift));
        if (fv > best) { best = fv; bi = k; }
    }
    return bi;
}

template <int NT_>
__device__ __forceinline__ void write_labels(Cx<NT_> &c, double shift) {
    const EStepParams &p = *c.p;
    uint8_t *out = p.labels + (size_t)c.b * p.N;
    const bool has_absent = c.P < p.C;
    const bool have_abs = has_absent && c.P == 1;          // suppressed absent maximum was kept
    const int nq = (c.n_t + 3) >> 2;
    for (int q = c.tid; q < nq; q += NT_) {
        const int idx = c.base + 4 * q;
        float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        float lowest[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
        int bi[4] = {0, 0, 0, 0};
        for (int pl = 0; pl < c.P; pl++) {
            const float4 v = __ldcg(reinterpret_cast<const float4 *>(c.planes + (size_t)pl * p.Npad + idx));
            float vv[4] = {v.x, v.y, v.z, v.w};
            const int hn = c.ms->hn[pl];
            for (int k = 0; k < hn; k++) {
                const float hb = c.ms->hbias[pl][k];
#pragma unroll
                for (int e = 0; e < 4; e++) vv[e] = __fadd_rn(vv[e], hb);
            }
            const int cls = c.ms->plist[pl];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float fv = __double2float_rn(__dadd_rn((double)vv[e], shift));
                if (fv > best[e]) { best[e] = fv; bi[e] = cls; }
                if (fv < lowest[e]) lowest[e] = fv;
            }
        }
        if (have_abs) {
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
                        bi[e] = label_slow(c.p, c.ms, c.planes, c.mask, c.b, idx + e, shift);
                }
            }
        } else if (has_absent) {
            // every suppressed absent score is <= the smallest ORIGINAL present score <= the smallest
            // present final score: an absent class can reach the maximum only if all present classes
            // round to the same float
#pragma unroll
            for (int e = 0; e < 4; e++)
                if (idx + e < p.N && lowest[e] >= best[e]) bi[e] = label_slow(c.p, c.ms, c.planes, c.mask, c.b, idx + e, shift);
        }
#pragma unroll
        for (int e = 0; e < 4; e++)
            if (idx + e < p.N) out[idx + e] = (uint8_t)bi[e];
    }
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
template <int NT_, int PPT>
__global__ void __launch_bounds__(NT_ + NPW * 32, 1) estep_v3_kernel(const __grid_constant__ EStepParams p) {
    typedef Layout<NT_> LY;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Cx<NT_> c;
    c.p = &p;
    c.tid = threadIdx.x;
    c.lane = c.tid & 31;
    c.warp = c.tid >> 5;
    c.G = p.G;
    c.rank = blockIdx.x % p.G;
    const int slot = blockIdx.x / p.G;
    unsigned char *staging;
    {
        unsigned char *q = smem_raw;
        c.s_E = reinterpret_cast<float *>(q);
        q += LY::A16((size_t)PPT * NT_ * 4);
        c.mbox = reinterpret_cast<uint32_t (*)[MAXG][MBW]>(q);
        q += LY::MBOX_BYTES;
        c.ms = reinterpret_cast<Misc *>(q);
        q += LY::A16(sizeof(Misc));
        c.uni = q;
        q += LY::uni_bytes();
        staging = q;
    }
    c.xround = 0;
    c.hround = 0;
    c.csyncs = 0;
    c.base = c.rank * p.TS;
    c.n_t = max(0, min(p.TS, p.N - c.base));
    if (c.tid == 0) {
        for (int d = 0; d < NST; d++) mbar_init(&c.ms->full_bar[d], 1);
        mbar_init(&c.ms->e0_bar, 1);
        for (int q = 0; q < 2; q++) {
            mbar_init(&c.ms->cmd_bar[q], 1);
            mbar_init(&c.ms->ready_bar[q], NPW);
            mbar_init(&c.ms->xbar[q], p.G);
        }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (c.warp == 0) tm_alloc<tm_cols(PPT)>(&c.ms->tmem_base);
    tm_fence_before();
    hw_cluster_sync();                                  // every CTA's barriers exist before any remote arrival
    tm_fence_after();
    c.tx = c.ms->tmem_base + ((uint32_t)(c.warp & 3) << 21) + (uint32_t)((c.warp >> 2) * PPT);   // lane 32*(warp%4), my columns
    if (c.tid >= NT_) {
        producer_main<NT_>(p, c.ms, staging, c.rank, slot, c.base, c.n_t);
        hw_cluster_sync();
        return;
    }
    const bool lead_thread = (c.rank == 0 && c.tid == 0);

    for (uint32_t seq = 0;; seq++) {
        // claim the image of sequence number seq + 1 (and seq when seq == 0) for the cluster and
        // hand it to the producers, which stream it while image seq is worked on
        trc(c, 0, (seq << 8) | 1u);
        if (c.rank == 0 && c.warp == 0) {
            int img0 = 0, img1 = 0;
            if (c.lane == 0) {
                if (seq == 0) {
                    const int t = atomicAdd(p.image_counter, 1);
                    img0 = (t < p.B) ? (p.sched ? p.sched[t] : t) : p.B;
                }
                const int t = atomicAdd(p.image_counter, 1);
                img1 = (t < p.B) ? (p.sched ? p.sched[t] : t) : p.B;
            }
            img0 = __shfl_sync(0xffffffffu, img0, 0);
            img1 = __shfl_sync(0xffffffffu, img1, 0);
            if (c.lane < c.G) {
                if (seq == 0) st_peer(peer_addr(&c.ms->cmd_img[0], c.lane), (uint32_t)img0);
                st_peer(peer_addr(&c.ms->cmd_img[(seq + 1) & 1u], c.lane), (uint32_t)img1);
            }
        }
        cluster_sync_compute(c);
        if (c.tid == 0) {
            if (seq == 0) mbar_arrive(&c.ms->cmd_bar[0]);
            mbar_arrive(&c.ms->cmd_bar[(seq + 1) & 1u]);
        }
        trc(c, 0, (seq << 8) | 2u);
        c.b = *reinterpret_cast<volatile int *>(&c.ms->cmd_img[seq & 1u]);
        if (c.b >= p.B) break;
        c.planes = p.planes + (size_t)(slot * 2 + (int)(seq & 1u)) * (size_t)(p.C + 2) * p.Npad;
        c.absmax = c.planes + (size_t)p.C * p.Npad;
        c.absidx = p.absent_idx + (size_t)(slot * 2 + (int)(seq & 1u)) * p.Npad;
        unsigned long long *prof = p.prof ? p.prof + (size_t)c.b * 8 : nullptr;
        if (prof && lead_thread) prof[0] = gtimer();
        const ImageMeta *im = p.meta + c.b;
        c.mask = im->mask;
        c.P = im->P;
        c.first_absent = im->first_absent;
        if (c.tid < MAXC) {
            c.ms->plist[c.tid] = im->plist[c.tid];
            c.ms->zlb[c.tid] = 0u;
            c.ms->hn[c.tid] = 0;
        }
        // 1. the planes of my tile are in the scratch (written by my producers together with E0,
        //    the zero-margin counts and the largest margin): E0 -> shared memory
        mbar_wait(&c.ms->ready_bar[seq & 1u], (seq >> 1) & 1u);
        trc(c, 0, (seq << 8) | 3u);
        trc(c, 1, (uint32_t)c.b);
        if (c.tid == 0) {
            const uint32_t bytes = (uint32_t)((c.n_t + 3) & ~3) * 4u;
            if (bytes) {
                asm volatile("fence.proxy.async;\n" ::: "memory");   // generic accesses to s_E (previous image) first
                mbar_expect_tx(&c.ms->e0_bar, bytes);
                const unsigned char *src = reinterpret_cast<const unsigned char *>(c.planes + (size_t)(p.C + 1) * p.Npad + c.base);
                for (uint32_t o = 0; o < bytes; o += 32768u)
                    bulk_g2s(reinterpret_cast<unsigned char *>(c.s_E) + o, src + o, min(32768u, bytes - o), &c.ms->e0_bar);
            } else {
                mbar_arrive(&c.ms->e0_bar);
            }
        }
        mbar_wait(&c.ms->e0_bar, seq & 1u);
        if (prof && lead_thread) prof[1] = gtimer();

        trc(c, 0, (seq << 8) | 4u);
        // 2. one exchange: zero counts, largest margin, tile totals (prediction for mean_max_0)
        double tile_total;
        double cs0;
        const double cpre0 = chunk_prefix(c, &tile_total, &cs0);
        if (c.tid < 5) {
            uint32_t v = 0u;
            for (int w = 0; w < NPW; w++) {
                const uint32_t t = c.ms->pstat[seq & 1u][w][c.tid];
                v = (c.tid == 4) ? max(v, t) : v + t;
            }
            c.ms->outbox[c.tid] = v;
        }
        if (c.tid == 0) {
            const unsigned long long tb = (unsigned long long)__double_as_longlong(tile_total);
            c.ms->outbox[5] = (uint32_t)tb;
            c.ms->outbox[6] = (uint32_t)(tb >> 32);
        }
        float dmax_bound;
        double before0 = 0.0;
        {
            const uint32_t (*box)[MBW] = exchange(c, 7);
            uint32_t mx = 0u;
            for (int q = 0; q < c.G; q++) {
                mx = max(mx, box[q][4]);
                if (q < c.rank)
                    before0 += __longlong_as_double((long long)((unsigned long long)box[q][5] | ((unsigned long long)box[q][6] << 32)));
            }
            dmax_bound = __uint_as_float(mx);
            if (c.tid < EAGER_P && c.tid < c.P) {
                uint32_t z = 0u;
                for (int q = 0; q < c.G; q++) z += box[q][c.tid];
                c.ms->zlb[c.tid] = z;
            }
        }
        csync<NT_>();

        trc(c, 0, (seq << 8) | 5u);
        // 3. mean_max_0 (weaklabel.cpp:69-80): float running sum in pixel order
        const float sum0 = exact_sum<float, NT_>(c, cpre0, cs0, before0);
        const float mean0 = __fdiv_rn(sum0, (float)p.N);
        if (prof && lead_thread) prof[2] = gtimer();

        trc(c, 0, (seq << 8) | 6u);
        // 4. iterations (weaklabel.cpp:82-101)
        // The zero-margin lower bounds only pay when quotas leave slack (a class that keeps more pixels
        // than its quota is skipped without a selection).  When the quotas of the label set add up to
        // the whole image every step selects anyway: the per-fold exchange that maintains them is dropped.
        bool track_zeros;
        {
            long long quota = 0;
            for (int k = 0; k < c.P; k++) quota += (c.ms->plist[k] == 0) ? p.nth_bg : p.nth_fg;
            track_zeros = quota < (long long)p.N - (p.N >> 6);
        }
        const uint8_t *ord = p.order + (size_t)c.b * p.num_iter * p.C;
        uint32_t nsel = 0;
        for (int j = 0; j < p.num_iter; j++) {
            for (int t = 0; t < c.P; t++) {
                const int plane = ord[j * p.C + t];
                const int cls = c.ms->plist[plane];
                const uint32_t k = (uint32_t)((cls == 0) ? p.nth_bg : p.nth_fg);      // weaklabel.cpp:88
                float bias = 0.f;
                if (k >= c.ms->zlb[plane]) {        // else: the rank falls among the zero margins, bias 0
                    nsel++;
                    trc(c, 0, (seq << 8) | 7u);
                    trc(c, 2, ((uint32_t)j << 8) | (uint32_t)t);
                    // the class plane, with the biases the class received so far re-applied in order
                    // (weaklabel.cpp:90-92), goes into tensor memory: element j of my thread = tile
                    // pixel j * NT + tid.  The three passes of the selection read it from there.
                    float *plc = c.planes + (size_t)plane * p.Npad + c.base;
                    const int hn = c.ms->hn[plane];
                    {
                        constexpr int NC = PPT / 8;
#pragma unroll
                        for (int ch = 0; ch < NC; ch++) {
                            float v[8];
#pragma unroll
                            for (int e = 0; e < 8; e++) {
                                const int i = (ch * 8 + e) * NT_ + c.tid;
                                v[e] = (i < c.n_t) ? __ldcg(plc + i) : 0.f;
                            }
                            for (int e2 = 0; e2 < hn; e2++) {
                                const float hb = c.ms->hbias[plane][e2];
#pragma unroll
                                for (int e = 0; e < 8; e++) v[e] = __fadd_rn(v[e], hb);
                            }
                            tm_st8(c.tx + ch * 8, v);
                        }
#pragma unroll
                        for (int j = NC * 8; j < PPT; j++) {
                            const int i = j * NT_ + c.tid;
                            float v = (i < c.n_t) ? __ldcg(plc + i) : 0.f;
                            for (int e2 = 0; e2 < hn; e2++) v = __fadd_rn(v, c.ms->hbias[plane][e2]);
                            tm_st1(c.tx + j, v);
                        }
                        tm_wait_st();
                    }
                    uint32_t e_top = (__float_as_uint(dmax_bound) >> 23) + 1u;
                    if (e_top > 254u) e_top = 254u;
                    const uint32_t lo_key = (e_top > (uint32_t)WIN_BINADES ? e_top - WIN_BINADES : 0u) << 23;
                    uint32_t zeros;
                    bias = select_margin<NT_, PPT>(c, k, lo_key, &zeros);
                    csync<NT_>();
                    if (bias != 0.f) {
                        // fold: x += bias (weaklabel.cpp:90-92), E = max(E, x) (:94-99)
                        const bool rewrite = (hn == HCAP);
                        uint32_t gained = 0u, znew = 0u;
                        for_each_x<PPT>(c.tx, [&](int q, float xv) {
                            const int i = q * NT_ + c.tid;
                            if (i < c.n_t) {
                                const float xn = __fadd_rn(xv, bias);
                                if (rewrite) __stcg(plc + i, xn);
                                const float E = c.s_E[i];
                                if (xn > E) { c.s_E[i] = xn; gained++; }
                                if (xn >= E) znew++;
                            }
                        });
                        if (track_zeros) {
                            block_reduce_to_outbox(c, gained, znew, 0u, 0u, 0u, 0);
                            const uint32_t (*box)[MBW] = exchange(c, 2);
                            uint32_t gsum = 0u, zsum = 0u;
                            for (int q = 0; q < c.G; q++) { gsum += box[q][0]; zsum += box[q][1]; }
                            if (c.tid < c.P) {
                                const uint32_t z = c.ms->zlb[c.tid];
                                c.ms->zlb[c.tid] = (c.tid == plane) ? zsum : (z > gsum ? z - gsum : 0u);
                            }
                        } else if (c.tid < c.P) {
                            c.ms->zlb[c.tid] = 0u;                      // a valid lower bound without the exchange
                        }
                        if (c.tid == 0) {
                            if (rewrite) { c.ms->hn[plane] = 0; }
                            else { c.ms->hbias[plane][hn] = bias; c.ms->hn[plane] = hn + 1; }
                        }
                        dmax_bound = __fadd_rn(dmax_bound, bias);
                    } else if (zeros != 0xffffffffu) {
                        if (c.tid == 0) c.ms->zlb[plane] = zeros;       // exact count of zero margins
                    }
                    csync<NT_>();
                }
                if (lead_thread) p.bias_out[((size_t)c.b * p.num_iter + j) * p.C + t] = bias;
            }
        }
        if (prof && lead_thread) prof[3] = gtimer();

        trc(c, 0, (seq << 8) | 8u);
        // 5. mean_max_1 (weaklabel.cpp:102-106), shift (:110), labels
        csync<NT_>();
        double before1 = 0.0;
        double cs1;
        const double cpre1 = chunk_prefix(c, &tile_total, &cs1);
        if (c.tid == 0) {
            const unsigned long long tb = (unsigned long long)__double_as_longlong(tile_total);
            c.ms->outbox[0] = (uint32_t)tb;
            c.ms->outbox[1] = (uint32_t)(tb >> 32);
        }
        {
            const uint32_t (*box)[MBW] = exchange(c, 2);
            for (int q = 0; q < c.rank; q++)
                before1 += __longlong_as_double((long long)((unsigned long long)box[q][0] | ((unsigned long long)box[q][1] << 32)));
        }
        const double sum1 = exact_sum<double, NT_>(c, cpre1, cs1, before1);
        const double mean1 = __ddiv_rn(sum1, (double)p.N);
        const double shift = __dsub_rn((double)mean0, mean1);
        if (lead_thread) {
            p.mean0[c.b] = mean0;
            p.mean1[c.b] = mean1;
            p.shift[c.b] = shift;
        }
        if (prof && lead_thread) prof[4] = gtimer();
        trc(c, 0, (seq << 8) | 9u);
        if (p.labels) write_labels(c, shift);
        csync<NT_>();
        if (prof && lead_thread) {
            prof[5] = gtimer();
            prof[6] = nsel;
            prof[7] = (unsigned long long)c.P;
        }
    }
    trc(c, 0, 0xEEEEu);
    csync<NT_>();
    if (c.warp == 0) tm_dealloc<tm_cols(PPT)>(c.ms->tmem_base);
    hw_cluster_sync();                                  // nobody leaves while a peer may still address its shared memory
    trc(c, 0, 0xFFFFu);
}

template <int NT_, int PPT>
static cudaError_t launch_t(const EStepParams &p, int n_clusters, cudaStream_t stream, int *max_clusters) {
    const size_t smem = Layout<NT_>::total(PPT);
    cudaError_t e = cudaFuncSetAttribute(estep_v3_kernel<NT_, PPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(estep_v3_kernel<NT_, PPT>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((max_clusters ? 64 : n_clusters) * p.G));
    cfg.blockDim = dim3(NT_ + NPW * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.G;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (max_clusters) {
        int n = 0;
        e = cudaOccupancyMaxActiveClusters(&n, estep_v3_kernel<NT_, PPT>, &cfg);
        *max_clusters = (e == cudaSuccess) ? n : 0;
        return e;
    }
    return cudaLaunchKernelEx(&cfg, estep_v3_kernel<NT_, PPT>, p);
}

static cudaError_t dispatch(const EStepParams &p, int ppt, int n_clusters, cudaStream_t stream, int *maxc) {
    if (ppt == V3_PPT_MAX) return launch_t<V3_NT, V3_PPT_MAX>(p, n_clusters, stream, maxc);
    if (ppt == V3_PPT_BIG) return launch_t<V3_NT, V3_PPT_BIG>(p, n_clusters, stream, maxc);
    if (ppt == V3_PPT_MID) return launch_t<V3_NT, V3_PPT_MID>(p, n_clusters, stream, maxc);
    if (ppt == V3_PPT_SMALL) return launch_t<V3_NT, V3_PPT_SMALL>(p, n_clusters, stream, maxc);
    return cudaErrorInvalidValue;
}

}  // namespace v3

// Geometry of the v3 kernel for an image of N pixels with C classes and a given cluster size.
static bool v3_try(int N, int C, int G, int tune_TS, V3Config *out) {
    const int nt = V3_NT;
    const int cap = V3_PPT_MAX * nt;
    if (G < 1 || G > V3_MAXG) return false;
    int TS = tune_TS ? tune_TS : (N + G - 1) / G;
    TS = (TS + 3) / 4 * 4;
    if (TS < 4) TS = 4;
    if (TS > cap || (long long)TS * G < N) return false;
    const int need = (TS + nt - 1) / nt;
    out->nt = nt;
    out->ppt = need <= V3_PPT_SMALL ? V3_PPT_SMALL : (need <= V3_PPT_MID ? V3_PPT_MID : (need <= V3_PPT_BIG ? V3_PPT_BIG : V3_PPT_MAX));
    out->G = G;
    out->TS = TS;
    out->L = need | 1;                                   // odd: conflict-free chunk walks
    int cpx = (V3_STAGE_BYTES - 32) / (4 * C);
    if (cpx < 1) return false;
    out->cpx = cpx;
    EStepParams p = {};
    p.G = G;
    int maxc = 0;
    if (v3::dispatch(p, out->ppt, 0, nullptr, &maxc) != cudaSuccess || maxc <= 0) {
        cudaGetLastError();
        return false;
    }
    out->max_clusters = maxc;
    return true;
}

// tune_G = 0: automatic -- the smallest power-of-two cluster that holds the image; large batches of
// large images also try clusters of 9 (non-portable size): on B200 as many of them are co-resident
// as clusters of 8 (15), so 135 instead of 120 SMs work and every thread holds fewer pixels.
bool v3_configure(int N, int C, int B, int sms, int tune_G, int tune_TS, V3Config *out) {
    if (tune_G) return v3_try(N, C, tune_G, tune_TS, out);
    const int cap = V3_PPT_MAX * V3_NT;
    int G = 1;
    while (G < 8 && (N + G - 1) / G > cap) G *= 2;
    while (G < 8 && (long long)B * G * 2 <= sms && N / (2 * G) >= 4096) G *= 2;
    if (!v3_try(N, C, G, tune_TS, out)) return false;
    if (G == 8 && tune_TS == 0) {
        V3Config alt;
        if (v3_try(N, C, 9, 0, &alt) && alt.max_clusters >= out->max_clusters && B >= alt.max_clusters) *out = alt;
    }
    return true;
}

cudaError_t launch_estep_v3(const EStepParams &p, const V3Config &cfg, int n_clusters, cudaStream_t stream) {
    return v3::dispatch(p, cfg.ppt, n_clusters, stream, nullptr);
}

}  // namespace wl
