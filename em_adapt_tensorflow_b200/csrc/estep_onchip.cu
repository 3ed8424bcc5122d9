// estep_onchip.cu -- the E-step kernel of the common configuration, fourth generation (sm_100a).
//
// Contract (unchanged from the third generation): one thread-block cluster per image, persistent,
// images claimed from a counter; suppress_others = true with a margin >= 0, contiguous pixels, the
// image fits the cluster.  Warp-specialised: NT compute threads + NPW producer warps per CTA.  The
// producers stream the 84-byte score vectors of image t+1 through a bulk-copy ring while the compute
// threads run the selection chain of image t (weaklabel.cpp:82-101).
//
// What is new, all of it on the compute side (the kernel was latency-bound there):
//   * QUAD OWNERSHIP.  A thread owns whole quads of four consecutive pixels (quad q = k * NT + tid):
//     the running max E ("extrenum", weaklabel.cpp:67-77) is read and written as float4 from shared
//     memory, a class plane moves as float4 between the L2 scratch and tensor memory, and four
//     consecutive TMEM columns are one tcgen05.ld/st.x4.  Padding pixels carry E = 0, x = -inf, so that
//     no pass needs a validity test per element (their margin is +inf: the largest key, never selected).
//   * TWO PLANES RESIDENT IN TENSOR MEMORY (512 columns = two slots): label sets of one or two classes
//     (63 % of VOC-style batches) never reload a plane and never replay a bias history; larger sets use
//     the two slots as a cache with write-back, so a plane in the scratch is always current.
//   * A BRANCH-FREE HISTOGRAM PASS (10 instructions per margin instead of 28): 2048 bins = zero bin,
//     one bin below the window, 128 bins per binade over the 15 binades under the largest possible
//     margin; zero margins go to a private counter per warp instead of one contended bin.
//   * EXCHANGES BY st.async: every CTA stores its words straight into the mailboxes of its peers with the
//     completion counted in bytes on the receiver's mbarrier -- no release fence, no second barrier.
//   * THE ZERO-MARGIN BOUND IS TIGHT: a fold only takes pixels away from the other classes where the
//     folded class was NOT the maximum before (plus a bound on exact ties), so a class that still owns
//     more than its quota is skipped without a selection (the previous bound counted every pixel the
//     folded class already owned and forced one or two useless selections per image).
//   * ORDER-EXACT SUMS BY DIGEST (weaklabel.cpp:69-80, 102-106): every CTA condenses its tile into a
//     few composite functions (seqsum.cuh) plus the terms of the rare binade-crossing chunks and pushes
//     this digest to CTA 0 with one bulk copy; two warps of CTA 0 walk the float and the double sum at
//     the same time out of their own shared memory.  mean_max_0's digest is formed at the start of the
//     image (E0 is still in shared memory) and costs no cluster round trip of its own.
//   * LABELS IN FLOAT: argmax over the final scores; only pixels whose best two scores are within the
//     rounding distance of the re-centring shift (weaklabel.cpp:108-112) take the exact double path.
//   * LABELS-ONLY MODE: the shift is bounded WITHOUT the two order-exact sums, every provably safe pixel is
//     labelled at once (pass A, out through shared memory in 16-byte stores), and the sums are formed only for
//     the images in which a pixel stays open (pass B decides those exactly).
// and on the producer side:
//   * a LIGHT path for label sets of one or two classes (both planes stay in tensor memory): the producers only
//     de-interleave, E0 and the statistics are formed by the compute threads while the planes go into TMEM;
//   * ring stages of 96 pixels (three full rounds of 32 lanes), every interior chunk ONE bulk copy of the
//     16-byte-aligned span around it (cp.async head / tail floats only at the two ends of the tensor);
//   * FUSED BILINEAR UP-SAMPLING (network.py:40): with EStepParams::up_src the producers stream nothing and
//     interpolate the present classes from the low-resolution tensor (prod_upsampled, upsample.cuh).
#include <float.h>
#include <math.h>
#include "estep_kernel.cuh"
#include "seqsum.cuh"
#include "upsample.cuh"

namespace wl {
namespace oc {

constexpr int NB = 2048;                // histogram bins
constexpr int NSB = 64;                 // super bins exchanged through the mailboxes
constexpr int SUBB = NB / NSB;          // 32 fine bins per super bin
constexpr int WIN_SH = 16;              // 128 bins per binade inside the window (256 bins over 7 binades was measured:
constexpr int WIN_BINADES = 15;         // slower -- more ranks fall below the window and cost a second round)
constexpr int GCAP = 1024;              // candidates gathered into every CTA
constexpr int RCMAX = 256;              // up to here the local selection ranks by counting
constexpr int MBW = 72;                 // mailbox words per CTA per exchange (a multiple of 4)
constexpr int HT = 512;                 // threads that scan a 2048-bin histogram (4 bins each)
constexpr int MAXG = OC_MAXG;
constexpr int NPW = 4;                  // producer warps per CTA
constexpr int NST = 8;                  // stages of the staging ring (a multiple of NPW: a stage always belongs
                                        // to the same warp, which then sees its mbarrier phases in order)
constexpr int STAGE_BYTES = 8096;              // 96 pixels of 21 classes + 32 bytes of alignment slack: three full rounds of 32 lanes
constexpr int EAGER_P = 4;              // planes whose zero counts are formed by the producers
constexpr int DIG_SEG = 8;              // segments of a pushed digest
constexpr int DIG_CROSS = 3;            // binade-crossing chunks whose terms travel with the digest
constexpr int DIG_L = 44;               // terms per crossing chunk (>= the longest chunk)
constexpr int BIG_SEG = 64;             // segments of the local digest
constexpr int ORD_S = 512;              // bytes of the order table kept in shared memory (num_iter * C <= 512, else global)
constexpr unsigned long long WAIT_LIMIT_NS = 4000000000ull;   // a wait gives up after 4 s (status + trap instead of a hung GPU)
static_assert(NST % NPW == 0, "a stage must always belong to the same producer warp");
static_assert(MBW % 4 == 0 && NSB <= MBW, "mailbox rows are moved as 16-byte vectors");

// ---------------------------------------------------------------------------------------------
// PTX helpers
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
__device__ __forceinline__ uint32_t ld_peer(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared::cluster.u32 %0, [%1];\n" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
// remote store whose completion is counted (in bytes) on an mbarrier of the SAME remote CTA
__device__ __forceinline__ void st_async_u32(uint32_t addr, uint32_t v, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.u32 [%0], %1, [%2];\n"
                 ::"r"(addr), "r"(v), "r"(bar) : "memory");
}
__device__ __forceinline__ void st_async_u64(uint32_t addr, unsigned long long v, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.u64 [%0], %1, [%2];\n"
                 ::"r"(addr), "l"(v), "r"(bar) : "memory");
}
__device__ __forceinline__ void st_async_v4(uint32_t addr, uint4 v, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.u32 [%0], {%1,%2,%3,%4}, [%5];\n"
                 ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(bar) : "memory");
}
// shared memory of this CTA -> shared memory of a peer, completion on the peer's mbarrier
__device__ __forceinline__ void bulk_s2peer(uint32_t peer_dst, const void *src, uint32_t bytes, uint32_t peer_bar) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                 ::"r"(peer_dst), "r"(smem_u32(src)), "r"(bytes), "r"(peer_bar) : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {     // arrive + expect
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {           // release.cta
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_peer(uint32_t cluster_addr) {   // release at cluster scope
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
    return t;
}
// Every wait is bounded: a lost arrival ends in a trap with a status code instead of a hung GPU.
__device__ __noinline__ void wait_failed(const EStepParams *p, int code) {
    atomicExch(p->status, code);
    if (p->trail) p->trail[blockIdx.x * 32 + 30] = (unsigned int)code | (threadIdx.x << 16);
    __threadfence_system();
    __trap();
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity, const EStepParams *status, int code) {  // acquire.cta
    const uint32_t a = smem_u32(bar);
    uint32_t ok;
    unsigned long long t0 = 0ull;
    for (;;) {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
        if (ok) break;
        const unsigned long long t = gtimer();          // a failed try_wait has already slept for a while
        if (t0 == 0ull) t0 = t;
        else if (t - t0 > WAIT_LIMIT_NS) wait_failed(status, code);
    }
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t *bar, uint32_t parity, const EStepParams *status, int code) {  // acquire.cluster
    const uint32_t a = smem_u32(bar);
    uint32_t ok;
    unsigned long long t0 = 0ull;
    for (;;) {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
        if (ok) break;
        const unsigned long long t = gtimer();
        if (t0 == 0ull) t0 = t;
        else if (t - t0 > WAIT_LIMIT_NS) wait_failed(status, code);
    }
}
// global -> shared bulk copy (16-byte aligned on both sides, size a multiple of 16), completion on bar
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;\n" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;\n"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}
// ---- tensor memory as a per-thread array: 32x32b shape, thread i of the warp <-> TMEM lane
// (32 * (warp % 4) + i), one 32-bit column per element; four consecutive columns = one quad of pixels.
__device__ __forceinline__ void tm_alloc512(uint32_t *smem_result) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(smem_u32(smem_result)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tm_dealloc512(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(taddr) : "memory");
}
__device__ __forceinline__ void tm_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tm_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ float4 tm_ld4(uint32_t taddr) {
    uint32_t r0, r1, r2, r3;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(taddr) : "memory");
    tm_wait_ld();
    return make_float4(__uint_as_float(r0), __uint_as_float(r1), __uint_as_float(r2), __uint_as_float(r3));
}
__device__ __forceinline__ void tm_ld8(uint32_t taddr, float4 &a, float4 &b) {
    uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7) : "r"(taddr) : "memory");
    tm_wait_ld();
    a = make_float4(__uint_as_float(r0), __uint_as_float(r1), __uint_as_float(r2), __uint_as_float(r3));
    b = make_float4(__uint_as_float(r4), __uint_as_float(r5), __uint_as_float(r6), __uint_as_float(r7));
}
__device__ __forceinline__ void tm_st4(uint32_t taddr, float4 v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};\n"
                 ::"r"(taddr), "r"(__float_as_uint(v.x)), "r"(__float_as_uint(v.y)), "r"(__float_as_uint(v.z)),
                   "r"(__float_as_uint(v.w)) : "memory");
}
// columns of one plane slot: 7 warps share a lane quarter, 4 * QPT columns each
__host__ __device__ constexpr int slot_cols(int qpt) { return 28 * qpt <= 256 ? 256 : 512; }
__host__ __device__ constexpr int n_slots(int qpt) { return 512 / slot_cols(qpt); }

// ---------------------------------------------------------------------------------------------
// shared memory
// ---------------------------------------------------------------------------------------------
// One segment of a tile's digest: the action on the running sum of the chunks [head, tail] of the tile
// (chunk t = terms [t * L, (t + 1) * L)).  kind 1: m -> m + a[m & 1] while the exponent field is e
// (seqsum.cuh); kind 0: add the terms one by one (cslot >= 0: they travel with the digest).
struct alignas(16) DigEntry {
    unsigned long long e;
    long long a0, a1;
    int kind;
    int cslot;
    int nterms;
    uint16_t head, tail;
    uint32_t pad[2];
};
static_assert(sizeof(DigEntry) == 48, "DigEntry layout");
struct alignas(16) Digest {                 // what a CTA pushes to CTA 0 (fixed size)
    int nseg, ncross, pad0, pad1;
    DigEntry seg[DIG_SEG];
    float cross[DIG_CROSS][DIG_L];
};
static_assert(sizeof(Digest) % 16 == 0, "digests move by bulk copy");
struct alignas(16) BigDigest {              // the local digest of a tile
    int nseg, total, pad0, pad1;
    DigEntry seg[BIG_SEG];
};

struct alignas(16) Misc {
    alignas(16) uint32_t outbox[2][MBW];    // alternating with the exchanges
    double wscr_d[32];
    uint32_t wscr_u[32];
    uint32_t wred[32][8];
    uint32_t zlb[MAXC];                 // zero-margin count per plane: exact or a lower bound
    uint64_t full_bar[NST];             // producers: stage filled (bulk-copy bytes landed)
    uint64_t e0_bar;                    // compute: E0 of my tile landed in shared memory
    uint64_t cmd_bar[2];                // compute -> producers: image of sequence parity q may be streamed
    uint64_t ready_bar[2];              // producers -> compute: planes of sequence parity q are in the scratch
    uint64_t xbar[2];                   // exchanges: bytes of the peers' st.async (alternating)
    uint64_t sbar[2];                   // plain cluster sync of the compute threads (alternating)
    uint64_t dig_bar;                   // CTA 0: the peers' digests landed
    uint64_t res_bar;                   // the two sums landed
    unsigned long long sumres[2];       // [0] float sum bits, [1] double sum bits
    uint32_t tmem_base;                 // tcgen05.alloc result
    uint32_t pstat[2][NPW][8];          // producers -> compute: zero counts [0..3], largest margin [4], ties [5], sum of E0 [6]
    int cmd_img[2];                     // image index of the command (>= B: stop)
    double bc_d;
    uint32_t bc_u;
    int sel_S;
    uint32_t sel_preS;
    int sel_bin;
    uint32_t sel_pre, sel_cnt;
    uint32_t sel_key;
    uint32_t lcount;
    uint32_t wcnt[32];
    uint32_t wfirst[32];
    uint8_t plist[MAXC];
    uint8_t pplist[NPW][MAXC];          // the producers' copy (image t+1)
    uint8_t ord_s[ORD_S];               // class order of the image in work (a skipped step must not cost a global load)
#ifdef WL_SELPROF
    unsigned long long tacc[10], tlast; // -DWL_SELPROF: where a selection's time goes (lead thread, ns)
#endif
};
// -DWL_SELPROF (development builds only, WEAKLABEL_NVCC_EXTRA=-DWL_SELPROF): lap timer of the lead thread inside
// select / fold (slot 9 = discard) and the producers' time per image; read by tools/phase_report.py
#ifdef WL_SELPROF
#define SELP(c, i) do { if ((c).p->prof && (c).rank == 0 && (c).tid == 0) { const unsigned long long t_ = gtimer(); \
        (c).msp()->tacc[i] += t_ - (c).msp()->tlast; (c).msp()->tlast = t_; } } while (0)
#else
#define SELP(c, i) do { } while (0)
#endif

constexpr int NBX = NB + 32;             // + one private zero-margin counter per warp
struct SelScratch {
    alignas(16) uint32_t hist[2][NBX];  // rotating per histogram round; peers read 32 fine bins
    uint32_t lhist[NB];                 // CTA-local selection among gathered candidates
    alignas(16) uint32_t gather[GCAP + 4];
    uint32_t mylist[GCAP];
    uint32_t fine[MAXG][SUBB];
};

template <typename T> struct ScanItem {     // what the segmented scan carries
    ChunkFn<T> f;
    int hidx;
    int flag;
};
// Aliases SelScratch.  Only `inbound` (written in CTA 0 at the END of an image) may overlap the two
// histograms, which the selections expect to be zero: they are cleared at the start of every image.
struct alignas(16) SumScratch {
    Digest inbound[MAXG][2];            // CTA 0: [tile][0 float, 1 double]
    BigDigest big;
    ScanItem<double> wagg[32];
    ChunkFn<double> wlast[32];
};
static_assert(sizeof(Digest) * MAXG * 2 >= sizeof(uint32_t) * 2 * NBX, "the local digest must not overlap the histograms");

template <int NT_> struct Layout {
    static __host__ __device__ constexpr size_t A16(size_t x) { return (x + 15) / 16 * 16; }
    static __host__ __device__ constexpr size_t uni_bytes() {
        size_t b = A16(sizeof(SelScratch)), c = A16(sizeof(SumScratch));
        return b > c ? b : c;
    }
    static constexpr size_t MBOX_BYTES = 2 * MAXG * MBW * 4;
    static constexpr size_t STAGING_BYTES = (size_t)NST * STAGE_BYTES;
    static size_t total(int se_floats) {
        return A16((size_t)se_floats * 4) + MBOX_BYTES + A16(sizeof(Misc)) + uni_bytes() + 2 * A16(sizeof(Digest)) + STAGING_BYTES;
    }
};

__device__ __forceinline__ unsigned char *smem0() {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    return smem_raw;
}

template <int NT_> struct Cx {
    const EStepParams *p;
    int tid, lane, warp, rank, G;
    float *planes;        // scratch of the image in work: [C+2][Npad]
    float *absmax;        // plane C of the scratch (single foreground class only)
    uint8_t *absidx;
    // Shared memory is addressed as (dynamic shared base + byte offset), never through pointers kept in
    // this struct: if the compiler leaves the struct in local memory, a pointer loaded back from there is
    // a GENERIC pointer and every LDS / ATOMS of the kernel degrades to a generic LD / ATOM.
    uint32_t o_mbox, o_ms, o_uni, o_outd;
    __device__ __forceinline__ float *sE() const { return reinterpret_cast<float *>(smem0()); }
    __device__ __forceinline__ uint32_t (*mboxp() const)[MAXG][MBW] { return reinterpret_cast<uint32_t (*)[MAXG][MBW]>(smem0() + o_mbox); }
    __device__ __forceinline__ Misc *msp() const { return reinterpret_cast<Misc *>(smem0() + o_ms); }
    __device__ __forceinline__ unsigned char *unip() const { return smem0() + o_uni; }
    __device__ __forceinline__ Digest *outdp() const { return reinterpret_cast<Digest *>(smem0() + o_outd); }     // [2]: my float and double digests (pushed to CTA 0)
    int base, n_t, TQ;    // my tile: pixels [base, base + n_t), TQ = quads that hold a pixel
    uint32_t xround;      // exchanges done (mailbox / outbox / barrier parity), identical in every CTA
    uint32_t hround;      // histogram rounds done (hist buffer parity)
    uint32_t csyncs;      // plain cluster syncs done
    uint32_t n_e0, n_dig, n_res;   // uses of the E0 / digest / result barriers (their parity; identical in every CTA)
    uint32_t tm;          // TMEM address of my lane quarter, my warp's first column of slot 0
    // per image
    int b, P;
    uint32_t mask;
    int first_absent;
    int res0, res1;       // plane resident in each TMEM slot (-1: none); scalars: an indexed array would
    bool dirty0, dirty1;  // push the whole context into local memory
    int mru;
};

// progress words of every CTA in mapped host memory (WEAKLABEL_TRAIL=1): a debugging aid of the bring-up,
// compiled in with -DWL_TRAIL only (the test for the pointer alone cost ~1.5 % of the issue slots)
template <int NT_> __device__ __forceinline__ void trc(const Cx<NT_> &c, int k, uint32_t v) {
#ifdef WL_TRAIL
    if (c.p->trail && c.tid == 0) c.p->trail[blockIdx.x * 32 + k] = v;
#endif
}
template <int NT_> __device__ __forceinline__ SelScratch *sel(Cx<NT_> &c) { return reinterpret_cast<SelScratch *>(c.unip()); }
template <int NT_> __device__ __forceinline__ SumScratch *sums(Cx<NT_> &c) { return reinterpret_cast<SumScratch *>(c.unip()); }

// Plain cluster-wide barrier of the COMPUTE threads: every CTA arrives on the barrier of every CTA
// (release at cluster scope after a CTA barrier), then waits on its own.  Two barriers alternate.
template <int NT_>
__device__ __forceinline__ void cluster_sync_compute(Cx<NT_> &c) {
    csync<NT_>();
    if (c.G > 1) {
        uint64_t *bar = &c.msp()->sbar[c.csyncs & 1u];
        if (c.tid < c.G) mbar_arrive_peer(peer_addr(bar, c.tid));
        mbar_wait_cluster(bar, (c.csyncs >> 1) & 1u, c.p, 401);
        c.csyncs++;
    }
}

// The outbox of the current exchange (written by any compute threads before exchange()).
template <int NT_> __device__ __forceinline__ uint32_t *outbox(Cx<NT_> &c) { return c.msp()->outbox[c.xround & 1u]; }

// Every CTA hands the first n words of its outbox to every CTA of the cluster (itself included);
// returns the mailbox that now holds them: box[p][w] = word w of CTA p.
//
// st.async path: thread 0 arms the barrier (one arrival + the bytes expected from all G senders) BEFORE
// the CTA barrier after which this CTA sends, so no peer can complete this exchange and send data of
// the next one (which goes to the other barrier / mailbox / outbox) before this barrier is armed.
template <int NT_>
__device__ __forceinline__ const uint32_t (*exchange(Cx<NT_> &c, int n))[MBW] {
    Misc *s = c.msp();
    const uint32_t par = c.xround & 1u;
    if (c.G > 1) {
        const int nv = (n + 3) >> 2;
        if (c.p->flags & OC_F_CLASSIC_XCHG) {
            csync<NT_>();
            for (int i = c.tid; i < c.G * n; i += NT_) {
                const int peer = i / n, w = i - peer * n;
                st_peer(peer_addr(&c.mboxp()[par][c.rank][w], peer), s->outbox[par][w]);
            }
            cluster_sync_compute(c);
        } else {
            if (c.tid == 0) mbar_expect_tx(&s->xbar[par], (uint32_t)(c.G * nv * 16));
            csync<NT_>();
            for (int i = c.tid; i < c.G * nv; i += NT_) {
                const int peer = i / nv, v = i - peer * nv;
                const uint4 w = *reinterpret_cast<const uint4 *>(&s->outbox[par][4 * v]);
                st_async_v4(peer_addr(&c.mboxp()[par][c.rank][4 * v], peer), w, peer_addr(&s->xbar[par], peer));
            }
            mbar_wait(&s->xbar[par], (c.xround >> 1) & 1u, c.p, 402);
        }
    } else {
        csync<NT_>();
        for (int w = c.tid; w < n; w += NT_) c.mboxp()[par][0][w] = s->outbox[par][w];
        csync<NT_>();
    }
    c.xround++;
    return c.mboxp()[par];
}

template <int NT_>
__device__ __forceinline__ uint32_t block_excl_scan_u32(Cx<NT_> &c, uint32_t v, uint32_t *total) {
    constexpr int NW = NT_ / 32;
    uint32_t inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (c.lane >= o) inc += t;
    }
    if (c.lane == 31) c.msp()->wscr_u[c.warp] = inc;
    csync<NT_>();
    if (c.warp == 0) {
        uint32_t w = c.lane < NW ? c.msp()->wscr_u[c.lane] : 0u, winc = w;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
            if (c.lane >= o) winc += t;
        }
        c.msp()->wscr_u[c.lane] = winc - w;          // exclusive warp offsets
        if (c.lane == 31) c.msp()->bc_u = winc;
    }
    csync<NT_>();
    uint32_t r = c.msp()->wscr_u[c.warp] + inc - v;
    *total = c.msp()->bc_u;
    csync<NT_>();
    return r;
}

// exclusive prefix of a double over the compute threads; *total = the sum (two CTA barriers)
template <int NT_>
__device__ __forceinline__ double block_excl_scan_f64(Cx<NT_> &c, double v, double *total) {
    constexpr int NW = NT_ / 32;
    double inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        double t = __shfl_up_sync(0xffffffffu, inc, o);
        if (c.lane >= o) inc += t;
    }
    if (c.lane == 31) c.msp()->wscr_d[c.warp] = inc;
    csync<NT_>();
    // every warp scans the warp totals itself (NW <= 32 values): no second hand-over
    double w = c.lane < NW ? c.msp()->wscr_d[c.lane] : 0.0, winc = w;
    for (int o = 1; o < 32; o <<= 1) {
        double t = __shfl_up_sync(0xffffffffu, winc, o);
        if (c.lane >= o) winc += t;
    }
    const double wex = __shfl_sync(0xffffffffu, winc - w, c.warp);
    *total = __shfl_sync(0xffffffffu, winc, 31);
    csync<NT_>();                                 // wscr_d may be rewritten
    return wex + (inc - v);
}

// ---------------------------------------------------------------------------------------------
// producers: HBM -> (L2 prefetch) -> staging ring (bulk copy) -> compact planes + E0 plane in the
// L2 scratch, zero-margin counts, tie count and the largest possible margin of my tile
// ---------------------------------------------------------------------------------------------
// A LIGHT image: one or two present classes, both planes resident in tensor memory for the whole image, no
// suppressed absent maximum to keep (that needs the whole pixel): see prod_chunk_light / plane_load_light.
__device__ __forceinline__ bool image_is_light(const EStepParams &p, int P, int cls0) {
    return p.light_ok && P <= 2 && !(P == 1 && P < p.C && cls0 != 0) && !(p.flags & OC_F_NO_RESIDENT);
}

struct Prod {
    const EStepParams *p;
    uint32_t o_ms, o_staging; // shared memory by offset (see Cx); NST stages shared by the producer warps
    __device__ __forceinline__ Misc *msp() const { return reinterpret_cast<Misc *>(smem0() + o_ms); }
    __device__ __forceinline__ unsigned char *stagingp() const { return smem0() + o_staging; }
    int pw, lane;
    int base, n_t;
    uint32_t nbase;           // chunks of earlier images (stage and parity of chunk q: from nbase + q)
    uint64_t pol;             // L2 evict-first policy of the score stream
#ifdef WL_SELPROF
    unsigned long long t_wait;
    uint32_t n_chunks;
#endif
};

__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void prod_prefetch(const Prod &w, const float *img, int q) {
    const EStepParams &p = *w.p;
    const int cpx = p.sub_px;
    const int npx = min(cpx, w.n_t - q * cpx);
    const uintptr_t a0 = (uintptr_t)(img + (long long)(w.base + q * cpx) * p.C);
    const uintptr_t lo = (a0 + 15u) & ~(uintptr_t)15u, hi = (a0 + (uintptr_t)npx * p.C * 4u) & ~(uintptr_t)15u;
    if (hi > lo)
        asm volatile("cp.async.bulk.prefetch.L2.global.L2::cache_hint [%0], %1, %2;\n" ::"l"(lo), "r"((uint32_t)(hi - lo)), "l"(w.pol) : "memory");
}

// Per-image constants of the chunk loop.  The producers are bound by the instructions AROUND the payload (a
// chunk is ~36 instructions of payload for a light image), so everything that does not change from chunk to
// chunk is formed once per image: when a chunk is a whole number of 16-byte units (cpx * C * 4 % 16 == 0) every
// full chunk has the same alignment, ragged head / tail and bulk size.
struct ChunkGeom {
    const float *gtile;       // scores of my tile's first pixel
    int cs;                   // floats per full chunk (cpx * C)
    int cpx, nchunks, n_t;
    int uniform;              // every full chunk shares lead / nh / bytes / t0
    int lead, nh, t0;
    uint32_t bytes;
    uint32_t stage_s, bar_s;  // shared-window addresses of stage 0 / full_bar[0]
    uint64_t pol;
    int pf_ahead;
    int whole_lo, whole_hi;   // chunks whole_lo .. whole_hi travel as ONE bulk copy of the 16-byte-aligned span around them
    uint32_t wbytes;          // its size for a full chunk
};
__device__ __forceinline__ void chunk_shape(const ChunkGeom &g, int C, int q, int &npx, int &lead, int &nh, uint32_t &bytes, int &t0, int &nfl) {
    if (g.uniform && (q + 1) * g.cpx <= g.n_t) {
        npx = g.cpx; nfl = g.cs; lead = g.lead; nh = g.nh; bytes = g.bytes; t0 = g.t0;
    } else {
        npx = min(g.cpx, g.n_t - q * g.cpx);
        nfl = npx * C;
        lead = (int)(((uintptr_t)(g.gtile + (long long)q * g.cs) & 15u) >> 2);
        nh = lead ? min(4 - lead, nfl) : 0;               // floats before the first aligned 16 bytes
        bytes = (uint32_t)((nfl - nh) * 4) & ~15u;          // aligned interior
        t0 = nh + (int)(bytes >> 2);                        // first float of the ragged tail
    }
}
// called by a whole warp: chunk q of the tile into ring slot d (its stage and barrier)
__device__ __forceinline__ void prod_issue2(const ChunkGeom &g, int C, int lane, int q, int d, const Prod &w, const float *img) {
    int npx, lead, nh, t0, nfl;
    uint32_t bytes;
    chunk_shape(g, C, q, npx, lead, nh, bytes, t0, nfl);
    const float *g0 = g.gtile + (long long)q * g.cs;
    const uint32_t stage_s = g.stage_s + (uint32_t)d * STAGE_BYTES + 4u * (uint32_t)lead, bar_s = g.bar_s + 8u * (uint32_t)d;
    if (q >= g.whole_lo && q <= g.whole_hi) {
        // The chunk sits inside the tensor with at least 12 bytes of it on either side: the copy engine moves the
        // 16-byte-aligned span AROUND it (up to 3 floats of the neighbouring chunks come along and are ignored), the
        // first pixel lands at the same stage offset as below.  No per-chunk cp.async, no group to commit.
        if (lane == 0) {
            const uint32_t wb = (g.uniform && nfl == g.cs) ? g.wbytes : (uint32_t)(((lead + nfl) * 4 + 15) & ~15);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar_s), "r"(wb) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;\n"
                         ::"r"(stage_s - 4u * (uint32_t)lead), "l"(g0 - lead), "r"(wb), "r"(bar_s), "l"(g.pol) : "memory");
        }
        return;
    }
    // the ragged head / tail floats travel as 4-byte asynchronous copies (one cp.async group per chunk and lane:
    // a plain load here would put a global-memory round trip on the issue path of EVERY chunk); lanes 0-2 the
    // head, lanes 3-5 the tail.  The stage is drained by this same warp, which waits for the group.
    {
        const int f = lane < 3 ? lane : t0 + lane - 3;
        const bool mine = lane < 3 ? lane < nh : (lane < 6 && f < nfl);
        if (mine) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(stage_s + 4u * (uint32_t)f), "l"(g0 + f) : "memory");
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    if (g.pf_ahead > 0) {                                   // warp-uniform: off by default
        if (lane == 7 && q + g.pf_ahead < g.nchunks) prod_prefetch(w, img, q + g.pf_ahead);
    }
    if (lane == 0) {
        if (bytes) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar_s), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;\n"
                         ::"r"(stage_s + 4u * (uint32_t)nh), "l"(g0 + nh), "r"(bytes), "r"(bar_s), "l"(g.pol) : "memory");
        } else {
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar_s) : "memory");
        }
    }
}
// The chunk loop of one image for one way of draining a stage: body(px0, npx, i0) gets the staged pixels of the
// chunk (px0 = first pixel, C floats each), their number and the tile-relative... image-relative index of the first.
template <typename BODY>
__device__ __forceinline__ void prod_loop(Prod &w, const ChunkGeom &g, int C, const float *img, int q0, BODY &&body) {
    const EStepParams &p = *w.p;
    const uint32_t full_bar0 = g.bar_s;
    for (int q = q0; q < g.nchunks; q += NPW) {
        const uint32_t n = w.nbase + (uint32_t)q;
        const int d = (int)(n % NST);
#ifdef WL_SELPROF
        const unsigned long long tw0 = gtimer();
#endif
        {   // mbar_wait on full_bar[d] (bounded)
            const uint32_t a = full_bar0 + 8u * (uint32_t)d, parity = (n / NST) & 1u;
            uint32_t ok;
            unsigned long long t0w = 0ull;
            for (;;) {
                asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                             : "=r"(ok) : "r"(a), "r"(parity) : "memory");
                if (ok) break;
                const unsigned long long t = gtimer();
                if (t0w == 0ull) t0w = t;
                else if (t - t0w > WAIT_LIMIT_NS) wait_failed(&p, 410);
            }
        }
        // my ragged floats of this chunk have landed too (the one younger group is the chunk issued into my
        // other stage; at the tail of the image there is none)
        static_assert(NST / NPW == 2, "one younger cp.async group at most");
        if (q + NPW < g.nchunks) asm volatile("cp.async.wait_group 1;\n" ::: "memory");
        else asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncwarp();
#ifdef WL_SELPROF
        w.t_wait += gtimer() - tw0;
        w.n_chunks++;
#endif
        int npx, lead, nh, t0, nfl;
        uint32_t bytes;
        chunk_shape(g, C, q, npx, lead, nh, bytes, t0, nfl);
        const float *px0 = reinterpret_cast<const float *>(w.stagingp() + (size_t)d * STAGE_BYTES) + lead;
        body(px0, npx, w.base + q * g.cpx);
        __syncwarp();                                  // stage drained by every lane
        if (q + NST < g.nchunks) prod_issue2(g, C, w.lane, q + NST, d, w, img);
    }
}

// One staged chunk of a label set of PP <= 4 classes (compile-time): the compact class planes, the floored
// running max E0 (weaklabel.cpp:67-77) and the statistics of the tile.  The producers are bound by the
// LATENCY of this loop (one warp per chunk; measured 1.6 us per 84-pixel chunk with 85 instructions per
// round), not by the stream, so: no branches, running pointers instead of 64-bit address arithmetic per
// store, and three rounds of 32 pixels in flight at a time.
struct ProdStats { uint32_t zc0, zc1, zc2, zc3, ties; float dmx, esum; };
template <int V> struct IntTag { static constexpr int value = V; };
template <int PP>
__device__ __forceinline__ void prod_chunk(const float *spx, int C, int npx, int lane, int pl0, int pl1, int pl2, int pl3,
                                           float *o0, float *o1, float *o2, float *o3, float *oe, ProdStats &st) {
    for (int g0 = 0; g0 < npx; g0 += 96) {
        float v[3][4];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const float *px = spx + min(g0 + 32 * r + lane, npx - 1) * C;        // clamped: always a staged pixel
            v[r][0] = px[pl0];
            if (PP > 1) v[r][1] = px[pl1];
            if (PP > 2) v[r][2] = px[pl2];
            if (PP > 3) v[r][3] = px[pl3];
        }
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const int g = g0 + 32 * r;
            const bool ok = g + lane < npx;
            const float v0 = v[r][0], v1 = PP > 1 ? v[r][1] : v0, v2 = PP > 2 ? v[r][2] : v0, v3 = PP > 3 ? v[r][3] : v0;
            float hi = v0, lo = v0;
            if (PP > 1) { hi = fmaxf(hi, v1); lo = fminf(lo, v1); }
            if (PP > 2) { hi = fmaxf(hi, v2); lo = fminf(lo, v2); }
            if (PP > 3) { hi = fmaxf(hi, v3); lo = fminf(lo, v3); }
            const float E = fmaxf(hi, FLT_MIN);                                   // the floor (weaklabel.cpp:71)
            if (ok) {
                __stcg(o0 + g, v0);
                if (PP > 1) __stcg(o1 + g, v1);
                if (PP > 2) __stcg(o2 + g, v2);
                if (PP > 3) __stcg(o3 + g, v3);
                __stcg(oe + g, E);
            }
            const uint32_t e0 = (ok && v0 == E) ? 1u : 0u, e1 = (PP > 1 && ok && v1 == E) ? 1u : 0u;
            const uint32_t e2 = (PP > 2 && ok && v2 == E) ? 1u : 0u, e3 = (PP > 3 && ok && v3 == E) ? 1u : 0u;
            st.zc0 += e0;
            if (PP > 1) st.zc1 += e1;
            if (PP > 2) st.zc2 += e2;
            if (PP > 3) st.zc3 += e3;
            if (PP > 1) st.ties += (e0 + e1 + e2 + e3 >= 2u) ? 1u : 0u;
            st.dmx = fmaxf(st.dmx, ok ? __fsub_rn(E, lo) : 0.f);
            st.esum += ok ? E : 0.f;
        }
    }
}

// The LIGHT variant for label sets of one or two classes whose planes both stay in tensor memory: the producers
// only de-interleave (PP loads + PP stores per pixel); E0 and the statistics are formed by the compute threads
// while they move the planes into tensor memory (plane_load_light) -- they wait for the stream of such images
// anyway, and the producers were the bottleneck (1.15 us per chunk with E0 and the statistics in here).
template <int PP>
__device__ __forceinline__ void prod_chunk_light(const float *spx, int C, int npx, int lane, int pl0, int pl1, float *o0, float *o1) {
    for (int g0 = 0; g0 < npx; g0 += 96) {
        float v[3][2];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const float *px = spx + min(g0 + 32 * r + lane, npx - 1) * C;
            v[r][0] = px[pl0];
            if (PP > 1) v[r][1] = px[pl1];
        }
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const int g = g0 + 32 * r;
            if (g + lane < npx) {
                __stcg(o0 + g, v[r][0]);
                if (PP > 1) __stcg(o1 + g, v[r][1]);
            }
        }
    }
}

// Fused up-sampling (network.py:40): nothing is streamed.  Each lane forms the scores of the PRESENT classes of its
// pixel from the four taps of the low-resolution tensor and leaves the same planes, E0, absent maximum and statistics
// as the streaming paths.  The low-resolution rows a tile needs (its H / G output rows map to ~ h / G + 2 of them) are
// copied into the idle staging ring first -- the L1 of this kernel is a few KB beside 220 KB of shared memory, taps out
// of the L2 made the loop latency-bound -- unless they do not fit (then the taps stay global).  Kept OUT of the kernel
// body: its registers must not disturb the allocation of the streaming paths (the kernel sits at the 64-register cap).
template <bool STAGED>
__device__ __forceinline__ void prod_upsampled_loop(const EStepParams &p, const float *src, int row0, int pw, int lane, int base, int n_t,
                                                    const uint8_t *pl, int P, int first_absent, int need_abs, float *planes,
                                                    uint8_t *absidx, ProdStats &st) {
    float *absmax = planes + (size_t)p.C * p.Npad;
    float *e0pl = planes + (size_t)(p.C + 1) * p.Npad;
    const float margin = p.margin;
    const int pl0 = pl[0];
    const int shift_o = row0 * p.up_w * p.C;           // staged copy: offsets are relative to low-resolution row row0
    for (int t0 = pw * 32; t0 < n_t; t0 += NPW * 32) {
        const int t = t0 + lane;
        if (t >= n_t) continue;
        const int i = base + t;
        const int y = i / p.up_W, x = i - y * p.up_W;
        UpTap tap = up_tap(y, x, p.up_h, p.up_w, p.C, p.up_sy, p.up_sx);
        if (STAGED) { tap.o00 -= shift_o; tap.o01 -= shift_o; tap.o10 -= shift_o; tap.o11 -= shift_o; }
        float E = FLT_MIN, mn = FLT_MAX;
        uint32_t at = 0u, eq = 0u;                 // classes that hold the running max: count, mask of the first four
        for (int k = 0; k < P; k++) {
            const float v = up_value<!STAGED>(src, tap, pl[k]);
            __stcg(planes + (size_t)k * p.Npad + i, v);
            if (v > E) { E = v; at = 1u; eq = (k < 4) ? (1u << k) : 0u; }
            else if (v == E) { at++; if (k < 4) eq |= 1u << k; }
            if (v < mn) mn = v;
        }
        __stcg(e0pl + i, E);
        if (need_abs) {
            const float xv = mn;                   // P == 1: the one present score
            const float low = __fsub_rn(xv, margin);
            float amax = -INFINITY, tf = -INFINITY;
            for (int k = 0; k < p.C; k++) {
                if (k == pl0) continue;
                float a = up_value<!STAGED>(src, tap, k);
                if (a > xv) a = low;
                amax = fmaxf(amax, a);
                if (k == first_absent) tf = a;
            }
            __stcg(absmax + i, amax);
            absidx[i] = (uint8_t)((tf == amax) ? first_absent : 254);
        }
        st.zc0 += eq & 1u; st.zc1 += (eq >> 1) & 1u; st.zc2 += (eq >> 2) & 1u; st.zc3 += (eq >> 3) & 1u;
        st.ties += (at >= 2u) ? 1u : 0u;
        st.dmx = fmaxf(st.dmx, __fsub_rn(E, mn));
        st.esum += E;
    }
}
// The SEPARABLE form of the same values: value = top + (bottom - top) * y_lerp, where top / bottom are the
// horizontally interpolated low-resolution rows y0 / y1 -- they depend on (row, x, class) only.  The producers form
// them once per tile for the rows it touches and the PRESENT classes (hrow[row][class][x], in the idle staging
// ring), and a pixel then costs two shared-memory loads and one lerp per class instead of four taps and three.
// Same operations on the same operands as up_value(): the same bits.
__device__ __forceinline__ void prod_upsampled_separable(const EStepParams &p, const float *src, float *hrow, int r0, int nrows, int pw,
                                                         int lane, int base, int n_t, const uint8_t *pl, int P, float *planes,
                                                         bool light, ProdStats &st) {
    const int W = p.up_W, w = p.up_w, C = p.C;
    const int total = nrows * P * W;
    for (int idx = pw * 32 + lane; idx < total; idx += NPW * 32) {
        const int rk = idx / W, x = idx - rk * W;
        const int r = rk / P, k = rk - r * P;
        const float fx = __fmul_rn((float)x, p.up_sx);
        const int x0 = (int)fx, x1 = min(x0 + 1, w - 1);
        const float lx = __fsub_rn(fx, (float)x0);
        const float *row = src + (size_t)(r0 + r) * w * C + pl[k];
        const float tl = __ldg(row + x0 * C), tr = __ldg(row + x1 * C);
        hrow[idx] = __fadd_rn(tl, __fmul_rn(__fsub_rn(tr, tl), lx));
    }
    asm volatile("bar.sync 2, %0;\n" ::"n"(NPW * 32) : "memory");           // the four producer warps: rows formed
    float *e0pl = planes + (size_t)(p.C + 1) * p.Npad;
    const int PW = P * W;
    for (int t0 = pw * 32; t0 < n_t; t0 += NPW * 32) {
        const int t = t0 + lane;
        if (t >= n_t) continue;
        const int i = base + t;
        const int y = i / W, x = i - y * W;
        const float fy = __fmul_rn((float)y, p.up_sy);
        const int y0 = (int)fy, y1 = min(y0 + 1, p.up_h - 1);
        const float ly = __fsub_rn(fy, (float)y0);
        const float *h0 = hrow + (y0 - r0) * PW + x, *h1 = hrow + (y1 - r0) * PW + x;
        float E = FLT_MIN, mn = FLT_MAX;
        uint32_t at = 0u, eq = 0u;
        for (int k = 0; k < P; k++) {
            const float top = h0[k * W], bot = h1[k * W];
            const float v = __fadd_rn(top, __fmul_rn(__fsub_rn(bot, top), ly));
            __stcg(planes + (size_t)k * p.Npad + i, v);
            if (v > E) { E = v; at = 1u; eq = (k < 4) ? (1u << k) : 0u; }
            else if (v == E) { at++; if (k < 4) eq |= 1u << k; }
            if (v < mn) mn = v;
        }
        if (light) continue;                       // LIGHT image: E0 and the statistics are formed by the compute threads
        __stcg(e0pl + i, E);
        st.zc0 += eq & 1u; st.zc1 += (eq >> 1) & 1u; st.zc2 += (eq >> 2) & 1u; st.zc3 += (eq >> 3) & 1u;
        st.ties += (at >= 2u) ? 1u : 0u;
        st.dmx = fmaxf(st.dmx, __fsub_rn(E, mn));
        st.esum += E;
    }
}
__device__ __noinline__ void prod_upsampled(const EStepParams *pp, int b, int pw, int lane, int base, int n_t, const uint8_t *pl,
                                            int P, int first_absent, int need_abs, float *planes, uint8_t *absidx,
                                            float *staging, ProdStats *out) {
    const EStepParams &p = *pp;
    const float *src = p.up_src + (long long)b * p.up_h * p.up_w * p.C;
    ProdStats st = *out;
    if (n_t > 0) {
        // low-resolution rows touched by the output rows of my tile: lower tap of the first, upper tap of the last
        const int ya = base / p.up_W, yb = (base + n_t - 1) / p.up_W;
        const int r0 = (int)__fmul_rn((float)ya, p.up_sy), r1 = min((int)__fmul_rn((float)yb, p.up_sy) + 1, p.up_h - 1);
        const int nfl = (r1 - r0 + 1) * p.up_w * p.C;
        if (!need_abs && (size_t)(r1 - r0 + 1) * P * p.up_W * sizeof(float) <= (size_t)NST * STAGE_BYTES) {
            prod_upsampled_separable(p, src, staging, r0, r1 - r0 + 1, pw, lane, base, n_t, pl, P, planes,
                                     image_is_light(p, P, pl[0]), st);
        } else if ((size_t)nfl * sizeof(float) <= (size_t)NST * STAGE_BYTES) {
            const float *g = src + (size_t)r0 * p.up_w * p.C;
            for (int f = pw * 32 + lane; f < nfl; f += NPW * 32) staging[f] = __ldg(g + f);
            asm volatile("bar.sync 2, %0;\n" ::"n"(NPW * 32) : "memory");       // the four producer warps: rows staged
            prod_upsampled_loop<true>(p, staging, r0, pw, lane, base, n_t, pl, P, first_absent, need_abs, planes, absidx, st);
        } else {
            asm volatile("bar.sync 2, %0;\n" ::"n"(NPW * 32) : "memory");       // (same barrier count on both paths)
            prod_upsampled_loop<false>(p, src, 0, pw, lane, base, n_t, pl, P, first_absent, need_abs, planes, absidx, st);
        }
    } else {
        asm volatile("bar.sync 2, %0;\n" ::"n"(NPW * 32) : "memory");
    }
    *out = st;
}

// one image: chunks q with (nbase + q) % NPW == pw of the CTA's tile
__device__ __forceinline__ void prod_image(Prod &w, int b, int par, float *planes, uint8_t *absidx) {
    const EStepParams &p = *w.p;
    const ImageMeta *im = p.meta + b;
    const int P = im->P;
    uint8_t *pl = w.msp()->pplist[w.pw];
    if (w.lane < MAXC) pl[w.lane] = im->plist[w.lane];
    __syncwarp();
    const int pl0 = pl[0], pl1 = pl[1], pl2 = pl[2], pl3 = pl[3];
    const int first_absent = im->first_absent;
    const float *img = p.map + (long long)b * p.s0;
    const int cpx = p.sub_px;
    const int nchunks = (w.n_t + cpx - 1) / cpx;
    // a single FOREGROUND class: an absent class of lower index (the background) can win a rounding tie,
    // so the suppressed absent maximum is kept.  A background-only image needs nothing: every absent class
    // has a higher index and a suppressed score <= the background's (weaklabel.cpp:45-65).
    const bool need_abs = (P == 1) && (P < p.C) && (pl0 != 0);
    const bool light = image_is_light(p, P, pl0);
    float *absmax = planes + (size_t)p.C * p.Npad;
    float *e0pl = planes + (size_t)(p.C + 1) * p.Npad;
    uint32_t zc0 = 0u, zc1 = 0u, zc2 = 0u, zc3 = 0u, ties = 0u;
    float dmx = 0.f, esum = 0.f;        // esum: plain float sum of my E0 values (a bound of the shift needs mean E0 to ~1 %)
    const int q0 = (int)((w.pw + NPW - (w.nbase % NPW)) % NPW);   // my first chunk
    ChunkGeom g;
    g.gtile = img + (long long)w.base * p.C;
    g.cpx = cpx; g.cs = cpx * p.C; g.nchunks = nchunks; g.n_t = w.n_t;
    g.uniform = ((g.cs & 3) == 0) ? 1 : 0;
    g.lead = (int)(((uintptr_t)g.gtile & 15u) >> 2);
    g.nh = g.lead ? min(4 - g.lead, g.cs) : 0;
    g.bytes = (uint32_t)((g.cs - g.nh) * 4) & ~15u;
    g.t0 = g.nh + (int)(g.bytes >> 2);
    g.stage_s = smem_u32(w.stagingp());
    g.bar_s = smem_u32(&w.msp()->full_bar[0]);
    g.pol = w.pol;
    g.pf_ahead = p.pf_ahead;
    // whole-span copies read up to 12 bytes before / 15 bytes after a chunk: everywhere but at the two ends of the tensor
    g.wbytes = (uint32_t)(((g.lead + g.cs) * 4 + 15) & ~15);
    g.whole_lo = (b == 0 && w.base == 0 && g.lead != 0) ? 1 : 0;
    g.whole_hi = (b == p.B - 1 && w.base + w.n_t == p.N) ? nchunks - 2 : nchunks - 1;
    if (p.flags & OC_F_RAGGED_COPIES) { g.whole_lo = 1; g.whole_hi = 0; }
    if (p.up_src) {
        ProdStats ust = {0u, 0u, 0u, 0u, 0u, 0.f, 0.f};
        prod_upsampled(&p, b, w.pw, w.lane, w.base, w.n_t, pl, P, first_absent, need_abs ? 1 : 0, planes, absidx,
                       reinterpret_cast<float *>(w.stagingp()), &ust);
        zc0 = ust.zc0; zc1 = ust.zc1; zc2 = ust.zc2; zc3 = ust.zc3; ties = ust.ties; dmx = ust.dmx; esum = ust.esum;
    } else {
    // every stage is free (the producers met after the previous image): fill the ring
    for (int q = q0; q < nchunks && q < NST; q += NPW) prod_issue2(g, p.C, w.lane, q, (int)((w.nbase + (uint32_t)q) % NST), w, img);
    const int C = p.C;
    const int lane = w.lane;
    const size_t npad = (size_t)p.Npad;
    if (need_abs) {
        const float margin = p.margin;
        prod_loop(w, g, C, img, q0, [&](const float *px0, int npx, int i0) {
            float *out = planes + i0;
            for (int gt = lane; gt < npx; gt += 32) {
                const float *px = px0 + gt * C;
                const float x = px[pl0];
                const float low = __fsub_rn(x, margin);
                float a0 = -INFINITY, a1 = -INFINITY;
                int k = 0;
                for (; k + 1 < pl0; k += 2) {
                    const float t0 = px[k], t1 = px[k + 1];
                    a0 = fmaxf(a0, t0 > x ? low : t0);
                    a1 = fmaxf(a1, t1 > x ? low : t1);
                }
                if (k < pl0) { const float t0 = px[k]; a0 = fmaxf(a0, t0 > x ? low : t0); }
                k = max(k, pl0 + 1);
                for (; k + 1 < C; k += 2) {
                    const float t0 = px[k], t1 = px[k + 1];
                    a0 = fmaxf(a0, t0 > x ? low : t0);
                    a1 = fmaxf(a1, t1 > x ? low : t1);
                }
                if (k < C) { const float t0 = px[k]; a0 = fmaxf(a0, t0 > x ? low : t0); }
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
                esum += E;
            }
        });
    } else if (light && P == 1) {
        prod_loop(w, g, C, img, q0, [&](const float *px0, int npx, int i0) {
            float *o0 = planes + i0 + lane;
            prod_chunk_light<1>(px0, C, npx, lane, pl0, pl1, o0, o0 + npad);
        });
    } else if (light) {
        prod_loop(w, g, C, img, q0, [&](const float *px0, int npx, int i0) {
            float *o0 = planes + i0 + lane;
            prod_chunk_light<2>(px0, C, npx, lane, pl0, pl1, o0, o0 + npad);
        });
    } else if (P <= 4) {
        ProdStats st = {zc0, zc1, zc2, zc3, ties, dmx, esum};
        auto run = [&](auto tag) {
            constexpr int PP = decltype(tag)::value;
            prod_loop(w, g, C, img, q0, [&](const float *px0, int npx, int i0) {
                float *o0 = planes + i0 + lane, *oe = e0pl + i0 + lane;
                float *o1 = o0 + npad, *o2 = o1 + npad, *o3 = o2 + npad;
                prod_chunk<PP>(px0, C, npx, lane, pl0, pl1, pl2, pl3, o0, o1, o2, o3, oe, st);
            });
        };
        if (P == 1) run(IntTag<1>());
        else if (P == 2) run(IntTag<2>());
        else if (P == 3) run(IntTag<3>());
        else run(IntTag<4>());
        zc0 = st.zc0; zc1 = st.zc1; zc2 = st.zc2; zc3 = st.zc3; ties = st.ties; dmx = st.dmx; esum = st.esum;
    } else {
        prod_loop(w, g, C, img, q0, [&](const float *px0, int npx, int i0) {
            float *out = planes + i0;
            for (int gt = lane; gt < npx; gt += 32) {
                const float *px = px0 + gt * C;
                float E = FLT_MIN, mn = FLT_MAX;
                for (int k = 0; k < P; k++) {
                    const float v = px[pl[k]];
                    __stcg(out + (size_t)k * npad + gt, v);
                    if (v > E) E = v;
                    if (v < mn) mn = v;
                }
                __stcg(e0pl + i0 + gt, E);
                uint32_t at = 0u;
                for (int k = 0; k < P; k++) at += (px[pl[k]] == E) ? 1u : 0u;
                zc0 += (px[pl0] == E) ? 1u : 0u;
                zc1 += (px[pl1] == E) ? 1u : 0u;
                zc2 += (px[pl2] == E) ? 1u : 0u;
                zc3 += (px[pl3] == E) ? 1u : 0u;
                ties += (at >= 2u) ? 1u : 0u;
                dmx = fmaxf(dmx, __fsub_rn(E, mn));
                esum += E;
            }
        });
    }
    w.nbase += (uint32_t)nchunks;
    }   // (streamed image)
    // my share of the tile's statistics
    uint32_t dm = __float_as_uint(dmx);               // margins are >= +0: the bit patterns order like the values
    for (int o = 16; o; o >>= 1) {
        zc0 += __shfl_xor_sync(0xffffffffu, zc0, o);
        zc1 += __shfl_xor_sync(0xffffffffu, zc1, o);
        zc2 += __shfl_xor_sync(0xffffffffu, zc2, o);
        zc3 += __shfl_xor_sync(0xffffffffu, zc3, o);
        ties += __shfl_xor_sync(0xffffffffu, ties, o);
        dm = max(dm, __shfl_xor_sync(0xffffffffu, dm, o));
        esum += __shfl_xor_sync(0xffffffffu, esum, o);
    }
    if (w.lane == 0) {
        uint32_t *st = w.msp()->pstat[par][w.pw];
        st[0] = zc0; st[1] = zc1; st[2] = zc2; st[3] = zc3; st[4] = dm; st[5] = ties; st[6] = __float_as_uint(esum);
    }
}

template <int NT_>
__device__ __forceinline__ void producer_main(const EStepParams &p, Misc *ms, unsigned char *staging, int slot, int base, int n_t) {
    Prod w;
    w.p = &p;
    w.o_ms = (uint32_t)(reinterpret_cast<unsigned char *>(ms) - smem0());
    w.pw = (int)(threadIdx.x - NT_) >> 5;
    w.lane = threadIdx.x & 31;
    w.o_staging = (uint32_t)(staging - smem0());
    w.base = base;
    w.n_t = n_t;
    w.nbase = 0;
    w.pol = l2_evict_first_policy();
    for (uint32_t s = 0;; s++) {
        const int q = s & 1u;
        mbar_wait(&ms->cmd_bar[q], (s >> 1) & 1u, &p, 411);
        const int b = *reinterpret_cast<volatile int *>(&ms->cmd_img[q]);
        if (b >= p.B) break;
        float *planes = p.planes + (size_t)(slot * 2 + q) * (size_t)(p.C + 2) * p.Npad;
        uint8_t *absidx = p.absent_idx + (size_t)(slot * 2 + q) * p.Npad;
        asm volatile("bar.sync 2, %0;\n" ::"n"(NPW * 32) : "memory");   // every stage of the ring is drained
#ifdef WL_SELPROF
        w.t_wait = 0ull; w.n_chunks = 0u;
        const unsigned long long tp0 = gtimer();
#endif
        prod_image(w, b, q, planes, absidx);
#ifdef WL_SELPROF
        if (p.prof && w.pw == 0 && w.lane == 0) {
            uint32_t rk;
            asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(rk));
            if (rk == 0) {
                unsigned long long *pr = p.prof + (size_t)b * 32;
                pr[13] = w.t_wait; pr[14] = gtimer() - tp0; pr[15] = w.n_chunks; pr[12] = tp0;
            }
        }
#endif
        __threadfence();                                        // my plane/E0 stores have reached the L2 (the
        __syncwarp();                                           // E0 tile is fetched from there by the copy engine)
        if (w.lane == 0) mbar_arrive(&ms->ready_bar[q]);        // release: planes and statistics of my chunks
    }
}

// ---------------------------------------------------------------------------------------------
// order-exact running sums (weaklabel.cpp:69-80 float, :102-106 double) by digest
// ---------------------------------------------------------------------------------------------
template <typename T> struct FnIO;
template <> struct FnIO<float> {
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
    static __device__ __forceinline__ ChunkFn<double> widen(const ChunkFn<float> &f) {
        ChunkFn<double> g;
        g.e = f.e; g.a0 = f.a0; g.a1 = f.a1; g.kind = f.kind;
        return g;
    }
    static __device__ __forceinline__ ChunkFn<float> narrow(const ChunkFn<double> &f) {
        ChunkFn<float> g;
        g.e = (uint32_t)f.e; g.a0 = (int32_t)f.a0; g.a1 = (int32_t)f.a1; g.kind = f.kind;
        return g;
    }
};
template <> struct FnIO<double> {
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
    static __device__ __forceinline__ ChunkFn<double> widen(const ChunkFn<double> &f) { return f; }
    static __device__ __forceinline__ ChunkFn<double> narrow(const ChunkFn<double> &f) { return f; }
};

// Tolerance (relative to the binade) inside which a predicted binade boundary is not trusted: the
// float running sum drifts from the real sum by accumulated rounding (a few hundred ulps at most
// here), the double one practically not at all.  Only speed depends on it, never the result.
template <typename T> struct drift_tol;
template <> struct drift_tol<float> { static constexpr double value = 1.0 / 8192.0; };
template <> struct drift_tol<double> { static constexpr double value = 1.0 / 1073741824.0; };

// per-thread chunk sums in double (the prediction of seqsum.cuh); returns my chunk's exclusive
// prefix inside the tile, *tile_total = the tile's sum.  s_E is zero beyond the tile.
template <int NT_>
__device__ __forceinline__ double chunk_prefix(Cx<NT_> &c, double *tile_total, double *chunk_sum) {
    const int L = c.p->L;
    const float *t = c.sE() + c.tid * L;
    double a0 = 0.0, a1 = 0.0;
    int i = 0;
    for (; i + 1 < L; i += 2) { a0 += (double)t[i]; a1 += (double)t[i + 1]; }
    if (i < L) a0 += (double)t[i];
    const double cs = a0 + a1;
    *chunk_sum = cs;
    return block_excl_scan_f64(c, cs, tile_total);
}

// The local digest of my tile for the sum of type T over the terms in s_E: maximal runs of chunks
// summarised for the same binade become one composite (segmented scan over the tile); chunks that
// contain, or are suspiciously close to, a binade crossing stay singletons that are added term by
// term.  `before` = the (double) sum of the tiles of lower rank; cpre / cs from chunk_prefix.
template <typename T, int NT_>
__device__ __forceinline__ void build_big_digest(Cx<NT_> &c, double cpre, double cs, double before) {
    typedef fp_traits<T> F;
    typedef FnIO<T> IO;
    constexpr int NW = NT_ / 32;
    Misc *ms = c.msp();
    SumScratch *ss = sums(c);
    BigDigest *big = &ss->big;
    const int L = c.p->L;
    const int c0 = c.tid * L;
    const int len = max(0, min(L, c.n_t - c0));
    const double pin = before + cpre;
    ChunkFn<T> f = (len > 0) ? chunk_eval<T>(c.sE() + c0, len, pin) : chunk_identity<T>();
    if (f.kind == 1) {
        // distrust predictions that pass close to a boundary of the binade
        const double lo = (double)F::from(f.e), tol = lo * drift_tol<T>::value;
        if (pin - lo < tol || (lo + lo) - (pin + cs) < tol) f.kind = 0;
    }
    if (c.p->flags & OC_F_NO_DIGEST) { if (f.kind == 1) f.kind = 0; }
    // segment heads: the chunk before me (warp edges through shared memory)
    ChunkFn<T> pv = IO::shfl_up(f, 1);
    if (c.lane == 31) ss->wlast[c.warp] = IO::widen(f);
    csync<NT_>();
    if (c.lane == 0 && c.warp > 0) pv = IO::narrow(ss->wlast[c.warp - 1]);
    bool head;
    if (f.kind == 2) head = false;
    else if (c.tid == 0 || f.kind == 0) head = true;
    else head = (pv.kind != 1) || (pv.e != f.e);
    // segmented inclusive scan inside the warp; hidx = chunk index of the head of my segment
    ChunkFn<T> x = f;
    int hidx = c.tid;
    bool flag = head;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const ChunkFn<T> y = IO::shfl_up(x, d);
        const int yh = __shfl_up_sync(0xffffffffu, hidx, d);
        const bool yf = __shfl_up_sync(0xffffffffu, flag ? 1 : 0, d) != 0;
        if (c.lane >= d && !flag) {
            if (y.kind != 2) hidx = yh;
            x = chunk_compose(y, x);
            flag = yf;
        }
    }
    if (c.lane == 31) { ss->wagg[c.warp].f = IO::widen(x); ss->wagg[c.warp].hidx = hidx; ss->wagg[c.warp].flag = flag ? 1 : 0; }
    if (c.lane == 0) ms->wfirst[c.warp] = head ? 1u : 0u;
    csync<NT_>();
    {
        // every warp scans the warp aggregates itself; carry INTO warp w = inclusive result of warp w-1
        ChunkFn<T> v = c.lane < NW ? IO::narrow(ss->wagg[c.lane].f) : chunk_identity<T>();
        int vh = c.lane < NW ? ss->wagg[c.lane].hidx : 0;
        bool vf = c.lane < NW ? ss->wagg[c.lane].flag != 0 : false;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const ChunkFn<T> y = IO::shfl_up(v, d);
            const int yh = __shfl_up_sync(0xffffffffu, vh, d);
            const bool yf = __shfl_up_sync(0xffffffffu, vf ? 1 : 0, d) != 0;
            if (c.lane >= d && !vf) {
                if (y.kind != 2) vh = yh;
                v = chunk_compose(y, v);
                vf = yf;
            }
        }
        const int src = c.warp > 0 ? c.warp - 1 : 0;
        ChunkFn<T> cin = IO::shfl(v, src);
        const int cinh = __shfl_sync(0xffffffffu, vh, src);
        if (c.warp == 0) cin = chunk_identity<T>();
        if (!flag) {                                  // no head at or before me inside my warp
            if (cin.kind != 2) hidx = cinh;
            x = chunk_compose(cin, x);
        }
    }
    // tails: the chunk after me starts a segment (or I am the last chunk)
    bool next_head = __shfl_down_sync(0xffffffffu, head ? 1 : 0, 1) != 0;
    if (c.lane == 31) next_head = (c.warp + 1 < NW) ? (ms->wfirst[c.warp + 1] != 0u) : true;
    const bool emit = next_head && x.kind != 2;
    const uint32_t bal = __ballot_sync(0xffffffffu, emit);
    if (c.lane == 0) ms->wcnt[c.warp] = (uint32_t)__popc(bal);
    csync<NT_>();
    {
        uint32_t cnt = c.lane < NW ? ms->wcnt[c.lane] : 0u, inc = cnt;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (c.lane >= o) inc += t;
        }
        const uint32_t woff = __shfl_sync(0xffffffffu, inc - cnt, c.warp);
        const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
        if (emit) {
            const uint32_t id = woff + (uint32_t)__popc(bal & ((1u << c.lane) - 1u));
            if (id < (uint32_t)BIG_SEG) {
                DigEntry e;
                e.e = (unsigned long long)x.e;
                e.a0 = (long long)x.a0;
                e.a1 = (long long)x.a1;
                e.kind = x.kind == 1 ? 1 : 0;
                e.cslot = -1;
                e.nterms = 0;
                e.head = (uint16_t)hidx;
                e.tail = (uint16_t)c.tid;
                e.pad[0] = e.pad[1] = 0u;
                if (id == (uint32_t)BIG_SEG - 1u && total > (uint32_t)BIG_SEG) {   // overflow: the rest of the tile, term by term
                    e.kind = 0;
                    e.tail = (uint16_t)(NT_ - 1);
                }
                big->seg[id] = e;
            }
        }
        if (c.tid == 0) { big->nseg = (int)min(total, (uint32_t)BIG_SEG); big->total = (int)total; }
    }
    csync<NT_>();
}

// warp 0: condense the local digest into the fixed-size digest that travels to CTA 0
template <int NT_>
__device__ __forceinline__ void compact_digest(Cx<NT_> &c, Digest *out) {
    const BigDigest *big = &sums(c)->big;
    const int L = c.p->L;
    const int n = big->nseg;
    int nseg = n;
    if (n > DIG_SEG) nseg = DIG_SEG;
    int ncross = 0;
    for (int i0 = 0; i0 < nseg; i0 += 32) {
        const int i = i0 + c.lane;
        DigEntry e;
        bool inl = false;
        if (i < nseg) {
            e = big->seg[i];
            if (i == DIG_SEG - 1 && n > DIG_SEG) { e.kind = 0; e.tail = (uint16_t)(NT_ - 1); }
            inl = (e.kind == 0) && (e.head == e.tail) && !(c.p->flags & OC_F_NO_DIGEST);
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, inl);
        if (i < nseg) {
            if (inl) {
                const int slot = ncross + __popc(bal & ((1u << c.lane) - 1u));
                if (slot < DIG_CROSS) {
                    e.cslot = slot * DIG_L;
                    e.nterms = max(0, min(L, c.n_t - (int)e.head * L));
                }
            }
            out->seg[i] = e;
        }
        ncross += __popc(bal);
    }
    if (c.lane == 0) { out->nseg = nseg; out->ncross = min(ncross, DIG_CROSS); out->pad0 = out->pad1 = 0; }
    __syncwarp();
    // the terms of the crossing chunks
    for (int i = 0; i < nseg; i++) {
        const int off = out->seg[i].cslot;
        if (off >= 0) {
            const int t0 = (int)out->seg[i].head * L, nt = out->seg[i].nterms;
            for (int j = c.lane; j < DIG_L; j += 32) out->cross[0][off + j] = (j < nt) ? c.sE()[t0 + j] : 0.f;
        }
    }
    __syncwarp();
}

// add terms one by one: `get(j)` delivers term j (any lane may be asked for any j, uniformly)
template <typename T, typename G>
__device__ __forceinline__ T add_range(T s, int i0, int i1, int lane, G &&get) {
    typedef fp_traits<T> F;
    for (int j0 = i0; j0 < i1; j0 += 32) {
        const int j = j0 + lane;
        const float v = (j < i1) ? get(j) : 0.f;                 // + 0 leaves a sum >= 0 unchanged
#pragma unroll
        for (int l = 0; l < 32; l++) s = F::add(s, (T)__shfl_sync(0xffffffffu, v, l));
    }
    return s;
}

// Walk the entries of one digest (all lanes of the warp hold the same state s).  `get(j)` = term j of
// the tile the digest describes (used for raw ranges and for composites that do not apply).
template <typename T, typename G>
__device__ __forceinline__ T walk_digest(T s, const DigEntry *seg, int nseg, const float *cross,
                                         int L, int n_t, int lane, const float *local, G &&get) {
    typedef fp_traits<T> F;
    for (int i = 0; i < nseg; i++) {
        const DigEntry e = seg[i];
        if (e.kind == 1) {
            ChunkFn<T> f;
            f.e = (typename F::U)e.e;
            f.a0 = (typename F::S)e.a0;
            f.a1 = (typename F::S)e.a1;
            f.kind = 1;
            if (chunk_apply(f, s)) continue;
        }
        if (e.kind == 0 && e.cslot >= 0 && cross != nullptr) {
            const float *tt = cross + e.cslot;                    // cslot: offset of the terms (a multiple of 4 floats)
            const float4 *t4 = reinterpret_cast<const float4 *>(tt);
            const int m = e.nterms;
            int j = 0;
            for (; j + 4 <= m; j += 4) {
                const float4 v = t4[j >> 2];
                s = F::add(s, (T)v.x);
                s = F::add(s, (T)v.y);
                s = F::add(s, (T)v.z);
                s = F::add(s, (T)v.w);
            }
            for (; j < m; j++) s = F::add(s, (T)tt[j]);
            continue;
        }
        const int i0 = (int)e.head * L, i1 = min(((int)e.tail + 1) * L, n_t);
        if (local != nullptr) {
            // my own tile: the terms come straight out of shared memory (the loads do not depend on the
            // running sum, so they are issued ahead of the add chain)
            int j = i0;
#pragma unroll 1
            for (; j + 8 <= i1; j += 8) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; u++) v[u] = local[j + u];
#pragma unroll
                for (int u = 0; u < 8; u++) s = F::add(s, (T)v[u]);
            }
            for (; j < i1; j++) s = F::add(s, (T)local[j]);
        } else {
            s = add_range<T>(s, i0, i1, lane, get);
        }
    }
    return s;
}

// One order-exact running sum (T = float: weaklabel.cpp:69-80, T = double: :102-106) of the terms in shared
// memory (s_E): a digest per tile, pushed to CTA 0 with one bulk copy, walked there by warp 0, the result
// stored into every CTA.  fallback(r, j): term j of tile r, for the rare ranges of a peer's digest whose
// terms did not travel with it.  Called by every compute thread of every CTA of the cluster.
template <typename T, int NT_, typename FB>
__device__ __forceinline__ T exact_sum(Cx<NT_> &c, int which, FB &&fallback) {
    typedef fp_traits<T> F;
    const EStepParams &p = *c.p;
    Misc *ms = c.msp();
    const int L = p.L;
    double tile_total, cs;
    const double cpre = chunk_prefix(c, &tile_total, &cs);
    if (c.tid == 32) {
        const unsigned long long tb = (unsigned long long)__double_as_longlong(tile_total);
        uint32_t *ob = outbox(c);
        ob[0] = (uint32_t)tb; ob[1] = (uint32_t)(tb >> 32); ob[2] = 0u; ob[3] = 0u;
    }
    double before = 0.0;                    // plain (double) sum of the tiles of lower rank: the binade prediction
    {
        const uint32_t (*box)[MBW] = exchange(c, 4);
        for (int q = 0; q < c.rank; q++)
            before += __longlong_as_double((long long)((unsigned long long)box[q][0] | ((unsigned long long)box[q][1] << 32)));
    }
    if (c.rank == 0 && c.tid == 0 && c.G > 1) mbar_expect_tx(&ms->dig_bar, (uint32_t)((c.G - 1) * sizeof(Digest)));
    if (c.tid == 32) mbar_expect_tx(&ms->res_bar, 8u);
    build_big_digest<T, NT_>(c, cpre, cs, before);
    if (c.rank != 0) {
        if (c.warp == 0) {
            compact_digest(c, &c.outdp()[which]);
            asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");   // my digest writes -> visible to the copy engine
            __syncwarp();
            if (c.lane == 0)
                bulk_s2peer(peer_addr(&sums(c)->inbound[c.rank][which], 0), &c.outdp()[which], (uint32_t)sizeof(Digest),
                            peer_addr(&ms->dig_bar, 0));
        }
    } else if (c.warp == 0) {
        SumScratch *ss = sums(c);
        const BigDigest *big = &ss->big;
        T s = walk_digest<T>((T)0, big->seg, big->nseg, nullptr, L, c.n_t, c.lane, c.sE(), [&](int j) { return c.sE()[j]; });
        if (c.G > 1) mbar_wait(&ms->dig_bar, c.n_dig & 1u, &p, 406);
        for (int r = 1; r < c.G; r++) {
            const int n_t_r = max(0, min(p.TS, p.N - r * p.TS));
            if (n_t_r == 0) break;
            const Digest *d = &ss->inbound[r][which];
            s = walk_digest<T>(s, d->seg, d->nseg, &d->cross[0][0], L, n_t_r, c.lane, nullptr, [&](int j) { return fallback(r, j); });
        }
        const unsigned long long bits = (unsigned long long)F::bits(s);
        if (c.lane < c.G) st_async_u64(peer_addr(&ms->sumres[which], c.lane), bits, peer_addr(&ms->res_bar, c.lane));
    }
    c.n_dig++;
    mbar_wait(&ms->res_bar, c.n_res & 1u, &p, 408);
    c.n_res++;
    return F::from((typename F::U)ms->sumres[which]);
}

// ---------------------------------------------------------------------------------------------
// binning of a margin key (bit pattern of a float >= +0) and the inverse
//   bin 0: zero margins; bin 1: below the window; bins 2 .. : 128 per binade from lo_key; 2047: above
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int win_bin(uint32_t key, int base) {
    int b = (int)(key >> WIN_SH) - base;
    b = max(b, (int)min(key, 1u));
    return min(b, NB - 1);
}
__device__ __forceinline__ void win_range(uint32_t bin, uint32_t lo_key, uint32_t &klo, uint32_t &khi) {
    if (bin == 0u) { klo = khi = 0u; return; }
    if (bin == 1u) { klo = 1u; khi = lo_key > 1u ? lo_key - 1u : 1u; return; }
    klo = lo_key + ((bin - 2u) << WIN_SH);
    khi = bin == (uint32_t)NB - 1u ? 0xffffffffu : klo + ((1u << WIN_SH) - 1u);
    if (klo == 0u) klo = 1u;                            // lo_key == 0: key 0 belongs to bin 0, not to the first window bin
}

struct Located {
    int bin;              // -1: inconsistent (NaN in the scores)
    uint32_t pre, cnt;    // keys of the histogram below the bin / inside it
};

// the quads of a thread: k-th quad = tile quad k * NT + tid, TMEM columns [4k, 4k+4) of the slot.
// tcgen05.ld/st are warp-collective: they are issued for EVERY k by every lane; quads beyond the
// tile (q >= TQ) hold -inf and f must not touch shared/global memory for them.
template <int NT_, int QPT, typename F>
__device__ __forceinline__ void for_each_quad(Cx<NT_> &c, uint32_t ta, F &&f) {
#pragma unroll
    for (int k = 0; k + 1 < QPT; k += 2) {
        float4 a, b;
        tm_ld8(ta + 4 * k, a, b);
        f(k, k * NT_ + c.tid, a);
        f(k + 1, (k + 1) * NT_ + c.tid, b);
    }
    if (QPT & 1) {
        constexpr int k = QPT - 1;
        f(k, k * NT_ + c.tid, tm_ld4(ta + 4 * k));
    }
}

// one histogram round over my tile's margins E - x: window bins.  Zero margins (usually the largest
// single bin by far) go to a private counter per warp (h[NB + warp]) instead of one contended bin.
template <int NT_, int QPT>
__device__ __forceinline__ void hist_pass_win(Cx<NT_> &c, uint32_t ta, int base, uint32_t *h) {
    const int zi = NB + c.warp;
    if (c.lane == 0) h[zi] = 0u;
    __syncwarp();
    for_each_quad<NT_, QPT>(c, ta, [&](int, int q, float4 x) {
        if (q < c.TQ) {
            const float4 E = *reinterpret_cast<const float4 *>(c.sE() + 4 * q);
            const uint32_t k0 = __float_as_uint(__fsub_rn(E.x, x.x)), k1 = __float_as_uint(__fsub_rn(E.y, x.y));
            const uint32_t k2 = __float_as_uint(__fsub_rn(E.z, x.z)), k3 = __float_as_uint(__fsub_rn(E.w, x.w));
            atomicAdd(&h[k0 ? win_bin(k0, base) : zi], 1u);
            atomicAdd(&h[k1 ? win_bin(k1, base) : zi], 1u);
            atomicAdd(&h[k2 ? win_bin(k2, base) : zi], 1u);
            atomicAdd(&h[k3 ? win_bin(k3, base) : zi], 1u);
        }
    });
    __syncwarp();
    if (c.lane == 0) { const uint32_t z = h[zi]; if (z) atomicAdd(&h[0], z); }
    csync<NT_>();
}
// bracketed bins (refinement rounds: a coarse bin or heavy ties)
template <int NT_, int QPT>
__device__ __forceinline__ void hist_pass_range(Cx<NT_> &c, uint32_t ta, uint32_t lo, uint32_t hi, int sh, uint32_t *h) {
    for_each_quad<NT_, QPT>(c, ta, [&](int, int q, float4 x) {
        if (q < c.TQ) {
            const float4 E = *reinterpret_cast<const float4 *>(c.sE() + 4 * q);
            const uint32_t k0 = __float_as_uint(__fsub_rn(E.x, x.x)), k1 = __float_as_uint(__fsub_rn(E.y, x.y));
            const uint32_t k2 = __float_as_uint(__fsub_rn(E.z, x.z)), k3 = __float_as_uint(__fsub_rn(E.w, x.w));
            if (k0 >= lo && k0 <= hi) atomicAdd(&h[(k0 - lo) >> sh], 1u);
            if (k1 >= lo && k1 <= hi) atomicAdd(&h[(k1 - lo) >> sh], 1u);
            if (k2 >= lo && k2 <= hi) atomicAdd(&h[(k2 - lo) >> sh], 1u);
            if (k3 >= lo && k3 <= hi) atomicAdd(&h[(k3 - lo) >> sh], 1u);
        }
    });
    csync<NT_>();
}

// agree, across the cluster, on the bin of the current histogram that holds rank r
template <int NT_>
__device__ __forceinline__ Located locate(Cx<NT_> &c, uint32_t r) {
    constexpr int BPT = NB / HT;                // bins per scanning thread (the first HT threads scan)
    constexpr int TPS = SUBB / BPT;             // threads per super bin
    Misc *s = c.msp();
    SelScratch *ss = sel(c);
    uint32_t *h = ss->hist[c.hround & 1u];
    if (c.tid < HT) {   // 64 super-bin sums of my histogram -> outbox
        const uint4 hv = *reinterpret_cast<const uint4 *>(&h[BPT * c.tid]);
        uint32_t v = hv.x + hv.y + hv.z + hv.w;
#pragma unroll
        for (int o = TPS / 2; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((c.tid & (TPS - 1)) == 0) outbox(c)[c.tid / TPS] = v;
    }
    if (c.tid == NT_ - 1) s->lcount = 0u;
    const uint32_t (*box)[MBW] = exchange(c, NSB);
    SELP(c, 1);
    // every warp finds the super bin itself (64 sums x G CTAs)
    int S = -1;
    uint32_t preS = 0u;
    {
        uint32_t s0 = 0, s1 = 0;
        for (int p = 0; p < c.G; p++) { const uint2 t = *reinterpret_cast<const uint2 *>(&box[p][2 * c.lane]); s0 += t.x; s1 += t.y; }
        uint32_t inc = s0 + s1;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (c.lane >= o) inc += t;
        }
        const uint32_t ex = inc - (s0 + s1);
        int mine = -1;
        uint32_t mpre = 0u;
        if (r >= ex && r < ex + s0) { mine = 2 * c.lane; mpre = ex; }
        else if (r >= ex + s0 && r < ex + s0 + s1) { mine = 2 * c.lane + 1; mpre = ex + s0; }
        const uint32_t bal = __ballot_sync(0xffffffffu, mine >= 0);
        if (bal) {
            const int src = __ffs(bal) - 1;
            S = __shfl_sync(0xffffffffu, mine, src);
            preS = __shfl_sync(0xffffffffu, mpre, src);
        }
    }
    Located L;
    L.bin = -1; L.pre = 0; L.cnt = 0;
    c.hround++;
    if (S < 0) return L;                               // uniform over the cluster
    // the 32 fine bins of super bin S from every CTA: warp p fetches CTA p's
    if (c.warp < c.G)
        ss->fine[c.warp][c.lane] = (c.warp == c.rank) ? h[S * SUBB + c.lane] : ld_peer(peer_addr(&h[S * SUBB + c.lane], c.warp));
    csync<NT_>();
    {
        uint32_t f = 0;
        for (int p = 0; p < c.G; p++) f += ss->fine[p][c.lane];
        uint32_t inc = f;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (c.lane >= o) inc += t;
        }
        const uint32_t ex = inc - f, rr = r - preS;
        const bool hit = rr >= ex && rr < ex + f;
        const uint32_t bal = __ballot_sync(0xffffffffu, hit);
        if (bal) {
            const int src = __ffs(bal) - 1;
            L.bin = S * SUBB + src;
            L.pre = preS + __shfl_sync(0xffffffffu, ex, src);
            L.cnt = __shfl_sync(0xffffffffu, f, src);
        }
    }
    return L;
}

// exact selection of rank r among cnt keys in [lo, hi] held in CTA-local shared memory: histogram rounds
template <int NT_>
__device__ __forceinline__ uint32_t local_select_hist(Cx<NT_> &c, const uint32_t *keys, uint32_t cnt, uint32_t r,
                                                      uint32_t lo, uint32_t hi) {
    constexpr int BPT = NB / HT;
    Misc *s = c.msp();
    uint32_t *h = sel(c)->lhist;
    for (int guard = 0; guard < 4; guard++) {
        if (lo == hi) break;
        int sh = 0;
        while (((hi - lo) >> sh) >= (uint32_t)NB) sh++;
        for (int i = c.tid; i < NB; i += NT_) h[i] = 0u;
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

// the same by counting: candidate i has rank [lt, lt + eq) -- one CTA barrier, no histogram.
// keys[cnt .. cnt+3] must hold 0xffffffff.
template <int NT_>
__device__ __forceinline__ uint32_t local_select_rank(Cx<NT_> &c, const uint32_t *keys, uint32_t cnt, uint32_t r) {
    if ((uint32_t)c.tid < cnt) {
        const uint32_t my = keys[c.tid];
        uint32_t lt = 0u, eq = 0u;
        const uint4 *k4 = reinterpret_cast<const uint4 *>(keys);
        for (uint32_t j = 0; j < cnt; j += 4) {
            const uint4 g = k4[j >> 2];
            lt += (g.x < my) + (g.y < my) + (g.z < my) + (g.w < my);
            eq += (g.x == my) + (g.y == my) + (g.z == my) + (g.w == my);
        }
        if (lt <= r && r < lt + eq) c.msp()->sel_key = my;
    }
    csync<NT_>();
    return c.msp()->sel_key;
}

// The value of rank k (0-based) among the margins E[i] - x[i] of the whole image; the plane is in the
// TMEM slot at ta.  *zeros receives the exact number of zero margins when the selection saw them
// (else 0xffffffff).
template <int NT_, int QPT>
__device__ __forceinline__ float select_margin(Cx<NT_> &c, uint32_t ta, uint32_t k, float dmax_bound, uint32_t *zeros) {
    SelScratch *ss = sel(c);
    uint32_t e_top = (__float_as_uint(dmax_bound) >> 23) + 1u;
    if (e_top > 254u) e_top = 254u;
    const uint32_t lo_key = (e_top > (uint32_t)WIN_BINADES ? e_top - WIN_BINADES : 0u) << 23;
    const int base = (int)(lo_key >> WIN_SH) - 2;
    uint32_t lo = lo_key, hi = 0u, r = k;
    int sh = 0;
    bool window = true;
    *zeros = 0xffffffffu;
    for (int guard = 0; guard < 6; guard++) {
        uint32_t *h = ss->hist[c.hround & 1u];          // zeroed since its last use
        SELP(c, 9);
        if (window) hist_pass_win<NT_, QPT>(c, ta, base, h);
        else hist_pass_range<NT_, QPT>(c, ta, lo, hi, sh, h);
        SELP(c, 0);
        const Located L = locate(c, r);
        SELP(c, 2);
        // the histogram of the PREVIOUS round is no longer read by any peer (they all took part in this
        // round's exchange after their fine-bin reads): clear it for the next round
        {
            uint4 *hz4 = reinterpret_cast<uint4 *>(ss->hist[c.hround & 1u]);
            for (int i = c.tid; i < NBX / 4; i += NT_) hz4[i] = make_uint4(0u, 0u, 0u, 0u);
        }
        if (L.bin < 0) {
            if (c.tid == 0) atomicExch(c.p->status, 200 + guard);
            csync<NT_>();
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
        if (klo == khi) { csync<NT_>(); return __uint_as_float(klo); }
        if (L.cnt <= (uint32_t)GCAP) {
            // gather the candidates of the bin into every CTA, in rank order of the CTAs
            for_each_quad<NT_, QPT>(c, ta, [&](int, int q, float4 x) {
                if (q >= c.TQ) return;
                const float4 E = *reinterpret_cast<const float4 *>(c.sE() + 4 * q);
                const uint32_t kk[4] = {__float_as_uint(__fsub_rn(E.x, x.x)), __float_as_uint(__fsub_rn(E.y, x.y)),
                                        __float_as_uint(__fsub_rn(E.z, x.z)), __float_as_uint(__fsub_rn(E.w, x.w))};
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    if (kk[e] - klo <= khi - klo) {
                        const uint32_t pos = atomicAdd(&c.msp()->lcount, 1u);
                        if (pos < (uint32_t)GCAP) ss->mylist[pos] = kk[e];
                    }
                }
            });
            SELP(c, 3);
            const int fb = L.bin & (SUBB - 1);
            const uint32_t mine = ss->fine[c.rank][fb];
            uint32_t off = 0;
            for (int p = 0; p < c.rank; p++) off += ss->fine[p][fb];
            if (c.tid < 4) ss->gather[L.cnt + c.tid] = 0xffffffffu;
            const uint32_t par = c.xround & 1u;
            if (c.G > 1 && !(c.p->flags & OC_F_CLASSIC_XCHG)) {
                if (c.tid == 0) mbar_expect_tx(&c.msp()->xbar[par], L.cnt * 4u);
                csync<NT_>();                            // my list is complete, the barrier is armed
                for (uint32_t i = c.tid; i < (uint32_t)c.G * mine; i += NT_) {
                    const uint32_t peer = i / mine, w = i - peer * mine;
                    st_async_u32(peer_addr(&ss->gather[off + w], (int)peer), ss->mylist[w], peer_addr(&c.msp()->xbar[par], (int)peer));
                }
                mbar_wait(&c.msp()->xbar[par], (c.xround >> 1) & 1u, c.p, 403);
                c.xround++;
                SELP(c, 4);
            } else if (c.G > 1) {
                csync<NT_>();
                for (uint32_t i = c.tid; i < (uint32_t)c.G * mine; i += NT_) {
                    const uint32_t peer = i / mine, w = i - peer * mine;
                    st_peer(peer_addr(&ss->gather[off + w], (int)peer), ss->mylist[w]);
                }
                cluster_sync_compute(c);
            } else {
                csync<NT_>();
                for (uint32_t w = c.tid; w < mine; w += NT_) ss->gather[w] = ss->mylist[w];
                csync<NT_>();
            }
            uint32_t key;
            if (L.cnt <= (uint32_t)RCMAX && !(c.p->flags & OC_F_NO_RANKCOUNT)) key = local_select_rank(c, ss->gather, L.cnt, r);
            else { key = local_select_hist(c, ss->gather, L.cnt, r, klo, khi); csync<NT_>(); }
            SELP(c, 5);
            return __uint_as_float(key);
        }
        // heavy ties or a coarse (out-of-window) bin: narrow the bracket and bin again
        window = false;
        lo = klo; hi = khi;
        sh = 0;
        while (((hi - lo) >> sh) >= (uint32_t)NB) sh++;
        csync<NT_>();                                   // the cleared histogram before the next round's atomics
    }
    if (c.tid == 0) atomicExch(c.p->status, 299);
    csync<NT_>();
    return 0.f;
}

// ---------------------------------------------------------------------------------------------
// class planes: L2 scratch <-> tensor memory
// ---------------------------------------------------------------------------------------------
template <int NT_, int QPT> __device__ __forceinline__ uint32_t slot_addr(const Cx<NT_> &c, int slot) {
    return c.tm + (uint32_t)(slot * slot_cols(QPT));
}
template <int NT_, int QPT>
__device__ __forceinline__ void plane_load(Cx<NT_> &c, int plane, int slot) {
    const float *src = c.planes + (size_t)plane * c.p->Npad + c.base;
    const uint32_t ta = slot_addr<NT_, QPT>(c, slot);
    // every load of the thread is in flight before the first value is used (the plane comes out of the L2)
    float4 v[QPT];
#pragma unroll
    for (int k = 0; k < QPT; k++) {
        const int q = k * NT_ + c.tid;
        v[k] = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);         // quads beyond the tile
        if (q < c.TQ) v[k] = __ldcg(reinterpret_cast<const float4 *>(src + 4 * q));
    }
#pragma unroll
    for (int k = 0; k < QPT; k++) {
        const int q = k * NT_ + c.tid;
        if (q < c.TQ && 4 * q + 3 >= c.n_t) {           // the last quad of the image: padding pixels
            if (4 * q + 1 >= c.n_t) v[k].y = -INFINITY;
            if (4 * q + 2 >= c.n_t) v[k].z = -INFINITY;
            v[k].w = -INFINITY;
        }
        tm_st4(ta + 4 * k, v[k]);                       // warp-collective: every lane, every k
    }
    tm_wait_st();
}
// LIGHT images (image_is_light): both planes go from the L2 scratch into the two tensor-memory slots and, on
// the way, the floored running max E0 (weaklabel.cpp:67-77) goes into shared memory and the statistics of my
// tile are formed -- zero-margin counts per plane, exact ties, the largest margin and the plain sum of E0 --
// exactly as the producers form them for the other images.  st: zc0, zc1, ties, dmx, esum (per thread).
struct LightStats { uint32_t zc0, zc1, ties; float dmx, esum; };
template <int NT_, int QPT, int PP>
__device__ __forceinline__ void plane_load_light(Cx<NT_> &c, LightStats &st) {
    const float *s0 = c.planes + c.base, *s1 = s0 + c.p->Npad;
    const uint32_t t0 = slot_addr<NT_, QPT>(c, 0), t1 = slot_addr<NT_, QPT>(c, 1 % n_slots(QPT));
    float *sE = c.sE();
    const float4 ninf = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
#pragma unroll
    for (int k0 = 0; k0 < QPT; k0 += 3) {
        float4 a[3], b[3];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const int q = (k0 + r) * NT_ + c.tid;
            a[r] = ninf; b[r] = ninf;
            if (k0 + r < QPT && q < c.TQ) {
                a[r] = __ldcg(reinterpret_cast<const float4 *>(s0 + 4 * q));
                if (PP > 1) b[r] = __ldcg(reinterpret_cast<const float4 *>(s1 + 4 * q));
            }
        }
#pragma unroll
        for (int r = 0; r < 3; r++) {
            if (k0 + r >= QPT) continue;
            const int k = k0 + r, q = k * NT_ + c.tid;
            float va[4] = {a[r].x, a[r].y, a[r].z, a[r].w}, vb[4] = {b[r].x, b[r].y, b[r].z, b[r].w};
            float E[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const bool ok = q < c.TQ && 4 * q + e < c.n_t;
                if (!ok) { va[e] = -INFINITY; vb[e] = -INFINITY; }             // padding pixels: x = -inf, E = 0
                const float v0 = va[e], v1 = PP > 1 ? vb[e] : v0;
                const float Ee = fmaxf(fmaxf(v0, v1), FLT_MIN);
                E[e] = ok ? Ee : 0.f;
                const uint32_t e0 = (ok && v0 == Ee) ? 1u : 0u, e1 = (PP > 1 && ok && v1 == Ee) ? 1u : 0u;
                st.zc0 += e0;
                st.zc1 += e1;
                st.ties += e0 & e1;
                st.dmx = fmaxf(st.dmx, ok ? __fsub_rn(Ee, fminf(v0, v1)) : 0.f);
                st.esum += ok ? Ee : 0.f;
            }
            tm_st4(t0 + 4 * k, make_float4(va[0], va[1], va[2], va[3]));          // warp-collective: every lane, every k
            if (PP > 1) tm_st4(t1 + 4 * k, make_float4(vb[0], vb[1], vb[2], vb[3]));
            if (q < c.TQ) *reinterpret_cast<float4 *>(sE + 4 * q) = make_float4(E[0], E[1], E[2], E[3]);
        }
    }
    tm_wait_st();
}
// E0 of a LIGHT image again (the float running sum, when it is needed after all): the planes in the scratch
// still hold the original scores -- label sets of <= 2 classes never write a plane back
template <int NT_, int PP>
__device__ __forceinline__ void recompute_E0_light(Cx<NT_> &c) {
    const float *s0 = c.planes + c.base, *s1 = s0 + c.p->Npad;
    float *sE = c.sE();
    for (int i = c.tid; i < c.n_t; i += NT_) {
        const float v0 = __ldcg(s0 + i), v1 = PP > 1 ? __ldcg(s1 + i) : v0;
        sE[i] = fmaxf(fmaxf(v0, v1), FLT_MIN);
    }
}

template <int NT_, int QPT>
__device__ __forceinline__ void plane_store(Cx<NT_> &c, int plane, int slot) {
    float *dst = c.planes + (size_t)plane * c.p->Npad + c.base;
    const uint32_t ta = slot_addr<NT_, QPT>(c, slot);
    for_each_quad<NT_, QPT>(c, ta, [&](int, int q, float4 x) {
        if (q < c.TQ) __stcg(reinterpret_cast<float4 *>(dst + 4 * q), x);   // padding lanes of the last quad: never read as scores
    });
}
// the TMEM slot that holds `plane` (loading it, and writing back the plane it replaces, if needed)
template <int NT_, int QPT>
__device__ __forceinline__ int plane_acquire(Cx<NT_> &c, int plane) {
    constexpr int NS = n_slots(QPT);
    int s;
    if (c.res0 == plane) s = 0;
    else if (NS > 1 && c.res1 == plane) s = 1;
    else {
        s = (NS > 1) ? (c.res0 < 0 ? 0 : (c.res1 < 0 ? 1 : (c.mru ^ 1))) : 0;
        if (s == 0) {
            if (c.res0 >= 0 && c.dirty0) plane_store<NT_, QPT>(c, c.res0, 0);
            plane_load<NT_, QPT>(c, plane, 0);
            c.res0 = plane;
            c.dirty0 = false;
        } else {
            if (c.res1 >= 0 && c.dirty1) plane_store<NT_, QPT>(c, c.res1, 1);
            plane_load<NT_, QPT>(c, plane, 1);
            c.res1 = plane;
            c.dirty1 = false;
        }
    }
    c.mru = s;
    return s;
}

// ---------------------------------------------------------------------------------------------
// labels (weaklabel.cpp:108-112 + deeplab.py:123)
// ---------------------------------------------------------------------------------------------
// Exact label of one pixel from the final scores of the present classes (vals[pl], float) and the
// original scores of the absent ones, exactly as the reference forms them: suppressed (weaklabel.cpp:
// 45-65), shifted in double, rounded to float (:110), first maximum (deeplab.py:123).
// The C original scores of pixel i of image b: a row of the map -- or, with fused up-sampling, the values the
// producers formed for it, interpolated again into loc (same taps, same operations: same bits).
__device__ __noinline__ void up_row(const EStepParams &p, int b, int i, float *loc) {
    const int y = i / p.up_W, x = i - y * p.up_W;
    const UpTap tap = up_tap(y, x, p.up_h, p.up_w, p.C, p.up_sy, p.up_sx);
    const float *src = p.up_src + (long long)b * p.up_h * p.up_w * p.C;
    for (int k = 0; k < p.C; k++) loc[k] = up_value<true>(src, tap, k);
}
template <typename V>
static __device__ __forceinline__ int label_exact(const EStepParams &p, uint32_t mask, int b, int i, double shift, V &&val) {
    float loc[MAXC];
    const float *px = p.map + (long long)b * p.s0 + (long long)i * p.s2;
    if (p.up_src) { up_row(p, b, i, loc); px = loc; }
    float m = FLT_MAX;
    for (int k = 0; k < p.C; k++)
        if ((mask >> k) & 1u) { const float t = px[k]; if (t < m) m = t; }
    const float low = __fsub_rn(m, p.margin);
    float best = -INFINITY;
    int bi = 0, pp = 0;
    for (int k = 0; k < p.C; k++) {
        float v;
        if ((mask >> k) & 1u) {
            v = val(pp);
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

// Two distinct positive floats lo < hi stay distinct (and ordered) after the shift s, rounded to float, for
// EVERY s in [-sneg, +spos] if no shifted value can land in a binade whose float spacing exceeds hi - lo:
// hi - lo >= ulp(lo) = 2^(e-23) with 2^e <= lo, and every y in [lo - sneg, hi + spos] with |y| < 2^(e+1) has
// ulp(y) <= 2^(e-23); two reals at least one grid step apart never round to the same grid point.  (The
// shift of the E-step is <= 0 up to the rounding of mean_max_0, so positive top scores only move towards
// finer spacing: the exact tie a selection leaves at its pivot -- or its 1-ulp neighbour -- is decided here.)
__device__ __forceinline__ bool pair_stays_distinct(float lo, float hi, float sneg, float spos) {
    if (!(lo > 0.f)) return false;
    const float top = __uint_as_float((__float_as_uint(lo) & 0x7f800000u) + 0x00800000u);   // 2^(e+1) (inf: never safe)
    return fmaxf(fabsf(__fsub_rd(lo, sneg)), __fadd_ru(hi, spos)) < top;
}

// The largest SUPPRESSED absent score of one pixel (weaklabel.cpp:45-65) out of the original map; -inf when
// every class is present.  For the rare pixels at which all present classes hold the same final score (every
// selection leaves such an exact tie at its pivot): an absent class can reach the maximum only there.
__device__ __noinline__ float absent_max(const float *px, uint32_t mask, int C, float margin) {
    float m = FLT_MAX;
    for (int k = 0; k < C; k++)
        if ((mask >> k) & 1u) { const float t = px[k]; if (t < m) m = t; }
    const float low = __fsub_rn(m, margin);
    float a = -INFINITY;
    for (int k = 0; k < C; k++)
        if (!((mask >> k) & 1u)) { float v = px[k]; if (v > m) v = low; a = fmaxf(a, v); }
    return a;
}

template <int NT_> __device__ __forceinline__ bool labels_via_smem(const Cx<NT_> &c) {
    return (size_t)c.n_t + 32 <= sizeof(SelScratch);
}

// Pass A -- runs while the two running sums are still being walked, i.e. WITHOUT the re-centring shift
// (weaklabel.cpp:108-112), of which only an upper bound `ashift` on the magnitude is known.  Two final
// scores a > b can round to the same float after the shift only if a - b is below the spacing of the
// floats around the shifted values: guard = 2^-22 * (|best| + |shift|) is at least twice that spacing.
// Every pixel whose first maximum is safe against that guard gets its label now; the others are
// returned as a bit per pixel of this thread (bit 4k + e) and decided by pass B with the exact shift.
template <int NT_, int QPT>
__device__ __forceinline__ unsigned long long labels_pass_a(Cx<NT_> &c, float ashift, float spos, int tid_q, uint32_t tm_q) {
    // (tid_q, tm_q): the thread whose quads are labelled and its tensor-memory columns -- this thread's own, or
    // those of a walking warp of CTA 0 (same lane quarter, lower columns) served by a proxy warp
    const EStepParams &p = *c.p;
    uint8_t *out = p.labels + (size_t)c.b * p.N + c.base;
    if (c.P == 1 && c.msp()->plist[0] == 0) {
        // background only: every absent class has a higher index and a score <= the background's
        for (int i = tid_q; i < c.n_t; i += NT_) out[i] = 0;
        return 0ull;
    }
    // the tile's labels are collected in the (now idle) selection scratch, one word per quad, and leave in
    // 16-byte stores (labels_copy_out) instead of one byte store per pixel
    const bool via_smem = labels_via_smem<NT_>(c);
    uint32_t *labs = reinterpret_cast<uint32_t *>(c.unip());
    const bool has_absent = c.P < p.C;
    const bool have_abs = has_absent && c.P == 1;          // suppressed absent maximum was kept
    const bool exact_all = (p.flags & OC_F_NO_FASTLABEL) != 0;
    const uint32_t t0 = tm_q, t1 = tm_q + (uint32_t)((1 % n_slots(QPT)) * slot_cols(QPT));
    const int r0 = c.res0, r1 = n_slots(QPT) > 1 ? c.res1 : -1;
    const int cls0 = r0 >= 0 ? c.msp()->plist[r0] : 0, cls1 = r1 >= 0 ? c.msp()->plist[r1] : 0;
    const bool two_resident = (c.P == 2 && r0 >= 0 && r1 >= 0 && r0 != r1);
    unsigned long long unsure = 0ull;
    // one loop per case (both planes resident / the general one): inside ONE loop the two bodies shared a register
    // allocation and the loop spilled around every quad
    auto rounds = [&](auto tag) {
    constexpr bool TWO = decltype(tag)::value != 0;
#pragma unroll 1
    for (int k = 0; k < QPT; k++) {
        // tcgen05.ld is warp-collective: both slots are read by every lane for every k
        const float4 s0v = tm_ld4(t0 + 4 * k);
        const float4 s1v = n_slots(QPT) > 1 ? tm_ld4(t1 + 4 * k) : s0v;
        const int q = k * NT_ + tid_q;
        if (q >= c.TQ) continue;
        const float a0[4] = {s0v.x, s0v.y, s0v.z, s0v.w};
        const float a1[4] = {s1v.x, s1v.y, s1v.z, s1v.w};
        uint32_t lab4 = 0u, un4 = 0u;
        if (TWO) {
            // both planes in tensor memory: the first maximum in CLASS order, safe unless the two scores
            // are distinct and within the guard, or equal (then an absent class may tie after rounding)
            const int lo_cls = r0 < r1 ? cls0 : cls1, hi_cls = r0 < r1 ? cls1 : cls0;
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float lo = r0 < r1 ? a0[e] : a1[e], hi = r0 < r1 ? a1[e] : a0[e];
                const float guard = __fmul_ru(__fadd_ru(fmaxf(fabsf(lo), fabsf(hi)), ashift), 2.4e-7f);   // > 2^-22
                const float d = fabsf(__fsub_rn(hi, lo));
                bool bad = exact_all || !(d > guard);                       // d <= guard, NaN
                if (bad && d > 0.f && !exact_all) bad = !pair_stays_distinct(fminf(lo, hi), fmaxf(lo, hi), ashift, spos);
                if (d == 0.f && !exact_all && 4 * q + e < c.n_t) {
                    // an exact tie of the two present classes goes to the lower index whatever the shift -- unless
                    // an absent class comes within the guard of them
                    // (fused up-sampling: there is no map row to look at -- the pixel is left to pass B, which interpolates it)
                    bad = has_absent && (p.up_src != nullptr ||
                                         !(__fsub_rn(lo, absent_max(p.map + (long long)c.b * p.s0 + (long long)(c.base + 4 * q + e) * p.s2,
                                                                    c.mask, p.C, p.margin)) > guard));
                }
                lab4 |= (uint32_t)(hi > lo ? hi_cls : lo_cls) << (8 * e);
                un4 |= (bad ? 1u : 0u) << e;
            }
        } else {
            float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
            float near_[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // largest score strictly below best
            int bi[4] = {0, 0, 0, 0};
            for (int pl = 0; pl < c.P; pl++) {
                float4 v;
                if (r0 == pl) v = s0v;
                else if (r1 == pl) v = s1v;
                else v = __ldcg(reinterpret_cast<const float4 *>(c.planes + (size_t)pl * p.Npad + c.base + 4 * q));
                const float vv[4] = {v.x, v.y, v.z, v.w};
                const int cls = c.msp()->plist[pl];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float t = vv[e];
                    if (t > best[e]) { near_[e] = best[e]; best[e] = t; bi[e] = cls; }
                    else if (t < best[e]) near_[e] = fmaxf(near_[e], t);
                }
            }
            float4 av = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
            if (have_abs) av = __ldcg(reinterpret_cast<const float4 *>(c.absmax + c.base + 4 * q));
            const float ava[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float guard = __fmul_ru(__fadd_ru(fabsf(best[e]), ashift), 2.4e-7f);       // > 2^-22
                // a distinct score within the guard; or (absent classes exist) every present score equal: a
                // suppressed absent score is <= the smallest present one and can reach the maximum only then;
                // or (one present class) the suppressed absent maximum within the guard
                bool bad = exact_all || !(__fsub_rn(best[e], near_[e]) > guard);
                if (bad && !exact_all && near_[e] < best[e]) bad = !pair_stays_distinct(near_[e], best[e], ashift, spos);
                if (has_absent && !have_abs && near_[e] == -INFINITY && !bad && 4 * q + e < c.n_t)
                    bad = p.up_src != nullptr ||
                          !(__fsub_rn(best[e], absent_max(p.map + (long long)c.b * p.s0 + (long long)(c.base + 4 * q + e) * p.s2,
                                                          c.mask, p.C, p.margin)) > guard);
                if (have_abs && !(__fsub_rn(best[e], ava[e]) > guard)) bad = true;
                lab4 |= (uint32_t)bi[e] << (8 * e);
                un4 |= (bad ? 1u : 0u) << e;
            }
        }
        const int i0 = 4 * q;
        if (i0 + 3 >= c.n_t) un4 &= (1u << (c.n_t - i0)) - 1u;      // padding pixels of the last quad are nobody's
        if (via_smem) {
            labs[q] = lab4;                     // open pixels too: pass B overwrites them in the label map
        } else {
#pragma unroll
            for (int e = 0; e < 4; e++)
                if (i0 + e < c.n_t && !((un4 >> e) & 1u)) out[i0 + e] = (uint8_t)(lab4 >> (8 * e));
        }
        unsure |= (unsigned long long)un4 << (4 * k);
    }
    };
    if (two_resident) rounds(IntTag<1>()); else rounds(IntTag<0>());
    return unsure;
}

// The tile's labels, left in the selection scratch by pass A, go to the label map: the tile starts at an
// arbitrary byte of the map (513^2 / 9 pixels per tile), so the words are re-aligned to the destination
// with funnel shifts and leave as 16-byte stores; head and tail bytes one by one.
template <int NT_>
__device__ __forceinline__ void labels_copy_out(Cx<NT_> &c) {
    const EStepParams &p = *c.p;
    if (!labels_via_smem<NT_>(c)) return;
    if (c.P == 1 && c.msp()->plist[0] == 0) return;              // background only: written directly
    uint8_t *dst = p.labels + (size_t)c.b * p.N + c.base;
    const uint32_t *S = reinterpret_cast<const uint32_t *>(c.unip());
    const uint8_t *Sb = reinterpret_cast<const uint8_t *>(S);
    const int n = c.n_t;
    const int head = min(n, (int)((16u - (uint32_t)((uintptr_t)dst & 15u)) & 15u));
    const int nvec = (n - head) >> 4;
    if (c.tid < head) dst[c.tid] = Sb[c.tid];
    for (int v = c.tid; v < nvec; v += NT_) {
        const int o = head + 16 * v;
        const uint32_t *w = S + (o >> 2);
        const uint32_t sh = (uint32_t)(o & 3) * 8u;
        const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = sh ? w[4] : 0u;
        uint4 r;
        r.x = __funnelshift_r(w0, w1, sh);
        r.y = __funnelshift_r(w1, w2, sh);
        r.z = __funnelshift_r(w2, w3, sh);
        r.w = __funnelshift_r(w3, w4, sh);
        *reinterpret_cast<uint4 *>(dst + o) = r;
    }
    const int t0 = head + 16 * nvec;
    if (c.tid < n - t0) dst[t0 + c.tid] = Sb[t0 + c.tid];
}

// Pass B -- the pixels pass A left open, now with the exact shift: the reference's own arithmetic
// (label_exact), one pixel at a time.  Rare, so it is kept OUT of the kernel body: a separate function
// with plain arguments (inlined, its size kept the compiler from promoting the kernel's context struct to
// registers, which turned every shared-memory access of the kernel into a generic one).
struct PassB {
    const EStepParams *p;
    const uint8_t *plist;
    const float *planes, *absmax;
    const uint8_t *absidx;
    uint8_t *out;                 // labels of my tile
    double shift;
    unsigned long long unsure;
    uint32_t t0, t1, mask;
    int r0, r1, b, base, n_t, tid, P, first_absent, nt, qpt, two_slots;
};
__device__ __noinline__ void labels_pass_b(const PassB a) {
    const EStepParams &p = *a.p;
    const bool have_abs = (a.P < p.C) && a.P == 1;
#pragma unroll 1
    for (int k = 0; k < a.qpt; k++) {
        const uint32_t nib = (uint32_t)(a.unsure >> (4 * k)) & 15u;
        if (!__any_sync(0xffffffffu, nib != 0u)) continue;           // warp-uniform: tcgen05.ld is collective
        const float4 s0v = tm_ld4(a.t0 + 4 * k);
        const float4 s1v = a.two_slots ? tm_ld4(a.t1 + 4 * k) : s0v;
        const int q = k * a.nt + a.tid;
#pragma unroll 1
        for (int e = 0; e < 4; e++) {
            const int i = 4 * q + e;
            if (!((nib >> e) & 1u) || i >= a.n_t) continue;
            const float v0 = e == 0 ? s0v.x : (e == 1 ? s0v.y : (e == 2 ? s0v.z : s0v.w));
            const float v1 = e == 0 ? s1v.x : (e == 1 ? s1v.y : (e == 2 ? s1v.z : s1v.w));
            int lab = -1;
            if (have_abs) {
                // one present class whose score is v0 (plane 0 is resident): when the lowest absent class holds
                // the suppressed absent maximum the two shifted floats decide, lowest index on a tie
                const float ava = __ldcg(a.absmax + a.base + i);
                const int aia = __ldcg(a.absidx + a.base + i);
                if (a.r0 == 0 && aia == a.first_absent) {
                    const float fa = __double2float_rn(__dadd_rn((double)ava, a.shift));
                    const float fb = __double2float_rn(__dadd_rn((double)v0, a.shift));
                    const int cls = a.plist[0];
                    lab = fa > fb ? aia : (fa == fb ? min(aia, cls) : cls);
                }
            }
            if (lab < 0)
                lab = label_exact(p, a.mask, a.b, a.base + i, a.shift, [&](int pp) {
                    if (a.r0 == pp) return v0;
                    if (a.r1 == pp) return v1;
                    return __ldcg(a.planes + (size_t)pp * p.Npad + a.base + i);
                });
            a.out[i] = (uint8_t)lab;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
template <int NT_, int QPT>
__global__ void __launch_bounds__(NT_ + NPW * 32, 1) estep_onchip_kernel(const __grid_constant__ EStepParams p) {
    typedef Layout<NT_> LY;
    constexpr int NW = NT_ / 32;
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
        size_t o = LY::A16((size_t)p.se_floats * 4);        // the running-max tile sits at offset 0
        c.o_mbox = (uint32_t)o;
        o += LY::MBOX_BYTES;
        c.o_ms = (uint32_t)o;
        o += LY::A16(sizeof(Misc));
        c.o_uni = (uint32_t)o;
        o += LY::uni_bytes();
        c.o_outd = (uint32_t)o;
        o += 2 * LY::A16(sizeof(Digest));
        staging = smem_raw + o;
    }
    c.xround = 0;
    c.hround = 0;
    c.csyncs = 0;
    c.n_e0 = c.n_dig = c.n_res = 0;
    c.base = c.rank * p.TS;
    c.n_t = max(0, min(p.TS, p.N - c.base));
    c.TQ = (c.n_t + 3) >> 2;
    Misc *const ms = c.msp();
    if (c.tid == 0) {
        for (int d = 0; d < NST; d++) mbar_init(&ms->full_bar[d], 1);
        mbar_init(&ms->e0_bar, 1);
        mbar_init(&ms->dig_bar, 1);
        mbar_init(&ms->res_bar, 1);
        for (int q = 0; q < 2; q++) {
            mbar_init(&ms->cmd_bar[q], 1);
            mbar_init(&ms->ready_bar[q], NPW);
            mbar_init(&ms->xbar[q], 1);
            mbar_init(&ms->sbar[q], p.G);
        }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (c.warp == 0) tm_alloc512(&ms->tmem_base);
    tm_fence_before();
    hw_cluster_sync();                                  // every CTA's barriers exist before any remote arrival
    tm_fence_after();
    c.tm = ms->tmem_base + ((uint32_t)(c.warp & 3) << 21) + (uint32_t)((c.warp >> 2) * 4 * QPT);
    if (c.tid >= NT_) {
        producer_main<NT_>(p, ms, staging, slot, c.base, c.n_t);
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
                if (seq == 0) st_peer(peer_addr(&ms->cmd_img[0], c.lane), (uint32_t)img0);
                st_peer(peer_addr(&ms->cmd_img[(seq + 1) & 1u], c.lane), (uint32_t)img1);
            }
        }
        cluster_sync_compute(c);
        if (c.tid == 0) {
            if (seq == 0) mbar_arrive(&ms->cmd_bar[0]);
            mbar_arrive(&ms->cmd_bar[(seq + 1) & 1u]);
        }
        c.b = *reinterpret_cast<volatile int *>(&ms->cmd_img[seq & 1u]);
        if (c.b >= p.B) break;
        c.planes = p.planes + (size_t)(slot * 2 + (int)(seq & 1u)) * (size_t)(p.C + 2) * p.Npad;
        c.absmax = c.planes + (size_t)p.C * p.Npad;
        c.absidx = p.absent_idx + (size_t)(slot * 2 + (int)(seq & 1u)) * p.Npad;
        unsigned long long *prof = p.prof ? p.prof + (size_t)c.b * 32 : nullptr;
        unsigned long long t_acq = 0ull, t_sel = 0ull, t_fold = 0ull;
        if (prof && lead_thread) prof[0] = gtimer();
#ifdef WL_SELPROF
        if (lead_thread) for (int i = 0; i < 10; i++) ms->tacc[i] = 0ull;
#endif
        const ImageMeta *im = p.meta + c.b;
        c.mask = im->mask;
        c.P = im->P;
        c.first_absent = im->first_absent;
        c.res0 = c.res1 = -1;
        c.dirty0 = c.dirty1 = false;
        c.mru = 0;
        if (c.tid < MAXC) {
            ms->plist[c.tid] = im->plist[c.tid];
            ms->zlb[c.tid] = 0u;
        }
        const bool ord_in_smem = p.num_iter * p.C <= ORD_S;
        if (ord_in_smem && c.tid >= 64 && c.tid - 64 < p.num_iter * p.C)
            ms->ord_s[c.tid - 64] = p.order[(size_t)c.b * p.num_iter * p.C + (c.tid - 64)];
        // 1. the planes of my tile are in the scratch (written by my producers together with E0, the
        //    zero-margin counts, the tie count and the largest margin): E0 -> shared memory by bulk
        //    copy, the first two planes -> tensor memory meanwhile
        mbar_wait(&ms->ready_bar[seq & 1u], (seq >> 1) & 1u, &p, 404);
        if (prof && lead_thread) prof[8] = gtimer();
        trc(c, 0, (seq << 8) | 3u);
        trc(c, 1, (uint32_t)c.b);
        const bool light = image_is_light(p, c.P, (int)im->plist[0]);
        LightStats lst = {0u, 0u, 0u, 0.f, 0.f};
        if (light) {
            // LIGHT image: the producers only de-interleaved; E0 and the statistics are formed here, on the way
            // of the planes into tensor memory
            if (c.P == 1) plane_load_light<NT_, QPT, 1>(c, lst);
            else plane_load_light<NT_, QPT, 2>(c, lst);
            c.res0 = 0;
            if (c.P > 1) { c.res1 = 1; c.mru = 1; }
        } else {
            const int n_t4 = (c.n_t + 3) & ~3;
            if (c.tid == 0) {
                const uint32_t bytes = (uint32_t)n_t4 * 4u;
                if (bytes) {
                    asm volatile("fence.proxy.async;\n" ::: "memory");   // generic accesses to s_E (previous image) first
                    mbar_expect_tx(&ms->e0_bar, bytes);
                    const unsigned char *src = reinterpret_cast<const unsigned char *>(c.planes + (size_t)(p.C + 1) * p.Npad + c.base);
                    for (uint32_t o = 0; o < bytes; o += 32768u)
                        bulk_g2s(reinterpret_cast<unsigned char *>(c.sE()) + o, src + o, min(32768u, bytes - o), &ms->e0_bar);
                } else {
                    mbar_arrive(&ms->e0_bar);
                }
            }
            if (!(p.flags & OC_F_NO_RESIDENT)) {
                plane_load<NT_, QPT>(c, 0, 0);
                c.res0 = 0;
                if (n_slots(QPT) > 1 && c.P > 1) { plane_load<NT_, QPT>(c, 1, 1); c.res1 = 1; c.mru = 1; }
            }
        }
        if (prof && lead_thread) prof[9] = gtimer();
        {   // both histograms start every image cleared (in CTA 0 the incoming digests of the previous image lay there)
            uint4 *h4 = reinterpret_cast<uint4 *>(sel(c)->hist);
            for (int i = c.tid; i < 2 * NBX / 4; i += NT_) h4[i] = make_uint4(0u, 0u, 0u, 0u);
        }
        if (light) {
            // my share of the tile's statistics -> one row per warp
            uint32_t dm = __float_as_uint(lst.dmx);
            for (int o = 16; o; o >>= 1) {
                lst.zc0 += __shfl_xor_sync(0xffffffffu, lst.zc0, o);
                lst.zc1 += __shfl_xor_sync(0xffffffffu, lst.zc1, o);
                lst.ties += __shfl_xor_sync(0xffffffffu, lst.ties, o);
                dm = max(dm, __shfl_xor_sync(0xffffffffu, dm, o));
                lst.esum += __shfl_xor_sync(0xffffffffu, lst.esum, o);
            }
            if (c.lane == 0) {
                uint32_t *r = ms->wred[c.warp];
                r[0] = lst.zc0; r[1] = lst.zc1; r[2] = 0u; r[3] = 0u; r[4] = dm; r[5] = lst.ties; r[6] = __float_as_uint(lst.esum);
            }
        } else {
            mbar_wait(&ms->e0_bar, c.n_e0 & 1u, &p, 405);
            c.n_e0++;
        }
        for (int i = c.n_t + c.tid; i < p.se_floats; i += NT_) c.sE()[i] = 0.f;      // padding pixels: E = 0
        if (prof && lead_thread) prof[1] = gtimer();
        csync<NT_>();

        trc(c, 0, (seq << 8) | 4u);
        // 2. one exchange: zero counts, largest margin, ties and the plain sum of E0 of every tile (as the
        //    producers formed them while streaming)
        if (c.tid < 7) {
            uint32_t v = 0u;
            float fs = 0.f;
            const int nrow = light ? NW : NPW;          // one row per compute warp (light) or per producer warp
            for (int w = 0; w < nrow; w++) {
                const uint32_t t = light ? ms->wred[w][c.tid] : ms->pstat[seq & 1u][w][c.tid];
                v = (c.tid == 4) ? max(v, t) : v + t;
                fs += __uint_as_float(t);
            }
            outbox(c)[c.tid] = (c.tid == 6) ? __float_as_uint(fs) : v;
        }
        if (c.tid == 7) outbox(c)[7] = 0u;
        float dmax_bound;
        double total0 = 0.0;                  // ~ sum of E0 over the image (float partial sums: good to ~1e-5)
        uint32_t tie_bound = 0u;
        {
            const uint32_t (*box)[MBW] = exchange(c, 8);
            uint32_t mx = 0u;
            for (int q = 0; q < c.G; q++) {
                mx = max(mx, box[q][4]);
                tie_bound += box[q][5];
                total0 += (double)__uint_as_float(box[q][6]);
            }
            dmax_bound = __uint_as_float(mx);
            if (c.tid < EAGER_P && c.tid < c.P) {
                uint32_t z = 0u;
                for (int q = 0; q < c.G; q++) z += box[q][c.tid];
                ms->zlb[c.tid] = z;
            }
        }
        csync<NT_>();                         // zlb
        if (prof && lead_thread) prof[2] = gtimer();

        trc(c, 0, (seq << 8) | 6u);
        // 4. iterations (weaklabel.cpp:82-101)
        // The zero-margin lower bounds only pay when quotas leave slack (a class that keeps more pixels
        // than its quota is skipped without a selection).  When the quotas of the label set add up to
        // the whole image every step selects anyway: the per-fold exchange that maintains them is dropped.
        bool track_zeros;
        {
            long long quota = 0;
            for (int k = 0; k < c.P; k++) quota += (ms->plist[k] == 0) ? p.nth_bg : p.nth_fg;
            track_zeros = quota < (long long)p.N - (p.N >> 6);
        }
        const uint8_t *ord_g = p.order + (size_t)c.b * p.num_iter * p.C;
        uint32_t nsel = 0;
        float bias_sum = 0.f;                 // every fold lifts the running max by at most its bias: bounds mean_max_1 - mean_max_0
        for (int j = 0; j < p.num_iter; j++) {
            for (int t = 0; t < c.P; t++) {
                const int plane = ord_in_smem ? ms->ord_s[j * p.C + t] : ord_g[j * p.C + t];
                const int cls = ms->plist[plane];
                const uint32_t k = (uint32_t)((cls == 0) ? p.nth_bg : p.nth_fg);      // weaklabel.cpp:88
                float bias = 0.f;
                if (k >= ms->zlb[plane]) {        // else: the rank falls among the zero margins, bias 0
                    nsel++;
                    trc(c, 0, (seq << 8) | 7u);
                    trc(c, 2, ((uint32_t)j << 8) | (uint32_t)t);
                    unsigned long long tq0 = 0ull, tq1 = 0ull, tq2 = 0ull;
                    if (prof && lead_thread) tq0 = gtimer();
                    const int sl = plane_acquire<NT_, QPT>(c, plane);
                    const uint32_t ta = slot_addr<NT_, QPT>(c, sl);
                    uint32_t zeros;
                    if (prof && lead_thread) tq1 = gtimer();
                    bias = select_margin<NT_, QPT>(c, ta, k, dmax_bound, &zeros);
                    if (prof && lead_thread) { tq2 = gtimer(); t_acq += tq1 - tq0; t_sel += tq2 - tq1; }
                    if (bias != 0.f) {
                        // fold: x += bias (weaklabel.cpp:90-92), E = max(E, x) (:94-99)
                        uint32_t gall = 0u, gnew = 0u, znew = 0u;
                        if (track_zeros) {
                            for_each_quad<NT_, QPT>(c, ta, [&](int kq, int q, float4 x) {
                                const float4 xn = make_float4(__fadd_rn(x.x, bias), __fadd_rn(x.y, bias), __fadd_rn(x.z, bias), __fadd_rn(x.w, bias));
                                tm_st4(ta + 4 * kq, xn);                  // warp-collective (quads beyond the tile stay -inf)
                                if (q >= c.TQ) return;
                                float4 *Ep = reinterpret_cast<float4 *>(c.sE() + 4 * q);
                                float4 E = *Ep;
                                gall += (xn.x > E.x) + (xn.y > E.y) + (xn.z > E.z) + (xn.w > E.w);
                                gnew += (xn.x > E.x && x.x < E.x) + (xn.y > E.y && x.y < E.y) + (xn.z > E.z && x.z < E.z) + (xn.w > E.w && x.w < E.w);
                                znew += (xn.x >= E.x) + (xn.y >= E.y) + (xn.z >= E.z) + (xn.w >= E.w);
                                E.x = fmaxf(E.x, xn.x); E.y = fmaxf(E.y, xn.y); E.z = fmaxf(E.z, xn.z); E.w = fmaxf(E.w, xn.w);
                                *Ep = E;
                            });
                        } else {
                            // every step of this image selects anyway: no zero-margin counts to maintain
                            for_each_quad<NT_, QPT>(c, ta, [&](int kq, int q, float4 x) {
                                const float4 xn = make_float4(__fadd_rn(x.x, bias), __fadd_rn(x.y, bias), __fadd_rn(x.z, bias), __fadd_rn(x.w, bias));
                                tm_st4(ta + 4 * kq, xn);
                                if (q >= c.TQ) return;
                                float4 *Ep = reinterpret_cast<float4 *>(c.sE() + 4 * q);
                                float4 E = *Ep;
                                E.x = fmaxf(E.x, xn.x); E.y = fmaxf(E.y, xn.y); E.z = fmaxf(E.z, xn.z); E.w = fmaxf(E.w, xn.w);
                                *Ep = E;
                            });
                        }
                        tm_wait_st();
                        SELP(c, 6);
                        if (sl) c.dirty1 = true; else c.dirty0 = true;
                        if (track_zeros) {
                            for (int o = 16; o; o >>= 1) {
                                gall += __shfl_xor_sync(0xffffffffu, gall, o);
                                gnew += __shfl_xor_sync(0xffffffffu, gnew, o);
                                znew += __shfl_xor_sync(0xffffffffu, znew, o);
                            }
                            if (c.lane == 0) { ms->wred[c.warp][0] = gall; ms->wred[c.warp][1] = gnew; ms->wred[c.warp][2] = znew; }
                            csync<NT_>();
                            if (c.warp == 0) {
                                uint32_t a0 = 0, a1 = 0, a2 = 0;
                                if (c.lane < NW) { a0 = ms->wred[c.lane][0]; a1 = ms->wred[c.lane][1]; a2 = ms->wred[c.lane][2]; }
                                for (int o = 16; o; o >>= 1) {
                                    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
                                    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
                                    a2 += __shfl_xor_sync(0xffffffffu, a2, o);
                                }
                                if (c.lane == 0) { uint32_t *ob = outbox(c); ob[0] = a0; ob[1] = a1; ob[2] = a2; ob[3] = 0u; }
                            }
                            const uint32_t (*box)[MBW] = exchange(c, 4);
                            SELP(c, 7);
                            uint32_t gall_s = 0u, gnew_s = 0u, znew_s = 0u;
                            for (int q = 0; q < c.G; q++) { gall_s += box[q][0]; gnew_s += box[q][1]; znew_s += box[q][2]; }
                            // Pixels that leave the zero-margin set of another class: those the folded class newly
                            // exceeds (it was below the max there) plus, at most, the pixels that ever held an
                            // exact tie at the max.  Folding creates a tie wherever the new score EQUALS the old max.
                            const uint32_t loss = (p.flags & OC_F_LOOSE_ZEROS) ? gall_s : gnew_s + tie_bound;
                            tie_bound += znew_s - gall_s;
                            if (c.tid < c.P) {
                                const uint32_t z = ms->zlb[c.tid];
                                ms->zlb[c.tid] = (c.tid == plane) ? znew_s : (z > loss ? z - loss : 0u);
                            }
                        } else if (c.tid < c.P) {
                            ms->zlb[c.tid] = 0u;                      // a valid lower bound without the exchange
                        }
                        dmax_bound = __fadd_rn(dmax_bound, bias);
                        bias_sum = __fadd_ru(bias_sum, bias);
                    } else if (zeros != 0xffffffffu) {
                        if (c.tid == 0) ms->zlb[plane] = zeros;       // exact count of zero margins
                    }
                    if (p.flags & OC_F_NO_RESIDENT) {                 // A/B: no residency across steps
                        if (sl) { if (c.dirty1) plane_store<NT_, QPT>(c, plane, 1); c.res1 = -1; c.dirty1 = false; }
                        else { if (c.dirty0) plane_store<NT_, QPT>(c, plane, 0); c.res0 = -1; c.dirty0 = false; }
                    }
                    csync<NT_>();
                    SELP(c, 8);
                    if (prof && lead_thread) t_fold += gtimer() - tq2;
                }
                if (lead_thread) p.bias_out[((size_t)c.b * p.num_iter + j) * p.C + t] = bias;
            }
        }
        if (prof && lead_thread) prof[3] = gtimer();

        trc(c, 0, (seq << 8) | 8u);
        if (prof && lead_thread) { prof[16] = t_acq; prof[17] = t_sel; prof[18] = t_fold; prof[20] = gtimer(); }
        // 5. labels, pass A.  The re-centring shift (weaklabel.cpp:108-112) is one constant per image added to
        //    every class: it can change an argmax only through the rounding of near-ties.  Its magnitude is
        //    bounded WITHOUT the two order-exact sums: mean_max_1 - mean_max_0 lies in [0, sum of the biases
        //    folded] (a fold lifts the running max by at most its bias), and the sequential float sum behind
        //    mean_max_0 is within N * 2^-24 < 2 % of the plain one.  Every pixel whose first maximum is safe
        //    against that bound is labelled now; NaN / overflow: no pixel is safe.
        float ashift, spos;                   // the shift lies in [-ashift, +spos]
        {
            const double m0 = total0 / (double)p.N;
            const double up = 0.0202 * fabs(m0) + 1e-30, ub = (double)bias_sum * 1.001 + up;
            ashift = (ub == ub && ub < 3.0e38) ? __double2float_ru(ub) : INFINITY;
            spos = (up == up && up < 3.0e38) ? __double2float_ru(up) : INFINITY;
        }
        unsigned long long unsure = 0ull;
        if (p.labels) unsure = labels_pass_a<NT_, QPT>(c, ashift, spos, c.tid, c.tm);
        // 6. does this image need the order-exact sums at all?  Only if the caller wants them (in-place map,
        //    probes) or a pixel was left open: one flag per CTA, one exchange.
        {
            const bool mine = __any_sync(0xffffffffu, unsure != 0ull);
            if (c.lane == 0) ms->wred[c.warp][0] = mine ? 1u : 0u;
            csync<NT_>();
            if (p.labels) labels_copy_out<NT_>(c);
            if (c.tid == 0) {
                uint32_t any = p.want_sums ? 1u : 0u;
                for (int w = 0; w < NW; w++) any |= ms->wred[w][0];
                uint32_t *ob = outbox(c);
                ob[0] = any; ob[1] = 0u; ob[2] = 0u; ob[3] = 0u;
            }
        }
        bool need_sums = false;
        {
            const uint32_t (*box)[MBW] = exchange(c, 4);
            for (int q = 0; q < c.G; q++) need_sums = need_sums || (box[q][0] != 0u);
        }
        if (prof && lead_thread) { prof[21] = gtimer(); prof[19] = need_sums ? 1ull : 0ull; }
        // (after this exchange no CTA of the cluster is inside a selection any more: the selection
        //  scratch of CTA 0, which the incoming digests alias, is free)
        float mean0 = __int_as_float(0x7fc00000);                                    // NaN: not computed for this image
        double mean1 = __longlong_as_double(0x7ff8000000000000ll), shift = mean1;
        if (need_sums) {
            // 7. mean_max_1 (weaklabel.cpp:102-106): the double running sum of the final E in shared memory;
            //    absent terms of a peer's digest come out of the peer's shared memory
            const double sum1 = exact_sum<double, NT_>(c, 1, [&](int r, int j) {
                return __uint_as_float(ld_peer(peer_addr(c.sE(), r) + 4u * (uint32_t)j));
            });
            if (prof && lead_thread) prof[22] = gtimer();
            //    mean_max_0 (weaklabel.cpp:69-80): the float running sum of E0, which the producers left in the
            //    L2 scratch -- the final E is dead now (every CTA has the double sum), E0 takes its place
            csync<NT_>();
            if (light) {
                if (c.P == 1) recompute_E0_light<NT_, 1>(c);
                else recompute_E0_light<NT_, 2>(c);
            } else if (c.tid == 0) {
                const uint32_t bytes = (uint32_t)((c.n_t + 3) & ~3) * 4u;
                if (bytes) {
                    asm volatile("fence.proxy.async;\n" ::: "memory");
                    mbar_expect_tx(&ms->e0_bar, bytes);
                    const unsigned char *src = reinterpret_cast<const unsigned char *>(c.planes + (size_t)(p.C + 1) * p.Npad + c.base);
                    for (uint32_t o = 0; o < bytes; o += 32768u)
                        bulk_g2s(reinterpret_cast<unsigned char *>(c.sE()) + o, src + o, min(32768u, bytes - o), &ms->e0_bar);
                } else {
                    mbar_arrive(&ms->e0_bar);
                }
            }
            if (!light) {
                mbar_wait(&ms->e0_bar, c.n_e0 & 1u, &p, 409);
                c.n_e0++;
            }
            for (int i = c.n_t + c.tid; i < p.se_floats; i += NT_) c.sE()[i] = 0.f;
            csync<NT_>();
            const float *e0 = c.planes + (size_t)(p.C + 1) * p.Npad;
            const float *pl0 = c.planes, *pl1 = c.planes + (c.P > 1 ? p.Npad : 0);
            const float sum0 = exact_sum<float, NT_>(c, 0, [&](int r, int j) {
                const size_t i = (size_t)r * p.TS + j;
                if (light) return fmaxf(fmaxf(__ldcg(pl0 + i), __ldcg(pl1 + i)), FLT_MIN);   // E0 was never written for it
                return __ldcg(e0 + i);
            });
            mean0 = __fdiv_rn(sum0, (float)p.N);
            mean1 = __ddiv_rn(sum1, (double)p.N);
            shift = __dsub_rn((double)mean0, mean1);
        }
        if (lead_thread) {
            p.mean0[c.b] = mean0;
            p.mean1[c.b] = mean1;
            p.shift[c.b] = shift;
        }
        if (prof && lead_thread) prof[4] = gtimer();
        trc(c, 0, (seq << 8) | 9u);
        //    pass B: the few pixels that need the exact shift
        if (p.labels && __any_sync(0xffffffffu, unsure != 0ull)) {
            PassB a;
            a.p = &p; a.plist = ms->plist; a.planes = c.planes; a.absmax = c.absmax; a.absidx = c.absidx;
            a.out = p.labels + (size_t)c.b * p.N + c.base;
            a.shift = shift; a.unsure = unsure;
            a.t0 = slot_addr<NT_, QPT>(c, 0); a.t1 = slot_addr<NT_, QPT>(c, 1 % n_slots(QPT)); a.mask = c.mask;
            a.r0 = c.res0; a.r1 = n_slots(QPT) > 1 ? c.res1 : -1;
            a.b = c.b; a.base = c.base; a.n_t = c.n_t; a.tid = c.tid; a.P = c.P; a.first_absent = c.first_absent;
            a.nt = NT_; a.qpt = QPT; a.two_slots = n_slots(QPT) > 1 ? 1 : 0;
            labels_pass_b(a);
        }
        csync<NT_>();
        if (prof && lead_thread) {
            prof[5] = gtimer();
            prof[6] = nsel;
#ifdef WL_SELPROF
            for (int i = 0; i < 9; i++) { prof[23 + i] = ms->tacc[i]; ms->tacc[i] = 0ull; }
#endif
            prof[7] = (unsigned long long)c.P;
        }
    }
    trc(c, 0, 0xEEEEu);
    csync<NT_>();
    if (c.warp == 0) tm_dealloc512(ms->tmem_base);
    hw_cluster_sync();                                  // nobody leaves while a peer may still address its shared memory
    trc(c, 0, 0xFFFFu);
}

template <int NT_, int QPT>
static cudaError_t launch_t(const EStepParams &p, int n_clusters, cudaStream_t stream, int *max_clusters) {
    const size_t smem = Layout<NT_>::total(p.se_floats);
    cudaError_t e = cudaFuncSetAttribute(estep_onchip_kernel<NT_, QPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(estep_onchip_kernel<NT_, QPT>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
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
        e = cudaOccupancyMaxActiveClusters(&n, estep_onchip_kernel<NT_, QPT>, &cfg);
        *max_clusters = (e == cudaSuccess) ? n : 0;
        return e;
    }
    return cudaLaunchKernelEx(&cfg, estep_onchip_kernel<NT_, QPT>, p);
}

static cudaError_t dispatch(const EStepParams &p0, int qpt, int n_clusters, cudaStream_t stream, int *maxc) {
    EStepParams p = p0;
    p.light_ok = n_slots(qpt) > 1 ? 1 : 0;
    if (qpt == 1) return launch_t<OC_NT, 1>(p, n_clusters, stream, maxc);
    if (qpt == 3) return launch_t<OC_NT, 3>(p, n_clusters, stream, maxc);
    if (qpt == 9) return launch_t<OC_NT, 9>(p, n_clusters, stream, maxc);
    if (qpt == 10) return launch_t<OC_NT, 10>(p, n_clusters, stream, maxc);
    return cudaErrorInvalidValue;
}

size_t smem_bytes(int se_floats) { return Layout<OC_NT>::total(se_floats); }

}  // namespace oc

// Geometry of the v4 kernel for an image of N pixels with C classes and a given cluster size.
static bool onchip_try(int N, int C, int G, int tune_TS, OnchipConfig *out) {
    const int nt = OC_NT;
    const int cap = 10 * nt * 4;
    if (G < 1 || G > OC_MAXG) return false;
    int TS = tune_TS ? tune_TS : (N + G - 1) / G;
    TS = (TS + 3) / 4 * 4;
    if (TS < 4) TS = 4;
    if (TS > cap || (long long)TS * G < N) return false;
    const int needq = (TS / 4 + nt - 1) / nt;
    out->nt = nt;
    out->qpt = needq <= 1 ? 1 : (needq <= 3 ? 3 : (needq <= 9 ? 9 : 10));
    out->G = G;
    out->TS = TS;
    out->L = ((TS + nt - 1) / nt) | 1;                   // odd: conflict-free chunk walks
    if (out->L > oc::DIG_L) return false;
    out->se_floats = ((TS > nt * out->L ? TS : nt * out->L) + 3) / 4 * 4;
    int cpx = (oc::STAGE_BYTES - 32) / (4 * C);
    if (cpx < 1) return false;
    out->cpx = cpx;
    if (oc::smem_bytes(out->se_floats) > 232448) return false;
    EStepParams p = {};
    p.G = G;
    p.se_floats = out->se_floats;
    int maxc = 0;
    if (oc::dispatch(p, out->qpt, 0, nullptr, &maxc) != cudaSuccess || maxc <= 0) {
        cudaGetLastError();
        return false;
    }
    out->max_clusters = maxc;
    return true;
}

// tune_G = 0: automatic -- the smallest power-of-two cluster that holds the image (two planes in tensor
// memory need at most 9 quads per thread); large batches of large images take clusters of 9 (non-portable
// size): on B200 as many of them are co-resident as clusters of 8 (15), so 135 instead of 120 SMs work.
bool onchip_configure(int N, int C, int B, int sms, int tune_G, int tune_TS, OnchipConfig *out) {
    if (tune_G) return onchip_try(N, C, tune_G, tune_TS, out);
    const int cap = 9 * OC_NT * 4;
    int G = 1;
    while (G < 8 && (N + G - 1) / G > cap) G *= 2;
    while (G < 8 && (long long)B * G * 2 <= sms && N / (2 * G) >= 4096) G *= 2;
    bool ok = onchip_try(N, C, G, tune_TS, out);
    if (G == 8 && tune_TS == 0) {
        OnchipConfig alt;
        if (onchip_try(N, C, 9, 0, &alt) && (!ok || (alt.max_clusters >= out->max_clusters && B >= alt.max_clusters) || alt.qpt < out->qpt)) { *out = alt; ok = true; }
    }
    return ok;
}

cudaError_t launch_estep_onchip(const EStepParams &p, const OnchipConfig &cfg, int n_clusters, cudaStream_t stream) {
    return oc::dispatch(p, cfg.qpt, n_clusters, stream, nullptr);
}

}  // namespace wl
