// estep_fast.cu -- the on-chip fast path of the EM-Adapt E-step for sm_100a (B200).
//
// Same contract as estep_kernel.cu (one thread-block cluster per image, persistent, images
// claimed from a counter) for the common configuration: suppress_others = true with a
// non-negative margin, contiguous pixels, and an image that fits G x 33792 pixels.  What
// changes is where the iteration state lives and how the cluster agrees on an order statistic:
//
//   * the floored running max E ("extrenum", estep/weaklabel.cpp:67-77, 94-99) of a CTA's pixel
//     tile stays in SHARED memory for the whole image; the margins of the class being visited are
//     recomputed from E and a REGISTER copy of that class's plane (one coalesced L2 read per
//     pixel per step); planes live in an L2-resident scratch and are rewritten only when a
//     non-zero bias is folded in (weaklabel.cpp:90-92);
//   * a step whose rank falls among the zero margins (the class already owns more than its
//     quota) is decided from per-class zero counts kept up to date by the fold pass -- no pass
//     over the pixels, no cluster synchronisation;
//   * otherwise ONE histogram pass bins the margins on a window of 31 binades x 64 bins below the
//     largest possible margin; the cluster merges a 64-entry summary through distributed shared
//     memory (one barrier.cluster), reads the 32 fine bins it needs from the peers' shared
//     memory, gathers the few candidates of the selected bin into every CTA (second barrier) and
//     finishes the exact selection locally.  Heavier ties fall back to further narrowing rounds.
//
// The value selected is the exact order statistic std::nth_element leaves at diff[nth]
// (weaklabel.cpp:85-89); sums, re-centring and labels are shared with the generic kernel.
#include "estep_device.cuh"

namespace wl {

constexpr int NB2 = 2048;               // fine bins per histogram round
constexpr int NB1 = 64;                 // super bins (32 fine bins each) exchanged through DSMEM
constexpr int SUBB = NB2 / NB1;
constexpr int GCAP = 1024;              // candidates gathered into every CTA
constexpr int MBW = 80;                 // mailbox words per CTA per exchange
constexpr int WIN_UNDER = 33;           // bin 0: zero margins; 1..32: below the window, by exponent group
constexpr int WIN_BINADES = 31;
constexpr int WIN_SH = 17;              // 64 bins per binade inside the window
constexpr int WIN_BINS = WIN_BINADES * 64;          // 1984
constexpr int WIN_OVER0 = WIN_UNDER + WIN_BINS;     // 2017 .. 2047: above the window
constexpr int EAGER_P = 4;              // label sets up to this size keep per-class zero counts

struct FastSmem {
    uint32_t hist[2][NB2];              // rotating per histogram round; peers read 32 fine bins
    uint32_t lhist[NB2];                // CTA-local selection among gathered candidates
    uint32_t mbox[2][MAX_CLUSTER][MBW]; // written by the peers, double buffered per exchange
    uint32_t outbox[MBW];
    uint32_t gather[GCAP];
    uint32_t mylist[GCAP];
    uint32_t fine[MAX_CLUSTER][SUBB];
    uint32_t wred[NWARPS][8];
    uint32_t lcount;
    int sel_S;
    uint32_t sel_preS;
    int pad;
    SharedMisc misc;
};

struct FCtx : Ctx {
    FastSmem *fs;
    float *s_E;           // [TS] floored running max of my tile
    int base, n_t;        // my tile: pixels [base, base + n_t) of the image
    uint32_t xround;      // exchanges done (mailbox parity), identical in every CTA
    uint32_t hround;      // histogram rounds done (hist buffer parity)
};

// ---------------------------------------------------------------------------------------------
// distributed shared memory
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t peer_addr(const void *local_smem, int rank) {
    uint32_t a = (uint32_t)__cvta_generic_to_shared(local_smem), r;
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

// Every CTA hands the first n words of its outbox to every CTA of the cluster (itself included);
// returns the mailbox that now holds them: box[p][w] = word w of CTA p.
__device__ __forceinline__ const uint32_t (*exchange(FCtx &c, int n))[MBW] {
    FastSmem *s = c.fs;
    const uint32_t par = c.xround & 1u;
    __syncthreads();                                   // outbox complete
    if (c.G > 1) {
        for (int i = c.tid; i < c.G * n; i += NT) {
            const int peer = i / n, w = i - peer * n;
            st_peer(peer_addr(&s->mbox[par][c.rank][w], peer), s->outbox[w]);
        }
    } else {
        for (int w = c.tid; w < n; w += NT) s->mbox[par][0][w] = s->outbox[w];
    }
    cluster_barrier(c.G);
    c.xround++;
    return s->mbox[par];
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

// ---------------------------------------------------------------------------------------------
// one histogram round: bin my tile's margins, agree on the bin that holds rank r
// ---------------------------------------------------------------------------------------------
struct Located {
    int bin;              // -1: inconsistent (NaN in the scores)
    uint32_t pre, cnt;    // keys of the histogram below the bin / inside it
    int hpar;
};

template <int PPT, bool WINDOW>
__device__ __forceinline__ void hist_pass(FCtx &c, const float (&x)[PPT], uint32_t lo, uint32_t hi, int sh,
                                          uint32_t *h) {
    h[2 * c.tid] = 0u;
    h[2 * c.tid + 1] = 0u;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < PPT; j++) {
        const int i = j * NT + c.tid;
        if (i < c.n_t) {
            const uint32_t key = __float_as_uint(__fsub_rn(c.s_E[i], x[j]));
            if (WINDOW) atomicAdd(&h[win_bin(key, lo)], 1u);
            else if (key >= lo && key <= hi) atomicAdd(&h[(key - lo) >> sh], 1u);
        }
    }
    __syncthreads();
}

__device__ __forceinline__ Located locate(FCtx &c, uint32_t r) {
    FastSmem *s = c.fs;
    const int hpar = (int)(c.hround & 1u);
    uint32_t *h = s->hist[hpar];
    {   // 64 super-bin sums of my histogram -> outbox
        uint32_t v = h[2 * c.tid] + h[2 * c.tid + 1];
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        if ((c.lane & 15) == 0) s->outbox[c.tid >> 4] = v;
    }
    const uint32_t (*box)[MBW] = exchange(c, NB1);
    if (c.tid == 0) s->sel_S = -1;
    __syncthreads();
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
    __syncthreads();
    const int S = s->sel_S;
    const uint32_t preS = s->sel_preS;
    Located L;
    L.hpar = hpar;
    L.bin = -1; L.pre = 0; L.cnt = 0;
    c.hround++;
    if (S < 0) return L;                               // uniform over the cluster
    if (c.tid < c.G * SUBB) {
        const int p = c.tid >> 5, i = c.tid & 31;
        s->fine[p][i] = (p == c.rank) ? h[S * SUBB + i] : ld_peer(peer_addr(&h[S * SUBB + i], p));
    }
    if (c.tid == 0) s->misc.sel_bin = -1;
    __syncthreads();
    if (c.warp == 0) {
        uint32_t f = 0;
        for (int p = 0; p < c.G; p++) f += s->fine[p][c.lane];
        uint32_t inc = f;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (c.lane >= o) inc += t;
        }
        const uint32_t ex = inc - f, rr = r - preS;
        if (rr >= ex && rr < ex + f) { s->misc.sel_bin = S * SUBB + c.lane; s->misc.sel_pre = preS + ex; s->misc.sel_cnt = f; }
    }
    __syncthreads();
    L.bin = s->misc.sel_bin;
    L.pre = s->misc.sel_pre;
    L.cnt = s->misc.sel_cnt;
    return L;
}

// exact selection of rank r among cnt keys in [lo, hi] held in CTA-local shared memory
__device__ __forceinline__ uint32_t local_select(FCtx &c, const uint32_t *keys, uint32_t cnt, uint32_t r,
                                                 uint32_t lo, uint32_t hi) {
    FastSmem *s = c.fs;
    uint32_t *h = s->lhist;
    for (int guard = 0; guard < 4; guard++) {
        if (lo == hi) break;
        int sh = 0;
        while (((hi - lo) >> sh) >= (uint32_t)NB2) sh++;
        h[2 * c.tid] = 0u;
        h[2 * c.tid + 1] = 0u;
        __syncthreads();
        for (uint32_t i = c.tid; i < cnt; i += NT) {
            const uint32_t key = keys[i];
            if (key >= lo && key <= hi) atomicAdd(&h[(key - lo) >> sh], 1u);
        }
        __syncthreads();
        const uint32_t h0 = h[2 * c.tid], h1 = h[2 * c.tid + 1];
        if (c.tid == 0) s->misc.sel_bin = -1;
        uint32_t total;
        const uint32_t ex = block_excl_scan_u32(c, h0 + h1, &total);
        if (r >= ex && r < ex + h0) { s->misc.sel_bin = 2 * c.tid; s->misc.sel_pre = ex; }
        else if (r >= ex + h0 && r < ex + h0 + h1) { s->misc.sel_bin = 2 * c.tid + 1; s->misc.sel_pre = ex + h0; }
        __syncthreads();
        const int bin = s->misc.sel_bin;
        const uint32_t pre = s->misc.sel_pre;
        __syncthreads();
        if (bin < 0) { if (c.tid == 0) atomicExch(c.p->status, 300 + guard); return lo; }
        r -= pre;
        lo += (uint32_t)bin << sh;
        const uint32_t top = lo + ((1u << sh) - 1u);
        if (top < hi) hi = top;
    }
    return lo;
}

// The value of rank k (0-based) among the margins E[i] - x[i] of the whole image.
template <int PPT>
__device__ __forceinline__ float select_margin_fast(FCtx &c, const float (&x)[PPT], uint32_t k, uint32_t lo_key) {
    FastSmem *s = c.fs;
    uint32_t lo = lo_key, hi = 0u, r = k;
    int sh = 0;
    bool window = true;
    for (int guard = 0; guard < 6; guard++) {
        uint32_t *h = s->hist[c.hround & 1u];
        if (window) hist_pass<PPT, true>(c, x, lo, hi, sh, h);
        else hist_pass<PPT, false>(c, x, lo, hi, sh, h);
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
        r -= L.pre;
        if (klo == khi) return __uint_as_float(klo);
        if (L.cnt <= (uint32_t)GCAP) {
            // gather the candidates of the bin into every CTA, in rank order of the CTAs
            if (c.tid == 0) s->lcount = 0u;
            __syncthreads();
#pragma unroll
            for (int j = 0; j < PPT; j++) {
                const int i = j * NT + c.tid;
                if (i < c.n_t) {
                    const uint32_t key = __float_as_uint(__fsub_rn(c.s_E[i], x[j]));
                    if (key >= klo && key <= khi) {
                        const uint32_t pos = atomicAdd(&s->lcount, 1u);
                        if (pos < (uint32_t)GCAP) s->mylist[pos] = key;
                    }
                }
            }
            __syncthreads();
            const int fb = L.bin & (SUBB - 1);
            const uint32_t mine = s->fine[c.rank][fb];
            uint32_t off = 0;
            for (int p = 0; p < c.rank; p++) off += s->fine[p][fb];
            if (c.G > 1) {
                for (uint32_t i = c.tid; i < (uint32_t)c.G * mine; i += NT) {
                    const uint32_t peer = i / mine, w = i - peer * mine;
                    st_peer(peer_addr(&s->gather[off + w], (int)peer), s->mylist[w]);
                }
            } else {
                for (uint32_t w = c.tid; w < mine; w += NT) s->gather[w] = s->mylist[w];
            }
            cluster_barrier(c.G);
            return __uint_as_float(local_select(c, s->gather, L.cnt, r, klo, khi));
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
// small block reductions into the outbox (5 values: four counters and one max)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void block_reduce_to_outbox(FCtx &c, uint32_t z0, uint32_t z1, uint32_t z2,
                                                       uint32_t z3, uint32_t mx) {
    FastSmem *s = c.fs;
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
    __syncthreads();
    if (c.warp == 0) {
        uint32_t a0 = s->wred[c.lane][0], a1 = s->wred[c.lane][1], a2 = s->wred[c.lane][2],
                 a3 = s->wred[c.lane][3], am = s->wred[c.lane][4];
        for (int o = 16; o; o >>= 1) {
            a0 += __shfl_xor_sync(0xffffffffu, a0, o);
            a1 += __shfl_xor_sync(0xffffffffu, a1, o);
            a2 += __shfl_xor_sync(0xffffffffu, a2, o);
            a3 += __shfl_xor_sync(0xffffffffu, a3, o);
            am = max(am, __shfl_xor_sync(0xffffffffu, am, o));
        }
        if (c.lane == 0) { s->outbox[0] = a0; s->outbox[1] = a1; s->outbox[2] = a2; s->outbox[3] = a3; s->outbox[4] = am; }
    }
}

// ---------------------------------------------------------------------------------------------
// order-exact running sum of E over the whole image (weaklabel.cpp:69-80 float, :102-106 double)
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T exact_sum_E(FCtx &c, ChunkFn<T> *s_fn, uint32_t &tot_parity) {
    typedef fp_traits<T> F;
    const EStepParams &p = *c.p;
    ChunkFn<T> *s_wfn = s_fn + NT;
    const float *terms = c.s_E;
    const int n_t = c.n_t;
    const int c0 = c.tid * p.L;
    const int len = max(0, min(p.L, n_t - c0));
    double cs = 0.0;
    for (int i = 0; i < len; i++) cs += (double)terms[c0 + i];
    double tile_total;
    const double cpre = block_excl_scan_f64(c, cs, &tile_total);
    double *tot = c.ctl->tot[tot_parity & 1u];
    if (c.tid == 0) tot[c.rank] = tile_total;
    cluster_barrier(c.G);
    double before = 0.0;
    for (int r = 0; r < c.rank; r++) before += __ldcg(tot + r);
    ChunkFn<T> f = chunk_eval<T>(terms + c0, len, before + cpre);
    ChunkFn<T> w = f;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        ChunkFn<T> o = FnShfl<T>::down(w, d);
        if ((c.lane & (2 * d - 1)) == 0) w = chunk_compose(w, o);
    }
    s_fn[c.tid] = f;
    if (c.lane == 0) s_wfn[c.warp] = w;
    __syncthreads();
    for (int r = 0; r < c.G; r++) {
        if (c.rank == r && c.tid == 0) {
            T s = (r == 0) ? (T)0 : F::from((typename F::U)__ldcg(&c.ctl->carry));
            for (int wv = 0; wv < NWARPS; wv++) {
                if (chunk_apply(s_wfn[wv], s)) continue;
                for (int l = 0; l < 32; l++) {
                    const int ch = wv * 32 + l;
                    if (chunk_apply(s_fn[ch], s)) continue;
                    const int cc0 = ch * p.L;
                    s = chunk_sequential<T>(terms + cc0, max(0, min(p.L, n_t - cc0)), s);
                }
            }
            c.ctl->carry = (unsigned long long)F::bits(s);
        }
        cluster_barrier(c.G);
    }
    tot_parity++;
    T res = F::from((typename F::U)__ldcg(&c.ctl->carry));
    __syncthreads();
    return res;
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
constexpr size_t FSMEM_FIXED = ((sizeof(FastSmem) + 15) / 16) * 16;
constexpr size_t FSMEM_STREAM = 2 * (size_t)STREAM_BUF_FLOATS * sizeof(float);

static size_t fast_smem_bytes(int ppt) {
    const size_t e_and_sum = (size_t)ppt * NT * sizeof(float) + (NT + NWARPS) * sizeof(ChunkFn<double>);
    return FSMEM_FIXED + (e_and_sum > FSMEM_STREAM ? e_and_sum : FSMEM_STREAM);
}

template <int PPT>
__global__ void __launch_bounds__(NT, 1) estep_fast_kernel(const __grid_constant__ EStepParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FCtx c;
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
    c.fs = reinterpret_cast<FastSmem *>(smem_raw);
    c.sm = &c.fs->misc;
    c.s_hist = c.fs->lhist;
    c.s_dyn = reinterpret_cast<float *>(smem_raw + FSMEM_FIXED);
    c.s_E = c.s_dyn;
    ChunkFn<double> *s_fn_d = reinterpret_cast<ChunkFn<double> *>(c.s_dyn + (size_t)PPT * NT);
    ChunkFn<float> *s_fn_f = reinterpret_cast<ChunkFn<float> *>(s_fn_d);
    c.pass_no = 0;
    c.xround = 0;
    c.hround = 0;
    c.base = c.rank * p.TS;
    c.n_t = max(0, min(p.TS, p.N - c.base));
    uint32_t tot_parity = 0;

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

        // 1. stream-in: HBM -> shared staging -> suppression -> compact planes in the L2 scratch
        stream_in(c);
        __syncthreads();          // my tile's planes were written by this CTA; the staging buffer
                                  // is about to be reused as E

        // 2. E0 = max(FLT_MIN, max over the present classes) (weaklabel.cpp:67-77; with suppression
        //    and margin >= 0 no absent class exceeds the smallest present one), zero-margin counts
        //    per class and the largest margin
        const bool eager = c.P <= EAGER_P;
        const float *pl = c.planes + c.base;
        uint32_t zc[EAGER_P] = {0u, 0u, 0u, 0u};
        uint32_t zero_cnt[EAGER_P] = {0u, 0u, 0u, 0u};
        float dmax_bound;
        {
            float dmx = 0.f;
#pragma unroll 1
            for (int j = 0; j < PPT; j++) {
                const int i = j * NT + c.tid;
                if (i < c.n_t) {
                    float E = FLT_MIN, mn = FLT_MAX;
                    if (eager) {
                        float v[EAGER_P];
#pragma unroll
                        for (int q = 0; q < EAGER_P; q++) {
                            v[q] = (q < c.P) ? __ldcg(pl + (size_t)q * p.Npad + i) : -FLT_MAX;
                            if (v[q] > E) E = v[q];
                            if (q < c.P && v[q] < mn) mn = v[q];
                        }
#pragma unroll
                        for (int q = 0; q < EAGER_P; q++) zc[q] += (v[q] == E) ? 1u : 0u;
                    } else {
                        for (int q = 0; q < c.P; q++) {
                            const float v = __ldcg(pl + (size_t)q * p.Npad + i);
                            if (v > E) E = v;
                            if (v < mn) mn = v;
                        }
                    }
                    c.s_E[i] = E;
                    dmx = fmaxf(dmx, __fsub_rn(E, mn));
                }
            }
            block_reduce_to_outbox(c, zc[0], zc[1], zc[2], zc[3], __float_as_uint(dmx));
            const uint32_t (*box)[MBW] = exchange(c, 5);
            uint32_t mx = 0u;
            for (int q = 0; q < c.G; q++) {
#pragma unroll
                for (int e = 0; e < EAGER_P; e++) zero_cnt[e] += box[q][e];
                mx = max(mx, box[q][4]);
            }
            dmax_bound = __uint_as_float(mx);
        }

        // 3. mean_max_0 (weaklabel.cpp:69-80): float running sum in pixel order
        const float sum0 = exact_sum_E<float>(c, s_fn_f, tot_parity);
        const float mean0 = __fdiv_rn(sum0, (float)p.N);

        // 4. iterations (weaklabel.cpp:82-101)
        const uint8_t *ord = p.order + (size_t)c.b * p.num_iter * p.C;
        for (int j = 0; j < p.num_iter; j++) {
            for (int t = 0; t < c.P; t++) {
                const int plane = ord[j * p.C + t];
                const int cls = c.sm->plist[plane];
                const uint32_t k = (uint32_t)((cls == 0) ? p.nth_bg : p.nth_fg);      // weaklabel.cpp:88
                float bias = 0.f;
                bool skip = false;
                if (eager) {
                    uint32_t z = zero_cnt[0];
                    if (plane == 1) z = zero_cnt[1];
                    if (plane == 2) z = zero_cnt[2];
                    if (plane == 3) z = zero_cnt[3];
                    skip = k < z;           // the rank falls among the zero margins: bias 0, nothing changes
                }
                if (!skip) {
                    float x[PPT];
                    float *plc = c.planes + (size_t)plane * p.Npad + c.base;
#pragma unroll
                    for (int q = 0; q < PPT; q++) {
                        const int i = q * NT + c.tid;
                        x[q] = (i < c.n_t) ? __ldcg(plc + i) : 0.f;
                    }
                    uint32_t e_top = (__float_as_uint(dmax_bound) >> 23) + 1u;
                    if (e_top > 254u) e_top = 254u;
                    const uint32_t lo_key = (e_top > (uint32_t)WIN_BINADES ? e_top - WIN_BINADES : 0u) << 23;
                    bias = select_margin_fast<PPT>(c, x, k, lo_key);
                    if (bias != 0.f) {
                        // fold: x += bias (weaklabel.cpp:90-92), E = max(E, x) (:94-99); recount zeros
                        uint32_t nz[EAGER_P] = {0u, 0u, 0u, 0u};
#pragma unroll
                        for (int q = 0; q < PPT; q++) {
                            const int i = q * NT + c.tid;
                            if (i < c.n_t) {
                                const float xn = __fadd_rn(x[q], bias);
                                __stcg(plc + i, xn);
                                float E = c.s_E[i];
                                if (xn > E) { E = xn; c.s_E[i] = E; }
                                if (eager) {
#pragma unroll
                                    for (int e = 0; e < EAGER_P; e++) {
                                        if (e < c.P) {
                                            const float v = (e == plane) ? xn : __ldcg(pl + (size_t)e * p.Npad + i);
                                            nz[e] += (v == E) ? 1u : 0u;
                                        }
                                    }
                                }
                            }
                        }
                        dmax_bound = __fadd_rn(dmax_bound, bias);
                        if (eager) {
                            block_reduce_to_outbox(c, nz[0], nz[1], nz[2], nz[3], 0u);
                            const uint32_t (*box)[MBW] = exchange(c, 4);
#pragma unroll
                            for (int e = 0; e < EAGER_P; e++) {
                                zero_cnt[e] = 0u;
                                for (int q = 0; q < c.G; q++) zero_cnt[e] += box[q][e];
                            }
                        } else {
                            __syncthreads();
                        }
                    }
                }
                if (c.rank == 0 && c.tid == 0)
                    p.bias_out[((size_t)c.b * p.num_iter + j) * p.C + t] = bias;
            }
        }

        // 5. mean_max_1 (weaklabel.cpp:102-106), shift (:110), labels
        __syncthreads();
        const double sum1 = exact_sum_E<double>(c, s_fn_d, tot_parity);
        const double mean1 = __ddiv_rn(sum1, (double)p.N);
        const double shift = __dsub_rn((double)mean0, mean1);
        if (c.rank == 0 && c.tid == 0) {
            p.mean0[c.b] = mean0;
            p.mean1[c.b] = mean1;
            p.shift[c.b] = shift;
        }
        if (p.labels) write_labels(c, shift);
        __syncthreads();
    }
}

template <int PPT>
static cudaError_t launch_fast_t(const EStepParams &p, int n_clusters, cudaStream_t stream) {
    const size_t smem = fast_smem_bytes(PPT);
    cudaError_t e = cudaFuncSetAttribute(estep_fast_kernel<PPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(estep_fast_kernel<PPT>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
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
    return cudaLaunchKernelEx(&cfg, estep_fast_kernel<PPT>, p);
}

template <int PPT>
static int fast_max_clusters_t(int G) {
    const size_t smem = fast_smem_bytes(PPT);
    if (cudaFuncSetAttribute(estep_fast_kernel<PPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
    cudaFuncSetAttribute(estep_fast_kernel<PPT>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
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
    if (cudaOccupancyMaxActiveClusters(&n, estep_fast_kernel<PPT>, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int fast_ppt_for(int tile_pixels) {
    const int need = (tile_pixels + NT - 1) / NT;
    if (need <= FAST_PPT_SMALL) return FAST_PPT_SMALL;
    if (need <= FAST_PPT_MAX) return FAST_PPT_MAX;
    return 0;
}

cudaError_t launch_estep_fast(const EStepParams &p, int n_clusters, cudaStream_t stream) {
    switch (fast_ppt_for(p.TS)) {
    case FAST_PPT_SMALL: return launch_fast_t<FAST_PPT_SMALL>(p, n_clusters, stream);
    case FAST_PPT_MAX: return launch_fast_t<FAST_PPT_MAX>(p, n_clusters, stream);
    default: return cudaErrorInvalidValue;
    }
}

int estep_fast_max_clusters(int G, int tile_pixels) {
    switch (fast_ppt_for(tile_pixels)) {
    case FAST_PPT_SMALL: return fast_max_clusters_t<FAST_PPT_SMALL>(G);
    case FAST_PPT_MAX: return fast_max_clusters_t<FAST_PPT_MAX>(G);
    default: return 0;
    }
}

}  // namespace wl
