// seqsum.cuh -- order-exact parallel evaluation of a sequential floating-point sum.
//
// The reference re-centres every image by (mean_max_0 - mean_max_1) where mean_max_0 is a
// FLOAT running sum over the pixels in index order (estep/weaklabel.cpp:69,78) and
// mean_max_1 a DOUBLE running sum (weaklabel.cpp:102-105).  Bit-exact maps (and, at
// rounding-merge ties, bit-exact labels) need those sums with the reference's rounding
// order, i.e. s_{i+1} = fl(s_i + e_i), which is not associative.
//
// All terms are positive (every term is >= FLT_MIN, weaklabel.cpp:71), so the running
// sum is monotone.  While it stays inside one binade [2^k, 2^(k+1)) with ulp u, the
// state is an integer m (s = m*u) and adding a term e = q*u + r gives
//     m' = m + q + [r > u/2] + [r == u/2 and (m+q) odd]          (round to nearest even)
// so a whole CHUNK of terms acts on the state as  m -> m + A[m & 1]  with two integers
// A[0], A[1] -- as long as the sum does not leave the binade inside the chunk.  Such
// maps compose associatively:  (F then G)[p] = F[p] + G[(p + F[p]) & 1].
//
// A[0] and A[1] are obtained with the hardware's own adder: run the chunk from the even
// state 2^k and from the odd state 2^k + u.  Which binade a chunk will be entered in is
// PREDICTED from an ordinary (double) prefix sum; the sequential walk over the chunk
// summaries VERIFIES the prediction (state exponent == chunk exponent before and after)
// and otherwise falls back to adding that chunk's terms one by one.  The result is
// therefore exact for any prediction; the prediction only decides how much is serial.
//
// Host-compilable (tests/cpu/seqsum_test.cpp drives it against a plain loop).
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define WL_HD __host__ __device__ __forceinline__
#else
#define WL_HD inline
#endif

namespace wl {

template <typename T> struct fp_traits;
template <> struct fp_traits<float> {
    typedef uint32_t U;
    typedef int32_t S;
    static constexpr U exp_mask = 0x7f800000u;
    static constexpr U min_exp = (U)40 << 23;   // below this the walk adds term by term
    static WL_HD U bits(float x) {
#if defined(__CUDA_ARCH__)
        return __float_as_uint(x);
#else
        U u; memcpy(&u, &x, 4); return u;
#endif
    }
    static WL_HD float from(U u) {
#if defined(__CUDA_ARCH__)
        return __uint_as_float(u);
#else
        float x; memcpy(&x, &u, 4); return x;
#endif
    }
    static WL_HD float add(float a, float b) {
#if defined(__CUDA_ARCH__)
        return __fadd_rn(a, b);
#else
        return a + b;
#endif
    }
};
template <> struct fp_traits<double> {
    typedef uint64_t U;
    typedef int64_t S;
    static constexpr U exp_mask = 0x7ff0000000000000ull;
    static constexpr U min_exp = (U)960 << 52;  // ~2^-63; smaller sums go term by term
    static WL_HD U bits(double x) {
#if defined(__CUDA_ARCH__)
        return (U)__double_as_longlong(x);
#else
        U u; memcpy(&u, &x, 8); return u;
#endif
    }
    static WL_HD double from(U u) {
#if defined(__CUDA_ARCH__)
        return __longlong_as_double((long long)u);
#else
        double x; memcpy(&x, &u, 8); return x;
#endif
    }
    static WL_HD double add(double a, double b) {
#if defined(__CUDA_ARCH__)
        return __dadd_rn(a, b);
#else
        return a + b;
#endif
    }
};

// Action of a chunk of terms on the running sum.  kind: 0 = unusable (walk must add the
// terms one by one), 1 = m -> m + a[m&1] valid while the state's exponent field == e,
// 2 = identity (empty chunk).
template <typename T> struct ChunkFn {
    typename fp_traits<T>::U e;
    typename fp_traits<T>::S a0, a1;
    int kind;
};

template <typename T> WL_HD ChunkFn<T> chunk_identity() {
    ChunkFn<T> f; f.e = 0; f.a0 = 0; f.a1 = 0; f.kind = 2; return f;
}

// Summarise terms[0..n) (stride 1) assuming the running sum enters the chunk near
// `predicted` (any finite non-negative double; a bad guess costs speed, not exactness).
template <typename T>
WL_HD ChunkFn<T> chunk_eval(const float *terms, int n, double predicted) {
    typedef fp_traits<T> F;
    typedef typename F::U U;
    typedef typename F::S S;
    ChunkFn<T> f;
    if (n <= 0) return chunk_identity<T>();
    f.kind = 0; f.a0 = 0; f.a1 = 0;
    U e = F::bits((T)predicted) & F::exp_mask;
    f.e = e;
    if (e < F::min_exp || e == F::exp_mask) return f;
    T s0 = F::from(e), s1 = F::from(e + 1);
    for (int i = 0; i < n; i++) {
        T t = (T)terms[i];
        s0 = F::add(s0, t);
        s1 = F::add(s1, t);
    }
    U b0 = F::bits(s0), b1 = F::bits(s1);
    if ((b0 & F::exp_mask) != e || (b1 & F::exp_mask) != e) return f;
    f.a0 = (S)(b0 - e);
    f.a1 = (S)(b1 - (e + 1));
    f.kind = 1;
    return f;
}

// H = "F first, then G".  Only meaningful when both are usable on the same binade.
template <typename T> WL_HD ChunkFn<T> chunk_compose(const ChunkFn<T> &f, const ChunkFn<T> &g) {
    if (f.kind == 2) return g;
    if (g.kind == 2) return f;
    ChunkFn<T> h;
    h.e = f.e;
    if (f.kind != 1 || g.kind != 1 || f.e != g.e) { h.kind = 0; h.a0 = 0; h.a1 = 0; return h; }
    h.kind = 1;
    h.a0 = f.a0 + ((f.a0 & 1) ? g.a1 : g.a0);
    h.a1 = f.a1 + (((1 + f.a1) & 1) ? g.a1 : g.a0);
    return h;
}

// Try to advance the running sum across a chunk.  Returns false when the summary does
// not apply to this state (caller then adds the chunk's terms sequentially).
template <typename T> WL_HD bool chunk_apply(const ChunkFn<T> &f, T &s) {
    typedef fp_traits<T> F;
    typedef typename F::U U;
    if (f.kind == 2) return true;
    if (f.kind != 1) return false;
    U b = F::bits(s);
    if ((b & F::exp_mask) != f.e) return false;
    U nb = b + (U)((b & 1) ? f.a1 : f.a0);
    if ((nb & F::exp_mask) != f.e) return false;
    s = F::from(nb);
    return true;
}

template <typename T> WL_HD T chunk_sequential(const float *terms, int n, T s) {
    typedef fp_traits<T> F;
    for (int i = 0; i < n; i++) s = F::add(s, (T)terms[i]);
    return s;
}

}  // namespace wl
