// tests/cpu/seqsum_test.cpp -- host check of csrc/seqsum.cuh: the chunked, predicted-binade
// evaluation must reproduce the plain sequential float / double running sums of
// estep/weaklabel.cpp:69-80 and :102-106 bit for bit, for any data and any chunking.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cfloat>
#include <vector>
#include <random>
#include "../../em_adapt_tensorflow_b200/csrc/seqsum.cuh"

template <typename T>
static T walk(const std::vector<float> &t, int L, int group, long *n_seq_chunks) {
    int n = (int)t.size();
    int nchunks = (n + L - 1) / L;
    std::vector<double> pre(nchunks + 1, 0.0);
    for (int c = 0; c < nchunks; c++) {
        double s = 0;
        for (int i = c * L; i < n && i < (c + 1) * L; i++) s += (double)t[i];
        pre[c + 1] = pre[c] + s;
    }
    std::vector<wl::ChunkFn<T>> fn(nchunks);
    for (int c = 0; c < nchunks; c++) {
        int len = std::min(L, n - c * L);
        fn[c] = wl::chunk_eval<T>(&t[c * L], len, pre[c]);
    }
    T s = 0;
    for (int g0 = 0; g0 < nchunks; g0 += group) {
        int g1 = std::min(nchunks, g0 + group);
        wl::ChunkFn<T> all = wl::chunk_identity<T>();
        for (int c = g0; c < g1; c++) all = wl::chunk_compose(all, fn[c]);
        if (wl::chunk_apply(all, s)) continue;
        for (int c = g0; c < g1; c++) {
            if (wl::chunk_apply(fn[c], s)) continue;
            int len = std::min(L, n - c * L);
            s = wl::chunk_sequential<T>(&t[c * L], len, s);
            (*n_seq_chunks)++;
        }
    }
    return s;
}

template <typename T> static T plain(const std::vector<float> &t) {
    T s = 0;
    for (float x : t) s += x;
    return s;
}

int main(int argc, char **argv) {
    int trials = argc > 1 ? atoi(argv[1]) : 200;
    std::mt19937_64 rng(12345);
    std::normal_distribution<float> nd(0.f, 3.f);
    long bad = 0, seqf = 0, seqd = 0, chunks = 0;
    for (int trial = 0; trial < trials; trial++) {
        int kind = trial % 10;
        int n = 1 + (int)(rng() % (trial % 7 == 0 ? 300000 : 20000));
        std::vector<float> t(n);
        for (int i = 0; i < n; i++) {
            float x = nd(rng);
            switch (kind) {
            case 0: x = std::max(FLT_MIN, x); break;                       // floored max, like ext
            case 1: x = std::max(FLT_MIN, x - 2.f); break;                 // mostly floor
            case 2: x = FLT_MIN; break;                                    // all floor
            case 3: x = std::fabs(x) * 1e-30f + FLT_MIN; break;            // tiny
            case 4: x = std::fabs(x) * 1e4f + 3e4f; break;                 // big
            case 5: x = (float)(1 + rng() % 7); break;                     // small integers (ties)
            case 6: x = std::ldexp(1.0f, (int)(rng() % 60) - 30); break;   // powers of two, wild range
            case 7: x = std::max(FLT_MIN, std::round(x * 8) / 8); break;   // coarse grid: many exact ties
            case 8: x = (i % 1000 == 0) ? 1e6f : std::max(FLT_MIN, x); break; // spikes
            default: x = std::fabs(x) * std::ldexp(1.0f, (int)(rng() % 40) - 20) + FLT_MIN; break;
            }
            t[i] = x;
        }
        int Ls[] = {1, 3, 33, 5, 64, 129};
        int L = Ls[trial % 6];
        int group = (trial % 3 == 0) ? 32 : (trial % 3 == 1 ? 1 : 7);
        float pf = plain<float>(t);
        double pd = plain<double>(t);
        long sf = 0, sd = 0;
        float wf = walk<float>(t, L, group, &sf);
        double wd = walk<double>(t, L, group, &sd);
        chunks += (n + L - 1) / L; seqf += sf; seqd += sd;
        if (memcmp(&pf, &wf, 4) || memcmp(&pd, &wd, 8)) {
            bad++;
            printf("MISMATCH trial %d kind %d n %d L %d: float %a vs %a, double %a vs %a\n",
                   trial, kind, n, L, pf, wf, pd, wd);
        }
    }
    printf("trials %d bad %ld chunks %ld sequential-fallback chunks: float %ld double %ld\n",
           trials, bad, chunks, seqf, seqd);
    return bad ? 1 : 0;
}
