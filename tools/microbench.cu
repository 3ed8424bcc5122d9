// tools/microbench.cu -- design-time measurements on the B200 (not part of the product):
//   L2-resident vs HBM read bandwidth, shared-memory histogram atomics, cluster barrier latency,
//   DSMEM push exchange, L2-atomic histogram merge, max resident clusters.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <stdio.h>
#include <stdint.h>
#include <vector>
namespace cg = cooperative_groups;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

// ---------------- 1. read bandwidth over a working set of `n4` float4, `reps` sweeps ----------------
__global__ void __launch_bounds__(1024) read_bw(const float4 *__restrict__ p, size_t n4, int reps, float *sink) {
    float acc = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; r++) {
        size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < n4; i += 4 * stride) {
            float4 a = __ldcg(p + i), b = __ldcg(p + i + stride), c = __ldcg(p + i + 2 * stride), d = __ldcg(p + i + 3 * stride);
            acc += a.x + b.y + c.z + d.w;
        }
        for (; i < n4; i += stride) acc += __ldcg(p + i).x;
    }
    if (acc == 123.456f) *sink = acc;
}

// ---------------- 2. smem histogram atomics ----------------
template <int KEYS>
__global__ void __launch_bounds__(1024) smem_hist(const uint32_t *__restrict__ keys, int nbins, int reps, unsigned long long *cycles, uint32_t *out) {
    extern __shared__ uint32_t hist[];
    for (int i = threadIdx.x; i < nbins; i += blockDim.x) hist[i] = 0;
    uint32_t k[KEYS];
#pragma unroll
    for (int j = 0; j < KEYS; j++) k[j] = keys[(size_t)blockIdx.x * blockDim.x * KEYS + j * blockDim.x + threadIdx.x] % nbins;
    __syncthreads();
    long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
#pragma unroll
        for (int j = 0; j < KEYS; j++) atomicAdd(&hist[k[j]], 1u);
        __syncthreads();
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (threadIdx.x < nbins && hist[threadIdx.x] == 0xffffffffu) out[0] = 1;
}

// only a fraction of the keys (those inside a bracket) do the atomic; the rest only count
template <int KEYS>
__global__ void __launch_bounds__(1024) smem_hist_bracket(const uint32_t *__restrict__ keys, int nbins, uint32_t lo, uint32_t hi, int reps,
                                                          unsigned long long *cycles, uint32_t *out) {
    extern __shared__ uint32_t hist[];
    for (int i = threadIdx.x; i < nbins; i += blockDim.x) hist[i] = 0;
    uint32_t k[KEYS];
#pragma unroll
    for (int j = 0; j < KEYS; j++) k[j] = keys[(size_t)blockIdx.x * blockDim.x * KEYS + j * blockDim.x + threadIdx.x];
    __syncthreads();
    uint32_t below = 0;
    long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
#pragma unroll
        for (int j = 0; j < KEYS; j++) {
            uint32_t key = k[j] + r;
            if (key < lo) below++;
            else if (key < hi) atomicAdd(&hist[(key - lo) % nbins], 1u);
        }
        __syncthreads();
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (below == 0xffffffffu) out[0] = below;
}

// ---------------- 3. cluster barrier latency ----------------
__global__ void __launch_bounds__(1024) cluster_bar(int reps, unsigned long long *cycles) {
    cg::cluster_group cl = cg::this_cluster();
    cl.sync();
    long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
        asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
}
// relaxed arrive + wait (no memory ordering)
__global__ void __launch_bounds__(1024) cluster_bar_relaxed(int reps, unsigned long long *cycles) {
    cg::cluster_group cl = cg::this_cluster();
    cl.sync();
    long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
        asm volatile("barrier.cluster.arrive.relaxed.aligned;\nbarrier.cluster.wait.aligned;\n" ::: "memory");
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
}

// ---------------- 4. DSMEM push exchange: every CTA writes `words` words into every peer, barrier, sums ----------------
__global__ void __launch_bounds__(1024) dsmem_push(int words, int reps, unsigned long long *cycles, uint32_t *out) {
    cg::cluster_group cl = cg::this_cluster();
    extern __shared__ uint32_t box[];          // [2][csize][words]
    const int G = cl.num_blocks(), me = cl.block_rank();
    for (int i = threadIdx.x; i < 2 * G * words; i += blockDim.x) box[i] = 0;
    cl.sync();
    uint32_t acc = 0;
    long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
        uint32_t *slot = box + (size_t)(r & 1) * G * words;
        for (int i = threadIdx.x; i < G * words; i += blockDim.x) {
            const int peer = i / words, w = i % words;
            uint32_t *dst = cl.map_shared_rank(slot + me * words + w, peer);
            *dst = (uint32_t)(r + w + me);
        }
        asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
        for (int i = threadIdx.x; i < G * words; i += blockDim.x) acc += slot[i];
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (acc == 0xdeadbeefu) out[0] = acc;
}

// ---------------- 5. L2-atomic histogram merge (what estep_kernel.cu v1 does per pass) ----------------
__global__ void __launch_bounds__(1024) l2_merge(uint32_t *ghist /*[clusters][3][2048]*/, int nbins, int nonzero, int reps,
                                                 unsigned long long *cycles, uint32_t *out) {
    cg::cluster_group cl = cg::this_cluster();
    const int G = cl.num_blocks();
    uint32_t *base = ghist + (size_t)(blockIdx.x / G) * 3 * nbins;
    cl.sync();
    uint32_t acc = 0;
    long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
        uint32_t *h = base + (size_t)(r % 3) * nbins;
        for (int i = threadIdx.x; i < nonzero; i += blockDim.x) atomicAdd(&h[(i * 7 + r) % nbins], 1u);
        asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
        for (int i = threadIdx.x; i < nbins; i += blockDim.x) acc += __ldcg(&h[i]);
        acc += __syncthreads_count(acc & 1);
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (acc == 0xdeadbeefu) out[0] = acc;
}

template <typename K, typename... Args>
static void launch_cluster(K kernel, int nclusters, int G, int threads, size_t smem, Args... args) {
    CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nclusters * G);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = G; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, kernel, args...));
}

template <typename K>
static int max_clusters(K kernel, int G, int threads, size_t smem) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(G * 64); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = G; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    int n = -1;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess) { cudaGetLastError(); return -1; }
    return n;
}

static double avg(const std::vector<unsigned long long> &v, int n) { double s = 0; for (int i = 0; i < n; i++) s += (double)v[i]; return s / n; }

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    printf("device %s sms %d l2 %d MB smem/blk optin %zu clock %d kHz\n", prop.name, prop.multiProcessorCount, prop.l2CacheSize >> 20, prop.sharedMemPerBlockOptin, prop.clockRate);
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float *sink; CK(cudaMalloc(&sink, 64));
    unsigned long long *dcy; CK(cudaMalloc(&dcy, 4096 * 8));
    uint32_t *dout; CK(cudaMalloc(&dout, 64));
    std::vector<unsigned long long> hcy(4096);

    // 1. bandwidth
    {
        size_t big = (size_t)4 << 30;
        float4 *buf; CK(cudaMalloc(&buf, big)); CK(cudaMemset(buf, 0, big));
        size_t sizes_mb[] = {16, 32, 48, 64, 96, 2048};
        for (size_t mb : sizes_mb) {
            size_t n4 = (mb << 20) / 16;
            int reps = mb >= 1024 ? 2 : 40;
            for (int grid : {148, 296}) {
                read_bw<<<grid, 1024>>>(buf, n4, 2, sink);            // warm
                CK(cudaEventRecord(e0));
                read_bw<<<grid, 1024>>>(buf, n4, reps, sink);
                CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
                float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
                printf("read_bw working_set %4zu MB grid %3d x1024: %8.1f GB/s\n", mb, grid, (double)(mb << 20) * reps / (ms * 1e-3) / 1e9);
            }
        }
        CK(cudaFree(buf));
    }
    // 2. smem histogram atomics
    {
        const int KEYS = 17;
        size_t n = (size_t)148 * 1024 * KEYS;
        std::vector<uint32_t> hk(n);
        uint64_t s = 88172645463325252ull;
        for (size_t i = 0; i < n; i++) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; hk[i] = (uint32_t)(s >> 20); }
        uint32_t *dk; CK(cudaMalloc(&dk, n * 4)); CK(cudaMemcpy(dk, hk.data(), n * 4, cudaMemcpyHostToDevice));
        for (int nbins : {2048, 256, 32}) {
            CK(cudaFuncSetAttribute(smem_hist<KEYS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2048 * 4));
            smem_hist<KEYS><<<148, 1024, 2048 * 4>>>(dk, nbins, 20, dcy, dout);
            CK(cudaMemcpy(hcy.data(), dcy, 148 * 8, cudaMemcpyDeviceToHost));
            printf("smem_hist uniform keys nbins %4d: %7.0f cycles per pass of 1024x%d atomics (%.2f cyc/atomic/SM)\n", nbins, avg(hcy, 148) / 20, KEYS,
                   avg(hcy, 148) / 20 / (1024.0 * KEYS));
        }
        for (double frac : {1.0, 0.25, 0.06, 0.015}) {
            uint32_t lo = 1u << 20, hi = lo + (uint32_t)(frac * (double)(1ull << 44 >> 20) * 0 + frac * 17592186044416.0 / 4096.0);
            // keys are 44-bit >> 20 -> 24-bit... keep simple: keys uniform in [0, 2^44>>20 = 2^24)
            hi = lo + (uint32_t)(frac * ((1u << 24) - lo));
            smem_hist_bracket<KEYS><<<148, 1024, 2048 * 4>>>(dk, 2048, lo, hi, 20, dcy, dout);
            CK(cudaMemcpy(hcy.data(), dcy, 148 * 8, cudaMemcpyDeviceToHost));
            printf("smem_hist_bracket in-bracket fraction %.3f: %7.0f cycles per pass of 1024x%d keys\n", frac * (1.0 - 1.0 / 16), avg(hcy, 148) / 20, KEYS);
        }
        CK(cudaFree(dk));
    }
    // 3. occupancy + cluster barrier
    for (int G : {2, 4, 8, 16}) {
        int m0 = max_clusters(cluster_bar, G, 1024, 0), m1 = max_clusters(dsmem_push, G, 1024, 200 * 1024);
        int m2 = max_clusters(dsmem_push, G, 512, 100 * 1024);
        printf("cluster size %2d: max active clusters 1024thr/0KB %d, 1024thr/200KB %d, 512thr/100KB %d\n", G, m0, m1, m2);
        if (m0 <= 0) continue;
        launch_cluster(cluster_bar, m0, G, 1024, 0, 200, dcy);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(hcy.data(), dcy, (size_t)m0 * G * 8, cudaMemcpyDeviceToHost));
        printf("  barrier.cluster release/acquire (all %d clusters busy): %.0f cycles\n", m0, avg(hcy, m0 * G) / 200);
        launch_cluster(cluster_bar_relaxed, m0, G, 1024, 0, 200, dcy);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(hcy.data(), dcy, (size_t)m0 * G * 8, cudaMemcpyDeviceToHost));
        printf("  barrier.cluster relaxed: %.0f cycles\n", avg(hcy, m0 * G) / 200);
        for (int words : {8, 64, 256}) {
            size_t smem = (size_t)2 * G * words * 4;
            int mc = max_clusters(dsmem_push, G, 1024, smem);
            launch_cluster(dsmem_push, mc, G, 1024, smem, words, 100, dcy, dout);
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(hcy.data(), dcy, (size_t)mc * G * 8, cudaMemcpyDeviceToHost));
            printf("  dsmem push %3d words to each of %d peers + barrier + local sum: %.0f cycles/round\n", words, G, avg(hcy, mc * G) / 100);
        }
        {
            uint32_t *gh; CK(cudaMalloc(&gh, (size_t)m0 * 3 * 2048 * 4)); CK(cudaMemset(gh, 0, (size_t)m0 * 3 * 2048 * 4));
            for (int nz : {2048, 256, 32}) {
                launch_cluster(l2_merge, m0, G, 1024, 0, gh, 2048, nz, 100, dcy, dout);
                CK(cudaDeviceSynchronize());
                CK(cudaMemcpy(hcy.data(), dcy, (size_t)m0 * G * 8, cudaMemcpyDeviceToHost));
                printf("  L2-atomic merge %4d nonzero bins/CTA + barrier + read 2048 bins: %.0f cycles/round\n", nz, avg(hcy, m0 * G) / 100);
            }
            CK(cudaFree(gh));
        }
    }
    printf("done\n");
    return 0;
}
