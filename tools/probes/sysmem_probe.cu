// sysmem_probe.cu -- measurement aid (not product code): how a RESIDENT kernel should read pinned host memory.
//   1. latency of 1 and of 4 independent 16-byte loads per lane, one warp and 200 warps, per load flavour;
//   2. freshness: the host rewrites a buffer between doorbells while the kernel stays resident; which flavours
//      ever return a stale value (a line kept in L1/L2 from the previous lap)?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sysmem_probe sysmem_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>

enum { PLAIN = 0, CG = 1, VOLATILE = 2, RELAXED_SYS = 3, CV = 4, NFLAV = 5 };
static const char *NAMES[] = {"ld.global (plain)", "ld.global.cg", "ld.volatile (generic)", "ld.relaxed.sys.global", "ld.global.cv"};

template <int F> __device__ __forceinline__ uint4 ld16(const uint4 *p) {
    uint4 v;
    if (F == PLAIN) asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (F == CG) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (F == VOLATILE) asm volatile("ld.volatile.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (F == RELAXED_SYS) asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (F == CV) asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned long long gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

// each warp: `reps` rounds of NL independent row loads (512 B per row per warp), rows never revisited
template <int F, int NL> __global__ void lat_kernel(const uint4 *buf, size_t rows, int reps, unsigned long long *out, uint32_t *sink) {
    const uint32_t lane = threadIdx.x & 31u, warp = blockIdx.x;
    uint32_t acc = 0;
    unsigned long long total = 0;
    for (int r = 0; r < reps; r++) {
        const size_t row0 = ((size_t)warp * reps + r) * NL;
        __syncwarp();
        const unsigned long long t0 = gt();
        uint4 v[NL];
#pragma unroll
        for (int i = 0; i < NL; i++) v[i] = ld16<F>(buf + ((row0 + i) % rows) * 32 + lane);
#pragma unroll
        for (int i = 0; i < NL; i++) acc += v[i].x ^ v[i].w;
        acc = __shfl_xor_sync(0xffffffffu, acc, 1) + acc;
        const unsigned long long t1 = gt();
        total += t1 - t0;
    }
    if (lane == 0) out[warp] = total;
    if (acc == 0x12345u) sink[0] = acc;
}

// resident reader: waits for flag == it+1, reads `n16` uint4 of buf with flavour F, counts words != it, acks
template <int F> __global__ void fresh_kernel(const uint4 *buf, uint32_t n16, volatile uint32_t *flag, volatile uint32_t *ack, uint32_t iters,
                                              uint32_t *stale) {
    uint32_t bad = 0;
    for (uint32_t it = 0; it < iters; it++) {
        while (*flag != it + 1u) {}
        for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x) {
            const uint4 v = ld16<F>(buf + i);
            bad += (v.x != it) + (v.y != it) + (v.z != it) + (v.w != it);
        }
        __syncthreads();
        if (threadIdx.x == 0) { __threadfence_system(); *ack = it + 1u; }
    }
    atomicAdd(stale, bad);
}

// GPU -> host stores: each warp writes `rows` rows; pattern 0: 16 B per lane, 512 contiguous bytes per instruction;
// 1: 16 B per lane at a 64-byte stride (2 KiB span per instruction); 2: 4 B per lane, 128 contiguous bytes;
// 3: pattern 0 but each row followed by a system-scope fence (what a per-tile completion costs)
template <int PAT> __global__ void wr_kernel(uint4 *buf, size_t rows_total, int rows, uint32_t seed) {
    const uint32_t lane = threadIdx.x & 31u, warp = blockIdx.x;
    for (int r = 0; r < rows; r++) {
        const size_t row = ((size_t)warp * rows + r) % rows_total;
        uint4 v = make_uint4(seed + r, lane, warp, 7u);
        if (PAT == 0 || PAT == 3) buf[row * 32 + lane] = v;
        if (PAT == 1) buf[(row * 32 + lane * 4) % (rows_total * 32)] = v;
        if (PAT == 2) reinterpret_cast<uint32_t *>(buf)[row * 128 + lane] = v.x;
        if (PAT == 3) asm volatile("fence.acq_rel.sys;" ::: "memory");
    }
    asm volatile("fence.acq_rel.sys;" ::: "memory");
}
template <int PAT> static void run_wr(uint4 *buf, size_t rows_total, int warps, int rows, const char *what) {
    wr_kernel<PAT><<<warps, 32>>>(buf, rows_total, 4, 1);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    wr_kernel<PAT><<<warps, 32>>>(buf, rows_total, rows, 2);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = (double)warps * rows * (PAT == 2 ? 128.0 : 512.0);
    printf("  %-52s %3d warps x %4d rows: %8.2f us, %.1f GB/s\n", what, warps, rows, ms * 1e3, bytes / (ms * 1e-3) / 1e9);
}

// Does a device-scope fence in one warp wait for ANOTHER warp's stores to host memory?  Warp 0 of block 0 times
// {store to device memory; fence.acq_rel.gpu} while (mode 1) warp 1 of the same block, or (mode 2) a warp of
// another block, streams 512-byte rows to pinned host memory; mode 0: nobody does.
__global__ void fence_kernel(uint4 *host, uint32_t *dev, int mode, unsigned long long *out, volatile uint32_t *stop) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const bool timer = blockIdx.x == 0 && warp == 0;
    const bool writer = (mode == 1 && blockIdx.x == 0 && warp == 1) || (mode == 2 && blockIdx.x == 1 && warp == 0);
    if (timer) {
        unsigned long long total = 0;
        for (int i = 0; i < 200; i++) {
            __nanosleep(2000);
            const unsigned long long t0 = gt();
            dev[lane] = i;
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            total += gt() - t0;
        }
        if (lane == 0) { out[0] = total; *stop = 1; }
    } else if (writer) {
        for (uint32_t r = 0; *stop == 0; r++) host[(size_t)(r & 4095u) * 32 + lane] = make_uint4(r, lane, 1, 2);
    }
}
static void run_fence(uint4 *buf, int mode, const char *what) {
    uint32_t *dev; unsigned long long *out; uint32_t *stop;
    cudaMalloc(&dev, 4096); cudaMalloc(&out, 8); cudaMalloc(&stop, 4); cudaMemset(stop, 0, 4);
    fence_kernel<<<2, 64>>>(buf, dev, mode, out, stop);
    cudaDeviceSynchronize();
    unsigned long long t; cudaMemcpy(&t, out, 8, cudaMemcpyDeviceToHost);
    printf("  device store + fence.acq_rel.gpu, %-44s %.2f us\n", what, (double)t / 200 * 1e-3);
    cudaFree(dev); cudaFree(out); cudaFree(stop);
}

// The relay's own read pattern: per tile 160 B of descriptors, then (dependent) a 640- or 960-byte payload window as
// 16 B/lane loads; `tiles` tiles per launch spread over `warps` resident warps.  Optionally a second kernel streams
// results to the host at the same time.
__global__ void tile_read_kernel(const uint4 *descs, const uint4 *payload, int tiles, uint32_t *sink) {
    const uint32_t lane = threadIdx.x & 31u;
    const int warps = gridDim.x * (blockDim.x >> 5), w0 = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    uint32_t acc = 0;
    for (int t = w0; t < tiles; t += warps) {
        uint4 d = make_uint4(0, 0, 0, 0);
        if (lane < 10) d = descs[(size_t)t * 10 + lane];
        const uint32_t off = __shfl_sync(0xffffffffu, d.y, 0) & 0u;          // dependent, but keep the address computable
        const int pieces = (t & 1) ? 60 : 40;                                   // 960 B (open tile) or 640 B (seal tile)
        const uint4 *win = payload + (size_t)t * 60 + off;
        uint4 a = make_uint4(0, 0, 0, 0), b = a;
        if ((int)lane < pieces) a = win[lane];
        if ((int)lane + 32 < pieces) b = win[lane + 32];
        acc += a.x ^ b.w ^ d.x;
    }
    if (acc == 0x1234567u) sink[0] = acc;
}
static void run_tiles(uint4 *buf, size_t rows, int warps, bool with_writer) {
    const int tiles = 205 * 64;
    uint32_t *sink; cudaMalloc(&sink, 4);
    cudaStream_t s1, s2; cudaStreamCreate(&s1); cudaStreamCreate(&s2);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const uint4 *descs = buf, *payload = buf + (size_t)tiles * 10 + 64;
    uint4 *wr = buf + rows * 16;   // second half of the buffer
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(e0, s1);
        tile_read_kernel<<<warps / 4, 128, 0, s1>>>(descs, payload, tiles, sink);
        cudaEventRecord(e1, s1);
        if (with_writer) wr_kernel<0><<<148, 32, 0, s2>>>(wr, rows / 2, 512, 3);   // ~38 MB of stores alongside
        cudaDeviceSynchronize();
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = (double)tiles * (160 + 800);
    printf("  relay read pattern (160 B descriptors -> 640/960 B window), %4d warps%s: %.2f us per 205 tiles, %.1f GB/s\n", warps,
           with_writer ? ", while another kernel streams stores to the host" : "", ms * 1e3 / 64, bytes / (ms * 1e-3) / 1e9);
    cudaFree(sink);
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

template <int F, int NL> static void run_lat(const uint4 *buf, size_t rows, int warps) {
    const int reps = 64;
    unsigned long long *d_out; uint32_t *sink;
    CK(cudaMalloc(&d_out, sizeof(unsigned long long) * warps)); CK(cudaMalloc(&sink, 4));
    lat_kernel<F, NL><<<warps, 32>>>(buf, rows, 4, d_out, sink);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    lat_kernel<F, NL><<<warps, 32>>>(buf, rows, reps, d_out, sink);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    unsigned long long *h = (unsigned long long *)malloc(sizeof(unsigned long long) * warps);
    CK(cudaMemcpy(h, d_out, sizeof(unsigned long long) * warps, cudaMemcpyDeviceToHost));
    double sum = 0; for (int i = 0; i < warps; i++) sum += (double)h[i];
    printf("  %-24s %d load(s)/lane, %3d warps: %.2f us per round, %.1f GB/s aggregate\n", NAMES[F], NL, warps,
           sum / warps / reps * 1e-3, (double)warps * reps * NL * 512.0 / (ms * 1e-3) / 1e9);
    free(h); cudaFree(d_out); cudaFree(sink);
}

template <int F> static void run_fresh(uint4 *buf, uint32_t n16, volatile uint32_t *flag, volatile uint32_t *ack) {
    uint32_t *d_stale; CK(cudaMalloc(&d_stale, 4)); CK(cudaMemset(d_stale, 0, 4));
    const uint32_t iters = 400;
    *flag = 0; *ack = 0;
    uint32_t *w = (uint32_t *)buf;
    for (uint32_t i = 0; i < n16 * 4; i++) w[i] = 0xffffffffu;
    fresh_kernel<F><<<1, 256>>>(buf, n16, flag, ack, iters, d_stale);
    for (uint32_t it = 0; it < iters; it++) {
        for (uint32_t i = 0; i < n16 * 4; i++) w[i] = it;
        std::atomic_thread_fence(std::memory_order_seq_cst);
        *flag = it + 1u;
        while (*ack != it + 1u) {}
    }
    CK(cudaDeviceSynchronize());
    uint32_t stale; CK(cudaMemcpy(&stale, d_stale, 4, cudaMemcpyDeviceToHost));
    printf("  %-24s %u stale words of %u read\n", NAMES[F], stale, iters * n16 * 4);
    cudaFree(d_stale);
}

int main() {
    const size_t rows = 1u << 17;   // 64 MiB of pinned host memory in 512-byte rows
    uint4 *buf; CK(cudaHostAlloc(&buf, rows * 512, cudaHostAllocMapped));
    memset(buf, 1, rows * 512);
    printf("latency of pinned-host loads from a kernel (rows of 512 B per warp, never revisited):\n");
    run_lat<PLAIN, 1>(buf, rows, 1); run_lat<CG, 1>(buf, rows, 1); run_lat<VOLATILE, 1>(buf, rows, 1); run_lat<RELAXED_SYS, 1>(buf, rows, 1); run_lat<CV, 1>(buf, rows, 1);
    run_lat<PLAIN, 4>(buf, rows, 1); run_lat<CG, 4>(buf, rows, 1); run_lat<VOLATILE, 4>(buf, rows, 1); run_lat<RELAXED_SYS, 4>(buf, rows, 1); run_lat<CV, 4>(buf, rows, 1);
    run_lat<PLAIN, 4>(buf, rows, 200); run_lat<CG, 4>(buf, rows, 200); run_lat<VOLATILE, 4>(buf, rows, 200); run_lat<RELAXED_SYS, 4>(buf, rows, 200); run_lat<CV, 4>(buf, rows, 200);
    run_lat<PLAIN, 2>(buf, rows, 600); run_lat<CG, 2>(buf, rows, 600); run_lat<VOLATILE, 2>(buf, rows, 600);
    printf("stores from a kernel to pinned host memory:\n");
    for (int w : {1, 32, 200, 592}) {
        run_wr<0>(buf, rows, w, 256, "16 B/lane, 512 contiguous bytes per instruction");
        run_wr<1>(buf, rows, w, 256, "16 B/lane at a 64-byte stride");
        run_wr<2>(buf, rows, w, 256, "4 B/lane, 128 contiguous bytes per instruction");
        run_wr<3>(buf, rows, w, 64, "512-byte rows, system fence after each");
    }
    {   // what the copy engine charges for the same bytes: one cudaMemcpyAsync D2H of 172 KB per batch
        void *d; cudaMalloc(&d, 1 << 20);
        cudaStream_t st; cudaStreamCreate(&st);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0, st);
            for (int i = 0; i < 200; i++) cudaMemcpyAsync((char *)buf + (size_t)i * (172 << 10), d, 172 << 10, cudaMemcpyDeviceToHost, st);
            cudaEventRecord(e1, st);
            cudaStreamSynchronize(st);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("  copy engine, 200 back-to-back D2H copies of 172 KiB: %.2f us each, %.1f GB/s\n", ms * 1e3 / 200, 200.0 * (172 << 10) / (ms * 1e-3) / 1e9);
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0, st);
            for (int i = 0; i < 200; i++) cudaMemcpyAsync(d, (char *)buf + (size_t)i * (198 << 10), 198 << 10, cudaMemcpyHostToDevice, st);
            cudaEventRecord(e1, st);
            cudaStreamSynchronize(st);
        }
        cudaEventElapsedTime(&ms, e0, e1);
        printf("  copy engine, 200 back-to-back H2D copies of 198 KiB: %.2f us each, %.1f GB/s\n", ms * 1e3 / 200, 200.0 * (198 << 10) / (ms * 1e-3) / 1e9);
    }
    for (int w : {148, 444, 888, 1776}) { run_tiles(buf, rows, w, false); run_tiles(buf, rows, w, true); }
    run_fence(buf, 0, "nobody writes to the host:");
    run_fence(buf, 1, "a warp of the SAME block streams host stores:");
    run_fence(buf, 2, "a warp of ANOTHER block streams host stores:");
    printf("freshness inside ONE resident kernel (host rewrites 64 KiB between doorbells, 400 laps):\n");
    volatile uint32_t *flags; CK(cudaHostAlloc((void **)&flags, 256, cudaHostAllocMapped));
    run_fresh<PLAIN>(buf, 4096, flags, flags + 32);
    run_fresh<CG>(buf, 4096, flags, flags + 32);
    run_fresh<VOLATILE>(buf, 4096, flags, flags + 32);
    run_fresh<RELAXED_SYS>(buf, 4096, flags, flags + 32);
    run_fresh<CV>(buf, 4096, flags, flags + 32);
    return 0;
}
