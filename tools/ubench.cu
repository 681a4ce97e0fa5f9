// Microbenchmarks that size the episode kernel's design on the B200 in front of us (nothing here ships):
//   1. FP64 pipe: dependent-issue latency of DFMA/DADD/DMUL and throughput vs (warps per SM sub-partition, ILP)
//   2. instruction cache: throughput of a straight-line FP64 loop body as its SASS grows from 4 KB to 256 KB,
//      with 2 / 3 / 4 warps per sub-partition walking it out of phase
//   3. co-issue: FP64 stream with interleaved integer / shared-memory instructions
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu ; run on the GPU box.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1);} } while (0)

// ---- 1. latency / throughput -------------------------------------------------------------------------------
template <int ILP, int OP>
__global__ void chain_kernel(double *out, long long *cyc, int iters, double a, double b)
{
    double v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) v[i] = a + i + threadIdx.x;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 16; r++) {
#pragma unroll
            for (int i = 0; i < ILP; i++) {
                if (OP == 0) v[i] = fma(v[i], a, b);
                else if (OP == 1) v[i] = v[i] + b;
                else v[i] = v[i] * a;
            }
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int ILP, int OP>
void run_chain(const char *name, int threads, double *out, long long *cyc)
{
    const int iters = 2000;
    chain_kernel<ILP, OP><<<148, threads>>>(out, cyc, iters, 1.0000001, 1e-9);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    chain_kernel<ILP, OP><<<148, threads>>>(out, cyc, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; CK(cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost));
    const double n_inst = (double)iters * 16 * ILP;           // per warp
    const double warps_per_smsp = threads / 32.0 / 4.0;
    printf("%-5s ILP %d warps/SMSP %.2f: %.2f cycles per warp-instruction, SMSP issue interval %.2f cycles/inst (%.3f ms)\n",
           name, ILP, warps_per_smsp, c / n_inst, c / (n_inst * (warps_per_smsp < 1 ? 1 : warps_per_smsp)), ms);
}

// ---- 2. instruction cache ---------------------------------------------------------------------------------
// Body: N DFMA on 8 accumulators, fully unrolled -> N * 16 bytes of SASS.
template <int N>
__global__ void __launch_bounds__(128) icache_kernel(double *out, long long *cyc, int iters, double a, double b, int skew)
{
    double v[8];
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = a + i + threadIdx.x;
    // warps start out of phase: warp w first spins for w * skew dependent DFMAs
    const int w = threadIdx.x >> 5;
    for (int s = 0; s < w * skew; s++) v[0] = fma(v[0], a, b);
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < N / 8; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = fma(v[i], a, b);
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int N>
void run_icache(double *out, long long *cyc, int blocks_per_sm)
{
    const int iters = (1 << 22) / N;           // ~4M DFMA per warp
    for (int skew = 0; skew <= 997; skew += 997) {
        icache_kernel<N><<<148 * blocks_per_sm, 128>>>(out, cyc, iters, 1.0000001, 1e-9, skew);
        CK(cudaDeviceSynchronize());
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        icache_kernel<N><<<148 * blocks_per_sm, 128>>>(out, cyc, iters, 1.0000001, 1e-9, skew);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double n_inst = (double)iters * N;
        const double total = n_inst * 148.0 * blocks_per_sm * 4;      // warp-instructions over the chip
        // FP64 peak: 1 warp-DFMA per 2 cycles per SMSP
        printf("body %4d KB  warps/SMSP %d  skew %3d: %.3f ms, %.2f cycles per DFMA per SMSP (2.00 = pipe peak)\n",
               N * 16 / 1024, blocks_per_sm, skew, ms, ms * 1e-3 * 1.965e9 / (total / (148.0 * 4)));
    }
}

// ---- 3. co-issue ------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(128) coissue_kernel(double *out, int iters, double a, double b)
{
    __shared__ double sm[128 * 4];
    double v[8];
    int q[4] = {1, 2, 3, 4};
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = a + i + threadIdx.x;
    sm[threadIdx.x] = a; sm[threadIdx.x + 128] = b; sm[threadIdx.x + 256] = a; sm[threadIdx.x + 384] = b;
    volatile double *vs = sm;
    double acc = 0;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                v[i] = fma(v[i], a, b);
                if (MODE == 1) q[i & 3] = q[i & 3] * 3 + r;                       // one integer op per DFMA
                if (MODE == 2 && (i & 1)) acc += vs[threadIdx.x + 128 * (i >> 1)]; // LDS.64 + DADD per 2 DFMA
                if (MODE == 3) { q[i & 3] = q[i & 3] * 3 + r; q[(i + 1) & 3] ^= q[i & 3]; } // two integer ops
            }
        }
    }
    double s = acc + q[0] + q[1] + q[2] + q[3];
#pragma unroll
    for (int i = 0; i < 8; i++) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run_coissue(const char *name, double *out)
{
    const int iters = 20000;
    coissue_kernel<MODE><<<148 * 2, 128>>>(out, iters, 1.0000001, 1e-9);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    coissue_kernel<MODE><<<148 * 2, 128>>>(out, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double dfma_per_smsp = (double)iters * 64 * 2;
    printf("co-issue %-28s: %.3f ms, %.2f cycles per DFMA per SMSP\n", name, ms, ms * 1e-3 * 1.965e9 / dfma_per_smsp);
}

// ---- 4. instruction supply --------------------------------------------------------------------------------
// Straight-line body of N independent instructions (8 accumulator chains, so no dependency waits): MIX 0 = FFMA only
// (issue limit 1 per cycle), MIX 1 = DFMA and FFMA alternating (FP64 pipe half busy, as in run_kernel).  Blocks of
// 128 threads = one warp per scheduler; with a body larger than the per-scheduler instruction buffer every 128-byte
// line (8 instructions) has to come from the SM's instruction cache.
template <int N, int MIX>
__global__ void __launch_bounds__(128) supply_kernel(double *out, long long *cyc, int iters, float a, float b, int skew)
{
    float v[8];
    double d[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { v[i] = a + i + threadIdx.x; d[i] = v[i]; }
    const int w = (threadIdx.x >> 5) + 4 * (blockIdx.x & 1);
    for (int s = 0; s < w * skew; s++) v[0] = fmaf(v[0], a, b);
    const double da = a, db = b;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < N / 8; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MIX == 1 && (i & 1)) d[i] = fma(d[i], da, db);
                else v[i] = fmaf(v[i], a, b);
            }
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += v[i] + d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int N, int MIX>
void run_supply(double *out, long long *cyc, int blocks_per_sm)
{
    const int iters = (1 << 22) / N;
    for (int skew = 0; skew <= 997; skew += 997) {
        supply_kernel<N, MIX><<<148 * blocks_per_sm, 128>>>(out, cyc, iters, 1.0000001f, 1e-9f, skew);
        CK(cudaDeviceSynchronize());
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        supply_kernel<N, MIX><<<148 * blocks_per_sm, 128>>>(out, cyc, iters, 1.0000001f, 1e-9f, skew);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double per_warp = (double)iters * N;
        printf("%s body %3d KB  warps/SMSP %d  skew %3d: %.3f ms, %.2f cycles per instruction per warp, %.1f per 128-byte line\n",
               MIX ? "DFMA+FFMA" : "FFMA     ", N * 16 / 1024, blocks_per_sm, skew, ms, ms * 1e-3 * 1.965e9 / per_warp,
               8 * ms * 1e-3 * 1.965e9 / per_warp);
    }
}

int main(int argc, char **argv)
{
    double *out; long long *cyc;
    CK(cudaMalloc(&out, sizeof(double) * 148 * 8 * 1024));
    CK(cudaMalloc(&cyc, 8));
    if (argc > 1 && argv[1][0] == '4') {
        printf("== 4. instruction supply: straight-line independent instructions, 128-thread blocks ==\n");
        for (int bps = 1; bps <= 3; bps++) {
            run_supply<128, 0>(out, cyc, bps);
            run_supply<512, 0>(out, cyc, bps);
            run_supply<1536, 0>(out, cyc, bps);
            run_supply<128, 1>(out, cyc, bps);
            run_supply<512, 1>(out, cyc, bps);
            run_supply<1536, 1>(out, cyc, bps);
        }
        return 0;
    }
    printf("== 1. FP64 latency / throughput (clock64 on one warp; all SMs busy) ==\n");
    run_chain<1, 0>("DFMA", 32, out, cyc);
    run_chain<1, 1>("DADD", 32, out, cyc);
    run_chain<1, 2>("DMUL", 32, out, cyc);
    run_chain<2, 0>("DFMA", 32, out, cyc);
    run_chain<4, 0>("DFMA", 32, out, cyc);
    run_chain<8, 0>("DFMA", 32, out, cyc);
    run_chain<1, 0>("DFMA", 128, out, cyc);
    run_chain<2, 0>("DFMA", 128, out, cyc);
    run_chain<4, 0>("DFMA", 128, out, cyc);
    run_chain<8, 0>("DFMA", 128, out, cyc);
    run_chain<1, 0>("DFMA", 256, out, cyc);
    run_chain<2, 0>("DFMA", 256, out, cyc);
    run_chain<4, 0>("DFMA", 256, out, cyc);
    run_chain<1, 0>("DFMA", 384, out, cyc);
    run_chain<2, 0>("DFMA", 384, out, cyc);
    run_chain<1, 0>("DFMA", 512, out, cyc);
    run_chain<2, 0>("DFMA", 512, out, cyc);
    run_chain<1, 0>("DFMA", 1024, out, cyc);
    printf("== 2. instruction cache: straight-line DFMA body, 128-thread blocks ==\n");
    for (int bps = 2; bps <= 4; bps++) {
        run_icache<256>(out, cyc, bps);
        run_icache<512>(out, cyc, bps);
        run_icache<1024>(out, cyc, bps);
        run_icache<1536>(out, cyc, bps);
        run_icache<2048>(out, cyc, bps);
        run_icache<3072>(out, cyc, bps);
        run_icache<4096>(out, cyc, bps);
        run_icache<6144>(out, cyc, bps);
        run_icache<8192>(out, cyc, bps);
        run_icache<16384>(out, cyc, bps);
    }
    printf("== 3. co-issue with the FP64 stream (2 warps per SMSP) ==\n");
    run_coissue<0>("DFMA only", out);
    run_coissue<1>("DFMA + 1 integer op", out);
    run_coissue<3>("DFMA + 2 integer ops", out);
    run_coissue<2>("DFMA + LDS.64/DADD per 2", out);
    return 0;
}
