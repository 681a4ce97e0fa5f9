// Post-hoc statistics of BG traces on the device: the three formulas of the
// reference's simglucose/analysis/report.py
//   percent_stats      :74-92    time in range percentages
//   risk_index_trace   :95-133   LBGI / HBGI / RI from the chunk-mean of f(BG), 60 ROWS per chunk
//   CVGA_analysis      :198-217  2.5 / 97.5 percentiles (numpy linear interpolation), clipped to [50, 400], zones
// One thread per episode, trace layout [row][episode] (row stride given), so every pass over the rows is a
// coalesced stream.  Percentiles are exact order statistics found by a bitwise radix select on the
// order-preserving integer image of the doubles (64 counting passes serve both percentiles).
#include "../../include/simglucose_b200.h"

#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

namespace {

__device__ __forceinline__ uint64_t ordered_key(double v)
{
    const uint64_t u = (uint64_t)__double_as_longlong(v);
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}

__device__ __forceinline__ double key_value(uint64_t k)
{
    const uint64_t u = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)u);
}

// numpy's _lerp (lib/_function_base_impl.py): a + (b - a) * t, computed from the far end when t >= 0.5
__device__ __forceinline__ double np_lerp(double a, double b, double t)
{
    const double d = b - a;
    double r = a + d * t;
    if (t >= 0.5) r = b - d * (1.0 - t);
    if (d == 0.0) r = a;
    return r;
}

struct ReportArgs {
    const double *bg;
    int64_t row_stride, n_rows, n;
    int32_t chunk_rows, n_chunks;
    double *percent;     // [5][n]: 70<=BG<=180, BG>180, BG<70, BG>250, BG<50 (percent of rows)
    double *ri;          // [3][n_chunks][n]: LBGI, HBGI, RI per chunk
    double *ri_mean;     // [3][n]: mean over chunks
    double *cvga;        // [2][n]: clipped 2.5 / 97.5 percentiles
};

__global__ void __launch_bounds__(128) report_kernel(const ReportArgs a)
{
    const int64_t e = (int64_t)blockIdx.x * 128 + threadIdx.x;
    if (e >= a.n) return;
    const double *col = a.bg + e;
    const double rows = (double)a.n_rows;

    if (a.percent) {
        int64_t hyper = 0, hypo = 0, normal = 0, v250 = 0, v50 = 0;
        for (int64_t r = 0; r < a.n_rows; r++) {
            const double b = col[r * a.row_stride];
            hyper += b > 180.0; hypo += b < 70.0; normal += (b >= 70.0) && (b <= 180.0);
            v250 += b > 250.0; v50 += b < 50.0;
        }
        a.percent[0 * a.n + e] = (double)normal / rows * 100.0;
        a.percent[1 * a.n + e] = (double)hyper / rows * 100.0;
        a.percent[2 * a.n + e] = (double)hypo / rows * 100.0;
        a.percent[3 * a.n + e] = (double)v250 / rows * 100.0;
        a.percent[4 * a.n + e] = (double)v50 / rows * 100.0;
    }

    if (a.ri) {
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
        double ml = 0.0, mh = 0.0, mr = 0.0;
        int valid = 0;                                           // pandas' mean skips NaN chunks (report.py:127)
        for (int c = 0; c < a.n_chunks; c++) {
            const int64_t r0 = (int64_t)c * a.chunk_rows;
            const int64_t r1 = (r0 + a.chunk_rows < a.n_rows) ? r0 + a.chunk_rows : a.n_rows;
            double sum = 0.0;
            int cnt = 0;
            for (int64_t r = r0; r < r1; r++) {
                const double b = col[r * a.row_stride];
                if (b > 0.0) {                                   // BG[BG > 0]; NaN results are skipped by the mean
                    const double f = 1.509 * (pow(log(b), 1.084) - 5.381);
                    if (f == f) { sum += f; cnt++; }
                }
            }
            const double f = cnt ? sum / (double)cnt : nan;
            const double fl = f * ((f < 0.0) ? 1.0 : 0.0), fh = f * ((f > 0.0) ? 1.0 : 0.0);
            const double lb = 10.0 * (fl * fl), hb = 10.0 * (fh * fh);
            a.ri[((int64_t)0 * a.n_chunks + c) * a.n + e] = lb;
            a.ri[((int64_t)1 * a.n_chunks + c) * a.n + e] = hb;
            a.ri[((int64_t)2 * a.n_chunks + c) * a.n + e] = lb + hb;
            if (cnt) { ml += lb; mh += hb; mr += lb + hb; valid++; }
        }
        if (a.ri_mean) {
            a.ri_mean[0 * a.n + e] = valid ? ml / (double)valid : nan;
            a.ri_mean[1 * a.n + e] = valid ? mh / (double)valid : nan;
            a.ri_mean[2 * a.n + e] = valid ? mr / (double)valid : nan;
        }
    }

    if (a.cvga) {
        // order statistics k and k+1 for q = 2.5 % and 97.5 %: virtual index (n-1) * q / 100
        const double pos_lo = (rows - 1.0) * (2.5 / 100.0), pos_hi = (rows - 1.0) * (97.5 / 100.0);
        const int64_t k_lo = (int64_t)floor(pos_lo), k_hi = (int64_t)floor(pos_hi);
        uint64_t pre_lo = 0, pre_hi = 0;            // prefixes found so far (bits above `bit`)
        int64_t rank_lo = k_lo, rank_hi = k_hi;     // rank inside the current prefix bucket
        for (int bit = 63; bit >= 0; bit--) {
            const uint64_t mask_hi = (bit == 63) ? 0ull : (~0ull << (bit + 1));
            int64_t zeros_lo = 0, zeros_hi = 0;
            for (int64_t r = 0; r < a.n_rows; r++) {
                const uint64_t k = ordered_key(col[r * a.row_stride]);
                const bool zero = ((k >> bit) & 1ull) == 0ull;
                zeros_lo += ((k & mask_hi) == pre_lo) && zero;
                zeros_hi += ((k & mask_hi) == pre_hi) && zero;
            }
            if (rank_lo >= zeros_lo) { rank_lo -= zeros_lo; pre_lo |= 1ull << bit; }
            if (rank_hi >= zeros_hi) { rank_hi -= zeros_hi; pre_hi |= 1ull << bit; }
        }
        // successors: the next order statistic is the same value if it is repeated beyond rank k, else the
        // smallest larger value
        int64_t le_lo = 0, le_hi = 0;
        uint64_t nx_lo = ~0ull, nx_hi = ~0ull;
        for (int64_t r = 0; r < a.n_rows; r++) {
            const uint64_t k = ordered_key(col[r * a.row_stride]);
            le_lo += k <= pre_lo; le_hi += k <= pre_hi;
            if (k > pre_lo && k < nx_lo) nx_lo = k;
            if (k > pre_hi && k < nx_hi) nx_hi = k;
        }
        const double a_lo = key_value(pre_lo), a_hi = key_value(pre_hi);
        const double b_lo = (le_lo > k_lo + 1 || nx_lo == ~0ull) ? a_lo : key_value(nx_lo);
        const double b_hi = (le_hi > k_hi + 1 || nx_hi == ~0ull) ? a_hi : key_value(nx_hi);
        double mn = np_lerp(a_lo, b_lo, pos_lo - (double)k_lo);
        double mx = np_lerp(a_hi, b_hi, pos_hi - (double)k_hi);
        mn = mn < 50.0 ? 50.0 : (mn > 400.0 ? 400.0 : mn);       // report.py:201-204
        mx = mx < 50.0 ? 50.0 : (mx > 400.0 ? 400.0 : mx);
        a.cvga[0 * a.n + e] = mn;
        a.cvga[1 * a.n + e] = mx;
    }
}

// CVGA zone counts (report.py:206-216): out[0..4] = number of episodes in A, A-or-B (the reference
// subtracts A afterwards), C, D, E.  One block, integer counters, warp shuffles then shared memory.
__global__ void __launch_bounds__(256) cvga_zone_kernel(const double *cvga, int64_t n, double *out)
{
    __shared__ unsigned long long part[8][5];
    unsigned long long c[5] = {0, 0, 0, 0, 0};
    for (int64_t e = threadIdx.x; e < n; e += 256) {
        const double mn = cvga[e], mx = cvga[n + e];
        c[0] += (mn > 90.0) && (mn <= 110.0) && (mx >= 110.0) && (mx < 180.0);
        c[1] += (mn > 70.0) && (mn <= 110.0) && (mx >= 110.0) && (mx < 300.0);
        c[2] += ((mn > 90.0) && (mn <= 110.0) && (mx >= 300.0)) || ((mn <= 70.0) && (mx >= 110.0) && (mx < 180.0));
        c[3] += ((mn > 70.0) && (mn <= 90.0) && (mx >= 300.0)) || ((mn <= 70.0) && (mx >= 180.0) && (mx < 300.0));
        c[4] += (mn <= 70.0) && (mx >= 300.0);
    }
#pragma unroll
    for (int i = 0; i < 5; i++) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) c[i] += __shfl_xor_sync(0xffffffffu, c[i], off);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int i = 0; i < 5; i++) part[threadIdx.x >> 5][i] = c[i];
    }
    __syncthreads();
    if (threadIdx.x < 5) {
        unsigned long long s = 0;
        for (int w = 0; w < 8; w++) s += part[w][threadIdx.x];
        out[threadIdx.x] = (double)s;
    }
}

}  // namespace

extern "C" {

int sg_report_stats(const double *bg, int64_t row_stride, int64_t n_rows, int64_t n_episodes, int32_t chunk_rows,
                    double *percent_out, double *ri_out, double *ri_mean_out, double *cvga_out, void *stream)
{
    if (!bg || n_rows <= 0 || n_episodes <= 0 || row_stride < n_episodes) return SG_ERR_ARG;
    if ((ri_out || ri_mean_out) && chunk_rows <= 0) return SG_ERR_ARG;
    if (ri_mean_out && !ri_out) return SG_ERR_ARG;
    ReportArgs a{};
    a.bg = bg; a.row_stride = row_stride; a.n_rows = n_rows; a.n = n_episodes;
    a.chunk_rows = chunk_rows > 0 ? chunk_rows : 1;
    a.n_chunks = (int32_t)((n_rows + a.chunk_rows - 1) / a.chunk_rows);
    a.percent = percent_out; a.ri = ri_out; a.ri_mean = ri_mean_out; a.cvga = cvga_out;
    report_kernel<<<(unsigned)((n_episodes + 127) / 128), 128, 0, (cudaStream_t)stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_cvga_zones(const double *cvga, int64_t n_episodes, double *counts_out, void *stream)
{
    if (!cvga || !counts_out || n_episodes <= 0) return SG_ERR_ARG;
    cvga_zone_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(cvga, n_episodes, counts_out);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

}  // extern "C"
