// MinKNOW-style event detection -- SURVEY.md section 8(f) rank 4, second half: raw current samples -> events
// (start, length, mean, stdv), the table events_to_features (ntb_events.cuh) standardises for the event model.
//
// Replaces minknow_event_detect (nanonet/nanonet/eventdetection/filters.py:299-322): compute_sum_sumsq (filters.c:57-72),
// compute_tstat for two window lengths (filters.c:130-195), short_long_peak_detector (filters.c:235-300, Detector state
// of filters.py:116-143) and _construct_events (filters.py:146-204).
//
// Bit-exact against the reference's own compiled C (oracle/_ref): the cumulative sums are the same SEQUENTIAL float64
// additions (a parallel scan would round differently), the t-statistic is the same un-fused expression, the peak
// detector is the same two-detector state machine.  The sequential stages run one WARP per read: the lanes load 32
// consecutive samples with one coalesced access and then step through them together via shuffles, every lane carrying
// the same state (lane k keeps the value of step k, so the store is coalesced too); throughput comes from thousands of
// reads in flight, the parallel stages (t-statistics, events) from one CTA per read.  Float64 IEEE operations only.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ntb {

__device__ __forceinline__ double ed_bcast(double v, int k) { return __shfl_sync(0xffffffffu, v, k); }

// K1: sum[i] = sum[i-1] + x[i], sumsq[i] = sumsq[i-1] + x[i]*x[i], in order (filters.c:57-72); one warp per read
__global__ void __launch_bounds__(128) k_ed_prefix(const double *__restrict__ raw, const int64_t *__restrict__ read_off,
                                                   int64_t n_reads, double *__restrict__ sums, double *__restrict__ sumsqs) {
    const int64_t r = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const int64_t a = read_off[r], b = read_off[r + 1];
    double s = 0.0, q = 0.0;
    for (int64_t base = a; base < b; base += 32) {
        const int64_t i = base + lane;
        const double x = i < b ? raw[i] : 0.0;
        const double xx = __dmul_rn(x, x);
        const int m = (int)min((int64_t)32, b - base);
        double my_s = 0.0, my_q = 0.0;
        for (int k = 0; k < m; ++k) {
            const double v = ed_bcast(x, k), vv = ed_bcast(xx, k);
            if (base == a && k == 0) { s = v; q = vv; }        // sum[0] = data[0]: no 0.0 + x (keeps -0.0)
            else { s = __dadd_rn(s, v); q = __dadd_rn(q, vv); }
            if (k == lane) { my_s = s; my_q = q; }
        }
        if (i < b) { sums[i] = my_s; sumsqs[i] = my_q; }
    }
}

// K2: windowed t-statistic of the two windows (filters.c:130-195, pooled = false); one CTA per read
__device__ __forceinline__ double ed_tstat_at(const double *__restrict__ sum, const double *__restrict__ sumsq, int64_t n,
                                              int64_t i, int w) {
    if (n < 2 * (int64_t)w || w < 2 || i < w || i > n - w) return 0.0;
    const double dw = (double)w;
    double sum1 = sum[i - 1], sumsq1 = sumsq[i - 1];
    if (i > w) {
        sum1 = __dsub_rn(sum1, sum[i - w - 1]);
        sumsq1 = __dsub_rn(sumsq1, sumsq[i - w - 1]);
    }
    const double sum2 = __dsub_rn(sum[i + w - 1], sum[i - 1]), sumsq2 = __dsub_rn(sumsq[i + w - 1], sumsq[i - 1]);
    const double mean1 = __ddiv_rn(sum1, dw), mean2 = __ddiv_rn(sum2, dw);
    double var1 = __dsub_rn(__ddiv_rn(sumsq1, dw), __dmul_rn(mean1, mean1));
    double var2 = __dsub_rn(__ddiv_rn(sumsq2, dw), __dmul_rn(mean2, mean2));
    var1 = fmax(var1, 1e-100);
    var2 = fmax(var2, 1e-100);
    const double delta = __dsub_rn(mean2, mean1);
    const double totvar = __dadd_rn(__ddiv_rn(var1, dw), __ddiv_rn(var2, dw));
    return fabs(__ddiv_rn(delta, __dsqrt_rn(totvar)));
}

__global__ void __launch_bounds__(256) k_ed_tstat(const double *__restrict__ sums, const double *__restrict__ sumsqs,
                                                  const int64_t *__restrict__ read_off, int w_short, int w_long,
                                                  double *__restrict__ t_short, double *__restrict__ t_long) {
    const int64_t a = read_off[blockIdx.x], n = read_off[blockIdx.x + 1] - a;
    for (int64_t i = threadIdx.x + (int64_t)blockIdx.y * 256; i < n; i += 256 * (int64_t)gridDim.y) {
        t_short[a + i] = ed_tstat_at(sums + a, sumsqs + a, n, i, w_short);
        t_long[a + i] = ed_tstat_at(sums + a, sumsqs + a, n, i, w_long);
    }
}

// K3: the two-detector peak state machine (filters.c:235-300); one warp per read, all lanes in lock step
struct EdDet {
    double threshold, peak_value;
    long long masked_to;
    int window, peak_pos;
    bool valid;
};

__global__ void __launch_bounds__(128) k_ed_peaks(const double *__restrict__ t_short, const double *__restrict__ t_long,
                                                  const int64_t *__restrict__ read_off, int64_t n_reads, int w_short,
                                                  int w_long, double th_short, double th_long, double peak_height,
                                                  int32_t *__restrict__ peaks, int64_t *__restrict__ n_events) {
    const int64_t r = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const int64_t a = read_off[r], b = read_off[r + 1];
    EdDet d[2];
    d[0] = EdDet{th_short, 1e100, 0, w_short, -1, false};
    d[1] = EdDet{th_long, 1e100, 0, w_long, -1, false};
    int32_t *out = peaks + a;
    int count = 0;
    for (int64_t base = a; base < b; base += 32) {
        const double xs = base + lane < b ? t_short[base + lane] : 0.0, xl = base + lane < b ? t_long[base + lane] : 0.0;
        const int m = (int)min((int64_t)32, b - base);
        for (int k = 0; k < m; ++k) {
            const long long i = base - a + k;
            const double cur2[2] = {ed_bcast(xs, k), ed_bcast(xl, k)};
#pragma unroll
            for (int z = 0; z < 2; ++z) {
                EdDet &t = d[z];
                if (t.masked_to >= i) continue;
                const double cur = cur2[z];
                if (t.peak_pos == -1) {
                    if (cur < t.peak_value) t.peak_value = cur;
                    else if (__dsub_rn(cur, t.peak_value) > peak_height) { t.peak_value = cur; t.peak_pos = (int)i; }
                } else {
                    if (cur > t.peak_value) { t.peak_value = cur; t.peak_pos = (int)i; }
                    if (z == 0 && t.peak_value > t.threshold) {
                        d[1].masked_to = (long long)t.peak_pos + t.window;
                        d[1].peak_pos = -1;
                        d[1].peak_value = 1e100;
                        d[1].valid = false;
                    }
                    if (__dsub_rn(t.peak_value, cur) > peak_height && t.peak_value > t.threshold) t.valid = true;
                    if (t.valid && (i - t.peak_pos) > t.window / 2) {
                        if (lane == 0) out[count] = t.peak_pos;
                        ++count;
                        t.peak_pos = -1;
                        t.peak_value = cur;
                        t.valid = false;
                    }
                }
            }
        }
    }
    if (lane == 0) n_events[r] = b > a ? count + 1 : 0;      // _construct_events: the edges plus the end of the read
}

// ------------------------------------------------------------------------------------------------------------------
// LANE-PER-READ forms of the two sequential stages.  One warp-instruction then advances R reads instead of one, which
// is what the sequential float64 chains need: the warp-per-read forms above spend ~60 issue slots per sample on a
// single read.  A warp owns R consecutive reads (R = 4, 8, 16 or 32: the host picks the largest R that still leaves
// enough warps to fill the schedulers); for every chunk of 32 samples it loads each read's chunk with one coalesced
// access of the whole warp into a padded shared-memory tile, and lanes 0 .. R-1 then walk their own read's row.
constexpr int ED_TILE = 33;     // row pitch in doubles: lane r reading column k hits bank (33 r + k) % 32, conflict-free

template <int R>
__global__ void __launch_bounds__(64) k_ed_prefix_t(const double *__restrict__ raw, const int64_t *__restrict__ read_off,
                                                     int64_t n_reads, double *__restrict__ sums, double *__restrict__ sumsqs) {
    __shared__ double tile[2][2][R * ED_TILE];
    const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t r0 = ((int64_t)blockIdx.x * 2 + wi) * R;
    if (r0 >= n_reads) return;
    double *tx = tile[wi][0], *tq = tile[wi][1];
    const int64_t my_r = min(r0 + (lane < R ? lane : R - 1), n_reads - 1);
    const int64_t my_a = read_off[my_r], my_n = (lane < R && r0 + lane < n_reads) ? read_off[my_r + 1] - my_a : 0;
    int64_t longest = my_n;
#pragma unroll
    for (int o = 16; o; o >>= 1) longest = max(longest, __shfl_xor_sync(0xffffffffu, longest, o));
    double s = 0.0, q = 0.0;
    for (int64_t base = 0; base < longest; base += 32) {
        for (int rr = 0; rr < R; ++rr) {                       // coalesced loads, one read at a time
            const int64_t a = __shfl_sync(0xffffffffu, my_a, rr), n = __shfl_sync(0xffffffffu, my_n, rr);
            if (base + lane < n) tx[rr * ED_TILE + lane] = raw[a + base + lane];
        }
        __syncwarp();
        const int m = (int)max((int64_t)0, min((int64_t)32, my_n - base));
        for (int k = 0; k < m; ++k) {
            const double v = tx[lane * ED_TILE + k], vv = __dmul_rn(v, v);
            if (base == 0 && k == 0) { s = v; q = vv; }
            else { s = __dadd_rn(s, v); q = __dadd_rn(q, vv); }
            tx[lane * ED_TILE + k] = s;
            tq[lane * ED_TILE + k] = q;
        }
        __syncwarp();
        for (int rr = 0; rr < R; ++rr) {                       // coalesced stores
            const int64_t a = __shfl_sync(0xffffffffu, my_a, rr), n = __shfl_sync(0xffffffffu, my_n, rr);
            if (base + lane < n) {
                sums[a + base + lane] = tx[rr * ED_TILE + lane];
                sumsqs[a + base + lane] = tq[rr * ED_TILE + lane];
            }
        }
        __syncwarp();
    }
}

template <int R>
__global__ void __launch_bounds__(64) k_ed_peaks_t(const double *__restrict__ t_short, const double *__restrict__ t_long,
                                                    const int64_t *__restrict__ read_off, int64_t n_reads, int w_short,
                                                    int w_long, double th_short, double th_long, double peak_height,
                                                    int32_t *__restrict__ peaks, int64_t *__restrict__ n_events) {
    __shared__ double tile[2][2][R * ED_TILE];
    const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t r0 = ((int64_t)blockIdx.x * 2 + wi) * R;
    if (r0 >= n_reads) return;
    double *ts = tile[wi][0], *tl = tile[wi][1];
    const int64_t my_r = min(r0 + (lane < R ? lane : R - 1), n_reads - 1);
    const int64_t my_a = read_off[my_r], my_n = (lane < R && r0 + lane < n_reads) ? read_off[my_r + 1] - my_a : 0;
    int64_t longest = my_n;
#pragma unroll
    for (int o = 16; o; o >>= 1) longest = max(longest, __shfl_xor_sync(0xffffffffu, longest, o));
    EdDet d[2];
    d[0] = EdDet{th_short, 1e100, 0, w_short, -1, false};
    d[1] = EdDet{th_long, 1e100, 0, w_long, -1, false};
    int32_t *out = peaks + my_a;
    int count = 0;
    for (int64_t base = 0; base < longest; base += 32) {
        for (int rr = 0; rr < R; ++rr) {
            const int64_t a = __shfl_sync(0xffffffffu, my_a, rr), n = __shfl_sync(0xffffffffu, my_n, rr);
            if (base + lane < n) {
                ts[rr * ED_TILE + lane] = t_short[a + base + lane];
                tl[rr * ED_TILE + lane] = t_long[a + base + lane];
            }
        }
        __syncwarp();
        const int m = (int)max((int64_t)0, min((int64_t)32, my_n - base));
        for (int k = 0; k < m; ++k) {
            const long long i = base + k;
            const double cur2[2] = {ts[lane * ED_TILE + k], tl[lane * ED_TILE + k]};
#pragma unroll
            for (int z = 0; z < 2; ++z) {
                EdDet &t = d[z];
                if (t.masked_to >= i) continue;
                const double cur = cur2[z];
                if (t.peak_pos == -1) {
                    if (cur < t.peak_value) t.peak_value = cur;
                    else if (__dsub_rn(cur, t.peak_value) > peak_height) { t.peak_value = cur; t.peak_pos = (int)i; }
                } else {
                    if (cur > t.peak_value) { t.peak_value = cur; t.peak_pos = (int)i; }
                    if (z == 0 && t.peak_value > t.threshold) {
                        d[1].masked_to = (long long)t.peak_pos + t.window;
                        d[1].peak_pos = -1;
                        d[1].peak_value = 1e100;
                        d[1].valid = false;
                    }
                    if (__dsub_rn(t.peak_value, cur) > peak_height && t.peak_value > t.threshold) t.valid = true;
                    if (t.valid && (i - t.peak_pos) > t.window / 2) {
                        out[count++] = t.peak_pos;
                        t.peak_pos = -1;
                        t.peak_value = cur;
                        t.valid = false;
                    }
                }
            }
        }
        __syncwarp();
    }
    if (lane < R && r0 + lane < n_reads) n_events[r0 + lane] = my_n > 0 ? count + 1 : 0;
}

// ------------------------------------------------------------------------------------------------------------------
// FUSED form of K1 + K2 + K3 (the default from 4000 reads and windows <= 31): cumulative sums, both t-statistics and the
// peak state machine in ONE kernel, R reads per warp.  The t-statistics never reach memory (2 x 8 bytes per sample
// written and read back before), the cumulative sums are written once, interleaved as (sum, sumsq) pairs so that the
// event kernel's two look-ups per edge share a sector, and are not read again by a t-statistic launch.
// Per chunk of 32 samples: the warp loads each read's chunk (coalesced) -> lanes 0..R-1 extend their read's cumulative
// sums in order into a ring of the last 96 positions (shared memory) -> all 32 lanes evaluate the t-statistics of the 32
// positions that lag the chunk by L = w_long samples (their look-ahead window is complete by then), one position per lane
// and read -> lanes 0..R-1 walk those positions through the two-detector state machine.  Same operations in the same order
// as the separate kernels: bit-identical events.
constexpr int ED_RING = 96, ED_RINGP = 97;      // ring positions per read / row pitch (conflict-free for R <= 8 lanes)

__device__ __forceinline__ double ed_tstat_ring(const double2 *__restrict__ ring, int64_t n, int64_t i, int w) {
    if (n < 2 * (int64_t)w || w < 2 || i < w || i > n - w) return 0.0;
    const double dw = (double)w;
    const double2 m1 = ring[(i - 1) % ED_RING], p2 = ring[(i + w - 1) % ED_RING];
    double sum1 = m1.x, sumsq1 = m1.y;
    if (i > w) {
        const double2 m0 = ring[(i - w - 1) % ED_RING];
        sum1 = __dsub_rn(sum1, m0.x);
        sumsq1 = __dsub_rn(sumsq1, m0.y);
    }
    const double sum2 = __dsub_rn(p2.x, m1.x), sumsq2 = __dsub_rn(p2.y, m1.y);
    const double mean1 = __ddiv_rn(sum1, dw), mean2 = __ddiv_rn(sum2, dw);
    double var1 = __dsub_rn(__ddiv_rn(sumsq1, dw), __dmul_rn(mean1, mean1));
    double var2 = __dsub_rn(__ddiv_rn(sumsq2, dw), __dmul_rn(mean2, mean2));
    var1 = fmax(var1, 1e-100);
    var2 = fmax(var2, 1e-100);
    const double delta = __dsub_rn(mean2, mean1);
    const double totvar = __dadd_rn(__ddiv_rn(var1, dw), __ddiv_rn(var2, dw));
    return fabs(__ddiv_rn(delta, __dsqrt_rn(totvar)));
}

template <int R>
__global__ void __launch_bounds__(64) k_ed_fused(const double *__restrict__ raw, const int64_t *__restrict__ read_off,
                                                 int64_t n_reads, int w_short, int w_long, double th_short, double th_long,
                                                 double peak_height, double2 *__restrict__ sq, int32_t *__restrict__ peaks,
                                                 int64_t *__restrict__ n_events) {
    __shared__ double2 ring_s[2][R * ED_RINGP];
    __shared__ double tile[2][3][R * ED_TILE];                  // raw chunk, t_short, t_long of the positions being walked
    const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t r0 = ((int64_t)blockIdx.x * 2 + wi) * R;
    if (r0 >= n_reads) return;
    double2 *ring = ring_s[wi];
    double *tx = tile[wi][0], *ts = tile[wi][1], *tl = tile[wi][2];
    const int64_t my_r = min(r0 + (lane < R ? lane : R - 1), n_reads - 1);
    const int64_t my_a = read_off[my_r], my_n = (lane < R && r0 + lane < n_reads) ? read_off[my_r + 1] - my_a : 0;
    int64_t longest = my_n;
#pragma unroll
    for (int o = 16; o; o >>= 1) longest = max(longest, __shfl_xor_sync(0xffffffffu, longest, o));
    const int L = w_long;                                        // the walk lags the cumulative sums by L positions
    double s = 0.0, q = 0.0;
    EdDet d[2];
    d[0] = EdDet{th_short, 1e100, 0, w_short, -1, false};
    d[1] = EdDet{th_long, 1e100, 0, w_long, -1, false};
    int32_t *out = peaks + my_a;
    int count = 0;
    for (int64_t base = 0; base < longest + L; base += 32) {
        const int slot0 = (int)((base / 32) % 3) * 32;           // ring positions of this chunk: base + k -> (base + k) % 96
        for (int rr = 0; rr < R; ++rr) {                          // coalesced loads, one read at a time
            const int64_t a = __shfl_sync(0xffffffffu, my_a, rr), n = __shfl_sync(0xffffffffu, my_n, rr);
            if (base + lane < n) tx[rr * ED_TILE + lane] = raw[a + base + lane];
        }
        __syncwarp();
        const int m = (int)max((int64_t)0, min((int64_t)32, my_n - base));
        for (int k = 0; k < m; ++k) {                            // (lanes >= R and finished reads: m = 0)
            const double v = tx[lane * ED_TILE + k], vv = __dmul_rn(v, v);
            if (base == 0 && k == 0) { s = v; q = vv; }          // sum[0] = data[0]: no 0.0 + x (keeps -0.0)
            else { s = __dadd_rn(s, v); q = __dadd_rn(q, vv); }
            ring[lane * ED_RINGP + slot0 + k] = make_double2(s, q);
        }
        __syncwarp();
        for (int rr = 0; rr < R; ++rr) {
            const int64_t a = __shfl_sync(0xffffffffu, my_a, rr), n = __shfl_sync(0xffffffffu, my_n, rr);
            if (base + lane < n) sq[a + base + lane] = ring[rr * ED_RINGP + slot0 + lane];      // coalesced, written once
            const int64_t i = base - L + lane;                   // the position this lane evaluates for read rr
            if (i >= 0 && i < n) {
                ts[rr * ED_TILE + lane] = ed_tstat_ring(ring + rr * ED_RINGP, n, i, w_short);
                tl[rr * ED_TILE + lane] = ed_tstat_ring(ring + rr * ED_RINGP, n, i, w_long);
            }
        }
        __syncwarp();
        for (int k = 0; k < 32; ++k) {
            const long long i = base - L + k;
            if (i < 0 || i >= my_n) continue;
            const double cur2[2] = {ts[lane * ED_TILE + k], tl[lane * ED_TILE + k]};
#pragma unroll
            for (int z = 0; z < 2; ++z) {
                EdDet &t = d[z];
                if (t.masked_to >= i) continue;
                const double cur = cur2[z];
                if (t.peak_pos == -1) {
                    if (cur < t.peak_value) t.peak_value = cur;
                    else if (__dsub_rn(cur, t.peak_value) > peak_height) { t.peak_value = cur; t.peak_pos = (int)i; }
                } else {
                    if (cur > t.peak_value) { t.peak_value = cur; t.peak_pos = (int)i; }
                    if (z == 0 && t.peak_value > t.threshold) {
                        d[1].masked_to = (long long)t.peak_pos + t.window;
                        d[1].peak_pos = -1;
                        d[1].peak_value = 1e100;
                        d[1].valid = false;
                    }
                    if (__dsub_rn(t.peak_value, cur) > peak_height && t.peak_value > t.threshold) t.valid = true;
                    if (t.valid && (i - t.peak_pos) > t.window / 2) {
                        out[count++] = t.peak_pos;
                        t.peak_pos = -1;
                        t.peak_value = cur;
                        t.valid = false;
                    }
                }
            }
        }
        __syncwarp();
    }
    if (lane < R && r0 + lane < n_reads) n_events[r0 + lane] = my_n > 0 ? count + 1 : 0;
}

// K4 for the fused form: the cumulative sums are (sum, sumsq) pairs
__global__ void __launch_bounds__(256) k_ed_events_sq(const double2 *__restrict__ sq, const int32_t *__restrict__ peaks,
                                                      const int64_t *__restrict__ read_off,
                                                      const int64_t *__restrict__ event_off, double sample_rate,
                                                      int64_t capacity, double *__restrict__ events) {
    const int64_t a = read_off[blockIdx.x], n = read_off[blockIdx.x + 1] - a;
    const int64_t e0 = event_off[blockIdx.x], ne = event_off[blockIdx.x + 1] - e0;
    const double2 *c = sq + a;
    const int32_t *pk = peaks + a;
    for (int64_t k = threadIdx.x; k < ne; k += 256) {
        if (e0 + k >= capacity) break;
        const int64_t e = k < ne - 1 ? pk[k] : n, s = k ? pk[k - 1] : 0;
        const double ln = (double)(e - s);
        const double2 ce = c[e - 1];
        const double2 cs = k ? c[s - 1] : make_double2(0.0, 0.0);
        const double mean = __ddiv_rn(k ? __dsub_rn(ce.x, cs.x) : ce.x, ln);
        const double v = __dsub_rn(__ddiv_rn(k ? __dsub_rn(ce.y, cs.y) : ce.y, ln), __dmul_rn(mean, mean));
        double *ev = events + (e0 + k) * 4;
        ev[0] = __ddiv_rn((double)s, sample_rate);
        ev[1] = __ddiv_rn(ln, sample_rate);
        ev[2] = mean;
        ev[3] = __dsqrt_rn(v > 0.0 ? v : 0.0);
    }
}

// K4: events from the edges (filters.py:180-204); one CTA per read; events [*, 4] = start, length, mean, stdv
__global__ void __launch_bounds__(256) k_ed_events(const double *__restrict__ sums, const double *__restrict__ sumsqs,
                                                   const int32_t *__restrict__ peaks, const int64_t *__restrict__ read_off,
                                                   const int64_t *__restrict__ event_off, double sample_rate,
                                                   int64_t capacity, double *__restrict__ events) {
    const int64_t a = read_off[blockIdx.x], n = read_off[blockIdx.x + 1] - a;
    const int64_t e0 = event_off[blockIdx.x], ne = event_off[blockIdx.x + 1] - e0;
    const double *sum = sums + a, *sumsq = sumsqs + a;
    const int32_t *pk = peaks + a;
    for (int64_t k = threadIdx.x; k < ne; k += 256) {
        if (e0 + k >= capacity) break;
        const int64_t e = k < ne - 1 ? pk[k] : n, s = k ? pk[k - 1] : 0;
        const double ln = (double)(e - s);
        const double sm = k ? sum[s - 1] : 0.0, smsq = k ? sumsq[s - 1] : 0.0;
        const double mean = __ddiv_rn(k ? __dsub_rn(sum[e - 1], sm) : sum[e - 1], ln);
        const double v = __dsub_rn(__ddiv_rn(k ? __dsub_rn(sumsq[e - 1], smsq) : sumsq[e - 1], ln), __dmul_rn(mean, mean));
        double *ev = events + (e0 + k) * 4;
        ev[0] = __ddiv_rn((double)s, sample_rate);
        ev[1] = __ddiv_rn(ln, sample_rate);
        ev[2] = mean;
        ev[3] = __dsqrt_rn(v > 0.0 ? v : 0.0);
    }
}

}  // namespace ntb
