// Consensus assembly of the per-window calls of a read -- SURVEY.md section 8(f) rank 1.
//
// Replaces process_output (nanotensor/dataset.py:421-432): index2base per window (nanotensor/trim_signal.py:116-122),
// simple_assembly + add_count (chiron/chiron/utils/easy_assembler.py:11-37), np.argmax, index2base.  Every window is
// placed against its predecessor by the offset of their longest common block (the reference asks difflib for
// max(get_matching_blocks(), key=size); for windows under 200 bases that is the longest common substring, ties broken by
// the scan order of difflib's dynamic programme: end in the previous window ascending, then end in this window), the
// offsets accumulate along the read, every base votes for its position, and the consensus is the first-max vote.
// All integer work: results are identical to the reference's, vote matrix included.
//
// Five small kernels, HBM/latency-bound integer work (no tensor cores): compact the label rows -> pairwise longest block
// (one warp per window pair, the DP row in shared memory, lanes across the second window) -> per-read prefix sum of the
// displacements and consensus length -> votes (int32 atomics) -> arg-max to ASCII.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ntb {

constexpr int ASM_AUTOJUNK = 200;          // difflib drops "popular" elements from 200 on: not restated, rejected
constexpr int ASM_WARPS = 4;               // warps (window pairs) per CTA in the pair kernel
constexpr int ASM_STATUS_LONG_WINDOW = 2, ASM_STATUS_CAPACITY = 1;

// K1: one warp per window: bases = the entries of the label row that are not -1 (index2base), blen = their count
__global__ void __launch_bounds__(128) k_asm_compact(const int32_t *__restrict__ labels, int64_t n_windows, int T,
                                                     uint8_t *__restrict__ bases, int32_t *__restrict__ blen,
                                                     int32_t *status) {
    const int64_t w = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (w >= n_windows) return;
    const int32_t *row = labels + w * T;
    uint8_t *dst = bases + w * T;
    int n = 0;
    for (int t0 = 0; t0 < T; t0 += 32) {
        const int t = t0 + lane;
        const int32_t v = t < T ? row[t] : -1;
        const unsigned keep = __ballot_sync(0xffffffffu, v != -1);
        if (v != -1) dst[n + __popc(keep & ((1u << lane) - 1))] = (uint8_t)v;
        n += __popc(keep);
    }
    if (lane == 0) {
        blen[w] = n;
        if (n >= ASM_AUTOJUNK) atomicMax(status, ASM_STATUS_LONG_WINDOW);
    }
}

// window -> read id (one CTA per read)
__global__ void k_asm_window_reads(const int64_t *__restrict__ read_off, int32_t *__restrict__ win_read) {
    const int r = blockIdx.x;
    for (int64_t w = read_off[r] + threadIdx.x; w < read_off[r + 1]; w += blockDim.x) win_read[w] = r;
}

// K2: one warp per window: displacement against the previous window of the same read (0 for the first)
__global__ void __launch_bounds__(ASM_WARPS * 32) k_asm_disp(const uint8_t *__restrict__ bases,
                                                             const int32_t *__restrict__ blen,
                                                             const int32_t *__restrict__ win_read,
                                                             const int64_t *__restrict__ read_off, int64_t n_windows,
                                                             int T, int32_t *__restrict__ disp) {
    extern __shared__ uint8_t asm_smem[];
    const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * ASM_WARPS + wi;
    if (w >= n_windows) return;
    if (w == read_off[win_read[w]]) {
        if (lane == 0) disp[w] = 0;
        return;
    }
    const int la = blen[w - 1], lb = blen[w];
    if (la == 0 || lb == 0 || la >= ASM_AUTOJUNK || lb >= ASM_AUTOJUNK) {   // no block: the sentinel (la, lb, 0)
        if (lane == 0) disp[w] = la - lb;
        return;
    }
    // per warp: a[T], b[T], two DP rows of lb + 1 run lengths (< 200, one byte each)
    uint8_t *a = asm_smem + (size_t)wi * (4 * (size_t)T + 8), *b = a + T, *prev = b + T, *cur = prev + T + 4;
    for (int t = lane; t < la; t += 32) a[t] = bases[(w - 1) * T + t];
    for (int t = lane; t < lb; t += 32) b[t] = bases[w * T + t];
    for (int t = lane; t <= lb; t += 32) prev[t] = 0;
    __syncwarp();
    int best_k = 0, best_i = 0, best_j = 0;
    for (int i = 0; i < la; ++i) {
        const uint8_t ai = a[i];
        for (int j = lane; j < lb; j += 32) {
            const int k = b[j] == ai ? prev[j] + 1 : 0;         // prev[j] = run ending at (i-1, j-1)
            cur[j + 1] = (uint8_t)k;
            if (k > best_k) { best_k = k; best_i = i; best_j = j; }
        }
        if (lane == 0) cur[0] = 0;
        __syncwarp();
        uint8_t *t = prev; prev = cur; cur = t;
    }
    // longest; then earliest end in a; then earliest end in b
    unsigned long long key = ((unsigned long long)best_k << 40) | ((unsigned long long)(0xFFFFF - best_i) << 20) |
                             (unsigned long long)(0xFFFFF - best_j);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    if (lane == 0) {
        const int k = (int)(key >> 40), i = 0xFFFFF - (int)((key >> 20) & 0xFFFFF), j = 0xFFFFF - (int)(key & 0xFFFFF);
        disp[w] = k ? i - j : la - lb;
    }
}

// K3: one CTA per read: start[w] = running sum of the displacements, bound = sum of window lengths (capacity the votes of
// the read can need), len = max over the windows after the first of start + blen (0 for a single window, like the
// reference, whose `length` is never raised by the first window)
__global__ void __launch_bounds__(256) k_asm_scan(const int32_t *__restrict__ disp, const int32_t *__restrict__ blen,
                                                  const int64_t *__restrict__ read_off, int32_t *__restrict__ start,
                                                  int64_t *__restrict__ bound, int64_t *__restrict__ len) {
    __shared__ int warp_sum[8];
    __shared__ int carry_s;
    __shared__ long long red_b[8];
    __shared__ int red_l[8];
    const int r = blockIdx.x, lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
    const int64_t w0 = read_off[r], w1 = read_off[r + 1];
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    long long tot = 0;
    int longest = 0;
    for (int64_t base = w0; base < w1; base += 256) {
        const int64_t w = base + threadIdx.x;
        int v = w < w1 ? disp[w] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += up;
        }
        if (lane == 31) warp_sum[wi] = v;
        __syncthreads();
        int pre = carry_s;
        for (int q = 0; q < wi; ++q) pre += warp_sum[q];
        v += pre;
        if (w < w1) {
            start[w] = v;
            tot += blen[w];
            if (w > w0) longest = max(longest, v + blen[w]);
        }
        __syncthreads();
        if (threadIdx.x == 255) carry_s = v;
        __syncthreads();
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        tot += __shfl_xor_sync(0xffffffffu, tot, o);
        longest = max(longest, __shfl_xor_sync(0xffffffffu, longest, o));
    }
    if (lane == 0) { red_b[wi] = tot; red_l[wi] = longest; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int q = 1; q < 8; ++q) { tot += red_b[q]; longest = max(longest, red_l[q]); }
        bound[r] = tot;
        len[r] = longest;
    }
}

// exclusive prefix sums over the reads (one CTA): bound -> bound_off, len -> cons_off; flags a capacity overflow.
// Each thread scans SCAN_PER consecutive entries serially, the per-thread totals are scanned across the block
// (warp shuffles + one shared-memory hop), 8192 entries per trip.  Also used by the event-detection and parse rows.
constexpr int SCAN_PER = 8;
__global__ void __launch_bounds__(1024) k_asm_read_offsets(const int64_t *__restrict__ bound, const int64_t *__restrict__ len,
                                                           int64_t n_reads, int64_t capacity, int64_t *__restrict__ bound_off,
                                                           int64_t *__restrict__ cons_off, int32_t *status) {
    __shared__ long long wb[32], wl[32];
    __shared__ long long cb, cl;
    const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
    if (threadIdx.x == 0) { cb = 0; cl = 0; }
    __syncthreads();
    for (int64_t base = 0; base < n_reads; base += 1024 * SCAN_PER) {
        const int64_t r0 = base + (int64_t)threadIdx.x * SCAN_PER;
        long long vb[SCAN_PER], vl[SCAN_PER], tb = 0, tl = 0;
#pragma unroll
        for (int k = 0; k < SCAN_PER; ++k) {
            vb[k] = r0 + k < n_reads ? bound[r0 + k] : 0;
            vl[k] = r0 + k < n_reads ? len[r0 + k] : 0;
            tb += vb[k];
            tl += vl[k];
        }
        long long ib = tb, il = tl;                       // inclusive scan of the per-thread totals
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long ub = __shfl_up_sync(0xffffffffu, ib, o), ul = __shfl_up_sync(0xffffffffu, il, o);
            if (lane >= o) { ib += ub; il += ul; }
        }
        if (lane == 31) { wb[wi] = ib; wl[wi] = il; }
        __syncthreads();
        long long pb = cb + ib - tb, pl = cl + il - tl;   // exclusive prefix of this thread's first entry
        for (int q = 0; q < wi; ++q) { pb += wb[q]; pl += wl[q]; }
#pragma unroll
        for (int k = 0; k < SCAN_PER; ++k) {
            if (r0 + k < n_reads) { bound_off[r0 + k] = pb; cons_off[r0 + k] = pl; }
            pb += vb[k];
            pl += vl[k];
        }
        __syncthreads();
        if (threadIdx.x == 1023) { cb = pb; cl = pl; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        bound_off[n_reads] = cb;
        cons_off[n_reads] = cl;
        if (cb > capacity) atomicMax(status, ASM_STATUS_CAPACITY);
    }
}

// K4: one warp per window: add_count -- base i of the window votes for position start + i (bases before position 0
// are dropped; positions at or beyond the read's consensus length are cut off by the reference's final slice)
__global__ void __launch_bounds__(128) k_asm_vote(const uint8_t *__restrict__ bases, const int32_t *__restrict__ blen,
                                                  const int32_t *__restrict__ start, const int32_t *__restrict__ win_read,
                                                  const int64_t *__restrict__ bound_off, const int64_t *__restrict__ len,
                                                  int64_t n_windows, int T, int64_t capacity, int32_t *__restrict__ votes) {
    const int64_t w = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (w >= n_windows) return;
    const int r = win_read[w];
    const int64_t off = bound_off[r], n = len[r];
    const int s = start[w], lb = blen[w];
    for (int i = threadIdx.x & 31; i < lb; i += 32) {
        const int64_t p = (int64_t)s + i;
        const int base = bases[w * T + i];                     // > 4 is an IndexError in the reference: no vote
        if (p >= 0 && p < n && off + p < capacity && base < 5) atomicAdd(votes + (int64_t)base * capacity + off + p, 1);
    }
}

// K5: consensus = first-max vote per position, as ASCII (index2base of np.argmax); optionally the vote matrix itself
__global__ void __launch_bounds__(256) k_asm_consensus(const int32_t *__restrict__ votes, const int64_t *__restrict__ bound_off,
                                                       const int64_t *__restrict__ cons_off, int64_t capacity,
                                                       uint8_t *__restrict__ cons, int32_t *__restrict__ counts_out,
                                                       int64_t counts_stride) {
    const int r = blockIdx.x;
    const int64_t off = bound_off[r], dst = cons_off[r], n = cons_off[r + 1] - dst;
    for (int64_t p = threadIdx.x; p < n; p += 256) {
        if (off + p >= capacity || dst + p >= capacity) break;
        int best = 0, bv = votes[off + p];
        if (counts_out) counts_out[dst + p] = bv;
#pragma unroll
        for (int c = 1; c < 5; ++c) {
            const int v = votes[(int64_t)c * capacity + off + p];
            if (counts_out) counts_out[(int64_t)c * counts_stride + dst + p] = v;
            if (v > bv) { bv = v; best = c; }
        }
        cons[dst + p] = (uint8_t)("ACGTE"[best]);
    }
}

inline size_t asm_disp_smem(int T) { return (size_t)ASM_WARPS * (4 * (size_t)T + 8); }

}  // namespace ntb
