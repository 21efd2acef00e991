// Consensus assembly of the per-window calls of a read -- SURVEY.md section 8(f) rank 1.
//
// Replaces process_output (nanotensor/dataset.py:421-432): index2base per window (nanotensor/trim_signal.py:116-122),
// simple_assembly + add_count (chiron/chiron/utils/easy_assembler.py:11-37), np.argmax, index2base.  Every window is
// placed against its predecessor by the offset of their longest common block (the reference asks difflib for
// max(get_matching_blocks(), key=size); for windows under 200 bases that is the longest common substring, ties broken by
// the scan order of difflib's dynamic programme: end in the previous window ascending, then end in this window; from 200
// bases on difflib's "autojunk" regime applies and k_asm_disp_long restates the whole procedure), the
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

constexpr int ASM_AUTOJUNK = 200;          // difflib drops "popular" elements from the match index from 200 on (k_asm_disp_long)
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
    if (lane == 0) blen[w] = n;
    (void)status;
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
    if (lb >= ASM_AUTOJUNK) return;                             // autojunk regime: k_asm_disp_long writes disp[w]
    if (la == 0 || lb == 0) {                                   // no block: the sentinel (la, lb, 0)
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

// K2': the windows k_asm_disp left out -- 200 and more called bases in the SECOND window of a pair, where
// difflib.SequenceMatcher switches "autojunk" on (__chain_b): every symbol occurring more than lb / 100 + 1 times in b is
// POPULAR and leaves the match index (with four bases that is normally every base).  The whole procedure is restated:
// find_longest_match = longest run of matches made of non-popular symbols (difflib's scan order: end in a ascending, then
// end in b, strict improvement), extended at both ends over equal symbols, popular ones included; get_matching_blocks
// recurses left and right of every block (a stack, as difflib pops its queue from the end), sorts the blocks, merges
// adjacent ones, and the reference takes the FIRST block of maximal size (the sentinel (la, lb, 0) when there is none).
// One warp per pair: the dynamic programme runs with lanes across b, lane 0 keeps the stack and the block list.
// Shared memory per warp: a[T], b[T], two rows of T + 2 run lengths (16 bit), stack 4 x T, blocks 3 x (T + 1) (16 bit).
__host__ __device__ inline size_t asm_long_smem_per_warp(int T) { return ((size_t)20 * T + 64 + 15) & ~(size_t)15; }

__global__ void __launch_bounds__(ASM_WARPS * 32) k_asm_disp_long(const uint8_t *__restrict__ bases,
                                                                  const int32_t *__restrict__ blen,
                                                                  const int32_t *__restrict__ win_read,
                                                                  const int64_t *__restrict__ read_off, int64_t n_windows,
                                                                  int T, int32_t *__restrict__ disp) {
    extern __shared__ __align__(16) uint8_t asm_long_smem[];
    const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *base = asm_long_smem + (size_t)wi * asm_long_smem_per_warp(T);
    uint8_t *a = base, *b = a + T;
    uint16_t *row0 = reinterpret_cast<uint16_t *>(base + ((2 * (size_t)T + 15) & ~(size_t)15));
    uint16_t *row1 = row0 + T + 2;
    int16_t *stk = reinterpret_cast<int16_t *>(row1 + T + 2);          // [T + 2][4]
    int16_t *blk = stk + 4 * (T + 2);                                   // [T + 1][3]
    for (int64_t w = (int64_t)blockIdx.x * ASM_WARPS + wi; w < n_windows; w += (int64_t)gridDim.x * ASM_WARPS) {
        const int lb = blen[w];
        if (lb < ASM_AUTOJUNK || w == read_off[win_read[w]]) continue;
        const int la = blen[w - 1];
        if (la == 0) { if (lane == 0) disp[w] = la - lb; continue; }
        __syncwarp();
        for (int t = lane; t < la; t += 32) a[t] = bases[(w - 1) * T + t];
        for (int t = lane; t < lb; t += 32) b[t] = bases[w * T + t];
        __syncwarp();
        // popular symbols of b
        uint32_t popular = 0;
        for (int sym = 0; sym < 5; ++sym) {
            int c = 0;
            for (int j0 = 0; j0 < lb; j0 += 32) c += __popc(__ballot_sync(0xffffffffu, j0 + lane < lb && b[j0 + lane] == sym));
            if (c > lb / 100 + 1) popular |= 1u << sym;
        }
        int n_stk = 1, n_blk = 0;
        if (lane == 0) { stk[0] = 0; stk[1] = (int16_t)la; stk[2] = 0; stk[3] = (int16_t)lb; }
        __syncwarp();
        while (n_stk > 0) {
            --n_stk;
            const int alo = stk[4 * n_stk], ahi = stk[4 * n_stk + 1], blo = stk[4 * n_stk + 2], bhi = stk[4 * n_stk + 3];
            __syncwarp();
            uint16_t *prev = row0, *cur = row1;                  // index j + 1 = run ending at b[j]; index blo = "before blo"
            for (int j = blo + lane; j <= bhi; j += 32) prev[j] = 0;
            __syncwarp();
            int best_k = 0, best_i = 0, best_j = 0;
            for (int i = alo; i < ahi; ++i) {
                const uint8_t ai = a[i];
                const bool pop = (popular >> ai) & 1u;
                for (int j = blo + lane; j < bhi; j += 32) {
                    const int k = (!pop && b[j] == ai) ? prev[j] + 1 : 0;
                    cur[j + 1] = (uint16_t)k;
                    if (k > best_k) { best_k = k; best_i = i; best_j = j; }
                }
                if (lane == 0) cur[blo] = 0;
                __syncwarp();
                uint16_t *tmp = prev; prev = cur; cur = tmp;
            }
            unsigned long long key = ((unsigned long long)best_k << 40) | ((unsigned long long)(0xFFFFF - best_i) << 20) |
                                     (unsigned long long)(0xFFFFF - best_j);
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
                key = other > key ? other : key;
            }
            int push = 0;
            if (lane == 0) {
                int k = (int)(key >> 40), bi = alo, bj = blo;
                if (k) { bi = 0xFFFFF - (int)((key >> 20) & 0xFFFFF) - k + 1; bj = 0xFFFFF - (int)(key & 0xFFFFF) - k + 1; }
                while (bi > alo && bj > blo && a[bi - 1] == b[bj - 1]) { --bi; --bj; ++k; }
                while (bi + k < ahi && bj + k < bhi && a[bi + k] == b[bj + k]) ++k;
                if (k) {
                    blk[3 * n_blk] = (int16_t)bi; blk[3 * n_blk + 1] = (int16_t)bj; blk[3 * n_blk + 2] = (int16_t)k;
                    push |= 4;
                    if (alo < bi && blo < bj) {
                        stk[4 * n_stk] = (int16_t)alo; stk[4 * n_stk + 1] = (int16_t)bi; stk[4 * n_stk + 2] = (int16_t)blo; stk[4 * n_stk + 3] = (int16_t)bj;
                        push |= 1;
                    }
                    if (bi + k < ahi && bj + k < bhi) {
                        const int at = n_stk + (push & 1);
                        stk[4 * at] = (int16_t)(bi + k); stk[4 * at + 1] = (int16_t)ahi; stk[4 * at + 2] = (int16_t)(bj + k); stk[4 * at + 3] = (int16_t)bhi;
                        push |= 2;
                    }
                }
            }
            push = __shfl_sync(0xffffffffu, push, 0);
            n_blk += (push >> 2) & 1;
            n_stk += (push & 1) + ((push >> 1) & 1);
            __syncwarp();
        }
        if (lane == 0) {
            // blocks sorted by (i, j), adjacent ones merged, first of maximal size; the sentinel (la, lb, 0) if none
            for (int x = 1; x < n_blk; ++x) {
                const int16_t bi = blk[3 * x], bj = blk[3 * x + 1], bk = blk[3 * x + 2];
                int y = x - 1;
                while (y >= 0 && (blk[3 * y] > bi || (blk[3 * y] == bi && blk[3 * y + 1] > bj))) {
                    blk[3 * y + 3] = blk[3 * y]; blk[3 * y + 4] = blk[3 * y + 1]; blk[3 * y + 5] = blk[3 * y + 2];
                    --y;
                }
                blk[3 * y + 3] = bi; blk[3 * y + 4] = bj; blk[3 * y + 5] = bk;
            }
            int best_i = la, best_j = lb, best_k = 0;
            int i1 = 0, j1 = 0, k1 = 0;
            for (int x = 0; x <= n_blk; ++x) {
                const bool end = x == n_blk;
                const int i2 = end ? -1 : blk[3 * x], j2 = end ? -1 : blk[3 * x + 1], k2 = end ? 0 : blk[3 * x + 2];
                if (!end && i1 + k1 == i2 && j1 + k1 == j2) { k1 += k2; continue; }
                if (k1 > best_k) { best_k = k1; best_i = i1; best_j = j1; }
                i1 = i2; j1 = j2; k1 = k2;
            }
            disp[w] = best_i - best_j;
        }
        __syncwarp();
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
