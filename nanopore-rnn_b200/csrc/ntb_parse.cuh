// `.signal` text -> int16 samples on the device -- SURVEY.md section 8(f) rank 2 (read ingestion).
//
// Replaces the parse in read_signal (chiron/chiron/chiron_input.py:220-226): for every line, int(x) for x in
// line.split().  The bytes of many files are concatenated; every file is cut into fixed chunks (table built on the
// host), a CTA owns one chunk: pass 1 counts the tokens that START in the chunk, a scan turns the counts into output
// offsets, pass 2 re-finds the starts, ranks them with a block scan and parses each token (a token may run past its
// chunk, never past its file).  Byte work, HBM-bound: two bytes per sample are written; the default since the end of round 2
// is ONE pass (k_parse_fused below: the text is read once), the two-pass pair stays for A/B runs.
// Accepted tokens: [+-]?[0-9]{1,6} (what a .signal file holds); anything else int() would also take (underscores,
// other Unicode digits) or reject raises status 2; a value outside int16 raises status 3.  Output feeds
// ntb_pack_windows directly (samples + per-read offsets).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ntb {

constexpr int PARSE_CHUNK = 4096, PARSE_THREADS = 256, PARSE_PER_THREAD = PARSE_CHUNK / PARSE_THREADS;
constexpr int PARSE_STATUS_CAPACITY = 1, PARSE_STATUS_BADCHAR = 2, PARSE_STATUS_RANGE = 3;

__device__ __forceinline__ bool parse_is_ws(uint8_t c) {       // str.split() whitespace in ASCII
    return c == ' ' || (c >= 9 && c <= 13) || (c >= 28 && c <= 31);
}

// Stage the chunk [lo, hi) plus the byte before it into shared memory with 16-byte loads: sh[1 + k] = text[lo + k],
// sh[0] = the byte before the chunk (a space at the start of the file).  text is 16-byte aligned (checked by the API).
__device__ __forceinline__ void parse_stage(const uint8_t *__restrict__ text, int64_t total, int64_t lo, int64_t hi,
                                            int64_t f_lo, uint8_t *sh /* [16 + PARSE_CHUNK + 16] */) {
    const int64_t a_lo = lo & ~(int64_t)15;                    // aligned start, <= lo
    const int shift = (int)(lo - a_lo);                        // sh_raw[shift + k] = text[lo + k]
    uint8_t *raw = sh + 16;                                    // raw[j] = text[a_lo + j]; 16 bytes of headroom in front
    for (int64_t j = (int64_t)threadIdx.x * 16; a_lo + j < hi; j += (int64_t)PARSE_THREADS * 16) {
        if (a_lo + j + 16 <= total) *reinterpret_cast<uint4 *>(raw + j) = *reinterpret_cast<const uint4 *>(text + a_lo + j);
        else for (int q = 0; q < 16 && a_lo + j + q < total; ++q) raw[j + q] = text[a_lo + j + q];
    }
    if (threadIdx.x == 0) raw[shift - 1] = lo > f_lo ? text[lo - 1] : (uint8_t)' ';
    __syncthreads();
}

// token starts among this thread's bytes [k0, k0 + PARSE_PER_THREAD) of the staged chunk of n bytes; c[-1] is valid
__device__ __forceinline__ int parse_count_starts(const uint8_t *c, int k0, int n, unsigned *mask) {
    unsigned m = 0;
    uint8_t prev = c[k0 - 1];
#pragma unroll
    for (int k = 0; k < PARSE_PER_THREAD; ++k) {
        const uint8_t ch = k0 + k < n ? c[k0 + k] : (uint8_t)' ';
        if (!parse_is_ws(ch) && parse_is_ws(prev)) m |= 1u << k;
        prev = ch;
    }
    *mask = m;
    return __popc(m);
}

// chunk table: chunk c covers bytes [chunk_lo[c], chunk_lo[c] + PARSE_CHUNK) clipped to its file [file_lo, file_hi)
__global__ void __launch_bounds__(PARSE_THREADS) k_parse_count(const uint8_t *__restrict__ text,
                                                               const int64_t *__restrict__ chunk_lo,
                                                               const int32_t *__restrict__ chunk_file,
                                                               const int64_t *__restrict__ file_off, int64_t total,
                                                               int64_t *__restrict__ counts) {
    __shared__ int warp_n[PARSE_THREADS / 32];
    __shared__ __align__(16) uint8_t sh[16 + PARSE_CHUNK + 32];
    const int c = blockIdx.x, f = chunk_file[c];
    const int64_t lo = chunk_lo[c], hi = min(lo + PARSE_CHUNK, file_off[f + 1]);
    parse_stage(text, total, lo, hi, file_off[f], sh);
    unsigned mask;
    int n = parse_count_starts(sh + 16 + (int)(lo & 15), threadIdx.x * PARSE_PER_THREAD, (int)(hi - lo), &mask);
#pragma unroll
    for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0) warp_n[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int q = 0; q < PARSE_THREADS / 32; ++q) t += warp_n[q];
        counts[c] = t;
    }
}

__global__ void __launch_bounds__(PARSE_THREADS) k_parse_emit(const uint8_t *__restrict__ text,
                                                              const int64_t *__restrict__ chunk_lo,
                                                              const int32_t *__restrict__ chunk_file,
                                                              const int64_t *__restrict__ file_off, int64_t total,
                                                              const int64_t *__restrict__ chunk_out, int64_t capacity,
                                                              int16_t *__restrict__ samples, int32_t *status) {
    __shared__ int warp_n[PARSE_THREADS / 32];
    __shared__ __align__(16) uint8_t sh[16 + PARSE_CHUNK + 32];
    const int c = blockIdx.x, f = chunk_file[c], lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
    const int64_t f_hi = file_off[f + 1];
    const int64_t lo = chunk_lo[c], hi = min(lo + PARSE_CHUNK, f_hi);
    const int64_t p = lo + (int64_t)threadIdx.x * PARSE_PER_THREAD;
    parse_stage(text, total, lo, hi, file_off[f], sh);
    const uint8_t *cs = sh + 16 + (int)(lo & 15);              // cs[k] = text[lo + k] for k < hi - lo
    unsigned mask;
    const int n = parse_count_starts(cs, threadIdx.x * PARSE_PER_THREAD, (int)(hi - lo), &mask);
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
    }
    if (lane == 31) warp_n[wi] = incl;
    __syncthreads();
    int before = incl - n;
    for (int q = 0; q < wi; ++q) before += warp_n[q];
    int64_t out = chunk_out[c] + before;
    while (mask) {
        const int k = __ffs(mask) - 1;
        mask &= mask - 1;
        int64_t i = p + k;
        auto at = [&](int64_t q) -> uint8_t { return q < hi ? cs[q - lo] : (q < f_hi ? text[q] : (uint8_t)' '); };
        uint8_t ch = at(i);
        const bool neg = ch == '-';
        if (ch == '-' || ch == '+') ch = at(++i);
        int v = 0, digits = 0;
        bool bad = false;
        while (!parse_is_ws(ch)) {
            if (ch < '0' || ch > '9' || digits >= 6) { bad = true; break; }
            v = v * 10 + (ch - '0');
            ++digits;
            ch = at(++i);
        }
        if (bad || digits == 0) atomicMax(status, PARSE_STATUS_BADCHAR);
        else {
            v = neg ? -v : v;
            if (v < -32768 || v > 32767) atomicMax(status, PARSE_STATUS_RANGE);
            else if (out < capacity) samples[out] = (int16_t)v;
        }
        ++out;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// ONE pass (the default): count, output offset and emit in the same CTA, the text is read once.  The chunk offsets come from
// a decoupled look-back over per-chunk state words instead of a separate scan launch: a CTA takes its chunk from an atomic
// ticket (so every lower chunk is already running or done -- no deadlock whatever order the CTAs are scheduled in),
// publishes {flag = aggregate, its token count}, warp 0 walks back over its predecessors' words (32 at a time) adding
// aggregates until it meets an inclusive prefix, then publishes {flag = prefix, sum}.  A state word is self-contained (flag
// in the top two bits, value below), so relaxed 64-bit accesses are enough.  state[] and the ticket must be zero at launch.
constexpr unsigned long long PARSE_FLAG_AGG = 1ull << 62, PARSE_FLAG_PFX = 2ull << 62, PARSE_VAL_MASK = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long parse_ld_state(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void parse_st_state(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void __launch_bounds__(PARSE_THREADS) k_parse_fused(const uint8_t *__restrict__ text,
                                                               const int64_t *__restrict__ chunk_lo,
                                                               const int32_t *__restrict__ chunk_file,
                                                               const int64_t *__restrict__ file_off, int64_t total,
                                                               int64_t n_chunks, unsigned long long *__restrict__ state,
                                                               unsigned int *__restrict__ ticket, int64_t capacity,
                                                               int64_t *__restrict__ chunk_out, int16_t *__restrict__ samples,
                                                               int32_t *status) {
    __shared__ int warp_n[PARSE_THREADS / 32];
    __shared__ __align__(16) uint8_t sh[16 + PARSE_CHUNK + 32];
    __shared__ unsigned int c_s;
    __shared__ long long excl_s;
    const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
    if (threadIdx.x == 0) c_s = atomicAdd(ticket, 1u);
    __syncthreads();
    const int64_t c = c_s;
    if (c >= n_chunks) return;
    const int f = chunk_file[c];
    const int64_t f_hi = file_off[f + 1];
    const int64_t lo = chunk_lo[c], hi = min(lo + PARSE_CHUNK, f_hi);
    const int64_t p = lo + (int64_t)threadIdx.x * PARSE_PER_THREAD;
    parse_stage(text, total, lo, hi, file_off[f], sh);
    const uint8_t *cs = sh + 16 + (int)(lo & 15);              // cs[k] = text[lo + k] for k < hi - lo
    unsigned mask;
    const int n = parse_count_starts(cs, threadIdx.x * PARSE_PER_THREAD, (int)(hi - lo), &mask);
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
    }
    if (lane == 31) warp_n[wi] = incl;
    __syncthreads();
    int before = incl - n, block_total = 0;
    for (int q = 0; q < PARSE_THREADS / 32; ++q) {
        if (q < wi) before += warp_n[q];
        block_total += warp_n[q];
    }
    if (wi == 0) {
        // publish the aggregate, look back, publish the inclusive prefix
        if (lane == 0) parse_st_state(state + c, (c == 0 ? PARSE_FLAG_PFX : PARSE_FLAG_AGG) | (unsigned long long)block_total);
        long long excl = 0;
        int64_t j = c - 1;                                     // lane i looks at chunk j - i
        unsigned spins = 0;
        while (j >= 0) {
            const int64_t mine = j - lane;
            unsigned long long v = mine >= 0 ? parse_ld_state(state + mine) : PARSE_FLAG_PFX;   // before chunk 0: prefix 0
            const unsigned ready = __ballot_sync(0xffffffffu, (v >> 62) != 0);
            // the leading run of ready words (from lane 0 = the nearest predecessor) up to and including the first prefix
            const unsigned pfx = __ballot_sync(0xffffffffu, (v >> 62) == 2);
            const int first_unready = __ffs(~ready) - 1;       // 32 if all ready (ffs(0) = 0 -> -1 handled below)
            const int n_ready = ready == 0xffffffffu ? 32 : first_unready;
            const int first_pfx = pfx ? __ffs(pfx) - 1 : 32;
            const int take = first_pfx < n_ready ? first_pfx + 1 : n_ready;     // lanes [0, take) are added
            long long part = lane < take ? (long long)(v & PARSE_VAL_MASK) : 0;
#pragma unroll
            for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
            excl += part;
            if (first_pfx < n_ready) break;                    // met an inclusive prefix: done
            j -= take;
            if (take == 0) {
                __nanosleep(20);
                if (++spins > (1u << 24)) __trap();            // protocol bug -> launch failure, not a hung GPU
            }
        }
        if (lane == 0) {
            if (c > 0) parse_st_state(state + c, PARSE_FLAG_PFX | (unsigned long long)(excl + block_total));
            excl_s = excl;
            chunk_out[c] = excl;
            if (c == n_chunks - 1) {
                chunk_out[n_chunks] = excl + block_total;
                if (excl + block_total > capacity) atomicMax(status, PARSE_STATUS_CAPACITY);
            }
        }
    }
    __syncthreads();
    int64_t out = excl_s + before;
    while (mask) {
        const int k = __ffs(mask) - 1;
        mask &= mask - 1;
        int64_t i = p + k;
        auto at = [&](int64_t q) -> uint8_t { return q < hi ? cs[q - lo] : (q < f_hi ? text[q] : (uint8_t)' '); };
        uint8_t ch = at(i);
        const bool neg = ch == '-';
        if (ch == '-' || ch == '+') ch = at(++i);
        int v = 0, digits = 0;
        bool bad = false;
        while (!parse_is_ws(ch)) {
            if (ch < '0' || ch > '9' || digits >= 6) { bad = true; break; }
            v = v * 10 + (ch - '0');
            ++digits;
            ch = at(++i);
        }
        if (bad || digits == 0) atomicMax(status, PARSE_STATUS_BADCHAR);
        else {
            v = neg ? -v : v;
            if (v < -32768 || v > 32767) atomicMax(status, PARSE_STATUS_RANGE);
            else if (out < capacity) samples[out] = (int16_t)v;
        }
        ++out;
    }
}

// samples per file from the scanned chunk offsets: read_off[f] = chunk_out[first chunk of f]
__global__ void k_parse_read_off(const int64_t *__restrict__ chunk_out, const int64_t *__restrict__ first_chunk,
                                 int64_t n_files, int64_t *__restrict__ read_off) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f <= n_files) read_off[f] = chunk_out[first_chunk[f]];
}

}  // namespace ntb
