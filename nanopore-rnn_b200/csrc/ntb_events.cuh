// Event-feature front end of the event (CrossEntropy) BLSTM model -- SURVEY.md section 8(f) rank 4.
//
// Replaces nanonet/nanonet/features.py::events_to_features (features.py:61-75) with
// SquiggleFeatureGenerator.__init__ (features.py:266-293, sloika_model=False), scale_array (features.py:32-58) and
// padded_offset_array (features.py:14-29): per read, the event fields mean / stdv / length are standardised over the
// read, delta = first difference of the standardised mean divided by its population std, and every event gets
// [mean, stdv, length, delta] of its left neighbour, itself and its right neighbour (zeros past the read ends).
//
// The result is BIT-IDENTICAL to the reference (float64 output), not merely close: every float64 operation is the
// IEEE round-to-nearest one NumPy performs (no FMA contraction: explicit __dadd_rn / __dmul_rn / __ddiv_rn), and the
// reductions follow NumPy's pairwise-summation tree (loops_utils.h.src @TYPE@_pairwise_sum: blocks of <= 128 values,
// eight interleaved running sums folded ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), halves split at n/2 rounded down to a
// multiple of 8).  The tree maps onto the GPU naturally: a leaf is eight independent running sums = eight lanes, a
// warp takes four leaves at a time, the fold is three xor-shuffles, and one lane per field adds the leaf sums in the
// tree's post-order.
//
// One CTA per read.  Two kernels: k_event_features_smem keeps the field being standardised in shared memory and updates
// it in place (reads up to ~27.9k events; the only HBM traffic is 24 B in and 12 x sizeof(out) B out per event);
// k_event_features is the any-length path that re-derives each reduction's operand from the raw fields in global memory.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ntb {

constexpr int EV_THREADS = 256;
constexpr int EV_PW_BLOCK = 128;          // numpy PW_BLOCKSIZE
constexpr int EV_MIN_LEAF = 64;           // a split node has n > 128, so both halves hold >= 64 values
constexpr double EV_ATOL = 1.0e-8;        // np.allclose(m, 0.0)

struct EvScale {                          // scalars of scale_array for one field, in the order they are found
    double mean0, m1, sm, s, m2;
    int c1, c2;                           // the two "re-centre" branches taken?
};

// value of a field entry after `stage` steps of scale_array: 0 raw, 1 centred, 2 re-centred, 3 squared deviation
// (np.std's operand), 4 divided by the std, 5 final
__device__ __forceinline__ double ev_stage(double x, const EvScale &p, int stage) {
    if (stage == 0) return x;
    double v = __dsub_rn(x, p.mean0);
    if (stage == 1) return v;
    if (p.c1) v = __dsub_rn(v, p.m1);
    if (stage == 2) return v;
    if (stage == 3) {
        const double d = __dsub_rn(v, p.sm);
        return __dmul_rn(d, d);
    }
    v = __ddiv_rn(v, p.s);
    if (stage == 4) return v;
    if (p.c2) v = __dsub_rn(v, p.m2);
    return v;
}

// The pairwise tree of n values as a leaf table: leaf k covers [leaf_off[k], leaf_off[k+1]); after pushing leaf k's sum
// the evaluator performs merges[k] additions of the two topmost partial sums (post-order).  Returns the leaf count.
__device__ inline int ev_build_tree(int n, int *leaf_off, uint8_t *merges) {
    struct Node { int off, n; };
    Node st[64];
    int sp = 0, L = 0;
    st[sp++] = Node{0, n};
    while (sp) {
        const Node e = st[--sp];
        if (e.n < 0) {                    // merge marker: both children of a node are done
            ++merges[L - 1];
            continue;
        }
        if (e.n <= EV_PW_BLOCK) {
            leaf_off[L] = e.off;
            merges[L] = 0;
            ++L;
            continue;
        }
        int n2 = e.n / 2;
        n2 -= n2 % 8;
        st[sp++] = Node{0, -1};
        st[sp++] = Node{e.off + n2, e.n - n2};
        st[sp++] = Node{e.off, n2};
    }
    leaf_off[L] = n;
    return L;
}

// result[f] = 0.0 + pairwise_sum_i val(f, i) for f < n_fields, in NumPy's order.  All threads of the CTA call this.
template <class F>
__device__ __forceinline__ void ev_reduce(int n_fields, const int *leaf_off, const uint8_t *merges, int L, double *sums,
                                          int l_stride, double *result, F val) {
    const int grp = threadIdx.x >> 3, j = threadIdx.x & 7;
    const int items = L * n_fields;
    for (int base = 0; base < items; base += EV_THREADS / 8) {
        const int it = base + grp;
        const bool on = it < items;
        const int f = on ? it / L : 0, k = on ? it - f * L : 0;
        const int a = leaf_off[k], len = on ? leaf_off[k + 1] - a : 0;
        const int full = len & ~7;
        double r = 0.0;
        if (len >= 8) {
            r = val(f, a + j);
            for (int i = 8 + j; i < full; i += 8) r = __dadd_rn(r, val(f, a + i));
        }
        r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 1));
        r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 2));
        r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 4));
        if (on && j == 0) {
            if (len < 8) {
                r = 0.0;
                for (int i = 0; i < len; ++i) r = __dadd_rn(r, val(f, a + i));
            } else {
                for (int i = full; i < len; ++i) r = __dadd_rn(r, val(f, a + i));
            }
            sums[f * l_stride + k] = r;
        }
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && (threadIdx.x >> 5) < n_fields) {
        const int f = threadIdx.x >> 5;
        double st[40];
        int sp = 0;
        for (int k = 0; k < L; ++k) {
            st[sp++] = sums[f * l_stride + k];
            for (int m = merges[k]; m; --m) {
                --sp;
                st[sp - 1] = __dadd_rn(st[sp - 1], st[sp]);
            }
        }
        result[f] = __dadd_rn(0.0, st[0]);
    }
    __syncthreads();
}

// GENERIC path (reads too long to hold one field in shared memory; any length up to the leaf-table limit).
// out [total_events, 12]; l_max >= max over reads of n / 64 + 1 sizes the dynamic shared memory
template <typename OutT>
__global__ void __launch_bounds__(EV_THREADS) k_event_features(const double *__restrict__ mean,
                                                               const double *__restrict__ stdv,
                                                               const double *__restrict__ length,
                                                               const int64_t *__restrict__ read_off, OutT *out,
                                                               int l_max, int n_lo) {
    extern __shared__ __align__(16) unsigned char ev_smem[];
    double *sums = reinterpret_cast<double *>(ev_smem);                   // [3][l_max]
    int *leaf_off = reinterpret_cast<int *>(sums + 3 * (size_t)l_max);    // [l_max + 1]
    uint8_t *merges = reinterpret_cast<uint8_t *>(leaf_off + l_max + 1);  // [l_max]
    __shared__ double red[3];
    __shared__ EvScale sc[3];
    __shared__ double d_sm, d_s;
    __shared__ int s_L;

    const int64_t a0 = read_off[blockIdx.x];
    const int n = (int)(read_off[blockIdx.x + 1] - a0);
    if (n <= 0 || n < n_lo) return;       // shorter reads belong to k_event_features_smem
    const double *xm = mean + a0, *xs = stdv + a0, *xl = length + a0;
    if (threadIdx.x == 0) s_L = ev_build_tree(n, leaf_off, merges);
    __syncthreads();
    const int L = s_L;
    const double dn = (double)n;

    // scale_array on the three fields at once: five dependent reductions
    for (int stage = 0; stage <= 4; ++stage) {
        ev_reduce(3, leaf_off, merges, L, sums, l_max, red, [&](int f, int i) {
            const double *x = f == 0 ? xm : (f == 1 ? xs : xl);
            return ev_stage(x[i], sc[f], stage);
        });
        if (threadIdx.x < 3) {
            EvScale &p = sc[threadIdx.x];
            const double m = __ddiv_rn(red[threadIdx.x], dn);
            if (stage == 0) p.mean0 = m;
            else if (stage == 1) { p.m1 = m; p.c1 = !(fabs(m) <= EV_ATOL); }
            else if (stage == 2) p.sm = m;
            else if (stage == 3) { const double s = __dsqrt_rn(m); p.s = (s == 0.0) ? 1.0 : s; }
            else { p.m2 = m; p.c2 = !(fabs(m) <= EV_ATOL); }
        }
        __syncthreads();
    }

    // delta = ediff1d(scaled mean, to_begin=0) / np.std(delta): two more reductions
    auto diff = [&](int i) {
        return i == 0 ? 0.0 : __dsub_rn(ev_stage(xm[i], sc[0], 5), ev_stage(xm[i - 1], sc[0], 5));
    };
    ev_reduce(1, leaf_off, merges, L, sums, l_max, red, [&](int, int i) { return diff(i); });
    if (threadIdx.x == 0) d_sm = __ddiv_rn(red[0], dn);
    __syncthreads();
    ev_reduce(1, leaf_off, merges, L, sums, l_max, red, [&](int, int i) {
        const double d = __dsub_rn(diff(i), d_sm);
        return __dmul_rn(d, d);
    });
    if (threadIdx.x == 0) {
        const double s = __dsqrt_rn(__ddiv_rn(red[0], dn));
        d_s = (s == 0.0) ? 1.0 : s;
    }
    __syncthreads();

    // centre columns 4..7 of every row, then the neighbours' copies (columns 0..3 from row i-1, 8..11 from row i+1)
    OutT *o = out + a0 * 12;
    for (int i = threadIdx.x; i < n; i += EV_THREADS) {
        OutT *row = o + (int64_t)i * 12 + 4;
        row[0] = (OutT)ev_stage(xm[i], sc[0], 5);
        row[1] = (OutT)ev_stage(xs[i], sc[1], 5);
        row[2] = (OutT)ev_stage(xl[i], sc[2], 5);
        row[3] = (OutT)__ddiv_rn(diff(i), d_s);
    }
    __syncthreads();
    for (int64_t e = threadIdx.x; e < (int64_t)n * 8; e += EV_THREADS) {
        const int i = (int)(e >> 3), c = (int)(e & 7);                    // c < 4: pos -1, else pos +1
        const int src = c < 4 ? i - 1 : i + 1;
        const OutT v = (src >= 0 && src < n) ? o[(int64_t)src * 12 + 4 + (c & 3)] : (OutT)0;
        o[(int64_t)i * 12 + (c < 4 ? c : c + 4)] = v;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// FAST path: reads short enough for one field to live in shared memory (n <= ~9.1k events at three CTAs per SM, ~13.9k
// at two, ~27.9k at one; the reference's own callers cap reads at 9000 events).  The field is loaded once and updated IN PLACE the way
// scale_array updates its array (x -= mean, x /= std ...), so every value is read from HBM once, divided once, and
// written once.  The pairwise tree is held in HEAP form (children of slot q are 2q+1 and 2q+2): every thread derives
// the span of its own slot from n alone, so the tree is built by the whole CTA in a few integer steps, the leaf sums land
// in their slots in any order, and one warp folds the levels bottom-up between two block barriers.
constexpr int EV_SM_THREADS = 384;

__host__ __device__ inline int ev_tree_depth(int n) {     // depth of the deepest leaf (the right child is never smaller)
    int d = 0;
    while (n > EV_PW_BLOCK) {
        int n2 = n / 2;
        n2 -= n2 % 8;
        n -= n2;
        ++d;
    }
    return d;
}

struct EvTreeSm {
    double *vals;            // [heap] partial sums, slot q
    int2 *leaf;              // [l_max] {offset << 8 | length, slot} of every leaf, in no particular order
    uint8_t *inner;          // [heap] 1 = slot q is a split node
    double *tot;             // [2] totals of the last two passes (double-buffered: no barrier between read and reuse)
    int *n_leaf;             // leaf counter
    int L, D, pass;
};

// all threads; heap = 2^(D+1) - 1 slots
__device__ __forceinline__ void ev_build_tree_sm(int n, EvTreeSm &t) {
    t.D = ev_tree_depth(n);
    t.pass = 0;
    if (threadIdx.x == 0) *t.n_leaf = 0;
    __syncthreads();
    const int heap = (2 << t.D) - 1;
    for (int q = threadIdx.x; q < heap; q += EV_SM_THREADS) {
        const int d = 31 - __clz(q + 1);
        int off = 0, m = n;
        bool exists = true;
        for (int s = d - 1; s >= 0; --s) {                // walk down from the root along the bits of q + 1
            if (m <= EV_PW_BLOCK) { exists = false; break; }
            int n2 = m / 2;
            n2 -= n2 % 8;
            if (((q + 1) >> s) & 1) { off += n2; m -= n2; } else m = n2;
        }
        const bool split = exists && m > EV_PW_BLOCK;
        t.inner[q] = split ? 1 : 0;
        if (exists && !split) t.leaf[atomicAdd(t.n_leaf, 1)] = make_int2((off << 8) | m, q);
    }
    __syncthreads();
    t.L = *t.n_leaf;
}

// One reduction pass over the field: every element takes v = f(src + i, i), is stored to dst[i] when WRITE, and
// contributes term(v); returns 0.0 + the pairwise sum to every thread.  A lane walks its column of the leaf (elements
// j, j + 8, ...) through a pointer, four rows per trip, so that the address arithmetic stays off the dependent DADD chain.
template <bool WRITE, class F, class T>
__device__ __forceinline__ double ev_pass_sm(EvTreeSm &t, const double *src, double *dst, F f, T term) {
    const int grp = threadIdx.x >> 3, j = threadIdx.x & 7;
    for (int base = 0; base < t.L; base += EV_SM_THREADS / 8) {
        const int k = base + grp;
        const bool on = k < t.L;
        const int2 lf = on ? t.leaf[k] : make_int2(0, 0);
        const int a = lf.x >> 8, len = lf.x & 255;
        const int cnt = len >> 3;                       // full rows of eight
        double r = 0.0;
        if (cnt) {
            int i = a + j;
            const double *ps = src + i;
            double *pd = dst + i;
            auto one = [&](int o) {
                const double v = f(ps + o, i + o);
                if (WRITE) pd[o] = v;
                return term(v);
            };
            r = one(0);
            int c = 1;
            for (; c + 4 <= cnt; c += 4) {
                const double v0 = one(8), v1 = one(16), v2 = one(24), v3 = one(32);
                r = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(r, v0), v1), v2), v3);
                ps += 32; pd += 32; i += 32;
            }
            for (; c < cnt; ++c) {
                r = __dadd_rn(r, one(8));
                ps += 8; pd += 8; i += 8;
            }
        }
        r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 1));
        r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 2));
        r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 4));
        if (on && j == 0) {
            if (len < 8) r = 0.0;
            for (int i = a + (cnt << 3); i < a + len; ++i) {          // the n % 8 tail (all of a leaf shorter than 8)
                const double v = f(src + i, i);
                if (WRITE) dst[i] = v;
                r = __dadd_rn(r, term(v));
            }
            t.vals[lf.y] = r;
        }
    }
    __syncthreads();
    const int par = t.pass++ & 1;
    if (threadIdx.x < 32) {
        for (int d = t.D - 1; d >= 0; --d) {
            const int first = (1 << d) - 1;
            for (int q = first + (int)threadIdx.x; q < 2 * first + 1; q += 32)
                if (t.inner[q]) t.vals[q] = __dadd_rn(t.vals[2 * q + 1], t.vals[2 * q + 2]);
            __syncwarp();
        }
        if (threadIdx.x == 0) t.tot[par] = __dadd_rn(0.0, t.vals[0]);
    }
    __syncthreads();
    return t.tot[par];
}

// scale_array(with_mean=True) of one field: g (global) -> x (shared) holding the final values
__device__ __forceinline__ void ev_scale_field_sm(const double *__restrict__ g, double *x, int n, EvTreeSm &t) {
    const double dn = (double)n;
    auto id = [](double v) { return v; };
    const double mean0 = __ddiv_rn(ev_pass_sm<true>(t, g, x, [](const double *p, int) { return *p; }, id), dn);
    const double m1 = __ddiv_rn(ev_pass_sm<true>(t, x, x, [&](const double *p, int) { return __dsub_rn(*p, mean0); }, id), dn);
    const bool c1 = !(fabs(m1) <= EV_ATOL);
    // np.std's own mean of the array: without the re-centring the array, the order and hence the sum are those of
    // the previous pass, bit for bit
    const double sm = !c1 ? m1 : __ddiv_rn(ev_pass_sm<true>(t, x, x, [&](const double *p, int) { return __dsub_rn(*p, m1); }, id), dn);
    const double var = __ddiv_rn(ev_pass_sm<false>(t, x, x, [](const double *p, int) { return *p; },
                                                   [&](double v) { const double d = __dsub_rn(v, sm); return __dmul_rn(d, d); }), dn);
    double s = __dsqrt_rn(var);
    if (s == 0.0) s = 1.0;
    const double m2 = __ddiv_rn(ev_pass_sm<true>(t, x, x, [&](const double *p, int) { return __ddiv_rn(*p, s); }, id), dn);
    if (!(fabs(m2) <= EV_ATOL)) {
        for (int i = threadIdx.x; i < n; i += EV_SM_THREADS) x[i] = __dsub_rn(x[i], m2);
    }
    __syncthreads();
}

// one thread's share of an output group as a streaming 16-byte store (the output must not push the scratch out of
// L2); p is 16-byte aligned: out is, and every group starts at a multiple of four values
__device__ __forceinline__ void ev_store_group(double *p, const double (&v)[2]) {
    __stcs(reinterpret_cast<double2 *>(p), make_double2(v[0], v[1]));
}
__device__ __forceinline__ void ev_store_group(float *p, const double (&v)[4]) {
    __stcs(reinterpret_cast<float4 *>(p), make_float4((float)v[0], (float)v[1], (float)v[2], (float)v[3]));
}
template <int N>
__device__ __forceinline__ void ev_store_zero(double *p) {
    __stcs(reinterpret_cast<double2 *>(p), make_double2(0.0, 0.0));
}
template <int N>
__device__ __forceinline__ void ev_store_zero(float *p) {
    __stcs(reinterpret_cast<float4 *>(p), make_float4(0.f, 0.f, 0.f, 0.f));
}

__device__ __forceinline__ void ev_prefetch_l2(const double *g, int n) {
    const char *p = reinterpret_cast<const char *>(g);
    for (int b = threadIdx.x * 128; b < n * 8; b += EV_SM_THREADS * 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p + b));
}

// handles the reads with n_lo <= n <= n_cap; x holds n_cap doubles, the heap 2^(depth(n_cap)+1) slots; scratch
// [2][total_events] doubles (field-major standardised stdv and length -- coalesced, L2-resident between its write and
// its read a few microseconds later)
template <typename OutT>
__global__ void __launch_bounds__(EV_SM_THREADS) k_event_features_smem(const double *__restrict__ mean,
                                                                       const double *__restrict__ stdv,
                                                                       const double *__restrict__ length,
                                                                       const int64_t *__restrict__ read_off, OutT *out,
                                                                       double *scratch, int64_t total_events, int l_max,
                                                                       int heap, int n_lo, int n_cap) {
    extern __shared__ __align__(16) unsigned char ev_smem[];
    __shared__ double tot[2];
    __shared__ int n_leaf;
    const int64_t a0 = read_off[blockIdx.x];
    const int n = (int)(read_off[blockIdx.x + 1] - a0);
    if (n <= 0 || n < n_lo || n > n_cap) return;
    ev_prefetch_l2(stdv + a0, n);
    ev_prefetch_l2(length + a0, n);
    ev_prefetch_l2(mean + a0, n);
    double *x = reinterpret_cast<double *>(ev_smem);                      // [n_cap]
    EvTreeSm t;
    t.vals = x + n_cap;
    t.leaf = reinterpret_cast<int2 *>(t.vals + heap);
    t.inner = reinterpret_cast<uint8_t *>(t.leaf + l_max);
    t.tot = tot;
    t.n_leaf = &n_leaf;
    ev_build_tree_sm(n, t);
    const double dn = (double)n;
    double *f_sd = scratch + a0, *f_len = scratch + total_events + a0;

    ev_scale_field_sm(stdv + a0, x, n, t);
    for (int i = threadIdx.x; i < n; i += EV_SM_THREADS) f_sd[i] = x[i];
    __syncthreads();
    ev_scale_field_sm(length + a0, x, n, t);
    for (int i = threadIdx.x; i < n; i += EV_SM_THREADS) f_len[i] = x[i];
    __syncthreads();
    // mean last: it stays in shared memory for delta and for the output pass
    ev_scale_field_sm(mean + a0, x, n, t);
    auto diff = [](const double *p, int i) { return i == 0 ? 0.0 : __dsub_rn(p[0], p[-1]); };
    auto idn = [](double v) { return v; };
    const double d_sm = __ddiv_rn(ev_pass_sm<false>(t, x, x, diff, idn), dn);
    const double d_var = __ddiv_rn(ev_pass_sm<false>(t, x, x, diff, [&](double v) { const double d = __dsub_rn(v, d_sm); return __dmul_rn(d, d); }), dn);
    double d_s = __dsqrt_rn(d_var);
    if (d_s == 0.0) d_s = 1.0;

    // Output: the features of event i go, as one aligned group of four, to columns 4..7 of row i, 0..3 of row i+1 and
    // 8..11 of row i-1 (zeros past the read ends).  float32: one thread per event, one 16-byte store per group;
    // float64: two threads per event (features 0-1 / 2-3), so that a lane pair fills one 32-byte sector.
    OutT *o = out + a0 * 12;
    constexpr int TPE = sizeof(OutT) == 8 ? 2 : 1, FPT = 4 / TPE;   // threads per event, features per thread
    constexpr int UB = 2;                                 // groups in flight per thread: the scratch reads are L2 / HBM latency
    const int h = TPE == 2 ? (threadIdx.x & 1) : 0;       // EV_SM_THREADS is even: a thread keeps its half
    for (int e0 = threadIdx.x; e0 < n * TPE; e0 += EV_SM_THREADS * UB) {
        double v[UB][FPT];
#pragma unroll
        for (int u = 0; u < UB; ++u) {
            const int e = e0 + u * EV_SM_THREADS, i = e / TPE;
            if (e < n * TPE) {
                if (TPE == 1) {
                    v[u][0] = x[i];
                    v[u][1] = __ldcg(f_sd + i);
                    v[u][FPT - 2] = __ldcg(f_len + i);
                    v[u][FPT - 1] = __ddiv_rn(diff(x + i, i), d_s);
                } else if (h == 0) {
                    v[u][0] = x[i];
                    v[u][1] = __ldcg(f_sd + i);
                } else {
                    v[u][0] = __ldcg(f_len + i);
                    v[u][1] = __ddiv_rn(diff(x + i, i), d_s);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < UB; ++u) {
            const int e = e0 + u * EV_SM_THREADS, i = e / TPE;
            if (e < n * TPE) {
                OutT *row = o + i * 12 + h * FPT;
                ev_store_group(row + 4, v[u]);
                if (i + 1 < n) ev_store_group(row + 12, v[u]); else ev_store_zero<FPT>(row + 8);
                if (i > 0) ev_store_group(row - 4, v[u]); else ev_store_zero<FPT>(row);
            }
        }
    }
}

inline size_t ev_smem_bytes_sm(int n_cap, int l_max) {
    const size_t heap = ((size_t)2 << ev_tree_depth(n_cap)) - 1;
    return sizeof(double) * ((size_t)n_cap + heap) + sizeof(int2) * (size_t)l_max + heap + 16;
}

inline size_t ev_smem_bytes(int l_max) {
    return 3 * sizeof(double) * (size_t)l_max + sizeof(int) * ((size_t)l_max + 1) + (size_t)l_max + 16;
}

}  // namespace ntb
