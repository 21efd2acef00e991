// the CTC decoders next to the fused readout.
//   k_ctc_greedy       tf.nn.ctc_greedy_decoder with BOTH outputs (nanotensor/network.py:528-533 keeps [0] and drops
//                      neg_sum_logits; chiron/chiron/chiron_rcnn_train.py:68): first-max class per frame, emitted iff not
//                      blank and (not merge_repeated or != previous frame's class); neg_sum_logits = -(sum of the frame
//                      maxima), accumulated in frame order in float32 like CTCGreedyDecoder::Decode.
//   k_ctc_beam_search  tf.nn.ctc_beam_search_decoder (chiron_rcnn_train.py:66-67: merge_repeated=False, top_paths=k,
//                      beam_width 100): ctc_beam_search.h::CTCBeamSearchDecoder::Step restated for one warp per sequence.
// The beam is a set of prefix-tree nodes.  TF keeps the tree as heap objects (BeamEntry) that SURVIVE leaving the beam: a
// node that is evicted and later proposed again by its parent is the same object, and its children still point at it.
// The same identity is kept here: the tree lives in a per-warp global scratch (node = parent, label, entry slot, child
// ids), the beam itself (<= 128 entries: log-probabilities blank / label / total, node id) in shared memory.
// Per frame:  input = logits - max (TF 1.4 uses this AS log-probability; log_softmax = 1 normalises it first);
//   every leaf: label' = LSE(label, parent active ? (same label ? parent.blank : parent.total) : -inf) + input[label],
//               blank' = total + input[blank], total' = LSE(blank', label');
//   every inactive child c of a leaf b: total' = label' = input[c] + (c == b.label ? b.blank : b.total), blank' = -inf;
//   the leaves are pushed back first (there are at most beam_width of them); then, leaf by leaf in descending OLD total,
//   a leaf proposes its children -- only while its old total still beats the current bottom of a full beam -- and a
//   child enters iff its total beats the current bottom, which it evicts (an evicted leaf loses its probabilities).
//   This is TF's order-dependent procedure, kept as it is: a leaf evicted before its parent's turn is proposed again as a
//   fresh child, loses, has its OLD probabilities reset too and therefore does not expand in this frame -- the result is
//   not always the beam_width best totals.  The warp walks the leaves sequentially; the (<= 7) children of a leaf are
//   evaluated by lanes, the current bottom is a warp arg-min over the live slots.
// TopPaths: the top_paths best leaves by total, log_probability = total; merge_repeated collapses the OUTPUT labels.
#pragma once
#include "ntb_common.cuh"

namespace ntb {

constexpr int CTC_MAXW = 128;        // beam width limit (TF default 100)
constexpr int CTC_MAXC = 8;          // classes incl. blank
constexpr int CTC_WARPS = 4;
constexpr int CTC_NODE_I32 = 12;     // parent, label, entry, pad, child[8]

__global__ void __launch_bounds__(256) k_ctc_greedy(const float *__restrict__ logits, const int32_t *__restrict__ lens,
                                                    int64_t n, int T, int C, int merge_repeated,
                                                    int32_t *__restrict__ labels, int32_t *__restrict__ label_len,
                                                    float *__restrict__ neg_sum) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (w >= n) return;
    int len = lens[w];
    len = len > T ? T : (len < 0 ? 0 : len);
    const float *row0 = logits + w * (int64_t)T * C;
    int prev_last = -1, cnt = 0;
    float acc = 0.0f;
    for (int t0 = 0; t0 < len; t0 += 32) {
        const int t = t0 + lane;
        int best = -1;
        float bv = 0.0f;
        if (t < len) {
            const float *r = row0 + (int64_t)t * C;
            best = 0; bv = r[0];
            for (int c = 1; c < C; ++c) { const float v = r[c]; if (v > bv) { bv = v; best = c; } }   // first max
        }
        int prev = __shfl_up_sync(0xffffffffu, best, 1);
        if (lane == 0) prev = prev_last;
        const bool emit = t < len && best != C - 1 && !(merge_repeated && best == prev);
        const uint32_t m = __ballot_sync(0xffffffffu, emit);
        if (emit) labels[w * T + cnt + __popc(m & ((1u << lane) - 1))] = best;
        cnt += __popc(m);
        const int last_lane = (len - t0 < 32 ? len - t0 : 32) - 1;
        prev_last = __shfl_sync(0xffffffffu, best, last_lane);
        // frame order, float32: acc -= max_t
        for (int k = 0; k <= last_lane; ++k) acc -= __shfl_sync(0xffffffffu, bv, k);
    }
    for (int i = cnt + lane; i < T; i += 32) labels[w * T + i] = -1;
    if (lane == 0) {
        if (label_len) label_len[w] = cnt;
        if (neg_sum) neg_sum[w] = acc;
    }
}

__device__ __forceinline__ float ctc_lse(float a, float b) {
    if (a == -INFINITY && b == -INFINITY) return -INFINITY;
    return a > b ? a + log1pf(expf(b - a)) : b + log1pf(expf(a - b));
}
__device__ __forceinline__ uint32_t ctc_key(float f) {   // order-preserving float -> uint
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct CtcBeamSmem {
    float nb[2][CTC_MAXW], nl[2][CTC_MAXW], nt[2][CTC_MAXW];    // ping-pong beam
    int32_t node[2][CTC_MAXW];
    int8_t label[2][CTC_MAXW];
    int16_t parent[CTC_MAXW];                                    // entry slot of the parent if it is in the beam, else -1
    int16_t plabel[CTC_MAXW];                                    // the parent's label
    float tb[CTC_MAXW], tl[CTC_MAXW];                            // updated blank / label of the leaves
    float tt[CTC_MAXW];                                          // updated total of the leaves
    float lt[CTC_MAXW];                                          // live slots of the new beam: total
    int32_t lsrc[CTC_MAXW];                                      // .. source: leaf slot j >= 0, or -(1 + b * 8 + c) for child c of leaf b
    int16_t ord[CTC_MAXW];                                       // leaves in descending old total
    uint8_t evicted[CTC_MAXW], disabled[CTC_MAXW];               // leaf left the beam in this frame / may not expand in this frame
    float inp[CTC_MAXC];
    uint8_t path[512];
};

// logits [n][T][C]; out_paths [n][top_paths][T] (-1 padded), out_len [n][top_paths], out_logp [n][top_paths];
// tree: CTC_NODE_I32 * (T * W + 1) int32 per warp of the grid.
__global__ void __launch_bounds__(CTC_WARPS * 32) k_ctc_beam_search(
    const float *__restrict__ logits, const int32_t *__restrict__ lens, int64_t n, int T, int C, int W, int top_paths,
    int merge_repeated, int log_softmax, int32_t *__restrict__ tree_all, int32_t *__restrict__ out_paths,
    int32_t *__restrict__ out_len, float *__restrict__ out_logp) {
    extern __shared__ __align__(16) uint8_t ctc_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    CtcBeamSmem &S = reinterpret_cast<CtcBeamSmem *>(ctc_smem)[warp];
    const int64_t gwarp = (int64_t)blockIdx.x * CTC_WARPS + warp, n_warps = (int64_t)gridDim.x * CTC_WARPS;
    int32_t *tree = tree_all + gwarp * (int64_t)CTC_NODE_I32 * ((int64_t)T * W + 1);
    const int blank = C - 1, NC = C - 1;
    for (int64_t w = gwarp; w < n; w += n_warps) {
        int len = lens[w];
        len = len > T ? T : (len < 0 ? 0 : len);
        // root: label -1, blank = total = 0 (log 1), label prob = log 0
        int cur = 0, n_ent = 1, n_nodes = 1;
        if (lane == 0) {
            S.nb[0][0] = 0.0f; S.nl[0][0] = -INFINITY; S.nt[0][0] = 0.0f; S.node[0][0] = 0; S.label[0][0] = -1;
            tree[0] = -1; tree[1] = -1; tree[2] = 0;
            for (int c = 0; c < 8; ++c) tree[4 + c] = -1;
        }
        __syncwarp();
        for (int t = 0; t < len; ++t) {
            // ---- input = logits - max (- log sum exp)
            {
                const float v = lane < C ? logits[(w * T + t) * (int64_t)C + lane] : -INFINITY;
                float mx = v;
                for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                float x = v - mx;
                if (log_softmax) {
                    float e = lane < C ? expf(x) : 0.0f, sum = 0.0f;
                    for (int c = 0; c < C; ++c) sum += __shfl_sync(0xffffffffu, e, c);      // class order, float32
                    x -= logf(sum);
                }
                if (lane < C) S.inp[lane] = x;
            }
            // ---- parent slots of the leaves (tree identity: the parent node's entry slot, -1 when not in the beam)
            for (int j = lane; j < n_ent; j += 32) {
                const int32_t pn = tree[(int64_t)S.node[cur][j] * CTC_NODE_I32];
                int ps = -1, pl = -2;
                if (pn >= 0) { ps = tree[(int64_t)pn * CTC_NODE_I32 + 2]; pl = tree[(int64_t)pn * CTC_NODE_I32 + 1]; }
                S.parent[j] = (int16_t)ps; S.plabel[j] = (int16_t)pl;
            }
            __syncwarp();
            // ---- leaves
            const float in_blank = S.inp[blank];
            for (int j = lane; j < n_ent; j += 32) {
                const int l = S.label[cur][j], ps = S.parent[j];
                float nlv = S.nl[cur][j];
                if (l >= 0) {
                    if (ps >= 0) nlv = ctc_lse(nlv, l == S.plabel[j] ? S.nb[cur][ps] : S.nt[cur][ps]);
                    nlv += S.inp[l];
                }
                const float nbv = S.nt[cur][j] + in_blank;
                S.tb[j] = nbv; S.tl[j] = nlv;
                S.tt[j] = ctc_lse(nbv, nlv);
            }
            __syncwarp();
            // ---- the leaves go back into the (new) beam first; order of expansion: descending old total (ties: slot)
            for (int j = lane; j < n_ent; j += 32) {
                S.lt[j] = S.tt[j]; S.lsrc[j] = j; S.evicted[j] = 0; S.disabled[j] = 0;
                const float me = S.nt[cur][j];
                int r = 0;
                for (int i = 0; i < n_ent; ++i) { const float o = S.nt[cur][i]; r += (o > me) || (o == me && i < j); }
                S.ord[r] = (int16_t)j;
            }
            __syncwarp();
            int n_live = n_ent;
            float bmin = INFINITY;
            int bslot = -1;
            auto find_bottom = [&]() {          // arg-min of the live totals (ties: the highest slot), whole warp
                float m = INFINITY; int ms = -1;
                for (int q = lane; q < n_live; q += 32) { const float v = S.lt[q]; if (v < m || (v == m && q > ms)) { m = v; ms = q; } }
                for (int o = 16; o > 0; o >>= 1) {
                    const float om = __shfl_xor_sync(0xffffffffu, m, o);
                    const int os = __shfl_xor_sync(0xffffffffu, ms, o);
                    if (os >= 0 && (ms < 0 || om < m || (om == m && os > ms))) { m = om; ms = os; }
                }
                bmin = m; bslot = ms;
            };
            if (n_live == W) find_bottom();
            for (int r = 0; r < n_ent; ++r) {
                const int b = S.ord[r];
                const float ot = S.nt[cur][b];
                if (S.disabled[b] || !(ot > -INFINITY) || (n_live == W && !(ot > bmin))) continue;    // is_candidate(b->oldp)
                // the children of b, one per lane
                float tot = -INFINITY;
                int eslot = -1;                        // slot the child's node held at the start of the frame (-1: not a leaf)
                if (lane < NC) {
                    const int32_t cn = tree[(int64_t)S.node[cur][b] * CTC_NODE_I32 + 4 + lane];
                    if (cn >= 0) eslot = tree[(int64_t)cn * CTC_NODE_I32 + 2];
                    tot = S.inp[lane] + (lane == S.label[cur][b] ? S.nb[cur][b] : S.nt[cur][b]);
                }
                for (int c = 0; c < NC; ++c) {
                    const float tc = __shfl_sync(0xffffffffu, tot, c);
                    const int es = __shfl_sync(0xffffffffu, eslot, c);
                    // active NOW: a leaf of this frame that has not been evicted so far (an earlier child of this very
                    // leaf may just have evicted it)
                    const bool ac = es >= 0 && !S.evicted[es];
                    const int rl = (es >= 0 && !ac) ? es : -1;     // a leaf that left the beam in this frame, proposed again
                    if (ac) continue;
                    if (tc > -INFINITY && (n_live < W || tc > bmin)) {
                        int slot;
                        if (n_live == W) {
                            slot = bslot;
                            const int src = S.lsrc[slot];
                            if (src >= 0 && lane == 0) S.evicted[src] = 1;          // the bottom leaves the beam
                        } else {
                            slot = n_live++;
                        }
                        if (lane == 0) { S.lt[slot] = tc; S.lsrc[slot] = -(1 + b * 8 + c); }
                        __syncwarp();
                        if (n_live == W) find_bottom();
                    } else if (rl >= 0) {
                        if (lane == 0) S.disabled[rl] = 1;                          // c.oldp.Reset(): no expansion in this frame
                        __syncwarp();
                    }
                }
            }
            __syncwarp();
            // ---- the new beam = the live slots
            const int nxt = cur ^ 1;
            for (int j = lane; j < n_ent; j += 32)
                if (S.evicted[j]) tree[(int64_t)S.node[cur][j] * CTC_NODE_I32 + 2] = -1;   // probabilities are dropped
            int n_created = 0;
            for (int s0 = 0; s0 < n_live; s0 += 32) {
                const int q = s0 + lane;
                int src = 0, bb = 0, cc = 0;
                int32_t cn = -1;
                bool child = false;
                if (q < n_live) {
                    src = S.lsrc[q];
                    child = src < 0;
                    if (child) { const int code = -src - 1; bb = code >> 3; cc = code & 7; cn = tree[(int64_t)S.node[cur][bb] * CTC_NODE_I32 + 4 + cc]; }
                }
                const bool need = child && cn < 0;
                const uint32_t mn = __ballot_sync(0xffffffffu, need);
                if (q < n_live) {
                    if (!child) {
                        S.nb[nxt][q] = S.tb[src]; S.nl[nxt][q] = S.tl[src]; S.nt[nxt][q] = S.lt[q];
                        S.node[nxt][q] = S.node[cur][src]; S.label[nxt][q] = S.label[cur][src];
                        tree[(int64_t)S.node[cur][src] * CTC_NODE_I32 + 2] = q;
                    } else {
                        if (need) {
                            cn = n_nodes + n_created + __popc(mn & ((1u << lane) - 1));
                            int32_t *nd = tree + (int64_t)cn * CTC_NODE_I32;
                            nd[0] = S.node[cur][bb]; nd[1] = cc;
                            for (int z = 0; z < 8; ++z) nd[4 + z] = -1;
                            tree[(int64_t)S.node[cur][bb] * CTC_NODE_I32 + 4 + cc] = cn;
                        }
                        tree[(int64_t)cn * CTC_NODE_I32 + 2] = q;
                        S.nb[nxt][q] = -INFINITY; S.nl[nxt][q] = S.lt[q]; S.nt[nxt][q] = S.lt[q];
                        S.node[nxt][q] = cn; S.label[nxt][q] = (int8_t)cc;
                    }
                }
                n_created += __popc(mn);
            }
            const int n_new = n_live;
            n_nodes += n_created;
            n_ent = n_new;
            cur = nxt;
            __threadfence_block();
            __syncwarp();
        }
        // ---- TopPaths: the top_paths best leaves by total (ties: lower slot), label sequences from the tree
        for (int k = 0; k < top_paths; ++k) {
            float best = -INFINITY;
            int bj = -1;
            for (int j = lane; j < n_ent; j += 32)
                if (S.nt[cur][j] > best) { best = S.nt[cur][j]; bj = j; }
            for (int o = 16; o > 0; o >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
                if (oj >= 0 && (ob > best || (ob == best && (bj < 0 || oj < bj)))) { best = ob; bj = oj; }
            }
            int32_t *op = out_paths + (w * top_paths + k) * (int64_t)T;
            int plen = 0;
            if (bj >= 0) {
                if (lane == 0) {
                    int cnt = 0;
                    for (int32_t nd = S.node[cur][bj]; nd > 0 && cnt < 512; nd = tree[(int64_t)nd * CTC_NODE_I32])
                        S.path[cnt++] = (uint8_t)tree[(int64_t)nd * CTC_NODE_I32 + 1];
                    int prev = -1;
                    for (int i = cnt - 1; i >= 0; --i) {              // root -> leaf order
                        const int c = S.path[i];
                        if (!(merge_repeated && c == prev)) op[plen++] = c;
                        prev = c;
                    }
                    S.nt[cur][bj] = -INFINITY;                         // taken
                }
                plen = __shfl_sync(0xffffffffu, plen, 0);
            }
            __syncwarp();
            for (int i = plen + lane; i < T; i += 32) op[i] = -1;
            if (lane == 0) {
                out_len[w * top_paths + k] = plen;
                out_logp[w * top_paths + k] = bj >= 0 ? best : -INFINITY;
            }
        }
        __syncwarp();
    }
}

}  // namespace ntb
