// K4 (tensor-core form)  BLSTM recurrence on tcgen05 (network.py:266-288: rnn.LSTMCell x2 under
// tf.nn.bidirectional_dynamic_rnn).  TF semantics restated in ntb_lstm.cuh; this file is the
// sm_100a implementation used when 4H + Hp fits the 512 TMEM columns (H <= 100).
//
// One persistent CTA per (tile of 128 windows, direction); forward and backward tiles are CTAs
// of one launch.  Per timestep the recurrent term is a [128 x Hp] x [Hp x 4H] GEMM:
//   * B = W_hh^T, fp16 hi + lo planes, resident in shared memory for all T steps in the
//     non-swizzled K-major core-matrix layout (2 * 4H * Hp * 2 bytes = 175 KB at H = 100);
//   * A = h_{t-1}, fp16 hi + lo, lives in TENSOR MEMORY next to the accumulator ("TS" MMA):
//     TMEM columns [0,4H) gate pre-activations (fp32), [4H, 4H+Hp/2) h hi, [.., 4H+Hp) h lo;
//     the epilogue writes h_t straight back with tcgen05.st, so h never touches shared memory;
//   * fp16x3: D = hhi.Whi + hlo.Whi + hhi.Wlo accumulated in fp32 (relative error ~2^-16; the
//     recurrence has to hold the 1e-3 logits budget over 100 dependent steps).
// Warp roles: warp 0 issues the MMAs (one thread) and prefetches the next step's projected gates
// towards L2; warp 2 owns the TMEM allocation; warps 4-11 are the gate epilogue: thread = one
// window (TMEM lane), two warps per lane quarter take alternate 32-column chunks (8 hidden
// units x i,j,f,o), add the projected gates (coalesced float4 loads, layout below), apply the
// fused cell update with the cell state in registers, and write h_t to TMEM (next step's A) and
// to the layer output.
// Both directions run on a uniform clock: step s handles t = s (fw) or t = T-1-s (bw) for every
// window; a window is inactive while t >= len (state stays zero for bw, frozen-and-unused for
// fw), which is equivalent to TF's reverse_sequence formulation.
//
// Projected-gates layout (written by the projection GEMM's epilogue, ntb_gemm_tc.cu):
//   xproj[dir][tile*T + t][unit][row 0..127] float4 = (i, j, f, o) pre-activations with bias and
//   forget_bias folded -- one timestep of one tile is a contiguous 128*H*16 bytes.
#pragma once
#include "ntb_common.cuh"
#include <cuda.h>

#include "ntb_lstm_cell.cuh"
#include "ntb_tc_ptx.cuh"

namespace ntb {

constexpr int LSTM_TC_ROWS = 128;       // windows per tile = TMEM lanes
constexpr int LSTM_TC_THREADS = 384;    // 4 control warps + 8 epilogue warps
constexpr int LSTM_TC_MAXCH = 13;       // chunks of 8 units (13 at H = 100); each epilogue warp takes half a chunk

struct LstmTcParams {
    const float4 *xproj;              // see layout above
    const op16_t *whh_hi;      // [2][Hp/8][4H][8]  (K-chunk major, row n = unit*4 + gate)
    const op16_t *whh_lo;
    const int32_t *lens;              // [n_windows]
    float *out;                       // [n_windows*T][ldo], column dir*H + unit
    int ldo;
    int64_t n_windows;
    int n_tiles;                      // ceil(n_windows / 128)
    int T, H, Hp;                     // Hp = H rounded up to 16
    // next_hi/next_lo != null: h_t is emitted as the NEXT layer's projection operand (fp16 hi/lo planes, time-major
    // 128-window tiles: plane row (tile*T + t)*128 + window, next_K columns) instead of the fp32 `out`
    op16_t *next_hi, *next_lo;
    int next_K;
    // next_ilv != null (default between tensor-core BLSTM layers): same rows, ONE plane of 2 * next_K columns whose
    // 8-element groups are (4 x hi, 4 x lo) of 4 consecutive units -- a thread's 4 units are one 16-byte store, the
    // same store count as the fp32 output, and the next projection consumes the plane directly (GemmWeights::ilv)
    op16_t *next_ilv;
    int flags;                        // debug/ablation: 2 = no slice overlap, 16 = L2 prefetch of the next step, 32 = warp-3 L2 prefetch two steps ahead, 64 = hi.hi pass only, 128 = no MMA (timing only), 256 = no projected-gates stream (timing only), 512 = no output stores (timing only)
};

static inline bool lstm_tc_supported(int H) {
    const int Hp = (int)align_up(H, 16);
    return H >= 16 && H % 4 == 0 && 4 * H + Hp <= 512 && (4 * H <= 256 || (4 * H - 256) % 16 == 0) &&
           (H + 7) / 8 <= LSTM_TC_MAXCH;
}
constexpr int LSTM_TC_XSLOTS = 3;                  // ring of projected-gate chunks (8 units x 128 rows x float4)
constexpr int LSTM_TC_XSLOT_BYTES = 8 * LSTM_TC_ROWS * 16;
static inline int64_t lstm_tc_smem_bytes(int H) {
    const int Hp = (int)align_up(H, 16);
    return align_up(2LL * 4 * H * Hp * 2, 128) + LSTM_TC_XSLOTS * LSTM_TC_XSLOT_BYTES + 128;
}
static inline int64_t lstm_tc_whh_elems(int H) { return 2LL * align_up(H, 16) * 4 * H; }   // per plane, both dirs

// TF LSTMCell kernel rows D..D+H-1 (the h part), gate column blocks i|j|f|o  ->  fp16 hi/lo
// planes [dir][k/8][n = u*4+g][k%8], zero for k >= H.
__global__ void k_pack_whh_tc(const float *__restrict__ kernel_fw, const float *__restrict__ kernel_bw, int D, int H,
                              int Hp, op16_t *__restrict__ hi, op16_t *__restrict__ lo) {
    const int N = 4 * H;
    const int64_t per_dir = (int64_t)Hp * N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < 2 * per_dir;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int dir = (int)(i / per_dir);
        const int64_t j = i - dir * per_dir;
        const int k8 = (int)(j % 8), n = (int)((j / 8) % N), kc = (int)(j / (8 * N));
        const int k = kc * 8 + k8, u = n >> 2, g = n & 3;
        const float *kern = dir ? kernel_bw : kernel_fw;
        // pre-scaled by -log2(e) (i, f, o) / -2 log2(e) (j): the MMA yields the ex2 arguments directly
        const float sc = g == 1 ? -2.885390081777927f : -1.4426950408889634f;
        const float v = k < H ? sc * kern[(int64_t)(D + k) * 4 * H + g * H + u] : 0.0f;
        op16_t h, l;
        ptx::split_op16(v, h, l);
        hi[i] = h;
        lo[i] = l;
    }
}

__global__ void __launch_bounds__(LSTM_TC_THREADS, 1) k_lstm_recurrence_tc(const __grid_constant__ CUtensorMap map_x,
                                                                               const LstmTcParams p) {
    using namespace ptx;
    extern __shared__ __align__(128) uint8_t smem_w[];          // [2 planes][Hp/8][4H][8] fp16, then the x ring
    __shared__ uint64_t d_full[2], h_ready, x_full[LSTM_TC_XSLOTS], x_empty[LSTM_TC_XSLOTS];
    __shared__ uint32_t tmem_base_smem;

    const int H = p.H, Hp = p.Hp, T = p.T, N = 4 * H;
    const int dir = blockIdx.y, tile = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t plane_bytes = (uint32_t)N * Hp * 2;
    uint8_t *xring = smem_w + ((2 * plane_bytes + 127u) & ~127u);
    const int n_chunks = (H + 7) >> 3;
    const int N1 = N < 256 ? N : 256, N2 = N - N1;
    const bool two = N2 > 0;                                   // second N slice (gate columns >= 256)

    // ---- one-time setup: W_hh^T hi/lo planes -> smem, barriers, TMEM
    {
        const uint4 *src_hi = reinterpret_cast<const uint4 *>(p.whh_hi + (int64_t)dir * N * Hp);
        const uint4 *src_lo = reinterpret_cast<const uint4 *>(p.whh_lo + (int64_t)dir * N * Hp);
        uint4 *dst = reinterpret_cast<uint4 *>(smem_w);
        const int n16 = (int)(plane_bytes >> 4);
        for (int i = threadIdx.x; i < n16; i += LSTM_TC_THREADS) {
            dst[i] = __ldg(src_hi + i);
            dst[n16 + i] = __ldg(src_lo + i);
        }
        fence_proxy_async_smem();          // generic-proxy stores -> visible to tcgen05.mma reads
    }
    if (threadIdx.x == 0) {
        prefetch_tmap(&map_x);
        mbar_init(&d_full[0], 1);
        mbar_init(&d_full[1], 1);
        mbar_init(&h_ready, 8);
        for (int i = 0; i < LSTM_TC_XSLOTS; ++i) { mbar_init(&x_full[i], 1); mbar_init(&x_empty[i], 8); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(&tmem_base_smem, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;
    const uint32_t a_hi_col = (uint32_t)N, a_lo_col = (uint32_t)(N + Hp / 2);
    const int64_t m_tiles = (int64_t)p.n_tiles * T;
    const int64_t step_f4 = (int64_t)H * LSTM_TC_ROWS;           // float4 per (tile, t, dir)
    const float4 *xtile = p.xproj + ((int64_t)dir * m_tiles + (int64_t)tile * T) * step_f4;

    if (warp == 0) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc1 = idesc_f16_f32(128, N1), idesc2 = idesc_f16_f32(128, two ? N2 : 16);
            const uint32_t lbo = (uint32_t)N * 16, sbo = 128;     // K-chunk stride, 8-row group stride
            const uint32_t w_hi = smem_u32(smem_w), w_lo = w_hi + plane_bytes;
            const int k16n = Hp / 16;
            for (int s = 0; s < T; ++s) {
                mbar_wait_backoff(&h_ready, (uint32_t)(s & 1), 1000 + s);
                tc_fence_after();
                for (int half = 0; half < 2; ++half) {
                    if (half && !two) break;
                    const uint32_t d = tmem_base + (uint32_t)(half * 256);
                    const uint32_t idesc = half ? idesc2 : idesc1;
                    const uint32_t row_off = (uint32_t)(half * 256 / 8) * 128;
                    for (int j = 0; j < k16n; ++j) {
                        const uint32_t koff = (uint32_t)j * 2 * lbo + row_off;
                        const uint64_t bh = smem_desc_nosw(w_hi + koff, lbo, sbo);
                        const uint64_t bl = smem_desc_nosw(w_lo + koff, lbo, sbo);
                        const uint32_t ah = tmem_base + a_hi_col + (uint32_t)(8 * j);
                        const uint32_t al = tmem_base + a_lo_col + (uint32_t)(8 * j);
                        if (p.flags & 128) continue;              // timing ablation: no MMA at all
                        umma_f16_ts(d, ah, bh, idesc, j != 0);
                        if (p.flags & 64) continue;               // timing ablation: hi.hi pass only
                        umma_f16_ts(d, al, bh, idesc, 1);
                        umma_f16_ts(d, ah, bl, idesc, 1);
                    }
                    if (!(p.flags & 2)) umma_commit(&d_full[half]);   // slice 0 goes to the epilogue while slice 1 still runs
                }
                if (p.flags & 2) { umma_commit(&d_full[0]); if (two) umma_commit(&d_full[1]); }
                // next step's projected gates towards L2 while this step computes
                if (s + 1 < T && (p.flags & 16)) {   // off by default: measured slower (the TMA ring already runs ahead)
                    const int tn = dir ? T - 2 - s : s + 1;
                    const char *g = reinterpret_cast<const char *>(xtile + (int64_t)tn * step_f4);
                    const uint32_t bytes = (uint32_t)step_f4 * 16, piece = (bytes / 4) & ~15u;
                    for (int q = 0; q < 4; ++q)
                        bulk_prefetch_l2(g + (size_t)q * piece, q == 3 ? bytes - 3 * piece : piece);
                }
            }
        }
    } else if (warp == 3) {
        // ===================== L2 prefetcher (flags & 32): projected gates of step s + 2 towards L2, paced by the MMAs
        if (lane == 0 && (p.flags & 32)) {
            const uint32_t bytes = (uint32_t)step_f4 * 16, piece = (bytes / 8) & ~15u;
            for (int s = 0; s + 2 < T; ++s) {
                if (s > 0) mbar_wait_backoff(&d_full[0], (uint32_t)((s - 1) & 1), 6000 + s);
                const int tn = dir ? T - 3 - s : s + 2;
                const char *g = reinterpret_cast<const char *>(xtile + (int64_t)tn * step_f4);
                for (int q = 0; q < 8; ++q)
                    bulk_prefetch_l2(g + (size_t)q * piece, q == 7 ? bytes - 7 * piece : piece);
            }
        }
    } else if (warp == 1) {
        // ===================== x producer: projected-gate chunks -> smem ring (TMA) ==========
        if (lane == 0 && !(p.flags & 256)) {
            uint32_t slot = 0, ph = 0;
            for (int s = 0; s < T; ++s) {
                const int t = dir ? T - 1 - s : s;
                const int64_t unit_t = ((int64_t)dir * m_tiles + (int64_t)tile * T + t) * H;
                for (int c = 0; c < n_chunks; ++c) {
                    mbar_wait_backoff(&x_empty[slot], ph ^ 1, 2000 + s);
                    // TMA box {256 floats, 2, 8 units}: always 16 KB (a short last chunk over-reads the next
                    // slab / zero-fills past the end; those units are never used)
                    mbar_expect_tx(&x_full[slot], (uint32_t)LSTM_TC_XSLOT_BYTES);
                    tma_load_3d(xring + slot * LSTM_TC_XSLOT_BYTES, &map_x, &x_full[slot], 0, 0, (int)(unit_t + c * 8));
                    if (++slot == LSTM_TC_XSLOTS) { slot = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp >= 4) {
        // ===================== gate epilogue =====================
        // thread = one window (TMEM lane); the two warps of a lane quarter split every 8-unit chunk in
        // halves (4 units = 16 accumulator columns each), so all 8 warps consume every ring slot in order
        // (parity waits stay unambiguous) and share the MUFU work evenly.  H % 4 == 0, so a warp's 4 units
        // are all real or all padding (padding: skipped; its A columns stay at their initial zero).
        const int q = warp & 3, half = (warp - 4) >> 2;
        const int r = q * 32 + lane;
        const int64_t w = (int64_t)tile * LSTM_TC_ROWS + r;
        const bool exists = w < p.n_windows;
        int len = exists ? p.lens[w] : 0;
        len = len > T ? T : (len < 0 ? 0 : len);    // dynamic_rnn's sequence_length, clamped to [0, T]
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t d_col = lane_base + (uint32_t)(16 * half);
        const uint32_t ah_col = lane_base + a_hi_col + (uint32_t)(2 * half), al_col = lane_base + a_lo_col + (uint32_t)(2 * half);
        float cst[LSTM_TC_MAXCH][4];
#pragma unroll
        for (int ci = 0; ci < LSTM_TC_MAXCH; ++ci)
#pragma unroll
            for (int u = 0; u < 4; ++u) cst[ci][u] = 0.0f;
        // h_0 = 0: clear the A columns (hi and lo, pad units included)
        if (half == 0)
            for (int c4 = 0; c4 < Hp; c4 += 4) tmem_st_32x4(lane_base + a_hi_col + (uint32_t)c4, 0u, 0u, 0u, 0u);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&h_ready);

        const float4 *xsm0 = reinterpret_cast<const float4 *>(xring) + half * 4 * LSTM_TC_ROWS + r;
        uint32_t slot = 0, xph = 0;                              // ring position
        float *orow = p.out + (w * T + (dir ? T - 1 : 0)) * (int64_t)p.ldo + dir * H + half * 4;
        const int64_t ostep = dir ? -(int64_t)p.ldo : (int64_t)p.ldo;
        const bool planes = p.next_hi != nullptr, ilv = p.next_ilv != nullptr;
        int64_t poff = (((int64_t)tile * T + (dir ? T - 1 : 0)) * LSTM_TC_ROWS + r) * p.next_K + dir * H + half * 4;
        const int64_t pstep = (dir ? -1 : 1) * (int64_t)LSTM_TC_ROWS * p.next_K;
        for (int s = 0; s < T; ++s, orow += ostep, poff += pstep) {
            const int t = dir ? T - 1 - s : s;
            const bool active = t < len;
            uint32_t dh[8][2], dl[8][2];                         // h of the slice-0 chunks, written once slice 1 is done
            mbar_wait(&d_full[0], (uint32_t)(s & 1), 3000 + s);
            tc_fence_after();
            uint32_t acc[16];
            // software pipeline: the TMEM load of chunk c + 1 is in flight during the math of chunk c
            tmem_ld_32x16(d_col, acc);
#pragma unroll
            for (int c = 0; c < LSTM_TC_MAXCH; ++c) {
                if (c == 8) {
                    // chunks >= 8 live in N slice 1; once it is complete no MMA reads h_{t-1} any more,
                    // so the deferred h_t of the slice-0 chunks may overwrite the A columns
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        if (k < n_chunks && 8 * k + 4 * half < H) {
                            tmem_st_32x2(ah_col + (uint32_t)(4 * k), dh[k][0], dh[k][1]);
                            tmem_st_32x2(al_col + (uint32_t)(4 * k), dl[k][0], dl[k][1]);
                        }
                    }
                }
                if (c < n_chunks) {
                    const bool real = c * 8 + half * 4 < H;     // warp-uniform
                    if (!(p.flags & 256)) mbar_wait(&x_full[slot], xph, 5000 + s);
                    float4 xv[4];
                    uint32_t a[16];
                    if (real) {
                        const float4 *xsm = xsm0 + slot * (LSTM_TC_XSLOT_BYTES / 16);
#pragma unroll
                        for (int u = 0; u < 4; ++u) xv[u] = xsm[u * LSTM_TC_ROWS];
                    }
                    tmem_ld_wait_dep(acc);
#pragma unroll
                    for (int i = 0; i < 16; ++i) a[i] = acc[i];
                    if (c + 1 < n_chunks) {
                        if (c + 1 == 8 && two) {                 // next chunk is the first of slice 1
                            mbar_wait(&d_full[1], (uint32_t)(s & 1), 4000 + s);
                            tc_fence_after();
                        }
                        tmem_ld_32x16(d_col + (uint32_t)(32 * (c + 1)), acc);
                    }
                    if (real) {
                        float hv[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            float cc = cst[c][u], hh;
                            lstm_cell_update(__uint_as_float(a[4 * u + 0]) + xv[u].x, __uint_as_float(a[4 * u + 1]) + xv[u].y,
                                             __uint_as_float(a[4 * u + 2]) + xv[u].z, __uint_as_float(a[4 * u + 3]) + xv[u].w,
                                             cc, hh);
                            cst[c][u] = active ? cc : cst[c][u];
                            hv[u] = active ? hh : 0.0f;
                        }
                        uint32_t ph2[2], pl2[2];
                        split_pack_op16x2(hv[0], hv[1], ph2[0], pl2[0]);
                        split_pack_op16x2(hv[2], hv[3], ph2[1], pl2[1]);
                        if (c < 8) {
                            dh[c][0] = ph2[0]; dh[c][1] = ph2[1]; dl[c][0] = pl2[0]; dl[c][1] = pl2[1];
                        } else {
                            tmem_st_32x2(ah_col + (uint32_t)(4 * c), ph2[0], ph2[1]);
                            tmem_st_32x2(al_col + (uint32_t)(4 * c), pl2[0], pl2[1]);
                        }
                        if (ilv) {          // rows of windows past the end are written too (zeros: len = 0)
                            // interleaved plane: twice the columns, same rows
                            *reinterpret_cast<uint4 *>(p.next_ilv + 2 * poff + 16 * c) = make_uint4(ph2[0], ph2[1], pl2[0], pl2[1]);
                        } else if (planes) {
                            *reinterpret_cast<uint2 *>(p.next_hi + poff + 8 * c) = make_uint2(ph2[0], ph2[1]);
                            *reinterpret_cast<uint2 *>(p.next_lo + poff + 8 * c) = make_uint2(pl2[0], pl2[1]);
                        } else if (exists && !(p.flags & 512)) {
                            *reinterpret_cast<float4 *>(orow + 8 * c) = make_float4(hv[0], hv[1], hv[2], hv[3]);
                        }
                    }
                    // every lane has consumed its x values (data dependence) -> the slot may be refilled
                    __syncwarp();
                    if (lane == 0 && !(p.flags & 256)) mbar_arrive(&x_empty[slot]);
                    if (++slot == LSTM_TC_XSLOTS) { slot = 0; xph ^= 1; }
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&h_ready);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace ntb
