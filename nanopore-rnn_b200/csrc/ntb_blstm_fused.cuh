// K3+K4 fused  One BLSTM layer (network.py:266-288: rnn.LSTMCell x2 under tf.nn.bidirectional_dynamic_rnn) as ONE
// persistent kernel: the input projection x_t . W_ih AND the recurrent term h_{t-1} . W_hh of every timestep are
// issued into the same TMEM accumulator, so the projected gates never exist in memory (the separate projection GEMM
// wrote 2.6 GB of them per 8192 windows and the recurrence read them back).  TF semantics as restated in ntb_lstm.cuh /
// ntb_lstm_tc.cuh (gate order i, j, f, o; forget_bias folded into the bias; dynamic_rnn masking; the backward direction
// runs t = T-1 .. 0 on the same clock with windows inactive while t >= len).
//
// Work item = (pair of 128-window tiles, direction), processed by a CTA PAIR (cluster of 2, tcgen05 cta_group::2,
// M = 256): each CTA owns one tile (its TMEM lanes = its windows) and HALF of every weight operand's rows, so W_ih
// streams from L2 once per 256 windows and W_hh (hi + lo planes, 175 KB per direction) is resident as 2 x 88 KB.
//   TMEM (512 columns per CTA): [0, NA) and [NA, 4H) gate pre-activations of the two N halves (fp32; NA = 208,
//     NB = 192 at H = 100: cta_group::2 needs N % 16 == 0, N <= 256), [4H, 4H + Hp/2) h hi, [.., 4H + Hp) h lo -- h_{t-1}
//     is the A operand of the recurrent MMAs straight from tensor memory ("TS"), written by the epilogue with
//     tcgen05.st: it never touches shared memory.
//   per step and N half:  D  = x_t . W_ih[:, half]   K = Kp in KC chunks of 16, operands streamed by TMA (ring of
//                               stages {x hi, x lo [128 x 16] of the CTA's own tile | W hi, W lo [N/4.. rows x 16]}),
//                         D += h_{t-1} . W_hh[:, half]   K = Hp from TMEM x resident smem,
//     every product as fp16x3 (hi.hi + lo.hi + hi.lo, fp32 accumulate).
//   Issue order of the leader's MMA thread per step s:  P_A(s) [as soon as the epilogue of step s-1 has read the A half
//     out of TMEM]  ->  R_A(s) [h_{s-1} complete]  ->  P_B(s)  ->  R_B(s);  the epilogue (8 warps per CTA, thread =
//     window, fused cell update with c in registers) works on half A while the tensor pipe runs P_B / R_B and on
//     half B while it already runs P_A(s+1): the tensor pipe is busy through the whole step.
//   x is streamed once per N half (2 x 64 KB per step and CTA) -- the price of starting P_A(s+1) under the epilogue.
//
// Operand images (all "K-major, no swizzle" UMMA core-matrix layout: 8 rows x 16 bytes contiguous, SBO = 128 B between
// 8-row groups, LBO = rows x 16 B between the two 8-element K groups of one MMA):
//   x planes  [tile][t][K8][128 windows][8] fp16 hi / lo, K8 = 2 KC  (gemm_tc_split_kmajor / gemm_tc_combine layout 3 /
//             this kernel's own output for the next layer): chunk kc of (tile, t) = 4 KB contiguous per plane
//   W_ih      [dir][rank][kc][plane][ half A: 2 x RA x 8 | half B: 2 x RB x 8 ],  RA = NA/2, RB = NB/2 rows per CTA
//   W_hh      [dir][rank][plane][half][Hp/8][R_half][8]
// Gate columns n = 4 u + g as everywhere; rows of half h held by CTA `rank`: n = h NA + rank R_h + row.
// Weights and bias are pre-scaled by -log2 e (i, f, o) / -2 log2 e (j): the MMAs produce ex2 arguments (ntb_lstm_cell.cuh).
#pragma once
#include "ntb_common.cuh"
#include <cuda.h>

#include "ntb_gemm_tc.cuh"
#include "ntb_lstm_cell.cuh"
#include "ntb_tc_ptx.cuh"

namespace ntb {

constexpr int BF_THREADS = 384;          // warps: 0 TMA producer, 1 MMA issuer (leader CTA), 2 TMEM, 3 idle, 4-11 epilogue
constexpr int BF_MAXI = 13;              // half-chunks (4 units = 16 accumulator columns) per epilogue warp: ceil(25 / 2)
constexpr int BF_MAXDEF = 7;             // of those, at most this many lie in N half A (their h_t is written late)
constexpr int BF_XBYTES = 128 * 16 * 2;  // one plane of an x chunk: 128 windows x 16 channels x fp16
constexpr int BF_SMEM_MAX = 227 * 1024;

struct BlstmGeom {
    int H, Hp, N, NA, NB, RA, RB;        // N = 4H; NA + NB = N; R = rows per CTA of each half
    int D, KC, K8;                       // input width, k16 chunks, 8-channel groups per (tile, t)
    int whh_bytes;                       // resident W_hh per CTA (both planes, both halves)
    int wst_bytes;                       // weight part of a stage: 2 planes x max(RA, RB) x 32 B
    int stage_bytes, stages;
    int64_t smem;
};

static inline bool blstm_fused_geom(int H, int D, BlstmGeom &g) {
    g.H = H; g.Hp = (int)align_up(H, 16); g.N = 4 * H; g.D = D;
    if (H < 16 || H % 4 || D % 8 || D < 16) return false;
    g.NA = 16 * ((g.N + 31) / 32);
    g.NB = g.N - g.NA;
    if (g.NA > 256 || g.NB < 16 || g.NB % 16 || g.N + g.Hp > 512) return false;
    if ((H / 4 + 1) / 2 > BF_MAXI || (g.NA / 16 + 1) / 2 > BF_MAXDEF) return false;
    g.RA = g.NA / 2; g.RB = g.NB / 2;
    g.KC = (D + 15) / 16; g.K8 = 2 * g.KC;
    g.whh_bytes = 2 * (g.RA + g.RB) * g.Hp * 2;
    g.wst_bytes = 2 * g.RA * 32;
    g.stage_bytes = 2 * BF_XBYTES + g.wst_bytes;
    const int fixed = g.whh_bytes + (int)align_up(g.N * 4, 128) + 1024;
    g.stages = (BF_SMEM_MAX - 1024 - fixed) / g.stage_bytes;
    if (g.stages > 12) g.stages = 12;
    g.smem = (int64_t)fixed + (int64_t)g.stages * g.stage_bytes;
    return g.stages >= 3;
}
static inline bool blstm_fused_supported(int H, int D) { BlstmGeom g; return blstm_fused_geom(H, D, g); }
static inline int64_t blstm_wih_img_elems(const BlstmGeom &g) { return 2LL * 2 * g.KC * 2 * (g.RA + g.RB) * 16; }
static inline int64_t blstm_whh_img_elems(const BlstmGeom &g) { return 2LL * 2 * 2 * (g.RA + g.RB) * g.Hp; }

struct BlstmFusedParams {
    const op16_t *whh_img;
    const float *bias;                   // [2][4H] pre-scaled, forget_bias folded into f
    const int32_t *lens;
    float *out; int ldo;                 // fp32 [n_windows*T][ldo], column dir*H + u, or null
    op16_t *next_hi, *next_lo;           // the next layer's x planes (K = 2H), or null
    int next_K8;
    int64_t n_windows;
    int n_pairs;                         // ceil(n_windows / 256)
    int T;
    BlstmGeom g;
};

// TF LSTMCell kernels [D+H, 4H] (column blocks i|j|f|o) + biases -> the three images above.
__global__ void k_pack_blstm_fused(const float *__restrict__ kernel_fw, const float *__restrict__ bias_fw,
                                   const float *__restrict__ kernel_bw, const float *__restrict__ bias_bw,
                                   float forget_bias, const BlstmGeom g, op16_t *__restrict__ wih_img,
                                   op16_t *__restrict__ whh_img, float *__restrict__ bias_out) {
    const int H = g.H, N = g.N;
    const int64_t n_ih = 2LL * N * g.KC * 16, n_hh = 2LL * N * g.Hp, n_b = 2LL * N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_ih + n_hh + n_b;
         i += (int64_t)gridDim.x * blockDim.x) {
        if (i >= n_ih + n_hh) {                                  // bias
            const int j = (int)(i - n_ih - n_hh), dir = j / N, n = j % N, u = n >> 2, gt = n & 3;
            const float sc = gt == 1 ? LSTM_PRESCALE_J : LSTM_PRESCALE_IFO;
            bias_out[j] = sc * ((dir ? bias_bw : bias_fw)[gt * H + u] + (gt == 2 ? forget_bias : 0.0f));
            continue;
        }
        const bool ih = i < n_ih;
        const int Kt = ih ? g.KC * 16 : g.Hp;                    // K extent of this operand
        const int64_t j = ih ? i : i - n_ih;
        const int dir = (int)(j / ((int64_t)N * Kt));
        const int64_t r = j - (int64_t)dir * N * Kt;
        const int n = (int)(r / Kt), k = (int)(r % Kt);
        const int u = n >> 2, gt = n & 3;
        const int half = n >= g.NA, nn = n - half * g.NA, R = half ? g.RB : g.RA;
        const int rank = nn >= R, row = nn - rank * R;
        const float *kern = dir ? kernel_bw : kernel_fw;
        const float sc = gt == 1 ? LSTM_PRESCALE_J : LSTM_PRESCALE_IFO;
        const int Kreal = ih ? g.D : H;
        float v = k < Kreal ? sc * kern[(int64_t)((ih ? 0 : g.D) + k) * N + gt * H + u] : 0.0f;
        // the recurrent operand has spare K rows when H is not a multiple of 16: row H carries the bias (forget_bias
        // folded), multiplied by a constant 1 the kernel keeps in the A operand's column of "unit H"
        if (!ih && k == H && g.Hp > H) v = sc * ((dir ? bias_bw : bias_fw)[gt * H + u] + (gt == 2 ? forget_bias : 0.0f));
        op16_t hv, lv;
        ptx::split_op16(v, hv, lv);
        const int RS = g.RA + g.RB;
        if (ih) {
            const int kc = k >> 4, k8 = (k >> 3) & 1, e = k & 7;
            const int64_t blk = (((int64_t)(dir * 2 + rank) * g.KC + kc) * 2) * RS * 16;      // plane 0 of this chunk
            const int64_t in = (half ? g.RA * 16 : 0) + ((int64_t)k8 * R + row) * 8 + e;
            wih_img[blk + in] = hv;
            wih_img[blk + (int64_t)RS * 16 + in] = lv;
        } else {
            const int k8 = k >> 3, e = k & 7;
            const int64_t per_plane = (int64_t)RS * g.Hp;
            const int64_t base = (int64_t)(dir * 2 + rank) * 2 * per_plane;
            const int64_t in = (half ? (int64_t)g.RA * g.Hp : 0) + ((int64_t)k8 * R + row) * 8 + e;
            whh_img[base + in] = hv;
            whh_img[base + per_plane + in] = lv;
        }
    }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(BF_THREADS, 1)
k_blstm_fused(const __grid_constant__ CUtensorMap map_x_hi, const __grid_constant__ CUtensorMap map_x_lo,
              const __grid_constant__ CUtensorMap map_w_a, const __grid_constant__ CUtensorMap map_w_b,
              const BlstmFusedParams p) {
    using namespace ptx;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    __shared__ uint64_t full_bar[12], empty_bar[12], d_full[2], drain_a, h_ready;
    __shared__ uint32_t tmem_base_smem;

    const BlstmGeom &g = p.g;
    const int H = g.H, Hp = g.Hp, T = p.T, N = g.N, NA = g.NA;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_cpairs = gridDim.x >> 1;
    const int n_items = 2 * p.n_pairs;                         // (tile pair, direction); this CTA pair takes item, item + n_cpairs, ..

    uint8_t *whh_s = smem_raw;                                 // [plane][half][Hp/8][R][8]
    float *bias_s = reinterpret_cast<float *>(smem_raw + g.whh_bytes);
    uint8_t *stage0 = smem_raw + g.whh_bytes + ((N * 4 + 127) & ~127);
    const uint32_t a_hi_col = (uint32_t)N, a_lo_col = (uint32_t)(N + Hp / 2);
    // the number of CTA pairs is even (host), so item = pair + k n_cpairs keeps its parity: a pair works on ONE direction
    const int dir = pair & 1;

    // ---- one-time setup: this CTA's W_hh rows (both planes) and the bias of its direction -> shared memory
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(p.whh_img) + (size_t)(dir * 2 + rank) * (g.whh_bytes / 16);
        uint4 *dst = reinterpret_cast<uint4 *>(whh_s);
        for (int i = threadIdx.x; i < g.whh_bytes / 16; i += BF_THREADS) dst[i] = __ldg(src + i);
        for (int i = threadIdx.x; i < N; i += BF_THREADS) bias_s[i] = p.bias[dir * N + i];
        fence_proxy_async_smem();          // generic-proxy stores -> visible to tcgen05.mma reads
    }
    if (threadIdx.x == 0) {
        prefetch_tmap(&map_x_hi); prefetch_tmap(&map_x_lo); prefetch_tmap(&map_w_a); prefetch_tmap(&map_w_b);
        for (int s = 0; s < g.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&d_full[0], 1); mbar_init(&d_full[1], 1);
        mbar_init(&drain_a, 16); mbar_init(&h_ready, 16);       // 8 epilogue warps of each CTA (leader's copies are used)
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_pair(&tmem_base_smem, 512);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    // shared-window addresses of the barriers and the stage ring, computed once (uniform registers in the loops below)
    const uint32_t full_u = smem_u32(&full_bar[0]), empty_u = smem_u32(&empty_bar[0]), dfull_u = smem_u32(&d_full[0]);
    const uint32_t drain_u = smem_u32(&drain_a), hready_u = smem_u32(&h_ready), stage_u = smem_u32(stage0);

    if (warp == 0) {
        // ===================== TMA producer (both CTAs: own x rows, own half of the weight rows) =====================
        // the whole warp walks the loop (warp-uniform counters), one elected lane issues
        int st = 0; uint32_t ph = 0;
        const int RS4 = (g.RA + g.RB) / 4;                     // 128-byte rows of one plane of a W_ih chunk
        const int wrow_base = (dir * 2 + rank) * g.KC * 2 * RS4;
        const uint32_t wlo_off = 2u * BF_XBYTES + (uint32_t)g.wst_bytes / 2;
        for (int item = pair; item < n_items; item += n_cpairs) {
            const int tile = 2 * (item >> 1) + rank;
            for (int s = 0; s < T; ++s) {
                const int t = dir ? T - 1 - s : s;
                const int xrow0 = (tile * T + t) * g.K8 * 16;
                for (int half = 0; half < 2; ++half) {
                    const uint32_t wb = (uint32_t)(half ? g.RB : g.RA) * 32;           // bytes per plane of this half's chunk
                    const CUtensorMap *mw = half ? &map_w_b : &map_w_a;
                    int wr = wrow_base + (half ? g.RA / 4 : 0);
                    for (int kc = 0; kc < g.KC; ++kc, wr += 2 * RS4) {
                        mbar_wait_u32(empty_u + 8u * st, ph ^ 1);
                        if (elect_one()) {
                            const uint32_t sp = stage_u + (uint32_t)st * g.stage_bytes, fb = full_u + 8u * st;
                            if (rank == 0) mbar_expect_tx_u32(fb, 2u * (2u * BF_XBYTES + 2u * wb));
                            tma_load_2d_pair_u32(sp, &map_x_hi, fb, 0, xrow0 + kc * 32);
                            tma_load_2d_pair_u32(sp + BF_XBYTES, &map_x_lo, fb, 0, xrow0 + kc * 32);
                            tma_load_2d_pair_u32(sp + 2 * BF_XBYTES, mw, fb, 0, wr);
                            tma_load_2d_pair_u32(sp + wlo_off, mw, fb, 0, wr + RS4);
                        }
                        __syncwarp();
                        if (++st == g.stages) { st = 0; ph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA; warp-uniform loop, one elected lane issues) =====================
        if (rank == 0) {
            const uint32_t idA = idesc_f16_f32(256, g.NA), idB = idesc_f16_f32(256, g.NB);
            const uint32_t whh_u = smem_u32(whh_s);
            const uint32_t plane_b = (uint32_t)(g.RA + g.RB) * Hp * 2;                  // bytes of one W_hh plane
            const int k16n = Hp / 16;
            const uint32_t xlbo = 2048u << 12;                                         // (LBO >> 4) << 16
            const uint32_t stage16 = (uint32_t)g.stage_bytes >> 4;
            int st = 0; uint32_t ph = 0, ph_drain = 0, ph_h = 0;
            for (int item = pair; item < n_items; item += n_cpairs) {
                for (int s = 0; s < T; ++s) {
                    for (int half = 0; half < 2; ++half) {
                        const uint32_t R = (uint32_t)(half ? g.RB : g.RA);
                        const uint32_t d = tmem_base + (uint32_t)(half ? NA : 0);
                        const uint32_t idesc = half ? idB : idA;
                        const uint32_t wlbo = R << 16;                                  // LBO = R * 16 bytes
                        if (half == 0) {
                            // the epilogue of the previous step has read the A half out of TMEM
                            mbar_wait_u32(drain_u, ph_drain);
                            ph_drain ^= 1;
                            tc_fence_after();
                        }
                        // ---- projection: D = x_t . W_ih[:, half]
                        for (int kc = 0; kc < g.KC; ++kc) {
                            mbar_wait_u32(full_u + 8u * st, ph);
                            tc_fence_after();
                            if (elect_one()) {
                                const uint32_t sp16 = ((stage_u & 0x3FFFFu) >> 4) + (uint32_t)st * stage16;
                                const uint32_t xh = sp16 | xlbo, xl = (sp16 + (BF_XBYTES >> 4)) | xlbo;
                                const uint32_t wh = (sp16 + (2 * BF_XBYTES >> 4)) | wlbo;
                                const uint32_t wl = (sp16 + (2 * BF_XBYTES >> 4) + ((uint32_t)g.wst_bytes >> 5)) | wlbo;
                                umma_f16_pair_w(d, xh, wh, DESC_NOSW_HI_SBO128, idesc, kc != 0);
                                umma_f16_pair_w(d, xl, wh, DESC_NOSW_HI_SBO128, idesc, 1);
                                umma_f16_pair_w(d, xh, wl, DESC_NOSW_HI_SBO128, idesc, 1);
                                umma_commit_pair_u32(empty_u + 8u * st, 3);   // stage free in both CTAs once these retire
                            }
                            __syncwarp();
                            if (++st == g.stages) { st = 0; ph ^= 1; }
                        }
                        if (half == 0) {
                            // h_{s-1} is complete in both CTAs' TMEM (and the B half has been read out)
                            mbar_wait_u32(hready_u, ph_h);
                            ph_h ^= 1;
                            tc_fence_after();
                        }
                        // ---- recurrent term: D += h_{s-1} . W_hh[:, half]
                        if (elect_one()) {
                            const uint32_t wb16 = ((whh_u & 0x3FFFFu) + (half ? (uint32_t)g.RA * Hp * 2 : 0u)) >> 4;
                            const uint32_t pl16 = plane_b >> 4, step16 = 2 * R;       // one k16 chunk = 2 K groups of R x 16 B
                            for (int j = 0; j < k16n; ++j) {
                                const uint32_t bh = (wb16 + (uint32_t)j * step16) | wlbo;
                                const uint32_t bl = (wb16 + pl16 + (uint32_t)j * step16) | wlbo;
                                const uint32_t ah = tmem_base + a_hi_col + (uint32_t)(8 * j);
                                const uint32_t al = tmem_base + a_lo_col + (uint32_t)(8 * j);
                                umma_f16_ts_pair_w(d, ah, bh, DESC_NOSW_HI_SBO128, idesc);
                                umma_f16_ts_pair_w(d, al, bh, DESC_NOSW_HI_SBO128, idesc);
                                umma_f16_ts_pair_w(d, ah, bl, DESC_NOSW_HI_SBO128, idesc);
                            }
                            umma_commit_pair_u32(dfull_u + 8u * half, 3);
                        }
                        __syncwarp();
                    }
                }
            }
        }
    } else if (warp >= 4) {
        // ===================== gate epilogue (both CTAs; thread = one window of the CTA's tile) =====================
        const int q = warp & 3, sub = (warp - 4) >> 2;
        const int r = q * 32 + lane;
        const int n_hc = H / 4, hc_a = NA / 16;                 // half-chunks in all / in N half A
        const int n_mine = (n_hc - sub + 1) / 2;                // this warp takes hc = 2 i + sub
        const int n_def = (hc_a - sub + 1) / 2;                 // .. of which the first n_def lie in half A
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t hi_w = lane_base + a_hi_col + (uint32_t)(2 * sub), lo_w = lane_base + a_lo_col + (uint32_t)(2 * sub);
        const bool bias_mma = Hp > H;                           // the bias rides in the recurrent MMA (see k_pack_blstm_fused)
        const uint32_t ready_u32 = smem_u32(&h_ready), drain_u32 = smem_u32(&drain_a);
        auto arrive = [&](uint64_t *bar) { if (rank == 0) mbar_arrive(bar); else mbar_arrive_cluster(bar, 0); };
        (void)ready_u32; (void)drain_u32;
        // store bases: a thread's 4 units of half-chunk hc = 2 i + sub sit at column dir H + 4 hc of the layer output;
        // in the next layer's planes that is group (dir H + 4 hc) / 8, element (dir H + 4 hc) % 8 -- with
        // m = sub + (dir H % 8) / 4 the offset is poff0 + i * 1024 elements: constant per warp, immediate per half-chunk
        const int cb = (dir * H) & 7, m = sub + (cb >> 2);
        const int64_t poff0 = ((int64_t)(((dir * H) >> 3) + (m >> 1))) * 1024 + (m & 1) * 4;
        // one-time: A-operand columns of the pad units (>= H) are zero for good, except unit H, a constant 1 whose W_hh
        // row holds the bias
        if (sub == 0) {
            for (int c4 = 0; c4 < Hp; c4 += 4) tmem_st_32x4(lane_base + a_hi_col + (uint32_t)c4, 0u, 0u, 0u, 0u);
            if (bias_mma) {
                const uint32_t one = (H & 1) ? 0x3C000000u : 0x00003C00u;       // fp16 1.0 in unit H's half of the column
                tmem_st_32x2(lane_base + a_hi_col + (uint32_t)(H >> 1), one, 0u);
            }
        }
        tmem_st_wait();
        tc_fence_before();
        asm volatile("bar.sync 1, 256;" ::: "memory");          // the 8 epilogue warps: sub 1 starts after sub 0's clears
        tc_fence_after();
        uint32_t ph_d = 0;
        bool first = true;
        for (int item = pair; item < n_items; item += n_cpairs) {
            const int tile = 2 * (item >> 1) + rank;
            const int64_t w = (int64_t)tile * 128 + r;
            const bool exists = w < p.n_windows;
            int len = exists ? p.lens[w] : 0;
            len = len > T ? T : (len < 0 ? 0 : len);
            float cst[BF_MAXI][4];
#pragma unroll
            for (int i = 0; i < BF_MAXI; ++i)
#pragma unroll
                for (int u = 0; u < 4; ++u) cst[i][u] = 0.0f;
            // h_0 = 0: every warp clears the columns it writes (the previous item's last R_B has retired: its d_full was
            // waited for), then tells the issuer that h is ready; at kernel start there is nothing to drain either
#pragma unroll
            for (int i = 0; i < BF_MAXI; ++i)
                if (i < n_mine) { tmem_st_32x2(hi_w + (uint32_t)(4 * i), 0u, 0u); tmem_st_32x2(lo_w + (uint32_t)(4 * i), 0u, 0u); }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (first) arrive(&drain_a);
                arrive(&h_ready);
            }
            first = false;
            float *orow = (p.out && exists) ? p.out + (w * T + (dir ? T - 1 : 0)) * (int64_t)p.ldo + dir * H + 4 * sub : nullptr;
            const int64_t ostep = dir ? -(int64_t)p.ldo : (int64_t)p.ldo;
            // next layer's planes: element offset of (tile, t, group, r, e) = (((tile T + t) K8n + group) 128 + r) 8 + e
            const int64_t pbase0 = (((int64_t)tile * T + (dir ? T - 1 : 0)) * p.next_K8 * 128 + r) * 8;
            const int64_t pstep = (dir ? -1 : 1) * (int64_t)p.next_K8 * 1024;
            op16_t *nh = p.next_hi ? p.next_hi + pbase0 + poff0 : nullptr;
            op16_t *nl = p.next_hi ? p.next_lo + pbase0 + poff0 : nullptr;
            const bool pad_writer = p.next_hi && dir == 1 && sub == ((n_hc - 1) & 1) && p.next_K8 > ((2 * H) >> 3);
            for (int s = 0; s < T; ++s) {
                const int t = dir ? T - 1 - s : s;
                const bool active = t < len;
                uint32_t dh[BF_MAXDEF][2], dl[BF_MAXDEF][2];
                mbar_wait(&d_full[0], ph_d, 500 + s);
                tc_fence_after();
                uint32_t acc[16];
                tmem_ld_32x16(lane_base + (uint32_t)(16 * sub), acc);
#pragma unroll
                for (int i = 0; i < BF_MAXI; ++i) {
                    if (i < n_mine) {
                        uint32_t a[16];
                        tmem_ld_wait_dep(acc);
#pragma unroll
                        for (int k = 0; k < 16; ++k) a[k] = acc[k];
                        if (i + 1 == n_def) {
                            // this was the last read of the A half: the next step's projection may overwrite it
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) arrive(&drain_a);
                        }
                        if (i + 1 < n_mine) {
                            if (i + 1 == n_def) {                // the next half-chunk is the first of the B half
                                mbar_wait(&d_full[1], ph_d, 600 + s);
                                tc_fence_after();
                            }
                            tmem_ld_32x16(lane_base + (uint32_t)(16 * sub + 32 * (i + 1)), acc);
                        }
                        if (i == n_def) {
                            // R_B(s) has retired: no MMA reads h_{s-1} any more -> the deferred h_s of the A half goes in
#pragma unroll
                            for (int k = 0; k < BF_MAXDEF; ++k)
                                if (k < n_def) {
                                    tmem_st_32x2(hi_w + (uint32_t)(4 * k), dh[k][0], dh[k][1]);
                                    tmem_st_32x2(lo_w + (uint32_t)(4 * k), dl[k][0], dl[k][1]);
                                }
                        }
                        float hv[4];
                        if (bias_mma) {
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                float cc = cst[i][u], hh;
                                lstm_cell_update(__uint_as_float(a[4 * u + 0]), __uint_as_float(a[4 * u + 1]),
                                                 __uint_as_float(a[4 * u + 2]), __uint_as_float(a[4 * u + 3]), cc, hh);
                                cst[i][u] = active ? cc : cst[i][u];
                                hv[u] = active ? hh : 0.0f;
                            }
                        } else {
                            const float4 *b4 = reinterpret_cast<const float4 *>(bias_s + 16 * (2 * i + sub));
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const float4 bb = b4[u];
                                float cc = cst[i][u], hh;
                                lstm_cell_update(__uint_as_float(a[4 * u + 0]) + bb.x, __uint_as_float(a[4 * u + 1]) + bb.y,
                                                 __uint_as_float(a[4 * u + 2]) + bb.z, __uint_as_float(a[4 * u + 3]) + bb.w, cc, hh);
                                cst[i][u] = active ? cc : cst[i][u];
                                hv[u] = active ? hh : 0.0f;
                            }
                        }
                        uint32_t ph2[2], pl2[2];
                        split_pack_op16x2(hv[0], hv[1], ph2[0], pl2[0]);
                        split_pack_op16x2(hv[2], hv[3], ph2[1], pl2[1]);
                        if (i < BF_MAXDEF && i < n_def) {
                            dh[i][0] = ph2[0]; dh[i][1] = ph2[1]; dl[i][0] = pl2[0]; dl[i][1] = pl2[1];
                        } else {
                            tmem_st_32x2(hi_w + (uint32_t)(4 * i), ph2[0], ph2[1]);
                            tmem_st_32x2(lo_w + (uint32_t)(4 * i), pl2[0], pl2[1]);
                        }
                        if (nh) {                // rows of windows past the end are written too (zeros: len = 0)
                            *reinterpret_cast<uint2 *>(nh + i * 1024) = make_uint2(ph2[0], ph2[1]);
                            *reinterpret_cast<uint2 *>(nl + i * 1024) = make_uint2(pl2[0], pl2[1]);
                        }
                        if (orow) *reinterpret_cast<float4 *>(orow + 8 * i) = make_float4(hv[0], hv[1], hv[2], hv[3]);
                    }
                }
                if (n_mine == n_def) {
                    // (H so small that this warp has no half-chunk in the B half) still has to order its h stores behind R_B
                    mbar_wait(&d_full[1], ph_d, 700 + s);
                    tc_fence_after();
#pragma unroll
                    for (int k = 0; k < BF_MAXDEF; ++k)
                        if (k < n_def) {
                            tmem_st_32x2(hi_w + (uint32_t)(4 * k), dh[k][0], dh[k][1]);
                            tmem_st_32x2(lo_w + (uint32_t)(4 * k), dl[k][0], dl[k][1]);
                        }
                }
                if (pad_writer) {                // zero the K pad groups of the next layer's operand (once per row and step)
                    op16_t *zh = nh - poff0, *zl = nl - poff0;
                    for (int gz = (2 * H) >> 3; gz < p.next_K8; ++gz) {
                        *reinterpret_cast<uint4 *>(zh + (int64_t)gz * 1024) = make_uint4(0, 0, 0, 0);
                        *reinterpret_cast<uint4 *>(zl + (int64_t)gz * 1024) = make_uint4(0, 0, 0, 0);
                    }
                }
                ph_d ^= 1;
                if (orow) orow += ostep;
                if (nh) { nh += pstep; nl += pstep; }
                if (s + 1 < T) {                 // after the last step the next item's reset (above) signals instead
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) arrive(&h_ready);
                }
            }
        }
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, 512);
    }
}

}  // namespace ntb
