// K2/K3  Dense contractions of the NanoTensor forward on the 5th-gen tensor cores:
//   conv_layer (network.py:312-329: 1x1 / 1x3 SAME convs as GEMMs over all rows = window*T + t),
//   the LSTM input projection x_t . kernel[0:D, :] (network.py:275-282) and fulconn_layer
//   (network.py:291-309) when the class count is large.
//
// The reference computes these in float32.  single 16-bit tensor-core operands alone miss the 1e-3 logits
// budget by two orders of magnitude (precision table), so every fp32 operand is split
// into fp16 hi + fp16 lo planes and the product is accumulated in fp32 TMEM as
//   A.B ~= Ahi.Bhi + Alo.Bhi + Ahi.Blo            (relative error ~2^-16, "fp16x3").
//
// Kernel: persistent, warp-specialised, one CTA per SM.
//   warp 0      TMA producer: cp.async.bulk.tensor loads of the A hi/lo [128 x 64] and
//               B hi/lo [BN x 64] tiles (128-byte swizzle) into a ring of smem stages
//   warp 1      MMA issuer: one thread issues tcgen05.mma (M=128, N=BN, K=16) x 3 passes per
//               k-step into one of two TMEM accumulators; tcgen05.commit frees the smem stage /
//               publishes the accumulator
//   warp 2      TMEM allocator (512 columns = 2 accumulators of up to 256 columns)
//   warps 4-11  epilogue: tcgen05.ld 32 lanes x 32 columns -> registers -> (+bias) -> global fp32,
//               overlapping the next tile's main loop through the second accumulator; two warps per TMEM lane
//               quarter take alternate 32-column chunks (with four warps the epilogue paced the conv GEMMs)
// 1x3 convs run as 3 taps accumulated into the same accumulator: the A planes use a row layout
// padded per window ([window][T+2][K], zero rows at both ends) so a tap is a row offset of the
// TMA coordinate and SAME zero padding falls out of the layout.
#include <cuda.h>
#include <stdlib.h>

#include "ntb_gemm_tc.cuh"
#include "ntb_lstm_cell.cuh"
#include "ntb_tc_ptx.cuh"

namespace ntb {

namespace {

constexpr int BM = 128;        // rows per tile (TMEM lanes)
constexpr int BK = 64;         // fp16 per smem row = one 128-byte swizzle span
constexpr int TC_THREADS_EW8 = 384;     // 4 control warps + 8 epilogue warps (two per TMEM lane quarter, alternate column chunks)
constexpr int TMEM_COLS = 512;
constexpr int A_TILE_BYTES = BM * BK * 2;     // 16 KB
constexpr int EPI_STAGE_BYTES = 4 * 2 * 4096;   // 4 epilogue warps x 2 buffers x [32 rows][32 fp32], TMA-store staging
constexpr int SMEM_BUDGET = 227 * 1024 - 1024 - EPI_STAGE_BYTES - 2048;   // operand ring

struct TcParams {
    int64_t m_tiles;       // ceil(plane_rows / 128)
    int n_tiles, BN, Np, N;
    int taps, k_blocks;    // k_blocks per tap = ceil(K / 64)
    int k16_last;          // k16 steps in the last k block of a tap
    int passes;            // 3 = fp16x3, 2 = interleaved-A form (one A plane, B' and B''), 1 = single pass (ablation only)
    int stages;
    int T, pad;            // pad = 1: plane rows are [window][T+2]
    int64_t out_rows;      // rows of `out`
    const float *bias;
    float *out;
    int ldo;
    int tma_store;         // plain row-major output through smem staging + cp.async.bulk.tensor stores
    int xproj_H;           // > 0: out is the recurrence's unit-major projected-gates layout (ntb_lstm_tc.cuh)
    int xproj_tma;         // projected-gates epilogue through smem staging + 3-D TMA stores (map_out = store map)
    int l2pf;              // opt-in experiment: the producer prefetches the NEXT work item's A tiles into L2 (measured slower)
    int abl;               // timing ablation (NTB_XPROJ_ABL): 1 = projected-gates epilogue without its global stores
    int epi_bytes;         // bytes reserved behind the operand ring (TMA-store staging, or just the bias vector)
    int bstat;             // B-stationary schedule (CG = 2, several N tiles, small K): every CTA pair keeps ONE weight
                           // tile (all k-blocks, hi + lo) resident in shared memory and streams M tiles through it;
                           // only A moves through the ring -- the operand feed from L2 drops from ~46 to ~26 B/clk/SM
    // fused batch-norm statistics (bn_part != null): every epilogue warp reduces its 32 rows x 32 columns to per-column
    // (sum, sum of squares), split at the window boundary inside the block (T >= 32: at most one), and
    // writes bn_part[(block * 2 + seg) * N + n]; k_bn_finish folds the partials of each BN group (ntb_kernels.cuh)
    float2 *bn_part;
    // dual output (n_split > 0): two GEMMs of the SAME A operand in one launch (conv1 | conv2a of residual_layer, weights
    // concatenated along N): N tiles >= n_split go to map_out2 / bn_part2 with column origin n_split * BN, and the BN
    // partial records of both halves are indexed with N = n_half channels.  A is read from DRAM once instead of twice.
    int n_split, n_half;
    float2 *bn_part2;
    // LSTM timestep mode (lstm_H > 0): this GEMM is h_{t-1} . W_hh of ONE timestep for all windows and both
    // directions (A rows = [dir][Mp windows], B rows = [dir][Np]); the epilogue adds the projected gates and runs
    // the fused cell update (see gemm_tc_lstm_layer)
    int lstm_H, lstm_s, lstm_mt_dir;   // hidden size, step index, M tiles per direction
    int64_t lstm_n, lstm_Mp;           // windows, windows padded to the M-tile pair
    const float *lstm_x; int lstm_ldx; // projected gates [n*T][ldx], column dir*4H + u*4 + g (bias folded, pre-scaled)
    float *lstm_c;                     // cell state [2 dirs][128-window tile][H/4][128][4]: a warp's float4 accesses are contiguous
    const int32_t *lstm_lens;
    op16_t *lstm_hi, *lstm_lo;         // h_t operand planes of the NEXT step, k-group-major [H/8][2 Mp rows][8] (a_kmajor)
    int64_t lstm_xmt;                  // M tiles per direction of the projected-gates buffer = 128-window tiles x T
    // lstm_kbx > 0: the input projection rides in the SAME GEMM -- the first lstm_kbx k-blocks of every timestep's
    // contraction are x_t . W_ih (A = the layer input's time-major operand planes through map_out / map_out2, one
    // [128 windows x 64] box per tile and k-block at row (tile T + t) 128), the rest h_{t-1} . W_hh; B = [W_ih ; W_hh]
    // along K.  No projected gates exist then (lstm_x == null): the epilogue adds lstm_bias [2][4H] instead.
    int lstm_kbx;
    const float *lstm_bias;
    // lstm_nx_hi != null: h_t goes out as the NEXT layer's time-major operand planes ([(tile T + t) 128 + window][lstm_nx_K],
    // column dir H + unit) instead of the fp32 layer output: no split pass between two stacked layers
    op16_t *lstm_nx_hi, *lstm_nx_lo;
    int lstm_nx_K;
    // amax != null: no output at all -- the epilogue keeps, per row, the best (value, column) of the tile's columns (+ bias) and
    // merges it into amax[row] with one atomicMax per warp and tile: key = orderable(value) << 32 | ~column, so the maximum
    // key is the largest value and, among equal values, the FIRST column (tf.argmax); the dense readout of a large class
    // count (fulconn_layer, network.py:291-309 + tf.argmax :208-211) then never materialises its logits
    unsigned long long *amax;
    int a_kmajor;                      // the A planes are [K/8][rows][8] (no swizzle; 3-D TMA box {8, 128, 8}) instead of
                                       // row-major rows with the 128-byte swizzle: the LSTM epilogue's h stores coalesce
};

// Column statistics of one staged [32 rows][32 cols] fp32 tile (128-byte swizzled rows): lane = column; rows
// [0, split) belong to the first window of the block (segment 0), [split, 32) to the next (segment 1); rows whose
// bit in vmask is clear (padding rows of the 1x3 layout, rows past the end) are skipped.
// A record is (S, M2) = (sum, sum of squared deviations from the SEGMENT's own mean) of <= 32 values, accumulated around a
// pivot (the column's value in row 0 of the tile): S = n p + sum(v - p), M2 = sum((v - p)^2) - sum(v - p)^2 / n.  The
// reference's tf.nn.moments is two-pass (network.py:339-349); sum / sum-of-squares records in fp32 lose
// mean^2 / var * 6e-8 of the variance, which a channel with |mean| ~ 100 std turns into 1e-3 -- centred records do not,
// and k_bn_finish merges them exactly (Chan et al.) in fp64.
__device__ __forceinline__ void bn_partials(const uint8_t *buf, int lane, uint32_t vmask, int split, float2 *part,
                                            int64_t rec0, int N, int n) {
    // fp32, fixed order -> bit-stable
    const int cj = lane >> 2, ci = (lane & 3) << 2;
    const float pv = *reinterpret_cast<const float *>(buf + (cj << 4) + ci);       // row 0
    float n0 = 32.0f, n1 = 0.0f;
    float sa, qa, sb = 0.0f, qb = 0.0f;
    if (vmask == 0xffffffffu && split == 32) {
        // common case (no window boundary, no padding row in the block): rows in pairs through the packed f32x2 pipe
        // (three instructions per two values), two interleaved chains of pairs
        float2 s2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)}, q2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        const float2 np2 = make_float2(-pv, -pv);
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
            const float2 v = make_float2(*reinterpret_cast<const float *>(buf + i * 128 + ((cj ^ (i & 7)) << 4) + ci),
                                         *reinterpret_cast<const float *>(buf + (i + 1) * 128 + ((cj ^ ((i + 1) & 7)) << 4) + ci));
            const float2 d = __fadd2_rn(v, np2);
            s2[(i >> 1) & 1] = __fadd2_rn(s2[(i >> 1) & 1], d);
            q2[(i >> 1) & 1] = __ffma2_rn(d, d, q2[(i >> 1) & 1]);
        }
        sa = (s2[0].x + s2[0].y) + (s2[1].x + s2[1].y);
        qa = (q2[0].x + q2[0].y) + (q2[1].x + q2[1].y);
    } else if (vmask == 0xffffffffu) {
        // a window boundary inside the block (with T = 100 every third block): the same packed loop over ALL rows plus, for
        // the pairs in front of the boundary (warp-uniform predicate), segment 0 on its own; segment 1 = all - segment 0
        // (sums of deviations from the pivot: the subtraction costs an ulp of a small number)
        float2 s2 = make_float2(0.f, 0.f), q2 = make_float2(0.f, 0.f), t2 = make_float2(0.f, 0.f), u2 = make_float2(0.f, 0.f);
        const float2 np2 = make_float2(-pv, -pv);
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
            const float2 v = make_float2(*reinterpret_cast<const float *>(buf + i * 128 + ((cj ^ (i & 7)) << 4) + ci),
                                         *reinterpret_cast<const float *>(buf + (i + 1) * 128 + ((cj ^ ((i + 1) & 7)) << 4) + ci));
            const float2 d = __fadd2_rn(v, np2);
            s2 = __fadd2_rn(s2, d);
            q2 = __ffma2_rn(d, d, q2);
            if (i < split) {
                const float2 d0 = make_float2(d.x, i + 1 < split ? d.y : 0.0f);
                t2 = __fadd2_rn(t2, d0);
                u2 = __ffma2_rn(d0, d0, u2);
            }
        }
        n0 = (float)split;
        n1 = (float)(32 - split);
        sa = t2.x + t2.y; qa = u2.x + u2.y;
        sb = (s2.x + s2.y) - sa; qb = fmaxf((q2.x + q2.y) - qa, 0.0f);
    } else {
        // padded plane rows / rows past the end in the block (the layouts without the fused operand transform): per-row
        // selects, four interleaved chains per segment (row i -> chain i & 3)
        float s0[4] = {0, 0, 0, 0}, q0[4] = {0, 0, 0, 0}, s1[4] = {0, 0, 0, 0}, q1[4] = {0, 0, 0, 0};
        const uint32_t low = split >= 32 ? 0xffffffffu : ((1u << split) - 1u);
        n0 = (float)__popc(vmask & low);
        n1 = (float)__popc(vmask & ~low);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const float f = *reinterpret_cast<const float *>(buf + i * 128 + ((cj ^ (i & 7)) << 4) + ci) - pv;
            const float v = ((vmask >> i) & 1u) ? f : 0.0f;
            if (i < split) {
                s0[i & 3] += v;
                q0[i & 3] = fmaf(v, v, q0[i & 3]);
            } else {
                s1[i & 3] += v;
                q1[i & 3] = fmaf(v, v, q1[i & 3]);
            }
        }
        sa = (s0[0] + s0[1]) + (s0[2] + s0[3]); qa = (q0[0] + q0[1]) + (q0[2] + q0[3]);
        sb = (s1[0] + s1[1]) + (s1[2] + s1[3]); qb = (q1[0] + q1[1]) + (q1[2] + q1[3]);
    }
    if (n < N) {
        part[rec0 * N + n] = make_float2(fmaf(n0, pv, sa), n0 > 0.0f ? fmaxf(qa - sa * sa / n0, 0.0f) : 0.0f);
        if (split < 32)
            part[(rec0 + 1) * N + n] = make_float2(fmaf(n1, pv, sb), n1 > 0.0f ? fmaxf(qb - sb * sb / n1, 0.0f) : 0.0f);
    }
}

// CG = 1: one CTA per tile (M = 128).  CG = 2: CTA pair (cluster of 2, tcgen05 cta_group::2, M = 256): each CTA loads
// its own 128 A rows and HALF of the B tile, the leader issues the MMAs for both SMs, so the B operand
// traffic per output halves -- these GEMMs are operand-bandwidth bound (hi + lo planes of both operands).
// LSTM = true: the per-timestep mode of gemm_tc_lstm_layer (its own instantiation, so that the cell-update epilogue and its
// prefetch registers do not weigh on the conv / projection GEMMs)
template <int CG, bool LSTM>
__global__ void __launch_bounds__(TC_THREADS_EW8, 1)
k_gemm_tc(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
          const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo,
          const __grid_constant__ CUtensorMap map_out, const __grid_constant__ CUtensorMap map_out2, const TcParams p) {
    using namespace ptx;
    constexpr int EW = 8;        // epilogue warps: the accumulator's column chunks are dealt round-robin to EW / 4 warps per lane quarter
    const int lstm_kbx = LSTM ? p.lstm_kbx : 0;
    const bool a_kmajor = LSTM && p.a_kmajor;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[8], empty_bar[8], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base_smem;

    const uint32_t raw = smem_u32(smem_raw);
    uint8_t *smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);     // 1024 B: swizzle-128B atoms
    const int b_tile_bytes = (p.BN / CG) * BK * 2;               // this CTA's share of the B tile
    const bool bstat = CG == 2 && p.bstat;
    const int stage_bytes = bstat ? 2 * A_TILE_BYTES
                          : p.passes == 2 ? A_TILE_BYTES + 2 * b_tile_bytes : (p.passes == 3 ? 2 : 1) * (A_TILE_BYTES + b_tile_bytes);
    const int bres_bytes = bstat ? p.k_blocks * 2 * b_tile_bytes : 0;   // resident B: [k block][hi | lo]
    uint8_t *stages_base = smem + bres_bytes;
    __shared__ uint64_t b_full;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = CG == 2 ? (int)cluster_ctarank() : 0;      // 0 = leader of the pair
    const int64_t m_groups = (p.m_tiles + CG - 1) / CG;         // a work item = CG consecutive M tiles x one N tile
    const int64_t total_tiles = m_groups * p.n_tiles;
    const int64_t P = blockIdx.x / CG, NP = gridDim.x / CG;
    const int kb_total = p.taps * p.k_blocks;
    // work item `it` of this CTA (pair) -> (M group, N tile).  B-stationary: the N tile is fixed, M groups are strided
    auto item = [&](int64_t it, int64_t &mg, int &nt) -> bool {
        if (bstat) {
            nt = (int)(P % p.n_tiles);
            mg = P / p.n_tiles + it * (NP / p.n_tiles);
            return mg < m_groups;
        }
        if (p.n_split > 0) {
            // dual output: one pair takes BOTH halves of an M group back to back, so the second pass over the A tiles is an
            // L2 hit (with the N tiles spread over neighbouring pairs ncu showed the A planes read from DRAM 1.84 times)
            nt = (int)(it % p.n_tiles);
            mg = P + (it / p.n_tiles) * NP;
            return mg < m_groups;
        }
        const int64_t tile = P + it * NP;
        nt = (int)(tile % p.n_tiles);
        mg = tile / p.n_tiles;
        return tile < total_tiles;
    };

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&map_a_hi); prefetch_tmap(&map_a_lo); prefetch_tmap(&map_b_hi); prefetch_tmap(&map_b_lo);
        if (p.tma_store || p.xproj_tma || lstm_kbx) prefetch_tmap(&map_out);
        if (p.n_split > 0 || lstm_kbx) prefetch_tmap(&map_out2);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], EW * CG); }
        mbar_init(&b_full, 1);
        fence_barrier_init();
    }
    if (warp == 2) {
        if (CG == 2) tmem_alloc_pair(&tmem_base_smem, TMEM_COLS); else tmem_alloc(&tmem_base_smem, TMEM_COLS);
    }
    // xproj mode: the (unused) TMA-store staging area holds the bias vector, so the epilogue's bias reads
    // are shared-memory broadcasts instead of global loads queued behind its own stores
    float *bias_s = reinterpret_cast<float *>(stages_base + (size_t)p.stages * stage_bytes + (p.xproj_tma ? EPI_STAGE_BYTES : 0));
    const bool bias_in_smem = p.xproj_H > 0 && p.bias && p.Np * 4 <= p.epi_bytes - (p.xproj_tma ? EPI_STAGE_BYTES : 0);
    if (bias_in_smem)
        for (int i = threadIdx.x; i < p.Np; i += (int)blockDim.x) bias_s[i] = i < p.N ? p.bias[i] : 0.0f;
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();     // peer barriers are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            const int left = (p.taps - 1) / 2;
            int64_t mg; int n_tile;
            if (bstat && item(0, mg, n_tile)) {
                // the resident weight tile of this pair: every k-block, hi and lo, each CTA its half of the rows
                if (rank == 0) mbar_expect_tx(&b_full, (uint32_t)(2 * bres_bytes));
                const int brow0 = n_tile * p.BN + rank * (p.BN / CG);
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    tma_load_2d_pair(smem + (size_t)(2 * kb) * b_tile_bytes, &map_b_hi, &b_full, kb * BK, brow0);
                    tma_load_2d_pair(smem + (size_t)(2 * kb + 1) * b_tile_bytes, &map_b_lo, &b_full, kb * BK, brow0);
                }
            }
            for (int64_t it = 0; item(it, mg, n_tile); ++it) {
                const int64_t m_tile = mg * CG + rank;
                const int dir_rows = (LSTM && m_tile >= p.lstm_mt_dir) ? p.Np : 0;   // W_hh of the bw direction
                if (p.l2pf) {
                    int64_t mg2; int nt2;
                    if (item(it + 1, mg2, nt2) && mg2 != mg) {
                        const int r2 = (int)((mg2 * CG + rank) * BM) - left;
                        for (int kb = 0; kb < p.k_blocks; ++kb) {
                            tma_prefetch_2d(&map_a_hi, kb * BK, r2);
                            if (p.passes == 3) tma_prefetch_2d(&map_a_lo, kb * BK, r2);
                        }
                    }
                }
                for (int tap = 0; tap < p.taps; ++tap) {
                    const int row0 = (int)(m_tile * BM) + tap - left;
                    const int brow0 = tap * p.Np + n_tile * p.BN + rank * (p.BN / CG) + dir_rows;
                    for (int kb = 0; kb < p.k_blocks; ++kb) {
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        uint8_t *s = stages_base + (size_t)stage * stage_bytes;
                        if (bstat) {
                            if (rank == 0) mbar_expect_tx(&full_bar[stage], (uint32_t)(2 * stage_bytes));
                            tma_load_2d_pair(s, &map_a_hi, &full_bar[stage], kb * BK, row0);
                            tma_load_2d_pair(s + A_TILE_BYTES, &map_a_lo, &full_bar[stage], kb * BK, row0);
                        } else if (CG == 1) {
                            mbar_expect_tx(&full_bar[stage], (uint32_t)stage_bytes);
                            tma_load_2d(s, &map_a_hi, &full_bar[stage], kb * BK, row0);
                            tma_load_2d(s + A_TILE_BYTES, &map_b_hi, &full_bar[stage], kb * BK, brow0);
                            if (p.passes == 3) {
                                tma_load_2d(s + A_TILE_BYTES + b_tile_bytes, &map_a_lo, &full_bar[stage], kb * BK, row0);
                                tma_load_2d(s + 2 * A_TILE_BYTES + b_tile_bytes, &map_b_lo, &full_bar[stage], kb * BK, brow0);
                            } else if (p.passes == 2) {
                                tma_load_2d(s + A_TILE_BYTES + b_tile_bytes, &map_b_lo, &full_bar[stage], kb * BK, brow0);
                            }
                        } else {
                            // both CTAs' bytes are credited to the LEADER's full barrier
                            if (rank == 0) mbar_expect_tx(&full_bar[stage], (uint32_t)(2 * stage_bytes));
                            const bool xk = lstm_kbx > 0 && kb < lstm_kbx;      // an x_t k-block of the fused step GEMM
                            int xrow = 0;
                            if (xk) {
                                const int dirx = m_tile >= p.lstm_mt_dir ? 1 : 0;
                                const int64_t tl = m_tile - (int64_t)dirx * p.lstm_mt_dir;
                                xrow = (int)((tl * p.T + (dirx ? p.T - 1 - p.lstm_s : p.lstm_s)) * BM);
                            }
                            if (xk) tma_load_2d_pair(s, &map_out, &full_bar[stage], kb * BK, xrow);
                            else if (a_kmajor) tma_load_3d_pair(s, &map_a_hi, &full_bar[stage], 0, row0, (kb - lstm_kbx) * (BK / 8));
                            else tma_load_2d_pair(s, &map_a_hi, &full_bar[stage], kb * BK, row0);
                            tma_load_2d_pair(s + A_TILE_BYTES, &map_b_hi, &full_bar[stage], kb * BK, brow0);
                            if (p.passes == 3) {
                                if (xk) tma_load_2d_pair(s + A_TILE_BYTES + b_tile_bytes, &map_out2, &full_bar[stage], kb * BK, xrow);
                                else if (a_kmajor) tma_load_3d_pair(s + A_TILE_BYTES + b_tile_bytes, &map_a_lo, &full_bar[stage], 0, row0, (kb - lstm_kbx) * (BK / 8));
                                else
                                tma_load_2d_pair(s + A_TILE_BYTES + b_tile_bytes, &map_a_lo, &full_bar[stage], kb * BK, row0);
                                tma_load_2d_pair(s + 2 * A_TILE_BYTES + b_tile_bytes, &map_b_lo, &full_bar[stage], kb * BK, brow0);
                            } else if (p.passes == 2) {
                                tma_load_2d_pair(s + A_TILE_BYTES + b_tile_bytes, &map_b_lo, &full_bar[stage], kb * BK, brow0);
                            }
                        }
                        if (++stage == p.stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0 && rank == 0) {
            const uint32_t idesc = idesc_f16_f32(BM * CG, p.BN);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            int64_t mg; int n_tile;
            if (bstat && item(0, mg, n_tile)) {
                mbar_wait(&b_full, 0);
                tc_fence_after();
            }
            for (int64_t it = 0; item(it, mg, n_tile); ++it) {
                mbar_wait(&acc_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
                for (int kb = 0; kb < kb_total; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t s = smem_u32(stages_base + (size_t)stage * stage_bytes);
                    const uint32_t sb = smem_u32(smem + (size_t)(2 * kb) * b_tile_bytes);   // resident B (bstat)
                    const uint64_t a_hi = smem_desc_sw128(s);
                    const uint64_t b_hi = smem_desc_sw128(bstat ? sb : s + A_TILE_BYTES);
                    const uint64_t a_lo = smem_desc_sw128(bstat ? s + A_TILE_BYTES : s + A_TILE_BYTES + b_tile_bytes);
                    const uint64_t b_lo = smem_desc_sw128(bstat ? sb + b_tile_bytes : s + 2 * A_TILE_BYTES + b_tile_bytes);
                    const int k16n = ((kb % p.k_blocks) == p.k_blocks - 1) ? p.k16_last : BK / 16;
                    for (int k = 0; k < k16n; ++k) {
                        if (CG == 1) {
                            umma_f16(d_tmem, desc_advance_k16(a_hi, k), desc_advance_k16(b_hi, k), idesc, (kb | k) != 0);
                            if (p.passes == 3) {
                                umma_f16(d_tmem, desc_advance_k16(a_lo, k), desc_advance_k16(b_hi, k), idesc, 1);
                                umma_f16(d_tmem, desc_advance_k16(a_hi, k), desc_advance_k16(b_lo, k), idesc, 1);
                            } else if (p.passes == 2) {     // a_lo's slot holds B'' here
                                umma_f16(d_tmem, desc_advance_k16(a_hi, k), desc_advance_k16(a_lo, k), idesc, 1);
                            }
                        } else if (a_kmajor && kb >= lstm_kbx) {
                            // A tile = [8 K groups][128 rows][8]: k16 step k starts at group 2k; LBO = one group (2 KB), SBO = 128 B
                            const uint64_t ah = smem_desc_nosw(s + (uint32_t)k * 4096u, 2048, 128);
                            const uint64_t al = smem_desc_nosw(s + A_TILE_BYTES + b_tile_bytes + (uint32_t)k * 4096u, 2048, 128);
                            umma_f16_pair(d_tmem, ah, desc_advance_k16(b_hi, k), idesc, (kb | k) != 0);
                            umma_f16_pair(d_tmem, al, desc_advance_k16(b_hi, k), idesc, 1);
                            umma_f16_pair(d_tmem, ah, desc_advance_k16(b_lo, k), idesc, 1);
                        } else {
                            umma_f16_pair(d_tmem, desc_advance_k16(a_hi, k), desc_advance_k16(b_hi, k), idesc, (kb | k) != 0);
                            if (p.passes == 3) {
                                umma_f16_pair(d_tmem, desc_advance_k16(a_lo, k), desc_advance_k16(b_hi, k), idesc, 1);
                                umma_f16_pair(d_tmem, desc_advance_k16(a_hi, k), desc_advance_k16(b_lo, k), idesc, 1);
                            } else if (p.passes == 2) {     // a_lo's slot holds B'' here
                                umma_f16_pair(d_tmem, desc_advance_k16(a_hi, k), desc_advance_k16(a_lo, k), idesc, 1);
                            }
                        }
                    }
                    // smem stage reusable (in both CTAs) once these MMAs retire
                    if (CG == 1) umma_commit(&empty_bar[stage]); else umma_commit_pair(&empty_bar[stage], 3);
                    if (++stage == p.stages) { stage = 0; phase ^= 1; }
                }
                if (CG == 1) umma_commit(&acc_full[acc]); else umma_commit_pair(&acc_full[acc], 3);   // accumulator complete
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int q = warp & 3;                                // TMEM lane quarter this warp may access
        int acc = 0; uint32_t acc_phase = 0;
        const bool vec_ok = ((p.ldo & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
        // staging: 4 warps -> two 4 KB buffers each (the TMA store of one chunk overlaps the next chunk), 8 warps -> one each
        const int eh = (warp - 4) >> 2;                         // which of the EW / 4 warps of this lane quarter
        uint8_t *stage_buf = stages_base + (size_t)p.stages * stage_bytes + (EW == 4 ? (size_t)q * 8192 : (size_t)(warp - 4) * 4096);   // 1024-B aligned
        int stage_it = 0;
        int64_t mg; int n_tile;
        for (int64_t it = 0; item(it, mg, n_tile); ++it) {
            const int64_t m_tile = mg * CG + rank;
            const int64_t prow = m_tile * BM + q * 32 + lane;       // row in the A-plane layout
            // LSTM timestep mode: what does not depend on the column chunk, and the first chunk's projected gates / cell
            // state requested BEFORE the wait for the accumulator (four epilogue warps have nothing else to hide a DRAM
            // latency behind; every later chunk is requested one chunk ahead, see below)
            int l_dir = 0, l_t = 0, l_wl = 0;
            int64_t l_w = 0, l_tile = 0;
            bool l_live = false, l_active = false;
            float4 xpre[8], cpre[2];
            const float4 *l_x = nullptr;
            float4 *l_c = nullptr;
            auto lstm_fetch = [&](int c0f) {       // projected gates + cell state of the chunk at column c0f of this N tile
                const int nbf = n_tile * p.BN + c0f;
                if (!l_live || nbf >= p.N) return;
                const int nuf = (p.N - nbf >= 32 && c0f + 32 <= p.BN) ? 8 : 4, u0f = nbf >> 2;
                const float4 *cgp = l_c + (int64_t)(u0f >> 2) * BM;
                if (p.lstm_x) {
                    const float4 *xg = l_x + (int64_t)u0f * BM;
#pragma unroll
                    for (int k = 0; k < 8; ++k) if (k < nuf) xpre[k] = __ldg(xg + k * BM);
                } else {                           // projection inside the GEMM: only the bias is left to add
                    const float4 *bg = reinterpret_cast<const float4 *>(p.lstm_bias) + ((l_dir * p.N) >> 2) + u0f;
#pragma unroll
                    for (int k = 0; k < 8; ++k) if (k < nuf) xpre[k] = __ldg(bg + k);
                }
                cpre[0] = cgp[0];
                if (nuf == 8) cpre[1] = cgp[BM];
            };
            if (LSTM) {
                l_dir = m_tile >= p.lstm_mt_dir ? 1 : 0;
                l_w = prow - l_dir * p.lstm_Mp;
                l_live = l_w < p.lstm_n;
                if (l_live) {
                    l_t = l_dir ? p.T - 1 - p.lstm_s : p.lstm_s;
                    int len = p.lstm_lens[l_w];
                    len = len > p.T ? p.T : (len < 0 ? 0 : len);
                    l_active = l_t < len;
                    l_tile = l_w >> 7; l_wl = (int)(l_w & 127);
                    const int64_t n_tiles = p.lstm_xmt / p.T;          // 128-window tiles of the job (not of the padded M)
                    if (p.lstm_x)
                        l_x = reinterpret_cast<const float4 *>(p.lstm_x) + ((l_dir * p.lstm_xmt + l_tile * p.T + l_t) * p.lstm_H) * (int64_t)BM + l_wl;
                    l_c = reinterpret_cast<float4 *>(p.lstm_c) + ((l_dir * n_tiles + l_tile) * (p.lstm_H >> 2)) * (int64_t)BM + l_wl;
                }
                lstm_fetch(((warp - 4) >> 2) * 32);
            }
            mbar_wait(&acc_full[acc], acc_phase);
            tc_fence_after();
            int64_t orow = prow;
            bool row_ok = true;
            if (p.pad) {
                const int64_t w = prow / (p.T + 2);
                const int tt = (int)(prow - w * (p.T + 2)) - 1;
                row_ok = tt >= 0 && tt < p.T;
                orow = w * p.T + tt;
            }
            row_ok = row_ok && orow < p.out_rows;
            float *orow_ptr = p.out + orow * (int64_t)p.ldo;
            const int n0 = n_tile * p.BN;
            // fused BN statistics: which of this warp's 32 rows are real output rows, and where the window changes
            uint32_t bn_vmask = 0;
            int bn_split = 32;
            const bool second = p.n_split > 0 && n_tile >= p.n_split;          // dual output: the second GEMM's tiles
            float2 *bn_dst = second ? p.bn_part2 : p.bn_part;
            const int bn_N = p.n_split > 0 ? p.n_half : p.N;
            const int n_org = second ? p.n_split * p.BN : 0;
            if (bn_dst) {
                const int Tp = p.pad ? p.T + 2 : p.T;
                const int64_t w_lane = prow / Tp, w_first = __shfl_sync(0xffffffffu, w_lane, 0);
                bn_vmask = __ballot_sync(0xffffffffu, row_ok);
                const uint32_t diff = __ballot_sync(0xffffffffu, w_lane != w_first);
                bn_split = diff ? __ffs(diff) - 1 : 32;
            }
            float am_v = -INFINITY;                            // argmax mode: this row's best value / column over the warp's chunks
            int am_i = -1;
            for (int c0 = eh * 32; c0 < p.BN; c0 += 32 * (EW / 4)) {
                uint32_t r[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256 + c0);
                if (c0 + 32 <= p.BN) {
                    tmem_ld_32x32(taddr, r);
                } else {                                           // BN % 32 == 16 tail
                    uint32_t h[16];
                    tmem_ld_32x16(taddr, h);
#pragma unroll
                    for (int i = 0; i < 16; ++i) { r[i] = h[i]; r[i + 16] = 0; }
                }
                tmem_ld_wait();
                if (p.tma_store) {
                    // registers -> 128B-swizzled [32 rows][32 cols] staging tile -> one TMA store per warp
                    uint8_t *buf = stage_buf + (EW == 4 ? (stage_it & 1) * 4096 : 0);
                    ++stage_it;
                    if (lane == 0) { if (EW == 4) bulk_wait_group_read<1>(); else bulk_wait_group_read<0>(); }   // the store that last read `buf` is done
                    __syncwarp();
                    const int nb = n0 + c0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                               __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                        if (p.bias) {
                            const int n = nb + 4 * j;
                            v.x += n + 0 < p.N ? p.bias[n + 0] : 0.0f; v.y += n + 1 < p.N ? p.bias[n + 1] : 0.0f;
                            v.z += n + 2 < p.N ? p.bias[n + 2] : 0.0f; v.w += n + 3 < p.N ? p.bias[n + 3] : 0.0f;
                        }
                        *reinterpret_cast<float4 *>(buf + lane * 128 + ((j ^ (lane & 7)) << 4)) = v;
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_2d(second ? &map_out2 : &map_out, buf, nb - n_org, (int)(m_tile * BM) + q * 32);
                        bulk_commit_group();
                    }
                    if (bn_dst) bn_partials(buf, lane, bn_vmask, bn_split, bn_dst, (m_tile * 4 + q) * 2, bn_N, nb - n_org + lane);
                } else if (LSTM) {
                    // thread = one window of one direction; this chunk = 8 hidden units x (i, j, f, o).  Every global access
                    // of a warp is contiguous: projected gates unit-major ([..][unit][128 windows] float4, the layout the
                    // projection's epilogue writes), cell state [..][unit / 4][128][4], h planes k-group-major -- with
                    // row-major buffers each warp-wide 16-byte access touched 32 lines and the L1 wavefronts (576 per chunk
                    // and warp) took three times as long as the tile's MMAs
                    const int H = p.lstm_H, nb = n0 + c0;
                    if (l_live && nb < p.N) {             // pad rows of the h planes stay at their initial zero
                        const int u0 = nb >> 2;                               // 8 units (4 in the 16-column tail chunk of a tile when BN % 32 == 16)
                        const int nu = (p.N - nb >= 32 && c0 + 32 <= p.BN) ? 8 : 4;
                        float4 xcur[8], ccur[2];
#pragma unroll
                        for (int k = 0; k < 8; ++k) xcur[k] = xpre[k];
                        ccur[0] = cpre[0]; ccur[1] = cpre[1];
                        float4 *cg = l_c + (int64_t)(u0 >> 2) * BM;
                        lstm_fetch(c0 + 32 * (EW / 4));   // the next chunk's operands fly during this chunk's cell updates
                        float hv[8];
#pragma unroll
                        for (int half = 0; half < 2; ++half) {
                            if (half * 4 >= nu) { hv[4] = hv[5] = hv[6] = hv[7] = 0.0f; break; }
                            float cc[4] = {ccur[half].x, ccur[half].y, ccur[half].z, ccur[half].w};
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const int i = 16 * half + 4 * u;
                                const float4 xv = xcur[4 * half + u];
                                float cn = cc[u], hh;
                                lstm_cell_update(__uint_as_float(r[i]) + xv.x, __uint_as_float(r[i + 1]) + xv.y,
                                                 __uint_as_float(r[i + 2]) + xv.z, __uint_as_float(r[i + 3]) + xv.w, cn, hh);
                                cc[u] = l_active ? cn : cc[u];
                                hv[4 * half + u] = l_active ? hh : 0.0f;
                            }
                            if (l_active) cg[half * BM] = make_float4(cc[0], cc[1], cc[2], cc[3]);
                        }
                        uint4 ph, pl;
                        split_pack_op16x2(hv[0], hv[1], ph.x, pl.x);
                        split_pack_op16x2(hv[2], hv[3], ph.y, pl.y);
                        split_pack_op16x2(hv[4], hv[5], ph.z, pl.z);
                        split_pack_op16x2(hv[6], hv[7], ph.w, pl.w);
                        if (p.lstm_nx_hi) {
                            const int64_t no = ((l_tile * p.T + l_t) * BM + l_wl) * (int64_t)p.lstm_nx_K + l_dir * H + u0;
                            if (nu == 8) {
                                *reinterpret_cast<uint4 *>(p.lstm_nx_hi + no) = ph;
                                *reinterpret_cast<uint4 *>(p.lstm_nx_lo + no) = pl;
                            } else {
                                *reinterpret_cast<uint2 *>(p.lstm_nx_hi + no) = make_uint2(ph.x, ph.y);
                                *reinterpret_cast<uint2 *>(p.lstm_nx_lo + no) = make_uint2(pl.x, pl.y);
                            }
                        } else {
                            float *og = p.out + (l_w * p.T + l_t) * (int64_t)p.ldo + l_dir * H + u0;
                            *reinterpret_cast<float4 *>(og) = make_float4(hv[0], hv[1], hv[2], hv[3]);
                            if (nu == 8) *reinterpret_cast<float4 *>(og + 4) = make_float4(hv[4], hv[5], hv[6], hv[7]);
                        }
                        if (nu == 8 && (u0 & 7) == 0) {                      // one whole K group: 16 bytes per plane
                            const int64_t po = ((int64_t)(u0 >> 3) * (2 * p.lstm_Mp) + prow) * 8;
                            *reinterpret_cast<uint4 *>(p.lstm_hi + po) = ph;
                            *reinterpret_cast<uint4 *>(p.lstm_lo + po) = pl;
                        } else {                                             // N tiles of 208 columns: chunks start mid-group
                            const int64_t p0 = ((int64_t)(u0 >> 3) * (2 * p.lstm_Mp) + prow) * 8 + (u0 & 7);
                            *reinterpret_cast<uint2 *>(p.lstm_hi + p0) = make_uint2(ph.x, ph.y);
                            *reinterpret_cast<uint2 *>(p.lstm_lo + p0) = make_uint2(pl.x, pl.y);
                            if (nu == 8) {
                                const int u1 = u0 + 4;
                                const int64_t p1 = ((int64_t)(u1 >> 3) * (2 * p.lstm_Mp) + prow) * 8 + (u1 & 7);
                                *reinterpret_cast<uint2 *>(p.lstm_hi + p1) = make_uint2(ph.z, ph.w);
                                *reinterpret_cast<uint2 *>(p.lstm_lo + p1) = make_uint2(pl.z, pl.w);
                            }
                        }
                    }
                } else if (!LSTM && p.amax) {
                    if (row_ok) {
                        const int nb = n0 + c0;
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            const int n = nb + i;
                            if (n < p.N && c0 + i < p.BN) {
                                const float v = __uint_as_float(r[i]) + (p.bias ? __ldg(p.bias + n) : 0.0f);
                                if (v > am_v) { am_v = v; am_i = n; }       // strict: the first maximum stays
                            }
                        }
                    }
                } else if (p.xproj_H > 0 && m_tile < p.m_tiles) {
                    if (p.abl & 1) continue;
                    // xproj[dir][m_tile][unit][row] float4 (i,j,f,o): a warp's 32 rows are contiguous (512 B per unit).
                    float4 *out4 = reinterpret_cast<float4 *>(p.out);
                    const int nb = n0 + c0, H = p.xproj_H;
                    const int d = (nb >> 2) / H, u = (nb >> 2) - d * H;
                    const int64_t dir_stride = p.m_tiles * (int64_t)H * BM;
                    float4 *dst = out4 + d * dir_stride + (m_tile * H + u) * (int64_t)BM + (q * 32 + lane);
                    int lim = (p.N < n0 + p.BN ? p.N : n0 + p.BN) - nb;       // valid columns of this chunk
                    lim = lim > 32 ? 32 : lim;
                    if (lim == 32 && u + 8 <= H && bias_in_smem && p.xproj_tma) {
                        // common case, TMA form: [8 units][32 rows] float4 staging tile (4 KB, conflict-free) -> ONE
                        // 3-D tensor store {128 floats, 1, 8 units} per warp and chunk
                        uint8_t *buf = stage_buf + (EW == 4 ? (stage_it & 1) * 4096 : 0);
                        ++stage_it;
                        if (lane == 0) { if (EW == 4) bulk_wait_group_read<1>(); else bulk_wait_group_read<0>(); }
                        __syncwarp();
                        const float4 *b4 = reinterpret_cast<const float4 *>(bias_s + nb);
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const float4 bb = b4[k];
                            *reinterpret_cast<float4 *>(buf + k * 512 + lane * 16) =
                                make_float4(__uint_as_float(r[4 * k]) + bb.x, __uint_as_float(r[4 * k + 1]) + bb.y,
                                            __uint_as_float(r[4 * k + 2]) + bb.z, __uint_as_float(r[4 * k + 3]) + bb.w);
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            tma_store_3d(&map_out, buf, (q & 1) * 128, q >> 1, (int)((d * p.m_tiles + m_tile) * H + u));
                            bulk_commit_group();
                        }
                    } else if (lim == 32 && u + 8 <= H && bias_in_smem) {
                        // common case: 8 whole units of one direction -> straight-line code, constant store offsets
                        const float4 *b4 = reinterpret_cast<const float4 *>(bias_s + nb);
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const float4 bb = b4[k];
                            dst[k * BM] = make_float4(__uint_as_float(r[4 * k]) + bb.x, __uint_as_float(r[4 * k + 1]) + bb.y,
                                                      __uint_as_float(r[4 * k + 2]) + bb.z, __uint_as_float(r[4 * k + 3]) + bb.w);
                        }
                    } else {
                        int uu = u;
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            if (i < lim) {
                                float4 v = make_float4(__uint_as_float(r[i]), __uint_as_float(r[i + 1]),
                                                       __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
                                if (p.bias) {
                                    const float4 bb = bias_in_smem ? *reinterpret_cast<const float4 *>(bias_s + nb + i)
                                                                   : __ldg(reinterpret_cast<const float4 *>(p.bias + nb + i));
                                    v.x += bb.x; v.y += bb.y; v.z += bb.z; v.w += bb.w;
                                }
                                *dst = v;
                            }
                            dst += BM;
                            if (++uu == H) { uu = 0; dst += dir_stride - (int64_t)H * BM; }
                        }
                    }
                } else {
                  if (p.bn_part) {
                    // statistics of the tile (direct-store path, e.g. the padded-row 1x3 conv): stage, reduce columns
                    uint8_t *buf = stage_buf + (EW == 4 ? (stage_it & 1) * 4096 : 0);
                    ++stage_it;
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<float4 *>(buf + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                            make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                        __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                    __syncwarp();
                    bn_partials(buf, lane, bn_vmask, bn_split, p.bn_part, (m_tile * 4 + q) * 2, p.N, n0 + c0 + lane);
                  }
                  if (row_ok) {
                    const int nb = n0 + c0;
                    if (vec_ok && nb + 32 <= p.N && c0 + 32 <= p.BN) {
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            float4 v = make_float4(__uint_as_float(r[i]), __uint_as_float(r[i + 1]),
                                                   __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
                            if (p.bias) {
                                const float4 b = *reinterpret_cast<const float4 *>(p.bias + nb + i);
                                v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
                            }
                            *reinterpret_cast<float4 *>(orow_ptr + nb + i) = v;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            const int n = nb + i;
                            if (n < p.N && c0 + i < p.BN)
                                orow_ptr[n] = __uint_as_float(r[i]) + (p.bias ? p.bias[n] : 0.0f);
                        }
                    }
                  }
                }
            }
            if (!LSTM && p.amax && am_i >= 0) {
                const uint32_t u = __float_as_uint(am_v + 0.0f);                            // (-0.0 and +0.0 compare equal: one key)
                const uint32_t ord = u ^ ((u >> 31) ? 0xffffffffu : 0x80000000u);          // monotonic in the float order
                atomicMax(p.amax + orow, ((unsigned long long)ord << 32) | (unsigned long long)(0xffffffffu - (uint32_t)am_i));
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {                                   // the LEADER's MMA warp owns the accumulators of both SMs
                if (CG == 1 || rank == 0) mbar_arrive(&acc_empty[acc]); else mbar_arrive_cluster(&acc_empty[acc], 0);
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if ((p.tma_store || p.xproj_tma) && lane == 0) bulk_wait_group<0>();   // stores complete before smem goes away
    }
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        if (CG == 2) tmem_dealloc_pair(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------
// K2b'  conv GEMM with the operand transform fused in (conv2b / conv2c of residual_layer, network.py:385-400):
//   out = conv( relu( y * a[g] + b[g] ) )        y = the previous conv's raw fp32 output, (a, b) its BN affine
// The standalone split pass (read y, write fp16 hi/lo planes, read them back) is gone: warp 0 TMA-loads raw fp32
// tiles of y, sixteen TRANSFORM warps apply affine + ReLU, split to fp16 hi/lo and write the 128-byte-swizzled UMMA
// operand tiles in shared memory, the MMA warp of the pair's leader consumes them.  A 1x3 conv loads ONE raw tile of
// 130 rows per k-block (rows r0-1 .. r0+128) and produces its three taps from it (zero rows outside the window =
// SAME padding), so the padded plane layout is not needed either and the output goes out through TMA stores with the
// BN partials (bn_partials) like the 1x1 convs.
// CTA pair (cta_group::2, M = 256): each CTA transforms its own 128 rows and loads half of the weight tile.
// Shared memory: 2 raw stages (33 KB) + 2 operand stages (A hi/lo 32 KB + B hi/lo <= 32 KB) + 16 KB staging.
// ------------------------------------------------------------------------------------------
constexpr int CF_THREADS = 768;                 // warps: 0 TMA, 1 MMA, 2 TMEM, 3 idle, 4-7 epilogue, 8-23 transform
constexpr int CF_TWARPS = 16;
constexpr int CF_RAW_STAGE = 33792;             // 130 rows x 64 fp32 = 33280, rounded to 1 KB
constexpr int CF_RAW_STAGES = 2, CF_OP_STAGES = 2;
constexpr int CF_EPI_BYTES = 4 * 4096;

struct ConvFusedParams {
    int64_t m_tiles, rows;
    int N, K, taps, T, k_blocks, relu;
    const float *aff_a, *aff_b;
    const int32_t *win_group;
    float2 *bn_part;
    const float *xs;   // != null: the producer conv had ONE input channel and is folded into (aff_a, aff_b) (k_k1_affine): the
                       // operand is relu(xs[row] * a[g][c] + b[g][c]) -- no raw tile exists, the transform warps make it from xs
    int abl;       // timing ablations (NTB_CF_ABL): 1 = no BN partials, 2 = no output store, 4 = transform warps skip their work
    int l2pf;      // the producer pulls the NEXT tile's raw rows into L2 while this tile runs: only one or two 33 KB raw
                   // stages are in flight per SM, so every k-block otherwise waits a full DRAM latency
};

__global__ void __launch_bounds__(CF_THREADS, 1)
k_conv_fused(const __grid_constant__ CUtensorMap map_raw, const __grid_constant__ CUtensorMap map_b_hi,
             const __grid_constant__ CUtensorMap map_b_lo, const __grid_constant__ CUtensorMap map_out,
             const ConvFusedParams p) {
    using namespace ptx;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t raw_full[CF_RAW_STAGES], raw_empty[CF_RAW_STAGES], op_full[CF_OP_STAGES], op_empty[CF_OP_STAGES],
        acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base_smem;

    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t *smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
    const int b_tile_bytes = (p.N / 2) * BK * 2;                       // this CTA's half of one B plane tile
    const int op_stage_bytes = 2 * A_TILE_BYTES + 2 * b_tile_bytes;    // A_hi | A_lo | B_hi | B_lo
    uint8_t *op_base = smem;
    uint8_t *raw_base = op_base + CF_OP_STAGES * op_stage_bytes;
    uint8_t *epi_base = raw_base + CF_RAW_STAGES * CF_RAW_STAGE;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int64_t pairs = (p.m_tiles + 1) / 2;
    const int64_t first = blockIdx.x / 2, step = gridDim.x / 2;
    const int left = (p.taps - 1) / 2;
    const int raw_rows = p.taps == 3 ? BM + 2 : BM;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&map_raw); prefetch_tmap(&map_b_hi); prefetch_tmap(&map_b_lo); prefetch_tmap(&map_out);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < CF_RAW_STAGES; ++s) { mbar_init(&raw_full[s], 1); mbar_init(&raw_empty[s], CF_TWARPS); }
        for (int s = 0; s < CF_OP_STAGES; ++s) { mbar_init(&op_full[s], 1 + 2 * CF_TWARPS); mbar_init(&op_empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 8); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_pair(&tmem_base_smem, TMEM_COLS);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == 0) {
        // ===================== TMA producer: raw fp32 activation tiles + weight tiles =====================
        if (lane == 0) {
            int rs = 0, os = 0; uint32_t rph = 0, oph = 0;
            for (int64_t it = first; it < pairs; it += step) {
                const int64_t m_tile = it * 2 + rank;
                if (p.l2pf && !p.xs && it + step < pairs) {
                    const int r2 = (int)(((it + step) * 2 + rank) * BM) - left;
                    for (int kb = 0; kb < p.k_blocks; ++kb) tma_prefetch_2d(&map_raw, kb * BK, r2);
                }
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    if (!p.xs) {
                        mbar_wait_backoff<128>(&raw_empty[rs], rph ^ 1, 100);
                        mbar_expect_tx(&raw_full[rs], (uint32_t)(raw_rows * BK * 4));
                        tma_load_2d(raw_base + rs * CF_RAW_STAGE, &map_raw, &raw_full[rs], kb * BK, (int)(m_tile * BM) - left);
                        if (++rs == CF_RAW_STAGES) { rs = 0; rph ^= 1; }
                    }
                    for (int tap = 0; tap < p.taps; ++tap) {
                        mbar_wait_backoff<128>(&op_empty[os], oph ^ 1, 101);
                        uint8_t *s = op_base + os * op_stage_bytes + 2 * A_TILE_BYTES;
                        const int brow0 = tap * p.N + rank * (p.N / 2);
                        if (rank == 0) mbar_expect_tx(&op_full[os], (uint32_t)(4 * b_tile_bytes));   // both CTAs, both planes
                        tma_load_2d_pair(s, &map_b_hi, &op_full[os], kb * BK, brow0);
                        tma_load_2d_pair(s + b_tile_bytes, &map_b_lo, &op_full[os], kb * BK, brow0);
                        if (++os == CF_OP_STAGES) { os = 0; oph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA) =====================
        if (lane == 0 && rank == 0) {
            const uint32_t idesc = idesc_f16_f32(2 * BM, p.N);
            int os = 0; uint32_t oph = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int64_t it = first; it < pairs; it += step) {
                mbar_wait_backoff<64>(&acc_empty[acc], acc_phase ^ 1, 102);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
                for (int kb = 0; kb < p.k_blocks * p.taps; ++kb) {
                    mbar_wait_backoff(&op_full[os], oph, 103);
                    tc_fence_after();
                    const uint32_t s = smem_u32(op_base + os * op_stage_bytes);
                    const uint64_t a_hi = smem_desc_sw128(s), a_lo = smem_desc_sw128(s + A_TILE_BYTES);
                    const uint64_t b_hi = smem_desc_sw128(s + 2 * A_TILE_BYTES);
                    const uint64_t b_lo = smem_desc_sw128(s + 2 * A_TILE_BYTES + b_tile_bytes);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        umma_f16_pair(d_tmem, desc_advance_k16(a_hi, k), desc_advance_k16(b_hi, k), idesc, (kb | k) != 0);
                        umma_f16_pair(d_tmem, desc_advance_k16(a_lo, k), desc_advance_k16(b_hi, k), idesc, 1);
                        umma_f16_pair(d_tmem, desc_advance_k16(a_hi, k), desc_advance_k16(b_lo, k), idesc, 1);
                    }
                    umma_commit_pair(&op_empty[os], 3);
                    if (++os == CF_OP_STAGES) { os = 0; oph ^= 1; }
                }
                umma_commit_pair(&acc_full[acc], 3);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 8) {
        // ===================== transform: raw fp32 -> affine + ReLU -> fp16 hi/lo swizzled operand tiles ==========
        // thread = channels 4j .. 4j+3 and 32+4j .. 32+4j+3 of the k-block for 4 rows of its warp's 16: the eight lanes of a
        // quarter warp read 128 CONTIGUOUS bytes of one raw row per 16-byte load (the row pitch is 256 B: a 32-byte lane
        // stride would put lanes j and j+4 on the same banks), and a half warp writes rows r and r+4, whose swizzle keys
        // differ in bit 2, so its sixteen 8-byte operand stores fall on 128 distinct bytes of the banks.
        constexpr int RPT = BM / (4 * CF_TWARPS);                 // rows per thread
        const int tw = warp - 8, j = lane & 7, rsel = lane >> 3;
        const int rlane = (rsel & 1) * 4 + (rsel >> 1);           // row offsets {0, 4, 1, 5}
        int rs = 0, os = 0; uint32_t rph = 0, oph = 0;
        for (int64_t it = first; it < pairs; it += step) {
            const int64_t m_tile = it * 2 + rank;
            int tpos[RPT], grp[RPT];                              // t of the row inside its window (-1000: no such row), BN group
#pragma unroll
            for (int i = 0; i < RPT; ++i) {
                const int64_t grow = m_tile * BM + tw * (4 * RPT) + (i >> 1) * 8 + (i & 1) * 2 + rlane;
                if (grow < p.rows) {
                    const int64_t w = grow / p.T;
                    tpos[i] = (int)(grow - w * p.T);
                    grp[i] = p.aff_a ? __ldg(p.win_group + w) : 0;
                } else {
                    tpos[i] = -1000; grp[i] = 0;
                }
            }
            float xv[RPT][3];                                     // scalar-input form: x of the row's taps, once per tile
            if (p.xs) {
#pragma unroll
                for (int i = 0; i < RPT; ++i)
#pragma unroll
                    for (int tap = 0; tap < 3; ++tap) {
                        const int64_t grow = m_tile * BM + tw * (4 * RPT) + (i >> 1) * 8 + (i & 1) * 2 + rlane + tap - left;
                        const int tt = tpos[i] + tap - left;
                        xv[i][tap] = (tap < p.taps && tt >= 0 && tt < p.T) ? __ldg(p.xs + grow) : 0.0f;
                    }
            }
            for (int kb = 0; kb < p.k_blocks; ++kb) {
                if (!p.xs) mbar_wait(&raw_full[rs], rph, 200);
                const uint8_t *raw = raw_base + rs * CF_RAW_STAGE;
                const int c = kb * BK + j * 4;                    // first of this thread's channels (second group: c + 32)
                int gcur = -1;                                    // (a, b) of group gcur for the 8 channels stay in registers
                float a[8], b[8];                                 //  across the taps (and rows) of this k-block
                for (int tap = 0; tap < p.taps; ++tap) {
                    mbar_wait(&op_empty[os], oph ^ 1, 201);
                    uint8_t *a_hi = op_base + os * op_stage_bytes, *a_lo = a_hi + A_TILE_BYTES;
#pragma unroll
                    for (int i = 0; i < RPT; ++i) {
                        if (p.abl & 4) break;
                        const int row = tw * (4 * RPT) + (i >> 1) * 8 + (i & 1) * 2 + rlane;
                        const int tt = tpos[i] + tap - left;
                        float v[8];
                        if (tt >= 0 && tt < p.T) {                // same window as the output row: SAME padding otherwise
                            if (p.xs) {
#pragma unroll
                                for (int q = 0; q < 8; ++q) v[q] = tap == 0 ? xv[i][0] : (tap == 1 ? xv[i][1] : xv[i][2]);
                            } else {
                                const float4 x0 = *reinterpret_cast<const float4 *>(raw + (row + tap) * (BK * 4) + j * 16);
                                const float4 x1 = *reinterpret_cast<const float4 *>(raw + (row + tap) * (BK * 4) + 128 + j * 16);
                                v[0] = x0.x; v[1] = x0.y; v[2] = x0.z; v[3] = x0.w; v[4] = x1.x; v[5] = x1.y; v[6] = x1.z; v[7] = x1.w;
                            }
                            if (p.aff_a) {
                                if (grp[i] != gcur) {
                                    gcur = grp[i];
                                    const float4 a0 = __ldg(reinterpret_cast<const float4 *>(p.aff_a + (int64_t)gcur * p.K + c));
                                    const float4 a1 = __ldg(reinterpret_cast<const float4 *>(p.aff_a + (int64_t)gcur * p.K + c + 32));
                                    const float4 b0 = __ldg(reinterpret_cast<const float4 *>(p.aff_b + (int64_t)gcur * p.K + c));
                                    const float4 b1 = __ldg(reinterpret_cast<const float4 *>(p.aff_b + (int64_t)gcur * p.K + c + 32));
                                    a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
                                    b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
                                }
#pragma unroll
                                for (int q = 0; q < 8; ++q) v[q] = fmaf(v[q], a[q], b[q]);
                            }
                            if (p.relu) {
#pragma unroll
                                for (int q = 0; q < 8; ++q) v[q] = fmaxf(v[q], 0.0f);
                            }
                        } else {
#pragma unroll
                            for (int q = 0; q < 8; ++q) v[q] = 0.0f;
                        }
                        uint4 h, l;
                        if (p.relu) split_op16x8<true>(v, h, l); else split_op16x8<false>(v, h, l);
                        // channels 4j..4j+3 -> half (j & 1) of 16-byte chunk j >> 1; channels 32+4j.. -> chunk 4 + (j >> 1)
                        const int off0 = row * 128 + (((j >> 1) ^ (row & 7)) << 4) + (j & 1) * 8;
                        const int off1 = row * 128 + (((4 + (j >> 1)) ^ (row & 7)) << 4) + (j & 1) * 8;
                        *reinterpret_cast<uint2 *>(a_hi + off0) = make_uint2(h.x, h.y);
                        *reinterpret_cast<uint2 *>(a_hi + off1) = make_uint2(h.z, h.w);
                        *reinterpret_cast<uint2 *>(a_lo + off0) = make_uint2(l.x, l.y);
                        *reinterpret_cast<uint2 *>(a_lo + off1) = make_uint2(l.z, l.w);
                    }
                    fence_proxy_async_smem();                     // generic-proxy stores -> visible to tcgen05.mma
                    __syncwarp();
                    if (lane == 0) { if (rank == 0) mbar_arrive(&op_full[os]); else mbar_arrive_cluster(&op_full[os], 0); }
                    if (++os == CF_OP_STAGES) { os = 0; oph ^= 1; }
                }
                if (!p.xs) {
                    __syncwarp();                                 // every lane is done reading the raw stage
                    if (lane == 0) mbar_arrive(&raw_empty[rs]);
                    if (++rs == CF_RAW_STAGES) { rs = 0; rph ^= 1; }
                }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue: TMEM -> swizzled staging -> TMA store, BN partials =====================
        const int q = warp & 3;
        int acc = 0; uint32_t acc_phase = 0;
        uint8_t *buf = epi_base + q * 4096;
        for (int64_t it = first; it < pairs; it += step) {
            const int64_t m_tile = it * 2 + rank;
            mbar_wait_backoff<128>(&acc_full[acc], acc_phase, 300);   // long wait next to the transform warps: sleep
            tc_fence_after();
            const int64_t prow = m_tile * BM + q * 32 + lane;
            uint32_t bn_vmask = 0;
            int bn_split = 32;
            if (p.bn_part) {
                const int64_t w_lane = prow / p.T, w_first = __shfl_sync(0xffffffffu, w_lane, 0);
                bn_vmask = __ballot_sync(0xffffffffu, prow < p.rows);
                const uint32_t diff = __ballot_sync(0xffffffffu, w_lane != w_first);
                bn_split = diff ? __ffs(diff) - 1 : 32;
            }
            for (int c0 = 0; c0 < p.N; c0 += 32) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256 + c0), r);
                tmem_ld_wait();
                if (lane == 0) bulk_wait_group_read<0>();          // the previous store has read the staging tile
                __syncwarp();
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    *reinterpret_cast<float4 *>(buf + lane * 128 + ((k ^ (lane & 7)) << 4)) =
                        make_float4(__uint_as_float(r[4 * k]), __uint_as_float(r[4 * k + 1]), __uint_as_float(r[4 * k + 2]),
                                    __uint_as_float(r[4 * k + 3]));
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0 && !(p.abl & 2)) {
                    tma_store_2d(&map_out, buf, c0, (int)(m_tile * BM) + q * 32);
                    bulk_commit_group();
                }
                if (p.bn_part && !(p.abl & 1)) bn_partials(buf, lane, bn_vmask, bn_split, p.bn_part, (m_tile * 4 + q) * 2, p.N, c0 + lane);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (rank == 0) mbar_arrive(&acc_empty[acc]); else mbar_arrive_cluster(&acc_empty[acc], 0); }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (lane == 0) bulk_wait_group<0>();
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------
// fp32 -> fp16 hi/lo operand planes, with the producer's BN affine (+ReLU) applied on the way
// (simple_global_bn + relu of the previous conv, network.py:330-349) and, for the 1x3 conv, the
// per-window padded row layout.  8 channels per thread: two float4 loads, two 16-byte stores.
// ------------------------------------------------------------------------------------------
struct SplitParams {
    const float *A; int lda;
    int64_t rows;            // source rows (= n_windows * T when pad)
    int K;                   // multiple of 8
    const float *aff_a, *aff_b; const int32_t *win_group;
    int T, relu, pad;
    op16_t *hi, *lo;  // [plane_rows][K]
    int64_t plane_rows;
    int perm;                // 1: plane row (tile*T + t)*128 + wl <- source row (tile*128 + wl)*T + t
                             // 2: same rows, k-group-major tiles [tile][t][K8p][128][8] (fused BLSTM kernel, ntb_blstm_fused.cuh)
    int K8p;                 // perm 2: 8-channel groups per (tile, t) = 2 * ceil(K / 16); groups >= K / 8 are zero
    int expand;              // 1: A is [rows] (a Cin = 1 conv folded into the affine): v[c] = A[r] * a[g][c] + b[g][c]
    int Ksrc;                // > 0: the source has only Ksrc (% 4 == 0) channels, the plane's other channels are zero
};

__global__ void __launch_bounds__(256) k_split_op16(const SplitParams p) {
    const int K8 = p.K >> 3;
    const int64_t total = p.plane_rows * K8;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t pr = i / K8;
        const int c = (int)(i - pr * K8) << 3;
        int64_t r = pr, w = 0;
        bool real = true;
        if (p.perm) {
            const int64_t mt = pr >> 7;
            const int wl = (int)(pr & 127);
            const int64_t tile = mt / p.T;
            const int t = (int)(mt - tile * p.T);
            w = tile * 128 + wl;
            r = w * p.T + t;
        } else if (p.pad) {
            w = pr / (p.T + 2);
            const int tt = (int)(pr - w * (p.T + 2)) - 1;
            real = tt >= 0 && tt < p.T;
            r = w * p.T + tt;
        } else if (p.aff_a) {
            w = pr / p.T;
        }
        real = real && r < p.rows;
        float v[8];
        if (real) {
            const int ks = p.Ksrc > 0 ? p.Ksrc : p.K;
            const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 x0 = c < ks ? *reinterpret_cast<const float4 *>(p.A + r * p.lda + c) : zero4;
            const float4 x1 = c + 4 < ks ? *reinterpret_cast<const float4 *>(p.A + r * p.lda + c + 4) : zero4;
            v[0] = x0.x; v[1] = x0.y; v[2] = x0.z; v[3] = x0.w; v[4] = x1.x; v[5] = x1.y; v[6] = x1.z; v[7] = x1.w;
            if (p.aff_a) {
                const int g = p.win_group[w];
                const float *a = p.aff_a + (int64_t)g * p.K + c, *b = p.aff_b + (int64_t)g * p.K + c;
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = v[j] * a[j] + b[j];
            }
            if (p.relu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.0f);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.0f;
        }
        uint4 h, l;
        ptx::split_op16x8(v, h, l);
        if (p.perm == 2) {
            const int64_t o = (((pr >> 7) * p.K8p + (c >> 3)) * 128 + (pr & 127)) * 8;
            *reinterpret_cast<uint4 *>(p.hi + o) = h;
            *reinterpret_cast<uint4 *>(p.lo + o) = l;
            if (c + 8 == p.K)
                for (int g = K8; g < p.K8p; ++g) {
                    const int64_t oz = (((pr >> 7) * p.K8p + g) * 128 + (pr & 127)) * 8;
                    *reinterpret_cast<uint4 *>(p.hi + oz) = make_uint4(0, 0, 0, 0);
                    *reinterpret_cast<uint4 *>(p.lo + oz) = make_uint4(0, 0, 0, 0);
                }
            continue;
        }
        *reinterpret_cast<uint4 *>(p.hi + pr * p.K + c) = h;
        if (p.lo) *reinterpret_cast<uint4 *>(p.lo + pr * p.K + c) = l;
    }
}

// Same transform with one CTA per window: the BN affine of the producer is constant over a window
// (window -> group -> [a, b] per channel), so every thread keeps its 8 channels' a/b in registers
// and walks the window's rows; no per-element index arithmetic.  Used for the conv operands.
__global__ void __launch_bounds__(256) k_split_window(const SplitParams p) {
    const int K8 = p.K >> 3;
    const int rows_per_it = blockDim.x / K8;
    const int cg = threadIdx.x % K8, rs = threadIdx.x / K8;
    if (rs >= rows_per_it) return;
    const int c = cg << 3;
    const int64_t w = blockIdx.x;
    float a[8], b[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { a[j] = 1.0f; b[j] = 0.0f; }
    if (p.aff_a) {
        const int g = p.win_group[w];
        const float4 a0 = *reinterpret_cast<const float4 *>(p.aff_a + (int64_t)g * p.K + c);
        const float4 a1 = *reinterpret_cast<const float4 *>(p.aff_a + (int64_t)g * p.K + c + 4);
        const float4 b0 = *reinterpret_cast<const float4 *>(p.aff_b + (int64_t)g * p.K + c);
        const float4 b1 = *reinterpret_cast<const float4 *>(p.aff_b + (int64_t)g * p.K + c + 4);
        a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
        b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
    }
    const int prow_n = p.pad ? p.T + 2 : p.T;              // plane rows of this window
    const int64_t pr0 = w * prow_n;
    const float *src = p.A + (w * p.T) * (int64_t)p.lda + c;
    for (int i = rs; i < prow_n; i += rows_per_it) {
        const int tt = p.pad ? i - 1 : i;
        float v[8];
        if (tt >= 0 && tt < p.T) {
            if (p.expand) {
                const float xv = __ldg(p.A + w * p.T + tt);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = xv;
            } else {
                const float4 x0 = *reinterpret_cast<const float4 *>(src + (int64_t)tt * p.lda);
                const float4 x1 = *reinterpret_cast<const float4 *>(src + (int64_t)tt * p.lda + 4);
                v[0] = x0.x; v[1] = x0.y; v[2] = x0.z; v[3] = x0.w; v[4] = x1.x; v[5] = x1.y; v[6] = x1.z; v[7] = x1.w;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                v[j] = fmaf(v[j], a[j], b[j]);
                if (p.relu) v[j] = fmaxf(v[j], 0.0f);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.0f;
        }
        uint4 h, l;
        ptx::split_op16x8(v, h, l);
        *reinterpret_cast<uint4 *>(p.hi + (pr0 + i) * p.K + c) = h;
        if (p.lo) *reinterpret_cast<uint4 *>(p.lo + (pr0 + i) * p.K + c) = l;
    }
}

// residual_layer tail (network.py:398-399)  out = relu( bn?(branch1) + bn(conv2c) )  emitted directly as
// the NEXT GEMM's fp16 hi/lo operand planes (plain rows for the next residual layer's 1x1 convs, time-major
// 128-window tiles for the LSTM projection) and/or as fp32.  8 channels per thread, one plane row per
// K/8 threads.  y1_scalar: branch1 is a Cin = 1 conv folded into its affine (y1[r] * a1[g][c] + b1[g][c]).
struct CombineParams {
    const float *y1; int y1_scalar;
    const float *a1, *b1;            // [n_groups][C] or null (branch1 without BN)
    const float *y2, *a2, *b2;
    const int32_t *win_group;
    int T, C;
    int64_t rows;                    // source rows = n_windows * T
    float *out;                      // fp32 [rows][C] or null
    op16_t *hi, *lo;          // planes or null
    int64_t plane_rows;              // = rows unless perm (whole 128-window tiles)
    int perm;                        // 1: time-major tiles, 2: k-group-major tiles (SplitParams::perm)
    int K8p;
};

// relu( bn?(branch1) + bn(conv2c) ) of 8 channels c..c+7 of source row r (window w)
__device__ __forceinline__ void combine_values(const CombineParams &p, int64_t r, int64_t w, int c, float (&v)[8]) {
    const int g = p.win_group[w];
    float a2[8], b2[8];
    {
        const float4 *pa = reinterpret_cast<const float4 *>(p.a2 + (int64_t)g * p.C + c);
        const float4 *pb = reinterpret_cast<const float4 *>(p.b2 + (int64_t)g * p.C + c);
        const float4 a0 = __ldg(pa), a1 = __ldg(pa + 1), b0 = __ldg(pb), b1 = __ldg(pb + 1);
        a2[0] = a0.x; a2[1] = a0.y; a2[2] = a0.z; a2[3] = a0.w; a2[4] = a1.x; a2[5] = a1.y; a2[6] = a1.z; a2[7] = a1.w;
        b2[0] = b0.x; b2[1] = b0.y; b2[2] = b0.z; b2[3] = b0.w; b2[4] = b1.x; b2[5] = b1.y; b2[6] = b1.z; b2[7] = b1.w;
    }
    const float4 y0 = *reinterpret_cast<const float4 *>(p.y2 + r * p.C + c);
    const float4 y1 = *reinterpret_cast<const float4 *>(p.y2 + r * p.C + c + 4);
    const float yy[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
    float u[8];
    if (p.y1_scalar) {
        const float xv = __ldg(p.y1 + r);
#pragma unroll
        for (int j = 0; j < 8; ++j) u[j] = xv;
    } else {
        const float4 u0 = *reinterpret_cast<const float4 *>(p.y1 + r * p.C + c);
        const float4 u1 = *reinterpret_cast<const float4 *>(p.y1 + r * p.C + c + 4);
        u[0] = u0.x; u[1] = u0.y; u[2] = u0.z; u[3] = u0.w; u[4] = u1.x; u[5] = u1.y; u[6] = u1.z; u[7] = u1.w;
    }
    if (p.a1) {
        const float4 *pa = reinterpret_cast<const float4 *>(p.a1 + (int64_t)g * p.C + c);
        const float4 *pb = reinterpret_cast<const float4 *>(p.b1 + (int64_t)g * p.C + c);
        const float4 a0 = __ldg(pa), a1 = __ldg(pa + 1), b0 = __ldg(pb), b1 = __ldg(pb + 1);
        const float aa[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) u[j] = u[j] * aa[j] + bb[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaxf(u[j] + (yy[j] * a2[j] + b2[j]), 0.0f);
    if (p.out) {
        *reinterpret_cast<float4 *>(p.out + r * p.C + c) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4 *>(p.out + r * p.C + c + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
}

__global__ void __launch_bounds__(256) k_combine_split(const CombineParams p) {
    const int K8 = p.C >> 3;
    const int64_t total = p.plane_rows * K8;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t pr = i / K8;
        const int c = (int)(i - pr * K8) << 3;
        int64_t r = pr, w;
        if (p.perm) {
            const int64_t mt = pr >> 7;
            const int wl = (int)(pr & 127);
            const int64_t tile = mt / p.T;
            const int t = (int)(mt - tile * p.T);
            w = tile * 128 + wl;
            r = w * p.T + t;
        } else {
            w = pr / p.T;
        }
        float v[8];
        if (r < p.rows) {
            combine_values(p, r, w, c, v);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.0f;
        }
        if (p.hi) {
            uint4 h, l;
            ptx::split_op16x8(v, h, l);
            if (p.perm == 2) {
                const int64_t o = (((pr >> 7) * p.K8p + (c >> 3)) * 128 + (pr & 127)) * 8;
                *reinterpret_cast<uint4 *>(p.hi + o) = h;
                *reinterpret_cast<uint4 *>(p.lo + o) = l;
                if (c + 8 == p.C)
                    for (int g = K8; g < p.K8p; ++g) {
                        const int64_t oz = (((pr >> 7) * p.K8p + g) * 128 + (pr & 127)) * 8;
                        *reinterpret_cast<uint4 *>(p.hi + oz) = make_uint4(0, 0, 0, 0);
                        *reinterpret_cast<uint4 *>(p.lo + oz) = make_uint4(0, 0, 0, 0);
                    }
            } else {
                *reinterpret_cast<uint4 *>(p.hi + pr * p.C + c) = h;
                *reinterpret_cast<uint4 *>(p.lo + pr * p.C + c) = l;
            }
        }
    }
}

// Same tail written as the fused BLSTM kernel's k-group-major planes [tile][t][K8p][128 windows][8] (C <= 256), through a
// shared-memory transpose: a work unit = (tile, t, slab of 32 windows); reads run along the channels of a source row
// (lane = 8-channel group, 1 KB contiguous per row), stores along the windows of a plane row group (lane = window,
// 512 B contiguous).  Without it every 16-byte store of a warp lands 2 KB from the next one.
__global__ void __launch_bounds__(256) k_combine_kmajor(const CombineParams p) {
    __shared__ uint4 s_hi[32][33], s_lo[32][33];           // [channel group][window of the slab], padded: conflict-free both ways
    const int K8 = p.C >> 3;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t n_tiles = p.plane_rows / ((int64_t)128 * p.T);
    const int64_t units = n_tiles * p.T * 4;
    for (int64_t u = blockIdx.x; u < units; u += gridDim.x) {
        const int64_t tile = u / (p.T * 4);
        const int rem = (int)(u - tile * (p.T * 4));
        const int t = rem >> 2, wl0 = (rem & 3) * 32;
        for (int jj = 0; jj < 4; ++jj) {
            const int j = 4 * warp + jj;
            const int64_t w = tile * 128 + wl0 + j, r = w * p.T + t;
            if (lane < K8) {
                float v[8];
                if (r < p.rows) {
                    combine_values(p, r, w, lane << 3, v);
                } else {
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[k] = 0.0f;
                }
                ptx::split_op16x8(v, s_hi[lane][j], s_lo[lane][j]);
            }
        }
        __syncthreads();
        const int64_t base = ((tile * p.T + t) * p.K8p) * 128 + wl0 + lane;
        for (int cg = warp; cg < p.K8p; cg += 8) {
            const uint4 z = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4 *>(p.hi + (base + (int64_t)cg * 128) * 8) = cg < K8 ? s_hi[cg][lane] : z;
            *reinterpret_cast<uint4 *>(p.lo + (base + (int64_t)cg * 128) * 8) = cg < K8 ? s_lo[cg][lane] : z;
        }
        __syncthreads();
    }
}

// TF weights [taps][K][N] fp32 -> K-major fp16 planes [taps][Np][Kp] (zero padded)
__global__ void k_pack_weights(const float *__restrict__ w, int taps, int K, int N, int Kp, int Np,
                               op16_t *__restrict__ hi, op16_t *__restrict__ lo) {
    const int64_t total = (int64_t)taps * Np * Kp;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % Kp);
        const int n = (int)((i / Kp) % Np);
        const int tap = (int)(i / ((int64_t)Kp * Np));
        const float v = (k < K && n < N) ? w[((int64_t)tap * K + k) * N + n] : 0.0f;
        op16_t h, l;
        ptx::split_op16(v, h, l);
        hi[i] = h;
        lo[i] = l;
    }
}

// Interleaved-A form of the same weights (GemmWeights::ilv): operand column k' = 8 * (k / 4) + 4 * slot + k % 4 of the
// A plane is the hi (slot 0) or lo (slot 1) part of source column k, so B'[n][k'] = hi(W[k][n]) and B''[n][k'] = lo(W[k][n])
// in BOTH slots.
__global__ void k_pack_weights_ilv(const float *__restrict__ w, int Ksrc, int N, int Kp, int Np,
                                   op16_t *__restrict__ hi, op16_t *__restrict__ lo) {
    const int64_t total = (int64_t)Np * Kp;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int kk = (int)(i % Kp);
        const int n = (int)(i / Kp);
        const int k = 4 * (kk >> 3) + (kk & 3);
        const float v = (k < Ksrc && n < N) ? w[(int64_t)k * N + n] : 0.0f;
        op16_t h, l;
        ptx::split_op16(v, h, l);
        hi[i] = h;
        lo[i] = l;
    }
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess) p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// fp16 [rows][K] row-major plane, box {64, box_rows}, 128-byte swizzle, zero fill out of bounds
int make_map(CUtensorMap *m, const void *base, int64_t rows, int K, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(NTB_ECUDA, "cuTensorMapEncodeTiled entry point not found%s", "");
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(NTB_ECUDA, "cuTensorMapEncodeTiled failed (%s%lld) rows %lld", "", (long long)r, (long long)rows);
    return NTB_OK;
}

// fp32 [rows][ldo] row-major output, box {32 cols, 32 rows}, 128-byte swizzle (TMA-store epilogue)
int make_map_out(CUtensorMap *m, const float *base, int64_t rows, int N, int ldo) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(NTB_ECUDA, "cuTensorMapEncodeTiled entry point not found%s", "");
    cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ldo * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(NTB_ECUDA, "cuTensorMapEncodeTiled (output) failed (%s%lld) rows %lld", "", (long long)r, (long long)rows);
    return NTB_OK;
}

// fp32 [rows][K] activation, box {64 floats, box_rows}, no swizzle (the transform warps read it with plain loads)
int make_map_raw(CUtensorMap *m, const float *base, int64_t rows, int K, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(NTB_ECUDA, "cuTensorMapEncodeTiled entry point not found%s", "");
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 4};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(NTB_ECUDA, "cuTensorMapEncodeTiled (raw activation) failed (%s%lld) rows %lld", "", (long long)r, (long long)rows);
    return NTB_OK;
}

int pick_bn(int N) {
    // fewest N tiles first (A is re-read once per N tile), then least padding; BN % 16 == 0, <= 256
    int best = 0, best_tiles = 1 << 30, best_pad = 1 << 30;
    for (int bn = 16; bn <= 256; bn += 16) {
        const int tiles = (N + bn - 1) / bn, padn = tiles * bn - N;
        if (tiles < best_tiles || (tiles == best_tiles && padn < best_pad)) { best = bn; best_tiles = tiles; best_pad = padn; }
    }
    return best;
}

int sm_count() {
    static int n = [] {
        int dev = 0, v = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        return v;
    }();
    return n;
}

int launch_split(cudaStream_t st, const SplitParams &sp, int64_t *launches) {
    const int K8 = sp.K >> 3;
    if (!sp.perm && sp.T > 1 && K8 <= 256 && sp.rows % sp.T == 0 && (sp.pad || sp.aff_a)) {
        const int threads = (256 / K8) * K8;
        k_split_window<<<(unsigned)(sp.rows / sp.T), threads, 0, st>>>(sp);
        if (launches) ++*launches;
        NTB_LAUNCH_CHECK();
        return NTB_OK;
    }
    const int64_t total = sp.plane_rows * (sp.K >> 3);
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    k_split_op16<<<(unsigned)blocks, 256, 0, st>>>(sp);
    if (launches) ++*launches;
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

int gemm_cta_group() {
    // NTB_GEMM_CG=1 selects the single-CTA kernel (A/B measurement); default is the CTA-pair kernel
    static int cg = [] {
        const char *e = ntb_env("NTB_GEMM_CG");
        return (e && e[0] == '1') ? 1 : 2;
    }();
    return cg;
}

// tile / ring geometry of one GEMM launch (everything but the operands' addresses)
int fill_params(TcParams &p, const GemmWeights &w, int64_t plane_rows, bool allow_bstat = true, bool small_epi = false,
                bool xproj_tma = false) {
    const int cg = gemm_cta_group();
    p.m_tiles = (plane_rows + BM - 1) / BM;
    p.BN = w.BN; p.Np = w.Np; p.N = w.N; p.n_tiles = w.Np / w.BN;
    p.taps = w.taps; p.k_blocks = (w.K + BK - 1) / BK;
    p.k16_last = ((w.K - (p.k_blocks - 1) * BK) + 15) / 16;
    p.passes = w.single_pass ? 1 : (w.ilv ? 2 : 3);
    const int b_tile_bytes = (w.BN / cg) * BK * 2;
    int stage_bytes = p.passes == 2 ? A_TILE_BYTES + 2 * b_tile_bytes : (p.passes == 3 ? 2 : 1) * (A_TILE_BYTES + b_tile_bytes);   // per CTA
    // B-stationary (opt-in, NTB_BSTAT=1): measured on the LSTM projections (K <= 256, N = 800) it ties with the
    // streaming schedule (26.7 vs 27.0 ms per 65 536 windows) -- these GEMMs are bound by the 2.6 GB projected-gates
    // write, not by the operand feed from L2 -- so streaming stays the default
    static const bool no_bstat = ntb_env("NTB_BSTAT") == nullptr;
    const int bres = p.k_blocks * 2 * b_tile_bytes;
    // the projected-gates epilogue stores straight from registers: its staging area only holds the bias vector
    p.epi_bytes = small_epi && w.Np * 4 <= 4096 ? 4096 : EPI_STAGE_BYTES;
    if (xproj_tma && w.Np * 4 <= 4096) p.epi_bytes = EPI_STAGE_BYTES + 4096;     // store staging + bias vector
    const int budget = SMEM_BUDGET + EPI_STAGE_BYTES - p.epi_bytes;
    p.bstat = (allow_bstat && small_epi && !no_bstat && cg == 2 && p.passes == 3 && p.taps == 1 && p.n_tiles > 1 && p.n_tiles <= sm_count() / 2 &&
               bres + 3 * 2 * A_TILE_BYTES <= budget) ? 1 : 0;
    if (p.bstat) {
        stage_bytes = 2 * A_TILE_BYTES;
        p.stages = (budget - bres) / stage_bytes;
    } else {
        p.stages = budget / stage_bytes;
    }
    if (p.stages > 8) p.stages = 8;
    if (p.stages < 2) return fail(NTB_EINVAL, "tcgen05 GEMM: tile does not fit shared memory%s", "");
    return NTB_OK;
}

int do_launch(cudaStream_t st, const CUtensorMap &ma_hi, const CUtensorMap &ma_lo, const CUtensorMap &mb_hi,
              const CUtensorMap &mb_lo, const CUtensorMap &mo, const TcParams &p, int64_t *launches,
              const CUtensorMap *mo2 = nullptr) {
    const int cg = gemm_cta_group();
    const int b_tile_bytes = (p.BN / cg) * BK * 2;
    const int stage_bytes = p.bstat ? 2 * A_TILE_BYTES
                          : p.passes == 2 ? A_TILE_BYTES + 2 * b_tile_bytes : (p.passes == 3 ? 2 : 1) * (A_TILE_BYTES + b_tile_bytes);
    const size_t smem = (size_t)p.stages * stage_bytes + (p.bstat ? p.k_blocks * 2 * b_tile_bytes : 0) + 1024 + p.epi_bytes;
    static bool attr_set = false;
    if (!attr_set) {
        NTB_CUDA(cudaFuncSetAttribute(k_gemm_tc<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET + 1024 + EPI_STAGE_BYTES));
        NTB_CUDA(cudaFuncSetAttribute(k_gemm_tc<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET + 1024 + EPI_STAGE_BYTES));
        NTB_CUDA(cudaFuncSetAttribute(k_gemm_tc<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET + 1024 + EPI_STAGE_BYTES));
        attr_set = true;
    }
    if (cg == 1) {
        const int64_t tiles = p.m_tiles * p.n_tiles;
        const int grid = (int)(tiles < sm_count() ? tiles : sm_count());
        if (p.lstm_H > 0) return fail(NTB_EINVAL, "tcgen05 GEMM: the LSTM timestep mode needs the CTA-pair kernel%s", "");
        k_gemm_tc<1, false><<<grid, TC_THREADS_EW8, smem, st>>>(ma_hi, ma_lo, mb_hi, mb_lo, mo, mo2 ? *mo2 : mo, p);
    } else {
        const int64_t items = ((p.m_tiles + 1) / 2) * p.n_tiles;            // pairs of M tiles x N tiles
        int64_t pairs_hw = sm_count() / 2;
        if (p.bstat) pairs_hw = pairs_hw / p.n_tiles * p.n_tiles;          // every pair owns one N tile for the whole launch
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(2 * (items < pairs_hw ? items : pairs_hw)));
        cfg.blockDim = dim3(TC_THREADS_EW8);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        if (p.lstm_H > 0) NTB_CUDA(cudaLaunchKernelEx(&cfg, k_gemm_tc<2, true>, ma_hi, ma_lo, mb_hi, mb_lo, mo, mo2 ? *mo2 : mo, p));
        else NTB_CUDA(cudaLaunchKernelEx(&cfg, k_gemm_tc<2, false>, ma_hi, ma_lo, mb_hi, mb_lo, mo, mo2 ? *mo2 : mo, p));
    }
    if (launches) ++*launches;
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

// store view of the projected-gates buffer (layout: make_map_xproj below): one warp's [8 units][32 rows] float4 tile
int make_map_xproj_store(CUtensorMap *m, const float *xproj, int64_t units_total) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(NTB_ECUDA, "cuTensorMapEncodeTiled entry point not found%s", "");
    cuuint64_t dims[3] = {256, 2, (cuuint64_t)units_total};
    cuuint64_t strides[2] = {1024, 2048};
    cuuint32_t box[3] = {128, 1, 8};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float *>(xproj), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(NTB_ECUDA, "cuTensorMapEncodeTiled (xproj store) failed (%s%lld) units %lld", "", (long long)r, (long long)units_total);
    return NTB_OK;
}

int launch_gemm(cudaStream_t st, const GemmWeights &w, const GemmScratch &gs, int64_t plane_rows, int T, int pad,
                int64_t out_rows, const float *bias, float *out, int ldo, int64_t *launches, int xproj_H = 0,
                float2 *bn_part = nullptr, float *out2 = nullptr, float2 *bn_part2 = nullptr, unsigned long long *amax = nullptr) {
    const int cg = gemm_cta_group();
    CUtensorMap ma_hi, ma_lo, mb_hi, mb_lo;
    NTB_TRY(make_map(&ma_hi, gs.a_hi, plane_rows, w.K, BM));
    if (w.ilv) ma_lo = ma_hi;      // one interleaved plane [rows][2 Ksrc] spanning both scratch planes
    else NTB_TRY(make_map(&ma_lo, gs.a_lo, plane_rows, w.K, BM));
    NTB_TRY(make_map(&mb_hi, w.hi, (int64_t)w.taps * w.Np, w.Kp, w.BN / cg));
    NTB_TRY(make_map(&mb_lo, w.lo, (int64_t)w.taps * w.Np, w.Kp, w.BN / cg));
    TcParams p{};
    // opt-in (NTB_XPROJ_TMA=1): measured SLOWER than the register-direct stores (projection 28.4 vs 23.8 ms per 65 536
    // windows, profiles/r01_final_ablations.txt) -- the staging round trip costs more than the store path saves
    static const bool xtma_on = ntb_env("NTB_XPROJ_TMA") && ntb_env("NTB_XPROJ_TMA")[0] == '1';
    const bool xtma = xproj_H > 0 && xtma_on && w.Np * 4 <= 4096 && bias;
    NTB_TRY(fill_params(p, w, plane_rows, true, xproj_H > 0, xtma));
    p.xproj_tma = xtma ? 1 : 0;
    p.T = T; p.pad = pad; p.out_rows = out_rows; p.bias = bias; p.out = out; p.ldo = ldo;
    p.xproj_H = xproj_H;
    p.bn_part = bn_part;
    static const int abl = ntb_env("NTB_XPROJ_ABL") ? atoi(ntb_env("NTB_XPROJ_ABL")) : 0;
    p.abl = abl;
    // opt-in (NTB_GEMM_L2PF=1): measured slower (projection GEMM 7.49 -> 8.93 ms per 65 536 windows, conv stage +2 ms)
    static const int l2pf = ntb_env("NTB_GEMM_L2PF") ? atoi(ntb_env("NTB_GEMM_L2PF")) : 0;
    p.l2pf = l2pf;
    // coalesced TMA-store epilogue for plain row-major outputs (N tiles of whole 32-column chunks)
    p.tma_store = (!amax && !pad && xproj_H == 0 && w.BN % 32 == 0 && (ldo & 3) == 0 &&
                   (reinterpret_cast<uintptr_t>(out) & 15) == 0) ? 1 : 0;
    p.amax = amax;
    CUtensorMap mo = ma_hi;
    if (p.tma_store) NTB_TRY(make_map_out(&mo, out, out_rows, w.N, ldo));
    if (p.xproj_tma) NTB_TRY(make_map_xproj_store(&mo, out, 2 * p.m_tiles * (int64_t)xproj_H));
    if (out2) {
        // dual output: w holds [W1 | W2] along N (N = 2 * n_half, whole N tiles per half), both outputs [rows][n_half]
        if (!p.tma_store || p.n_tiles % 2 || w.N % 2 || (w.N / 2) % w.BN)
            return fail(NTB_EINVAL, "tcgen05 dual GEMM: unsupported tiling (N %s%lld)", "", (long long)w.N);
        CUtensorMap mo2;
        p.n_split = p.n_tiles / 2; p.n_half = w.N / 2; p.bn_part2 = bn_part2;
        NTB_TRY(make_map_out(&mo, out, out_rows, w.N / 2, ldo));
        NTB_TRY(make_map_out(&mo2, out2, out_rows, w.N / 2, ldo));
        return do_launch(st, ma_hi, ma_lo, mb_hi, mb_lo, mo, p, launches, &mo2);
    }
    return do_launch(st, ma_hi, ma_lo, mb_hi, mb_lo, mo, p, launches);
}

}  // namespace

// ------------------------------------------------------------------------------------------ interface
// Projected gates as a 3-D fp32 tensor {256 floats, 2, units_total}: one (dir, m_tile, unit) slab of 128 rows x float4
// is 2 KB = {256, 2}; a box {256, 2, 8} is one 16 KB chunk of 8 units (ntb_lstm_tc.cuh x ring), no swizzle.
int make_map_xproj(void *map_out, const float *xproj, int64_t units_total) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(NTB_ECUDA, "cuTensorMapEncodeTiled entry point not found%s", "");
    cuuint64_t dims[3] = {256, 2, (cuuint64_t)units_total};
    cuuint64_t strides[2] = {1024, 2048};
    cuuint32_t box[3] = {256, 2, 8};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn((CUtensorMap *)map_out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float *>(xproj), dims, strides,
                    box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(NTB_ECUDA, "cuTensorMapEncodeTiled (xproj) failed (%s%lld) units %lld", "", (long long)r, (long long)units_total);
    return NTB_OK;
}

int make_map_rows128(void *map_out, const void *base, int64_t rows, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(NTB_ECUDA, "cuTensorMapEncodeTiled entry point not found%s", "");
    cuuint64_t dims[2] = {128, (cuuint64_t)rows};
    cuuint64_t strides[1] = {128};
    cuuint32_t box[2] = {128, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn((CUtensorMap *)map_out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(NTB_ECUDA, "cuTensorMapEncodeTiled (128-byte rows) failed (%s%lld) rows %lld", "", (long long)r, (long long)rows);
    return NTB_OK;
}

int gemm_tc_split_kmajor(cudaStream_t st, const float *A, int lda, int64_t n_windows, int T, int K, op16_t *hi, op16_t *lo,
                         int64_t *launches) {
    if (K % 8 || (lda & 3) || (reinterpret_cast<uintptr_t>(A) & 15) || !hi || !lo)
        return fail(NTB_EINVAL, "k-group-major split: operand not supported (K %s%lld)", "", (long long)K);
    SplitParams sp{};
    sp.A = A; sp.lda = lda; sp.rows = n_windows * T; sp.K = K; sp.T = T; sp.relu = 0; sp.pad = 0; sp.perm = 2;
    sp.K8p = 2 * ((K + 15) / 16);
    sp.hi = hi; sp.lo = lo; sp.plane_rows = align_up(n_windows, 2 * BM) * T;
    return launch_split(st, sp, launches);
}

void gemm_tc_plan_weights(GemmWeights &w, int taps, int K, int N, Arena &ar, bool ilv) {
    w = GemmWeights();
    w.taps = taps; w.K = K; w.N = N;
    if (ilv && taps == 1 && K % 4 == 0) { w.ilv = true; w.Ksrc = K; w.K = 2 * K; K = 2 * K; }
    // the tensor-core path needs 16-byte rows in the A planes (K % 8) and a worthwhile shape;
    // Cin = 1 convs, len_x = 12 / H = 5 event shapes and the tiny CTC readout stay on the SIMT path
    if (K % 8 != 0 || K < 64 || N < 32) return;
    w.Kp = (int)align_up(K, BK);
    w.BN = pick_bn(N);
    w.Np = (int)align_up(N, w.BN);
    const int64_t n = (int64_t)taps * w.Np * w.Kp;
    w.hi = ar.take<op16_t>(n);
    w.lo = ar.take<op16_t>(n);
    w.valid = true;
}

int gemm_tc_pack_weights(cudaStream_t st, const float *w_kn, GemmWeights &w) {
    if (!w.valid) return NTB_OK;
    const int64_t total = (int64_t)w.taps * w.Np * w.Kp;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 1024) blocks = 1024;
    if (w.ilv) k_pack_weights_ilv<<<blocks, 256, 0, st>>>(w_kn, w.Ksrc, w.N, w.Kp, w.Np, w.hi, w.lo);
    else k_pack_weights<<<blocks, 256, 0, st>>>(w_kn, w.taps, w.K, w.N, w.Kp, w.Np, w.hi, w.lo);
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

void gemm_tc_plan_scratch(GemmScratch &gs, int64_t rows, int T, int maxk, Arena &ar, int64_t min_rows) {
    gs = GemmScratch();
    if (maxk < 64 || rows <= 0) return;
    // padded layout upper bound: (T + 2) / T more rows, rounded up to whole tiles
    const int64_t windows = T > 0 ? (rows + T - 1) / T : 0;
    int64_t cap = rows + 2 * windows;
    cap = align_up(cap, BM) + BM;
    if (cap < min_rows) cap = min_rows;
    gs.rows_cap = cap;
    gs.kp_cap = (int)align_up(maxk, 8);
    gs.a_hi = ar.take<op16_t>(cap * gs.kp_cap);
    gs.a_lo = ar.take<op16_t>(cap * gs.kp_cap);
    gs.valid = true;
}

int gemm_tc_run(cudaStream_t st, const float *A, int lda, int64_t rows, const GemmWeights &w, const float *bias,
                float *out, int ldo, const GemmScratch &gs, int64_t *launches) {
    if (!w.valid || !gs.valid || rows > gs.rows_cap || w.K > gs.kp_cap || (lda & 3) ||
        (reinterpret_cast<uintptr_t>(A) & 15))
        return fail(NTB_EINVAL, "tcgen05 GEMM: operand not supported (rows %s%lld, K %lld)", "", (long long)rows,
                    (long long)w.K);
    if (w.taps != 1) return fail(NTB_EINVAL, "gemm_tc_run: taps must be 1%s", "");
    SplitParams sp{};
    sp.A = A; sp.lda = lda; sp.rows = rows; sp.K = w.K; sp.T = 1; sp.relu = 0; sp.pad = 0;
    sp.hi = gs.a_hi; sp.lo = w.single_pass ? nullptr : gs.a_lo; sp.plane_rows = rows;
    NTB_TRY(launch_split(st, sp, launches));
    return launch_gemm(st, w, gs, rows, 1, 0, rows, bias, out, ldo, launches);
}

int gemm_tc_run_argmax(cudaStream_t st, const float *A, int lda, int64_t rows, const GemmWeights &w, const float *bias,
                       unsigned long long *keys, const GemmScratch &gs, int64_t *launches) {
    if (!w.valid || !gs.valid || rows > gs.rows_cap || w.K > gs.kp_cap || (lda & 3) || w.taps != 1 || w.single_pass || w.ilv ||
        (reinterpret_cast<uintptr_t>(A) & 15) || !keys)
        return fail(NTB_EINVAL, "tcgen05 GEMM (argmax epilogue): operand not supported (rows %s%lld, K %lld)", "", (long long)rows,
                    (long long)w.K);
    NTB_CUDA(cudaMemsetAsync(keys, 0, sizeof(unsigned long long) * (size_t)rows, st));
    SplitParams sp{};
    sp.A = A; sp.lda = lda; sp.rows = rows; sp.K = w.K; sp.T = 1; sp.relu = 0; sp.pad = 0;
    sp.hi = gs.a_hi; sp.lo = gs.a_lo; sp.plane_rows = rows;
    NTB_TRY(launch_split(st, sp, launches));
    return launch_gemm(st, w, gs, rows, 1, 0, rows, bias, nullptr, 0, launches, 0, nullptr, nullptr, nullptr, keys);
}

int gemm_tc_run_xproj(cudaStream_t st, const float *A, int lda, int64_t n_windows, int T, const GemmWeights &w,
                      const float *bias, float *xproj, int H, const GemmScratch &gs, int64_t *launches,
                      bool planes_ready, void (*before_gemm)(void *), void *hook_arg) {
    const int64_t n_tiles = (n_windows + BM - 1) / BM;
    const int64_t plane_rows = n_tiles * BM * T;
    if (w.ilv) {
        // the previous recurrence wrote the interleaved plane into the scratch region (both planes, contiguous)
        if (!w.valid || !gs.valid || !planes_ready || !gs.contiguous() || w.N != 8 * H ||
            plane_rows * (int64_t)w.K > 2 * gs.rows_cap * (int64_t)gs.kp_cap)
            return fail(NTB_EINVAL, "tcgen05 projection: interleaved operand not available (windows %s%lld, K %lld)", "",
                        (long long)n_windows, (long long)w.K);
        if (before_gemm) before_gemm(hook_arg);
        return launch_gemm(st, w, gs, plane_rows, 1, 0, plane_rows, bias, xproj, 0, launches, H);
    }
    if (!w.valid || !gs.valid || plane_rows > gs.rows_cap || w.K > gs.kp_cap || (lda & 3) || w.taps != 1 ||
        w.N != 8 * H || (reinterpret_cast<uintptr_t>(A) & 15))
        return fail(NTB_EINVAL, "tcgen05 projection: operand not supported (windows %s%lld, K %lld)", "",
                    (long long)n_windows, (long long)w.K);
    SplitParams sp{};
    sp.A = A; sp.lda = lda; sp.rows = n_windows * T; sp.K = w.K; sp.T = T; sp.relu = 0; sp.pad = 0; sp.perm = 1;
    sp.hi = gs.a_hi; sp.lo = w.single_pass ? nullptr : gs.a_lo; sp.plane_rows = plane_rows;
    if (!planes_ready) NTB_TRY(launch_split(st, sp, launches));   // else: the producer wrote the planes itself
    if (before_gemm) before_gemm(hook_arg);                       // profiling mark: the GEMM kernel alone
    return launch_gemm(st, w, gs, plane_rows, 1, 0, plane_rows, bias, xproj, 0, launches, H);
}

int gemm_tc_conv(cudaStream_t st, const float *A, int K, int64_t rows, int T, const float *aff_a,
                 const float *aff_b, const int32_t *win_group, int relu_in, const GemmWeights &w, float *out,
                 int ldo, const GemmScratch &gs, int64_t *launches, bool reuse_planes, bool expand, float2 *bn_part) {
    const int pad = w.taps == 3;
    const int64_t windows = rows / T;
    const int64_t plane_rows = pad ? windows * (T + 2) : rows;
    if (!w.valid || !gs.valid || plane_rows > gs.rows_cap || K != w.K || w.K > gs.kp_cap || rows % T)
        return fail(NTB_EINVAL, "tcgen05 conv: operand not supported (rows %s%lld, K %lld)", "", (long long)rows,
                    (long long)K);
    SplitParams sp{};
    sp.A = A; sp.lda = K; sp.rows = rows; sp.K = K; sp.aff_a = aff_a; sp.aff_b = aff_b; sp.win_group = win_group;
    sp.T = T; sp.relu = relu_in; sp.pad = pad; sp.expand = expand ? 1 : 0;
    if (expand) sp.lda = 1;
    sp.hi = gs.a_hi; sp.lo = w.single_pass ? nullptr : gs.a_lo; sp.plane_rows = plane_rows;
    if (!reuse_planes) NTB_TRY(launch_split(st, sp, launches));   // else: the previous conv split the same input
    if (bn_part && T < 32) return fail(NTB_EINVAL, "tcgen05 conv: fused BN statistics need T >= 32%s", "");
    return launch_gemm(st, w, gs, plane_rows, T, pad, rows, nullptr, out, ldo, launches, 0, bn_part);
}

int gemm_tc_conv_dual(cudaStream_t st, const float *A, int K, int64_t rows, int T, const GemmWeights &w12, float *out1,
                      float *out2, int C, const GemmScratch &gs, int64_t *launches, bool reuse_planes, float2 *part1,
                      float2 *part2) {
    if (!w12.valid || !gs.valid || w12.taps != 1 || w12.N != 2 * C || rows > gs.rows_cap || K != w12.K || K > gs.kp_cap)
        return fail(NTB_EINVAL, "tcgen05 dual conv: operand not supported (rows %s%lld, K %lld)", "", (long long)rows,
                    (long long)K);
    SplitParams sp{};
    sp.A = A; sp.lda = K; sp.rows = rows; sp.K = K; sp.T = T; sp.relu = 0; sp.pad = 0;
    sp.hi = gs.a_hi; sp.lo = gs.a_lo; sp.plane_rows = rows;
    if (!reuse_planes) NTB_TRY(launch_split(st, sp, launches));
    if ((part1 || part2) && T < 32) return fail(NTB_EINVAL, "tcgen05 conv: fused BN statistics need T >= 32%s", "");
    return launch_gemm(st, w12, gs, rows, T, 0, rows, nullptr, out1, C, launches, 0, part1, out2, part2);
}

bool gemm_tc_conv_fused_supported(const GemmWeights &w, int K, int T, int ldo) {
    static const bool off = ntb_env("NTB_NO_FUSED_CONV") != nullptr;      // A/B measurement switch
    return !off && w.valid && !w.single_pass && gemm_cta_group() == 2 && w.K == K && K % BK == 0 && w.Np == w.N &&
           w.BN == w.N && w.N % 32 == 0 && w.N <= 256 && (w.taps == 1 || w.taps == 3) && T >= 32 && (ldo & 3) == 0;
}

int gemm_tc_conv_fused(cudaStream_t st, const float *y, int K, int64_t rows, int T, const float *aff_a,
                       const float *aff_b, const int32_t *win_group, int relu_in, const GemmWeights &w, float *out,
                       int ldo, int64_t *launches, float2 *bn_part, const float *x_scalar) {
    if (x_scalar) y = out;      // no raw activation exists: the tensor map only needs a valid base, it is never used
    if (!gemm_tc_conv_fused_supported(w, K, T, ldo) || rows % T || (reinterpret_cast<uintptr_t>(y) & 15) ||
        (reinterpret_cast<uintptr_t>(out) & 15) || (aff_a && !win_group) || (x_scalar && !aff_a))
        return fail(NTB_EINVAL, "tcgen05 fused conv: operand not supported (rows %s%lld, K %lld)", "", (long long)rows,
                    (long long)K);
    CUtensorMap m_raw, mb_hi, mb_lo, mo;
    NTB_TRY(make_map_raw(&m_raw, y, rows, K, w.taps == 3 ? BM + 2 : BM));
    NTB_TRY(make_map(&mb_hi, w.hi, (int64_t)w.taps * w.Np, w.Kp, w.BN / 2));
    NTB_TRY(make_map(&mb_lo, w.lo, (int64_t)w.taps * w.Np, w.Kp, w.BN / 2));
    NTB_TRY(make_map_out(&mo, out, rows, w.N, ldo));
    ConvFusedParams p{};
    p.m_tiles = (rows + BM - 1) / BM; p.rows = rows; p.N = w.N; p.K = K; p.taps = w.taps; p.T = T;
    p.k_blocks = K / BK; p.relu = relu_in; p.aff_a = aff_a; p.aff_b = aff_b; p.win_group = win_group; p.bn_part = bn_part;
    p.xs = x_scalar;
    // opt-in: measured neutral to slightly slower (conv stage 56.5 vs 55.8 ms per 65 536 windows)
    static const int cf_l2pf = ntb_env("NTB_CF_L2PF") ? atoi(ntb_env("NTB_CF_L2PF")) : 0;
    p.l2pf = cf_l2pf;
    static const int cf_abl = ntb_env("NTB_CF_ABL") ? atoi(ntb_env("NTB_CF_ABL")) : 0;
    p.abl = cf_abl;
    const size_t smem = (size_t)CF_OP_STAGES * (2 * A_TILE_BYTES + 2 * (w.N / 2) * BK * 2) +
                        (size_t)CF_RAW_STAGES * CF_RAW_STAGE + CF_EPI_BYTES + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        NTB_CUDA(cudaFuncSetAttribute(k_conv_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048));
        attr_set = true;
    }
    const int64_t items = (p.m_tiles + 1) / 2, pairs_hw = sm_count() / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * (items < pairs_hw ? items : pairs_hw)));
    cfg.blockDim = dim3(CF_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    NTB_CUDA(cudaLaunchKernelEx(&cfg, k_conv_fused, m_raw, mb_hi, mb_lo, mo, p));
    if (launches) ++*launches;
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

int64_t gemm_tc_bn_part_elems(const GemmScratch &gs, int C) {
    return gs.valid ? (gs.rows_cap / 32 + 1) * 2 * (int64_t)C : 0;   // double2 records
}

int gemm_tc_combine(cudaStream_t st, const float *y1, bool y1_scalar, const float *a1, const float *b1,
                    const float *y2, const float *a2, const float *b2, const int32_t *win_group, int T, int C,
                    int64_t rows, float *out, const GemmScratch *planes, int layout, int64_t *launches,
                    op16_t *kmajor_hi, op16_t *kmajor_lo) {
    if (C % 8 || rows % T) return fail(NTB_EINVAL, "combine: C %% 8 and whole windows required%s", "");
    CombineParams p{};
    p.y1 = y1; p.y1_scalar = y1_scalar ? 1 : 0; p.a1 = a1; p.b1 = b1; p.y2 = y2; p.a2 = a2; p.b2 = b2;
    p.win_group = win_group; p.T = T; p.C = C; p.rows = rows; p.out = out;
    p.plane_rows = rows;
    if (layout == 3) {
        if (!kmajor_hi || !kmajor_lo) return fail(NTB_EINVAL, "combine: k-group-major planes missing%s", "");
        p.perm = 2; p.K8p = 2 * ((C + 15) / 16);
        p.plane_rows = align_up(rows / T, 2 * BM) * T;
        p.hi = kmajor_hi; p.lo = kmajor_lo;
    } else if (planes && layout) {
        p.perm = layout == 2;
        if (p.perm) p.plane_rows = align_up(rows / T, BM) * T;
        if (!planes->valid || p.plane_rows > planes->rows_cap || C > planes->kp_cap)
            return fail(NTB_EINVAL, "combine: operand planes too small (rows %s%lld)", "", (long long)p.plane_rows);
        p.hi = planes->a_hi; p.lo = planes->a_lo;
    }
    const int64_t total = p.plane_rows * (C >> 3);
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (p.perm == 2 && C <= 256) {
        int64_t units = p.plane_rows / 128 * 4;
        const int64_t ucap = (int64_t)sm_count() * 8;
        k_combine_kmajor<<<(unsigned)(units < ucap ? (units > 0 ? units : 1) : ucap), 256, 0, st>>>(p);
    } else
    k_combine_split<<<(unsigned)blocks, 256, 0, st>>>(p);
    if (launches) ++*launches;
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

// ------------------------------------------------------------------------------------------
// BLSTM recurrence for hidden sizes the persistent kernel cannot hold (4H + H/2 > 512 TMEM columns, W_hh
// larger than one SM's shared memory): ONE tcgen05 GEMM launch per timestep,
//   z_t = h_{t-1} . W_hh   for every window and both directions (M = 2 * windows, N = 4H, K = H, fp16x3),
// whose epilogue adds the projected gates, applies the fused cell update and writes h_t three ways: fp32 layer
// output, and the fp16 hi/lo operand planes of the next step (ping-pong buffers: other CTAs still read h_{t-1}).
// W_hh hi/lo (1 MB at H = 256) stays L2-resident across the T launches; c lives in HBM as fp32.
// ------------------------------------------------------------------------------------------
__global__ void k_pack_whh_kn(const float *__restrict__ kernel_fw, const float *__restrict__ kernel_bw, int D, int H,
                              float *__restrict__ out) {   // out [2][H][4H], column u*4 + g, pre-scaled
    const int N = 4 * H;
    const int64_t total = 2LL * H * N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int n = (int)(i % N), k = (int)((i / N) % H), dir = (int)(i / ((int64_t)N * H));
        const int u = n >> 2, g = n & 3;
        const float *kern = dir ? kernel_bw : kernel_fw;
        out[i] = (g == 1 ? LSTM_PRESCALE_J : LSTM_PRESCALE_IFO) * kern[(int64_t)(D + k) * N + g * H + u];
    }
}

bool gemm_tc_lstm_supported(int H) { return H >= 64 && H % 8 == 0; }

void gemm_tc_plan_whh(GemmWeights &w, int H, Arena &ar, float **staging) {
    gemm_tc_plan_weights(w, 2, H, 4 * H, ar);        // "taps" = the two directions
    *staging = w.valid ? ar.take<float>(2LL * H * 4 * H) : nullptr;
}

int gemm_tc_pack_whh(cudaStream_t st, const float *kernel_fw, const float *kernel_bw, int D, int H, float *staging,
                     GemmWeights &w) {
    if (!w.valid) return NTB_OK;
    k_pack_whh_kn<<<512, 256, 0, st>>>(kernel_fw, kernel_bw, D, H, staging);
    NTB_LAUNCH_CHECK();
    return gemm_tc_pack_weights(st, staging, w);
}

int64_t gemm_tc_lstm_state_floats(int64_t n_windows, int H) { return 2 * align_up(n_windows, BM) * H; }

// [W_ih ; W_hh] along K for the step GEMM that carries the input projection: out [2 dirs][D + H][4H], column u*4 + g,
// pre-scaled; wih is k_repack_lstm's [D][8H] (prescale = 1), rows D.. come from the TF kernels like k_pack_whh_kn
// (D = the layer's input width, Dp = D padded to the k-block of 64: the padding rows are zero)
__global__ void k_pack_wcat_kn(const float *__restrict__ wih, const float *__restrict__ kernel_fw,
                               const float *__restrict__ kernel_bw, int D, int Dp, int H, float *__restrict__ out) {
    const int N = 4 * H, K = Dp + H;
    const int64_t total = 2LL * K * N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int n = (int)(i % N), k = (int)((i / N) % K), dir = (int)(i / ((int64_t)N * K));
        if (k < Dp) { out[i] = k < D ? wih[(int64_t)k * 2 * N + dir * N + n] : 0.0f; continue; }
        const int u = n >> 2, g = n & 3;
        const float *kern = dir ? kernel_bw : kernel_fw;
        out[i] = (g == 1 ? LSTM_PRESCALE_J : LSTM_PRESCALE_IFO) * kern[(int64_t)(D + k - Dp) * N + g * H + u];
    }
}

// any input width that is a multiple of 4 floats: the planes and W_ih are zero-padded to whole 64-wide k-blocks
bool gemm_tc_lstm_fused_supported(int H, int D) { return gemm_tc_lstm_supported(H) && D % 4 == 0 && D >= 4 && H % BK == 0; }
int gemm_tc_lstm_fused_dpad(int D) { return (int)align_up(D, BK); }

void gemm_tc_plan_wcat(GemmWeights &w, int H, int D, Arena &ar, float **staging) {
    const int Dp = gemm_tc_lstm_fused_dpad(D);
    gemm_tc_plan_weights(w, 2, Dp + H, 4 * H, ar);
    *staging = w.valid ? ar.take<float>(2LL * (Dp + H) * 4 * H) : nullptr;
}

int gemm_tc_pack_wcat(cudaStream_t st, const float *wih, const float *kernel_fw, const float *kernel_bw, int D, int H,
                      float *staging, GemmWeights &w) {
    if (!w.valid) return NTB_OK;
    k_pack_wcat_kn<<<1024, 256, 0, st>>>(wih, kernel_fw, kernel_bw, D, gemm_tc_lstm_fused_dpad(D), H, staging);
    NTB_LAUNCH_CHECK();
    return gemm_tc_pack_weights(st, staging, w);
}

int gemm_tc_split_timemajor(cudaStream_t st, const float *A, int lda, int64_t n_windows, int T, int Ksrc, int K, op16_t *hi,
                            op16_t *lo, int64_t *launches) {
    if (K % 8 || Ksrc % 4 || Ksrc > K || (lda & 3) || (reinterpret_cast<uintptr_t>(A) & 15) || !hi || !lo)
        return fail(NTB_EINVAL, "time-major split: operand not supported (K %s%lld)", "", (long long)K);
    SplitParams sp{};
    sp.A = A; sp.lda = lda; sp.rows = n_windows * T; sp.K = K; sp.Ksrc = Ksrc; sp.T = T; sp.relu = 0; sp.pad = 0; sp.perm = 1;
    sp.hi = hi; sp.lo = lo; sp.plane_rows = align_up(n_windows, BM) * T;
    return launch_split(st, sp, launches);
}

// 3-D view of a k-group-major operand plane [K/8][rows][8] fp16: box {8, 128 rows, 8 groups} = one [128 x 64] A tile
static int make_map_kmajor(CUtensorMap *m, const void *base, int64_t rows, int K) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(NTB_ECUDA, "cuTensorMapEncodeTiled entry point not found%s", "");
    cuuint64_t dims[3] = {8, (cuuint64_t)rows, (cuuint64_t)(K / 8)};
    cuuint64_t strides[2] = {16, (cuuint64_t)rows * 16};
    cuuint32_t box[3] = {8, (cuuint32_t)BM, (cuuint32_t)(BK / 8)};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void *>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(NTB_ECUDA, "cuTensorMapEncodeTiled (k-group-major plane) failed (%s%lld) rows %lld", "", (long long)r, (long long)rows);
    return NTB_OK;
}

// xproj: the unit-major projected-gates buffer gemm_tc_run_xproj writes ([dir][128-window tile x T + t][unit][128] float4)
int gemm_tc_lstm_layer(cudaStream_t st, const GemmWeights &whh, const float *xproj, int ldx, const int32_t *lens,
                       int64_t n_windows, int T, int H, float *cstate, float *out, int ldo, const GemmScratch &gs,
                       int64_t *launches, int D_fused, const op16_t *x_hi, const op16_t *x_lo, const float *bias,
                       op16_t *next_hi, op16_t *next_lo) {
    const int cg = gemm_cta_group();
    const int Kw = H + D_fused;                                      // K of the weight planes: [W_ih ;] W_hh
    if (D_fused && (!x_hi || !x_lo || !bias || D_fused % BK || H % BK))
        return fail(NTB_EINVAL, "tcgen05 LSTM step with the projection inside: operand not supported (D %s%lld)", "", (long long)D_fused);
    const int64_t Mp = align_up(n_windows, 2 * BM);                 // a CTA pair never straddles the directions
    const int64_t plane_elems = 2 * Mp * H;                          // one parity: [2 dirs][Mp][H]
    if (!whh.valid || !gs.valid || whh.K != Kw || whh.N != 4 * H || whh.BN % 16 || cg != 2 ||
        2 * plane_elems > gs.rows_cap * (int64_t)gs.kp_cap || (ldo & 3) || H % 8)
        return fail(NTB_EINVAL, "tcgen05 LSTM step: operand not supported (windows %s%lld, H %lld)", "",
                    (long long)n_windows, (long long)H);
    // h_0 = 0 (both parities: pad rows are never written), c_0 = 0
    NTB_CUDA(cudaMemsetAsync(gs.a_hi, 0, 2 * plane_elems * sizeof(op16_t), st));
    NTB_CUDA(cudaMemsetAsync(gs.a_lo, 0, 2 * plane_elems * sizeof(op16_t), st));
    NTB_CUDA(cudaMemsetAsync(cstate, 0, gemm_tc_lstm_state_floats(n_windows, H) * sizeof(float), st));
    CUtensorMap ma_hi[2], ma_lo[2], mb_hi, mb_lo;
    for (int par = 0; par < 2; ++par) {
        NTB_TRY(make_map_kmajor(&ma_hi[par], gs.a_hi + par * plane_elems, 2 * Mp, H));
        NTB_TRY(make_map_kmajor(&ma_lo[par], gs.a_lo + par * plane_elems, 2 * Mp, H));
    }
    NTB_TRY(make_map(&mb_hi, whh.hi, 2LL * whh.Np, whh.Kp, whh.BN / cg));
    NTB_TRY(make_map(&mb_lo, whh.lo, 2LL * whh.Np, whh.Kp, whh.BN / cg));
    GemmWeights w1 = whh;
    w1.taps = 1;                                                     // the second "tap" is the bw direction's W_hh
    TcParams p{};
    NTB_TRY(fill_params(p, w1, 2 * Mp, false));   // the weight tile depends on the direction of the M tile
    p.T = T; p.pad = 0; p.out_rows = n_windows * T; p.bias = nullptr; p.out = out; p.ldo = ldo;
    p.lstm_H = H; p.lstm_mt_dir = (int)(Mp / BM); p.lstm_n = n_windows; p.lstm_Mp = Mp;
    p.lstm_x = xproj; p.lstm_ldx = ldx; p.lstm_c = cstate; p.lstm_lens = lens;
    p.lstm_xmt = (align_up(n_windows, BM) / BM) * T; p.a_kmajor = 1;
    CUtensorMap mx_hi = ma_hi[0], mx_lo = ma_lo[0];
    if (D_fused) {
        const int64_t xrows = align_up(n_windows, BM) * T;           // time-major planes: row (tile T + t) 128 + window
        NTB_TRY(make_map(&mx_hi, x_hi, xrows, D_fused, BM));
        NTB_TRY(make_map(&mx_lo, x_lo, xrows, D_fused, BM));
        p.lstm_kbx = D_fused / BK; p.lstm_bias = bias; p.lstm_x = nullptr;
    }
    if (next_hi && next_lo) { p.lstm_nx_hi = next_hi; p.lstm_nx_lo = next_lo; p.lstm_nx_K = 2 * H; }
    for (int s = 0; s < T; ++s) {
        const int rd = s & 1, wr = rd ^ 1;
        p.lstm_s = s;
        p.lstm_hi = gs.a_hi + wr * plane_elems;
        p.lstm_lo = gs.a_lo + wr * plane_elems;
        NTB_TRY(do_launch(st, ma_hi[rd], ma_lo[rd], mb_hi, mb_lo, mx_hi, p, launches, &mx_lo));
    }
    return NTB_OK;
}

}  // namespace ntb
