// tcgen05 / TMA / TMEM GEMM (fp16x3 split operands) -- interface; implementation in ntb_gemm_tc.cu.
#pragma once
#include "ntb_common.cuh"

namespace ntb {

// K-major fp16 hi/lo planes of a [taps][K][N] fp32 weight tensor, padded for the MMA tiles.
struct GemmWeights {
    bool valid = false;
    bool single_pass = false;   // hi*hi only (single 16-bit pass; measurement / ablation only)
    // Interleaved-A form (LSTM projections fed straight by the previous recurrence, ntb_lstm_tc.cuh): the A operand is ONE
    // plane [rows][2*Ksrc] whose 8-element groups are (4 x hi, 4 x lo) of 4 consecutive source columns; K = 2*Ksrc here,
    // plane `hi` holds W_hi in both slots and plane `lo` holds W_lo in both slots, so two passes A'.B' + A'.B'' give
    // a_hi.W_hi + a_lo.W_hi + a_hi.W_lo (+ a_lo.W_lo) -- the fp16x3 product without a separate operand-split pass.
    bool ilv = false;
    int Ksrc = 0;
    int taps = 1, K = 0, N = 0, Kp = 0, Np = 0, BN = 0;
    op16_t *hi = nullptr, *lo = nullptr;   // [taps][Np][Kp], Kp = K rounded up to 64
};
// fp16 hi/lo planes of the A operand for one GEMM (rewritten by every call).
struct GemmScratch {
    bool valid = false;
    int64_t rows_cap = 0;
    int kp_cap = 0;
    op16_t *a_hi = nullptr, *a_lo = nullptr;   // [plane rows][K] of the GEMM in flight
    // the two planes are adjacent in the workspace: an interleaved-A operand [rows][2K] uses them as one region
    bool contiguous() const { return a_lo == a_hi + rows_cap * (int64_t)kp_cap; }
};

void gemm_tc_plan_weights(GemmWeights &w, int taps, int K, int N, Arena &ar, bool ilv = false);
int gemm_tc_pack_weights(cudaStream_t st, const float *w_kn, GemmWeights &w);
void gemm_tc_plan_scratch(GemmScratch &gs, int64_t rows, int T, int maxk, Arena &ar, int64_t min_rows = 0);
int gemm_tc_run(cudaStream_t st, const float *A, int lda, int64_t rows, const GemmWeights &w, const float *bias,
                float *out, int ldo, const GemmScratch &gs, int64_t *launches);
// Same GEMM with an argmax epilogue instead of an output: keys[row] = max over columns of (orderable(out + bias) << 32 |
// ~column) -- the row's largest value and, among equals, its FIRST column (tf.argmax); the logits are never written.
// keys [rows] uint64 (zeroed by the call); column = 0xffffffff - (key & 0xffffffff), key 0 = no finite value (all NaN).
int gemm_tc_run_argmax(cudaStream_t st, const float *A, int lda, int64_t rows, const GemmWeights &w, const float *bias,
                       unsigned long long *keys, const GemmScratch &gs, int64_t *launches);
// LSTM input projection for the tensor-core recurrence: A rows are permuted to time-major tiles
// ([tile][t][128 windows]) and the output is the unit-major projected-gates layout of ntb_lstm_tc.cuh
// (xproj must hold ceil(n_windows/128)*128*T*8H floats).
int gemm_tc_run_xproj(cudaStream_t st, const float *A, int lda, int64_t n_windows, int T, const GemmWeights &w,
                      const float *bias, float *xproj, int H, const GemmScratch &gs, int64_t *launches,
                      bool planes_ready = false, void (*before_gemm)(void *) = nullptr, void *hook_arg = nullptr);
// CUtensorMap (128 bytes, 64-byte aligned) over the projected-gates buffer for the recurrence's TMA ring.
int make_map_xproj(void *map_out, const float *xproj, int64_t units_total);
// CUtensorMap over a byte region seen as [rows][128 bytes] (no swizzle), box = {128 B, box_rows}: the contiguous operand
// images of the fused BLSTM kernel (ntb_blstm_fused.cuh) are loaded as runs of whole 128-byte rows.
int make_map_rows128(void *map_out, const void *base, int64_t rows, int box_rows);
// fp32 activation [n_windows*T][lda] -> the fused BLSTM kernel's operand planes (fp16 hi / lo):
// [tile of 128 windows][t][K8 groups of 8 channels][128 windows][8], K8 = 2*ceil(K/16) (pad groups zero), whole tile
// PAIRS (windows padded to 256 with zeros).  Plane capacity: blstm_plane_elems().
int gemm_tc_split_kmajor(cudaStream_t st, const float *A, int lda, int64_t n_windows, int T, int K, op16_t *hi, op16_t *lo,
                         int64_t *launches);
static inline int64_t blstm_plane_elems(int64_t n_windows, int T, int K) {
    return align_up(n_windows, 256) * T * align_up(K, 16);
}
int gemm_tc_conv(cudaStream_t st, const float *A, int K, int64_t rows, int T, const float *aff_a,
                 const float *aff_b, const int32_t *win_group, int relu_in, const GemmWeights &w, float *out,
                 int ldo, const GemmScratch &gs, int64_t *launches, bool reuse_planes = false, bool expand = false,
                 float2 *bn_part = nullptr);
// conv1 and conv2a of a residual layer (two 1x1 convs of the same input) as ONE launch: w12 = [W1 | W2] along N
// (gemm_tc_plan_weights with N = 2C, packed from a [K][2C] staging matrix), outputs out1 / out2 [rows][C], BN partials
// part1 (may be null: branch1 without BN) / part2.  The A planes are streamed from DRAM once instead of twice.
int gemm_tc_conv_dual(cudaStream_t st, const float *A, int K, int64_t rows, int T, const GemmWeights &w12, float *out1,
                      float *out2, int C, const GemmScratch &gs, int64_t *launches, bool reuse_planes, float2 *part1,
                      float2 *part2);
// bn_part (optional): the GEMM epilogue also emits per-(32-row block, window segment, channel) (sum, sum of
// squares) partials of the conv output -- [(block * 2 + seg) * C + c], block = plane row / 32, plane rows per
// window = T (1x1) or T + 2 (1x3) -- which k_bn_finish folds per BN group (no extra pass over the output).
int64_t gemm_tc_bn_part_elems(const GemmScratch &gs, int C);
// Same conv with the operand transform (BN affine + ReLU + fp16 hi/lo split of the producer's raw fp32 output y)
// done by transform warps inside the GEMM: no operand planes, no padded layout, output rows = input rows.
bool gemm_tc_conv_fused_supported(const GemmWeights &w, int K, int T, int ldo);
int gemm_tc_conv_fused(cudaStream_t st, const float *y, int K, int64_t rows, int T, const float *aff_a,
                       const float *aff_b, const int32_t *win_group, int relu_in, const GemmWeights &w, float *out,
                       int ldo, int64_t *launches, float2 *bn_part, const float *x_scalar = nullptr);
// x_scalar != null: the producer was a conv of ONE input channel folded into its BN affine (k_k1_affine): the operand is
// relu(x_scalar[row] * aff_a[g][c] + aff_b[g][c]), made by the transform warps from 4 bytes per row; `y` is ignored.
// residual_layer tail relu(bn?(branch1) + bn(conv2c)) written as fp32 (`out`, may be null) and/or straight into the
// operand planes of the next GEMM: layout 0 = none, 1 = plain rows (next 1x1 convs), 2 = time-major tiles (projection),
// 3 = the fused BLSTM kernel's k-group-major tiles (gemm_tc_split_kmajor's layout; `kmajor_hi/lo` name the planes).
// y1_scalar: branch1 is a Cin = 1 conv folded into (a1, b1): value = y1[r] * a1[g][c] + b1[g][c].
int gemm_tc_combine(cudaStream_t st, const float *y1, bool y1_scalar, const float *a1, const float *b1,
                    const float *y2, const float *a2, const float *b2, const int32_t *win_group, int T, int C,
                    int64_t rows, float *out, const GemmScratch *planes, int layout, int64_t *launches,
                    op16_t *kmajor_hi = nullptr, op16_t *kmajor_lo = nullptr);

// BLSTM recurrence as one tcgen05 GEMM launch per timestep (hidden sizes beyond the persistent kernel, e.g. H = 256):
// see ntb_gemm_tc.cu.  whh = [2 dirs][4H][H] hi/lo planes (plan/pack below); xproj row-major [n*T][ldx] with column
// dir*4H + u*4 + g, pre-scaled like whh (k_repack_lstm prescale = 1); cstate = gemm_tc_lstm_state_floats() floats.
bool gemm_tc_lstm_supported(int H);
void gemm_tc_plan_whh(GemmWeights &w, int H, Arena &ar, float **staging);
int gemm_tc_pack_whh(cudaStream_t st, const float *kernel_fw, const float *kernel_bw, int D, int H, float *staging,
                     GemmWeights &w);
int64_t gemm_tc_lstm_state_floats(int64_t n_windows, int H);
// D_fused > 0: the input projection rides in the same GEMM (whh = [W_ih ; W_hh] planes from gemm_tc_plan_wcat /
// gemm_tc_pack_wcat, x_hi / x_lo = the layer input as time-major operand planes from gemm_tc_split_timemajor, bias
// [2][4H] pre-scaled with the forget bias folded); xproj is then unused.
int gemm_tc_lstm_layer(cudaStream_t st, const GemmWeights &whh, const float *xproj, int ldx, const int32_t *lens,
                       int64_t n_windows, int T, int H, float *cstate, float *out, int ldo, const GemmScratch &gs,
                       int64_t *launches, int D_fused = 0, const op16_t *x_hi = nullptr, const op16_t *x_lo = nullptr,
                       const float *bias = nullptr, op16_t *next_hi = nullptr, op16_t *next_lo = nullptr);
// next_hi / next_lo: h_t is written as the next stacked layer's time-major operand planes (K = 2H, must be a multiple of
// 64) instead of the fp32 `out`.
bool gemm_tc_lstm_fused_supported(int H, int D);
void gemm_tc_plan_wcat(GemmWeights &w, int H, int D, Arena &ar, float **staging);
int gemm_tc_pack_wcat(cudaStream_t st, const float *wih, const float *kernel_fw, const float *kernel_bw, int D, int H,
                      float *staging, GemmWeights &w);
int gemm_tc_lstm_fused_dpad(int D);   // input width padded to whole k-blocks (64): D_fused of gemm_tc_lstm_layer
// fp32 [n_windows*T][lda] (Ksrc channels) -> time-major fp16 hi/lo planes [(tile T + t) 128 + window][K], K >= Ksrc
// (channels beyond Ksrc zero; whole 128-window tiles)
int gemm_tc_split_timemajor(cudaStream_t st, const float *A, int lda, int64_t n_windows, int T, int Ksrc, int K, op16_t *hi,
                            op16_t *lo, int64_t *launches);

}  // namespace ntb
