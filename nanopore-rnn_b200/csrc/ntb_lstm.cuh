// K4  BLSTM recurrence (network.py:266-288: rnn.LSTMCell x2 under tf.nn.bidirectional_dynamic_rnn).
//
// One persistent CTA per (tile of 64 windows, direction); forward and backward tiles are
// independent CTAs of the same launch, so both directions run concurrently.  W_hh for the
// direction stays resident in shared memory for all T steps (fp32, 16*H*Hp bytes: 160 KB at
// H = 100), h_t is double-buffered in shared memory (one __syncthreads per step), the cell
// state c and all four gate pre-activations of a (window, unit) pair live in ONE thread's
// registers, so sigmoid/tanh and the cell update are fused with the mat-vec and never leave
// the register file.  The mat-vec runs on packed fp32 FMAs (FFMA2, 2 MACs per issue slot):
// the recurrence must be fp32-exact to hold the 1e-3 logits budget over 100 dependent steps
// (precision table), so it does not go through bf16 tensor-core operands.
//
// TF semantics restated (rnn_cell_impl.LSTMCell.call, rnn.dynamic_rnn/_rnn_step,
// bidirectional_dynamic_rnn + array_ops.reverse_sequence):
//   z = x_t W_ih + b (precomputed by the projection GEMM, forget_bias folded into f) + h W_hh
//   c = sigmoid(f) c + sigmoid(i) tanh(j) ; h = sigmoid(o) tanh(c) ; c_0 = h_0 = 0
//   t >= len[b]: output 0, state frozen (never read again since steps only move forward)
//   backward: step s consumes t = len-1-s and writes its output at that same t.
#pragma once
#include "ntb_common.cuh"

namespace ntb {

struct LstmParams {
    const float *xproj;   // [rows, ldx]: col = dir*4Hp + u*4 + g, bias folded
    int ldx;
    const float *whh;     // [2][H][4][UG][4] (see k_repack_lstm)
    const int32_t *lens;  // [n_windows]
    float *out;           // [rows, ldo]: col = dir*H + u
    int ldo;
    int64_t n_windows;
    int T, H, Hp;
};

__device__ __forceinline__ void ffma2(unsigned long long &d, float a, unsigned long long b) {
    unsigned long long aa;
    asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(aa), "l"(b));
}
__device__ __forceinline__ void unpack2(unsigned long long v, float &x, float &y) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v));
}

constexpr int LSTM_WPT = 8;    // windows per thread
constexpr int LSTM_UPT = 4;    // hidden units per thread (all 4 gates each)
constexpr int LSTM_WG = 8;     // window groups per CTA
constexpr int LSTM_BT = LSTM_WPT * LSTM_WG;   // 64 windows per CTA

static inline int lstm_threads(int Hp) { return (int)align_up((Hp / LSTM_UPT) * LSTM_WG, 32); }
static inline int64_t lstm_smem_bytes(int H, int Hp, bool w_in_smem) {
    return (int64_t)sizeof(float) * ((w_in_smem ? (int64_t)H * Hp * 4 : 0) + 2LL * H * LSTM_BT);
}

template <bool W_IN_SMEM>
__global__ void __launch_bounds__(256, 1) k_lstm_recurrence(LstmParams p) {
    extern __shared__ __align__(16) float smem[];
    const int H = p.H, Hp = p.Hp, UG = Hp / LSTM_UPT, T = p.T;
    const int dir = blockIdx.y;
    const int64_t w_tile = (int64_t)blockIdx.x * LSTM_BT;
    const float *whh_g = p.whh + (int64_t)dir * H * Hp * 4;
    float *Ws = smem;                                          // [H][4][UG] float4
    float *hs = smem + (W_IN_SMEM ? (int64_t)H * Hp * 4 : 0);  // [2][H][BT]

    const int tid = threadIdx.x;
    if (W_IN_SMEM) {
        const float4 *src = (const float4 *)whh_g;
        float4 *dst = (float4 *)Ws;
        for (int i = tid; i < H * Hp; i += blockDim.x) dst[i] = src[i];
    }
    for (int i = tid; i < 2 * H * LSTM_BT; i += blockDim.x) hs[i] = 0.0f;
    __syncthreads();

    const int ug = tid % UG, wg = tid / UG;
    const bool worker = wg < LSTM_WG;
    const int wl0 = wg * LSTM_WPT;          // first local window of this thread
    const int u0 = ug * LSTM_UPT;           // first unit of this thread

    int len[LSTM_WPT];
    float c[LSTM_WPT][LSTM_UPT];
#pragma unroll
    for (int i = 0; i < LSTM_WPT; ++i) {
        const int64_t w = w_tile + wl0 + i;
        len[i] = (worker && w < p.n_windows) ? max(p.lens[w], 0) : -1;   // -1: window does not exist
        if (len[i] > T) len[i] = T;
#pragma unroll
        for (int q = 0; q < LSTM_UPT; ++q) c[i][q] = 0.0f;
    }
    const float4 *Wv = W_IN_SMEM ? (const float4 *)Ws : (const float4 *)whh_g;

    for (int s = 0; s < T; ++s) {
        const float *hcur = hs + (s & 1) * H * LSTM_BT;
        float *hnxt = hs + ((s & 1) ^ 1) * H * LSTM_BT;
        // acc[i][q][0] = (i, j) pre-activations, acc[i][q][1] = (f, o) of unit u0+q, window i
        unsigned long long acc[LSTM_WPT][LSTM_UPT][2];
        if (worker) {
#pragma unroll
            for (int i = 0; i < LSTM_WPT; ++i) {
                const bool active = s < len[i];
                const int t = dir ? len[i] - 1 - s : s;
                const int64_t row = (w_tile + wl0 + i) * T + (active ? t : 0);
                const ulonglong2 *src =
                    (const ulonglong2 *)(p.xproj + row * p.ldx + (int64_t)dir * 4 * Hp + u0 * 4);
#pragma unroll
                for (int q = 0; q < LSTM_UPT; ++q) {
                    ulonglong2 v = make_ulonglong2(0ull, 0ull);
                    if (active) {
                        v = __ldg(src + q);
                        // pull the next step's gates towards L2 while this step computes
                        if (s + 1 < len[i])
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(
                                (const char *)(src + q) + (dir ? -(int64_t)p.ldx : (int64_t)p.ldx) * 4));
                    }
                    acc[i][q][0] = v.x;
                    acc[i][q][1] = v.y;
                }
            }
            // z += h_{t-1} W_hh : k-major rank-1 updates, 64 FFMA2 per k per thread
#pragma unroll 2
            for (int k = 0; k < H; ++k) {
                const float4 a0 = *(const float4 *)(hcur + k * LSTM_BT + wl0);
                const float4 a1 = *(const float4 *)(hcur + k * LSTM_BT + wl0 + 4);
                const float a[LSTM_WPT] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                ulonglong2 b[LSTM_UPT];
#pragma unroll
                for (int q = 0; q < LSTM_UPT; ++q) {
                    const float4 *bp = Wv + ((int64_t)k * 4 + q) * UG + ug;
                    float4 t4 = W_IN_SMEM ? *bp : __ldg(bp);
                    b[q] = *(ulonglong2 *)&t4;
                }
#pragma unroll
                for (int i = 0; i < LSTM_WPT; ++i)
#pragma unroll
                    for (int q = 0; q < LSTM_UPT; ++q) {
                        ffma2(acc[i][q][0], a[i], b[q].x);
                        ffma2(acc[i][q][1], a[i], b[q].y);
                    }
            }
            // gates + cell update, h_t to smem (next step's operand) and to the layer output
#pragma unroll
            for (int i = 0; i < LSTM_WPT; ++i) {
                if (len[i] < 0) continue;
                const bool active = s < len[i];
                const int t = active ? (dir ? len[i] - 1 - s : s) : s;
                float hv[LSTM_UPT];
#pragma unroll
                for (int q = 0; q < LSTM_UPT; ++q) {
                    float gi, gj, gf, go;
                    unpack2(acc[i][q][0], gi, gj);
                    unpack2(acc[i][q][1], gf, go);
                    const float cn = sigmoid_acc(gf) * c[i][q] + sigmoid_acc(gi) * tanh_acc(gj);
                    const float hn = sigmoid_acc(go) * tanh_acc(cn);
                    if (active) c[i][q] = cn;
                    hv[q] = active ? hn : 0.0f;
                    if (u0 + q < H) hnxt[(u0 + q) * LSTM_BT + wl0 + i] = hv[q];
                }
                float *dst = p.out + ((w_tile + wl0 + i) * T + t) * (int64_t)p.ldo + dir * H + u0;
                if (u0 + LSTM_UPT <= H && ((p.ldo | H) & 3) == 0) {
                    *(float4 *)dst = make_float4(hv[0], hv[1], hv[2], hv[3]);
                } else {
#pragma unroll
                    for (int q = 0; q < LSTM_UPT; ++q)
                        if (u0 + q < H) dst[q] = hv[q];
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace ntb
