// TEST INFRASTRUCTURE -- not part of the product path.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this library.
//
// Plain C++ (OpenMP) restatement of the reference's forward: a third, independently written
// statement of the same arithmetic next to oracle_np.py (NumPy) and oracle_torch.py (torch CPU), and the fastest CPU
// form of it -- the one bench.py times as the host-core baseline.  PARITY UNPINNED: the reference's arithmetic is
// TensorFlow 1.4 (absent) and its tests hold no golden outputs; this file is checked against the
// other two restatements and the committed goldens (tests/test_oracle.py).
//
// Each function cites the reference lines it follows (paths relative to /root/reference).
// Model description: the product's own C descriptor (include/nanotensor_b200.h: ntb_model_desc with HOST pointers
// to the TF-layout variables), so both sides are driven by the same bytes.
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include <algorithm>
#include <vector>

#include "../include/nanotensor_b200.h"

namespace {

// out[r][n] += sum_k A[r][k] * W[k][n]   (row-major, 4 x 32 register block so that gcc keeps the accumulators in
// AVX2 registers; plain loops, no intrinsics).  The order of the k-sum per output is ascending k, as in the naive form.
void gemm_acc(const float *A, int lda, int rows, int K, const float *W, int ldw, int N, float *out, int ldo) {
    constexpr int RB = 4, NB = 32;
    for (int r0 = 0; r0 < rows; r0 += RB) {
        const int rn = std::min(RB, rows - r0);
        for (int n0 = 0; n0 < N; n0 += NB) {
            const int nn = std::min(NB, N - n0);
            float acc[RB][NB];
            for (int i = 0; i < RB; ++i)
                for (int j = 0; j < NB; ++j) acc[i][j] = (i < rn && j < nn) ? out[(size_t)(r0 + i) * ldo + n0 + j] : 0.0f;
            if (rn == RB && nn == NB) {
                for (int k = 0; k < K; ++k) {
                    const float *w = W + (size_t)k * ldw + n0;
                    for (int i = 0; i < RB; ++i) {
                        const float a = A[(size_t)(r0 + i) * lda + k];
                        for (int j = 0; j < NB; ++j) acc[i][j] += a * w[j];
                    }
                }
            } else {
                for (int k = 0; k < K; ++k) {
                    const float *w = W + (size_t)k * ldw + n0;
                    for (int i = 0; i < rn; ++i) {
                        const float a = A[(size_t)(r0 + i) * lda + k];
                        for (int j = 0; j < nn; ++j) acc[i][j] += a * w[j];
                    }
                }
            }
            for (int i = 0; i < rn; ++i)
                for (int j = 0; j < nn; ++j) out[(size_t)(r0 + i) * ldo + n0 + j] = acc[i][j];
        }
    }
}

// nanotensor/network.py:312-329 conv_layer: tf.nn.conv2d NHWC on [B,1,T,Cin], filter [1,k,Cin,Cout], stride 1, SAME,
// no bias: cross-correlation, (k-1)/2 zeros on the left.
void conv(const float *v, int B, int T, int Cin, const float *W, int k, int Cout, float *out) {
    const int left = (k - 1) / 2;
    std::fill(out, out + (size_t)B * T * Cout, 0.0f);
    for (int b = 0; b < B; ++b)
        for (int dk = 0; dk < k; ++dk) {
            // output rows t with 0 <= t + dk - left < T read input rows t + dk - left: one contiguous row range per tap
            const int sh = dk - left;
            const int t0 = std::max(0, -sh), t1 = std::min(T, T - sh);
            if (t1 <= t0) continue;
            gemm_acc(v + ((size_t)b * T + t0 + sh) * Cin, Cin, t1 - t0, Cin, W + (size_t)dk * Cin * Cout, Cout, Cout,
                     out + ((size_t)b * T + t0) * Cout, Cout);
        }
}

// network.py:339-349 simple_global_bn: tf.nn.moments over [0,1,2] of THIS batch (also at inference, padded positions
// included), then tf.nn.batch_normalization: inv = rsqrt(var + 1e-5) * scale ; y * inv + (offset - mean * inv).
void bn(float *y, int rows, int C, const float *scale, const float *offset, bool relu) {
    std::vector<double> s(C, 0.0), q(C, 0.0);
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < C; ++c) s[c] += y[(size_t)r * C + c];
    for (int c = 0; c < C; ++c) s[c] /= rows;
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < C; ++c) {
            const double d = y[(size_t)r * C + c] - s[c];
            q[c] += d * d;
        }
    std::vector<float> a(C), b(C);
    for (int c = 0; c < C; ++c) {
        const float var = (float)(q[c] / rows), mean = (float)s[c];
        a[c] = (1.0f / sqrtf(var + 1e-5f)) * scale[c];
        b[c] = offset[c] - mean * a[c];
    }
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < C; ++c) {
            float v = y[(size_t)r * C + c] * a[c] + b[c];
            y[(size_t)r * C + c] = relu ? std::max(v, 0.0f) : v;
        }
}

// network.py:385-400 residual_layer: branch1 = conv1x1 (+BN iff "batchnorm", no ReLU);
// branch2 = ReLU(BN(conv1x1)) -> ReLU(BN(conv1x3)) -> BN(conv1x1); out = ReLU(branch1 + branch2).
void residual(const ntb_layer &L, const float *v, int B, int T, std::vector<float> &out) {
    const int Cin = L.in_dim, C = L.size, rows = B * T;
    std::vector<float> a((size_t)rows * C), b1((size_t)rows * C), b2((size_t)rows * C);
    conv(v, B, T, Cin, L.w[0], 1, C, a.data());
    if (L.batchnorm && L.w[1]) bn(a.data(), rows, C, L.w[1], L.w[2], false);
    conv(v, B, T, Cin, L.w[3], 1, C, b1.data());
    bn(b1.data(), rows, C, L.w[4], L.w[5], true);
    conv(b1.data(), B, T, C, L.w[6], 3, C, b2.data());
    bn(b2.data(), rows, C, L.w[7], L.w[8], true);
    conv(b2.data(), B, T, C, L.w[9], 1, C, b1.data());
    bn(b1.data(), rows, C, L.w[10], L.w[11], false);
    out.resize((size_t)rows * C);
    for (size_t i = 0; i < out.size(); ++i) out[i] = std::max(a[i] + b1[i], 0.0f);
}

inline float sigmoidf(float x) { return 1.0f / (1.0f + expf(-x)); }

// network.py:266-288 blstm: TF 1.4 rnn.LSTMCell (no peepholes / projection / clipping) under
// tf.nn.bidirectional_dynamic_rnn with sequence_length.  kernel [D+H,4H] (rows 0..D-1 <-> x, D..D+H-1 <-> h), gate
// column blocks i|j|f|o; c' = sigmoid(f + forget_bias) c + sigmoid(i) tanh(j); h' = sigmoid(o) tanh(c'); forget_bias
// is added at run time.  Steps t >= len: output 0, state carried.  Backward direction: t = len-1 .. 0 (reverse_sequence
// over the first len steps), zeros at t >= len.  In-tree statement of the cell: chiron/chiron/utils/lstm.py:37-46.
void lstm_dir(const float *v, int B, int T, int D, int H, const int32_t *lens, const float *kernel, const float *bias,
              float fb, bool reverse, float *out, int ldo_row) {
    // all B windows of the batch advance together (step s: t = s forward, t = len-1-s backward), so the recurrent term
    // is one small blocked GEMM [B x H] . [H x 4H] per step; rows whose sequence has ended are computed and ignored
    const int N = 4 * H;
    // the x part of z = [x_t, h] . kernel + bias does not depend on the recurrence: all rows in one blocked GEMM
    std::vector<float> zx((size_t)B * T * N);
    for (size_t r = 0; r < (size_t)B * T; ++r)
        for (int n = 0; n < N; ++n) zx[r * N + n] = bias[n];
    gemm_acc(v, D, B * T, D, kernel, N, N, zx.data(), N);
    std::vector<float> h((size_t)B * H, 0.0f), c((size_t)B * H, 0.0f), z((size_t)B * N);
    int maxlen = 0;
    for (int b = 0; b < B; ++b) maxlen = std::max(maxlen, std::min<int>(lens[b], T));
    for (int s = 0; s < maxlen; ++s) {
        for (int b = 0; b < B; ++b) {
            const int len = std::min<int>(lens[b], T);
            const int t = s < len ? (reverse ? len - 1 - s : s) : 0;
            const float *src = zx.data() + ((size_t)b * T + t) * N;
            for (int n = 0; n < N; ++n) z[(size_t)b * N + n] = src[n];
        }
        gemm_acc(h.data(), H, B, H, kernel + (size_t)D * N, N, N, z.data(), N);
        for (int b = 0; b < B; ++b) {
            const int len = std::min<int>(lens[b], T);
            if (s >= len) continue;                       // dynamic_rnn: state carried, output stays 0
            const int t = reverse ? len - 1 - s : s;
            const float *zb = z.data() + (size_t)b * N;
            float *o = out + ((size_t)b * T + t) * ldo_row;
            for (int u = 0; u < H; ++u) {
                const float cn = sigmoidf(zb[2 * H + u] + fb) * c[(size_t)b * H + u] + sigmoidf(zb[u]) * tanhf(zb[H + u]);
                const float hn = sigmoidf(zb[3 * H + u]) * tanhf(cn);
                c[(size_t)b * H + u] = cn;
                h[(size_t)b * H + u] = hn;
                o[u] = hn;
            }
        }
    }
}

// network.py:208-211 tf.argmax: first index on ties.
inline int argmax_first(const float *row, int C) {
    int best = 0;
    for (int c = 1; c < C; ++c)
        if (row[c] > row[best]) best = c;
    return best;
}

// One batch (= one BN group) through the whole network; logits [B*T, C].
void forward_group(const ntb_model_desc &d, const float *x, const int32_t *lens, int B, float *logits) {
    const int T = d.seq_len;
    std::vector<float> cur(x, x + (size_t)B * T * d.len_x), nxt;
    int dim = d.len_x;
    for (int li = 0; li < d.n_layers; ++li) {
        const ntb_layer &L = d.layers[li];
        if (L.type == NTB_LAYER_RESIDUAL) {
            residual(L, cur.data(), B, T, nxt);
            dim = L.size;
        } else if (L.type == NTB_LAYER_BLSTM) {
            const int H = L.size;
            nxt.assign((size_t)B * T * 2 * H, 0.0f);     // tf.concat(outputs, 2): fw then bw (network.py:285)
            lstm_dir(cur.data(), B, T, dim, H, lens, L.w[0], L.w[1], L.forget_bias, false, nxt.data(), 2 * H);
            lstm_dir(cur.data(), B, T, dim, H, lens, L.w[2], L.w[3], L.forget_bias, true, nxt.data() + H, 2 * H);
            dim = 2 * H;
        } else {   // network.py:164-175 chiron_fnn: x[...,0,h] w[0,h] + x[...,1,h] w[1,h] + b[h]
            const int H = dim / 2;
            nxt.resize((size_t)B * T * H);
            for (int r = 0; r < B * T; ++r)
                for (int u = 0; u < H; ++u)
                    nxt[(size_t)r * H + u] = cur[(size_t)r * dim + u] * L.w[0][u] + cur[(size_t)r * dim + H + u] * L.w[0][H + u] + L.w[1][u];
            dim = H;
        }
        cur.swap(nxt);
    }
    // network.py:291-309 fulconn_layer / :114-118, 507-513 create_output_layer: logits = flat . W + b, every position
    const int C = d.n_classes;
    for (int r = 0; r < B * T; ++r)
        for (int c = 0; c < C; ++c) logits[(size_t)r * C + c] = d.out_b[c];
    gemm_acc(cur.data(), dim, B * T, dim, d.out_w, C, C, logits, C);
}

}  // namespace

extern "C" {

// logits [n*T, C] (may be NULL -> internal scratch); labels: CrossEntropy [n*T] argmax, CTC [n, T] greedy-decoded
// padded with -1 (network.py:528-533: tf.nn.ctc_greedy_decoder(merge_repeated=True) + sparse_tensor_to_dense(-1));
// label_len [n] decoded lengths (CTC, may be NULL).  group_off [n_groups+1] window offsets of the reference's batches
// (NULL: batches of `batch_size` consecutive windows).  Returns 0, or -1 on a bad argument.
__attribute__((visibility("default")))
int oracle_cpu_forward(const ntb_model_desc *d, const float *x, const int32_t *lens, const int32_t *group_off,
                       int64_t n, int64_t n_groups, int32_t batch_size, float *logits, int32_t *labels,
                       int32_t *label_len, int32_t n_threads) {
    if (!d || !x || !lens || !labels || n < 0) return -1;
    const int T = d->seq_len, C = d->n_classes;
    std::vector<int32_t> goff;
    if (!group_off) {
        if (batch_size <= 0) batch_size = 64;
        for (int64_t w = 0; w < n; w += batch_size) goff.push_back((int32_t)w);
        goff.push_back((int32_t)n);
        group_off = goff.data();
        n_groups = (int64_t)goff.size() - 1;
    }
    std::vector<float> own;
    if (!logits) { own.resize((size_t)n * T * C); logits = own.data(); }
#ifdef _OPENMP
    const int nt = n_threads > 0 ? n_threads : 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nt > 0 ? nt : omp_get_max_threads())
#endif
    for (int64_t g = 0; g < n_groups; ++g) {
        const int64_t w0 = group_off[g], w1 = group_off[g + 1];
        if (w1 <= w0) continue;
        forward_group(*d, x + (size_t)w0 * T * d->len_x, lens + w0, (int)(w1 - w0), logits + (size_t)w0 * T * C);
    }
    if (d->graph == NTB_GRAPH_CROSS_ENTROPY) {
        for (int64_t r = 0; r < n * T; ++r) labels[r] = argmax_first(logits + (size_t)r * C, C);
    } else {
        for (int64_t w = 0; w < n; ++w) {
            int cnt = 0, prev = -1;
            const int len = std::min<int>(lens[w], T);
            for (int t = 0; t < len; ++t) {
                const int k = argmax_first(logits + ((size_t)w * T + t) * C, C);
                if (k != C - 1 && k != prev) labels[w * T + cnt++] = k;   // blank = C-1; prev updated every step
                prev = k;
            }
            for (int i = cnt; i < T; ++i) labels[w * T + i] = -1;
            if (label_len) label_len[w] = cnt;
        }
    }
    return 0;
}

__attribute__((visibility("default")))
int oracle_cpu_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
}
