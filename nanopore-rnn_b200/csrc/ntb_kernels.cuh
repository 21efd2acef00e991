// Streaming / small-shape kernels of the NanoTensor forward: window packer, fp32 SIMT conv-GEMM
// (odd shapes: Cin = 1, len_x = 12, H = 5), batch-statistics BN, residual combine, readout.
// Reference semantics are cited per kernel (paths relative to the reference repository).
#pragma once
#include "ntb_common.cuh"

namespace ntb {

// ------------------------------------------------------------------------------------------
// window -> BN group id   (group = one reference batch: network.py:339-349 takes moments over it)
// ------------------------------------------------------------------------------------------
// group_off comes from the caller's device memory on the device-pointer entry (the host entries validate it on the
// CPU first): offsets are clamped to [0, n_windows] and win_group is zero-filled beforehand, so a malformed table can
// give wrong BN statistics but never an out-of-bounds index into the [n_groups][C] affine arrays.
__global__ void k_window_groups(const int32_t *__restrict__ group_off, int n_groups, int64_t n_windows,
                                int32_t *__restrict__ win_group) {
    int g = blockIdx.x;
    if (g >= n_groups) return;
    int64_t w0 = group_off[g], w1 = group_off[g + 1];
    w0 = w0 < 0 ? 0 : w0;
    w1 = w1 > n_windows ? n_windows : w1;
    for (int64_t w = w0 + threadIdx.x; w < w1; w += blockDim.x) win_group[w] = g;
}

// ------------------------------------------------------------------------------------------
// K1  read z-score + windowing + zero pad
//     dataset.py:435-445 (FullSignalSequence.load_data_inference), chiron_input.py:220-230
//     (read_signal: (x - mean) / np.std over the WHOLE read, float64), dataset.py:157-169 (padding)
// ------------------------------------------------------------------------------------------
// pass 1: per-read mean and population std in float64, fixed reduction order (bit-stable)
__global__ void __launch_bounds__(256) k_read_moments(const int16_t *__restrict__ raw,
                                                      const int64_t *__restrict__ read_off,
                                                      double *__restrict__ mean_std) {
    __shared__ double sh[256];
    const int r = blockIdx.x;
    const int64_t s0 = read_off[r], s1 = read_off[r + 1];
    const int64_t n = s1 - s0;
    double acc = 0.0;
    for (int64_t i = s0 + threadIdx.x; i < s1; i += 256) acc += (double)raw[i];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    const double mean = n > 0 ? sh[0] / (double)n : 0.0;
    __syncthreads();
    acc = 0.0;
    for (int64_t i = s0 + threadIdx.x; i < s1; i += 256) {
        double d = (double)raw[i] - mean;
        acc += d * d;
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        mean_std[2 * r] = mean;
        mean_std[2 * r + 1] = n > 0 ? sqrt(sh[0] / (double)n) : 1.0;
    }
}

// pass 2: one warp per window, coalesced reads of int16, float32 stores
__global__ void __launch_bounds__(256) k_pack_windows(const int16_t *__restrict__ raw,
                                                      const int64_t *__restrict__ read_off,
                                                      const int64_t *__restrict__ win_off,
                                                      const double *__restrict__ mean_std,
                                                      int64_t n_reads, int64_t n_windows, int seq_len,
                                                      int step, int start_index,
                                                      float *__restrict__ x, int32_t *__restrict__ lens) {
    const int64_t w = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    const int lane = threadIdx.x & 31;
    if (w >= n_windows) return;
    // binary search: read r with win_off[r] <= w < win_off[r+1]
    int64_t lo = 0, hi = n_reads;
    while (hi - lo > 1) {
        int64_t mid = (lo + hi) >> 1;
        if (win_off[mid] <= w) lo = mid; else hi = mid;
    }
    const int64_t r = lo;
    const int64_t s0 = read_off[r] + start_index, s1 = read_off[r + 1];
    const int64_t first = s0 + (w - win_off[r]) * (int64_t)step;
    int64_t valid = s1 - first;
    valid = valid < 0 ? 0 : (valid > seq_len ? seq_len : valid);
    const double mean = mean_std[2 * r], sd = mean_std[2 * r + 1];
    for (int t = lane; t < seq_len; t += 32) {
        float v = 0.0f;
        if (t < valid) v = (float)(((double)raw[first + t] - mean) / sd);  // f64 then cast, as the tf feed does
        x[w * seq_len + t] = v;
    }
    if (lane == 0) lens[w] = (int32_t)valid;
}

// ------------------------------------------------------------------------------------------
// fp32 SIMT conv-as-GEMM with the producer's BN(+ReLU) applied on load
//     out[r, n] = sum_tap sum_c f(A[r + tap - left, c]) * W[tap][c][n] (+ bias[n])
//     f(v) = relu?(v * aff_a[g, c] + aff_b[g, c]) with g = BN group of the source row, and 0
//     outside the window (tf.nn.conv2d SAME zero-pads the activated tensor, network.py:327-329)
// Used for Cin = 1 convs, len_x = 12 / H = 5 event shapes and as the GPU-side check of the
// tcgen05 path.
// ------------------------------------------------------------------------------------------
struct ConvGemmParams {
    const float *A; int lda;          // [rows, K]
    const float *W;                   // [taps, K, N]
    const float *bias;                // [N] or null
    const float *aff_a, *aff_b;       // [n_groups, K] or null
    const int32_t *win_group;         // [n_windows] (only with aff)
    float *out; int ldo;              // [rows, N]
    int64_t rows; int K, N, taps, T, relu_in;
};

template <int BM, int BN, int BK, int TM, int TN>
__global__ void __launch_bounds__((BM / TM) * (BN / TN)) k_conv_gemm_simt(ConvGemmParams p) {
    constexpr int NT = (BM / TM) * (BN / TN);
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN];
    const int tid = threadIdx.x;
    const int tx = tid % (BN / TN), ty = tid / (BN / TN);
    const int64_t m0 = (int64_t)blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;
    const int left = (p.taps - 1) / 2;
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.0f;

    for (int tap = 0; tap < p.taps; ++tap) {
        for (int k0 = 0; k0 < p.K; k0 += BK) {
            for (int idx = tid; idx < BM * BK; idx += NT) {
                const int m = idx / BK, kk = idx % BK;
                const int64_t r = m0 + m;
                const int c = k0 + kk;
                float v = 0.0f;
                if (r < p.rows && c < p.K) {
                    const int t = (int)(r % p.T) + tap - left;
                    if (t >= 0 && t < p.T) {
                        const int64_t rr = r + tap - left;
                        v = p.A[rr * p.lda + c];
                        if (p.aff_a) {
                            const int g = p.win_group[rr / p.T];
                            v = v * p.aff_a[(int64_t)g * p.K + c] + p.aff_b[(int64_t)g * p.K + c];
                        }
                        if (p.relu_in) v = fmaxf(v, 0.0f);
                    }
                }
                As[kk][m] = v;
            }
            for (int idx = tid; idx < BK * BN; idx += NT) {
                const int kk = idx / BN, n = idx % BN;
                const int c = k0 + kk;
                float v = 0.0f;
                if (c < p.K && n0 + n < p.N) v = p.W[((int64_t)tap * p.K + c) * p.N + n0 + n];
                Bs[kk][n] = v;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < BK; ++kk) {
                float a[TM], b[TN];
#pragma unroll
                for (int i = 0; i < TM; ++i) a[i] = As[kk][ty * TM + i];
#pragma unroll
                for (int j = 0; j < TN; ++j) b[j] = Bs[kk][tx * TN + j];
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int64_t r = m0 + ty * TM + i;
        if (r >= p.rows) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int n = n0 + tx * TN + j;
            if (n < p.N) p.out[r * p.ldo + n] = acc[i][j] + (p.bias ? p.bias[n] : 0.0f);
        }
    }
}

// ------------------------------------------------------------------------------------------
// simple_global_bn statistics (network.py:339-349): tf.nn.moments over axes [0,1,2] of ONE
// reference batch (= BN group; padded positions included), then the affine form
// tf.nn.batch_normalization uses:  a = rsqrt(var + eps) * scale ; b = offset - mean * a.
// ONE pass over the conv output: sum x and sum x^2 accumulate in float64 (fp32 inputs and their squares
// are exact in fp64, so var = E[x^2] - mean^2 loses nothing visible in fp32), fixed order ->
// bit-stable, no atomics.  Block = 64 channels x 4 row slices (a warp reads 128 contiguous
// bytes of one row); the slices are combined through shared memory in slice order.
// ------------------------------------------------------------------------------------------
constexpr int BN_CH4 = 64, BN_SLICES = 4;   // block = 64 float4 columns (256 channels) x 4 row slices
__global__ void __launch_bounds__(BN_CH4 * BN_SLICES) k_bn_stats(const float *__restrict__ y, int C, int T,
                                                                 const int32_t *__restrict__ group_off,
                                                                 const float *__restrict__ scale,
                                                                 const float *__restrict__ offset, float eps,
                                                                 float *__restrict__ aff_a, float *__restrict__ aff_b) {
    __shared__ double sh[2][BN_SLICES][BN_CH4 * 4];
    const int g = blockIdx.x;
    const int cl = threadIdx.x % BN_CH4, sl = threadIdx.x / BN_CH4;
    const int c = (blockIdx.y * BN_CH4 + cl) * 4;          // first of this thread's 4 channels
    const int64_t r0 = (int64_t)group_off[g] * T, r1 = (int64_t)group_off[g + 1] * T;
    const int64_t n = r1 - r0;
    double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0}, s2[4] = {0, 0, 0, 0}, q2[4] = {0, 0, 0, 0};
    if (c < C) {
        const float *p = y + r0 * C + c;
        int64_t i = sl;
        for (; i + BN_SLICES < n; i += 2 * BN_SLICES) {
            const float4 u = *reinterpret_cast<const float4 *>(p + i * C);
            const float4 v = *reinterpret_cast<const float4 *>(p + (i + BN_SLICES) * C);
            const double u0 = u.x, u1 = u.y, u2 = u.z, u3 = u.w, v0 = v.x, v1 = v.y, v2 = v.z, v3 = v.w;
            s[0] += u0; q[0] = fma(u0, u0, q[0]); s[1] += u1; q[1] = fma(u1, u1, q[1]);
            s[2] += u2; q[2] = fma(u2, u2, q[2]); s[3] += u3; q[3] = fma(u3, u3, q[3]);
            s2[0] += v0; q2[0] = fma(v0, v0, q2[0]); s2[1] += v1; q2[1] = fma(v1, v1, q2[1]);
            s2[2] += v2; q2[2] = fma(v2, v2, q2[2]); s2[3] += v3; q2[3] = fma(v3, v3, q2[3]);
        }
        if (i < n) {
            const float4 u = *reinterpret_cast<const float4 *>(p + i * C);
            const double u0 = u.x, u1 = u.y, u2 = u.z, u3 = u.w;
            s[0] += u0; q[0] = fma(u0, u0, q[0]); s[1] += u1; q[1] = fma(u1, u1, q[1]);
            s[2] += u2; q[2] = fma(u2, u2, q[2]); s[3] += u3; q[3] = fma(u3, u3, q[3]);
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        sh[0][sl][cl * 4 + k] = s[k] + s2[k];
        sh[1][sl][cl * 4 + k] = q[k] + q2[k];
    }
    __syncthreads();
    // 256 threads finish the block's 256 channels: slices combined in slice order (bit-stable)
    const int ch = blockIdx.y * BN_CH4 * 4 + threadIdx.x;
    if (ch < C) {
        double ss = 0, qq = 0;
#pragma unroll
        for (int k = 0; k < BN_SLICES; ++k) { ss += sh[0][k][threadIdx.x]; qq += sh[1][k][threadIdx.x]; }
        const double mean = n > 0 ? ss / (double)n : 0.0;
        double var = n > 0 ? qq / (double)n - mean * mean : 0.0;
        var = var < 0.0 ? 0.0 : var;
        const float a = (1.0f / sqrtf((float)var + eps)) * scale[ch];
        aff_a[(int64_t)g * C + ch] = a;
        aff_b[(int64_t)g * C + ch] = offset[ch] - (float)mean * a;
    }
}

// Same affine from the partial sums the tcgen05 conv epilogue emitted (ntb_gemm_tc.cu, bn_partials): record
// (block, seg) holds the fp32 (sum, centred sum of squares) of the rows of plane-row block `block` (32 rows) that belong to
// the block's first window (seg 0) or to the next one (seg 1); a window occupies Tp plane rows (T, or T + 2 in the
// padded 1x3 layout whose padding rows were left out).  One thread per channel, records folded in fp64 in block order -> bit-stable.
// 1 / nb for 1 <= nb <= 32 without an fp64 division: float reciprocal + one Newton step (relative error < 1e-14, exact for
// powers of two; deterministic)
__device__ __forceinline__ double bn_rcp(int nb) {
    const double r0 = (double)__frcp_rn((float)nb);
    return r0 * (2.0 - (double)nb * r0);
}

__global__ void __launch_bounds__(256) k_bn_finish(const float2 *__restrict__ part, int C, int T, int Tp,
                                                   const int32_t *__restrict__ group_off,
                                                   const float *__restrict__ scale, const float *__restrict__ offset,
                                                   float eps, float *__restrict__ aff_a, float *__restrict__ aff_b) {
    const int g = blockIdx.x;
    const int ch = blockIdx.y * 256 + threadIdx.x;
    if (ch >= C) return;
    const int64_t wa = group_off[g], wb = group_off[g + 1];
    const int64_t n = (wb - wa) * T;
    // records are (S_b, M2_b) of n_b <= 32 values (bn_partials): exact merge
    //   M2 = sum M2_b + sum S_b^2 / n_b - (sum S_b)^2 / n        (fp64: the cancellation costs mean^2 / var * 1e-16)
    double ss = 0.0, m2 = 0.0, sq = 0.0;
    if (wb > wa) {
        const int64_t b0 = (wa * Tp) >> 5, b1 = (wb * Tp - 1) >> 5;
        // first window of block b and the offset of the block's first row inside it, advanced without divisions
        int64_t w_first = (b0 << 5) / Tp;
        int rem = (int)((b0 << 5) - w_first * Tp);
        const int lo = Tp != T ? 1 : 0, hi = lo + T;               // real rows of a window inside its Tp plane rows
        const float2 *rec = part + (b0 * 2) * C + ch;
        // unrolled: the (predicated) record loads of 8 blocks are in flight together; the fp64 fold order is unchanged
#pragma unroll 8
        for (int64_t b = b0; b <= b1; ++b, rec += 2 * C) {
            const bool two = rem + 31 >= Tp;                       // the window changes inside this block (Tp >= 32)
            if (w_first >= wa && w_first < wb) {
                const int e = rem + 32 < Tp ? rem + 32 : Tp;
                const int nb = (e < hi ? e : hi) - (rem > lo ? rem : lo);
                const float2 v = rec[0];
                if (nb > 0) { const double S = (double)v.x; ss += S; m2 += (double)v.y; sq = fma(S * S, nb == 32 ? 0.03125 : bn_rcp(nb), sq); }
            }
            if (two && w_first + 1 >= wa && w_first + 1 < wb) {
                const int e = rem + 32 - Tp;                       // rows [0, e) of the next window
                const int nb = (e < hi ? e : hi) - lo;
                const float2 v = rec[C];
                if (nb > 0) { const double S = (double)v.x; ss += S; m2 += (double)v.y; sq = fma(S * S, nb == 32 ? 0.03125 : bn_rcp(nb), sq); }
            }
            rem += 32;
            if (rem >= Tp) { rem -= Tp; ++w_first; }
        }
    }
    const double mean = n > 0 ? ss / (double)n : 0.0;
    double var = n > 0 ? (m2 + (sq - ss * ss / (double)n)) / (double)n : 0.0;
    var = var < 0.0 ? 0.0 : var;
    const float a = (1.0f / sqrtf((float)var + eps)) * scale[ch];
    aff_a[(int64_t)g * C + ch] = a;
    aff_b[(int64_t)g * C + ch] = offset[ch] - (float)mean * a;
}

// Cin = 1 convs (the first residual layer of the CtcLoss graphs, network.py:515-518 reshapes x to [B,T,1]):
// a 1x1 conv of a single channel is the outer product y[r][n] = x[r] * w[n], so it is never materialised.
// Its batch moments are those of x scaled: mean_y = w mean_x, var_y = w^2 var_x, and BN(y) = A x + B with
//   a = rsqrt(w^2 var_x + eps) * scale ; A = a * w ; B = offset - (w mean_x) * a        (no BN: A = w, B = 0).
// One block per BN group: fp64 moments of x in a fixed order (bit-stable), then the per-channel (A, B) of up
// to two convs of the same input (conv1 and conv2a).
struct K1Conv {
    const float *w, *scale, *offset;   // w [C]; scale/offset null = conv without BN
    float *A, *B;                      // [n_groups][C]
};
__global__ void __launch_bounds__(256) k_k1_affine(const float *__restrict__ x, int T,
                                                   const int32_t *__restrict__ group_off, int C, float eps,
                                                   K1Conv c0, K1Conv c1) {
    __shared__ double sh[2][256];
    const int g = blockIdx.x;
    const int64_t r0 = (int64_t)group_off[g] * T, r1 = (int64_t)group_off[g + 1] * T;
    const int64_t n = r1 - r0;
    double s = 0.0, q = 0.0;
    for (int64_t i = r0 + threadIdx.x; i < r1; i += 256) {
        const double v = x[i];
        s += v;
        q = fma(v, v, q);
    }
    sh[0][threadIdx.x] = s;
    sh[1][threadIdx.x] = q;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            sh[0][threadIdx.x] += sh[0][threadIdx.x + o];
            sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
        }
        __syncthreads();
    }
    const double mean = n > 0 ? sh[0][0] / (double)n : 0.0;
    double var = n > 0 ? sh[1][0] / (double)n - mean * mean : 0.0;
    var = var < 0.0 ? 0.0 : var;
    for (int ch = threadIdx.x; ch < C; ch += 256) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const K1Conv &cv = k ? c1 : c0;
            if (!cv.w) continue;
            const double w = cv.w[ch];
            float A = (float)w, B = 0.0f;
            if (cv.scale) {
                const float a = (1.0f / sqrtf((float)(w * w * var) + eps)) * cv.scale[ch];
                A = a * (float)w;
                B = cv.offset[ch] - (float)(w * mean) * a;
            }
            cv.A[(int64_t)g * C + ch] = A;
            cv.B[(int64_t)g * C + ch] = B;
        }
    }
}

// out = relu?( y*a + b )   -- materialised BN(+ReLU) (op-level entry point / odd shapes)
__global__ void __launch_bounds__(256) k_bn_apply(const float *y, const float *__restrict__ a,
                                                  const float *__restrict__ b,
                                                  const int32_t *__restrict__ win_group, int T, int C,
                                                  int64_t total, int relu, float *out) {   // out may alias y
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < total; i += stride) {
        const int64_t r = i / C;
        const int c = (int)(i % C);
        float v = y[i];
        if (a) {
            const int g = win_group[r / T];
            v = v * a[(int64_t)g * C + c] + b[(int64_t)g * C + c];
        }
        out[i] = relu ? fmaxf(v, 0.0f) : v;
    }
}

// residual_layer tail (network.py:398-399): relu( bn?(branch1) + bn(conv2c) ), float4 streaming
__global__ void __launch_bounds__(256) k_residual_combine(
    const float4 *__restrict__ y1, const float *__restrict__ a1, const float *__restrict__ b1,
    const float4 *__restrict__ y2, const float *__restrict__ a2, const float *__restrict__ b2,
    const int32_t *__restrict__ win_group, int T, int C, int64_t total4, float4 *__restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int C4 = C / 4;
    for (; i < total4; i += stride) {
        const int64_t r = i / C4;
        const int c = (int)(i % C4) * 4;
        const int g = win_group[r / T];
        float4 u = y1[i], v = y2[i];
        const float4 va = *(const float4 *)(a2 + (int64_t)g * C + c);
        const float4 vb = *(const float4 *)(b2 + (int64_t)g * C + c);
        v.x = v.x * va.x + vb.x; v.y = v.y * va.y + vb.y; v.z = v.z * va.z + vb.z; v.w = v.w * va.w + vb.w;
        if (a1) {
            const float4 ua = *(const float4 *)(a1 + (int64_t)g * C + c);
            const float4 ub = *(const float4 *)(b1 + (int64_t)g * C + c);
            u.x = u.x * ua.x + ub.x; u.y = u.y * ua.y + ub.y; u.z = u.z * ua.z + ub.z; u.w = u.w * ua.w + ub.w;
        }
        float4 o;
        o.x = fmaxf(u.x + v.x, 0.0f); o.y = fmaxf(u.y + v.y, 0.0f);
        o.z = fmaxf(u.z + v.z, 0.0f); o.w = fmaxf(u.w + v.w, 0.0f);
        out[i] = o;
    }
}

// ------------------------------------------------------------------------------------------
// chiron_fnn (network.py:164-175): out[r,h] = v[r,h]*w[0,h] + v[r,H+h]*w[1,h] + b[h]
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_chiron_fnn(const float *__restrict__ v, const float *__restrict__ w,
                                                    const float *__restrict__ b, int H, int64_t rows,
                                                    float *__restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, total = rows * H;
    for (; i < total; i += stride) {
        const int64_t r = i / H;
        const int h = (int)(i % H);
        out[i] = v[r * 2 * H + h] * w[h] + v[r * 2 * H + H + h] * w[H + h] + b[h];
    }
}

// ------------------------------------------------------------------------------------------
// Fused readout for small C (the CTC configs: C = 4..6): one warp per window walks t,
//   [chiron_fnn] -> dense (network.py:291-309) -> argmax (first max) -> CTC greedy merge
//   (tf.nn.ctc_greedy_decoder merge_repeated=True, blank = C-1, prev updated every step;
//    sparse_tensor_to_dense default -1; network.py:528-533) or plain argmax (network.py:208-211).
// Dense applies to ALL positions (padded ones too) so the optional logits match the reference's
// output_layer; decoding stops at len[b].
// ------------------------------------------------------------------------------------------
template <int MAXC>
__global__ void __launch_bounds__(256) k_readout_small(
    const float *__restrict__ v, int F_in,                       // [rows, F_in]; F_in = 2H with fnn, else F
    const float *__restrict__ fnn_w, const float *__restrict__ fnn_b,   // null = no fnn
    const float *__restrict__ out_w, const float *__restrict__ out_b,   // [F, C], [C]
    const int32_t *__restrict__ lens, int64_t n_windows, int T, int C, int graph,
    float *__restrict__ logits, int32_t *__restrict__ labels, int32_t *__restrict__ label_len) {
    extern __shared__ float sm[];
    const int F = fnn_w ? F_in / 2 : F_in;
    float *sw = sm;                 // [F][C]
    float *sf = sw + F * C;         // fnn_w [2][F], fnn_b [F]
    for (int i = threadIdx.x; i < F * C; i += blockDim.x) sw[i] = out_w[i];
    if (fnn_w) {
        for (int i = threadIdx.x; i < 2 * F; i += blockDim.x) sf[i] = fnn_w[i];
        for (int i = threadIdx.x; i < F; i += blockDim.x) sf[2 * F + i] = fnn_b[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (w >= n_windows) return;
    const int len = lens[w];
    int prev = -1, cnt = 0;
    for (int t = 0; t < T; ++t) {
        const float *row = v + (w * T + t) * (int64_t)F_in;
        float part[MAXC];
#pragma unroll
        for (int c = 0; c < MAXC; ++c) part[c] = 0.0f;
        for (int f = lane; f < F; f += 32) {
            float y = fnn_w ? row[f] * sf[f] + row[F + f] * sf[F + f] + sf[2 * F + f] : row[f];
#pragma unroll
            for (int c = 0; c < MAXC; ++c)
                if (c < C) part[c] = fmaf(y, sw[f * C + c], part[c]);
        }
#pragma unroll
        for (int c = 0; c < MAXC; ++c) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part[c] += __shfl_xor_sync(0xffffffffu, part[c], o);
        }
        int best = 0;
        float bv = 0.0f;
#pragma unroll
        for (int c = 0; c < MAXC; ++c) {
            if (c < C) {
                const float lg = part[c] + out_b[c];
                if (logits && lane == c % 32) logits[(w * T + t) * (int64_t)C + c] = lg;
                if (c == 0 || lg > bv) { bv = lg; best = c; }   // strict > : first max wins
            }
        }
        if (graph == NTB_GRAPH_CTC) {
            if (t < len) {
                if (best != C - 1 && best != prev) {
                    if (lane == 0) labels[w * T + cnt] = best;
                    ++cnt;
                }
                prev = best;
            }
        } else if (lane == 0) {
            labels[w * T + t] = best;
        }
    }
    if (graph == NTB_GRAPH_CTC) {
        for (int i = cnt + lane; i < T; i += 32) labels[w * T + i] = -1;
        if (lane == 0 && label_len) label_len[w] = cnt;
    }
}

// Same readout for F <= 128 (the shipped CRNN configs: F = 100): every lane owns up to 4 features, keeps their
// dense / chiron_fnn weights in registers and loads row t + 1 while it reduces row t (one warp walks its window's
// T rows in order -- without the prefetch each row costs a full memory latency).
template <int MAXC>
__global__ void __launch_bounds__(256) k_readout_small_pipe(
    const float *__restrict__ v, int F_in, const float *__restrict__ fnn_w, const float *__restrict__ fnn_b,
    const float *__restrict__ out_w, const float *__restrict__ out_b, const int32_t *__restrict__ lens,
    int64_t n_windows, int T, int C, int graph, float *__restrict__ logits, int32_t *__restrict__ labels,
    int32_t *__restrict__ label_len) {
    const int F = fnn_w ? F_in / 2 : F_in;
    const int lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (w >= n_windows) return;
    float wd[4][MAXC], fw0[4], fw1[4], fb[4], ob[MAXC];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int f = lane + 32 * k;
        const bool ok = f < F;
#pragma unroll
        for (int c = 0; c < MAXC; ++c) wd[k][c] = (ok && c < C) ? out_w[f * C + c] : 0.0f;
        fw0[k] = (ok && fnn_w) ? fnn_w[f] : 1.0f;
        fw1[k] = (ok && fnn_w) ? fnn_w[F + f] : 0.0f;
        fb[k] = (ok && fnn_w) ? fnn_b[f] : 0.0f;
    }
#pragma unroll
    for (int c = 0; c < MAXC; ++c) ob[c] = c < C ? out_b[c] : 0.0f;
    const int len = lens[w];
    const float *row = v + w * T * (int64_t)F_in;
    float a[4], b[4], na[4], nb[4];
    auto load = [&](const float *r, float (&x)[4], float (&y)[4]) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int f = lane + 32 * k;
            x[k] = f < F ? __ldg(r + f) : 0.0f;
            y[k] = (f < F && fnn_w) ? __ldg(r + F + f) : 0.0f;
        }
    };
    load(row, a, b);
    int prev = -1, cnt = 0;
    for (int t = 0; t < T; ++t) {
        if (t + 1 < T) load(row + (int64_t)(t + 1) * F_in, na, nb);
        float part[MAXC];
#pragma unroll
        for (int c = 0; c < MAXC; ++c) part[c] = 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            // same operation order as the generic kernel: (v0 w0 + v1 w1) + b, then the dense FMAs over f ascending
            const float y = fnn_w ? a[k] * fw0[k] + b[k] * fw1[k] + fb[k] : a[k];
#pragma unroll
            for (int c = 0; c < MAXC; ++c) part[c] = fmaf(y, wd[k][c], part[c]);
        }
#pragma unroll
        for (int c = 0; c < MAXC; ++c) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part[c] += __shfl_xor_sync(0xffffffffu, part[c], o);
        }
        int best = 0;
        float bv = 0.0f;
#pragma unroll
        for (int c = 0; c < MAXC; ++c) {
            if (c < C) {
                const float lg = part[c] + ob[c];
                if (logits && lane == c % 32) logits[(w * T + t) * (int64_t)C + c] = lg;
                if (c == 0 || lg > bv) { bv = lg; best = c; }   // strict > : first max wins
            }
        }
        if (graph == NTB_GRAPH_CTC) {
            if (t < len) {
                if (best != C - 1 && best != prev) {
                    if (lane == 0) labels[w * T + cnt] = best;
                    ++cnt;
                }
                prev = best;
            }
        } else if (lane == 0) {
            labels[w * T + t] = best;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) { a[k] = na[k]; b[k] = nb[k]; }
    }
    if (graph == NTB_GRAPH_CTC) {
        for (int i = cnt + lane; i < T; i += 32) labels[w * T + i] = -1;
        if (lane == 0 && label_len) label_len[w] = cnt;
    }
}

// Vectorised form for F % 4 == 0, F <= 128 (F = 100 in the shipped CRNN configs): lane l owns features 4l .. 4l+3
// (one 16-byte load per direction and row, row t + 1 in flight under the reduction of row t), and the C class sums
// are reduced across the warp TRANSPOSED -- every butterfly step halves the number of classes a lane carries
// (MAXC/2 + MAXC/4 + ... shuffles instead of 5 per class) -- then gathered for the first-max argmax.
template <int MAXC>
__global__ void __launch_bounds__(256) k_readout_small_vec(
    const float *__restrict__ v, int F_in, const float *__restrict__ fnn_w, const float *__restrict__ fnn_b,
    const float *__restrict__ out_w, const float *__restrict__ out_b, const int32_t *__restrict__ lens,
    int64_t n_windows, int T, int C, int graph, float *__restrict__ logits, int32_t *__restrict__ labels,
    int32_t *__restrict__ label_len) {
    static_assert(MAXC == 4 || MAXC == 8, "class tile");
    const int F = fnn_w ? F_in / 2 : F_in;
    const int lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (w >= n_windows) return;
    const bool own = 4 * lane < F;
    float wd[4][MAXC], fw0[4], fw1[4], fb[4], ob[MAXC];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int f = 4 * lane + k;
#pragma unroll
        for (int c = 0; c < MAXC; ++c) wd[k][c] = (own && c < C) ? out_w[f * C + c] : 0.0f;
        fw0[k] = (own && fnn_w) ? fnn_w[f] : 1.0f;
        fw1[k] = (own && fnn_w) ? fnn_w[F + f] : 0.0f;
        fb[k] = (own && fnn_w) ? fnn_b[f] : 0.0f;
    }
#pragma unroll
    for (int c = 0; c < MAXC; ++c) ob[c] = c < C ? out_b[c] : 0.0f;
    const int len = lens[w];
    const float *row = v + w * T * (int64_t)F_in + 4 * lane;
    const float4 z4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    float4 a = own ? __ldg(reinterpret_cast<const float4 *>(row)) : z4;
    float4 b = (own && fnn_w) ? __ldg(reinterpret_cast<const float4 *>(row + F)) : z4;
    // lane that ends up holding class c after the transposed reduction (bits 4, 3[, 2] of the lane select the class)
    int prev = -1, cnt = 0;
    for (int t = 0; t < T; ++t) {
        float4 na = z4, nb = z4;
        if (t + 1 < T && own) {
            na = __ldg(reinterpret_cast<const float4 *>(row + (int64_t)(t + 1) * F_in));
            if (fnn_w) nb = __ldg(reinterpret_cast<const float4 *>(row + (int64_t)(t + 1) * F_in + F));
        }
        const float av[4] = {a.x, a.y, a.z, a.w}, bv4[4] = {b.x, b.y, b.z, b.w};
        float part[MAXC];
#pragma unroll
        for (int c = 0; c < MAXC; ++c) part[c] = 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float y = fnn_w ? av[k] * fw0[k] + bv4[k] * fw1[k] + fb[k] : av[k];
#pragma unroll
            for (int c = 0; c < MAXC; ++c) part[c] = fmaf(y, wd[k][c], part[c]);
        }
        // transposed butterfly: after the step with offset o a lane carries half as many classes
        float r4[4];
        if (MAXC == 8) {
            const bool up = lane & 16;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float send = up ? part[i] : part[i + 4];
                const float keep = up ? part[i + 4] : part[i];
                r4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) r4[i] = part[i];
        }
        float r2[2], r1;
        {
            const int o = MAXC == 8 ? 8 : 16;
            const bool up = lane & o;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const float send = up ? r4[i] : r4[i + 2];
                const float keep = up ? r4[i + 2] : r4[i];
                r2[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
            }
        }
        {
            const int o = MAXC == 8 ? 4 : 8;
            const bool up = lane & o;
            const float send = up ? r2[0] : r2[1];
            const float keep = up ? r2[1] : r2[0];
            r1 = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
#pragma unroll
        for (int o = (MAXC == 8 ? 2 : 4); o > 0; o >>= 1) r1 += __shfl_xor_sync(0xffffffffu, r1, o);
        // class c lives in the lanes whose top bits spell c (MAXC = 8: bits 4,3,2; MAXC = 4: bits 4,3)
        int best = 0;
        float bvv = 0.0f;
#pragma unroll
        for (int c = 0; c < MAXC; ++c) {
            if (c < C) {
                const int src = MAXC == 8 ? (((c >> 2) & 1) << 4) | (((c >> 1) & 1) << 3) | ((c & 1) << 2)
                                          : (((c >> 1) & 1) << 4) | ((c & 1) << 3);
                const float lg = __shfl_sync(0xffffffffu, r1, src) + ob[c];
                if (logits && lane == c) logits[(w * T + t) * (int64_t)C + c] = lg;
                if (c == 0 || lg > bvv) { bvv = lg; best = c; }   // strict > : first max wins
            }
        }
        if (graph == NTB_GRAPH_CTC) {
            if (t < len) {
                if (best != C - 1 && best != prev) {
                    if (lane == 0) labels[w * T + cnt] = best;
                    ++cnt;
                }
                prev = best;
            }
        } else if (lane == 0) {
            labels[w * T + t] = best;
        }
        a = na; b = nb;
    }
    if (graph == NTB_GRAPH_CTC) {
        for (int i = cnt + lane; i < T; i += 32) labels[w * T + i] = -1;
        if (lane == 0 && label_len) label_len[w] = cnt;
    }
}

// ------------------------------------------------------------------------------------------
// Decode for large C (event model, C = 1025): logits already materialised by the dense GEMM.
// argmax: one warp per row; CTC: one warp per window walking t.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int warp_argmax_first(const float *row, int C, int lane) {
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
        const float x = row[c];
        if (x > bv || (x == bv && c < bi)) { bv = x; bi = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    return bi == 0x7fffffff ? 0 : bi;   // all-NaN row -> 0 like a first-max scan
}

// ld = floats between consecutive logits rows (C for the caller's dense [rows][C]; the model's own scratch pads it to a
// multiple of 4 so that the dense layer's epilogue stores 16 bytes at a time)
__global__ void __launch_bounds__(256) k_decode(const float *__restrict__ logits,
                                                const int32_t *__restrict__ lens, int64_t n_windows,
                                                int T, int C, int graph, int32_t *__restrict__ labels,
                                                int32_t *__restrict__ label_len, int ld) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (graph == NTB_GRAPH_CROSS_ENTROPY) {
        if (gw >= n_windows * T) return;
        const int best = warp_argmax_first(logits + gw * (int64_t)ld, C, lane);
        if (lane == 0) labels[gw] = best;
        return;
    }
    if (gw >= n_windows) return;
    const int len = lens[gw];
    int prev = -1, cnt = 0;
    for (int t = 0; t < len && t < T; ++t) {
        const int best = warp_argmax_first(logits + (gw * T + t) * (int64_t)ld, C, lane);
        if (best != C - 1 && best != prev) {
            if (lane == 0) labels[gw * T + cnt] = best;
            ++cnt;
        }
        prev = best;
    }
    for (int i = cnt + lane; i < T; i += 32) labels[gw * T + i] = -1;
    if (lane == 0 && label_len) label_len[gw] = cnt;
}

// Same decode from the argmax keys of the dense GEMM's epilogue (gemm_tc_run_argmax): no logits exist.
// argmax: one thread per row; CTC: one thread per window walking t (reads of 8 bytes per step).
__global__ void __launch_bounds__(256) k_decode_keys(const unsigned long long *__restrict__ keys,
                                                     const int32_t *__restrict__ lens, int64_t n_windows, int T, int C,
                                                     int graph, int32_t *__restrict__ labels, int32_t *__restrict__ label_len) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    auto col = [&](int64_t row) -> int {
        const unsigned long long k = keys[row];
        return k ? (int)(0xffffffffu - (uint32_t)(k & 0xffffffffull)) : 0;       // all-NaN row -> 0 like a first-max scan
    };
    if (graph == NTB_GRAPH_CROSS_ENTROPY) {
        if (i < n_windows * T) labels[i] = col(i);
        return;
    }
    if (i >= n_windows) return;
    const int len = lens[i];
    int prev = -1, cnt = 0;
    for (int t = 0; t < len && t < T; ++t) {
        const int best = col(i * T + t);
        if (best != C - 1 && best != prev) labels[i * T + cnt++] = best;
        prev = best;
    }
    for (int k = cnt; k < T; ++k) labels[i * T + k] = -1;
    if (label_len) label_len[i] = cnt;
}

// ------------------------------------------------------------------------------------------
// LSTM input projection for inputs the tensor-core GEMM does not take (len_x = 12: K % 8 != 0 or K < 64), written
// straight in the unit-major projected-gates layout the per-timestep recurrence reads:
//   xproj[dir][tile * T + t][unit][128 windows] float4 (i, j, f, o) = bias + x[w, t, :] . W[:, dir 4H + 4 unit + g]
// thread = window of the tile; a block walks 32 units: stores are 2 KB contiguous per unit, the weights are warp
// broadcasts.  FMA-bound on paper, store-bound in practice (K is tiny).  wih [K][8H], bias [8H] as k_repack_lstm packs them.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_xproj_unit_major(const float *__restrict__ x, const float *__restrict__ wih,
                                                          const float *__restrict__ bias, int64_t n_windows, int T, int K,
                                                          int H, float4 *__restrict__ xproj) {
    const int64_t mt = blockIdx.x;                    // tile * T + t
    const int dir = blockIdx.y, u0 = blockIdx.z * 32;
    const int64_t tile = mt / T;
    const int t = (int)(mt - tile * T);
    const int64_t w = tile * 128 + threadIdx.x;
    const int64_t xmt = (int64_t)gridDim.x;
    float xv[16];
    const bool exists = w < n_windows;
    for (int k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
        for (int k = 0; k < 16; ++k) xv[k] = (exists && k0 + k < K) ? x[(w * T + t) * (int64_t)K + k0 + k] : 0.0f;
        for (int u = u0; u < u0 + 32 && u < H; ++u) {
            const int col = dir * 4 * H + 4 * u;
            float4 *dst = xproj + ((dir * xmt + mt) * H + u) * 128 + threadIdx.x;
            float4 acc = k0 == 0 ? __ldg(reinterpret_cast<const float4 *>(bias + col)) : *dst;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                if (k0 + k < K) {
                    const float4 wv = __ldg(reinterpret_cast<const float4 *>(wih + (int64_t)(k0 + k) * 8 * H + col));
                    acc.x = fmaf(xv[k], wv.x, acc.x); acc.y = fmaf(xv[k], wv.y, acc.y);
                    acc.z = fmaf(xv[k], wv.z, acc.z); acc.w = fmaf(xv[k], wv.w, acc.w);
                }
            }
            *dst = acc;
        }
    }
}

// ------------------------------------------------------------------------------------------
// Weight repack (device side, once per model): TF LSTMCell kernel [D+H, 4H] with gate column
// blocks i|j|f|o (network.py:275-282; rnn_cell_impl.LSTMCell) ->
//   wih  [D, 2*4*Hp]  projection B operand, column = dir*4Hp + u*4 + g  (gate-interleaved)
//   bias [2*4*Hp]     bias with forget_bias folded into the f slot (TF adds it at run time)
//   whh  [2][H][4][UG][4]  recurrence tile: float4 (i,j,f,o) of unit u = ug*4 + q at [k][q][ug]
// Pad units (u >= H) get zero weights and zero bias.  With prescale the SIMT tile `whh` is left unscaled
// (it is not used by the tensor-core recurrence).
// ------------------------------------------------------------------------------------------
__global__ void k_repack_lstm(const float *__restrict__ kernel_fw, const float *__restrict__ bias_fw,
                              const float *__restrict__ kernel_bw, const float *__restrict__ bias_bw,
                              float forget_bias, int D, int H, int Hp, float *__restrict__ wih,
                              float *__restrict__ bias, float *__restrict__ whh, int prescale) {
    const int N = 2 * 4 * Hp;
    const int UG = Hp / 4;
    const int64_t total = (int64_t)(D + 1 + H) * N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int row = (int)(i / N), col = (int)(i % N);
        const int dir = col / (4 * Hp), u = (col % (4 * Hp)) / 4, g = col % 4;
        const float *kern = dir ? kernel_bw : kernel_fw;
        const float *bs = dir ? bias_bw : bias_fw;
        const bool real = u < H;
        // prescale (tensor-core recurrence): pre-activations are produced as -log2(e) z (i, f, o) and
        // -2 log2(e) z (j), the arguments of the ex2 in sigmoid / tanh (ntb_lstm_tc.cuh)
        const float sc = prescale ? (g == 1 ? -2.885390081777927f : -1.4426950408889634f) : 1.0f;
        if (row < D) {
            wih[(int64_t)row * N + col] = real ? sc * kern[(int64_t)row * 4 * H + g * H + u] : 0.0f;
        } else if (row == D) {
            bias[col] = real ? sc * (bs[g * H + u] + (g == 2 ? forget_bias : 0.0f)) : 0.0f;
        } else {
            const int k = row - D - 1;
            const int ug = u / 4, q = u % 4;
            whh[(((int64_t)dir * H + k) * 4 + q) * UG * 4 + ug * 4 + g] =
                real ? kern[(int64_t)(D + k) * 4 * H + g * H + u] : 0.0f;
        }
    }
}

}  // namespace ntb
