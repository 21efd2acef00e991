// Fused LSTM cell update shared by the tensor-core recurrences (ntb_lstm_tc.cuh: persistent kernel for H <= 100;
// ntb_gemm_tc.cu: one tcgen05 GEMM per timestep for larger H).  rnn_cell_impl.LSTMCell.call restated:
//   c' = sigmoid(f + forget_bias) c + sigmoid(i) tanh(j) ;  h' = sigmoid(o) tanh(c')
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace ntb {

// scale factors folded into W_ih, W_hh and the bias at pack time: the GEMMs produce the ex2 arguments
constexpr float LSTM_PRESCALE_IFO = -1.4426950408889634f;   // -log2(e)   : sigmoid(z) = 1 / (1 + 2^(y))
constexpr float LSTM_PRESCALE_J = -2.885390081777927f;      // -2 log2(e) : tanh(z) = (1 - 2^y) / (1 + 2^y)

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// c' = sigmoid(f) c + sigmoid(i) tanh(j) ; h' = sigmoid(o) tanh(c')  with 5 ex2 + 2 rcp (7 MUFU
// instead of 10): the three sigmoid/tanh denominators of c' share one reciprocal.
//   sigmoid(z) = 1/(1+2^y), y = -log2(e) z ;  tanh(z) = (1-2^y)/(1+2^y), y = -2 log2(e) z.
// The inputs ARE those y (weights and bias are pre-scaled at pack time).  y is clamped at 38.95
// (z >= -27 resp. -13.5: sigmoid(-27) ~ 2e-12, far below the fp32 noise) so that the product of
// denominators stays below 2^120.
__device__ __forceinline__ void lstm_cell_update(float yi, float yj, float yf, float yo, float &c, float &h) {
    const float CL = 38.95f;
    const float A = ex2_approx(fminf(yi, CL));
    const float B = ex2_approx(fminf(yj, CL));
    const float F = ex2_approx(fminf(yf, CL));
    const float pf = 1.0f + F;
    const float P = (1.0f + A) * (1.0f + B);
    const float cn = fmaf(c, P, (1.0f - B) * pf) * rcp_approx(P * pf);
    const float O = ex2_approx(fminf(yo, CL));
    const float C2 = ex2_approx(fminf(-2.885390081777927f * cn, CL));
    h = (1.0f - C2) * rcp_approx((1.0f + O) * (1.0f + C2));
    c = cn;
}

// two floats -> packed fp16 pair (a -> low half = even k) and the packed pair of their fp16 remainders
__device__ __forceinline__ void split_pack_op16x2(float a, float b, uint32_t &hi2, uint32_t &lo2) {
    const __half2 h = __floats2half2_rn(a, b);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(a - hf.x, b - hf.y);
    hi2 = *reinterpret_cast<const uint32_t *>(&h);
    lo2 = *reinterpret_cast<const uint32_t *>(&l);
}

}  // namespace ntb
