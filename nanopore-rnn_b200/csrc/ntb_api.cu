// C ABI of nanotensor_b200 (include/nanotensor_b200.h): model life cycle, the chunk forward and
// the op-level entry points.  Host-side orchestration only; the kernels live in the .cuh files.
#include <algorithm>
#include <new>
#include <stdlib.h>
#include <vector>

#include "ntb_common.cuh"
#include "ntb_assembly.cuh"
#include "ntb_eventdetect.cuh"
#include "ntb_parse.cuh"
#include "ntb_events.cuh"
#include "ntb_gemm_tc.cuh"
#include "ntb_kernels.cuh"
#include "ntb_lstm.cuh"
#include "ntb_lstm_tc.cuh"
#include "ntb_blstm_fused.cuh"
#include "ntb_ctc.cuh"

namespace ntb {

thread_local char g_err[512] = "";

static int check_device() {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(NTB_ENODEVICE, "no CUDA device: %s", cudaGetErrorString(e));
    int major = 0;
    e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess || major != 10)
        return fail(NTB_ENODEVICE, "nanotensor_b200 needs an sm_100 device (found cc major %s%lld)", "",
                    (long long)major);
    return NTB_OK;
}

struct LayerDev {
    int type = 0, in_dim = 0, size = 0, batchnorm = 0;
    float forget_bias = 0.f;
    // residual: TF filters [taps][Cin][C] and BN scale/offset, raw on the device
    const float *conv[4] = {nullptr, nullptr, nullptr, nullptr};      // conv1, conv2a, conv2b, conv2c
    const float *bn_scale[4] = {nullptr, nullptr, nullptr, nullptr};
    const float *bn_offset[4] = {nullptr, nullptr, nullptr, nullptr};
    GemmWeights convw[4];   // tensor-core tiles (fp16 hi/lo, K-major) where the shape allows
    GemmWeights conv12w;    // [conv1 | conv2a] along N: both 1x1 convs of the layer input as one dual-output launch
    float *conv12_stage = nullptr;   // [Cin][2C] fp32 staging the tiles are packed from
    // blstm
    int Hp = 0;
    float *wih = nullptr, *bias = nullptr, *whh = nullptr;
    GemmWeights wihw;
    const float *kernel_fw = nullptr, *kernel_bw = nullptr;   // raw TF kernels on the device
    op16_t *whh_hi = nullptr, *whh_lo = nullptr;       // tensor-core recurrence planes (ntb_lstm_tc.cuh)
    bool fused = false;       // projection + recurrence in ONE persistent CTA-pair kernel (ntb_blstm_fused.cuh; H <= 100)
    BlstmGeom fg;
    op16_t *f_wih = nullptr, *f_whh = nullptr;   // its operand images
    float *f_bias = nullptr;
    bool tc_rec = false;      // persistent tensor-core recurrence (H <= 100)
    bool step_rec = false;    // one tcgen05 GEMM per timestep (larger H; ntb_gemm_tc.cu)
    GemmWeights whhw;         // [2 dirs][4H][H] planes for step_rec
    float *whh_stage = nullptr;
    bool step_fused = false;  // step_rec with the input projection inside the per-timestep GEMM ([W_ih ; W_hh], D % 64 == 0)
    GemmWeights wcatw;        // [2 dirs][4H][D + H] planes
    float *wcat_stage = nullptr;
    // chiron_fnn
    const float *fnn_w = nullptr, *fnn_b = nullptr;
};

}  // namespace ntb

struct ntb_model {
    int graph, T, len_x, C, F_final;
    std::vector<ntb::LayerDev> layers;
    const float *out_w = nullptr, *out_b = nullptr;
    ntb::GemmWeights outw;
    int max_res_c = 0, max_act_c = 0, max_proj_n = 0, max_step_h = 0;
    int fused_maxk = 0, fused_next_k = 0;   // widest input of a fused BLSTM layer / widest hand-over between two of them (16-padded)
    bool has_res = false, any_tc_rec = false;
    int64_t launches = 0;
    int profile = 0;
    float stage_ms[5] = {0, 0, 0, 0, 0};
    // profile mode: (stage, event) marks of every forward since the last ntb_model_stage_ms();
    // recorded without synchronising so that profiling does not perturb the timed region
    std::vector<std::pair<int, cudaEvent_t>> marks;
    int sm_count = 148;
};

namespace ntb {

// ------------------------------------------------------------------ launch helpers
struct Ctx {
    cudaStream_t st;
    ntb_model *m;   // may be null (op-level calls)
    int64_t launches = 0;
    void count() { ++launches; if (m) ++m->launches; }
};

static int launch_conv_gemm(Ctx &cx, const ConvGemmParams &p) {
    dim3 grid(cdiv(p.rows, 64), cdiv(p.N, 64));
    k_conv_gemm_simt<64, 64, 16, 4, 4><<<grid, 256, 0, cx.st>>>(p);
    cx.count();
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

static int launch_bn_stats(Ctx &cx, const float *y, int C, int T, const int32_t *group_off, int64_t n_groups,
                           const float *scale, const float *offset, float *a, float *b) {
    dim3 grid((unsigned)n_groups, cdiv(C, BN_CH4 * 4));
    k_bn_stats<<<grid, BN_CH4 * BN_SLICES, 0, cx.st>>>(y, C, T, group_off, scale, offset, 1e-5f, a, b);
    cx.count();
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

static int elementwise_grid(int64_t total, int sm_count) {
    int64_t blocks = (total + 255) / 256;
    int64_t cap = (int64_t)sm_count * 16;
    return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

static int launch_lstm(Ctx &cx, const LstmParams &p, int sm_count) {
    (void)sm_count;
    const int UG = p.Hp / LSTM_UPT;
    if (UG * LSTM_WG > 256)
        return fail(NTB_EINVAL, "blstm: hidden size %s%lld is outside the SIMT recurrence (H <= 128); larger H must be a "
                    "multiple of 8 to take the per-timestep tcgen05 path", "", (long long)p.H);
    const bool w_smem = lstm_smem_bytes(p.H, p.Hp, true) <= 227 * 1024;
    const int64_t smem = lstm_smem_bytes(p.H, p.Hp, w_smem);
    dim3 grid(cdiv(p.n_windows, LSTM_BT), 2);
    const int threads = lstm_threads(p.Hp);
    if (w_smem) {
        NTB_CUDA(cudaFuncSetAttribute(k_lstm_recurrence<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_lstm_recurrence<true><<<grid, threads, smem, cx.st>>>(p);
    } else {
        NTB_CUDA(cudaFuncSetAttribute(k_lstm_recurrence<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_lstm_recurrence<false><<<grid, threads, smem, cx.st>>>(p);
    }
    cx.count();
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

static int launch_lstm_tc(Ctx &cx, const LstmTcParams &p) {
    const int64_t smem = lstm_tc_smem_bytes(p.H);
    static bool attr_set = false;
    if (!attr_set) {
        NTB_CUDA(cudaFuncSetAttribute(k_lstm_recurrence_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
        attr_set = true;
    }
    dim3 grid((unsigned)p.n_tiles, 2);
    LstmTcParams pp = p;
    if (const char *e = ntb_env("NTB_LSTM_FLAGS")) pp.flags = atoi(e);   // ablation switches (ntb_lstm_tc.cuh)
    alignas(64) CUtensorMap map_x;
    NTB_TRY(make_map_xproj(&map_x, reinterpret_cast<const float *>(p.xproj), 2LL * p.n_tiles * p.T * p.H));
    k_lstm_recurrence_tc<<<grid, LSTM_TC_THREADS, smem, cx.st>>>(map_x, pp);
    cx.count();
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

// One fused BLSTM layer: x planes (k-group-major, fp16 hi/lo) -> h as fp32 rows (`out`) or as the next layer's planes.
static int launch_blstm_fused(Ctx &cx, const BlstmGeom &g, const op16_t *x_hi, const op16_t *x_lo, const op16_t *wih_img,
                              const op16_t *whh_img, const float *bias, const int32_t *lens, int64_t n, int T, float *out,
                              int ldo, op16_t *next_hi, op16_t *next_lo, int sm_count) {
    BlstmFusedParams p{};
    p.whh_img = whh_img; p.bias = bias; p.lens = lens; p.out = out; p.ldo = ldo;
    p.next_hi = next_hi; p.next_lo = next_lo; p.next_K8 = 2 * ((2 * g.H + 15) / 16);
    p.n_windows = n; p.n_pairs = cdiv(n, 256); p.T = T; p.g = g;
    alignas(64) CUtensorMap mx_hi, mx_lo, mw_a, mw_b;
    const int64_t xrows = (int64_t)p.n_pairs * 2 * T * g.K8 * 16;           // 128-byte rows of one x plane
    const int64_t wrows = 2LL * 2 * g.KC * 2 * ((g.RA + g.RB) / 4);
    NTB_TRY(make_map_rows128(&mx_hi, x_hi, xrows, 32));
    NTB_TRY(make_map_rows128(&mx_lo, x_lo, xrows, 32));
    NTB_TRY(make_map_rows128(&mw_a, wih_img, wrows, g.RA / 4));
    NTB_TRY(make_map_rows128(&mw_b, wih_img, wrows, g.RB / 4));
    static bool attr_set = false;
    if (!attr_set) {
        NTB_CUDA(cudaFuncSetAttribute(k_blstm_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, BF_SMEM_MAX - 1024));
        attr_set = true;
    }
    const int items = 2 * p.n_pairs;                  // even; with <= sm_count/2 (even) CTA pairs a pair keeps ONE direction
    int pairs_hw = (sm_count / 2) & ~1;
    if (pairs_hw < 2) pairs_hw = 2;
    const int cpairs = items < pairs_hw ? items : pairs_hw;
    k_blstm_fused<<<2 * cpairs, BF_THREADS, (size_t)g.smem, cx.st>>>(mx_hi, mx_lo, mw_a, mw_b, p);
    cx.count();
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

static inline int pad_hidden(int H) { return (int)align_up(H, LSTM_UPT); }

// Input projection of a BLSTM layer on the per-timestep recurrence (H > 100), in the unit-major layout its epilogue reads:
// the tcgen05 GEMM with the projected-gates epilogue when the input width allows, an FMA kernel otherwise (len_x = 12).
static int step_rec_projection(Ctx &cx, const float *in, int D, int64_t n, int T, int H, const float *wih, const float *bias,
                               const GemmWeights &wihw, float *xproj, const GemmScratch &gs) {
    if (wihw.valid && gs.valid)
        return gemm_tc_run_xproj(cx.st, in, D, n, T, wihw, bias, xproj, H, gs, &cx.launches, false);
    dim3 grid((unsigned)(align_up(n, 128) / 128 * T), 2, (unsigned)cdiv(H, 32));
    k_xproj_unit_major<<<grid, 128, 0, cx.st>>>(in, wih, bias, n, T, D, H, reinterpret_cast<float4 *>(xproj));
    cx.count();
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

// Dense [rows,K] x [K,N] (+bias): tensor-core path when the packed tiles exist, SIMT otherwise.
static int dense(Ctx &cx, const float *A, int lda, int64_t rows, int K, int N, const float *W_kn,
                 const GemmWeights *tc, const float *bias, float *out, int ldo, GemmScratch *scratch) {
    if (tc && tc->valid && scratch && scratch->valid) {
        return gemm_tc_run(cx.st, A, lda, rows, *tc, bias, out, ldo, *scratch, &cx.launches);
    }
    ConvGemmParams p{};
    p.A = A; p.lda = lda; p.W = W_kn; p.bias = bias; p.out = out; p.ldo = ldo;
    p.rows = rows; p.K = K; p.N = N; p.taps = 1; p.T = 1; p.relu_in = 0;
    return launch_conv_gemm(cx, p);
}

}  // namespace ntb

using namespace ntb;

// ====================================================================== misc
extern "C" const char *ntb_last_error(void) { return g_err; }
extern "C" const char *ntb_version(void) { return "0.1.0 (sm_100a)"; }
extern "C" int ntb_device_ok(void) { return check_device() == NTB_OK ? 1 : 0; }

// ====================================================================== model life cycle
static int validate_desc(const ntb_model_desc *d) {
    if (!d || !d->layers || d->n_layers <= 0) return fail(NTB_EINVAL, "model desc: no layers%s", "");
    if (d->graph != NTB_GRAPH_CROSS_ENTROPY && d->graph != NTB_GRAPH_CTC)
        return fail(NTB_EINVAL, "model desc: graph must be NTB_GRAPH_CROSS_ENTROPY or NTB_GRAPH_CTC%s", "");
    if (d->seq_len <= 0 || d->len_x <= 0 || d->n_classes <= 0 || !d->out_w || !d->out_b)
        return fail(NTB_EINVAL, "model desc: bad seq_len/len_x/n_classes/out weights%s", "");
    int dim = d->len_x;
    for (int i = 0; i < d->n_layers; ++i) {
        const ntb_layer &l = d->layers[i];
        if (l.in_dim != dim)
            return fail(NTB_EINVAL, "layer %s%lld: in_dim %lld does not match the previous layer's output", "",
                        (long long)i, (long long)l.in_dim);
        if (l.type == NTB_LAYER_RESIDUAL) {
            if (l.size <= 0 || l.size % 4) return fail(NTB_EINVAL, "residual layer: out_channel must be a multiple of 4%s", "");
            for (int k : {0, 3, 4, 5, 6, 7, 8, 9, 10, 11})
                if (!l.w[k]) return fail(NTB_EINVAL, "residual layer %s%lld: missing tensor %lld", "", (long long)i, (long long)k);
            if (l.batchnorm && (!l.w[1] || !l.w[2])) return fail(NTB_EINVAL, "residual layer: batchnorm set but conv1_bn tensors missing%s", "");
            dim = l.size;
        } else if (l.type == NTB_LAYER_BLSTM) {
            if (l.size <= 0) return fail(NTB_EINVAL, "blstm layer: size must be positive%s", "");
            for (int k = 0; k < 4; ++k)
                if (!l.w[k]) return fail(NTB_EINVAL, "blstm layer %s%lld: missing tensor %lld", "", (long long)i, (long long)k);
            dim = 2 * l.size;
        } else if (l.type == NTB_LAYER_CHIRON_FNN) {
            if (dim % 2 || !l.w[0] || !l.w[1]) return fail(NTB_EINVAL, "chiron_fnn layer: needs an even input width and 2 tensors%s", "");
            dim = dim / 2;
        } else {
            return fail(NTB_EINVAL, "layer %s%lld: unsupported type %lld", "", (long long)i, (long long)l.type);
        }
    }
    return NTB_OK;
}

static int64_t model_weight_plan(const ntb_model_desc *d, ntb_model *m, Arena &ar, cudaStream_t st, bool upload) {
    // Walks the layer list once; with upload=false only sizes the arena.
    auto put = [&](const float *host, int64_t n) -> const float * {
        float *dst = ar.take<float>(n);
        if (upload && ar.ok()) cudaMemcpyAsync(dst, host, n * sizeof(float), cudaMemcpyHostToDevice, st);
        return dst;
    };
    int dim = d->len_x;
    bool m_prev_tc = false;      // the previous layer is a BLSTM on the tensor-core recurrence
    for (int i = 0; i < d->n_layers; ++i) {
        const ntb_layer &l = d->layers[i];
        LayerDev ld;
        ld.type = l.type; ld.in_dim = l.in_dim; ld.size = l.size; ld.batchnorm = l.batchnorm;
        ld.forget_bias = l.forget_bias;
        if (l.type == NTB_LAYER_RESIDUAL) {
            const int C = l.size, Cin = dim;
            const int taps[4] = {1, 1, 3, 1};
            const int kin[4] = {Cin, Cin, C, C};
            const int widx[4] = {0, 3, 6, 9};
            for (int c = 0; c < 4; ++c) {
                ld.conv[c] = put(l.w[widx[c]], (int64_t)taps[c] * kin[c] * C);
                if (c > 0 || l.batchnorm) {
                    ld.bn_scale[c] = put(l.w[widx[c] + 1], C);
                    ld.bn_offset[c] = put(l.w[widx[c] + 2], C);
                }
                gemm_tc_plan_weights(ld.convw[c], taps[c], kin[c], C, ar);
            }
            if (ld.convw[0].valid && ld.convw[1].valid && C % 32 == 0 && !ntb_env("NTB_NO_DUAL_CONV")) {
                gemm_tc_plan_weights(ld.conv12w, 1, Cin, 2 * C, ar);
                if (ld.conv12w.valid && ld.conv12w.BN <= C && C % ld.conv12w.BN == 0) ld.conv12_stage = ar.take<float>((int64_t)Cin * 2 * C);
                else ld.conv12w = GemmWeights();
            }
            dim = C;
        } else if (l.type == NTB_LAYER_BLSTM) {
            const int H = l.size, D = dim, Hp = pad_hidden(H);
            ld.Hp = Hp;
            const float *kf = put(l.w[0], (int64_t)(D + H) * 4 * H);
            const float *bf = put(l.w[1], 4 * H);
            const float *kb = put(l.w[2], (int64_t)(D + H) * 4 * H);
            const float *bb = put(l.w[3], 4 * H);
            ld.wih = ar.take<float>((int64_t)D * 8 * Hp);
            ld.bias = ar.take<float>(8 * Hp);
            ld.whh = ar.take<float>(2LL * H * Hp * 4);
            // a BLSTM layer fed by a tensor-core BLSTM layer can take its input as the interleaved hi/lo plane the previous
            // recurrence writes (no operand-split pass, no fp32 hand-over).  Opt-in (NTB_ILV=1): it removes k_split_op16 but
            // the projection GEMM then runs 2 passes over 2K columns (7 instead of 4 k-blocks at K = 200) and loses exactly
            // what the split cost (projection stage 26.8 vs 26.5 ms per 65 536 windows, profiles/r01_final_ablations.txt)
            static const bool no_ilv = !(ntb_env("NTB_ILV") && ntb_env("NTB_ILV")[0] == '1');
            const bool prev_tc = i > 0 && d->layers[i - 1].type == NTB_LAYER_BLSTM && m_prev_tc;
            const bool can_tc = Hp == H && lstm_tc_supported(H) && D % 8 == 0 && D >= 64;
            gemm_tc_plan_weights(ld.wihw, 1, D, 8 * Hp, ar, prev_tc && can_tc && !no_ilv);
            ld.kernel_fw = kf; ld.kernel_bw = kb;
            ld.tc_rec = ld.wihw.valid && Hp == H && lstm_tc_supported(H);
            // default for these shapes: projection and recurrence fused (NTB_NO_FUSED_BLSTM=1 in an ablation build
            // selects the round-1 pair of kernels for A/B runs)
            ld.fused = ld.tc_rec && !ntb_env("NTB_NO_FUSED_BLSTM") && blstm_fused_geom(H, D, ld.fg);
            if (ld.fused) {
                ld.f_wih = ar.take<op16_t>(blstm_wih_img_elems(ld.fg));
                ld.f_whh = ar.take<op16_t>(blstm_whh_img_elems(ld.fg));
                ld.f_bias = ar.take<float>(8 * H);
            }
            m_prev_tc = ld.tc_rec;
            if (ld.tc_rec) {
                ld.whh_hi = ar.take<op16_t>(lstm_tc_whh_elems(H));
                ld.whh_lo = ar.take<op16_t>(lstm_tc_whh_elems(H));
            } else if (Hp == H && gemm_tc_lstm_supported(H)) {
                gemm_tc_plan_whh(ld.whhw, H, ar, &ld.whh_stage);
                ld.step_rec = ld.whhw.valid;
                if (ld.step_rec && gemm_tc_lstm_fused_supported(H, D) && !ntb_env("NTB_NO_FUSED_STEP")) {
                    gemm_tc_plan_wcat(ld.wcatw, H, D, ar, &ld.wcat_stage);
                    ld.step_fused = ld.wcatw.valid;
                }
            }
            if (upload && ar.ok()) {
                k_repack_lstm<<<256, 256, 0, st>>>(kf, bf, kb, bb, l.forget_bias, D, H, Hp, ld.wih, ld.bias, ld.whh,
                                                   (ld.tc_rec || ld.step_rec) ? 1 : 0);
                if (ld.tc_rec)
                    k_pack_whh_tc<<<256, 256, 0, st>>>(kf, kb, D, H, (int)align_up(H, 16), ld.whh_hi, ld.whh_lo);
                if (ld.step_rec) gemm_tc_pack_whh(st, kf, kb, D, H, ld.whh_stage, ld.whhw);
                if (ld.step_fused) gemm_tc_pack_wcat(st, ld.wih, kf, kb, D, H, ld.wcat_stage, ld.wcatw);
                if (ld.fused)
                    k_pack_blstm_fused<<<512, 256, 0, st>>>(kf, bf, kb, bb, l.forget_bias, ld.fg, ld.f_wih, ld.f_whh, ld.f_bias);
            }
            dim = 2 * H;
        } else {
            ld.fnn_w = put(l.w[0], dim);
            ld.fnn_b = put(l.w[1], dim / 2);
            dim = dim / 2;
        }
        if (m) m->layers.push_back(ld);
    }
    const float *ow = put(d->out_w, (int64_t)dim * d->n_classes);
    const float *ob = put(d->out_b, d->n_classes);
    GemmWeights outw;
    gemm_tc_plan_weights(outw, 1, dim, d->n_classes, ar);
    if (m) { m->out_w = ow; m->out_b = ob; m->F_final = dim; m->outw = outw; }
    return ar.off;
}

extern "C" int64_t ntb_model_weight_bytes(const ntb_model_desc *desc) {
    if (validate_desc(desc) != NTB_OK) return NTB_EINVAL;
    Arena ar(nullptr, 0);
    return model_weight_plan(desc, nullptr, ar, nullptr, false);
}

extern "C" int ntb_model_create(const ntb_model_desc *desc, void *weights_dev, int64_t weights_bytes,
                                void *stream, ntb_model **out) {
    if (!out) return fail(NTB_EINVAL, "ntb_model_create: out is NULL%s", "");
    *out = nullptr;
    NTB_TRY(validate_desc(desc));
    NTB_TRY(check_device());
    if (!weights_dev) return fail(NTB_EINVAL, "ntb_model_create: weights_dev is NULL%s", "");
    cudaStream_t st = (cudaStream_t)stream;
    ntb_model *m = new (std::nothrow) ntb_model();
    if (!m) return fail(NTB_ENOMEM, "out of host memory%s", "");
    m->graph = desc->graph; m->T = desc->seq_len; m->len_x = desc->len_x; m->C = desc->n_classes;
    Arena ar(weights_dev, weights_bytes);
    model_weight_plan(desc, m, ar, st, true);
    if (!ar.ok()) {
        delete m;
        return fail(NTB_ENOMEM, "weights buffer too small: need %s%lld bytes, have %lld", "", (long long)ar.off,
                    (long long)weights_bytes);
    }
    // tensor-core tiles: split the fp32 TF tensors into K-major fp16 hi/lo planes
    for (auto &ld : m->layers) {
        if (ld.type == NTB_LAYER_RESIDUAL) {
            for (int c = 0; c < 4; ++c)
                if (ld.convw[c].valid) gemm_tc_pack_weights(st, ld.conv[c], ld.convw[c]);
            if (ld.conv12w.valid) {
                const size_t cb = (size_t)ld.size * sizeof(float);
                cudaMemcpy2DAsync(ld.conv12_stage, 2 * cb, ld.conv[0], cb, cb, ld.in_dim, cudaMemcpyDeviceToDevice, st);
                cudaMemcpy2DAsync(ld.conv12_stage + ld.size, 2 * cb, ld.conv[1], cb, cb, ld.in_dim, cudaMemcpyDeviceToDevice, st);
                gemm_tc_pack_weights(st, ld.conv12_stage, ld.conv12w);
            }
            m->has_res = true;
            if (ld.size > m->max_res_c) m->max_res_c = ld.size;
            if (ld.size > m->max_act_c) m->max_act_c = ld.size;
        } else if (ld.type == NTB_LAYER_BLSTM) {
            if (ld.wihw.valid) gemm_tc_pack_weights(st, ld.wih, ld.wihw);
            if (2 * ld.size > m->max_act_c) m->max_act_c = 2 * ld.size;
            if (!ld.fused && 8 * ld.Hp > m->max_proj_n) m->max_proj_n = 8 * ld.Hp;
            if (ld.tc_rec && !ld.fused) m->any_tc_rec = true;
            if (ld.fused) {
                const int kin = (int)align_up(ld.in_dim, 16);
                if (kin > m->fused_maxk) m->fused_maxk = kin;
                const size_t at = &ld - &m->layers[0];
                if (at + 1 < m->layers.size() && m->layers[at + 1].type == NTB_LAYER_BLSTM && m->layers[at + 1].fused) {
                    const int kn = (int)align_up(2 * ld.size, 16);
                    if (kn > m->fused_next_k) m->fused_next_k = kn;
                }
            }
            if (ld.step_rec && ld.size > m->max_step_h) m->max_step_h = ld.size;
        } else {
            if (ld.in_dim / 2 > m->max_act_c) m->max_act_c = ld.in_dim / 2;
        }
    }
    if (m->outw.valid) gemm_tc_pack_weights(st, m->out_w, m->outw);
    cudaError_t e = cudaStreamSynchronize(st);   // host tensors may be freed by the caller after return
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
        delete m;
        return fail(NTB_ECUDA, "ntb_model_create: %s", cudaGetErrorString(e));
    }
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&m->sm_count, cudaDevAttrMultiProcessorCount, dev);
    *out = m;
    return NTB_OK;
}

static void drop_marks(ntb_model *m);
extern "C" void ntb_model_destroy(ntb_model *m) {
    if (!m) return;
    drop_marks(m);
    delete m;
}

// ====================================================================== workspace plan
namespace ntb {
struct Workspace {
    int32_t *win_group;
    float *act[2];
    float *y1, *y2a, *y2b;
    float *aff[8];
    float *xproj;
    float2 *bn_part, *bn_part2;
    float *cstate;
    float *logits;
    GemmScratch gs;
    op16_t *fx_hi = nullptr, *fx_lo = nullptr;   // second set of x planes of the fused BLSTM layers (the first is gs.a_hi / a_lo)
};

static int64_t plan_workspace(const ntb_model *m, int64_t n, int64_t n_groups, bool own_logits, Arena &ar,
                              Workspace &ws) {
    const int64_t rows = n * m->T;
    ws.win_group = ar.take<int32_t>(n > 0 ? n : 1);
    const int actc = m->max_act_c > 0 ? m->max_act_c : 1;
    ws.act[0] = ar.take<float>(rows * actc);
    ws.act[1] = ar.take<float>(rows * actc);
    if (m->has_res) {
        ws.y1 = ar.take<float>(rows * m->max_res_c);
        ws.y2a = ar.take<float>(rows * m->max_res_c);
        ws.y2b = ar.take<float>(rows * m->max_res_c);
        for (auto &a : ws.aff) a = ar.take<float>((n_groups > 0 ? n_groups : 1) * (int64_t)m->max_res_c);
    } else {
        ws.y1 = ws.y2a = ws.y2b = nullptr;
        for (auto &a : ws.aff) a = nullptr;
    }
    // the tensor-core recurrence works on whole 128-window tiles (time-major), tail tile padded
    const int64_t tile_rows = (m->any_tc_rec || m->max_step_h) ? align_up(n, LSTM_TC_ROWS) * m->T : 0;
    ws.xproj = m->max_proj_n ? ar.take<float>((tile_rows > rows ? tile_rows : rows) * m->max_proj_n) : nullptr;
    ws.bn_part = nullptr;   // set below, once the operand planes are sized
    ws.cstate = m->max_step_h ? ar.take<float>(gemm_tc_lstm_state_floats(n, m->max_step_h)) : nullptr;
    const bool small_c = m->C <= 8;
    ws.logits = (!small_c && own_logits) ? ar.take<float>(rows * align_up(m->C, 4)) : nullptr;   // row stride padded to 16 bytes
    int maxk = m->len_x > m->max_step_h ? m->len_x : m->max_step_h;
    for (auto &ld : m->layers) {
        if (ld.in_dim > maxk) maxk = ld.in_dim;
        if (ld.type == NTB_LAYER_RESIDUAL && ld.size > maxk) maxk = ld.size;
    }
    if (m->F_final > maxk) maxk = m->F_final;
    // the per-timestep recurrence keeps its h operand planes (2 parities x 2 dirs x padded windows x H) in the same scratch
    const int64_t step_rows = m->max_step_h ? cdiv(4 * align_up(n, 256) * m->max_step_h, align_up(maxk, 8)) : 0;
    if (m->fused_maxk > maxk) maxk = m->fused_maxk;
    // fused BLSTM layers read their input as k-group-major planes of whole 256-window tile pairs
    const int64_t fused_rows = m->fused_maxk ? cdiv(blstm_plane_elems(n, m->T, m->fused_maxk), align_up(maxk, 8)) : 0;
    int64_t min_rows = tile_rows > step_rows ? tile_rows : step_rows;
    if (fused_rows > min_rows) min_rows = fused_rows;
    gemm_tc_plan_scratch(ws.gs, rows, m->T, maxk, ar, min_rows);
    if (m->fused_next_k) {
        ws.fx_hi = ar.take<op16_t>(blstm_plane_elems(n, m->T, m->fused_next_k));
        ws.fx_lo = ar.take<op16_t>(blstm_plane_elems(n, m->T, m->fused_next_k));
    }
    ws.bn_part2 = nullptr;
    if (m->has_res && ws.gs.valid && m->T >= 32) {
        ws.bn_part = ar.take<float2>(gemm_tc_bn_part_elems(ws.gs, m->max_res_c));
        ws.bn_part2 = ar.take<float2>(gemm_tc_bn_part_elems(ws.gs, m->max_res_c));
    }
    return ar.off;
}
}  // namespace ntb

extern "C" int64_t ntb_model_workspace_bytes(const ntb_model *m, int64_t n_windows, int64_t n_groups) {
    if (!m || n_windows < 0) return NTB_EINVAL;
    Arena ar(nullptr, 0);
    Workspace ws;
    return plan_workspace(m, n_windows, n_groups, true, ar, ws);
}

// ====================================================================== the forward
// in_planes: the operand planes already hold this layer's input (written by the previous layer's combine).
// want_layout: operand-plane layout the NEXT layer consumes (0 none, 1 plain rows, 2 time-major tiles); when the
// layer can emit it, the fp32 output is skipped and *got_layout is set.
static int residual_forward(Ctx &cx, ntb_model *m, const LayerDev &ld, const float *in, int Cin, int64_t n,
                            const int32_t *group_off, int64_t n_groups, Workspace &ws, float *out, bool in_planes,
                            int want_layout, int *got_layout) {
    const int C = ld.size, T = m->T;
    const int64_t rows = n * T;
    float *ys[4] = {ws.y1, ws.y2a, ws.y2b, ws.y2a};   // conv2c reuses conv2a's buffer
    const int taps[4] = {1, 1, 3, 1};
    const int src_of[4] = {-1, -1, 1, 2};             // conv2b reads bn+relu(conv2a), conv2c reads bn+relu(conv2b)
    float *aa[4] = {ws.aff[0], ws.aff[2], ws.aff[4], ws.aff[6]};
    float *bb[4] = {ws.aff[1], ws.aff[3], ws.aff[5], ws.aff[7]};
    const bool tc_tail = ws.gs.valid && ld.convw[2].valid && ld.convw[3].valid && C % 8 == 0;
    // Cin = 1: conv1 / conv2a are outer products folded into per-group affines of x (k_k1_affine)
    const bool k1 = Cin == 1 && tc_tail;
    if (k1) {
        K1Conv c0{ld.conv[0], ld.bn_scale[0], ld.bn_offset[0], aa[0], bb[0]};
        K1Conv c1{ld.conv[1], ld.bn_scale[1], ld.bn_offset[1], aa[1], bb[1]};
        k_k1_affine<<<(unsigned)n_groups, 256, 0, cx.st>>>(in, T, group_off, C, 1e-5f, c0, c1);
        cx.count();
        NTB_LAUNCH_CHECK();
    }
    int c_first = k1 ? 2 : 0;
    if (!k1 && ld.conv12w.valid && ws.gs.valid && ws.bn_part && ws.bn_part2 && ld.bn_scale[1]) {
        // conv1 | conv2a: one dual-output launch over the shared operand planes, then the two BN folds
        float2 *p1 = ld.bn_scale[0] ? ws.bn_part : nullptr;
        NTB_TRY(gemm_tc_conv_dual(cx.st, in, Cin, rows, T, ld.conv12w, ys[0], ys[1], C, ws.gs, &cx.launches, in_planes, p1,
                                  ws.bn_part2));
        for (int c = 0; c < 2; ++c) {
            if (!ld.bn_scale[c]) continue;
            dim3 grid((unsigned)n_groups, cdiv(C, 256));
            k_bn_finish<<<grid, 256, 0, cx.st>>>(c ? ws.bn_part2 : ws.bn_part, C, T, T, group_off, ld.bn_scale[c],
                                                 ld.bn_offset[c], 1e-5f, aa[c], bb[c]);
            cx.count();
            NTB_LAUNCH_CHECK();
        }
        c_first = 2;
    }
    for (int c = c_first; c < 4; ++c) {
        if (c == 3) { aa[3] = ws.aff[2]; bb[3] = ws.aff[3]; }   // conv2a's affine is dead by now
        const float *A = src_of[c] < 0 ? in : ys[src_of[c]];
        const int K = src_of[c] < 0 ? Cin : C;
        const float *fa = src_of[c] < 0 ? nullptr : aa[src_of[c]];
        const float *fb = src_of[c] < 0 ? nullptr : bb[src_of[c]];
        if (ld.convw[c].valid && ws.gs.valid) {
            // conv1 and conv2a are both 1x1 convs of the layer input: one operand split serves both
            const bool reuse = (c == 1 && ld.convw[0].valid) || (c == 0 && in_planes);
            const bool expand = k1 && c == 2;
            // BN statistics ride in the GEMM epilogue (partials) and are folded per group right after
            float2 *part = ld.bn_scale[c] ? ws.bn_part : nullptr;
            // conv2b / conv2c: the producer's BN + ReLU + operand split run inside the GEMM (transform warps)
            // (a Cin = 1 layer's conv2b makes its operand from x itself: 4 bytes per row instead of padded planes)
            const bool fused = src_of[c] >= 0 && gemm_tc_conv_fused_supported(ld.convw[c], K, T, C);
            int tp = taps[c] == 3 ? T + 2 : T;
            if (fused) {
                NTB_TRY(gemm_tc_conv_fused(cx.st, A, K, rows, T, fa, fb, ws.win_group, 1, ld.convw[c], ys[c], C, &cx.launches, part,
                                           expand ? in : nullptr));
                tp = T;
            } else {
                NTB_TRY(gemm_tc_conv(cx.st, expand ? in : A, K, rows, T, fa, fb, ws.win_group, src_of[c] >= 0, ld.convw[c],
                                     ys[c], C, ws.gs, &cx.launches, reuse, expand, part));
            }
            if (part) {
                dim3 grid((unsigned)n_groups, cdiv(C, 256));
                k_bn_finish<<<grid, 256, 0, cx.st>>>(part, C, T, tp, group_off, ld.bn_scale[c],
                                                     ld.bn_offset[c], 1e-5f, aa[c], bb[c]);
                cx.count();
                NTB_LAUNCH_CHECK();
                continue;
            }
        } else {
            ConvGemmParams p{};
            p.A = A; p.lda = K; p.W = ld.conv[c]; p.bias = nullptr;
            p.aff_a = fa; p.aff_b = fb; p.win_group = ws.win_group;
            p.out = ys[c]; p.ldo = C; p.rows = rows; p.K = K; p.N = C; p.taps = taps[c]; p.T = T;
            p.relu_in = src_of[c] >= 0;
            NTB_TRY(launch_conv_gemm(cx, p));
        }
        if (ld.bn_scale[c])
            NTB_TRY(launch_bn_stats(cx, ys[c], C, T, group_off, n_groups, ld.bn_scale[c], ld.bn_offset[c], aa[c], bb[c]));
    }
    *got_layout = 0;
    if (tc_tail) {
        const bool bn1 = k1 || ld.bn_scale[0];
        *got_layout = want_layout;
        return gemm_tc_combine(cx.st, k1 ? in : ws.y1, k1, bn1 ? aa[0] : nullptr, bn1 ? bb[0] : nullptr, ys[3], aa[3], bb[3],
                               ws.win_group, T, C, rows, want_layout ? nullptr : out, &ws.gs, want_layout, &cx.launches,
                               ws.gs.a_hi, ws.gs.a_lo);
    }
    const int64_t total4 = rows * C / 4;
    k_residual_combine<<<elementwise_grid(total4, m->sm_count), 256, 0, cx.st>>>(
        (const float4 *)ws.y1, ld.bn_scale[0] ? aa[0] : nullptr, ld.bn_scale[0] ? bb[0] : nullptr,
        (const float4 *)ys[3], aa[3], bb[3], ws.win_group, T, C, total4, (float4 *)out);
    cx.count();
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

static int forward_impl(ntb_model *m, const float *x, const int32_t *lens, const int32_t *group_off, int64_t n,
                        int64_t n_groups, Workspace &ws, float *logits_out, int32_t *labels, int32_t *label_len,
                        cudaStream_t st) {
    Ctx cx{st, m};
    m->launches = 0;
    const int T = m->T;
    const int64_t rows = n * T;
    if (n == 0) return NTB_OK;
    int stage = -1;   // 0 conv, 1 projection (operand split), 2 recurrence, 3 readout, 4 projection GEMM kernel
    auto &marks = m->marks;
    auto mark = [&](int s) -> int {
        if (!m->profile) return NTB_OK;
        if (s == stage) return NTB_OK;
        stage = s;
        cudaEvent_t e;
        NTB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDefault));
        NTB_CUDA(cudaEventRecord(e, st));
        marks.push_back({s, e});
        return NTB_OK;
    };
    if (m->has_res) {
        if (!group_off || n_groups <= 0) return fail(NTB_EINVAL, "forward: residual layers need group_off (BN groups)%s", "");
        NTB_CUDA(cudaMemsetAsync(ws.win_group, 0, sizeof(int32_t) * (size_t)n, st));
        k_window_groups<<<(unsigned)n_groups, 128, 0, st>>>(group_off, (int)n_groups, n, ws.win_group);
        cx.count();
        NTB_LAUNCH_CHECK();
    }
    const float *cur = x;
    int dim = m->len_x;
    int flip = 0;
    const size_t nl = m->layers.size();
    const bool small_c = m->C <= 8;
    const LayerDev *fused_fnn = nullptr;
    int planes = 0;   // the GEMM operand planes hold the current activation: 1 plain rows, 2 time-major tiles,
                      // 3 k-group-major tiles in the scratch planes, 4 the same in the fx planes (fused BLSTM layers),
                      // 5 / 6 time-major planes in region 0 / 1 of the projected-gates buffer (H > 100 layers)
    for (size_t li = 0; li < nl; ++li) {
        const LayerDev &ld = m->layers[li];
        float *out = ws.act[flip];
        if (ld.type == NTB_LAYER_RESIDUAL) {
            NTB_TRY(mark(0));
            int want = 0;
            if (li + 1 < nl && ws.gs.valid) {
                const LayerDev &nx = m->layers[li + 1];
                if (nx.type == NTB_LAYER_RESIDUAL && nx.convw[0].valid && nx.convw[1].valid) want = 1;
                else if (nx.type == NTB_LAYER_BLSTM && nx.fused) want = 3;
                else if (nx.type == NTB_LAYER_BLSTM && nx.tc_rec) want = 2;
            }
            int got = 0;
            NTB_TRY(residual_forward(cx, m, ld, cur, dim, n, group_off, n_groups, ws, out, planes == 1, want, &got));
            planes = got;
            dim = ld.size;
        } else if (ld.type == NTB_LAYER_BLSTM) {
            NTB_TRY(mark(1));
            const int N = 8 * ld.Hp;
            if (ld.fused) {
                if (!ws.gs.valid) return fail(NTB_EINVAL, "blstm: operand planes missing for the fused tensor-core path%s", "");
                // x planes: set 0 = the GEMM scratch planes, set 1 = fx; a layer reads one set and hands h to the other
                op16_t *xin_hi = ws.gs.a_hi, *xin_lo = ws.gs.a_lo;
                if (planes == 4) { xin_hi = ws.fx_hi; xin_lo = ws.fx_lo; }
                else if (planes != 3) NTB_TRY(gemm_tc_split_kmajor(cx.st, cur, dim, n, T, dim, xin_hi, xin_lo, &cx.launches));
                NTB_TRY(mark(2));
                const bool feed = li + 1 < nl && m->layers[li + 1].type == NTB_LAYER_BLSTM && m->layers[li + 1].fused;
                op16_t *nx_hi = nullptr, *nx_lo = nullptr;
                if (feed) { nx_hi = planes == 4 ? ws.gs.a_hi : ws.fx_hi; nx_lo = planes == 4 ? ws.gs.a_lo : ws.fx_lo; }
                NTB_TRY(launch_blstm_fused(cx, ld.fg, xin_hi, xin_lo, ld.f_wih, ld.f_whh, ld.f_bias, lens, n, T,
                                           feed ? nullptr : out, 2 * ld.size, nx_hi, nx_lo, m->sm_count));
                dim = 2 * ld.size;
                planes = feed ? (planes == 4 ? 3 : 4) : 0;
                cur = out;
                flip ^= 1;
                continue;
            }
            if (ld.tc_rec) {
                // the packed projection weights are pre-scaled for this recurrence: no SIMT detour from here
                if (!ws.gs.valid) return fail(NTB_EINVAL, "blstm: operand planes missing for the tensor-core path%s", "");
                struct Hook { decltype(mark) *mk; static void call(void *a) { (*static_cast<Hook *>(a)->mk)(4); } } hook{&mark};
                NTB_TRY(gemm_tc_run_xproj(cx.st, cur, dim, n, T, ld.wihw, ld.bias, ws.xproj, ld.size, ws.gs, &cx.launches,
                                          planes == 2, &Hook::call, &hook));
                NTB_TRY(mark(2));
                LstmTcParams p{};
                p.xproj = (const float4 *)ws.xproj; p.whh_hi = ld.whh_hi; p.whh_lo = ld.whh_lo; p.lens = lens;
                p.out = out; p.ldo = 2 * ld.size; p.n_windows = n; p.n_tiles = cdiv(n, LSTM_TC_ROWS);
                p.T = T; p.H = ld.size; p.Hp = (int)align_up(ld.size, 16);
                // a following tensor-core BLSTM layer can take h straight as its projection operand planes.  Opt-in
                // (NTB_REC_PLANES=1): the two 8-byte row-strided stores per chunk cost the recurrence more (24.3 -> 30.0 ms
                // per 65 536 windows) than the saved split pass gives back (32.2 -> 23.8 ms incl. the TMA-store epilogue,
                // 26.8 -> 23.8 without it): its scattered stores are the recurrence's L1 bottleneck
                static const bool rec_planes = ntb_env("NTB_REC_PLANES") && ntb_env("NTB_REC_PLANES")[0] == '1';
                const bool feed_next = rec_planes && li + 1 < nl && m->layers[li + 1].type == NTB_LAYER_BLSTM &&
                                       m->layers[li + 1].tc_rec && 2 * ld.size <= ws.gs.kp_cap && (2 * ld.size) % 8 == 0;
                if (feed_next) { p.next_hi = ws.gs.a_hi; p.next_lo = ws.gs.a_lo; p.next_K = 2 * ld.size; }
                const bool feed_ilv = li + 1 < nl && m->layers[li + 1].type == NTB_LAYER_BLSTM && m->layers[li + 1].wihw.ilv;
                if (feed_ilv) {
                    if (!ws.gs.contiguous() ||
                        (int64_t)p.n_tiles * LSTM_TC_ROWS * T * 4 * ld.size > 2 * ws.gs.rows_cap * (int64_t)ws.gs.kp_cap)
                        return fail(NTB_EINVAL, "blstm: operand scratch too small for the interleaved hand-over%s", "");
                    p.next_hi = p.next_lo = nullptr;
                    p.next_ilv = ws.gs.a_hi; p.next_K = 2 * ld.size;
                }
                NTB_TRY(launch_lstm_tc(cx, p));
                dim = 2 * ld.size;
                planes = (feed_next || feed_ilv) ? 2 : 0;
                cur = out;
                flip ^= 1;
                continue;
            } else if (ld.step_rec) {
                if (!ws.gs.valid) return fail(NTB_EINVAL, "blstm: operand planes missing for the tensor-core path%s", "");
                if (ld.step_fused) {
                    // the projection rides in the per-timestep GEMM: the layer input goes to time-major operand planes (they
                    // live where the projected gates would have been) and no projected gates exist
                    const int Dp = gemm_tc_lstm_fused_dpad(dim);       // len_x = 12 -> one zero-padded k-block of 64
                    // two plane regions inside the (otherwise unused) projected-gates buffer: a layer reads one and, when the
                    // next layer is of the same kind, writes h_t straight into the other (no fp32 hand-over, no split pass)
                    const int64_t trows = align_up(n, LSTM_TC_ROWS) * T;
                    const int64_t region = 2 * trows * (int64_t)(m->max_proj_n / 4);       // elements; 8H/4 = 2H >= any Dp here
                    op16_t *base = reinterpret_cast<op16_t *>(ws.xproj);
                    const int rin = planes == 6 ? 1 : 0;
                    op16_t *x_hi = base + rin * region, *x_lo = x_hi + trows * (int64_t)Dp;
                    if (planes != 5 && planes != 6)
                        NTB_TRY(gemm_tc_split_timemajor(cx.st, cur, dim, n, T, dim, Dp, x_hi, x_lo, &cx.launches));
                    NTB_TRY(mark(2));
                    const bool feed = li + 1 < nl && m->layers[li + 1].type == NTB_LAYER_BLSTM && m->layers[li + 1].step_fused &&
                                      (2 * ld.size) % 64 == 0 && 2 * ld.size <= m->max_proj_n / 4 && Dp <= m->max_proj_n / 4;
                    op16_t *nx_hi = feed ? base + (1 - rin) * region : nullptr;
                    op16_t *nx_lo = feed ? nx_hi + trows * (int64_t)(2 * ld.size) : nullptr;
                    NTB_TRY(gemm_tc_lstm_layer(cx.st, ld.wcatw, nullptr, N, lens, n, T, ld.size, ws.cstate, out, 2 * ld.size,
                                               ws.gs, &cx.launches, Dp, x_hi, x_lo, ld.bias, nx_hi, nx_lo));
                    dim = 2 * ld.size;
                    planes = feed ? (rin ? 5 : 6) : 0;
                    cur = out;
                    flip ^= 1;
                    continue;
                } else {
                NTB_TRY(step_rec_projection(cx, cur, dim, n, T, ld.size, ld.wih, ld.bias, ld.wihw, ws.xproj, ws.gs));
                NTB_TRY(mark(2));
                NTB_TRY(gemm_tc_lstm_layer(cx.st, ld.whhw, ws.xproj, N, lens, n, T, ld.size, ws.cstate, out, 2 * ld.size,
                                           ws.gs, &cx.launches));
                }
            } else {
                NTB_TRY(dense(cx, cur, dim, rows, dim, N, ld.wih, &ld.wihw, ld.bias, ws.xproj, N, &ws.gs));
                NTB_TRY(mark(2));
                LstmParams p{};
                p.xproj = ws.xproj; p.ldx = N; p.whh = ld.whh; p.lens = lens; p.out = out; p.ldo = 2 * ld.size;
                p.n_windows = n; p.T = T; p.H = ld.size; p.Hp = ld.Hp;
                NTB_TRY(launch_lstm(cx, p, m->sm_count));
            }
            dim = 2 * ld.size;
            planes = 0;
        } else {   // chiron_fnn
            NTB_TRY(mark(3));
            if (li + 1 == nl && small_c) { fused_fnn = &ld; continue; }   // fused into the readout
            const int64_t total = rows * (dim / 2);
            k_chiron_fnn<<<elementwise_grid(total, m->sm_count), 256, 0, st>>>(cur, ld.fnn_w, ld.fnn_b, dim / 2, rows, out);
            cx.count();
            NTB_LAUNCH_CHECK();
            dim = dim / 2;
        }
        cur = out;
        flip ^= 1;
    }
    NTB_TRY(mark(3));
    if (small_c) {
        const int F = fused_fnn ? dim / 2 : dim;
        const size_t smem = sizeof(float) * ((size_t)F * m->C + (fused_fnn ? 3 * F : 0));
        if (smem > 200 * 1024) return fail(NTB_EINVAL, "readout: dense weights do not fit shared memory%s", "");
        const bool vec = F <= 128 && F % 4 == 0 && (dim % 4) == 0 && (reinterpret_cast<uintptr_t>(cur) & 15) == 0;
        if (vec && m->C <= 4) {
            k_readout_small_vec<4><<<cdiv(n, 8), 256, 0, st>>>(cur, dim, fused_fnn ? fused_fnn->fnn_w : nullptr,
                                                               fused_fnn ? fused_fnn->fnn_b : nullptr, m->out_w, m->out_b,
                                                               lens, n, T, m->C, m->graph, logits_out, labels, label_len);
        } else if (vec) {
            k_readout_small_vec<8><<<cdiv(n, 8), 256, 0, st>>>(cur, dim, fused_fnn ? fused_fnn->fnn_w : nullptr,
                                                               fused_fnn ? fused_fnn->fnn_b : nullptr, m->out_w, m->out_b,
                                                               lens, n, T, m->C, m->graph, logits_out, labels, label_len);
        } else if (F <= 128) {
            k_readout_small_pipe<8><<<cdiv(n, 8), 256, 0, st>>>(cur, dim, fused_fnn ? fused_fnn->fnn_w : nullptr,
                                                               fused_fnn ? fused_fnn->fnn_b : nullptr, m->out_w, m->out_b,
                                                               lens, n, T, m->C, m->graph, logits_out, labels, label_len);
        } else {
            NTB_CUDA(cudaFuncSetAttribute(k_readout_small<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_readout_small<8><<<cdiv(n, 8), 256, smem, st>>>(cur, dim, fused_fnn ? fused_fnn->fnn_w : nullptr,
                                                              fused_fnn ? fused_fnn->fnn_b : nullptr, m->out_w, m->out_b,
                                                              lens, n, T, m->C, m->graph, logits_out, labels, label_len);
        }
        cx.count();
        NTB_LAUNCH_CHECK();
    } else {
        float *lg = logits_out ? logits_out : ws.logits;
        if (!lg) return fail(NTB_ENOMEM, "forward: no logits buffer for a large-C readout%s", "");
        if (!logits_out && m->outw.valid && ws.gs.valid && !ntb_env("NTB_NO_ARGMAX_EPILOGUE")) {
            // labels only: dense layer + argmax in ONE GEMM epilogue (the logits are never written; the scratch holds 8 bytes
            // per row of (value, column) keys instead of C floats), then the collapse / copy of the winners
            unsigned long long *keys = reinterpret_cast<unsigned long long *>(ws.logits);
            NTB_TRY(gemm_tc_run_argmax(cx.st, cur, dim, rows, m->outw, m->out_b, keys, ws.gs, &cx.launches));
            const int64_t threads = m->graph == NTB_GRAPH_CROSS_ENTROPY ? rows : n;
            k_decode_keys<<<cdiv(threads, 256), 256, 0, st>>>(keys, lens, n, T, m->C, m->graph, labels, label_len);
            cx.count();
            NTB_LAUNCH_CHECK();
        } else {
        // the caller's logits are dense [rows][C]; the internal scratch (labels only) pads the row stride to a multiple of 4
        // floats: with C = 1025 the dense layer's epilogue otherwise falls back to scalar stores
        const int ldl = logits_out ? m->C : (int)align_up(m->C, 4);
        NTB_TRY(dense(cx, cur, dim, rows, dim, m->C, m->out_w, &m->outw, m->out_b, lg, ldl, &ws.gs));
        const int64_t warps = m->graph == NTB_GRAPH_CROSS_ENTROPY ? rows : n;
        k_decode<<<cdiv(warps, 8), 256, 0, st>>>(lg, lens, n, T, m->C, m->graph, labels, label_len, ldl);
        cx.count();
        NTB_LAUNCH_CHECK();
        }
    }
    if (m->profile) {
        cudaEvent_t e;
        NTB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDefault));
        NTB_CUDA(cudaEventRecord(e, st));
        marks.push_back({-1, e});
    }
    m->launches = cx.launches;
    return NTB_OK;
}

extern "C" int ntb_model_forward(ntb_model *m, const float *x_dev, const int32_t *len_dev,
                                 const int32_t *group_off_dev, int64_t n_windows, int64_t n_groups,
                                 void *workspace_dev, int64_t workspace_bytes, float *logits_dev,
                                 int32_t *labels_dev, int32_t *label_len_dev, void *stream) {
    if (!m || !x_dev || !len_dev || !labels_dev || n_windows < 0)
        return fail(NTB_EINVAL, "ntb_model_forward: NULL argument%s", "");
    NTB_TRY(check_device());
    Arena ar(workspace_dev, workspace_bytes);
    Workspace ws;
    plan_workspace(m, n_windows, n_groups, logits_dev == nullptr, ar, ws);
    if (!workspace_dev || !ar.ok())
        return fail(NTB_ENOMEM, "workspace too small: need %s%lld bytes, have %lld", "", (long long)ar.off,
                    (long long)workspace_bytes);
    return forward_impl(m, x_dev, len_dev, group_off_dev, n_windows, n_groups, ws, logits_dev, labels_dev,
                        label_len_dev, (cudaStream_t)stream);
}

namespace ntb {
struct HostStage {
    float *x;
    int32_t *len, *group_off, *labels, *label_len;
    float *logits;
};
static int64_t plan_host_stage(const ntb_model *m, int64_t n, int64_t n_groups, int want_logits, Arena &ar,
                               HostStage &hs) {
    const int64_t rows = n * m->T;
    hs.x = ar.take<float>(rows * m->len_x);
    hs.len = ar.take<int32_t>(n > 0 ? n : 1);
    hs.group_off = ar.take<int32_t>(n_groups + 1);
    hs.labels = ar.take<int32_t>(rows > 0 ? rows : 1);
    hs.label_len = ar.take<int32_t>(n > 0 ? n : 1);
    hs.logits = want_logits ? ar.take<float>(rows * m->C) : nullptr;
    return ar.off;
}
}  // namespace ntb

extern "C" int64_t ntb_model_workspace_bytes_host(const ntb_model *m, int64_t n_windows, int64_t n_groups,
                                                  int want_logits) {
    if (!m || n_windows < 0) return NTB_EINVAL;
    Arena ar(nullptr, 0);
    HostStage hs;
    plan_host_stage(m, n_windows, n_groups, want_logits, ar, hs);
    Workspace ws;
    return plan_workspace(m, n_windows, n_groups, !want_logits, ar, ws);
}

static int forward_host_impl(ntb_model *m, const float *x_host, const int32_t *len_host,
                             const int32_t *group_off_host, int64_t n_windows, int64_t n_groups,
                             void *workspace_dev, int64_t workspace_bytes, float *logits_host,
                             int32_t *labels_host, int32_t *label_len_host, void *stream, bool sync) {
    if (!m || !x_host || !len_host || !labels_host || n_windows < 0)
        return fail(NTB_EINVAL, "ntb_model_forward_host: NULL argument%s", "");
    NTB_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    Arena ar(workspace_dev, workspace_bytes);
    HostStage hs;
    plan_host_stage(m, n_windows, n_groups, logits_host != nullptr, ar, hs);
    Workspace ws;
    plan_workspace(m, n_windows, n_groups, logits_host == nullptr, ar, ws);
    if (!workspace_dev || !ar.ok())
        return fail(NTB_ENOMEM, "workspace too small: need %s%lld bytes, have %lld", "", (long long)ar.off,
                    (long long)workspace_bytes);
    const int64_t rows = n_windows * m->T;
    if (n_windows == 0) return NTB_OK;
    if (group_off_host && n_groups > 0) {   // the BN group table must tile [0, n_windows) exactly (header contract)
        bool ok = group_off_host[0] == 0 && group_off_host[n_groups] == n_windows;
        for (int64_t g = 0; ok && g < n_groups; ++g) ok = group_off_host[g] <= group_off_host[g + 1];
        if (!ok)
            return fail(NTB_EINVAL, "ntb_model_forward_host: group_off must ascend from 0 to n_windows (%s%lld)", "",
                        (long long)n_windows);
    }
    for (int64_t w = 0; w < n_windows; ++w)
        if (len_host[w] < 0 || len_host[w] > m->T)
            return fail(NTB_EINVAL, "ntb_model_forward_host: len[%s%lld] = %lld outside [0, seq_len]", "", (long long)w,
                        (long long)len_host[w]);
    NTB_CUDA(cudaMemcpyAsync(hs.x, x_host, rows * m->len_x * sizeof(float), cudaMemcpyHostToDevice, st));
    NTB_CUDA(cudaMemcpyAsync(hs.len, len_host, n_windows * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    if (group_off_host && n_groups > 0)
        NTB_CUDA(cudaMemcpyAsync(hs.group_off, group_off_host, (n_groups + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    NTB_TRY(forward_impl(m, hs.x, hs.len, (group_off_host && n_groups > 0) ? hs.group_off : nullptr, n_windows,
                         n_groups, ws, hs.logits, hs.labels, hs.label_len, st));
    NTB_CUDA(cudaMemcpyAsync(labels_host, hs.labels, rows * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    if (label_len_host && m->graph == NTB_GRAPH_CTC)
        NTB_CUDA(cudaMemcpyAsync(label_len_host, hs.label_len, n_windows * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    if (logits_host)
        NTB_CUDA(cudaMemcpyAsync(logits_host, hs.logits, rows * m->C * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (sync) NTB_CUDA(cudaStreamSynchronize(st));
    return NTB_OK;
}

extern "C" int ntb_model_forward_host(ntb_model *m, const float *x_host, const int32_t *len_host,
                                      const int32_t *group_off_host, int64_t n_windows, int64_t n_groups,
                                      void *workspace_dev, int64_t workspace_bytes, float *logits_host,
                                      int32_t *labels_host, int32_t *label_len_host, void *stream) {
    return forward_host_impl(m, x_host, len_host, group_off_host, n_windows, n_groups, workspace_dev, workspace_bytes,
                             logits_host, labels_host, label_len_host, stream, true);
}

extern "C" int ntb_model_forward_host_async(ntb_model *m, const float *x_host, const int32_t *len_host,
                                            const int32_t *group_off_host, int64_t n_windows, int64_t n_groups,
                                            void *workspace_dev, int64_t workspace_bytes, float *logits_host,
                                            int32_t *labels_host, int32_t *label_len_host, void *stream) {
    return forward_host_impl(m, x_host, len_host, group_off_host, n_windows, n_groups, workspace_dev, workspace_bytes,
                             logits_host, labels_host, label_len_host, stream, false);
}

extern "C" int64_t ntb_model_last_launches(const ntb_model *m) { return m ? m->launches : 0; }
static void drop_marks(ntb_model *m) {
    for (auto &mk : m->marks) cudaEventDestroy(mk.second);
    m->marks.clear();
}
extern "C" int ntb_model_set_profile(ntb_model *m, int enable) {
    if (!m) return NTB_EINVAL;
    m->profile = enable ? 1 : 0;
    drop_marks(m);
    return NTB_OK;
}
extern "C" int ntb_model_stage_ms(ntb_model *m, float *ms_out, int max_stages) {
    if (!m || !ms_out) return 0;
    float acc[5] = {0, 0, 0, 0, 0};
    auto &marks = m->marks;
    if (!marks.empty()) cudaEventSynchronize(marks.back().second);
    for (size_t i = 0; i + 1 < marks.size(); ++i) {
        if (marks[i].first < 0 || marks[i].first > 4) continue;   // -1 = end of one forward
        float ms = 0;
        if (cudaEventElapsedTime(&ms, marks[i].second, marks[i + 1].second) == cudaSuccess) acc[marks[i].first] += ms;
    }
    drop_marks(m);
    for (int i = 0; i < 5; ++i) m->stage_ms[i] = acc[i];
    if (max_stages < 5) m->stage_ms[1] += acc[4];   // callers of the 4-stage form: projection = split + GEMM
    int k = max_stages < 5 ? max_stages : 5;
    for (int i = 0; i < k; ++i) ms_out[i] = m->stage_ms[i];
    return k;
}

// ====================================================================== op-level entry points
extern "C" int ntb_pack_windows(const int16_t *raw_dev, const int64_t *read_off_dev, const int64_t *win_off_dev,
                                int64_t n_reads, int64_t n_windows, int32_t seq_len, int32_t step,
                                int32_t start_index, float *x_dev, int32_t *len_dev, void *stream) {
    // needs a per-read (mean, std) scratch: carved from the tail of x_dev is not possible -> caller
    // passes nothing; we keep a small stream-ordered allocation instead.
    if (!raw_dev || !read_off_dev || !win_off_dev || !x_dev || !len_dev || seq_len <= 0 || step <= 0 || start_index < 0)
        return fail(NTB_EINVAL, "ntb_pack_windows: bad argument%s", "");
    NTB_TRY(check_device());
    if (n_reads == 0 || n_windows == 0) return NTB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    double *ms = nullptr;
    NTB_CUDA(cudaMallocAsync((void **)&ms, sizeof(double) * 2 * n_reads, st));
    k_read_moments<<<(unsigned)n_reads, 256, 0, st>>>(raw_dev, read_off_dev, ms);
    k_pack_windows<<<cdiv(n_windows, 8), 256, 0, st>>>(raw_dev, read_off_dev, win_off_dev, ms, n_reads, n_windows,
                                                       seq_len, step, start_index, x_dev, len_dev);
    cudaError_t e = cudaGetLastError();
    cudaFreeAsync(ms, st);
    if (e != cudaSuccess) return fail(NTB_ECUDA, "ntb_pack_windows: %s", cudaGetErrorString(e));
    return NTB_OK;
}

// ---- event features (nanonet/nanonet/features.py:61-75; kernels in ntb_events.cuh) ----
namespace ntb {
// Which kernels a read table needs: k_event_features_smem at three, two or one CTA per SM by read length, the any-length
// kernel beyond that.
struct EventPlan {
    int sm[3] = {0, 0, 0};             // longest read of each shared-memory class (0 = class empty)
    int lng = 0;                       // longest read of the any-length class
};

static int ev_cap(size_t budget) {     // largest n whose field + tree fit `budget` bytes of shared memory
    int lo = 0, hi = 1 << 16;
    while (lo < hi) {
        const int m = (lo + hi + 1) / 2;
        if (ev_smem_bytes_sm(m, m / EV_MIN_LEAF + 1) <= budget) lo = m; else hi = m - 1;
    }
    return lo;
}

static int plan_events(const int64_t *read_off, int64_t n_reads, int64_t total_events, EventPlan &pl) {
    static const int cap[3] = {ev_cap((228 * 1024) / 3 - 1024 - 128), ev_cap((228 * 1024) / 2 - 1024 - 128),
                               ev_cap(227 * 1024 - 128)};
    if (read_off[0] != 0 || read_off[n_reads] != total_events)
        return fail(NTB_EINVAL, "ntb_event_features: read_off must run from 0 to total_events%s", "");
    for (int64_t r = 0; r < n_reads; ++r) {
        const int64_t n = read_off[r + 1] - read_off[r];
        if (n < 0) return fail(NTB_EINVAL, "ntb_event_features: read_off not monotonic at read %s%lld", "", (long long)r);
        if (n > (1 << 30) || ev_smem_bytes((int)(n / EV_MIN_LEAF + 1)) > 227 * 1024)
            return fail(NTB_EINVAL, "ntb_event_features: read of %s%lld events exceeds the per-read limit", "", (long long)n);
        int &slot = n <= cap[0] ? pl.sm[0] : (n <= cap[1] ? pl.sm[1] : (n <= cap[2] ? pl.sm[2] : pl.lng));
        slot = (int)n > slot ? (int)n : slot;
    }
    return NTB_OK;
}

template <typename OutT>
static int launch_event_features_t(const double *mean, const double *stdv, const double *length, const int64_t *read_off_dev,
                                   int64_t n_reads, int64_t total_events, const EventPlan &pl, double *scratch, OutT *out,
                                   cudaStream_t st) {
    int n_lo = 1;
    for (int cls = 0; cls < 3; ++cls) {
        const int longest = pl.sm[cls];
        if (!longest) continue;
        const int n_cap = (longest + 7) & ~7, l_max = n_cap / EV_MIN_LEAF + 1;
        const size_t smem = ev_smem_bytes_sm(n_cap, l_max);
        cudaFuncSetAttribute(k_event_features_smem<OutT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_event_features_smem<OutT><<<(unsigned)n_reads, EV_SM_THREADS, smem, st>>>(mean, stdv, length, read_off_dev, out,
                                                                                   scratch, total_events, l_max,
                                                                                   (2 << ev_tree_depth(n_cap)) - 1, n_lo, n_cap);
        n_lo = n_cap + 1;
    }
    if (pl.lng) {
        const int l_max = pl.lng / EV_MIN_LEAF + 1;
        const size_t smem = ev_smem_bytes(l_max);
        cudaFuncSetAttribute(k_event_features<OutT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_event_features<OutT><<<(unsigned)n_reads, EV_THREADS, smem, st>>>(mean, stdv, length, read_off_dev, out, l_max, n_lo);
    }
    const cudaError_t e = cudaGetLastError();      // also reports a failed cudaFuncSetAttribute above
    if (e != cudaSuccess) return fail(NTB_ECUDA, "ntb_event_features: %s", cudaGetErrorString(e));
    return NTB_OK;
}

static int launch_event_features(const double *mean, const double *stdv, const double *length, const int64_t *read_off_dev,
                                 int64_t n_reads, int64_t total_events, const EventPlan &pl, double *scratch, int out_f64,
                                 void *out, cudaStream_t st) {
    return out_f64 ? launch_event_features_t<double>(mean, stdv, length, read_off_dev, n_reads, total_events, pl, scratch, (double *)out, st)
                   : launch_event_features_t<float>(mean, stdv, length, read_off_dev, n_reads, total_events, pl, scratch, (float *)out, st);
}
}  // namespace ntb

// workspace: [read_off: n_reads + 1 int64][scratch: 2 * total_events doubles]
extern "C" int64_t ntb_event_features_workspace_bytes(int64_t n_reads, int64_t total_events) {
    if (n_reads < 0 || total_events < 0) return 0;
    return align_up((int64_t)sizeof(int64_t) * (n_reads + 1), 256) + (int64_t)sizeof(double) * 2 * total_events;
}

extern "C" int ntb_event_features(const double *mean_dev, const double *stdv_dev, const double *length_dev,
                                  const int64_t *read_off_host, int64_t n_reads, int64_t total_events, int32_t out_f64,
                                  void *workspace_dev, int64_t workspace_bytes, void *out_dev, void *stream) {
    if (!read_off_host || n_reads < 0 || total_events < 0)
        return fail(NTB_EINVAL, "ntb_event_features: bad argument%s", "");
    if (total_events > 0 && (!mean_dev || !stdv_dev || !length_dev || !out_dev))
        return fail(NTB_EINVAL, "ntb_event_features: null buffer%s", "");
    if (((uintptr_t)out_dev & 15) || ((uintptr_t)workspace_dev & 15))
        return fail(NTB_EINVAL, "ntb_event_features: out_dev and workspace_dev must be 16-byte aligned%s", "");
    EventPlan pl;
    NTB_TRY(plan_events(read_off_host, n_reads, total_events, pl));
    NTB_TRY(check_device());
    if (n_reads == 0 || total_events == 0) return NTB_OK;
    const int64_t need = ntb_event_features_workspace_bytes(n_reads, total_events);
    if (!workspace_dev || workspace_bytes < need)
        return fail(NTB_ENOMEM, "ntb_event_features: workspace too small (need %s%lld)", "", (long long)need);
    cudaStream_t st = (cudaStream_t)stream;
    int64_t *off = (int64_t *)workspace_dev;
    double *scratch = (double *)((char *)workspace_dev + align_up((int64_t)sizeof(int64_t) * (n_reads + 1), 256));
    NTB_CUDA(cudaMemcpyAsync(off, read_off_host, sizeof(int64_t) * (n_reads + 1), cudaMemcpyHostToDevice, st));
    return launch_event_features(mean_dev, stdv_dev, length_dev, off, n_reads, total_events, pl, scratch, out_f64, out_dev, st);
}

extern "C" int ntb_event_features_host(const double *mean_host, const double *stdv_host, const double *length_host,
                                       const int64_t *read_off_host, int64_t n_reads, int64_t total_events,
                                       int32_t out_f64, void *out_host, void *stream) {
    if (!read_off_host || n_reads < 0 || total_events < 0)
        return fail(NTB_EINVAL, "ntb_event_features_host: bad argument%s", "");
    if (total_events > 0 && (!mean_host || !stdv_host || !length_host || !out_host))
        return fail(NTB_EINVAL, "ntb_event_features_host: null buffer%s", "");
    EventPlan pl;
    NTB_TRY(plan_events(read_off_host, n_reads, total_events, pl));
    NTB_TRY(check_device());
    if (n_reads == 0 || total_events == 0) return NTB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t in_b = (size_t)align_up((int64_t)sizeof(double) * total_events, 256);   // keeps every sub-buffer aligned
    const size_t out_b = (size_t)align_up((int64_t)(out_f64 ? 8 : 4) * 12 * total_events, 256);
    const size_t off_b = sizeof(int64_t) * (size_t)(n_reads + 1);
    unsigned char *buf = nullptr;
    NTB_CUDA(cudaMallocAsync((void **)&buf, 5 * in_b + out_b + off_b, st));
    double *m = (double *)buf, *s = (double *)(buf + in_b), *l = (double *)(buf + 2 * in_b), *scratch = (double *)(buf + 3 * in_b);
    unsigned char *o = buf + 5 * in_b;
    int64_t *off = (int64_t *)(o + out_b);
    const size_t in_n = sizeof(double) * (size_t)total_events, out_n = (out_f64 ? 8 : 4) * 12 * (size_t)total_events;
    cudaError_t e = cudaMemcpyAsync(m, mean_host, in_n, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(s, stdv_host, in_n, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(l, length_host, in_n, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(off, read_off_host, off_b, cudaMemcpyHostToDevice, st);
    int rc = e == cudaSuccess ? launch_event_features(m, s, l, off, n_reads, total_events, pl, scratch, out_f64, o, st)
                              : fail(NTB_ECUDA, "ntb_event_features_host: %s", cudaGetErrorString(e));
    if (rc == NTB_OK) {
        e = cudaMemcpyAsync(out_host, o, out_n, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) rc = fail(NTB_ECUDA, "ntb_event_features_host: %s", cudaGetErrorString(e));
    }
    cudaFreeAsync(buf, st);
    return rc;
}

// ---- consensus assembly (nanotensor/dataset.py:421-432; kernels in ntb_assembly.cuh) ----
namespace ntb {
struct AsmBufs {
    int64_t *read_off, *bound, *len, *bound_off;
    uint8_t *bases;
    int32_t *blen, *disp, *start, *win_read, *votes, *status;
};
static AsmBufs asm_carve(Arena &ar, int64_t n_windows, int64_t n_reads, int32_t T, int64_t capacity) {
    AsmBufs b;
    b.read_off = ar.take<int64_t>(n_reads + 1);
    b.bound = ar.take<int64_t>(n_reads);
    b.len = ar.take<int64_t>(n_reads);
    b.bound_off = ar.take<int64_t>(n_reads + 1);
    b.bases = ar.take<uint8_t>(n_windows * T);
    b.blen = ar.take<int32_t>(n_windows);
    b.disp = ar.take<int32_t>(n_windows);
    b.start = ar.take<int32_t>(n_windows);
    b.win_read = ar.take<int32_t>(n_windows);
    b.votes = ar.take<int32_t>(5 * capacity);
    b.status = ar.take<int32_t>(1);
    return b;
}
static int asm_check_table(const int64_t *read_off, int64_t n_reads, int64_t n_windows) {
    if (read_off[0] != 0 || read_off[n_reads] != n_windows)
        return fail(NTB_EINVAL, "ntb_assemble_reads: read_off must run from 0 to n_windows%s", "");
    for (int64_t r = 0; r < n_reads; ++r)
        if (read_off[r + 1] < read_off[r])
            return fail(NTB_EINVAL, "ntb_assemble_reads: read_off not monotonic at read %s%lld", "", (long long)r);
    return NTB_OK;
}
}  // namespace ntb

extern "C" int64_t ntb_assemble_workspace_bytes(int64_t n_windows, int64_t n_reads, int32_t T, int64_t capacity) {
    if (n_windows < 0 || n_reads < 0 || T <= 0 || capacity < 0) return 0;
    Arena ar(nullptr, 0);
    asm_carve(ar, n_windows, n_reads, T, capacity);
    return ar.off;
}

extern "C" int ntb_assemble_reads(const int32_t *labels_dev, const int64_t *read_off_host, int64_t n_reads,
                                  int64_t n_windows, int32_t T, int64_t capacity, void *workspace_dev,
                                  int64_t workspace_bytes, uint8_t *cons_dev, int64_t *cons_off_dev, int32_t *counts_dev,
                                  int32_t *status_dev, void *stream) {
    if (!read_off_host || !cons_off_dev || !status_dev || n_reads < 0 || n_windows < 0 || T <= 0 || capacity < 0 ||
        (n_windows > 0 && !labels_dev) || (capacity > 0 && !cons_dev))
        return fail(NTB_EINVAL, "ntb_assemble_reads: bad argument%s", "");
    NTB_TRY(asm_check_table(read_off_host, n_reads, n_windows));
    NTB_TRY(check_device());
    Arena ar(workspace_dev, workspace_bytes);
    AsmBufs b = asm_carve(ar, n_windows, n_reads, T, capacity);
    if (!workspace_dev || !ar.ok())
        return fail(NTB_ENOMEM, "ntb_assemble_reads: workspace too small (need %s%lld)", "", (long long)ar.off);
    cudaStream_t st = (cudaStream_t)stream;
    NTB_CUDA(cudaMemsetAsync(status_dev, 0, sizeof(int32_t), st));
    NTB_CUDA(cudaMemsetAsync(cons_off_dev, 0, sizeof(int64_t) * (n_reads + 1), st));
    if (n_reads == 0) return NTB_OK;
    NTB_CUDA(cudaMemcpyAsync(b.read_off, read_off_host, sizeof(int64_t) * (n_reads + 1), cudaMemcpyHostToDevice, st));
    NTB_CUDA(cudaMemsetAsync(b.votes, 0, sizeof(int32_t) * 5 * capacity, st));
    if (n_windows > 0) {
        const unsigned wgrid = (unsigned)cdiv(n_windows, 4);
        k_asm_compact<<<wgrid, 128, 0, st>>>(labels_dev, n_windows, T, b.bases, b.blen, status_dev);
        k_asm_window_reads<<<(unsigned)n_reads, 128, 0, st>>>(b.read_off, b.win_read);
        const size_t smem = asm_disp_smem(T);
        if (smem > 48 * 1024) cudaFuncSetAttribute(k_asm_disp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_asm_disp<<<(unsigned)cdiv(n_windows, ASM_WARPS), ASM_WARPS * 32, smem, st>>>(b.bases, b.blen, b.win_read, b.read_off,
                                                                                      n_windows, T, b.disp);
        if (T >= ASM_AUTOJUNK) {
            // windows with 200 and more called bases exist only from T = 200 on: difflib's autojunk regime, restated in full
            const size_t lsmem = asm_long_smem_per_warp(T) * ASM_WARPS;
            if (lsmem > 200 * 1024) return fail(NTB_EINVAL, "ntb_assemble_reads: T too large for the long-window kernel%s", "");
            if (lsmem > 48 * 1024) cudaFuncSetAttribute(k_asm_disp_long, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lsmem);
            const int64_t want = cdiv(n_windows, ASM_WARPS);
            k_asm_disp_long<<<(unsigned)(want < 148 * 8 ? want : 148 * 8), ASM_WARPS * 32, lsmem, st>>>(
                b.bases, b.blen, b.win_read, b.read_off, n_windows, T, b.disp);
        }
    }
    k_asm_scan<<<(unsigned)n_reads, 256, 0, st>>>(b.disp, b.blen, b.read_off, b.start, b.bound, b.len);
    k_asm_read_offsets<<<1, 1024, 0, st>>>(b.bound, b.len, n_reads, capacity, b.bound_off, cons_off_dev, status_dev);
    if (n_windows > 0)
        k_asm_vote<<<(unsigned)cdiv(n_windows, 4), 128, 0, st>>>(b.bases, b.blen, b.start, b.win_read, b.bound_off, b.len,
                                                                 n_windows, T, capacity, b.votes);
    k_asm_consensus<<<(unsigned)n_reads, 256, 0, st>>>(b.votes, b.bound_off, cons_off_dev, capacity, cons_dev, counts_dev, capacity);
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

extern "C" int ntb_assemble_reads_host(const int32_t *labels_host, const int64_t *read_off_host, int64_t n_reads,
                                       int64_t n_windows, int32_t T, int64_t capacity, uint8_t *cons_host,
                                       int64_t *cons_off_host, int32_t *counts_host, void *stream) {
    if (!read_off_host || !cons_off_host || n_reads < 0 || n_windows < 0 || T <= 0 || capacity < 0 ||
        (n_windows > 0 && !labels_host) || (capacity > 0 && !cons_host))
        return fail(NTB_EINVAL, "ntb_assemble_reads_host: bad argument%s", "");
    NTB_TRY(asm_check_table(read_off_host, n_reads, n_windows));
    NTB_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t ws_b = ntb_assemble_workspace_bytes(n_windows, n_reads, T, capacity);
    const int64_t lab_b = align_up((int64_t)sizeof(int32_t) * n_windows * T, 256), cons_b = align_up(capacity, 256);
    const int64_t off_b = align_up((int64_t)sizeof(int64_t) * (n_reads + 1), 256);
    const int64_t cnt_b = counts_host ? align_up((int64_t)sizeof(int32_t) * 5 * capacity, 256) : 0;
    char *buf = nullptr;
    NTB_CUDA(cudaMallocAsync((void **)&buf, ws_b + lab_b + cons_b + off_b + cnt_b + 256, st));
    int32_t *labels = (int32_t *)(buf + ws_b);
    uint8_t *cons = (uint8_t *)(buf + ws_b + lab_b);
    int64_t *cons_off = (int64_t *)(buf + ws_b + lab_b + cons_b);
    int32_t *counts = (int32_t *)(buf + ws_b + lab_b + cons_b + off_b);
    int32_t *status = (int32_t *)(buf + ws_b + lab_b + cons_b + off_b + cnt_b);
    int rc = NTB_OK;
    cudaError_t e = cudaMemcpyAsync(labels, labels_host, sizeof(int32_t) * n_windows * T, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess)
        rc = ntb_assemble_reads(labels, read_off_host, n_reads, n_windows, T, capacity, buf, ws_b, cons, cons_off,
                                counts_host ? counts : nullptr, status, st);
    int32_t status_h = 0;
    if (e == cudaSuccess && rc == NTB_OK) {
        e = cudaMemcpyAsync(cons_off_host, cons_off, sizeof(int64_t) * (n_reads + 1), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(&status_h, status, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e == cudaSuccess && status_h == 0) {
            const int64_t total = cons_off_host[n_reads];
            e = cudaMemcpyAsync(cons_host, cons, (size_t)total, cudaMemcpyDeviceToHost, st);
            for (int c = 0; counts_host && c < 5 && e == cudaSuccess; ++c)     // compact [5, total] like the reference's matrix
                e = cudaMemcpyAsync(counts_host + c * total, counts + c * capacity, sizeof(int32_t) * total,
                                    cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
    }
    cudaFreeAsync(buf, st);
    if (e != cudaSuccess) return fail(NTB_ECUDA, "ntb_assemble_reads_host: %s", cudaGetErrorString(e));
    if (rc != NTB_OK) return rc;
    if (status_h == ASM_STATUS_LONG_WINDOW)
        return fail(NTB_EINVAL, "ntb_assemble_reads: a window has >= 200 called bases (difflib autojunk regime, not built)%s", "");
    if (status_h == ASM_STATUS_CAPACITY)
        return fail(NTB_ENOMEM, "ntb_assemble_reads: capacity smaller than the number of called bases%s", "");
    return NTB_OK;
}

// ---- event detection (nanonet/nanonet/eventdetection/filters.py:299-322; kernels in ntb_eventdetect.cuh) ----
namespace ntb {
struct EdBufs {
    int64_t *read_off, *n_events, *tmp_off;
    double *sums, *sumsqs, *t_short, *t_long;
    int32_t *peaks;
};
static EdBufs ed_carve(Arena &ar, int64_t n_reads, int64_t total) {
    EdBufs b;
    b.read_off = ar.take<int64_t>(n_reads + 1);
    b.n_events = ar.take<int64_t>(n_reads);
    b.tmp_off = ar.take<int64_t>(n_reads + 1);
    b.sums = ar.take<double>(2 * total);       // sums | sumsqs in one block: the fused form keeps (sum, sumsq) pairs there
    b.sumsqs = b.sums + total;
    b.t_short = ar.take<double>(total);
    b.t_long = ar.take<double>(total);
    b.peaks = ar.take<int32_t>(total);
    return b;
}
static int ed_check(const int64_t *read_off, int64_t n_reads, int64_t total, const int32_t *w, const char *who) {
    if (read_off[0] != 0 || read_off[n_reads] != total) return fail(NTB_EINVAL, "%s: read_off must run from 0 to total_samples", who);
    for (int64_t r = 0; r < n_reads; ++r) {
        if (read_off[r + 1] < read_off[r]) return fail(NTB_EINVAL, "%s: read_off not monotonic", who);
        if (read_off[r + 1] - read_off[r] > 0x7fffffff) return fail(NTB_EINVAL, "%s: read longer than 2^31 samples", who);
    }
    if (w[0] < 1 || w[1] < 1) return fail(NTB_EINVAL, "%s: window lengths must be positive", who);
    return NTB_OK;
}
}  // namespace ntb

extern "C" int64_t ntb_event_detect_workspace_bytes(int64_t n_reads, int64_t total_samples) {
    if (n_reads < 0 || total_samples < 0) return 0;
    Arena ar(nullptr, 0);
    ed_carve(ar, n_reads, total_samples);
    return ar.off;
}

// warps the 8-reads-per-warp form needs to beat the 4-reads form (profiles/r02_ed_lanes_sweep.txt: 32 768 ... 131 072 reads
// -> R = 8 at 25-26 G samples/s, ahead of 16 and 32; 16 384 and 8192 reads -> R = 4 at 20.7 / 15.8 G; 1024 reads -> one warp per read)
static const int64_t ED_MIN_WARPS = 3000;

extern "C" int ntb_event_detect(const double *raw_dev, const int64_t *read_off_host, int64_t n_reads, int64_t total_samples,
                                double sample_rate, const int32_t *window_lengths, const double *thresholds,
                                double peak_height, void *workspace_dev, int64_t workspace_bytes, int64_t capacity,
                                double *events_dev, int64_t *event_off_dev, int32_t *status_dev, void *stream) {
    if (!read_off_host || !window_lengths || !thresholds || !event_off_dev || !status_dev || n_reads < 0 ||
        total_samples < 0 || capacity < 0 || (total_samples > 0 && !raw_dev) || (capacity > 0 && !events_dev))
        return fail(NTB_EINVAL, "ntb_event_detect: bad argument%s", "");
    NTB_TRY(ed_check(read_off_host, n_reads, total_samples, window_lengths, "ntb_event_detect"));
    NTB_TRY(check_device());
    Arena ar(workspace_dev, workspace_bytes);
    EdBufs b = ed_carve(ar, n_reads, total_samples);
    if (!workspace_dev || !ar.ok())
        return fail(NTB_ENOMEM, "ntb_event_detect: workspace too small (need %s%lld)", "", (long long)ar.off);
    cudaStream_t st = (cudaStream_t)stream;
    NTB_CUDA(cudaMemsetAsync(status_dev, 0, sizeof(int32_t), st));
    NTB_CUDA(cudaMemsetAsync(event_off_dev, 0, sizeof(int64_t) * (n_reads + 1), st));
    if (n_reads == 0) return NTB_OK;
    // the shorter window is the dominant detector, whatever the order given (filters.py:232-233)
    const int z = window_lengths[0] > window_lengths[1] ? 1 : 0;
    const int w_short = window_lengths[z], w_long = window_lengths[1 - z];
    const double th_short = thresholds[z], th_long = thresholds[1 - z];
    NTB_CUDA(cudaMemcpyAsync(b.read_off, read_off_host, sizeof(int64_t) * (n_reads + 1), cudaMemcpyHostToDevice, st));
    const unsigned wgrid = (unsigned)cdiv(n_reads, 4);
    // sequential stages: a lane per read once there are enough reads to fill the machine that way (32 reads per warp:
    // measured 26 ms vs 39 ms at 32 768 reads, but 78 ms vs 41 ms at 8192), else a warp per read
    // (NTB_ED_LANES=0/1 forces one form: A/B and tests)
    static const char *ed_env = getenv("NTB_ED_LANES");
    // reads per warp of the sequential stages: 0 = one warp per read, else R lanes of a warp walk R reads.  Default: the
    // 8 when that still leaves ED_MIN_WARPS warps, 4 from 4000 reads (measured, tools/sweep_ed_lanes.sh); NTB_ED_LANES=0/1/4/8/16/32 forces
    // a form (1 = 32): A/B runs and tests
    int R = 0;
    if (ed_env) {
        R = atoi(ed_env);
        R = R == 1 ? 32 : (R == 4 || R == 8 || R == 16 || R == 32 ? R : 0);
    } else {
        R = n_reads / 8 >= ED_MIN_WARPS ? 8 : (n_reads >= 4000 ? 4 : 0);
    }
    const unsigned lgrid = (unsigned)cdiv(n_reads, 2 * (R ? R : 1));
    // Opt-in (NTB_ED_FUSED=1, with 4 or 8 reads per warp and windows that fit the ring): cumulative sums, t-statistics and
    // peaks in ONE kernel.  It moves 2.4x fewer bytes (the t-statistics never reach memory) but measured SLOWER than the
    // separate kernels (27.2 vs 25.9 ms at 8192 reads, 26.9 vs 16.2 ms at 32 768): the t-statistics' float64 divisions run
    // at the fused kernel's low occupancy and in series with the sequential stages, so the default stays unfused.
    static const char *fused_env = getenv("NTB_ED_FUSED");
    if ((R == 4 || R == 8) && w_long <= 31 && fused_env && fused_env[0] == '1') {
        double2 *sq = reinterpret_cast<double2 *>(b.sums);     // sums | sumsqs are adjacent in the workspace: one array of pairs
        if (R == 8)
            k_ed_fused<8><<<lgrid, 64, 0, st>>>(raw_dev, b.read_off, n_reads, w_short, w_long, th_short, th_long, peak_height, sq,
                                                b.peaks, b.n_events);
        else
            k_ed_fused<4><<<lgrid, 64, 0, st>>>(raw_dev, b.read_off, n_reads, w_short, w_long, th_short, th_long, peak_height, sq,
                                                b.peaks, b.n_events);
        k_asm_read_offsets<<<1, 1024, 0, st>>>(b.n_events, b.n_events, n_reads, capacity, b.tmp_off, event_off_dev, status_dev);
        k_ed_events_sq<<<(unsigned)n_reads, 256, 0, st>>>(sq, b.peaks, b.read_off, event_off_dev, sample_rate, capacity, events_dev);
        NTB_LAUNCH_CHECK();
        return NTB_OK;
    }
    switch (R) {
        case 32: k_ed_prefix_t<32><<<lgrid, 64, 0, st>>>(raw_dev, b.read_off, n_reads, b.sums, b.sumsqs); break;
        case 16: k_ed_prefix_t<16><<<lgrid, 64, 0, st>>>(raw_dev, b.read_off, n_reads, b.sums, b.sumsqs); break;
        case 8: k_ed_prefix_t<8><<<lgrid, 64, 0, st>>>(raw_dev, b.read_off, n_reads, b.sums, b.sumsqs); break;
        case 4: k_ed_prefix_t<4><<<lgrid, 64, 0, st>>>(raw_dev, b.read_off, n_reads, b.sums, b.sumsqs); break;
        default: k_ed_prefix<<<wgrid, 128, 0, st>>>(raw_dev, b.read_off, n_reads, b.sums, b.sumsqs);
    }
    const int64_t per_read = total_samples / n_reads + 1;
    const unsigned ychunks = (unsigned)std::min<int64_t>(std::max<int64_t>(per_read / 4096, 1), 64);
    k_ed_tstat<<<dim3((unsigned)n_reads, ychunks), 256, 0, st>>>(b.sums, b.sumsqs, b.read_off, w_short, w_long, b.t_short, b.t_long);
#define NTB_ED_PEAKS(RR) k_ed_peaks_t<RR><<<lgrid, 64, 0, st>>>(b.t_short, b.t_long, b.read_off, n_reads, w_short, w_long, \
                                                                th_short, th_long, peak_height, b.peaks, b.n_events)
    switch (R) {
        case 32: NTB_ED_PEAKS(32); break;
        case 16: NTB_ED_PEAKS(16); break;
        case 8: NTB_ED_PEAKS(8); break;
        case 4: NTB_ED_PEAKS(4); break;
        default:
            k_ed_peaks<<<wgrid, 128, 0, st>>>(b.t_short, b.t_long, b.read_off, n_reads, w_short, w_long, th_short, th_long,
                                              peak_height, b.peaks, b.n_events);
    }
#undef NTB_ED_PEAKS
    k_asm_read_offsets<<<1, 1024, 0, st>>>(b.n_events, b.n_events, n_reads, capacity, b.tmp_off, event_off_dev, status_dev);
    k_ed_events<<<(unsigned)n_reads, 256, 0, st>>>(b.sums, b.sumsqs, b.peaks, b.read_off, event_off_dev, sample_rate, capacity,
                                                   events_dev);
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

extern "C" int ntb_event_detect_host(const double *raw_host, const int64_t *read_off_host, int64_t n_reads,
                                     int64_t total_samples, double sample_rate, const int32_t *window_lengths,
                                     const double *thresholds, double peak_height, int64_t capacity, double *events_host,
                                     int64_t *event_off_host, void *stream) {
    if (!read_off_host || !window_lengths || !thresholds || !event_off_host || n_reads < 0 || total_samples < 0 ||
        capacity < 0 || (total_samples > 0 && !raw_host) || (capacity > 0 && !events_host))
        return fail(NTB_EINVAL, "ntb_event_detect_host: bad argument%s", "");
    NTB_TRY(ed_check(read_off_host, n_reads, total_samples, window_lengths, "ntb_event_detect_host"));
    NTB_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t ws_b = ntb_event_detect_workspace_bytes(n_reads, total_samples);
    const int64_t raw_b = align_up((int64_t)sizeof(double) * total_samples, 256), ev_b = align_up((int64_t)sizeof(double) * 4 * capacity, 256);
    const int64_t off_b = align_up((int64_t)sizeof(int64_t) * (n_reads + 1), 256);
    char *buf = nullptr;
    NTB_CUDA(cudaMallocAsync((void **)&buf, ws_b + raw_b + ev_b + off_b + 256, st));
    double *raw = (double *)(buf + ws_b), *ev = (double *)(buf + ws_b + raw_b);
    int64_t *ev_off = (int64_t *)(buf + ws_b + raw_b + ev_b);
    int32_t *status = (int32_t *)(buf + ws_b + raw_b + ev_b + off_b);
    int rc = NTB_OK;
    int32_t status_h = 0;
    cudaError_t e = cudaMemcpyAsync(raw, raw_host, sizeof(double) * total_samples, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess)
        rc = ntb_event_detect(raw, read_off_host, n_reads, total_samples, sample_rate, window_lengths, thresholds, peak_height,
                              buf, ws_b, capacity, ev, ev_off, status, st);
    if (e == cudaSuccess && rc == NTB_OK) {
        e = cudaMemcpyAsync(event_off_host, ev_off, sizeof(int64_t) * (n_reads + 1), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(&status_h, status, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e == cudaSuccess && status_h == 0) {
            e = cudaMemcpyAsync(events_host, ev, sizeof(double) * 4 * event_off_host[n_reads], cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
    }
    cudaFreeAsync(buf, st);
    if (e != cudaSuccess) return fail(NTB_ECUDA, "ntb_event_detect_host: %s", cudaGetErrorString(e));
    if (rc != NTB_OK) return rc;
    if (status_h) return fail(NTB_ENOMEM, "ntb_event_detect: capacity smaller than the number of events%s", "");
    return NTB_OK;
}

// ---- .signal text parse (chiron/chiron/chiron_input.py:220-226; kernels in ntb_parse.cuh) ----
namespace ntb {
struct ParseBufs {
    int64_t *file_off, *first_chunk, *chunk_lo, *counts, *chunk_out, *tmp;
    int32_t *chunk_file;
};
static ParseBufs parse_carve(Arena &ar, int64_t n_files, int64_t max_chunks) {
    ParseBufs b;
    b.file_off = ar.take<int64_t>(n_files + 1);
    b.first_chunk = ar.take<int64_t>(n_files + 1);
    b.chunk_lo = ar.take<int64_t>(max_chunks);
    b.counts = ar.take<int64_t>(max_chunks);
    b.chunk_out = ar.take<int64_t>(max_chunks + 1);
    b.tmp = ar.take<int64_t>(max_chunks + 1);
    b.chunk_file = ar.take<int32_t>(max_chunks);
    return b;
}
static int64_t parse_max_chunks(int64_t total_bytes, int64_t n_files) { return total_bytes / PARSE_CHUNK + n_files + 1; }
}  // namespace ntb

extern "C" int64_t ntb_parse_signal_workspace_bytes(int64_t total_bytes, int64_t n_files) {
    if (total_bytes < 0 || n_files < 0) return 0;
    Arena ar(nullptr, 0);
    parse_carve(ar, n_files, parse_max_chunks(total_bytes, n_files));
    return ar.off;
}

extern "C" int ntb_parse_signal(const uint8_t *text_dev, const int64_t *file_off_host, int64_t n_files, int64_t total_bytes,
                                void *workspace_dev, int64_t workspace_bytes, int64_t capacity, int16_t *samples_dev,
                                int64_t *read_off_dev, int32_t *status_dev, void *stream) {
    if (!file_off_host || !read_off_dev || !status_dev || n_files < 0 || total_bytes < 0 || capacity < 0 ||
        (total_bytes > 0 && !text_dev) || (capacity > 0 && !samples_dev))
        return fail(NTB_EINVAL, "ntb_parse_signal: bad argument%s", "");
    if (file_off_host[0] != 0 || file_off_host[n_files] != total_bytes)
        return fail(NTB_EINVAL, "ntb_parse_signal: file_off must run from 0 to total_bytes%s", "");
    if ((uintptr_t)text_dev & 15) return fail(NTB_EINVAL, "ntb_parse_signal: text_dev must be 16-byte aligned%s", "");
    std::vector<int64_t> chunk_lo, first_chunk((size_t)n_files + 1);
    std::vector<int32_t> chunk_file;
    for (int64_t f = 0; f < n_files; ++f) {
        if (file_off_host[f + 1] < file_off_host[f])
            return fail(NTB_EINVAL, "ntb_parse_signal: file_off not monotonic at file %s%lld", "", (long long)f);
        first_chunk[f] = (int64_t)chunk_lo.size();
        for (int64_t b = file_off_host[f]; b < file_off_host[f + 1]; b += PARSE_CHUNK) {
            chunk_lo.push_back(b);
            chunk_file.push_back((int32_t)f);
        }
    }
    const int64_t n_chunks = (int64_t)chunk_lo.size();
    first_chunk[n_files] = n_chunks;
    NTB_TRY(check_device());
    Arena ar(workspace_dev, workspace_bytes);
    ParseBufs b = parse_carve(ar, n_files, parse_max_chunks(total_bytes, n_files));
    if (!workspace_dev || !ar.ok())
        return fail(NTB_ENOMEM, "ntb_parse_signal: workspace too small (need %s%lld)", "", (long long)ar.off);
    cudaStream_t st = (cudaStream_t)stream;
    NTB_CUDA(cudaMemsetAsync(status_dev, 0, sizeof(int32_t), st));
    NTB_CUDA(cudaMemsetAsync(read_off_dev, 0, sizeof(int64_t) * (n_files + 1), st));
    if (n_chunks == 0) return NTB_OK;
    // the tables are small (16 + 4 bytes per 4 KiB of text); pageable copies complete before the vectors go away
    NTB_CUDA(cudaMemcpyAsync(b.file_off, file_off_host, sizeof(int64_t) * (n_files + 1), cudaMemcpyHostToDevice, st));
    NTB_CUDA(cudaMemcpyAsync(b.first_chunk, first_chunk.data(), sizeof(int64_t) * (n_files + 1), cudaMemcpyHostToDevice, st));
    NTB_CUDA(cudaMemcpyAsync(b.chunk_lo, chunk_lo.data(), sizeof(int64_t) * n_chunks, cudaMemcpyHostToDevice, st));
    NTB_CUDA(cudaMemcpyAsync(b.chunk_file, chunk_file.data(), sizeof(int32_t) * n_chunks, cudaMemcpyHostToDevice, st));
    NTB_CUDA(cudaStreamSynchronize(st));
    static const bool two_pass = ntb_env("NTB_PARSE_2PASS") != nullptr;      // ablation builds: the round-1 pair of passes
    if (two_pass) {
        k_parse_count<<<(unsigned)n_chunks, PARSE_THREADS, 0, st>>>(text_dev, b.chunk_lo, b.chunk_file, b.file_off, total_bytes, b.counts);
        k_asm_read_offsets<<<1, 1024, 0, st>>>(b.counts, b.counts, n_chunks, capacity, b.tmp, b.chunk_out, status_dev);
        k_parse_emit<<<(unsigned)n_chunks, PARSE_THREADS, 0, st>>>(text_dev, b.chunk_lo, b.chunk_file, b.file_off, total_bytes,
                                                                   b.chunk_out, capacity, samples_dev, status_dev);
    } else {
        // one pass: state words (in `counts`) and the ticket (in `tmp`) start at zero
        NTB_CUDA(cudaMemsetAsync(b.counts, 0, sizeof(int64_t) * n_chunks, st));
        NTB_CUDA(cudaMemsetAsync(b.tmp, 0, sizeof(int64_t), st));
        k_parse_fused<<<(unsigned)n_chunks, PARSE_THREADS, 0, st>>>(text_dev, b.chunk_lo, b.chunk_file, b.file_off, total_bytes, n_chunks,
                                                                   reinterpret_cast<unsigned long long *>(b.counts),
                                                                   reinterpret_cast<unsigned int *>(b.tmp), capacity, b.chunk_out,
                                                                   samples_dev, status_dev);
    }
    k_parse_read_off<<<(unsigned)cdiv(n_files + 1, 256), 256, 0, st>>>(b.chunk_out, b.first_chunk, n_files, read_off_dev);
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

extern "C" int ntb_parse_signal_host(const uint8_t *text_host, const int64_t *file_off_host, int64_t n_files,
                                     int64_t total_bytes, int64_t capacity, int16_t *samples_host, int64_t *read_off_host,
                                     void *stream) {
    if (!file_off_host || !read_off_host || n_files < 0 || total_bytes < 0 || capacity < 0 || (total_bytes > 0 && !text_host) ||
        (capacity > 0 && !samples_host))
        return fail(NTB_EINVAL, "ntb_parse_signal_host: bad argument%s", "");
    NTB_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t ws_b = ntb_parse_signal_workspace_bytes(total_bytes, n_files);
    const int64_t txt_b = align_up(total_bytes, 256), smp_b = align_up((int64_t)sizeof(int16_t) * capacity, 256);
    const int64_t off_b = align_up((int64_t)sizeof(int64_t) * (n_files + 1), 256);
    char *buf = nullptr;
    NTB_CUDA(cudaMallocAsync((void **)&buf, ws_b + txt_b + smp_b + off_b + 256, st));
    uint8_t *text = (uint8_t *)(buf + ws_b);
    int16_t *samples = (int16_t *)(buf + ws_b + txt_b);
    int64_t *read_off = (int64_t *)(buf + ws_b + txt_b + smp_b);
    int32_t *status = (int32_t *)(buf + ws_b + txt_b + smp_b + off_b);
    int rc = NTB_OK;
    int32_t status_h = 0;
    cudaError_t e = cudaMemcpyAsync(text, text_host, (size_t)total_bytes, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess)
        rc = ntb_parse_signal(text, file_off_host, n_files, total_bytes, buf, ws_b, capacity, samples, read_off, status, st);
    if (e == cudaSuccess && rc == NTB_OK) {
        e = cudaMemcpyAsync(read_off_host, read_off, sizeof(int64_t) * (n_files + 1), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(&status_h, status, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e == cudaSuccess && status_h == 0) {
            e = cudaMemcpyAsync(samples_host, samples, sizeof(int16_t) * read_off_host[n_files], cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
    }
    cudaFreeAsync(buf, st);
    if (e != cudaSuccess) return fail(NTB_ECUDA, "ntb_parse_signal_host: %s", cudaGetErrorString(e));
    if (rc != NTB_OK) return rc;
    if (status_h == PARSE_STATUS_CAPACITY) return fail(NTB_ENOMEM, "ntb_parse_signal: capacity smaller than the number of samples%s", "");
    if (status_h == PARSE_STATUS_BADCHAR) return fail(NTB_EINVAL, "ntb_parse_signal: a token is not [+-]?[0-9]{1,6}%s", "");
    if (status_h == PARSE_STATUS_RANGE) return fail(NTB_EINVAL, "ntb_parse_signal: a sample does not fit int16%s", "");
    return NTB_OK;
}

extern "C" int64_t ntb_conv_bn_workspace_bytes(int64_t n_windows, int64_t n_groups, int32_t T, int32_t Cin,
                                               int32_t Cout, int32_t ksize) {
    (void)Cin; (void)ksize;
    Arena ar(nullptr, 0);
    ar.take<int32_t>(n_windows);
    ar.take<float>(n_windows * T * (int64_t)Cout);
    ar.take<float>(n_groups * (int64_t)Cout);
    ar.take<float>(n_groups * (int64_t)Cout);
    return ar.off;
}

extern "C" int ntb_conv_bn(const float *in_dev, const float *filt_dev, const float *scale_dev,
                           const float *offset_dev, const int32_t *group_off_dev, int64_t n_windows,
                           int64_t n_groups, int32_t T, int32_t Cin, int32_t Cout, int32_t ksize, int32_t relu,
                           void *workspace_dev, int64_t workspace_bytes, float *out_dev, void *stream) {
    if (!in_dev || !filt_dev || !out_dev || (ksize != 1 && ksize != 3) || T <= 0 || Cin <= 0 || Cout <= 0)
        return fail(NTB_EINVAL, "ntb_conv_bn: bad argument%s", "");
    if ((scale_dev == nullptr) != (offset_dev == nullptr)) return fail(NTB_EINVAL, "ntb_conv_bn: scale/offset must both be set%s", "");
    if (scale_dev && (!group_off_dev || n_groups <= 0)) return fail(NTB_EINVAL, "ntb_conv_bn: BN needs group_off%s", "");
    if (scale_dev && (Cout % 4)) return fail(NTB_EINVAL, "ntb_conv_bn: BN needs Cout %% 4 == 0%s", "");
    NTB_TRY(check_device());
    Ctx cx{(cudaStream_t)stream, nullptr};
    Arena ar(workspace_dev, workspace_bytes);
    int32_t *win_group = ar.take<int32_t>(n_windows);
    float *y = ar.take<float>(n_windows * T * (int64_t)Cout);
    float *a = ar.take<float>(n_groups * (int64_t)Cout);
    float *b = ar.take<float>(n_groups * (int64_t)Cout);
    if (!workspace_dev || !ar.ok()) return fail(NTB_ENOMEM, "ntb_conv_bn: workspace too small (need %s%lld)", "", (long long)ar.off);
    const int64_t rows = n_windows * T;
    if (rows == 0) return NTB_OK;
    ConvGemmParams p{};
    p.A = in_dev; p.lda = Cin; p.W = filt_dev; p.out = scale_dev ? y : out_dev; p.ldo = Cout;
    p.rows = rows; p.K = Cin; p.N = Cout; p.taps = ksize; p.T = T;
    NTB_TRY(launch_conv_gemm(cx, p));
    if (scale_dev) {
        NTB_CUDA(cudaMemsetAsync(win_group, 0, sizeof(int32_t) * (size_t)n_windows, cx.st));
        k_window_groups<<<(unsigned)n_groups, 128, 0, cx.st>>>(group_off_dev, (int)n_groups, n_windows, win_group);
        NTB_TRY(launch_bn_stats(cx, y, Cout, T, group_off_dev, n_groups, scale_dev, offset_dev, a, b));
        k_bn_apply<<<elementwise_grid(rows * Cout, 148), 256, 0, cx.st>>>(y, a, b, win_group, T, Cout, rows * Cout, relu, out_dev);
    } else if (relu) {
        k_bn_apply<<<elementwise_grid(rows * Cout, 148), 256, 0, cx.st>>>(out_dev, nullptr, nullptr, nullptr, T, Cout, rows * Cout, 1, out_dev);
    }
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

namespace ntb {
struct BlstmOpBufs {
    float *wih, *bias, *whh, *xproj;
    bool tc;
    bool step = false;
    GemmWeights whhw;
    float *whh_stage = nullptr, *cstate = nullptr;
    GemmWeights wihw;
    bool step_fused = false;
    GemmWeights wcatw;
    float *wcat_stage = nullptr;
    GemmScratch gs;
    op16_t *whh_hi, *whh_lo;
    bool fused = false;           // the model's default for these shapes: ONE fused kernel (ntb_blstm_fused.cuh)
    BlstmGeom fg;
    op16_t *f_wih = nullptr, *f_whh = nullptr, *fx_hi = nullptr, *fx_lo = nullptr;
    float *f_bias = nullptr;
};
static int64_t plan_blstm_op(Arena &ar, int64_t n, int T, int D, int H, BlstmOpBufs &b) {
    const int Hp = pad_hidden(H);
    b.fused = Hp == H && lstm_tc_supported(H) && D % 8 == 0 && D >= 64 && !ntb_env("NTB_NO_FUSED_BLSTM") &&
              blstm_fused_geom(H, D, b.fg);
    if (b.fused) {
        b.f_wih = ar.take<op16_t>(blstm_wih_img_elems(b.fg));
        b.f_whh = ar.take<op16_t>(blstm_whh_img_elems(b.fg));
        b.f_bias = ar.take<float>(8 * H);
        b.fx_hi = ar.take<op16_t>(blstm_plane_elems(n, T, D));
        b.fx_lo = ar.take<op16_t>(blstm_plane_elems(n, T, D));
        return ar.off;
    }
    b.wih = ar.take<float>((int64_t)D * 8 * Hp);
    b.bias = ar.take<float>(8 * Hp);
    b.whh = ar.take<float>(2LL * H * Hp * 4);
    gemm_tc_plan_weights(b.wihw, 1, D, 8 * Hp, ar);
    b.tc = b.wihw.valid && Hp == H && lstm_tc_supported(H);
    const int64_t rows = n * T, tile_rows = align_up(n, LSTM_TC_ROWS) * T;
    b.xproj = ar.take<float>((tile_rows > rows ? tile_rows : rows) * 8LL * Hp);
    b.whh_hi = b.whh_lo = nullptr;
    b.gs = GemmScratch();
    if (b.tc) {
        b.whh_hi = ar.take<op16_t>(lstm_tc_whh_elems(H));
        b.whh_lo = ar.take<op16_t>(lstm_tc_whh_elems(H));
        gemm_tc_plan_scratch(b.gs, rows, T, D, ar, tile_rows);
    } else if (Hp == H && gemm_tc_lstm_supported(H)) {
        gemm_tc_plan_whh(b.whhw, H, ar, &b.whh_stage);
        b.cstate = ar.take<float>(gemm_tc_lstm_state_floats(n, H));
        const int kcap = D > H ? D : H;
        const int64_t step_rows = cdiv(4 * align_up(n, 256) * H, align_up(kcap, 8));
        gemm_tc_plan_scratch(b.gs, rows, T, kcap, ar, step_rows > tile_rows ? step_rows : tile_rows);
        b.step = b.whhw.valid && b.gs.valid;
        if (b.step && gemm_tc_lstm_fused_supported(H, D) && !ntb_env("NTB_NO_FUSED_STEP")) {
            gemm_tc_plan_wcat(b.wcatw, H, D, ar, &b.wcat_stage);
            b.step_fused = b.wcatw.valid;
        }
    }
    return ar.off;
}
}  // namespace ntb

extern "C" int64_t ntb_blstm_workspace_bytes(int64_t n_windows, int32_t T, int32_t D, int32_t H) {
    Arena ar(nullptr, 0);
    BlstmOpBufs b;
    return plan_blstm_op(ar, n_windows, T, D, H, b);
}

extern "C" int ntb_blstm(const float *in_dev, const int32_t *len_dev, const float *kernel_fw_dev,
                         const float *bias_fw_dev, const float *kernel_bw_dev, const float *bias_bw_dev,
                         float forget_bias, int64_t n_windows, int32_t T, int32_t D, int32_t H,
                         void *workspace_dev, int64_t workspace_bytes, float *out_dev, void *stream) {
    if (!in_dev || !len_dev || !kernel_fw_dev || !bias_fw_dev || !kernel_bw_dev || !bias_bw_dev || !out_dev ||
        T <= 0 || D <= 0 || H <= 0)
        return fail(NTB_EINVAL, "ntb_blstm: bad argument%s", "");
    NTB_TRY(check_device());
    Ctx cx{(cudaStream_t)stream, nullptr};
    const int Hp = pad_hidden(H);
    Arena ar(workspace_dev, workspace_bytes);
    BlstmOpBufs b;
    plan_blstm_op(ar, n_windows, T, D, H, b);
    if (!workspace_dev || !ar.ok()) return fail(NTB_ENOMEM, "ntb_blstm: workspace too small (need %s%lld)", "", (long long)ar.off);
    if (n_windows == 0) return NTB_OK;
    if (b.fused) {
        k_pack_blstm_fused<<<512, 256, 0, cx.st>>>(kernel_fw_dev, bias_fw_dev, kernel_bw_dev, bias_bw_dev, forget_bias, b.fg,
                                                   b.f_wih, b.f_whh, b.f_bias);
        NTB_LAUNCH_CHECK();
        NTB_TRY(gemm_tc_split_kmajor(cx.st, in_dev, D, n_windows, T, D, b.fx_hi, b.fx_lo, &cx.launches));
        int sms = 148;
        { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
        return launch_blstm_fused(cx, b.fg, b.fx_hi, b.fx_lo, b.f_wih, b.f_whh, b.f_bias, len_dev, n_windows, T, out_dev,
                                  2 * H, nullptr, nullptr, sms);
    }
    k_repack_lstm<<<256, 256, 0, cx.st>>>(kernel_fw_dev, bias_fw_dev, kernel_bw_dev, bias_bw_dev, forget_bias, D, H, Hp,
                                          b.wih, b.bias, b.whh, ((b.tc && b.gs.valid) || b.step) ? 1 : 0);
    NTB_LAUNCH_CHECK();
    if (b.tc && b.gs.valid) {
        // same kernels the model uses for H = 100: tcgen05 projection + tensor-core recurrence
        NTB_TRY(gemm_tc_pack_weights(cx.st, b.wih, b.wihw));
        k_pack_whh_tc<<<256, 256, 0, cx.st>>>(kernel_fw_dev, kernel_bw_dev, D, H, (int)align_up(H, 16), b.whh_hi, b.whh_lo);
        NTB_LAUNCH_CHECK();
        NTB_TRY(gemm_tc_run_xproj(cx.st, in_dev, D, n_windows, T, b.wihw, b.bias, b.xproj, H, b.gs, &cx.launches));
        LstmTcParams p{};
        p.xproj = (const float4 *)b.xproj; p.whh_hi = b.whh_hi; p.whh_lo = b.whh_lo; p.lens = len_dev;
        p.out = out_dev; p.ldo = 2 * H; p.n_windows = n_windows; p.n_tiles = cdiv(n_windows, LSTM_TC_ROWS);
        p.T = T; p.H = H; p.Hp = (int)align_up(H, 16);
        return launch_lstm_tc(cx, p);
    }
    if (b.step) {
        // H beyond the persistent kernel: one tcgen05 GEMM per timestep (same path the model takes)
        if (b.wihw.valid) NTB_TRY(gemm_tc_pack_weights(cx.st, b.wih, b.wihw));
        NTB_TRY(gemm_tc_pack_whh(cx.st, kernel_fw_dev, kernel_bw_dev, D, H, b.whh_stage, b.whhw));
        if (b.step_fused) {
            NTB_TRY(gemm_tc_pack_wcat(cx.st, b.wih, kernel_fw_dev, kernel_bw_dev, D, H, b.wcat_stage, b.wcatw));
            const int Dp = gemm_tc_lstm_fused_dpad(D);
            op16_t *x_hi = reinterpret_cast<op16_t *>(b.xproj);
            op16_t *x_lo = x_hi + align_up(n_windows, LSTM_TC_ROWS) * T * (int64_t)Dp;
            NTB_TRY(gemm_tc_split_timemajor(cx.st, in_dev, D, n_windows, T, D, Dp, x_hi, x_lo, &cx.launches));
            return gemm_tc_lstm_layer(cx.st, b.wcatw, nullptr, 8 * Hp, len_dev, n_windows, T, H, b.cstate, out_dev, 2 * H, b.gs,
                                      &cx.launches, Dp, x_hi, x_lo, b.bias);
        }
        NTB_TRY(step_rec_projection(cx, in_dev, D, n_windows, T, H, b.wih, b.bias, b.wihw, b.xproj, b.gs));
        return gemm_tc_lstm_layer(cx.st, b.whhw, b.xproj, 8 * Hp, len_dev, n_windows, T, H, b.cstate, out_dev, 2 * H, b.gs,
                                  &cx.launches);
    }
    NTB_TRY(dense(cx, in_dev, D, n_windows * T, D, 8 * Hp, b.wih, nullptr, b.bias, b.xproj, 8 * Hp, nullptr));
    LstmParams p{};
    p.xproj = b.xproj; p.ldx = 8 * Hp; p.whh = b.whh; p.lens = len_dev; p.out = out_dev; p.ldo = 2 * H;
    p.n_windows = n_windows; p.T = T; p.H = H; p.Hp = Hp;
    return launch_lstm(cx, p, 148);
}

extern "C" int ntb_decode(const float *logits_dev, const int32_t *len_dev, int64_t n_windows, int32_t T, int32_t C,
                          int32_t graph, int32_t *labels_dev, int32_t *label_len_dev, void *stream) {
    if (!logits_dev || !labels_dev || T <= 0 || C <= 0 || (graph == NTB_GRAPH_CTC && !len_dev))
        return fail(NTB_EINVAL, "ntb_decode: bad argument%s", "");
    NTB_TRY(check_device());
    if (n_windows == 0) return NTB_OK;
    const int64_t warps = graph == NTB_GRAPH_CROSS_ENTROPY ? n_windows * T : n_windows;
    k_decode<<<cdiv(warps, 8), 256, 0, (cudaStream_t)stream>>>(logits_dev, len_dev, n_windows, T, C, graph, labels_dev, label_len_dev, C);
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

// ---- CTC decoders (kernels in ntb_ctc.cuh) ----
extern "C" int ntb_ctc_greedy(const float *logits_dev, const int32_t *len_dev, int64_t n_windows, int32_t T, int32_t C,
                              int32_t merge_repeated, int32_t *labels_dev, int32_t *label_len_dev,
                              float *neg_sum_logits_dev, void *stream) {
    if (!logits_dev || !len_dev || !labels_dev || n_windows < 0 || T <= 0 || C < 2)
        return fail(NTB_EINVAL, "ntb_ctc_greedy: bad argument%s", "");
    NTB_TRY(check_device());
    if (n_windows == 0) return NTB_OK;
    k_ctc_greedy<<<cdiv(n_windows, 8), 256, 0, (cudaStream_t)stream>>>(logits_dev, len_dev, n_windows, T, C,
                                                                      merge_repeated ? 1 : 0, labels_dev, label_len_dev,
                                                                      neg_sum_logits_dev);
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

static int ctc_beam_blocks(int64_t n) {
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t want = (n + CTC_WARPS - 1) / CTC_WARPS, cap = (int64_t)sms * 3;
    return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

extern "C" int64_t ntb_ctc_beam_workspace_bytes(int64_t n_windows, int32_t T, int32_t beam_width) {
    if (n_windows < 0 || T <= 0 || beam_width <= 0) return 0;
    if (check_device() != NTB_OK) return 0;
    return (int64_t)ctc_beam_blocks(n_windows) * CTC_WARPS * CTC_NODE_I32 * ((int64_t)T * beam_width + 1) * (int64_t)sizeof(int32_t);
}

extern "C" int ntb_ctc_beam_search(const float *logits_dev, const int32_t *len_dev, int64_t n_windows, int32_t T, int32_t C,
                                   int32_t beam_width, int32_t top_paths, int32_t merge_repeated, int32_t log_softmax,
                                   void *workspace_dev, int64_t workspace_bytes, int32_t *paths_dev,
                                   int32_t *path_len_dev, float *log_prob_dev, void *stream) {
    if (!logits_dev || !len_dev || !paths_dev || !path_len_dev || !log_prob_dev || n_windows < 0 || T <= 0 || T > 512 ||
        C < 2 || C > CTC_MAXC || beam_width < 1 || beam_width > CTC_MAXW || top_paths < 1 || top_paths > beam_width)
        return fail(NTB_EINVAL, "ntb_ctc_beam_search: bad argument (2 <= C <= 8, beam_width <= 128, T <= 512, top_paths <= beam_width)%s", "");
    NTB_TRY(check_device());
    if (n_windows == 0) return NTB_OK;
    const int64_t need = ntb_ctc_beam_workspace_bytes(n_windows, T, beam_width);
    if (!workspace_dev || workspace_bytes < need)
        return fail(NTB_ENOMEM, "ntb_ctc_beam_search: workspace too small (need %s%lld)", "", (long long)need);
    const size_t smem = sizeof(CtcBeamSmem) * CTC_WARPS;
    static bool attr_set = false;
    if (!attr_set) {
        NTB_CUDA(cudaFuncSetAttribute(k_ctc_beam_search, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    k_ctc_beam_search<<<ctc_beam_blocks(n_windows), CTC_WARPS * 32, smem, (cudaStream_t)stream>>>(
        logits_dev, len_dev, n_windows, T, C, beam_width, top_paths, merge_repeated ? 1 : 0, log_softmax ? 1 : 0,
        (int32_t *)workspace_dev, paths_dev, path_len_dev, log_prob_dev);
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

// ---------------------------------------------------------------------- allocation-free variant (no reference counterpart)
// ntb_ctc_beam_search with its scratch (the prefix trees of the sequences in flight) in a file-scope device array instead of
// a caller workspace: for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
// Every GPU that loads the library holds the array until exit.  All calls share it: calls that overlap on the GPU race.
// Nothing has to start at zero -- the kernel writes every tree node (root included) before it reads it -- so the previous
// call's contents are simply overwritten.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NTB_NOALLOC_SCRATCH_BYTES];

extern "C" int ntb_ctc_beam_search_noalloc(const float *logits_dev, const int32_t *len_dev, int64_t n_windows, int32_t T,
                                           int32_t C, int32_t beam_width, int32_t top_paths, int32_t merge_repeated,
                                           int32_t log_softmax, int32_t *paths_dev, int32_t *path_len_dev,
                                           float *log_prob_dev, void *stream) {
    if (!logits_dev || !len_dev || !paths_dev || !path_len_dev || !log_prob_dev || n_windows < 0 || T <= 0 || T > 512 ||
        C < 2 || C > CTC_MAXC || beam_width < 1 || beam_width > CTC_MAXW || top_paths < 1 || top_paths > beam_width)
        return fail(NTB_EINVAL, "ntb_ctc_beam_search_noalloc: bad argument (2 <= C <= 8, beam_width <= 128, T <= 512, top_paths <= beam_width)%s", "");
    // one block (CTC_WARPS sequences in flight) must fit the static scratch
    const int64_t per_block = (int64_t)CTC_WARPS * CTC_NODE_I32 * ((int64_t)T * beam_width + 1) * (int64_t)sizeof(int32_t);
    if (per_block > (int64_t)NTB_NOALLOC_SCRATCH_BYTES)
        return fail(NTB_EINVAL, "ntb_ctc_beam_search_noalloc: seq_len * beam_width = %s%lld exceeds the static scratch (limit %lld); "
                    "use ntb_ctc_beam_search with a caller workspace", "", (long long)T * beam_width,
                    (long long)NTB_NOALLOC_MAX_T_TIMES_BEAM);
    NTB_TRY(check_device());
    if (n_windows == 0) return NTB_OK;
    // device address of the array, once per device with that device current (the symbol's name is only a host-side handle)
    static int32_t *scratch_of[64] = {};
    int dev = 0;
    NTB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(NTB_EINVAL, "ntb_ctc_beam_search_noalloc: device ordinal %s%lld out of range", "", (long long)dev);
    const size_t smem = sizeof(CtcBeamSmem) * CTC_WARPS;
    if (!scratch_of[dev]) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return fail(NTB_ECUDA, "ntb_ctc_beam_search_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        NTB_CUDA(cudaFuncSetAttribute(k_ctc_beam_search, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        scratch_of[dev] = (int32_t *)ptr;
    }
    int64_t blocks = (int64_t)NTB_NOALLOC_SCRATCH_BYTES / per_block;       // the grid strides over the sequences
    const int64_t want = ctc_beam_blocks(n_windows);
    if (blocks > want) blocks = want;
    k_ctc_beam_search<<<(unsigned)blocks, CTC_WARPS * 32, smem, (cudaStream_t)stream>>>(
        logits_dev, len_dev, n_windows, T, C, beam_width, top_paths, merge_repeated ? 1 : 0, log_softmax ? 1 : 0,
        scratch_of[dev], paths_dev, path_len_dev, log_prob_dev);
    NTB_LAUNCH_CHECK();
    return NTB_OK;
}

extern "C" int64_t ntb_gemm_workspace_bytes(int64_t M, int32_t N, int32_t K, int32_t impl) {
    if (impl == 0) return 256;
    Arena ar(nullptr, 0);
    GemmWeights w;
    gemm_tc_plan_weights(w, 1, K, N, ar);
    GemmScratch gs;
    gemm_tc_plan_scratch(gs, M, 0, K, ar);
    return ar.off + 256;
}

extern "C" int ntb_gemm(const float *a_dev, const float *b_dev, const float *bias_dev, int64_t M, int32_t N,
                        int32_t K, int32_t impl, void *workspace_dev, int64_t workspace_bytes, float *out_dev,
                        void *stream) {
    if (!a_dev || !b_dev || !out_dev || M < 0 || N <= 0 || K <= 0) return fail(NTB_EINVAL, "ntb_gemm: bad argument%s", "");
    NTB_TRY(check_device());
    if (M == 0) return NTB_OK;
    Ctx cx{(cudaStream_t)stream, nullptr};
    if (impl == 0) return dense(cx, a_dev, K, M, K, N, b_dev, nullptr, bias_dev, out_dev, N, nullptr);
    Arena ar(workspace_dev, workspace_bytes);
    GemmWeights w;
    gemm_tc_plan_weights(w, 1, K, N, ar);
    GemmScratch gs;
    gemm_tc_plan_scratch(gs, M, 0, K, ar);
    if (!w.valid || !gs.valid) return fail(NTB_EINVAL, "ntb_gemm: shape not supported by the tcgen05 path (K %% 16, N %% 16)%s", "");
    if (!workspace_dev || !ar.ok()) return fail(NTB_ENOMEM, "ntb_gemm: workspace too small (need %s%lld)", "", (long long)ar.off);
    w.single_pass = impl == 2;
    NTB_TRY(gemm_tc_pack_weights(cx.st, b_dev, w));
    return gemm_tc_run(cx.st, a_dev, K, M, w, bias_dev, out_dev, N, gs, &cx.launches);
}
