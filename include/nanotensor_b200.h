/*
 * nanotensor_b200.h -- C ABI of the B200-native NanoTensor BLSTM basecall forward.
 *
 * The reference (BD2KGenomics/nanopore-RNN) has NO FFI on this path: its forward is a chain of
 * TensorFlow-1.4 op calls built in nanotensor/network.py and driven by
 * nanotensor/run_nanotensor.py.  Each entry point below replaces one of those call sites (cited
 * per function); INTEGRATION.md shows the ctypes binding a maintainer would add.
 *
 * Conventions
 *  - plain pointers and sizes only; no torch / TF types.  `stream` is a cudaStream_t passed as
 *    void* (NULL = legacy default stream).  Pointers suffixed _dev are device pointers on the
 *    current CUDA device, _host are host pointers.
 *  - the device-pointer entry points never allocate device memory: callers size buffers with the
 *    *_bytes() queries and pass them in (the Python host side uses torch's caching allocator).
 *  - every function returns 0 on success, a negative NTB_E* code otherwise;
 *    ntb_last_error() returns a thread-local message.  There is no CPU fallback: without an
 *    sm_100 device every compute entry point returns NTB_ENODEVICE.
 *  - all arithmetic is float32 in / float32 accumulate; the dense contractions on the tensor
 *    cores use fp16x3 split operands (hi*hi + lo*hi + hi*lo), see DESIGN.md.
 *  - activations are row-major [window][t][channel]; "row" = window * seq_len + t.
 */
#ifndef NANOTENSOR_B200_H
#define NANOTENSOR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define NTB_API __attribute__((visibility("default")))
#else
#define NTB_API
#endif

#define NTB_OK 0
#define NTB_EINVAL (-1)    /* bad argument / unsupported shape */
#define NTB_ENODEVICE (-2) /* no sm_100 CUDA device */
#define NTB_ECUDA (-3)     /* CUDA runtime error (message has the cudaError string) */
#define NTB_ENOMEM (-4)    /* caller-provided buffer too small */

#define NTB_LAYER_RESIDUAL 1   /* network.py:153-163, 385-400 */
#define NTB_LAYER_BLSTM 2      /* network.py:131-137, 266-288 */
#define NTB_LAYER_CHIRON_FNN 3 /* network.py:164-175 */

#define NTB_GRAPH_CROSS_ENTROPY 0 /* network.py:452-473 : argmax labels      */
#define NTB_GRAPH_CTC 1           /* network.py:476-533 : ctc greedy labels  */

#define NTB_MAX_LAYER_TENSORS 12

/* One entry of the config's "network" list plus HOST pointers to its variables in the TF
 * checkpoint layouts (tf.train.Saver.restore, run_nanotensor.py:337-338):
 *  RESIDUAL   w[0] conv1  [1,1,Cin,C]   w[1] conv1_bn_scale [C] (NULL unless batchnorm) w[2] conv1_bn_offset
 *             w[3] conv2a [1,1,Cin,C]   w[4] conv2a_bn_scale    w[5] conv2a_bn_offset
 *             w[6] conv2b [1,3,C,C]     w[7] conv2b_bn_scale    w[8] conv2b_bn_offset
 *             w[9] conv2c [1,1,C,C]     w[10] conv2c_bn_scale   w[11] conv2c_bn_offset
 *  BLSTM      w[0] fw kernel [D+H,4H] (gate column blocks i|j|f|o)  w[1] fw bias [4H]
 *             w[2] bw kernel [D+H,4H]                               w[3] bw bias [4H]
 *  CHIRON_FNN w[0] weights [2,H]  w[1] bias [H]
 */
typedef struct ntb_layer {
    int32_t type;        /* NTB_LAYER_*                                  */
    int32_t in_dim;      /* channels entering the layer                  */
    int32_t size;        /* "size" (H) or "out_channel" (C)              */
    int32_t batchnorm;   /* residual: "batchnorm" (BN on branch1)        */
    float forget_bias;   /* blstm: the layer's "bias" key (network.py:135) */
    const float *w[NTB_MAX_LAYER_TENSORS];
} ntb_layer;

typedef struct ntb_model_desc {
    int32_t graph;       /* NTB_GRAPH_*                                          */
    int32_t seq_len;     /* T                                                    */
    int32_t len_x;       /* input features per timestep (1 for CtcLoss graphs)   */
    int32_t n_classes;   /* C (CtcLoss: len_y + 1, blank = C-1, network.py:495)  */
    int32_t n_layers;
    const ntb_layer *layers;
    const float *out_w;  /* host [F, C]  fulconn_layer weights (network.py:291-309) */
    const float *out_b;  /* host [C]                                              */
} ntb_model_desc;

typedef struct ntb_model ntb_model; /* opaque */

NTB_API const char *ntb_last_error(void);
/* "major.minor.patch"; also reports the gencode the kernels were built for. */
NTB_API const char *ntb_version(void);
/* 1 if the current device is compute capability 10.x, else 0 (never errors). */
NTB_API int ntb_device_ok(void);

/* ---- model life cycle (replaces Saver.restore + graph construction, network.py:32-103) ---- */
/* Device bytes needed for the repacked, GPU-resident weight tiles of `desc`. */
NTB_API int64_t ntb_model_weight_bytes(const ntb_model_desc *desc);
/* Uploads the TF-layout host tensors and repacks them on the device into `weights_dev`
 * (gate-interleaved fp32 recurrence tiles, K-major fp16 hi/lo GEMM tiles).  The buffer must
 * stay alive until ntb_model_destroy. */
NTB_API int ntb_model_create(const ntb_model_desc *desc, void *weights_dev, int64_t weights_bytes,
                     void *stream, ntb_model **out);
NTB_API void ntb_model_destroy(ntb_model *m);
/* Scratch bytes for one forward over `n_windows` windows split in `n_groups` BN groups. */
NTB_API int64_t ntb_model_workspace_bytes(const ntb_model *m, int64_t n_windows, int64_t n_groups);

/* ---- the forward (replaces sess.run([prediction]) of run_nanotensor.py:345) -------------
 * x_dev        [n_windows, T, len_x] float32
 * len_dev      [n_windows] int32  (dynamic_rnn sequence_length)
 * group_off_dev[n_groups+1] int32 : window offsets of the BN groups (= the reference's
 *              batches: moments are taken over each batch, network.py:339-349); ascending,
 *              group_off[0]=0, group_off[n_groups]=n_windows.  May be NULL when the network
 *              has no residual layer.
 * logits_dev   optional [n_windows*T, C] float32 ("output_layer", network.py:114-118/507-513)
 * labels_dev   CrossEntropy: [n_windows*T] int32 argmax (first max);
 *              CTC: [n_windows, T] int32 greedy-decoded, padded with -1
 * label_len_dev CTC: [n_windows] int32 decoded length (NULL allowed for CrossEntropy)
 */
NTB_API int ntb_model_forward(ntb_model *m, const float *x_dev, const int32_t *len_dev,
                      const int32_t *group_off_dev, int64_t n_windows, int64_t n_groups,
                      void *workspace_dev, int64_t workspace_bytes, float *logits_dev,
                      int32_t *labels_dev, int32_t *label_len_dev, void *stream);

/* Same call with HOST buffers (pinned or pageable): copies inputs host->device into the
 * workspace, runs the forward, copies labels (and logits if non-NULL) back and synchronises
 * the stream.  This is the end-to-end entry the reference-facing plugin uses. */
NTB_API int ntb_model_forward_host(ntb_model *m, const float *x_host, const int32_t *len_host,
                           const int32_t *group_off_host, int64_t n_windows, int64_t n_groups,
                           void *workspace_dev, int64_t workspace_bytes, float *logits_host,
                           int32_t *labels_host, int32_t *label_len_host, void *stream);
/* Same, WITHOUT the final synchronisation: copies and kernels are only enqueued on `stream`.  The host buffers must be
 * page-locked, and they, the workspace and the inputs must stay untouched until the caller has synchronised the
 * stream.  Alternating two streams and two workspaces over the chunks of a job overlaps the host<->device copies of
 * one chunk with the kernels of the other (what engine.Model.forward_host does; the reference's counterpart is the
 * tf.data prefetch in front of sess.run, dataset.py:357-380). */
NTB_API int ntb_model_forward_host_async(ntb_model *m, const float *x_host, const int32_t *len_host,
                                 const int32_t *group_off_host, int64_t n_windows, int64_t n_groups,
                                 void *workspace_dev, int64_t workspace_bytes, float *logits_host,
                                 int32_t *labels_host, int32_t *label_len_host, void *stream);
/* Workspace bytes for ntb_model_forward_host (forward scratch + staged inputs/outputs). */
NTB_API int64_t ntb_model_workspace_bytes_host(const ntb_model *m, int64_t n_windows, int64_t n_groups,
                                       int want_logits);

/* Number of kernel launches the last ntb_model_forward on this model issued. */
NTB_API int64_t ntb_model_last_launches(const ntb_model *m);
/* Per-stage device time (ms), summed over every forward issued since profiling was enabled with
 * ntb_model_set_profile(m, 1) (or since the previous ntb_model_stage_ms call): stages are
 * 0 conv/BN front-end, 1 LSTM input projections, 2 LSTM recurrences, 3 readout and -- only when
 * max_stages >= 5 -- 4 = the projection GEMM kernel alone (stage 1 then holds just the operand
 * split passes; with max_stages == 4 stage 1 is split + GEMM).  Stage marks are
 * CUDA events recorded on the forward's stream WITHOUT synchronising; this call waits for the
 * last one.  Returns the number of stages written. */
NTB_API int ntb_model_set_profile(ntb_model *m, int enable);
NTB_API int ntb_model_stage_ms(ntb_model *m, float *ms_out, int max_stages);

/* ---- op-level entry points (kernel-boundary parity tests; same kernels the model uses) ---- */

/* FullSignalSequence.load_data_inference + read_signal(normalize=True)
 * (dataset.py:435-445, chiron/chiron/chiron_input.py:220-230): z-score every read with the
 * mean / population std of the WHOLE read (float64 accumulation), drop start_index samples,
 * cut windows of seq_len every `step`, zero pad.
 * raw_dev [total_samples] int16; read_off_dev [n_reads+1] int64 sample offsets;
 * win_off_dev [n_reads+1] int64 window offsets (host computes ceil((len-start)/step));
 * x_dev [n_windows, seq_len] float32; len_dev [n_windows] int32. */
NTB_API int ntb_pack_windows(const int16_t *raw_dev, const int64_t *read_off_dev,
                     const int64_t *win_off_dev, int64_t n_reads, int64_t n_windows,
                     int32_t seq_len, int32_t step, int32_t start_index, float *x_dev,
                     int32_t *len_dev, void *stream);

/* events_to_features (nanonet/nanonet/features.py:61-75) with SquiggleFeatureGenerator.__init__
 * (features.py:266-293, sloika_model=False), scale_array (features.py:32-58) and padded_offset_array
 * (features.py:14-29), window [-1, 0, 1]: the feature front end of the event (CrossEntropy) model
 * (nanonetcall.py:204 feeds features.astype(nn.dtype) to the network).  Per read: mean / stdv / length
 * standardised over the read, delta = diff of the standardised mean / its population std; per event the
 * 12 columns [mean, stdv, length, delta] x (left neighbour, self, right neighbour), zeros past the ends.
 * float64 arithmetic in NumPy's own operation and pairwise-summation order: with out_f64 = 1 the result is
 * bit-identical to the reference's; out_f64 = 0 rounds once to float32 (the .astype at the call site).
 * mean/stdv/length [total_events] float64, reads concatenated; read_off_host [n_reads+1] int64 is a HOST
 * pointer (event offsets, 0 .. total_events); out [total_events, 12] float64 or float32.
 * NTB_EINVAL for a read too long for the per-read leaf table in shared memory (> ~500k events; the
 * reference's own callers drop reads above max_len = 9000 events, features.py:78).
 * workspace_dev: ntb_event_features_workspace_bytes() bytes on the device (the read table and the
 * standardised stdv / length columns between the kernel's passes); nothing is allocated by the call.
 * out_dev and workspace_dev must be 16-byte aligned (vector stores).
 * _host: all buffers on the host; allocates, copies in, runs, copies out and synchronises `stream`. */
NTB_API int64_t ntb_event_features_workspace_bytes(int64_t n_reads, int64_t total_events);
NTB_API int ntb_event_features(const double *mean_dev, const double *stdv_dev, const double *length_dev,
                       const int64_t *read_off_host, int64_t n_reads, int64_t total_events,
                       int32_t out_f64, void *workspace_dev, int64_t workspace_bytes, void *out_dev,
                       void *stream);
NTB_API int ntb_event_features_host(const double *mean_host, const double *stdv_host,
                            const double *length_host, const int64_t *read_off_host, int64_t n_reads,
                            int64_t total_events, int32_t out_f64, void *out_host, void *stream);

/* process_output (nanotensor/dataset.py:421-432): SignalLabel.index2base per window
 * (nanotensor/trim_signal.py:116-122), simple_assembly + add_count
 * (chiron/chiron/utils/easy_assembler.py:11-37), np.argmax, index2base -- the consensus call of every read
 * from the label rows the CTC readout produced.  labels [n_windows, T] int32, -1 padded (ntb_model_forward's
 * labels_dev); read_off_host [n_reads+1] int64 window offsets (HOST pointer).  Each window is placed against
 * its predecessor by the offset of the block the reference takes, max(SequenceMatcher.get_matching_blocks(), key=size):
 * the longest common block with difflib's tie-break for windows under 200 called bases, difflib's whole "autojunk"
 * procedure (popular symbols out of the match index, extension, recursion, merge) from 200 on, bases vote per position, the
 * consensus is the first-max vote.  Integer work: identical to the reference, vote matrix included.
 * cons_dev [capacity] uint8 ASCII "ACGTE", reads back to back; cons_off_dev [n_reads+1] int64;
 * counts_dev optional [5, capacity] int32 vote matrix (row stride = capacity); capacity >= the number of
 * label entries that are not -1 (n_windows * T always suffices; status 1 otherwise);
 * status_dev int32: 0 ok.  _host: host buffers, allocates, synchronises, returns the status as an error;
 * counts_host optional [5, cons_off[n_reads]] compact. */
NTB_API int64_t ntb_assemble_workspace_bytes(int64_t n_windows, int64_t n_reads, int32_t T, int64_t capacity);
NTB_API int ntb_assemble_reads(const int32_t *labels_dev, const int64_t *read_off_host, int64_t n_reads,
                       int64_t n_windows, int32_t T, int64_t capacity, void *workspace_dev,
                       int64_t workspace_bytes, uint8_t *cons_dev, int64_t *cons_off_dev, int32_t *counts_dev,
                       int32_t *status_dev, void *stream);
NTB_API int ntb_assemble_reads_host(const int32_t *labels_host, const int64_t *read_off_host, int64_t n_reads,
                            int64_t n_windows, int32_t T, int64_t capacity, uint8_t *cons_host,
                            int64_t *cons_off_host, int32_t *counts_host, void *stream);

/* minknow_event_detect (nanonet/nanonet/eventdetection/filters.py:299-322): compute_sum_sumsq
 * (filters.c:57-72), compute_tstat for the two windows (filters.c:130-195), short_long_peak_detector
 * (filters.c:235-300) and _construct_events (filters.py:146-204): raw current samples -> events, the table
 * ntb_event_features standardises.  raw [total_samples] float64, reads concatenated; read_off_host
 * [n_reads+1] int64 sample offsets (HOST pointer); window_lengths[2] / thresholds[2] HOST arrays in the
 * reference's argument order (the shorter window dominates, as in the reference); events [capacity, 4]
 * float64 = start, length (both / sample_rate), mean, stdv, reads back to back; event_off [n_reads+1].
 * Same float64 operations in the same order as the reference's C (sequential cumulative sums included):
 * start / length / mean bit-identical, stdv within 1 ulp-of-variance (the reference squares the mean with
 * libm pow).  capacity >= number of events (total_samples always suffices); status_dev: 0 ok, 1 capacity. */
NTB_API int64_t ntb_event_detect_workspace_bytes(int64_t n_reads, int64_t total_samples);
NTB_API int ntb_event_detect(const double *raw_dev, const int64_t *read_off_host, int64_t n_reads,
                     int64_t total_samples, double sample_rate, const int32_t *window_lengths,
                     const double *thresholds, double peak_height, void *workspace_dev,
                     int64_t workspace_bytes, int64_t capacity, double *events_dev, int64_t *event_off_dev,
                     int32_t *status_dev, void *stream);
NTB_API int ntb_event_detect_host(const double *raw_host, const int64_t *read_off_host, int64_t n_reads,
                          int64_t total_samples, double sample_rate, const int32_t *window_lengths,
                          const double *thresholds, double peak_height, int64_t capacity,
                          double *events_host, int64_t *event_off_host, void *stream);

/* The text parse of read_signal (chiron/chiron/chiron_input.py:220-226): `.signal` files hold whitespace
 * separated integers; signal += [int(x) for x in line.split()].  text [total_bytes] = the bytes of n_files
 * files back to back, file_off_host [n_files+1] byte offsets (HOST pointer; a token never spans two files);
 * samples [capacity] int16 and read_off [n_files+1] int64 (samples per file, cumulative) are exactly the raw_dev /
 * read_off_dev inputs of ntb_pack_windows, which does the z-score and the windows.  capacity >= number of
 * samples (total_bytes / 2 + n_files always suffices).  status_dev: 0 ok, 1 capacity, 2 a token that is not
 * [+-]?[0-9]{1,6}, 3 a value outside int16.  Synchronises the stream once (chunk table upload). */
NTB_API int64_t ntb_parse_signal_workspace_bytes(int64_t total_bytes, int64_t n_files);
NTB_API int ntb_parse_signal(const uint8_t *text_dev, const int64_t *file_off_host, int64_t n_files,
                     int64_t total_bytes, void *workspace_dev, int64_t workspace_bytes, int64_t capacity,
                     int16_t *samples_dev, int64_t *read_off_dev, int32_t *status_dev, void *stream);
NTB_API int ntb_parse_signal_host(const uint8_t *text_host, const int64_t *file_off_host, int64_t n_files,
                          int64_t total_bytes, int64_t capacity, int16_t *samples_host,
                          int64_t *read_off_host, void *stream);

/* conv_layer + simple_global_bn (+ReLU) (network.py:312-349): y = conv(in) (k = 1 or 3, SAME,
 * no bias), then per-group batch moments and normalisation.  in_dev [n_windows*T, Cin],
 * filt_dev = TF filter [1,k,Cin,Cout] on the device; scale/offset NULL = no BN.
 * out_dev [n_windows*T, Cout]. */
NTB_API int ntb_conv_bn(const float *in_dev, const float *filt_dev, const float *scale_dev,
                const float *offset_dev, const int32_t *group_off_dev, int64_t n_windows,
                int64_t n_groups, int32_t T, int32_t Cin, int32_t Cout, int32_t ksize, int32_t relu,
                void *workspace_dev, int64_t workspace_bytes, float *out_dev, void *stream);
NTB_API int64_t ntb_conv_bn_workspace_bytes(int64_t n_windows, int64_t n_groups, int32_t T, int32_t Cin,
                                    int32_t Cout, int32_t ksize);

/* blstm (network.py:266-288): both directions of tf.nn.bidirectional_dynamic_rnn over
 * rnn.LSTMCell, fw|bw concatenated.  kernels/biases are DEVICE pointers in the TF layouts.
 * in_dev [n_windows*T, D] -> out_dev [n_windows*T, 2H]. */
NTB_API int ntb_blstm(const float *in_dev, const int32_t *len_dev, const float *kernel_fw_dev,
              const float *bias_fw_dev, const float *kernel_bw_dev, const float *bias_bw_dev,
              float forget_bias, int64_t n_windows, int32_t T, int32_t D, int32_t H,
              void *workspace_dev, int64_t workspace_bytes, float *out_dev, void *stream);
NTB_API int64_t ntb_blstm_workspace_bytes(int64_t n_windows, int32_t T, int32_t D, int32_t H);

/* tf.nn.ctc_greedy_decoder(merge_repeated=True) + sparse_tensor_to_dense(-1)
 * (network.py:528-533) / tf.argmax (network.py:208-211) on given logits [n_windows*T, C]. */
NTB_API int ntb_decode(const float *logits_dev, const int32_t *len_dev, int64_t n_windows, int32_t T,
               int32_t C, int32_t graph, int32_t *labels_dev, int32_t *label_len_dev,
               void *stream);

/* tf.nn.ctc_greedy_decoder(logits, sequence_length, merge_repeated) with BOTH outputs (nanotensor/network.py:528-533 keeps
 * decoded[0] and drops neg_sum_logits; chiron/chiron/chiron_rcnn_train.py:68).  logits [n_windows, T, C] float32 (the
 * "output_layer" of the CTC graphs; blank = class C-1); labels [n_windows, T] int32 padded with -1; label_len [n_windows]
 * (may be NULL); neg_sum_logits [n_windows] float32 = -(sum over t < len of max_c logits), accumulated in frame order in
 * float32 as TF's CTCGreedyDecoder does (may be NULL).  Integer outputs identical to ntb_decode(graph = CTC) when
 * merge_repeated = 1. */
NTB_API int ntb_ctc_greedy(const float *logits_dev, const int32_t *len_dev, int64_t n_windows, int32_t T, int32_t C,
                   int32_t merge_repeated, int32_t *labels_dev, int32_t *label_len_dev, float *neg_sum_logits_dev,
                   void *stream);

/* tf.nn.ctc_beam_search_decoder(logits, sequence_length, beam_width, top_paths, merge_repeated)
 * (chiron/chiron/chiron_rcnn_train.py:66-67: merge_repeated=False, top_paths=k; chiron_cnn_train.py:32, train_test.py:43,83;
 * beam_width = TF's default 100): TensorFlow 1.4's CTCBeamSearchDecoder::Step / TopPaths restated (prefix tree of beam
 * entries that keep their identity when they leave and re-enter the beam; per frame logits - max are used AS
 * log-probabilities, which is what TF 1.4 does; log_softmax = 1 normalises them first, the later TF behaviour).
 * 2 <= C <= 8, beam_width <= 128, T <= 512, top_paths <= beam_width.  paths [n_windows, top_paths, T] int32 padded with
 * -1, path_len [n_windows, top_paths], log_prob [n_windows, top_paths] float32 (-inf for a missing path).
 * workspace: ntb_ctc_beam_workspace_bytes() bytes on the device (the prefix trees of the sequences in flight). */
NTB_API int64_t ntb_ctc_beam_workspace_bytes(int64_t n_windows, int32_t T, int32_t beam_width);
NTB_API int ntb_ctc_beam_search(const float *logits_dev, const int32_t *len_dev, int64_t n_windows, int32_t T, int32_t C,
                        int32_t beam_width, int32_t top_paths, int32_t merge_repeated, int32_t log_softmax,
                        void *workspace_dev, int64_t workspace_bytes, int32_t *paths_dev, int32_t *path_len_dev,
                        float *log_prob_dev, void *stream);

/* Allocation-free variant of ntb_ctc_beam_search.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same kernel, same arguments and results as ntb_ctc_beam_search, but the scratch (the
 * prefix trees of the sequences in flight) is a file-scope `static __device__` array of NTB_NOALLOC_SCRATCH_BYTES inside
 * the library instead of a caller workspace; the call allocates nothing (no cudaMalloc / cudaFree, no library routine that
 * allocates) and every GPU that loads the library holds the array until the process exits.
 * Limit: one block of 4 sequences must fit the array, i.e. seq_len * beam_width <= NTB_NOALLOC_MAX_T_TIMES_BEAM
 * (T = 300 x beam 100 fits, T = 512 x beam 128 does not); larger shapes return NTB_EINVAL with a message that names the
 * limit -- use ntb_ctc_beam_search with a caller workspace for those.  n_windows is not limited: the grid is sized to the
 * array (fewer sequences in flight than with a caller workspace, so large jobs run slower).
 * All calls on a device share the one array: calls that overlap on the GPU (different streams, or different host threads)
 * race on it -- serialise them.  The array keeps the previous call's contents; nothing in it has to start at zero (the
 * kernel writes every tree node before reading it), so no clearing happens between calls. */
#define NTB_NOALLOC_SCRATCH_BYTES (8 * 1024 * 1024)
#define NTB_NOALLOC_MAX_T_TIMES_BEAM (NTB_NOALLOC_SCRATCH_BYTES / (4 * 12 * 4) - 1)
NTB_API int ntb_ctc_beam_search_noalloc(const float *logits_dev, const int32_t *len_dev, int64_t n_windows, int32_t T,
                                int32_t C, int32_t beam_width, int32_t top_paths, int32_t merge_repeated,
                                int32_t log_softmax, int32_t *paths_dev, int32_t *path_len_dev, float *log_prob_dev,
                                void *stream);

/* Dense GEMM used by the conv / projection stages, exposed for roofline measurement and
 * parity tests: out[M,N] = A[M,K] @ B[K,N] (+ bias[N]), float32 in/out.
 * impl 0 = fp32 SIMT, 1 = tcgen05 fp16x3 (TMA + TMEM), 2 = tcgen05 single-pass fp16. */
NTB_API int ntb_gemm(const float *a_dev, const float *b_dev, const float *bias_dev, int64_t M, int32_t N,
             int32_t K, int32_t impl, void *workspace_dev, int64_t workspace_bytes,
             float *out_dev, void *stream);
NTB_API int64_t ntb_gemm_workspace_bytes(int64_t M, int32_t N, int32_t K, int32_t impl);

#ifdef __cplusplus
}
#endif
#endif /* NANOTENSOR_B200_H */
