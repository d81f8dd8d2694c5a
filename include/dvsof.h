/* dvsof.h -- C ABI of libdvsof.so: the B200 (sm_100a) CUDA hot path of
 * e-sha/dvs_of_training_framework (event-image encoding, event transforms, multiscale
 * warp loss forward/backward).
 *
 * Conventions
 *  - every pointer marked [dev] is a device pointer owned by the caller; the library
 *    allocates nothing and keeps no global state (re-entrant, stream-ordered) -- the one exception
 *    is dvsof_encode_event_image_static below, which says so;
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *  - return value: 0 = DVSOF_OK, negative = host-side argument / launch error, text
 *    available from dvsof_last_error() (thread-local);
 *  - data-dependent input validation happens ON DEVICE: kernels OR flag bits into the
 *    caller's `status` word [dev, uint32, zero-initialised by the caller]; the host binding
 *    reads it when it chooses (no forced sync) and raises the reference's error type.
 *
 * Each entry point cites the reference interface (relative to the reference repo) it
 * replaces.  No torch types appear here; PyTorch/ctypes bindings live in
 * dvs_of_training_framework_b200/_lib.py, reference-side stubs in INTEGRATION.md.
 */
#ifndef DVSOF_H_
#define DVSOF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DVSOF_OK 0
#define DVSOF_ERR_INVALID_ARGUMENT (-1)
#define DVSOF_ERR_CUDA (-2)
#define DVSOF_ERR_WORKSPACE (-3)

/* status bits written by kernels */
#define DVSOF_ST_BAD_POLARITY 0x1u     /* |p| != 1            (EV_FlowNet/net.py:124)      */
#define DVSOF_ST_NOT_GROUPED 0x2u      /* sample_index not ascending (net.py:131-134)      */
#define DVSOF_ST_BAD_SAMPLE 0x4u       /* sample_index outside [0, B)                      */
#define DVSOF_ST_BAD_PIXEL 0x8u        /* x/y outside the image (reference: index error)   */
#define DVSOF_ST_NO_MATCH 0x10u        /* image<->flow timestamp match != 1 (utils/loss.py:193) */
#define DVSOF_ST_BAD_LUT 0x20u         /* map_events LUT entry outside [0, H*W)            */
#define DVSOF_ST_BAD_COUNTS 0x40u      /* batch codec: run lengths do not cover the events / element outside the batch */

const char* dvsof_last_error(void);
int dvsof_version(void);

/* ---------------------------------------------------------------------------------------
 * Event-image encoder.  Replaces EV_FlowNet/net.py:88-150 `compute_event_image` (called from
 * Model.quantize, net.py:197-205) including the torch_scatter.scatter_sum / scatter_max calls
 * at net.py:147-149.
 *   out[b, c, y, x]   = #events with polarity c (c = (1-p)/2),
 *   out[b, c+2, y, x] = max(0, max (t - start[b]) / dt[b]),  start[b] = timestamps[2b],
 *   dt[b] = max(t[last event of sample b] - start[b], 1e-9).
 * Events must be grouped by ascending sample_index.  `out` (B,4,H,W) fp32 is fully
 * overwritten.  workspace: dvsof_encode_workspace_bytes(N, B, H, W) bytes; pass its size as
 * `workspace_bytes`.  Two device paths with identical results: the direct one (two L2 reductions
 * per event) and, for large streams when the workspace has room for it, the two-pass one that
 * bins the events by row band and reduces them in shared memory (csrc/encode_binned.cuh).
 * A workspace of dvsof_encode_workspace_bytes(0, B, H, W) bytes selects the direct path.
 * ------------------------------------------------------------------------------------- */
size_t dvsof_encode_workspace_bytes(int64_t n_events, int64_t n_samples, int32_t H, int32_t W);
int dvsof_encode_event_image(const int64_t* x, const int64_t* y, const float* t,
                             const int64_t* polarity, const int64_t* sample_index, /* [dev] N */
                             int64_t n_events,
                             const float* timestamps, /* [dev] 2*B */
                             int64_t n_samples, int32_t H, int32_t W,
                             float* out,        /* [dev] B*4*H*W */
                             void* workspace,   /* [dev] */
                             size_t workspace_bytes,
                             uint32_t* status,  /* [dev] */
                             void* stream);

/* The two stream-ordered phases of dvsof_encode_event_image, exported separately so a caller
 * (bench.py) can put CUDA events around the scatter kernel alone:
 *   prepare: zero `out`, per-sample (start, dt) into `workspace`;   scatter: the event stream. */
int dvsof_encode_prepare(const float* t, const int64_t* sample_index, int64_t n_events,
                         const float* timestamps, int64_t n_samples, int32_t H, int32_t W,
                         float* out, void* workspace, void* stream);
int dvsof_encode_scatter(const int64_t* x, const int64_t* y, const float* t,
                         const int64_t* polarity, const int64_t* sample_index, int64_t n_events,
                         int64_t n_samples, int32_t H, int32_t W, float* out,
                         const void* workspace, uint32_t* status, void* stream);
/* Which path dvsof_encode_event_image takes for these sizes and this workspace: 0 = direct
 * scatter (L2 atomics), 1 = binned (counting sort into records + per-slab reduction). */
int dvsof_encode_path(int64_t n_events, int64_t n_samples, int32_t H, int32_t W, size_t workspace_bytes);
/* One pass of the binned path alone (kernel-level timing): which = 1 the counting sort of the
 * event chunks into 6-byte records, 2 the per-slab reduction of the records into the image.
 * `workspace` must hold the per-sample bounds (dvsof_encode_prepare) and, for pass 2, the records
 * of an earlier pass 1; DVSOF_ERR_WORKSPACE when the sizes do not admit the binned path. */
int dvsof_encode_binned_pass(const int64_t* x, const int64_t* y, const float* t,
                             const int64_t* polarity, const int64_t* sample_index, int64_t n_events,
                             int64_t n_samples, int32_t H, int32_t W, float* out, void* workspace,
                             size_t workspace_bytes, uint32_t* status, int32_t which, void* stream);

/* Compact storage-form entry point: consumes the on-disk/encoded batch layout directly
 * (utils/dataset.py:286-289 encode_batch: x,y int16, polarity bool, timestamp fp32) plus
 * per-sample event counts instead of the int64 training form produced by decode_batch
 * (utils/dataset.py:335-373).  Same output as dvsof_encode_event_image on the decoded batch.
 * sample_begin: [dev] B+1 int64 exclusive prefix of events per sample. */
int dvsof_encode_event_image_compact(const int16_t* x, const int16_t* y, const float* t,
                                     const uint8_t* polarity01, /* 1 = positive */
                                     const int64_t* sample_begin, int64_t n_events,
                                     const float* timestamps, int64_t n_samples,
                                     int32_t H, int32_t W, float* out, void* workspace,
                                     size_t workspace_bytes, /* dvsof_encode_workspace_bytes(n_events, ...) admits
                                                              * the binned path for dense streams; (0, ...) = direct only */
                                     uint32_t* status, void* stream);

/* EXTENSION, NO COUNTERPART IN THE REFERENCE (whose network input is the 4-channel event image
 * above): the temporal-bilinear voxel grid that BASELINE.json's north_star and configs ("5 / 9 / 15
 * bins") name, by its published definition (Zhu et al., CVPR 2019, eq. 1-3):
 *   t* = (n_bins - 1) * (t - start[b]) / dt[b]        start, dt as for the event image (net.py:136-141)
 *   out[b, k, y, x] = sum of polarity * max(0, 1 - |t* - k|) over the events of pixel (x, y) of sample b.
 * out (B, n_bins, H, W) fp32 is fully overwritten; same validation bits, same two device paths
 * (direct: two fp32 reductions per event; binned: pass 1 of the event-image path, then a
 * [n_bins][rows x W] shared-memory slab per (sample, row band)).  fp32 sums in no fixed order:
 * within ~1e-7 of the absolute sum of a voxel's weights of the exact result.  1 <= n_bins <= 64.
 * workspace: dvsof_voxel_workspace_bytes(N, B, H, W, n_bins); (0, ...) selects the direct path.
 * dvsof_voxel_path: 0 = direct, 1 = binned for these sizes and this workspace.  Parity unpinned. */
size_t dvsof_voxel_workspace_bytes(int64_t n_events, int64_t n_samples, int32_t H, int32_t W, int32_t n_bins);
int dvsof_voxel_path(int64_t n_events, int64_t n_samples, int32_t H, int32_t W, int32_t n_bins,
                     size_t workspace_bytes);
int dvsof_encode_voxel_grid(const int64_t* x, const int64_t* y, const float* t,
                            const int64_t* polarity, const int64_t* sample_index, /* [dev] N */
                            int64_t n_events, const float* timestamps /* [dev] 2*B */, int64_t n_samples,
                            int32_t n_bins, int32_t H, int32_t W, float* out /* [dev] B*n_bins*H*W */,
                            void* workspace, size_t workspace_bytes, uint32_t* status, void* stream);

/* dvsof_encode_event_image without a caller workspace and without any allocation.  NO COUNTERPART
 * IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and the same (bit-identical) result as dvsof_encode_event_image on its direct
 * path; the per-sample scratch (16 B per sample) lives in a static __device__ array of
 * DVSOF_STATIC_MAX_SAMPLES * 16 B = 64 KiB that every GPU loading the library holds until exit.
 * n_samples > DVSOF_STATIC_MAX_SAMPLES: DVSOF_ERR_WORKSPACE (use the workspace form).  The call
 * makes no cudaMalloc / cudaFree / allocating library call.  All calls on one device share the
 * array (it keeps the last call's contents; each call rewrites every word it reads): calls that
 * overlap on the GPU, e.g. on two streams, race -- keep them stream-ordered. */
#define DVSOF_STATIC_MAX_SAMPLES 4096
int64_t dvsof_encode_static_max_samples(void);
int dvsof_encode_event_image_static(const int64_t* x, const int64_t* y, const float* t,
                                    const int64_t* polarity, const int64_t* sample_index, /* [dev] N */
                                    int64_t n_events,
                                    const float* timestamps, /* [dev] 2*B */
                                    int64_t n_samples, int32_t H, int32_t W,
                                    float* out,        /* [dev] B*4*H*W */
                                    uint32_t* status,  /* [dev] */
                                    void* stream);

/* ---------------------------------------------------------------------------------------
 * Event transforms.
 * ------------------------------------------------------------------------------------- */
/* utils/data.py:120-136 get_count_image: out[y, x] += 1 per event; out (H,W) uint64 is zeroed
 * inside the call. */
int dvsof_count_image(const int64_t* x, const int64_t* y, int64_t n_events, int32_t H, int32_t W,
                      unsigned long long* out, uint32_t* status, void* stream);

/* utils/data.py:24-42 EventCrop.__call__: keep rows with x in [x0,x0+w), y in [y0,y0+h), shift
 * by (x0,y0), order preserved.  events / out are row-major (N, E) of `elem_size`-byte floats
 * (4 = fp32, 8 = fp64); out must have room for N rows; *n_out [dev int64] receives N'. */
size_t dvsof_compact_workspace_bytes(int64_t n_rows);
int dvsof_event_crop(const void* events, int64_t n_events, int32_t n_cols, int32_t elem_size,
                     double y0, double x0, double h, double w, void* out, int64_t* n_out,
                     void* workspace, void* stream);

/* utils/modules_to_build/transformations/src/transformation.cpp:35-85 `map_events`
 * (python: utils.transformation.map, src/module.cpp:18-40), split in three stream-ordered
 * steps because the output size is data dependent:
 *   build: LUT (src[k] -> dst[k], k in LUT order) -> per-pixel CSR lists in `lut_ws`;
 *   count: per-event fan-out + scan -> *n_out [dev int64];
 *   write: rows in input order x LUT order; columns >= 2 copied verbatim. */
size_t dvsof_map_lut_workspace_bytes(int32_t H, int32_t W, int64_t lut_size);
int dvsof_map_events_build(const uint64_t* src, const uint64_t* dst, int64_t lut_size, int32_t H,
                           int32_t W, void* lut_ws, uint32_t* status, void* stream);
int dvsof_map_events_count(const float* events, int64_t n_events, int32_t n_cols, int32_t H,
                           int32_t W, const void* lut_ws, void* ev_ws /* compact_workspace(N) */,
                           int64_t* n_out, void* stream);
int dvsof_map_events_write(const float* events, int64_t n_events, int32_t n_cols, int32_t H,
                           int32_t W, const void* lut_ws, const void* ev_ws, float* out,
                           void* stream);

/* The event side of the augmentation block of utils/dataset.py:753-769 in one count + scan +
 * write over the (N, E) fp32 rows: optional horizontal flip x <- W - x - 1 (:757), optional
 * rotation remap through a LUT built by dvsof_map_events_build (utils/data.py:213-216; lut_ws ==
 * NULL: none), optional crop to box = [y0, x0, h, w] with shift (utils/data.py:24-42 EventCrop; box
 * [host, 4 doubles] == NULL: none).  Same rows, same order as the three reference steps run in
 * sequence.  ev_ws: dvsof_compact_workspace_bytes(N); *n_out [dev int64] = number of output rows. */
int dvsof_augment_events_count(const float* events, int64_t n_events, int32_t n_cols, int32_t H,
                               int32_t W, int32_t flip, const void* lut_ws, const double* box,
                               void* ev_ws, int64_t* n_out, void* stream);
int dvsof_augment_events_write(const float* events, int64_t n_events, int32_t n_cols, int32_t H,
                               int32_t W, int32_t flip, const void* lut_ws, const double* box,
                               const void* ev_ws, float* out, void* stream);

/* Batch codec, event part.  Storage form = utils/dataset.py:286-302: x, y
 * int16, polarity bool (1 = positive), events_per_element [int64, E] (events of every element, the
 * elements of all samples concatenated), elements_per_sample [int64, B]; the fp32 timestamps are
 * the same array in both forms.  Training form = the int64 arrays of utils/dataset.py:362-372.
 *   decode (dataset.py:335-373 `decode_batch`): widening, polarity*2-1 and the run-length expansion
 *     into element_index / sample_index in one pass (5 B read, 40 B written per event).  Sets
 *     DVSOF_ST_BAD_COUNTS when the run lengths do not add up to n_events.
 *   encode (dataset.py:283-304 `encode_batch`): narrowing with torch's wrap-around, polarity
 *     ((p+1)/2).to(bool), events_per_element[E] = histogram of element_index + shift[sample_index]
 *     (zeroed inside).  n_elements is the caller's `int(new_e[-1]) + 1` (dataset.py:298).
 * workspace: dvsof_codec_workspace_bytes(n_elements, n_samples). */
size_t dvsof_codec_workspace_bytes(int64_t n_elements, int64_t n_samples);
int dvsof_decode_events(const int16_t* x, const int16_t* y, const uint8_t* polarity01,
                        int64_t n_events, const int64_t* events_per_element, int64_t n_elements,
                        const int64_t* elements_per_sample, int64_t n_samples, int64_t* out_x,
                        int64_t* out_y, int64_t* out_polarity, int64_t* out_element_index,
                        int64_t* out_sample_index, void* workspace, uint32_t* status, void* stream);
int dvsof_encode_events(const int64_t* x, const int64_t* y, const int64_t* polarity,
                        const int64_t* element_index, const int64_t* sample_index, int64_t n_events,
                        const int64_t* elements_per_sample, int64_t n_samples, int64_t n_elements,
                        int16_t* out_x, int16_t* out_y, uint8_t* out_polarity01,
                        int64_t* events_per_element, void* workspace, uint32_t* status, void* stream);

/* utils/data.py:139-152 frame_generator: out[k] = np.searchsorted(t, bounds[k], side='right') for
 * every window boundary (frames.ravel()); t sorted ascending, fp32 (elem_size 4) or fp64 (8),
 * compared in fp64 as numpy does.  The (start, stop) offset pairs slice the event arrays per
 * evaluation window (utils/testing.py:64) without leaving the device. */
int dvsof_window_offsets(const void* t, int32_t elem_size, int64_t n_events, const double* bounds,
                         int64_t n_bounds, int64_t* out, void* stream);

/* utils/eval.py:6-50 flow_error_dense (per evaluation window, utils/testing.py:84-92): endpoint
 * error of flow_pred against flow_gt over the first `rows` rows (all rows, or 190 with is_car) of
 * (H, W, 2) row-major flows [dev; elem size 4 = fp32, 8 = fp64], restricted to pixels with
 * count_image[y, x] > 0 (the uint64 image of dvsof_count_image; ignored when is_dense) whose GT is
 * finite and non-zero.  out [dev] 3 doubles: sum of EE, number of points, number with EE < 3 --
 * AEE = out[0] / out[1], percent_AEE = out[2] / (out[1] + 1e-5).  Deterministic.
 * workspace: dvsof_flow_error_workspace_bytes(). */
size_t dvsof_flow_error_workspace_bytes(void);
int dvsof_flow_error_dense(const void* flow_gt, int32_t gt_elem_size, const void* flow_pred,
                           int32_t pred_elem_size, const unsigned long long* count_image,
                           int32_t rows, int32_t W, int32_t is_dense, double* out, void* workspace,
                           void* stream);

/* ---------------------------------------------------------------------------------------
 * Multiscale warp loss.  Replaces utils/loss.py:121-171 (Loss.__call__), :179-214
 * (Losses.__call__) and autograd's backward of both.
 * ------------------------------------------------------------------------------------- */
/* utils/loss.py:182-206: for prediction p and k in {start, stop}: the unique d with
 * timestamps[d] == flow_ts[p,k] (exact fp32) and sample_idx[d] == flow_sample_idx[p].
 * match_idx: [dev] int32 2*P (first P = start indices, next P = stop indices). */
int dvsof_match_timestamps(const float* flow_ts, const int64_t* flow_sample_idx, int64_t P,
                           const float* timestamps, const int64_t* sample_idx, int64_t D,
                           int32_t* match_idx, uint32_t* status, void* stream);

/* utils/loss.py:20-21 interpolate (bilinear, align_corners=True), one level of the chained
 * pyramid of utils/loss.py:207-210: src (D,Hin,Win) -> dst (D,Hout,Wout). */
int dvsof_resize_bilinear(const float* src, int64_t D, int32_t Hin, int32_t Win, float* dst,
                          int32_t Hout, int32_t Wout, void* stream);

#define DVSOF_MAX_SCALES 8
typedef struct {
    int32_t n_scales;
    int32_t n_pred;                          /* P */
    int32_t n_images;                        /* D: frames in every images[s]; 0 = unknown (the level tiles
                                              * are then staged with plain loads instead of TMA; the
                                              * multiscale calls take it from `frames`) */
    int32_t reserved_;
    int32_t H[DVSOF_MAX_SCALES], W[DVSOF_MAX_SCALES];
    const float* flow[DVSOF_MAX_SCALES];     /* [dev] (P,2,H,W) contiguous */
    const float* images[DVSOF_MAX_SCALES];   /* [dev] (D,H,W) pyramid level */
    float* grad_flow[DVSOF_MAX_SCALES];      /* [dev] (P,2,H,W), backward only */
} dvsof_loss_scales_t;

/* forward: terms [dev] 3*n_scales fp32 laid out [smoothness[s]..., photometric[s]...,
 * outborder[s]...]; n_bad [dev] int32 n_scales*P (out-of-border pixel count n_i per scale and
 * prediction, kept for backward).  workspace: dvsof_loss_workspace_bytes(&scales). */
size_t dvsof_loss_workspace_bytes(const dvsof_loss_scales_t* scales);
int dvsof_warp_loss_forward(const dvsof_loss_scales_t* scales, const int32_t* match_idx,
                            float* terms, int32_t* n_bad, void* workspace, void* stream);
/* backward: grad_terms [dev] 3*n_scales fp32 (same layout as terms); writes grad_flow. */
int dvsof_warp_loss_backward(const dvsof_loss_scales_t* scales, const int32_t* match_idx,
                             const float* grad_terms, const int32_t* n_bad, void* stream);

/* utils/loss.py:179-214 `Losses.__call__` forward in ONE call: timestamp matching (:182-206),
 * the chained pyramid (:207-210; level s is resized from level s-1, level 0 from `images`) written
 * into the caller's level buffers scales->images[s] (D,H_s,W_s), and the per-scale terms.
 * match_idx [dev int32 2P], terms, n_bad as in dvsof_warp_loss_forward (all outputs, kept by the
 * caller for dvsof_warp_loss_backward). */
typedef struct {
    const float* images;            /* [dev] (D,H0,W0) full-resolution frames, fp32 */
    int32_t n_images, H0, W0;       /* D */
    const float* flow_ts;           /* [dev] (P,2) */
    const int64_t* flow_sample_idx; /* [dev] (P) */
    const float* timestamps;        /* [dev] (D) */
    const int64_t* sample_idx;      /* [dev] (D) */
} dvsof_loss_frames_t;
int dvsof_multiscale_loss_forward(const dvsof_loss_scales_t* scales,
                                  const dvsof_loss_frames_t* frames, int32_t* match_idx,
                                  float* terms, int32_t* n_bad, void* workspace, uint32_t* status,
                                  void* stream);

/* The training pass of utils/training.py:12-24 + :186 (`combined_loss` then `loss.backward()`):
 * everything dvsof_multiscale_loss_forward does PLUS, in the same sweep over the flow,
 *   loss       = sum_k sum_s coef[k*n_scales + s] * term[k][s]      (coef [host] = w_k / n_scales)
 *   grad_flow  = d(loss) / d(flow[s])                               (scales->grad_flow, all written)
 * i.e. autograd's backward of utils/loss.py:121-171 for that fixed linear combination, without a
 * second pass over flow and frames: 16 B/pixel read, 8 B/pixel written.  Two launches: the
 * prelude (timestamp matching, all pyramid levels, and a count of the out-of-border pixels of
 * every tile, which the border term's gradient needs as n_i) and the sweep.  n_bad [dev int32
 * n_scales*P] receives the per-(scale, prediction) counts.  loss [dev] 1 float.  Deterministic. */
int dvsof_multiscale_loss_fused(const dvsof_loss_scales_t* scales,
                                const dvsof_loss_frames_t* frames, const float* coef,
                                int32_t* match_idx, float* terms, float* loss, int32_t* n_bad,
                                void* workspace, uint32_t* status, void* stream);

/* The same fused sweep on ready-made pyramid levels scales->images[s] and a given match_idx
 * (single-scale utils/loss.py:121-171 callers; kernel-level timing).  count_pass != 0: run the
 * out-of-border counting pre-pass into n_bad first; 0: n_bad already holds the counts. */
int dvsof_warp_loss_fused(const dvsof_loss_scales_t* scales, const int32_t* match_idx,
                          const float* coef, float* terms, float* loss, int32_t* n_bad,
                          void* workspace, int32_t count_pass, void* stream);

/* grad_flow[s] *= *factor [dev, 1 float] for every scale, in place, one launch; returns without
 * touching memory when *factor == 1 (backward of the fused pass: the incoming d/d(loss)). */
int dvsof_scale_gradients(const dvsof_loss_scales_t* scales, const float* factor, void* stream);

/* ---------------------------------------------------------------------------------------
 * Train-step glue (the network's convolutions stay on cuDNN).
 * ------------------------------------------------------------------------------------- */
/* EV_FlowNet/net.py:72-85, the head of every decoder block: nearest x2 upsampling followed by
 * ReflectionPad2d(1), as one gather.  x [dev] fp32: (N, C, H, W) contiguous, or (N, H, W, C) when
 * channels_last != 0; y likewise with 2H+2, 2W+2.  backward: grad_x = the adjoint, fixed
 * summation order. */
int dvsof_upsample2x_reflect_pad_forward(const float* x, int64_t N, int32_t C, int32_t H, int32_t W,
                                         float* y, int32_t channels_last, void* stream);
int dvsof_upsample2x_reflect_pad_backward(const float* grad_y, int64_t N, int32_t C, int32_t H,
                                          int32_t W, float* grad_x, int32_t channels_last,
                                          void* stream);

/* The same with the channel concatenation that builds the decoder input folded in
 * (EV_FlowNet/net.py:246-262 `cat((h, skip), 1)`): parts [host array of n_parts <= 4 device
 * pointers], each (N, H, W, part_channels[k]) channels_last fp32; pad_channels zero channels are
 * appended (channel count of y = sum + pad).  backward writes one contiguous gradient per part
 * (grad_parts[k] may be NULL: not needed).  channels_last only. */
int dvsof_upsample2x_reflect_pad_cat_forward(const float* const* parts, const int32_t* part_channels,
                                             int32_t n_parts, int32_t pad_channels, int64_t N, int32_t H,
                                             int32_t W, float* y, void* stream);
int dvsof_upsample2x_reflect_pad_cat_backward(const float* grad_y, float* const* grad_parts,
                                              const int32_t* part_channels, int32_t n_parts,
                                              int32_t pad_channels, int64_t N, int32_t H, int32_t W,
                                              void* stream);

/* EV_FlowNet/net.py:36,54, the tail of every conv block: (convolution bias,) ReLU and
 * BatchNorm2d in training mode: y = BN(relu(z + conv_bias)), batch statistics over (N, H, W),
 * biased variance; running statistics updated with `momentum`, unbiased variance, either may be
 * NULL.  Stats + apply passes that spread a channel over many CTAs.  z, y, grad_* [dev] fp32,
 * 16-byte aligned: (N, C, HW) contiguous, or (N, HW, C) when channels_last != 0; conv_bias [dev] C
 * floats or NULL (z is then the output of a bias-free convolution: the bias add and its gradient
 * reduction are folded in here); weight / bias [dev] C floats or NULL; save_mean / save_invstd
 * [dev] C floats (kept for backward).  backward: grad_z = d/dz, grad_weight / grad_bias /
 * grad_conv_bias [dev] C floats (or NULL).  Deterministic.
 * workspace: dvsof_relu_bn_workspace_bytes(N, C, HW, channels_last)  (channels_last: C % 4 == 0,
 * C <= 1024, else 0). */
size_t dvsof_relu_bn_workspace_bytes(int64_t N, int32_t C, int64_t HW, int32_t channels_last);
int dvsof_relu_bn_forward(const float* z, const float* conv_bias, int64_t N, int32_t C, int64_t HW,
                          const float* weight, const float* bias, float* running_mean,
                          float* running_var, float momentum, float eps, float* y, float* save_mean,
                          float* save_invstd, void* workspace, int32_t channels_last, void* stream);
int dvsof_relu_bn_backward(const float* z, const float* conv_bias, const float* grad_y, int64_t N,
                           int32_t C, int64_t HW, const float* weight, const float* save_mean,
                           const float* save_invstd, float* grad_z, float* grad_weight,
                           float* grad_bias, float* grad_conv_bias, void* workspace,
                           int32_t channels_last, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DVSOF_H_ */
