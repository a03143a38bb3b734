/* b200cnn.h -- C ABI of the B200-native cifar10cnn.py training step.
 *
 * The reference (Huzo/Distributed-Machine-Learning-using-CNN-CIFAR-10-dataset-, one file,
 * cifar10cnn.py) has no FFI or plugin interface: its hot path is reached through four Python
 * functions (create_cnn :94, cifar_loss :150, train_step :159, batch_accuracy :166) and the
 * loop line `mon_sess.run([train_op, loss])` (:230).  This header is the boundary a
 * maintainer binds instead of TensorFlow for that path; every entry point names the
 * reference lines it replaces.  INTEGRATION.md shows the ctypes stub.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no C++/torch types.
 *   - every call returns int: 0 = ok, negative = error (B200_E_*); b200_last_error() gives a
 *     thread-local message.  No C++ exception crosses the boundary.
 *   - all tensor pointers are DEVICE pointers unless the name says `host`.  The caller owns
 *     every buffer (parameters, gradients, workspace, inputs, outputs); the library
 *     owns only its opaque handle.  No call allocates device memory.
 *   - every call is asynchronous on the `stream` argument (a cudaStream_t passed as void*),
 *     does no host synchronisation and can be captured in a CUDA graph.
 *   - layouts are TensorFlow's: activations NHWC, conv filters HWIO, FC weights [in,out],
 *     labels int32, records uint8 CHW (cifar10cnn.py:57-62).
 *   - one handle per process / GPU, not thread-safe (the reference loop is one Python thread).
 */
#ifndef B200CNN_H_
#define B200CNN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK            0
#define B200_E_INVALID    -1   /* bad argument / unsupported shape */
#define B200_E_CUDA       -2   /* a CUDA runtime call or launch failed */
#define B200_E_STATE      -3   /* call order violated (e.g. backward before forward) */
#define B200_E_NODEVICE   -4   /* no usable sm_100 CUDA device */

#define B200_MODE_FP32  0      /* SIMT fp32 everywhere: the validation mode */
#define B200_MODE_BF16  1      /* tcgen05 bf16 operands, fp32 accumulation, fp32 master weights */

#define B200_NUM_PARAMS 10     /* variables of create_cnn in creation order, cifar10cnn.py:106-144 */

typedef struct b200_model* b200_handle;

/* Model / batch geometry.  Defaults of the script: batch 128 (:13), 32x32x3 records (:15-19)
 * centre-cropped to 24x24 (:17-18, :68), 64/64 conv channels (:106,:117), FC 384/192 (:131,:137),
 * 10 classes (:20). */
typedef struct {
  int32_t batch;
  int32_t img_h, img_w;      /* record image size */
  int32_t crop_h, crop_w;    /* what the CNN sees */
  int32_t c_in;              /* 3 */
  int32_t c1, c2;            /* conv1 / conv2 output channels */
  int32_t f1, f2;            /* full1 / full2 widths */
  int32_t n_cls;             /* <= 16 */
  int32_t mode;              /* B200_MODE_* */
} b200_config;

const char* b200_last_error(void);
int b200_version(void);

/* ---- handle ------------------------------------------------------------------------------ */
int b200_create(const b200_config* cfg, b200_handle* out);
int b200_destroy(b200_handle h);

/* Flat parameter / gradient buffers: the ten tensors of cifar10cnn.py:106-144 in creation
 * order, each starting at a multiple of 64 floats.  b200_param_layout fills offsets[10] and
 * sizes[10] (in floats) and the padded total.  Padding floats must be zero. */
int b200_param_layout(b200_handle h, int64_t* offsets, int64_t* sizes, int64_t* total);
int b200_bind_params(b200_handle h, float* params_flat);
int b200_bind_grads(b200_handle h, float* grads_flat);
int b200_workspace_bytes(b200_handle h, int64_t* nbytes);
int b200_bind_workspace(b200_handle h, void* workspace, int64_t nbytes);

/* create_cnn (:94-147) + cifar_loss (:150-157) + batch_accuracy (:166-176) on one batch of
 * records.  images: uint8 [batch,3,img_h,img_w]; labels: int32 [batch] (or [batch,1]).
 * The crop/transposes/cast of read_cifar_files (:57-68) happen inside.  Results stay in the
 * workspace; b200_result_ptrs exposes them. */
int b200_forward(b200_handle h, const uint8_t* images, const int32_t* labels, void* stream);
/* the gradient graph built by minimize() (:162-163); fills the bound gradient buffer.  The `images`
 * buffer given to the last b200_forward must still be intact (the conv1 gradient re-reads the records).
 * In bf16 mode independent branches run on two internal side streams forked from / joined to `stream` with
 * events; those streams are created by the first call, which must therefore happen before any capture.
 * One backward per forward: a second b200_backward without a new b200_forward returns B200_E_STATE. */
int b200_backward(b200_handle h, void* stream);
/* b200_forward + b200_backward as one call -- what sess.run(train_op) makes of :147-163 before the update.  Same
 * results as the two calls; knowing that the backward follows lets the bf16 path leave the FC2/FC3 parameter
 * gradients (and the loss / accuracy sums) on a side stream until the backward's final join, so they overlap FC1's
 * input gradient.  Everything is complete on `stream` when the call's work is. */
int b200_forward_backward(b200_handle h, const uint8_t* images, const int32_t* labels, void* stream);
/* ApplyGradientDescent x10 (:162-163): params -= lr * grad_scale * grads.  grad_scale = 1 for a
 * single worker or a SUM allreduce, 1/world for a MEAN allreduce. */
int b200_sgd_apply(b200_handle h, float lr, float grad_scale, void* stream);
/* The ps push/apply for world > 1 as ONE kernel per GPU over NVLink peer memory: all-reduce (sum) of the flat
 * gradient buffers of all workers fused with the update params -= lr * grad_scale * sum, written into every
 * worker's parameter buffer.  peer_grads / peer_params / peer_pads: host arrays of `world` device pointers
 * (entry i = worker i; entry `rank` = the buffers bound to this handle); all must be peer-mapped symmetric
 * allocations of the padded parameter size (pads: >= 64 B of zero-initialised uint32 flags per worker).
 * mc_grads / mc_params: NVSwitch multicast mappings of the same buffers, or NULL (then plain peer loads /
 * stores).  state: 16 zero-initialised bytes of local device memory.  Every worker must launch it once per
 * step; the kernels wait for one another, so all `world` GPUs must be running.  The wait is bounded: a worker
 * that has not arrived after B200_P2P_TIMEOUT_S seconds (environment variable, read once; default 120) makes every
 * waiting kernel trap, which leaves a sticky error on those CUDA contexts -- set it above the longest pause a
 * worker may take between two steps (checkpoint, evaluation, data stall). */
int b200_p2p_sgd_apply(b200_handle h, const void* const* peer_grads, void* const* peer_params, void* const* peer_pads,
                       const void* mc_grads, void* mc_params, uint32_t* state, int32_t rank, int32_t world, float lr,
                       float grad_scale, void* stream);
/* One `mon_sess.run([train_op, loss])` (:230) of one worker for world > 1: forward + backward with the exchange of
 * b200_p2p_sgd_apply split in two and overlapped -- the FC gradients (full_weight_1 onward, 90 % of the buffer) are
 * all-reduced and applied on an internal stream as soon as they exist, beside the conv layers' backward kernels;
 * the conv gradients follow at the end.  Same arguments and result as b200_forward_backward followed by
 * b200_p2p_sgd_apply, except: state = 32 zero-initialised bytes, pads >= 128 B per worker.  bf16 mode only. */
int b200_train_step_p2p(b200_handle h, const uint8_t* images, const int32_t* labels, const void* const* peer_grads,
                        void* const* peer_params, void* const* peer_pads, const void* mc_grads, void* mc_params,
                        uint32_t* state, int32_t rank, int32_t world, float lr, float grad_scale, void* stream);
/* Optional inbox for b200_train_step_p2p's SECOND channel (the conv gradients, exposed at the end of the step): with
 * it that channel runs as a one-shot exchange -- every worker pushes its conv gradients into a slot of every worker's
 * inbox (multimem.st through inbox_mc when given, peer stores otherwise), publishes one flag, waits for the others'
 * flags and sums the slots in rank order into its OWN parameters: one flag wait instead of two barriers, no remote
 * reads.  inbox_local: this worker's inbox, inbox_floats >= 2 * world * (offset of full_weight_1) floats of symmetric
 * memory, zero-initialised; inbox_peers: host array of `world` device pointers to every worker's inbox.  NULL
 * inbox_local switches back to the two-barrier kernel.  No counterpart in the reference (its gRPC runtime did this). */
int b200_p2p_set_inbox(b200_handle h, void* inbox_local, void* inbox_mc, void* const* inbox_peers, int32_t world,
                       int64_t inbox_floats);
/* The same step with the exchange as ONE kernel at the end (b200_forward_backward + b200_p2p_sgd_apply in one call;
 * state = 16 zero-initialised bytes, pads >= 64 B): what eight workers use, where a second pair of flag barriers costs
 * more than overlapping the FC gradients saves.  conv1's gradient partials are summed inside the exchange kernel. */
int b200_train_step_p2p_single(b200_handle h, const uint8_t* images, const int32_t* labels, const void* const* peer_grads,
                               void* const* peer_params, void* const* peer_pads, const void* mc_grads, void* mc_params,
                               uint32_t* state, int32_t rank, int32_t world, float lr, float grad_scale, void* stream);
/* forward + backward + sgd_apply: one `mon_sess.run([train_op, loss])` (:230) for world == 1. */
int b200_train_step(b200_handle h, const uint8_t* images, const int32_t* labels, float lr, void* stream);
/* The same call with HOST (pinned) inputs and a host loss: copies the records and labels to
 * the device, runs the step, copies the scalar loss back -- all on `stream`.  The caller
 * synchronises the stream before reading *loss_host. */
int b200_train_step_host(b200_handle h, const uint8_t* images_host, const int32_t* labels_host,
                         float lr, float* loss_host, void* stream);
/* The same with the input copy taken off the critical path (what an input pipeline with prefetch does for the loop of
 * :228-230): two staging slots in the workspace.  b200_stage_host_batch copies a pinned host batch into `slot`
 * (0 or 1) on `stream`; b200_train_step_host_prefetch trains on the batch already in `slot` while the NEXT pinned
 * host batch (may be NULL at the end of the data) is copied into the other slot on an internal copy stream, forked
 * from and joined to `stream` with events (created by the first call: make it before any capture), and copies the
 * scalar loss back.  The caller alternates `slot` and synchronises the stream before reading *loss_host. */
int b200_stage_host_batch(b200_handle h, int32_t slot, const uint8_t* images_host, const int32_t* labels_host, void* stream);
int b200_train_step_host_prefetch(b200_handle h, int32_t slot, const uint8_t* next_images_host,
                                  const int32_t* next_labels_host, float lr, float* loss_host, void* stream);

/* One eager b200_train_step with a CUDA event recorded on `stream` before every launch site
 * (a site = the kernels of one logical op: "conv2_fwd", "full1_bwd", "sgd_apply" ...).  Blocks
 * until the step is done and returns the device time of each site in milliseconds; the bench
 * uses it for the live roofline of the dominant kernel.  Not capturable (it synchronises). */
int b200_profile_step(b200_handle h, const uint8_t* images, const int32_t* labels, float lr, void* stream,
                      float* ms_per_site, int32_t max_sites, int32_t* n_sites);
const char* b200_site_name(b200_handle h, int32_t i);

/* device pointers to the results of the last forward: logits float[batch,n_cls],
 * loss float[1] (mean), n_correct int32[1], pred int32[batch].  Any out pointer may be NULL. */
int b200_result_ptrs(b200_handle h, float** logits, float** loss, int32_t** n_correct, int32_t** pred);
/* named layer-boundary tensor inside the workspace (tests): "x","conv1","pool1","arg1","conv2",
 * "pool2","arg2","full1","full2","logits","dlogits","dfull2","dfull1","dpool2","dconv2",
 * "dpool1","dconv1".  Returns the device pointer, its byte size and element size. */
int b200_tensor_ptr(b200_handle h, const char* name, void** ptr, int64_t* nbytes, int32_t* elem_size);
/* how many kernels one b200_train_step launches (for the bench's gpu_launches claim). */
int b200_launches_per_step(b200_handle h, int32_t* n);

/* ---- per-op entry points (kernel-level parity tests; same kernels the handle path uses) ------ */
/* read_cifar_files crop/cast, cifar10cnn.py:57-68 */
int b200_crop_cast_f32(const uint8_t* images, float* out, int32_t batch, int32_t img_h, int32_t img_w,
                       int32_t crop_h, int32_t crop_w, int32_t c_in, void* stream);
/* tf.nn.conv2d SAME 5x5 s1 + bias_add + relu, :107-110 / :118-121.  bias may be NULL, relu 0/1. */
int b200_conv5x5_fwd_f32(const float* x, const float* w, const float* bias, float* y,
                         int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out, int32_t relu,
                         void* stream);
/* Conv2DBackpropInput (autodiff of :118).  w_scratch: c_in*c_out*25 floats of scratch. */
int b200_conv5x5_dgrad_f32(const float* dy, const float* w, float* w_scratch, float* dx,
                           int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out, void* stream);
/* Conv2DBackpropFilter + BiasAddGrad (autodiff of :107-109 / :118-120).  scratch holds the
 * split-K partial sums; b200_conv5x5_wgrad_scratch_bytes gives its size.  db may be NULL. */
int b200_conv5x5_wgrad_scratch_bytes(int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out,
                                     int64_t* nbytes);
int b200_conv5x5_wgrad_f32(const float* x, const float* dy, float* dw, float* db, void* scratch,
                           int64_t scratch_bytes, int32_t batch, int32_t h, int32_t wd, int32_t c_in,
                           int32_t c_out, void* stream);
/* tf.nn.max_pool k3 s2 SAME (:113 / :123); arg = kh*3+kw of the first maximal cell. */
int b200_maxpool_fwd_f32(const float* x, float* y, uint8_t* arg, int32_t batch, int32_t h, int32_t wd,
                         int32_t c, void* stream);
/* MaxPoolGrad fused with the ReluGrad of the layer that fed the pool: dx = scatter(dy) * [x > 0]
 * (relu_mask = 1) or plain MaxPoolGrad (relu_mask = 0). */
int b200_maxpool_bwd_f32(const float* dy, const uint8_t* arg, const float* x, float* dx, int32_t batch,
                         int32_t h, int32_t wd, int32_t c, int32_t relu_mask, void* stream);
/* relu(x @ w + b), :133 / :139 / :145 */
int b200_fc_fwd_f32(const float* x, const float* w, const float* b, float* y, int32_t batch, int32_t k,
                    int32_t n, int32_t relu, void* stream);
/* gradients of the above given dy (already masked by this layer's ReluGrad):
 * dw = x^T dy, db = colsum(dy), dx = (dy w^T) * [x > 0] if mask_x else dy w^T.
 * dx may be NULL.  scratch: b200_colsum_scratch_bytes(batch, n). */
int b200_colsum_scratch_bytes(int32_t rows, int32_t cols, int64_t* nbytes);
int b200_fc_bwd_f32(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                    void* scratch, int64_t scratch_bytes, int32_t batch, int32_t k, int32_t n,
                    int32_t mask_x, void* stream);
/* The whole first block of create_cnn on the tensor cores, one image per CTA: record crop/cast (:57-68),
 * conv1 + bias + ReLU (:107-110) and pool1 with argmax (:113); only pool and arg reach HBM (conv_dbg,
 * if not NULL, also receives the un-pooled [batch,crop_h,crop_w,64] activations for tests).  3 -> 64
 * channels only; crops up to about 24x24 (TMEM / shared-memory limit). */
int b200_conv1_pool_fwd_bf16(const uint8_t* images, const float* w, const float* bias, float* pool, uint8_t* arg,
                             float* conv_dbg, int32_t batch, int32_t img_h, int32_t img_w, int32_t crop_h,
                             int32_t crop_w, void* stream);
/* Its gradient: MaxPoolGrad + ReluGrad of pool1 (autodiff of :113, :110) fused with
 * Conv2DBackpropFilter + BiasAddGrad of :107-109, from the gradient w.r.t. pool1. */
int b200_conv1_wgrad_bf16_scratch_bytes(int32_t batch, int64_t* nbytes);
int b200_conv1_wgrad_bf16(const uint8_t* images, const float* dpool, const float* pool, const uint8_t* arg, float* dw,
                          float* db, void* scratch, int64_t scratch_bytes, int32_t batch, int32_t img_h, int32_t img_w,
                          int32_t crop_h, int32_t crop_w, void* stream);
/* conv2d / Conv2DBackpropInput / Conv2DBackpropFilter of :118 on the tensor cores (tcgen05 implicit
 * GEMM over a zero-padded shared-memory strip; bf16 operands rounded from the fp32 tensors, fp32
 * accumulation).  c_in % 16 == 0 and c_out % 64 == 0 (wgrad: 64 -> 64 only).  The wgrad scratch
 * holds per-CTA partial sums that a second kernel adds in a fixed order.
 * forward and dgrad take the filter PACKED: b200_conv5x5_pack_weights turns the fp32 HWIO tensor into
 * bf16 shared-memory images (b200_conv5x5_packed_bytes bytes) that the kernels fetch with bulk copies;
 * one packed copy serves both directions. */
/* b200_conv1_pool_fwd_bf16 / b200_conv1_wgrad_bf16 for a conv1 with c1 output channels (a multiple of 64; the
 * width-scaled model of BASELINE.json has 256): w [5,5,3,c1], bias [c1], pool / arg / dpool [batch,OH,OW,c1], dw
 * [5,5,3,c1], db [c1]; one CTA per image and 64-channel slice.  scratch for the gradient: 19456 bytes x min(batch, 148)
 * x c1 / 64.  Replace :107-113 and their gradients like the 64-channel entry points. */
int b200_conv1_pool_fwd_bf16_c(const uint8_t* images, const float* w, const float* bias, float* pool, uint8_t* arg,
                               float* conv_dbg, int32_t batch, int32_t img_h, int32_t img_w, int32_t crop_h,
                               int32_t crop_w, int32_t c1, void* stream);
int b200_conv1_wgrad_bf16_c(const uint8_t* images, const float* dpool, const float* pool, const uint8_t* arg, float* dw,
                            float* db, void* scratch, int64_t scratch_bytes, int32_t batch, int32_t img_h, int32_t img_w,
                            int32_t crop_h, int32_t crop_w, int32_t c1, void* stream);
/* diagnostics: when given a device buffer of 32 int64 (conv stamps in [0,8), FC-tail stamps in [16,25)), CTA 0 of the conv2 forward / dgrad kernel stores
 * clock64() at its phase boundaries (start, strip zeroed, strip filled, MMA start, MMA issued,
 * accumulators ready, epilogue done); NULL switches it off. */
int b200_debug_conv_timeline(int64_t* device_buf32);
/* diagnostics / tests: the conv2 forward and input-gradient kernels come in two shapes -- one CTA per SM with
 * filter stages of five taps (single-wave batches), and two CTAs per SM with single-tap stages (chosen automatically
 * once the batch exceeds the 148 SMs, so that one CTA's MMAs overlap the other's strip fill and epilogue).
 * 0 / 1 force the first / second shape wherever it is legal, -1 restores the automatic choice.  Process-wide.
 * No counterpart in the reference. */
int b200_debug_conv_variant(int32_t variant);
/* diagnostics, only in a library built with B200_TRACE=1 (else B200_E_STATE): every kernel of the bf16 step stores
 * %globaltimer (ns) of its first CTA's start / last CTA's end into device_buf[2*k], device_buf[2*k+1] (atomic min /
 * max: preset them to ~0 / 0), k = pack, W2 image, conv1 fwd, conv2 fwd, FC1 fwd, tail rows, tail grads, FC1 dW,
 * FC1 dX, conv2 dgrad, conv2 wgrad, wgrad reduce, conv1 wgrad, SGD.  No counterpart in the reference. */
int b200_debug_trace(uint64_t* device_buf);
int b200_conv5x5_packed_bytes(int32_t c_in, int32_t c_out, int64_t* nbytes);
int b200_conv5x5_pack_weights(const float* w, void* w_packed, int32_t c_in, int32_t c_out, void* stream);
int b200_conv5x5_fwd_bf16(const float* x, const void* w_packed, const float* bias, float* y, int32_t batch, int32_t h,
                          int32_t wd, int32_t c_in, int32_t c_out, int32_t relu, void* stream);
/* conv + bias + ReLU (:118-121) fused with the 3x3/2 SAME max-pool that follows (:123): pool [batch,ceil(h/2),
 * ceil(wd/2),64] and its argmax; y (un-pooled) may be NULL.  64 output channels, one image per CTA. */
int b200_conv5x5_pool_fwd_bf16(const float* x, const void* w_packed, const float* bias, float* y, float* pool,
                               uint8_t* arg, int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out,
                               void* stream);
/* b200_conv5x5_pool_fwd_bf16 with the input activations already in bf16 (dense [B,H,W,c_in] bf16, what conv1's pool
 * epilogue leaves behind in the step): the strip is then written by tensor-map TMA boxes with out-of-bounds zero
 * fill instead of loads through registers.  Replaces :118-123 likewise. */
int b200_conv5x5_pool_fwd_bf16x(const void* x_bf16, const void* w_packed, const float* bias, float* y, float* pool,
                                uint8_t* arg, int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out,
                                void* stream);
int b200_conv5x5_dgrad_bf16(const float* dy, const void* w_packed, float* dx, int32_t batch, int32_t h, int32_t wd,
                            int32_t c_in, int32_t c_out, void* stream);
int b200_conv5x5_wgrad_bf16_scratch_bytes(int32_t batch, int64_t* nbytes);
/* scratch of b200_conv5x5_wgrad_bf16 for c_in -> c_out channels (multiples of 64; the kernel works on 64 x 64 channel
 * blocks): b200_conv5x5_wgrad_bf16_scratch_bytes is the 64 -> 64 case.  Conv2DBackpropFilter of :118. */
int b200_conv5x5_wgrad_bf16_scratch_bytes_c(int32_t batch, int32_t c_in, int32_t c_out, int64_t* nbytes);
int b200_conv5x5_wgrad_bf16(const float* x, const float* dy, float* dw, void* scratch, int64_t scratch_bytes,
                            int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out, void* stream);
/* Conv2DBackpropInput / Conv2DBackpropFilter fed by the POOLED gradient: MaxPoolGrad (:123) and the ReluGrad
 * of :121 are applied while the kernels build their shared-memory strips, so the gradient w.r.t. the conv
 * output never exists in HBM.  pool / arg are the forward pool output and argmax, dpool its gradient, all
 * [batch, ceil(h/2), ceil(wd/2), 64].  dbias (optional, must be zero on entry) receives BiasAddGrad (:120). */
int b200_conv5x5_dgrad_pool_bf16(const float* dpool, const float* pool, const uint8_t* arg, const void* w_packed,
                                 float* dx, float* dbias, int32_t batch, int32_t h, int32_t wd, int32_t c_in,
                                 int32_t c_out, void* stream);
int b200_conv5x5_wgrad_pool_bf16(const float* x, const float* dpool, const float* pool, const uint8_t* arg, float* dw,
                                 void* scratch, int64_t scratch_bytes, int32_t batch, int32_t h, int32_t wd,
                                 int32_t c_in, int32_t c_out, void* stream);
/* The same two products on the tensor cores (tcgen05, bf16 operands rounded from the fp32 tensors
 * on the fly, fp32 accumulation in TMEM).  batch, k and n must be multiples of 8. */
int b200_fc_fwd_bf16(const float* x, const float* w, const float* b, float* y, int32_t batch, int32_t k, int32_t n,
                     int32_t relu, void* stream);
int b200_fc_bwd_bf16(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                     void* scratch, int64_t scratch_bytes, int32_t batch, int32_t k, int32_t n,
                     int32_t mask_x, void* stream);
/* b200_fc_bwd_f32 without the scratch arguments, for callers that cannot allocate device memory:
 * plug-ins of a framework whose caching pool already owns all free HBM, operators that are only handed
 * tensors and have no allocator of their own, and step loops that may not make a device-synchronising
 * call such as cudaMalloc.  It has no counterpart in
 * the reference (TensorFlow allocates for the script).  The column-sum scratch is a fixed
 * B200_STATIC_SCRATCH_BYTES array inside the library image: each GPU that loads the library holds it until
 * the process exits, and a request needing more (see b200_colsum_scratch_bytes(batch, n)) is refused
 * with B200_E_INVALID.  The first call on a device resolves the array's address; make it before any
 * stream capture.  All calls share the array (contents persist, nothing needs zeroing); calls that
 * overlap on the GPU race on it.  Allocates nothing. */
#define B200_STATIC_SCRATCH_BYTES (1 << 20)
int b200_fc_bwd_f32_static(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                           int32_t batch, int32_t k, int32_t n, int32_t mask_x, void* stream);
/* Everything above full1 in two launches: full2 (:139), logits (:145), cifar_loss (:150-157),
 * batch_accuracy (:166-176) and all their gradients (dfull1 = gradient w.r.t. full1 after its ReluGrad,
 * dW2, db2, dW3, db3, db1 = colsum(dfull1)).  fp32; sums over the batch in a fixed order.  full1 is the
 * [batch,f1] input (post-ReLU); f1 % 32 == 0, f2 % 4 == 0, f2 <= 256. */
int b200_fc_tail_scratch_bytes(int32_t batch, int32_t f1, int32_t f2, int32_t n_cls, int64_t* nbytes);
int b200_fc_tail_f32(float* full1, const float* w2, const float* b2, const float* w3, const float* b3,
                     const int32_t* labels, float* full2, float* logits, float* dlogits, float* dfull2, float* dfull1,
                     int32_t* pred, float* loss, int32_t* n_correct, int32_t* n_bad, float* dw2, float* db2, float* dw3,
                     float* db3, float* db1, void* scratch, int64_t scratch_bytes, int32_t batch, int32_t f1, int32_t f2,
                     int32_t n_cls, void* stream);
/* cifar_loss (:150-157), batch_accuracy (:166-176) and the first backward step: dlogits =
 * (softmax - onehot)/batch * [logits > 0] (ReLU on the logits, :145).  Labels outside
 * [0,n_cls) set *n_bad (int32, device) > 0 and contribute nothing. */
int b200_head_f32(const float* logits, const int32_t* labels, float* loss_mean, float* dlogits,
                  int32_t* pred, int32_t* n_correct, int32_t* n_bad, int32_t batch, int32_t n_cls,
                  void* stream);
/* ApplyGradientDescent (:162-163): p -= lr * scale * g over n floats. */
/* tf.train.shuffle_batch(capacity 5384, min_after_dequeue 5000) over the file list of input_pipeline (:72-91) with
 * the record set resident in HBM: images uint8 [n_records][image_bytes] (the 3072 CHW payload bytes of each 3073-byte
 * record, :57-62), labels int32 [n_records].  One call = one dequeue_many(batch) + the reader's refill:
 *   - slots_in [capacity] holds the record ids in the queue; `batch` of them are drawn uniformly without replacement
 *     (partial Fisher-Yates, draw i = splitmix64(seed, step, i): counter-based, reproducible on the host);
 *   - the drawn records are copied to images_out [batch][image_bytes] / labels_out [batch];
 *   - slots_out (a DIFFERENT buffer) receives the queue with the freed slots refilled by stream positions
 *     stream_pos .. stream_pos+batch-1 of the sequential reader: position p is record
 *     file_order[(p / per_file) % n_order] * per_file + p % per_file (file_order = the per-epoch shuffled file lists,
 *     flattened, as string_input_producer produces them).
 * The caller alternates the two slot buffers and advances step by 1 and stream_pos by batch. */
int b200_shuffle_batch(const uint8_t* images, const int32_t* labels, int64_t n_records, const int32_t* file_order,
                       int32_t n_order, int32_t per_file, const int32_t* slots_in, int32_t* slots_out, int32_t capacity,
                       uint64_t seed, int64_t step, int64_t stream_pos, uint8_t* images_out, int32_t* labels_out,
                       int32_t batch, int32_t image_bytes, void* stream);
int b200_sgd_apply_f32(float* p, const float* g, int64_t n, float lr, float scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200CNN_H_ */
