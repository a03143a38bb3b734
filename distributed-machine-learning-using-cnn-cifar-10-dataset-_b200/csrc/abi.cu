// C ABI of include/b200cnn.h: the opaque handle, workspace carving and the orchestration of
// one training step (cifar10cnn.py:230) out of the kernels in f32_kernels.cu / bf16_*.cu.
#include <stdarg.h>
#include <string.h>

#include <new>
#include <string>
#include <vector>

#include <algorithm>

#include "launch.h"

namespace b200 {
thread_local bool g_pdl_next = false;

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

struct Tensor {
  std::string name;
  int64_t elems;
  int32_t esize;
  int64_t offset;   // bytes into the workspace
};

}  // namespace b200

using namespace b200;

struct b200_model {
  b200_config cfg;
  int h1, w1, h2, w2, flat;             // pool1 / pool2 spatial sizes and the FC1 fan-in (:127)
  int64_t p_off[B200_NUM_PARAMS], p_size[B200_NUM_PARAMS], p_total;
  float* params = nullptr;
  float* grads = nullptr;
  char* ws = nullptr;
  int64_t ws_bytes = 0;
  std::vector<Tensor> tensors;
  bool forward_done = false;
  // side streams / events for the independent branches of the backward pass (created at the first
  // bf16 backward, which must therefore run once eagerly before any stream capture)
  cudaStream_t copy_stream = nullptr;      // host-batch prefetch (b200_train_step_host_prefetch)
  cudaEvent_t ev_copy[2] = {nullptr, nullptr};
  cudaStream_t side[2] = {nullptr, nullptr};
  cudaEvent_t ev_fork[2] = {nullptr, nullptr}, ev_join[2] = {nullptr, nullptr}, ev_aux[2] = {nullptr, nullptr};
  const uint8_t* last_images = nullptr;   // records of the last forward (the fused conv1 gradient re-reads them)
  // inbox of the one-shot conv-gradient exchange (b200_p2p_set_inbox), or null
  float* inbox_local = nullptr; void* inbox_mc = nullptr; void* inbox_peers[8] = {}; int64_t inbox_floats = 0;
  // per-launch-site timing (b200_profile_step): events are created lazily, once
  bool profiling = false;
  std::vector<std::string> site_names;
  std::vector<cudaEvent_t> site_events;
  int site_cursor = 0;

  int64_t add(const char* name, int64_t elems, int32_t esize) {
    Tensor t{name, elems, esize, ws_bytes};
    tensors.push_back(t);
    ws_bytes += round_up(elems * esize, 256);
    return t.offset;
  }
  const Tensor* find(const char* name) const {
    for (auto& t : tensors)
      if (t.name == name) return &t;
    return nullptr;
  }
  template <class T>
  T* ptr(const char* name) const {
    const Tensor* t = find(name);
    return t ? reinterpret_cast<T*>(ws + t->offset) : nullptr;
  }
  float* P(int i) const { return params + p_off[i]; }
  float* G(int i) const { return grads + p_off[i]; }
};

enum { P_W1 = 0, P_B1, P_W2, P_B2, P_FW1, P_FB1, P_FW2, P_FB2, P_FW3, P_FB3 };

// bf16 mode has two schedules.  The script's own widths (64 / 64 conv channels, FC <= 256 wide) run the fused,
// multi-stream step tuned for them (conv + pool in one kernel, FC tail in two kernels, gradients gathered through the
// pools).  Any other width (BASELINE.json's width-scaled model: 256 / 256 channels, FC 768 / 384) runs the layer-by-
// layer schedule below it: the contractions -- conv2 forward / input gradient / filter gradient, FC1, FC2 -- on the
// same tcgen05 kernels (channel-blocked), pooling, the 3-channel conv1, the 10-wide FC3 and the loss on the fp32
// kernels of the validation mode.
static bool fused_bf16_shapes(const b200_config& c) {
  return c.c1 == 64 && c.c2 == 64 && c.f1 % 32 == 0 && c.f2 % 32 == 0 && c.f2 <= 256;
}

extern "C" {

const char* b200_last_error(void) { return g_err; }
int b200_version(void) { return 100; }

int b200_create(const b200_config* c, b200_handle* out) {
  B200_REQUIRE(c && out, "b200_create: null argument");
  B200_REQUIRE(c->batch > 0 && c->batch <= (1 << 16), "b200_create: batch must be in 1..65536");
  B200_REQUIRE(c->crop_h >= 4 && c->crop_w >= 4 && c->crop_h <= c->img_h && c->crop_w <= c->img_w,
               "b200_create: crop %dx%d must fit the %dx%d record and be >= 4", c->crop_h, c->crop_w,
               c->img_h, c->img_w);
  B200_REQUIRE(c->c_in > 0 && c->c1 > 0 && c->c2 > 0 && c->f1 > 0 && c->f2 > 0, "b200_create: bad widths");
  B200_REQUIRE(c->c1 % 4 == 0 && c->c2 % 4 == 0 && c->f1 % 4 == 0 && c->f2 % 4 == 0,
               "b200_create: channel / FC widths must be multiples of 4");
  B200_REQUIRE(c->n_cls > 0 && c->n_cls <= 16, "b200_create: n_cls must be in 1..16");
  B200_REQUIRE(c->mode == B200_MODE_FP32 || c->mode == B200_MODE_BF16, "b200_create: unknown mode %d", c->mode);
  if (c->mode == B200_MODE_BF16)
    B200_REQUIRE(c->batch % 8 == 0 && c->c1 % 64 == 0 && c->c2 % 64 == 0 && c->c1 <= 256 && c->c2 <= 256 &&
                     c->f1 % 8 == 0 && c->f2 % 8 == 0,
                 "b200_create: bf16 mode needs batch %% 8 == 0, conv channels in {64, 128, 192, 256}, f1 %% 8 == 0, f2 %% 8 == 0");
  b200_model* h = new (std::nothrow) b200_model();
  B200_REQUIRE(h, "b200_create: out of host memory");
  h->cfg = *c;
  h->h1 = (c->crop_h + 1) / 2; h->w1 = (c->crop_w + 1) / 2;
  h->h2 = (h->h1 + 1) / 2;     h->w2 = (h->w1 + 1) / 2;
  h->flat = h->h2 * h->w2 * c->c2;
  const int64_t sz[B200_NUM_PARAMS] = {25ll * c->c_in * c->c1, c->c1, 25ll * c->c1 * c->c2, c->c2,
                                       (int64_t)h->flat * c->f1, c->f1, (int64_t)c->f1 * c->f2, c->f2,
                                       (int64_t)c->f2 * c->n_cls, c->n_cls};
  int64_t off = 0;
  for (int i = 0; i < B200_NUM_PARAMS; ++i) {
    h->p_off[i] = off; h->p_size[i] = sz[i];
    off += round_up(sz[i], 64);
  }
  h->p_total = off;

  const int64_t B = c->batch;
  const int64_t n_x = B * c->crop_h * c->crop_w * c->c_in, n_c1 = B * c->crop_h * c->crop_w * c->c1,
                n_p1 = B * h->h1 * h->w1 * c->c1, n_c2 = B * h->h1 * h->w1 * c->c2,
                n_p2 = B * h->flat;
  h->add("x", n_x, 4);
  h->add("conv1", n_c1, 4);   h->add("pool1", n_p1, 4);  h->add("arg1", n_p1, 1);
  h->add("conv2", n_c2, 4);   h->add("pool2", n_p2, 4);  h->add("arg2", n_p2, 1);
  h->add("full1", B * c->f1, 4); h->add("full2", B * c->f2, 4); h->add("logits", B * c->n_cls, 4);
  h->add("dlogits", B * c->n_cls, 4); h->add("dfull2", B * c->f2, 4); h->add("dfull1", B * c->f1, 4);
  h->add("dpool2", n_p2, 4);  h->add("dconv2", n_c2, 4);
  h->add("pool1_bf16", n_p1, 2);  h->add("dconv2_bf16", n_c2, 2);      // bf16 mode: operand copies for conv2 fwd / wgrad
  h->add("dpool1", n_p1, 4);  h->add("dconv1", n_c1, 4);
  h->add("wrot", 25ll * c->c1 * c->c2, 4);
  int64_t scratch = conv5x5_wgrad_scratch_bytes(c->batch, c->crop_h, c->crop_w, c->c_in, c->c1);
  scratch = std::max(scratch, conv5x5_wgrad_scratch_bytes(c->batch, h->h1, h->w1, c->c1, c->c2));
  scratch = std::max(scratch, colsum_scratch_bytes(c->batch, std::max(c->f1, std::max(c->f2, c->n_cls))));
  scratch = std::max(scratch, conv1_wgrad_tc_scratch_bytes(c->batch, c->c1 % 64 == 0 ? c->c1 : 64));
  if (c->c1 % 64 == 0 && c->c2 % 64 == 0)
    scratch = std::max(scratch, conv5x5_wgrad_tc_scratch_bytes(c->batch, c->c1, c->c2));
  scratch = std::max(scratch, colsum_scratch_bytes(c->batch * h->h1 * h->w1, c->c2));
  h->add("scratch", scratch, 1);
  h->add("scratch2", conv5x5_wgrad_tc_scratch_bytes(c->batch), 1);      // conv2 wgrad (64 -> 64) runs beside conv1 wgrad
  h->add("db2_part", B * 64, 4);      // conv2's bias gradient: one partial per dgrad CTA, summed in a fixed order
  h->add("w2_packed", conv5x5_packed_bytes(c->c1, c->c2), 1);
  h->add("fc1_part", (int64_t)fc_fwd_tc_splits(c->batch, h->flat, c->f1) * B * c->f1, 4);
  h->add("tail_scratch", fc_tail_scratch_bytes(c->batch, c->f1, c->f2, c->n_cls), 1);
  // FC1's bf16 operand images (bf16 mode, shapes permitting): W1 packed once per step, pool2 written by conv2's pool
  // epilogue, dfull1 written by the FC tail -- the three FC1 products fetch them with bulk copies
  h->add("w1_img", fc_weight_image_bytes(h->flat, c->f1), 1);
  h->add("pool2_img", (int64_t)h->flat * B, 2);
  h->add("dfull1_img", (int64_t)c->f1 * B, 2);
  h->add("loss", 1, 4); h->add("n_correct", 1, 4); h->add("n_bad", 1, 4); h->add("pred", B, 4);
  h->add("stage_images", B * c->c_in * c->img_h * c->img_w, 1);
  h->add("stage_labels", B, 4);
  h->add("stage_images1", B * c->c_in * c->img_h * c->img_w, 1);      // second staging slot (b200_train_step_host_prefetch)
  h->add("stage_labels1", B, 4);
  *out = h;
  return B200_OK;
}

int b200_destroy(b200_handle h) {
  if (h) {
    for (cudaEvent_t ev : h->site_events) cudaEventDestroy(ev);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    for (int i = 0; i < 2; ++i)
      if (h->ev_copy[i]) cudaEventDestroy(h->ev_copy[i]);
    for (int i = 0; i < 2; ++i) {
      if (h->ev_fork[i]) cudaEventDestroy(h->ev_fork[i]);
      if (h->ev_join[i]) cudaEventDestroy(h->ev_join[i]);
      if (h->ev_aux[i]) cudaEventDestroy(h->ev_aux[i]);
      if (h->side[i]) cudaStreamDestroy(h->side[i]);
    }
  }
  delete h;
  return B200_OK;
}

int b200_param_layout(b200_handle h, int64_t* offsets, int64_t* sizes, int64_t* total) {
  B200_REQUIRE(h, "null handle");
  for (int i = 0; i < B200_NUM_PARAMS; ++i) {
    if (offsets) offsets[i] = h->p_off[i];
    if (sizes) sizes[i] = h->p_size[i];
  }
  if (total) *total = h->p_total;
  return B200_OK;
}
int b200_bind_params(b200_handle h, float* p) {
  B200_REQUIRE(h && p && ((uintptr_t)p & 255) == 0, "bind_params: need a 256-byte aligned device pointer");
  h->params = p;
  return B200_OK;
}
int b200_bind_grads(b200_handle h, float* g) {
  B200_REQUIRE(h && g && ((uintptr_t)g & 255) == 0, "bind_grads: need a 256-byte aligned device pointer");
  h->grads = g;
  return B200_OK;
}
int b200_workspace_bytes(b200_handle h, int64_t* n) {
  B200_REQUIRE(h && n, "null argument");
  *n = h->ws_bytes;
  return B200_OK;
}
int b200_bind_workspace(b200_handle h, void* ws, int64_t nbytes) {
  B200_REQUIRE(h && ws && ((uintptr_t)ws & 255) == 0, "bind_workspace: need a 256-byte aligned device pointer");
  B200_REQUIRE(nbytes >= h->ws_bytes, "bind_workspace: %lld bytes given, %lld needed", (long long)nbytes,
               (long long)h->ws_bytes);
  h->ws = (char*)ws;
  h->forward_done = false;
  return B200_OK;
}

#define B200_READY(h)                                                                              \
  do {                                                                                             \
    B200_REQUIRE(h, "null handle");                                                                \
    if (!(h)->params || !(h)->grads || !(h)->ws) {                                                 \
      set_error("handle not ready: bind params, grads and workspace first");                       \
      return B200_E_STATE;                                                                         \
    }                                                                                              \
  } while (0)
#define TRY(expr) do { int rc__ = (expr); if (rc__) return rc__; } while (0)

// side streams / events for the independent branches of a bf16 step (lazily, at the first eager call)
static int ensure_side_streams(b200_model* h) {
  if (h->side[0]) return B200_OK;
  int prio_lo = 0, prio_hi = 0;      // stream 0 (FC1's dW: 108 big CTAs that the next main-stream kernel must not wait for)
  B200_CUDA_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));      // gets its CTAs placed first
  for (int i = 0; i < 2; ++i) {
    B200_CUDA_TRY(cudaStreamCreateWithPriority(&h->side[i], cudaStreamNonBlocking, i == 0 ? prio_hi : prio_lo));
    B200_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_fork[i], cudaEventDisableTiming));
    B200_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_join[i], cudaEventDisableTiming));
    B200_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_aux[i], cudaEventDisableTiming));
  }
  return B200_OK;
}

// A launch site: one logical kernel group of the step.  When b200_profile_step is active an
// event is recorded on the stream before each site (and one after the last).
static int site_mark(b200_model* h, const char* name, cudaStream_t st) {
  if (!h->profiling) return B200_OK;
  int i = h->site_cursor++;
  if ((int)h->site_events.size() <= i) {
    cudaEvent_t ev;
    B200_CUDA_TRY(cudaEventCreate(&ev));
    h->site_events.push_back(ev);
    h->site_names.push_back(name);
  } else {
    h->site_names[i] = name;
  }
  B200_CUDA_TRY(cudaEventRecord(h->site_events[i], st));
  return B200_OK;
}
#define SITE(name, expr) do { TRY(site_mark(h, name, st)); TRY(expr); } while (0)
// a site whose (first) kernel follows its producer on the same stream: launched with programmatic stream
// serialization so its prologue overlaps the producer's tail (see common.cuh); ordinary while profiling per site.
// Used where the early-resident consumer does not take SMs away from a side-stream kernel: conv2's dgrad and
// conv1's wgrad (128 CTAs each) stay ordinary launches -- resident and waiting they starved FC1's dW and conv2's
// wgrad on the side streams (measured with the B200_TRACE schedule).
#define PSITE(name, expr)                    \
  do {                                       \
    TRY(site_mark(h, name, st));             \
    g_pdl_next = !h->profiling;              \
    const int rc__ = (expr);                 \
    g_pdl_next = false;                      \
    if (rc__) return rc__;                   \
  } while (0)

// defer_tail: the caller runs b200_backward right after on the same stream, so the tail's parameter-gradient
// kernel (which also writes loss / n_correct) may stay on side stream 0 and be joined by the backward's final join
static int forward_impl(b200_handle h, const uint8_t* images, const int32_t* labels, void* stream, bool defer_tail) {
  B200_READY(h);
  B200_REQUIRE(images && labels, "forward: null input");
  cudaStream_t st = (cudaStream_t)stream;
  const b200_config& c = h->cfg;
  const bool tc = c.mode == B200_MODE_BF16;
  float* x = h->ptr<float>("x");
  float *conv1 = h->ptr<float>("conv1"), *pool1 = h->ptr<float>("pool1");
  float *conv2 = h->ptr<float>("conv2"), *pool2 = h->ptr<float>("pool2");
  float *full1 = h->ptr<float>("full1"), *full2 = h->ptr<float>("full2"), *logits = h->ptr<float>("logits");
  const bool fused1 = tc && conv1_tc_supported(c.crop_h, c.crop_w, c.c_in, c.c1);
  h->last_images = images;
  if (tc && !fused_bf16_shapes(c)) {
    // ---- layer-by-layer bf16 schedule (widths other than the script's) ------------------------------------------
    if (fused1) {      // crop + conv1 + bias + ReLU + pool1 on tcgen05, one CTA per image and 64-channel slice
      SITE("conv1_pool1_fwd", conv1_pool_fwd_tc(images, h->P(P_W1), h->P(P_B1), pool1, h->ptr<uint8_t>("arg1"), nullptr,
                                                c.batch, c.img_h, c.img_w, c.crop_h, c.crop_w, st,
                                                h->ptr<void>("pool1_bf16"), c.c1));
    } else {
      SITE("crop_cast", crop_cast_f32(images, x, c.batch, c.img_h, c.img_w, c.crop_h, c.crop_w, c.c_in, st));
      SITE("conv1_fwd", conv5x5_fwd_f32(x, h->P(P_W1), h->P(P_B1), conv1, c.batch, c.crop_h, c.crop_w, c.c_in, c.c1, 1, st));
      SITE("pool1_fwd", maxpool_fwd_f32(conv1, pool1, h->ptr<uint8_t>("arg1"), c.batch, c.crop_h, c.crop_w, c.c1, st));
    }
    SITE("weights_prep", pack_conv_weights(h->P(P_W2), h->ptr<void>("w2_packed"), c.c1, c.c2, st));
    SITE("conv2_fwd", conv5x5_fwd_tc(pool1, h->ptr<void>("w2_packed"), h->P(P_B2), conv2, c.batch, h->h1, h->w1, c.c1, c.c2, 1, st,
                                     fused1 ? h->ptr<void>("pool1_bf16") : nullptr));
    SITE("pool2_fwd", maxpool_fwd_f32(conv2, pool2, h->ptr<uint8_t>("arg2"), c.batch, h->h1, h->w1, c.c2, st));
    // FC1 (M = batch is one tile row, K = 9216 in the wide model): split-K over the SMs + a bias / ReLU pass
    SITE("full1_fwd", fc_fwd_tc_splitk(pool2, h->P(P_FW1), h->ptr<float>("fc1_part"), c.batch, h->flat, c.f1, st));
    g_pdl_next = !h->profiling;
    TRY(splitk_finish_f32(h->ptr<float>("fc1_part"), fc_fwd_tc_splits(c.batch, h->flat, c.f1), h->P(P_FB1), full1, c.batch,
                          c.f1, 1, st));
    SITE("full2_fwd", fc_fwd_tc(full1, h->P(P_FW2), h->P(P_FB2), full2, c.batch, c.f1, c.f2, 1, st));
    SITE("full3_fwd", fc_fwd_f32(full2, h->P(P_FW3), h->P(P_FB3), logits, c.batch, c.f2, c.n_cls, 1, st));
    SITE("head", head_f32(logits, labels, h->ptr<float>("loss"), h->ptr<float>("dlogits"), h->ptr<int32_t>("pred"),
                          h->ptr<int32_t>("n_correct"), h->ptr<int32_t>("n_bad"), c.batch, c.n_cls, st));
    h->forward_done = true;
    return B200_OK;
  }
  // bf16 mode: the weight images the tensor-core kernels stream (conv2 filter, W2) depend only on the
  // parameters, so they are rebuilt on a side stream while conv1 runs (every step: the caller or the
  // last update may have changed the parameters).
  const bool fork = tc && !h->profiling;
  const bool fcimg = tc && fc_images_supported(c.batch, h->flat, c.f1);
  // With the fused conv1 kernel the three image kernels go BEHIND conv1 on the same stream as programmatic
  // dependents that never wait: conv1 releases its dependents at its first instruction, so they run beside it on
  // the SMs it leaves free, and conv2 -- an ordinary launch after them -- waits for all four.  No fork / join: the
  // event join in front of conv2 cost ~1.5 us more than a plain kernel-to-kernel dependency.
  const bool prep_behind_conv1 = fork && fused1;
  if (fork) TRY(ensure_side_streams(h));
  if (tc && !prep_behind_conv1) {
    if (fork) {
      TRY(ensure_side_streams(h));
      B200_CUDA_TRY(cudaEventRecord(h->ev_fork[0], st));
      B200_CUDA_TRY(cudaStreamWaitEvent(h->side[0], h->ev_fork[0], 0));
    }
    cudaStream_t st = fork ? h->side[0] : (cudaStream_t)stream;
    SITE("weights_prep", pack_conv_weights(h->P(P_W2), h->ptr<void>("w2_packed"), c.c1, c.c2, st));
    TRY(fc_tail_prepare_w2(h->P(P_FW2), h->ptr<void>("tail_scratch"), c.batch, c.f1, c.f2, c.n_cls, st));
    if (fcimg) TRY(pack_fc_weights(h->P(P_FW1), h->ptr<void>("w1_img"), h->flat, c.f1, st));
    if (fork) B200_CUDA_TRY(cudaEventRecord(h->ev_join[0], st));
  }
  if (fused1) {
    SITE("conv1_pool1_fwd", conv1_pool_fwd_tc(images, h->P(P_W1), h->P(P_B1), pool1, h->ptr<uint8_t>("arg1"), nullptr,
                                              c.batch, c.img_h, c.img_w, c.crop_h, c.crop_w, st, h->ptr<void>("pool1_bf16")));
  } else {
    SITE("crop_cast", crop_cast_f32(images, x, c.batch, c.img_h, c.img_w, c.crop_h, c.crop_w, c.c_in, st));
    SITE("conv1_fwd", conv5x5_fwd_f32(x, h->P(P_W1), h->P(P_B1), conv1, c.batch, c.crop_h, c.crop_w, c.c_in, c.c1, 1, st));
    SITE("pool1_fwd", maxpool_fwd_f32(conv1, pool1, h->ptr<uint8_t>("arg1"), c.batch, c.crop_h, c.crop_w, c.c1, st));
  }
  if (prep_behind_conv1) {
    g_pdl_next = true;
    TRY(pack_conv_weights(h->P(P_W2), h->ptr<void>("w2_packed"), c.c1, c.c2, st));
    g_pdl_next = true;
    TRY(fc_tail_prepare_w2(h->P(P_FW2), h->ptr<void>("tail_scratch"), c.batch, c.f1, c.f2, c.n_cls, st));
    if (fcimg) {
      g_pdl_next = true;
      TRY(pack_fc_weights(h->P(P_FW1), h->ptr<void>("w1_img"), h->flat, c.f1, st));
    }
    g_pdl_next = false;
  }
  if (tc) {
    if (fork && !prep_behind_conv1) B200_CUDA_TRY(cudaStreamWaitEvent(st, h->ev_join[0], 0));
    // conv2 + bias + ReLU + pool2 in one kernel; the un-pooled activations are not needed again (the
    // backward gathers through pool2 / arg2)
    SITE("conv2_pool2_fwd", conv5x5_pool_fwd_tc(pool1, h->ptr<void>("w2_packed"), h->P(P_B2), nullptr, pool2,
                                                h->ptr<uint8_t>("arg2"), c.batch, h->h1, h->w1, c.c1, c.c2, st,
                                                fused1 ? h->ptr<void>("pool1_bf16") : nullptr,
                                                fcimg ? h->ptr<void>("pool2_img") : nullptr));
  }
  else
    SITE("conv2_fwd", conv5x5_fwd_f32(pool1, h->P(P_W2), h->P(P_B2), conv2, c.batch, h->h1, h->w1, c.c1, c.c2, 1, st));
  if (!tc) SITE("pool2_fwd", maxpool_fwd_f32(conv2, pool2, h->ptr<uint8_t>("arg2"), c.batch, h->h1, h->w1, c.c2, st));
  if (tc) {
    // FC1 as a split-K tensor-core product; the tail kernels add the partials, bias and ReLU, run
    // FC2 / FC3 / loss / accuracy and already produce every gradient above FC1 (dfull1, dW2, db2, dW3,
    // db3, db1) -- they depend on nothing else, so b200_backward starts at FC1's weight gradient.
    float* part = h->ptr<float>("fc1_part");
    if (fcimg)
      PSITE("full1_fwd", fc_fwd_tc_splitk_img(h->ptr<void>("pool2_img"), h->ptr<void>("w1_img"), part, c.batch, h->flat,
                                              c.f1, st));
    else
      PSITE("full1_fwd", fc_fwd_tc_splitk(pool2, h->P(P_FW1), part, c.batch, h->flat, c.f1, st));
    SITE("fc_tail", fc_tail(part, fc_fwd_tc_splits(c.batch, h->flat, c.f1), h->P(P_FB1), full1, h->P(P_FW2),
                            h->P(P_FB2), h->P(P_FW3), h->P(P_FB3), labels, full2, logits, h->ptr<float>("dlogits"),
                            h->ptr<float>("dfull2"), h->ptr<float>("dfull1"), h->ptr<int32_t>("pred"),
                            h->ptr<float>("loss"), h->ptr<int32_t>("n_correct"), h->ptr<int32_t>("n_bad"),
                            h->G(P_FW2), h->G(P_FB2), h->G(P_FW3), h->G(P_FB3), h->G(P_FB1),
                            h->ptr<void>("tail_scratch"), nullptr, 0, c.batch, c.f1, c.f2,
                            c.n_cls, st, true, fork ? h->side[1] : nullptr, h->ev_fork[1], !h->profiling,
                            fcimg ? h->ptr<void>("dfull1_img") : nullptr));
    // (side stream 1, not 0: the backward puts FC1's dW on stream 0 and wants it to start with FC1's dX, not behind
    //  this kernel -- 36 + 108 CTAs fit the machine together)
    if (fork && !defer_tail) {
      B200_CUDA_TRY(cudaEventRecord(h->ev_join[1], h->side[1]));
      B200_CUDA_TRY(cudaStreamWaitEvent(st, h->ev_join[1], 0));
    }
  } else {
    PSITE("full1_fwd", fc_fwd_f32(pool2, h->P(P_FW1), h->P(P_FB1), full1, c.batch, h->flat, c.f1, 1, st));
    SITE("full2_fwd", fc_fwd_f32(full1, h->P(P_FW2), h->P(P_FB2), full2, c.batch, c.f1, c.f2, 1, st));
    SITE("full3_fwd", fc_fwd_f32(full2, h->P(P_FW3), h->P(P_FB3), logits, c.batch, c.f2, c.n_cls, 1, st));
    SITE("head", head_f32(logits, labels, h->ptr<float>("loss"), h->ptr<float>("dlogits"), h->ptr<int32_t>("pred"),
                          h->ptr<int32_t>("n_correct"), h->ptr<int32_t>("n_bad"), c.batch, c.n_cls, st));
  }
  h->forward_done = true;
  return B200_OK;
}

int b200_forward(b200_handle h, const uint8_t* images, const int32_t* labels, void* stream) {
  return forward_impl(h, images, labels, stream, false);
}

int b200_forward_backward(b200_handle h, const uint8_t* images, const int32_t* labels, void* stream) {
  TRY(forward_impl(h, images, labels, stream, true));
  return b200_backward(h, stream);
}

// the peer tables of b200_p2p_sgd_apply, for the variant of the backward pass that overlaps the exchange
struct P2PCtx {
  const void* const* peer_grads; void* const* peer_params; void* const* peer_pads;
  const void* mc_grads; void* mc_params; uint32_t* state;
  int rank, world;
  float lr, scale;
};

static int backward_impl(b200_handle h, void* stream, const P2PCtx* px, bool defer_conv1_reduce = false) {
  B200_READY(h);
  if (!h->forward_done) {
    set_error("backward: call b200_forward first");
    return B200_E_STATE;
  }
  h->forward_done = false;      // one backward per forward (buffers cleared by the forward are consumed here)
  cudaStream_t st = (cudaStream_t)stream;
  const b200_config& c = h->cfg;
  const bool tc = c.mode == B200_MODE_BF16;
  float* scratch = h->ptr<float>("scratch");
  float *dlogits = h->ptr<float>("dlogits"), *dfull2 = h->ptr<float>("dfull2"), *dfull1 = h->ptr<float>("dfull1");
  float *dpool2 = h->ptr<float>("dpool2"), *dconv2 = h->ptr<float>("dconv2");
  float *dpool1 = h->ptr<float>("dpool1"), *dconv1 = h->ptr<float>("dconv1");
  if (tc && !fused_bf16_shapes(c)) {
    // ---- layer-by-layer bf16 schedule (see fused_bf16_shapes) ----------------------------------------------------
    B200_REQUIRE(!px, "the overlapped exchange needs the script's own layer widths");
    SITE("full3_bwd", fc_bwd_f32(h->ptr<float>("full2"), h->P(P_FW3), dlogits, dfull2, h->G(P_FW3), h->G(P_FB3),
                                 scratch, c.batch, c.f2, c.n_cls, 1, st));
    SITE("full2_bwd", fc_bwd_tc(h->ptr<float>("full1"), h->P(P_FW2), dfull2, dfull1, h->G(P_FW2), c.batch, c.f1, c.f2, 1, st));
    TRY(colsum_f32(dfull2, h->G(P_FB2), scratch, c.batch, c.f2, st));
    SITE("full1_bwd", fc_bwd_tc(h->ptr<float>("pool2"), h->P(P_FW1), dfull1, dpool2, h->G(P_FW1), c.batch, h->flat, c.f1, 0, st));
    TRY(colsum_f32(dfull1, h->G(P_FB1), scratch, c.batch, c.f1, st));
    const bool f1tc = conv1_tc_supported(c.crop_h, c.crop_w, c.c_in, c.c1);      // then pool1 also exists in bf16
    SITE("pool2_bwd", maxpool_bwd_f32(dpool2, h->ptr<uint8_t>("arg2"), h->ptr<float>("conv2"), dconv2, c.batch, h->h1,
                                      h->w1, c.c2, 1, st, nullptr, f1tc ? h->ptr<void>("dconv2_bf16") : nullptr));
    SITE("conv2_wgrad", conv5x5_wgrad_tc(h->ptr<float>("pool1"), dconv2, h->G(P_W2), scratch, c.batch, h->h1, h->w1, c.c1,
                                         c.c2, st, f1tc ? h->ptr<void>("pool1_bf16") : nullptr,
                                         f1tc ? h->ptr<void>("dconv2_bf16") : nullptr));
    TRY(colsum_f32(dconv2, h->G(P_B2), scratch, c.batch * h->h1 * h->w1, c.c2, st));
    SITE("conv2_dgrad", conv5x5_dgrad_tc(dconv2, h->ptr<void>("w2_packed"), dpool1, c.batch, h->h1, h->w1, c.c1, c.c2, st,
                                         f1tc ? h->ptr<void>("dconv2_bf16") : nullptr));
    if (conv1_tc_supported(c.crop_h, c.crop_w, c.c_in, c.c1)) {
      SITE("conv1_wgrad", conv1_wgrad_tc(h->last_images, dpool1, h->ptr<float>("pool1"), h->ptr<uint8_t>("arg1"),
                                         h->G(P_W1), h->G(P_B1), scratch, c.batch, c.img_h, c.img_w, c.crop_h,
                                         c.crop_w, st, false, c.c1));
    } else {
      SITE("pool1_bwd", maxpool_bwd_f32(dpool1, h->ptr<uint8_t>("arg1"), h->ptr<float>("conv1"), dconv1, c.batch,
                                        c.crop_h, c.crop_w, c.c1, 1, st));
      SITE("conv1_wgrad", conv5x5_wgrad_f32(h->ptr<float>("x"), dconv1, h->G(P_W1), h->G(P_B1), scratch, c.batch,
                                            c.crop_h, c.crop_w, c.c_in, c.c1, st));
    }
    return B200_OK;
  }
  if (tc) {
    // Dependency graph of the bf16 backward (dfull1 and everything above it came from fc_tail):
    //   main : FC1 dX -> pool2 bwd -> conv2 dgrad -> conv1 wgrad
    //   side0: FC1 dW                      (needs dfull1 only)
    //   side1: conv2 wgrad                 (needs dconv2 only)
    // The side branches are forked with events so they overlap the main chain, eagerly and inside a
    // captured CUDA graph alike.  While profiling per site everything stays on one stream.
    const bool fork = !h->profiling;
    if (fork) TRY(ensure_side_streams(h));
    cudaStream_t s0 = fork ? h->side[0] : st, s1 = fork ? h->side[1] : st;
    const bool fused1 = conv1_tc_supported(c.crop_h, c.crop_w, c.c_in, c.c1);
    const bool fcimg = fc_images_supported(c.batch, h->flat, c.f1);
    if (fork) {
      B200_CUDA_TRY(cudaEventRecord(h->ev_fork[0], st));
      B200_CUDA_TRY(cudaStreamWaitEvent(s0, h->ev_fork[0], 0));
    }
    {
      cudaStream_t st = s0;   // SITE() launches on `st`
      if (fcimg)
        SITE("full1_dw", fc_dw_tc_img(h->ptr<void>("pool2_img"), h->ptr<void>("dfull1_img"), h->G(P_FW1), c.batch, h->flat,
                                      c.f1, st));
      else
        SITE("full1_dw", fc_dw_tc(h->ptr<float>("pool2"), dfull1, h->G(P_FW1), c.batch, h->flat, c.f1, st));
    }
    if (fcimg)
      PSITE("full1_dx", fc_dx_tc_img(h->ptr<void>("dfull1_img"), h->ptr<void>("w1_img"), dpool2, c.batch, h->flat, c.f1, st));
    else
      PSITE("full1_dx", fc_dx_tc(h->ptr<float>("pool2"), h->P(P_FW1), dfull1, dpool2, c.batch, h->flat, c.f1, 0, st));
    if (px && fork) {
      // Channel 0 of the exchange: the FC gradients (everything from full_weight_1 on, 90 % of the buffer) are
      // complete once FC1's dW (stream 0) and the tail-gradient kernel (stream 1) are; their all-reduce + update
      // runs on stream 0 beside conv2's dgrad and the conv filter gradients, capped at 16 CTAs so that it lives on
      // the SMs those 128-CTA kernels leave free.  It may not start before FC1's dX has read W1 (peers write the
      // updated W1 into this GPU's buffer).
      B200_CUDA_TRY(cudaEventRecord(h->ev_aux[0], st));
      B200_CUDA_TRY(cudaEventRecord(h->ev_aux[1], h->side[1]));
      B200_CUDA_TRY(cudaStreamWaitEvent(s0, h->ev_aux[0], 0));
      B200_CUDA_TRY(cudaStreamWaitEvent(s0, h->ev_aux[1], 0));
      const int64_t off = h->p_off[P_FW1];
      TRY(p2p_allreduce_sgd(px->peer_grads, px->peer_params, px->peer_pads, px->mc_grads, px->mc_params, px->state,
                            px->rank, px->world, h->p_total - off, px->lr, px->scale, s0, off, 0, 16));
    }
    // pool2's backward is not a kernel of its own: conv2's dgrad gathers MaxPoolGrad * ReluGrad from (dpool2,
    // pool2, arg2) while it builds its strip, adds the bias gradient (atomics into the buffer the tail kernel
    // cleared) and leaves dconv2 behind for the filter gradient, which then overlaps conv1's gradient.
    int n_db2 = 0;
    SITE("conv2_dgrad", conv5x5_dgrad_pool_tc(dpool2, h->ptr<float>("pool2"), h->ptr<uint8_t>("arg2"),
                                              h->ptr<void>("w2_packed"), dpool1, nullptr,
                                              fused1 ? nullptr : dconv2, c.batch, h->h1, h->w1, c.c1, c.c2, st,
                                              fused1 ? h->ptr<void>("dconv2_bf16") : nullptr, h->ptr<float>("db2_part"),
                                              &n_db2));
    if (fork) {
      B200_CUDA_TRY(cudaEventRecord(h->ev_fork[1], st));
      B200_CUDA_TRY(cudaStreamWaitEvent(s1, h->ev_fork[1], 0));
    }
    {
      cudaStream_t st = s1;
      // (x and dY as the bf16 copies conv1's forward and conv2's dgrad left behind: the five filter-row CTAs of an
      //  image group each re-read both strips, so halving the bytes is what this L2-bound kernel needs)
      PSITE("conv2_wgrad", conv5x5_wgrad_tc(h->ptr<float>("pool1"), dconv2, h->G(P_W2), h->ptr<float>("scratch2"),
                                           c.batch, h->h1, h->w1, c.c1, c.c2, st,
                                           fused1 ? h->ptr<void>("pool1_bf16") : nullptr,
                                           fused1 ? h->ptr<void>("dconv2_bf16") : nullptr,
                                           // conv2's bias gradient: the dgrad CTAs' partial sums, in a fixed order, in the
                                           // launch that sums the filter-gradient partials
                                           h->ptr<float>("db2_part"), h->G(P_B2), n_db2));
    }
    if (fused1) {
      // conv1's filter / bias gradient leaves one partial per CTA in `scratch`; they are summed in a fixed order by a
      // small kernel -- or, when the update follows at once (b200_train_step), inside the update kernel
      SITE("conv1_wgrad", conv1_wgrad_tc(h->last_images, dpool1, h->ptr<float>("pool1"), h->ptr<uint8_t>("arg1"),
                                         h->G(P_W1), h->G(P_B1), scratch, c.batch, c.img_h, c.img_w, c.crop_h,
                                         c.crop_w, st, (defer_conv1_reduce || px) && !h->profiling));
    } else {
      SITE("pool1_bwd", maxpool_bwd_f32(dpool1, h->ptr<uint8_t>("arg1"), h->ptr<float>("conv1"), dconv1, c.batch,
                                        c.crop_h, c.crop_w, c.c1, 1, st));
      SITE("conv1_wgrad", conv5x5_wgrad_f32(h->ptr<float>("x"), dconv1, h->G(P_W1), h->G(P_B1), scratch, c.batch,
                                            c.crop_h, c.crop_w, c.c_in, c.c1, st));
    }
    if (fork) {
      for (int i = 0; i < 2; ++i) {
        B200_CUDA_TRY(cudaEventRecord(h->ev_join[i], h->side[i]));
        B200_CUDA_TRY(cudaStreamWaitEvent(st, h->ev_join[i], 0));
      }
    }
    if (px) {      // channel 1: the conv gradients (10 % of the buffer); both channels here when nothing was forked
      const int64_t off = h->p_off[P_FW1];
      if (!fork)
        TRY(p2p_allreduce_sgd(px->peer_grads, px->peer_params, px->peer_pads, px->mc_grads, px->mc_params, px->state,
                              px->rank, px->world, h->p_total - off, px->lr, px->scale, st, off, 0, 0));
      // (conv1's gradient partials are summed by the exchange kernel's first CTAs: no reduction launch in between)
      const bool red = fused1 && !h->profiling && h->p_off[P_W1] == 0 && h->p_off[P_B1] == 4800;
      if (h->inbox_local && h->inbox_floats >= 2 * px->world * off && fork)
        // one-shot: push the conv gradients into every worker's inbox, one flag wait, local rank-ordered sum + update
        SITE("p2p_conv_update", p2p_push_sum_sgd(h->grads, h->params, h->inbox_local, h->inbox_mc, h->inbox_peers,
                                                 px->peer_pads, px->state, px->rank, px->world, off, px->lr, px->scale, st,
                                                 1, red ? scratch : nullptr, red ? conv1_wgrad_tc_partials(c.batch) : 0,
                                                 red ? 4864 : 0));
      else
      SITE("p2p_conv_update", p2p_allreduce_sgd(px->peer_grads, px->peer_params, px->peer_pads, px->mc_grads,
                                                px->mc_params, px->state, px->rank, px->world, off, px->lr, px->scale,
                                                st, 0, 1, 0, red ? scratch : nullptr,
                                                red ? conv1_wgrad_tc_partials(c.batch) : 0, red ? 4864 : 0));
    }
    return B200_OK;
  }
  B200_REQUIRE(!px, "the overlapped exchange needs the bf16 mode");
  SITE("full3_bwd", fc_bwd_f32(h->ptr<float>("full2"), h->P(P_FW3), dlogits, dfull2, h->G(P_FW3), h->G(P_FB3),
                               scratch, c.batch, c.f2, c.n_cls, 1, st));
  SITE("full2_bwd", fc_bwd_f32(h->ptr<float>("full1"), h->P(P_FW2), dfull2, dfull1, h->G(P_FW2), h->G(P_FB2), scratch,
                               c.batch, c.f1, c.f2, 1, st));
  SITE("full1_bwd", fc_bwd_f32(h->ptr<float>("pool2"), h->P(P_FW1), dfull1, dpool2, h->G(P_FW1), h->G(P_FB1), scratch,
                               c.batch, h->flat, c.f1, 0, st));
  SITE("pool2_bwd", maxpool_bwd_f32(dpool2, h->ptr<uint8_t>("arg2"), h->ptr<float>("conv2"), dconv2, c.batch, h->h1,
                                    h->w1, c.c2, 1, st));
  PSITE("conv2_wgrad", conv5x5_wgrad_f32(h->ptr<float>("pool1"), dconv2, h->G(P_W2), h->G(P_B2), scratch, c.batch,
                                        h->h1, h->w1, c.c1, c.c2, st));
  SITE("conv2_dgrad_prep", conv5x5_rot180_f32(h->P(P_W2), h->ptr<float>("wrot"), c.c1, c.c2, st));
  SITE("conv2_dgrad", conv5x5_fwd_f32(dconv2, h->ptr<float>("wrot"), nullptr, dpool1, c.batch, h->h1, h->w1, c.c2,
                                      c.c1, 0, st));
  SITE("pool1_bwd", maxpool_bwd_f32(dpool1, h->ptr<uint8_t>("arg1"), h->ptr<float>("conv1"), dconv1, c.batch,
                                    c.crop_h, c.crop_w, c.c1, 1, st));
  SITE("conv1_wgrad", conv5x5_wgrad_f32(h->ptr<float>("x"), dconv1, h->G(P_W1), h->G(P_B1), scratch, c.batch,
                                        c.crop_h, c.crop_w, c.c_in, c.c1, st));
  return B200_OK;
}

int b200_backward(b200_handle h, void* stream) { return backward_impl(h, stream, nullptr); }

int b200_train_step_p2p(b200_handle h, const uint8_t* images, const int32_t* labels, const void* const* peer_grads,
                        void* const* peer_params, void* const* peer_pads, const void* mc_grads, void* mc_params,
                        uint32_t* state, int32_t rank, int32_t world, float lr, float grad_scale, void* stream) {
  B200_READY(h);
  B200_REQUIRE(peer_grads && peer_params && peer_pads && state && rank >= 0 && rank < world && world <= 8,
               "train_step_p2p: bad argument");
  B200_REQUIRE(peer_grads[rank] == h->grads && peer_params[rank] == h->params,
               "train_step_p2p: entry `rank` of the peer tables must be the buffers bound to this handle");
  B200_REQUIRE(h->p_off[P_FW1] % (4 * world) == 0 && (h->p_total - h->p_off[P_FW1]) % (4 * world) == 0,
               "train_step_p2p: the parameter ranges do not split into 16-byte shards for %d workers", world);
  TRY(forward_impl(h, images, labels, stream, true));
  const P2PCtx px{peer_grads, peer_params, peer_pads, mc_grads, mc_params, state, rank, world, lr, grad_scale};
  return backward_impl(h, stream, &px);
}

int b200_p2p_set_inbox(b200_handle h, void* inbox_local, void* inbox_mc, void* const* inbox_peers, int32_t world,
                       int64_t inbox_floats) {
  B200_REQUIRE(h, "null handle");
  if (!inbox_local) { h->inbox_local = nullptr; h->inbox_floats = 0; return B200_OK; }
  B200_REQUIRE(inbox_peers && world >= 2 && world <= 8 && inbox_floats > 0 && ((uintptr_t)inbox_local & 15) == 0,
               "p2p_set_inbox: bad argument");
  for (int p = 0; p < world; ++p) B200_REQUIRE(inbox_peers[p], "p2p_set_inbox: null peer pointer for rank %d", p);
  for (int p = 0; p < world; ++p) h->inbox_peers[p] = inbox_peers[p];
  h->inbox_local = (float*)inbox_local; h->inbox_mc = inbox_mc; h->inbox_floats = inbox_floats;
  return B200_OK;
}

int b200_train_step_p2p_single(b200_handle h, const uint8_t* images, const int32_t* labels, const void* const* peer_grads,
                               void* const* peer_params, void* const* peer_pads, const void* mc_grads, void* mc_params,
                               uint32_t* state, int32_t rank, int32_t world, float lr, float grad_scale, void* stream) {
  B200_READY(h);
  B200_REQUIRE(peer_grads && peer_params && peer_pads && state && rank >= 0 && rank < world && world <= 8,
               "train_step_p2p_single: bad argument");
  B200_REQUIRE(peer_grads[rank] == h->grads && peer_params[rank] == h->params,
               "train_step_p2p_single: entry `rank` of the peer tables must be the buffers bound to this handle");
  const b200_config& c = h->cfg;
  const bool red = c.mode == B200_MODE_BF16 && fused_bf16_shapes(c) && !h->profiling &&
                   conv1_tc_supported(c.crop_h, c.crop_w, c.c_in, c.c1) && h->p_off[P_W1] == 0 && h->p_off[P_B1] == 4800;
  TRY(forward_impl(h, images, labels, stream, true));
  TRY(backward_impl(h, stream, nullptr, red));
  cudaStream_t st = (cudaStream_t)stream;
  SITE("p2p_sgd_apply", p2p_allreduce_sgd(peer_grads, peer_params, peer_pads, mc_grads, mc_params, state, rank, world,
                                          h->p_total, lr, grad_scale, st, 0, 0, 0, red ? h->ptr<float>("scratch") : nullptr,
                                          red ? conv1_wgrad_tc_partials(c.batch) : 0, red ? 4864 : 0));
  return B200_OK;
}

int b200_sgd_apply(b200_handle h, float lr, float grad_scale, void* stream) {
  B200_READY(h);
  cudaStream_t st = (cudaStream_t)stream;
  PSITE("sgd_apply", sgd_apply_f32(h->params, h->grads, h->p_total, lr, grad_scale, st));
  return B200_OK;
}

int b200_p2p_sgd_apply(b200_handle h, const void* const* peer_grads, void* const* peer_params, void* const* peer_pads,
                       const void* mc_grads, void* mc_params, uint32_t* state, int32_t rank, int32_t world, float lr,
                       float grad_scale, void* stream) {
  B200_READY(h);
  B200_REQUIRE(peer_grads && peer_params && rank >= 0 && rank < world && world <= 8, "p2p_sgd_apply: bad argument");
  B200_REQUIRE(peer_grads[rank] == h->grads && peer_params[rank] == h->params,
               "p2p_sgd_apply: entry `rank` of the peer tables must be the buffers bound to this handle");
  cudaStream_t st = (cudaStream_t)stream;
  SITE("p2p_sgd_apply", p2p_allreduce_sgd(peer_grads, peer_params, peer_pads, mc_grads, mc_params, state, rank, world,
                                          h->p_total, lr, grad_scale, st));
  return B200_OK;
}

int b200_train_step(b200_handle h, const uint8_t* images, const int32_t* labels, float lr, void* stream) {
  B200_READY(h);
  const b200_config& c = h->cfg;
  const bool fuse = c.mode == B200_MODE_BF16 && fused_bf16_shapes(c) && !h->profiling &&
                    conv1_tc_supported(c.crop_h, c.crop_w, c.c_in, c.c1) &&
                    h->p_off[P_W1] == 0 && h->p_off[P_B1] == 4800;
  if (!fuse) {
    TRY(b200_forward_backward(h, images, labels, stream));
    return b200_sgd_apply(h, lr, 1.f, stream);
  }
  // the fixed-order sum of conv1's gradient partials rides in the update kernel (no reduction launch at the very end
  // of the step); the gradient buffer is complete when the step is
  TRY(forward_impl(h, images, labels, stream, true));
  TRY(backward_impl(h, stream, nullptr, true));
  cudaStream_t st = (cudaStream_t)stream;
  g_pdl_next = true;
  return sgd_apply_reduce_f32(h->params, h->grads, h->p_total, lr, 1.f, h->ptr<float>("scratch"),
                              conv1_wgrad_tc_partials(c.batch), 4864, st);
}

int b200_train_step_host(b200_handle h, const uint8_t* images_host, const int32_t* labels_host, float lr,
                         float* loss_host, void* stream) {
  B200_READY(h);
  B200_REQUIRE(images_host && labels_host && loss_host, "train_step_host: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  const b200_config& c = h->cfg;
  uint8_t* si = h->ptr<uint8_t>("stage_images");
  int32_t* sl = h->ptr<int32_t>("stage_labels");
  B200_CUDA_TRY(cudaMemcpyAsync(si, images_host, (size_t)c.batch * c.c_in * c.img_h * c.img_w,
                                cudaMemcpyHostToDevice, st));
  B200_CUDA_TRY(cudaMemcpyAsync(sl, labels_host, (size_t)c.batch * 4, cudaMemcpyHostToDevice, st));
  TRY(b200_train_step(h, si, sl, lr, stream));
  B200_CUDA_TRY(cudaMemcpyAsync(loss_host, h->ptr<float>("loss"), 4, cudaMemcpyDeviceToHost, st));
  return B200_OK;
}

// ---- host batches with prefetch: two staging slots, the copy of the NEXT batch overlaps the step on the current one ----
static int stage_ptrs(b200_model* h, int slot, uint8_t** si, int32_t** sl) {
  B200_REQUIRE(slot == 0 || slot == 1, "staging slot must be 0 or 1");
  *si = h->ptr<uint8_t>(slot ? "stage_images1" : "stage_images");
  *sl = h->ptr<int32_t>(slot ? "stage_labels1" : "stage_labels");
  return B200_OK;
}
int b200_stage_host_batch(b200_handle h, int32_t slot, const uint8_t* images_host, const int32_t* labels_host, void* stream) {
  B200_READY(h);
  B200_REQUIRE(images_host && labels_host, "stage_host_batch: null argument");
  uint8_t* si; int32_t* sl;
  TRY(stage_ptrs(h, slot, &si, &sl));
  const b200_config& c = h->cfg;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA_TRY(cudaMemcpyAsync(si, images_host, (size_t)c.batch * c.c_in * c.img_h * c.img_w, cudaMemcpyHostToDevice, st));
  B200_CUDA_TRY(cudaMemcpyAsync(sl, labels_host, (size_t)c.batch * 4, cudaMemcpyHostToDevice, st));
  return B200_OK;
}
int b200_train_step_host_prefetch(b200_handle h, int32_t slot, const uint8_t* next_images_host,
                                  const int32_t* next_labels_host, float lr, float* loss_host, void* stream) {
  B200_READY(h);
  B200_REQUIRE(loss_host, "train_step_host_prefetch: null loss pointer");
  B200_REQUIRE((next_images_host == nullptr) == (next_labels_host == nullptr),
               "train_step_host_prefetch: give both next-batch pointers or none");
  uint8_t *si, *ni; int32_t *sl, *nl;
  TRY(stage_ptrs(h, slot, &si, &sl));
  TRY(stage_ptrs(h, slot ^ 1, &ni, &nl));
  const b200_config& c = h->cfg;
  cudaStream_t st = (cudaStream_t)stream;
  if (!h->copy_stream) {      // created by the first (eager) call, like the side streams
    B200_CUDA_TRY(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) B200_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_copy[i], cudaEventDisableTiming));
  }
  if (next_images_host) {     // fork: the next batch travels while this one is trained on
    B200_CUDA_TRY(cudaEventRecord(h->ev_copy[0], st));
    B200_CUDA_TRY(cudaStreamWaitEvent(h->copy_stream, h->ev_copy[0], 0));
    B200_CUDA_TRY(cudaMemcpyAsync(ni, next_images_host, (size_t)c.batch * c.c_in * c.img_h * c.img_w,
                                  cudaMemcpyHostToDevice, h->copy_stream));
    B200_CUDA_TRY(cudaMemcpyAsync(nl, next_labels_host, (size_t)c.batch * 4, cudaMemcpyHostToDevice, h->copy_stream));
    B200_CUDA_TRY(cudaEventRecord(h->ev_copy[1], h->copy_stream));
  }
  TRY(b200_train_step(h, si, sl, lr, stream));
  B200_CUDA_TRY(cudaMemcpyAsync(loss_host, h->ptr<float>("loss"), 4, cudaMemcpyDeviceToHost, st));
  if (next_images_host) B200_CUDA_TRY(cudaStreamWaitEvent(st, h->ev_copy[1], 0));      // join
  return B200_OK;
}

int b200_profile_step(b200_handle h, const uint8_t* images, const int32_t* labels, float lr, void* stream,
                      float* ms_per_site, int32_t max_sites, int32_t* n_sites) {
  B200_READY(h);
  B200_REQUIRE(ms_per_site && n_sites && max_sites > 0, "profile_step: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  h->profiling = true;
  h->site_cursor = 0;
  int rc = b200_train_step(h, images, labels, lr, stream);
  if (!rc) rc = site_mark(h, "end", st);
  h->profiling = false;
  if (rc) return rc;
  B200_CUDA_TRY(cudaStreamSynchronize(st));
  int n = h->site_cursor - 1;
  B200_REQUIRE(n <= max_sites, "profile_step: %d sites, room for %d", n, max_sites);
  for (int i = 0; i < n; ++i)
    B200_CUDA_TRY(cudaEventElapsedTime(&ms_per_site[i], h->site_events[i], h->site_events[i + 1]));
  *n_sites = n;
  return B200_OK;
}
const char* b200_site_name(b200_handle h, int32_t i) {
  if (!h || i < 0 || i >= (int)h->site_names.size()) return "";
  return h->site_names[i].c_str();
}

int b200_result_ptrs(b200_handle h, float** logits, float** loss, int32_t** n_correct, int32_t** pred) {
  B200_REQUIRE(h && h->ws, "result_ptrs: bind a workspace first");
  if (logits) *logits = h->ptr<float>("logits");
  if (loss) *loss = h->ptr<float>("loss");
  if (n_correct) *n_correct = h->ptr<int32_t>("n_correct");
  if (pred) *pred = h->ptr<int32_t>("pred");
  return B200_OK;
}

int b200_tensor_ptr(b200_handle h, const char* name, void** ptr, int64_t* nbytes, int32_t* esize) {
  B200_REQUIRE(h && h->ws && name, "tensor_ptr: bind a workspace first");
  const Tensor* t = h->find(name);
  B200_REQUIRE(t, "tensor_ptr: no tensor named '%s'", name);
  if (ptr) *ptr = h->ws + t->offset;
  if (nbytes) *nbytes = t->elems * t->esize;
  if (esize) *esize = t->esize;
  return B200_OK;
}

int b200_launches_per_step(b200_handle h, int32_t* n) {
  B200_REQUIRE(h && n, "null argument");
  // forward 9; fc_bwd 3 x (dw, colsum x2, dx); pool bwd x2; conv wgrad 2 x (gemm, reduce, colsum x2);
  // dgrad (rot180 + gemm); sgd
  *n = 9 + 12 + 2 + 8 + 2 + 1;
  if (h->cfg.mode == B200_MODE_BF16 && !fused_bf16_shapes(h->cfg)) {
    // layer-by-layer schedule: forward 8 (10 with the SIMT conv1) | FC3 bwd 4, FC2 / FC1 bwd 2 + 2 each, pool2 bwd, conv2
    // wgrad 2 + colsum 2, dgrad, conv1 wgrad 2 (pool1 bwd + 4 with the SIMT conv1) | sgd
    const bool f1 = conv1_tc_supported(h->cfg.crop_h, h->cfg.crop_w, h->cfg.c_in, h->cfg.c1);
    *n = (f1 ? 8 : 10) + 1 + 4 + 4 + 4 + 1 + 4 + 1 + (f1 ? 2 : 5) + 1;      // (+1: FC1's split-K bias / ReLU pass)
  } else if (h->cfg.mode == B200_MODE_BF16) {
    // conv1 (3 or 1), conv2, pool2, FC1 split-K, tail x2 | FC1 bwd x2, pool2 bwd, conv2 wgrad x2, dgrad, conv1 wgrad (2 or 1+4), sgd
    const bool f1 = conv1_tc_supported(h->cfg.crop_h, h->cfg.crop_w, h->cfg.c_in, h->cfg.c1);
    *n = (f1 ? 1 : 3) + 2 + 1 + 3 + 2 + 2 + 1 + (f1 ? 1 : 5) + 1 +
         (fc_images_supported(h->cfg.batch, h->flat, h->cfg.f1) ? 1 : 0);      // + W1 image pack
  }
  return B200_OK;
}

// ---- per-op entry points --------------------------------------------------------------------
int b200_crop_cast_f32(const uint8_t* images, float* out, int32_t batch, int32_t img_h, int32_t img_w,
                       int32_t crop_h, int32_t crop_w, int32_t c_in, void* stream) {
  B200_REQUIRE(images && out, "null pointer");
  return crop_cast_f32(images, out, batch, img_h, img_w, crop_h, crop_w, c_in, (cudaStream_t)stream);
}
int b200_conv5x5_fwd_f32(const float* x, const float* w, const float* bias, float* y, int32_t batch, int32_t h,
                         int32_t wd, int32_t c_in, int32_t c_out, int32_t relu, void* stream) {
  B200_REQUIRE(x && w && y, "null pointer");
  return conv5x5_fwd_f32(x, w, bias, y, batch, h, wd, c_in, c_out, relu, (cudaStream_t)stream);
}
int b200_conv5x5_dgrad_f32(const float* dy, const float* w, float* w_scratch, float* dx, int32_t batch,
                           int32_t h, int32_t wd, int32_t c_in, int32_t c_out, void* stream) {
  B200_REQUIRE(dy && w && w_scratch && dx, "null pointer");
  TRY(conv5x5_rot180_f32(w, w_scratch, c_in, c_out, (cudaStream_t)stream));
  return conv5x5_fwd_f32(dy, w_scratch, nullptr, dx, batch, h, wd, c_out, c_in, 0, (cudaStream_t)stream);
}
int b200_conv5x5_wgrad_scratch_bytes(int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out,
                                     int64_t* nbytes) {
  B200_REQUIRE(nbytes && batch > 0 && h > 0 && wd > 0 && c_in > 0 && c_out > 0, "bad argument");
  *nbytes = conv5x5_wgrad_scratch_bytes(batch, h, wd, c_in, c_out);
  return B200_OK;
}
int b200_conv5x5_wgrad_f32(const float* x, const float* dy, float* dw, float* db, void* scratch,
                           int64_t scratch_bytes, int32_t batch, int32_t h, int32_t wd, int32_t c_in,
                           int32_t c_out, void* stream) {
  B200_REQUIRE(x && dy && dw && scratch, "null pointer");
  B200_REQUIRE(batch > 0 && h > 0 && wd > 0 && c_in > 0 && c_out > 0, "bad shape");
  B200_REQUIRE(scratch_bytes >= conv5x5_wgrad_scratch_bytes(batch, h, wd, c_in, c_out), "scratch too small");
  return conv5x5_wgrad_f32(x, dy, dw, db, (float*)scratch, batch, h, wd, c_in, c_out, (cudaStream_t)stream);
}
int b200_maxpool_fwd_f32(const float* x, float* y, uint8_t* arg, int32_t batch, int32_t h, int32_t wd, int32_t c,
                         void* stream) {
  B200_REQUIRE(x && y && arg, "null pointer");
  return maxpool_fwd_f32(x, y, arg, batch, h, wd, c, (cudaStream_t)stream);
}
int b200_maxpool_bwd_f32(const float* dy, const uint8_t* arg, const float* x, float* dx, int32_t batch, int32_t h,
                         int32_t wd, int32_t c, int32_t relu_mask, void* stream) {
  B200_REQUIRE(dy && arg && dx, "null pointer");
  return maxpool_bwd_f32(dy, arg, x, dx, batch, h, wd, c, relu_mask, (cudaStream_t)stream);
}
int b200_fc_fwd_f32(const float* x, const float* w, const float* b, float* y, int32_t batch, int32_t k, int32_t n,
                    int32_t relu, void* stream) {
  B200_REQUIRE(x && w && y, "null pointer");
  return fc_fwd_f32(x, w, b, y, batch, k, n, relu, (cudaStream_t)stream);
}
int b200_colsum_scratch_bytes(int32_t rows, int32_t cols, int64_t* nbytes) {
  B200_REQUIRE(nbytes && rows > 0 && cols > 0, "bad argument");
  *nbytes = colsum_scratch_bytes(rows, cols);
  return B200_OK;
}
int b200_fc_bwd_f32(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                    void* scratch, int64_t scratch_bytes, int32_t batch, int32_t k, int32_t n, int32_t mask_x,
                    void* stream) {
  B200_REQUIRE(x && w && dy && dw, "null pointer");
  B200_REQUIRE(batch > 0 && k > 0 && n > 0, "bad shape");
  B200_REQUIRE(!db || (scratch && scratch_bytes >= colsum_scratch_bytes(batch, n)), "scratch too small");
  return fc_bwd_f32(x, w, dy, dx, dw, db, (float*)scratch, batch, k, n, mask_x, (cudaStream_t)stream);
}
int b200_conv1_pool_fwd_bf16(const uint8_t* images, const float* w, const float* bias, float* pool, uint8_t* arg,
                             float* conv_dbg, int32_t batch, int32_t img_h, int32_t img_w, int32_t crop_h,
                             int32_t crop_w, void* stream) {
  B200_REQUIRE(images && w && bias && pool && arg, "null pointer");
  return conv1_pool_fwd_tc(images, w, bias, pool, arg, conv_dbg, batch, img_h, img_w, crop_h, crop_w,
                           (cudaStream_t)stream);
}
int b200_conv1_wgrad_bf16_scratch_bytes(int32_t batch, int64_t* nbytes) {
  B200_REQUIRE(nbytes && batch > 0, "bad argument");
  *nbytes = conv1_wgrad_tc_scratch_bytes(batch);
  return B200_OK;
}
int b200_conv1_wgrad_bf16(const uint8_t* images, const float* dpool, const float* pool, const uint8_t* arg, float* dw,
                          float* db, void* scratch, int64_t scratch_bytes, int32_t batch, int32_t img_h, int32_t img_w,
                          int32_t crop_h, int32_t crop_w, void* stream) {
  B200_REQUIRE(images && dpool && pool && arg && dw && scratch, "null pointer");
  B200_REQUIRE(batch > 0 && scratch_bytes >= conv1_wgrad_tc_scratch_bytes(batch), "scratch too small");
  return conv1_wgrad_tc(images, dpool, pool, arg, dw, db, (float*)scratch, batch, img_h, img_w, crop_h, crop_w,
                        (cudaStream_t)stream);
}
int b200_conv1_pool_fwd_bf16_c(const uint8_t* images, const float* w, const float* bias, float* pool, uint8_t* arg,
                               float* conv_dbg, int32_t batch, int32_t img_h, int32_t img_w, int32_t crop_h,
                               int32_t crop_w, int32_t c1, void* stream) {
  B200_REQUIRE(images && w && bias && pool && arg, "null pointer");
  return conv1_pool_fwd_tc(images, w, bias, pool, arg, conv_dbg, batch, img_h, img_w, crop_h, crop_w,
                           (cudaStream_t)stream, nullptr, c1);
}
int b200_conv1_wgrad_bf16_c(const uint8_t* images, const float* dpool, const float* pool, const uint8_t* arg, float* dw,
                            float* db, void* scratch, int64_t scratch_bytes, int32_t batch, int32_t img_h, int32_t img_w,
                            int32_t crop_h, int32_t crop_w, int32_t c1, void* stream) {
  B200_REQUIRE(images && dpool && pool && arg && dw && scratch, "null pointer");
  B200_REQUIRE(batch > 0 && c1 > 0 && c1 % 64 == 0 && scratch_bytes >= conv1_wgrad_tc_scratch_bytes(batch, c1),
               "bad width or scratch too small");
  return conv1_wgrad_tc(images, dpool, pool, arg, dw, db, (float*)scratch, batch, img_h, img_w, crop_h, crop_w,
                        (cudaStream_t)stream, false, c1);
}
int b200_debug_conv_timeline(int64_t* device_buf8) {
  conv5x5_set_timeline(reinterpret_cast<long long*>(device_buf8));
  fc_tail_set_timeline(device_buf8 ? reinterpret_cast<long long*>(device_buf8) + 16 : nullptr);
  conv1_set_timeline(reinterpret_cast<long long*>(device_buf8));
  gemm_set_timeline(reinterpret_cast<long long*>(device_buf8));
  return B200_OK;
}
int b200_debug_conv_variant(int32_t variant) {
  B200_REQUIRE(variant >= -1 && variant <= 1, "debug_conv_variant: -1 (automatic), 0 or 1");
  conv5x5_force_variant(variant);
  return B200_OK;
}
int b200_debug_trace(uint64_t* device_buf) {
#ifdef B200_TRACE
  unsigned long long* p = reinterpret_cast<unsigned long long*>(device_buf);
  trace_set_fc_tail(p); trace_set_tc_conv(p); trace_set_tc_conv1(p); trace_set_tc_gemm(p); trace_set_f32(p); trace_set_p2p(p);
  B200_LAUNCH_CHECK();
  return B200_OK;
#else
  (void)device_buf;
  set_error("b200_debug_trace: this library was built without B200_TRACE");
  return B200_E_STATE;
#endif
}
int b200_conv5x5_packed_bytes(int32_t c_in, int32_t c_out, int64_t* nbytes) {
  B200_REQUIRE(nbytes && c_in > 0 && c_out > 0, "bad argument");
  *nbytes = conv5x5_packed_bytes(c_in, c_out);
  return B200_OK;
}
int b200_conv5x5_pack_weights(const float* w, void* w_packed, int32_t c_in, int32_t c_out, void* stream) {
  B200_REQUIRE(w && w_packed, "null pointer");
  return pack_conv_weights(w, w_packed, c_in, c_out, (cudaStream_t)stream);
}
int b200_conv5x5_fwd_bf16(const float* x, const void* w_packed, const float* bias, float* y, int32_t batch, int32_t h,
                          int32_t wd, int32_t c_in, int32_t c_out, int32_t relu, void* stream) {
  B200_REQUIRE(x && w_packed && y, "null pointer");
  return conv5x5_fwd_tc(x, w_packed, bias, y, batch, h, wd, c_in, c_out, relu, (cudaStream_t)stream);
}
int b200_conv5x5_pool_fwd_bf16(const float* x, const void* w_packed, const float* bias, float* y, float* pool,
                               uint8_t* arg, int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out,
                               void* stream) {
  B200_REQUIRE(x && w_packed && bias && pool && arg, "null pointer");
  return conv5x5_pool_fwd_tc(x, w_packed, bias, y, pool, arg, batch, h, wd, c_in, c_out, (cudaStream_t)stream);
}
int b200_conv5x5_pool_fwd_bf16x(const void* x_bf16, const void* w_packed, const float* bias, float* y, float* pool,
                                uint8_t* arg, int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out,
                                void* stream) {
  B200_REQUIRE(x_bf16 && w_packed && bias && pool && arg, "null pointer");
  return conv5x5_pool_fwd_tc(nullptr, w_packed, bias, y, pool, arg, batch, h, wd, c_in, c_out, (cudaStream_t)stream, x_bf16);
}
int b200_conv5x5_dgrad_bf16(const float* dy, const void* w_packed, float* dx, int32_t batch, int32_t h, int32_t wd,
                            int32_t c_in, int32_t c_out, void* stream) {
  B200_REQUIRE(dy && w_packed && dx, "null pointer");
  return conv5x5_dgrad_tc(dy, w_packed, dx, batch, h, wd, c_in, c_out, (cudaStream_t)stream);
}
int b200_conv5x5_wgrad_bf16_scratch_bytes(int32_t batch, int64_t* nbytes) {
  B200_REQUIRE(nbytes && batch > 0, "bad argument");
  *nbytes = conv5x5_wgrad_tc_scratch_bytes(batch);
  return B200_OK;
}
int b200_conv5x5_wgrad_bf16_scratch_bytes_c(int32_t batch, int32_t c_in, int32_t c_out, int64_t* nbytes) {
  B200_REQUIRE(nbytes && batch > 0 && c_in > 0 && c_out > 0 && c_in % 64 == 0 && c_out % 64 == 0, "bad argument");
  *nbytes = conv5x5_wgrad_tc_scratch_bytes(batch, c_in, c_out);
  return B200_OK;
}
int b200_conv5x5_wgrad_bf16(const float* x, const float* dy, float* dw, void* scratch, int64_t scratch_bytes,
                            int32_t batch, int32_t h, int32_t wd, int32_t c_in, int32_t c_out, void* stream) {
  B200_REQUIRE(x && dy && dw && scratch, "null pointer");
  B200_REQUIRE(batch > 0 && c_in > 0 && c_out > 0 && c_in % 64 == 0 && c_out % 64 == 0, "channels must be multiples of 64");
  B200_REQUIRE(scratch_bytes >= conv5x5_wgrad_tc_scratch_bytes(batch, c_in, c_out), "scratch too small");
  return conv5x5_wgrad_tc(x, dy, dw, (float*)scratch, batch, h, wd, c_in, c_out, (cudaStream_t)stream);
}
int b200_conv5x5_dgrad_pool_bf16(const float* dpool, const float* pool, const uint8_t* arg, const void* w_packed,
                                 float* dx, float* dbias, int32_t batch, int32_t h, int32_t wd, int32_t c_in,
                                 int32_t c_out, void* stream) {
  B200_REQUIRE(dpool && pool && arg && w_packed && dx, "null pointer");
  return conv5x5_dgrad_pool_tc(dpool, pool, arg, w_packed, dx, dbias, nullptr, batch, h, wd, c_in, c_out,
                               (cudaStream_t)stream);
}
int b200_conv5x5_wgrad_pool_bf16(const float* x, const float* dpool, const float* pool, const uint8_t* arg, float* dw,
                                 void* scratch, int64_t scratch_bytes, int32_t batch, int32_t h, int32_t wd,
                                 int32_t c_in, int32_t c_out, void* stream) {
  B200_REQUIRE(x && dpool && pool && arg && dw && scratch, "null pointer");
  B200_REQUIRE(batch > 0 && scratch_bytes >= conv5x5_wgrad_tc_scratch_bytes(batch), "scratch too small");
  return conv5x5_wgrad_pool_tc(x, dpool, pool, arg, dw, (float*)scratch, batch, h, wd, c_in, c_out,
                               (cudaStream_t)stream);
}
int b200_fc_fwd_bf16(const float* x, const float* w, const float* b, float* y, int32_t batch, int32_t k, int32_t n,
                     int32_t relu, void* stream) {
  B200_REQUIRE(x && w && y, "null pointer");
  return fc_fwd_tc(x, w, b, y, batch, k, n, relu, (cudaStream_t)stream);
}
int b200_fc_bwd_bf16(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                     void* scratch, int64_t scratch_bytes, int32_t batch, int32_t k, int32_t n, int32_t mask_x,
                     void* stream) {
  B200_REQUIRE(x && w && dy && dw, "null pointer");
  B200_REQUIRE(batch > 0 && k > 0 && n > 0, "bad shape");
  B200_REQUIRE(!db || (scratch && scratch_bytes >= colsum_scratch_bytes(batch, n)), "scratch too small");
  TRY(fc_bwd_tc(x, w, dy, dx, dw, batch, k, n, mask_x, (cudaStream_t)stream));
  if (db) return colsum_f32(dy, db, (float*)scratch, batch, n, (cudaStream_t)stream);
  return B200_OK;
}
int b200_fc_tail_f32(float* full1, const float* w2, const float* b2, const float* w3, const float* b3,
                     const int32_t* labels, float* full2, float* logits, float* dlogits, float* dfull2, float* dfull1,
                     int32_t* pred, float* loss, int32_t* n_correct, int32_t* n_bad, float* dw2, float* db2, float* dw3,
                     float* db3, float* db1, void* scratch, int64_t scratch_bytes, int32_t batch, int32_t f1, int32_t f2,
                     int32_t n_cls, void* stream) {
  B200_REQUIRE(full1 && w2 && b2 && w3 && b3 && labels && full2 && logits && dlogits && dfull2 && dfull1 && pred &&
                   loss && n_correct && n_bad && dw2 && db2 && dw3 && db3 && db1 && scratch,
               "null pointer");
  B200_REQUIRE(batch > 0 && f1 > 0 && f2 > 0 && n_cls > 0 && scratch_bytes >= fc_tail_scratch_bytes(batch, f1, f2, n_cls),
               "scratch too small");
  return fc_tail(nullptr, 0, nullptr, full1, w2, b2, w3, b3, labels, full2, logits, dlogits, dfull2, dfull1, pred, loss,
                 n_correct, n_bad, dw2, db2, dw3, db3, db1, scratch, nullptr, 0, batch, f1, f2, n_cls,
                 (cudaStream_t)stream);
}
int b200_fc_tail_scratch_bytes(int32_t batch, int32_t f1, int32_t f2, int32_t n_cls, int64_t* nbytes) {
  B200_REQUIRE(nbytes && batch > 0 && f1 > 0 && f2 > 0 && n_cls > 0, "bad argument");
  *nbytes = fc_tail_scratch_bytes(batch, f1, f2, n_cls);
  return B200_OK;
}
int b200_head_f32(const float* logits, const int32_t* labels, float* loss_mean, float* dlogits, int32_t* pred,
                  int32_t* n_correct, int32_t* n_bad, int32_t batch, int32_t n_cls, void* stream) {
  B200_REQUIRE(logits && labels, "null pointer");
  return head_f32(logits, labels, loss_mean, dlogits, pred, n_correct, n_bad, batch, n_cls,
                  (cudaStream_t)stream);
}
int b200_shuffle_batch(const uint8_t* images, const int32_t* labels, int64_t n_records, const int32_t* file_order,
                       int32_t n_order, int32_t per_file, const int32_t* slots_in, int32_t* slots_out, int32_t capacity,
                       uint64_t seed, int64_t step, int64_t stream_pos, uint8_t* images_out, int32_t* labels_out,
                       int32_t batch, int32_t image_bytes, void* stream) {
  return shuffle_batch(images, labels, n_records, file_order, n_order, per_file, slots_in, slots_out, capacity, seed, step,
                       stream_pos, images_out, labels_out, batch, image_bytes, (cudaStream_t)stream);
}
int b200_sgd_apply_f32(float* p, const float* g, int64_t n, float lr, float scale, void* stream) {
  B200_REQUIRE(p && g, "null pointer");
  return sgd_apply_f32(p, g, n, lr, scale, (cudaStream_t)stream);
}

}  // extern "C"
