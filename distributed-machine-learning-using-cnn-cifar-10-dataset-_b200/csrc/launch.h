// Internal C++ launch API shared by abi.cu and the kernel files.  Every function enqueues
// on `st`, returns a B200_* code and never synchronises.
#pragma once
#include "common.cuh"

namespace b200 {

// ---- fp32 validation kernels (f32_kernels.cu) ----
int crop_cast_f32(const uint8_t* img, float* out, int B, int ih, int iw, int ch, int cw, int cin, cudaStream_t st);
int conv5x5_fwd_f32(const float* x, const float* w, const float* bias, float* y, int B, int H, int W,
                    int Cin, int Cout, int relu, cudaStream_t st);
int conv5x5_rot180_f32(const float* w, float* wr, int Cin, int Cout, cudaStream_t st);
int64_t conv5x5_wgrad_scratch_bytes(int B, int H, int W, int Cin, int Cout);
int conv5x5_wgrad_f32(const float* x, const float* dy, float* dw, float* db, float* scratch, int B, int H,
                      int W, int Cin, int Cout, cudaStream_t st);
int maxpool_fwd_f32(const float* x, float* y, uint8_t* arg, int B, int H, int W, int C, cudaStream_t st);
// colsum_atomic (optional, must be zero on entry): += sum of dx over (B,H,W) per channel, i.e. the
// BiasAddGrad of the layer that fed the pool
int maxpool_bwd_f32(const float* dy, const uint8_t* arg, const float* x, float* dx, int B, int H, int W,
                    int C, int relu_mask, cudaStream_t st, float* colsum_atomic = nullptr,
                    void* dx_bf16 = nullptr);      // (optional) dx also rounded to bf16, same layout
// y = relu?(sum of nsplit partials [nsplit][B][N] + bias): closes fc_fwd_tc_splitk
int splitk_finish_f32(const float* part, int nsplit, const float* bias, float* y, int B, int N, int relu, cudaStream_t st);
int fc_fwd_f32(const float* x, const float* w, const float* b, float* y, int B, int K, int N, int relu,
               cudaStream_t st);
int64_t colsum_scratch_bytes(int rows, int cols);
int colsum_f32(const float* a, float* out, float* scratch, int rows, int cols, cudaStream_t st);
int fc_bwd_f32(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
               float* scratch, int B, int K, int N, int mask_x, cudaStream_t st);
int head_f32(const float* logits, const int32_t* labels, float* loss_mean, float* dlogits, int32_t* pred,
             int32_t* n_correct, int32_t* n_bad, int B, int ncls, cudaStream_t st);
int sgd_apply_f32(float* p, const float* g, int64_t n, float lr, float scale, cudaStream_t st);
// the update with the first n_part gradients still in S partials [S][n_part] (summed in a fixed order, stored to g)
int sgd_apply_reduce_f32(float* p, float* g, int64_t n, float lr, float scale, const float* part, int S, int n_part,
                         cudaStream_t st);
// the same sum in the order sgd_apply_reduce_f32 uses (one warp per 16 bytes): what conv1's gradient goes through when
// the update does not follow at once, so that both routes give the same bits
int reduce_partials_warp(const float* part, float* out, int n, int S, int stride, cudaStream_t st);
// out[i] = sum over z < S of part[z * stride + i], fixed order (tc_conv.cu); n, stride multiples of 4 (stride 0 = n)
int reduce_partials(const float* part, float* out, int n, int S, cudaStream_t st, int stride = 0);


// tf.train.shuffle_batch (:82-90) over a record set resident in HBM (shuffle_batch.cu)
int shuffle_batch(const uint8_t* images, const int32_t* labels, int64_t n_records, const int32_t* file_order,
                  int n_order, int per_file, const int32_t* slots_in, int32_t* slots_out, int capacity, uint64_t seed,
                  int64_t step, int64_t stream_pos, uint8_t* images_out, int32_t* labels_out, int batch, int image_bytes,
                  cudaStream_t st);

// ---- bf16 tensor-core kernels (tc_*.cu): fp32 tensors in HBM, bf16 operands, fp32 accumulation ----
int fc_fwd_tc(const float* x, const float* w, const float* b, float* y, int B, int K, int N, int relu,
              cudaStream_t st);
int fc_bwd_tc(const float* x, const float* w, const float* dy, float* dx, float* dw, int B, int K, int N,
              int mask_x, cudaStream_t st);
int fc_dw_tc(const float* x, const float* dy, float* dw, int B, int K, int N, cudaStream_t st);
int fc_dx_tc(const float* x, const float* w, const float* dy, float* dx, int B, int K, int N, int mask_x,
             cudaStream_t st);
// wp: the filter packed once per step by pack_conv_weights (conv5x5_packed_bytes bytes)
void conv5x5_set_timeline(long long* dev_buf);   // diagnostics: CTA 0 of conv5x5_tc_kernel stamps clock64 per phase
void conv5x5_force_variant(int v);               // tests: 0 = one CTA per SM, 1 = two per SM (where the shape allows), -1 = automatic
int64_t conv5x5_packed_bytes(int Cin, int Cout);
int pack_conv_weights(const float* w, void* wp, int Cin, int Cout, cudaStream_t st);
int conv5x5_fwd_tc(const float* x, const void* wp, const float* bias, float* y, int B, int H, int W, int Cin,
                   int Cout, int relu, cudaStream_t st, const void* x_bf16 = nullptr);      // (optional) x already in bf16
// x_bf16 (optional): the same activations already rounded to bf16, [B,H,W,Cin]; read instead of x (half the bytes)
int conv5x5_pool_fwd_tc(const float* x, const void* wp, const float* bias, float* y, float* pool, uint8_t* arg, int B,
                        int H, int W, int Cin, int Cout, cudaStream_t st, const void* x_bf16 = nullptr,
                        void* pool_img = nullptr);   // (optional) pooled output as FC1's bf16 operand image [H'W'Cout/8][B][8]
int conv5x5_dgrad_tc(const float* dy, const void* wp, float* dx, int B, int H, int W, int Cin, int Cout,
                     cudaStream_t st, const void* dy_bf16 = nullptr);      // (optional) dy already in bf16: TMA-filled strip
// the same two gradients fed by the POOLED gradient: dconv = MaxPoolGrad(dpool, arg) * [pool > 0] is gathered
// inside the kernels (64 channels); dbias (optional, zero on entry) receives sum(dconv) through atomics;
// dconv_out (optional) receives dconv as a dense [B,H,W,64] fp32 tensor
int conv5x5_dgrad_pool_tc(const float* dpool, const float* pool, const uint8_t* arg, const void* wp, float* dx,
                          float* dbias, float* dconv_out, int B, int H, int W, int Cin, int Cout, cudaStream_t st,
                          void* dconv_out_bf16 = nullptr,       // (optional) dconv as a dense bf16 tensor
                          float* dbias_part = nullptr,          // (optional, instead of dbias) per-CTA bias-gradient sums
                          int* n_part = nullptr);               //   [*n_part][64] for a fixed-order reduction (<= B of them)
int conv5x5_wgrad_pool_tc(const float* x, const float* dpool, const float* pool, const uint8_t* arg, float* dw,
                          float* scratch, int B, int H, int W, int Cin, int Cout, cudaStream_t st);
int64_t conv5x5_wgrad_tc_scratch_bytes(int B, int Cin = 64, int Cout = 64);
// x_bf16 / dy_bf16 (optional, both or none): bf16 copies of x / dy, read instead of the fp32 tensors
// db_part / db / n_db_part (optional): the per-CTA bias-gradient sums [n_db_part][64] of conv5x5_dgrad_pool_tc, reduced
// to db [64] in the same launch as the filter-gradient partials
int conv5x5_wgrad_tc(const float* x, const float* dy, float* dw, float* scratch, int B, int H, int W, int Cin,
                     int Cout, cudaStream_t st, const void* x_bf16 = nullptr, const void* dy_bf16 = nullptr,
                     const float* db_part = nullptr, float* db = nullptr, int n_db_part = 0);
int fc_fwd_tc_splits(int B, int K, int N);
// FC1's three products on bf16 operand IMAGES ([chunk of 8][rows][8], the shared-memory layout of a stage: fetched
// with bulk copies, nothing converted): x_img [K/8][B][8] (written by conv2's pool epilogue), w_img [N/8][K][8]
// (pack_fc_weights, once per step), dy_img [N/8][B][8] (written by the FC tail).  Same arithmetic as the fp32-source
// entry points (operands rounded to bf16 either way).
bool fc_images_supported(int B, int K, int N);
int64_t fc_weight_image_bytes(int K, int N);
int pack_fc_weights(const float* w, void* img, int K, int N, cudaStream_t st);
int fc_fwd_tc_splitk_img(const void* x_img, const void* w_img, float* part, int B, int K, int N, cudaStream_t st);
int fc_dw_tc_img(const void* x_img, const void* dy_img, float* dw, int B, int K, int N, cudaStream_t st);
int fc_dx_tc_img(const void* dy_img, const void* w_img, float* dx, int B, int K, int N, cudaStream_t st);
int fc_fwd_tc_splitk(const float* x, const float* w, float* part, int B, int K, int N, cudaStream_t st);
void fc_tail_set_timeline(long long* dev_buf);
void conv1_set_timeline(long long* dev_buf);
void gemm_set_timeline(long long* dev_buf);      // slots [0,6), shared with conv5x5_tc_kernel: the last kernel that ran wins
// diagnostics build (B200_TRACE): hand the %globaltimer trace buffer to each translation unit
void trace_set_fc_tail(unsigned long long*);
void trace_set_tc_conv(unsigned long long*);
void trace_set_tc_conv1(unsigned long long*);
void trace_set_tc_gemm(unsigned long long*);
void trace_set_f32(unsigned long long*);
void trace_set_p2p(unsigned long long*);   // diagnostics (see conv5x5_set_timeline)
int64_t fc_tail_scratch_bytes(int B, int F1, int F2, int NC);
int fc_tail(const float* fc1_part, int n_part, const float* b1, float* full1, const float* w2, const float* b2,
            const float* w3, const float* b3, const int32_t* labels, float* full2, float* logits, float* dlogits,
            float* dfull2, float* dfull1, int32_t* pred, float* loss, int32_t* n_correct, int32_t* n_bad, float* dw2,
            float* db2, float* dw3, float* db3, float* db1, void* scratch, float* zero_ptr, int zero_n, int B, int F1,
            int F2, int NC, cudaStream_t st, bool w2_image_ready = false, cudaStream_t grads_stream = nullptr,
            cudaEvent_t rows_done = nullptr, bool overlap_producer = false,
            void* dfull1_img = nullptr);     // (optional) dfull1 also as FC1's bf16 operand image [F1/8][B][8]
int fc_tail_prepare_w2(const float* w2, void* scratch, int B, int F1, int F2, int NC, cudaStream_t st);
// allreduce + SGD in one kernel over NVLink peer memory (p2p_update.cu)
int p2p_allreduce_sgd(const void* const* peer_grads, void* const* peer_params, void* const* peer_pads,
                      const void* mc_grads, void* mc_params, uint32_t* state, int rank, int world, int64_t total,
                      float lr, float scale, cudaStream_t st,
                      int64_t offset = 0, int channel = 0, int max_ctas = 0,
                      // (optional) the first red_n floats of this rank's range still sit in red_S partials [red_S][red_n]:
                      // summed (fixed order) by the kernel's first CTAs before it tells the peers "ready"
                      const float* red_part = nullptr, int red_S = 0, int red_n = 0);
// one-shot exchange of a small range: push into every rank's inbox, one flag wait, local rank-ordered sum + update
// (p2p_update.cu); inbox = [2][world][n] floats of symmetric memory per rank
int p2p_push_sum_sgd(float* grads, float* params, const float* inbox_local, void* inbox_mc, void* const* inbox_peers,
                     void* const* peer_pads, uint32_t* state, int rank, int world, int64_t n, float lr, float scale,
                     cudaStream_t st, int channel, const float* red_part = nullptr, int red_S = 0, int red_n = 0);
bool conv1_tc_supported(int H, int W, int c_in, int c1);
int conv1_pool_fwd_tc(const uint8_t* img, const float* w, const float* bias, float* pool, uint8_t* arg,
                      float* conv_dbg, int B, int ih, int iw, int H, int W, cudaStream_t st,
                      void* pool_bf16 = nullptr,       // (optional) second copy of pool, rounded to bf16
                      int C1 = 64);                    // output channels (a multiple of 64: one CTA per image and 64-channel slice)
int64_t conv1_wgrad_tc_scratch_bytes(int B, int C1 = 64);
// the kernel leaves one partial [4864] (dw [25][3][64], then db [64]) per CTA in scratch and a fixed-order reduction
// writes dw / db -- unless defer_reduce, in which case the caller sums the conv1_wgrad_tc_partials(B) partials itself
// (sgd_apply_reduce_f32 does, at the end of a step; dw and db are then adjacent in the flat gradient buffer)
int conv1_wgrad_tc_partials(int B);
int conv1_wgrad_tc(const uint8_t* img, const float* dpool, const float* pool, const uint8_t* arg, float* dw, float* db,
                   float* scratch, int B, int ih, int iw, int H, int W, cudaStream_t st, bool defer_reduce = false,
                   int C1 = 64);

}  // namespace b200
