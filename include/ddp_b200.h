/* ddp_b200 — C ABI of the native library behind the B200 data-parallel trainer.
 *
 * Every function enqueues work on the CUDA stream passed as `stream` (a cudaStream_t cast to
 * void*; NULL = the legacy default stream) and returns 0 on success, a cudaError_t value, or
 * one of the DDPB_ERR_* codes below.  ddpb_error_string() renders any of them.
 * Pointers are device pointers unless a comment says "host".  Nothing here allocates device
 * memory unless the name says so (ddpb_comm_*): callers own all buffers, including scratch,
 * and query scratch sizes with the *_scratch_bytes() functions.
 *
 * Layouts
 *   parameters / gradients : fp32, the reference's state_dict layouts
 *                            (conv weight [Cout][Cin][3][3], linear weight [Nout][C*H*W] with
 *                            k = c*H*W + h*W + w, reference model.py:8-16).
 *   activations            : bf16 "padded planar-chunk" planes, see csrc/layout.cuh:
 *                            C/8 planes x ddpb_plane_rows(batch_cap,H,W) rows x 8 channels;
 *                            pixel (b,h,w) is row ddpb_guard_rows() + b*(H+2)*(W+2) + (h+1)*(W+2) + (w+1).
 *                            Buffers must be zero-filled once after allocation.
 *   `batch` is the number of images in this step, `batch_cap` the capacity the activation
 *   buffers were sized for (batch <= batch_cap), so a short tail batch reuses the same buffers
 *   (reference data.py:21-25 keeps the last short batch: drop_last=False).
 *
 * The reference (zahmedy/PyTorch-Distributed-Data-Parallel-DDP-Trainer) has no native code; the
 * comments name the reference line whose delegated library call each entry point replaces.
 */
#ifndef DDP_B200_H_
#define DDP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDPB_ERR_SHAPE 9001 /* unsupported shape / argument */
#define DDPB_ERR_STATE 9002 /* call made in the wrong state */
#define DDPB_ERR_LIMIT 9003 /* size beyond a documented fixed limit */

const char* ddpb_error_string(int code);
/* Programmatic dependent launch for the step kernels: 1 on, 0 off, -1 = DDPB_PDL env (default on).
 * The setting is per host thread (it applies to the launches that thread makes). */
void ddpb_set_pdl(int on);
int ddpb_get_pdl(void);
int ddpb_num_sms(void);
/* Tuning hooks (tools/kprof.py, tools/mma_bench.py): SM-clock phase stamps of the conv kernels into
 * buf[block*16 + slot] (NULL = off), and a tcgen05.mma issue-cost microbenchmark. */
void ddpb_set_prof_buffer(void* buf);
int ddpb_mma_microbench(int config, int a_shift16, int b_shift16, int outer, unsigned lbo_a,
                        unsigned sbo_a, unsigned lbo_b, unsigned sbo_b,
                        unsigned long long* out_dev, void* stream);

/* ---- geometry of the activation layout */
long long ddpb_plane_rows(int batch_cap, int H, int W);
int ddpb_guard_rows(void);
int ddpb_tile_rows(void);

/* ---- input pipeline: x[i] = images[idx[i]]/255 (fp32 [batch][pix]), y[i] = labels[idx[i]].
 * Replaces torchvision ToTensor + DataLoader collate, reference data.py:11-14,21-25. */
int ddpb_gather_u8(const void* images_u8, const void* labels_i64, const void* idx_i64, float* x,
                   void* y_i64, int batch, int pix, void* stream);

/* ---- conv1 (Cin=1 -> 32), bias + ReLU fused; x is fp32 NCHW.  reference model.py:9-10 */
int ddpb_conv1_fwd(const float* x, const float* w, const float* b, void* out_planes, int cin,
                   int cout, int batch, int batch_cap, int H, int W, void* stream);
/* dw_db: cout*cin*9 weight gradients immediately followed by cout bias gradients. */
size_t ddpb_conv1_wgrad_scratch_bytes(int cin, int cout, int batch_cap, int H, int W);
int ddpb_conv1_wgrad(const void* dact_planes, const float* x, float* dw_db, float* scratch,
                     int cin, int cout, int batch, int batch_cap, int H, int W, void* stream);

/* Tensor-core versions (the step engine uses these): forward = im2col rows built in shared memory
 * as TF32 + tcgen05.mma kind::tf32; wgrad = dact planes as MN-major A, im2col rows as two bf16
 * terms (hi + lo) + a ones column for the bias gradient. */
int ddpb_conv1_fwd_tc(const float* x, const float* w, const float* b, void* out_planes, int cin,
                      int cout, int batch, int batch_cap, int H, int W, void* stream);
size_t ddpb_conv1_wgrad_tc_scratch_bytes(int cin, int cout);
int ddpb_conv1_wgrad_tc(const void* dact_planes, const float* x, float* dw_db, float* scratch,
                        int cin, int cout, int batch, int batch_cap, int H, int W, void* stream);
/* Same kernel, but the split-K partials stay in scratch[nparts][cout*cin*9 + cout] and *nparts_out
 * (host) receives their count: feed them to ddpb_reduce_allreduce_sgd. */
int ddpb_conv1_wgrad_tc_partials(const void* dact_planes, const float* x, float* scratch,
                                 int* nparts_out, int cin, int cout, int batch, int batch_cap, int H,
                                 int W, void* stream);
/* out[i] = sum_p part[p*n + i], fixed order. */
int ddpb_reduce_partials(const float* part, int nparts, long long n, float* out, void* stream);

/* The same gradient as ddpb_conv1_wgrad up to rounding (the batch is split into a different number
 * of partial blocks, so the summation order differs), for callers that cannot pass scratch because
 * they cannot allocate device memory (a host that gave all of it to its own pool, or a plug-in barred from
 * cudaMalloc/cudaFree because those stall every stream on the device).  The partial sums live in
 * an 80 KB array inside the library's device image, held by every GPU that loads the library.
 * Limit: 64 partial blocks = batches of up to ddpb_conv1_wgrad_noalloc_max_batch(H,W) images
 * (147 for 28x28); larger batches return DDPB_ERR_LIMIT and launch nothing.  Allocates nothing.
 * The first call on a device resolves the array's address (make it before any stream capture).
 * Calls share the array: calls that overlap on the GPU race; nothing in it needs zeroing.
 * This entry point has no counterpart in the reference. */
int ddpb_conv1_wgrad_noalloc(const void* dact_planes, const float* x, float* dw_db, int cin,
                             int cout, int batch, int batch_cap, int H, int W, void* stream);
int ddpb_conv1_wgrad_noalloc_max_batch(int H, int W);

/* ---- 3x3 pad-1 convolution on tcgen05 tensor cores (bf16 operands, fp32 accumulate in TMEM),
 * operands fetched with bulk async copies.  reference model.py:11-12 (forward),
 * train_ddp.py:199 (backward through cuDNN). Supported (cin,cout): (32,64). */
int ddpb_conv3x3_fwd(const void* in_planes, const float* w, const float* bias, void* out_planes,
                     int cin, int cout, int batch, int batch_cap, int H, int W, void* stream);
/* din = relu'(mask) * conv_transpose(dout, w); mask = the forward activation that fed the conv. */
int ddpb_conv3x3_dgrad(const void* dout_planes, const float* w, const void* mask_planes,
                       void* din_planes, int cin, int cout, int batch, int batch_cap, int H, int W,
                       void* stream);
size_t ddpb_conv3x3_wgrad_scratch_bytes(int cin, int cout);
int ddpb_conv3x3_wgrad(const void* dout_planes, const void* in_planes, float* dw, float* db,
                       float* scratch, int cin, int cout, int batch, int batch_cap, int H, int W,
                       void* stream);
/* Same tensor-core kernel, but the per-CTA accumulator images stay in scratch[nparts][cout*(9*cin+8)]
 * and *nparts_out (host) receives their count: feed them to ddpb_reduce_allreduce_sgd, layout 1. */
int ddpb_conv3x3_wgrad_partials(const void* dout_planes, const void* in_planes, float* scratch,
                                int* nparts_out, int cin, int cout, int batch, int batch_cap, int H,
                                int W, void* stream);

/* ---- reference-precision mode ("bf16x3"): what torch runs for the reference is fp32 storage, TF32
 * convolutions and an fp32 linear layer (reference model.py:8-16 through cuDNN/cuBLAS).  Here:
 *   split planes  = every value as two bf16 terms x = hi + lo (C/8 hi planes then C/8 lo planes,
 *                   4 bytes per value, 16-bit significand); GEMMs over split operands issue three
 *                   tcgen05.mma per K step (hi.hi + lo.hi + hi.lo, fp32 accumulation);
 *   fp32 planes   = 4 fp32 channels per 16-byte row, for tensors the fp32 CUDA-core ops read.
 * packed = the split bf16 tile images of the conv weight [fwd hi | fwd lo | dgrad hi | dgrad lo]
 * (ddpb_conv3x3_pack_split, or rewritten in place by ddpb_reduce_allreduce_sgd_pack).
 * Measured error of every op against an fp64 oracle, next to torch's own path: profiles/precision.md. */
int ddpb_conv1_fwd_tc_split(const float* x, const float* w, const float* b, void* out_split_planes,
                            int cin, int cout, int batch, int batch_cap, int H, int W, void* stream);
size_t ddpb_conv3x3_pack_split_bytes(int cin, int cout);
int ddpb_conv3x3_pack_split(const float* w, void* packed, int cin, int cout, void* stream);
/* out: fp32 planes when out_f32 != 0, else split planes. */
int ddpb_conv3x3_fwd_split(const void* in_split_planes, const void* packed, const float* bias,
                           void* out_planes, int cin, int cout, int batch, int batch_cap, int H, int W,
                           int out_f32, void* stream);
int ddpb_conv3x3_dgrad_split(const void* dout_split_planes, const void* packed,
                             const void* mask_split_planes, void* din_planes, int cin, int cout,
                             int batch, int batch_cap, int H, int W, int out_f32, void* stream);
/* bias_part/nsplit/nbx (optional): exact-fp32 bias-gradient parts left by ddpb_linear_bwd_f32;
 * db is then their fixed-order sum instead of the tensor-core ones-column. */
int ddpb_conv3x3_wgrad_split(const void* dout_split_planes, const void* in_split_planes, float* dw,
                             float* db, float* scratch, const float* bias_part, int nsplit, int nbx,
                             int cin, int cout, int batch, int batch_cap, int H, int W, void* stream);
int ddpb_conv3x3_wgrad_split_partials(const void* dout_split_planes, const void* in_split_planes,
                                      float* scratch, int* nparts_out, int cin, int cout, int batch,
                                      int batch_cap, int H, int W, void* stream);
/* conv1 weight gradient from split dact planes on the tensor cores (two passes: hi, lo). */
int ddpb_conv1_wgrad_tc_split(const void* dact_split_planes, const float* x, float* dw_db,
                              float* scratch, int cin, int cout, int batch, int batch_cap, int H, int W,
                              void* stream);
int ddpb_conv1_wgrad_tc_split_partials(const void* dact_split_planes, const float* x, float* scratch,
                                       int* nparts_out, int cin, int cout, int batch, int batch_cap,
                                       int H, int W, void* stream);
/* conv1 weight gradient in exact fp32 on the CUDA cores from fp32 dact planes (cuDNN runs this
 * Cin = 1 layer in plain fp32 too); partials have the layout of ddpb_conv1_wgrad_tc_partials. */
size_t ddpb_conv1_wgrad_f32_scratch_bytes(int cin, int cout);
int ddpb_conv1_wgrad_f32(const void* dact_f32_planes, const float* x, float* dw_db, float* scratch,
                         int cin, int cout, int batch, int batch_cap, int H, int W, void* stream);
int ddpb_conv1_wgrad_f32_partials(const void* dact_f32_planes, const float* x, float* scratch,
                                  int* nparts_out, int cin, int cout, int batch, int batch_cap, int H,
                                  int W, void* stream);
/* fp32 linear layer (FMA on the CUDA cores, like cuBLAS with TF32 off) on fp32 activation planes.
 * bwd: dact = split planes (split_out != 0) or fp32 planes rounded to tf32; bias_part (optional,
 * ddpb_linear_bwd_f32_bias_part_bytes) receives exact-fp32 per-channel partial sums of dact and
 * *nsplit_out / *nbx_out (host) their extents. */
size_t ddpb_linear_fwd_f32_scratch_bytes(int nout, int C, int batch_cap, int H, int W);
int ddpb_linear_fwd_f32(const void* act_f32_planes, const float* W_fl, float* partial, int* S_out,
                        int nout, int C, int batch, int batch_cap, int H, int W, void* stream);
/* Forward with the loss fused into its tail (one launch instead of ddpb_linear_fwd_f32 +
 * ddpb_softmax_xent): the last block to finish adds the slices and computes loss / dlogits / dbias.
 * batch <= ddpb_linear_fwd_f32_xent_max_batch() (128), else DDPB_ERR_LIMIT.  Calls on one device must
 * not overlap.  reference model.py:16,19 + train_ddp.py:40,198. */
int ddpb_linear_fwd_f32_xent_max_batch(void);
int ddpb_linear_fwd_f32_xent(const void* act_f32_planes, const float* W_fl, float* partial,
                             const float* bias, const void* labels_i64, float* logits, float* dlogits,
                             float* dbias, float* loss, int nout, int C, int batch, int batch_cap, int H,
                             int W, void* stream);
size_t ddpb_linear_bwd_f32_bias_part_bytes(int C, int batch_cap, int H, int W);
int ddpb_linear_bwd_f32(const void* act_f32_planes, const float* dlogits, const float* W_fl,
                        void* dact_planes, float* dW, float* scratch, int nout, int C, int batch,
                        int batch_cap, int H, int W, int split_out, float* bias_part, int* nsplit_out,
                        int* nbx_out, void* stream);
/* The same in two halves: mode 1 = dact + bias parts only, mode 2 = the dW half only (mode 0 = both).
 * The step engine runs mode 1 on its main stream and mode 2 on the fl bucket's side stream, where it
 * overlaps the convolution backward. */
int ddpb_linear_bwd_f32_mode(const void* act_f32_planes, const float* dlogits, const float* W_fl,
                             void* dact_planes, float* dW, float* scratch, int nout, int C, int batch,
                             int batch_cap, int H, int W, int split_out, float* bias_part,
                             int* nsplit_out, int* nbx_out, int mode, void* stream);
/* tf32 building blocks on fp32 planes (forward / data gradient only: tcgen05 defines an MN-major tf32
 * operand — the weight gradient's view of a plane — only for the 128B_BASE32B swizzled layout). */
int ddpb_conv1_fwd_tc_f32(const float* x, const float* w, const float* b, void* out_f32_planes,
                          int cin, int cout, int batch, int batch_cap, int H, int W, void* stream);
size_t ddpb_conv3x3_pack_tf32_bytes(int cin, int cout);
int ddpb_conv3x3_pack_tf32(const float* w, void* packed, int cin, int cout, void* stream);
int ddpb_conv3x3_fwd_tf32(const void* in_f32_planes, const void* packed, const float* bias,
                          void* out_f32_planes, int cin, int cout, int batch, int batch_cap, int H,
                          int W, int round_out, void* stream);
int ddpb_conv3x3_dgrad_tf32(const void* dout_f32_planes, const void* packed, const void* mask_f32_planes,
                            void* din_f32_planes, int cin, int cout, int batch, int batch_cap, int H,
                            int W, int round_out, void* stream);

/* ---- linear layer over the flattened activation (NCHW flatten order).  reference model.py:16,19.
 * fwd writes per-slice partial logits partial[S][batch][nout]; *S_out (host) receives S. */
size_t ddpb_linear_fwd_scratch_bytes(int nout, int C, int batch_cap, int H, int W);
int ddpb_linear_fwd(const void* act_planes, const float* W_fl, float* partial, int* S_out,
                    int nout, int C, int batch, int batch_cap, int H, int W, void* stream);
/* loss = mean CE; dlogits = (softmax - onehot)/batch; dbias = column sums of dlogits;
 * logits [batch][nout] is required (it is also the kernel's staging buffer).
 * Calls on one device must not overlap.  reference train_ddp.py:40,198-199. */
int ddpb_softmax_xent(const float* partial, int S, const float* bias, const void* labels_i64,
                      float* logits, float* dlogits, float* dbias, float* loss, int nout,
                      int batch, void* stream);
/* dact = relu'(act) * dlogits.W ; dW = dlogits^T.act (deterministic split + reduce). */
size_t ddpb_linear_bwd_scratch_bytes(int nout, int C, int batch_cap, int H, int W);
int ddpb_linear_bwd(const void* act_planes, const float* dlogits, const float* W_fl,
                    void* dact_planes, float* dW, float* scratch, int nout, int C, int batch,
                    int batch_cap, int H, int W, void* stream);

/* ---- linear layer on the tensor cores (what the step engine uses).
 * ddpb_linear_pack turns the fp32 weight [nout][C*H*W] into bf16 UMMA tiles (forward tile as
 * hi+lo pairs, backward tile likewise); call it whenever the weight changed.
 * Forward: 128-image tiles gathered with cp.async into UMMA core matrices, K split over the SMs,
 * partial[S][batch][nout] as above.  Backward: one persistent kernel, no scratch: dX with the
 * ReLU mask fused, dW accumulated in TMEM over all image tiles (one writer per element).
 * dl_packed = the bf16 hi/lo dlogits tiles ddpb_softmax_xent_packed writes
 * ((batch_cap rounded up to 128) * 64 bytes). */
size_t ddpb_linear_pack_bytes(int nout, int C, int H, int W);
int ddpb_linear_pack(const float* W_fl, void* packed, int nout, int C, int H, int W, void* stream);
size_t ddpb_linear_fwd_tc_scratch_bytes(int nout, int C, int batch_cap, int H, int W);
int ddpb_linear_fwd_tc(const void* act_planes, const void* packed, float* partial, int* S_out,
                       int nout, int C, int batch, int batch_cap, int H, int W, void* stream);
int ddpb_softmax_xent_packed(const float* partial, int S, const float* bias, const void* labels_i64,
                             float* logits, float* dlogits, float* dbias, float* loss, int nout,
                             int batch, void* dl_packed, void* stream);
int ddpb_linear_bwd_tc(const void* act_planes, const void* dl_packed, const void* packed,
                       void* dact_planes, float* dW, int nout, int C, int batch, int batch_cap,
                       int H, int W, void* stream);
int ddpb_zero_tail_rows(void* planes, int nplanes, int batch, int batch_cap, int H, int W,
                        void* stream);

/* ---- symmetric-memory communicator (csrc/comm.cu): one process per GPU, world 1..8.
 * ddpb_comm_create cudaMallocs this rank's buffer (header + payload, zero-filled);
 * ddpb_comm_get_handle / ddpb_comm_connect exchange CUDA IPC handles (host side, any transport).
 * Offsets are byte offsets into the payload, identical on all ranks, 16-byte aligned.
 * algo: 0 by size, 1 one-shot, 2 two-shot, 4 two-shot through the NVSwitch (multimem; needs the
 * multicast communicator below).  slot: independent flag channel in [0, 8).
 * Replaces ProcessGroupNCCL + the DDP Reducer + torch.optim.SGD, reference train_ddp.py:34,199-200. */
size_t ddpb_comm_header_bytes(void);
int ddpb_comm_handle_bytes(void);
int ddpb_comm_max_world(void);   /* 8 */
int ddpb_comm_num_slots(void);   /* 8 */
int ddpb_comm_create(int rank, int world, size_t payload_bytes, void** comm_out);
int ddpb_comm_get_handle(void* comm, void* handle_out /* host */);
int ddpb_comm_connect(void* comm, const void* all_handles /* host, world * handle_bytes */);
void* ddpb_comm_local_ptr(void* comm);
void* ddpb_comm_peer_ptr(void* comm, int rank);
size_t ddpb_comm_payload_bytes(void* comm);
void ddpb_comm_set_timeout_ms(void* comm, long long ms);
int ddpb_comm_status(void* comm, unsigned int* status_out /* host */);
int ddpb_comm_destroy(void* comm);
/* VMM + multicast (NVLS) variant of the communicator: the buffer is a cuMemCreate allocation shared
 * as POSIX file descriptors (the caller carries them between the rank processes, e.g. AF_UNIX +
 * SCM_RIGHTS) and bound to one multicast object, so multimem.ld_reduce / multimem.st reduce and
 * broadcast inside the switch.  Call order on every rank: create_vmm -> export_fd, import_peer_fd for
 * every peer -> rank 0 mc_create (others mc_import with rank 0's descriptor) -> mc_add_device ->
 * [all ranks agree it succeeded] -> mc_bind -> set_connected.  Every call returns 0 or an error code
 * and never aborts, so the ranks can agree to fall back to ddpb_comm_create.
 * ddpb_multicast_supported: 1 / 0, or < 0 when the driver API cannot be resolved. */
int ddpb_multicast_supported(void);
int ddpb_comm_create_vmm(int rank, int world, size_t payload_bytes, void** comm_out);
int ddpb_comm_export_fd(void* comm, int* fd_out);
int ddpb_comm_import_peer_fd(void* comm, int rank, int fd);
int ddpb_comm_mc_create(void* comm, int* fd_out);
int ddpb_comm_mc_import(void* comm, int fd);
int ddpb_comm_mc_add_device(void* comm);
int ddpb_comm_mc_bind(void* comm);
int ddpb_comm_set_connected(void* comm);
int ddpb_comm_has_multicast(void* comm);
void* ddpb_comm_multicast_ptr(void* comm);
/* params[slice] -= lr/world * sum_over_ranks(grads[slice]) in ONE kernel (in-kernel flags). */
int ddpb_allreduce_sgd(void* comm, size_t grad_off, size_t param_off, long long n_floats,
                       const float* lr_dev, int algo, int slot, void* stream);
/* Split-K reduce + all-reduce + SGD in ONE kernel for a latency-bound bucket (n % 4 == 0):
 * grads[slice] = fixed-order sum of nparts local partial vectors of n floats, then
 * params[slice] -= lr/world * sum_over_ranks(grads[slice]).  algo 0/3 = push exchange (8-byte
 * {epoch, value} words stored straight into the peers' staging lines, one one-way NVLink latency;
 * n <= 19,456), 1 = flag round followed by peer loads (n <= 5,120).  layout 0: partial element i
 * is slice element i; layout 1 (push only): the partials are the accumulator images written by
 * ddpb_conv3x3_wgrad_partials (n = 18,944) and the slice is net.2.weight followed by net.2.bias. */
int ddpb_reduce_allreduce_sgd(void* comm, const float* partials, int nparts, long long n,
                              size_t grad_off, size_t param_off, const float* lr_dev, int algo,
                              int layout, int slot, void* stream);
/* Same; with layout 1 it can also (a) rewrite `pack` — the split bf16 tile images of net.2.weight,
 * ddpb_conv3x3_pack_split_bytes(32, 64) bytes — from the updated weights, and (b) take the net.2.bias
 * gradient from the exact-fp32 parts of ddpb_linear_bwd_f32 (bias_part, extents bp_nsplit x bp_nbx,
 * bp_hw = H*W) instead of the accumulator images' bias column.  Either pointer may be NULL. */
int ddpb_reduce_allreduce_sgd_pack(void* comm, const float* partials, int nparts, long long n,
                                   size_t grad_off, size_t param_off, const float* lr_dev, int algo,
                                   int layout, int slot, void* pack, const float* bias_part,
                                   int bp_nsplit, int bp_nbx, int bp_hw, void* stream);
/* plain sum all-reduce, dtype 0 = fp32, 1 = bf16; out_off must not overlap in_off. */
int ddpb_allreduce(void* comm, size_t in_off, size_t out_off, long long nbytes, int dtype, int algo,
                   int slot, void* stream);
int ddpb_broadcast(void* comm, size_t off, long long nbytes, int root, int slot, void* stream);

/* ---- fully connected layers of any shape on the tensor cores (csrc/gemm_tc.cu); replaces the
 * cuBLAS GEMMs behind nn.Linear (reference model.py:16, train_ddp.py:199) for shapes beyond the
 * 10-class head of the step engine (BASELINE.json config 4: fc 32768 -> 750 -> 10).
 * A matrix of `rows` x `k` fp32 values is packed once into "split row planes" (x = hi + lo, both
 * bf16; [2 sets][k/8 planes][rows][8], rows and k zero-padded to ddpb_gemm_pad() elements); the same
 * packed tensor is a K-major operand (reduce over k) or an MN-major operand (reduce over rows) of
 * tcgen05.mma, so x, W and dy are packed once each for the layer's three GEMMs.  Every K step issues
 * hi.hi + lo.hi + hi.lo ("bf16x3": above TF32 operand precision), fp32 accumulation.
 *   pack_rows_split: `packed` must be zero-filled when it is allocated; `mask` (optional, same shape
 *     and ld as src) zeroes the elements whose mask value is <= 0 (ReLU backward fused in the pack).
 *   gemm_split: out[m][n] (or out[n][m] when transposed) = sum_k A[m][k] B[n][k]; a_rows/a_k and
 *     b_rows/b_k are the shapes the operands were packed from, a_mn/b_mn = 1 reads an operand
 *     MN-major.  Supported majors: (0,0), (1,0), (1,1).  k_splits > 1 writes that many partial images
 *     split_stride floats apart (sum them, add a bias, apply ReLU with ddpb_splitk_reduce).
 *   colsum: out[c] = sum_r src[r][c] (masked like the pack): the bias gradient.  row_splits > 1 writes
 *     that many partial sums [row_splits][cols] over consecutive row ranges (sum them with a second call). */
int ddpb_gemm_pad(void);
size_t ddpb_pack_rows_split_bytes(int rows, int k);
int ddpb_pack_rows_split(const float* src, const float* mask, long long ld, int rows, int k, void* packed,
                         void* stream);
int ddpb_gemm_split(const void* a, int a_rows, int a_k, int a_mn, const void* b, int b_rows, int b_k, int b_mn,
                    int M, int N, int K, float* out, long long ld, int transposed, int k_splits,
                    long long split_stride, void* stream);
int ddpb_gemm_split_suggest_splits(int M, int N, int K);
int ddpb_splitk_reduce(const float* partial, int nsplit, long long split_stride, int rows, int cols, long long ld,
                       const float* bias, int relu, float* out, long long ld_out, void* stream);
int ddpb_colsum(const float* src, const float* mask, int rows, int cols, long long ld, float* out, int row_splits,
                void* stream);
/* ---- fused (bias +) ReLU + 2x2 max pool over NCHW fp32 (what nn.ReLU -> nn.MaxPool2d(2) compute,
 * one pass); `code` [batch][C][H/2][W/2] bytes: index of the winning input (first maximum, row-major)
 * or 4 where the pooled value is <= 0.  H even, W a multiple of 4.  No reference counterpart (its
 * model has no pooling layer, model.py:8-16); BASELINE.json names the op for the widened CNN. */
int ddpb_relu_maxpool2_fwd(const float* in, const float* bias, float* out, unsigned char* code, int batch, int C,
                           int H, int W, void* stream);
int ddpb_relu_maxpool2_bwd(const float* dout, const unsigned char* code, float* din, int batch, int C, int H, int W,
                           void* stream);

/* ---- channels-last (NHWC) fp32 CUDA-core kernels for the widened CNN's first convolution and pooling
 * (csrc/wide_ops.cu).  No reference counterpart (model.py:8-16 has neither a 3-channel input nor a pool).
 *   small_cin_fwd:   out NHWC = relu(conv3x3_pad1(x NCHW, w[cout][cin][3][3]) + bias); cin 3, cout 64, W 32.
 *   small_cin_wgrad: dw_db[cout][28] (27 taps in [ci][dy][dx] order, then the bias gradient) from dact /
 *                    act (NHWC; act = forward output used as ReLU mask, may be NULL) and x; `partial` =
 *                    scratch of ddpb_conv3x3_small_cin_wgrad_parts() * cout * 28 floats.
 *   relu_maxpool2_nhwc_*: as ddpb_relu_maxpool2_* on NHWC tensors (C a multiple of 4). */
int ddpb_conv3x3_small_cin_fwd(const float* x, const float* w, const float* bias, float* out, int cin, int cout,
                               int batch, int H, int W, void* stream);
int ddpb_conv3x3_small_cin_wgrad_parts(int batch, int H);
int ddpb_conv3x3_small_cin_wgrad(const float* dact, const float* act, const float* x, float* partial,
                                 float* dw_db, int cin, int cout, int batch, int H, int W, void* stream);
int ddpb_relu_maxpool2_nhwc_fwd(const float* in, const float* bias, float* out, unsigned char* code, int batch,
                                int C, int H, int W, void* stream);
int ddpb_relu_maxpool2_nhwc_bwd(const float* dout, const unsigned char* code, float* din, int batch, int C, int H,
                                int W, void* stream);

/* ---- linear layers with at most ddpb_small_linear_max_out() (16) outputs — a classifier head of any
 * input width — in exact fp32 on the CUDA cores: one launch forward, three backward (csrc/wide_ops.cu).
 * What nn.Linear's cuBLAS calls do for such a layer (reference model.py:16 is the 10-output case the step
 * engine specialises further).  x [batch][k], w [nout][k], y / dy [batch][nout], all row-major fp32;
 * scratch = ddpb_small_linear_bwd_scratch_floats() floats; dx and db may be NULL. */
int ddpb_small_linear_max_out(void);
int ddpb_small_linear_fwd(const float* x, const float* w, const float* bias, float* y, int batch, int k, int nout,
                          void* stream);
size_t ddpb_small_linear_bwd_scratch_floats(int batch, int k, int nout);
int ddpb_small_linear_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw, float* db,
                          float* scratch, int batch, int k, int nout, void* stream);

/* ---- p -= lr*g with lr read from device memory (single-GPU path; multi-GPU runs fuse the
 * update into the all-reduce).  reference train_ddp.py:41,200. */
int ddpb_sgd(float* p, const float* grad, const float* lr_dev, long long n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DDP_B200_H_ */
