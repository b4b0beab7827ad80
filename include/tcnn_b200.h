/* tcnn_b200.h — C ABI of libtcnn_b200.so: the B200-native (sm_100a) HashGrid -> FullyFusedMLP hot path.
 *
 * This is the drop-in boundary.  Every entry point replaces a specific piece of the reference
 * (leejaeyong7/tiny-cuda-nn; paths relative to the reference root) and is what the reference-side bindings
 * (C++ facade include/tiny-cuda-nn/*.h in this repo, bindings/torch via ctypes) call.
 *
 * Conventions:
 *   - plain C types only: opaque plan handle, caller-owned DEVICE pointers, explicit sizes, a cudaStream_t passed as void*;
 *   - every function returns 0 on success, non-zero on error; tcnn_b200_last_error() returns the thread-local message
 *     (the C++/Python facades turn it into std::runtime_error / RuntimeError like common_host.h:71-110);
 *   - nothing synchronises the stream, nothing allocates device memory behind the caller's back (the caller passes a
 *     workspace of tcnn_b200_workspace_bytes()); one plan is used from one host thread at a time;
 *   - batch sizes must be multiples of 256 (BATCH_SIZE_GRANULARITY, include/tiny-cuda-nn/common.h:235; object.h:130);
 *   - layouts: input x is [B][n_input_dims] fp32 (== GPUMatrix<float, CM> [D x B]); network output is [B][16] fp16
 *     (== CM [16 x B], src/cpp_api.cu:76); params are fp16 [n_params] = [MLP weights | grid tables]
 *     (network_with_input_encoding.h:115-130); param gradients fp16, same layout, loss-scaled.
 *   - there is NO CPU fallback: every compute entry point fails if no sm_100 device is present.
 */
#ifndef TCNN_B200_H
#define TCNN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tcnn_b200_plan tcnn_b200_plan;

/* Hyper-parameters resolved from the JSON config (defaults as in the reference: encodings/grid.h:1145-1200,
 * src/network.cu:107-113, optimizers/adam.h:312-326, src/loss.cu:86). */
typedef struct tcnn_b200_info {
	uint32_t n_input_dims;        /* D */
	uint32_t n_output_dims;       /* O (unpadded) */
	uint32_t padded_output_width; /* 16 */
	uint32_t n_levels, n_features_per_level, log2_hashmap_size, base_resolution;
	float per_level_scale;
	uint32_t encoded_width;       /* L*F padded to a multiple of 16 = MLP input width */
	uint32_t n_neurons, n_hidden_layers;
	uint64_t n_params;            /* P = n_params_mlp + n_params_grid */
	uint64_t n_params_mlp;        /* "matrix" weights (come first) */
	uint64_t n_params_grid;
	int loss_type;                /* 0 = L2, 1 = RelativeL2 */
	float learning_rate, beta1, beta2, epsilon, l2_reg, relative_decay, absolute_decay;
	uint32_t level_offset[33];    /* entry offsets per level (GridOffsetTable, grid_interface.h:84-88) */
	float level_scale[32];        /* grid_scale per level, evaluated with CUDA exp2f when a device is present */
	uint32_t level_resolution[32];
} tcnn_b200_info;

/* Error text of the last failing call on this thread. */
const char* tcnn_b200_last_error(void);
/* Library/ABI version, and the SM architecture the kernels were compiled for ("sm_100a"). */
uint32_t tcnn_b200_abi_version(void);
const char* tcnn_b200_arch(void);
/* batch_size_granularity() of cpp_api.h (256). */
uint32_t tcnn_b200_batch_size_granularity(void);

/* create_from_config / create_network_with_input_encoding (config.h:53-63, src/cpp_api.cu:147-149).
 * json_utf8 = {"encoding": {...}, "network": {...}, "loss": {...}?, "optimizer": {...}?}.
 * With "network" alone the plan is create_network (Identity encoding + network, src/cpp_api.cu:151-153); with "encoding"
 * alone it is create_encoding (src/cpp_api.cu:156-165): inference/forward then write the encoded batch, fp16
 * [B][next multiple of 8 of n_levels * n_features_per_level], and backward takes dL/d(encoded) of that shape.  The training
 * entry points need both objects.
 * A stand-alone grid encoding may carry "precision": "fp32" at the top level (create_encoding(n, encoding, Precision::Fp32),
 * cpp_api.h:71-75,115): parameters, encoded output, dL/d(encoded) and the parameter gradient of tcnn_b200_inference /
 * _forward / _backward are then fp32 (same shapes; the `*_fp16` argument names of those calls then carry fp32 buffers). */
int tcnn_b200_plan_create(const char* json_utf8, uint32_t n_input_dims, uint32_t n_output_dims, tcnn_b200_plan** plan);
void tcnn_b200_plan_destroy(tcnn_b200_plan* plan);
int tcnn_b200_plan_info(const tcnn_b200_plan* plan, tcnn_b200_info* info);
/* GridEncoding::set_max_level (encodings/grid_interface.h:101-107; default 1000 = all levels): levels with
 * level >= max_level * n_levels + 1e-3 encode to zero and receive no gradient (grid.h:69-82, 236-243).  Takes effect for
 * every later call on this plan.  The per-sample variant (set_max_level_gpu) is not implemented. */
int tcnn_b200_plan_set_max_level(tcnn_b200_plan* plan, float max_level);
/* GridEncoding::set_max_level_gpu (encodings/grid_interface.h:113-115): `max_level_per_sample` is a device array with one value
 * per sample of every batch passed from now on (sample i sees the levels l < max_level[i] * n_levels + 1e-3, grid.h:69-75); NULL
 * switches it off.  The pointer is kept, not copied.  Plans with a per-sample max level, with a hash other than CoherentPrime
 * ("hash": Prime / ReversedPrime / Rng, common_device.h:631-688) or with n_features_per_level = 1 (gradients accumulated in fp32
 * and cast, grid.h:655-666, 877-885) run general one-thread-per-(sample, level) kernels instead of the specialised ones. */
int tcnn_b200_plan_set_max_level_gpu(tcnn_b200_plan* plan, const float* max_level_per_sample);
/* Trainer::update_hyperparams (trainer.h:213-216): JSON object with optimizer keys (learning_rate, beta1, ...). */
int tcnn_b200_plan_update_optimizer(tcnn_b200_plan* plan, const char* json_utf8);

/* Scratch device memory one call on a batch of B needs (any entry point below; max over them). */
size_t tcnn_b200_workspace_bytes(const tcnn_b200_plan* plan, uint32_t batch_size);

/* Trainer::initialize_params (trainer.h:68-87) when trainer_seeding != 0 (std::seed_seq{seed} -> pcg32), or
 * cpp::Module::initialize_params (src/cpp_api.cu:124-127) when 0 (pcg32{seed}).  Writes fp32 [n_params]. */
int tcnn_b200_init_params(const tcnn_b200_plan* plan, void* stream, uint64_t seed, int trainer_seeding, float* params_fp32);
/* fp32 -> fp16 copy of the parameters (trainer.h:83-85). */
int tcnn_b200_cast_params(void* stream, uint64_t n, const float* params_fp32, void* params_fp16);

/* generate_random_uniform<float>(stream, rng, n, out, lower, upper) of random.h:58-71 with the reference's element -> draw mapping
 * (random.h:39-56), for a pcg32 generator given by its (state, inc); the caller advances its generator by n afterwards. */
int tcnn_b200_generate_random_uniform(void* stream, uint64_t rng_state, uint64_t rng_inc, uint64_t n, float* out, float lower, float upper);

/* GridEncoding forward, kernel_grid (encodings/grid.h:48-212, 732-807).  encoded: fp16 [B][encoded_width]. */
int tcnn_b200_encode_forward(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const void* grid_params_fp16, void* encoded_fp16);
/* kernel_grid_backward (grid.h:214-320, 809-900).  Adds into grid_grads_fp16 [n_params_grid]; zero_first != 0
 * clears it first (the reference's cudaMemsetAsync, grid.h:857-859).  workspace may be NULL (then the coarse levels
 * are scattered without the contention-avoiding private copies). */
int tcnn_b200_encode_backward(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const void* dL_dencoded_fp16, void* grid_grads_fp16, int zero_first, void* workspace);
/* tcnn_b200_encode_backward without a workspace and without any device allocation.  This entry point has no counterpart
 * in the reference; it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch of the scatter (the private gradient
 * copies of the coarse levels, used for batches >= 2^14) lives in a file-scope device array of
 * TCNN_B200_NOALLOC_SCRATCH_BYTES inside the library, which every GPU that loads the library holds until exit.  The call
 * itself calls no cudaMalloc / cudaFree / allocating routine; a grid whose coarse levels need more scratch than the array
 * holds is rejected with an error (use tcnn_b200_encode_backward), as is a plan created without a CUDA device or with
 * profiling enabled.  All calls share the one array: each call clears the part it uses on its stream first, and calls
 * that overlap on the GPU (different streams, different host threads) race on it - serialise them.  fp16 grid encodings. */
#define TCNN_B200_NOALLOC_SCRATCH_BYTES (2u << 20)
int tcnn_b200_encode_backward_noalloc(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const void* dL_dencoded_fp16, void* grid_grads_fp16, int zero_first);
/* Parity probe: entry index of every interpolation corner, uint32 [n_levels][B][2^D], computed by the same device
 * functions as the two kernels above (grid_index, common_device.h:690-707). */
int tcnn_b200_grid_indices(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, uint32_t* indices);

/* FullyFusedMLP::inference_mixed_precision_impl (src/fully_fused_mlp.cu:680-705).  in: fp16 [B][encoded_width];
 * out: fp16 [B][16]. */
int tcnn_b200_mlp_forward(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const void* input_fp16, const void* mlp_params_fp16, void* output_fp16);
/* FullyFusedMLP::backward_impl (src/fully_fused_mlp.cu:735-836).  The forward activations are recomputed on chip from
 * `input` instead of being re-read from HBM.  dL_dparams (fp16 [n_params_mlp], overwritten) and dL_dinput
 * (fp16 [B][encoded_width]) may each be NULL. */
int tcnn_b200_mlp_backward(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const void* input_fp16, const void* mlp_params_fp16,
                           const void* dL_doutput_fp16, void* dL_dparams_fp16, void* dL_dinput_fp16, void* workspace);

/* cpp::Module::inference (src/cpp_api.cu:72-82) / Network::inference (object.h:147-176): encoding + MLP.
 * output_fp16 [B][16] and/or output_fp32 [B][n_output_dims] (either may be NULL). */
int tcnn_b200_inference(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const void* params_fp16,
                        void* output_fp16, float* output_fp32, void* workspace);
/* cpp::Module::forward (src/cpp_api.cu:84-94): like inference, and leaves what backward needs in the workspace
 * (the encoded batch), which therefore plays the role of the reference's Context. */
int tcnn_b200_forward(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const void* params_fp16, void* output_fp16, void* workspace);
/* cpp::Module::backward (src/cpp_api.cu:96-109), GradientMode::Overwrite.  dL_dparams fp16 [n_params] may be NULL.
 * dL_dinput (fp32 [B][D], may be NULL): identity_backward (encodings/identity.h:68-83) for network-only plans, otherwise
 * kernel_grid_backward_input (grid.h:322-349) with dy/dx (grid.h:171-211) recomputed on the fly instead of stored. */
int tcnn_b200_backward(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const void* params_fp16,
                       const void* dL_doutput_fp16, void* dL_dparams_fp16, float* dL_dinput, void* workspace);
/* The same with an explicit GradientMode (object.h:115-119): 0 = Ignore (no parameter gradients), 1 = Overwrite, 2 = Accumulate:
 * the weight gradients are added to what dL_dparams holds (the reference's GEMMs with beta = 1, src/fully_fused_mlp.cu:769) and
 * the grid gradients are scattered without the memset in front (grid.h:857-859). */
int tcnn_b200_backward_mode(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const void* params_fp16,
                            const void* dL_doutput_fp16, void* dL_dparams_fp16, float* dL_dinput, void* workspace, int gradient_mode);

/* DifferentiableObject::backward_backward_input (object.h:208-240) for the stand-alone grid encoding
 * (encodings/grid.h:351-650, 902-1026): the backward pass of dL/dx = sum_k dL/dy_k * dy_k/dx.  Given dL_ddLdinput
 * (fp32 [B][D]) and the dL_doutput (fp16 [B][encoded_width]) of the first backward pass, writes any of
 *   dL_ddLdoutput fp16 [B][encoded_width]  (kernel_grid_backward_input_backward_dLdoutput),
 *   dL_dparams    fp16 [n_params], overwritten (kernel_grid_backward_input_backward_grid),
 *   dL_dinput     fp32 [B][D]               (kernel_grid_backward_input_backward_input).
 * Encoding-only plans; like the reference, the fused MLP has no second order. */
int tcnn_b200_backward_backward_input(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const void* params_fp16,
                                      const float* dL_ddLdinput, const void* dL_doutput_fp16, void* dL_ddLdoutput_fp16, void* dL_dparams_fp16,
                                      float* dL_dinput, void* workspace);

/* Trainer::training_step minus the optimizer (trainer.h:163-184): forward, loss (losses/l2.h, relative_l2.h),
 * backward.  target fp32 [B][n_output_dims]; grads fp16 [n_params]: the MLP part is overwritten, the grid part is
 * accumulated into (zero_grid_grads != 0 clears it first; tcnn_b200_adam_step re-zeroes it for the next step).
 * loss_out: device float receiving the mean loss (Trainer::loss, trainer.h:205-207) or NULL;
 * output_fp16: [B][16] or NULL. */
int tcnn_b200_training_fwd_bwd(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const float* target,
                               const void* params_fp16, void* grads_fp16, float loss_scale, int zero_grid_grads,
                               float* loss_out, void* output_fp16, void* workspace);

/* AdamOptimizer::step (optimizers/adam.h:150-188) over [MLP | grid] parameters.  grad_divisor = loss_scale *
 * world_size.  adam_step_count is the optimizer's global step after this update (only used by adabound). */
int tcnn_b200_adam_step(const tcnn_b200_plan* plan, void* stream, float grad_divisor, uint32_t adam_step_count,
                        float* params_fp32, void* params_fp16, void* grads_fp16, float* first_moments, float* second_moments,
                        uint32_t* param_steps, int zero_grid_grads);

/* tcnn_b200_adam_step restricted to parameters [param_begin, param_end) (param_begin a multiple of 4). */
int tcnn_b200_adam_step_range(const tcnn_b200_plan* plan, void* stream, float grad_divisor, uint32_t adam_step_count,
                              float* params_fp32, void* params_fp16, void* grads_fp16, float* first_moments, float* second_moments,
                              uint32_t* param_steps, int zero_grid_grads, uint64_t param_begin, uint64_t param_end);

/* Data-parallel training step over NVLink peer memory (no reference counterpart).  The fp16 parameter and gradient
 * buffers of all ranks (one process per GPU, at most 8 = one NVSwitch domain) are symmetric allocations mapped into every
 * process (e.g. torch.distributed._symmetric_memory); peer_*[r] is rank r's buffer as seen from the calling process.
 * After the local forward/backward, ONE kernel sums the gradient shard [param_begin, param_end) of this rank over all
 * ranks, applies Adam to it and stores the new fp16 parameters into every rank; both inter-GPU barriers are inside that
 * kernel (flag blocks of tcnn_b200_dp_flag_bytes() zeroed bytes per rank; `epoch` = 1, 2, 3, ... on every rank).  With
 * the multicast mappings set, reduction and broadcast go through the NVSwitch (multimem.ld_reduce / multimem.st).
 * fp32 master weights, Adam moments and step counters are passed for the rank's OWN shard only (element 0 = parameter
 * param_begin of tcnn_b200_dp_shard).  Grid gradients must be zero on entry and are zero on return. */
typedef struct tcnn_b200_dp_comm {
  uint32_t rank, world_size, epoch, reserved;
  void* peer_grads[8];
  void* peer_params[8];
  void* peer_flags[8];
  void* grads_multicast;  /* or NULL: peer loads */
  void* params_multicast; /* or NULL: peer stores */
} tcnn_b200_dp_comm;
int tcnn_b200_dp_shard(const tcnn_b200_plan* plan, uint32_t rank, uint32_t world_size, uint64_t* param_begin, uint64_t* param_end);
uint32_t tcnn_b200_dp_flag_bytes(void);
/* Ranks may enter a step minutes apart (evaluation on one rank, checkpointing, a profiler): the in-kernel waits give up only
 * after TCNN_B200_DP_TIMEOUT_S seconds (environment, default 600).  A rank that gave up does not trap; it leaves its
 * parameters untouched and stores the epoch into 32-bit word tcnn_b200_dp_flag_error_word() of its OWN flag block, which the
 * host reads (after synchronising the stream) and turns into an error. */
uint32_t tcnn_b200_dp_flag_error_word(void);
int tcnn_b200_training_step_dp(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const float* target,
                               float loss_scale, uint32_t adam_step_count, float* params_fp32_shard, float* first_moments_shard,
                               float* second_moments_shard, uint32_t* param_steps_shard, float* loss_out, void* output_fp16,
                               void* workspace, const tcnn_b200_dp_comm* comm);

/* Trainer::training_step with run_optimizer = true (trainer.h:163-190) as ONE call for a single GPU: the work of
 * tcnn_b200_training_fwd_bwd followed by tcnn_b200_adam_step, scheduled so that the Adam update of a finished group of
 * grid levels overlaps the gradient scatter of the next group (internal second stream, joined before returning to the
 * caller's stream; capturable into a CUDA graph).  The grid part of grads_fp16 must be zero on entry and is zero again
 * on exit.  Data-parallel training uses the two separate calls with the gradient all-reduce in between. */
int tcnn_b200_training_step(const tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* x, const float* target,
                            float loss_scale, uint32_t adam_step_count, float* params_fp32, void* params_fp16, void* grads_fp16,
                            float* first_moments, float* second_moments, uint32_t* param_steps, float* loss_out, void* output_fp16,
                            void* workspace);

/* Static CUDA graphs behind tcnn_b200_training_step (what include/tiny-cuda-nn/cuda_graph.h:65-153 does for the reference's
 * Trainer::training_step).  The launches of one step are captured once per (batch size, buffer addresses, loss scale, plan
 * settings) and replayed with one cudaGraphLaunch on the caller's stream; the per-step Adam scalars are patched into the
 * optimizer kernel nodes before each replay.  Up to 8 buffer sets are cached per plan; a caller that passes new buffers on
 * every call is no longer captured after 32 misses in a row.  A step issued while the CALLER's stream is being captured is
 * launched plainly so that it lands in the caller's graph.  mode: 1 = graphs, 0 = plain launches, -1 = environment
 * TCNN_B200_GRAPH ("0" disables; default graphs).  Changing the mode drops the cached graphs. */
int tcnn_b200_plan_set_graph_mode(tcnn_b200_plan* plan, int mode);
int tcnn_b200_plan_graph_stats(const tcnn_b200_plan* plan, uint32_t* captures, uint32_t* replays);


/* Batches that live in host memory (no reference counterpart: the reference's hosts create their batches on the device, e.g.
 * samples/mlp_learning_an_image.cu:240-243).  begin/end bracket ONE call of any step entry point above:
 *   begin  stages this step's batch in one of two device slots inside `staging` (skipped when an earlier begin prefetched the
 *          same host pointers as its `next_*`), starts the copy of the next batch into the other slot on an internal copy
 *          stream, makes `stream` wait for this step's data and returns device pointers for x, target and a loss word;
 *   end    marks the slot reusable once the work enqueued on `stream` so far is done and, if host_loss_out is not NULL, copies
 *          the loss word to it on a third stream (join != 0: `stream` also waits for that copy).
 * Host buffers should be pinned (cudaHostAlloc / torch pin_memory) or the copies serialise with the host.  host_loss_out is
 * valid after the caller synchronises (device, or `stream` after an end with join).  Nothing is allocated: staging is caller
 * memory of tcnn_b200_host_staging_bytes(plan, batch_size) bytes. */
size_t tcnn_b200_host_staging_bytes(const tcnn_b200_plan* plan, uint32_t batch_size);
int tcnn_b200_host_batch_begin(tcnn_b200_plan* plan, void* stream, uint32_t batch_size, const float* host_x, const float* host_target,
                               const float* next_host_x, const float* next_host_target, void* staging,
                               float** device_x, float** device_target, float** device_loss);
int tcnn_b200_host_batch_end(tcnn_b200_plan* plan, void* stream, float* host_loss_out, int join);

/* Per-kernel device timing for bench.py's roofline line (no reference counterpart): when enabled, every kernel launch
 * of this plan is bracketed by CUDA events on its launch stream.  profile_read synchronises them and returns the summed
 * milliseconds and launch counts for kind 0..4 = grid forward, MLP, MLP finalize, grid backward (+memset), Adam. */
int tcnn_b200_profile_enable(tcnn_b200_plan* plan, int enable);
int tcnn_b200_profile_read(tcnn_b200_plan* plan, float* ms_sum, uint32_t* count, uint32_t n_kinds);

#ifdef __cplusplus
}
#endif
#endif /* TCNN_B200_H */
