/*
 * dppo_b200.h - C ABI of the B200-native PPO update path (libdppo_b200.so).
 *
 * The reference (kibeomKim/Pytorch-DPPO) is pure Python on torch CPU and has no
 * FFI of its own; every entry point below replaces an inline block of the
 * reference, cited as file:line.  The Python modules in pytorch-dppo_b200/
 * (model.py, train.py, chief.py, ppo.py, utils.py - same names as the
 * reference's) bind these with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - the library owns no device memory: the caller allocates inputs, outputs
 *     and scratch (sizes from the *_scratch_bytes queries) and keeps them
 *     alive until the stream has drained;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*)
 *     and may be captured in a CUDA graph;
 *   - return value: 0 = ok, < 0 = DPPO_E_* argument error, > 0 = cudaError_t;
 *   - all arithmetic is IEEE fp32 unless stated (no fast-math), indices int32;
 *   - one host thread per device.
 */
#ifndef DPPO_B200_H
#define DPPO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DPPO_E_BADARG    (-1)   /* null pointer, negative size, unknown flavour */
#define DPPO_E_TOOLARGE  (-2)   /* a dimension beyond what the kernels are built for */
#define DPPO_E_SCRATCH   (-3)   /* scratch buffer smaller than *_scratch_bytes says */
#define DPPO_E_ALIGN     (-4)   /* pointer not aligned as documented */
#define DPPO_E_COMM      (-5)   /* communicator misuse */
#define DPPO_E_NOALLOC_LIMIT (-6) /* size beyond the built-in scratch of a *_noalloc entry point */

#define DPPO_FLAVOUR_DPPO 0     /* train.py:140-166: per-dim pdf ratio, clipped value loss */
#define DPPO_FLAVOUR_PPO  1     /* ppo.py:147-172 : log-prob ratio, plain value loss, entropy */

#define DPPO_N_PARAMS 13        /* model.py:14-22 */

/* Network shape.  Policy trunk O->H1->H2->A, value trunk O->G1->G2->1
 * (model.py:11-22 hard-codes H1=H2=100, G1=500, G2=100). */
typedef struct dppo_dims {
    int O, A, H1, H2, G1, G2;
} dppo_dims;

int         dppo_version(void);
const char* dppo_error_string(int code);

/* Flat parameter bucket.  The 13 tensors are laid out in named_parameters()
 * order (log_std, p_fc1.{weight,bias}, p_fc2.*, v_fc1.*, v_fc2.*, mu.*, v.*;
 * model.py:14-22), weights row-major (out,in) as nn.Linear stores them, each
 * tensor starting on a 16-byte boundary (padding floats stay 0).  Fills
 * offsets_host[13] / sizes_host[13] (in floats) and returns the padded length
 * P_pad, or a DPPO_E_* code. */
int dppo_param_layout(const dppo_dims* d, int* offsets_host, int* sizes_host);

/* ---- K1: GAE / returns -------------------------------------------------
 * Replaces the reverse Python loop train.py:113-122 / ppo.py:124-133 for all
 * segments of a rollout buffer at once.
 *   rewards[T], values[T]   r_i (already clipped, train.py:96) and V(s_i)
 *   seg_end[T] (uint8)      1 on the last step of a segment (done, or num_steps
 *                           reached); element T-1 is always treated as one
 *   boot[T]                 read only where seg_end: V_{i+1} there (0 if done,
 *                           else the bootstrap value, train.py:109-112)
 *   adv[T], ret[T]          A_i = td_i + (gamma*lam) A_{i+1},  R_i = A_i + V_i
 * scratch: dppo_gae_scratch_bytes(T) bytes; contents need not be initialised. */
size_t dppo_gae_scratch_bytes(int64_t T);
int dppo_gae_scan(const float* rewards, const float* values, const float* boot, const uint8_t* seg_end,
                  float* adv, float* ret, int64_t T, double gamma, double lam,
                  void* scratch, size_t scratch_bytes, void* stream);

/* dppo_gae_scan for callers that cannot allocate scratch (no counterpart in the
 * reference; added for host programs that own no device allocator or must not call
 * cudaMalloc while running).  Same inputs, outputs and results as dppo_gae_scan; the
 * look-back chain lives in a ~12 KB array inside the library image, resident on every
 * GPU that loads the library.  T <= DPPO_GAE_NOALLOC_MAX_T, otherwise
 * DPPO_E_NOALLOC_LIMIT.  The first call on a device resolves the array's device address:
 * make it before any stream capture.  Allocates and frees nothing.  The array is used
 * only when the scan has to chain its tiles (gamma*lam >= ~0.987; with a faster decay,
 * e.g. the reference's 0.99*0.95, tiles are independent and need no scratch).  All calls
 * on a device share the array (in the chained mode its ticket and tile flags are
 * re-cleared on `stream` each call): calls that may overlap on the GPU race and must be
 * serialised by the caller. */
#define DPPO_GAE_NOALLOC_MAX_T 4194304
int dppo_gae_scan_noalloc(const float* rewards, const float* values, const float* boot, const uint8_t* seg_end,
                          float* adv, float* ret, int64_t T, double gamma, double lam, void* stream);

/* ---- K4: running observation statistics --------------------------------
 * Replaces Shared_obs_stats.observes + normalize (model.py:68-80) applied row
 * by row to a rollout of T observations: row t is normalised with the
 * statistics that include rows 0..t, then clamped to [-5,5].
 *   obs[T*O] row-major raw observations, out[T*O] normalised (may alias obs)
 *   n[O], mean[O], mean_diff[O], var[O]: the four state vectors, updated in
 *   place to the state after row T-1 (n is a vector in the reference; all its
 *   entries are equal and n[0] is the one read).  Faithful while n < 2^24.
 * Prefix sums are carried in fp64, the normalisation itself is fp32. */
size_t dppo_welford_scratch_bytes(int64_t T, int O);
int dppo_welford(const float* obs, float* out, int64_t T, int O,
                 float* n, float* mean, float* mean_diff, float* var,
                 void* scratch, size_t scratch_bytes, void* stream);

/* Statistics shared by W workers (the reference has ONE Shared_obs_stats object in shared memory that every
 * worker updates, main.py:54, train.py:84-85).  Here every worker holds a replica and the rollouts of one outer
 * iteration enter in RANK ORDER: worker r's rows are normalised with the statistics of everything before this
 * iteration, plus the whole rollouts of workers 0..r-1, plus its own rows so far; afterwards every replica holds
 * the statistics of all W rollouts.  Three steps per rollout:
 *   dppo_welford_totals  this worker's (rows, S1[O], S2[O]) relative to the current mean -> totals[1+2*O] (doubles)
 *   all-gather of the W vectors in rank order (dppo_comm_allgather, or any collective) -> totals_all[W][1+2*O]
 *   dppo_welford_shared  as dppo_welford, with the other workers' totals taken into account
 * world == 1 reproduces dppo_welford exactly in the normalised rows; the final state comes from the totals. */
size_t dppo_welford_totals_scratch_bytes(int64_t T, int O);
int dppo_welford_totals(const float* obs, int64_t T, int O, const float* mean, double* totals,
                        void* scratch, size_t scratch_bytes, void* stream);
int dppo_welford_shared(const float* obs, float* out, int64_t T, int O,
                        float* n, float* mean, float* mean_diff, float* var,
                        const double* totals_all, int rank, int world,
                        void* scratch, size_t scratch_bytes, void* stream);

/* ---- K2: fused minibatch gather + actor-critic forward/backward + losses --
 * Replaces ReplayMemory.sample (train.py:31-33) + train.py:140-166 (flavour
 * DPPO) or ppo.py:147-172 (flavour PPO) + Model.forward (model.py:35-45) and
 * the autograd backward.  One call = one minibatch ("epoch" in the
 * reference's wording).
 *   obs[T*O], act[T*A], adv[T], ret[T], logp_old[T] (PPO flavour only)
 *                         the rollout buffer, struct-of-arrays
 *   idx[B]                positions to gather (Python random.sample order)
 *   theta, theta_old      flat parameter buckets (theta_old: DPPO flavour only,
 *                         the lagging copy train.py:128-129,164)
 *   partials              OUT [nparts][P_pad] per-CTA gradient partial sums,
 *                         nparts = dppo_update_nparts(B); their sum in index
 *                         order is the gradient (done by dppo_reduce_adam)
 *   loss_partials         OUT [nparts][4] per-CTA sums of the loss terms
 *                         (clip, value, entropy, unused), already scaled
 * Deterministic: no atomics, fixed reduction order. */
/* Large minibatches of nets whose hidden layers all run on the tensor-core routines take two launches: the update
 * kernel leaves every layer's input and dz feature-major in an internal per-device buffer, a second pass forms all
 * weight gradients from it as batch-level GEMMs over a few ranges of samples - dppo_update_nparts is then that number
 * of ranges (<= 8) instead of one partial per CTA.  The buffer is allocated by dppo_update_prepare (or by the first plain call; a call that
 * finds it missing while its stream is being captured returns DPPO_E_SCRATCH); it is shared by all calls on a device,
 * which must therefore be ordered on one stream, and it only grows (captured graphs keep pointing at the buffer they
 * were captured with).  DPPO_K2_DWG=0 in the environment or dppo_update_set_batch_dw(0) keeps the single-launch form -
 * the choice changes dppo_update_nparts, so make it before buffers are sized.  No reference counterpart (autograd forms dW inside loss.backward(), train.py:162). */
int    dppo_update_prepare(const dppo_dims* d, int B);
int    dppo_update_set_batch_dw(int on);                  /* returns the previous setting */
int    dppo_update_launches(const dppo_dims* d, int B);   /* kernels one dppo_update_fused call launches: 1 or 2 */
int    dppo_update_nparts(const dppo_dims* d, int B);
size_t dppo_update_partials_bytes(const dppo_dims* d, int B);
int dppo_update_fused(int flavour, const dppo_dims* d,
                      const float* obs, const float* act, const float* adv, const float* ret,
                      const float* logp_old, const int32_t* idx, int B,
                      const float* theta, const float* theta_old,
                      float clip, float ent_coeff,
                      float* partials, float* loss_partials, void* stream);
/* ---- the rollout's forward: batched Model.forward + action sampling ------------------------------------------
 * Replaces, for N environments stepped together, the per-step block of the rollout loop train.py:84-90 /
 * ppo.py:90-98 after the observation has been normalised (K4): ``mu, sigma_sq, v = model(state)``,
 * ``action = mu + sigma_sq.sqrt()*eps`` (train.py:89; std_is_variance != 0) or ``mu + sigma_sq*eps`` (ppo.py:93), and
 * the log-probability ppo.py:95-98.  Same kernel and layer routines as dppo_update_fused, forward only.
 *   states[N*O] normalised observations, eps[N*A] standard-normal draws (NULL: action = mu)
 *   outputs (each nullable, at least one given): mu_out[.*A], v_out[.], act_out[.*A], logp_out[.], states_out[.*O];
 *   batch row r is written to row  out_row0 + r*out_row_stride  of every output - with the columns of a rollout
 *   buffer, stride = steps per environment and row0 = the step, the rows land where K1/K2 read them. */
int dppo_policy_forward(const dppo_dims* d, const float* states, int N, const float* theta, const float* eps,
                        int std_is_variance, float* mu_out, float* v_out, float* act_out, float* logp_out,
                        float* states_out, int64_t out_row0, int64_t out_row_stride, void* stream);

/* Which layer routines dppo_update_fused uses: DPPO_K2_PATH_AUTO (default) puts every eligible layer on the
 * tensor cores (tcgen05, 3xTF32) and the rest on the FP32 pipe; DPPO_K2_PATH_FP32 forces the FP32-pipe
 * routines for all layers (A/B runs, tests).  Process-wide; the environment variable DPPO_K2_NO_TS=1, read
 * once when the library is first used, selects DPPO_K2_PATH_FP32 as the initial value.  Returns the
 * previous setting, or DPPO_E_BADARG. */
#define DPPO_K2_PATH_AUTO 0
#define DPPO_K2_PATH_FP32 1
int dppo_update_set_path(int path);

/* ---- K3: gradient sum + [global-norm clip] + Adam ------------------------
 * Replaces Shared_grad_buffers.add_gradient/reset (model.py:53-59), the
 * chief's Adam step on the summed gradient (chief.py:13-20), and for the PPO
 * flavour clip_grad_norm + optimizer.step (ppo.py:173-174).
 *   G = gshared + sum_parts partials[p]      (fixed order; gshared may be NULL)
 *   if max_grad_norm > 0: G *= min(1, max_grad_norm / (||G||_2 + 1e-6))
 *   Adam exactly as torch 2.11 _single_tensor_adam (lerp / addcmul / addcdiv)
 *   with the 1-based update count `step`; gshared is reset to 0 (chief.py:19).
 *   theta_prev (nullable) receives theta before the update (the next epoch's
 *   theta_old); grad_out (nullable) receives G (after clipping);
 *   loss_partials/loss_out (nullable): loss_out[4] = sum of the K2 partials.
 *   scalars_dev (nullable): 6 floats in DEVICE memory as written by
 *   dppo_adam_scalars_host for the update count / learning rate to use; when
 *   given, step/lr/beta/eps arguments are ignored.  This lets a captured CUDA
 *   graph be replayed with a new update count (the host refreshes the 24 bytes).
 * scratch: dppo_reduce_adam_scratch_bytes(P_pad) (only needed when clipping). */
size_t dppo_reduce_adam_scratch_bytes(int P_pad);
int dppo_adam_scalars_host(int step, double lr, double beta1, double beta2, double eps, float* out6_host);
int dppo_reduce_adam(const float* partials, int nparts, float* gshared,
                     float* theta, float* theta_prev, float* m, float* v, int P_pad,
                     int step, double lr, double beta1, double beta2, double eps,
                     const float* scalars_dev, double max_grad_norm, float* grad_out,
                     const float* loss_partials, float* loss_out,
                     void* scratch, size_t scratch_bytes, void* stream);

/* ---- K3 across GPUs: gradient SUM over workers + Adam in one kernel over NVLink -------
 * Replaces, for W workers on W GPUs of one node, the shared-memory gradient bucket that
 * every worker adds into (model.py:53-55), the Counter/TrafficLight barrier (utils.py:4-39,
 * train.py:171-175, chief.py:9-13) and the chief's Adam step + reset (chief.py:15-20).
 * One process per GPU.  Every rank calls dppo_allreduce_adam once per epoch with its own K2
 * partials; the kernel pushes the rank's summed slice into every peer's receive buffer
 * (peer-mapped memory), waits for all ranks' slices and applies
 *     G = gshared + sum_{r=0..W-1} g_r   (rank order),  Adam(G),  gshared = 0
 * so all replicas stay bit-identical.  Arguments as dppo_reduce_adam (no norm clipping).
 * Life cycle: dppo_comm_create allocates the peer-visible buffer (the only allocation the
 * library makes), dppo_comm_local_handle exports it (dppo_comm_handle_bytes() bytes, to be
 * all-gathered by the host, e.g. torch.distributed), dppo_comm_open_peers maps the others.
 * The *_loopback pair runs W virtual ranks on ONE GPU in a cooperative launch (tests).
 * dppo_reduce_partials only sums the K2 partials (used by the NCCL baseline path). */
typedef struct dppo_comm dppo_comm;
int dppo_comm_create(int rank, int world, int P_pad, dppo_comm** out);
int dppo_comm_handle_bytes(void);
int dppo_comm_local_handle(dppo_comm* c, void* handle_host);
int dppo_comm_open_peers(dppo_comm* c, const void* handles_host);
int dppo_comm_create_loopback(int world, int P_pad, dppo_comm** out_array);
int dppo_comm_destroy(dppo_comm* c);
/* Exchange algorithm of dppo_allreduce_adam: PUSH_ALL sends every rank's bucket to every peer (one NVLink hop; default
 * for W <= 4), OWNER sends each column block to one owning rank that adds the W slices in rank order and sends the sum
 * back to all (two hops, a quarter of the bytes at W = 8; default for W > 4).  Same results bit for bit.  All ranks must
 * choose the same algorithm, between epochs. */
#define DPPO_COMM_ALGO_PUSH_ALL 1
#define DPPO_COMM_ALGO_OWNER    2
int dppo_comm_set_algorithm(dppo_comm* c, int algo);
int dppo_comm_get_algorithm(dppo_comm* c);
/* A rank that waits for peers which never push (a dead process, unequal numbers of calls) gives up after a few
 * seconds of polling and records it; dppo_comm_check returns DPPO_E_COMM from then on (0 otherwise).  It
 * synchronises the device.  The reference has no such detection: a dead worker deadlocks the chief
 * (chief.py:13) and the other workers (train.py:174-175). */
int dppo_comm_check(dppo_comm* c);
int dppo_allreduce_adam(dppo_comm* c, const float* partials, int nparts, float* gshared,
                        float* theta, float* theta_prev, float* m, float* v, int P_pad,
                        int step, double lr, double beta1, double beta2, double eps,
                        const float* scalars_dev, float* grad_out,
                        const float* loss_partials, float* loss_out, void* stream);
int dppo_allreduce_adam_loopback(dppo_comm** comms, int world, const float* const* partials, const int* nparts,
                                 float* const* gshared, float* const* theta, float* const* theta_prev,
                                 float* const* m, float* const* v, int P_pad,
                                 int step, double lr, double beta1, double beta2, double eps,
                                 float* const* grad_out, void* stream);
int dppo_reduce_partials(const float* partials, int nparts, float* out, int P_pad, void* stream);
/* Small all-gather over the same peer channel (no NCCL): every rank contributes `count` doubles
 * (count <= 8193), dst[W][count] receives all ranks' vectors in rank order.  Every rank must call it the same
 * number of times.  Used for the shared observation statistics above (Shared_obs_stats in shared memory,
 * main.py:54, model.py:61-66). */
int dppo_comm_allgather(dppo_comm* c, const double* src, double* dst, int count, void* stream);
int dppo_comm_allgather_loopback(dppo_comm** comms, int world, const double* const* src, double* const* dst,
                                 int count, void* stream);

/* ---- a4: minibatch positions (HOST function, no GPU work) ------------------
 * Replaces ``random.sample(self.memory, batch_size)`` (train.py:31-33 /
 * ppo.py:58-60): `rounds` consecutive samples of k distinct positions out of
 * range(n), bit-identical to CPython 3.12 random.sample driven by the same
 * Mersenne Twister.  mt_state624_host / mt_index_host are the 624 state words
 * and the position taken from ``random.getstate()[1]``; both are advanced in
 * place so the caller can ``random.setstate`` them back.  out_host[rounds*k]. */
int dppo_sample_indices_host(uint32_t* mt_state624_host, int* mt_index_host, int64_t n, int k, int rounds,
                             int32_t* out_host);

/* The same on the DEVICE: the Mersenne Twister state (624 words) and position live in device
 * memory, the positions are written to device memory where K2 reads them - no index block crosses PCIe.  Bit-identical
 * to random.sample for both of CPython's branches (pool / set), state and position advanced in place exactly as
 * CPython would leave them.  scratch: dppo_sample_scratch_bytes(n, k) bytes.  One CTA; asynchronous on `stream`. */
size_t dppo_sample_scratch_bytes(int64_t n, int k);
int dppo_sample_indices(uint32_t* mt_state624, int* mt_index, int64_t n, int k, int rounds, int32_t* out,
                        void* scratch, size_t scratch_bytes, void* stream);

/* ---- bench utility: FP32 FMA throughput probe (no reference counterpart) ---
 * Launches 8 CTAs per SM of dependent-chain FMAs; *flops_host = FLOPs issued. */
int dppo_fp32_probe(int iters, float* out, double* flops_host, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPPO_B200_H */
