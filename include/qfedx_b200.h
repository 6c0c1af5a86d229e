/* qfedx_b200 C ABI - hand-written sm_100a kernels behind the Python package.
 *
 * Every entry point returns 0 on success, a positive cudaError_t value if a CUDA call
 * failed, or a negative QFX_ERR_* code for a rejected argument; qfx_last_error() gives the
 * message.  All pointers are device pointers unless a comment says otherwise.  Kernels are
 * enqueued on `stream` and do not synchronise.  The library is built by
 * `python -m qfedx_b200.build` into qfedx_b200/lib/libqfedx_b200.so.
 *
 * Citations name the part of the reference (/root/reference) or of its ROADMAP that the
 * entry point provides; the reference itself contains no native code.
 */
#ifndef QFEDX_B200_H
#define QFEDX_B200_H

#include <cuda_runtime_api.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QFX_ERR_BAD_ARG (-1)
#define QFX_ERR_UNSUPPORTED (-2)
#define QFX_ERR_TOO_LARGE (-3)

/* gate kinds of the large-state engine (K5/K6) */
#define QFX_GATE_RX 0
#define QFX_GATE_RY 1
#define QFX_GATE_RZ 2
#define QFX_GATE_CNOT 3
#define QFX_GATE_CZ 4
#define QFX_GATE_RXRZ 5        /* RZ(p1) * RX(p0), produced by the compiler's fusion */
#define QFX_GATE_RY_FEATURE 6  /* RY(angles[sample][p0]): the feature map */

/* basis: 0 rx, 1 ry, 2 rz (feature-map rotation, /root/reference/src/QFed/qAngle.py:43-50)
 * entangler: 0 CNOT ring, 1 CZ ring, 2 none (/root/reference/ROADMAP.md:127) */

const char* qfx_last_error(void);
/* out[j] = keyed Feistel permutation of [0, n) (the in-kernel sample shuffle); test hook, host + device agree. */
int qfx_debug_feistel(int n, unsigned long long key, int* out_dev, cudaStream_t stream);
int qfx_version(void);

/* K1+K3: feature map (angles [B,n]) or amplitude-encoded real state (init_state [B,2^n]),
 * L ansatz layers theta [G,L,n,2] (gid [B] picks the group, NULL = group 0), <Z_q> for every
 * qubit -> expvals [B,n] (nullable).  state_out [B,2^n] complex64 interleaved (nullable).
 * n <= 10.  Ansatz: /root/reference/ROADMAP.md:126-128; encoders: src/QFed/qAngle.py:27-51,
 * src/QFed/qAmplitude.py:25-46. */
int qfx_sv_forward(int n_qubits, int n_layers, int basis, int entangler, const float* angles,
                   const float* init_state, const float* theta, const int* gid, int batch,
                   float* expvals, float* state_out, cudaStream_t stream);

/* K4a: adjoint differentiation of sum_q cot[b,q] <Z_q>; ACCUMULATES into grad_theta [G,L,n,2]
 * (zero it first); grad_angles [B,n] nullable.  /root/reference/ROADMAP.md:23,27. */
int qfx_sv_adjoint_backward(int n_qubits, int n_layers, int basis, int entangler, const float* angles,
                            const float* init_state, const float* theta, const int* gid, int batch,
                            const float* cot, float* grad_theta, float* grad_angles, cudaStream_t stream);

/* K4b: the same gradient by the two-term parameter-shift rule, 2P shifted circuits per sample
 * as an extra batch axis; ACCUMULATES into grad_theta.  /root/reference/ROADMAP.md:130-135. */
int qfx_sv_param_shift(int n_qubits, int n_layers, int basis, int entangler, const float* angles,
                       const float* init_state, const float* theta, const int* gid, int batch,
                       const float* cot, float* grad_theta, cudaStream_t stream);

/* Whole-client local training in one launch: one thread-block cluster (cluster_size 1..8 CTAs)
 * per selected client walks `epochs` x ceil(n_k/batch_size) minibatches with forward, loss,
 * adjoint (grad_mode 0) or parameter-shift (grad_mode 1) gradients and the optimizer
 * (0 SGD-momentum: beta1 = momentum; 1 Adam) on-chip.  Parameters are the flat vector
 * [theta (2nL) | a (C) | b (C)], logits_j = a_j <Z_j> + b_j, C = 1 -> BCE, C > 1 -> softmax CE.
 * angles [N,n], labels [N], offsets [K+1], perm [epochs,N] (nullable), client_ids [n_selected],
 * params_out [n_selected, D], stats_out [n_selected, 2] (mean loss, accuracy of the last epoch).
 * lane_mode 0 keeps most amplitudes per lane (throughput), 1 spreads a circuit over most lanes (latency).
 * Sample order per epoch: perm if given; else a keyed Feistel permutation when shuffle_key != 0; else sequential.
 * Counterpart of client_update, /root/reference/src/CFed/Classical_FL.py:40-64, for the VQC of
 * /root/reference/ROADMAP.md:29-43. */
int qfx_vqc_local_train(int n_qubits, int n_layers, int basis, int entangler, int n_classes,
                        const float* angles, const int* labels, const int* offsets, const int* perm,
                        int n_total, const float* params_in, const int* client_ids, int n_selected,
                        float* params_out, float* stats_out, int epochs, int batch_size, int optimizer,
                        float lr, float beta1, float beta2, float eps, int grad_mode, int cluster_size,
                        int lane_mode, unsigned long long shuffle_key, cudaStream_t stream);

/* Phase timers of the VQC training kernel: n_selected x 12 int64 cycle counters (thread 0 of each cluster's first
 * CTA; caller-zeroed) accumulated by the following launches on this device; null switches them off (default). */
int qfx_vqc_train_profile(long long* prof);

/* Evaluation in one launch: out2[0] = sum of losses, out2[1] = number correct.
 * Counterpart of evaluate_model, /root/reference/src/CFed/Classical_FL.py:83-102. */
int qfx_vqc_eval(int n_qubits, int n_layers, int basis, int entangler, int n_classes, const float* angles,
                 const int* labels, const float* params, int n_samples, float* out2, cudaStream_t stream);

/* Noisy variants (noise models of /root/reference/ROADMAP.md:64-73 in the K3 epilogue): the exact <Z> of every logit
 * qubit goes through E' = noise_alpha * E + noise_beta (depolarizing on the measured qubit, amplitude damping, readout
 * confusion - all affine on <Z>) and, when shots > 0, is replaced by the mean of `shots` sampled +-1 outcomes (counter RNG
 * keyed by noise_key, epoch, sample, qubit); gradients flow through the affine part.  alpha 1 / beta 0 / shots 0 = exact. */
int qfx_vqc_local_train_noisy(int n_qubits, int n_layers, int basis, int entangler, int n_classes,
                              const float* angles, const int* labels, const int* offsets, const int* perm,
                              int n_total, const float* params_in, const int* client_ids, int n_selected,
                              float* params_out, float* stats_out, int epochs, int batch_size, int optimizer,
                              float lr, float beta1, float beta2, float eps, int grad_mode, int cluster_size,
                              int lane_mode, unsigned long long shuffle_key, float noise_alpha, float noise_beta,
                              int shots, unsigned long long noise_key, cudaStream_t stream);
int qfx_vqc_eval_noisy(int n_qubits, int n_layers, int basis, int entangler, int n_classes, const float* angles,
                       const int* labels, const float* params, int n_samples, float* out2, float noise_alpha,
                       float noise_beta, int shots, unsigned long long noise_key, cudaStream_t stream);
/* the same readout noise as a stand-alone pass over exact expectation values [batch, n_qubits] (in place allowed) */
int qfx_readout_noise(const float* expvals, float* out, int batch, int n_qubits, float noise_alpha,
                      float noise_beta, int shots, unsigned long long noise_key, cudaStream_t stream);

/* K7: fused FedAvg + per-client L2 clip + Gaussian noise + pairwise-mask secure aggregation +
 * server update, one cooperative launch.  client_params [K,D]; weights [K] sample counts;
 * participate [K] 0/1 (nullable); mode 0 = average weights (the reference's federated_averaging,
 * /root/reference/src/CFed/Classical_FL.py:66-81), 1 = average deltas theta_k - theta_g with the
 * first n_wrap entries wrapped to [-pi,pi) (/root/reference/ROADMAP.md:37); clip <= 0 disables
 * clipping; sigma is the noise multiplier (std = sigma*clip per client, ROADMAP.md:50-51);
 * secagg != 0 sums fixed-point messages carrying pairwise masks (ROADMAP.md:52-55).
 * norms_out [K] and masked_out [K,D] are nullable; scratch is K floats of workspace. */
int qfx_fedavg_fused(const float* client_params, const float* global_in, float* global_out,
                     const float* weights, const int* participate, int n_clients, int dim, int mode,
                     int n_wrap, float clip, float sigma, float server_lr, unsigned long long seed, int round,
                     int secagg, float* norms_out, long long* masked_out, float* scratch, cudaStream_t stream);

/* Allocation-free variant of qfx_fedavg_fused.  It has no counterpart in the reference; it exists for
 * callers that cannot provide (or allocate) the workspace: C/C++ servers using this ABI without a
 * tensor library's allocator, processes whose device memory is owned by another framework's arena,
 * and latency-critical loops where a cudaMalloc/cudaFree would synchronise the device.  Arguments are
 * those of qfx_fedavg_fused without `scratch`; the norm accumulators live in a static device array of
 * QFX_FEDAVG_STATIC_MAX_CLIENTS floats (16 KiB held on every GPU that loads the library).  More clients
 * than that -> QFX_ERR_TOO_LARGE.  The first call on a device resolves the array's address: make it at
 * start-up, before any stream capture.  Calls share the array: consecutive calls on one stream are
 * safe (the used part is cleared first), calls that overlap on one GPU race. */
#define QFX_FEDAVG_STATIC_MAX_CLIENTS 4096
int qfx_fedavg_fused_static(const float* client_params, const float* global_in, float* global_out,
                            const float* weights, const int* participate, int n_clients, int dim, int mode,
                            int n_wrap, float clip, float sigma, float server_lr, unsigned long long seed,
                            int round, int secagg, float* norms_out, long long* masked_out, cudaStream_t stream);

/* K7 across GPUs: this rank's k_local clients (global ids client_gids) are reduced locally,
 * the partial sum is pushed into slot [rank] of every rank's receive buffer with peer stores over
 * NVLink, flags are released/acquired at system scope, and every rank applies the same update.
 * recv_slots[r] / flags[r] are device-visible pointers to rank r's buffers ([world][dim] int64 and
 * [world] u32), each slot array 16-byte aligned.  epoch must increase every call; alternate two slot sets
 * between calls. */
int qfx_fedavg_fused_multi(const float* client_params, const float* global_in, float* global_out,
                           const float* weights, const int* participate, const int* client_gids, int k_local,
                           int k_global, int dim, int mode, int n_wrap, float clip, float sigma, float server_lr,
                           unsigned long long seed, int round, int secagg, float* norms_out, float* scratch,
                           int rank, int world, long long* const* recv_slots, unsigned int* const* flags,
                           unsigned int epoch, cudaStream_t stream);

/* Symmetric heap (no reference counterpart - the reference has no communication layer):
 * allocate + zero a device buffer and export its 64-byte CUDA IPC handle; map a peer's handle; unmap;
 * free.  Handles are exchanged by the caller (torch.distributed object all-gather). */
int qfx_symm_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int qfx_symm_open(const unsigned char* handle64, void** ptr);
int qfx_symm_close(void* ptr);
int qfx_symm_free(void* ptr);
int qfx_enable_peer(int dev, int peer);

/* K7 across GPUs with an explicit exchange algorithm.  two_shot 0: every rank pushes its partial sum into every peer's
 * slot and reduces W slots locally (one cross-GPU barrier; recv_slots: world x dim 64-bit words per rank) - the choice for
 * latency-bound vectors.  two_shot 1: reduce-scatter by peer stores during the sweep over the client rows (slice o of the
 * partial goes straight into rank o's buffer), local reduce + update of the own slice, then every rank gathers the other
 * slices by peer loads (two barriers; recv_slots: (world + 1) * (ceil(dim / world) rounded up to 4) words per rank):
 * 2 (W-1)/W D instead of (W-1) D words over NVLink per rank.  pair_keys nullable; flags >= 32 words per rank. */
int qfx_fedavg_fused_multi_ex(const float* client_params, const float* global_in, float* global_out,
                              const float* weights, const int* participate, const int* client_gids,
                              int k_local, int k_global, int dim, int mode, int n_wrap, float clip,
                              float sigma, float server_lr, unsigned long long seed, int round, int secagg,
                              const unsigned long long* pair_keys, float* norms_out, float* scratch,
                              int rank, int world, long long* const* recv_slots,
                              unsigned int* const* flags, unsigned int epoch, int two_shot, cudaStream_t stream);

/* K7 across GPUs through the NVSwitch reduction (NVLS): every rank's partial stays in its own row `partial` of a buffer
 * bound to a multicast object over the job's GPUs; the owner of a slice reads the sum of all W rows with
 * multimem.ld_reduce on `partial_mc` (the multicast address of the same row), applies the server update and broadcasts
 * the slice with multimem.st on `out_mc`.  partial: dim words of 4 bytes (8 with secagg); out: dim floats; global_out
 * may be `out` itself (no copy).  flags as for qfx_fedavg_fused_multi_ex. */
int qfx_fedavg_fused_multi_nvls(const float* client_params, const float* global_in, float* global_out,
                                const float* weights, const int* participate, const int* client_gids,
                                int k_local, int k_global, int dim, int mode, int n_wrap, float clip,
                                float sigma, float server_lr, unsigned long long seed, int round, int secagg,
                                const unsigned long long* pair_keys, float* norms_out, float* scratch,
                                int rank, int world, void* partial, const void* partial_mc, float* out,
                                float* out_mc, unsigned int* const* flags, unsigned int epoch, cudaStream_t stream);

/* Loopback world for tests on a single GPU: `world` logical ranks of K7's multi-GPU path in ONE cooperative launch
 * (blockIdx.y = rank; every "rank's" receive slots and flag words live in this GPU's memory).  client_params, client_gids,
 * k_local, global_out and scratch are HOST arrays with one entry per rank; args_dev is world * 512 bytes of device scratch. */
int qfx_fedavg_fused_loopback(int world, const float* const* client_params, const int* const* client_gids,
                              const int* k_local, float* const* global_out, float* const* scratch,
                              const float* global_in, const float* weights, const int* participate,
                              int k_global, int dim, int mode, int n_wrap, float clip, float sigma,
                              float server_lr, unsigned long long seed, int round, int secagg,
                              const unsigned long long* pair_keys, long long* const* recv_slots,
                              unsigned int* const* flags, unsigned int epoch, int two_shot, void* args_dev,
                              cudaStream_t stream);

/* K7 with per-pair mask secrets: pair_keys [k_global, k_global] holds at [lo * k_global + hi] the 64-bit secret clients
 * lo < hi agreed on at registration (qfedx_b200/fl/secagg.py: KeyAgreement - the "seed exchange API" planned in
 * /root/reference/ROADMAP.md:52-55); each pair's mask stream is keyed by its secret instead of the job seed. */
int qfx_fedavg_fused_pairkeys(const float* client_params, const float* global_in, float* global_out,
                              const float* weights, const int* participate, int n_clients, int dim, int mode,
                              int n_wrap, float clip, float sigma, float server_lr, unsigned long long seed,
                              int round, const unsigned long long* pair_keys, float* norms_out,
                              long long* masked_out, float* scratch, cudaStream_t stream);
int qfx_fedavg_fused_multi_pairkeys(const float* client_params, const float* global_in, float* global_out,
                                    const float* weights, const int* participate, const int* client_gids,
                                    int k_local, int k_global, int dim, int mode, int n_wrap, float clip,
                                    float sigma, float server_lr, unsigned long long seed, int round,
                                    const unsigned long long* pair_keys, float* norms_out, float* scratch,
                                    int rank, int world, long long* const* recv_slots,
                                    unsigned int* const* flags, unsigned int epoch, cudaStream_t stream);

/* K8: TinyCNN (the reference's model, /root/reference/src/CFed/Classical_FL.py:21-38) trained for
 * all selected clients at once.  One call = one SGD-momentum step of client_update
 * (/root/reference/src/CFed/Classical_FL.py:56-62) for every client: 8 kernels on `stream`, meant to
 * be captured in a CUDA graph.  params/moment are [n_selected, nd] rows in state_dict order;
 * batch_size <= 32; clients whose data is exhausted at (epoch, batch_index) are skipped.
 * round keys the dropout masks; round_dev (nullable device int) overrides it so a captured graph can be
 * replayed for later rounds.  conv2_mode 0: fp32 CUDA-core conv2 kernels; 1: tcgen05 bf16 implicit-GEMM conv2 (fp32 accumulate, fp32 master weights).
 * workspace: qfx_tinycnn_workspace_bytes; its tail (qfx_tinycnn_stats_offset) holds per-client
 * (loss sum, #correct) accumulators that the caller zeroes per epoch. */
int qfx_tinycnn_workspace_bytes(int n_selected, size_t* bytes);
int qfx_tinycnn_stats_offset(int n_selected, size_t* offset);
/* once per round, after the global parameters were copied into every row of `params` */
int qfx_tinycnn_begin_round(const float* params, int nd, int n_selected, void* workspace, cudaStream_t stream);
int qfx_tinycnn_train_step(const float* X, const int* labels, const int* offsets, const int* perm, int n_total,
                           const int* client_ids, int n_selected, int epoch, int batch_index, int batch_size,
                           float* params, float* moment, int nd, int n_classes, float lr, float momentum,
                           float p_drop, unsigned long long seed, int round, const int* round_dev,
                           void* workspace, int conv2_mode, unsigned long long shuffle_key, cudaStream_t stream);
/* Same step, pipelined across steps: the conv2 parameter update stays on the library's side stream behind the conv2
 * weight gradient, and the NEXT call's conv1 forward runs beside it (conv1 output alternates between two buffers);
 * the owed work is joined in front of the next conv2 forward.  Call qfx_tinycnn_finish_steps(stream) after the last
 * step of a sequence - before reading parameters or statistics, and before ending a stream capture. */
int qfx_tinycnn_train_step_overlapped(const float* X, const int* labels, const int* offsets, const int* perm, int n_total,
                           const int* client_ids, int n_selected, int epoch, int batch_index, int batch_size,
                           float* params, float* moment, int nd, int n_classes, float lr, float momentum,
                           float p_drop, unsigned long long seed, int round, const int* round_dev,
                           void* workspace, int conv2_mode, unsigned long long shuffle_key, cudaStream_t stream);
int qfx_tinycnn_finish_steps(cudaStream_t stream);
/* Evaluation (evaluate_model, /root/reference/src/CFed/Classical_FL.py:83-102) in three launches:
 * out2 = (loss sum, #correct); scratch needs n * 23,520 + 102,400 bytes. */
int qfx_tinycnn_eval(const float* X, const int* labels, int n, const float* params, int n_classes,
                     void* scratch, float* out2, cudaStream_t stream);

/* K8p: the same local training as ONE persistent launch per round (csrc/tinycnn_persist.cu): G CTAs per client stay
 * resident, walk epochs x minibatches themselves and meet at five group barriers per step; conv2 forward / data
 * gradient / weight gradient run on tcgen05 with a two-term bf16 split of both operands (3 MMAs, fp32 accumulate:
 * products exact to ~2^-16).  Restates client_update, /root/reference/src/CFed/Classical_FL.py:40-64.
 * params [K,nd] = global model in every row, moment [K,nd] = 0 on entry; workspace from
 * qfx_tinycnn_round_workspace_bytes (its tail at *stats_offset holds the last epoch's (loss sum, #correct) per client).
 * ctas_per_client 0 = min(32, #SM / K).  qfx_tinycnn_round_status reports a group barrier that gave up waiting. */
int qfx_tinycnn_round_workspace_bytes(int n_selected, size_t* bytes, size_t* stats_offset);
int qfx_tinycnn_train_round(const float* X, const int* labels, const int* offsets, const int* perm, int n_total,
                            const int* client_ids, int n_selected, int epochs, int n_batches, int batch_size,
                            float* params, float* moment, int nd, int n_classes, float lr, float momentum,
                            float p_drop, unsigned long long seed, int round, const int* round_dev,
                            void* workspace, unsigned long long shuffle_key, int ctas_per_client, cudaStream_t stream);
int qfx_tinycnn_round_status(int n_selected, void* workspace, cudaStream_t stream, int* timed_out);
/* evaluation with conv2 on the same tensor-core path and precision as qfx_tinycnn_train_round;
 * scratch: 102,400 + n * 7,840 bytes */
int qfx_tinycnn_eval_tc(const float* X, const int* labels, int n, const float* params, int n_classes,
                        void* scratch, float* out2, cudaStream_t stream);

/* K2: batched circuit unitaries as a complex GEMM on tcgen05 / TMEM fed by TMA (no reference
 * counterpart; accelerates the planned VQC forward, /root/reference/ROADMAP.md:126-128).
 *   qfx_unitary_pack: U^T as complex64 rows (row b = U|b>, e.g. qfx_sv_forward's state_out for the
 *     2^n basis states) -> real embedding [[Ur,-Ui],[Ui,Ur]] split into tf32 hi / lo parts,
 *     each [2*2^n, 2*2^n] fp32.
 *   qfx_unitary_pack_product_states: feature-map angles [B,n] -> B operand rows [re | im], hi / lo,
 *     [batch_padded, 2*2^n] (batch_padded multiple of 64, padding rows zero).
 *   qfx_unitary_split_states: arbitrary input states [B, 2*2^n] -> hi / lo.
 *   qfx_unitary_gemm: out_states [batch_padded, 2*2^n] (nullable) and/or <Z_q> expvals
 *     [batch_padded, n] (nullable, fused in the epilogue); precision 3 = 3xTF32, 1 = TF32. */
int qfx_unitary_pack(const float* u_transposed_c64, int n_qubits, float* a_hi, float* a_lo, cudaStream_t stream);
int qfx_unitary_pack_product_states(const float* angles, int n_qubits, int basis, int batch, int batch_padded,
                                    float* b_hi, float* b_lo, cudaStream_t stream);
int qfx_unitary_split_states(const float* states, int n_qubits, int batch, int batch_padded, float* b_hi,
                             float* b_lo, cudaStream_t stream);
int qfx_unitary_gemm(const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, int n_qubits,
                     int batch_padded, int precision, float* out_states, float* expvals, cudaStream_t stream);

/* K5/K6 large-state engine (no reference counterpart: BASELINE.json configs 4 and 5; the reference's
 * only simulator call is a 4-qubit Statevector, /root/reference/src/QFed/qAmplitude.py:44-46).
 *
 * qfx_circuit_compile (HOST arrays): schedule a gate list (n_gates x {kind, q0, q1, param}) into passes
 *   that each touch <= tile_bits qubits including qubits 0..low_bits-1; returns the pass count.
 * qfx_sv_large_pass: run one pass on `batch` states [batch][2^n] complex64; tile_qubits_host is a HOST
 *   array, ops_dev a DEVICE array of n_ops x 5 ints from the compiler; theta [theta_rows][n_params],
 *   angles [angle_rows][n_feature] (row = batch index modulo rows); shift_param/shift_value [batch]
 *   add a per-state shift to one parameter (parameter-shift rule, /root/reference/ROADMAP.md:130-135).
 *   plan_dev (nullable, tile_bits == 12 only): segment plan for the register-resident engine (DEVICE ints:
 *   n_seg [| 1 << 16 when a trailer follows], then per segment pos_to_bit[12], n_ops, n_ops x {kind, posA, posB,
 *   op index}; optional trailer: one mode per op of the pass (0 general, 1 RXRZ without its scalar phase, 2 the
 *   carrier gate that absorbs those phases) and the carrier's op index (-1 none)); see ops/large.py.
 * qfx_sv_large_expz: <Z_q> for all n_total qubits -> out [batch][n_total]; qubits >= n_local are
 *   read from high_bits (the shard's rank).
 * qfx_sv_large_expz_masks: parity readout out[b][q] = sum_i |psi_i|^2 (-1)^parity(idx & masks_host[q]) with
 *   idx = high_bits << n_local | i (masks_host: HOST array of n_out <= 32 words).  <Z_q> behind a trailing CNOT
 *   ring is such a parity in front of it, so the ring is never simulated (ops/large.py: ring_readout_masks).
 * qfx_sv_global_swap_fused: sharded state, world = 2^g_bits ranks: exchange the g global qubits with
 *   the top g local qubits by pulling slices from every peer's shard (src_table: device array of
 *   world peer pointers) and apply n_ops <= 32 gates on the swapped-in qubits (local bits 0..g-1 of the
 *   slice index) in registers before storing to dst.  flags_table (nullable) gives the in-kernel
 *   cross-GPU barrier; epoch must increase per call. */
int qfx_circuit_compile(int n_qubits, int n_gates, const int* gates_in, int tile_bits, int low_bits, int max_passes,
                        int max_ops, int* pass_tile_qubits, int* pass_op_begin, int* ops_out);
int qfx_sv_large_pass(float* states, int n_qubits, int batch, int tile_bits, int low_bits,
                      const int* tile_qubits_host, const int* ops_dev, int n_ops, const float* theta, int theta_rows,
                      int n_params, const float* angles, int angle_rows, int n_feature, const int* shift_param,
                      const float* shift_value, int init_zero_state, const int* plan_dev, int plan_len,
                      cudaStream_t stream);
int qfx_sv_large_expz(const float* states, int n_local, int n_total, unsigned int high_bits, int batch, float* out,
                      cudaStream_t stream);
int qfx_sv_large_expz_masks(const float* states, int n_local, int n_out, unsigned int high_bits,
                            const unsigned long long* masks_host, int batch, float* out, cudaStream_t stream);
int qfx_sv_global_swap_fused(const float* const* src_table, float* dst, int n_local, int g_bits, int rank, int world,
                             const int* ops_dev, int n_ops, const float* theta, int n_params, const float* angles,
                             int n_feature, unsigned int* const* flags_table, unsigned int epoch,
                             cudaStream_t stream);

/* TinyCNN conv2 on tcgen05 as a zero-copy implicit GEMM (csrc/tinycnn_umma.cu); standalone entry used by
 * the tests: forward conv + bias + ReLU + 2x2 max-pool of n images.  a1 [n,16,14,14], weight [32,16,5,5],
 * bias [32] fp32; a2 [n,32,7,7] fp32, i2 [n,32,7,7] u8 (argmax 0..3); scratch 51,200 bytes.
 * Model: /root/reference/src/CFed/Classical_FL.py:25,33-34. */
int qfx_tinycnn_conv2_fwd_umma(const float* a1, const float* weight, const float* bias, int n, float* a2,
                               unsigned char* i2, void* scratch, cudaStream_t stream);
/* same, backward: data gradient da1 [n,16,14,14] from g2 [n,32,14,14]; per-image weight/bias gradient
 * partials part [n,13248] (weights at offset 416 in [o][c][tap] order, bias at 13216). */
int qfx_tinycnn_conv2_dgrad_umma(const float* g2, const float* weight, int n, float* da1, void* scratch,
                                 cudaStream_t stream);
int qfx_tinycnn_conv2_wgrad_umma(const float* g2, const float* a1, int n, float* part, cudaStream_t stream);

/* Adjoint differentiation on the pass engine (K4a for large states; the "adjoint" option of /root/reference/ROADMAP.md:23,27):
 * qfx_sv_large_parity_obs builds lambda = (sum_q cot[b][q] Z-string(masks[q])) psi after the forward passes;
 * qfx_sv_large_adjoint_pass walks one compiled pass BACKWARDS over psi and lambda (both end as they were before the pass)
 * and accumulates d/d theta into grad [theta_rows][n_params] (caller-zeroed).  Call it for the passes in reverse order. */
int qfx_sv_large_parity_obs(const float* psi, float* lam, int n_local, int n_out, const unsigned long long* masks_host,
                            const float* cot, int batch, cudaStream_t stream);
int qfx_sv_large_adjoint_pass(float* psi, float* lam, int n_qubits, int batch, int tile_bits, int low_bits,
                              const int* tile_qubits_host, const int* ops_dev, int n_ops, const float* theta,
                              int theta_rows, int n_params, const float* angles, int angle_rows, int n_feature,
                              float* grad, cudaStream_t stream);

/* tuning knob of K6: CTAs per SM of the fused swap's grid on the current device (0 = default 4) */
int qfx_sv_swap_set_ctas_per_sm(int ctas_per_sm);

/* NVLink bandwidth probes on the symmetric heap (csrc/p2p_probe.cu), run beside K6 so that its roofline is measured in the
 * same job: mode 0 ring pull (peer copy), 1 all-to-all pull (the swap's pattern without gates), 2 all-to-all push.
 * Plain loads / stores on peer pointers; the host orders the launches.  No reference counterpart. */
int qfx_p2p_probe(const float* const* src_table, float* const* dst_table, size_t shard_bytes, int rank, int world,
                  int mode, int unroll, int ctas_per_sm, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* QFEDX_B200_H */
