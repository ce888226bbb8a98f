/* srnn_b200.h -- C ABI of libsrnn_b200.so, the B200 (sm_100a) population engine.
 *
 * The reference (illiumst/self-replicating-neural-networks) has no FFI: its "operator API" is the
 * Python class surface of code/network.py / code/soup.py / code/experiment.py.  Each entry point
 * below replaces the inner loops of the cited reference methods for a whole population at once;
 * the Python facade (self-replicating-neural-networks_b200/network.py, soup.py, experiment.py)
 * binds them with ctypes and keeps the reference's class names and call signatures.
 *
 * Conventions
 *  - plain pointers and sizes only; every `float* w` is a DEVICE pointer to a population arena of
 *    records (include/srnn_types.h): w[i*ld + k].  Index arrays, outputs and counters are DEVICE
 *    pointers too unless the comment says "host".
 *  - `stream` is a cudaStream_t passed as void* (NULL = default stream).  Calls are asynchronous on
 *    that stream; the library allocates nothing and never frees caller memory.
 *  - return value: SRNN_OK (0) or an SRNN_ERR_* code; srnn_last_error_string() gives the text.
 *    NaN / Inf in the data is a domain state ("divergent"), never an error.
 *  - arithmetic: FP32, flush-to-zero, operation order and FMA placement pinned to the reference's
 *    committed Keras trajectories (DESIGN.md "Pinned arithmetic").
 *  - net shapes: the benchmarked ones are compile-time instantiated (fast path); every other shape with at most
 *    SRNN_MAX_WEIGHTS weights runs the per-net operations (init, apply, forward, self_apply, classify, train) on
 *    a generic run-time-shape path, and so do srnn_kfv, srnn_soup_sync, srnn_soup_sync_peers and srnn_soup_sync_run
 *    (one generic generation per call); srnn_soup_ordered, srnn_soup_exact* and srnn_soup_chain_peers exist for
 *    the instantiated shapes only (SRNN_ERR_SHAPE).
 *    srnn_shape_supported(): 1 = instantiated, 2 = generic path, 0 = invalid.
 */
#ifndef SRNN_B200_H
#define SRNN_B200_H

#include "srnn_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* library / device plumbing (no reference counterpart) */
int         srnn_version(void);
const char *srnn_last_error_string(void);
int         srnn_set_device(int device);
/* fills d->n_weights from (kind, width, depth, aggregates): the Keras layer stacks of
 * code/network.py:226-230 (WW), :329-333 (AGG), :531-535 (RNN) */
int         srnn_desc_init(srnn_net_desc *d);
int         srnn_shape_supported(const srnn_net_desc *d);

/* NeuralNetwork.apply_to_network / attack / fuck (code/network.py:112-122) over a list of pairs, with
 * WeightwiseNeuralNetwork.apply_to_weights :265-279, AggregatingNeuralNetwork.apply_to_weights :359-386,
 * RecurrentNeuralNetwork.apply_to_weights :544-564 as the operator:
 *     out[out_idx[p]] <- f_{op[op_idx[p]]}( in[in_idx[p]] ),  p = 0..n-1     (NULL idx = identity p)
 * attack(a -> v): op=a, in=out=v.  fuck: op=self, in=other, out=self.  Pairs run in parallel: within
 * one call no record may be written by one pair and read or written by another. */
int srnn_apply(const srnn_net_desc *d, const float *op_w, int64_t op_ld, const int32_t *op_idx,
               const float *in_w, int64_t in_ld, const int32_t *in_idx,
               float *out_w, int64_t out_ld, const int32_t *out_idx, int64_t n, void *stream);

/* NeuralNetwork.apply(*inputs) = model.predict on one input row (WeightwiseNeuralNetwork.apply
 * code/network.py:235-237, AggregatingNeuralNetwork.apply :355-357, RecurrentNeuralNetwork.apply :540-542):
 * WW / AGG: x = float[n][io] -> y = float[n][out] (io = 4 / aggregates, out = 1 / aggregates);
 * RNN: x = float[n][n_weights] (a sequence as long as the weight list) -> y = float[n][n_weights].
 * Row p is evaluated by the net in record idx[p] (NULL = p).  x, y: device pointers. */
int srnn_forward(const srnn_net_desc *d, const float *w, int64_t ld, const int32_t *idx, const float *x, float *y,
                 int64_t n, void *stream);

/* NeuralNetwork.self_attack(n_steps) (code/network.py:124-127) for every net, weights register-resident
 * for the whole run; mode SRNN_RUN_UNTIL adds FixpointExperiment.run_net's stop rule
 * (code/experiment.py:70-77: stop when is_diverged() or is_fixpoint()).
 *   steps_out  (nullable) int32[n]: self-applications executed per net
 *   cls_out    (nullable) uint8[n]: FixpointExperiment.count class of the final weights (experiment.py:79-91)
 *   counters   (nullable) int64[5]: class histogram, ACCUMULATED into (caller zeroes)
 *   traj       (nullable) float[n_steps][n][ld]: state after each executed step (ParticleDecorator.save_state,
 *              network.py:193-198); slots of steps never executed are left untouched */
int srnn_self_apply(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, int32_t n_steps, int32_t mode,
                    double eps, int32_t with_fix_sec, int32_t *steps_out, uint8_t *cls_out, int64_t *counters,
                    float *traj, void *stream);

/* Soup.count / FixpointExperiment.count decision list (code/soup.py:89-103, code/experiment.py:79-91):
 * is_diverged (network.py:133) -> is_fixpoint(1) (:140-157) -> is_zero (:136) -> is_fixpoint(2).
 * with_fix_sec = 0 gives the setups' soup-level count without the fix_sec branch (mixed-soup.py:27-53).
 * flags_out (nullable) uint8[n]: bit0 is_diverged, bit1 is_zero, bit2 is_fixpoint(1), bit3 is_fixpoint(2). */
int srnn_classify(const srnn_net_desc *d, const float *w, int64_t ld, int64_t n, double eps, int32_t with_fix_sec,
                  uint8_t *cls_out, uint8_t *flags_out, int64_t *counters, void *stream);

/* TrainingNeuralNetworkDecorator.train (code/network.py:613-618) when teacher_w == NULL, learn_from
 * (:620-626) otherwise: n_epochs Keras epochs of SGD(lr) on MSE with batch_size 1, fused forward +
 * backward + update, weights register-resident.  Samples = compute_samples() of the net itself
 * (re-frozen at the start of every epoch, :615) or of teacher record teacher_idx[i] (NULL = i).
 *   perm_mode  SRNN_PERM_NONE | SRNN_PERM_EXPLICIT (perm = uint8[n_epochs][n_samples], device) |
 *              SRNN_PERM_PHILOX (per (slot0+i, epoch0+e) Fisher-Yates)
 *   loss_out   (nullable) double[n]: Keras History 'loss' of the LAST epoch */
int srnn_train(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, int32_t n_epochs, float lr,
               int32_t perm_mode, const uint8_t *perm, uint64_t seed, uint64_t slot0, uint64_t epoch0,
               const float *teacher_w, int64_t teacher_ld, const int32_t *teacher_idx, double *loss_out,
               void *stream);

/* Soup.generate_particle (code/soup.py:37-40) = fresh Keras model: Dense kernels glorot_uniform, SimpleRNN
 * recurrent kernels orthogonal (Keras defaults implied by network.py:227-230, :330-333, :532-535).
 * Deterministic: record idx[p] (NULL = p) is a pure function of (seed, slot0 + index, generation). */
int srnn_init(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, const int32_t *idx, uint64_t seed,
              uint64_t slot0, uint32_t generation, void *stream);

/* code/setups/known-fixpoint-variation.py: vary() :37-46 + trial loop :60-89 for n trials at one noise
 * level.  base = the fixpoint's n_weights floats (HOST pointer, copied by value into the launch).
 *   time_to_vergence / time_as_fixpoint (nullable) int32[n]      = the script's ys / zs
 *   hist (nullable) int64[2][max_steps+1]: histograms of ys and zs, ACCUMULATED into
 *   w_out (nullable) float[n][ld]: final weights */
int srnn_kfv(const srnn_net_desc *d, const float *base_host, int64_t n, double scale, uint32_t level,
             int32_t max_steps, double eps, uint64_t seed, uint64_t slot0, int32_t *time_to_vergence,
             int32_t *time_as_fixpoint, int64_t *hist, float *w_out, int64_t ld, void *stream);

/* Soup.evolve (code/soup.py:51-87), exact order: one generation, particles visited in index order,
 * every write visible to later particles.  att / lrn = int64[n] target index or -1 per particle (drawn
 * by the host in the reference's order, soup.py:56-63).  Outputs (nullable): action uint8[n]
 * (SRNN_ACT_*), loss double[n] (last train epoch, NaN if none), logged float[n][ld] = the weights
 * particle.save_state (soup.py:87) logs for the old particle.  Respawn (generate_particle, soup.py:78,83):
 * if `spares` (float[n_spares][spare_ld], nullable) still holds unused records -- fresh nets the host's
 * generator() produced ahead of time -- the next one is consumed, in particle order; otherwise the record is
 * re-initialised like srnn_init(seed, slot, gen).  *spares_used (device int32, nullable) = records consumed.
 * Meant for the setups' soups (10-100 nets). */
int srnn_soup_ordered(const srnn_net_desc *d, const srnn_soup_params *p, float *w, int64_t ld, int64_t n,
                      const int64_t *att, const int64_t *lrn, uint64_t seed, uint64_t gen, uint8_t *action,
                      double *loss, float *logged, const float *spares, int64_t spare_ld, int32_t n_spares,
                      int32_t *spares_used, void *stream);

/* Soup.evolve (code/soup.py:51-87) with the reference's EXACT sequential semantics at population scale: the result is
 * bit-identical to visiting the particles one by one in index order (= srnn_soup_ordered), but the generation runs
 * as a dependency-levelled wavefront, in place on `w` (DESIGN.md "Soup modes"): ONE cooperative launch whose CTAs meet
 * at grid barriers; after the first pass only the not-yet-runnable events are revisited, and the device itself notices
 * when none is left.
 *   att / lrn  int64[n] device arrays of target indices or -1 (host-drawn, e.g. Python `random` in the reference's
 *              order); both NULL = Philox draws of (seed, particle, gen), the same function srnn_soup_sync uses;
 *              exactly one of them NULL is an error
 *   scratch    int32[SRNN_SOUP_EXACT_SCRATCH(n)], no initialisation needed
 *   action uint8[n], loss double[n], dead uint8[n] (all nullable)
 *   rounds_out (HOST int, nullable) = depth of the dependency DAG; asking for it makes the call wait for the stream
 *              (one 4-byte read-back) -- pass NULL to stay asynchronous
 * Respawn = srnn_init(seed, slot, gen).  Unsharded soups only; 3 n < 2^31. */
#define SRNN_SOUP_EXACT_SCRATCH(n) (11 * (int64_t)(n) + ((int64_t)(n) + 1) / 2 + 64)
int srnn_soup_exact(const srnn_net_desc *d, const srnn_soup_params *p, float *w, int64_t ld, int64_t n, const int64_t *att,
                    const int64_t *lrn, uint64_t seed, uint64_t gen, int32_t *scratch, uint8_t *action, double *loss,
                    uint8_t *dead, int32_t *rounds_out, void *stream);

/* n_gens exact-order generations (gen0, gen0 + 1, ...) with Philox pairings in ONE launch, fully asynchronous:
 * the multi-generation form of Soup.evolve(iterations) (code/soup.py:51-53).  action / dead describe the LAST
 * generation; rounds_dev (DEVICE int32[n_gens], nullable) receives every generation's DAG depth.  Same scratch. */
int srnn_soup_exact_run(const srnn_net_desc *d, const srnn_soup_params *p, float *w, int64_t ld, int64_t n, uint64_t seed,
                        uint64_t gen0, int32_t n_gens, int32_t *scratch, uint8_t *action, uint8_t *dead,
                        int32_t *rounds_dev, void *stream);

/* Soup.evolve, snapshot-synchronous generation for large / sharded soups (DESIGN.md "Soup modes"): every
 * partner is read from `snap`, the start-of-generation arena of ALL n_total particles (after the
 * per-generation all-gather when sharded); this call owns particles [slot0, slot0+n) and writes their
 * new records to w_out (w_out must not alias snap).  Pairings are Philox draws of (seed, slot, gen) in the
 * reference's per-particle order; a victim hit several times applies attackers in ascending index.
 *   scratch  int32[n_total + 4*n + 16]: attack lists (head[n] | next[n_total]) and, for the opt-in variants
 *            (params->engine_path 2 = multi-pass, 3 = persistent), work lists, flags, counters and the grid barrier
 *            word; needs no initialisation
 *   action (nullable) uint8[n], loss (nullable) double[n], dead (nullable) uint8[n] (1 = respawned) */
int srnn_soup_sync(const srnn_net_desc *d, const srnn_soup_params *p, const float *snap, int64_t snap_ld,
                   int64_t n_total, float *w_out, int64_t out_ld, int64_t slot0, int64_t n, uint64_t seed,
                   uint64_t gen, int32_t *scratch, uint8_t *action, double *loss, uint8_t *dead, void *stream);

/* Soup.evolve, snapshot-synchronous generation of a SHARDED soup without a gather: the exchange step is fused
 * into the generation kernel.  shards (HOST array of n_shards DEVICE pointers) are the ranks' start-of-generation
 * arenas, each holding shard_len particles (rank r owns global particles [r*shard_len, (r+1)*shard_len)); peers'
 * arenas must be mapped into this process (CUDA peer / symmetric memory over NVLink).  The kernel reads its own
 * records from shards[my_shard] and fetches only the partner records it needs (attackers, teachers: ~20 % of the
 * particles) straight from the owning GPU.  Everything else as srnn_soup_sync with n = shard_len.  The caller
 * must order generations across ranks (a barrier after each one) and ping-pong arenas: w_out of generation g is
 * the shard of generation g+1. */
int srnn_soup_sync_peers(const srnn_net_desc *d, const srnn_soup_params *p, const float *const *shards, int32_t n_shards,
                         int64_t shard_len, int32_t my_shard, int64_t ld, float *w_out, int64_t out_ld, uint64_t seed,
                         uint64_t gen, int32_t *scratch, uint8_t *action, double *loss, uint8_t *dead, void *stream);

/* The same generation of a sharded soup as srnn_soup_sync_peers, with the attack lists built by the ATTACKER's rank
 * (no counterpart in the reference: code/soup.py:56-61 is one process; this is what keeps the per-generation cost of a
 * rank proportional to ITS particles).  Every rank holds, in peer-mapped memory, two parities of
 *     head  int32[shard_len]   last attacker pushed onto each of ITS particles, -1 = none
 *     next  int32[n_total]     previous head, stored by attacker g at index g
 * heads_cur / nexts_cur (HOST arrays of n_shards device pointers) are the arrays that hold THIS generation's lists (only
 * [my_shard] is read, and its heads are reset to -1 as they are consumed); heads_nxt / nexts_nxt (both NULL: skip)
 * are the other parity: the kernel pushes generation gen+1's attacks of this rank's particles onto the victim owners'
 * lists -- one system-scope exchange + one store over NVLink per attack.  Protocol: all heads -1, barrier across ranks,
 * srnn_soup_push_lists(gen0) on every rank, barrier, then per generation: this call, barrier, swap parities and arenas.
 * The per-generation barrier can be left to the library: flags (HOST array of n_shards device pointers, nullable) points
 * at every rank's int32[n_shards + 1] flag array in peer-mapped memory (zeroed once, before the first generation and
 * behind a barrier).  Behind the generation a one-warp kernel stores gen into word my_shard of every rank's array;
 * wait_gen > 0 makes the call wait -- on the device, ahead of the generation -- until this rank's array shows every
 * rank at wait_gen or later (pass gen - 1 from the second generation of a run on; generations count up, < 2^31).
 * scratch (local int32[shard_len + 16], nullable, no initialisation): lets the recurrent kind hand its learn_from epochs
 * to a dense second kernel; without it everything runs inside the generation kernel. */
int srnn_soup_chain_peers(const srnn_net_desc *d, const srnn_soup_params *p, const float *const *shards, int32_t n_shards,
                          int64_t shard_len, int32_t my_shard, int64_t ld, float *w_out, int64_t out_ld, uint64_t seed,
                          uint64_t gen, int32_t *const *heads_cur, int32_t *const *nexts_cur, int32_t *const *heads_nxt,
                          int32_t *const *nexts_nxt, int32_t *const *flags, int64_t wait_gen, int32_t *scratch,
                          uint8_t *action, double *loss, uint8_t *dead, void *stream);

/* First generation of such a run: this rank's particles push their generation-`gen` attacks (Philox draws of
 * (seed, global particle, gen), code/soup.py:56-59) onto the owners' lists.  heads must be -1 on every rank beforehand. */
int srnn_soup_push_lists(const srnn_soup_params *p, int32_t n_shards, int64_t shard_len, int32_t my_shard, uint64_t seed,
                         uint64_t gen, int32_t *const *heads, int32_t *const *nexts, void *stream);

/* Trajectory logging at population scale (particle.save_state at the end of every particle's turn, code/soup.py:87,
 * for a chosen subset): writes ONE ROW of the caller's device ring buffer for the n_watch particles whose LOCAL indices
 * are in `watch`, after generation `gen` has run on arena w (and before the respawned slots get their new uids):
 *   ring_w float[n_watch][n_weights]   record after the generation (the fresh net where the particle was respawned)
 *   ring_dying float[n_watch][n_weights]  for particles that died: the record they had when the cull test fired (the
 *                                      state the reference logs for the old particle), taken from the dying side channel
 *                                      of `p` (include/srnn_types.h); NaN if not dead or the channel overflowed
 *   ring_action uint8[n_watch] (SRNN_ACT_*), ring_uid int64[n_watch] (uid the particle ran under), ring_loss double[n_watch]
 *   ring_counterpart int64[n_watch]    GLOBAL INDEX of the teacher if it learned, else of the attack victim, else -1
 * p, seed, gen must be those of the generation call (the pairing is re-derived, not stored).  loss may be NULL. */
int srnn_soup_log_watched(const srnn_net_desc *d, const srnn_soup_params *p, const float *w, int64_t ld, int64_t n_total,
                          int64_t slot0, const int32_t *watch, int32_t n_watch, uint64_t seed, uint64_t gen,
                          const uint8_t *action, const double *loss, const int64_t *uid_local, float *ring_w,
                          float *ring_dying, uint8_t *ring_action, int64_t *ring_uid, int64_t *ring_counterpart,
                          double *ring_loss, void *stream);

/* n_gens consecutive snapshot-synchronous generations (gen0, gen0+1, ...) of an UNSHARDED soup in one call, so a
 * small soup is not bound by the host's per-call overhead: generation gen0+g reads w_a and writes w_b for even g,
 * the other way round for odd g -- the newest arena is w_b if n_gens is odd, else w_a.  Same arguments and
 * semantics as srnn_soup_sync with n_total = n, slot0 = 0; action / dead (nullable) hold the LAST generation's.
 * params->engine_path = 3 runs all generations as ONE persistent cooperative kernel (n <= 2^20); same results. */
int srnn_soup_sync_run(const srnn_net_desc *d, const srnn_soup_params *p, float *w_a, float *w_b, int64_t ld, int64_t n,
                       uint64_t seed, uint64_t gen0, int32_t n_gens, int32_t *scratch, uint8_t *action, uint8_t *dead,
                       void *stream);

/* ParticleDecorator uid allocation for respawned particles (code/network.py:167-171, soup.py:78,83):
 * uid[i] <- uid_base + (number of dead[j] != 0 with j < i) for every dead i; *n_dead_out (device int64,
 * nullable) receives the total.  block_tmp = int64[ceil(n/1024) + 1] scratch. */
int srnn_assign_uids(const uint8_t *dead, int64_t n, int64_t uid_base, int64_t *uid, int64_t *block_tmp,
                     int64_t *n_dead_out, void *stream);

/* Same result as srnn_assign_uids, for callers that cannot allocate device memory.  (No counterpart in the
 * reference: it exists for hosts that only pass through device pointers owned by another runtime and link no
 * allocator themselves, and for soup loops whose device memory was handed out once at start-up and that must
 * not call cudaMalloc -- which synchronises the whole device -- between generations.)  The block-count scratch
 * is a fixed array inside the library image: 128 KiB on every GPU that loads libsrnn_b200.so, kept until the
 * process exits.  n may be at most SRNN_UIDS_STATIC_MAX; larger n is refused with SRNN_ERR_LIMIT.  The call
 * makes no allocation of any kind.  The first call on a device looks up the array's device address (a host-side
 * query with that device current); make that first call before starting any stream capture.  `stream` must belong
 * to the device that is current at the time of the call -- the scratch address is the current device's -- and a
 * stream of another device is refused with SRNN_ERR_ARG.  All calls share
 * the one array per GPU: it needs no clearing between calls (every counter is rewritten before it is read),
 * but two calls whose kernels overlap on the same GPU -- different streams or different host threads --
 * race on it, so keep these calls ordered on one stream per device. */
#define SRNN_UIDS_STATIC_MAX (1 << 24)
int srnn_assign_uids_static(const uint8_t *dead, int64_t n, int64_t uid_base, int64_t *uid, int64_t *n_dead_out,
                            void *stream);

/* Communicator family for hosts without torch.distributed: the all-gather of the record arenas that the
 * `allgather` exchange of a sharded soup needs (soup.py PopulationSoup does the same through torch.distributed; the
 * peer-memory exchanges srnn_soup_sync_peers / srnn_soup_chain_peers need no communicator at all).  NCCL is looked up at
 * run time (the libnccl.so.2 already loaded in the process, else the system's); without it the calls return
 * SRNN_ERR_CUDA with a message.  One communicator per process and device:
 *   rank 0: srnn_comm_unique_id(id)  ->  the host broadcasts the 128 bytes by its own means  ->
 *   every rank, its GPU current: srnn_comm_init(id, rank, world, &comm)   (collective)
 *   srnn_comm_allgather(comm, send, recv, count, stream): recv[r * count .. (r + 1) * count) = rank r's send[0 .. count),
 *   floats, asynchronous on `stream`;  srnn_comm_destroy(comm). */
#define SRNN_COMM_ID_BYTES 128
int srnn_comm_unique_id(uint8_t *id_out);
int srnn_comm_init(const uint8_t *id, int32_t rank, int32_t world, void **comm_out);
int srnn_comm_allgather(void *comm, const float *send, float *recv, int64_t count, void *stream);
int srnn_comm_destroy(void *comm);

/* FP32 issue-rate probe: an FFMA-only kernel (independent chains, operands in registers) used as the
 * measured roofline denominator for the register-resident kernels (MEASURED_PEAKS.json has no FP32 entry).
 * Runs |iters| x 64 FMA per thread over grid x block threads; out = float[grid*block] sink.  iters > 0: three-register
 * FFMA; -2^24 < iters < 0: FFMA with an immediate operand; iters <= -2^24: packed FFMA2 (|iters| - 2^24 iterations). */
int srnn_fp32_probe(float *out, int32_t grid, int32_t block, int32_t iters, void *stream);

#ifdef __cplusplus
}
#endif
#endif
