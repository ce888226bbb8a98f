/* srnn_types.h -- plain-C types shared by the B200 library (libsrnn_b200.so) and the CPU oracle.
 *
 * Vocabulary follows the reference (illiumst/self-replicating-neural-networks, code/network.py,
 * code/soup.py): a *net* is one tiny linear, bias-free Keras model; a *particle* is a net with a uid;
 * a *soup* is a population of particles; a *self-application* replaces a net's weights by the output
 * of the net applied to its own position-encoded weight list.
 *
 * Population arena layout (both libraries): one contiguous FP32 record per net,
 *     w[i * ld + k]      i = net index (0 .. n_nets-1), k = flat weight index (0 .. n_weights-1),
 *                        ld = record stride in floats (>= n_weights; n_weights when packed)
 * i.e. a row-major (n_nets, ld) matrix whose rows are get_weights_flat() vectors.  Records keep a
 * partner's weights in 2-3 DRAM sectors for the attack / learn_from gathers and make the per-generation
 * all-gather one contiguous block per rank; streaming kernels stage record tiles through shared memory.
 * "flat weight index" is the order of NeuralNetwork.get_weights_flat() (code/network.py:103-104):
 * Keras kernels (in, out) concatenated row-major, layer after layer; for the recurrent net the
 * kernel list is [K0, U0, K1, U1, ...] (input kernel, recurrent kernel per SimpleRNN layer).
 */
#ifndef SRNN_TYPES_H
#define SRNN_TYPES_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* net kinds -- code/network.py:213 (Weightwise), :292 (Aggregating), :524 (Recurrent) */
enum { SRNN_KIND_WW = 0, SRNN_KIND_AGG = 1, SRNN_KIND_RNN = 2 };

/* trajectory classes, in the reference's decision-list order
 * (code/experiment.py:79-91, code/soup.py:89-103) */
enum { SRNN_CLS_DIVERGENT = 0, SRNN_CLS_FIX_ZERO = 1, SRNN_CLS_FIX_OTHER = 2,
       SRNN_CLS_FIX_SEC = 3, SRNN_CLS_OTHER = 4 };

/* SGD sample order inside one Keras epoch (code/network.py:613-626 -> model.fit, batch_size=1) */
enum { SRNN_PERM_NONE = 0,      /* shuffle=False: samples in flat order                              */
       SRNN_PERM_EXPLICIT = 1,  /* caller supplies uint8 perm[n_epochs][n_samples], shared by all nets */
       SRNN_PERM_PHILOX = 2 };  /* per-(net, epoch) Fisher-Yates from Philox4x32-10                  */

/* self-application stop rules */
enum { SRNN_RUN_FIXED = 0,      /* self_attack(n_steps)                  code/network.py:124-127      */
       SRNN_RUN_UNTIL = 1 };    /* FixpointExperiment.run_net stop rule  code/experiment.py:70-77     */

/* soup-event action codes; precedence = later overwrites earlier, as in code/soup.py:55-87 */
enum { SRNN_ACT_NONE = 0, SRNN_ACT_ATTACK = 1, SRNN_ACT_LEARN = 2, SRNN_ACT_TRAIN = 3,
       SRNN_ACT_DIV_DEAD = 4, SRNN_ACT_ZERO_DEAD = 5 };

typedef struct srnn_net_desc {
    int32_t kind;        /* SRNN_KIND_*                                                    */
    int32_t width;       /* units per hidden layer                                         */
    int32_t depth;       /* number of hidden layers                                        */
    int32_t aggregates;  /* AGG only: number of aggregates (= net input = net output size) */
    int32_t n_weights;   /* total weights; filled by srnn_desc_init()                      */
} srnn_net_desc;

/* Soup.params (code/soup.py:17) snapshotted into a POD */
typedef struct srnn_soup_params {
    double  attacking_rate;      /* < 0 disables                                  */
    double  learn_from_rate;     /* < 0 disables                                  */
    int32_t train;               /* self-train epochs per particle per generation */
    int32_t learn_from_severity; /* learn_from epochs per event                   */
    int32_t remove_divergent;    /* bool                                          */
    int32_t remove_zero;         /* bool                                          */
    double  epsilon;             /* particles' params['epsilon']                  */
    float   lr;                  /* Keras SGD default 0.01                        */
    int32_t perm_mode;           /* SRNN_PERM_NONE or SRNN_PERM_PHILOX            */
    int32_t engine_path;         /* srnn_soup_sync*: 0 auto, 1 fused kernel per generation, 2 multi-pass, 3 persistent
                                    cooperative kernel for srnn_soup_sync_run (all give the same results) */
    /* trajectory side channel (optional; all zero = off): the generation kernels append every particle that dies -- its
     * GLOBAL index and the record it had when the cull test fired, i.e. what particle.save_state would have logged for
     * the old particle (code/soup.py:77-87) -- before the slot is respawned.  Used by srnn_soup_log_watched. */
    int32_t dying_cap;           /* capacity of the two arrays below, in records                                  */
    float   *dying_w;            /* device float[dying_cap][n_weights]                                            */
    int64_t *dying_idx;          /* device int64[dying_cap]                                                       */
    int32_t *dying_count;        /* device int32, zeroed by the caller; may end above dying_cap (excess dropped)   */
} srnn_soup_params;

#define SRNN_MAX_WEIGHTS 64      /* per-net weight count supported by the generic code paths */
#define SRNN_MAX_WIDTH 8
#define SRNN_MAX_LAYERS 8
#define SRNN_MAX_SHARDS 16       /* ranks whose arenas one soup generation can read directly (peer memory) */

/* status codes (0 = ok); NaN/Inf in the data is never an error */
enum { SRNN_OK = 0, SRNN_ERR_ARG = 1, SRNN_ERR_SHAPE = 2, SRNN_ERR_CUDA = 3, SRNN_ERR_LIMIT = 4 };

#ifdef __cplusplus
}
#endif
#endif
