/* ll_gpu.h — C ABI of libll_gpu.so: the B200 (sm_100a) implementation of lunar-lander's neuro-evolution
 * hot path (population evaluation + GA select/mutate).
 *
 * The reference has no C ABI; this path sits behind three C++ interfaces.  Each entry point
 * below names the reference interface it replaces (file:line under azimonti/lunar-lander).  The host-side
 * mirrors of those interfaces (C++: lunar_lander_b200/host/, Python: the lunar_lander_b200 package) are thin
 * wrappers over exactly these functions; INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions: extern "C"; every function returns 0 on success and a non-zero LL_E* code otherwise, with a
 * message retrievable through ll_last_error(); plain pointers and sizes only; the caller owns every host
 * buffer; the context owns all device memory, streams and events; no exception crosses the boundary; one host
 * thread per context; calls are synchronous from the caller's point of view unless stated.  There is no CPU
 * fallback: without a CUDA device ll_create fails with LL_ENODEVICE.
 */
#ifndef LL_GPU_H_
#define LL_GPU_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LL_ABI_VERSION 1

enum
{
    LL_OK        = 0,
    LL_EINVAL    = 1, /* bad argument */
    LL_ENODEVICE = 2, /* no usable CUDA device / driver */
    LL_ECUDA     = 3, /* a CUDA runtime call failed (message has the CUDA error string) */
    LL_EIO       = 4, /* weight-file read/write failed */
    LL_ENOTOWNED = 5  /* the member lives on another rank's shard */
};

enum
{
    LL_MODE_EXACT = 0, /* same operation order as the CPU reference path: separate multiply/add, sequential-k
                          dot products, deterministic tanh -> bit-exact actions / steps / flags / fitness */
    LL_MODE_FAST  = 1  /* fp32 only: fused multiply-add, tanh.approx.f32, free summation order */
};

enum
{
    LL_ACT_SIGMOID = 0,
    LL_ACT_TANH    = 1 /* config.txt:40 NNConfig.activation_id */
};

enum
{
    LL_STRATEGY_ALL_FROM_ELITES             = 1,
    LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION = 2 /* nn::PopulationStrategy, src/nn_engine_trainer.cpp:93 */
};

/* Every configuration value GameLogicCpp<T>'s constructor reads (src/game_logic.cpp:48-95), as doubles; the
 * library converts to T with static_cast<T>() exactly as the reference does.  One instance replaces the
 * `pop` per-simulator copies the reference keeps (src/nn_engine_trainer.cpp:325-330). */
typedef struct ll_env_cfg
{
    double pad_y1, terrain_y, max_vx, max_vy;                                /* GameCfg.*            :48-56 */
    double g, mu_x, mu_y;                                                    /* PlanetCfg.*          :58-60 */
    double lander_width, lander_height, max_fuel;                            /* LanderCfg.*          :62-64 */
    double lpad_x1, lpad_width, display_width, display_height, spad_width;   /* Game/DisplayCfg      :66-72 */
    double x0, y0, vx0, vy0, ax0, ay0;                                       /* GameCfg.x0/v0/a0     :74-83 */
    double sp_dist_factor, sp_action_factor;                                 /* NNTrainingCfg.sp_*   :86-87 */
    double tp_steps_factor, tp_dist_factor, tp_landed_bonus, tp_landed_lr_bonus, tp_fuel_bonus_factor,
        tp_crashed_penalty, tp_crash_v_mag_factor, tp_no_fuel_penalty;       /* NNTrainingCfg.tp_*   :88-95 */
} ll_env_cfg;

/* Arguments of nn::ANN_MLP_GA<T>(sizes, seed, pop, top, act) (src/nn_engine_trainer.cpp:87-89) plus what a
 * GPU build has to know. */
typedef struct ll_net_cfg
{
    const uint64_t* layer_sizes; /* {5, h..., 4}: src/nn_engine_trainer.cpp:45-48 */
    int32_t n_layers;            /* number of sizes, 2..16 */
    int32_t activation_id;       /* LL_ACT_* */
    uint64_t population;         /* GLOBAL population size */
    uint64_t top;                /* elites kept per generation (NNConfig.top_individuals) */
    uint64_t seed;               /* NNConfig.seed */
    int32_t precision;           /* 32 or 64  (main_train_f / main_train_d, src/main_train.cpp:24-28) */
    int32_t mode;                /* LL_MODE_* ; LL_MODE_FAST requires precision 32 */
    int32_t rank, world;         /* this process' shard: members [rank*pop/world, (rank+1)*pop/world) */
    int32_t device;              /* CUDA ordinal, -1 = current device */
    int32_t reserved;
    double sigma_init;           /* <= 0 -> 1.0 */
    double sigma_mut;            /* <= 0 -> 0.1 */
} ll_net_cfg;

/* Per-generation reductions (what train() prints, src/nn_engine_trainer.cpp:413-435), over the LOCAL shard. */
typedef struct ll_gen_stats
{
    double best_fitness;   /* min over the shard */
    uint64_t best_index;   /* GLOBAL index of that member; lowest index among ties */
    double sum_fitness;    /* sequential-order independent pairwise sum, informational */
    double sum_steps;      /* = number of lander steps executed (exact: integers < 2^53) */
    double best_lr;        /* last_gen_lr of best_index */
    uint64_t members;      /* shard size */
} ll_gen_stats;

/* One step of an episode as seen by the trainer loop (src/nn_engine_trainer.cpp:501-510). */
typedef struct ll_step_rec
{
    double action;   /* 0 noop, 1 up, 2 left, 3 right (src/game_logic.cpp:165-181) */
    double state[5]; /* get_state() AFTER the update (src/game_logic.cpp:275-293) */
    double x, y, vx, vy, fuel;
} ll_step_rec;

typedef struct ll_episode_result
{
    int32_t steps;
    int32_t flags;             /* bit0 landed, bit1 crashed, bit2 landed_successfully */
    double step_penalty_sum;   /* sum of calculate_step_penalty (double accumulator, :507-508) */
    double terminal_penalty;   /* calculate_terminal_penalty(steps, dir, bonus) (:514-515) */
    double bonus_lr[2];
} ll_episode_result;

typedef struct ll_cand
{
    double fitness;
    uint64_t index; /* GLOBAL member index */
} ll_cand;

typedef struct ll_info
{
    uint64_t params;        /* P = sum(in*out + out) */
    uint64_t params_padded; /* device row length in elements */
    uint64_t pop_global, pop_local, first_global, top;
    int32_t precision, mode, rank, world, device, sm_count;
    char device_name[64];
    char kernel[64];        /* name of the episode-kernel variant selected for this topology */
} ll_info;

typedef struct ll_kernel_times
{
    double episode_ms;  uint64_t episode_launches;  /* K1  fused episode kernel */
    double select_ms;   uint64_t select_launches;   /* K2  fitness reduce + top-k */
    double gather_ms;   uint64_t gather_launches;   /* K3  elite gather */
    double mutate_ms;   uint64_t mutate_launches;   /* K4/K0 children / init */
    uint64_t other_launches;                        /* layout transposes, trace, feedforward */
} ll_kernel_times;

/* The public face of one GameLogicCpp<T> object (src/game_logic.h:17-21 members + is_done() + get_state()). */
typedef struct ll_sim_view
{
    double x, y, vx, vy, fuel;
    double landing_pad_center_x, landing_pad_y;
    int32_t landed, crashed, landed_successfully, time_step, last_action, done;
    double state[5]; /* get_state() of the current state */
} ll_sim_view;

typedef struct ll_ctx ll_ctx;

int ll_abi_version(void);
/* thread-local message of the last failing call on this thread (ctx may be NULL for ll_create failures) */
const char* ll_last_error(const ll_ctx* ctx);

/* ---- life cycle ---------------------------------------------------------------------------------------
 * replaces  nn::ANN_MLP_GA<T>::ANN_MLP_GA(sizes, seed, pop, top, act)  src/nn_engine_trainer.cpp:87-89
 *      and  the pop x GameLogicCpp<T>(no_print=true) constructions      src/nn_engine_trainer.cpp:325-330 */
int ll_create(ll_ctx** out, const ll_env_cfg* env, const ll_net_cfg* net);
void ll_destroy(ll_ctx* ctx);
int ll_get_info(const ll_ctx* ctx, ll_info* out);
/* run all work of this context on an existing cudaStream_t (so callers can time it with their own events) */
int ll_set_stream(ll_ctx* ctx, void* cuda_stream);
int ll_synchronize(ll_ctx* ctx);

/* ---- GA -----------------------------------------------------------------------------------------------
 * replaces  SetPopulationStrategy(MIXED_WITH_RANDOM_INJECTION, ratio)   src/nn_engine_trainer.cpp:93 */
int ll_set_population_strategy(ll_ctx* ctx, int strategy, double random_injection_ratio);
/* replaces  CreatePopulation(elitism)   src/nn_engine_trainer.cpp:94 (initial) and :405 (per generation).
 * Every draw is a function of (seed, generation, GLOBAL member, position): the children do not depend on the GPU count.  In
 * LL_MODE_EXACT the normals are reproducible bit for bit on the host (csrc/ll_philox.h, what the oracle runs); in LL_MODE_FAST
 * (fp32) they come from the GPU's special-function unit: deterministic on the device, judged statistically. */
int ll_create_population(ll_ctx* ctx, int elitism);
/* standard deviations of the initial / injected individuals and of the mutation noise (values <= 0 keep the current
 * ones; ma-libs' own constants are not recoverable, SURVEY.md App. C — defaults 1.0 and 0.1) */
int ll_set_sigmas(ll_ctx* ctx, double sigma_init, double sigma_mut);
/* replaces  UpdateEpochs(n) / GetEpochs()   src/nn_engine_trainer.cpp:406-407 */
int ll_update_epochs(ll_ctx* ctx, uint64_t n);
int ll_get_epochs(const ll_ctx* ctx, uint64_t* out);
int ll_set_epochs(ll_ctx* ctx, uint64_t epochs);

/* replaces  std::sort of indices by fitness + UpdateWeightsAndBiases(sorted_indices)
 *           src/nn_engine_trainer.cpp:376-379, 404        (world == 1, or any world once ll_comm_init has been called:
 *           then every rank makes the same call and sorted_idx must be the same on all of them)
 * sorted_idx: `n_sorted` (>= top) GLOBAL indices in ascending-fitness order as the reference computes them on
 * the host, or NULL: then the elites are chosen on the device from the last ll_evaluate, ordered by
 * (fitness, index).  elite_idx_out (may be NULL) receives the `top` chosen indices. */
int ll_select(ll_ctx* ctx, const uint64_t* sorted_idx, uint64_t n_sorted, uint64_t* elite_idx_out);

/* multi-rank selection, one process per GPU:
 *   1. ll_local_topk        -> this shard's `top` best (fitness, global index), ascending
 *   2. caller all-gathers the candidates (NCCL) and merges them by (fitness, index)
 *   3. ll_gather_elites     -> writes a [top][params_padded] matrix of T into caller-provided DEVICE memory:
 *                              row r = parameters of elite_idx[r] if this rank owns it, zero bits otherwise
 *   4. caller all-reduces that buffer (integer SUM over the bit patterns = exact broadcast from the owners)
 *   5. ll_set_elites        -> installs the matrix as the parent set of the next ll_create_population */
int ll_local_topk(ll_ctx* ctx, ll_cand* out, uint64_t* n_out);
int ll_gather_elites(ll_ctx* ctx, const uint64_t* elite_idx, uint64_t n, void* device_buf);
int ll_set_elites(ll_ctx* ctx, const void* device_buf, uint64_t n);
uint64_t ll_elite_buffer_bytes(const ll_ctx* ctx);

/* ---- the exchange between GPUs, owned by the context --------------------------------------------------
 * replaces  nothing in the reference (it is single-process); it is what lets std::sort + UpdateWeightsAndBiases
 *           (src/nn_engine_trainer.cpp:376-379, 404) and the thread-pool fan-out (:460-538) span several GPUs.
 * One context per GPU (one host thread or one process each), created with rank/world.  Rank 0 obtains an id, the host hands its
 * LL_COMM_ID_BYTES bytes to every rank (threads: memcpy; processes: MPI / a torch store / a file), every rank calls ll_comm_init
 * (collective: ncclCommInitRank).  From then on ll_step_generation and ll_select work on shards: K2 local top-k -> NCCL all-gather
 * of top x 16 B -> one-block merge by (fitness, index) -> K3 row gather -> NCCL all-reduce of the [top][params_padded] parent
 * matrix, enqueued on the context's stream with no host synchronisation in between; ll_gen_stats then describe the WHOLE population
 * on every rank.  NCCL is dlopen'ed (libnccl.so.2); without it these calls fail with LL_ENODEVICE. */
#define LL_COMM_ID_BYTES 128
int ll_comm_get_unique_id(void* id_out);
int ll_comm_init(ll_ctx* ctx, const void* unique_id);
int ll_comm_destroy(ll_ctx* ctx);

/* ---- evaluation: the hot path ------------------------------------------------------------------------------
 * replaces  NNEngineTrainer<T>::train_generation(layouts, layout_directions, population_size, fitness_scores,
 *           last_gen_lr, all_member_steps, game_sims)                        src/nn_engine_trainer.cpp:443-539
 * spad_x1/lpad_x1/directions: n_layouts entries (LayoutInfo::spad_x1/lpad_x1, layout_directions[i] in {0,1}).
 * fitness/last_gen_lr/steps: pop_local doubles each, or NULL to skip that copy.  stats may be NULL. */
int ll_evaluate(ll_ctx* ctx, const int32_t* spad_x1, const int32_t* lpad_x1, const uint64_t* directions,
                int32_t n_layouts, int32_t max_steps, double* fitness, double* last_gen_lr, double* steps,
                ll_gen_stats* stats);
/* the same split in three, so the episode kernel can be timed with inputs already resident in HBM */
int ll_set_layouts(ll_ctx* ctx, const int32_t* spad_x1, const int32_t* lpad_x1, const uint64_t* directions,
                   int32_t n_layouts);
int ll_evaluate_device(ll_ctx* ctx, int32_t max_steps); /* asynchronous: enqueues K1 (+K2 stats) on the stream */
int ll_read_results(ll_ctx* ctx, double* fitness, double* last_gen_lr, double* steps, ll_gen_stats* stats);
/* per-episode outcomes of the last evaluation, [pop_local][n_layouts] each (either may be NULL):
 *   meta    = steps | landed << 28 | crashed << 29 | landed_successfully << 30   (src/nn_engine_trainer.cpp:499-511)
 *   fitness = terminal penalty + accumulated step penalties of that layout        (:514-517) */
int ll_read_episodes(ll_ctx* ctx, int32_t* meta, double* fitness);

/* one whole generation without a host round trip in the middle (world == 1, or shards with a communicator):
 * evaluate -> device top-k -> elite gather -> children -> epochs += 1   (src/nn_engine_trainer.cpp:373-407) */
int ll_step_generation(ll_ctx* ctx, int32_t max_steps, int elitism, ll_gen_stats* stats);

/* parity aid: one (member, layout) episode with every step recorded (body of sim_task, :491-517) */
int ll_trace_episode(ll_ctx* ctx, uint64_t member, int32_t spad_x1, int32_t lpad_x1, uint64_t direction,
                     int32_t max_steps, ll_step_rec* rec, int32_t cap, ll_episode_result* result);

/* replaces  feedforward(const T*, size_t, size_t member) -> action   src/nn_engine_trainer.cpp:501-502
 *      and  cpp_nn_py feedforwardIndex(inputs, member)               mod_nn_play.py:56
 * state: 5 values of T (float when precision == 32, double when 64). */
int ll_feedforward_index(ll_ctx* ctx, const void* state, uint64_t member, uint64_t* action);

/* ---- parameters ---------------------------------------------------------------------------------------------
 * canonical order per member: for each layer l: W_l[out][in] row-major, then b_l[out]  (the order of the
 * weight file).  Buffers hold T. */
int ll_get_member(ll_ctx* ctx, uint64_t member, void* params);
int ll_set_member(ll_ctx* ctx, uint64_t member, const void* params);
/* bulk upload of `count` consecutive members starting at GLOBAL index `first` (must lie in this shard) */
int ll_set_members(ll_ctx* ctx, uint64_t first, uint64_t count, const void* params);
int ll_get_members(ll_ctx* ctx, uint64_t first, uint64_t count, void* params);
/* marks slots [0, top) as a valid parent set (what UpdateWeightsAndBiases / Deserialize establish) */
int ll_mark_elites(ll_ctx* ctx);

/* replaces  Serialize(path) / Deserialize(path)   src/nn_engine_trainer.cpp:172, 126  (format: App. B) */
int ll_serialize(ll_ctx* ctx, const char* path, const char* name, int32_t digits);
/* header only: fills sizes (cap entries), pop, top, epochs, act, strategy, ratio — so a caller can build the
 * matching context before ll_deserialize */
int ll_peek_weight_file(const char* path, uint64_t* sizes, int32_t cap, int32_t* n_sizes, uint64_t* pop,
                        uint64_t* top, uint64_t* epochs, int32_t* act, int32_t* strategy, double* ratio,
                        char* name, int32_t name_cap);
int ll_deserialize(ll_ctx* ctx, const char* path);

/* ---- host-side inputs of the path (kept on the CPU: 50 calls per run) --------------------------------------
 * replaces  generate_random_pad_positions(...)   src/game_utils.cpp:12-74 */
int ll_generate_random_pad_positions(double screen_width, int32_t spad_width, int32_t lpad_width, int32_t has_seed,
                                     int32_t seed, int32_t force_left_to_right, int32_t force_right_to_left,
                                     int32_t* spad_x1, int32_t* lpad_x1);
/* replaces  NNEngineTrainer<T>::balance_layouts   src/nn_engine_trainer.cpp:196-260
 * spad_x1/lpad_x1/spad_center/lpad_center/directions are in/out arrays of n entries. */
int ll_balance_layouts(int32_t* spad_x1, int32_t* lpad_x1, double* spad_center, double* lpad_center,
                       uint64_t* directions, int32_t n, double target_ltr_ratio, double screen_width,
                       double spad_width, double lpad_width);

/* ---- one stepped lander: the GameLogicCpp<T> object of nn_play / play mode ----------------------------------------
 * replaces  GameLogicCpp<T>::reset() / reset(T spad_x1, T lpad_x1)     src/game_logic.cpp:124-154
 *           update(int action, T* state, size_t n) -> done             src/game_logic.cpp:295-310
 *           get_state / is_done / the public members                   src/game_logic.cpp:253-293, src/game_logic.h:17-21
 *           calculate_step_penalty(action)                             src/game_logic.cpp:312-320
 *           calculate_terminal_penalty(steps) / (steps, dir, bonus&)   src/game_logic.cpp:322-388
 *      and  the lunar_lander_cpp bindings of those                     src/lunar_lander_cpp.cpp:21-93
 * The lander lives on the device and is advanced by the same device functions as the episode kernel, in the
 * context's precision with the exact arithmetic.  `out` (may be NULL) receives the state after the call. */
int ll_sim_reset(ll_ctx* ctx, int32_t use_pads, double spad_x1, double lpad_x1, ll_sim_view* out);
int ll_sim_update(ll_ctx* ctx, int32_t action, ll_sim_view* out); /* out->done = the reference's return value */
int ll_sim_get(ll_ctx* ctx, ll_sim_view* out);
/* writes x, y, vx, vy, fuel, landing_pad_center_x, the three flags and last_action (the def_readwrite members) */
int ll_sim_set(ll_ctx* ctx, const ll_sim_view* in, ll_sim_view* out);
int ll_sim_step_penalty(ll_ctx* ctx, int32_t action, double* penalty);
/* direction < 0: the one-argument overload; 0/1: the three-argument one, bonus_lr[direction] += tp_landed_lr_bonus on a
 * successful landing (bonus_lr: 2 doubles, in/out, may be NULL) */
int ll_sim_terminal_penalty(ll_ctx* ctx, int32_t steps_taken, int32_t direction, double* penalty, double* bonus_lr);

/* ---- measurement ---------------------------------------------------------------------------------------------*/
int ll_profile(ll_ctx* ctx, int enable);                 /* CUDA-event timing of every kernel class */
int ll_kernel_times_get(ll_ctx* ctx, ll_kernel_times* out, int reset);
/* lander steps (= feedforward calls, the reference's Sum all_member_steps, src/nn_engine_trainer.cpp:519,530) executed by
 * all evaluations since the last reset, accumulated on the device so that timed loops need no per-generation read */
int ll_step_counter(ll_ctx* ctx, double* total_steps, uint64_t* evaluations, int reset);
/* FMA-pipe roofline denominator measured on this device: dependency-free FMA chains for ~`seconds`, MEAN over the loop.
 * precision 32 -> FFMA, 322 -> packed FFMA2 (fma.rn.f32x2), 64 -> DFMA, 197 -> mma.sync.m16n8k16 f16 with rotating operands at one
 * warp per SM sub-partition (the legacy tensor pipe of the mma / mmat episode kernels).  Result in TFLOP/s (2 flop per FMA);
 * sm_mhz_est = SM cycles per nanosecond of one thread over the same loop (clock64 / %globaltimer). */
int ll_microbench_fma(int32_t device, int32_t precision, double seconds, double* tflops, double* sm_mhz_est);
/* tcgen05 (5th-generation tensor core) known-answer tests and issue-rate / hop-latency microbenchmarks on `device`:
 * args[16] = {test, m_instr, n, k, a_mn_major, b_mn_major, swap_lbo_sbo, iters, n_acc, busy_warps, busy_iters, mufu_every,
 * group_threads, all_sms, a_from_tmem, lane_off}; d_out (test 0 only): [128][n] floats, the raw TMEM dump; result[16] per test, see
 * csrc/ll_tc05_probe.cu.  Fails with LL_ENODEVICE on anything but sm_100. */
int ll_tc05_probe(int32_t device, const int32_t* args, float* d_out, double* result);
/* HBM copy bandwidth (read+write bytes / time), GB/s */
int ll_microbench_copy(int32_t device, uint64_t bytes, int32_t iters, double* gbps);

/* ---- allocation-free top-k (no counterpart in the reference) -------------
 * For callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it):
 * the radix select of ll_local_topk / ll_select over the caller's own DEVICE array d_fitness[n], enqueued on `cuda_stream`
 * (NULL = the default stream) of `device` (< 0 = the current one), returning immediately.  d_out[0..k) (DEVICE memory) receives
 * the k smallest (fitness, first_global + position) pairs in ascending (fitness, index) order, bit-identical to ll_local_topk on
 * the same values.  It calls no cudaMalloc / cudaFree / allocating library routine: its scratch (select state, candidates, one
 * 64-bit key per member) is a file-scope device array of LL_TOPK_STATIC_SCRATCH_BYTES that every GPU which loads the library
 * holds until the process exits.  Limits: 1 <= n <= LL_TOPK_STATIC_MAX, 1 <= k <= min(n, 4096); anything larger is rejected with
 * LL_EINVAL and a message naming the limit.  All calls on one GPU share the array (it keeps the last call's contents; the state
 * that must start at zero is rewritten by the first kernel of every call), so calls that overlap on one GPU — different
 * streams, or different threads — race: serialise them. */
#define LL_TOPK_STATIC_MAX (1u << 20)
#define LL_TOPK_STATIC_SCRATCH_BYTES (2048u + 4096u * 16u + 8u * LL_TOPK_STATIC_MAX) /* 8.07 MB per GPU */
int ll_topk_static(int32_t device, void* cuda_stream, const double* d_fitness, uint64_t n, uint64_t first_global, uint32_t k,
                   ll_cand* d_out);

#ifdef __cplusplus
}
#endif
#endif /* LL_GPU_H_ */
