// ll_internal.h — private types shared by the translation units of libll_gpu.so (not part of the ABI).
#ifndef LL_INTERNAL_H_
#define LL_INTERNAL_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/ll_gpu.h"

#define LL_MAX_LAYERS 16 /* weight layers (n_sizes - 1) */
#define LL_MAX_WIDTH 1024
#define LL_MAX_LAYOUTS 4096
#define LL_MAX_TOP 4096

// Network topology in the device ("forward") layout.  A member row holds, for each layer l:
//   Wt_l[k][out_pad_l]  k = 0..in_l-1   (k-major = transposed w.r.t. the weight file's [out][in]; pad columns are 0)
//   b_l[out_pad_l]
// out_pad_l = out_l rounded up to a multiple of 8, so every k-row starts on a 32-byte boundary: two LDS.128 fetch
// the 8 consecutive outputs of a register tile in the fast kernel (float), and every tile width of the exact
// kernel (16 / 8 / 4 outputs) divides a row.
struct ll_topo
{
    int n_layers;
    int max_width; // max over layers of out_pad (and the input width)
    int p_canon;   // P = sum(in*out + out)
    int p_dev;     // elements per member row in HBM / shared memory (multiple of 4)
    int in[LL_MAX_LAYERS], out[LL_MAX_LAYERS], out_pad[LL_MAX_LAYERS];
    int w_off[LL_MAX_LAYERS], b_off[LL_MAX_LAYERS]; // element offsets inside a device row
    int activation;                                  // LL_ACT_*
};

// Every constant GameLogicCpp<T> derives from the configuration (src/game_logic.cpp:48-122), already
// converted to T on the host with the same static_cast<T>(double) the reference applies.
template <typename T> struct ll_env
{
    T pad_y, ground, max_vx, max_vy, ay, one_minus_mu_x, one_minus_mu_y;
    T lander_w, lpad_w, width, height, max_fuel;
    T y0, vx0, vy0;
    T sp_dist, sp_action;
    T tp_steps, tp_dist, tp_landed_bonus, tp_landed_lr_bonus, tp_fuel_bonus, tp_crashed, tp_crash_v_mag, tp_no_fuel;
    // fast mode only: reciprocals so that get_state() multiplies instead of dividing
    T inv_max_vx, inv_max_vy, inv_width, inv_height, inv_max_fuel;
};

// Per-layout values reset(spad_x1, lpad_x1) derives (src/game_logic.cpp:143-154, :101-122), in T.
template <typename T> struct ll_layout
{
    T init_x, lpad_x1, lpad_right, center_x;
    int dir;
    int next; // schedule hint (ll_set_layouts): entry i holds the index of the layout to START i-th, longest expected flight first
};

struct ll_select_state
{
    unsigned long long prefix_hi;
    unsigned int prefix_lo;
    unsigned int k_rem;
    unsigned int ticket;
    unsigned int n_cand;
    unsigned int hist[256];
};

struct ll_stats_dev
{
    double best_fitness, sum_fitness, sum_steps, best_lr;
    unsigned long long best_index;
    double steps_accum;               // running total of lander steps over evaluations (ll_step_counter)
    unsigned long long evals_accum;   // number of evaluations accumulated
};

struct ll_ctx
{
    // configuration
    ll_env_cfg env_cfg;
    std::vector<uint64_t> sizes;
    ll_topo topo;
    uint64_t pop_global = 0, pop_local = 0, first_global = 0, top = 0, seed = 0, epochs = 0;
    int precision = 32, mode = LL_MODE_EXACT, rank = 0, world = 1, device = 0, sm_count = 0;
    int strategy = LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION;
    double ratio = 0.0, sigma_init = 1.0, sigma_mut = 0.1;
    bool has_elites = false, have_results = false;
    size_t elt = 4; // sizeof(T)
    char device_name[64] = {0};
    char kernel_name[64] = {0};

    // device memory
    void* d_params   = nullptr; // [pop_local][p_dev] T
    void* d_elites   = nullptr; // [top][p_dev] T
    int* d_dev2canon = nullptr; // [p_dev]
    unsigned int* d_breed = nullptr; // [pop_local] per-member breeding decision of K4 (ll_pop_meta_kernel): elite copy / fresh / parent slot
    void* d_layouts  = nullptr; // [n_layouts] ll_layout<T>
    int n_layouts = 0, layouts_cap = 0;
    double *d_fitness = nullptr, *d_lr = nullptr, *d_steps = nullptr; // [pop_local]
    int* d_ep_steps = nullptr;                                        // [pop_local][n_layouts] steps | flags<<24.. (see kernel)
    double* d_ep_fit = nullptr;                                       // [pop_local][n_layouts]
    size_t ep_cap   = 0;
    unsigned int* d_queue   = nullptr;
    ll_stats_dev* d_stats   = nullptr;
    double* d_partials      = nullptr; // stats partials
    ll_select_state* d_sel  = nullptr;
    unsigned long long* d_keys = nullptr; // [pop_local] orderable fitness keys
    ll_cand* d_cand         = nullptr;    // [LL_MAX_TOP] sorted candidates
    unsigned long long* d_elite_idx = nullptr; // [top] global indices of the current elites
    ll_step_rec* d_trace    = nullptr;
    int trace_cap = 0;
    void* d_scratch = nullptr; // small scratch (feedforward state, episode result)
    void* d_part    = nullptr; // ll_episode_mma.cu: per-(member, chunk) partial sums + arrival counters (small populations)
    size_t part_cap = 0;
    void* d_resume  = nullptr; // ll_episode_tc05.cu: hand-over records + per-member counts + member list
    size_t resume_cap = 0;
    void* d_sim     = nullptr; // ll_sim.cu: the single stepped lander (state, view, penalties), allocated on first use
    bool sim_ready = false, sim_custom = false;
    double sim_lpad_x1 = 0.0, sim_init_x = 0.0;

    // pinned host staging
    void* h_pin = nullptr;
    size_t h_pin_bytes = 0;

    cudaStream_t stream = nullptr;
    bool own_stream = false;

    // profiling
    bool profile = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    ll_kernel_times times = {};

    // host copy of dev2canon and the CUDA-event profiling state live WITH the context (one host thread per context, any number of
    // contexts per process: the multi-GPU trainer drives one context per thread)
    std::vector<int> d2c;
    struct ll_prof_holder* prof = nullptr;

    // multi-rank exchange owned by the context (ll_comm.cu): NCCL communicator + scratch
    void* comm = nullptr;            // ncclComm_t
    ll_cand* d_cand_all = nullptr;   // [world * top] gathered candidates
    double* d_red = nullptr;         // [4] all-reduce scratch: sum fitness, sum steps, lr of the best (owner only), unused
    bool stats_global = false;       // d_stats holds whole-population statistics (after a communicator-backed selection)

    std::string err;
};

// multi-rank selection behind the ABI (ll_comm.cu)
int ll_comm_select(ll_ctx* ctx, bool indices_given); // local top-k -> all-gather -> merge -> gather rows -> all-reduce; fills d_elites
void ll_comm_release(ll_ctx* ctx);

// ---- error plumbing ------------------------------------------------------------------------------------------
void ll_set_error(ll_ctx* ctx, const std::string& msg);
int ll_fail(ll_ctx* ctx, int code, const std::string& msg);
#define LL_CUDA(ctx, call)                                                                                           \
    do {                                                                                                             \
        cudaError_t e__ = (call);                                                                                    \
        if (e__ != cudaSuccess)                                                                                      \
            return ll_fail((ctx), LL_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e__));                    \
    } while (0)

// ---- kernel launchers (one per translation unit) ---------------------------------------------------------------
// episode kernels (ll_episode.cu)
struct ll_episode_args
{
    const void* params;     // [members][p_dev]
    const void* layouts;    // ll_layout<T>[n_layouts]
    int n_layouts;
    int max_steps;
    uint64_t n_members;     // members in `params`
    double *fitness, *lr, *steps; // [n_members]
    int* ep_steps;          // [n_members][n_layouts] or null
    double* ep_fit;         // [n_members][n_layouts] or null
    unsigned int* queue;    // work counter (zeroed by the launcher)
    ll_step_rec* trace;     // non-null: record every step of (member 0 of params, layout 0)
    int trace_cap;
    ll_episode_result* trace_result;
};
int ll_launch_episode(ll_ctx* ctx, const ll_episode_args& a);
int ll_launch_episode_fast(ll_ctx* ctx, const ll_episode_args& a); // ll_episode_fast.cu
bool ll_fast_applicable(const ll_ctx* ctx, int n_layouts);
int ll_launch_episode_mma(ll_ctx* ctx, const ll_episode_args& a); // ll_episode_mma.cu
bool ll_mma_applicable(const ll_ctx* ctx, int n_layouts);
int ll_launch_episode_mmat(ll_ctx* ctx, const ll_episode_args& a); // ll_episode_mmat.cu
bool ll_mmat_applicable(const ll_ctx* ctx, int n_layouts);
// An episode handed from the tcgen05 tile kernel to the 16-row mma.sync kernel when only a few of a member's episodes are still
// running (ll_episode_tc05.cu -> ll_episode_mma.cu): everything sim_task keeps per episode (src/nn_engine_trainer.cpp:486-511).
struct ll_resume_rec
{
    float x, y, vx, vy, fuel;
    int layout, steps, pad_;
    double acc;
};
#define LL_RESUME_SLOTS 16 /* rows of the straggler kernel's tile */
#define LL_RESUME_MAX 64   /* most episodes a member can hand over (they refill the 16 rows as episodes end) */
struct ll_resume_args
{
    ll_resume_rec* rec;            // [n_members][stride]
    int stride;                    // records per member (= the hand-over threshold)
    unsigned int* count;           // [n_members] records of the member (0: nothing handed over)
    unsigned int* list;            // members with records, in arrival order
    unsigned int* n;               // number of entries of `list`
};
int ll_launch_episode_mma_resume(ll_ctx* ctx, const ll_episode_args& a, const ll_resume_args& r); // ll_episode_mma.cu: 16-row tile
int ll_launch_episode_mmat_resume(ll_ctx* ctx, const ll_episode_args& a, const ll_resume_args& r); // ll_episode_mmat.cu: 8-column tile
int ll_launch_episode_tc05(ll_ctx* ctx, const ll_episode_args& a); // ll_episode_tc05.cu
bool ll_tc05_applicable(const ll_ctx* ctx, int n_layouts);
// default choice of the tcgen05 kernel: where it applies (17..64 layouts, the instantiated topologies) and the shard fills its 6 member
// slots per SM; measured 1.18-1.25 x the mma.sync kernel on the benchmark workload together with the straggler hand-over
#define LL_TC05_DEFAULT(ctx, n_layouts) ((ctx)->pop_local >= 6ull * (unsigned long long)(ctx)->sm_count)
int ll_launch_episode_exactw(ll_ctx* ctx, const ll_episode_args& a); // ll_episode_exactw.cu: exact mode, one warp per episode (small populations)
bool ll_exactw_applicable(const ll_ctx* ctx, const ll_episode_args& a);
int ll_launch_episode_warp(ll_ctx* ctx, const ll_episode_args& a); // ll_episode_warp.cu
bool ll_warp_applicable(const ll_ctx* ctx, int n_layouts);
bool ll_warp_per_episode(const ll_ctx* ctx, unsigned long long n_members, int n_layouts, bool have_episode_arrays); // small populations: warp per episode
#define LL_WARP_MAX_LAYOUTS 1 /* default choice of the warp-per-member kernel: a single layout per member (configs[1]) */
int ll_launch_feedforward(ll_ctx* ctx, const void* member_row, const void* d_state, unsigned long long* d_action);
const char* ll_episode_kernel_name(const ll_ctx* ctx, int n_layouts);

// GA kernels (ll_ga.cu)
// generation key `gen`: epochs for the initial population / a reload, epochs + 1 for the children of a generation
int ll_launch_create_population_gen(ll_ctx* ctx, int elitism, uint64_t gen);
size_t ll_stats_partial_bytes(unsigned long long pop_local);
int ll_launch_stats(ll_ctx* ctx);
int ll_launch_topk(ll_ctx* ctx, unsigned int k); // fills d_cand[0..k) sorted by (fitness, global index)
int ll_launch_gather(ll_ctx* ctx, const unsigned long long* d_idx, unsigned int n, void* d_out, int zero_unowned);

// host helpers (ll_host.cpp)
template <typename T> void ll_make_env(const ll_env_cfg& c, int fast, ll_env<T>* out);
template <typename T> void ll_make_layout(const ll_env_cfg& c, int spad_x1, int lpad_x1, int dir, ll_layout<T>* out);
void ll_build_topo(const std::vector<uint64_t>& sizes, int activation, ll_topo* t, std::vector<int>* dev2canon);

#endif
