// ll_api.cu — the C ABI of libll_gpu.so (include/ll_gpu.h): context life cycle, parameter transfer, evaluation,
// selection and population creation.  Every entry point is a thin, synchronous-looking wrapper that enqueues the
// kernels of ll_episode.cu / ll_ga.cu on the context's stream.  No exception leaves this file.
#include <algorithm>
#include <cstring>
#include <new>

#include "ll_internal.h"

size_t ll_stats_partial_bytes(unsigned long long pop_local);
int ll_file_write(ll_ctx* ctx, const char* path, const char* name, int digits, const std::vector<double>& elites_canon);
int ll_file_read(const char* path, std::vector<uint64_t>* sizes, uint64_t* pop, uint64_t* top, uint64_t* epochs, int* act, int* strategy,
                 double* ratio, std::string* name, std::vector<double>* elites_canon, std::string* err, uint64_t* generator_word0 = nullptr);

namespace
{
    thread_local std::string g_last_error;

    struct timer_pair
    {
        cudaEvent_t a, b;
        int cls;
    };
    // pending CUDA-event pairs of a profiling context, resolved lazily by ll_kernel_times_get
    struct prof_state
    {
        std::vector<timer_pair> pending;
        std::vector<timer_pair> pool;
    };
    prof_state* prof_of(ll_ctx* ctx);

    enum
    {
        CLS_EPISODE = 0,
        CLS_SELECT  = 1,
        CLS_GATHER  = 2,
        CLS_MUTATE  = 3,
        CLS_OTHER   = 4
    };

    struct scope
    {
        ll_ctx* ctx;
        int cls;
        timer_pair tp{};
        bool on = false;
        scope(ll_ctx* c, int cl, uint64_t launches);
        ~scope();
    };
} // namespace

struct ll_prof_holder
{
    prof_state st;
};

namespace
{
    prof_state* prof_of(ll_ctx* ctx)
    {
        if (!ctx->prof) ctx->prof = new ll_prof_holder();
        return &ctx->prof->st;
    }
    void prof_drop(ll_ctx* ctx)
    {
        if (!ctx->prof) return;
        for (auto& t : ctx->prof->st.pending) cudaEventDestroy(t.a), cudaEventDestroy(t.b);
        for (auto& t : ctx->prof->st.pool) cudaEventDestroy(t.a), cudaEventDestroy(t.b);
        delete ctx->prof;
        ctx->prof = nullptr;
    }

    scope::scope(ll_ctx* c, int cl, uint64_t launches) : ctx(c), cls(cl)
    {
        switch (cl)
        {
        case CLS_EPISODE: c->times.episode_launches += launches; break;
        case CLS_SELECT: c->times.select_launches += launches; break;
        case CLS_GATHER: c->times.gather_launches += launches; break;
        case CLS_MUTATE: c->times.mutate_launches += launches; break;
        default: c->times.other_launches += launches; break;
        }
        if (!c->profile) return;
        prof_state* ps = prof_of(c);
        if (!ps->pool.empty())
        {
            tp = ps->pool.back();
            ps->pool.pop_back();
        }
        else if (cudaEventCreate(&tp.a) != cudaSuccess || cudaEventCreate(&tp.b) != cudaSuccess) return;
        tp.cls = cl;
        on     = cudaEventRecord(tp.a, c->stream) == cudaSuccess;
    }
    scope::~scope()
    {
        if (!on) return;
        cudaEventRecord(tp.b, ctx->stream);
        prof_of(ctx)->pending.push_back(tp);
    }

    template <typename T> void canon_to_dev(const ll_ctx* ctx, const std::vector<int>& d2c, const T* canon, T* dev)
    {
        for (int d = 0; d < ctx->topo.p_dev; ++d) dev[d] = d2c[(size_t)d] >= 0 ? canon[d2c[(size_t)d]] : T(0);
    }
    template <typename T> void dev_to_canon(const ll_ctx* ctx, const std::vector<int>& d2c, const T* dev, T* canon)
    {
        for (int d = 0; d < ctx->topo.p_dev; ++d)
            if (d2c[(size_t)d] >= 0) canon[d2c[(size_t)d]] = dev[d];
    }

    int ensure_pinned(ll_ctx* ctx, size_t bytes)
    {
        if (ctx->h_pin_bytes >= bytes) return LL_OK;
        if (ctx->h_pin) cudaFreeHost(ctx->h_pin);
        ctx->h_pin = nullptr, ctx->h_pin_bytes = 0;
        LL_CUDA(ctx, cudaMallocHost(&ctx->h_pin, bytes));
        ctx->h_pin_bytes = bytes;
        return LL_OK;
    }

    bool owns(const ll_ctx* ctx, uint64_t g) { return g >= ctx->first_global && g - ctx->first_global < ctx->pop_local; }

    __global__ void ll_cand_to_idx_kernel(const ll_cand* c, unsigned int n, unsigned long long* idx)
    {
        const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n) idx[i] = c[i].index;
    }
} // namespace

static const std::vector<int>& d2c_of(const ll_ctx* ctx) { return ctx->d2c; }

void ll_set_error(ll_ctx* ctx, const std::string& msg)
{
    g_last_error = msg;
    if (ctx) ctx->err = msg;
}
int ll_fail(ll_ctx* ctx, int code, const std::string& msg)
{
    ll_set_error(ctx, msg);
    return code;
}

extern "C" int ll_abi_version(void) { return LL_ABI_VERSION; }
extern "C" const char* ll_last_error(const ll_ctx* ctx) { return ctx && !ctx->err.empty() ? ctx->err.c_str() : g_last_error.c_str(); }

#define LL_CHECK_CTX(ctx)                                                                                            \
    do {                                                                                                             \
        if (!(ctx)) return ll_fail(nullptr, LL_EINVAL, "null context");                                              \
        cudaSetDevice((ctx)->device);                                                                                \
    } while (0)
// kernels of ll_launch_topk: keys + 12 radix passes + collect + sort, or candidates + sort for populations of at most LL_MAX_TOP members
static int topk_launches(const ll_ctx* ctx) { return ctx->pop_local <= LL_MAX_TOP ? 2 : 15; }

#define LL_TRY(call)                                                                                                 \
    do {                                                                                                             \
        int rc__ = (call);                                                                                           \
        if (rc__ != LL_OK) return rc__;                                                                              \
    } while (0)

extern "C" void ll_destroy(ll_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    prof_drop(ctx);
    ll_comm_release(ctx);
    void* ptrs[] = {ctx->d_params, ctx->d_elites,   ctx->d_dev2canon, ctx->d_layouts, ctx->d_fitness, ctx->d_lr,        ctx->d_steps,
                    ctx->d_ep_steps, ctx->d_ep_fit, ctx->d_queue,     ctx->d_stats,   ctx->d_partials, ctx->d_sel,      ctx->d_keys,
                    ctx->d_cand,   ctx->d_elite_idx, ctx->d_trace,    ctx->d_scratch, ctx->d_sim, ctx->d_part, ctx->d_resume, ctx->d_breed};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (ctx->h_pin) cudaFreeHost(ctx->h_pin);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int ll_create(ll_ctx** out, const ll_env_cfg* env, const ll_net_cfg* net)
{
    if (!out || !env || !net || !net->layer_sizes) return ll_fail(nullptr, LL_EINVAL, "ll_create: null argument");
    *out = nullptr;
    if (net->n_layers < 2 || net->n_layers > LL_MAX_LAYERS + 1) return ll_fail(nullptr, LL_EINVAL, "ll_create: n_layers must be 2..17");
    if (net->precision != 32 && net->precision != 64) return ll_fail(nullptr, LL_EINVAL, "ll_create: precision must be 32 or 64");
    if (net->mode != LL_MODE_EXACT && net->mode != LL_MODE_FAST) return ll_fail(nullptr, LL_EINVAL, "ll_create: bad mode");
    if (net->mode == LL_MODE_FAST && net->precision != 32) return ll_fail(nullptr, LL_EINVAL, "ll_create: LL_MODE_FAST requires precision 32");
    if (net->activation_id != LL_ACT_SIGMOID && net->activation_id != LL_ACT_TANH) return ll_fail(nullptr, LL_EINVAL, "ll_create: bad activation_id");
    if (net->layer_sizes[0] != 5) return ll_fail(nullptr, LL_EINVAL, "ll_create: the input layer must have 5 units (get_state)");
    for (int i = 0; i < net->n_layers; ++i)
        if (net->layer_sizes[i] == 0 || net->layer_sizes[i] > LL_MAX_WIDTH) return ll_fail(nullptr, LL_EINVAL, "ll_create: layer width must be 1..1024");
    if (net->world < 1 || net->rank < 0 || net->rank >= net->world) return ll_fail(nullptr, LL_EINVAL, "ll_create: bad rank/world");
    if (net->population == 0 || net->top == 0 || net->top > net->population || net->top > LL_MAX_TOP)
        return ll_fail(nullptr, LL_EINVAL, "ll_create: need 1 <= top <= min(population, 4096)");
    if (net->population >= (1ull << 32)) return ll_fail(nullptr, LL_EINVAL, "ll_create: population must be < 2^32");

    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    {
        cudaGetLastError();
        return ll_fail(nullptr, LL_ENODEVICE, "ll_create: no CUDA device (this library has no CPU fallback)");
    }
    int device = net->device;
    if (device < 0 && cudaGetDevice(&device) != cudaSuccess) return ll_fail(nullptr, LL_ENODEVICE, "ll_create: cudaGetDevice failed");
    if (device >= n_dev) return ll_fail(nullptr, LL_ENODEVICE, "ll_create: device ordinal out of range");

    ll_ctx* ctx = new (std::nothrow) ll_ctx();
    if (!ctx) return ll_fail(nullptr, LL_EINVAL, "ll_create: out of host memory");
    ctx->device = device;
    auto bail   = [&](int rc) {
        std::string msg = ctx->err;
        ll_destroy(ctx);
        g_last_error = msg;
        return rc;
    };
#define LL_CUDA_B(call)                                                                                              \
    do {                                                                                                             \
        cudaError_t e__ = (call);                                                                                    \
        if (e__ != cudaSuccess)                                                                                      \
        {                                                                                                            \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e__);                                          \
            return bail(LL_ECUDA);                                                                                   \
        }                                                                                                            \
    } while (0)
    LL_CUDA_B(cudaSetDevice(device));
    cudaDeviceProp prop;
    LL_CUDA_B(cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    std::snprintf(ctx->device_name, sizeof(ctx->device_name), "%.63s", prop.name);

    ctx->env_cfg = *env;
    ctx->sizes.assign(net->layer_sizes, net->layer_sizes + net->n_layers);
    ctx->pop_global = net->population, ctx->top = net->top, ctx->seed = net->seed;
    ctx->precision = net->precision, ctx->mode = net->mode, ctx->rank = net->rank, ctx->world = net->world;
    ctx->elt = net->precision == 64 ? 8 : 4;
    if (net->sigma_init > 0.0) ctx->sigma_init = net->sigma_init;
    if (net->sigma_mut > 0.0) ctx->sigma_mut = net->sigma_mut;
    // contiguous shards [rank*pop/world, (rank+1)*pop/world)
    ctx->first_global = net->population * (uint64_t)net->rank / (uint64_t)net->world;
    ctx->pop_local    = net->population * (uint64_t)(net->rank + 1) / (uint64_t)net->world - ctx->first_global;
    if (ctx->pop_local == 0)
    {
        ctx->err = "ll_create: empty shard";
        return bail(LL_EINVAL);
    }
    std::vector<int> d2c;
    ll_build_topo(ctx->sizes, net->activation_id, &ctx->topo, &d2c);
    std::snprintf(ctx->kernel_name, sizeof(ctx->kernel_name), "%s", ll_episode_kernel_name(ctx, 0));

    LL_CUDA_B(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    ctx->own_stream  = true;
    const size_t row = (size_t)ctx->topo.p_dev * ctx->elt;
    LL_CUDA_B(cudaMalloc(&ctx->d_params, row * ctx->pop_local));
    LL_CUDA_B(cudaMemsetAsync(ctx->d_params, 0, row * ctx->pop_local, ctx->stream));
    LL_CUDA_B(cudaMalloc(&ctx->d_elites, row * ctx->top));
    LL_CUDA_B(cudaMemsetAsync(ctx->d_elites, 0, row * ctx->top, ctx->stream));
    LL_CUDA_B(cudaMalloc(&ctx->d_dev2canon, sizeof(int) * d2c.size()));
    LL_CUDA_B(cudaMalloc(&ctx->d_breed, sizeof(unsigned int) * ctx->pop_local));
    LL_CUDA_B(cudaMemcpy(ctx->d_dev2canon, d2c.data(), sizeof(int) * d2c.size(), cudaMemcpyHostToDevice));
    LL_CUDA_B(cudaMalloc(&ctx->d_fitness, sizeof(double) * ctx->pop_local));
    LL_CUDA_B(cudaMalloc(&ctx->d_lr, sizeof(double) * ctx->pop_local));
    LL_CUDA_B(cudaMalloc(&ctx->d_steps, sizeof(double) * ctx->pop_local));
    LL_CUDA_B(cudaMalloc(&ctx->d_queue, 64));
    LL_CUDA_B(cudaMalloc(&ctx->d_stats, sizeof(ll_stats_dev)));
    LL_CUDA_B(cudaMemsetAsync(ctx->d_stats, 0, sizeof(ll_stats_dev), ctx->stream));
    LL_CUDA_B(cudaMalloc(&ctx->d_partials, ll_stats_partial_bytes(ctx->pop_local)));
    LL_CUDA_B(cudaMalloc(&ctx->d_sel, sizeof(ll_select_state)));
    LL_CUDA_B(cudaMalloc(&ctx->d_keys, sizeof(unsigned long long) * ctx->pop_local));
    LL_CUDA_B(cudaMalloc(&ctx->d_cand, sizeof(ll_cand) * LL_MAX_TOP));
    LL_CUDA_B(cudaMalloc(&ctx->d_elite_idx, sizeof(unsigned long long) * LL_MAX_TOP));
    LL_CUDA_B(cudaMalloc(&ctx->d_scratch, 1024));
    LL_CUDA_B(cudaMemsetAsync(ctx->d_scratch, 0, 1024, ctx->stream));
    LL_CUDA_B(cudaStreamSynchronize(ctx->stream));
#undef LL_CUDA_B
    ctx->d2c = std::move(d2c);
    *out = ctx;
    return LL_OK;
}

extern "C" int ll_get_info(const ll_ctx* ctx, ll_info* out)
{
    if (!ctx || !out) return ll_fail(nullptr, LL_EINVAL, "ll_get_info: null argument");
    std::memset(out, 0, sizeof(*out));
    out->params = (uint64_t)ctx->topo.p_canon, out->params_padded = (uint64_t)ctx->topo.p_dev;
    out->pop_global = ctx->pop_global, out->pop_local = ctx->pop_local, out->first_global = ctx->first_global, out->top = ctx->top;
    out->precision = ctx->precision, out->mode = ctx->mode, out->rank = ctx->rank, out->world = ctx->world;
    out->device = ctx->device, out->sm_count = ctx->sm_count;
    std::snprintf(out->device_name, sizeof(out->device_name), "%s", ctx->device_name);
    std::snprintf(out->kernel, sizeof(out->kernel), "%s", ctx->kernel_name);
    return LL_OK;
}

extern "C" int ll_set_stream(ll_ctx* ctx, void* cuda_stream)
{
    LL_CHECK_CTX(ctx);
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    ctx->stream     = static_cast<cudaStream_t>(cuda_stream);
    ctx->own_stream = false;
    return LL_OK;
}

extern "C" int ll_synchronize(ll_ctx* ctx)
{
    LL_CHECK_CTX(ctx);
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LL_OK;
}

// ---- GA ---------------------------------------------------------------------------------------------------------
extern "C" int ll_set_population_strategy(ll_ctx* ctx, int strategy, double ratio)
{
    LL_CHECK_CTX(ctx);
    if (strategy != LL_STRATEGY_ALL_FROM_ELITES && strategy != LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION)
        return ll_fail(ctx, LL_EINVAL, "ll_set_population_strategy: unknown strategy");
    if (!(ratio >= 0.0 && ratio <= 1.0)) return ll_fail(ctx, LL_EINVAL, "ll_set_population_strategy: ratio must be in [0,1]");
    ctx->strategy = strategy, ctx->ratio = ratio;
    return LL_OK;
}

extern "C" int ll_set_sigmas(ll_ctx* ctx, double sigma_init, double sigma_mut)
{
    LL_CHECK_CTX(ctx);
    if (sigma_init > 0.0) ctx->sigma_init = sigma_init;
    if (sigma_mut >= 0.0) ctx->sigma_mut = sigma_mut;
    return LL_OK;
}

int ll_launch_create_population_gen(ll_ctx* ctx, int elitism, uint64_t gen);

extern "C" int ll_create_population(ll_ctx* ctx, int elitism)
{
    LL_CHECK_CTX(ctx);
    scope s(ctx, CLS_MUTATE, ctx->precision == 32 ? 2 : 1); // float: breeding decisions + the flat kernel
    LL_TRY(ll_launch_create_population_gen(ctx, elitism, ctx->has_elites ? ctx->epochs + 1 : ctx->epochs));
    ctx->have_results = false;
    return LL_OK;
}

extern "C" int ll_update_epochs(ll_ctx* ctx, uint64_t n)
{
    LL_CHECK_CTX(ctx);
    ctx->epochs += n;
    return LL_OK;
}
extern "C" int ll_get_epochs(const ll_ctx* ctx, uint64_t* out)
{
    if (!ctx || !out) return ll_fail(nullptr, LL_EINVAL, "ll_get_epochs: null argument");
    *out = ctx->epochs;
    return LL_OK;
}
extern "C" int ll_set_epochs(ll_ctx* ctx, uint64_t epochs)
{
    LL_CHECK_CTX(ctx);
    ctx->epochs = epochs;
    return LL_OK;
}

static int install_elites_from_idx(ll_ctx* ctx, unsigned int n)
{ // d_elite_idx[0..n) -> d_elites (world == 1: every index is local)
    scope s(ctx, CLS_GATHER, 1);
    LL_TRY(ll_launch_gather(ctx, ctx->d_elite_idx, n, ctx->d_elites, 1));
    ctx->has_elites = true;
    return LL_OK;
}

static int device_topk(ll_ctx* ctx, unsigned int k)
{
    if (!ctx->have_results) return ll_fail(ctx, LL_EINVAL, "selection needs the results of ll_evaluate");
    scope s(ctx, CLS_SELECT, topk_launches(ctx) + 1);
    LL_TRY(ll_launch_topk(ctx, k));
    ll_cand_to_idx_kernel<<<(k + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_cand, k, ctx->d_elite_idx);
    LL_CUDA(ctx, cudaGetLastError());
    return LL_OK;
}

extern "C" int ll_select(ll_ctx* ctx, const uint64_t* sorted_idx, uint64_t n_sorted, uint64_t* elite_idx_out)
{
    LL_CHECK_CTX(ctx);
    if (ctx->world != 1 && !ctx->comm)
        return ll_fail(ctx, LL_EINVAL, "ll_select on a shard needs a communicator (ll_comm_init), or use ll_local_topk/ll_gather_elites/ll_set_elites");
    const unsigned int k = (unsigned int)ctx->top;
    if (sorted_idx)
    {
        if (n_sorted < k) return ll_fail(ctx, LL_EINVAL, "ll_select: need at least `top` sorted indices");
        for (unsigned int i = 0; i < k; ++i)
            if (sorted_idx[i] >= ctx->pop_global) return ll_fail(ctx, LL_EINVAL, "ll_select: index out of range");
        LL_TRY(ensure_pinned(ctx, sizeof(uint64_t) * k));
        std::memcpy(ctx->h_pin, sorted_idx, sizeof(uint64_t) * k);
        LL_CUDA(ctx, cudaMemcpyAsync(ctx->d_elite_idx, ctx->h_pin, sizeof(uint64_t) * k, cudaMemcpyHostToDevice, ctx->stream));
        LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    else if (ctx->world == 1) { LL_TRY(device_topk(ctx, k)); }
    if (ctx->world == 1) { LL_TRY(install_elites_from_idx(ctx, k)); }
    else
    { // every rank makes the same call: all-gather of the local top-k + merge (no indices given), row gather + all-reduce
        if (!sorted_idx && !ctx->have_results) return ll_fail(ctx, LL_EINVAL, "selection needs the results of ll_evaluate");
        scope s(ctx, CLS_SELECT, sorted_idx ? 1 : topk_launches(ctx) + 4);
        LL_TRY(ll_comm_select(ctx, sorted_idx != nullptr));
    }
    if (elite_idx_out)
    {
        LL_TRY(ensure_pinned(ctx, sizeof(uint64_t) * k));
        LL_CUDA(ctx, cudaMemcpyAsync(ctx->h_pin, ctx->d_elite_idx, sizeof(uint64_t) * k, cudaMemcpyDeviceToHost, ctx->stream));
        LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        std::memcpy(elite_idx_out, ctx->h_pin, sizeof(uint64_t) * k);
    }
    return LL_OK;
}

extern "C" int ll_local_topk(ll_ctx* ctx, ll_cand* out, uint64_t* n_out)
{
    LL_CHECK_CTX(ctx);
    if (!out || !n_out) return ll_fail(ctx, LL_EINVAL, "ll_local_topk: null argument");
    const unsigned int k = (unsigned int)std::min<uint64_t>(ctx->top, ctx->pop_local);
    LL_TRY(device_topk(ctx, k));
    LL_TRY(ensure_pinned(ctx, sizeof(ll_cand) * k));
    LL_CUDA(ctx, cudaMemcpyAsync(ctx->h_pin, ctx->d_cand, sizeof(ll_cand) * k, cudaMemcpyDeviceToHost, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    std::memcpy(out, ctx->h_pin, sizeof(ll_cand) * k);
    *n_out = k;
    return LL_OK;
}

extern "C" uint64_t ll_elite_buffer_bytes(const ll_ctx* ctx) { return ctx ? (uint64_t)ctx->top * ctx->topo.p_dev * ctx->elt : 0; }

extern "C" int ll_gather_elites(ll_ctx* ctx, const uint64_t* elite_idx, uint64_t n, void* device_buf)
{
    LL_CHECK_CTX(ctx);
    if (!elite_idx || !device_buf || n == 0 || n > ctx->top) return ll_fail(ctx, LL_EINVAL, "ll_gather_elites: bad argument");
    LL_TRY(ensure_pinned(ctx, sizeof(uint64_t) * n));
    std::memcpy(ctx->h_pin, elite_idx, sizeof(uint64_t) * n);
    LL_CUDA(ctx, cudaMemcpyAsync(ctx->d_elite_idx, ctx->h_pin, sizeof(uint64_t) * n, cudaMemcpyHostToDevice, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // h_pin is reused by later calls
    scope s(ctx, CLS_GATHER, 1);
    LL_TRY(ll_launch_gather(ctx, ctx->d_elite_idx, (unsigned int)n, device_buf, 1));
    return LL_OK;
}

extern "C" int ll_set_elites(ll_ctx* ctx, const void* device_buf, uint64_t n)
{
    LL_CHECK_CTX(ctx);
    if (!device_buf || n != ctx->top) return ll_fail(ctx, LL_EINVAL, "ll_set_elites: need exactly `top` rows");
    LL_CUDA(ctx, cudaMemcpyAsync(ctx->d_elites, device_buf, ll_elite_buffer_bytes(ctx), cudaMemcpyDeviceToDevice, ctx->stream));
    ctx->has_elites = true;
    return LL_OK;
}

// ---- evaluation ---------------------------------------------------------------------------------------------------
extern "C" int ll_set_layouts(ll_ctx* ctx, const int32_t* spad_x1, const int32_t* lpad_x1, const uint64_t* directions, int32_t n)
{
    LL_CHECK_CTX(ctx);
    if (!spad_x1 || !lpad_x1 || !directions || n < 1 || n > LL_MAX_LAYOUTS) return ll_fail(ctx, LL_EINVAL, "ll_set_layouts: need 1..4096 layouts");
    for (int i = 0; i < n; ++i)
        if (directions[i] > 1) return ll_fail(ctx, LL_EINVAL, "ll_set_layouts: direction must be 0 or 1"); // assert at src/game_logic.cpp:356
    const size_t lay_bytes = ctx->precision == 64 ? sizeof(ll_layout<double>) : sizeof(ll_layout<float>);
    if (n > ctx->layouts_cap)
    {
        LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (ctx->d_layouts) cudaFree(ctx->d_layouts);
        ctx->d_layouts = nullptr, ctx->layouts_cap = 0;
        LL_CUDA(ctx, cudaMalloc(&ctx->d_layouts, lay_bytes * (size_t)n));
        ctx->layouts_cap = n;
    }
    const size_t need = (size_t)ctx->pop_local * (size_t)n;
    if (need > ctx->ep_cap)
    {
        LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (ctx->d_ep_steps) cudaFree(ctx->d_ep_steps);
        if (ctx->d_ep_fit) cudaFree(ctx->d_ep_fit);
        ctx->d_ep_steps = nullptr, ctx->d_ep_fit = nullptr, ctx->ep_cap = 0;
        if (need * 12 <= (size_t)8 << 30)
        {
            LL_CUDA(ctx, cudaMalloc(&ctx->d_ep_steps, need * sizeof(int)));
            LL_CUDA(ctx, cudaMalloc(&ctx->d_ep_fit, need * sizeof(double)));
            ctx->ep_cap = need;
        }
    }
    LL_TRY(ensure_pinned(ctx, lay_bytes * (size_t)n));
    for (int i = 0; i < n; ++i)
    {
        if (ctx->precision == 64) ll_make_layout<double>(ctx->env_cfg, spad_x1[i], lpad_x1[i], (int)directions[i], static_cast<ll_layout<double>*>(ctx->h_pin) + i);
        else ll_make_layout<float>(ctx->env_cfg, spad_x1[i], lpad_x1[i], (int)directions[i], static_cast<ll_layout<float>*>(ctx->h_pin) + i);
    }
    {
        // schedule hint for the slot-refilling kernels: entry i names the layout to start i-th, longest expected flight
        // (horizontal distance between the pads) first — a member's episodes are then packed onto its slots LPT-style
        std::vector<int> order((size_t)n);
        std::vector<double> dist((size_t)n);
        for (int i = 0; i < n; ++i) order[(size_t)i] = i, dist[(size_t)i] = std::abs((double)spad_x1[i] - (double)lpad_x1[i]);
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return dist[(size_t)x] > dist[(size_t)y]; });
        for (int i = 0; i < n; ++i)
        {
            if (ctx->precision == 64) static_cast<ll_layout<double>*>(ctx->h_pin)[i].next = order[(size_t)i];
            else static_cast<ll_layout<float>*>(ctx->h_pin)[i].next = order[(size_t)i];
        }
    }
    LL_CUDA(ctx, cudaMemcpyAsync(ctx->d_layouts, ctx->h_pin, lay_bytes * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->n_layouts = n;
    std::snprintf(ctx->kernel_name, sizeof(ctx->kernel_name), "%s", ll_episode_kernel_name(ctx, n));
    return LL_OK;
}

extern "C" int ll_evaluate_device(ll_ctx* ctx, int32_t max_steps)
{
    LL_CHECK_CTX(ctx);
    if (ctx->n_layouts < 1) return ll_fail(ctx, LL_EINVAL, "ll_evaluate_device: call ll_set_layouts first");
    if (max_steps < 1 || max_steps >= (1 << 28)) return ll_fail(ctx, LL_EINVAL, "ll_evaluate_device: max_steps must be in [1, 2^28)");
    ll_episode_args a{};
    a.params = ctx->d_params, a.layouts = ctx->d_layouts, a.n_layouts = ctx->n_layouts, a.max_steps = max_steps;
    a.n_members = ctx->pop_local, a.fitness = ctx->d_fitness, a.lr = ctx->d_lr, a.steps = ctx->d_steps;
    const bool ep = ctx->ep_cap >= (size_t)ctx->pop_local * (size_t)ctx->n_layouts;
    a.ep_steps = ep ? ctx->d_ep_steps : nullptr, a.ep_fit = ep ? ctx->d_ep_fit : nullptr;
    a.queue = ctx->d_queue;
    {
        scope s(ctx, CLS_EPISODE, 1);
        LL_TRY(ll_launch_episode(ctx, a));
    }
    {
        scope s(ctx, CLS_SELECT, 2);
        LL_TRY(ll_launch_stats(ctx));
    }
    ctx->have_results = true;
    ctx->stats_global = false;
    return LL_OK;
}

// the tcgen05 episode kernel bounds every tensor-pipe wait and raises this flag (the word next to the work counter) instead of hanging
static int check_kernel_flag(ll_ctx* ctx)
{
    unsigned int q[2] = {0u, 0u};
    LL_CUDA(ctx, cudaMemcpyAsync(q, ctx->d_queue, sizeof(q), cudaMemcpyDeviceToHost, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (q[1] != 0u) return ll_fail(ctx, LL_ECUDA, "episode kernel: a tensor-pipe wait timed out (results are invalid)");
    return LL_OK;
}

extern "C" int ll_read_results(ll_ctx* ctx, double* fitness, double* last_gen_lr, double* steps, ll_gen_stats* stats)
{
    LL_CHECK_CTX(ctx);
    if (!ctx->have_results) return ll_fail(ctx, LL_EINVAL, "ll_read_results: nothing evaluated yet");
    LL_TRY(check_kernel_flag(ctx));
    const size_t n = ctx->pop_local * sizeof(double);
    LL_TRY(ensure_pinned(ctx, 3 * n + sizeof(ll_stats_dev)));
    char* h = static_cast<char*>(ctx->h_pin);
    if (fitness) LL_CUDA(ctx, cudaMemcpyAsync(h, ctx->d_fitness, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (last_gen_lr) LL_CUDA(ctx, cudaMemcpyAsync(h + n, ctx->d_lr, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (steps) LL_CUDA(ctx, cudaMemcpyAsync(h + 2 * n, ctx->d_steps, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (stats) LL_CUDA(ctx, cudaMemcpyAsync(h + 3 * n, ctx->d_stats, sizeof(ll_stats_dev), cudaMemcpyDeviceToHost, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (fitness) std::memcpy(fitness, h, n);
    if (last_gen_lr) std::memcpy(last_gen_lr, h + n, n);
    if (steps) std::memcpy(steps, h + 2 * n, n);
    if (stats)
    {
        ll_stats_dev sd;
        std::memcpy(&sd, h + 3 * n, sizeof(sd));
        stats->best_fitness = sd.best_fitness, stats->best_index = sd.best_index, stats->sum_fitness = sd.sum_fitness;
        stats->sum_steps = sd.sum_steps, stats->best_lr = sd.best_lr, stats->members = ctx->stats_global ? ctx->pop_global : ctx->pop_local;
    }
    return LL_OK;
}

extern "C" int ll_read_episodes(ll_ctx* ctx, int32_t* meta, double* fitness)
{
    LL_CHECK_CTX(ctx);
    if (!ctx->have_results) return ll_fail(ctx, LL_EINVAL, "ll_read_episodes: nothing evaluated yet");
    const size_t n = (size_t)ctx->pop_local * (size_t)ctx->n_layouts;
    if (ctx->ep_cap < n) return ll_fail(ctx, LL_EINVAL, "ll_read_episodes: per-episode buffers were not allocated (population x layouts too large)");
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (meta) LL_CUDA(ctx, cudaMemcpy(meta, ctx->d_ep_steps, n * sizeof(int), cudaMemcpyDeviceToHost));
    if (fitness) LL_CUDA(ctx, cudaMemcpy(fitness, ctx->d_ep_fit, n * sizeof(double), cudaMemcpyDeviceToHost));
    return LL_OK;
}

extern "C" int ll_evaluate(ll_ctx* ctx, const int32_t* spad_x1, const int32_t* lpad_x1, const uint64_t* directions, int32_t n_layouts,
                           int32_t max_steps, double* fitness, double* last_gen_lr, double* steps, ll_gen_stats* stats)
{
    LL_TRY(ll_set_layouts(ctx, spad_x1, lpad_x1, directions, n_layouts));
    LL_TRY(ll_evaluate_device(ctx, max_steps));
    return ll_read_results(ctx, fitness, last_gen_lr, steps, stats);
}

extern "C" int ll_step_generation(ll_ctx* ctx, int32_t max_steps, int elitism, ll_gen_stats* stats)
{
    LL_CHECK_CTX(ctx);
    if (ctx->world != 1 && !ctx->comm) return ll_fail(ctx, LL_EINVAL, "ll_step_generation on a shard needs a communicator (ll_comm_init)");
    LL_TRY(ll_evaluate_device(ctx, max_steps));
    if (ctx->world == 1)
    {
        LL_TRY(device_topk(ctx, (unsigned int)ctx->top));
        LL_TRY(install_elites_from_idx(ctx, (unsigned int)ctx->top));
    }
    else
    { // K2 -> NCCL all-gather -> merge -> K3 -> NCCL all-reduce, all on the context's stream: no host round trip
        scope s(ctx, CLS_SELECT, topk_launches(ctx) + 4);
        LL_TRY(ll_comm_select(ctx, false));
    }
    {
        scope s(ctx, CLS_MUTATE, ctx->precision == 32 ? 2 : 1); // float: breeding decisions + the flat kernel
        LL_TRY(ll_launch_create_population_gen(ctx, elitism, ctx->epochs + 1));
    }
    ctx->epochs += 1;
    if (stats) LL_TRY(ll_read_results(ctx, nullptr, nullptr, nullptr, stats));
    return LL_OK;
}

extern "C" int ll_trace_episode(ll_ctx* ctx, uint64_t member, int32_t spad_x1, int32_t lpad_x1, uint64_t direction, int32_t max_steps,
                                ll_step_rec* rec, int32_t cap, ll_episode_result* result)
{
    LL_CHECK_CTX(ctx);
    if (!owns(ctx, member)) return ll_fail(ctx, LL_ENOTOWNED, "ll_trace_episode: member is not in this shard");
    if (direction > 1 || max_steps < 1 || cap < 0 || !result) return ll_fail(ctx, LL_EINVAL, "ll_trace_episode: bad argument");
    if (cap > ctx->trace_cap)
    {
        LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (ctx->d_trace) cudaFree(ctx->d_trace);
        ctx->d_trace = nullptr, ctx->trace_cap = 0;
        LL_CUDA(ctx, cudaMalloc(&ctx->d_trace, sizeof(ll_step_rec) * (size_t)cap));
        ctx->trace_cap = cap;
    }
    char* sc = static_cast<char*>(ctx->d_scratch);
    LL_TRY(ensure_pinned(ctx, 256));
    if (ctx->precision == 64) ll_make_layout<double>(ctx->env_cfg, spad_x1, lpad_x1, (int)direction, static_cast<ll_layout<double>*>(ctx->h_pin));
    else ll_make_layout<float>(ctx->env_cfg, spad_x1, lpad_x1, (int)direction, static_cast<ll_layout<float>*>(ctx->h_pin));
    LL_CUDA(ctx, cudaMemcpyAsync(sc + 512, ctx->h_pin, 64, cudaMemcpyHostToDevice, ctx->stream));
    ll_episode_args a{};
    a.params    = static_cast<char*>(ctx->d_params) + (member - ctx->first_global) * (size_t)ctx->topo.p_dev * ctx->elt;
    a.layouts   = sc + 512;
    a.n_layouts = 1, a.max_steps = max_steps, a.n_members = 1;
    a.fitness = reinterpret_cast<double*>(sc + 256), a.lr = reinterpret_cast<double*>(sc + 264), a.steps = reinterpret_cast<double*>(sc + 272);
    a.queue = reinterpret_cast<unsigned int*>(sc + 640);
    a.trace = cap > 0 ? ctx->d_trace : reinterpret_cast<ll_step_rec*>(sc + 768) /* non-null marker, cap 0 */, a.trace_cap = cap;
    a.trace_result = reinterpret_cast<ll_episode_result*>(sc + 128);
    {
        scope s(ctx, CLS_OTHER, 1);
        LL_TRY(ll_launch_episode(ctx, a));
    }
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    LL_CUDA(ctx, cudaMemcpy(result, sc + 128, sizeof(ll_episode_result), cudaMemcpyDeviceToHost));
    const int n = std::min(cap, result->steps);
    if (rec && n > 0) LL_CUDA(ctx, cudaMemcpy(rec, ctx->d_trace, sizeof(ll_step_rec) * (size_t)n, cudaMemcpyDeviceToHost));
    return LL_OK;
}

extern "C" int ll_feedforward_index(ll_ctx* ctx, const void* state, uint64_t member, uint64_t* action)
{
    LL_CHECK_CTX(ctx);
    if (!state || !action) return ll_fail(ctx, LL_EINVAL, "ll_feedforward_index: null argument");
    if (!owns(ctx, member)) return ll_fail(ctx, LL_ENOTOWNED, "ll_feedforward_index: member is not in this shard");
    char* sc = static_cast<char*>(ctx->d_scratch);
    LL_CUDA(ctx, cudaMemcpyAsync(sc, state, 5 * ctx->elt, cudaMemcpyHostToDevice, ctx->stream));
    const void* row = static_cast<char*>(ctx->d_params) + (member - ctx->first_global) * (size_t)ctx->topo.p_dev * ctx->elt;
    {
        scope s(ctx, CLS_OTHER, 1);
        LL_TRY(ll_launch_feedforward(ctx, row, sc, reinterpret_cast<unsigned long long*>(sc + 64)));
    }
    unsigned long long a = 0;
    LL_CUDA(ctx, cudaMemcpyAsync(&a, sc + 64, 8, cudaMemcpyDeviceToHost, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *action = a;
    return LL_OK;
}

// ---- parameters -----------------------------------------------------------------------------------------------------
static int xfer_members(ll_ctx* ctx, uint64_t first, uint64_t count, void* params, bool to_device)
{
    if (!params) return ll_fail(ctx, LL_EINVAL, "null parameter buffer");
    if (count == 0) return LL_OK;
    if (!owns(ctx, first) || !owns(ctx, first + count - 1)) return ll_fail(ctx, LL_ENOTOWNED, "member range is not inside this shard");
    const std::vector<int>& d2c = d2c_of(ctx);
    const size_t pd = (size_t)ctx->topo.p_dev, pc = (size_t)ctx->topo.p_canon;
    const uint64_t chunk = std::max<uint64_t>(1, ((size_t)32 << 20) / (pd * ctx->elt));
    LL_TRY(ensure_pinned(ctx, std::min<uint64_t>(chunk, count) * pd * ctx->elt));
    for (uint64_t done = 0; done < count; done += chunk)
    {
        const uint64_t n = std::min<uint64_t>(chunk, count - done);
        char* dptr       = static_cast<char*>(ctx->d_params) + (first - ctx->first_global + done) * pd * ctx->elt;
        if (to_device)
        {
            for (uint64_t m = 0; m < n; ++m)
            {
                if (ctx->precision == 64) canon_to_dev<double>(ctx, d2c, static_cast<const double*>(params) + (done + m) * pc, static_cast<double*>(ctx->h_pin) + m * pd);
                else canon_to_dev<float>(ctx, d2c, static_cast<const float*>(params) + (done + m) * pc, static_cast<float*>(ctx->h_pin) + m * pd);
            }
            LL_CUDA(ctx, cudaMemcpyAsync(dptr, ctx->h_pin, n * pd * ctx->elt, cudaMemcpyHostToDevice, ctx->stream));
            LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        }
        else
        {
            LL_CUDA(ctx, cudaMemcpyAsync(ctx->h_pin, dptr, n * pd * ctx->elt, cudaMemcpyDeviceToHost, ctx->stream));
            LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            for (uint64_t m = 0; m < n; ++m)
            {
                if (ctx->precision == 64) dev_to_canon<double>(ctx, d2c, static_cast<const double*>(ctx->h_pin) + m * pd, static_cast<double*>(params) + (done + m) * pc);
                else dev_to_canon<float>(ctx, d2c, static_cast<const float*>(ctx->h_pin) + m * pd, static_cast<float*>(params) + (done + m) * pc);
            }
        }
    }
    if (to_device) ctx->have_results = false;
    return LL_OK;
}

extern "C" int ll_get_member(ll_ctx* ctx, uint64_t member, void* params)
{
    LL_CHECK_CTX(ctx);
    return xfer_members(ctx, member, 1, params, false);
}
extern "C" int ll_set_member(ll_ctx* ctx, uint64_t member, const void* params)
{
    LL_CHECK_CTX(ctx);
    return xfer_members(ctx, member, 1, const_cast<void*>(params), true);
}
extern "C" int ll_set_members(ll_ctx* ctx, uint64_t first, uint64_t count, const void* params)
{
    LL_CHECK_CTX(ctx);
    return xfer_members(ctx, first, count, const_cast<void*>(params), true);
}
extern "C" int ll_get_members(ll_ctx* ctx, uint64_t first, uint64_t count, void* params)
{
    LL_CHECK_CTX(ctx);
    return xfer_members(ctx, first, count, params, false);
}

extern "C" int ll_mark_elites(ll_ctx* ctx)
{
    LL_CHECK_CTX(ctx);
    if (ctx->world == 1 || ctx->rank == 0)
    {
        if (ctx->pop_local < ctx->top) return ll_fail(ctx, LL_EINVAL, "ll_mark_elites: slots [0, top) must lie in rank 0's shard");
        LL_CUDA(ctx, cudaMemcpyAsync(ctx->d_elites, ctx->d_params, ll_elite_buffer_bytes(ctx), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    else return ll_fail(ctx, LL_ENOTOWNED, "ll_mark_elites: only the rank owning slots [0, top) can do this; broadcast with ll_set_elites");
    ctx->has_elites = true;
    return LL_OK;
}

// ---- weight file ---------------------------------------------------------------------------------------------------------
extern "C" int ll_serialize(ll_ctx* ctx, const char* path, const char* name, int32_t digits)
{
    LL_CHECK_CTX(ctx);
    if (!path) return ll_fail(ctx, LL_EINVAL, "ll_serialize: null path");
    if (!ctx->has_elites) return ll_fail(ctx, LL_EINVAL, "ll_serialize: no elite set yet (run a selection or ll_mark_elites)");
    // the file stores the parent set (slots 0..top-1 after UpdateWeightsAndBiases), which every rank holds in d_elites
    const size_t pd = (size_t)ctx->topo.p_dev, pc = (size_t)ctx->topo.p_canon;
    std::vector<char> host(ll_elite_buffer_bytes(ctx));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    LL_CUDA(ctx, cudaMemcpy(host.data(), ctx->d_elites, host.size(), cudaMemcpyDeviceToHost));
    const std::vector<int>& d2c = d2c_of(ctx);
    std::vector<double> canon(ctx->top * pc);
    for (uint64_t m = 0; m < ctx->top; ++m)
        for (size_t d = 0; d < pd; ++d)
            if (d2c[d] >= 0)
                canon[m * pc + (size_t)d2c[d]] = ctx->precision == 64 ? reinterpret_cast<const double*>(host.data())[m * pd + d]
                                                                      : (double)reinterpret_cast<const float*>(host.data())[m * pd + d];
    return ll_file_write(ctx, path, name ? name : "lunar_lander", digits > 0 ? digits : 6, canon);
}

extern "C" int ll_peek_weight_file(const char* path, uint64_t* sizes, int32_t cap, int32_t* n_sizes, uint64_t* pop, uint64_t* top,
                                   uint64_t* epochs, int32_t* act, int32_t* strategy, double* ratio, char* name, int32_t name_cap)
{
    if (!path) return ll_fail(nullptr, LL_EINVAL, "ll_peek_weight_file: null path");
    std::vector<uint64_t> s;
    uint64_t p = 0, t = 0, e = 0;
    int a = 1, st = 2;
    double r = 0.0;
    std::string nm, err;
    const int rc = ll_file_read(path, &s, &p, &t, &e, &a, &st, &r, &nm, nullptr, &err);
    if (rc != LL_OK) return ll_fail(nullptr, rc, err);
    if (n_sizes) *n_sizes = (int32_t)s.size();
    if (sizes)
        for (int i = 0; i < cap && i < (int)s.size(); ++i) sizes[i] = s[(size_t)i];
    if (pop) *pop = p;
    if (top) *top = t;
    if (epochs) *epochs = e;
    if (act) *act = a;
    if (strategy) *strategy = st;
    if (ratio) *ratio = r;
    if (name && name_cap > 0) std::snprintf(name, (size_t)name_cap, "%s", nm.c_str());
    return LL_OK;
}

extern "C" int ll_deserialize(ll_ctx* ctx, const char* path)
{
    LL_CHECK_CTX(ctx);
    if (!path) return ll_fail(ctx, LL_EINVAL, "ll_deserialize: null path");
    std::vector<uint64_t> s;
    uint64_t p = 0, t = 0, e = 0;
    int a = 1, st = 2;
    double r = 0.0;
    std::string nm, err;
    std::vector<double> canon;
    uint64_t gen0 = ~0ull;
    const int rc = ll_file_read(path, &s, &p, &t, &e, &a, &st, &r, &nm, &canon, &err, &gen0);
    if (rc != LL_OK) return ll_fail(ctx, rc, err);
    if (s != ctx->sizes) return ll_fail(ctx, LL_EINVAL, "ll_deserialize: the file's vSize does not match this context's topology");
    if (t != ctx->top) return ll_fail(ctx, LL_EINVAL, "ll_deserialize: the file's nTop does not match this context");
    if (p != ctx->pop_global) return ll_fail(ctx, LL_EINVAL, "ll_deserialize: the file's nPopSize does not match this context's population");
    if (a != ctx->topo.activation) return ll_fail(ctx, LL_EINVAL, "ll_deserialize: the file's act does not match this context");
    const size_t pd = (size_t)ctx->topo.p_dev, pc = (size_t)ctx->topo.p_canon;
    const std::vector<int>& d2c = d2c_of(ctx);
    std::vector<char> host(ll_elite_buffer_bytes(ctx), 0);
    for (uint64_t m = 0; m < ctx->top; ++m)
        for (size_t d = 0; d < pd; ++d)
            if (d2c[d] >= 0)
            {
                const double v = canon[m * pc + (size_t)d2c[d]];
                if (ctx->precision == 64) reinterpret_cast<double*>(host.data())[m * pd + d] = v;
                else reinterpret_cast<float*>(host.data())[m * pd + d] = (float)v;
            }
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    LL_CUDA(ctx, cudaMemcpy(ctx->d_elites, host.data(), host.size(), cudaMemcpyHostToDevice));
    ctx->has_elites = true, ctx->epochs = e, ctx->strategy = st, ctx->ratio = r;
    if (gen0 != ~0ull) ctx->seed = gen0 & 0xffffffffull; // the generator line carries the GA key of the run that wrote the file
    // the file holds the parent set only; slots [top, pop) are rebuilt exactly as the CreatePopulation that
    // preceded the save built them (generation key = stored epoch count)
    scope sc(ctx, CLS_MUTATE, ctx->precision == 32 ? 2 : 1);
    LL_TRY(ll_launch_create_population_gen(ctx, 1, e));
    ctx->have_results = false;
    return LL_OK;
}

// ---- measurement ---------------------------------------------------------------------------------------------------------
extern "C" int ll_step_counter(ll_ctx* ctx, double* total_steps, uint64_t* evaluations, int reset)
{
    LL_CHECK_CTX(ctx);
    ll_stats_dev sd;
    LL_CUDA(ctx, cudaMemcpyAsync(&sd, ctx->d_stats, sizeof(sd), cudaMemcpyDeviceToHost, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (total_steps) *total_steps = sd.steps_accum;
    if (evaluations) *evaluations = sd.evals_accum;
    if (reset)
    {
        sd.steps_accum = 0.0, sd.evals_accum = 0ull;
        LL_CUDA(ctx, cudaMemcpyAsync(ctx->d_stats, &sd, sizeof(sd), cudaMemcpyHostToDevice, ctx->stream));
        LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return LL_OK;
}

extern "C" int ll_profile(ll_ctx* ctx, int enable)
{
    LL_CHECK_CTX(ctx);
    ctx->profile = enable != 0;
    return LL_OK;
}

extern "C" int ll_kernel_times_get(ll_ctx* ctx, ll_kernel_times* out, int reset)
{
    LL_CHECK_CTX(ctx);
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    prof_state* ps = prof_of(ctx);
    for (auto& t : ps->pending)
    {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess)
        {
            switch (t.cls)
            {
            case CLS_EPISODE: ctx->times.episode_ms += ms; break;
            case CLS_SELECT: ctx->times.select_ms += ms; break;
            case CLS_GATHER: ctx->times.gather_ms += ms; break;
            case CLS_MUTATE: ctx->times.mutate_ms += ms; break;
            default: break;
            }
        }
        ps->pool.push_back(t);
    }
    ps->pending.clear();
    if (out) *out = ctx->times;
    if (reset) ctx->times = ll_kernel_times{};
    return LL_OK;
}
