// ll_comm.cu — the per-generation exchange between the GPUs of one box, owned by the context:
//   K2 local top-k  ->  ncclAllGather of top x (fitness, global index)  ->  one-block merge by (fitness, index)
//   K3 row gather (owned rows, zeros elsewhere)  ->  ncclAllReduce over the bit patterns (= exact broadcast from the owners)
//   + an all-reduce of three doubles so that every rank holds the whole population's statistics (what train() logs)
// all enqueued on the context's stream: no host synchronisation between evaluate -> select -> exchange -> breed.
// Replaces, across GPUs, std::sort + UpdateWeightsAndBiases(sorted) (src/nn_engine_trainer.cpp:376-379, 404); the reference's own
// parallelism is the thread pool over members (:460-538), which the population shards map onto.
// NCCL is loaded at run time (dlopen "libnccl.so.2": the copy already mapped by the process, e.g. torch's, else the system's), so
// libll_gpu.so has no link-time dependency on it and single-GPU hosts never touch it.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <mutex>

#include "ll_ops.cuh"

int ll_launch_topk(ll_ctx* ctx, unsigned int k);

namespace
{
    struct nccl_api
    {
        void* handle = nullptr;
        ncclResult_t (*GetUniqueId)(ncclUniqueId*)                                                                       = nullptr;
        ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int)                                                = nullptr;
        ncclResult_t (*CommDestroy)(ncclComm_t)                                                                          = nullptr;
        ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t)                  = nullptr;
        ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t)     = nullptr;
        const char* (*GetErrorString)(ncclResult_t)                                                                      = nullptr;
        std::string why;
    };
    nccl_api g_nccl;
    std::once_flag g_nccl_once;

    const nccl_api* nccl()
    {
        std::call_once(g_nccl_once, [] {
            const char* names[] = {"libnccl.so.2", "libnccl.so"};
            for (const char* n : names)
                if ((g_nccl.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL)) != nullptr) break;
            if (!g_nccl.handle)
            {
                g_nccl.why = std::string("cannot load libnccl.so.2: ") + dlerror();
                return;
            }
#define LL_SYM(field, name)                                                                        \
    if (!(g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(g_nccl.handle, name))))    \
    {                                                                                              \
        g_nccl.why = std::string("libnccl.so.2 lacks ") + name;                                    \
        return;                                                                                    \
    }
            LL_SYM(GetUniqueId, "ncclGetUniqueId");
            LL_SYM(CommInitRank, "ncclCommInitRank");
            LL_SYM(CommDestroy, "ncclCommDestroy");
            LL_SYM(AllGather, "ncclAllGather");
            LL_SYM(AllReduce, "ncclAllReduce");
            LL_SYM(GetErrorString, "ncclGetErrorString");
#undef LL_SYM
        });
        return g_nccl.why.empty() ? &g_nccl : nullptr;
    }

#define LL_NCCL(ctx, call)                                                                                          \
    do {                                                                                                            \
        ncclResult_t r__ = (call);                                                                                  \
        if (r__ != ncclSuccess) return ll_fail((ctx), LL_ECUDA, std::string(#call) + ": " + g_nccl.GetErrorString(r__)); \
    } while (0)

    __device__ __forceinline__ unsigned long long orderable(double d)
    { // ascending order-preserving map, -0.0 folded onto +0.0 (the same key ll_ga.cu's radix select uses)
        if (d == 0.0) d = 0.0;
        const unsigned long long u = (unsigned long long)__double_as_longlong(d);
        return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
    }

    // entries [k, top) of a short shard's candidate block: sort last
    __global__ void ll_cand_pad_kernel(ll_cand* c, unsigned int k, unsigned int top)
    {
        const unsigned int i = k + blockIdx.x * blockDim.x + threadIdx.x;
        if (i < top) c[i].fitness = __longlong_as_double(0x7ff0000000000000ll), c[i].index = ~0ull;
    }

    // One block: bitonic sort of the n = world x top gathered candidates by (fitness, global index); the first `top` become the
    // global elites — the order a single shard's radix select produces, so the result does not depend on the GPU count.
    __global__ void __launch_bounds__(1024) ll_cand_merge_kernel(const ll_cand* __restrict__ all, unsigned int n, unsigned int npow2, unsigned int top,
                                                                  ll_cand* __restrict__ out, unsigned long long* __restrict__ idx)
    {
        extern __shared__ __align__(16) unsigned char smem_raw[];
        unsigned long long* sk = reinterpret_cast<unsigned long long*>(smem_raw);
        unsigned long long* si = sk + npow2;
        for (unsigned int i = threadIdx.x; i < npow2; i += blockDim.x)
        {
            sk[i] = i < n ? orderable(all[i].fitness) : ~0ull;
            si[i] = i < n ? all[i].index : ~0ull;
        }
        __syncthreads();
        for (unsigned int size = 2; size <= npow2; size <<= 1)
            for (unsigned int stride = size >> 1; stride > 0; stride >>= 1)
            {
                for (unsigned int t = threadIdx.x; t < npow2 / 2; t += blockDim.x)
                {
                    const unsigned int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                    const bool up = (lo & size) == 0;
                    const bool gt = sk[lo] > sk[hi] || (sk[lo] == sk[hi] && si[lo] > si[hi]);
                    if (gt == up)
                    {
                        const unsigned long long tk = sk[lo], ti = si[lo];
                        sk[lo] = sk[hi], si[lo] = si[hi];
                        sk[hi] = tk, si[hi] = ti;
                    }
                }
                __syncthreads();
            }
        for (unsigned int i = threadIdx.x; i < top; i += blockDim.x)
        {
            const unsigned long long k = sk[i];
            const unsigned long long u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
            out[i].fitness = __longlong_as_double((long long)u);
            out[i].index   = si[i];
            idx[i]         = si[i];
        }
    }

    // red[0..2] = this shard's sum of fitness, sum of steps, and the lr of the global best if this shard owns it (else 0)
    __global__ void ll_stats_pack_kernel(const ll_stats_dev* st, const ll_cand* best, const double* lr, unsigned long long first_global,
                                         unsigned long long pop_local, double* red)
    {
        if (threadIdx.x != 0 || blockIdx.x != 0) return;
        const unsigned long long g = best[0].index;
        red[0] = st->sum_fitness, red[1] = st->sum_steps;
        red[2] = (g >= first_global && g - first_global < pop_local) ? lr[g - first_global] : 0.0;
        red[3] = 0.0;
    }
    __global__ void ll_stats_unpack_kernel(ll_stats_dev* st, const ll_cand* best, const double* red)
    {
        if (threadIdx.x != 0 || blockIdx.x != 0) return;
        st->best_fitness = best[0].fitness, st->best_index = best[0].index;
        st->sum_fitness = red[0], st->sum_steps = red[1], st->best_lr = red[2];
    }
} // namespace

extern "C" int ll_comm_get_unique_id(void* id_out)
{
    if (!id_out) return ll_fail(nullptr, LL_EINVAL, "ll_comm_get_unique_id: null argument");
    if (!nccl()) return ll_fail(nullptr, LL_ENODEVICE, g_nccl.why);
    ncclUniqueId id;
    LL_NCCL(nullptr, g_nccl.GetUniqueId(&id));
    std::memcpy(id_out, &id, sizeof(id));
    return LL_OK;
}

extern "C" int ll_comm_init(ll_ctx* ctx, const void* unique_id)
{
    if (!ctx || !unique_id) return ll_fail(ctx, LL_EINVAL, "ll_comm_init: null argument");
    if (ctx->world < 2) return ll_fail(ctx, LL_EINVAL, "ll_comm_init: the context was created with world == 1");
    if (ctx->comm) return ll_fail(ctx, LL_EINVAL, "ll_comm_init: the context already owns a communicator");
    if ((uint64_t)ctx->world * ctx->top > LL_MAX_TOP) return ll_fail(ctx, LL_EINVAL, "ll_comm_init: world x top exceeds 4096 candidates");
    if (!nccl()) return ll_fail(ctx, LL_ENODEVICE, g_nccl.why);
    LL_CUDA(ctx, cudaSetDevice(ctx->device));
    ncclUniqueId id;
    std::memcpy(&id, unique_id, sizeof(id));
    ncclComm_t comm = nullptr;
    LL_NCCL(ctx, g_nccl.CommInitRank(&comm, ctx->world, id, ctx->rank));
    ctx->comm = comm;
    LL_CUDA(ctx, cudaMalloc(&ctx->d_cand_all, sizeof(ll_cand) * (size_t)ctx->world * ctx->top));
    LL_CUDA(ctx, cudaMalloc(&ctx->d_red, 4 * sizeof(double)));
    return LL_OK;
}

void ll_comm_release(ll_ctx* ctx)
{
    if (ctx->comm && nccl()) g_nccl.CommDestroy(static_cast<ncclComm_t>(ctx->comm));
    ctx->comm = nullptr;
    if (ctx->d_cand_all) cudaFree(ctx->d_cand_all);
    if (ctx->d_red) cudaFree(ctx->d_red);
    ctx->d_cand_all = nullptr, ctx->d_red = nullptr;
}

extern "C" int ll_comm_destroy(ll_ctx* ctx)
{
    if (!ctx) return ll_fail(nullptr, LL_EINVAL, "null context");
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    ll_comm_release(ctx);
    return LL_OK;
}

// indices_given: d_elite_idx[0..top) already holds the global elite indices (host order, ll_select(sorted_idx)); otherwise they are
// chosen here from the last evaluation.  Leaves the parent matrix in d_elites on every rank.
int ll_comm_select(ll_ctx* ctx, bool indices_given)
{
    if (!ctx->comm) return ll_fail(ctx, LL_EINVAL, "multi-rank selection needs a communicator: call ll_comm_init first");
    ncclComm_t comm        = static_cast<ncclComm_t>(ctx->comm);
    const unsigned int top = (unsigned int)ctx->top;
    if (!indices_given)
    {
        const unsigned int k = (unsigned int)std::min<uint64_t>(ctx->top, ctx->pop_local);
        int rc               = ll_launch_topk(ctx, k);
        if (rc != LL_OK) return rc;
        if (k < top) ll_cand_pad_kernel<<<(top - k + 127) / 128, 128, 0, ctx->stream>>>(ctx->d_cand, k, top);
        LL_NCCL(ctx, g_nccl.AllGather(ctx->d_cand, ctx->d_cand_all, sizeof(ll_cand) * top, ncclChar, comm, ctx->stream));
        const unsigned int n = (unsigned int)ctx->world * top;
        unsigned int np2     = 2;
        while (np2 < n) np2 <<= 1;
        const size_t smem = (size_t)np2 * 16;
        LL_CUDA(ctx, cudaFuncSetAttribute(ll_cand_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ll_cand_merge_kernel<<<1, np2 / 2 < 1024 ? (np2 / 2 < 32 ? 32 : np2 / 2) : 1024, smem, ctx->stream>>>(ctx->d_cand_all, n, np2, top, ctx->d_cand,
                                                                                                             ctx->d_elite_idx);
        // whole-population statistics on every rank
        ll_stats_pack_kernel<<<1, 32, 0, ctx->stream>>>(ctx->d_stats, ctx->d_cand, ctx->d_lr, ctx->first_global, ctx->pop_local, ctx->d_red);
        LL_NCCL(ctx, g_nccl.AllReduce(ctx->d_red, ctx->d_red, 4, ncclDouble, ncclSum, comm, ctx->stream));
        ll_stats_unpack_kernel<<<1, 32, 0, ctx->stream>>>(ctx->d_stats, ctx->d_cand, ctx->d_red);
        ctx->stats_global = true;
    }
    int rc = ll_launch_gather(ctx, ctx->d_elite_idx, top, ctx->d_elites, 1);
    if (rc != LL_OK) return rc;
    const size_t words = (size_t)top * ctx->topo.p_dev;
    LL_NCCL(ctx, g_nccl.AllReduce(ctx->d_elites, ctx->d_elites, words, ctx->precision == 64 ? ncclUint64 : ncclUint32, ncclSum, comm, ctx->stream));
    LL_CUDA(ctx, cudaGetLastError());
    ctx->has_elites = true;
    return LL_OK;
}
