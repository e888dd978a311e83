// ll_ga.cu — the GA side of a generation, all on the device:
//   K0/K4  ll_create_population_kernel : initial Gaussians / elite copy + mutated children + random injection
//   K2     ll_stats_*                  : best / sum fitness / sum steps (what train() logs)
//   K2     ll_keys_kernel, ll_select_hist_kernel, ll_select_collect_kernel, ll_select_sort_kernel :
//          exact top-k of (fitness, global index) by a 12-digit radix select — no full sort of the population
//   K3     ll_gather_rows_kernel       : winners' rows -> parent matrix
//
// Replaces (reference call sites; the ma-libs sources are not in the tree):
//   std::sort of indices by fitness                          src/nn_engine_trainer.cpp:376-379
//   nn::ANN_MLP_GA<T>::UpdateWeightsAndBiases(sorted)        src/nn_engine_trainer.cpp:404
//   nn::ANN_MLP_GA<T>::CreatePopulation(elitism)             src/nn_engine_trainer.cpp:94, 405
// Random numbers are Philox4x32-10 counters keyed by (seed, generation, GLOBAL member, canonical parameter pair)
// (ll_philox.h), so children do not depend on the GPU count, the shard or the launch geometry.
// All of these kernels are HBM-bound streaming kernels: coalesced row reads/writes, one pass each.
#include <algorithm>
#include <mutex>

#include "ll_ops.cuh"
#include "ll_philox.h"

namespace
{
    // ------------------------------------------------------------------------------------------------ K0 / K4
    struct pop_args
    {
        void* params;        // [pop_local][p_dev]
        const void* elites;  // [top][p_dev]
        const int* dev2canon;
        unsigned long long first_global, pop_global, pop_local;
        unsigned int seed, gen, top;
        unsigned long long first_child;  // global slots below this are elite copies
        unsigned long long n_children, n_inject; // ll_is_injected(slot - first_child, ...) marks the fresh random individuals
        int from_elites;
        int p_canon, p_dev;
        double sigma_init, sigma_mut;
    };

    template <typename T> struct exact_ops;
    template <> struct exact_ops<float> : ll_ops<float, LL_MODE_EXACT> {};
    template <> struct exact_ops<double> : ll_ops<double, LL_MODE_EXACT> {};

    // one block per member row
    template <typename T> __global__ void __launch_bounds__(256) ll_create_population_kernel(const __grid_constant__ pop_args a)
    {
        using O = exact_ops<T>;
        extern __shared__ __align__(16) unsigned char smem_raw[];
        T* g = reinterpret_cast<T*>(smem_raw); // canonical-order N(0,1) draws of this member, already cast to T
        for (unsigned long long ml = blockIdx.x; ml < a.pop_local; ml += gridDim.x)
        {
            const unsigned long long mg = a.first_global + ml;
            T* row                      = static_cast<T*>(a.params) + ml * (size_t)a.p_dev;
            if (a.from_elites && mg < a.first_child)
            { // elitism: slot r keeps elite r unchanged
                const T* src = static_cast<const T*>(a.elites) + mg * (size_t)a.p_dev;
                for (int d = threadIdx.x; d < a.p_dev; d += blockDim.x) row[d] = src[d];
                continue;
            }
            const int n_pairs = (a.p_canon + 1) / 2;
            for (int p = threadIdx.x; p < n_pairs; p += blockDim.x)
            {
                double n0, n1;
                ll_normal_pair(a.seed, a.gen, (uint32_t)mg, (uint32_t)p, &n0, &n1);
                g[2 * p] = (T)n0;
                if (2 * p + 1 < a.p_canon) g[2 * p + 1] = (T)n1;
            }
            __syncthreads();
            const bool inject = !a.from_elites || ll_is_injected(mg - a.first_child, a.n_children, a.n_inject);
            const T s_init = (T)a.sigma_init, s_mut = (T)a.sigma_mut;
            const T* parent = nullptr;
            if (!inject) parent = static_cast<const T*>(a.elites) + (size_t)ll_parent_pick(a.seed, a.gen, (uint32_t)mg, a.top) * a.p_dev;
            for (int d = threadIdx.x; d < a.p_dev; d += blockDim.x)
            {
                const int c = a.dev2canon[d];
                T v         = T(0);
                if (c >= 0) v = inject ? O::mul(s_init, g[c]) : O::add(parent[d], O::mul(s_mut, g[c]));
                row[d] = v;
            }
            __syncthreads();
        }
    }

    // float populations, two launches:
    //   ll_pop_meta_kernel: one thread per member decides how its row is bred — copy of elite r (elitism), fresh N(0, sigma_init^2)
    //     individual, or child of ONE uniformly picked elite (ll_parent_pick: a Philox block of its own) — into a 4-byte word;
    //   ll_create_population_f32_kernel: one WARP per member row (persistent warps stride over the rows).  ONE THREAD = 4 consecutive
    //     device-order parameters = one Philox block = four Box-Muller variates in float, the parent's float4 (L2-resident), one 16-byte
    //     store; a warp writes 512 contiguous bytes per iteration.  No shared memory and no barrier (a block per row needed two
    //     __syncthreads per row to broadcast the parent — a quarter of the warp samples sat there — and left 4.5 % of the lanes idle in the
    //     ragged last iteration of a 2 203-quad row on 256 threads; flattening (member, quad) over the grid instead cost 25 more
    //     instructions of 64-bit index arithmetic per piece and was slower).
    // FAST (LL_MODE_FAST): the normals come from the special-function unit (ll_normal4_f32_fast) instead of the host-reproducible routine.
    constexpr unsigned int BREED_COPY = 0x80000000u, BREED_FRESH = 0x40000000u; // low bits: elite slot (copy source / parent)

    __global__ void __launch_bounds__(256) ll_pop_meta_kernel(const __grid_constant__ pop_args a, unsigned int* __restrict__ breed)
    {
        const unsigned long long ml = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
        if (ml >= a.pop_local) return;
        const unsigned long long mg = a.first_global + ml;
        unsigned int w;
        if (a.from_elites && mg < a.first_child) w = BREED_COPY | (unsigned int)mg; // elitism: slot r keeps elite r unchanged
        else if (!a.from_elites || ll_is_injected(mg - a.first_child, a.n_children, a.n_inject)) w = BREED_FRESH;
        else w = ll_parent_pick(a.seed, a.gen, (uint32_t)mg, a.top);
        breed[ml] = w;
    }

    template <bool FAST>
    __global__ void __launch_bounds__(256) ll_create_population_f32_kernel(const __grid_constant__ pop_args a, const unsigned int* __restrict__ breed)
    {
        using O = exact_ops<float>;
        const int qpr        = a.p_dev >> 2; // quads per row (p_dev is a multiple of 8)
        const int4* d2c4     = reinterpret_cast<const int4*>(a.dev2canon);
        const float4* elites = reinterpret_cast<const float4*>(a.elites);
        const float s_init = (float)a.sigma_init, s_mut = (float)a.sigma_mut;
        const int lane = threadIdx.x & 31;
        const unsigned long long n_warps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
        for (unsigned long long ml = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ml < a.pop_local; ml += n_warps)
        {
            const unsigned int w = breed[ml];
            const uint32_t mg    = (uint32_t)(a.first_global + ml);
            float4* row          = reinterpret_cast<float4*>(static_cast<float*>(a.params) + ml * (size_t)a.p_dev);
            const float4* src    = elites + (size_t)(w & 0x3fffffffu) * qpr; // copy source or parent (unused for fresh individuals)
            if (w & BREED_COPY)
            {
                for (int qd = lane; qd < qpr; qd += 32) row[qd] = src[qd];
                continue;
            }
            const bool fresh = (w & BREED_FRESH) != 0u;
#pragma unroll 2
            for (int qd = lane; qd < qpr; qd += 32)
            {
                float g[4];
                if constexpr (FAST) ll_normal4_f32_fast(a.seed, a.gen, mg, (uint32_t)qd, g);
                else ll_normal4_f32(a.seed, a.gen, mg, (uint32_t)qd, g);
                const int4 c = d2c4[qd];
                float4 v;
                if (fresh) v = make_float4(O::mul(s_init, g[0]), O::mul(s_init, g[1]), O::mul(s_init, g[2]), O::mul(s_init, g[3]));
                else
                {
                    const float4 pr = src[qd];
                    v = make_float4(O::add(pr.x, O::mul(s_mut, g[0])), O::add(pr.y, O::mul(s_mut, g[1])), O::add(pr.z, O::mul(s_mut, g[2])),
                                    O::add(pr.w, O::mul(s_mut, g[3])));
                }
                if (c.x < 0) v.x = 0.f; // padding columns of the forward layout stay zero
                if (c.y < 0) v.y = 0.f;
                if (c.z < 0) v.z = 0.f;
                if (c.w < 0) v.w = 0.f;
                row[qd] = v;
            }
        }
    }

    // ------------------------------------------------------------------------------------------------ K2 stats
    // Deterministic two-level reduction: block b owns elements [b*CH, (b+1)*CH); partials are combined in block
    // order by the last kernel.  Ties on the minimum resolve to the lowest index.
    constexpr int ST_THREADS = 256;
    constexpr int ST_CHUNK   = 4096;

    struct st_partial
    {
        double sum_f, sum_s, best;
        unsigned long long idx;
    };

    __global__ void __launch_bounds__(ST_THREADS) ll_stats_partial_kernel(const double* __restrict__ fit, const double* __restrict__ steps,
                                                                          unsigned long long n, st_partial* out)
    {
        __shared__ double s_f[ST_THREADS], s_s[ST_THREADS], s_b[ST_THREADS];
        __shared__ unsigned long long s_i[ST_THREADS];
        const unsigned long long base = (unsigned long long)blockIdx.x * ST_CHUNK;
        double f = 0.0, s = 0.0, best = __longlong_as_double(0x7ff0000000000000ll);
        unsigned long long bi = ~0ull;
        for (int i = threadIdx.x; i < ST_CHUNK; i += ST_THREADS)
        {
            const unsigned long long e = base + i;
            if (e < n)
            {
                const double v = fit[e];
                f += v, s += steps[e];
                if (v < best || bi == ~0ull) best = v, bi = e;
            }
        }
        s_f[threadIdx.x] = f, s_s[threadIdx.x] = s, s_b[threadIdx.x] = best, s_i[threadIdx.x] = bi;
        __syncthreads();
        for (int w = ST_THREADS / 2; w > 0; w >>= 1)
        {
            if (threadIdx.x < w)
            {
                s_f[threadIdx.x] += s_f[threadIdx.x + w];
                s_s[threadIdx.x] += s_s[threadIdx.x + w];
                const double ob             = s_b[threadIdx.x + w];
                const unsigned long long oi = s_i[threadIdx.x + w];
                if (oi != ~0ull && (s_i[threadIdx.x] == ~0ull || ob < s_b[threadIdx.x] || (ob == s_b[threadIdx.x] && oi < s_i[threadIdx.x])))
                    s_b[threadIdx.x] = ob, s_i[threadIdx.x] = oi;
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) out[blockIdx.x] = st_partial{s_f[0], s_s[0], s_b[0], s_i[0]};
    }

    __global__ void ll_stats_final_kernel(const st_partial* __restrict__ p, unsigned int n_part, const double* __restrict__ lr,
                                          unsigned long long first_global, ll_stats_dev* out)
    {
        if (threadIdx.x != 0 || blockIdx.x != 0) return;
        double f = 0.0, s = 0.0, best = 0.0;
        unsigned long long bi = ~0ull;
        for (unsigned int i = 0; i < n_part; ++i)
        {
            f += p[i].sum_f, s += p[i].sum_s;
            if (p[i].idx != ~0ull && (bi == ~0ull || p[i].best < best)) best = p[i].best, bi = p[i].idx;
        }
        out->sum_fitness = f, out->sum_steps = s, out->best_fitness = best;
        out->best_index = bi == ~0ull ? 0ull : first_global + bi;
        out->best_lr    = bi == ~0ull ? 0.0 : lr[bi];
        out->steps_accum += s;
        out->evals_accum += 1ull;
    }

    // ------------------------------------------------------------------------------------------------ K2 top-k
    // order-preserving map double -> u64 (ascending); -0.0 is folded onto +0.0 so that it ties with it, as the
    // reference's `<` comparator does.
    __device__ __forceinline__ unsigned long long ll_orderable(double d)
    {
        if (d == 0.0) d = 0.0;
        const unsigned long long u = (unsigned long long)__double_as_longlong(d);
        return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
    }
    __device__ __forceinline__ double ll_from_orderable(unsigned long long k)
    {
        const unsigned long long u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
        return __longlong_as_double((long long)u);
    }

    __global__ void ll_keys_kernel(const double* __restrict__ fit, unsigned long long n, unsigned long long* __restrict__ keys,
                                   ll_select_state* st, unsigned int k)
    {
        const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n) keys[i] = ll_orderable(fit[i]);
        if (i == 0)
        {
            st->prefix_hi = 0ull, st->prefix_lo = 0u, st->k_rem = k, st->ticket = 0u, st->n_cand = 0u;
        }
        if (blockIdx.x == 0 && threadIdx.x < 256) st->hist[threadIdx.x] = 0u;
    }

    // composite key = (orderable fitness : 64, local index : 32), most significant digit first, 8 bits per pass
    __device__ __forceinline__ bool ll_sel_match(int pass, unsigned long long hi, unsigned int lo, unsigned long long phi, unsigned int plo,
                                                 unsigned int* digit)
    {
        if (pass < 8)
        {
            const int sh = 56 - 8 * pass;
            *digit       = (unsigned int)(hi >> sh) & 255u;
            return pass == 0 || (hi >> (sh + 8)) == (phi >> (sh + 8));
        }
        const int sh = 24 - 8 * (pass - 8);
        *digit       = (lo >> sh) & 255u;
        if (hi != phi) return false;
        return pass == 8 || (lo >> (sh + 8)) == (plo >> (sh + 8));
    }

    __global__ void __launch_bounds__(256) ll_select_hist_kernel(const unsigned long long* __restrict__ keys, unsigned long long n, int pass,
                                                                 ll_select_state* st)
    {
        __shared__ unsigned int h[256];
        __shared__ bool last;
        h[threadIdx.x] = 0u;
        __syncthreads();
        const unsigned long long phi = st->prefix_hi;
        const unsigned int plo       = st->prefix_lo;
        for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x)
        {
            unsigned int d;
            if (ll_sel_match(pass, keys[i], (unsigned int)i, phi, plo, &d)) atomicAdd(&h[d], 1u);
        }
        __syncthreads();
        if (h[threadIdx.x]) atomicAdd(&st->hist[threadIdx.x], h[threadIdx.x]);
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) last = atomicAdd(&st->ticket, 1u) == gridDim.x - 1;
        __syncthreads();
        if (!last) return;
        __threadfence();
        if (threadIdx.x == 0)
        { // pick the digit in which the k-th smallest composite key falls
            unsigned int k = st->k_rem, d = 0;
            volatile unsigned int* gh = st->hist;
            for (; d < 255; ++d)
            {
                const unsigned int c = gh[d];
                if (k <= c) break;
                k -= c;
            }
            st->k_rem = k;
            if (pass < 8) st->prefix_hi |= (unsigned long long)d << (56 - 8 * pass);
            else st->prefix_lo |= d << (24 - 8 * (pass - 8));
            st->ticket = 0u;
        }
        __syncthreads();
        st->hist[threadIdx.x] = 0u;
    }

    // small populations (n <= LL_MAX_TOP): every member is a candidate, the single-block sort below does the whole selection
    // (2 launches instead of 15: the reference's own population sizes are launch-latency territory)
    __global__ void ll_all_candidates_kernel(const double* __restrict__ fit, unsigned int n, unsigned long long first_global, ll_cand* __restrict__ cand)
    {
        const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n) return;
        cand[i].fitness = __longlong_as_double((long long)ll_orderable(fit[i])); // orderable bits, decoded by the sort kernel
        cand[i].index   = first_global + i;
    }

    // after 12 passes (prefix_hi, prefix_lo) IS the k-th smallest composite key: collect everything <= it
    __global__ void ll_select_collect_kernel(const unsigned long long* __restrict__ keys, unsigned long long n, unsigned long long first_global,
                                             ll_select_state* st, ll_cand* cand, unsigned int cap)
    {
        const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n) return;
        const unsigned long long hi = keys[i], phi = st->prefix_hi;
        if (hi < phi || (hi == phi && (unsigned int)i <= st->prefix_lo))
        {
            const unsigned int slot = atomicAdd(&st->n_cand, 1u);
            if (slot < cap)
            {
                cand[slot].fitness = __longlong_as_double((long long)hi); // orderable bits for now; decoded by the sort kernel
                cand[slot].index   = first_global + i;
            }
        }
    }

    // single block bitonic sort of the (<= LL_MAX_TOP) candidates by (key, index)
    __global__ void __launch_bounds__(1024) ll_select_sort_kernel(ll_cand* cand, unsigned int k, unsigned int npow2)
    {
        extern __shared__ __align__(16) unsigned char smem_raw[];
        unsigned long long* sk = reinterpret_cast<unsigned long long*>(smem_raw);
        unsigned long long* si = sk + npow2;
        for (unsigned int i = threadIdx.x; i < npow2; i += blockDim.x)
        {
            sk[i] = i < k ? (unsigned long long)__double_as_longlong(cand[i].fitness) : ~0ull;
            si[i] = i < k ? cand[i].index : ~0ull;
        }
        __syncthreads();
        for (unsigned int size = 2; size <= npow2; size <<= 1)
            for (unsigned int stride = size >> 1; stride > 0; stride >>= 1)
            {
                for (unsigned int t = threadIdx.x; t < npow2 / 2; t += blockDim.x)
                {
                    const unsigned int lo = 2 * t - (t & (stride - 1));
                    const unsigned int hi = lo + stride;
                    const bool up         = (lo & size) == 0;
                    const bool gt         = sk[lo] > sk[hi] || (sk[lo] == sk[hi] && si[lo] > si[hi]);
                    if (gt == up)
                    {
                        const unsigned long long tk = sk[lo], ti = si[lo];
                        sk[lo] = sk[hi], si[lo] = si[hi];
                        sk[hi] = tk, si[hi] = ti;
                    }
                }
                __syncthreads();
            }
        for (unsigned int i = threadIdx.x; i < k; i += blockDim.x)
        {
            cand[i].fitness = ll_from_orderable(sk[i]);
            cand[i].index   = si[i];
        }
    }

    // ------------------------------------------------------------------------------------------------ K3
    // out[r] = row of global member idx[r] if this shard owns it; otherwise zeros (zero_unowned) or untouched.
    template <typename T>
    __global__ void ll_gather_rows_kernel(const T* __restrict__ params, const unsigned long long* __restrict__ idx, unsigned int n,
                                          unsigned long long first_global, unsigned long long pop_local, int p_dev, T* __restrict__ out,
                                          int zero_unowned)
    {
        const unsigned int r = blockIdx.y;
        if (r >= n) return;
        const unsigned long long g = idx[r];
        const bool owned           = g >= first_global && g - first_global < pop_local;
        const int d                = blockIdx.x * blockDim.x + threadIdx.x;
        if (d >= p_dev) return;
        if (owned) out[(size_t)r * p_dev + d] = params[(g - first_global) * (size_t)p_dev + d];
        else if (zero_unowned) out[(size_t)r * p_dev + d] = T(0);
    }
} // namespace

int ll_launch_create_population_gen(ll_ctx* ctx, int elitism, uint64_t gen)
{
    pop_args a;
    a.params = ctx->d_params, a.elites = ctx->d_elites, a.dev2canon = ctx->d_dev2canon;
    a.first_global = ctx->first_global, a.pop_global = ctx->pop_global, a.pop_local = ctx->pop_local;
    a.seed = (unsigned int)ctx->seed, a.top = (unsigned int)ctx->top;
    a.p_canon = ctx->topo.p_canon, a.p_dev = ctx->topo.p_dev;
    a.sigma_init = ctx->sigma_init, a.sigma_mut = ctx->sigma_mut;
    a.from_elites = ctx->has_elites ? 1 : 0;
    if (!ctx->has_elites)
    { // first population: every slot is a fresh Gaussian individual
        a.gen = (unsigned int)gen, a.first_child = 0, a.n_children = ctx->pop_global, a.n_inject = ctx->pop_global;
    }
    else
    {
        a.gen         = (unsigned int)gen;
        a.first_child = elitism ? ctx->top : 0;
        const unsigned long long n_children = ctx->pop_global - a.first_child;
        unsigned long long n_inject         = 0;
        if (ctx->strategy == LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION)
        {
            const double v = ctx->ratio * (double)n_children + 0.5;
            n_inject       = v <= 0.0 ? 0ull : (unsigned long long)v; // floor(x + 0.5) for x >= 0
        }
        if (n_inject > n_children) n_inject = n_children;
        a.n_children = n_children, a.n_inject = n_inject;
    }
    const size_t smem = ((size_t)ctx->topo.p_canon + 1) * ctx->elt;
    unsigned int grid = (unsigned int)(ctx->pop_local < (1u << 20) ? ctx->pop_local : (1u << 20));
    if (grid == 0) return LL_OK;
    if (ctx->precision == 64)
    {
        LL_CUDA(ctx, cudaFuncSetAttribute(ll_create_population_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ll_create_population_kernel<double><<<grid, 256, smem, ctx->stream>>>(a);
    }
    else
    {
        ll_pop_meta_kernel<<<(unsigned int)((ctx->pop_local + 255) / 256), 256, 0, ctx->stream>>>(a, ctx->d_breed);
        const unsigned int blocks = (unsigned int)std::min<unsigned long long>((ctx->pop_local + 7) / 8, (unsigned long long)ctx->sm_count * 8);
        if (ctx->mode == LL_MODE_FAST) ll_create_population_f32_kernel<true><<<blocks, 256, 0, ctx->stream>>>(a, ctx->d_breed);
        else ll_create_population_f32_kernel<false><<<blocks, 256, 0, ctx->stream>>>(a, ctx->d_breed);
    }
    LL_CUDA(ctx, cudaGetLastError());
    return LL_OK;
}

int ll_launch_stats(ll_ctx* ctx)
{
    const unsigned int n_part = (unsigned int)((ctx->pop_local + ST_CHUNK - 1) / ST_CHUNK);
    st_partial* part          = reinterpret_cast<st_partial*>(ctx->d_partials);
    ll_stats_partial_kernel<<<n_part, ST_THREADS, 0, ctx->stream>>>(ctx->d_fitness, ctx->d_steps, ctx->pop_local, part);
    ll_stats_final_kernel<<<1, 32, 0, ctx->stream>>>(part, n_part, ctx->d_lr, ctx->first_global, ctx->d_stats);
    LL_CUDA(ctx, cudaGetLastError());
    return LL_OK;
}

size_t ll_stats_partial_bytes(unsigned long long pop_local) { return ((pop_local + ST_CHUNK - 1) / ST_CHUNK + 1) * sizeof(st_partial); }

int ll_launch_topk(ll_ctx* ctx, unsigned int k)
{
    const unsigned long long n = ctx->pop_local;
    if (k == 0 || k > n || k > LL_MAX_TOP) return ll_fail(ctx, LL_EINVAL, "top-k: k out of range");
    const unsigned int blocks = (unsigned int)((n + 255) / 256);
    unsigned int n_sort       = k; // entries the sort kernel orders
    if (n <= LL_MAX_TOP)
    {
        ll_all_candidates_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_fitness, (unsigned int)n, ctx->first_global, ctx->d_cand);
        n_sort = (unsigned int)n;
    }
    else
    {
        ll_keys_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_fitness, n, ctx->d_keys, ctx->d_sel, k);
        unsigned int hb = blocks < (unsigned int)(ctx->sm_count * 8) ? blocks : (unsigned int)(ctx->sm_count * 8);
        for (int pass = 0; pass < 12; ++pass) ll_select_hist_kernel<<<hb, 256, 0, ctx->stream>>>(ctx->d_keys, n, pass, ctx->d_sel);
        ll_select_collect_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_keys, n, ctx->first_global, ctx->d_sel, ctx->d_cand, LL_MAX_TOP);
    }
    unsigned int np2 = 2;
    while (np2 < n_sort) np2 <<= 1;
    const size_t smem = (size_t)np2 * 16;
    LL_CUDA(ctx, cudaFuncSetAttribute(ll_select_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ll_select_sort_kernel<<<1, np2 / 2 < 1024 ? (np2 / 2 < 32 ? 32 : np2 / 2) : 1024, smem, ctx->stream>>>(ctx->d_cand, n_sort, np2);
    LL_CUDA(ctx, cudaGetLastError());
    return LL_OK;
}

int ll_launch_gather(ll_ctx* ctx, const unsigned long long* d_idx, unsigned int n, void* d_out, int zero_unowned)
{
    if (n == 0) return LL_OK;
    dim3 grid((ctx->topo.p_dev + 255) / 256, n);
    if (ctx->precision == 64)
        ll_gather_rows_kernel<double><<<grid, 256, 0, ctx->stream>>>(static_cast<const double*>(ctx->d_params), d_idx, n, ctx->first_global,
                                                                     ctx->pop_local, ctx->topo.p_dev, static_cast<double*>(d_out), zero_unowned);
    else
        ll_gather_rows_kernel<float><<<grid, 256, 0, ctx->stream>>>(static_cast<const float*>(ctx->d_params), d_idx, n, ctx->first_global,
                                                                    ctx->pop_local, ctx->topo.p_dev, static_cast<float*>(d_out), zero_unowned);
    LL_CUDA(ctx, cudaGetLastError());
    return LL_OK;
}

// ---------------------------------------------------------------------------------------------------- K2 without allocations
// ll_topk_static: the same radix select as ll_launch_topk over a caller's DEVICE fitness array, for callers that may not allocate
// (CUDA graph capture, tight serving loops).  No counterpart in the reference.
// The scratch (select state, candidate list, orderable keys) lives in ONE file-scope device array: every GPU that loads the library
// holds LL_TOPK_STATIC_SCRATCH bytes until the process exits, every call on a device shares them, and the array keeps the last
// call's contents — ll_keys_kernel rewrites the select state (prefix, k, ticket, candidate count, histogram) at the start of every
// call and the keys are overwritten for [0, n), which is all the later kernels read.  Calls that overlap on one GPU race.
namespace
{
    constexpr size_t TS_STATE_OFF = 0;                                        // ll_select_state (1048 bytes)
    constexpr size_t TS_CAND_OFF  = 2048;                                     // ll_cand[LL_MAX_TOP]
    constexpr size_t TS_KEYS_OFF  = TS_CAND_OFF + sizeof(ll_cand) * LL_MAX_TOP; // u64[LL_TOPK_STATIC_MAX]
    constexpr size_t TS_BYTES     = TS_KEYS_OFF + sizeof(unsigned long long) * (size_t)LL_TOPK_STATIC_MAX;
    static_assert(sizeof(ll_select_state) <= TS_CAND_OFF && TS_KEYS_OFF % 256 == 0 && TS_BYTES == LL_TOPK_STATIC_SCRATCH_BYTES, "scratch layout");

    static __device__ __align__(256) unsigned char ll_topk_static_scratch[TS_BYTES];

    struct ts_device
    {
        unsigned char* scratch = nullptr; // DEVICE address of ll_topk_static_scratch on this GPU (the name itself is a host-side stand-in)
        int sm_count           = 0;
    };
    constexpr int TS_MAX_DEVICES = 64;
    ts_device g_ts[TS_MAX_DEVICES];
    std::mutex g_ts_mutex;

    // once per device, with that device current
    int ts_setup(int device, ts_device* out)
    {
        std::lock_guard<std::mutex> lock(g_ts_mutex);
        ts_device& d = g_ts[device];
        if (!d.scratch)
        {
            void* p = nullptr;
            LL_CUDA(nullptr, cudaGetSymbolAddress(&p, ll_topk_static_scratch));
            int sms = 0;
            LL_CUDA(nullptr, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
            int smem_max = 0;
            LL_CUDA(nullptr, cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
            if ((size_t)smem_max < (size_t)LL_MAX_TOP * 16) return ll_fail(nullptr, LL_ENODEVICE, "ll_topk_static: too little shared memory per block");
            d.sm_count = sms, d.scratch = static_cast<unsigned char*>(p);
        }
        *out = d;
        return LL_OK;
    }

    int ts_run(int device, cudaStream_t stream, const double* d_fitness, unsigned long long n, unsigned long long first_global, unsigned int k,
               ll_cand* d_out)
    {
        ts_device dev;
        if (const int rc = ts_setup(device, &dev)) return rc;
        ll_select_state* st      = reinterpret_cast<ll_select_state*>(dev.scratch + TS_STATE_OFF);
        ll_cand* cand            = reinterpret_cast<ll_cand*>(dev.scratch + TS_CAND_OFF);
        unsigned long long* keys = reinterpret_cast<unsigned long long*>(dev.scratch + TS_KEYS_OFF);
        const unsigned int blocks = (unsigned int)((n + 255) / 256);
        ll_keys_kernel<<<blocks, 256, 0, stream>>>(d_fitness, n, keys, st, k);
        const unsigned int hb = blocks < (unsigned int)(dev.sm_count * 8) ? blocks : (unsigned int)(dev.sm_count * 8);
        for (int pass = 0; pass < 12; ++pass) ll_select_hist_kernel<<<hb, 256, 0, stream>>>(keys, n, pass, st);
        ll_select_collect_kernel<<<blocks, 256, 0, stream>>>(keys, n, first_global, st, cand, LL_MAX_TOP);
        unsigned int np2 = 2;
        while (np2 < k) np2 <<= 1;
        // per call, not at set-up: ll_launch_topk sets this attribute of the same kernel to its own (possibly smaller) size
        LL_CUDA(nullptr, cudaFuncSetAttribute(ll_select_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(np2 * 16)));
        ll_select_sort_kernel<<<1, np2 / 2 < 1024 ? (np2 / 2 < 32 ? 32 : np2 / 2) : 1024, (size_t)np2 * 16, stream>>>(cand, k, np2);
        LL_CUDA(nullptr, cudaGetLastError());
        LL_CUDA(nullptr, cudaMemcpyAsync(d_out, cand, sizeof(ll_cand) * k, cudaMemcpyDeviceToDevice, stream));
        return LL_OK;
    }
} // namespace

extern "C" int ll_topk_static(int32_t device, void* cuda_stream, const double* d_fitness, uint64_t n, uint64_t first_global, uint32_t k,
                              ll_cand* d_out)
{
    if (!d_fitness || !d_out) return ll_fail(nullptr, LL_EINVAL, "ll_topk_static: null argument");
    if (n == 0 || n > LL_TOPK_STATIC_MAX)
        return ll_fail(nullptr, LL_EINVAL,
                       "ll_topk_static: n = " + std::to_string(n) + " is outside 1.." + std::to_string((unsigned long long)LL_TOPK_STATIC_MAX) +
                           " (LL_TOPK_STATIC_MAX, the size of the static scratch array); use ll_local_topk on a context for larger shards");
    if (k == 0 || k > n || k > LL_MAX_TOP)
        return ll_fail(nullptr, LL_EINVAL, "ll_topk_static: k = " + std::to_string(k) + " is outside 1..min(n, " + std::to_string(LL_MAX_TOP) + ")");
    int n_dev = 0, prev = -1;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    {
        cudaGetLastError();
        return ll_fail(nullptr, LL_ENODEVICE, "ll_topk_static: no CUDA device");
    }
    LL_CUDA(nullptr, cudaGetDevice(&prev));
    if (device < 0) device = prev;
    if (device >= n_dev || device >= TS_MAX_DEVICES) return ll_fail(nullptr, LL_EINVAL, "ll_topk_static: no such device");
    if (device != prev) LL_CUDA(nullptr, cudaSetDevice(device));
    const int rc = ts_run(device, static_cast<cudaStream_t>(cuda_stream), d_fitness, n, first_global, k, d_out);
    if (device != prev) cudaSetDevice(prev);
    return rc;
}
