// ll_tc05.cuh — inline-PTX wrappers for the Blackwell (sm_100a) tensor path: tcgen05.mma / TMEM / mbarrier / bulk copy.
// Shared by ll_episode_tc05.cu (K1 fast on 5th-generation tensor cores) and ll_tc05_probe.cu (descriptor known-answer tests
// and issue-rate microbenchmarks).  Bit layouts follow the PTX ISA "tcgen05" chapter: shared-memory matrix descriptor
// (start address, leading / stride byte offsets, version, swizzle mode) and the kind::f16 instruction descriptor.
#ifndef LL_TC05_CUH_
#define LL_TC05_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

namespace tc05
{
    __device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

    // ---- descriptors -----------------------------------------------------------------------------------------------
    // Shared-memory matrix descriptor, SWIZZLE_NONE.  The operand is made of 8 x 16-byte "core matrices" (128 contiguous bytes):
    //   K-major  : core matrix = 8 rows (M or N index) x 8 consecutive k (f16); row r at +16 r bytes;
    //              LBO = byte distance between core matrices adjacent in K, SBO = between 8-row groups adjacent in M / N;
    //   MN-major : core matrix = 8 k x 8 consecutive rows; k at +16 k bytes;
    //              LBO = byte distance between 8-k groups, SBO = between 8-row groups.
    __device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes)
    {
        uint64_t d = 0;
        d |= (uint64_t)((saddr >> 4) & 0x3fffu);
        d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
        d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
        d |= (uint64_t)1 << 46; // descriptor version 1 (sm_100)
        return d;               // base offset 0, LBO mode 0, layout type 0 = no swizzle
    }

    // Instruction descriptor, kind::f16: A, B = f16, D = f32, dense, no negate.  a_mn / b_mn: 1 = MN-major operand.
    __host__ __device__ constexpr uint32_t idesc_f16(int M, int N, int a_mn, int b_mn)
    {
        return (1u << 4) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    }

    // ---- tcgen05.mma (issued by ONE thread for the CTA) ---------------------------------------------------------------
    __device__ __forceinline__ void mma_f16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
    {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                     :
                     : "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
                     : "memory");
    }
    // A operand from TMEM (lane = row, two consecutive k per 32-bit column), B from shared memory
    __device__ __forceinline__ void mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
    {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                     :
                     : "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
                     : "memory");
    }
    // arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed (implies fence::before_thread_sync)
    __device__ __forceinline__ void commit(uint32_t mbar_saddr)
    {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" : : "r"(mbar_saddr) : "memory");
    }
    __device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
    __device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
    // generic-proxy writes to shared memory (st.shared) -> visible to the async proxy (tcgen05.mma operand reads, TMA)
    __device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

    // ---- TMEM ----------------------------------------------------------------------------------------------------------
    // one full warp allocates `cols` (power of two >= 32) columns; the base address lands in shared memory
    __device__ __forceinline__ void tmem_alloc(uint32_t dst_saddr, uint32_t cols)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" : : "r"(dst_saddr), "r"(cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    __device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols)
    {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" : : "r"(taddr), "r"(cols) : "memory");
    }
    __device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
    __device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

    // tcgen05.ld.32x32b: thread i of the warp receives `N` consecutive 32-bit columns of TMEM lane (lane field of taddr) + i.
    // A warp can only touch the 32 lanes of its own sub-partition: lane field = 32 * (warp index in the CTA % 4).
    __device__ __forceinline__ void ld_x4(uint32_t taddr, uint32_t (&v)[4])
    {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr));
    }
    __device__ __forceinline__ void ld_x8(uint32_t taddr, uint32_t (&v)[8])
    {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                     : "r"(taddr));
    }
    __device__ __forceinline__ void ld_x16(uint32_t taddr, uint32_t (&v)[16])
    {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                       "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                     : "r"(taddr));
    }
    __device__ __forceinline__ void ld_x32(uint32_t taddr, uint32_t (&v)[32])
    {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                     "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                       "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
                       "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
                       "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(taddr));
    }
    __device__ __forceinline__ void st_x4(uint32_t taddr, const uint32_t (&v)[4])
    {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" : : "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]) : "memory");
    }

    __device__ __forceinline__ void st_x8(uint32_t taddr, const uint32_t (&v)[8])
    {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                     :
                     : "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                     : "memory");
    }

    __device__ __forceinline__ void st_x16(uint32_t taddr, const uint32_t (&v)[16])
    {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                     :
                     : "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
                       "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
                     : "memory");
    }

    // ---- mbarrier ------------------------------------------------------------------------------------------------------
    __device__ __forceinline__ void mbar_init(uint32_t saddr, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" : : "r"(saddr), "r"(count) : "memory"); }
    __device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __device__ __forceinline__ void mbar_arrive(uint32_t saddr)
    {
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" : : "r"(saddr) : "memory");
    }
    __device__ __forceinline__ void mbar_expect_tx(uint32_t saddr, uint32_t bytes)
    {
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" : : "r"(saddr), "r"(bytes) : "memory");
    }
    __device__ __forceinline__ uint32_t mbar_try_wait(uint32_t saddr, uint32_t parity)
    {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(saddr), "r"(parity)
                     : "memory");
        return ok;
    }
    // Bounded wait: a tensor-pipe or descriptor mistake must end the kernel with an error flag, never hang the GPU.
    // Returns false on time-out (~ 2^26 polls, seconds).
    __device__ __forceinline__ bool mbar_wait(uint32_t saddr, uint32_t parity)
    {
#pragma unroll 1
        for (uint32_t spin = 0; spin < (1u << 26); ++spin)
            if (mbar_try_wait(saddr, parity)) return true;
        return false;
    }

    // ---- TMA bulk copy global -> shared, completion on an mbarrier (SASS UBLKCP) -----------------------------------------
    __device__ __forceinline__ void bulk_g2s(uint32_t dst_saddr, const void* src, uint32_t bytes, uint32_t mbar_saddr)
    {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :
                     : "r"(dst_saddr), "l"(src), "r"(bytes), "r"(mbar_saddr)
                     : "memory");
    }

    // warp-uniform election of one lane (the MMA issue loop stays non-divergent: descriptors live in uniform registers)
    __device__ __forceinline__ bool elect_one()
    {
        uint32_t pred;
        asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
        return pred != 0;
    }

    __device__ __forceinline__ void named_bar_sync(int id, int threads) { asm volatile("bar.sync %0, %1;" : : "r"(id), "r"(threads) : "memory"); }
} // namespace tc05

#endif
