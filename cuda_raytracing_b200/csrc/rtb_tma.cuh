// rtb_tma.cuh -- top tree levels staged into shared memory by a TMA bulk copy (cp.async.bulk + mbarrier).
//
// The builder emits wide nodes level by level, so the first N nodes of a node array ARE its top levels.  A
// persistent CTA copies the TLAS prefix and the prefix of one BLAS (the mesh most instances use) into shared
// memory once, at kernel start: one elected thread arms an mbarrier with the byte count and issues one 1-D
// bulk copy per array (SASS UBLKCP), every thread then waits on the barrier's phase.  Traversal reads a node
// from shared memory when its index is below the staged count (rtb_trace_state.cuh, SMEM variants).
#pragma once
#include "rtb_internal.cuh"

namespace rtb {

struct TopLevels {                  // CTA-uniform view of what was staged
    const float4* s_tlas;           // shared-memory copies (generic addresses)
    const float4* s_blas;
    const float4* g_blas;           // the global node array s_blas mirrors
    int n_tlas, n_blas;             // staged node counts
};

XF unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// smem: dynamic shared memory, 128-byte aligned, (n_tlas + n_blas) * NODE_F4 * 16 + 16 bytes; called by every thread of the CTA
XF void stage_top_levels(const SceneDev& sc, float4* smem, TopLevels& T)
{
    T.n_tlas = sc.top_tlas_n; T.n_blas = sc.top_blas_n; T.g_blas = sc.top_blas;
    T.s_tlas = smem; T.s_blas = smem + (size_t)T.n_tlas * NODE_F4;
    const unsigned bytes_t = (unsigned)T.n_tlas * (NODE_F4 * 16u), bytes_b = (unsigned)T.n_blas * (NODE_F4 * 16u);
    if (bytes_t + bytes_b == 0u) return;
    unsigned long long* mbar = (unsigned long long*)(smem + (size_t)(T.n_tlas + T.n_blas) * NODE_F4);
    const unsigned mb = smem_u32(mbar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mb));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mb), "r"(bytes_t + bytes_b) : "memory");
        if (bytes_t)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         :: "r"(smem_u32(T.s_tlas)), "l"(sc.tlas_nodes), "r"(bytes_t), "r"(mb) : "memory");
        if (bytes_b)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         :: "r"(smem_u32(T.s_blas)), "l"(sc.top_blas), "r"(bytes_b), "r"(mb) : "memory");
    }
    unsigned done = 0;
    while (!done) {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(mb) : "memory");
    }
}

}  // namespace rtb
