// L2 residency hints.  One frame touches ~100 MB of scene (wide nodes + triangle records: the data every ray needs again)
// and streams ~1 GB of per-pixel scratch (hit records, pending terms, queue entries: written once, read once) through the
// same 126 MB L2.  The scene fetches carry an evict_last policy, the scratch accesses evict_first, so the stream does not
// push the tree out.  createpolicy folds to a constant in a uniform register: the hint rides in the LDG/STG descriptor
// operand and costs no per-thread register (checked with cuobjdump -sass).
#pragma once
#include "rtb_internal.cuh"

namespace rtb {

#ifndef RTB_L2_HINTS
#define RTB_L2_HINTS 1
#endif

XF unsigned long long pol_keep() { unsigned long long p; asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
XF unsigned long long pol_stream() { unsigned long long p; asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }

// read-only scene data (ld.global.nc), kept in L2
XF float4 ld_keep(const float4* a)
{
#if RTB_L2_HINTS
    float4 v;
    asm("ld.global.nc.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(a), "l"(pol_keep()));
    return v;
#else
    return __ldg(a);
#endif
}

// per-pixel scratch written by another lane / SM earlier in the launch: L2 only (ld.cg), first to go
XF float4 ld_stream(const float4* a)
{
#if RTB_L2_HINTS
    float4 v;
    asm volatile("ld.global.cg.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(a), "l"(pol_stream()) : "memory");
    return v;
#else
    return __ldcg(a);
#endif
}

XF void st_stream(float4* a, float4 v)
{
#if RTB_L2_HINTS
    asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" :: "l"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol_stream()) : "memory");
#else
    *a = v;
#endif
}

}  // namespace rtb
