// Minimal asynchronous-copy wrappers (inline PTX, sm_100a).
//  * mbarrier + TMA bulk copy: `cp.async.bulk` (1-D, no tensor map) shows up in SASS as UBLKCP, the
//    transaction barrier as SYNCS.ARRIVE.TRANS64 / SYNCS.PHASECHK.  Used where a CTA pulls ONE long
//    contiguous run (onesweep partitions, 16 KB + 8 KB per CTA).
//  * cp.async (LDGSTS) + commit/wait groups: used where every lane gathers its own 128-byte row
//    (compositors).  A per-lane bulk copy is compiled into an ELECT-serialised loop, one UBLKCP and ~9
//    dependent instructions per row; measured on B200 that loop was 10.7 % / 4.6 % of the forward /
//    backward compositor's stall samples (profiles/r01_v13), and replacing it by eight LDGSTS per lane, all
//    lanes at once, took 7.6 % / 5.6 % off the two kernels.
#pragma once
#include <stdint.h>

namespace guava {
namespace tma {

__device__ __forceinline__ uint32_t smem_addr(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}

__device__ __forceinline__ void fence_mbar_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_addr(bar)),
        "r"(parity)
        : "memory");
}

// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_addr(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_addr(bar))
        : "memory");
}

// TMA tensor copy, sm_100 gather form: FOUR rows of a 2-D tensor, given by their row coordinates, land in
// consecutive box-sized pieces of shared memory (here: four 128-byte feature rows -> 512 bytes) with one
// instruction from one lane (SASS: UTMALDG...GATHER4); completion on the mbarrier's transaction count.  `tmap` is the
// generic address of a CUtensorMap kernel parameter (__grid_constant__); the box of the map is {row length, 1}.
__device__ __forceinline__ void tensor_gather4_g2s(void* smem_dst, const void* tmap, int col, int r0, int r1, int r2, int r3,
                                                   uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::
            "r"(smem_addr(smem_dst)), "l"(tmap), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_addr(bar))
        : "memory");
}

// 16-byte asynchronous copy global -> shared (LDGSTS), completion by commit/wait groups of the issuing thread
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

}  // namespace tma
}  // namespace guava
