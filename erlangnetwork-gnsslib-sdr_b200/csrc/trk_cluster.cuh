/* trk_cluster.cuh - cluster / distributed-shared-memory primitives of the tracking kernel.
 *
 * One thread-block cluster tracks one channel (src/sdrtrk.c:15-54 per code period):
 * the control warps of cluster rank 0 replay the NCOs and run the loop filters, the worker
 * teams of the cluster's CTAs correlate whole code periods.  They talk through each other's
 * shared memory only:
 *   control -> worker : NCO segment tables + a slot header written with remote stores,
 *                       then one mbarrier arrive (release.cluster) per writing lane;
 *   worker -> control : the tap sums written with remote stores, one arrive per lane.
 * Waiting is mbarrier.try_wait (acquire.cluster), which sleeps in hardware.
 */
#ifndef SDRB200_TRK_CLUSTER_CUH
#define SDRB200_TRK_CLUSTER_CUH

#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t cl_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cl_rank()
{
    uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ uint32_t cl_size()
{
    uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r;
}
/* shared::cluster address of the same variable in CTA `rank` of the cluster */
__device__ __forceinline__ uint32_t cl_map(const void *local, uint32_t rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cl_smem_u32(local)), "r"(rank));
    return r;
}
__device__ __forceinline__ void cl_st_u32(uint32_t addr, uint32_t v)
{
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void cl_st_v4(uint32_t addr, uint4 v)
{
    asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void cl_mbar_init(void *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(cl_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void cl_fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
/* arrive on an mbarrier of another (or the own) CTA; orders this thread's earlier writes before it */
__device__ __forceinline__ void cl_mbar_arrive_remote(uint32_t cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
/* the same without the cluster-scope release: for "I am done READING your shared memory" signals,
 * where nothing this thread wrote has to become visible to the other side.  The instruction carries
 * no dependency on the loads it follows and ptxas schedules it right behind them, before their data
 * has returned: the caller passes a count (1) COMPUTED FROM the loaded data                        */
__device__ __forceinline__ void cl_mbar_arrive_remote_light(uint32_t cluster_addr, uint32_t count)
{
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_addr), "r"(count) : "memory");
}
__device__ __forceinline__ int cl_mbar_try_wait(void *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(cl_smem_u32(bar)), "r"(parity) : "memory");
    return (int)ok;
}
__device__ __forceinline__ void cl_mbar_wait(void *bar, uint32_t parity)
{
    while (!cl_mbar_try_wait(bar, parity)) {}
}
__device__ __forceinline__ void cl_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
/* flags between warps of ONE CTA (monotonic counters in shared memory) */
__device__ __forceinline__ void cl_flag_set(volatile int *f, int v)
{
    __threadfence_block();
    *f = v;
}
__device__ __forceinline__ void cl_flag_wait_ge(volatile int *f, int v)
{
    while (*f < v) { __nanosleep(20); }
    __threadfence_block();
}

#endif
