// srt_wavefront_variant.cu — one instantiation of the sorted wavefront kernel (srt_wavefront.cuh);
// compiled SRT_N_WF_VARIANTS times with -DSRT_WF_VARIANT=<n> (srt_variants.h).  The counting variant
// is the instrumented build of the same kernel: it is what srt_render_counted replays.
#ifndef SRT_WF_VARIANT
#error "compile with -DSRT_WF_VARIANT=<n>"
#endif

#include "srt_variants.h"
#ifndef SRT_WFP_THREADS
#define SRT_WFP_THREADS SRT_WF_VARIANT_PERSIST_THREADS(SRT_WF_VARIANT)
#endif
#ifndef SRT_WFP_MINBLOCKS
#define SRT_WFP_MINBLOCKS SRT_WF_VARIANT_CTAS_PER_SM(SRT_WF_VARIANT)
#endif
#include "srt_wavefront.cuh"

#define SRT_CAT2(a, b) a##b
#define SRT_CAT(a, b) SRT_CAT2(a, b)

namespace {
constexpr uint32_t kMask = SRT_WF_VARIANT_MASK(SRT_WF_VARIANT);
constexpr bool kCount = SRT_WF_VARIANT_COUNTING(SRT_WF_VARIANT);
}

cudaError_t SRT_CAT(srt_wavefront_variant_, SRT_WF_VARIANT)(const WfLaunch& L, WfVariantInfo* info) {
    auto sorted = srt::k_wf_sorted<kMask, kCount>;
    cudaError_t e;
    if (info) {  // runs before the first launch with a given shared-memory size (srt_api.cu: run_wavefront)
        cudaFuncAttributes fa;
        info->features = kMask;
        // static + dynamic shared memory (hot tables + the index queues) exceeds the 48 KB default
        e = cudaFuncSetAttribute((const void*)sorted, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
        if (e != cudaSuccess) return e;
        info->threads = SRT_WFP_THREADS;
        info->ctas_per_sm = SRT_WFP_MINBLOCKS;
        info->slot_bytes = srt::SlotQuads<kMask>::value * 16;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&info->blocks_per_sm, sorted, SRT_WFP_THREADS, L.smem_bytes);
        if (e != cudaSuccess) return e;
        e = cudaFuncGetAttributes(&fa, sorted);
        if (e != cudaSuccess) return e;
        info->registers = fa.numRegs;
        return cudaSuccess;
    }
    const srt::WfState& st = *static_cast<const srt::WfState*>(L.state);
    sorted<<<L.grid, L.threads > 0 ? L.threads : SRT_WFP_THREADS, L.smem_bytes, L.stream>>>(*L.scene, *L.args, st, L.sobol, L.out, L.counters, L.per_thread);
    return cudaGetLastError();
}
