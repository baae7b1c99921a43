// srt_wavefront_variant.cu — one instantiation of the wavefront kernels (srt_wavefront.cuh);
// compiled SRT_N_WF_VARIANTS times with -DSRT_WF_VARIANT=<n> (srt_variants.h).  The counting variant
// instantiates the instrumented k_wf_sorted only (stage 3): it is what srt_render_counted replays.
#ifndef SRT_WF_VARIANT
#error "compile with -DSRT_WF_VARIANT=<n>"
#endif

#include "srt_variants.h"
#ifndef SRT_WF_TRACE_MINBLOCKS
#define SRT_WF_TRACE_MINBLOCKS SRT_WF_VARIANT_TRACE_MINBLOCKS(SRT_WF_VARIANT)
#endif
#ifndef SRT_WF_SHADE_MINBLOCKS
#define SRT_WF_SHADE_MINBLOCKS SRT_WF_VARIANT_SHADE_MINBLOCKS(SRT_WF_VARIANT)
#endif
#ifndef SRT_WFP_THREADS
#define SRT_WFP_THREADS SRT_WF_VARIANT_PERSIST_THREADS(SRT_WF_VARIANT)
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
    cudaFuncAttributes fa;
#if !SRT_WF_VARIANT_COUNTING(SRT_WF_VARIANT)
    auto trace = srt::k_wf_trace<kMask>;
    auto shade = srt::k_wf_shade<kMask>;
    auto persist = srt::k_wf_persist<kMask>;
#endif
    if (info) {  // runs before the first launch with a given shared-memory size (srt_api.cu: run_wavefront)
        info->features = kMask;
        // static (k_wf_sorted's queues) + dynamic shared memory may exceed the 48 KB default
        e = cudaFuncSetAttribute((const void*)sorted, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
        if (e != cudaSuccess) return e;
#if !SRT_WF_VARIANT_COUNTING(SRT_WF_VARIANT)
        const void* fns[3] = {(const void*)trace, (const void*)shade, (const void*)persist};
        for (int i = 0; i < 3; i++) {
            e = cudaFuncSetAttribute(fns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
            if (e != cudaSuccess) return e;
        }
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&info->trace_blocks_per_sm, trace, srt::kWfThreads, L.smem_bytes);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&info->shade_blocks_per_sm, shade, srt::kWfThreads, L.smem_bytes);
        if (e != cudaSuccess) return e;
        e = cudaFuncGetAttributes(&fa, trace);
        if (e != cudaSuccess) return e;
        info->trace_registers = fa.numRegs;
        e = cudaFuncGetAttributes(&fa, shade);
        if (e != cudaSuccess) return e;
        info->shade_registers = fa.numRegs;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&info->persist_blocks_per_sm, persist, SRT_WFP_THREADS,
                                                          L.smem_bytes);
        if (e != cudaSuccess) return e;
        e = cudaFuncGetAttributes(&fa, persist);
        if (e != cudaSuccess) return e;
        info->persist_registers = fa.numRegs;
#else
        info->trace_blocks_per_sm = info->shade_blocks_per_sm = 1;
        info->trace_registers = info->shade_registers = info->persist_registers = 0;
        info->persist_blocks_per_sm = 1;
#endif
        info->persist_threads = SRT_WFP_THREADS;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&info->sorted_blocks_per_sm, sorted, SRT_WFP_THREADS, L.smem_bytes);
        if (e != cudaSuccess) return e;
        e = cudaFuncGetAttributes(&fa, sorted);
        if (e != cudaSuccess) return e;
        info->sorted_registers = fa.numRegs;
        return cudaSuccess;
    }
    const srt::WfState& st = *static_cast<const srt::WfState*>(L.state);
    if (L.stage == 3) {
        sorted<<<L.grid, SRT_WFP_THREADS, L.smem_bytes, L.stream>>>(*L.scene, *L.args, st, L.sobol, L.out, L.counters,
                                                                    L.per_thread);
        return cudaGetLastError();
    }
#if !SRT_WF_VARIANT_COUNTING(SRT_WF_VARIANT)
    if (L.stage == 2)
        persist<<<L.grid, SRT_WFP_THREADS, L.smem_bytes, L.stream>>>(*L.scene, *L.args, st, L.sobol, L.out, L.counters,
                                                                     L.per_thread);
    else if (L.stage == 0)
        trace<<<L.grid, srt::kWfThreads, L.smem_bytes, L.stream>>>(*L.scene, *L.args, st, L.live, L.pass);
    else
        shade<<<L.grid, srt::kWfThreads, L.smem_bytes, L.stream>>>(*L.scene, *L.args, st, L.sobol, L.out, L.counters, L.live,
                                                                   L.pass);
    return cudaGetLastError();
#else
    return cudaErrorInvalidValue;  // the counting build exists for the sorted scheduler only
#endif
}
