// srt_render_variant.cu — one instantiation of k_render; compiled SRT_N_VARIANTS times with
// -DSRT_VARIANT=<n> (see srt_variants.h and __graft_entry__.build_backend).
#ifndef SRT_VARIANT
#error "compile with -DSRT_VARIANT=<n>"
#endif

#include "srt_variants.h"
#ifndef SRT_MINBLOCKS
#define SRT_MINBLOCKS SRT_VARIANT_MINBLOCKS(SRT_VARIANT)
#endif
#include "srt_kernels.cuh"

#define SRT_CAT2(a, b) a##b
#define SRT_CAT(a, b) SRT_CAT2(a, b)

namespace {
constexpr uint32_t kMask = SRT_VARIANT_MASK(SRT_VARIANT);
constexpr bool kCount = SRT_VARIANT_COUNTING(SRT_VARIANT);
}  // namespace

cudaError_t SRT_CAT(srt_render_variant_, SRT_VARIANT)(const RenderLaunch& L, RenderVariantInfo* info) {
    auto kernel = srt::k_render<kMask, kCount>;
    cudaError_t e;
    if (L.smem_bytes > 48 * 1024) {
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
        if (e != cudaSuccess) return e;
    }
    if (info) {
        info->features = kMask;
        info->counting = kCount ? 1 : 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&info->blocks_per_sm, kernel, srt::kRenderThreads, L.smem_bytes);
        if (e != cudaSuccess) return e;
        cudaFuncAttributes fa;
        e = cudaFuncGetAttributes(&fa, kernel);
        if (e != cudaSuccess) return e;
        info->registers = fa.numRegs;
    }
    if (L.grid <= 0) return cudaSuccess;
    kernel<<<L.grid, srt::kRenderThreads, L.smem_bytes, L.stream>>>(*L.scene, *L.args, L.sobol, L.out, L.counters);
    return cudaGetLastError();
}
