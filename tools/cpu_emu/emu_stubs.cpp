// TEST INFRASTRUCTURE ONLY: what api.cu links against besides the emulated kernel files (README.md).  The SW rerank kernels
// (s16x2 video intrinsics) and the GPU builder are not emulated: their launchers refuse.  launch_classify is restated on the host
// because every reference handle needs it (base classes 0..3 = A,C,G,T, 4 = N, 7 = any other byte; csrc/sw_kernel.cu).
#include "common.h"

namespace drm {

static int na(const char *what) { return fail(DRM_EUNSUPPORTED, std::string(what) + " is not available in the host-thread emulation"); }

int launch_classify(const uint8_t *raw, uint8_t *cls, int64_t n, int *has_other_dev, cudaStream_t)
{
    for (int64_t i = 0; i < n; ++i) {
        uint8_t c;
        switch (raw[i]) {
        case 'A': c = 0; break;
        case 'C': c = 1; break;
        case 'G': c = 2; break;
        case 'T': c = 3; break;
        case 'N': c = 4; break;
        default: c = 7; *has_other_dev = 1; break;
        }
        cls[i] = c;
    }
    return DRM_OK;
}

int launch_rerank(const DeviceRef &, const RerankArgs &, cudaStream_t) { return na("the SW rerank kernel"); }
int launch_sw_pairs(const char *, const int64_t *, const char *, const int64_t *, int64_t, int, int32_t *, cudaStream_t) { return na("the SW pair kernel"); }

}  // namespace drm

// C-ABI entry points that live in build_gpu.cu / sw_kernel.cu
extern "C" {
int drm_sw_pairs_device_noalloc(int, const char *, const int64_t *, const char *, const int64_t *, int64_t, int, int32_t *, void *) { return drm::na("the SW pair kernel"); }
int drm_build_index_gpu(const float *, int64_t, int, const float *, int64_t, int, int, int, int, int, uint32_t, const char *, double *)
{
    return drm::fail(DRM_EUNSUPPORTED, "the GPU index builder is not available in the host-thread emulation");
}
int drm_build_index_gpu_device(const float *, int64_t, int, int64_t, int, int, int, int, int, uint32_t, const char *, double *)
{
    return drm::fail(DRM_EUNSUPPORTED, "the GPU index builder is not available in the host-thread emulation");
}
}

// C-ABI entry points of shards.cu (NCCL, peer memory): not emulated
extern "C" {
static int no_shards() { return drm::fail(DRM_EUNSUPPORTED, "the position-sharded data plane (NCCL / peer memory) is not available in the host-thread emulation"); }
int drm_shards_open(const char *const *, const int64_t *, const int *, int, drm_shards **) { return no_shards(); }
int drm_shards_unique_id(void *) { return no_shards(); }
int drm_shards_open_rank(const char *, int64_t, int, int, int, const void *, drm_shards **) { return no_shards(); }
void drm_shards_close(drm_shards *) {}
int drm_shards_info(const drm_shards *, int64_t *) { return no_shards(); }
int drm_shards_search_rerank(drm_shards *, drm_ref *, int, const float *, int64_t, int, int, int, const char *, const int64_t *, int64_t, int64_t, int64_t, int,
                             int64_t, float *, int64_t *, int32_t *, uint64_t *, drm_shard_stats *) { return no_shards(); }
int drm_shards_step_device(drm_shards *, int, const float *, int64_t, int, int, int, int, void *, const float **, const int64_t **, int64_t *, int64_t *, void **)
{
    return no_shards();
}
int drm_shards_stats(drm_shards *, int, int, drm_shard_stats *) { return no_shards(); }
}
