// wavefront.cuh — wavefront pipeline (generate -> extend -> shade -> accumulate) with SoA ray/hit queues in HBM.
// Placeholder state until the pipeline lands; the megakernel is the only variant wired up so far.
#pragma once

#include "kernels.cuh"

struct WavefrontState
{
    void* buffers = nullptr;
    size_t capacity_paths = 0;
};

static inline void wavefront_free(WavefrontState& w)
{
    if (w.buffers) cudaFree(w.buffers);
    w.buffers = nullptr;
    w.capacity_paths = 0;
}
