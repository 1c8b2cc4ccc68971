// Wavefront (queued, multi-kernel) variant of the render loop.  Placeholder until it lands: selecting
// PT_KERNEL_WAVEFRONT reports PT_ERR_UNSUPPORTED rather than silently running the megakernel.
#pragma once
#include <string>

#include "pt_kernels.cuh"

struct PtWavefront {
    void release() {}
    int render(cudaStream_t, const PtDeviceScene &, const PtRenderArgs &, uint32_t *, std::string &err) {
        err = "wavefront kernels not built into this library yet";
        return PT_ERR_UNSUPPORTED;
    }
};
