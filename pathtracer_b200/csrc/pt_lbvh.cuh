// Device LBVH build (Morton codes -> radix sort -> Karras hierarchy -> refit).  Placeholder until the
// device builder lands: selecting PT_BVH_DEVICE_LBVH reports PT_ERR_UNSUPPORTED rather than silently
// using the host builder.
#pragma once
#include <string>
#include <vector>

#include "pt_api.h"
#include "pt_bvh.h"

static inline int pt_lbvh_build(cudaStream_t, const std::vector<ptbvh::Box> &, int, ptbvh::Result &, std::string &err) {
    err = "device LBVH builder not built into this library yet";
    return PT_ERR_UNSUPPORTED;
}
