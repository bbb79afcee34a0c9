// Definitions of the opaque handles of the hierarchy API, shared by hierarchy.cu and solver.cu.
#pragma once

#include <memory>
#include <vector>

#include "stages.h"

using namespace gmg;

struct gmg_input {
    gmg_ctx* ctx = nullptr;
    int32_t n = 0;
    int64_t nnz = 0;
    DevBuf<double> pts;
    DevBuf<int32_t> colptr, rowidx;
    DevBuf<double> a_val;
    bool has_a = false;
    // gmg_input_set_values_async: the values are still on their way on the copy stream until this event has passed
    cudaEvent_t a_ready = nullptr;
    ~gmg_input() { if (a_ready) cudaEventDestroy(a_ready); }
};
// Whoever reads in->a_val on the context's stream calls this first.
inline void await_values(gmg_ctx* c, const gmg_input* in) {
    if (in && in->a_ready) GMG_CUDA(cudaStreamWaitEvent(c->stream, in->a_ready, 0));
}

struct Level {
    gmg_level_info info{};
    DevBuf<int32_t> samples, parents;
    DeviceCsc coarse;
    DevBuf<double> coarse_pts;
    Triangles tri;
    DeviceCsr P;
    DeviceCsc Ac;
};

struct gmg_hierarchy {
    gmg_ctx* ctx = nullptr;
    DevBuf<double> fine_val;  // level-0 toEdgeDistanceMatrix values
    int64_t fine_nnz = 0;
    std::vector<std::unique_ptr<Level>> levels;
};

