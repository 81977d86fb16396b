// Host-build stub (test tooling): g++ cannot digest the reference's CUDA-11-only cuda_graph.h
// (SURVEY.md fact 8).  Only current_capture()/schedule_synchronize() are referenced by gpu_memory.h.
#pragma once
#include <tiny-cuda-nn/common.h>
namespace tcnn {
struct CudaGraph { void schedule_synchronize() {} };
inline CudaGraph*& current_capture() { static CudaGraph* p = nullptr; return p; }
}
