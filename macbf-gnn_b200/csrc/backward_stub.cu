#include "common.cuh"
namespace macbf {
size_t cbf_backward_scratch_bytes(int, int, int) { return 256; }
int cbf_backward(const float*, const float*, const int32_t*, const float*, const float*, int, int, int, int, double,
                 double, float*, float*, void*, size_t, cudaStream_t) { set_error("cbf_backward: not built yet"); return 3; }
size_t controller_backward_scratch_bytes(int, int, int) { return 256; }
int controller_backward(const float*, const float*, const float*, const int32_t*, const float*, const float*,
                        const float*, int, int, int, int, double, double, float*, float*, float*, void*, size_t,
                        cudaStream_t) { set_error("controller_backward: not built yet"); return 3; }
}
