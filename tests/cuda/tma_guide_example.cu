// tma_guide_example.cu -- the CUDA C++ Programming Guide's 2-D `int` tensor-map example ("Using TMA to transfer
// multi-dimensional arrays"), written with libcu++'s cuda::barrier / cuda::device::experimental wrappers exactly as the
// guide does, plus a host check.  Purpose: a KNOWN-GOOD tensor-map load to tell "tensor-map TMA does not work on this
// pool" apart from "tests/cuda/tma_probe.cu is wrong" (housekeeping item 10).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 tma_guide_example.cu -o tma_guide_example
// Exit code 0 and "tma guide example: OK" = tensor-map loads and stores work.
#include <cuda.h>   // CUtensorMap
#include <cuda/barrier>
#include <cudaTypedefs.h>   // PFN_cuTensorMapEncodeTiled
#include <cuda_runtime.h>

#include <cstdio>
#include <utility>
#include <vector>

using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;

constexpr uint32_t GMEM_WIDTH = 1024, GMEM_HEIGHT = 1024;
constexpr uint32_t SMEM_WIDTH = 32, SMEM_HEIGHT = 8;

__global__ void kernel(const __grid_constant__ CUtensorMap tensor_map, int x, int y) {
    // The destination shared memory buffer of a bulk tensor operation should be 128 byte aligned.
    __shared__ alignas(128) int smem_buffer[SMEM_HEIGHT][SMEM_WIDTH];
#pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier bar;
    if (threadIdx.x == 0) {
        init(&bar, blockDim.x);
        cde::fence_proxy_async_shared_cta();
    }
    __syncthreads();
    barrier::arrival_token token;
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_2d_global_to_shared(&smem_buffer, &tensor_map, x, y, bar);
        token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(smem_buffer));
    } else {
        token = bar.arrive();
    }
    bar.wait(std::move(token));
    smem_buffer[0][threadIdx.x] += threadIdx.x;
    cde::fence_proxy_async_shared_cta();
    __syncthreads();
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_2d_shared_to_global(&tensor_map, x, y, &smem_buffer);
        cde::cp_async_bulk_commit_group();
        cde::cp_async_bulk_wait_group_read<0>();
    }
    if (threadIdx.x == 0) (&bar)->~barrier();
}

#define CK(x)                                                          \
    do {                                                               \
        cudaError_t e = (x);                                           \
        if (e != cudaSuccess) {                                        \
            printf("%s: %s\n", #x, cudaGetErrorString(e));             \
            return 1;                                                  \
        }                                                              \
    } while (0)

int main() {
    std::vector<int> h((size_t)GMEM_WIDTH * GMEM_HEIGHT);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (int)i;
    int* d = nullptr;
    CK(cudaMalloc(&d, h.size() * sizeof(int)));
    CK(cudaMemcpy(d, h.data(), h.size() * sizeof(int), cudaMemcpyHostToDevice));
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    auto encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fp);
    CUtensorMap tensor_map{};
    constexpr uint32_t rank = 2;
    uint64_t size[rank] = {GMEM_WIDTH, GMEM_HEIGHT};
    uint64_t stride[rank - 1] = {GMEM_WIDTH * sizeof(int)};
    uint32_t box_size[rank] = {SMEM_WIDTH, SMEM_HEIGHT};
    uint32_t elem_stride[rank] = {1, 1};
    CUresult res = encode(&tensor_map, CU_TENSOR_MAP_DATA_TYPE_INT32, rank, d, size, stride, box_size, elem_stride,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("cuTensorMapEncodeTiled -> %d\n", (int)res);
    if (res != CUDA_SUCCESS) return 1;
    const int x = 64, y = 16;
    kernel<<<1, SMEM_WIDTH>>>(tensor_map, x, y);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<int> o(h.size());
    CK(cudaMemcpy(o.data(), d, o.size() * sizeof(int), cudaMemcpyDeviceToHost));
    int bad = 0;
    for (uint32_t j = 0; j < GMEM_HEIGHT; ++j)
        for (uint32_t i = 0; i < GMEM_WIDTH; ++i) {
            int want = h[(size_t)j * GMEM_WIDTH + i];
            if (j == (uint32_t)y && i >= (uint32_t)x && i < x + SMEM_WIDTH) want += (int)(i - x);
            if (o[(size_t)j * GMEM_WIDTH + i] != want) ++bad;
        }
    printf("tma guide example: %s (%d mismatches)\n", bad ? "WRONG DATA" : "OK", bad);
    return bad != 0;
}
