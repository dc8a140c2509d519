// lat_probe.cu -- what the fixed costs of a short host call are on this box: launch, dependent launches,
// synchronisation, small copies, mapped-pinned reads and writes, a completion flag polled by the host,
// programmatic dependent launch, graph launch.  The numbers decide how the single-call paths of libslamgpu are
// put together.   nvcc -O2 -gencode arch=compute_100a,code=sm_100a tools/lat_probe.cu -o tools/lat_probe
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#include <cuda_runtime.h>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            std::fprintf(stderr, "%s: %s (line %d)\n", #x, cudaGetErrorString(e_), __LINE__); \
            std::exit(1);                                                                       \
        }                                                                                       \
    } while (0)

__global__ void null_kernel(int* p)
{
    if (p && threadIdx.x == 0 && blockIdx.x == 0x7fffffff) *p = 1;
}

__global__ void null_kernel_pdl(int* p)
{
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
    cudaGridDependencySynchronize();
#endif
    if (p && threadIdx.x == 0 && blockIdx.x == 0x7fffffff) *p = 1;
}

__global__ void flag_kernel(volatile unsigned int* host_flag, unsigned int value)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        __threadfence_system();
        *host_flag = value;
    }
}

// many CTAs; the last one to finish raises the flag
__global__ void flag_last_kernel(unsigned int* counter, volatile unsigned int* host_flag, unsigned int value, int* sink)
{
    if (sink && blockIdx.x == 0x7fffffff) *sink = 1;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int done = atomicAdd(counter, 1u);
        if (done == gridDim.x - 1) {
            *counter = 0;
            __threadfence_system();
            *host_flag = value;
        }
    }
}

__global__ void pull_kernel(const uint4* __restrict__ host_src, uint4* dst, int n)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = host_src[i];
}

__global__ void push_kernel(const uint4* __restrict__ src, uint4* host_dst, int n)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) host_dst[i] = src[i];
}

__global__ void push32_kernel(const uint32_t* __restrict__ src, uint32_t* host_dst, int n)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) host_dst[i] = src[i];
}

static double med(const std::function<void()>& fn, int reps = 400)
{
    for (int i = 0; i < 30; i++) fn();
    std::vector<double> t(reps);
    for (int i = 0; i < reps; i++) {
        const auto a = std::chrono::steady_clock::now();
        fn();
        const auto b = std::chrono::steady_clock::now();
        t[i] = std::chrono::duration<double, std::micro>(b - a).count();
    }
    std::sort(t.begin(), t.end());
    return t[reps / 2];
}

int main()
{
    CK(cudaSetDevice(0));
    cudaStream_t s;
    CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    int* d_int;
    CK(cudaMalloc(&d_int, 4));
    const size_t kb128 = 128 << 10;
    uint8_t *h_pin, *d_buf, *h_page = static_cast<uint8_t*>(malloc(kb128));
    CK(cudaHostAlloc(&h_pin, 1 << 20, cudaHostAllocMapped));
    CK(cudaMalloc(&d_buf, 1 << 20));
    volatile unsigned int* h_flag = reinterpret_cast<volatile unsigned int*>(h_pin + (512 << 10));
    unsigned int* d_counter;
    CK(cudaMalloc(&d_counter, 4));
    CK(cudaMemset(d_counter, 0, 4));
    unsigned int gen = 0;

    std::printf("{");
    std::printf("\"launch_1_sync\": %.2f", med([&] { null_kernel<<<1, 32, 0, s>>>(d_int); CK(cudaStreamSynchronize(s)); }));
    std::printf(", \"launch_148x256_sync\": %.2f", med([&] { null_kernel<<<148, 256, 0, s>>>(d_int); CK(cudaStreamSynchronize(s)); }));
    for (int k : {2, 3, 4, 6}) {
        std::printf(", \"launch_%d_dependent_sync\": %.2f", k, med([&] {
                        for (int i = 0; i < k; i++) null_kernel<<<148, 256, 0, s>>>(d_int);
                        CK(cudaStreamSynchronize(s));
                    }));
    }
    std::printf(", \"launch_only_no_sync_x1\": %.2f", med([&] { null_kernel<<<148, 256, 0, s>>>(d_int); }));
    CK(cudaStreamSynchronize(s));
    std::printf(", \"memset_then_launch_sync\": %.2f", med([&] {
                    CK(cudaMemsetAsync(d_int, 0, 4, s));
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"h2d_128k_pinned_then_launch_sync\": %.2f", med([&] {
                    CK(cudaMemcpyAsync(d_buf, h_pin, kb128, cudaMemcpyHostToDevice, s));
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"h2d_16k_pinned_then_launch_sync\": %.2f", med([&] {
                    CK(cudaMemcpyAsync(d_buf, h_pin, 16 << 10, cudaMemcpyHostToDevice, s));
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"host_memcpy_128k_then_h2d_then_launch_sync\": %.2f", med([&] {
                    memcpy(h_pin, h_page, kb128);
                    CK(cudaMemcpyAsync(d_buf, h_pin, kb128, cudaMemcpyHostToDevice, s));
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"pull_kernel_128k_then_launch_sync\": %.2f", med([&] {
                    pull_kernel<<<32, 256, 0, s>>>(reinterpret_cast<const uint4*>(h_pin), reinterpret_cast<uint4*>(d_buf), int(kb128 / 16));
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"pull_kernel_16k_then_launch_sync\": %.2f", med([&] {
                    pull_kernel<<<4, 256, 0, s>>>(reinterpret_cast<const uint4*>(h_pin), reinterpret_cast<uint4*>(d_buf), (16 << 10) / 16);
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"launch_then_d2h_16k_sync\": %.2f", med([&] {
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaMemcpyAsync(h_pin, d_buf, 16 << 10, cudaMemcpyDeviceToHost, s));
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"push_kernel_16k_sync\": %.2f", med([&] {
                    push_kernel<<<4, 256, 0, s>>>(reinterpret_cast<const uint4*>(d_buf), reinterpret_cast<uint4*>(h_pin), (16 << 10) / 16);
                    CK(cudaStreamSynchronize(s));
                }));
    // completion flag in mapped pinned memory, polled by the host
    std::printf(", \"launch_1_flag_poll\": %.2f", med([&] {
                    ++gen;
                    flag_kernel<<<1, 32, 0, s>>>(h_flag, gen);
                    while (*h_flag != gen) {}
                }));
    std::printf(", \"launch_148_last_cta_flag_poll\": %.2f", med([&] {
                    ++gen;
                    flag_last_kernel<<<148, 256, 0, s>>>(d_counter, h_flag, gen, d_int);
                    while (*h_flag != gen) {}
                }));
    std::printf(", \"launch_3_dependent_last_flag_poll\": %.2f", med([&] {
                    ++gen;
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    flag_last_kernel<<<148, 256, 0, s>>>(d_counter, h_flag, gen, d_int);
                    while (*h_flag != gen) {}
                }));
    CK(cudaStreamSynchronize(s));
    // programmatic dependent launch: the next kernel's launch overlaps the previous one's tail
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(148);
        cfg.blockDim = dim3(256);
        cfg.stream = s;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        for (int k : {3, 6}) {
            std::printf(", \"launch_%d_pdl_sync\": %.2f", k, med([&] {
                            for (int i = 0; i < k; i++) CK(cudaLaunchKernelEx(&cfg, null_kernel_pdl, d_int));
                            CK(cudaStreamSynchronize(s));
                        }));
        }
    }
    // a graph of 3 dependent kernels
    {
        cudaGraph_t g;
        cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < 3; i++) null_kernel<<<148, 256, 0, s>>>(d_int);
        CK(cudaStreamEndCapture(s, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        std::printf(", \"graph_3_kernels_sync\": %.2f", med([&] { CK(cudaGraphLaunch(ge, s)); CK(cudaStreamSynchronize(s)); }));
    }
    // graphs: size, per-call parameter updates, re-capture + whole-graph update
    {
        for (int k : {1, 5}) {
            cudaGraph_t g;
            cudaGraphExec_t ge;
            CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
            for (int i = 0; i < k; i++) null_kernel<<<148, 256, 0, s>>>(d_int);
            CK(cudaStreamEndCapture(s, &g));
            CK(cudaGraphInstantiate(&ge, g, 0));
            std::printf(", \"graph_%d_kernels_sync\": %.2f", k, med([&] { CK(cudaGraphLaunch(ge, s)); CK(cudaStreamSynchronize(s)); }));
        }
        // explicit nodes so that their parameters can be set per call
        cudaGraph_t g;
        CK(cudaGraphCreate(&g, 0));
        cudaGraphNode_t nodes[3];
        int* arg_ptr = d_int;
        void* args[1] = {&arg_ptr};
        cudaKernelNodeParams kp = {};
        kp.func = reinterpret_cast<void*>(null_kernel);
        kp.gridDim = dim3(148);
        kp.blockDim = dim3(256);
        kp.kernelParams = args;
        for (int i = 0; i < 3; i++) CK(cudaGraphAddKernelNode(&nodes[i], g, i ? &nodes[i - 1] : nullptr, i ? 1 : 0, &kp));
        cudaGraphExec_t ge;
        CK(cudaGraphInstantiate(&ge, g, 0));
        int flip = 0;
        std::printf(", \"graph_3_setparams_each_sync\": %.2f", med([&] {
                        kp.gridDim = dim3(148 + (flip ^= 1));
                        for (int i = 0; i < 3; i++) CK(cudaGraphExecKernelNodeSetParams(ge, nodes[i], &kp));
                        CK(cudaGraphLaunch(ge, s));
                        CK(cudaStreamSynchronize(s));
                    }));
        std::printf(", \"graph_setparams_one_node_cpu_only\": %.2f", med([&] {
                        kp.gridDim = dim3(148 + (flip ^= 1));
                        CK(cudaGraphExecKernelNodeSetParams(ge, nodes[0], &kp));
                    }, 2000));
        cudaEvent_t ev;
        CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        std::printf(", \"graph_3_setparams_event_sync\": %.2f", med([&] {
                        kp.gridDim = dim3(148 + (flip ^= 1));
                        for (int i = 0; i < 3; i++) CK(cudaGraphExecKernelNodeSetParams(ge, nodes[i], &kp));
                        CK(cudaGraphLaunch(ge, s));
                        CK(cudaEventRecord(ev, s));
                        CK(cudaEventSynchronize(ev));
                    }));
        std::printf(", \"graph_3_recapture_execupdate_sync\": %.2f", med([&] {
                        cudaGraph_t g2;
                        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
                        for (int i = 0; i < 3; i++) null_kernel<<<148 + flip, 256, 0, s>>>(d_int);
                        CK(cudaStreamEndCapture(s, &g2));
                        cudaGraphExecUpdateResultInfo info;
                        CK(cudaGraphExecUpdate(ge, g2, &info));
                        CK(cudaGraphLaunch(ge, s));
                        CK(cudaStreamSynchronize(s));
                        CK(cudaGraphDestroy(g2));
                        flip ^= 1;
                    }));
        // a copy node in front: 128 KB from pinned memory, then two kernels
        cudaGraph_t g3;
        cudaGraphExec_t ge3;
        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        CK(cudaMemcpyAsync(d_buf, h_pin, kb128, cudaMemcpyHostToDevice, s));
        null_kernel<<<148, 256, 0, s>>>(d_int);
        null_kernel<<<148, 256, 0, s>>>(d_int);
        CK(cudaStreamEndCapture(s, &g3));
        CK(cudaGraphInstantiate(&ge3, g3, 0));
        std::printf(", \"graph_h2d_128k_2_kernels_sync\": %.2f", med([&] { CK(cudaGraphLaunch(ge3, s)); CK(cudaStreamSynchronize(s)); }));
        cudaGraph_t g4;
        cudaGraphExec_t ge4;
        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        pull_kernel<<<32, 256, 0, s>>>(reinterpret_cast<const uint4*>(h_pin), reinterpret_cast<uint4*>(d_buf), int(kb128 / 16));
        null_kernel<<<148, 256, 0, s>>>(d_int);
        null_kernel<<<148, 256, 0, s>>>(d_int);
        CK(cudaStreamEndCapture(s, &g4));
        CK(cudaGraphInstantiate(&ge4, g4, 0));
        std::printf(", \"graph_pull_128k_2_kernels_sync\": %.2f", med([&] { CK(cudaGraphLaunch(ge4, s)); CK(cudaStreamSynchronize(s)); }));
    }
    std::printf(", \"h2d_2x64k_pageable_then_launch_sync\": %.2f", med([&] {
                    CK(cudaMemcpyAsync(d_buf, h_page, 64 << 10, cudaMemcpyHostToDevice, s));
                    CK(cudaMemcpyAsync(d_buf + (64 << 10), h_page + (64 << 10), 64 << 10, cudaMemcpyHostToDevice, s));
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"h2d_128k_pageable_then_launch_sync\": %.2f", med([&] {
                    CK(cudaMemcpyAsync(d_buf, h_page, kb128, cudaMemcpyHostToDevice, s));
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"pull_kernel_512k_then_launch_sync\": %.2f", med([&] {
                    pull_kernel<<<128, 256, 0, s>>>(reinterpret_cast<const uint4*>(h_pin), reinterpret_cast<uint4*>(d_buf), (512 << 10) / 16);
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    std::printf(", \"h2d_512k_pinned_then_launch_sync\": %.2f", med([&] {
                    CK(cudaMemcpyAsync(d_buf, h_pin, 512 << 10, cudaMemcpyHostToDevice, s));
                    null_kernel<<<148, 256, 0, s>>>(d_int);
                    CK(cudaStreamSynchronize(s));
                }));
    // pageable source -> device, the ways a host-buffer call can do it
    {
        uint8_t* h_wc;
        CK(cudaHostAlloc(&h_wc, 1 << 20, cudaHostAllocMapped | cudaHostAllocWriteCombined));
        std::printf(", \"stage_cached_pinned_then_pull_128k_launch_sync\": %.2f", med([&] {
                        memcpy(h_pin, h_page, kb128);
                        pull_kernel<<<32, 256, 0, s>>>(reinterpret_cast<const uint4*>(h_pin), reinterpret_cast<uint4*>(d_buf), int(kb128 / 16));
                        null_kernel<<<148, 256, 0, s>>>(d_int);
                        CK(cudaStreamSynchronize(s));
                    }));
        std::printf(", \"stage_wc_pinned_then_h2d_128k_launch_sync\": %.2f", med([&] {
                        memcpy(h_wc, h_page, kb128);
                        CK(cudaMemcpyAsync(d_buf, h_wc, kb128, cudaMemcpyHostToDevice, s));
                        null_kernel<<<148, 256, 0, s>>>(d_int);
                        CK(cudaStreamSynchronize(s));
                    }));
        std::printf(", \"stage_wc_pinned_then_pull_128k_launch_sync\": %.2f", med([&] {
                        memcpy(h_wc, h_page, kb128);
                        pull_kernel<<<32, 256, 0, s>>>(reinterpret_cast<const uint4*>(h_wc), reinterpret_cast<uint4*>(d_buf), int(kb128 / 16));
                        null_kernel<<<148, 256, 0, s>>>(d_int);
                        CK(cudaStreamSynchronize(s));
                    }));
        std::printf(", \"stage_wc_memcpy_128k_cpu_only\": %.2f", med([&] { memcpy(h_wc, h_page, kb128); }));
        std::printf(", \"stage_cached_memcpy_128k_cpu_only\": %.2f", med([&] { memcpy(h_pin, h_page, kb128); }));
        // the kernel that needs the data pulls it itself: launch first, stage while the launch is in flight?  (not
        // possible: the data must be staged before the kernel reads it) -- instead: stage half, launch pull of half, ...
        std::printf(", \"stage_halves_overlapped_pull_128k_launch_sync\": %.2f", med([&] {
                        memcpy(h_pin, h_page, kb128 / 2);
                        pull_kernel<<<16, 256, 0, s>>>(reinterpret_cast<const uint4*>(h_pin), reinterpret_cast<uint4*>(d_buf), int(kb128 / 32));
                        memcpy(h_pin + kb128 / 2, h_page + kb128 / 2, kb128 / 2);
                        pull_kernel<<<16, 256, 0, s>>>(reinterpret_cast<const uint4*>(h_pin + kb128 / 2), reinterpret_cast<uint4*>(d_buf + kb128 / 2), int(kb128 / 32));
                        null_kernel<<<148, 256, 0, s>>>(d_int);
                        CK(cudaStreamSynchronize(s));
                    }));
    }
    // results written straight to mapped pinned memory: how the 16 KB are spread over CTAs matters
    {
        const int shapes[][2] = {{1, 1024}, {4, 256}, {16, 64}, {32, 32}, {128, 8}};
        for (auto& sh : shapes) {
            std::printf(", \"push_16k_uint4_%dx%d_sync\": %.2f", sh[0], sh[1], med([&] {
                            push_kernel<<<sh[0], sh[1], 0, s>>>(reinterpret_cast<const uint4*>(d_buf), reinterpret_cast<uint4*>(h_pin), (16 << 10) / 16);
                            CK(cudaStreamSynchronize(s));
                        }));
        }
        const int shapes4[][2] = {{4, 1024}, {16, 256}, {64, 64}, {128, 32}};
        for (auto& sh : shapes4) {
            std::printf(", \"push_16k_uint32_%dx%d_sync\": %.2f", sh[0], sh[1], med([&] {
                            push32_kernel<<<sh[0], sh[1], 0, s>>>(reinterpret_cast<const uint32_t*>(d_buf), reinterpret_cast<uint32_t*>(h_pin), (16 << 10) / 4);
                            CK(cudaStreamSynchronize(s));
                        }));
        }
        std::printf(", \"store_16k_device_16x256_sync\": %.2f", med([&] {
                        push32_kernel<<<16, 256, 0, s>>>(reinterpret_cast<const uint32_t*>(d_buf), reinterpret_cast<uint32_t*>(d_buf + (64 << 10)), (16 << 10) / 4);
                        CK(cudaStreamSynchronize(s));
                    }));
    }
    // bookkeeping calls
    {
        cudaPointerAttributes a;
        std::printf(", \"pointer_get_attributes_pageable\": %.2f", med([&] { cudaPointerGetAttributes(&a, h_page); cudaGetLastError(); }, 2000));
        std::printf(", \"pointer_get_attributes_pinned\": %.2f", med([&] { cudaPointerGetAttributes(&a, h_pin); }, 2000));
        std::printf(", \"set_device\": %.2f", med([&] { cudaSetDevice(0); }, 2000));
        std::printf(", \"stream_query_idle\": %.2f", med([&] { cudaStreamQuery(s); }, 2000));
        cudaEvent_t ev;
        CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        std::printf(", \"launch_event_record_event_sync\": %.2f", med([&] {
                        null_kernel<<<148, 256, 0, s>>>(d_int);
                        CK(cudaEventRecord(ev, s));
                        CK(cudaEventSynchronize(ev));
                    }));
    }
    std::printf(", \"unit\": \"us, median of 400, steady_clock\"}\n");
    return 0;
}
