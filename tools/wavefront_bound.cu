// wavefront_bound.cu — upper bound for a WAVEFRONT formulation of the integrator (measurement aid, not product).
//
// A wavefront design keeps the per-job path state in global memory (SoA) and runs one kernel per stage over
// compacted index lists (warp ballot + prefix sum). Every HDDA step / free-flight draw of a lane then costs at
// least one read and one write of the state that stage touches. This program measures exactly that floor on the
// B200: a kernel that only streams a W-word SoA state in and out through an index list (one trivial add so the
// traffic is not optimised away), for a state that fits in L2 and one that does not. Divide the lane-steps/s by
// the counted steps per pixel-sample (dense 512^3: 183 HDDA steps + 155 draws, bench.py roofline.per_pixel_sample)
// to get the pixel-samples/s no wavefront design can exceed, before ANY arithmetic, compaction or launch cost.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/wavefront_bound.cu -o build/wavefront_bound
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

template <int W>
__global__ void stream_state(float* __restrict__ state, const int* __restrict__ index, int n, int stride)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int k = index[i]; // compacted list: consecutive lanes -> consecutive (coalesced) state slots
    float v[W];
#pragma unroll
    for (int w = 0; w < W; ++w) v[w] = state[size_t(w) * stride + k];
#pragma unroll
    for (int w = 0; w < W; ++w) state[size_t(w) * stride + k] = v[w] + 1.0f;
}

template <int W>
static void run(int n, const char* label, double steps_per_sample)
{
    float* state; int* index;
    cudaMalloc(&state, size_t(n) * W * sizeof(float));
    cudaMemset(state, 0, size_t(n) * W * sizeof(float));
    std::vector<int> h(n);
    for (int i = 0; i < n; ++i) h[i] = i;
    cudaMalloc(&index, size_t(n) * sizeof(int));
    cudaMemcpy(index, h.data(), size_t(n) * sizeof(int), cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 5; ++i) stream_state<W><<<(n + 255) / 256, 256>>>(state, index, n, n);
    const int reps = 50;
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) stream_state<W><<<(n + 255) / 256, 256>>>(state, index, n, n);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double lane_steps = double(n) * reps / (ms * 1e-3);
    printf("%-28s W=%2d words  n=%9d  state %7.1f MB : %7.2f G lane-steps/s  %7.1f GB/s (r+w)  => <= %6.1f M pixel-samples/s\n", label, W, n,
           double(n) * W * 4 / 1e6, lane_steps / 1e9, lane_steps * W * 8 / 1e9, lane_steps / steps_per_sample / 1e6);
    cudaFree(state); cudaFree(index);
}

int main(int argc, char** argv)
{
    const double steps = argc > 1 ? atof(argv[1]) : 338.0; // HDDA steps + draws per pixel-sample
    // 24 words = ray (9) + HDDA (10) + segment (3) + RNG (2); 40 words = + radiance, throughput, path and pixel bookkeeping
    run<24>(1 << 18, "fits L2 (hot stages only)", steps);
    run<40>(1 << 18, "fits L2 (full state)", steps);
    run<24>(1 << 22, "HBM (hot stages only)", steps);
    run<40>(1 << 22, "HBM (full state)", steps);
    return 0;
}
