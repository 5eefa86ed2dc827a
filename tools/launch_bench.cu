// Diagnostics: what one dependent kernel node of a CUDA graph costs on this GPU as a function of cluster size, block
// size and dynamic shared memory -- the floor under the single-launch batch-norm passes (8 MiB tensors: the body of
// a pass is ~5 us of SM-active time, profiles/r2_r2c_bn_fused.md).
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/bin/launch_bench tools/launch_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void body(int mode, float* sink) {
    // mode 0: nothing.  1: one cluster barrier.  2: two cluster barriers
    if (mode >= 1) {
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    if (mode >= 2) {
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    if (sink != nullptr && threadIdx.x == 9999) *sink = 1.f;
}

static float run(int grid, int block, int cluster, size_t smem, int mode, int nodes, int reps, int coop = 0) {
    cudaStream_t s;
    cudaStreamCreate(&s);
    cudaFuncSetAttribute(body, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(body, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = cluster > 1 ? 1 : 0;
    if (coop) { attr[0].id = cudaLaunchAttributeCooperative; attr[0].val.cooperative = 1; cfg.numAttrs = 1; }
    float* sink = nullptr;
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < nodes; ++i)
        if (cudaLaunchKernelEx(&cfg, body, mode, sink) != cudaSuccess) { cudaStreamEndCapture(s, &g); return -1.f; }
    if (cudaStreamEndCapture(s, &g) != cudaSuccess) return -1.f;
    if (cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) return -1.f;
    cudaGraphLaunch(ge, s);
    cudaStreamSynchronize(s);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, s);
    for (int r = 0; r < reps; ++r) cudaGraphLaunch(ge, s);
    cudaEventRecord(e1, s);
    cudaStreamSynchronize(s);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaGraphExecDestroy(ge); cudaGraphDestroy(g); cudaStreamDestroy(s);
    if (cudaGetLastError() != cudaSuccess) return -1.f;
    return ms * 1e3f / (float)(nodes * reps);
}

int main() {
    const int clusters[] = {1, 2, 4, 8, 16};
    const size_t smems[] = {0, 64 * 1024, 200 * 1024};
    printf("us per dependent graph node (100 nodes x 20 replays)\n");
    for (int block : {256, 512, 1024})
        for (size_t smem : smems)
            for (int cl : clusters) {
                const int grid = (cl == 16) ? 112 : 128;            // what fits in one wave on B200
                const float t0 = run(grid, block, cl, smem, 0, 100, 20);
                const float t1 = cl > 1 ? run(grid, block, cl, smem, 1, 100, 20) : 0.f;
                const float t2 = cl > 1 ? run(grid, block, cl, smem, 2, 100, 20) : 0.f;
                printf("block %4d smem %6zu cluster %2d grid %3d: empty %.2f  +1 cluster barrier %.2f  +2 %.2f\n", block, smem,
                       cl, grid, t0, t1, t2);
            }
    for (int block : {256, 544})
        for (size_t smem : smems)
            printf("cooperative launch: block %4d smem %6zu grid 148: empty %.2f us per node (plain launch %.2f)\n", block, smem,
                   run(148, block, 1, smem, 0, 100, 20, 1), run(148, block, 1, smem, 0, 100, 20, 0));
    return 0;
}
