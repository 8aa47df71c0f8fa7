// Probe: chained IF conditional graph nodes against plain empty launches inside a CUDA graph (profiles/r02_sweeps.txt R).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o cond_probe cond_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include <chrono>
// chain of IF nodes: body b does work and enables IF b+1 when flag[b] != 0 (set from inside a body graph for a sibling node)
__global__ void stage(unsigned *out, const unsigned *flag, int b, cudaGraphConditionalHandle next)
{
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        atomicAdd(out, 1u);
        if (next && flag[b]) cudaGraphSetConditional(next, 1);
    }
}
int main()
{
    cudaStream_t s, s2; cudaStreamCreate(&s); cudaStreamCreate(&s2);
    unsigned *d, *flag; cudaMalloc(&d, 8); cudaMalloc(&flag, 64 * 4);
    const int depth = 10, grid = 1184, threads = 128;
    for (int mode = 0; mode < 2; mode++)  // 0: plain chain of 3 kernels per bounce, 1: IF nodes
    {
        cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed);
        cudaStreamCaptureStatus st; cudaGraph_t g; const cudaGraphNode_t *deps; size_t nd;
        cudaStreamGetCaptureInfo(s, &st, nullptr, &g, &deps, &nd);
        std::vector<cudaGraphConditionalHandle> h(depth + 1, 0);
        if (mode) for (int b = 1; b < depth; b++) if (cudaGraphConditionalHandleCreate(&h[b], g, 0, cudaGraphCondAssignDefault) != cudaSuccess) printf("create failed\n");
        for (int b = 0; b < depth; b++)
        {
            cudaStream_t t = s;
            if (mode && b > 0)
            {
                cudaStreamGetCaptureInfo(s, &st, nullptr, &g, &deps, &nd);
                cudaGraphNodeParams cp = {cudaGraphNodeTypeConditional};
                cp.conditional.handle = h[b]; cp.conditional.type = cudaGraphCondTypeIf; cp.conditional.size = 1;
                cudaGraphNode_t cn;
                if (cudaGraphAddNode(&cn, g, deps, nd, &cp) != cudaSuccess) printf("add failed\n");
                cudaStreamUpdateCaptureDependencies(s, &cn, 1, cudaStreamSetCaptureDependencies);
                if (cudaStreamBeginCaptureToGraph(s2, cp.conditional.phGraph_out[0], nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed) != cudaSuccess) printf("begin failed\n");
                t = s2;
            }
            stage<<<grid, threads, 0, t>>>(d, flag, b, 0);
            stage<<<grid, threads, 0, t>>>(d, flag, b, 0);
            stage<<<grid, threads, 0, t>>>(d, flag, b, mode ? h[b + 1] : 0);
            if (t == s2 && cudaStreamEndCapture(s2, nullptr) != cudaSuccess) printf("end body failed\n");
        }
        cudaGraph_t graph; if (cudaStreamEndCapture(s, &graph) != cudaSuccess) printf("end failed\n");
        cudaGraphExec_t ex; if (cudaGraphInstantiate(&ex, graph, 0) != cudaSuccess) printf("inst failed %s\n", cudaGetErrorString(cudaGetLastError()));
        for (int live : {1, 3, 10})
        {
            std::vector<unsigned> f(64, 0);
            for (int b = 0; b + 1 < live; b++) f[b] = 1;  // bounce b leaves survivors
            cudaMemcpy(flag, f.data(), 64 * 4, cudaMemcpyHostToDevice);
            cudaMemset(d, 0, 4);
            cudaGraphLaunch(ex, s); cudaStreamSynchronize(s);
            unsigned v; cudaMemcpy(&v, d, 4, cudaMemcpyDeviceToHost);
            cudaGraphLaunch(ex, s); cudaStreamSynchronize(s);
            auto t0 = std::chrono::steady_clock::now();
            for (int i = 0; i < 200; i++) cudaGraphLaunch(ex, s);
            cudaStreamSynchronize(s);
            double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / 200;
            printf("mode %d live bounces %2d: kernels run %u (IF expects %d), %.1f us per graph launch, err %s\n", mode, live, v, mode ? 3 * live : 30, us,
                   cudaGetErrorString(cudaGetLastError()));
        }
    }
    return 0;
}
