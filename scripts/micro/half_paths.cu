// Microbenchmark: are there slow paths in the packed-half instructions the wide node step uses (HFMA2 / HMNMX2 / VHMNMX / HSET2 / F2FP)
// for denormal, infinite or NaN operands?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o half_paths half_paths.cu
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <cmath>
__device__ __forceinline__ __half2 hb(uint32_t u) { return *reinterpret_cast<const __half2 *>(&u); }
__device__ __forceinline__ uint32_t ub(__half2 h) { return *reinterpret_cast<const uint32_t *>(&h); }

template <int OP>
__global__ void k(const uint32_t *in, uint32_t *out, int iters)
{
    // operands come from memory so that nothing folds; 8 independent chains
    uint32_t q[8];
    for (int i = 0; i < 8; i++) q[i] = in[i] + threadIdx.x % 3;
    const __half2 A = hb(in[8]), B = hb(in[9]), C = hb(in[10]);
    const float fa = __uint_as_float(in[11]), fb = __uint_as_float(in[12]);
    uint32_t acc = 0;
    for (int it = 0; it < iters; it++)
    {
#pragma unroll
        for (int i = 0; i < 8; i++)
        {
            uint32_t r;
            if (OP == 0) r = ub(__hfma2(hb(q[i]), A, B));
            if (OP == 1) r = ub(__hmax2(hb(q[i]), B));
            if (OP == 2) r = ub(__hmax2(__hmax2(hb(q[i]), A), B));  // VHMNMX
            if (OP == 3) r = __hle2_mask(hb(q[i]), B);
            if (OP == 4) r = ub(__floats2half2_rn(__uint_as_float(q[i]), fa));
            if (OP == 5) r = __float_as_uint(__fmaf_rn(__uint_as_float(q[i]), fa, fb));
            acc ^= r;
            q[i] += (OP == 4 || OP == 5) ? 0x00000100u : 0x00010001u;  // stays in the same class of values for the iteration counts used
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int OP>
float run(const uint32_t *h, const char *name)
{
    uint32_t *din, *dout;
    cudaMalloc(&din, 64 * 4);
    cudaMalloc(&dout, 148 * 8 * 256 * 4);
    cudaMemcpy(din, h, 64 * 4, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 20000;
    k<OP><<<148 * 8, 256>>>(din, dout, 100);
    cudaEventRecord(e0);
    k<OP><<<148 * 8, 256>>>(din, dout, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double inst = 148.0 * 8 * 8 * iters * 8;  // warp instructions of the op under test
    printf("%-58s %8.3f ms  %6.2f warp-inst/clk/SM (at 1.965 GHz)\n", name, ms, inst / (ms * 1e-3) / 1.965e9 / 148);
    cudaFree(din);
    cudaFree(dout);
    return ms;
}

static uint32_t h2(float lo, float hi)
{
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&v);
}
static uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

int main()
{
    uint32_t h[64] = {0};
    const uint32_t HINF = 0x7c007c00u, HNINF = 0xfc00fc00u, HNAN = 0x7e007e00u;
    auto setq = [&](uint32_t base) { for (int i = 0; i < 8; i++) h[i] = base + 0x00020002u * i; };
    // ---- HFMA2
    setq(h2(1.5f, 2.5f)); h[8] = h2(3.f, 3.f); h[9] = h2(0.25f, 0.25f);
    run<0>(h, "HFMA2 normal * normal + normal");
    setq(0x00100020u); h[8] = h2(20000.f, 20000.f); h[9] = h2(0.25f, 0.25f);
    run<0>(h, "HFMA2 denormal q * normal A + normal B -> normal");
    setq(0x00100020u); h[8] = h2(16.f, 16.f); h[9] = h2(1e-6f, 1e-6f);
    run<0>(h, "HFMA2 denormal q * normal A + denormal B -> denormal");
    setq(0x00100020u); h[8] = h2(3e-6f, 3e-6f); h[9] = h2(0.25f, 0.25f);
    run<0>(h, "HFMA2 denormal q * denormal A + normal B");
    setq(0x00100020u); h[8] = h2(20000.f, 20000.f); h[9] = HNINF;
    run<0>(h, "HFMA2 denormal q * normal A + (-inf)");
    setq(0x00100020u); h[8] = h2(20000.f, 20000.f); h[9] = HINF;
    run<0>(h, "HFMA2 denormal q * normal A + (+inf)");
    setq(0x00100020u); h[8] = h2(20000.f, 20000.f); h[9] = HNAN;
    run<0>(h, "HFMA2 denormal q * normal A + NaN");
    setq(0x00000000u); h[8] = HINF; h[9] = h2(0.25f, 0.25f);
    run<0>(h, "HFMA2 (0..) * inf + normal (NaN results)");
    // ---- HMNMX2 / VHMNMX / HSET2
    setq(h2(1.5f, 2.5f)); h[8] = h2(3.f, 3.f); h[9] = h2(0.25f, 0.25f);
    run<1>(h, "HMNMX2 normal");
    h[9] = HINF;  run<1>(h, "HMNMX2 with +inf");
    h[9] = HNINF; run<1>(h, "HMNMX2 with -inf");
    h[9] = HNAN;  run<1>(h, "HMNMX2 with NaN");
    setq(0x00100020u); h[9] = h2(1e-6f, 1e-6f); run<1>(h, "HMNMX2 denormals");
    setq(h2(1.5f, 2.5f)); h[8] = h2(3.f, 3.f); h[9] = h2(0.25f, 0.25f);
    run<2>(h, "VHMNMX (3-input) normal");
    h[9] = HNINF; run<2>(h, "VHMNMX with -inf");
    h[9] = HINF; run<2>(h, "VHMNMX with +inf");
    h[9] = HNAN; run<2>(h, "VHMNMX with NaN");
    setq(0x00100020u); h[8] = h2(2e-6f, 2e-6f); h[9] = h2(1e-6f, 1e-6f); run<2>(h, "VHMNMX denormals");
    setq(h2(1.5f, 2.5f)); h[9] = h2(0.25f, 0.25f);
    run<3>(h, "HSET2.LE normal");
    h[9] = HINF; run<3>(h, "HSET2.LE with +inf");
    h[9] = HNAN; run<3>(h, "HSET2.LE with NaN");
    setq(0x00100020u); h[9] = h2(1e-6f, 1e-6f); run<3>(h, "HSET2.LE denormals");
    // ---- F2FP
    setq(fbits(1.5f)); h[11] = fbits(2.5f);
    run<4>(h, "F2FP normal");
    setq(fbits(1e-6f)); h[11] = fbits(2e-6f);
    run<4>(h, "F2FP -> denormal halves");
    setq(fbits(1e9f)); h[11] = fbits(-1e9f);
    run<4>(h, "F2FP overflow -> inf");
    setq(fbits(1e-12f)); h[11] = fbits(1e-12f);
    run<4>(h, "F2FP underflow -> 0");
    setq(fbits(INFINITY)); h[11] = fbits(NAN);
    run<4>(h, "F2FP inf / NaN");
    // ---- FFMA with float denormals (the scaled constants)
    setq(fbits(1.5f)); h[11] = fbits(2.5f); h[12] = fbits(0.5f);
    run<5>(h, "FFMA normal");
    setq(fbits(1e-39f)); h[11] = fbits(0.5f); h[12] = fbits(1e-40f);
    run<5>(h, "FFMA float denormals");
    return 0;
}
