// probe_ffma.cu -- FP32 issue-rate probes for the inner loop of k_integrate (Horner rows from
// shared memory): scalar FFMA vs packed FFMA2 (fma.rn.f32x2), register-bank effects.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/probes/probe_ffma scripts/probes/probe_ffma.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c)
{
    unsigned long long ra, rb, rc, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}

#define ROWS 1024
extern __shared__ float4 s_tab[];

template <int G, int MODE>
__global__ void __launch_bounds__(128) k_probe(const float4 *tab, float *out, int reps, float seed)
{
    for (int i = threadIdx.x; i < ROWS * 2; i += blockDim.x) s_tab[i] = tab[i];
    __syncthreads();
    float r[G];
#pragma unroll
    for (int g = 0; g < G; ++g) r[g] = 0.9f + 1e-4f * (threadIdx.x + g) + seed;
    float s = 0.f;
    if (MODE == 0) {                       // scalar, 7 columns
        float P[G][7];
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int c = 0; c < 7; ++c) P[g][c] = 0.f;
        for (int it = 0; it < reps; ++it) {
#pragma unroll 32
            for (int j = 0; j < ROWS; ++j) {
                float4 t0 = s_tab[2 * j], t1 = s_tab[2 * j + 1];
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    P[g][0] = fmaf(P[g][0], r[g], t0.x); P[g][1] = fmaf(P[g][1], r[g], t0.y);
                    P[g][2] = fmaf(P[g][2], r[g], t0.z); P[g][3] = fmaf(P[g][3], r[g], t0.w);
                    P[g][4] = fmaf(P[g][4], r[g], t1.x); P[g][5] = fmaf(P[g][5], r[g], t1.y);
                    P[g][6] = fmaf(P[g][6], r[g], t1.z);
                }
            }
        }
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int c = 0; c < 7; ++c) s += P[g][c];
    } else if (MODE == 1) {                // packed, 4 pairs (8 columns)
        float2 P[G][4];
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int c = 0; c < 4; ++c) P[g][c] = make_float2(0.f, 0.f);
        for (int it = 0; it < reps; ++it) {
#pragma unroll 32
            for (int j = 0; j < ROWS; ++j) {
                float4 t0 = s_tab[2 * j], t1 = s_tab[2 * j + 1];
                float2 t[4] = {make_float2(t0.x, t0.y), make_float2(t0.z, t0.w), make_float2(t1.x, t1.y), make_float2(t1.z, t1.w)};
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    float2 rr = make_float2(r[g], r[g]);
#pragma unroll
                    for (int c = 0; c < 4; ++c) P[g][c] = ffma2(P[g][c], rr, t[c]);
                }
            }
        }
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int c = 0; c < 4; ++c) s += P[g][c].x + P[g][c].y;
    } else if (MODE == 3) {                // packed over pairs of (sub-beam gate, species) pairs: 7 columns, no padding
        float2 P[G / 2][7];
#pragma unroll
        for (int g = 0; g < G / 2; ++g)
#pragma unroll
            for (int c = 0; c < 7; ++c) P[g][c] = make_float2(0.f, 0.f);
        for (int it = 0; it < reps; ++it) {
#pragma unroll 32
            for (int j = 0; j < ROWS; ++j) {
                float4 t0 = s_tab[2 * j], t1 = s_tab[2 * j + 1];
                const float t[7] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z};
#pragma unroll
                for (int g = 0; g < G / 2; ++g) {
                    float2 rr = make_float2(r[2 * g], r[2 * g + 1]);
#pragma unroll
                    for (int c = 0; c < 7; ++c) P[g][c] = ffma2(P[g][c], rr, make_float2(t[c], t[c]));
                }
            }
        }
#pragma unroll
        for (int g = 0; g < G / 2; ++g)
#pragma unroll
            for (int c = 0; c < 7; ++c) s += P[g][c].x + P[g][c].y;
    } else {                               // packed 3 pairs + 1 scalar (7 columns)
        float2 P[G][3];
        float Q[G];
#pragma unroll
        for (int g = 0; g < G; ++g) {
            Q[g] = 0.f;
#pragma unroll
            for (int c = 0; c < 3; ++c) P[g][c] = make_float2(0.f, 0.f);
        }
        for (int it = 0; it < reps; ++it) {
#pragma unroll 32
            for (int j = 0; j < ROWS; ++j) {
                float4 t0 = s_tab[2 * j], t1 = s_tab[2 * j + 1];
                float2 t[3] = {make_float2(t0.x, t0.y), make_float2(t0.z, t0.w), make_float2(t1.x, t1.y)};
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    float2 rr = make_float2(r[g], r[g]);
#pragma unroll
                    for (int c = 0; c < 3; ++c) P[g][c] = ffma2(P[g][c], rr, t[c]);
                    Q[g] = fmaf(Q[g], r[g], t1.z);
                }
            }
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
            s += Q[g];
#pragma unroll
            for (int c = 0; c < 3; ++c) s += P[g][c].x + P[g][c].y;
        }
    }
    if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int G, int MODE>
void run(const char *name, const float4 *tab, float *out, int cols_useful)
{
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    cudaFuncSetAttribute(k_probe<G, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, ROWS * 32);
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_probe<G, MODE>, 128, ROWS * 32);
    occ = occ > 4 ? 4 : occ;
    const int reps = 64;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int i = 0; i < 5; ++i) {
        cudaEventRecord(a);
        k_probe<G, MODE><<<sms * occ, 128, ROWS * 32>>>(tab, out, reps, 0.f);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (i && ms < best) best = ms;
    }
    double fma = (double)sms * occ * 128 * G * cols_useful * (double)ROWS * reps;
    double per_clk_sm = fma / (best * 1e-3) / (clk * 1e3) / sms;
    printf("%-34s G=%d occ=%d  %.3f ms  useful FMA/clk/SM %.1f (of 128)  err=%s\n", name, G, occ, best, per_clk_sm,
           cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    float4 *tab; float *out;
    cudaMalloc(&tab, ROWS * 32); cudaMemset(tab, 0, ROWS * 32);
    cudaMalloc(&out, 1 << 24);
    run<4, 0>("scalar FFMA 7 cols", tab, out, 7);
    run<4, 1>("packed FFMA2 4 pairs (7 useful)", tab, out, 7);
    run<4, 2>("packed 3 pairs + 1 scalar", tab, out, 7);
    run<8, 1>("packed FFMA2 4 pairs G=8", tab, out, 7);
    run<2, 0>("scalar FFMA 7 cols", tab, out, 7);
    run<8, 0>("scalar FFMA 7 cols", tab, out, 7);
    run<6, 2>("packed 3 pairs + 1 scalar G=6", tab, out, 7);
    run<4, 3>("packed over pairs, 7 cols", tab, out, 7);
    run<8, 3>("packed over pairs, 7 cols", tab, out, 7);
    run<6, 3>("packed over pairs, 7 cols", tab, out, 7);
    return 0;
}
