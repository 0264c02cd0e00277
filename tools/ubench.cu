// ubench.cu -- FP32 pipe micro-benchmarks for sm_100a: issue cost of packed FFMA2/FMUL2/FADD2 vs scalar
// FFMA under different register-operand patterns and warps per SM sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench tools/ubench.cu && ./ubench
#include <cstdio>
#include <cuda_runtime.h>

constexpr int U = 12;
constexpr int ITER = 32768;

template <int MODE>
__global__ void __launch_bounds__(128) k(const float2* __restrict__ in, float* out, long long* cyc) {
    float2 x[U], a[U], b[U];
    float sc[U];
#pragma unroll
    for (int i = 0; i < U; ++i) sc[i] = reinterpret_cast<const float*>(in)[threadIdx.x + 128 * (6 * U + i)];
#pragma unroll
    for (int i = 0; i < U; ++i) {
        x[i] = in[threadIdx.x + 128 * i];
        a[i] = in[threadIdx.x + 128 * (U + i)];
        b[i] = in[threadIdx.x + 128 * (2 * U + i)];
    }
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < U; ++i) {
            if (MODE == 0) x[i] = __ffma2_rn(a[i], b[i], x[i]);               // 3 distinct register pairs
            if (MODE == 1) x[i] = __ffma2_rn(a[i], b[0], x[i]);               // one operand shared
            if (MODE == 2) x[i] = __ffma2_rn(x[i], a[0], b[0]);               // chain form, two shared
            if (MODE == 3) { x[i].x = fmaf(a[i].x, b[i].x, x[i].x); x[i].y = fmaf(a[i].y, b[i].y, x[i].y); }
            if (MODE == 4) x[i] = __fmul2_rn(x[i], a[i]);
            if (MODE == 5) x[i] = __fadd2_rn(x[i], a[i]);
            if (MODE == 6) { x[i].x = fmaf(x[i].x, a[0].x, b[0].x); x[i].y = fmaf(x[i].y, a[0].x, b[0].x); }
            if (MODE == 7) x[i] = __ffma2_rn(a[i], b[i], x[(i + 1) % U]);      // writes a different pair
            if (MODE == 8) x[i] = __ffma2_rn(a[0], b[i], x[i]);               // slot-A operand shared
            if (MODE == 9) x[i] = __ffma2_rn(a[i], a[i], x[i]);               // same pair twice
            if (MODE == 10) { x[i] = __ffma2_rn(a[i], b[i], x[i]); a[i] = __fadd2_rn(a[i], b[(i + 5) % U]); }
            if (MODE == 11) { x[i] = __ffma2_rn(a[i], b[i], x[i]); if (i % 6 == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b[i].x)); }
            if (MODE == 12) { x[i] = __ffma2_rn(a[i / 4], b[i], x[i]); }      // slot-A operand shared in runs of 4
            // round 2: 32-bit broadcast operand forms (SASS: Rn.F32)
            if (MODE == 13) x[i] = __ffma2_rn(a[i], make_float2(sc[i], sc[i]), x[i]);          // pair * scalar + pair, all distinct
            if (MODE == 14) x[i] = __ffma2_rn(a[i], make_float2(sc[0], sc[0]), x[i]);          // scalar shared
            if (MODE == 15) x[i] = __ffma2_rn(a[0], make_float2(sc[i], sc[i]), x[i]);          // pair shared, scalar distinct
            if (MODE == 16) x[i] = __ffma2_rn(x[i], a[i], make_float2(1.25f, 1.25f));          // 2 pairs + constant
            if (MODE == 17) x[i] = __fmul2_rn(x[i], make_float2(sc[i], sc[i]));                // FMUL2 pair * scalar
            if (MODE == 18) x[i] = __ffma2_rn(a[i], make_float2(sc[i], sc[i]), make_float2(sc[(i + 1) % U], sc[(i + 1) % U]));  // pair * scalar + scalar -> pair
            if (MODE == 19) { x[i] = __ffma2_rn(a[i], make_float2(sc[i], sc[i]), x[i]); if (i % 6 == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b[i].x)); }
            if (MODE == 20) { x[i] = __ffma2_rn(a[i], make_float2(sc[i], sc[i]), x[i]); if (i % 3 == 0) asm volatile("min.f32 %0, %0, %1;" : "+f"(b[i].x) : "f"(b[i].y)); }
            if (MODE == 21) x[i] = __ffma2_rn(a[i / 3], make_float2(sc[i], sc[i]), x[i]);      // the statistics pattern: g shared in runs of 3
            if (MODE == 22) x[i] = __ffma2_rn(a[i], make_float2(sc[i / 3], sc[i / 3]), x[i]);  // the z-column pattern: x shared in runs of 3
        }
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < U; ++i) s += x[i].x + x[i].y;
    if (s == 123.456f) out[0] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, const float2* in, float* out, long long* cyc, int sms) {
    for (int bps = 1; bps <= 4; ++bps) {
        int grid = sms * bps;
        k<MODE><<<grid, 128>>>(in, out, cyc);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        k<MODE><<<grid, 128>>>(in, out, cyc);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long c[1]; cudaMemcpy(c, cyc, 8, cudaMemcpyDeviceToHost);
        // warp-instructions per SMSP: bps warps (one per block per SMSP) x ITER x U (x2 for scalar modes)
        double inst = (double)bps * ITER * U * ((MODE == 3 || MODE == 6) ? 2 : 1);
        double lane_flops = (double)grid * 128 * ITER * U * 2 * ((MODE == 4 || MODE == 5 || MODE == 17) ? 1 : (MODE == 10 ? 3 : 2));
        printf("%-34s warps/SMSP=%d  cycles/inst/SMSP=%.3f  %.1f TFLOP/s  (%.3f ms)\n", name, bps, c[0] / inst,
               lane_flops / (ms * 1e-3) / 1e12, ms);
    }
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float2* in; float* out; long long* cyc;
    cudaMalloc(&in, 128 * 4 * U * sizeof(float2));
    cudaMemset(in, 0, 128 * 4 * U * sizeof(float2));
    cudaMalloc(&out, 16);
    cudaMalloc(&cyc, 8 * sms * 4);
    run<0>("FFMA2 a[i]*b[i]+x[i] (3 distinct)", in, out, cyc, sms);
    run<1>("FFMA2 a[i]*b0+x[i]", in, out, cyc, sms);
    run<2>("FFMA2 x[i]*a0+b0 (chain)", in, out, cyc, sms);
    run<3>("FFMA  a[i]*b[i]+x[i] (3 distinct)", in, out, cyc, sms);
    run<6>("FFMA  x[i]*a0+b0 (chain)", in, out, cyc, sms);
    run<4>("FMUL2 x[i]*a[i]", in, out, cyc, sms);
    run<5>("FADD2 x[i]+a[i]", in, out, cyc, sms);
    run<7>("FFMA2 a[i]*b[i]+x[i+1] -> x[i]", in, out, cyc, sms);
    run<8>("FFMA2 a0*b[i]+x[i]", in, out, cyc, sms);
    run<9>("FFMA2 a[i]*a[i]+x[i]", in, out, cyc, sms);
    run<10>("FFMA2 3-distinct + FADD2 1:1", in, out, cyc, sms);
    run<11>("FFMA2 3-distinct + MUFU 1:6", in, out, cyc, sms);
    run<12>("FFMA2 a[i/4]*b[i]+x[i]", in, out, cyc, sms);
    run<13>("FFMA2 a[i]*s[i].F32+x[i]", in, out, cyc, sms);
    run<14>("FFMA2 a[i]*s0.F32+x[i]", in, out, cyc, sms);
    run<15>("FFMA2 a0*s[i].F32+x[i]", in, out, cyc, sms);
    run<16>("FFMA2 x[i]*a[i]+const", in, out, cyc, sms);
    run<17>("FMUL2 x[i]*s[i].F32", in, out, cyc, sms);
    run<18>("FFMA2 a[i]*s[i].F32+s[i+1].F32", in, out, cyc, sms);
    run<19>("FFMA2 pair*F32+pair + MUFU 1:6", in, out, cyc, sms);
    run<20>("FFMA2 pair*F32+pair + FMNMX 1:3", in, out, cyc, sms);
    run<21>("FFMA2 a[i/3]*s[i].F32+x[i]", in, out, cyc, sms);
    run<22>("FFMA2 a[i]*s[i/3].F32+x[i]", in, out, cyc, sms);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
