// Integer-pipe micro-benchmarks for the 64-bit modular butterfly on sm_100a (run on the GPU box):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/int_pipe_probe tools/int_pipe_probe.cu && gpurun_out/int_pipe_probe
// Measures (1) raw issue rates of IMAD / IMAD.WIDE / IADD3 and their overlap, (2) butterflies/s of several
// formulations of the lazy Shoup butterfly, all register-resident, every SM busy.  The numbers set the cost model
// used in DESIGN.md (IMAD.WIDE = 2 fmaheavy slots) and the denominator of the integer roofline.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;

#define REP16(X) X X X X X X X X X X X X X X X X

// ---------------- raw instruction streams: 8 independent chains per thread ----------------
template <int MODE>
__global__ void __launch_bounds__(256) k_raw(int rounds, u64 *sink, uint32_t m0, uint32_t m1) {
    uint32_t a[8]; u64 w[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 8 + i + m0; w[i] = (u64)a[i] * m1 + i; }
    for (int it = 0; it < rounds; it++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) a[i] = a[i] * m0 + m1;                                   // IMAD
                if (MODE == 1) w[i] = (u64)(uint32_t)w[i] * m0 + w[i];                  // IMAD.WIDE.U32 with 64-bit addend
                if (MODE == 2) a[i] = (a[i] + m0) ^ m1;                                 // IADD3 + LOP3 (alu pipe)
                if (MODE == 3) { w[i] = (u64)(uint32_t)w[i] * m0 + w[i]; a[i] = (a[i] + m0) ^ m1; }   // WIDE + 2 alu
                if (MODE == 4) { w[i] = (u64)(uint32_t)w[i] * m0 + w[i]; a[i] = a[i] * m0 + m1; }     // WIDE + IMAD
                if (MODE == 5) { a[i] = a[i] * m0 + m1; a[i] = (a[i] + m0) ^ m1; }      // IMAD + 2 alu (dependent)
            }
        }
    }
    u64 x = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) x ^= a[i] ^ w[i];
    if (x == 0x123456789ull) sink[0] = x;
}

// ---------------- butterfly formulations ----------------
__device__ __forceinline__ u64 shoup_exact(u64 y, u64 w, u64 ws, u64 q) { return y * w - __umul64hi(y, ws) * q; }
// quotient estimate without the low x low product: hi' in {hi-1, hi}  ->  result in [0, 3q)
__device__ __forceinline__ u64 shoup_approx(u64 y, u64 w, u64 ws, u64 nq) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 yl, yh, wl, wh, sl, sh, nl, nh, alo, ahi, clo, chi, mlo, mhi, mc, hl, hh, tlo, thi;\n\t"
        ".reg .u64 a, c, ad, hi, t;\n\t"
        "mov.b64 {yl, yh}, %1;\n\t" "mov.b64 {wl, wh}, %2;\n\t" "mov.b64 {sl, sh}, %3;\n\t" "mov.b64 {nl, nh}, %4;\n\t"
        "mul.wide.u32 a, yl, sh;\n\t" "mul.wide.u32 c, yh, sl;\n\t"
        "mov.b64 {alo, ahi}, a;\n\t" "mov.b64 {clo, chi}, c;\n\t"
        "add.cc.u32 mlo, alo, clo;\n\t" "addc.cc.u32 mhi, ahi, chi;\n\t" "addc.u32 mc, 0, 0;\n\t"
        "mov.b64 ad, {mhi, mc};\n\t"
        "mad.wide.u32 hi, yh, sh, ad;\n\t"
        "mov.b64 {hl, hh}, hi;\n\t"
        "mul.wide.u32 t, yl, wl;\n\t" "mad.wide.u32 t, hl, nl, t;\n\t"
        "mov.b64 {tlo, thi}, t;\n\t"
        "mad.lo.u32 thi, yl, wh, thi;\n\t" "mad.lo.u32 thi, yh, wl, thi;\n\t"
        "mad.lo.u32 thi, hl, nh, thi;\n\t" "mad.lo.u32 thi, hh, nl, thi;\n\t"
        "mov.b64 %0, {tlo, thi};\n\t"
        "}" : "=l"(r) : "l"(y), "l"(w), "l"(ws), "l"(nq));
    return r;
}
// quotient from the high words only: hi' in {hi-3 .. hi}  ->  result in [0, 5q)  (2 wide products for the quotient)
__device__ __forceinline__ u64 shoup_top(u64 y, u64 w, u64 ws, u64 q) {
    const uint32_t yh = (uint32_t)(y >> 32), sh = (uint32_t)(ws >> 32), yl = (uint32_t)y, sl = (uint32_t)ws;
    const u64 hi = (u64)yh * sh + (((u64)yh * sl) >> 32) + (u64)__umulhi(yl, sh);
    return y * w - hi * q;
}

template <int V>
__device__ __forceinline__ void bfly(u64 &x, u64 &y, const ulonglong2 w, u64 q, u64 nq, u64 qk) {
    u64 t;
    if (V == 0) t = shoup_exact(y, w.x, w.y, q);
    if (V == 1) t = shoup_approx(y, w.x, w.y, nq);
    if (V == 2) t = shoup_top(y, w.x, w.y, q);
    y = x - t + qk;
    x = x + t;
}
template <int V>
__global__ void __launch_bounds__(256) k_bfly(int rounds, u64 *sink, u64 q, u64 nq, u64 qk, const ulonglong2 *tw) {
    u64 a[8];
#pragma unroll
    for (int r = 0; r < 8; r++) a[r] = (u64)(threadIdx.x * 8 + r + blockIdx.x) % q;
    ulonglong2 w[7];
#pragma unroll
    for (int r = 0; r < 7; r++) w[r] = tw[r];
    for (int it = 0; it < rounds; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) bfly<V>(a[r], a[r + 4], w[0], q, nq, qk);
        bfly<V>(a[0], a[2], w[1], q, nq, qk); bfly<V>(a[1], a[3], w[1], q, nq, qk);
        bfly<V>(a[4], a[6], w[2], q, nq, qk); bfly<V>(a[5], a[7], w[2], q, nq, qk);
#pragma unroll
        for (int r = 0; r < 4; r++) bfly<V>(a[2 * r], a[2 * r + 1], w[3 + r], q, nq, qk);
        // values wrap mod 2^64 after a few rounds; the instruction stream (what is measured) does not depend on them
    }
    u64 x = 0;
#pragma unroll
    for (int r = 0; r < 8; r++) x ^= a[r];
    if (x == 0x1234567) sink[0] = x;
}

template <typename F>
static double time_ms(F launch) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(8);
    cudaEventRecord(e0); launch(2048); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int sms = pr.multiProcessorCount;
    printf("device %s, %d SMs, clock %d MHz\n", pr.name, sms, clk_khz / 1000);
    u64 *sink; cudaMalloc(&sink, 8);
    ulonglong2 htw[7]; const u64 q = 288230376150630401ull;
    for (int i = 0; i < 7; i++) { htw[i].x = (q / 3 + 977 * i) % q; htw[i].y = (u64)(((unsigned __int128)htw[i].x << 64) / q); }
    ulonglong2 *tw; cudaMalloc(&tw, sizeof(htw)); cudaMemcpy(tw, htw, sizeof(htw), cudaMemcpyHostToDevice);
    const double hz = clk_khz * 1e3;
    for (int wpb = 1; wpb <= 2; wpb++) {
        const int blocks = sms * 4 * wpb, threads = 256;   // 32 or 64 warps per SM
        printf("--- %d warps per SM ---\n", blocks * threads / 32 / sms);
        const char *names[6] = {"IMAD", "IMAD.WIDE(+64b addend)", "IADD3+LOP3", "WIDE + 2 alu", "WIDE + IMAD", "IMAD + 2 alu"};
        const int per[6] = {1, 1, 2, 3, 2, 3};
#define RAW(M) { double ms = time_ms([&](int r) { k_raw<M><<<blocks, threads>>>(r, sink, 12345u, 777u); });                      \
            double inst = (double)blocks * threads / 32 * 2048.0 * 16 * 8 * per[M];                                                \
            printf("raw %-24s %8.3f ms  %.3f warp-inst/clk/SM (%.2f per SMSP)\n", names[M], ms, inst / (ms * 1e-3) / hz / sms,    \
                   inst / (ms * 1e-3) / hz / sms / 4); }
        RAW(0) RAW(1) RAW(2) RAW(3) RAW(4) RAW(5)
        const char *bn[3] = {"exact (compiler __umul64hi)", "approx 3-wide quotient [0,3q)", "top-words quotient [0,5q)"};
#define BF(V) { double ms = time_ms([&](int r) { k_bfly<V><<<blocks, threads>>>(r, sink, q, 0 - q, (V == 0 ? 2 : V == 1 ? 3 : 5) * q, tw); });                               \
            double bf = (double)blocks * threads * 2048.0 * 12;                                                                     \
            printf("bfly %-32s %8.3f ms  %.1f Gbfly/s  %.2f bfly/clk/SM\n", bn[V], ms, bf / (ms * 1e-3) / 1e9, bf / (ms * 1e-3) / hz / sms); }
        BF(0) BF(1) BF(2)
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
