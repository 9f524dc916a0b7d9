// Micro-benchmark (not part of the product): issue rate of the integer dot-product instruction next to LOP3 / IMAD / FFMA on
// one SM, as a function of resident warps.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dp4a dp4a.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(int iters, int* out, long long* cyc) {
    int a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const int b = threadIdx.x * 0x01010101 + 0x03020100, c = 0x0F0F0F0F ^ threadIdx.x;
    float f0 = a0, f1 = a1, f2 = a2, f3 = a3, f4 = a4, f5 = a5, f6 = a6, f7 = a7; const float g = 1.0001f, h = 0.5f;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        if (OP == 0) { a0 = __dp4a(b, c, a0); a1 = __dp4a(b, c, a1); a2 = __dp4a(b, c, a2); a3 = __dp4a(b, c, a3); a4 = __dp4a(b, c, a4); a5 = __dp4a(b, c, a5); a6 = __dp4a(b, c, a6); a7 = __dp4a(b, c, a7); }
        if (OP == 1) { a0 = (a0 & b) ^ c; a1 = (a1 & b) ^ c; a2 = (a2 & b) ^ c; a3 = (a3 & b) ^ c; a4 = (a4 & b) ^ c; a5 = (a5 & b) ^ c; a6 = (a6 & b) ^ c; a7 = (a7 & b) ^ c; }
        if (OP == 2) { a0 = a0 * b + c; a1 = a1 * b + c; a2 = a2 * b + c; a3 = a3 * b + c; a4 = a4 * b + c; a5 = a5 * b + c; a6 = a6 * b + c; a7 = a7 * b + c; }
        if (OP == 3) { f0 = f0 * g + h; f1 = f1 * g + h; f2 = f2 * g + h; f3 = f3 * g + h; f4 = f4 * g + h; f5 = f5 * g + h; f6 = f6 * g + h; f7 = f7 * g + h; }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + (int)(f0 + f1 + f2 + f3 + f4 + f5 + f6 + f7);
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    int* out; long long* cyc; cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 8 * 1024);
    const char* names[] = {"IDP.4A", "LOP3", "IMAD", "FFMA"};
    for (int op = 0; op < 4; op++)
        for (int warps : {1, 4, 8, 16, 32}) {
            const int iters = 4096; long long h = 0;
            for (int rep = 0; rep < 2; rep++) {
                if (op == 0) k<0><<<1, warps * 32>>>(iters, out, cyc); if (op == 1) k<1><<<1, warps * 32>>>(iters, out, cyc);
                if (op == 2) k<2><<<1, warps * 32>>>(iters, out, cyc); if (op == 3) k<3><<<1, warps * 32>>>(iters, out, cyc);
                cudaDeviceSynchronize(); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
            }
            printf("%-7s %2d warps on one SM: %.3f warp-instructions / clk / SM\n", names[op], warps, (double)iters * 8 * warps / (double)h);
        }
    return 0;
}
