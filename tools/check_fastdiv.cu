// Developer tool (GPU): DivBy / div_fast of scn_march.cuh against __fdiv_rn, bit for bit, over the operand ranges the march kernels
// feed it (divisors: cell sizes and the 2.5^k level scales; dividends: coordinates of any magnitude inside the guard, both signs).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I sparse_conv_gpis_tungsten_b200/csrc -o gpurun_out/check_fastdiv tools/check_fastdiv.cu && gpurun_out/check_fastdiv
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "scn_march.cuh"
using namespace scn;
__device__ uint32_t mix(uint32_t a) { a ^= a >> 16; a *= 0x7feb352du; a ^= a >> 15; a *= 0x846ca68bu; a ^= a >> 16; return a; }
__global__ void k(unsigned long long *bad, unsigned long long *slow, unsigned long long n_per_thread, int mode) {
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long b = 0, sl = 0;
    for (unsigned long long i = 0; i < n_per_thread; ++i) {
        const uint32_t h1 = mix((uint32_t)(tid * n_per_thread + i) ^ 0x9e3779b9u), h2 = mix(h1 + 0x85ebca6bu), h3 = mix(h2 ^ 0xc2b2ae35u);
        float c, x;
        if (mode == 0) { // divisor 2.5^k (float pow like the host table), k in [-20, 20], or an arbitrary cell size in [1e-3, 1e3]
            c = (h3 & 1) ? powf(2.5f, (float)((int)(h3 >> 1) % 41 - 20)) : exp2f(((h3 >> 1) & 0xffff) * (20.0f / 65536.f) - 10.f);
            x = __uint_as_float((h1 & 0x807fffffu) | ((uint32_t)(127 - 40 + (h2 % 80)) << 23)); // |x| in 2^-40 .. 2^40, random mantissa and sign
        } else { // any bit patterns (includes zeros, denormals, infinities, NaNs): the guard must route them to __fdiv_rn
            c = __uint_as_float(h3 & 0x7fffffffu); x = __uint_as_float(h1);
        }
        const DivBy d = div_by(c);
        const f3 q = xdiv3_by(F3(x, x, x), d);
        const float want = __fdiv_rn(x, c);
        const bool same = __float_as_uint(q.x) == __float_as_uint(want) || (want != want && q.x != q.x);
        b += same ? 0 : 1;
        sl += (fabsf(x) > d.lo && fabsf(x) < 1e18f) ? 0 : 1;
    }
    atomicAdd(bad, b); atomicAdd(slow, sl);
}
int main() {
    unsigned long long *bad, *slow, hb, hs;
    cudaMalloc(&bad, 8); cudaMalloc(&slow, 8);
    for (int mode = 0; mode < 2; ++mode) {
        cudaMemset(bad, 0, 8); cudaMemset(slow, 0, 8);
        const unsigned long long per = 4096; const int blocks = 148 * 8, threads = 256;
        k<<<blocks, threads>>>(bad, slow, per, mode);
        cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&hs, slow, 8, cudaMemcpyDeviceToHost);
        printf("mode %d (%s): %llu quotients, %llu through the __fdiv_rn guard, %llu differ from __fdiv_rn\n", mode, mode == 0 ? "lattice-range operands" : "arbitrary bit patterns",
               per * blocks * threads, hs, hb);
        if (cudaGetLastError() != cudaSuccess) { printf("CUDA error\n"); return 2; }
    }
    return 0;
}
