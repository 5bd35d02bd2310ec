// Probe: which (lane, column) of TMEM does each register of tcgen05.ld.16x256b / tcgen05.st.16x128b touch?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe/tmem_layout tools/probe/tmem_layout.cu ; run on a B200.
#include "../../avion_b200/csrc/common.cuh"
#include <cstdio>
using namespace avn;
void avn_set_error(const char*, ...) {}
__global__ void probe(uint32_t* out_ld, uint32_t* out_st) {
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(&tptr, 128);
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tb = tptr;
    const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
    // fill columns 0..63 with 32x32b stores: value = row*1000 + col
    uint32_t v[32];
    for (int c = 0; c < 2; ++c) {
        for (int i = 0; i < 32; ++i) v[i] = (warp * 32 + lane) * 1000 + c * 32 + i;
        tmem_st32(tb + lane_addr + c * 32, v);
    }
    tmem_st_wait();
    __syncthreads();
    // 16x256b.x8 load of lanes [warp*32 + 16*lg, +16), columns 0..63
    for (int lg = 0; lg < 2; ++lg) {
        uint32_t r[32];
        asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
              "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
              "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
              "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(tb + lane_addr + ((uint32_t)(16 * lg) << 16)));
        tmem_ld_wait();
        for (int i = 0; i < 32; ++i) out_ld[((warp * 2 + lg) * 32 + lane) * 32 + i] = r[i];
    }
    __syncthreads();
    // 16x128b.x8 store into columns 64..95: value = (lg, lane, reg) code; read back with 32x32b
    for (int lg = 0; lg < 2; ++lg) {
        uint32_t r[16];
        for (int i = 0; i < 16; ++i) r[i] = 100000 * (lg + 1) + lane * 100 + i;
        asm volatile("tcgen05.st.sync.aligned.16x128b.x8.b32 [%0], "
            "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
            ::"r"(tb + lane_addr + ((uint32_t)(16 * lg) << 16) + 64), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]),
              "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
    }
    tmem_st_wait();
    __syncthreads();
    tmem_ld32(tb + lane_addr + 64, v);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) out_st[(warp * 32 + lane) * 32 + i] = v[i];
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 128);
}
int main() {
    uint32_t *d_ld, *d_st;
    cudaMalloc(&d_ld, 4 * 2 * 32 * 32 * 4); cudaMalloc(&d_st, 128 * 32 * 4);
    probe<<<1, 128>>>(d_ld, d_st);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
    static uint32_t ld[4 * 2 * 32 * 32], st[128 * 32];
    cudaMemcpy(ld, d_ld, sizeof(ld), cudaMemcpyDeviceToHost); cudaMemcpy(st, d_st, sizeof(st), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int w = 0; w < 4; ++w) for (int lg = 0; lg < 2; ++lg) for (int T = 0; T < 32; ++T) for (int i = 0; i < 32; ++i) {
        const int j = i / 4, k = i % 4;
        const int row = w * 32 + 16 * lg + T / 4 + 8 * (k / 2), col = 8 * j + 2 * (T % 4) + (k % 2);
        if (ld[((w * 2 + lg) * 32 + T) * 32 + i] != (uint32_t)(row * 1000 + col)) ++bad;
    }
    printf("16x256b.x8 ld: expected mapping reg[4j+k] = (row T/4 + 8*(k/2), col 8j + 2*(T%%4) + k%%2): %s (%d mismatches)\n", bad ? "NO" : "YES", bad);
    if (bad) for (int i = 0; i < 8; ++i) printf("  warp0 lg0 T=1 reg %d -> %u\n", i, ld[(0 * 32 + 1) * 32 + i]);
    bad = 0;
    for (int row = 0; row < 128; ++row) for (int c = 0; c < 32; ++c) {
        const int lg = (row % 32) / 16, rr = row % 16, j = c / 4;
        const int T = (rr % 8) * 4 + c % 4, reg = 2 * j + rr / 8;
        if (st[row * 32 + c] != (uint32_t)(100000 * (lg + 1) + T * 100 + reg)) ++bad;
    }
    printf("16x128b.x8 st: expected mapping reg[2j+k] = (row T/4 + 8k, col 4j + T%%4): %s (%d mismatches)\n", bad ? "NO" : "YES", bad);
    if (bad) for (int c = 0; c < 8; ++c) printf("  row 1 col %d <- %u\n", c, st[1 * 32 + c]);
    return 0;
}
