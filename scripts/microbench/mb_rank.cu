// Microbenchmarks that size the ranking primitives on B200 (sm_100a): cycles per warp-level op.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define ITERS 256
__device__ __forceinline__ uint32_t hash(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int MODE>
__global__ void __launch_bounds__(512) k(uint32_t* out, int distinct_mask)
{
    __shared__ uint32_t s[16 * 256 + 32 * 256];
    for (int i = threadIdx.x; i < 16 * 256 + 32 * 256; i += blockDim.x) s[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t x = hash(threadIdx.x + blockIdx.x * blockDim.x + 1);
    uint32_t acc = 0;
    uint32_t* my = s + warp * 256;
    const uint32_t lt = (1u << lane) - 1;
#pragma unroll 4
    for (int it = 0; it < ITERS; ++it) {
        x = x * 1664525u + 1013904223u;
        uint32_t d = (x >> 24) & distinct_mask;
        if (MODE == 0) { // hardware match.any
            acc += __popc(__match_any_sync(0xffffffffu, d) & lt);
        } else if (MODE == 1) { // ballot loop
            uint32_t peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; ++b) { bool p = (d >> b) & 1; uint32_t m = __ballot_sync(0xffffffffu, p); peers &= p ? m : ~m; }
            acc += __popc(peers & lt);
        } else if (MODE == 2) { // smem atomic no return, warp-private 256 bins
            atomicAdd(&my[d], 1u);
        } else if (MODE == 3) { // smem atomic with return
            acc += atomicAdd(&my[d], 1u);
        } else if (MODE == 4) { // smem atomic, conflict-free layout digit*32+lane (block-shared)
            atomicAdd(&s[16 * 256 + d * 32 + lane], 1u);
        } else if (MODE == 5) { // plain LDS random + STS leader
            uint32_t c = my[d]; acc += c; __syncwarp(); if ((x & 3) == 0) my[d] = c + 1; __syncwarp();
        } else if (MODE == 6) { // LDS random only
            acc += my[d];
        } else if (MODE == 7) { // empty loop baseline
            acc += d;
        } else if (MODE == 9) { // ballot match + LDS/STS leader counters (does VOTE overlap the smem pipe?)
            uint32_t peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; ++b) { bool p = (d >> b) & 1; uint32_t m = __ballot_sync(0xffffffffu, p); peers &= p ? m : ~m; }
            uint32_t c = my[d]; uint32_t lower = __popc(peers & lt); acc += c + lower; __syncwarp(); if (lower == 0) my[d] = c + __popc(peers); __syncwarp();
        } else if (MODE == 10) { // 6 shuffles
            uint32_t v = d;
#pragma unroll
            for (int b = 0; b < 6; ++b) v += __shfl_xor_sync(0xffffffffu, v, 1 << (b % 5));
            acc += v;
        } else if (MODE == 11) { // 2 random ATOMS + 2 random LDS (is an ATOMS wavefront 2 pipe cycles?)
            atomicAdd(&my[d], 1u); atomicOr(&my[(d * 7 + 3) & 255], 1u << lane); acc += my[(d + 17) & 255]; acc += my[(d * 5 + 1) & 255];
        } else if (MODE == 12) { // 4 random LDS
            acc += my[d]; acc += my[(d * 7 + 3) & 255]; acc += my[(d + 17) & 255]; acc += my[(d * 5 + 1) & 255];
        } else if (MODE == 13) { // 4 random ATOMS
            atomicAdd(&my[d], 1u); atomicOr(&my[(d * 7 + 3) & 255], 1u << lane); atomicAdd(&my[(d + 17) & 255], 1u); atomicAdd(&my[(d * 5 + 1) & 255], 1u);
        } else if (MODE == 14) { // ballot match + 4 random LDS
            uint32_t peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; ++b) { bool p = (d >> b) & 1; uint32_t m = __ballot_sync(0xffffffffu, p); peers &= p ? m : ~m; }
            acc += __popc(peers & lt);
            acc += my[d]; acc += my[(d * 7 + 3) & 255]; acc += my[(d + 17) & 255]; acc += my[(d * 5 + 1) & 255];
        } else if (MODE == 15) { // 4 random STS
            my[d] = x; my[(d * 7 + 3) & 255] = x; my[(d + 17) & 255] = x; my[(d * 5 + 1) & 255] = x;
        } else if (MODE == 16) { // 4 random STS.64
            uint2* m2 = (uint2*)s + warp * 256;
            m2[d] = make_uint2(x, x); m2[(d * 7 + 3) & 255] = make_uint2(x, x); m2[(d + 17) & 255] = make_uint2(x, x); m2[(d * 5 + 1) & 255] = make_uint2(x, x);
        } else if (MODE == 17) { // 4 random LDS.64
            uint2* m2 = (uint2*)s + warp * 256;
            uint2 a = m2[d], b = m2[(d * 7 + 3) & 255], c = m2[(d + 17) & 255], e = m2[(d * 5 + 1) & 255];
            acc += a.x + a.y + b.x + b.y + c.x + c.y + e.x + e.y;
        } else if (MODE == 18) { // 4 conflict-free STS
            my[lane] = x; my[lane + 32] = x; my[lane + 64] = x; my[lane + 96] = x;
        } else if (MODE == 19) { // 4 conflict-free LDS
            acc += my[lane] + my[lane + 32 + (d & 32)] + my[lane + 64 + (d & 64)] + my[lane + 128 + (d & 96)];
        } else if (MODE == 20) { // 4 random STS over 8192 words (staging-like, block shared)
            uint32_t h = hash(x); s[h & 8191] = x; s[(h >> 8) & 8191] = x; s[(h >> 16) & 8191] = x; s[(h >> 19) & 8191] = x;
        } else if (MODE == 21) { // 4 random LDS over 8192 words
            uint32_t h = hash(x); acc += s[h & 8191] + s[(h >> 8) & 8191] + s[(h >> 16) & 8191] + s[(h >> 19) & 8191];
        } else if (MODE == 30) { // 8 independent ballots, no match logic (VOTE throughput)
            uint32_t v = 0;
#pragma unroll
            for (int b = 0; b < 8; ++b) v ^= __ballot_sync(0xffffffffu, (d >> b) & 1);
            acc += v;
        } else if (MODE == 31) { // counter-return ranking: ATOMS ret + LDS, groups fixed by a ballot loop
            uint32_t old = atomicAdd(&my[d], 1u);
            __syncwarp();
            uint32_t cur = my[d];
            uint32_t rank = old;
            uint32_t multi = __ballot_sync(0xffffffffu, cur != old + 1u);
            while (multi) {
                int L = __ffs(multi) - 1;
                uint32_t dL = __shfl_sync(0xffffffffu, d, L);
                uint32_t grp = __ballot_sync(0xffffffffu, d == dL);
                if (d == dL) rank = cur - __popc(grp) + __popc(grp & lt);
                multi &= ~grp;
            }
            acc += rank;
            __syncwarp();
        } else if (MODE == 32) { // MODE 31 + base lookup + 64-bit staging store
            uint32_t old = atomicAdd(&my[d], 1u);
            __syncwarp();
            uint32_t cur = my[d];
            uint32_t rank = old;
            uint32_t multi = __ballot_sync(0xffffffffu, cur != old + 1u);
            while (multi) {
                int L = __ffs(multi) - 1;
                uint32_t dL = __shfl_sync(0xffffffffu, d, L);
                uint32_t grp = __ballot_sync(0xffffffffu, d == dL);
                if (d == dL) rank = cur - __popc(grp) + __popc(grp & lt);
                multi &= ~grp;
            }
            __syncwarp();
            uint32_t base = s[16 * 256 + warp * 256 + ((d * 5 + 1) & 255)];
            uint2* st = (uint2*)(s + 16 * 256);
            st[(hash(x) + base + rank) & 4095] = make_uint2(x, rank);
        } else if (MODE == 33) { // present kernel's step: ATOMS count + atomicOr mask + LDS.64 + leader STS.64 + 64-bit staging store
            uint2* m2 = (uint2*)s + warp * 256;
            atomicAdd(&m2[(d * 7 + 3) & 255].y, 1u);
            atomicOr(&m2[d].x, 1u << lane);
            __syncwarp();
            uint2 e = m2[d];
            uint32_t lower = __popc(e.x & lt);
            __syncwarp();
            if (lower == 0) m2[d] = make_uint2(0u, e.y + __popc(e.x));
            __syncwarp();
            uint2* st = (uint2*)(s + 16 * 256);
            st[(hash(x) + e.y + lower) & 4095] = make_uint2(x, lower);
        } else if (MODE == 34) { // 8-ballot match + leader ATOMS ret + shfl broadcast (CUB-like) + staging store
            uint32_t peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; ++b) { bool p = (d >> b) & 1; uint32_t m = __ballot_sync(0xffffffffu, p); peers &= p ? m : ~m; }
            uint32_t lower = __popc(peers & lt);
            uint32_t base = 0;
            if (lower == 0) base = atomicAdd(&my[d], __popc(peers));
            base = __shfl_sync(0xffffffffu, base, __ffs(peers) - 1);
            uint2* st = (uint2*)(s + 16 * 256);
            st[(hash(x) + base + lower) & 4095] = make_uint2(x, lower);
        } else if (MODE == 35) { // staging store only (64-bit random into 4096 slots)
            uint2* st = (uint2*)(s + 16 * 256);
            st[hash(x) & 4095] = make_uint2(x, d);
        } else if (MODE == 36) { // counter-return step: ATOMS count + ATOMS.ADD ret position + 64-bit staging store
            atomicAdd(&my[(d * 7 + 3) & 255], 1u);
            uint32_t pos = atomicAdd(&my[d], 1u);
            uint2* st = (uint2*)(s + 16 * 256);
            st[(hash(x) + pos) & 4095] = make_uint2(x, pos);
        } else if (MODE == 8) { // atomicOr match-mask (reference's scheme)
            atomicOr(&my[d], 1u << lane); __syncwarp(); uint32_t m = my[d]; acc += __popc(m & lt); __syncwarp(); if ((m & lt) == 0) my[d] = 0; __syncwarp();
        }
    }
    if (acc == 0x12345678u) out[0] = acc + s[threadIdx.x];
    __syncthreads(); if ((MODE >= 2 && MODE <= 4) || MODE >= 15 || MODE == 8) { if (blockIdx.x == 0) out[2 + (threadIdx.x & 7)] = s[threadIdx.x * 7] + s[threadIdx.x + 4096]; }
}

template <int MODE>
void run(const char* name, int mask, uint32_t* out)
{
    int blocks = 148 * 2;
    k<MODE><<<blocks, 512>>>(out, mask);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    for (int r = 0; r < 10; ++r) k<MODE><<<blocks, 512>>>(out, mask);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
    double warp_ops = (double)blocks * 16 * ITERS;
    // per SM: warp_ops/148 ops in ms; report ns per warp-op per SM and cycles at 1.9 GHz
    double ns = ms * 1e6 / (warp_ops / 148);
    printf("%-38s mask=0x%02x  %.3f ms  %.2f ns/warp-op/SM  (~%.1f cyc @1.9GHz)\n", name, mask, ms, ns, ns * 1.9);
}

int main()
{
    uint32_t* out; cudaMalloc(&out, 64);
    for (int mask : {0xff}) {
        run<7>("baseline loop", mask, out);
        run<0>("match.any hw", mask, out);
        run<1>("ballot-loop match", mask, out);
        run<2>("ATOMS add noret warp-private", mask, out);
        run<3>("ATOMS add ret warp-private", mask, out);
        run<4>("ATOMS add noret digit*32+lane", mask, out);
        run<5>("LDS+STS(leader) counters", mask, out);
        run<6>("LDS random", mask, out);
        run<8>("atomicOr match-mask scheme", mask, out);
        run<15>("4 STS random", mask, out);
        run<16>("4 STS.64 random", mask, out);
        run<17>("4 LDS.64 random", mask, out);
        run<18>("4 STS conflict-free", mask, out);
        run<19>("4 LDS conflict-free", mask, out);
        run<20>("4 STS random 8192 words", mask, out);
        run<21>("4 LDS random 8192 words", mask, out);
        run<9>("ballot match + LDS/STS counters", mask, out);
        run<10>("6 x SHFL.XOR", mask, out);
        run<11>("2 ATOMS + 2 LDS random", mask, out);
        run<12>("4 LDS random", mask, out);
        run<13>("4 ATOMS random", mask, out);
        run<14>("ballot match + 4 LDS random", mask, out);
        run<30>("8 ballots only", mask, out);
        run<31>("counter-return rank + group ballots", mask, out);
        run<32>("counter-return rank + base + stage", mask, out);
        run<33>("present step (2 ATOMS+LDS64+STS64+stage)", mask, out);
        run<34>("8-ballot + leader ATOMS + shfl + stage", mask, out);
        run<35>("stage only (STS.64 random)", mask, out);
        run<36>("counter-return step (2 ATOMS + stage)", mask, out);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
