// Floor of the claim kernel's access pattern: one thread per item, a coalesced 32-byte
// record load followed by one dependent 32-byte load (and optionally a 128-bit CAS) at a
// pseudo-random table entry.  Prints microseconds per launch of 1.38 M items.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
struct __align__(64) Entry { uint64_t k1, k2, mask, first, h, r[3]; };
struct __align__(32) Rec { uint64_t k1, k2, h, req; };
__device__ __forceinline__ uint64_t mix(uint64_t x) { x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL; x ^= x >> 27; x *= 0x94D049BB133111EBULL; return x ^ (x >> 31); }
__global__ void fill(Rec *r, int64_t n, uint64_t salt) { int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; if (g < n) { r[g].k1 = mix(g + salt); r[g].k2 = mix(g * 3 + salt); r[g].h = g; r[g].req = g / 21; } }
// MODE 0: rec only; 1: rec + entry load; 2: rec + entry load + CAS when empty; 3: rec + CAS first
template <int MODE, int HOTPCT, int STRIDE = 64, bool STH = true, bool DORED = true>
__global__ void __launch_bounds__(256) probe(Entry *tab, int cap_log2, const Rec *recs, int64_t n, uint32_t *out)
{
    int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (g >= n) return;
    uint64_t k1, k2, h, req;
    asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(k1), "=l"(k2), "=l"(h), "=l"(req) : "l"(recs + g) : "memory");
    uint32_t s = (uint32_t)(k1 >> (64 - cap_log2));
    if (HOTPCT && (g % 21) < (21 * HOTPCT) / 100) s = (uint32_t)(req & 63) * 16 + (uint32_t)(g % 21);   // hot: 1024 entries
    uint32_t res = s;
    if (MODE >= 1) {
        Entry *e = (Entry *)((char *)tab + (size_t)s * STRIDE);
        uint64_t a = 0, b = 0, m = 0, f = 0;
        if (MODE != 3) asm volatile("ld.global.ca.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(m), "=l"(f) : "l"(e) : "memory");
        if (MODE >= 2 && a == 0 && b == 0) {
            asm volatile("{\n\t.reg .b128 c, nw, old;\n\tmov.b128 c, {%2, %3};\n\tmov.b128 nw, {%4, %5};\n\t"
                         "atom.global.relaxed.gpu.cas.b128 old, [%6], c, nw;\n\tmov.b128 {%0, %1}, old;\n\t}\n"
                         : "=l"(a), "=l"(b) : "l"(0ULL), "l"(0ULL), "l"(k1 | 1), "l"(k2 | 1), "l"(e) : "memory");
            if (a == 0 && b == 0) { if (STH) e->h = h; if (DORED) atomicMax((unsigned long long *)&e->first, (unsigned long long)req); }
        }
        res ^= (uint32_t)(a + b + m + f);
    }
    out[g] = res;
}
template <int MODE, int HOT, int STRIDE = 64, bool STH = true, bool DORED = true>
static void run(const char *name, Entry *tab, int cap_log2, Rec *recs, int64_t n, uint32_t *out, bool clear)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9, sum = 0; const int reps = 6;
    for (int r = 0; r < reps; r++) {
        if (clear) { cudaMemset(tab, 0, sizeof(Entry) << cap_log2); cudaMemset(tab, 1, sizeof(Entry) * 1024); }   // hot entries exist
        fill<<<(n + 255) / 256, 256>>>(recs, n, 1000003ULL * (r + 1));
        cudaDeviceSynchronize();
        cudaEventRecord(a);
        probe<MODE, HOT, STRIDE, STH, DORED><<<(n + 255) / 256, 256>>>(tab, cap_log2, recs, n, out);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (r) { sum += ms; if (ms < best) best = ms; }
    }
    printf("%-58s %7.1f us avg  %7.1f us best  (%s)\n", name, sum / (reps - 1) * 1e3, best * 1e3, cudaGetErrorString(cudaGetLastError()));
}
int main()
{
    const int cap_log2 = 25; const int64_t n = 1380000;
    Entry *tab; Rec *recs; uint32_t *out;
    cudaMalloc(&tab, sizeof(Entry) << cap_log2); cudaMalloc(&recs, n * sizeof(Rec)); cudaMalloc(&out, n * 4);
    cudaMemset(tab, 0, sizeof(Entry) << cap_log2);
    run<0, 0>("rec load + store only", tab, cap_log2, recs, n, out, false);
    run<1, 0>("rec + random entry load (all DRAM)", tab, cap_log2, recs, n, out, true);
    run<1, 75>("rec + entry load, 75% hot (1024 entries)", tab, cap_log2, recs, n, out, true);
    run<1, 100>("rec + entry load, 100% hot", tab, cap_log2, recs, n, out, true);
    run<2, 0>("rec + load + CAS + st h + RED, all new", tab, cap_log2, recs, n, out, true);
    run<2, 75>("rec + load + CAS + st h + RED, 75% hot (like C3)", tab, cap_log2, recs, n, out, true);
    run<2, 0, 64, false, false>("load + CAS, all new", tab, cap_log2, recs, n, out, true);
    run<2, 0, 64, false, true>("load + CAS + RED, all new", tab, cap_log2, recs, n, out, true);
    run<2, 0, 64, true, false>("load + CAS + st h, all new", tab, cap_log2, recs, n, out, true);
    run<2, 0, 32, false, true>("32-byte entries: load + CAS + RED, all new", tab, cap_log2, recs, n, out, true);
    run<2, 75, 32, false, true>("32-byte entries: load + CAS + RED, 75% hot", tab, cap_log2, recs, n, out, true);
    run<2, 75, 64, false, true>("64-byte entries: load + CAS + RED (no st h), 75% hot", tab, cap_log2, recs, n, out, true);
    run<1, 0, 32>("32-byte entries: random entry load (all DRAM)", tab, cap_log2, recs, n, out, true);
    return 0;
}
