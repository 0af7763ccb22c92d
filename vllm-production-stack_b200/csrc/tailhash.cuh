// XXH64 (seed 0) of a message of 1..127 bytes held in sixteen 64-bit little-endian words,
// with static register indexing only (no local memory): the ragged last chunk of a prompt.
// Bytes at and beyond `len` are never looked at.  The includer supplies
//   B200R_TH_MUL(C, x)   x * C mod 2^64      B200R_TH_ROTL(R, x)   rotate left by R
// (device: mulc<> / rotl<> of common.cuh; the host unit test: plain C++), and
// B200R_TH_FN, the function qualifiers.
#ifndef B200R_TAILHASH_CUH
#define B200R_TAILHASH_CUH
#include <cstdint>

namespace b200r_th {
constexpr uint64_t P1 = 0x9E3779B185EBCA87ULL, P2 = 0xC2B2AE3D27D4EB4FULL, P3 = 0x165667B19E3779F9ULL,
                   P4 = 0x85EBCA77C2B2AE63ULL, P5 = 0x27D4EB2F165667C5ULL;

B200R_TH_FN uint64_t round64(uint64_t acc, uint64_t in)
{
    return B200R_TH_MUL(P1, B200R_TH_ROTL(31, acc + B200R_TH_MUL(P2, in)));
}
B200R_TH_FN uint64_t merge64(uint64_t h, uint64_t v)
{
    return B200R_TH_MUL(P1, h ^ round64(0, v)) + P4;
}

B200R_TH_FN uint64_t xxh64_tail_words(const uint64_t (&w)[16], int len)
{
    const int s = len >> 5;                       // full 32-byte stripes: 0..3
    uint64_t h;
    if (s > 0) {
        uint64_t v1 = P1 + P2, v2 = P2, v3 = 0, v4 = 0 - P1;
#pragma unroll
        for (int j = 0; j < 3; j++)
            if (j < s) {
                v1 = round64(v1, w[4 * j]);     v2 = round64(v2, w[4 * j + 1]);
                v3 = round64(v3, w[4 * j + 2]); v4 = round64(v4, w[4 * j + 3]);
            }
        h = B200R_TH_ROTL(1, v1) + B200R_TH_ROTL(7, v2) + B200R_TH_ROTL(12, v3) + B200R_TH_ROTL(18, v4);
        h = merge64(h, v1); h = merge64(h, v2); h = merge64(h, v3); h = merge64(h, v4);
    } else {
        h = P5;
    }
    h += (uint64_t)len;
    // the < 32 bytes after the stripes live in words 4s .. 4s+3
    const uint64_t t0 = s == 0 ? w[0] : s == 1 ? w[4] : s == 2 ? w[8] : w[12];
    const uint64_t t1 = s == 0 ? w[1] : s == 1 ? w[5] : s == 2 ? w[9] : w[13];
    const uint64_t t2 = s == 0 ? w[2] : s == 1 ? w[6] : s == 2 ? w[10] : w[14];
    const uint64_t t3 = s == 0 ? w[3] : s == 1 ? w[7] : s == 2 ? w[11] : w[15];
    const int r = len & 31, nw = r >> 3;
    if (nw > 0) { h ^= round64(0, t0); h = B200R_TH_MUL(P1, B200R_TH_ROTL(27, h)) + P4; }
    if (nw > 1) { h ^= round64(0, t1); h = B200R_TH_MUL(P1, B200R_TH_ROTL(27, h)) + P4; }
    if (nw > 2) { h ^= round64(0, t2); h = B200R_TH_MUL(P1, B200R_TH_ROTL(27, h)) + P4; }
    uint64_t u = nw == 0 ? t0 : nw == 1 ? t1 : nw == 2 ? t2 : t3;
    int rb = r & 7;
    if (rb >= 4) {
        h ^= B200R_TH_MUL(P1, u & 0xFFFFFFFFULL);
        h = B200R_TH_MUL(P2, B200R_TH_ROTL(23, h)) + P3;
        u >>= 32; rb -= 4;
    }
#pragma unroll
    for (int k = 0; k < 3; k++)
        if (k < rb) {
            h ^= B200R_TH_MUL(P5, u & 0xFFULL);
            h = B200R_TH_MUL(P1, B200R_TH_ROTL(11, h));
            u >>= 8;
        }
    h ^= h >> 33; h = B200R_TH_MUL(P2, h); h ^= h >> 29; h = B200R_TH_MUL(P3, h); h ^= h >> 32;
    return h;
}
}  // namespace b200r_th
#endif
