#include <cstdio>
#include <cstdlib>
#include <cstring>
#define B200R_TH_FN static inline
#define B200R_TH_MUL(C, x) ((uint64_t)(x) * (uint64_t)(C))
#define B200R_TH_ROTL(R, x) ((((uint64_t)(x)) << (R)) | (((uint64_t)(x)) >> (64 - (R))))
#include "tailhash.cuh"
extern "C" uint64_t orc_xxh64(const uint8_t *p, int64_t len, uint64_t seed);
int main() {
    srand(7); int bad = 0;
    for (int rep = 0; rep < 50; rep++)
        for (int len = 1; len < 128; len++) {
            uint8_t buf[128]; for (int i = 0; i < 128; i++) buf[i] = (uint8_t)rand();
            uint64_t w[16]; memcpy(w, buf, 128);
            if (b200r_th::xxh64_tail_words(w, len) != orc_xxh64(buf, len, 0)) { bad++; if (bad < 5) printf("mismatch len %d\n", len); }
        }
    printf("bad = %d\n", bad); return bad != 0;
}
