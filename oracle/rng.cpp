// ORACLE — TEST INFRASTRUCTURE ONLY.
// glibc rand()/srand() (TYPE_3 additive feedback generator, degree 31,
// separation 3: r[i] = r[i-3] + r[i-31], output r[i] >> 1) and
// Random::randomInt (System/Random.cpp:15-20).  The reference seeds with
// time(NULL) (System/Random.cpp:7-13); the oracle takes the seed explicitly.
// Pinned against this image's glibc 2.39: srand(42) -> 71876166, 708592740,
// 1483128881, ... (tests/test_oracle_rng.py calls libc through ctypes).
#include "oracle.h"

namespace orc {

void GlibcRand::seed(unsigned s)
{
    // ring of 34 words: position i of the linear recurrence lives at i % 34
    uint32_t lin[344];
    if (s == 0)
        s = 1;
    lin[0] = s;
    for (int i = 1; i < 31; i++) {
        int64_t w = (16807LL * (int32_t)lin[i - 1]) % 2147483647;
        if (w < 0)
            w += 2147483647;
        lin[i] = (uint32_t)w;
    }
    for (int i = 31; i < 34; i++)
        lin[i] = lin[i - 31];
    for (int i = 34; i < 344; i++)
        lin[i] = lin[i - 31] + lin[i - 3];
    // keep the last 34 values; next output index is 344
    for (int i = 344 - 34; i < 344; i++)
        r[i % 34] = (int32_t)lin[i];
    f = 344; // linear index of the next value to produce
    b = 0;
}

int GlibcRand::next()
{
    uint32_t v = (uint32_t)r[(f - 31) % 34] + (uint32_t)r[(f - 3) % 34];
    r[f % 34] = (int32_t)v;
    f++;
    if (f >= 34 * 1000000)
        f -= 34 * 999990; // keep the index bounded; 34-periodic positions preserved
    return (int)(v >> 1);
}

int GlibcRand::random_int(int mn, int mx)
{
    int d = mx - mn + 1;
    return int(((double)next() / ((double)2147483647 + 1.0)) * d) + mn;
}

} // namespace orc

extern "C" {
void orc_rng_seed(orc::GlibcRand* g, unsigned s) { g->seed(s); }
int orc_rng_next(orc::GlibcRand* g) { return g->next(); }
int orc_rng_random_int(orc::GlibcRand* g, int mn, int mx) { return g->random_int(mn, mx); }
int orc_rng_sizeof() { return (int)sizeof(orc::GlibcRand); }
}
