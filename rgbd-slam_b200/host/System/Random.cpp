// Random over the context's rand() stream (reference: System/Random.cpp:5-20).
#include "Random.h"

#include "../rgbdfe_host.h"

#include <ctime>

bool Random::SET_RAND = false;

void Random::initSeed()
{
    // srand(time(NULL)) once per process (System/Random.cpp:7-13)
    if (!SET_RAND) {
        SET_RAND = true;
        if (!rgbdfe_host::seeded()) // a caller (a test) may have pinned the stream already
            rgbdfe_host::seed((unsigned int)time(NULL));
    }
}

int Random::nextRand()
{
    int v = 0;
    rgbdfe_host::report("rgbdfe_rng_next", rgbdfe_rng_next(rgbdfe_host::ctx(), &v));
    return v;
}

int Random::randomInt(int min, int max)
{
    initSeed();
    int d = max - min + 1;
    return int(((double)nextRand() / ((double)2147483647 + 1.0)) * d) + min;
}
