// glibc rand()/srand() TYPE_3 stream on the device (System/Random.cpp:7-20 draws from it).
//
// rand() is the additive lagged generator r[i] = r[i-3] + r[i-31] (mod 2^32), output r[i] >> 1.
// RngState keeps the last 34 words as a ring indexed by the linear output index.  The kernels
// never step it one call at a time: they unroll the ring into a linear window in shared memory
// (`lin[0..34)` = the 34 most recent words, oldest first), extend the window by a block of
// outputs with one thread (one LDS + IADD + STS per output, the three interleaved chains keep
// their newest word in registers), and store the window back at whatever position the
// sequential reference would have reached.
#pragma once
#include "common.cuh"

namespace rgbdfe {

#define RNG_HIST 34

// lin[0..34) <- the ring, oldest word first.  Call with one thread (or lanes 0..33 with lane = t).
__device__ __forceinline__ void rng_unroll(const RngState& s, unsigned* lin)
{
    const int base = s.idx % RNG_HIST; // word written 34 steps ago sits at idx % 34
    for (int t = 0; t < RNG_HIST; t++)
        lin[t] = (unsigned)s.r[(base + t) % RNG_HIST];
}

// Extend lin[34 .. 34+n) from lin[0..34).  One thread.
__device__ __forceinline__ void rng_extend(unsigned* lin, int n)
{
    unsigned a = lin[RNG_HIST - 3], b = lin[RNG_HIST - 2], c = lin[RNG_HIST - 1];
    int i = RNG_HIST;
    const int end = RNG_HIST + n;
    for (; i + 3 <= end; i += 3) {
        a += lin[i - 31];
        b += lin[i - 30];
        c += lin[i - 29];
        lin[i] = a;
        lin[i + 1] = b;
        lin[i + 2] = c;
    }
    for (; i < end; i++)
        lin[i] = lin[i - 3] + lin[i - 31];
}

// The state after `consumed` outputs of the window: ring <- lin[consumed .. consumed+34).
__device__ __forceinline__ void rng_roll(RngState& s, const unsigned* lin, int consumed)
{
    int idx = s.idx + consumed;
    while (idx >= RNG_HIST * 1000000)
        idx -= RNG_HIST * 999990; // keeps idx % 34
    const int base = idx % RNG_HIST;
    for (int t = 0; t < RNG_HIST; t++)
        s.r[(base + t) % RNG_HIST] = (int)lin[consumed + t];
    s.idx = idx;
}

// rand() value of window word i, and Random::randomInt(0, d-1) of it (System/Random.cpp:15-20):
// int(rand() / (RAND_MAX + 1.0) * d); the division by 2^31 is exact, so it is a multiply here.
__device__ __forceinline__ int rng_value(unsigned w) { return (int)(w >> 1); }
__device__ __forceinline__ int rng_random_int0(unsigned w, int d) { return int(((double)rng_value(w) * (1.0 / 2147483648.0)) * (double)d); }

// srandom_r(seed) for TYPE_3: 31-word LCG fill, then 310 discarded outputs.  One thread.
__device__ inline void rng_seed_state(RngState& s, unsigned seed, unsigned* lin /* >= 34 + 310 words of scratch */)
{
    if (seed == 0)
        seed = 1;
    lin[0] = seed;
    for (int i = 1; i < 31; i++) {
        long long w = (16807LL * (int)lin[i - 1]) % 2147483647;
        if (w < 0)
            w += 2147483647;
        lin[i] = (unsigned)w;
    }
    for (int i = 31; i < 34; i++)
        lin[i] = lin[i - 31];
    rng_extend(lin, 310);
    for (int t = 0; t < RNG_HIST; t++)
        s.r[(344 + t) % RNG_HIST] = (int)lin[310 + t];
    s.idx = 344;
}

// Window <-> ring helpers on a stream that lives in global memory (one thread).
__device__ __forceinline__ int rng_window_load(const RngState* st, unsigned* lin)
{
    const int idx = st->idx, base = idx % RNG_HIST;
    for (int t = 0; t < RNG_HIST; t++)
        lin[t] = (unsigned)st->r[(base + t) % RNG_HIST];
    return idx;
}
__device__ __forceinline__ void rng_window_store(RngState* st, const unsigned* lin, int idx)
{
    while (idx >= RNG_HIST * 1000000)
        idx -= RNG_HIST * 999990; // glibc never wraps; this only keeps idx % 34 and the int range
    const int base = idx % RNG_HIST;
    for (int t = 0; t < RNG_HIST; t++)
        st->r[(base + t) % RNG_HIST] = (int)lin[t];
    st->idx = idx;
}

// Advance a stream by `n` rand() calls without using their values; one thread.
// lin: 34 + block words of scratch.
__device__ inline void rng_advance(unsigned* lin, RngState* st, int n, int block)
{
    const int idx0 = rng_window_load(st, lin);
    int left = n;
    while (left > 0) {
        const int nb = min(left, block);
        rng_extend(lin, nb);
        for (int t = 0; t < RNG_HIST; t++)
            lin[t] = lin[nb + t];
        left -= nb;
    }
    rng_window_store(st, lin, idx0 + n);
}

} // namespace rgbdfe
