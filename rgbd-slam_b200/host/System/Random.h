// Random — same interface as the reference's System/Random.h:6-50.  The reference draws from
// the process-wide glibc rand(); here the stream is the glibc-compatible generator owned by the
// calling thread's librgbdfe context, so that RansacSE3 / Ransac (which sample on the device)
// and host callers of Random consume ONE stream in the reference's order.
#ifndef RANDOM_H
#define RANDOM_H

#include <cmath>
#include <random>

class Random {
public:
    static void initSeed();

    // Returns a random int in the range [min..max]
    static int randomInt(int min, int max);

    // rand() of the context stream (RAND_MAX = 2147483647 as in glibc)
    static int nextRand();

    // Returns a random number in the range [0..1]
    template <class T>
    static T randomValue()
    {
        initSeed();
        return (T)nextRand() / (T)2147483647;
    }

    // Returns a random number in the range [min..max]
    template <class T>
    static T randomValue(T min, T max)
    {
        initSeed();
        return randomValue<T>() * (max - min) + min;
    }

    // Returns a random number from a guassian distribution
    template <class T>
    static T randomGaussianValue(T mean = 0., T sigma = 1.)
    {
        initSeed();
        T x1, x2, w, y1;

        do {
            x1 = (T)2. * randomValue<T>() - (T)1.;
            x2 = (T)2. * randomValue<T>() - (T)1.;
            w = x1 * x1 + x2 * x2;
        } while (w >= (T)1. || w == (T)0.);

        w = sqrt(((T)-2.0 * log(w)) / w);
        y1 = x1 * w;

        return (mean + y1 * sigma);
    }

protected:
    static bool SET_RAND;
};

#endif // RANDOM_H
