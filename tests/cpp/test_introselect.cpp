// Host check: rgbdfe::std_nth_element / std_partition (the device-side restatement of libstdc++'s introselect and
// bidirectional partition behind cv::KeyPointsFilter::retainBest) leave arrays in the same order as the library.
#include "../../rgbd-slam_b200/csrc/introselect.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
struct E {
    float r;
    int id;
};
struct Greater {
    bool operator()(const E& a, const E& b) const { return a.r > b.r; }
};
static long run(std::vector<E> a, int n_points, long& bad)
{
    std::vector<E> b = a;
    const int n = (int)a.size();
    std::nth_element(a.begin(), a.begin() + n_points - 1, a.end(), Greater());
    const float amb = a[n_points - 1].r;
    auto e1 = std::partition(a.begin() + n_points, a.end(), [amb](const E& k) { return k.r >= amb; });
    rgbdfe::std_nth_element(b.data(), n_points - 1, n, Greater());
    const float amb2 = b[n_points - 1].r;
    const int e2 = rgbdfe::std_partition(b.data(), n_points, n, [amb2](const E& k) { return k.r >= amb2; });
    if (e2 != (int)(e1 - a.begin()) || memcmp(a.data(), b.data(), (size_t)n * sizeof(E)) != 0)
        bad++;
    return 1;
}
int main()
{
    srand(11);
    long cases = 0, bad = 0;
    for (int trial = 0; trial < 4000; trial++) {
        const int n = 2 + rand() % 6000;
        const int n_points = 1 + rand() % (n - 1);
        const int range = 1 + rand() % (trial % 3 == 0 ? 4 : (trial % 3 == 1 ? 235 : 100000));
        std::vector<E> a(n);
        for (int i = 0; i < n; i++)
            a[i] = E { (float)(trial % 7 == 0 ? (i % range) : (rand() % range)), i };
        if (trial % 11 == 0)
            std::sort(a.begin(), a.end(), [](const E& x, const E& y) { return x.r < y.r; });
        cases += run(a, n_points, bad);
    }
    // patterns that exhaust the depth limit (heap_select fallback)
    for (int n = 50; n < 6000; n += 61) {
        std::vector<E> a(n);
        for (int i = 0; i < n; i++)
            a[i] = E { (float)((i % 2) ? (n - i) : i), i };
        cases += run(a, n / 2, bad);
        for (int i = 0; i < n; i++)
            a[i] = E { (float)(i < n / 2 ? 2 * i : 2 * (n - i) + 1), i };
        cases += run(a, n / 3 + 1, bad);
    }
    printf("cases %ld bad %ld\n", cases, bad);
    return bad != 0;
}
