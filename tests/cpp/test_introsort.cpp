// Host check: rgbdfe::std_sort (the device-side restatement of libstdc++ std::sort)
// yields the same permutation as std::sort for DMatch arrays with many distance ties.
#include "../../rgbd-slam_b200/csrc/introsort.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
int main()
{
    srand(7);
    long cases = 0, bad = 0;
    for (int trial = 0; trial < 4000; trial++) {
        int n = rand() % 1500;
        int range = 1 + rand() % (trial % 3 == 0 ? 4 : (trial % 3 == 1 ? 60 : 257));
        std::vector<rgbdfe_dmatch> a(n), b;
        for (int i = 0; i < n; i++) {
            a[i].queryIdx = i;
            a[i].trainIdx = rand() % 5000;
            a[i].imgIdx = 0;
            a[i].distance = (float)(trial % 7 == 0 ? (i % range) : (rand() % range));
        }
        if (trial % 11 == 0) std::sort(a.begin(), a.end(), [](const rgbdfe_dmatch& x, const rgbdfe_dmatch& y) { return x.distance > y.distance; });
        b = a;
        std::sort(a.begin(), a.end(), [](const rgbdfe_dmatch& x, const rgbdfe_dmatch& y) { return x.distance < y.distance; });
        rgbdfe::std_sort(b.data(), b.data() + n, rgbdfe::DmLess());
        cases++;
        if (n && memcmp(a.data(), b.data(), n * sizeof(rgbdfe_dmatch)) != 0) bad++;
    }
    // adversarial: organ-pipe / median-of-3 killer-ish patterns to reach the heapsort fallback
    for (int n = 17; n < 3000; n += 97) {
        std::vector<rgbdfe_dmatch> a(n), b;
        for (int i = 0; i < n; i++) {
            a[i].queryIdx = i; a[i].trainIdx = 0; a[i].imgIdx = 0;
            a[i].distance = (float)((i % 2) ? (n - i) : i);
        }
        b = a;
        std::sort(a.begin(), a.end(), [](const rgbdfe_dmatch& x, const rgbdfe_dmatch& y) { return x.distance < y.distance; });
        rgbdfe::std_sort(b.data(), b.data() + n, rgbdfe::DmLess());
        cases++;
        if (memcmp(a.data(), b.data(), n * sizeof(rgbdfe_dmatch)) != 0) bad++;
    }
    printf("introsort cases %ld bad %ld\n", cases, bad);
    return bad != 0;
}
