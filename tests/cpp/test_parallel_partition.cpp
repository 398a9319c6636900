// Checks the closed form the device octree sort uses for libstdc++'s __unguarded_partition
// (extractor.cu: oct_introsort_warp) and "final insertion sort == stable sort" against std::sort.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../orb_slam3_v1.0_b200/csrc/stdsort_emul.h"
typedef unsigned long long E;
struct Less { bool operator()(const E& a, const E& b) const { return (a >> 16) < (b >> 16); } };
static int partition_closed_form(std::vector<E>& a, int first, int last) {   // range (first,last), pivot a[first]
    Less less; const E pv = a[first];
    std::vector<int> L, R;
    for (int p = first + 1; p < last; ++p) { if (!less(a[p], pv)) L.push_back(p); if (!less(pv, a[p])) R.push_back(p); }
    const int totL = (int)L.size(), totR = (int)R.size(), kmax = std::min(totL, totR);
    int K = 0;
    while (K < kmax && L[K] < R[totR - 1 - K]) ++K;
    const int cut = std::min(K < totL ? L[K] : last, K > 0 ? R[totR - K] : last);
    for (int k = 0; k < K; ++k) std::swap(a[L[k]], a[R[totR - 1 - k]]);
    return cut;
}
static void sort_device_style(std::vector<E>& a) {
    const int n = (int)a.size(); Less less;
    if (n > 16) {
        int lg = 0; for (int t = n; t > 1; t >>= 1) ++lg;
        std::vector<int> F{0}, Lst{n}, D{2 * lg};
        while (!F.empty()) {
            int first = F.back(), last = Lst.back(), depth = D.back(); F.pop_back(); Lst.pop_back(); D.pop_back();
            while (last - first > 16) {
                if (depth == 0) { orbx_stdsort::heap_sort_(a.data() + first, last - first, less); break; }
                --depth;
                orbx_stdsort::move_median_to_first_(a.data(), first, first + 1, first + (last - first) / 2, last - 1, less);
                const int cut = partition_closed_form(a, first, last);
                F.push_back(cut); Lst.push_back(last); D.push_back(depth);
                last = cut;
            }
        }
    }
    std::vector<E> out(n);
    for (int i = 0; i < n; ++i) {
        int r = 0; const E ci = (a[i] & ~0xFFFFull) | (E)i;
        for (int j = 0; j < n; ++j) r += ((a[j] & ~0xFFFFull) | (E)j) < ci;
        out[r] = a[i];
    }
    a = out;
}
int main(int argc, char** argv) {
    const int iters = argc > 1 ? atoi(argv[1]) : 20000;
    srand(7); long cases = 0;
    for (int it = 0; it < iters; ++it) {
        const int n = 1 + rand() % (it % 50 == 0 ? 3000 : 400);
        const int ksz = 1 + rand() % 6, kx = 1 + rand() % 8;       // few distinct keys -> many ties
        std::vector<E> a(n);
        const int mode = rand() % 4;
        for (int i = 0; i < n; ++i) {
            unsigned size = 2 + rand() % ksz, ulx = (rand() % kx) * 37;
            if (mode == 1) size = 2 + i % ksz; if (mode == 2) size = 2 + (n - i) / 8; if (mode == 3) size = 2 + rand() % 3000;
            a[i] = ((E)size << 32) | ((E)ulx << 16) | (E)i;
        }
        std::vector<E> b = a, c = a;
        std::sort(b.begin(), b.end(), Less());
        sort_device_style(c);
        if (b != c) { printf("MISMATCH n=%d it=%d\n", n, it); return 1; }
        ++cases;
    }
    // organ-pipe / adversarial inputs that reach the heapsort fallback
    for (int n = 17; n < 600; ++n) {
        std::vector<E> a(n);
        for (int i = 0; i < n; ++i) { unsigned v = i < n / 2 ? i : n - i; a[i] = ((E)(2 + v % 5) << 32) | ((E)(v % 3) << 16) | (E)i; }
        std::vector<E> b = a, c = a; std::sort(b.begin(), b.end(), Less()); sort_device_style(c);
        if (b != c) { printf("MISMATCH organ n=%d\n", n); return 1; }
        ++cases;
    }
    printf("OK %ld cases\n", cases); return 0;
}
