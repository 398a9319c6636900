/*
 * stdsort_emul.h -- step-for-step emulation of libstdc++'s std::sort (GCC 13
 * bits/stl_algo.h: __introsort_loop + __final_insertion_sort, heap fallback from
 * bits/stl_heap.h) for a strict-weak comparator that is NOT a total order.
 *
 * Why: DistributeOctTree sorts (size, node*) pairs with compareNodes
 * (reference src/ORBextractor.cc:209-224, call at :373).  Elements that tie on
 * size and UL.x are "equivalent", so the order std::sort leaves them in decides
 * which quadtree nodes are split before the `break` at :419-420 -- i.e. it is part
 * of the reference's keypoint selection.  To reproduce that selection exactly on the
 * device the sort has to make the same comparisons and swaps as libstdc++.
 *
 * Usable from host and device code.  Elements are opaque `T`, compared by `less(a,b)`.
 */
#ifndef ORBX_STDSORT_EMUL_H
#define ORBX_STDSORT_EMUL_H

#if defined(__CUDACC__)
#define ORBX_HD __host__ __device__ __forceinline__
#else
#define ORBX_HD inline
#endif

namespace orbx_stdsort {

template <typename T> ORBX_HD void swap_(T& a, T& b) { T t = a; a = b; b = t; }

/* ---- heap pieces (bits/stl_heap.h) ---- */
template <typename T, typename Less>
ORBX_HD void push_heap_(T* first, int holeIndex, int topIndex, T value, Less less) {
    int parent = (holeIndex - 1) / 2;
    while (holeIndex > topIndex && less(first[parent], value)) {
        first[holeIndex] = first[parent];
        holeIndex = parent;
        parent = (holeIndex - 1) / 2;
    }
    first[holeIndex] = value;
}

template <typename T, typename Less>
ORBX_HD void adjust_heap_(T* first, int holeIndex, int len, T value, Less less) {
    const int topIndex = holeIndex;
    int secondChild = holeIndex;
    while (secondChild < (len - 1) / 2) {
        secondChild = 2 * (secondChild + 1);
        if (less(first[secondChild], first[secondChild - 1])) secondChild--;
        first[holeIndex] = first[secondChild];
        holeIndex = secondChild;
    }
    if ((len & 1) == 0 && secondChild == (len - 2) / 2) {
        secondChild = 2 * (secondChild + 1);
        first[holeIndex] = first[secondChild - 1];
        holeIndex = secondChild - 1;
    }
    push_heap_(first, holeIndex, topIndex, value, less);
}

template <typename T, typename Less>
ORBX_HD void make_heap_(T* first, int len, Less less) {
    if (len < 2) return;
    int parent = (len - 2) / 2;
    while (true) {
        T value = first[parent];
        adjust_heap_(first, parent, len, value, less);
        if (parent == 0) return;
        parent--;
    }
}

/* std::__partial_sort(first, last, last): __heap_select degenerates to make_heap, then __sort_heap */
template <typename T, typename Less>
ORBX_HD void heap_sort_(T* first, int len, Less less) {
    make_heap_(first, len, less);
    int last = len;
    while (last > 1) {
        --last;
        T value = first[last];          /* __pop_heap(first, last, last) */
        first[last] = first[0];
        adjust_heap_(first, 0, last, value, less);
    }
}

template <typename T, typename Less>
ORBX_HD void move_median_to_first_(T* a, int result, int ia, int ib, int ic, Less less) {
    if (less(a[ia], a[ib])) {
        if (less(a[ib], a[ic])) swap_(a[result], a[ib]);
        else if (less(a[ia], a[ic])) swap_(a[result], a[ic]);
        else swap_(a[result], a[ia]);
    } else if (less(a[ia], a[ic])) swap_(a[result], a[ia]);
    else if (less(a[ib], a[ic])) swap_(a[result], a[ic]);
    else swap_(a[result], a[ib]);
}

template <typename T, typename Less>
ORBX_HD int unguarded_partition_(T* a, int first, int last, int pivot, Less less) {
    while (true) {
        while (less(a[first], a[pivot])) ++first;
        --last;
        while (less(a[pivot], a[last])) --last;
        if (!(first < last)) return first;
        swap_(a[first], a[last]);
        ++first;
    }
}

template <typename T, typename Less>
ORBX_HD void unguarded_linear_insert_(T* a, int last, Less less) {
    T val = a[last];
    int next = last - 1;
    while (less(val, a[next])) {
        a[last] = a[next];
        last = next;
        --next;
    }
    a[last] = val;
}

template <typename T, typename Less>
ORBX_HD void insertion_sort_(T* a, int first, int last, Less less) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (less(a[i], a[first])) {
            T val = a[i];
            for (int k = i; k > first; --k) a[k] = a[k - 1];   /* move_backward */
            a[first] = val;
        } else {
            unguarded_linear_insert_(a, i, less);
        }
    }
}

/* std::sort(a, a+n, less).  Recursion of __introsort_loop is replaced by an explicit
 * stack of (first,last,depth) -- the recursive call handles [cut,last) FIRST and the
 * loop continues with [first,cut), which the stack order reproduces (the two ranges are
 * disjoint, so the order of processing does not change any comparison). */
template <typename T, typename Less>
ORBX_HD void sort(T* a, int n, Less less) {
    if (n <= 1) return;
    int lg = 0;
    for (int t = n; t > 1; t >>= 1) ++lg;            /* std::__lg(n) */
    const int kThreshold = 16;
    int stFirst[64], stLast[64], stDepth[64];
    int sp = 0;
    stFirst[0] = 0; stLast[0] = n; stDepth[0] = 2 * lg; sp = 1;
    while (sp > 0) {
        --sp;
        int first = stFirst[sp], last = stLast[sp], depth = stDepth[sp];
        while (last - first > kThreshold) {
            if (depth == 0) {
                heap_sort_(a + first, last - first, less);
                break;
            }
            --depth;
            const int mid = first + (last - first) / 2;
            move_median_to_first_(a, first, first + 1, mid, last - 1, less);
            const int cut = unguarded_partition_(a, first + 1, last, first, less);
            /* recurse on [cut,last) */
            stFirst[sp] = cut; stLast[sp] = last; stDepth[sp] = depth; ++sp;
            last = cut;
        }
    }
    if (n > kThreshold) {
        insertion_sort_(a, 0, kThreshold, less);
        for (int i = kThreshold; i != n; ++i) unguarded_linear_insert_(a, i, less);
    } else {
        insertion_sort_(a, 0, n, less);
    }
}

} // namespace orbx_stdsort
#endif
