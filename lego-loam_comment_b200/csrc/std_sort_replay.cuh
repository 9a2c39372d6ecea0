// Exact replay of the permutation libstdc++'s std::sort produces (GCC 13, bits/stl_algo.h introsort).
//
// Why: extractFeatures sorts each sector with std::sort and a comparator that looks only at the curvature value
// (reference utility.h:139-148, featureAssociation.cpp:699).  std::sort is not stable and equal curvatures are
// common (SURVEY.md hard part 4 / quirk Q17: 14 % of sectors), and the order inside a run of equal keys decides
// which points become features.  Bit-exact feature indices therefore need the same permutation, which depends on
// the algorithm's exact sequence of comparisons and moves.  This header restates that published algorithm
// (introsort loop with median-of-three to *first*, unguarded Hoare partition, heap-sort fallback when the depth
// budget 2*floor(log2 n) is exhausted, final guarded-16 + unguarded insertion sort) for a single thread working
// on an array in shared or local memory.  tests/native/std_sort_replay_host.cpp checks it against the real
// std::sort on tie-heavy, presorted and adversarial inputs.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define SSR_HD __host__ __device__ __forceinline__
#else
#define SSR_HD static inline
#endif

namespace ssr {

struct Item { float value; int ind; };

SSR_HD bool less(const Item& a, const Item& b) { return a.value < b.value; }
SSR_HD void swap_items(Item& a, Item& b) { Item t = a; a = b; b = t; }

SSR_HD void push_heap(Item* first, int hole, int top, Item value) {
    int parent = (hole - 1) / 2;
    while (hole > top && less(first[parent], value)) { first[hole] = first[parent]; hole = parent; parent = (hole - 1) / 2; }
    first[hole] = value;
}

SSR_HD void adjust_heap(Item* first, int hole, int len, Item value) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (less(first[child], first[child - 1])) --child;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    push_heap(first, hole, top, value);
}

// partial_sort(first, last, last): make_heap then sort_heap (the heap_select scan over [middle,last) is empty).
SSR_HD void heap_sort(Item* first, int len) {
    if (len >= 2) {
        int parent = (len - 2) / 2;
        while (true) {
            Item v = first[parent];
            adjust_heap(first, parent, len, v);
            if (parent == 0) break;
            --parent;
        }
    }
    int last = len;
    while (last > 1) {
        --last;
        Item v = first[last];
        first[last] = first[0];
        adjust_heap(first, 0, last, v);
    }
}

SSR_HD void unguarded_linear_insert(Item* a, int last) {
    Item v = a[last];
    int next = last - 1;
    while (less(v, a[next])) { a[last] = a[next]; last = next; --next; }
    a[last] = v;
}

SSR_HD void insertion_sort(Item* a, int first, int last) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (less(a[i], a[first])) {
            Item v = a[i];
            for (int k = i; k > first; --k) a[k] = a[k - 1];
            a[first] = v;
        } else {
            unguarded_linear_insert(a, i);
        }
    }
}

SSR_HD int floor_log2(int n) { int r = 0; while (n > 1) { n >>= 1; ++r; } return r; }

// Sorts a[0..n) ascending by value with exactly std::sort's permutation.
SSR_HD void std_sort(Item* a, int n) {
    if (n <= 0) return;
    // The introsort loop recurses on the right part and iterates on the left; an explicit stack of
    // (first, last, depth) replaces the recursion.  Depth budget <= 2*log2(n) bounds the stack.
    int stk_first[64], stk_last[64], stk_depth[64];
    int sp = 0;
    stk_first[0] = 0; stk_last[0] = n; stk_depth[0] = 2 * floor_log2(n); sp = 1;
    while (sp > 0) {
        --sp;
        int first = stk_first[sp], last = stk_last[sp], depth = stk_depth[sp];
        // pending right-hand parts must be processed before the continuation of the left iteration to keep the
        // original order of operations; ranges are disjoint, so the order does not change the result, but we keep
        // it anyway: process right part first by pushing the left continuation below it.
        while (last - first > 16) {
            if (depth == 0) { heap_sort(a + first, last - first); break; }
            --depth;
            // median of (first+1, mid, last-1) moved to first
            const int mid = first + (last - first) / 2;
            const int ia = first + 1, ib = mid, ic = last - 1;
            if (less(a[ia], a[ib])) {
                if (less(a[ib], a[ic])) swap_items(a[first], a[ib]);
                else if (less(a[ia], a[ic])) swap_items(a[first], a[ic]);
                else swap_items(a[first], a[ia]);
            } else if (less(a[ia], a[ic])) swap_items(a[first], a[ia]);
            else if (less(a[ib], a[ic])) swap_items(a[first], a[ic]);
            else swap_items(a[first], a[ib]);
            // unguarded partition of [first+1, last) around a[first]
            int lo = first + 1, hi = last;
            while (true) {
                while (less(a[lo], a[first])) ++lo;
                --hi;
                while (less(a[first], a[hi])) --hi;
                if (!(lo < hi)) break;
                swap_items(a[lo], a[hi]);
                ++lo;
            }
            const int cut = lo;
            // recurse on [cut, last) with the decremented depth, continue with [first, cut)
            stk_first[sp] = cut; stk_last[sp] = last; stk_depth[sp] = depth; ++sp;
            last = cut;
        }
    }
    if (n > 16) {
        insertion_sort(a, 0, 16);
        for (int i = 16; i < n; ++i) unguarded_linear_insert(a, i);
    } else {
        insertion_sort(a, 0, n);
    }
}

}  // namespace ssr
