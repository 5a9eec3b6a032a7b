// introsort_emul.h — step-exact emulation of libstdc++'s std::sort on a small array of 32-bit keys.
//
// Why: the reference orders the octree's size-sorted expansion with
//     sort(vPrevSizeAndPointerToNode.begin(), vPrevSizeAndPointerToNode.end(), compareNodes)
// (/root/reference/MCSlam/src/ORBextractor.cpp:700, comparator :537-552).  compareNodes orders by (size, UL.x) and
// leaves elements that tie on both keys in whatever order libstdc++'s introsort happens to produce.  That order decides
// which node is split first, hence the keypoint ORDER and, at the `>= N` break (:745), the keypoint SET.  Bit-exact
// parity therefore needs the same permutation, so this file restates the published libstdc++ algorithm (bits/stl_algo.h,
// bits/stl_heap.h: introsort_loop with 2*floor(log2 n) depth limit, median-of-three to first, unguarded partition,
// threshold 16, final insertion sort, heap-sort fallback) for use inside a CUDA kernel.  tests/cpp/test_introsort.cpp
// checks it against the real std::sort on tie-heavy inputs.
//
// Elements are `uint32_t key` + `uint16_t payload` kept in two parallel arrays; comparison is on `key` only (the caller packs
// (size << 16 | ULx) so that key order == compareNodes order and equal keys == compareNodes ties).
#ifndef MCS_INTROSORT_EMUL_H
#define MCS_INTROSORT_EMUL_H

#include <stdint.h>

#ifdef __CUDACC__
#define MCS_HD __host__ __device__ __forceinline__
#else
#define MCS_HD inline
#endif

namespace mcs_sort {

struct Arr {
    uint32_t* k;
    uint16_t* p;
};

MCS_HD void swap_at(Arr a, int i, int j) {
    uint32_t tk = a.k[i];
    a.k[i] = a.k[j];
    a.k[j] = tk;
    uint16_t tp = a.p[i];
    a.p[i] = a.p[j];
    a.p[j] = tp;
}

// std::__unguarded_linear_insert
MCS_HD void unguarded_linear_insert(Arr a, int last) {
    uint32_t vk = a.k[last];
    uint16_t vp = a.p[last];
    int next = last - 1;
    while (vk < a.k[next]) {
        a.k[last] = a.k[next];
        a.p[last] = a.p[next];
        last = next;
        --next;
    }
    a.k[last] = vk;
    a.p[last] = vp;
}

// std::__insertion_sort on [first, last)
MCS_HD void insertion_sort(Arr a, int first, int last) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (a.k[i] < a.k[first]) {
            uint32_t vk = a.k[i];
            uint16_t vp = a.p[i];
            for (int j = i; j > first; --j) {  // move_backward(first, i, i + 1)
                a.k[j] = a.k[j - 1];
                a.p[j] = a.p[j - 1];
            }
            a.k[first] = vk;
            a.p[first] = vp;
        } else {
            unguarded_linear_insert(a, i);
        }
    }
}

// std::__adjust_heap + std::__push_heap on the heap rooted at `first`
MCS_HD void adjust_heap(Arr a, int first, int hole, int len, uint32_t vk, uint16_t vp) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (a.k[first + child] < a.k[first + child - 1]) child--;
        a.k[first + hole] = a.k[first + child];
        a.p[first + hole] = a.p[first + child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        a.k[first + hole] = a.k[first + child - 1];
        a.p[first + hole] = a.p[first + child - 1];
        hole = child - 1;
    }
    int parent = (hole - 1) / 2;
    while (hole > top && a.k[first + parent] < vk) {
        a.k[first + hole] = a.k[first + parent];
        a.p[first + hole] = a.p[first + parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    a.k[first + hole] = vk;
    a.p[first + hole] = vp;
}

// std::__partial_sort(first, last, last) = make_heap + sort_heap (heap_select's second loop is empty when middle == last)
MCS_HD void heap_sort(Arr a, int first, int last) {
    const int len = last - first;
    if (len >= 2) {
        int parent = (len - 2) / 2;
        while (true) {
            adjust_heap(a, first, parent, len, a.k[first + parent], a.p[first + parent]);
            if (parent == 0) break;
            parent--;
        }
    }
    int l = last;
    while (l - first > 1) {
        --l;
        uint32_t vk = a.k[l];
        uint16_t vp = a.p[l];
        a.k[l] = a.k[first];
        a.p[l] = a.p[first];
        adjust_heap(a, first, 0, l - first, vk, vp);
    }
}

// std::__move_median_to_first(result, a, b, c)
MCS_HD void move_median_to_first(Arr arr, int result, int a, int b, int c) {
    const uint32_t ka = arr.k[a], kb = arr.k[b], kc = arr.k[c];
    if (ka < kb) {
        if (kb < kc) swap_at(arr, result, b);
        else if (ka < kc) swap_at(arr, result, c);
        else swap_at(arr, result, a);
    } else if (ka < kc) swap_at(arr, result, a);
    else if (kb < kc) swap_at(arr, result, c);
    else swap_at(arr, result, b);
}

// std::__unguarded_partition(first, last, pivot)
MCS_HD int unguarded_partition(Arr a, int first, int last, int pivot) {
    while (true) {
        while (a.k[first] < a.k[pivot]) ++first;
        --last;
        while (a.k[pivot] < a.k[last]) --last;
        if (!(first < last)) return first;
        swap_at(a, first, last);
        ++first;
    }
}

// std::sort(k, k + n) with operator< on k; p carried along.  Recursion of __introsort_loop is unrolled onto a small stack
// (it recurses on the right part and loops on the left, so right parts are processed first, depth <= 2*log2(n)).
MCS_HD void sort(uint32_t* k, uint16_t* p, int n) {
    if (n <= 0) return;
    Arr a{k, p};
    int lg = 0;
    for (int t = n; t > 1; t >>= 1) lg++;
    // explicit stack of (first, last, depth_limit)
    int sf[64], sl[64], sd[64];
    int sp = 0;
    sf[0] = 0;
    sl[0] = n;
    sd[0] = lg * 2;
    sp = 1;
    while (sp > 0) {
        --sp;
        int first = sf[sp], last = sl[sp], depth = sd[sp];
        while (last - first > 16) {
            if (depth == 0) {
                heap_sort(a, first, last);
                break;
            }
            --depth;
            const int mid = first + (last - first) / 2;
            move_median_to_first(a, first, first + 1, mid, last - 1);
            const int cut = unguarded_partition(a, first + 1, last, first);
            // recursive call on [cut, last) happens BEFORE the loop continues on [first, cut): emulate by pushing the
            // left part (to be done later) and continuing with the right part now.
            sf[sp] = first;
            sl[sp] = cut;
            sd[sp] = depth;
            ++sp;
            first = cut;
        }
    }
    // __final_insertion_sort
    if (n > 16) {
        insertion_sort(a, 0, 16);
        for (int i = 16; i != n; ++i) unguarded_linear_insert(a, i);
    } else {
        insertion_sort(a, 0, n);
    }
}

}  // namespace mcs_sort
#endif
