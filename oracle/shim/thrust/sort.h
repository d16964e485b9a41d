// thrust/sort.h shim — TEST INFRASTRUCTURE. The reference calls thrust::sort
// from device code (hitable/bvh.h:86-92), which runs sequentially in the
// calling thread. Its comparator is not a strict weak ordering
// (bvh.h:36-41), so std::sort would be undefined behaviour; a hand-written
// insertion/merge that only ever asks cmp(a, b) is well defined for any cmp.
#pragma once
#include <vector>
#include <cstddef>
namespace thrust {
template <class It, class Cmp>
void sort(It first, It last, Cmp cmp)
{
    typedef typename std::iterator_traits<It>::value_type T;
    std::ptrdiff_t n = last - first;
    if (n < 2) return;
    std::vector<T> a(first, last), b(n);
    for (std::ptrdiff_t w = 1; w < n; w *= 2) {
        for (std::ptrdiff_t lo = 0; lo < n; lo += 2 * w) {
            std::ptrdiff_t mid = lo + w < n ? lo + w : n;
            std::ptrdiff_t hi = lo + 2 * w < n ? lo + 2 * w : n;
            std::ptrdiff_t i = lo, j = mid, k = lo;
            while (i < mid && j < hi) {
                if (cmp(a[j], a[i])) b[k++] = a[j++];
                else b[k++] = a[i++];
            }
            while (i < mid) b[k++] = a[i++];
            while (j < hi) b[k++] = a[j++];
        }
        a.swap(b);
    }
    for (std::ptrdiff_t i = 0; i < n; i++) first[i] = a[i];
}
}  // namespace thrust
