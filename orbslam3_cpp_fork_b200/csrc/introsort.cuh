// Deterministic re-implementation of the algorithm libstdc++'s std::sort runs (introsort:
// median-of-3 quicksort with unguarded partition, depth limit 2*floor(log2 n) -> heapsort,
// then one final insertion sort; threshold 16).  DistributeOctTree sorts (nKeys, UL.x) pairs
// with a comparator that has ties (src/ORBextractor.cc:542-554, 689) and std::sort is not stable,
// so where tied nodes land decides which node is split last before the quota is reached.  To be
// bit-exact with the reference built against GCC's libstdc++, the device runs the same sequence of
// comparisons and moves.  tests/test_introsort.py checks this header against std::sort on the
// host (same element permutation, ties included).
#pragma once

#ifdef __CUDACC__
#define ISORT_HD __host__ __device__ __forceinline__
#else
#define ISORT_HD inline
#endif

namespace cuorb_sort {

// T: trivially copyable element; Less: bool operator()(const T&, const T&) const
template <typename T, typename Less>
struct Introsort {
  enum { kThreshold = 16 };

  ISORT_HD static void swp(T& a, T& b) { T t = a; a = b; b = t; }

  ISORT_HD static void move_median_to_first(T* result, T* a, T* b, T* c, const Less& lt) {
    if (lt(*a, *b)) {
      if (lt(*b, *c)) swp(*result, *b);
      else if (lt(*a, *c)) swp(*result, *c);
      else swp(*result, *a);
    } else if (lt(*a, *c)) swp(*result, *a);
    else if (lt(*b, *c)) swp(*result, *c);
    else swp(*result, *b);
  }

  ISORT_HD static T* unguarded_partition(T* first, T* last, T* pivot, const Less& lt) {
    for (;;) {
      while (lt(*first, *pivot)) ++first;
      --last;
      while (lt(*pivot, *last)) --last;
      if (!(first < last)) return first;
      swp(*first, *last);
      ++first;
    }
  }

  ISORT_HD static void push_heap(T* first, long hole, long top, T value, const Less& lt) {
    long parent = (hole - 1) / 2;
    while (hole > top && lt(first[parent], value)) {
      first[hole] = first[parent];
      hole = parent;
      parent = (hole - 1) / 2;
    }
    first[hole] = value;
  }

  ISORT_HD static void adjust_heap(T* first, long hole, long len, T value, const Less& lt) {
    const long top = hole;
    long child = hole;
    while (child < (len - 1) / 2) {
      child = 2 * (child + 1);
      if (lt(first[child], first[child - 1])) child--;
      first[hole] = first[child];
      hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
      child = 2 * (child + 1);
      first[hole] = first[child - 1];
      hole = child - 1;
    }
    push_heap(first, hole, top, value, lt);
  }

  // std::__partial_sort(first, last, last): make_heap + sort_heap
  ISORT_HD static void heap_sort(T* first, T* last, const Less& lt) {
    const long len = last - first;
    if (len >= 2) {
      long parent = (len - 2) / 2;
      for (;;) {
        T value = first[parent];
        adjust_heap(first, parent, len, value, lt);
        if (parent == 0) break;
        parent--;
      }
    }
    while (last - first > 1) {
      --last;
      T value = *last;
      *last = *first;
      adjust_heap(first, 0, last - first, value, lt);
    }
  }

  ISORT_HD static void unguarded_linear_insert(T* last, const Less& lt) {
    T val = *last;
    T* next = last - 1;
    while (lt(val, *next)) {
      *last = *next;
      last = next;
      --next;
    }
    *last = val;
  }

  ISORT_HD static void insertion_sort(T* first, T* last, const Less& lt) {
    if (first == last) return;
    for (T* i = first + 1; i != last; ++i) {
      if (lt(*i, *first)) {
        T val = *i;
        for (T* p = i; p != first; --p) *p = *(p - 1);
        *first = val;
      } else {
        unguarded_linear_insert(i, lt);
      }
    }
  }

  ISORT_HD static int lg(long n) {
    int k = 0;
    while (n > 1) { n >>= 1; k++; }
    return k;
  }

  // std::sort(first, last, lt).  The recursion of __introsort_loop (recurse on the right part,
  // iterate on the left) is unrolled with an explicit stack of (first, last, depth) records.
  ISORT_HD static void sort(T* first, T* last, const Less& lt) {
    if (first == last) return;
    struct Rec { T* f; T* l; int d; };
    Rec stack[64];
    int sp = 0;
    stack[sp++] = Rec{first, last, 2 * lg(last - first)};
    while (sp > 0) {
      Rec r = stack[--sp];
      T* f = r.f; T* l = r.l; int depth = r.d;
      // pending left parts must run AFTER the right part finishes, as in the recursion; since the
      // parts are disjoint ranges the order of processing does not change the result.
      while (l - f > kThreshold) {
        if (depth == 0) { heap_sort(f, l, lt); break; }
        --depth;
        T* mid = f + (l - f) / 2;
        move_median_to_first(f, f + 1, mid, l - 1, lt);
        T* cut = unguarded_partition(f + 1, l, f, lt);
        stack[sp++] = Rec{cut, l, depth};
        l = cut;
      }
    }
    // __final_insertion_sort
    if (last - first > kThreshold) {
      insertion_sort(first, first + kThreshold, lt);
      for (T* i = first + kThreshold; i != last; ++i) unguarded_linear_insert(i, lt);
    } else {
      insertion_sort(first, last, lt);
    }
  }
};

}  // namespace cuorb_sort
