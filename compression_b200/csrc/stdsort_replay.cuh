// Replays the decision sequence of libstdc++'s std::sort (GCC 13 bits/stl_algo.h: introsort
// with threshold 16, median-of-three moved to the front, unguarded Hoare partition, depth
// limit 2*floor(log2 n) with heap-sort fallback, final insertion sort) on an array of
// (key, payload) items addressed through an accessor, so that items with EQUAL keys end up in
// exactly the order the reference's `std::sort(queue.begin(), queue.end())` leaves them in
// (cc/kernels/range_coding_helper_kernels.cc:172, :189 — SURVEY.md H1).
//
// Written from the published algorithm; usable from host code too (tests/ compile it with g++
// and compare against the real std::sort).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TFC_HD __host__ __device__ __forceinline__
#else
#define TFC_HD inline
#endif

namespace tfc {

struct SortItem {
  double key;
  int32_t idx;
  int32_t pad;
};

// Acc: item load/store by position.  DESC: compare with '>' (std::greater) instead of '<'.
template <typename Acc, bool DESC>
struct StdSortReplay {
  Acc a;
  TFC_HD explicit StdSortReplay(Acc acc) : a(acc) {}

  TFC_HD static bool lt(double x, double y) { return DESC ? x > y : x < y; }
  TFC_HD bool less(int i, int j) const { return lt(a.get(i).key, a.get(j).key); }
  TFC_HD void swap(int i, int j) {
    const SortItem t = a.get(i);
    a.set(i, a.get(j));
    a.set(j, t);
  }

  TFC_HD void median_to_first(int result, int x, int y, int z) {
    if (less(x, y)) {
      if (less(y, z)) swap(result, y);
      else if (less(x, z)) swap(result, z);
      else swap(result, x);
    } else if (less(x, z)) swap(result, x);
    else if (less(y, z)) swap(result, z);
    else swap(result, y);
  }

  TFC_HD int partition(int first, int last, int pivot) {
    for (;;) {
      while (less(first, pivot)) ++first;
      --last;
      while (less(pivot, last)) --last;
      if (!(first < last)) return first;
      swap(first, last);
      ++first;
    }
  }

  // ---- heap sort fallback (positions relative to `base`) ----
  TFC_HD void push_heap(int base, int hole, int top, const SortItem& v) {
    int parent = (hole - 1) / 2;
    while (hole > top && lt(a.get(base + parent).key, v.key)) {
      a.set(base + hole, a.get(base + parent));
      hole = parent;
      parent = (hole - 1) / 2;
    }
    a.set(base + hole, v);
  }
  TFC_HD void adjust_heap(int base, int hole, int len, const SortItem& v) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
      child = 2 * (child + 1);
      if (lt(a.get(base + child).key, a.get(base + child - 1).key)) --child;
      a.set(base + hole, a.get(base + child));
      hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
      child = 2 * (child + 1);
      a.set(base + hole, a.get(base + child - 1));
      hole = child - 1;
    }
    push_heap(base, hole, top, v);
  }
  TFC_HD void heap_sort(int first, int last) {
#ifdef TFC_SORT_TRACE
    ++tfc_sort_trace_heap_calls;  // host tests only: proves the fallback was exercised
#endif
    const int len = last - first;
    if (len >= 2) {
      for (int parent = (len - 2) / 2;; --parent) {
        const SortItem v = a.get(first + parent);
        adjust_heap(first, parent, len, v);
        if (parent == 0) break;
      }
    }
    for (int end = last; end - first > 1;) {
      --end;
      const SortItem v = a.get(end);
      a.set(end, a.get(first));
      adjust_heap(first, 0, end - first, v);
    }
  }

  TFC_HD void linear_insert(int last) {  // unguarded
    const SortItem v = a.get(last);
    int next = last - 1;
    while (lt(v.key, a.get(next).key)) {
      a.set(last, a.get(next));
      last = next;
      --next;
    }
    a.set(last, v);
  }
  TFC_HD void insertion_sort(int first, int last) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
      if (less(i, first)) {
        const SortItem v = a.get(i);
        for (int k = i; k > first; --k) a.set(k, a.get(k - 1));
        a.set(first, v);
      } else {
        linear_insert(i);
      }
    }
  }

  // ---- decomposition used by the warp-cooperative sort (pmf_to_cdf.cu) ----------------------
  // std::sort's result is fixed by (a) the sequence of partitions, each of which only touches its
  // own range, and (b) the final insertion sort, which never moves an item across a partition
  // boundary (left <= pivot <= right and the insertion compares strictly), i.e. it is a stable
  // insertion sort of every leaf range (<= 16 items) on its own.  So disjoint ranges can be
  // finished independently and in any order: `sort_range` is the serial introsort loop on one
  // range with the leaves insertion-sorted on the spot, and `partition_ranked` is the same
  // unguarded Hoare partition computed from candidate lists instead of two walking pointers.

  // [first, last) with `depth` levels left before the heap-sort fallback
  TFC_HD void sort_range(int first0, int last0, int depth0) {
    int sf[48], sl[48], sd[48];
    int sp = 0;
    sf[0] = first0; sl[0] = last0; sd[0] = depth0; sp = 1;
    while (sp > 0) {
      --sp;
      int first = sf[sp], last = sl[sp], depth = sd[sp];
      bool heap = false;
      while (last - first > 16) {
        if (depth == 0) { heap_sort(first, last); heap = true; break; }
        --depth;
        const int mid = first + (last - first) / 2;
        median_to_first(first, first + 1, mid, last - 1);
        const int cut = partition(first + 1, last, first);
        sf[sp] = cut; sl[sp] = last; sd[sp] = depth; ++sp;
        last = cut;
      }
      if (!heap) insertion_sort(first, last);
    }
  }

  // The partition of `partition(first + 1, last, first)` without walking: in the ORIGINAL range
  // let L[0] < L[1] < ... be the positions in [first + 1, last) whose item is not less than the
  // pivot and R[0] > R[1] > ... the positions in [first, last) whose item the pivot is not less
  // than (position `first`, the pivot itself, is the right scan's sentinel).  The two pointers of
  // the serial loop only ever read untouched positions, so its k-th swap is (L[k], R[k]) for as
  // long as L[k] < R[k] (K swaps; the predicate is monotone in k), and it returns the next left
  // stop: L[K] unless the swapped-in item at R[K-1] comes first.  `lpos` / `rasc` are scratch for
  // the candidate positions (rasc ascending: R[k] = rasc[nr - 1 - k]).
  template <typename Idx>
  TFC_HD int partition_ranked(int first, int last, Idx* lpos, Idx* rasc) {
    const double pk = a.get(first).key;
    int nl = 0, nr = 0;
    for (int i = first; i < last; ++i) {
      const double k = a.get(i).key;
      if (i > first && !lt(k, pk)) lpos[nl++] = (Idx)i;
      if (!lt(pk, k)) rasc[nr++] = (Idx)i;
    }
    int K = 0;
    while (K < nl && K < nr && (int)lpos[K] < (int)rasc[nr - 1 - K]) ++K;
    for (int k = 0; k < K; ++k) swap((int)lpos[k], (int)rasc[nr - 1 - k]);
    int cut = 0x7FFFFFFF;
    if (K < nl) cut = (int)lpos[K];
    if (K > 0 && (int)rasc[nr - K] < cut) cut = (int)rasc[nr - K];
    return cut;
  }

  // serial model of the warp algorithm: ranges above `coop` items are split with
  // partition_ranked (the warp does that step together), smaller ones go to sort_range (one
  // lane each).  Host tests compare it with the real std::sort.
  template <typename Idx>
  TFC_HD void sort_decomposed(int n, int coop, Idx* lpos, Idx* rasc) {
    if (n <= 0) return;
    int lg = 0;
    for (int m = n; m > 1; m >>= 1) ++lg;
    int sf[64], sl[64], sd[64];
    int sp = 0;
    sf[0] = 0; sl[0] = n; sd[0] = 2 * lg; sp = 1;
    while (sp > 0) {
      --sp;
      const int first = sf[sp], last = sl[sp];
      int depth = sd[sp];
      if (last - first <= coop || last - first <= 16) { sort_range(first, last, depth); continue; }
      if (depth == 0) { heap_sort(first, last); continue; }
      --depth;
      const int mid = first + (last - first) / 2;
      median_to_first(first, first + 1, mid, last - 1);
      const int cut = partition_ranked(first, last, lpos, rasc);
      sf[sp] = cut; sl[sp] = last; sd[sp] = depth; ++sp;
      sf[sp] = first; sl[sp] = cut; sd[sp] = depth; ++sp;
    }
  }

  TFC_HD void sort(int n) {
    if (n <= 0) return;
    int lg = 0;
    for (int m = n; m > 1; m >>= 1) ++lg;
    // explicit stack of pending (first, last, depth) ranges; sub-ranges are disjoint, so the
    // order they are processed in does not change the outcome
    int sf[64], sl[64], sd[64];
    int sp = 0;
    sf[0] = 0; sl[0] = n; sd[0] = 2 * lg; sp = 1;
    while (sp > 0) {
      --sp;
      int first = sf[sp], last = sl[sp], depth = sd[sp];
      while (last - first > 16) {
        if (depth == 0) { heap_sort(first, last); break; }
        --depth;
        const int mid = first + (last - first) / 2;
        median_to_first(first, first + 1, mid, last - 1);
        const int cut = partition(first + 1, last, first);
        sf[sp] = cut; sl[sp] = last; sd[sp] = depth; ++sp;
        last = cut;
      }
    }
    if (n > 16) {
      insertion_sort(0, 16);
      for (int i = 16; i != n; ++i) linear_insert(i);
    } else {
      insertion_sort(0, n);
    }
  }
};

}  // namespace tfc
