// Host check of compression_b200/csrc/stdsort_replay.cuh against the real libstdc++
// std::sort: the permutation of items with equal keys must be identical.
// Includes McIlroy's adaptive adversary ("A Killer Adversary for Quicksort", 1999) to drive
// std::sort into its heap-sort fallback.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <random>
#include <vector>

static long tfc_sort_trace_heap_calls = 0;
#define TFC_SORT_TRACE 1
#include "stdsort_replay.cuh"

struct VecAcc {
  tfc::SortItem* p;
  tfc::SortItem get(int i) const { return p[i]; }
  void set(int i, const tfc::SortItem& v) const { p[i] = v; }
};

struct Item { int* ptr; double mass; double key; };  // same shape as the reference's items

template <bool DESC>
static bool run_case(const std::vector<double>& keys) {
  const int n = (int)keys.size();
  std::vector<int> dummy(n);
  std::vector<Item> ref(n);
  std::vector<tfc::SortItem> mine(n);
  for (int i = 0; i < n; ++i) {
    ref[i] = Item{&dummy[i], 0.0, keys[i]};
    mine[i].key = keys[i]; mine[i].idx = i; mine[i].pad = 0;
  }
  if (DESC) std::sort(ref.begin(), ref.end(), [](const Item& a, const Item& b) { return a.key > b.key; });
  else std::sort(ref.begin(), ref.end(), [](const Item& a, const Item& b) { return a.key < b.key; });
  const std::vector<tfc::SortItem> unsorted = mine;
  tfc::StdSortReplay<VecAcc, DESC> s(VecAcc{mine.data()});
  s.sort(n);
  for (int i = 0; i < n; ++i)
    if ((int)(ref[i].ptr - dummy.data()) != mine[i].idx) return false;
  // the decomposition the warp-cooperative device sort uses (candidate-list partitions above
  // `coop` items, independent serial ranges below) must give the same permutation
  std::vector<uint16_t> lpos(n + 1), rasc(n + 1);
  for (int coop : {16, 24, 64, 200}) {
    std::vector<tfc::SortItem> again = unsorted;
    tfc::StdSortReplay<VecAcc, DESC> d(VecAcc{again.data()});
    d.sort_decomposed(n, coop, lpos.data(), rasc.data());
    for (int i = 0; i < n; ++i)
      if (again[i].idx != mine[i].idx) return false;
  }
  return true;
}

// McIlroy adversary: builds, by answering std::sort's comparisons adaptively, a key array on
// which that very std::sort goes quadratic (=> depth limit => heap sort).
static std::vector<double> killer(int n) {
  std::vector<int> val(n), ptr(n);
  const int gas = n - 1;
  for (int i = 0; i < n; ++i) { ptr[i] = i; val[i] = gas; }
  int nsolid = 0, candidate = 0;
  auto cmp = [&](int x, int y) {
    if (val[x] == gas && val[y] == gas) {
      if (x == candidate) val[x] = nsolid++; else val[y] = nsolid++;
    }
    if (val[x] == gas) candidate = x; else if (val[y] == gas) candidate = y;
    return val[x] < val[y];
  };
  std::sort(ptr.begin(), ptr.end(), cmp);
  std::vector<double> keys(n);
  for (int i = 0; i < n; ++i) keys[i] = (double)val[i];
  return keys;
}

int main() {
  std::mt19937_64 rng(12345);
  long cases = 0, bad = 0;
  auto check = [&](const std::vector<double>& k) {
    ++cases;
    if (!run_case<false>(k)) { ++bad; std::printf("ASC mismatch n=%zu\n", k.size()); }
    if (!run_case<true>(k)) { ++bad; std::printf("DESC mismatch n=%zu\n", k.size()); }
  };
  const int sizes[] = {0, 1, 2, 3, 5, 15, 16, 17, 18, 22, 31, 32, 33, 64, 100, 257, 1000, 1481, 4096};
  for (int n : sizes) {
    for (int rep = 0; rep < 40; ++rep) {
      const int distinct = 1 + (int)(rng() % (rep < 10 ? 3 : (rep < 25 ? 17 : 100000)));
      std::vector<double> k(n);
      for (auto& x : k) x = (double)(rng() % distinct);
      if (rep % 7 == 3) for (int i = 0; i < n; ++i) k[i] = std::abs(i - n / 2) / 3;  // symmetric
      if (rep % 7 == 4) for (int i = 0; i < n; ++i) k[i] = (i % 5 == 0) ? 1.0 / 0.0 : k[i];
      if (rep % 7 == 5) std::sort(k.begin(), k.end());
      if (rep % 7 == 6) std::sort(k.begin(), k.end(), std::greater<double>());
      check(k);
    }
  }
  for (int n : {40, 100, 500, 1481, 3000}) {
    std::vector<double> k = killer(n);
    check(k);
    for (auto& x : k) x = (double)((long)x / 3);  // same adversarial shape, now with ties
    check(k);
    std::vector<double> r(k.rbegin(), k.rend());
    check(r);
  }
  std::printf("cases=%ld bad=%ld heap_sort_calls=%ld\n", cases, bad, tfc_sort_trace_heap_calls);
  return (bad || tfc_sort_trace_heap_calls == 0) ? 1 : 0;
}
