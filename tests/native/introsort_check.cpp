// Host check of csrc/introsort.cuh against libstdc++ std::sort: same comparator with ties
// (compareNodes: (size, UL.x)), same input order -> the permutation of tied elements must match.
#include <algorithm>
#include <cstdint>
#include <random>
#include <vector>

#include "../../orbslam3_cpp_fork_b200/csrc/introsort.cuh"

struct Elem { int size; int ulx; int id; };
struct Less {
  bool operator()(const Elem& a, const Elem& b) const {
    if (a.size < b.size) return true;
    if (a.size > b.size) return false;
    return a.ulx < b.ulx;
  }
};
static bool cmp_ref(Elem& a, Elem& b) { return Less()(a, b); }

extern "C" int introsort_check(int n_trials, int max_n, int seed, int mode) {
  std::mt19937 rng(seed);
  int bad = 0;
  for (int t = 0; t < n_trials; t++) {
    int n = 1 + rng() % max_n;
    if (t < 40) n = t;  // includes 0, 1, 16, 17
    std::vector<Elem> a(n);
    int ksz = 1 + rng() % 6, kx = 1 + rng() % 8;
    for (int i = 0; i < n; i++) {
      a[i] = Elem{2 + (int)(rng() % ksz), (int)(rng() % kx) * 7, i};
      if (mode == 1) a[i].size = i;             // sorted input
      if (mode == 2) a[i].size = n - i;         // reversed
      if (mode == 3) { a[i].size = 2; a[i].ulx = 0; }  // all equal
      if (mode == 4) { a[i].size = (i * 7919) % 13; }  // adversarial-ish pattern
    }
    std::vector<Elem> b = a;
    std::sort(a.begin(), a.end(), cmp_ref);
    cuorb_sort::Introsort<Elem, Less>::sort(b.data(), b.data() + n, Less());
    for (int i = 0; i < n; i++)
      if (a[i].id != b[i].id) { bad++; break; }
  }
  return bad;
}

// organ-pipe / median-of-3 killer style input to reach the heapsort fallback
extern "C" int introsort_check_killer(int n) {
  std::vector<Elem> a(n);
  // classic median-of-3 killer sequence (Musser)
  int k = n / 2;
  for (int i = 1; i <= k; i++) {
    if (i % 2 == 1) { a[i - 1].size = i; a[i].size = k + i; }
    a[k + i - 1].size = 2 * i;
  }
  for (int i = 0; i < n; i++) { a[i].ulx = 0; a[i].id = i; a[i].size /= 3; }
  std::vector<Elem> b = a;
  std::sort(a.begin(), a.end(), cmp_ref);
  cuorb_sort::Introsort<Elem, Less>::sort(b.data(), b.data() + n, Less());
  int bad = 0;
  for (int i = 0; i < n; i++) bad += a[i].id != b[i].id;
  return bad;
}
