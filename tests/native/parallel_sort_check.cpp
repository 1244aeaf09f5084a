// Host model of the CTA-parallel std::sort in csrc/octree.cu (block_std_sort / warp_partition):
// the same data-parallel formulation (stopper sequences, disjoint swaps, cut rule, stable rank per
// final part), evaluated sequentially, must reproduce std::sort's permutation (GCC libstdc++)
// on tie-heavy inputs.  Exit code 0 = all cases identical.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <random>
#include <vector>

struct Ent { int size; short ulx; short node; };
static bool less_(const Ent& a, const Ent& b) {
  if (a.size < b.size) return true;
  if (a.size > b.size) return false;
  return a.ulx < b.ulx;
}
static uint64_t key(const Ent& e) { return ((uint64_t)(uint32_t)e.size << 16) | (uint16_t)e.ulx; }

static void median_to_first(Ent* r, Ent* a, Ent* b, Ent* c) {
  if (less_(*a, *b)) {
    if (less_(*b, *c)) std::swap(*r, *b);
    else if (less_(*a, *c)) std::swap(*r, *c);
    else std::swap(*r, *a);
  } else if (less_(*a, *c)) std::swap(*r, *a);
  else if (less_(*b, *c)) std::swap(*r, *c);
  else std::swap(*r, *b);
}

static int partition_model(std::vector<Ent>& a, int first, int last) {
  median_to_first(&a[first], &a[first + 1], &a[first + (last - first) / 2], &a[last - 1]);
  const uint64_t pk = key(a[first]);
  std::vector<int> L, R;
  for (int q = first + 1; q < last; q++) if (key(a[q]) >= pk) L.push_back(q);
  for (int q = last - 1; q > first; q--) if (key(a[q]) <= pk) R.push_back(q);
  const int nmin = (int)std::min(L.size(), R.size());
  int K = 0;
  for (int k = 0; k < nmin; k++) K += L[k] < R[k];
  for (int k = 0; k < K; k++) std::swap(a[L[k]], a[R[k]]);
  int cut = 0x7fffffff;
  if (K < (int)L.size()) cut = L[K];
  if (K > 0) cut = std::min(cut, R[K - 1]);
  return cut;
}

static bool sort_model(std::vector<Ent>& a) {
  const int n = (int)a.size();
  struct Seg { int f, l, d; };
  int lg = 0; for (int m = n; m > 1; m >>= 1) lg++;
  std::vector<Seg> segs{{0, n, 2 * lg}};
  for (bool more = n > 16; more;) {
    more = false;
    const size_t ns = segs.size();
    for (size_t si = 0; si < ns; si++) {
      Seg s = segs[si];
      if (s.l - s.f <= 16) continue;
      if (s.d == 0) return false;
      const int cut = partition_model(a, s.f, s.l);
      segs[si] = {s.f, cut, s.d - 1};
      segs.push_back({cut, s.l, s.d - 1});
      if (cut - s.f > 16 || s.l - cut > 16) more = true;
    }
  }
  std::vector<Ent> tmp(n);
  for (const Seg& s : segs)
    for (int q = s.f; q < s.l; q++) {
      int r = s.f;
      for (int j = s.f; j < s.l; j++) r += key(a[j]) < key(a[q]) || (key(a[j]) == key(a[q]) && j < q);
      tmp[r] = a[q];
    }
  a = tmp;
  return true;
}

int main() {
  std::mt19937 rng(12345);
  int cases = 0, fallbacks = 0;
  for (int it = 0; it < 20000; it++) {
    const int n = 1 + rng() % (it % 7 == 0 ? 2000 : 300);
    const int ksz = 1 + rng() % 6, kx = 1 + rng() % 40;
    std::vector<Ent> a(n);
    for (int i = 0; i < n; i++) a[i] = Ent{2 + (int)(rng() % ksz), (short)(rng() % kx * 7), (short)i};
    if (it % 11 == 0) std::sort(a.begin(), a.end(), less_);           // presorted
    if (it % 13 == 0) std::reverse(a.begin(), a.end());
    std::vector<Ent> ref = a, got = a;
    std::sort(ref.begin(), ref.end(), less_);
    if (!sort_model(got)) { fallbacks++; continue; }
    cases++;
    for (int i = 0; i < n; i++)
      if (ref[i].node != got[i].node) { std::printf("MISMATCH it=%d n=%d at %d\n", it, n, i); return 1; }
  }
  std::printf("ok cases=%d fallbacks=%d\n", cases, fallbacks);
  return 0;
}
