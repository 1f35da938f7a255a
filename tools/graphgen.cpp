// graphgen.cpp — synthetic directed graphs of the shapes BASELINE.json names (SURVEY.md
// Appendix A).  Stateless SplitMix64-style hash so that every implementation agrees:
//   edge i->j (i != j)  iff  u(h(seed,i,j)) < p(i,j)   (strict <, f64)
// Edge order = row-major (i ascending, then j ascending), the order a .flag file would list.
// This is input generation for tests/bench; it is neither the oracle nor the product.
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
typedef uint64_t u64;
typedef uint32_t u32;

static inline u64 mix(u64 z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
static inline u64 h(u64 seed, u64 i, u64 j) {
  return mix(mix(seed + 0x9E3779B97F4A7C15ULL * (i + 1)) ^ (0xD1B54A32D192ED03ULL * (j + 1)));
}
static inline double u01(u64 x) { return (double)(x >> 11) * (1.0 / 9007199254740992.0); }

extern "C" {
// kind 0: ER        params = {p}
// kind 1: Chung-Lu  params = {c, alpha}          p = min(1, c * (i+1)^-alpha * (j+1)^-alpha)
// kind 2: geometric params = {A, lambda, bx, by, bz}   p = min(1, A exp(-dist/lambda))
// returns number of edges; *edges_out = malloc'd u32[2*E] (free with gg_free)
u64 gg_generate(int kind, u32 n, u64 seed, const double* params, int nthreads, u32** edges_out) {
  if (nthreads < 1) nthreads = 1;
  std::vector<double> w, px, py, pz;
  if (kind == 1) { w.resize(n); for (u32 i = 0; i < n; i++) w[i] = std::pow((double)(i + 1), -params[1]); }
  if (kind == 2) {
    px.resize(n); py.resize(n); pz.resize(n);
    for (u32 i = 0; i < n; i++) {
      px[i] = u01(h(seed ^ 0xABCDEFULL, i, 0)) * params[2];
      py[i] = u01(h(seed ^ 0xABCDEFULL, i, 1)) * params[3];
      pz[i] = u01(h(seed ^ 0xABCDEFULL, i, 2)) * params[4];
    }
  }
  std::vector<std::vector<u32>> rows(n);
  auto work = [&](int t) {
    for (u32 i = t; i < n; i += nthreads) {
      std::vector<u32>& r = rows[i];
      for (u32 j = 0; j < n; j++) {
        if (i == j) continue;
        double p;
        if (kind == 0) p = params[0];
        else if (kind == 1) { p = params[0] * w[i] * w[j]; if (p > 1.0) p = 1.0; }
        else {
          double dx = px[i] - px[j], dy = py[i] - py[j], dz = pz[i] - pz[j];
          double dist = std::sqrt(dx * dx + dy * dy + dz * dz);
          // cheap reject: p <= A*exp(-dist/lambda); skip the hash when p is astronomically small
          p = params[0] * std::exp(-dist / params[1]);
          if (p > 1.0) p = 1.0;
        }
        if (u01(h(seed, i, j)) < p) r.push_back(j);
      }
    }
  };
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++) th.emplace_back(work, t);
  for (auto& t : th) t.join();
  u64 E = 0;
  for (auto& r : rows) E += r.size();
  u32* out = (u32*)malloc(sizeof(u32) * 2 * (E ? E : 1));
  u64 k = 0;
  for (u32 i = 0; i < n; i++) for (u32 j : rows[i]) { out[2 * k] = i; out[2 * k + 1] = j; k++; }
  *edges_out = out;
  return E;
}
void gg_copy(const u32* src, u64 nedges, u32* dst) { memcpy(dst, src, sizeof(u32) * 2 * nedges); }
void gg_free(u32* p) { free(p); }
}
