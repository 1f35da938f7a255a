// directed_scm.hpp — C++ host side of the hot path, above the C ABI of fcm.h (header only).
//
// The reference is compiled code (Rust crate `directed-scm`); rustc is not in this image, so the host
// logic that stays on the host is mirrored here in C++ with the reference's names and meaning:
//   io::read_flag_file                                  src/io.rs:18-35
//   util::{all_le, factorial, binomial, calc_relax_de}   src/util.rs:53-105
//   Bounds::{calculate, check}                          src/lib.rs:119-160
//   initialize_new_sampler (target bounds, distance)    src/bin/sample.rs:80-105
//   MCMCSampler::{next, acceptance_ratio}               src/lib.rs:181-197
// Everything that computes on the graph goes through libfcm.so (no CPU fallback); a failed call
// throws directed_scm::Error where the reference panics.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "fcm.h"

namespace directed_scm {

using Node = uint32_t;                                   // src/flagser.rs:9
using Edge = std::array<Node, 2>;                        // (from, to)
using Counts = std::vector<uint64_t>;                    // Vec<usize>

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
inline void check(int rc) {
  if (rc != FCM_OK) throw Error(rc, std::string("libfcm: ") + fcm_last_error());
}

// ---- io.rs ------------------------------------------------------------------------------------------
// line 1 skipped, the token count of line 2 is the number of vertices, line 3 skipped, then "i j [w]"
// Tokens are split on ' ' ONLY, empty tokens dropped (src/io.rs:25,29: `split(' ').filter(|s| *s != "")`), lines as
// Rust's str::lines() yields them (split on '\n', one trailing '\r' removed).  A token that is not a u32 - e.g. two
// numbers joined by a tab - throws where the reference's `parse().unwrap()` panics (:31).
inline std::pair<uint32_t, std::vector<Edge>> read_flag_file(const std::string& fname) {
  std::ifstream f(fname, std::ios::binary);
  if (!f) throw Error(FCM_EINVAL, "cannot open " + fname);
  std::string line;
  auto next_line = [&](std::string& out) {
    if (!std::getline(f, out, '\n')) return false;
    if (!out.empty() && out.back() == '\r') out.pop_back();
    return true;
  };
  auto tokens = [](const std::string& s) {
    std::vector<std::string> t;
    size_t i = 0;
    while (i <= s.size()) {
      const size_t j = std::min(s.find(' ', i), s.size());
      if (j > i) t.push_back(s.substr(i, j - i));
      i = j + 1;
    }
    return t;
  };
  auto parse_node = [&](const std::string& w) {             // Rust u32::from_str: optional '+', then decimal digits only
    size_t i = (!w.empty() && w[0] == '+') ? 1 : 0;
    if (i >= w.size()) throw Error(FCM_EINVAL, "not a vertex id: '" + w + "' in " + fname);
    uint64_t v = 0;
    for (; i < w.size(); i++) {
      if (w[i] < '0' || w[i] > '9') throw Error(FCM_EINVAL, "not a vertex id: '" + w + "' in " + fname);
      v = v * 10 + (uint64_t)(w[i] - '0');
      if (v > 0xffffffffull) throw Error(FCM_EINVAL, "vertex id out of range: '" + w + "' in " + fname);
    }
    return (Node)v;
  };
  next_line(line);                                           // skip "dim 0:"
  if (!next_line(line)) throw Error(FCM_EINVAL, "no vertex line in " + fname);   // reference: unwrap() on None panics
  const uint32_t n = (uint32_t)tokens(line).size();
  next_line(line);                                           // skip "dim 1:"
  std::vector<Edge> edges;
  while (next_line(line)) {
    const auto t = tokens(line);
    if (t.size() >= 2) edges.push_back({parse_node(t[0]), parse_node(t[1])});
  }
  return {n, edges};
}

// ---- util.rs ----------------------------------------------------------------------------------------
inline bool all_le(const Counts& a, const Counts& b, uint64_t z = 0) {      // :53-63, shorter side padded with z
  const size_t m = std::max(a.size(), b.size());
  for (size_t i = 0; i < m; i++)
    if ((i < a.size() ? a[i] : z) > (i < b.size() ? b[i] : z)) return false;
  return true;
}
inline uint64_t factorial(uint64_t x) {                  // :65-71: `for i in 1..x` => (x-1)!, kept as it is
  uint64_t r = 1;
  for (uint64_t i = 1; i < x; i++) r *= i;
  return r;
}
inline uint64_t binomial(uint64_t n, uint64_t k) {       // :73-77 on that factorial, integer division
  if (k > n) throw Error(FCM_EASSERT, "binomial: n - k underflows (reference panics)");
  const uint64_t den = factorial(k) * factorial(n - k);
  if (den == 0) throw Error(FCM_EASSERT, "binomial: divisor wrapped to zero (reference panics)");
  return factorial(n) / den;
}
// OEIS A058298 (triangle n!/(n-k), 1 <= k < n, read by rows), the 64 entries of src/util.rs:98-105
inline const std::vector<uint64_t>& oeis_a058298() {
  static const std::vector<uint64_t> t = [] {
    std::vector<uint64_t> v;
    uint64_t fact = 1;
    for (uint64_t n = 2; v.size() < 64; n++) {
      fact *= n;                                         // n!
      for (uint64_t k = 1; k < n && v.size() < 64; k++) v.push_back(fact / (n - k));
    }
    return v;
  }();
  return t;
}
inline Counts calc_relax_de(const Counts& sc) {          // :79-93
  const auto& T = oeis_a058298();
  Counts out;
  for (size_t d = 0; d < sc.size(); d++) {
    size_t ind = 1;
    uint64_t best = 0;
    bool any = false;
    for (;;) {
      if (ind >= T.size()) throw Error(FCM_EASSERT, "calc_relax_de: OEIS_A058298 index out of range (reference panics)");
      if (!(T[ind] < sc[d])) break;
      best = std::max(best, T[ind] - T[ind - 1]);
      any = true;
      ind++;
    }
    out.push_back(std::min<uint64_t>(any ? best : 1, factorial(d + 1)));
  }
  return out;
}
inline uint64_t f64_as_usize(double x) {                 // Rust `as usize`: saturating, NaN -> 0
  if (!(x > 0.0)) return 0;
  if (x >= 18446744073709551616.0) return UINT64_MAX;
  return (uint64_t)x;
}

// ---- Bounds (src/lib.rs:113-161) ---------------------------------------------------------------------
struct Bounds {
  Counts flag_count_min, flag_count_max;

  // src/bin/sample.rs:89-92: exact for d < 2, floor(s * (1 -+ t)) in f64 for d >= 2
  static Bounds target(const Counts& flag_count, double target_relaxation) {
    Bounds b;
    for (size_t d = 0; d < flag_count.size(); d++) {
      const uint64_t s = flag_count[d];
      b.flag_count_min.push_back(d < 2 ? s : f64_as_usize(std::floor((double)s * (1. - target_relaxation))));
      b.flag_count_max.push_back(d < 2 ? s : f64_as_usize(std::floor((double)s * (1. + target_relaxation))));
    }
    return b;
  }
  // src/lib.rs:119-156; ncliques = flagser_count of the normalised graph (:125-130)
  static Bounds calculate(const Counts& flag_count, const Counts& ncliques, uint64_t n_undirected_edges, const Bounds& target) {
    if (flag_count.size() > 1 && n_undirected_edges == flag_count[1]) return Bounds{target.flag_count_min, ncliques};   // SEO case
    Bounds b = target;
    const Counts relax_de = calc_relax_de(flag_count);
    const size_t len = flag_count.size();
    for (size_t d = 2; d < len; d++) {
      const uint64_t relax = relax_de[d] * binomial(len - 2, d - 1);
      b.flag_count_max[d] = std::max(b.flag_count_min[d] + relax, b.flag_count_max[d]);
      if (b.flag_count_max[d] < relax) throw Error(FCM_EASSERT, "flag_count_max[d] - relax underflows (reference panics in debug)");
      b.flag_count_min[d] = std::min(b.flag_count_max[d] - relax, b.flag_count_min[d]);
    }
    if (b.flag_count_max.size() <= 2) throw Error(FCM_EASSERT, "flag_count_max[2] out of range (reference panics)");
    b.flag_count_max[2] = UINT64_MAX;                    // :151
    b.flag_count_max.push_back(10);                      // :152
    return b;
  }
  bool check(const Counts& flag_count) const {           // :157-160
    return all_le(flag_count_min, flag_count) && all_le(flag_count, flag_count_max);
  }
};

// ---- device plumbing -----------------------------------------------------------------------------------
class Context {
 public:
  explicit Context(int device = 0) { check(fcm_ctx_create(device, &h_)); }
  ~Context() { if (h_) fcm_ctx_destroy(h_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  fcm_ctx* handle() const { return h_; }
  void synchronize() { check(fcm_ctx_synchronize(h_)); }
 private:
  fcm_ctx* h_ = nullptr;
};

inline Counts trim(const uint64_t* c, int len) { return Counts(c, c + len); }

// graph.flagser_count() (src/lib.rs:51) / the count of the normalised graph (:125-130)
inline Counts flagser_count(Context& ctx, uint32_t n, const std::vector<Edge>& edges, bool normalised = false,
                            uint64_t* nedges_of_graph = nullptr) {
  fcm_graph* g = nullptr;
  const auto* e = reinterpret_cast<const uint32_t(*)[2]>(edges.data());
  check(normalised ? fcm_graph_create_normalised(ctx.handle(), n, edges.size(), e, &g)
                   : fcm_graph_create(ctx.handle(), n, edges.size(), e, &g));
  uint64_t counts[FCM_MAXD];
  int len = 0;
  const int rc = fcm_graph_count(g, 0, 1, counts, &len);
  if (nedges_of_graph) fcm_graph_info(g, nullptr, nedges_of_graph, nullptr, nullptr);
  fcm_graph_destroy(g);
  check(rc);
  return trim(counts, len);
}

// graph.compute_maximal_cliques() (external crate, src/lib.rs:41): vertices ascending, cliques lexicographic.  Host code.
inline std::vector<std::vector<Node>> compute_maximal_cliques(uint32_t n, const std::vector<Edge>& edges) {
  const auto* e = reinterpret_cast<const uint32_t(*)[2]>(edges.data());
  uint64_t nc = 0, nv = 0;
  check(fcm_maximal_cliques(n, edges.size(), e, &nc, &nv, nullptr, nullptr));
  std::vector<uint32_t> sizes(nc ? nc : 1), verts(nv ? nv : 1);
  check(fcm_maximal_cliques(n, edges.size(), e, &nc, &nv, sizes.data(), verts.data()));
  std::vector<std::vector<Node>> out;
  out.reserve(nc);
  size_t at = 0;
  for (uint64_t c = 0; c < nc; c++) { out.emplace_back(verts.begin() + at, verts.begin() + at + sizes[c]); at += sizes[c]; }
  return out;
}
// State::new's filing by order (src/lib.rs:42-49): cliques_by_order[k - 1] = the maximal cliques of k vertices
inline std::vector<std::vector<std::vector<Node>>> cliques_by_order(uint32_t n, const std::vector<Edge>& edges) {
  std::vector<std::vector<std::vector<Node>>> by;
  for (auto& c : compute_maximal_cliques(n, edges)) {
    if (c.size() > by.size()) by.resize(c.size());
    by[c.size() - 1].push_back(std::move(c));
  }
  return by;
}

const double MOVE_DISTRIBUTION_SIMPLE[4] = {0.5, 0.5, 0.0, 0.0};   // src/bin/sample.rs:16
const double MOVE_DISTRIBUTION[4] = {0.1, 0.1, 0.6, 0.2};          // src/bin/sample.rs:17

struct ChainView {                                       // what sample.rs reads from `&State` after next()
  Counts flag_count;
  std::vector<Edge> edges;                               // state.graph.edges(), container order
  uint64_t sampled = 0, accepted = 0;
  double acceptance_ratio() const { return sampled ? (double)accepted / (double)sampled : NAN; }   // src/lib.rs:195-197
};

// MCMCSampler (src/lib.rs:164-198) for seeds.size() independent chains on one device
class MCMCSampler {
 public:
  MCMCSampler(Context& ctx, uint32_t n, const std::vector<Edge>& edges, const std::vector<uint64_t>& seeds,
              const Bounds& bounds, const Counts& flag_count0, const double move_distribution[4], uint64_t sample_distance)
      : bounds(bounds), sample_distance(sample_distance), nedges_(edges.size()) {
    check(fcm_ensemble_create(ctx.handle(), n, edges.size(), reinterpret_cast<const uint32_t(*)[2]>(edges.data()),
                              (uint32_t)seeds.size(), seeds.data(), flag_count0.data(), (int)flag_count0.size(),
                              bounds.flag_count_min.data(), (int)bounds.flag_count_min.size(),
                              bounds.flag_count_max.data(), (int)bounds.flag_count_max.size(), move_distribution, &h_));
  }
  ~MCMCSampler() { if (h_) fcm_ensemble_destroy(h_); }
  MCMCSampler(MCMCSampler&& o) noexcept : bounds(std::move(o.bounds)), sample_distance(o.sample_distance), h_(o.h_), nedges_(o.nedges_) { o.h_ = nullptr; }
  MCMCSampler(const MCMCSampler&) = delete;
  MCMCSampler& operator=(const MCMCSampler&) = delete;

  void step(uint64_t nproposals) { check(fcm_ensemble_step(h_, nproposals)); }
  void next() { step(sample_distance); }                 // src/lib.rs:181-194
  ChainView state(uint32_t chain) {
    ChainView v;
    uint64_t fc[FCM_MAXD], ne = 0, nd = 0;
    int len = 0;
    std::vector<uint32_t> e(2 * 2 * nedges_ + 2);         // a chain never holds more than both directions of every pair
    check(fcm_ensemble_read(h_, chain, fc, &len, &v.sampled, &v.accepted, e.data(), &ne, &nd, nullptr));
    v.flag_count = trim(fc, len);
    for (uint64_t i = 0; i < ne; i++) v.edges.push_back({e[2 * i], e[2 * i + 1]});
    return v;
  }
  fcm_ensemble* handle() const { return h_; }

  Bounds bounds;
  uint64_t sample_distance;

 private:
  fcm_ensemble* h_ = nullptr;
  size_t nedges_;
};

// src/bin/sample.rs:80-105 for one or many seeds
inline MCMCSampler initialize_new_sampler(Context& ctx, uint32_t n, const std::vector<Edge>& edges,
                                          const std::vector<uint64_t>& seeds, double target_relaxation = 0.01,
                                          bool simple = false, uint64_t sample_distance = 0) {
  const Counts flag_count = flagser_count(ctx, n, edges);                       // State::new
  uint64_t n_und = 0;
  const Counts ncliques = flagser_count(ctx, n, edges, true, &n_und);           // Bounds::calculate's normalised graph
  const Bounds bounds = Bounds::calculate(flag_count, ncliques, n_und, Bounds::target(flag_count, target_relaxation));
  if (sample_distance == 0 && flag_count.size() > 1) {                          // :102
    const double E = (double)flag_count[1];
    sample_distance = f64_as_usize(std::ceil(2. * E * std::log2(E)));
  }
  return MCMCSampler(ctx, n, edges, seeds, bounds, flag_count, simple ? MOVE_DISTRIBUTION_SIMPLE : MOVE_DISTRIBUTION, sample_distance);
}

}  // namespace directed_scm
