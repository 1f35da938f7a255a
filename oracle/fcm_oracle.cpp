// fcm_oracle.cpp — CPU restatement of the flag-complex-mcmc hot path.
//
// THIS IS TEST INFRASTRUCTURE.  It is the checker the CUDA path is compared with, and the
// "C++ restatement of reference CPU path" timed as the CPU baseline.  Nothing in the product
// (libfcm.so, the Python/C++ host mirror) links, imports or falls back to it.
//
// Parity status
//   * simplex counts / deltas : pinned by mathematics and by the known answers in
//     tests/golden/ (four reference fixtures + Testcases.pdf toy graphs, each computed by three
//     independent methods).
//   * seeded trajectories     : PARITY UNPINNED.  The reference cannot be built here (no rustc,
//     no network) and the code that defines the trajectory is in un-vendored, un-pinned
//     third-party crates: `flag-complex` (git HEAD, Cargo.toml:26), `rand = "*"` and
//     `rand_xoshiro = "*"` (Cargo.toml:13,23).  The restatement follows the published
//     algorithms of rand 0.8.x / rand_xoshiro 0.6.x and ONE documented assumption about the
//     edge container (struct EdgeSet below).  The CUDA path is bit-exact against THIS file.
//
// What is restated, with the reference lines followed:
//   State::new / apply_transition / revert_transition / edgeset_neighborhood   src/lib.rs:38-111
//   Bounds::calculate / Bounds::check                                        src/lib.rs:119-160
//   MCMCSampler::next / acceptance_ratio                                     src/lib.rs:181-197
//   Transition::random_move / single_edge_flip / double_edge_move            src/lib.rs:207-212,292-325
//   compute_edge_neighborhoods                                               src/lib.rs:331-356
//   util::{intersect_sorted, all_le, factorial, binomial, calc_relax_de}     src/util.rs:5-26,53-105
//   initialize_new_sampler (target bounds, sample distance, move mix)        src/bin/sample.rs:80-105
//   io::read_flag_file                                                       src/io.rs:18-35
//   flagser_count (external crate; algorithm = luetge/flagser directed-clique recursion,
//   credited in README.md:29)                                                SURVEY.md §3.3
//
// Build: see oracle/Makefile (g++ -O3 -march=native -shared).  C ABI at the bottom (orc_*).

#include <algorithm>
#include <unordered_set>
#include <tuple>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

typedef uint64_t u64;
typedef uint32_t u32;
typedef u32 Node;                      // src/flagser.rs:9, src/bin/distance_matrix.rs:44
struct Edge { Node a, b; };            // = [Node;2]  (from, to)

// ---------------------------------------------------------------------------------------------
// rand 0.8 / rand_xoshiro 0.6 primitives (published algorithms; see header: unpinned)
// ---------------------------------------------------------------------------------------------
struct SplitMix64 {                    // rand_xoshiro::SplitMix64
  u64 x;
  u64 next_u64() {
    x += 0x9e3779b97f4a7c15ULL;
    u64 z = x;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
  }
};

struct Xoshiro256SS {                  // rand_xoshiro::Xoshiro256StarStar
  u64 s[4];
  static inline u64 rotl(u64 x, int k) { return (x << k) | (x >> (64 - k)); }
  u64 next_u64() {
    u64 result = rotl(s[1] * 5, 7) * 9;
    u64 t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
    s[2] ^= t;
    s[3] = rotl(s[3], 45);
    return result;
  }
  // seed_from_u64: four successive SplitMix64 outputs (src/bin/sample.rs:86 calls this)
  static Xoshiro256SS seed_from_u64(u64 seed) {
    SplitMix64 sm{seed};
    Xoshiro256SS r;
    for (int i = 0; i < 4; i++) r.s[i] = sm.next_u64();
    return r;
  }
};

// rand 0.8 UniformInt<usize>::sample_single(0, range) on a 64-bit target (gen_range(0..range)).
static inline u64 gen_range_u64(Xoshiro256SS& rng, u64 range) {
  u64 zone = (range << __builtin_clzll(range)) - 1;
  for (;;) {
    u64 v = rng.next_u64();
    unsigned __int128 m = (unsigned __int128)v * range;
    u64 hi = (u64)(m >> 64), lo = (u64)m;
    if (lo <= zone) return hi;
  }
}

// rand 0.8 UniformInt<u32>::sample_single(0, range): next_u32 = upper half of next_u64
// (rand_xoshiro 0.6 Xoshiro256StarStar::next_u32), 32x32 widening multiply, rejection zone.
static inline u32 gen_range_u32(Xoshiro256SS& rng, u32 range) {
  u32 zone = (range << __builtin_clz(range)) - 1;
  for (;;) {
    u32 v = (u32)(rng.next_u64() >> 32);
    u64 m = (u64)v * range;
    if ((u32)m <= zone) return (u32)(m >> 32);
  }
}
// rand 0.8 seq::gen_index: the u32 sampler when the bound fits u32 (SliceRandom::choose / shuffle)
static inline size_t gen_index(Xoshiro256SS& rng, size_t ubound) {
  if (ubound <= 0xffffffffULL) return gen_range_u32(rng, (u32)ubound);
  return gen_range_u64(rng, ubound);
}
// util::random_perm (src/util.rs:28-32): (l..h) then SliceRandom::shuffle =
// for i in (1..len).rev() { swap(i, gen_index(rng, i + 1)) }
static inline std::vector<size_t> random_perm(size_t l, size_t h, Xoshiro256SS& rng) {
  std::vector<size_t> perm;
  for (size_t x = l; x < h; x++) perm.push_back(x);
  for (size_t i = perm.size(); i-- > 1;) std::swap(perm[i], perm[gen_index(rng, i + 1)]);
  return perm;
}

// rand 0.8 Bernoulli: p_int = (p * 2^64) as u64 ; sample = next_u64 < p_int  (p < 1)
static inline bool gen_bool(Xoshiro256SS& rng, double p) {
  if (p >= 1.0) return true;                      // ALWAYS_TRUE consumes no draw
  u64 p_int = (u64)(p * 18446744073709551616.0);
  return rng.next_u64() < p_int;
}

// rand 0.8 WeightedIndex<f64>: cumulative sums without the last weight, Uniform<f64>(0,total)
struct WeightedIndexF64 {
  std::vector<double> cum;
  double total = 0, scale = 0;
  bool init(const double* w, size_t n) {
    if (n == 0) return false;
    cum.clear();
    total = w[0];
    if (!(w[0] >= 0)) return false;
    for (size_t i = 1; i < n; i++) {
      if (!(w[i] >= 0)) return false;
      cum.push_back(total);
      total += w[i];
    }
    if (!(total > 0)) return false;
    // UniformFloat::new(0, total): shrink scale until 0 + scale*max_rand < total
    const double max_rand = 1.0 - 2.220446049250313e-16;   // (u64::MAX>>12 | exp 0) - 1
    scale = total;
    while (!(scale * max_rand + 0.0 < total)) {
      u64 b; memcpy(&b, &scale, 8); b -= 1; memcpy(&scale, &b, 8);
    }
    return true;
  }
  size_t sample(Xoshiro256SS& rng) const {
    u64 bits = (rng.next_u64() >> 12) | 0x3ff0000000000000ULL;
    double v12; memcpy(&v12, &bits, 8);
    volatile double v01 = v12 - 1.0;
    volatile double prod = v01 * scale;             // no FMA contraction
    double chosen = prod + 0.0;
    size_t i = 0;
    while (i < cum.size() && cum[i] <= chosen) i++;  // partition_point(|w| w <= chosen)
    return i;
  }
};

// ---------------------------------------------------------------------------------------------
// Graph container.  ASSUMED policy of the external `flag-complex` crate (cannot be checked):
//   edges live in an insertion-ordered index set (indexmap::IndexSet, a leftover dependency in
//   Cargo.toml:16-17); sample_edge = element at gen_range(0..len); remove = swap_remove (last
//   element moves into the hole); add = append.  Double edges live in a second index set keyed
//   by the normalised pair [min,max]; sample_double_edge returns that [min,max].
//   An empty set returns None WITHOUT consuming a draw.
// ---------------------------------------------------------------------------------------------
static int g_sample_policy = 0;          // see Graph::at (process-wide: tests set it, run, and set it back)
static inline u64 ekey(Node a, Node b) { return ((u64)a << 32) | b; }

struct EdgeSet {
  std::vector<Edge> items;
  std::unordered_map<u64, u32> index;
  bool contains(Node a, Node b) const { return index.find(ekey(a, b)) != index.end(); }
  bool insert(Node a, Node b) {
    auto r = index.emplace(ekey(a, b), (u32)items.size());
    if (!r.second) return false;
    items.push_back({a, b});
    return true;
  }
  bool swap_remove(Node a, Node b) {
    auto it = index.find(ekey(a, b));
    if (it == index.end()) return false;
    u32 i = it->second;
    index.erase(it);
    Edge last = items.back();
    items.pop_back();
    if (i < items.size()) { items[i] = last; index[ekey(last.a, last.b)] = i; }
    return true;
  }
};

struct Graph {
  u32 n = 0;
  EdgeSet edges, dbl;
  static Graph new_disconnected(u32 n) { Graph g; g.n = n; return g; }
  u32 nnodes() const { return n; }
  bool has_edge(Node f, Node t) const { return edges.contains(f, t); }
  void add_edge(Node f, Node t) {
    edges.insert(f, t);
    if (has_edge(t, f)) dbl.insert(std::min(f, t), std::max(f, t));
  }
  void remove_edge(Node f, Node t) {
    edges.swap_remove(f, t);
    dbl.swap_remove(std::min(f, t), std::max(f, t));
  }
  void set_edge(Node f, Node t, bool add) { if (add) add_edge(f, t); else remove_edge(f, t); }
  // THE POLICY SEAM: which element a draw r = gen_range(0..len) names.  0 = element r (the documented assumption),
  // 1 = element len-1-r (exists to show that the CUDA path follows the oracle when the policy changes; orc_set_policy)
  static size_t at(size_t r, size_t len) { return g_sample_policy ? len - 1 - r : r; }
  bool sample_edge(Xoshiro256SS& rng, Edge& out) const {
    if (edges.items.empty()) return false;
    out = edges.items[at(gen_range_u64(rng, edges.items.size()), edges.items.size())];
    return true;
  }
  bool sample_double_edge(Xoshiro256SS& rng, Edge& out) const {
    if (dbl.items.empty()) return false;
    out = dbl.items[at(gen_range_u64(rng, dbl.items.size()), dbl.items.size())];
    return true;
  }
  // [a,b] with a > b, once per adjacent pair (asserted at src/lib.rs:344)
  std::vector<Edge> undirected_edges() const {
    std::vector<Edge> r;
    for (const Edge& e : edges.items) {
      if (e.a > e.b) r.push_back(e);
      else if (!has_edge(e.b, e.a)) r.push_back({e.b, e.a});
    }
    return r;
  }
  // Graph::subgraph(&g, nodes): induced subgraph, vertex nodes[i] -> i  (K^2 has_edge probes,
  // fresh heap-allocated container: the reference's shape, src/lib.rs:63,71)
  static Graph subgraph(const Graph& g, const std::vector<Node>& nodes) {
    Graph s = new_disconnected((u32)nodes.size());
    for (u32 i = 0; i < nodes.size(); i++)
      for (u32 j = 0; j < nodes.size(); j++)
        if (i != j && g.has_edge(nodes[i], nodes[j])) s.add_edge(i, j);
    return s;
  }
};

// ---------------------------------------------------------------------------------------------
// flagser_count: per-dimension number of directed simplices
// ---------------------------------------------------------------------------------------------
struct BitRows {
  u32 n, W;
  std::vector<u64> w;
  BitRows(u32 n_) : n(n_), W((n_ + 63) / 64), w((size_t)n_ * ((n_ + 63) / 64), 0) {}
  void set(u32 a, u32 b) { w[(size_t)a * W + (b >> 6)] |= 1ULL << (b & 63); }
  const u64* row(u32 a) const { return &w[(size_t)a * W]; }
};

static void count_rec(const BitRows& m, const u64* P, int d, std::vector<u64>& cnt,
                      std::vector<std::vector<u64>>& scratch) {
  // every w in P closes a d-simplex; recurse into P ∩ out(w)
  while ((int)scratch.size() <= d) scratch.emplace_back(m.W);
  for (u32 k = 0; k < m.W; k++) {
    u64 bits = P[k];
    while (bits) {
      u32 w = k * 64 + __builtin_ctzll(bits);
      bits &= bits - 1;
      if ((int)cnt.size() <= d) cnt.resize(d + 1, 0);
      cnt[d]++;
      const u64* rw = m.row(w);
      u64 any = 0;
      u64* Q = scratch[d].data();
      for (u32 j = 0; j < m.W; j++) { Q[j] = P[j] & rw[j]; any |= Q[j]; }
      if (any) { count_rec(m, Q, d + 1, cnt, scratch); Q = scratch[d].data(); }
    }
  }
}

static std::vector<u64> count_rows(const BitRows& m, int nthreads) {
  std::vector<u64> total;
  if (m.n == 0) return total;
  if (nthreads <= 1) {
    total.assign(1, 0);
    std::vector<std::vector<u64>> scratch;
    for (u32 v = 0; v < m.n; v++) { total[0]++; count_rec(m, m.row(v), 1, total, scratch); }
    return total;
  }
  std::vector<std::vector<u64>> part(nthreads);
  std::atomic<u32> next(0);
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++)
    th.emplace_back([&, t]() {
      std::vector<u64>& c = part[t];
      c.assign(1, 0);
      std::vector<std::vector<u64>> scratch;
      for (;;) {
        u32 v = next.fetch_add(1);
        if (v >= m.n) break;
        c[0]++;
        count_rec(m, m.row(v), 1, c, scratch);
      }
    });
  for (auto& t : th) t.join();
  for (auto& c : part) {
    if (c.size() > total.size()) total.resize(c.size(), 0);
    for (size_t i = 0; i < c.size(); i++) total[i] += c[i];
  }
  return total;
}

static std::vector<u64> flagser_count(const Graph& g, int nthreads = 1) {
  BitRows m(g.n);
  for (const Edge& e : g.edges.items) m.set(e.a, e.b);
  return count_rows(m, nthreads);
}

// ---------------------------------------------------------------------------------------------
// util.rs
// ---------------------------------------------------------------------------------------------
static std::vector<Node> intersect_sorted(const std::vector<Node>& a, const std::vector<Node>& b) {
  std::vector<Node> out;                                        // src/util.rs:5-26
  size_t ai = 0, bi = 0;
  while (ai < a.size() && bi < b.size()) {
    if (a[ai] == b[bi]) { out.push_back(a[ai]); ai++; bi++; }
    else if (a[ai] < b[bi]) ai++;
    else bi++;
  }
  return out;
}

static bool all_le(const std::vector<u64>& a, const std::vector<u64>& b) {   // src/util.rs:53-63, z = 0
  size_t maxlen = std::max(a.size(), b.size());
  for (size_t i = 0; i < maxlen; i++) {
    u64 l = i < a.size() ? a[i] : 0, r = i < b.size() ? b[i] : 0;
    if (l > r) return false;
  }
  return true;
}

static u64 factorial(u64 x) {          // src/util.rs:65-71: loop 1..x  => (x-1)!  (quirk kept)
  u64 res = 1;
  for (u64 i = 1; i < x; i++) res *= i;
  return res;
}
static u64 binomial(u64 n, u64 k) { return factorial(n) / (factorial(k) * factorial(n - k)); }  // :73-77

static const u64 OEIS_A058298[64] = {                                       // src/util.rs:98-105
    2, 3, 6, 8, 12, 24, 30, 40, 60, 120, 144, 180, 240, 360, 720, 840, 1008, 1260, 1680, 2520, 5040,
    5760, 6720, 8064, 10080, 13440, 20160, 40320, 45360, 51840, 60480, 72576, 90720, 120960, 181440,
    362880, 403200, 453600, 518400, 604800, 725760, 907200, 1209600, 1814400, 3628800, 3991680,
    4435200, 4989600, 5702400, 6652800, 7983360, 9979200, 13305600, 19958400, 39916800, 43545600,
    47900160, 53222400, 59875200, 68428800, 79833600, 95800320, 119750400, 159667200};

// returns false where the reference would panic (table index out of range)
static bool calc_relax_de(const std::vector<u64>& sc, std::vector<u64>& relax_de) {  // src/util.rs:79-93
  relax_de.clear();
  for (size_t d = 0; d < sc.size(); d++) {
    size_t ind = 1;
    u64 best = 0; bool any = false;
    for (;;) {
      if (ind >= 64) return false;
      if (!(OEIS_A058298[ind] < sc[d])) break;
      u64 lost = OEIS_A058298[ind] - OEIS_A058298[ind - 1];
      best = any ? std::max(best, lost) : lost; any = true;
      ind++;
    }
    u64 a = any ? best : 1;
    u64 b = factorial(d + 1);
    relax_de.push_back(std::min(a, b));
  }
  return true;
}

// ---------------------------------------------------------------------------------------------
// State / Transition / Bounds / MCMCSampler  (src/lib.rs)
// ---------------------------------------------------------------------------------------------
// Graph::compute_maximal_cliques of the external flag-complex crate (called src/lib.rs:41): maximal
// cliques of the underlying undirected graph.  ASSUMED output order (the crate is not on disk, parity
// unpinned): every clique's vertices ascending, the cliques in lexicographic order.  Bron-Kerbosch
// with a pivot; the order is imposed by the final sort, not by the enumeration.
static void bron_kerbosch(const std::vector<std::vector<Node>>& adj, std::vector<Node>& R, std::vector<Node> P,
                          std::vector<Node> X, std::vector<std::vector<Node>>& out) {
  if (P.empty()) { if (X.empty()) { std::vector<Node> c = R; std::sort(c.begin(), c.end()); out.push_back(c); } return; }
  Node pivot = P[0]; size_t best = 0; bool first = true;
  auto common = [&](const std::vector<Node>& S, Node u) { size_t c = 0; for (Node v : S) c += std::binary_search(adj[u].begin(), adj[u].end(), v); return c; };
  for (const std::vector<Node>* S : {&P, &X})
    for (Node u : *S) { size_t c = common(P, u); if (first || c > best) { best = c; pivot = u; first = false; } }
  std::vector<Node> cand;
  for (Node v : P) if (!std::binary_search(adj[pivot].begin(), adj[pivot].end(), v)) cand.push_back(v);
  for (Node v : cand) {
    std::vector<Node> P2, X2;
    for (Node w : P) if (std::binary_search(adj[v].begin(), adj[v].end(), w)) P2.push_back(w);
    for (Node w : X) if (std::binary_search(adj[v].begin(), adj[v].end(), w)) X2.push_back(w);
    R.push_back(v);
    bron_kerbosch(adj, R, P2, X2, out);
    R.pop_back();
    P.erase(std::find(P.begin(), P.end(), v));
    X.push_back(v);
  }
}
static std::vector<std::vector<Node>> compute_maximal_cliques(const Graph& g) {
  std::vector<std::vector<Node>> adj(g.nnodes());
  for (const Edge& e : g.edges.items) { adj[e.a].push_back(e.b); adj[e.b].push_back(e.a); }
  for (auto& v : adj) { std::sort(v.begin(), v.end()); v.erase(std::unique(v.begin(), v.end()), v.end()); }
  std::vector<Node> R, P, X;
  for (Node v = 0; v < g.nnodes(); v++) P.push_back(v);
  std::vector<std::vector<Node>> out;
  bron_kerbosch(adj, R, P, X, out);
  std::sort(out.begin(), out.end());
  return out;
}

struct Change { Edge e; bool add; };
struct Transition { std::vector<Change> change_edges; };

struct State {
  std::vector<std::vector<std::vector<Node>>> cliques_by_order;   // [order-1] -> cliques (src/lib.rs:41-49)
  std::unordered_map<u64, std::vector<Node>> edge_neighborhood;   // key [big,small]
  Graph graph;
  std::vector<u64> flag_count;

  static std::unordered_map<u64, std::vector<Node>> compute_edge_neighborhoods(const Graph& g) {
    std::vector<std::vector<Node>> adj(g.nnodes());               // src/lib.rs:331-356
    for (const Edge& e : g.edges.items) { adj[e.a].push_back(e.b); adj[e.b].push_back(e.a); }
    for (auto& v : adj) std::sort(v.begin(), v.end());
    std::unordered_map<u64, std::vector<Node>> m;
    auto und = g.undirected_edges();
    m.reserve(und.size());
    for (const Edge& e : und) m.emplace(ekey(e.a, e.b), intersect_sorted(adj[e.a], adj[e.b]));
    return m;
  }
  static State create(Graph g, int nthreads, bool with_cliques = false) {                    // State::new, src/lib.rs:38-58
    State s;
    if (with_cliques)                                             // src/lib.rs:41-49
      for (auto& c : compute_maximal_cliques(g)) {
        if (c.size() > s.cliques_by_order.size()) s.cliques_by_order.resize(c.size());
        s.cliques_by_order[c.size() - 1].push_back(c);
      }
    s.flag_count = flagser_count(g, nthreads);
    s.edge_neighborhood = compute_edge_neighborhoods(g);
    s.graph = std::move(g);
    return s;
  }
  std::vector<Node> edgeset_neighborhood(const std::vector<Edge>& edges) const {   // :99-111
    std::vector<Node> av;
    for (const Edge& e : edges) {
      Node big = std::max(e.a, e.b), small = std::min(e.a, e.b);
      auto it = edge_neighborhood.find(ekey(big, small));
      if (it == edge_neighborhood.end()) { fprintf(stderr, "oracle: pair not adjacent (reference would panic)\n"); abort(); }
      av.insert(av.end(), it->second.begin(), it->second.end());
      av.push_back(e.a);
      av.push_back(e.b);
    }
    std::sort(av.begin(), av.end());
    av.erase(std::unique(av.begin(), av.end()), av.end());
    return av;
  }
  void apply_transition(const Transition& t, std::vector<u64>& pre, std::vector<u64>& post,
                        u32* K_out = nullptr) {                   // :61-79
    std::vector<Edge> pairs;
    for (const Change& c : t.change_edges) pairs.push_back({std::max(c.e.a, c.e.b), std::min(c.e.a, c.e.b)});
    std::vector<Node> nbhd = edgeset_neighborhood(pairs);
    if (K_out) *K_out = (u32)nbhd.size();
    pre = flagser_count(Graph::subgraph(graph, nbhd));
    for (size_t i = 0; i < pre.size() && i < flag_count.size(); i++) {
      if (flag_count[i] < pre[i]) { fprintf(stderr, "oracle: assert s >= p failed\n"); abort(); }
      flag_count[i] -= pre[i];
    }
    for (const Change& c : t.change_edges) graph.set_edge(c.e.a, c.e.b, c.add);
    post = flagser_count(Graph::subgraph(graph, nbhd));
    if (post.size() > flag_count.size()) flag_count.resize(post.size(), 0);
    for (size_t i = 0; i < post.size(); i++) flag_count[i] += post[i];
  }
  void revert_transition(const Transition& t, const std::vector<u64>& pre, const std::vector<u64>& post) {  // :81-95
    for (const Change& c : t.change_edges) graph.set_edge(c.e.a, c.e.b, !c.add);
    for (size_t i = 0; i < post.size() && i < flag_count.size(); i++) {
      if (flag_count[i] < post[i]) { fprintf(stderr, "oracle: assert s >= p failed\n"); abort(); }
      flag_count[i] -= post[i];
    }
    if (pre.size() > flag_count.size()) flag_count.resize(pre.size(), 0);
    for (size_t i = 0; i < pre.size(); i++) flag_count[i] += pre[i];
  }
};

struct Bounds {
  std::vector<u64> flag_count_min, flag_count_max;
  // returns 0 ok, <0 where the reference would panic
  static int calculate(const State& initial, const Bounds& target, Bounds& out, int nthreads,
                       std::vector<u64>* ncliques_out = nullptr) {                  // :119-156
    auto und = initial.graph.undirected_edges();
    Graph norm = Graph::new_disconnected(initial.graph.nnodes());
    for (const Edge& e : und) norm.add_edge(e.a, e.b);
    std::vector<u64> ncliques = flagser_count(norm, nthreads);
    if (ncliques_out) *ncliques_out = ncliques;
    if (initial.flag_count.size() < 2) return -1;
    if (und.size() == initial.flag_count[1]) {                                     // SEO case
      out.flag_count_min = target.flag_count_min;
      out.flag_count_max = ncliques;
      return 0;
    }
    std::vector<u64> mn = target.flag_count_min, mx = target.flag_count_max;
    std::vector<u64> relax_de;
    if (!calc_relax_de(initial.flag_count, relax_de)) return -2;
    size_t len = initial.flag_count.size();
    for (size_t d = 2; d < len; d++) {
      u64 dense = binomial(len - 2, d - 1);
      u64 relax = relax_de[d] * dense;
      mx[d] = std::max(mn[d] + relax, mx[d]);
      mn[d] = std::min(mx[d] - relax, mn[d]);
    }
    if (mx.size() < 3) return -3;
    mx[2] = UINT64_MAX;
    mx.push_back(10);
    out.flag_count_min = mn;
    out.flag_count_max = mx;
    return 0;
  }
  bool check(const State& s) const {                                               // :157-160
    return all_le(flag_count_min, s.flag_count) && all_le(s.flag_count, flag_count_max);
  }
};

// trace record shared with the CUDA path (include/fcm.h: fcm_trace_rec)
struct TraceRec { u32 flags, a, b, dm0, dm1, K, draws, pad; };
enum { TR_KIND_MASK = 3, TR_ACCEPT = 1 << 4, TR_EMPTY = 1 << 5 };

struct Sampler {
  Xoshiro256SS rng;
  State state;
  Bounds bounds;
  WeightedIndexF64 move_distribution, clique_order_distribution;
  u64 sample_distance = 0, sampled = 0, accepted = 0;
  u64 draws = 0;

  // Transition::single_edge_flip  src/lib.rs:292-299
  Transition single_edge_flip(TraceRec* tr) {
    Edge e;
    if (state.graph.sample_edge(rng, e)) {
      if (tr) { tr->a = e.a; tr->b = e.b; }
      if (!state.graph.has_edge(e.b, e.a))
        return Transition{{{{e.a, e.b}, false}, {{e.b, e.a}, true}}};
    }
    return Transition{};
  }
  // Transition::double_edge_move  src/lib.rs:304-325
  Transition double_edge_move(TraceRec* tr) {
    Edge de;
    if (state.graph.sample_double_edge(rng, de)) {
      Edge s{0, 0};
      for (;;) {
        state.graph.sample_edge(rng, s);
        if (!state.graph.has_edge(s.b, s.a)) break;
      }
      Edge delme = gen_bool(rng, 0.5) ? de : Edge{de.b, de.a};
      if (tr) { tr->a = s.a; tr->b = s.b; tr->dm0 = delme.a; tr->dm1 = delme.b; }
      return Transition{{{{s.b, s.a}, true}, {delme, false}}};
    }
    return Transition{};
  }
  // util::vec_intersect / vec_setminus (src/util.rs:34-50): order of the first argument is kept
  static std::vector<Node> vec_intersect(const std::vector<Node>& xs, const std::vector<Node>& ys) {
    std::vector<Node> r;
    for (Node x : xs) if (std::find(ys.begin(), ys.end(), x) != ys.end()) r.push_back(x);
    return r;
  }
  static std::vector<Node> vec_setminus(const std::vector<Node>& xs, const std::vector<Node>& ys) {
    std::vector<Node> r;
    for (Node x : xs) if (std::find(ys.begin(), ys.end(), x) == ys.end()) r.push_back(x);
    return r;
  }
  // Transition::clique_permute  src/lib.rs:214-233
  Transition clique_permute(TraceRec* tr) {
    const size_t ord = clique_order_distribution.sample(rng);
    const auto& cliques = state.cliques_by_order[ord];
    if (cliques.empty()) { fprintf(stderr, "oracle: choose() on an empty clique list (reference unwrap panics)\n"); abort(); }
    const size_t ci = gen_index(rng, cliques.size());             // SliceRandom::choose
    const std::vector<Node>& cl = cliques[ci];
    std::vector<size_t> perm = random_perm(0, cl.size(), rng);
    Transition t;
    for (size_t i = 0; i < cl.size(); i++)
      for (size_t j = 0; j < cl.size(); j++) {
        bool pre = state.graph.has_edge(cl[perm[i]], cl[perm[j]]);
        bool post = state.graph.has_edge(cl[i], cl[j]);
        if (pre != post) t.change_edges.push_back({{cl[perm[i]], cl[perm[j]]}, post});
      }
    if (tr) { tr->a = (u32)ci; tr->b = 0; tr->dm0 = (u32)ord; tr->dm1 = (u32)t.change_edges.size(); }
    return t;
  }
  // Transition::clique_swap  src/lib.rs:235-290
  Transition clique_swap(TraceRec* tr) {
    const size_t ord = clique_order_distribution.sample(rng);
    const auto& cliques = state.cliques_by_order[ord];
    if (cliques.empty()) { fprintf(stderr, "oracle: choose() on an empty clique list (reference unwrap panics)\n"); abort(); }
    const size_t i1 = gen_index(rng, cliques.size());
    const size_t i2 = gen_index(rng, cliques.size());
    const std::vector<Node>&m1 = cliques[i1], &m2 = cliques[i2];
    std::vector<Node> c = vec_intersect(m1, m2);
    std::vector<Node> d = c;
    for (Node x : vec_setminus(m1, c)) d.push_back(x);
    for (Node x : vec_setminus(m2, c)) d.push_back(x);
    const size_t n_c = c.size(), n_d = d.size(), n_a = m1.size() - n_c;
    std::vector<size_t> perm_c = random_perm(0, n_c, rng);
    std::vector<size_t> perm_a = random_perm(n_c, n_c + n_a, rng);
    std::vector<size_t> perm_b = random_perm(n_c + n_a, n_d, rng);
    std::vector<size_t> perm_d = perm_c;
    perm_d.insert(perm_d.end(), perm_b.begin(), perm_b.end());
    perm_d.insert(perm_d.end(), perm_a.begin(), perm_a.end());
    std::vector<std::pair<Node, Node>> new_edges, old_edges;
    auto visit = [&](const std::vector<size_t>& idx) {
      for (size_t i : idx)
        for (size_t j : idx)
          if (state.graph.has_edge(d[i], d[j])) {
            new_edges.push_back({d[perm_d[i]], d[perm_d[j]]});
            old_edges.push_back({d[i], d[j]});
          }
    };
    std::vector<size_t> first, second;
    for (size_t i = 0; i < n_c + n_a; i++) first.push_back(i);
    for (size_t i = 0; i < n_c; i++) second.push_back(i);
    for (size_t i = n_c + n_a; i < n_d; i++) second.push_back(i);
    visit(first);
    visit(second);
    std::sort(new_edges.begin(), new_edges.end());
    new_edges.erase(std::unique(new_edges.begin(), new_edges.end()), new_edges.end());
    std::sort(old_edges.begin(), old_edges.end());
    old_edges.erase(std::unique(old_edges.begin(), old_edges.end()), old_edges.end());
    Transition t;
    for (const auto& ne : new_edges) {
      auto it = std::find(old_edges.begin(), old_edges.end(), ne);
      if (it != old_edges.end()) old_edges.erase(it);             // retain(|e| e != ne): the list is deduplicated
      else t.change_edges.push_back({{ne.first, ne.second}, true});
    }
    for (const auto& oe : old_edges) t.change_edges.push_back({{oe.first, oe.second}, false});
    if (tr) { tr->a = (u32)i1; tr->b = (u32)i2; tr->dm0 = (u32)ord; tr->dm1 = (u32)t.change_edges.size(); }
    return t;
  }
  // one iteration of MCMCSampler::next  src/lib.rs:182-192
  int step(TraceRec* tr) {
    if (tr) memset(tr, 0, sizeof(*tr));
    size_t kind = move_distribution.sample(rng);                  // random_move :207-212
    Transition t;
    if (kind == 0) t = single_edge_flip(tr);
    else if (kind == 1) t = double_edge_move(tr);
    else if (state.cliques_by_order.empty()) return -1;           // built without cliques
    else if (kind == 2) t = clique_permute(tr);
    else t = clique_swap(tr);
    std::vector<u64> pre, post;
    u32 K = 0;
    state.apply_transition(t, pre, post, &K);
    sampled++;
    bool ok = bounds.check(state);
    if (ok) accepted++;
    else state.revert_transition(t, pre, post);
    if (tr) {
      tr->flags = (u32)kind | (ok ? TR_ACCEPT : 0) | (t.change_edges.empty() ? TR_EMPTY : 0);
      tr->K = K;
    }
    return 0;
  }
};

// io::read_flag_file  src/io.rs:18-35 : line 1 skipped, line 2 token count = n, line 3 skipped,
// then "i j [w]" (weight ignored), lines with < 2 tokens ignored.  Lines as str::lines() (:23) gives them
// ('\n' ends a line, one '\r' before it is dropped); tokens split on ' ' only (:25,:29); a token that is not a
// u32 makes the reference panic (`parse().unwrap()`, :31) -> false here.
static bool parse_u32_token(const std::string& w, Node& out) {
  size_t i = (!w.empty() && w[0] == '+') ? 1 : 0;
  if (i >= w.size()) return false;
  u64 v = 0;
  for (; i < w.size(); i++) {
    if (w[i] < '0' || w[i] > '9') return false;
    v = v * 10 + (u64)(w[i] - '0');
    if (v > 0xffffffffULL) return false;
  }
  out = (Node)v;
  return true;
}
static bool read_flag_file(const char* path, u32& n, std::vector<Edge>& edges) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::string line;
  auto next_line = [&](std::string& out) {
    if (!std::getline(f, out, '\n')) return false;
    if (!out.empty() && out.back() == '\r') out.pop_back();
    return true;
  };
  auto tokens = [](const std::string& s) {
    std::vector<std::string> t; std::string cur;
    for (char c : s) { if (c == ' ') { if (!cur.empty()) { t.push_back(cur); cur.clear(); } } else cur.push_back(c); }
    if (!cur.empty()) t.push_back(cur);
    return t;
  };
  if (!next_line(line)) return false;
  if (!next_line(line)) return false;
  n = (u32)tokens(line).size();
  next_line(line);
  edges.clear();
  while (next_line(line)) {
    auto t = tokens(line);
    if (t.size() < 2) continue;
    Edge e;
    if (!parse_u32_token(t[0], e.a) || !parse_u32_token(t[1], e.b)) return false;
    edges.push_back(e);
  }
  return true;
}

// =============================================================================================
// C ABI for tests/ and bench.py (ctypes)
// =============================================================================================
#define ORC_MAXD 32
extern "C" {

int orc_read_flag_header(const char* path, u32* n, u64* nedges) {
  std::vector<Edge> e; u32 nn;
  if (!read_flag_file(path, nn, e)) return -1;
  *n = nn; *nedges = e.size(); return 0;
}
int orc_read_flag(const char* path, u32* edges_out /* 2*nedges */) {
  std::vector<Edge> e; u32 nn;
  if (!read_flag_file(path, nn, e)) return -1;
  for (size_t i = 0; i < e.size(); i++) { edges_out[2 * i] = e[i].a; edges_out[2 * i + 1] = e[i].b; }
  return 0;
}

static Graph graph_from(u32 n, u64 ne, const u32* edges) {
  Graph g = Graph::new_disconnected(n);
  g.edges.items.reserve(ne);
  g.edges.index.reserve(ne);
  for (u64 i = 0; i < ne; i++) g.add_edge(edges[2 * i], edges[2 * i + 1]);
  return g;
}

// flagser_count of an edge list; dense-row recursion, parallel over source vertices
int orc_count(u32 n, u64 ne, const u32* edges, int nthreads, u64* out /*ORC_MAXD*/, int* len) {
  BitRows m(n);
  for (u64 i = 0; i < ne; i++) if (edges[2 * i] != edges[2 * i + 1]) m.set(edges[2 * i], edges[2 * i + 1]);
  // duplicates collapse in the bit matrix exactly as in an edge *set*; count[1] = number of set bits
  std::vector<u64> c = count_rows(m, nthreads);
  if ((int)c.size() > ORC_MAXD) return -1;
  for (size_t i = 0; i < c.size(); i++) out[i] = c[i];
  *len = (int)c.size();
  return 0;
}

// count restricted to the given source vertices (first vertex of the simplex): what one shard of a
// sharded recount computes; the shards' vectors sum to orc_count's.
int orc_count_sources(u32 n, u64 ne, const u32* edges, const u32* sources, u32 nsrc, u64* out /*ORC_MAXD*/) {
  BitRows m(n);
  for (u64 i = 0; i < ne; i++) if (edges[2 * i] != edges[2 * i + 1]) m.set(edges[2 * i], edges[2 * i + 1]);
  std::vector<u64> c(1, 0);
  std::vector<std::vector<u64>> scratch;
  for (u32 i = 0; i < nsrc; i++) { c[0]++; count_rec(m, m.row(sources[i]), 1, c, scratch); }
  if ((int)c.size() > ORC_MAXD) return -1;
  for (int i = 0; i < ORC_MAXD; i++) out[i] = i < (int)c.size() ? c[i] : 0;
  return 0;
}

// same, sources shared dynamically between nthreads host threads (bounded CPU-baseline sample of a
// large recount: bench.py counts every 4th source of the degree-descending order)
int orc_count_sources_mt(u32 n, u64 ne, const u32* edges, const u32* sources, u32 nsrc, int nthreads, u64* out /*ORC_MAXD*/) {
  BitRows m(n);
  for (u64 i = 0; i < ne; i++) if (edges[2 * i] != edges[2 * i + 1]) m.set(edges[2 * i], edges[2 * i + 1]);
  if (nthreads < 1) nthreads = 1;
  std::vector<std::vector<u64>> part(nthreads);
  std::atomic<u32> next(0);
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++)
    th.emplace_back([&, t]() {
      std::vector<u64>& c = part[t];
      c.assign(1, 0);
      std::vector<std::vector<u64>> scratch;
      for (;;) {
        u32 i = next.fetch_add(1);
        if (i >= nsrc) break;
        c[0]++;
        count_rec(m, m.row(sources[i]), 1, c, scratch);
      }
    });
  for (auto& t : th) t.join();
  for (int i = 0; i < ORC_MAXD; i++) out[i] = 0;
  for (auto& c : part) {
    if ((int)c.size() > ORC_MAXD) return -1;
    for (size_t i = 0; i < c.size(); i++) out[i] += c[i];
  }
  return 0;
}

// --- rand known-answer hooks --------------------------------------------------------------
void orc_rng_seed(u64 seed, u64* s4) { Xoshiro256SS r = Xoshiro256SS::seed_from_u64(seed); memcpy(s4, r.s, 32); }
u64 orc_rng_next(u64* s4) { Xoshiro256SS r; memcpy(r.s, s4, 32); u64 v = r.next_u64(); memcpy(s4, r.s, 32); return v; }
u64 orc_rng_gen_range(u64* s4, u64 range) { Xoshiro256SS r; memcpy(r.s, s4, 32); u64 v = gen_range_u64(r, range); memcpy(s4, r.s, 32); return v; }
int orc_rng_gen_bool(u64* s4, double p) { Xoshiro256SS r; memcpy(r.s, s4, 32); int v = gen_bool(r, p); memcpy(s4, r.s, 32); return v; }
int orc_weighted_index(const double* w, int n, u64* s4, double* cum_out, double* scale_out) {
  WeightedIndexF64 wi; if (!wi.init(w, n)) return -1;
  if (cum_out) for (size_t i = 0; i < wi.cum.size(); i++) cum_out[i] = wi.cum[i];
  if (scale_out) *scale_out = wi.scale;
  Xoshiro256SS r; memcpy(r.s, s4, 32); int v = (int)wi.sample(r); memcpy(s4, r.s, 32); return v;
}

// --- util known-answer hooks ----------------------------------------------------------------
u64 orc_factorial(u64 x) { return factorial(x); }
u64 orc_binomial(u64 n, u64 k) { return binomial(n, k); }
int orc_calc_relax_de(const u64* sc, int len, u64* out) {
  std::vector<u64> v(sc, sc + len), r; if (!calc_relax_de(v, r)) return -1;
  for (int i = 0; i < len; i++) out[i] = r[i];
  return 0;
}
int orc_all_le(const u64* a, int na, const u64* b, int nb) {
  return all_le(std::vector<u64>(a, a + na), std::vector<u64>(b, b + nb));
}
int orc_intersect_sorted(const u32* a, int na, const u32* b, int nb, u32* out) {
  auto r = intersect_sorted(std::vector<Node>(a, a + na), std::vector<Node>(b, b + nb));
  for (size_t i = 0; i < r.size(); i++) out[i] = r[i];
  return (int)r.size();
}

// Rust `f64 as usize`: saturating, NaN -> 0 (the casts of src/bin/sample.rs:90-91)
static inline u64 f64_as_usize(double x) {
  if (!(x > 0.0)) return 0;
  if (x >= 18446744073709551616.0) return ~0ULL;
  return (u64)x;
}

// --- sampler ----------------------------------------------------------------------------------
// mirrors initialize_new_sampler (src/bin/sample.rs:80-105).  move_weights = 4 doubles
// ([0.5,0.5,0,0] for --simple).  sample_distance 0 => ceil(2 E log2 E).
void* orc_sampler_new(u32 n, u64 ne, const u32* edges, u64 seed, double target_relaxation,
                      const double* move_weights, u64 sample_distance, int nthreads, int* err) {
  *err = 0;
  Sampler* s = new Sampler();
  const bool with_cliques = move_weights[2] != 0.0 || move_weights[3] != 0.0;
  s->state = State::create(graph_from(n, ne, edges), nthreads, with_cliques);
  if (with_cliques) {                                             // src/bin/sample.rs:87-88
    std::vector<double> w;
    for (const auto& cs : s->state.cliques_by_order) w.push_back(std::pow((double)cs.size(), 0.2));
    if (!s->clique_order_distribution.init(w.data(), w.size())) { *err = -11; delete s; return nullptr; }
  }
  s->rng = Xoshiro256SS::seed_from_u64(seed);
  Bounds target;
  for (size_t d = 0; d < s->state.flag_count.size(); d++) {
    u64 sc = s->state.flag_count[d];
    target.flag_count_min.push_back(d < 2 ? sc : f64_as_usize(std::floor((double)sc * (1. - target_relaxation))));
    target.flag_count_max.push_back(d < 2 ? sc : f64_as_usize(std::floor((double)sc * (1. + target_relaxation))));
  }
  int rc = Bounds::calculate(s->state, target, s->bounds, nthreads);
  if (rc) { *err = rc; delete s; return nullptr; }
  if (!s->move_distribution.init(move_weights, 4)) { *err = -10; delete s; return nullptr; }
  if (sample_distance == 0) {
    double E = (double)s->state.flag_count[1];
    double sd = std::ceil(2. * E * std::log2(E));
    sample_distance = (sd != sd || sd < 0) ? 0 : (u64)sd;
  }
  s->sample_distance = sample_distance;
  return s;
}
void orc_sampler_free(void* h) { delete (Sampler*)h; }
void orc_set_policy(int policy) { g_sample_policy = policy ? 1 : 0; }
void orc_sampler_set_bounds(void* h, const u64* mn, int nmin, const u64* mx, int nmax) {
  Sampler* s = (Sampler*)h;
  s->bounds.flag_count_min.assign(mn, mn + nmin);
  s->bounds.flag_count_max.assign(mx, mx + nmax);
}
// nsteps proposals; trace may be null (else nsteps records)
int orc_sampler_steps(void* h, u64 nsteps, TraceRec* trace) {
  Sampler* s = (Sampler*)h;
  for (u64 i = 0; i < nsteps; i++) { int rc = s->step(trace ? trace + i : nullptr); if (rc) return rc; }
  return 0;
}
// MCMCSampler::next: sample_distance proposals
int orc_sampler_next(void* h) { Sampler* s = (Sampler*)h; return orc_sampler_steps(h, s->sample_distance, nullptr); }
void orc_sampler_info(void* h, u64* sample_distance, u64* sampled, u64* accepted, u64* nedges, u64* ndouble,
                      int* fc_len, int* min_len, int* max_len) {
  Sampler* s = (Sampler*)h;
  *sample_distance = s->sample_distance; *sampled = s->sampled; *accepted = s->accepted;
  *nedges = s->state.graph.edges.items.size(); *ndouble = s->state.graph.dbl.items.size();
  *fc_len = (int)s->state.flag_count.size();
  *min_len = (int)s->bounds.flag_count_min.size(); *max_len = (int)s->bounds.flag_count_max.size();
}
// buffers sized by orc_sampler_info; any may be null
void orc_sampler_read(void* h, u64* flag_count, u64* bmin, u64* bmax, u32* edges /*2*E, container order*/,
                      u32* dbl /*2*ND*/, u64* rng4) {
  Sampler* s = (Sampler*)h;
  if (flag_count) for (size_t i = 0; i < s->state.flag_count.size(); i++) flag_count[i] = s->state.flag_count[i];
  if (bmin) for (size_t i = 0; i < s->bounds.flag_count_min.size(); i++) bmin[i] = s->bounds.flag_count_min[i];
  if (bmax) for (size_t i = 0; i < s->bounds.flag_count_max.size(); i++) bmax[i] = s->bounds.flag_count_max[i];
  if (edges) for (size_t i = 0; i < s->state.graph.edges.items.size(); i++) {
    edges[2 * i] = s->state.graph.edges.items[i].a; edges[2 * i + 1] = s->state.graph.edges.items[i].b; }
  if (dbl) for (size_t i = 0; i < s->state.graph.dbl.items.size(); i++) {
    dbl[2 * i] = s->state.graph.dbl.items[i].a; dbl[2 * i + 1] = s->state.graph.dbl.items[i].b; }
  if (rng4) memcpy(rng4, s->rng.s, 32);
}
// State::apply_transition / revert_transition with explicit edits (a,b,add) triples
int orc_apply(void* h, const u32* edits3, int nedits, u64* pre, int* prelen, u64* post, int* postlen) {
  Sampler* s = (Sampler*)h;
  Transition t;
  for (int i = 0; i < nedits; i++) t.change_edges.push_back({{edits3[3 * i], edits3[3 * i + 1]}, edits3[3 * i + 2] != 0});
  std::vector<u64> p, q;
  s->state.apply_transition(t, p, q);
  for (size_t i = 0; i < p.size(); i++) pre[i] = p[i];
  for (size_t i = 0; i < q.size(); i++) post[i] = q[i];
  *prelen = (int)p.size(); *postlen = (int)q.size();
  return 0;
}
int orc_revert(void* h, const u32* edits3, int nedits, const u64* pre, int prelen, const u64* post, int postlen) {
  Sampler* s = (Sampler*)h;
  Transition t;
  for (int i = 0; i < nedits; i++) t.change_edges.push_back({{edits3[3 * i], edits3[3 * i + 1]}, edits3[3 * i + 2] != 0});
  s->state.revert_transition(t, std::vector<u64>(pre, pre + prelen), std::vector<u64>(post, post + postlen));
  return 0;
}
int orc_bounds_check(void* h) { Sampler* s = (Sampler*)h; return s->bounds.check(s->state); }
// maximal cliques in the assumed order: sizes_out[ncliques], verts_out[sum sizes] (either may be null) -> ncliques
u64 orc_maximal_cliques(u32 n, u64 ne, const u32* edges, u32* sizes_out, u32* verts_out, u64* total_verts) {
  auto cl = compute_maximal_cliques(graph_from(n, ne, edges));
  u64 tot = 0;
  for (size_t i = 0; i < cl.size(); i++) {
    if (sizes_out) sizes_out[i] = (u32)cl[i].size();
    for (Node v : cl[i]) { if (verts_out) verts_out[tot] = v; tot++; }
  }
  if (total_verts) *total_verts = tot;
  return cl.size();
}

// undirected clique counts of the normalised graph (Bounds::calculate's `ncliques`)
int orc_ncliques(u32 n, u64 ne, const u32* edges, int nthreads, u64* out, int* len) {
  Graph g = graph_from(n, ne, edges);
  auto und = g.undirected_edges();
  BitRows m(n);
  for (const Edge& e : und) m.set(e.a, e.b);
  auto c = count_rows(m, nthreads);
  for (size_t i = 0; i < c.size(); i++) out[i] = c[i];
  *len = (int)c.size();
  return 0;
}
// common-neighbour list sizes (for K statistics in bench): returns number of pairs
u64 orc_nbhd_stats(void* h, u64* sum_len, u32* max_len) {
  Sampler* s = (Sampler*)h; u64 sum = 0; u32 mx = 0;
  for (auto& kv : s->state.edge_neighborhood) { sum += kv.second.size(); mx = std::max<u32>(mx, kv.second.size()); }
  *sum_len = sum; *max_len = mx; return s->state.edge_neighborhood.size();
}

// io::BitOutput (src/io.rs:128-212).  new(): the slot list = edges of the initial graph plus their
// reverses, sorted by the key (max(a,b), min(a,b), a < b), deduplicated (:152-159).  save(): one bit
// per slot, set iff the current graph has that directed edge, packed LSB-first into bytes, a last
// partial byte is written as it stands (:180-194).
// orc_edgebits_slots: slots_out[2*nslots] (pass null to get the count only) -> nslots
u64 orc_edgebits_slots(u64 ne0, const u32* edges0, u32* slots_out) {
  std::vector<std::array<u32, 2>> ed;
  ed.reserve(2 * ne0);
  for (u64 i = 0; i < ne0; i++) ed.push_back({edges0[2 * i], edges0[2 * i + 1]});
  for (u64 i = 0; i < ne0; i++) ed.push_back({edges0[2 * i + 1], edges0[2 * i]});
  auto key = [](const std::array<u32, 2>& e) { return std::make_tuple(std::max(e[0], e[1]), std::min(e[0], e[1]), e[0] < e[1]); };
  std::sort(ed.begin(), ed.end(), [&](const std::array<u32, 2>& x, const std::array<u32, 2>& y) { return key(x) < key(y); });
  ed.erase(std::unique(ed.begin(), ed.end()), ed.end());
  if (slots_out) for (size_t i = 0; i < ed.size(); i++) { slots_out[2 * i] = ed[i][0]; slots_out[2 * i + 1] = ed[i][1]; }
  return ed.size();
}
// orc_edgebits_save: one sample; out holds ceil(nslots/8) bytes
void orc_edgebits_save(u64 nslots, const u32* slots, u64 ne, const u32* edges, unsigned char* out) {
  std::unordered_set<u64> have;
  have.reserve(ne * 2);
  for (u64 i = 0; i < ne; i++) have.insert(((u64)edges[2 * i] << 32) | edges[2 * i + 1]);
  unsigned char bit = 1, byte = 0;
  u64 o = 0;
  for (u64 i = 0; i < nslots; i++) {
    if (have.count(((u64)slots[2 * i] << 32) | slots[2 * i + 1])) byte |= bit;
    bit = (unsigned char)((bit << 1) | (bit >> 7));     // u8::rotate_left(1)
    if (bit == 1) { out[o++] = byte; byte = 0; }
  }
  if (bit != 1) out[o++] = byte;
}

// flag_complex::for_each_cell(&G, &mut F, min, max) (external crate; used src/bin/all_cxs.rs:60-64,79-83):
// the flagser recursion calling F with the vertex list of every cell of dimension min..max, sources
// and candidates in ascending order (ASSUMED callback order: the crate is not on disk).
// rows of (max + 1) ids padded with 0xffffffff; returns the number of cells (rows beyond cap dropped)
static void cells_rec(const BitRows& m, const u64* P, std::vector<u32>& cur, int dmin, int dmax, u32* out, u64 cap, u64& n) {
  const int d = (int)cur.size() - 1;
  if (d >= dmin && d <= dmax) {
    if (n < cap) { for (int t = 0; t <= dmax; t++) out[n * (dmax + 1) + t] = t <= d ? cur[t] : 0xffffffffu; }
    n++;
  }
  if (d >= dmax) return;
  std::vector<u64> Q(m.W);
  for (u32 k = 0; k < m.W; k++) {
    u64 bits = P[k];
    while (bits) {
      u32 w = k * 64 + __builtin_ctzll(bits);
      bits &= bits - 1;
      const u64* rw = m.row(w);
      for (u32 j = 0; j < m.W; j++) Q[j] = P[j] & rw[j];
      cur.push_back(w);
      cells_rec(m, Q.data(), cur, dmin, dmax, out, cap, n);
      cur.pop_back();
    }
  }
}
u64 orc_cells(u32 n, u64 ne, const u32* edges, int dmin, int dmax, u32* out, u64 cap) {
  Graph g = graph_from(n, ne, edges);
  BitRows m(g.n);
  for (const Edge& e : g.edges.items) m.set(e.a, e.b);
  u64 cnt = 0;
  std::vector<u32> cur;
  for (u32 v = 0; v < n; v++) { cur.assign(1, v); cells_rec(m, m.row(v), cur, dmin, dmax, out, cap, cnt); }
  return cnt;
}

// State::compute_edge_neighborhoods (src/lib.rs:331-356) for a graph: pairs [big, small] sorted, CSR lists.
// pairs_out[2*npairs], off_out[npairs+1], list_out[total] (any may be null) -> npairs; *total_out = list length
u64 orc_edge_neighborhoods(u32 n, u64 ne, const u32* edges, u32* pairs_out, u64* off_out, u32* list_out, u64* total_out) {
  Graph g = graph_from(n, ne, edges);
  auto m = State::compute_edge_neighborhoods(g);
  std::vector<u64> keys;
  for (auto& kv : m) keys.push_back(kv.first);
  std::sort(keys.begin(), keys.end());
  u64 tot = 0;
  for (size_t q = 0; q < keys.size(); q++) {
    const auto& l = m[keys[q]];
    if (pairs_out) { pairs_out[2 * q] = (u32)(keys[q] >> 32); pairs_out[2 * q + 1] = (u32)keys[q]; }
    if (off_out) off_out[q] = tot;
    for (Node v : l) { if (list_out) list_out[tot] = v; tot++; }
  }
  if (off_out) off_out[keys.size()] = tot;
  if (total_out) *total_out = tot;
  return keys.size();
}

}  // extern "C"
