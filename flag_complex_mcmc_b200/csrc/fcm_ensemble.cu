// fcm_ensemble.cu — host side of the chain ensemble: builds the static structures
// (undirected pairs, common-neighbour CSR = compute_edge_neighborhoods, src/lib.rs:331-356), the
// initial edge containers in insertion order, uploads them and replicates them per chain.
#include "fcm_internal.cuh"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <thread>

#define REQUIRE(cond, msg)                                 \
  do {                                                     \
    if (!(cond)) { fcm_set_error("%s", msg); return FCM_EINVAL; } \
  } while (0)

static inline u64 pkey(u32 big, u32 small) { return ((u64)big << 32) | small; }

// rand 0.8 SplitMix64 seeding of xoshiro256** (Xoshiro256StarStar::seed_from_u64, src/bin/sample.rs:86)
static void seed_from_u64(u64 seed, u64 s[4]) {
  u64 x = seed;
  for (int i = 0; i < 4; i++) {
    x += 0x9e3779b97f4a7c15ULL;
    u64 z = x;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    s[i] = z ^ (z >> 31);
  }
}

// rand 0.8 WeightedIndex<f64>::new + Uniform<f64>::new(0, total)
static int weighted_index_init(const double w[4], double cum[3], double* scale) {
  double total = w[0];
  for (int i = 0; i < 4; i++) if (!(w[i] >= 0.0) || !std::isfinite(w[i])) { fcm_set_error("move weight %d is invalid", i); return FCM_EINVAL; }
  for (int i = 1; i < 4; i++) { cum[i - 1] = total; total += w[i]; }
  if (!(total > 0.0)) { fcm_set_error("all move weights are zero"); return FCM_EINVAL; }
  const double max_rand = 1.0 - 2.220446049250313e-16;
  double sc = total;
  for (;;) {
    volatile double prod = sc * max_rand;
    if (prod + 0.0 < total) break;
    u64 b; memcpy(&b, &sc, 8); b -= 1; memcpy(&sc, &b, 8);
  }
  *scale = sc;
  return FCM_OK;
}

// Keep the static clique index resident in L2: the chains' own state (GBs of random rows) otherwise competes with the
// ~50 MB index for the 126 MB L2 (measured on the C5 shard: 363.1 -> 372.6 M proposals/s, profiles/README.md).
// FCM_L2_PERSIST=<hit ratio scale, 0 < x <= 1> tunes it, FCM_L2_PERSIST=0 turns it off; default 1.
// The window is handed to every launch of the chain kernel as a launch attribute (fcm_chain.cu), never installed on a
// stream: a caller-owned stream (fcm_ctx_set_stream) is left as it was, and the window follows the context wherever
// it launches.  Process-global side effect: cudaLimitPersistingL2CacheSize is raised to cover the index;
// fcm_ensemble_destroy resets the persisting lines (cudaCtxResetPersistingL2Cache).
static void index_l2_window(fcm_ensemble* e, void* base, size_t bytes) {
  fcm_ctx* ctx = e->ctx;
  memset(&e->l2_window, 0, sizeof(e->l2_window));
  const char* env = getenv("FCM_L2_PERSIST");
  const double scale = env ? atof(env) : 1.0;
  if (!(scale > 0.0) || bytes == 0) return;
  int max_persist = 0, max_win = 0;
  cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, ctx->device);
  cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, ctx->device);
  if (max_persist <= 0 || max_win <= 0) return;
  const size_t carve = std::min(bytes, (size_t)max_persist), win = std::min(bytes, (size_t)max_win);
  if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) != cudaSuccess) { cudaGetLastError(); return; }
  e->l2_window.base_ptr = base;
  e->l2_window.num_bytes = win;
  e->l2_window.hitRatio = (float)std::min(1.0, scale * (double)carve / (double)win);
  e->l2_window.hitProp = cudaAccessPropertyPersisting;
  e->l2_window.missProp = cudaAccessPropertyStreaming;
}

template <typename T>
static int dev_alloc(fcm_ensemble* e, T** p, size_t count) {
  void* q = nullptr;
  cudaError_t err = cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T));
  if (err != cudaSuccess) {
    fcm_set_error("cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(err));
    return err == cudaErrorMemoryAllocation ? FCM_ENOMEM : FCM_ECUDA;
  }
  e->dev_allocs.push_back(q);
  *p = (T*)q;
  return FCM_OK;
}

static int copy_bounds(EnsParams& p, const u64* mn, int nmin, const u64* mx, int nmax) {
  REQUIRE(nmin >= 0 && nmin <= FCM_MAXB && nmax >= 0 && nmax <= FCM_MAXB, "bounds longer than FCM_MAXD+1");
  REQUIRE((nmin == 0 || mn) && (nmax == 0 || mx), "bounds pointer is null");
  memset(p.bmin, 0, sizeof(p.bmin));
  memset(p.bmax, 0, sizeof(p.bmax));
  for (int i = 0; i < nmin; i++) p.bmin[i] = mn[i];
  for (int i = 0; i < nmax; i++) p.bmax[i] = mx[i];
  p.nmin = nmin; p.nmax = nmax;
  return FCM_OK;
}

// rand 0.8 WeightedIndex<f64>::new for any number of weights (clique_order_distribution,
// src/bin/sample.rs:87-88): cum[i] = w[0] + .. + w[i], i < n - 1
static int weighted_index_init_n(const std::vector<double>& w, double* cum, double* scale) {
  if (w.empty()) { fcm_set_error("no clique orders"); return FCM_EINVAL; }
  double total = w[0];
  for (size_t i = 1; i < w.size(); i++) { cum[i - 1] = total; total += w[i]; }
  if (!(total > 0.0)) { fcm_set_error("the graph has no maximal cliques to draw from"); return FCM_EINVAL; }
  const double max_rand = 1.0 - 2.220446049250313e-16;
  double sc = total;
  for (;;) {
    volatile double prod = sc * max_rand;
    if (prod + 0.0 < total) break;
    u64 b; memcpy(&b, &sc, 8); b -= 1; memcpy(&sc, &b, 8);
  }
  *scale = sc;
  return FCM_OK;
}

// Graph::compute_maximal_cliques (external crate, called src/lib.rs:41): maximal cliques of the
// underlying undirected graph.  Enumeration: for every vertex v the cliques whose SMALLEST vertex is
// v, by pivoted expansion over sorted candidate lists with bit-row adjacency tests (one host thread
// per stripe of v).  Order handed to the chains: each clique ascending, cliques lexicographic - the
// documented assumption of oracle/fcm_oracle.cpp (the crate is not on disk: parity unpinned).
// Maximal cliques as one flat list: clique c = verts[off[c] .. off[c + 1]), ascending; cliques in lexicographic order.
struct CliqueList {
  std::vector<u32> verts;
  std::vector<u64> off{0};
  size_t size() const { return off.size() - 1; }
  const u32* begin(size_t c) const { return verts.data() + off[c]; }
  u32 order(size_t c) const { return (u32)(off[c + 1] - off[c]); }
};
// Bron-Kerbosch with pivoting inside the neighbourhood of one vertex v, everything as bit sets over the local ids
// 0..d-1 of und[v] (LW 64-bit words): P / X / the local adjacency rows never leave the stack frame.
template <int LW>
struct LocalBK {
  u64 row[LW * 64][LW];
  u32 R[FCM_MAXD + 2];
  std::vector<u32>* out;                                // flat: order, then the local ids of the clique (without v)
  bool too_deep = false;
  static int popc(const u64* a, const u64* b) { int c = 0; for (int w = 0; w < LW; w++) c += __builtin_popcountll(a[w] & b[w]); return c; }
  void expand(int depth, u64* P, u64* X) {
    bool pe = true, xe = true;
    for (int w = 0; w < LW; w++) { pe &= P[w] == 0; xe &= X[w] == 0; }
    if (pe) {
      if (xe) { out->push_back((u32)depth); for (int t = 0; t < depth; t++) out->push_back(R[t]); }
      return;
    }
    if (depth > FCM_MAXD) { too_deep = true; return; }   // a clique beyond FCM_MAXD + 1 vertices: reported by the caller
    int pivot = -1, best = -1;
    for (int w = 0; w < LW; w++) {
      u64 m = P[w] | X[w];
      while (m) {
        const int u = w * 64 + __builtin_ctzll(m);
        m &= m - 1;
        const int c = popc(P, row[u]);
        if (c > best) { best = c; pivot = u; }
      }
    }
    u64 cand[LW];
    for (int w = 0; w < LW; w++) cand[w] = P[w] & ~row[pivot][w];
    for (int w = 0; w < LW; w++) {
      u64 m = cand[w];
      while (m) {
        const int u = w * 64 + __builtin_ctzll(m);
        m &= m - 1;
        u64 P2[LW], X2[LW];
        for (int y = 0; y < LW; y++) { P2[y] = P[y] & row[u][y]; X2[y] = X[y] & row[u][y]; }
        R[depth] = (u32)u;
        expand(depth + 1, P2, X2);
        P[u >> 6] &= ~(1ULL << (u & 63));
        X[u >> 6] |= 1ULL << (u & 63);
      }
    }
  }
};
// the general form (neighbourhoods beyond 256 vertices): sorted candidate lists, bit-row adjacency tests
struct CliqueEnum {
  const std::vector<std::vector<u32>>& und;
  const std::vector<u64>& bits;          // n x W undirected bit rows
  u32 W;
  std::vector<u32>* out;                 // flat: order, then the clique's vertices (global ids, v included, unsorted)
  bool adj(u32 a, u32 b) const { return (bits[(size_t)a * W + (b >> 6)] >> (b & 63)) & 1ULL; }
  void expand(std::vector<u32>& R, std::vector<u32>& P, std::vector<u32>& X) {
    if (P.empty()) { if (X.empty()) { out->push_back((u32)R.size()); out->insert(out->end(), R.begin(), R.end()); } return; }
    u32 pivot = P[0]; size_t best = 0; bool first = true;
    for (int which = 0; which < 2; which++)
      for (u32 u : (which ? X : P)) {
        size_t c = 0;
        for (u32 w : P) c += adj(u, w);
        if (first || c > best) { best = c; pivot = u; first = false; }
      }
    std::vector<u32> cand;
    for (u32 w : P) if (!adj(pivot, w)) cand.push_back(w);
    for (u32 v : cand) {
      std::vector<u32> P2, X2;
      for (u32 w : P) if (adj(v, w)) P2.push_back(w);
      for (u32 w : X) if (adj(v, w)) X2.push_back(w);
      R.push_back(v);
      expand(R, P2, X2);
      R.pop_back();
      P.erase(std::find(P.begin(), P.end(), v));
      X.push_back(v);
    }
  }
};
static std::vector<u64> undirected_bit_rows(u32 n, const std::vector<std::vector<u32>>& und) {
  const u32 W = (n + 63) / 64;
  std::vector<u64> bits((size_t)n * W, 0);
  for (u32 v = 0; v < n; v++) for (u32 w : und[v]) bits[(size_t)v * W + (w >> 6)] |= 1ULL << (w & 63);
  return bits;
}
// The cliques whose smallest vertex is v, appended to `raw` as (order, vertices ascending)* in lexicographic order.
// `local_of` (n entries, thread-private) holds the local id of every neighbour of v while the call runs; `scratch` and
// `keys` are reused between calls.
template <int LW>
static bool cliques_of_vertex_bits(u32 v, const std::vector<u32>& nb, const std::vector<u64>& bits, u32 W, LocalBK<LW>& bk,
                                   std::vector<u32>& local_of, std::vector<u32>& scratch, std::vector<u64>& keys, std::vector<u32>& raw) {
  const int d = (int)nb.size();
  for (int i = 0; i < d; i++) local_of[nb[i]] = (u32)i;
  // local adjacency rows: row(nb[i]) & row(v), word by word, the surviving bits renamed through local_of
  const u64* gv = &bits[(size_t)v * W];
  for (int i = 0; i < d; i++) {
    for (int w = 0; w < LW; w++) bk.row[i][w] = 0;
    const u64* gi = &bits[(size_t)nb[i] * W];
    for (u32 w = 0; w < W; w++) {
      u64 m = gi[w] & gv[w];
      while (m) {
        const u32 l = local_of[w * 64 + (u32)__builtin_ctzll(m)];
        m &= m - 1;
        bk.row[i][l >> 6] |= 1ULL << (l & 63);
      }
    }
  }
  u64 P[LW], X[LW];
  for (int w = 0; w < LW; w++) P[w] = X[w] = 0;
  for (int i = 0; i < d; i++) (nb[i] > v ? P : X)[i >> 6] |= 1ULL << (i & 63);
  scratch.clear();
  bk.out = &scratch;
  bk.too_deep = false;
  bk.expand(0, P, X);
  if (bk.too_deep) return false;
  // Lexicographic order of the group.  und[v] is ascending, so local ids order like vertices: a clique of at most 7
  // members besides v packs into one u64 (9 bits per member, local id + 1, first member in the top bits, 0 = "no more":
  // a prefix sorts before its extensions), and the group is a sort of integers.  Larger cliques: compare the lists.
  bool packable = true;
  for (size_t x = 0; x < scratch.size(); x += scratch[x] + 1) {
    const u32 o = scratch[x];
    if (o > 7) packable = false;
    std::sort(scratch.begin() + x + 1, scratch.begin() + x + 1 + o);
  }
  if (packable) {
    keys.clear();
    for (size_t x = 0; x < scratch.size(); x += scratch[x] + 1) {
      u64 k = 0;
      for (u32 t = 0; t < 7; t++) k = (k << 9) | (t < scratch[x] ? (u64)(scratch[x + 1 + t] + 1) : 0ULL);
      keys.push_back(k);
    }
    std::sort(keys.begin(), keys.end());
    for (u64 k : keys) {
      const size_t at = raw.size();
      raw.push_back(0);
      raw.push_back(v);
      for (int t = 6; t >= 0; t--) { const u32 l = (u32)((k >> (9 * t)) & 511u); if (l) raw.push_back(nb[l - 1]); }
      raw[at] = (u32)(raw.size() - at - 1);
    }
  } else {
    std::vector<u32> idx;
    for (size_t x = 0; x < scratch.size(); x += scratch[x] + 1) idx.push_back((u32)x);
    std::sort(idx.begin(), idx.end(), [&](u32 a, u32 b) {
      return std::lexicographical_compare(scratch.begin() + a + 1, scratch.begin() + a + 1 + scratch[a],
                                          scratch.begin() + b + 1, scratch.begin() + b + 1 + scratch[b]);
    });
    for (u32 a : idx) {
      raw.push_back(scratch[a] + 1);
      raw.push_back(v);
      for (u32 t = 0; t < scratch[a]; t++) raw.push_back(nb[scratch[a + 1 + t]]);
    }
  }
  return true;
}
// *too_big: a clique of more than FCM_MAXD + 1 vertices was met (the list is then incomplete)
static CliqueList maximal_cliques(u32 n, const std::vector<std::vector<u32>>& und, const std::vector<u64>& bits, int nthreads, bool* too_big) {
  const u32 W = (n + 63) / 64;
  // per vertex: where its group sits in its thread's raw buffer, how many cliques / vertices it holds
  struct Group { u64 at = 0, ncl = 0, nv = 0; };
  std::vector<Group> grp(n);
  std::vector<std::vector<u32>> raws(nthreads);
  std::vector<int> deep(nthreads, 0);
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++)
    th.emplace_back([&, t]() {
      std::vector<u32>& raw = raws[t];
      std::vector<u32> scratch, grp_raw, idx, local_of(n, 0);
      std::vector<u64> keys;
      auto bk1 = std::make_unique<LocalBK<1>>();
      auto bk2 = std::make_unique<LocalBK<2>>();
      auto bk4 = std::make_unique<LocalBK<4>>();
      for (u32 v = t; v < n; v += nthreads) {
        const std::vector<u32>& nb = und[v];
        Group& g = grp[v];
        g.at = raw.size();
        bool ok = true;
        if (nb.size() <= 64) ok = cliques_of_vertex_bits<1>(v, nb, bits, W, *bk1, local_of, scratch, keys, raw);
        else if (nb.size() <= 128) ok = cliques_of_vertex_bits<2>(v, nb, bits, W, *bk2, local_of, scratch, keys, raw);
        else if (nb.size() <= 256) ok = cliques_of_vertex_bits<4>(v, nb, bits, W, *bk4, local_of, scratch, keys, raw);
        else {
          CliqueEnum ce{und, bits, W, &scratch};
          scratch.clear();
          std::vector<u32> R{v}, P, X;
          for (u32 w : nb) (w > v ? P : X).push_back(w);
          ce.expand(R, P, X);
          grp_raw.clear();
          for (size_t x = 0; x < scratch.size();) {
            const u32 o = scratch[x++];
            grp_raw.push_back(o);
            const size_t at = grp_raw.size();
            for (u32 k = 0; k < o; k++) grp_raw.push_back(scratch[x++]);
            std::sort(grp_raw.begin() + at, grp_raw.end());
          }
          idx.clear();
          for (size_t x = 0; x < grp_raw.size(); x += grp_raw[x] + 1) idx.push_back((u32)x);
          std::sort(idx.begin(), idx.end(), [&](u32 a, u32 b) {
            return std::lexicographical_compare(grp_raw.begin() + a + 1, grp_raw.begin() + a + 1 + grp_raw[a],
                                                grp_raw.begin() + b + 1, grp_raw.begin() + b + 1 + grp_raw[b]);
          });
          for (u32 a : idx) raw.insert(raw.end(), grp_raw.begin() + a, grp_raw.begin() + a + 1 + grp_raw[a]);
        }
        if (!ok) { deep[t] = 1; continue; }
        // what the group holds: walk it once
        g.ncl = 0; g.nv = 0;
        for (size_t x = g.at; x < raw.size(); x += raw[x] + 1) { g.ncl++; g.nv += raw[x]; }
      }
    });
  for (auto& t : th) t.join();
  CliqueList cl;
  *too_big = false;
  for (int d : deep) if (d) *too_big = true;
  if (*too_big) return cl;
  // groups in ascending v = all cliques in lexicographic order; every thread copies its own groups into place
  std::vector<u64> cbase(n + 1, 0), vbase(n + 1, 0);
  for (u32 v = 0; v < n; v++) { cbase[v + 1] = cbase[v] + grp[v].ncl; vbase[v + 1] = vbase[v] + grp[v].nv; }
  cl.off.assign(cbase[n] + 1, 0);
  cl.verts.resize(vbase[n]);
  th.clear();
  for (int t = 0; t < nthreads; t++)
    th.emplace_back([&, t]() {
      const std::vector<u32>& raw = raws[t];
      for (u32 v = t; v < n; v += nthreads) {
        u64 x = grp[v].at, c = cbase[v], o = vbase[v];
        for (u64 k = 0; k < grp[v].ncl; k++) {
          const u32 ord = raw[x++];
          cl.off[c++] = o;
          for (u32 y = 0; y < ord; y++) cl.verts[o++] = raw[x++];
        }
      }
    });
  for (auto& t : th) t.join();
  cl.off[cbase[n]] = vbase[n];
  return cl;
}

// ---- compute_edge_neighborhoods (src/lib.rs:331-356) on the device: for every adjacent pair (big, small)
//      the ascending list of common undirected neighbours = the set bits of und[big] & und[small].
//      One warp per pair: a counting pass, offsets by prefix sum, a compaction pass (lane l owns word
//      l of every 32-word chunk; warp scan of the popcounts gives each lane its place in the list). ----
__global__ void und_bits_kernel(const uint2* __restrict__ pairs, u64 np, u32 W, u64* __restrict__ und) {
  for (u64 q = (u64)blockIdx.x * blockDim.x + threadIdx.x; q < np; q += (u64)gridDim.x * blockDim.x) {
    const uint2 ab = pairs[q];
    atomicOr((unsigned long long*)&und[(u64)ab.x * W + (ab.y >> 6)], 1ULL << (ab.y & 63));
    atomicOr((unsigned long long*)&und[(u64)ab.y * W + (ab.x >> 6)], 1ULL << (ab.x & 63));
  }
}
// dbl (may be null): bit rows of the DOUBLE edges.  The reference builds its adjacency lists from the
// directed edges without deduplication (src/lib.rs:333-336), so a common neighbour w joined to both a
// and b by double edges appears TWICE in the list of (a, b); with dbl given the lists carry that
// multiplicity (the API-visible map), without it they are sets (what the chains use).
__global__ void nbhd_count_kernel(const uint2* __restrict__ pairs, u64 np, u32 W, const u64* __restrict__ und,
                                  const u64* __restrict__ dbl, u32* __restrict__ cnt) {
  const int lane = threadIdx.x & 31;
  const u64 nwarps = ((u64)gridDim.x * blockDim.x) >> 5;
  for (u64 q = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < np; q += nwarps) {
    const uint2 ab = pairs[q];
    const u64 *ra = und + (u64)ab.x * W, *rb = und + (u64)ab.y * W;
    u32 c = 0;
    for (u32 w = lane; w < W; w += 32) {
      c += __popcll(__ldg(&ra[w]) & __ldg(&rb[w]));
      if (dbl) c += __popcll(__ldg(&dbl[(u64)ab.x * W + w]) & __ldg(&dbl[(u64)ab.y * W + w]));
    }
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) cnt[q] = c;
  }
}
__global__ void nbhd_fill_kernel(const uint2* __restrict__ pairs, u64 np, u32 W, const u64* __restrict__ und,
                                 const u64* __restrict__ dbl, const u64* __restrict__ off, u32* __restrict__ list) {
  const int lane = threadIdx.x & 31;
  const u64 nwarps = ((u64)gridDim.x * blockDim.x) >> 5;
  for (u64 q = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < np; q += nwarps) {
    const uint2 ab = pairs[q];
    const u64 *ra = und + (u64)ab.x * W, *rb = und + (u64)ab.y * W;
    u64 base = off[q];
    for (u32 w0 = 0; w0 < W; w0 += 32) {
      const u32 w = w0 + lane;
      u64 bits = w < W ? (__ldg(&ra[w]) & __ldg(&rb[w])) : 0ULL;
      const u64 twice = (dbl && w < W) ? (__ldg(&dbl[(u64)ab.x * W + w]) & __ldg(&dbl[(u64)ab.y * W + w])) : 0ULL;
      const u32 c = __popcll(bits) + __popcll(twice);
      u32 incl = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const u32 t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
      u64 o = base + incl - c;
      while (bits) {
        const int b = __ffsll((long long)bits) - 1;
        bits &= bits - 1;
        list[o++] = w * 64 + b;
        if ((twice >> b) & 1ULL) list[o++] = w * 64 + b;
      }
      base += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
}

// pairs (big, small) -> nb_off[npairs + 1] (exclusive prefix), nb_list, maxc.  Temporary device memory only.
static int edge_neighborhoods_device(fcm_ctx* ctx, u32 n, const std::vector<uint2>& pair_ab,
                                     std::vector<u64>& nb_off, std::vector<u32>& nb_list, u32* maxc_out,
                                     const std::vector<uint2>* double_pairs = nullptr) {
  const u64 np = pair_ab.size();
  nb_off.assign(np + 1, 0);
  nb_list.assign(1, 0);
  *maxc_out = 0;
  if (np == 0) return FCM_OK;
  const u32 W = ((n + 63) / 64 + 1) & ~1u;
  cudaStream_t s = ctx->stream;
  uint2* d_pairs = nullptr; u64* d_und = nullptr; u32* d_cnt = nullptr; u64* d_off = nullptr; u32* d_list = nullptr;
  uint2* d_dpairs = nullptr; u64* d_dbl = nullptr;
  auto cleanup = [&]() { cudaFree(d_pairs); cudaFree(d_und); cudaFree(d_cnt); cudaFree(d_off); cudaFree(d_list); cudaFree(d_dpairs); cudaFree(d_dbl); };
#define N_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { fcm_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); cleanup(); return e__ == cudaErrorMemoryAllocation ? FCM_ENOMEM : FCM_ECUDA; } } while (0)
  N_CUDA(cudaMalloc(&d_pairs, np * 8));
  N_CUDA(cudaMalloc(&d_und, (size_t)n * W * 8));
  N_CUDA(cudaMalloc(&d_cnt, np * 4));
  N_CUDA(cudaMemcpyAsync(d_pairs, pair_ab.data(), np * 8, cudaMemcpyHostToDevice, s));
  N_CUDA(cudaMemsetAsync(d_und, 0, (size_t)n * W * 8, s));
  const int tb = 256;
  const int gb = (int)std::min<u64>((np + tb - 1) / tb, (u64)ctx->sm_count * 16);
  und_bits_kernel<<<gb, tb, 0, s>>>(d_pairs, np, W, d_und);
  if (double_pairs) {
    N_CUDA(cudaMalloc(&d_dbl, (size_t)n * W * 8));
    N_CUDA(cudaMemsetAsync(d_dbl, 0, (size_t)n * W * 8, s));
    if (!double_pairs->empty()) {
      N_CUDA(cudaMalloc(&d_dpairs, double_pairs->size() * 8));
      N_CUDA(cudaMemcpyAsync(d_dpairs, double_pairs->data(), double_pairs->size() * 8, cudaMemcpyHostToDevice, s));
      und_bits_kernel<<<gb, tb, 0, s>>>(d_dpairs, double_pairs->size(), W, d_dbl);
      ctx->launches++;
    }
  }
  const int gw = (int)std::min<u64>((np * 32 + tb - 1) / tb, (u64)ctx->sm_count * 32);
  nbhd_count_kernel<<<gw, tb, 0, s>>>(d_pairs, np, W, d_und, d_dbl, d_cnt);
  ctx->launches += 2;
  N_CUDA(cudaGetLastError());
  std::vector<u32> cnt(np);
  N_CUDA(cudaMemcpyAsync(cnt.data(), d_cnt, np * 4, cudaMemcpyDeviceToHost, s));
  N_CUDA(cudaStreamSynchronize(s));
  u32 maxc = 0;
  for (u64 q = 0; q < np; q++) { maxc = std::max(maxc, cnt[q]); nb_off[q + 1] = nb_off[q] + cnt[q]; }
  *maxc_out = maxc;
  const u64 total = nb_off[np];
  nb_list.assign(std::max<u64>(total, 1), 0);
  if (total) {
    N_CUDA(cudaMalloc(&d_off, (np + 1) * 8));
    N_CUDA(cudaMalloc(&d_list, total * 4));
    N_CUDA(cudaMemcpyAsync(d_off, nb_off.data(), (np + 1) * 8, cudaMemcpyHostToDevice, s));
    nbhd_fill_kernel<<<gw, tb, 0, s>>>(d_pairs, np, W, d_und, d_dbl, d_off, d_list);
    ctx->launches++;
    N_CUDA(cudaGetLastError());
    N_CUDA(cudaMemcpyAsync(nb_list.data(), d_list, total * 4, cudaMemcpyDeviceToHost, s));
    N_CUDA(cudaStreamSynchronize(s));
  }
#undef N_CUDA
  cleanup();
  return FCM_OK;
}

// State::compute_edge_neighborhoods (src/lib.rs:331-356) as a call of its own (src/bin/edgeset_nbhd.rs
// prints these lists): pairs in order of first appearance in the edge list, key [big, small].
extern "C" int fcm_edge_neighborhoods(fcm_ctx* ctx, uint32_t n, uint64_t nedges, const uint32_t (*edges)[2],
                                      uint64_t* npairs_out, uint64_t* list_len_out, uint32_t* pairs_out,
                                      uint64_t* off_out, uint32_t* list_out) {
  REQUIRE(ctx && npairs_out && list_len_out, "ctx/npairs_out/list_len_out is null");
  REQUIRE(nedges == 0 || edges, "edges is null");
  FCM_CUDA(cudaSetDevice(ctx->device));
  std::vector<uint2> pair_ab, double_pairs;
  std::unordered_map<u64, u32> seen;                 // pair -> directions met (bit 0: big -> small, bit 1: small -> big)
  seen.reserve(nedges * 2);
  for (u64 i = 0; i < nedges; i++) {
    const u32 a = edges[i][0], b = edges[i][1];
    if (a >= n || b >= n || a == b) { fcm_set_error("edge %llu = (%u,%u): vertex id >= n or self loop", (unsigned long long)i, a, b); return FCM_EINVAL; }
    const u32 big = std::max(a, b), small = std::min(a, b), dir = a > b ? 1u : 2u;
    auto ins = seen.emplace(pkey(big, small), dir);
    if (ins.second) pair_ab.push_back(make_uint2(big, small));
    else if ((ins.first->second | dir) == 3u && ins.first->second != 3u) { ins.first->second = 3u; double_pairs.push_back(make_uint2(big, small)); }
  }
  std::vector<u64> off; std::vector<u32> list; u32 maxc = 0;
  int rc = edge_neighborhoods_device(ctx, n, pair_ab, off, list, &maxc, &double_pairs);
  if (rc) return rc;
  *npairs_out = pair_ab.size();
  *list_len_out = off[pair_ab.size()];
  if (pairs_out) memcpy(pairs_out, pair_ab.data(), pair_ab.size() * 8);
  if (off_out) memcpy(off_out, off.data(), off.size() * 8);
  if (list_out && off[pair_ab.size()]) memcpy(list_out, list.data(), off[pair_ab.size()] * 4);
  return FCM_OK;
}

// Graph::compute_maximal_cliques (src/lib.rs:41) as a call of its own: host threads, no device (include/fcm.h)
extern "C" int fcm_maximal_cliques(uint32_t n, uint64_t nedges, const uint32_t (*edges)[2], uint64_t* ncliques_out,
                                   uint64_t* nverts_out, uint32_t* sizes_out, uint32_t* verts_out) {
  REQUIRE(ncliques_out && nverts_out, "ncliques/nverts is null");
  REQUIRE(nedges == 0 || edges, "edges is null");
  REQUIRE((sizes_out == nullptr) == (verts_out == nullptr), "sizes_out and verts_out go together");
  std::vector<std::vector<u32>> und(n);
  for (u64 i = 0; i < nedges; i++) {
    const u32 a = edges[i][0], b = edges[i][1];
    if (a >= n || b >= n || a == b) { fcm_set_error("edge %llu = (%u,%u): vertex id >= n or self loop", (unsigned long long)i, a, b); return FCM_EINVAL; }
    und[a].push_back(b); und[b].push_back(a);
  }
  for (auto& l : und) { std::sort(l.begin(), l.end()); l.erase(std::unique(l.begin(), l.end()), l.end()); }
  const int nthreads = (int)std::max(1u, std::min(std::thread::hardware_concurrency(), 32u));
  bool too_big = false;
  const CliqueList cl = maximal_cliques(n, und, undirected_bit_rows(n, und), nthreads, &too_big);
  if (too_big) { fcm_set_error("a maximal clique of order > %d", FCM_MAXD); return FCM_EDIM; }
  for (size_t c = 0; c < cl.size(); c++)
    if (cl.order(c) > (u32)FCM_MAXD) { fcm_set_error("a maximal clique of order %u (> %d)", cl.order(c), FCM_MAXD); return FCM_EDIM; }
  if (sizes_out) {
    REQUIRE(*ncliques_out >= cl.size() && *nverts_out >= cl.verts.size(), "buffers are smaller than the first call reported");
    for (size_t c = 0; c < cl.size(); c++) sizes_out[c] = cl.order(c);
    if (!cl.verts.empty()) memcpy(verts_out, cl.verts.data(), cl.verts.size() * 4);
  }
  *ncliques_out = cl.size();
  *nverts_out = cl.verts.size();
  return FCM_OK;
}

extern "C" int fcm_ensemble_destroy(fcm_ensemble* e) {
  if (!e) return FCM_OK;
  if (e->ctx) { cudaSetDevice(e->ctx->device); cudaStreamSynchronize(e->ctx->stream); }
  for (void* q : e->dev_allocs) cudaFree(q);
  if (e->l2_window.num_bytes) cudaCtxResetPersistingL2Cache();   // the index's persisting lines go back to normal
  if (e->trace_dev) cudaFree(e->trace_dev);
  if (e->eb_out_dev) cudaFree(e->eb_out_dev);
  delete e;
  return FCM_OK;
}

extern "C" int fcm_ensemble_create(fcm_ctx* ctx, uint32_t n, uint64_t nedges, const uint32_t (*edges)[2],
                                   uint32_t nchains, const uint64_t* seeds,
                                   const uint64_t* flag_count0, int len0,
                                   const uint64_t* bounds_min, int nmin, const uint64_t* bounds_max, int nmax,
                                   const double move_weights[4], fcm_ensemble** out) {
  REQUIRE(ctx && out, "ctx/out is null");
  *out = nullptr;
  REQUIRE(nedges == 0 || edges, "edges is null");
  REQUIRE(nchains >= 1 && seeds, "need at least one chain and its seed");
  REQUIRE(flag_count0 && len0 >= 0 && len0 <= FCM_MAXD, "flag_count0 missing or longer than FCM_MAXD");
  REQUIRE(move_weights, "move_weights is null");
  const bool want_cliques = move_weights[2] != 0.0 || move_weights[3] != 0.0;
  FCM_CUDA(cudaSetDevice(ctx->device));
  const bool timing = getenv("FCM_TIMING") != nullptr;   // FCM_TIMING=1: the phases of this call on stderr
  auto t_last = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!timing) return;
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[fcm] ensemble_create: %-44s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
    t_last = now;
  };
  fcm_ensemble* e = new fcm_ensemble();
  e->ctx = ctx;
  EnsParams& p = e->p;
  memset(&p, 0, sizeof(p));
  int rc;
  auto fail = [&](int code) { fcm_ensemble_destroy(e); return code; };
  if ((rc = copy_bounds(p, bounds_min, nmin, bounds_max, nmax))) return fail(rc);
  if ((rc = weighted_index_init(move_weights, p.cum, &p.scale))) return fail(rc);
  p.n = n; p.nchains = nchains;
  e->deep_hint = len0 > 6;                           // flag_count0 reaches dimension 6: see fcm_chain_launch_step
  p.W = ((n + 63) / 64 + 1) & ~1u;
  if (p.W == 0) p.W = 2;

  // ---- pairs, slots, containers: replay Graph::add_edge in edge-list order (src/io.rs:26-33) ----
  // Pair ids are an internal numbering (the API speaks vertices).  They are dealt in (big, small) order: the
  // pairs {x, u} with x < u then sit side by side, 2 orientation bits each, so the orientation words a
  // proposal looks up for the triangles and 4-cliques of (u, v) fall into few 32-byte sectors - range(u),
  // range(v) and range(w) of the common neighbours w > v - instead of one random sector per entry
  // (measured on the C5 shard, profiles/README.md r02_b).  FCM_PAIR_ORDER=0 numbers them in order of first
  // appearance in the edge list instead (the round-1 layout, kept for the A/B measurement).
  std::vector<uint2> pair_ab;
  std::vector<u32> items, pos, dbl, dblpos;
  std::vector<u64> adj((size_t)n * p.W, 0);
  e->pairmap.reserve(nedges * 2);
  {
    const char* po = getenv("FCM_PAIR_ORDER");
    const bool sorted_ids = !(po && po[0] == '0');
    for (u64 i = 0; i < nedges; i++) {
      const u32 a = edges[i][0], b = edges[i][1];
      if (a >= n || b >= n || a == b) { fcm_set_error("edge %llu = (%u,%u): vertex id >= n or self loop", (unsigned long long)i, a, b); return fail(FCM_EINVAL); }
      const u32 big = std::max(a, b), small = std::min(a, b);
      if (e->pairmap.emplace(pkey(big, small), (u32)pair_ab.size()).second) pair_ab.push_back(make_uint2(big, small));
    }
    if (sorted_ids) {
      std::sort(pair_ab.begin(), pair_ab.end(), [](const uint2& x, const uint2& y) { return pkey(x.x, x.y) < pkey(y.x, y.y); });
      for (size_t q = 0; q < pair_ab.size(); q++) e->pairmap[pkey(pair_ab[q].x, pair_ab[q].y)] = (u32)q;
    }
    pos.assign(2 * pair_ab.size(), FCM_NONE);
    dblpos.assign(pair_ab.size(), FCM_NONE);
  }
  for (u64 i = 0; i < nedges; i++) {
    const u32 a = edges[i][0], b = edges[i][1];
    const u32 pair = e->pairmap.find(pkey(std::max(a, b), std::min(a, b)))->second;
    const u32 slot = 2 * pair + (a > b ? 0u : 1u);
    if (pos[slot] != FCM_NONE) continue;               // duplicate line: IndexSet insert is a no-op
    pos[slot] = (u32)items.size();
    items.push_back(slot);
    adj[(size_t)a * p.W + (b >> 6)] |= 1ULL << (b & 63);
    if (pos[slot ^ 1] != FCM_NONE) { dblpos[pair] = (u32)dbl.size(); dbl.push_back(pair); }
  }
  p.npairs = pair_ab.size();
  e->ne0 = items.size(); e->nd0 = dbl.size();
  if (p.npairs > 0x7fffffffULL) { fcm_set_error("too many adjacent pairs"); return fail(FCM_EINVAL); }
  if (len0 >= 2 && flag_count0[1] != e->ne0) { fcm_set_error("flag_count0[1] = %llu but the graph has %llu edges", (unsigned long long)flag_count0[1], (unsigned long long)e->ne0); return fail(FCM_EINVAL); }

  // ---- compute_edge_neighborhoods (src/lib.rs:331-356): sorted common undirected neighbours ----
  std::vector<std::vector<u32>> und(n);
  for (const uint2& ab : pair_ab) { und[ab.x].push_back(ab.y); und[ab.y].push_back(ab.x); }
  int nthreads = (int)std::max(1u, std::min(std::thread::hardware_concurrency(), 32u));
  {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back([&, t]() { for (u32 v = t; v < n; v += nthreads) std::sort(und[v].begin(), und[v].end()); });
    for (auto& t : th) t.join();
  }
  std::vector<u64> nb_off;                    // 64-bit while summing; uploaded as u32
  std::vector<u32> nb_list;
  u32 maxc = 0;
  lap("pairs, containers, undirected lists (host)");
  if ((rc = edge_neighborhoods_device(ctx, n, pair_ab, nb_off, nb_list, &maxc))) return fail(rc);
  p.maxc = maxc;
  lap("edge neighbourhoods (device)");

  // ---- static clique index (hot path of chain_step_kernel): for every pair the slots of the edges
  //      between its common neighbours and its endpoints (triangles) and of the edges among its
  //      common neighbours (4-cliques).  The undirected graph never changes under the moves
  //      (src/lib.rs:292-325), so a proposal only has to look up the ORIENTATION of these slots. ----
  std::vector<uint4> pinfo;
  std::vector<uint2> tri;
  std::vector<u32> k4;
  std::vector<u64> und_bits;                  // n x ceil(n/64) undirected bit rows (host), built when the index or the cliques need them
  {
    const u64 nslots_ = 2 * p.npairs;
    u64 work = 0;
    for (u64 q = 0; q < p.npairs; q++) { u64 c = nb_off[q + 1] - nb_off[q]; if (c <= 32) work += c * c; }
    const char* off_env = getenv("FCM_NO_CLIQUE_INDEX");
    bool build = p.npairs > 0 && nslots_ < (1ULL << 22) && work < (1ULL << 32) && !(off_env && off_env[0] == '1');
    if (build) {
      // pair id of (v, und[v][k]), parallel to the sorted neighbour lists
      std::vector<std::vector<u32>> undpid(n);
      for (u32 v = 0; v < n; v++) undpid[v].resize(und[v].size());
      for (u64 q = 0; q < p.npairs; q++) {
        const uint2 ab = pair_ab[q];
        undpid[ab.x][std::lower_bound(und[ab.x].begin(), und[ab.x].end(), ab.y) - und[ab.x].begin()] = (u32)q;
        undpid[ab.y][std::lower_bound(und[ab.y].begin(), und[ab.y].end(), ab.x) - und[ab.y].begin()] = (u32)q;
      }
      auto slot_of = [&](u32 from, u32 to) -> u32 {      // FCM_NONE when not adjacent
        const std::vector<u32>& L = und[from];
        auto it = std::lower_bound(L.begin(), L.end(), to);
        if (it == L.end() || *it != to) return FCM_NONE;
        return 2 * undpid[from][it - L.begin()] + (from > to ? 0u : 1u);
      };
      und_bits = undirected_bit_rows(n, und);
      const u32 Wb = (n + 63) / 64;
      tri.resize(nb_list.size());
      std::vector<u64> k4_off(p.npairs + 1, 0);
      std::vector<std::vector<u32>> k4_part(nthreads);
      std::vector<std::vector<u32>> k4_cnt(nthreads);
      // contiguous ranges of pairs per thread so that the per-thread entry lists concatenate in pair order
      auto range = [&](int t) { return std::make_pair(p.npairs * t / nthreads, p.npairs * (t + 1) / nthreads); };
      std::vector<std::thread> th;
      for (int t = 0; t < nthreads; t++) th.emplace_back([&, t]() {
        auto r = range(t);
        for (u64 q = r.first; q < r.second; q++) {
          const uint2 ab = pair_ab[q];
          const u64 o = nb_off[q];
          const u32 c = (u32)(nb_off[q + 1] - o);
          for (u32 i = 0; i < c; i++) tri[o + i] = make_uint2(slot_of(nb_list[o + i], ab.x), slot_of(nb_list[o + i], ab.y));
          u32 m = 0;
          if (c <= 32)
            for (u32 i = 0; i < c; i++)
              for (u32 j = i + 1; j < c; j++) {
                const u32 wi = nb_list[o + i], wj = nb_list[o + j];
                if (!((und_bits[(size_t)wi * Wb + (wj >> 6)] >> (wj & 63)) & 1ULL)) continue;   // nine in ten are not adjacent: no search
                const u32 sl = slot_of(wi, wj);
                if (sl != FCM_NONE) { k4_part[t].push_back((sl << 10) | (j << 5) | i); m++; }
              }
          k4_cnt[t].push_back(m);
        }
      });
      for (auto& t : th) t.join();
      u64 tot = 0;
      for (int t = 0; t < nthreads; t++) {
        auto r = range(t);
        for (u64 q = r.first; q < r.second; q++) { k4_off[q] = tot; tot += k4_cnt[t][q - r.first]; }
      }
      k4_off[p.npairs] = tot;
      if (tot < 0xffffffffULL) {
        k4.reserve(tot + 1);
        for (int t = 0; t < nthreads; t++) k4.insert(k4.end(), k4_part[t].begin(), k4_part[t].end());
        pinfo.resize(2 * p.npairs);
        for (u64 q = 0; q < p.npairs; q++) {
          pinfo[2 * q] = make_uint4(pair_ab[q].x, pair_ab[q].y, (u32)nb_off[q], (u32)k4_off[q]);
          pinfo[2 * q + 1] = make_uint4((u32)(nb_off[q + 1] - nb_off[q]), (u32)(k4_off[q + 1] - k4_off[q]), 0, 0);
        }
        e->have_index = 1;
      } else tri.clear();
    }
  }
  lap("clique index (host threads)");
  u32 kneed = 2 * maxc + 4;                            // double-edge move: two pairs + four endpoints
  // ---- clique moves: maximal cliques by order, the order distribution, the sorted pair keys ----
  std::vector<u32> cq_verts;
  std::vector<u64> pkeys(p.npairs);
  std::vector<u32> pkey_pair(p.npairs);
  std::vector<u32> pkey_begin((size_t)n + 1, 0);
  {
    std::vector<u32> rk(p.npairs);
    for (u64 q = 0; q < p.npairs; q++) rk[q] = (u32)q;
    auto key_less = [&](u32 x, u32 y) { return pkey(pair_ab[x].x, pair_ab[x].y) < pkey(pair_ab[y].x, pair_ab[y].y); };
    if (!std::is_sorted(rk.begin(), rk.end(), key_less)) std::sort(rk.begin(), rk.end(), key_less);   // (pair ids are dealt in key order by default)
    for (u64 r = 0; r < p.npairs; r++) { pkeys[r] = pkey(pair_ab[rk[r]].x, pair_ab[rk[r]].y); pkey_pair[r] = rk[r]; }
    for (u64 r = 0; r < p.npairs; r++) pkey_begin[(size_t)(pkeys[r] >> 32) + 1]++;
    for (u32 v = 0; v < n; v++) pkey_begin[v + 1] += pkey_begin[v];
  }
  if (want_cliques) {
    if (und_bits.empty()) und_bits = undirected_bit_rows(n, und);
    bool too_big = false;
    const CliqueList cl = maximal_cliques(n, und, und_bits, nthreads, &too_big);
    lap("maximal cliques (host threads)");
    size_t maxo = 0;
    for (size_t ci = 0; ci < cl.size(); ci++) maxo = std::max<size_t>(maxo, cl.order(ci));
    if (too_big || maxo > FCM_MAXD) { fcm_set_error("a maximal clique of order %s%d", too_big ? "> " : "", too_big ? FCM_MAXD : (int)maxo); return fail(FCM_EDIM); }
    std::vector<u64> cnt(maxo, 0);
    for (size_t ci = 0; ci < cl.size(); ci++) cnt[cl.order(ci) - 1]++;
    u64 base = 0;
    for (size_t o = 0; o < maxo; o++) {
      if (cnt[o] > 0xffffffffULL || base > 0xffffffffULL) { fcm_set_error("too many maximal cliques"); return fail(FCM_ELIMIT); }
      p.cq_count[o] = (u32)cnt[o]; p.cq_base[o] = (u32)base; base += cnt[o] * (o + 1);
    }
    cq_verts.resize(std::max<u64>(base, 1));
    std::vector<u64> fill(maxo, 0);
    for (size_t ci = 0; ci < cl.size(); ci++) {        // lexicographic order is kept inside every order
      const size_t o = cl.order(ci) - 1;
      std::copy(cl.begin(ci), cl.begin(ci) + o + 1, cq_verts.begin() + p.cq_base[o] + fill[o] * (o + 1));
      fill[o]++;
    }
    // |edgeset_neighborhood| of a permutation of a clique: its vertices + the common neighbours of its pairs; the largest
    // one sizes the full path's local matrix (host threads, one mark array each; the pair of two vertices by binary
    // search in the key run of the bigger one)
    u32 kunion = 0;
    {
      std::vector<u32> best(nthreads, 0);
      std::vector<std::thread> th;
      for (int t = 0; t < nthreads; t++) th.emplace_back([&, t]() {
        std::vector<u32> mark(n, 0);
        u32 stamp = 0, mx = 0;
        const size_t c0 = cl.size() * t / nthreads, c1 = cl.size() * (t + 1) / nthreads;
        for (size_t ci = c0; ci < c1; ci++) {
          const u32* c = cl.begin(ci);
          const u32 ord = cl.order(ci);
          stamp++;
          u32 usz = 0;
          for (u32 i = 0; i < ord; i++) if (mark[c[i]] != stamp) { mark[c[i]] = stamp; usz++; }
          for (u32 i = 0; i < ord; i++)
            for (u32 j = i + 1; j < ord; j++) {
              const u64 key = pkey(c[j], c[i]);          // ascending clique: c[j] is the bigger vertex
              const u64* lo = std::lower_bound(pkeys.data() + pkey_begin[c[j]], pkeys.data() + pkey_begin[c[j] + 1], key);
              const u32 q = pkey_pair[lo - pkeys.data()];
              for (u64 x = nb_off[q]; x < nb_off[q + 1]; x++) if (mark[nb_list[x]] != stamp) { mark[nb_list[x]] = stamp; usz++; }
            }
          mx = std::max(mx, usz);
        }
        best[t] = mx;
      });
      for (auto& t : th) t.join();
      for (u32 b : best) kunion = std::max(kunion, b);
    }
    std::vector<double> w(maxo);
    for (size_t o = 0; o < maxo; o++) w[o] = std::pow((double)cnt[o], 0.2);   // src/bin/sample.rs:87
    if ((rc = weighted_index_init_n(w, p.cq_cum, &p.cq_scale))) return fail(rc);
    p.cq_norders = (int)maxo;
    e->cliques = 1;
    kneed = std::max<u32>(kneed, std::min<u32>(n, 2 * kunion));   // clique_swap: two cliques
  }
  lap("pair keys + clique tables");
  p.Kcap = std::min<u32>(std::max<u32>(kneed, 8), 512);
  u32 kw = (p.Kcap + 63) / 64;
  p.KWcap = kw <= 1 ? 1 : kw <= 2 ? 2 : kw <= 4 ? 4 : 8;
  e->smem_per_warp = fcm_chain_smem_per_warp(p.Kcap, p.KWcap, e->cliques != 0);
  e->warps_per_block = 4;
  while (e->warps_per_block > 1 && e->smem_per_warp * e->warps_per_block > ctx->smem_optin) e->warps_per_block >>= 1;
  if (e->smem_per_warp > ctx->smem_optin) { fcm_set_error("neighbourhood tile does not fit shared memory"); return fail(FCM_ELIMIT); }

  // ---- device memory ----
  const u64 nslots = 2 * p.npairs;
  const u64 adjw = (u64)n * p.W;
  e->bytes_per_chain = adjw * 8 + nslots * 8 + p.npairs * 8 + (nslots + 7) / 8 + sizeof(ChainState);   // before padding
  uint2* d_pair; u32* d_off; u32* d_list;
  if (nb_off[p.npairs] > 0xffffffffULL) { fcm_set_error("common-neighbour lists hold %llu entries (> 2^32)", (unsigned long long)nb_off[p.npairs]); return fail(FCM_ELIMIT); }
  std::vector<u32> nb_off32(nb_off.begin(), nb_off.end());
  if ((rc = dev_alloc(e, &d_pair, p.npairs))) return fail(rc);
  if ((rc = dev_alloc(e, &d_off, p.npairs + 1))) return fail(rc);
  if ((rc = dev_alloc(e, &d_list, nb_list.size()))) return fail(rc);
  {
    // one block per chain: [adj | items | pos | dbl | dblpos], block stride padded to 128 bytes
    const u64 w_ori = (nslots + 31) / 32 + 1;
    const u64 w_adj = adjw * 2, w_total = w_adj + 2 * nslots + 2 * p.npairs + w_ori;
    if (w_total > 0xfffffff0ULL) { fcm_set_error("graph too large: a chain's state exceeds 2^32 words"); return fail(FCM_ELIMIT); }
    p.stride32 = (w_total + 31) & ~31ULL;
    p.off_items = (u32)w_adj; p.off_pos = (u32)(w_adj + nslots);
    p.off_dbl = (u32)(w_adj + 2 * nslots); p.off_dblpos = (u32)(w_adj + 2 * nslots + p.npairs);
    u32* blk = nullptr;
    if ((rc = dev_alloc(e, &blk, p.stride32 * nchains))) return fail(rc);
    p.adj = (u64*)blk; p.items = blk + p.off_items; p.pos = blk + p.off_pos;
    p.off_ori = (u32)(w_adj + 2 * nslots + 2 * p.npairs);
    p.dbl = blk + p.off_dbl; p.dblpos = blk + p.off_dblpos; p.ori = blk + p.off_ori;
    std::vector<u32> ori(w_ori, 0);
    for (u64 sl = 0; sl < nslots; sl++) if (pos[sl] != FCM_NONE) ori[sl >> 5] |= 1u << (sl & 31);
    cudaError_t ce = cudaMemcpyAsync(p.ori, ori.data(), w_ori * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);   // `ori` is a local
    if (ce != cudaSuccess) { fcm_set_error("upload of the orientation bits failed: %s", cudaGetErrorString(ce)); return fail(FCM_ECUDA); }
  }
  if (e->have_index) {
    // pinfo | tri | k4 share one slab (256-byte aligned parts) so that a single L2 access-policy window covers the index
    const size_t b_pinfo = pinfo.size() * sizeof(uint4), b_tri = tri.size() * sizeof(uint2), b_k4 = k4.size() * 4;
    const size_t o_tri = (b_pinfo + 255) & ~(size_t)255, o_k4 = (o_tri + b_tri + 255) & ~(size_t)255;
    unsigned char* slab = nullptr;
    if ((rc = dev_alloc(e, &slab, o_k4 + b_k4))) return fail(rc);
    uint4* d_pinfo = (uint4*)slab; uint2* d_tri = (uint2*)(slab + o_tri); u32* d_k4 = (u32*)(slab + o_k4);
    index_l2_window(e, slab, o_k4 + b_k4);
    cudaError_t ce = cudaMemcpyAsync(d_pinfo, pinfo.data(), pinfo.size() * sizeof(uint4), cudaMemcpyHostToDevice, ctx->stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(d_tri, tri.data(), tri.size() * sizeof(uint2), cudaMemcpyHostToDevice, ctx->stream);
    if (ce == cudaSuccess && !k4.empty()) ce = cudaMemcpyAsync(d_k4, k4.data(), k4.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
    if (ce != cudaSuccess) { fcm_set_error("upload of the clique index failed: %s", cudaGetErrorString(ce)); return fail(FCM_ECUDA); }
    p.pinfo = d_pinfo; p.tri = d_tri; p.k4 = d_k4;
    e->index_bytes = pinfo.size() * sizeof(uint4) + tri.size() * sizeof(uint2) + k4.size() * 4;
  }
  {
    u64* d_keys; u32* d_kp; u32* d_cq; u32* d_kb;
    if ((rc = dev_alloc(e, &d_keys, pkeys.size()))) return fail(rc);
    if ((rc = dev_alloc(e, &d_kp, pkey_pair.size()))) return fail(rc);
    if ((rc = dev_alloc(e, &d_cq, cq_verts.size()))) return fail(rc);
    if ((rc = dev_alloc(e, &d_kb, pkey_begin.size()))) return fail(rc);
    cudaError_t ce = cudaSuccess;
    if (!pkeys.empty()) ce = cudaMemcpyAsync(d_keys, pkeys.data(), pkeys.size() * 8, cudaMemcpyHostToDevice, ctx->stream);
    if (ce == cudaSuccess && !pkey_pair.empty()) ce = cudaMemcpyAsync(d_kp, pkey_pair.data(), pkey_pair.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (ce == cudaSuccess && !cq_verts.empty()) ce = cudaMemcpyAsync(d_cq, cq_verts.data(), cq_verts.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(d_kb, pkey_begin.data(), pkey_begin.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
    if (ce != cudaSuccess) { fcm_set_error("upload of the clique tables failed: %s", cudaGetErrorString(ce)); return fail(FCM_ECUDA); }
    p.pkeys = d_keys; p.pkey_pair = d_kp; p.cq_verts = d_cq; p.pkey_begin = d_kb;
  }
  if ((rc = dev_alloc(e, &p.st, (size_t)nchains))) return fail(rc);
  if ((rc = dev_alloc(e, &e->edit_dev, (size_t)FCM_MAXEDIT * 2))) return fail(rc);
  if ((rc = dev_alloc(e, &e->io_dev, (size_t)2 * FCM_MAXD + 8))) return fail(rc);
  p.pair_ab = d_pair; p.nb_off = d_off; p.nb_list = d_list;
  cudaStream_t s = ctx->stream;
#define E_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { fcm_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); return fail(FCM_ECUDA); } } while (0)
  items.resize(nslots, 0); dbl.resize(p.npairs, 0);
  if (p.npairs) {
    E_CUDA(cudaMemcpyAsync(d_pair, pair_ab.data(), p.npairs * 8, cudaMemcpyHostToDevice, s));
    E_CUDA(cudaMemcpyAsync(p.items, items.data(), nslots * 4, cudaMemcpyHostToDevice, s));
    E_CUDA(cudaMemcpyAsync(p.pos, pos.data(), nslots * 4, cudaMemcpyHostToDevice, s));
    E_CUDA(cudaMemcpyAsync(p.dbl, dbl.data(), p.npairs * 4, cudaMemcpyHostToDevice, s));
    E_CUDA(cudaMemcpyAsync(p.dblpos, dblpos.data(), p.npairs * 4, cudaMemcpyHostToDevice, s));
  }
  E_CUDA(cudaMemcpyAsync(d_off, nb_off32.data(), (p.npairs + 1) * 4, cudaMemcpyHostToDevice, s));
  E_CUDA(cudaMemcpyAsync(d_list, nb_list.data(), nb_list.size() * 4, cudaMemcpyHostToDevice, s));
  if (adjw) E_CUDA(cudaMemcpyAsync(p.adj, adj.data(), adjw * 8, cudaMemcpyHostToDevice, s));
  std::vector<ChainState> st(nchains);
  for (u32 k = 0; k < nchains; k++) {
    memset(&st[k], 0, sizeof(ChainState));
    seed_from_u64(seeds[k], st[k].rng);
    for (int d = 0; d < len0; d++) st[k].flag_count[d] = flag_count0[d];
    st[k].fc_len = len0;
    st[k].E = (u32)e->ne0; st[k].ND = (u32)e->nd0;
  }
  E_CUDA(cudaMemcpyAsync(p.st, st.data(), sizeof(ChainState) * nchains, cudaMemcpyHostToDevice, s));
  if ((rc = fcm_chain_launch_replicate(e))) return fail(rc);
  E_CUDA(cudaStreamSynchronize(s));
  lap("device allocation, uploads, replication");
#undef E_CUDA
  e->pair_ab_host = std::move(pair_ab);
  *out = e;
  return FCM_OK;
}

extern "C" int fcm_ensemble_set_bounds(fcm_ensemble* e, const uint64_t* mn, int nmin, const uint64_t* mx, int nmax) {
  REQUIRE(e, "ensemble is null");
  return copy_bounds(e->p, mn, nmin, mx, nmax);
}

extern "C" int fcm_ensemble_set_kernel(fcm_ensemble* e, int kind) {
  REQUIRE(e, "ensemble is null");
  REQUIRE(kind >= 0 && kind <= 2, "kernel kind must be 0 (auto), 1 (full pre/post counts) or 2 (delta path from the adjacency rows)");
  e->force_full = kind == 1;
  e->force_rows = kind == 2;
  return FCM_OK;
}

extern "C" int fcm_ensemble_set_policy(fcm_ensemble* e, int policy) {
  REQUIRE(e, "ensemble is null");
  REQUIRE(policy == 0 || policy == 1, "sampler policy must be 0 (element r) or 1 (element len-1-r)");
  e->p.sample_policy = (u32)policy;
  return FCM_OK;
}

extern "C" int fcm_ensemble_set_stats(fcm_ensemble* e, int on) {
  REQUIRE(e, "ensemble is null");
  e->stats = on != 0;
  return FCM_OK;
}

extern "C" int fcm_ensemble_index_info(const fcm_ensemble* e, int* have_index, uint64_t* index_bytes) {
  REQUIRE(e, "ensemble is null");
  if (have_index) *have_index = e->have_index;
  if (index_bytes) *index_bytes = e->index_bytes;
  return FCM_OK;
}

extern "C" int fcm_ensemble_info(const fcm_ensemble* e, uint32_t* n, uint64_t* npairs, uint32_t* nchains,
                                 uint32_t* maxc, uint64_t* bytes_per_chain) {
  REQUIRE(e, "ensemble is null");
  if (n) *n = e->p.n;
  if (npairs) *npairs = e->p.npairs;
  if (nchains) *nchains = e->p.nchains;
  if (maxc) *maxc = e->p.maxc;
  if (bytes_per_chain) *bytes_per_chain = e->bytes_per_chain;
  return FCM_OK;
}

extern "C" int fcm_ensemble_step(fcm_ensemble* e, uint64_t nproposals) {
  REQUIRE(e, "ensemble is null");
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  if (nproposals == 0) return FCM_OK;
  return fcm_chain_launch_step(e, nproposals, nullptr, 0);
}

static const char* status_text(int st) {
  switch (st) {
    case FCM_EDIM: return "a simplex of dimension >= FCM_MAXD was met";
    case FCM_EASSERT: return "flag_count would go negative (reference: assert!(*s >= *p), src/lib.rs:65,86)";
    case FCM_ELIMIT: return "neighbourhood larger than the local-matrix capacity (512 vertices)";
    case FCM_EINVAL: return "double_edge_move on a graph without single edges (reference loops forever, src/lib.rs:307)";
    case FCM_EUNSUPPORTED: return "clique move drawn";
    default: return "unknown device status";
  }
}

extern "C" int fcm_ensemble_step_trace(fcm_ensemble* e, uint64_t nproposals, uint32_t chain, fcm_trace_rec* trace_out) {
  REQUIRE(e && trace_out, "ensemble/trace_out is null");
  REQUIRE(chain < e->p.nchains, "chain out of range");
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  if (nproposals == 0) return FCM_OK;
  if (nproposals > e->trace_cap) {
    if (e->trace_dev) FCM_CUDA(cudaFree(e->trace_dev));
    e->trace_dev = nullptr; e->trace_cap = 0;
    FCM_CUDA(cudaMalloc(&e->trace_dev, nproposals * sizeof(fcm_trace_rec)));
    e->trace_cap = nproposals;
  }
  FCM_CUDA(cudaMemsetAsync(e->trace_dev, 0, nproposals * sizeof(fcm_trace_rec), e->ctx->stream));
  int rc = fcm_chain_launch_step(e, nproposals, e->trace_dev, chain);
  if (rc) return rc;
  FCM_CUDA(cudaMemcpyAsync(trace_out, e->trace_dev, nproposals * sizeof(fcm_trace_rec), cudaMemcpyDeviceToHost, e->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(e->ctx->stream));
  return fcm_ensemble_status(e);
}

static int fetch_states(fcm_ensemble* e, std::vector<ChainState>& st, u32 first, u32 count) {
  st.resize(count);
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  FCM_CUDA(cudaMemcpyAsync(st.data(), e->p.st + first, sizeof(ChainState) * count, cudaMemcpyDeviceToHost, e->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(e->ctx->stream));
  return FCM_OK;
}

extern "C" int fcm_ensemble_status(fcm_ensemble* e) {
  REQUIRE(e, "ensemble is null");
  std::vector<ChainState> st;
  int rc = fetch_states(e, st, 0, e->p.nchains);
  if (rc) return rc;
  for (u32 k = 0; k < e->p.nchains; k++)
    if (st[k].err) { fcm_set_error("chain %u: %s", k, status_text(st[k].err)); return st[k].err; }
  return FCM_OK;
}

extern "C" int fcm_ensemble_read(fcm_ensemble* e, uint32_t chain, uint64_t* fc, int* len, uint64_t* sampled,
                                 uint64_t* accepted, uint32_t* edges_out, uint64_t* nedges_out,
                                 uint64_t* ndouble_out, uint64_t* rng_out) {
  REQUIRE(e, "ensemble is null");
  REQUIRE(chain < e->p.nchains, "chain out of range");
  std::vector<ChainState> st;
  int rc = fetch_states(e, st, chain, 1);
  if (rc) return rc;
  const ChainState& c = st[0];
  if (fc) for (int d = 0; d < FCM_MAXD; d++) fc[d] = c.flag_count[d];
  if (len) *len = c.fc_len;
  if (sampled) *sampled = c.sampled;
  if (accepted) *accepted = c.accepted;
  if (nedges_out) *nedges_out = c.E;
  if (ndouble_out) *ndouble_out = c.ND;
  if (rng_out) memcpy(rng_out, c.rng, 32);
  if (edges_out && c.E) {
    std::vector<u32> items(c.E);
    FCM_CUDA(cudaMemcpyAsync(items.data(), e->p.items + (u64)chain * e->p.stride32, (size_t)c.E * 4, cudaMemcpyDeviceToHost, e->ctx->stream));
    FCM_CUDA(cudaStreamSynchronize(e->ctx->stream));
    for (u32 i = 0; i < c.E; i++) {
      uint2 ab = e->pair_ab_host[items[i] >> 1];
      bool up = items[i] & 1;                          // dir 1: small -> big
      edges_out[2 * i] = up ? ab.y : ab.x;
      edges_out[2 * i + 1] = up ? ab.x : ab.y;
    }
  }
  if (c.err) { fcm_set_error("chain %u: %s", chain, status_text(c.err)); return c.err; }
  return FCM_OK;
}

extern "C" int fcm_ensemble_read_all(fcm_ensemble* e, uint64_t* fcs, int* lens, uint64_t* sampled,
                                     uint64_t* accepted, uint64_t* stats) {
  REQUIRE(e, "ensemble is null");
  std::vector<ChainState> st;
  int rc = fetch_states(e, st, 0, e->p.nchains);
  if (rc) return rc;
  u64 sk = 0, ne = 0;
  int err = 0; u32 errchain = 0;
  for (u32 k = 0; k < e->p.nchains; k++) {
    if (fcs) memcpy(fcs + (size_t)k * FCM_MAXD, st[k].flag_count, FCM_MAXD * 8);
    if (lens) lens[k] = st[k].fc_len;
    if (sampled) sampled[k] = st[k].sampled;
    if (accepted) accepted[k] = st[k].accepted;
    sk += st[k].sumK;
    ne += st[k].nonempty;
    if (st[k].err && !err) { err = st[k].err; errchain = k; }
  }
  if (stats) { stats[0] = sk; stats[1] = ne; }
  if (err) { fcm_set_error("chain %u: %s", errchain, status_text(err)); return err; }
  return FCM_OK;
}

// ---- io::BitOutput (src/io.rs:128-212): one packed bit per directed slot of the underlying
//      undirected graph, slots sorted by (max, min, a < b) = pairs in (big, small) order, big -> small
//      first.  A chain's orientation bits already hold exactly these bits in PAIR-ID order, so a sample
//      is a static permutation of 2-bit fields: one thread per output word (16 pairs), packed on the
//      device, and only ceil(2*npairs/8) bytes per chain cross PCIe instead of the edge list. ----
__global__ void edgebits_pack_kernel(const EnsParams p, const u32* __restrict__ rank2pair, u32 chain_first,
                                     u32 nchains, u32 words, u32* __restrict__ out) {
  const u64 total = (u64)nchains * words;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x) {
    const u32 ch = (u32)(i / words), w = (u32)(i - (u64)ch * words);
    const u32* ori = p.ori + (u64)(chain_first + ch) * p.stride32;
    u32 v = 0;
#pragma unroll 4
    for (u32 t = 0; t < 16; t++) {
      const u64 r = (u64)w * 16 + t;
      if (r < p.npairs) {
        const u32 pair = __ldg(&rank2pair[r]);
        v |= ((__ldcg(&ori[pair >> 4]) >> (2 * (pair & 15))) & 3u) << (2 * t);
      }
    }
    out[i] = v;
  }
}

static int edgebits_prepare(fcm_ensemble* e) {
  if (e->eb_rank2pair_dev || e->p.npairs == 0) return FCM_OK;
  const u64 np = e->p.npairs;
  e->eb_rank2pair.resize(np);
  for (u64 q = 0; q < np; q++) e->eb_rank2pair[q] = (u32)q;
  const std::vector<uint2>& ab = e->pair_ab_host;
  std::sort(e->eb_rank2pair.begin(), e->eb_rank2pair.end(), [&](u32 x, u32 y) {
    return ab[x].x != ab[y].x ? ab[x].x < ab[y].x : ab[x].y < ab[y].y;
  });
  int rc = dev_alloc(e, &e->eb_rank2pair_dev, np);
  if (rc) return rc;
  FCM_CUDA(cudaMemcpyAsync(e->eb_rank2pair_dev, e->eb_rank2pair.data(), np * 4, cudaMemcpyHostToDevice, e->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(e->ctx->stream));
  return FCM_OK;
}

extern "C" int fcm_ensemble_edgebits_info(const fcm_ensemble* e, uint64_t* nslots, uint64_t* nbytes) {
  REQUIRE(e, "ensemble is null");
  if (nslots) *nslots = 2 * e->p.npairs;
  if (nbytes) *nbytes = (2 * e->p.npairs + 7) / 8;
  return FCM_OK;
}

extern "C" int fcm_ensemble_edgebits_slots(fcm_ensemble* e, uint32_t* slots_out) {
  REQUIRE(e && slots_out, "ensemble/slots_out is null");
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  int rc = edgebits_prepare(e);
  if (rc) return rc;
  for (u64 r = 0; r < e->p.npairs; r++) {
    const uint2 ab = e->pair_ab_host[e->eb_rank2pair[r]];
    slots_out[4 * r] = ab.x; slots_out[4 * r + 1] = ab.y;       // a > b: key (.., false) sorts first
    slots_out[4 * r + 2] = ab.y; slots_out[4 * r + 3] = ab.x;
  }
  return FCM_OK;
}

extern "C" int fcm_ensemble_edgebits(fcm_ensemble* e, uint32_t chain_first, uint32_t nchains, uint8_t* out) {
  REQUIRE(e && out, "ensemble/out is null");
  REQUIRE(nchains >= 1 && (u64)chain_first + nchains <= e->p.nchains, "chain range out of bounds");
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  const u64 nbytes = (2 * e->p.npairs + 7) / 8;
  if (nbytes == 0) return FCM_OK;
  int rc = edgebits_prepare(e);
  if (rc) return rc;
  const u32 words = (u32)((nbytes + 3) / 4);
  if (nchains > e->eb_out_cap) {
    if (e->eb_out_dev) FCM_CUDA(cudaFree(e->eb_out_dev));
    e->eb_out_dev = nullptr; e->eb_out_cap = 0;
    FCM_CUDA(cudaMalloc(&e->eb_out_dev, (size_t)nchains * words * 4));
    e->eb_out_cap = nchains;
  }
  cudaStream_t s = e->ctx->stream;
  const u64 total = (u64)nchains * words;
  const int blocks = (int)std::min<u64>((total + 255) / 256, (u64)e->ctx->sm_count * 8);
  edgebits_pack_kernel<<<blocks, 256, 0, s>>>(e->p, e->eb_rank2pair_dev, chain_first, nchains, words, e->eb_out_dev);
  e->ctx->launches++;
  FCM_CUDA(cudaGetLastError());
  // rows of `words` 32-bit words on the device, rows of nbytes bytes for the caller (little endian: LSB-first bytes)
  FCM_CUDA(cudaMemcpy2DAsync(out, nbytes, e->eb_out_dev, (size_t)words * 4, nbytes, nchains, cudaMemcpyDeviceToHost, s));
  FCM_CUDA(cudaStreamSynchronize(s));
  return FCM_OK;
}

// ---- checkpoint / resume of one chain: the device analogue of io::save_state / load_state
//      (src/io.rs:51-62).  What a chain IS beyond the static structures: its two edge containers in
//      order (that order decides every later sample_edge), the RNG state and the counters. ----
extern "C" int fcm_ensemble_read_doubles(fcm_ensemble* e, uint32_t chain, uint32_t* pairs_out, uint64_t* ndouble_out) {
  REQUIRE(e && ndouble_out, "ensemble/ndouble_out is null");
  REQUIRE(chain < e->p.nchains, "chain out of range");
  std::vector<ChainState> st;
  int rc = fetch_states(e, st, chain, 1);
  if (rc) return rc;
  const u32 nd = st[0].ND;
  *ndouble_out = nd;
  if (pairs_out && nd) {
    std::vector<u32> dbl(nd);
    FCM_CUDA(cudaMemcpyAsync(dbl.data(), e->p.dbl + (u64)chain * e->p.stride32, (size_t)nd * 4, cudaMemcpyDeviceToHost, e->ctx->stream));
    FCM_CUDA(cudaStreamSynchronize(e->ctx->stream));
    for (u32 i = 0; i < nd; i++) {                     // [min, max] as sample_double_edge returns it
      const uint2 ab = e->pair_ab_host[dbl[i]];
      pairs_out[2 * i] = ab.y; pairs_out[2 * i + 1] = ab.x;
    }
  }
  return FCM_OK;
}

extern "C" int fcm_ensemble_restore(fcm_ensemble* e, uint32_t chain, const uint32_t* edges, uint64_t nedges,
                                    const uint32_t* doubles, uint64_t ndouble, const uint64_t* rng4,
                                    const uint64_t* flag_count, int len, uint64_t sampled, uint64_t accepted) {
  REQUIRE(e && rng4 && flag_count, "ensemble/rng4/flag_count is null");
  REQUIRE(chain < e->p.nchains, "chain out of range");
  REQUIRE((nedges == 0 || edges) && (ndouble == 0 || doubles), "edges/doubles is null");
  REQUIRE(len >= 0 && len <= FCM_MAXD, "flag_count longer than FCM_MAXD");
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  const EnsParams& p = e->p;
  const u64 nslots = 2 * p.npairs;
  std::vector<u32> blk(p.stride32, 0);
  u32* items = blk.data() + p.off_items;
  u32* pos = blk.data() + p.off_pos;
  u32* dbl = blk.data() + p.off_dbl;
  u32* dblpos = blk.data() + p.off_dblpos;
  u32* ori = blk.data() + p.off_ori;
  u64* adj = (u64*)blk.data();
  for (u64 i = 0; i < nslots; i++) pos[i] = FCM_NONE;
  for (u64 i = 0; i < p.npairs; i++) dblpos[i] = FCM_NONE;
  REQUIRE(nedges <= nslots, "more edges than directed slots");
  for (u64 i = 0; i < nedges; i++) {
    const u32 a = edges[2 * i], b = edges[2 * i + 1];
    auto it = a == b ? e->pairmap.end() : e->pairmap.find(pkey(std::max(a, b), std::min(a, b)));
    if (it == e->pairmap.end()) { fcm_set_error("restore: edge %llu = (%u,%u) is not a pair of the underlying undirected graph", (unsigned long long)i, a, b); return FCM_ENOTADJ; }
    const u32 slot = 2 * it->second + (a > b ? 0u : 1u);
    if (pos[slot] != FCM_NONE) { fcm_set_error("restore: edge (%u,%u) listed twice", a, b); return FCM_EINVAL; }
    pos[slot] = (u32)i; items[i] = slot;
    adj[(size_t)a * p.W + (b >> 6)] |= 1ULL << (b & 63);
    ori[slot >> 5] |= 1u << (slot & 31);
  }
  for (u64 q = 0; q < p.npairs; q++)
    if (pos[2 * q] == FCM_NONE && pos[2 * q + 1] == FCM_NONE) { fcm_set_error("restore: pair (%u,%u) has lost both directions", e->pair_ab_host[q].x, e->pair_ab_host[q].y); return FCM_EINVAL; }
  for (u64 i = 0; i < ndouble; i++) {
    const u32 a = doubles[2 * i], b = doubles[2 * i + 1];
    auto it = a == b ? e->pairmap.end() : e->pairmap.find(pkey(std::max(a, b), std::min(a, b)));
    if (it == e->pairmap.end() || pos[2 * it->second] == FCM_NONE || pos[2 * it->second + 1] == FCM_NONE || dblpos[it->second] != FCM_NONE) {
      fcm_set_error("restore: double-edge entry %llu = (%u,%u) is not a double edge of the edge list (or listed twice)", (unsigned long long)i, a, b);
      return FCM_EINVAL;
    }
    dblpos[it->second] = (u32)i; dbl[i] = it->second;
  }
  u64 have_double = 0;
  for (u64 q = 0; q < p.npairs; q++) have_double += pos[2 * q] != FCM_NONE && pos[2 * q + 1] != FCM_NONE;
  if (have_double != ndouble) { fcm_set_error("restore: the edge list has %llu double edges, the double-edge list %llu", (unsigned long long)have_double, (unsigned long long)ndouble); return FCM_EINVAL; }
  ChainState cs;
  memset(&cs, 0, sizeof(cs));
  memcpy(cs.rng, rng4, 32);
  cs.sampled = sampled; cs.accepted = accepted;
  for (int d = 0; d < len; d++) cs.flag_count[d] = flag_count[d];
  cs.fc_len = len; cs.E = (u32)nedges; cs.ND = (u32)ndouble;
  cudaStream_t s = e->ctx->stream;
  FCM_CUDA(cudaStreamSynchronize(s));
  FCM_CUDA(cudaMemcpyAsync((u32*)p.adj + (u64)chain * p.stride32, blk.data(), p.stride32 * 4, cudaMemcpyHostToDevice, s));
  FCM_CUDA(cudaMemcpyAsync(p.st + chain, &cs, sizeof(cs), cudaMemcpyHostToDevice, s));
  FCM_CUDA(cudaStreamSynchronize(s));
  return FCM_OK;
}

// ---- explicit-edit API: State::apply_transition / revert_transition (src/lib.rs:61-95) ----
static int upload_edits(fcm_ensemble* e, const uint32_t* edits3, int nedits) {
  REQUIRE(nedits >= 0 && nedits <= FCM_MAXEDIT, "too many edits (FCM_MAXEDIT = 64)");
  REQUIRE(nedits == 0 || edits3, "edits is null");
  u32 buf[2 * FCM_MAXEDIT];
  for (int i = 0; i < nedits; i++) {
    u32 a = edits3[3 * i], b = edits3[3 * i + 1];
    auto it = a == b ? e->pairmap.end() : e->pairmap.find(pkey(std::max(a, b), std::min(a, b)));
    if (it == e->pairmap.end()) {
      fcm_set_error("edit %d: vertices %u and %u are not adjacent in the underlying undirected graph", i, a, b);
      return FCM_ENOTADJ;
    }
    buf[2 * i] = 2 * it->second + (a > b ? 0u : 1u);
    buf[2 * i + 1] = edits3[3 * i + 2] ? 1u : 0u;
  }
  if (nedits) FCM_CUDA(cudaMemcpyAsync(e->edit_dev, buf, (size_t)nedits * 8, cudaMemcpyHostToDevice, e->ctx->stream));
  return FCM_OK;
}

extern "C" int fcm_apply(fcm_ensemble* e, uint32_t chain, const uint32_t* edits3, int nedits,
                         uint64_t* pre, int* prelen, uint64_t* post, int* postlen) {
  REQUIRE(e && pre && prelen && post && postlen, "null argument");
  REQUIRE(chain < e->p.nchains, "chain out of range");
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  int rc = upload_edits(e, edits3, nedits);
  if (rc) return rc;
  if ((rc = fcm_chain_launch_apply(e, chain, nedits, 0))) return rc;
  u64 io[2 * FCM_MAXD + 8];
  FCM_CUDA(cudaMemcpyAsync(io, e->io_dev, sizeof(io), cudaMemcpyDeviceToHost, e->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(e->ctx->stream));
  int st = (int)(long long)io[66];
  if (st) { fcm_set_error("fcm_apply: %s", status_text(st)); return st; }
  memcpy(pre, io, FCM_MAXD * 8);
  memcpy(post, io + FCM_MAXD, FCM_MAXD * 8);
  *prelen = (int)io[64]; *postlen = (int)io[65];
  return FCM_OK;
}

extern "C" int fcm_revert(fcm_ensemble* e, uint32_t chain, const uint32_t* edits3, int nedits,
                          const uint64_t* pre, int prelen, const uint64_t* post, int postlen) {
  REQUIRE(e, "ensemble is null");
  REQUIRE(chain < e->p.nchains, "chain out of range");
  REQUIRE(prelen >= 0 && prelen <= FCM_MAXD && postlen >= 0 && postlen <= FCM_MAXD, "pre/post longer than FCM_MAXD");
  REQUIRE((prelen == 0 || pre) && (postlen == 0 || post), "pre/post is null");
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  int rc = upload_edits(e, edits3, nedits);
  if (rc) return rc;
  u64 io[2 * FCM_MAXD + 8];
  memset(io, 0, sizeof(io));
  for (int d = 0; d < prelen; d++) io[d] = pre[d];
  for (int d = 0; d < postlen; d++) io[FCM_MAXD + d] = post[d];
  io[64] = prelen; io[65] = postlen;
  FCM_CUDA(cudaMemcpyAsync(e->io_dev, io, sizeof(io), cudaMemcpyHostToDevice, e->ctx->stream));
  if ((rc = fcm_chain_launch_apply(e, chain, nedits, 1))) return rc;
  FCM_CUDA(cudaMemcpyAsync(io, e->io_dev, sizeof(io), cudaMemcpyDeviceToHost, e->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(e->ctx->stream));
  int st = (int)(long long)io[66];
  if (st) { fcm_set_error("fcm_revert: %s", status_text(st)); return st; }
  return FCM_OK;
}

extern "C" int fcm_bounds_check(fcm_ensemble* e, uint32_t chain, int* ok) {
  REQUIRE(e && ok, "null argument");
  REQUIRE(chain < e->p.nchains, "chain out of range");
  FCM_CUDA(cudaSetDevice(e->ctx->device));
  int rc = fcm_chain_launch_apply(e, chain, 0, 2);
  if (rc) return rc;
  u64 io[2 * FCM_MAXD + 8];
  FCM_CUDA(cudaMemcpyAsync(io, e->io_dev, sizeof(io), cudaMemcpyDeviceToHost, e->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(e->ctx->stream));
  *ok = (int)io[67];
  return FCM_OK;
}
