// fcm_chain.cu — MCMC chains on sm_100a: one warp per chain, many chains per launch.
//
// Replaces, for the simple moves, the loop body of MCMCSampler::next (src/lib.rs:182-192):
//   Transition::random_move / single_edge_flip / double_edge_move   src/lib.rs:207-212,292-325
//   State::apply_transition / revert_transition / edgeset_neighborhood   src/lib.rs:61-111
//   Bounds::check / util::all_le                                    src/lib.rs:157-160, src/util.rs:53-63
// plus the rand 0.8 / rand_xoshiro 0.6 primitives the proposals draw from and the edge container
// policy documented in oracle/fcm_oracle.cpp (struct EdgeSet): insertion-ordered index set,
// sample = element at gen_range(0..len), remove = swap_remove, add = append.
//
// Per proposal a warp
//   1. draws the move (uniform across lanes: every lane runs the same xoshiro256** stream),
//   2. builds the neighbourhood = common undirected neighbours of every changed pair + endpoints
//      (static CSR, the undirected graph is invariant under all moves),
//   3. gathers the K x K induced adjacency from the chain's packed 64-bit rows in HBM: lane j probes
//      column verts[j] of row verts[i], one ballot per 32 columns,
//   4. counts the local directed simplices before and after the edits (one DFS root per lane,
//      popc on K-bit candidate sets held as KW 64-bit words),
//   5. updates flag_count, tests the bounds, and commits or reverts the container/bit edits.
#include "fcm_internal.cuh"

#include <algorithm>

#define FULL 0xffffffffu
#define MAXL_CHAIN (FCM_MAXD - 1)     // level l holds candidates for vertex index l+1
static_assert(FCM_MAXD == 32, "flag_count is held one entry per lane");

struct Rng {
  u64 s0, s1, s2, s3;
  __device__ __forceinline__ static u64 rotl(u64 x, int k) { return (x << k) | (x >> (64 - k)); }
  __device__ __forceinline__ u64 next() {              // xoshiro256**
    u64 r = rotl(s1 * 5, 7) * 9;
    u64 t = s1 << 17;
    s2 ^= s0; s3 ^= s1; s1 ^= s2; s0 ^= s3;
    s2 ^= t;
    s3 = rotl(s3, 45);
    return r;
  }
};

// rand 0.8 UniformInt<u64>::sample_single(0, range): widening multiply, rejection zone
__device__ __forceinline__ u64 gen_range(Rng& g, u64 range) {
  u64 zone = (range << __clzll((long long)range)) - 1;
  for (;;) {
    u64 v = g.next();
    u64 hi = __umul64hi(v, range), lo = v * range;
    if (lo <= zone) return hi;
  }
}

struct Chain {                        // pointers of one chain + live scalars
  u64* adj;
  u32 *items, *pos, *dbl, *dblpos;
  u32 E, ND;
};

__device__ __forceinline__ uint2 slot_edge(const EnsParams& p, u32 slot) {   // (from, to)
  uint2 ab = p.pair_ab[slot >> 1];                     // (big, small)
  return (slot & 1) ? make_uint2(ab.y, ab.x) : ab;     // dir 0: big -> small, dir 1: small -> big
}

// Graph::add_edge / remove_edge under the container policy (all lanes run it; lane 0 stores)
__device__ __forceinline__ void c_add_edge(Chain& c, u32 slot, int lane) {
  if (c.pos[slot] == FCM_NONE) {
    if (lane == 0) { c.items[c.E] = slot; c.pos[slot] = c.E; }
    c.E++;
  }
  if (c.pos[slot ^ 1] != FCM_NONE) {
    u32 pair = slot >> 1;
    if (c.dblpos[pair] == FCM_NONE) {
      if (lane == 0) { c.dbl[c.ND] = pair; c.dblpos[pair] = c.ND; }
      c.ND++;
    }
  }
  __syncwarp();
}
__device__ __forceinline__ void c_remove_edge(Chain& c, u32 slot, int lane) {
  u32 q = c.pos[slot];
  if (q != FCM_NONE) {
    u32 last = c.items[c.E - 1];
    if (lane == 0) { c.items[q] = last; c.pos[last] = q; c.pos[slot] = FCM_NONE; }
    c.E--;
  }
  u32 pair = slot >> 1;
  u32 r = c.dblpos[pair];
  if (r != FCM_NONE) {
    u32 lastp = c.dbl[c.ND - 1];
    if (lane == 0) { c.dbl[r] = lastp; c.dblpos[lastp] = r; c.dblpos[pair] = FCM_NONE; }
    c.ND--;
  }
  __syncwarp();
}
__device__ __forceinline__ void c_set_bit(const EnsParams& p, Chain& c, u32 slot, bool add, int lane) {
  if (lane == 0) {
    uint2 e = slot_edge(p, slot);
    u64* w = &c.adj[(u64)e.x * p.W + (e.y >> 6)];
    u64 m = 1ULL << (e.y & 63);
    *w = add ? (*w | m) : (*w & ~m);
  }
}

// index of vertex x in verts[0..K) or -1 (warp-uniform)
__device__ __forceinline__ int find_local(const u32* verts, int K, u32 x, int lane) {
  for (int base = 0; base < K; base += 32) {
    int j = base + lane;
    u32 b = __ballot_sync(FULL, j < K && verts[j] == x);
    if (b) return base + __ffs(b) - 1;
  }
  return -1;
}

// State::edgeset_neighborhood (src/lib.rs:99-111) without the sort: counts are label-invariant,
// only the SET matters.  Returns false when the set does not fit Kcap.
__device__ bool build_nbhd(const EnsParams& p, const u32* edits, int nedit, u32* verts, int& K, int lane) {
  K = 0;
  for (int e = 0; e < nedit; e++) {
    const u32 pair = edits[2 * e] >> 1;
    bool seen = false;
    for (int f = 0; f < e; f++) seen |= (edits[2 * f] >> 1) == pair;
    if (seen) continue;                                // same list, same endpoints: nothing new
    const u64 o = p.nb_off[pair];
    const int c = (int)(p.nb_off[pair + 1] - o);
    const int K0 = K;
    for (int base = 0; base < c; base += 32) {
      int t = base + lane;
      bool keep = t < c;
      u32 x = keep ? __ldg(&p.nb_list[o + t]) : 0;
      for (int s = 0; s < K0; s++) keep &= verts[s] != x;   // broadcast reads
      u32 b = __ballot_sync(FULL, keep);
      int add = __popc(b);
      if (K + add > (int)p.Kcap) return false;
      if (keep) verts[K + __popc(b & ((1u << lane) - 1))] = x;
      K += add;
    }
    __syncwarp();
    uint2 ab = p.pair_ab[pair];
    for (int q = 0; q < 2; q++) {
      u32 x = q ? ab.y : ab.x;
      if (find_local(verts, K, x, lane) < 0) {
        if (K + 1 > (int)p.Kcap) return false;
        if (lane == 0) verts[K] = x;
        K++;
        __syncwarp();
      }
    }
  }
  return true;
}

// K x K induced adjacency of the chain's graph on verts -> M[i*KW + w] (bit j of row i = edge i->j)
__device__ void gather_local(const EnsParams& p, const u64* adj, const u32* verts, int K, int KW,
                             u64* M, int lane) {
  u32* M32 = (u32*)M;
  const int nch = (K + 31) >> 5;
  if (nch < 2 * KW) {                                  // columns beyond the last ballot chunk
    for (int i = lane; i < K * KW; i += 32) M[i] = 0;
    __syncwarp();
  }
  for (int c = 0; c < nch; c++) {
    const int j = c * 32 + lane;
    const bool have = j < K;
    const u32 gv = have ? verts[j] : 0;
    const u32 gw = gv >> 6, gs = gv & 63;
#pragma unroll 4
    for (int i = 0; i < K; i++) {
      const u64* row = adj + (u64)verts[i] * p.W;
      bool bit = have && ((row[gw] >> gs) & 1ULL);
      u32 b = __ballot_sync(FULL, bit);
      if (lane == 0) M32[i * 2 * KW + c] = b;
    }
  }
  __syncwarp();
}

// flagser_count of the K-vertex local graph: out[0..FCM_MAXD) zero padded, returns length.
template <int KW>
__device__ int local_count(const u64* M, int K, int lane, u64* out, bool& dim_overflow) {
  u64 cnt[FCM_MAXD];
#pragma unroll
  for (int d = 0; d < FCM_MAXD; d++) cnt[d] = 0;
  int maxd = 0;
  for (int v0 = lane; v0 < K; v0 += 32) {
    u64 P[MAXL_CHAIN][KW];
    int cw[MAXL_CHAIN];
    u64 cb[MAXL_CHAIN];
    u32 pc = 0;
#pragma unroll
    for (int x = 0; x < KW; x++) { u64 q = M[v0 * KW + x]; P[0][x] = q; pc += __popcll(q); }
    if (!pc) continue;
    cnt[1] += pc;
    if (maxd < 1) maxd = 1;
    int l = 0, w = 0;
    u64 bits = P[0][0];
    for (;;) {
      while (bits == 0) { if (++w == KW) break; bits = P[l][w]; }
      if (w == KW) {
        if (l == 0) break;
        --l; w = cw[l]; bits = cb[l];
        continue;
      }
      const int j = w * 64 + (__ffsll((long long)bits) - 1);
      bits &= bits - 1;
      const u64* Mj = M + j * KW;
      u32 cpc = 0;
      if (l + 1 < MAXL_CHAIN) {
#pragma unroll
        for (int x = 0; x < KW; x++) { u64 q = P[l][x] & Mj[x]; P[l + 1][x] = q; cpc += __popcll(q); }
        if (cpc) {
          cnt[l + 2] += cpc;
          if (maxd < l + 2) maxd = l + 2;
          cw[l] = w; cb[l] = bits; ++l; w = 0; bits = P[l][0];
        }
      } else {
#pragma unroll
        for (int x = 0; x < KW; x++) cpc += __popcll(P[l][x] & Mj[x]);
        if (cpc) dim_overflow = true;
      }
    }
  }
  const int wmax = __reduce_max_sync(FULL, maxd);
  out[lane] = 0;                                       // FCM_MAXD == 32 == warp size
  __syncwarp();
  for (int d = 1; d <= wmax; d++) {
    u64 v = cnt[d];
    u32 lo = (u32)v, hi = (u32)(v >> 32);
    u64 tot = (u64)__reduce_add_sync(FULL, lo & 0xffffu) + ((u64)__reduce_add_sync(FULL, lo >> 16) << 16);
    if (__any_sync(FULL, hi != 0))
      tot += ((u64)__reduce_add_sync(FULL, hi & 0xffffu) << 32) + ((u64)__reduce_add_sync(FULL, hi >> 16) << 48);
    if (lane == 0) out[d] = tot;
  }
  if (lane == 0) out[0] = (u64)K;
  __syncwarp();
  return K ? wmax + 1 : 0;
}

__device__ int local_count_any(const u64* M, int K, int KW, int lane, u64* out, bool& dim_overflow) {
  switch (KW) {
    case 1: return local_count<1>(M, K, lane, out, dim_overflow);
    case 2: return local_count<2>(M, K, lane, out, dim_overflow);
    case 4: return local_count<4>(M, K, lane, out, dim_overflow);
    default: return local_count<8>(M, K, lane, out, dim_overflow);
  }
}

struct WarpSmem {
  u64 *M, *fc, *pre, *post;
  u32 *verts, *edits;
};
__host__ __device__ inline size_t warp_smem_bytes(u32 Kcap, u32 KWcap) {
  return (size_t)Kcap * KWcap * 8 + 3 * FCM_MAXD * 8 + (size_t)Kcap * 4 + FCM_MAXEDIT * 2 * 4;
}
size_t fcm_chain_smem_per_warp(u32 Kcap, u32 KWcap) { return (warp_smem_bytes(Kcap, KWcap) + 15) & ~(size_t)15; }

__device__ __forceinline__ WarpSmem carve(unsigned char* base, u32 Kcap, u32 KWcap) {
  WarpSmem s;
  s.M = (u64*)base;
  s.fc = s.M + (size_t)Kcap * KWcap;
  s.pre = s.fc + FCM_MAXD;
  s.post = s.pre + FCM_MAXD;
  s.verts = (u32*)(s.post + FCM_MAXD);
  s.edits = s.verts + Kcap;
  return s;
}

// neighbourhood + pre/post local counts of the edits in s.edits (slot, add) — the two
// flagser_count calls of State::apply_transition (src/lib.rs:62-71).  0 or an FCM_E* code.
__device__ int transition_counts(const EnsParams& p, const Chain& c, const WarpSmem& s, int nedit,
                                 int lane, int& K, int& prelen, int& postlen) {
  if (!build_nbhd(p, s.edits, nedit, s.verts, K, lane)) return FCM_ELIMIT;
  int KW = (K + 63) >> 6;
  KW = KW <= 1 ? 1 : KW <= 2 ? 2 : KW <= 4 ? 4 : 8;
  gather_local(p, c.adj, s.verts, K, KW, s.M, lane);
  bool dimov = false;
  prelen = local_count_any(s.M, K, KW, lane, s.pre, dimov);
  for (int e = 0; e < nedit; e++) {
    uint2 ft = slot_edge(p, s.edits[2 * e]);
    int la = find_local(s.verts, K, ft.x, lane), lb = find_local(s.verts, K, ft.y, lane);
    if (lane == 0) {
      u64 m = 1ULL << (lb & 63);
      u64* w = &s.M[la * KW + (lb >> 6)];
      *w = s.edits[2 * e + 1] ? (*w | m) : (*w & ~m);
    }
    __syncwarp();
  }
  postlen = local_count_any(s.M, K, KW, lane, s.post, dimov);
  return dimov ? FCM_EDIM : FCM_OK;
}

// Bounds::check (src/lib.rs:157-160) on the flag_count held one entry per lane
__device__ __forceinline__ bool bounds_ok(const EnsParams& p, u64 f, int lane) {
  u64 lo = lane < p.nmin ? p.bmin[lane] : 0, hi = lane < p.nmax ? p.bmax[lane] : 0;
  bool bad = lo > f || f > hi;
  if (lane == 0 && p.nmin > FCM_MAXD && p.bmin[FCM_MAXD] > 0) bad = true;   // entry 32: count is 0 there
  return !__any_sync(FULL, bad);
}

__global__ void __launch_bounds__(128)
chain_step_kernel(const EnsParams p, u64 nsteps, fcm_trace_rec* __restrict__ trace, u32 trace_chain,
                  u32 smem_per_warp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const u32 chain = blockIdx.x * (blockDim.x >> 5) + wib;
  if (chain >= p.nchains) return;
  WarpSmem s = carve(smem_raw + (size_t)wib * smem_per_warp, p.Kcap, p.KWcap);
  ChainState* st = &p.st[chain];
  if (st->err) return;
  Chain c;
  const u64 nslots = 2 * p.npairs;
  c.adj = p.adj + (u64)chain * p.n * p.W;
  c.items = p.items + (u64)chain * nslots;
  c.pos = p.pos + (u64)chain * nslots;
  c.dbl = p.dbl + (u64)chain * p.npairs;
  c.dblpos = p.dblpos + (u64)chain * p.npairs;
  c.E = st->E; c.ND = st->ND;
  Rng g{st->rng[0], st->rng[1], st->rng[2], st->rng[3]};
  u64 fc = st->flag_count[lane];                       // lane d owns flag_count[d]
  int len = st->fc_len;
  u64 sampled = st->sampled, accepted = st->accepted, sumK = st->sumK, nonempty = st->nonempty;
  int err = 0;
  const bool tracing = trace != nullptr && chain == trace_chain;

  for (u64 step = 0; step < nsteps; step++) {
    // ---- Transition::random_move: WeightedIndex<f64>::sample (src/lib.rs:210) ----
    u64 r = g.next();
    double v = __longlong_as_double((long long)((r >> 12) | 0x3ff0000000000000ULL)) - 1.0;
    double chosen = __dadd_rn(__dmul_rn(v, p.scale), 0.0);
    int kind = (p.cum[0] <= chosen) + (p.cum[1] <= chosen) + (p.cum[2] <= chosen);
    int nedit = 0;
    u32 ta = 0, tb = 0, tdm0 = 0, tdm1 = 0;
    if (kind == 0) {                                   // single_edge_flip  src/lib.rs:292-299
      if (c.E) {
        u32 slot = c.items[gen_range(g, c.E)];
        uint2 ft = slot_edge(p, slot);
        ta = ft.x; tb = ft.y;
        if (c.pos[slot ^ 1] == FCM_NONE) {
          if (lane == 0) { s.edits[0] = slot; s.edits[1] = 0; s.edits[2] = slot ^ 1; s.edits[3] = 1; }
          nedit = 2;
        }
      }
    } else if (kind == 1) {                            // double_edge_move  src/lib.rs:304-325
      if (c.ND) {
        if (2ULL * c.ND >= c.E) { err = FCM_EINVAL; break; }   // reference spins forever (FIXME :307)
        u32 dpair = c.dbl[gen_range(g, c.ND)];
        u32 slot;
        for (;;) {
          slot = c.items[gen_range(g, c.E)];
          if (c.pos[slot ^ 1] == FCM_NONE) break;
        }
        bool first = g.next() < 0x8000000000000000ULL;  // gen_bool(0.5)
        u32 delme = 2 * dpair + (first ? 1u : 0u);      // first: [min,max] = small -> big
        uint2 ft = slot_edge(p, slot), dm = slot_edge(p, delme);
        ta = ft.x; tb = ft.y; tdm0 = dm.x; tdm1 = dm.y;
        if (lane == 0) { s.edits[0] = slot ^ 1; s.edits[1] = 1; s.edits[2] = delme; s.edits[3] = 0; }
        nedit = 2;
      }
    } else { err = FCM_EUNSUPPORTED; break; }
    __syncwarp();

    // ---- State::apply_transition  src/lib.rs:61-79 ----
    int K = 0, prelen = 0, postlen = 0;
    u64 fnew = fc;
    if (nedit) {
      int rc = transition_counts(p, c, s, nedit, lane, K, prelen, postlen);
      if (rc) { err = rc; break; }
      u64 pr = lane < min(prelen, len) ? s.pre[lane] : 0;
      if (__any_sync(FULL, fc < pr)) { err = FCM_EASSERT; break; }
      len = max(len, postlen);
      fnew = fc - pr + (lane < postlen ? s.post[lane] : 0);
      for (int e = 0; e < nedit; e++) {
        if (s.edits[2 * e + 1]) c_add_edge(c, s.edits[2 * e], lane); else c_remove_edge(c, s.edits[2 * e], lane);
      }
      sumK += K;
      nonempty++;
    } else {
      // empty transition: pre = post = count of the empty graph (length 0), state unchanged
    }
    sampled++;
    const bool ok = bounds_ok(p, fnew, lane);
    if (ok) {
      accepted++;
      if (nedit) {
        for (int e = 0; e < nedit; e++) c_set_bit(p, c, s.edits[2 * e], s.edits[2 * e + 1] != 0, lane);
        __syncwarp();
        fc = fnew;
      }
    } else if (nedit) {                                // revert_transition  src/lib.rs:81-95
      for (int e = 0; e < nedit; e++) {
        if (s.edits[2 * e + 1]) c_remove_edge(c, s.edits[2 * e], lane); else c_add_edge(c, s.edits[2 * e], lane);
      }
      len = max(len, prelen);
    }
    if (tracing && lane == 0) {
      fcm_trace_rec tr;
      tr.flags = (u32)kind | (ok ? FCM_TR_ACCEPT : 0u) | (nedit ? 0u : FCM_TR_EMPTY);
      tr.a = ta; tr.b = tb; tr.dm0 = tdm0; tr.dm1 = tdm1; tr.K = (u32)K; tr.draws = 0; tr.pad = 0;
      trace[step] = tr;
    }
  }
  __syncwarp();
  st->flag_count[lane] = fc;
  if (lane == 0) {
    st->rng[0] = g.s0; st->rng[1] = g.s1; st->rng[2] = g.s2; st->rng[3] = g.s3;
    st->E = c.E; st->ND = c.ND; st->fc_len = len;
    st->sampled = sampled; st->accepted = accepted; st->sumK = sumK; st->nonempty = nonempty;
    st->err = err;
  }
}

// State::apply_transition / revert_transition with explicit edits on one chain (one warp).
// io: [0..32) pre, [32..64) post, [64] prelen, [65] postlen, [66] status, [67] bounds ok
__global__ void chain_apply_kernel(const EnsParams p, u32 chain, const u32* __restrict__ edits_in,
                                   int nedit, int mode, u64* __restrict__ io) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31;
  WarpSmem s = carve(smem_raw, p.Kcap, p.KWcap);
  ChainState* st = &p.st[chain];
  Chain c;
  const u64 nslots = 2 * p.npairs;
  c.adj = p.adj + (u64)chain * p.n * p.W;
  c.items = p.items + (u64)chain * nslots;
  c.pos = p.pos + (u64)chain * nslots;
  c.dbl = p.dbl + (u64)chain * p.npairs;
  c.dblpos = p.dblpos + (u64)chain * p.npairs;
  c.E = st->E; c.ND = st->ND;
  u64 fc = st->flag_count[lane];
  int len = st->fc_len;
  if (mode == 2) {                                     // Bounds::check only
    bool ok = bounds_ok(p, fc, lane);
    if (lane == 0) { io[66] = 0; io[67] = ok; }
    return;
  }
  for (int i = lane; i < 2 * nedit; i += 32) s.edits[i] = edits_in[i];
  __syncwarp();
  int status = 0;
  if (mode == 0) {                                     // apply
    int K = 0, prelen = 0, postlen = 0;
    if (nedit) status = transition_counts(p, c, s, nedit, lane, K, prelen, postlen);
    if (!status) {
      u64 pr = lane < min(prelen, len) ? s.pre[lane] : 0;
      if (__any_sync(FULL, fc < pr)) status = FCM_EASSERT;
      else {
        len = max(len, postlen);
        fc = fc - pr + (lane < postlen ? s.post[lane] : 0);
        for (int e = 0; e < nedit; e++) {
          bool add = s.edits[2 * e + 1] != 0;
          if (add) c_add_edge(c, s.edits[2 * e], lane); else c_remove_edge(c, s.edits[2 * e], lane);
          c_set_bit(p, c, s.edits[2 * e], add, lane);
        }
        io[lane] = nedit ? s.pre[lane] : 0;
        io[32 + lane] = nedit ? s.post[lane] : 0;
        if (lane == 0) { io[64] = prelen; io[65] = postlen; }
      }
    }
  } else {                                             // revert
    int prelen = (int)io[64], postlen = (int)io[65];
    for (int e = 0; e < nedit; e++) {
      bool add = s.edits[2 * e + 1] == 0;              // !add
      if (add) c_add_edge(c, s.edits[2 * e], lane); else c_remove_edge(c, s.edits[2 * e], lane);
      c_set_bit(p, c, s.edits[2 * e], add, lane);
    }
    u64 po = lane < min(postlen, len) ? io[32 + lane] : 0;
    if (__any_sync(FULL, fc < po)) status = FCM_EASSERT;
    else {
      len = max(len, prelen);
      fc = fc - po + (lane < prelen ? io[lane] : 0);
    }
  }
  __syncwarp();
  if (!status) {
    st->flag_count[lane] = fc;
    if (lane == 0) { st->E = c.E; st->ND = c.ND; st->fc_len = len; }
  }
  if (lane == 0) io[66] = (u64)(long long)status;
}

// copy chain 0's arrays to chains 1..nchains-1
template <typename T>
__global__ void replicate_kernel(T* base, u64 per_chain, u32 nchains) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  u64 stride = (u64)gridDim.x * blockDim.x;
  for (; i < per_chain; i += stride) {
    T v = base[i];
    for (u32 ch = 1; ch < nchains; ch++) base[(u64)ch * per_chain + i] = v;
  }
}

int fcm_chain_launch_replicate(fcm_ensemble* e) {
  const EnsParams& p = e->p;
  cudaStream_t s = e->ctx->stream;
  if (p.nchains <= 1) return FCM_OK;
  auto blocks = [&](u64 cnt) { return (int)std::min<u64>((cnt + 255) / 256, (u64)e->ctx->sm_count * 8); };
  u64 nslots = 2 * p.npairs;
  if ((u64)p.n * p.W) { replicate_kernel<u64><<<blocks((u64)p.n * p.W), 256, 0, s>>>(p.adj, (u64)p.n * p.W, p.nchains); e->ctx->launches++; }
  if (nslots) {
    replicate_kernel<u32><<<blocks(nslots), 256, 0, s>>>(p.items, nslots, p.nchains);
    replicate_kernel<u32><<<blocks(nslots), 256, 0, s>>>(p.pos, nslots, p.nchains);
    replicate_kernel<u32><<<blocks(p.npairs), 256, 0, s>>>(p.dbl, p.npairs, p.nchains);
    replicate_kernel<u32><<<blocks(p.npairs), 256, 0, s>>>(p.dblpos, p.npairs, p.nchains);
    e->ctx->launches += 4;
  }
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

int fcm_chain_launch_step(fcm_ensemble* e, u64 nsteps, fcm_trace_rec* trace_dev, u32 trace_chain) {
  const int wpb = e->warps_per_block;
  size_t smem = e->smem_per_warp * wpb;
  FCM_CUDA(cudaFuncSetAttribute(chain_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  u32 blocks = (e->p.nchains + wpb - 1) / wpb;
  chain_step_kernel<<<blocks, wpb * 32, smem, e->ctx->stream>>>(e->p, nsteps, trace_dev, trace_chain,
                                                              (u32)e->smem_per_warp);
  e->ctx->launches++;
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

int fcm_chain_launch_apply(fcm_ensemble* e, u32 chain, int nedits, int mode) {
  size_t smem = e->smem_per_warp;
  FCM_CUDA(cudaFuncSetAttribute(chain_apply_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  chain_apply_kernel<<<1, 32, smem, e->ctx->stream>>>(e->p, chain, e->edit_dev, nedits, mode, e->io_dev);
  e->ctx->launches++;
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}
