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
// Two ways to obtain `flag_count += post - pre` (src/lib.rs:63-77), identical results:
//
//   DELTA path (the hot one).  post - pre only involves simplices that contain a changed edge; all
//   other simplices of the neighbourhood graph cancel.  A simplex through the edge u->v is
//   (x.., u, y.., v, z..) with every other vertex w a common neighbour of u and v, tagged by where it
//   may stand: X (w->u, w->v), Y (u->w, w->v) or Z (u->w, v->w).  The warp
//     1. reads the static common-neighbour list C of the pair (c <= 64 entries),
//     2. gathers, from the chain's packed adjacency rows in HBM, the four relation masks of C
//        against u and v and the c x c adjacency among C (lane j probes column C[j] of row C[i],
//        one ballot per row; 32 lanes read one 256-byte row => coalesced),
//     3. counts the ordered cliques of the tagged graph by dimension: one DFS root per lane,
//        candidate sets = three 64-bit masks (X,Y,Z), extension = AND with a shared-memory row +
//        popc; the first levels live in registers (template recursion), deeper ones in a generic
//        loop.  A flip costs two counts on ONE gather (pre: X,Y,Z; post: X,Y',Z with Y' = the
//        middle set of v->u); a double-edge move is a removal count followed by an addition count.
//   flag_count's length grows exactly as in the reference: a simplex of a dimension the complex did
//   not have before must contain the new edge, so it is met by the addition count.
//
//   FULL path (neighbourhoods with a common-neighbour list > 64, and fcm_apply where pre/post are
//   API-visible, src/bin/seo_bt_flip_only_once.rs:51-74): builds the K-vertex neighbourhood, gathers
//   its K x K adjacency and counts all its simplices before and after the edits.
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
  // delta path
  u64 *A, *deep;                       // A[64]: adjacency among the common neighbours; deep[FCM_MAXD]
  u32 *cv, *cv2;                       // cv[64], cv2[64]: common-neighbour lists (second: K of a double move)
  int* dflag;                          // [0] deep counts used, [1] highest deep dimension of an added simplex, [2] error
};
#define FAST_C 64
__host__ __device__ inline size_t warp_smem_bytes(u32 Kcap, u32 KWcap) {
  size_t full = (size_t)Kcap * KWcap * 8 + 3 * FCM_MAXD * 8 + (size_t)Kcap * 4 + FCM_MAXEDIT * 2 * 4;
  full = (full + 7) & ~(size_t)7;
  return full + FAST_C * 8 + FCM_MAXD * 8 + 2 * FAST_C * 4 + 16;
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
  size_t off = (size_t)((unsigned char*)(s.edits + FCM_MAXEDIT * 2) - base);
  off = (off + 7) & ~(size_t)7;
  s.A = (u64*)(base + off);
  s.deep = s.A + FAST_C;
  s.cv = (u32*)(s.deep + FCM_MAXD);
  s.cv2 = s.cv + FAST_C;
  s.dflag = (int*)(s.cv2 + FAST_C);
  return s;
}

// neighbourhood + pre/post local counts of the edits in s.edits (slot, add) — the two
// flagser_count calls of State::apply_transition (src/lib.rs:62-71).  0 or an FCM_E* code.
__device__ __noinline__ int transition_counts(const EnsParams& p, const Chain& c, const WarpSmem& s, int nedit,
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

// ---------------------------------------------------------------------------------------------------
// DELTA path
// ---------------------------------------------------------------------------------------------------
struct P3 { u64 x, y, z; };            // candidates by tag: before u / between u and v / after v
__device__ __forceinline__ int popc3(const P3& p) { return __popcll(p.x) + __popcll(p.y) + __popcll(p.z); }
// candidates left after appending vertex w (out-row a) with tag q: tags are non-decreasing along a simplex
__device__ __forceinline__ P3 child3(const P3& p, u64 a, int q) {
  P3 r;
  r.x = q == 0 ? (p.x & a) : 0ULL;
  r.y = q <= 1 ? (p.y & a) : 0ULL;
  r.z = p.z & a;
  return r;
}

// Common neighbours of the pair and their adjacency, gathered from the chain's rows.
//   cv[j] = C[j];  A[i] = { j : C[i] -> C[j] };  in*/out* = { j : C[j] -> u|v } / { j : u|v -> C[j] }
// myA0 / myA1 = A[lane] / A[lane + 32].  Returns false when the list does not fit (c > 64).
struct Gathered { int c; u64 inU, outU, inV, outV, myA0, myA1; };
__device__ __forceinline__ bool gather_pair(const EnsParams& p, const u64* adj, u32 pair, u32 u, u32 v,
                                            u32* cv, u64* A, int lane, Gathered& g) {
  const u64 o = __ldg(&p.nb_off[pair]);
  const int c = (int)(__ldg(&p.nb_off[pair + 1]) - o);
  g.c = c;
  if (c > FAST_C) return false;
  const bool v0 = lane < c, v1 = lane + 32 < c;
  const u32 x0 = v0 ? __ldg(&p.nb_list[o + lane]) : 0u;
  const u32 x1 = v1 ? __ldg(&p.nb_list[o + 32 + lane]) : 0u;
  cv[lane] = x0;
  cv[lane + 32] = x1;
  const u32* a32 = (const u32*)adj;                     // the same rows as 32-bit words (little endian)
  const u32 W32 = 2 * p.W;
  const u32* ru = a32 + (u64)u * W32;
  const u32* rv = a32 + (u64)v * W32;
  const u32* r0 = a32 + (u64)x0 * W32;
  const u32 w0 = x0 >> 5, s0 = x0 & 31;
  // relations of slot 0 (loads first, ballots after the row loop is in flight)
  const u32 lu0 = ru[w0], lv0 = rv[w0], li0u = r0[u >> 5], li0v = r0[v >> 5];
  u32 lu1 = 0, lv1 = 0, li1u = 0, li1v = 0, w1 = 0, s1 = 0;
  const bool two = c > 32;                              // warp-uniform
  if (two) {
    const u32* r1 = a32 + (u64)x1 * W32;
    w1 = x1 >> 5; s1 = x1 & 31;
    lu1 = ru[w1]; lv1 = rv[w1]; li1u = r1[u >> 5]; li1v = r1[v >> 5];
  }
  __syncwarp();
  g.myA0 = 0; g.myA1 = 0;
  for (int i0 = 0; i0 < c; i0 += 4) {                   // 4 rows in flight per iteration
    u32 t0[4], t1[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int i = min(i0 + k, c - 1);
      const u32* ri = a32 + (u64)cv[i] * W32;
      t0[k] = ri[w0];
      t1[k] = two ? ri[w1] : 0u;
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int i = i0 + k;
      if (i < c) {
        u64 row = __ballot_sync(FULL, v0 && ((t0[k] >> s0) & 1u));
        if (two) row |= (u64)__ballot_sync(FULL, v1 && ((t1[k] >> s1) & 1u)) << 32;
        if (lane == 0) A[i] = row;
        if (i == lane) g.myA0 = row;
        if (i == lane + 32) g.myA1 = row;
      }
    }
  }
  g.outU = __ballot_sync(FULL, v0 && ((lu0 >> s0) & 1u));
  g.outV = __ballot_sync(FULL, v0 && ((lv0 >> s0) & 1u));
  g.inU = __ballot_sync(FULL, v0 && ((li0u >> (u & 31)) & 1u));
  g.inV = __ballot_sync(FULL, v0 && ((li0v >> (v & 31)) & 1u));
  if (two) {
    g.outU |= (u64)__ballot_sync(FULL, v1 && ((lu1 >> s1) & 1u)) << 32;
    g.outV |= (u64)__ballot_sync(FULL, v1 && ((lv1 >> s1) & 1u)) << 32;
    g.inU |= (u64)__ballot_sync(FULL, v1 && ((li1u >> (u & 31)) & 1u)) << 32;
    g.inV |= (u64)__ballot_sync(FULL, v1 && ((li1v >> (v & 31)) & 1u)) << 32;
  }
  __syncwarp();
  return true;
}

// tuples of >= DELTA_SD+1 extra vertices: generic DFS, counts into the warp's shared deep[] (by
// flag_count index).  p0 = candidates of a tuple of `size0` extra vertices (non-empty).
#define DELTA_SD 3
__device__ __noinline__ void deep_dfs(P3 p0, int size0, const u64* A, int sign, u64* deep, int* dflag) {
  P3 P[FCM_MAXD];
  u64 R[FCM_MAXD];
  int Q[FCM_MAXD];
  dflag[0] = 1;
  int l = 0;
  P[0] = p0;
  for (;;) {
    // entering level l: the tuple has size0 + l extra vertices; its children are simplices of
    // size0 + l + 1 extra vertices + u + v  =>  flag_count index size0 + l + 2
    const int d = size0 + l + 2;
    if (d >= FCM_MAXD) { dflag[2] = FCM_EDIM; return; }
    atomicAdd((unsigned long long*)&deep[d], (unsigned long long)(long long)(sign * popc3(P[l])));
    if (sign > 0) atomicMax(&dflag[1], d);
    Q[l] = 0;
    R[l] = P[l].x;
    for (;;) {                                          // next child of level l, or pop
      while (R[l] == 0 && Q[l] < 2) { Q[l]++; R[l] = Q[l] == 1 ? P[l].y : P[l].z; }
      if (R[l] == 0) {
        if (l == 0) return;
        --l;
        continue;
      }
      const int w = __ffsll((long long)R[l]) - 1;
      R[l] &= R[l] - 1;
      const P3 ch = child3(P[l], A[w], Q[l]);
      if (ch.x | ch.y | ch.z) { P[++l] = ch; break; }
    }
  }
}

// p = candidates of a tuple of L extra vertices (non-empty); cnt[k] is the number of simplices with
// k + 2 extra vertices (flag_count index k + 3), k = 0..DELTA_SD-1, this lane's share
template <int L>
__device__ __forceinline__ void dfs_static(const P3& p, const u64* A, int sign, long long (&cnt)[DELTA_SD],
                                           u64* deep, int* dflag) {
  cnt[L - 1] += sign * popc3(p);
#pragma unroll 1
  for (int q = 0; q < 3; q++) {
    u64 R = q == 0 ? p.x : q == 1 ? p.y : p.z;
    while (R) {
      const int w = __ffsll((long long)R) - 1;
      R &= R - 1;
      const P3 ch = child3(p, A[w], q);
      if (ch.x | ch.y | ch.z) {
        if constexpr (L < DELTA_SD) dfs_static<L + 1>(ch, A, sign, cnt, deep, dflag);
        else deep_dfs(ch, L + 1, A, sign, deep, dflag);
      }
    }
  }
}

// Number of simplices through the edge u->v by dimension, times `sign`, added to this lane's cnt[]
// (and the warp's deep[]): X/Y/Z = the three tag masks over the gathered common neighbours.
__device__ __forceinline__ void count_through_edge(const P3& p0, const Gathered& g, const u64* A, int lane, int sign,
                                                   long long (&cnt)[DELTA_SD], u64* deep, int* dflag) {
#pragma unroll 1
  for (int slot = 0; slot < 2; slot++) {
    const int w = lane + 32 * slot;
    if (w >= g.c) break;
    const u64 a = slot ? g.myA1 : g.myA0;
    const u64 bit = 1ULL << w;
#pragma unroll 1
    for (int q = 0; q < 3; q++) {
      const u64 m = q == 0 ? p0.x : q == 1 ? p0.y : p0.z;
      if (!(m & bit)) continue;
      const P3 ch = child3(p0, a, q);
      if (ch.x | ch.y | ch.z) dfs_static<1>(ch, A, sign, cnt, deep, dflag);
    }
  }
}

// warp sum of one signed 64-bit value per lane
__device__ __forceinline__ long long warp_sum_i64(long long v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

// Adds, to the flag_count entry this lane owns (index = lane), the warp totals of cnt[] (indices 3..)
// and of the uniform triangle term tri (index 2).
__device__ __forceinline__ void fold_counts(long long (&cnt)[DELTA_SD], long long tri, int lane, long long& delta) {
  bool big = false;
#pragma unroll
  for (int k = 0; k < DELTA_SD; k++) big |= ((u64)(cnt[k] + (1LL << 25)) >> 26) != 0;
  if (!__any_sync(FULL, big)) {
#pragma unroll
    for (int k = 0; k < DELTA_SD; k++) {
      const int t = __reduce_add_sync(FULL, (int)cnt[k]);
      if (lane == k + 3) delta += t;
    }
  } else {
#pragma unroll
    for (int k = 0; k < DELTA_SD; k++) {
      const long long t = warp_sum_i64(cnt[k]);
      if (lane == k + 3) delta += t;
    }
  }
  if (lane == 2) delta += tri;
}

__device__ __forceinline__ bool in_sorted(const u32* s, int n, u32 key) {
  int lo = 0, hi = n;
  while (lo < hi) { int mid = (lo + hi) >> 1; if (s[mid] < key) lo = mid + 1; else hi = mid; }
  return lo < n && s[lo] == key;
}
__device__ __forceinline__ bool in_list_warp(const u32* s, int n, u32 key, int lane) {
  return __any_sync(FULL, (lane < n && s[lane] == key) || (lane + 32 < n && s[lane + 32] == key));
}
// |C1 u C2 u {a,b,x,y}|: State::edgeset_neighborhood's size for a double-edge move (src/lib.rs:99-111)
__device__ __forceinline__ int union_size(const u32* l1, int c1, const u32* l2, int c2, u32 a, u32 b, u32 x, u32 y, int lane) {
  int both = __popc(__ballot_sync(FULL, lane < c2 && in_sorted(l1, c1, l2[lane])));
  if (c2 > 32) both += __popc(__ballot_sync(FULL, lane + 32 < c2 && in_sorted(l1, c1, l2[lane + 32])));
  int K = c1 + c2 - both;
  // endpoints: a, b are never in l1 (their own common neighbours), x, y never in l2
  K += !in_list_warp(l2, c2, a, lane);
  K += !in_list_warp(l2, c2, b, lane);
  if (x != a && x != b) K += !in_list_warp(l1, c1, x, lane);
  if (y != a && y != b) K += !in_list_warp(l1, c1, y, lane);
  return K;
}

// Bounds::check (src/lib.rs:157-160) on the flag_count held one entry per lane
__device__ __forceinline__ bool bounds_ok(const EnsParams& p, u64 f, int lane) {
  u64 lo = lane < p.nmin ? p.bmin[lane] : 0, hi = lane < p.nmax ? p.bmax[lane] : 0;
  bool bad = lo > f || f > hi;
  if (lane == 0 && p.nmin > FCM_MAXD && p.bmin[FCM_MAXD] > 0) bad = true;   // entry 32: count is 0 there
  return !__any_sync(FULL, bad);
}

__device__ __forceinline__ void set_adj_bit(const EnsParams& p, u64* adj, u32 from, u32 to, bool add, int lane) {
  if (lane == 0) {
    u64* w = &adj[(u64)from * p.W + (to >> 6)];
    const u64 m = 1ULL << (to & 63);
    *w = add ? (*w | m) : (*w & ~m);
  }
  __syncwarp();
}

__global__ void __launch_bounds__(128, 7)
chain_step_kernel(const EnsParams p, u64 nsteps, fcm_trace_rec* __restrict__ trace, u32 trace_chain,
                  u32 smem_per_warp, int force_full) {
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
  s.deep[lane] = 0;
  if (lane < 3) s.dflag[lane] = 0;
  __syncwarp();

  for (u64 step = 0; step < nsteps; step++) {
    // ---- Transition::random_move: WeightedIndex<f64>::sample (src/lib.rs:210) ----
    u64 r = g.next();
    double v = __longlong_as_double((long long)((r >> 12) | 0x3ff0000000000000ULL)) - 1.0;
    double chosen = __dadd_rn(__dmul_rn(v, p.scale), 0.0);
    int kind = (p.cum[0] <= chosen) + (p.cum[1] <= chosen) + (p.cum[2] <= chosen);
    int nedit = 0;
    u32 ta = 0, tb = 0, tdm0 = 0, tdm1 = 0;
    u32 slot = 0, delme = 0;
    if (kind == 0) {                                   // single_edge_flip  src/lib.rs:292-299
      if (c.E) {
        slot = c.items[gen_range(g, c.E)];
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
        for (;;) {
          slot = c.items[gen_range(g, c.E)];
          if (c.pos[slot ^ 1] == FCM_NONE) break;
        }
        bool first = g.next() < 0x8000000000000000ULL;  // gen_bool(0.5)
        delme = 2 * dpair + (first ? 1u : 0u);          // first: [min,max] = small -> big
        uint2 ft = slot_edge(p, slot), dm = slot_edge(p, delme);
        ta = ft.x; tb = ft.y; tdm0 = dm.x; tdm1 = dm.y;
        if (lane == 0) { s.edits[0] = slot ^ 1; s.edits[1] = 1; s.edits[2] = delme; s.edits[3] = 0; }
        nedit = 2;
      }
    } else { err = FCM_EUNSUPPORTED; break; }
    __syncwarp();

    // ---- State::apply_transition  src/lib.rs:61-79 ----
    int K = 0, prelen = 0;
    u64 fnew = fc;
    bool bit_cleared = false;                          // double move, delta path: delme's bit is already 0
    if (nedit) {
      bool done = false;
      if (!force_full) {
        long long cnt[DELTA_SD];
#pragma unroll
        for (int k = 0; k < DELTA_SD; k++) cnt[k] = 0;
        long long tri = 0, delta = 0;
        int top = 0;                                   // highest flag_count index holding an added simplex
        Gathered ga;
        if (kind == 0) {
          // a flip ta->tb => tb->ta: one gather, the counts through ta->tb leave, those through tb->ta come
          if (gather_pair(p, c.adj, slot >> 1, ta, tb, s.cv, s.A, lane, ga)) {
            const u64 X = ga.inU & ga.inV, Z = ga.outU & ga.outV;
            const P3 post{X, ga.inU & ga.outV, Z}, pre{X, ga.outU & ga.inV, Z};
            count_through_edge(post, ga, s.A, lane, +1, cnt, s.deep, s.dflag);
#pragma unroll
            for (int k = 0; k < DELTA_SD; k++) if (__any_sync(FULL, cnt[k] != 0)) top = k + 3;
            if (post.x | post.y | post.z) top = max(top, 2);
            count_through_edge(pre, ga, s.A, lane, -1, cnt, s.deep, s.dflag);
            tri = (long long)__popcll(post.y) - (long long)__popcll(pre.y);
            K = ga.c + 2;
            top = max(top, 1);
            done = true;
          }
        } else {
          // double move: delme = tdm0->tdm1 leaves (counted in the graph before any edit), then
          // tb->ta comes (counted in the final graph: delme's bit is cleared in between)
          const u64 o2 = __ldg(&p.nb_off[delme >> 1]);
          const int c2 = (int)(__ldg(&p.nb_off[(delme >> 1) + 1]) - o2);
          const u64 o1 = __ldg(&p.nb_off[slot >> 1]);
          const int c1 = (int)(__ldg(&p.nb_off[(slot >> 1) + 1]) - o1);
          if (c1 <= FAST_C && c2 <= FAST_C) {
            gather_pair(p, c.adj, delme >> 1, tdm0, tdm1, s.cv2, s.A, lane, ga);
            const P3 gone{ga.inU & ga.inV, ga.outU & ga.inV, ga.outU & ga.outV};
            count_through_edge(gone, ga, s.A, lane, -1, cnt, s.deep, s.dflag);
            tri = -(long long)popc3(gone);
            set_adj_bit(p, c.adj, tdm0, tdm1, false, lane);
            bit_cleared = true;
            long long cntp[DELTA_SD];
#pragma unroll
            for (int k = 0; k < DELTA_SD; k++) cntp[k] = 0;
            gather_pair(p, c.adj, slot >> 1, tb, ta, s.cv, s.A, lane, ga);
            const P3 come{ga.inU & ga.inV, ga.outU & ga.inV, ga.outU & ga.outV};
            count_through_edge(come, ga, s.A, lane, +1, cntp, s.deep, s.dflag);
#pragma unroll
            for (int k = 0; k < DELTA_SD; k++) { if (__any_sync(FULL, cntp[k] != 0)) top = k + 3; cnt[k] += cntp[k]; }
            if (come.x | come.y | come.z) top = max(top, 2);
            tri += (long long)popc3(come);
            K = union_size(s.cv, c1, s.cv2, c2, ta, tb, tdm0, tdm1, lane);
            top = max(top, 1);
            done = true;
          }
        }
        if (done) {
          fold_counts(cnt, tri, lane, delta);
          __syncwarp();
          if (__any_sync(FULL, s.dflag[0] != 0)) {     // some lane went below the static levels
            if (s.dflag[2]) { err = s.dflag[2]; break; }
            delta += (long long)s.deep[lane];
            top = max(top, s.dflag[1]);
            __syncwarp();
            s.deep[lane] = 0;
            if (lane < 3) s.dflag[lane] = 0;
            __syncwarp();
          }
          fnew = fc + (u64)delta;
          if (__any_sync(FULL, (long long)fnew < 0)) { err = FCM_EASSERT; break; }
          prelen = 0;                                  // the reverted length never exceeds len (see header)
          len = max(len, top + 1);
        }
      }
      if (!done) {
        int postlen = 0;
        int rc = transition_counts(p, c, s, nedit, lane, K, prelen, postlen);
        if (rc) { err = rc; break; }
        u64 pr = lane < min(prelen, len) ? s.pre[lane] : 0;
        if (__any_sync(FULL, fc < pr)) { err = FCM_EASSERT; break; }
        len = max(len, postlen);
        fnew = fc - pr + (lane < postlen ? s.post[lane] : 0);
      }
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
      if (bit_cleared) set_adj_bit(p, c.adj, tdm0, tdm1, true, lane);
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
                                                              (u32)e->smem_per_warp, e->force_full);
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
