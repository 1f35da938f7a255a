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
#include <cstdlib>
#include <mutex>

#ifndef FCM_CHAIN_MINBLOCKS
#define FCM_CHAIN_MINBLOCKS 7              // 7 blocks x 4 warps = 28 chains per SM: 4096 chains fill 148 SMs in one wave (72 registers per thread)
#endif
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

// THE EDGE-SAMPLER POLICY SEAM ("isolated behind one small policy").  Graph::sample_edge / sample_double_edge of
// the external flag-complex crate pick an element of the insertion-ordered container from one gen_range(0..len) draw r.
// Policy 0 (the documented assumption, oracle/fcm_oracle.cpp Graph::sample_edge): element r.  Policy 1: element len-1-r -
// not a guess about the crate, it exists so that tests can show GPU and oracle change together when the policy changes
// (tests/test_gpu_chain.py::test_sampler_policy_seam); the crate's real rule, once known, goes here and in the oracle.
__device__ __forceinline__ u32 sample_at(const EnsParams& p, u32 r, u32 len) { return p.sample_policy ? len - 1u - r : r; }

struct Chain {                        // pointers of one chain + live scalars
  u64* adj;
  u32 *items, *pos, *dbl, *dblpos, *ori;
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
    u32* o = &c.ori[slot >> 5];
    const u32 om = 1u << (slot & 31);
    *o = add ? (*o | om) : (*o & ~om);
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
    const u32 o = p.nb_off[pair];
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
      bool bit = have && ((__ldcg(&row[gw]) >> gs) & 1ULL);
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

// Per-warp shared memory.  The delta path's arrays sit at fixed offsets (the hot loop keeps one base
// register); behind them one region serves either as the cp.async staging tile or as the full
// path's arrays (a proposal takes one path or the other).
#define FAST_C 64                      // longest common-neighbour list the delta path takes
#define NARROW_C 32                    // staged gather: one column per lane
#define TILE_WORDS 1504                // >= hot_tile_words(32) = 1152; 7 blocks of 4 warps x (SM_TILE + 5.9 KB) + 1 KB fit one SM
#define SM_A 0                         // u64[FAST_C + 4]  adjacency among the common neighbours
#define SM_DEEP 544                    // u64[FCM_MAXD]    counts below the register levels, by flag_count index
#define SM_ROWOFF 800                  // u32[64]          row offsets of the two lists
#define SM_CV 1056                     // u32[64]          common neighbours of the coming pair
#define SM_CV2 1312                    // u32[64]          ... of the going pair (double move)
#define SM_FLAG 1568                   // int[12]          see FL_*
#define SM_AT 1616                     // u32[32]          hot path: rows of the tagged graph
#define SM_SRC 1744                    // u32[32]          hot path: source vertex of each extra tagged copy
#define SM_PARK 1872                   // u32[32]          the sampled proposal (update stage, cold_step), see PK_*
#define SM_RNG 2000                    // u64[4]           xoshiro state between the sampling stages
#define SM_FC 2032                     // u64[32]          flag_count, one entry per lane
#define SM_TILE 2288
enum { FL_DEEP = 0, FL_TOP = 1, FL_ERR = 2, FL_RC = 3, FL_K = 4, FL_LEN = 5, FL_OVF = 6, FL_E = 8, FL_ND = 9, FL_ACC = 10, FL_FCLEN = 11 };
enum { PK_KIND = 0, PK_SLOT, PK_DELME, PK_IDX, PK_DIDX, PK_DPAIR, PK_LAST, PK_QPOS, PK_TA, PK_TB, PK_TDM0, PK_TDM1,
       PK_CA, PK_CD, PK_OA, PK_OD, PK_ORIA, PK_ORID };
#define CQ_SCRATCH 160                  // u32 words behind the header: clique_swap's rank / row tables (128) + perm_a while it is drawn (32)
#define CQ_BYTES (3 * FCM_MAXCQ * 4 + 64 + CQ_SCRATCH * 4)
// staging block of the index hot path (idx_proposal), laid over the full path's arrays behind SM_TILE
#ifndef FCM_CQ_PREFETCH
#define FCM_CQ_PREFETCH 0              // clique moves: L2 prefetch of the next edit's orientation words - measured SLOWER (53.5 vs 56.8 M
                                       // proposals/s, default mix on C3): the path is bound by instruction fetch / issue, not by the round trips
#endif
#ifndef FCM_LA
#define FCM_LA 0                       // look-ahead in the 7-blocks-per-SM instance: measured a loss there (28 warps per SM hide the latency
                                       // themselves: 502.6 vs 511 M), off; the 4-blocks-per-SM instance of small ensembles has it on (LA below)
#endif
#ifndef FCM_LA2
#define FCM_LA2 1
#endif
#define ST_STAGE 0                     // u32[2][6][32]  per pair (A, D): staged orientation words wu, wv, w0, w1, then the lanes' packed entries
#define ST_ROWS 6
#define ST_PP 1536                     // u32[16]        per pair: list offset, c, k4 offset, m | uBig << 16; [8] [9] root sets of the tagged graph
#define ST_LA 1600                     // u32[LA_WORDS]  the look-ahead of the next proposal
#define ST_RNGLA 1664                  // u64[4]         RNG state behind the look-ahead draws
#define ST_PINFO 1696                  // uint4[4]       records of the next proposal's pairs: A (2), D (2)
#define ST_BYTES 1760
struct WarpSmem {
  u64 *M, *fc, *pre, *post;
  u32 *verts, *edits, *cq;
  u64 *A, *deep;
  u32 *cv, *cv2, *rowoff;
  int* dflag;
  u32* tile;
};
// tile = false: the hot path reads the clique index, no staging tile is needed
// long_edits: room for FCM_MAXEDIT edits (clique moves, fcm_apply); the step kernel of an ensemble without clique
// moves only ever holds the two edits of a simple move
__host__ __device__ inline size_t warp_smem_bytes(u32 Kcap, u32 KWcap, bool tile, bool cq, bool long_edits = true) {
  size_t full = (size_t)Kcap * KWcap * 8 + 3 * FCM_MAXD * 8 + (size_t)(Kcap + (Kcap & 1)) * 4 + (long_edits || cq ? FCM_MAXEDIT * 2 * 4 : 32) + (cq ? CQ_BYTES : 0);
  if (full < ST_BYTES) full = ST_BYTES;                 // the index hot path's staging block overlays these arrays
  if (tile && full < TILE_WORDS * 4) full = TILE_WORDS * 4;
  return SM_TILE + full;
}
size_t fcm_chain_smem_per_warp(u32 Kcap, u32 KWcap, bool cliques) { return (warp_smem_bytes(Kcap, KWcap, true, cliques) + 15) & ~(size_t)15; }
static size_t chain_smem_per_warp_mode(u32 Kcap, u32 KWcap, bool tile, bool cq) { return (warp_smem_bytes(Kcap, KWcap, tile, cq, false) + 15) & ~(size_t)15; }

__device__ __forceinline__ WarpSmem carve(unsigned char* base, u32 Kcap, u32 KWcap) {
  WarpSmem s;
  s.A = (u64*)(base + SM_A);
  s.deep = (u64*)(base + SM_DEEP);
  s.rowoff = (u32*)(base + SM_ROWOFF);
  s.cv = (u32*)(base + SM_CV);
  s.cv2 = (u32*)(base + SM_CV2);
  s.dflag = (int*)(base + SM_FLAG);
  s.tile = (u32*)(base + SM_TILE);
  s.M = (u64*)(base + SM_TILE);
  s.fc = s.M + (size_t)Kcap * KWcap;
  s.pre = s.fc + FCM_MAXD;
  s.post = s.pre + FCM_MAXD;
  s.verts = (u32*)(s.post + FCM_MAXD);
  s.edits = s.verts + Kcap + (Kcap & 1);
  s.cq = s.edits + 2 * FCM_MAXEDIT;                     // clique-move scratch (only sized when the ensemble has clique moves)
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
#define DELTA_SD 3                     // tuple sizes 1..3 are walked in registers, deeper ones generically
__device__ __forceinline__ int popc_m(u32 x) { return __popc(x); }
__device__ __forceinline__ int popc_m(u64 x) { return __popcll(x); }
__device__ __forceinline__ int low_m(u32 x) { return __ffs((int)x) - 1; }
__device__ __forceinline__ int low_m(u64 x) { return __ffsll((long long)x) - 1; }

// Common neighbours C of the pair and their adjacency, gathered from the chain's rows (32-bit words,
// L2-cached loads: a row is touched once per proposal and is rewritten by RED atomics on accept).
//   list[j] = C[j];  A[i] = { j : C[i] -> C[j] };  in*/out* = { j : C[j] -> u|v } / { j : u|v -> C[j] }
// The caller has checked c <= FAST_C.
struct Gathered { int c; u64 inU, outU, inV, outV; };
__device__ __forceinline__ void gather_pair(const EnsParams& p, const u32* a32, u32 o, int c, u32 u, u32 v,
                                            u32* list, u32* rowoff, u64* A, int lane, Gathered& g) {
  const u32 W32 = 2 * p.W;
  const int ncs = (c + 31) >> 5;                        // column slots: lanes hold C[lane], C[lane + 32]
  const int cpad = (c + 3) & ~3;                        // rows are walked four at a time
  g.c = c; g.inU = g.outU = g.inV = g.outV = 0;
  for (int cs = 0; cs < ncs; cs++) {
    const int j = cs * 32 + lane;
    const u32 x = j < c ? __ldg(&p.nb_list[o + j]) : 0u;
    list[j] = x;
    rowoff[j] = x * W32;                                // word offset of row x (n*W*2 < 2^32 is checked at create)
  }
  __syncwarp();
  const u32* ru = a32 + u * W32;
  const u32* rv = a32 + v * W32;
#pragma unroll 1
  for (int cs = 0; cs < ncs; cs++) {
    const int j = cs * 32 + lane;
    const u32 x = list[j];
    const u32 w = x >> 5;
    const u32 m = j < c ? 1u << (x & 31) : 0u;          // 0 for idle lanes: never votes
    const u32* rx = a32 + rowoff[j];
    const u32 lu = __ldcg(&ru[w]), lv = __ldcg(&rv[w]);
    const u32 xu = __ldcg(&rx[u >> 5]), xv = __ldcg(&rx[v >> 5]);
    u32* A32 = (u32*)A + cs;
#pragma unroll 1
    for (int i0 = 0; i0 < cpad; i0 += 4) {
      const uint4 ro = *(const uint4*)&rowoff[i0];      // rows past c read offset 0 (row 0): never used
      const u32 t0 = __ldcg(&a32[ro.x + w]), t1 = __ldcg(&a32[ro.y + w]);
      const u32 t2 = __ldcg(&a32[ro.z + w]), t3 = __ldcg(&a32[ro.w + w]);
      const u32 b0 = __ballot_sync(FULL, t0 & m), b1 = __ballot_sync(FULL, t1 & m);
      const u32 b2 = __ballot_sync(FULL, t2 & m), b3 = __ballot_sync(FULL, t3 & m);
      if (lane == 0) {
        if (cs == 0) { A[i0] = b0; A[i0 + 1] = b1; A[i0 + 2] = b2; A[i0 + 3] = b3; }
        else { A32[2 * i0] = b0; A32[2 * i0 + 2] = b1; A32[2 * i0 + 4] = b2; A32[2 * i0 + 6] = b3; }
      }
    }
    const int sh = 32 * cs;
    g.outU |= (u64)__ballot_sync(FULL, lu & m) << sh;
    g.outV |= (u64)__ballot_sync(FULL, lv & m) << sh;
    g.inU |= (u64)__ballot_sync(FULL, m && ((xu >> (u & 31)) & 1u)) << sh;
    g.inV |= (u64)__ballot_sync(FULL, m && ((xv >> (v & 31)) & 1u)) << sh;
  }
  __syncwarp();
}

// local index of vertex x in list[0..c) or -1 (warp-uniform)
__device__ __forceinline__ int list_index(const u32* list, int c, u32 x, int lane) {
  u32 b = __ballot_sync(FULL, lane < c && list[lane] == x);
  if (b) return __ffs(b) - 1;
  if (c > 32) { b = __ballot_sync(FULL, lane + 32 < c && list[lane + 32] == x); if (b) return 32 + __ffs(b) - 1; }
  return -1;
}

// tuples of more than DELTA_SD extra vertices: generic DFS, counts into the warp's shared deep[] (by
// flag_count index).  (px,py,pz) = candidates of a tuple of `size0` extra vertices (non-empty).
__device__ __noinline__ void deep_dfs(u64 px, u64 py, u64 pz, int size0, const u64* A, int sign, u64* deep, int* dflag) {
  u64 PX[FCM_MAXD], PY[FCM_MAXD], PZ[FCM_MAXD], R[FCM_MAXD];
  int Q[FCM_MAXD];
  dflag[0] = 1;
  int l = 0;
  PX[0] = px; PY[0] = py; PZ[0] = pz;
  for (;;) {
    // entering level l: the tuple has size0 + l extra vertices; each candidate closes a simplex of
    // size0 + l + 1 extra vertices + u + v  =>  flag_count index size0 + l + 2
    const int d = size0 + l + 2;
    if (d >= FCM_MAXD) { dflag[2] = FCM_EDIM; return; }
    const int pc = __popcll(PX[l]) + __popcll(PY[l]) + __popcll(PZ[l]);
    atomicAdd((unsigned long long*)&deep[d], (unsigned long long)(long long)(sign * pc));
    if (sign > 0) atomicMax(&dflag[1], d);
    Q[l] = 0;
    R[l] = PX[l];
    for (;;) {                                          // next child of level l, or pop
      while (R[l] == 0 && Q[l] < 2) { Q[l]++; R[l] = Q[l] == 1 ? PY[l] : PZ[l]; }
      if (R[l] == 0) {
        if (l == 0) return;
        --l;
        continue;
      }
      const int w = __ffsll((long long)R[l]) - 1;
      R[l] &= R[l] - 1;
      const u64 a = A[w];
      const u64 cx = Q[l] == 0 ? (PX[l] & a) : 0ULL, cy = Q[l] <= 1 ? (PY[l] & a) : 0ULL, cz = PZ[l] & a;
      if (cx | cy | cz) { ++l; PX[l] = cx; PY[l] = cy; PZ[l] = cz; break; }
    }
  }
}

// (px,py,pz) = candidates (by tag) of a tuple of L extra vertices, non-empty.  cnt[k] = this lane's
// signed number of simplices with k + 2 extra vertices (flag_count index k + 3).  |cnt| cannot
// overflow: at most 64^3 tuples at the register levels, each adding <= 192.
template <int L, typename M>
__device__ __forceinline__ void walk(M px, M py, M pz, const u64* A, int sign, int (&cnt)[DELTA_SD], u64* deep, int* dflag) {
  cnt[L - 1] += sign * (popc_m(px) + popc_m(py) + popc_m(pz));
#pragma unroll 1
  for (int q = 0; q < 3; q++) {
    M R = q == 0 ? px : q == 1 ? py : pz;
    const M kx = q == 0 ? px : (M)0, ky = q <= 1 ? py : (M)0;   // tags never decrease along a simplex
    while (R) {
      const int w = low_m(R);
      R &= R - 1;
      const M a = (M)A[w];
      const M cx = kx & a, cy = ky & a, cz = pz & a;
      if (cx | cy | cz) {
        if constexpr (L < DELTA_SD) walk<L + 1, M>(cx, cy, cz, A, sign, cnt, deep, dflag);
        else deep_dfs(cx, cy, cz, L + 1, A, sign, deep, dflag);
      }
    }
  }
}

// Simplices through the edge u->v with >= 2 extra vertices, by dimension, times `sign`: one root
// (extra vertex) per lane; (X,Y,Z) = who may stand before u / between u and v / after v.
template <typename M>
__device__ __forceinline__ void count_through_edge(M X, M Y, M Z, int c, const u64* A, int lane, int sign,
                                                   int (&cnt)[DELTA_SD], u64* deep, int* dflag) {
#pragma unroll 1
  for (int w = lane; w < c; w += 32) {
    const M a = (M)A[w];
    const M bit = (M)1 << w;
    const M zx = Z & a, yx = Y & a, xx = X & a;
#pragma unroll 1
    for (int q = 0; q < 3; q++) {                       // the tags this root may carry
      const M m = q == 0 ? X : q == 1 ? Y : Z;
      if (!(m & bit)) continue;
      const M cx = q == 0 ? xx : (M)0, cy = q <= 1 ? yx : (M)0;
      if (cx | cy | zx) walk<1, M>(cx, cy, zx, A, sign, cnt, deep, dflag);
    }
  }
}

__device__ __forceinline__ bool in_sorted(const u32* s, int n, u32 key) {
  int lo = 0, hi = n;
  while (lo < hi) { int mid = (lo + hi) >> 1; if (s[mid] < key) lo = mid + 1; else hi = mid; }
  return lo < n && s[lo] == key;
}
// |C1 u C2 u {a,b,x,y}|: State::edgeset_neighborhood's size for a double-edge move (src/lib.rs:99-111)
__device__ __forceinline__ int union_size(const u32* l1, int c1, const u32* l2, int c2, u32 a, u32 b, u32 x, u32 y, int lane) {
  int both = __popc(__ballot_sync(FULL, lane < c2 && in_sorted(l1, c1, l2[lane])));
  if (c2 > 32) both += __popc(__ballot_sync(FULL, lane + 32 < c2 && in_sorted(l1, c1, l2[lane + 32])));
  int K = c1 + c2 - both;
  // endpoints: a, b are never in l1 (their own common neighbours), x, y never in l2
  K += list_index(l2, c2, a, lane) < 0;
  K += list_index(l2, c2, b, lane) < 0;
  if (x != a && x != b) K += list_index(l1, c1, x, lane) < 0;
  if (y != a && y != b) K += list_index(l1, c1, y, lane) < 0;
  return K;
}

// Bounds::check (src/lib.rs:157-160) on the flag_count held one entry per lane
__device__ __forceinline__ bool bounds_ok(const EnsParams& p, u64 f, int lane) {
  u64 lo = lane < p.nmin ? p.bmin[lane] : 0, hi = lane < p.nmax ? p.bmax[lane] : 0;
  bool bad = lo > f || f > hi;
  if (lane == 0 && p.nmin > FCM_MAXD && p.bmin[FCM_MAXD] > 0) bad = true;   // entry 32: count is 0 there
  return !__any_sync(FULL, bad);
}

// adjacency bit from->to of the chain, no return value needed (RED at L2; rows are read with ld.cg)
__device__ __forceinline__ void red_adj_bit(const EnsParams& p, u32* a32, u32 from, u32 to, bool add) {
  u32* w = a32 + (u64)from * (2 * p.W) + (to >> 5);
  const u32 m = 1u << (to & 31);
  if (add) atomicOr(w, m); else atomicAnd(w, ~m);
}


// L2 eviction priorities (build switches, measurements in profiles/README.md): FCM_IDX_KEEP = the static clique index
// (pinfo / tri / k4, shared by every chain) is loaded with an evict_last policy; FCM_STATE_STREAM = the per-chain
// orientation words are fetched with evict_first (a chain's words are not touched again before 4095 other chains'
// traffic has passed through L2).
#ifndef FCM_IDX_KEEP
#define FCM_IDX_KEEP 0
#endif
#ifndef FCM_STATE_STREAM
#define FCM_STATE_STREAM 0
#endif
__device__ __forceinline__ u64 l2_policy_keep() { u64 pol; asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol)); return pol; }
__device__ __forceinline__ u64 l2_policy_stream() { u64 pol; asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol)); return pol; }
__device__ __forceinline__ u32 ldg_idx(const u32* g) {
#if FCM_IDX_KEEP
  u32 v; asm("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(g), "l"(l2_policy_keep())); return v;
#else
  return __ldg(g);
#endif
}
__device__ __forceinline__ uint2 ldg_idx(const uint2* g) {
#if FCM_IDX_KEEP
  uint2 v; asm("ld.global.nc.L2::cache_hint.v2.u32 {%0, %1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(g), "l"(l2_policy_keep())); return v;
#else
  return __ldg(g);
#endif
}
__device__ __forceinline__ uint4 ldg_idx(const uint4* g) {
#if FCM_IDX_KEEP
  uint4 v; asm("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(g), "l"(l2_policy_keep())); return v;
#else
  return __ldg(g);
#endif
}
__device__ __forceinline__ void cp_async4(u32 saddr, const void* g) {
#if FCM_STATE_STREAM
  asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 4, %2;" ::"r"(saddr), "l"(g), "l"(l2_policy_stream()) : "memory");
#else
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(saddr), "l"(g) : "memory");
#endif
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// the double move's count of the coming edge u->v sees the graph without delme = d0->d1: take that bit
// out of the gathered view (list/A/ga are the coming pair's)
__device__ __forceinline__ void drop_edge_from_view(const u32* list, int c, u32 u, u32 v, u32 d0, u32 d1, u64* A,
                                                    int lane, Gathered& ga) {
  const int ip = d0 == u ? -2 : d0 == v ? -3 : list_index(list, c, d0, lane);
  const int iq = d1 == u ? -2 : d1 == v ? -3 : list_index(list, c, d1, lane);
  if (ip == -1 || iq == -1) return;
  if (ip >= 0 && iq >= 0) { if (lane == 0) A[ip] &= ~(1ULL << iq); __syncwarp(); }
  else if (ip >= 0) { if (iq == -2) ga.inU &= ~(1ULL << ip); else ga.inV &= ~(1ULL << ip); }
  else if (iq >= 0) { if (ip == -2) ga.outU &= ~(1ULL << iq); else ga.outV &= ~(1ULL << iq); }
}

// ---- hot path (lists of <= 32 common neighbours, tagged graph of <= 32 vertices) ----------------------
// hot_issue: every word the pair needs is requested with cp.async, nothing is held in registers:
//   tile rows 0..cpad-1 : row C[i], the word holding column C[lane]
//   tile rows cpad..+3  : row u and row v at that word; row C[lane] at the words of u and of v
__device__ __forceinline__ int hot_tile_words(int c) { return (((c + 3) & ~3) + 4) * c; }
__device__ __forceinline__ void hot_issue(const u32* cb, u32 W32, u32 x, int c, u32 u, u32 v, u32* list, u32* rowoff,
                                          u32* tile, int lane) {
  list[lane] = x;
  rowoff[lane] = x * W32;                               // lanes past c hold row 0: copied for the padding rows only
  __syncwarp();
  if (lane < c) {
    u32 dst = (u32)__cvta_generic_to_shared(tile + lane);
    const u32* src = cb + (x >> 5);
    const u32 pitch = 4u * (u32)c;
    const int cpad = (c + 3) & ~3;
#pragma unroll 1
    for (int i0 = 0; i0 < cpad; i0 += 4) {
      const uint4 ro = *(const uint4*)&rowoff[i0];
      cp_async4(dst, src + ro.x); dst += pitch;
      cp_async4(dst, src + ro.y); dst += pitch;
      cp_async4(dst, src + ro.z); dst += pitch;
      cp_async4(dst, src + ro.w); dst += pitch;
    }
    const u32* rx = cb + x * W32;
    cp_async4(dst, src + u * W32); dst += pitch;
    cp_async4(dst, src + v * W32); dst += pitch;
    cp_async4(dst, rx + (u >> 5)); dst += pitch;
    cp_async4(dst, rx + (v >> 5));
  }
  cp_async_commit();
}
// hot_finish: tile -> A32[i] = { j : C[i] -> C[j] } and the four relation masks
struct G32 { u32 inU, outU, inV, outV; };
__device__ __forceinline__ G32 hot_finish(const u32* tile, int c, const u32* list, u32 u, u32 v, u32* A32, int lane) {
  const u32 x = list[lane];
  const u32 m = lane < c ? 1u << (x & 31) : 0u;         // idle lanes never vote
  const u32* col = tile + min(lane, max(c - 1, 0));
  const int cpad = (c + 3) & ~3;
#pragma unroll 1
  for (int i = 0; i < cpad; i += 4) {                   // the padding rows give rows that nobody reads
    const u32 t0 = col[i * c], t1 = col[(i + 1) * c], t2 = col[(i + 2) * c], t3 = col[(i + 3) * c];
    uint4 b;
    b.x = __ballot_sync(FULL, t0 & m); b.y = __ballot_sync(FULL, t1 & m);
    b.z = __ballot_sync(FULL, t2 & m); b.w = __ballot_sync(FULL, t3 & m);
    if (lane == 0) *(uint4*)&A32[i] = b;
  }
  const u32* rel = col + cpad * c;
  const u32 lu = rel[0], lv = rel[c], xu = rel[2 * c], xv = rel[3 * c];
  G32 g;
  g.outU = __ballot_sync(FULL, lu & m);
  g.outV = __ballot_sync(FULL, lv & m);
  g.inU = __ballot_sync(FULL, m && ((xu >> (u & 31)) & 1u));
  g.inV = __ballot_sync(FULL, m && ((xv >> (v & 31)) & 1u));
  __syncwarp();
  return g;
}
// the double move's coming edge u->v is counted in the graph without delme = d0->d1
__device__ __forceinline__ void hot_drop_edge(const u32* list, int c, u32 u, u32 v, u32 d0, u32 d1, u32* A32, int lane, G32& g) {
  const int ip = d0 == u ? -2 : d0 == v ? -3 : list_index(list, c, d0, lane);
  const int iq = d1 == u ? -2 : d1 == v ? -3 : list_index(list, c, d1, lane);
  if (ip == -1 || iq == -1) return;
  if (ip >= 0 && iq >= 0) { if (lane == 0) A32[ip] &= ~(1u << iq); __syncwarp(); }
  else if (ip >= 0) { if (iq == -2) g.inU &= ~(1u << ip); else g.inV &= ~(1u << ip); }
  else if (iq >= 0) { if (ip == -2) g.outU &= ~(1u << iq); else g.outV &= ~(1u << iq); }
}
// ---- hot path from the static clique index: the pair's triangles and 4-cliques are known slots, a
//      proposal only looks up their orientation bits (two bits per adjacent pair, 32-bit words) ----
struct OriReq { u32 sU, sV, wu, wv, e0, w0, e1, w1; };
// issues every load of the pair (static entries, then the orientation words they name); u is the
// pair's bigger vertex iff uBig
__device__ __forceinline__ void idx_issue(const EnsParams& p, const u32* ori, u32 o, int c, u32 k4o, int m, bool uBig,
                                          int lane, OriReq& r) {
  r.sU = r.sV = r.wu = r.wv = r.e0 = r.w0 = r.e1 = r.w1 = 0;
  if (lane < c) { const uint2 t = ldg_idx(&p.tri[o + lane]); r.sU = uBig ? t.x : t.y; r.sV = uBig ? t.y : t.x; }
  if (lane < m) r.e0 = ldg_idx(&p.k4[k4o + lane]);
  if (lane + 32 < m) r.e1 = ldg_idx(&p.k4[k4o + 32 + lane]);
  if (lane < c) { r.wu = __ldcg(&ori[r.sU >> 5]); r.wv = __ldcg(&ori[r.sV >> 5]); }
  if (lane < m) r.w0 = __ldcg(&ori[r.e0 >> 15]);
  if (lane + 32 < m) r.w1 = __ldcg(&ori[r.e1 >> 15]);
}
// entry e = slot(w_i -> w_j) << 10 | j << 5 | i; w = the orientation word of that slot
__device__ __forceinline__ void k4_scatter(u32 e, u32 w, u32* A32) {
  const u32 s = e >> 10, i = e & 31, j = (e >> 5) & 31;
  const u32 two = w >> (s & 30), d = s & 1;
  if ((two >> d) & 1u) atomicOr(&A32[i], 1u << j);
  if ((two >> (d ^ 1)) & 1u) atomicOr(&A32[j], 1u << i);
}
__device__ __forceinline__ G32 idx_finish(const EnsParams& p, const u32* ori, u32 k4o, int m, const OriReq& r, u32* A32,
                                          int lane) {
  A32[lane] = 0;
  __syncwarp();
  if (lane < m) k4_scatter(r.e0, r.w0, A32);
  if (lane + 32 < m) k4_scatter(r.e1, r.w1, A32);
  for (int t = 64 + lane; t < m; t += 32) {
    const u32 e = ldg_idx(&p.k4[k4o + t]);
    k4_scatter(e, __ldcg(&ori[e >> 15]), A32);
  }
  const u32 tu = r.wu >> (r.sU & 30), tv = r.wv >> (r.sV & 30);   // lanes past c hold 0: never vote
  const u32 du = r.sU & 1, dv = r.sV & 1;
  G32 g;
  g.inU = __ballot_sync(FULL, (tu >> du) & 1u);
  g.outU = __ballot_sync(FULL, (tu >> (du ^ 1)) & 1u);
  g.inV = __ballot_sync(FULL, (tv >> dv) & 1u);
  g.outV = __ballot_sync(FULL, (tv >> (dv ^ 1)) & 1u);
  __syncwarp();
  return g;
}

// Tagged graph.  A common neighbour w of u and v can stand before u (tag X: w->u, w->v), between them
// (Y: u->w->v; Y2: v->w->u, the other direction of the same pair, used by a flip) or after v (Z).
// A vertex gets one bit per tag it carries: its first tag on its own bit (< c), further tags on
// extra bits c.. (vertices doubly joined to u or v).  Along a simplex tags never decrease and Y, Y2
// never meet, so with AT[t] = (out-neighbours of t's vertex, expanded to all their tagged copies) &
// (copies whose tag may follow t's) the simplices through the edge are the ordered cliques of AT:
// plain single-mask candidate sets.  Returns false when there are more than 32 tagged copies.
struct Tagged { u32 come, go, row; };                  // root sets of the two counts; this lane's row
__device__ __forceinline__ bool hot_tag(const G32& g, int c, bool flip, const u32* A32, u32* AT, u32* src, int lane,
                                        Tagged& t) {
  const u32 X = g.inU & g.inV, Y = g.outU & g.inV, Y2 = flip ? (g.inU & g.outV) : 0u, Z = g.outU & g.outV;
  const u32 eY = Y & X, eY2 = Y2 & (X | Y), eZ = Z & (X | Y | Y2);
  const int nY = __popc(eY), nY2 = __popc(eY2), nZ = __popc(eZ), ne = nY + nY2 + nZ;
  if (c + ne > 32) return false;
  const u32 below = (1u << lane) - 1, bit = 1u << lane;
  u32 TY = Y & ~X, TY2 = Y2 & ~(X | Y), TZ = Z & ~(X | Y | Y2);
  u32 a = A32[lane];                                    // rows past c are padding (never selected below)
  if (ne) {
    if (eY & bit) src[__popc(eY & below)] = lane;
    if (eY2 & bit) src[nY + __popc(eY2 & below)] = lane;
    if (eZ & bit) src[nY + nY2 + __popc(eZ & below)] = lane;
    TY |= ((1u << nY) - 1) << c;
    TY2 |= ((1u << nY2) - 1) << (c + nY);
    TZ |= ((1u << nZ) - 1) << (c + nY + nY2);
    __syncwarp();
    if (lane >= c && lane < c + ne) a = A32[src[lane - c]];
#pragma unroll 1                                         // (unrolled it is 350 instructions at every inlined copy: the loops that call this live in the instruction cache)
    for (int k = 0; k < ne; k++) a |= ((a >> src[k]) & 1u) << (c + k);
  }
  const u32 allow = (X & bit) ? (X | TY | TY2 | TZ) : (TY & bit) ? (TY | TZ) : (TY2 & bit) ? (TY2 | TZ) : (TZ & bit) ? TZ : 0u;
  t.row = a & allow;
  AT[lane] = t.row;
  t.come = X | TY | TZ;
  t.go = X | TY2 | TZ;
  __syncwarp();
  return true;
}
// Ordered cliques of the tagged graph with >= 2 vertices, rooted at this lane's tagged vertex:
// cnt[k] += sign * (number with k + 2 vertices) (flag_count index k + 3).  Sets *ovf when a clique of
// more than DELTA_SD + 1 vertices exists (the caller runs hot_count_deep for those, or hands the proposal to cold_step).
__device__ __forceinline__ void hot_count(u32 roots, u32 row, const u32* AT, int lane, int sign, int (&cnt)[DELTA_SD], int* ovf) {
  if (!((roots >> lane) & 1u)) return;
  const u32 P1 = row & roots;
  cnt[0] += sign * __popc(P1);
  for (u32 R1 = P1; R1; R1 &= R1 - 1) {
    const u32 P2 = P1 & AT[__ffs((int)R1) - 1];
    if (!P2) continue;
    cnt[1] += sign * __popc(P2);
    for (u32 R2 = P2; R2; R2 &= R2 - 1) {
      const u32 P3 = P2 & AT[__ffs((int)R2) - 1];
      if (!P3) continue;
      cnt[2] += sign * __popc(P3);
      for (u32 R3 = P3; R3; R3 &= R3 - 1)
        if (P3 & AT[__ffs((int)R3) - 1]) *ovf = 1;
    }
  }
}

// Cliques of the tagged graph beyond the register levels: P0 = candidates (non-empty) for the next vertex of a clique of
// nv0 tagged vertices.  Generic walk with a per-lane stack; the counts go to the warp's shared deep[] (by flag_count
// index: a clique of v tagged vertices + u + v is a simplex of dimension v + 1), the highest index met by an addition to
// dflag[FL_TOP].  Same bookkeeping as deep_dfs, but on the single-mask tagged graph the hot path has already built.
__device__ __noinline__ void deep32(u32 P0, int nv0, const u32* AT, int sign, u64* deep, int* dflag) {
  u32 P[FCM_MAXD], R[FCM_MAXD];
  dflag[FL_DEEP] = 1;
  int l = 0;
  P[0] = P0;
  for (;;) {
    const int d = nv0 + l + 2;                            // flag_count index of the cliques the candidates of level l close
    if (d >= FCM_MAXD) { dflag[FL_ERR] = FCM_EDIM; return; }
    atomicAdd((unsigned long long*)&deep[d], (unsigned long long)(long long)(sign * __popc(P[l])));
    if (sign > 0) atomicMax(&dflag[FL_TOP], d);
    R[l] = P[l];
    for (;;) {                                            // next child of level l, or pop
      if (R[l] == 0) {
        if (l == 0) return;
        --l;
        continue;
      }
      const int w = __ffs((int)R[l]) - 1;
      R[l] &= R[l] - 1;
      const u32 c = P[l] & AT[w];
      if (c) { ++l; P[l] = c; break; }
    }
  }
}
// hot_count's walk once more for the cliques it flagged and did not count: those of more than DELTA_SD + 1 tagged
// vertices, through deep32 into deep[] / dflag[] (the caller folds them in).
__device__ __noinline__ void hot_count_deep(u32 roots, u32 row, const u32* AT, int lane, int sign, u64* deep, int* dflag) {
  if (!((roots >> lane) & 1u)) return;
  const u32 P1 = row & roots;
  for (u32 R1 = P1; R1; R1 &= R1 - 1) {
    const u32 P2 = P1 & AT[__ffs((int)R1) - 1];
    for (u32 R2 = P2; R2; R2 &= R2 - 1) {
      const u32 P3 = P2 & AT[__ffs((int)R2) - 1];
      for (u32 R3 = P3; R3; R3 &= R3 - 1) {
        const u32 P4 = P3 & AT[__ffs((int)R3) - 1];
        if (P4) deep32(P4, DELTA_SD + 1, AT, sign, deep, dflag);
      }
    }
  }
}

// Delta of a proposal whose lists need two column slots (32 < c <= 64): register gathers, one pair
// after the other.  Cold path: the warp totals go to deep[] / dflag[] like the counts of the deep
// levels (the caller folds them in), K to dflag[FL_K].
__device__ __noinline__ void delta_wide(const EnsParams& p, const u32* a32, unsigned char* wb, int kind, u32 ta, u32 tb,
                                        u32 tdm0, u32 tdm1, u32 o0, int c0, u32 o1, int c1, int lane) {
  const WarpSmem s = carve(wb, p.Kcap, p.KWcap);
  int cnt[DELTA_SD];
#pragma unroll
  for (int k = 0; k < DELTA_SD; k++) cnt[k] = 0;
  int tri = 0, top = 1;
  Gathered ga;
#pragma unroll 1
  for (int ph = 0; ph < 2; ph++) {
    const int sign = ph ? -1 : 1;
    if (ph == 0 || kind == 1)
      gather_pair(p, a32, ph ? o1 : o0, ph ? c1 : c0, ph ? tdm0 : tb, ph ? tdm1 : ta, ph ? s.cv2 : s.cv, s.rowoff, s.A, lane, ga);
    else {                                              // the flip's going edge: same pair, roles of u and v swapped
      const u64 t = ga.inU; ga.inU = ga.inV; ga.inV = t;
      const u64 t2 = ga.outU; ga.outU = ga.outV; ga.outV = t2;
    }
    if (ph == 0 && kind == 1) drop_edge_from_view(s.cv, c0, tb, ta, tdm0, tdm1, s.A, lane, ga);
    const u64 X = ga.inU & ga.inV, Y = ga.outU & ga.inV, Z = ga.outU & ga.outV;
    tri += sign * (__popcll(X) + __popcll(Y) + __popcll(Z));
    count_through_edge<u64>(X, Y, Z, ga.c, s.A, lane, sign, cnt, s.deep, s.dflag);
    if (ph == 0) {
      const int mytop = cnt[2] ? 5 : cnt[1] ? 4 : cnt[0] ? 3 : (X | Y | Z) ? 2 : 1;
      top = __reduce_max_sync(FULL, mytop);
    }
    __syncwarp();
  }
  const int K = kind == 0 ? c0 + 2 : union_size(s.cv, c0, s.cv2, c1, ta, tb, tdm0, tdm1, lane);
#pragma unroll
  for (int k = 0; k < DELTA_SD; k++) {
    const int t = __reduce_add_sync(FULL, cnt[k]);
    if (lane == 0) s.deep[k + 3] += (u64)(long long)t;
  }
  if (lane == 0) {
    s.deep[2] += (u64)(long long)tri;
    s.dflag[FL_DEEP] = 1;
    s.dflag[FL_TOP] = max(s.dflag[FL_TOP], top);
    s.dflag[FL_K] = K;
  }
  __syncwarp();
}

// Full pre/post counts of the neighbourhood graph (lists longer than FAST_C, or forced): returns this
// lane's new flag_count entry; dflag[FL_RC] = status, dflag[FL_K] = |nbhd|, dflag[FL_LEN] = new length.
__device__ __noinline__ u64 full_path(const EnsParams& p, u32* cb, unsigned char* wb, u32 slot0, u32 add0, u32 slot1,
                                      u32 add1, u64 fc, int len, int lane) {
  const WarpSmem s = carve(wb, p.Kcap, p.KWcap);
  Chain c;
  c.adj = (u64*)cb; c.items = cb + p.off_items; c.pos = cb + p.off_pos; c.dbl = cb + p.off_dbl; c.dblpos = cb + p.off_dblpos;
  c.ori = cb + p.off_ori;
  c.E = 0; c.ND = 0;
  if (lane == 0) { s.edits[0] = slot0; s.edits[1] = add0; s.edits[2] = slot1; s.edits[3] = add1; }
  __syncwarp();
  int K = 0, prelen = 0, postlen = 0;
  int rc = transition_counts(p, c, s, 2, lane, K, prelen, postlen);
  u64 fnew = fc;
  if (!rc) {
    const u64 pr = lane < min(prelen, len) ? s.pre[lane] : 0;
    if (__any_sync(FULL, fc < pr)) rc = FCM_EASSERT;
    else fnew = fc - pr + (lane < postlen ? s.post[lane] : 0);
  }
  if (lane == 0) { s.dflag[FL_RC] = rc; s.dflag[FL_K] = K; s.dflag[FL_LEN] = max(len, max(postlen, prelen)); }
  __syncwarp();
  return fnew;
}

// The edge containers after apply (+ revert), replayed in closed form (both moves keep |E| and the
// number of double pairs).  Lane 0 stores.
//   flip : remove slot (swap_remove: the last element moves into its hole), append slot^1; a revert
//          appends slot again and removes slot^1: either way `fin` ends up last.
//   move : append add = slot^1 (its pair turns double), swap_remove delme: add moves into delme's hole;
//          dpair leaves the double set and add's pair, just appended there, moves into its hole.  A revert
//          removes add again (the old last elements move into the holes) and appends delme / dpair.
__device__ __forceinline__ void update_containers(const EnsParams& p, u32* cb, int kind, bool ok, u32 slot, u32 delme, u32 idx,
                                                  u32 didx, u32 dpair, u32 last, u32 qpos, u32 E, u32 ND, int lane) {
  u32* const items = cb + p.off_items;
  u32* const pos = cb + p.off_pos;
  u32* const dbl = cb + p.off_dbl;
  u32* const dblpos = cb + p.off_dblpos;
  if (kind == 0) {
    const u32 fin = ok ? slot ^ 1 : slot;
    if (lane == 0) {
      items[idx] = last; pos[last] = idx;
      pos[fin ^ 1] = FCM_NONE;
      items[E - 1] = fin; pos[fin] = E - 1;
    }
  } else if (ok) {
    const u32 add = slot ^ 1, pairA = slot >> 1;
    if (lane == 0) {
      items[qpos] = add; pos[add] = qpos; pos[delme] = FCM_NONE;
      dbl[didx] = pairA; dblpos[pairA] = didx; dblpos[dpair] = FCM_NONE;
    }
  } else {
    const u32 last0 = items[E - 1], lastd0 = dbl[ND - 1];
    if (lane == 0) {
      items[qpos] = last0; pos[last0] = qpos;
      items[E - 1] = delme; pos[delme] = E - 1;
      dbl[didx] = lastd0; dblpos[lastd0] = didx;
      dbl[ND - 1] = dpair; dblpos[dpair] = ND - 1;
    }
  }
}
// adjacency bits of an accepted move (cold path): the edge of slot `sgo` = gu->gv leaves, `scome` = cu->cv
// comes; the orientation bits follow.  red = false: plain stores (the row-gather hot path reads the rows
// through this SM's L1, which its own stores keep coherent); red = true: L2 atomics, as the index hot
// path updates the rows (every reader of the rows then uses ld.cg).
__device__ __forceinline__ void update_adjacency(const EnsParams& p, u32* cb, u32 W32, u32 gu, u32 gv, u32 cu, u32 cv,
                                                 u32 sgo, u32 scome, bool red, int lane) {
  if (lane == 0) {
    const u32 igo = gu * W32 + (gv >> 5), icome = cu * W32 + (cv >> 5);
    const u32 mgo = 1u << (gv & 31), mcome = 1u << (cv & 31);
    if (red) { atomicAnd(&cb[igo], ~mgo); atomicOr(&cb[icome], mcome); }
    else if (igo == icome) cb[igo] = (cb[igo] & ~mgo) | mcome;
    else { cb[igo] &= ~mgo; cb[icome] |= mcome; }
    u32* const ori = cb + p.off_ori;
    const u32 wg = sgo >> 5, wc = scome >> 5;
    const u32 og = __ldcg(&ori[wg]) & ~(1u << (sgo & 31));
    if (wg == wc) ori[wg] = og | (1u << (scome & 31));
    else { ori[wg] = og; ori[wc] = __ldcg(&ori[wc]) | (1u << (scome & 31)); }
  }
}

// ---------------------------------------------------------------------------------------------------
// clique moves: Transition::clique_permute / clique_swap (src/lib.rs:214-290).  Cold path: lane 0
// makes the draws (util::random_perm, vec_intersect / vec_setminus), the warp gathers the has_edge bits
// of d x d, builds the edit list in the reference's push order (lane i = row i) and counts the edits
// one after the other as single-edge deltas.
// ---------------------------------------------------------------------------------------------------
// rand 0.8 UniformInt<u32>::sample_single(0, range) as used by seq::gen_index (choose / shuffle):
// next_u32 = upper half of next_u64 (rand_xoshiro 0.6), 32 x 32 widening multiply, rejection zone
__device__ __forceinline__ u32 gen_range_u32(Rng& g, u32 range) {
  const u32 zone = (range << __clz((int)range)) - 1u;
  for (;;) {
    const u32 v = (u32)(g.next() >> 32);
    const u64 m = (u64)v * range;
    if ((u32)m <= zone) return (u32)(m >> 32);
  }
}
// util::random_perm(l, h) (src/util.rs:28-32): out[0..h-l) = shuffle of l..h
__device__ void random_perm_dev(Rng& g, u32 l, u32 h, u32* out) {
  const u32 len = h - l;
#pragma unroll 1
  for (u32 i = 0; i < len; i++) out[i] = l + i;
#pragma unroll 1
  for (u32 i = len; i-- > 1;) {
    const u32 j = gen_range_u32(g, i + 1);
    const u32 t = out[i]; out[i] = out[j]; out[j] = t;
  }
}
// pair id of the adjacent pair {a, b} or FCM_NONE: binary search among the keys of the bigger vertex
__device__ u32 find_pair(const EnsParams& p, u32 a, u32 b) {
  const u32 big = max(a, b);
  const u64 key = ((u64)big << 32) | min(a, b);
  u32 lo = __ldg(&p.pkey_begin[big]);
  const u32 end = __ldg(&p.pkey_begin[big + 1]);
  u32 hi = end;
  while (lo < hi) { const u32 mid = (lo + hi) >> 1; if (__ldg(&p.pkeys[mid]) < key) lo = mid + 1; else hi = mid; }
  if (lo < end && __ldg(&p.pkeys[lo]) == key) return __ldg(&p.pkey_pair[lo]);
  return FCM_NONE;
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// The edits of a clique move are counted one after the other, and each one waits a DRAM round trip for the orientation
// words its pair's triangles / 4-cliques name.  While edit t's words are in flight, the words edit t+1 will ask for are
// pulled into L2 (the static entries that name them are L2 hits): edit t+1 then finds them there - with edit t's own
// change already in them, since that is written through L2 too.  Nothing is held in registers or shared memory.
__device__ __forceinline__ void clique_prefetch_edit(const EnsParams& p, const u32* cb, const u32* ori, u32 slot, u32 W32, int lane) {
  const uint4 i0 = ldg_idx(&p.pinfo[2 * (slot >> 1)]), i1 = ldg_idx(&p.pinfo[2 * (slot >> 1) + 1]);
  const int c = (int)i1.x, m = (int)i1.y;
  if (c > NARROW_C) return;
  if (lane < c) { const uint2 t = ldg_idx(&p.tri[i0.z + lane]); prefetch_l2(&ori[t.x >> 5]); prefetch_l2(&ori[t.y >> 5]); }
  if (lane < m) prefetch_l2(&ori[ldg_idx(&p.k4[i0.w + lane]) >> 15]);
  if (lane + 32 < m) prefetch_l2(&ori[ldg_idx(&p.k4[i0.w + 32 + lane]) >> 15]);
  if (lane == 0) prefetch_l2(&ori[slot >> 5]);             // the word the edit rewrites
}

// kind 2 = clique_permute, 3 = clique_swap.  The chain's RNG state (after the move-kind draw) and all
// other state are in HBM (the hot stretch has been closed).  Returns 0 or an FCM_E* code.
__device__ __noinline__ int clique_step(const EnsParams& p, u32 chain, unsigned char* wb, int lane, int kind,
                                        fcm_trace_rec* tr_out, int force_full, bool red, bool instr) {
  const WarpSmem s = carve(wb, p.Kcap, p.KWcap);
  int* const sflag = (int*)(wb + SM_FLAG);
  ChainState* const st = &p.st[chain];
  u32* const cb = (u32*)p.adj + (u64)chain * p.stride32;
  const u32 W32 = 2 * p.W;
  u32* const dv = s.cq;                                 // d = c ++ a ++ b            (src/lib.rs:240)
  u32* const perm = s.cq + FCM_MAXCQ;                   // perm_d (swap) / perm (permute)
  u32* const H = s.cq + 2 * FCM_MAXCQ;                  // H[i] bit j = has_edge(d[i], d[j])
  u32* const hdr = s.cq + 3 * FCM_MAXCQ;                // [0] n_d [1] n_c [2] n_a [3] status [4] i1 [5] i2 [6] ord [7] nedit
  u32* const R = hdr + 16;                              // clique_swap: [0,32) rank of d[i], [32,64) index in d of a rank, [64,96) new rows, [96,128) old rows
  // ---- the draws and the vertex lists (lane 0) ----
  if (lane == 0) {
    Rng g; g.s0 = st->rng[0]; g.s1 = st->rng[1]; g.s2 = st->rng[2]; g.s3 = st->rng[3];
    int status = 0;
    const u64 r = g.next();                             // clique_order_distribution.sample(rng)
    const double v = __longlong_as_double((long long)((r >> 12) | 0x3ff0000000000000ULL)) - 1.0;
    const double chosen = __dadd_rn(__dmul_rn(v, p.cq_scale), 0.0);
    int ord = 0;
    while (ord < p.cq_norders - 1 && p.cq_cum[ord] <= chosen) ord++;
    const u32 ncl = p.cq_count[ord], k = (u32)ord + 1;
    u32 n_d = 0, n_c = 0, n_a = 0, i1 = 0, i2 = 0;
    if (ncl == 0) status = FCM_EINVAL;                  // reference: choose() -> None -> unwrap panics
    else if (kind == 2) {
      i1 = gen_range_u32(g, ncl);
      if (k > FCM_MAXCQ) status = FCM_ELIMIT;
      else {
        const u32* cl = p.cq_verts + p.cq_base[ord] + (u64)i1 * k;
#pragma unroll 1
        for (u32 i = 0; i < k; i++) dv[i] = __ldg(&cl[i]);
        n_d = k; n_c = k;
        random_perm_dev(g, 0, k, perm);
      }
    } else {
      i1 = gen_range_u32(g, ncl);
      i2 = gen_range_u32(g, ncl);
      const u32* m1 = p.cq_verts + p.cq_base[ord] + (u64)i1 * k;
      const u32* m2 = p.cq_verts + p.cq_base[ord] + (u64)i2 * k;
      if (k > FCM_MAXCQ) status = FCM_ELIMIT;
      else {
        // c = vec_intersect(m1, m2) in m1's order; d = c ++ (m1 - c) ++ (m2 - c)
        u32 inm2 = 0, inm1 = 0;                         // bit i: m1[i] is in m2 / m2[i] is in m1
#pragma unroll 1
        for (u32 i = 0; i < k; i++) {
          const u32 x = __ldg(&m1[i]);
#pragma unroll 1
          for (u32 j = 0; j < k; j++)
            if (x == __ldg(&m2[j])) { inm2 |= 1u << i; inm1 |= 1u << j; }
        }
        n_c = (u32)__popc(inm2); n_a = k - n_c;
        if (2 * k - n_c > FCM_MAXCQ) status = FCM_ELIMIT;
        else {
#pragma unroll 1
          for (u32 i = 0; i < k; i++) if ((inm2 >> i) & 1u) dv[n_d++] = __ldg(&m1[i]);
#pragma unroll 1
          for (u32 i = 0; i < k; i++) if (!((inm2 >> i) & 1u)) dv[n_d++] = __ldg(&m1[i]);
#pragma unroll 1
          for (u32 i = 0; i < k; i++) if (!((inm1 >> i) & 1u)) dv[n_d++] = __ldg(&m2[i]);
          // perm_d = perm_c ++ perm_b ++ perm_a, drawn in the order c, a, b (src/lib.rs:247-251)
          u32* const tmp = R + 128;
          random_perm_dev(g, 0, n_c, perm);
          random_perm_dev(g, n_c, n_c + n_a, tmp);      // perm_a
          random_perm_dev(g, n_c + n_a, n_d, perm + n_c);   // perm_b sits right behind perm_c
#pragma unroll 1
          for (u32 i = 0; i < n_a; i++) perm[n_c + (n_d - n_c - n_a) + i] = tmp[i];
        }
      }
    }
    st->rng[0] = g.s0; st->rng[1] = g.s1; st->rng[2] = g.s2; st->rng[3] = g.s3;
    hdr[0] = n_d; hdr[1] = n_c; hdr[2] = n_a; hdr[3] = (u32)status; hdr[4] = i1; hdr[5] = i2; hdr[6] = (u32)ord; hdr[7] = 0;
  }
  __syncwarp();
  if (hdr[3]) { const int e = (int)hdr[3]; if (lane == 0) st->err = e; return e; }
  const int n_d = (int)hdr[0], n_c = (int)hdr[1], n_a = (int)hdr[2];
  // ---- has_edge among d: lane j probes column d[j] of row d[i] ----
  {
    // every (i, j) probe is an independent load: all of them are in flight together
    if (lane < n_d) H[lane] = 0;
    __syncwarp();
    for (int t = lane; t < n_d * n_d; t += 32) {
      const int i = t / n_d, j = t - i * n_d;
      const u32 col = dv[j];
      if ((__ldcg(&cb[dv[i] * W32 + (col >> 5)]) >> (col & 31)) & 1u) atomicOr(&H[i], 1u << j);
    }
  }
  __syncwarp();
  // ---- the transition: change_edges as (from, to, add) in s.edits, then as (slot, add).  Lane i owns
  //      row i of d; every lane writes its own edits at the position a warp prefix sum gives it, which
  //      is the reference's push order ----
  {
    u32* const ed = s.edits;                            // 2 words per edit: (from << 16 | to) indices into d, add
    u32 first = 0, second = 0, addbits = 0;             // this lane's edits: bits of `first`, then of `second`
    u32 from = 0;
    if (kind == 2) {
      // clique_permute (src/lib.rs:214-232): for i, for j: perm[i] -> perm[j] takes the state of i -> j
      if (lane < n_d) {
        const u32 hp = H[perm[lane]];
        u32 pre = 0;
#pragma unroll 1
        for (int j = 0; j < n_d; j++) pre |= ((hp >> perm[j]) & 1u) << j;
        addbits = H[lane];
        first = pre ^ addbits;
        from = perm[lane];
      }
    } else {
      // clique_swap (src/lib.rs:234-290): new_edges / old_edges are sorted and deduplicated as [Node; 2],
      // i.e. by (from, to) vertex ids: with vertices replaced by their rank among d both sets are 32 x 32
      // bit matrices; the pushes are new \ old in key order (add), then old \ new in key order (remove)
      u32 rk = 0;
      if (lane < n_d) {
        const u32 x = dv[lane];
#pragma unroll 1
        for (int j = 0; j < n_d; j++) rk += dv[j] < x;
        R[lane] = rk; R[32 + rk] = (u32)lane;
      }
      R[64 + lane] = 0; R[96 + lane] = 0;
      __syncwarp();
      if (lane < n_d) {
        const int nca = n_c + n_a;
        const u32 mCA = nca >= 32 ? ~0u : (1u << nca) - 1u, mC = n_c >= 32 ? ~0u : (1u << n_c) - 1u;
        const u32 mAll = n_d >= 32 ? ~0u : (1u << n_d) - 1u, mCB = mC | (mAll & ~mCA);
        u32 h = H[lane] & (lane < n_c ? mAll : lane < nca ? mCA : mCB);
        u32 ow = 0, nw = 0;
        while (h) {
          const int j = __ffs((int)h) - 1;
          h &= h - 1;
          ow |= 1u << R[j];
          nw |= 1u << R[perm[j]];
        }
        R[96 + rk] = ow;
        R[64 + R[perm[lane]]] = nw;
      }
      __syncwarp();
      const u32 nw = R[64 + lane], ow = R[96 + lane];   // lane = rank of the from vertex
      first = nw & ~ow; second = ow & ~nw; addbits = first;
      from = R[32 + lane];
    }
    // positions: all `first` edits of the warp in lane order, then all `second` edits
    const int c1 = __popc(first), c2 = __popc(second);
    int p1 = c1, p2 = c2;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int a1 = __shfl_up_sync(FULL, p1, d), a2 = __shfl_up_sync(FULL, p2, d);
      if (lane >= d) { p1 += a1; p2 += a2; }
    }
    const int n1 = __shfl_sync(FULL, p1, 31), n2 = __shfl_sync(FULL, p2, 31);
    const int ne = n1 + n2;
    if (ne <= FCM_MAXEDIT) {
      int at = p1 - c1;
      while (first) {
        const int j = __ffs((int)first) - 1;
        first &= first - 1;
        const u32 to = kind == 2 ? perm[j] : R[32 + j];
        ed[2 * at] = (from << 16) | to; ed[2 * at + 1] = (addbits >> j) & 1u; at++;
      }
      at = n1 + p2 - c2;
      while (second) {
        const int j = __ffs((int)second) - 1;
        second &= second - 1;
        ed[2 * at] = (from << 16) | R[32 + j]; ed[2 * at + 1] = 0u; at++;
      }
    }
    if (lane == 0) { hdr[3] = ne <= FCM_MAXEDIT ? 0u : (u32)FCM_ELIMIT; hdr[7] = (u32)ne; }
  }
  __syncwarp();
  if (hdr[3]) { const int e = (int)hdr[3]; if (lane == 0) st->err = e; return e; }
  const int nedit = (int)hdr[7];
  bool fast = !force_full;
  {
    bool bad = false, narrow = true;
    for (int t = lane; t < nedit; t += 32) {            // (from, to) -> slot; is every list short enough for the delta form?
      const u32 ft = s.edits[2 * t], from = dv[ft >> 16], to = dv[ft & 0xffffu];
      const u32 pair = find_pair(p, from, to);
      if (pair == FCM_NONE) bad = true;
      else {
        s.edits[2 * t] = 2 * pair + (from > to ? 0u : 1u);
        narrow &= __ldg(&p.nb_off[pair + 1]) - __ldg(&p.nb_off[pair]) <= FAST_C;
      }
    }
    if (__any_sync(FULL, bad)) { if (lane == 0) st->err = FCM_ENOTADJ; return FCM_ENOTADJ; }
    fast = fast && __all_sync(FULL, narrow);
  }
  __syncwarp();
  // ---- State::apply_transition (src/lib.rs:61-79) ----
  // (the chain's containers and flag_count are only picked up behind the counts: the per-edit loop is
  // short of registers under the hot loop's cap)
  u32* const ori = cb + p.off_ori;
  int len = st->fc_len, K = 0, prelen = 0, postlen = 0, rc = 0;
  u64 fnew = 0;
  // DELTA form: post - pre = sum over the edits, applied one after the other, of +-(simplices through
  // the edited edge in the graph as it stands then) - every term is the single-edge count of the flip
  // path, on the static common-neighbour list of that pair.  Lists beyond FAST_C (or kernel kind 1)
  // take the reference's two counts of the whole union neighbourhood instead; same flag_count.
  bool bits_applied = false;
  if (nedit && fast) {
    if (instr && !build_nbhd(p, s.edits, nedit, s.verts, K, lane)) rc = FCM_ELIMIT;
    if (lane < 8) s.dflag[lane] = 0;
    s.deep[lane] = 0;
    __syncwarp();
    int cnt[DELTA_SD];
#pragma unroll
    for (int k = 0; k < DELTA_SD; k++) cnt[k] = 0;
    int tri = 0, edges = 0;
#pragma unroll 1
    for (int t = 0; t < nedit && !rc; t++) {
      const u32 slot = s.edits[2 * t];
      const bool add = s.edits[2 * t + 1] != 0;
      const int sign = add ? 1 : -1;
      // index path: the pair's static entry holds its endpoints, list offset and length (one load, not three)
      uint4 i0 = make_uint4(0, 0, 0, 0), i1 = i0;
      if (red) { i0 = ldg_idx(&p.pinfo[2 * (slot >> 1)]); i1 = ldg_idx(&p.pinfo[2 * (slot >> 1) + 1]); }
      // the edge u -> v; dir 0 of a slot = big -> small
      const uint2 ft = red ? ((slot & 1) ? make_uint2(i0.y, i0.x) : make_uint2(i0.x, i0.y)) : slot_edge(p, slot);
      const u32 o = red ? i0.z : __ldg(&p.nb_off[slot >> 1]);
      const int cc = red ? (int)i1.x : (int)(__ldg(&p.nb_off[(slot >> 1) + 1]) - o);
      edges += sign;
      if (red && cc <= NARROW_C) {
        // the pair's triangles and 4-cliques from the static clique index, their orientation from the
        // chain's orientation bits (kept current edit by edit below): two round trips instead of c rows
        u32* const sA32 = (u32*)(wb + SM_A);
        OriReq r;
        idx_issue(p, ori, i0.z, cc, i0.w, (int)i1.y, !(slot & 1), lane, r);
        if (FCM_CQ_PREFETCH && t + 1 < nedit) clique_prefetch_edit(p, cb, ori, s.edits[2 * (t + 1)], W32, lane);
        const G32 g = idx_finish(p, ori, i0.w, (int)i1.y, r, sA32, lane);
        // first the flip path's tagged single-mask graph (<= 32 tagged copies, <= DELTA_SD + 1 vertices);
        // when it does not fit, the general tagged count on the same gathered rows
        Tagged tg;
        bool done = false;
        if (hot_tag(g, cc, false, sA32, (u32*)(wb + SM_AT), (u32*)(wb + SM_SRC), lane, tg)) {
          hot_count(tg.come, tg.row, (const u32*)(wb + SM_AT), lane, sign, cnt, &sflag[FL_OVF]);
          __syncwarp();
          tri += sign * __popc(tg.come);
          if (sflag[FL_OVF]) {                            // cliques beyond the register levels, on the same tagged graph
            hot_count_deep(tg.come, tg.row, (const u32*)(wb + SM_AT), lane, sign, s.deep, s.dflag);
            __syncwarp();
            if (lane == 0) sflag[FL_OVF] = 0;
            __syncwarp();
          }
          done = true;
        }
        if (!done) {
          const u32 a32v = sA32[lane];
          __syncwarp();
          s.A[lane] = a32v;                             // count_through_edge reads 64-bit rows
          __syncwarp();
          const u32 X = g.inU & g.inV, Y = g.outU & g.inV, Z = g.outU & g.outV;
          tri += sign * (__popc(X) + __popc(Y) + __popc(Z));
          count_through_edge<u32>(X, Y, Z, cc, s.A, lane, sign, cnt, s.deep, s.dflag);
        }
      } else {
        Gathered ga;
        gather_pair(p, cb, o, cc, ft.x, ft.y, s.cv, s.rowoff, s.A, lane, ga);
        const u64 X = ga.inU & ga.inV, Y = ga.outU & ga.inV, Z = ga.outU & ga.outV;
        tri += sign * (__popcll(X) + __popcll(Y) + __popcll(Z));
        count_through_edge<u64>(X, Y, Z, cc, s.A, lane, sign, cnt, s.deep, s.dflag);
      }
      __syncwarp();
      if (lane == 0) {                                  // the next edit sees this one: row bit + orientation bit
        u32* const w = &cb[ft.x * W32 + (ft.y >> 5)];
        const u32 m = 1u << (ft.y & 31);
        if (red) { if (add) atomicOr(w, m); else atomicAnd(w, ~m); }
        else { const u32 wv = __ldcg(w); *w = add ? (wv | m) : (wv & ~m); }
        u32* const ow = &ori[slot >> 5];
        const u32 om = 1u << (slot & 31);
        const u32 ov = __ldcg(ow);
        *ow = add ? (ov | om) : (ov & ~om);
        __threadfence_block();
      }
      __syncwarp();
    }
    bits_applied = true;
    if (!rc && s.dflag[FL_ERR]) rc = s.dflag[FL_ERR];
    if (!rc) {
#pragma unroll
      for (int k = 0; k < DELTA_SD; k++) {
        const int tot = __reduce_add_sync(FULL, cnt[k]);
        if (lane == 0) s.deep[k + 3] += (u64)(long long)tot;
      }
      if (lane == 0) { s.deep[2] += (u64)(long long)tri; s.deep[1] += (u64)(long long)edges; }
      __syncwarp();
      fnew = st->flag_count[lane] + s.deep[lane];
      if (__any_sync(FULL, (long long)fnew < 0)) rc = FCM_EASSERT;
      // post.len() exceeds the old length iff a simplex of a new dimension exists afterwards (src/lib.rs:72-74)
      const u32 nz = __ballot_sync(FULL, fnew != 0);
      len = max(len, nz ? 32 - __clz((int)nz) : 0);
    }
    __syncwarp();
    s.deep[lane] = 0;
    if (lane < 8) s.dflag[lane] = 0;
    __syncwarp();
  } else if (nedit) {
    Chain cf;
    cf.adj = (u64*)cb; cf.items = cf.pos = cf.dbl = cf.dblpos = nullptr; cf.ori = ori; cf.E = cf.ND = 0;
    rc = transition_counts(p, cf, s, nedit, lane, K, prelen, postlen);
    const u64 fc = st->flag_count[lane];
    if (!rc) {
      const u64 pr = lane < min(prelen, len) ? s.pre[lane] : 0;
      if (__any_sync(FULL, fc < pr)) rc = FCM_EASSERT;
      else fnew = fc - pr + (lane < postlen ? s.post[lane] : 0);
      len = max(len, max(postlen, prelen));
    }
  }
  else fnew = st->flag_count[lane];
  if (rc) { if (lane == 0) st->err = rc; return rc; }
  const bool ok = bounds_ok(p, fnew, lane);
  Chain c;
  c.adj = (u64*)cb; c.items = cb + p.off_items; c.pos = cb + p.off_pos; c.dbl = cb + p.off_dbl; c.dblpos = cb + p.off_dblpos;
  c.ori = ori;
  c.E = st->E; c.ND = st->ND;
  // containers: the edits in order; a rejected transition is applied and reverted like the reference
  // does (src/lib.rs:186-191), which leaves the containers in a different ORDER than before
  {
    // the serial container replay below walks pos / dblpos of every edited slot: touch them all at once
    // first so that the replay finds them in cache
    u32 touch = 0;
    for (int t = lane; t < nedit; t += 32) {
      const u32 slot = s.edits[2 * t];
      touch += c.pos[slot] + c.pos[slot ^ 1] + c.dblpos[slot >> 1];
    }
    if (touch == 0x9e3779b9u && lane == 33) sflag[FL_OVF] = 0;    // keeps the loads alive, never true
  }
  for (int pass = 0; pass < (ok ? 1 : 2); pass++)
    for (int t = 0; t < nedit; t++) {
      const u32 slot = s.edits[2 * t];
      const bool add = (s.edits[2 * t + 1] != 0) != (pass == 1);
      if (add) c_add_edge(c, slot, lane); else c_remove_edge(c, slot, lane);
    }
  if (lane == 0 && ok != bits_applied)
    for (int t = 0; t < nedit; t++) {                   // rows + orientation bits: set after the full counts, or taken back after a rejected delta pass
      const u32 slot = s.edits[2 * t];
      const bool add = (s.edits[2 * t + 1] != 0) == ok;
      const uint2 ft = slot_edge(p, slot);
      u32* const w = &cb[ft.x * W32 + (ft.y >> 5)];
      const u32 m = 1u << (ft.y & 31);
      if (red) { if (add) atomicOr(w, m); else atomicAnd(w, ~m); }
      else { const u32 wv = __ldcg(w); *w = add ? (wv | m) : (wv & ~m); }
      u32* const o = &c.ori[slot >> 5];
      const u32 om = 1u << (slot & 31);
      const u32 ov = __ldcg(o);
      *o = add ? (ov | om) : (ov & ~om);
    }
  __syncwarp();
  if (ok) st->flag_count[lane] = fnew;
  if (lane == 0) {
    st->E = c.E; st->ND = c.ND;
    st->fc_len = len;
    st->sampled += 1; st->accepted += ok;
    if (instr && nedit) { st->sumK += (u64)K; st->nonempty += 1; }
    if (tr_out) {
      fcm_trace_rec tr;
      tr.flags = (u32)kind | (ok ? FCM_TR_ACCEPT : 0u) | (nedit ? 0u : FCM_TR_EMPTY);
      tr.a = hdr[4]; tr.b = kind == 3 ? hdr[5] : 0u; tr.dm0 = hdr[6]; tr.dm1 = (u32)nedit; tr.K = (u32)K; tr.draws = 0; tr.pad = 0;
      *tr_out = tr;
    }
    sflag[FL_E] = (int)c.E; sflag[FL_ND] = (int)c.ND;   // the hot stretch keeps both in shared memory
  }
  __syncwarp();
  return 0;
}

// One sampled proposal (parked in shared memory by the hot loop) finished by the general code:
// two-slot lists (32 < c <= 64), more than 32 tagged copies, simplices beyond the register levels,
// lists beyond FAST_C (full path) or a forced full path.  Reads and writes the chain's state in HBM.
__device__ __noinline__ int cold_step(const EnsParams& p, u32 chain, unsigned char* wb, int lane, fcm_trace_rec* tr_out,
                                      int force_full, bool red, bool instr) {
  const u32* park = (const u32*)(wb + SM_PARK);
  u64* const sdeep = (u64*)(wb + SM_DEEP);
  int* const sflag = (int*)(wb + SM_FLAG);
  const int kind = (int)park[PK_KIND];
  if (kind >= 2) return clique_step(p, chain, wb, lane, kind, tr_out, force_full, red, instr);
  const u32 slot = park[PK_SLOT], delme = park[PK_DELME], idx = park[PK_IDX], didx = park[PK_DIDX], dpair = park[PK_DPAIR];
  const u32 last = park[PK_LAST], qpos = park[PK_QPOS], ta = park[PK_TA], tb = park[PK_TB], tdm0 = park[PK_TDM0], tdm1 = park[PK_TDM1];
  const int cA = (int)park[PK_CA], cD = (int)park[PK_CD];
  const u32 oA = park[PK_OA], oD = park[PK_OD];
  ChainState* const st = &p.st[chain];
  u32* const cb = (u32*)p.adj + (u64)chain * p.stride32;
  const u32 W32 = 2 * p.W;
  const u64 fc = st->flag_count[lane];
  int len = st->fc_len;
  const u32 E = st->E, ND = st->ND;
  int err = 0, K = 0;
  u64 fnew = fc;
  const int cB = kind == 0 ? 0 : cD;
  if (lane < 8) sflag[lane] = 0;
  sdeep[lane] = 0;
  __syncwarp();
  if (!force_full && cA <= FAST_C && cB <= FAST_C) {
    delta_wide(p, cb, wb, kind, ta, tb, tdm0, tdm1, oA, cA, oD, cD, lane);
    if (sflag[FL_ERR]) err = sflag[FL_ERR];
    fnew = fc + sdeep[lane];
    if (__any_sync(FULL, (long long)fnew < 0)) err = FCM_EASSERT;
    len = max(len, sflag[FL_TOP] + 1);
    K = sflag[FL_K];
    __syncwarp();
    sdeep[lane] = 0;
    if (lane < 8) sflag[lane] = 0;
    __syncwarp();
  } else {
    const u32 s0 = kind == 0 ? slot : slot ^ 1, s1 = kind == 0 ? slot ^ 1 : delme;
    fnew = full_path(p, cb, wb, s0, kind == 0 ? 0u : 1u, s1, kind == 0 ? 1u : 0u, fc, len, lane);
    if (sflag[FL_RC]) err = sflag[FL_RC];
    K = sflag[FL_K];
    len = sflag[FL_LEN];
  }
  if (err) { if (lane == 0) st->err = err; return err; }
  const bool ok = bounds_ok(p, fnew, lane);
  update_containers(p, cb, kind, ok, slot, delme, idx, didx, dpair, last, qpos, E, ND, lane);
  if (ok) update_adjacency(p, cb, W32, kind == 0 ? ta : tdm0, kind == 0 ? tb : tdm1, tb, ta, kind == 0 ? slot : delme, slot ^ 1, red, lane);
  __syncwarp();
  if (ok) st->flag_count[lane] = fnew;
  if (lane == 0) {
    st->fc_len = len;
    st->sampled += 1; st->accepted += ok;
    if (instr) { st->sumK += (u64)K; st->nonempty += 1; }
    if (tr_out) {
      fcm_trace_rec tr;
      tr.flags = (u32)kind | (ok ? FCM_TR_ACCEPT : 0u);
      tr.a = ta; tr.b = tb; tr.dm0 = tdm0; tr.dm1 = tdm1; tr.K = (u32)K; tr.draws = 0; tr.pad = 0;
      *tr_out = tr;
    }
  }
  __syncwarp();
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// Index hot path, one proposal (flip or double-edge move on lists of <= 32 common neighbours).
//
// Built for the 72 registers a thread has when 28 chains share an SM:
//   * the two move kinds are separate code (a flip never carries the second pair's state);
//   * the orientation words a pair's triangles / 4-cliques name are copied global -> shared with cp.async
//     (LDGSTS: the data never sits in registers while the warp waits); between issue and finish a lane keeps
//     two packed words per pair (the bit positions of its entries), nothing else;
//   * LOOK-AHEAD.  Simple moves keep |E| and the number of double pairs, so once proposal t has made its last
//     draw, the first draws of proposal t+1 (move kind, container index) are known.  They are made right
//     there and the container entries they name are fetched with cp.async while t is being counted (level 1);
//     when those have landed, the pair records and the orientation word of the sampled slot follow (level 2).
//     t's own container / orientation updates are patched into the staged values by the update stage (or the
//     stage is dropped when a patched entry invalidates level 2), so the look-ahead is exact, never a guess.
//     The RNG state with the look-ahead draws lives beside the committed one; leaving the hot stretch (end of
//     launch, cold proposal) drops the look-ahead and keeps the committed state.
// Returns 0 (proposal done), 1 (parked for cold_step) or an FCM_E* code.
// ---------------------------------------------------------------------------------------------------
enum { LA_VALID = 0, LA_L2 = 1, LA_SLOT = 2, LA_DPAIR = 3, LA_KIND = 4, LA_IDX = 5, LA_DIDX = 6, LA_ORIA = 8, LA_ORID = 9, LA_WORDS = 16 };   // [LA_KIND..+3] is written as one uint4
static_assert((SM_TILE + ST_PP) % 16 == 0 && (SM_TILE + ST_PINFO) % 16 == 0 && (SM_TILE + ST_RNGLA) % 16 == 0, "staging block alignment");

__device__ __forceinline__ void cp_async16(u32 saddr, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(g) : "memory");
}
// move kind of a proposal from one draw: WeightedIndex<f64>::sample (src/lib.rs:210)
__device__ __forceinline__ int draw_kind(const EnsParams& p, Rng& g) {
  const u64 r = g.next();
  const double v = __longlong_as_double((long long)((r >> 12) | 0x3ff0000000000000ULL)) - 1.0;
  const double chosen = __dadd_rn(__dmul_rn(v, p.scale), 0.0);
  return (p.cum[0] <= chosen) + (p.cum[1] <= chosen) + (p.cum[2] <= chosen);
}
// every load of a pair: the static entries pass through registers, the orientation words they name go straight into
// shared memory; what the finish stage needs of the entries (bit positions only) is left beside them, packed:
// pk0 = (sU & 31) | (sV & 31) << 5, pk1 = low 15 bits of e0 | of e1 << 15.
__device__ __forceinline__ void stage_issue(const EnsParams& p, const u32* ori, u32 o, int c, u32 k4o, int m, bool uBig,
                                            int lane, u32* stage) {
  u32 pk0 = 0, pk1 = 0;
  const u32 sbase = (u32)__cvta_generic_to_shared(stage + lane);
  if (lane < c) {
    const uint2 t = ldg_idx(&p.tri[o + lane]);
    const u32 sU = uBig ? t.x : t.y, sV = uBig ? t.y : t.x;
    cp_async4(sbase, &ori[sU >> 5]);
    cp_async4(sbase + 128, &ori[sV >> 5]);
    pk0 = (sU & 31u) | ((sV & 31u) << 5);
  }
  if (lane < m) {
    const u32 e0 = ldg_idx(&p.k4[k4o + lane]);
    cp_async4(sbase + 256, &ori[e0 >> 15]);
    pk1 = e0 & 0x7fffu;
  }
  if (lane + 32 < m) {
    const u32 e1 = ldg_idx(&p.k4[k4o + 32 + lane]);
    cp_async4(sbase + 384, &ori[e1 >> 15]);
    pk1 |= (e1 & 0x7fffu) << 15;
  }
  stage[128 + lane] = pk0;
  stage[160 + lane] = pk1;
}
// packed entry pe = (slot & 31) << 10 | j << 5 | i, w = the orientation word of that slot
__device__ __forceinline__ void k4_scatter_packed(u32 pe, u32 w, u32* A32) {
  const u32 sb = pe >> 10, i = pe & 31, j = (pe >> 5) & 31;
  const u32 two = w >> (sb & 30), d = sb & 1;
  if ((two >> d) & 1u) atomicOr(&A32[i], 1u << j);
  if ((two >> (d ^ 1)) & 1u) atomicOr(&A32[j], 1u << i);
}
// staged words -> A32[i] = { j : C[i] -> C[j] } and the four relation masks.  dropslot: a directed slot the view
// must not contain (the double move's leaving edge while the coming edge is counted), FCM_NONE for none.
__device__ __forceinline__ G32 stage_finish(const EnsParams& p, const u32* ori, int c, u32 k4o, int m, const u32* stage,
                                            u32 dropslot, u32 o, bool uBig, u32* A32, int lane) {
  A32[lane] = 0;
  const u32 pk0 = stage[128 + lane], pk1 = stage[160 + lane];
  __syncwarp();
  // the staged words do not say which word they are: a lane finds out from its entries only when a drop is asked for
  u32 wu = 0, wv = 0;
  if (lane < c) { wu = stage[lane]; wv = stage[32 + lane]; }
  u32 w0 = lane < m ? stage[64 + lane] : 0u, w1 = lane + 32 < m ? stage[96 + lane] : 0u;
  if (dropslot != FCM_NONE) {
    const u32 dw = dropslot >> 5, dm = ~(1u << (dropslot & 31));
    if (lane < c) {
      const uint2 t = ldg_idx(&p.tri[o + lane]);
      const u32 sU = uBig ? t.x : t.y, sV = uBig ? t.y : t.x;
      if ((sU >> 5) == dw) wu &= dm;
      if ((sV >> 5) == dw) wv &= dm;
    }
    if (lane < m && (ldg_idx(&p.k4[k4o + lane]) >> 15) == dw) w0 &= dm;
    if (lane + 32 < m && (ldg_idx(&p.k4[k4o + 32 + lane]) >> 15) == dw) w1 &= dm;
  }
  if (lane < m) k4_scatter_packed(pk1 & 0x7fffu, w0, A32);
  if (lane + 32 < m) k4_scatter_packed(pk1 >> 15, w1, A32);
  for (int t = 64 + lane; t < m; t += 32) {
    const u32 e = ldg_idx(&p.k4[k4o + t]);
    u32 w = __ldcg(&ori[e >> 15]);
    if (dropslot != FCM_NONE && (e >> 15) == (dropslot >> 5)) w &= ~(1u << (dropslot & 31));
    k4_scatter(e, w, A32);
  }
  const u32 bu = pk0 & 31u, bv = pk0 >> 5;
  const u32 tu = wu >> (bu & 30), tv = wv >> (bv & 30);               // lanes past c hold 0: never vote
  const u32 du = bu & 1, dv = bv & 1;
  G32 g;
  g.inU = __ballot_sync(FULL, (tu >> du) & 1u);
  g.outU = __ballot_sync(FULL, (tu >> (du ^ 1)) & 1u);
  g.inV = __ballot_sync(FULL, (tv >> dv) & 1u);
  g.outV = __ballot_sync(FULL, (tv >> (dv ^ 1)) & 1u);
  __syncwarp();
  return g;
}

// LA: the look-ahead is compiled in (ensembles of at most 16 chains per SM: +6 % at 1024 chains, +4 % at 64)
template <bool INSTR, bool DEEP, bool LA>
__device__ __forceinline__ int idx_proposal(const EnsParams& p, u32* const cb, unsigned char* const wb, const int lane,
                                            const u32 step, const u32 nsteps, fcm_trace_rec* trace, int force_full,
                                            u32& sumK, u32& nonempty) {
  u32* const sA32 = (u32*)(wb + SM_A);
  u32* const scv = (u32*)(wb + SM_CV);
  u32* const scv2 = (u32*)(wb + SM_CV2);
  int* const sflag = (int*)(wb + SM_FLAG);
  u32* const sAT = (u32*)(wb + SM_AT);
  u32* const ssrc = (u32*)(wb + SM_SRC);
  u32* const spark = (u32*)(wb + SM_PARK);
  u64* const srng = (u64*)(wb + SM_RNG);
  u64* const sfc = (u64*)(wb + SM_FC);
  u32* const stage = (u32*)(wb + SM_TILE + ST_STAGE);
  u32* const sla = (u32*)(wb + SM_TILE + ST_LA);
  u64* const srngla = (u64*)(wb + SM_TILE + ST_RNGLA);
  const uint4* const spinfo = (const uint4*)(wb + SM_TILE + ST_PINFO);
  u32* const ori = cb + p.off_ori;
  const u32 E = (u32)sflag[FL_E], ND = (u32)sflag[FL_ND];
#define ITEMS(i) cb[p.off_items + (i)]
#define POS(i) cb[p.off_pos + (i)]
#define DBL(i) cb[p.off_dbl + (i)]

  // ---- the draws (Transition::random_move, src/lib.rs:207-212) and the sampled container entries ----
  Rng g;
  int kind;
  u32 idx = 0, didx = 0, slot = 0, dpair = 0;
  const bool la = sla[LA_VALID] != 0;
  const bool la2 = la && sla[LA_L2] != 0;
  if (la) {
    cp_async_wait<0>();
    __syncwarp();
    { const ulonglong2 a = *(const ulonglong2*)&srngla[0], b = *(const ulonglong2*)&srngla[2]; g.s0 = a.x; g.s1 = a.y; g.s2 = b.x; g.s3 = b.y; }
    kind = (int)sla[LA_KIND]; idx = sla[LA_IDX]; didx = sla[LA_DIDX]; slot = sla[LA_SLOT]; dpair = sla[LA_DPAIR];
  } else {
    { const ulonglong2 a = *(const ulonglong2*)&srng[0], b = *(const ulonglong2*)&srng[2]; g.s0 = a.x; g.s1 = a.y; g.s2 = b.x; g.s3 = b.y; }
    kind = draw_kind(p, g);
    if (kind > 1) {
      // clique move: the general code takes it from here (the kind draw is consumed)
      if (p.cq_norders == 0) return FCM_EUNSUPPORTED;
      if (lane == 0) {
        *(ulonglong2*)&srng[0] = make_ulonglong2(g.s0, g.s1); *(ulonglong2*)&srng[2] = make_ulonglong2(g.s2, g.s3);
        spark[PK_KIND] = (u32)kind;
      }
      __syncwarp();
      return 1;
    }
    if (kind == 0) {
      if (E) { idx = sample_at(p, (u32)gen_range(g, E), E); slot = ITEMS(idx); }
    } else if (ND) {
      if (2ULL * ND >= E) return FCM_EINVAL;              // reference spins forever (FIXME :307)
      didx = sample_at(p, (u32)gen_range(g, ND), ND);
      idx = sample_at(p, (u32)gen_range(g, E), E);                         // the next draw does not depend on the loaded pair
      dpair = DBL(didx);
      slot = ITEMS(idx);
    }
  }
  bool live = false, sampled_edge = false;
  u32 delme = 0, last = 0, qpos = 0;
  u32 oA = 0, oD = 0, kA = 0, kD = 0, oriA = 0, oriD = 0;
  int cA = 0, cD = 0, mA = 0, mD = 0;
  uint2 abA = make_uint2(0, 0), abD = make_uint2(0, 0);
  if (kind == 0) {
    if (E) {
      // single_edge_flip (src/lib.rs:292-299)
      last = ITEMS(E - 1);
      uint4 i0, i1;
      if (la2) { i0 = spinfo[0]; i1 = spinfo[1]; oriA = sla[LA_ORIA]; }
      else { oriA = __ldcg(&ori[slot >> 5]); i0 = ldg_idx(&p.pinfo[2 * (slot >> 1)]); i1 = ldg_idx(&p.pinfo[2 * (slot >> 1) + 1]); }
      abA = make_uint2(i0.x, i0.y); oA = i0.z; kA = i0.w; cA = (int)i1.x; mA = (int)i1.y;
      live = !((oriA >> ((slot ^ 1) & 31)) & 1u);         // the reverse edge exists: empty transition
      sampled_edge = true;
    }
  } else if (ND) {
    // double_edge_move (src/lib.rs:304-325)
    {
      uint4 i0, i1;
      if (la2) { i0 = spinfo[2]; i1 = spinfo[3]; oriD = sla[LA_ORID]; }
      else { oriD = __ldcg(&ori[dpair >> 4]); i0 = ldg_idx(&p.pinfo[2 * dpair]); i1 = ldg_idx(&p.pinfo[2 * dpair + 1]); }
      abD = make_uint2(i0.x, i0.y); oD = i0.z; kD = i0.w; cD = (int)i1.x; mD = (int)i1.y;
    }
    bool first_try = true;
    for (;;) {
      uint4 i0, i1;
      if (la2 && first_try) { i0 = spinfo[0]; i1 = spinfo[1]; oriA = sla[LA_ORIA]; }
      else { oriA = __ldcg(&ori[slot >> 5]); i0 = ldg_idx(&p.pinfo[2 * (slot >> 1)]); i1 = ldg_idx(&p.pinfo[2 * (slot >> 1) + 1]); }
      abA = make_uint2(i0.x, i0.y); oA = i0.z; kA = i0.w; cA = (int)i1.x; mA = (int)i1.y;
      if (!((oriA >> ((slot ^ 1) & 31)) & 1u)) break;     // a single edge
      first_try = false;
      idx = sample_at(p, (u32)gen_range(g, E), E);
      slot = ITEMS(idx);
    }
    const bool first = g.next() < 0x8000000000000000ULL;  // gen_bool(0.5)
    delme = 2 * dpair + (first ? 1u : 0u);                // first: [min,max] = small -> big
    qpos = POS(delme);                                    // needed when the containers are updated
    live = true;
    sampled_edge = true;
  }
  // (ta,tb) = the sampled single edge, (tdm0,tdm1) = delme; dir 0 of a slot = big -> small
  u32 ta = 0, tb = 0, tdm0 = 0, tdm1 = 0;
  if (sampled_edge) { ta = (slot & 1) ? abA.y : abA.x; tb = (slot & 1) ? abA.x : abA.y; }
  if (kind == 1 && live) { tdm0 = (delme & 1) ? abD.y : abD.x; tdm1 = (delme & 1) ? abD.x : abD.y; }
  // this proposal has made its last draw: commit the RNG
  if (lane == 0) { *(ulonglong2*)&srng[0] = make_ulonglong2(g.s0, g.s1); *(ulonglong2*)&srng[2] = make_ulonglong2(g.s2, g.s3); }
  const bool cold = live && (force_full || cA > NARROW_C || (kind == 1 && cD > NARROW_C));
  // the proposal is parked in shared memory: the update stage (lane 0) and cold_step read it back, so
  // that its scalars do not occupy registers while the simplices are counted
  if (live && lane == 0) {
    *(uint4*)&spark[PK_KIND] = make_uint4((u32)kind, slot, delme, idx);
    *(uint4*)&spark[PK_DIDX] = make_uint4(didx, dpair, last, qpos);
    *(uint4*)&spark[PK_TA] = make_uint4(ta, tb, tdm0, tdm1);
    *(uint4*)&spark[PK_CA] = make_uint4((u32)cA, (u32)cD, oA, oD);
    *(uint2*)&spark[PK_ORIA] = make_uint2(oriA, oriD);
  }
  if (cold) {
    if (lane == 0) sla[LA_VALID] = 0;
    __syncwarp();
    return 1;
  }
  // ---- look-ahead, level 1: the next proposal's kind and container indices; its entries start travelling ----
  {
    bool nla = false;
    if (LA && step + 1 < nsteps) {
      Rng h = g;
      const int k2 = draw_kind(p, h);
      if (k2 == 0 ? E != 0 : (k2 == 1 && ND != 0 && 2ULL * ND < E)) {
        u32 d2 = 0;
        if (k2 == 1) d2 = sample_at(p, (u32)gen_range(h, ND), ND);
        const u32 i2 = sample_at(p, (u32)gen_range(h, E), E);
        if (lane == 0) {
          *(ulonglong2*)&srngla[0] = make_ulonglong2(h.s0, h.s1); *(ulonglong2*)&srngla[2] = make_ulonglong2(h.s2, h.s3);
          *(uint4*)&sla[LA_KIND] = make_uint4((u32)k2, i2, d2, 0u);
          cp_async4((u32)__cvta_generic_to_shared(&sla[LA_SLOT]), &ITEMS(i2));
          if (k2 == 1) cp_async4((u32)__cvta_generic_to_shared(&sla[LA_DPAIR]), &DBL(d2));
        }
        nla = true;
      }
    }
    if (lane == 0) { sla[LA_VALID] = nla ? 1u : 0u; sla[LA_L2] = 0u; }
    cp_async_commit();
  }

  // ---- State::apply_transition  src/lib.rs:61-79: flag_count += post - pre ----
  int K = 0;
  u64 fnew;
  if (live) {
    int cnt[DELTA_SD];
#pragma unroll
    for (int k = 0; k < DELTA_SD; k++) cnt[k] = 0;
    int tri = 0, top = 1;
    bool overflow = false;
    // one copy of every stage, walked by a loop (the hot loop has to stay inside the 32 KB instruction cache; the
    // second pair's parameters wait in shared memory, not in registers): a flip is one pair and two counts on one
    // gather (coming edge tb -> ta, going edge ta -> tb); a double-edge move is the coming edge tb -> ta counted in
    // the graph without delme, then delme's own count
    u32* const spp = (u32*)(wb + SM_TILE + ST_PP);
    if (lane == 0) {
      *(uint4*)&spp[0] = make_uint4(oA, (u32)cA, kA, (u32)mA | (tb == abA.x ? 1u << 16 : 0u));
      *(uint4*)&spp[4] = make_uint4(oD, (u32)cD, kD, (u32)mD | (tdm0 == abD.x ? 1u << 16 : 0u));
    }
    stage_issue(p, ori, oA, cA, kA, mA, tb == abA.x, lane, stage);
    if (kind == 1) {
      stage_issue(p, ori, oD, cD, kD, mD, tdm0 == abD.x, lane, stage + ST_ROWS * 32);
      if constexpr (INSTR) {
        scv[lane] = lane < cA ? __ldg(&p.nb_list[oA + lane]) : 0u;
        scv2[lane] = lane < cD ? __ldg(&p.nb_list[oD + lane]) : 0u;
      }
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncwarp();
    const u32 dropA = kind == 1 ? delme : FCM_NONE;
#pragma unroll 1
    for (int ph = 0; ph < 2; ph++) {
      if (ph == 0 || kind == 1) {
        const uint4 q = *(const uint4*)&spp[4 * ph];      // this pair's parameters, back from shared memory
        const G32 gq = stage_finish(p, ori, (int)q.y, q.z, (int)(q.w & 0xffffu), stage + ph * (ST_ROWS * 32), ph ? FCM_NONE : dropA,
                                    q.x, (q.w >> 16) != 0, sA32, lane);
        Tagged t;
        if (!hot_tag(gq, (int)q.y, kind == 0, sA32, sAT, ssrc, lane, t)) { overflow = true; break; }
        if (lane == 0) { spp[8] = t.come; spp[9] = t.go; }   // root sets and rows wait in shared memory, not in registers
        __syncwarp();
      }
      const u32 roots = spp[(ph == 1 && kind == 0) ? 9 : 8];
      const int sign = ph ? -1 : 1;
      tri += sign * __popc(roots);
      hot_count(roots, sAT[lane], sAT, lane, sign, cnt, &sflag[FL_OVF]);
      if (ph == 0) {
        const int mytop = cnt[2] ? 5 : cnt[1] ? 4 : cnt[0] ? 3 : roots ? 2 : 1;
        top = __reduce_max_sync(FULL, mytop);
      }
      __syncwarp();
      if constexpr (DEEP) {
        // the kernel of graphs with simplices beyond dimension DELTA_SD + 2 (BASELINE config 2 reaches 7): cliques
        // beyond the register levels are counted right here, on the tagged graph that is in place
        if (sflag[FL_OVF]) {
          hot_count_deep(roots, sAT[lane], sAT, lane, sign, (u64*)(wb + SM_DEEP), sflag);
          __syncwarp();
          if (lane == 0) sflag[FL_OVF] = 0;
          __syncwarp();
        }
      }
    }
    if constexpr (INSTR) K = kind == 0 ? cA + 2 : union_size(scv, cA, scv2, cD, ta, tb, tdm0, tdm1, lane);
    __syncwarp();
    if (sflag[FL_OVF]) { overflow = true; __syncwarp(); if (lane == 0) sflag[FL_OVF] = 0; }
    if (overflow) {
      // the parked proposal goes to the general code (nothing of the chain was modified yet)
      cp_async_wait<0>();
      if (lane == 0) sla[LA_VALID] = 0;
      __syncwarp();
      return 1;
    }
    long long delta = 0;
#pragma unroll
    for (int k = 0; k < DELTA_SD; k++) {
      const int tot = __reduce_add_sync(FULL, cnt[k]);
      if (lane == k + 3) delta = tot;
    }
    if (lane == 2) delta = tri;
    if constexpr (DEEP) {
      if (sflag[FL_DEEP]) {
        // cliques beyond the register levels were met (deep32): their counts sit in the warp's deep[] by flag_count index
        u64* const sdeep = (u64*)(wb + SM_DEEP);
        if (sflag[FL_ERR]) return sflag[FL_ERR];
        delta += (long long)sdeep[lane];
        top = max(top, sflag[FL_TOP]);
        __syncwarp();
        sdeep[lane] = 0;
        if (lane < 3) sflag[lane] = 0;
        __syncwarp();
      }
    }
    fnew = sfc[lane] + (u64)delta;
    if (__any_sync(FULL, (long long)fnew < 0)) return FCM_EASSERT;
    if (lane == 0) sflag[FL_FCLEN] = max(sflag[FL_FCLEN], top + 1);
    if constexpr (INSTR) { sumK += K; nonempty++; }
  } else fnew = sfc[lane];

  // ---- look-ahead, level 2: the entries have landed - the next proposal's pair records and orientation words ----
  cp_async_wait<0>();
  __syncwarp();
  if (LA && FCM_LA2 && sla[LA_VALID]) {
    if (lane < 2) {
      const u32 s2 = sla[LA_SLOT];
      cp_async16((u32)__cvta_generic_to_shared(&spinfo[lane]), &p.pinfo[2 * (s2 >> 1) + lane]);
      if (lane == 0) cp_async4((u32)__cvta_generic_to_shared(&sla[LA_ORIA]), &ori[s2 >> 5]);
    } else if (lane < 4 && sla[LA_KIND] == 1u) {
      const u32 d2 = sla[LA_DPAIR];
      cp_async16((u32)__cvta_generic_to_shared(&spinfo[lane]), &p.pinfo[2 * d2 + (lane - 2)]);
      if (lane == 2) cp_async4((u32)__cvta_generic_to_shared(&sla[LA_ORID]), &ori[d2 >> 4]);
    }
  }
  cp_async_commit();

  // ---- Bounds::check, then keep or revert_transition (src/lib.rs:186-191, 81-95) ----
  const bool ok = bounds_ok(p, fnew, lane);
  if (lane == 0) sflag[FL_ACC] += ok;
  if (live) {
    if (lane == 0) {                                      // the parked proposal, read back
      const uint4 q0 = *(const uint4*)&spark[PK_KIND], q1 = *(const uint4*)&spark[PK_DIDX];
      const u32 uslot = q0.y, udelme = q0.z, uidx = q0.w, udidx = q1.x, udpair = q1.y, ulast = q1.z, uqpos = q1.w;
      // what this update writes into the containers, mirrored into the look-ahead's staged entries (exact look-ahead)
      const bool nla = sla[LA_VALID] != 0;
      const u32 nidx = sla[LA_IDX], ndidx = sla[LA_DIDX];
      const bool nmove = sla[LA_KIND] == 1u;
      u32 nslot = sla[LA_SLOT], ndp = sla[LA_DPAIR];
      if (nla) {
        if (kind == 0) {
          const u32 fin = ok ? uslot ^ 1 : uslot;
          if (nidx == uidx) nslot = ulast;
          if (nidx == E - 1) nslot = fin;
        } else if (ok) {
          if (nidx == uqpos) nslot = uslot ^ 1;
          if (nmove && ndidx == udidx) ndp = uslot >> 1;
        } else {
          const u32 last0 = ITEMS(E - 1), lastd0 = DBL(ND - 1);
          if (nidx == uqpos) nslot = last0;
          if (nidx == E - 1) nslot = udelme;
          if (nmove) { if (ndidx == udidx) ndp = lastd0; if (ndidx == ND - 1) ndp = udpair; }
        }
      }
      update_containers(p, cb, kind, ok, uslot, udelme, uidx, udidx, udpair, ulast, uqpos, E, ND, 0);
      u32 wa = FCM_NONE, wd = FCM_NONE;                   // orientation words this update rewrites
      if (ok) {
        // adjacency bits: gu->gv leaves, tb->ta comes
        const uint4 q2 = *(const uint4*)&spark[PK_TA];
        const u32 gu = kind == 0 ? q2.x : q2.z, gv = kind == 0 ? q2.y : q2.w;
        const u32 W32 = 2 * p.W;
        const u32 igo = gu * W32 + (gv >> 5), icome = q2.y * W32 + (q2.x >> 5);
        const u32 mgo = 1u << (gv & 31), mcome = 1u << (q2.x & 31);
        // rows: fire-and-forget L2 atomics (only the cold path reads them, with ld.cg); orientation
        // bits: plain stores of the words read while sampling
        const uint2 q4 = *(const uint2*)&spark[PK_ORIA];
        atomicAnd(&cb[igo], ~mgo); atomicOr(&cb[icome], mcome);
        wa = uslot >> 5;
        if (kind == 0) ori[wa] = q4.x ^ (3u << (uslot & 30));
        else {
          wd = udelme >> 5;
          const u32 ma = 1u << ((uslot ^ 1) & 31), md = 1u << (udelme & 31);
          if (wa == wd) ori[wa] = (q4.x | ma) & ~md;
          else { ori[wa] = q4.x | ma; ori[wd] = q4.y & ~md; }
        }
      }
      if (nla) {
        // level 2 was fetched for the entries as they stood: drop it when an entry changed under it or when one of
        // its orientation words was just rewritten (the next proposal then loads them itself)
        const bool changed = nslot != sla[LA_SLOT] || ndp != sla[LA_DPAIR];
        const bool oriclash = (nslot >> 5) == wa || (nslot >> 5) == wd || (nmove && ((ndp >> 4) == wa || (ndp >> 4) == wd));
        sla[LA_SLOT] = nslot; sla[LA_DPAIR] = ndp;
        sla[LA_L2] = (!LA || !FCM_LA2 || changed || oriclash) ? 0u : 1u;
      }
    }
    if (ok) sfc[lane] = fnew;
    __syncwarp();
  } else if (LA && FCM_LA2 && lane == 0 && sla[LA_VALID]) sla[LA_L2] = 1u;
  if constexpr (INSTR) {
    if (trace != nullptr && lane == 0) {
      fcm_trace_rec tr;
      tr.flags = (u32)kind | (ok ? FCM_TR_ACCEPT : 0u) | (live ? 0u : FCM_TR_EMPTY);
      tr.a = ta; tr.b = tb; tr.dm0 = tdm0; tr.dm1 = tdm1; tr.K = (u32)K; tr.draws = 0; tr.pad = 0;
      trace[step] = tr;
    }
  }
  __syncwarp();
  return 0;
#undef ITEMS
#undef POS
#undef DBL
}

// IDX = true: the hot path reads the static clique index + the chain's orientation bits;
// IDX = false: it gathers from the chain's adjacency rows (graphs too large for the index).
// INSTR = true: per-proposal trace and the neighbourhood-size statistic (sumK / nonempty) are produced.
// MINB = resident blocks per SM the kernel is compiled for: 7 (28 chains per SM, 72 registers per thread: 4096 chains in one
// wave) or 4 (16 chains per SM, 128 registers, no spills) for ensembles that four blocks per SM already hold in one wave.
template <bool IDX, bool INSTR, bool DEEP, int MINB>
__global__ void __launch_bounds__(128, MINB)
chain_step_kernel(const __grid_constant__ EnsParams p, u32 nsteps, fcm_trace_rec* __restrict__ trace, u32 trace_chain,
                  u32 smem_per_warp, int force_full) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31;
  const u32 chain = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (chain >= p.nchains) return;
  unsigned char* const wb = smem_raw + (size_t)(threadIdx.x >> 5) * smem_per_warp;
  u32* const sA32 = (u32*)(wb + SM_A);
  u32* const srowoff = (u32*)(wb + SM_ROWOFF);
  u32* const scv = (u32*)(wb + SM_CV);
  u32* const scv2 = (u32*)(wb + SM_CV2);
  int* const sflag = (int*)(wb + SM_FLAG);
  u32* const sAT = (u32*)(wb + SM_AT);
  u32* const ssrc = (u32*)(wb + SM_SRC);
  u32* const spark = (u32*)(wb + SM_PARK);
  u32* const stile = (u32*)(wb + SM_TILE);
  ChainState* const st = &p.st[chain];
  if (st->err) return;
  // the chain's block: adjacency rows (as 32-bit words) | items | pos | dbl | dblpos | ori
  u32* const cb = (u32*)p.adj + (u64)chain * p.stride32;
#define ITEMS(i) cb[p.off_items + (i)]
#define POS(i) cb[p.off_pos + (i)]
#define DBL(i) cb[p.off_dbl + (i)]
#define ORI(i) cb[p.off_ori + (i)]
  const u32 W32 = 2 * p.W;
  if (lane < 12) sflag[lane] = lane == FL_E ? (int)st->E : lane == FL_ND ? (int)st->ND : 0;   // both moves keep |E| and #double pairs
  ((u64*)(wb + SM_DEEP))[lane] = 0;
  __syncwarp();
  if (!INSTR || chain != trace_chain) trace = nullptr;

  u64* const srng = (u64*)(wb + SM_RNG);
  u64* const sfc = (u64*)(wb + SM_FC);
  u32 step = 0;
  while (step < nsteps) {
    // ---- hot stretch: the chain's scalars live in shared memory (registers are the scarce resource:
    //      28 resident warps per SM leave 72 per thread) until a proposal needs the general code ----
    if (lane < 4) srng[lane] = st->rng[lane];
    sfc[lane] = st->flag_count[lane];                  // lane d owns flag_count[d]
    if (lane == 0) { sflag[FL_ACC] = 0; sflag[FL_FCLEN] = st->fc_len; }
    __syncwarp();
    u32 sumK = 0, nonempty = 0;                        // per stretch (nsteps <= 2^24: the host splits longer runs)
    const u32 step0 = step;
    int err = 0;
    bool cold = false;
    if constexpr (IDX) { if (lane == 0) ((u32*)(wb + SM_TILE + ST_LA))[LA_VALID] = 0; __syncwarp(); }
    for (; step < nsteps; step++) {
      if constexpr (IDX) {
        const int rc = idx_proposal<INSTR, DEEP, (MINB < FCM_CHAIN_MINBLOCKS) || FCM_LA>(p, cb, wb, lane, step, nsteps, trace, force_full, sumK, nonempty);
        if (rc == 1) { cold = true; break; }
        if (rc) { err = rc; break; }
        continue;
      }
      const u32 E = (u32)sflag[FL_E], ND = (u32)sflag[FL_ND];
      Rng g;
      { const ulonglong2 a = *(const ulonglong2*)&srng[0], b = *(const ulonglong2*)&srng[2]; g.s0 = a.x; g.s1 = a.y; g.s2 = b.x; g.s3 = b.y; }
      // ---- Transition::random_move: WeightedIndex<f64>::sample (src/lib.rs:210) ----
      const u64 r = g.next();
      const double v = __longlong_as_double((long long)((r >> 12) | 0x3ff0000000000000ULL)) - 1.0;
      const double chosen = __dadd_rn(__dmul_rn(v, p.scale), 0.0);
      const int kind = (p.cum[0] <= chosen) + (p.cum[1] <= chosen) + (p.cum[2] <= chosen);
      if (kind > 1) {
        // clique move: the general code takes it from here (the kind draw is consumed)
        if (p.cq_norders == 0) { err = FCM_EUNSUPPORTED; break; }
        if (lane == 0) {
          *(ulonglong2*)&srng[0] = make_ulonglong2(g.s0, g.s1); *(ulonglong2*)&srng[2] = make_ulonglong2(g.s2, g.s3);
          spark[PK_KIND] = (u32)kind;
        }
        cold = true;
        __syncwarp();
        break;
      }
      // ---- single_edge_flip (src/lib.rs:292-299) / double_edge_move (:304-325): the draws.  Loads are
      //      written in the order they become issuable so that independent ones overlap. ----
      bool live = false;                                // a non-empty transition
      bool sampled_edge = false;
      u32 idx = 0, slot = 0, didx = 0, dpair = 0, delme = 0, last = 0, qpos = 0;
      u32 oA = 0, oD = 0, kA = 0, kD = 0, oriA = 0, oriD = 0;
      int cA = 0, cD = 0, mA = 0, mD = 0;
      uint2 abA = make_uint2(0, 0), abD = make_uint2(0, 0);
      if (kind == 0) {
        if (E) {
          idx = sample_at(p, (u32)gen_range(g, E), E);
          slot = ITEMS(idx);
          last = ITEMS(E - 1);
          if constexpr (IDX) {
            oriA = __ldcg(&ORI(slot >> 5));
            const uint4 i0 = ldg_idx(&p.pinfo[2 * (slot >> 1)]), i1 = ldg_idx(&p.pinfo[2 * (slot >> 1) + 1]);
            abA = make_uint2(i0.x, i0.y); oA = i0.z; kA = i0.w; cA = (int)i1.x; mA = (int)i1.y;
            live = !((oriA >> ((slot ^ 1) & 31)) & 1u);     // the reverse edge exists: empty transition
          } else {
            const u32 rev = POS(slot ^ 1);
            abA = __ldg(&p.pair_ab[slot >> 1]);
            oA = __ldg(&p.nb_off[slot >> 1]);
            cA = (int)(__ldg(&p.nb_off[(slot >> 1) + 1]) - oA);
            live = rev == FCM_NONE;
          }
          sampled_edge = true;
        }
      } else if (ND) {
        if (2ULL * ND >= E) { err = FCM_EINVAL; break; }       // reference spins forever (FIXME :307)
        didx = sample_at(p, (u32)gen_range(g, ND), ND);
        idx = sample_at(p, (u32)gen_range(g, E), E);                     // the next draw does not depend on the loaded pair
        dpair = DBL(didx);
        slot = ITEMS(idx);
        if constexpr (IDX) {
          oriD = __ldcg(&ORI(dpair >> 4));
          const uint4 i0 = ldg_idx(&p.pinfo[2 * dpair]), i1 = ldg_idx(&p.pinfo[2 * dpair + 1]);
          abD = make_uint2(i0.x, i0.y); oD = i0.z; kD = i0.w; cD = (int)i1.x; mD = (int)i1.y;
        } else {
          abD = __ldg(&p.pair_ab[dpair]);
          oD = __ldg(&p.nb_off[dpair]);
          cD = (int)(__ldg(&p.nb_off[dpair + 1]) - oD);
        }
        for (;;) {
          bool single;
          if constexpr (IDX) {
            oriA = __ldcg(&ORI(slot >> 5));
            const uint4 i0 = ldg_idx(&p.pinfo[2 * (slot >> 1)]), i1 = ldg_idx(&p.pinfo[2 * (slot >> 1) + 1]);
            abA = make_uint2(i0.x, i0.y); oA = i0.z; kA = i0.w; cA = (int)i1.x; mA = (int)i1.y;
            single = !((oriA >> ((slot ^ 1) & 31)) & 1u);
          } else {
            const u32 rev = POS(slot ^ 1);
            abA = __ldg(&p.pair_ab[slot >> 1]);
            oA = __ldg(&p.nb_off[slot >> 1]);
            cA = (int)(__ldg(&p.nb_off[(slot >> 1) + 1]) - oA);
            single = rev == FCM_NONE;
          }
          if (single) break;
          idx = sample_at(p, (u32)gen_range(g, E), E);
          slot = ITEMS(idx);
        }
        const bool first = g.next() < 0x8000000000000000ULL;   // gen_bool(0.5)
        delme = 2 * dpair + (first ? 1u : 0u);                 // first: [min,max] = small -> big
        qpos = POS(delme);                                     // needed when the containers are updated
        live = true;
        sampled_edge = true;
      }
      // (ta,tb) = the sampled single edge, (tdm0,tdm1) = delme; dir 0 of a slot = big -> small
      u32 ta = 0, tb = 0, tdm0 = 0, tdm1 = 0;
      if (sampled_edge) { ta = (slot & 1) ? abA.y : abA.x; tb = (slot & 1) ? abA.x : abA.y; }
      if (kind == 1 && live) { tdm0 = (delme & 1) ? abD.y : abD.x; tdm1 = (delme & 1) ? abD.x : abD.y; }
      if (lane == 0) { *(ulonglong2*)&srng[0] = make_ulonglong2(g.s0, g.s1); *(ulonglong2*)&srng[2] = make_ulonglong2(g.s2, g.s3); }
      // the proposal is parked in shared memory: the update stage (lane 0) and cold_step read it back, so
      // that its scalars do not occupy registers while the simplices are counted
      if (live && lane == 0) {
        *(uint4*)&spark[PK_KIND] = make_uint4((u32)kind, slot, delme, idx);
        *(uint4*)&spark[PK_DIDX] = make_uint4(didx, dpair, last, qpos);
        *(uint4*)&spark[PK_TA] = make_uint4(ta, tb, tdm0, tdm1);
        *(uint4*)&spark[PK_CA] = make_uint4((u32)cA, (u32)cD, oA, oD);
        *(uint2*)&spark[PK_ORIA] = make_uint2(oriA, oriD);
      }

      // ---- State::apply_transition  src/lib.rs:61-79: flag_count += post - pre ----
      int K = 0;
      u64 fnew;
      u32 vgo = 0, vcome = 0;                           // row-gather path: the two row words an accept rewrites
      if (live) {
        const int cB = kind == 0 ? 0 : cD;
        cold = force_full || cA > NARROW_C || cB > NARROW_C;
        int cnt[DELTA_SD];
#pragma unroll
        for (int k = 0; k < DELTA_SD; k++) cnt[k] = 0;
        int tri = 0, top = 1;
        if (!cold) {
          Tagged t;
          if constexpr (IDX) {
            // every load of both pairs is in flight before the first one is used
            u32* const ori = cb + p.off_ori;
            OriReq rA, rD;
            idx_issue(p, ori, oA, cA, kA, mA, tb == abA.x, lane, rA);
            if (kind == 1) {
              idx_issue(p, ori, oD, cD, kD, mD, tdm0 == abD.x, lane, rD);
              // the vertex lists are only needed to take delme out of the coming pair's view and for K
              scv[lane] = lane < cA ? __ldg(&p.nb_list[oA + lane]) : 0u;
              if constexpr (INSTR) scv2[lane] = lane < cD ? __ldg(&p.nb_list[oD + lane]) : 0u;
            }
#pragma unroll 1
            for (int ph = 0; ph < 2 && !cold; ph++) {
              const int sign = ph ? -1 : 1;
              if (ph == 0 || kind == 1) {
                G32 ga = ph == 0 ? idx_finish(p, ori, kA, mA, rA, sA32, lane) : idx_finish(p, ori, kD, mD, rD, sA32, lane);
                if (ph == 0 && kind == 1) hot_drop_edge(scv, cA, tb, ta, tdm0, tdm1, sA32, lane, ga);
                if (!hot_tag(ga, ph ? cD : cA, kind == 0, sA32, sAT, ssrc, lane, t)) { cold = true; break; }
              }
              const u32 roots = (ph == 1 && kind == 0) ? t.go : t.come;
              tri += sign * __popc(roots);
              hot_count(roots, t.row, sAT, lane, sign, cnt, &sflag[FL_OVF]);
              if (ph == 0) {
                const int mytop = cnt[2] ? 5 : cnt[1] ? 4 : cnt[0] ? 3 : roots ? 2 : 1;
                top = __reduce_max_sync(FULL, mytop);
              }
              __syncwarp();
            }
          } else {
            const int wordsA = hot_tile_words(cA);
            // the edge that comes is tb->ta for both moves; the edge that goes is ta->tb (flip) or delme
            vgo = cb[(kind == 0 ? ta : tdm0) * W32 + ((kind == 0 ? tb : tdm1) >> 5)];
            vcome = cb[tb * W32 + (ta >> 5)];             // early: rewritten on accept without a stall
            // both gathers are in flight together when the tile holds both, else one after the other
            const bool both = wordsA + hot_tile_words(cB) <= TILE_WORDS;
            u32* const tileD = both ? stile + wordsA : stile;
            const u32 xA = lane < cA ? __ldg(&p.nb_list[oA + lane]) : 0u;
            const u32 xD = lane < cB ? __ldg(&p.nb_list[oD + lane]) : 0u;
            hot_issue(cb, W32, xA, cA, tb, ta, scv, srowoff, stile, lane);
            if (kind == 1 && both) hot_issue(cb, W32, xD, cD, tdm0, tdm1, scv2, srowoff + 32, tileD, lane);
#pragma unroll 1
            for (int ph = 0; ph < 2 && !cold; ph++) {
              const int sign = ph ? -1 : 1;
              if (ph == 0 || kind == 1) {
                if (ph == 1 && !both) hot_issue(cb, W32, xD, cD, tdm0, tdm1, scv2, srowoff + 32, tileD, lane);
                if (ph == 0 && kind == 1 && both) cp_async_wait<1>(); else cp_async_wait<0>();
                __syncwarp();
                G32 ga = ph == 0 ? hot_finish(stile, cA, scv, tb, ta, sA32, lane)
                                 : hot_finish(tileD, cD, scv2, tdm0, tdm1, sA32, lane);
                if (ph == 0 && kind == 1) hot_drop_edge(scv, cA, tb, ta, tdm0, tdm1, sA32, lane, ga);
                if (!hot_tag(ga, ph ? cD : cA, kind == 0, sA32, sAT, ssrc, lane, t)) { cold = true; break; }
              }
              const u32 roots = (ph == 1 && kind == 0) ? t.go : t.come;
              tri += sign * __popc(roots);
              hot_count(roots, t.row, sAT, lane, sign, cnt, &sflag[FL_OVF]);
              if (ph == 0) {
                const int mytop = cnt[2] ? 5 : cnt[1] ? 4 : cnt[0] ? 3 : roots ? 2 : 1;
                top = __reduce_max_sync(FULL, mytop);
              }
              __syncwarp();
            }
            cp_async_wait<0>();
          }
          if (sflag[FL_OVF]) { cold = true; __syncwarp(); if (lane == 0) sflag[FL_OVF] = 0; }
        }
        if (cold) {
          // the parked proposal goes to the general code (nothing of the chain was modified yet)
          __syncwarp();
          break;
        }
        if constexpr (INSTR) K = kind == 0 ? cA + 2 : union_size(scv, cA, scv2, cD, ta, tb, tdm0, tdm1, lane);
        long long delta = 0;
#pragma unroll
        for (int k = 0; k < DELTA_SD; k++) {
          const int tot = __reduce_add_sync(FULL, cnt[k]);
          if (lane == k + 3) delta = tot;
        }
        if (lane == 2) delta = tri;
        fnew = sfc[lane] + (u64)delta;
        if (__any_sync(FULL, (long long)fnew < 0)) { err = FCM_EASSERT; break; }
        if (lane == 0) sflag[FL_FCLEN] = max(sflag[FL_FCLEN], top + 1);
        if constexpr (INSTR) { sumK += K; nonempty++; }
      } else fnew = sfc[lane];
      // ---- Bounds::check, then keep or revert_transition (src/lib.rs:186-191, 81-95) ----
      const bool ok = bounds_ok(p, fnew, lane);
      if (lane == 0) sflag[FL_ACC] += ok;
      if (live) {
        if (lane == 0) {                                // the parked proposal, read back
          const uint4 q0 = *(const uint4*)&spark[PK_KIND], q1 = *(const uint4*)&spark[PK_DIDX];
          const u32 uslot = q0.y, udelme = q0.z;
          update_containers(p, cb, kind, ok, uslot, udelme, q0.w, q1.x, q1.y, q1.z, q1.w, E, ND, 0);
          if (ok) {
            // adjacency bits: gu->gv leaves, tb->ta comes
            const uint4 q2 = *(const uint4*)&spark[PK_TA];
            const u32 gu = kind == 0 ? q2.x : q2.z, gv = kind == 0 ? q2.y : q2.w;
            const u32 igo = gu * W32 + (gv >> 5), icome = q2.y * W32 + (q2.x >> 5);
            const u32 mgo = 1u << (gv & 31), mcome = 1u << (q2.x & 31);
            if constexpr (IDX) {
              // rows: fire-and-forget L2 atomics (only the cold path reads them, with ld.cg); orientation
              // bits: plain stores of the words read while sampling
              const uint2 q4 = *(const uint2*)&spark[PK_ORIA];
              atomicAnd(&cb[igo], ~mgo); atomicOr(&cb[icome], mcome);
              if (kind == 0) ORI(uslot >> 5) = q4.x ^ (3u << (uslot & 30));
              else {
                const u32 wa = uslot >> 5, wd = udelme >> 5;
                const u32 ma = 1u << ((uslot ^ 1) & 31), md = 1u << (udelme & 31);
                if (wa == wd) ORI(wa) = (q4.x | ma) & ~md;
                else { ORI(wa) = q4.x | ma; ORI(wd) = q4.y & ~md; }
              }
            } else {
              // plain stores of the words read above: the rows are read through this SM's L1, which its
              // own stores keep coherent
              if (igo == icome) cb[igo] = (vgo & ~mgo) | mcome;
              else { cb[igo] = vgo & ~mgo; cb[icome] = vcome | mcome; }
              u32* const ori = cb + p.off_ori;
              const u32 sg = kind == 0 ? uslot : udelme, sc = uslot ^ 1;
              const u32 og = ori[sg >> 5] & ~(1u << (sg & 31));
              if ((sg >> 5) == (sc >> 5)) ori[sg >> 5] = og | (1u << (sc & 31));
              else { ori[sg >> 5] = og; ori[sc >> 5] |= 1u << (sc & 31); }
            }
          }
        }
        if (ok) sfc[lane] = fnew;
        __syncwarp();
      }
      if constexpr (INSTR) {
        if (trace != nullptr && lane == 0) {
          fcm_trace_rec tr;
          tr.flags = (u32)kind | (ok ? FCM_TR_ACCEPT : 0u) | (live ? 0u : FCM_TR_EMPTY);
          tr.a = ta; tr.b = tb; tr.dm0 = tdm0; tr.dm1 = tdm1; tr.K = (u32)K; tr.draws = 0; tr.pad = 0;
          trace[step] = tr;
        }
      }
    }
    // ---- end of the stretch: the chain's state goes back to HBM ----
    if constexpr (IDX) cp_async_wait<0>();                // a look-ahead in flight lands before its buffer is reused
    __syncwarp();
    st->flag_count[lane] = sfc[lane];
    if (lane < 4) st->rng[lane] = srng[lane];
    if (lane == 0) {
      st->fc_len = sflag[FL_FCLEN];
      st->sampled += step - step0; st->accepted += (u32)sflag[FL_ACC];
      if constexpr (INSTR) { st->sumK += sumK; st->nonempty += nonempty; }
      if (err) st->err = err;
    }
    __syncwarp();
    if (err || !cold) break;
    if (cold_step(p, chain, wb, lane, trace ? trace + step : nullptr, force_full, IDX, INSTR)) break;
    step++;
  }
#undef ITEMS
#undef POS
#undef DBL
#undef ORI
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
  c.adj = (u64*)((u32*)p.adj + (u64)chain * p.stride32);
  c.items = p.items + (u64)chain * p.stride32;
  c.pos = p.pos + (u64)chain * p.stride32;
  c.dbl = p.dbl + (u64)chain * p.stride32;
  c.dblpos = p.dblpos + (u64)chain * p.stride32;
  c.ori = p.ori + (u64)chain * p.stride32;
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
  const u64 words = p.stride32;                         // chain 0's block -> every other chain's
  int blocks = (int)std::min<u64>((words + 255) / 256, (u64)e->ctx->sm_count * 8);
  replicate_kernel<u32><<<blocks, 256, 0, s>>>((u32*)p.adj, words, p.nchains);
  e->ctx->launches++;
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

template <bool IDX, bool INSTR, bool DEEP, int MINB>
static int launch_step_mode(fcm_ensemble* e, u64 nsteps, fcm_trace_rec* trace_dev, u32 trace_chain) {
  EnsParams p = e->p;
  // ensembles with clique moves size the local matrix M for the union neighbourhood of two cliques; when
  // every common-neighbour list fits the delta path and the full counts are not forced, M is never
  // touched: one word per row keeps the per-warp tile small and more chains resident
  if (e->cliques && !e->force_full && p.maxc <= FAST_C) {
    p.KWcap = 1;
    if (!INSTR) p.Kcap = 8;                             // the vertex list is only built for the statistic K
  }
  const size_t per_warp = chain_smem_per_warp_mode(p.Kcap, p.KWcap, !IDX, e->cliques != 0);
  // warps per block: the choice that keeps most warps resident on one SM (big neighbourhood tiles of
  // ensembles with clique moves fit better in smaller blocks)
  int wpb = 4;
  {
    size_t best = 0;
    for (int w = 4; w >= 1; w >>= 1) {
      if (per_warp * w > e->ctx->smem_optin) continue;
      // blocks per SM: registers (72 per thread, 64 K per SM => 28 warps) and shared memory
      const size_t resident = std::min<size_t>((MINB * 4) / w, (size_t)(227 * 1024) / (per_warp * w + 1024)) * w;
      if (resident > best) { best = resident; wpb = w; }
    }
  }
  if (per_warp * wpb > e->ctx->smem_optin) { fcm_set_error("neighbourhood tile does not fit shared memory"); return FCM_ELIMIT; }
  const size_t smem = per_warp * wpb;
  // a kernel's attributes are state of the function, not of the launch: two host threads stepping ensembles of different
  // shapes (src/bin/all_cxs.rs runs a private State per OS thread) must not interleave "set the size" and "launch"
  static std::mutex launch_mu;
  std::lock_guard<std::mutex> launch_lock(launch_mu);
  FCM_CUDA(cudaFuncSetAttribute(chain_step_kernel<IDX, INSTR, DEEP, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    // shared-memory carve-out: what the resident blocks need and no more - the rest of the 256 KB stays L1 (spilled
    // registers, the static index entries a proposal re-reads).  FCM_CARVEOUT=<percent> overrides (measurements).
    const size_t per_sm = std::min<size_t>((MINB * 4) / wpb, (size_t)(227 * 1024) / (smem + 1024));
    int pct = (int)((per_sm * (smem + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024));
    if (const char* cv = getenv("FCM_CARVEOUT")) pct = atoi(cv);
    pct = std::max(0, std::min(100, pct));
    FCM_CUDA(cudaFuncSetAttribute(chain_step_kernel<IDX, INSTR, DEEP, MINB>, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
  }
  const u32 blocks = (p.nchains + wpb - 1) / wpb;
  const u64 chunk = 1ULL << 24;                         // the kernel's per-launch counters are 32-bit
  for (u64 done = 0; done < nsteps; done += chunk) {
    const u32 now = (u32)std::min<u64>(chunk, nsteps - done);
    // the launch carries the L2 access-policy window of the clique index (persisting hits, streaming misses)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(wpb * 32); cfg.dynamicSmemBytes = smem; cfg.stream = e->ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeAccessPolicyWindow;
    attr[0].val.accessPolicyWindow = e->l2_window;
    cfg.attrs = attr; cfg.numAttrs = e->l2_window.num_bytes ? 1 : 0;
    FCM_CUDA(cudaLaunchKernelEx(&cfg, chain_step_kernel<IDX, INSTR, DEEP, MINB>, p, now, trace_dev ? trace_dev + done : (fcm_trace_rec*)nullptr,
                                trace_chain, (u32)per_warp, (int)e->force_full));
    e->ctx->launches++;
  }
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

template <int MINB>
static int launch_step_pick(fcm_ensemble* e, u64 nsteps, fcm_trace_rec* trace_dev, u32 trace_chain) {
  const bool idx = e->have_index && !e->force_rows, instr = trace_dev != nullptr || e->stats;
  // graphs that start with simplices beyond dimension DELTA_SD + 2 run the index kernel that counts deep cliques in
  // place; the others keep the leaner loop (a deep clique met later sends that proposal through cold_step: same result)
  if (idx && e->deep_hint)
    return instr ? launch_step_mode<true, true, true, MINB>(e, nsteps, trace_dev, trace_chain)
                 : launch_step_mode<true, false, true, MINB>(e, nsteps, trace_dev, trace_chain);
  if (idx) return instr ? launch_step_mode<true, true, false, MINB>(e, nsteps, trace_dev, trace_chain)
                        : launch_step_mode<true, false, false, MINB>(e, nsteps, trace_dev, trace_chain);
  return instr ? launch_step_mode<false, true, false, MINB>(e, nsteps, trace_dev, trace_chain)
               : launch_step_mode<false, false, false, MINB>(e, nsteps, trace_dev, trace_chain);
}
int fcm_chain_launch_step(fcm_ensemble* e, u64 nsteps, fcm_trace_rec* trace_dev, u32 trace_chain) {
  // ensembles that 4 blocks x 4 warps per SM hold in one wave get the 128-register instance (no spills: +5 % at
  // 1024 / 2048 chains on C3, profiles/README.md); larger ones need 7 blocks per SM to stay in one wave
  const bool small = (u64)e->p.nchains <= (u64)e->ctx->sm_count * 16;
  return small ? launch_step_pick<4>(e, nsteps, trace_dev, trace_chain)
               : launch_step_pick<FCM_CHAIN_MINBLOCKS>(e, nsteps, trace_dev, trace_chain);
}

int fcm_chain_launch_apply(fcm_ensemble* e, u32 chain, int nedits, int mode) {
  size_t smem = e->smem_per_warp;
  static std::mutex launch_mu;
  std::lock_guard<std::mutex> launch_lock(launch_mu);
  FCM_CUDA(cudaFuncSetAttribute(chain_apply_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  chain_apply_kernel<<<1, 32, smem, e->ctx->stream>>>(e->p, chain, e->edit_dev, nedits, mode, e->io_dev);
  e->ctx->launches++;
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}
