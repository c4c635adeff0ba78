// Fused attention forward for sm_100a:  O = softmax(Q K^T * scale [+ key mask]) V, online softmax.
//
// One CTA per (kQTiles x 128 query rows, head, sample); the query tiles share every K/V tile.  One CTA per SM
// (it owns all 512 TMEM columns); 64 + kQTiles*128 threads:
//   warp 0 (one lane)  TMA producer: Q tiles once, then K and V tiles of 64 keys in two separate rings
//                      (64-column boxes, SWIZZLE_128B; the head is a column slice of the [B, N, ld] buffer)
//   warp 1 (one lane)  tcgen05.mma issuer, event driven: when tile t publishes P_t(j) it issues O_t += P_t(j) V_j
//                      (A operand = P straight from TMEM, V is the MN-major B operand from its [key][d] layout) and
//                      then S_t(j+1) = Q_t K_{j+1}^T over the same TMEM columns (the tensor pipe keeps issue order);
//                      whichever tile is ready first goes first
//   warps 2 + 4t + q   softmax of query tile t, TMEM lane quarter q, one query row per thread: tcgen05.ld S ->
//                      registers, running max with lazy rescaling of O in TMEM (only when the max grows by > 2^8),
//                      exp2, row sum, P -> bf16 pairs -> tcgen05.st over the first 32 columns of the row's own S
//                      (P never touches shared memory: the smem port carries Q/K/V operand reads only);
//                      final O / l -> bf16 -> global.
// Head dim 40 is bound by the MUFU unit (one ex2 per 80 tensor FLOPs), not by the tensor cores, so the kernel is
// organised around keeping MUFU fed: kQTiles = 4 puts four independent softmax warps on every SM sub-partition,
// so that while some wait for scores, take the row max or publish P, others are in their exponential phase.  The
// exponential phase is FFMA2 (scale and max folded) -> MUFU.EX2 -> FADD2 -> F2FP, the row max FMNMX3 on raw scores;
// one pair of exponentials in three runs on the FMA pipe instead (degree-3 minimax exp2).
// Two more configurations of the same kernel (AttnCfg): head dim 80 keeps TWO S buffers per query tile so that
// S(j+1) is computed while the softmax warps work on S(j) (kSDouble), and short key sequences (the 77-token context)
// run as one-query-tile CTAs with 128 / 256 TMEM columns, several per SM (kSmall).
//
// Replaces attention.py:198-242 (einsum QK^T, masked_fill, softmax, einsum PV, head merge) without ever
// materialising the [8B, N, N] score tensor.
#include "common.cuh"
#include "../../include/adaface_b200.h"

namespace afb {
void count_launch();
namespace {

constexpr int kBlockQ = 128;                // rows of one query tile (= TMEM lanes)
constexpr int kBlockKV = 64;                // keys per tile (= one 128-byte swizzle atom of P)
constexpr float kMaskedScore = -1.0e30f;   // finite: a fully masked row degrades to a uniform average,
                                           // like masked_fill(-finfo.max) + softmax in the reference
constexpr float kRescaleThreshold = 8.0f;  // log2 units

struct AttnDev {
  __nv_bfloat16* out;
  float* lse;               // optional [B, heads, q_len]: log2-domain log-sum-exp of the scaled scores (for the backward)
  const uint8_t* key_mask;
  int q_len, kv_len, heads;
  long long ldo;
  float scale_log2;
#ifdef AFB_ATTN_PROFILE
  long long* prof;          // [prof_blocks][16 softmax warps][8 phases] clock64 sums (tools/attn_probe.cu)
  int prof_blocks;
#endif
};

#ifdef AFB_ATTN_PROFILE
#define AFB_PROF(idx)                   \
  do {                                  \
    const long long _now = clock64();   \
    prof_acc[idx] += _now - prof_prev;  \
    prof_prev = _now;                   \
  } while (0)
#else
#define AFB_PROF(idx) do { } while (0)
#endif

#ifndef AFB_ATTN_STAGGER
#define AFB_ATTN_STAGGER 0
#endif

// P through tensor memory: the softmax warps write P_j (bf16) over the first 32 columns of their own S_j with
// tcgen05.st and O += P_j V_j takes its A operand from TMEM.  The shared-memory port then carries only Q/K/V operand
// reads (it was ~75 % busy with the P round trip: 64 KB of P written and 64 KB re-read by the MMAs per 64 keys on top of
// 88 KB of Q/K/V reads).  S_{j+1} can no longer be issued before P_j has been consumed: it follows O += P_j V_j in
// the (in-order) tensor pipe.  AFB_ATTN_PTMEM=0 builds the shared-memory P path.
#ifndef AFB_ATTN_PTMEM
#define AFB_ATTN_PTMEM 1
#endif

// Row sums on the tensor pipe: column D of every V tile is overwritten with ones (by the otherwise idle lanes of the
// MMA warp, once the tile has landed), so O's column D accumulates sum_j P_j -- of the bf16-rounded probabilities, the
// same values the numerator uses -- and the softmax warps drop one FADD2 per pair of scores (10 % of their
// instructions; d = 40 only: 560 -> 546 us).  AFB_ATTN_ONES=0 keeps the row sum in registers.
#ifndef AFB_ATTN_ONES
#define AFB_ATTN_ONES 1
#endif
// AFB_ATTN_SDOUBLE: 0 = one S buffer per query tile everywhere, 1 = two at d = 80 (default), 2 = also at d = 40.
#ifndef AFB_ATTN_SDOUBLE
#define AFB_ATTN_SDOUBLE 1
#endif
#ifndef AFB_ATTN_HALF
#define AFB_ATTN_HALF 0
#endif
// Column-split softmax (d = 40, long key sequences), OFF: TWO warps per TMEM lane quarter and query tile, each taking 32
// of the round's 64 score columns of the same 32 rows (the row max crosses between them through shared memory and a
// 64-thread named barrier), with three query tiles per CTA: 24 softmax warps = six per SM sub-partition instead of
// four, each with half the serial chain (load -> max -> exponentials -> store -> MMA round trip).  The idea came from
// ncu on the four-warp version - issue slots 60 %, XU 63 %, top stall long_scoreboard (TMEM loads, barrier waits) -
// read as latency that more warps would hide.  MEASURED SLOWER (B200, 4096 x 4096, batch 16, in the step): 703 us with
// one S buffer (-DAFB_ATTN_COLSPLIT=1), 715 us with two (=2, 3 x (128 + 40) TMEM columns, row sums back in registers)
// against 556 us: per query tile a 64-key round takes ~670 instead of ~600 cycles.  Every softmax warp pays a fixed
// per-round cost (barrier waits, fences, the publish, loop control, and now the max exchange) that the split doubles
// while the exponentials stay the same, i.e. the kernel is bound by the instruction stream the softmax warps issue, not
// by exposed latency: fewer instructions per score is the only lever (finding 5 of DESIGN.md section 4).
#ifndef AFB_ATTN_COLSPLIT
#define AFB_ATTN_COLSPLIT 0
#endif

// kSmall: the variant for a short key sequence (the 77-token context, the 8x8 image's self-attention): such a CTA is
// one serial chain (set-up, loads, two score tiles, epilogue) with nothing to pipeline, so instead of four query tiles
// behind one chain it takes ONE query tile, 192 threads and as many TMEM columns as S and O need, and several CTAs
// share the SM and overlap their chains.
template <int D, bool kSmall = false>
struct AttnCfg {
  static constexpr bool kPTmem = AFB_ATTN_PTMEM != 0;
  // kSDouble: two S buffers per query tile, so S_{j+1} = Q K_{j+1}^T is computed while the softmax warps still work on
  // S_j: the MMA round trip (publish P_j -> MMA warp reacts -> tensor pipe -> commit -> softmax warp wakes, ~800 cycles
  // plus queueing behind the other tiles' MMAs) leaves the critical path.  P_j overwrites the first half of ITS S
  // buffer; S_{j+2} is issued into that buffer behind O += P_j V_j.  d = 80 fits as it is (2 x (128 + 80) = 416 TMEM
  // columns): 82.2 -> 74.5 us at 1024 x 1024.  d = 40 would pay with one query tile (3 instead of 4: 3 x (128 + 40) =
  // 504 columns, no room for the ones column) and measured SLOWER, 563 -> 604 us at 4096 x 4096: per query tile the same
  // ~560 cycles per 64 keys with or without the round trip (three warps per sub-partition at ~60 % MUFU and ~63 % issue
  // occupancy, i.e. bound by the exponentials' instruction stream, not by the MMA latency), plus an 8 % tail from
  // 11 CTAs of 384 rows per head (1408 CTAs = 9.5 waves).  AFB_ATTN_SDOUBLE=2 builds it for d = 40 as well.
  // kHalf (d = 40): the two S buffers are the two 32-key HALVES of the tile's 64 score columns, so run-ahead costs no
  // tensor memory: a round is 32 keys; while the softmax warps work on half h, the tensor pipe runs O += P V and the next
  // scores of half 1 - h.  All four query tiles and the ones column stay (4 x (64 + 48) = 448 columns).
  // MEASURED SLOWER (B200, 4096 x 4096, batch 16): 759 us against 560 us for the single-buffered 64-key rounds.  A
  // tcgen05.mma of 128 x N x 16 costs ~30 clocks of the tensor pipe's issue time whatever N <= 64 is, so ten
  // instructions per 64 keys (4 for P V, 2 x 3 for the two 32-key score halves) instead of seven take 1258 instead of 812
  // clocks per 64 keys of the four tiles, and the softmax warps wait for the pipe (2100 instead of 800 clocks per 64
  // keys) instead of for the round trip.  Off by default; -DAFB_ATTN_HALF=1 builds it.
  static constexpr bool kHalf = AFB_ATTN_HALF != 0 && kPTmem && !kSmall && D <= 40 && AFB_ATTN_SDOUBLE < 2;
  // (AFB_ATTN_COLSPLIT: 1 = column split over one S buffer, 2 = column split over two S buffers)
  static constexpr bool kColSplit = AFB_ATTN_COLSPLIT != 0 && kPTmem && !kSmall && D <= 40 && AFB_ATTN_HALF == 0;
  static constexpr bool kSDouble = kHalf || (AFB_ATTN_SDOUBLE != 0 && kPTmem && !kSmall && D <= 80 &&
                                             (D > 40 || AFB_ATTN_SDOUBLE >= 2 || (kColSplit && AFB_ATTN_COLSPLIT >= 2)));
  static constexpr int kSKeys = kHalf ? 32 : kBlockKV;          // keys per S buffer = keys per softmax round
  static constexpr int kQTiles = kSmall ? 1 : ((D <= 40) ? (((kSDouble && !kHalf) || kColSplit) ? 3 : 4) : 2);   // query tiles per CTA
  static constexpr int kWarpsPerTile = kColSplit ? 8 : 4;        // softmax warps of one query tile
  static constexpr int kThreads = 64 + kQTiles * kWarpsPerTile * 32;
  static constexpr int kCols = kColSplit ? kSKeys / 2 : kSKeys;  // score columns a softmax thread holds per round
  static constexpr int kAtoms = (D + 63) / 64;                 // 64-column (128-byte) swizzle atoms per row
  static constexpr int kDPadK = (D + 15) / 16 * 16;            // K extent of S = Q K^T (UMMA K step 16)
  static constexpr bool kOnes = AFB_ATTN_ONES != 0 && D <= 40 && (!kSDouble || kHalf);   // d = 80 is tensor-bound: the wider P V costs more there
  static_assert(D % 8 == 0, "head dim");
  static constexpr int kDPadN = kOnes ? D + 8 : D;             // N extent of O += P V (UMMA N step 8); column D = row sum
  static constexpr int kQTileBytes = kAtoms * kBlockQ * 128;   // one query tile: atom a at a * 16 KB
  static constexpr int kKVTileBytes = kAtoms * kBlockKV * 128;  // one of K or V
  static constexpr int kPBytes = kBlockQ * 128;                 // one query tile's P (64 keys = one atom)
  // (column split: three query tiles leave room for a fifth K/V stage, and keep the CTA above half the SM's shared memory)
  static constexpr int kStages = kSmall ? 2 : ((D <= 40) ? (kColSplit ? 5 : 4) : (D <= 80 ? 3 : 2));
#ifndef AFB_ATTN_POLY_EVERY
#define AFB_ATTN_POLY_EVERY 3
#endif
// d = 80: no FMA-pipe exponentials (measured 74.7 us without, 75.3 / 76.3 us with one pair in three / four: the
// kernel is latency-, not MUFU-bound there)
#ifndef AFB_ATTN_POLY_EVERY_80
#define AFB_ATTN_POLY_EVERY_80 0
#endif
  static constexpr int kPolyEvery = (D <= 40) ? AFB_ATTN_POLY_EVERY : (D <= 80 ? AFB_ATTN_POLY_EVERY_80 : 0);          // every kPolyEvery-th pair of exponentials on the FMA pipe
  static constexpr int kSmemBytes =
      kQTiles * (kQTileBytes + ((kSmall && kPTmem) ? 0 : kPBytes)) + kStages * 2 * kKVTileBytes;
  static constexpr int kSStride = kSDouble ? 2 * kSKeys : kBlockKV;       // TMEM columns of one tile's S buffer(s)
  static constexpr int kTmemNeed = kQTiles * (kSStride + kDPadN);
  // the full-size CTA owns all of TMEM (one per SM); a small one takes the power of two that holds its S and O
  static constexpr int kTmemCols = !kSmall ? 512 : (kTmemNeed <= 128 ? 128 : (kTmemNeed <= 256 ? 256 : 512));
  static constexpr int kMinBlocks = !kSmall ? 1 : (D <= 40 ? 4 : (D <= 80 ? 2 : 1));
  static constexpr int kTmemS = 0;                               // S_t: kSStride columns each (buffer b at + 64 b)
  static constexpr int kTmemO = kQTiles * kSStride;              // O_t: kDPadN columns each
  static_assert(kTmemNeed <= kTmemCols, "TMEM budget");
  static_assert(kSmemBytes <= 226 * 1024, "shared memory budget");
  static_assert(kSmall || kSmemBytes > 114 * 1024, "a second resident CTA would only wait for the 512 TMEM columns");
  static_assert(kMinBlocks * kTmemCols <= 512 && kMinBlocks * (kSmemBytes + 2048) <= 227 * 1024, "co-residency budget");
};

// 2^x for a packed pair on the FMA / ALU pipes instead of MUFU (x <= ~8; anything below -126 flushes towards 2^-126):
// round-to-nearest split x = n + f, |f| <= 0.5, via the 1.5 * 2^23 magic constant; 2^f by a degree-3 minimax
// polynomial (relative error 7.5e-5, far below the bf16 rounding of P); 2^n by adding n to the exponent field.
// The d = 40 kernel is MUFU-bound (one ex2 per 80 tensor FLOPs), so a fixed share of the exponentials goes this way.
__device__ __forceinline__ uint64_t exp2_poly_x2(uint64_t x2) {
  float lo, hi;
  unpack_f32x2(x2, lo, hi);
  x2 = pack_f32x2(fmaxf(lo, -126.0f), fmaxf(hi, -126.0f));
  const uint64_t magic = pack_f32x2(12582912.0f, 12582912.0f);
  const uint64_t t2 = fadd2(x2, magic);
  const uint64_t n2 = fadd2(t2, pack_f32x2(-12582912.0f, -12582912.0f));
  const uint64_t f2 = ffma2(n2, pack_f32x2(-1.0f, -1.0f), x2);
  uint64_t p2 = ffma2(f2, pack_f32x2(0.0551716648f, 0.0551716648f), pack_f32x2(0.2426111251f, 0.2426111251f));
  p2 = ffma2(p2, f2, pack_f32x2(0.6932609677f, 0.6932609677f));
  p2 = ffma2(p2, f2, pack_f32x2(0.9999280572f, 0.9999280572f));
  float tl, th, pl, ph;
  unpack_f32x2(t2, tl, th);
  unpack_f32x2(p2, pl, ph);
  const float rl = __int_as_float(__float_as_int(pl) + (__float_as_int(tl) << 23));
  const float rh = __int_as_float(__float_as_int(ph) + (__float_as_int(th) << 23));
  return pack_f32x2(rl, rh);
}

// 576 threads = 18 warps put five warps on two of the SM sub-partitions (16 K registers each): 96 registers per thread
// is the hardware ceiling for the d = 40 configuration, which ptxas derives from the launch bounds
template <int D, bool kHasMask, bool kSmall>
__global__ void __launch_bounds__(AttnCfg<D, kSmall>::kThreads, AttnCfg<D, kSmall>::kMinBlocks)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                     const __grid_constant__ CUtensorMap tm_v, const AttnDev p) {
  using C = AttnCfg<D, kSmall>;
  constexpr int BKV = kBlockKV;
  constexpr int NS = C::kStages;
  constexpr int NT = C::kQTiles;
  extern __shared__ __align__(1024) uint8_t smem[];     // SWIZZLE_128B atoms need 1024-byte alignment
  __shared__ __align__(8) uint64_t bars[2 + 6 * NT + 4 * NS + 1];
  __shared__ float s_mx[C::kColSplit ? 2 * NT * 2 * kBlockQ : 1];      // column split: partial row maxima [parity][tile][half][row]
  __shared__ float s_lsum[(C::kColSplit && !C::kOnes) ? NT * 2 * kBlockQ : 1];   // ... and partial row sums (epilogue)
  static_assert(!C::kColSplit || C::kPTmem, "column split: P through tensor memory");
  uint8_t* sQ = smem;                                   // [NT] query tiles
  uint8_t* sK = sQ + NT * C::kQTileBytes;               // [NS] K tiles
  uint8_t* sV = sK + NS * C::kKVTileBytes;              // [NS] V tiles
  uint8_t* sP = sV + NS * C::kKVTileBytes;              // [NT] P tiles
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint64_t* q_full = bars + 0;
  uint64_t* q_ready = bars + 1;
  uint64_t* s_full = bars + 2;                           // [2t + buf] S_j of tile t is in TMEM (buf = j & 1 if kSDouble, else 0)
  uint64_t* s_free = s_full + 2 * NT;                    // [t] S_j is in registers: the S columns may be overwritten
  uint64_t* p_ready = s_free + NT;                       // [2t + buf] P_j is published (and O rescaled if needed)
  uint64_t* pv_done = p_ready + 2 * NT;                  // [t] O += P_j V_j finished: P buffer and O are free
  uint64_t* k_full = pv_done + NT;
  uint64_t* k_empty = k_full + NS;
  uint64_t* v_full = k_empty + NS;
  uint64_t* v_empty = v_full + NS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(v_empty + NS);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q_block = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
  const int q0 = q_block * (NT * kBlockQ);
  const int num_kv_tiles = (p.kv_len + BKV - 1) / BKV;
  const int col0 = head * D;                             // first column of this head in the row
  const int n_tiles = min(NT, (p.q_len - q0 + kBlockQ - 1) / kBlockQ);   // query tiles with at least one valid row

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    mbar_init(q_full, 1);
    mbar_init(q_ready, C::kWarpsPerTile * n_tiles);
    for (int t = 0; t < NT; ++t) {
      mbar_init(&s_full[2 * t], 1);
      mbar_init(&s_full[2 * t + 1], 1);
      mbar_init(&s_free[t], C::kSDouble ? 1 : 4);
      mbar_init(&p_ready[2 * t], C::kWarpsPerTile);
      mbar_init(&p_ready[2 * t + 1], C::kWarpsPerTile);
      mbar_init(&pv_done[t], 1);
    }
    for (int i = 0; i < 4 * NS; ++i) mbar_init(&k_full[i], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<C::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();        // Q/K/V come from the previous kernel; the set-up above overlapped its tail
  pdl_trigger();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, n_tiles * C::kQTileBytes);
      for (int t = 0; t < n_tiles; ++t)
#pragma unroll
        for (int a = 0; a < C::kAtoms; ++a)     // rows past q_len are zero-filled by TMA
          tma_load_3d(&tm_q, q_full, sQ + t * C::kQTileBytes + a * (kBlockQ * 128), col0 + a * 64,
                      q0 + t * kBlockQ, b, kEvictFirst);
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < num_kv_tiles; ++j) {
        // K and V travel in separate rings: K_j is released as soon as S_j exists (the score MMA runs one tile
        // ahead of the softmax), V_j only after O += P_j V_j.
        mbar_wait(&k_empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&k_full[stage], C::kKVTileBytes);
#pragma unroll
        for (int a = 0; a < C::kAtoms; ++a)
          tma_load_3d(&tm_k, &k_full[stage], sK + stage * C::kKVTileBytes + a * BKV * 128, col0 + a * 64, j * BKV, b,
                      kEvictLast);
        mbar_wait(&v_empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&v_full[stage], C::kKVTileBytes);
#pragma unroll
        for (int a = 0; a < C::kAtoms; ++a)
          tma_load_3d(&tm_v, &v_full[stage], sV + stage * C::kKVTileBytes + a * BKV * 128, col0 + a * 64, j * BKV, b,
                      kEvictLast);
        if (++stage == NS) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // The whole warp runs this loop converged (barrier polls are made warp-uniform); one elected lane issues the
    // tcgen05 instructions.  Converged control flow lets the descriptors live in uniform registers and avoids the
    // per-instruction election loop the compiler emits around tcgen05.mma in divergent code.
    {
      constexpr int BS = C::kSKeys;                         // keys per round
      constexpr int H = BKV / BS;                           // rounds per K/V tile
      constexpr uint32_t idesc_qk = umma_idesc_bf16(kBlockQ, BS, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(kBlockQ, C::kDPadN, 0, 1);   // B (= V) is MN-major
      const uint64_t q_desc0 = umma_desc_kmajor_sw128(smem_u32(sQ));
      const uint32_t p_addr = smem_u32(sP);
      const bool leader = elect_one();
      // S_t = Q_t K_j^T : K-steps of 16 over the (padded) head dim
      auto issue_qk = [&](int t, int stage, int buf = 0, int half = 0) {
        if (leader) {
          // one descriptor per operand tile; the K-step offsets are added to its 16-byte address field
          // (half: keys [half * BS, half * BS + BS) of the tile = rows of the K-major K tile, 128 bytes each)
          const uint64_t kd = umma_desc_kmajor_sw128(smem_u32(sK + stage * C::kKVTileBytes)) +
                              static_cast<uint64_t>((half * BS * 128) >> 4);
          const uint64_t qd = q_desc0 + static_cast<uint64_t>(t * (C::kQTileBytes >> 4));
          const uint32_t tmem_s = tmem_base + C::kTmemS + t * C::kSStride + buf * BS;
#pragma unroll
          for (int kk = 0; kk < C::kDPadK / 16; ++kk) {
            const uint64_t qoff = static_cast<uint64_t>(((kk >> 2) * (kBlockQ * 128) + (kk & 3) * 32) >> 4);
            const uint64_t koff = static_cast<uint64_t>(((kk >> 2) * (BKV * 128) + (kk & 3) * 32) >> 4);
            umma_ss(tmem_s, qd + qoff, kd + koff, idesc_qk, kk != 0);
          }
          umma_commit(&s_full[2 * t + buf]);
        }
      };
      // O_t += P_t V_j : K-steps of 16 keys; V tile is [key][64 cols] per atom -> MN-major, LBO = atom stride
      auto issue_pv = [&](int t, int stage, bool first, bool last, int buf = 0, int half = 0) {
        if (leader) {
          const uint64_t vd = umma_desc_mnmajor_sw128(smem_u32(sV + stage * C::kKVTileBytes), BKV * 128);
          const uint32_t tmem_o = tmem_base + C::kTmemO + t * C::kDPadN;
#pragma unroll
          for (int kk = 0; kk < BS / 16; ++kk) {
            const uint64_t voff = static_cast<uint64_t>(((half * (BS / 16) + kk) * (16 * 128)) >> 4);
            if constexpr (C::kPTmem) {
              // A = P_t from TMEM: 16 keys = 8 columns of packed bf16 pairs, same lanes as the S rows
              umma_ts(tmem_o, tmem_base + C::kTmemS + t * C::kSStride + buf * BS + kk * 8, vd + voff, idesc_pv,
                      !(first && kk == 0));
            } else {
              const uint32_t pa = p_addr + t * C::kPBytes + kk * 32;
              umma_ss(tmem_o, umma_desc_kmajor_sw128(pa), vd + voff, idesc_pv, !(first && kk == 0));
            }
          }
          // double-buffered S: pv_done[t] completes every round (the rare O rescale waits for round j-1 on it), and
          // the epilogue gets its own one-shot barrier (s_free[t], otherwise unused with P in TMEM): a softmax warp can
          // finish P(n-1) before round n-2 has completed, and a parity wait cannot tell two phases apart
          if (!C::kPTmem || C::kSDouble || last) umma_commit(&pv_done[t]);
          if (C::kSDouble && last) umma_commit(&s_free[t]);
        }
      };
#ifdef AFB_ATTN_PROFILE
      long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      long long prof_prev = clock64();
#endif
      mbar_wait(q_full, 0);
      mbar_wait(q_ready, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      for (int t = 0; t < n_tiles; ++t) issue_qk(t, 0);
      if (H == 1 && leader) umma_commit(&k_empty[0]);      // half rounds: K tile 0 also feeds S(1), released below
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t all = (1u << n_tiles) - 1u;
      if constexpr (C::kSDouble) {
        // Run-ahead schedule over ROUNDS of BS keys (H rounds per K/V tile): S(0) and S(1) of every tile up front; in
        // round r, when tile t publishes P_t(r) (in S buffer r & 1): O_t += P_t(r) V(r), then S_t(r+2) into that same
        // buffer (the tensor pipe keeps issue order, so the overwrite lands behind the read).  The softmax of round r+1
        // never waits for an MMA.
        static_assert(NS >= 3, "K ring: the tiles of rounds r+1 and r+2 are live while the next one loads");
        const int num_rounds = (p.kv_len + BS - 1) / BS;
        if (num_rounds > 1) {
          if (H == 1) {
            mbar_wait(&k_full[1], 0);
            tc_fence_after();
          }
          for (int t = 0; t < n_tiles; ++t) issue_qk(t, H == 1 ? 1 : 0, 1, H == 1 ? 0 : 1);
          if (H == 1 && leader) umma_commit(&k_empty[1]);
        }
        if (H == 2 && leader) umma_commit(&k_empty[0]);      // (H == 1: committed above, after S(0))
        for (int r = 0; r < num_rounds; ++r) {
          const int j = r / H, h = r % H, st = j % NS;
          const int r2 = r + 2, j2 = r2 / H, h2 = r2 % H, st2 = j2 % NS;
          const bool more = r2 < num_rounds;
          AFB_PROF(3);
          if (more && h2 == 0) mbar_wait(&k_full[st2], (j2 / NS) & 1);      // first round that reads K tile j2
          if (h == 0) {
            mbar_wait(&v_full[st], (j / NS) & 1);
            if constexpr (C::kOnes) {
              // ones into column D of the V tile (atom D / 64, 16-byte chunk (D % 64) / 8 at its swizzled position)
              uint8_t* vt = sV + st * C::kKVTileBytes + (D / 64) * (BKV * 128);
#pragma unroll
              for (int k = lane; k < BKV; k += 32)
                *reinterpret_cast<uint16_t*>(vt + k * 128 + (((((D % 64) / 8) ^ (k & 7))) << 4)) = 0x3F80;   // bf16 1.0
              fence_proxy_async_smem();
              __syncwarp();
            }
          }
          AFB_PROF(0);
          const int buf = r & 1;
          const uint32_t par = (r >> 1) & 1;
          uint32_t pending = all;
          while (pending != 0u) {
#pragma unroll
            for (int t = 0; t < NT; ++t) {
              if ((pending >> t) & 1u) {
                if (__all_sync(0xffffffffu, mbar_test_wait(&p_ready[2 * t + buf], par))) {
                  AFB_PROF(1);
                  tc_fence_after();
                  issue_pv(t, st, r == 0, r + 1 == num_rounds, buf, h);
                  if (more) issue_qk(t, st2, buf, h2);
                  pending &= ~(1u << t);
                  AFB_PROF(2);
                }
              }
            }
          }
          if (leader) {
            if (h == H - 1 || r + 1 == num_rounds) umma_commit(&v_empty[st]);                    // last round on V tile j
            if (more && (h2 == H - 1 || r2 + 1 == num_rounds)) umma_commit(&k_empty[st2]);      // last scores from K tile j2
          }
        }
      } else
      for (int j = 0; j < num_kv_tiles; ++j) {
        int nstage = stage + 1;
        uint32_t nphase = phase;
        if (nstage == NS) {
          nstage = 0;
          nphase ^= 1;
        }
        const bool more = j + 1 < num_kv_tiles;
        AFB_PROF(3);
        if (more) mbar_wait(&k_full[nstage], nphase);
        mbar_wait(&v_full[stage], phase);
        if constexpr (C::kOnes) {
          // ones into column D of the V tile (atom D / 64, 16-byte chunk (D % 64) / 8 at its swizzled position)
          uint8_t* vt = sV + stage * C::kKVTileBytes + (D / 64) * (BKV * 128);
#pragma unroll
          for (int k = lane; k < BKV; k += 32)
            *reinterpret_cast<uint16_t*>(vt + k * 128 + (((((D % 64) / 8) ^ (k & 7))) << 4)) = 0x3F80;   // bf16 1.0
          fence_proxy_async_smem();
          __syncwarp();
        }
        AFB_PROF(0);                                      // wait for K/V tiles
        // events of this round: bit t = "S_t(j+1) may be issued" (scores j are in registers),
        //                       bit NT + t = "O_t += P_t(j) V_j may be issued" (P_t(j) is published)
        uint32_t pending = (more && !C::kPTmem ? all : 0u) | (all << NT);
        const uint32_t par = j & 1;
        while (pending != 0u) {
#pragma unroll
          for (int t = 0; t < NT; ++t) {
            if constexpr (C::kPTmem) {
              // one event per tile and round: P_t(j) sits in TMEM -> O_t += P_t(j) V_j, then S_t(j+1) = Q_t K_{j+1}^T
              // over the same columns (the tensor pipe runs them in issue order)
              if ((pending >> (NT + t)) & 1u) {
                if (__all_sync(0xffffffffu, mbar_test_wait(&p_ready[2 * t], par))) {
                  AFB_PROF(1);
                  tc_fence_after();
                  issue_pv(t, stage, j == 0, !more);
                  if (more) issue_qk(t, nstage);
                  pending &= ~(1u << (NT + t));
                  AFB_PROF(2);
                }
              }
              continue;
            }
            if ((pending >> t) & 1u) {
              if (__all_sync(0xffffffffu, mbar_test_wait(&s_free[t], par))) {
                AFB_PROF(1);                              // polling
                tc_fence_after();
                issue_qk(t, nstage);
                pending &= ~(1u << t);
                if ((pending & all) == 0u && leader) umma_commit(&k_empty[nstage]);
                AFB_PROF(2);                              // issuing
              }
            }
            if ((pending >> (NT + t)) & 1u) {
              if (__all_sync(0xffffffffu, mbar_test_wait(&p_ready[2 * t], par))) {
                AFB_PROF(1);
                tc_fence_after();
                issue_pv(t, stage, j == 0, !more);
                pending &= ~(1u << (NT + t));
                AFB_PROF(2);
              }
            }
          }
        }
        if (leader) {
          umma_commit(&v_empty[stage]);
          if (C::kPTmem && more) umma_commit(&k_empty[nstage]);
        }
        stage = nstage;
        phase = nphase;
      }
#ifdef AFB_ATTN_PROFILE
      {
        const int blk = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
        if (lane == 0 && blk < p.prof_blocks)
          for (int i = 0; i < 8; ++i) p.prof[(blk * 17 + 16) * 8 + i] = prof_acc[i];
      }
#endif
    }
  } else if (warp < 2 + C::kWarpsPerTile * n_tiles) {
    // ------------------------------------------------------------------ softmax / correction / epilogue
    const int t = (warp - 2) / C::kWarpsPerTile;          // query tile
    const int quarter = warp & 3;                         // TMEM lane quarter this warp may access (= its SMSP)
    // column split: the tile's warps 0-3 take score columns 0-31 of a round, warps 4-7 columns 32-63 (8 t = 0 mod 4, so
    // both halves see every lane quarter)
    const int ch = C::kColSplit ? (((warp - 2) % C::kWarpsPerTile) >> 2) : 0;
    constexpr int CN = C::kCols;                          // score columns per thread and round
    const int row = quarter * 32 + lane;                  // query row inside the tile
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t tmem_s0 = tmem_base + lane_addr + C::kTmemS + t * C::kSStride;
    const uint32_t tmem_o = tmem_base + lane_addr + C::kTmemO + t * C::kDPadN;
    uint8_t* prow = sP + t * C::kPBytes + row * 128;
#ifdef AFB_ATTN_PROFILE
    long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long prof_prev = clock64();
#endif

    mbar_wait(q_full, 0);
    if constexpr (D % 16 != 0) {
      // head_dim 40: K-step 2 covers columns 32..47; zero this row's columns 40..47 of Q (one 16-byte chunk,
      // swizzled position) so that whatever K holds there (the next head's data) contributes nothing.
      constexpr int chunk = D / 8;                        // 16-byte chunk index inside the 128-byte row
      uint8_t* dst = sQ + t * C::kQTileBytes + (C::kAtoms - 1) * (kBlockQ * 128) + row * 128 +
                     (((chunk & 7) ^ (row & 7)) << 4);
      if (ch == 0) *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
      fence_proxy_async_smem();
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(q_ready);

    // Scores stay raw (unscaled) in registers; the softmax scale (times log2 e) is folded into the one FFMA that
    // also subtracts the reference max:  p = 2^(s * scale_log2 - m_used).
    float m_used = 0.f;      // max (scaled log2 domain) the exponentials are currently referenced to
    float l_run = 0.f;       // running row sum
    const float sc = p.scale_log2;
    const uint64_t sc2 = pack_f32x2(sc, sc);
    const uint8_t* mask_row = kHasMask ? p.key_mask + static_cast<long long>(b) * p.kv_len : nullptr;

#if AFB_ATTN_STAGGER > 0
    // De-phase the query tiles: the four softmax warps of an SM sub-partition (one per tile) would otherwise walk
    // the KV loop in step -- all in the exponential phase (MUFU saturated) or all outside it (MUFU idle).
    if (t > 0) {
      const long long t0 = clock64();
      while (clock64() - t0 < static_cast<long long>(t) * AFB_ATTN_STAGGER) {}
    }
#endif
    constexpr int BS = C::kSKeys;                         // keys per round (32: half rounds, else 64)
    const int num_rounds = (p.kv_len + BS - 1) / BS;
    for (int j = 0; j < num_rounds; ++j) {                // j counts ROUNDS of BS keys
      AFB_PROF(7);
      const int sbuf = C::kSDouble ? (j & 1) : 0;
      const uint32_t tmem_s = tmem_s0 + sbuf * BS;
      mbar_wait(&s_full[2 * t + sbuf], C::kSDouble ? ((j >> 1) & 1) : (j & 1));
      tc_fence_after();
      AFB_PROF(0);                                        // wait for S
      uint32_t r[CN];
#pragma unroll
      for (int c = 0; c < CN / 32; ++c) tmem_ld32p(tmem_s + ch * CN + c * 32, &r[c * 32]);
      tmem_ld_wait();
      if constexpr (!C::kPTmem) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_free[t]);           // the MMA warp may start S_{j+1}
      }
      AFB_PROF(1);                                        // TMEM load

      const int kv0 = j * BS + ch * CN;                   // first key of this thread's columns
      if (kHasMask) {
        // masked keys get a large finite negative IN THE SCALED DOMAIN (like the reference's masked_fill after the
        // scaling): a fully masked row then cancels exactly to exp2(0) and averages uniformly
#pragma unroll
        for (int i = 0; i < CN; ++i) {
          float v = __uint_as_float(r[i]) * sc;
          if (kv0 + i < p.kv_len && mask_row[kv0 + i] == 0) v = kMaskedScore;
          r[i] = __float_as_uint(v);
        }
      }
      if (kv0 + CN > p.kv_len) {
#pragma unroll
        for (int i = 0; i < CN; ++i)
          if (kv0 + i >= p.kv_len) r[i] = 0xff800000u;    // -inf
      }
      float mx[4] = {__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3])};
#pragma unroll
      for (int i = 4; i < CN; i += 8) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          if (i + 2 * c + 1 < CN)
            mx[c] = fmax3(mx[c], __uint_as_float(r[i + 2 * c]), __uint_as_float(r[i + 2 * c + 1]));
      }
      float m_raw = fmaxf(fmax3(mx[0], mx[1], mx[2]), mx[3]);
      if constexpr (C::kColSplit) {
        // the row's other 32 columns belong to the partner warp (same lane quarter, other column half): swap partial
        // maxima through shared memory (double-buffered by round parity) across a 64-thread named barrier.  The barrier
        // also orders the partner's tcgen05.ld of ITS score columns before this warp's P store, which overwrites the
        // first 32 columns of the S buffer.
        float* slot = s_mx + (((j & 1) * NT + t) * 2) * kBlockQ + row;
        slot[ch * kBlockQ] = m_raw;
        tc_fence_before();
        asm volatile("bar.sync %0, 64;" ::"r"(1 + t * 4 + quarter) : "memory");
        tc_fence_after();
        m_raw = fmaxf(m_raw, slot[(ch ^ 1) * kBlockQ]);
      }
      const float m_tile = kHasMask ? m_raw : m_raw * sc;   // scale > 0

      const bool need = (j == 0) || (m_tile > m_used + kRescaleThreshold);
      if (j == 0) {
        m_used = m_tile;
      } else if (__any_sync(0xffffffffu, need)) {
        // rare after the first tiles: O must be rescaled, which needs O += P_{j-1} V_{j-1} to have finished
        // (P through TMEM: S_j was issued behind it and has completed, so it has)
        // (double-buffered S: S_j sits behind O += P_{j-2} V_{j-2} only, so round j-1 is waited for explicitly)
        if constexpr (!C::kPTmem || C::kSDouble) {
          mbar_wait(&pv_done[t], (j - 1) & 1);
          tc_fence_after();
        }
        const float m_new = need ? m_tile : m_used;
        const float alpha = ex2_approx(m_used - m_new);  // 1 for rows that keep their reference max
        m_used = m_new;
        l_run *= alpha;
        // (column split: the two warps of a row rescale half of its O columns each)
        constexpr int kOChunks = C::kDPadN / 8, kOHalf = (kOChunks + 1) / 2;
#pragma unroll
        for (int c = 0; c < kOChunks; ++c) {
          if (C::kColSplit && (c < kOHalf) != (ch == 0)) continue;
          uint32_t o[8];
          tmem_ld8(tmem_o + c * 8, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st8(tmem_o + c * 8, o);
        }
        tmem_st_wait();
      }
      AFB_PROF(2);                                        // max (+ rare rescale)

      // exponentials: FFMA2 (scale, subtract max) -> 2 x MUFU.EX2 -> FADD2 (row sum) -> F2FP (bf16x2), in place
      const uint64_t nm2 = pack_f32x2(-m_used, -m_used);
      uint64_t acc0 = 0ull, acc1 = 0ull;                 // two packed partial sums
#pragma unroll
      for (int i = 0; i < CN / 2; ++i) {
        uint64_t x = pack_f32x2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
        x = kHasMask ? fadd2(x, nm2) : ffma2(x, sc2, nm2);
        float lo, hi;
        uint64_t e;
        if (C::kPolyEvery > 0 && (i % (C::kPolyEvery > 0 ? C::kPolyEvery : 1)) == 1) {
          e = exp2_poly_x2(x);                            // FMA-pipe exponentials: relieves MUFU
          unpack_f32x2(e, lo, hi);
        } else {
          unpack_f32x2(x, lo, hi);
          lo = ex2_approx(lo);
          hi = ex2_approx(hi);
          e = pack_f32x2(lo, hi);
        }
        if constexpr (!C::kOnes) {
          if (i & 1) acc1 = fadd2(acc1, e); else acc0 = fadd2(acc0, e);
        }
        r[i] = cvt_bf16x2(lo, hi);
      }
      if constexpr (!C::kOnes) {
        float a, bb;
        unpack_f32x2(fadd2(acc0, acc1), a, bb);
        l_run += a + bb;
      }
      AFB_PROF(4);                                        // exponentials

      if constexpr (C::kPTmem) {
        // P_j (bf16 pairs, BS / 2 columns) over the first half of this round's S columns
        if constexpr (CN == 64) {
          tmem_st32p(tmem_s, r);
        } else {
          // 32 scores -> 16 columns of bf16 pairs (column split: the second warp's keys follow the first warp's)
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) pk[i] = r[i];
          tmem_st16(tmem_s + ch * (CN / 2), pk);
        }
        tmem_st_wait();
      } else {
        // the P buffer is free once O += P_{j-1} V_{j-1} has completed (long ago in steady state)
        if (j > 0) mbar_wait(&pv_done[t], (j - 1) & 1);
        AFB_PROF(5);                                        // wait for the P buffer
        // P -> canonical K-major SW128 layout: row pitch 128 B, 16-byte chunk c at c ^ (row & 7)
#pragma unroll
        for (int c = 0; c < 8; ++c)
          *reinterpret_cast<uint4*>(prow + ((c ^ (row & 7)) << 4)) =
              make_uint4(r[c * 4], r[c * 4 + 1], r[c * 4 + 2], r[c * 4 + 3]);
        fence_proxy_async_smem();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_ready[2 * t + sbuf]);
      AFB_PROF(6);                                        // P store
    }
#ifdef AFB_ATTN_PROFILE
    {
      const int blk = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
      if (lane == 0 && blk < p.prof_blocks)
        for (int i = 0; i < 8; ++i) p.prof[(blk * 17 + (warp - 2)) * 8 + i] = prof_acc[i];
    }
#endif

    // ---- epilogue: O / l -> bf16 -> out[b, q, head*D : head*D + D]
    if constexpr (C::kSDouble) mbar_wait(&s_free[t], 0);
    else mbar_wait(&pv_done[t], C::kPTmem ? 0 : ((num_kv_tiles - 1) & 1));
    tc_fence_after();
    const int q_idx = q0 + t * kBlockQ + row;
    if constexpr (C::kOnes) {
      uint32_t o[8];
      tmem_ld8(tmem_o + D, o);                              // column D of O: the row sum
      tmem_ld_wait();
      l_run = __uint_as_float(o[0]);
    }
    if constexpr (C::kColSplit && !C::kOnes) {
      // row sum in registers: each of the row's two warps holds the sum over its own score columns
      float* slot = s_lsum + (t * 2) * kBlockQ + row;
      slot[ch * kBlockQ] = l_run;
      asm volatile("bar.sync %0, 64;" ::"r"(1 + t * 4 + quarter) : "memory");
      l_run += slot[(ch ^ 1) * kBlockQ];
    }
    const float inv_l = 1.0f / l_run;
    if (p.lse != nullptr && q_idx < p.q_len && ch == 0)      // P = 2^(s * scale * log2(e) - lse)
      p.lse[(static_cast<long long>(b) * p.heads + head) * p.q_len + q_idx] = m_used + log2f(l_run);
    __nv_bfloat16* orow = p.out + (static_cast<long long>(b) * p.q_len + q_idx) * p.ldo + col0;
#pragma unroll
    for (int c = 0; c < D / 8; ++c) {
      if (C::kColSplit && (c < (D / 8 + 1) / 2) != (ch == 0)) continue;     // each of the row's two warps writes its half
      uint32_t o[8];
      tmem_ld8(tmem_o + c * 8, o);
      tmem_ld_wait();
      if (q_idx < p.q_len) {
        const uint4 pk = make_uint4(
            pack_bf16x2(__uint_as_float(o[0]) * inv_l, __uint_as_float(o[1]) * inv_l),
            pack_bf16x2(__uint_as_float(o[2]) * inv_l, __uint_as_float(o[3]) * inv_l),
            pack_bf16x2(__uint_as_float(o[4]) * inv_l, __uint_as_float(o[5]) * inv_l),
            pack_bf16x2(__uint_as_float(o[6]) * inv_l, __uint_as_float(o[7]) * inv_l));
        *reinterpret_cast<uint4*>(orow + c * 8) = pk;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<C::kTmemCols>(tmem_base);
}

#ifdef AFB_ATTN_PROFILE
long long* g_attn_prof = nullptr;
int g_attn_prof_blocks = 0;
#endif

int make_qkv_map(CUtensorMap* map, const void* ptr, int width, int len, int batch, long long ld, int box_rows) {
  const uint64_t dims[3] = {static_cast<uint64_t>(width), static_cast<uint64_t>(len), static_cast<uint64_t>(batch)};
  const uint64_t strides[2] = {static_cast<uint64_t>(ld) * 2, static_cast<uint64_t>(ld) * 2 * static_cast<uint64_t>(len)};
  const uint32_t box[3] = {64, static_cast<uint32_t>(box_rows), 1};
  return make_tensor_map(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, ptr, dims, strides, box, Swizzle::B128);
}

template <int D, bool kHasMask, bool kSmall>
int launch_attn_kernel(const CUtensorMap& tm_q, const CUtensorMap& tm_k, const CUtensorMap& tm_v, const AttnDev& p,
                       dim3 grid, cudaStream_t stream) {
  using C = AttnCfg<D, kSmall>;
  auto kern = attention_fwd_kernel<D, kHasMask, kSmall>;
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    attr_set = true;
  }
  AFB_CHECK_CUDA(launch_pdl(kern, grid, dim3(C::kThreads), C::kSmemBytes, stream, 1, tm_q, tm_k, tm_v, p));
  return 0;
}

// AFB_ATTN_SMALL=0 sends short key sequences through the full-size configuration too (A/B switch)
inline bool attn_small_enabled() {
  static const bool on = [] {
    const char* e = std::getenv("AFB_ATTN_SMALL");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

template <int D, bool kSmall>
int launch_attn(const afb_attn_args& a, cudaStream_t stream) {
  using C = AttnCfg<D, kSmall>;
  CUtensorMap tm_q, tm_k, tm_v;
  const int width = a.heads * a.head_dim;
  if (int rc = make_qkv_map(&tm_q, a.q, width, a.q_len, a.batch, a.ldq, kBlockQ)) return rc;
  if (int rc = make_qkv_map(&tm_k, a.k, width, a.kv_len, a.batch, a.ldk, kBlockKV)) return rc;
  if (int rc = make_qkv_map(&tm_v, a.v, width, a.kv_len, a.batch, a.ldv, kBlockKV)) return rc;
  AttnDev p;
  p.out = reinterpret_cast<__nv_bfloat16*>(a.out);
  p.key_mask = a.key_mask;
  p.lse = a.lse;
  p.q_len = a.q_len;
  p.kv_len = a.kv_len;
  p.heads = a.heads;
  p.ldo = a.ldo;
  p.scale_log2 = a.scale * 1.4426950408889634f;
#ifdef AFB_ATTN_PROFILE
  p.prof = g_attn_prof;
  p.prof_blocks = g_attn_prof_blocks;
#endif
  constexpr int rows_per_cta = C::kQTiles * kBlockQ;
  dim3 grid((a.q_len + rows_per_cta - 1) / rows_per_cta, a.heads, a.batch);
  const int rc = a.key_mask != nullptr ? launch_attn_kernel<D, true, kSmall>(tm_q, tm_k, tm_v, p, grid, stream)
                                       : launch_attn_kernel<D, false, kSmall>(tm_q, tm_k, tm_v, p, grid, stream);
  if (rc) return rc;
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace
}  // namespace afb

extern "C" int afb_attention_fwd(const afb_attn_args* args, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(args != nullptr, "afb_attention_fwd: args is NULL");
  const afb_attn_args& a = *args;
  AFB_REQUIRE(a.q && a.k && a.v && a.out, "afb_attention_fwd: NULL operand");
  AFB_REQUIRE(a.batch > 0 && a.heads > 0 && a.q_len > 0 && a.kv_len > 0, "afb_attention_fwd: empty problem");
  AFB_REQUIRE(a.batch <= 65535 && a.heads <= 65535, "afb_attention_fwd: batch/heads exceed the grid limits");
  const int width = a.heads * a.head_dim;
  AFB_REQUIRE(a.ldq >= width && a.ldk >= width && a.ldv >= width && a.ldo >= width,
              "afb_attention_fwd: leading dimension smaller than heads*head_dim");
  AFB_REQUIRE(a.ldq % 8 == 0 && a.ldk % 8 == 0 && a.ldv % 8 == 0 && a.ldo % 8 == 0,
              "afb_attention_fwd: leading dimensions must be multiples of 8 elements");
  const uintptr_t all = reinterpret_cast<uintptr_t>(a.q) | reinterpret_cast<uintptr_t>(a.k) |
                        reinterpret_cast<uintptr_t>(a.v) | reinterpret_cast<uintptr_t>(a.out);
  AFB_REQUIRE((all & 15) == 0, "afb_attention_fwd: q/k/v/out must be 16-byte aligned");
  const bool small = a.kv_len <= 2 * kBlockKV && attn_small_enabled();
  switch (a.head_dim) {
    case 40: return small ? launch_attn<40, true>(a, stream) : launch_attn<40, false>(a, stream);
    case 80: return small ? launch_attn<80, true>(a, stream) : launch_attn<80, false>(a, stream);
    // d = 160: one small CTA fills the SM's shared memory, so nothing overlaps and two query tiles per chain win
    // (measured 15.5 vs 17.3-18.8 us at 256 queries x 77 keys)
    case 160: return launch_attn<160, false>(a, stream);
    default:
      AFB_FAIL(2, "afb_attention_fwd: head_dim %d not supported (SD-1.5 uses 40, 80, 160)", a.head_dim);
  }
}
