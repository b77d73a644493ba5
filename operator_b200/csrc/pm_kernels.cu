// sm_100a kernels of the log-scan path.  Everything here is HBM-bound integer/byte work
// (north_star: "tensor cores are not used because this is not a contraction"), so the design
// rules are: read every log byte exactly once with 128-bit coalesced loads, keep the only
// random-access structure of the hot loop (the gram filter bitmap) in shared memory, one
// persistent CTA per SM, no block-wide synchronisation inside the streaming loop.
//
//   k_scan        a10 + a11 (hot): newline counts per 2 KiB tile and gram-filter candidates
//   k_prefix_*    a10: exclusive prefix of the tile counts (the two-level newline index)
//   k_pods        a10: per-pod line bases
//   k_confirm     a11: candidate -> exact gram -> literal -> verify DFA on the line -> raw event
//   k_dfa_scan    a11: combined DFA in shared memory for patterns without a usable literal
//   sort/unique   ordering + de-duplication of raw events (CUB radix sort, not a hot path)
//   k_plan / k_check / k_check_long / k_finalize   a12 + a13: line numbers, context ranges,
//                 secondary proximity, fp64 score
//   k_line_starts a10: materialised line table of one pod (ballot/prefix-scan scatter)
#include <atomic>
#include <cub/cub.cuh>

#include "pm_kernels.h"

namespace pm {

thread_local void (*g_stage_mark)(void *, const char *) = nullptr;
thread_local void *g_stage_mark_arg = nullptr;

namespace {

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int kQueue = 64;   // per-warp candidate staging entries

__device__ __forceinline__ uint4 ld_stream16(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// Streaming load that is also first in line for eviction from L2: the log text is read once,
// while the library tables and the small index arrays are re-read by every later kernel and
// should survive the 1 GiB stream in the 126 MB L2.
__device__ __forceinline__ unsigned long long l2_evict_first_policy() {
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint4 ld_stream16_ef(const uint4 *p, unsigned long long pol) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p), "l"(pol));
    return r;
}

// bit 7 of every byte of the result is set iff that byte of w is NOT '\n' (exact, carry-free)
__device__ __forceinline__ uint32_t not_newline_flags(uint32_t w) {
    uint32_t t = (w ^ 0x0A0A0A0Au) & 0x7F7F7F7Fu;
    return ((t + 0x7F7F7F7Fu) | w) & 0x80808080u;
}

__device__ __forceinline__ uint32_t newline_mask4(uint32_t w) {   // bit i set iff byte i == '\n'
    // flags at bit 0 of each byte, gathered into bits 21..24 by one multiply (no two partial
    // products share a bit position, so there are no carries)
    uint32_t x = (~not_newline_flags(w) >> 7) & 0x01010101u;
    return ((x * 0x00204081u) >> 21) & 0xFu;
}

__device__ __forceinline__ uint32_t newline_mask16(const uint4 &v) {
    return newline_mask4(v.x) | (newline_mask4(v.y) << 4) | (newline_mask4(v.z) << 8) | (newline_mask4(v.w) << 12);
}

__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// ---- warp time stamps (-DPM_TRACE builds only; tools_warp_trace.py): where the time of the small post-scan
// kernels goes.  Region r holds 8 stamps for each of the first kTraceWarps warps of one kernel.
#ifdef PM_TRACE
constexpr unsigned kTraceWarps = 16384, kTraceRegions = 4;
__device__ unsigned long long g_trace[kTraceRegions * kTraceWarps * 8];
__device__ __forceinline__ void trace_stamp(unsigned region, unsigned long long warp, unsigned slot) {
    if ((threadIdx.x & 31) == 0 && warp < kTraceWarps) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_trace[((unsigned long long)region * kTraceWarps + warp) * 8 + slot] = t;
    }
}
__device__ __forceinline__ void trace_value(unsigned region, unsigned long long warp, unsigned slot, unsigned long long v) {
    if ((threadIdx.x & 31) == 0 && warp < kTraceWarps) g_trace[((unsigned long long)region * kTraceWarps + warp) * 8 + slot] = v;
}
#define PM_TR(region, warp, slot) trace_stamp(region, warp, slot)
#define PM_TV(region, warp, slot, v) trace_value(region, warp, slot, v)
#else
#define PM_TR(region, warp, slot) ((void)0)
#define PM_TV(region, warp, slot, v) ((void)0)
#endif

// ------------------------------------------------------------------------------------ confirm
// What a warp needs to turn one filter hit into verify tasks.
struct ConfirmParams {
    const uint8_t *text;
    unsigned long long total;
    const uint32_t *gkeys, *gfirst, *gcount;
    const Probe *probes;
    const uint8_t *litpool;
    uint32_t log2_cap, fold;
    unsigned long long *tasks;
    unsigned long long task_cap;
};

// One filter hit, the whole warp: (16-byte group, mask of its 4 aligned words) -> exact gram lookup
// (warp-uniform) -> every lane takes one probe of the gram's probe list and compares its literal
// -> (position, pattern) verify tasks.
__device__ __forceinline__ void confirm_candidate(const ConfirmParams &cp, ScanCounters *ctr, unsigned long long group, uint32_t m4, uint32_t step) {
    const int lane = threadIdx.x & 31;
    const uint32_t cap_mask = (1u << cp.log2_cap) - 1u;
    while (m4) {
        const int k = __ffs(m4) - 1;
        m4 &= m4 - 1;
        const unsigned long long p = group * 16 + k * step;
        // (at granule 2 the word may sit at an offset that is only 2-aligned: two aligned halves)
        uint32_t g = (uint32_t)*reinterpret_cast<const unsigned short *>(cp.text + p) | ((uint32_t)*reinterpret_cast<const unsigned short *>(cp.text + p + 2) << 16);
        if (cp.fold) g |= 0x20202020u;
        uint32_t slot = gram_table_slot(g, cp.log2_cap);
        uint32_t key;
        while ((key = __ldg(cp.gkeys + slot)) != g && key != kGramEmpty) slot = (slot + 1) & cap_mask;
        if (key != g) continue;   // filter collision (warp-uniform)
        const uint32_t q0 = __ldg(cp.gfirst + slot), qn = __ldg(cp.gcount + slot);
        for (uint32_t qb = 0; qb < qn; qb += 32) {
            const uint32_t q = qb + lane;
            bool eq = false;
            Probe pr;
            if (q < qn) {
                pr = cp.probes[q0 + q];
                if (p >= kHeaderBytes + pr.gram_off && p - pr.gram_off + pr.lit_len <= cp.total) {
                    // word-wise: literal and case mask are 4-byte aligned and 0xFF-padded (pm_library.cc),
                    // the text word at an arbitrary offset comes from two aligned words and a funnel shift
                    const unsigned long long st = p - pr.gram_off;
                    const uint32_t n_words = (pr.lit_len + 3u) >> 2, sh = (uint32_t)(st & 3) * 8;
                    const uint32_t *lit = reinterpret_cast<const uint32_t *>(cp.litpool + pr.lit_off), *msk = lit + n_words;
                    const uint32_t *tw = reinterpret_cast<const uint32_t *>(cp.text + (st & ~3ull));
                    uint32_t lo = tw[0];
                    eq = true;
                    for (uint32_t j = 0; j < n_words; j++) {
                        const uint32_t hi = tw[j + 1];
                        if ((__funnelshift_r(lo, hi, sh) | msk[j]) != lit[j]) {
                            eq = false;
                            break;
                        }
                        lo = hi;
                    }
                }
            }
            if (eq) {
                unsigned long long at = atomicAdd(&ctr->n_verified, 1ull);
                if (at < cp.task_cap)
                    cp.tasks[at] = (p << kPatternBits) | pr.pattern;
                else
                    ctr->raw_overflow = 1;
            }
        }
    }
}

// ------------------------------------------------------------------------------------ k_scan
// ---- hot-loop primitives.  ncu (profiles/r01_*) showed the first version of this loop bound by
// the ALU pipe (LOP3/SHF/IADD: 69 % of its instructions, pipe ~89 % busy) while the FMA pipe,
// which executes IMAD, idled.  The loop below therefore expresses shifts and adds as integer
// multiply-adds wherever that is exact and never materialises a bit index.  The constants arrive
// as a kernel parameter (constant bank) on purpose: ptxas folds literal powers of two back into
// LEA/VIADD/SHF, i.e. back onto the ALU pipe.
//   (w ^ 0x0A..) & 0x7F..            one LOP3 (register constants)
//   t + 0x7F..                       IMAD  t * 1 + c           (FMA pipe)
//   (u | w) & 0x80..                 one LOP3  -> bit 7 of each byte: "not a newline"
//   counter += 0x80 * popcount       IDP.4A   dot(f, 0x01010101) + acc (one instruction, no shift/popc)
//   bitmap word index                IMAD.HI  hi(h * 2^(BITS-5))   (top bits of the hash)
//   shared address                   IMAD     idx * 4 + base
//   bit test                         SHF.R.W  word >> (h & 31): the shifter wraps, no extraction
__device__ __forceinline__ uint32_t lop3_xor_and(uint32_t a, uint32_t b, uint32_t c) {   // (a ^ b) & c
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x28;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t lop3_or_and(uint32_t a, uint32_t b, uint32_t c) {    // (a | b) & c
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xA8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t lop3_or3(uint32_t a, uint32_t b, uint32_t c) {       // a | b | c
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xFE;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t mad_lo(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t mad_hi(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t shr_wrap(uint32_t v, uint32_t amount) {              // v >> (amount & 31)
    uint32_t r;
    asm("shf.r.wrap.b32 %0, %1, 0, %2;" : "=r"(r) : "r"(v), "r"(amount));
    return r;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t saddr) {
    uint32_t r;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r) : "r"(saddr));
    return r;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
    uint32_t r;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(saddr));
    return r;
}

struct ScanConsts {
    uint32_t c_nl, c_7f, c_80, c_20, one, four, two25, idx_mul, mult, bm_base;
};

// One aligned text word: updates the per-byte "not newline" counters, returns the word's filter
// bit in bit 0 of the result (other bits are garbage).
template <bool FOLD>
__device__ __forceinline__ uint32_t scan_word(uint32_t w, const ScanConsts &k, uint32_t &nz_acc) {
    const uint32_t t = lop3_xor_and(w, k.c_nl, k.c_7f);
    const uint32_t u = mad_lo(t, k.one, k.c_7f);
    const uint32_t f = lop3_or_and(u, w, k.c_80);
    nz_acc = __dp4a(f, k.two25, nz_acc);   // += 0x80 * (number of non-newline bytes); k.two25 holds 0x01010101
    const uint32_t g = FOLD ? (w | k.c_20) : w;
    const uint32_t h = g * k.mult;
    const uint32_t idx = mad_hi(h, k.idx_mul, 0u);
    const uint32_t word = lds_u32(mad_lo(idx, k.four, k.bm_base));
    return shr_wrap(word, h);
}

// Confirming the hits inside this kernel (so that k_confirm drops off the post-scan chain) was measured
// twice in round 2 and not adopted: in the streaming loop every confirmation delayed its warp's last
// tile (warps own a fixed set of tiles; 179 -> 204 us), and as a per-CTA phase after the loop, one thread
// per queued hit, it took 33 us plus 14 us from the changed loop code (179 -> 226 us) - against the
// 17 us of the separate launch, whose ~16 k warps run one hit each.
template <int BITS, bool FOLD>
__global__ void __launch_bounds__(kScanThreads, 1)
k_scan(const uint4 *__restrict__ text, unsigned long long n_tiles, const uint4 *__restrict__ bitmap_g, const uint4 *__restrict__ bitmap2_g,
       uint32_t bits2, uint32_t *__restrict__ tile_nl, unsigned long long *__restrict__ cand, unsigned long long cand_cap,
       ScanCounters *__restrict__ ctr, const ScanConsts kc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int kBitmapWords = (1 << BITS) / 32;
    uint32_t *bm = reinterpret_cast<uint32_t *>(smem_raw);
    const int bm2_words = (1 << bits2) / 32;
    uint32_t *bm2 = bm + kBitmapWords;
    unsigned long long *queues = reinterpret_cast<unsigned long long *>(smem_raw + (size_t)(kBitmapWords + bm2_words) * 4);
    const uint32_t shift2 = 32 - bits2;

    // stage the gram filter: 2^BITS bits (+ the smaller second-level bitmap), 128-bit copies
    for (int i = threadIdx.x; i < kBitmapWords / 4; i += blockDim.x) reinterpret_cast<uint4 *>(bm)[i] = bitmap_g[i];
    for (int i = threadIdx.x; i < bm2_words / 4; i += blockDim.x) reinterpret_cast<uint4 *>(bm2)[i] = bitmap2_g[i];
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long *wq = queues + warp * kQueue;
    int qn = 0;
    const unsigned long long n_warps = (unsigned long long)gridDim.x * (kScanThreads / 32);
    const unsigned lt = lanemask_lt();
    const unsigned long long pol = l2_evict_first_policy();
    ScanConsts k = kc;
    k.bm_base = (uint32_t)__cvta_generic_to_shared(bm);

    // software pipeline: the next tile's four 128-bit loads are in flight while this tile is
    // processed (a warp alone would otherwise alternate "wait ~1 us for HBM" and "compute")
    auto process_tile = [&](unsigned long long tt, uint4 (&v)[4]) {
        uint32_t nz_acc = 0;   // 0x80 per non-newline byte of this lane's 64 bytes
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t t0 = scan_word<FOLD>(v[j].x, k, nz_acc), t1 = scan_word<FOLD>(v[j].y, k, nz_acc);
            const uint32_t t2 = scan_word<FOLD>(v[j].z, k, nz_acc), t3 = scan_word<FOLD>(v[j].w, k, nz_acc);
            const uint32_t any = lop3_or_and(lop3_or3(t0, t1, t2), t3, k.one);
            unsigned b = __ballot_sync(FULL, any != 0);
            if (b) {   // warp-uniform, rare.  Second-level filter first: an independent hash into a small
                       // bitmap removes the first level's collisions before anything is queued
                uint32_t m4 = (t0 & 1u) | ((t1 & 1u) << 1) | ((t2 & 1u) << 2) | ((t3 & 1u) << 3);
                if (m4) {
                    const uint32_t w4[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        if (m4 & (1u << q)) {
                            uint32_t g = FOLD ? (w4[q] | 0x20202020u) : w4[q];
                            uint32_t bit2 = (g * 0xC2B2AE3Du) >> shift2;
                            if (!((bm2[bit2 >> 5] >> (bit2 & 31u)) & 1u)) m4 &= ~(1u << q);
                        }
                }
                b = __ballot_sync(FULL, m4 != 0);
                if (b) {   // stage (group, mask) entries, flush 32 at a time
                    if (m4) wq[qn + __popc(b & lt)] = ((tt * (kTileBytes / 16) + j * 32 + lane) << 4) | m4;
                    qn += __popc(b);
                    __syncwarp();
                    if (qn >= 32) {
                        unsigned long long at = 0;
                        if (lane == 0) at = atomicAdd(&ctr->n_cand, 32ull);
                        at = __shfl_sync(FULL, at, 0);
                        if (at + 32 <= cand_cap)
                            cand[at + lane] = wq[lane];
                        else if (lane == 0)
                            ctr->cand_overflow = 1;
                        unsigned long long carry = (lane + 32 < qn) ? wq[lane + 32] : 0ull;
                        __syncwarp();
                        qn -= 32;
                        if (lane < qn) wq[lane] = carry;
                        __syncwarp();
                    }
                }
            }
        }
        // horizontal sum of the four byte counters, then newlines = 64 - non-newline bytes
        const uint32_t nz = nz_acc >> 7;
        uint32_t nl = __reduce_add_sync(FULL, 64u - nz);
        if (lane == 0) tile_nl[tt] = nl;
    };
    // ping-pong between two register sets so that no copy is needed between rounds
    unsigned long long t = (unsigned long long)blockIdx.x * (kScanThreads / 32) + warp;
    uint4 va[4], vb[4];
    auto load_tile = [&](unsigned long long tt, uint4 (&dst)[4]) {
        const unsigned long long tc = tt < n_tiles ? tt : n_tiles - 1;   // past the end: harmless re-load
        const uint4 *nb = text + tc * (kTileBytes / 16) + lane;
#pragma unroll
        for (int j = 0; j < 4; j++) dst[j] = ld_stream16_ef(nb + j * 32, pol);
    };
    if (t < n_tiles) load_tile(t, va);
    while (t < n_tiles) {
        load_tile(t + n_warps, vb);
        process_tile(t, va);
        t += n_warps;
        if (t >= n_tiles) break;
        load_tile(t + n_warps, va);
        process_tile(t, vb);
        t += n_warps;
    }
    if (qn > 0) {
        unsigned long long at = 0;
        if (lane == 0) at = atomicAdd(&ctr->n_cand, (unsigned long long)qn);
        at = __shfl_sync(FULL, at, 0);
        if (at + qn <= cand_cap) {
            if (lane < qn) cand[at + lane] = wq[lane];
            if (lane + 32 < qn) cand[at + lane + 32] = wq[lane + 32];
        } else if (lane == 0) {
            ctr->cand_overflow = 1;
        }
    }
}

// ------------------------------------------------------------------------------------ k_scan_half
// Granule 2 (pm_blob.h): the filter also tests the words that straddle two aligned ones, i.e. the 4-byte word at
// every EVEN text offset, so that a required literal of 5 bytes is enough (it always contains a word that starts
// at an even offset).  Same structure as k_scan - a kernel of its own so that k_scan's loop, which sits exactly at
// 64 registers, is not touched.  Per aligned word one funnel shift, one more hash and one more bitmap lookup; the
// word behind a lane's 16 bytes comes from the next lane (the last lane: from lane 0 of the next 512-byte row, or
// from the tile behind).  Hits leave as (16-byte group << 8) | mask of the 8 even offsets.
template <bool FOLD>
__device__ __forceinline__ uint32_t filter_word(uint32_t w, const ScanConsts &k) {
    const uint32_t g = FOLD ? (w | k.c_20) : w;
    const uint32_t h = g * k.mult;
    const uint32_t idx = mad_hi(h, k.idx_mul, 0u);
    const uint32_t word = lds_u32(mad_lo(idx, k.four, k.bm_base));
    return shr_wrap(word, h);
}

constexpr int kHalfThreads = 1024;

template <int BITS, bool FOLD>
__global__ void __launch_bounds__(kHalfThreads, 1)
k_scan_half(const uint4 *__restrict__ text, unsigned long long n_tiles, const uint4 *__restrict__ bitmap_g, const uint4 *__restrict__ bitmap2_g,
            uint32_t bits2, uint32_t *__restrict__ tile_nl, unsigned long long *__restrict__ cand, unsigned long long cand_cap,
            ScanCounters *__restrict__ ctr, const ScanConsts kc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int kBitmapWords = (1 << BITS) / 32;
    uint32_t *bm = reinterpret_cast<uint32_t *>(smem_raw);
    const int bm2_words = (1 << bits2) / 32;
    uint32_t *bm2 = bm + kBitmapWords;
    unsigned long long *queues = reinterpret_cast<unsigned long long *>(smem_raw + (size_t)(kBitmapWords + bm2_words) * 4);
    const uint32_t shift2 = 32 - bits2;
    for (int i = threadIdx.x; i < kBitmapWords / 4; i += blockDim.x) reinterpret_cast<uint4 *>(bm)[i] = bitmap_g[i];
    for (int i = threadIdx.x; i < bm2_words / 4; i += blockDim.x) reinterpret_cast<uint4 *>(bm2)[i] = bitmap2_g[i];
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long *wq = queues + warp * kQueue;
    int qn = 0;
    const unsigned long long n_warps = (unsigned long long)gridDim.x * (kHalfThreads / 32);
    const unsigned lt = lanemask_lt();
    const unsigned long long pol = l2_evict_first_policy();
    ScanConsts k = kc;
    k.bm_base = (uint32_t)__cvta_generic_to_shared(bm);
    const uint32_t *text_w = reinterpret_cast<const uint32_t *>(text);

    auto process_tile = [&](unsigned long long tt, uint4 (&v)[4], uint32_t behind) {   // behind: first word of the tile after this one
        uint32_t nz_acc = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            // the word that follows this lane's 16 bytes
            const uint32_t from_next_lane = __shfl_down_sync(FULL, v[j].x, 1);
            const uint32_t from_next_row = j < 3 ? __shfl_sync(FULL, v[j < 3 ? j + 1 : 3].x, 0) : behind;
            const uint32_t nx = lane == 31 ? from_next_row : from_next_lane;
            const uint32_t t0 = scan_word<FOLD>(v[j].x, k, nz_acc), t1 = scan_word<FOLD>(v[j].y, k, nz_acc);
            const uint32_t t2 = scan_word<FOLD>(v[j].z, k, nz_acc), t3 = scan_word<FOLD>(v[j].w, k, nz_acc);
            const uint32_t s0 = filter_word<FOLD>(__funnelshift_r(v[j].x, v[j].y, 16), k), s1 = filter_word<FOLD>(__funnelshift_r(v[j].y, v[j].z, 16), k);
            const uint32_t s2 = filter_word<FOLD>(__funnelshift_r(v[j].z, v[j].w, 16), k), s3 = filter_word<FOLD>(__funnelshift_r(v[j].w, nx, 16), k);
            const uint32_t any = lop3_or_and(lop3_or3(t0, t1, t2) | lop3_or3(s0, s1, s2), t3 | s3, k.one);
            unsigned b = __ballot_sync(FULL, any != 0);
            if (b) {   // warp-uniform, rare: second-level bitmap, then the queue
                uint32_t m8 = (t0 & 1u) | ((s0 & 1u) << 1) | ((t1 & 1u) << 2) | ((s1 & 1u) << 3) | ((t2 & 1u) << 4) | ((s2 & 1u) << 5) | ((t3 & 1u) << 6) |
                              ((s3 & 1u) << 7);
                if (m8) {
                    const uint32_t w8[8] = {v[j].x, __funnelshift_r(v[j].x, v[j].y, 16), v[j].y, __funnelshift_r(v[j].y, v[j].z, 16),
                                            v[j].z, __funnelshift_r(v[j].z, v[j].w, 16), v[j].w, __funnelshift_r(v[j].w, nx, 16)};
#pragma unroll
                    for (int q = 0; q < 8; q++)
                        if (m8 & (1u << q)) {
                            uint32_t g = FOLD ? (w8[q] | 0x20202020u) : w8[q];
                            uint32_t bit2 = (g * 0xC2B2AE3Du) >> shift2;
                            if (!((bm2[bit2 >> 5] >> (bit2 & 31u)) & 1u)) m8 &= ~(1u << q);
                        }
                }
                b = __ballot_sync(FULL, m8 != 0);
                if (b) {
                    if (m8) wq[qn + __popc(b & lt)] = ((tt * (kTileBytes / 16) + j * 32 + lane) << 8) | m8;
                    qn += __popc(b);
                    __syncwarp();
                    if (qn >= 32) {
                        unsigned long long at = 0;
                        if (lane == 0) at = atomicAdd(&ctr->n_cand, 32ull);
                        at = __shfl_sync(FULL, at, 0);
                        if (at + 32 <= cand_cap)
                            cand[at + lane] = wq[lane];
                        else if (lane == 0)
                            ctr->cand_overflow = 1;
                        unsigned long long carry = (lane + 32 < qn) ? wq[lane + 32] : 0ull;
                        __syncwarp();
                        qn -= 32;
                        if (lane < qn) wq[lane] = carry;
                        __syncwarp();
                    }
                }
            }
        }
        const uint32_t nz = nz_acc >> 7;
        uint32_t nl = __reduce_add_sync(FULL, 64u - nz);
        if (lane == 0) tile_nl[tt] = nl;
    };
    unsigned long long t = (unsigned long long)blockIdx.x * (kHalfThreads / 32) + warp;
    uint4 va[4];
    uint32_t na = 0x0A0A0A0Au;
    auto load_tile = [&](unsigned long long tt, uint4 (&dst)[4], uint32_t &behind) {
        const unsigned long long tc = tt < n_tiles ? tt : n_tiles - 1;   // past the end: harmless re-load
        const uint4 *nb = text + tc * (kTileBytes / 16) + lane;
#pragma unroll
        for (int j = 0; j < 4; j++) dst[j] = ld_stream16_ef(nb + j * 32, pol);
        // (only the last lane uses it; behind the last tile there is nothing: the buffer ends in '\n' padding)
        behind = (lane == 31 && tc + 1 < n_tiles) ? __ldg(text_w + (tc + 1) * (kTileBytes / 4)) : 0x0A0A0A0Au;
    };
    // One register set, no software pipeline: this kernel is bound by instruction issue, not by HBM, and the 32 warps of
    // the CTA hide each other's load latency.  (With k_scan's two register sets it needed 80 registers: 768 threads,
    // 0.43 ms per GiB; this form fits 1024 threads x 64 registers and takes 0.32 ms.)
    for (; t < n_tiles; t += n_warps) {
        load_tile(t, va, na);
        process_tile(t, va, na);
    }
    if (qn > 0) {
        unsigned long long at = 0;
        if (lane == 0) at = atomicAdd(&ctr->n_cand, (unsigned long long)qn);
        at = __shfl_sync(FULL, at, 0);
        if (at + qn <= cand_cap) {
            if (lane < qn) cand[at + lane] = wq[lane];
            if (lane + 32 < qn) cand[at + lane + 32] = wq[lane + 32];
        } else if (lane == 0) {
            ctr->cand_overflow = 1;
        }
    }
}

// Newline-only variant: libraries without any filtered pattern still need the tile counts.
__global__ void __launch_bounds__(kScanThreads, 1)
k_scan_nl_only(const uint4 *__restrict__ text, unsigned long long n_tiles, uint32_t *__restrict__ tile_nl) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long n_warps = (unsigned long long)gridDim.x * (kScanThreads / 32);
    for (unsigned long long t = (unsigned long long)blockIdx.x * (kScanThreads / 32) + warp; t < n_tiles; t += n_warps) {
        const uint4 *base = text + t * (kTileBytes / 16) + lane;
        uint4 v[4];
#pragma unroll
        for (int j = 0; j < 4; j++) v[j] = ld_stream16(base + j * 32);
        uint32_t nz = 0;
#pragma unroll
        for (int j = 0; j < 4; j++)
            nz += __popc(not_newline_flags(v[j].x) | (not_newline_flags(v[j].y) >> 1) | (not_newline_flags(v[j].z) >> 2) |
                         (not_newline_flags(v[j].w) >> 3));
        uint32_t nl = __reduce_add_sync(FULL, 64u - nz);
        if (lane == 0) tile_nl[t] = nl;
    }
}

// ------------------------------------------------------------------------------------ prefix
// 256-thread CTAs on purpose: these kernels run on a side stream while k_verify (1536 resident threads
// per SM) owns the GPU, and a 1024-thread CTA would find no SM to start on until that kernel drains.
constexpr int kPrefThreads = 256, kPrefItems = 16, kPrefBlock = kPrefThreads * kPrefItems;

// Exclusive scan over the CTA (any multiple of 32 threads up to 1024); *total receives the sum.
__device__ __forceinline__ unsigned long long block_exclusive_scan(unsigned long long v, unsigned long long *total) {
    __shared__ unsigned long long wsum[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    unsigned long long inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long y = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += y;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        unsigned long long w = lane < n_warps ? wsum[lane] : 0ull, winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(FULL, winc, o);
            if (lane >= o) winc += y;
        }
        wsum[lane] = winc - w;
        if (lane == 31 && total) *total = winc;
    }
    __syncthreads();
    unsigned long long r = inc - v + wsum[warp];
    __syncthreads();
    return r;
}

// Sum of one block of kPrefBlock tile counts.
__global__ void __launch_bounds__(kPrefThreads) k_prefix_partial(const uint32_t *__restrict__ in, unsigned long long n, unsigned long long *__restrict__ block_sums) {
    const unsigned long long i0 = (unsigned long long)blockIdx.x * kPrefBlock + threadIdx.x * 4;
    unsigned long long s = 0;
#pragma unroll
    for (int r = 0; r < kPrefItems / 4; r++) {   // coalesced 16-byte loads (n_tiles counts are padded to a multiple of 4 by the caller's buffer)
        const unsigned long long i = i0 + (unsigned long long)r * kPrefThreads * 4;
        if (i + 4 <= n) {
            const uint4 x = *reinterpret_cast<const uint4 *>(in + i);
            s += (unsigned long long)x.x + x.y + x.z + x.w;
        } else {
            for (int k = 0; k < 4; k++)
                if (i + k < n) s += in[i + k];
        }
    }
    __shared__ unsigned long long tot;
    block_exclusive_scan(s, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// Tile prefix of one block: every CTA first adds up the sums of the blocks in front of it (a few
// hundred values even for a full device), then scans its own counts.  One launch instead of a
// single-CTA scan of the block sums followed by the per-block pass.
__global__ void __launch_bounds__(kPrefThreads) k_prefix_final(const uint32_t *__restrict__ in, unsigned long long n, const unsigned long long *__restrict__ block_sums,
                                                                unsigned long long *__restrict__ out) {
    __shared__ unsigned long long base_s;
    unsigned long long mine = 0;
    for (unsigned b = threadIdx.x; b < blockIdx.x; b += kPrefThreads) mine += block_sums[b];
    block_exclusive_scan(mine, &base_s);
    // thread t owns kPrefItems consecutive counts
    const unsigned long long i0 = (unsigned long long)blockIdx.x * kPrefBlock + (unsigned long long)threadIdx.x * kPrefItems;
    uint32_t x[kPrefItems];
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < kPrefItems; k += 4) {
        if (i0 + k + 4 <= n) {
            const uint4 v = *reinterpret_cast<const uint4 *>(in + i0 + k);
            x[k] = v.x, x[k + 1] = v.y, x[k + 2] = v.z, x[k + 3] = v.w;
        } else {
            for (int q = 0; q < 4; q++) x[k + q] = (i0 + k + q < n) ? in[i0 + k + q] : 0;
        }
        s += (unsigned long long)x[k] + x[k + 1] + x[k + 2] + x[k + 3];
    }
    unsigned long long ex = block_exclusive_scan(s, nullptr) + base_s;
#pragma unroll
    for (int k = 0; k < kPrefItems; k += 2) {
        if (i0 + k + 2 <= n) {
            *reinterpret_cast<ulonglong2 *>(out + i0 + k) = make_ulonglong2(ex, ex + x[k]);
        } else if (i0 + k < n) {
            out[i0 + k] = ex;
        }
        ex += (unsigned long long)x[k] + x[k + 1];
    }
}

// ------------------------------------------------------------------------------------ helpers
// '\n' bytes in text[0, pos): tile prefix + a warp-cooperative recount inside pos's tile.
__device__ unsigned long long warp_newlines_before(const uint8_t *text, const unsigned long long *tile_prefix, unsigned long long pos) {
    const int lane = threadIdx.x & 31;
    unsigned long long tile = pos / kTileBytes, base = tile * kTileBytes;
    unsigned cnt = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        unsigned long long a = base + k * 512 + lane * 16;
        if (a < pos) {
            uint4 v = *reinterpret_cast<const uint4 *>(text + a);
            uint32_t m = newline_mask16(v);
            unsigned long long r = pos - a;
            if (r < 16) m &= (1u << r) - 1u;
            cnt += __popc(m);
        }
    }
    cnt = __reduce_add_sync(FULL, cnt);
    return tile_prefix[tile] + cnt;
}

struct UnitTables {
    const uint16_t *tab;
    const uint16_t *wide;   // 256-column table or nullptr
    const uint8_t *cmap;
    uint32_t stride, accept, dead, mid;
};

__device__ __forceinline__ UnitTables unit_tables(const uint8_t *blob, const BlobHeader &h, uint32_t unit) {
    const UnitDesc *u = reinterpret_cast<const UnitDesc *>(blob + h.off_units) + unit;
    UnitTables t;
    t.tab = reinterpret_cast<const uint16_t *>(blob + h.off_tables) + u->table_off;
    t.cmap = blob + h.off_classmaps + u->classmap_off;
    t.wide = u->wide_off != 0xFFFFFFFFu ? reinterpret_cast<const uint16_t *>(blob + h.off_tables) + u->wide_off : nullptr;
    t.stride = u->stride;
    t.accept = u->accept_state;
    t.dead = u->dead_state;
    t.mid = u->mid_state;
    return t;
}

// One thread walks one line [a, b) through a verify DFA, then feeds END-OF-LINE.
__device__ bool line_matches(const uint8_t *text, unsigned long long a, unsigned long long b, const UnitTables &t) {
    uint32_t s = 0;
    if (t.wide) {
        // 256-column table in global memory (units whose table does not fit k_check's shared-memory budget): one dependent
        // lookup per byte, the text in aligned 16-byte pieces.  The accepting state is sticky and the dead state absorbing,
        // so both are tested once per piece.
        const uint16_t *w = t.wide;
        for (unsigned long long p = a & ~15ull; p < b && s != t.accept && s != t.dead; p += 16) {
            const uint4 v = *reinterpret_cast<const uint4 *>(text + p);
            const uint32_t x[4] = {v.x, v.y, v.z, v.w};
            // one code path for whole and partial pieces (see walk_staged): bytes outside the line are looked up and dropped
            const int lo = p >= a ? 0 : (int)(a - p), hi = p + 16 <= b ? 16 : (int)(b - p);
            const uint32_t valid = (0xFFFFu >> (16 - hi)) & (0xFFFFu << lo);
#pragma unroll
            for (int j = 0; j < 16; j++) {
                const uint32_t nx = w[(s << 8) | ((x[j >> 2] >> (8 * (j & 3))) & 0xFFu)];
                s = ((valid >> j) & 1u) ? nx : s;
            }
        }
        if (s == t.accept) return true;
        if (s == t.dead) return false;
        s = t.tab[s * t.stride + t.stride - 1];
        return s == t.accept;
    }
    for (unsigned long long p = a; p < b; p++) {
        s = t.tab[s * t.stride + t.cmap[text[p]]];
        if (s == t.accept) return true;
        if (s == t.dead) return false;   // anchored regexes give up after a few bytes
    }
    s = t.tab[s * t.stride + t.stride - 1];
    return s == t.accept;
}

// Is `le` (the position of a '\n') the first padding byte after some pod, i.e. not a separator
// that belongs to the log?  Only called when the batch has a pod ending in a lone '\r'.
__device__ bool is_pod_raw_end(const PodTable &pt, unsigned long long le) {
    int lo = 0, hi = pt.n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pt.pods[mid].start <= le) lo = mid; else hi = mid - 1;
    }
    return pt.n > 0 && pt.pods[lo].raw_end == le;
}

// End of the line whose '\n' sits at `le`: the '\r' of a "\r\n" separator is not part of the line.
__device__ __forceinline__ unsigned long long strip_cr(const uint8_t *text, const PodTable &pt, unsigned long long ls, unsigned long long le) {
    if (le > ls && text[le - 1] == '\r' && !(pt.any_trailing_cr && is_pod_raw_end(pt, le))) return le - 1;
    return le;
}

// ------------------------------------------------------------------------------------ k_pad / k_pods
__global__ void k_pad(uint8_t *text, const PodDev *pods, int n_pods, unsigned long long total) {
    int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w > n_pods) return;
    // gap w: [raw_end[w-1], start[w]) ; gap 0 is the header, gap n_pods runs to the end of the buffer
    unsigned long long a = w == 0 ? 0 : pods[w - 1].raw_end, b = w == n_pods ? total : pods[w].start;
    for (unsigned long long p = a + lane; p < b; p += 32) text[p] = '\n';
}

__global__ void k_pods(const uint8_t *text, const unsigned long long *tile_prefix, PodDev *pods, int n_pods) {
    int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= n_pods) return;
    unsigned long long st = pods[w].start, ee = pods[w].eff_end;
    unsigned long long c0 = warp_newlines_before(text, tile_prefix, st);
    unsigned long long c1 = warp_newlines_before(text, tile_prefix, ee);
    if (lane == 0) {
        pods[w].nl_before = c0;
        pods[w].n_lines = ee > st ? (int)(c1 - c0 + 1) : 0;
    }
}

// ------------------------------------------------------------------------------------ k_confirm
// Stage 1 of the confirmation: candidate -> exact gram -> literal.  Survivors become verify
// tasks (position, pattern).  One WARP per candidate: the gram lookup is uniform, then every
// lane takes one probe of the gram's probe list and compares its literal.  (One thread per
// candidate made each warp serialise 32 unrelated probe/compare loops.)
__global__ void __launch_bounds__(256) k_confirm(const uint8_t *__restrict__ text, unsigned long long total, FilterView fv,
                                                 const unsigned long long *__restrict__ cand, unsigned long long cand_cap,
                                                 unsigned long long *__restrict__ tasks, unsigned long long task_cap, ScanCounters *ctr) {
    if (ctr->cand_overflow) return;   // the host reruns the scan with a larger queue
    const unsigned long long n = ctr->n_cand < cand_cap ? ctr->n_cand : cand_cap;
    const unsigned long long n_warps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    ConfirmParams cp;
    cp.text = text, cp.total = total, cp.gkeys = fv.gkeys, cp.gfirst = fv.gfirst, cp.gcount = fv.gcount, cp.probes = fv.probes, cp.litpool = fv.litpool;
    cp.log2_cap = fv.log2_cap, cp.fold = fv.fold, cp.tasks = tasks, cp.task_cap = task_cap;
    const bool half = fv.granule == 2;   // (group << 8) | 8 even offsets, else (group << 4) | 4 aligned words
    PM_TR(0, ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, 0);
    for (unsigned long long i = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += n_warps) {
        const unsigned long long e = cand[i];
        PM_TR(0, ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, 1);
        if (half)
            confirm_candidate(cp, ctr, e >> 8, (uint32_t)(e & 255u), 2u);
        else
            confirm_candidate(cp, ctr, e >> 4, (uint32_t)(e & 15u), 4u);
    }
    PM_TR(0, ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, 2);
}

// One warp walks one line [a, b) through a unit DFA and returns the state after the last byte
// (END-OF-LINE not yet fed), or `acc` / `dead` as soon as one of them is reached (accepting is
// sticky, dead is absorbing).  A walk is a chain of dependent table lookups, so everything that
// does not depend on the state is taken off the chain first: the warp loads the line 512 bytes at
// a time (16 aligned bytes per lane), maps the bytes to character classes in parallel and leaves
// the classes in shared memory.  What remains per byte is one shared-memory read that can be
// issued ahead and one table read through L1 (the caller has touched the table's cache lines, see
// warp_touch_table: a line meets its pattern's table for the first time, and an L1 miss in the
// chain costs ~300 ns).
//   - pieces up to kWalkSerial bytes: every lane runs the same chain (broadcast reads, no divergence);
//   - longer pieces: lane i guesses the state at the start of its 1/32 of the piece (the DFA's
//     mid-line idle state warmed up over the 16 bytes in front), the guesses are validated in lane
//     order and a wrong one re-walks that chunk from the true state: exact, and on text that
//     leaves the DFA idle ~16x shorter than the serial chain.  Matching lines do not leave the DFA
//     idle (after the first half of "A.*B" every later guess is wrong), which is why short lines,
//     the common case, are not speculated on at all.
// gcm: the unit's 256-byte class map (global); cmap_s / cls: 256 / kWalkPiece bytes of shared
// memory owned by the warp.
constexpr uint32_t kWalkPiece = 512, kWalkSerial = 192;

// Brings a table of `bytes` bytes into L1 with one load per 32-byte sector (L1 fills sector by
// sector), all of them in flight at once.
__device__ __forceinline__ void warp_touch_table(const uint16_t *tab, uint32_t bytes) {
    const int lane = threadIdx.x & 31;
    const uint8_t *base = reinterpret_cast<const uint8_t *>(tab);
    uint32_t sink = 0;
#pragma unroll 4
    for (uint32_t off = lane * 32u; off < bytes; off += 32u * 32u) sink ^= __ldg(reinterpret_cast<const uint32_t *>(base + off));
    asm volatile("" ::"r"(sink));   // keeps the loads; nothing waits for them here
}

// TAB_SHARED: `tab` is the warp's copy of the table in shared memory (k_verify stages tables that fit
// its per-warp budget: a lookup then costs one shared-memory read, ~25 cycles, and never misses; through
// L1 the same walk saw a 56 % hit rate, 48 warps per SM with a table each, and ~70 % of the kernel's
// stall samples sat on these loads).
// TAB: 0 = `tab` is the table in global memory, 1 = the warp's copy in shared memory, 2 = the same with one byte per
// entry (units of at most 256 states: twice the tables fit the warp's budget).
template <int TAB>
__device__ __forceinline__ uint32_t warp_walk_line(const uint8_t *__restrict__ text, unsigned long long a, unsigned long long b, const uint16_t *__restrict__ tab,
                                                   uint32_t stride, uint32_t acc, uint32_t dead, uint32_t mid, const uint8_t *__restrict__ gcm,
                                                   uint8_t *cmap_s, uint8_t *cls) {
    const int lane = threadIdx.x & 31;
    uint32_t s = 0;
    const uint32_t tab_saddr = TAB ? (uint32_t)__cvta_generic_to_shared(tab) : 0u;
    auto next = [&](uint32_t st, uint32_t c) -> uint32_t {
        if (TAB == 2) {
            uint32_t v;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(tab_saddr + (st * stride + c)));
            return v;
        }
        if (TAB == 1) {
            unsigned short v;
            asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(tab_saddr + 2u * (st * stride + c)));
            return v;
        }
        return __ldg(tab + (st * stride + c));
    };
    auto walk = [&](uint32_t s0, uint32_t p0, uint32_t p1) {
        uint32_t k = p0;
        for (; k + 4 <= p1 && s0 != acc && s0 != dead; k += 4) {
            const uint32_t c0 = cls[k], c1 = cls[k + 1], c2 = cls[k + 2], c3 = cls[k + 3];
            s0 = next(s0, c0);
            s0 = next(s0, c1);
            s0 = next(s0, c2);
            s0 = next(s0, c3);
        }
        for (; k < p1 && s0 != acc && s0 != dead; k++) s0 = next(s0, cls[k]);
        return s0;
    };
    bool have_cmap = false;
    for (unsigned long long base = a & ~15ull; base < b && s != acc && s != dead; base += kWalkPiece) {
        __syncwarp();
        const unsigned long long mine = base + 16ull * lane;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (mine < b) v = *reinterpret_cast<const uint4 *>(text + mine);
        if (!have_cmap) {   // the class map travels together with the first piece of the line
            const uint32_t m0 = __ldg(reinterpret_cast<const uint32_t *>(gcm) + lane), m1 = __ldg(reinterpret_cast<const uint32_t *>(gcm) + 32 + lane);
            reinterpret_cast<uint32_t *>(cmap_s)[lane] = m0;
            reinterpret_cast<uint32_t *>(cmap_s)[32 + lane] = m1;
            have_cmap = true;
            __syncwarp();
        }
        if (mine < b) {
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
            uint32_t o[4];
#pragma unroll
            for (int q = 0; q < 4; q++)
                o[q] = (uint32_t)cmap_s[w[q] & 0xFFu] | ((uint32_t)cmap_s[(w[q] >> 8) & 0xFFu] << 8) | ((uint32_t)cmap_s[(w[q] >> 16) & 0xFFu] << 16) |
                       ((uint32_t)cmap_s[w[q] >> 24] << 24);
            *reinterpret_cast<uint4 *>(cls + 16 * lane) = make_uint4(o[0], o[1], o[2], o[3]);
        }
        __syncwarp();
        const uint32_t lo = a > base ? (uint32_t)(a - base) : 0u;
        const uint32_t hi = b - base < kWalkPiece ? (uint32_t)(b - base) : kWalkPiece;
        const uint32_t n = hi - lo;
        if (n <= kWalkSerial) {
            s = walk(s, lo, hi);
            continue;
        }
        const uint32_t c = (n + 31) / 32;
        const uint32_t ca = min(hi, lo + lane * c), cb = min(hi, ca + c);
        const uint32_t in_spec = (lane == 0 || ca - lo <= 16) ? walk(s, lo, ca) : walk(mid, ca - 16, ca);
        const uint32_t out_spec = walk(in_spec, ca, cb);
        uint32_t truth = s;   // true state at the start of chunk i
        for (int i = 0; i < 32; i++) {
            uint32_t in_i = __shfl_sync(FULL, in_spec, i), out_i = __shfl_sync(FULL, out_spec, i);
            if (i > 0 && in_i != truth) {   // warp-uniform; chunk 0 starts from the true state by construction
                uint32_t redo = 0;
                if (lane == i) redo = walk(truth, ca, cb);
                out_i = __shfl_sync(FULL, redo, i);
            }
            truth = out_i;
            if (truth == acc || truth == dead) break;
        }
        s = truth;
    }
    __syncwarp();
    return s;
}

// Stage 2: one warp per verify task.  The chain "find the line, fetch the pattern's tables, walk
// the line" is a string of dependent cache misses when one thread does it byte by byte, so the
// warp fetches in bulk instead: 512 bytes around the literal in one round trip (newline search by
// ballot), the pattern's class map into shared memory, then warp_walk_line.
constexpr int kVerifyWarps = 4;

__global__ void __launch_bounds__(kVerifyWarps * 32, 12) k_verify(const uint8_t *__restrict__ text, unsigned long long total, const uint8_t *__restrict__ blob, BlobHeader h, PodTable pt,
                                                              const unsigned long long *__restrict__ tasks, unsigned long long task_cap,
                                                              unsigned long long *__restrict__ raw, unsigned long long raw_cap, ScanCounters *ctr,
                                                              uint32_t table_cap) {
    extern __shared__ __align__(16) unsigned char s_tables[];   // kVerifyWarps x table_cap bytes: each warp's copy of its task's table
    __shared__ __align__(16) uint8_t s_cmap[kVerifyWarps][256];
    __shared__ __align__(16) uint8_t s_cls[kVerifyWarps][kWalkPiece];
    if (ctr->cand_overflow || ctr->raw_overflow) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long n = ctr->n_verified < task_cap ? ctr->n_verified : task_cap;
    const PatternDesc *pd = reinterpret_cast<const PatternDesc *>(blob + h.off_patterns);
    const unsigned long long n_warps = (unsigned long long)gridDim.x * kVerifyWarps;
    PM_TR(1, (unsigned long long)blockIdx.x * kVerifyWarps + warp, 0);
    for (unsigned long long i = (unsigned long long)blockIdx.x * kVerifyWarps + warp; i < n; i += n_warps) {
        const unsigned long long t = tasks[i], p = t >> kPatternBits;
        const uint32_t pat = (uint32_t)(t & ((1u << kPatternBits) - 1u));
        PM_TR(1, i, 1);
        // ---- line bounds: 16 lanes look back, 16 lanes look ahead, 256 bytes each, one round trip;
        //      the pattern's descriptors are fetched under it, and the table's lines after them
        const unsigned long long p16 = p & ~15ull;
        unsigned long long a = lane < 16 ? p16 - 16ull * (lane + 1) : p16 + 16ull * (lane - 16);
        uint4 around = make_uint4(0x20202020u, 0x20202020u, 0x20202020u, 0x20202020u);
        if (lane >= 16 || p16 >= 16ull * (lane + 1)) around = *reinterpret_cast<const uint4 *>(text + a);
        const UnitDesc ud = reinterpret_cast<const UnitDesc *>(blob + h.off_units)[pd[pat].unit];
        const uint16_t *gtab = reinterpret_cast<const uint16_t *>(blob + h.off_tables) + ud.table_off;
        // (opaque to the compiler from here on: otherwise it re-derives the pointer from the uniform base
        // and the 32-bit table offset inside every lookup, a 64-bit carry chain of five instructions)
        asm volatile("" : "+l"(gtab));
        const uint32_t entries = (uint32_t)ud.nstates * ud.stride, tbytes = entries * 2u;
        // (warp-uniform) one byte per entry where the state numbers allow it: most verify tables are 2-7 KiB, the budget 4 KiB
        const bool staged8 = ud.nstates <= 256 && entries <= table_cap;
        const bool staged = !staged8 && tbytes <= table_cap;
        unsigned char *stab = s_tables + (size_t)warp * table_cap;
        // a pattern that is nothing but its literal(s): k_confirm's comparison was the match, only the line start is
        // still needed (warp-uniform)
        const bool lit = ud.literal_is_match != 0;
        if (lit) {
        } else if (staged8) {   // 8 entries per lane and step, all loads of the warp in flight together
            __syncwarp();
            const uint4 *gt = reinterpret_cast<const uint4 *>(gtab);
#pragma unroll 4
            for (uint32_t e = lane * 8u; e < entries; e += 256u) {
                const uint4 q = __ldg(gt + (e >> 3));
                uint2 o;
                o.x = __byte_perm(q.x, q.y, 0x6420);
                o.y = __byte_perm(q.z, q.w, 0x6420);
                *reinterpret_cast<uint2 *>(stab + e) = o;
            }
        } else if (staged) {   // tables start 16-byte aligned in the blob and are followed by at least their padding
            __syncwarp();
            for (uint32_t off = lane * 16u; off < tbytes; off += 512u)
                *reinterpret_cast<uint4 *>(stab + off) = __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned char *>(gtab) + off));
        } else {
            warp_touch_table(gtab, tbytes);
        }
        uint32_t m = newline_mask16(around);
        PM_TR(1, i, 2);
        if (lane == 16) m &= ~((1u << (p - p16)) - 1u);          // forward search starts at p
        unsigned has = __ballot_sync(FULL, m != 0);
        unsigned long long ls, le;
        unsigned fwd = has >> 16, back = has & 0xFFFFu;
        if (fwd) {
            int src = 16 + __ffs(fwd) - 1;
            uint32_t mm = __shfl_sync(FULL, m, src);
            le = p16 + 16ull * (src - 16) + (__ffs(mm) - 1);
        } else if (lit) {
            le = p;   // not needed
        } else {   // a long line: keep looking, 512 bytes per round trip (the text is '\n'-padded up to `total`)
            le = total;
            for (unsigned long long q = p16 + 256; q < total; q += 512) {
                const unsigned long long mine = q + 16ull * lane;
                const uint32_t mq = mine < total ? newline_mask16(*reinterpret_cast<const uint4 *>(text + mine)) : 0u;
                const unsigned hq = __ballot_sync(FULL, mq != 0);
                if (hq) {
                    const int src = __ffs(hq) - 1;
                    le = q + 16ull * src + (__ffs(__shfl_sync(FULL, mq, src)) - 1);
                    break;
                }
            }
        }
        // backward: the nearest newline strictly before p (bytes of p16's own chunk below p belong to lane 16's chunk)
        uint32_t m_own = __shfl_sync(FULL, lane == 16 ? newline_mask16(*reinterpret_cast<const uint4 *>(text + p16)) & ((1u << (p - p16)) - 1u) : 0u, 16);
        if (m_own) {
            ls = p16 + (31 - __clz(m_own)) + 1;
        } else if (back) {
            int src = __ffs(back) - 1;                           // lane 0 is the nearest chunk
            uint32_t mm = __shfl_sync(FULL, m, src);
            ls = p16 - 16ull * (src + 1) + (31 - __clz(mm)) + 1;
        } else {
            ls = 0;
            for (unsigned long long q = p16 >= 256 ? p16 - 256 : 0; q > 0; q = q >= 512 ? q - 512 : 0) {   // q: exclusive upper end, 16-aligned
                const uint32_t mq = q >= 16ull * (lane + 1) ? newline_mask16(*reinterpret_cast<const uint4 *>(text + q - 16ull * (lane + 1))) : 0u;
                const unsigned hq = __ballot_sync(FULL, mq != 0);
                if (hq) {
                    const int src = __ffs(hq) - 1;                   // lane 0 holds the nearest chunk
                    ls = q - 16ull * (src + 1) + (31 - __clz(__shfl_sync(FULL, mq, src))) + 1;
                    break;
                }
            }
        }
        const uint32_t stride = ud.stride, acc = ud.accept_state;
        PM_TR(1, i, 3);
        if (lit) {
            PM_TR(1, i, 5);
            if (lane == 0) {
                unsigned long long at = atomicAdd(&ctr->n_raw, 1ull);
                if (at < raw_cap)
                    raw[at] = (ls << kPatternBits) | pat;
                else
                    ctr->raw_overflow = 1;
            }
            __syncwarp();
            continue;
        }
        const unsigned long long lend = strip_cr(text, pt, ls, le);
        // ---- the walk (class map and first piece of the line travel together)
        const uint8_t *gcm = blob + h.off_classmaps + ud.classmap_off;
        uint32_t st = staged8  ? warp_walk_line<2>(text, ls, lend, reinterpret_cast<const uint16_t *>(stab), stride, acc, ud.dead_state, ud.mid_state, gcm,
                                                   s_cmap[warp], s_cls[warp])
                      : staged ? warp_walk_line<1>(text, ls, lend, reinterpret_cast<const uint16_t *>(stab), stride, acc, ud.dead_state, ud.mid_state, gcm,
                                                   s_cmap[warp], s_cls[warp])
                               : warp_walk_line<0>(text, ls, lend, gtab, stride, acc, ud.dead_state, ud.mid_state, gcm, s_cmap[warp], s_cls[warp]);
        if (st != acc) st = __ldg(gtab + (st * stride + stride - 1));
        PM_TR(1, i, 4);
        PM_TV(1, i, 6, lend - ls);
        PM_TV(1, i, 7, ((unsigned long long)tbytes << 1) | ((staged || staged8) ? 1u : 0u));
        if (lane == 0) {
            if (st == acc) {
                unsigned long long at = atomicAdd(&ctr->n_raw, 1ull);
                if (at < raw_cap)
                    raw[at] = (ls << kPatternBits) | pat;
                else
                    ctr->raw_overflow = 1;
            }
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------ k_gram_profile
// Occurrence counts of literal grams in a text sample.  Streams the sample like k_scan (one
// aligned word per filter probe position) and looks every word up in an open-addressing table
// that lives in L2; hits are rare, so the atomics are cheap.
__global__ void __launch_bounds__(256) k_gram_profile(const uint4 *__restrict__ text, unsigned long long n_groups, const uint32_t *__restrict__ keys,
                                                      uint32_t log2_cap, uint32_t fold, unsigned int *__restrict__ counts) {
    const uint32_t mask = (1u << log2_cap) - 1u;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_groups; i += (unsigned long long)gridDim.x * blockDim.x) {
        uint4 v = ld_stream16(text + i);
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            uint32_t g = fold ? (w[k] | 0x20202020u) : w[k];
            uint32_t slot = gram_table_slot(g, log2_cap), key;
            while ((key = __ldg(keys + slot)) != g && key != kGramEmpty) slot = (slot + 1) & mask;
            if (key == g) atomicAdd(counts + slot, 1u);
        }
    }
}

// ------------------------------------------------------------------------------------ k_dfa_scan
// Combined search DFA of the patterns that have no 7-byte literal (the kernel north_star
// sketches: multi-pattern DFA, transition table in shared memory, overlapping chunks).
//
//  * One persistent CTA per SM owns a contiguous range of 64 KiB tiles and streams them through
//    a double buffer with cp.async (16-byte pieces, 128-byte rows XOR-swizzled so that the 32
//    lanes of a warp, each walking its own 128-byte row, read conflict-free 128-bit pieces).
//  * The table has END-OF-LINE folded into the '\n' byte (pm_library.cc, build_scan_table), so
//    the walk has no newline branch and a thread's chunk may start anywhere in a line.
//  * A thread does not know the DFA state at the start of its chunk.  It guesses: mid-line idle
//    state, warmed up over the 16 bytes in front of the chunk (the overlap; 32 cost 10 % more time
//    and bought no fewer re-walks on log text).  After the walk the
//    guesses are validated against the true exit state of the previous chunk, in order; a wrong
//    guess re-walks that one chunk.  Exact, and on log text the guesses are almost always right.
//  * The hot walk only tracks "was any accepting state seen".  Chunks where one was seen (rare)
//    are walked once more from their validated entry state to emit (line_start, pattern) events.
//  * CTA ranges are cut at line starts (a CTA finishes the line that crosses its upper bound), so
//    the first chunk of a range starts in the exact start-of-line state.
constexpr int kDfaThreads = 512, kDfaChunk = 128, kDfaTile = kDfaThreads * kDfaChunk, kDfaWarm = 16;

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;"); }

// byte `off` of row `row` in a swizzled tile buffer
__device__ __forceinline__ uint32_t dfa_tile_byte(const uint8_t *buf, int row, int off) {
    return buf[row * kDfaChunk + ((((off >> 4) ^ (row & 7))) << 4) + (off & 15)];
}

struct DfaCtx {
    const uint16_t *tab;      // shared memory, entries premultiplied by the row stride
    const uint8_t *cmap;      // shared memory (compressed form only)
    const uint8_t *text;      // global, for the rare look-ahead / look-behind
    unsigned long long tile_start;
    uint32_t first_accept_p;  // first accepting state, premultiplied
    PodTable pt;
};

template <bool WIDE>
__device__ __forceinline__ uint32_t dfa_next(const DfaCtx &c, uint32_t s, uint32_t b) {
    return WIDE ? c.tab[s + b] : c.tab[s + c.cmap[b]];
}

// '\r' directly in front of a separator '\n' is not part of the line: feed END-OF-LINE ('\n') for it.
__device__ __noinline__ uint32_t dfa_cr_fix(const DfaCtx &c, uint32_t b, unsigned long long pos) {
    if (b == '\r' && c.text[pos + 1] == '\n' && !(c.pt.any_trailing_cr && is_pod_raw_end(c.pt, pos + 1))) return '\n';
    return b;
}

// 16 bytes of one 128-bit piece through the DFA.  The per-byte "is this the '\r' of a \r\n" test
// is hoisted to one SIMD-in-register check per piece; pieces that do hold a '\r' (none in LF
// logs) take an out-of-line byte loop, which keeps the hot loop small enough for the I-cache.
__device__ __forceinline__ bool piece_has_cr(const uint4 &v) {
    uint32_t x0 = v.x ^ 0x0D0D0D0Du, x1 = v.y ^ 0x0D0D0D0Du, x2 = v.z ^ 0x0D0D0D0Du, x3 = v.w ^ 0x0D0D0D0Du;
    uint32_t z = ((x0 - 0x01010101u) & ~x0) | ((x1 - 0x01010101u) & ~x1) | ((x2 - 0x01010101u) & ~x2) | ((x3 - 0x01010101u) & ~x3);
    return (z & 0x80808080u) != 0;
}

template <bool WIDE>
__device__ __noinline__ void dfa_piece_cr(const DfaCtx &c, uint4 v, unsigned long long pos0, uint32_t &s, uint32_t &smax) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    for (int j = 0; j < 16; j++) {
        uint32_t b = (w[j >> 2] >> ((j & 3) * 8)) & 0xFFu;
        if (b == '\r') b = dfa_cr_fix(c, b, pos0 + j);
        s = dfa_next<WIDE>(c, s, b);
        smax = max(smax, s);
    }
}

template <bool WIDE>
__device__ __forceinline__ void dfa_piece(const DfaCtx &c, const uint4 &v, unsigned long long pos0, uint32_t &s, uint32_t &smax) {
    if (piece_has_cr(v)) {
        dfa_piece_cr<WIDE>(c, v, pos0, s, smax);
        return;
    }
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 16; j++) {
        const uint32_t b = __byte_perm(w[j >> 2], 0u, 0x4440u | (uint32_t)(j & 3));   // one PRMT instead of shift + mask
        s = dfa_next<WIDE>(c, s, b);
        smax = max(smax, s);
    }
}

// Walks bytes [lo, hi) of the thread's row; returns the exit state, accumulates the largest state seen.
// A full row is walked as TWO independent dependency chains (the lookups are latency-bound): the
// first half from `s`, the second half from a guessed entry state (mid-line idle state warmed up
// over the last 16 bytes of the first half).  If the guess differs from the first half's true
// exit state, the second half is walked again -- same speculation as between rows, resolved
// inside the thread.
template <bool WIDE>
__device__ __forceinline__ uint32_t dfa_walk(const DfaCtx &c, const uint8_t *buf, int row, int lo, int hi, uint32_t s, uint32_t &smax, uint32_t mid_p) {
    if (lo == 0 && hi == kDfaChunk) {
        const unsigned long long pos = c.tile_start + (unsigned long long)row * kDfaChunk;
        const uint8_t *rp = buf + row * kDfaChunk;
        const int x = row & 7;
        uint32_t sa = s, sb = mid_p, ma = 0, mb = 0, dummy = 0;
        // chain A: pieces 0..3; chain B beside it: warm-up over piece 3, then pieces 4..6 - and 7 alone
        uint32_t guess_b = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint4 va = *reinterpret_cast<const uint4 *>(rp + ((k ^ x) << 4));
            const uint4 vb = *reinterpret_cast<const uint4 *>(rp + (((k + 3) ^ x) << 4));
            dfa_piece<WIDE>(c, va, pos + k * 16, sa, ma);
            if (k == 0) {
                dfa_piece<WIDE>(c, vb, pos + 3 * 16, sb, dummy);
                guess_b = sb;
            } else {
                dfa_piece<WIDE>(c, vb, pos + (k + 3) * 16, sb, mb);
            }
        }
        {
            const uint4 vb = *reinterpret_cast<const uint4 *>(rp + ((7 ^ x) << 4));
            dfa_piece<WIDE>(c, vb, pos + 7 * 16, sb, mb);
        }
        if (guess_b != sa) {   // mis-speculated half: walk it again from the truth
            sb = sa;
            mb = 0;
            for (int k = 4; k < 8; k++) {
                const uint4 vb = *reinterpret_cast<const uint4 *>(rp + ((k ^ x) << 4));
                dfa_piece<WIDE>(c, vb, pos + k * 16, sb, mb);
            }
        }
        smax = max(smax, max(ma, mb));
        return sb;
    }
    if (((lo | hi) & 15) == 0) {   // whole pieces (the warm-up over the tail of the row in front)
        const unsigned long long pos = c.tile_start + (unsigned long long)row * kDfaChunk;
        const uint8_t *rp = buf + row * kDfaChunk;
        for (int k = lo >> 4; k < (hi >> 4); k++) {
            const uint4 v = *reinterpret_cast<const uint4 *>(rp + ((k ^ (row & 7)) << 4));
            dfa_piece<WIDE>(c, v, pos + k * 16, s, smax);
        }
        return s;
    }
    for (int off = lo; off < hi; off++) {
        uint32_t b = dfa_tile_byte(buf, row, off);
        if (b == '\r') b = dfa_cr_fix(c, b, c.tile_start + (unsigned long long)row * kDfaChunk + off);
        s = dfa_next<WIDE>(c, s, b);
        smax = max(smax, s);
    }
    return s;
}

template <bool WIDE>
__global__ void __launch_bounds__(kDfaThreads, 1) k_dfa_scan(const uint8_t *__restrict__ text, unsigned long long total, const uint8_t *__restrict__ blob, BlobHeader h,
                                                             PodTable pt, int group, unsigned long long *__restrict__ raw, unsigned long long raw_cap,
                                                             ScanCounters *ctr) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint16_t exit_s[kDfaThreads];
    __shared__ unsigned long long own_s[2];
    __shared__ uint32_t carry_s;
    const GroupDesc gd = reinterpret_cast<const GroupDesc *>(blob + h.off_groups)[group];
    const uint16_t *gtab = reinterpret_cast<const uint16_t *>(blob + h.off_tables) + gd.table_off;
    const uint32_t *acc = reinterpret_cast<const uint32_t *>(blob + h.off_group_accept) + gd.accept_off;
    const uint32_t n_entries = (uint32_t)gd.nstates * gd.stride;
    // shared memory: [tile buffer 0][tile buffer 1][table][class map]
    uint8_t *bufs[2] = {smem_raw, smem_raw + kDfaTile};
    uint16_t *tab = reinterpret_cast<uint16_t *>(smem_raw + 2 * kDfaTile);
    uint8_t *cmap = reinterpret_cast<uint8_t *>(tab) + (((size_t)n_entries * 2 + 15) & ~(size_t)15);
    const int tid = threadIdx.x;
    for (uint32_t i = tid; i < (n_entries + 7) / 8; i += kDfaThreads)
        reinterpret_cast<uint4 *>(tab)[i] = reinterpret_cast<const uint4 *>(gtab)[i];   // table start is 16-byte aligned in the blob
    if (!WIDE && tid < 64) reinterpret_cast<uint32_t *>(cmap)[tid] = reinterpret_cast<const uint32_t *>(blob + h.off_classmaps + gd.classmap_off)[tid];

    // ---- this CTA's range of tiles, cut at line starts
    const unsigned long long n_tiles = (total + kDfaTile - 1) / kDfaTile;
    const unsigned long long per = (n_tiles + gridDim.x - 1) / gridDim.x;
    const unsigned long long seg_lo = min((unsigned long long)blockIdx.x * per * kDfaTile, total);
    const unsigned long long seg_hi = min(((unsigned long long)blockIdx.x + 1) * per * kDfaTile, total);
    if (tid < 2) {
        // first line start at or after `x` (x = 0: the buffer starts with '\n' padding, position 0 is fine)
        unsigned long long x = tid == 0 ? seg_lo : seg_hi, q = x;
        if (x > 0 && x < total) {
            q = x - 1;
            while (q < total && text[q] != '\n') q++;
            q = min(q + 1, total);
        }
        own_s[tid] = q;
    }
    if (tid == 0) carry_s = 0;
    __syncthreads();
    const unsigned long long own_lo = own_s[0], own_hi = own_s[1];
    if (own_lo >= own_hi) return;

    DfaCtx c;
    c.tab = tab;
    c.cmap = cmap;
    c.text = text;
    c.pt = pt;
    c.first_accept_p = (uint32_t)gd.first_accept * gd.stride;
    const uint32_t mid_p = (uint32_t)gd.mid_state * gd.stride;

    const unsigned long long t_first = own_lo / kDfaTile, t_last = (own_hi - 1) / kDfaTile;
    auto issue = [&](unsigned long long t, uint8_t *buf) {   // 8 x 16-byte pieces per thread, coalesced, swizzled on the way in
        const unsigned long long ts = t * kDfaTile;
#pragma unroll
        for (int i = 0; i < kDfaTile / 16 / kDfaThreads; i++) {
            const int q = i * kDfaThreads + tid, r = q >> 3, j = q & 7;
            const unsigned long long g = ts + (unsigned long long)q * 16;
            if (g + 16 <= total) cp_async16(buf + r * kDfaChunk + ((j ^ (r & 7)) << 4), text + g);
        }
        cp_async_commit();
    };
    issue(t_first, bufs[0]);
    for (unsigned long long t = t_first; t <= t_last; t++) {
        uint8_t *buf = bufs[(t - t_first) & 1];
        cp_async_wait_all();
        __syncthreads();                                     // tile t landed; everyone is done with the other buffer
        if (t < t_last) issue(t + 1, bufs[(t - t_first + 1) & 1]);
        c.tile_start = t * kDfaTile;
        // ---- phase A: speculative walk of this thread's chunk, clipped to the CTA's range
        const unsigned long long cs = c.tile_start + (unsigned long long)tid * kDfaChunk;
        const unsigned long long a = max(cs, own_lo), b = min(cs + kDfaChunk, own_hi);
        const bool active = a < b;
        const bool exact = active && a == own_lo;            // starts at a line start: state 0 is the truth
        const int lo = active ? (int)(a - cs) : 0, hi = active ? (int)(b - cs) : 0;
        uint32_t guess = 0, ex = 0, smax = 0;
        if (active) {
            if (!exact) {
                uint32_t s = mid_p, dummy = 0;
                if (lo > 0) {                                // clipped chunk that is not the range start cannot happen; keep it safe
                    s = dfa_walk<WIDE>(c, buf, tid, max(0, lo - kDfaWarm), lo, s, dummy, mid_p);
                } else if (tid > 0) {
                    s = dfa_walk<WIDE>(c, buf, tid - 1, kDfaChunk - kDfaWarm, kDfaChunk, s, dummy, mid_p);
                } else {                                     // row 0: the bytes in front of it are in the previous tile
                    for (unsigned long long p = cs - kDfaWarm; p < cs; p++) {
                        uint32_t bb = text[p];
                        if (bb == '\r') bb = dfa_cr_fix(c, bb, p);
                        s = dfa_next<WIDE>(c, s, bb);
                    }
                }
                guess = s;
            }
            ex = dfa_walk<WIDE>(c, buf, tid, lo, hi, guess, smax, mid_p);
        }
        // ---- phase B: validate the guesses in chunk order
        const int t0 = (int)((max(c.tile_start, own_lo) - c.tile_start) / kDfaChunk);   // first active row
        exit_s[tid] = (uint16_t)ex;
        __syncthreads();
        for (;;) {
            bool wrong = false;
            uint32_t truth = 0;
            if (active && !exact) {
                truth = tid == t0 ? carry_s : exit_s[tid - 1];
                wrong = truth != guess;
            }
            if (!__syncthreads_or(wrong)) break;
            if (wrong) {
                guess = truth;
                smax = 0;
                ex = dfa_walk<WIDE>(c, buf, tid, lo, hi, guess, smax, mid_p);
                exit_s[tid] = (uint16_t)ex;
            }
            __syncthreads();
        }
        // last active row's exit state is the next tile's carry
        const int t1 = (int)((min(c.tile_start + (unsigned long long)kDfaTile, own_hi) - 1 - c.tile_start) / kDfaChunk);
        __syncthreads();
        if (tid == t1) carry_s = ex;
        // ---- phase C: emit (rare): walk again from the validated entry state
        if (active && smax >= c.first_accept_p) {
            uint32_t s = guess, last_emit = 0xFFFFFFFFu;
            long long last_nl = -1;                          // position of the last '\n' before the current byte, if seen in this chunk
            for (int off = lo; off < hi; off++) {
                const unsigned long long pos = cs + off;
                uint32_t bb = dfa_tile_byte(buf, tid, off);
                if (bb == '\r') bb = dfa_cr_fix(c, bb, pos);
                s = dfa_next<WIDE>(c, s, bb);
                if (s >= c.first_accept_p && s != last_emit) {
                    last_emit = s;
                    unsigned long long ls;
                    if (last_nl >= 0) {
                        ls = (unsigned long long)last_nl + 1;
                    } else {
                        ls = pos;                            // line start lies before this chunk: look behind in the text
                        while (text[ls - 1] != '\n') ls--;
                    }
                    const uint32_t sid = s / gd.stride - gd.first_accept;
                    for (uint32_t k = acc[sid], ke = acc[sid + 1]; k < ke; k++) {
                        unsigned long long at = atomicAdd(&ctr->n_raw, 1ull);
                        if (at < raw_cap)
                            raw[at] = (ls << kPatternBits) | acc[k];
                        else
                            ctr->raw_overflow = 1;
                    }
                }
                if (bb == '\n') {
                    last_nl = (long long)pos;
                    if (text[pos] == '\r') last_nl = (long long)pos + 1;   // CRLF: the '\r' stood in for the separator
                    last_emit = 0xFFFFFFFFu;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------ unique
__global__ void k_mark_unique(const unsigned long long *sorted, unsigned long long n, unsigned char *flags) {
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = (sorted[i] != ~0ull) && (i == 0 || sorted[i] != sorted[i - 1]);
}

// Ordering + de-duplication of the raw events without host involvement (<= kMediumSort keys; hits
// are ~1e-3 of lines, and 1e-2 of the lines of a 1 GiB batch still fit).  Every launch reads its size
// from the device counters, so the host need not learn the event count before it can enqueue the
// rest of the pipeline:
//   k_sort_chunks   one CTA per 1024-key chunk that holds keys: bitonic sort in shared memory
//   k_sort_merge    n <= kSmallSort: every key finds its global rank by binary search in the other
//                   (at most 15) chunks, all of them in shared memory - one launch
//   k_merge_all     n >  kSmallSort: seven rounds of pairwise run merges (1024 -> 131072 keys) in one launch, every
//                   key finds its place in the merged run by one binary search in the sibling run
//   k_sort_unique   one CTA drops duplicates and counts
// The launches of the path that does not apply return at once.
constexpr int kSortChunk = 1024, kSortChunksSmall = kSmallSort / kSortChunk, kSortChunksMedium = kMediumSort / kSortChunk, kMergePasses = 7;
static_assert((kSortChunk << kMergePasses) == kMediumSort, "merge passes must take 1024-key runs to kMediumSort");

__device__ __forceinline__ bool sort_too_big(const ScanCounters *ctr, unsigned long long raw_cap) {
    return ctr->cand_overflow || ctr->raw_overflow || ctr->n_raw > (unsigned long long)kMediumSort || ctr->n_raw > raw_cap;
}

__global__ void __launch_bounds__(kSortChunk / 2) k_sort_chunks(const unsigned long long *__restrict__ raw, const ScanCounters *ctr, unsigned long long raw_cap,
                                                                unsigned long long *__restrict__ chunks) {
    __shared__ unsigned long long keys[kSortChunk];
    if (sort_too_big(ctr, raw_cap)) return;
    const unsigned long long n = ctr->n_raw, base = (unsigned long long)blockIdx.x * kSortChunk;
    if (base >= n) return;   // no key in this chunk: the later stages stop at the last chunk that has one
    for (unsigned i = threadIdx.x; i < kSortChunk; i += blockDim.x) keys[i] = base + i < n ? raw[base + i] : ~0ull;
    __syncthreads();
    const unsigned t = threadIdx.x;
    for (unsigned k = 2; k <= kSortChunk; k <<= 1)
        for (unsigned j = k >> 1; j > 0; j >>= 1) {
            unsigned i = 2 * t - (t & (j - 1)), l = i | j;
            unsigned long long a = keys[i], b = keys[l];
            if ((a > b) == ((i & k) == 0)) {
                keys[i] = b;
                keys[l] = a;
            }
            __syncthreads();
        }
    for (unsigned i = threadIdx.x; i < kSortChunk; i += blockDim.x) chunks[base + i] = keys[i];
}

__global__ void __launch_bounds__(1024) k_sort_merge(const unsigned long long *__restrict__ chunks, const ScanCounters *ctr, unsigned long long raw_cap,
                                                    unsigned long long *__restrict__ sorted) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned long long *sk = reinterpret_cast<unsigned long long *>(smem_raw);   // all sorted chunks (<= 128 KiB)
    if (sort_too_big(ctr, raw_cap)) return;
    const unsigned long long n = ctr->n_raw;
    if (n > (unsigned long long)kSmallSort) return;   // k_merge_all takes it
    const unsigned n_chunks = (unsigned)((n + kSortChunk - 1) / kSortChunk);
    if (blockIdx.x * blockDim.x >= n_chunks * kSortChunk) return;
    for (unsigned k = threadIdx.x; k < n_chunks * kSortChunk / 2; k += blockDim.x) reinterpret_cast<uint4 *>(sk)[k] = reinterpret_cast<const uint4 *>(chunks)[k];
    __syncthreads();
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned c = i / kSortChunk;
    const unsigned long long key = sk[i];
    // rank = own index + for every other chunk o the number of its keys that sort before this one
    // (strictly smaller, or equal with o < c: stable); binary searches in shared memory
    unsigned rank = i % kSortChunk;
    for (unsigned o = 0; o < n_chunks; o++) {
        if (o == c) continue;
        const unsigned long long *q = sk + (size_t)o * kSortChunk;
        unsigned lo = 0, hi = kSortChunk;
        while (lo < hi) {
            unsigned mid = (lo + hi) >> 1;
            unsigned long long v = q[mid];
            if (v < key || (v == key && o < c)) lo = mid + 1; else hi = mid;
        }
        rank += lo;
    }
    sorted[rank] = key;
}

// All seven rounds of pairwise run merges (1024 -> 131072 keys) in ONE launch, for 16384 < n <= 131072 keys; with
// fewer keys the launch returns at once.  (As seven launches the rounds cost the common, small case 10 us: the sort
// runs beside k_plan / k_check and is joined before k_finalize, and every launch on the busy GPU, even one with
// nothing to do, adds its latency to that chain.)  A round: key i of run r goes to (start of the pair) + (its index in
// r) + (number of keys of the sibling run that sort before it); ties go to the left run, so equal keys never collide.
// The P = ceil(n / 1024) * 1024 keys that exist take part (the tail of the last chunk is +inf padding).  Rounds are
// separated by a grid-wide barrier on a counter in the ScanCounters block (zeroed with it before every scan); the
// grid is 128 CTAs of 256 threads, far below what the GPU keeps resident, and the kernels it shares the GPU with
// finish on their own, so every CTA of this grid gets to run while the others wait.
constexpr int kMergeCtas = 128, kMergeThreads = 256;

__global__ void __launch_bounds__(kMergeThreads) k_merge_all(unsigned long long *__restrict__ buf_a, unsigned long long *__restrict__ buf_b, ScanCounters *ctr,
                                                           unsigned long long raw_cap) {
    if (sort_too_big(ctr, raw_cap)) return;
    const unsigned long long n = ctr->n_raw;
    if (n <= (unsigned long long)kSmallSort) return;   // k_sort_merge took it
    const unsigned P = (unsigned)((n + kSortChunk - 1) / kSortChunk) * kSortChunk;
    unsigned int *barrier = &ctr->merge_barrier;
    for (int pass = 0; pass < kMergePasses; pass++) {
        const unsigned long long *src = pass & 1 ? buf_b : buf_a;
        unsigned long long *dst = pass & 1 ? buf_a : buf_b;
        const unsigned run = (unsigned)kSortChunk << pass;
        for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
            const unsigned long long key = __ldcg(src + i);   // written by other CTAs in the round before: read through L2
            const unsigned r = i / run, sib = (r ^ 1u) * run;
            if (sib >= P) {   // no sibling run: this run moves as it is
                dst[i] = key;
                continue;
            }
            const unsigned sib_end = min(sib + run, P);
            unsigned lo = sib, hi = sib_end;
            if (r & 1u) {   // right run: sibling keys <= key come first
                while (lo < hi) {
                    const unsigned mid = (lo + hi) >> 1;
                    if (__ldcg(src + mid) <= key) lo = mid + 1; else hi = mid;
                }
            } else {        // left run: sibling keys < key come first
                while (lo < hi) {
                    const unsigned mid = (lo + hi) >> 1;
                    if (__ldcg(src + mid) < key) lo = mid + 1; else hi = mid;
                }
            }
            dst[(r & ~1u) * run + (i - r * run) + (lo - sib)] = key;
        }
        // grid-wide barrier
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(barrier, 1u);
            const unsigned want = (unsigned)(pass + 1) * gridDim.x;
            while (atomicAdd(barrier, 0u) < want) __nanosleep(64);
            __threadfence();
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(1024) k_sort_unique(const unsigned long long *__restrict__ sorted, const ScanCounters *ctr, unsigned long long raw_cap,
                                                      unsigned long long *__restrict__ uniq, unsigned long long *n_unique, unsigned int *big_flag) {
    if (sort_too_big(ctr, raw_cap)) {
        if (threadIdx.x == 0) {
            *n_unique = 0;
            *big_flag = 1;
        }
        return;
    }
    const unsigned n = (unsigned)ctr->n_raw;
    // thread t owns `per` consecutive keys: one block scan of the per-thread counts, then each thread writes its survivors
    const unsigned per = (n + 1023u) / 1024u;
    const unsigned i0 = threadIdx.x * per;
    unsigned long long prev = (i0 > 0 && i0 - 1 < n) ? sorted[i0 - 1] : ~0ull;
    const unsigned long long first_prev = prev;
    unsigned cnt = 0;
    for (unsigned r = 0; r < per && i0 + r < n; r++) {
        const unsigned long long k = sorted[i0 + r];
        cnt += (k != ~0ull && (i0 + r == 0 || k != prev));
        prev = k;
    }
    __shared__ unsigned long long tot;
    unsigned long long at = block_exclusive_scan(cnt, &tot);
    prev = first_prev;
    for (unsigned r = 0; r < per && i0 + r < n; r++) {
        const unsigned long long k = sorted[i0 + r];
        if (k != ~0ull && (i0 + r == 0 || k != prev)) uniq[at++] = k;
        prev = k;
    }
    if (threadIdx.x == 0) {
        *n_unique = tot;
        *big_flag = 0;
    }
}

// Rank of every raw key in the sorted, de-duplicated array.  Runs on the sort's stream, beside
// k_plan / k_check, so that k_finalize (on the main chain) reads its row index instead of doing a
// 13-step binary search through L2 per key.
__global__ void __launch_bounds__(256) k_rank(const unsigned long long *__restrict__ raw, const ScanCounters *ctr, unsigned long long raw_cap,
                                              const unsigned long long *__restrict__ uniq, const unsigned long long *n_unique_ptr, uint32_t *__restrict__ rank) {
    if (sort_too_big(ctr, raw_cap)) return;
    const unsigned n = (unsigned)ctr->n_raw, nu = (unsigned)*n_unique_ptr;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned long long key = raw[i];
        unsigned lo = 0, hi = nu;   // lower bound; the key is present
        while (lo < hi) {
            const unsigned mid = (lo + hi) >> 1;
            if (uniq[mid] < key) lo = mid + 1; else hi = mid;
        }
        rank[i] = lo;
    }
}

// ------------------------------------------------------------------------------------ resolve (a12 + a13)
// Three kernels, so that the expensive part -- walking context lines through the STACK and
// secondary-pattern DFAs -- is balanced over the whole GPU instead of serialised per event:
//   k_plan       warp per event: line number, context geometry, one work item per
//                (event, unit, line) that needs a DFA verdict
//   k_check      thread per item (short lines); k_check_long: warp per item with a speculative
//                chunked walk (every lane guesses its entry state, guesses are validated in order)
//   k_finalize   thread per event: STACK run length, nearest secondary hit, fp64 score
constexpr int kPlanWarps = 4;
constexpr int kCtxSlots = kMaxCtxLines + 2;
constexpr uint32_t kItemInvalid = 0xFFFFFFFFu;
constexpr uint32_t kLongLine = 256;      // lines above this are walked by a whole warp

struct Item {
    unsigned long long a;    // line start in text
    uint32_t len;            // kItemInvalid: the line does not exist (outside the pod)
    uint32_t unit;
};

struct Group {                           // a run of items that share a unit: the STACK run of an event, or one secondary's window
    unsigned long long first;            // index of the run's first item
    uint32_t count;
    uint32_t unit;
};

struct EvAux {                           // what k_plan learns about one key, indexed like keys[] (64 bytes)
    unsigned long long item_base;
    int32_t n_stack, na;
    int32_t pod, line_no, line_len, ctx_first, ctx_last, pad;
    int64_t line_start, ctx_start, ctx_end;
};

// Collects positions of the next `need` '\n' bytes at positions >= from into out[0..need).
// Four 512-byte steps are loaded per round trip (the text around a hit is cold in L2/DRAM).
__device__ void warp_collect_forward(const uint8_t *text, unsigned long long from, int need, unsigned long long *out) {
    const int lane = threadIdx.x & 31;
    unsigned long long a0 = from & ~15ull;
    int got = 0;
    while (got < need) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) v[u] = *reinterpret_cast<const uint4 *>(text + a0 + u * 512 + (unsigned long long)lane * 16);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (got >= need) break;
            unsigned long long a = a0 + u * 512 + (unsigned long long)lane * 16;
            uint32_t m = newline_mask16(v[u]);
            if (a < from) m &= (from - a >= 16) ? 0u : ~((1u << (from - a)) - 1u);
            int c = __popc(m);
            int inc = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(FULL, inc, o);
                if (lane >= o) inc += y;
            }
            int rank = got + inc - c;
            while (m && rank < need) {
                int bpos = __ffs(m) - 1;
                m &= m - 1;
                out[rank++] = a + bpos;
            }
            got += __shfl_sync(FULL, inc, 31);
        }
        a0 += 2048;
    }
    __syncwarp();
}

// Collects positions of the previous `need` '\n' bytes at positions < before into out[0..need),
// nearest first.
__device__ void warp_collect_backward(const uint8_t *text, unsigned long long before, int need, unsigned long long *out) {
    const int lane = threadIdx.x & 31;
    unsigned long long top = (before + 15) & ~15ull;   // first 16-byte boundary >= before
    int got = 0;
    while (got < need) {
        long long a = (long long)top - 16ll * (lane + 1);
        uint32_t m = 0;
        if (a >= 0) {
            uint4 v = *reinterpret_cast<const uint4 *>(text + a);
            m = newline_mask16(v);
            if ((unsigned long long)a + 16 > before) m &= (1u << (before - (unsigned long long)a)) - 1u;
        }
        int c = __popc(m);
        int inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += y;
        }
        int rank = got + inc - c;
        while (m && rank < need) {
            int bpos = 31 - __clz(m);
            m &= ~(1u << bpos);
            out[rank++] = (unsigned long long)a + bpos;
        }
        got += __shfl_sync(FULL, inc, 31);
        if (top < 512) break;
        top -= 512;
    }
    __syncwarp();
}

// (StackFirst: pm_kernels.h - the bytes a STACK line can start with)
__global__ void __launch_bounds__(kPlanWarps * 32) k_plan(const uint8_t *__restrict__ text, const uint8_t *__restrict__ blob, BlobHeader h,
                                                          const unsigned long long *__restrict__ tile_prefix, PodTable pt,
                                                          const unsigned long long *__restrict__ keys, const unsigned long long *n_keys_ptr,
                                                          unsigned long long key_cap,
                                                          Item *__restrict__ items, unsigned long long item_cap, uint32_t *__restrict__ long_list,
                                                          unsigned long long long_cap, EvAux *__restrict__ aux, ResolveCounters *ctr, int slots,
                                                          const StackFirst stack_first, Group *__restrict__ groups, unsigned long long group_cap) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long *fwd = reinterpret_cast<unsigned long long *>(smem_raw) + (size_t)warp * 2 * slots;   // fwd[i]: '\n' that ends line L+i
    unsigned long long *bwd = fwd + slots;                                                                   // bwd[i]: '\n' that precedes line L-i
    const PatternDesc *pds = reinterpret_cast<const PatternDesc *>(blob + h.off_patterns);
    const SecDesc *sds = reinterpret_cast<const SecDesc *>(blob + h.off_secondaries);
    const PodDev *pods = pt.pods;
    // keys[] may be the raw (unsorted, possibly duplicated) event queue: a duplicate plans the
    // same work twice and k_finalize writes the same row twice.  More keys than key_cap means
    // the caller will come back through the sorted path: nothing to do here.
    const unsigned long long n_events = *n_keys_ptr;
    if (n_events > key_cap) {
        if (threadIdx.x == 0 && blockIdx.x == 0) ctr->overflow = 1;
        return;
    }

    PM_TR(2, (unsigned long long)blockIdx.x * kPlanWarps + warp, 0);
    for (unsigned long long ev = (unsigned long long)blockIdx.x * kPlanWarps + warp; ev < n_events; ev += (unsigned long long)gridDim.x * kPlanWarps) {
        PM_TR(2, ev, 1);
        const unsigned long long key = keys[ev], ls = key >> kPatternBits;
        const uint32_t pat = (uint32_t)(key & ((1u << kPatternBits) - 1u));
        const PatternDesc pd = pds[pat];
        int lo = 0, hi = pt.n - 1;   // pod = last pod with start <= ls
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (pods[mid].start <= ls) lo = mid; else hi = mid - 1;
        }
        const PodDev pod = pods[lo];
        PM_TR(2, ev, 2);
        const long long L = (long long)(warp_newlines_before(text, tile_prefix, ls) - pod.nl_before) + 1;
        PM_TR(2, ev, 3);
        const long long n_lines = pod.n_lines;
        const int nb = (int)min((long long)pd.lines_before, L - 1);
        const int na = (int)min((long long)pd.lines_after, n_lines - L);
        int stack_cap = pd.include_stack ? (int)min((long long)kStackMax, n_lines - L) : 0;
        int need_f = na, need_b = nb;
        unsigned long long n_items = 0;
        for (uint32_t j = 0; j < pd.sec_count; j++) {
            int w = sds[pd.sec_first + j].window;
            need_f = max(need_f, (int)min((long long)w, n_lines - L));
            need_b = max(need_b, (int)min((long long)w, L - 1));
            n_items += 2ull * w + 1;
        }
        warp_collect_forward(text, ls, need_f + 1, fwd);
        if (stack_cap > need_f) {
            // The STACK extension (SPEC.md section 6) wants up to 50 more lines, ~7 KB of text per event - but only while
            // they are stack-trace lines.  Its regex is anchored at the line start, so the first byte of the first candidate
            // line decides whether the run can start at all: most hit lines are not followed by a frame, and then neither
            // the lines are collected nor the 50 work items created (half of k_plan's instructions and of k_check's items).
            __syncwarp();
            const unsigned long long q = fwd[na] + 1;   // start of line L + na + 1 (it exists: stack_cap > na)
            const uint32_t b0 = text[q];
            if ((stack_first.bits[b0 >> 5] >> (b0 & 31u)) & 1u) {
                need_f = stack_cap;
                __syncwarp();
                warp_collect_forward(text, ls, need_f + 1, fwd);
            } else {
                stack_cap = 0;
            }
        }
        const int n_stack = max(0, stack_cap - na);
        n_items += (unsigned long long)n_stack;
        PM_TR(2, ev, 4);
        if (lane == 0) bwd[0] = ls - 1;
        if (need_b > 0) warp_collect_backward(text, ls - 1, need_b, bwd + 1);
        __syncwarp();
        PM_TR(2, ev, 5);

        auto line_span = [&](int rel, unsigned long long &a, unsigned long long &b) {   // rel: line L+rel
            if (rel >= 0) {
                a = rel == 0 ? ls : fwd[rel - 1] + 1;
                b = fwd[rel];
            } else {
                a = bwd[-rel] + 1;
                b = bwd[-rel - 1];
            }
            if (b > a && text[b - 1] == '\r' && !(pt.any_trailing_cr && b == pod.raw_end)) b--;
        };

        unsigned long long base = 0;
        if (lane == 0 && n_items) base = atomicAdd(&ctr->n_items, n_items);
        base = __shfl_sync(FULL, base, 0);
        // the items come in runs that share a unit (the STACK run, one window per secondary): k_check takes a warp per run
        // (at most 32 items per run, one per lane of k_check's warp: longer runs are cut)
        const uint32_t stack_grp = ((uint32_t)n_stack + 31u) / 32u;
        uint32_t n_grp = stack_grp;
        for (uint32_t j = 0; j < pd.sec_count; j++) n_grp += (2u * (uint32_t)sds[pd.sec_first + j].window + 1u + 31u) / 32u;
        unsigned gbase = 0;
        if (lane == 0 && n_grp) gbase = atomicAdd(&ctr->n_groups, n_grp);
        gbase = __shfl_sync(FULL, gbase, 0);
        const bool fits = base + n_items <= item_cap && (unsigned long long)gbase + n_grp <= group_cap;
        if (!fits && lane == 0) ctr->overflow = 1;

        if (fits) {
            if ((uint32_t)lane < stack_grp) groups[gbase + lane] = Group{base + 32ull * lane, min(32u, (uint32_t)n_stack - 32u * lane), h.stack_unit};
            unsigned gnext = gbase + stack_grp;
            // STACK items: lines L+na+1 .. L+stack_cap
            for (int k = lane; k < n_stack; k += 32) {
                unsigned long long a, b;
                line_span(na + 1 + k, a, b);
                Item it{a, (uint32_t)(b - a), h.stack_unit};
                items[base + k] = it;
                if (it.len > kLongLine) {
                    unsigned long long at = atomicAdd(&ctr->n_long, 1ull);
                    if (at < long_cap) long_list[at] = (uint32_t)(base + k); else ctr->overflow = 1;
                }
            }
            // secondary items: idx 0..2w, distance d = (idx+1)/2, odd idx = before, even = after
            unsigned long long off = base + n_stack;
            for (uint32_t j = 0; j < pd.sec_count; j++) {
                const SecDesc sd = sds[pd.sec_first + j];
                const int cnt = 2 * sd.window + 1;
                const uint32_t pieces = ((uint32_t)cnt + 31u) / 32u;
                for (uint32_t q = lane; q < pieces; q += 32) groups[gnext + q] = Group{off + 32ull * q, min(32u, (uint32_t)cnt - 32u * q), sd.unit};
                gnext += pieces;
                for (int idx = lane; idx < cnt; idx += 32) {
                    int d = (idx + 1) >> 1;
                    int rel = (idx & 1) ? -d : d;
                    bool valid = rel >= 0 ? (d <= n_lines - L) : (d <= L - 1);
                    Item it{0, kItemInvalid, sd.unit};
                    if (valid) {
                        unsigned long long a, b;
                        line_span(rel, a, b);
                        it.a = a;
                        it.len = (uint32_t)(b - a);
                    }
                    items[off + idx] = it;
                    if (valid && it.len > kLongLine) {
                        unsigned long long at = atomicAdd(&ctr->n_long, 1ull);
                        if (at < long_cap) long_list[at] = (uint32_t)(off + idx); else ctr->overflow = 1;
                    }
                }
                off += cnt;
            }
        }
        if (lane == 0) {
            unsigned long long la, lb, ca, cb, ea, eb;
            line_span(0, la, lb);
            line_span(-nb, ca, cb);
            line_span(na, ea, eb);
            EvAux ax;
            ax.item_base = base;
            ax.n_stack = n_stack;
            ax.na = na;
            ax.pod = lo;
            ax.line_no = (int32_t)L;
            ax.line_len = (int32_t)(lb - la);
            ax.ctx_first = (int32_t)(L - nb);
            ax.ctx_last = (int32_t)(L + na);                 // k_finalize adds the STACK run
            ax.pad = 0;
            ax.line_start = (int64_t)(ls - pod.start);
            ax.ctx_start = (int64_t)(ca - pod.start);
            ax.ctx_end = (int64_t)(eb - pod.start);
            aux[ev] = ax;
        }
        __syncwarp();
    }
}

// One thread walks one line [a, b) through the warp's shared-memory copy of a unit's table (one byte per entry) and class
// map, then feeds END-OF-LINE.  The text comes in aligned 16-byte pieces; the class of every byte of a piece is looked up
// before the walk, so that what remains per byte is one multiply-add and one shared-memory read.
__device__ __forceinline__ uint32_t walk_staged(const uint8_t *__restrict__ text, unsigned long long a, unsigned long long b, uint32_t s, const uint8_t *tab,
                                                const uint8_t *cm, uint32_t stride, uint32_t acc, uint32_t dead, unsigned long long tr_group = ~0ull) {
    const uint32_t tab_s = (uint32_t)__cvta_generic_to_shared(tab);
    unsigned long long p = a & ~15ull;
#ifdef PM_TRACE
    const long long tr_c0 = clock64();
    const unsigned long long tr_p0 = p;
#endif
    {   // The lanes walk 32 different lines that were last touched by k_plan: a piece load of the warp waits for the slowest of
        // 32 separate L2/DRAM accesses, ~2 000 cycles where the 16 lookups it feeds take 650 (tools_warp_trace.py).  So every
        // lane asks for its whole range first (at most nine 32-byte sectors), all at once and none waited for.
        uint32_t sink = 0;
        for (unsigned long long q = (a & ~31ull) + 32; q < b; q += 32) sink ^= __ldg(reinterpret_cast<const uint32_t *>(text + q));
        asm volatile("" ::"r"(sink));
    }
    uint4 v = *reinterpret_cast<const uint4 *>(text + p);
    for (; p < b && s != acc && s != dead; p += 16) {
        const uint32_t x[4] = {v.x, v.y, v.z, v.w};
        // the next piece is requested before this one is walked: the lanes walk 32 different lines and cross into new cache
        // lines at different steps, so nearly every piece load of the warp has some lane waiting for L2 or DRAM
        if (p + 16 < b) v = *reinterpret_cast<const uint4 *>(text + p + 16);
        uint32_t c[16];   // the column of each byte as a shared-memory address: the row offset is all the walk adds
#pragma unroll
        for (int j = 0; j < 16; j++) c[j] = tab_s + cm[(x[j >> 2] >> (8 * (j & 3))) & 0xFFu];
        // ONE code path for whole and partial pieces (bytes outside the range are looked up like the others and the result
        // dropped): the lanes are at different places of their lines, and with a separate path for the first and last piece
        // the warp ran both paths in nearly every iteration - 2 000 cycles per piece instead of 700
        const int lo = p >= a ? 0 : (int)(a - p), hi = p + 16 <= b ? 16 : (int)(b - p);
        const uint32_t valid = (0xFFFFu >> (16 - hi)) & (0xFFFFu << lo);
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const uint32_t nx = lds_u8(s * stride + c[j]);
            s = ((valid >> j) & 1u) ? nx : s;
        }
    }
#ifdef PM_TRACE
    if ((threadIdx.x & 31) == 0 && tr_group != ~0ull) {
        PM_TV(3, tr_group, 4, (unsigned long long)(clock64() - tr_c0));
        PM_TV(3, tr_group, 5, (p - tr_p0) / 16);
    }
#endif
    return s;
}

__device__ __forceinline__ bool line_matches_staged(const uint8_t *__restrict__ text, unsigned long long a, unsigned long long b, const uint8_t *tab,
                                                    const uint8_t *cm, uint32_t stride, uint32_t acc, uint32_t dead, unsigned long long tr_group = ~0ull) {
    const uint32_t s = walk_staged(text, a, b, 0, tab, cm, stride, acc, dead, tr_group);
    if (s == acc) return true;
    if (s == dead) return false;
    return tab[s * stride + stride - 1] == acc;
}

// The warp's copy of a unit's table (one byte per entry) and class map in shared memory; false if the unit has more than 256
// states or more entries than the budget.  Tables start 16-byte aligned in the blob and are followed by at least their padding.
constexpr uint32_t kCheckTabCap = 1536;   // table entries a warp can stage (the context units of a crash-log library have 150-1400)
__device__ __forceinline__ bool warp_stage_unit(const uint8_t *__restrict__ blob, const BlobHeader &h, const UnitDesc &ud, uint8_t *cm_s, uint8_t *tb_s) {
    const int lane = threadIdx.x & 31;
    const uint32_t entries = (uint32_t)ud.nstates * ud.stride;
    if (ud.nstates > 256 || entries > kCheckTabCap) return false;   // warp-uniform
    const uint32_t *gcm = reinterpret_cast<const uint32_t *>(blob + h.off_classmaps + ud.classmap_off);
    reinterpret_cast<uint32_t *>(cm_s)[lane] = __ldg(gcm + lane);
    reinterpret_cast<uint32_t *>(cm_s)[32 + lane] = __ldg(gcm + 32 + lane);
    const uint4 *gt = reinterpret_cast<const uint4 *>(reinterpret_cast<const uint16_t *>(blob + h.off_tables) + ud.table_off);
#pragma unroll 2
    for (uint32_t e = lane * 8u; e < entries; e += 256u) {   // 8 entries per lane and step
        const uint4 q = __ldg(gt + (e >> 3));
        uint2 o;
        o.x = __byte_perm(q.x, q.y, 0x6420);
        o.y = __byte_perm(q.z, q.w, 0x6420);
        *reinterpret_cast<uint2 *>(tb_s + e) = o;
    }
    __syncwarp();
    return true;
}

// The context lines of the events: one WARP per run of items that share a unit (k_plan's groups), one lane per line.
// A line walk is a chain of dependent table reads; with the tables in global memory a step cost ~130 cycles (the 32
// lanes of a warp stand in different rows and columns of a 256-column table: several L1 lines per step, and the warps
// of an SM do not share tables), 15 us for the longest line of a warp (tools_warp_trace.py).  So the warp first copies its
// unit's table - one byte per entry, which every unit with at most 256 states allows - and class map into shared memory.
constexpr int kCheckWarps = 8;

__global__ void __launch_bounds__(kCheckWarps * 32) k_check(const uint8_t *__restrict__ text, const uint8_t *__restrict__ blob, BlobHeader h,
                                                           const Item *__restrict__ items, const Group *__restrict__ groups, unsigned long long group_cap,
                                                           unsigned char *__restrict__ results, const ResolveCounters *ctr) {
    __shared__ __align__(16) uint8_t s_cm[kCheckWarps][256];
    __shared__ __align__(16) uint8_t s_tb[kCheckWarps][kCheckTabCap];
    if (ctr->overflow) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long n = ctr->n_groups < group_cap ? ctr->n_groups : group_cap;
    const UnitDesc *units = reinterpret_cast<const UnitDesc *>(blob + h.off_units);
    PM_TR(3, (unsigned long long)blockIdx.x * kCheckWarps + warp, 0);
    for (unsigned long long g = (unsigned long long)blockIdx.x * kCheckWarps + warp; g < n; g += (unsigned long long)gridDim.x * kCheckWarps) {
        const Group gr = groups[g];
        Item it{0, kItemInvalid, 0};
        if ((uint32_t)lane < gr.count) it = items[gr.first + lane];   // under the table copy
        const UnitDesc ud = units[gr.unit];
        PM_TR(3, g, 1);
        const uint32_t entries = (uint32_t)ud.nstates * ud.stride;
        (void)entries;
        __syncwarp();   // the previous run's walks are over
        const bool staged = warp_stage_unit(blob, h, ud, s_cm[warp], s_tb[warp]);   // warp-uniform
        PM_TR(3, g, 2);
        if ((uint32_t)lane < gr.count) {   // a run has at most 32 items (k_plan)
            const unsigned long long i = gr.first + lane;
            if (it.len == kItemInvalid) {
                results[i] = 0;
            } else if (it.len <= kLongLine) {
                if (staged) {
                    results[i] = line_matches_staged(text, it.a, it.a + it.len, s_tb[warp], s_cm[warp], ud.stride, ud.accept_state, ud.dead_state, g);
                } else {
                    UnitTables ut = unit_tables(blob, h, it.unit);
                    results[i] = line_matches(text, it.a, it.a + it.len, ut);
                }
            }
        }
        PM_TR(3, g, 3);
        PM_TV(3, g, 6, ((unsigned long long)entries << 1) | (staged ? 1u : 0u));
        PM_TV(3, g, 7, gr.count);
    }
}

// One warp walks one long line.  Lane i owns bytes [a + i*c, a + (i+1)*c).  Lane 0 starts in
// the exact start-of-line state; every other lane GUESSES its entry state by warming up from
// the DFA's mid-line idle state over the 32 bytes in front of its chunk.  The guesses are then
// validated in lane order: where a guess equals the true exit state of the previous chunk the
// lane's walk stands, otherwise that one chunk is walked again from the true state.  The result
// is exact; speculation only shortens the critical path (log text leaves search DFAs idle).
__global__ void __launch_bounds__(128) k_check_long(const uint8_t *__restrict__ text, const uint8_t *__restrict__ blob, BlobHeader h, const Item *__restrict__ items,
                                                    const uint32_t *__restrict__ long_list, unsigned long long long_cap, unsigned char *__restrict__ results,
                                                    const ResolveCounters *ctr) {
    __shared__ __align__(16) uint8_t s_cm[4][256];
    __shared__ __align__(16) uint8_t s_tb[4][kCheckTabCap];
    if (ctr->overflow) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long n = ctr->n_long < long_cap ? ctr->n_long : long_cap;
    const unsigned long long n_warps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
    for (unsigned long long k = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); k < n; k += n_warps) {
        const uint32_t idx = long_list[k];
        const Item it = items[idx];
        UnitTables ut = unit_tables(blob, h, it.unit);
        __syncwarp();   // the previous line's walks are over
        // like k_check: the unit's table and class map in shared memory where they fit (one byte per entry)
        const bool staged = warp_stage_unit(blob, h, reinterpret_cast<const UnitDesc *>(blob + h.off_units)[it.unit], s_cm[warp], s_tb[warp]);
        const uint32_t mid = ut.mid;   // mid-line idle state (pm_library.cc)
        const unsigned long long a = it.a, b = it.a + it.len;
        const unsigned long long c = (it.len + 31) / 32;
        const unsigned long long ca = min(b, a + lane * c), cb = min(b, ca + c);
        auto walk = [&](uint32_t s, unsigned long long p0, unsigned long long p1) {
            if (staged) return walk_staged(text, p0, p1, s, s_tb[warp], s_cm[warp], ut.stride, ut.accept, ut.dead);
            if (ut.wide) {
                // text as aligned 32-bit words: one load per four lookups, and off the dependent chain (byte loads in the
                // chain were 74 % of this kernel's stall samples, and k_finalize waited 12 us for it); the accepting state is
                // sticky and the dead state absorbing, so both are tested once per word
                const uint16_t *w = ut.wide;
                unsigned long long p = p0;
                for (; p < p1 && (p & 3); p++) s = w[(s << 8) | text[p]];
                for (; p + 4 <= p1 && s != ut.accept && s != ut.dead; p += 4) {
                    const uint32_t x = *reinterpret_cast<const uint32_t *>(text + p);
                    s = w[(s << 8) | (x & 0xFFu)];
                    s = w[(s << 8) | ((x >> 8) & 0xFFu)];
                    s = w[(s << 8) | ((x >> 16) & 0xFFu)];
                    s = w[(s << 8) | (x >> 24)];
                }
                for (; p < p1 && s != ut.accept && s != ut.dead; p++) s = w[(s << 8) | text[p]];
            } else {
                for (unsigned long long p = p0; p < p1 && s != ut.accept && s != ut.dead; p++) s = ut.tab[s * ut.stride + ut.cmap[text[p]]];
            }
            return s;
        };
        uint32_t in_spec;
        if (lane == 0 || ca - a <= 32) {
            in_spec = walk(0, a, ca);                       // exact: walked from the line start
        } else {
            in_spec = walk(mid, ca - 32, ca);
        }
        uint32_t out_spec = walk(in_spec, ca, cb);
        uint32_t truth = 0;                                 // true state at the start of chunk i
        for (int i = 0; i < 32; i++) {
            uint32_t in_i = __shfl_sync(FULL, in_spec, i), out_i = __shfl_sync(FULL, out_spec, i);
            if (in_i != truth) {                            // mis-speculated chunk: redo it (warp-uniform branch)
                uint32_t redo = 0;
                if (lane == i) redo = walk(truth, ca, cb);
                out_i = __shfl_sync(FULL, redo, i);
            }
            truth = out_i;
            if (truth == ut.accept || truth == ut.dead) break;
        }
        if (truth != ut.accept) truth = ut.tab[truth * ut.stride + ut.stride - 1];
        if (lane == 0) results[idx] = (truth == ut.accept);
    }
}

// One thread per key.  The row of a key goes to out[rank], rank = position of the key in the
// sorted, de-duplicated uniq[]; so keys[] can be the raw queue, planned and checked while the sort
// was still running on another stream (duplicates write identical rows).
__global__ void __launch_bounds__(512) k_finalize(const uint8_t *__restrict__ blob, BlobHeader h, const unsigned long long *__restrict__ keys,
                                                  const unsigned long long *n_keys_ptr, const unsigned long long *__restrict__ uniq,
                                                  const unsigned long long *n_unique_ptr, unsigned long long ev_cap, EventOut out,
                                                  const Item *__restrict__ items, const unsigned char *__restrict__ results,
                                                  const EvAux *__restrict__ aux, const PodTable pt, ResolveCounters *ctr, const uint32_t *__restrict__ rank) {
    if (ctr->overflow) return;
    const unsigned long long n_keys = *n_keys_ptr, n_unique = *n_unique_ptr;
    if (n_unique > ev_cap) {
        if (threadIdx.x == 0 && blockIdx.x == 0) ctr->overflow = 1;
        return;
    }
    const PatternDesc *pds = reinterpret_cast<const PatternDesc *>(blob + h.off_patterns);
    const SecDesc *sds = reinterpret_cast<const SecDesc *>(blob + h.off_secondaries);
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_keys; i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long key = keys[i];
        unsigned long long ev;
        if (rank) {                                          // keys[] is the raw queue and k_rank has placed it (launch_sort_small)
            ev = rank[i];
        } else {                                             // lower bound; the key is present
            unsigned long long lo = 0, hi = n_unique;
            while (lo < hi) {
                unsigned long long mid = (lo + hi) >> 1;
                if (uniq[mid] < key) lo = mid + 1; else hi = mid;
            }
            ev = lo;
        }
        if (ev >= n_unique || uniq[ev] != key) continue;
        const uint32_t pat = (uint32_t)(key & ((1u << kPatternBits) - 1u));
        const PatternDesc pd = pds[pat];
        const EvAux ax = aux[i];
        // context: leading run of STACK lines after the lines_after window
        int run = 0;
        // (verdict bytes are read eight at a time, independent loads: one latency instead of one per line)
        for (int base = 0; base < ax.n_stack; base += 8) {
            unsigned miss = 0;
#pragma unroll
            for (int q = 0; q < 8; q++)
                if (base + q < ax.n_stack && !results[ax.item_base + base + q]) miss |= 1u << q;
            if (miss) {
                run = base + __ffs(miss) - 1;
                break;
            }
            run = min(base + 8, ax.n_stack);
        }
        int64_t ctx_end = ax.ctx_end;
        if (run > 0) {
            const Item it = items[ax.item_base + run - 1];
            ctx_end = (int64_t)(it.a + it.len - pt.pods[ax.pod].start);
        }
        // score: secondaries in order, nearest matching line (SPEC.md section 5)
        double acc = 0.0;
        unsigned long long off = ax.item_base + ax.n_stack;
        for (uint32_t j = 0; j < pd.sec_count; j++) {
            const SecDesc sd = sds[pd.sec_first + j];
            const int cnt = 2 * sd.window + 1;
            int best = -1;
            for (int base = 0; base < cnt && best < 0; base += 8) {
                unsigned hit = 0;
#pragma unroll
                for (int q = 0; q < 8; q++)
                    if (base + q < cnt && results[off + base + q]) hit |= 1u << q;
                if (hit) best = (base + __ffs(hit)) >> 1;   // first hit idx = base + ffs - 1; distance (idx + 1) >> 1; idx ascending = distance non-decreasing
            }
            if (best >= 0) {
                double term = __dmul_rn(sd.weight, __dsub_rn(1.0, __ddiv_rn((double)best, (double)(sd.window + 1))));
                acc = __dadd_rn(acc, term);
            }
            off += cnt;
        }
        out.pod[ev] = ax.pod;
        out.pattern[ev] = (int32_t)pat;
        out.line_no[ev] = ax.line_no;
        out.line_start[ev] = ax.line_start;
        out.line_len[ev] = ax.line_len;
        out.ctx_first[ev] = ax.ctx_first;
        out.ctx_last[ev] = ax.ctx_last + run;
        out.ctx_start[ev] = ax.ctx_start;
        out.ctx_end[ev] = ctx_end;
        out.score[ev] = __dmul_rn(__dmul_rn(pd.confidence, pd.mult), __dadd_rn(1.0, acc));
    }
}

// ------------------------------------------------------------------------------------ k_line_starts
// Materialised newline index of one pod (a10): out[r] = offset of line r relative to the pod.
// One warp per 2 KiB tile: 16-bit newline masks per lane, warp prefix scan of the popcounts,
// scatter at tile_prefix[tile] - nl_before(pod) + rank + 1.
__global__ void __launch_bounds__(256) k_line_starts(const uint8_t *__restrict__ text, const unsigned long long *__restrict__ tile_prefix, unsigned long long nl_before,
                                                     int n_lines, unsigned long long pod_start, unsigned long long pod_eff_end, long long *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const unsigned long long first_tile = pod_start / kTileBytes, last_tile = (pod_eff_end + kTileBytes - 1) / kTileBytes;
    const unsigned long long n_warps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
    if (blockIdx.x == 0 && threadIdx.x == 0 && n_lines > 0) out[0] = 0;
    for (unsigned long long t = first_tile + (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < last_tile; t += n_warps) {
        unsigned long long base = t * kTileBytes;
        long long rank0 = (long long)tile_prefix[t] - (long long)nl_before;   // newlines of the pod before this tile
#pragma unroll
        for (int k = 0; k < 4; k++) {
            unsigned long long a = base + k * 512 + lane * 16;
            uint4 v = *reinterpret_cast<const uint4 *>(text + a);
            uint32_t m = newline_mask16(v);
            int c = __popc(m), inc = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(FULL, inc, o);
                if (lane >= o) inc += y;
            }
            long long r = rank0 + inc - c;
            while (m) {
                int bpos = __ffs(m) - 1;
                m &= m - 1;
                unsigned long long pos = a + bpos;
                // newline number r (0-based within the pod) ends line r and starts line r + 1
                if (pos >= pod_start && r + 1 < n_lines && r >= 0) out[r + 1] = (long long)(pos + 1 - pod_start);
                r++;
            }
            rank0 += __shfl_sync(FULL, inc, 31);
        }
    }
}

template <int BITS, bool FOLD>
cudaError_t launch_scan_one(const uint8_t *text, unsigned long long n_tiles, const uint4 *bitmap, const uint4 *bitmap2, uint32_t bits2, uint32_t *tile_nl,
                            unsigned long long *cand, unsigned long long cand_cap, ScanCounters *ctr, int num_sms, cudaStream_t s) {
    // shared memory: the two bitmaps and the per-warp candidate queues
    const size_t smem = (size_t)(1 << BITS) / 8 + (size_t)(1u << bits2) / 8 + (size_t)(kScanThreads / 32) * kQueue * 8;
    ScanConsts kc;
    kc.c_nl = 0x0A0A0A0Au;
    kc.c_7f = 0x7F7F7F7Fu;
    kc.c_80 = 0x80808080u;
    kc.c_20 = 0x20202020u;
    kc.one = 1u;
    kc.four = 4u;
    kc.two25 = 0x01010101u;   // dp4a weights
    kc.idx_mul = 1u << (BITS - 5);
    kc.mult = kFilterMult0;
    kc.bm_base = 0;
    cudaError_t e = cudaFuncSetAttribute(k_scan<BITS, FOLD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_scan<BITS, FOLD><<<num_sms, kScanThreads, smem, s>>>(reinterpret_cast<const uint4 *>(text), n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, kc);
    return cudaGetLastError();
}

template <int BITS, bool FOLD>
cudaError_t launch_scan_half_one(const uint8_t *text, unsigned long long n_tiles, const uint4 *bitmap, const uint4 *bitmap2, uint32_t bits2, uint32_t *tile_nl,
                                 unsigned long long *cand, unsigned long long cand_cap, ScanCounters *ctr, int num_sms, cudaStream_t s) {
    const size_t smem = (size_t)(1 << BITS) / 8 + (size_t)(1u << bits2) / 8 + (size_t)(kHalfThreads / 32) * kQueue * 8;
    ScanConsts kc;
    kc.c_nl = 0x0A0A0A0Au;
    kc.c_7f = 0x7F7F7F7Fu;
    kc.c_80 = 0x80808080u;
    kc.c_20 = 0x20202020u;
    kc.one = 1u;
    kc.four = 4u;
    kc.two25 = 0x01010101u;
    kc.idx_mul = 1u << (BITS - 5);
    kc.mult = kFilterMult0;
    kc.bm_base = 0;
    cudaError_t e = cudaFuncSetAttribute(k_scan_half<BITS, FOLD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_scan_half<BITS, FOLD><<<num_sms, kHalfThreads, smem, s>>>(reinterpret_cast<const uint4 *>(text), n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr,
                                                               kc);
    return cudaGetLastError();
}

template <int BITS>
cudaError_t launch_scan_half_bits(bool fold, const uint8_t *text, unsigned long long n_tiles, const uint4 *bitmap, const uint4 *bitmap2, uint32_t bits2,
                                  uint32_t *tile_nl, unsigned long long *cand, unsigned long long cand_cap, ScanCounters *ctr, int num_sms, cudaStream_t s) {
    return fold ? launch_scan_half_one<BITS, true>(text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s)
                : launch_scan_half_one<BITS, false>(text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
}

template <int BITS>
cudaError_t launch_scan_bits(bool fold, const uint8_t *text, unsigned long long n_tiles, const uint4 *bitmap, const uint4 *bitmap2, uint32_t bits2,
                             uint32_t *tile_nl, unsigned long long *cand, unsigned long long cand_cap, ScanCounters *ctr, int num_sms, cudaStream_t s) {
    return fold ? launch_scan_one<BITS, true>(text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s)
                : launch_scan_one<BITS, false>(text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
}

}  // namespace

cudaError_t launch_pad(uint8_t *text, const PodDev *pods, int n_pods, unsigned long long total_bytes, cudaStream_t s) {
    int warps = n_pods + 1, threads = 128, blocks = (warps * 32 + threads - 1) / threads;
    k_pad<<<blocks, threads, 0, s>>>(text, pods, n_pods, total_bytes);
    return cudaGetLastError();
}

cudaError_t launch_scan(const uint8_t *text, unsigned long long n_tiles, const DeviceLibrary &lib, uint32_t *tile_nl, unsigned long long *cand,
                        unsigned long long cand_cap, ScanCounters *ctr, int num_sms, cudaStream_t s) {
    const FilterView &fv = lib.fv;
    if (fv.log2_bits == 0) {
        k_scan_nl_only<<<num_sms, kScanThreads, 0, s>>>(reinterpret_cast<const uint4 *>(text), n_tiles, tile_nl);
        return cudaGetLastError();
    }
    const uint4 *bitmap = reinterpret_cast<const uint4 *>(fv.bitmap), *bitmap2 = reinterpret_cast<const uint4 *>(fv.bitmap2);
    const uint32_t bits2 = fv.log2_bits2;
    const bool fold = fv.fold != 0;
    if (fv.granule == 2) {   // every even offset is tested (5-byte literals)
        switch (fv.log2_bits) {
            case 15: return launch_scan_half_bits<15>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
            case 16: return launch_scan_half_bits<16>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
            case 17: return launch_scan_half_bits<17>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
            case 18: return launch_scan_half_bits<18>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
            case 19: return launch_scan_half_bits<19>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
            case 20: return launch_scan_half_bits<20>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
            default: return cudaErrorInvalidValue;
        }
    }
    switch (fv.log2_bits) {
        case 15: return launch_scan_bits<15>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
        case 16: return launch_scan_bits<16>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
        case 17: return launch_scan_bits<17>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
        case 18: return launch_scan_bits<18>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
        case 19: return launch_scan_bits<19>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
        case 20: return launch_scan_bits<20>(fold, text, n_tiles, bitmap, bitmap2, bits2, tile_nl, cand, cand_cap, ctr, num_sms, s);
        default: return cudaErrorInvalidValue;
    }
}

// One L1/shared-memory split for k_check and k_check_long, which run side by side.  An SM cannot hold CTAs of two kernels that want different
// splits: it drains one kind before it takes the other.  k_check wants five CTAs of 14 KiB per SM; with the driver's own choice
// per kernel, k_check beside k_check_long (no shared memory) started half of its CTAs 11 us late (tools_warp_trace.py).  Set once
// per device.  (k_verify with its tables in shared memory asks for the maximum instead, see launch_confirm.  The same split for
// ALL the small kernels was measured too: 4.29 against 4.41 TB/s with five batches in flight, 0.340 against 0.332 ms alone.)
static void tail_carveout() {
    static std::atomic<unsigned long long> done{0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
    if (done.load(std::memory_order_relaxed) & (1ull << dev)) return;
    static const int pct = getenv("PM_TAIL_CARVEOUT") ? atoi(getenv("PM_TAIL_CARVEOUT")) : 40;
    cudaFuncSetAttribute(k_check, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    cudaFuncSetAttribute(k_check_long, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    done.fetch_or(1ull << dev, std::memory_order_relaxed);
}

cudaError_t launch_prefix(const uint32_t *tile_nl, unsigned long long *tile_prefix, unsigned long long n_tiles, unsigned long long *scratch, cudaStream_t s) {
    unsigned long long n_blocks = (n_tiles + kPrefBlock - 1) / kPrefBlock;
    if (n_blocks == 0) return cudaSuccess;
    k_prefix_partial<<<(unsigned)n_blocks, kPrefThreads, 0, s>>>(tile_nl, n_tiles, scratch);
    k_prefix_final<<<(unsigned)n_blocks, kPrefThreads, 0, s>>>(tile_nl, n_tiles, scratch, tile_prefix);
    return cudaGetLastError();
}

cudaError_t launch_pods(const uint8_t *text, const unsigned long long *tile_prefix, PodDev *pods, int n_pods, cudaStream_t s) {
    if (n_pods == 0) return cudaSuccess;
    int threads = 128, blocks = (n_pods * 32 + threads - 1) / threads;
    k_pods<<<blocks, threads, 0, s>>>(text, tile_prefix, pods, n_pods);
    return cudaGetLastError();
}

cudaError_t launch_confirm(const uint8_t *text, unsigned long long total_bytes, const DeviceLibrary &lib, PodTable pt, const unsigned long long *cand,
                           unsigned long long cand_cap, unsigned long long *tasks, unsigned long long *raw, unsigned long long raw_cap, ScanCounters *ctr,
                           int num_sms, cudaStream_t s, bool tables_in_smem) {
    if (lib.fv.log2_bits == 0) return cudaSuccess;
    tail_carveout();
    static const int confirm_ctas = getenv("PM_CONFIRM_CTAS") ? atoi(getenv("PM_CONFIRM_CTAS")) : 8;
    k_confirm<<<num_sms * confirm_ctas, 256, 0, s>>>(text, total_bytes, lib.fv, cand, cand_cap, tasks, raw_cap, ctr);
    stage_mark("k_confirm");
    // per-warp shared-memory budget for the task's table: the largest verify table of the library, as long as
    // 12 CTAs still fit an SM (the register limit of this kernel); larger tables are walked through L1.  Only with
    // one batch at a time on the device (tables_in_smem): the step alone gains 7 us, but with five batches in flight the
    // 19 KiB per CTA cost 5 % of throughput (3.96 -> 3.77 TB/s) - the tails of several batches share the SMs then
    const uint32_t table_cap = tables_in_smem ? lib.verify_table_cap : 0u;
    const size_t vsm = (size_t)kVerifyWarps * table_cap;
    if (vsm > 40 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_verify, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vsm);
        if (e != cudaSuccess) return e;
    }
    // with tables in shared memory the kernel wants all the shared memory an SM has (6 CTAs of 35 KiB); without the hint the
    // driver's L1/shared split left room for fewer and the kernel ran in waves
    cudaFuncSetAttribute(k_verify, cudaFuncAttributePreferredSharedMemoryCarveout, vsm ? (int)cudaSharedmemCarveoutMaxShared : (int)cudaSharedmemCarveoutDefault);
    k_verify<<<num_sms * 16, kVerifyWarps * 32, vsm, s>>>(text, total_bytes, lib.blob, lib.h, pt, tasks, raw_cap, raw, raw_cap, ctr, table_cap);
    return cudaGetLastError();
}

cudaError_t launch_dfa_scan(const uint8_t *text, unsigned long long total_bytes, const DeviceLibrary &lib, PodTable pt, int group, bool wide,
                            size_t table_bytes, unsigned long long *raw, unsigned long long raw_cap, ScanCounters *ctr, int num_sms, cudaStream_t s) {
    size_t smem = 2 * (size_t)kDfaTile + ((table_bytes + 15) & ~(size_t)15) + 256;
    cudaError_t e;
    if (wide) {
        e = cudaFuncSetAttribute(k_dfa_scan<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        k_dfa_scan<true><<<num_sms, kDfaThreads, smem, s>>>(text, total_bytes, lib.blob, lib.h, pt, group, raw, raw_cap, ctr);
    } else {
        e = cudaFuncSetAttribute(k_dfa_scan<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        k_dfa_scan<false><<<num_sms, kDfaThreads, smem, s>>>(text, total_bytes, lib.blob, lib.h, pt, group, raw, raw_cap, ctr);
    }
    return cudaGetLastError();
}

size_t sort_temp_bytes(unsigned long long n) {
    size_t a = 0, b = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, a, (const unsigned long long *)nullptr, (unsigned long long *)nullptr, (long long)n);
    cub::DeviceSelect::Flagged(nullptr, b, (const unsigned long long *)nullptr, (const unsigned char *)nullptr, (unsigned long long *)nullptr,
                               (unsigned long long *)nullptr, (long long)n);
    return ((a > b ? a : b) + 255) / 256 * 256 + ((n + 255) / 256) * 256;   // + flag bytes
}

cudaError_t launch_sort_unique(unsigned long long *raw, unsigned long long *sorted, unsigned long long *uniq, unsigned long long n, unsigned long long *n_unique,
                               void *temp, size_t temp_bytes, cudaStream_t s) {
    if (n == 0) return cudaMemsetAsync(n_unique, 0, sizeof(unsigned long long), s);
    size_t flag_bytes = ((n + 255) / 256) * 256;
    unsigned char *flags = reinterpret_cast<unsigned char *>(temp);
    void *cub_temp = flags + flag_bytes;
    size_t cub_bytes = temp_bytes - flag_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortKeys(cub_temp, cub_bytes, raw, sorted, (long long)n, 0, 64, s);
    if (e != cudaSuccess) return e;
    k_mark_unique<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(sorted, n, flags);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return cub::DeviceSelect::Flagged(cub_temp, cub_bytes, sorted, flags, uniq, n_unique, (long long)n, s);
}

cudaError_t launch_gram_profile(const uint8_t *text, unsigned long long sample_bytes, const uint32_t *keys, uint32_t log2_cap, uint32_t fold,
                                unsigned int *counts, int num_sms, cudaStream_t s) {
    k_gram_profile<<<num_sms * 8, 256, 0, s>>>(reinterpret_cast<const uint4 *>(text), sample_bytes / 16, keys, log2_cap, fold, counts);
    return cudaGetLastError();
}

size_t resolve_item_bytes() { return sizeof(Item); }
size_t resolve_aux_bytes() { return sizeof(EvAux); }
size_t resolve_group_bytes() { return sizeof(Group); }

cudaError_t launch_resolve(const uint8_t *text, const DeviceLibrary &lib, const unsigned long long *tile_prefix, PodTable pt,
                           const unsigned long long *keys, const unsigned long long *n_keys_dev, unsigned long long key_cap,
                           const unsigned long long *uniq, const unsigned long long *n_unique_dev, unsigned long long ev_cap, EventOut out,
                           const ResolveScratch &rs, int num_sms, cudaStream_t s, cudaStream_t side, cudaEvent_t fork_ev, cudaEvent_t join_ev,
                           cudaEvent_t uniq_ready, const uint32_t *rank) {
    if (ev_cap == 0 || key_cap == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(rs.ctr, 0, sizeof(ResolveCounters), s);
    if (e != cudaSuccess) return e;
    unsigned blocks = (unsigned)((key_cap + kPlanWarps - 1) / kPlanWarps);
    if (blocks > (unsigned)num_sms * 16u) blocks = (unsigned)num_sms * 16u;
    Item *items = reinterpret_cast<Item *>(rs.items);
    int slots = lib.ctx_slots > 0 ? lib.ctx_slots : kCtxSlots;
    size_t plan_smem = (size_t)kPlanWarps * 2 * slots * 8;
    if (plan_smem > 48 * 1024) {
        e = cudaFuncSetAttribute(k_plan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan_smem);
        if (e != cudaSuccess) return e;
    }
    k_plan<<<blocks, kPlanWarps * 32, plan_smem, s>>>(text, lib.blob, lib.h, tile_prefix, pt, keys, n_keys_dev, key_cap, items, rs.item_cap,
                                                      rs.long_list, rs.long_cap, reinterpret_cast<EvAux *>(rs.aux), rs.ctr, slots, lib.stack_first,
                                                      reinterpret_cast<Group *>(rs.groups), rs.group_cap);
    stage_mark("k_plan");
    // short and long lines are independent: walk them concurrently when a side stream is given.  (Both in one
    // launch, the long lines on warp_walk_line in the first CTAs of the grid, was measured in round 2: 47 us
    // against 26 + 12 us for this pair.)
    const bool fork = side && fork_ev && join_ev;
    if (fork) {
        cudaEventRecord(fork_ev, s);
        cudaStreamWaitEvent(side, fork_ev, 0);
    }
    tail_carveout();
    k_check_long<<<num_sms * 16, 128, 0, fork ? side : s>>>(text, lib.blob, lib.h, items, rs.long_list, rs.long_cap, rs.results, rs.ctr);
    if (fork) cudaEventRecord(join_ev, side);
    k_check<<<num_sms * 5, kCheckWarps * 32, 0, s>>>(text, lib.blob, lib.h, items, reinterpret_cast<const Group *>(rs.groups), rs.group_cap, rs.results, rs.ctr);
    stage_mark("k_check");
    if (fork) cudaStreamWaitEvent(s, join_ev, 0);
    stage_mark("join k_check_long");
    if (uniq_ready) cudaStreamWaitEvent(s, uniq_ready, 0);   // the sort ran beside plan/check
    stage_mark("join sort");
    // few, full CTAs: with several batches in flight every SM a small kernel sits on is an SM the next
    // batch's scan CTA (a whole SM each) has to wait for
    // (sized for the common case, <= kSmallSort keys; the kernel strides over more)
    unsigned fb = (unsigned)((std::min<unsigned long long>(key_cap, (unsigned long long)kSmallSort) + 511) / 512);
    if (fb > (unsigned)num_sms * 2u) fb = (unsigned)num_sms * 2u;
    k_finalize<<<fb, 512, 0, s>>>(lib.blob, lib.h, keys, n_keys_dev, uniq, n_unique_dev, ev_cap, out, items, rs.results,
                                  reinterpret_cast<const EvAux *>(rs.aux), pt, rs.ctr, uniq_ready ? rank : nullptr);
    return cudaGetLastError();
}

// Result rows and counters to pinned host memory, written by the device: the first *n_rows entries of
// each of the ten arrays (coalesced 16-byte stores over PCIe) and the counter block.  Replaces two
// cudaMemcpyAsync whose sizes the host would have to know - the event count never leaves the device
// before the batch is done, so the copy is exactly as large as the result.
__global__ void __launch_bounds__(256) k_export(EventOut src, EventOut dst, const unsigned long long *n_rows_ptr, unsigned long long cap,
                                                const uint4 *__restrict__ ctr_src, uint4 *__restrict__ ctr_dst, int ctr_vec) {
    const unsigned long long n = *n_rows_ptr <= cap ? *n_rows_ptr : 0;   // more rows than room: the host re-runs the resolve stage
    const unsigned long long n8 = (n * 8 + 15) / 16, n4 = (n * 4 + 15) / 16;   // 16-byte pieces per array
    const unsigned long long per = 4 * n8 + 6 * n4, stride = (unsigned long long)gridDim.x * blockDim.x;
    const uint4 *s8[4] = {reinterpret_cast<const uint4 *>(src.line_start), reinterpret_cast<const uint4 *>(src.ctx_start),
                          reinterpret_cast<const uint4 *>(src.ctx_end), reinterpret_cast<const uint4 *>(src.score)};
    uint4 *d8[4] = {reinterpret_cast<uint4 *>(dst.line_start), reinterpret_cast<uint4 *>(dst.ctx_start), reinterpret_cast<uint4 *>(dst.ctx_end),
                    reinterpret_cast<uint4 *>(dst.score)};
    const uint4 *s4[6] = {reinterpret_cast<const uint4 *>(src.pod), reinterpret_cast<const uint4 *>(src.pattern), reinterpret_cast<const uint4 *>(src.line_no),
                          reinterpret_cast<const uint4 *>(src.line_len), reinterpret_cast<const uint4 *>(src.ctx_first), reinterpret_cast<const uint4 *>(src.ctx_last)};
    uint4 *d4[6] = {reinterpret_cast<uint4 *>(dst.pod), reinterpret_cast<uint4 *>(dst.pattern), reinterpret_cast<uint4 *>(dst.line_no),
                    reinterpret_cast<uint4 *>(dst.line_len), reinterpret_cast<uint4 *>(dst.ctx_first), reinterpret_cast<uint4 *>(dst.ctx_last)};
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += stride) {
        if (i < 4 * n8) {
            const unsigned a = (unsigned)(i / n8);
            const unsigned long long k = i - a * n8;
            d8[a][k] = s8[a][k];
        } else {
            const unsigned long long j = i - 4 * n8;
            const unsigned a = (unsigned)(j / n4);
            const unsigned long long k = j - a * n4;
            d4[a][k] = s4[a][k];
        }
    }
    if (blockIdx.x == 0 && (int)threadIdx.x < ctr_vec) ctr_dst[threadIdx.x] = ctr_src[threadIdx.x];
}

cudaError_t launch_export(EventOut src, EventOut dst_host, const unsigned long long *n_rows_dev, unsigned long long cap, const void *ctr_dev, void *ctr_host,
                          size_t ctr_bytes, int num_sms, cudaStream_t s) {
    // a handful of CTAs saturates the PCIe link; more would only sit on SMs that other batches' scans wait for
    static const unsigned max_blocks = getenv("PM_EXPORT_CTAS") ? (unsigned)atoi(getenv("PM_EXPORT_CTAS")) : 8u;
    unsigned blocks = (unsigned)std::min<unsigned long long>(max_blocks, (cap * 64 / 16 + 255) / 256 + 1);
    (void)num_sms;
    k_export<<<blocks, 256, 0, s>>>(src, dst_host, n_rows_dev, cap, reinterpret_cast<const uint4 *>(ctr_dev), reinterpret_cast<uint4 *>(ctr_host),
                                    (int)((ctr_bytes + 15) / 16));
    return cudaGetLastError();
}

cudaError_t launch_sort_small(const unsigned long long *raw, ScanCounters *ctr, unsigned long long raw_cap, unsigned long long *scratch,
                              unsigned long long *uniq, unsigned long long *n_unique, unsigned int *big_flag, int num_sms, cudaStream_t s) {
    unsigned long long *buf_a = scratch, *buf_b = scratch + kMediumSort;
    k_sort_chunks<<<kSortChunksMedium, kSortChunk / 2, 0, s>>>(raw, ctr, raw_cap, buf_a);
    cudaError_t e = cudaFuncSetAttribute(k_sort_merge, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmallSort * 8);
    if (e != cudaSuccess) return e;
    k_sort_merge<<<kSortChunksSmall, 1024, kSmallSort * 8, s>>>(buf_a, ctr, raw_cap, buf_b);           // n <= kSmallSort, else returns at once
    // every CTA of this grid waits for all the others at a barrier: never more CTAs than the SMs it may run on can hold at
    // once (a green-context partition can be as small as 8 SMs)
    const int merge_ctas = std::max(1, std::min(kMergeCtas, num_sms * 4));
    k_merge_all<<<merge_ctas, kMergeThreads, 0, s>>>(buf_a, buf_b, ctr, raw_cap);                       // n >  kSmallSort, else returns at once
    k_sort_unique<<<1, 1024, 0, s>>>(buf_b, ctr, raw_cap, uniq, n_unique, big_flag);                   // an odd number of passes: the result is in b
    k_rank<<<64, 256, 0, s>>>(raw, ctr, raw_cap, uniq, n_unique, reinterpret_cast<uint32_t *>(buf_a));   // a is free again
    return cudaGetLastError();
}

cudaError_t launch_line_starts(const uint8_t *text, const unsigned long long *tile_prefix, unsigned long long nl_before, int n_lines,
                                  unsigned long long pod_start, unsigned long long pod_eff_end, long long *out, int num_sms, cudaStream_t s) {
    k_line_starts<<<num_sms * 2, 256, 0, s>>>(text, tile_prefix, nl_before, n_lines, pod_start, pod_eff_end, out);
    return cudaGetLastError();
}

}  // namespace pm

#ifdef PM_TRACE
// Debug builds only: copies the warp time stamps of the last launches to the host and clears them.
extern "C" int pm_debug_trace_read(unsigned long long *out, unsigned long long n_words) {
    const unsigned long long all = (unsigned long long)pm::kTraceRegions * pm::kTraceWarps * 8;
    if (n_words > all) n_words = all;
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(out, pm::g_trace, n_words * 8);
    if (e == cudaSuccess) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, pm::g_trace);
        if (e == cudaSuccess) e = cudaMemset(p, 0, all * 8);
    }
    return (int)e;
}
#endif
