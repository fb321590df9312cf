// smsgate_b200 — CUDA kernels (sm_100a) and the C ABI declared in
// include/smsgate_b200.h.  No torch types, no host fallback: every entry
// point launches kernels or fails with an error code.
//
// Kernels
//   k_scan_keys     one 640-thread CTA per SM over tiles that a single cp.async.bulk
//                   (TMA 1-D) copy stages in shared memory: skip filters, body date,
//                   clean+mask rewrite, SHA-1, SHA-256.  Integer-ALU bound, see DESIGN.md.
//   k_cache_build   inserts the cache entries into an open-addressing table.
//   k_join_route    probe + decimal canonicalisation + date/card + routing (+ the
//                   upsert-table insert when called from the whole-path entries).
//   k_dedup_insert / k_dedup_resolve   upsert-by-msg_id.
//   k_export_* / k_merge_* / k_apply_*   global dedup through an all-to-all (NCCL).
//   k_bits_* / k_p2p_*                   the same through NVLink peer memory.
//   k_legacy_parse, k_dt_str_batch       the legacy offline replay and its record builders.
//   k_xml_* (xml_extract.cuh)            XML backup tokenizer.
//   k_md5_keys, k_xsrc_probe             device-source keys, cross-source report.
//   k_decimal_batch / k_strict_date_batch / k_pack_records / k_pack_verdict   helpers.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include <new>
#include <cstdlib>
#include <string>

#include "../../include/smsgate_b200.h"
#include "join_core.cuh"
#include "legacy_core.cuh"
#include "scan_core.cuh"

// ------------------------------------------------------------------ PTX helpers (TMA 1-D bulk copy + mbarrier)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// ask L2 for [src, src + bytes) ahead of the bulk copy that will read it; src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void l2_prefetch_bulk(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// ------------------------------------------------------------------ stage 1: keys + scans
#ifndef K1_THREADS_N
#define K1_THREADS_N 640
#endif
constexpr int K1_THREADS = K1_THREADS_N;   // 20 warps: one CTA per SM
#ifndef K1_CTAS_PER_SM_N
#define K1_CTAS_PER_SM_N 1
#endif
constexpr int K1_CTAS_PER_SM = K1_CTAS_PER_SM_N;
constexpr int K1_TILE = 2 * K1_THREADS;      // messages per tile: every thread owns one short and one long body
#ifndef K1_STAGE_KB
#define K1_STAGE_KB 176
#endif
constexpr int K1_STAGE = K1_STAGE_KB * 1024;  // staged body bytes per tile
constexpr int K1_PAD = 32;                // slack read by the unaligned word reader
#ifndef K1_BUCKET_SHIFT
#define K1_BUCKET_SHIFT 2
#endif
#ifndef K1_BUCKETS_N
#define K1_BUCKETS_N 256
#endif
constexpr int K1_BUCKETS = K1_BUCKETS_N;   // length buckets of 4 bytes for the in-tile sort; a multiple of 32

struct __align__(16) K1Smem {
    DeviceTables tab;
    uint8_t stage[K1_STAGE + K1_PAD];
    uint64_t mbar;
    int32_t m_off[K1_TILE];               // start of message t inside `stage`
    int32_t m_len[K1_TILE];               // body length
    int32_t hist[K1_BUCKETS];
    int32_t cursor[K1_BUCKETS];
    uint16_t order[K1_TILE];              // tile-local message ids sorted by length bucket
    uint16_t work[K1_TILE];               // hashing order: rewritten bodies first, each group longest first
    uint8_t mod[K1_TILE];                 // first-pass flags of the body (SG_F_MODIFIED, SG_F_MAYBE, ...)
    uint16_t rlist[K1_TILE];              // bodies the first pass flagged and clean/mask leaves alone: filter re-scan
    int32_t pcnt[2 * K1_THREADS / 32];    // per warp and half: rewritten | untouched << 16
    int32_t next;                         // next 32-message chunk to hash
    int32_t next_s;                       // next 32-message chunk to scan
    int32_t nr;                           // entries of rlist
};

struct ScanOut {
    uint8_t* msg_id;
    uint8_t* cache_key;
    uint8_t* flags;
    int32_t* body_date;
    int32_t* fixed_len;
    uint8_t* fixed_bytes;
};

__device__ __forceinline__ void k1_store_scan(const ScanOut& o, int64_t i, const ScanResult& sr) {
    if (o.flags) o.flags[i] = (uint8_t)sr.flags;
    if (o.body_date) o.body_date[i] = sr.body_date;
}
__device__ __forceinline__ void k1_store_id(const ScanOut& o, int64_t i, const uint32_t h1[5]) {
    if (o.msg_id) {
        uint32_t* p = (uint32_t*)(o.msg_id + i * 20);
#pragma unroll
        for (int k = 0; k < 5; ++k) p[k] = __byte_perm(h1[k], 0, 0x0123);
    }
}
__device__ __forceinline__ void k1_store_key(const ScanOut& o, int64_t i, const uint32_t h2[8], int flen) {
    if (o.cache_key) {
        uint4* p = (uint4*)(o.cache_key + i * 32);
        p[0] = make_uint4(__byte_perm(h2[0], 0, 0x0123), __byte_perm(h2[1], 0, 0x0123), __byte_perm(h2[2], 0, 0x0123),
                          __byte_perm(h2[3], 0, 0x0123));
        p[1] = make_uint4(__byte_perm(h2[4], 0, 0x0123), __byte_perm(h2[5], 0, 0x0123), __byte_perm(h2[6], 0, 0x0123),
                          __byte_perm(h2[7], 0, 0x0123));
    }
    if (o.fixed_len) o.fixed_len[i] = flen;
}

// A body that does not fit the staging buffer: one thread walks it in global
// memory.  Kept out of line so the hot loop's instruction footprint stays small.
__device__ __noinline__ void k1_giant(const uint8_t* bytes, int64_t lo, int64_t hi, int64_t msg, const ScanTables T,
                                      const ScanOut out) {
    const uint8_t* m = bytes + lo;
    const int len = (int)(hi - lo);
    ScanResult sr = sg_scan_body(m, len, T);
    uint32_t h1[5], h2[8];
    sg_sha1_outofline(m, len, h1);
    int flen = len;
    const uint8_t* fm = m;
    if ((sr.flags & SG_F_MODIFIED) && out.fixed_bytes) {
        uint8_t* w = out.fixed_bytes + lo;
        for (int k = 0; k < len; ++k) w[k] = m[k];
        flen = sg_rewrite_fixed(w, len, T);
        fm = w;
    }
    sg_sha256_outofline(fm, flen, h2);
    k1_store_scan(out, msg, sr);
    k1_store_id(out, msg, h1);
    k1_store_key(out, msg, h2, flen);
}

// -DK1_PHASE_TIMERS (variant build, scripts/time_variants.py): thread 0 of every CTA adds up the clocks between the
// tile's barriers - staging + sort, scan phase, partition, hash phase - into g_k1_phase; never in the shipped build.
#ifdef K1_PHASE_TIMERS
__device__ unsigned long long g_k1_phase[8];
#define K1_T(slot) do { if (tid == 0) { const long long now_ = clock64(); atomicAdd(&g_k1_phase[slot], (unsigned long long)(now_ - t_prev)); t_prev = now_; } } while (0)
extern "C" int smsgate_debug_scan_phases(unsigned long long* h_out8, int reset) {
    if (h_out8 && cudaMemcpyFromSymbol(h_out8, g_k1_phase, sizeof(g_k1_phase)) != cudaSuccess) return -1;
    if (reset) { unsigned long long z[8] = {0}; if (cudaMemcpyToSymbol(g_k1_phase, z, sizeof(z)) != cudaSuccess) return -1; }
    return 0;
}
#else
#define K1_T(slot) do { } while (0)
#endif

__global__ void __launch_bounds__(K1_THREADS, K1_CTAS_PER_SM)
k_scan_keys(const uint8_t* __restrict__ bytes, const int64_t* __restrict__ offsets, int64_t n,
            const DeviceTables* __restrict__ gtab, ScanOut out) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    K1Smem& S = *reinterpret_cast<K1Smem*>(smem_raw);
    const int tid = threadIdx.x;

    {   // tables -> shared, 16 bytes per thread per step
        const uint4* src = reinterpret_cast<const uint4*>(gtab);
        uint4* dst = reinterpret_cast<uint4*>(&S.tab);
        for (int k = tid; k < (int)(sizeof(DeviceTables) / 16); k += K1_THREADS) dst[k] = src[k];
    }
    if (tid == 0) {
        mbar_init(&S.mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    ScanTables T;
    T.t = &S.tab;

    uint32_t parity = 0;
#ifdef K1_PHASE_TIMERS
    long long t_prev = clock64();
#endif
    // every CTA takes an equal contiguous share of the messages and walks it in equal tiles of at most
    // K1_TILE: all CTAs do the same number of tiles, no tail wave
    const int64_t r0 = n * (int64_t)blockIdx.x / gridDim.x, r1 = n * ((int64_t)blockIdx.x + 1) / gridDim.x;
    // ... of at most K1_TILE messages AND about 90 % of the staging buffer in bytes (long bodies: mixed-locale and
    // multi-part messages), so that a tile is staged in one piece instead of a full sub-batch plus a sliver
    const int64_t nt_cnt = (r1 - r0 + K1_TILE - 1) / K1_TILE;
    const int64_t offsets_end = r1 > r0 ? offsets[r1] : 0;
    const int64_t share_bytes = r1 > r0 ? offsets_end - offsets[r0] : 0;
    const int64_t nt_bytes = (share_bytes + (K1_STAGE / 10) * 9 - 1) / ((K1_STAGE / 10) * 9);
    const int64_t nt = nt_cnt > nt_bytes ? nt_cnt : nt_bytes;
    // (full tiles of K1_TILE plus one short tile, i.e. exactly two chunks per warp and phase, measured slower:
    // 0.495 vs 0.484 ms per 1 M messages, profiles/r2t_tile_variants.json)
    const int64_t per = nt ? (r1 - r0 + nt - 1) / nt : 1;
    for (int64_t t0 = r0; t0 < r1; t0 += per) {
        const int64_t t1 = (t0 + per < r1) ? t0 + per : r1;
        int64_t s = t0;
        while (s < t1) {
            // ---- how many messages starting at s fit the staging buffer (offsets are monotone: a prefix)
            const int64_t lo = offsets[s];
            const int64_t base = lo & ~(int64_t)15;
            int nfit = 0;
            int64_t my_lo[2], my_hi[2];
            bool fits[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int64_t mine = s + tid + h * K1_THREADS;
                fits[h] = false;
                my_lo[h] = my_hi[h] = 0;
                if (mine < t1) {
                    my_lo[h] = offsets[mine];
                    my_hi[h] = offsets[mine + 1];
                    fits[h] = (my_hi[h] - base) <= K1_STAGE;
                }
                nfit += fits[h] ? 1 : 0;
            }
            for (int k = tid; k < K1_BUCKETS; k += K1_THREADS) S.hist[k] = 0;
            if (tid == 0) { S.next = 0; S.next_s = 0; S.nr = 0; }
            // the first half is a prefix of the tile, so the count of fitting messages is exact
            // only if the second half is counted after the first: do two block-wide counts
            const int cnt0 = __syncthreads_count(fits[0]);
            const int cnt1 = (cnt0 == K1_THREADS) ? __syncthreads_count(fits[1]) : 0;
            const int cnt = cnt0 + cnt1;
            if (cnt == 0) {
                if (tid == 0) k1_giant(bytes, my_lo[0], my_hi[0], s, T, out);
                s += 1;
                __syncthreads();
                continue;
            }
            const int64_t e = s + cnt;
            const int64_t hi = offsets[e];
            // ---- stage [base, hi): one bulk copy for the 16-byte-aligned part, plain loads for the tail
            const int64_t bulk_end = hi & ~(int64_t)15;
            const uint32_t bulk = (uint32_t)(bulk_end > base ? bulk_end - base : 0);
            if (tid == 0 && bulk) {
                fence_proxy_async();                         // generic-proxy writes of the last tile come first
                mbar_expect_tx(&S.mbar, bulk);
                tma_load_1d(S.stage, bytes + base, bulk, &S.mbar);
            }
            {
                const int64_t tail0 = (bulk_end > base) ? bulk_end : base;
                if (tid < 16 && tail0 + tid < hi) S.stage[tail0 - base + tid] = bytes[tail0 + tid];
            }
            // ---- while the copy is in flight: sort the tile's messages by length (4-byte buckets) so that
            //      the 32 lanes of a warp walk bodies of about the same length and SHA block count
            int key[2] = {0, 0};
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int loc = tid + h * K1_THREADS;
                if (loc < cnt) {
                    const int len = (int)(my_hi[h] - my_lo[h]);
                    S.m_off[loc] = (int)(my_lo[h] - base);
                    S.m_len[loc] = len;
                    int k = len >> K1_BUCKET_SHIFT;
                    k = k < K1_BUCKETS - 1 ? k : K1_BUCKETS - 1;
                    key[h] = k;
                    atomicAdd(&S.hist[k], 1);
                }
            }
            __syncthreads();
            if (tid < 32) {                                  // exclusive scan of the bucket counts, K1_BUCKETS / 32 per lane
                constexpr int PER = K1_BUCKETS / 32;
                int c[PER], sum = 0;
#pragma unroll
                for (int k = 0; k < PER; ++k) { c[k] = S.hist[PER * tid + k]; sum += c[k]; }
                int incl = sum;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, d);
                    if (tid >= d) incl += v;
                }
                int run = incl - sum;
#pragma unroll
                for (int k = 0; k < PER; ++k) { S.cursor[PER * tid + k] = run; run += c[k]; }
            }
            __syncthreads();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int loc = tid + h * K1_THREADS;
                if (loc < cnt) S.order[atomicAdd(&S.cursor[key[h]], 1)] = (uint16_t)loc;
            }
            if (bulk) {
                mbar_wait(&S.mbar, parity);
                parity ^= 1;
            }
            __syncthreads();
            K1_T(0);                                         // offsets, staging copy, sort

            // ---- phase S: the scans.  Warps pull 32-message chunks of the length-sorted order, longest first:
            //      the lanes of a warp walk bodies of about the same length, and a tile that mixes 100-byte and
            //      1 200-byte bodies still ends with all warps busy (a fixed short+long pairing did not).
            const int lane = tid & 31, wrp = tid >> 5;
#pragma unroll 1
            for (;;) {
                int c = 0;
                if (lane == 0) c = atomicAdd(&S.next_s, 1);
                c = __shfl_sync(0xffffffffu, c, 0);
                if (c * 32 >= cnt) break;
                const int r = cnt - 1 - (c * 32 + lane);
                if (r < 0) continue;
                const int t = S.order[r];
                // first pass (scan_core.cuh): two automata; the few flagged bodies get their filter flags later
                ScanResult sr = sg_scan_first(S.stage + S.m_off[t], S.m_len[t], T);
                S.mod[t] = (uint8_t)sr.flags;
                if ((sr.flags & SG_F_MAYBE) && !(sr.flags & SG_F_MODIFIED)) S.rlist[atomicAdd(&S.nr, 1)] = (uint16_t)t;
                sr.flags &= ~(uint32_t)SG_F_MAYBE;
                k1_store_scan(out, s + t, sr);
            }
            __syncthreads();
            K1_T(1);                                         // scan phase
            // ---- hashing order: the bodies clean/mask changes first (they cost SHA-1 + rewrite + SHA-256 of
            //      the result), then the untouched ones (both digests in one walk), each group longest first,
            //      so that a 32-message chunk is one kind of work and the dearest chunks start first
            int wt[2];
            uint32_t balM[2], balU[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int q = tid + h * K1_THREADS;
                const bool valid = q < cnt;
                wt[h] = valid ? S.order[cnt - 1 - q] : 0;
                const bool isM = valid && (S.mod[wt[h]] & SG_F_MODIFIED);
                balM[h] = __ballot_sync(0xffffffffu, isM);
                balU[h] = __ballot_sync(0xffffffffu, valid && !isM);
                if (lane == 0) S.pcnt[h * (K1_THREADS / 32) + wrp] = __popc(balM[h]) | (__popc(balU[h]) << 16);
            }
            __syncthreads();
            int nM = 0;
            {
                int pre[2] = {0, 0};
#pragma unroll
                for (int k = 0; k < 2 * K1_THREADS / 32; ++k) {
                    const int c = S.pcnt[k];
                    if (k < wrp) pre[0] += c;
                    if (k < K1_THREADS / 32 + wrp) pre[1] += c;
                    nM += c & 0xFFFF;
                }
                const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int q = tid + h * K1_THREADS;
                    if (q < cnt) {
                        const bool isM = (balM[h] >> lane) & 1u;
                        const int pos = isM ? (pre[h] & 0xFFFF) + __popc(balM[h] & lt)
                                            : nM + (pre[h] >> 16) + __popc(balU[h] & lt);
                        S.work[pos] = (uint16_t)wt[h];
                    }
                }
            }
            __syncthreads();
            K1_T(2);                                         // partition
            // ---- phase H: warps pull chunks until none is left
#ifndef K1_NO_L2_PREFETCH
            // the hash phase is seven tenths of a tile and touches no global memory but its results: have L2 fetch
            // the next tile's bytes and offsets meanwhile, so that its bulk copy and its first loads find them there
            // (0.476 vs 0.485 ms per 1 M messages, profiles/r2u_prefetch_variants.json)
            if (tid == 0 && e < r1) {
                const int64_t from = hi & ~(int64_t)15;
                int64_t want = hi - base;                                  // as many bytes again as this tile had
                const int64_t left = (offsets_end & ~(int64_t)15) - from;
                if (want > left) want = left;
                want &= ~(int64_t)15;
                if (want > 0) l2_prefetch_bulk(bytes + from, (uint32_t)want);
                const uintptr_t o_lo = ((uintptr_t)(offsets + e) + 15) & ~(uintptr_t)15;
                uintptr_t o_hi = (uintptr_t)(offsets + e + cnt + 2);
                if (o_hi > (uintptr_t)(offsets + n + 1)) o_hi = (uintptr_t)(offsets + n + 1);
                o_hi &= ~(uintptr_t)15;
                if (o_hi > o_lo) l2_prefetch_bulk((const void*)o_lo, (uint32_t)(o_hi - o_lo));
            }
#endif
            const int nchM = (nM + 31) >> 5, nchU = (cnt - nM + 31) >> 5;
            const int nR = S.nr, nchR = (nR + 31) >> 5;
#pragma unroll 1
            for (;;) {
                int c = 0;
                if (lane == 0) c = atomicAdd(&S.next, 1);
                c = __shfl_sync(0xffffffffu, c, 0);
                if (c >= nchM + nchU + nchR) break;
                if (c >= nchM + nchU) {
                    // filter re-scan of flagged bodies that are not rewritten (those that are re-scan themselves
                    // below, before the rewrite destroys the original): the stored flags gain the two filter bits
                    const int idx = (c - nchM - nchU) * 32 + lane;
                    if (idx < nR) {
                        const int t = S.rlist[idx];
                        const uint32_t fl = (S.mod[t] & ~(uint32_t)SG_F_MAYBE) | sg_scan_filters(S.stage + S.m_off[t], S.m_len[t], T);
                        if (out.flags) out.flags[s + t] = (uint8_t)fl;
                    }
                    continue;
                }
                const bool isM = c < nchM;
                const int idx = isM ? c * 32 + lane : nM + (c - nchM) * 32 + lane;
                const bool act = idx < (isM ? nM : cnt);
                const int t = act ? S.work[idx] : 0;
                const int off = S.m_off[t];
                const int len = act ? S.m_len[t] : 0;
                uint8_t* m = S.stage + off;
                uint32_t h1[5], h2[8];
                int cur = len;
                const int npass = isM ? 2 : 1;
#pragma unroll 1
                for (int pass = 0; pass < npass; ++pass) {   // one copy of the hashing loop, see sg_hash_pass
                    const uint32_t mode = isM ? (pass ? 2u : 1u) : 3u;
                    if (pass && act) {
                        if (S.mod[t] & SG_F_MAYBE) {         // flagged AND rewritten (rare): filters first, on the original
                            const uint32_t fl = (S.mod[t] & ~(uint32_t)SG_F_MAYBE) | sg_scan_filters(m, len, T);
                            if (out.flags) out.flags[s + t] = (uint8_t)fl;
                        }
                        cur = sg_rewrite_fixed(m, len, T);
                    }
                    if (act) sg_hash_pass(m, cur, mode, h1, h2);
                }
                if (act) {
                    k1_store_id(out, s + t, h1);
                    k1_store_key(out, s + t, h2, cur);
                }
                if (isM && out.fixed_bytes) {                // rewritten bodies -> fixed_bytes at their old offsets
                    __syncwarp();
                    const int live = min(32, nM - c * 32);
                    for (int j = 0; j < live; ++j) {
                        const int o = __shfl_sync(0xffffffffu, off, j), fl = __shfl_sync(0xffffffffu, cur, j);
                        // stage and fixed_bytes + base are both 16-byte aligned: same word phase on both sides
                        const uint8_t* src = S.stage + o;
                        uint8_t* dst = out.fixed_bytes + base + o;
                        const bool words = (((uintptr_t)dst ^ (uintptr_t)src) & 3) == 0;
                        const int head = words ? min(fl, (4 - (o & 3)) & 3) : fl;
                        for (int k = lane; k < head; k += 32) dst[k] = src[k];
                        const int nw = (fl - head) >> 2;
                        for (int k = lane; k < nw; k += 32)
                            *reinterpret_cast<uint32_t*>(dst + head + 4 * k) =
                                *reinterpret_cast<const uint32_t*>(src + head + 4 * k);
                        const int done = head + 4 * nw;
                        if (lane < fl - done) dst[done + lane] = src[done + lane];
                    }
                }
            }
            __syncthreads();
            K1_T(3);                                         // hash phase
            s = e;
        }
    }
}

// ------------------------------------------------------------------ cache table
// slots: int2 {entry index or -1, 32-bit fingerprint}; keys stay in their own array.
__device__ __forceinline__ uint64_t key_hash(const uint8_t* k) { return *reinterpret_cast<const uint64_t*>(k); }
__device__ __forceinline__ uint32_t key_fp(const uint8_t* k) { return *reinterpret_cast<const uint32_t*>(k + 8); }

__global__ void k_cache_build(const uint8_t* __restrict__ keys, int64_t m, int2* __restrict__ table, uint64_t mask,
                              int* __restrict__ dup_count) {
    const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (e >= m) return;
    const uint8_t* k = keys + e * 32;
    uint64_t pos = key_hash(k) & mask;
    const uint32_t fp = key_fp(k);
    for (;;) {
        int* slot = &table[pos].x;
        const int prev = atomicCAS(slot, -1, (int)e);
        if (prev == -1) {
            table[pos].y = (int)fp;
            return;
        }
        // duplicate key in the input: keep the later entry out, count it
        const uint4* a = reinterpret_cast<const uint4*>(keys + (int64_t)prev * 32);
        const uint4* b = reinterpret_cast<const uint4*>(k);
        const uint4 a0 = a[0], a1 = a[1], b0 = b[0], b1 = b[1];
        if (a0.x == b0.x && a0.y == b0.y && a0.z == b0.z && a0.w == b0.w && a1.x == b1.x && a1.y == b1.y &&
            a1.z == b1.z && a1.w == b1.w) {
            atomicAdd(dup_count, 1);
            return;
        }
        pos = (pos + 1) & mask;
    }
}
// fingerprints are written after the CAS; a second tiny pass makes them visible for every slot
__global__ void k_cache_fix_fp(const uint8_t* __restrict__ keys, int2* __restrict__ table, uint64_t cap) {
    const uint64_t p = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (p >= cap) return;
    const int e = table[p].x;
    if (e >= 0) table[p].y = (int)key_fp(keys + (int64_t)e * 32);
}

// Per cache entry, next to the table: where its four strings start, how long each is and the loader's
// flags, in ONE 16-byte record - a hit reads one sector here instead of one each from val_offsets (20 bytes
// straddling two), eflags and date_packed.  A length of 0xFFFF means "look it up in val_offsets" (a field
// of 64 KiB or more); date_packed is read only for the entries whose flags say it is used.
struct __align__(16) CacheSide {
    int32_t val_off;
    uint16_t eflags;
    uint16_t len[4];     // str(amount), str(balance), date, card
    uint16_t pad;
};
static_assert(sizeof(CacheSide) == 16, "one side record is 16 bytes");

__global__ void k_cache_side(int64_t m, const int32_t* __restrict__ val_offsets, const uint16_t* __restrict__ eflags,
                             CacheSide* __restrict__ side) {
    const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (e >= m) return;
    const int32_t* o = val_offsets + e * 4;
    CacheSide sd;
    sd.val_off = o[0];
    sd.eflags = eflags[e];
    bool wide = false;
    for (int k = 0; k < 4; ++k) wide |= (o[k + 1] - o[k]) >= 0xFFFF;
    for (int k = 0; k < 4; ++k) sd.len[k] = wide ? (uint16_t)0xFFFF : (uint16_t)(o[k + 1] - o[k]);
    sd.pad = 0;
    side[e] = sd;
}

struct CacheView {
    const uint8_t* keys;
    const int2* table;
    uint64_t mask;
    const int32_t* val_offsets;
    const uint8_t* val_bytes;
    const CacheSide* side;
    const int64_t* date_packed;
    int64_t m;
};

struct JoinOut {
    int32_t* slot;
    uint8_t* status;
    int64_t* amount_coef;
    int64_t* balance_coef;
    int32_t* amount_exp;
    int32_t* balance_exp;
    uint8_t* dec_flags;
    int64_t* date;
    uint8_t* card;
    uint8_t* card_len;
    int64_t* amount_cents;
    int64_t* balance_cents;
};

// ------------------------------------------------------------------ dedup (upsert-by-msg_id)
// One 8-byte slot per key: {first, last} = lowest / highest local index of the messages that carry it.
// `first` doubles as the slot's key: any message holding the key identifies it, so no separate owner word
// is needed, and the CAS that claims a slot, the atomicMin and the atomicMax all land in one 32-byte sector.
// Empty = all bits set (first = 0xFFFFFFFF as unsigned, last = -1), so one memset clears the table.
struct DedupScratch {
    int2* tab;         // [cap] x = first (unsigned, 0xFFFFFFFF = empty), y = last
    int32_t* slot_of;  // [n]   slot of message i, -1 = not an upsert
    uint64_t mask;
};

__device__ __forceinline__ bool id20_equal(const uint8_t* a, const uint8_t* b) {
    const uint32_t* x = reinterpret_cast<const uint32_t*>(a);
    const uint32_t* y = reinterpret_cast<const uint32_t*>(b);
    return x[0] == y[0] && x[1] == y[1] && x[2] == y[2] && x[3] == y[3] && x[4] == y[4];
}

// Claim / find the slot of message i's key and fold i into the slot's first / last occurrence.
//  * The slot is read first; only a slot that looks free is claimed, with ONE 64-bit CAS that puts {first, last} =
//    {i, i} in place: a key seen once costs a load and an atomic.
//  * A taken slot's `first` is some message with the slot's key (a concurrent atomicMin only swaps it for another
//    one), so equality is decided on the full 20 bytes of THAT message's key.
//  * `first` only ever decreases and `last` only ever increases, so the values just read are enough to tell that an
//    atomicMin / atomicMax would change nothing.  Hot keys (50 % Zipf(1.2) duplicates put 90 000 of a million
//    messages on one slot) would otherwise queue a CAS and two more atomics each on the same eight bytes.
//  * The key to compare with - this message's own when the slot looks free - is fetched BEFORE the CAS is issued:
//    in a warp some lanes claim and others compare, the two paths run one after the other, and this way both wait
//    for memory at the same time (measured: no difference to fetching inside the comparing path).
// Measured forms (three atomics, 32-bit CAS + skipped atomics, 64-bit CAS, read first, one lane per key and warp):
// profiles/r2p_dedup_variants.json, r2q_dedup_variants.json, r2r_dedup_variants.json.
__device__ __forceinline__ void dedup_insert_one(int64_t i, const uint8_t* __restrict__ msg_id, const DedupScratch& D) {
    const uint32_t* kw = reinterpret_cast<const uint32_t*>(msg_id + i * 20);
    const uint32_t k0 = kw[0], k1 = kw[1], k2 = kw[2], k3 = kw[3], k4 = kw[4];
    const unsigned me = (unsigned)i;
    const unsigned long long empty = ~0ull;                  // {0xFFFFFFFF, -1}: the table is cleared with 0xFF bytes
    const unsigned long long mine = ((unsigned long long)me << 32) | me;                // .x = first is the low word
    uint64_t pos = (((uint64_t)k1 << 32) | k0) & D.mask;
    for (;;) {
        unsigned long long* slot = reinterpret_cast<unsigned long long*>(&D.tab[pos]);
        unsigned long long cur = __ldcg(slot);
        const uint32_t* ow = reinterpret_cast<const uint32_t*>(msg_id + (int64_t)(cur == empty ? me : (unsigned)cur) * 20);
        uint32_t o0 = ow[0], o1 = ow[1], o2 = ow[2], o3 = ow[3], o4 = ow[4];
        if (cur == empty) {
            cur = atomicCAS(slot, empty, mine);
            if (cur == empty) break;
            ow = reinterpret_cast<const uint32_t*>(msg_id + (int64_t)(unsigned)cur * 20);   // claimed in between
            o0 = ow[0]; o1 = ow[1]; o2 = ow[2]; o3 = ow[3]; o4 = ow[4];
        }
        if ((((o0 ^ k0) | (o1 ^ k1)) | ((o2 ^ k2) | (o3 ^ k3)) | (o4 ^ k4)) == 0) {
            if (me < (unsigned)cur) atomicMin(reinterpret_cast<unsigned*>(&D.tab[pos].x), me);
            if ((int)me > (int)(cur >> 32)) atomicMax(&D.tab[pos].y, (int)me);
            break;
        }
        pos = (pos + 1) & D.mask;
    }
    D.slot_of[i] = (int32_t)pos;
}

#ifndef K2_MIN_BLOCKS
#define K2_MIN_BLOCKS 5      // 48 registers, no spills; 6 (40 registers) is as fast but spills 48 bytes per thread to local
                             // memory since the one-pass decimal parser (+30 MB of DRAM writes per 1 M messages), 4 is slower
                             // (profiles/r2i_scan_variants.json: join5 / join4 against base)
#endif
__global__ void __launch_bounds__(256, K2_MIN_BLOCKS)
k_join_route(int64_t n, const uint8_t* __restrict__ cache_key, const uint8_t* __restrict__ flags,
             const int32_t* __restrict__ body_date, int64_t now_packed, CacheView C, JoinOut out,
             DedupScratch D, int64_t dedup_i0, const uint8_t* __restrict__ dedup_ids,
             const int64_t* __restrict__ phone_dt, uint32_t* __restrict__ key_bits, int64_t key_bits_words) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* kp = reinterpret_cast<const uint4*>(cache_key + i * 32);
    const uint4 k0 = kp[0], k1 = kp[1];
    int e = -1;
    uint4 sdw = make_uint4(0, 0, 0, 0);
    if (C.m > 0) {
        uint64_t pos = (((uint64_t)k0.y << 32) | k0.x) & C.mask;
        const int fp = (int)k0.z;
        for (;;) {
            const int2 sl = C.table[pos];
            if (sl.x < 0) break;
            if (sl.y == fp) {
                const uint4* a = reinterpret_cast<const uint4*>(C.keys + (int64_t)sl.x * 32);
                const uint4 a0 = a[0], a1 = a[1];
                sdw = *reinterpret_cast<const uint4*>(C.side + sl.x);      // in flight together with the key
                if (a0.x == k0.x && a0.y == k0.y && a0.z == k0.z && a0.w == k0.w && a1.x == k1.x && a1.y == k1.y &&
                    a1.z == k1.z && a1.w == k1.w) {
                    e = sl.x;
                    break;
                }
            }
            pos = (pos + 1) & C.mask;
        }
    }
    EntryView ev;
    ev.amount = ev.balance = ev.date = ev.card = nullptr;
    ev.amount_len = ev.balance_len = ev.date_len = ev.card_len = 0;
    ev.eflags = 0;
    ev.date_packed = 0;
    if (e >= 0) {
        // side record: x = val_off, y = eflags | len[0] << 16, z = len[1] | len[2] << 16, w = len[3]
        int32_t o0 = (int32_t)sdw.x, l0 = (int32_t)(sdw.y >> 16), l1 = (int32_t)(sdw.z & 0xFFFF),
                l2 = (int32_t)(sdw.z >> 16), l3 = (int32_t)(sdw.w & 0xFFFF);
        if (l0 == 0xFFFF) {                              // a field of 64 KiB or more: the full offsets
            const int32_t* o = C.val_offsets + (int64_t)e * 4;
            o0 = o[0]; l0 = o[1] - o[0]; l1 = o[2] - o[1]; l2 = o[3] - o[2]; l3 = o[4] - o[3];
        }
        ev.amount = C.val_bytes + o0; ev.amount_len = l0;
        ev.balance = ev.amount + l0; ev.balance_len = l1;
        ev.date = ev.balance + l1; ev.date_len = l2;
        ev.card = ev.date + l2; ev.card_len = l3;
        ev.eflags = sdw.y & 0xFFFF;
        if (ev.eflags & (SG_E_DATE_HOST | SG_E_DATE_NOBODY)) ev.date_packed = C.date_packed[e];
    }
    const RouteResult r = sg_route(flags[i], body_date[i], e >= 0, ev, now_packed,
                                   phone_dt ? phone_dt[dedup_i0 + i] : 0);   // dedup_i0: this launch's first message in the batch
    if (out.slot) out.slot[i] = e;
    if (out.status) out.status[i] = (uint8_t)r.status;
    if (out.amount_coef) out.amount_coef[i] = r.amount.coef;
    if (out.balance_coef) out.balance_coef[i] = r.balance.coef;
    if (out.amount_exp) out.amount_exp[i] = r.amount.exp;
    if (out.balance_exp) out.balance_exp[i] = r.balance.exp;
    if (out.dec_flags)
        out.dec_flags[i] = (uint8_t)(r.amount.err | (r.balance.err << 2) | (r.amount.sign << 4) | (r.balance.sign << 5) |
                                     (r.cents_overflow << 6));
    if (out.amount_cents) out.amount_cents[i] = r.amount_cents;
    if (out.balance_cents) out.balance_cents[i] = r.balance_cents;
    if (out.date) out.date[i] = r.date;
    if (out.card) {
        uint4 c;
        memcpy(&c, r.card, 16);
        *reinterpret_cast<uint4*>(out.card + i * 16) = c;
    }
    if (out.card_len) out.card_len[i] = (uint8_t)r.card_bytes;
    // fused upsert-table insert (the whole-path calls): message i of this launch is message dedup_i0 + i of
    // the batch the table was cleared for; saves a launch and a second pass over msg_id / status
    if (D.tab) {
        if (r.status == SG_ST_UPSERT) {
            dedup_insert_one(dedup_i0 + i, dedup_ids, D);
            // multi-GPU prefilter: mark the key in this rank's bitmap right here (k_bits_build's work, same bit
            // choice as key_bit()) instead of in another pass over msg_id / slot_of on the exchange stream
            if (key_bits) {
                const uint32_t h = reinterpret_cast<const uint32_t*>(dedup_ids + (dedup_i0 + i) * 20)[2] &
                                   (uint32_t)(key_bits_words * 32 - 1);
                atomicOr(&key_bits[h >> 5], 1u << (h & 31));
            }
        } else D.slot_of[dedup_i0 + i] = -1;
    }
}

// ------------------------------------------------------------------ dedup (upsert-by-msg_id): kernels
__global__ void __launch_bounds__(256)
k_dedup_insert(int64_t i0, int64_t n, const uint8_t* __restrict__ msg_id, const uint8_t* __restrict__ status, DedupScratch D) {
    const int64_t i = i0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;   // messages [i0, n) of the batch
    if (i >= n) return;
    if (status && status[i] != SG_ST_UPSERT) {              // status == NULL: every message takes part
        D.slot_of[i] = -1;
        return;
    }
    dedup_insert_one(i, msg_id, D);
}

__global__ void __launch_bounds__(256)
k_dedup_resolve(int64_t n, int64_t base_index, DedupScratch D, uint8_t* __restrict__ survivor,
                uint8_t* __restrict__ first_seen, int64_t* __restrict__ last_idx) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t sl = D.slot_of[i];
    uint8_t sv = 0, fs = 0;
    int64_t li = -1;
    if (sl >= 0) {
        const int2 fl = D.tab[sl];
        const int32_t l = fl.y, f = fl.x;
        sv = (l == (int32_t)i);
        fs = (f == (int32_t)i);
        li = base_index + l;
    }
    if (survivor) survivor[i] = sv;
    if (first_seen) first_seen[i] = fs;
    if (last_idx) last_idx[i] = li;
}

// ------------------------------------------------------------------ batch helpers
__global__ void k_decimal_batch(int64_t n, const int32_t* __restrict__ offsets, const uint8_t* __restrict__ bytes,
                                int64_t* coef, int32_t* exp, uint8_t* flags) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const DecResult r = sg_parse_decimal(bytes + offsets[i], offsets[i + 1] - offsets[i]);
    coef[i] = r.coef;
    exp[i] = r.exp;
    flags[i] = (uint8_t)(r.err | (r.sign << 4));
}
__global__ void k_strict_date_batch(int64_t n, const int32_t* __restrict__ offsets, const uint8_t* __restrict__ bytes,
                                    int64_t* packed) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    packed[i] = sg_parse_strict_datetime(bytes + offsets[i], offsets[i + 1] - offsets[i]);
}

// ------------------------------------------------------------------ device-source keys
// MD5 of every body: the msg_id the API gateway gives a device-sourced message.  One thread per message
// straight from global memory (this is not on the bulk XML path: 64 rounds per block, no staging).
__global__ void __launch_bounds__(256)
k_md5_keys(const uint8_t* __restrict__ bytes, const int64_t* __restrict__ offsets, int64_t n, uint8_t* __restrict__ md5) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t lo = offsets[i];
    uint32_t h[4];
    sg_md5(bytes + lo, (int)(offsets[i + 1] - lo), h);
    *reinterpret_cast<uint4*>(md5 + i * 16) = make_uint4(h[0], h[1], h[2], h[3]);
}
// Which device-sourced messages carry a body that the XML backup also holds?  The reference keys the two
// sources differently (MD5 vs SHA-1 of the same body), so such a body ends up stored twice.  The XML side's
// SHA-1s are in the dedup table (k_dedup_insert with status == NULL); this probes it with the SHA-1 of each
// device-sourced body and reports the first XML message with the same body, or -1.
__global__ void __launch_bounds__(256)
k_xsrc_probe(int64_t n_dev, const uint8_t* __restrict__ dev_sha1, const uint8_t* __restrict__ xml_sha1, DedupScratch D,
             int32_t* __restrict__ match) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_dev) return;
    const uint8_t* k = dev_sha1 + i * 20;
    const uint32_t* kw = reinterpret_cast<const uint32_t*>(k);
    uint64_t pos = (((uint64_t)kw[1] << 32) | kw[0]) & D.mask;
    int32_t found = -1;
    for (;;) {
        const int first = D.tab[pos].x;                  // lowest XML index with the slot's key, -1 = empty
        if (first == -1) break;
        if (id20_equal(xml_sha1 + (int64_t)first * 20, k)) { found = first; break; }
        pos = (pos + 1) & D.mask;
    }
    match[i] = found;
}

// the date string of every cache entry (string 4e+2 of the loader's layout) through the strict parser
__global__ void k_strict_date_entries(int64_t m, const int32_t* __restrict__ val_offsets, const uint8_t* __restrict__ val_bytes,
                                      int64_t* packed) {
    const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (e >= m) return;
    const int32_t a = val_offsets[4 * e + 2], b = val_offsets[4 * e + 3];
    packed[e] = b > a ? sg_parse_strict_datetime(val_bytes + a, b - a) : -1;
}

__global__ void k_dt_str_batch(int64_t n, const int32_t* __restrict__ offsets, const uint8_t* __restrict__ bytes,
                               const DeviceTables* __restrict__ tab, int64_t* packed) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    ScanTables T;
    T.t = tab;
    packed[i] = sg_dt_str(bytes + offsets[i], offsets[i + 1] - offsets[i], T);
}

// ------------------------------------------------------------------ legacy offline replay (process_cached.py)
struct LegacyOut {
    uint8_t* status; uint8_t* lflags; int32_t* strip; int32_t* caps; double* amount; double* balance; uint8_t* body_hash;
};
__global__ void __launch_bounds__(128)
k_legacy_parse(int64_t n, const int64_t* __restrict__ offsets, const uint8_t* __restrict__ bytes, LegacyTables L,
               LegacyOut out, uint32_t mode) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const LegacyResult r = sg_legacy_process(bytes + offsets[i], (int)(offsets[i + 1] - offsets[i]), L, mode);
    out.status[i] = (uint8_t)r.status;
    out.lflags[i] = (uint8_t)r.lflags;
    out.strip[2 * i] = r.strip_lo;
    out.strip[2 * i + 1] = r.strip_hi;
    for (int k = 0; k < 2 * RX_MAX_GROUPS; ++k) out.caps[i * 2 * RX_MAX_GROUPS + k] = r.caps[k];
    out.amount[i] = r.amount;
    out.balance[i] = r.balance;
    uint32_t* h = reinterpret_cast<uint32_t*>(out.body_hash + i * 32);
    for (int k = 0; k < 8; ++k) h[k] = __byte_perm(r.digest[k], 0, 0x0123);
}

// ================================================================== host side / C ABI
struct smsgate_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    // smsgate_replay_host overlaps host->device copies, kernels and device->host copies chunk by chunk
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
    static const int MAX_CHUNKS = 16;
    cudaEvent_t ev_in[MAX_CHUNKS] = {nullptr}, ev_k[MAX_CHUNKS] = {nullptr};
    cudaEvent_t ev_start = nullptr, ev_dedup = nullptr, ev_done = nullptr;
    int64_t host_n = 0, host_total = 0;      // batch of the pending smsgate_replay_host_begin
    int host_chunks = 0;                     // chunks of the last smsgate_replay_host_begin (timeline)
    int64_t view_n = 0, view_total = 0;      // batch whose columns are still in s_out / s_fixed (smsgate_fetch_columns)
    // pending XML extraction (smsgate_xml_scan .. smsgate_xml_emit)
    uint8_t* xml_slab = nullptr;
    int64_t xml_n = 0, xml_T = 0, xml_nx = 0;
    const uint8_t* xml_x = nullptr;
    int64_t *xml_tag_pos = nullptr, *xml_end = nullptr, *xml_sms_tag = nullptr, *xml_body_off = nullptr, *xml_sender_off = nullptr,
            *xml_date_ms = nullptr;
    int32_t *xml_name_off = nullptr, *xml_name_len = nullptr;
    uint8_t *xml_kind = nullptr, *xml_mflags = nullptr;
    uint8_t* xml_attr_slab = nullptr;        // smsgate_xml_attrs_sizes .. _emit: per-message attribute offsets
    int64_t *xml_attr_off = nullptr, *xml_attr_nbase = nullptr, *xml_attr_vbase = nullptr;
    int64_t xml_attr_total = -1;
    int sm_count = 148;
    int scan_ctas = 0;                       // CTAs of k_scan_keys, 0 = one per SM (smsgate_set_scan_ctas)
    int host_div = 3, host_offsets_first = 1;    // smsgate_set_host_pipeline: how smsgate_replay_host* cuts the batch
    int64_t host_min_chunk = 131072;             // (measured best of profiles/r2k_host_pipeline_sweep.json)
    int64_t fuse_bits_words = 0;             // > 0: the join's fused insert also marks the prefilter bitmap (smsgate_p2p_fuse_filter)
    bool bits_ready = false;                 // ... and has done so for the batch in the dedup table
    const int64_t* phone_dt = nullptr;       // per-message phone timestamps for the next join (smsgate_set_phone_dates)
    bool cache_shared = false;               // the cache arrays belong to another context (smsgate_cache_share)
    std::string err;
    int64_t launches = 0;
    DeviceTables* d_tables = nullptr;
    uint8_t* d_legacy = nullptr;            // regex programs + Unicode tables of the legacy path, one slab
    LegacyTables legacy;
    // cache
    int64_t cache_m = 0;
    uint8_t* c_keys = nullptr;
    int2* c_table = nullptr;
    uint64_t c_cap = 0;
    int32_t* c_val_offsets = nullptr;
    uint8_t* c_val_bytes = nullptr;
    uint16_t* c_eflags = nullptr;
    CacheSide* c_side = nullptr;
    int64_t* c_date_packed = nullptr;
    // scratch sized by reserve()
    int64_t cap_msgs = 0, cap_bytes = 0;
    int2* d_dtab = nullptr;                             // [dd_cap] upsert table slots {first, last}
    int32_t* d_slot_of = nullptr;
    uint64_t dd_cap = 0;
    int64_t *d_gfirst = nullptr, *d_glast = nullptr;   // [dd_cap] global first / last index per local slot
    int* d_counts = nullptr;                            // [SMSGATE_MAX_WORLD] export counters / cursors
    int* d_cursor = nullptr;                            // [SMSGATE_MAX_WORLD] peer-exchange cursors
    // peer-memory exchange (smsgate_p2p_reserve)
    int64_t p2p_capacity = 0;
    struct P2PSlot* p2p_table = nullptr;
    int32_t* p2p_slot_of = nullptr;
    int* p2p_overflow = nullptr;
    uint32_t* p2p_bits = nullptr;                       // this rank's key bitmap (prefilter)
    int64_t p2p_bits_words = 0;
    // owner-side merge table (smsgate_reserve_merge)
    int64_t mg_records = 0;
    uint64_t mg_cap = 0;
    int32_t *mg_owner = nullptr, *mg_slot_of = nullptr;
    unsigned long long *mg_first = nullptr, *mg_last = nullptr;
    // device mirror of a host batch (smsgate_replay_host)
    uint8_t* s_bytes = nullptr;
    int64_t* s_offsets = nullptr;
    uint8_t* s_out = nullptr;     // one slab carved into the output columns
    uint8_t* s_fixed = nullptr;
    smsgate_record* s_records = nullptr;   // packed results of smsgate_replay_host_records
    int64_t* s_verdict = nullptr;
};

static thread_local std::string g_create_error;

static int fail(smsgate_ctx* ctx, int code, const std::string& msg) {
    if (ctx) ctx->err = msg; else g_create_error = msg;
    return code;
}
#define CUDA_TRY(ctx, expr)                                                                       \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            return fail(ctx, SMSGATE_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
    } while (0)

static uint64_t pow2_at_least(uint64_t x) {
    uint64_t p = 1;
    while (p < x) p <<= 1;
    return p;
}

#include "xml_extract.cuh"
#include "cache_reader.h"

// ---- the on-disk response cache, read natively (cache_reader.h)
struct smsgate_dcache {
    sgcache::Table T;
};
extern "C" int smsgate_dcache_open(const char* cache_dir, double now, smsgate_dcache** out, char* errbuf, int errbuf_len) {
    if (!cache_dir || !out) return SMSGATE_E_ARG;
    *out = nullptr;
    smsgate_dcache* h = new (std::nothrow) smsgate_dcache();
    if (!h) return SMSGATE_E_NOMEM;
    bool ok = false;
    try {
        const char* forced = getenv("SMSGATE_DCACHE_THREADS");           // tests force the multi-connection path on small caches
        ok = sgcache::read_diskcache(cache_dir, now, &h->T, forced ? atoi(forced) : 0);
    } catch (const std::bad_alloc&) {
        h->T.err = "out of host memory";
    }
    if (!ok) {
        if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "%s", h->T.err.c_str());
        delete h;
        return SMSGATE_E_STATE;
    }
    *out = h;
    return SMSGATE_OK;
}
extern "C" int smsgate_dcache_sizes(const smsgate_dcache* h, int64_t* m, int64_t* value_bytes, int64_t* text_bytes,
                                    int64_t* skipped_keys, int64_t* needs_host) {
    if (!h) return SMSGATE_E_ARG;
    if (m) *m = h->T.m;
    if (value_bytes) *value_bytes = (int64_t)h->T.val_bytes.size();
    if (text_bytes) *text_bytes = (int64_t)h->T.text_bytes.size();
    if (skipped_keys) *skipped_keys = h->T.skipped;
    if (needs_host) *needs_host = h->T.fallbacks;
    return SMSGATE_OK;
}
extern "C" int smsgate_dcache_fill(const smsgate_dcache* h, uint8_t* keys, int32_t* val_offsets, uint8_t* val_bytes,
                                   uint16_t* eflags, int64_t* rowids, uint8_t* needs_host, int64_t* text_offsets,
                                   uint8_t* text_bytes, uint8_t* text_kind) {
    if (!h) return SMSGATE_E_ARG;
    const sgcache::Table& T = h->T;
    if (T.val_bytes.size() >= 0x7fffffffull) return SMSGATE_E_LIMIT;       // int32 offsets, as smsgate_cache_load takes them
    if (keys && T.m) memcpy(keys, T.keys.data(), T.keys.size());
    if (val_offsets) for (size_t k = 0; k < T.val_offsets.size(); ++k) val_offsets[k] = (int32_t)T.val_offsets[k];
    if (val_bytes && !T.val_bytes.empty()) memcpy(val_bytes, T.val_bytes.data(), T.val_bytes.size());
    if (eflags && T.m) memcpy(eflags, T.eflags.data(), T.eflags.size() * 2);
    if (rowids && T.m) memcpy(rowids, T.rowids.data(), T.rowids.size() * 8);
    if (needs_host && T.m) memcpy(needs_host, T.needs_host.data(), T.needs_host.size());
    if (text_offsets) memcpy(text_offsets, T.text_offsets.data(), T.text_offsets.size() * 8);
    if (text_bytes && !T.text_bytes.empty()) memcpy(text_bytes, T.text_bytes.data(), T.text_bytes.size());
    if (text_kind && T.m) memcpy(text_kind, T.text_kind.data(), T.text_kind.size());
    return SMSGATE_OK;
}
extern "C" void smsgate_dcache_close(smsgate_dcache* h) { delete h; }

// ---- bulk emit of sms.parsed payloads / PocketBase records (emit_json.h)
#include "emit_json.h"
extern "C" int smsgate_emit_json(const smsgate_emit_input* in, int mode, int64_t n_sel, const int64_t* sel, int64_t* out_offsets,
                                 uint8_t* out_bytes, int64_t capacity, uint8_t* fallback, int64_t* needed) {
    if (!in || n_sel < 0 || (n_sel > 0 && !sel) || !out_offsets || (mode != 0 && mode != 1)) return SMSGATE_E_ARG;
    sgemit::In I;
    I.msg_id = in->msg_id; I.status = in->status; I.flags = in->flags; I.dec_flags = in->dec_flags; I.slot = in->slot;
    I.amount_coef = in->amount_coef; I.balance_coef = in->balance_coef; I.amount_exp = in->amount_exp; I.balance_exp = in->balance_exp;
    I.date = in->date; I.card = in->card; I.card_len = in->card_len; I.body_offsets = in->body_offsets; I.body_bytes = in->body_bytes;
    I.fixed_len = in->fixed_len; I.fixed_bytes = in->fixed_bytes; I.sender_offsets = in->sender_offsets; I.sender_bytes = in->sender_bytes;
    I.device_id = in->device_id; I.text_offsets = in->text_offsets; I.text_bytes = in->text_bytes; I.text_kind = in->text_kind;
    I.record_stride = in->records ? (int64_t)sizeof(smsgate_record) : 0;
    if (in->records) I.msg_id = reinterpret_cast<const uint8_t*>(in->records);
    if (!I.msg_id || !I.card || !I.card_len || !I.body_offsets || !I.sender_offsets || !I.text_offsets || !I.text_kind ||
        (!in->records && (!I.flags || !I.dec_flags || !I.slot || !I.amount_coef || !I.balance_coef || !I.amount_exp || !I.balance_exp || !I.date)))
        return SMSGATE_E_ARG;
    int threads = 0;
    if (const char* forced = getenv("SMSGATE_EMIT_THREADS")) threads = atoi(forced) > 0 ? atoi(forced) : 1;
    return sgemit::emit_many(I, mode, n_sel, sel, out_offsets, out_bytes, capacity, fallback, needed, threads) ? SMSGATE_OK
                                                                                                               : SMSGATE_E_LIMIT;
}

extern "C" int smsgate_abi_version(void) { return SMSGATE_ABI_VERSION; }

extern "C" const char* smsgate_last_error(const smsgate_ctx* ctx) {
    return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

extern "C" int64_t smsgate_launch_count(const smsgate_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" int64_t smsgate_cache_size(const smsgate_ctx* ctx) { return ctx ? ctx->cache_m : 0; }

extern "C" int smsgate_create(int device, void* cuda_stream, smsgate_ctx** out) {
    if (!out) return fail(nullptr, SMSGATE_E_ARG, "out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, SMSGATE_E_CUDA,
                    std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "count is 0") +
                        " (this library has no CPU path)");
    if (device < 0 || device >= count) return fail(nullptr, SMSGATE_E_ARG, "device index out of range");
    smsgate_ctx* ctx = new (std::nothrow) smsgate_ctx();
    if (!ctx) return fail(nullptr, SMSGATE_E_NOMEM, "out of host memory");
    ctx->device = device;
    auto bail = [&](const char* what, cudaError_t ce) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(ce);
        delete ctx;
        return SMSGATE_E_CUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail("cudaGetDeviceProperties", e);
    if (prop.major < 10) {
        g_create_error = "device is not sm_100 class (compute capability " + std::to_string(prop.major) + "." +
                         std::to_string(prop.minor) + "); this library carries sm_100a code only";
        delete ctx;
        return SMSGATE_E_CUDA;
    }
    ctx->sm_count = prop.multiProcessorCount;
    ctx->stream = (cudaStream_t)cuda_stream;          // NULL is CUDA's default stream
    if ((e = cudaStreamCreateWithFlags(&ctx->s_h2d, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
    if ((e = cudaStreamCreateWithFlags(&ctx->s_d2h, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
    for (int k = 0; k < smsgate_ctx::MAX_CHUNKS; ++k) {
        if ((e = cudaEventCreate(&ctx->ev_in[k])) != cudaSuccess) return bail("cudaEventCreate", e);
        if ((e = cudaEventCreate(&ctx->ev_k[k])) != cudaSuccess) return bail("cudaEventCreate", e);
    }
    if ((e = cudaEventCreate(&ctx->ev_start)) != cudaSuccess) return bail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&ctx->ev_dedup)) != cudaSuccess) return bail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&ctx->ev_done)) != cudaSuccess) return bail("cudaEventCreate", e);
    DeviceTables* h = new DeviceTables();
    sg_fill_tables(h);
    if ((e = cudaMalloc(&ctx->d_tables, sizeof(DeviceTables))) != cudaSuccess) { delete h; return bail("cudaMalloc tables", e); }
    e = cudaMemcpy(ctx->d_tables, h, sizeof(DeviceTables), cudaMemcpyHostToDevice);
    delete h;
    if (e != cudaSuccess) return bail("cudaMemcpy tables", e);
    {   // legacy-path tables: one slab, 16-byte aligned pieces
        const size_t sz[6] = {sizeof(H_RX_DEBIT), sizeof(H_RX_CREDIT), sizeof(H_RX_WORD_LO), sizeof(H_RX_WORD_HI),
                              sizeof(H_RX_CI_EXC), sizeof(H_LEG_OTP_NEXT)};
        const void* src[6] = {H_RX_DEBIT, H_RX_CREDIT, H_RX_WORD_LO, H_RX_WORD_HI, H_RX_CI_EXC, H_LEG_OTP_NEXT};
        size_t off[6], total = 0;
        for (int k = 0; k < 6; ++k) { off[k] = total; total += (sz[k] + 15) & ~(size_t)15; }
        if ((e = cudaMalloc(&ctx->d_legacy, total)) != cudaSuccess) return bail("cudaMalloc legacy tables", e);
        for (int k = 0; k < 6; ++k)
            if ((e = cudaMemcpy(ctx->d_legacy + off[k], src[k], sz[k], cudaMemcpyHostToDevice)) != cudaSuccess)
                return bail("cudaMemcpy legacy tables", e);
        ctx->legacy.debit = reinterpret_cast<const uint32_t*>(ctx->d_legacy + off[0]);
        ctx->legacy.credit = reinterpret_cast<const uint32_t*>(ctx->d_legacy + off[1]);
        ctx->legacy.word_lo = reinterpret_cast<const uint32_t*>(ctx->d_legacy + off[2]);
        ctx->legacy.word_hi = reinterpret_cast<const uint32_t*>(ctx->d_legacy + off[3]);
        ctx->legacy.ci_exc = reinterpret_cast<const uint32_t*>(ctx->d_legacy + off[4]);
        ctx->legacy.otp_next = ctx->d_legacy + off[5];
        ctx->legacy.scan.t = ctx->d_tables;
    }
    if ((e = cudaFuncSetAttribute(k_scan_keys, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K1Smem))) != cudaSuccess)
        return bail("cudaFuncSetAttribute(k_scan_keys)", e);
    *out = ctx;
    return SMSGATE_OK;
}

static void free_cache(smsgate_ctx* c) {
    if (c->cache_shared) {
        c->c_keys = nullptr; c->c_table = nullptr; c->c_val_offsets = nullptr; c->c_val_bytes = nullptr;
        c->c_eflags = nullptr; c->c_date_packed = nullptr; c->c_side = nullptr;
        c->cache_m = 0; c->c_cap = 0; c->cache_shared = false;
        return;
    }
    cudaFree(c->c_keys); cudaFree(c->c_table); cudaFree(c->c_val_offsets); cudaFree(c->c_val_bytes);
    cudaFree(c->c_eflags); cudaFree(c->c_date_packed); cudaFree(c->c_side);
    c->c_keys = nullptr; c->c_table = nullptr; c->c_val_offsets = nullptr; c->c_val_bytes = nullptr;
    c->c_eflags = nullptr; c->c_date_packed = nullptr; c->c_side = nullptr;
    c->cache_m = 0; c->c_cap = 0;
}
static void free_scratch(smsgate_ctx* c) {
    cudaFree(c->d_dtab); cudaFree(c->d_slot_of);
    cudaFree(c->d_gfirst); cudaFree(c->d_glast); cudaFree(c->d_counts); cudaFree(c->d_cursor);
    c->d_gfirst = c->d_glast = nullptr; c->d_counts = nullptr; c->d_cursor = nullptr;
    cudaFree(c->s_bytes); cudaFree(c->s_offsets); cudaFree(c->s_out); cudaFree(c->s_fixed); cudaFree(c->s_records); cudaFree(c->s_verdict);
    c->d_dtab = nullptr; c->d_slot_of = nullptr;
    c->s_bytes = nullptr; c->s_offsets = nullptr; c->s_out = nullptr; c->s_fixed = nullptr; c->s_records = nullptr; c->s_verdict = nullptr;
    c->cap_msgs = c->cap_bytes = 0; c->dd_cap = 0;
}

extern "C" void smsgate_destroy(smsgate_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    smsgate_xml_release(ctx);                            // a scan that was never emitted still holds the document
    free_cache(ctx);
    free_scratch(ctx);
    cudaFree(ctx->mg_owner); cudaFree(ctx->mg_first); cudaFree(ctx->mg_last); cudaFree(ctx->mg_slot_of);
    cudaFree(ctx->p2p_table); cudaFree(ctx->p2p_slot_of); cudaFree(ctx->p2p_overflow); cudaFree(ctx->p2p_bits);
    cudaFree(ctx->d_legacy);
    cudaFree(ctx->d_tables);
    if (ctx->s_h2d) { cudaStreamSynchronize(ctx->s_h2d); cudaStreamDestroy(ctx->s_h2d); }
    if (ctx->s_d2h) { cudaStreamSynchronize(ctx->s_d2h); cudaStreamDestroy(ctx->s_d2h); }
    for (int k = 0; k < smsgate_ctx::MAX_CHUNKS; ++k) {
        if (ctx->ev_in[k]) cudaEventDestroy(ctx->ev_in[k]);
        if (ctx->ev_k[k]) cudaEventDestroy(ctx->ev_k[k]);
    }
    if (ctx->ev_start) cudaEventDestroy(ctx->ev_start);
    if (ctx->ev_dedup) cudaEventDestroy(ctx->ev_dedup);
    if (ctx->ev_done) cudaEventDestroy(ctx->ev_done);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

// bytes of the output slab per message (smsgate_replay_host): see carve_slab()
static const int64_t SLAB_PER_MSG = 20 + 32 + 1 + 4 + 4 + 4 + 1 + 8 + 8 + 4 + 4 + 1 + 8 + 16 + 1 + 1 + 1 + 8 + 8 + 8;

extern "C" int smsgate_reserve(smsgate_ctx* ctx, int64_t max_msgs, int64_t max_bytes) {
    if (!ctx) return SMSGATE_E_ARG;
    if (max_msgs < 0 || max_bytes < 0) return fail(ctx, SMSGATE_E_ARG, "negative size");
    if (max_msgs > 0x40000000ll) return fail(ctx, SMSGATE_E_LIMIT, "a batch holds at most 2^30 messages");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (max_msgs <= ctx->cap_msgs && max_bytes <= ctx->cap_bytes) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    free_scratch(ctx);
    ctx->view_n = ctx->view_total = 0;
    const uint64_t cap = pow2_at_least((uint64_t)(max_msgs > 8 ? max_msgs : 8) * 2);
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_dtab, cap * sizeof(int2)));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_slot_of, (size_t)(max_msgs + 1) * 4));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_gfirst, cap * 8));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_glast, cap * 8));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_counts, sizeof(int) * SMSGATE_MAX_WORLD));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_cursor, sizeof(int) * SMSGATE_MAX_WORLD));
    CUDA_TRY(ctx, cudaMalloc(&ctx->s_bytes, (size_t)max_bytes + 64));
    CUDA_TRY(ctx, cudaMalloc(&ctx->s_fixed, (size_t)max_bytes + 64));
    CUDA_TRY(ctx, cudaMalloc(&ctx->s_offsets, (size_t)(max_msgs + 1) * 8));
    CUDA_TRY(ctx, cudaMalloc(&ctx->s_out, (size_t)(max_msgs + 16) * SLAB_PER_MSG + 65536));
    CUDA_TRY(ctx, cudaMalloc(&ctx->s_records, (size_t)(max_msgs + 1) * sizeof(smsgate_record)));
    CUDA_TRY(ctx, cudaMalloc(&ctx->s_verdict, (size_t)(max_msgs + 1) * 8));
    ctx->dd_cap = cap;
    ctx->cap_msgs = max_msgs;
    ctx->cap_bytes = max_bytes;
    return SMSGATE_OK;
}

extern "C" int smsgate_cache_load(smsgate_ctx* ctx, int64_t m, const uint8_t* h_keys, const int32_t* h_val_offsets,
                                  const uint8_t* h_val_bytes, const uint16_t* h_eflags, const int64_t* h_date_packed) {
    if (!ctx) return SMSGATE_E_ARG;
    if (m < 0 || (m > 0 && (!h_keys || !h_val_offsets || !h_eflags || !h_date_packed)))
        return fail(ctx, SMSGATE_E_ARG, "cache_load: NULL array");
    if (m > 0x3FFFFFFFll) return fail(ctx, SMSGATE_E_LIMIT, "cache holds at most 2^30 entries");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    free_cache(ctx);
    if (m == 0) return SMSGATE_OK;
    const int64_t vbytes = h_val_offsets[4 * m];
    if (vbytes < 0 || (vbytes > 0 && !h_val_bytes)) return fail(ctx, SMSGATE_E_ARG, "cache_load: bad value offsets");
    const uint64_t cap = pow2_at_least((uint64_t)m * 2);
    CUDA_TRY(ctx, cudaMalloc(&ctx->c_keys, (size_t)m * 32));
    CUDA_TRY(ctx, cudaMalloc(&ctx->c_table, cap * sizeof(int2)));
    CUDA_TRY(ctx, cudaMalloc(&ctx->c_val_offsets, (size_t)(4 * m + 1) * 4));
    CUDA_TRY(ctx, cudaMalloc(&ctx->c_val_bytes, (size_t)vbytes + 16));
    CUDA_TRY(ctx, cudaMalloc(&ctx->c_eflags, (size_t)m * 2));
    CUDA_TRY(ctx, cudaMalloc(&ctx->c_date_packed, (size_t)m * 8));
    CUDA_TRY(ctx, cudaMalloc(&ctx->c_side, (size_t)m * sizeof(CacheSide)));
    cudaStream_t st = ctx->stream;
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->c_keys, h_keys, (size_t)m * 32, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->c_val_offsets, h_val_offsets, (size_t)(4 * m + 1) * 4, cudaMemcpyHostToDevice, st));
    if (vbytes) CUDA_TRY(ctx, cudaMemcpyAsync(ctx->c_val_bytes, h_val_bytes, (size_t)vbytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->c_eflags, h_eflags, (size_t)m * 2, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->c_date_packed, h_date_packed, (size_t)m * 8, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->c_table, 0xFF, cap * sizeof(int2), st));
    int* d_dup = nullptr;
    CUDA_TRY(ctx, cudaMalloc(&d_dup, 4));
    CUDA_TRY(ctx, cudaMemsetAsync(d_dup, 0, 4, st));
    k_cache_build<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(ctx->c_keys, m, ctx->c_table, cap - 1, d_dup);
    k_cache_fix_fp<<<(unsigned)((cap + 255) / 256), 256, 0, st>>>(ctx->c_keys, ctx->c_table, cap);
    k_cache_side<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(m, ctx->c_val_offsets, ctx->c_eflags, ctx->c_side);
    ctx->launches += 3;
    int dup = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&dup, d_dup, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    cudaFree(d_dup);
    CUDA_TRY(ctx, cudaGetLastError());
    if (dup) {
        free_cache(ctx);
        return fail(ctx, SMSGATE_E_ARG, "cache_load: " + std::to_string(dup) + " duplicate key(s)");
    }
    ctx->cache_m = m;
    ctx->c_cap = cap;
    return SMSGATE_OK;
}

extern "C" int smsgate_cache_share(smsgate_ctx* ctx, const smsgate_ctx* owner) {
    if (!ctx || !owner || ctx == owner) return ctx ? fail(ctx, SMSGATE_E_ARG, "cache_share: bad argument") : SMSGATE_E_ARG;
    if (owner->device != ctx->device) return fail(ctx, SMSGATE_E_ARG, "cache_share: the contexts are on different devices");
    if (owner->cache_shared) return fail(ctx, SMSGATE_E_ARG, "cache_share: share with the context that loaded the cache");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    free_cache(ctx);
    ctx->c_keys = owner->c_keys; ctx->c_table = owner->c_table; ctx->c_val_offsets = owner->c_val_offsets;
    ctx->c_val_bytes = owner->c_val_bytes; ctx->c_eflags = owner->c_eflags; ctx->c_date_packed = owner->c_date_packed;
    ctx->c_side = owner->c_side; ctx->cache_m = owner->cache_m; ctx->c_cap = owner->c_cap;
    ctx->cache_shared = owner->cache_m > 0;
    return SMSGATE_OK;
}

extern "C" int smsgate_set_stream(smsgate_ctx* ctx, void* cuda_stream) {
    if (!ctx) return SMSGATE_E_ARG;
    ctx->stream = (cudaStream_t)cuda_stream;
    return SMSGATE_OK;
}

extern "C" int smsgate_set_phone_dates(smsgate_ctx* ctx, const int64_t* d_phone_dt) {
    if (!ctx) return SMSGATE_E_ARG;
    ctx->phone_dt = d_phone_dt;
    return SMSGATE_OK;
}

extern "C" int smsgate_set_host_pipeline(smsgate_ctx* ctx, int divisor, int64_t min_chunk, int offsets_first) {
    if (!ctx) return SMSGATE_E_ARG;
    if (divisor < 1 || divisor > 16 || min_chunk < 1024) return fail(ctx, SMSGATE_E_ARG, "set_host_pipeline: divisor in 1..16, min_chunk >= 1024");
    ctx->host_div = divisor;
    ctx->host_min_chunk = min_chunk;
    ctx->host_offsets_first = offsets_first ? 1 : 0;
    return SMSGATE_OK;
}

extern "C" int smsgate_set_scan_ctas(smsgate_ctx* ctx, int ctas) {
    if (!ctx) return SMSGATE_E_ARG;
    if (ctas < 0 || ctas > ctx->sm_count) return fail(ctx, SMSGATE_E_ARG, "set_scan_ctas: between 0 (default) and the SM count");
    ctx->scan_ctas = ctas;
    return SMSGATE_OK;
}

// ---- batch helpers
extern "C" int smsgate_decimal_batch(smsgate_ctx* ctx, int64_t n, const int32_t* h_offsets, const uint8_t* h_bytes,
                                     int64_t* h_coef, int32_t* h_exp, uint8_t* h_flags) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || (n > 0 && (!h_offsets || !h_coef || !h_exp || !h_flags))) return fail(ctx, SMSGATE_E_ARG, "decimal_batch: NULL array");
    if (n == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int64_t nb = h_offsets[n];
    int32_t* d_off = nullptr; uint8_t* d_b = nullptr; int64_t* d_c = nullptr; int32_t* d_e = nullptr; uint8_t* d_f = nullptr;
    cudaStream_t st = ctx->stream;
    int rc = SMSGATE_OK;
    cudaError_t e;
#define HB_TRY(expr) if ((e = (expr)) != cudaSuccess) { rc = fail(ctx, SMSGATE_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e)); goto done; }
    HB_TRY(cudaMalloc(&d_off, (size_t)(n + 1) * 4));
    HB_TRY(cudaMalloc(&d_b, (size_t)nb + 16));
    HB_TRY(cudaMalloc(&d_c, (size_t)n * 8));
    HB_TRY(cudaMalloc(&d_e, (size_t)n * 4));
    HB_TRY(cudaMalloc(&d_f, (size_t)n));
    HB_TRY(cudaMemcpyAsync(d_off, h_offsets, (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, st));
    if (nb) HB_TRY(cudaMemcpyAsync(d_b, h_bytes, (size_t)nb, cudaMemcpyHostToDevice, st));
    k_decimal_batch<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, d_off, d_b, d_c, d_e, d_f);
    ctx->launches += 1;
    HB_TRY(cudaMemcpyAsync(h_coef, d_c, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    HB_TRY(cudaMemcpyAsync(h_exp, d_e, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    HB_TRY(cudaMemcpyAsync(h_flags, d_f, (size_t)n, cudaMemcpyDeviceToHost, st));
    HB_TRY(cudaStreamSynchronize(st));
    HB_TRY(cudaGetLastError());
done:
    cudaFree(d_off); cudaFree(d_b); cudaFree(d_c); cudaFree(d_e); cudaFree(d_f);
    return rc;
}

extern "C" int smsgate_strict_date_batch(smsgate_ctx* ctx, int64_t n, const int32_t* h_offsets, const uint8_t* h_bytes,
                                         int64_t* h_packed) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || (n > 0 && (!h_offsets || !h_packed))) return fail(ctx, SMSGATE_E_ARG, "strict_date_batch: NULL array");
    if (n == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int64_t nb = h_offsets[n];
    int32_t* d_off = nullptr; uint8_t* d_b = nullptr; int64_t* d_p = nullptr;
    cudaStream_t st = ctx->stream;
    int rc = SMSGATE_OK;
    cudaError_t e;
    HB_TRY(cudaMalloc(&d_off, (size_t)(n + 1) * 4));
    HB_TRY(cudaMalloc(&d_b, (size_t)nb + 16));
    HB_TRY(cudaMalloc(&d_p, (size_t)n * 8));
    HB_TRY(cudaMemcpyAsync(d_off, h_offsets, (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, st));
    if (nb) HB_TRY(cudaMemcpyAsync(d_b, h_bytes, (size_t)nb, cudaMemcpyHostToDevice, st));
    k_strict_date_batch<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, d_off, d_b, d_p);
    ctx->launches += 1;
    HB_TRY(cudaMemcpyAsync(h_packed, d_p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    HB_TRY(cudaStreamSynchronize(st));
    HB_TRY(cudaGetLastError());
done:
    cudaFree(d_off); cudaFree(d_b); cudaFree(d_p);
    return rc;
}

extern "C" int smsgate_strict_date_entries(smsgate_ctx* ctx, int64_t m, const int32_t* h_val_offsets, const uint8_t* h_val_bytes,
                                           int64_t* h_packed) {
    if (!ctx) return SMSGATE_E_ARG;
    if (m < 0 || (m > 0 && (!h_val_offsets || !h_packed))) return fail(ctx, SMSGATE_E_ARG, "strict_date_entries: NULL array");
    if (m == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int64_t nb = h_val_offsets[4 * m];
    int32_t* d_off = nullptr; uint8_t* d_b = nullptr; int64_t* d_p = nullptr;
    cudaStream_t st = ctx->stream;
    int rc = SMSGATE_OK;
    cudaError_t e;
    HB_TRY(cudaMalloc(&d_off, (size_t)(4 * m + 1) * 4));
    HB_TRY(cudaMalloc(&d_b, (size_t)nb + 16));
    HB_TRY(cudaMalloc(&d_p, (size_t)m * 8));
    HB_TRY(cudaMemcpyAsync(d_off, h_val_offsets, (size_t)(4 * m + 1) * 4, cudaMemcpyHostToDevice, st));
    if (nb) HB_TRY(cudaMemcpyAsync(d_b, h_val_bytes, (size_t)nb, cudaMemcpyHostToDevice, st));
    k_strict_date_entries<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(m, d_off, d_b, d_p);
    ctx->launches += 1;
    HB_TRY(cudaMemcpyAsync(h_packed, d_p, (size_t)m * 8, cudaMemcpyDeviceToHost, st));
    HB_TRY(cudaStreamSynchronize(st));
    HB_TRY(cudaGetLastError());
done:
    cudaFree(d_off); cudaFree(d_b); cudaFree(d_p);
    return rc;
}

extern "C" int smsgate_dt_str_batch(smsgate_ctx* ctx, int64_t n, const int32_t* h_offsets, const uint8_t* h_bytes,
                                         int64_t* h_packed) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || (n > 0 && (!h_offsets || !h_packed))) return fail(ctx, SMSGATE_E_ARG, "dt_str_batch: NULL array");
    if (n == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int64_t nb = h_offsets[n];
    int32_t* d_off = nullptr; uint8_t* d_b = nullptr; int64_t* d_p = nullptr;
    cudaStream_t st = ctx->stream;
    int rc = SMSGATE_OK;
    cudaError_t e;
    HB_TRY(cudaMalloc(&d_off, (size_t)(n + 1) * 4));
    HB_TRY(cudaMalloc(&d_b, (size_t)nb + 16));
    HB_TRY(cudaMalloc(&d_p, (size_t)n * 8));
    HB_TRY(cudaMemcpyAsync(d_off, h_offsets, (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, st));
    if (nb) HB_TRY(cudaMemcpyAsync(d_b, h_bytes, (size_t)nb, cudaMemcpyHostToDevice, st));
    k_dt_str_batch<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, d_off, d_b, ctx->d_tables, d_p);
    ctx->launches += 1;
    HB_TRY(cudaMemcpyAsync(h_packed, d_p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    HB_TRY(cudaStreamSynchronize(st));
    HB_TRY(cudaGetLastError());
done:
    cudaFree(d_off); cudaFree(d_b); cudaFree(d_p);
    return rc;
#undef HB_TRY
}

extern "C" int smsgate_legacy_parse(smsgate_ctx* ctx, int64_t n, const int64_t* h_offsets, const uint8_t* h_bytes,
                                    const smsgate_legacy_outputs* h_out) {
    return smsgate_legacy_parse_mode(ctx, n, h_offsets, h_bytes, 0, h_out);
}

extern "C" int smsgate_legacy_parse_mode(smsgate_ctx* ctx, int64_t n, const int64_t* h_offsets, const uint8_t* h_bytes,
                                         uint32_t mode, const smsgate_legacy_outputs* h_out) {
    if (!ctx) return SMSGATE_E_ARG;
    if (mode > 0xF) return fail(ctx, SMSGATE_E_ARG, "legacy_parse: unknown mode bits");
    if (n < 0 || !h_out || (n > 0 && (!h_offsets || !h_out->status || !h_out->lflags || !h_out->strip || !h_out->caps ||
                                      !h_out->amount || !h_out->balance || !h_out->body_hash)))
        return fail(ctx, SMSGATE_E_ARG, "legacy_parse: NULL array");
    if (n == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int64_t nb = h_offsets[n] - h_offsets[0];
    if (nb < 0 || (nb > 0 && !h_bytes)) return fail(ctx, SMSGATE_E_ARG, "legacy_parse: bad offsets");
    cudaStream_t st = ctx->stream;
    uint8_t* slab = nullptr;
    const size_t per = 1 + 1 + 8 + 8 * RX_MAX_GROUPS + 8 + 8 + 32;
    const size_t need = (size_t)(n + 1) * 8 + (size_t)nb + 64 + (size_t)n * per + 8 * 256;
    int rc = SMSGATE_OK;
    cudaError_t e;
#define LG_TRY(expr) if ((e = (expr)) != cudaSuccess) { rc = fail(ctx, SMSGATE_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e)); goto done; }
    LG_TRY(cudaMalloc(&slab, need));
    {
        uint8_t* p = slab;
        auto take = [&](size_t b) { uint8_t* q = p; p += (b + 255) & ~(size_t)255; return q; };
        int64_t* d_off = (int64_t*)take((size_t)(n + 1) * 8);
        uint8_t* d_bytes = take((size_t)nb + 64);
        LegacyOut o;
        o.amount = (double*)take((size_t)n * 8);
        o.balance = (double*)take((size_t)n * 8);
        o.strip = (int32_t*)take((size_t)n * 8);
        o.caps = (int32_t*)take((size_t)n * 8 * RX_MAX_GROUPS);
        o.body_hash = take((size_t)n * 32);
        o.status = take((size_t)n);
        o.lflags = take((size_t)n);
        LG_TRY(cudaMemcpyAsync(d_off, h_offsets, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, st));
        if (nb) LG_TRY(cudaMemcpyAsync(d_bytes, h_bytes + h_offsets[0], (size_t)nb, cudaMemcpyHostToDevice, st));
        // the kernel indexes bytes with the caller's absolute offsets
        k_legacy_parse<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(n, d_off, d_bytes - h_offsets[0], ctx->legacy, o, mode);
        ctx->launches += 1;
        LG_TRY(cudaMemcpyAsync(h_out->status, o.status, (size_t)n, cudaMemcpyDeviceToHost, st));
        LG_TRY(cudaMemcpyAsync(h_out->lflags, o.lflags, (size_t)n, cudaMemcpyDeviceToHost, st));
        LG_TRY(cudaMemcpyAsync(h_out->strip, o.strip, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
        LG_TRY(cudaMemcpyAsync(h_out->caps, o.caps, (size_t)n * 8 * RX_MAX_GROUPS, cudaMemcpyDeviceToHost, st));
        LG_TRY(cudaMemcpyAsync(h_out->amount, o.amount, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
        LG_TRY(cudaMemcpyAsync(h_out->balance, o.balance, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
        LG_TRY(cudaMemcpyAsync(h_out->body_hash, o.body_hash, (size_t)n * 32, cudaMemcpyDeviceToHost, st));
        LG_TRY(cudaStreamSynchronize(st));
        LG_TRY(cudaGetLastError());
    }
done:
    cudaFree(slab);
    return rc;
#undef LG_TRY
}

// ---- XML extraction
// exclusive prefix sum of n int32 values into n + 1 int64 values (out[n] = total); tmp holds ceil(n/1024)+1 int64
static void xml_scan_i32(cudaStream_t st, const int32_t* in, int64_t n, int64_t* tmp, int64_t* out) {
    const int64_t nb = (n + SCAN_ITEMS - 1) / SCAN_ITEMS;
    k_scan_sums<<<(unsigned)nb, 256, 0, st>>>(in, n, tmp);
    k_scan_top<<<1, 1024, 0, st>>>(tmp, nb, nullptr);
    k_scan_apply<<<(unsigned)nb, 256, 0, st>>>(in, n, tmp, out);
}

extern "C" int smsgate_xml_release(smsgate_ctx* ctx) {
    if (!ctx) return SMSGATE_E_ARG;
    if (ctx->xml_slab) {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        cudaFree(ctx->xml_slab);
        cudaFree((void*)ctx->xml_x);
        cudaFree(ctx->xml_attr_slab);
    }
    ctx->xml_attr_slab = nullptr;
    ctx->xml_attr_total = -1;
    ctx->xml_slab = nullptr;
    ctx->xml_x = nullptr;
    ctx->xml_n = ctx->xml_T = ctx->xml_nx = 0;
    return SMSGATE_OK;
}

extern "C" int smsgate_xml_scan(smsgate_ctx* ctx, const uint8_t* h_xml, int64_t nbytes, int64_t* n_out,
                                int64_t* body_bytes, int64_t* sender_bytes, int64_t* err_pos) {
    if (!ctx) return SMSGATE_E_ARG;
    if (nbytes < 0 || !n_out || !body_bytes || !sender_bytes || (nbytes > 0 && !h_xml))
        return fail(ctx, SMSGATE_E_ARG, "xml_scan: bad argument");
    smsgate_xml_release(ctx);
    *n_out = *body_bytes = *sender_bytes = 0;
    if (err_pos) *err_pos = -1;
    if (nbytes == 0) { if (err_pos) *err_pos = 0; return fail(ctx, SMSGATE_E_XML, "xml: no element found"); }
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int64_t nchunks = (nbytes + XML_CHUNK - 1) / XML_CHUNK;
    // stage A: the document and the '<' positions
    uint8_t* d_x = nullptr;
    int32_t* d_cnt = nullptr;
    int64_t *d_off = nullptr, *d_tmp = nullptr;
    int rc = SMSGATE_OK;
    cudaError_t e;
    int64_t T = 0, n = 0, nblocks = 0, tot_body = 0, tot_sender = 0;
    unsigned long long h_err = ~0ull;
    int h_roots = 0;
    uint8_t* slab = nullptr;
#define XT(expr) if ((e = (expr)) != cudaSuccess) { rc = fail(ctx, SMSGATE_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e)); goto done; }
    XT(cudaMalloc(&d_x, (size_t)nbytes + 64));
    XT(cudaMalloc(&d_cnt, (size_t)nchunks * 4));
    XT(cudaMalloc(&d_off, (size_t)(nchunks + 1) * 8));
    XT(cudaMalloc(&d_tmp, (size_t)(nbytes / SCAN_ITEMS + 1024) * 8));
    XT(cudaMemcpyAsync(d_x, h_xml, (size_t)nbytes, cudaMemcpyHostToDevice, st));
    XT(cudaMemsetAsync(d_x + nbytes, 0, 64, st));
    k_xml_count_lt<<<(unsigned)nchunks, 256, 0, st>>>(d_x, nbytes, d_cnt);
    xml_scan_i32(st, d_cnt, nchunks, d_tmp, d_off);
    XT(cudaMemcpyAsync(&T, d_off + nchunks, 8, cudaMemcpyDeviceToHost, st));
    XT(cudaStreamSynchronize(st));
    ctx->launches += 4;
    if (T == 0) { if (err_pos) *err_pos = 0; rc = fail(ctx, SMSGATE_E_XML, "xml: not well-formed (no markup)"); goto done; }
    {
        // one slab for everything sized by T
        auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
        const size_t need = al((T + 1) * 8) * 5 + al(T * 4) * 5 + al(T) + al((T + 1) * 8) * 2 + 4096;
        XT(cudaMalloc(&slab, need));
        uint8_t* p = slab;
        auto take = [&](size_t b) { uint8_t* q = p; p += al(b); return q; };
        int64_t* tag_pos = (int64_t*)take((T + 1) * 8);
        int64_t* tag_end = (int64_t*)take((T + 1) * 8);
        int64_t* depth = (int64_t*)take((T + 1) * 8);
        int64_t* sel_off = (int64_t*)take((T + 1) * 8);
        int64_t* blk_off = (int64_t*)take((T + 1) * 8);
        int32_t* name_off = (int32_t*)take(T * 4);
        int32_t* name_len = (int32_t*)take(T * 4);
        int32_t* delta = (int32_t*)take(T * 4);
        int32_t* is_sms = (int32_t*)take(T * 4);
        int32_t* is_block = (int32_t*)take(T * 4);
        uint8_t* kind = take(T);
        int64_t* sms_tag = (int64_t*)take((T + 1) * 8);
        int64_t* blocks = (int64_t*)take((T + 1) * 8);
        unsigned long long* d_err = (unsigned long long*)take(16);
        int* d_roots = (int*)take(16);
        XT(cudaMemsetAsync(d_err, 0xFF, 8, st));
        XT(cudaMemsetAsync(d_roots, 0, 4, st));
        k_xml_write_lt<<<(unsigned)nchunks, 256, 0, st>>>(d_x, nbytes, d_off, tag_pos);
        XmlTags G{tag_pos, tag_end, name_off, name_len, kind, delta, is_sms, is_block};
        const unsigned gt = (unsigned)((T + 127) / 128);
        k_xml_parse_tags<<<gt, 128, 0, st>>>(d_x, nbytes, T, G, d_err);
        xml_scan_i32(st, delta, T, d_tmp, depth);
        k_xml_structure<<<gt, 128, 0, st>>>(d_x, nbytes, T, G, depth, d_err, d_roots);
        xml_scan_i32(st, is_sms, T, d_tmp, sel_off);
        xml_scan_i32(st, is_block, T, d_tmp, blk_off);
        k_xml_compact<<<(unsigned)((T + 255) / 256), 256, 0, st>>>(T, is_sms, sel_off, sms_tag);
        k_xml_compact<<<(unsigned)((T + 255) / 256), 256, 0, st>>>(T, is_block, blk_off, blocks);
        XT(cudaMemcpyAsync(&n, sel_off + T, 8, cudaMemcpyDeviceToHost, st));
        XT(cudaMemcpyAsync(&nblocks, blk_off + T, 8, cudaMemcpyDeviceToHost, st));
        XT(cudaStreamSynchronize(st));
        k_xml_match_names<<<1, 32, 0, st>>>(d_x, G, blocks, nblocks, d_err);
        XT(cudaMemcpyAsync(&h_err, d_err, 8, cudaMemcpyDeviceToHost, st));
        XT(cudaMemcpyAsync(&h_roots, d_roots, 4, cudaMemcpyDeviceToHost, st));
        XT(cudaStreamSynchronize(st));
        ctx->launches += 16;
        if (h_err != ~0ull) {
            if (err_pos) *err_pos = (int64_t)(h_err >> 2);
            const bool unsup = (h_err & 3) == XML_UNSUPPORTED;
            rc = fail(ctx, unsup ? SMSGATE_E_UNSUPPORTED : SMSGATE_E_XML,
                      std::string(unsup ? "xml: construct outside the supported dialect at byte " : "xml: not well-formed at byte ") +
                          std::to_string((long long)(h_err >> 2)));
            goto done;
        }
        if (h_roots != 1) { if (err_pos) *err_pos = 0; rc = fail(ctx, SMSGATE_E_XML, "xml: not exactly one root element"); goto done; }
        // stage B: per-message sizes, offsets
        uint8_t* slab2 = nullptr;
        const size_t need2 = al((n + 1) * 4) * 2 + al((n + 1) * 8) * 3 + al(n + 1) + 1024;
        XT(cudaMalloc(&slab2, need2 + need));            // what smsgate_xml_emit needs outlives this call
        ctx->xml_slab = slab2;
        {
            uint8_t* q = slab2;
            auto take2 = [&](size_t b) { uint8_t* r = q; q += al(b); return r; };
            int32_t* body_len = (int32_t*)take2((n + 1) * 4);
            int32_t* sender_len = (int32_t*)take2((n + 1) * 4);
            int64_t* date_ms = (int64_t*)take2((n + 1) * 8);
            int64_t* body_off = (int64_t*)take2((n + 1) * 8);
            int64_t* sender_off = (int64_t*)take2((n + 1) * 8);
            uint8_t* mflags = take2(n + 1);
            // carry over the tag arrays emit needs
            int64_t* k_pos = (int64_t*)take2((T + 1) * 8);
            int64_t* k_end = (int64_t*)take2((T + 1) * 8);
            int64_t* k_sms = (int64_t*)take2((n + 1) * 8);
            int32_t* k_noff = (int32_t*)take2(T * 4);
            int32_t* k_nlen = (int32_t*)take2(T * 4);
            uint8_t* k_kind = take2(T);
            XT(cudaMemcpyAsync(k_pos, tag_pos, T * 8, cudaMemcpyDeviceToDevice, st));
            XT(cudaMemcpyAsync(k_end, tag_end, T * 8, cudaMemcpyDeviceToDevice, st));
            XT(cudaMemcpyAsync(k_sms, sms_tag, (n > 0 ? n : 1) * 8, cudaMemcpyDeviceToDevice, st));
            XT(cudaMemcpyAsync(k_noff, name_off, T * 4, cudaMemcpyDeviceToDevice, st));
            XT(cudaMemcpyAsync(k_nlen, name_len, T * 4, cudaMemcpyDeviceToDevice, st));
            XT(cudaMemcpyAsync(k_kind, kind, T, cudaMemcpyDeviceToDevice, st));
            XmlTags K{k_pos, k_end, k_noff, k_nlen, k_kind, nullptr, nullptr, nullptr};
            XmlSms S{body_len, sender_len, date_ms, mflags};
            if (n > 0) {
                k_xml_sms_sizes<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_x, n, k_sms, K, ctx->d_tables, S);
                xml_scan_i32(st, body_len, n, d_tmp, body_off);
                xml_scan_i32(st, sender_len, n, d_tmp, sender_off);
                XT(cudaMemcpyAsync(&tot_body, body_off + n, 8, cudaMemcpyDeviceToHost, st));
                XT(cudaMemcpyAsync(&tot_sender, sender_off + n, 8, cudaMemcpyDeviceToHost, st));
                ctx->launches += 7;
            }
            XT(cudaStreamSynchronize(st));
            XT(cudaGetLastError());
            ctx->xml_n = n; ctx->xml_T = T; ctx->xml_nx = nbytes;
            ctx->xml_tag_pos = k_pos; ctx->xml_end = k_end; ctx->xml_sms_tag = k_sms; ctx->xml_name_off = k_noff;
            ctx->xml_name_len = k_nlen; ctx->xml_kind = k_kind; ctx->xml_body_off = body_off; ctx->xml_sender_off = sender_off;
            ctx->xml_date_ms = date_ms; ctx->xml_mflags = mflags;
            // the document stays resident for emit: park it behind the slab pointer bookkeeping
            ctx->xml_x = d_x;
            d_x = nullptr;
        }
        *n_out = n; *body_bytes = tot_body; *sender_bytes = tot_sender;
    }
done:
    cudaFree(slab); cudaFree(d_cnt); cudaFree(d_off); cudaFree(d_tmp); cudaFree(d_x);
    if (rc) { cudaFree(ctx->xml_slab); ctx->xml_slab = nullptr; ctx->xml_n = 0; }
    return rc;
#undef XT
}

extern "C" int smsgate_xml_emit(smsgate_ctx* ctx, uint8_t* d_body_bytes, int64_t* d_body_offsets, uint8_t* d_sender_bytes,
                                int64_t* d_sender_offsets, int64_t* d_date_ms, uint8_t* d_mflags) {
    if (!ctx) return SMSGATE_E_ARG;
    if (!ctx->xml_slab) return fail(ctx, SMSGATE_E_STATE, "xml_emit: no pending smsgate_xml_scan");
    const int64_t n = ctx->xml_n;
    if (!d_body_offsets || !d_sender_offsets || !d_date_ms || !d_mflags) return fail(ctx, SMSGATE_E_ARG, "xml_emit: NULL array");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (n > 0) {
        XmlTags K{ctx->xml_tag_pos, ctx->xml_end, ctx->xml_name_off, ctx->xml_name_len, ctx->xml_kind, nullptr, nullptr, nullptr};
        k_xml_sms_emit<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(ctx->xml_x, n, ctx->xml_sms_tag, K, ctx->xml_body_off,
                                                                    ctx->xml_sender_off, d_body_bytes, d_sender_bytes);
        ctx->launches += 1;
        CUDA_TRY(ctx, cudaMemcpyAsync(d_body_offsets, ctx->xml_body_off, (size_t)(n + 1) * 8, cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(d_sender_offsets, ctx->xml_sender_off, (size_t)(n + 1) * 8, cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(d_date_ms, ctx->xml_date_ms, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(d_mflags, ctx->xml_mflags, (size_t)n, cudaMemcpyDeviceToDevice, st));
    } else {
        CUDA_TRY(ctx, cudaMemsetAsync(d_body_offsets, 0, 8, st));
        CUDA_TRY(ctx, cudaMemsetAsync(d_sender_offsets, 0, 8, st));
    }
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    CUDA_TRY(ctx, cudaGetLastError());
    return smsgate_xml_release(ctx);
}

extern "C" int smsgate_xml_attrs_sizes(smsgate_ctx* ctx, int64_t* n_attrs, int64_t* name_bytes, int64_t* value_bytes) {
    if (!ctx) return SMSGATE_E_ARG;
    if (!n_attrs || !name_bytes || !value_bytes) return fail(ctx, SMSGATE_E_ARG, "xml_attrs_sizes: NULL argument");
    if (!ctx->xml_slab) return fail(ctx, SMSGATE_E_STATE, "xml_attrs_sizes: no pending smsgate_xml_scan");
    *n_attrs = *name_bytes = *value_bytes = 0;
    const int64_t n = ctx->xml_n;
    ctx->xml_attr_total = 0;
    if (n == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    cudaFree(ctx->xml_attr_slab);
    ctx->xml_attr_slab = nullptr;
    auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t need = al((n + 1) * 4) * 3 + al((n + 1) * 8) * 3 + al((size_t)(n / SCAN_ITEMS + 1024) * 8);
    CUDA_TRY(ctx, cudaMalloc(&ctx->xml_attr_slab, need));
    uint8_t* q = ctx->xml_attr_slab;
    auto take = [&](size_t b) { uint8_t* r = q; q += al(b); return r; };
    int32_t* cnt = (int32_t*)take((n + 1) * 4);
    int32_t* nb = (int32_t*)take((n + 1) * 4);
    int32_t* vb = (int32_t*)take((n + 1) * 4);
    ctx->xml_attr_off = (int64_t*)take((n + 1) * 8);
    ctx->xml_attr_nbase = (int64_t*)take((n + 1) * 8);
    ctx->xml_attr_vbase = (int64_t*)take((n + 1) * 8);
    int64_t* tmp = (int64_t*)take((size_t)(n / SCAN_ITEMS + 1024) * 8);
    XmlTags K{ctx->xml_tag_pos, ctx->xml_end, ctx->xml_name_off, ctx->xml_name_len, ctx->xml_kind, nullptr, nullptr, nullptr};
    k_xml_attr_sizes<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(ctx->xml_x, n, ctx->xml_sms_tag, K, cnt, nb, vb);
    xml_scan_i32(st, cnt, n, tmp, ctx->xml_attr_off);
    xml_scan_i32(st, nb, n, tmp, ctx->xml_attr_nbase);
    xml_scan_i32(st, vb, n, tmp, ctx->xml_attr_vbase);
    ctx->launches += 10;
    int64_t tot[3] = {0, 0, 0};
    CUDA_TRY(ctx, cudaMemcpyAsync(&tot[0], ctx->xml_attr_off + n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(&tot[1], ctx->xml_attr_nbase + n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(&tot[2], ctx->xml_attr_vbase + n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->xml_attr_total = tot[0];
    *n_attrs = tot[0]; *name_bytes = tot[1]; *value_bytes = tot[2];
    return SMSGATE_OK;
}

extern "C" int smsgate_xml_attrs_emit(smsgate_ctx* ctx, int64_t* d_attr_offsets, int64_t* d_name_offsets, uint8_t* d_name_bytes,
                                      int64_t* d_value_offsets, uint8_t* d_value_bytes) {
    if (!ctx) return SMSGATE_E_ARG;
    if (!ctx->xml_slab || ctx->xml_attr_total < 0) return fail(ctx, SMSGATE_E_STATE, "xml_attrs_emit: call smsgate_xml_attrs_sizes first");
    if (!d_attr_offsets || !d_name_offsets || !d_value_offsets) return fail(ctx, SMSGATE_E_ARG, "xml_attrs_emit: NULL array");
    const int64_t n = ctx->xml_n;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (n == 0) {
        CUDA_TRY(ctx, cudaMemsetAsync(d_attr_offsets, 0, 8, st));
        CUDA_TRY(ctx, cudaMemsetAsync(d_name_offsets, 0, 8, st));
        CUDA_TRY(ctx, cudaMemsetAsync(d_value_offsets, 0, 8, st));
    } else {
        XmlTags K{ctx->xml_tag_pos, ctx->xml_end, ctx->xml_name_off, ctx->xml_name_len, ctx->xml_kind, nullptr, nullptr, nullptr};
        k_xml_attr_emit<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(ctx->xml_x, n, ctx->xml_sms_tag, K, ctx->xml_attr_off,
                                                                     ctx->xml_attr_nbase, ctx->xml_attr_vbase, d_name_offsets,
                                                                     d_name_bytes, d_value_offsets, d_value_bytes);
        ctx->launches += 1;
        CUDA_TRY(ctx, cudaMemcpyAsync(d_attr_offsets, ctx->xml_attr_off, (size_t)(n + 1) * 8, cudaMemcpyDeviceToDevice, st));
    }
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

// ---- stages
extern "C" int smsgate_scan_keys(smsgate_ctx* ctx, const uint8_t* d_bytes, const int64_t* d_offsets, int64_t n,
                                 const smsgate_outputs* out) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || !out || (n > 0 && (!d_bytes || !d_offsets))) return fail(ctx, SMSGATE_E_ARG, "scan_keys: NULL argument");
    if (((uintptr_t)d_bytes & 15) != 0) return fail(ctx, SMSGATE_E_ARG, "scan_keys: bytes must be 16-byte aligned");
    if (n == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    ScanOut so{out->msg_id, out->cache_key, out->flags, out->body_date, out->fixed_len, out->fixed_bytes};
    // one CTA per SM as soon as there are 64 messages for each: a small batch is latency, not throughput
    const int64_t want = (n + 63) / 64;
    const int64_t maxgrid = ctx->scan_ctas > 0 ? ctx->scan_ctas : (int64_t)ctx->sm_count * K1_CTAS_PER_SM;
    const unsigned grid = (unsigned)(want < maxgrid ? want : maxgrid);
    k_scan_keys<<<grid, K1_THREADS, sizeof(K1Smem), ctx->stream>>>(d_bytes, d_offsets, n, ctx->d_tables, so);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

static int join_route_impl(smsgate_ctx* ctx, int64_t n, int64_t now_packed, const smsgate_outputs* out,
                           const DedupScratch* D, int64_t dedup_i0, const uint8_t* dedup_ids) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || !out) return fail(ctx, SMSGATE_E_ARG, "join_route: NULL argument");
    if (n == 0) return SMSGATE_OK;
    if (!out->cache_key || !out->flags || !out->body_date)
        return fail(ctx, SMSGATE_E_ARG, "join_route needs the cache_key, flags and body_date columns of stage 1");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CacheView C{ctx->c_keys, ctx->c_table, ctx->c_cap ? ctx->c_cap - 1 : 0, ctx->c_val_offsets, ctx->c_val_bytes,
                ctx->c_side, ctx->c_date_packed, ctx->cache_m};
    JoinOut jo{out->slot, out->status, out->amount_coef, out->balance_coef, out->amount_exp, out->balance_exp,
               out->dec_flags, out->date, out->card, out->card_len, out->amount_cents, out->balance_cents};
    const DedupScratch none{nullptr, nullptr, 0};
    k_join_route<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(n, out->cache_key, out->flags, out->body_date,
                                                                        now_packed, C, jo, D ? *D : none, dedup_i0,
                                                                        dedup_ids,
                                                                        ctx->phone_dt,
                                                                        (D && ctx->fuse_bits_words) ? ctx->p2p_bits : nullptr,
                                                                        ctx->fuse_bits_words);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    if (D && ctx->fuse_bits_words > 0 && ctx->p2p_bits) ctx->bits_ready = true;
    return SMSGATE_OK;
}

extern "C" int smsgate_join_route(smsgate_ctx* ctx, int64_t n, int64_t now_packed, const smsgate_outputs* out) {
    return join_route_impl(ctx, n, now_packed, out, nullptr, 0, nullptr);
}

// An empty dedup table for a batch of n messages (power of two >= 2n) inside the reserved arrays.
static int dedup_prepare(smsgate_ctx* ctx, int64_t n, DedupScratch* D) {
    uint64_t cap = pow2_at_least((uint64_t)(n > 8 ? n : 8) * 2);
    if (cap > ctx->dd_cap) cap = ctx->dd_cap;
    cudaStream_t st = ctx->stream;
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_dtab, 0xFF, cap * sizeof(int2), st));
    ctx->bits_ready = false;
    if (ctx->fuse_bits_words > 0 && ctx->p2p_bits && ctx->fuse_bits_words <= ctx->p2p_bits_words)
        CUDA_TRY(ctx, cudaMemsetAsync(ctx->p2p_bits, 0, (size_t)ctx->fuse_bits_words * 4, st));
    *D = DedupScratch{ctx->d_dtab, ctx->d_slot_of, cap - 1};
    return SMSGATE_OK;
}

extern "C" int smsgate_dedup(smsgate_ctx* ctx, int64_t n, int64_t base_index, const smsgate_outputs* out) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || !out) return fail(ctx, SMSGATE_E_ARG, "dedup: NULL argument");
    if (n == 0) return SMSGATE_OK;
    if (!out->msg_id || !out->status) return fail(ctx, SMSGATE_E_ARG, "dedup needs the msg_id and status columns");
    if (n > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "dedup: batch larger than smsgate_reserve() sized the scratch for");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    DedupScratch D;
    const int prc = dedup_prepare(ctx, n, &D);
    if (prc) return prc;
    const unsigned grid = (unsigned)((n + 255) / 256);
    k_dedup_insert<<<grid, 256, 0, st>>>(0, n, out->msg_id, out->status, D);
    k_dedup_resolve<<<grid, 256, 0, st>>>(n, base_index, D, out->survivor, out->first_seen, out->last_idx);
    ctx->launches += 2;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_device_keys(smsgate_ctx* ctx, const uint8_t* d_bytes, const int64_t* d_offsets, int64_t n,
                                   uint8_t* d_md5) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || (n > 0 && (!d_bytes || !d_offsets || !d_md5))) return fail(ctx, SMSGATE_E_ARG, "device_keys: NULL argument");
    if (((uintptr_t)d_md5 & 15) != 0) return fail(ctx, SMSGATE_E_ARG, "device_keys: md5 must be 16-byte aligned");
    if (n == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    k_md5_keys<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_bytes, d_offsets, n, d_md5);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_cross_source(smsgate_ctx* ctx, int64_t n_xml, const uint8_t* d_xml_sha1, int64_t n_dev,
                                    const uint8_t* d_dev_sha1, int32_t* d_match) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n_xml < 0 || n_dev < 0 || (n_xml > 0 && !d_xml_sha1) || (n_dev > 0 && (!d_dev_sha1 || !d_match)))
        return fail(ctx, SMSGATE_E_ARG, "cross_source: NULL argument");
    if (n_dev == 0) return SMSGATE_OK;
    if (n_xml > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "cross_source: more XML messages than smsgate_reserve() sized the scratch for");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (n_xml == 0 || !ctx->d_dtab) {
        CUDA_TRY(ctx, cudaMemsetAsync(d_match, 0xFF, (size_t)n_dev * 4, st));
        return SMSGATE_OK;
    }
    DedupScratch D;
    const int rc = dedup_prepare(ctx, n_xml, &D);
    if (rc) return rc;
    k_dedup_insert<<<(unsigned)((n_xml + 255) / 256), 256, 0, st>>>(0, n_xml, d_xml_sha1, nullptr, D);
    k_xsrc_probe<<<(unsigned)((n_dev + 255) / 256), 256, 0, st>>>(n_dev, d_dev_sha1, d_xml_sha1, D, d_match);
    ctx->launches += 2;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

// stages 2 and 3 in one call: the upsert table is cleared first, the join kernel inserts as it routes, one
// pass resolves - one launch and one pass over msg_id / status less than smsgate_join_route + smsgate_dedup
extern "C" int smsgate_join_route_dedup(smsgate_ctx* ctx, int64_t n, int64_t now_packed, int64_t base_index,
                                        const smsgate_outputs* out) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || !out) return fail(ctx, SMSGATE_E_ARG, "join_route_dedup: NULL argument");
    if (n == 0) return SMSGATE_OK;
    if (!out->msg_id || !out->status) return fail(ctx, SMSGATE_E_ARG, "join_route_dedup needs the msg_id and status columns");
    if (n > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "join_route_dedup: batch larger than smsgate_reserve() sized the scratch for");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    DedupScratch D;
    int rc = dedup_prepare(ctx, n, &D);
    if (rc) return rc;
    rc = join_route_impl(ctx, n, now_packed, out, &D, 0, out->msg_id);
    if (rc) return rc;
    k_dedup_resolve<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(n, base_index, D, out->survivor, out->first_seen,
                                                                          out->last_idx);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_replay_device(smsgate_ctx* ctx, const uint8_t* d_bytes, const int64_t* d_offsets, int64_t n,
                                     int64_t now_packed, int64_t base_index, const smsgate_outputs* out) {
    const int rc = smsgate_scan_keys(ctx, d_bytes, d_offsets, n, out);
    if (rc) return rc;
    return smsgate_join_route_dedup(ctx, n, now_packed, base_index, out);
}

// carve the context's device slab into the output columns (alignment: widest first)
static void carve_slab(uint8_t* slab, int64_t n, uint8_t* fixed, smsgate_outputs* d) {
    auto take = [&](size_t bytes) {
        uint8_t* p = slab;
        slab += (bytes + 255) & ~(size_t)255;
        return p;
    };
    d->amount_coef = (int64_t*)take((size_t)n * 8);
    d->balance_coef = (int64_t*)take((size_t)n * 8);
    d->date = (int64_t*)take((size_t)n * 8);
    d->last_idx = (int64_t*)take((size_t)n * 8);
    d->amount_cents = (int64_t*)take((size_t)n * 8);
    d->balance_cents = (int64_t*)take((size_t)n * 8);
    d->cache_key = take((size_t)n * 32);
    d->msg_id = take((size_t)n * 20);
    d->card = take((size_t)n * 16);
    d->body_date = (int32_t*)take((size_t)n * 4);
    d->fixed_len = (int32_t*)take((size_t)n * 4);
    d->slot = (int32_t*)take((size_t)n * 4);
    d->amount_exp = (int32_t*)take((size_t)n * 4);
    d->balance_exp = (int32_t*)take((size_t)n * 4);
    d->flags = take((size_t)n);
    d->status = take((size_t)n);
    d->dec_flags = take((size_t)n);
    d->card_len = take((size_t)n);
    d->survivor = take((size_t)n);
    d->first_seen = take((size_t)n);
    d->fixed_bytes = fixed;
}

// column pointers shifted to message `lo` (fixed_bytes is addressed by absolute byte offsets and stays)
static smsgate_outputs shift_columns(const smsgate_outputs& d, int64_t lo) {
    smsgate_outputs o = d;
#define SH(col, w) if (o.col) o.col = o.col + lo * (w)
    SH(msg_id, 20); SH(cache_key, 32); SH(flags, 1); SH(body_date, 1); SH(fixed_len, 1); SH(slot, 1); SH(status, 1);
    SH(amount_coef, 1); SH(balance_coef, 1); SH(amount_exp, 1); SH(balance_exp, 1); SH(dec_flags, 1); SH(date, 1);
    SH(card, 16); SH(card_len, 1); SH(survivor, 1); SH(first_seen, 1); SH(last_idx, 1);
    SH(amount_cents, 1); SH(balance_cents, 1);
#undef SH
    return o;
}

extern "C" int smsgate_device_view(smsgate_ctx* ctx, int64_t n, smsgate_outputs* d_out) {
    if (!ctx || !d_out) return SMSGATE_E_ARG;
    if (n < 0 || n > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "device_view: n exceeds the reserved scratch");
    carve_slab(ctx->s_out, n, ctx->s_fixed, d_out);
    return SMSGATE_OK;
}

// ---- packed results.  The host call returns ~10 columns; copied back chunk by chunk that is ~10 small
// transfers per chunk, and on PCIe the set-up time of small device->host copies both exceeds their transfer
// time and slows the concurrent copy-in down.  smsgate_replay_host_records returns the same values as one
// 64-byte record per message instead: one transfer per chunk.
static_assert(sizeof(smsgate_record) == 64, "smsgate_record is 64 bytes");
__global__ void __launch_bounds__(256)
k_pack_records(int64_t lo, int64_t hi, const smsgate_outputs d, smsgate_record* __restrict__ rec) {
    const int64_t i = lo + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= hi) return;
    const uint32_t* id = reinterpret_cast<const uint32_t*>(d.msg_id + i * 20);
    const uint32_t small = (uint32_t)d.status[i] | ((uint32_t)d.flags[i] << 8) | ((uint32_t)d.dec_flags[i] << 16) |
                           ((uint32_t)d.card_len[i] << 24);
    const int64_t ac = d.amount_coef[i], bc = d.balance_coef[i], dt = d.date[i];
    uint4* o = reinterpret_cast<uint4*>(rec + i);
    o[0] = make_uint4(id[0], id[1], id[2], id[3]);
    o[1] = make_uint4(id[4], small, (uint32_t)d.slot[i], (uint32_t)d.amount_exp[i]);
    o[2] = make_uint4((uint32_t)d.balance_exp[i], (uint32_t)d.body_date[i], (uint32_t)ac, (uint32_t)((uint64_t)ac >> 32));
    o[3] = make_uint4((uint32_t)bc, (uint32_t)((uint64_t)bc >> 32), (uint32_t)dt, (uint32_t)((uint64_t)dt >> 32));
}
__global__ void __launch_bounds__(256)
k_pack_verdict(int64_t n, const uint8_t* __restrict__ survivor, const uint8_t* __restrict__ first_seen,
               const int64_t* __restrict__ last_idx, int64_t* __restrict__ verdict) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t l = last_idx[i];
    verdict[i] = l < 0 ? -1 : ((l << 2) | ((int64_t)(first_seen[i] & 1) << 1) | (int64_t)(survivor[i] & 1));
}

// one byte per message instead of eight: 0x80 = the message is upserted, bit 0 survivor, bit 1 first_seen; the
// index of the surviving occurrence stays on the device (smsgate_fetch_columns: last_idx) for whoever wants it
__global__ void __launch_bounds__(256)
k_pack_verdict8(int64_t n, const uint8_t* __restrict__ survivor, const uint8_t* __restrict__ first_seen,
                const int64_t* __restrict__ last_idx, uint8_t* __restrict__ verdict) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    verdict[i] = last_idx[i] < 0 ? 0 : (uint8_t)(0x80 | ((first_seen[i] & 1) << 1) | (survivor[i] & 1));
}

static int replay_host_begin_impl(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                                  int64_t now_packed, int64_t base_index, const smsgate_outputs* h_cols,
                                  smsgate_record* h_records) {
    if (!ctx) return SMSGATE_E_ARG;
    static const smsgate_outputs no_columns = {};
    const smsgate_outputs* h_out = h_cols ? h_cols : &no_columns;
    if (n < 0 || (!h_cols && !h_records) || (n > 0 && (!h_offsets))) return fail(ctx, SMSGATE_E_ARG, "replay_host: NULL argument");
    ctx->host_n = 0;
    ctx->host_total = 0;
    ctx->view_n = ctx->view_total = 0;
    if (n == 0) return SMSGATE_OK;
    const int64_t first = h_offsets[0], total = h_offsets[n];
    if (first != 0) return fail(ctx, SMSGATE_E_ARG, "replay_host: offsets[0] must be 0");
    if (total < 0 || (total > 0 && !h_bytes)) return fail(ctx, SMSGATE_E_ARG, "replay_host: bad offsets");
    if (n > ctx->cap_msgs || total > ctx->cap_bytes)
        return fail(ctx, SMSGATE_E_STATE, "replay_host: batch larger than smsgate_reserve() sized the scratch for");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream, sin = ctx->s_h2d, sout = ctx->s_d2h;
    smsgate_outputs d;
    carve_slab(ctx->s_out, n, ctx->s_fixed, &d);
    // Chunks: copy-in of chunk c+1 and copy-out of chunk c-1 overlap the kernels of chunk c (PCIe is full
    // duplex and the copies, not the kernels, are what the end-to-end time is made of).  Every copy has a
    // few microseconds of set-up, so chunks are large where that is free - each takes a fifth of what is
    // left - and small only at the end, where the size of the last chunk is what remains to be done after
    // the last input byte has arrived.
    int64_t bound[smsgate_ctx::MAX_CHUNKS + 1];
    int nchunks = 0;
    {
        const int64_t min_chunk = ctx->host_min_chunk;
        int64_t lo = 0;
        bound[0] = 0;
        while (lo < n) {
            const int64_t rem = n - lo;
            int64_t sz = rem / ctx->host_div > min_chunk ? rem / ctx->host_div : min_chunk;
            if (nchunks == smsgate_ctx::MAX_CHUNKS - 1 || rem - sz < min_chunk / 2) sz = rem;
            lo += sz;
            bound[++nchunks] = lo;
        }
    }
    // nothing of this call may start before what the caller queued on its stream earlier
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_start, st));
    CUDA_TRY(ctx, cudaStreamWaitEvent(sin, ctx->ev_start, 0));
    CUDA_TRY(ctx, cudaStreamWaitEvent(sout, ctx->ev_start, 0));
    if (ctx->host_offsets_first)                       // all offsets in ONE copy ahead of the bodies instead of one small copy per chunk
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->s_offsets, h_offsets, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, sin));
    for (int c = 0; c < nchunks; ++c) {
        const int64_t lo = bound[c], hi = bound[c + 1];
        const int64_t b0 = h_offsets[lo], b1 = h_offsets[hi];
        if (!ctx->host_offsets_first)
            CUDA_TRY(ctx, cudaMemcpyAsync(ctx->s_offsets + lo, h_offsets + lo, (size_t)(hi - lo + 1) * 8, cudaMemcpyHostToDevice, sin));
        if (b1 > b0) CUDA_TRY(ctx, cudaMemcpyAsync(ctx->s_bytes + b0, h_bytes + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, sin));
        CUDA_TRY(ctx, cudaEventRecord(ctx->ev_in[c], sin));
    }
    // the dedup table takes each chunk's keys as soon as they exist; only the verdicts wait for the end
    DedupScratch D;
    int rc = dedup_prepare(ctx, n, &D);
    if (rc) return rc;
    int64_t back_lo = 0;                                // results of [back_lo, ...) are still on the device
    for (int c = 0; c < nchunks; ++c) {
        const int64_t lo = bound[c], hi = bound[c + 1];
        const int64_t m = hi - lo;
        CUDA_TRY(ctx, cudaStreamWaitEvent(st, ctx->ev_in[c], 0));
        const smsgate_outputs dc = shift_columns(d, lo);
        rc = smsgate_scan_keys(ctx, ctx->s_bytes, ctx->s_offsets + lo, m, &dc);
        if (rc) return rc;
        rc = join_route_impl(ctx, m, now_packed, &dc, &D, lo, d.msg_id);    // routes and inserts into the upsert table
        if (rc) return rc;
        if (h_records) {
            k_pack_records<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(lo, hi, d, ctx->s_records);
            ctx->launches += 1;
            CUDA_TRY(ctx, cudaGetLastError());
        }
        CUDA_TRY(ctx, cudaEventRecord(ctx->ev_k[c], st));
        // Results go back in few, large transfers (a quarter of the batch at a time): every column is its own
        // copy, small copies are mostly set-up time and they also slow the copy-in down.  Only the last two
        // chunks go back one by one, so that little is left to do after the last kernel.
        if (c < nchunks - 2 && (hi - back_lo) * 4 < n) continue;
        CUDA_TRY(ctx, cudaStreamWaitEvent(sout, ctx->ev_k[c], 0));
        const int64_t glo = back_lo, gm = hi - back_lo;
        back_lo = hi;
        if (h_records)
            CUDA_TRY(ctx, cudaMemcpyAsync(h_records + glo, ctx->s_records + glo, (size_t)gm * sizeof(smsgate_record), cudaMemcpyDeviceToHost, sout));
#define BACK(col, w) \
    if (h_out->col) CUDA_TRY(ctx, cudaMemcpyAsync(h_out->col + glo * (w), d.col + glo * (w), (size_t)(gm * (w)) * sizeof(*d.col), cudaMemcpyDeviceToHost, sout))
        BACK(msg_id, 20); BACK(cache_key, 32); BACK(flags, 1); BACK(body_date, 1); BACK(fixed_len, 1); BACK(slot, 1);
        BACK(status, 1); BACK(amount_coef, 1); BACK(balance_coef, 1); BACK(amount_exp, 1); BACK(balance_exp, 1);
        BACK(dec_flags, 1); BACK(date, 1); BACK(card, 16); BACK(card_len, 1); BACK(amount_cents, 1); BACK(balance_cents, 1);
#undef BACK
        if (h_out->fixed_bytes) {
            const int64_t b0 = h_offsets[glo], b1 = h_offsets[hi];
            if (b1 > b0) CUDA_TRY(ctx, cudaMemcpyAsync(h_out->fixed_bytes + b0, d.fixed_bytes + b0, (size_t)(b1 - b0), cudaMemcpyDeviceToHost, sout));
        }
    }
    k_dedup_resolve<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, base_index, D, d.survivor, d.first_seen, d.last_idx);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->host_n = n;
    ctx->host_total = total;
    ctx->host_chunks = nchunks;
    ctx->view_n = n;
    ctx->view_total = total;
    return SMSGATE_OK;
}

static int replay_host_finish_impl(smsgate_ctx* ctx, const smsgate_outputs* h_cols, int64_t* h_verdict, uint8_t* h_verdict8 = nullptr) {
    if (!ctx || (!h_cols && !h_verdict && !h_verdict8)) return SMSGATE_E_ARG;
    static const smsgate_outputs no_columns = {};
    const smsgate_outputs* h_out = h_cols ? h_cols : &no_columns;
    const int64_t n = ctx->host_n;
    if (n == 0) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream, sout = ctx->s_d2h;
    smsgate_outputs d;
    carve_slab(ctx->s_out, n, ctx->s_fixed, &d);
    if (h_verdict) {
        k_pack_verdict<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, d.survivor, d.first_seen, d.last_idx, ctx->s_verdict);
        ctx->launches += 1;
        CUDA_TRY(ctx, cudaGetLastError());
    }
    if (h_verdict8) {
        k_pack_verdict8<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, d.survivor, d.first_seen, d.last_idx,
                                                                     reinterpret_cast<uint8_t*>(ctx->s_verdict));
        ctx->launches += 1;
        CUDA_TRY(ctx, cudaGetLastError());
    }
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_dedup, st));
    CUDA_TRY(ctx, cudaStreamWaitEvent(sout, ctx->ev_dedup, 0));
    if (h_verdict) CUDA_TRY(ctx, cudaMemcpyAsync(h_verdict, ctx->s_verdict, (size_t)n * 8, cudaMemcpyDeviceToHost, sout));
    if (h_verdict8) CUDA_TRY(ctx, cudaMemcpyAsync(h_verdict8, ctx->s_verdict, (size_t)n, cudaMemcpyDeviceToHost, sout));
    if (h_out->survivor) CUDA_TRY(ctx, cudaMemcpyAsync(h_out->survivor, d.survivor, (size_t)n, cudaMemcpyDeviceToHost, sout));
    if (h_out->first_seen) CUDA_TRY(ctx, cudaMemcpyAsync(h_out->first_seen, d.first_seen, (size_t)n, cudaMemcpyDeviceToHost, sout));
    if (h_out->last_idx) CUDA_TRY(ctx, cudaMemcpyAsync(h_out->last_idx, d.last_idx, (size_t)n * 8, cudaMemcpyDeviceToHost, sout));
    // the caller's stream is complete only when the copy-out stream is
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_done, sout));
    CUDA_TRY(ctx, cudaStreamWaitEvent(st, ctx->ev_done, 0));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->s_h2d));
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->host_n = 0;
    return SMSGATE_OK;
}

extern "C" int smsgate_replay_host_begin(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                                         int64_t now_packed, int64_t base_index, const smsgate_outputs* h_out) {
    if (!h_out) return ctx ? fail(ctx, SMSGATE_E_ARG, "replay_host: NULL argument") : SMSGATE_E_ARG;
    return replay_host_begin_impl(ctx, h_bytes, h_offsets, n, now_packed, base_index, h_out, nullptr);
}
extern "C" int smsgate_replay_host_finish(smsgate_ctx* ctx, const smsgate_outputs* h_out) {
    if (!h_out) return SMSGATE_E_ARG;
    return replay_host_finish_impl(ctx, h_out, nullptr);
}
extern "C" int smsgate_replay_host_records_begin(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets,
                                                 int64_t n, int64_t now_packed, int64_t base_index,
                                                 smsgate_record* h_records) {
    if (!h_records) return ctx ? fail(ctx, SMSGATE_E_ARG, "replay_host_records: NULL argument") : SMSGATE_E_ARG;
    return replay_host_begin_impl(ctx, h_bytes, h_offsets, n, now_packed, base_index, nullptr, h_records);
}
extern "C" int smsgate_replay_host_records_finish(smsgate_ctx* ctx, int64_t* h_verdict) {
    if (!h_verdict) return SMSGATE_E_ARG;
    return replay_host_finish_impl(ctx, nullptr, h_verdict);
}
extern "C" int smsgate_replay_host_records8_finish(smsgate_ctx* ctx, uint8_t* h_verdict8) {
    if (!h_verdict8) return SMSGATE_E_ARG;
    return replay_host_finish_impl(ctx, nullptr, nullptr, h_verdict8);
}
extern "C" int smsgate_replay_host_records8(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                                            int64_t now_packed, int64_t base_index, smsgate_record* h_records,
                                            uint8_t* h_verdict8) {
    const int rc = smsgate_replay_host_records_begin(ctx, h_bytes, h_offsets, n, now_packed, base_index, h_records);
    if (rc) return rc;
    return smsgate_replay_host_records8_finish(ctx, h_verdict8);
}
extern "C" int smsgate_replay_host_records(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                                           int64_t now_packed, int64_t base_index, smsgate_record* h_records,
                                           int64_t* h_verdict) {
    const int rc = smsgate_replay_host_records_begin(ctx, h_bytes, h_offsets, n, now_packed, base_index, h_records);
    if (rc) return rc;
    return smsgate_replay_host_records_finish(ctx, h_verdict);
}

extern "C" int smsgate_fetch_columns(smsgate_ctx* ctx, int64_t n, const smsgate_outputs* h_out) {
    if (!ctx || !h_out) return SMSGATE_E_ARG;
    if (n <= 0 || n != ctx->view_n || ctx->host_n != 0)
        return fail(ctx, SMSGATE_E_STATE, "fetch_columns: the device no longer holds the columns of that batch "
                                          "(a later replay or reserve replaced them, or the call is still pending)");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    smsgate_outputs d;
    carve_slab(ctx->s_out, n, ctx->s_fixed, &d);
#define FETCH(col, w) \
    if (h_out->col) CUDA_TRY(ctx, cudaMemcpyAsync(h_out->col, d.col, (size_t)(n * (w)) * sizeof(*d.col), cudaMemcpyDeviceToHost, st))
    FETCH(msg_id, 20); FETCH(cache_key, 32); FETCH(flags, 1); FETCH(body_date, 1); FETCH(fixed_len, 1); FETCH(slot, 1);
    FETCH(status, 1); FETCH(amount_coef, 1); FETCH(balance_coef, 1); FETCH(amount_exp, 1); FETCH(balance_exp, 1);
    FETCH(dec_flags, 1); FETCH(date, 1); FETCH(card, 16); FETCH(card_len, 1); FETCH(survivor, 1); FETCH(first_seen, 1);
    FETCH(last_idx, 1); FETCH(amount_cents, 1); FETCH(balance_cents, 1);
#undef FETCH
    if (h_out->fixed_bytes && ctx->view_total > 0)
        CUDA_TRY(ctx, cudaMemcpyAsync(h_out->fixed_bytes, d.fixed_bytes, (size_t)ctx->view_total, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return SMSGATE_OK;
}

extern "C" int smsgate_replay_host_timeline(smsgate_ctx* ctx, float* ms, int capacity) {
    if (!ctx || !ms) return SMSGATE_E_ARG;
    const int nc = ctx->host_chunks;
    if (capacity < 2 * nc + 2) return fail(ctx, SMSGATE_E_ARG, "replay_host_timeline: capacity too small");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    for (int c = 0; c < nc; ++c) {
        CUDA_TRY(ctx, cudaEventElapsedTime(&ms[c], ctx->ev_start, ctx->ev_in[c]));
        CUDA_TRY(ctx, cudaEventElapsedTime(&ms[nc + c], ctx->ev_start, ctx->ev_k[c]));
    }
    CUDA_TRY(ctx, cudaEventElapsedTime(&ms[2 * nc], ctx->ev_start, ctx->ev_dedup));
    CUDA_TRY(ctx, cudaEventElapsedTime(&ms[2 * nc + 1], ctx->ev_start, ctx->ev_done));
    return nc;
}

extern "C" int smsgate_replay_host(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                                   int64_t now_packed, int64_t base_index, const smsgate_outputs* h_out) {
    const int rc = smsgate_replay_host_begin(ctx, h_bytes, h_offsets, n, now_packed, base_index, h_out);
    if (rc) return rc;
    return smsgate_replay_host_finish(ctx, h_out);
}

// ================================================================== global dedup across GPUs
// Messages are sharded by index; after the local dedup every locally distinct
// msg_id becomes one 32-byte record {key[20], first48, last48} (global indices)
// routed to the rank that owns its key prefix.  The owner merges the records of
// all ranks and answers each with the global first / last index.
struct __align__(16) KeyRecord {
    uint32_t key[5];
    uint32_t first_lo;
    uint32_t last_lo;
    uint16_t first_hi;
    uint16_t last_hi;
};
static_assert(sizeof(KeyRecord) == 32, "exchange records are 32 bytes");

__device__ __forceinline__ int key_owner(const uint8_t* k, int world) {
    // the digest's first four bytes, big-endian, are its top 32 bits: range-partition them over `world`
    const uint32_t p = ((uint32_t)k[0] << 24) | ((uint32_t)k[1] << 16) | ((uint32_t)k[2] << 8) | k[3];
    return (int)(((uint64_t)p * (uint64_t)world) >> 32);
}

// pass 1: count representatives (the message that claimed its slot) per destination.
// Counts are gathered per block in shared memory first: with 2-8 destinations a
// global atomic per record would serialise on a handful of addresses.
__global__ void __launch_bounds__(256)
k_export_count(int64_t n, const uint8_t* __restrict__ msg_id, DedupScratch D, int world, int* __restrict__ counts) {
    __shared__ int hist[SMSGATE_MAX_WORLD];
    if (threadIdx.x < SMSGATE_MAX_WORLD) hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) {
        const int32_t sl = D.slot_of[i];
        if (sl >= 0 && D.tab[sl].x == (int32_t)i) atomicAdd(&hist[key_owner(msg_id + i * 20, world)], 1);
    }
    __syncthreads();
    if (threadIdx.x < world && hist[threadIdx.x]) atomicAdd(&counts[threadIdx.x], hist[threadIdx.x]);
}
// pass 2: scatter the records into destination-contiguous order (one global atomic per block and destination)
__global__ void __launch_bounds__(256)
k_export_scatter(int64_t n, int64_t base_index, const uint8_t* __restrict__ msg_id, DedupScratch D, int world,
                 int* __restrict__ cursor, KeyRecord* __restrict__ records, int32_t* __restrict__ rec_slot) {
    __shared__ int hist[SMSGATE_MAX_WORLD];
    __shared__ int base[SMSGATE_MAX_WORLD];
    if (threadIdx.x < SMSGATE_MAX_WORLD) hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int32_t sl = -1;
    int2 fl = make_int2(0, 0);
    int dest = 0, local = 0;
    if (i < n) {
        sl = D.slot_of[i];
        if (sl >= 0 && (fl = D.tab[sl]).x != (int32_t)i) sl = -1;   // one record per key: its first message sends it
        if (sl >= 0) {
            dest = key_owner(msg_id + i * 20, world);
            local = atomicAdd(&hist[dest], 1);
        }
    }
    __syncthreads();
    if (threadIdx.x < world && hist[threadIdx.x]) base[threadIdx.x] = atomicAdd(&cursor[threadIdx.x], hist[threadIdx.x]);
    __syncthreads();
    if (sl < 0) return;
    const int r = base[dest] + local;
    const uint32_t* kw = reinterpret_cast<const uint32_t*>(msg_id + i * 20);
    const uint64_t f = (uint64_t)(base_index + fl.x), l = (uint64_t)(base_index + fl.y);
    KeyRecord rec;
    rec.key[0] = kw[0]; rec.key[1] = kw[1]; rec.key[2] = kw[2]; rec.key[3] = kw[3]; rec.key[4] = kw[4];
    rec.first_lo = (uint32_t)f; rec.last_lo = (uint32_t)l;
    rec.first_hi = (uint16_t)(f >> 32); rec.last_hi = (uint16_t)(l >> 32);
    records[r] = rec;
    rec_slot[r] = sl;
}

struct MergeScratch {
    int32_t* owner;              // [cap] record index that claimed the slot
    unsigned long long* first;   // [cap]
    unsigned long long* last;    // [cap]
    int32_t* slot_of;            // [n_recv]
    uint64_t mask;
};
__global__ void __launch_bounds__(256)
k_merge_insert(int64_t n, const KeyRecord* __restrict__ recs, MergeScratch M) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const KeyRecord r = recs[i];
    uint64_t pos = (((uint64_t)r.key[1] << 32) | r.key[0]) & M.mask;
    for (;;) {
        int prev = M.owner[pos];
        if (prev == -1) prev = atomicCAS(&M.owner[pos], -1, (int)i);
        if (prev == -1 || prev == (int)i) break;
        const KeyRecord o = recs[prev];
        if (o.key[0] == r.key[0] && o.key[1] == r.key[1] && o.key[2] == r.key[2] && o.key[3] == r.key[3] &&
            o.key[4] == r.key[4]) break;
        pos = (pos + 1) & M.mask;
    }
    atomicMin(&M.first[pos], ((unsigned long long)r.first_hi << 32) | r.first_lo);
    atomicMax(&M.last[pos], ((unsigned long long)r.last_hi << 32) | r.last_lo);
    M.slot_of[i] = (int32_t)pos;
}
__global__ void __launch_bounds__(256)
k_merge_verdict(int64_t n, MergeScratch M, int64_t* __restrict__ verdict /* [n][2] */) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t sl = M.slot_of[i];
    verdict[2 * i] = (int64_t)M.first[sl];
    verdict[2 * i + 1] = (int64_t)M.last[sl];
}
// back home: verdict r belongs to local slot rec_slot[r]
__global__ void __launch_bounds__(256)
k_apply_scatter(int64_t n_rec, const int32_t* __restrict__ rec_slot, const int64_t* __restrict__ verdict,
                int64_t* __restrict__ gfirst, int64_t* __restrict__ glast) {
    const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (r >= n_rec) return;
    const int32_t sl = rec_slot[r];
    gfirst[sl] = verdict[2 * r];
    glast[sl] = verdict[2 * r + 1];
}
__global__ void __launch_bounds__(256)
k_apply_resolve(int64_t n, int64_t base_index, const int32_t* __restrict__ slot_of,
                const int64_t* __restrict__ gfirst, const int64_t* __restrict__ glast,
                uint8_t* __restrict__ survivor, uint8_t* __restrict__ first_seen, int64_t* __restrict__ last_idx) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t sl = slot_of[i];
    uint8_t sv = 0, fs = 0;
    int64_t li = -1;
    if (sl >= 0) {
        const int64_t g = base_index + i;
        li = glast[sl];
        sv = (li == g);
        fs = (gfirst[sl] == g);
    }
    if (survivor) survivor[i] = sv;
    if (first_seen) first_seen[i] = fs;
    if (last_idx) last_idx[i] = li;
}

extern "C" int smsgate_dedup_export(smsgate_ctx* ctx, int64_t n, int64_t base_index, int world,
                                    const smsgate_outputs* out, uint8_t* d_records, int32_t* d_rec_slot,
                                    int64_t* h_counts) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || world < 1 || world > SMSGATE_MAX_WORLD || !out || !h_counts)
        return fail(ctx, SMSGATE_E_ARG, "dedup_export: bad argument");
    for (int w = 0; w < world; ++w) h_counts[w] = 0;
    if (n == 0) return SMSGATE_OK;
    if (n > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "dedup_export: run smsgate_dedup on this batch first");
    if (!out->msg_id || !d_records || !d_rec_slot) return fail(ctx, SMSGATE_E_ARG, "dedup_export: NULL array");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    uint64_t cap = pow2_at_least((uint64_t)(n > 8 ? n : 8) * 2);
    if (cap > ctx->dd_cap) cap = ctx->dd_cap;
    DedupScratch D{ctx->d_dtab, ctx->d_slot_of, cap - 1};
    const unsigned grid = (unsigned)((n + 255) / 256);
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_counts, 0, sizeof(int) * SMSGATE_MAX_WORLD, st));
    k_export_count<<<grid, 256, 0, st>>>(n, out->msg_id, D, world, ctx->d_counts);
    int counts[SMSGATE_MAX_WORLD];
    CUDA_TRY(ctx, cudaMemcpyAsync(counts, ctx->d_counts, sizeof(int) * world, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));        // the split sizes are needed on the host for the all-to-all
    int cursor[SMSGATE_MAX_WORLD];
    int run = 0;
    for (int w = 0; w < world; ++w) { cursor[w] = run; run += counts[w]; h_counts[w] = counts[w]; }
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->d_counts, cursor, sizeof(int) * world, cudaMemcpyHostToDevice, st));
    k_export_scatter<<<grid, 256, 0, st>>>(n, base_index, out->msg_id, D, world, ctx->d_counts,
                                           reinterpret_cast<KeyRecord*>(d_records), d_rec_slot);
    ctx->launches += 2;
    CUDA_TRY(ctx, cudaGetLastError());
    // `cursor` is read by an async copy: make sure it has been consumed before this frame goes away
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return SMSGATE_OK;
}

extern "C" int smsgate_reserve_merge(smsgate_ctx* ctx, int64_t max_records) {
    if (!ctx) return SMSGATE_E_ARG;
    if (max_records < 0 || max_records > 0x40000000ll) return fail(ctx, SMSGATE_E_LIMIT, "reserve_merge: at most 2^30 records");
    if (max_records <= ctx->mg_records) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->mg_owner); cudaFree(ctx->mg_first); cudaFree(ctx->mg_last); cudaFree(ctx->mg_slot_of);
    ctx->mg_owner = nullptr; ctx->mg_first = nullptr; ctx->mg_last = nullptr; ctx->mg_slot_of = nullptr;
    ctx->mg_records = 0;
    const uint64_t cap = pow2_at_least((uint64_t)(max_records > 8 ? max_records : 8) * 2);
    CUDA_TRY(ctx, cudaMalloc(&ctx->mg_owner, cap * 4));
    CUDA_TRY(ctx, cudaMalloc(&ctx->mg_first, cap * 8));
    CUDA_TRY(ctx, cudaMalloc(&ctx->mg_last, cap * 8));
    CUDA_TRY(ctx, cudaMalloc(&ctx->mg_slot_of, (size_t)(max_records + 1) * 4));
    ctx->mg_cap = cap;
    ctx->mg_records = max_records;
    return SMSGATE_OK;
}

extern "C" int smsgate_dedup_merge(smsgate_ctx* ctx, int64_t n_recv, const uint8_t* d_records, int64_t* d_verdicts) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n_recv < 0) return fail(ctx, SMSGATE_E_ARG, "dedup_merge: negative count");
    if (n_recv == 0) return SMSGATE_OK;
    if (!d_records || !d_verdicts) return fail(ctx, SMSGATE_E_ARG, "dedup_merge: NULL array");
    if (n_recv > ctx->mg_records) return fail(ctx, SMSGATE_E_STATE, "dedup_merge: more records than smsgate_reserve_merge() sized the table for");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    uint64_t cap = pow2_at_least((uint64_t)(n_recv > 8 ? n_recv : 8) * 2);
    if (cap > ctx->mg_cap) cap = ctx->mg_cap;
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->mg_owner, 0xFF, cap * 4, st));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->mg_first, 0x7F, cap * 8, st));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->mg_last, 0x00, cap * 8, st));
    MergeScratch M{ctx->mg_owner, ctx->mg_first, ctx->mg_last, ctx->mg_slot_of, cap - 1};
    const unsigned grid = (unsigned)((n_recv + 255) / 256);
    k_merge_insert<<<grid, 256, 0, st>>>(n_recv, reinterpret_cast<const KeyRecord*>(d_records), M);
    k_merge_verdict<<<grid, 256, 0, st>>>(n_recv, M, d_verdicts);
    ctx->launches += 2;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_dedup_apply(smsgate_ctx* ctx, int64_t n, int64_t base_index, int64_t n_rec,
                                   const int32_t* d_rec_slot, const int64_t* d_verdicts, const smsgate_outputs* out) {
    if (!ctx) return SMSGATE_E_ARG;
    if (n < 0 || n_rec < 0 || !out) return fail(ctx, SMSGATE_E_ARG, "dedup_apply: bad argument");
    if (n == 0) return SMSGATE_OK;
    if (n > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "dedup_apply: run smsgate_dedup on this batch first");
    if (n_rec > 0 && (!d_rec_slot || !d_verdicts)) return fail(ctx, SMSGATE_E_ARG, "dedup_apply: NULL array");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (n_rec > 0) {
        k_apply_scatter<<<(unsigned)((n_rec + 255) / 256), 256, 0, st>>>(n_rec, d_rec_slot, d_verdicts, ctx->d_gfirst,
                                                                         ctx->d_glast);
        ctx->launches += 1;
    }
    k_apply_resolve<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, base_index, ctx->d_slot_of, ctx->d_gfirst,
                                                                 ctx->d_glast, out->survivor, out->first_seen,
                                                                 out->last_idx);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

// ================================================================== the same exchange over peer memory
// One process per GPU, every rank's inbox / counters / verdict arrays mapped into
// every other rank (NVLink peer access; the Python layer gets the mappings from
// torch's symmetric memory).  The export kernel stores its records straight into
// the owner's inbox, the owner's merge stores each verdict straight back into the
// home rank's per-slot arrays: no count exchange, no host synchronisation, no
// staging copies; the only cross-GPU synchronisation is two barriers (plus two for the
// optional bitmap prefilter that keeps keys held by one rank only out of the exchange).
struct __align__(16) P2PRecord {          // three 16-byte stores over NVLink
    uint32_t key[5];
    uint32_t first_lo;
    uint32_t last_lo;
    uint16_t first_hi;
    uint16_t last_hi;
    uint32_t home_slot;
    uint32_t src_rank;
    uint32_t pad[2];
};
static_assert(sizeof(P2PRecord) == SMSGATE_P2P_RECORD_BYTES, "peer records are 48 bytes");

struct __align__(32) P2PSlot {
    int32_t owner;               // record index that claimed the slot, -1 = empty
    int32_t pad;
    unsigned long long first;    // min global first index
    unsigned long long last;     // max global last index
    unsigned long long pad2;
};

struct P2PView {                 // smsgate_p2p, by value, for the kernels
    int world, rank;
    P2PRecord* inbox[SMSGATE_MAX_WORLD];
    int32_t* in_count[SMSGATE_MAX_WORLD];
    longlong2* gfl[SMSGATE_MAX_WORLD];    // per local slot {global first, global last}
    int64_t capacity;
    uint32_t* stage[SMSGATE_MAX_WORLD];   // prefilter: staged bitmap slices, one region per sending rank
    uint32_t* collide[SMSGATE_MAX_WORLD]; // prefilter: bits set on two or more ranks
    int64_t bitmap_words;                 // 0 = no prefilter
    int32_t* flags[SMSGATE_MAX_WORLD];    // per rank [SMSGATE_MAX_WORLD] arrival flags of smsgate_p2p_barrier (may be NULL)
    uint32_t* ovr[SMSGATE_MAX_WORLD];     // per rank [2 * ovr_words]: bit per local slot "an owner overrode this slot's verdict", one half per parity
    int64_t ovr_words;
};

// ---- prefilter.  Only a key that occurs on two or more ranks needs the exchange; for every other key the
// local dedup already holds the global answer.  Each rank marks its keys in a bitmap (32 bits of the SHA-1
// that neither the table hash nor the owner prefix use), the bitmap is cut into one slice per rank, every
// rank stores slice w into rank w's staging area, rank w ORs/ANDs the slices into "seen on >= 2 ranks" and
// stores that slice back into everybody's `collide` bitmap.  A set bit can be a false positive (that key
// is exchanged for nothing), never a false negative.  Cost: two bitmaps over NVLink instead of one 48-byte
// record per distinct key - for a batch without cross-rank duplicates almost nothing is left to exchange.
// The peer-exchange kernels run in blocks of 128 threads with at most 32 registers each: one such block fits
// on an SM NEXT TO the keys kernel's CTA (640 threads x 96 registers leave 4 096 registers and ~20 KB of shared
// memory), so the exchange of batch k proceeds on every SM while the keys kernel of batch k+1 occupies them all.
constexpr int XCH_THREADS = 128;
constexpr int XCH_MIN_BLOCKS = 16;       // 65 536 registers / (16 x 128 threads) = 32 per thread

__device__ __forceinline__ int64_t slice_lo(int64_t words, int w, int world) { return (words * w / world) & ~(int64_t)3; }
__device__ __forceinline__ int64_t slice_cap(int64_t words, int world) { return ((words + world - 1) / world + 8) & ~(int64_t)3; }
__device__ __forceinline__ uint32_t key_bit(const uint8_t* key, int64_t words) {
    return reinterpret_cast<const uint32_t*>(key)[2] & (uint32_t)(words * 32 - 1);
}

__global__ void __launch_bounds__(XCH_THREADS, XCH_MIN_BLOCKS)
k_bits_build(int64_t n, const uint8_t* __restrict__ msg_id, const int32_t* __restrict__ slot_of, uint32_t* __restrict__ bits,
             int64_t words) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n || slot_of[i] < 0) return;                 // every upsert marks its key; duplicates mark the same bit
    const uint32_t h = key_bit(msg_id + i * 20, words);
    atomicOr(&bits[h >> 5], 1u << (h & 31));
}
__global__ void __launch_bounds__(XCH_THREADS, XCH_MIN_BLOCKS)
k_bits_push(const uint32_t* __restrict__ bits, P2PView P) {
    // slice w of this rank's bitmap -> rank w's staging row for this rank.  One block-stride loop per slice, two
    // 16-byte loads in flight per thread before the two peer stores.
    const int64_t cap = slice_cap(P.bitmap_words, P.world);
    for (int w = 0; w < P.world; ++w) {
        const int64_t lo = slice_lo(P.bitmap_words, w, P.world);
        const int64_t hi = w + 1 == P.world ? P.bitmap_words : slice_lo(P.bitmap_words, w + 1, P.world);
        const uint4* src = reinterpret_cast<const uint4*>(bits + lo);
        uint4* dst = reinterpret_cast<uint4*>(P.stage[w] + P.rank * cap);               // peer memory
        const int64_t nq = (hi - lo) >> 2, step = (int64_t)gridDim.x * blockDim.x;
        for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < nq; q += 2 * step) {
            uint4 v0 = src[q], v1 = make_uint4(0, 0, 0, 0);
            const bool second = q + step < nq;
            if (second) v1 = src[q + step];
            dst[q] = v0;
            if (second) dst[q + step] = v1;
        }
    }
}
__global__ void __launch_bounds__(XCH_THREADS, XCH_MIN_BLOCKS)
k_bits_reduce(P2PView P) {
    const int64_t lo = slice_lo(P.bitmap_words, P.rank, P.world);
    const int64_t hi = P.rank + 1 == P.world ? P.bitmap_words : slice_lo(P.bitmap_words, P.rank + 1, P.world);
    const int64_t cap = slice_cap(P.bitmap_words, P.world), nq = (hi - lo) >> 2;
    const uint32_t* mine = P.stage[P.rank];
    for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < nq; q += (int64_t)gridDim.x * blockDim.x) {
        uint4 one = make_uint4(0, 0, 0, 0), two = make_uint4(0, 0, 0, 0);
        for (int r0 = 0; r0 < P.world; r0 += 4) {                                       // four rows' loads in flight
            uint4 x[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                x[k] = r0 + k < P.world ? *reinterpret_cast<const uint4*>(mine + (r0 + k) * cap + (q << 2)) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                two.x |= one.x & x[k].x; two.y |= one.y & x[k].y; two.z |= one.z & x[k].z; two.w |= one.w & x[k].w;
                one.x |= x[k].x; one.y |= x[k].y; one.z |= x[k].z; one.w |= x[k].w;
            }
        }
        for (int r = 0; r < P.world; ++r) *reinterpret_cast<uint4*>(P.collide[r] + lo + (q << 2)) = two;   // peer stores
    }
}

// Records of this rank's distinct keys go straight into their owners' inboxes.  A block counts its records
// per destination in shared memory and reserves its ranges with one remote atomic per destination; with the
// prefilter only keys whose bit is set in `collide` are sent at all.
constexpr int XCH_ITEMS = 4;             // messages per thread in the per-message exchange kernels: four independent loads in
                                         // flight per thread - these kernels mostly run squeezed next to the keys kernel, one
                                         // small block per SM, where memory parallelism per thread is all the parallelism there is
__global__ void __launch_bounds__(XCH_THREADS, XCH_MIN_BLOCKS)
k_p2p_scatter(int64_t n, int64_t base_index, int parity, const uint8_t* __restrict__ msg_id, DedupScratch D, P2PView P,
              const uint32_t* __restrict__ collide, int* __restrict__ overflow) {
    __shared__ int hist[SMSGATE_MAX_WORLD];
    __shared__ int base[SMSGATE_MAX_WORLD];
    if (threadIdx.x < SMSGATE_MAX_WORLD) hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t i0 = (int64_t)blockIdx.x * (XCH_THREADS * XCH_ITEMS) + threadIdx.x;
    // pass 1 (all messages, cheap): is the message an upsert whose key some other rank may hold?
    int32_t sl[XCH_ITEMS];
    uint32_t hb[XCH_ITEMS];
#pragma unroll
    for (int k = 0; k < XCH_ITEMS; ++k) {
        const int64_t i = i0 + (int64_t)k * XCH_THREADS;
        sl[k] = i < n ? D.slot_of[i] : -1;
        hb[k] = (i < n && collide) ? key_bit(msg_id + i * 20, P.bitmap_words) : 0;
    }
    uint32_t cand = 0;
#pragma unroll
    for (int k = 0; k < XCH_ITEMS; ++k)
        if (sl[k] >= 0 && (!collide || ((collide[hb[k] >> 5] >> (hb[k] & 31)) & 1u))) cand |= 1u << k;
    // pass 2 (the few candidates): one record per key - its first message sends it
    uint32_t dest_packed = 0;                            // 8 bits per item: the owner rank (< SMSGATE_MAX_WORLD = 64)
#pragma unroll
    for (int k = 0; k < XCH_ITEMS; ++k) {
        if (!((cand >> k) & 1u)) continue;
        const int64_t i = i0 + (int64_t)k * XCH_THREADS;
        if (D.tab[sl[k]].x != (int32_t)i) { cand &= ~(1u << k); continue; }
        const int dest = key_owner(msg_id + i * 20, P.world);
        hb[k] = (uint32_t)atomicAdd(&hist[dest], 1);     // the record's place inside this block's range at that owner
        dest_packed |= (uint32_t)dest << (8 * k);
    }
    __syncthreads();
    if (threadIdx.x < P.world && hist[threadIdx.x]) {
        const int c = hist[threadIdx.x];
        int b = atomicAdd_system(&P.in_count[threadIdx.x][parity], c);       // remote when the owner is a peer
        if ((int64_t)b + c > P.capacity) { atomicExch(overflow, 1); b = -1; }
        base[threadIdx.x] = b;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < XCH_ITEMS; ++k) {
        if (!((cand >> k) & 1u)) continue;
        const int dest = (int)((dest_packed >> (8 * k)) & 255u);
        if (base[dest] < 0) continue;
        const int64_t i = i0 + (int64_t)k * XCH_THREADS;
        const int2 fl = D.tab[sl[k]];
        const uint64_t f = (uint64_t)(base_index + fl.x), l = (uint64_t)(base_index + fl.y);
        const uint32_t* kw = reinterpret_cast<const uint32_t*>(msg_id + i * 20);
        P2PRecord rec;
        rec.key[0] = kw[0]; rec.key[1] = kw[1]; rec.key[2] = kw[2]; rec.key[3] = kw[3]; rec.key[4] = kw[4];
        rec.first_lo = (uint32_t)f; rec.last_lo = (uint32_t)l;
        rec.first_hi = (uint16_t)(f >> 32); rec.last_hi = (uint16_t)(l >> 32);
        rec.home_slot = (uint32_t)sl[k];
        rec.src_rank = (uint32_t)P.rank;
        rec.pad[0] = rec.pad[1] = 0;
        const uint4* src = reinterpret_cast<const uint4*>(&rec);
        uint4* dst = reinterpret_cast<uint4*>(&P.inbox[dest][base[dest] + (int)hb[k]]);   // peer stores over NVLink when dest != rank
        dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
    }
}

__device__ __forceinline__ int p2p_count(const int32_t* count_ptr, int64_t capacity) {
    const int c = *count_ptr;                   // a refused reservation still advanced the counter
    return (int64_t)c > capacity ? (int)capacity : c;
}
__device__ __forceinline__ uint64_t p2p_table_mask(int count) {
    uint64_t want = (uint64_t)(count > 8 ? count : 8) * 2, cap = 16;
    while (cap < want) cap <<= 1;
    return cap - 1;
}
__global__ void __launch_bounds__(XCH_THREADS, XCH_MIN_BLOCKS)
k_p2p_clear(const int32_t* __restrict__ count_ptr, int64_t capacity, P2PSlot* __restrict__ table) {
    const uint64_t cap = p2p_table_mask(p2p_count(count_ptr, capacity)) + 1;
    for (uint64_t s = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; s < cap; s += (uint64_t)gridDim.x * blockDim.x) {
        P2PSlot e;
        e.owner = -1; e.pad = 0; e.first = ~0ull; e.last = 0; e.pad2 = 0;
        table[s] = e;
    }
}
__global__ void __launch_bounds__(XCH_THREADS, XCH_MIN_BLOCKS)
k_p2p_insert(const int32_t* __restrict__ count_ptr, int64_t capacity, const P2PRecord* __restrict__ recs,
             P2PSlot* __restrict__ table, int32_t* __restrict__ slot_of) {
    const int count = p2p_count(count_ptr, capacity);
    const uint64_t mask = p2p_table_mask(count);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const P2PRecord r = recs[i];
        uint64_t pos = (((uint64_t)r.key[1] << 32) | r.key[0]) & mask;
        for (;;) {
            int prev = table[pos].owner;
            if (prev == -1) prev = atomicCAS(&table[pos].owner, -1, (int)i);
            if (prev == -1 || prev == (int)i) break;
            const P2PRecord o = recs[prev];
            if (o.key[0] == r.key[0] && o.key[1] == r.key[1] && o.key[2] == r.key[2] && o.key[3] == r.key[3] &&
                o.key[4] == r.key[4]) break;
            pos = (pos + 1) & mask;
        }
        atomicMin(&table[pos].first, ((unsigned long long)r.first_hi << 32) | r.first_lo);
        atomicMax(&table[pos].last, ((unsigned long long)r.last_hi << 32) | r.last_lo);
        slot_of[i] = (int32_t)pos;
    }
}
__global__ void __launch_bounds__(XCH_THREADS, XCH_MIN_BLOCKS)
k_p2p_verdict(const int32_t* __restrict__ count_ptr, const P2PRecord* __restrict__ recs,
              const P2PSlot* __restrict__ table, const int32_t* __restrict__ slot_of, P2PView P, int parity) {
    const int count = p2p_count(count_ptr, P.capacity);
    // the arrival counter of the NEXT batch: nobody scatters into it before the barrier that follows this kernel
    if (blockIdx.x == 0 && threadIdx.x == 0) P.in_count[P.rank][parity ^ 1] = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const P2PSlot e = table[slot_of[i]];
        const P2PRecord r = recs[i];
        const unsigned long long f = ((unsigned long long)r.first_hi << 32) | r.first_lo;
        const unsigned long long l = ((unsigned long long)r.last_hi << 32) | r.last_lo;
        // only a key that another rank also holds changes the home rank's view: one 16-byte peer store
        if (e.first != f || e.last != l) {
            P.gfl[r.src_rank][r.home_slot] = make_longlong2((long long)e.first, (long long)e.last);
            atomicOr_system(&P.ovr[r.src_rank][(int64_t)parity * P.ovr_words + (r.home_slot >> 5)], 1u << (r.home_slot & 31));
        }
    }
}

__global__ void __launch_bounds__(XCH_THREADS, XCH_MIN_BLOCKS)
k_p2p_resolve(int64_t n, int64_t base_index, const int32_t* __restrict__ slot_of, const longlong2* __restrict__ gfl,
              uint32_t* __restrict__ ovr, int64_t ovr_words, int parity,
              uint8_t* __restrict__ survivor, uint8_t* __restrict__ first_seen, int64_t* __restrict__ last_idx) {
    // The columns already hold this rank's local verdicts (k_dedup_resolve).  Only a key that another rank also
    // holds AND whose global first / last differs from the local one was answered by its owner: those slots have
    // their bit set in this parity's half of `ovr`, and only their messages are rewritten.  The other half is
    // cleared here for the next batch (its owners write it after the next batch's barriers).
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    uint32_t* next_half = ovr + (int64_t)(parity ^ 1) * ovr_words;
    for (int64_t w = t; w < ovr_words; w += (int64_t)gridDim.x * blockDim.x) next_half[w] = 0;
    const uint32_t* half = ovr + (int64_t)parity * ovr_words;
    const int64_t i0 = (int64_t)blockIdx.x * (XCH_THREADS * XCH_ITEMS) + threadIdx.x;
    int32_t sl[XCH_ITEMS];
#pragma unroll
    for (int k = 0; k < XCH_ITEMS; ++k) {
        const int64_t i = i0 + (int64_t)k * XCH_THREADS;
        sl[k] = i < n ? slot_of[i] : -1;
    }
#pragma unroll
    for (int k = 0; k < XCH_ITEMS; ++k) {
        if (sl[k] < 0 || !((half[sl[k] >> 5] >> (sl[k] & 31)) & 1u)) continue;
        const int64_t i = i0 + (int64_t)k * XCH_THREADS;
        const longlong2 v = gfl[sl[k]];
        const int64_t g = base_index + i;
        if (survivor) survivor[i] = (v.y == g);
        if (first_seen) first_seen[i] = (v.x == g);
        if (last_idx) last_idx[i] = v.y;
    }
}

// Cross-GPU barrier from inside the stream, without a host round trip and small enough (one warp-sized block, a
// dozen registers) to be scheduled next to anything: thread w stores `epoch` into rank w's flag for this rank
// (a system-scope release after the fence: everything this rank's earlier kernels stored into peer memory is
// visible before the flag is), then waits until rank w's flag for it has reached `epoch`.  Epochs only grow, the
// flags are never reset.  The wait gives up after about ten seconds and raises the overflow flag (value 2)
// rather than hang the GPU: ranks must reach a barrier within that time of each other.
__global__ void __launch_bounds__(SMSGATE_MAX_WORLD)
k_p2p_barrier(P2PView P, int epoch, int* __restrict__ overflow) {
    const int w = threadIdx.x;
    if (w >= P.world) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(P.flags[w] + P.rank), "r"(epoch) : "memory");
    const int32_t* mine = P.flags[P.rank] + w;
    const long long t0 = clock64();
    for (;;) {
        int v;
        asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
        if (v - epoch >= 0) break;
        if (clock64() - t0 > 20000000000ll) { atomicExch(overflow, 2); break; }    // ~10 s at 2 GHz
        __nanosleep(200);
    }
}

static int p2p_view(smsgate_ctx* ctx, const smsgate_p2p* p, P2PView* v) {
    if (!p || p->world < 1 || p->world > SMSGATE_MAX_WORLD || p->rank < 0 || p->rank >= p->world || p->capacity < 1)
        return fail(ctx, SMSGATE_E_ARG, "p2p: bad descriptor");
    v->world = p->world; v->rank = p->rank; v->capacity = p->capacity;
    for (int w = 0; w < p->world; ++w) {
        if (!p->inbox[w] || !p->in_count[w] || !p->gfl[w]) return fail(ctx, SMSGATE_E_ARG, "p2p: NULL peer pointer");
        v->inbox[w] = reinterpret_cast<P2PRecord*>(p->inbox[w]);
        v->in_count[w] = p->in_count[w];
        v->gfl[w] = reinterpret_cast<longlong2*>(p->gfl[w]);
    }
    for (int w = 0; w < p->world; ++w) {
        v->flags[w] = p->flags[w];
        if (!p->ovr[w] || p->ovr_words < 1) return fail(ctx, SMSGATE_E_ARG, "p2p: NULL override bitmap");
        v->ovr[w] = p->ovr[w];
    }
    v->ovr_words = p->ovr_words;
    v->bitmap_words = p->bitmap_words;
    if (p->bitmap_words) {
        if (p->bitmap_words < 64 || (p->bitmap_words & (p->bitmap_words - 1)) || p->bitmap_words > (1ll << 26))
            return fail(ctx, SMSGATE_E_ARG, "p2p: bitmap_words must be a power of two in [64, 2^26]");
        for (int w = 0; w < p->world; ++w) {
            if (!p->stage[w] || !p->collide[w]) return fail(ctx, SMSGATE_E_ARG, "p2p: NULL prefilter pointer");
            v->stage[w] = p->stage[w];
            v->collide[w] = p->collide[w];
        }
    }
    return SMSGATE_OK;
}

extern "C" int smsgate_p2p_reserve(smsgate_ctx* ctx, int64_t capacity) {
    if (!ctx) return SMSGATE_E_ARG;
    if (capacity < 0 || capacity > 0x40000000ll) return fail(ctx, SMSGATE_E_LIMIT, "p2p_reserve: at most 2^30 records");
    if (capacity <= ctx->p2p_capacity) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->p2p_table); cudaFree(ctx->p2p_slot_of); cudaFree(ctx->p2p_overflow);
    ctx->p2p_table = nullptr; ctx->p2p_slot_of = nullptr; ctx->p2p_overflow = nullptr; ctx->p2p_capacity = 0;
    const uint64_t cap = pow2_at_least((uint64_t)(capacity > 8 ? capacity : 8) * 2);
    CUDA_TRY(ctx, cudaMalloc(&ctx->p2p_table, cap * sizeof(P2PSlot)));
    CUDA_TRY(ctx, cudaMalloc(&ctx->p2p_slot_of, (size_t)(capacity + 1) * 4));
    CUDA_TRY(ctx, cudaMalloc(&ctx->p2p_overflow, 4));
    CUDA_TRY(ctx, cudaMemset(ctx->p2p_overflow, 0, 4));
    ctx->p2p_capacity = capacity;
    return SMSGATE_OK;
}

extern "C" int smsgate_p2p_scatter(smsgate_ctx* ctx, int64_t n, int64_t base_index, int parity,
                                   const smsgate_outputs* out, const smsgate_p2p* p) {
    if (!ctx) return SMSGATE_E_ARG;
    P2PView V;
    int rc = p2p_view(ctx, p, &V);
    if (rc) return rc;
    if (n < 0 || !out || (n > 0 && !out->msg_id)) return fail(ctx, SMSGATE_E_ARG, "p2p_scatter: bad argument");
    if (n > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "p2p_scatter: run smsgate_dedup on this batch first");
    if (!ctx->p2p_overflow) return fail(ctx, SMSGATE_E_STATE, "p2p_scatter: call smsgate_p2p_reserve first");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    uint64_t cap = pow2_at_least((uint64_t)(n > 8 ? n : 8) * 2);
    if (cap > ctx->dd_cap) cap = ctx->dd_cap;
    if (n > 0 && (p->gfl_slots <= 0 || cap > (uint64_t)p->gfl_slots))
        return fail(ctx, SMSGATE_E_LIMIT, "p2p_scatter: the batch needs " + std::to_string((unsigned long long)cap) +
                                              " verdict slots, the descriptor's gfl arrays hold " +
                                              std::to_string((long long)p->gfl_slots));
    if (n > 0 && (uint64_t)p->ovr_words * 32 < cap)
        return fail(ctx, SMSGATE_E_LIMIT, "p2p_scatter: the descriptor's override bitmaps are smaller than the batch's table");
    if (n > 0 && n > p->capacity / p->world)
        return fail(ctx, SMSGATE_E_LIMIT, "p2p_scatter: the batch exceeds capacity / world messages per rank");
    DedupScratch D{ctx->d_dtab, ctx->d_slot_of, cap - 1};
    if (n > 0) {
        k_p2p_scatter<<<(unsigned)((n + XCH_THREADS * XCH_ITEMS - 1) / (XCH_THREADS * XCH_ITEMS)), XCH_THREADS, 0, st>>>(n, base_index, parity & 1, out->msg_id, D, V,
                                                                   V.bitmap_words ? V.collide[V.rank] : nullptr,
                                                                   ctx->p2p_overflow);
        ctx->launches += 1;
    }
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_p2p_filter_push(smsgate_ctx* ctx, int64_t n, const smsgate_outputs* out, const smsgate_p2p* p) {
    if (!ctx) return SMSGATE_E_ARG;
    P2PView V;
    int rc = p2p_view(ctx, p, &V);
    if (rc) return rc;
    if (!V.bitmap_words) return fail(ctx, SMSGATE_E_ARG, "p2p_filter_push: the descriptor has no prefilter bitmaps");
    if (n < 0 || !out || (n > 0 && !out->msg_id)) return fail(ctx, SMSGATE_E_ARG, "p2p_filter_push: bad argument");
    if (n > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "p2p_filter_push: run smsgate_dedup on this batch first");
    if (V.bitmap_words > ctx->p2p_bits_words) return fail(ctx, SMSGATE_E_STATE, "p2p_filter_push: call smsgate_p2p_reserve_filter first");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const bool fused = ctx->bits_ready && ctx->fuse_bits_words == V.bitmap_words;    // the join kernel marked the keys already
    ctx->bits_ready = false;
    if (!fused) {
        CUDA_TRY(ctx, cudaMemsetAsync(ctx->p2p_bits, 0, (size_t)V.bitmap_words * 4, st));
        if (n > 0) k_bits_build<<<(unsigned)((n + XCH_THREADS - 1) / XCH_THREADS), XCH_THREADS, 0, st>>>(n, out->msg_id, ctx->d_slot_of, ctx->p2p_bits, V.bitmap_words);
    }
    k_bits_push<<<(unsigned)ctx->sm_count * 4, XCH_THREADS, 0, st>>>(ctx->p2p_bits, V);
    ctx->launches += (n > 0 && !fused) ? 2 : 1;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_p2p_filter_reduce(smsgate_ctx* ctx, const smsgate_p2p* p) {
    if (!ctx) return SMSGATE_E_ARG;
    P2PView V;
    int rc = p2p_view(ctx, p, &V);
    if (rc) return rc;
    if (!V.bitmap_words) return fail(ctx, SMSGATE_E_ARG, "p2p_filter_reduce: the descriptor has no prefilter bitmaps");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    k_bits_reduce<<<(unsigned)ctx->sm_count * 2, XCH_THREADS, 0, ctx->stream>>>(V);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_p2p_barrier(smsgate_ctx* ctx, const smsgate_p2p* p, int epoch) {
    if (!ctx) return SMSGATE_E_ARG;
    P2PView V;
    int rc = p2p_view(ctx, p, &V);
    if (rc) return rc;
    for (int w = 0; w < V.world; ++w)
        if (!V.flags[w]) return fail(ctx, SMSGATE_E_ARG, "p2p_barrier: the descriptor carries no flags arrays");
    if (!ctx->p2p_overflow) return fail(ctx, SMSGATE_E_STATE, "p2p_barrier: call smsgate_p2p_reserve first");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    k_p2p_barrier<<<1, SMSGATE_MAX_WORLD, 0, ctx->stream>>>(V, epoch, ctx->p2p_overflow);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_p2p_fuse_filter(smsgate_ctx* ctx, int64_t bitmap_words) {
    if (!ctx) return SMSGATE_E_ARG;
    if (bitmap_words < 0 || bitmap_words > ctx->p2p_bits_words)
        return fail(ctx, SMSGATE_E_STATE, "p2p_fuse_filter: call smsgate_p2p_reserve_filter for that many words first");
    ctx->fuse_bits_words = bitmap_words;
    ctx->bits_ready = false;
    return SMSGATE_OK;
}

// The whole exchange of one batch queued by ONE call (the steps are the entry points above, in the order
// dist.GlobalDedupP2P used to issue them from Python): a multi-GPU step is launch-bound on the host otherwise.
extern "C" int smsgate_p2p_exchange(smsgate_ctx* ctx, int64_t n, int64_t base_index, int parity, int epoch_base,
                                    const smsgate_outputs* out, const smsgate_p2p* p) {
    if (!ctx || !p) return SMSGATE_E_ARG;
    int rc, e = epoch_base;
    if (p->bitmap_words) {
        if ((rc = smsgate_p2p_filter_push(ctx, n, out, p))) return rc;
        if ((rc = smsgate_p2p_barrier(ctx, p, ++e))) return rc;
        if ((rc = smsgate_p2p_filter_reduce(ctx, p))) return rc;
        if ((rc = smsgate_p2p_barrier(ctx, p, ++e))) return rc;
    }
    if ((rc = smsgate_p2p_scatter(ctx, n, base_index, parity, out, p))) return rc;
    if ((rc = smsgate_p2p_barrier(ctx, p, ++e))) return rc;
    if ((rc = smsgate_p2p_merge(ctx, parity, p))) return rc;
    if ((rc = smsgate_p2p_barrier(ctx, p, ++e))) return rc;
    if ((rc = smsgate_p2p_resolve(ctx, n, base_index, parity, out, p, nullptr))) return rc;
    return e - epoch_base;                       // barriers used: the caller's next epoch_base is this much further
}

extern "C" int smsgate_p2p_reserve_filter(smsgate_ctx* ctx, int64_t bitmap_words) {
    if (!ctx) return SMSGATE_E_ARG;
    if (bitmap_words < 0 || bitmap_words > (1ll << 26)) return fail(ctx, SMSGATE_E_LIMIT, "p2p_reserve_filter: at most 2^26 words");
    if (bitmap_words <= ctx->p2p_bits_words) return SMSGATE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->p2p_bits);
    ctx->p2p_bits = nullptr; ctx->p2p_bits_words = 0;
    CUDA_TRY(ctx, cudaMalloc(&ctx->p2p_bits, (size_t)bitmap_words * 4));
    ctx->p2p_bits_words = bitmap_words;
    return SMSGATE_OK;
}

extern "C" int smsgate_p2p_merge(smsgate_ctx* ctx, int parity, const smsgate_p2p* p) {
    if (!ctx) return SMSGATE_E_ARG;
    P2PView V;
    int rc = p2p_view(ctx, p, &V);
    if (rc) return rc;
    if (V.capacity > ctx->p2p_capacity) return fail(ctx, SMSGATE_E_STATE, "p2p_merge: inbox larger than smsgate_p2p_reserve() sized the table for");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int32_t* cnt = V.in_count[V.rank] + (parity & 1);
    const unsigned grid = (unsigned)ctx->sm_count * 8;
    k_p2p_clear<<<grid, XCH_THREADS, 0, st>>>(cnt, V.capacity, ctx->p2p_table);
    k_p2p_insert<<<grid, XCH_THREADS, 0, st>>>(cnt, V.capacity, V.inbox[V.rank], ctx->p2p_table, ctx->p2p_slot_of);
    k_p2p_verdict<<<grid, XCH_THREADS, 0, st>>>(cnt, V.inbox[V.rank], ctx->p2p_table, ctx->p2p_slot_of, V, parity & 1);
    ctx->launches += 3;
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

extern "C" int smsgate_p2p_resolve(smsgate_ctx* ctx, int64_t n, int64_t base_index, int parity, const smsgate_outputs* out,
                                   const smsgate_p2p* p, int* h_overflow) {
    if (!ctx) return SMSGATE_E_ARG;
    P2PView V;
    int rc = p2p_view(ctx, p, &V);
    if (rc) return rc;
    if (n < 0 || !out) return fail(ctx, SMSGATE_E_ARG, "p2p_resolve: bad argument");
    if (n > ctx->cap_msgs) return fail(ctx, SMSGATE_E_STATE, "p2p_resolve: run smsgate_dedup on this batch first");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (n > 0) {
        k_p2p_resolve<<<(unsigned)((n + XCH_THREADS * XCH_ITEMS - 1) / (XCH_THREADS * XCH_ITEMS)), XCH_THREADS, 0, st>>>(n, base_index, ctx->d_slot_of, V.gfl[V.rank],
                                                                   V.ovr[V.rank], V.ovr_words, parity & 1,
                                                                   out->survivor, out->first_seen, out->last_idx);
        ctx->launches += 1;
    }
    if (h_overflow) {       // optional: costs a synchronisation
        CUDA_TRY(ctx, cudaMemcpyAsync(h_overflow, ctx->p2p_overflow, 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
    }
    CUDA_TRY(ctx, cudaGetLastError());
    return SMSGATE_OK;
}

// ================================================================== dedup without any allocation
// Some callers of this library are not allowed to allocate device memory at
// the point where they need the upsert-by-msg_id verdicts: a replay loop whose
// memory budget was fixed (and fully handed out) at start-up, a process in
// which another component owns the CUDA allocator, or code that must not
// touch the allocator because an allocation there can stall or fail.  For
// them the hash table that k_dedup_insert / k_dedup_resolve need lives in the
// file-scope array below instead of in a context's cudaMalloc'ed scratch.
// The price: every GPU that loads this library carries the array
// (SG_STATIC_BYTES, 20 MiB) until the process exits, and a batch is
// limited to SMSGATE_DEDUP_STATIC_MAX messages.
constexpr int64_t SG_STATIC_MAX_N = SMSGATE_DEDUP_STATIC_MAX;              // 1 Mi messages
constexpr uint64_t SG_STATIC_CAP = 2ull * SMSGATE_DEDUP_STATIC_MAX;        // table slots (power of two)
constexpr uint64_t SG_STATIC_WORDS = 2 * SG_STATIC_CAP + SG_STATIC_MAX_N;  // slots {first, last} | slot_of
constexpr uint64_t SG_STATIC_BYTES = SG_STATIC_WORDS * 4;
static_assert((SG_STATIC_CAP & (SG_STATIC_CAP - 1)) == 0, "table capacity must be a power of two");
static_assert(SG_STATIC_BYTES <= (32ull << 20), "the built-in table is meant to stay modest (<= 32 MiB per GPU)");

static __device__ __align__(256) int32_t g_dedup_static[SG_STATIC_WORDS];

// In host code `g_dedup_static` is only a handle for the runtime, not a device
// address; the address each device gave the array is looked up once per device.
static int32_t* g_dedup_static_addr[64] = {nullptr};
static std::mutex g_dedup_static_mutex;

static int static_fail(char* errbuf, int errbuf_len, int code, const char* what, const char* detail) {
    if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "smsgate_dedup_static: %s%s%s", what,
                                           detail ? ": " : "", detail ? detail : "");
    return code;
}

extern "C" int smsgate_dedup_static(int device, void* cuda_stream, int64_t n, int64_t base_index,
                                    const uint8_t* d_msg_id, const uint8_t* d_status, uint8_t* d_survivor,
                                    uint8_t* d_first_seen, int64_t* d_last_idx, char* errbuf, int errbuf_len) {
    if (n < 0 || (n > 0 && (!d_msg_id || !d_status)))
        return static_fail(errbuf, errbuf_len, SMSGATE_E_ARG, "msg_id and status are required", nullptr);
    if (n > SG_STATIC_MAX_N) {
        char detail[160];
        snprintf(detail, sizeof detail, "%lld messages, the built-in table holds at most %lld per call; "
                 "split the batch or use smsgate_dedup with a reserved context",
                 (long long)n, (long long)SG_STATIC_MAX_N);
        return static_fail(errbuf, errbuf_len, SMSGATE_E_LIMIT, "batch too large", detail);
    }
    if (device < 0 || device >= 64) return static_fail(errbuf, errbuf_len, SMSGATE_E_ARG, "device index out of range", nullptr);
    if (n == 0) return SMSGATE_OK;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return static_fail(errbuf, errbuf_len, SMSGATE_E_CUDA, "cudaSetDevice", cudaGetErrorString(e));
    int32_t* base = nullptr;
    {
        // first call on this device: ask the runtime where the array lives here (that device is current)
        std::lock_guard<std::mutex> lock(g_dedup_static_mutex);
        if (!g_dedup_static_addr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_dedup_static);
            if (e != cudaSuccess || !p)
                return static_fail(errbuf, errbuf_len, SMSGATE_E_CUDA, "cudaGetSymbolAddress",
                                   e != cudaSuccess ? cudaGetErrorString(e) : "null address");
            g_dedup_static_addr[device] = static_cast<int32_t*>(p);
        }
        base = g_dedup_static_addr[device];
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const uint64_t cap = pow2_at_least((uint64_t)(n > 8 ? n : 8) * 2);       // <= SG_STATIC_CAP
    DedupScratch D;
    D.tab = reinterpret_cast<int2*>(base);
    D.slot_of = base + 2 * SG_STATIC_CAP;
    D.mask = cap - 1;
    // the array still holds the previous call's table: reset the part this call uses (all bits set = empty)
    if ((e = cudaMemsetAsync(D.tab, 0xFF, cap * sizeof(int2), st)) != cudaSuccess)
        return static_fail(errbuf, errbuf_len, SMSGATE_E_CUDA, "cudaMemsetAsync", cudaGetErrorString(e));
    const unsigned grid = (unsigned)((n + 255) / 256);
    k_dedup_insert<<<grid, 256, 0, st>>>(0, n, d_msg_id, d_status, D);
    k_dedup_resolve<<<grid, 256, 0, st>>>(n, base_index, D, d_survivor, d_first_seen, d_last_idx);
    if ((e = cudaGetLastError()) != cudaSuccess)
        return static_fail(errbuf, errbuf_len, SMSGATE_E_CUDA, "kernel launch", cudaGetErrorString(e));
    return SMSGATE_OK;
}
